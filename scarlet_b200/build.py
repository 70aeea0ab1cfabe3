"""Build libscarlet_b200.so (hand-written sm_100a CUDA + the C ABI) in-tree."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "api.cu")
LIB = os.path.join(HERE, "lib", "libscarlet_b200.so")
SOURCES = ["api.cu", "api_batch.inc", "api_noalloc.inc", "api_pybind.inc", "api_f64.inc", "k_f64.cu", "k_forward.cu", "k_grad.cu", "k_prox.cu", "k_prox_split.cu", "k_lowres.cu", "k_scene.cu", "k_psf.cu", "k_products.cu", "k_update.cu", "prox_dev.cuh", "fft_smem.cuh",
           "fft_reg.cuh", "common.cuh"]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(HERE, "csrc", f) for f in SOURCES] + [os.path.join(HERE, "..", "include", "scarlet_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
           "-Xcompiler", "-fPIC", "-shared", "-o", LIB, SRC]
    cmd[1:1] = os.environ.get("SCL_NVCC_FLAGS", "").split()       # experiments: -DSCL_PROX_WIDE=256 ...
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.check_call(cmd, cwd=os.path.join(HERE, "csrc"))
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force=True, verbose="-v" in sys.argv))
