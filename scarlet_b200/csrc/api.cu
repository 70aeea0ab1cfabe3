// C ABI of libscarlet_b200.so (see include/scarlet_b200.h).  Single translation
// unit: the kernels are included so that one nvcc invocation builds the library.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <map>
#include <vector>
#include <utility>
#include <mutex>

#include "../../include/scarlet_b200.h"
#include "k_forward.cu"
#include "k_grad.cu"
#include "k_update.cu"
#include "k_prox.cu"
#include "k_prox_split.cu"
#include "k_lowres.cu"
#include "k_scene.cu"
#include "k_psf.cu"
#include "k_products.cu"
#include "k_f64.cu"

static_assert(SCL_FFT == SCL_FFT_EDGE && SCL_RS == SCL_SPEC_COLS, "header / kernel constants diverged");
static_assert(SCL_F_SYMMETRIC == SCL_SYMMETRIC && SCL_F_MONOTONIC == SCL_MONOTONIC && SCL_F_FIX_SED == SCL_FIX_SED &&
              SCL_F_FIX_MORPH == SCL_FIX_MORPH && SCL_F_KSPACE == SCL_KSPACE && SCL_F_CONSTRAINED == SCL_CONSTRAINED &&
              SCL_F_SCHEDULE == SCL_SCHEDULE && SCL_SCHED_MAX == SCL_SCHED_OPS,
              "flag constants diverged");

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, const char* a = "", const char* b = "") {
    snprintf(g_err, sizeof(g_err), fmt, a, b);
    return code;
}

#define CK(call)                                                                        \
    do {                                                                                \
        cudaError_t e_ = (call);                                                        \
        if (e_ != cudaSuccess) return fail(SCL_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

constexpr size_t kFwdSmem = (size_t)(SCL_FFT * SCL_RS + 256 + 2 * SCL_RS) * sizeof(float2) + 32 * sizeof(double);

int next_fast_len(int target) {     // smallest 2^a 3^b 5^c >= target (scipy.fftpack.next_fast_len, fft.py:99)
    if (target <= 6) return target;
    int best = 1 << 30;
    for (long p5 = 1; p5 < 2L * target; p5 *= 5)
        for (long p35 = p5; p35 < 2L * target; p35 *= 3) {
            long n = p35;
            while (n < target) n *= 2;
            if (n < best) best = (int)n;
        }
    return best;
}

struct Tables {
    float* wtab = nullptr;       // [9][(2Ny-1)(2Nx-1)][8]
    double* ctab = nullptr;      // [(2Ny-1)(2Nx-1)][8]
    unsigned* plist = nullptr;   // sweep plans of every centre (prox_dev.cuh), 32 entries per pass
    unsigned* pfirst = nullptr;  // [N + 1] first pass of each centre
    int pbits = 0;               // 0: no plan (shape too large): generic monotonic path
    float4* wlev = nullptr;      // [8][wlevE] level-ordered four-tap weights of the wide sweep (prox_dev.cuh); nullptr: none
    int wlevE = 0;
    SclMonoPlan plan() const { SclMonoPlan p; p.list = plist; p.first = pfirst; p.pbits = pbits; return p; }
    void release() { cudaFree(wtab); cudaFree(ctab); cudaFree(plist); cudaFree(pfirst); cudaFree(wlev); }
};
constexpr size_t kWlevBudget = (size_t)128 << 20;    // bytes of level-ordered weights per shape
constexpr size_t kPlanBudget = (size_t)768 << 20;    // bytes of sweep plans per shape
constexpr int kNflN = 4096;
constexpr int kMaxDev = 64;
// Everything the library keeps on a device between calls -- operator tables, next_fast_len tables, the kernels'
// shared-memory opt-ins -- belongs to ONE device: the state is kept per CUDA device and looked up through the
// calling thread's current device, so a process that switches devices (scl_device_set) gets that device's own copy.
struct DevState {
    std::map<std::pair<int, int>, Tables> tables;
    int* nfl = nullptr;          // device [kNflN]
    int* nfl_even = nullptr;     // device [kNflN]
    bool attr_done = false;
};
std::mutex g_mu;
DevState g_dev[kMaxDev];

DevState* dev_state() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= kMaxDev) return nullptr;
    return &g_dev[d];
}
int* cur_nfl() { DevState* st = dev_state(); return st ? st->nfl : nullptr; }
int* cur_nfl_even() { DevState* st = dev_state(); return st ? st->nfl_even : nullptr; }

int ensure_globals() {
    DevState* st = dev_state();
    if (!st) return fail(SCL_E_CUDA, "no current CUDA device (or device index above 63)%s%s");
    std::lock_guard<std::mutex> lk(g_mu);
    if (!st->nfl) {
        std::vector<int> a(kNflN), b(kNflN);
        for (int n = 0; n < kNflN; n++) {
            a[n] = next_fast_len(n);
            int e = a[n];
            while (e % 2 != 0) e = next_fast_len(e + 1);      // fft.py:103-105
            b[n] = e;
        }
        int *pa = nullptr, *pb = nullptr;
        CK(cudaMalloc(&pa, kNflN * sizeof(int)));
        CK(cudaMalloc(&pb, kNflN * sizeof(int)));
        CK(cudaMemcpy(pa, a.data(), kNflN * sizeof(int), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(pb, b.data(), kNflN * sizeof(int), cudaMemcpyHostToDevice));
        st->nfl = pa;
        st->nfl_even = pb;
    }
    if (!st->attr_done) {
        CK(cudaFuncSetAttribute(scl_k_forward_adjoint<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFwdSmem));
        CK(cudaFuncSetAttribute(scl_k_forward_adjoint<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFwdSmem));
        CK(cudaFuncSetAttribute(scl_k_convolve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFwdSmem));
        CK(cudaFuncSetAttribute(scl_k_scene_conv<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFwdSmem));
        CK(cudaFuncSetAttribute(scl_k_scene_conv<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFwdSmem));
        CK(cudaFuncSetAttribute(scl_k_scene_conv<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFwdSmem));
        const int big = 200 * 1024;
        CK(cudaFuncSetAttribute(scl_k_prox, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CK(cudaFuncSetAttribute(scl_k_source_center, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        // scene-mode variants: symmetry tables only, but those pass 48 KB once an image edge reaches ~1480 pixels
        CK(cudaFuncSetAttribute(scl_k_prox_wide, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CK(cudaFuncSetAttribute(scl_k_source_center_wide, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CK(cudaFuncSetAttribute(scl_k_symmetry, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CK(cudaFuncSetAttribute(scl_k_measure, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CK(cudaFuncSetAttribute(scl_k_monotonic<float, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CK(cudaFuncSetAttribute(scl_k_monotonic<double, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CK(cudaFuncSetAttribute(scl_k_monotonic<double, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        // set once to the limit psf_prepare checks against (a per-call value raced between host threads)
        CK(cudaFuncSetAttribute(scl_k_match_psfs, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        CK(cudaFuncSetAttribute(scl_k_mono_plan, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        st->attr_done = true;
    }
    return SCL_OK;
}

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t count) {
        n = count;
        return cudaMalloc(&p, (count ? count : 1) * sizeof(T));
    }
    cudaError_t up(const T* h, cudaStream_t s) { return cudaMemcpyAsync(p, h, n * sizeof(T), cudaMemcpyHostToDevice, s); }
    cudaError_t down(T* h, cudaStream_t s) const { return cudaMemcpyAsync(h, p, n * sizeof(T), cudaMemcpyDeviceToHost, s); }
};

// Sweep plans of every centre of the shape (prox_dev.cuh "precomputed per-centre plan"): count, prefix on the
// host, emit.  Shapes whose plans would not fit the entry format or the budget keep pbits = 0.
size_t prox_smem(int Ny, int Nx);
int build_plans(int Ny, int Nx, Tables* t) {
    const int N = Ny * Nx;
    if (Ny < 2 || Nx < 2 || Ny > 256 || Nx > 256) return SCL_OK;
    if (prox_smem(Ny, Nx) > 200 * 1024) return SCL_OK;       // scene-mode shapes: generic level loop
    int pbits = 1;
    while ((1 << pbits) < N) pbits++;
    const unsigned long long rows = 9ull * (2 * Ny - 1) * (2 * Nx - 1);
    if (rows >= (1ull << (32 - pbits)) - 1) return SCL_OK;
    const int mx = Ny > Nx ? Ny : Nx;
    const size_t smem = scl_mono_work_bytes(N, 3 * mx) + (3 * mx + 3) * sizeof(int) + 16;
    DevBuf<unsigned> cnt;
    CK(cnt.alloc(N));
    scl_k_mono_plan<<<N, SCL_THREADS, smem>>>(Ny, Nx, pbits, cnt.p, nullptr, nullptr);
    CK(cudaGetLastError());
    std::vector<unsigned> h(N), first(N + 1);
    CK(cudaMemcpy(h.data(), cnt.p, N * sizeof(unsigned), cudaMemcpyDeviceToHost));
    unsigned long long run = 0;
    for (int i = 0; i < N; i++) { first[i] = (unsigned)run; run += h[i]; }
    first[N] = (unsigned)run;
    if (run * 128 > kPlanBudget || run >= (1ull << 32)) return SCL_OK;
    CK(cudaMalloc(&t->pfirst, (N + 1) * sizeof(unsigned)));
    CK(cudaMalloc(&t->plist, (size_t)(run + 4) * 128));          // + 4 passes: the sweep copies entries four passes at a time
    CK(cudaMemcpy(t->pfirst, first.data(), (N + 1) * sizeof(unsigned), cudaMemcpyHostToDevice));
    scl_k_mono_plan<<<N, SCL_THREADS, smem>>>(Ny, Nx, pbits, nullptr, t->pfirst, t->plist);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    t->pbits = pbits;
    return SCL_OK;
}

int get_tables(int Ny, int Nx, Tables* t) {
    DevState* st = dev_state();
    if (!st) return fail(SCL_E_CUDA, "no current CUDA device%s%s");
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = st->tables.find({Ny, Nx});
    if (it == st->tables.end())
        return fail(SCL_E_ARG, "no monotonic tables registered for this shape on the current device (call scl_set_monotonic_tables)%s%s");
    *t = it->second;
    return SCL_OK;
}

size_t prox_smem(int Ny, int Nx) {
    // pad [(Nx+4)&~3] | X [N] (rounded to 16 bytes) | Z [Ny][(Nx+3)&~3] | tables; Z + tables >= SCL_MONO_STREAM_BYTES
    int mx = Ny > Nx ? Ny : Nx;
    size_t n4 = ((size_t)Ny * Nx + 3) & ~(size_t)3;
    size_t zt = ((size_t)Ny * ((Nx + 3) & ~3) + 3 * (2 * mx - 1) + 2 * Nx + 32) * sizeof(float);
    if (zt < SCL_MONO_STREAM_BYTES) zt = SCL_MONO_STREAM_BYTES;
    return (((Nx + 4) & ~3) + n4) * sizeof(float) + zt;
}

// shared memory of the constraint kernels in scene mode: the symmetry tables only
size_t prox_smem_scene(int Ny, int Nx) {
    int mx = Ny > Nx ? Ny : Nx;
    return (size_t)(3 * (2 * mx - 1) + 2 * Nx + 32) * sizeof(float);
}
// floats of HBM scratch per CTA in scene mode (layout of prox_smem without the tables, + a back pad)
size_t prox_scratch_floats(int Ny, int Nx) {
    size_t n4 = ((size_t)Ny * Nx + 3) & ~(size_t)3;
    return (size_t)((Nx + 4) & ~3) + n4 + (size_t)Ny * ((Nx + 3) & ~3) + (size_t)((Nx + 8) & ~3);
}
bool fits_smem(int Ny, int Nx) { return prox_smem(Ny, Nx) <= 200 * 1024; }

int check_shape(int Ny, int Nx, bool allow_big = false) {
    if (Ny < 1 || Nx < 1) return fail(SCL_E_ARG, "empty image shape%s%s");
    if (2 * Ny + 10 >= kNflN || 2 * Nx + 10 >= kNflN) return fail(SCL_E_LIMIT, "image edge above 2042 pixels%s%s");
    if (allow_big) return SCL_OK;
    if (prox_smem(Ny, Nx) > 200 * 1024)
        return fail(SCL_E_LIMIT, "image shape exceeds the shared-memory resident limit%s%s");
    return SCL_OK;
}


// device core: x [n][Ny][Nx] and centers [n][2] are DEVICE arrays; stream-ordered, no copies, no synchronisation
template <typename T, bool NEAREST>
static int launch_monotonic(T* dx, int n, int Ny, int Nx, const int32_t* dcenters, T thresh, cudaStream_t s) {
    int rc = check_shape(Ny, Nx, true);
    if (rc) return rc;
    if ((rc = ensure_globals())) return rc;
    const size_t msm = (size_t)Ny * Nx * sizeof(T) + (Nx + 1) * sizeof(T) + 32 + SCL_MONO_STREAM_BYTES;
    const int in_place = msm > 200 * 1024;
    Tables t;
    if ((rc = get_tables(Ny, Nx, &t))) return rc;
    if (in_place)
        scl_k_monotonic<T, NEAREST><<<n, SCL_THREADS, 0, s>>>(dx, dcenters, Ny, Nx, thresh, t.wtab, t.ctab, t.plan(), 1);
    else
        scl_k_monotonic<T, NEAREST><<<n, SCL_PROX_THREADS, msm, s>>>(dx, dcenters, Ny, Nx, thresh, t.wtab, t.ctab, t.plan(), 0);
    CK(cudaGetLastError());
    return SCL_OK;
}

template <typename T, bool NEAREST>
static int run_monotonic(T* x, int n, int Ny, int Nx, const int32_t* centers, T thresh, void* stream) {
    if (n == 0) return SCL_OK;
    if (!x || !centers || n < 0) return fail(SCL_E_ARG, "bad argument%s%s");
    int rc = check_shape(Ny, Nx, true);
    if (rc) return rc;
    for (int i = 0; i < n; i++)
        if (centers[2 * i] < 0 || centers[2 * i] >= Ny || centers[2 * i + 1] < 0 || centers[2 * i + 1] >= Nx)
            return fail(SCL_E_ARG, "center outside the image%s%s");
    cudaStream_t s = (cudaStream_t)stream;
    size_t N = (size_t)Ny * Nx;
    DevBuf<T> dx;
    DevBuf<int> dc;
    CK(dx.alloc(n * N));
    CK(dc.alloc(2 * (size_t)n));
    CK(dx.up(x, s));
    CK(dc.up(centers, s));
    if ((rc = launch_monotonic<T, NEAREST>(dx.p, n, Ny, Nx, dc.p, thresh, s))) return rc;
    CK(dx.down(x, s));
    CK(cudaStreamSynchronize(s));
    return SCL_OK;
}

}  // namespace

extern "C" {

int scl_version(void) { return 100; }
const char* scl_last_error(void) { return g_err; }
int scl_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int scl_device_get(int* device) {
    if (!device) return fail(SCL_E_ARG, "null device pointer%s%s");
    CK(cudaGetDevice(device));
    return SCL_OK;
}
int scl_device_set(int device) {
    CK(cudaSetDevice(device));
    return SCL_OK;
}
int scl_stream_create(void** stream) {
    if (!stream) return fail(SCL_E_ARG, "null stream pointer%s%s");
    cudaStream_t s;
    CK(cudaStreamCreate(&s));
    *stream = (void*)s;
    return SCL_OK;
}
int scl_stream_destroy(void* stream) {
    if (stream) CK(cudaStreamDestroy((cudaStream_t)stream));
    return SCL_OK;
}

// Level-ordered four-tap weights of the wide float32 sweep (prox_dev.cuh: scl_k_build_wlev), derived on the device
// from the interior table just uploaded.  Left out (the sweep then reads the eight-tap table) when they would exceed
// kWlevBudget or when the table carries weight outside an octant's four candidate taps.
static int build_wlev(int Ny, int Nx, Tables* t) {
    const int amax_tab = (Ny > Nx ? Ny : Nx) - 1;
    if (amax_tab < 1) return SCL_OK;
    const int nv = 3 * amax_tab;
    std::vector<int> start(nv + 2, 0);
    for (int v = 1; v <= nv; v++) start[v + 1] = start[v] + scl_wlev_count(v, amax_tab);
    const int E = start[nv + 1];
    if ((size_t)E * 8 * sizeof(float4) > kWlevBudget) return SCL_OK;
    DevBuf<int> dstart, dbad;
    CK(dstart.alloc(start.size()));
    CK(dbad.alloc(1));
    CK(cudaMemcpy(dstart.p, start.data(), start.size() * sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMemset(dbad.p, 0, sizeof(int)));
    CK(cudaMalloc(&t->wlev, (size_t)E * 8 * sizeof(float4)));
    scl_k_build_wlev<<<dim3((amax_tab + 128) / 128, nv, 8), 128>>>(t->wtab, dstart.p, t->wlev, E, Ny, Nx, amax_tab, dbad.p);
    CK(cudaGetLastError());
    int bad = 0;
    CK(cudaMemcpy(&bad, dbad.p, sizeof(int), cudaMemcpyDeviceToHost));
    if (bad) { cudaFree(t->wlev); t->wlev = nullptr; return SCL_OK; }
    t->wlevE = E;
    return SCL_OK;
}

int scl_set_monotonic_tables(int Ny, int Nx, const float* wtab, const double* ctab) {
    int rc = check_shape(Ny, Nx, true);
    if (rc) return rc;
    if (!wtab || !ctab) return fail(SCL_E_ARG, "null table%s%s");
    if ((rc = ensure_globals())) return rc;
    DevState* st = dev_state();
    size_t n = (size_t)(2 * Ny - 1) * (2 * Nx - 1) * 8;
    Tables t;
    CK(cudaMalloc(&t.wtab, 9 * n * sizeof(float)));
    CK(cudaMalloc(&t.ctab, n * sizeof(double)));
    CK(cudaMemcpy(t.wtab, wtab, 9 * n * sizeof(float), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(t.ctab, ctab, n * sizeof(double), cudaMemcpyHostToDevice));
    if ((rc = build_plans(Ny, Nx, &t))) { t.release(); return rc; }
    if ((rc = build_wlev(Ny, Nx, &t))) { t.release(); return rc; }
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = st->tables.find({Ny, Nx});
    if (it != st->tables.end()) it->second.release();
    st->tables[{Ny, Nx}] = t;
    return SCL_OK;
}

int scl_prox_weighted_monotonic_f32(float* x, int n, int Ny, int Nx, const int32_t* c, float thresh, void* stream) {
    return run_monotonic<float, false>(x, n, Ny, Nx, c, thresh, stream);
}
int scl_prox_weighted_monotonic_f64(double* x, int n, int Ny, int Nx, const int32_t* c, double thresh, void* stream) {
    return run_monotonic<double, false>(x, n, Ny, Nx, c, thresh, stream);
}
int scl_prox_monotonic_f64(double* x, int n, int Ny, int Nx, const int32_t* c, double thresh, void* stream) {
    return run_monotonic<double, true>(x, n, Ny, Nx, c, thresh, stream);
}

int scl_prox_symmetry_f32(float* x, int n, int Ny, int Nx, const int32_t* centers, const double* shifts,
                          int mode, float strength, void* stream) {
    if (n == 0) return SCL_OK;
    if (!x || !centers || n < 0 || mode < 0 || mode > 2) return fail(SCL_E_ARG, "bad argument%s%s");
    if (mode == 0 && !shifts) return fail(SCL_E_ARG, "kspace symmetry needs shifts%s%s");
    int rc = check_shape(Ny, Nx, true);
    if (rc) return rc;
    if ((rc = ensure_globals())) return rc;
    for (int i = 0; i < n; i++) {
        if (centers[2 * i] == -1 && centers[2 * i + 1] == -1) continue;      // whole-array window
        if (centers[2 * i] < 0 || centers[2 * i] >= Ny || centers[2 * i + 1] < 0 || centers[2 * i + 1] >= Nx)
            return fail(SCL_E_ARG, "center outside the image%s%s");
    }
    cudaStream_t s = (cudaStream_t)stream;
    size_t N = (size_t)Ny * Nx;
    DevBuf<float> dx, dz;
    DevBuf<int> dc;
    DevBuf<double> ds;
    const bool big = !fits_smem(Ny, Nx);     // in place in HBM; the tile loops may read (never use) up to 8 rows past an image
    CK(dx.alloc(n * N + (big ? 8 * (size_t)Nx + 16 : 0)));
    CK(dc.alloc(2 * (size_t)n));
    CK(ds.alloc(2 * (size_t)n));
    CK(cudaMemcpyAsync(dx.p, x, n * N * sizeof(float), cudaMemcpyHostToDevice, s));
    if (big) CK(cudaMemsetAsync(dx.p + n * N, 0, (8 * (size_t)Nx + 16) * sizeof(float), s));
    CK(dc.up(centers, s));
    if (shifts) CK(ds.up(shifts, s)); else CK(cudaMemsetAsync(ds.p, 0, 2 * (size_t)n * sizeof(double), s));
    if (big) {
        CK(dz.alloc((size_t)n * Ny * ((Nx + 3) & ~3) + 16));
        scl_k_symmetry<<<n, SCL_PROX_THREADS, prox_smem_scene(Ny, Nx), s>>>(dx.p, dc.p, ds.p, Ny, Nx, mode, strength,
                                                                           cur_nfl(), cur_nfl_even(), dz.p);
    } else
        scl_k_symmetry<<<n, SCL_PROX_THREADS, prox_smem(Ny, Nx), s>>>(dx.p, dc.p, ds.p, Ny, Nx, mode, strength, cur_nfl(),
                                                                     cur_nfl_even(), nullptr);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(x, dx.p, n * N * sizeof(float), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return SCL_OK;
}

int scl_measure_f32(const float* x, int n, int Ny, int Nx, int32_t* centers, double* shifts,
                    const float* psf, int ph, int pw, int do_max, int do_centroid, void* stream) {
    if (n == 0) return SCL_OK;
    if (!x || !centers || !shifts || n < 0 || (do_centroid && (!psf || ph < 1 || pw < 1)))
        return fail(SCL_E_ARG, "bad argument%s%s");
    int rc = check_shape(Ny, Nx, true);
    if (rc) return rc;
    if ((rc = ensure_globals())) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    size_t N = (size_t)Ny * Nx;
    const int in_place = N * sizeof(float) > 200 * 1024;
    DevBuf<float> dx, dp;
    DevBuf<int> dc;
    DevBuf<double> ds;
    CK(dx.alloc(n * N));
    CK(dp.alloc(psf ? (size_t)ph * pw : 1));
    CK(dc.alloc(2 * (size_t)n));
    CK(ds.alloc(2 * (size_t)n));
    CK(cudaMemcpyAsync(dx.p, x, n * N * sizeof(float), cudaMemcpyHostToDevice, s));
    if (psf) CK(dp.up(psf, s));
    CK(dc.up(centers, s));
    scl_k_measure<<<n, SCL_THREADS, in_place ? 0 : N * sizeof(float), s>>>(dx.p, dc.p, ds.p, Ny, Nx, dp.p, ph, pw, do_max,
                                                                            do_centroid, in_place);
    CK(cudaGetLastError());
    CK(dc.down(centers, s));
    CK(ds.down(shifts, s));
    CK(cudaStreamSynchronize(s));
    return SCL_OK;
}

int scl_convolve_f32(const float* model, float* out, int n, int C, int Ny, int Nx, const float* spec, int n_ker,
                     const int32_t* ker_idx, int adjoint, int ker_half, void* stream) {
    if (n == 0) return SCL_OK;
    if (!model || !out || !ker_idx || n < 0 || C < 1 || (n_ker > 0 && !spec)) return fail(SCL_E_ARG, "bad argument%s%s");
    if (C > SCL_MAX_BANDS) return fail(SCL_E_LIMIT, "1 <= C <= 8 bands supported%s%s");
    for (int i = 0; i < n * C; i++)
        if (ker_idx[i] >= n_ker) return fail(SCL_E_ARG, "ker_idx out of range%s%s");
    // frames that do not fit one 128-point period together with the kernel: overlap-save tiles (k_scene.cu)
    const bool tiled = ker_half > 0 && (Ny + ker_half > SCL_FFT || Nx + ker_half > SCL_FFT);
    if (!tiled && (Ny > SCL_FFT || Nx > SCL_FFT))
        return fail(SCL_E_LIMIT, "frame larger than the 128x128 convolution tile: pass ker_half%s%s");
    if (tiled && 2 * ker_half >= SCL_FFT - 8) return fail(SCL_E_LIMIT, "difference kernel too wide for the 128-point tile%s%s");
    int rc = ensure_globals();
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    size_t tot = (size_t)n * C * Ny * Nx;
    DevBuf<float> dm, dout;
    DevBuf<float2> dspec;
    DevBuf<int> dk;
    CK(dm.alloc(tot));
    CK(dout.alloc(tot));
    CK(dspec.alloc((size_t)(n_ker > 0 ? n_ker : 1) * SCL_FFT * SCL_RS));
    CK(dk.alloc((size_t)n * C));
    CK(cudaMemcpyAsync(dm.p, model, tot * sizeof(float), cudaMemcpyHostToDevice, s));
    if (n_ker > 0) CK(cudaMemcpyAsync(dspec.p, spec, dspec.n * sizeof(float2), cudaMemcpyHostToDevice, s));
    CK(dk.up(ker_idx, s));
    if (tiled) {
        const int T = SCL_FFT - 2 * ker_half, ty = (Ny + T - 1) / T, tx = (Nx + T - 1) / T;
        SclPlanes pl;
        pl.P = C;
        for (int c = 0; c < SCL_MAX_BANDS; c++) pl.band[c] = c;
        DevBuf<int> act;
        CK(act.alloc(n));
        std::vector<int> ones(n, 1);
        CK(cudaMemcpyAsync(act.p, ones.data(), n * sizeof(int), cudaMemcpyHostToDevice, s));
        if (adjoint)
            scl_k_scene_conv<false, true><<<dim3(ty * tx, C, n), SCL_THREADS, kFwdSmem, s>>>(
                dm.p, dspec.p, dk.p, nullptr, nullptr, act.p, dout.p, nullptr, C, pl, Ny, Nx, ker_half, tx, C, 0);
        else
            scl_k_scene_conv<false, false><<<dim3(ty * tx, C, n), SCL_THREADS, kFwdSmem, s>>>(
                dm.p, dspec.p, dk.p, nullptr, nullptr, act.p, dout.p, nullptr, C, pl, Ny, Nx, ker_half, tx, C, 0);
        CK(cudaGetLastError());
        CK(dout.down(out, s));
        CK(cudaStreamSynchronize(s));      // `ones` is a host temporary
        return SCL_OK;
    }
    scl_k_convolve<<<dim3(C, n), SCL_THREADS, kFwdSmem, s>>>(dm.p, dspec.p, dk.p, dout.p, C, Ny, Nx, adjoint);
    CK(cudaGetLastError());
    CK(dout.down(out, s));
    CK(cudaStreamSynchronize(s));
    return SCL_OK;
}

}  // extern "C"

int scl_lowres_render_f32(const float* model, float* out, int n, int C, int N, int n_lr, const float* op, void* stream) {
    if (n == 0) return SCL_OK;
    if (!model || !out || !op || n < 0 || C < 1 || N < 1 || n_lr < 1) return fail(SCL_E_ARG, "bad argument%s%s");
    int rc = ensure_globals();
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    DevBuf<float> dm, dop, dout;
    CK(dm.alloc((size_t)n * C * N)); CK(dop.alloc((size_t)C * n_lr * N)); CK(dout.alloc((size_t)n * C * n_lr));
    CK(dm.up(model, s)); CK(dop.up(op, s));
    scl_k_lowres_render<<<dim3(n_lr, C, n), 128, 0, s>>>(dm.p, dop.p, dout.p, C, N, n_lr);
    CK(cudaGetLastError());
    CK(dout.down(out, s));
    CK(cudaStreamSynchronize(s));
    return SCL_OK;
}

// Shapes, placement and shared-memory size of one scl_k_match_psfs launch (fft.py:87-106 for the FFT lengths of the
// match, fft.convolve's padding for the placement of the truncated kernel on the 128-point grid).
int psf_prepare(SclPsfArgs& a, int C, int ky, int kx, int kym, int kxm, int n_model, int padding, bool spec, int Ny, int Nx,
                size_t* smem, int* ker_half) {
    a.C = C; a.ky = ky; a.kx = kx; a.kym = kym; a.kxm = kxm; a.n_model = n_model;
    a.Fy = next_fast_len(ky + kym + padding);
    a.Fx = next_fast_len(kx + kxm + padding);
    while (a.Fx % 2 != 0) a.Fx = next_fast_len(a.Fx + 1);
    a.oy = ky; a.ox = kx;                                              // fft.py:297-300 (the stack with more bands)
    a.offy = a.offx = 0;
    if (spec) {
        int Fcy = next_fast_len(Ny + ky + 3), Fcx = next_fast_len(Nx + kx + 3);
        while (Fcx % 2 != 0) Fcx = next_fast_len(Fcx + 1);
        a.offy = scl_psf_pos(Fcy, ky);
        a.offx = scl_psf_pos(Fcx, kx);
        int half = -a.offy;
        if (ky - 1 + a.offy > half) half = ky - 1 + a.offy;
        if (-a.offx > half) half = -a.offx;
        if (kx - 1 + a.offx > half) half = kx - 1 + a.offx;
        if (2 * half >= SCL_FFT_EDGE - 8) return fail(SCL_E_LIMIT, "difference kernel too wide for the 128-point tile%s%s");
        if (ker_half) *ker_half = half;
    }
    const SclPsfLayout L = scl_psf_layout(ky, kx, kym, kxm, a.Fy, a.Fx, a.oy, a.ox, SCL_PSF_THREADS / 32);
    *smem = L.total * sizeof(double);
    if (*smem > 227 * 1024) return fail(SCL_E_LIMIT, "PSF stamps too large for the shared-memory transform%s%s");
    return SCL_OK;
}

int scl_match_psfs_f32(const float* psf_obs, int n, int C, int ky, int kx, const float* psf_model, int n_model,
                       int kym, int kxm, int padding, float* diff, float* spec, int Ny, int Nx, void* stream) {
    if (n == 0) return SCL_OK;
    if (!psf_obs || !psf_model || n < 0 || C < 1 || ky < 1 || kx < 1 || kym < 1 || kxm < 1 || padding < 0 ||
        (n_model != 1 && n_model != n) || (!diff && !spec))
        return fail(SCL_E_ARG, "bad argument%s%s");
    if (spec && (Ny < 1 || Nx < 1)) return fail(SCL_E_ARG, "the spectrum needs the frame shape%s%s");
    int rc = ensure_globals();
    if (rc) return rc;
    SclPsfArgs a;
    size_t smem = 0;
    rc = psf_prepare(a, C, ky, kx, kym, kxm, n_model, padding, spec != nullptr, Ny, Nx, &smem, nullptr);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    DevBuf<float> d1, d2, dd, ds;
    CK(d1.alloc((size_t)n * C * ky * kx)); CK(d2.alloc((size_t)n_model * kym * kxm));
    CK(d1.up(psf_obs, s)); CK(d2.up(psf_model, s));
    if (diff) CK(dd.alloc((size_t)n * C * ky * kx));
    if (spec) CK(ds.alloc((size_t)n * C * SCL_FFT_EDGE * SCL_SPEC_COLS * 2));
    a.psf_obs = d1.p; a.psf_model = d2.p; a.diff = diff ? dd.p : nullptr; a.spec = spec ? (float2*)ds.p : nullptr;
    scl_k_match_psfs<<<dim3(C, n), SCL_PSF_THREADS, smem, s>>>(a);
    CK(cudaGetLastError());
    if (diff) CK(dd.down(diff, s));
    if (spec) CK(ds.down(spec, s));
    CK(cudaStreamSynchronize(s));
    return SCL_OK;
}

int scl_trim_f32(const float* x, int n, int Ny, int Nx, float min_value, int32_t* bounds, int32_t* edge, double* sum,
                 void* stream) {
    if (n == 0) return SCL_OK;
    if (!x || n < 0 || Ny < 1 || Nx < 1) return fail(SCL_E_ARG, "bad argument%s%s");
    int rc = ensure_globals();
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    DevBuf<float> dx;
    DevBuf<int> db, de;
    DevBuf<double> dsum;
    CK(dx.alloc((size_t)n * Ny * Nx)); CK(db.alloc((size_t)4 * n)); CK(de.alloc(n)); CK(dsum.alloc(n));
    CK(dx.up(x, s));
    scl_k_trim<<<n, 256, 0, s>>>(dx.p, Ny, Nx, min_value, db.p, de.p, dsum.p);
    CK(cudaGetLastError());
    if (bounds) CK(db.down(bounds, s));
    if (edge) CK(de.down(edge, s));
    if (sum) CK(dsum.down(sum, s));
    CK(cudaStreamSynchronize(s));
    return SCL_OK;
}

int scl_prox_sparse_f32(float* x, long long n, float thresh, int mode, void* stream) {
    if (n == 0) return SCL_OK;
    if (!x || n < 0 || (mode != 0 && mode != 1)) return fail(SCL_E_ARG, "bad argument%s%s");
    int rc = ensure_globals();
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    DevBuf<float> dx;
    CK(dx.alloc((size_t)n));
    CK(dx.up(x, s));
    scl_k_sparse<<<(unsigned)(((size_t)n + 255) / 256), 256, 0, s>>>(dx.p, (size_t)n, thresh, mode);
    CK(cudaGetLastError());
    CK(dx.down(x, s));
    CK(cudaStreamSynchronize(s));
    return SCL_OK;
}

int scl_threshold_f32(float* x, int n, int Ny, int Nx, int apply, double* thresh, int32_t* bins, int32_t* bounds,
                      void* stream) {
    if (n == 0) return SCL_OK;
    if (!x || n < 0 || Ny < 1 || Nx < 1) return fail(SCL_E_ARG, "bad argument%s%s");
    int rc = ensure_globals();
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    DevBuf<float> dx;
    DevBuf<double> dt;
    DevBuf<int> db, dbo;
    CK(dx.alloc((size_t)n * Ny * Nx)); CK(dt.alloc(n)); CK(db.alloc(n)); CK(dbo.alloc((size_t)4 * n));
    CK(dx.up(x, s));
    scl_k_threshold<<<n, 256, 0, s>>>(dx.p, Ny, Nx, apply, dt.p, db.p, dbo.p);
    CK(cudaGetLastError());
    if (apply) CK(dx.down(x, s));
    if (thresh) CK(dt.down(thresh, s));
    if (bins) CK(db.down(bins, s));
    if (bounds && apply) CK(dbo.down(bounds, s));
    CK(cudaStreamSynchronize(s));
    return SCL_OK;
}

int scl_resample_f32(const float* x, float* out, int n, int Ny, int Nx, const double* ky, const int32_t* y0, int Ly,
                     const double* kx, const int32_t* x0, int Lx, void* stream) {
    if (n == 0) return SCL_OK;
    if (!x || !out || !ky || !kx || !y0 || !x0 || n < 0 || Ny < 1 || Nx < 1 || Ly < 1 || Lx < 1)
        return fail(SCL_E_ARG, "bad argument%s%s");
    if (Ly > SCL_RESAMPLE_MAX_TAPS || Lx > SCL_RESAMPLE_MAX_TAPS) return fail(SCL_E_LIMIT, "at most 16 taps per axis%s%s");
    int rc = ensure_globals();
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t N = (size_t)Ny * Nx;
    DevBuf<float> dx, dout;
    DevBuf<double> dky, dkx;
    DevBuf<int> dy0, dx0;
    CK(dx.alloc(n * N)); CK(dout.alloc(n * N)); CK(dky.alloc((size_t)n * Ly)); CK(dkx.alloc((size_t)n * Lx));
    CK(dy0.alloc(n)); CK(dx0.alloc(n));
    CK(dx.up(x, s)); CK(dky.up(ky, s)); CK(dkx.up(kx, s)); CK(dy0.up(y0, s)); CK(dx0.up(x0, s));
    scl_k_resample<<<dim3((unsigned)((N + 255) / 256), n), 256, 0, s>>>(dx.p, dout.p, Ny, Nx, dky.p, dy0.p, Ly, dkx.p, dx0.p, Lx);
    CK(cudaGetLastError());
    CK(dout.down(out, s));
    CK(cudaStreamSynchronize(s));
    return SCL_OK;
}

// ---- the array-level operators on caller-owned DEVICE arrays: stream-ordered, no host<->device copies, no
// synchronisation (SURVEY 8(b) "Ownership").  Centres live on the device and cannot be validated here: the kernels
// leave an image whose centre is outside the frame untouched.
extern "C" {

int scl_prox_weighted_monotonic_f32_dev(float* x_dev, int n, int Ny, int Nx, const int32_t* centers_dev, float thresh, void* stream) {
    if (n == 0) return SCL_OK;
    if (!x_dev || !centers_dev || n < 0) return fail(SCL_E_ARG, "bad argument%s%s");
    return launch_monotonic<float, false>(x_dev, n, Ny, Nx, centers_dev, thresh, (cudaStream_t)stream);
}
int scl_prox_weighted_monotonic_f64_dev(double* x_dev, int n, int Ny, int Nx, const int32_t* centers_dev, double thresh, void* stream) {
    if (n == 0) return SCL_OK;
    if (!x_dev || !centers_dev || n < 0) return fail(SCL_E_ARG, "bad argument%s%s");
    return launch_monotonic<double, false>(x_dev, n, Ny, Nx, centers_dev, thresh, (cudaStream_t)stream);
}

int scl_measure_f32_dev(const float* x_dev, int n, int Ny, int Nx, int32_t* centers_dev, double* shifts_dev,
                        const float* psf_dev, int ph, int pw, int do_max, int do_centroid, void* stream) {
    if (n == 0) return SCL_OK;
    if (!x_dev || !centers_dev || !shifts_dev || n < 0 || (do_centroid && (!psf_dev || ph < 1 || pw < 1)))
        return fail(SCL_E_ARG, "bad argument%s%s");
    int rc = check_shape(Ny, Nx, true);
    if (rc) return rc;
    if ((rc = ensure_globals())) return rc;
    const size_t N = (size_t)Ny * Nx;
    const int in_place = N * sizeof(float) > 200 * 1024;
    scl_k_measure<<<n, SCL_THREADS, in_place ? 0 : N * sizeof(float), (cudaStream_t)stream>>>(
        x_dev, centers_dev, shifts_dev, Ny, Nx, psf_dev, ph, pw, do_max, do_centroid, in_place);
    CK(cudaGetLastError());
    return SCL_OK;
}

int scl_convolve_f32_dev(const float* model_dev, float* out_dev, int n, int C, int Ny, int Nx, const float* spec_dev,
                         int n_ker, const int32_t* ker_idx_dev, int adjoint, void* stream) {
    if (n == 0) return SCL_OK;
    if (!model_dev || !out_dev || !ker_idx_dev || n < 0 || C < 1 || (n_ker > 0 && !spec_dev))
        return fail(SCL_E_ARG, "bad argument%s%s");
    if (C > SCL_MAX_BANDS) return fail(SCL_E_LIMIT, "1 <= C <= 8 bands supported%s%s");
    if (Ny > SCL_FFT || Nx > SCL_FFT)
        return fail(SCL_E_LIMIT, "frame larger than the 128x128 convolution tile: use scl_convolve_f32 with ker_half%s%s");
    int rc = ensure_globals();
    if (rc) return rc;
    scl_k_convolve<<<dim3(C, n), SCL_THREADS, kFwdSmem, (cudaStream_t)stream>>>(
        model_dev, reinterpret_cast<const float2*>(spec_dev), ker_idx_dev, out_dev, C, Ny, Nx, adjoint);
    CK(cudaGetLastError());
    return SCL_OK;
}

}  // extern "C"

#include "api_batch.inc"
#include "api_noalloc.inc"
#include "api_pybind.inc"
#include "api_f64.inc"
