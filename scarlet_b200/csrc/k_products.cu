// Post-fit products of the components (the step after the loop, what a caller such as LSST
// meas_deblender reads out of a fitted blend).
//
// Replaces (reference file:line):
//   bbox.trim                      bbox.py:170-191        bounds of the pixels above min_value
//   bbox.flux_at_edge              bbox.py:194-210        any border pixel above min_value
//   Component.get_flux             component.py:172-175   morph.sum() * sed
//   BlendFlag.EDGE_PIXELS / NO_VALID_PIXELS   component.py:13-36   (defined by the reference, never set by it)
//   Observation.render(component.get_model())  observation.py:198-220, component.py:148-170   per-component model
//                                                          in the observed frame (expand here, convolve with K1's
//                                                          standalone kernels)
#include "common.cuh"

struct SclTrim {
    double sum;
    int y0, y1, x0, x1;      // bottom, top, left, right of the pixels > min_value (y0 > y1: none)
    float edge;              // max over the four border lines
};

// Block-wide scan of one image; the result is valid in thread 0.  red: >= 7 * 32 doubles of shared memory.
__device__ __forceinline__ SclTrim scl_trim_block(const float* __restrict__ x, int Ny, int Nx, float min_value, double* red)
{
    SclTrim t;
    t.sum = 0.0; t.y0 = Ny; t.y1 = -1; t.x0 = Nx; t.x1 = -1; t.edge = -INFINITY;
    const int N = Ny * Nx;
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        const float v = x[p];
        const int y = p / Nx, xx = p - y * Nx;
        t.sum += (double)v;
        if (v > min_value) {
            t.y0 = min(t.y0, y); t.y1 = max(t.y1, y);
            t.x0 = min(t.x0, xx); t.x1 = max(t.x1, xx);
        }
        if (y == 0 || y == Ny - 1 || xx == 0 || xx == Nx - 1) t.edge = fmaxf(t.edge, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        t.sum += __shfl_xor_sync(0xffffffffu, t.sum, o);
        t.y0 = min(t.y0, __shfl_xor_sync(0xffffffffu, t.y0, o));
        t.y1 = max(t.y1, __shfl_xor_sync(0xffffffffu, t.y1, o));
        t.x0 = min(t.x0, __shfl_xor_sync(0xffffffffu, t.x0, o));
        t.x1 = max(t.x1, __shfl_xor_sync(0xffffffffu, t.x1, o));
        t.edge = fmaxf(t.edge, __shfl_xor_sync(0xffffffffu, t.edge, o));
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    int* ired = reinterpret_cast<int*>(red + 32);
    float* fred = reinterpret_cast<float*>(ired + 4 * 32);
    if (lane == 0) {
        red[warp] = t.sum;
        ired[warp] = t.y0; ired[32 + warp] = t.y1; ired[64 + warp] = t.x0; ired[96 + warp] = t.x1;
        fred[warp] = t.edge;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < nw; w++) {              // fixed order: the same sum on every run
            t.sum += red[w];
            t.y0 = min(t.y0, ired[w]); t.y1 = max(t.y1, ired[32 + w]);
            t.x0 = min(t.x0, ired[64 + w]); t.x1 = max(t.x1, ired[96 + w]);
            t.edge = fmaxf(t.edge, fred[w]);
        }
    }
    return t;
}

// Array level: x [n][Ny][Nx] -> bounds [n][4] (bottom, top, left, right; -1 when no pixel is above min_value),
// edge [n] (0 / 1), sum [n].  grid: n, block 256.
extern "C" __global__ void __launch_bounds__(256)
scl_k_trim(const float* __restrict__ x, int Ny, int Nx, float min_value, int* __restrict__ bounds, int* __restrict__ edge,
           double* __restrict__ sum)
{
    __shared__ double red[7 * 32];
    const SclTrim t = scl_trim_block(x + (size_t)blockIdx.x * Ny * Nx, Ny, Nx, min_value, red);
    if (threadIdx.x == 0) {
        const bool any = t.y1 >= t.y0;
        int* b = bounds + 4 * (size_t)blockIdx.x;
        b[0] = any ? t.y0 : -1; b[1] = any ? t.y1 : -1; b[2] = any ? t.x0 : -1; b[3] = any ? t.x1 : -1;
        edge[blockIdx.x] = t.edge > min_value;
        sum[blockIdx.x] = t.sum;
    }
}

// Batch level, one CTA per component of the fitted state: flux [Ktot][C] = float32(sum morph) * sed, bounds as
// above, flags = convergence bits | 4 (EDGE_PIXELS) | 8 (NO_VALID_PIXELS).
extern "C" __global__ void __launch_bounds__(256)
scl_k_products(const float* __restrict__ A, const float* __restrict__ S, const int* __restrict__ conv, int C, int Ny, int Nx,
               float min_value, float* __restrict__ flux, int* __restrict__ bounds, int* __restrict__ flags)
{
    __shared__ double red[7 * 32];
    const int k = blockIdx.x;
    const SclTrim t = scl_trim_block(S + (size_t)k * Ny * Nx, Ny, Nx, min_value, red);
    if (threadIdx.x == 0) {
        const bool any = t.y1 >= t.y0;
        if (bounds) {
            int* b = bounds + 4 * (size_t)k;
            b[0] = any ? t.y0 : -1; b[1] = any ? t.y1 : -1; b[2] = any ? t.x0 : -1; b[3] = any ? t.x1 : -1;
        }
        if (flags) flags[k] = (conv[k] & 3) | (t.edge > min_value ? 4 : 0) | (any ? 0 : 8);
        if (flux) {
            const float s = (float)t.sum;
            for (int c = 0; c < C; c++) flux[(size_t)k * C + c] = s * A[(size_t)k * C + c];
        }
    }
}

// Component.get_model on the planes of the batch: M[j][p][pix] = A[k0+j][band(p)] * S[k0+j][pix], and the kernel
// index of every (component, plane) from the blend that owns it.  grid: (ceil(N/256), P, nk)
extern "C" __global__ void __launch_bounds__(256)
scl_k_expand_components(const float* __restrict__ A, const float* __restrict__ S, const int* __restrict__ comp_blend,
                        const int* __restrict__ ker_idx, float* __restrict__ M, int* __restrict__ kidx,
                        int k0, int Cm, SclPlanes pl, int N)
{
    const int j = blockIdx.z, p = blockIdx.y, k = k0 + j, P = pl.P;
    int cm = 0;
#pragma unroll
    for (int q = 0; q < SCL_MAX_BANDS; q++) cm = (q == p) ? pl.band[q] : cm;
    const float a = A[(size_t)k * Cm + cm];
    const int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix < N) M[((size_t)j * P + p) * N + pix] = a * S[(size_t)k * N + pix];
    if (blockIdx.x == 0 && threadIdx.x == 0) kidx[j * P + p] = ker_idx[comp_blend[k] * P + p];
}
