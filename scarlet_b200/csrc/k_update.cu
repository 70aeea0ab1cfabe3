// Constraints the reference exports for custom update() schedules but does not use in its own sources (SURVEY 8(f)
// row N4), as array-level device operators.
//
// Replaces (reference file:line):
//   update.sparse_l0 / sparse_l1        update.py:71-82     proxmin.operators.prox_hard / prox_soft with
//                                                           thresh * step  (proxmin 0.6 operators.py: hard:
//                                                           X[|X| < t] = 0; soft: sign(X) * max(|X| - t, 0))
//   measurement.threshold               measurement.py:96-113   cut at the last empty bin of the log10 histogram
//   update.threshold                    update.py:85-103    morph[morph < thresh] = 0, bbox.trim of the rest
#include "common.cuh"

// mode 0: hard threshold, mode 1: soft threshold; one thread per element
extern "C" __global__ void __launch_bounds__(256)
scl_k_sparse(float* __restrict__ x, size_t n, float t, int mode)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float v = x[i], av = fabsf(v);
    if (mode == 0) {
        if (av < t) x[i] = 0.f;
    } else {
        const float m = fmaxf(av - t, 0.f);
        x[i] = (v > 0.f) ? m : (v < 0.f ? -m : 0.f * m);
    }
}

// One CTA per image [Ny][Nx] (in place).  measurement.threshold: the positive pixels' log10 go into `bins` equal
// bins between their minimum and maximum (numpy.histogram: left-closed bins, the last one closed on both sides);
// 50 bins, or count/10 below 500 pixels (a single bin: no cut); the cut is 10^(left edge of the last empty bin), 0
// when every bin is occupied.  Then update.threshold: pixels below the cut are zeroed and the box of what is left
// is reported (bounds -1 when nothing is).
// all threads of the CTA; X: one image; outputs (may be nullptr) written by thread 0
__device__ void scl_threshold_body(float* X, int Ny, int Nx, int apply, double* thresh_out, int* bins_out, int* bb)
{
    __shared__ float fred[2][32];
    __shared__ int ired[4][32];
    __shared__ int hist[64];
    __shared__ float s_lo, s_hi;
    __shared__ int s_cnt;
    __shared__ double s_thresh;
    const int N = Ny * Nx, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    __syncthreads();            // the shared arrays may still be read by a previous call

    float lo = INFINITY, hi = -INFINITY;
    int cnt = 0;
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        const float v = X[p];
        if (v > 0.f) { const float l = log10f(v); lo = fminf(lo, l); hi = fmaxf(hi, l); cnt++; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    if (lane == 0) { fred[0][warp] = lo; fred[1][warp] = hi; ired[0][warp] = cnt; }
    if (threadIdx.x < 64) hist[threadIdx.x] = 0;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < nw; w++) { lo = fminf(lo, fred[0][w]); hi = fmaxf(hi, fred[1][w]); cnt += ired[0][w]; }
        s_lo = lo; s_hi = hi; s_cnt = cnt;
    }
    __syncthreads();
    cnt = s_cnt;
    int bins = 50;
    if (cnt < 500) bins = max(cnt / 10, 1);
    double thresh = 0.0;
    if (bins > 1) {
        double a = (double)s_lo, b = (double)s_hi;
        if (a == b) { a -= 0.5; b += 0.5; }                       // numpy.histogram widens an empty range
        const double step = (b - a) / bins;
        for (int p = threadIdx.x; p < N; p += blockDim.x) {
            const float v = X[p];
            if (!(v > 0.f)) continue;
            const double l = (double)log10f(v);
            int i = (int)((l - a) / (b - a) * bins);
            if (i >= bins) i = bins - 1;
            if (i < 0) i = 0;
            // settle on the edges themselves: edge(i) <= l < edge(i + 1)
            if (l < a + i * step && i > 0) i--;
            else if (i < bins - 1 && l >= a + (i + 1) * step) i++;
            atomicAdd(&hist[i], 1);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int last = -1;
            for (int i = 0; i < bins; i++) if (hist[i] == 0) last = i;
            s_thresh = last < 0 ? 0.0 : pow(10.0, a + last * step);
        }
        __syncthreads();
        thresh = s_thresh;
    }
    if (threadIdx.x == 0) { if (thresh_out) *thresh_out = thresh; if (bins_out) *bins_out = bins; }
    if (!apply) return;

    int y0 = Ny, y1 = -1, x0 = Nx, x1 = -1;
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        float v = X[p];
        if ((double)v < thresh) { v = 0.f; X[p] = 0.f; }
        if (v > 0.f) {
            const int y = p / Nx, xx = p - y * Nx;
            y0 = min(y0, y); y1 = max(y1, y); x0 = min(x0, xx); x1 = max(x1, xx);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        y0 = min(y0, __shfl_xor_sync(0xffffffffu, y0, o)); y1 = max(y1, __shfl_xor_sync(0xffffffffu, y1, o));
        x0 = min(x0, __shfl_xor_sync(0xffffffffu, x0, o)); x1 = max(x1, __shfl_xor_sync(0xffffffffu, x1, o));
    }
    __syncthreads();
    if (lane == 0) { ired[0][warp] = y0; ired[1][warp] = y1; ired[2][warp] = x0; ired[3][warp] = x1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < nw; w++) {
            y0 = min(y0, ired[0][w]); y1 = max(y1, ired[1][w]); x0 = min(x0, ired[2][w]); x1 = max(x1, ired[3][w]);
        }
        const bool any = y1 >= y0;
        if (bb) { bb[0] = any ? y0 : -1; bb[1] = any ? y1 : -1; bb[2] = any ? x0 : -1; bb[3] = any ? x1 : -1; }
    }
    __syncthreads();
}

extern "C" __global__ void __launch_bounds__(256)
scl_k_threshold(float* __restrict__ x, int Ny, int Nx, int apply, double* __restrict__ thresh_out, int* __restrict__ bins_out,
                int* __restrict__ bounds)
{
    scl_threshold_body(x + (size_t)blockIdx.x * Ny * Nx, Ny, Nx, apply, thresh_out + blockIdx.x, bins_out + blockIdx.x,
                       bounds + 4 * (size_t)blockIdx.x);
}

// interpolation.fft_resample / update.translation (interpolation.py:408-448, update.py:159-167): the image convolved
// with a separable interpolation kernel of Ly x Lx taps (2 .. 7 per axis) whose first tap sits at offset (y0, x0):
//   out[y][x] = sum_i sum_j ky[i] kx[j] img[y - y0 - i][x - x0 - j],   zero outside the image.
// The reference multiplies in Fourier space on a box padded by the kernel, which is this linear convolution.
// grid: (ceil(N / 256), n); taps float64 [n][L], accumulated in float64.
#define SCL_RESAMPLE_MAX_TAPS 16
extern "C" __global__ void __launch_bounds__(256)
scl_k_resample(const float* __restrict__ x, float* __restrict__ out, int Ny, int Nx, const double* __restrict__ ky,
               const int* __restrict__ y0, int Ly, const double* __restrict__ kx, const int* __restrict__ x0, int Lx)
{
    __shared__ double sky[SCL_RESAMPLE_MAX_TAPS], skx[SCL_RESAMPLE_MAX_TAPS];
    const int img = blockIdx.y, N = Ny * Nx;
    if (threadIdx.x < Ly) sky[threadIdx.x] = ky[(size_t)img * Ly + threadIdx.x];
    if (threadIdx.x >= 32 && threadIdx.x - 32 < Lx) skx[threadIdx.x - 32] = kx[(size_t)img * Lx + threadIdx.x - 32];
    __syncthreads();
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= N) return;
    const int y = p / Nx, xx = p - y * Nx;
    const int oy = y0[img], ox = x0[img];
    const float* src = x + (size_t)img * N;
    double acc = 0.0;
    for (int i = 0; i < Ly; i++) {
        const int sy = y - oy - i;
        if (sy < 0 || sy >= Ny) continue;
        double row = 0.0;
        for (int j = 0; j < Lx; j++) {
            const int sx = xx - ox - j;
            if (sx >= 0 && sx < Nx) row += skx[j] * (double)src[(size_t)sy * Nx + sx];
        }
        acc += sky[i] * row;
    }
    out[(size_t)img * N + p] = (float)acc;
}
