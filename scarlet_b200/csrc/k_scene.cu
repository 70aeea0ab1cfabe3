// Scene mode: the same iteration for frames larger than one 128x128 convolution tile and blends with more than 8
// components (BASELINE config 4: one 6-band 512x512 scene, ~340 components).  K1 is split into three kernels that
// hand the model and the residual through HBM; K3 / K2b run unchanged with their working set in an HBM scratch
// (k_prox.cu, SclProxArgs::xscratch).
//
// Replaces (reference file:line), per iteration:
//   ComponentTree.get_model                component.py:321-347   scl_k_scene_render
//   Observation.render + get_loss          observation.py:198-239 scl_k_scene_conv<residual>   (overlap-save tiles)
//   autograd adjoint of the convolution    blend.py:105-119       scl_k_scene_conv<adjoint>
//   Blend._set_lipschitz (approximate_L)   blend.py:189-202       scl_k_scene_norms + scl_k_scene_lipschitz
//
// Overlap-save: an output tile of T = 128 - 2h pixels per axis (h = half-width of the truncated difference kernel)
// needs the input on the tile grown by h on every side, which is exactly one 128-point period: the circular
// convolution of that window with the kernel (centre tap at index 0) is the linear convolution on the tile's
// interior.  Pixels outside the frame are zero, as in the reference's zero padding (fft.py:38-65).
#include "common.cuh"

// M[b][c][p] = sum_k A[k][c] S[k][p], k over the components of blend b in order.  grid: (ceil(N/256), B)
extern "C" __global__ void __launch_bounds__(256)
scl_k_scene_render(const float* __restrict__ A, const float* __restrict__ S, const int* __restrict__ comp_off,
                   const int* __restrict__ active, float* __restrict__ M, int C, int N)
{
    const int b = blockIdx.y;
    if (!active[b]) return;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const int k0 = comp_off[b], k1 = comp_off[b + 1];
    float m[SCL_MAX_BANDS];
#pragma unroll
    for (int c = 0; c < SCL_MAX_BANDS; c++) m[c] = 0.f;
    if (p < N) {
        for (int k = k0; k < k1; k++) {
            const float s = S[(size_t)k * N + p];
            const float* a = A + (size_t)k * C;
#pragma unroll
            for (int c = 0; c < SCL_MAX_BANDS; c++)
                if (c < C) m[c] += a[c] * s;
        }
#pragma unroll
        for (int c = 0; c < SCL_MAX_BANDS; c++)
            if (c < C) M[((size_t)b * C + c) * N + p] = m[c];
    }
}

// One overlap-save tile of one plane.  grid: (tiles_y * tiles_x, P, B); block SCL_THREADS; smem kFwdSmem.
//   RESIDUAL: src = model M [B][C][N] (band pl.band[plane]);  dst = r = W^2 (src (*) d - I)  [B][P][N];
//             loss_tile[b][plane][tile] = 0.5 sum (W (R - I))^2 over the tile
//   else    : src = r [B][P][N];  dst = G = r (star) d (correlation, conj spectrum)  [B][P][N]
template <bool RESIDUAL>
__global__ void __launch_bounds__(SCL_THREADS)
scl_k_scene_conv(const float* __restrict__ src, const float2* __restrict__ Dhat, const int* __restrict__ ker_idx,
                 const float* __restrict__ images, const float* __restrict__ weights, const int* __restrict__ active,
                 float* __restrict__ dst, double* __restrict__ loss_tile,
                 int Cm, SclPlanes pl, int Ny, int Nx, int h, int tiles_x)
{
    extern __shared__ __align__(16) unsigned char smem_sc[];
    float2* buf = reinterpret_cast<float2*>(smem_sc);
    float2* tw = buf + F * RS;
    double* red = reinterpret_cast<double*>(tw + 256 + 2 * RS);
    const int tile = blockIdx.x, plane = blockIdx.y, b = blockIdx.z, P = pl.P;
    if (!active[b]) return;
    const int ntile = gridDim.x;
    const int T = F - 2 * h;
    const int oy = (tile / tiles_x) * T - h, ox = (tile % tiles_x) * T - h;     // frame coordinates of window (0, 0)
    const int N = Ny * Nx;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, t = lane & 15, half = lane >> 4;
    const int nwarp = blockDim.x >> 5;
    const float inv = 1.0f / (float)(F * F);
    const int kid = ker_idx[b * P + plane];
    const float* in = RESIDUAL ? src + ((size_t)b * Cm + pl.band[plane]) * N : src + ((size_t)b * P + plane) * N;
    float* out = dst + ((size_t)b * P + plane) * N;
    const float* img = RESIDUAL ? images + ((size_t)b * P + plane) * N : nullptr;
    const float* wgt = (RESIDUAL && weights) ? weights + ((size_t)b * P + plane) * N : nullptr;
    const float2* D = Dhat + (size_t)(kid < 0 ? 0 : kid) * F * RS;
    double loss = 0.0;

    if (kid >= 0) {
        fill_twiddles128(tw);
        __syncthreads();
        for (int j0 = warp * 2; j0 < F / 2; j0 += nwarp * 2) {
            const int j = j0 + half;
            float2* line = buf + j * LS;
            const int y0 = oy + 2 * j, y1 = y0 + 1;
            for (int x = t; x < F; x += 16) {
                const int gx = ox + x;
                const bool xin = gx >= 0 && gx < Nx;
                line[x] = make_float2((xin && y0 >= 0 && y0 < Ny) ? in[(size_t)y0 * Nx + gx] : 0.f,
                                      (xin && y1 >= 0 && y1 < Ny) ? in[(size_t)y1 * Nx + gx] : 0.f);
            }
            __syncwarp();
            fft_line16<128>(line, 1, t, tw, F, true);
            untangle_line(line, t, true);
        }
        __syncthreads();
        if (RESIDUAL) column_phase<false>(buf, tw, D, F); else column_phase<true>(buf, tw, D, F);
        __syncthreads();
    }
    for (int j0 = warp * 2; j0 < F / 2; j0 += nwarp * 2) {
        const int j = j0 + half;
        float2* line = buf + j * LS;
        if (kid >= 0) {
            retangle_line(line, t, true);
            fft_line16<128>(line, 1, t, tw, F, true);
        }
        for (int x = h + t; x < F - h; x += 16) {
            const int gx = ox + x;
            if (gx >= Nx) continue;
#pragma unroll
            for (int r = 0; r < 2; r++) {
                const int wy = 2 * j + r, gy = oy + wy;
                if (wy < h || wy >= F - h || gy >= Ny) continue;
                const size_t p = (size_t)gy * Nx + gx;
                float v;
                if (kid >= 0) { const float2 z = line[x]; v = (r == 0 ? z.y : z.x) * inv; }
                else v = in[p];                                   // identical PSFs: no convolution
                if (RESIDUAL) {
                    const float w = wgt ? wgt[p] : 1.f;
                    const float d = w * (v - img[p]);
                    loss += (double)d * (double)d;
                    v = w * d;
                }
                out[p] = v;
            }
        }
    }
    if (RESIDUAL) {
        const double tot = block_sum<double>(loss, red);
        if (threadIdx.x == 0) loss_tile[((size_t)b * P + plane) * ntile + tile] = 0.5 * tot;
    }
}
template __global__ void scl_k_scene_conv<true>(const float*, const float2*, const int*, const float*, const float*,
                                                const int*, float*, double*, int, SclPlanes, int, int, int, int);
template __global__ void scl_k_scene_conv<false>(const float*, const float2*, const int*, const float*, const float*,
                                                 const int*, float*, double*, int, SclPlanes, int, int, int, int);

// Per-component squared norms for approximate_L (blend.py:189-202): snorm[k] = sum_p S[k][p]^2.  grid: Ktot
extern "C" __global__ void __launch_bounds__(256)
scl_k_scene_norms(const float* __restrict__ S, double* __restrict__ snorm, int N)
{
    __shared__ double red[32];
    const float* s = S + (size_t)blockIdx.x * N;
    double acc = 0.0;
    for (int p = threadIdx.x; p < N; p += blockDim.x) { const double v = s[p]; acc += v * v; }
    acc = block_sum<double>(acc, red);
    if (threadIdx.x == 0) snorm[blockIdx.x] = acc;
}

// Loss of the iteration + approximate Lipschitz constants, one thread per blend (any number of components):
//   logL = sum of the tile losses in (plane, tile) order;  L_sed = n_obs sum_k |S_k|^2, L_morph = n_obs sum_k |A_k|^2,
//   both doubled when the loss rose (blend.py:189-202,219-220).
extern "C" __global__ void scl_k_scene_lipschitz(const float* __restrict__ A, const int* __restrict__ comp_off,
                                                 const double* __restrict__ snorm, const double* __restrict__ loss_tile,
                                                 const int* __restrict__ active, const int* __restrict__ it0,
                                                 const double* __restrict__ prev_loss, double* __restrict__ logL,
                                                 float* __restrict__ lips, int B, int C, int n_loss, int step,
                                                 int logL_cap, int n_obs)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B || !active[b]) return;
    double total = 0.0;
    for (int i = 0; i < n_loss; i++) total += loss_tile[(size_t)b * n_loss + i];
    double LA = 0.0, LS = 0.0;
    for (int k = comp_off[b]; k < comp_off[b + 1]; k++) {
        LA += snorm[k];
        for (int c = 0; c < C; c++) { const double a = A[(size_t)k * C + c]; LS += a * a; }
    }
    const int it = it0[b] + step;
    if (it >= 1) {
        const double prev = (step >= 1) ? logL[(size_t)b * logL_cap + step - 1] : prev_loss[b];
        if (total > prev) { LA *= 2; LS *= 2; }
    }
    if (step < logL_cap) logL[(size_t)b * logL_cap + step] = total;
    lips[2 * b] = (float)(LA * n_obs);
    lips[2 * b + 1] = (float)(LS * n_obs);
}
