// K1-LR: the image planes of a low-resolution observation.
//
// Replaces (reference file:line):
//   LowResObservation._render      observation.py:524-558   resample + convolve the model into the LR frame
//   LowResObservation.get_loss     observation.py:580-599   0.5 sum (W (render - images[window]))^2
//   autograd.grad of that loss     blend.py:44-45,105-119   the transpose of `_render`
//
// `_render` is linear in the model: per band a fixed map L_c from the Ny*Nx model pixels to the n_lr pixels of the
// LR window (sinc shift along one axis, contraction with the shifted difference kernel along the other, final flip;
// all folded into L_c once per geometry by LowResObservation.match on the host).  Per (blend, LR plane):
//   M   = sum_k A[k, band] S_k                    (ComponentTree.get_model, component.py:321-347)
//   R   = L_c M ;  d = W (R - I) ; logL += d^2/2 ; r = W d
//   G_p = L_c^T r                                                                  -> plane P0 + c of G
// The operator is streamed twice per plane and iteration (row-wise for R, column-wise for G); blocking it over
// blends is the next step for large batches of this kind.
#include "common.cuh"

// Three kernels per iteration, all with enough CTAs to stream the operator at HBM speed:
//   scl_k_lowres_model    M[b][c][p] = sum_k A[k, band] S_k[p]            grid (N/256, Plr, nB)   (skipped in scene
//                                                                          mode, which already has the model in HBM)
//   scl_k_lowres_forward  one warp per low-resolution pixel q: R = L_c[q] . M, d = W (R - I), r[q] = W d, d2[q] = d^2
//                                                                          grid (n_lr/8, Plr, nB), 256 threads
//   scl_k_lowres_adjoint  G[p] = sum_q L_c[q][p] r[q] (coalesced over p, q loop unrolled); the CTA of the first
//                         pixel block also sums d2 in q order -> loss         grid (N/256, Plr, nB)
__device__ __forceinline__ int scl_lr_band(const SclPlanes& pl, int idx) {
    int cm = 0;
#pragma unroll
    for (int p = 0; p < SCL_MAX_BANDS; p++) cm = (p == idx) ? pl.band[p] : cm;
    return cm;
}

extern "C" __global__ void __launch_bounds__(256)
scl_k_lowres_model(const float* __restrict__ A, const float* __restrict__ S, const int* __restrict__ comp_off,
                   const int* __restrict__ active, const int* __restrict__ blend_list, float* __restrict__ Mlr,
                   int Cm, SclPlanes pl, int P0, int Plr, int N)
{
    const int c = blockIdx.y, b = blend_list[blockIdx.z];
    if (!active[b]) return;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= N) return;
    const int k0 = comp_off[b], K = comp_off[b + 1] - k0, cm = scl_lr_band(pl, P0 + c);
    float m = 0.f;
    for (int k = 0; k < K; k++) m += A[(size_t)(k0 + k) * Cm + cm] * S[(size_t)(k0 + k) * N + p];
    Mlr[((size_t)b * Plr + c) * N + p] = m;
}

// M: [B][m_planes][N], the plane of LR band c is m_off(c): scene mode passes the full model (m_planes = Cm, plane =
// model band), the fused mode the scl_k_lowres_model output (m_planes = Plr, plane = c)
extern "C" __global__ void __launch_bounds__(256)
scl_k_lowres_forward(const float* __restrict__ M, int m_planes, int m_is_model,
                     const float* __restrict__ lr_images, const float* __restrict__ lr_weights,
                     const float* __restrict__ lr_op, const int* __restrict__ lr_op_idx,
                     const int* __restrict__ active, const int* __restrict__ blend_list,
                     float* __restrict__ r_lr, float* __restrict__ d2_lr, SclPlanes pl, int P0, int Plr, int N, int n_lr)
{
    const int c = blockIdx.y, b = blend_list[blockIdx.z];
    if (!active[b]) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * (blockDim.x >> 5) + warp;
    if (q >= n_lr) return;
    const int mp = m_is_model ? scl_lr_band(pl, P0 + c) : c;
    const float* m = M + ((size_t)b * m_planes + mp) * N;
    const float* Lq = lr_op + (((size_t)lr_op_idx[b] * Plr + c) * n_lr + q) * (size_t)N;
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
    int p = lane;
    for (; p + 96 < N; p += 128) {
        acc0 += Lq[p] * m[p]; acc1 += Lq[p + 32] * m[p + 32];
        acc2 += Lq[p + 64] * m[p + 64]; acc3 += Lq[p + 96] * m[p + 96];
    }
    for (; p < N; p += 32) acc0 += Lq[p] * m[p];
    const float acc = warp_sum<float>((acc0 + acc1) + (acc2 + acc3));
    if (lane == 0) {
        const size_t i = ((size_t)b * Plr + c) * n_lr + q;
        const float w = lr_weights ? lr_weights[i] : 1.f;
        const float d = w * (acc - lr_images[i]);
        r_lr[i] = w * d;
        d2_lr[i] = d * d;
    }
}

extern "C" __global__ void __launch_bounds__(256)
scl_k_lowres_adjoint(const float* __restrict__ r_lr, const float* __restrict__ d2_lr,
                     const float* __restrict__ lr_op, const int* __restrict__ lr_op_idx,
                     const int* __restrict__ active, const int* __restrict__ blend_list,
                     float* __restrict__ G, double* __restrict__ loss_out,
                     int P0, int Plr, int PG, int N, int n_lr, int loss_stride, int loss_off)
{
    __shared__ double red[32];
    const int c = blockIdx.y, b = blend_list[blockIdx.z];
    if (!active[b]) return;
    const float* L = lr_op + ((size_t)lr_op_idx[b] * Plr + c) * (size_t)n_lr * N;
    const float* r = r_lr + ((size_t)b * Plr + c) * n_lr;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < N) {
        float g0 = 0.f, g1 = 0.f, g2 = 0.f, g3 = 0.f;
        int q = 0;
        for (; q + 3 < n_lr; q += 4) {
            g0 += L[(size_t)q * N + p] * r[q];
            g1 += L[(size_t)(q + 1) * N + p] * r[q + 1];
            g2 += L[(size_t)(q + 2) * N + p] * r[q + 2];
            g3 += L[(size_t)(q + 3) * N + p] * r[q + 3];
        }
        for (; q < n_lr; q++) g0 += L[(size_t)q * N + p] * r[q];
        G[((size_t)b * PG + P0 + c) * N + p] = (g0 + g1) + (g2 + g3);
    }
    if (blockIdx.x == 0) {
        const float* d2 = d2_lr + ((size_t)b * Plr + c) * n_lr;
        double loss = 0.0;
        for (int q = threadIdx.x; q < n_lr; q += blockDim.x) loss += (double)d2[q];
        loss = block_sum<double>(loss, red);
        if (threadIdx.x == 0) loss_out[(size_t)b * loss_stride + loss_off + c] = 0.5 * loss;
    }
}

// Array-level operator: out[n][c][q] = sum_p op[c][q][p] model[n][c][p]  (LowResObservation._render for display /
// get_loss outside the loop).  grid: (n_lr, C, n); block 128.
extern "C" __global__ void __launch_bounds__(128)
scl_k_lowres_render(const float* __restrict__ model, const float* __restrict__ op, float* __restrict__ out,
                    int C, int N, int n_lr)
{
    __shared__ float red[32];
    const int q = blockIdx.x, c = blockIdx.y, n = blockIdx.z;
    const float* Lq = op + ((size_t)c * n_lr + q) * N;
    const float* m = model + ((size_t)n * C + c) * N;
    float acc = 0.f;
    for (int p = threadIdx.x; p < N; p += blockDim.x) acc += Lq[p] * m[p];
    acc = block_sum<float>(acc, red);
    if (threadIdx.x == 0) out[((size_t)n * C + c) * n_lr + q] = acc;
}

// ---- blocked over blends ----------------------------------------------------------------------------------------------
// Many blends that share one geometry share the operator: R[q][b] = sum_p L[q][p] M[p][b] and
// G[p][b] = sum_q L[q][p] r[q][b] are matrix products with the blends as their third dimension, so an operator tile
// fetched once serves 64 blends (the per-blend kernels above stream the whole operator for every blend: 379 GB per
// iteration for 1000 blends of config 5).  Plain fp32 FMA on 64 x 64 x 16 tiles, 4 x 4 outputs per thread (fp32 parity:
// no reduced-precision tensor path); k ascending inside a tile, so results do not depend on scheduling.
//   ADJ = false: m = low-resolution pixel q, k = model pixel p;  epilogue: d = W (R - I), r = W d, d2 = d^2
//   ADJ = true:  m = model pixel p, k = low-resolution pixel q;  epilogue: G[b][P0 + c][p]
// grid: (ceil(M / 64), Plr, ceil(nB / 64)), 256 threads.  lr_op: ONE geometry [Plr][n_lr][N].
template <bool ADJ>
__global__ void __launch_bounds__(256)
scl_k_lowres_gemm(const float* __restrict__ lr_op, const float* __restrict__ X, int x_planes, int x_is_model,
                  const float* __restrict__ lr_images, const float* __restrict__ lr_weights,
                  const int* __restrict__ active, const int* __restrict__ blend_list, int nB,
                  float* __restrict__ r_lr, float* __restrict__ d2_lr, float* __restrict__ G,
                  SclPlanes pl, int P0, int Plr, int PG, int N, int n_lr)
{
    __shared__ __align__(16) float As[16][64 + 4];
    __shared__ __align__(16) float Bs[16][64 + 4];
    __shared__ int sb[64];
    const int c = blockIdx.y, m0 = blockIdx.x * 64, n0 = blockIdx.z * 64;
    const int Mdim = ADJ ? N : n_lr, Kdim = ADJ ? n_lr : N;
    const float* L = lr_op + (size_t)c * n_lr * N;
    if (threadIdx.x < 64) sb[threadIdx.x] = (n0 + threadIdx.x < nB) ? blend_list[n0 + threadIdx.x] : -1;
    __syncthreads();
    // the operand that carries the blends: forward the model plane of LR band c, adjoint the residual r
    const int xp = ADJ ? c : (x_is_model ? scl_lr_band(pl, P0 + c) : c);
    const size_t xlen = ADJ ? (size_t)n_lr : (size_t)N;              // length of one blend's vector along k
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
    // loaders: 64 x 16 elements per tile and operand, 4 per thread
    const int lr_ = threadIdx.x >> 2, lc = (threadIdx.x & 3) * 4;     // (row 0..63, k offset 0,4,8,12): rows contiguous in k
    const int ar = threadIdx.x >> 4, ac = (threadIdx.x & 15) * 4;     // adjoint operator tile: (k 0..15, m offset): contiguous in m
    const bool vec = (N & 3) == 0;                                    // 16-byte operator loads
    for (int k0 = 0; k0 < Kdim; k0 += 16) {
        if (!ADJ) {
            // As[k][m] = L[m0 + m][k0 + k]
            const int q = m0 + lr_, k = k0 + lc;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (q < n_lr) {
                if (vec && k + 3 < N) v = *reinterpret_cast<const float4*>(L + (size_t)q * N + k);
                else {
                    if (k < N) v.x = L[(size_t)q * N + k];
                    if (k + 1 < N) v.y = L[(size_t)q * N + k + 1];
                    if (k + 2 < N) v.z = L[(size_t)q * N + k + 2];
                    if (k + 3 < N) v.w = L[(size_t)q * N + k + 3];
                }
            }
            As[lc][lr_] = v.x; As[lc + 1][lr_] = v.y; As[lc + 2][lr_] = v.z; As[lc + 3][lr_] = v.w;
        } else {
            // As[k][m] = L[k0 + k][m0 + m]
            const int q = k0 + ar, p = m0 + ac;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (q < n_lr) {
                if (vec && p + 3 < N) v = *reinterpret_cast<const float4*>(L + (size_t)q * N + p);
                else {
                    if (p < N) v.x = L[(size_t)q * N + p];
                    if (p + 1 < N) v.y = L[(size_t)q * N + p + 1];
                    if (p + 2 < N) v.z = L[(size_t)q * N + p + 2];
                    if (p + 3 < N) v.w = L[(size_t)q * N + p + 3];
                }
            }
            *reinterpret_cast<float4*>(&As[ar][ac]) = v;
        }
        {
            // Bs[k][n] = x_b[k0 + k] of blend sb[n]
            const int b = sb[lr_];
            const float* xb = b >= 0 ? X + ((size_t)b * x_planes + xp) * xlen : nullptr;
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int k = k0 + lc + u;
                Bs[lc + u][lr_] = (xb && k < Kdim) ? xb[k] : 0.f;
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 bq = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {bq.x, bq.y, bq.z, bq.w};
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int b = sb[tx * 4 + j];
        if (b < 0 || !active[b]) continue;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int m = m0 + ty * 4 + i;
            if (m >= Mdim) continue;
            if (!ADJ) {
                const size_t idx = ((size_t)b * Plr + c) * n_lr + m;
                const float w = lr_weights ? lr_weights[idx] : 1.f;
                const float d = w * (acc[i][j] - lr_images[idx]);
                r_lr[idx] = w * d;
                d2_lr[idx] = d * d;
            } else {
                G[((size_t)b * PG + P0 + c) * N + m] = acc[i][j];
            }
        }
    }
}

// loss of every (blend, LR plane): 0.5 sum_q d2, q ascending per thread, fixed combination.  grid (Plr, nB)
extern "C" __global__ void __launch_bounds__(256)
scl_k_lowres_loss(const float* __restrict__ d2_lr, const int* __restrict__ active, const int* __restrict__ blend_list,
                  double* __restrict__ loss_out, int Plr, int n_lr, int loss_stride, int loss_off)
{
    __shared__ double red[32];
    const int c = blockIdx.x, b = blend_list[blockIdx.y];
    if (!active[b]) return;
    const float* d2 = d2_lr + ((size_t)b * Plr + c) * n_lr;
    double loss = 0.0;
    for (int q = threadIdx.x; q < n_lr; q += blockDim.x) loss += (double)d2[q];
    loss = block_sum<double>(loss, red);
    if (threadIdx.x == 0) loss_out[(size_t)b * loss_stride + loss_off + c] = 0.5 * loss;
}
