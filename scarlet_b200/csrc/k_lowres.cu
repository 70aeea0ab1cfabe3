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
