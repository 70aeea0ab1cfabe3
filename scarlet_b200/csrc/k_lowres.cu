// K1-LR: the image planes of a low-resolution observation.
//
// Replaces (reference file:line):
//   LowResObservation._render      observation.py:524-558   resample + convolve the model into the LR frame
//   LowResObservation.get_loss     observation.py:580-599   0.5 sum (W (render - images[window]))^2
//   autograd.grad of that loss     blend.py:44-45,105-119   the transpose of `_render`
//
// `_render` is linear in the model: per band a fixed map L_c from the Ny*Nx model pixels to the n_lr pixels of the
// LR window (sinc shift along one axis, contraction with the shifted difference kernel along the other, final flip;
// all folded into L_c once per geometry by LowResObservation.match on the host).  One CTA per (blend, LR plane):
//   M   = sum_k A[k, band] S_k                    (ComponentTree.get_model, component.py:321-347)  -> shared memory
//   R   = L_c M ;  d = W (R - I) ; logL += d^2/2 ; r = W d                                          -> shared memory
//   G_p = L_c^T r                                                                                    -> HBM plane
// The operator is read from L2/HBM twice per CTA (row-wise for R, column-wise for G): a first correct path; the
// blocked form (tile of blends per CTA) is the next step for batches of this kind.
#include "common.cuh"

// grid: (Plr * n_listed_blends); block: SCL_THREADS; dynamic smem: (N + n_lr) floats + 32 doubles
extern "C" __global__ void __launch_bounds__(SCL_THREADS)
scl_k_lowres(const float* __restrict__ A,          // [Ktot, Cm]
             const float* __restrict__ S,          // [Ktot, N]
             const int* __restrict__ comp_off,     // [B+1]
             const float* __restrict__ lr_images,  // [B, Plr, n_lr]   LR window of every band
             const float* __restrict__ lr_weights, // [B, Plr, n_lr] or nullptr (W = 1)
             const float* __restrict__ lr_op,      // [n_op, Plr, n_lr, N]
             const int* __restrict__ lr_op_idx,    // [B]
             const int* __restrict__ active,       // [B]
             const int* __restrict__ blend_list,   // [gridDim.x / Plr]
             float* __restrict__ G,                // [B, PG, N]; this kernel fills planes P0 .. P0 + Plr - 1
             double* __restrict__ loss_out,        // [B, loss_stride]; this kernel fills entries loss_off .. + Plr - 1
             int Cm, SclPlanes pl, int P0, int Plr, int PG, int N, int n_lr, int loss_stride, int loss_off)
{
    extern __shared__ __align__(16) unsigned char smem_lr[];
    double* red = reinterpret_cast<double*>(smem_lr);
    float* M = reinterpret_cast<float*>(red + 32);
    float* r = M + N;

    const int c = blockIdx.x % Plr, b = blend_list[blockIdx.x / Plr];
    if (!active[b]) return;
    const int k0 = comp_off[b], K = comp_off[b + 1] - k0;
    int cm = 0;
#pragma unroll
    for (int p = 0; p < SCL_MAX_BANDS; p++) cm = (p == P0 + c) ? pl.band[p] : cm;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
    const float* L = lr_op + ((size_t)lr_op_idx[b] * Plr + c) * (size_t)n_lr * N;
    const float* img = lr_images + ((size_t)b * Plr + c) * n_lr;
    const float* wgt = lr_weights ? lr_weights + ((size_t)b * Plr + c) * n_lr : nullptr;
    float* gout = G + ((size_t)b * PG + P0 + c) * N;

    for (int p = threadIdx.x; p < N; p += blockDim.x) {          // any number of components
        float m = 0.f;
        for (int k = 0; k < K; k++) m += A[(size_t)(k0 + k) * Cm + cm] * S[(size_t)(k0 + k) * N + p];
        M[p] = m;
    }
    __syncthreads();

    double loss = 0.0;
    for (int q = warp; q < n_lr; q += nwarp) {
        const float* Lq = L + (size_t)q * N;
        float acc = 0.f;
        for (int p = lane; p < N; p += 32) acc += Lq[p] * M[p];
        acc = warp_sum<float>(acc);
        if (lane == 0) {
            const float w = wgt ? wgt[q] : 1.f;
            const float d = w * (acc - img[q]);
            loss += (double)d * (double)d;
            r[q] = w * d;
        }
    }
    __syncthreads();
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        float g = 0.f;
        for (int q = 0; q < n_lr; q++) g += L[(size_t)q * N + p] * r[q];
        gout[p] = g;
    }
    const double tot = block_sum<double>(loss, red);
    if (threadIdx.x == 0) loss_out[(size_t)b * loss_stride + loss_off + c] = 0.5 * tot;
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
