// K2: analytic SED / morphology gradients fused with the Lipschitz constants
// and the gradient step.  One CTA per blend.
//
// Replaces (reference file:line):
//   chain rule through Component.get_model      component.py:166-167 / blend.py:105-119
//       dA[k,c] = sum_p G[c,p] S[k,p]     dS[k,p] = sum_c A[k,c] G[c,p]
//   Blend._set_lipschitz                         blend.py:186-223
//       exact:  L_sed = n_obs lambda_max(S S^T),  L_morph = n_obs lambda_max(A^T A)
//       approx: L_sed = n_obs sum|S|^2, L_morph = n_obs sum|A|^2, x2 if the loss rose
//   the step                                     blend.py:87-96, component.py:198-210
//   Blend.mse bookkeeping                        blend.py:138 (loss BEFORE the step)
#include "common.cuh"

namespace {

// largest eigenvalue of a symmetric n x n matrix (row-major, double) by cyclic
// Jacobi; destroys `a`.  n is tiny (K components or C bands).
__device__ double jacobi_lambda_max(double* a, int n) {
    for (int sweep = 0; sweep < 30; sweep++) {
        double off = 0.0, diag = 0.0;
        for (int i = 0; i < n; i++) {
            diag += a[i * n + i] * a[i * n + i];
            for (int j = i + 1; j < n; j++) off += a[i * n + j] * a[i * n + j];
        }
        if (off <= 1e-30 * diag || off == 0.0) break;
        for (int p = 0; p < n - 1; p++)
            for (int q = p + 1; q < n; q++) {
                double apq = a[p * n + q];
                if (apq == 0.0) continue;
                double theta = (a[q * n + q] - a[p * n + p]) / (2.0 * apq);
                double tt = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                double cs = 1.0 / sqrt(tt * tt + 1.0), sn = tt * cs;
                for (int r = 0; r < n; r++) {
                    double arp = a[r * n + p], arq = a[r * n + q];
                    a[r * n + p] = cs * arp - sn * arq;
                    a[r * n + q] = sn * arp + cs * arq;
                }
                for (int r = 0; r < n; r++) {
                    double apr = a[p * n + r], aqr = a[q * n + r];
                    a[p * n + r] = cs * apr - sn * aqr;
                    a[q * n + r] = sn * apr + cs * aqr;
                }
            }
    }
    double m = a[0];
    for (int i = 1; i < n; i++) m = fmax(m, a[i * n + i]);
    return m;
}

}  // namespace

// KMAX: compile-time bound on components per blend (register accumulators).
// grid: (n_blends); block: SCL_THREADS.
template <int KMAX>
__global__ void __launch_bounds__(SCL_THREADS, 1)
scl_k_grad_step(const float* __restrict__ A,          // [Ktot, C]  current SEDs
                const float* __restrict__ S,          // [Ktot, N]  current morphologies
                float* __restrict__ A_new,            // [Ktot, C]  stepped SEDs
                float* __restrict__ S_new,            // [Ktot, N]  stepped morphologies
                const int* __restrict__ comp_off,     // [B+1]
                const unsigned* __restrict__ cflags,  // [Ktot]
                const float* __restrict__ G,          // [B, C, N]
                const double* __restrict__ loss_part, // [B, C]
                const int* __restrict__ active,       // [B]
                const int* __restrict__ it0,          // [B] iterations done before this fit call
                const double* __restrict__ prev_loss, // [B] last loss of an earlier fit call (if it0 > 0)
                double* __restrict__ logL,            // [B, logL_cap] loss of step s of THIS fit call
                float* __restrict__ lips,             // [B, 2] (L_sed, L_morph) for inspection
                int C, int N, int step, int logL_cap, int n_obs, int approximate_L)
{
    const int b = blockIdx.x;
    if (!active[b]) return;
    const int k0 = comp_off[b], K = comp_off[b + 1] - k0;
    constexpr int NG = KMAX * (KMAX + 1) / 2;
    __shared__ float sA[KMAX * SCL_MAX_BANDS];
    __shared__ double sred[8][KMAX * SCL_MAX_BANDS + NG];   // per-warp partials
    __shared__ double sdA[KMAX * SCL_MAX_BANDS];
    __shared__ double sgram[KMAX * KMAX];
    __shared__ double sata[SCL_MAX_BANDS * SCL_MAX_BANDS];
    __shared__ float sL[2];

    for (int i = threadIdx.x; i < K * C; i += blockDim.x) sA[i] = A[(size_t)k0 * C + i];
    __syncthreads();

    // ---- phase 1: dA and Gram(S) in one pass over the pixels
    float accA[KMAX][SCL_MAX_BANDS];
    float accG[NG];
#pragma unroll
    for (int k = 0; k < KMAX; k++)
#pragma unroll
        for (int c = 0; c < SCL_MAX_BANDS; c++) accA[k][c] = 0.f;
#pragma unroll
    for (int i = 0; i < NG; i++) accG[i] = 0.f;

    const float* Gb = G + (size_t)b * C * N;
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        float s[KMAX], g[SCL_MAX_BANDS];
#pragma unroll
        for (int k = 0; k < KMAX; k++) s[k] = (k < K) ? S[(size_t)(k0 + k) * N + p] : 0.f;
#pragma unroll
        for (int c = 0; c < SCL_MAX_BANDS; c++) g[c] = (c < C) ? Gb[(size_t)c * N + p] : 0.f;
#pragma unroll
        for (int k = 0; k < KMAX; k++)
#pragma unroll
            for (int c = 0; c < SCL_MAX_BANDS; c++) accA[k][c] += g[c] * s[k];
        int gi = 0;
#pragma unroll
        for (int k = 0; k < KMAX; k++)
#pragma unroll
            for (int l = k; l < KMAX; l++) accG[gi++] += s[k] * s[l];
    }
    {
        const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
#pragma unroll
        for (int k = 0; k < KMAX; k++)
#pragma unroll
            for (int c = 0; c < SCL_MAX_BANDS; c++) {
                double v = warp_sum<double>((double)accA[k][c]);
                if (l == 0) sred[w][k * SCL_MAX_BANDS + c] = v;
            }
#pragma unroll
        for (int i = 0; i < NG; i++) {
            double v = warp_sum<double>((double)accG[i]);
            if (l == 0) sred[w][KMAX * SCL_MAX_BANDS + i] = v;
        }
    }
    __syncthreads();
    const int nw = blockDim.x >> 5;
    for (int i = threadIdx.x; i < KMAX * SCL_MAX_BANDS + NG; i += blockDim.x) {
        double v = 0.0;
        for (int w = 0; w < nw; w++) v += sred[w][i];
        if (i < KMAX * SCL_MAX_BANDS) {
            sdA[i] = v;
        } else {
            // unpack the upper triangle
            int gi = i - KMAX * SCL_MAX_BANDS, k = 0, rem = gi;
            while (rem >= KMAX - k) { rem -= KMAX - k; k++; }
            int l = k + rem;
            sgram[k * KMAX + l] = v;
            sgram[l * KMAX + k] = v;
        }
    }
    __syncthreads();

    // ---- phase 2: loss bookkeeping + Lipschitz constants (one thread; tiny)
    if (threadIdx.x == 0) {
        double total = 0.0;
        for (int c = 0; c < C; c++) total += loss_part[b * C + c];
        const int it = it0[b] + step;              // index of this loss in Blend.mse
        if (step < logL_cap) logL[(size_t)b * logL_cap + step] = total;
        double LA, LS;
        if (approximate_L) {
            LA = 0.0; LS = 0.0;
            for (int k = 0; k < K; k++) LA += sgram[k * KMAX + k];
            for (int i = 0; i < K * C; i++) LS += (double)sA[i] * (double)sA[i];
            if (it >= 1) {
                // blend.py:199: len(mse) > 1 (the loss just appended counts) and mse[-1] > mse[-2]
                double prev = (step >= 1) ? logL[(size_t)b * logL_cap + step - 1] : prev_loss[b];
                if (total > prev) { LA *= 2; LS *= 2; }
            }
        } else {
            double m[KMAX * KMAX];
            for (int k = 0; k < K; k++)
                for (int l = 0; l < K; l++) m[k * K + l] = sgram[k * KMAX + l];
            LA = jacobi_lambda_max(m, K);
            for (int c = 0; c < C; c++)
                for (int d = 0; d < C; d++) {
                    double v = 0.0;
                    for (int k = 0; k < K; k++) v += (double)sA[k * C + c] * (double)sA[k * C + d];
                    sata[c * C + d] = v;
                }
            LS = jacobi_lambda_max(sata, C);
        }
        sL[0] = (float)(LA * n_obs);               // L_sed   (blend.py:219-222: the morph Gram bounds the SED step)
        sL[1] = (float)(LS * n_obs);               // L_morph
        lips[b * 2] = sL[0];
        lips[b * 2 + 1] = sL[1];
    }
    __syncthreads();
    const float step_sed = 1.0f / sL[0], step_morph = 1.0f / sL[1];

    // ---- phase 3: the step (out of place, blend.py:91-96)
    for (int i = threadIdx.x; i < K * C; i += blockDim.x) {
        int k = i / C;
        float a = sA[i];
        if (!(cflags[k0 + k] & SCL_F_FIX_SED)) a = a - step_sed * (float)sdA[k * SCL_MAX_BANDS + (i - k * C)];
        A_new[(size_t)k0 * C + i] = a;
    }
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        float g[SCL_MAX_BANDS];
#pragma unroll
        for (int c = 0; c < SCL_MAX_BANDS; c++) g[c] = (c < C) ? Gb[(size_t)c * N + p] : 0.f;
        for (int k = 0; k < K; k++) {
            float s = S[(size_t)(k0 + k) * N + p];
            if (!(cflags[k0 + k] & SCL_F_FIX_MORPH)) {
                float d = 0.f;
                for (int c = 0; c < C; c++) d += sA[k * C + c] * g[c];
                s = s - step_morph * d;
            }
            S_new[(size_t)(k0 + k) * N + p] = s;
        }
    }
}

template __global__ void scl_k_grad_step<4>(const float*, const float*, float*, float*, const int*, const unsigned*,
                                            const float*, const double*, const int*, const int*, const double*, double*, float*,
                                            int, int, int, int, int, int);
template __global__ void scl_k_grad_step<8>(const float*, const float*, float*, float*, const int*, const unsigned*,
                                            const float*, const double*, const int*, const int*, const double*, double*, float*,
                                            int, int, int, int, int, int);
