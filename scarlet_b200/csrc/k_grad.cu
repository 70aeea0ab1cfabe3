// K2: per-blend scalar work between the likelihood kernel (K1) and the constraint kernel (K3):
// loss bookkeeping and Lipschitz constants.  Two threads per blend (one per Gram matrix), everything
// in registers.
//
// Replaces (reference file:line):
//   Blend.mse bookkeeping        blend.py:138 (loss BEFORE the step)
//   Blend._set_lipschitz         blend.py:186-223
//       exact:  L_sed = n_obs lambda_max(S S^T),  L_morph = n_obs lambda_max(A^T A)
//       approx: L_sed = n_obs sum|S|^2, L_morph = n_obs sum|A|^2, x2 if the loss rose
// Inputs produced by K1: loss_part[b,c], gram[b] = S S^T (double 8x8).
// Both gradient steps (blend.py:91-96) are applied by K3 with the constants written here.
//
// lambda_max: top eigenvector by repeated squaring in f32 on the register-resident 8x8 matrix (zero padded), then
// the Rayleigh quotient of that vector in f64 against the f64 matrix: an eigenvector error e enters the eigenvalue
// as e^2 (power8_top_vector has the bound).  (The reference's own value comes from LAPACK sgeev on an f32 Gram
// matrix, i.e. is itself only f32-accurate.)
#include "common.cuh"

namespace {

constexpr int JN = 8;      // padded matrix edge == max components per blend == SCL_MAX_BANDS

// In: a (symmetric positive semi-definite, full storage).  Out: vec ~ eigenvector of the largest eigenvalue.
// Repeated squaring, a <- a a / trace(a a): after s squarings a is proportional to sum_i (l_i / l_1)^(2^s) v_i v_i^T,
// so any non-zero column is the top eigenvector up to e = (l_2 / l_1)^(2^s).  The caller's f64 Rayleigh quotient
// then misses l_1 by at most (l_1 - l_2) e^2 <= l_1 max_g g exp(-2^(s+1) g) = l_1 / (e 2^(s+1)): 2.8e-6 l_1 for
// s = 16, attained only when the two largest eigenvalues agree to 1e-5 (where either is as good a bound), and
// far below that otherwise.  Unlike cyclic Jacobi (serial rotations, ~100 k dependent cycles per matrix) the 36
// dot products of a squaring are independent, so a thread pipelines them at full rate.
__device__ __forceinline__ void power8_top_vector(float (&a)[JN][JN], float (&vec)[JN]) {
    float tr = 0.f;
#pragma unroll
    for (int i = 0; i < JN; i++) tr += a[i][i];
    if (!(tr > 0.f)) {
#pragma unroll
        for (int i = 0; i < JN; i++) vec[i] = (i == 0) ? 1.f : 0.f;
        return;
    }
    {
        const float s = 1.f / tr;
#pragma unroll
        for (int i = 0; i < JN; i++)
#pragma unroll
            for (int j = 0; j < JN; j++) a[i][j] *= s;
    }
#pragma unroll 1
    for (int it = 0; it < 16; it++) {
        float b[JN][JN];
        float t2 = 0.f;
#pragma unroll
        for (int i = 0; i < JN; i++)
#pragma unroll
            for (int j = i; j < JN; j++) {
                float d = 0.f;
#pragma unroll
                for (int k = 0; k < JN; k++) d += a[i][k] * a[j][k];
                b[i][j] = d;
                if (i == j) t2 += d;
            }
        const float s = 1.f / t2;          // t2 = sum l_i^2 / (sum l_i)^2 >= 1/8 of the previous normalisation
#pragma unroll
        for (int i = 0; i < JN; i++)
#pragma unroll
            for (int j = i; j < JN; j++) { const float v = b[i][j] * s; a[i][j] = v; a[j][i] = v; }
    }
    int best = 0;
    float top = a[0][0];
#pragma unroll
    for (int i = 1; i < JN; i++)
        if (a[i][i] > top) { top = a[i][i]; best = i; }
#pragma unroll
    for (int r = 0; r < JN; r++) {
        float x = a[r][0];
#pragma unroll
        for (int j = 1; j < JN; j++) x = (best == j) ? a[r][j] : x;
        vec[r] = x;
    }
}

}  // namespace

// grid: ceil(2B / 128); block: 128 threads; thread 2b handles S S^T of blend b, thread 2b+1 handles A^T A.
extern "C" __global__ void __launch_bounds__(128)
scl_k_lipschitz(const float* __restrict__ A,          // [Ktot, C]  current SEDs
                const int* __restrict__ comp_off,     // [B+1]
                const double* __restrict__ gram,      // [B, 8, 8]  S S^T (K1), rows/cols >= K unspecified
                const double* __restrict__ loss_part, // [B, P]  one entry per image plane
                const int* __restrict__ active,       // [B]
                const int* __restrict__ blend_list,   // [nB] compacted active list
                const int* __restrict__ it0,          // [B] iterations done before this fit call
                const double* __restrict__ prev_loss, // [B] last loss of an earlier fit call (if it0 > 0)
                double* __restrict__ logL,            // [B, logL_cap] loss of step s of THIS fit call
                float* __restrict__ lips,             // [B, 2] (L_sed, L_morph)
                int nB, int C, int P, int step, int logL_cap, int n_obs, int approximate_L)
{
    const int gt = blockIdx.x * blockDim.x + threadIdx.x;
    const int which = gt & 1;
    const int b = (gt >> 1) < nB ? blend_list[gt >> 1] : -1;
    const bool live = b >= 0 && active[b];
    const int k0 = live ? comp_off[b] : 0, K = live ? comp_off[b + 1] - k0 : 0;
    const int n = which ? C : K;

    double lam = 0.0;
    if (live) {
        if (approximate_L) {
            if (which == 0) { for (int k = 0; k < K; k++) lam += gram[(size_t)b * 64 + k * 8 + k]; }
            else { for (int i = 0; i < K * C; i++) { double x = (double)A[(size_t)k0 * C + i]; lam += x * x; } }
        } else {
            float a[JN][JN];
            if (which == 0) {
#pragma unroll
                for (int i = 0; i < JN; i++)
#pragma unroll
                    for (int j = 0; j < JN; j++)
                        a[i][j] = (i < K && j < K) ? (float)gram[(size_t)b * 64 + (i < j ? i * 8 + j : j * 8 + i)] : 0.f;
            } else {
#pragma unroll
                for (int i = 0; i < JN; i++)
#pragma unroll
                    for (int j = 0; j < JN; j++) a[i][j] = 0.f;
                for (int k = 0; k < K; k++) {
                    float row[JN];
#pragma unroll
                    for (int c = 0; c < JN; c++) row[c] = (c < C) ? A[(size_t)(k0 + k) * C + c] : 0.f;
#pragma unroll
                    for (int i = 0; i < JN; i++)
#pragma unroll
                        for (int j = 0; j < JN; j++) a[i][j] += row[i] * row[j];
                }
            }
            float vec[JN];
            power8_top_vector(a, vec);
            // f64 Rayleigh quotient v^T M v / v^T v
            double num = 0.0, den = 0.0;
#pragma unroll
            for (int i = 0; i < JN; i++) den += (double)vec[i] * (double)vec[i];
            if (which == 0) {
#pragma unroll
                for (int i = 0; i < JN; i++) {
                    if (i < K) {
                        double w = 0.0;
#pragma unroll
                        for (int j = 0; j < JN; j++)
                            if (j < K) w += gram[(size_t)b * 64 + (i < j ? i * 8 + j : j * 8 + i)] * (double)vec[j];
                        num += w * (double)vec[i];
                    }
                }
            } else {
                for (int k = 0; k < K; k++) {
                    double w = 0.0;
#pragma unroll
                    for (int c = 0; c < JN; c++)
                        if (c < C) w += (double)A[(size_t)(k0 + k) * C + c] * (double)vec[c];
                    num += w * w;
                }
            }
            lam = (den > 0.0 && n > 0) ? num / den : 0.0;
        }
    }
    // exchange: lane pair (2b, 2b+1) -> both know LA (from S S^T) and LS (from A^T A)
    const double other = __shfl_xor_sync(0xffffffffu, lam, 1);
    if (!live) return;
    double LA = which ? other : lam, LS = which ? lam : other;
    double total = 0.0;
    for (int c = 0; c < P; c++) total += loss_part[b * P + c];      // Blend._loss: sum over observations
    if (approximate_L) {
        const int it = it0[b] + step;              // index of this loss in Blend.mse
        if (it >= 1) {
            // blend.py:199: len(mse) > 1 (the loss just appended counts) and mse[-1] > mse[-2]
            const double prev = (step >= 1) ? logL[(size_t)b * logL_cap + step - 1] : prev_loss[b];
            if (total > prev) { LA *= 2; LS *= 2; }
        }
    }
    const float L_sed = (float)(LA * n_obs);       // blend.py:219-222: the morph Gram bounds the SED step
    const float L_morph = (float)(LS * n_obs);
    if (which == 0) {
        if (step < logL_cap) logL[(size_t)b * logL_cap + step] = total;
        lips[b * 2] = L_sed;
        lips[b * 2 + 1] = L_morph;
    }
}
