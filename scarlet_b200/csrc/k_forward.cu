// K1: scene render -> PSF-matching convolution -> weighted residual / logL ->
// adjoint convolution, fused per (blend, band) in one CTA with the 128x128
// half-spectrum held in shared memory.
//
// Replaces, for one band of one observation (reference file:line):
//   ComponentTree.get_model            component.py:321-347  (M_c = sum_k A[k,c] S_k)
//   Observation.render / _convolve     observation.py:198-220 -> fft.convolve fft.py:304-317
//   Observation.get_loss               observation.py:222-239 (0.5 sum (W (R - I))^2)
//   the autograd adjoint of the above  blend.py:44-45,105-119 (G_c = conv^T(W^2 (R - I)))
//
// Convolution layout: the model occupies [0,Ny)x[0,Nx) of a 128x128 periodic
// grid and the kernel spectrum D is that of the truncated real-space difference
// kernel with its centre tap wrapped to index 0, so the circular convolution
// restricted to [0,Ny)x[0,Nx) equals the reference's pad/ifftshift/crop result
// for any FFT length >= N + k/2 (DESIGN.md "FFT length").
//
// Real-input packing: rows (2j, 2j+1) are transformed as one complex line
// (line j = smem rows 2j and 2j+1, 130 slots), untangled into 64 columns with
// the DC and Nyquist columns packed as re/im of column 0; column FFTs follow.
// Every step is owned by a half-warp per line, so the kernel needs only four
// block barriers.
#include "fft_reg.cuh"

namespace {

constexpr int F = SCL_FFT;
constexpr int RS = SCL_RS;
constexpr int LS = 2 * SCL_RS;   // packed-line stride (two smem rows)

// untangle line j (after the row FFT) in place: Z[k], Z[128-k] -> A[k] (row 2j), B[k] (row 2j+1)
__device__ __forceinline__ void untangle_line(float2* line, int t, bool live) {
    float2 A[4], B[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int k = t + 16 * i;
        float2 z1 = make_float2(0.f, 0.f), z2 = z1;
        if (live) {
            z1 = line[k];
            z2 = line[k == 0 ? 64 : 128 - k];
        }
        if (k == 0) {
            A[i] = make_float2(z1.x, z2.x);
            B[i] = make_float2(z1.y, z2.y);
        } else {
            A[i] = make_float2(0.5f * (z1.x + z2.x), 0.5f * (z1.y - z2.y));
            B[i] = make_float2(0.5f * (z1.y + z2.y), 0.5f * (z2.x - z1.x));
        }
    }
    __syncwarp();
    if (live) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            int k = t + 16 * i;
            line[k] = A[i];
            line[RS + k] = B[i];
        }
    }
    __syncwarp();
}

// inverse of untangle_line, leaving swap(Z) in the packed line (ready for the
// forward FFT that implements the inverse transform).  Input rows hold swap(u).
__device__ __forceinline__ void retangle_line(float2* line, int t, bool live) {
    float2 Zk[4], Zm[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int k = t + 16 * i;
        float2 a = make_float2(0.f, 0.f), b = a;
        if (live) {
            a = cswap(line[k]);
            b = cswap(line[RS + k]);
        }
        if (k == 0) {
            Zk[i] = make_float2(a.x, b.x);      // Z[0]  = Re a0 + i Re b0
            Zm[i] = make_float2(a.y, b.y);      // Z[64] = Im a0 + i Im b0
        } else {
            Zk[i] = make_float2(a.x - b.y, a.y + b.x);      // A + iB
            Zm[i] = make_float2(a.x + b.y, b.x - a.y);      // conj(A) + i conj(B)
        }
    }
    __syncwarp();
    if (live) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            int k = t + 16 * i;
            line[k] = cswap(Zk[i]);
            line[k == 0 ? 64 : 128 - k] = cswap(Zm[i]);
        }
    }
    __syncwarp();
}

// column phase: forward column FFT, multiply by D (or conj D), inverse column FFT (swapped domain)
template <bool CONJ>
__device__ __forceinline__ void column_phase(float2* buf, const float2* tw, const float2* __restrict__ D, int NR) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, t = lane & 15, half = lane >> 4;
    const int nwarp = blockDim.x >> 5;
    for (int c0 = warp * 2; c0 < 64; c0 += nwarp * 2) {
        int kx = c0 + half;
        float2* col = buf + kx;
        fft_line16<128>(col, RS, t, tw, NR, true);
        if (kx > 0) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                int ky = t + 16 * i;
                float2 d = D[ky * RS + kx];
                float2 v = col[ky * RS];
                v = CONJ ? cmulc(v, d) : cmul(v, d);
                col[ky * RS] = cswap(v);
            }
        } else {
            // column 0 carries C = Col0 + i ColN; pairs (ky, 128-ky), ky = 0..64
            for (int ky = t; ky <= 64; ky += 16) {
                int k2 = (128 - ky) & 127;
                float2 c1 = col[ky * RS], c2 = col[k2 * RS];
                float2 p0 = make_float2(0.5f * (c1.x + c2.x), 0.5f * (c1.y - c2.y));     // Col0[ky]
                float2 pn = make_float2(0.5f * (c1.y + c2.y), 0.5f * (c2.x - c1.x));     // ColN[ky]
                float2 q0 = make_float2(p0.x, -p0.y);                                    // Col0[k2] = conj
                float2 qn = make_float2(pn.x, -pn.y);                                    // ColN[k2] = conj
                float2 d0 = D[ky * RS], dn = D[ky * RS + 64];
                float2 e0 = D[k2 * RS], en = D[k2 * RS + 64];
                p0 = CONJ ? cmulc(p0, d0) : cmul(p0, d0);
                pn = CONJ ? cmulc(pn, dn) : cmul(pn, dn);
                q0 = CONJ ? cmulc(q0, e0) : cmul(q0, e0);
                qn = CONJ ? cmulc(qn, en) : cmul(qn, en);
                col[ky * RS] = cswap(make_float2(p0.x - pn.y, p0.y + pn.x));
                col[k2 * RS] = cswap(make_float2(q0.x - qn.y, q0.y + qn.x));
            }
        }
        __syncwarp();
        fft_line16<128>(col, RS, t, tw, 128, true);
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Register-FFT version of the fused kernel (fft_reg.cuh): 8 lanes per 128-point line, one exchange per
// transform, conflict-free shared-memory access.
namespace {

// Row line (pair of image rows) and spectrum column owned by this thread's 8-lane group.  The two groups of
// a half-warp get lines 4 apart (4 * LS = 8 mod 16 float2) and columns 8 apart: conflict-free 64-bit accesses.
__device__ __forceinline__ int grp_row_line() {
    const int g = threadIdx.x >> 3, w = g >> 2, a = g & 1, b = (g >> 1) & 1;
    return ((w & 1) * 2 + b) + 4 * a + 8 * (w >> 1);
}
__device__ __forceinline__ int grp_column() {
    const int g = threadIdx.x >> 3, w = g >> 2, a = g & 1, b = (g >> 1) & 1;
    return ((w & 3) * 2 + b) + 8 * a + 16 * (w >> 2);
}

// freq-distributed spectrum Z of a packed row pair -> half spectra A (row0) and B (row1), columns 0..63,
// with (DC, Nyquist) of each packed as re/im of column 0.
__device__ __forceinline__ void untangle_store(const float2 (&v)[16], float2* row0, int t, bool live) {
    float2 r[16];
    fft128_partner<0, 4>(v, r, t);
    if (!live) return;
#pragma unroll
    for (int s = 0; s < 2; s++)
#pragma unroll
        for (int k2 = 0; k2 < 4; k2++) {
            const int k = t + 8 * s + 16 * k2;
            const float2 z1 = v[8 * s + k2];
            float2 z2 = r[8 * s + k2];
            float2 A, B;
            if (s == 0 && k2 == 0 && t == 0) {
                z2 = v[4];                                    // Z[64]
                A = make_float2(z1.x, z2.x);
                B = make_float2(z1.y, z2.y);
            } else {
                A = make_float2(0.5f * (z1.x + z2.x), 0.5f * (z1.y - z2.y));
                B = make_float2(0.5f * (z1.y + z2.y), 0.5f * (z2.x - z1.x));
            }
            row0[k] = A;
            row0[RS + k] = B;
        }
}

// inverse of untangle_store: rows hold swap(A'), swap(B') (the column phase leaves its result swapped);
// returns swap(Z') freq-distributed, ready for fft128_f2t acting as the inverse transform.
__device__ __forceinline__ void load_retangle(float2 (&v)[16], const float2* row0, int t, bool live) {
    float2 w[16], r[16];
    float2 z64 = make_float2(0.f, 0.f);
#pragma unroll
    for (int s = 0; s < 2; s++)
#pragma unroll
        for (int k2 = 0; k2 < 4; k2++) {
            const int k = t + 8 * s + 16 * k2;
            float2 a = make_float2(0.f, 0.f), b = a;
            if (live) { a = cswap(row0[k]); b = cswap(row0[RS + k]); }
            if (s == 0 && k2 == 0 && t == 0) {
                v[0] = make_float2(a.x, b.x);                 // Z[0]
                z64 = make_float2(a.y, b.y);                  // Z[64]
                w[0] = v[0];
            } else {
                v[8 * s + k2] = make_float2(a.x - b.y, a.y + b.x);      // A + iB
                w[8 * s + k2] = make_float2(a.x + b.y, b.x - a.y);      // conj(A) + i conj(B) = Z[128 - k]
            }
        }
#pragma unroll
    for (int i = 0; i < 16; i++) if ((i & 7) >= 4) w[i] = make_float2(0.f, 0.f);
    fft128_partner<4, 8>(w, r, t);
#pragma unroll
    for (int s = 0; s < 2; s++)
#pragma unroll
        for (int k2 = 4; k2 < 8; k2++) v[8 * s + k2] = r[8 * s + k2];
    if (t == 0) v[4] = z64;
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = cswap(v[i]);
}

// forward column FFT, x D (or conj D), inverse column FFT; result left swapped in rows [0, NR).
// Dt: transposed kernel spectrum [65][128] (Dt[kx*128 + ky]).
template <bool HALF>
__device__ __forceinline__ void column_phase_reg(float2* buf, const float2* tw, const float2* __restrict__ Dt, int NR,
                                                 bool conj) {
    const float cs = conj ? -1.f : 1.f;          // multiply by D or by conj(D)
    const int t = threadIdx.x & 7;
#pragma unroll 1
    for (int round = 0; round < 2; round++) {
        const int kx = 32 * round + grp_column();
        float2* col = buf + kx;
        float2 v[16];
#pragma unroll
        for (int n1 = 0; n1 < 16; n1++) {
            const int y = 8 * n1 + t;
            v[n1] = (y < NR && (!HALF || n1 < 8)) ? col[y * RS] : make_float2(0.f, 0.f);
        }
        fft128_t2f<HALF>(v, col, RS, t, tw);
        if (kx > 0) {
            const float2* Dc = Dt + kx * 128;
#pragma unroll
            for (int s = 0; s < 2; s++)
#pragma unroll
                for (int k2 = 0; k2 < 8; k2++) {
                    float2 d = Dc[t + 8 * s + 16 * k2];
                    d.y *= cs;
                    v[8 * s + k2] = cswap(cmul(v[8 * s + k2], d));
                }
        }
        if (__any_sync(0xffffffffu, kx == 0)) {
            // column 0 carries C = Col0 + i ColN (kx = 0 and kx = 64): split, multiply each, recombine
            float2 r[16];
            fft128_partner<0, 8>(v, r, t);
            if (kx == 0) {
#pragma unroll
                for (int s = 0; s < 2; s++)
#pragma unroll
                    for (int k2 = 0; k2 < 8; k2++) {
                        const int ky = t + 8 * s + 16 * k2;
                        const float2 c1 = v[8 * s + k2], c2 = r[8 * s + k2];
                        float2 p0 = make_float2(0.5f * (c1.x + c2.x), 0.5f * (c1.y - c2.y));
                        float2 pn = make_float2(0.5f * (c1.y + c2.y), 0.5f * (c2.x - c1.x));
                        float2 d0 = Dt[ky], dn = Dt[64 * 128 + ky];
                        d0.y *= cs; dn.y *= cs;
                        p0 = cmul(p0, d0);
                        pn = cmul(pn, dn);
                        v[8 * s + k2] = cswap(make_float2(p0.x - pn.y, p0.y + pn.x));
                    }
            }
        }
        fft128_f2t<HALF>(v, col, RS, t, tw);
#pragma unroll
        for (int n1 = 0; n1 < (HALF ? 8 : 16); n1++) {
            const int y = 8 * n1 + t;
            if (y < NR) col[y * RS] = v[n1];
        }
    }
}

}  // namespace

// grid: (P * n_listed_blends), P = pl.P image planes per blend (one per band of every observation); block: SCL_THREADS; dynamic smem: kFwdSmem (api.cu)
//   buf [128][65] float2 | tw [256] float2 | dummy line [130] float2 | red [32] double
// HALF: Nx <= 64 and Ny <= 64 -- the zero-padded halves of every line are pruned from the transforms.
template <bool HALF>
__global__ void __launch_bounds__(SCL_THREADS, 2)
scl_k_forward_adjoint(const float* __restrict__ A,        // [Ktot, C]
                      const float* __restrict__ S,        // [Ktot, N]
                      const int* __restrict__ comp_off,   // [B+1]
                      const float* __restrict__ images,   // [B, P, N]
                      const float* __restrict__ weights,  // [B, P, N] or nullptr (W = 1)
                      const float2* __restrict__ Dhat_t,  // [n_ker, 65, 128] transposed kernel spectra
                      const int* __restrict__ ker_idx,    // [B*P], -1 = identity (no PSF matching)
                      const int* __restrict__ active,     // [B]
                      const int* __restrict__ blend_list, // [gridDim.x / P] blends to process (compacted active list)
                      float* __restrict__ G,              // [B, PG, N]  adjoint image of every plane (PG >= P: the planes
                                                          //             of low-resolution observations follow, k_lowres.cu)
                      double* __restrict__ loss_part,     // [B, PG]
                      double* __restrict__ gram,          // [B, 8, 8]  rows r = plane, plane+P, .. of S S^T
                      int Cm, SclPlanes pl, int PG, int Ny, int Nx)
{
    extern __shared__ __align__(16) unsigned char smem_fwd[];
    float2* buf = reinterpret_cast<float2*>(smem_fwd);
    float2* tw = buf + F * RS;
    float2* dummy = tw + 256;
    double* red = reinterpret_cast<double*>(dummy + LS);
    __shared__ float sA[SCL_MAX_COMP];
    __shared__ double sred[SCL_THREADS / 32][2 * SCL_MAX_COMP];

    // c: this CTA's plane; C: planes per blend (work split of the Gram rows); cm: the model band the plane sees
    const int C = pl.P, c = blockIdx.x % C, b = blend_list[blockIdx.x / C];
    int cm = 0;                      // pl.band[c] without indexing the parameter struct dynamically
#pragma unroll
    for (int p = 0; p < SCL_MAX_BANDS; p++) cm = (p == c) ? pl.band[p] : cm;
    if (!active[b]) return;
    const int k0 = comp_off[b], K = comp_off[b + 1] - k0;
    const int N = Ny * Nx;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, t = threadIdx.x & 7;
    const int nwarp = blockDim.x >> 5;
    const int NR = (Ny + 1) & ~1, NP = NR >> 1;
    const float inv = 1.0f / (float)(F * F);
    const int kid = ker_idx[b * C + c];
    const float2* Dt = Dhat_t + (size_t)(kid < 0 ? 0 : kid) * RS * F;
    const float* img = images + ((size_t)b * C + c) * N;
    const float* wgt = weights ? weights + ((size_t)b * C + c) * N : nullptr;
    float* gout = G + ((size_t)b * PG + c) * N;

    // the residual pass reads the band's image and weights one whole transform later: start them towards L2 now
    for (int i = threadIdx.x * 32; i < N; i += blockDim.x * 32) {
        asm volatile("prefetch.global.L2 [%0];\n" :: "l"(img + i));
        if (wgt) asm volatile("prefetch.global.L2 [%0];\n" :: "l"(wgt + i));
    }

    fill_twiddles_reg(tw);
    if (threadIdx.x < SCL_MAX_COMP) sA[threadIdx.x] = (threadIdx.x < K) ? A[(size_t)(k0 + threadIdx.x) * Cm + cm] : 0.f;
    __syncthreads();

    double loss = 0.0;
    // Gram rows r0 = c and r1 = c + C of S S^T (Blend._set_lipschitz, blend.py:205-217) ride on the render loads;
    // further rows (only when K > 2C) take the separate pass at the end
    float gr0[SCL_MAX_COMP], gr1[SCL_MAX_COMP];
#pragma unroll
    for (int k = 0; k < SCL_MAX_COMP; k++) { gr0[k] = 0.f; gr1[k] = 0.f; }
    const int r0 = c, r1 = c + C;

    if (kid < 0) {
        // identical PSFs: R = M, G = W^2 (M - I)   (observation.py:216-220)
        for (int p = threadIdx.x; p < N; p += blockDim.x) {
            float m = 0.f;
            for (int k = 0; k < K; k++) m += sA[k] * S[(size_t)(k0 + k) * N + p];
            float w = wgt ? wgt[p] : 1.f;
            float d = w * (m - img[p]);
            loss += (double)d * (double)d;
            gout[p] = w * d;
        }
        for (int p = threadIdx.x; p < N; p += blockDim.x) {
            float sv[SCL_MAX_COMP];
#pragma unroll
            for (int k = 0; k < SCL_MAX_COMP; k++) sv[k] = (k < K) ? S[(size_t)(k0 + k) * N + p] : 0.f;
            float a0 = 0.f, a1 = 0.f;
#pragma unroll
            for (int k = 0; k < SCL_MAX_COMP; k++) { a0 = (k == r0) ? sv[k] : a0; a1 = (k == r1) ? sv[k] : a1; }
#pragma unroll
            for (int k = 0; k < SCL_MAX_COMP; k++) { gr0[k] += a0 * sv[k]; gr1[k] += a1 * sv[k]; }
        }
        // Gram rows c and c + C: reduce now so that the accumulators die before the column phase
        {
#pragma unroll
            for (int l = 0; l < SCL_MAX_COMP; l++) {
                double v0 = warp_sum<double>((double)gr0[l]), v1 = warp_sum<double>((double)gr1[l]);
                if (lane == 0) { sred[warp][l] = v0; sred[warp][SCL_MAX_COMP + l] = v1; }
            }
            __syncthreads();
            if (threadIdx.x < 2 * SCL_MAX_COMP) {
                const int l = threadIdx.x & (SCL_MAX_COMP - 1), r = (threadIdx.x < SCL_MAX_COMP) ? r0 : r1;
                if (r < K && l < K) {
                    double v = 0.0;
                    for (int w = 0; w < nwarp; w++) v += sred[w][threadIdx.x];
                    gram[(size_t)b * 64 + r * 8 + l] = v;
                }
            }
        }
    } else {
        const int jg = grp_row_line();
        // Three row passes share one copy of the code (the unrolled FFT bodies are large; one copy keeps the kernel
        // inside the instruction cache):
        //   pass 0: render row pairs (ComponentTree.get_model)            -> row FFT -> columns x D
        //   pass 1: inverse rows -> R; r = W^2 (R - I), logL              -> row FFT -> columns x conj(D)
        //   pass 2: inverse rows -> G
#pragma unroll 1
        for (int pass = 0; pass < 3; pass++) {
            for (int j0 = 0; j0 < NP; j0 += 32) {
                const int j = j0 + jg;
                const bool live = j < NP;
                float2* line = live ? buf + j * LS : dummy;
                const int y0 = 2 * j, y1 = 2 * j + 1;
                float2 v[16];
                if (pass > 0) {
                    load_retangle(v, line, t, live);
                    fft128_f2t<HALF>(v, line, 1, t, tw);
                }
                if (pass == 0) {
#pragma unroll
                    for (int n1 = (HALF ? 8 : 16); n1 < 16; n1++) v[n1] = make_float2(0.f, 0.f);
#pragma unroll
                    for (int n1 = 0; n1 < (HALF ? 8 : 16); n1++) {
                        const int x = 8 * n1 + t;
                        float m0 = 0.f, m1 = 0.f;
                        if (live && x < Nx) {
                            float s0[SCL_MAX_COMP], s1[SCL_MAX_COMP];
#pragma unroll
                            for (int k = 0; k < SCL_MAX_COMP; k++) {
                                const float* Sk = S + (size_t)(k0 + k) * N;
                                s0[k] = (k < K) ? Sk[y0 * Nx + x] : 0.f;
                                s1[k] = (k < K && y1 < Ny) ? Sk[y1 * Nx + x] : 0.f;
                            }
                            float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;      // S[r0], S[r1] at the two pixels
#pragma unroll
                            for (int k = 0; k < SCL_MAX_COMP; k++) {
                                m0 += sA[k] * s0[k];
                                m1 += sA[k] * s1[k];
                                a0 = (k == r0) ? s0[k] : a0; b0 = (k == r0) ? s1[k] : b0;
                                a1 = (k == r1) ? s0[k] : a1; b1 = (k == r1) ? s1[k] : b1;
                            }
#pragma unroll
                            for (int k = 0; k < SCL_MAX_COMP; k++) {
                                gr0[k] += a0 * s0[k] + b0 * s1[k];
                                gr1[k] += a1 * s0[k] + b1 * s1[k];
                            }
                        }
                        v[n1] = make_float2(m0, m1);
                    }
                } else if (pass == 1) {
#pragma unroll
                    for (int n1 = (HALF ? 8 : 16); n1 < 16; n1++) v[n1] = make_float2(0.f, 0.f);
#pragma unroll
                    for (int n1 = 0; n1 < (HALF ? 8 : 16); n1++) {
                        const int x = 8 * n1 + t;
                        float o0 = 0.f, o1 = 0.f;
                        if (live && x < Nx) {
                            const float rr0 = v[n1].y * inv, rr1 = v[n1].x * inv;       // swapped: (im, re)
                            const int p0 = y0 * Nx + x, p1 = y1 * Nx + x;
                            const float w0 = wgt ? wgt[p0] : 1.f;
                            const float d0 = w0 * (rr0 - img[p0]);
                            loss += (double)d0 * (double)d0;
                            o0 = w0 * d0;
                            if (y1 < Ny) {
                                const float w1 = wgt ? wgt[p1] : 1.f;
                                const float d1 = w1 * (rr1 - img[p1]);
                                loss += (double)d1 * (double)d1;
                                o1 = w1 * d1;
                            }
                        }
                        v[n1] = make_float2(o0, o1);
                    }
                } else {
#pragma unroll
                    for (int n1 = 0; n1 < (HALF ? 8 : 16); n1++) {
                        const int x = 8 * n1 + t;
                        if (live && x < Nx) {
                            gout[y0 * Nx + x] = v[n1].y * inv;
                            if (y1 < Ny) gout[y1 * Nx + x] = v[n1].x * inv;
                        }
                    }
                }
                if (pass < 2) {
                    fft128_t2f<HALF>(v, line, 1, t, tw);
                    untangle_store(v, line, t, live);
                }
            }
            if (pass == 2) break;
            if (pass == 0) {
                // Gram rows c and c + C: reduce now so that the accumulators die before the column phase
#pragma unroll
                for (int l = 0; l < SCL_MAX_COMP; l++) {
                    double v0 = warp_sum<double>((double)gr0[l]), v1 = warp_sum<double>((double)gr1[l]);
                    if (lane == 0) { sred[warp][l] = v0; sred[warp][SCL_MAX_COMP + l] = v1; }
                }
                __syncthreads();
                if (threadIdx.x < 2 * SCL_MAX_COMP) {
                    const int l = threadIdx.x & (SCL_MAX_COMP - 1), r = (threadIdx.x < SCL_MAX_COMP) ? r0 : r1;
                    if (r < K && l < K) {
                        double v = 0.0;
                        for (int w = 0; w < nwarp; w++) v += sred[w][threadIdx.x];
                        gram[(size_t)b * 64 + r * 8 + l] = v;
                    }
                }
            }
            __syncthreads();
            column_phase_reg<HALF>(buf, tw, Dt, NR, pass == 1);
            __syncthreads();
        }
    }
    double tot = block_sum<double>(loss, red);
    if (threadIdx.x == 0) loss_part[b * PG + c] = 0.5 * tot;

    // ---- remaining rows r = c + 2C, ... (K > 2C only)
    for (int r = c + 2 * C; r < K; r += C) {
        float acc[SCL_MAX_COMP];
#pragma unroll
        for (int l = 0; l < SCL_MAX_COMP; l++) acc[l] = 0.f;
        const float* Sr = S + (size_t)(k0 + r) * N;
        for (int p = threadIdx.x; p < N; p += blockDim.x) {
            const float sr = Sr[p];
#pragma unroll
            for (int l = 0; l < SCL_MAX_COMP; l++)
                if (l < K) acc[l] += sr * S[(size_t)(k0 + l) * N + p];
        }
        __syncthreads();
#pragma unroll
        for (int l = 0; l < SCL_MAX_COMP; l++) {
            double v = warp_sum<double>((double)acc[l]);
            if (lane == 0) sred[warp][l] = v;
        }
        __syncthreads();
        if (threadIdx.x < K) {
            double v = 0.0;
            for (int w = 0; w < nwarp; w++) v += sred[w][threadIdx.x];
            gram[(size_t)b * 64 + r * 8 + threadIdx.x] = v;
        }
    }
}

#define SCL_FWD_ARGS const float*, const float*, const int*, const float*, const float*, const float2*, const int*, \
    const int*, const int*, float*, double*, double*, int, SclPlanes, int, int, int
template __global__ void scl_k_forward_adjoint<true>(SCL_FWD_ARGS);
template __global__ void scl_k_forward_adjoint<false>(SCL_FWD_ARGS);

// [n][128][65] -> [n][65][128]; grid: n * 65 blocks of 128 threads
extern "C" __global__ void scl_k_transpose_spec(const float2* __restrict__ in, float2* __restrict__ out) {
    const int n = blockIdx.x / RS, kx = blockIdx.x % RS, ky = threadIdx.x;
    out[((size_t)n * RS + kx) * F + ky] = in[((size_t)n * F + ky) * RS + kx];
}

// Standalone render: observation.render(model) for host callers (drop-in
// Observation.render, observation.py:203-220).  model [B,C,N] -> out [B,C,N].
extern "C" __global__ void __launch_bounds__(SCL_THREADS)
scl_k_convolve(const float* __restrict__ model, const float2* __restrict__ Dhat,
               const int* __restrict__ ker_idx, float* __restrict__ out, int C, int Ny, int Nx, int conj)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* buf = reinterpret_cast<float2*>(smem_raw);
    float2* tw = buf + F * RS;
    const int c = blockIdx.x, b = blockIdx.y;
    const int N = Ny * Nx;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, t = lane & 15, half = lane >> 4;
    const int nwarp = blockDim.x >> 5;
    const int NR = (Ny + 1) & ~1, NP = NR >> 1;
    const float inv = 1.0f / (float)(F * F);
    const int kid = ker_idx[b * C + c];
    const float* src = model + ((size_t)b * C + c) * N;
    float* dst = out + ((size_t)b * C + c) * N;
    if (kid < 0) {
        for (int p = threadIdx.x; p < N; p += blockDim.x) dst[p] = src[p];
        return;
    }
    const float2* D = Dhat + (size_t)kid * F * RS;
    fill_twiddles128(tw);
    __syncthreads();
    for (int j0 = warp * 2; j0 < NP; j0 += nwarp * 2) {
        int j = j0 + half;
        bool live = j < NP;
        float2* line = buf + j * LS;
        if (live)
            for (int x = t; x < Nx; x += 16)
                line[x] = make_float2(src[2 * j * Nx + x], (2 * j + 1 < Ny) ? src[(2 * j + 1) * Nx + x] : 0.f);
        __syncwarp();
        fft_line16<128>(line, 1, t, tw, Nx, live);
        untangle_line(line, t, live);
    }
    __syncthreads();
    if (conj) column_phase<true>(buf, tw, D, NR); else column_phase<false>(buf, tw, D, NR);
    __syncthreads();
    for (int j0 = warp * 2; j0 < NP; j0 += nwarp * 2) {
        int j = j0 + half;
        bool live = j < NP;
        float2* line = buf + j * LS;
        retangle_line(line, t, live);
        fft_line16<128>(line, 1, t, tw, 128, live);
        if (live)
            for (int x = t; x < Nx; x += 16) {
                float2 z = line[x];
                dst[2 * j * Nx + x] = z.y * inv;
                if (2 * j + 1 < Ny) dst[(2 * j + 1) * Nx + x] = z.x * inv;
            }
    }
}
