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
#include "fft_smem.cuh"

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

// grid: (C * n_listed_blends); block: SCL_THREADS; dynamic smem: (128*65 + 128) float2 + 32 doubles
extern "C" __global__ void __launch_bounds__(SCL_THREADS)
scl_k_forward_adjoint(const float* __restrict__ A,        // [Ktot, C]
                      const float* __restrict__ S,        // [Ktot, N]
                      const int* __restrict__ comp_off,   // [B+1]
                      const float* __restrict__ images,   // [B, C, N]
                      const float* __restrict__ weights,  // [B, C, N] or nullptr (W = 1)
                      const float2* __restrict__ Dhat,    // [n_ker, 128, 65]
                      const int* __restrict__ ker_idx,    // [B*C], -1 = identity (no PSF matching)
                      const int* __restrict__ active,     // [B]
                      const int* __restrict__ blend_list, // [gridDim.x / C] blends to process (compacted active list)
                      float* __restrict__ G,              // [B, C, N]
                      double* __restrict__ loss_part,     // [B, C]
                      float* __restrict__ dA,             // [Ktot, C]  sum_p G[c,p] S[k,p]   (K <= 8)
                      double* __restrict__ gram,          // [B, 8, 8]  rows r = c, c+C, .. of S S^T
                      int C, int Ny, int Nx)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* buf = reinterpret_cast<float2*>(smem_raw);
    float2* tw = buf + F * RS;
    double* red = reinterpret_cast<double*>(tw + 128);
    __shared__ float sA[64];
    __shared__ double sred[SCL_THREADS / 32][2 * SCL_MAX_COMP];

    const int c = blockIdx.x % C, b = blend_list[blockIdx.x / C];
    if (!active[b]) return;
    const int k0 = comp_off[b], K = comp_off[b + 1] - k0;
    const int N = Ny * Nx;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, t = lane & 15, half = lane >> 4;
    const int nwarp = blockDim.x >> 5;
    const int NR = (Ny + 1) & ~1, NP = NR >> 1;
    const float inv = 1.0f / (float)(F * F);
    const int kid = ker_idx[b * C + c];
    const float2* D = Dhat + (size_t)(kid < 0 ? 0 : kid) * F * RS;
    const float* img = images + ((size_t)b * C + c) * N;
    const float* wgt = weights ? weights + ((size_t)b * C + c) * N : nullptr;
    float* gout = G + ((size_t)b * C + c) * N;

    fill_twiddles128(tw);
    for (int k = threadIdx.x; k < K && k < 64; k += blockDim.x) sA[k] = A[(size_t)(k0 + k) * C + c];
    __syncthreads();

    double loss = 0.0;
    float accA[SCL_MAX_COMP];            // this thread's share of dA[k, c]
#pragma unroll
    for (int k = 0; k < SCL_MAX_COMP; k++) accA[k] = 0.f;

    if (kid < 0) {
        // identical PSFs: R = M, G = W^2 (M - I)   (observation.py:216-220)
        for (int p = threadIdx.x; p < N; p += blockDim.x) {
            float m = 0.f;
            for (int k = 0; k < K; k++) m += A[(size_t)(k0 + k) * C + c] * S[(size_t)(k0 + k) * N + p];
            float w = wgt ? wgt[p] : 1.f;
            float d = w * (m - img[p]);
            loss += (double)d * (double)d;
            const float g = w * d;
            gout[p] = g;
#pragma unroll
            for (int k = 0; k < SCL_MAX_COMP; k++)
                if (k < K) accA[k] += g * S[(size_t)(k0 + k) * N + p];
        }
    } else {
        // ---- phase A: render rows, row FFT, untangle
        for (int j0 = warp * 2; j0 < NP; j0 += nwarp * 2) {
            int j = j0 + half;
            bool live = j < NP;
            float2* line = buf + j * LS;
            if (live) {
                int y0 = 2 * j, y1 = 2 * j + 1;
                for (int x = t; x < Nx; x += 16) {
                    float m0 = 0.f, m1 = 0.f;
                    for (int k = 0; k < K; k++) {
                        const float* Sk = S + (size_t)(k0 + k) * N;
                        float a = (k < 64) ? sA[k] : A[(size_t)(k0 + k) * C + c];
                        m0 += a * Sk[y0 * Nx + x];
                        if (y1 < Ny) m1 += a * Sk[y1 * Nx + x];
                    }
                    line[x] = make_float2(m0, m1);
                }
            }
            __syncwarp();
            fft_line16<128>(line, 1, t, tw, Nx, live);
            untangle_line(line, t, live);
        }
        __syncthreads();
        // ---- phase B: columns, x D
        column_phase<false>(buf, tw, D, NR);
        __syncthreads();
        // ---- phase C: inverse rows -> R; residual; forward rows of r
        for (int j0 = warp * 2; j0 < NP; j0 += nwarp * 2) {
            int j = j0 + half;
            bool live = j < NP;
            float2* line = buf + j * LS;
            retangle_line(line, t, live);
            fft_line16<128>(line, 1, t, tw, 128, live);
            if (live) {
                int y0 = 2 * j, y1 = 2 * j + 1;
                for (int x = t; x < Nx; x += 16) {
                    float2 z = line[x];                    // swapped: (im, re)
                    float r0 = z.y * inv, r1 = z.x * inv;
                    int p0 = y0 * Nx + x, p1 = y1 * Nx + x;
                    float w0 = wgt ? wgt[p0] : 1.f;
                    float d0 = w0 * (r0 - img[p0]);
                    loss += (double)d0 * (double)d0;
                    float o0 = w0 * d0, o1 = 0.f;
                    if (y1 < Ny) {
                        float w1 = wgt ? wgt[p1] : 1.f;
                        float d1 = w1 * (r1 - img[p1]);
                        loss += (double)d1 * (double)d1;
                        o1 = w1 * d1;
                    }
                    line[x] = make_float2(o0, o1);
                }
            }
            __syncwarp();
            fft_line16<128>(line, 1, t, tw, Nx, live);
            untangle_line(line, t, live);
        }
        __syncthreads();
        // ---- phase D: columns, x conj(D)
        column_phase<true>(buf, tw, D, NR);
        __syncthreads();
        // ---- phase E: inverse rows -> G
        for (int j0 = warp * 2; j0 < NP; j0 += nwarp * 2) {
            int j = j0 + half;
            bool live = j < NP;
            float2* line = buf + j * LS;
            retangle_line(line, t, live);
            fft_line16<128>(line, 1, t, tw, 128, live);
            if (live) {
                int y0 = 2 * j, y1 = 2 * j + 1;
                for (int x = t; x < Nx; x += 16) {
                    float2 z = line[x];
                    const float g0 = z.y * inv, g1 = (y1 < Ny) ? z.x * inv : 0.f;
                    const int p0 = y0 * Nx + x, p1 = (y1 < Ny) ? y1 * Nx + x : p0;
                    gout[p0] = g0;
                    if (y1 < Ny) gout[p1] = g1;
#pragma unroll
                    for (int k = 0; k < SCL_MAX_COMP; k++)
                        if (k < K) {
                            const float* Sk = S + (size_t)(k0 + k) * N;
                            accA[k] += g0 * Sk[p0] + g1 * Sk[p1];
                        }
                }
            }
        }
    }
    double tot = block_sum<double>(loss, red);
    if (threadIdx.x == 0) loss_part[b * C + c] = 0.5 * tot;

    // ---- SED gradient of this band: dA[k, c]
    {
#pragma unroll
        for (int k = 0; k < SCL_MAX_COMP; k++) {
            double v = warp_sum<double>((double)accA[k]);
            if (lane == 0) sred[warp][k] = v;
        }
        __syncthreads();
        if (threadIdx.x < K) {
            double v = 0.0;
            for (int w = 0; w < nwarp; w++) v += sred[w][threadIdx.x];
            dA[(size_t)(k0 + threadIdx.x) * C + c] = (float)v;
        }
    }
    // ---- rows r = c, c + C, ... of the Gram matrix S S^T (Blend._set_lipschitz, blend.py:205-217)
    for (int r = c; r < K; r += C) {
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
