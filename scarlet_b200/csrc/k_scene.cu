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
//   else    : src = r [B][P][N];  dst = r (star) d (CONJ: correlation, conj spectrum) or r (*) d  [B][P][N]
template <bool RESIDUAL, bool CONJ>
__global__ void __launch_bounds__(SCL_THREADS)
scl_k_scene_conv(const float* __restrict__ src, const float2* __restrict__ Dhat, const int* __restrict__ ker_idx,
                 const float* __restrict__ images, const float* __restrict__ weights, const int* __restrict__ active,
                 float* __restrict__ dst, double* __restrict__ loss_tile,
                 int Cm, SclPlanes pl, int Ny, int Nx, int h, int tiles_x, int dst_planes, int loss_stride)
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
    float* out = dst + ((size_t)b * dst_planes + plane) * N;      // dst_planes >= P (G also holds low-resolution planes)
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
        column_phase<CONJ>(buf, tw, D, F);
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
        if (threadIdx.x == 0) loss_tile[(size_t)b * loss_stride + plane * ntile + tile] = 0.5 * tot;
    }
}
#define SCN_CONV_ARGS const float*, const float2*, const int*, const float*, const float*, const int*, float*, double*, \
    int, SclPlanes, int, int, int, int, int, int
template __global__ void scl_k_scene_conv<true, false>(SCN_CONV_ARGS);
template __global__ void scl_k_scene_conv<false, true>(SCN_CONV_ARGS);
template __global__ void scl_k_scene_conv<false, false>(SCN_CONV_ARGS);

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

// ---------------------------------------------------------------- exact Lipschitz constants for any K
// Blend._set_lipschitz (blend.py:203-217): L_sed = lambda_max(S S^T), L_morph = lambda_max(A^T A).
// Gram matrix: gram[b][i][j] = sum_p S[k0+i][p] S[k0+j][p], Kp x Kp (Kp = padded component count, zero beyond K),
// float32 products accumulated per CTA over a chunk of pixels, chunks combined with float64 atomics.
// grid: (tiles * tiles, chunks, B), 64x64 output tile per CTA, 256 threads (4x4 outputs each).
#define SCN_GT 64
#define SCN_GP 32
extern "C" __global__ void __launch_bounds__(256)
scl_k_scene_gram(const float* __restrict__ S, const int* __restrict__ comp_off, const int* __restrict__ active,
                 float* __restrict__ part, int Kp, int N, int chunk)      // part [B][chunks][Kp][Kp]
{
    // pixel-major tiles [pixel][component]; a thread owns the 4x4 outputs (4 ty + u, 4 tx + v) and reads its four
    // components of a pixel as ONE 16-byte shared load per operand (2 LDS.128 feed 16 FMAs).  The 4-float groups of a
    // pixel row are XOR-swizzled with the pixel index so that the transposed stores spread over the banks.
    __shared__ __align__(16) float sa[2][SCN_GP][SCN_GT], sb[2][SCN_GP][SCN_GT];      // double buffered (cp.async)
    const int b = blockIdx.z;
    if (!active[b]) return;
    const int k0 = comp_off[b], K = comp_off[b + 1] - k0;
    const int nt = (Kp + SCN_GT - 1) / SCN_GT, ti = blockIdx.x / nt, tj = blockIdx.x % nt;
    if (tj < ti || ti * SCN_GT >= K || tj * SCN_GT >= K) return;          // upper triangle, mirrored on write
    const int p_lo = blockIdx.y * chunk, p_hi = min(N, p_lo + chunk);
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
    float acc[4][4];
#pragma unroll
    for (int u = 0; u < 4; u++)
#pragma unroll
        for (int v = 0; v < 4; v++) acc[u][v] = 0.f;
    // one pixel block (SCN_GP pixels of the 2 x 64 components) into buffer `buf`: 4-byte async copies, zero filled
    // past K or past the chunk
    auto stage = [&](int buf, int p0) {
        for (int i = threadIdx.x; i < SCN_GT * SCN_GP; i += blockDim.x) {
            const int row = i / SCN_GP, col = i % SCN_GP, p = p0 + col;
            const int ka = ti * SCN_GT + row, kb = tj * SCN_GT + row;
            const int phys = (((row >> 2) ^ (col & 15)) << 2) | (row & 3);
            const bool pin = p < p_hi;
            const float* ga = S + (size_t)(k0 + (ka < K ? ka : 0)) * N + (pin ? p : p_lo);
            const float* gb = S + (size_t)(k0 + (kb < K ? kb : 0)) * N + (pin ? p : p_lo);
            const unsigned da = (unsigned)__cvta_generic_to_shared(&sa[buf][col][phys]);
            const unsigned db = (unsigned)__cvta_generic_to_shared(&sb[buf][col][phys]);
            const int na = (ka < K && pin) ? 4 : 0, nb = (kb < K && pin) ? 4 : 0;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" :: "r"(da), "l"(ga), "r"(na));
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" :: "r"(db), "l"(gb), "r"(nb));
        }
        asm volatile("cp.async.commit_group;\n" ::);
    };
    stage(0, p_lo);
    int buf = 0;
    for (int p0 = p_lo; p0 < p_hi; p0 += SCN_GP, buf ^= 1) {
        if (p0 + SCN_GP < p_hi) {
            stage(buf ^ 1, p0 + SCN_GP);
            asm volatile("cp.async.wait_group 1;\n" ::);
        } else {
            asm volatile("cp.async.wait_group 0;\n" ::);
        }
        __syncthreads();
#pragma unroll 8
        for (int q = 0; q < SCN_GP; q++) {
            const float4 a4 = *reinterpret_cast<const float4*>(&sa[buf][q][(ty ^ (q & 15)) << 2]);
            const float4 b4 = *reinterpret_cast<const float4*>(&sb[buf][q][(tx ^ (q & 15)) << 2]);
            const float a[4] = {a4.x, a4.y, a4.z, a4.w}, bb[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
            for (int u = 0; u < 4; u++)
#pragma unroll
                for (int v = 0; v < 4; v++) acc[u][v] += a[u] * bb[v];
        }
        __syncthreads();
    }
    // per-chunk partial sums, combined in chunk order by scl_k_scene_gram_reduce (5 M float64 atomics cost more than
    // the products themselves)
    float* g = part + ((size_t)b * gridDim.y + blockIdx.y) * Kp * Kp;
#pragma unroll
    for (int u = 0; u < 4; u++)
#pragma unroll
        for (int v = 0; v < 4; v++) {
            const int i = ti * SCN_GT + 4 * ty + u, j = tj * SCN_GT + 4 * tx + v;
            if (i < K && j < K) g[(size_t)i * Kp + j] = acc[u][v];
        }
}

// gram[b][i][j] = gram[b][j][i] = sum over chunks of part[b][chunk][i][j], j >= i.  grid: (K rows, B)
extern "C" __global__ void __launch_bounds__(128)
scl_k_scene_gram_reduce(const float* __restrict__ part, const int* __restrict__ comp_off,
                        const int* __restrict__ active, double* __restrict__ gram, int Kp, int chunks)
{
    const int b = blockIdx.y, i = blockIdx.x;
    if (!active[b]) return;
    const int K = comp_off[b + 1] - comp_off[b];
    if (i >= K) return;
    for (int j = i + threadIdx.x; j < K; j += blockDim.x) {
        double v = 0.0;
        for (int c = 0; c < chunks; c++) v += (double)part[(((size_t)b * chunks + c) * Kp + i) * Kp + j];
        gram[((size_t)b * Kp + i) * Kp + j] = v;
        gram[((size_t)b * Kp + j) * Kp + i] = v;
    }
}

// f32 working copy of the Gram matrix, scaled by 1/trace.  grid: (Kp, B)
extern "C" __global__ void scl_k_scene_eig_init(const double* __restrict__ gram, float* __restrict__ M,
                                                double* __restrict__ fro, int Kp)
{
    __shared__ double red[32];
    const double* g = gram + (size_t)blockIdx.y * Kp * Kp;
    double tr = 0.0;
    for (int i = threadIdx.x; i < Kp; i += blockDim.x) tr += g[(size_t)i * Kp + i];
    tr = block_sum<double>(tr, red);
    const double s = tr > 0.0 ? 1.0 / tr : 0.0;
    float* m = M + ((size_t)blockIdx.y * Kp + blockIdx.x) * Kp;
    double f2 = 0.0;
    for (int j = threadIdx.x; j < Kp; j += blockDim.x) {
        const float v = (float)(g[(size_t)blockIdx.x * Kp + j] * s);
        m[j] = v;
        f2 += (double)v * (double)v;
    }
    f2 = block_sum<double>(f2, red);
    if (threadIdx.x == 0 && f2 != 0.0) atomicAdd(fro + blockIdx.y, f2);      // trace(M M) of the first squaring
}

// fro[b] = sum_ij M_ij^2 = trace(M M) for symmetric M.  grid: B
extern "C" __global__ void __launch_bounds__(1024)
scl_k_scene_frob(const float* __restrict__ M, double* __restrict__ fro, int Kp)
{
    __shared__ double red[32];
    const float* m = M + (size_t)blockIdx.x * Kp * Kp;
    double acc = 0.0;
    for (int i = threadIdx.x; i < Kp * Kp; i += blockDim.x) { const double v = m[i]; acc += v * v; }
    acc = block_sum<double>(acc, red);
    if (threadIdx.x == 0) fro[blockIdx.x] = acc;
}

// Normalised squaring out = in * in / trace(in * in): the top eigenvector's share of every column doubles its
// exponent per call (K2 does the same in registers for K <= 8).  grid: (Kp/32, Kp/32, B); block (32, 8)
extern "C" __global__ void __launch_bounds__(256)
scl_k_scene_square(const float* __restrict__ in, float* __restrict__ out, const double* __restrict__ fro,
                   double* __restrict__ fro_next, int Kp)
{
    __shared__ float ta[32][33], tb[32][33];
    __shared__ double red[32];
    const int b = blockIdx.z;
    const float* A = in + (size_t)b * Kp * Kp;
    const int i0 = blockIdx.y * 32, j0 = blockIdx.x * 32, tx = threadIdx.x, ty = threadIdx.y;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    // the next pair of tiles is fetched into registers while this one is multiplied: the kernel is 11 dependent steps
    // of a matrix that sits in L2, and paid one L2 round trip per step (21 us a launch, 16 launches per iteration)
    float ra[4], rb[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
        ra[u] = A[(size_t)(i0 + ty + 8 * u) * Kp + tx];
        rb[u] = A[(size_t)(ty + 8 * u) * Kp + j0 + tx];
    }
    for (int l0 = 0; l0 < Kp; l0 += 32) {
#pragma unroll
        for (int u = 0; u < 4; u++) { ta[ty + 8 * u][tx] = ra[u]; tb[ty + 8 * u][tx] = rb[u]; }
        __syncthreads();
        if (l0 + 32 < Kp) {
#pragma unroll
            for (int u = 0; u < 4; u++) {
                ra[u] = A[(size_t)(i0 + ty + 8 * u) * Kp + l0 + 32 + tx];
                rb[u] = A[(size_t)(l0 + 32 + ty + 8 * u) * Kp + j0 + tx];
            }
        }
#pragma unroll
        for (int l = 0; l < 32; l++) {
            const float bv = tb[l][tx];
#pragma unroll
            for (int u = 0; u < 4; u++) acc[u] += ta[ty + 8 * u][l] * bv;
        }
        __syncthreads();
    }
    const float s = fro[b] > 0.0 ? (float)(1.0 / fro[b]) : 0.f;
#pragma unroll
    double f2 = 0.0;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const float v = acc[u] * s;
        out[(size_t)b * Kp * Kp + (size_t)(i0 + ty + 8 * u) * Kp + j0 + tx] = v;
        f2 += (double)v * (double)v;
    }
    // trace(out out) for the next squaring, folded in (a separate one-CTA reduction cost 59 us per squaring)
    f2 = warp_sum<double>(f2);
    if (tx == 0) red[ty] = f2;
    __syncthreads();
    if (tx == 0 && ty == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; w++) t += red[w];
        if (t != 0.0) atomicAdd(fro_next + b, t);
    }
}

// Rayleigh quotient of the heaviest column of the squared matrix against the float64 Gram matrix -> lambda_max(S S^T);
// lambda_max(A^T A) (C x C, float64 squarings by one warp); loss sum; lips.  grid: B, block: a power of two <= 1024
extern "C" __global__ void __launch_bounds__(1024)
scl_k_scene_eig_finish(const float* __restrict__ M, const double* __restrict__ gram, const float* __restrict__ A,
                       const int* __restrict__ comp_off, const double* __restrict__ loss_tile,
                       const int* __restrict__ active, double* __restrict__ logL, float* __restrict__ lips,
                       int Kp, int C, int n_loss, int step, int logL_cap, int n_obs)
{
    __shared__ double red[32];
    __shared__ int sbest;
    __shared__ double sLS;
    const int b = blockIdx.x;
    if (!active[b]) return;
    const int k0 = comp_off[b], K = comp_off[b + 1] - k0;
    const float* m = M + (size_t)b * Kp * Kp;
    const double* g = gram + (size_t)b * Kp * Kp;
    {
        // heaviest column: largest diagonal entry, lowest index on ties (block-wide arg-max)
        __shared__ float sval[1024];
        __shared__ int sidx[1024];
        float bv = -1.f; int bi = 0;
        for (int i = threadIdx.x; i < K; i += blockDim.x) {
            const float v = m[(size_t)i * Kp + i];
            if (v > bv) { bv = v; bi = i; }
        }
        sval[threadIdx.x] = bv; sidx[threadIdx.x] = bi;
        __syncthreads();
        for (int o = blockDim.x >> 1; o > 0; o >>= 1) {
            if (threadIdx.x < o) {
                const float v = sval[threadIdx.x + o]; const int i = sidx[threadIdx.x + o];
                if (v > sval[threadIdx.x] || (v == sval[threadIdx.x] && i < sidx[threadIdx.x])) {
                    sval[threadIdx.x] = v; sidx[threadIdx.x] = i;
                }
            }
            __syncthreads();
        }
        if (threadIdx.x == 0) sbest = sidx[0];
    }
    // A^T A: one (i, j) entry per thread of the first 64
    __shared__ double sATA[SCL_MAX_BANDS][SCL_MAX_BANDS];
    if (threadIdx.x < SCL_MAX_BANDS * SCL_MAX_BANDS) {
        const int i = threadIdx.x / SCL_MAX_BANDS, j = threadIdx.x % SCL_MAX_BANDS;
        double s = 0.0;
        if (i < C && j < C)
            for (int k = 0; k < K; k++) s += (double)A[(size_t)(k0 + k) * C + i] * (double)A[(size_t)(k0 + k) * C + j];
        sATA[i][j] = s;
    }
    __syncthreads();
    if ((threadIdx.x >> 5) == 1) {
        // lambda_max(A^T A): C x C, normalised squarings in float64, Rayleigh quotient of the heaviest column.  One
        // warp, an entry or two per lane (the serial version by one thread cost 0.1 ms of dependent DFMAs per
        // iteration); every sum runs in the order the serial loops had.
        __shared__ double sT[SCL_MAX_BANDS][SCL_MAX_BANDS], sQ[SCL_MAX_BANDS][SCL_MAX_BANDS];
        const int lane = threadIdx.x & 31;
        for (int e = lane; e < SCL_MAX_BANDS * SCL_MAX_BANDS; e += 32)
            sT[e / SCL_MAX_BANDS][e % SCL_MAX_BANDS] = sATA[e / SCL_MAX_BANDS][e % SCL_MAX_BANDS];
        for (int it = 0; it < 24; it++) {
            __syncwarp();
            for (int e = lane; e < SCL_MAX_BANDS * SCL_MAX_BANDS; e += 32) {
                const int i = e / SCL_MAX_BANDS, j = e % SCL_MAX_BANDS;
                if (i < C && j < C) {
                    double s = 0.0;
                    for (int l = 0; l < C; l++) s += sT[i][l] * sT[l][j];
                    sQ[i][j] = s;
                }
            }
            __syncwarp();
            double tr = 0.0;
            for (int i = 0; i < C; i++) tr += sQ[i][i];
            const double sc = tr > 0.0 ? 1.0 / tr : 0.0;
            for (int e = lane; e < SCL_MAX_BANDS * SCL_MAX_BANDS; e += 32) {
                const int i = e / SCL_MAX_BANDS, j = e % SCL_MAX_BANDS;
                if (i < C && j < C) sT[i][j] = sQ[i][j] * sc;
            }
        }
        __syncwarp();
        if (lane == 0) {
            int bc = 0;
            for (int i = 1; i < C; i++) if (sT[i][i] > sT[bc][bc]) bc = i;
            double num = 0.0, den = 0.0;
            for (int i = 0; i < C; i++) {
                double w = 0.0;
                for (int j = 0; j < C; j++) w += sATA[i][j] * sT[j][bc];
                num += sT[i][bc] * w; den += sT[i][bc] * sT[i][bc];
            }
            sLS = den > 0.0 ? num / den : 0.0;
        }
    }
    __syncthreads();
    const int best = sbest;
    double num = 0.0, den = 0.0;
    for (int i = threadIdx.x; i < K; i += blockDim.x) {
        double w = 0.0;
        // (the Gram matrix is symmetric bit for bit: read column i, consecutive threads on consecutive addresses)
#pragma unroll 8
        for (int j = 0; j < K; j++) w += g[(size_t)j * Kp + i] * (double)m[(size_t)j * Kp + best];
        const double vi = (double)m[(size_t)i * Kp + best];
        num += vi * w; den += vi * vi;
    }
    num = block_sum<double>(num, red);
    den = block_sum<double>(den, red);
    // loss of the iteration: the tile losses added in (plane, tile) order by one thread, fetched by all of them
    __shared__ double sloss[512];
    double total = 0.0;
    for (int i0 = 0; i0 < n_loss; i0 += 512) {
        __syncthreads();
        for (int i = threadIdx.x; i < 512 && i0 + i < n_loss; i += blockDim.x) sloss[i] = loss_tile[(size_t)b * n_loss + i0 + i];
        __syncthreads();
        if (threadIdx.x == 0)
            for (int i = 0; i < 512 && i0 + i < n_loss; i++) total += sloss[i];
    }
    if (threadIdx.x == 0) {
        if (step < logL_cap) logL[(size_t)b * logL_cap + step] = total;
        const double LA = den > 0.0 ? num / den : 0.0;
        lips[2 * b] = (float)(LA * n_obs);
        lips[2 * b + 1] = (float)(sLS * n_obs);
    }
}

// ---- exact L through the components' footprints ------------------------------------------------------------------
// After positivity a morphology is exactly zero outside the region the monotonic sweep left positive, so
// (S S^T)[k][l] only needs the pixels where BOTH footprints are non-zero: the intersection of their bounding boxes.
// On the config-4 scene the boxes cover ~9 % of the frame on average and a third of the pairs do not overlap at all;
// the same products as the dense contraction, the rest are exact zeros.
//
// box[k] = (y0, y1, x0, x1) inclusive bounds of the non-zero pixels of S_k, y1 < y0 when there is none.  grid: Ktot
extern "C" __global__ void __launch_bounds__(256)
scl_k_scene_bbox(const float* __restrict__ S, const int* __restrict__ comp_blend, const int* __restrict__ active,
                 int Ny, int Nx, int* __restrict__ box)
{
    __shared__ int red[4][8];
    const int k = blockIdx.x, N = Ny * Nx;
    if (!active[comp_blend[k]]) return;
    const float* s = S + (size_t)k * N;
    int y0 = Ny, y1 = -1, x0 = Nx, x1 = -1;
    if ((Nx & 3) == 0) {
        const float4* s4 = reinterpret_cast<const float4*>(s);
        for (int i = threadIdx.x; i < (N >> 2); i += blockDim.x) {
            const float4 v = s4[i];
            if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) {
                const int p = i << 2, y = p / Nx, x = p - y * Nx;
                y0 = min(y0, y); y1 = max(y1, y);
                if (v.x != 0.f) { x0 = min(x0, x); x1 = max(x1, x); }
                if (v.y != 0.f) { x0 = min(x0, x + 1); x1 = max(x1, x + 1); }
                if (v.z != 0.f) { x0 = min(x0, x + 2); x1 = max(x1, x + 2); }
                if (v.w != 0.f) { x0 = min(x0, x + 3); x1 = max(x1, x + 3); }
            }
        }
    } else {
        for (int p = threadIdx.x; p < N; p += blockDim.x)
            if (s[p] != 0.f) {
                const int y = p / Nx, x = p - y * Nx;
                y0 = min(y0, y); y1 = max(y1, y); x0 = min(x0, x); x1 = max(x1, x);
            }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        y0 = min(y0, __shfl_xor_sync(0xffffffffu, y0, o)); y1 = max(y1, __shfl_xor_sync(0xffffffffu, y1, o));
        x0 = min(x0, __shfl_xor_sync(0xffffffffu, x0, o)); x1 = max(x1, __shfl_xor_sync(0xffffffffu, x1, o));
    }
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { red[0][w] = y0; red[1][w] = y1; red[2][w] = x0; red[3][w] = x1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < (int)(blockDim.x >> 5); i++) {
            y0 = min(y0, red[0][i]); y1 = max(y1, red[1][i]); x0 = min(x0, red[2][i]); x1 = max(x1, red[3][i]);
        }
        box[4 * k] = y0; box[4 * k + 1] = y1; box[4 * k + 2] = x0; box[4 * k + 3] = x1;
    }
}

#ifndef SCL_GRAM_SPLIT
#define SCL_GRAM_SPLIT 128
#endif
// gram[b][i][j] = gram[b][j][i] = sum over box_i ^ box_j of S_i S_j for j = i + blockIdx.y, i + blockIdx.y + gridDim.y, ...
// grid: (Kmax, SPLIT, B); one warp per row of the intersection, lanes along x; float products and per-thread float
// sums, combined in a fixed order in float64.
extern "C" __global__ void __launch_bounds__(256)
scl_k_scene_gram_boxes(const float* __restrict__ S, const int* __restrict__ comp_off, const int* __restrict__ active,
                       const int* __restrict__ box, double* __restrict__ gram, int Kp, int Ny, int Nx)
{
    __shared__ double red[32];
    const int b = blockIdx.z, i = blockIdx.x;
    if (!active[b]) return;
    const int k0 = comp_off[b], K = comp_off[b + 1] - k0;
    if (i >= K) return;
    const int N = Ny * Nx;
    const int iy0 = box[4 * (k0 + i)], iy1 = box[4 * (k0 + i) + 1], ix0 = box[4 * (k0 + i) + 2], ix1 = box[4 * (k0 + i) + 3];
    const float* si = S + (size_t)(k0 + i) * N;
    double* g = gram + (size_t)b * Kp * Kp;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int j = i + blockIdx.y; j < K; j += gridDim.y) {
        const int y0 = max(iy0, box[4 * (k0 + j)]), y1 = min(iy1, box[4 * (k0 + j) + 1]);
        const int x0 = max(ix0, box[4 * (k0 + j) + 2]), x1 = min(ix1, box[4 * (k0 + j) + 3]);
        double v = 0.0;
        if (y0 <= y1 && x0 <= x1) {                      // block-uniform
            const float* sj = S + (size_t)(k0 + j) * N;
            float acc = 0.f;
            for (int y = y0 + warp; y <= y1; y += nw) {
                const float* ri = si + (size_t)y * Nx;
                const float* rj = sj + (size_t)y * Nx;
                for (int x = x0 + lane; x <= x1; x += 32) acc = fmaf(ri[x], rj[x], acc);
            }
            v = block_sum<double>((double)acc, red);
        }
        if (threadIdx.x == 0) { g[(size_t)i * Kp + j] = v; g[(size_t)j * Kp + i] = v; }
    }
}
