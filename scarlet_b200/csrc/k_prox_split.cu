// Scene mode in stages: the constraint schedule of k_prox.cu cut where its parallelism changes.
//
// One CTA per component (scl_k_prox_wide) leaves the k-space symmetry of a frame-sized window -- two dense products
// with the Dirichlet operators, ny*nx*(ny+nx) multiply-adds, 250 M for a source near the middle of a 512x512 scene --
// to a single SM: measured 4.1 ms for the longest one, with the whole kernel waiting for it (profiles/README.md).
// Here the iteration's update is a chain of kernels over the same scratch slots (X, Z in HBM as before):
//   scl_k_wide_step            one CTA per component: gradient steps, max_pixel, centroid          (source.py:402-421)
//   scl_k_wide_source_center   one CTA per multi-component source: centre of the summed morphology  (source.py:605-627)
//   scl_k_wide_prep            one CTA per component: symmetry window, Dirichlet tables, Nyquist-row vectors
//   scl_k_wide_sym_rows/_cols  one CTA per 64x128 tile of a window: Z = X Tx,  X <- (X <= 0) ? 0 : (X + Ty Z)/2
//   scl_k_wide_finish          one CTA per component: mirror symmetry, sweep, positivity, normalisation, write-back
// The arithmetic of every stage is the monolithic kernel's (same routines, same thread partition where a reduction
// order depends on it); the two products sum over l in ascending order.
#include "prox_dev.cuh"

#define SCL_SYM_BM 64        // tile rows
#define SCL_SYM_BN 128       // tile columns
#define SCL_SYM_KC 32        // l values staged per round
#define SCL_SYM_THREADS 256  // 8 row groups of 8 rows x 32 column groups of 4 columns

__host__ __device__ inline int scl_sym_tiles(int Ny, int Nx) {
    return ((Ny + SCL_SYM_BM - 1) / SCL_SYM_BM) * ((Nx + SCL_SYM_BN - 1) / SCL_SYM_BN);
}

__device__ __forceinline__ float* scl_wide_x(const SclProxArgs& a, int slot) {
    return a.xscratch + (size_t)slot * a.xstride + ((a.Nx + 4) & ~3);
}
__device__ __forceinline__ float* scl_wide_z(const SclProxArgs& a, int slot) {
    return scl_wide_x(a, slot) + (((size_t)a.Ny * a.Nx + 3) & ~(size_t)3);
}

// ---- stage 1: gradient steps of every listed component, centre of those that own one
extern "C" __global__ void __launch_bounds__(SCL_PROX_WIDE, SCL_WIDE_MIN)
scl_k_wide_step(SclProxArgs a)
{
    const int k = a.comp_list[blockIdx.x];
    const int b = a.comp_blend[k];
    SclWideRec& r = a.wrec[blockIdx.x];
    if (!a.active[b]) {
        if (threadIdx.x == 0) r.k = -1;
        return;
    }
    const int Ny = a.Ny, Nx = a.Nx;
    float* X = scl_wide_x(a, blockIdx.x);
    __shared__ double red[32];
    __shared__ int sc[2];
    __shared__ double ss[2];
    __shared__ double sdA[SCL_PROX_WIDE / 32][SCL_MAX_BANDS];
    __shared__ float sAnew[SCL_MAX_BANDS], sAold[SCL_MAX_BANDS];
    __shared__ float sstep;
    const unsigned fl = a.cflags[k];
    const int it = a.it_override >= 0 ? a.it_override : a.it0[b] + a.step + 1;
    long long t0 = a.phase_clk ? clock64() : 0;
    scl_component_step<true>(a, k, b, fl, X, sdA, sAnew, sAold, &sstep);
    if (threadIdx.x == 0) {
        sc[0] = a.center[2 * k]; sc[1] = a.center[2 * k + 1];
        ss[0] = a.shift[2 * k]; ss[1] = a.shift[2 * k + 1];
    }
    __syncthreads();
    if (a.phase_clk && threadIdx.x == 0) { const long long t1 = clock64(); atomicAdd(a.phase_clk + 0, (unsigned long long)(t1 - t0)); t0 = t1; }
    if ((fl & SCL_F_CONSTRAINED) && !(fl & SCL_F_SHARED_CENTER)) {
        scl_max_pixel(X, Ny, Nx, sc[0], sc[1], sc);
        __syncthreads();
        if ((fl & SCL_F_SYMMETRIC) && it % 5 == 0) {
            const float* psf = a.cw + (size_t)a.cw_idx[b] * a.ph * a.pw;
            const int cy = sc[0], cx = sc[1];
            __syncthreads();
            scl_centroid(X, Ny, Nx, psf, a.ph, a.pw, cy, cx, sc, ss, red);
        }
    }
    __syncthreads();
    if (threadIdx.x < SCL_MAX_BANDS) { r.sAnew[threadIdx.x] = sAnew[threadIdx.x]; r.sAold[threadIdx.x] = sAold[threadIdx.x]; }
    if (threadIdx.x == 0) {
        r.k = k; r.mode = 0; r.cy = sc[0]; r.cx = sc[1]; r.dy = ss[0]; r.dx = ss[1]; r.step = sstep;
        a.slot_of[k] = blockIdx.x;
        if (a.phase_clk) atomicAdd(a.phase_clk + 1, (unsigned long long)(clock64() - t0));
    }
}

// ---- stage 2: MultiComponentSource.update, first half (source.py:605-627) on the stepped components of stage 1
// (scl_k_source_center recomputes the steps; here they are read back).  The summed morphology is built in the Z region
// of the first component's slot, which nothing uses before the symmetry kernels.
extern "C" __global__ void __launch_bounds__(SCL_PROX_WIDE, SCL_WIDE_MIN)
scl_k_wide_source_center(SclProxArgs a, const int* __restrict__ msrc)
{
    const int k0 = msrc[2 * blockIdx.x], n = msrc[2 * blockIdx.x + 1];
    const int b = a.comp_blend[k0];
    if (!a.active[b]) return;
    const int Ny = a.Ny, Nx = a.Nx, C = a.C;
    __shared__ double red[32];
    __shared__ int sc[2];
    __shared__ double ss[2];
    const int it = a.it_override >= 0 ? a.it_override : a.it0[b] + a.step + 1;
    const unsigned fl0 = a.cflags[k0];
    float* M = scl_wide_z(a, a.slot_of[k0]);
    // max_pixel looks at the 5x5 pixels around the previous centre and the centroid at a PSF-sized window around the
    // pixel it finds (measurement.py:3-29,32-94): the summed morphology is needed on that neighbourhood only, not on
    // the frame (the same products and sums, pixel by pixel; the full-frame pass cost 0.2 ms per iteration)
    const int rad = a.pw / 2 + 2;
    const int cy0 = a.center[2 * k0], cx0 = a.center[2 * k0 + 1];
    const int wy0 = max(cy0 - rad, 0), wy1 = min(cy0 + rad, Ny - 1), wx0 = max(cx0 - rad, 0), wx1 = min(cx0 + rad, Nx - 1);
    const int wnx = wx1 - wx0 + 1, wn = (wy1 - wy0 + 1) * wnx;
    for (int j = 0; j < n; j++) {
        const int slot = a.slot_of[k0 + j];
        const float* Xk = scl_wide_x(a, slot);
        const SclWideRec& rj = a.wrec[slot];
        float ssum = 0.f;                                    // c.sed.sum(), float32, band order
        for (int c = 0; c < C; c++) ssum = __fadd_rn(ssum, rj.sAnew[c]);
        for (int i = threadIdx.x; i < wn; i += blockDim.x) {
            const int y = wy0 + i / wnx, x = wx0 + i % wnx, p = y * Nx + x;
            const float t = __fmul_rn(Xk[p], ssum);
            M[p] = (j == 0) ? t : __fadd_rn(M[p], t);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        sc[0] = a.center[2 * k0]; sc[1] = a.center[2 * k0 + 1];
        ss[0] = a.shift[2 * k0]; ss[1] = a.shift[2 * k0 + 1];
    }
    __syncthreads();
    scl_max_pixel(M, Ny, Nx, sc[0], sc[1], sc);
    __syncthreads();
    if ((fl0 & SCL_F_SYMMETRIC) && it % 5 == 0) {
        const float* psf = a.cw + (size_t)a.cw_idx[b] * a.ph * a.pw;
        const int cy = sc[0], cx = sc[1];
        __syncthreads();
        scl_centroid(M, Ny, Nx, psf, a.ph, a.pw, cy, cx, sc, ss, red);
    }
    __syncthreads();
    if (threadIdx.x < n) {
        SclWideRec& rj = a.wrec[a.slot_of[k0 + threadIdx.x]];
        rj.cy = sc[0]; rj.cx = sc[1]; rj.dy = ss[0]; rj.dx = ss[1];
    }
}

// ---- stage 3: which symmetry the component takes, on which window, and the k-space operator's tables
// aux layout per slot: Ty [2ny-1] | Tx [2nx-1] | Hx [2nx-1] | bvec [nx] | vvec [nx]
extern "C" __global__ void __launch_bounds__(256)
scl_k_wide_prep(SclProxArgs a)
{
    SclWideRec& r = a.wrec[blockIdx.x];
    const int k = r.k;
    if (k < 0) return;
    const unsigned fl = a.cflags[k];
    // only the components that take the k-space symmetry (mirror pairs are taken by the finishing kernel); mode stays 0
    if (!(fl & SCL_F_CONSTRAINED) || !(fl & SCL_F_SYMMETRIC) || !(fl & SCL_F_KSPACE) || (fl & SCL_F_SCHEDULE)) return;
    const int Ny = a.Ny, Nx = a.Nx;
    const SclWindow w = scl_window(Ny, Nx, r.cy, r.cx);
    // update.symmetric discards the return value of the centred case (update.py:193-198)
    const int mode = w.centered ? 0 : 1;
    if (threadIdx.x == 0) { r.mode = mode; r.y0 = w.y0; r.x0 = w.x0; r.ny = w.ny; r.nx = w.nx; }
    if (mode != 1) return;
    const int ny = w.ny, nx = w.nx;
    const int Fy = a.nfl[2 * ny + 10], Fx = a.nfl_even[2 * nx + 10];
    float* Ty = a.waux + (size_t)blockIdx.x * a.waux_stride;
    float* Tx = Ty + (2 * ny - 1);
    float* Hx = Tx + (2 * nx - 1);
    float* bvec = Hx + (2 * nx - 1);
    float* vvec = bvec + nx;
    const double dy = r.dy, dx = r.dx;
    scl_dirichlet(Ty, nullptr, ny, Fy, 2.0 * dy, false);
    scl_dirichlet(Tx, Hx, nx, Fx, 2.0 * dx, true);
    const bool nyq = (Fy % 2 == 0);
    const int oy = (Fy - ny + 1) / 2 - Fy / 2;
    if (threadIdx.x == 0) { r.nyq = nyq; r.oy = oy; }
    if (!nyq) return;
    const float* Xw = scl_wide_x(a, blockIdx.x) + w.y0 * Nx + w.x0;
    // b[l] = sum_i (-1)^{p_i} Xw[i][l]
    for (int l = threadIdx.x; l < nx; l += blockDim.x) {
        float s = 0.f;
        // (same order of additions as the one-CTA kernel; unrolled so that eight loads are in flight: the column is
        // read from L2 by one thread and the loop was a chain of ny round trips)
#pragma unroll 8
        for (int i = 0; i < ny; i++) s += (((i + oy) & 1) ? -1.f : 1.f) * Xw[i * Nx + l];
        bvec[l] = s;
    }
    __syncthreads();
    double sphi, cphi;
    sincospi(2.0 * dy, &sphi, &cphi);
    const float cf = (float)(cphi / Fy), sf = (float)(sphi / Fy);
    for (int j = threadIdx.x; j < nx; j += blockDim.x) {
        float s1 = 0.f, s2 = 0.f;
#pragma unroll 8
        for (int l = 0; l < nx; l++) { s1 += Tx[j + l] * bvec[l]; s2 += Hx[j + l] * bvec[l]; }
        vvec[j] = cf * s1 + sf * s2;
    }
}

// ---- stage 4a: Z[i][j] = sum_l Xw[i][l] Tx[j + l] on one 64x128 tile.  grid (tiles of the frame, slots); a CTA whose
// slot takes no k-space symmetry, or whose tile lies outside the slot's window, returns at once.
// Per thread 8 rows x 4 columns; the l loop reads the rows of a staged l (stored l-major, so that a row pair is one
// 64-bit operand of FFMA2) and slides four table values along.
extern "C" __global__ void __launch_bounds__(SCL_SYM_THREADS)
scl_k_wide_sym_rows(SclProxArgs a)
{
    const SclWideRec& r = a.wrec[blockIdx.y];
    if (r.k < 0 || r.mode != 1) return;
    const int ny = r.ny, nx = r.nx, Nx = a.Nx;
    const int ntx = (nx + SCL_SYM_BN - 1) / SCL_SYM_BN, nty = (ny + SCL_SYM_BM - 1) / SCL_SYM_BM;
    if ((int)blockIdx.x >= ntx * nty) return;
    const int I0 = (blockIdx.x / ntx) * SCL_SYM_BM, J0 = (blockIdx.x % ntx) * SCL_SYM_BN;
    const float* Xw = scl_wide_x(a, blockIdx.y) + r.y0 * Nx + r.x0;
    float* Z = scl_wide_z(a, blockIdx.y);
    const float* Tx = a.waux + (size_t)blockIdx.y * a.waux_stride + (2 * ny - 1);
    const int Zs = (nx + 3) & ~3;
    constexpr int XS = SCL_SYM_BM + 4;                       // row stride of the staged l-major block (16-byte rows)
    __shared__ __align__(16) float Xs[SCL_SYM_KC * XS];
    __shared__ __align__(16) float Ts[SCL_SYM_BN + SCL_SYM_KC + 8];
    const int ty = threadIdx.x >> 5, tx = threadIdx.x & 31;
    unsigned long long acc2[4][4];
#pragma unroll
    for (int q = 0; q < 4; q++)
#pragma unroll
        for (int c = 0; c < 4; c++) acc2[q][c] = 0ull;
    for (int l0 = 0; l0 < nx; l0 += SCL_SYM_KC) {
        __syncthreads();
        // rows I0.., columns l0..l0+31 of the window, transposed; zeros outside the window
#pragma unroll
        for (int q = 0; q < (SCL_SYM_BM * SCL_SYM_KC) / SCL_SYM_THREADS; q++) {
            const int idx = threadIdx.x + q * SCL_SYM_THREADS, rr = idx >> 5, cc = idx & 31;
            const int i = I0 + rr, l = l0 + cc;
            Xs[cc * XS + rr] = (i < ny && l < nx) ? Xw[i * Nx + l] : 0.f;
        }
        for (int i = threadIdx.x; i < SCL_SYM_BN + SCL_SYM_KC + 8; i += SCL_SYM_THREADS) {
            const int m = J0 + l0 + i;
            Ts[i] = (m < 2 * nx - 1) ? Tx[m] : 0.f;
        }
        __syncthreads();
        float4 tl = *reinterpret_cast<const float4*>(Ts + 4 * tx);
#pragma unroll
        for (int lg = 0; lg < SCL_SYM_KC; lg += 4) {
            const float4 th = *reinterpret_cast<const float4*>(Ts + 4 * tx + lg + 4);
            const float t[8] = {tl.x, tl.y, tl.z, tl.w, th.x, th.y, th.z, th.w};
#pragma unroll
            for (int ll = 0; ll < 4; ll++) {
                const ulonglong2 xa = *reinterpret_cast<const ulonglong2*>(Xs + (lg + ll) * XS + 8 * ty);
                const ulonglong2 xb = *reinterpret_cast<const ulonglong2*>(Xs + (lg + ll) * XS + 8 * ty + 4);
                const unsigned long long xp[4] = {xa.x, xa.y, xb.x, xb.y};
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const unsigned long long tt = scl_pk(t[ll + c], t[ll + c]);
#pragma unroll
                    for (int q = 0; q < 4; q++) acc2[q][c] = scl_fma2(xp[q], tt, acc2[q][c]);
                }
            }
            tl = th;
        }
    }
    const int i0 = I0 + 8 * ty, j0 = J0 + 4 * tx;
    if (j0 < Zs) {
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const float2 c0 = scl_upk(acc2[q][0]), c1 = scl_upk(acc2[q][1]), c2 = scl_upk(acc2[q][2]), c3 = scl_upk(acc2[q][3]);
            if (i0 + 2 * q < ny)
                *reinterpret_cast<float4*>(Z + (size_t)(i0 + 2 * q) * Zs + j0) = make_float4(c0.x, c1.x, c2.x, c3.x);
            if (i0 + 2 * q + 1 < ny)
                *reinterpret_cast<float4*>(Z + (size_t)(i0 + 2 * q + 1) * Zs + j0) = make_float4(c0.y, c1.y, c2.y, c3.y);
        }
    }
}

// ---- stage 4b: Y[i][j] = sum_l Ty[i + l] Z[l][j] (+ the Nyquist-row term);  X <- (X <= 0) ? 0 : (X + Y)/2
extern "C" __global__ void __launch_bounds__(SCL_SYM_THREADS)
scl_k_wide_sym_cols(SclProxArgs a)
{
    const SclWideRec& r = a.wrec[blockIdx.y];
    if (r.k < 0 || r.mode != 1) return;
    const int ny = r.ny, nx = r.nx, Nx = a.Nx;
    const int ntx = (nx + SCL_SYM_BN - 1) / SCL_SYM_BN, nty = (ny + SCL_SYM_BM - 1) / SCL_SYM_BM;
    if ((int)blockIdx.x >= ntx * nty) return;
    const int I0 = (blockIdx.x / ntx) * SCL_SYM_BM, J0 = (blockIdx.x % ntx) * SCL_SYM_BN;
    float* Xw = scl_wide_x(a, blockIdx.y) + r.y0 * Nx + r.x0;
    const float* Z = scl_wide_z(a, blockIdx.y);
    const float* Ty = a.waux + (size_t)blockIdx.y * a.waux_stride;
    const float* vvec = Ty + (2 * ny - 1) + 2 * (2 * nx - 1) + nx;
    const int Zs = (nx + 3) & ~3;
    __shared__ __align__(16) float Zc[SCL_SYM_KC * SCL_SYM_BN];
    __shared__ __align__(16) float Ts[SCL_SYM_BM + SCL_SYM_KC + 8];
    const int ty = threadIdx.x >> 5, tx = threadIdx.x & 31;
    unsigned long long acc2[8][2];
#pragma unroll
    for (int rr = 0; rr < 8; rr++) { acc2[rr][0] = 0ull; acc2[rr][1] = 0ull; }
    for (int l0 = 0; l0 < ny; l0 += SCL_SYM_KC) {
        __syncthreads();
        // rows l0..l0+31 of Z, columns J0..J0+127 (16-byte loads: Zs and J0 are multiples of 4); zeros outside
#pragma unroll
        for (int q = 0; q < (SCL_SYM_KC * SCL_SYM_BN / 4) / SCL_SYM_THREADS; q++) {
            const int idx = threadIdx.x + q * SCL_SYM_THREADS, rr = idx >> 5, c4 = (idx & 31) * 4;
            const int l = l0 + rr, j = J0 + c4;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (l < ny && j < Zs) v = *reinterpret_cast<const float4*>(Z + (size_t)l * Zs + j);
            *reinterpret_cast<float4*>(Zc + rr * SCL_SYM_BN + c4) = v;
        }
        for (int i = threadIdx.x; i < SCL_SYM_BM + SCL_SYM_KC + 8; i += SCL_SYM_THREADS) {
            const int m = I0 + l0 + i;
            Ts[i] = (m < 2 * ny - 1) ? Ty[m] : 0.f;
        }
        __syncthreads();
        // t[rr + ll] = Ty[i0 + rr + l]: eleven values per four l, one 16-byte load slides in per group
        float4 ta = *reinterpret_cast<const float4*>(Ts + 8 * ty);
        float4 tb = *reinterpret_cast<const float4*>(Ts + 8 * ty + 4);
#pragma unroll
        for (int lg = 0; lg < SCL_SYM_KC; lg += 4) {
            const float4 tc = *reinterpret_cast<const float4*>(Ts + 8 * ty + lg + 8);
            const float t[12] = {ta.x, ta.y, ta.z, ta.w, tb.x, tb.y, tb.z, tb.w, tc.x, tc.y, tc.z, tc.w};
#pragma unroll
            for (int ll = 0; ll < 4; ll++) {
                const ulonglong2 zv = *reinterpret_cast<const ulonglong2*>(Zc + (lg + ll) * SCL_SYM_BN + 4 * tx);
#pragma unroll
                for (int rr = 0; rr < 8; rr++) {
                    const unsigned long long tt = scl_pk(t[rr + ll], t[rr + ll]);
                    acc2[rr][0] = scl_fma2(tt, zv.x, acc2[rr][0]);
                    acc2[rr][1] = scl_fma2(tt, zv.y, acc2[rr][1]);
                }
            }
            ta = tb; tb = tc;
        }
    }
    const int i0 = I0 + 8 * ty, j0 = J0 + 4 * tx;
    const bool nyq = r.nyq;
    const int oy = r.oy;
#pragma unroll
    for (int rr = 0; rr < 8; rr++) {
        const float2 p = scl_upk(acc2[rr][0]), q = scl_upk(acc2[rr][1]);
        const float yv[4] = {p.x, p.y, q.x, q.y};
        const int i = i0 + rr;
        if (i >= ny) continue;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const int j = j0 + c;
            if (j < nx) {
                float y = yv[c];
                if (nyq) y += (((i + oy) & 1) ? -1.f : 1.f) * vvec[j];
                const float x = Xw[i * Nx + j];
                Xw[i * Nx + j] = (x <= 0.f) ? 0.f : 0.5f * (x + y);
            }
        }
    }
}

// ---- stage 5: what is left of the schedule (mirror symmetry, sweep, positivity, normalisation, write-back, convergence)
extern "C" __global__ void __launch_bounds__(SCL_PROX_WIDE, SCL_WIDE_MIN)
scl_k_wide_finish(SclProxArgs a) { scl_prox_body<SCL_PROX_WIDE, 2>(a); }
