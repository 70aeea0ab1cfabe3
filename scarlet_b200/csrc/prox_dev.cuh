// Device building blocks of the constraint chain run on one morphology image
// held in shared memory (row-major [Ny][Nx] floats) by one CTA.
//
// Reference (file:line):
//   max_pixel                 measurement.py:3-29
//   psf_weighted_centroid     measurement.py:32-94
//   uncentered_operator       operator.py:175-228 (window geometry)
//   prox_kspace_symmetry      operator.py:253-288, interpolation.mk_shifter interpolation.py:302-340
//   prox_soft_symmetry        operator.py:242-251;  prox_sdss_symmetry operator.py:231-239
//   prox_weighted_monotonic   operators_pybind11.cc:27-50 with weights of operator.py:540-621
#pragma once
#include "common.cuh"

struct SclWindow { int y0, x0, ny, nx; bool centered; };

// operator.py:197-220: largest window in which (py,px) is the //2 centre
__device__ __forceinline__ SclWindow scl_window(int Ny, int Nx, int py, int px) {
    SclWindow w;
    int cy = Ny / 2, cx = Nx / 2;
    w.centered = (py == cy && px == cx);
    int dy = 2 * (py - cy) + ((Ny & 1) ? 0 : 1);
    int dx = 2 * (px - cx) + ((Nx & 1) ? 0 : 1);
    if (w.centered) { dy = 0; dx = 0; }
    if (dy < 0) { w.y0 = 0; w.ny = Ny + dy; } else { w.y0 = dy; w.ny = Ny - dy; }
    if (dx < 0) { w.x0 = 0; w.nx = Nx + dx; } else { w.x0 = dx; w.nx = Nx - dx; }
    return w;
}

// ---------------------------------------------------------------- max_pixel
// first row-major argmax of the 5x5 window around (cy,cx); executed by warp 0,
// result broadcast through `out` (shared int[2]); caller syncs.
__device__ __forceinline__ void scl_max_pixel(const float* X, int Ny, int Nx, int cy, int cx, int* out) {
    if (threadIdx.x < 32) {
        int l = threadIdx.x;
        int ys = max(cy - 2, 0), ye = min(cy + 3, Ny), xs = max(cx - 2, 0), xe = min(cx + 3, Nx);
        int wy = ye - ys, wx = xe - xs;
        float v = -INFINITY;
        int idx = 0x7fffffff;
        if (l < wy * wx) {
            int y = ys + l / wx, x = xs + l % wx;
            v = X[y * Nx + x];
            idx = l;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float v2 = __shfl_xor_sync(0xffffffffu, v, o);
            int i2 = __shfl_xor_sync(0xffffffffu, idx, o);
            if (v2 > v || (v2 == v && i2 < idx)) { v = v2; idx = i2; }
        }
        if (l == 0) {
            if (idx == 0x7fffffff) { out[0] = cy; out[1] = cx; }
            else { out[0] = ys + idx / wx; out[1] = xs + idx % wx; }
        }
    }
}

// ---------------------------------------------------------------- centroid
// all threads call; results (new centre, shift) written by thread 0 into outc/outs (shared)
__device__ __forceinline__ void scl_centroid(const float* X, int Ny, int Nx, const float* __restrict__ psf,
                                             int ph, int pw, int cy, int cx, int* outc, double* outs, double* red) {
    int rad = pw / 2;
    int ry = min(min(cy, Ny - 1 - cy), rad), rx = min(min(cx, Nx - 1 - cx), rad);
    int wy = 2 * ry + 1, wx = 2 * rx + 1;
    double sw = 0.0, sy = 0.0, sx = 0.0;
    for (int i = threadIdx.x; i < wy * wx; i += blockDim.x) {
        int iy = i / wx, ix = i % wx;
        float w = X[(cy - ry + iy) * Nx + (cx - rx + ix)] * psf[(rad - ry + iy) * pw + (rad - rx + ix)];
        sw += (double)w;
        sy += (double)iy * (double)w;
        sx += (double)ix * (double)w;
    }
    sw = block_sum<double>(sw, red);
    sy = block_sum<double>(sy, red);
    sx = block_sum<double>(sx, red);
    if (threadIdx.x == 0) {
        double wsum = (double)(float)sw;         // the reference sums the f32 products in f32
        double my = sy / wsum, mx = sx / wsum;
        double wyr = rint(my), wxr = rint(mx);   // np.round: half to even
        outs[0] = wyr - my;
        outs[1] = wxr - mx;
        outc[0] = (int)(wyr + (double)(cy - ry));
        outc[1] = (int)(wxr + (double)(cx - rx));
    }
    __syncthreads();
}

// ---------------------------------------------------------------- symmetry
// Dirichlet-kernel tables of the k-space symmetry operator in real space
// (DESIGN.md "symmetry"): T[m], H[m] for m = 0..2n-2 (index p_i + p_j - 2o).
//   T = (1/F)[1 + 2 sum_{k=1}^{M} cos(k th)] (+ (-1)^m cos(pi a)/F on the rfft axis)
//   H = (1/F)[ -2 sum_{k=1}^{M} sin(k th)]   (+ (-1)^m sin(pi a)/F on the rfft axis)
// th = 2 pi (m + 2o - a)/F, a = 2*shift, M = (F-1)/2, o = (F-n+1)/2 - F/2.
__device__ __forceinline__ void scl_dirichlet(float* T, float* H, int n, int F, double a, bool rfft_axis) {
    int o = (F - n + 1) / 2 - F / 2;
    int M = (F - 1) / 2;
    for (int i = threadIdx.x; i < 2 * n - 1; i += blockDim.x) {
        int m = i + 2 * o;
        double th = 2.0 * M_PI * ((double)m - a) / (double)F;
        double s2, c2;
        sincos(0.5 * th, &s2, &c2);
        double D, Ssum;
        if (fabs(s2) < 1e-14) { D = 2.0 * M + 1.0; Ssum = 0.0; }
        else {
            double sm, cm;
            sincos(((double)M + 0.5) * th, &sm, &cm);
            D = sm / s2;
            Ssum = (c2 - cm) / (2.0 * s2);
        }
        double t = D / F, h = -2.0 * Ssum / F;
        if (rfft_axis && (F % 2 == 0)) {
            double sg = (m & 1) ? -1.0 : 1.0;
            double sa, ca;
            sincospi(a, &sa, &ca);
            t += sg * ca / F;
            h += sg * sa / F;
        }
        T[i] = (float)t;
        if (H) H[i] = (float)h;
    }
}

// k-space symmetry on the window w of X (in place).  Z: 16-byte aligned scratch of ny*((nx+3)&~3) floats;
// tab: scratch of >= 3*(2*max(ny,nx)-1) + 2*nx + 16 floats.  nfl: next_fast_len table
// (nfl[n] = smallest 5-smooth >= n), nfl_even likewise but even.
__device__ void scl_kspace_symmetry(float* X, int Nx, SclWindow w, double dy, double dx,
                                    float* Z, float* tab, const int* __restrict__ nfl,
                                    const int* __restrict__ nfl_even) {
    const int ny = w.ny, nx = w.nx;
    const int Fy = nfl[2 * ny + 10], Fx = nfl_even[2 * nx + 10];
    float* Ty = tab;
    float* Tx = Ty + (2 * ny - 1);
    float* Hx = Tx + (2 * nx - 1);
    float* bvec = Hx + (2 * nx - 1);      // nx
    float* vvec = bvec + nx;              // nx
    scl_dirichlet(Ty, nullptr, ny, Fy, 2.0 * dy, false);
    scl_dirichlet(Tx, Hx, nx, Fx, 2.0 * dx, true);
    __syncthreads();
    float* Xw = X + w.y0 * Nx + w.x0;
    // Z[i][j] = sum_l Xw[i][l] Tx[j + l]: 8x4 register tiles, the four table values slide along l (one new
    // load per step).  Rows/columns of a tile that fall outside the window read in-bounds garbage whose
    // products only reach outputs that are never used (Z's pad columns, unsaved rows).
    const int Zs = (nx + 3) & ~3;
    const int nty = (ny + 7) / 8, ntx = Zs / 4;
    for (int tile = threadIdx.x; tile < nty * ntx; tile += blockDim.x) {
        const int i0 = (tile / ntx) * 8, j0 = (tile % ntx) * 4;
        // accumulators as packed pairs over two rows: one FFMA2 updates (row 2q, row 2q+1) of a column
        unsigned long long acc2[4][4];
#pragma unroll
        for (int q = 0; q < 4; q++)
#pragma unroll
            for (int c = 0; c < 4; c++) acc2[q][c] = 0ull;
        const float* xr = Xw + i0 * Nx;
        const float* tp = Tx + j0;
        // The two half-warps work on different row groups (same bank whenever Nx is a multiple of 32): the upper
        // half runs its l loop rotated by one element, so that its broadcast load hits the neighbouring bank.
        const int rot = (nx > 1) ? ((threadIdx.x >> 4) & 1) : 0;
#pragma unroll 1
        for (int seg = 0; seg < 2; seg++) {
            const int lb = seg ? 0 : rot, le = seg ? rot : nx;
            if (lb >= le) continue;
            float t0 = tp[lb], t1 = tp[lb + 1], t2 = tp[lb + 2];
#pragma unroll 4
            for (int l = lb; l < le; l++) {
                const float t3 = tp[l + 3];
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const unsigned long long xp = scl_pk(xr[(2 * q) * Nx + l], xr[(2 * q + 1) * Nx + l]);
                    acc2[q][0] = scl_fma2(xp, scl_pk(t0, t0), acc2[q][0]);
                    acc2[q][1] = scl_fma2(xp, scl_pk(t1, t1), acc2[q][1]);
                    acc2[q][2] = scl_fma2(xp, scl_pk(t2, t2), acc2[q][2]);
                    acc2[q][3] = scl_fma2(xp, scl_pk(t3, t3), acc2[q][3]);
                }
                t0 = t1; t1 = t2; t2 = t3;
            }
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const float2 c0 = scl_upk(acc2[q][0]), c1 = scl_upk(acc2[q][1]), c2 = scl_upk(acc2[q][2]), c3 = scl_upk(acc2[q][3]);
            if (i0 + 2 * q < ny)
                *reinterpret_cast<float4*>(Z + (i0 + 2 * q) * Zs + j0) = make_float4(c0.x, c1.x, c2.x, c3.x);
            if (i0 + 2 * q + 1 < ny)
                *reinterpret_cast<float4*>(Z + (i0 + 2 * q + 1) * Zs + j0) = make_float4(c0.y, c1.y, c2.y, c3.y);
        }
    }
    const bool nyq = (Fy % 2 == 0);
    const int oy = (Fy - ny + 1) / 2 - Fy / 2;
    if (nyq) {
        // b[l] = sum_i (-1)^{p_i} Xw[i][l]
        for (int l = threadIdx.x; l < nx; l += blockDim.x) {
            float s = 0.f;
            for (int i = 0; i < ny; i++) s += (((i + oy) & 1) ? -1.f : 1.f) * Xw[i * Nx + l];
            bvec[l] = s;
        }
    }
    __syncthreads();
    if (nyq) {
        double sphi, cphi;
        sincospi(2.0 * dy, &sphi, &cphi);
        float cf = (float)(cphi / Fy), sf = (float)(sphi / Fy);
        for (int j = threadIdx.x; j < nx; j += blockDim.x) {
            float s1 = 0.f, s2 = 0.f;
            for (int l = 0; l < nx; l++) { s1 += Tx[j + l] * bvec[l]; s2 += Hx[j + l] * bvec[l]; }
            vvec[j] = cf * s1 + sf * s2;
        }
        __syncthreads();
    }
    // Y[i][j] = sum_l Ty[i + l] Z[l][j];  X <- (X <= 0) ? 0 : (X + Y)/2       (8x4 tiles, eight sliding table values)
    for (int tile = threadIdx.x; tile < nty * ntx; tile += blockDim.x) {
        const int i0 = (tile / ntx) * 8, j0 = (tile % ntx) * 4;
        // accumulators as packed pairs over two columns
        unsigned long long acc2[8][2];
#pragma unroll
        for (int r = 0; r < 8; r++) { acc2[r][0] = 0ull; acc2[r][1] = 0ull; }
        const float* tp = Ty + i0;
        float t[8];
#pragma unroll
        for (int r = 0; r < 7; r++) t[r] = tp[r];
#pragma unroll 8
        for (int l = 0; l < ny; l++) {
            t[7] = tp[l + 7];
            const float4 zv = *reinterpret_cast<const float4*>(Z + l * Zs + j0);
            const unsigned long long z01 = scl_pk(zv.x, zv.y), z23 = scl_pk(zv.z, zv.w);
#pragma unroll
            for (int r = 0; r < 8; r++) {
                const unsigned long long tt = scl_pk(t[r], t[r]);
                acc2[r][0] = scl_fma2(tt, z01, acc2[r][0]);
                acc2[r][1] = scl_fma2(tt, z23, acc2[r][1]);
            }
#pragma unroll
            for (int r = 0; r < 7; r++) t[r] = t[r + 1];
        }
        float acc[8][4];
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const float2 a = scl_upk(acc2[r][0]), b = scl_upk(acc2[r][1]);
            acc[r][0] = a.x; acc[r][1] = a.y; acc[r][2] = b.x; acc[r][3] = b.y;
        }
#pragma unroll
        for (int r = 0; r < 8; r++)
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const int i = i0 + r, j = j0 + c;
                if (i < ny && j < nx) {
                    float y = acc[r][c];
                    if (nyq) y += (((i + oy) & 1) ? -1.f : 1.f) * vvec[j];
                    const float x = Xw[i * Nx + j];
                    Xw[i * Nx + j] = (x <= 0.f) ? 0.f : 0.5f * (x + y);
                }
            }
    }
    __syncthreads();
}

// soft (strength s) or sdss symmetry on window w, in place.  Pairs (p, flip(p)).
__device__ __forceinline__ void scl_flip_symmetry(float* X, int Nx, SclWindow w, int sdss, float strength) {
    int n = w.ny * w.nx;
    float* Xw = X + w.y0 * Nx + w.x0;
    for (int i = threadIdx.x; i < (n + 1) / 2; i += blockDim.x) {
        int j = n - 1 - i;
        int iy = i / w.nx, ix = i % w.nx, jy = j / w.nx, jx = j % w.nx;
        float a = Xw[iy * Nx + ix], b = Xw[jy * Nx + jx];
        float na, nb;
        if (sdss) { na = fminf(a, b); nb = na; }
        else {
            na = 0.5f * strength * (a + b) + (1.f - strength) * a;
            nb = 0.5f * strength * (b + a) + (1.f - strength) * b;
        }
        Xw[iy * Nx + ix] = na;
        Xw[jy * Nx + jx] = nb;
    }
    __syncthreads();
}

// ---------------------------------------------------------------- monotonicity
// Radius-ordered wavefront of the sequential sweep of operators_pybind11.cc:27-50.
// Pixel at offset (Y,X) from the peak only reads strictly closer 8-neighbours
// (operator.py:569-571), all of which lie on a lower level
//     level(Y,X) = 2 max(|Y|,|X|) + min(|Y|,|X|) - 1,
// so processing level by level (any order inside a level) reproduces the
// radius-sorted sequential result bit for bit (tests/test_wavefront_order.py).
//
// Weights: wtab  f32 [(2Ny-1)(2Nx-1)][8]  normalised weights of an interior pixel
//          ctab  f64 [(2Ny-1)(2Nx-1)][8]  un-normalised cosines (0 where the tap is not
//                                          strictly closer), for pixels on the image border
// both indexed by (Y + Ny-1, X + Nx-1) and built on the host with the reference's
// own formula (scarlet_b200/operator.py:monotonic_tables).
__device__ __forceinline__ float scl_mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double scl_mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float scl_add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double scl_add_rn(double a, double b) { return __dadd_rn(a, b); }

// T = float: the f32 overload (weights rounded to f32 first, operators_pybind11.cc:82-83);
// T = double: the f64 overload.  NEAREST: prox_monotonic (operators_pybind11.cc:11-25) with
// the single best-aligned closer neighbour as reference (operator.py:602-611).
template <typename T, bool NEAREST>
__device__ __forceinline__ void scl_mono_pixel(T* X, int Ny, int Nx, int y, int x, int Yo, int Xo,
                                               const float* __restrict__ wtab, const double* __restrict__ ctab,
                                               T one_minus_thresh) {
    const int ti = ((Yo + Ny - 1) * (2 * Nx - 1) + (Xo + Nx - 1)) * 8;
    T w[8];
    const bool border = (y == 0) | (y == Ny - 1) | (x == 0) | (x == Nx - 1);
    if (sizeof(T) == 4 && !NEAREST && !border) {
        const float4 w0 = *reinterpret_cast<const float4*>(wtab + ti);
        const float4 w1 = *reinterpret_cast<const float4*>(wtab + ti + 4);
        w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w;
        w[4] = w1.x; w[5] = w1.y; w[6] = w1.z; w[7] = w1.w;
    } else {
        double c[8], s = 0.0;
        int best = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            int oy = (i < 3) ? -1 : (i < 5 ? 0 : 1);
            int ox = (i == 0 || i == 3 || i == 5) ? -1 : ((i == 1 || i == 6) ? 0 : 1);
            bool in = (y + oy >= 0) & (y + oy < Ny) & (x + ox >= 0) & (x + ox < Nx);
            c[i] = in ? ctab[ti + i] : 0.0;
            s = s + c[i];
            if (c[i] > c[best]) best = i;
        }
        if (NEAREST) {
#pragma unroll
            for (int i = 0; i < 8; i++) w[i] = (i == best) ? (T)1 : (T)0;
        } else {
            if (s == 0.0) s = 1.0;
#pragma unroll
            for (int i = 0; i < 8; i++) w[i] = (T)(c[i] / s);
        }
    }
    T ref = 0;
    const int p = y * Nx + x;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        int oy = (i < 3) ? -1 : (i < 5 ? 0 : 1);
        int ox = (i == 0 || i == 3 || i == 5) ? -1 : ((i == 1 || i == 6) ? 0 : 1);
        if (w[i] > (T)0) {
            if (NEAREST) ref = X[p + oy * Nx + ox];
            else ref = scl_add_rn(ref, scl_mul_rn(X[p + oy * Nx + ox], w[i]));
        }
    }
    T r = scl_mul_rn(ref, one_minus_thresh);
    T cur = X[p];
    X[p] = (r < cur) ? r : cur;
}

// ---- fast f32 wavefront (the in-loop path) ---------------------------------
// wt9: f32 [9][(2Ny-1)(2Nx-1)][8] normalised weights per border type
//   bt = by*3 + bx, by = 0 interior / 1 first row / 2 last row (bx likewise), built on the host in f64 with the
//   reference's formula and rounded to f32 (operators_pybind11.cc:82-83) -- no arithmetic on weights here.
// NT threads (tid = 0..NT-1, all of them must call) sweep the levels; `sync` separates levels.
// Neighbour values of pixel p.  All nine loads are issued together and unconditionally: X must be readable from
// X[-Nx-1] to X[N+Nx] (a front pad, and whatever scratch follows X in shared memory); a tap outside the image
// has zero weight and its value is never used.
struct SclMonoNb { float x0, x1, x2, x3, x4, x5, x6, x7, cur; };
__device__ __forceinline__ SclMonoNb scl_mono_load(const float* X, int Nx, int p) {
    const float* q = X + p;
    SclMonoNb n;
    n.x0 = q[-Nx - 1]; n.x1 = q[-Nx]; n.x2 = q[-Nx + 1]; n.x3 = q[-1];
    n.x4 = q[1]; n.x5 = q[Nx - 1]; n.x6 = q[Nx]; n.x7 = q[Nx + 1];
    n.cur = q[0];
    return n;
}
// The sum runs in tap order over the positive weights only (operators_pybind11.cc:39-47), un-fused so that it
// rounds like the reference; returns min(cur, ref (1 - thresh)).
__device__ __forceinline__ float scl_mono_value(const SclMonoNb& n, const float4& w0, const float4& w1, float omt) {
    float ref = 0.f;
    if (w0.x > 0.f) ref = __fadd_rn(ref, __fmul_rn(n.x0, w0.x));
    if (w0.y > 0.f) ref = __fadd_rn(ref, __fmul_rn(n.x1, w0.y));
    if (w0.z > 0.f) ref = __fadd_rn(ref, __fmul_rn(n.x2, w0.z));
    if (w0.w > 0.f) ref = __fadd_rn(ref, __fmul_rn(n.x3, w0.w));
    if (w1.x > 0.f) ref = __fadd_rn(ref, __fmul_rn(n.x4, w1.x));
    if (w1.y > 0.f) ref = __fadd_rn(ref, __fmul_rn(n.x5, w1.y));
    if (w1.z > 0.f) ref = __fadd_rn(ref, __fmul_rn(n.x6, w1.z));
    if (w1.w > 0.f) ref = __fadd_rn(ref, __fmul_rn(n.x7, w1.w));
    const float r = __fmul_rn(ref, omt);
    return (r < n.cur) ? r : n.cur;
}

// ---- single-warp wavefront over a precomputed per-centre plan -----------------
// The sweep has ~2 max(H,W) + min(H,W) dependent levels with a few dozen pixels each; the level a pixel belongs
// to depends only on its offset from the centre.  When the tables of a shape are registered, a one-off kernel
// sorts the pixels by level for EVERY possible centre (counting sort in shared memory; the order inside a level
// is irrelevant, see above) and stores the result in global memory as "passes" of 32 entries
//     entry = table_row << pbits | pixel,     0xffffffff = lane idle,
// each level padded to whole passes (table_row = index of the pixel's eight weights in wt9, border type
// included).  In the loop ONE warp per component streams its centre's passes: pass i's entries are copied to a
// ring in shared memory (cp.async) 2R passes ahead, its weights (wt9, L2) R passes ahead, so that the dependent
// chain of a level touches shared memory only and costs ~60 instructions per pass.  The other warps of the CTA
// wait at a barrier and issue nothing.
#ifndef SCL_MONO_RING
#define SCL_MONO_RING 8           // power of two
#endif
#ifndef SCL_STREAM_INLINE
#define SCL_STREAM_INLINE __forceinline__
#endif
#ifndef SCL_EXP
#define SCL_EXP 0
#endif
struct SclMonoPlan {
    const unsigned* list;       // all centres' passes, 32 entries each
    const unsigned* first;      // [N + 1] first pass of centre cy * Nx + cx
    int pbits;                  // bits of the pixel field; 0 = no plan for this shape (generic path)
};
// scratch of the streamed sweep: weights ring [R][2][32] float4, entry ring [SCL_MONO_ENTRIES][32]
#define SCL_MONO_ENTRIES 32     // passes of plan entries held in shared memory (power of two, multiple of 4, > R + 4)
#define SCL_MONO_STREAM_BYTES (SCL_MONO_RING * 1024 + SCL_MONO_ENTRIES * 128)

// Work areas of the plan builder (one CTA per centre)
struct SclMonoWork {
    int* start;                 // [vmax + 3] first list index of each level
    int* fill;                  // [vmax + 3]
    unsigned short* list;       // [N] packed (y << 8 | x), level-ordered
};
__host__ __device__ inline size_t scl_mono_work_bytes(int N, int vmax) {
    return (((size_t)8 * (vmax + 3) + 2 * (size_t)N + 15) & ~(size_t)15);
}
__device__ __forceinline__ SclMonoWork scl_mono_work(void* base, int N, int vmax) {
    SclMonoWork w;
    unsigned char* b = reinterpret_cast<unsigned char*>(base);
    w.start = reinterpret_cast<int*>(b);
    w.fill = w.start + (vmax + 3);
    w.list = reinterpret_cast<unsigned short*>(w.fill + (vmax + 3));
    (void)N;
    return w;
}

__device__ __forceinline__ int scl_mono_level(int Yo, int Xo) {
    const int a = abs(Yo), b = abs(Xo);
    return a > b ? 2 * a + b : 2 * b + a;
}

// all threads of the CTA; needs Ny, Nx <= 256
__device__ void scl_mono_build(const SclMonoWork& wk, int Ny, int Nx, int cy, int cx, int vmax) {
    const int N = Ny * Nx;
    for (int i = threadIdx.x; i < vmax + 3; i += blockDim.x) wk.start[i] = 0;
    __syncthreads();
    const int dyb = blockDim.x / Nx, dxb = blockDim.x - dyb * Nx;      // (y, x) advance of one stride
    {
        int y = threadIdx.x / Nx, x = threadIdx.x - y * Nx;
        for (int p = threadIdx.x; p < N; p += blockDim.x) {
            const int v = scl_mono_level(y - cy, x - cx);
            if (v > 0) atomicAdd(&wk.start[v + 1], 1);
            y += dyb; x += dxb;
            if (x >= Nx) { x -= Nx; y++; }
        }
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        // inclusive scan of the counts (level v is counted at v + 1) -> start[v] = first list index of level v
        const int n = vmax + 3, per = (n + 31) / 32, lane = threadIdx.x;
        int sum = 0;
        for (int i = lane * per; i < min(n, (lane + 1) * per); i++) sum += wk.start[i];
        int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        int run = incl - sum;
        for (int i = lane * per; i < min(n, (lane + 1) * per); i++) {
            run += wk.start[i];
            wk.start[i] = run;
            wk.fill[i] = run;
        }
    }
    __syncthreads();
    {
        int y = threadIdx.x / Nx, x = threadIdx.x - y * Nx;
        for (int p = threadIdx.x; p < N; p += blockDim.x) {
            const int v = scl_mono_level(y - cy, x - cx);
            if (v > 0) wk.list[atomicAdd(&wk.fill[v], 1)] = (unsigned short)((y << 8) | x);
            y += dyb; x += dxb;
            if (x >= Nx) { x -= Nx; y++; }
        }
    }
    __syncthreads();
}

// Called by the 32 lanes of ONE warp.  X readable on [-Nx-1, N+Nx] (scl_mono_load); `passes` = this centre's
// first entry, `scratch` = SCL_MONO_STREAM_BYTES of 16-byte aligned shared memory.
// (Processing the passes of one level two at a time -- two pixels per lane -- was measured and is not faster: the
// warp is bound by its instruction count, not by the number of dependent steps.)
#define SCL_MONO_IDLE 0xffffffffu
// R: weight-ring depth (passes of look-ahead), RL: entry-ring depth (power of two, multiple of 4, > R + 4); scratch =
// R * 1024 + RL * 128 bytes.
template <int R = SCL_MONO_RING, int RL = SCL_MONO_ENTRIES>
__device__ SCL_STREAM_INLINE void scl_monotonic_f32_stream(float* X, int Nx, const unsigned* __restrict__ passes, int npass,
                                                      int pbits, const float* __restrict__ wt9, float thresh, void* scratch) {
    static_assert((RL & (RL - 1)) == 0 && RL % 4 == 0 && RL > R + 4, "entry ring too small");
    const float omt = 1.f - thresh;
    const int lane = threadIdx.x & 31;
    const unsigned pmask = (1u << pbits) - 1u;
    float4* ringW = reinterpret_cast<float4*>(scratch);
    unsigned* ringL = reinterpret_cast<unsigned*>(ringW + R * 64);
    const unsigned aW = (unsigned)__cvta_generic_to_shared(ringW) + lane * 16;
    const unsigned aL = (unsigned)__cvta_generic_to_shared(ringL) + lane * 16;
    // Entries of passes q .. q+3 (q a multiple of 4) as ONE 16-byte copy per lane: the plan stores passes back to back,
    // so the 512 bytes of four passes map 1:1 onto four consecutive ring slots.  (A copy instruction costs the warp
    // ~100 cycles whatever it moves, so fewer, wider copies; .cg = past L1, measurably cheaper than .ca here.)  The
    // quad may reach up to 3 passes beyond npass: the next centre's plan or the slack at the end of the list.
    auto copy_entries4 = [&](int q) {
        if (q < npass)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n"
                         :: "r"(aL + (q & (RL - 1)) * 128), "l"(passes + (size_t)q * 32 + lane * 4));
    };
    // `dep` makes the compiler keep the copy after the ring reads it overwrites (no memory clobber: the copies may
    // otherwise be scheduled freely among the arithmetic)
    auto copy_weights = [&](unsigned e, int slotW, float dep0, float dep1) {
        if (e != SCL_MONO_IDLE) {
            const float* wp = wt9 + (size_t)(e >> pbits) * 8;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" :: "r"(aW + slotW * 1024), "l"(wp), "f"(dep0));
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" :: "r"(aW + slotW * 1024 + 512), "l"(wp + 4), "f"(dep1));
        }
    };
    // prologue: entries of passes 0 .. RL-1, then weights of passes 0 .. R-1
#pragma unroll
    for (int q = 0; q < RL; q += 4) copy_entries4(q);
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
#pragma unroll
    for (int i = 0; i < R; i++) {
        if (i < npass) copy_weights(ringL[i * 32 + lane], i, 0.f, 0.f);
        asm volatile("cp.async.commit_group;\n" ::);
    }
    unsigned e = ringL[lane];               // entry of the next pass, read one pass early
    for (int j = 0; j < npass; j++) {
        const int sW = j & (R - 1);         // weight slot of pass j
        // everything but the R-1 newest groups has landed: weights of pass j (entries land >= 20 passes early)
        asm volatile("cp.async.wait_group %0;\n" :: "n"(R - 1) : "memory");
        const float4 w0 = ringW[sW * 64 + lane], w1 = ringW[sW * 64 + 32 + lane];
        const bool has = e != SCL_MONO_IDLE;
        const int p = has ? (int)(e & pmask) : 0;           // idle lanes run on pixel 0 and store nothing
        const SclMonoNb nb = scl_mono_load(X, Nx, p);
        const unsigned en = ringL[((j + 1) & (RL - 1)) * 32 + lane];      // stale past the end, unused then
        // what a processed pass owes the pipeline: the weights of pass j + R into its weight slot, every fourth
        // pass the entries of four passes into the slots of the four just consumed, one commit group
        const unsigned eR = (j + R < npass) ? ringL[((j + R) & (RL - 1)) * 32 + lane] : SCL_MONO_IDLE;
        copy_weights(eR, sW, w0.x, w1.x);
        if ((j & 3) == 0 && j >= 4) copy_entries4(j + RL - 4);
        asm volatile("cp.async.commit_group;\n" ::);
        const float v = scl_mono_value(nb, w0, w1, omt);
        if (has) X[p] = v;
        __syncwarp();                       // a level may end here: its stores must be visible to the next pass
        e = en;
    }
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

// Plan builder, all threads of the CTA after scl_mono_build(wk, ...) for centre (cy, cx): number of passes
// (out == nullptr) or the passes themselves.  `pbase` = vmax + 3 ints of shared memory.
__device__ int scl_mono_emit(const SclMonoWork& wk, int Ny, int Nx, int cy, int cx, int vmax, int pbits,
                             int* pbase, unsigned* out) {
    const int H2 = 2 * Ny - 1, W2 = 2 * Nx - 1;
    if (threadIdx.x == 0) {
        int run = 0;
        for (int v = 0; v <= vmax; v++) { pbase[v] = run; run += (wk.start[v + 1] - wk.start[v] + 31) >> 5; }
        pbase[vmax + 1] = run;
    }
    __syncthreads();
    const int total = pbase[vmax + 1];
    if (out) {
        for (int v = 1; v <= vmax; v++) {
            const int s0 = wk.start[v], cnt = wk.start[v + 1] - s0, np = pbase[v + 1] - pbase[v];
            for (int s = threadIdx.x; s < np * 32; s += blockDim.x) {
                unsigned e = SCL_MONO_IDLE;
                if (s < cnt) {
                    const int yx = wk.list[s0 + s], y = yx >> 8, x = yx & 255;
                    const int by = (y == 0) ? 1 : ((y == Ny - 1) ? 2 : 0), bx = (x == 0) ? 1 : ((x == Nx - 1) ? 2 : 0);
                    const unsigned row = (unsigned)(((by * 3 + bx) * H2 + (y - cy + Ny - 1)) * W2 + (x - cx + Nx - 1));
                    e = (row << pbits) | (unsigned)(y * Nx + x);
                }
                out[(size_t)pbase[v] * 32 + s] = e;
            }
        }
    }
    return total;
}

template <typename T, bool NEAREST>
__device__ void scl_monotonic(T* X, int Ny, int Nx, int cy, int cx,
                              const float* __restrict__ wtab, const double* __restrict__ ctab, T thresh) {
    const T omt = (T)1 - thresh;
    const int ay = max(cy, Ny - 1 - cy), ax = max(cx, Nx - 1 - cx);
    const int vmax = 2 * max(ay, ax) + min(ay, ax);       // v = level + 1 = 2a + b
    const int slot = threadIdx.x >> 3, oct = threadIdx.x & 7;
    const int nslot = blockDim.x >> 3;
    const bool sb = oct & 1, sa = (oct >> 1) & 1, sw = (oct >> 2) & 1;
    __syncthreads();
    if (NEAREST && threadIdx.x == 0) {
        // the peak refers to itself (operator.py:641-647): x = min(x, x (1 - thresh))
        T cur = X[cy * Nx + cx], r = scl_mul_rn(cur, omt);
        X[cy * Nx + cx] = (r < cur) ? r : cur;
    }
    __syncthreads();
    for (int v = 1; v <= vmax; v++) {
        const int amin = (v + 2) / 3, amax = v / 2;
        for (int a = amin + slot; a <= amax; a += nslot) {
            const int bb = v - 2 * a;
            // drop duplicate images of the same pixel
            if ((bb == 0 && sb) || (a == bb && sw)) continue;
            const int pa = sa ? -a : a, pb = sb ? -bb : bb;
            const int Yo = sw ? pb : pa, Xo = sw ? pa : pb;
            const int y = cy + Yo, x = cx + Xo;
            if (y < 0 || y >= Ny || x < 0 || x >= Nx) continue;
            scl_mono_pixel<T, NEAREST>(X, Ny, Nx, y, x, Yo, Xo, wtab, ctab, omt);
        }
        __syncthreads();
    }
}

// ---- the same level loop for images that live in HBM / L2 (scene mode, array operators beyond the shared-memory
// limit).  Same results bit for bit; written for memory latency instead of shared-memory latency: neighbour loads
// unconditional, border pixels from their pre-normalised tables, a level's pixels computed before any is stored.
// Kept apart from scl_monotonic so that the register allocation of the shared-memory kernels does not change.
// returns the new value of pixel (y, x); the caller stores it (so that several pixels of one level can have their
// loads in flight together)
template <typename T, bool NEAREST>
__device__ __forceinline__ T scl_mono_pixel_value_hbm(const T* X, int Ny, int Nx, int y, int x, int Yo, int Xo,
                                                  const float* __restrict__ wtab, const double* __restrict__ ctab,
                                                  T one_minus_thresh) {
    const int ti = ((Yo + Ny - 1) * (2 * Nx - 1) + (Xo + Nx - 1)) * 8;
    T w[8];
    const bool border = (y == 0) | (y == Ny - 1) | (x == 0) | (x == Nx - 1);
    const int p = y * Nx + x;
    // the eight neighbours are read unconditionally (taps outside the image read the pixel itself; their weight is 0),
    // so that these loads and the weight loads below are in flight together: one memory round trip per level when X
    // lives in HBM/L2 (scene mode)
    T nb[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int oy = (i < 3) ? -1 : (i < 5 ? 0 : 1);
        const int ox = (i == 0 || i == 3 || i == 5) ? -1 : ((i == 1 || i == 6) ? 0 : 1);
        const bool in = (y + oy >= 0) & (y + oy < Ny) & (x + ox >= 0) & (x + ox < Nx);
        nb[i] = X[in ? p + oy * Nx + ox : p];
    }
    if (sizeof(T) == 4 && !NEAREST && (!border || (Ny >= 2 && Nx >= 2))) {
        // border pixels take their pre-normalised table (wt9, one per border type) like the streamed sweep: the f64
        // renormalisation below costs ~2000 instructions, and a warp pays it whenever ONE of its lanes sits on the
        // frame edge -- at 512x512 nearly every level touches an edge (3 us per level, 77 % of the scene-mode kernel)
        const int by = (y == 0) ? 1 : ((y == Ny - 1) ? 2 : 0), bx = (x == 0) ? 1 : ((x == Nx - 1) ? 2 : 0);
        const float* wt = wtab + (size_t)(by * 3 + bx) * (size_t)(2 * Ny - 1) * (size_t)(2 * Nx - 1) * 8;
        const float4 w0 = *reinterpret_cast<const float4*>(wt + ti);
        const float4 w1 = *reinterpret_cast<const float4*>(wt + ti + 4);
        w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w;
        w[4] = w1.x; w[5] = w1.y; w[6] = w1.z; w[7] = w1.w;
    } else {
        double c[8], s = 0.0;
        int best = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            int oy = (i < 3) ? -1 : (i < 5 ? 0 : 1);
            int ox = (i == 0 || i == 3 || i == 5) ? -1 : ((i == 1 || i == 6) ? 0 : 1);
            bool in = (y + oy >= 0) & (y + oy < Ny) & (x + ox >= 0) & (x + ox < Nx);
            c[i] = in ? ctab[ti + i] : 0.0;
            s = s + c[i];
            if (c[i] > c[best]) best = i;
        }
        if (NEAREST) {
#pragma unroll
            for (int i = 0; i < 8; i++) w[i] = (i == best) ? (T)1 : (T)0;
        } else {
            if (s == 0.0) s = 1.0;
#pragma unroll
            for (int i = 0; i < 8; i++) w[i] = (T)(c[i] / s);
        }
    }
    T ref = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        if (w[i] > (T)0) {
            if (NEAREST) ref = nb[i];
            else ref = scl_add_rn(ref, scl_mul_rn(nb[i], w[i]));
        }
    }
    T r = scl_mul_rn(ref, one_minus_thresh);
    T cur = X[p];
    return (r < cur) ? r : cur;
}


// exit_level != nullptr (the in-loop caller, whose next step is prox_plus): the sweep stops once three consecutive
// levels hold no positive pixel and *exit_level receives the last level it processed; every pixel of a higher level
// would come out <= 0 -- its strictly closer neighbours (operator.py:569-571) lie 1, 2 or 3 levels below it, only
// positive weights enter the sum (operators_pybind11.cc:43), so ref <= 0 and min(x, ref (1 - thresh)) <= 0 by
// induction over the levels -- and prox_plus (update.py:27-32) turns it into 0: the caller zeroes those pixels
// instead.  Bit-identical to the full sweep followed by prox_plus (tests/test_gpu_fit.py::test_scene_early_exit_*).
// *exit_level = 0x7fffffff when the sweep ran to the end.  exit_level is shared memory written by thread 0.
template <typename T, bool NEAREST>
__device__ void scl_monotonic_hbm(T* X, int Ny, int Nx, int cy, int cx,
                              const float* __restrict__ wtab, const double* __restrict__ ctab, T thresh,
                              int* exit_level = nullptr) {
    const T omt = (T)1 - thresh;
    const int ay = max(cy, Ny - 1 - cy), ax = max(cx, Nx - 1 - cx);
    const int vmax = 2 * max(ay, ax) + min(ay, ax);       // v = level + 1 = 2a + b
    // An octant per warp (blocks of a multiple of eight warps): of a source near the frame edge whole octants lie
    // outside the frame on the far levels, and their warps then issue nothing; with the octants spread over the lanes
    // of every warp each warp ran the full pixel path for one or two live lanes.
    const bool by_warp = (blockDim.x & 255) == 0;
    const int oct = by_warp ? ((threadIdx.x >> 5) & 7) : (threadIdx.x & 7);
    const int slot = by_warp ? ((threadIdx.x & 31) | ((threadIdx.x >> 8) << 5)) : (threadIdx.x >> 3);
    const int nslot = blockDim.x >> 3;
    const bool sb = oct & 1, sa = (oct >> 1) & 1, sw = (oct >> 2) & 1;
    __syncthreads();
    if (NEAREST && threadIdx.x == 0) {
        // the peak refers to itself (operator.py:641-647): x = min(x, x (1 - thresh))
        T cur = X[cy * Nx + cx], r = scl_mul_rn(cur, omt);
        X[cy * Nx + cx] = (r < cur) ? r : cur;
    }
    __syncthreads();
    // The pixels of a level are independent: each thread takes up to MONO_ILP of them per round and computes all
    // their values before storing any, so their loads overlap (with X in HBM/L2 a round is one memory round trip).
    constexpr int MONO_ILP = 2;
    // This thread's octant reaches the frame edge after Ba steps along its long axis and Bb along its short one: of a
    // level's offsets (a, v - 2a) only a in [(v - Bb + 1) / 2, Ba] lie inside the frame.  Far levels of a source near
    // an edge are mostly outside; enumerating only the inside keeps their cost proportional to what they update (the
    // sweep of one component runs on one SM and is bound by the instructions it issues).
    const int Ba = sw ? (sa ? cx : Nx - 1 - cx) : (sa ? cy : Ny - 1 - cy);
    const int Bb = sw ? (sb ? cy : Ny - 1 - cy) : (sb ? cx : Nx - 1 - cx);
    int dry = 0, vlast = 0x7fffffff;    // consecutive levels without a positive pixel (block-uniform); last level done
    for (int v = 1; v <= vmax; v++) {
        const int amin = max((v + 2) / 3, (v - Bb + 1) >> 1), amax = min(v / 2, Ba);
        int anypos = 0;
        for (int a0 = amin + slot; a0 <= amax; a0 += MONO_ILP * nslot) {
            T val[MONO_ILP];
            int pos[MONO_ILP];
#pragma unroll
            for (int u = 0; u < MONO_ILP; u++) {
                const int a = a0 + u * nslot;
                pos[u] = -1;
                if (a > amax) continue;
                const int bb = v - 2 * a;
                // drop duplicate images of the same pixel
                if ((bb == 0 && sb) || (a == bb && sw)) continue;
                const int pa = sa ? -a : a, pb = sb ? -bb : bb;
                const int Yo = sw ? pb : pa, Xo = sw ? pa : pb;
                const int y = cy + Yo, x = cx + Xo;
                pos[u] = y * Nx + x;
                val[u] = scl_mono_pixel_value_hbm<T, NEAREST>(X, Ny, Nx, y, x, Yo, Xo, wtab, ctab, omt);
            }
#pragma unroll
            for (int u = 0; u < MONO_ILP; u++)
                if (pos[u] >= 0) { X[pos[u]] = val[u]; anypos |= !(val[u] <= (T)0); }
        }
        if (exit_level) {
            // the level barrier doubles as the vote (a level without pixels inside the frame counts as dry)
            dry = __syncthreads_or(anypos) ? 0 : dry + 1;
            if (dry >= 3 && v >= 3) { vlast = v; break; }
        } else {
            __syncthreads();
        }
    }
    if (exit_level && threadIdx.x == 0) *exit_level = vlast;
}

// ---- the float32 weighted sweep of scene mode, written for its instruction count.
// One component's sweep runs on one SM, a level after the other, and a frame-sized component has ~1000 pixels on each
// of ~1000 levels: measured, the generic loop above spent ~450 instructions on a pixel (sign/swap selects, 64-bit
// address arithmetic and in-frame tests per tap, a branch per weight) and a level took 1.7 us with every load removed.
// Here the octant is a property of the warp, pixel and table indices are affine in (a, v - 2a) with per-warp constant
// steps, a pixel reads the four taps that can be strictly closer at fixed offsets, and the weights select instead of
// branch.  Same operations on the same values in the same order as scl_mono_pixel_value_hbm (products rounded, added
// in tap order, taps of weight 0 skipped), hence the same bits.

// Level-ordered weights of the wide sweep.  Of the eight taps of a pixel at offset a (ya, xa) + bb (yb, xb), a >= bb >= 0,
// only four can be strictly closer to the peak: the three one step back along the long axis and the one a step back
// along the short axis; the others always carry weight 0 (the builder checks it on the table it is given).  wlev holds,
// per octant, those four weights of every offset in the order the sweep visits them -- level v = 2a + bb after level,
// a ascending inside a level -- so that the lanes of a warp (consecutive a) read consecutive 16-byte entries instead
// of 32-byte rows scattered over a table of (2Ny-1)(2Nx-1) offsets.  Component j of an entry is the weight of the
// octant's j-th candidate tap in ascending tap index (the order the reference adds them in, operators_pybind11.cc:38-44).
struct SclOctant {
    int ya, xa, yb, xb;      // offset (Yo, Xo) = a (ya, xa) + bb (yb, xb)
    int tap[4];              // the candidate taps (index into the 3x3 neighbourhood without its centre), ascending
    int oy[4], ox[4];        // their steps
};
__host__ __device__ inline int scl_tap_index(int oy, int ox) {
    const int i = (oy + 1) * 3 + (ox + 1);
    return i > 4 ? i - 1 : i;
}
__host__ __device__ inline SclOctant scl_octant(int oct) {
    const bool sb = oct & 1, sa = (oct >> 1) & 1, sw = (oct >> 2) & 1;
    SclOctant o;
    o.ya = sw ? 0 : (sa ? -1 : 1); o.xa = sw ? (sa ? -1 : 1) : 0;
    o.yb = sw ? (sb ? -1 : 1) : 0; o.xb = sw ? 0 : (sb ? -1 : 1);
    const int da[4] = {-1, -1, -1, 0}, db[4] = {-1, 0, 1, -1};
    for (int j = 0; j < 4; j++) {
        o.oy[j] = da[j] * o.ya + db[j] * o.yb;
        o.ox[j] = da[j] * o.xa + db[j] * o.xb;
        o.tap[j] = scl_tap_index(o.oy[j], o.ox[j]);
    }
    for (int i = 1; i < 4; i++)                   // insertion sort by tap index
        for (int j = i; j > 0 && o.tap[j - 1] > o.tap[j]; j--) {
            int t;
            t = o.tap[j]; o.tap[j] = o.tap[j - 1]; o.tap[j - 1] = t;
            t = o.oy[j]; o.oy[j] = o.oy[j - 1]; o.oy[j - 1] = t;
            t = o.ox[j]; o.ox[j] = o.ox[j - 1]; o.ox[j - 1] = t;
        }
    return o;
}
// entries of level v in a table built for offsets up to amax_tab along the long axis
__host__ __device__ inline int scl_wlev_count(int v, int amax_tab) {
    const int lo = (v + 2) / 3, hi = (v >> 1) < amax_tab ? (v >> 1) : amax_tab;
    return hi >= lo ? hi - lo + 1 : 0;
}

// grid (ceil((amax_tab + 1) / 128), 3 * amax_tab, 8 octants); start[v]: first entry of level v; bad: count of offsets
// with a non-zero weight on a tap outside the octant's four (must stay 0 for the table to be used)
__global__ void scl_k_build_wlev(const float* __restrict__ wtab, const int* __restrict__ start, float4* __restrict__ wlev,
                                 int E, int Ny, int Nx, int amax_tab, int* __restrict__ bad) {
    const int v = blockIdx.y + 1, oct = blockIdx.z;
    const int ai = blockIdx.x * blockDim.x + threadIdx.x;
    if (ai >= scl_wlev_count(v, amax_tab)) return;
    const int a = (v + 2) / 3 + ai, bb = v - 2 * a;
    const SclOctant o = scl_octant(oct);
    const int Yo = a * o.ya + bb * o.yb, Xo = a * o.xa + bb * o.xb;
    float4 out = make_float4(0.f, 0.f, 0.f, 0.f);
    if (abs(Yo) <= Ny - 1 && abs(Xo) <= Nx - 1) {
        const float* w = wtab + ((size_t)(Yo + Ny - 1) * (2 * Nx - 1) + (Xo + Nx - 1)) * 8;
        out = make_float4(w[o.tap[0]], w[o.tap[1]], w[o.tap[2]], w[o.tap[3]]);
        int extra = 0;
        for (int i = 0; i < 8; i++)
            if (i != o.tap[0] && i != o.tap[1] && i != o.tap[2] && i != o.tap[3] && w[i] != 0.f) extra++;
        if (extra) atomicAdd(bad, 1);
    }
    wlev[(size_t)oct * E + start[v] + ai] = out;
}

// The pixels of level v in one octant, after dropping the duplicate images on the octant boundaries (bb == 0 belongs to
// the octants with sb == 0, a == bb to those with sw == 0) and what lies outside the frame: a in [lo, hi].
// a3 = (v + 2) / 3, v3 = v % 3, vh = v / 2 are carried along the level loop (no divisions per level).
__device__ __forceinline__ void scl_wide_range(int v, int a3, int v3, int vh, int Ba, int Bb, bool sb, bool sw,
                                               int& lo, int& hi) {
    const int lo0 = a3 + ((sw && v3 == 0) ? 1 : 0);
    const int hi0 = vh - ((sb && !(v & 1)) ? 1 : 0);
    lo = max(lo0, (v - Bb + 1) >> 1);
    hi = min(hi0, Ba);
}

__device__ void scl_monotonic_wide_f32(float* X, int Ny, int Nx, int cy, int cx,
                                       const float* __restrict__ wtab, const double* __restrict__ ctab, float thresh,
                                       int* exit_level = nullptr, const float4* __restrict__ wlev = nullptr, int wlevE = 0) {
    if ((blockDim.x & 511) != 0 || Ny < 2 || Nx < 2 || !wlev) {
        scl_monotonic_hbm<float, false>(X, Ny, Nx, cy, cx, wtab, ctab, thresh, exit_level);
        return;
    }
    const float omt = 1.f - thresh;
    const bool early = exit_level != nullptr;    // (tested once: exit_level points to shared memory through a generic pointer)
    const int ay = max(cy, Ny - 1 - cy), ax = max(cx, Nx - 1 - cx);
    const int vmax = 2 * max(ay, ax) + min(ay, ax);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const int nwm = nw - 1;                      // warps that take interior pixels; the last warp takes the frame edge
    const bool edge_warp = warp == nwm;
    const int W2 = 2 * Nx - 1;
    const int p0 = cy * Nx + cx, t0 = (Ny - 1) * W2 + (Nx - 1);
    const size_t tsize = (size_t)(2 * Ny - 1) * (size_t)W2 * 8;
    const int amax_tab = max(Ny, Nx) - 1;
    // Measured on the frame-sized sweeps that end the kernel (clock64 per warp): a level takes as long as the warp that
    // issues the most instructions for it, at ~10 cycles per instruction; waiting for the loads is 7 % of it.  So:
    // (1) every SCL_WIDE_CHUNK levels the interior warps are dealt to the octants in proportion to the pixels they
    // hold over those levels (a source near an edge has one or two octants that reach across the frame and six that
    // left it: with a fixed octant per warp two warps ran 1 067 rounds while others ran 27); (2) the interior loop is
    // straight-line code over the four candidate taps, its range already stripped of duplicates and edge pixels, the
    // level counters carried instead of divided; (3) the at most sixteen pixels of a level that sit on the frame edge
    // (border-type weights, taps to clamp) belong to one warp of their own, a lane each, which runs beside the others.
    constexpr int SCL_WIDE_CHUNK = 32;
    __shared__ int s_pre[9], s_cnt[8];
    // per-octant values of this warp: set at every deal (interior warps), once (a lane pair of the edge warp per octant)
    int oct = -1, slot = 0, nslot = 32;
    bool sb = false, sw = false;
    int ya = 0, xa = 0, yb = 0, xb = 0, Ba = -1, Bb = -1, pA = 0, pB = 0, tA = 0, tB = 0;
    int d0 = 0, d1 = 0, d2 = 0, d3 = 0;
    const float4* wl = wlev;
    auto set_octant = [&](int o_) {
        const bool sa = (o_ >> 1) & 1;
        oct = o_; sb = o_ & 1; sw = (o_ >> 2) & 1;
        ya = sw ? 0 : (sa ? -1 : 1); xa = sw ? (sa ? -1 : 1) : 0;
        yb = sw ? (sb ? -1 : 1) : 0; xb = sw ? 0 : (sb ? -1 : 1);
        Ba = sw ? (sa ? cx : Nx - 1 - cx) : (sa ? cy : Ny - 1 - cy);      // steps to the frame edge, long axis
        Bb = sw ? (sb ? cy : Ny - 1 - cy) : (sb ? cx : Nx - 1 - cx);      // ... short axis
        pA = ya * Nx + xa; pB = yb * Nx + xb;      // steps of the pixel index per unit of a / bb
        tA = ya * W2 + xa; tB = yb * W2 + xb;      // steps of the table entry
    };
    // edge warp: the four candidate taps of the lane's octant -- offsets in d0..d3, the frame sides each needs (bit 0 up,
    // 1 down, 2 left, 3 right; four bits per tap) and their indices into a row of the weight table (three bits per tap)
    int emask = 0, etap = 0;
    if (edge_warp && lane < 16) {
        set_octant(lane >> 1);
        const SclOctant o = scl_octant(lane >> 1);
        d0 = o.oy[0] * Nx + o.ox[0]; d1 = o.oy[1] * Nx + o.ox[1];
        d2 = o.oy[2] * Nx + o.ox[2]; d3 = o.oy[3] * Nx + o.ox[3];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int m = (o.oy[j] < 0 ? 1 : 0) | (o.oy[j] > 0 ? 2 : 0) | (o.ox[j] < 0 ? 4 : 0) | (o.ox[j] > 0 ? 8 : 0);
            emask |= m << (4 * j);
            etap |= o.tap[j] << (3 * j);
        }
    }
    int estart = 0;                  // first level-ordered entry of the current level
    int a3 = 0, v3 = 0;              // (v + 2) / 3 and v % 3 of the current level
    __syncthreads();
    int dry = 0, vlast = 0x7fffffff;
    for (int v = 1; v <= vmax; v++) {
        if (++v3 == 3) v3 = 0;
        if (v3 == 1) a3++;
        const int vh = v >> 1;
        if (((v - 1) & (SCL_WIDE_CHUNK - 1)) == 0) {
            // (every thread is past the barrier that ended level v - 1)
            if (warp == 0) {
                int n = 0;
                if (lane < 8) {
                    const bool qb = lane & 1, qa = (lane >> 1) & 1, qw = (lane >> 2) & 1;
                    const int qBa = qw ? (qa ? cx : Nx - 1 - cx) : (qa ? cy : Ny - 1 - cy);
                    const int qBb = qw ? (qb ? cy : Ny - 1 - cy) : (qb ? cx : Nx - 1 - cx);
                    const int ve = min(v + SCL_WIDE_CHUNK - 1, vmax);
                    for (int u = v; u <= ve; u++) {
                        int lo, hi;
                        scl_wide_range(u, (u + 2) / 3, u % 3, u >> 1, qBa, qBb, qb, qw, lo, hi);
                        n += max(hi - lo + 1, 0);
                    }
                }
                int tot = n;
#pragma unroll
                for (int o = 1; o < 8; o <<= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
                const int nz = __popc(__ballot_sync(0xffffffffu, n > 0));
                int w = 0;
                if (n > 0) w = 1 + (int)(((long long)(nwm - nz) * n) / tot);
                int incl = w;
#pragma unroll
                for (int o = 1; o < 8; o <<= 1) {
                    const int t = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += t;
                }
                if (lane < 8) { s_pre[lane] = incl - w; s_cnt[lane] = w; }
                if (lane == 7) s_pre[8] = incl;
            }
            __syncthreads();
            if (!edge_warp) {
                int o_ = -1;
#pragma unroll
                for (int o = 0; o < 8; o++)
                    if (warp >= s_pre[o] && warp < s_pre[o] + s_cnt[o]) o_ = o;
                oct = -1; Ba = -1; Bb = -1;
                if (o_ >= 0) {
                    set_octant(o_);
                    slot = (warp - s_pre[o_]) * 32 + lane;
                    nslot = s_cnt[o_] * 32;
                    wl = wlev + (size_t)o_ * wlevE;
                    const SclOctant o = scl_octant(o_);
                    d0 = o.oy[0] * Nx + o.ox[0]; d1 = o.oy[1] * Nx + o.ox[1];
                    d2 = o.oy[2] * Nx + o.ox[2]; d3 = o.oy[3] * Nx + o.ox[3];
                }
            }
        }
        int lo, hi;
        scl_wide_range(v, a3, v3, vh, Ba, Bb, sb, sw, lo, hi);
        int anypos = 0;
        if (!edge_warp) {
            // interior pixels: p = p0 + a pA + (v - 2a) pB; weights at entry estart + a - (v + 2) / 3 of the octant.
            // The pixels on the frame edge -- a == Ba (then hi == Ba) and bb == Bb (then 2 lo == v - Bb) -- are not ours.
            const int ilo = lo + ((2 * lo == v - Bb) ? 1 : 0), ihi = hi - ((hi == Ba) ? 1 : 0);
            float* xv = X + (p0 + v * pB);
            const int pS = pA - 2 * pB;
            const float4* wv = wl + (estart - a3);
            for (int a = ilo + slot; a <= ihi; a += nslot) {
                float* xr = xv + a * pS;
                const float4 w = wv[a];
                const float n0 = xr[d0], n1 = xr[d1], n2 = xr[d2], n3 = xr[d3], cur = xr[0];
                float r = 0.f;
                r = (w.x > 0.f) ? __fadd_rn(r, __fmul_rn(n0, w.x)) : r;
                r = (w.y > 0.f) ? __fadd_rn(r, __fmul_rn(n1, w.y)) : r;
                r = (w.z > 0.f) ? __fadd_rn(r, __fmul_rn(n2, w.z)) : r;
                r = (w.w > 0.f) ? __fadd_rn(r, __fmul_rn(n3, w.w)) : r;
                r = __fmul_rn(r, omt);
                const float val = (r < cur) ? r : cur;
                xr[0] = val;
                anypos |= !(val <= 0.f);
            }
        } else if (lane < 16) {
            // the frame edge: even lanes the pixel with a == Ba of their octant, odd lanes the one with bb == Bb
            const bool isA = !(lane & 1);
            const bool have = isA ? (hi == Ba && lo <= hi) : (2 * lo == v - Bb && lo <= hi && lo != Ba);
            if (have) {
                const int a = isA ? Ba : lo, bb = v - 2 * a;
                float* xr = X + (p0 + a * pA + bb * pB);
                const int y = cy + ya * a + yb * bb, x = cx + xa * a + xb * bb;
                const bool up = y > 0, dn = y < Ny - 1, lf = x > 0, rt = x < Nx - 1;
                // the table of the border type (out-of-frame taps weigh 0 there and read the pixel itself); of its
                // eight weights only the octant's four candidates can be non-zero
                const float* wt = wtab + (size_t)(t0 + a * tA + bb * tB) * 8
                                  + (size_t)((up ? (dn ? 0 : 2) : 1) * 3 + (lf ? (rt ? 0 : 2) : 1)) * tsize;
                const int missing = emask & ~((up ? 0x1111 : 0) | (dn ? 0x2222 : 0) | (lf ? 0x4444 : 0) | (rt ? 0x8888 : 0));
                const float w0 = wt[etap & 7], w1 = wt[(etap >> 3) & 7], w2 = wt[(etap >> 6) & 7], w3 = wt[(etap >> 9) & 7];
                const float n0 = xr[(missing & 0x000f) ? 0 : d0], n1 = xr[(missing & 0x00f0) ? 0 : d1];
                const float n2 = xr[(missing & 0x0f00) ? 0 : d2], n3 = xr[(missing & 0xf000) ? 0 : d3];
                const float cur = xr[0];
                float r = 0.f;
                r = (w0 > 0.f) ? __fadd_rn(r, __fmul_rn(n0, w0)) : r;
                r = (w1 > 0.f) ? __fadd_rn(r, __fmul_rn(n1, w1)) : r;
                r = (w2 > 0.f) ? __fadd_rn(r, __fmul_rn(n2, w2)) : r;
                r = (w3 > 0.f) ? __fadd_rn(r, __fmul_rn(n3, w3)) : r;
                r = __fmul_rn(r, omt);
                const float val = (r < cur) ? r : cur;
                xr[0] = val;
                anypos |= !(val <= 0.f);
            }
        }
        estart += max(min(vh, amax_tab) - a3 + 1, 0);
        if (early) {
            dry = __syncthreads_or(anypos) ? 0 : dry + 1;
            if (dry >= 3 && v >= 3) { vlast = v; break; }
        } else {
            __syncthreads();
        }
    }
    if (early && threadIdx.x == 0) *exit_level = vlast;
}
