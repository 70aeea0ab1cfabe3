// The Blend.fit loop in float64 (compatibility path: the reference's own loop tests build Frame(dtype=float64),
// /root/reference/tests/test_blend.py:59-63,79-83,100-104).  One blend per call, every array in global memory,
// one plain kernel per step of blend.py:79-102 -- written for agreement with the reference's float64 arithmetic,
// not for throughput (the float32 kernels K1-K4 are the measured path).  The difference-kernel convolution is a
// direct real-space sum: on the frame it equals the reference's zero-padded FFT product (fft.py:264-317) exactly,
// without an FFT's rounding.
//
//   k64_render        component.py:148-170,321-347     M[c] = sum_k A[k,c] S_k
//   k64_forward       observation.py:198-239           R = M (*) d, r = W^2 (R - I), loss partials
//   k64_adjoint       blend.py:105-119 (autograd)      G = r (star) d
//   k64_lipschitz     blend.py:138,186-223             mse.append, L_sed / L_morph (exact: normalised squarings in f64)
//   k64_step          blend.py:87-96                   gradient steps -> X[k], Anew[k]
//   k64_source_center source.py:605-627                joint centre of a MultiComponentSource
//   k64_constrain     source.py:402-440,628-641, update.py:27-68,106-199, blend.py:141-184
#include "prox_dev.cuh"

struct Scl64Args {
    int K, C, N, Ny, Nx, P, n_obs;
    int band[SCL_MAX_BANDS];       // model band of each plane
    int ker_on[SCL_MAX_BANDS];     // plane has a difference kernel (else identical PSFs, observation.py:191-194)
    int ky, kx, offy, offx;        // tap (j, i) of a kernel acts at displacement (j + offy, i + offx)
    double* A;                     // [K][C]
    double* S;                     // [K][N]
    const double* images;          // [P][N]
    const double* weights;         // [P][N] or nullptr
    const double* kernels;         // [P][ky][kx]
    double* M;                     // [C][N]
    double* R;                     // [P][N]  residual r = W^2 (R - I)
    double* G;                     // [P][N]
    double* part;                  // [P][nblk] loss partials
    int nblk;
    double* X;                     // [K][N]  stepped / constrained morphologies
    double* Z;                     // [K][N]  symmetry scratch
    double* tab;                   // [K][tab_stride] Dirichlet tables
    int tab_stride;
    double* Anew;                  // [K][C]
    double* lips;                  // [2] (L_sed, L_morph)
    double* logL;                  // [max_iter]
    double* gram;                  // [2][Kp][Kp] + [2][8][8] squaring buffers
    int Kp;
    const unsigned* cflags;
    int* center;                   // [K][2]
    double* shift;                 // [K][2]
    int* conv;                     // [K]
    const int* msrc;               // [n_msrc][2]
    const double* cw;              // [ph][pw] centroid weights
    int ph, pw;
    const float* wtab;             // monotonic tables of the shape (prox_dev.cuh)
    const double* ctab;
    const int* nfl;
    const int* nfl_even;
    int it0, step, approximate_L;
    double prev_loss, e_rel2;
    int* done;                     // [1] set when every component converged and it > 1
};

// M[c][p] = sum_k A[k][c] S[k][p], components added in order onto zeros (ComponentTree.get_model)
__global__ void k64_render(Scl64Args a) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.N) return;
    double m[SCL_MAX_BANDS];
#pragma unroll
    for (int c = 0; c < SCL_MAX_BANDS; c++) m[c] = 0.0;
    for (int k = 0; k < a.K; k++) {
        const double s = a.S[(size_t)k * a.N + p];
#pragma unroll
        for (int c = 0; c < SCL_MAX_BANDS; c++)
            if (c < a.C) m[c] = m[c] + a.A[k * a.C + c] * s;
    }
#pragma unroll
    for (int c = 0; c < SCL_MAX_BANDS; c++)
        if (c < a.C) a.M[(size_t)c * a.N + p] = m[c];
}

// out[y][x] = sum_{j,i} d[j][i] in[y - j - offy][x - i - offx]   (ADJ: in[y + j + offy][x + i + offx]), zero outside
template <bool ADJ>
__device__ __forceinline__ double k64_conv_pixel(const double* __restrict__ in, const double* __restrict__ d,
                                                 int Ny, int Nx, int ky, int kx, int offy, int offx, int y, int x) {
    double acc = 0.0;
    for (int j = 0; j < ky; j++) {
        const int yy = ADJ ? y + j + offy : y - j - offy;
        if (yy < 0 || yy >= Ny) continue;
        const double* row = in + (size_t)yy * Nx;
        const double* dr = d + j * kx;
        for (int i = 0; i < kx; i++) {
            const int xx = ADJ ? x + i + offx : x - i - offx;
            if (xx >= 0 && xx < Nx) acc += dr[i] * row[xx];
        }
    }
    return acc;
}

// grid (nblk, P), 128 threads
__global__ void k64_forward(Scl64Args a) {
    __shared__ double red[32];
    const int pl = blockIdx.y, p = blockIdx.x * blockDim.x + threadIdx.x;
    double half = 0.0;
    if (p < a.N) {
        const double* Mc = a.M + (size_t)a.band[pl] * a.N;
        const int y = p / a.Nx, x = p - y * a.Nx;
        const double r = a.ker_on[pl] ? k64_conv_pixel<false>(Mc, a.kernels + (size_t)pl * a.ky * a.kx, a.Ny, a.Nx,
                                                              a.ky, a.kx, a.offy, a.offx, y, x)
                                      : Mc[p];
        const double w = a.weights ? a.weights[(size_t)pl * a.N + p] : 1.0;
        const double d = w * (r - a.images[(size_t)pl * a.N + p]);
        half = 0.5 * d * d;
        a.R[(size_t)pl * a.N + p] = w * d;
    }
    half = block_sum<double>(half, red);
    if (threadIdx.x == 0) a.part[pl * a.nblk + blockIdx.x] = half;
}

__global__ void k64_adjoint(Scl64Args a) {
    const int pl = blockIdx.y, p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.N) return;
    const double* r = a.R + (size_t)pl * a.N;
    const int y = p / a.Nx, x = p - y * a.Nx;
    a.G[(size_t)pl * a.N + p] = a.ker_on[pl] ? k64_conv_pixel<true>(r, a.kernels + (size_t)pl * a.ky * a.kx, a.Ny, a.Nx,
                                                                    a.ky, a.kx, a.offy, a.offx, y, x)
                                             : r[p];
}

// lambda_max of the symmetric positive semi-definite n x n matrix Min (row stride ld) by normalised squarings
// B <- B B / tr(B B) (40 of them: the top eigenvector up to (l2/l1)^(2^40)) and the Rayleigh quotient of the
// heaviest column against Min.  All threads of the CTA; W0 / W1: n*n doubles of scratch each.
__device__ double k64_lambda_max(const double* Min, int n, int ld, double* W0, double* W1, double* red) {
    __shared__ double s_tr;
    __shared__ int s_col;
    double tr = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) tr += Min[i * ld + i];
    tr = block_sum<double>(tr, red);
    if (!(tr > 0.0)) return 0.0;
    for (int i = threadIdx.x; i < n * n; i += blockDim.x) W0[i] = Min[(i / n) * ld + (i % n)] / tr;
    __syncthreads();
    double *B = W0, *Bn = W1;
    for (int q = 0; q < 40; q++) {
        double t = 0.0;
        for (int i = threadIdx.x; i < n * n; i += blockDim.x) {
            const int r = i / n, c = i % n;
            double acc = 0.0;
            for (int l = 0; l < n; l++) acc += B[r * n + l] * B[l * n + c];
            Bn[i] = acc;
            if (r == c) t += acc;
        }
        t = block_sum<double>(t, red);
        if (!(t > 0.0)) break;
        for (int i = threadIdx.x; i < n * n; i += blockDim.x) Bn[i] = Bn[i] / t;
        __syncthreads();
        double* tmp = B; B = Bn; Bn = tmp;
    }
    if (threadIdx.x == 0) {
        int best = 0;
        for (int i = 1; i < n; i++) if (B[i * n + i] > B[best * n + best]) best = i;
        s_col = best;
    }
    __syncthreads();
    const int col = s_col;
    // Rayleigh quotient v^T Min v / v^T v with v = column `col`
    double num = 0.0, den = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        double mv = 0.0;
        for (int l = 0; l < n; l++) mv += Min[i * ld + l] * B[l * n + col];
        num += B[i * n + col] * mv;
        den += B[i * n + col] * B[i * n + col];
    }
    num = block_sum<double>(num, red);
    den = block_sum<double>(den, red);
    if (threadIdx.x == 0) s_tr = num / den;
    __syncthreads();
    return s_tr;
}

// one CTA of 256 threads
__global__ void k64_lipschitz(Scl64Args a) {
    __shared__ double red[32];
    __shared__ double ata[SCL_MAX_BANDS * SCL_MAX_BANDS], w0[SCL_MAX_BANDS * SCL_MAX_BANDS], w1[SCL_MAX_BANDS * SCL_MAX_BANDS];
    // Blend._loss: total loss of the observations (plane order, block order), appended to mse (blend.py:135-138)
    double loss = 0.0;
    if (threadIdx.x == 0) {
        for (int i = 0; i < a.P * a.nblk; i++) loss += a.part[i];
        a.logL[a.step] = loss;
    }
    const int K = a.K, C = a.C, N = a.N, Kp = a.Kp;
    double LA, LS;
    if (a.approximate_L) {
        double sa = 0.0, ss = 0.0;
        for (int i = threadIdx.x; i < K * C; i += blockDim.x) sa += a.A[i] * a.A[i];
        for (size_t i = threadIdx.x; i < (size_t)K * N; i += blockDim.x) ss += a.S[i] * a.S[i];
        LS = block_sum<double>(sa, red);
        LA = block_sum<double>(ss, red);
        if (threadIdx.x == 0) {
            const int it = a.it0 + a.step + 1;                      // len(mse) after the append
            const double prev = a.step > 0 ? a.logL[a.step - 1] : a.prev_loss;
            if (it > 1 && loss > prev) { LS *= 2.0; LA *= 2.0; }    // blend.py:199-202
            a.lips[0] = LA * a.n_obs;
            a.lips[1] = LS * a.n_obs;
        }
        return;
    }
    double* gram = a.gram;
    for (int pair = 0; pair < K * K; pair++) {
        const int i = pair / K, j = pair % K;
        if (j < i) continue;
        double acc = 0.0;
        const double *si = a.S + (size_t)i * N, *sj = a.S + (size_t)j * N;
        for (int p = threadIdx.x; p < N; p += blockDim.x) acc += si[p] * sj[p];
        acc = block_sum<double>(acc, red);
        if (threadIdx.x == 0) { gram[i * Kp + j] = acc; gram[j * Kp + i] = acc; }
    }
    for (int i = threadIdx.x; i < C * C; i += blockDim.x) {
        double acc = 0.0;
        for (int k = 0; k < K; k++) acc += a.A[k * C + i / C] * a.A[k * C + i % C];
        ata[i] = acc;
    }
    __syncthreads();
    LA = k64_lambda_max(gram, K, Kp, gram + (size_t)Kp * Kp, gram + 2 * (size_t)Kp * Kp, red);
    LS = k64_lambda_max(ata, C, C, w0, w1, red);
    if (threadIdx.x == 0) { a.lips[0] = LA * a.n_obs; a.lips[1] = LS * a.n_obs; }
}

// ---------------------------------------------------------------- constraint helpers in float64
// measurement.max_pixel (measurement.py:3-29): first row-major argmax of the 5x5 window; warp 0, result in out[2]
__device__ __forceinline__ void k64_max_pixel(const double* X, int Ny, int Nx, int cy, int cx, int* out) {
    if (threadIdx.x < 32) {
        const int l = threadIdx.x;
        const int ys = max(cy - 2, 0), ye = min(cy + 3, Ny), xs = max(cx - 2, 0), xe = min(cx + 3, Nx);
        const int wy = ye - ys, wx = xe - xs;
        double v = -INFINITY;
        int idx = 0x7fffffff;
        if (l < wy * wx) { v = X[(ys + l / wx) * Nx + xs + l % wx]; idx = l; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double v2 = __shfl_xor_sync(0xffffffffu, v, o);
            const int i2 = __shfl_xor_sync(0xffffffffu, idx, o);
            if (v2 > v || (v2 == v && i2 < idx)) { v = v2; idx = i2; }
        }
        if (l == 0) {
            if (idx == 0x7fffffff) { out[0] = cy; out[1] = cx; }
            else { out[0] = ys + idx / wx; out[1] = xs + idx % wx; }
        }
    }
}

// measurement.psf_weighted_centroid (measurement.py:32-94); all threads; thread 0 writes outc / outs
__device__ __forceinline__ void k64_centroid(const double* X, int Ny, int Nx, const double* __restrict__ psf, int ph, int pw,
                                             int cy, int cx, int* outc, double* outs, double* red) {
    const int rad = pw / 2;
    const int ry = min(min(cy, Ny - 1 - cy), rad), rx = min(min(cx, Nx - 1 - cx), rad);
    const int wy = 2 * ry + 1, wx = 2 * rx + 1;
    double sw = 0.0, sy = 0.0, sx = 0.0;
    for (int i = threadIdx.x; i < wy * wx; i += blockDim.x) {
        const int iy = i / wx, ix = i % wx;
        const double w = X[(cy - ry + iy) * Nx + (cx - rx + ix)] * psf[(rad - ry + iy) * pw + (rad - rx + ix)];
        sw += w; sy += iy * w; sx += ix * w;
    }
    sw = block_sum<double>(sw, red);
    sy = block_sum<double>(sy, red);
    sx = block_sum<double>(sx, red);
    if (threadIdx.x == 0) {
        const double my = sy / sw, mx = sx / sw;
        const double wyr = rint(my), wxr = rint(mx);           // np.round: half to even
        outs[0] = wyr - my; outs[1] = wxr - mx;
        outc[0] = (int)(wyr + (double)(cy - ry));
        outc[1] = (int)(wxr + (double)(cx - rx));
    }
    __syncthreads();
}

// Dirichlet-kernel tables of the k-space symmetry operator (scl_dirichlet of prox_dev.cuh, kept in float64)
__device__ __forceinline__ void k64_dirichlet(double* T, double* H, int n, int F, double a, bool rfft_axis) {
    const int o = (F - n + 1) / 2 - F / 2, M = (F - 1) / 2;
    for (int i = threadIdx.x; i < 2 * n - 1; i += blockDim.x) {
        const int m = i + 2 * o;
        const double th = 2.0 * M_PI * ((double)m - a) / (double)F;
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
            const double sg = (m & 1) ? -1.0 : 1.0;
            double sa, ca;
            sincospi(a, &sa, &ca);
            t += sg * ca / F;
            h += sg * sa / F;
        }
        T[i] = t;
        if (H) H[i] = h;
    }
}

// operator.prox_kspace_symmetry (operator.py:253-288) on window w of X, as y = (x + T_y x T_x^T) / 2 with the
// Dirichlet-kernel operators of the reference's own FFT lengths (DESIGN.md "Symmetry"); Z: N doubles; tab: tables
__device__ void k64_kspace_symmetry(double* X, int Nx, SclWindow w, double dy, double dx, double* Z, double* tab,
                                    const int* __restrict__ nfl, const int* __restrict__ nfl_even) {
    const int ny = w.ny, nx = w.nx;
    const int Fy = nfl[2 * ny + 10], Fx = nfl_even[2 * nx + 10];
    double* Ty = tab;
    double* Tx = Ty + (2 * ny - 1);
    double* Hx = Tx + (2 * nx - 1);
    double* bvec = Hx + (2 * nx - 1);
    double* vvec = bvec + nx;
    k64_dirichlet(Ty, nullptr, ny, Fy, 2.0 * dy, false);
    k64_dirichlet(Tx, Hx, nx, Fx, 2.0 * dx, true);
    __syncthreads();
    double* Xw = X + w.y0 * Nx + w.x0;
    for (int e = threadIdx.x; e < ny * nx; e += blockDim.x) {
        const int i = e / nx, j = e - i * nx;
        double acc = 0.0;
        for (int l = 0; l < nx; l++) acc += Xw[i * Nx + l] * Tx[j + l];
        Z[e] = acc;
    }
    const bool nyq = (Fy % 2 == 0);
    const int oy = (Fy - ny + 1) / 2 - Fy / 2;
    if (nyq)
        for (int l = threadIdx.x; l < nx; l += blockDim.x) {
            double s = 0.0;
            for (int i = 0; i < ny; i++) s += (((i + oy) & 1) ? -1.0 : 1.0) * Xw[i * Nx + l];
            bvec[l] = s;
        }
    __syncthreads();
    if (nyq) {
        double sphi, cphi;
        sincospi(2.0 * dy, &sphi, &cphi);
        const double cf = cphi / Fy, sf = sphi / Fy;
        for (int j = threadIdx.x; j < nx; j += blockDim.x) {
            double s1 = 0.0, s2 = 0.0;
            for (int l = 0; l < nx; l++) { s1 += Tx[j + l] * bvec[l]; s2 += Hx[j + l] * bvec[l]; }
            vvec[j] = cf * s1 + sf * s2;
        }
        __syncthreads();
    }
    for (int e = threadIdx.x; e < ny * nx; e += blockDim.x) {
        const int i = e / nx, j = e - i * nx;
        double y = 0.0;
        for (int l = 0; l < ny; l++) y += Ty[i + l] * Z[l * nx + j];
        if (nyq) y += (((i + oy) & 1) ? -1.0 : 1.0) * vvec[j];
        const double x = Xw[i * Nx + j];
        Xw[i * Nx + j] = (x <= 0.0) ? 0.0 : 0.5 * (x + y);       // each thread reads and writes its own pixel only
    }
    __syncthreads();
}

// soft symmetry, strength 1 (operator.py:242-251)
__device__ __forceinline__ void k64_flip_symmetry(double* X, int Nx, SclWindow w) {
    const int n = w.ny * w.nx;
    double* Xw = X + w.y0 * Nx + w.x0;
    for (int i = threadIdx.x; i < (n + 1) / 2; i += blockDim.x) {
        const int j = n - 1 - i;
        const int iy = i / w.nx, ix = i % w.nx, jy = j / w.nx, jx = j % w.nx;
        const double a = Xw[iy * Nx + ix], b = Xw[jy * Nx + jx];
        Xw[iy * Nx + ix] = 0.5 * (a + b);
        Xw[jy * Nx + jx] = 0.5 * (b + a);
    }
    __syncthreads();
}

// gradient steps of component k = blockIdx.x (blend.py:87-96) -> X[k], Anew[k]
__global__ void k64_step(Scl64Args a) {
    __shared__ double red[32];
    const int k = blockIdx.x, N = a.N, C = a.C;
    const unsigned fl = a.cflags[k];
    const double* Sk = a.S + (size_t)k * N;
    double* Xk = a.X + (size_t)k * N;
    double ak[SCL_MAX_BANDS], dA[SCL_MAX_BANDS];
    for (int pl = 0; pl < SCL_MAX_BANDS; pl++) { ak[pl] = pl < a.P ? a.A[k * C + a.band[pl]] : 0.0; dA[pl] = 0.0; }
    if (!a.G) {
        // construction-time pass (self.update() with it = 0, source.py:400,492): no gradient step
        for (int p = threadIdx.x; p < N; p += blockDim.x) Xk[p] = Sk[p];
        if (threadIdx.x < C) a.Anew[k * C + threadIdx.x] = a.A[k * C + threadIdx.x];
        return;
    }
    const double step_morph = 1.0 / a.lips[1], step_sed = 1.0 / a.lips[0];
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        const double s = Sk[p];
        double d = 0.0;
        for (int pl = 0; pl < a.P; pl++) {
            const double g = a.G[(size_t)pl * N + p];
            d += ak[pl] * g;
            dA[pl] += g * s;
        }
        Xk[p] = (fl & SCL_F_FIX_MORPH) ? s : s - step_morph * d;
    }
    for (int pl = 0; pl < a.P; pl++) dA[pl] = block_sum<double>(dA[pl], red);
    if (threadIdx.x < C) {
        double g = 0.0;
        for (int pl = 0; pl < a.P; pl++) if (a.band[pl] == (int)threadIdx.x) g += dA[pl];
        const double old = a.A[k * C + threadIdx.x];
        a.Anew[k * C + threadIdx.x] = (fl & SCL_F_FIX_SED) ? old : old - step_sed * g;
    }
}

// MultiComponentSource.update, first half (source.py:605-627): centre of the source from sum_k morph_k sum(sed_k)
__global__ void k64_source_center(Scl64Args a) {
    __shared__ double red[32];
    __shared__ int sc[2];
    __shared__ double ss[2];
    const int k0 = a.msrc[2 * blockIdx.x], n = a.msrc[2 * blockIdx.x + 1], N = a.N, C = a.C;
    double* Mx = a.Z + (size_t)k0 * N;
    const int it = a.it0 + a.step + 1;
    for (int j = 0; j < n; j++) {
        double ssum = 0.0;
        for (int c = 0; c < C; c++) ssum += a.Anew[(k0 + j) * C + c];
        const double* Xk = a.X + (size_t)(k0 + j) * N;
        for (int p = threadIdx.x; p < N; p += blockDim.x) Mx[p] = (j == 0 ? 0.0 : Mx[p]) + Xk[p] * ssum;
    }
    if (threadIdx.x == 0) { sc[0] = a.center[2 * k0]; sc[1] = a.center[2 * k0 + 1]; ss[0] = a.shift[2 * k0]; ss[1] = a.shift[2 * k0 + 1]; }
    __syncthreads();
    k64_max_pixel(Mx, a.Ny, a.Nx, sc[0], sc[1], sc);
    __syncthreads();
    if ((a.cflags[k0] & SCL_F_SYMMETRIC) && it % 5 == 0) {
        const int cy = sc[0], cx = sc[1];
        __syncthreads();
        k64_centroid(Mx, a.Ny, a.Nx, a.cw, a.ph, a.pw, cy, cx, sc, ss, red);
    }
    __syncthreads();
    if (threadIdx.x < n) {
        const int k = k0 + threadIdx.x;
        a.center[2 * k] = sc[0]; a.center[2 * k + 1] = sc[1];
        a.shift[2 * k] = ss[0]; a.shift[2 * k + 1] = ss[1];
    }
}

// PointSource.update / the per-component half of MultiComponentSource.update (source.py:402-440,628-641), then
// Blend._check_convergence for this component (blend.py:160-174).  One CTA per component.
__global__ void k64_constrain(Scl64Args a) {
    __shared__ double red[32];
    __shared__ int sc[2];
    __shared__ double ss[2];
    __shared__ double smx;
    const int k = blockIdx.x, N = a.N, Ny = a.Ny, Nx = a.Nx, C = a.C;
    const unsigned fl = a.cflags[k];
    double* X = a.X + (size_t)k * N;
    const int it = a.it0 + a.step + 1;
    if (threadIdx.x == 0) { sc[0] = a.center[2 * k]; sc[1] = a.center[2 * k + 1]; ss[0] = a.shift[2 * k]; ss[1] = a.shift[2 * k + 1]; smx = 1.0; }
    __syncthreads();
    const bool constrained = fl & SCL_F_CONSTRAINED;
    if (constrained) {
        const bool own = !(fl & SCL_F_SHARED_CENTER);
        if (own) k64_max_pixel(X, Ny, Nx, sc[0], sc[1], sc);
        __syncthreads();
        if (fl & SCL_F_SYMMETRIC) {
            if (own && it % 5 == 0) {
                const int cy = sc[0], cx = sc[1];
                __syncthreads();
                k64_centroid(X, Ny, Nx, a.cw, a.ph, a.pw, cy, cx, sc, ss, red);
            }
            const SclWindow w = scl_window(Ny, Nx, sc[0], sc[1]);
            if (fl & SCL_F_KSPACE) {
                // update.symmetric discards the result of the centred case (update.py:193-198): no-op there
                if (!w.centered)
                    k64_kspace_symmetry(X, Nx, w, ss[0], ss[1], a.Z + (size_t)k * N, a.tab + (size_t)k * a.tab_stride,
                                        a.nfl, a.nfl_even);
            } else {
                k64_flip_symmetry(X, Nx, w);
            }
        }
        if (fl & SCL_F_MONOTONIC) scl_monotonic_hbm<double, false>(X, Ny, Nx, sc[0], sc[1], a.wtab, a.ctab, 0.0);
        __syncthreads();
        // positive (update.py:27-32) + morph_max (update.py:62-65)
        double mx = -INFINITY;
        for (int p = threadIdx.x; p < N; p += blockDim.x) {
            double v = X[p];
            if (v < 0.0) { v = 0.0; X[p] = 0.0; }
            mx = fmax(mx, v);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        __syncthreads();
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
        __syncthreads();
        if (threadIdx.x == 0) {
            double v = -INFINITY;
            for (int w = 0; w < (int)(blockDim.x >> 5); w++) v = fmax(v, red[w]);
            smx = v;
        }
        __syncthreads();
    }
    const double norm = smx;
    double d2 = 0.0, n2 = 0.0;
    double* Sout = a.S + (size_t)k * N;
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        const double v = constrained ? X[p] / norm : X[p];
        const double df = Sout[p] - v;
        d2 += df * df;
        n2 += v * v;
        Sout[p] = v;
    }
    d2 = block_sum<double>(d2, red);
    n2 = block_sum<double>(n2, red);
    if (threadIdx.x == 0) {
        double ad = 0.0, an = 0.0;
        for (int c = 0; c < C; c++) {
            double v = a.Anew[k * C + c];
            if (constrained) { if (v < 0.0) v = 0.0; v = v * norm; }
            const double df = a.A[k * C + c] - v;
            ad += df * df;
            an += v * v;
            a.A[k * C + c] = v;
        }
        if (it > 1) {
            int f = 0;
            if (!(ad <= a.e_rel2 * an)) f |= SCL_SED_NOT_CONVERGED;
            if (!(d2 <= a.e_rel2 * n2)) f |= SCL_MORPH_NOT_CONVERGED;
            a.conv[k] = f;
        }
        a.center[2 * k] = sc[0]; a.center[2 * k + 1] = sc[1];
        a.shift[2 * k] = ss[0]; a.shift[2 * k + 1] = ss[1];
    }
}

// blend.py:157-176: converged when it > 1 and no component carries a flag
__global__ void k64_converge(Scl64Args a) {
    bool done = (a.it0 + a.step + 1) > 1;
    for (int k = 0; k < a.K; k++) done = done && (a.conv[k] == 0);
    *a.done = done ? 1 : 0;
}

// Observation.render for a stack of models (observation.py:203-220) in float64: one thread per pixel
__global__ void k64_convolve(const double* __restrict__ in, double* __restrict__ out, const double* __restrict__ kernels,
                             int C, int Ny, int Nx, int ky, int kx, int offy, int offx, int adjoint) {
    const int c = blockIdx.y, n = blockIdx.z, N = Ny * Nx, p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= N) return;
    const double* src = in + ((size_t)n * C + c) * N;
    const double* d = kernels + (size_t)c * ky * kx;
    const int y = p / Nx, x = p - y * Nx;
    out[((size_t)n * C + c) * N + p] = adjoint ? k64_conv_pixel<true>(src, d, Ny, Nx, ky, kx, offy, offx, y, x)
                                               : k64_conv_pixel<false>(src, d, Ny, Nx, ky, kx, offy, offx, y, x);
}
