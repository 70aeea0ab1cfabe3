// K0: PSF difference kernels and their spectra for many PSF sets at once (the step before the
// loop, once per blend).
//
// Replaces (reference file:line):
//   fft.match_psfs / _kspace_operation   fft.py:264-301     irfft2( rfft2(psf_obs) / rfft2(psf_model) ), centred crop
//   Fourier.fft / from_fft               fft.py:156-211     pad -> ifftshift -> rfftn ;  irfftn -> fftshift -> centre
//   _pad / _centered                     fft.py:20-65       left pad (dS+1)//2, crop start (cur-new+1)//2
//   Observation.match                    observation.py:155-196   one difference kernel per band
// and the spectrum of the truncated kernel on the 128-point convolution grid (what K1 multiplies with, DESIGN "FFT
// length").
//
// The FFT lengths of the reference (next_fast_len(k_obs + k_model + 3), 5-smooth, e.g. 90) are not powers of two and
// the arrays are tiny (41x41 in a 90x90 box), so the transforms are evaluated as separable DFTs in float64 with
// twiddle tables in shared memory: only the k non-zero samples enter the forward sums and only the cropped outputs
// are formed by the inverse ones.  In FFT coordinates an array of length k in a box of F sits at
//   t_j = j + (F-k+1)//2 - F//2      (pad left (F-k+1)//2, ifftshift by F//2)
// and the crop of the fftshift-ed inverse reads exactly the same coordinates with k = the output length.
//
// One CTA per (PSF set, band), 512 threads (one CTA per SM by shared memory; 16 warps hide the shared-load latency):
//   1  T1[i][v] = sum_j psf_obs[i][j]   w_x^(-v t_j)     T2 likewise for the model PSF          (v = 0..Fx/2)
//   2  one warp per column v:  P1[u] = sum_i T1[i][v] w_y^(-u t_i), P2[u];  Q[u] = P1/P2;
//                               U[y][v] = sum_u Q[u] w_y^(+u r_y)                               (y = cropped rows)
//   3  d[y][x] = (Re U[y][0] + 2 sum_v Re(U[y][v] w_x^(v r_x)) + Re U[y][Fx/2] (-1)^(r_x)) / (Fy Fx)   (c2r, Fx even)
//   4  (optional) spectrum of float32(d) on the 128 grid, tap (i, j) at ((i+offy) mod 128, (j+offx) mod 128):
//      V[i][v] = sum_j d[i][j] w^(-v (j+offx));  spec[u][v] = sum_i V[i][v] w^(-u (i+offy))  -> complex64 [128][65]
#include "common.cuh"

struct SclPsfArgs {
    const float* psf_obs;     // [n][C][ky][kx]
    const float* psf_model;   // [n_model][kym][kxm]   (n_model == 1: shared by every set)
    float* diff;              // [n][C][oy][ox]        (may be nullptr)
    float2* spec;             // [n][C][128][65]       (may be nullptr)
    int C, ky, kx, kym, kxm, n_model, Fy, Fx, oy, ox, offy, offx;
};

__host__ __device__ __forceinline__ int scl_psf_pos(int F, int k) { return (F - k + 1) / 2 - F / 2; }
__device__ __forceinline__ int scl_mod(int a, int F) { a %= F; return a < 0 ? a + F : a; }

__device__ __forceinline__ void scl_cfma(double2& acc, double2 a, double2 b) {
    acc.x += a.x * b.x - a.y * b.y;
    acc.y += a.x * b.y + a.y * b.x;
}

// shared-memory layout in doubles
struct SclPsfLayout {
    size_t twy, twx, tw128, T1, T2, U, col, d, p1, p2, total;
};
__host__ __device__ inline SclPsfLayout scl_psf_layout(int ky, int kx, int kym, int kxm, int Fy, int Fx, int oy, int ox,
                                                       int nwarps) {
    SclPsfLayout L;
    const size_t Fh = Fx / 2 + 1;
    size_t o = 0;
    L.twy = o; o += 2 * (size_t)Fy;
    L.twx = o; o += 2 * (size_t)Fx;
    L.tw128 = o; o += 2 * 128;
    const size_t t1 = (size_t)ky * Fh, v1 = (size_t)oy * 65;
    L.T1 = o; o += 2 * (t1 > v1 ? t1 : v1);
    L.T2 = o; o += 2 * (size_t)kym * Fh;
    L.U = o; o += 2 * (size_t)oy * Fh;
    L.col = o; o += 2 * (size_t)nwarps * Fy;
    L.d = o; o += ((size_t)oy * ox + 1) / 2;            // float32
    L.p1 = o; o += ((size_t)ky * kx + 1) / 2;           // float32
    L.p2 = o; o += ((size_t)kym * kxm + 1) / 2;         // float32
    L.total = o;
    return L;
}

#define SCL_PSF_THREADS 512
extern "C" __global__ void __launch_bounds__(SCL_PSF_THREADS)
scl_k_match_psfs(SclPsfArgs a)
{
    extern __shared__ double smem_psf[];
    double* sm = smem_psf;
    const int tid = threadIdx.x, nthr = blockDim.x, warp = tid >> 5, lane = tid & 31, nwarps = nthr >> 5;
    const int c = blockIdx.x, i_set = blockIdx.y;
    const int ky = a.ky, kx = a.kx, kym = a.kym, kxm = a.kxm, Fy = a.Fy, Fx = a.Fx, oy = a.oy, ox = a.ox;
    const int Fh = Fx / 2 + 1;
    const SclPsfLayout L = scl_psf_layout(ky, kx, kym, kxm, Fy, Fx, oy, ox, nwarps);
    double2* twy = (double2*)(sm + L.twy);
    double2* twx = (double2*)(sm + L.twx);
    double2* tw128 = (double2*)(sm + L.tw128);
    double2* T1 = (double2*)(sm + L.T1);
    double2* T2 = (double2*)(sm + L.T2);
    double2* U = (double2*)(sm + L.U);
    double2* col = (double2*)(sm + L.col) + (size_t)warp * Fy;
    float* d = (float*)(sm + L.d);
    float* p1 = (float*)(sm + L.p1);
    float* p2 = (float*)(sm + L.p2);

    // ---- tables (cos, sin)(2 pi t / F) and the two PSFs
    for (int t = tid; t < Fy; t += nthr) { double s, co; sincospi(2.0 * t / Fy, &s, &co); twy[t] = make_double2(co, s); }
    for (int t = tid; t < Fx; t += nthr) { double s, co; sincospi(2.0 * t / Fx, &s, &co); twx[t] = make_double2(co, s); }
    for (int t = tid; t < 128; t += nthr) { double s, co; sincospi(t / 64.0, &s, &co); tw128[t] = make_double2(co, s); }
    const float* g1 = a.psf_obs + ((size_t)i_set * a.C + c) * ky * kx;
    const float* g2 = a.psf_model + (size_t)(a.n_model == 1 ? 0 : i_set) * kym * kxm;
    for (int t = tid; t < ky * kx; t += nthr) p1[t] = g1[t];
    for (int t = tid; t < kym * kxm; t += nthr) p2[t] = g2[t];
    __syncthreads();

    // ---- 1: row transforms of the non-zero rows (forward sign: conjugate of the table)
    {
        const int ax1 = scl_psf_pos(Fx, kx), ax2 = scl_psf_pos(Fx, kxm);
        for (int e = tid; e < (ky + kym) * Fh; e += nthr) {
            const bool second = e >= ky * Fh;
            const int e2 = second ? e - ky * Fh : e;
            const int i = e2 / Fh, v = e2 - i * Fh;
            const float* row = second ? p2 + i * kxm : p1 + i * kx;
            const int k = second ? kxm : kx;
            int idx = scl_mod(v * (second ? ax2 : ax1), Fx);
            double re = 0.0, im = 0.0;
            for (int j = 0; j < k; j++) {
                const double2 w = twx[idx];
                const double x = (double)row[j];
                re += x * w.x; im -= x * w.y;
                idx += v; if (idx >= Fx) idx -= Fx;
            }
            (second ? T2 : T1)[e2] = make_double2(re, im);
        }
    }
    __syncthreads();

    // ---- 2: per spectrum column v: column transforms, division, inverse column transform at the cropped rows
    {
        const int ay1 = scl_psf_pos(Fy, ky), ay2 = scl_psf_pos(Fy, kym), ayo = scl_psf_pos(Fy, oy);
        for (int v = warp; v < Fh; v += nwarps) {
            for (int u = lane; u < Fy; u += 32) {
                double2 P1 = make_double2(0.0, 0.0), P2 = make_double2(0.0, 0.0);
                int idx = scl_mod(-u * ay1, Fy);                       // w^(-u t_i): table index (-u t_i) mod Fy
                const int stepu = scl_mod(-u, Fy);
                for (int i = 0; i < ky; i++) {
                    scl_cfma(P1, T1[i * Fh + v], twy[idx]);
                    idx += stepu; if (idx >= Fy) idx -= Fy;
                }
                idx = scl_mod(-u * ay2, Fy);
                for (int i = 0; i < kym; i++) {
                    scl_cfma(P2, T2[i * Fh + v], twy[idx]);
                    idx += stepu; if (idx >= Fy) idx -= Fy;
                }
                const double den = P2.x * P2.x + P2.y * P2.y;
                col[u] = make_double2((P1.x * P2.x + P1.y * P2.y) / den, (P1.y * P2.x - P1.x * P2.y) / den);
            }
            __syncwarp();
            for (int y = lane; y < oy; y += 32) {
                const int r = scl_mod(y + ayo, Fy);
                double2 acc = make_double2(0.0, 0.0);
                int idx = 0;
                for (int u = 0; u < Fy; u++) {
                    scl_cfma(acc, col[u], twy[idx]);
                    idx += r; if (idx >= Fy) idx -= Fy;
                }
                U[y * Fh + v] = acc;
            }
            __syncwarp();
        }
    }
    __syncthreads();

    // ---- 3: complex-to-real along x at the cropped columns
    {
        const int axo = scl_psf_pos(Fx, ox);
        const double scale = 1.0 / ((double)Fy * (double)Fx);
        for (int e = tid; e < oy * ox; e += nthr) {
            const int y = e / ox, x = e - y * ox;
            const int r = scl_mod(x + axo, Fx);
            const double2* Uy = U + (size_t)y * Fh;
            double acc = Uy[0].x;
            int idx = r;
            for (int v = 1; v < Fh - 1; v++) {
                const double2 w = twx[idx];
                acc += 2.0 * (Uy[v].x * w.x - Uy[v].y * w.y);
                idx += r; if (idx >= Fx) idx -= Fx;
            }
            acc += Uy[Fh - 1].x * ((r & 1) ? -1.0 : 1.0);
            const float val = (float)(acc * scale);
            d[e] = val;
            if (a.diff) a.diff[((size_t)i_set * a.C + c) * oy * ox + e] = val;
        }
    }
    if (!a.spec) return;
    __syncthreads();

    // ---- 4: spectrum of the truncated float32 kernel on the 128-point grid
    double2* V = T1;
    for (int e = tid; e < oy * 65; e += nthr) {
        const int i = e / 65, v = e - i * 65;
        int idx = scl_mod(v * a.offx, 128);
        double re = 0.0, im = 0.0;
        for (int j = 0; j < ox; j++) {
            const double2 w = tw128[idx];
            const double x = (double)d[i * ox + j];
            re += x * w.x; im -= x * w.y;
            idx = (idx + v) & 127;
        }
        V[e] = make_double2(re, im);
    }
    __syncthreads();
    float2* out = a.spec + ((size_t)i_set * a.C + c) * 128 * 65;
    for (int e = tid; e < 128 * 65; e += nthr) {
        const int u = e / 65, v = e - u * 65;
        int idx = scl_mod(-u * a.offy, 128);
        const int stepu = (128 - u) & 127;
        double2 acc = make_double2(0.0, 0.0);
        for (int i = 0; i < oy; i++) {
            scl_cfma(acc, V[i * 65 + v], tw128[idx]);
            idx = (idx + stepu) & 127;
        }
        out[e] = make_float2((float)acc.x, (float)acc.y);
    }
}
