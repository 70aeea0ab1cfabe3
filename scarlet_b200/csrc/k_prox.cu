// K3: the per-source constraint schedule + convergence norms, one CTA per
// component with the morphology resident in shared memory.
// K4: per-blend convergence decision.
//
// Replaces (reference file:line):
//   PointSource.update / ExtendedSource.update   source.py:402-440
//       max_pixel -> (it % 5 == 0) psf_weighted_centroid -> symmetric(kspace)
//       -> monotonic(thresh=0) -> positive -> normalized(morph_max)
//   update.symmetric / monotonic / positive / normalized   update.py:27-68,106-199
//   Blend._check_convergence                     blend.py:141-184
#include "prox_dev.cuh"

struct SclWideRec;
struct SclProxArgs {
    const float* G;           // [B, P, N]  adjoint image of every plane (K1); nullptr: no gradient step (construction pass)
    const float* lips;        // [B, 2]     (L_sed, L_morph) (K2)
    float* A;                 // [Ktot, C]  in: previous iterate; out: new iterate
    float* S;                 // [Ktot, N]  in: previous iterate; out: new iterate
    const int* comp_list;     // [gridDim.x] components to process (compacted active list)
    const int* comp_blend;    // [Ktot]
    const unsigned* cflags;   // [Ktot]
    int* center;              // [Ktot, 2]  (y, x) pixel_center, updated
    double* shift;            // [Ktot, 2]  (dy, dx), updated every 5th iteration
    int* conv;                // [Ktot]     BlendFlag bits, updated when it > 1
    const int* active;        // [B]
    const int* it0;           // [B]
    const float* cw;          // [n_cw, ph, pw] centroid weights (model PSF image)
    const int* cw_idx;        // [B]
    const float* wtab;        // monotonic tables (see prox_dev.cuh)
    const double* ctab;
    SclMonoPlan plan;         // per-centre sweep plans of this shape (pbits == 0: none)
    const float4* wlev;       // level-ordered four-tap weights of the wide sweep (nullptr: none), wlevE entries per octant
    int wlevE;
    const int* nfl;           // next_fast_len tables
    const int* nfl_even;
    int C, Ny, Nx, ph, pw, step;
    SclPlanes pl;             // image planes -> model bands (d loss / d model[c] = sum of G over the planes of band c)
    int it_override;          // >= 0: the iteration number update() sees (0 = source construction)
    double e_rel2;
    unsigned long long* phase_clk;   // optional [8]: clock64 per phase, [0..4] summed over all CTAs, [5..7] maxima (profiling; nullptr = off)
    // Frames too large for shared memory (scene mode, DESIGN "large frames"): the working copy X and the symmetry
    // scratch Z of CTA i live at xscratch + i * xstride in HBM/L2 (same layout as the shared-memory one); only the
    // symmetry tables stay in shared memory.  nullptr: everything in shared memory.
    float* xscratch;
    size_t xstride;
    int no_early_exit;        // tests: run the scene-mode sweep over every level (scl_monotonic_hbm, exit_level)
    // Declared schedules (components flagged SCL_F_SCHEDULE: a custom update() built from scarlet.update /
    // scarlet.measurement calls, traced once on the host) and priors (component.py:39-67,177-187); both only in the
    // wide (scene-mode) kernels, which a batch that uses either is routed to.
    const int* sched_ops;     // [Ktot][SCL_SCHED_MAX] op codes, 0 ends the list
    const float* sched_args;  // [Ktot][SCL_SCHED_MAX]
    const float* prior_par;   // [Ktot][4] (a_sed, a_morph, L_sed, L_morph); nullptr: no prior in the batch
    const float* prior_b_sed; // [Ktot][C]
    const int* prior_idx;     // [Ktot] row of prior_b_morph, -1 = none
    const float* prior_b_morph;   // [n][N]
    // Scene mode in stages (k_prox_split.cu): one record per scratch slot carries a component from the step kernel to
    // the finishing one; waux + i * waux_stride: the symmetry tables of slot i; slot_of[k]: the slot of component k.
    int* box;                 // [Ktot][4] scene mode: footprint box (y0, y1, x0, x1) of the morphology written back, for the
                              // next iteration's Gram matrix (k_scene.cu: scl_k_scene_gram_boxes); nullptr: not wanted
    int part;                 // staged scene mode: 0 every listed component, 1 only those that take the k-space symmetry,
                              // 2 only the others (the two groups finish on different streams)
    SclWideRec* wrec;
    float* waux;
    size_t waux_stride;
    int* slot_of;
};

// What the stages of an iteration hand to each other about the component in one scratch slot.
struct SclWideRec {
    int k;                    // component, -1: nothing to do in this slot (blend no longer active)
    int mode;                 // 1: the k-space operator is applied to the window below by the symmetry kernels; 0: not
    int cy, cx;               // pixel_center after max_pixel / centroid
    int y0, x0, ny, nx;       // window of the symmetry (scl_window)
    int nyq, oy;              // k-space operator: Nyquist row term present, row parity offset
    double dy, dx;            // shift
    float step;               // 1 / L_morph of the component
    float sAnew[SCL_MAX_BANDS], sAold[SCL_MAX_BANDS];
};

#define SCL_SCHED_MAX 16
#ifndef SCL_OP_POSITIVE_SED
// op codes of a declared schedule (host mirror: scarlet_b200/schedule.py)
#define SCL_OP_POSITIVE_SED 1      // update.positive_sed            update.py:13-17
#define SCL_OP_POSITIVE_MORPH 2    // update.positive_morph          update.py:20-24
#define SCL_OP_NORM_SED 3          // update.normalized('sed')       update.py:52-55
#define SCL_OP_NORM_MORPH 4        // update.normalized('morph')     update.py:56-59
#define SCL_OP_NORM_MORPH_MAX 5    // update.normalized('morph_max') update.py:60-63
#define SCL_OP_SPARSE_L0 6         // update.sparse_l0(thresh)       update.py:71-75   arg = thresh
#define SCL_OP_SPARSE_L1 7         // update.sparse_l1(thresh)       update.py:78-82   arg = thresh
#define SCL_OP_THRESHOLD 8         // update.threshold               update.py:85-103
#define SCL_OP_MONOTONIC 9         // update.monotonic(pixel_center, thresh)   update.py:106-156   arg = thresh
#define SCL_OP_SYM_KSPACE 10       // update.symmetric(pixel_center, "kspace") with a `.shift`   update.py:170-199
#define SCL_OP_SYM_SOFT 11         // ... "soft" (or kspace without a shift: strength 1)         arg = strength
#define SCL_OP_SYM_SDSS 12         // ... "sdss"
#define SCL_OP_MAX_PIXEL 13        // pixel_center = measurement.max_pixel(morph, pixel_center)  measurement.py:3-29
#define SCL_OP_CENTROID 14         // pixel_center, shift = measurement.psf_weighted_centroid(...) measurement.py:32-94
#endif

// dynamic smem: X[N] + Z[N] + tab[3*(2*max(Ny,Nx)-1) + 2*Nx] floats
// block: SCL_PROX_THREADS
// Gradient step of one component over its N pixels (the band count is a template parameter so that the band
// loops unroll without predicates and every load address is one 64-bit add):
//   X[p] = S[p] - step_morph * sum_c A[c] G[c][p]   (S[p] if fixm),   dAc[c] += sum_p G[c][p] S[p]
// Four pixels per thread and round: all their loads (S and CC bands of G) are issued before any use.
// PRIOR: the prior's morphology gradient am * S + bm[p] is added to the likelihood's (component.py:181-183)
template <int CC, bool PRIOR = false>
__device__ __forceinline__ void scl_step_loop(const float* __restrict__ Sin, const float* __restrict__ Gb, float* X, int N,
                                              const float (&ak)[SCL_MAX_BANDS], float (&dAc)[SCL_MAX_BANDS],
                                              bool fixm, float step_morph, float am = 0.f,
                                              const float* __restrict__ bm = nullptr) {
    constexpr int PX = 4;
    const int nt = blockDim.x;
    int p0 = threadIdx.x;
    for (; p0 + (PX - 1) * nt < N; p0 += PX * nt) {
        float sv[PX], gv[PX][CC];
#pragma unroll
        for (int u = 0; u < PX; u++) {
            const float* gp = Gb + (p0 + u * nt);
            sv[u] = Sin[p0 + u * nt];
#pragma unroll
            for (int c = 0; c < CC; c++) gv[u][c] = gp[c * N];
        }
#pragma unroll
        for (int u = 0; u < PX; u++) {
            float d = 0.f;
#pragma unroll
            for (int c = 0; c < CC; c++) {
                d += ak[c] * gv[u][c];
                dAc[c] += gv[u][c] * sv[u];
            }
            if (PRIOR) d += am * sv[u] + (bm ? bm[p0 + u * nt] : 0.f);
            X[p0 + u * nt] = fixm ? sv[u] : sv[u] - step_morph * d;
        }
    }
    for (; p0 < N; p0 += nt) {
        const float s = Sin[p0];
        float g[CC], d = 0.f;
#pragma unroll
        for (int c = 0; c < CC; c++) g[c] = Gb[c * N + p0];
#pragma unroll
        for (int c = 0; c < CC; c++) {
            d += ak[c] * g[c];
            dAc[c] += g[c] * s;
        }
        if (PRIOR) d += am * s + (bm ? bm[p0] : 0.f);
        X[p0] = fixm ? s : s - step_morph * d;
    }
}

// Gradients of one component (chain rule through Component.get_model, component.py:166-167)
//   dS[p] = sum_c A[k,c] G[c,p]      dA[c] = sum_p G[c,p] S[k,p]
// and both gradient steps (blend.py:91-96): X = S[k] - (1/L_morph) dS, sAnew = A[k] - (1/L_sed) dA (honouring
// fix_morph / fix_sed); sAold = A[k] as read.  Without G (construction-time pass) X = S[k], sAnew = A[k].  All
// threads of the CTA; the caller synchronises before reading sAnew / sAold.
// PRIOR (wide kernels only): Component.backward_prior (component.py:177-187) for priors of the affine form
// grad_sed = a_s sed + b_s, grad_morph = a_m morph + b_m: the prior's gradients join the likelihood's and its
// Lipschitz constants the blend's, per component and only for parameters that are not fixed.  step_out (optional,
// shared): the component's step_morph = 1 / L_morph, which the sparsity operators scale their threshold with.
template <bool PRIOR = false>
__device__ __forceinline__ void scl_component_step(const SclProxArgs& a, int k, int b, unsigned fl, float* X,
                                                   double (*sdA)[SCL_MAX_BANDS], float* sAnew, float* sAold,
                                                   float* step_out = nullptr) {
    const int N = a.Ny * a.Nx, C = a.C, P = a.pl.P;
    {
        const float* Sin = a.S + (size_t)k * N;
        if (a.G && !(SCL_EXP & 16)) {
            const bool fixm = fl & SCL_F_FIX_MORPH;
            float Lm = a.lips[2 * b + 1], Ls = a.lips[2 * b], a_s = 0.f, a_m = 0.f;
            const float* bm = nullptr;
            const bool has_prior = PRIOR && a.prior_par && a.prior_par[4 * k + 2] >= 0.f;
            if (has_prior) {
                a_s = a.prior_par[4 * k]; a_m = a.prior_par[4 * k + 1];
                if (!(fl & SCL_F_FIX_SED)) Ls = Ls + a.prior_par[4 * k + 2];
                if (!fixm) Lm = Lm + a.prior_par[4 * k + 3];
                if (a.prior_idx[k] >= 0) bm = a.prior_b_morph + (size_t)a.prior_idx[k] * N;
            }
            const float step_morph = 1.0f / Lm;
            if (step_out && threadIdx.x == 0) *step_out = step_morph;
            const float* Gb = a.G + (size_t)b * P * N;
            // per plane: the SED entry of the band it sees; dAc[p] = sum_px G_p S, folded per band below
            float ak[SCL_MAX_BANDS], dAc[SCL_MAX_BANDS];
#pragma unroll
            for (int c = 0; c < SCL_MAX_BANDS; c++) { ak[c] = (c < P) ? a.A[(size_t)k * C + a.pl.band[c]] : 0.f; dAc[c] = 0.f; }
            switch (P) {
#define SCL_STEP_CASE(CC) case CC: if (has_prior) scl_step_loop<CC, PRIOR>(Sin, Gb, X, N, ak, dAc, fixm, step_morph, a_m, bm); \
                               else scl_step_loop<CC>(Sin, Gb, X, N, ak, dAc, fixm, step_morph); break;
                SCL_STEP_CASE(1) SCL_STEP_CASE(2) SCL_STEP_CASE(3) SCL_STEP_CASE(4)
                SCL_STEP_CASE(5) SCL_STEP_CASE(6) SCL_STEP_CASE(7) SCL_STEP_CASE(8)
#undef SCL_STEP_CASE
            }
#pragma unroll
            for (int c = 0; c < SCL_MAX_BANDS; c++) {
                const double v = warp_sum<double>((double)dAc[c]);
                if ((threadIdx.x & 31) == 0) sdA[threadIdx.x >> 5][c] = v;
            }
            __syncthreads();
            if (threadIdx.x < C) {
                // SED step (blend.py:91-93): A[k] -= (1/L_sed) dA
                double v = 0.0;
#pragma unroll
                for (int p = 0; p < SCL_MAX_BANDS; p++)
                    if (p < P && a.pl.band[p] == (int)threadIdx.x)
                        for (int w = 0; w < (int)(blockDim.x >> 5); w++) v += sdA[w][p];
                float an = a.A[(size_t)k * C + threadIdx.x];
                sAold[threadIdx.x] = an;
                float g = (float)v;
                if (has_prior) g = g + (a_s * an + a.prior_b_sed[(size_t)k * C + threadIdx.x]);
                if (!(fl & SCL_F_FIX_SED)) an = an - (1.0f / Ls) * g;
                sAnew[threadIdx.x] = an;
            }
        } else {
            for (int p = threadIdx.x; p < N; p += blockDim.x) X[p] = Sin[p];
            if (threadIdx.x < C) sAnew[threadIdx.x] = sAold[threadIdx.x] = a.A[(size_t)k * C + threadIdx.x];
            if (step_out && threadIdx.x == 0) *step_out = 1.f;
        }
    }
}

// A declared schedule: the update() of a custom source, as the ordered list of scarlet.update / scarlet.measurement
// calls it makes (traced once on the host, scarlet_b200/schedule.py), run on the stepped component.  All threads of
// the CTA; X / Z: the component's working copy and symmetry scratch; sA: its stepped SED (shared); sc / ss: its
// pixel_center and shift (shared, live: ops that move them are seen by the ops after them).
__device__ void scl_run_schedule(const SclProxArgs& a, int k, int b, float* X, float* Z, float* tab, float* sA,
                                 int* sc, double* ss, double* red, float step_morph)
{
    __shared__ float s_n;
    const int N = a.Ny * a.Nx, Ny = a.Ny, Nx = a.Nx, C = a.C;
    const int* ops = a.sched_ops + (size_t)k * SCL_SCHED_MAX;
    const float* args = a.sched_args + (size_t)k * SCL_SCHED_MAX;
    __syncthreads();
    for (int i = 0; i < SCL_SCHED_MAX; i++) {
        const int op = ops[i];
        if (op == 0) break;
        const float arg = args[i];
        switch (op) {
        case SCL_OP_POSITIVE_SED:
            if (threadIdx.x < C && sA[threadIdx.x] < 0.f) sA[threadIdx.x] = 0.f;
            break;
        case SCL_OP_POSITIVE_MORPH:
            for (int p = threadIdx.x; p < N; p += blockDim.x) if (X[p] < 0.f) X[p] = 0.f;
            break;
        case SCL_OP_NORM_SED: case SCL_OP_NORM_MORPH: case SCL_OP_NORM_MORPH_MAX: {
            float n;
            if (op == SCL_OP_NORM_SED) {
                if (threadIdx.x == 0) { float t = 0.f; for (int c = 0; c < C; c++) t = __fadd_rn(t, sA[c]); s_n = t; }
                __syncthreads();
                n = s_n;
            } else if (op == SCL_OP_NORM_MORPH) {
                double t = 0.0;
                for (int p = threadIdx.x; p < N; p += blockDim.x) t += (double)X[p];
                n = (float)block_sum<double>(t, red);
            } else {
                float mx = -INFINITY;
                for (int p = threadIdx.x; p < N; p += blockDim.x) mx = fmaxf(mx, X[p]);
                mx = warp_max(mx);
                __syncthreads();
                if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = (double)mx;
                __syncthreads();
                if (threadIdx.x == 0) {
                    float v = -INFINITY;
                    for (int w = 0; w < (int)(blockDim.x >> 5); w++) v = fmaxf(v, (float)red[w]);
                    s_n = v;
                }
                __syncthreads();
                n = s_n;
            }
            const bool sed_down = (op == SCL_OP_NORM_SED);        // 'sed': sed / n, morph * n; else the reverse
            if (threadIdx.x < C) sA[threadIdx.x] = sed_down ? __fdiv_rn(sA[threadIdx.x], n) : __fmul_rn(sA[threadIdx.x], n);
            for (int p = threadIdx.x; p < N; p += blockDim.x) X[p] = sed_down ? __fmul_rn(X[p], n) : __fdiv_rn(X[p], n);
            break;
        }
        case SCL_OP_SPARSE_L0: case SCL_OP_SPARSE_L1: {
            const float t = arg * step_morph;                      // proxmin: thresh * step
            for (int p = threadIdx.x; p < N; p += blockDim.x) {
                const float v = X[p], av = fabsf(v);
                if (op == SCL_OP_SPARSE_L0) { if (av < t) X[p] = 0.f; }
                else { const float m = fmaxf(av - t, 0.f); X[p] = (v > 0.f) ? m : (v < 0.f ? -m : 0.f * m); }
            }
            break;
        }
        case SCL_OP_THRESHOLD:
            scl_threshold_body(X, Ny, Nx, 1, nullptr, nullptr, nullptr);
            break;
        case SCL_OP_MONOTONIC:
            scl_monotonic_hbm<float, false>(X, Ny, Nx, sc[0], sc[1], a.wtab, a.ctab, arg);
            break;
        case SCL_OP_SYM_KSPACE: case SCL_OP_SYM_SOFT: case SCL_OP_SYM_SDSS: {
            const SclWindow w = scl_window(Ny, Nx, sc[0], sc[1]);
            if (op == SCL_OP_SYM_KSPACE) {
                // the centred case returns a new array that update.symmetric drops (update.py:193-198)
                if (!w.centered) scl_kspace_symmetry(X, Nx, w, ss[0], ss[1], Z, tab, a.nfl, a.nfl_even);
            } else {
                scl_flip_symmetry(X, Nx, w, op == SCL_OP_SYM_SDSS, arg);
            }
            break;
        }
        case SCL_OP_MAX_PIXEL:
            scl_max_pixel(X, Ny, Nx, sc[0], sc[1], sc);
            break;
        case SCL_OP_CENTROID: {
            const float* psf = a.cw + (size_t)a.cw_idx[b] * a.ph * a.pw;
            const int cy = sc[0], cx = sc[1];
            __syncthreads();
            scl_centroid(X, Ny, Nx, psf, a.ph, a.pw, cy, cx, sc, ss, red);
            break;
        }
        default: break;
        }
        __syncthreads();
    }
}

// STAGE 0: the whole update of a component.  STAGE 2 (scene mode in stages): the component arrives stepped, centred and
// made symmetric by the kernels of k_prox_split.cu (record a.wrec[slot]); only mirror-pair symmetry is left to do here.
template <int THREADS, int STAGE = 0>
__device__ __forceinline__ void scl_prox_body(const SclProxArgs& a)
{
    extern __shared__ __align__(16) float smem_f[];
    const int k = a.comp_list[blockIdx.x];
    const int b = a.comp_blend[k];
    if (!a.active[b]) return;
    const int N = a.Ny * a.Nx, Ny = a.Ny, Nx = a.Nx, C = a.C;
    // the narrow variant keeps everything in shared memory (a compile-time fact: X must stay an ld.shared pointer)
    constexpr bool SCENE = THREADS != SCL_PROX_THREADS;
    float* xbase = SCENE ? a.xscratch + (size_t)blockIdx.x * a.xstride : smem_f;
    float* X = xbase + ((Nx + 4) & ~3);         // front pad: the monotonic sweep reads X[-Nx-1 ..] unconditionally
    float* Z = X + ((N + 3) & ~3);
    float* tab = SCENE ? smem_f : Z + Ny * ((Nx + 3) & ~3);
    __shared__ double red[32];
    __shared__ int sc[2];
    __shared__ double ss[2];
    __shared__ float sm;
    __shared__ int sexit;               // last level the scene-mode sweep processed (0x7fffffff: not stopped early)
    if (threadIdx.x == 0) sexit = 0x7fffffff;
    const unsigned fl = a.cflags[k];
    if (STAGE == 2 && a.part) {
        const unsigned ks = SCL_F_CONSTRAINED | SCL_F_SYMMETRIC | SCL_F_KSPACE;
        if ((a.part == 1) != ((fl & ks) == ks)) return;
    }
    long long tprev = a.phase_clk ? clock64() : 0;
    const long long tstart = tprev;
    // slots 0..4: clocks summed over CTAs per phase; 5, 6, 7: the longest symmetry phase, sweep and CTA of the launch
#define SCL_PHASE(i) do { if (a.phase_clk && threadIdx.x == 0) { long long t_ = clock64(); \
        const unsigned long long d_ = (unsigned long long)(t_ - tprev); atomicAdd(a.phase_clk + (i), d_); \
        if ((i) == 1) atomicMax(a.phase_clk + 5, d_); \
        if ((i) == 3) atomicMax(a.phase_clk + 6, d_); \
        if ((i) == 4) atomicMax(a.phase_clk + 7, (unsigned long long)(t_ - tstart)); \
        tprev = t_; } } while (0)
    const int it = a.it_override >= 0 ? a.it_override : a.it0[b] + a.step + 1;   // Blend.it seen by update()

    // gradient steps of this component
    __shared__ double sdA[THREADS / 32][SCL_MAX_BANDS];
    __shared__ float sAnew[SCL_MAX_BANDS], sAold[SCL_MAX_BANDS];
    __shared__ float sstep;
    if (STAGE == 0) {
        scl_component_step<SCENE>(a, k, b, fl, X, sdA, sAnew, sAold, SCENE ? &sstep : nullptr);
        if (threadIdx.x == 0) {
            sc[0] = a.center[2 * k]; sc[1] = a.center[2 * k + 1];
            ss[0] = a.shift[2 * k]; ss[1] = a.shift[2 * k + 1];
        }
    } else {
        const SclWideRec& r = a.wrec[blockIdx.x];
        if (threadIdx.x == 0) { sc[0] = r.cy; sc[1] = r.cx; ss[0] = r.dy; ss[1] = r.dx; }
        if (threadIdx.x < SCL_MAX_BANDS) { sAnew[threadIdx.x] = r.sAnew[threadIdx.x]; sAold[threadIdx.x] = r.sAold[threadIdx.x]; }
    }
    __syncthreads();
    SCL_PHASE(0);      // gradient step

    bool scheduled = false;
    if (STAGE == 0 && SCENE && (fl & SCL_F_SCHEDULE)) {
        scheduled = true;
        scl_run_schedule(a, k, b, X, Z, tab, sAnew, sc, ss, red, sstep);
        if (threadIdx.x == 0) sm = 1.f;
    } else
    if (fl & SCL_F_CONSTRAINED) {
        // components of a multi-component source: centre and shift were measured on the source's combined
        // morphology by scl_k_source_center just before this launch
        const bool own_center = !(fl & SCL_F_SHARED_CENTER);
        if (STAGE == 2) {
            // the k-space symmetry was applied by the kernels before this one; mirror pairs are taken here, on the window
            // of the centre the record carries (a multi-component source's was measured after its components' steps)
            if ((fl & SCL_F_SYMMETRIC) && !(fl & SCL_F_KSPACE))
                scl_flip_symmetry(X, Nx, scl_window(Ny, Nx, sc[0], sc[1]), 0, 1.0f);
        } else {
        if (own_center) scl_max_pixel(X, Ny, Nx, sc[0], sc[1], sc);
        __syncthreads();
        if (fl & SCL_F_SYMMETRIC) {
            if (own_center && it % 5 == 0) {
                const float* psf = a.cw + (size_t)a.cw_idx[b] * a.ph * a.pw;
                int cy = sc[0], cx = sc[1];
                __syncthreads();
                scl_centroid(X, Ny, Nx, psf, a.ph, a.pw, cy, cx, sc, ss, red);
            }
            SclWindow w = scl_window(Ny, Nx, sc[0], sc[1]);
            if (fl & SCL_F_KSPACE) {
                // update.symmetric discards the return value of the centred case (update.py:193-198)
                if (!w.centered) scl_kspace_symmetry(X, Nx, w, ss[0], ss[1], Z, tab, a.nfl, a.nfl_even);
            } else {
                scl_flip_symmetry(X, Nx, w, 0, 1.0f);
            }
        }
        }
        SCL_PHASE(1);  // max_pixel, centroid, symmetry
        if (fl & SCL_F_MONOTONIC) {
            if (a.plan.pbits) {
                // Z and the tables are dead after the symmetry step: scratch of the streamed sweep
                if (threadIdx.x >= 32) {
                    // the waiting warps pull the previous morphology (read again by the convergence sums below)
                    // back into L2 while the sweep runs
                    const float* So = a.S + (size_t)k * N;
                    for (int i = (threadIdx.x - 32) * 32; i < N; i += (blockDim.x - 32) * 32)
                        asm volatile("prefetch.global.L2 [%0];\n" :: "l"(So + i));
                }
                if (threadIdx.x < 32) {
                    const int ctr = sc[0] * Nx + sc[1];
                    const unsigned f0 = a.plan.first[ctr], f1 = a.plan.first[ctr + 1];
                    scl_monotonic_f32_stream(X, Nx, a.plan.list + (size_t)f0 * 32, (int)(f1 - f0), a.plan.pbits,
                                             a.wtab, 0.f, Z);
                }
                __syncthreads();
            } else if (SCENE) {
                scl_monotonic_wide_f32(X, Ny, Nx, sc[0], sc[1], a.wtab, a.ctab, 0.f, a.no_early_exit ? nullptr : &sexit,
                                       a.wlev, a.wlevE);
                __syncthreads();
            } else {
                scl_monotonic<float, false>(X, Ny, Nx, sc[0], sc[1], a.wtab, a.ctab, 0.f);
            }
        }
        SCL_PHASE(3);  // monotonic sweep
        // positive (update.py:27-32) + morph_max (update.py:62-65)
        float mx = -INFINITY;
        if (SCENE && (Nx & 3) == 0) {
            // X lives in HBM / L2 here: 16-byte accesses, two per thread in flight (the scalar loop had one 4-byte load
            // per thread outstanding and ran at a tenth of the memory bandwidth)
            const int vx = sexit, cy = sc[0], cx = sc[1], n4 = N >> 2, nt = blockDim.x;
            float4* X4 = reinterpret_cast<float4*>(X);
            for (int i0 = threadIdx.x; i0 < n4; i0 += 2 * nt) {
                float4 q[2];
                const bool has1 = i0 + nt < n4;
                q[0] = X4[i0];
                q[1] = has1 ? X4[i0 + nt] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int u = 0; u < 2; u++) {
                    if (u == 1 && !has1) break;
                    const int p = (i0 + u * nt) << 2, y = p / Nx, x = p - y * Nx;
                    float* e = reinterpret_cast<float*>(&q[u]);
                    bool dirty = false;
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        // beyond the last level an early-stopped sweep processed, pixels are <= 0 without having been computed
                        if (e[j] < 0.f || (vx != 0x7fffffff && scl_mono_level(y - cy, x + j - cx) > vx)) { e[j] = 0.f; dirty = true; }
                        mx = fmaxf(mx, e[j]);
                    }
                    if (dirty) X4[i0 + u * nt] = q[u];
                }
            }
        } else
        if (SCENE && sexit != 0x7fffffff) {
            // the sweep stopped early: pixels beyond its last level are <= 0 without having been computed
            const int vx = sexit, cy = sc[0], cx = sc[1];
            for (int p = threadIdx.x; p < N; p += blockDim.x) {
                const int y = p / Nx, x = p - y * Nx;
                float v = X[p];
                if (v < 0.f || scl_mono_level(y - cy, x - cx) > vx) { v = 0.f; X[p] = 0.f; }
                mx = fmaxf(mx, v);
            }
        } else
        for (int p = threadIdx.x; p < N; p += blockDim.x) {
            float v = X[p];
            if (v < 0.f) { v = 0.f; X[p] = 0.f; }
            mx = fmaxf(mx, v);
        }
        mx = warp_max(mx);
        __shared__ float smax[32];
        if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = mx;
        __syncthreads();
        if (threadIdx.x < 32) {
            float v = (threadIdx.x < (blockDim.x >> 5)) ? smax[threadIdx.x] : -INFINITY;
            v = warp_max(v);
            if (threadIdx.x == 0) sm = v;
        }
        __syncthreads();
    } else if (threadIdx.x == 0) {
        sm = 1.f;
    }
    __syncthreads();
    const float norm = sm;
    const bool constrained = (fl & SCL_F_CONSTRAINED) && !scheduled;

    // write back + convergence sums (blend.py:160-174)
    double d2 = 0.0, n2 = 0.0;
    float* Sout = a.S + (size_t)k * N;
    // scene mode: the footprint box of what is written (scl_k_scene_bbox spent a pass over every morphology on it)
    const bool want_box = SCENE && a.box != nullptr;
    int by0 = Ny, by1 = -1, bx0 = Nx, bx1 = -1;
#define SCL_BOX_ADD(p_, v_) do { if (want_box && (v_) != 0.f) { const int y_ = (p_) / Nx, x_ = (p_) - y_ * Nx; \
        by0 = min(by0, y_); by1 = max(by1, y_); bx0 = min(bx0, x_); bx1 = max(bx1, x_); } } while (0)
    if (SCENE && (N & 3) == 0) {
        const int n4 = N >> 2, nt = blockDim.x;
        const float4* X4 = reinterpret_cast<const float4*>(X);
        float4* S4 = reinterpret_cast<float4*>(Sout);
        for (int i0 = threadIdx.x; i0 < n4; i0 += 2 * nt) {
            const bool has1 = i0 + nt < n4;
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            float4 o[2], v[2];
            o[0] = S4[i0]; v[0] = X4[i0];
            o[1] = has1 ? S4[i0 + nt] : z; v[1] = has1 ? X4[i0 + nt] : z;
#pragma unroll
            for (int u = 0; u < 2; u++) {
                if (u == 1 && !has1) break;
                float* e = reinterpret_cast<float*>(&v[u]);
                const float* oe = reinterpret_cast<const float*>(&o[u]);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if (constrained) e[j] = e[j] / norm;
                    const double df = (double)oe[j] - (double)e[j];
                    d2 += df * df;
                    n2 += (double)e[j] * (double)e[j];
                    if ((Nx & 3) != 0) SCL_BOX_ADD(((i0 + u * nt) << 2) + j, e[j]);
                }
                if (want_box && (Nx & 3) == 0 && (e[0] != 0.f || e[1] != 0.f || e[2] != 0.f || e[3] != 0.f)) {
                    // the four pixels share a row: one division for the vector
                    const int p_ = (i0 + u * nt) << 2, y_ = p_ / Nx, x_ = p_ - y_ * Nx;
                    const int jf = e[0] != 0.f ? 0 : (e[1] != 0.f ? 1 : (e[2] != 0.f ? 2 : 3));
                    const int jl = e[3] != 0.f ? 3 : (e[2] != 0.f ? 2 : (e[1] != 0.f ? 1 : 0));
                    by0 = min(by0, y_); by1 = max(by1, y_); bx0 = min(bx0, x_ + jf); bx1 = max(bx1, x_ + jl);
                }
                S4[i0 + u * nt] = v[u];
            }
        }
    } else
    for (int p0 = threadIdx.x; p0 < N; p0 += 8 * blockDim.x) {
        float old[8];
#pragma unroll
        for (int u = 0; u < 8; u++) { const int p = p0 + u * blockDim.x; old[u] = (p < N) ? Sout[p] : 0.f; }
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int p = p0 + u * blockDim.x;
            if (p < N) {
                const float v = constrained ? X[p] / norm : X[p];
                const double df = (double)old[u] - (double)v;
                d2 += df * df;
                n2 += (double)v * (double)v;
                Sout[p] = v;
                SCL_BOX_ADD(p, v);
            }
        }
    }
#undef SCL_BOX_ADD
    if (want_box) {
        __shared__ int sbox[4][32];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            by0 = min(by0, __shfl_xor_sync(0xffffffffu, by0, o)); by1 = max(by1, __shfl_xor_sync(0xffffffffu, by1, o));
            bx0 = min(bx0, __shfl_xor_sync(0xffffffffu, bx0, o)); bx1 = max(bx1, __shfl_xor_sync(0xffffffffu, bx1, o));
        }
        if ((threadIdx.x & 31) == 0) {
            sbox[0][threadIdx.x >> 5] = by0; sbox[1][threadIdx.x >> 5] = by1;
            sbox[2][threadIdx.x >> 5] = bx0; sbox[3][threadIdx.x >> 5] = bx1;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int i = 1; i < (int)(blockDim.x >> 5); i++) {
                by0 = min(by0, sbox[0][i]); by1 = max(by1, sbox[1][i]); bx0 = min(bx0, sbox[2][i]); bx1 = max(bx1, sbox[3][i]);
            }
            a.box[4 * k] = by0; a.box[4 * k + 1] = by1; a.box[4 * k + 2] = bx0; a.box[4 * k + 3] = bx1;
        }
    }
    d2 = block_sum<double>(d2, red);
    n2 = block_sum<double>(n2, red);
    if (threadIdx.x < 32) {
        // SED: positivity + morph_max normalisation (update.py:27-32,62-65) and its convergence sums, one lane per
        // band (the previous SED is still in shared memory); lane 0 adds the bands in band order
        const int c = threadIdx.x;
        double ad = 0.0, an = 0.0;
        if (c < C) {
            float v = sAnew[c];
            if (constrained) {
                if (v < 0.f) v = 0.f;
                v = v * norm;
            }
            const double df = (double)sAold[c] - (double)v;
            ad = df * df;
            an = (double)v * (double)v;
            a.A[(size_t)k * C + c] = v;
        }
        double ad2 = 0.0, an2 = 0.0;
        for (int cc = 0; cc < C; cc++) {
            ad2 += __shfl_sync(0xffffffffu, ad, cc);
            an2 += __shfl_sync(0xffffffffu, an, cc);
        }
        if (threadIdx.x == 0) {
            if (it > 1) {
                int f = 0;
                if (!(ad2 <= a.e_rel2 * an2)) f |= SCL_SED_NOT_CONVERGED;
                if (!(d2 <= a.e_rel2 * n2)) f |= SCL_MORPH_NOT_CONVERGED;
                a.conv[k] = f;
            }
            a.center[2 * k] = sc[0]; a.center[2 * k + 1] = sc[1];
            a.shift[2 * k] = ss[0]; a.shift[2 * k + 1] = ss[1];
        }
    }
    SCL_PHASE(4);      // positivity, normalisation, write-back, convergence
#undef SCL_PHASE
}

extern "C" __global__ void __launch_bounds__(SCL_PROX_THREADS, 6)
scl_k_prox(SclProxArgs a) { scl_prox_body<SCL_PROX_THREADS>(a); }

// scene mode (working set in HBM, hundreds of thousands of pixels per component): wider CTAs
extern "C" __global__ void __launch_bounds__(SCL_PROX_WIDE, SCL_WIDE_MIN)
scl_k_prox_wide(SclProxArgs a) { scl_prox_body<SCL_PROX_WIDE>(a); }

// K2b: MultiComponentSource.update, first half (source.py:605-627).  One CTA per multi-component source:
//   _morph = sum_k morph_k * sum(sed_k)  over the source's components, AFTER this iteration's gradient steps
//   pixel_center = max_pixel(_morph, pixel_center);  every 5th iteration (symmetric sources):
//   pixel_center, shift = psf_weighted_centroid(_morph, weight, pixel_center)
// and writes (center, shift) to every component of the source, where K3 picks them up (SCL_F_SHARED_CENTER).
// The stepped morphologies/SEDs are recomputed here with K3's own routine (same thread partition, same
// reduction order, hence the same bits), since K3 keeps them in shared memory only.
// dynamic smem: as scl_k_prox (M [N] in the X slot, the stepped component in the Z slot)
template <int THREADS>
__device__ __forceinline__ void scl_source_center_body(const SclProxArgs& a, const int* __restrict__ msrc)
{
    extern __shared__ __align__(16) float smem_s[];
    const int k0 = msrc[2 * blockIdx.x], n = msrc[2 * blockIdx.x + 1];
    const int b = a.comp_blend[k0];
    if (!a.active[b]) return;
    const int N = a.Ny * a.Nx, Ny = a.Ny, Nx = a.Nx, C = a.C;
    constexpr bool SCENE = THREADS != SCL_PROX_THREADS;
    float* xbase = SCENE ? a.xscratch + (size_t)blockIdx.x * a.xstride : smem_s;
    float* M = xbase + ((Nx + 4) & ~3);
    float* Xk = M + ((N + 3) & ~3);
    __shared__ double red[32];
    __shared__ int sc[2];
    __shared__ double ss[2];
    __shared__ double sdA[THREADS / 32][SCL_MAX_BANDS];
    __shared__ float sAnew[SCL_MAX_BANDS], sAold[SCL_MAX_BANDS];
    const int it = a.it_override >= 0 ? a.it_override : a.it0[b] + a.step + 1;
    const unsigned fl0 = a.cflags[k0];
    for (int j = 0; j < n; j++) {
        const int k = k0 + j;
        scl_component_step<SCENE>(a, k, b, a.cflags[k], Xk, sdA, sAnew, sAold);
        __syncthreads();
        float ssum = 0.f;                                    // c.sed.sum(), float32, band order
        for (int c = 0; c < C; c++) ssum = __fadd_rn(ssum, sAnew[c]);
        for (int p = threadIdx.x; p < N; p += blockDim.x) {
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
        int cy = sc[0], cx = sc[1];
        __syncthreads();
        scl_centroid(M, Ny, Nx, psf, a.ph, a.pw, cy, cx, sc, ss, red);
    }
    __syncthreads();
    if (threadIdx.x < n) {
        const int k = k0 + threadIdx.x;
        a.center[2 * k] = sc[0]; a.center[2 * k + 1] = sc[1];
        a.shift[2 * k] = ss[0]; a.shift[2 * k + 1] = ss[1];
    }
}

extern "C" __global__ void __launch_bounds__(SCL_PROX_THREADS)
scl_k_source_center(SclProxArgs a, const int* __restrict__ msrc) { scl_source_center_body<SCL_PROX_THREADS>(a, msrc); }
extern "C" __global__ void __launch_bounds__(SCL_PROX_WIDE, SCL_WIDE_MIN)
scl_k_source_center_wide(SclProxArgs a, const int* __restrict__ msrc) { scl_source_center_body<SCL_PROX_WIDE>(a, msrc); }

// K4: one thread per listed blend.  blend.py:157-176: never converged before it > 1.
extern "C" __global__ void scl_k_converge(const int* __restrict__ comp_off, const int* __restrict__ conv,
                                          const int* __restrict__ it0, int* __restrict__ active,
                                          int* __restrict__ n_iter, const int* __restrict__ blend_list,
                                          int nB, int step)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nB) return;
    const int b = blend_list[i];
    if (!active[b]) return;
    n_iter[b] = step + 1;
    int it = it0[b] + step + 1;
    bool done = it > 1;
    for (int k = comp_off[b]; k < comp_off[b + 1]; k++) done = done && (conv[k] == 0);
    if (done) active[b] = 0;
}

// K5: ordered compaction of the active blends and of their components (one CTA of 1024 threads).
// counts[0] = active blends, counts[1] = their components.  Lists keep ascending index order, so the
// work a launch does never depends on scheduling.
extern "C" __global__ void __launch_bounds__(1024)
scl_k_compact(const int* __restrict__ active, const int* __restrict__ comp_off, int B,
              int* __restrict__ blend_list, int* __restrict__ comp_list, int* __restrict__ counts)
{
    __shared__ int wsum[2][32];
    __shared__ int base[2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { base[0] = 0; base[1] = 0; }
    __syncthreads();
    for (int start = 0; start < B; start += blockDim.x) {
        const int b = start + threadIdx.x;
        const int flag = (b < B && active[b]) ? 1 : 0;
        const int k0 = flag ? comp_off[b] : 0, nk = flag ? comp_off[b + 1] - k0 : 0;
        int xb = flag, xk = nk;                          // inclusive warp scans
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int yb = __shfl_up_sync(0xffffffffu, xb, o), yk = __shfl_up_sync(0xffffffffu, xk, o);
            if (lane >= o) { xb += yb; xk += yk; }
        }
        if (lane == 31) { wsum[0][warp] = xb; wsum[1][warp] = xk; }
        __syncthreads();
        int ob = base[0], ok = base[1];
        for (int w = 0; w < warp; w++) { ob += wsum[0][w]; ok += wsum[1][w]; }
        if (flag) {
            blend_list[ob + xb - 1] = b;
            int dst = ok + xk - nk;
            for (int j = 0; j < nk; j++) comp_list[dst + j] = k0 + j;
        }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) { base[0] = ob + xb; base[1] = ok + xk; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { counts[0] = base[0]; counts[1] = base[1]; }
}

// ---------------------------------------------------------------- standalone operators
// (drop-in device paths for scarlet.operator.prox_* / operators_pybind11.*)
// one CTA per image; images [n, Ny, Nx] float32 in global memory, updated in place.
// dynamic smem: front pad (scl_mono_load) | image [N] T | SCL_MONO_STREAM_BYTES
template <typename T, bool NEAREST>
__global__ void __launch_bounds__(SCL_THREADS)
scl_k_monotonic(T* __restrict__ imgs, const int* __restrict__ centers, int Ny, int Nx, T thresh,
                const float* __restrict__ wtab, const double* __restrict__ ctab, SclMonoPlan plan, int in_place)
{
    extern __shared__ __align__(16) unsigned char smem_b[];
    const int N = Ny * Nx;
    const size_t pad = (((size_t)Nx + 1) * sizeof(T) + 15) & ~(size_t)15;
    T* g = imgs + (size_t)blockIdx.x * N;
    {   // the device-pointer entry points cannot validate centres on the host: an image with a centre outside stays as is
        const int cy_ = centers[2 * blockIdx.x], cx_ = centers[2 * blockIdx.x + 1];
        if (cy_ < 0 || cy_ >= Ny || cx_ < 0 || cx_ >= Nx) return;
    }
    if (in_place) {
        // images beyond the shared-memory limit: the level loop works on the image where it lies (HBM / L2)
        scl_monotonic_hbm<T, NEAREST>(g, Ny, Nx, centers[2 * blockIdx.x], centers[2 * blockIdx.x + 1], wtab, ctab, thresh);
        return;
    }
    T* X = reinterpret_cast<T*>(smem_b + pad);
    for (int p = threadIdx.x; p < N; p += blockDim.x) X[p] = g[p];
    __syncthreads();
    const int cy = centers[2 * blockIdx.x], cx = centers[2 * blockIdx.x + 1];
    if (sizeof(T) == 4 && !NEAREST && plan.pbits) {
        if (threadIdx.x < 32) {
            const unsigned f0 = plan.first[cy * Nx + cx], f1 = plan.first[cy * Nx + cx + 1];
            scl_monotonic_f32_stream(reinterpret_cast<float*>(X), Nx, plan.list + (size_t)f0 * 32, (int)(f1 - f0),
                                     plan.pbits, wtab, (float)thresh,
                                     smem_b + pad + (((size_t)N * sizeof(T) + 15) & ~(size_t)15));
        }
        __syncthreads();
    } else
        scl_monotonic<T, NEAREST>(X, Ny, Nx, cy, cx, wtab, ctab, thresh);
    for (int p = threadIdx.x; p < N; p += blockDim.x) g[p] = X[p];
}
template __global__ void scl_k_monotonic<float, false>(float*, const int*, int, int, float, const float*, const double*, SclMonoPlan, int);
template __global__ void scl_k_monotonic<double, false>(double*, const int*, int, int, double, const float*, const double*, SclMonoPlan, int);
template __global__ void scl_k_monotonic<double, true>(double*, const int*, int, int, double, const float*, const double*, SclMonoPlan, int);

// Plan builder (see prox_dev.cuh): one CTA per centre c = cy * Nx + cx.  With list == nullptr only the number of
// passes is written (npass[c]); otherwise the passes themselves, at first[c].
// dynamic smem: scl_mono_work_bytes(N, vmax_max) + (vmax_max + 3) ints
extern "C" __global__ void __launch_bounds__(SCL_THREADS)
scl_k_mono_plan(int Ny, int Nx, int pbits, unsigned* __restrict__ npass, const unsigned* __restrict__ first,
                unsigned* __restrict__ list)
{
    extern __shared__ __align__(16) unsigned char smem_p[];
    const int N = Ny * Nx, cy = blockIdx.x / Nx, cx = blockIdx.x % Nx;
    const int ay = max(cy, Ny - 1 - cy), ax = max(cx, Nx - 1 - cx);
    const int vmax = 2 * max(ay, ax) + min(ay, ax);
    const SclMonoWork wk = scl_mono_work(smem_p, N, vmax);
    int* pbase = reinterpret_cast<int*>(smem_p + scl_mono_work_bytes(N, vmax));
    scl_mono_build(wk, Ny, Nx, cy, cx, vmax);
    const int total = scl_mono_emit(wk, Ny, Nx, cy, cx, vmax, pbits, pbase,
                                    list ? list + (size_t)first[blockIdx.x] * 32 : nullptr);
    if (npass && threadIdx.x == 0) npass[blockIdx.x] = (unsigned)total;
}

// N1: init_extended_source on the device (source.py:101-180), one CTA per candidate source, in the frame's
// float32 (what the reference computes in when images, sed and bg_rms are float32):
//   detect = sum_c w[k,c] images[b,c] / jacobian[k]       (build_detection_coadd, source.py:101-136; w = sed/bg^2 on
//            the bands with sed > 0 and 0 elsewhere; bands added in order, un-fused like np.einsum's loop)
//   detect = min(detect, flip(detect)) on the largest window centred on the source      (prox_sdss_symmetry)
//   monotonic sweep with thresh 0.1                                                      (source.py:165-166)
//   morph = detect where detect > cutoff[k], else 0; no pixel left -> ok[k] = 0 (SourceInitError, source.py:171-173)
//   morph /= morph[centre]                                                               (source.py:177-179)
// and writes the morphology into S[k].  list (optional): the components to initialise (grid = their number).
// dynamic smem: as scl_k_prox; SCENE: the working copy lives at xscratch + blockIdx.x * xstride (HBM / L2).
template <bool SCENE>
__device__ __forceinline__ void scl_init_extended_body(
    const float* __restrict__ images, const int* __restrict__ comp_blend, const int* __restrict__ center,
    const float* __restrict__ cw, const float* __restrict__ jac, const float* __restrict__ cutoff, float* __restrict__ S,
    int* __restrict__ ok, int C, int Ny, int Nx, int symmetric, int monotonic, const float* __restrict__ wtab,
    const double* __restrict__ ctab, SclMonoPlan plan, const int* __restrict__ list, float* xscratch, size_t xstride)
{
    extern __shared__ __align__(16) float smem_i[];
    const int k = list ? list[blockIdx.x] : blockIdx.x, b = comp_blend[k], N = Ny * Nx;
    float* xbase = SCENE ? xscratch + (size_t)blockIdx.x * xstride : smem_i;
    float* X = xbase + ((Nx + 4) & ~3);
    float* Z = X + ((N + 3) & ~3);
    __shared__ int any;
    const int cy = center[2 * k], cx = center[2 * k + 1];
    const float* img = images + (size_t)b * C * N;
    float w[SCL_MAX_BANDS];
#pragma unroll
    for (int c = 0; c < SCL_MAX_BANDS; c++) w[c] = (c < C) ? cw[(size_t)k * C + c] : 0.f;
    const float j = jac[k];
    if (threadIdx.x == 0) any = 0;
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < SCL_MAX_BANDS; c++)
            if (c < C && w[c] != 0.f) d = __fadd_rn(d, __fmul_rn(w[c], img[(size_t)c * N + p]));
        X[p] = __fdiv_rn(d, j);
    }
    __syncthreads();
    if (symmetric) scl_flip_symmetry(X, Nx, scl_window(Ny, Nx, cy, cx), 1, 0.f);
    if (monotonic) {
        if (SCENE) {
            scl_monotonic_hbm<float, false>(X, Ny, Nx, cy, cx, wtab, ctab, 0.1f);
        } else if (plan.pbits) {
            if (threadIdx.x < 32) {
                const unsigned f0 = plan.first[cy * Nx + cx], f1 = plan.first[cy * Nx + cx + 1];
                scl_monotonic_f32_stream(X, Nx, plan.list + (size_t)f0 * 32, (int)(f1 - f0), plan.pbits, wtab, 0.1f, Z);
            }
            __syncthreads();
        } else {
            scl_monotonic<float, false>(X, Ny, Nx, cy, cx, wtab, ctab, 0.1f);
        }
    }
    __syncthreads();
    const float cut = cutoff[k];
    int mine = 0;
    for (int p = threadIdx.x; p < N; p += blockDim.x) mine |= (X[p] > cut);
    if (mine) any = 1;
    __syncthreads();
    if (threadIdx.x == 0) ok[k] = any;
    const float peak = (X[cy * Nx + cx] > cut) ? X[cy * Nx + cx] : 0.f;
    float* out = S + (size_t)k * N;
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        const float v = (X[p] > cut) ? X[p] : 0.f;
        out[p] = any ? __fdiv_rn(v, peak) : 0.f;
    }
}

#define SCL_INIT_EXT_PARAMS const float* __restrict__ images, const int* __restrict__ comp_blend, const int* __restrict__ center, \
    const float* __restrict__ cw, const float* __restrict__ jac, const float* __restrict__ cutoff, float* __restrict__ S, \
    int* __restrict__ ok, int C, int Ny, int Nx, int symmetric, int monotonic, const float* __restrict__ wtab, \
    const double* __restrict__ ctab, SclMonoPlan plan, const int* __restrict__ list, float* xscratch, size_t xstride
#define SCL_INIT_EXT_ARGS images, comp_blend, center, cw, jac, cutoff, S, ok, C, Ny, Nx, symmetric, monotonic, wtab, ctab, plan, \
    list, xscratch, xstride
extern "C" __global__ void __launch_bounds__(SCL_PROX_THREADS)
scl_k_init_extended(SCL_INIT_EXT_PARAMS) { scl_init_extended_body<false>(SCL_INIT_EXT_ARGS); }
extern "C" __global__ void __launch_bounds__(SCL_PROX_WIDE, SCL_WIDE_MIN)
scl_k_init_extended_wide(SCL_INIT_EXT_PARAMS) { scl_init_extended_body<true>(SCL_INIT_EXT_ARGS); }

// N1: init_multicomponent_source (source.py:242-291) after the extended initialisation of the source's first
// component: the base morphology is cut into K flux layers at the percentile thresholds (each layer peak-normalised)
// and the K SEDs are the least-squares solution of images = A^T layers (get_best_fit_seds, source.py:74-98).
// One CTA per multi-component source; msrc [n][2] = (first component, K <= 4); perc [K-1] ascending.
extern "C" __global__ void __launch_bounds__(256)
scl_k_init_multi(const float* __restrict__ images, const int* __restrict__ comp_blend, const int* __restrict__ msrc,
                 const float* __restrict__ perc, float* __restrict__ S, float* __restrict__ A, const int* __restrict__ ok,
                 int C, int N)
{
    __shared__ double red[32];
    __shared__ float smax[5];
    __shared__ double sg[4][4], sp[4][SCL_MAX_BANDS];
    const int k0 = msrc[2 * blockIdx.x], K = msrc[2 * blockIdx.x + 1], b = comp_blend[k0];
    if (!ok[k0] || K < 2 || K > 4) return;
    float* base = S + (size_t)k0 * N;
    const float* img = images + (size_t)b * C * N;
    float mx = -INFINITY;
    for (int p = threadIdx.x; p < N; p += blockDim.x) mx = fmaxf(mx, base[p]);
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = (double)mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        float v = -INFINITY;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) v = fmaxf(v, (float)red[w]);
        smax[4] = v;
    }
    __syncthreads();
    const float max_flux = smax[4];
    // perc * max_flux / 100 with an integer percentile and a float32 maximum is float64 arithmetic in numpy, and so are
    // the comparison and the differences below before they are stored as float32 (source.py:263-270)
    double t[4] = {0.0, 0.0, 0.0, 0.0};
    for (int k = 1; k < K; k++) t[k] = (double)perc[k - 1] * (double)max_flux / 100.0;
    // layers (source.py:263-270), written in place: component k0 + k holds layer k
    float lmax[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        const float v = base[p];
        float out[4] = {v, 0.f, 0.f, 0.f};
        for (int k = 1; k < K; k++)
            if ((double)v > t[k]) { out[k - 1] = (float)(t[k] - t[k - 1]); out[k] = (float)((double)v - t[k]); }
        for (int k = 0; k < K; k++) { S[(size_t)(k0 + k) * N + p] = out[k]; lmax[k] = fmaxf(lmax[k], out[k]); }
    }
    for (int k = 0; k < K; k++) {
        const float v = warp_max(lmax[k]);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = (double)v;
        __syncthreads();
        if (threadIdx.x == 0) {
            float m = -INFINITY;
            for (int w = 0; w < (int)(blockDim.x >> 5); w++) m = fmaxf(m, (float)red[w]);
            smax[k] = m;
        }
    }
    __syncthreads();
    // peak-normalise (source.py:273-277) and accumulate the normal equations
    double g[4][4], pr[4][SCL_MAX_BANDS];
    for (int i = 0; i < 4; i++) { for (int j = 0; j < 4; j++) g[i][j] = 0.0; for (int c = 0; c < SCL_MAX_BANDS; c++) pr[i][c] = 0.0; }
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        float l[4] = {0.f, 0.f, 0.f, 0.f};
        for (int k = 0; k < K; k++) {
            l[k] = __fdiv_rn(S[(size_t)(k0 + k) * N + p], smax[k]);
            S[(size_t)(k0 + k) * N + p] = l[k];
        }
        for (int i = 0; i < K; i++) {
            for (int j = i; j < K; j++) g[i][j] += (double)l[i] * (double)l[j];
            for (int c = 0; c < C; c++) pr[i][c] += (double)l[i] * (double)img[(size_t)c * N + p];
        }
    }
    for (int i = 0; i < K; i++) {
        for (int j = i; j < K; j++) { const double v = block_sum<double>(g[i][j], red); if (threadIdx.x == 0) { sg[i][j] = v; sg[j][i] = v; } }
        for (int c = 0; c < C; c++) { const double v = block_sum<double>(pr[i][c], red); if (threadIdx.x == 0) sp[i][c] = v; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // seds = inv(G) P by Gauss-Jordan elimination with partial pivoting (K <= 4, float64)
        double a[4][4 + SCL_MAX_BANDS];
        for (int i = 0; i < K; i++) { for (int j = 0; j < K; j++) a[i][j] = sg[i][j]; for (int c = 0; c < C; c++) a[i][K + c] = sp[i][c]; }
        for (int col = 0; col < K; col++) {
            int piv = col;
            for (int r = col + 1; r < K; r++) if (fabs(a[r][col]) > fabs(a[piv][col])) piv = r;
            for (int j = 0; j < K + C; j++) { const double tmp = a[col][j]; a[col][j] = a[piv][j]; a[piv][j] = tmp; }
            const double d = a[col][col];
            for (int j = 0; j < K + C; j++) a[col][j] /= d;
            for (int r = 0; r < K; r++) {
                if (r == col) continue;
                const double f = a[r][col];
                for (int j = 0; j < K + C; j++) a[r][j] -= f * a[col][j];
            }
        }
        for (int i = 0; i < K; i++) for (int c = 0; c < C; c++) A[(size_t)(k0 + i) * C + c] = (float)a[i][K + c];
    }
}

// N1: PointSource.__init__ (source.py:355-375): the model PSF image projected onto the frame around the source's
// pixel (interpolation.get_projection_slices).  One CTA per listed component.
extern "C" __global__ void __launch_bounds__(256)
scl_k_init_point(const int* __restrict__ list, const int* __restrict__ comp_blend, const int* __restrict__ center,
                 const float* __restrict__ cw, const int* __restrict__ cw_idx, int ph, int pw, float* __restrict__ S,
                 int Ny, int Nx)
{
    const int k = list[blockIdx.x], b = comp_blend[k], N = Ny * Nx;
    const float* psf = cw + (size_t)cw_idx[b] * ph * pw;
    const int py = center[2 * k], px = center[2 * k + 1];
    const int bottom = py - (Ny - 1) / 2 - (ph - 1) / 2 + (Ny >> 1), left = px - (Nx - 1) / 2 - (pw - 1) / 2 + (Nx >> 1);
    float* out = S + (size_t)k * N;
    for (int p = threadIdx.x; p < N; p += blockDim.x) {
        const int y = p / Nx, x = p - y * Nx, iy = y - bottom, ix = x - left;
        out[p] = (iy >= 0 && iy < ph && ix >= 0 && ix < pw) ? psf[iy * pw + ix] : 0.f;
    }
}

// mode: 0 = kspace (shift required), 1 = soft(strength), 2 = sdss
extern "C" __global__ void __launch_bounds__(SCL_THREADS)
scl_k_symmetry(float* __restrict__ imgs, const int* __restrict__ centers, const double* __restrict__ shifts,
               int Ny, int Nx, int mode, float strength, const int* __restrict__ nfl,
               const int* __restrict__ nfl_even, float* __restrict__ zscratch)
{
    extern __shared__ __align__(16) float smem_f[];
    const int N = Ny * Nx;
    float* g = imgs + (size_t)blockIdx.x * N;
    // zscratch != nullptr (images beyond the shared-memory limit): work on the image in place, Z in HBM scratch
    float* X = zscratch ? g : smem_f;
    float* Z = zscratch ? zscratch + (size_t)blockIdx.x * Ny * ((Nx + 3) & ~3) : X + ((N + 3) & ~3);
    float* tab = zscratch ? smem_f : Z + Ny * ((Nx + 3) & ~3);
    if (!zscratch) for (int p = threadIdx.x; p < N; p += blockDim.x) X[p] = g[p];
    __syncthreads();
    SclWindow w;
    if (centers[2 * blockIdx.x] < 0) { w.y0 = 0; w.x0 = 0; w.ny = Ny; w.nx = Nx; w.centered = false; }   // whole array
    else w = scl_window(Ny, Nx, centers[2 * blockIdx.x], centers[2 * blockIdx.x + 1]);
    if (mode == 0) scl_kspace_symmetry(X, Nx, w, shifts[2 * blockIdx.x], shifts[2 * blockIdx.x + 1], Z, tab, nfl, nfl_even);
    else scl_flip_symmetry(X, Nx, w, mode == 2, strength);
    if (!zscratch) for (int p = threadIdx.x; p < N; p += blockDim.x) g[p] = X[p];
}

// max_pixel + psf_weighted_centroid for a stack of images (measurement.py)
extern "C" __global__ void __launch_bounds__(SCL_THREADS)
scl_k_measure(const float* __restrict__ imgs, int* __restrict__ centers, double* __restrict__ shifts,
              int Ny, int Nx, const float* __restrict__ psf, int ph, int pw, int do_max, int do_centroid, int in_place)
{
    extern __shared__ __align__(16) float smem_f[];
    __shared__ double red[32];
    __shared__ int sc[2];
    __shared__ double ss[2];
    const int N = Ny * Nx;
    const float* g = imgs + (size_t)blockIdx.x * N;
    const float* X = in_place ? g : smem_f;            // in_place: images beyond the shared-memory limit
    if (!in_place) for (int p = threadIdx.x; p < N; p += blockDim.x) smem_f[p] = g[p];
    if (threadIdx.x == 0) { sc[0] = centers[2 * blockIdx.x]; sc[1] = centers[2 * blockIdx.x + 1]; ss[0] = ss[1] = 0.0; }
    __syncthreads();
    if (sc[0] < 0 || sc[0] >= Ny || sc[1] < 0 || sc[1] >= Nx) return;       // see scl_k_monotonic
    if (do_max) { scl_max_pixel(X, Ny, Nx, sc[0], sc[1], sc); __syncthreads(); }
    if (do_centroid) { int cy = sc[0], cx = sc[1]; __syncthreads(); scl_centroid(X, Ny, Nx, psf, ph, pw, cy, cx, sc, ss, red); }
    if (threadIdx.x == 0) {
        centers[2 * blockIdx.x] = sc[0]; centers[2 * blockIdx.x + 1] = sc[1];
        shifts[2 * blockIdx.x] = ss[0]; shifts[2 * blockIdx.x + 1] = ss[1];
    }
}
