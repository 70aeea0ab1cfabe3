/* scarlet_b200 -- C ABI of the B200-native deblending hot path.
 *
 * This is the drop-in boundary for the reference's native module
 * `scarlet.operators_pybind11` (/root/reference/scarlet/operators_pybind11.cc:72-91,
 * built by setup.py:66-77) widened to the whole `Blend.fit` loop
 * (/root/reference/scarlet/blend.py:65-102).  Plain pointers and sizes only; every
 * function returns 0 on success or a negative SCL_E_* code (scl_last_error() gives
 * the message).  No exception crosses the boundary and there is no CPU fallback:
 * without a CUDA device every compute entry point returns SCL_E_CUDA.
 *
 * Unless a parameter says "device", pointers are HOST pointers; the library does the
 * host<->device copies on `stream` (a cudaStream_t passed as void*, NULL = default).
 */
#ifndef SCARLET_B200_H
#define SCARLET_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define SCL_OK 0
#define SCL_E_ARG (-1)      /* bad argument (maps to ValueError)      */
#define SCL_E_CUDA (-2)     /* CUDA runtime error (maps to RuntimeError) */
#define SCL_E_LIMIT (-3)    /* shape outside the compiled limits (NotImplementedError) */

#define SCL_FFT_EDGE 128    /* convolution FFT edge; needs Ny + ky/2 <= 128, Nx + kx/2 <= 128 */
#define SCL_SPEC_COLS 65    /* columns of a kernel spectrum: rfft2 layout [128][65] complex64 */

/* component flags */
#define SCL_SYMMETRIC 1u    /* source.symmetric (source.py:427) */
#define SCL_MONOTONIC 2u    /* source.monotonic (source.py:435) */
#define SCL_FIX_SED 4u      /* component.fix_sed (blend.py:91) */
#define SCL_FIX_MORPH 8u    /* component.fix_morph (blend.py:94) */
#define SCL_KSPACE 16u      /* owner has `.shift`: k-space symmetry; else soft symmetry, strength 1 (operator.py:337-339) */
#define SCL_CONSTRAINED 32u /* owner runs PointSource.update (source.py:402-440); plain Component.update is a no-op */
#define SCL_SHARED_CENTER 64u /* component of a MultiComponentSource (source.py:605-641): max_pixel / centroid are
                               * measured once per source on sum_k morph_k * sum(sed_k) and shared by its components
                               * (which carry no `.shift`: soft symmetry); needs scl_batch_desc.src_first */

int scl_version(void);
const char* scl_last_error(void);
int scl_device_count(void);
/* The CUDA device of the calling host thread (cudaGetDevice / cudaSetDevice).  The device is a per-thread setting:
 * a host thread that drives a batch created by another thread must select that thread's device first. */
int scl_device_get(int* device);
int scl_device_set(int device);
/* A CUDA stream for the `void* stream` arguments below (for callers without their own CUDA binding, e.g. ctypes).
 * Blocking flavour: ordered against the legacy default stream, so events recorded there bracket its work. */
int scl_stream_create(void** stream);
int scl_stream_destroy(void* stream);

/* ---- operator tables --------------------------------------------------------------
 * Monotonicity weights depend only on the pixel's offset from the peak (plus the image
 * border).  The host builds them once per image shape with the reference's formula
 * (operator.py:540-621) and registers them here:
 *   wtab  float32 [9][(2Ny-1)*(2Nx-1)*8]  normalised weights per border type bt = by*3+bx (0 = interior pixel,
 *                                          by/bx: 1 = first row/column, 2 = last): taps leaving the image dropped
 *   ctab  float64 [(2Ny-1)*(2Nx-1)*8]     un-normalised cosines, 0 where the tap is not strictly closer
 * indexed by ((Y+Ny-1)*(2Nx-1) + (X+Nx-1))*8 + tap, (Y,X) = pixel - peak. */
int scl_set_monotonic_tables(int Ny, int Nx, const float* wtab, const double* ctab);

/* ---- array-level operators (host arrays, updated in place) -------------------------
 * n images of shape [Ny][Nx]; centers int32 [n][2] = (y, x) of the peak. */

/* operators_pybind11.prox_weighted_monotonic, f32 / f64 overloads (operators_pybind11.cc:27-50,82-85)
 * with the weights, offsets and radius ordering of operator.prox_strict_monotonic(use_nearest=False)
 * (operator.py:116-122) generated on the device. */
int scl_prox_weighted_monotonic_f32(float* x, int n, int Ny, int Nx, const int32_t* centers, float thresh, void* stream);
int scl_prox_weighted_monotonic_f64(double* x, int n, int Ny, int Nx, const int32_t* centers, double thresh, void* stream);
/* operators_pybind11.prox_monotonic (operators_pybind11.cc:11-25), f64 only, nearest-neighbour reference
 * pixels of operator.prox_strict_monotonic(use_nearest=True) (operator.py:105-115). */
int scl_prox_monotonic_f64(double* x, int n, int Ny, int Nx, const int32_t* centers, double thresh, void* stream);
/* operator.prox_uncentered_symmetry (operator.py:291-350).  mode 0 = "kspace" (shifts float64 [n][2] = (dy,dx),
 * operator.py:253-288), 1 = "soft" (strength), 2 = "sdss".  The window of operator.uncentered_operator
 * (operator.py:175-228) with fill=None is applied on the device; a center of (-1,-1) selects the whole array. */
int scl_prox_symmetry_f32(float* x, int n, int Ny, int Nx, const int32_t* centers, const double* shifts,
                          int mode, float strength, void* stream);
/* measurement.max_pixel (measurement.py:3-29) and/or measurement.psf_weighted_centroid (measurement.py:32-94).
 * centers in/out; shifts out (float64 [n][2]); psf float32 [ph][pw]. */
int scl_measure_f32(const float* x, int n, int Ny, int Nx, int32_t* centers, double* shifts,
                    const float* psf, int ph, int pw, int do_max_pixel, int do_centroid, void* stream);
/* Observation.render (observation.py:203-220): out = convolve(model, diff kernel) per band.
 * model/out float32 [n][C][Ny][Nx]; spec complex64 [n_ker][128][65] (rfft2 of the truncated difference kernel
 * with its centre tap at index 0); ker_idx int32 [n*C] (-1 = identity).  adjoint != 0 applies the transpose.
 * ker_half: reach (k/2) of the kernels; frames with Ny + ker_half > 128 run as overlap-save tiles (0 = not given:
 * single-tile frames only). */
int scl_convolve_f32(const float* model, float* out, int n, int C, int Ny, int Nx, const float* spec, int n_ker,
                     const int32_t* ker_idx, int adjoint, int ker_half, void* stream);
/* LowResObservation._render (observation.py:524-558) as its matrix: out[i][c][q] = sum_p op[c][q][p] model[i][c][p].
 * model float32 [n][C][N] (the observation's bands of the model, N = Ny*Nx); op float32 [C][n_lr][N], the
 * resample-and-convolve operator of one geometry (built once by LowResObservation.match); out float32 [n][C][n_lr]. */
int scl_lowres_render_f32(const float* model, float* out, int n, int C, int N, int n_lr, const float* op, void* stream);
/* fft.match_psfs (fft.py:282-301) as Observation.match calls it (observation.py:155-196), for n PSF sets at once:
 * diff[i][c] = centred crop of fftshift(irfft2(rfft2(psf_obs[i][c]) / rfft2(psf_model[i or 0]))) at the FFT shape
 * next_fast_len(k_obs + k_model + padding) per axis (last axis even), float64 arithmetic on the device.
 * psf_obs float32 [n][C][ky][kx]; psf_model float32 [n_model][kym][kxm], n_model = 1 (shared) or n;
 * diff float32 [n][C][ky][kx] or NULL; spec complex64 [n][C][128][65] or NULL: the spectrum of the truncated float32
 * kernel placed for an Ny x Nx frame (what scl_batch_desc.kernel_spec and scl_convolve_f32 take). */
int scl_match_psfs_f32(const float* psf_obs, int n, int C, int ky, int kx, const float* psf_model, int n_model,
                       int kym, int kxm, int padding, float* diff, float* spec, int Ny, int Nx, void* stream);
/* bbox.trim (bbox.py:170-191), bbox.flux_at_edge (bbox.py:194-210) and the pixel sum of n images at once:
 * x float32 [n][Ny][Nx]; bounds int32 [n][4] = (bottom, top, left, right) of the pixels > min_value, -1 when there
 * is none (the reference raises ValueError there); edge int32 [n] = any border pixel > min_value; sum float64 [n].
 * Any output may be NULL. */
int scl_trim_f32(const float* x, int n, int Ny, int Nx, float min_value, int32_t* bounds, int32_t* edge, double* sum,
                 void* stream);
/* update.sparse_l0 / sparse_l1 (update.py:71-82; proxmin prox_hard / prox_soft) in place on n float32 values:
 * mode 0: x[|x| < thresh] = 0;  mode 1: x = sign(x) * max(|x| - thresh, 0).  thresh = the caller's thresh * step. */
int scl_prox_sparse_f32(float* x, long long n, float thresh, int mode, void* stream);
/* measurement.threshold (measurement.py:96-113) and update.threshold (update.py:85-103) for n images at once:
 * thresh float64 [n] = 10^(left edge of the last empty bin of the log10 histogram of the positive pixels), 0 when no
 * bin is empty; bins int32 [n] = the bin count used (50, or count/10 below 500 positive pixels).  apply != 0 also
 * zeroes the pixels below the cut in x (float32 [n][Ny][Nx], in place) and returns the box of what is left in
 * bounds int32 [n][4] (bottom, top, left, right; -1 = nothing). */
int scl_threshold_f32(float* x, int n, int Ny, int Nx, int apply, double* thresh, int32_t* bins, int32_t* bounds,
                      void* stream);
/* interpolation.fft_resample (interpolation.py:408-448) / update.translation (update.py:159-167) for n images at once:
 * out[i][y][x] = sum_a sum_b ky[i][a] kx[i][b] x[i][y - y0[i] - a][x - x0[i] - b], zero outside the image: the
 * separable interpolation kernel (lanczos, bilinear, cubic / quintic splines; Ly, Lx <= 16 taps, float64) whose first
 * tap sits at window offset (y0, x0).  x, out float32 [n][Ny][Nx]. */
int scl_resample_f32(const float* x, float* out, int n, int Ny, int Nx, const double* ky, const int32_t* y0, int Ly,
                     const double* kx, const int32_t* x0, int Lx, void* stream);
/* The same operators on caller-owned DEVICE arrays (SURVEY 8(b) "Ownership": no hidden host<->device copies): every
 * pointer is a device pointer, the call is ordered on `stream` and returns without synchronising.  Centres cannot be
 * validated on the host; an image whose centre lies outside the frame is left untouched.  Single-tile frames for the
 * convolution (Ny, Nx <= 128 - ker_half); the kernel spectra are the device copy of what scl_convolve_f32 takes. */
int scl_prox_weighted_monotonic_f32_dev(float* x_dev, int n, int Ny, int Nx, const int32_t* centers_dev, float thresh, void* stream);
int scl_prox_weighted_monotonic_f64_dev(double* x_dev, int n, int Ny, int Nx, const int32_t* centers_dev, double thresh, void* stream);
int scl_measure_f32_dev(const float* x_dev, int n, int Ny, int Nx, int32_t* centers_dev, double* shifts_dev,
                        const float* psf_dev, int ph, int pw, int do_max_pixel, int do_centroid, void* stream);
int scl_convolve_f32_dev(const float* model_dev, float* out_dev, int n, int C, int Ny, int Nx, const float* spec_dev,
                         int n_ker, const int32_t* ker_idx_dev, int adjoint, void* stream);
/* (the symmetry operator on device arrays: scl_prox_symmetry_noalloc_f32 below) */

/* No counterpart in the reference: added for callers that may not allocate
 * device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * scl_prox_symmetry_f32 on caller-owned DEVICE arrays, stream-ordered, no host<->device copy, no synchronisation and no
 * cudaMalloc / cudaFree / allocating library call: the working copies and the contraction scratch live in one
 * file-scope `static __device__` array of SCL_NOALLOC_MAX_BYTES that every GPU which loads the library holds until
 * exit.  x_dev float32 [n][Ny][Nx] in place; centers_dev int32 [n][2] ((-1,-1) = whole array); shifts_dev float64
 * [n][2] (mode 0 only, else may be NULL); modes as scl_prox_symmetry_f32.  An image whose centre lies outside the
 * frame is left untouched and n_bad_dev (device int32, may be NULL; the caller zeroes it) is incremented -- the host
 * cannot validate device-resident centres without a copy.  Returns SCL_E_LIMIT when
 *   4 * (8256 + n * (roundup4(Ny*Nx + 8*Nx + 16) + roundup4(Ny*roundup4(Nx) + 16))) > SCL_NOALLOC_MAX_BYTES
 * (e.g. up to 479 images of 64x64, 24 of 300x280, one of 1024x1024).  The first call on a device looks up the array's
 * device address (cudaGetSymbolAddress) and copies two small tables into it: make that call once outside any capture.
 * Every call shares the one array, which keeps the previous call's contents (the kernel re-initialises what it
 * reads): calls that may overlap on the GPU -- different streams, or several host threads -- race; serialise them. */
#define SCL_NOALLOC_MAX_BYTES (16u << 20)
int scl_prox_symmetry_noalloc_f32(float* x_dev, int n, int Ny, int Nx, const int32_t* centers_dev,
                                  const double* shifts_dev, int mode, float strength, int32_t* n_bad_dev, void* stream);
/* ---- the reference's native module, entry for entry, on the CALLER's tables -------------------------------------
 * operators_pybind11.prox_weighted_monotonic(flat_img, step, weights, offsets, dist_idx, thresh)
 * (operators_pybind11.cc:27-50,82-85; caller operator.py:32-37): x [N] in place; weights [n_taps][N] row-major in x's
 * type (the pybind layer casts the caller's table to it, operators_pybind11.cc:82-85); offsets int32 [n_taps <= 8];
 * dist_idx int32 [M], walked in the given order: x[d] = min(x[d], (1 - thresh) * sum_{i: w[i,d] > 0} x[d + off[i]] w[i,d]),
 * accumulated in tap order in x's type.  Unlike the C++ (no bounds checks) an index that leaves the image is SCL_E_ARG. */
int scl_pb_prox_weighted_monotonic_f32(float* x, long long N, const float* weights, const int32_t* offsets, int n_taps,
                                       const int32_t* dist_idx, long long M, float thresh, void* stream);
int scl_pb_prox_weighted_monotonic_f64(double* x, long long N, const double* weights, const int32_t* offsets, int n_taps,
                                       const int32_t* dist_idx, long long M, double thresh, void* stream);
/* operators_pybind11.prox_monotonic(X, step, ref_idx, dist_idx, thresh) (operators_pybind11.cc:11-25; caller
 * operator.py:24-29): x[d] = min(x[d], x[ref_idx[d]] * (1 - thresh)) for d in dist_idx, float64 only. */
int scl_pb_prox_monotonic_f64(double* x, long long N, const int32_t* ref_idx, long long n_ref, const int32_t* dist_idx,
                              long long M, double thresh, void* stream);
/* operators_pybind11.apply_filter(image, values, y_start, y_end, x_start, x_end, result) (operators_pybind11.cc:52-70;
 * no caller in the reference): result = sum_n values[n] * image[y_end[n] : y_end[n] + r_n, x_end[n] : x_end[n] + c_n]
 * added into result[y_start[n] : y_start[n] + r_n, x_start[n] : x_start[n] + c_n], r_n = rows - y_start[n] - y_end[n],
 * c_n likewise; image / result [rows][cols] row-major. */
int scl_apply_filter_f32(const float* image, int rows, int cols, const float* values, const int32_t* y_start,
                         const int32_t* y_end, const int32_t* x_start, const int32_t* x_end, int n, float* result,
                         void* stream);
int scl_apply_filter_f64(const double* image, int rows, int cols, const double* values, const int32_t* y_start,
                         const int32_t* y_end, const int32_t* x_start, const int32_t* x_end, int n, double* result,
                         void* stream);

/* ---- Blend.fit on a batch of independent blends ------------------------------------ */
typedef struct scl_batch scl_batch;

typedef struct {
    int32_t n_blends, n_comp;       /* B, Ktot */
    int32_t C, Ny, Nx;              /* frame shape shared by the batch */
    const int32_t* comp_off;        /* [B+1] component range of each blend */
    const float* sed;               /* [Ktot][C] */
    const float* morph;             /* [Ktot][Ny*Nx]; NULL = zeros (morphologies come from scl_batch_init_extended) */
    const uint32_t* cflags;         /* [Ktot] SCL_* flags */
    const int32_t* center;          /* [Ktot][2] pixel_center (y, x) */
    const double* shift;            /* [Ktot][2] (dy, dx) */
    const int32_t* conv;            /* [Ktot] BlendFlag bits (component.py:32-36) */
    const float* images;            /* [B][P][Ny*Nx]; P = n_planes (C when n_planes == 0) */
    const float* weights;           /* [B][P][Ny*Nx] or NULL (weights = 1, observation.py:148-151) */
    int32_t n_ker;                  /* number of distinct kernel spectra */
    const float* kernel_spec;       /* complex64 [n_ker][128][65] */
    const int32_t* ker_idx;         /* [B*P] index into kernel_spec, -1 = identical PSFs (observation.py:191-194) */
    int32_t n_cw, ph, pw;           /* centroid weight images (source.py:389-397) */
    const float* cw;                /* [n_cw][ph][pw] */
    const int32_t* cw_idx;          /* [B] */
    const int32_t* it0;             /* [B] Blend.it before this fit (len(mse)) */
    const double* prev_loss;        /* [B] mse[-1] before this fit (ignored where it0 == 0) */
    const int32_t* src_first;       /* [Ktot] or NULL: index of the first component of the source that owns component
                                     * k (the components of a source are contiguous, ComponentTree order); only read
                                     * for components flagged SCL_SHARED_CENTER.  NULL = no such component */
    /* Several observations per blend (Blend._loss sums them, blend.py:135-139; each sees the model bands of its
     * `_band_slice`, observation.py:184-189,216): the image planes of all observations are stacked, plane p of
     * every blend looks at model band plane_band[p].  n_planes == 0: one observation covering all C bands
     * (P = C, plane_band[p] = p, n_obs = 1).  n_obs scales both Lipschitz constants (blend.py:214-218). */
    int32_t n_planes;               /* P <= 8 */
    const int32_t* plane_band;      /* [P] model band of each plane */
    int32_t n_obs;                  /* number of observations the planes come from (>= 1) */
    /* Planes of low-resolution observations (LowResObservation, observation.py:242-599), after the n_planes
     * full-resolution ones: plane_band then has n_planes + n_lr_planes entries (<= 8 together; at least one
     * full-resolution plane).  Each sees its model band through a linear resample-and-convolve operator. */
    int32_t n_lr_planes;            /* Plr (0 = none) */
    int32_t lr_n;                   /* pixels of the low-resolution window, n_lr */
    const float* lr_images;         /* [B][Plr][n_lr] */
    const float* lr_weights;        /* [B][Plr][n_lr] or NULL */
    int32_t n_lr_op;                /* distinct geometries */
    const float* lr_op;             /* [n_lr_op][Plr][n_lr][Ny*Nx] */
    const int32_t* lr_op_idx;       /* [B] */
    /* Scene mode: frames larger than one 128x128 convolution tile, or blends with more than 8 components
     * (BASELINE config 4).  The model and the residual go through HBM, the convolution runs as overlap-save tiles of
     * 128 - 2*ker_half pixels, and the constraint kernels keep their working copies in an HBM scratch.  Chosen
     * automatically when the shape or a component count needs it; scene_mode = 1 forces it (tests).  Needs
     * ker_half when any ker_idx >= 0; no low-resolution planes yet. */
    int32_t ker_half;               /* half-width (k/2) of the truncated difference kernels; 0 = not given */
    int32_t scene_mode;             /* 0 = automatic, 1 = force scene mode */
    /* PSFs instead of spectra (Observation.match inside the batch, observation.py:155-196): when psf_obs is not NULL
     * kernel_spec is ignored and the n_ker = psf_sets * P spectra are built on the device (as scl_match_psfs_f32 does)
     * and stay there: spectrum s * P + p is psf_obs[s][p] matched to psf_model[s] (or [0] when psf_model_sets == 1).
     * ker_half is derived from the stamp size when it is 0. */
    const float* psf_obs;           /* [psf_sets][P][psf_ky][psf_kx] or NULL */
    const float* psf_model;         /* [psf_model_sets][psf_kym][psf_kxm] */
    int32_t psf_sets, psf_model_sets, psf_ky, psf_kx, psf_kym, psf_kxm;
    /* Declared schedules: a component flagged SCL_SCHEDULE runs, instead of the built-in
     * PointSource.update, the ordered list of scarlet.update / scarlet.measurement calls its source's own update()
     * makes: sched_ops [Ktot][16] int32 op codes (SCL_OP_*, 0 ends the list), sched_args [Ktot][16] float32 (the
     * op's parameter).  Either pointer NULL = no scheduled component.  Such batches run in scene mode. */
    const int32_t* sched_ops;
    const float* sched_args;
    /* Priors (component.py:39-67,177-187) of the affine form grad_sed = a_s sed + b_s, grad_morph = a_m morph + b_m:
     * prior_par float32 [Ktot][4] = (a_s, a_m, L_sed, L_morph), L_sed < 0 = the component has no prior;
     * prior_b_sed float32 [Ktot][C]; prior_idx int32 [Ktot] row of prior_b_morph float32 [n_prior_morph][Ny*Nx]
     * (-1 = b_m = 0).  prior_par == NULL = no priors.  Such batches run in scene mode. */
    const float* prior_par;
    const float* prior_b_sed;
    const int32_t* prior_idx;
    const float* prior_b_morph;
    int32_t n_prior_morph;
} scl_batch_desc;

/* op codes of a declared schedule (reference file:line of each in csrc/k_prox.cu) */
#define SCL_SCHEDULE 128u
#define SCL_SCHED_OPS 16
#define SCL_OP_POSITIVE_SED 1
#define SCL_OP_POSITIVE_MORPH 2
#define SCL_OP_NORM_SED 3
#define SCL_OP_NORM_MORPH 4
#define SCL_OP_NORM_MORPH_MAX 5
#define SCL_OP_SPARSE_L0 6
#define SCL_OP_SPARSE_L1 7
#define SCL_OP_THRESHOLD 8
#define SCL_OP_MONOTONIC 9
#define SCL_OP_SYM_KSPACE 10
#define SCL_OP_SYM_SOFT 11
#define SCL_OP_SYM_SDSS 12
#define SCL_OP_MAX_PIXEL 13
#define SCL_OP_CENTROID 14

/* Allocate device state for `desc`'s sizes and upload it. */
int scl_batch_create(const scl_batch_desc* desc, scl_batch** out, void* stream);
/* Re-upload inputs of identical sizes into an existing batch (no allocation).  kernel_spec == NULL and
 * psf_obs == NULL (with the same n_ker) keep the kernel spectra the batch already holds. */
int scl_batch_upload(scl_batch* batch, const scl_batch_desc* desc, void* stream);
/* Blend.fit(max_iter, e_rel, approximate_L) (blend.py:65-102) for every blend of the batch.
 * check_every: how often (iterations) the host polls the device-side active count to stop early.
 * device_ms (optional): device time of the whole loop by CUDA events on `stream`.
 * kernel_ms (optional, float[4]): summed device time of each of the four kernels of an iteration
 *   [0] render/convolve/residual/adjoint  [1] gradients + Lipschitz + step  [2] constraint chain  [3] convergence
 * by CUDA events between the launches (only measured when profile != 0); steps_run: iterations launched
 * (= launches of each kernel); launches: total kernel launches. */
int scl_batch_fit(scl_batch* batch, int max_iter, double e_rel, int approximate_L, int check_every,
                  int profile, float* device_ms, float* kernel_ms, int* steps_run, int* launches, void* stream);
/* The constraint pass every source runs once at construction (`self.update()` with it = 0, source.py:400,492):
 * max_pixel, centroid + shift, symmetry, monotonicity, positivity, normalisation on the batch's current state. */
int scl_batch_constrain(scl_batch* batch, void* stream);
/* init_extended_source (source.py:139-180) for every component of the batch, on the device, from the batch's own
 * images and centres, in float32: detection coadd with per-component band weights cw [Ktot][C] (sed/bg_rms^2 on the
 * bands with sed > 0, else 0) and jacobian jac [Ktot] (source.py:101-136), sdss symmetry, monotonicity (thresh .1),
 * noise cut at cutoff [Ktot] and peak normalisation; overwrites the batch's morphologies.  ok [Ktot] (host, out):
 * 0 where no pixel survives the cut (the reference raises SourceInitError, source.py:171-173). */
int scl_batch_init_extended(scl_batch* batch, const float* cw, const float* jac, const float* cutoff,
                            int symmetric, int monotonic, int32_t* ok, void* stream);
/* The same for a mix of source kinds, kind int32 [Ktot] per component (NULL = all 0):
 *   0  ExtendedSource: as scl_batch_init_extended.
 *   1  a component of a MultiComponentSource (init_multicomponent_source, source.py:242-291): the source's first
 *      component takes the extended initialisation, the base morphology is then cut into len(percentiles) + 1 flux
 *      layers at percentiles[i] % of its maximum (float32 [n_perc], ascending, 1 <= n_perc <= 3; the reference's
 *      default is [25]), each layer peak-normalised, and the SEDs of the layers are the least-squares solution of
 *      images = A^T layers (get_best_fit_seds, source.py:74-98), written over the batch's SEDs.  The components of
 *      such a source are the SCL_SHARED_CENTER group of src_first; cw / jac / cutoff are read at its first component.
 *   2  PointSource (source.py:355-375): the model PSF image (the blend's centroid-weight image cw) projected onto
 *      the frame around pixel_center; the SED as uploaded.
 * Works in scene mode too (the working copy of the coadd then lives in the HBM scratch). */
int scl_batch_init_sources(scl_batch* batch, const int32_t* kind, const float* cw, const float* jac, const float* cutoff,
                           int symmetric, int monotonic, const float* percentiles, int n_perc, int32_t* ok, void* stream);
/* Copy results back.  Any pointer may be NULL.  logL: [B][max_iter of the last fit] (Blend.mse entries appended
 * by that fit; n_iter[b] of them valid); active[b] == 0 <=> blend b converged. */
int scl_batch_download(scl_batch* batch, float* sed, float* morph, int32_t* center, double* shift, int32_t* conv,
                       int32_t* n_iter, int32_t* active, double* logL, float* lipschitz, void* stream);
/* Keep / bring back a device-side copy of the model state (sed, morph, center, shift, flags), so that a fit can be
 * repeated from the same starting point without another host upload (stream-ordered device-to-device copies). */
int scl_batch_snapshot(scl_batch* batch, void* stream);
int scl_batch_restore(scl_batch* batch, void* stream);
/* Post-fit products of every component, from the device state:
 *   flux   float32 [Ktot][C]  morph.sum() * sed                       (Component.get_flux, component.py:172-175)
 *   bounds int32   [Ktot][4]  as scl_trim_f32 on the morphology        (bbox.trim, update.py:97-102)
 *   flags  int32   [Ktot]     BlendFlag bits (component.py:13-36): 1 SED_NOT_CONVERGED, 2 MORPH_NOT_CONVERGED as left by
 *                             the fit, 4 EDGE_PIXELS = flux_at_edge(morph, min_value), 8 NO_VALID_PIXELS = no pixel
 *                             above min_value.  Any output may be NULL. */
int scl_batch_products(scl_batch* batch, float min_value, float* flux, int32_t* bounds, int32_t* flags, void* stream);
/* Observation.render(component.get_model()) (observation.py:198-220, component.py:148-170) for the components
 * [k0, k0 + nk): out float32 [nk][n_planes][Ny][Nx], plane p = sed[k, plane_band[p]] * morph_k convolved with the
 * difference kernel of that plane of the component's blend.  Summed over the components of a blend this is the
 * rendered model of the loop. */
int scl_batch_render_components(scl_batch* batch, int k0, int nk, float* out, void* stream);
/* Profiling aid: while enabled, the constraint kernel adds the clock64 ticks each of its phases took (summed over
 * CTAs: 0 gradient step, 1 centre + symmetry, 2 level list, 3 monotonic sweep, 4 normalise/write-back) to a device
 * counter; slots 5, 6, 7 hold the longest symmetry phase, sweep and CTA seen; the call returns the values so far in
 * out8 (may be NULL) and clears them. */
int scl_batch_phase_clocks(scl_batch* batch, int enable, unsigned long long* out8);
/* How the host waits for the batch's stream (polls of the active count, end of fit, copies): 0 (default) spins on
 * the stream -- lowest latency for one batch driven by one thread; 1 sleeps on a blocking-sync event, for hosts where
 * many ranks and chunk threads share the cores (a spinning waiter per stream starves the staging copies). */
int scl_batch_wait_mode(scl_batch* batch, int blocking);
/* Test switches (no counterpart in the reference).  bit 0: the scene-mode monotonic sweep visits every level instead of
 * stopping at the first three consecutive levels without a positive pixel (the results are bit-identical; the
 * tests compare the two).  bit 1: scene mode forms S S^T by the dense tiled contraction instead of per pair of
 * overlapping footprint boxes (same products; the sums differ in their last bits).  bit 2 / bit 3: low-resolution planes
 * always by the per-blend kernels / always by the kernels blocked over 64 blends (default: blocked from 16 blends of
 * one geometry on).  bit 4: scene mode runs a component's whole constraint schedule in one CTA instead of the chain of
 * kernels that spreads its k-space symmetry products over the GPU (csrc/k_prox_split.cu).  bit 5: the scene-mode sweep
 * reads every pixel's eight weights from the offset-indexed table instead of the four that can be non-zero from the
 * level-ordered copy (bit-identical results; the tests compare the two).  bit 6: scene mode with exact constants runs
 * the Gram chain in line on the caller's stream instead of beside the render / convolution kernels on a stream of the
 * batch's own, and the constraint schedule as one chain of kernels instead of two (components with / without the
 * k-space symmetry). */
int scl_batch_debug_flags(scl_batch* batch, int flags);
void scl_batch_destroy(scl_batch* batch);

/* ---- Blend.fit in float64 (compatibility path) -----------------------------------------------------------------------
 * The reference's own loop tests build Frame(dtype=float64) (tests/test_blend.py:59-63,79-83,100-104).  One blend per
 * call, host arrays in and out, every step of blend.py:79-102 as a plain float64 kernel (csrc/k_f64.cu); the
 * difference-kernel convolution is the direct real-space sum the reference's zero-padded FFT product equals
 * (fft.py:264-317).  Same flags, planes and schedules as scl_batch_desc; no low-resolution planes. */
typedef struct {
    int32_t n_comp, C, Ny, Nx;
    const double* sed;              /* [K][C] */
    const double* morph;            /* [K][Ny*Nx] */
    const uint32_t* cflags;         /* [K] SCL_* flags */
    const int32_t* center;          /* [K][2] */
    const double* shift;            /* [K][2] */
    const int32_t* conv;            /* [K] BlendFlag bits */
    const int32_t* src_first;       /* [K] or NULL (SCL_SHARED_CENTER components) */
    int32_t n_planes;               /* P <= 8 image planes (bands of all observations stacked) */
    const int32_t* plane_band;      /* [P] */
    int32_t n_obs;
    const double* images;           /* [P][Ny*Nx] */
    const double* weights;          /* [P][Ny*Nx] or NULL */
    const int32_t* ker_on;          /* [P] 0 = identical PSFs (observation.py:191-194) */
    const double* kernels;          /* [P][ky][kx] truncated real-space difference kernels (Observation._diff_kernels) */
    int32_t ky, kx;
    int32_t offy, offx;             /* kernel tap (j, i) acts at displacement (j + offy, i + offx):
                                     * off = (F - k + 1) / 2 - F / 2 with F the reference's FFT length (fft.py:60-65,87-106) */
    const double* cw;               /* [ph][pw] centroid weights (source.py:389-397) or NULL */
    int32_t ph, pw;
    int32_t it0;                    /* Blend.it before this fit */
    double prev_loss;               /* mse[-1] before this fit */
} scl_fit64_desc;
/* Blend.fit(max_iter, e_rel, approximate_L) (blend.py:65-102).  Outputs (any may be NULL except n_iter): final sed / morph /
 * center / shift / conv, n_iter = iterations run, converged, logL [max_iter] (n_iter valid), lipschitz [2]. */
int scl_fit_f64(const scl_fit64_desc* desc, int max_iter, double e_rel, int approximate_L, double* sed, double* morph,
                int32_t* center, double* shift, int32_t* conv, int32_t* n_iter, int32_t* converged, double* logL,
                double* lipschitz, void* stream);
/* The constraint pass every source runs once at construction (`self.update()` with it = 0, source.py:400,492) in
 * float64, for the components of desc (images / weights / kernels are not read and may be NULL; n_planes >= 1). */
int scl_constrain_f64(const scl_fit64_desc* desc, double* sed, double* morph, int32_t* center, double* shift, void* stream);
/* Observation.render (observation.py:203-220) in float64: model / out [n][C][Ny][Nx], kernels [C][ky][kx] with the
 * placement (offy, offx) of scl_fit64_desc; adjoint != 0 applies the transpose. */
int scl_convolve_f64(const double* model, double* out, int n, int C, int Ny, int Nx, const double* kernels, int ky, int kx,
                     int offy, int offx, int adjoint, void* stream);

#ifdef __cplusplus
}
#endif
#endif
