/* Oracle (TEST INFRASTRUCTURE ONLY): plain-C restatement of the two sequential
 * monotonicity sweeps of the reference's native module.
 *
 *   orc_prox_monotonic_f64            /root/reference/scarlet/operators_pybind11.cc:11-25
 *   orc_prox_weighted_monotonic_{f32,f64}   operators_pybind11.cc:27-50 (f32/f64
 *                                     overloads registered at :82-85)
 *
 * The reference file itself needs Eigen + pybind11's removed PYBIND11_PLUGIN
 * macro and cannot be compiled here, hence this restatement.
 * Build: `make -C oracle` -> oracle/_build/liborc_monotonic.so.
 * Compile WITHOUT -ffast-math / FMA contraction (-ffp-contract=off): the
 * reference accumulates `ref += x*w` as a separate multiply and add in type T.
 */
#include <stddef.h>

void orc_prox_monotonic_f64(double *x, const int *ref_idx, const int *dist_idx,
                            int n, double thresh)
{
    for (int d = 0; d < n; d++) {
        int p = dist_idx[d];
        double r = x[ref_idx[p]] * (1 - thresh);
        x[p] = (r < x[p]) ? r : x[p];      /* std::min(x, r) */
    }
}

#define WEIGHTED(NAME, T)                                                       \
void NAME(T *x, const T *w, long n_pix, const int *offsets, int n_off,          \
          const int *dist_idx, int n, T thresh)                                 \
{                                                                               \
    for (int d = 0; d < n; d++) {                                               \
        int p = dist_idx[d];                                                    \
        T ref = 0;                                                              \
        for (int i = 0; i < n_off; i++) {                                       \
            T wi = w[(size_t)i * n_pix + p];                                    \
            if (wi > 0) {                                                       \
                T prod = x[offsets[i] + p] * wi;                                \
                ref += prod;                                                    \
            }                                                                   \
        }                                                                       \
        T r = ref * (1 - thresh);                                               \
        x[p] = (r < x[p]) ? r : x[p];                                           \
    }                                                                           \
}

WEIGHTED(orc_prox_weighted_monotonic_f32, float)
WEIGHTED(orc_prox_weighted_monotonic_f64, double)
