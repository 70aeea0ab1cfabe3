// Shared helpers for the scarlet_b200 CUDA kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define SCL_MAX_BANDS 8          // C <= 8 spectral channels per frame
#define SCL_MAX_COMP 8           // K <= 8 components per blend (register accumulators, 8x8 Gram)
#define SCL_FFT 128              // smem FFT edge (power of two); Ny+ky/2 <= 128, Nx+kx/2 <= 128
#define SCL_RS 65                // complex row stride of the smem spectrum tile (odd: conflict-free columns)
#define SCL_THREADS 256
#define SCL_PROX_THREADS 128     // constraint kernel: one CTA of 4 warps per component
#ifndef SCL_PROX_WIDE
#define SCL_PROX_WIDE 512        // the same kernel in scene mode (working set in HBM)
#endif
#ifndef SCL_WIDE_MIN
#define SCL_WIDE_MIN 2           // resident CTAs per SM the wide variant is compiled for
#endif

// component flags (host mirror: scarlet_b200/_native.py)
#define SCL_F_SYMMETRIC 1u       // run the symmetry prox (source.py:427-432)
#define SCL_F_MONOTONIC 2u       // run the monotonicity prox (source.py:435-436)
#define SCL_F_FIX_SED   4u       // skip the SED gradient step (blend.py:91)
#define SCL_F_FIX_MORPH 8u       // skip the morphology gradient step (blend.py:94)
#define SCL_F_KSPACE    16u      // component owner has a `.shift` -> k-space symmetry; else soft, strength 1
#define SCL_F_CONSTRAINED 32u    // run max_pixel / positive / normalized (PointSource.update); plain Components: no-op
#define SCL_F_SCHEDULE 128u      // the component runs a declared schedule (SclProxArgs::sched_ops) instead of the built-in one
#define SCL_F_SHARED_CENTER 64u  // component of a multi-component source: centre + shift come from the source's
                                 // flux-weighted morphology (MultiComponentSource.update, source.py:605-627)

// Image planes of a blend: the bands of all its observations stacked; plane p sees model band band[p]
// (Observation._band_slice, observation.py:184-189).  Passed to the kernels by value.
struct SclPlanes {
    int P;                       // planes per blend (<= SCL_MAX_BANDS)
    int band[SCL_MAX_BANDS];
};

// BlendFlag bits (component.py:32-36)
#define SCL_SED_NOT_CONVERGED 1
#define SCL_MORPH_NOT_CONVERGED 2

// Complex arithmetic on the packed-f32x2 instructions of sm_100 (FADD2 / FMUL2 / FFMA2: one issue slot for both
// halves; ptxas folds scalar broadcasts and half swaps into operand modifiers).  add/sub round exactly like the
// scalar forms; the products use the same mul-then-fma pairing the compiler contracted the scalar forms to.
#ifndef SCL_PACKED
#define SCL_PACKED 1
#endif
__device__ __forceinline__ unsigned long long scl_pk(float a, float b) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float2 scl_upk(unsigned long long v) {
    float2 r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
    return r;
}
__device__ __forceinline__ unsigned long long scl_fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
#if SCL_PACKED
// a * b = a.x (b.x, b.y) + a.y (-b.y, b.x)
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    unsigned long long t, r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(t) : "l"(scl_pk(a.x, a.x)), "l"(scl_pk(b.x, b.y)));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(scl_pk(a.y, a.y)), "l"(scl_pk(-b.y, b.x)), "l"(t));
    return scl_upk(r);
}
// a * conj(b) = a.x (b.x, -b.y) + a.y (b.y, b.x)
__device__ __forceinline__ float2 cmulc(float2 a, float2 b) {
    unsigned long long t, r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(t) : "l"(scl_pk(a.y, a.y)), "l"(scl_pk(b.y, b.x)));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(scl_pk(a.x, a.x)), "l"(scl_pk(b.x, -b.y)), "l"(t));
    return scl_upk(r);
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) {
    unsigned long long r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(scl_pk(a.x, a.y)), "l"(scl_pk(b.x, b.y)));
    return scl_upk(r);
}
__device__ __forceinline__ float2 csub(float2 a, float2 b) {
    unsigned long long r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(scl_pk(a.x, a.y)), "l"(scl_pk(b.x, b.y)));
    return scl_upk(r);
}
#else
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cmulc(float2 a, float2 b) {   // a * conj(b)
    return make_float2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y);
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
#endif
__device__ __forceinline__ float2 cswap(float2 a) { return make_float2(a.y, a.x); }

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum (all threads get the result).  `red` holds >= 32 T.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* red) {
    v = warp_sum(v);
    int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (l == 0) red[w] = v;
    __syncthreads();
    T r = (l < nw) ? red[l] : T(0);
    r = warp_sum(r);
    return r;
}
