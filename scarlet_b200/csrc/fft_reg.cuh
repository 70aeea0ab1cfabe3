// Register-resident 128-point FFT lines for the PSF-matching convolution.
//
// Eight lanes (t = lane & 7; four such groups per warp) own one complex line of 128 elements.
// Two data distributions, 16 complex registers per lane each:
//   time:  lane t holds x[8*n1 + t],                 n1 = 0..15   -> register n1
//   freq:  lane t holds X[k1 + 16*k2], k1 = t + 8*s, k2 = 0..7    -> register 8*s + k2
// Forward (time -> freq):   16-point DFT over n1 in registers, twiddle w128^(t*k1), ONE exchange through
// shared memory, 8-point DFT over n2 in registers.  The mirrored routine (freq -> time) runs the same DFT
// with the roles of the index pairs swapped: 8-point DFT over k2, twiddle, the reverse exchange, 16-point DFT
// over k1.  Both compute the forward sign; an inverse transform is obtained by swapping re/im on the way in
// and out (IFFT(x) = swap(FFT(swap(x)))), which is free in registers.
//
// Exchange element (k1, n2) lives in slot 8*k1 + (n2 ^ (k1 & 7)) of the line's own 128 slots (element stride
// `es`: 1 for a row line, RS for a column).  With 64-bit accesses a wavefront serves 16 lanes = two groups; the
// callers map the two groups of a half-warp to lines whose base addresses differ by 8 (mod 16) float2, which
// makes every access of both exchange directions conflict-free (DESIGN.md "FFT core").
#pragma once
#include "fft_smem.cuh"

// 4-point DFT whose inputs 2 and 3 are zero
__device__ __forceinline__ void dft4_in2(float2& a0, float2& a1, float2& a2, float2& a3) {
    const float2 x0 = a0, x1 = a1, m = make_float2(x1.y, -x1.x);       // -i * x1
    a0 = cadd(x0, x1);
    a1 = cadd(x0, m);
    a2 = csub(x0, x1);
    a3 = csub(x0, m);
}
// 4-point DFT of which only outputs 0 and 1 are wanted (a2, a3 left undefined)
__device__ __forceinline__ void dft4_out2(float2& a0, float2& a1, float2& a2, float2& a3) {
    const float2 s02 = cadd(a0, a2), d02 = csub(a0, a2), s13 = cadd(a1, a3), d13 = csub(a1, a3);
    a0 = cadd(s02, s13);
    a1 = cadd(d02, make_float2(d13.y, -d13.x));
}

// 16-point forward DFT, natural order in and out.
// IN8: inputs a[8..15] are zero (zero-padded signal); OUT8: only X[0..7] are wanted (cropped result).
template <bool IN8, bool OUT8>
__device__ __forceinline__ void dft16(float2 (&a)[16]) {
    // step 1: 4-point DFTs over n1' of x[n0 + 4 n1'] -> y[n0][k0] kept at a[n0 + 4 k0]
#pragma unroll
    for (int n0 = 0; n0 < 4; n0++) {
        if (IN8) dft4_in2(a[n0], a[n0 + 4], a[n0 + 8], a[n0 + 12]);
        else dft4(a[n0], a[n0 + 4], a[n0 + 8], a[n0 + 12]);
    }
    // twiddles w16^(n0*k0)
    const float c1 = 0.92387953251128675613f, s1 = 0.38268343236508977173f, h = 0.70710678118654752440f;
    a[1 + 4] = cmul(a[1 + 4], make_float2(c1, -s1));          // w^1
    a[1 + 8] = cmul(a[1 + 8], make_float2(h, -h));            // w^2
    a[1 + 12] = cmul(a[1 + 12], make_float2(s1, -c1));        // w^3
    a[2 + 4] = cmul(a[2 + 4], make_float2(h, -h));            // w^2
    a[2 + 8] = make_float2(a[2 + 8].y, -a[2 + 8].x);          // w^4 = -i
    a[2 + 12] = cmul(a[2 + 12], make_float2(-h, -h));         // w^6
    a[3 + 4] = cmul(a[3 + 4], make_float2(s1, -c1));          // w^3
    a[3 + 8] = cmul(a[3 + 8], make_float2(-h, -h));           // w^6
    a[3 + 12] = cmul(a[3 + 12], make_float2(-c1, s1));        // w^9
    // step 2: 4-point DFTs over n0 -> X[k0 + 4 k1'] at a[4 k0 + k1']
#pragma unroll
    for (int k0 = 0; k0 < 4; k0++) {
        if (OUT8) dft4_out2(a[4 * k0], a[4 * k0 + 1], a[4 * k0 + 2], a[4 * k0 + 3]);
        else dft4(a[4 * k0], a[4 * k0 + 1], a[4 * k0 + 2], a[4 * k0 + 3]);
    }
    // natural order: X[k] sits at a[4 (k & 3) + (k >> 2)]
    float2 b[16];
#pragma unroll
    for (int k = 0; k < 16; k++) b[k] = a[4 * (k & 3) + (k >> 2)];
#pragma unroll
    for (int k = 0; k < 16; k++) a[k] = b[k];
}

// Twiddle tables in shared memory (256 float2): twA[k1*8 + t] = w128^(t*k1) for the forward routine,
// twB[n2*16 + k1] = w128^(k1*n2) for the mirrored one; both are read with consecutive lanes.
__device__ __forceinline__ void fill_twiddles_reg(float2* tw) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        int m;
        if (i < 128) m = (i >> 3) * (i & 7);            // k1 = i >> 3, t = i & 7
        else { int q = i - 128; m = (q >> 4) * (q & 15); }   // n2 = q >> 4, k1 = q & 15
        float s, c;
        sincospif(-(float)(m & 127) / 64.0f, &s, &c);
        tw[i] = make_float2(c, s);
    }
}

// time -> freq.  v in: x[8*n1 + t]; v out: X[t + 8*s + 16*k2] at v[8*s + k2].
// All 32 lanes of the warp call together; `line` may be any valid address for lanes whose data is ignored.
// IN64: x[n] = 0 for n >= 64 (registers 8..15 are ignored).
template <bool IN64>
__device__ __forceinline__ void fft128_t2f(float2 (&v)[16], float2* line, int es, int t, const float2* tw) {
    dft16<IN64, false>(v);
#pragma unroll
    for (int k1 = 1; k1 < 16; k1++) v[k1] = cmul(v[k1], tw[k1 * 8 + t]);
#pragma unroll
    for (int k1 = 0; k1 < 16; k1++) line[(8 * k1 + (t ^ (k1 & 7))) * es] = v[k1];
    __syncwarp();
#pragma unroll
    for (int s = 0; s < 2; s++)
#pragma unroll
        for (int n2 = 0; n2 < 8; n2++) v[8 * s + n2] = line[(8 * (t + 8 * s) + (n2 ^ t)) * es];
    __syncwarp();
    {
        float2 lo[8], hi[8];
#pragma unroll
        for (int i = 0; i < 8; i++) { lo[i] = v[i]; hi[i] = v[8 + i]; }
        dft8(lo);
        dft8(hi);
#pragma unroll
        for (int i = 0; i < 8; i++) { v[i] = lo[i]; v[8 + i] = hi[i]; }
    }
}

// freq -> time (same forward-sign DFT).  v in: X[t + 8*s + 16*k2] at v[8*s + k2]; v out: x[8*n1 + t] at v[n1].
// OUT64: only x[n], n < 64 are wanted (registers 8..15 come back undefined).
template <bool OUT64>
__device__ __forceinline__ void fft128_f2t(float2 (&v)[16], float2* line, int es, int t, const float2* tw) {
    {
        float2 lo[8], hi[8];
#pragma unroll
        for (int i = 0; i < 8; i++) { lo[i] = v[i]; hi[i] = v[8 + i]; }
        dft8(lo);
        dft8(hi);
#pragma unroll
        for (int i = 0; i < 8; i++) { v[i] = lo[i]; v[8 + i] = hi[i]; }
    }
    const float2* twB = tw + 128;
#pragma unroll
    for (int s = 0; s < 2; s++)
#pragma unroll
        for (int n2 = 1; n2 < 8; n2++) v[8 * s + n2] = cmul(v[8 * s + n2], twB[n2 * 16 + t + 8 * s]);
#pragma unroll
    for (int s = 0; s < 2; s++)
#pragma unroll
        for (int n2 = 0; n2 < 8; n2++) line[(8 * (t + 8 * s) + (n2 ^ t)) * es] = v[8 * s + n2];
    __syncwarp();
#pragma unroll
    for (int k1 = 0; k1 < 16; k1++) v[k1] = line[(8 * k1 + (t ^ (k1 & 7))) * es];
    __syncwarp();
    dft16<false, OUT64>(v);
}

// Partner values for the conjugate-symmetry split of a freq-distributed line: r[i] = X[(128 - k) & 127] for the
// k of register i = 8*s + k2, k2 in [K2A, K2B).  X[128 - k] sits in lane (8 - t) & 7 at register
// 8*(1-s) + 7-k2; lane 0 pairs with itself at (8 - k2) & 7 for s = 0 and at 8 + 7-k2 for s = 1.
template <int K2A, int K2B>
__device__ __forceinline__ void fft128_partner(const float2 (&v)[16], float2 (&r)[16], int t) {
    const int src = (threadIdx.x & 24) | ((8 - t) & 7);
#pragma unroll
    for (int s = 0; s < 2; s++)
#pragma unroll
        for (int k2 = K2A; k2 < K2B; k2++) {
            const float2 send = v[8 * (1 - s) + 7 - k2];
            float2 got;
            got.x = __shfl_sync(0xffffffffu, send.x, src);
            got.y = __shfl_sync(0xffffffffu, send.y, src);
            if (t == 0) got = (s == 0) ? v[(8 - k2) & 7] : v[8 + 7 - k2];
            r[8 * s + k2] = got;
        }
}
