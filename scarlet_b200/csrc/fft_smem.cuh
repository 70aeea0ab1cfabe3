// Shared-memory FFT lines for the PSF-matching convolution.
//
// A "line" is a length-N (N = 64 or 128) complex sequence living in shared
// memory with element stride `es`.  16 threads (half a warp) own one line and
// run every Stockham radix-8/4 pass on it; because no other thread touches the
// line, the only synchronisation between the read and the write half of a pass
// is __syncwarp().  Natural order in, natural order out; forward sign only —
// the inverse transform is obtained by swapping re/im at the boundaries
// (IFFT(x) = swap(FFT(swap(x)))).
#pragma once
#include "common.cuh"

__device__ __forceinline__ void dft4(float2& a0, float2& a1, float2& a2, float2& a3) {
    // forward (exp(-2 pi i jk/4)) 4-point DFT, outputs in natural order
    float2 s02 = cadd(a0, a2), d02 = csub(a0, a2);
    float2 s13 = cadd(a1, a3), d13 = csub(a1, a3);
    float2 md13 = make_float2(d13.y, -d13.x);           // -i * d13
    a0 = cadd(s02, s13);
    a2 = csub(s02, s13);
    a1 = cadd(d02, md13);
    a3 = csub(d02, md13);
}

__device__ __forceinline__ void dft8(float2 (&a)[8]) {
    // 8-point forward DFT as 2 x dft4 (even/odd) + twiddles
    float2 e0 = a[0], e1 = a[2], e2 = a[4], e3 = a[6];
    float2 o0 = a[1], o1 = a[3], o2 = a[5], o3 = a[7];
    dft4(e0, e1, e2, e3);
    dft4(o0, o1, o2, o3);
    const float h = 0.70710678118654752440f;
    // o_k *= exp(-2 pi i k / 8)
    o1 = make_float2(h * (o1.x + o1.y), h * (o1.y - o1.x));
    o2 = make_float2(o2.y, -o2.x);
    o3 = make_float2(h * (o3.y - o3.x), -h * (o3.x + o3.y));
    a[0] = cadd(e0, o0); a[4] = csub(e0, o0);
    a[1] = cadd(e1, o1); a[5] = csub(e1, o1);
    a[2] = cadd(e2, o2); a[6] = csub(e2, o2);
    a[3] = cadd(e3, o3); a[7] = csub(e3, o3);
}

// One radix-4 Stockham butterfly: a_r = x[q + s(p + m r)], y[q + s(4p + k)] = b_k w_n^{pk}.
// `twstep` = N/n so that w_n^{p} = tw[p * twstep]; tw has N entries exp(-2 pi i j / N).
struct Bfly4 { float2 b0, b1, b2, b3; int out; };

__device__ __forceinline__ Bfly4 r4_read(const float2* line, int es, int p, int q, int m, int s,
                                         const float2* tw, int twstep, bool live) {
    Bfly4 r;
    int i0 = q + s * p;
    float2 z = make_float2(0.f, 0.f);
    r.b0 = live ? line[(i0) * es] : z;
    r.b1 = live ? line[(i0 + s * m) * es] : z;
    r.b2 = live ? line[(i0 + 2 * s * m) * es] : z;
    r.b3 = live ? line[(i0 + 3 * s * m) * es] : z;
    dft4(r.b0, r.b1, r.b2, r.b3);
    if (twstep > 0 && p > 0) {
        int t = p * twstep;
        r.b1 = cmul(r.b1, tw[t]);
        r.b2 = cmul(r.b2, tw[2 * t]);
        r.b3 = cmul(r.b3, tw[3 * t]);
    }
    r.out = q + s * 4 * p;
    return r;
}
__device__ __forceinline__ void r4_write(float2* line, int es, const Bfly4& r, int s, bool live) {
    if (!live) return;
    line[(r.out) * es] = r.b0;
    line[(r.out + s) * es] = r.b1;
    line[(r.out + 2 * s) * es] = r.b2;
    line[(r.out + 3 * s) * es] = r.b3;
}

// Forward FFT of one line by 16 cooperating threads (t = 0..15, all 32 lanes of
// the warp must call this together; `live` = this half-warp has a real line).
// Inputs with index >= nz are taken as zero (zero-padding never materialised).
template <int N>
__device__ __forceinline__ void fft_line16(float2* line, int es, int t, const float2* tw, int nz, bool live);

template <>
__device__ __forceinline__ void fft_line16<128>(float2* line, int es, int t, const float2* tw, int nz, bool live) {
    // pass 1: n=128, s=1, R=8, m=16; butterfly p=t
    float2 a[8];
#pragma unroll
    for (int r = 0; r < 8; r++) {
        int i = t + 16 * r;
        a[r] = (live && i < nz) ? line[i * es] : make_float2(0.f, 0.f);
    }
    dft8(a);
#pragma unroll
    for (int k = 1; k < 8; k++) a[k] = cmul(a[k], tw[(t * k) & 127]);
    __syncwarp();
    if (live) {
#pragma unroll
        for (int k = 0; k < 8; k++) line[(8 * t + k) * es] = a[k];
    }
    __syncwarp();
    // pass 2: n=16, s=8, R=4, m=4; 32 butterflies, two per thread
    {
        int i0 = t, i1 = t + 16;
        Bfly4 u = r4_read(line, es, i0 >> 3, i0 & 7, 4, 8, tw, 8, live);
        Bfly4 v = r4_read(line, es, i1 >> 3, i1 & 7, 4, 8, tw, 8, live);
        __syncwarp();
        r4_write(line, es, u, 8, live);
        r4_write(line, es, v, 8, live);
        __syncwarp();
    }
    // pass 3: n=4, s=32, R=4, m=1; p=0, q = 0..31
    {
        Bfly4 u = r4_read(line, es, 0, t, 1, 32, tw, 0, live);
        Bfly4 v = r4_read(line, es, 0, t + 16, 1, 32, tw, 0, live);
        __syncwarp();
        r4_write(line, es, u, 32, live);
        r4_write(line, es, v, 32, live);
        __syncwarp();
    }
}

__device__ __forceinline__ void fill_twiddles128(float2* tw) {
    for (int j = threadIdx.x; j < 128; j += blockDim.x) {
        float s, c;
        sincospif(-(float)j / 64.0f, &s, &c);
        tw[j] = make_float2(c, s);
    }
}
