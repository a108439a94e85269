// fft16k3.cuh — third-generation one-CTA-per-transform 16384-point FFT: the v2 decomposition
// (512 threads, N = 32 * 32 * 16, two shared-memory exchanges; reference
// src/dsp/fft_process.cpp:62-97) re-expressed for Blackwell's packed FP32 pipe (FADD2 / FMUL2 /
// FFMA2: two lanes of one 64-bit register pair per instruction).
//
// v2 was ISSUE-bound (profiles/r01_fft16k2_*: 70 % issue-active, FMA pipe 58 %): 86 % of its
// instructions were scalar FP32.  A packed instruction costs the FP32 pipe the same as two scalar
// ones (profiles/microbench/f32x2_tput.cu) but only one issue slot, so the kernel moves to the
// FP32-pipe bound.  Packing is free only when both halves of a register pair go through the same
// instruction stream, so every value is kept as a "two worlds" complex number
//     re = (re_A, re_B), im = (im_A, im_B)
// and the worlds are chosen per phase so that no shuffle is ever needed:
//   phase 1  thread (b, c), DFT32 over a, decimation in time:  worlds = (a even, a odd).
//            Packed DFT16 gives (E[k], O[k]); scalar tail X[k] = E + W32^k O, X[k+16] = E - W32^k O
//            written into the pair (X[k], X[k+16]); packed twiddle (W_N^(t k), W_N^(t (k+16)));
//            scalar 4-byte stores into the pair layout phase 2 wants.
//   phase 2  thread (ka, c), DFT32 over b, worlds = (b even, b odd): same structure; its output
//            pair (kb, kb + 16) IS phase 3's pair of worlds, so it stores packed, in place.
//   phase 3  thread (ka, kb'), DFT16 over c for the two columns kb', kb' + 16 = the two worlds;
//            packed |X|^2; fftshift folded into the store index.
// Exchange buffer: RE plane and IM plane, each 32 x 257 units of 8 bytes (unit = one pair);
// unit(ka, row, c) = ka * 257 + row * 16 + c: conflict-free for every access pattern above
// (audited by tests/cpp/fft_sim3.cpp).
//
// The phase functions are __host__ __device__ so the CPU simulator can run them thread by thread.
#pragma once
#include <stdint.h>

#include "fft16k2.cuh"
#include "fft_core.cuh"

namespace sdrb {

constexpr int F3_N = 16384;
constexpr int F3_THREADS = 512;
constexpr int F3_PITCH = 257;                       // units per ka plane
constexpr int F3_PLANE_UNITS = 32 * F3_PITCH;       // 8224 units = 65792 B per plane
constexpr int F3_E_BYTES = 2 * F3_PLANE_UNITS * 8;  // 131584 B
constexpr int F3_TP = 5 * 512;                      // cpx [j][t]  W_N^(t 2^j)
constexpr int F3_T2 = 16 * 16;                      // pw4 [k][c]  (W_512^(c k), W_512^(c (k + 16)))
constexpr int F3_TABLE_BYTES = F3_TP * 8 + F3_T2 * 16;

// one 64-bit register pair
struct alignas(8) pf {
    float x, y;
};
struct alignas(16) pw4 {  // packed twiddle pair: (re_lo, re_hi, im_lo, im_hi)
    pf re, im;
};
struct c2 {               // two-worlds complex value
    pf re, im;
};

SDRB_HD pf mkp(float x, float y) { pf r; r.x = x; r.y = y; return r; }
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ float2 as_f2(pf a) { return make_float2(a.x, a.y); }
__device__ __forceinline__ pf as_pf(float2 a) { pf r; r.x = a.x; r.y = a.y; return r; }
__device__ __forceinline__ pf padd(pf a, pf b) { return as_pf(__fadd2_rn(as_f2(a), as_f2(b))); }
__device__ __forceinline__ pf psub(pf a, pf b) { return as_pf(__fadd2_rn(as_f2(a), make_float2(-b.x, -b.y))); }
__device__ __forceinline__ pf pmul(pf a, pf b) { return as_pf(__fmul2_rn(as_f2(a), as_f2(b))); }
__device__ __forceinline__ pf pfma(pf a, pf b, pf c) { return as_pf(__ffma2_rn(as_f2(a), as_f2(b), as_f2(c))); }
#else
inline pf padd(pf a, pf b) { return mkp(a.x + b.x, a.y + b.y); }
inline pf psub(pf a, pf b) { return mkp(a.x - b.x, a.y - b.y); }
inline pf pmul(pf a, pf b) { return mkp(a.x * b.x, a.y * b.y); }
inline pf pfma(pf a, pf b, pf c) { return mkp(a.x * b.x + c.x, a.y * b.y + c.y); }
#endif
SDRB_HD pf pneg(pf a) { return mkp(-a.x, -a.y); }
SDRB_HD pf pdup(float v) { return mkp(v, v); }

SDRB_HD c2 c2add(c2 a, c2 b) { c2 r; r.re = padd(a.re, b.re); r.im = padd(a.im, b.im); return r; }
SDRB_HD c2 c2sub(c2 a, c2 b) { c2 r; r.re = psub(a.re, b.re); r.im = psub(a.im, b.im); return r; }
SDRB_HD c2 c2mul_mi(c2 a) { c2 r; r.re = a.im; r.im = pneg(a.re); return r; }
// (a.re + i a.im) * (w.re + i w.im), every factor a pair
SDRB_HD c2 c2mul(c2 a, pf wre, pf wim) {
    c2 r;
    r.re = pfma(a.re, wre, pneg(pmul(a.im, wim)));
    r.im = pfma(a.re, wim, pmul(a.im, wre));
    return r;
}
SDRB_HD c2 c2mulc(c2 a, float wr, float wi) { return c2mul(a, pdup(wr), pdup(wi)); }
// a * (1 - i)/sqrt2 and a * (-1 - i)/sqrt2
SDRB_HD c2 c2mul_w8_1(c2 a) {
    c2 r; r.re = pmul(padd(a.re, a.im), pdup(SDRB_SQRT1_2)); r.im = pmul(psub(a.im, a.re), pdup(SDRB_SQRT1_2)); return r;
}
SDRB_HD c2 c2mul_w8_3(c2 a) {
    c2 r; r.re = pmul(psub(a.im, a.re), pdup(SDRB_SQRT1_2)); r.im = pmul(padd(a.re, a.im), pdup(-SDRB_SQRT1_2)); return r;
}

SDRB_HD void c2fft4(c2& a0, c2& a1, c2& a2, c2& a3) {
    const c2 s02 = c2add(a0, a2), d02 = c2sub(a0, a2);
    const c2 s13 = c2add(a1, a3), d13 = c2mul_mi(c2sub(a1, a3));
    a0 = c2add(s02, s13);
    a2 = c2sub(s02, s13);
    a1 = c2add(d02, d13);
    a3 = c2sub(d02, d13);
}

// 16-point forward DFT of both worlds at once; X[k] is left in v[rev16(k)] like fft16().
SDRB_HD void c2fft16(c2* v) {
    c2fft4(v[0], v[4], v[8], v[12]);
    c2fft4(v[1], v[5], v[9], v[13]);
    c2fft4(v[2], v[6], v[10], v[14]);
    c2fft4(v[3], v[7], v[11], v[15]);
    v[5] = c2mulc(v[5], SDRB_C1_16, -SDRB_S1_16);
    v[9] = c2mul_w8_1(v[9]);
    v[13] = c2mulc(v[13], SDRB_S1_16, -SDRB_C1_16);
    v[6] = c2mul_w8_1(v[6]);
    v[10] = c2mul_mi(v[10]);
    v[14] = c2mul_w8_3(v[14]);
    v[7] = c2mulc(v[7], SDRB_S1_16, -SDRB_C1_16);
    v[11] = c2mul_w8_3(v[11]);
    v[15] = c2mulc(v[15], -SDRB_C1_16, SDRB_S1_16);
    c2fft4(v[0], v[1], v[2], v[3]);
    c2fft4(v[4], v[5], v[6], v[7]);
    c2fft4(v[8], v[9], v[10], v[11]);
    c2fft4(v[12], v[13], v[14], v[15]);
}

// Decimation-in-time tail of a 32-point DFT.  In: v[rev16(k)] = (E[k], O[k]).  Out: z[k] =
// (X[k], X[k + 16]) with X[k] = E[k] + W32^k O[k], X[k + 16] = E[k] - W32^k O[k].  Scalar: it
// crosses the two halves of a pair.
SDRB_HD void f3_dit_tail(const c2* v, c2* z) {
    const cpx w[16] = {mk(1.f, 0.f),
                       mk(SDRB_C1_32, -SDRB_S1_32), mk(SDRB_C1_16, -SDRB_S1_16), mk(SDRB_C3_32, -SDRB_S3_32),
                       mk(SDRB_SQRT1_2, -SDRB_SQRT1_2),
                       mk(SDRB_S3_32, -SDRB_C3_32), mk(SDRB_S1_16, -SDRB_C1_16), mk(SDRB_S1_32, -SDRB_C1_32),
                       mk(0.f, -1.f),
                       mk(-SDRB_S1_32, -SDRB_C1_32), mk(-SDRB_S1_16, -SDRB_C1_16), mk(-SDRB_S3_32, -SDRB_C3_32),
                       mk(-SDRB_SQRT1_2, -SDRB_SQRT1_2),
                       mk(-SDRB_C3_32, -SDRB_S3_32), mk(-SDRB_C1_16, -SDRB_S1_16), mk(-SDRB_C1_32, -SDRB_S1_32)};
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const c2 p = v[rev16(k)];
        const cpx e = mk(p.re.x, p.im.x);
        cpx o = mk(p.re.y, p.im.y);
        if (k == 4) o = mul_w8_1(o);
        else if (k == 8) o = mul_mi(o);
        else if (k == 12) o = mul_w8_3(o);
        else if (k != 0) o = cmul(o, w[k]);
        z[k].re = mkp(e.x + o.x, e.x - o.x);
        z[k].im = mkp(e.y + o.y, e.y - o.y);
    }
}

// byte-free index helpers: unit (8 bytes) and float (4 bytes) addressing of one plane
SDRB_HD int f3_unit(int ka, int row, int c) { return ka * F3_PITCH + row * 16 + c; }

// ---- phase 1: thread t = 16 b + c.  v[j] = (x[a = 2j], x[a = 2j + 1]) on entry (loaded,
// converted, optionally windowed by the caller); on exit the exchange buffer holds, for every
// ka < 32, X1[ka][b][c] W_N^(t ka) as HALF (b & 1) of unit (ka, b >> 1, c).
// Split in two so that the kernel can put its "exchange buffer is free" barrier between the
// arithmetic (which overlaps other warps' phase 3) and the stores.
SDRB_HD void f3_phase1_compute(int t, c2* v, const cpx* TP, c2* y) {
    c2fft16(v);
    c2 z[16];
    f3_dit_tail(v, z);
    // twiddle pairs q[k] = (W_N^(t k), W_N^(t (k + 16))) from P[j] = W_N^(t 2^j)
    cpx P[5];
#pragma unroll
    for (int j = 0; j < 5; j++) P[j] = TP[j * 512 + t];
    pf qre[16], qim[16];
    qre[0] = mkp(1.0f, P[4].x);
    qim[0] = mkp(0.0f, P[4].y);
#pragma unroll
    for (int k = 1; k < 16; k++) {
        const int low = k & (-k);
        const int j = low == 1 ? 0 : (low == 2 ? 1 : (low == 4 ? 2 : 3));
        if (k == low) {
            const cpx h = cmul(P[j], P[4]);
            qre[k] = mkp(P[j].x, h.x);
            qim[k] = mkp(P[j].y, h.y);
        } else {
            c2 a;
            a.re = qre[k - low];
            a.im = qim[k - low];
            a = c2mulc(a, P[j].x, P[j].y);
            qre[k] = a.re;
            qim[k] = a.im;
        }
    }
#pragma unroll
    for (int k = 0; k < 16; k++) y[k] = c2mul(z[k], qre[k], qim[k]);
}
// The same phase 1 cut for the kernel: arithmetic up to the DIT tail (no shared memory), then twiddle and
// exchange stores interleaved value by value, so the stores spread over the twiddle arithmetic.
SDRB_HD void f3_phase1_fft(c2* v, c2* z) {
    c2fft16(v);
    f3_dit_tail(v, z);
}
SDRB_HD void f3_phase1_twiddle_store(int t, const c2* z, const cpx* TP, float* RE, float* IM) {
    cpx P[5];
#pragma unroll
    for (int j = 0; j < 5; j++) P[j] = TP[j * 512 + t];
    const int b = t >> 4, c = t & 15;
    const int base = ((b >> 1) * 16 + c) * 2 + (b & 1);
    pf qre[16], qim[16];
    qre[0] = mkp(1.0f, P[4].x);
    qim[0] = mkp(0.0f, P[4].y);
#pragma unroll
    for (int k = 0; k < 16; k++) {
        if (k > 0) {
            const int low = k & (-k);
            const int j = low == 1 ? 0 : (low == 2 ? 1 : (low == 4 ? 2 : 3));
            if (k == low) {
                const cpx h = cmul(P[j], P[4]);
                qre[k] = mkp(P[j].x, h.x);
                qim[k] = mkp(P[j].y, h.y);
            } else {
                c2 a;
                a.re = qre[k - low];
                a.im = qim[k - low];
                a = c2mulc(a, P[j].x, P[j].y);
                qre[k] = a.re;
                qim[k] = a.im;
            }
        }
        const c2 y = c2mul(z[k], qre[k], qim[k]);
        RE[k * (2 * F3_PITCH) + base] = y.re.x;
        IM[k * (2 * F3_PITCH) + base] = y.im.x;
        RE[(k + 16) * (2 * F3_PITCH) + base] = y.re.y;
        IM[(k + 16) * (2 * F3_PITCH) + base] = y.im.y;
    }
}
SDRB_HD void f3_phase1_store(int t, const c2* y, float* RE, float* IM) {
    const int b = t >> 4, c = t & 15;
    const int base = ((b >> 1) * 16 + c) * 2 + (b & 1);   // float index inside a ka plane
#pragma unroll
    for (int k = 0; k < 16; k++) {
        RE[k * (2 * F3_PITCH) + base] = y[k].re.x;
        IM[k * (2 * F3_PITCH) + base] = y[k].im.x;
        RE[(k + 16) * (2 * F3_PITCH) + base] = y[k].re.y;
        IM[(k + 16) * (2 * F3_PITCH) + base] = y[k].im.y;
    }
}
SDRB_HD void f3_phase1(int t, c2* v, const cpx* TP, float* RE, float* IM) {
    c2 y[16];
    f3_phase1_compute(t, v, TP, y);
    f3_phase1_store(t, y, RE, IM);
}

// ---- phase 2: thread t = 16 ka + c, DFT32 over b, worlds (b even, b odd); in place: unit (ka, j, c) in, unit (ka, k, c) = (X2[kb = k], X2[kb = k + 16]) out.
SDRB_HD void f3_phase2(int t, pf* RE, pf* IM) {
    const int ka = t >> 4, c = t & 15;
    pf* re = RE + f3_unit(ka, 0, c);
    pf* im = IM + f3_unit(ka, 0, c);
    c2 v[16];
#pragma unroll
    for (int j = 0; j < 16; j++) { v[j].re = re[j * 16]; v[j].im = im[j * 16]; }
    c2fft16(v);
    c2 z[16];
    f3_dit_tail(v, z);
#pragma unroll
    for (int k = 0; k < 16; k++) {   // the twiddle W_512^(c kb) is applied by phase 3 on load
        re[k * 16] = z[k].re;
        im[k * 16] = z[k].im;
    }
}

// ---- phase 3: thread t: ka = t & 31, kb' = t >> 5; DFT16 over c of the columns kb' (world A)
// and kb' + 16 (world B); re*re + im*im (src/dsp/fft_process.cpp:85); bin k lands at
// (k + N/2) mod N (:92-97).
// tw(kb', c) returns the pair (W_512^(c kb'), W_512^(c (kb' + 16))); kb' is uniform over a warp, so the device
// functor reads the constant bank (broadcast) instead of shared memory.
template <class TW>
SDRB_HD void f3_phase3(int t, const pf* RE, const pf* IM, TW tw, float* __restrict__ out) {
    const int ka = t & 31, kbp = t >> 5;
    const pf* re = RE + f3_unit(ka, kbp, 0);
    const pf* im = IM + f3_unit(ka, kbp, 0);
    c2 v[16];
#pragma unroll
    for (int c = 0; c < 16; c++) { v[c].re = re[c]; v[c].im = im[c]; }
#pragma unroll
    for (int c = 1; c < 16; c++) {
        const pw4 w = tw(kbp, c);
        v[c] = c2mul(v[c], w.re, w.im);
    }
    c2fft16(v);
    const int k0 = ka + 32 * kbp;
#pragma unroll
    for (int kc = 0; kc < 16; kc++) {
        const c2 x = v[rev16(kc)];
        const pf p = pfma(x.im, x.im, pmul(x.re, x.re));
        out[(k0 + 1024 * kc + 8192) & 16383] = p.x;
        out[(k0 + 512 + 1024 * kc + 8192) & 16383] = p.y;
    }
}

// Two raw samples -> one two-worlds value (world A = r0, world B = r1); same arithmetic as
// Ingest2 (exactly convertIQ, src/ssb/ssb_demod_opt.cpp:37-41), two lanes per instruction.
template <int FMT>
struct Ingest3 {
    typedef typename Ingest2<FMT>::raw_t raw_t;
    SDRB_HD static c2 cvt2(raw_t r0, raw_t r1) {
        const cpx a = Ingest2<FMT>::cvt(r0), b = Ingest2<FMT>::cvt(r1);
        c2 v;
        v.re = mkp(a.x, b.x);
        v.im = mkp(a.y, b.y);
        return v;
    }
};
#if defined(__CUDA_ARCH__)
template <>
struct Ingest3<FMT_CU8> {
    typedef uint16_t raw_t;
    __device__ __forceinline__ static c2 cvt2(uint16_t r0, uint16_t r1) {
        // byte at bits 16..23 of a float word = b * 2^-133 exactly; fma(., 2^126, -127.4f/128) rounds once (fft16k2.cuh)
        const pf mi = mkp(__uint_as_float(__byte_perm((unsigned)r0, 0u, 0x4044)),
                          __uint_as_float(__byte_perm((unsigned)r1, 0u, 0x4044)));
        const pf mq = mkp(__uint_as_float(__byte_perm((unsigned)r0, 0u, 0x4144)),
                          __uint_as_float(__byte_perm((unsigned)r1, 0u, 0x4144)));
        const pf k = pdup(8.50705917302346158658e37f), off = pdup(-127.4f * 0.0078125f);
        c2 v;
        v.re = pfma(mi, k, off);
        v.im = pfma(mq, k, off);
        return v;
    }
};
template <>
struct Ingest3<FMT_CS16> {
    typedef uint32_t raw_t;
    __device__ __forceinline__ static c2 cvt2(uint32_t r0, uint32_t r1) {
        unsigned i0, q0, i1, q1;
        asm("prmt.b32 %0, %1, 0, 0x9910;" : "=r"(i0) : "r"(r0));
        asm("prmt.b32 %0, %1, 0, 0xBB32;" : "=r"(q0) : "r"(r0));
        asm("prmt.b32 %0, %1, 0, 0x9910;" : "=r"(i1) : "r"(r1));
        asm("prmt.b32 %0, %1, 0, 0xBB32;" : "=r"(q1) : "r"(r1));
        const pf k = pdup(3.0517578125e-5f), off = pdup(-384.0f);
        c2 v;
        v.re = pfma(mkp(__uint_as_float(0x4B400000u + i0), __uint_as_float(0x4B400000u + i1)), k, off);
        v.im = pfma(mkp(__uint_as_float(0x4B400000u + q0), __uint_as_float(0x4B400000u + q1)), k, off);
        return v;
    }
};
#endif

// Load of phase 1: v[j] = (x[512 (2j) + t], x[512 (2j + 1) + t]).
template <int FMT>
SDRB_HD void f3_phase1_load(int t, const typename Ingest2<FMT>::raw_t* in, c2* v) {
#pragma unroll
    for (int j = 0; j < 16; j++) v[j] = Ingest3<FMT>::cvt2(in[1024 * j + t], in[1024 * j + 512 + t]);
}

}  // namespace sdrb
