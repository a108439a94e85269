// fft16k4.cuh — fourth-generation 16384-point FFT: every transform is cut into TWO independent
// 8192-point jobs by one decimation-in-frequency radix-2 step taken at ingest,
//     even bins  X[2k']   = DFT_8192{ x[n] + x[n+8192] }[k']
//     odd bins   X[2k'+1] = DFT_8192{ (x[n] - x[n+8192]) W_16384^n }[k']
// (reference: src/dsp/fft_process.cpp:62-97 — copy, fftwf_execute, power, fftshift), so that a job is
// 256 threads x 32 points with a 64 KiB exchange buffer and TWO CTAs fit on one SM.  The two
// CTAs run out of phase: while one is in a shared-memory exchange the other is in arithmetic,
// which the one-CTA-per-SM kernels (fft16k2/3) could not do because a whole transform fills half
// of the register file.
//
//   8192-point job:  m = 256 a + 16 b + c        k' = ka + 32 kb + 512 kc    (a, ka < 32; b, c, kb, kc < 16)
//   phase 1: ingest + radix-2 combine, DFT32 over a        thread t = 16 b + c     twiddle W_8192^(t ka)
//            odd job: the factor W_16384^m = W_64^a W_16384^t is folded into the DFT32's first stage
//            (fft32_odd) and into the twiddle (W_16384^(t (2 ka + 1)))
//   phase 2: DFT16 over b for two (ka, c) columns q = t, t + 256: ka = q >> 4, c = q & 15   (in place)
//   phase 3: twiddle W_256^(c kb) on load (warp-uniform kb: constant-cache broadcast), DFT16 over c
//            for two (ka, kb) pairs q = t, t + 256: ka = q & 31, kb = q >> 5; |X|^2; bin
//            k = 2 k' + parity stored at (k + N/2) mod N
// Exchange buffer E[ka][row][c], plane pitch 257 complex (odd): all 8-byte accesses of a half-warp
// fall into 16 distinct bank pairs (audited by tests/cpp/fft_sim4.cpp).
// The per-thread twiddle seeds W_8192^(t 2^j), j < 5, and W_16384^t live in registers for the
// whole kernel.
#pragma once
#include <stdint.h>

#include "fft16k2.cuh"

namespace sdrb {

constexpr int F4_N = 16384;
constexpr int F4_H = 8192;
constexpr int F4_THREADS = 256;
constexpr int F4_PITCH = 257;
constexpr int F4_E_ELEMS = 32 * F4_PITCH;   // 8224 complex = 65792 B
constexpr int F4_TQ = 6 * 256;              // [j][t]: j < 5 -> W_8192^(t 2^j), j = 5 -> W_16384^t
constexpr int F4_T3 = 16 * 16;              // [kb][c]  W_256^(c kb)

// W_64^j, j < 48
SDRB_HD cpx w64(int j) {
    const cpx w[48] = {
        mk(1.0f, 0.0f),  // 0
        mk(0.99518472667219693f, -0.098017140329560604f),  // 1
        mk(0.98078528040323043f, -0.19509032201612825f),  // 2
        mk(0.95694033573220882f, -0.29028467725446233f),  // 3
        mk(0.92387953251128674f, -0.38268343236508978f),  // 4
        mk(0.88192126434835505f, -0.47139673682599764f),  // 5
        mk(0.83146961230254524f, -0.55557023301960218f),  // 6
        mk(0.77301045336273699f, -0.63439328416364549f),  // 7
        mk(0.70710678118654757f, -0.70710678118654746f),  // 8
        mk(0.63439328416364549f, -0.77301045336273699f),  // 9
        mk(0.55557023301960229f, -0.83146961230254524f),  // 10
        mk(0.47139673682599781f, -0.88192126434835494f),  // 11
        mk(0.38268343236508984f, -0.92387953251128674f),  // 12
        mk(0.29028467725446233f, -0.95694033573220894f),  // 13
        mk(0.19509032201612833f, -0.98078528040323043f),  // 14
        mk(0.09801714032956077f, -0.99518472667219682f),  // 15
        mk(0.0f, -1.0f),  // 16
        mk(-0.098017140329560645f, -0.99518472667219693f),  // 17
        mk(-0.19509032201612819f, -0.98078528040323043f),  // 18
        mk(-0.29028467725446216f, -0.95694033573220894f),  // 19
        mk(-0.38268343236508973f, -0.92387953251128674f),  // 20
        mk(-0.4713967368259977f, -0.88192126434835505f),  // 21
        mk(-0.55557023301960196f, -0.83146961230254546f),  // 22
        mk(-0.63439328416364538f, -0.7730104533627371f),  // 23
        mk(-0.70710678118654746f, -0.70710678118654757f),  // 24
        mk(-0.77301045336273699f, -0.63439328416364549f),  // 25
        mk(-0.83146961230254535f, -0.55557023301960218f),  // 26
        mk(-0.88192126434835494f, -0.47139673682599786f),  // 27
        mk(-0.92387953251128674f, -0.38268343236508989f),  // 28
        mk(-0.95694033573220882f, -0.29028467725446239f),  // 29
        mk(-0.98078528040323043f, -0.19509032201612861f),  // 30
        mk(-0.99518472667219682f, -0.098017140329560826f),  // 31
        mk(-1.0f, 0.0f),  // 32
        mk(-0.99518472667219693f, 0.09801714032956059f),  // 33
        mk(-0.98078528040323043f, 0.19509032201612836f),  // 34
        mk(-0.95694033573220894f, 0.29028467725446211f),  // 35
        mk(-0.92387953251128685f, 0.38268343236508967f),  // 36
        mk(-0.88192126434835505f, 0.47139673682599764f),  // 37
        mk(-0.83146961230254546f, 0.55557023301960196f),  // 38
        mk(-0.7730104533627371f, 0.63439328416364527f),  // 39
        mk(-0.70710678118654768f, 0.70710678118654746f),  // 40
        mk(-0.63439328416364593f, 0.77301045336273666f),  // 41
        mk(-0.55557023301960218f, 0.83146961230254524f),  // 42
        mk(-0.47139673682599786f, 0.88192126434835494f),  // 43
        mk(-0.38268343236509034f, 0.92387953251128652f),  // 44
        mk(-0.29028467725446244f, 0.95694033573220882f),  // 45
        mk(-0.19509032201612866f, 0.98078528040323032f),  // 46
        mk(-0.098017140329560451f, 0.99518472667219693f),  // 47
    };
    return w[j];
}

// "Odd" 32-point DFT: Y[k] = sum_a v[a] W_64^(a (2k+1)), natural-order input, Y[k] left in
// v[slot32(k)].  First stage: u[i] = v[i] W64^i, u[i+16] = v[i+16] W64^i (-i), so
//   s_i = W64^i  (v[i] + (-i) v[i+16])          -> even k
//   d_i = W64^3i (v[i] - (-i) v[i+16])          -> odd k
SDRB_HD void fft32_odd(cpx* v) {
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const cpx r = mul_mi(v[i + 16]);
        const cpx s = cadd(v[i], r);
        const cpx d = csub(v[i], r);
        if (i == 0) { v[0] = s; v[16] = d; }
        else if (i == 8) { v[8] = mul_w8_1(s); v[24] = mul_w8_3(d); }
        else { v[i] = cmul(s, w64(i)); v[i + 16] = cmul(d, w64(3 * i)); }
    }
    fft16(v);
    fft16(v + 16);
}

// Ingest for this kernel.  CU8 on the device is ONE instruction per component: the byte is placed
// at bits 16..23 of a float word, whose value is then b * 2^-133 exactly (sub-normal or first
// binade), and fma(that, 2^126, -127.4f/128) = round((b - 127.4f)/128) = convertIQ's value
// (src/ssb/ssb_demod_opt.cpp:37-41: scaling by 2^-7 commutes with the rounding).
template <int FMT>
struct Ingest4 : Ingest2<FMT> {};
template <>
struct Ingest4<FMT_CU8> {
    typedef uint16_t raw_t;
    SDRB_HD static cpx cvt(uint16_t r) {
#if defined(__CUDA_ARCH__)
        const float fi = __uint_as_float(__byte_perm((unsigned)r, 0u, 0x4044));
        const float fq = __uint_as_float(__byte_perm((unsigned)r, 0u, 0x4144));
        return mk(__fmaf_rn(fi, 8.50705917302346158658e37f, -127.4f * 0.0078125f),
                  __fmaf_rn(fq, 8.50705917302346158658e37f, -127.4f * 0.0078125f));
#else
        return Ingest2<FMT_CU8>::cvt(r);
#endif
    }
};

// ---- phase 1
// load + convert + radix-2 combine.  P = parity of the job (0: even bins, 1: odd bins).
template <int FMT, int P>
SDRB_HD void f4_phase1_load(int t, const typename Ingest4<FMT>::raw_t* in, cpx* v) {
#pragma unroll
    for (int a = 0; a < 32; a++) {
        const cpx x0 = Ingest4<FMT>::cvt(in[256 * a + t]);
        const cpx x1 = Ingest4<FMT>::cvt(in[F4_H + 256 * a + t]);
        v[a] = P ? csub(x0, x1) : cadd(x0, x1);
    }
}
// tq[j] = W_8192^(t 2^j) (j < 5), tq[5] = W_16384^t
template <int P>
SDRB_HD void f4_phase1_compute(const cpx* tq, cpx* v) {
    if (P) fft32_odd(v); else fft32(v);
    cpx w[32];
    if (P) { w[0] = tq[5]; v[0] = cmul(v[0], w[0]); }
#pragma unroll
    for (int k = 1; k < 32; k++) {
        const int low = k & (-k);
        const int j = low == 1 ? 0 : (low == 2 ? 1 : (low == 4 ? 2 : (low == 8 ? 3 : 4)));
        if (P) w[k] = cmul(w[k - low], tq[j]);
        else w[k] = (k == low) ? tq[j] : cmul(w[k - low], tq[j]);
        v[slot32(k)] = cmul(v[slot32(k)], w[k]);
    }
}
SDRB_HD void f4_phase1_store(int t, cpx* E, const cpx* v) {
#pragma unroll
    for (int k = 0; k < 32; k++) E[k * F4_PITCH + t] = v[slot32(k)];
}

// ---- phase 2: DFT16 over b for the columns (ka, c) of q = t and t + 256, in place, no twiddle
SDRB_HD void f4_phase2(int t, cpx* E) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int q = t + 256 * h;
        cpx* base = E + (q >> 4) * F4_PITCH + (q & 15);
        cpx v[16];
#pragma unroll
        for (int b = 0; b < 16; b++) v[b] = base[b * 16];
        fft16(v);
#pragma unroll
        for (int kb = 0; kb < 16; kb++) base[kb * 16] = v[rev16(kb)];
    }
}

// ---- phase 3: twiddle on load, DFT16 over c for two (ka, kb) pairs, power, fftshift, store.
// TW3(kb, c) returns W_256^(c kb); kb is uniform over a warp.
template <int P, class TW3>
SDRB_HD void f4_phase3(int t, const cpx* E, TW3 tw3, float* __restrict__ out) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int q = t + 256 * h;
        const int ka = q & 31, kb = q >> 5;
        const cpx* base = E + ka * F4_PITCH + kb * 16;
        int kbt = kb;
#if defined(__CUDA_ARCH__)
        // opaque copy: keeps the constant-bank twiddle loads below the preceding barrier (hoisted into
        // phase 1 they cost 30 registers and spill the transform)
        asm volatile("" : "+r"(kbt) :: "memory");
#endif
        cpx v[16];
        v[0] = base[0];
#pragma unroll
        for (int c = 1; c < 16; c++) v[c] = cmul(base[c], tw3(kbt, c));
        fft16(v);
        const int k0 = 2 * (ka + 32 * kb) + P;
#pragma unroll
        for (int kc = 0; kc < 16; kc++) {
            const cpx x = v[rev16(kc)];
            // re*re + im*im (src/dsp/fft_process.cpp:85); bin k lands at (k + N/2) mod N (:92-97)
            out[(k0 + 1024 * kc + 8192) & 16383] = x.x * x.x + x.y * x.y;
        }
    }
}

// ---- the same phases 2 and 3 cut into load / compute / store pieces (both columns in 32 registers), so
// that a kernel can group its shared-memory accesses into explicit exchange sections (fft16k5.cu)
SDRB_HD void f4_p2_load(int t, const cpx* E, cpx* v) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int q = t + 256 * h;
        const cpx* base = E + (q >> 4) * F4_PITCH + (q & 15);
#pragma unroll
        for (int b = 0; b < 16; b++) v[16 * h + b] = base[b * 16];
    }
}
SDRB_HD void f4_p2_compute(cpx* v) {
    fft16(v);
    fft16(v + 16);
}
SDRB_HD void f4_p2_store(int t, cpx* E, const cpx* v) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int q = t + 256 * h;
        cpx* base = E + (q >> 4) * F4_PITCH + (q & 15);
#pragma unroll
        for (int kb = 0; kb < 16; kb++) base[kb * 16] = v[16 * h + rev16(kb)];
    }
}
SDRB_HD void f4_p3_load(int t, const cpx* E, cpx* v) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int q = t + 256 * h;
        const cpx* base = E + (q & 31) * F4_PITCH + (q >> 5) * 16;
#pragma unroll
        for (int c = 0; c < 16; c++) v[16 * h + c] = base[c];
    }
}
template <int P, class TW3>
SDRB_HD void f4_p3_finish(int t, cpx* v, TW3 tw3, float* __restrict__ out) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int q = t + 256 * h;
        const int ka = q & 31, kb = q >> 5;
        cpx* w = v + 16 * h;
#pragma unroll
        for (int c = 1; c < 16; c++) w[c] = cmul(w[c], tw3(kb, c));
        fft16(w);
        const int k0 = 2 * (ka + 32 * kb) + P;
#pragma unroll
        for (int kc = 0; kc < 16; kc++) {
            const cpx x = w[rev16(kc)];
            out[(k0 + 1024 * kc + 8192) & 16383] = x.x * x.x + x.y * x.y;
        }
    }
}

}  // namespace sdrb
