// fft_core.cuh — register-level FFT building blocks shared by every FFT kernel.
//
// Replaces fftwf_plan_dft_1d/fftwf_execute at src/dsp/fft_process.cpp:77-79 of the
// reference (forward transform, sign -1, unnormalised).  Everything here is
// __host__ __device__ so that tests/cpp/fft_sim.cpp can run the exact index and
// butterfly logic of the kernels on the CPU, thread by thread, without a GPU.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define SDRB_HD __host__ __device__ __forceinline__
#else
#define SDRB_HD inline
#endif

namespace sdrb {

struct alignas(8) cpx {
    float x, y;
};

SDRB_HD cpx mk(float x, float y) { cpx r; r.x = x; r.y = y; return r; }
SDRB_HD cpx cadd(cpx a, cpx b) { return mk(a.x + b.x, a.y + b.y); }
SDRB_HD cpx csub(cpx a, cpx b) { return mk(a.x - b.x, a.y - b.y); }
SDRB_HD cpx cmul(cpx a, cpx b) { return mk(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
// multiply by -i  (forward-transform quarter turn)
SDRB_HD cpx mul_mi(cpx a) { return mk(a.y, -a.x); }

#define SDRB_SQRT1_2 0.70710678118654752440f
#define SDRB_C1_16 0.92387953251128675613f /* cos(pi/8) */
#define SDRB_S1_16 0.38268343236508977173f /* sin(pi/8) */

// a * W8^1 = a * (1 - i)/sqrt2 ; a * W8^3 = a * (-1 - i)/sqrt2
SDRB_HD cpx mul_w8_1(cpx a) { return mk((a.x + a.y) * SDRB_SQRT1_2, (a.y - a.x) * SDRB_SQRT1_2); }
SDRB_HD cpx mul_w8_3(cpx a) { return mk((a.y - a.x) * SDRB_SQRT1_2, -(a.x + a.y) * SDRB_SQRT1_2); }

// 4-point forward DFT in place: (a0,a1,a2,a3) -> (X0,X1,X2,X3).
SDRB_HD void fft4(cpx& a0, cpx& a1, cpx& a2, cpx& a3) {
    cpx s02 = cadd(a0, a2), d02 = csub(a0, a2);
    cpx s13 = cadd(a1, a3), d13 = mul_mi(csub(a1, a3));
    a0 = cadd(s02, s13);
    a2 = csub(s02, s13);
    a1 = cadd(d02, d13);
    a3 = csub(d02, d13);
}

// 2-point
SDRB_HD void fft2(cpx& a0, cpx& a1) {
    cpx s = cadd(a0, a1), d = csub(a0, a1);
    a0 = s;
    a1 = d;
}

// 16-point forward DFT, input v[n] in natural order.  Output X[k] is left in
// v[rev16(k)] with rev16(k) = 4*(k&3) + (k>>2)  (base-4 digit reversal).
SDRB_HD constexpr int rev16(int k) { return 4 * (k & 3) + (k >> 2); }

SDRB_HD void fft16(cpx* v) {
    // stage 1: four radix-4 butterflies over stride 4; result z[n0][k1] -> v[n0 + 4*k1]
    fft4(v[0], v[4], v[8], v[12]);
    fft4(v[1], v[5], v[9], v[13]);
    fft4(v[2], v[6], v[10], v[14]);
    fft4(v[3], v[7], v[11], v[15]);
    // twiddles W16^(n0*k1)
    v[5] = cmul(v[5], mk(SDRB_C1_16, -SDRB_S1_16));    // W16^1
    v[9] = mul_w8_1(v[9]);                              // W16^2
    v[13] = cmul(v[13], mk(SDRB_S1_16, -SDRB_C1_16));  // W16^3
    v[6] = mul_w8_1(v[6]);                              // W16^2
    v[10] = mul_mi(v[10]);                              // W16^4
    v[14] = mul_w8_3(v[14]);                            // W16^6
    v[7] = cmul(v[7], mk(SDRB_S1_16, -SDRB_C1_16));    // W16^3
    v[11] = mul_w8_3(v[11]);                            // W16^6
    v[15] = cmul(v[15], mk(-SDRB_C1_16, SDRB_S1_16));  // W16^9
    // stage 2: radix-4 over n0 for each k1; X[k1 + 4*k0] -> v[4*k1 + k0]
    fft4(v[0], v[1], v[2], v[3]);
    fft4(v[4], v[5], v[6], v[7]);
    fft4(v[8], v[9], v[10], v[11]);
    fft4(v[12], v[13], v[14], v[15]);
}

// ---------------------------------------------------------------- ingest formats
enum : int { FMT_CU8 = 0, FMT_CS16 = 1, FMT_CF32 = 2 };

// One complex sample of the wire format -> float pair (include/sdr_b200.h sdrb_format).
// CU8 follows convertIQ (src/ssb/ssb_demod_opt.cpp:37-41) except that the exact
// power-of-two scale 1/128 is deferred: the FFT is linear and scaling by 2^-7 commutes
// exactly with every float operation, so |X|^2 is multiplied by 2^-14 at the end and
// the result is bit-identical to scaling the input first.
template <int FMT>
struct Ingest;
template <>
struct Ingest<FMT_CU8> {
    typedef uint16_t raw_t;  // two bytes: I, Q
    static constexpr float kPowerScale = 1.0f / 16384.0f;
    SDRB_HD static cpx cvt(uint16_t r) {
        return mk((float)(int)(r & 0xffu) - 127.4f, (float)(int)(r >> 8) - 127.4f);
    }
};
template <>
struct Ingest<FMT_CS16> {
    typedef uint32_t raw_t;  // two int16: I, Q   (x * 2^-15, deferred the same way)
    static constexpr float kPowerScale = 1.0f / 1073741824.0f;
    SDRB_HD static cpx cvt(uint32_t r) {
        return mk((float)(int)(int16_t)(r & 0xffffu), (float)(int)(int16_t)(r >> 16));
    }
};
template <>
struct Ingest<FMT_CF32> {
    typedef cpx raw_t;
    static constexpr float kPowerScale = 1.0f;
    SDRB_HD static cpx cvt(cpx r) { return r; }
};

}  // namespace sdrb
