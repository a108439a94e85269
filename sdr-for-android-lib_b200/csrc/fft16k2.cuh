// fft16k2.cuh — second-generation one-CTA-per-transform 16384-point FFT: 512 threads, 32 points
// per thread, N = 32 * 32 * 16, TWO shared-memory exchanges instead of three
// (reference: src/dsp/fft_process.cpp:62-97 — copy, fftwf_execute, power, fftshift).
//
//   n = 512 a + 16 b + c           k = ka + 32 kb + 1024 kc        (a, b, ka, kb < 32; c, kc < 16)
//   phase 1: DFT32 over a   thread t = 16 b + c          twiddle W_N^(t ka)
//   phase 2: DFT32 over b   thread t = 16 ka + c         (twiddle W_512^(c kb) applied by phase 3 on load)
//   phase 3: DFT16 over c   thread t -> (ka, kb) pairs p = t and t + 512, ka = p & 31, kb = p >> 5
//            then |X|^2, fftshift folded into the store index
// Exchange buffer E[ka][row][c] with plane pitch 513 complex (odd pitch mod 16): every phase's
// 8-byte accesses cover all 32 banks exactly once per half-warp (audited in tests/cpp/fft_sim.cpp).
// Phase-1 twiddles W_N^(t ka), ka = 1..31, are built per thread from the five table values
// W_N^(t 2^j) by a product tree (depth <= 4), which costs fewer shared-memory wavefronts and no
// more FP32 work than a two-table product.
//
// The register codelets are the decimation-in-time, FMA-form ones of fft_dit.cuh (fft_dit<5>, fft_dit<4>); the
// decimation-in-frequency fft32 below stays for the four-step kernels (fft_big.cuh).
// The phase functions are __host__ __device__ so the CPU simulator can run them thread by thread.
#pragma once
#include <stdint.h>

#include "fft_core.cuh"
#include "fft_dit.cuh"

namespace sdrb {

constexpr int F2_N = 16384;
constexpr int F2_THREADS = 512;
constexpr int F2_PITCH = 513;                       // complex elements per ka plane
constexpr int F2_E_ELEMS = 32 * F2_PITCH;           // 16416 complex = 131328 B
constexpr int F2_TP = 5 * 512;                      // [j][t]   W_N^(t 2^j)
constexpr int F2_T2 = 32 * 16;                      // [kb][c]  W_512^(c kb)
constexpr int F2_TW_ELEMS = F2_TP + F2_T2;

SDRB_HD constexpr int slot32(int k) { return (k & 1) * 16 + rev16(k >> 1); }

// cos/sin(2 pi i / 32)
#define SDRB_C1_32 0.98078528040323044913f
#define SDRB_S1_32 0.19509032201612826785f
#define SDRB_C3_32 0.83146961230254523708f
#define SDRB_S3_32 0.55557023301960222474f

// 32-point forward DFT, natural-order input, X[k] left in v[slot32(k)].
SDRB_HD void fft32(cpx* v) {
    const cpx w[16] = {mk(1.f, 0.f),
                       mk(SDRB_C1_32, -SDRB_S1_32), mk(SDRB_C1_16, -SDRB_S1_16), mk(SDRB_C3_32, -SDRB_S3_32),
                       mk(SDRB_SQRT1_2, -SDRB_SQRT1_2),
                       mk(SDRB_S3_32, -SDRB_C3_32), mk(SDRB_S1_16, -SDRB_C1_16), mk(SDRB_S1_32, -SDRB_C1_32),
                       mk(0.f, -1.f),
                       mk(-SDRB_S1_32, -SDRB_C1_32), mk(-SDRB_S1_16, -SDRB_C1_16), mk(-SDRB_S3_32, -SDRB_C3_32),
                       mk(-SDRB_SQRT1_2, -SDRB_SQRT1_2),
                       mk(-SDRB_C3_32, -SDRB_S3_32), mk(-SDRB_C1_16, -SDRB_S1_16), mk(-SDRB_C1_32, -SDRB_S1_32)};
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const cpx s = cadd(v[i], v[i + 16]);
        const cpx d = csub(v[i], v[i + 16]);
        v[i] = s;
        if (i == 0) v[16] = d;
        else if (i == 4) v[20] = mul_w8_1(d);
        else if (i == 8) v[24] = mul_mi(d);
        else if (i == 12) v[28] = mul_w8_3(d);
        else v[i + 16] = cmul(d, w[i]);
    }
    fft16(v);
    fft16(v + 16);
}

// Ingest of one raw sample for this kernel: the wire format's power-of-two scale is applied
// here with a fused multiply-add, which rounds exactly like convertIQ's subtract-then-scale
// (src/ssb/ssb_demod_opt.cpp:37-41: scaling by 2^-7 commutes with rounding), so no scale is
// needed at the output.
template <int FMT>
struct Ingest2;
template <>
struct Ingest2<FMT_CU8> {
    typedef uint16_t raw_t;
    SDRB_HD static cpx cvt(uint16_t r) {
#if defined(__CUDA_ARCH__)
        // the byte placed at bits 16..23 of a float word is b * 2^-133 exactly (sub-normal or first binade);
        // fma(that, 2^126, -127.4f/128) = round((b - 127.4f) / 128): ONE instruction per component
        const float fi = __uint_as_float(__byte_perm((unsigned)r, 0u, 0x4044));
        const float fq = __uint_as_float(__byte_perm((unsigned)r, 0u, 0x4144));
        return mk(__fmaf_rn(fi, 8.50705917302346158658e37f, -127.4f * 0.0078125f),
                  __fmaf_rn(fq, 8.50705917302346158658e37f, -127.4f * 0.0078125f));
#else
        const float fi = (float)(int)(r & 0xffu), fq = (float)(int)(r >> 8);
        return mk((fi - 127.4f) * 0.0078125f, (fq - 127.4f) * 0.0078125f);
#endif
    }
};
template <>
struct Ingest2<FMT_CS16> {
    typedef uint32_t raw_t;
    SDRB_HD static cpx cvt(uint32_t r) {
#if defined(__CUDA_ARCH__)
        // 0x4B400000 + s is the float 1.5 * 2^23 + s; (that) * 2^-15 - 384 = s * 2^-15 exactly
        // prmt in its default mode replicates the sign of a byte when bit 3 of the selector nibble
        // is set (the __byte_perm intrinsic ignores that bit): one instruction per sign extension
        unsigned si, sq;
        asm("prmt.b32 %0, %1, 0, 0x9910;" : "=r"(si) : "r"(r));
        asm("prmt.b32 %0, %1, 0, 0xBB32;" : "=r"(sq) : "r"(r));
        const unsigned bi = 0x4B400000u + si, bq = 0x4B400000u + sq;
        return mk(__fmaf_rn(__uint_as_float(bi), 3.0517578125e-5f, -384.0f),
                  __fmaf_rn(__uint_as_float(bq), 3.0517578125e-5f, -384.0f));
#else
        return mk((float)(int)(int16_t)(r & 0xffffu) * 3.0517578125e-5f,
                  (float)(int)(int16_t)(r >> 16) * 3.0517578125e-5f);
#endif
    }
};
template <>
struct Ingest2<FMT_CF32> {
    typedef cpx raw_t;
    SDRB_HD static cpx cvt(cpx r) { return r; }
};

SDRB_HD int f2_slot(int plane, int row, int c) { return plane * F2_PITCH + row * 16 + c; }

// ---- phase 1: load + convert, then DFT32 over a and twiddle.  v[dit_slot<5>(ka)] holds the result.
template <int FMT>
SDRB_HD void f2_phase1_load(int t, const typename Ingest2<FMT>::raw_t* in, cpx* v) {
#pragma unroll
    for (int a = 0; a < 32; a++) v[a] = Ingest2<FMT>::cvt(in[512 * a + t]);
}
SDRB_HD void f2_phase1_compute(int t, const cpx* TP, cpx* v) {
    fft_dit<5>(v);
    // W_N^(t ka) from W_N^(t 2^j): w[k] = w[k with lowest set bit cleared] * P[ctz k]
    cpx P[5];
#pragma unroll
    for (int j = 0; j < 5; j++) P[j] = TP[j * 512 + t];
    cpx w[32];
#pragma unroll
    for (int k = 1; k < 32; k++) {
        const int low = k & (-k);
        const int j = low == 1 ? 0 : (low == 2 ? 1 : (low == 4 ? 2 : (low == 8 ? 3 : 4)));
        w[k] = (k == low) ? P[j] : cmul(w[k - low], P[j]);
        v[dit_slot<5>(k)] = cmul(v[dit_slot<5>(k)], w[k]);
    }
}
// The same twiddle pass with the exchange store folded in: each value is stored as soon as it is final, so the
// 32 stores spread over the twiddle arithmetic instead of bursting into the load/store queue after it.
SDRB_HD void f2_phase1_twiddle_store(int t, const cpx* TP, cpx* v, cpx* E) {
    cpx P[5];
#pragma unroll
    for (int j = 0; j < 5; j++) P[j] = TP[j * 512 + t];
    cpx w[32];
    E[t] = v[0];
#pragma unroll
    for (int k = 1; k < 32; k++) {
        const int low = k & (-k);
        const int j = low == 1 ? 0 : (low == 2 ? 1 : (low == 4 ? 2 : (low == 8 ? 3 : 4)));
        w[k] = (k == low) ? P[j] : cmul(w[k - low], P[j]);
        E[k * F2_PITCH + t] = cmul(v[dit_slot<5>(k)], w[k]);
    }
}
template <int FMT>
SDRB_HD void f2_phase1(int t, const typename Ingest2<FMT>::raw_t* in, const cpx* TP, cpx* v) {
    f2_phase1_load<FMT>(t, in, v);
    f2_phase1_compute(t, TP, v);
}
SDRB_HD void f2_phase1_store(int t, cpx* E, const cpx* v) {
#pragma unroll
    for (int k = 0; k < 32; k++) E[k * F2_PITCH + t] = v[dit_slot<5>(k)];
}

// ---- phase 2: DFT32 over b for (ka, c), in place (its twiddle W_512^(c kb) is applied by phase 3 on load)
SDRB_HD void f2_phase2(int t, cpx* E) {
    const int ka = t >> 4, c = t & 15;
    cpx* base = E + ka * F2_PITCH + c;
    cpx v[32];
#pragma unroll
    for (int b = 0; b < 32; b++) v[b] = base[b * 16];
    fft_dit<5>(v);
#pragma unroll
    for (int k = 0; k < 32; k++) base[k * 16] = v[dit_slot<5>(k)];
}

// ---- phase 3: twiddle W_512^(c kb) on load (kb is uniform over a warp: the device functor reads the
// constant bank, one broadcast per value instead of a 256-byte shared-memory wavefront pair), DFT16 over c
// for two (ka, kb) pairs, power, fftshift, store
template <class TW>
SDRB_HD void f2_phase3(int t, const cpx* E, TW tw, float* __restrict__ out) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int p = t + 512 * h;
        const int ka = p & 31, kb = p >> 5;
        const cpx* base = E + ka * F2_PITCH + kb * 16;
        cpx v[16];
        v[0] = base[0];
#pragma unroll
        for (int c = 1; c < 16; c++) v[c] = cmul(base[c], tw(kb, c));
        fft_dit<4>(v);
        const int k0 = ka + 32 * kb;
#pragma unroll
        for (int kc = 0; kc < 16; kc++) {
            const cpx x = v[dit_slot<4>(kc)];
            // re*re + im*im (src/dsp/fft_process.cpp:85); bin k lands at (k + N/2) mod N (:92-97)
            out[(k0 + 1024 * kc + 8192) & 16383] = x.x * x.x + x.y * x.y;
        }
    }
}

}  // namespace sdrb
