// fft16k.cuh — one-CTA-per-transform 16384-point FFT, fused with the ingest
// convert, |X|^2 and fftshift (reference: src/dsp/fft_process.cpp:62-97).
//
// Decomposition (decimation in frequency), N = 16 * 16 * 16 * 4, 1024 threads,
// 16 points per thread held in registers, three in-place exchanges through one
// 128 KiB shared-memory buffer:
//   n = 1024 a + 64 b + 4 c + d        k = ka + 16 kb + 256 kc + 4096 kd
//   phase 1: DFT16 over a   (thread = r = 64b+4c+d),  twiddle W_N^(r ka)
//   phase 2: DFT16 over b   (thread = (ka, s=4c+d)),  twiddle W_1024^(s kb)
//   phase 3: DFT16 over c   (thread = (ka, kb, d)),   twiddle W_64^(d kc)
//   phase 4: DFT4  over d   (thread = (ka, kb, kc mod 4); four butterflies)
// Exchange buffer E[ka][kb|b][slot], row pitch S1 = 1026 float2 per ka, and inside
// each 64-entry (ka, b) row the slot of s is rotated by 4*b so that every phase
// reads and writes conflict-free (see DESIGN.md "FFT16K shared-memory layout").
// Each thread writes back exactly the slots it read, so one barrier per phase suffices.
//
// The phase functions are __host__ __device__: tests/cpp/fft_sim.cpp executes them
// thread by thread on the CPU to validate the index algebra.
#pragma once
#include <stdint.h>

#include "fft_core.cuh"

namespace sdrb {

constexpr int F16K_N = 16384;
constexpr int F16K_THREADS = 1024;
constexpr int F16K_S1 = 1026;                     // float2 pitch of one ka plane
constexpr int F16K_E_ELEMS = 16 * F16K_S1;        // float2 elements of the exchange buffer
// twiddle tables (cpx) appended after E in shared memory
constexpr int F16K_TA = 32 * 16;                  // [warp][ka]   W_N^(32 w ka)
constexpr int F16K_TB = 16 * 32;                  // [ka][lane]   W_N^(lane ka)
constexpr int F16K_T2 = 16 * 64;                  // [kb][s]      W_1024^(s kb)
constexpr int F16K_T3 = 16 * 4;                   // [kc][d]      W_64^(d kc)
constexpr int F16K_TW_ELEMS = F16K_TA + F16K_TB + F16K_T2 + F16K_T3;
constexpr int F16K_SMEM_BYTES = (F16K_E_ELEMS + F16K_TW_ELEMS) * 8;

SDRB_HD int f16k_slot(int plane, int row, int s) {
    // plane = ka, row = b or kb (0..15), s = 0..63 logical position inside the row
    return plane * F16K_S1 + row * 64 + ((s + 4 * row) & 63);
}

// ---- phase 1 -------------------------------------------------------------
template <int FMT>
SDRB_HD void f16k_phase1(int t, const typename Ingest<FMT>::raw_t* __restrict__ in, cpx* E, const cpx* TA,
                         const cpx* TB) {
    cpx v[16];
#pragma unroll
    for (int a = 0; a < 16; a++) v[a] = Ingest<FMT>::cvt(in[1024 * a + t]);
    fft16(v);
    const int w = t >> 5, l = t & 31;
#pragma unroll
    for (int k = 1; k < 16; k++) {
        cpx tw = cmul(TA[w * 16 + k], TB[k * 32 + l]);
        v[rev16(k)] = cmul(v[rev16(k)], tw);
    }
    const int b = t >> 6, s = t & 63;
#pragma unroll
    for (int k = 0; k < 16; k++) E[f16k_slot(k, b, s)] = v[rev16(k)];
}

// ---- phase 2 -------------------------------------------------------------
SDRB_HD void f16k_phase2(int t, cpx* E, const cpx* T2) {
    const int ka = t >> 6, s = t & 63;
    cpx v[16];
#pragma unroll
    for (int b = 0; b < 16; b++) v[b] = E[f16k_slot(ka, b, s)];
    fft16(v);
#pragma unroll
    for (int k = 1; k < 16; k++) v[rev16(k)] = cmul(v[rev16(k)], T2[k * 64 + s]);
#pragma unroll
    for (int k = 0; k < 16; k++) E[f16k_slot(ka, k, s)] = v[rev16(k)];
}

// ---- phase 3 -------------------------------------------------------------
SDRB_HD void f16k_phase3(int t, cpx* E, const cpx* T3) {
    const int d = t & 3, kb = (t >> 2) & 15, ka = t >> 6;
    cpx v[16];
#pragma unroll
    for (int c = 0; c < 16; c++) v[c] = E[f16k_slot(ka, kb, 4 * c + d)];
    fft16(v);
#pragma unroll
    for (int k = 1; k < 16; k++) v[rev16(k)] = cmul(v[rev16(k)], T3[k * 4 + d]);
#pragma unroll
    for (int k = 0; k < 16; k++) E[f16k_slot(ka, kb, 4 * k + d)] = v[rev16(k)];
}

// ---- phase 4: radix-4 over d, power, fftshift, store ----------------------
// out points at this transform's N floats; bin k of the DFT lands at (k + N/2) mod N
// (src/dsp/fft_process.cpp:92-97).
SDRB_HD void f16k_phase4(int t, const cpx* E, float* __restrict__ out, float power_scale) {
    const int ka = t & 15, kb = (t >> 4) & 15, kc0 = t >> 8;
#pragma unroll
    for (int m = 0; m < 4; m++) {
        const int kc = kc0 + 4 * m;
        const int base = f16k_slot(ka, kb, 4 * kc);  // d = 0..3 are contiguous, 16-byte aligned pairs
        cpx x0 = E[base], x1 = E[base + 1], x2 = E[base + 2], x3 = E[base + 3];
        fft4(x0, x1, x2, x3);
        const int klo = ka + 16 * kb + 256 * kc;
        // power: re*re + im*im  (src/dsp/fft_process.cpp:85)
        out[klo + 8192] = (x0.x * x0.x + x0.y * x0.y) * power_scale;   // kd = 0
        out[klo + 12288] = (x1.x * x1.x + x1.y * x1.y) * power_scale;  // kd = 1
        out[klo] = (x2.x * x2.x + x2.y * x2.y) * power_scale;          // kd = 2
        out[klo + 4096] = (x3.x * x3.x + x3.y * x3.y) * power_scale;   // kd = 3
    }
}

}  // namespace sdrb
