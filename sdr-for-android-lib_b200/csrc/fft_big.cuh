// fft_big.cuh — multi-CTA four-step FFT for N = L * L, L = R * R  (R = 16: N = 65536; R = 32:
// N = 1048576).  Replaces fftwf_execute at src/dsp/fft_process.cpp:77-79 for the large block
// lengths of BASELINE configs C3 (64K) and C5 (1M); ingest convert, |X|^2 and fftshift are fused
// like in the 16K kernel (reference :62-97).
//
//   n = L n1 + n2,  k = k1 + L k2                      (n1, n2, k1, k2 < L)
//   kernel A ("columns"): for a tile of 16 columns n2: L-point DFT over n1, times W_N^(n2 k1),
//                         written to the intermediate A'
//   kernel B ("rows")   : for a group of 16 rows k1: L-point DFT over n2, |X|^2, fftshift, store
//
// Each L-point DFT is two radix-R passes in registers (R complex values per thread, 16 R threads
// per CTA, 16 interleaved transforms per CTA) with ONE in-place exchange through shared memory.
// Intermediate layout  A'[k1 / 16][n2][k1 % 16]  (complex): kernel A's final stores run along
// k1 % 16 (128-byte segments) and kernel B's whole 16-row group is one contiguous 128 L byte range
// that lands in shared memory, already in working layout [n2][row], by a single bulk TMA copy.
//
// All phase functions are __host__ __device__ and take the thread index explicitly so that
// tests/cpp/fft_big_sim.cpp can execute them thread by thread on the CPU against a double DFT.
#pragma once
#include <stdint.h>

#include "fft16k2.cuh"
#include "fft_core.cuh"

namespace sdrb {

template <int R>
struct RadixReg;
template <>
struct RadixReg<16> {
    static constexpr int kLog2 = 4;
    SDRB_HD static void fft(cpx* v) { fft_dit<4>(v); }   // FMA-form decimation in time (fft_dit.cuh)
    SDRB_HD static constexpr int slot(int k) { return dit_slot<4>(k); }
};
template <>
struct RadixReg<32> {
    static constexpr int kLog2 = 5;
    SDRB_HD static void fft(cpx* v) { fft_dit<5>(v); }
    SDRB_HD static constexpr int slot(int k) { return dit_slot<5>(k); }
};

template <int R>
struct FbGeom {
    static constexpr int L = R * R;                 // points per pass-level transform
    static constexpr int kThreads = 16 * R;         // 16 interleaved transforms x R butterflies
    static constexpr int kBlockPitchA = R * 16 + 1; // kernel A: odd pitch per b1 block (complex)
    static constexpr int kElemsA = R * kBlockPitchA;
    static constexpr int kElemsB = L * 16;          // kernel B: dense [n2][row], bulk-copied
    // twiddle tables, device layout (complex):  TL | TLO | THI | TB
    static constexpr int kTL = L;                   // W_L^i
    static constexpr int kTLO = 256;                // W_N^i,        i < 256
    static constexpr int kTHI = L * R / 256;        // W_N^(256 i),  i < L R / 256   (m = n2 b1 < L R)
    static constexpr int kTB = RadixReg<R>::kLog2 * L;  // [j][n2]  W_N^(n2 R 2^j)
    static constexpr int kTables = kTL + kTLO + kTHI + kTB;
};

// ---- kernel A, pass 0: column c = t % 16 of the tile, butterfly q = t / 16.
// Loads x[L (q + R r) + n2] for r < R straight from global memory (ingest fused), DFT_R,
// twiddle W_L^(q k), stores Y[k] to S[b1 = k][q][c].
template <int R, int FMT>
SDRB_HD void fbA_pass0(int t, const typename Ingest2<FMT>::raw_t* frame, int n2_0, const cpx* TL, cpx* S) {
    typedef FbGeom<R> G;
    const int c = t & 15, q = t >> 4;
    cpx v[R];
#pragma unroll
    for (int r = 0; r < R; r++) v[r] = Ingest2<FMT>::cvt(frame[(size_t)G::L * (q + R * r) + n2_0 + c]);
    RadixReg<R>::fft(v);
#pragma unroll
    for (int k = 0; k < R; k++) {
        cpx y = v[RadixReg<R>::slot(k)];
        if (k > 0) y = cmul(y, TL[q * k]);
        S[k * G::kBlockPitchA + q * 16 + c] = y;
    }
}

// ---- kernel A, pass 1: block b1 = t % R, column c = t / R.  DFT_R over q, then the four-step
// twiddle W_N^(n2 k1), k1 = b1 + R k', and the store to A'[k1 / 16][n2][k1 % 16].
template <int R>
SDRB_HD void fbA_pass1(int t, const cpx* S, const cpx* TLO, const cpx* THI, const cpx* TB, int n2_0,
                       cpx* __restrict__ Aframe) {
    typedef FbGeom<R> G;
    const int b1 = t % R, c = t / R, n2 = n2_0 + c;
    cpx v[R];
#pragma unroll
    for (int q = 0; q < R; q++) v[q] = S[b1 * G::kBlockPitchA + q * 16 + c];
    RadixReg<R>::fft(v);
    // w[k'] = W_N^(n2 b1) * (W_N^(n2 R))^k' by a product tree over the binary digits of k'
    const int m = n2 * b1;
    cpx P[RadixReg<R>::kLog2];
#pragma unroll
    for (int j = 0; j < RadixReg<R>::kLog2; j++) P[j] = TB[j * G::L + n2];
    cpx w[R];
    w[0] = cmul(TLO[m & 255], THI[m >> 8]);
#pragma unroll
    for (int k = 1; k < R; k++) {
        const int low = k & (-k);
        const int j = low == 1 ? 0 : (low == 2 ? 1 : (low == 4 ? 2 : (low == 8 ? 3 : 4)));
        w[k] = cmul(w[k - low], P[j]);
    }
#pragma unroll
    for (int k = 0; k < R; k++) {
        const int k1 = b1 + R * k;
        Aframe[((size_t)(k1 >> 4) * G::L + n2) * 16 + (k1 & 15)] = cmul(v[RadixReg<R>::slot(k)], w[k]);
    }
}

// ---- kernel B, pass 0: row rho = t % 16 of the group, butterfly q = t / 16; in place.
template <int R>
SDRB_HD void fbB_pass0(int t, cpx* W, const cpx* TL) {
    const int rho = t & 15, q = t >> 4;
    cpx v[R];
#pragma unroll
    for (int r = 0; r < R; r++) v[r] = W[(q + R * r) * 16 + rho];
    RadixReg<R>::fft(v);
#pragma unroll
    for (int k = 0; k < R; k++) {
        cpx y = v[RadixReg<R>::slot(k)];
        if (k > 0) y = cmul(y, TL[q * k]);
        W[(q + R * k) * 16 + rho] = y;
    }
}

// ---- kernel B, pass 1: row rho = t % 16, block b1 = t / 16.  DFT_R over q gives
// X[k1 + L k2], k2 = b1 + R k'; power (:85) and fftshift (:92-97) folded into the store.
template <int R>
SDRB_HD void fbB_pass1_load(int t, const cpx* W, cpx* v) {
    const int rho = t & 15, b1 = t >> 4;
#pragma unroll
    for (int q = 0; q < R; q++) v[q] = W[(q + R * b1) * 16 + rho];
}
template <int R>
SDRB_HD void fbB_pass1_finish(int t, cpx* v, int k1_0, float* __restrict__ out_frame) {
    typedef FbGeom<R> G;
    constexpr int N = G::L * G::L;
    const int rho = t & 15, b1 = t >> 4;
    RadixReg<R>::fft(v);
#pragma unroll
    for (int k = 0; k < R; k++) {
        const cpx x = v[RadixReg<R>::slot(k)];
        const int bin = (k1_0 + rho) + G::L * (b1 + R * k);
        out_frame[(bin + N / 2) & (N - 1)] = x.x * x.x + x.y * x.y;
    }
}
template <int R>
SDRB_HD void fbB_pass1(int t, const cpx* W, int k1_0, float* __restrict__ out_frame) {
    cpx v[R];
    fbB_pass1_load<R>(t, W, v);
    fbB_pass1_finish<R>(t, v, k1_0, out_frame);
}

}  // namespace sdrb
