// fft_small.cuh — one-CTA-per-transform shared-memory FFT for N = 256 C, C in {1, 2, 4, 8, 16} (256 .. 4096 points), and
// for 8192 points (Fs8k below):
// the 16K kernel's structure for the other power-of-two block lengths up to 8192
// (reference: src/dsp/fft_process.cpp:62-97 — copy, fftwf_execute of samplesPerReading points, power, fftshift).
//
//   n = 16 C a + C b + c          k = ka + 16 kb + 256 kc          (a, b, ka, kb < 16; c, kc < C)
//   phase 1: DFT16 over a   thread t = C b + c  (T = 16 C threads)   twiddle W_N^(t ka)
//   phase 2: DFT16 over b   thread t = C ka + c                      (twiddle W_16C^(c kb) applied by phase 3 on load)
//   phase 3: DFT_C over c   pairs p = ka + 16 kb: thread u takes p = u + T j, j < 16 / C, then |X|^2, fftshift folded
//            into the store index: for a fixed kc a warp's 32 pairs are 32 consecutive bins
// Exchange buffer E[ka][16 C + 1]: the 8-byte accesses of a half-warp fall on 16 distinct bank pairs in every phase
// and for every C (phase 3 reads 16 consecutive planes: the pitch is odd; phase 2 picks its planes accordingly);
// audited in tests/cpp/fft_small_sim.cpp.
// Every thread holds 16 points in every phase; small transforms share a CTA (256 / T of them).
// The register codelets are the FMA-form decimation-in-time ones of fft_dit.cuh.
// Phase functions are __host__ __device__ so that tests/cpp/fft_small_sim.cpp runs them thread by thread.
#pragma once
#include <stdint.h>

#include "fft16k2.cuh"
#include "fft_core.cuh"
#include "fft_dit.cuh"

namespace sdrb {

template <int C>
struct FsGeom {
    static constexpr int N = 256 * C;
    static constexpr int T = 16 * C;                     // threads per transform
    static constexpr int kPitch = 16 * C + 1;            // complex elements per ka plane
    static constexpr int kE = 16 * kPitch;               // exchange buffer of one transform
    static constexpr int kPerCta = T >= 256 ? 1 : 256 / T;   // transforms per CTA
    static constexpr int kThreads = T * kPerCta;         // 256
    static constexpr int kTP = 4 * T;                    // [j][t]   W_N^(t 2^j)
    static constexpr int kT2 = 16 * C;                   // [kb][c]  W_16C^(c kb)
    static constexpr int kTables = kTP + kT2;
    static constexpr int kLogC = C == 1 ? 0 : C == 2 ? 1 : C == 4 ? 2 : C == 8 ? 3 : 4;
    static_assert(C == 1 || C == 2 || C == 4 || C == 8 || C == 16, "256 .. 4096 points");
};

// ---- phase 1: load + convert, DFT16 over a, twiddle W_N^(t ka), store E[ka][t]
// (in two halves so that a kernel can put the extension front end - block mean, window - between them)
template <int FMT, int C>
SDRB_HD void fs_phase1_load(int t, const typename Ingest2<FMT>::raw_t* in, cpx* v) {
    typedef FsGeom<C> G;
#pragma unroll
    for (int a = 0; a < 16; a++) v[a] = Ingest2<FMT>::cvt(in[G::T * a + t]);
}
template <int C>
SDRB_HD void fs_phase1_finish(int t, cpx* v, const cpx* TP, cpx* E) {
    typedef FsGeom<C> G;
    fft_dit<4>(v);
    // W_N^(t ka) from W_N^(t 2^j): w[k] = w[k with lowest set bit cleared] * P[ctz k]
    cpx P[4];
#pragma unroll
    for (int j = 0; j < 4; j++) P[j] = TP[j * G::T + t];
    cpx w[16];
    E[t] = v[0];
#pragma unroll
    for (int k = 1; k < 16; k++) {
        const int low = k & (-k);
        const int j = low == 1 ? 0 : (low == 2 ? 1 : (low == 4 ? 2 : 3));
        w[k] = (k == low) ? P[j] : cmul(w[k - low], P[j]);
        E[k * G::kPitch + t] = cmul(v[dit_slot<4>(k)], w[k]);
    }
}

template <int FMT, int C>
SDRB_HD void fs_phase1(int t, const typename Ingest2<FMT>::raw_t* in, const cpx* TP, cpx* E) {
    cpx v[16];
    fs_phase1_load<FMT, C>(t, in, v);
    fs_phase1_finish<C>(t, v, TP, E);
}

// ---- phase 2: DFT16 over b for (ka, c), in place
template <int C>
SDRB_HD void fs_phase2(int t, cpx* E) {
    typedef FsGeom<C> G;
    // which (ka, c) a thread takes is free (the pass is in place per pair).  For C < 16 a half-warp spans 16 / C planes:
    // lane l of half-warp h takes c = l % C and plane ka = (l / C) C + h, so that (ka + c) mod 16 = (l + h) mod 16 runs
    // over all 16 bank pairs (consecutive planes would fall on top of each other: four-way conflicts at C = 4)
    int ka, c;
    if (C < 16) {
        const int h = t >> 4, l = t & 15;
        c = l % C;
        ka = (l / C) * C + h;
    } else {
        ka = t / C;
        c = t % C;
    }
    cpx* base = E + ka * G::kPitch + c;
    cpx v[16];
#pragma unroll
    for (int b = 0; b < 16; b++) v[b] = base[b * C];
    fft_dit<4>(v);
#pragma unroll
    for (int k = 0; k < 16; k++) base[k * C] = v[dit_slot<4>(k)];
}

// ---- phase 3: twiddle W_16C^(c kb) on load, DFT_C over c, power, fftshift, store
// Where bin k of the transform goes.  Plain: the fft-shifted spectrum of an N-point frame.  Strided (the rows pass of
// the two-pass scheme below): the transform is row `off` of `stride` rows, its bin k is bin off + stride k of the big
// frame of stride N points.
struct FsOutPlain {
    SDRB_HD int operator()(int k, int n) const { return (k + n / 2) & (n - 1); }
};
struct FsOutStrided {
    int stride, off;
    SDRB_HD int operator()(int k, int n) const { return (off + stride * k + (stride * n) / 2) & (stride * n - 1); }
};

struct FsOutKc {   // 4096 points, one pair per thread: bin p + 256 kc goes to slot kc of the thread's own array
    SDRB_HD int operator()(int k, int) const { return k >> 8; }
};

template <int C, class OUT>
SDRB_HD void fs_phase3_pair(int p, const cpx* E, const cpx* T2, float* __restrict__ out, OUT where) {
    typedef FsGeom<C> G;
    const int ka = p & 15, kb = p >> 4;
    const cpx* base = E + ka * G::kPitch + kb * C;
    cpx v[C];
    v[0] = base[0];
#pragma unroll
    for (int c = 1; c < C; c++) v[c] = cmul(base[c], T2[kb * C + c]);
    if constexpr (G::kLogC > 0) fft_dit<G::kLogC>(v);
#pragma unroll
    for (int kc = 0; kc < C; kc++) {
        const cpx x = v[dit_slot<G::kLogC>(kc)];
        // re*re + im*im (src/dsp/fft_process.cpp:85); bin k lands at (k + N/2) mod N (:92-97)
        out[where(p + 256 * kc, G::N)] = x.x * x.x + x.y * x.y;
    }
}
template <int C, class OUT>
SDRB_HD void fs_phase3(int t, const cpx* E, const cpx* T2, float* __restrict__ out, OUT where) {
    typedef FsGeom<C> G;
#pragma unroll
    for (int j = 0; j < 16 / C; j++) fs_phase3_pair<C>(t + G::T * j, E, T2, out, where);
}

// ---- 8192 points with every thread busy in every phase: N = 32 * 16 * 16, 256 threads x 32 points (the 16K kernel's
// shape with a 16-point middle phase; 512 threads x 16 points would leave half of them idle in a 32-point last phase).
//   n = 256 a + 16 b + c          k = ka + 32 kb + 512 kc          (a, ka < 32; b, c, kb, kc < 16)
//   phase 1: DFT32 over a   thread t = 16 b + c               twiddle W_N^(t ka), E[ka][t] (pitch 257)
//   phase 2: DFT16 over b   thread t -> (ka, c) and (ka + 16, c), ka = t / 16, c = t % 16, in place
//   phase 3: DFT16 over c   pairs p = t and t + 256: ka = p & 31, kb = p >> 5, twiddle W_256^(c kb) on load
struct Fs8k {
    static constexpr int N = 8192, T = 256, kPitch = 257, kE = 32 * kPitch;
    static constexpr int kTP = 5 * 256;      // [j][t]   W_N^(t 2^j)
    static constexpr int kT2 = 16 * 16;      // [kb][c]  W_256^(c kb)
    static constexpr int kTables = kTP + kT2;
};
template <int FMT>
SDRB_HD void fs8k_phase1_load(int t, const typename Ingest2<FMT>::raw_t* in, cpx* v) {
#pragma unroll
    for (int a = 0; a < 32; a++) v[a] = Ingest2<FMT>::cvt(in[256 * a + t]);
}
SDRB_HD void fs8k_phase1_finish(int t, cpx* v, const cpx* TP, cpx* E) {
    fft_dit<5>(v);
    cpx P[5];
#pragma unroll
    for (int j = 0; j < 5; j++) P[j] = TP[j * 256 + t];
    cpx w[32];
    E[t] = v[0];
#pragma unroll
    for (int k = 1; k < 32; k++) {
        const int low = k & (-k);
        const int j = low == 1 ? 0 : (low == 2 ? 1 : (low == 4 ? 2 : (low == 8 ? 3 : 4)));
        w[k] = (k == low) ? P[j] : cmul(w[k - low], P[j]);
        E[k * Fs8k::kPitch + t] = cmul(v[dit_slot<5>(k)], w[k]);
    }
}
template <int FMT>
SDRB_HD void fs8k_phase1(int t, const typename Ingest2<FMT>::raw_t* in, const cpx* TP, cpx* E) {
    cpx v[32];
    fs8k_phase1_load<FMT>(t, in, v);
    fs8k_phase1_finish(t, v, TP, E);
}
SDRB_HD void fs8k_phase2(int t, cpx* E) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
        cpx* base = E + ((t >> 4) + 16 * h) * Fs8k::kPitch + (t & 15);
        cpx v[16];
#pragma unroll
        for (int b = 0; b < 16; b++) v[b] = base[b * 16];
        fft_dit<4>(v);
#pragma unroll
        for (int k = 0; k < 16; k++) base[k * 16] = v[dit_slot<4>(k)];
    }
}
template <class OUT>
SDRB_HD void fs8k_phase3(int t, const cpx* E, const cpx* T2, float* __restrict__ out, OUT where) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int p = t + 256 * h;
        const int ka = p & 31, kb = p >> 5;
        const cpx* base = E + ka * Fs8k::kPitch + kb * 16;
        cpx v[16];
        v[0] = base[0];
#pragma unroll
        for (int c = 1; c < 16; c++) v[c] = cmul(base[c], T2[kb * 16 + c]);
        fft_dit<4>(v);
#pragma unroll
        for (int kc = 0; kc < 16; kc++) {
            const cpx x = v[dit_slot<4>(kc)];
            out[where(p + 512 * kc, Fs8k::N)] = x.x * x.x + x.y * x.y;
        }
    }
}

// ---- two-pass scheme for N = A B, B = 4096 or 8192, A = 2^LOGA <= 32 (32768 .. 262144 points):
//   n = B a + b,  k = ka + A kb        X[ka + A kb] = sum_b W_B^(b kb) [ W_N^(b ka) sum_a x[B a + b] W_A^(a ka) ]
//   columns pass: one thread per (frame, b): DFT_A over a in registers, twiddle W_N^(b ka), Y[frame][ka][b] (cf32)
//   rows pass   : the one-CTA kernel above on the A rows of every frame, bin kb of row ka stored at ka + A kb
template <int FMT, int LOGA>
SDRB_HD void fs_cols(int b, int B, const typename Ingest2<FMT>::raw_t* frame, const cpx* TPc, cpx* __restrict__ Yframe) {
    constexpr int A = 1 << LOGA;
    cpx v[A];
#pragma unroll
    for (int a = 0; a < A; a++) v[a] = Ingest2<FMT>::cvt(frame[(size_t)B * a + b]);
    fft_dit<LOGA>(v);
    cpx P[LOGA];
#pragma unroll
    for (int j = 0; j < LOGA; j++) P[j] = TPc[(size_t)j * B + b];
    cpx w[A];
    Yframe[b] = v[0];
#pragma unroll
    for (int k = 1; k < A; k++) {
        const int low = k & (-k);
        const int j = low == 1 ? 0 : (low == 2 ? 1 : (low == 4 ? 2 : (low == 8 ? 3 : 4)));
        w[k] = (k == low) ? P[j] : cmul(w[k - low], P[j]);
        Yframe[(size_t)k * B + b] = cmul(v[dit_slot<LOGA>(k)], w[k]);
    }
}

}  // namespace sdrb
