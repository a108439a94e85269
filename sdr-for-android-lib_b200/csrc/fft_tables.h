// fft_tables.h — host-side twiddle table generation (double precision, rounded once).
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>

#include "fft16k2.cuh"
#include "fft_big.cuh"

namespace sdrb {

inline cpx twiddle(long long num, long long den) {
    // W_den^num = exp(-2 pi i num / den), reduced first so the angle stays small and exact
    num %= den;
    const double a = -2.0 * 3.14159265358979323846264338327950288 * (double)num / (double)den;
    return mk((float)std::cos(a), (float)std::sin(a));
}

// Layout: TP | T2  (F2_TW_ELEMS entries), see fft16k2.cuh
inline std::vector<cpx> make_f2_tables() {
    std::vector<cpx> t(F2_TW_ELEMS);
    cpx* TP = t.data();
    cpx* T2 = TP + F2_TP;
    for (int j = 0; j < 5; j++)
        for (int r = 0; r < 512; r++) TP[j * 512 + r] = twiddle((long long)r << j, F2_N);
    for (int kb = 0; kb < 32; kb++)
        for (int c = 0; c < 16; c++) T2[kb * 16 + c] = twiddle((long long)c * kb, 512);
    return t;
}

// Small one-CTA kernel (fft_small.cuh), N = 256 C: TP[j][t] = W_N^(t 2^j) (j < 4, t < 16 C), then T2[kb][c] = W_16C^(c kb)
inline std::vector<cpx> make_fs_tables(int n) {
    if (n == 8192) {   // the 32 * 16 * 16 kernel (Fs8k): TP[j][t] = W_N^(t 2^j), j < 5, t < 256; T2[kb][c] = W_256^(c kb)
        std::vector<cpx> t(5 * 256 + 256);
        for (int j = 0; j < 5; j++)
            for (int r = 0; r < 256; r++) t[(size_t)j * 256 + r] = twiddle((long long)r << j, n);
        for (int kb = 0; kb < 16; kb++)
            for (int c = 0; c < 16; c++) t[5 * 256 + kb * 16 + c] = twiddle((long long)c * kb, 256);
        return t;
    }
    const int C = n / 256, T = 16 * C;
    std::vector<cpx> t((size_t)4 * T + 16 * C);
    for (int j = 0; j < 4; j++)
        for (int r = 0; r < T; r++) t[(size_t)j * T + r] = twiddle((long long)r << j, n);
    for (int kb = 0; kb < 16; kb++)
        for (int c = 0; c < C; c++) t[(size_t)4 * T + kb * C + c] = twiddle((long long)c * kb, 16 * C);
    return t;
}

// Two-pass scheme (fft_small.cu), N = A B: TPc[j][b] = W_N^(b 2^j), j < log2 A, then the small kernel's tables for B points
inline std::vector<cpx> make_tp_tables(int n, int B, int loga) {
    std::vector<cpx> t((size_t)loga * B);
    for (int j = 0; j < loga; j++)
        for (int b = 0; b < B; b++) t[(size_t)j * B + b] = twiddle((long long)b << j, n);
    const std::vector<cpx> s = make_fs_tables(B);
    t.insert(t.end(), s.begin(), s.end());
    return t;
}

// Layout: TL | TLO | THI | TB  (FbGeom<R>::kTables entries), see fft_big.cuh
template <int R>
inline std::vector<cpx> make_fb_tables() {
    typedef FbGeom<R> G;
    const long long N = (long long)G::L * G::L;
    std::vector<cpx> t(G::kTables);
    cpx* TL = t.data();
    cpx* TLO = TL + G::kTL;
    cpx* THI = TLO + G::kTLO;
    cpx* TB = THI + G::kTHI;
    for (int i = 0; i < G::kTL; i++) TL[i] = twiddle(i, G::L);
    for (int i = 0; i < G::kTLO; i++) TLO[i] = twiddle(i, N);
    for (int i = 0; i < G::kTHI; i++) THI[i] = twiddle(256LL * i, N);
    for (int j = 0; j < RadixReg<R>::kLog2; j++)
        for (int n2 = 0; n2 < G::L; n2++) TB[j * G::L + n2] = twiddle(((long long)n2 * R) << j, N);
    return t;
}

// Bluestein tables for length n, convolution length m (a power of two >= 2n - 1): chirp c[k] = exp(-i pi k^2 / n)
// with k^2 reduced mod 2n in integers, and bf = FFT_m of b, b[j] = b[m - j] = conj(c[j]) for j < n, in double.
inline void make_bluestein_tables(int n, int m, std::vector<cpx>& chirp, std::vector<cpx>& bf) {
    const double pi = 3.14159265358979323846264338327950288;
    std::vector<double> cr((size_t)n), ci((size_t)n);
    chirp.resize((size_t)n);
    for (long long k = 0; k < n; k++) {
        const long long q = (k * k) % (2LL * n);
        const double a = -pi * (double)q / (double)n;
        cr[(size_t)k] = std::cos(a);
        ci[(size_t)k] = std::sin(a);
        chirp[(size_t)k] = mk((float)cr[(size_t)k], (float)ci[(size_t)k]);
    }
    std::vector<double> re((size_t)m, 0.0), im((size_t)m, 0.0);
    for (int j = 0; j < n; j++) {
        re[(size_t)j] = cr[(size_t)j];
        im[(size_t)j] = -ci[(size_t)j];
        if (j) { re[(size_t)(m - j)] = cr[(size_t)j]; im[(size_t)(m - j)] = -ci[(size_t)j]; }
    }
    // in-place iterative radix-2 FFT in double (bit reversal + butterflies)
    int bits = 0;
    while ((1 << bits) < m) bits++;
    for (int i = 0; i < m; i++) {
        unsigned r = 0;
        for (int b = 0; b < bits; b++) if (i & (1 << b)) r |= 1u << (bits - 1 - b);
        if ((int)r > i) { std::swap(re[(size_t)i], re[r]); std::swap(im[(size_t)i], im[r]); }
    }
    for (int len = 2; len <= m; len <<= 1) {
        const int half = len >> 1;
        for (int j = 0; j < half; j++) {
            const double a = -2.0 * pi * (double)j / (double)len;
            const double wr = std::cos(a), wi = std::sin(a);
            for (int base = 0; base < m; base += len) {
                const size_t u = (size_t)(base + j), v = u + (size_t)half;
                const double tr = re[v] * wr - im[v] * wi, ti = re[v] * wi + im[v] * wr;
                re[v] = re[u] - tr; im[v] = im[u] - ti;
                re[u] += tr; im[u] += ti;
            }
        }
    }
    bf.resize((size_t)m);
    for (int i = 0; i < m; i++) bf[(size_t)i] = mk((float)re[(size_t)i], (float)im[(size_t)i]);
}

}  // namespace sdrb
