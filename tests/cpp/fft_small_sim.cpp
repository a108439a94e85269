// tests/cpp/fft_small_sim.cpp — CPU execution, thread by thread, of the small one-CTA FFT kernel's phase functions
// (sdr-for-android-lib_b200/csrc/fft_small.cuh) against a double-precision FFT, for every supported length
// (256 .. 8192) and for the two-pass scheme built on it (32768 points: radix-8 columns + 4096-point rows); plus an
// audit of the shared-memory bank pairs touched by a half-warp in every phase (8-byte accesses: 16 lanes per wavefront).
// Exits 0 when every spectrum matches within the float-FFT bound and the conflict degrees are the documented ones.
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../sdr-for-android-lib_b200/csrc/fft_small.cuh"
#include "../../sdr-for-android-lib_b200/csrc/fft_tables.h"

using namespace sdrb;

static void fft_double(std::vector<std::complex<double>>& a) {
    const size_t n = a.size();
    for (size_t i = 1, j = 0; i < n; i++) {
        size_t bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) std::swap(a[i], a[j]);
    }
    for (size_t len = 2; len <= n; len <<= 1) {
        const double ang = -2.0 * M_PI / (double)len;
        for (size_t i = 0; i < n; i += len)
            for (size_t k = 0; k < len / 2; k++) {
                const std::complex<double> w(std::cos(ang * (double)k), std::sin(ang * (double)k));
                const auto u = a[i + k], v = a[i + k + len / 2] * w;
                a[i + k] = u + v;
                a[i + k + len / 2] = u - v;
            }
    }
}

static std::vector<cpx> signal(int N, unsigned seed) {
    std::vector<cpx> in(N);
    unsigned s = seed;
    for (int i = 0; i < N; i++) {
        s = s * 1664525u + 1013904223u;
        const float a = ((int)(s >> 16) % 2001 - 1000) * 1e-3f;
        s = s * 1664525u + 1013904223u;
        const float b = ((int)(s >> 16) % 2001 - 1000) * 1e-3f;
        in[i] = mk(a + 0.3f * (float)std::cos(2 * M_PI * 37.25 * i / N), b + 0.3f * (float)std::sin(2 * M_PI * 37.25 * i / N));
    }
    return in;
}

static int compare(const std::vector<cpx>& in, const std::vector<float>& out, const char* what) {
    const int N = (int)in.size();
    std::vector<std::complex<double>> ref(N);
    double energy = 0;
    for (int i = 0; i < N; i++) { ref[i] = {in[i].x, in[i].y}; energy += std::norm(ref[i]); }
    fft_double(ref);
    const double total = energy * N;
    double worst = 0;
    int bad = 0;
    for (int k = 0; k < N; k++) {
        const double p = std::norm(ref[k]);
        const double got = out[(k + N / 2) & (N - 1)];
        const double tol = 1e-5 * p + 2.0 * std::sqrt(p) * 3e-7 * std::sqrt(total) + 1e-9 * total / N;
        const double e = std::fabs(got - p) / tol;
        if (e > worst) worst = e;
        if (e > 1.0 && bad++ < 5) printf("  bin %d: got %.6g want %.6g\n", k, got, p);
    }
    printf("%s N=%d worst error/tolerance %.3f, %d bad bins\n", what, N, worst, bad);
    return bad;
}

// largest number of lanes of a half-warp (16 consecutive threads) that share an 8-byte bank pair (16 pairs per 128 bytes)
template <class ADDR>
static int conflict_degree(int threads, int accesses, ADDR addr) {
    int worst = 0;
    for (int h = 0; h + 16 <= threads; h += 16)
        for (int i = 0; i < accesses; i++) {
            int cnt[16] = {0};
            for (int l = 0; l < 16; l++) {
                const int a = addr(h + l, i);
                if (a >= 0) cnt[a & 15]++;
            }
            for (int b = 0; b < 16; b++) worst = cnt[b] > worst ? cnt[b] : worst;
        }
    return worst;
}

template <int C>
static int run() {
    typedef FsGeom<C> G;
    const std::vector<cpx> in = signal(G::N, 777u + C);
    const std::vector<cpx> tab = make_fs_tables(G::N);
    const cpx* TP = tab.data();
    const cpx* T2 = TP + G::kTP;
    if ((int)tab.size() != G::kTables) { printf("table size %zu != %d\n", tab.size(), G::kTables); return 1; }
    std::vector<cpx> E(G::kE);
    std::vector<float> out(G::N, -1.0f);
    for (int t = 0; t < G::T; t++) fs_phase1<FMT_CF32, C>(t, in.data(), TP, E.data());
    for (int t = 0; t < G::T; t++) fs_phase2<C>(t, E.data());
    for (int t = 0; t < G::T; t++) fs_phase3<C>(t, E.data(), T2, out.data(), FsOutPlain());
    int bad = compare(in, out, "small");
    for (int k = 0; k < G::N; k++)
        if (out[k] < 0.0f) { printf("  bin %d never written\n", k); bad++; break; }
    // bank audit (T >= 16 threads)
    const int d1 = conflict_degree(G::T, 16, [](int t, int k) { return k * G::kPitch + t; });
    const int d2 = conflict_degree(G::T, 16, [](int t, int b) {
        const int ka = C < 16 ? ((t & 15) / C) * C + (t >> 4) : t / C, c = C < 16 ? (t & 15) % C : t % C;
        return ka * G::kPitch + c + b * C;
    });
    const int pairs = 16 / C;
    const int d3 = conflict_degree(G::T, pairs * C, [&](int t, int i) {
        const int p = t + G::T * (i / C), c = i % C;
        return (p & 15) * G::kPitch + (p >> 4) * C + c;
    });
    printf("  bank pairs shared by a half-warp: phase 1 stores %d, phase 2 %d, phase 3 loads %d\n", d1, d2, d3);
    if (d1 != 1 || d2 != 1 || d3 != 1) { printf("  unexpected conflict degree\n"); bad++; }
    return bad;
}

// 8192 points: the 32 * 16 * 16 kernel
static int run8k() {
    const std::vector<cpx> in = signal(Fs8k::N, 8192u);
    const std::vector<cpx> tab = make_fs_tables(Fs8k::N);
    if ((int)tab.size() != Fs8k::kTables) { printf("8k table size %zu != %d\n", tab.size(), Fs8k::kTables); return 1; }
    const cpx* TP = tab.data();
    const cpx* T2 = TP + Fs8k::kTP;
    std::vector<cpx> E(Fs8k::kE);
    std::vector<float> out(Fs8k::N, -1.0f);
    for (int t = 0; t < 256; t++) fs8k_phase1<FMT_CF32>(t, in.data(), TP, E.data());
    for (int t = 0; t < 256; t++) fs8k_phase2(t, E.data());
    for (int t = 0; t < 256; t++) fs8k_phase3(t, E.data(), T2, out.data(), FsOutPlain());
    int bad = compare(in, out, "8k");
    for (int k = 0; k < Fs8k::N; k++)
        if (out[k] < 0.0f) { printf("  bin %d never written\n", k); bad++; break; }
    const int d1 = conflict_degree(256, 32, [](int t, int k) { return k * Fs8k::kPitch + t; });
    const int d2 = conflict_degree(256, 32, [](int t, int i) { return ((t >> 4) + 16 * (i / 16)) * Fs8k::kPitch + (t & 15) + (i % 16) * 16; });
    const int d3 = conflict_degree(256, 32, [](int t, int i) {
        const int p = t + 256 * (i / 16);
        return (p & 31) * Fs8k::kPitch + (p >> 5) * 16 + (i % 16);
    });
    printf("  bank pairs shared by a half-warp: phase 1 stores %d, phase 2 %d, phase 3 loads %d\n", d1, d2, d3);
    if (d1 != 1 || d2 != 1 || d3 != 1) { printf("  unexpected conflict degree\n"); bad++; }
    return bad;
}

// two-pass scheme, 32768 points: radix-8 columns, then 4096-point rows with interleaved output
static int run_twopass() {
    constexpr int LOGA = 3, A = 8, B = 4096, N = A * B;
    const std::vector<cpx> in = signal(N, 4242u);
    const std::vector<cpx> tab = make_tp_tables(N, B, LOGA);
    const cpx* TPc = tab.data();
    const cpx* TP = TPc + (size_t)LOGA * B;
    const cpx* T2 = TP + FsGeom<16>::kTP;
    std::vector<cpx> Y(N), E(FsGeom<16>::kE);
    std::vector<float> out(N, -1.0f);
    for (int b = 0; b < B; b++) fs_cols<FMT_CF32, LOGA>(b, B, in.data(), TPc, Y.data());
    for (int ka = 0; ka < A; ka++) {
        for (int t = 0; t < 256; t++) fs_phase1<FMT_CF32, 16>(t, Y.data() + (size_t)ka * B, TP, E.data());
        for (int t = 0; t < 256; t++) fs_phase2<16>(t, E.data());
        const FsOutStrided where = {A, ka};
        for (int t = 0; t < 256; t++) fs_phase3<16>(t, E.data(), T2, out.data(), where);
    }
    return compare(in, out, "two-pass");
}

int main() {
    int bad = 0;
    bad += run<1>();
    bad += run<2>();
    bad += run<4>();
    bad += run<8>();
    bad += run<16>();
    bad += run8k();
    bad += run_twopass();
    printf("%s\n", bad ? "FAILED" : "all sizes ok");
    return bad ? 1 : 0;
}
