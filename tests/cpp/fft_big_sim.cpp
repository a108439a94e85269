// tests/cpp/fft_big_sim.cpp — CPU execution, thread by thread, of the four-step FFT kernels' phase
// functions (sdr-for-android-lib_b200/csrc/fft_big.cuh) against a double-precision FFT.
// Usage: fft_big_sim <R: 16|32>   exits 0 when the spectrum matches within the float-FFT bound.
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <vector>

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

template <int R>
static int run() {
    typedef FbGeom<R> G;
    const int L = G::L, N = L * L;
    std::vector<cpx> in(N);
    unsigned s = 12345u;
    for (int i = 0; i < N; i++) {
        s = s * 1664525u + 1013904223u;
        const float a = ((int)(s >> 16) % 2001 - 1000) * 1e-3f;
        s = s * 1664525u + 1013904223u;
        const float b = ((int)(s >> 16) % 2001 - 1000) * 1e-3f;
        in[i] = mk(a + 0.3f * (float)std::cos(2 * M_PI * 1234.25 * i / N), b + 0.3f * (float)std::sin(2 * M_PI * 1234.25 * i / N));
    }
    const std::vector<cpx> tab = make_fb_tables<R>();
    const cpx* TL = tab.data();
    const cpx* TLO = TL + G::kTL;
    const cpx* THI = TLO + G::kTLO;
    const cpx* TB = THI + G::kTHI;
    std::vector<cpx> A(N), S(G::kElemsA), W(G::kElemsB);
    std::vector<float> out(N, -1.0f);
    for (int tile = 0; tile < L / 16; tile++) {       // kernel A: one CTA per 16 columns
        for (int t = 0; t < G::kThreads; t++) fbA_pass0<R, FMT_CF32>(t, in.data(), tile * 16, TL, S.data());
        for (int t = 0; t < G::kThreads; t++) fbA_pass1<R>(t, S.data(), TLO, THI, TB, tile * 16, A.data());
    }
    for (int g = 0; g < L / 16; g++) {                // kernel B: one CTA per 16 rows
        for (int i = 0; i < G::kElemsB; i++) W[i] = A[(size_t)g * G::kElemsB + i];   // the bulk copy
        // pass 0 is in place per thread (reads and writes the same positions): sequential is faithful
        for (int t = 0; t < G::kThreads; t++) fbB_pass0<R>(t, W.data(), TL);
        for (int t = 0; t < G::kThreads; t++) fbB_pass1<R>(t, W.data(), g * 16, out.data());
    }
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
    printf("R=%d N=%d worst error/tolerance %.3f, %d bad bins\n", R, N, worst, bad);
    return bad ? 1 : 0;
}

int main(int argc, char** argv) {
    const int R = argc > 1 ? atoi(argv[1]) : 16;
    return R == 32 ? run<32>() : run<16>();
}
