// tests/cpp/fft_sim3.cpp — CPU execution, thread by thread, of the packed-math 16K kernel's phase
// functions (sdr-for-android-lib_b200/csrc/fft16k3.cuh) against a double-precision FFT, plus a
// shared-memory bank audit of every access pattern.  Exit code 0 = pass.
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <set>
#include <vector>

#include "fft_tables_exp.h"

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

// worst number of distinct 4-byte words mapping to one bank within a group of lanes
static int worst_conflict(const std::vector<int>& word_addr) {
    int worst = 0;
    for (int bank = 0; bank < 32; bank++) {
        std::set<int> words;
        for (int w : word_addr) if ((w & 31) == bank) words.insert(w);
        if ((int)words.size() > worst) worst = (int)words.size();
    }
    return worst;
}

static int audit() {
    int bad = 0;
    for (int warp = 0; warp < 16; warp++) {
        // phase 1 store: 4-byte, all 32 lanes in one wavefront
        std::vector<int> a;
        for (int lane = 0; lane < 32; lane++) {
            const int t = warp * 32 + lane, b = t >> 4, c = t & 15;
            a.push_back(((b >> 1) * 16 + c) * 2 + (b & 1));
        }
        if (worst_conflict(a) != 1) { printf("phase-1 store conflict, warp %d\n", warp); bad++; }
        // phase 2 load/store: 8-byte, half-warp per wavefront
        for (int half = 0; half < 2; half++) {
            std::vector<int> w2, w3;
            for (int lane = half * 16; lane < half * 16 + 16; lane++) {
                const int t = warp * 32 + lane;
                const int u2 = f3_unit(t >> 4, 3, t & 15), u3 = f3_unit(t & 31, t >> 5, 5);
                w2.push_back(2 * u2); w2.push_back(2 * u2 + 1);
                w3.push_back(2 * u3); w3.push_back(2 * u3 + 1);
            }
            if (worst_conflict(w2) != 1) { printf("phase-2 conflict, warp %d\n", warp); bad++; }
            if (worst_conflict(w3) != 1) { printf("phase-3 conflict, warp %d\n", warp); bad++; }
        }
    }
    return bad;
}

int main() {
    const int N = F3_N;
    std::vector<cpx> in(N);
    unsigned s = 777u;
    for (int i = 0; i < N; i++) {
        s = s * 1664525u + 1013904223u;
        const float a = ((int)(s >> 16) % 2001 - 1000) * 1e-3f;
        s = s * 1664525u + 1013904223u;
        const float b = ((int)(s >> 16) % 2001 - 1000) * 1e-3f;
        in[i] = mk(a + 0.4f * (float)std::cos(2 * M_PI * 321.5 * i / N), b + 0.4f * (float)std::sin(2 * M_PI * 321.5 * i / N));
    }
    const std::vector<unsigned char> tab = make_f3_tables();
    const cpx* TP = reinterpret_cast<const cpx*>(tab.data());
    const pw4* T2 = reinterpret_cast<const pw4*>(tab.data() + F3_TP * 8);
    std::vector<pf> RE(F3_PLANE_UNITS), IM(F3_PLANE_UNITS);
    std::vector<float> out(N, -1.0f);
    for (int t = 0; t < F3_THREADS; t++) {
        c2 v[16];
        f3_phase1_load<FMT_CF32>(t, in.data(), v);
        c2 z[16];   // the kernel's sequence: DFT, then twiddle + store
        f3_phase1_fft(v, z);
        f3_phase1_twiddle_store(t, z, TP, reinterpret_cast<float*>(RE.data()), reinterpret_cast<float*>(IM.data()));
    }
    for (int t = 0; t < F3_THREADS; t++) f3_phase2(t, RE.data(), IM.data());   // in place per thread
    for (int t = 0; t < F3_THREADS; t++) f3_phase3(t, RE.data(), IM.data(), [T2](int kbp, int c) { return T2[kbp * 16 + c]; }, out.data());
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
    const int conflicts = audit();
    printf("fft16k3: worst error/tolerance %.3f, %d bad bins, %d bank conflicts\n", worst, bad, conflicts);
    return (bad || conflicts) ? 1 : 0;
}
