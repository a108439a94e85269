// tests/cpp/fft_sim4.cpp — CPU execution of the fourth-generation 16K kernel's phase functions
// (two 8192-point jobs per transform, fft16k4.cuh), thread by thread, against a double-precision
// DFT, plus the shared-memory bank audit of its exchange layout.
// Build: g++ -O2 -std=c++17 -I<csrc> fft_sim4.cpp -o fft_sim4 ; exit code 0 = pass.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "fft_tables_exp.h"

using namespace sdrb;

static void dft_ref(const std::vector<double>& re, const std::vector<double>& im, std::vector<double>& Xr,
                    std::vector<double>& Xi) {
    int n = (int)re.size(), bits = 0;
    while ((1 << bits) < n) bits++;
    Xr.assign(n, 0); Xi.assign(n, 0);
    for (int i = 0; i < n; i++) {
        unsigned r = 0;
        for (int b = 0; b < bits; b++) if (i & (1 << b)) r |= 1u << (bits - 1 - b);
        Xr[r] = re[i]; Xi[r] = im[i];
    }
    for (int len = 2; len <= n; len <<= 1) {
        int half = len / 2;
        for (int base = 0; base < n; base += len)
            for (int j = 0; j < half; j++) {
                double a = -2.0 * M_PI * j / len, c = cos(a), s = sin(a);
                double xr = Xr[base + j + half], xi = Xi[base + j + half];
                double tr = xr * c - xi * s, ti = xr * s + xi * c;
                Xr[base + j + half] = Xr[base + j] - tr; Xi[base + j + half] = Xi[base + j] - ti;
                Xr[base + j] += tr; Xi[base + j] += ti;
            }
    }
}

struct HostTw3 {
    const cpx* T3;
    cpx operator()(int kb, int c) const { return T3[kb * 16 + c]; }
};

template <int P>
static void run_job(const uint16_t* raw, const cpx* TQ, const cpx* T3, std::vector<cpx>& E, float* out) {
    for (int t = 0; t < F4_THREADS; t++) {
        cpx v[32], tq[6];
        for (int j = 0; j < 6; j++) tq[j] = TQ[j * 256 + t];
        f4_phase1_load<FMT_CU8, P>(t, raw, v);
        f4_phase1_compute<P>(tq, v);
        f4_phase1_store(t, E.data(), v);
    }
    for (int t = 0; t < F4_THREADS; t++) f4_phase2(t, E.data());
    HostTw3 tw{T3};
    for (int t = 0; t < F4_THREADS; t++) f4_phase3<P>(t, E.data(), tw, out);
}

int main() {
    const int N = F4_N;
    std::vector<uint16_t> raw(N);
    std::vector<double> re(N), im(N);
    unsigned lcg = 12345;
    for (int i = 0; i < N; i++) {
        lcg = lcg * 1664525u + 1013904223u;
        int nI = (int)((lcg >> 24) & 7) - 3, nQ = (int)((lcg >> 16) & 7) - 3;
        int I = 127 + (int)lrint(38 * cos(2 * M_PI * 1000.25 * i / N)) + nI;
        int Q = 127 + (int)lrint(38 * sin(2 * M_PI * 1000.25 * i / N)) + nQ;
        raw[i] = (uint16_t)(I | (Q << 8));
        re[i] = ((double)(float)(I - 127.4f)) / 128.0;
        im[i] = ((double)(float)(Q - 127.4f)) / 128.0;
    }
    std::vector<double> Xr, Xi;
    dft_ref(re, im, Xr, Xi);
    double total = 0;
    for (int k = 0; k < N; k++) total += Xr[k] * Xr[k] + Xi[k] * Xi[k];

    std::vector<cpx> tw = make_f4_tables();
    const cpx* TQ = tw.data();
    const cpx* T3 = TQ + F4_TQ;
    std::vector<cpx> E(F4_E_ELEMS, mk(0, 0));
    std::vector<float> out(N, -1.0f);
    run_job<0>(raw.data(), TQ, T3, E, out.data());
    run_job<1>(raw.data(), TQ, T3, E, out.data());

    int bad = 0;
    double maxrel = 0;
    for (int k = 0; k < N; k++) {
        double p = Xr[k] * Xr[k] + Xi[k] * Xi[k];
        double g = out[(k + N / 2) % N];
        double err = fabs(g - p);
        double tol = 1e-5 * p + 2.0 * sqrt(p) * 3e-7 * sqrt(total) + 1e-9 * total / N;
        if (err > tol) { if (bad < 10) printf("bin %d ref %.6g got %.6g\n", k, p, g); bad++; }
        if (p > 1e-3 * total / N) maxrel = fmax(maxrel, err / p);
    }
    // bank audit: 8-byte accesses, each half-warp must cover 16 distinct 8-byte bank pairs
    int conflicts = 0;
    auto audit = [&](auto addr_of, int iters, const char* name) {
        for (int w = 0; w < F4_THREADS / 32; w++)
            for (int j = 0; j < iters; j++)
                for (int h = 0; h < 2; h++) {
                    int seen[16] = {0};
                    for (int l = 0; l < 16; l++) {
                        int t = w * 32 + h * 16 + l;
                        int bank = addr_of(t, j) & 15;
                        if (seen[bank]++) { conflicts++; if (conflicts < 5) printf("conflict %s w%d j%d\n", name, w, j); }
                    }
                }
    };
    audit([](int t, int k) { return k * F4_PITCH + t; }, 32, "p1 store");
    for (int hh = 0; hh < 2; hh++) {
        audit([hh](int t, int b) { int q = t + 256 * hh; return (q >> 4) * F4_PITCH + b * 16 + (q & 15); }, 16, "p2 ld/st");
        audit([hh](int t, int c) { int q = t + 256 * hh; return (q & 31) * F4_PITCH + (q >> 5) * 16 + c; }, 16, "p3 ld");
    }
    printf("fft16k4 sim: %d bad bins, %d bank conflicts, maxrel=%.3g\n", bad, conflicts, maxrel);
    return (bad || conflicts) ? 1 : 0;
}
