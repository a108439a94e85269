// tests/cpp/fft_sim.cpp — CPU execution of the FFT kernels' phase functions, thread by
// thread, against a double-precision DFT.  Validates index algebra, twiddle tables and the
// shared-memory layout (including a bank-conflict audit) without a GPU.
// Build: g++ -O2 -std=c++17 -I<csrc> fft_sim.cpp -o fft_sim ; exit code 0 = pass.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "fft_tables.h"

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

int main() {
    const int N = F2_N;

    // u8 input: tone + pseudo noise
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

    // ---------------------------------------------------------------- second-generation kernel
    int bad2 = 0, conflicts2 = 0;
    {
        std::vector<cpx> tw2 = make_f2_tables();
        const cpx* TP = tw2.data();
        const cpx* T2b = TP + F2_TP;
        std::vector<cpx> E2(F2_E_ELEMS, mk(0, 0));
        std::vector<float> out2(N, -1.0f);
        for (int t = 0; t < F2_THREADS; t++) {
            cpx v[32];
            f2_phase1_load<FMT_CU8>(t, raw.data(), v);   // the kernel's sequence: load, DFT32, twiddle + store
            fft_dit<5>(v);
            f2_phase1_twiddle_store(t, TP, v, E2.data());
        }
        for (int t = 0; t < F2_THREADS; t++) f2_phase2(t, E2.data());
        auto tw = [T2b](int kb, int c) { return T2b[kb * 16 + c]; };
        for (int t = 0; t < F2_THREADS; t++) f2_phase3(t, E2.data(), tw, out2.data());
        double maxrel2 = 0;
        for (int k = 0; k < N; k++) {
            double p = Xr[k] * Xr[k] + Xi[k] * Xi[k];
            double g = out2[(k + N / 2) % N];
            double err = fabs(g - p);
            double tol = 1e-5 * p + 2.0 * sqrt(p) * 3e-7 * sqrt(total) + 1e-9 * total / N;
            if (err > tol) { if (bad2 < 10) printf("v2 bin %d ref %.6g got %.6g\n", k, p, g); bad2++; }
            if (p > 1e-3 * total / N) maxrel2 = fmax(maxrel2, err / p);
        }
        printf("fft16k2 sim: bad=%d maxrel=%.3g\n", bad2, maxrel2);
        // bank audit: 8-byte accesses, each half-warp must cover 16 distinct 8-byte bank pairs
        auto audit2 = [&](auto addr_of, int iters, const char* name) {
            for (int w = 0; w < F2_THREADS / 32; w++)
                for (int j = 0; j < iters; j++)
                    for (int h = 0; h < 2; h++) {
                        int seen[16] = {0};
                        for (int l = 0; l < 16; l++) {
                            int t = w * 32 + h * 16 + l;
                            int bank = addr_of(t, j) & 15;
                            if (seen[bank]++) { conflicts2++; if (conflicts2 < 5) printf("conflict %s w%d j%d\n", name, w, j); }
                        }
                    }
        };
        audit2([](int t, int k) { return k * F2_PITCH + t; }, 32, "v2 p1 store");
        audit2([](int t, int b) { return (t >> 4) * F2_PITCH + b * 16 + (t & 15); }, 32, "v2 p2 ld/st");
        audit2([](int t, int c) { return (t & 31) * F2_PITCH + (t >> 5) * 16 + c; }, 16, "v2 p3 ld h0");
        audit2([](int t, int c) { return ((t + 512) & 31) * F2_PITCH + ((t + 512) >> 5) * 16 + c; }, 16, "v2 p3 ld h1");
        printf("v2 bank conflicts: %d\n", conflicts2);
    }
    return (bad2 || conflicts2) ? 1 : 0;
}
