// fft_dit.cuh — decimation-in-time register codelets with FMA-form butterflies (16- and 32-point forward DFT).
//
// Replaces fftwf_execute at src/dsp/fft_process.cpp:77-79 together with the kernels that call these codelets.
// In a decimation-in-time butterfly the twiddle multiply comes BEFORE the add / subtract, so it fuses:
//   w = m (1 + i t):  u = b.x - t b.y,  v = b.y + t b.x,  y0 = a + m (u + i v),  y1 = a - m (u + i v)   6 FMAs
//   w = m (t + i)  :  u = t b.x - b.y,  v = t b.y + b.x,  same tail (used where |Im w| > |Re w|)         6 FMAs
// against 4 (complex multiply) + 4 (add, subtract) for the decimation-in-frequency codelets of fft_core.cuh, whose
// multiply follows the subtract and cannot fuse; butterflies with w = 1 or -i are 4 adds.  32 points: 388 instructions
// instead of ~460; 16 points: 148 instead of ~170.  The bit reversal of the input is free: the values are in registers
// with static indices, the codelet simply works on v[bitrev(p)], and leaves X[k] in v[bitrev(k)]
// (dit_slot<LOG>(k)).  Twiddles are W_32^j as literals: (m, t) per the two forms above.
// __host__ __device__ so the CPU simulator (tests/cpp/fft_sim.cpp) runs the kernels' own code.
#pragma once
#include <math.h>

#include "fft_core.cuh"

namespace sdrb {

template <int LOG>
SDRB_HD constexpr int dit_slot(int k) {
    int r = 0;
    for (int b = 0; b < LOG; b++) r |= ((k >> b) & 1) << (LOG - 1 - b);
    return r;
}

// one butterfly (a, b) <- (a + w b, a - w b), w = W_32^J, J in [0, 16)
template <int J>
SDRB_HD void dit_bf(cpx& a, cpx& b) {
    constexpr int kind[16] = {0, 2, 2, 2, 2, 3, 3, 3, 1, 3, 3, 3, 3, 2, 2, 2};
    constexpr float M[16] = {1.000000000e+00f, 9.807852507e-01f, 9.238795042e-01f, 8.314695954e-01f, 7.071067691e-01f, -8.314695954e-01f, -9.238795042e-01f, -9.807852507e-01f, 1.000000000e+00f, -9.807852507e-01f, -9.238795042e-01f, -8.314695954e-01f, -7.071067691e-01f, -8.314695954e-01f, -9.238795042e-01f, -9.807852507e-01f};
    constexpr float T[16] = {0.000000000e+00f, -1.989123672e-01f, -4.142135680e-01f, -6.681786180e-01f, -1.000000000e+00f, -6.681786180e-01f, -4.142135680e-01f, -1.989123672e-01f, 0.000000000e+00f, 1.989123672e-01f, 4.142135680e-01f, 6.681786180e-01f, 1.000000000e+00f, 6.681786180e-01f, 4.142135680e-01f, 1.989123672e-01f};
    if (kind[J] == 0) {
        const cpx s = cadd(a, b), d = csub(a, b);
        a = s; b = d;
    } else if (kind[J] == 1) {   // w = -i: w b = (b.y, -b.x)
        const cpx s = mk(a.x + b.y, a.y - b.x), d = mk(a.x - b.y, a.y + b.x);
        a = s; b = d;
    } else {
        float u, v;
        if (kind[J] == 2) { u = fmaf(-T[J], b.y, b.x); v = fmaf(T[J], b.x, b.y); }
        else { u = fmaf(T[J], b.x, -b.y); v = fmaf(T[J], b.y, b.x); }
        const cpx s = mk(fmaf(M[J], u, a.x), fmaf(M[J], v, a.y));
        const cpx d = mk(fmaf(-M[J], u, a.x), fmaf(-M[J], v, a.y));
        a = s; b = d;
    }
}

template <int LOG, int SPAN, int BASE, int J>
struct DitStage {
    SDRB_HD static void run(cpx* v) {
        constexpr int N = 1 << LOG;
        dit_bf<J * (16 / SPAN)>(v[dit_slot<LOG>(BASE + J)], v[dit_slot<LOG>(BASE + J + SPAN)]);
        if constexpr (J + 1 < SPAN) DitStage<LOG, SPAN, BASE, J + 1>::run(v);
        else if constexpr (BASE + 2 * SPAN < N) DitStage<LOG, SPAN, BASE + 2 * SPAN, 0>::run(v);
        else if constexpr (2 * SPAN < N) DitStage<LOG, 2 * SPAN, 0, 0>::run(v);
    }
};

// N = 2^LOG point forward DFT (LOG = 4 or 5), natural-order input in v[n], X[k] left in v[dit_slot<LOG>(k)].
// (Decimation in time works on the bit-reversed sequence: position p of it is v[bitrev(p)] = v[dit_slot(p)].)
template <int LOG>
SDRB_HD void fft_dit(cpx* v) { DitStage<LOG, 1, 0, 0>::run(v); }

}  // namespace sdrb
