// fft_tables.h — host-side twiddle table generation (double precision, rounded once).
#pragma once
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

}  // namespace sdrb
