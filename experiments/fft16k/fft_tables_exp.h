// fft_tables_exp.h — twiddle tables of the experimental 16K kernels (fft16k3/4/5), split from csrc/fft_tables.h
// when the experiments left the product library.
#pragma once
#include "fft_tables.h"
#include "fft16k3.cuh"
#include "fft16k4.cuh"

namespace sdrb {

// Layout (bytes): TP as cpx[5][512] | T2 as pw4[16][16], see fft16k3.cuh
inline std::vector<unsigned char> make_f3_tables() {
    std::vector<unsigned char> raw(F3_TABLE_BYTES);
    cpx* TP = reinterpret_cast<cpx*>(raw.data());
    pw4* T2 = reinterpret_cast<pw4*>(raw.data() + F3_TP * 8);
    for (int j = 0; j < 5; j++)
        for (int r = 0; r < 512; r++) TP[j * 512 + r] = twiddle((long long)r << j, F3_N);
    for (int k = 0; k < 16; k++)
        for (int c = 0; c < 16; c++) {
            const cpx lo = twiddle((long long)c * k, 512), hi = twiddle((long long)c * (k + 16), 512);
            T2[k * 16 + c].re = mkp(lo.x, hi.x);
            T2[k * 16 + c].im = mkp(lo.y, hi.y);
        }
    return raw;
}

// Layout: TQ as cpx[6][256] | T3 as cpx[16][16], see fft16k4.cuh
inline std::vector<cpx> make_f4_tables() {
    std::vector<cpx> t(F4_TQ + F4_T3);
    for (int j = 0; j < 5; j++)
        for (int r = 0; r < 256; r++) t[j * 256 + r] = twiddle((long long)r << j, F4_H);
    for (int r = 0; r < 256; r++) t[5 * 256 + r] = twiddle(r, F4_N);
    for (int kb = 0; kb < 16; kb++)
        for (int c = 0; c < 16; c++) t[F4_TQ + kb * 16 + c] = twiddle((long long)c * kb, 256);
    return t;
}

}  // namespace sdrb
