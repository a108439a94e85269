// host_design.h — host-side parameter design for the GPU path: bin geometry of the estimator,
// RBJ filter coefficients, FIR taps and the block-to-block state map of the SSB chain.
//
// These are once-per-configuration scalars; they are evaluated on the host in the reference's
// float expressions (x86-64 baseline, no FMA contraction, glibc sinf/cosf/logf) so that every
// bin bound and every coefficient is bit-identical to what the reference computes:
//   geometry     src/dsp/fft_process.cpp:125-139,160,191-216,232,281-285,299
//   low-pass     src/ssb/ssb_demod_opt.cpp:60-73     HP/BP  :148-175
//   FIR taps     src/ssb/ssb_demod_opt.cpp:122-134   mode table :230-255
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>

#include "internal.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

namespace sdrb {

inline void design_geometry(uint32_t sampleRate, int n, int focusKhz, EstGeometry& g) {
    g = EstGeometry();
    g.n = n;
    const float freqPerBin = static_cast<float>(sampleRate) / static_cast<float>(n);
    const float X_hz = focusKhz * 1000.0f;
    const float nyquist = sampleRate / 2.0f;
    auto offToBin = [&](float off) -> int { return static_cast<int>((off + nyquist) / freqPerBin); };
    g.freqPerBin = freqPerBin;
    g.nyquist = nyquist;
    g.sampleRateF = static_cast<float>(sampleRate);
    g.focusLo = std::max(0, offToBin(-X_hz));
    g.focusHi = std::min(n - 1, offToBin(+X_hz) - 1);
    g.focusLen = g.focusHi - g.focusLo + 1;
    g.segOff[0] = 0;
    if (g.focusLen <= 0) { g.segOff[1] = 0; return; }
    g.winBins1k = std::max(1, static_cast<int>(std::ceil(1000.0f / freqPerBin)));
    g.segOff[1] = g.focusLen;
    for (int k = 1; k <= 5; k++) {
        const float nearX = (4 * k - 2) * X_hz;
        const float farX = 4 * k * X_hz;
        if (farX >= nyquist) break;
        const int lo[2] = {std::max(0, offToBin(+nearX)), std::max(0, offToBin(-farX))};
        const int hi[2] = {std::min(n - 1, offToBin(+farX) - 1), std::min(n - 1, offToBin(-nearX) - 1)};
        for (int side = 0; side < 2; side++) {
            if (hi[side] <= lo[side]) continue;
            g.winLo[g.nWin] = lo[side];
            g.winHi[g.nWin] = hi[side];
            g.segOff[g.nWin + 2] = g.segOff[g.nWin + 1] + (hi[side] - lo[side] + 1);
            g.nWin++;
        }
    }
    g.totalBins = g.segOff[g.nWin + 1];
    g.nBottom = std::max(1, static_cast<int>(g.nWin * 0.4f));
    const float logN = std::log(static_cast<float>(g.focusLen));
    g.sqrt2logN = std::sqrt(2.0f * logN);
    g.gumbelDen = std::sqrt(6.0f) * g.sqrt2logN;
    g.sqrtWin1k = std::sqrt(static_cast<float>(g.winBins1k));
}

inline BiquadCoef design_rf_lowpass(float fs, float fc, float Q) {
    float w0 = 2.0f * M_PI * fc / fs;
    float cosw0 = cosf(w0), sinw0 = sinf(w0);
    float alpha = sinw0 / (2.0f * Q);
    float norm = 1.0f / (1.0f + alpha);
    BiquadCoef c;
    c.a0 = (1.0f - cosw0) / 2.0f * norm;
    c.a1 = (1.0f - cosw0) * norm;
    c.a2 = c.a0;
    c.b1 = -2.0f * cosw0 * norm;
    c.b2 = (1.0f - alpha) * norm;
    return c;
}

inline BiquadCoef design_highpass(float fs, float f0, float Q) {
    float w0 = 2.0f * M_PI * f0 / fs;
    float cosw0 = cosf(w0), sinw0 = sinf(w0);
    float alpha = sinw0 / (2.0f * Q);
    float b0 = (1 + cosw0) / 2.0f, b1 = -(1 + cosw0), b2 = (1 + cosw0) / 2.0f;
    float a0 = 1 + alpha, a1 = -2 * cosw0, a2 = 1 - alpha;
    BiquadCoef c = {b0 / a0, b1 / a0, b2 / a0, a1 / a0, a2 / a0};
    return c;
}

inline BiquadCoef design_bandpass(float fs, float f0, float Q) {
    float w0 = 2.0f * M_PI * f0 / fs;
    float alpha = sinf(w0) / (2.0f * Q);
    float cosw0 = cosf(w0);
    float b0 = alpha, b1 = 0.0f, b2 = -alpha;
    float a0 = 1.0f + alpha, a1 = -2.0f * cosw0, a2 = 1.0f - alpha;
    BiquadCoef c = {b0 / a0, b1 / a0, b2 / a0, a1 / a0, a2 / a0};
    return c;
}

// Hann-windowed sinc of simpleFIRDecimate; returns the tap count (255, or n|1 for short blocks).
inline int design_fir(int n_in, int decim, float cutoffRel, float* h) {
    int N = 255;
    if (N > n_in) N = n_in | 1;
    int M = N - 1;
    float fc = cutoffRel / decim;
    for (int n = 0; n < N; n++) {
        int k = n - M / 2;
        float sinc = (k == 0) ? 2.0f * M_PI * fc : sinf(2.0f * M_PI * fc * k) / (float)k;
        float w = 0.5f - 0.5f * cosf(2.0f * M_PI * n / M);
        h[n] = (sinc / M_PI) * w;
    }
    float sum = 0.0f;
    for (int n = 0; n < N; n++) sum += h[n];
    if (sum != 0.0f) for (int n = 0; n < N; n++) h[n] /= sum;
    return N;
}

struct SsbModeParams { float agcTarget, agcFast, gain, lowpassBd, lowpassQ, transientCoeff; };
inline SsbModeParams ssb_mode(int mode) {
    // any other mode value leaves the file-static defaults in place, which equal mode 1 (:17-26)
    if (mode == 2) return {0.45f, 0.008f, 4.5f, 2200.0f, 1.2f, 0.7f};
    if (mode == 0) return {0.45f, 0.008f, 10.0f, 2200.0f, 1.2f, 0.7f};
    return {0.35f, 0.006f, 0.5f, 3200.0f, 0.9f, 0.55f};
}

// rfMode / rfFs = the mode and the sample rate of the FIRST block (the reference's static rfFilter is initialised
// once, with that call's sampleRate and parameters, :261-263).
// trueSideband (extension): the low-pass is designed at B/2 for the Weaver sideband filter.
inline void design_ssb(uint32_t fs, int n, int rfMode, uint32_t rfFs, int mode, int upper, bool trueSideband,
                       SsbParams& P, float* taps255) {
    const SsbModeParams mp = ssb_mode(mode), rp = ssb_mode(rfMode);
    P.n = n;
    P.decim = std::max(1, static_cast<int>(fs / 48000.0f));
    P.nTaps = n > 0 ? design_fir(n, P.decim, 0.45f, taps255) : 0;
    P.m = 0;
    if (P.nTaps > 0 && n >= P.nTaps) P.m = (n - P.nTaps) / P.decim + 1;
    P.upper = upper ? 1 : 0;
    P.dcAlpha = 0.9995f;
    P.agcTarget = mp.agcTarget;
    P.agcFast = mp.agcFast;
    P.agcSlow = 0.00035f;
    P.gain = mp.gain;
    P.transientCoeff = mp.transientCoeff;
    P.rf = design_rf_lowpass(static_cast<float>(rfFs ? rfFs : fs), trueSideband ? rp.lowpassBd * 0.5f : rp.lowpassBd, rp.lowpassQ);
    P.hp = design_highpass(48000.0f, 1200.0f, 0.7f);
    P.bp = design_bandpass(48000.0f, 2400.0f, 0.6f);
}

// [decim][smax] slot-major table for k_ssb_main: slot s' holds tap r + (s' - (smax - S)) * decim.
inline std::vector<float> slot_major_taps(const float* h, int nTaps, int decim, int smax) {
    const int S = nTaps > 0 ? (nTaps - 1) / decim + 1 : 1;
    std::vector<float> t((size_t)decim * smax, 0.0f);
    for (int r = 0; r < decim; r++)
        for (int sp = smax - S; sp < smax; sp++) {
            const int k = r + (sp - (smax - S)) * decim;
            if (k < nTaps) t[(size_t)r * smax + sp] = h[k];
        }
    return t;
}

}  // namespace sdrb
