/* oracle/sdr_oracle_ext.c — CPU definition of the EXTENSION stages of the hot path.
 * TEST INFRASTRUCTURE ONLY (see sdr_oracle.h).
 *
 * PARITY UNPINNED BY CONSTRUCTION: the reference has no code for these stages (SURVEY.md 8c
 * "Extension stages with no reference code at all": FFT window, FFT-path DC removal, NCO mix,
 * multi-channel SSB, true sideband selection).  This file IS their definition; the CUDA path
 * is checked against it.  Every knob defaults to identity, and with all knobs at identity the
 * functions below reduce to the pinned restatement in sdr_oracle.c (tested).
 *
 * Arithmetic is written with explicit single operations (and explicit fmaf where a fused
 * multiply-add is meant) so that the CUDA kernels can reproduce it bit for bit. */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "sdr_oracle.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ------------------------------------------------------------------ FFT front end */

/* Symmetric Hann window, the same family as the reference's FIR window
 * (ssb_demod_opt.cpp:131, 0.5 - 0.5 cos(2 pi n / M), M = N - 1), evaluated in double and
 * rounded once.  kind 0 = rectangular (reference behaviour). */
void sdr_oracle_window(int kind, int n, float* w) {
    for (int i = 0; i < n; i++) {
        if (kind == 1 && n > 1) w[i] = (float)(0.5 - 0.5 * cos(2.0 * M_PI * (double)i / (double)(n - 1)));
        else w[i] = 1.0f;
    }
}

/* x'[i] = (x[i] - mean) * w[i]; mean = block mean in double rounded to float (remove_dc != 0). */
void sdr_oracle_fft_front(const float* iq, int n, int window_kind, int remove_dc, float* out) {
    float mr = 0.0f, mi = 0.0f;
    if (remove_dc) {
        double sr = 0.0, si = 0.0;
        for (int i = 0; i < n; i++) { sr += iq[2 * i]; si += iq[2 * i + 1]; }
        mr = (float)(sr / n);
        mi = (float)(si / n);
    }
    float* w = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
    sdr_oracle_window(window_kind, n, w);
    for (int i = 0; i < n; i++) {
        out[2 * i] = (iq[2 * i] - mr) * w[i];
        out[2 * i + 1] = (iq[2 * i + 1] - mi) * w[i];
    }
    free(w);
}

/* ------------------------------------------------------------------ NCO */

/* cos/sin of a 32-bit phase (turns * 2^32): quadrant reduction to [-pi/4, pi/4], then the
 * Cephes single-precision minimax polynomials evaluated with fused multiply-adds. */
void sdr_oracle_nco(uint32_t phase, float* c_out, float* s_out) {
    const uint32_t q = (phase + 0x20000000u) >> 30;
    const int32_t r = (int32_t)(phase - (q << 30));
    const float x = (float)r * 1.4629180792671596e-9f; /* 2 pi / 2^32 */
    const float z = x * x;
    float sp = fmaf(z, -1.9515295891e-4f, 8.3321608736e-3f);
    sp = fmaf(z, sp, -1.6666654611e-1f);
    const float s = fmaf(x * z, sp, x);
    float cp = fmaf(z, 2.443315711809948e-5f, -1.388731625493765e-3f);
    cp = fmaf(z, cp, 4.166664568298827e-2f);
    cp = fmaf(z, cp, -0.5f);
    const float c = fmaf(z, cp, 1.0f);
    switch (q & 3u) {
        case 0: *c_out = c; *s_out = s; break;
        case 1: *c_out = -s; *s_out = c; break;
        case 2: *c_out = -c; *s_out = -s; break;
        default: *c_out = s; *s_out = -c; break;
    }
}

/* Phase increment of f_hz at sample rate fs, in turns * 2^32 (two's complement for f < 0). */
uint32_t sdr_oracle_nco_inc(double f_hz, uint32_t fs) {
    return (uint32_t)(int64_t)llround(f_hz / (double)fs * 4294967296.0);
}

/* ------------------------------------------------------------------ extended SSB chain */

void sdr_oracle_ssbx_init(sdr_oracle_ssbx_state* s) {
    memset(s, 0, sizeof(*s));
    sdr_oracle_ssb_init(&s->base);
}

/* One channel of the extended chain on one block.  Order of operations:
 *   removeDC on the raw complex IQ (reference :49-55, reset per block)
 *   NCO 1: multiply by exp(-j theta1), theta1 = inc1 * k, k = samples since reset
 *   IIR2 low-pass on I (reference :75-84) and, with true_sideband, on Q with its own state
 *   true_sideband: Weaver second mix, a = 2 (yI c2 -/+ yQ s2)   (USB: -, LSB: +)
 *   else          : the reference's demodSSB on {y, y} (:89-96): USB = y + y, LSB = y - y
 *   adaptiveAGC, FIR decimate, HP, BP, transientBoost, floatToPCM exactly as sdr_oracle.c
 * With true_sideband the mixers are  inc1 = (offset +/- B/2) / fs,  inc2 = (B/2) / fs  and the
 * low-pass is designed at B/2 (B = the mode's lowpassBd), else inc1 = offset / fs. */
int sdr_oracle_process_ssbx(sdr_oracle_ssbx_state* st, const float* iq_in, int n, uint32_t fs, int upper,
                            int mode, int true_sideband, double offset_hz, int16_t* pcm_out, int cap) {
    sdr_oracle_ssb_state* s = &st->base;
    if (mode == 2) {
        s->agcTarget = 0.45f; s->agcFast = 0.008f; s->gain = 4.5f;
        s->lowpassBd = 2200.0f; s->lowpassQ = 1.2f; s->transientCoeff = 0.7f;
    } else if (mode == 0) {
        s->agcTarget = 0.45f; s->agcFast = 0.008f; s->gain = 10.0f;
        s->lowpassBd = 2200.0f; s->lowpassQ = 1.2f; s->transientCoeff = 0.7f;
    } else if (mode == 1) {
        s->agcTarget = 0.35f; s->agcFast = 0.006f; s->agcSlow = 0.00035f; s->gain = 0.5f;
        s->lowpassBd = 3200.0f; s->lowpassQ = 0.9f; s->transientCoeff = 0.55f;
    }
    if (!s->initialised) {
        const float cut = true_sideband ? s->lowpassBd * 0.5f : s->lowpassBd;
        sdr_oracle_iir2_init_lowpass(&s->rf, (float)fs, cut, s->lowpassQ);
        st->rfq = s->rf;
        sdr_oracle_biquad_init_highpass(&s->hp, 48000.0f, 1200.0f, 0.7f);
        sdr_oracle_biquad_init_bandpass(&s->bp, 48000.0f, 2400.0f, 0.6f);
        st->halfBand = (double)s->lowpassBd * 0.5;
        s->initialised = 1;
        s->sampCount = n;
    }
    const double sgn = upper ? 1.0 : -1.0;
    const uint32_t inc1 = sdr_oracle_nco_inc(true_sideband ? offset_hz + sgn * st->halfBand : offset_hz, fs);
    const uint32_t inc2 = true_sideband ? sdr_oracle_nco_inc(st->halfBand, fs) : 0u;
    float* audio = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
    const float alpha = 0.9995f, oma = 1.0f - alpha;
    float dcr = 0.0f, dci = 0.0f;
    float iz1 = s->rf.z1, iz2 = s->rf.z2, qz1 = st->rfq.z1, qz2 = st->rfq.z2;
    const sdr_oracle_biquad* f = &s->rf;
    for (int i = 0; i < n; i++) {
        const uint64_t k = st->samples + (uint64_t)i;
        dcr = alpha * dcr + oma * iq_in[2 * i];
        dci = alpha * dci + oma * iq_in[2 * i + 1];
        const float xr = iq_in[2 * i] - dcr, xi = iq_in[2 * i + 1] - dci;
        float c1, s1;
        sdr_oracle_nco((uint32_t)(inc1 * k), &c1, &s1);
        const float mr = xr * c1 + xi * s1;
        const float mi = xi * c1 - xr * s1;
        float yi = f->a0 * mr + f->a1 * iz1 + f->a2 * iz2 - f->b1 * iz1 - f->b2 * iz2;
        iz2 = iz1; iz1 = yi;
        float a;
        if (true_sideband) {
            float yq = f->a0 * mi + f->a1 * qz1 + f->a2 * qz2 - f->b1 * qz1 - f->b2 * qz2;
            qz2 = qz1; qz1 = yq;
            float c2, s2;
            sdr_oracle_nco((uint32_t)(inc2 * k), &c2, &s2);
            const float p = yi * c2, q = yq * s2;
            const float w = upper ? p - q : p + q;
            a = w + w;
        } else {
            a = upper ? yi + yi : yi - yi;
        }
        audio[i] = a;
    }
    s->rf.z1 = iz1; s->rf.z2 = iz2; st->rfq.z1 = qz1; st->rfq.z2 = qz2;
    st->samples += (uint64_t)n;
    sdr_oracle_adaptive_agc(audio, n, s->agcTarget, s->agcFast, 0.00035f);
    int decim = (int)(fs / 48000.0f); if (decim < 1) decim = 1;
    int mcap = n / decim + 4;
    float* a48 = (float*)malloc(sizeof(float) * (size_t)mcap);
    int m = sdr_oracle_fir_decimate(audio, n, decim, 0.45f, a48, mcap);
    if (m > 0) {
        sdr_oracle_biquad_process(&s->hp, a48, m);
        sdr_oracle_biquad_process(&s->bp, a48, m);
        sdr_oracle_transient_boost(a48, m, s->transientCoeff);
    }
    int16_t* pcm = (int16_t*)malloc(sizeof(int16_t) * (size_t)(m > 0 ? m : 1));
    sdr_oracle_float_to_pcm(a48, m, s->gain, pcm);
    memcpy(pcm_out, pcm, sizeof(int16_t) * (size_t)(m < cap ? m : cap));
    free(pcm); free(a48); free(audio);
    return m;
}
