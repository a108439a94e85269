// synth.cuh — deterministic synthetic IQ source, integer-only, bit-identical on host and device.
//
// Replaces Soapy/USB capture (src/sdr-bridge-java-soapy.cpp:532-539 readStream) on a box
// without a dongle.  Per stream s (global id), seed = 0x5D120000 + s:
//   * approximately Gaussian noise: sum of the four bytes of a counter hash of (seed, i)
//   * tone 1: on/off keyed CW inside the default +-5 kHz focus band, offset in [-4, +4] kHz,
//     30 % full scale, period 1.75 s / 10 % duty (the detectors' tTargetInit,
//     src/dsp/spectral_pulse_detector.h:25); streams with s % 4 == 0 start keyed on
//   * tone 2: static carrier at +-(12.5 .. 16.5) kHz, 5 % full scale, lands inside the first
//     reference window of evaluateSignalStrength (src/dsp/fft_process.cpp:191-216)
//   * CU8 : offset binary around 127 with a 0.4 LSB mean dither so that the 127.4 offset of
//     convertIQ (src/ssb/ssb_demod_opt.cpp:37) leaves no DC spike
//   * CS16: same construction at 16-bit amplitude;  CF32 = CS16 * 2^-15 (exact)
// Everything is integer arithmetic on a 4096-entry Q15 sine table built once on the host.
#pragma once
#include <stdint.h>

#include "fft_core.cuh"

namespace sdrb {

constexpr uint32_t SYNTH_SEED0 = 0x5D120000u;
constexpr int SYNTH_SINE_N = 4096;

struct SynthStream {
    uint32_t seed;
    uint32_t step1, step2;    // 32-bit phase increments (two's complement for negative offsets)
    uint32_t keyPeriod;       // samples per keying period
    uint32_t keyOn;           // samples the tone is on at the start of each period
    uint32_t keyOffset;       // phase of the keying pattern at sample 0
    int32_t toneHz1, toneHz2;
};

SDRB_HD uint32_t synth_mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du;
    x ^= x >> 15; x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}

SDRB_HD SynthStream synth_stream(int stream_id, uint32_t fs) {
    SynthStream p;
    p.seed = SYNTH_SEED0 + (uint32_t)stream_id;
    const uint32_t h1 = synth_mix32(p.seed ^ 0xA5A5A5A5u);
    const uint32_t h2 = synth_mix32(h1 + 0x9E3779B9u);
    const uint32_t h3 = synth_mix32(h2 + 0x9E3779B9u);
    p.toneHz1 = -4000 + (int32_t)(h1 % 8001u);
    p.toneHz2 = (12500 + 1000 * (int32_t)(h2 % 5u)) * (((h2 >> 16) & 1u) ? -1 : 1);
    p.step1 = (uint32_t)(int32_t)(((long long)p.toneHz1 * 4294967296LL) / (long long)fs);
    p.step2 = (uint32_t)(int32_t)(((long long)p.toneHz2 * 4294967296LL) / (long long)fs);
    const unsigned long long per = ((unsigned long long)fs * 7ull) / 4ull;
    p.keyPeriod = (uint32_t)(per ? per : 1ull);
    p.keyOn = p.keyPeriod / 10u;
    p.keyOffset = (stream_id % 4 == 0) ? 0u : (h3 % p.keyPeriod);
    return p;
}

// Q15 lookup: cos / sin of a 32-bit phase
SDRB_HD int synth_cos(const int16_t* tab, uint32_t ph) { return tab[((ph >> 20) + 1024u) & 4095u]; }
SDRB_HD int synth_sin(const int16_t* tab, uint32_t ph) { return tab[(ph >> 20) & 4095u]; }

// One complex sample as 16-bit-scale integers (cs16) or offset-binary bytes (cu8).
// i = absolute sample index of the stream, key_pos = (i + keyOffset) % keyPeriod.
template <int FMT>
SDRB_HD void synth_sample(const SynthStream& p, const int16_t* tab, long long i, uint32_t key_pos, int& I, int& Q) {
    const uint32_t lo = (uint32_t)i, hi = (uint32_t)((unsigned long long)i >> 32);
    const uint32_t hI = synth_mix32(lo * 0x9E3779B1u ^ p.seed * 0x85EBCA77u ^ hi * 0xC2B2AE3Du);
    const uint32_t hQ = synth_mix32(hI ^ 0x68E31DA4u);
    const int nI = (int)((hI & 255u) + ((hI >> 8) & 255u) + ((hI >> 16) & 255u) + (hI >> 24)) - 510;
    const int nQ = (int)((hQ & 255u) + ((hQ >> 8) & 255u) + ((hQ >> 16) & 255u) + (hQ >> 24)) - 510;
    const uint32_t ph1 = lo * p.step1 + hi * 0u;  // phase wraps mod 2^32; only the low word matters
    const uint32_t ph2 = lo * p.step2;
    const bool on = key_pos < p.keyOn;
    if (FMT == 0) {
        const uint32_t hD = synth_mix32(hQ + 0x9E3779B9u);
        int vi = 127 + (((hD & 255u) < 102u) ? 1 : 0) + ((nI * 7 + 128) >> 8) + ((6 * synth_cos(tab, ph2) + 16384) >> 15);
        int vq = 127 + ((((hD >> 8) & 255u) < 102u) ? 1 : 0) + ((nQ * 7 + 128) >> 8) + ((6 * synth_sin(tab, ph2) + 16384) >> 15);
        if (on) {
            vi += (38 * synth_cos(tab, ph1) + 16384) >> 15;
            vq += (38 * synth_sin(tab, ph1) + 16384) >> 15;
        }
        I = vi < 0 ? 0 : (vi > 255 ? 255 : vi);
        Q = vq < 0 ? 0 : (vq > 255 ? 255 : vq);
    } else {
        int vi = nI * 2 + ((1638 * synth_cos(tab, ph2) + 16384) >> 15);
        int vq = nQ * 2 + ((1638 * synth_sin(tab, ph2) + 16384) >> 15);
        if (on) {
            vi += (9830 * synth_cos(tab, ph1) + 16384) >> 15;
            vq += (9830 * synth_sin(tab, ph1) + 16384) >> 15;
        }
        I = vi < -32767 ? -32767 : (vi > 32767 ? 32767 : vi);
        Q = vq < -32767 ? -32767 : (vq > 32767 ? 32767 : vq);
    }
}

}  // namespace sdrb
