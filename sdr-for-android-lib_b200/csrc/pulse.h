// pulse.h — batched beacon-pulse detectors: device state and launchers.
//   SpectralPulseDetector  src/dsp/spectral_pulse_detector.{h,cpp}   one detector per stream
//   AudioPulseDetector     src/ssb/audio_pulse_detector.{h,cpp}      one detector per (stream, channel)
// Both are sequential per-detector state machines over a <= 10 s energy history; the unit of
// parallelism is the detector (one thread each), frames / PCM samples are walked in order.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sdr_b200.h"

namespace sdrb {

constexpr int PULSE_MAX_ROIS = 128;   // the reference's deque is trimmed to 20 s of ROIs spaced >= ~0.5 s apart
constexpr int PULSE_FREQ_HIST = 30;   // kFreqHistoryMax, spectral_pulse_detector.h:62

// Resolved configuration of one detector family (host-computed, passed by value to the kernels).
struct PulseCfg {
    float fsEnergy, zDefaultS, tTargetInit, dtTolS, snrMin, snrRhythm, snrStrong, dispersionMax;
    int sumNMax;
    float liveWindowT, liveDivisor;
    int noiseRefFar, noiseRefNear;
    int maxBuf;          // int(10.f * fsEnergy): energy history length
    int ringCap;         // maxBuf + 1
    int audio;           // 0 spectral, 1 audio
    // audio front end (audio_pulse_detector.cpp:13-22): band = HP(fMin) + LP(fMax) at sampleRate,
    // low = LP(smoothCutoff) at fsEnergy; b0,b1,b2,a1,a2 designed on the host (tanf of glibc, as the reference)
    int frameSamples;
    float band[2][5], low[5];
};

// Carried state of one detector.  Scalars first, then the small deques as rings.
struct PulseState {
    int head, n;               // energy ring: index of eBuf_[0], eBuf_.size()
    float t0;                  // eBufT0_
    int roiHead, nRois;
    float tLastRoi;
    int lastScan;
    int fhHead, nFh;
    int locked;
    float tTarget;
    float last3[3];
    int nLast3;
    float histDt[5];
    int histN[5];
    int nHistDt, nHistN;
    int liveEtat, level;
    float lastSnr;
    int overflow;
    // audio front end
    float bz[2][2], lz[2];
    int frameCount;
    float frameAcc;
    float roiT[PULSE_MAX_ROIS];
    int roiEtat[PULSE_MAX_ROIS];
    float fhT[PULSE_FREQ_HIST], fhF[PULSE_FREQ_HIST];
};

void pulse_cfg_spectral(PulseCfg* c, uint32_t sample_rate, int samples_per_reading);
void pulse_cfg_audio(PulseCfg* c, const sdrb_audio_pulse_config* a);
void pulse_state_init(PulseState* s, const PulseCfg& c);

// One thread per stream: feeds results[s][b].best1kHzSnrSigma / best1kHzCenterFreqHz, b = 0..n_blocks-1.
// out record of (stream s, reading b) is out[(s * channels) * n_blocks + b].
cudaError_t launch_pulse_spectral(const sdrb_frame_result* d_results, PulseState* d_state, float* d_ring_e,
                                  float* d_ring_f, const PulseCfg& cfg, sdrb_pulse_result* d_out, int n_streams,
                                  int n_blocks, int channels, cudaStream_t st);
// One thread per unit (stream x channel): walks pcm[u][b][0..m) in order; out[u * n_blocks + b].
cudaError_t launch_pulse_audio(const int16_t* d_pcm, PulseState* d_state, float* d_ring_e, const PulseCfg& cfg,
                               sdrb_pulse_result* d_out, int n_units, int n_blocks, int m, cudaStream_t st);

}  // namespace sdrb
