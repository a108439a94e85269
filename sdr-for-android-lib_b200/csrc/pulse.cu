// pulse.cu — batched SpectralPulseDetector / AudioPulseDetector, sm_100a.
// Reference: src/dsp/spectral_pulse_detector.cpp, src/ssb/audio_pulse_detector.cpp.
//
// One thread per detector walks its frames (spectral) or PCM samples (audio) in order; the reference's
// deques are fixed-capacity rings in global memory.  Every float expression keeps the reference's
// evaluation order; this file is compiled with -fmad=false so no multiply-add is contracted, which makes
// the detectors bit-identical to the reference built with g++ -O2 on x86-64 (no FMA at baseline).
//
// What is whose: the batching (one struct serving both detectors, deques as rings with an overflow counter,
// biquad state in registers, host-side filter design) is this repository's; `Det::detect_rois` and `Det::est_freq`
// are statement-for-statement transliterations of the reference's detectRois (spectral_pulse_detector.cpp:47-146,
// same body in audio_pulse_detector.cpp:145-268) and estimatedFreqHz (:148-196) -- a sequential state machine whose
// every branch and float expression has to come out bit-identical leaves no freedom in its control flow.
#include "pulse.h"

#include <cmath>
#include <cstring>

namespace sdrb {

// ---------------------------------------------------------------- host: configuration
void pulse_cfg_spectral(PulseCfg* c, uint32_t sample_rate, int samples_per_reading) {
    // SpectralPulseDetector::Config defaults (spectral_pulse_detector.h:22-34) with the frame rate the
    // bridge configures (src/sdr-bridge-java-soapy.cpp:1131-1137)
    memset(c, 0, sizeof *c);
    c->fsEnergy = samples_per_reading > 0 ? (float)sample_rate / (float)samples_per_reading : 20.f;
    c->zDefaultS = 0.666f; c->tTargetInit = 1.75f; c->dtTolS = 0.150f;
    c->snrMin = 1.5f; c->snrRhythm = 2.5f; c->snrStrong = 4.0f; c->dispersionMax = 1.3f;
    c->sumNMax = 7; c->liveWindowT = 4.0f; c->liveDivisor = 3.0f;
    c->maxBuf = (int)(10.f * c->fsEnergy);
    c->ringCap = c->maxBuf + 1;
    c->audio = 0;
}

namespace {
// makeLP2 / makeHP2 (audio_pulse_detector.cpp:27-51): b0, b1, b2, a1, a2
void make_lp2(float fs, float fc, float q, float* o) {
    const float kPi = 3.14159265358979f;
    const float K = std::tan(kPi * fc / fs), K2 = K * K, norm = K2 + K / q + 1.f;
    o[0] = K2 / norm; o[1] = 2.f * K2 / norm; o[2] = K2 / norm;
    o[3] = 2.f * (K2 - 1.f) / norm; o[4] = (K2 - K / q + 1.f) / norm;
}
void make_hp2(float fs, float fc, float q, float* o) {
    const float kPi = 3.14159265358979f;
    const float K = std::tan(kPi * fc / fs), K2 = K * K, norm = K2 + K / q + 1.f;
    o[0] = 1.f / norm; o[1] = -2.f / norm; o[2] = 1.f / norm;
    o[3] = 2.f * (K2 - 1.f) / norm; o[4] = (K2 - K / q + 1.f) / norm;
}
}  // namespace

void pulse_cfg_audio(PulseCfg* c, const sdrb_audio_pulse_config* a) {
    memset(c, 0, sizeof *c);
    c->fsEnergy = a->fsEnergy; c->zDefaultS = a->zDefaultS; c->tTargetInit = a->tTargetInit; c->dtTolS = a->dtTolS;
    c->snrMin = a->snrMin; c->snrRhythm = a->snrRhythm; c->snrStrong = a->snrStrong;
    c->dispersionMax = a->dispersionMax; c->sumNMax = a->sumNMax;
    c->liveWindowT = a->liveWindowT; c->liveDivisor = a->liveDivisor;
    c->noiseRefFar = a->noiseRefFar; c->noiseRefNear = a->noiseRefNear;
    c->maxBuf = (int)(10.f * c->fsEnergy);
    c->ringCap = c->maxBuf + 1;
    c->audio = 1;
    // audio_pulse_detector.cpp:13-22
    make_hp2(a->sampleRate, a->fMin, 0.7071f, c->band[0]);
    make_lp2(a->sampleRate, a->fMax, 0.7071f, c->band[1]);
    make_lp2(a->fsEnergy, a->smoothCutoff, 0.7071f, c->low);
    const int fs = (int)(a->sampleRate / a->fsEnergy);
    c->frameSamples = fs > 1 ? fs : 1;
}

void pulse_state_init(PulseState* s, const PulseCfg& c) {
    // reset(): spectral_pulse_detector.cpp:200-216, audio_pulse_detector.cpp:281-297
    memset(s, 0, sizeof *s);
    s->tLastRoi = -1.f;
    s->tTarget = c.tTargetInit;
}

// ---------------------------------------------------------------- device: the shared state machine
namespace {

struct Det {
    PulseState* s;
    float* e;       // energy ring [ringCap]
    float* fq;      // frequency ring (spectral) or nullptr
    const PulseCfg& c;

    __device__ __forceinline__ float ev(int i) const {   // eBuf_[i]
        int k = s->head + i;
        if (k >= c.ringCap) k -= c.ringCap;
        return e[k];
    }
    // timeOfIdx: spectral_pulse_detector.cpp:17-19
    __device__ __forceinline__ float time_of_idx(int i) const { return s->t0 + (float)i / c.fsEnergy; }

    // noiseRef: audio_pulse_detector.cpp:76-88
    __device__ float noise_ref(int i) const {
        int far_ = i - c.noiseRefFar, near_ = i - c.noiseRefNear;
        if (near_ <= 0 || far_ >= near_) return -1.f;
        if (far_ < 0) far_ = 0;
        if (near_ > s->n) near_ = s->n;
        if (far_ >= near_) return -1.f;
        float sum = 0.f;
        for (int j = far_; j < near_; j++) sum += ev(j);
        return sum / (float)(near_ - far_);
    }

    __device__ __forceinline__ bool last3_tight() const {
        float mn = fminf(fminf(s->last3[0], s->last3[1]), s->last3[2]);
        float mx = fmaxf(fmaxf(s->last3[0], s->last3[1]), s->last3[2]);
        return mx - mn < c.dtTolS;
    }

    __device__ void detect_rois();
    __device__ void live_etat();
    __device__ void on_energy(float v, float freq_hz);
    __device__ float est_freq() const;
};

// detectRois: spectral_pulse_detector.cpp:47-146, audio_pulse_detector.cpp:152-258
__device__ void Det::detect_rois() {
    const int n = s->n;
    const float z_s = s->locked ? 0.75f * s->tTarget : c.zDefaultS;
    int idx_z = (int)(z_s * c.fsEnergy);
    if (idx_z < 1) idx_z = 1;
    const int safe_idx = n - idx_z;
    if (safe_idx <= idx_z) return;
    const int scan_from = idx_z > s->lastScan ? idx_z : s->lastScan;

    for (int i = scan_from; i < safe_idx; i++) {
        const float val = ev(i);
        bool is_max = true;
        for (int j = i - idx_z; j <= i + idx_z && is_max; j++)
            if (j != i && ev(j) >= val) is_max = false;
        if (!is_max) continue;

        float snr;
        if (c.audio) {
            const float noise = noise_ref(i);
            if (noise <= 0.f) continue;
            snr = val / noise;
        } else {
            snr = val;
        }
        if (snr < c.snrMin) continue;

        const float t_roi = time_of_idx(i);
        const float dt = (s->tLastRoi >= 0.f) ? t_roi - s->tLastRoi : 0.f;
        int n_cycles = 1;
        float norm_dt = dt;
        if (dt > 0.f) {
            n_cycles = (int)roundf(dt / s->tTarget);
            if (n_cycles < 1) n_cycles = 1;
            if (n_cycles > 1 && fabsf(dt - n_cycles * s->tTarget) > c.dtTolS) n_cycles = 1;
            norm_dt = dt / (float)n_cycles;
        }
        const bool dans_rythme = (dt > 0.f) && (fabsf(norm_dt - s->tTarget) < c.dtTolS);
        const bool admis = (snr >= c.snrStrong) || (snr >= c.snrRhythm && dans_rythme);
        if (!admis) continue;

        if (dt > 0.f) {
            if (s->nLast3 == 3) { s->last3[0] = s->last3[1]; s->last3[1] = s->last3[2]; s->nLast3 = 2; }
            s->last3[s->nLast3++] = norm_dt;
            if (s->nLast3 == 3 && last3_tight()) {
                s->locked = 1;
                s->tTarget = (s->last3[0] + s->last3[1] + s->last3[2]) / 3.f;
            }
            if (s->nHistDt == 5) { for (int k = 0; k < 4; k++) s->histDt[k] = s->histDt[k + 1]; s->nHistDt = 4; }
            s->histDt[s->nHistDt++] = norm_dt;
            if (s->nHistN == 5) { for (int k = 0; k < 4; k++) s->histN[k] = s->histN[k + 1]; s->nHistN = 4; }
            s->histN[s->nHistN++] = n_cycles;
        }

        int base;
        if (c.audio)   // audio_pulse_detector.cpp:215-218
            base = (snr >= 2.0f) ? 5 : (snr >= 1.5f) ? 4 : (snr >= 1.2f) ? 3 : (snr >= 1.1f) ? 2 : 1;
        else           // spectral_pulse_detector.cpp:103-106
            base = (snr >= c.snrStrong) ? 5 : (snr >= 3.0f) ? 4 : (snr >= c.snrRhythm) ? 3 : (snr >= 2.0f) ? 2 : 1;
        const int pen_rythme = (dt > 0.f && !dans_rythme) ? 2 : 0;
        int pen_confusion = 0;
        if (s->nHistDt >= 4) {
            float disp = 0.f;
            for (int j = 1; j < s->nHistDt; j++) disp += fabsf(s->histDt[j] - s->histDt[j - 1]);
            int sum_n = 0;
            for (int j = 0; j < s->nHistN; j++) sum_n += s->histN[j];
            if (disp > c.dispersionMax || sum_n > c.sumNMax) pen_confusion = 2;
        }
        if (s->nLast3 == 3 && last3_tight()) pen_confusion = 0;
        int etat = base - pen_rythme - pen_confusion;
        if (etat < 0) etat = 0;

        // rois_.push_back
        if (s->nRois == PULSE_MAX_ROIS) {   // never with sane rates; drop the oldest and count it
            s->roiHead = (s->roiHead + 1) % PULSE_MAX_ROIS;
            s->nRois--;
            s->overflow++;
        }
        {
            const int k = (s->roiHead + s->nRois) % PULSE_MAX_ROIS;
            s->roiT[k] = t_roi;
            s->roiEtat[k] = etat;
            s->nRois++;
        }
        s->tLastRoi = t_roi;
        s->lastSnr = snr;
        if (!c.audio) {   // freqHistory_, spectral_pulse_detector.cpp:133-134
            if (s->nFh == PULSE_FREQ_HIST) { s->fhHead = (s->fhHead + 1) % PULSE_FREQ_HIST; s->nFh--; }
            const int k = (s->fhHead + s->nFh) % PULSE_FREQ_HIST;
            int r = s->head + i;
            if (r >= c.ringCap) r -= c.ringCap;
            s->fhT[k] = t_roi;
            s->fhF[k] = fq[r];
            s->nFh++;
        }
        const float cutoff = t_roi - 20.f;
        while (s->nRois > 0 && s->roiT[s->roiHead] < cutoff) {
            s->roiHead = (s->roiHead + 1) % PULSE_MAX_ROIS;
            s->nRois--;
        }
        i += idx_z;
        s->lastScan = i + 1;
    }
    if (s->lastScan < safe_idx) s->lastScan = safe_idx;
}

// computeLiveEtat + toLevel: spectral_pulse_detector.cpp:150-169, audio_pulse_detector.cpp:262-277
__device__ void Det::live_etat() {
    int etat = 0;
    if (s->nRois > 0) {
        const float now = time_of_idx(s->n - 1);
        const float window_start = now - c.liveWindowT * s->tTarget;
        float sum = 0.f;
        for (int k = 0; k < s->nRois; k++) {
            const int r = (s->roiHead + k) % PULSE_MAX_ROIS;
            if (s->roiT[r] >= window_start) sum += (float)s->roiEtat[r];
        }
        etat = (int)floorf(sum / c.liveDivisor);
        if (etat > 5) etat = 5;
    }
    s->liveEtat = etat;
    s->level = etat >= 5 ? 3 : etat >= 3 ? 2 : etat >= 1 ? 1 : 0;
}

// onEnergyFrame: spectral_pulse_detector.cpp:28-43, audio_pulse_detector.cpp:131-148 (after smoothing)
__device__ void Det::on_energy(float v, float freq_hz) {
    if (s->n == 0) s->t0 = 0.f;
    {
        int k = s->head + s->n;
        if (k >= c.ringCap) k -= c.ringCap;
        e[k] = v;
        if (fq) fq[k] = freq_hz;
        s->n++;
    }
    while (s->n > c.maxBuf) {
        s->head = s->head + 1 == c.ringCap ? 0 : s->head + 1;
        s->n--;
        s->t0 += 1.f / c.fsEnergy;
        if (s->lastScan > 0) s->lastScan--;
    }
    detect_rois();
    live_etat();
}

// estimatedFreqHz: spectral_pulse_detector.cpp:173-196 (float products, double sums)
__device__ float Det::est_freq() const {
    const int n = s->nFh;
    if (n < 2) return 0.f;
    const float t_now = time_of_idx(s->n - 1);
    double sum_t = 0, sum_f = 0, sum_tt = 0, sum_tf = 0;
    for (int k = 0; k < n; k++) {
        const int r = (s->fhHead + k) % PULSE_FREQ_HIST;
        const float t = s->fhT[r], f = s->fhF[r];
        sum_t += t;
        sum_f += f;
        sum_tt += t * t;
        sum_tf += t * f;
    }
    const double denom = n * sum_tt - sum_t * sum_t;
    if (fabs(denom) < 1e-9) return (float)(sum_f / n);
    const double a = (n * sum_tf - sum_t * sum_f) / denom;
    const double b = (sum_f - a * sum_t) / n;
    return (float)(a * t_now + b);
}

__global__ void __launch_bounds__(64)
k_pulse_spectral(const sdrb_frame_result* __restrict__ results, PulseState* __restrict__ state,
                 float* __restrict__ ring_e, float* __restrict__ ring_f, const PulseCfg cfg,
                 sdrb_pulse_result* __restrict__ out, int n_streams, int n_blocks, int channels) {
    const int sidx = blockIdx.x * blockDim.x + threadIdx.x;
    if (sidx >= n_streams) return;
    Det d{state + sidx, ring_e + (size_t)sidx * cfg.ringCap, ring_f + (size_t)sidx * cfg.ringCap, cfg};
    for (int b = 0; b < n_blocks; b++) {
        const sdrb_frame_result& r = results[(size_t)sidx * n_blocks + b];
        // spectralPulseDetector.process(best1kHzSnrSigma, best1kFreqHz)  src/sdr-bridge-java-soapy.cpp:478-479
        const float sigma = r.best1kHzSnrSigma, freq = r.best1kHzCenterFreqHz;
        d.on_energy(sigma, freq);
        sdrb_pulse_result& o = out[((size_t)sidx * channels) * n_blocks + b];
        const float ef = d.est_freq();
        o.spectralSnrSigma = sigma;
        o.spectralLiveEtat = d.s->liveEtat;
        o.spectralEstFreqHzF = ef;
        o.spectralEstFreqHz = llroundf(ef);   // std::llround(float), src/sdr-bridge-java-soapy.cpp:481-482
        o.spectralLevel = d.s->level;
        o.spectralLocked = d.s->locked;
        o.spectralLastSnr = d.s->lastSnr;
        o.spectralTTarget = d.s->tTarget;
        o.overflow = d.s->overflow;
    }
}

// applyChain for one second-order section, Direct Form II transposed (audio_pulse_detector.cpp:55-63)
__device__ __forceinline__ float sos_step(const float* k, float* z, float x) {
    const float y = k[0] * x + z[0];
    z[0] = k[1] * x - k[3] * y + z[1];
    z[1] = k[2] * x - k[4] * y;
    return y;
}

__global__ void __launch_bounds__(64)
k_pulse_audio(const int16_t* __restrict__ pcm, PulseState* __restrict__ state, float* __restrict__ ring_e,
              const PulseCfg cfg, sdrb_pulse_result* __restrict__ out, int n_units, int n_blocks, int m) {
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= n_units) return;
    Det d{state + u, ring_e + (size_t)u * cfg.ringCap, nullptr, cfg};
    PulseState* s = d.s;
    // the per-sample filter state lives in registers inside a block
    float z0[2] = {s->bz[0][0], s->bz[0][1]}, z1[2] = {s->bz[1][0], s->bz[1][1]};
    float lz[2] = {s->lz[0], s->lz[1]};
    int frameCount = s->frameCount;
    float frameAcc = s->frameAcc;
    constexpr float inv = 1.f / 32767.f;
    for (int b = 0; b < n_blocks; b++) {
        const int16_t* p = pcm + ((size_t)u * n_blocks + b) * m;
        // AudioPulseDetector::process(const std::vector<int16_t>&)  audio_pulse_detector.cpp:112-125
        for (int i = 0; i < m; i++) {
            float x = (float)p[i] * inv;
            x = sos_step(cfg.band[0], z0, x);
            x = sos_step(cfg.band[1], z1, x);
            frameAcc += x * x;
            frameCount++;
            if (frameCount >= cfg.frameSamples) {
                const float rms = sqrtf(frameAcc / (float)cfg.frameSamples);
                d.on_energy(sos_step(cfg.low, lz, rms), 0.f);
                frameAcc = 0.f;
                frameCount = 0;
            }
        }
        sdrb_pulse_result& o = out[(size_t)u * n_blocks + b];
        o.audioLastSnr = s->lastSnr;
        o.audioLiveEtat = s->liveEtat;
        o.audioLevel = s->level;
        o.audioLocked = s->locked;
        o.audioTTarget = s->tTarget;
        o.overflow += s->overflow;
    }
    s->bz[0][0] = z0[0]; s->bz[0][1] = z0[1]; s->bz[1][0] = z1[0]; s->bz[1][1] = z1[1];
    s->lz[0] = lz[0]; s->lz[1] = lz[1];
    s->frameCount = frameCount;
    s->frameAcc = frameAcc;
}

}  // namespace

cudaError_t launch_pulse_spectral(const sdrb_frame_result* d_results, PulseState* d_state, float* d_ring_e,
                                  float* d_ring_f, const PulseCfg& cfg, sdrb_pulse_result* d_out, int n_streams,
                                  int n_blocks, int channels, cudaStream_t st) {
    if (n_streams <= 0 || n_blocks <= 0) return cudaSuccess;
    k_pulse_spectral<<<(n_streams + 63) / 64, 64, 0, st>>>(d_results, d_state, d_ring_e, d_ring_f, cfg, d_out,
                                                             n_streams, n_blocks, channels);
    return cudaGetLastError();
}

cudaError_t launch_pulse_audio(const int16_t* d_pcm, PulseState* d_state, float* d_ring_e, const PulseCfg& cfg,
                               sdrb_pulse_result* d_out, int n_units, int n_blocks, int m, cudaStream_t st) {
    if (n_units <= 0 || n_blocks <= 0 || m <= 0) return cudaSuccess;
    k_pulse_audio<<<(n_units + 63) / 64, 64, 0, st>>>(d_pcm, d_state, d_ring_e, cfg, d_out, n_units, n_blocks, m);
    return cudaGetLastError();
}

}  // namespace sdrb
