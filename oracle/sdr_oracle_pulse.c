/* oracle/sdr_oracle_pulse.c — plain-C restatement of SpectralPulseDetector and AudioPulseDetector.
 * TEST INFRASTRUCTURE ONLY (see sdr_oracle_pulse.h).  Every function cites the reference lines it
 * follows; float expressions keep the reference's evaluation order (compile with -ffp-contract=off). */
#include "sdr_oracle_pulse.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define MAX_ROIS 4096
#define FREQ_HIST_MAX 30 /* spectral_pulse_detector.h:62 */

struct sdr_oracle_pulse {
    sdr_oracle_pulse_cfg cfg;
    int audio;
    /* energy buffer (std::deque<float> eBuf_, freqBuf_) */
    float *e, *fq;
    int n, cap;
    float t0;
    /* ROI store */
    float roi_t[MAX_ROIS];
    int roi_etat[MAX_ROIS];
    int n_rois;
    float t_last_roi;
    int last_scan;
    /* frequency history (spectral) */
    float fh_t[FREQ_HIST_MAX], fh_f[FREQ_HIST_MAX];
    int n_fh;
    /* phase lock */
    int locked;
    float t_target;
    float last3[3]; int n_last3;
    float hist_dt[5]; int hist_n[5]; int n_hist_dt, n_hist_n;
    /* outputs */
    int live_etat, level;
    float last_snr;
    /* audio front end */
    sdr_oracle_sos band[2], low[1];
    int frame_samples, frame_count;
    float frame_acc;
};

void sdr_oracle_pulse_cfg_spectral(sdr_oracle_pulse_cfg* c, float fs_energy) {
    /* spectral_pulse_detector.h:22-34; fsEnergy from sdr-bridge-java-soapy.cpp:1132-1137 */
    memset(c, 0, sizeof *c);
    c->fsEnergy = fs_energy; c->zDefaultS = 0.666f; c->tTargetInit = 1.75f; c->dtTolS = 0.150f;
    c->snrMin = 1.5f; c->snrRhythm = 2.5f; c->snrStrong = 4.0f; c->dispersionMax = 1.3f;
    c->sumNMax = 7; c->liveWindowT = 4.0f; c->liveDivisor = 3.0f;
}

void sdr_oracle_pulse_cfg_audio(sdr_oracle_pulse_cfg* c) {
    /* audio_pulse_detector.h:20-38 */
    memset(c, 0, sizeof *c);
    c->sampleRate = 48000.f; c->fMin = 1500.f; c->fMax = 4000.f; c->fsEnergy = 100.f; c->smoothCutoff = 5.f;
    c->zDefaultS = 0.666f; c->tTargetInit = 1.75f; c->dtTolS = 0.150f;
    c->snrMin = 1.0f; c->snrRhythm = 1.1f; c->snrStrong = 2.0f; c->dispersionMax = 1.3f;
    c->sumNMax = 7; c->liveWindowT = 4.0f; c->liveDivisor = 3.0f; c->noiseRefFar = 80; c->noiseRefNear = 40;
}

/* audio_pulse_detector.cpp:27-38 (kPi is the float literal of :4) */
void sdr_oracle_pulse_lp2(float fs, float fc, float q, sdr_oracle_sos* s) {
    const float kPi = 3.14159265358979f;
    const float K = tanf(kPi * fc / fs), K2 = K * K, norm = K2 + K / q + 1.f;
    s->b0 = K2 / norm; s->b1 = 2.f * K2 / norm; s->b2 = K2 / norm;
    s->a1 = 2.f * (K2 - 1.f) / norm; s->a2 = (K2 - K / q + 1.f) / norm;
    s->z1 = s->z2 = 0.f;
}
/* audio_pulse_detector.cpp:40-51 */
void sdr_oracle_pulse_hp2(float fs, float fc, float q, sdr_oracle_sos* s) {
    const float kPi = 3.14159265358979f;
    const float K = tanf(kPi * fc / fs), K2 = K * K, norm = K2 + K / q + 1.f;
    s->b0 = 1.f / norm; s->b1 = -2.f / norm; s->b2 = 1.f / norm;
    s->a1 = 2.f * (K2 - 1.f) / norm; s->a2 = (K2 - K / q + 1.f) / norm;
    s->z1 = s->z2 = 0.f;
}

sdr_oracle_pulse* sdr_oracle_pulse_create(const sdr_oracle_pulse_cfg* cfg, int audio) {
    sdr_oracle_pulse* p = (sdr_oracle_pulse*)calloc(1, sizeof *p);
    p->cfg = *cfg;
    p->audio = audio;
    p->cap = 64;
    p->e = (float*)malloc(sizeof(float) * p->cap);
    p->fq = (float*)malloc(sizeof(float) * p->cap);
    p->t_last_roi = -1.f;
    p->t_target = cfg->tTargetInit;   /* spectral_pulse_detector.cpp:3-4, audio_pulse_detector.cpp:10-11 */
    if (audio) {
        /* audio_pulse_detector.cpp:13-22 */
        sdr_oracle_pulse_hp2(cfg->sampleRate, cfg->fMin, 0.7071f, &p->band[0]);
        sdr_oracle_pulse_lp2(cfg->sampleRate, cfg->fMax, 0.7071f, &p->band[1]);
        sdr_oracle_pulse_lp2(cfg->fsEnergy, cfg->smoothCutoff, 0.7071f, &p->low[0]);
        const int fsamp = (int)(cfg->sampleRate / cfg->fsEnergy);
        p->frame_samples = fsamp > 1 ? fsamp : 1;
    }
    return p;
}

void sdr_oracle_pulse_destroy(sdr_oracle_pulse* p) {
    if (!p) return;
    free(p->e); free(p->fq); free(p);
}

void sdr_oracle_pulse_configure(sdr_oracle_pulse* p, const sdr_oracle_pulse_cfg* cfg) { p->cfg = *cfg; }

/* ---- helpers ------------------------------------------------------------------------------ */
/* timeOfIdx: spectral_pulse_detector.cpp:17-19, audio_pulse_detector.cpp:66-68 */
static float time_of_idx(const sdr_oracle_pulse* p, int i) { return p->t0 + (float)i / p->cfg.fsEnergy; }

/* noiseRef: audio_pulse_detector.cpp:76-88 */
static float noise_ref(const sdr_oracle_pulse* p, int i) {
    int far_ = i - p->cfg.noiseRefFar, near_ = i - p->cfg.noiseRefNear;
    if (near_ <= 0 || far_ >= near_) return -1.f;
    if (far_ < 0) far_ = 0;
    if (near_ > p->n) near_ = p->n;
    if (far_ >= near_) return -1.f;
    float sum = 0.f;
    for (int j = far_; j < near_; j++) sum += p->e[j];
    return sum / (float)(near_ - far_);
}

static void push3(float* a, int* n, int cap, float v) {   /* push_back + pop_front when over cap */
    if (*n == cap) { memmove(a, a + 1, sizeof(float) * (cap - 1)); (*n)--; }
    a[(*n)++] = v;
}
static void pushi(int* a, int* n, int cap, int v) {
    if (*n == cap) { memmove(a, a + 1, sizeof(int) * (cap - 1)); (*n)--; }
    a[(*n)++] = v;
}

/* detectRois: spectral_pulse_detector.cpp:47-146, audio_pulse_detector.cpp:152-258 */
static void detect_rois(sdr_oracle_pulse* p) {
    const sdr_oracle_pulse_cfg* c = &p->cfg;
    const int n = p->n;
    const float z_s = p->locked ? 0.75f * p->t_target : c->zDefaultS;
    int idx_z = (int)(z_s * c->fsEnergy);
    if (idx_z < 1) idx_z = 1;
    const int safe_idx = n - idx_z;
    if (safe_idx <= idx_z) return;
    int scan_from = idx_z > p->last_scan ? idx_z : p->last_scan;

    for (int i = scan_from; i < safe_idx; i++) {
        const float val = p->e[i];
        int is_max = 1;
        for (int j = i - idx_z; j <= i + idx_z && is_max; j++)
            if (j != i && p->e[j] >= val) is_max = 0;
        if (!is_max) continue;

        float snr;
        if (p->audio) {
            const float noise = noise_ref(p, i);
            if (noise <= 0.f) continue;
            snr = val / noise;
        } else {
            snr = val;
        }
        if (snr < c->snrMin) continue;

        const float t_roi = time_of_idx(p, i);
        const float dt = (p->t_last_roi >= 0.f) ? t_roi - p->t_last_roi : 0.f;
        int n_cycles = 1;
        float norm_dt = dt;
        if (dt > 0.f) {
            n_cycles = (int)roundf(dt / p->t_target);
            if (n_cycles < 1) n_cycles = 1;
            if (n_cycles > 1 && fabsf(dt - n_cycles * p->t_target) > c->dtTolS) n_cycles = 1;
            norm_dt = dt / (float)n_cycles;
        }
        const int dans_rythme = (dt > 0.f) && (fabsf(norm_dt - p->t_target) < c->dtTolS);
        const int admis = (snr >= c->snrStrong) || (snr >= c->snrRhythm && dans_rythme);
        if (!admis) continue;

        if (dt > 0.f) {
            push3(p->last3, &p->n_last3, 3, norm_dt);
            if (p->n_last3 == 3) {
                float mn = p->last3[0], mx = p->last3[0];
                for (int k = 1; k < 3; k++) { if (p->last3[k] < mn) mn = p->last3[k]; if (p->last3[k] > mx) mx = p->last3[k]; }
                if (mx - mn < c->dtTolS) {
                    p->locked = 1;
                    p->t_target = (p->last3[0] + p->last3[1] + p->last3[2]) / 3.f;
                }
            }
            push3(p->hist_dt, &p->n_hist_dt, 5, norm_dt);
            pushi(p->hist_n, &p->n_hist_n, 5, n_cycles);
        }

        int base;
        if (p->audio)   /* audio_pulse_detector.cpp:215-218 (literal thresholds) */
            base = (snr >= 2.0f) ? 5 : (snr >= 1.5f) ? 4 : (snr >= 1.2f) ? 3 : (snr >= 1.1f) ? 2 : 1;
        else            /* spectral_pulse_detector.cpp:103-106 */
            base = (snr >= c->snrStrong) ? 5 : (snr >= 3.0f) ? 4 : (snr >= c->snrRhythm) ? 3 : (snr >= 2.0f) ? 2 : 1;

        const int pen_rythme = (dt > 0.f && !dans_rythme) ? 2 : 0;
        int pen_confusion = 0;
        if (p->n_hist_dt >= 4) {
            float disp = 0.f;
            for (int j = 1; j < p->n_hist_dt; j++) disp += fabsf(p->hist_dt[j] - p->hist_dt[j - 1]);
            int sum_n = 0;
            for (int j = 0; j < p->n_hist_n; j++) sum_n += p->hist_n[j];
            if (disp > c->dispersionMax || sum_n > c->sumNMax) pen_confusion = 2;
        }
        if (p->n_last3 == 3) {
            float mn = p->last3[0], mx = p->last3[0];
            for (int k = 1; k < 3; k++) { if (p->last3[k] < mn) mn = p->last3[k]; if (p->last3[k] > mx) mx = p->last3[k]; }
            if (mx - mn < c->dtTolS) pen_confusion = 0;
        }
        int etat = base - pen_rythme - pen_confusion;
        if (etat < 0) etat = 0;

        if (p->n_rois < MAX_ROIS) { p->roi_t[p->n_rois] = t_roi; p->roi_etat[p->n_rois] = etat; p->n_rois++; }
        p->t_last_roi = t_roi;
        p->last_snr = snr;
        if (!p->audio) {   /* spectral_pulse_detector.cpp:133-134 */
            if (p->n_fh == FREQ_HIST_MAX) {
                memmove(p->fh_t, p->fh_t + 1, sizeof(float) * (FREQ_HIST_MAX - 1));
                memmove(p->fh_f, p->fh_f + 1, sizeof(float) * (FREQ_HIST_MAX - 1));
                p->n_fh--;
            }
            p->fh_t[p->n_fh] = t_roi; p->fh_f[p->n_fh] = p->fq[i]; p->n_fh++;
        }
        const float cutoff = t_roi - 20.f;
        int drop = 0;
        while (drop < p->n_rois && p->roi_t[drop] < cutoff) drop++;
        if (drop) {
            memmove(p->roi_t, p->roi_t + drop, sizeof(float) * (p->n_rois - drop));
            memmove(p->roi_etat, p->roi_etat + drop, sizeof(int) * (p->n_rois - drop));
            p->n_rois -= drop;
        }
        i += idx_z;
        p->last_scan = i + 1;
    }
    if (p->last_scan < safe_idx) p->last_scan = safe_idx;
}

/* computeLiveEtat + toLevel: spectral_pulse_detector.cpp:150-169, audio_pulse_detector.cpp:262-277 */
static void live_etat(sdr_oracle_pulse* p) {
    int etat = 0;
    if (p->n_rois > 0) {
        const float now = time_of_idx(p, p->n - 1);
        const float window_start = now - p->cfg.liveWindowT * p->t_target;
        float sum = 0.f;
        for (int k = 0; k < p->n_rois; k++)
            if (p->roi_t[k] >= window_start) sum += (float)p->roi_etat[k];
        etat = (int)floorf(sum / p->cfg.liveDivisor);
        if (etat > 5) etat = 5;
    }
    p->live_etat = etat;
    p->level = etat >= 5 ? 3 : etat >= 3 ? 2 : etat >= 1 ? 1 : 0;
}

/* onEnergyFrame after the smoothing: spectral_pulse_detector.cpp:28-43, audio_pulse_detector.cpp:131-148 */
static void on_energy(sdr_oracle_pulse* p, float v, float freq_hz) {
    if (p->n == 0) p->t0 = 0.f;
    if (p->n == p->cap) {
        p->cap *= 2;
        p->e = (float*)realloc(p->e, sizeof(float) * p->cap);
        p->fq = (float*)realloc(p->fq, sizeof(float) * p->cap);
    }
    p->e[p->n] = v; p->fq[p->n] = freq_hz; p->n++;
    const int max_buf = (int)(10.f * p->cfg.fsEnergy);
    while (p->n > max_buf) {
        memmove(p->e, p->e + 1, sizeof(float) * (p->n - 1));
        memmove(p->fq, p->fq + 1, sizeof(float) * (p->n - 1));
        p->n--;
        p->t0 += 1.f / p->cfg.fsEnergy;
        if (p->last_scan > 0) p->last_scan--;
    }
    detect_rois(p);
    live_etat(p);
}

void sdr_oracle_pulse_spectral(sdr_oracle_pulse* p, float snr_sigma, float freq_hz) { on_energy(p, snr_sigma, freq_hz); }

/* applyChain: audio_pulse_detector.cpp:55-63 (Direct Form II transposed) */
static float apply_chain(sdr_oracle_sos* chain, int n, float x) {
    for (int k = 0; k < n; k++) {
        sdr_oracle_sos* s = &chain[k];
        const float y = s->b0 * x + s->z1;
        s->z1 = s->b1 * x - s->a1 * y + s->z2;
        s->z2 = s->b2 * x - s->a2 * y;
        x = y;
    }
    return x;
}

void sdr_oracle_pulse_audio_pcm(sdr_oracle_pulse* p, const int16_t* pcm, int n) {
    const float inv = 1.f / 32767.f;
    for (int i = 0; i < n; i++) {
        const float filtered = apply_chain(p->band, 2, pcm[i] * inv);
        p->frame_acc += filtered * filtered;
        p->frame_count++;
        if (p->frame_count >= p->frame_samples) {
            const float rms = sqrtf(p->frame_acc / p->frame_samples);
            on_energy(p, apply_chain(p->low, 1, rms), 0.f);
            p->frame_acc = 0.f;
            p->frame_count = 0;
        }
    }
}

int   sdr_oracle_pulse_live_etat(const sdr_oracle_pulse* p) { return p->live_etat; }
int   sdr_oracle_pulse_level(const sdr_oracle_pulse* p) { return p->level; }
int   sdr_oracle_pulse_locked(const sdr_oracle_pulse* p) { return p->locked; }
float sdr_oracle_pulse_last_snr(const sdr_oracle_pulse* p) { return p->last_snr; }
float sdr_oracle_pulse_t_target(const sdr_oracle_pulse* p) { return p->t_target; }
int   sdr_oracle_pulse_n_rois(const sdr_oracle_pulse* p) { return p->n_rois; }

/* estimatedFreqHz: spectral_pulse_detector.cpp:173-196 (float products, double sums) */
float sdr_oracle_pulse_est_freq(const sdr_oracle_pulse* p) {
    const int n = p->n_fh;
    if (n < 2) return 0.f;
    const float t_now = time_of_idx(p, p->n - 1);
    double sum_t = 0, sum_f = 0, sum_tt = 0, sum_tf = 0;
    for (int k = 0; k < n; k++) {
        sum_t += p->fh_t[k];
        sum_f += p->fh_f[k];
        sum_tt += p->fh_t[k] * p->fh_t[k];
        sum_tf += p->fh_t[k] * p->fh_f[k];
    }
    const double denom = n * sum_tt - sum_t * sum_t;
    if (fabs(denom) < 1e-9) return (float)(sum_f / n);
    const double a = (n * sum_tf - sum_t * sum_f) / denom;
    const double b = (sum_f - a * sum_t) / n;
    return (float)(a * t_now + b);
}
