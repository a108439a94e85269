/* oracle/sdr_oracle_pulse.h — CPU restatement of the reference's two beacon-pulse detectors
 * (SURVEY.md 8(f) rows 2 and 3).  TEST INFRASTRUCTURE ONLY: only tests/, __graft_entry__.smoke()
 * and bench.py's CPU legs may load this; the product path never does.
 *
 *   SpectralPulseDetector   src/dsp/spectral_pulse_detector.{h,cpp}
 *   AudioPulseDetector      src/ssb/audio_pulse_detector.{h,cpp}
 *
 * Pinned bit-exactly against the reference's own sources compiled into oracle/_ref/libsdr_ref.so
 * (tests/test_oracle_pulse.py).  Paths cited file:line are relative to the reference repository. */
#ifndef SDR_ORACLE_PULSE_H
#define SDR_ORACLE_PULSE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* SpectralPulseDetector::Config (spectral_pulse_detector.h:22-34) and the shared tail of
 * AudioPulseDetector::Config (audio_pulse_detector.h:20-38). */
typedef struct {
    float fsEnergy, zDefaultS, tTargetInit, dtTolS, snrMin, snrRhythm, snrStrong, dispersionMax;
    int32_t sumNMax;
    float liveWindowT, liveDivisor;
    /* audio only */
    float sampleRate, fMin, fMax, smoothCutoff;
    int32_t noiseRefFar, noiseRefNear;
} sdr_oracle_pulse_cfg;

void sdr_oracle_pulse_cfg_spectral(sdr_oracle_pulse_cfg* c, float fs_energy);
void sdr_oracle_pulse_cfg_audio(sdr_oracle_pulse_cfg* c);

typedef struct { float b0, b1, b2, a1, a2, z1, z2; } sdr_oracle_sos;   /* audio_pulse_detector.h:60 */

typedef struct sdr_oracle_pulse sdr_oracle_pulse;

/* audio = 0: SpectralPulseDetector, audio = 1: AudioPulseDetector */
sdr_oracle_pulse* sdr_oracle_pulse_create(const sdr_oracle_pulse_cfg* cfg, int audio);
void sdr_oracle_pulse_destroy(sdr_oracle_pulse* p);
/* SpectralPulseDetector::configure (spectral_pulse_detector.cpp:6-8): replaces the config only */
void sdr_oracle_pulse_configure(sdr_oracle_pulse* p, const sdr_oracle_pulse_cfg* cfg);
/* SpectralPulseDetector::process (spectral_pulse_detector.cpp:10-13) */
void sdr_oracle_pulse_spectral(sdr_oracle_pulse* p, float snr_sigma, float freq_hz);
/* AudioPulseDetector::process(const std::vector<int16_t>&) (audio_pulse_detector.cpp:112-125) */
void sdr_oracle_pulse_audio_pcm(sdr_oracle_pulse* p, const int16_t* pcm, int n);
/* outputs: liveEtat, level, locked, lastSnr, tTarget, estimatedFreqHz (spectral only) */
int   sdr_oracle_pulse_live_etat(const sdr_oracle_pulse* p);
int   sdr_oracle_pulse_level(const sdr_oracle_pulse* p);
int   sdr_oracle_pulse_locked(const sdr_oracle_pulse* p);
float sdr_oracle_pulse_last_snr(const sdr_oracle_pulse* p);
float sdr_oracle_pulse_t_target(const sdr_oracle_pulse* p);
float sdr_oracle_pulse_est_freq(const sdr_oracle_pulse* p);   /* spectral_pulse_detector.cpp:173-196 */
int   sdr_oracle_pulse_n_rois(const sdr_oracle_pulse* p);
/* biquad designs (audio_pulse_detector.cpp:27-53) */
void sdr_oracle_pulse_lp2(float fs, float fc, float q, sdr_oracle_sos* s);
void sdr_oracle_pulse_hp2(float fs, float fc, float q, sdr_oracle_sos* s);

#ifdef __cplusplus
}
#endif
#endif
