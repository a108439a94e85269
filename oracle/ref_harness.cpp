/* oracle/ref_harness.cpp — C-ABI harness around the UNMODIFIED reference sources.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path may link or call this.
 * It is compiled by oracle/Makefile together with the reference's own files, in
 * place under /root/reference (never copied):
 *     src/dsp/fft_process.cpp, src/ssb/ssb_demod_opt.cpp
 * and linked with the reference's vendored x86_64 libfftw3f.a, into
 * oracle/_ref/libsdr_ref.so.  The only alteration of the reference's behaviour
 * is its time source (oracle/clock_shim.h).
 *
 * Private members of FFTProcessor (the tracking candidate) are read through the
 * `#define private public` trick, applied only after the standard headers have
 * been included by the force-included shim. */
#include "clock_shim.h"
#include "jni.h"

#define private public
#include "dsp/fft_process.h"
#undef private
#include "ssb/ssb_demod_opt.h"
#include "sdr-bridge-internal.h"

#include <cstring>

/* Globals normally defined in src/sdr-bridge-java-soapy.cpp:337-340. */
namespace sdr_bridge_internal {
std::atomic<bool> isCenterFrequencyChanged{false};
JavaVM* gJavaVM = nullptr;
}

static int64_t g_now_ns = 0;
std::chrono::sdr_frame_clock::time_point std::chrono::sdr_frame_clock::now() noexcept {
    return time_point(duration(g_now_ns));
}

extern "C" {

/* ------------------------------ FFT side ------------------------------ */
void* sdr_ref_fft_create(void) { return new FFTProcessor(); }
void sdr_ref_fft_destroy(void* h) { delete static_cast<FFTProcessor*>(h); }

void sdr_ref_fft_configure(void* h, uint32_t fc, uint32_t fs, int n, int focus_khz) {
    FftProcessorConfig c;
    c.centerFrequency = fc; c.sampleRate = fs; c.samplesPerReading = n; c.freqFocusRangeKhz = focus_khz;
    static_cast<FFTProcessor*>(h)->configure(c);
}

void sdr_ref_set_center_frequency_changed(int flag) {
    sdr_bridge_internal::isCenterFrequencyChanged = (flag != 0);
}

/* iq: interleaved re,im float32; now_ns: frame clock for this call. */
void sdr_ref_fft_process(void* h, const float* iq, uint32_t n, int64_t now_ns) {
    g_now_ns = now_ns;
    static_cast<FFTProcessor*>(h)->process(reinterpret_cast<const std::complex<float>*>(iq), n);
}

/* out[0..12]: meanSnrDb, meanSnrSigma, peakAboveNoiseMeanDb, maxBinSnrDb, maxBinSnrSigma,
 * best1kHzSnrDb, best1kHzSnrSigma, best1kHzCenterFreqHz, perBinMean,
 * trackingFrequency(float, private), candidate peak dB, candidate peak Hz, (unused). */
void sdr_ref_fft_results(void* h, float* out, int64_t* tracking_long, int* detection_flag) {
    FFTProcessor* p = static_cast<FFTProcessor*>(h);
    out[0] = p->getMeanSnrDb();
    out[1] = p->getMeanSnrSigma();
    out[2] = p->getPeakAboveNoiseMeanDb();
    out[3] = p->getMaxBinSnrDb();
    out[4] = p->getMaxBinSnrSigma();
    out[5] = p->getBest1kHzSnrDb();
    out[6] = p->getBest1kHzSnrSigma();
    out[7] = p->getBest1kHzCenterFreqHz();
    out[8] = p->getPerBinMean();
    out[9] = p->trackingFrequency;
    out[10] = p->maxPeakAndFrequency.size() > 0 ? p->maxPeakAndFrequency[0] : 0.0f;
    out[11] = p->maxPeakAndFrequency.size() > 1 ? p->maxPeakAndFrequency[1] : 0.0f;
    out[12] = 0.0f;
    *tracking_long = static_cast<int64_t>(p->getTrackingFrequency());
    *detection_flag = p->getDetectionFlag();
}

int sdr_ref_fft_power(void* h, float* out, int cap) {
    const std::vector<float>& v = static_cast<FFTProcessor*>(h)->getPowerSpectrum();
    int n = static_cast<int>(v.size());
    if (n > cap) n = cap;
    std::memcpy(out, v.data(), sizeof(float) * n);
    return static_cast<int>(v.size());
}

/* ------------------------------ SSB stages ------------------------------ */
void sdr_ref_convertIQ(const unsigned char* buf, int len_bytes, float* iq_out) {
    std::vector<std::complex<float>> iq;
    convertIQ(buf, len_bytes, iq);
    std::memcpy(iq_out, iq.data(), sizeof(std::complex<float>) * iq.size());
}

void sdr_ref_removeDC(float* iq, int n, float alpha) {
    std::vector<std::complex<float>> v(reinterpret_cast<std::complex<float>*>(iq),
                                       reinterpret_cast<std::complex<float>*>(iq) + n);
    removeDC(v, alpha);
    std::memcpy(iq, v.data(), sizeof(std::complex<float>) * n);
}

/* cs[7] = a0,a1,a2,b1,b2,z1,z2 (layout of struct IIR2 / struct Biquad). */
void sdr_ref_iir2_init_lowpass(float* cs, float fs, float fc, float q) {
    IIR2 f; iir2InitLowpass(f, fs, fc, q); std::memcpy(cs, &f, sizeof(float) * 7);
}
void sdr_ref_iir2_process(float* cs, float* iq, int n) {
    IIR2 f; std::memcpy(&f, cs, sizeof(float) * 7);
    std::vector<std::complex<float>> v(reinterpret_cast<std::complex<float>*>(iq),
                                       reinterpret_cast<std::complex<float>*>(iq) + n);
    iir2Process(f, v);
    std::memcpy(iq, v.data(), sizeof(std::complex<float>) * n);
    std::memcpy(cs, &f, sizeof(float) * 7);
}
void sdr_ref_demodSSB(const float* iq, int n, int upper, float* audio) {
    std::vector<std::complex<float>> v(reinterpret_cast<const std::complex<float>*>(iq),
                                       reinterpret_cast<const std::complex<float>*>(iq) + n);
    std::vector<float> a;
    demodSSB(v, a, upper != 0);
    std::memcpy(audio, a.data(), sizeof(float) * n);
}
void sdr_ref_adaptiveAGC(float* audio, int n, float target, float fast, float slow) {
    std::vector<float> a(audio, audio + n);
    adaptiveAGC(a, target, fast, slow);
    std::memcpy(audio, a.data(), sizeof(float) * n);
}
int sdr_ref_fir_decimate(const float* in, int n, int decim, float cutoff, float* out, int cap) {
    std::vector<float> a(in, in + n);
    std::vector<float> o = simpleFIRDecimate(a, decim, cutoff);
    int m = static_cast<int>(o.size());
    std::memcpy(out, o.data(), sizeof(float) * (m < cap ? m : cap));
    return m;
}
void sdr_ref_biquad_init_highpass(float* cs, float fs, float f0, float q) {
    Biquad f; biquadInitHighpass(f, fs, f0, q); std::memcpy(cs, &f, sizeof(float) * 7);
}
void sdr_ref_biquad_init_bandpass(float* cs, float fs, float f0, float q) {
    Biquad f; biquadInitBandpass(f, fs, f0, q); std::memcpy(cs, &f, sizeof(float) * 7);
}
void sdr_ref_biquad_process(float* cs, float* x, int n) {
    Biquad f; std::memcpy(&f, cs, sizeof(float) * 7);
    std::vector<float> v(x, x + n);
    biquadProcess(f, v);
    std::memcpy(x, v.data(), sizeof(float) * n);
    std::memcpy(cs, &f, sizeof(float) * 7);
}
void sdr_ref_transient_boost(float* x, int n, float coeff) {
    std::vector<float> v(x, x + n);
    transientBoost(v, coeff);
    std::memcpy(x, v.data(), sizeof(float) * n);
}
void sdr_ref_float_to_pcm(const float* in, int n, float gain, int16_t* out) {
    std::vector<float> v(in, in + n);
    std::vector<int16_t> p = floatToPCM(v, gain);
    std::memcpy(out, p.data(), sizeof(int16_t) * n);
}

/* Whole pipeline.  processSSB_opt keeps its filter state in function statics
 * (src/ssb/ssb_demod_opt.cpp:224-225,261,277): ONE stream per process, and the
 * first call freezes block length, sample rate and low-pass mode. */
int sdr_ref_process_ssb(const float* iq, int n, uint32_t fs, int upper, int mode, int16_t* out, int cap) {
    std::vector<std::complex<float>> v(reinterpret_cast<const std::complex<float>*>(iq),
                                       reinterpret_cast<const std::complex<float>*>(iq) + n);
    std::vector<int16_t> pcm;
    bool pulse = false;
    processSSB_opt(v, fs, upper != 0, pcm, pulse, mode);
    int m = static_cast<int>(pcm.size());
    std::memcpy(out, pcm.data(), sizeof(int16_t) * (m < cap ? m : cap));
    return m;
}

/* Timed loop for the CPU baseline (bench.py --impl reference / cpu_baseline):
 * runs `frames` readings of n samples through FFTProcessor::process (and, when
 * do_ssb, processSSB_opt), exactly the two calls soapyCallback makes per reading
 * (src/sdr-bridge-java-soapy.cpp:441-444).  Returns a checksum so the work
 * cannot be optimised away.  iq holds `frames_in_buf` distinct readings that are
 * cycled through. */
double sdr_ref_run_frames(void* h, const float* iq, int n, int frames_in_buf, int frames,
                          uint32_t fs, int do_fft, int do_ssb, int mode) {
    FFTProcessor* p = static_cast<FFTProcessor*>(h);
    double acc = 0.0;
    std::vector<int16_t> pcm;
    for (int f = 0; f < frames; f++) {
        const float* src = iq + static_cast<size_t>(f % frames_in_buf) * n * 2;
        g_now_ns += static_cast<int64_t>(1e9 * n / fs);
        if (do_ssb) {
            std::vector<std::complex<float>> v(reinterpret_cast<const std::complex<float>*>(src),
                                               reinterpret_cast<const std::complex<float>*>(src) + n);
            bool pulse = false;
            processSSB_opt(std::move(v), fs, true, pcm, pulse, mode);
            if (!pcm.empty()) acc += pcm[pcm.size() / 2];
        }
        if (do_fft) {
            p->process(reinterpret_cast<const std::complex<float>*>(src), n);
            acc += p->getMeanSnrDb() + p->getPowerSpectrum()[n / 2];
        }
    }
    return acc;
}

} /* extern "C" */

/* ------------------------------ pulse detectors ------------------------------ */
#include "dsp/spectral_pulse_detector.h"
#include "ssb/audio_pulse_detector.h"
#include <vector>

extern "C" {

/* SpectralPulseDetector as the bridge uses it: default-constructed global, then configure() with
 * fsEnergy = sampleRate / samplesPerReading (src/sdr-bridge-java-soapy.cpp:105,1131-1137). */
void* sdr_ref_spd_create(float fs_energy) {
    SpectralPulseDetector* d = new SpectralPulseDetector();
    SpectralPulseDetector::Config c;
    c.fsEnergy = fs_energy;
    d->configure(c);
    return d;
}
void sdr_ref_spd_destroy(void* h) { delete static_cast<SpectralPulseDetector*>(h); }
/* out: liveEtat, level, locked, lastSnr, tTarget, estimatedFreqHz */
void sdr_ref_spd_process(void* h, float snr_sigma, float freq_hz, float* out) {
    SpectralPulseDetector* d = static_cast<SpectralPulseDetector*>(h);
    const int level = static_cast<int>(d->process(snr_sigma, freq_hz));
    out[0] = (float)d->liveEtat(); out[1] = (float)level; out[2] = d->isLocked() ? 1.f : 0.f;
    out[3] = d->lastPulseStrength(); out[4] = d->lockedPeriodS(); out[5] = d->estimatedFreqHz();
}

/* AudioPulseDetector with its default Config (src/ssb/ssb_processor.h member pulseDetector_). */
void* sdr_ref_apd_create(void) { return new AudioPulseDetector(); }
void sdr_ref_apd_destroy(void* h) { delete static_cast<AudioPulseDetector*>(h); }
/* out: liveEtat, level, locked, lastSnr, tTarget */
void sdr_ref_apd_process(void* h, const int16_t* pcm, int n, float* out) {
    AudioPulseDetector* d = static_cast<AudioPulseDetector*>(h);
    std::vector<int16_t> v(pcm, pcm + n);
    const int level = static_cast<int>(d->process(v));
    out[0] = (float)d->liveEtat(); out[1] = (float)level; out[2] = d->isLocked() ? 1.f : 0.f;
    out[3] = d->lastPulseStrength(); out[4] = d->lockedPeriodS();
}

}  /* extern "C" */

/* ------------------------------ Airspy real->IQ front end ------------------------------ */
extern "C" {
#include "iqconverter_float.h"
}
#include "filters.h"

extern "C" {
void* sdr_ref_airspy_create(void) { return iqconverter_float_create(HB_KERNEL_FLOAT, HB_KERNEL_FLOAT_LEN); }
void sdr_ref_airspy_destroy(void* h) { iqconverter_float_free(static_cast<iqconverter_float_t*>(h)); }
/* convert_samples_float is static in airspy.c (which needs libusb): its one-line body is restated here
 * (third_party/libairspy/libairspy/src/airspy.c:311-321, SAMPLE_SCALE = 1/2048 from :58-62); the converter
 * itself is the reference's own iqconverter_float.c. */
void sdr_ref_airspy_process(void* h, const uint16_t* raw, int n_real, float* out) {
    for (int i = 0; i < n_real; i++) out[i] = (raw[i] - 2048) * (1.0f / (1 << 11));
    iqconverter_float_process(static_cast<iqconverter_float_t*>(h), out, n_real);
}
}
