// host/fft_process.h — header-compatible front end of the reference's FFTProcessor
// (src/dsp/fft_process.h:20-55) over the C-ABI of include/sdr_b200.h.
//
// Same type names, same public member functions, same getter semantics (values stay valid
// until the next process() call), so src/sdr-bridge-java-soapy.cpp:101-108,441-455 compiles
// against this header unchanged.  One FFTProcessor = one GPU context with one stream and one
// reading per call (the batch-of-one case of sdrb_process); servers that hold many streams
// should call the batched C-ABI directly.
//
// Differences that are deliberate and visible:
//  * there is no CPU path: construction is lazy, the first process() creates the CUDA context
//    and throws std::runtime_error when no device / an unsupported length is given (the
//    reference has no error channel at all: SURVEY.md 8b);
//  * the tracking gate runs on the frame clock t = (frame+1)*N/fs instead of steady_clock
//    (src/dsp/fft_process.cpp:349-361), like the oracle;
//  * input_len is the FFT length, as in the reference (:47); it must be a power of two.
#pragma once
#include <complex>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/sdr_b200.h"

struct FftProcessOutput {            // src/dsp/fft_process.h:12-18 (unused by the reference itself)
    std::vector<float> powerSpectrum;
    float meanSnrDb;
    float meanSnrSigma;
    long trackingFrequency;
    int detectionFlag;
};

struct FftProcessorConfig {          // src/dsp/fft_process.h:20-26
    uint32_t centerFrequency;
    uint32_t sampleRate;
    int samplesPerReading;
    int freqFocusRangeKhz;
};

class FFTProcessor {
public:
    FFTProcessor();
    ~FFTProcessor();
    FFTProcessor(const FFTProcessor&) = delete;
    FFTProcessor& operator=(const FFTProcessor&) = delete;

    void configure(const FftProcessorConfig& config);                          // :20-39
    void process(const std::complex<float>* input_buf, uint32_t input_len);    // :42-105

    const std::vector<float>& getPowerSpectrum() const { return spectrum_; }
    float getMeanSnrDb() const { return last_.meanSnrDb; }
    float getMeanSnrSigma() const { return last_.meanSnrSigma; }
    long getTrackingFrequency() const { return static_cast<long>(last_.trackingFrequency); }
    int getDetectionFlag() const { return last_.detectionFlag; }
    float getPeakAboveNoiseMeanDb() const { return last_.peakAboveNoiseMeanDb; }
    float getMaxBinSnrDb() const { return last_.maxBinSnrDb; }
    float getMaxBinSnrSigma() const { return last_.maxBinSnrSigma; }
    float getBest1kHzSnrDb() const { return last_.best1kHzSnrDb; }
    float getBest1kHzSnrSigma() const { return last_.best1kHzSnrSigma; }
    float getBest1kHzCenterFreqHz() const { return last_.best1kHzCenterFreqHz; }
    float getPerBinMean() const { return last_.perBinMean; }

    // additive: the whole record of the last frame (frame-local internals included)
    const sdrb_frame_result& lastRecord() const { return last_; }
    // additive: SpectralPulseDetector state after the last frame.  The bridge's
    //   spectralPulseDetector.process(best1kHzSnrSigma, best1kFreqHz); liveEtat(); estimatedFreqHz()
    // (src/sdr-bridge-java-soapy.cpp:477-487) becomes a read of this record: the detector runs on the GPU
    // in the same call (SDRB_STAGE_SPECTRAL_PULSE).
    const sdrb_pulse_result& lastPulse() const { return pulse_; }
    // additive: CUDA device of the context created by the first process() (default 0)
    void setDevice(int device) { device_ = device; }
    // process() never throws (the reference's is void and runs on the bridge's processing thread): a failed call
    // keeps the previous outputs and is recorded here
    const std::string& lastError() const { return last_error_; }
    unsigned long failedCalls() const { return failed_calls_; }

private:
    void ensureContext(int n);
    FftProcessorConfig cfg_{};
    bool have_cfg_ = false;
    int device_ = 0;
    int ctx_n_ = 0;                  // FFT length the context is configured for
    sdrb_ctx* ctx_ = nullptr;
    sdrb_frame_result last_{};
    sdrb_pulse_result pulse_{};
    std::string last_error_;
    unsigned long failed_calls_ = 0;
    std::vector<float> spectrum_;
};
