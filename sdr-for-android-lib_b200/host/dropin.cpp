// host/dropin.cpp — FFTProcessor / processSSB_opt / SSBProcessor over the C-ABI (batch of one).
// Only include/sdr_b200.h entry points are used: this file is what a maintainer of the reference
// would link instead of src/dsp/fft_process.cpp, src/ssb/ssb_demod_opt.cpp and
// src/ssb/ssb_processor.cpp.
#include <atomic>
#include <stdexcept>
#include <string>

#include "fft_process.h"
#include "ssb_demod_opt.h"
#include "ssb_processor.h"
#ifdef SDRB_USE_BRIDGE_CONFIG
#include "bridge-config.h"
#endif

// The bridge owns this flag (src/sdr-bridge-java-soapy.cpp:337-340, src/sdr-bridge-internal.h:9);
// a weak definition lets the library link on its own.
namespace sdr_bridge_internal {
__attribute__((weak)) std::atomic<bool> isCenterFrequencyChanged{false};
}

namespace {
[[noreturn]] void raise(const sdrb_ctx* c, const char* what, int rc) {
    throw std::runtime_error(std::string(what) + " failed (" + std::to_string(rc) + "): " + sdrb_last_error(c));
}
#define SDRB_TRY(ctx, call)                      \
    do {                                         \
        const int rc_ = (call);                  \
        if (rc_ != SDRB_OK) raise(ctx, #call, rc_); \
    } while (0)

sdrb_config make_config(uint32_t fc, uint32_t fs, int n, int focus, int mode, bool upper) {
    sdrb_config c = {};
    c.centerFrequency = fc;
    c.sampleRate = fs;
    c.samplesPerReading = n;
    c.freqFocusRangeKhz = focus;
    c.soundMode = mode;
    c.inputFormat = SDRB_CF32;            // what SoapySDR hands soapyCallback ("CF32")
    c.upperSideband = upper ? 1 : 0;
    return c;
}
}  // namespace

// ------------------------------------------------------------------ FFTProcessor
FFTProcessor::FFTProcessor() = default;

FFTProcessor::~FFTProcessor() {
    if (ctx_) sdrb_destroy(ctx_);
}

void FFTProcessor::configure(const FftProcessorConfig& config) {
    const bool relayout = !have_cfg_ || config.sampleRate != cfg_.sampleRate ||
                          config.samplesPerReading != cfg_.samplesPerReading;
    const bool retune = have_cfg_ && config.centerFrequency != cfg_.centerFrequency;
    const bool refocus = have_cfg_ && config.freqFocusRangeKhz != cfg_.freqFocusRangeKhz;
    cfg_ = config;
    have_cfg_ = true;
    if (!ctx_) return;                    // applied when the first block arrives
    if (relayout) { ctx_n_ = 0; return; } // new geometry: the next process() re-plans; carried state survives (:20-39)
    // setFrequency / setFrequencyFocusRange push a whole config (src/sdr-bridge-java-soapy.cpp:897-905,
    // 1031-1036).  The bridge raises isCenterFrequencyChanged together with every retune, so the
    // retune is forwarded as sdrb_set_center_frequency (new centre + tracking reset, :336-339).
    if (refocus) SDRB_TRY(ctx_, sdrb_set_focus_range(ctx_, cfg_.freqFocusRangeKhz));
    if (retune) SDRB_TRY(ctx_, sdrb_set_center_frequency(ctx_, -1, cfg_.centerFrequency));
}

void FFTProcessor::ensureContext(int n) {
    if (!ctx_) SDRB_TRY(nullptr, sdrb_create(device_, &ctx_));
    if (ctx_n_ == n) return;
    // the reference keeps tracking / detection memory across configure(); carry it over a re-plan
    std::vector<unsigned char> blob(ctx_n_ ? sdrb_state_bytes(ctx_) : 0);
    if (!blob.empty()) SDRB_TRY(ctx_, sdrb_save_state(ctx_, blob.data(), blob.size()));
    const sdrb_config c = make_config(cfg_.centerFrequency, cfg_.sampleRate, n, cfg_.freqFocusRangeKhz, 1, true);
    SDRB_TRY(ctx_, sdrb_configure(ctx_, &c, 1, 1));
    if (!blob.empty()) {
        SDRB_TRY(ctx_, sdrb_load_state(ctx_, blob.data(), blob.size()));
        SDRB_TRY(ctx_, sdrb_set_center_frequency(ctx_, -1, cfg_.centerFrequency));
    }
    ctx_n_ = n;
    spectrum_.assign(static_cast<size_t>(n), 0.0f);
}

// The reference's process() is void and never throws (bad geometry is a silent early return that leaves the previous
// outputs in place, src/dsp/fft_process.cpp:139); it runs on the bridge's processing thread, where an exception would
// end the process.  Failures (no device, out of memory, a CUDA error) are therefore recorded -- lastError(),
// failedCalls() -- and the previous outputs stay.
void FFTProcessor::process(const std::complex<float>* input_buf, uint32_t input_len) {
    if (!input_buf || input_len == 0) return;
    try {
        if (!have_cfg_) throw std::runtime_error("FFTProcessor::process before configure");
        ensureContext(static_cast<int>(input_len));
        if (sdr_bridge_internal::isCenterFrequencyChanged.exchange(false))
            SDRB_TRY(ctx_, sdrb_set_center_frequency(ctx_, -1, cfg_.centerFrequency));
        SDRB_TRY(ctx_, sdrb_process_host(ctx_, input_buf, 1, SDRB_STAGE_FFT | SDRB_STAGE_SPECTRAL_PULSE, &last_,
                                         spectrum_.data(), nullptr));
        SDRB_TRY(ctx_, sdrb_get_pulse_results(ctx_, &pulse_, 1));
    } catch (const std::exception& e) {
        last_error_ = e.what();
        failed_calls_++;
    }
}

// ------------------------------------------------------------------ processSSB_opt
namespace {
struct SsbSingleton {
    std::mutex mu;
    sdrb_ctx* ctx = nullptr;
    int device = 0;
    int n = 0;
    uint32_t fs = 0;
    int mode = -1;
    bool upper = true;
    ~SsbSingleton() { if (ctx) sdrb_destroy(ctx); }
};
SsbSingleton& ssb_one() {
    static SsbSingleton s;
    return s;
}
}  // namespace

void sdrbSsbSetDevice(int device) { ssb_one().device = device; }

void sdrbSsbReset() {
    SsbSingleton& S = ssb_one();
    std::lock_guard<std::mutex> lock(S.mu);
    if (S.ctx) { sdrb_destroy(S.ctx); S.ctx = nullptr; }
    S.n = 0; S.fs = 0; S.mode = -1;
}

bool lastBlockContainsPulse() { return false; }

namespace {
void ssb_run(std::vector<std::complex<float>>& iq, uint32_t sampleRate, bool upperSideband,
             std::vector<int16_t>& pcmOut, int mode, bool with_pulse, const sdrb_audio_pulse_config* new_cfg,
             sdrb_pulse_result* pulse_out);
}

void processSSB_opt(std::vector<std::complex<float>> iq, uint32_t sampleRate, bool upperSideband,
                    std::vector<int16_t>& pcmOut, bool& pulse, int mode) {
    (void)pulse;                                    // never written: src/ssb/ssb_demod_opt.cpp:221-296
    ssb_run(iq, sampleRate, upperSideband, pcmOut, mode, false, nullptr, nullptr);
}

void sdrbSsbProcessWithPulse(std::vector<std::complex<float>> iq, uint32_t sampleRate, bool upperSideband,
                             std::vector<int16_t>& pcmOut, int mode, const sdrb_audio_pulse_config* new_cfg,
                             sdrb_pulse_result* pulse_out) {
    ssb_run(iq, sampleRate, upperSideband, pcmOut, mode, true, new_cfg, pulse_out);
}

namespace {
void ssb_run(std::vector<std::complex<float>>& iq, uint32_t sampleRate, bool upperSideband,
             std::vector<int16_t>& pcmOut, int mode, bool with_pulse, const sdrb_audio_pulse_config* new_cfg,
             sdrb_pulse_result* pulse_out) {
    SsbSingleton& S = ssb_one();
    std::lock_guard<std::mutex> lock(S.mu);         // the reference is non-re-entrant (statics)
    pcmOut.clear();
    if (iq.empty() && S.n == 0) return;
    if (!S.ctx) SDRB_TRY(nullptr, sdrb_create(S.device, &S.ctx));
    // `static size_t sampCount = iq.size(); iq.resize(sampCount);` (src/ssb/ssb_demod_opt.cpp:225-227): the block
    // length of the FIRST call is frozen; later blocks are cut or zero-extended to it.  sdrbSsbReset() starts over.
    if (S.n != 0 && static_cast<int>(iq.size()) != S.n) iq.resize(static_cast<size_t>(S.n));
    const int n = static_cast<int>(iq.size());
    if (S.n == 0) {
        const sdrb_config c = make_config(0, sampleRate, n, 5, mode, upperSideband);
        SDRB_TRY(S.ctx, sdrb_configure(S.ctx, &c, 1, 1));
        S.n = n; S.fs = sampleRate; S.upper = upperSideband; S.mode = mode;
    } else if (sampleRate != S.fs || upperSideband != S.upper) {
        // a new sample rate (decimation factor) or sideband: re-plan with every carried state kept -- the three
        // filters are function statics in the reference and the pulse detector a member of SSBProcessor
        const sdrb_config c = make_config(0, sampleRate, n, 5, mode, upperSideband);
        SDRB_TRY(S.ctx, sdrb_replan(S.ctx, &c, 1));
        S.fs = sampleRate; S.upper = upperSideband; S.mode = mode;
    } else if (mode != S.mode) {
        SDRB_TRY(S.ctx, sdrb_set_sound_mode(S.ctx, mode));   // gains/AGC follow, RF low-pass keeps the first mode
        S.mode = mode;
    }
    const int m = sdrb_pcm_per_block(S.ctx);
    if (m < 0) raise(S.ctx, "sdrb_pcm_per_block", m);
    pcmOut.resize(static_cast<size_t>(m));
    if (new_cfg) SDRB_TRY(S.ctx, sdrb_set_pulse_config(S.ctx, new_cfg));
    const unsigned stages = SDRB_STAGE_SSB | (with_pulse ? SDRB_STAGE_AUDIO_PULSE : 0u);
    SDRB_TRY(S.ctx, sdrb_process_host(S.ctx, iq.data(), 1, stages, nullptr, nullptr, m ? pcmOut.data() : nullptr));
    if (with_pulse && pulse_out) SDRB_TRY(S.ctx, sdrb_get_pulse_results(S.ctx, pulse_out, 1));
}
}  // namespace

// ------------------------------------------------------------------ SSBProcessor
SSBProcessor::SSBProcessor() : mode_src_([] {
#ifdef SDRB_USE_BRIDGE_CONFIG
    return BridgeConfig::getInstance().getSoundMode();
#else
    return 1;
#endif
}) {}

SSBProcessor::~SSBProcessor() { stopProcessing(); }

void SSBProcessor::setSoundModeSource(std::function<int()> src) {
    std::lock_guard<std::mutex> lock(mu_);
    if (src) mode_src_ = std::move(src);
}

void SSBProcessor::startProcessing(PcmDataCallback pcm_cb) {
    if (running_.exchange(true)) return;            // already running: ignored like :28-31
    on_pcm_ = std::move(pcm_cb);
    on_pulse_ = nullptr;                            // pulseCallback_ = nullptr, :33
    worker_ = std::thread([this] { workerLoop(); });
}

void SSBProcessor::startProcessing(PcmDataCallback pcm_cb, std::function<void(float, int)> pulse_cb) {
    if (running_.exchange(true)) return;
    on_pcm_ = std::move(pcm_cb);
    on_pulse_ = std::move(pulse_cb);                // :36-49
    worker_ = std::thread([this] { workerLoop(); });
}

void SSBProcessor::setPulseConfig(const AudioPulseDetector::Config& c) {
    std::lock_guard<std::mutex> lock(mu_);           // applied by the worker before the next block, :69-74,90-96
    pending_cfg_ = sdrb_audio_pulse_config{c.sampleRate, c.fMin, c.fMax, c.fsEnergy, c.smoothCutoff, c.zDefaultS,
                                           c.tTargetInit, c.dtTolS, c.snrMin, c.snrRhythm, c.snrStrong,
                                           c.dispersionMax, c.sumNMax, c.liveWindowT, c.liveDivisor, c.noiseRefFar,
                                           c.noiseRefNear};
    pending_cfg_set_ = true;
}

void SSBProcessor::stopProcessing() {
    if (!running_.exchange(false)) return;
    cv_.notify_all();
    if (worker_.joinable()) worker_.join();
}

void SSBProcessor::enqueueData(std::vector<std::complex<float>>&& iq_data, uint32_t sample_rate) {
    if (!running_) return;                           // :52
    {
        std::lock_guard<std::mutex> lock(mu_);
        while (pending_.size() >= 3) {               // drop-oldest, :58-60
            pending_.pop_front();
            dropped_++;
        }
        pending_.push_back(SSB_Data{std::move(iq_data), sample_rate});
    }
    cv_.notify_one();
}

void SSBProcessor::workerLoop() {
    for (;;) {
        SSB_Data job;
        int mode;
        bool have_cfg = false;
        sdrb_audio_pulse_config new_cfg{};
        {
            std::unique_lock<std::mutex> lock(mu_);
            cv_.wait(lock, [this] { return !pending_.empty() || !running_; });
            if (pending_.empty()) {
                if (!running_) return;
                continue;
            }
            job = std::move(pending_.front());
            pending_.pop_front();
            mode = mode_src_();
            have_cfg = pending_cfg_set_;
            if (have_cfg) { new_cfg = pending_cfg_; pending_cfg_set_ = false; }
        }
        std::vector<int16_t> pcm;
        sdrb_pulse_result pr{};
        // processSSB_opt (:103) + pulseDetector_.process(pcm) (:109) in one GPU call.  The reference's worker cannot
        // fail; here a failed call (no memory, a CUDA error) drops the block and is recorded instead of ending the
        // process through an exception that nothing on this thread would catch.
        try {
            sdrbSsbProcessWithPulse(std::move(job.iq), job.sampleRate, true, pcm, mode, have_cfg ? &new_cfg : nullptr, &pr);
        } catch (const std::exception& e) {
            std::lock_guard<std::mutex> lock(mu_);
            last_error_ = e.what();
            failed_++;
            continue;
        }
        if (on_pcm_ && !pcm.empty()) on_pcm_(pcm);                                      // :105-107
        if (on_pulse_) on_pulse_(pr.audioLastSnr, pr.audioLiveEtat);                    // :110-113
    }
}
