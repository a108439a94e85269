// host/ssb_processor.h — header-compatible front end of the reference's SSBProcessor
// (src/ssb/ssb_processor.h:16-58, src/ssb/ssb_processor.cpp:22-115): a bounded queue (3 entries,
// drop-oldest, :58-60) feeding a worker thread that runs processSSB_opt(upper = true,
// mode = BridgeConfig sound mode) and hands the PCM to the callback (:99-107).
//
// The audio pulse detector that the reference runs after the callback (:109-113) runs on the GPU in
// the same call (SDRB_STAGE_AUDIO_PULSE); its (lastPulseStrength, liveEtat) pair goes to the pulse
// callback exactly like :110-113, and setPulseConfig rebuilds it like :69-74,90-96.  The sound mode comes from setSoundModeSource(); when this header is
// compiled inside the reference tree (SDRB_USE_BRIDGE_CONFIG defined) it defaults to
// BridgeConfig::getInstance().getSoundMode() exactly like src/ssb/ssb_processor.cpp:101.
#pragma once
#include <atomic>
#include <complex>
#include <condition_variable>
#include <cstdint>
#include <deque>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/sdr_b200.h"
#include "audio_pulse_detector.h"

struct SSB_Data {
    std::vector<std::complex<float>> iq;
    uint32_t sampleRate;
};

using PcmDataCallback = std::function<void(const std::vector<int16_t>&)>;

class SSBProcessor {
public:
    SSBProcessor();
    ~SSBProcessor();

    void startProcessing(PcmDataCallback pcm_cb);
    void startProcessing(PcmDataCallback pcm_cb, std::function<void(float, int)> pulse_cb);   // :23, :36-49
    void setPulseConfig(const AudioPulseDetector::Config& cfg);                               // :69-74
    float getAmbientEnergy() const { return 0.f; }                                            // src/ssb/ssb_processor.h:34
    void stopProcessing();
    void enqueueData(std::vector<std::complex<float>>&& iq_data, uint32_t sample_rate);
    float getCurrentRatio() const { return 0.f; }          // src/ssb/ssb_processor.h:35

    // additive
    void setSoundModeSource(std::function<int()> src);       // default: mode 1 ("normal")
    uint64_t droppedBlocks() const { return dropped_.load(); }
    uint64_t failedBlocks() const { return failed_.load(); }   // blocks whose GPU call failed (dropped, worker kept alive)
    std::string lastError() {
        std::lock_guard<std::mutex> lock(mu_);
        return last_error_;
    }

private:
    void workerLoop();
    PcmDataCallback on_pcm_;
    std::function<void(float, int)> on_pulse_;
    sdrb_audio_pulse_config pending_cfg_{};
    bool pending_cfg_set_ = false;
    std::function<int()> mode_src_;
    std::deque<SSB_Data> pending_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::thread worker_;
    std::atomic<bool> running_{false};
    std::atomic<uint64_t> dropped_{0};
    std::atomic<uint64_t> failed_{0};
    std::string last_error_;
};
