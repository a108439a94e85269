// host/ssb_processor.h — header-compatible front end of the reference's SSBProcessor
// (src/ssb/ssb_processor.h:16-58, src/ssb/ssb_processor.cpp:22-115): a bounded queue (3 entries,
// drop-oldest, :58-60) feeding a worker thread that runs processSSB_opt(upper = true,
// mode = BridgeConfig sound mode) and hands the PCM to the callback (:99-107).
//
// The audio pulse detector that the reference runs after the callback (:109-113) is a
// SURVEY.md 8(f) "next" row; until it is built here the pulse callback overload is not offered
// (nothing is faked).  The sound mode comes from setSoundModeSource(); when this header is
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
#include <thread>
#include <vector>

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
    void stopProcessing();
    void enqueueData(std::vector<std::complex<float>>&& iq_data, uint32_t sample_rate);
    float getCurrentRatio() const { return 0.f; }          // src/ssb/ssb_processor.h:35

    // additive
    void setSoundModeSource(std::function<int()> src);       // default: mode 1 ("normal")
    uint64_t droppedBlocks() const { return dropped_.load(); }

private:
    void workerLoop();
    PcmDataCallback on_pcm_;
    std::function<int()> mode_src_;
    std::deque<SSB_Data> pending_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::thread worker_;
    std::atomic<bool> running_{false};
    std::atomic<uint64_t> dropped_{0};
};
