// host/iq_chunker.h — bridge-side chunking and back-pressure, header-only C++.
//
// What the reference's reader thread does between SoapySDR::readStream and soapyCallback
// (src/sdr-bridge-java-soapy.cpp:496-575): reads of at most min(MTU, samplesPerReading) samples are
// appended to an accumulator (:541-546), cut into EXACT samplesPerReading blocks (:549-573) and pushed
// into a queue of at most RX_QUEUE_MAX = 20 blocks that drops the OLDEST block when full (:556-563, :121);
// the processing thread waits up to 60 ms for a block (:595-605).  Here the sample type is a template
// parameter, so raw wire bytes (rtl_tcp-style u8 pairs, cs16, Airspy u16 words) are chunked without the
// float conversion the reference's Soapy layer does on the CPU: the conversion is fused into the GPU kernels.
//
//   IqChunker<T>      one stream: push(ptr, count) / pop(block, timeout)
//   IqFileSource<T>   replay of a raw IQ capture file in MTU-sized reads (stands in for readStream; the box
//                     has no dongle), optional looping
//   StreamBatcher<T>  S chunkers -> one [S][B][N] batch for sdrb_process_host once every stream holds B blocks
#pragma once
#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <deque>
#include <mutex>
#include <string>
#include <vector>

namespace sdrb_host {

constexpr size_t kRxQueueMax = 20;   // RX_QUEUE_MAX, src/sdr-bridge-java-soapy.cpp:121

// T = one complex sample of the wire format (e.g. uint16_t for u8 IQ, uint32_t for cs16, a float pair).
template <class T>
class IqChunker {
public:
    explicit IqChunker(size_t samples_per_reading, size_t queue_max = kRxQueueMax)
        : n_(samples_per_reading), queue_max_(queue_max) {}

    // BridgeConfig::setSamplesPerReading while streaming: the accumulator keeps its samples, later blocks
    // are cut at the new length (the reference re-reads the size every loop, :508-509).
    void setSamplesPerReading(size_t n) {
        std::lock_guard<std::mutex> lock(mu_);
        n_ = n;
    }
    size_t samplesPerReading() const { return n_; }

    // One readStream result (ret > 0 samples).  Returns the number of whole blocks queued by this call.
    size_t push(const T* samples, size_t count) {
        size_t made = 0;
        {
            std::lock_guard<std::mutex> lock(mu_);
            acc_.insert(acc_.end(), samples, samples + count);                       // :541-546
            while (n_ > 0 && acc_.size() >= n_) {                                    // :549
                if (queue_.size() >= queue_max_) {                                   // :559-560 drop-oldest
                    queue_.pop_front();
                    dropped_++;
                }
                queue_.emplace_back(acc_.begin(), acc_.begin() + (std::ptrdiff_t)n_);  // :551-554, :562
                acc_.erase(acc_.begin(), acc_.begin() + (std::ptrdiff_t)n_);           // :566-569
                made++;
            }
        }
        if (made) cv_.notify_all();                                                  // :571
        return made;
    }

    // The processing thread's wait (:595-605): true and one block, or false after timeout_ms.
    bool pop(std::vector<T>& block, int timeout_ms = 60) {
        std::unique_lock<std::mutex> lock(mu_);
        cv_.wait_for(lock, std::chrono::milliseconds(timeout_ms), [this] { return !queue_.empty() || stopped_; });
        if (queue_.empty()) return false;
        block = std::move(queue_.front());
        queue_.pop_front();
        return true;
    }

    void stop() {
        { std::lock_guard<std::mutex> lock(mu_); stopped_ = true; }
        cv_.notify_all();
    }
    size_t queued() const { std::lock_guard<std::mutex> lock(mu_); return queue_.size(); }
    size_t pendingSamples() const { std::lock_guard<std::mutex> lock(mu_); return acc_.size(); }
    uint64_t droppedBlocks() const { std::lock_guard<std::mutex> lock(mu_); return dropped_; }

private:
    size_t n_, queue_max_;
    std::vector<T> acc_;
    std::deque<std::vector<T>> queue_;
    mutable std::mutex mu_;
    std::condition_variable cv_;
    uint64_t dropped_ = 0;
    bool stopped_ = false;
};

// Raw capture replay: read() hands out at most min(mtu, samples_per_reading) samples per call, like the
// reader thread's readStream (:510-539).  Returns 0 at end of file unless looping.
template <class T>
class IqFileSource {
public:
    IqFileSource(const std::string& path, size_t mtu, bool loop = false) : mtu_(mtu), loop_(loop) {
        f_ = std::fopen(path.c_str(), "rb");
    }
    ~IqFileSource() { if (f_) std::fclose(f_); }
    IqFileSource(const IqFileSource&) = delete;
    IqFileSource& operator=(const IqFileSource&) = delete;
    bool ok() const { return f_ != nullptr; }

    size_t read(T* out, size_t samples_per_reading) {
        if (!f_) return 0;
        const size_t want = std::min(mtu_, samples_per_reading);   // :510
        size_t got = std::fread(out, sizeof(T), want, f_);
        if (got == 0 && loop_) {
            std::rewind(f_);
            got = std::fread(out, sizeof(T), want, f_);
        }
        return got;
    }

private:
    std::FILE* f_ = nullptr;
    size_t mtu_;
    bool loop_;
};

// S independent streams feeding one batched GPU call: tryAssemble() fills `batch` with [S][B][N] samples and
// returns true once EVERY stream has B blocks queued (streams that run ahead keep queueing, and drop their
// oldest blocks beyond the queue bound exactly like a single stream does).
template <class T>
class StreamBatcher {
public:
    StreamBatcher(size_t n_streams, size_t samples_per_reading, size_t blocks_per_batch)
        : n_(samples_per_reading), b_(blocks_per_batch) {
        for (size_t s = 0; s < n_streams; s++) ch_.emplace_back(new IqChunker<T>(samples_per_reading,
                                                                                 std::max(kRxQueueMax, blocks_per_batch)));
    }
    ~StreamBatcher() { for (auto* c : ch_) delete c; }
    StreamBatcher(const StreamBatcher&) = delete;
    StreamBatcher& operator=(const StreamBatcher&) = delete;

    IqChunker<T>& stream(size_t s) { return *ch_[s]; }
    size_t streams() const { return ch_.size(); }

    bool tryAssemble(std::vector<T>& batch) {
        for (auto* c : ch_) if (c->queued() < b_) return false;
        batch.resize(ch_.size() * b_ * n_);
        std::vector<T> block;
        for (size_t s = 0; s < ch_.size(); s++)
            for (size_t b = 0; b < b_; b++) {
                if (!ch_[s]->pop(block, 0) || block.size() != n_) return false;   // only with a concurrent consumer
                std::copy(block.begin(), block.end(), batch.begin() + (std::ptrdiff_t)((s * b_ + b) * n_));
            }
        return true;
    }

private:
    size_t n_, b_;
    std::vector<IqChunker<T>*> ch_;
};

}  // namespace sdrb_host
