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
//   IqTcpSource       rtl_tcp client (third_party/rtl-sdr/src/rtl_tcp.c: 12-byte "RTL0" greeting, then raw u8
//                     I,Q pairs; 5-byte big-endian commands back): a live network source in place of readStream
//   StreamBatcher<T>  S chunkers -> one [S][B][N] batch for sdrb_process_host once every stream holds B blocks
#pragma once
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <cstring>
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

// rtl_tcp client.  Wire format of the server in the reference tree (third_party/rtl-sdr/src/rtl_tcp.c): on connect
// it sends dongle_info_t = { char magic[4] = "RTL0"; uint32 tuner_type; uint32 tuner_gain_count; } in network byte
// order (:79-83, :618-629), then the dongle's raw samples, unsigned 8-bit I,Q pairs, as they arrive; the client
// may send struct command { uint8 cmd; uint32 param; } packed, param big-endian (:281-284): 0x01 centre frequency,
// 0x02 sample rate, 0x03 gain mode, 0x04 gain (:316-331).  read() has the shape of IqFileSource::read, so the same
// loop feeds an IqChunker<uint16_t> (one uint16_t = one u8 I,Q pair, the SDRB_CU8 sample).
class IqTcpSource {
public:
    IqTcpSource(const std::string& host_ipv4, uint16_t port, size_t mtu) : mtu_(mtu) {
        fd_ = ::socket(AF_INET, SOCK_STREAM, 0);
        if (fd_ < 0) return;
        sockaddr_in a{};
        a.sin_family = AF_INET;
        a.sin_port = htons(port);
        if (::inet_pton(AF_INET, host_ipv4.c_str(), &a.sin_addr) != 1 ||
            ::connect(fd_, reinterpret_cast<sockaddr*>(&a), sizeof a) != 0) { close(); return; }
        int one = 1;
        ::setsockopt(fd_, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
        unsigned char hello[12];
        if (!readAll(hello, sizeof hello) || std::memcmp(hello, "RTL0", 4) != 0) { close(); return; }
        uint32_t be;
        std::memcpy(&be, hello + 4, 4); tuner_type_ = ntohl(be);
        std::memcpy(&be, hello + 8, 4); gain_count_ = ntohl(be);
    }
    ~IqTcpSource() { close(); }
    IqTcpSource(const IqTcpSource&) = delete;
    IqTcpSource& operator=(const IqTcpSource&) = delete;
    bool ok() const { return fd_ >= 0; }
    uint32_t tunerType() const { return tuner_type_; }
    uint32_t tunerGainCount() const { return gain_count_; }

    bool setCenterFrequency(uint32_t hz) { return command(0x01, hz); }
    bool setSampleRate(uint32_t hz) { return command(0x02, hz); }
    bool setGainMode(uint32_t manual) { return command(0x03, manual); }
    bool setGain(uint32_t tenth_db) { return command(0x04, tenth_db); }

    // At most min(mtu, samples_per_reading) whole I,Q pairs per call, whatever the segment boundaries were (an odd
    // byte is kept for the next call).  0 = the server closed the connection.
    size_t read(uint16_t* out, size_t samples_per_reading) {
        if (fd_ < 0) return 0;
        const size_t want = std::min(mtu_, samples_per_reading) * 2;   // bytes
        unsigned char* dst = reinterpret_cast<unsigned char*>(out);
        size_t have = 0;
        if (has_odd_) { dst[0] = odd_; have = 1; has_odd_ = false; }
        while (have < 2) {                                // block until one whole pair is there
            const ssize_t r = ::recv(fd_, dst + have, want - have, 0);
            if (r <= 0) return 0;
            have += static_cast<size_t>(r);
        }
        if (have & 1) { odd_ = dst[have - 1]; has_odd_ = true; have--; }
        return have / 2;
    }

private:
    bool readAll(unsigned char* p, size_t n) {
        while (n) {
            const ssize_t r = ::recv(fd_, p, n, 0);
            if (r <= 0) return false;
            p += r; n -= static_cast<size_t>(r);
        }
        return true;
    }
    bool command(unsigned char cmd, uint32_t param) {
        if (fd_ < 0) return false;
        unsigned char b[5];
        const uint32_t be = htonl(param);
        b[0] = cmd;
        std::memcpy(b + 1, &be, 4);
        return ::send(fd_, b, sizeof b, MSG_NOSIGNAL) == (ssize_t)sizeof b;
    }
    void close() { if (fd_ >= 0) { ::close(fd_); fd_ = -1; } }
    int fd_ = -1;
    size_t mtu_;
    uint32_t tuner_type_ = 0, gain_count_ = 0;
    unsigned char odd_ = 0;
    bool has_odd_ = false;
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
