// tests/cpp/chunker_test.cpp — CPU check of host/iq_chunker.h against the behaviour of the reference's
// reader thread (src/sdr-bridge-java-soapy.cpp:496-575): exact block cutting from ragged reads, ordering,
// drop-oldest at 20 queued blocks, 60 ms consumer wait, file replay in MTU-sized reads, batch assembly.
#include <cstdio>
#include <cstdlib>
#include <thread>

#include "iq_chunker.h"

using namespace sdrb_host;

#define CHECK(c) do { if (!(c)) { printf("FAIL line %d: %s\n", __LINE__, #c); return 1; } } while (0)

int main(int argc, char** argv) {
    const size_t N = 1000;
    {   // ragged reads -> exact blocks, in order
        IqChunker<uint16_t> ch(N);
        std::vector<uint16_t> src(10 * N + 123);
        for (size_t i = 0; i < src.size(); i++) src[i] = (uint16_t)(i * 7 + 1);
        size_t pos = 0, made = 0, k = 0;
        const size_t reads[] = {1, 999, 1, 1500, 37, 4096, 2};
        while (pos < src.size()) {
            size_t r = std::min(reads[k++ % 7], src.size() - pos);
            made += ch.push(src.data() + pos, r);
            pos += r;
        }
        CHECK(made == 10 && ch.queued() == 10 && ch.pendingSamples() == 123 && ch.droppedBlocks() == 0);
        std::vector<uint16_t> blk;
        for (size_t b = 0; b < 10; b++) {
            CHECK(ch.pop(blk, 0) && blk.size() == N);
            for (size_t i = 0; i < N; i++) CHECK(blk[i] == src[b * N + i]);
        }
        auto t0 = std::chrono::steady_clock::now();
        CHECK(!ch.pop(blk, 60));   // empty: waits about 60 ms like rx_cv.wait_for
        auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
        CHECK(ms >= 50 && ms < 1000);
    }
    {   // drop-oldest beyond RX_QUEUE_MAX = 20
        IqChunker<uint32_t> ch(N);
        std::vector<uint32_t> src(25 * N);
        for (size_t i = 0; i < src.size(); i++) src[i] = (uint32_t)(i / N);   // block id
        ch.push(src.data(), src.size());
        CHECK(ch.queued() == 20 && ch.droppedBlocks() == 5);
        std::vector<uint32_t> blk;
        CHECK(ch.pop(blk, 0) && blk[0] == 5 && blk[N - 1] == 5);   // blocks 0..4 were dropped
    }
    {   // producer / consumer threads
        IqChunker<uint16_t> ch(N);
        std::thread prod([&] {
            std::vector<uint16_t> buf(333);
            for (int i = 0; i < 300; i++) { for (auto& v : buf) v = (uint16_t)i; ch.push(buf.data(), buf.size()); }
            ch.stop();
        });
        size_t got = 0;
        std::vector<uint16_t> blk;
        while (got + ch.droppedBlocks() < 99) if (ch.pop(blk, 60)) got++;
        prod.join();
        CHECK(got + ch.droppedBlocks() == 99);
    }
    {   // file replay + batch assembly
        const char* path = argc > 1 ? argv[1] : "/tmp/chunker_test.iq";
        const size_t S = 3, B = 2, n = 256;
        std::vector<uint16_t> cap(S * B * n + 17);
        for (size_t i = 0; i < cap.size(); i++) cap[i] = (uint16_t)(i ^ 0x5a5a);
        FILE* f = fopen(path, "wb");
        CHECK(f && fwrite(cap.data(), 2, cap.size(), f) == cap.size());
        fclose(f);
        StreamBatcher<uint16_t> bat(S, n, B);
        std::vector<uint16_t> batch, rd(n);
        CHECK(!bat.tryAssemble(batch));
        for (size_t s = 0; s < S; s++) {
            IqFileSource<uint16_t> srcf(path, 100);   // MTU 100 < n: several reads per block
            CHECK(srcf.ok());
            size_t fed = 0, got;
            while (fed < (s + 1) * B * n && (got = srcf.read(rd.data(), n)) > 0) {
                CHECK(got <= 100);
                const size_t lo = s * B * n, hi = (s + 1) * B * n;   // stream s replays its own slice of the capture
                for (size_t i = 0; i < got; i++)
                    if (fed + i >= lo && fed + i < hi) bat.stream(s).push(&rd[i], 1);
                fed += got;
            }
        }
        CHECK(bat.tryAssemble(batch) && batch.size() == S * B * n);
        for (size_t i = 0; i < batch.size(); i++) CHECK(batch[i] == cap[i]);
        CHECK(!bat.tryAssemble(batch));
        remove(path);
    }
    {   // rtl_tcp client against a loopback server speaking the wire format of third_party/rtl-sdr/src/rtl_tcp.c
        const int ls = socket(AF_INET, SOCK_STREAM, 0);
        CHECK(ls >= 0);
        sockaddr_in a{};
        a.sin_family = AF_INET;
        a.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
        a.sin_port = 0;
        CHECK(bind(ls, (sockaddr*)&a, sizeof a) == 0 && listen(ls, 1) == 0);
        socklen_t al = sizeof a;
        CHECK(getsockname(ls, (sockaddr*)&a, &al) == 0);
        const uint16_t port = ntohs(a.sin_port);
        const size_t n = 512, blocks = 5;
        std::vector<unsigned char> wire(2 * n * blocks + 2 * 77);
        for (size_t i = 0; i < wire.size(); i++) wire[i] = (unsigned char)(i * 13 + 5);
        unsigned char cmd[10] = {0};
        std::thread server([&] {
            const int c = accept(ls, nullptr, nullptr);
            unsigned char hello[12] = {'R', 'T', 'L', '0', 0, 0, 0, 5, 0, 0, 0, 29};   // R820T, 29 gains
            send(c, hello, sizeof hello, 0);
            size_t got = 0;
            while (got < sizeof cmd) { ssize_t r = recv(c, cmd + got, sizeof cmd - got, 0); if (r <= 0) break; got += (size_t)r; }
            const size_t seg[] = {1, 3, 1001, 2, 4095, 7};   // odd segment sizes: pairs straddle segments
            size_t pos = 0, k = 0;
            while (pos < wire.size()) {
                const size_t w = std::min(seg[k++ % 6], wire.size() - pos);
                send(c, wire.data() + pos, w, 0);
                pos += w;
            }
            close(c);
        });
        IqTcpSource src("127.0.0.1", port, 300);
        CHECK(src.ok() && src.tunerType() == 5 && src.tunerGainCount() == 29);
        CHECK(src.setCenterFrequency(430000000u) && src.setSampleRate(2000000u));
        IqChunker<uint16_t> ch(n);
        std::vector<uint16_t> rd(n);
        size_t got, total = 0;
        while ((got = src.read(rd.data(), n)) > 0) { CHECK(got <= 300); ch.push(rd.data(), got); total += got; }
        server.join();
        close(ls);
        CHECK(total == wire.size() / 2 && ch.queued() == blocks && ch.pendingSamples() == 77);
        CHECK(cmd[0] == 0x01 && cmd[1] == 0x19 && cmd[2] == 0xA1 && cmd[3] == 0x47 && cmd[4] == 0x80);   // 430 000 000 big-endian
        CHECK(cmd[5] == 0x02 && cmd[6] == 0x00 && cmd[7] == 0x1E && cmd[8] == 0x84 && cmd[9] == 0x80);   // 2 000 000
        std::vector<uint16_t> blk;
        for (size_t b = 0; b < blocks; b++) {
            CHECK(ch.pop(blk, 0));
            CHECK(std::memcmp(blk.data(), wire.data() + 2 * n * b, 2 * n) == 0);   // I at the low byte, Q at the high byte
        }
    }
    printf("chunker OK\n");
    return 0;
}
