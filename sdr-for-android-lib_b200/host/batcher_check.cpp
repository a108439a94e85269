// host/batcher_check.cpp — GPU check of the bridge-side chunking in front of the batched C-ABI call
// (SURVEY.md 8(f) row 4): every stream's raw u8 IQ arrives in ragged reads (sizes unrelated to the block length,
// like readStream results, src/sdr-bridge-java-soapy.cpp:510-546), goes through its IqChunker, and as soon as EVERY
// stream holds B blocks the StreamBatcher's [S][B][N] batch is handed to sdrb_process_host.  The outputs of each
// batch are written to a file; tests/test_gpu_dropin.py compares them with a direct sdrb_process call on the same
// samples.
//   batcher_check <out.bin> <streams> <blocks_per_batch> <n> <batches> <fs>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../include/sdr_b200.h"
#include "iq_chunker.h"

#define TRY(call)                                                                         \
    do {                                                                                  \
        const int rc_ = (call);                                                           \
        if (rc_ != SDRB_OK) { fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, sdrb_last_error(ctx)); return 2; } \
    } while (0)

int main(int argc, char** argv) {
    if (argc < 7) { fprintf(stderr, "usage: batcher_check out streams blocks n batches fs\n"); return 1; }
    const int S = atoi(argv[2]), B = atoi(argv[3]), n = atoi(argv[4]), batches = atoi(argv[5]);
    const uint32_t fs = (uint32_t)atol(argv[6]);
    sdrb_ctx* ctx = nullptr;
    if (sdrb_create(0, &ctx) != SDRB_OK) { fprintf(stderr, "sdrb_create: %s\n", sdrb_last_error(nullptr)); return 2; }
    sdrb_config cfg = {};
    cfg.centerFrequency = 430000000u; cfg.sampleRate = fs; cfg.samplesPerReading = n; cfg.freqFocusRangeKhz = 5;
    cfg.soundMode = 1; cfg.inputFormat = SDRB_CU8; cfg.upperSideband = 1;
    TRY(sdrb_configure(ctx, &cfg, S, B));
    const int m = sdrb_pcm_per_block(ctx);
    // the "capture": each stream's whole run of samples from the deterministic source
    const size_t total = (size_t)batches * B * n + 333;   // a tail that never completes a block
    std::vector<std::vector<uint16_t>> cap((size_t)S, std::vector<uint16_t>(total));
    for (int s = 0; s < S; s++) TRY(sdrb_synth_host(SDRB_CU8, fs, s, 1, 0, (int64_t)total, cap[s].data()));
    sdrb_host::StreamBatcher<uint16_t> bat((size_t)S, (size_t)n, (size_t)B);
    std::vector<size_t> pos((size_t)S, 0);
    std::vector<uint16_t> batch;
    std::vector<sdrb_frame_result> rec((size_t)S * B);
    std::vector<float> power((size_t)S * B * n);
    std::vector<int16_t> pcm((size_t)S * B * (m > 0 ? m : 1));
    FILE* f = fopen(argv[1], "wb");
    if (!f) return 1;
    const int hdr[4] = {S, B, n, m};
    fwrite(hdr, sizeof hdr, 1, f);
    int done = 0;
    unsigned lcg = 12345u;
    while (done < batches) {
        // one ragged read per stream per round; streams advance at different speeds
        bool any = false;
        for (int s = 0; s < S; s++) {
            lcg = lcg * 1664525u + 1013904223u;
            size_t r = 1 + (lcg >> 8) % (size_t)(3 * n / 2 + 7 * s);
            if (pos[s] + r > total) r = total - pos[s];
            if (!r) continue;
            bat.stream((size_t)s).push(cap[s].data() + pos[s], r);
            pos[s] += r;
            any = true;
        }
        while (done < batches && bat.tryAssemble(batch)) {
            TRY(sdrb_process_host(ctx, batch.data(), B, SDRB_STAGE_FFT | SDRB_STAGE_SSB, rec.data(), power.data(),
                                  m > 0 ? pcm.data() : nullptr));
            fwrite(rec.data(), sizeof(sdrb_frame_result), rec.size(), f);
            fwrite(power.data(), sizeof(float), power.size(), f);
            if (m > 0) fwrite(pcm.data(), sizeof(int16_t), (size_t)S * B * m, f);
            done++;
        }
        if (!any) break;
    }
    fclose(f);
    sdrb_destroy(ctx);
    if (done != batches) { fprintf(stderr, "only %d of %d batches assembled\n", done, batches); return 3; }
    printf("batcher OK: %d batches of %d x %d x %d\n", done, S, B, n);
    return 0;
}
