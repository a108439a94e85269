// host/dropin_check.cpp — drives the header-compatible front end exactly the way soapyCallback
// drives the reference objects (src/sdr-bridge-java-soapy.cpp:424-493): one configure, then per
// reading fftProcessor.process + the getters, and ssbProcessor.enqueueData -> PCM callback.
// Writes everything it read back to a flat binary file that tests/test_gpu_dropin.py compares
// with the oracle.  Usage: dropin_check.bin <out.bin> <n> <fs> <blocks> <focus_khz> <mode>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

#include "fft_process.h"
#include "ssb_demod_opt.h"
#include "ssb_processor.h"

namespace sdr_bridge_internal { extern std::atomic<bool> isCenterFrequencyChanged; }

static void put(FILE* f, const void* p, size_t bytes) { if (fwrite(p, 1, bytes, f) != bytes) { perror("fwrite"); exit(3); } }

int main(int argc, char** argv) {
    if (argc < 7) { fprintf(stderr, "usage: %s out.bin n fs blocks focus_khz mode\n", argv[0]); return 2; }
    const int n = atoi(argv[2]);
    const uint32_t fs = (uint32_t)atol(argv[3]);
    const int B = atoi(argv[4]), focus = atoi(argv[5]), mode = atoi(argv[6]);
    const uint32_t fc = 430000000u;
    std::vector<float> raw((size_t)B * n * 2);
    if (sdrb_synth_host(SDRB_CF32, fs, 3, 1, 0, (int64_t)B * n, raw.data()) != SDRB_OK) return 4;
    const auto* iq = reinterpret_cast<const std::complex<float>*>(raw.data());

    FILE* f = fopen(argv[1], "wb");
    if (!f) { perror(argv[1]); return 3; }
    try {
        FFTProcessor fft;
        fft.configure(FftProcessorConfig{fc, fs, n, focus});
        // pass 1: direct processSSB_opt per block (fresh statics)
        std::vector<std::vector<int16_t>> direct(B);
        bool pulse = false;
        for (int b = 0; b < B; b++)
            processSSB_opt(std::vector<std::complex<float>>(iq + (size_t)b * n, iq + (size_t)(b + 1) * n), fs, true,
                           direct[b], pulse, mode);
        const int32_t m = (int32_t)direct[0].size();
        const int32_t hdr[4] = {n, B, m, (int32_t)pulse};
        put(f, hdr, sizeof hdr);
        for (int b = 0; b < B; b++) {
            if (b == B / 2) {   // a retune half way, the way setFrequency does it (:897-905)
                fft.configure(FftProcessorConfig{fc + 25000u, fs, n, focus});
                sdr_bridge_internal::isCenterFrequencyChanged = true;
            }
            fft.process(iq + (size_t)b * n, (uint32_t)n);
            const float g[9] = {fft.getMeanSnrDb(), fft.getMeanSnrSigma(), fft.getPeakAboveNoiseMeanDb(),
                                fft.getMaxBinSnrDb(), fft.getMaxBinSnrSigma(), fft.getBest1kHzSnrDb(),
                                fft.getBest1kHzSnrSigma(), fft.getBest1kHzCenterFreqHz(), fft.getPerBinMean()};
            const int64_t tr = fft.getTrackingFrequency();
            const int32_t flag = fft.getDetectionFlag(), peak = fft.lastRecord().peakBin;
            put(f, g, sizeof g);
            put(f, &tr, sizeof tr);
            put(f, &flag, sizeof flag);
            put(f, &peak, sizeof peak);
            put(f, fft.getPowerSpectrum().data(), (size_t)n * sizeof(float));
            put(f, direct[b].data(), (size_t)m * sizeof(int16_t));
        }
        // pass 2: the queued SSBProcessor from fresh filter state must deliver the same PCM, in order
        sdrbSsbReset();
        SSBProcessor ssb;
        ssb.setSoundModeSource([mode] { return mode; });
        std::vector<std::vector<int16_t>> queued;
        std::mutex qm;
        int pulse_calls = 0, pulse_bad = 0;
        ssb.setPulseConfig(AudioPulseDetector::Config{});
        ssb.startProcessing([&](const std::vector<int16_t>& pcm) { std::lock_guard<std::mutex> l(qm); queued.push_back(pcm); },
                            [&](float strength, int etat) {   // (FI)V payload, src/ssb/ssb_processor.cpp:110-113
                                pulse_calls++;
                                if (!(strength >= 0.f) || etat < 0 || etat > 5) pulse_bad++;
                            });
        for (int b = 0; b < B; b++) {
            ssb.enqueueData(std::vector<std::complex<float>>(iq + (size_t)b * n, iq + (size_t)(b + 1) * n), fs);
            for (int spin = 0; spin < 20000; spin++) {   // pace like a live source: never overflow the 3-deep queue
                { std::lock_guard<std::mutex> l(qm); if ((int)queued.size() == b + 1) break; }
                std::this_thread::sleep_for(std::chrono::microseconds(500));
            }
        }
        ssb.stopProcessing();
        int32_t same = (int)queued.size() == B && ssb.droppedBlocks() == 0 && pulse_calls == B && pulse_bad == 0;
        // the spectral detector saw every frame's best-1 kHz sigma
        same = same && fft.lastPulse().spectralSnrSigma == fft.getBest1kHzSnrSigma() && fft.lastPulse().spectralLiveEtat >= 0;
        for (int b = 0; same && b < B; b++) same = queued[b] == direct[b];
        if (fft.failedCalls() || ssb.failedBlocks()) {   // the front end records failures instead of throwing
            fprintf(stderr, "dropin_check: %lu failed FFT calls (%s), %lu failed SSB blocks (%s)\n", fft.failedCalls(),
                    fft.lastError().c_str(), (unsigned long)ssb.failedBlocks(), ssb.lastError().c_str());
            same = 0;
        }
        put(f, &same, sizeof same);
        // pass 3: `static size_t sampCount = iq.size(); iq.resize(sampCount);` (src/ssb/ssb_demod_opt.cpp:225-227): the
        // first call's block length is frozen; a shorter block is zero-extended, a longer one cut.  Blocks 0, 1, 2 with
        // n, n / 2 and (when there is one more block to take from) 2 n samples, and a sound mode outside 0..2 on the
        // last call (the file statics then keep the previous mode's parameters, :230-255)
        if (B >= 4) {
            sdrbSsbReset();
            std::vector<int16_t> q0, q1, q2;
            processSSB_opt(std::vector<std::complex<float>>(iq, iq + n), fs, true, q0, pulse, mode);
            processSSB_opt(std::vector<std::complex<float>>(iq + (size_t)n, iq + (size_t)n + n / 2), fs, true, q1, pulse, mode);
            processSSB_opt(std::vector<std::complex<float>>(iq + (size_t)2 * n, iq + (size_t)4 * n), fs, true, q2, pulse, 7);
            const int32_t sizes[3] = {(int32_t)q0.size(), (int32_t)q1.size(), (int32_t)q2.size()};
            put(f, sizes, sizeof sizes);
            put(f, q0.data(), q0.size() * sizeof(int16_t));
            put(f, q1.data(), q1.size() * sizeof(int16_t));
            put(f, q2.data(), q2.size() * sizeof(int16_t));
        }
    } catch (const std::exception& e) {
        fprintf(stderr, "dropin_check: %s\n", e.what());
        fclose(f);
        return 1;
    }
    fclose(f);
    return 0;
}
