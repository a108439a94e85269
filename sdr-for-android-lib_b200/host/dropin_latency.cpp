// host/dropin_latency.cpp — per-reading latency of the header-compatible front end on the GPU, the way
// soapyCallback pays it (src/sdr-bridge-java-soapy.cpp:441-466): FFTProcessor::process + all getters, and
// processSSB_opt, one reading at a time with host buffers (batch of one: H2D, kernels, D2H inside the call).
// The reference's budget at its defaults (2.5 MS/s, 16384 samples per reading) is 6.55 ms per reading.
//   dropin_latency.bin [n = 16384] [fs = 2500000] [readings = 300]      -> one JSON line
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "fft_process.h"
#include "ssb_demod_opt.h"

static double pct(std::vector<double> v, double p) {
    std::sort(v.begin(), v.end());
    return v[(size_t)(p * (v.size() - 1))];
}

int main(int argc, char** argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 16384;
    const uint32_t fs = argc > 2 ? (uint32_t)atol(argv[2]) : 2500000u;
    const int R = argc > 3 ? atoi(argv[3]) : 300;
    const int nbuf = 16;
    std::vector<float> raw((size_t)nbuf * n * 2);
    if (sdrb_synth_host(SDRB_CF32, fs, 0, 1, 0, (int64_t)nbuf * n, raw.data()) != SDRB_OK) return 2;
    const auto* iq = reinterpret_cast<const std::complex<float>*>(raw.data());
    FFTProcessor fft;
    fft.configure(FftProcessorConfig{430000000u, fs, n, 5});
    std::vector<int16_t> pcm;
    bool pulse = false;
    std::vector<double> t_fft, t_ssb, t_both;
    double sink = 0;
    for (int r = -20; r < R; r++) {   // 20 untimed readings: context creation, first-use allocations
        const std::complex<float>* blk = iq + (size_t)((r + 20) % nbuf) * n;
        const auto t0 = std::chrono::steady_clock::now();
        fft.process(blk, (uint32_t)n);
        sink += fft.getMeanSnrDb() + fft.getMeanSnrSigma() + (double)fft.getTrackingFrequency() + fft.getDetectionFlag() +
                fft.getPeakAboveNoiseMeanDb() + fft.getMaxBinSnrDb() + fft.getMaxBinSnrSigma() + fft.getBest1kHzSnrDb() +
                fft.getBest1kHzSnrSigma() + fft.getBest1kHzCenterFreqHz() + fft.getPerBinMean() +
                fft.getPowerSpectrum()[(size_t)n / 2];
        const auto t1 = std::chrono::steady_clock::now();
        processSSB_opt(std::vector<std::complex<float>>(blk, blk + n), fs, true, pcm, pulse, 1);
        const auto t2 = std::chrono::steady_clock::now();
        if (r < 0) continue;
        t_fft.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
        t_ssb.push_back(std::chrono::duration<double, std::milli>(t2 - t1).count());
        t_both.push_back(std::chrono::duration<double, std::milli>(t2 - t0).count());
    }
    if (fft.failedCalls()) { fprintf(stderr, "FFTProcessor: %s\n", fft.lastError().c_str()); return 1; }
    printf("{\"n\": %d, \"sample_rate\": %u, \"readings\": %d, \"budget_ms_per_reading\": %.3f, "
           "\"fft_process_plus_getters_ms\": {\"median\": %.3f, \"p99\": %.3f, \"max\": %.3f}, "
           "\"processSSB_opt_ms\": {\"median\": %.3f, \"p99\": %.3f, \"max\": %.3f}, "
           "\"both_ms\": {\"median\": %.3f, \"p99\": %.3f, \"max\": %.3f}, \"pcm_per_reading\": %zu, \"checksum\": %.3f}\n",
           n, fs, R, 1e3 * n / (double)fs, pct(t_fft, 0.5), pct(t_fft, 0.99), pct(t_fft, 1.0), pct(t_ssb, 0.5),
           pct(t_ssb, 0.99), pct(t_ssb, 1.0), pct(t_both, 0.5), pct(t_both, 0.99), pct(t_both, 1.0), pcm.size(), sink);
    return 0;
}
