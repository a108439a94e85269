// synth.cu — deterministic synthetic IQ source: host loop and device kernel over the same
// integer sample function (synth.cuh).  Stands in for SoapySDR readStream
// (src/sdr-bridge-java-soapy.cpp:532-539 of the reference) on a box without a dongle.
#include <cmath>
#include <mutex>

#include "internal.h"
#include "synth.cuh"

namespace sdrb {

const int16_t* synth_sine_table() {
    static int16_t tab[SYNTH_SINE_N];
    static std::once_flag once;
    std::call_once(once, [] {
        for (int i = 0; i < SYNTH_SINE_N; i++) {
            const double a = 2.0 * 3.14159265358979323846264338327950288 * (double)i / (double)SYNTH_SINE_N;
            tab[i] = (int16_t)std::lrint(32767.0 * std::sin(a));
        }
    });
    return tab;
}

template <int FMT>
SDRB_HD void synth_store(void* out, size_t idx, int I, int Q) {
    if (FMT == FMT_CU8) {
        static_cast<uint16_t*>(out)[idx] = (uint16_t)((unsigned)I | ((unsigned)Q << 8));
    } else if (FMT == FMT_CS16) {
        static_cast<uint32_t*>(out)[idx] = ((uint32_t)(uint16_t)(int16_t)I) | ((uint32_t)(uint16_t)(int16_t)Q << 16);
    } else {
        static_cast<cpx*>(out)[idx] = mk((float)I * (1.0f / 32768.0f), (float)Q * (1.0f / 32768.0f));
    }
}

template <int FMT>
static void synth_host_fmt(uint32_t fs, int first_stream, int n_streams, int64_t first_sample, int64_t n_samples,
                           void* out) {
    const int16_t* tab = synth_sine_table();
    for (int s = 0; s < n_streams; s++) {
        const SynthStream p = synth_stream(first_stream + s, fs);
        uint32_t key_pos = (uint32_t)(((uint64_t)first_sample + p.keyOffset) % p.keyPeriod);
        for (int64_t j = 0; j < n_samples; j++) {
            int I, Q;
            synth_sample<FMT>(p, tab, first_sample + j, key_pos, I, Q);
            synth_store<FMT>(out, (size_t)s * (size_t)n_samples + (size_t)j, I, Q);
            if (++key_pos == p.keyPeriod) key_pos = 0;
        }
    }
}

void synth_host(int fmt, uint32_t fs, int first_stream, int n_streams, int64_t first_sample, int64_t n_samples,
                void* out) {
    switch (fmt) {
        case FMT_CU8: synth_host_fmt<FMT_CU8>(fs, first_stream, n_streams, first_sample, n_samples, out); break;
        case FMT_CS16: synth_host_fmt<FMT_CS16>(fs, first_stream, n_streams, first_sample, n_samples, out); break;
        default: synth_host_fmt<FMT_CF32>(fs, first_stream, n_streams, first_sample, n_samples, out); break;
    }
}

template <int FMT>
__global__ void __launch_bounds__(256)
k_synth(uint32_t fs, int first_stream, int64_t first_sample, int64_t n_samples, void* __restrict__ out,
        const int16_t* __restrict__ sine) {
    __shared__ int16_t tab[SYNTH_SINE_N];
    for (int i = threadIdx.x; i < SYNTH_SINE_N; i += blockDim.x) tab[i] = sine[i];
    __syncthreads();
    const int s = blockIdx.y;
    const SynthStream p = synth_stream(first_stream + s, fs);
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_samples;
         j += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = first_sample + j;
        const uint32_t key_pos = (uint32_t)(((uint64_t)i + p.keyOffset) % p.keyPeriod);
        int I, Q;
        synth_sample<FMT>(p, tab, i, key_pos, I, Q);
        synth_store<FMT>(out, (size_t)s * (size_t)n_samples + (size_t)j, I, Q);
    }
}

cudaError_t launch_synth(int fmt, uint32_t fs, int first_stream, int n_streams, int64_t first_sample,
                         int64_t n_samples, void* d_out, const int16_t* d_sine, cudaStream_t st) {
    if (n_streams <= 0 || n_samples <= 0) return cudaSuccess;
    int64_t bx = (n_samples + 256 * 8 - 1) / (256 * 8);
    if (bx > 4096) bx = 4096;
    if (bx < 1) bx = 1;
    // gridDim.y is limited to 65535 streams per launch
    for (int s0 = 0; s0 < n_streams; s0 += 32768) {
        const int ns = n_streams - s0 < 32768 ? n_streams - s0 : 32768;
        dim3 grid((unsigned)bx, (unsigned)ns);
        const size_t bytes_per = fmt == FMT_CU8 ? 2 : (fmt == FMT_CS16 ? 4 : 8);
        void* o = static_cast<unsigned char*>(d_out) + (size_t)s0 * (size_t)n_samples * bytes_per;
        switch (fmt) {
            case FMT_CU8: k_synth<FMT_CU8><<<grid, 256, 0, st>>>(fs, first_stream + s0, first_sample, n_samples, o, d_sine); break;
            case FMT_CS16: k_synth<FMT_CS16><<<grid, 256, 0, st>>>(fs, first_stream + s0, first_sample, n_samples, o, d_sine); break;
            case FMT_CF32: k_synth<FMT_CF32><<<grid, 256, 0, st>>>(fs, first_stream + s0, first_sample, n_samples, o, d_sine); break;
            default: return cudaErrorInvalidValue;
        }
    }
    return cudaGetLastError();
}

}  // namespace sdrb
