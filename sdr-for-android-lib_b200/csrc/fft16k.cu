// fft16k.cu — persistent one-CTA-per-transform 16384-point FFT kernel for sm_100a.
// Fuses K1-K4 of SURVEY.md section 2.3: ingest convert, FFT, |X|^2, fftshift
// (reference: src/dsp/fft_process.cpp:62-97).  See fft16k.cuh for the decomposition.
#include "fft16k.cuh"
#include "internal.h"

namespace sdrb {

template <int FMT>
__global__ void __launch_bounds__(F16K_THREADS, 1)
k_fft16k(const typename Ingest<FMT>::raw_t* __restrict__ in, float* __restrict__ out,
         const cpx* __restrict__ tables, int n_frames) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cpx* E = reinterpret_cast<cpx*>(smem_raw);
    cpx* TA = E + F16K_E_ELEMS;
    cpx* TB = TA + F16K_TA;
    cpx* T2 = TB + F16K_TB;
    cpx* T3 = T2 + F16K_T2;
    const int t = threadIdx.x;

    for (int i = t; i < F16K_TW_ELEMS; i += F16K_THREADS) TA[i] = tables[i];
    __syncthreads();

    for (int f = blockIdx.x; f < n_frames; f += gridDim.x) {
        const typename Ingest<FMT>::raw_t* src = in + (size_t)f * F16K_N;
        float* dst = out + (size_t)f * F16K_N;
        f16k_phase1<FMT>(t, src, E, TA, TB);
        __syncthreads();
        f16k_phase2(t, E, T2);
        __syncthreads();
        f16k_phase3(t, E, T3);
        __syncthreads();
        f16k_phase4(t, E, dst, Ingest<FMT>::kPowerScale);
        __syncthreads();
    }
}

cudaError_t fft16k_prepare() {
    cudaError_t e;
    e = cudaFuncSetAttribute(k_fft16k<FMT_CU8>, cudaFuncAttributeMaxDynamicSharedMemorySize, F16K_SMEM_BYTES);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_fft16k<FMT_CS16>, cudaFuncAttributeMaxDynamicSharedMemorySize, F16K_SMEM_BYTES);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_fft16k<FMT_CF32>, cudaFuncAttributeMaxDynamicSharedMemorySize, F16K_SMEM_BYTES);
    return e;
}

cudaError_t launch_fft16k(int fmt, const void* iq, float* power, const cpx* tables, int n_frames,
                          int sm_count, cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    const int grid = n_frames < sm_count ? n_frames : sm_count;
    switch (fmt) {
        case FMT_CU8:
            k_fft16k<FMT_CU8><<<grid, F16K_THREADS, F16K_SMEM_BYTES, st>>>(
                static_cast<const uint16_t*>(iq), power, tables, n_frames);
            break;
        case FMT_CS16:
            k_fft16k<FMT_CS16><<<grid, F16K_THREADS, F16K_SMEM_BYTES, st>>>(
                static_cast<const uint32_t*>(iq), power, tables, n_frames);
            break;
        case FMT_CF32:
            k_fft16k<FMT_CF32><<<grid, F16K_THREADS, F16K_SMEM_BYTES, st>>>(
                static_cast<const cpx*>(iq), power, tables, n_frames);
            break;
        default:
            return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace sdrb
