// fft16k2.cu — persistent one-CTA-per-SM 16384-point FFT kernel, second generation (sm_100a).
// Fuses K1-K4 of SURVEY.md 2.3 (ingest convert, FFT, |X|^2, fftshift; reference
// src/dsp/fft_process.cpp:62-97).  See fft16k2.cuh for the decomposition and layout.
//
// The raw IQ of the NEXT transform is staged into shared memory by one bulk asynchronous copy
// (cp.async.bulk, completion on an mbarrier) while the current transform is in phases 2 and 3,
// so the global-load latency of phase 1 is off the critical path (u8: 32 KiB, cs16: 64 KiB per
// transform; cf32 does not fit beside the 128 KiB exchange buffer and is loaded directly).
#include "fft16k2.cuh"
#include "internal.h"

namespace sdrb {

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(void* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, void* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(void* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

template <int FMT>
struct F2Smem {
    static constexpr bool kStaged = FMT != FMT_CF32;
    static constexpr int kStageBytes = kStaged ? F2_N * (int)sizeof(typename Ingest2<FMT>::raw_t) : 0;
    static constexpr int kBytes = (F2_E_ELEMS + F2_TW_ELEMS) * 8 + kStageBytes + 16;
};

template <int FMT>
__global__ void __launch_bounds__(F2_THREADS, 1)
k_fft16k2(const typename Ingest2<FMT>::raw_t* __restrict__ in, float* __restrict__ out,
          const cpx* __restrict__ tables, int n_frames) {
    typedef typename Ingest2<FMT>::raw_t raw_t;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx* E = reinterpret_cast<cpx*>(smem_raw);
    cpx* TP = E + F2_E_ELEMS;
    cpx* T2 = TP + F2_TP;
    raw_t* stage = reinterpret_cast<raw_t*>(T2 + F2_T2);
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(reinterpret_cast<unsigned char*>(stage) +
                                                                    F2Smem<FMT>::kStageBytes);
    constexpr bool kStaged = F2Smem<FMT>::kStaged;
    const int t = threadIdx.x;

    for (int i = t; i < F2_TW_ELEMS; i += F2_THREADS) TP[i] = tables[i];
    int f = blockIdx.x;
    if (kStaged && t == 0) {
        mbar_init(bar, 1);
        if (f < n_frames) {
            mbar_expect_tx(bar, F2Smem<FMT>::kStageBytes);
            bulk_g2s(stage, in + (size_t)f * F2_N, F2Smem<FMT>::kStageBytes, bar);
        }
    }
    __syncthreads();

    unsigned parity = 0;
    for (; f < n_frames; f += gridDim.x) {
        cpx v[32];
        if (kStaged) {
            mbar_wait(bar, parity);
            parity ^= 1u;
            f2_phase1<FMT>(t, stage, TP, v);
        } else {
            f2_phase1<FMT>(t, in + (size_t)f * F2_N, TP, v);
        }
        __syncthreads();  // E is free (previous phase 3 done) and the stage has been consumed
        if (kStaged && t == 0 && f + (int)gridDim.x < n_frames) {
            mbar_expect_tx(bar, F2Smem<FMT>::kStageBytes);
            bulk_g2s(stage, in + (size_t)(f + gridDim.x) * F2_N, F2Smem<FMT>::kStageBytes, bar);
        }
        f2_phase1_store(t, E, v);
        __syncthreads();
        f2_phase2(t, E, T2);
        __syncthreads();
        f2_phase3(t, E, out + (size_t)f * F2_N);
    }
}

cudaError_t fft16k2_prepare() {
    cudaError_t e;
    e = cudaFuncSetAttribute(k_fft16k2<FMT_CU8>, cudaFuncAttributeMaxDynamicSharedMemorySize, F2Smem<FMT_CU8>::kBytes);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_fft16k2<FMT_CS16>, cudaFuncAttributeMaxDynamicSharedMemorySize, F2Smem<FMT_CS16>::kBytes);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(k_fft16k2<FMT_CF32>, cudaFuncAttributeMaxDynamicSharedMemorySize, F2Smem<FMT_CF32>::kBytes);
}

cudaError_t launch_fft16k2(int fmt, const void* iq, float* power, const cpx* tables, int n_frames, int sm_count,
                           cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    const int grid = n_frames < sm_count ? n_frames : sm_count;
    switch (fmt) {
        case FMT_CU8:
            k_fft16k2<FMT_CU8><<<grid, F2_THREADS, F2Smem<FMT_CU8>::kBytes, st>>>(
                static_cast<const uint16_t*>(iq), power, tables, n_frames);
            break;
        case FMT_CS16:
            k_fft16k2<FMT_CS16><<<grid, F2_THREADS, F2Smem<FMT_CS16>::kBytes, st>>>(
                static_cast<const uint32_t*>(iq), power, tables, n_frames);
            break;
        case FMT_CF32:
            k_fft16k2<FMT_CF32><<<grid, F2_THREADS, F2Smem<FMT_CF32>::kBytes, st>>>(
                static_cast<const cpx*>(iq), power, tables, n_frames);
            break;
        default:
            return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace sdrb
