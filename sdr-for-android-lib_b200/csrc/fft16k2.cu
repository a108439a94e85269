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

__constant__ cpx c_f2_t2[F2_T2];   // W_512^(c kb), [kb][c]
struct F2DevTw {
    __device__ __forceinline__ cpx operator()(int kb, int c) const { return c_f2_t2[kb * 16 + c]; }
};

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
    static constexpr int kBytes = (F2_E_ELEMS + F2_TP) * 8 + kStageBytes + 16;
};

// Extension front end fused into phase 1 (no reference code: FFTProcessor::process feeds the raw
// block to FFTW, src/dsp/fft_process.cpp:62-79).  FRONT bit 1: subtract the block mean (tree sum
// over the CTA); bit 0: multiply by the window table w[n].  Definition: oracle/sdr_oracle_ext.c.
template <int FRONT>
__device__ __forceinline__ void f2_front(int t, cpx* v, const float* __restrict__ win, cpx* red) {
    if (FRONT & 2) {
        cpx s = mk(0.f, 0.f);
#pragma unroll
        for (int a = 0; a < 32; a++) s = cadd(s, v[a]);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            s.x += __shfl_xor_sync(0xffffffffu, s.x, d);
            s.y += __shfl_xor_sync(0xffffffffu, s.y, d);
        }
        if ((t & 31) == 0) red[t >> 5] = s;
        __syncthreads();
        cpx m = mk(0.f, 0.f);
#pragma unroll
        for (int w = 0; w < F2_THREADS / 32; w++) m = cadd(m, red[w]);
        m.x *= 1.0f / F2_N;
        m.y *= 1.0f / F2_N;
#pragma unroll
        for (int a = 0; a < 32; a++) v[a] = csub(v[a], m);
    }
    if (FRONT & 1) {
#pragma unroll
        for (int a = 0; a < 32; a++) {
            const float w = __ldg(win + 512 * a + t);
            v[a].x *= w;
            v[a].y *= w;
        }
    }
}

template <int FMT, int FRONT>
__global__ void __launch_bounds__(F2_THREADS, 1)
k_fft16k2(const typename Ingest2<FMT>::raw_t* __restrict__ in, float* __restrict__ out,
          const cpx* __restrict__ tables, const float* __restrict__ win, int n_frames, int* __restrict__ ticket) {
    typedef typename Ingest2<FMT>::raw_t raw_t;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx* E = reinterpret_cast<cpx*>(smem_raw);
    cpx* TP = E + F2_E_ELEMS;
    raw_t* stage = reinterpret_cast<raw_t*>(TP + F2_TP);
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(reinterpret_cast<unsigned char*>(stage) +
                                                                    F2Smem<FMT>::kStageBytes);
    constexpr bool kStaged = F2Smem<FMT>::kStaged;
    __shared__ cpx red[F2_THREADS / 32];
    __shared__ int s_next;
    const int t = threadIdx.x;

    for (int i = t; i < F2_TP; i += F2_THREADS) TP[i] = tables[i];
    int f = blockIdx.x;
    if (kStaged && t == 0) {
        mbar_init(bar, 1);
        if (f < n_frames) {
            mbar_expect_tx(bar, F2Smem<FMT>::kStageBytes);
            bulk_g2s(stage, in + (size_t)f * F2_N, F2Smem<FMT>::kStageBytes, bar);
        }
    }
    __syncthreads();

    // Which transform a CTA takes next: static round-robin (ticket == nullptr), or the next ticket of a global
    // counter when other kernels share the SMs (the SSB state pass beside the FFT): a CTA slowed by its
    // co-tenants then simply takes fewer transforms instead of finishing last.
    unsigned parity = 0;
    int fn;
    // Rotated loop (no front end): phase 3 of transform f and the load + DFT32 of the NEXT transform sit in one basic
    // block, so the scheduler can put the exchange loads and global stores of phase 3 under the FP work of the next
    // transform's phase 1 (measured on the B200, 37888 u8 transforms: 1.658 -> 1.640 ms per step, still 96 registers).
    if constexpr (FRONT == 0) {
        if (f >= n_frames) return;
        cpx v[32];
        if (kStaged) { mbar_wait(bar, parity); parity ^= 1u; f2_phase1_load<FMT>(t, stage, v); }
        else f2_phase1_load<FMT>(t, in + (size_t)f * F2_N, v);
        fft_dit<5>(v);
        for (;;) {
            __syncthreads();  // E is free (previous phase 3 has read it) and the stage has been consumed
            if (t == 0) {
                const int nx = ticket ? (int)gridDim.x + atomicAdd(ticket, 1) : f + (int)gridDim.x;
                s_next = nx;
                if (kStaged && nx < n_frames) {
                    mbar_expect_tx(bar, F2Smem<FMT>::kStageBytes);
                    bulk_g2s(stage, in + (size_t)nx * F2_N, F2Smem<FMT>::kStageBytes, bar);
                }
            }
            f2_phase1_twiddle_store(t, TP, v, E);
            __syncthreads();
            fn = s_next;
            f2_phase2(t, E);
            __syncthreads();
            if (fn >= n_frames) break;
            if (kStaged) { mbar_wait(bar, parity); parity ^= 1u; f2_phase1_load<FMT>(t, stage, v); }
            else f2_phase1_load<FMT>(t, in + (size_t)fn * F2_N, v);
            f2_phase3(t, E, F2DevTw(), out + (size_t)f * F2_N);
            fft_dit<5>(v);
            f = fn;
        }
        f2_phase3(t, E, F2DevTw(), out + (size_t)f * F2_N);
    } else {
    for (; f < n_frames; f = fn) {
        cpx v[32];
        if (kStaged) {
            mbar_wait(bar, parity);
            parity ^= 1u;
            f2_phase1_load<FMT>(t, stage, v);
        } else {
            f2_phase1_load<FMT>(t, in + (size_t)f * F2_N, v);
        }
        if (FRONT) f2_front<FRONT>(t, v, win, red);
        fft_dit<5>(v);
        __syncthreads();  // E is free (previous phase 3 has read it) and the stage has been consumed
        if (t == 0) {
            const int nx = ticket ? (int)gridDim.x + atomicAdd(ticket, 1) : f + (int)gridDim.x;
            s_next = nx;
            if (kStaged && nx < n_frames) {
                mbar_expect_tx(bar, F2Smem<FMT>::kStageBytes);
                bulk_g2s(stage, in + (size_t)nx * F2_N, F2Smem<FMT>::kStageBytes, bar);
            }
        }
        f2_phase1_twiddle_store(t, TP, v, E);   // twiddle W_N^(t ka) and exchange store, interleaved
        __syncthreads();
        fn = s_next;   // written before this barrier; rewritten only after the next iteration's first barrier
        f2_phase2(t, E);
        __syncthreads();
        f2_phase3(t, E, F2DevTw(), out + (size_t)f * F2_N);
    }
    }
}

template <int FMT, int FRONT>
static cudaError_t prep_one() {
    cudaError_t e = cudaFuncSetAttribute(k_fft16k2<FMT, FRONT>, cudaFuncAttributeMaxDynamicSharedMemorySize, F2Smem<FMT>::kBytes);
    if (e != cudaSuccess) return e;
    // all of the SM's L1 / shared memory as shared memory: what this CTA leaves free (~46 KB for u8) is then there for a
    // co-resident CTA of the SSB pass instead of sitting in an L1 this kernel does not use
    return cudaFuncSetAttribute(k_fft16k2<FMT, FRONT>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                (int)cudaSharedmemCarveoutMaxShared);
}
template <int FMT>
static cudaError_t prep_fmt() {
    cudaError_t e;
    if ((e = prep_one<FMT, 0>()) != cudaSuccess) return e;
    if ((e = prep_one<FMT, 1>()) != cudaSuccess) return e;
    if ((e = prep_one<FMT, 2>()) != cudaSuccess) return e;
    return prep_one<FMT, 3>();
}

cudaError_t fft16k2_prepare(const cpx* host_tables) {
    cudaError_t e;
    if ((e = cudaMemcpyToSymbol(c_f2_t2, host_tables + F2_TP, F2_T2 * sizeof(cpx))) != cudaSuccess) return e;
    if ((e = prep_fmt<FMT_CU8>()) != cudaSuccess) return e;
    if ((e = prep_fmt<FMT_CS16>()) != cudaSuccess) return e;
    return prep_fmt<FMT_CF32>();
}

template <int FMT>
static void launch_fmt(int front, const void* iq, float* power, const cpx* tables, const float* win, int n_frames,
                       int grid, cudaStream_t st, int* ticket) {
    typedef typename Ingest2<FMT>::raw_t raw_t;
    const raw_t* in = static_cast<const raw_t*>(iq);
    const size_t smem = F2Smem<FMT>::kBytes;
    switch (front & 3) {
        case 0: k_fft16k2<FMT, 0><<<grid, F2_THREADS, smem, st>>>(in, power, tables, win, n_frames, ticket); break;
        case 1: k_fft16k2<FMT, 1><<<grid, F2_THREADS, smem, st>>>(in, power, tables, win, n_frames, ticket); break;
        case 2: k_fft16k2<FMT, 2><<<grid, F2_THREADS, smem, st>>>(in, power, tables, win, n_frames, ticket); break;
        default: k_fft16k2<FMT, 3><<<grid, F2_THREADS, smem, st>>>(in, power, tables, win, n_frames, ticket); break;
    }
}

// front: bit 0 = window (win must hold F2_N floats), bit 1 = subtract the block mean
// ticket: device int for dynamic transform scheduling (zeroed here, on the stream), or nullptr for round-robin
cudaError_t launch_fft16k2(int fmt, const void* iq, float* power, const cpx* tables, int n_frames, int sm_count,
                           cudaStream_t st, int front, const float* win, int* ticket) {
    if (n_frames <= 0) return cudaSuccess;
    if ((front & 1) && !win) return cudaErrorInvalidValue;
    const int grid = n_frames < sm_count ? n_frames : sm_count;
    if (ticket) {
        cudaError_t e = cudaMemsetAsync(ticket, 0, sizeof(int), st);
        if (e != cudaSuccess) return e;
    }
    switch (fmt) {
        case FMT_CU8: launch_fmt<FMT_CU8>(front, iq, power, tables, win, n_frames, grid, st, ticket); break;
        case FMT_CS16: launch_fmt<FMT_CS16>(front, iq, power, tables, win, n_frames, grid, st, ticket); break;
        case FMT_CF32: launch_fmt<FMT_CF32>(front, iq, power, tables, win, n_frames, grid, st, ticket); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace sdrb
