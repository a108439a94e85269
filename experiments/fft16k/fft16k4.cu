// fft16k4.cu — fourth-generation 16384-point FFT kernel (sm_100a): two 8192-point jobs per
// transform, 256 threads per CTA, TWO resident CTAs per SM running out of phase.  Fuses K1-K4 of
// SURVEY.md 2.3 (ingest convert, FFT, |X|^2, fftshift; reference src/dsp/fft_process.cpp:62-97).
// See fft16k4.cuh for the decomposition and layout.
//
// A CTA takes whole frames (persistent, stride gridDim.x) and runs the even-bin job and then the
// odd-bin job of each, so the raw frame is fetched once: u8 frames (32 KiB) are staged into shared
// memory by one bulk asynchronous copy (cp.async.bulk + mbarrier, SASS UBLKCP) issued as soon as
// the odd job has consumed the previous frame; cs16 / cf32 frames do not fit twice per SM beside
// the exchange buffers and are read straight from global memory (the sibling CTA covers the latency).
#include "fft16k4.cuh"
#include "internal.h"

namespace sdrb {

__constant__ cpx c_f4_t3[F4_T3];

namespace {

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
        "WAIT_LOOP4:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE4;\n"
        "bra WAIT_LOOP4;\n"
        "DONE4:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

struct DevTw3 {
    __device__ __forceinline__ cpx operator()(int kb, int c) const { return c_f4_t3[kb * 16 + c]; }
};

template <int FMT>
struct F4Smem {
    static constexpr bool kStaged = FMT == FMT_CU8;
    static constexpr int kStageBytes = kStaged ? F4_N * (int)sizeof(typename Ingest4<FMT>::raw_t) : 0;
    static constexpr int kBytes = (F4_E_ELEMS + F4_TQ) * 8 + kStageBytes + 16;
};

// Extension front end (no reference code; definition: oracle/sdr_oracle_ext.c).  FRONT bit 1: subtract
// the block mean; bit 0: multiply by the window table.  Both act on the raw samples before the
// radix-2 combine, so the load is done here instead of f4_phase1_load.
template <int FMT, int P, int FRONT>
__device__ __forceinline__ void f4_front_load(int t, const typename Ingest4<FMT>::raw_t* in, cpx* v,
                                              const float* __restrict__ win, cpx* red) {
    cpx m = mk(0.f, 0.f);
    if (FRONT & 2) {
        cpx s = mk(0.f, 0.f);
#pragma unroll 8
        for (int a = 0; a < 64; a++) s = cadd(s, Ingest4<FMT>::cvt(in[256 * a + t]));
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            s.x += __shfl_xor_sync(0xffffffffu, s.x, d);
            s.y += __shfl_xor_sync(0xffffffffu, s.y, d);
        }
        __syncthreads();  // red is free (previous job's readers are done)
        if ((t & 31) == 0) red[t >> 5] = s;
        __syncthreads();
#pragma unroll
        for (int w = 0; w < F4_THREADS / 32; w++) m = cadd(m, red[w]);
        m.x *= 1.0f / F4_N;
        m.y *= 1.0f / F4_N;
    }
#pragma unroll
    for (int a = 0; a < 32; a++) {
        cpx x0 = csub(Ingest4<FMT>::cvt(in[256 * a + t]), m);
        cpx x1 = csub(Ingest4<FMT>::cvt(in[F4_H + 256 * a + t]), m);
        if (FRONT & 1) {
            const float w0 = __ldg(win + 256 * a + t), w1 = __ldg(win + F4_H + 256 * a + t);
            x0.x *= w0; x0.y *= w0;
            x1.x *= w1; x1.y *= w1;
        }
        v[a] = P ? csub(x0, x1) : cadd(x0, x1);
    }
}

template <int FMT, int FRONT, int P>
__device__ __forceinline__ void f4_job(int t, const typename Ingest4<FMT>::raw_t* src, cpx* E, const cpx* TQ,
                                       const float* __restrict__ win, cpx* red, float* __restrict__ out,
                                       bool issue_next, void* bar, void* stage, const void* next_src) {
    cpx v[32];
    __syncthreads();  // E is free: the previous job's phase 3 has read it (also keeps ptxas from merging
                      // that phase 3 with this phase 1, which spills the transform)
    if (FRONT) f4_front_load<FMT, P, FRONT>(t, src, v, win, red);
    else f4_phase1_load<FMT, P>(t, src, v);
    {
        cpx tq[6];
#pragma unroll
        for (int j = 0; j < 6; j++) tq[j] = TQ[j * 256 + t];
        f4_phase1_compute<P>(tq, v);
    }
    f4_phase1_store(t, E, v);
    __syncthreads();
    // after the odd job's loads every thread is done with the staged frame: fetch the next one
    if (F4Smem<FMT>::kStaged && P == 1 && issue_next && t == 0) {
        mbar_expect_tx(bar, F4Smem<FMT>::kStageBytes);
        bulk_g2s(stage, next_src, F4Smem<FMT>::kStageBytes, bar);
    }
    f4_phase2(t, E);
    __syncthreads();
    f4_phase3<P>(t, E, DevTw3(), out);
}

template <int FMT, int FRONT>
__global__ void __launch_bounds__(F4_THREADS, 2)
k_fft16k4(const typename Ingest4<FMT>::raw_t* __restrict__ in, float* __restrict__ out,
          const cpx* __restrict__ tables, const float* __restrict__ win, int n_frames) {
    typedef typename Ingest4<FMT>::raw_t raw_t;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx* E = reinterpret_cast<cpx*>(smem_raw);
    cpx* TQ = E + F4_E_ELEMS;
    raw_t* stage = reinterpret_cast<raw_t*>(TQ + F4_TQ);
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(reinterpret_cast<unsigned char*>(stage) +
                                                                    F4Smem<FMT>::kStageBytes);
    constexpr bool kStaged = F4Smem<FMT>::kStaged;
    __shared__ cpx red[F4_THREADS / 32];
    const int t = threadIdx.x;

    for (int i = t; i < F4_TQ; i += F4_THREADS) TQ[i] = tables[i];

    int f = blockIdx.x;
    if (kStaged && t == 0) {
        mbar_init(bar, 1);
        if (f < n_frames) {
            mbar_expect_tx(bar, F4Smem<FMT>::kStageBytes);
            bulk_g2s(stage, in + (size_t)f * F4_N, F4Smem<FMT>::kStageBytes, bar);
        }
    }
    __syncthreads();

    unsigned parity = 0;
    for (; f < n_frames; f += gridDim.x) {
        const raw_t* src = in + (size_t)f * F4_N;
        if (kStaged) {
            mbar_wait(bar, parity);
            parity ^= 1u;
            src = stage;
        }
        float* dst = out + (size_t)f * F4_N;
        const bool more = f + (int)gridDim.x < n_frames;
        f4_job<FMT, FRONT, 0>(t, src, E, TQ, win, red, dst, false, bar, stage, nullptr);
        f4_job<FMT, FRONT, 1>(t, src, E, TQ, win, red, dst, more, bar, stage, in + (size_t)(f + gridDim.x) * F4_N);
    }
}

template <int FMT, int FRONT>
cudaError_t prep_one() {
    return cudaFuncSetAttribute(k_fft16k4<FMT, FRONT>, cudaFuncAttributeMaxDynamicSharedMemorySize, F4Smem<FMT>::kBytes);
}
template <int FMT>
cudaError_t prep_fmt() {
    cudaError_t e;
    if ((e = prep_one<FMT, 0>()) != cudaSuccess) return e;
    if ((e = prep_one<FMT, 1>()) != cudaSuccess) return e;
    if ((e = prep_one<FMT, 2>()) != cudaSuccess) return e;
    return prep_one<FMT, 3>();
}

template <int FMT>
void launch_fmt(int front, const void* iq, float* power, const cpx* tables, const float* win, int n_frames, int grid,
                cudaStream_t st) {
    typedef typename Ingest4<FMT>::raw_t raw_t;
    const raw_t* in = static_cast<const raw_t*>(iq);
    const size_t smem = F4Smem<FMT>::kBytes;
    switch (front & 3) {
        case 0: k_fft16k4<FMT, 0><<<grid, F4_THREADS, smem, st>>>(in, power, tables, win, n_frames); break;
        case 1: k_fft16k4<FMT, 1><<<grid, F4_THREADS, smem, st>>>(in, power, tables, win, n_frames); break;
        case 2: k_fft16k4<FMT, 2><<<grid, F4_THREADS, smem, st>>>(in, power, tables, win, n_frames); break;
        default: k_fft16k4<FMT, 3><<<grid, F4_THREADS, smem, st>>>(in, power, tables, win, n_frames); break;
    }
}

}  // namespace

// tables: the make_f4_tables() layout in device memory (TQ | T3); T3 is also placed in constant memory here.
cudaError_t fft16k4_prepare(const cpx* host_tables) {
    cudaError_t e;
    if ((e = cudaMemcpyToSymbol(c_f4_t3, host_tables + F4_TQ, F4_T3 * sizeof(cpx))) != cudaSuccess) return e;
    if ((e = prep_fmt<FMT_CU8>()) != cudaSuccess) return e;
    if ((e = prep_fmt<FMT_CS16>()) != cudaSuccess) return e;
    return prep_fmt<FMT_CF32>();
}

// front: bit 0 = window (win must hold F4_N floats), bit 1 = subtract the block mean
cudaError_t launch_fft16k4(int fmt, const void* iq, float* power, const cpx* tables, int n_frames, int sm_count,
                           cudaStream_t st, int front, const float* win) {
    if (n_frames <= 0) return cudaSuccess;
    if ((front & 1) && !win) return cudaErrorInvalidValue;
    const int grid = n_frames < 2 * sm_count ? n_frames : 2 * sm_count;
    switch (fmt) {
        case FMT_CU8: launch_fmt<FMT_CU8>(front, iq, power, tables, win, n_frames, grid, st); break;
        case FMT_CS16: launch_fmt<FMT_CS16>(front, iq, power, tables, win, n_frames, grid, st); break;
        case FMT_CF32: launch_fmt<FMT_CF32>(front, iq, power, tables, win, n_frames, grid, st); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace sdrb
