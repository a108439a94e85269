// fft16k5.cu — fifth-generation 16384-point FFT kernel (sm_100a): the two 8192-point jobs of a transform
// (fft16k4.cuh: even bins / odd bins after one radix-2 step at ingest) run in ONE 512-thread CTA as two
// 256-thread warp groups that PING-PONG: a group's shared-memory exchange sections (store, group barrier,
// load) are serialised against the other group's by a pair of named barriers (bar.sync / bar.arrive), so an
// exchange of one group always runs under the arithmetic of the other.  Fuses K1-K4 of SURVEY.md 2.3
// (reference src/dsp/fft_process.cpp:62-97).
//
// Why: fft16k2/3 take FP32-pipe time PLUS shared-memory time because all warps of the CTA are in the same
// phase; two free-running CTAs per SM (fft16k4) lock into the same phase and serialise as well (measured,
// profiles/r01_fft16k4_*).  The hand-off below forces the anti-phase schedule
//     group 0:  F  S1 f  S2 F  S1 f  S2 ...        F = phase 3 + next phase 1 arithmetic, f = phase 2
//     group 1:  .. F  S1 f  S2 F  S1 f ...         S = exchange section (holds the token)
// Group g takes parity g of every frame; both read the same staged raw frame (u8: two 32 KiB buffers
// filled by bulk asynchronous copies, cp.async.bulk + mbarrier; cs16 / cf32: straight from global memory).
#include "fft16k4.cuh"
#include "internal.h"

namespace sdrb {

__constant__ cpx c_f5_t3[F4_T3];

namespace {

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count));
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
        "WAIT_LOOP5:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE5;\n"
        "bra WAIT_LOOP5;\n"
        "DONE5:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// named barriers: 1 = "group 0 left its exchange", 2 = "group 1 left its exchange", 3 + g = group-local
__device__ __forceinline__ void nb_sync(int id, int count) {
    asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void nb_arrive(int id, int count) {
    asm volatile("bar.arrive %0, %1;\n" ::"r"(id), "r"(count) : "memory");
}

struct DevTw5 {
    __device__ __forceinline__ cpx operator()(int kb, int c) const { return c_f5_t3[kb * 16 + c]; }
};

template <int FMT>
struct F5Smem {
    static constexpr bool kStaged = FMT == FMT_CU8;
    static constexpr int kStageBytes = kStaged ? F4_N * (int)sizeof(typename Ingest4<FMT>::raw_t) : 0;
    // two exchange buffers, the per-thread twiddle seeds, two stage buffers, two mbarriers
    static constexpr int kBytes = (2 * F4_E_ELEMS + F4_TQ) * 8 + 2 * kStageBytes + 16;
};

// One group's whole life: parity G of every frame this CTA owns.
template <int FMT, int G>
__device__ __forceinline__ void f5_group(int t, const typename Ingest4<FMT>::raw_t* __restrict__ in,
                                         float* __restrict__ out, cpx* E, const cpx* TQ,
                                         typename Ingest4<FMT>::raw_t* stage, unsigned long long* bars, int n_frames) {
    typedef typename Ingest4<FMT>::raw_t raw_t;
    constexpr bool kStaged = F5Smem<FMT>::kStaged;
    constexpr int kStageElems = F4_N;
    constexpr int WAIT_ID = G == 0 ? 2 : 1, DONE_ID = G == 0 ? 1 : 2, GROUP_ID = 3 + G;
    if (G == 1) nb_arrive(2, 512);   // primes the hand-off: group 0 takes the first exchange
    int i = 0;                       // this CTA's frame counter
    for (int f = blockIdx.x; f < n_frames; f += gridDim.x, i++) {
        cpx v[32];
        // ---- arithmetic: ingest + radix-2 combine + DFT32 + twiddle
        const raw_t* src = in + (size_t)f * F4_N;
        if (kStaged) {
            mbar_wait(bars + (i & 1), (unsigned)(i >> 1) & 1u);
            src = stage + (size_t)(i & 1) * kStageElems;
        }
        f4_phase1_load<FMT, G>(t, src, v);
        {
            cpx tq[6];
#pragma unroll
            for (int j = 0; j < 6; j++) tq[j] = TQ[j * 256 + t];
            f4_phase1_compute<G>(tq, v);
        }
        // ---- exchange 1 (token held)
        nb_sync(WAIT_ID, 512);
        f4_phase1_store(t, E, v);
        nb_sync(GROUP_ID, 256);
        if (kStaged && G == 1 && t == 0) {
            // both groups are done with stage buffer i & 1 (group 0's exchange 1 of this frame came first)
            const int f2 = f + 2 * (int)gridDim.x;
            if (f2 < n_frames) {
                mbar_expect_tx(bars + (i & 1), F5Smem<FMT>::kStageBytes);
                bulk_g2s(stage + (size_t)(i & 1) * kStageElems, in + (size_t)f2 * F4_N, F5Smem<FMT>::kStageBytes,
                         bars + (i & 1));
            }
        }
        f4_p2_load(t, E, v);
        nb_arrive(DONE_ID, 512);
        // ---- arithmetic: two DFT16 over b
        f4_p2_compute(v);
        // ---- exchange 2 (token held)
        nb_sync(WAIT_ID, 512);
        f4_p2_store(t, E, v);
        nb_sync(GROUP_ID, 256);
        f4_p3_load(t, E, v);
        nb_arrive(DONE_ID, 512);
        // ---- arithmetic: twiddle, two DFT16 over c, |X|^2, shifted store
        f4_p3_finish<G>(t, v, DevTw5(), out + (size_t)f * F4_N);
    }
    if (G == 0) nb_sync(2, 512);     // absorbs group 1's last arrival
}

template <int FMT>
__global__ void __launch_bounds__(512, 1)
k_fft16k5(const typename Ingest4<FMT>::raw_t* __restrict__ in, float* __restrict__ out,
          const cpx* __restrict__ tables, int n_frames) {
    typedef typename Ingest4<FMT>::raw_t raw_t;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx* E = reinterpret_cast<cpx*>(smem_raw);
    cpx* TQ = E + 2 * F4_E_ELEMS;
    raw_t* stage = reinterpret_cast<raw_t*>(TQ + F4_TQ);
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(reinterpret_cast<unsigned char*>(stage) +
                                                                     2 * F5Smem<FMT>::kStageBytes);
    constexpr bool kStaged = F5Smem<FMT>::kStaged;
    const int tid = threadIdx.x;
    for (int k = tid; k < F4_TQ; k += 512) TQ[k] = tables[k];
    if (kStaged && tid == 0) {
        mbar_init(bars, 1);
        mbar_init(bars + 1, 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
        for (int b = 0; b < 2; b++) {
            const int f = blockIdx.x + b * (int)gridDim.x;
            if (f < n_frames) {
                mbar_expect_tx(bars + b, F5Smem<FMT>::kStageBytes);
                bulk_g2s(stage + (size_t)b * F4_N, in + (size_t)f * F4_N, F5Smem<FMT>::kStageBytes, bars + b);
            }
        }
    }
    __syncthreads();
    if (tid < 256) f5_group<FMT, 0>(tid, in, out, E, TQ, stage, bars, n_frames);
    else f5_group<FMT, 1>(tid - 256, in, out, E + F4_E_ELEMS, TQ, stage, bars, n_frames);
}

template <int FMT>
cudaError_t prep_one() {
    return cudaFuncSetAttribute(k_fft16k5<FMT>, cudaFuncAttributeMaxDynamicSharedMemorySize, F5Smem<FMT>::kBytes);
}

template <int FMT>
void launch_fmt(const void* iq, float* power, const cpx* tables, int n_frames, int grid, cudaStream_t st) {
    typedef typename Ingest4<FMT>::raw_t raw_t;
    k_fft16k5<FMT><<<grid, 512, F5Smem<FMT>::kBytes, st>>>(static_cast<const raw_t*>(iq), power, tables, n_frames);
}

}  // namespace

// host_tables: make_f4_tables() layout (TQ | T3)
cudaError_t fft16k5_prepare(const cpx* host_tables) {
    cudaError_t e;
    if ((e = cudaMemcpyToSymbol(c_f5_t3, host_tables + F4_TQ, F4_T3 * sizeof(cpx))) != cudaSuccess) return e;
    if ((e = prep_one<FMT_CU8>()) != cudaSuccess) return e;
    if ((e = prep_one<FMT_CS16>()) != cudaSuccess) return e;
    return prep_one<FMT_CF32>();
}

// No extension front end here (window / mean removal): the caller uses fft16k2 for those.
cudaError_t launch_fft16k5(int fmt, const void* iq, float* power, const cpx* tables, int n_frames, int sm_count,
                           cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    const int grid = n_frames < sm_count ? n_frames : sm_count;
    switch (fmt) {
        case FMT_CU8: launch_fmt<FMT_CU8>(iq, power, tables, n_frames, grid, st); break;
        case FMT_CS16: launch_fmt<FMT_CS16>(iq, power, tables, n_frames, grid, st); break;
        case FMT_CF32: launch_fmt<FMT_CF32>(iq, power, tables, n_frames, grid, st); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace sdrb
