// fft16k3.cu — persistent one-CTA-per-SM 16384-point FFT kernel, third generation (sm_100a):
// packed FP32 (FADD2 / FMUL2 / FFMA2) arithmetic on "two worlds" register pairs.  Fuses ingest
// convert, optional window / block-mean removal, FFT, |X|^2 and fftshift (reference
// src/dsp/fft_process.cpp:62-97).  See fft16k3.cuh for the decomposition and the layout.
//
// As in v2 the raw IQ of the NEXT transform is staged into shared memory by one bulk asynchronous
// copy (cp.async.bulk, completion on an mbarrier; SASS UBLKCP) while the current transform is in
// phases 2 and 3 (u8: 32 KiB, cs16: 64 KiB per transform; cf32 is loaded directly).
#include "fft16k3.cuh"
#include "internal.h"

namespace sdrb {

__constant__ pw4 c_f3_t2[F3_T2];   // [kb'][c]  (W_512^(c kb'), W_512^(c (kb' + 16)))
struct F3DevTw {
    __device__ __forceinline__ pw4 operator()(int kbp, int c) const { return c_f3_t2[kbp * 16 + c]; }
};

namespace {
__device__ __forceinline__ unsigned f3_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void f3_mbar_init(void* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(f3_smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
}
__device__ __forceinline__ void f3_mbar_expect_tx(void* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(f3_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void f3_bulk_g2s(void* dst, const void* src, unsigned bytes, void* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                     f3_smem_u32(dst)), "l"(src), "r"(bytes), "r"(f3_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void f3_mbar_wait(void* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "F3_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra F3_DONE;\n"
        "bra F3_WAIT;\n"
        "F3_DONE:\n"
        "}\n" ::"r"(f3_smem_u32(bar)), "r"(parity) : "memory");
}
}  // namespace

template <int FMT>
struct F3Smem {
    static constexpr bool kStaged = FMT != FMT_CF32;
    static constexpr int kStageBytes = kStaged ? F3_N * (int)sizeof(typename Ingest2<FMT>::raw_t) : 0;
    static constexpr int kBytes = F3_E_BYTES + F3_TABLE_BYTES + kStageBytes + 16;
};

// Extension front end fused into phase 1 (no reference code: FFTProcessor::process feeds the raw
// block to FFTW, :62-79).  FRONT bit 1: subtract the block mean (tree sum over the CTA); bit 0:
// multiply by the window table w[n].  Definition: oracle/sdr_oracle_ext.c.
template <int FRONT>
__device__ __forceinline__ void f3_front(int t, c2* v, const float* __restrict__ win, cpx* red) {
    if (FRONT & 2) {
        c2 s = v[0];
#pragma unroll
        for (int j = 1; j < 16; j++) s = c2add(s, v[j]);
        cpx a = mk(s.re.x + s.re.y, s.im.x + s.im.y);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            a.x += __shfl_xor_sync(0xffffffffu, a.x, d);
            a.y += __shfl_xor_sync(0xffffffffu, a.y, d);
        }
        if ((t & 31) == 0) red[t >> 5] = a;
        __syncthreads();
        cpx m = mk(0.f, 0.f);
#pragma unroll
        for (int w = 0; w < F3_THREADS / 32; w++) m = cadd(m, red[w]);
        const pf mr = pdup(m.x * (1.0f / F3_N)), mi = pdup(m.y * (1.0f / F3_N));
#pragma unroll
        for (int j = 0; j < 16; j++) { v[j].re = psub(v[j].re, mr); v[j].im = psub(v[j].im, mi); }
    }
    if (FRONT & 1) {
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const pf w = mkp(__ldg(win + 1024 * j + t), __ldg(win + 1024 * j + 512 + t));
            v[j].re = pmul(v[j].re, w);
            v[j].im = pmul(v[j].im, w);
        }
    }
}

template <int FMT, int FRONT>
__global__ void __launch_bounds__(F3_THREADS, 1)
k_fft16k3(const typename Ingest2<FMT>::raw_t* __restrict__ in, float* __restrict__ out,
          const unsigned char* __restrict__ tables, const float* __restrict__ win, int n_frames) {
    typedef typename Ingest2<FMT>::raw_t raw_t;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    pf* RE = reinterpret_cast<pf*>(smem_raw);
    pf* IM = RE + F3_PLANE_UNITS;
    cpx* TP = reinterpret_cast<cpx*>(smem_raw + F3_E_BYTES);
    raw_t* stage = reinterpret_cast<raw_t*>(smem_raw + F3_E_BYTES + F3_TABLE_BYTES);
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem_raw + F3_E_BYTES + F3_TABLE_BYTES +
                                                                    F3Smem<FMT>::kStageBytes);
    constexpr bool kStaged = F3Smem<FMT>::kStaged;
    __shared__ cpx red[F3_THREADS / 32];
    const int t = threadIdx.x;

    {
        const uint4* src = reinterpret_cast<const uint4*>(tables);
        uint4* dst = reinterpret_cast<uint4*>(smem_raw + F3_E_BYTES);
        for (int i = t; i < F3_TABLE_BYTES / 16; i += F3_THREADS) dst[i] = src[i];
    }
    int f = blockIdx.x;
    if (kStaged && t == 0) {
        f3_mbar_init(bar, 1);
        if (f < n_frames) {
            f3_mbar_expect_tx(bar, F3Smem<FMT>::kStageBytes);
            f3_bulk_g2s(stage, in + (size_t)f * F3_N, F3Smem<FMT>::kStageBytes, bar);
        }
    }
    __syncthreads();

    unsigned parity = 0;
    for (; f < n_frames; f += gridDim.x) {
        c2 v[16];
        if (kStaged) {
            f3_mbar_wait(bar, parity);
            parity ^= 1u;
            f3_phase1_load<FMT>(t, stage, v);
        } else {
            f3_phase1_load<FMT>(t, in + (size_t)f * F3_N, v);
        }
        if (FRONT) f3_front<FRONT>(t, v, win, red);
        // phase-1 arithmetic needs no shared memory: it overlaps the slower warps' phase 3 of the
        // previous transform; only the stores wait for the exchange buffer to be free
        c2 z[16];
        f3_phase1_fft(v, z);
        __syncthreads();  // exchange buffer free (previous phase 3 done), stage consumed by every thread
        if (kStaged && t == 0 && f + (int)gridDim.x < n_frames) {
            f3_mbar_expect_tx(bar, F3Smem<FMT>::kStageBytes);
            f3_bulk_g2s(stage, in + (size_t)(f + gridDim.x) * F3_N, F3Smem<FMT>::kStageBytes, bar);
        }
        f3_phase1_twiddle_store(t, z, TP, reinterpret_cast<float*>(RE), reinterpret_cast<float*>(IM));
        __syncthreads();
        f3_phase2(t, RE, IM);
        __syncthreads();
        f3_phase3(t, RE, IM, F3DevTw(), out + (size_t)f * F3_N);
    }
}

template <int FMT, int FRONT>
static cudaError_t f3_prep_one() {
    return cudaFuncSetAttribute(k_fft16k3<FMT, FRONT>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3Smem<FMT>::kBytes);
}
template <int FMT>
static cudaError_t f3_prep_fmt() {
    cudaError_t e;
    if ((e = f3_prep_one<FMT, 0>()) != cudaSuccess) return e;
    if ((e = f3_prep_one<FMT, 1>()) != cudaSuccess) return e;
    if ((e = f3_prep_one<FMT, 2>()) != cudaSuccess) return e;
    return f3_prep_one<FMT, 3>();
}

cudaError_t fft16k3_prepare(const unsigned char* host_tables) {
    cudaError_t e;
    if ((e = cudaMemcpyToSymbol(c_f3_t2, host_tables + F3_TP * 8, F3_T2 * sizeof(pw4))) != cudaSuccess) return e;
    if ((e = f3_prep_fmt<FMT_CU8>()) != cudaSuccess) return e;
    if ((e = f3_prep_fmt<FMT_CS16>()) != cudaSuccess) return e;
    return f3_prep_fmt<FMT_CF32>();
}

template <int FMT>
static void f3_launch_fmt(int front, const void* iq, float* power, const unsigned char* tables, const float* win,
                          int n_frames, int grid, cudaStream_t st) {
    typedef typename Ingest2<FMT>::raw_t raw_t;
    const raw_t* in = static_cast<const raw_t*>(iq);
    const size_t smem = F3Smem<FMT>::kBytes;
    switch (front & 3) {
        case 0: k_fft16k3<FMT, 0><<<grid, F3_THREADS, smem, st>>>(in, power, tables, win, n_frames); break;
        case 1: k_fft16k3<FMT, 1><<<grid, F3_THREADS, smem, st>>>(in, power, tables, win, n_frames); break;
        case 2: k_fft16k3<FMT, 2><<<grid, F3_THREADS, smem, st>>>(in, power, tables, win, n_frames); break;
        default: k_fft16k3<FMT, 3><<<grid, F3_THREADS, smem, st>>>(in, power, tables, win, n_frames); break;
    }
}

// front: bit 0 = window (win must hold F3_N floats), bit 1 = subtract the block mean
cudaError_t launch_fft16k3(int fmt, const void* iq, float* power, const unsigned char* tables, int n_frames,
                           int sm_count, cudaStream_t st, int front, const float* win) {
    if (n_frames <= 0) return cudaSuccess;
    if ((front & 1) && !win) return cudaErrorInvalidValue;
    const int grid = n_frames < sm_count ? n_frames : sm_count;
    switch (fmt) {
        case FMT_CU8: f3_launch_fmt<FMT_CU8>(front, iq, power, tables, win, n_frames, grid, st); break;
        case FMT_CS16: f3_launch_fmt<FMT_CS16>(front, iq, power, tables, win, n_frames, grid, st); break;
        case FMT_CF32: f3_launch_fmt<FMT_CF32>(front, iq, power, tables, win, n_frames, grid, st); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace sdrb
