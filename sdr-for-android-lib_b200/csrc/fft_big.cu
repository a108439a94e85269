// fft_big.cu — multi-CTA four-step FFT kernels for N = 65536 and N = 1048576 (sm_100a).
// See fft_big.cuh for the decomposition, the intermediate layout and the phase functions.
//
//   k_fbA<R, FMT> : grid = frames x (L / 16) column tiles, 16 R threads.  Raw IQ is read straight
//                   into registers (16 columns x 128-byte-or-less segments per row), one in-place
//                   exchange through shared memory, result to the L2-resident intermediate A'.
//   k_fbB<R>      : grid = frames x (L / 16) row groups.  The group's 16 rows are ONE contiguous
//                   128 L-byte range of A', brought into shared memory by a single bulk TMA copy
//                   (cp.async.bulk, completion on an mbarrier; SASS UBLKCP), then two register
//                   passes, |X|^2 and the shifted store (64-byte segments).
// Frames are processed in groups sized so that A' (8 N bytes per frame) stays within the 126 MB
// L2 between the two kernels.
#include "fft_big.cuh"
#include "internal.h"

namespace sdrb {

namespace {
__device__ __forceinline__ unsigned fb_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void fb_mbar_init(void* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(fb_smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
}
__device__ __forceinline__ void fb_mbar_expect_tx(void* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(fb_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void fb_bulk_g2s(void* dst, const void* src, unsigned bytes, void* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                     fb_smem_u32(dst)), "l"(src), "r"(bytes), "r"(fb_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fb_mbar_wait(void* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "FB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra FB_DONE;\n"
        "bra FB_WAIT;\n"
        "FB_DONE:\n"
        "}\n" ::"r"(fb_smem_u32(bar)), "r"(parity) : "memory");
}
}  // namespace

template <int R, int FMT>
__global__ void __launch_bounds__(FbGeom<R>::kThreads)
k_fbA(const typename Ingest2<FMT>::raw_t* __restrict__ in, cpx* __restrict__ A, const cpx* __restrict__ tables) {
    typedef FbGeom<R> G;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx* S = reinterpret_cast<cpx*>(smem_raw);
    cpx* TL = S + G::kElemsA;
    const int t = threadIdx.x;
    for (int i = t; i < G::kTL; i += G::kThreads) TL[i] = tables[i];
    const cpx* TLO = tables + G::kTL;
    const cpx* THI = TLO + G::kTLO;
    const cpx* TB = THI + G::kTHI;
    const size_t frame = blockIdx.y;
    const int n2_0 = blockIdx.x * 16;
    constexpr size_t N = (size_t)G::L * G::L;
    __syncthreads();
    fbA_pass0<R, FMT>(t, in + frame * N, n2_0, TL, S);
    __syncthreads();
    fbA_pass1<R>(t, S, TLO, THI, TB, n2_0, A + frame * N);
}

// Persistent: a CTA walks row groups (tile = frame * L/16 + group; the groups of A' are contiguous, so the tile's
// source is A + tile * kElemsB).  Once pass 1 has pulled its values into registers the shared-memory tile is free,
// so the bulk copy of the NEXT tile is issued there and lands under pass 1's arithmetic and global stores.
template <int R>
__global__ void __launch_bounds__(FbGeom<R>::kThreads)
k_fbB(const cpx* __restrict__ A, float* __restrict__ out, const cpx* __restrict__ tables, int n_tiles) {
    typedef FbGeom<R> G;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx* W = reinterpret_cast<cpx*>(smem_raw);
    cpx* TL = W + G::kElemsB;
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(TL + G::kTL);
    const int t = threadIdx.x;
    constexpr size_t N = (size_t)G::L * G::L;
    constexpr int kGroups = G::L / 16;
    constexpr unsigned kBytes = G::kElemsB * sizeof(cpx);
    int tile = blockIdx.x;
    if (t == 0) {
        fb_mbar_init(bar, 1);
        if (tile < n_tiles) {
            fb_mbar_expect_tx(bar, kBytes);
            fb_bulk_g2s(W, A + (size_t)tile * G::kElemsB, kBytes, bar);
        }
    }
    for (int i = t; i < G::kTL; i += G::kThreads) TL[i] = tables[i];
    __syncthreads();
    unsigned parity = 0;
    for (; tile < n_tiles; tile += gridDim.x) {
        fb_mbar_wait(bar, parity);
        parity ^= 1u;
        fbB_pass0<R>(t, W, TL);
        __syncthreads();
        cpx v[R];
        fbB_pass1_load<R>(t, W, v);
        __syncthreads();   // every thread holds its values: the tile buffer is free
        const int next = tile + (int)gridDim.x;
        if (t == 0 && next < n_tiles) {
            // pass 0 wrote the tile through the generic proxy; order those writes before the async-proxy refill
            asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
            fb_mbar_expect_tx(bar, kBytes);
            fb_bulk_g2s(W, A + (size_t)next * G::kElemsB, kBytes, bar);
        }
        const int frame = tile / kGroups, g = tile % kGroups;
        fbB_pass1_finish<R>(t, v, g * 16, out + (size_t)frame * N);
    }
}

template <int R>
struct FbSmem {
    static constexpr int kA = (FbGeom<R>::kElemsA + FbGeom<R>::kTL) * 8;
    static constexpr int kB = (FbGeom<R>::kElemsB + FbGeom<R>::kTL) * 8 + 16;
};

template <int R>
static cudaError_t fb_prepare_r() {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_fbA<R, FMT_CU8>, cudaFuncAttributeMaxDynamicSharedMemorySize, FbSmem<R>::kA)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fbA<R, FMT_CS16>, cudaFuncAttributeMaxDynamicSharedMemorySize, FbSmem<R>::kA)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fbA<R, FMT_CF32>, cudaFuncAttributeMaxDynamicSharedMemorySize, FbSmem<R>::kA)) != cudaSuccess) return e;
    return cudaFuncSetAttribute(k_fbB<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, FbSmem<R>::kB);
}

cudaError_t fft_big_prepare() {
    cudaError_t e = fb_prepare_r<16>();
    if (e != cudaSuccess) return e;
    return fb_prepare_r<32>();
}

bool fft_big_supports(int n) { return n == 65536 || n == 1048576; }
int fft_big_table_elems(int n) { return n == 65536 ? FbGeom<16>::kTables : (n == 1048576 ? FbGeom<32>::kTables : 0); }

template <int R>
static cudaError_t fb_launch(int fmt, const void* iq, float* power, cpx* A, const cpx* tables, int n_frames,
                             cudaStream_t st) {
    typedef FbGeom<R> G;
    const dim3 grid(G::L / 16, (unsigned)n_frames);
    switch (fmt) {
        case FMT_CU8: k_fbA<R, FMT_CU8><<<grid, G::kThreads, FbSmem<R>::kA, st>>>(static_cast<const uint16_t*>(iq), A, tables); break;
        case FMT_CS16: k_fbA<R, FMT_CS16><<<grid, G::kThreads, FbSmem<R>::kA, st>>>(static_cast<const uint32_t*>(iq), A, tables); break;
        case FMT_CF32: k_fbA<R, FMT_CF32><<<grid, G::kThreads, FbSmem<R>::kA, st>>>(static_cast<const cpx*>(iq), A, tables); break;
        default: return cudaErrorInvalidValue;
    }
    // persistent rows kernel: as many CTAs as fit (R = 32: one 128 KiB tile per SM; R = 16: 32 KiB tiles, several per SM)
    static int ctas_per_sm = 0, sms = 0;
    if (!ctas_per_sm) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, k_fbB<R>, G::kThreads, FbSmem<R>::kB) != cudaSuccess ||
            ctas_per_sm < 1)
            ctas_per_sm = 1;
    }
    const int n_tiles = (G::L / 16) * n_frames;
    const int gridB = n_tiles < ctas_per_sm * sms ? n_tiles : ctas_per_sm * sms;
    k_fbB<R><<<gridB, G::kThreads, FbSmem<R>::kB, st>>>(A, power, tables, n_tiles);
    return cudaGetLastError();
}

// Frames per group (<= max_group, the L2 budget) whose row-group count fills whole waves of the persistent
// rows kernel best: the kernel runs ceil(tiles / slots) rounds, so a group just over a multiple of `slots`
// wastes most of a round (1M points, 12 frames: 768 tiles on 148 slots = 5.19 -> 6 rounds; 9 frames: 3.89 -> 4).
template <int R>
static int fb_best_group(int max_group) {
    typedef FbGeom<R> G;
    int dev = 0, sms = 148, cps = 1;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cps, k_fbB<R>, G::kThreads, FbSmem<R>::kB) != cudaSuccess || cps < 1) cps = 1;
    const int slots = sms * cps, per_frame = G::L / 16;
    int best = max_group;
    double best_eff = 0.0;
    for (int g = max_group; g >= 1 && g >= max_group / 2; g--) {
        const int tiles = g * per_frame, rounds = (tiles + slots - 1) / slots;
        const double eff = (double)tiles / ((double)rounds * slots);
        if (eff > best_eff + 1e-9) { best_eff = eff; best = g; }
    }
    return best;
}
int fft_big_best_group(int n, int max_group) {
    if (max_group <= 1) return 1;
    return n == 65536 ? fb_best_group<16>(max_group) : (n == 1048576 ? fb_best_group<32>(max_group) : max_group);
}

// n_frames <= the number of frames the intermediate A holds (the caller loops over groups).
cudaError_t launch_fft_big(int fmt, const void* iq, float* power, cpx* A, const cpx* tables, int n, int n_frames,
                           cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    if (n_frames > 65535) return cudaErrorInvalidValue;
    if (n == 65536) return fb_launch<16>(fmt, iq, power, A, tables, n_frames, st);
    if (n == 1048576) return fb_launch<32>(fmt, iq, power, A, tables, n_frames, st);
    return cudaErrorInvalidValue;
}

}  // namespace sdrb
