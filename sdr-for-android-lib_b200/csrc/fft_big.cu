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
#include <cstdlib>

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

// ---- fused persistent variant: both passes in ONE kernel, the intermediate in a small ring that stays in L2.
// Work items are handed out in order by a ticket counter: step s = the column tiles of frame s (pass A), then the
// row groups of frame s - delay (pass B).  A row group waits until all column tiles of its frame have been
// written (a counter per frame, release / acquire at gpu scope); a column tile of frame f waits until every row
// group of frame f - slots, which used the same ring slot, has been read.  Both waits point at EARLIER tickets,
// and a ticket is only ever held by a running CTA, so the kernel cannot deadlock whatever the number of CTAs
// resident.  With the delay at about the number of frames in flight the waits are almost never taken; an
// intermediate line is read back `delay` steps after it was written (tens of MB of traffic later, not the
// whole batch), so it is still in L2, and the ring is rewritten in place instead of streaming 8 B per sample
// out to DRAM and back.  The row-group tile comes in with 16-byte cp.async (generic proxy, L2 only): the
// writers are other CTAs of the same kernel, so the ordering is plain release / acquire.
template <int R>
struct FbFused {
    typedef FbGeom<R> G;
    static constexpr int kTiles = G::L / 16;                                     // per frame and pass
    static constexpr int kElems = G::kElemsA > G::kElemsB ? G::kElemsA : G::kElemsB;
    static constexpr int kSmem = (kElems + G::kTL) * 8 + 16;
};
__device__ __forceinline__ unsigned fb_ld_acquire(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fb_red_release(unsigned* p) {
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p) : "memory");
}

template <int R, int FMT>
__global__ void __launch_bounds__(FbGeom<R>::kThreads, R == 16 ? 4 : 1)
k_fb_fused(const typename Ingest2<FMT>::raw_t* __restrict__ in, cpx* __restrict__ ring, float* __restrict__ out,
           const cpx* __restrict__ tables, unsigned* __restrict__ ctl, int n_frames, int delay, int slots) {
    typedef FbGeom<R> G;
    typedef FbFused<R> F;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx* S = reinterpret_cast<cpx*>(smem_raw);
    cpx* TL = S + F::kElems;
    int* s_item = reinterpret_cast<int*>(TL + G::kTL);
    const int t = threadIdx.x;
    constexpr size_t N = (size_t)G::L * G::L;
    for (int i = t; i < G::kTL; i += G::kThreads) TL[i] = tables[i];
    const cpx* TLO = tables + G::kTL;
    const cpx* THI = TLO + G::kTLO;
    const cpx* TB = THI + G::kTHI;
    unsigned* adone = ctl + 1;
    unsigned* bdone = ctl + 1 + n_frames;
    const int n_items = (n_frames + delay) * 2 * F::kTiles;
    // thread 0 draws the NEXT ticket while the current item runs (the atomic's round trip hides under the work)
    if (t == 0) *s_item = (int)atomicAdd(ctl, 1u);
    for (;;) {
        __syncthreads();                         // the ticket is there; the previous item is done with S
        const int item = *s_item;
        if (item >= n_items) break;
        unsigned nxt = 0;
        if (t == 0) nxt = atomicAdd(ctl, 1u);
        __syncthreads();                         // everyone has read the ticket
        struct Publish {                         // the next ticket goes to shared memory when the item ends
            int* slot; unsigned v; bool on;
            __device__ ~Publish() { if (on) *slot = (int)v; }
        } publish{s_item, nxt, t == 0};
        const int step = item / (2 * F::kTiles), r = item - step * 2 * F::kTiles;
        if (r < F::kTiles) {
            const int frame = step;
            if (frame >= n_frames) continue;
            if (frame >= slots) {                // the ring slot's previous user has been read
                if (t == 0)
                    while (fb_ld_acquire(bdone + frame - slots) < (unsigned)F::kTiles) __nanosleep(64);
                __syncthreads();
            }
            const int n2_0 = r * 16;
            fbA_pass0<R, FMT>(t, in + (size_t)frame * N, n2_0, TL, S);
            __syncthreads();
            fbA_pass1<R>(t, S, TLO, THI, TB, n2_0, ring + (size_t)(frame % slots) * N);
            __syncthreads();                     // every thread's stores are issued ...
            if (t == 0) {
                __threadfence();                 // ... and ordered before the count (cumulativity through the barrier)
                fb_red_release(adone + frame);
            }
        } else {
            const int frame = step - delay, g = r - F::kTiles;
            if (frame < 0 || frame >= n_frames) continue;
            if (t == 0)
                while (fb_ld_acquire(adone + frame) < (unsigned)F::kTiles) __nanosleep(64);
            __syncthreads();
            const cpx* src = ring + (size_t)(frame % slots) * N + (size_t)g * G::kElemsB;
            for (int i = t; i < G::kElemsB / 2; i += G::kThreads)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(fb_smem_u32(S + 2 * i)), "l"(src + 2 * i) : "memory");
            asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
            __syncthreads();
            if (t == 0 && frame + slots < n_frames) fb_red_release(bdone + frame);   // the slot may be rewritten
            fbB_pass0<R>(t, S, TL);
            __syncthreads();
            fbB_pass1<R>(t, S, g * 16, out + (size_t)frame * N);
        }
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
    if ((e = cudaFuncSetAttribute(k_fbB<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, FbSmem<R>::kB)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fb_fused<R, FMT_CU8>, cudaFuncAttributeMaxDynamicSharedMemorySize, FbFused<R>::kSmem)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fb_fused<R, FMT_CS16>, cudaFuncAttributeMaxDynamicSharedMemorySize, FbFused<R>::kSmem)) != cudaSuccess) return e;
    return cudaFuncSetAttribute(k_fb_fused<R, FMT_CF32>, cudaFuncAttributeMaxDynamicSharedMemorySize, FbFused<R>::kSmem);
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

// ---- fused variant: ring of `slots` frames, ctl = 1 + 2 n_frames counters (zeroed here, on the stream)
template <int R, int FMT>
static cudaError_t fb_launch_fused_f(const void* iq, float* power, cpx* ring, int slots, unsigned* ctl, const cpx* tables,
                                     int n_frames, int delay, cudaStream_t st) {
    typedef FbGeom<R> G;
    typedef FbFused<R> F;
    static int slots_hw = 0;
    if (!slots_hw) {
        int dev = 0, sms = 148, cps = 1;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cps, k_fb_fused<R, FMT>, G::kThreads, F::kSmem) != cudaSuccess || cps < 1) cps = 1;
        slots_hw = sms * cps;
    }
    const long long items = (long long)(n_frames + delay) * 2 * F::kTiles;
    const int grid = (int)(items < slots_hw ? items : slots_hw);
    cudaError_t e = cudaMemsetAsync(ctl, 0, (1 + 2 * (size_t)n_frames) * sizeof(unsigned), st);
    if (e != cudaSuccess) return e;
    k_fb_fused<R, FMT><<<grid, G::kThreads, F::kSmem, st>>>(static_cast<const typename Ingest2<FMT>::raw_t*>(iq), ring, power,
                                                             tables, ctl, n_frames, delay, slots);
    return cudaGetLastError();
}
template <int R>
static cudaError_t fb_launch_fused(int fmt, const void* iq, float* power, cpx* ring, int slots, unsigned* ctl,
                                   const cpx* tables, int n_frames, cudaStream_t st) {
    // delay: about the frames in flight (R = 16: 888 CTAs over 32 tiles per step; R = 32: 148 over 128)
    const char* de = getenv("SDRB_FB_DELAY");   // test hook: a short delay and a small ring force both waits
    const int delay_env = de ? atoi(de) : 0;
    int delay = delay_env > 0 ? delay_env : (R == 16 ? 32 : 3);
    if (delay > n_frames) delay = n_frames;
    if (delay > slots - 1) delay = slots - 1;      // a column tile may only wait on an EARLIER ticket
    if (delay < 1 || slots < 2) return cudaErrorInvalidValue;
    switch (fmt) {
        case FMT_CU8: return fb_launch_fused_f<R, FMT_CU8>(iq, power, ring, slots, ctl, tables, n_frames, delay, st);
        case FMT_CS16: return fb_launch_fused_f<R, FMT_CS16>(iq, power, ring, slots, ctl, tables, n_frames, delay, st);
        case FMT_CF32: return fb_launch_fused_f<R, FMT_CF32>(iq, power, ring, slots, ctl, tables, n_frames, delay, st);
        default: return cudaErrorInvalidValue;
    }
}
int fft_big_ring_slots(int n) {
    const char* e = getenv("SDRB_FB_SLOTS");
    const int env = e ? atoi(e) : 0;
    if (env >= 2) return env;
    return n == 65536 ? 80 : 8;
}
cudaError_t launch_fft_big_fused(int fmt, const void* iq, float* power, cpx* ring, int slots, unsigned* ctl,
                                 const cpx* tables, int n, int n_frames, cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    if (n == 65536) return fb_launch_fused<16>(fmt, iq, power, ring, slots, ctl, tables, n_frames, st);
    if (n == 1048576) return fb_launch_fused<32>(fmt, iq, power, ring, slots, ctl, tables, n_frames, st);
    return cudaErrorInvalidValue;
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
