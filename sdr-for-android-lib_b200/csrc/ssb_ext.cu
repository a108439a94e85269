// ssb_ext.cu — the EXTENDED SSB chain: NCO channel bank and true sideband selection.
//
// Extension stages (sdrb_config.ssbChannels / ssbChannelOffsetHz / ssbChannelSpacingHz /
// ssbTrueSideband).  The reference demodulates one channel at the tuned centre and its USB/LSB
// switch degenerates to USB = 2y, LSB = 0 (src/ssb/ssb_demod_opt.cpp:82,93); these kernels add
//   * a per-channel NCO mix (carried 32-bit fixed-point phase = inc * samples-since-reset), and
//   * Weaver sideband selection: mix to the sideband centre, low-pass I and Q at B/2, mix back
//     by +/- B/2 and take the real part,
// in front of / around the reference's own chain (removeDC, IIR2, AGC, FIR decimate, HP/BP,
// transient boost, PCM), whose arithmetic and order are kept.  Definition and checker:
// oracle/sdr_oracle_ext.c:sdr_oracle_process_ssbx.  With every knob at identity the library
// never comes here (ssb.cu is the reference-exact path).
//
// Structure (round 2).  Only the IIR2 carries state from sample to sample across blocks; everything in front of it is
// either per block (removeDC, shared by every channel of a stream) or closed-form per sample (the NCO phase is
// inc * samples-since-reset), so it is hoisted out of the sequential lane:
//   k_ssbx_dc    : one lane per (stream, block): removeDC (complex, a 2-operation chain) -> X [stream][block][n] cf32
//   k_ssbx_mix   : one thread per 4 samples of a (stream, channel): NCO + first mix -> M [stream][channel][block][n]
//                  (re only; with true sideband planar: the re row of a unit, then its im row)
//   k_ssbx_state : one lane per CHAIN (the I filter of a (stream, channel), and its Q filter with true sideband),
//                  blocks in order: IIR2 on M only (9 operations per sample, a 20-cycle chain) -- it was 64
//                  instructions and 89 cycles per sample with DC, NCO and mix on the lane -- records the exact filter
//                  state at every block start
//   k_ssbx_agc   : one lane per (stream, channel, block): IIR2 from that state, second mix (true sideband), demod,
//                  AGC in sweeps -> post-AGC samples; k_ssb_fir (ssb.cu) decimates; k_ssb_audio: HP/BP, boost, PCM
// PCM layout: [stream][channel][block][m].  The block length must be a multiple of 32 samples.
#include <cstdlib>

#include "internal.h"
#include "nco.cuh"
#include "ssb_common.cuh"

namespace sdrb {

// Warp-level streamer whose 32 rows are arbitrary byte ranges (one pointer per lane's row).
template <int NST>
struct TileStreamX {
    const unsigned char* const* rowptr;   // this warp's 32 row pointers (shared memory)
    unsigned char* tile;                  // this warp's tiles
    int lane;
    const unsigned char* rp[4];           // the four rows this lane copies, in registers (bind())
    size_t stride = SSB_ROW_BYTES;        // source bytes between consecutive chunks of a row
    // Fetching the pointers from shared memory at every chunk put four dependent LDS.64 -> LDGSTS pairs (all through
    // one register, so in series) in front of every chunk: ~300 cycles on the lone in-order warp of a state pass.
    __device__ __forceinline__ void bind() {
        const int sub = lane & 3, r0 = lane >> 2;
#pragma unroll
        for (int k = 0; k < 4; k++) rp[k] = rowptr[r0 + 8 * k] + sub * 16;
    }
    __device__ __forceinline__ void issue(int chunk, int stage) {
        const int sub = lane & 3, r0 = lane >> 2;
        unsigned char* t = tile + stage * SSB_TILE_BYTES + sub * 16;
#pragma unroll
        for (int k = 0; k < 4; k++)
            cp_async16(t + (r0 + 8 * k) * SSB_ROW_PITCH, rp[k] + (size_t)chunk * stride);
        cp_async_commit();
    }
    __device__ __forceinline__ const uint4* row(int stage) const {
        return reinterpret_cast<const uint4*>(tile + stage * SSB_TILE_BYTES + lane * SSB_ROW_PITCH);
    }
    __device__ __forceinline__ void prologue(int nchunks) {
        for (int k = 0; k < NST - 1; k++) {
            if (k < nchunks) issue(k, k);
            else cp_async_commit();
        }
    }
    __device__ __forceinline__ void advance(int c, int nchunks) {
        const int ahead = c + NST - 1;
        if (ahead < nchunks) issue(ahead, ahead % NST);
        else cp_async_commit();
        cp_async_wait<NST - 1>();
    }
    __device__ __forceinline__ const uint4* cur(int c) const { return row(c % NST); }
};

constexpr int SSBX_WARPS = 4;
// cp.async stages of the state pass.  One warp per CTA and per SM pulls 2 KB per chunk and spends only 160-320 cycles
// of arithmetic on it (8 or 16 float samples per chain and chunk, against 32 for u8 input in ssb.cu): with eight stages
// the 14 KB in flight per SM did not cover the DRAM round trip and the walk ran at ~336 cycles per chunk whatever the
// chunk held (measured: 41 cycles per sample without, 62 with true sideband, against a 20-cycle chain).
constexpr int SSBX_STATE_STAGES = 32;   // a power of two: the stage index is a mask, not a division (80 KiB)

// A 16-byte shared-memory load that stays where it is written: left to the compiler, the loads of an iteration sink
// to their first use and the single warp of a state-pass CTA then waits a shared-memory round trip every few samples
// (measured: 36 cycles per sample against the chain's 20).
__device__ __forceinline__ float4 lds128_now(const float4* p) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "r"((unsigned)__cvta_generic_to_shared(p)));
    return v;
}

__device__ __forceinline__ void named_bar_sync_x(int id, int count) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void named_bar_arrive_x(int id, int count) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// iir2Process :75-84 (previous OUTPUTS as state), one sample
__device__ __forceinline__ float ssbx_iir(const BiquadCoef& c, float x, float& z1, float& z2) {
    float y = __fmul_rn(c.a0, x);
    y = __fadd_rn(y, __fmul_rn(c.a1, z1));
    y = __fadd_rn(y, __fmul_rn(c.a2, z2));
    y = __fsub_rn(y, __fmul_rn(c.b1, z1));
    y = __fsub_rn(y, __fmul_rn(c.b2, z2));
    z2 = z1;
    z1 = y;
    return y;
}

// ---------------------------------------------------------------- pre-pass 1: removeDC per (stream, block)
// removeDC (src/ssb/ssb_demod_opt.cpp:49-55) on the complex sample, restarted per block; lanes = frames.
template <int FMT>
__global__ void __launch_bounds__(SSBX_WARPS * 32)
k_ssbx_dc(const void* __restrict__ iq, const __grid_constant__ SsbParams P, float2* __restrict__ X, int n_frames) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int NS = 4 * SsbIngest<FMT>::kPerVec;
    constexpr int OP = 2 * NS + 4;                    // output tile row pitch (floats)
    __shared__ __align__(16) float s_out[SSBX_WARPS][32 * OP];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    TileStreamT<SSB_MAIN_STAGES> ts;
    ts.lane = lane;
    ts.base = static_cast<const unsigned char*>(iq);
    ts.row_bytes = (size_t)P.n * sizeof(typename Ingest<FMT>::raw_t);
    ts.first_row = ((size_t)blockIdx.x * SSBX_WARPS + warp) * 32;
    ts.last_row = (size_t)n_frames - 1;
    ts.tile = smem + warp * SSB_MAIN_STAGES * SSB_TILE_BYTES;
    if (ts.first_row >= (size_t)n_frames) return;
    const int nchunks = (int)(ts.row_bytes / SSB_ROW_BYTES);
    const float oma = __fsub_rn(1.0f, P.dcAlpha);
    float dcr = 0.0f, dci = 0.0f;
    float* tile = s_out[warp];
    constexpr int LPR = 2 * NS / 4, RPI = 32 / LPR;   // lanes per output row, rows per store instruction
    ts.prologue(nchunks);
    for (int c = 0; c < nchunks; c++) {
        ts.advance(c, nchunks);
        __syncwarp();
        const uint4* row = ts.cur(c);
        float4* mine = reinterpret_cast<float4*>(tile + lane * OP);
#pragma unroll
        for (int v = 0; v < 4; v++) {
            const uint4 d = row[v];
            float o[2 * SsbIngest<FMT>::kPerVec];
#pragma unroll
            for (int j = 0; j < SsbIngest<FMT>::kPerVec; j++) {
                float I, Q;
                SsbIngest<FMT>::get(d, j, I, Q);
                dcr = __fadd_rn(__fmul_rn(P.dcAlpha, dcr), __fmul_rn(oma, I));
                dci = __fadd_rn(__fmul_rn(P.dcAlpha, dci), __fmul_rn(oma, Q));
                o[2 * j] = __fsub_rn(I, dcr);
                o[2 * j + 1] = __fsub_rn(Q, dci);
            }
#pragma unroll
            for (int k = 0; k < SsbIngest<FMT>::kPerVec / 2; k++)
                mine[v * (SsbIngest<FMT>::kPerVec / 2) + k] = make_float4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < LPR; i++) {
            const int r = i * RPI + lane / LPR, col = lane % LPR;
            const size_t fr = ts.first_row + r;
            if (fr <= ts.last_row)
                *reinterpret_cast<float4*>(reinterpret_cast<float*>(X) + (fr * (size_t)P.n + (size_t)c * NS) * 2 + 4 * col) =
                    *reinterpret_cast<const float4*>(tile + r * OP + 4 * col);
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------- pre-pass 2: NCO + first mix, sample-parallel
// m = (x_r + j x_i) * exp(-j theta): re = x_r c + x_i s, im = x_i c - x_r s, theta = inc1 * (samples since reset);
// the expressions and their order are those of oracle/sdr_oracle_ext.c.  span = n_blocks * n samples per stream in
// this call.
// A thread takes 4 consecutive samples of a stream and runs them through a group of up to SSBX_MIX_CPT channels: the
// samples are loaded once, a channel's phase is one 32-bit multiply (inc * samples-since-reset, modulo 2^32), and no
// index is ever divided (the first version derived (unit, sample) from a flat index with two 64-bit divisions per
// four samples - more instructions than the NCO itself).  grid.x = (stream, channel group), grid.y strides over the
// span in blocks of 1024 samples.
constexpr int SSBX_MIX_CPT = 16;
template <bool TSB>
__global__ void __launch_bounds__(256)
k_ssbx_mix(const float2* __restrict__ X, const uint32_t* __restrict__ inc1_tab, float* __restrict__ M, int channels,
           size_t span, int groups, unsigned long long base_samples) {
    const unsigned s = blockIdx.x / (unsigned)groups, cg = blockIdx.x - s * (unsigned)groups;
    const int c0 = (int)cg * SSBX_MIX_CPT, c1 = min(channels, c0 + SSBX_MIX_CPT);
    for (size_t k = ((size_t)blockIdx.y * 256 + threadIdx.x) * 4; k < span; k += (size_t)gridDim.y * 1024) {
        const float4* x4 = reinterpret_cast<const float4*>(X + (size_t)s * span + k);
        const float4 a = x4[0], b = x4[1];
        const float xr[4] = {a.x, a.z, b.x, b.z}, xi[4] = {a.y, a.w, b.y, b.w};
        const uint32_t t0 = (uint32_t)(base_samples + k);
        for (int c = c0; c < c1; c++) {
            const uint32_t inc1 = inc1_tab[c];
            uint32_t ph = inc1 * t0;
            const size_t v = (size_t)s * channels + c;
            float mr[4], mi[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                float cs, sn;
                nco_cs(ph, cs, sn);
                ph += inc1;
                mr[j] = __fadd_rn(__fmul_rn(xr[j], cs), __fmul_rn(xi[j], sn));
                if (TSB) mi[j] = __fsub_rn(__fmul_rn(xi[j], cs), __fmul_rn(xr[j], sn));
            }
            if (TSB) {   // planar: the I row of a unit, then its Q row -- every chain of the state pass owns a contiguous row
                *reinterpret_cast<float4*>(M + (2 * v) * span + k) = make_float4(mr[0], mr[1], mr[2], mr[3]);
                *reinterpret_cast<float4*>(M + (2 * v + 1) * span + k) = make_float4(mi[0], mi[1], mi[2], mi[3]);
            } else {
                *reinterpret_cast<float4*>(M + v * span + k) = make_float4(mr[0], mr[1], mr[2], mr[3]);
            }
        }
    }
}

// ---------------------------------------------------------------- pass 1: the carried IIR2 state, block by block
// One lane per CHAIN: a (stream, channel) has one (I) or, with true sideband, two (I and Q: two lanes reading the same
// row of pairs) -- a warp issues in order, so two chains in one lane would queue behind each other's dependent
// operations.  The lane walks M[v][0 .. n_blocks * n) in order, one 64-byte chunk (16 or 8 samples of its chain) per
// iteration, the chunk's four shared-memory loads ahead of the arithmetic.
template <bool TSB>
__global__ void __launch_bounds__(32)
k_ssbx_state(const float* __restrict__ M, const __grid_constant__ SsbParams P, float4* __restrict__ start,
             SsbStreamState* __restrict__ state, size_t n_units, int n_blocks) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ const unsigned char* rows[32];
    constexpr int E = TSB ? 2 : 1;
    const int lane = threadIdx.x;
    const size_t n_chains = n_units * E;
    const size_t ch = (size_t)blockIdx.x * 32 + lane;
    const bool live = ch < n_chains;
    const size_t chc = live ? ch : n_chains - 1;
    const size_t v = chc / E;
    const int comp = (int)(chc % E);
    const size_t block_bytes = (size_t)P.n * sizeof(float);                 // of one chain: M is planar (I row, Q row)
    rows[lane] = reinterpret_cast<const unsigned char*>(M) + chc * block_bytes * n_blocks;
    __syncwarp();
    TileStreamX<SSBX_STATE_STAGES> ts = {rows, smem, lane};
    ts.bind();
    const int chunks_per_block = (int)(block_bytes / SSB_ROW_BYTES);
    const int nchunks = chunks_per_block * n_blocks;
    float z1 = comp ? state[v].pad0 : state[v].rf_z1, z2 = comp ? state[v].pad1 : state[v].rf_z2;
    // float4 per (unit, block) = (I z1, I z2, Q z1, Q z2); the pointer walks along: fetching `start` from the constant
    // bank and forming the address under a predicate cost ~40 cycles in every chunk (ncu), stored or not
    float2* sp = reinterpret_cast<float2*>(start) + (v * n_blocks) * 2 + comp;
    // software pipeline: the registers of chunk c + 1 are loaded while chunk c is in the chain (a load that is only
    // used in the next iteration cannot sink to its first use, which is where ptxas puts the loads of a chunk otherwise)
    ts.prologue(nchunks);
    int cb = 0;
    float4 nx[4];
    ts.advance(0, nchunks);
    __syncwarp();
    {
        const float4* row = reinterpret_cast<const float4*>(ts.cur(0));
#pragma unroll
        for (int q = 0; q < 4; q++) nx[q] = lds128_now(row + q);
    }
    __syncwarp();
    for (int c = 0; c < nchunks; c++) {
        float4 d[4];
#pragma unroll
        for (int q = 0; q < 4; q++) d[q] = nx[q];
        if (c + 1 < nchunks) {
            ts.advance(c + 1, nchunks);     // reuses the stage of chunk c, whose values are in registers already
            __syncwarp();
            const float4* row = reinterpret_cast<const float4*>(ts.cur(c + 1));
#pragma unroll
            for (int q = 0; q < 4; q++) nx[q] = lds128_now(row + q);
        }
        if (cb == 0) {            // warp-uniform: a branch, taken once per block
            if (live) {
                sp[0] = make_float2(z1, z2);
                if (!TSB) sp[1] = make_float2(0.0f, 0.0f);
            }
            sp += 2;
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
            ssbx_iir(P.rf, d[q].x, z1, z2); ssbx_iir(P.rf, d[q].y, z1, z2);
            ssbx_iir(P.rf, d[q].z, z1, z2); ssbx_iir(P.rf, d[q].w, z1, z2);
        }
        __syncwarp();
        if (++cb == chunks_per_block) cb = 0;
    }
    if (live) {
        if (comp) { state[v].pad0 = z1; state[v].pad1 = z2; }
        else { state[v].rf_z1 = z1; state[v].rf_z2 = z2; }
    }
}

// ---------------------------------------------------------------- pass 1 as a producer / chain warp pair
// The same walk with the roles of ssb.cu's k_ssb_state: warp 0 only moves data (cp.async from M into a ring of G groups
// of K chunks), warp 1 only runs the chains (a group's K x 64 bytes per lane into registers, then 9 operations per
// sample).  ncu on the one-warp version above: the chain itself at its 20 cycles per sample, plus ~17 cycles per
// sample of copy issue, waits and loop overhead that a lone in-order warp cannot hide.  Named barriers hand the ring
// slots over (full: ids 1..G, empty: ids 1+G..2G; bar.arrive by the side that hands over, bar.sync by the side that
// waits).  Used when the chunk count is a multiple of K.
constexpr int SSBX_RING_G = 4;      // ring slots (groups)
constexpr int SSBX_RING_D = 2;      // groups the producer keeps in flight behind the one it is issuing

template <bool TSB, int K>
__global__ void __launch_bounds__(64)
k_ssbx_state2(const float* __restrict__ M, const __grid_constant__ SsbParams P, float4* __restrict__ start,
              SsbStreamState* __restrict__ state, size_t n_units, int n_blocks) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int E = TSB ? 2 : 1, G = SSBX_RING_G, D = SSBX_RING_D;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t n_chains = n_units * E;
    const size_t ch0 = (size_t)blockIdx.x * 32;
    const size_t block_bytes = (size_t)P.n * sizeof(float);                 // of one chain: M is planar (I row, Q row)
    const int chunks_per_block = (int)(block_bytes / SSB_ROW_BYTES);
    const int ngroups = chunks_per_block * n_blocks / K;
    if (warp == 0) {
        // ---- producer: rows r0 + 8 j of the tile belong to chains ch0 + r0 + 8 j
        const int sub = lane & 3, r0 = lane >> 2;
        const unsigned char* rp[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            size_t ch = ch0 + r0 + 8 * j;
            if (ch >= n_chains) ch = n_chains - 1;
            rp[j] = reinterpret_cast<const unsigned char*>(M) + ch * block_bytes * n_blocks + sub * 16;
        }
        for (int gi = 0; gi < ngroups + D; gi++) {
            if (gi < ngroups) {
                const int slot = gi & (G - 1);
                if (gi >= G) named_bar_sync_x(1 + G + slot, 64);            // the chain warp has emptied the slot
                unsigned char* t = smem + (size_t)slot * K * SSB_TILE_BYTES + sub * 16;
#pragma unroll
                for (int k = 0; k < K; k++)
#pragma unroll
                    for (int j = 0; j < 4; j++)
                        cp_async16(t + k * SSB_TILE_BYTES + (r0 + 8 * j) * SSB_ROW_PITCH,
                                   rp[j] + ((size_t)gi * K + k) * SSB_ROW_BYTES);
            }
            cp_async_commit();                                              // one commit group per gi, empty past the end
            if (gi >= D) {
                cp_async_wait<D>();                                         // group gi - D has landed
                __syncwarp();
                __threadfence_block();
                named_bar_arrive_x(1 + ((gi - D) & (G - 1)), 64);           // slot full
            }
        }
    } else {
        // ---- chain.  Everything that is not one of the nine operations per sample costs this in-order warp issue
        // cycles the chain cannot use (ncu, round 2: 9.5 of its 28.4 cycles per sample were the per-chunk block-start
        // test with its convergence barriers, the re / im select per sample, constant-bank reloads of the
        // coefficients and the wait for the group's shared-memory loads).  So: blocks are the outer loop (a block is a
        // whole number of group pairs: checked by the launcher), M is planar so a Q chain reads its own row (no select,
        // and a scalar load per sample would cost ~4 issue cycles each), the NEXT group's values are loaded before the
        // current group's chain runs
        // (two register sets, loop unrolled by two), and the coefficients come from shared memory once.
        __shared__ float s_coef[5];
        if (lane == 0) { s_coef[0] = P.rf.a0; s_coef[1] = P.rf.a1; s_coef[2] = P.rf.a2; s_coef[3] = P.rf.b1; s_coef[4] = P.rf.b2; }
        __syncwarp();
        const BiquadCoef rf = {s_coef[0], s_coef[1], s_coef[2], s_coef[3], s_coef[4]};
        const size_t ch = ch0 + lane;
        const bool live = ch < n_chains;
        const size_t chc = live ? ch : n_chains - 1;
        const size_t v = chc / E;
        const int comp = (int)(chc % E);
        float z1 = comp ? state[v].pad0 : state[v].rf_z1, z2 = comp ? state[v].pad1 : state[v].rf_z2;
        float2* sp = reinterpret_cast<float2*>(start) + (v * n_blocks) * 2 + comp;
        constexpr int NV = K * 16;                                          // samples of a group on this lane
        const unsigned lane_base = (unsigned)__cvta_generic_to_shared(smem) + lane * SSB_ROW_PITCH;
        auto take = [&](int gi, float (&x)[NV]) {                           // group gi: wait, load, hand the slot back
            const int slot = gi & (G - 1);
            named_bar_sync_x(1 + slot, 64);                                 // slot full
            const unsigned t = lane_base + (unsigned)slot * K * SSB_TILE_BYTES;
#pragma unroll
            for (int k = 0; k < K; k++) {
#pragma unroll
                for (int q = 0; q < 4; q++)
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                                 : "=f"(x[16 * k + 4 * q]), "=f"(x[16 * k + 4 * q + 1]), "=f"(x[16 * k + 4 * q + 2]),
                                   "=f"(x[16 * k + 4 * q + 3])
                                 : "r"(t + k * SSB_TILE_BYTES + 16 * q));
            }
            if (gi + G < ngroups) {
                __threadfence_block();                                      // the loads are done before the slot is handed back
                named_bar_arrive_x(1 + G + slot, 64);
            }
        };
        auto run = [&](const float (&x)[NV]) {
#pragma unroll
            for (int i = 0; i < NV; i++) ssbx_iir(rf, x[i], z1, z2);
        };
        const int pairs_per_block = chunks_per_block / (2 * K);
        float xa[NV], xb[NV];
        int gi = 0;
        take(0, xa);
        for (int b = 0; b < n_blocks; b++) {
            if (live) {
                sp[0] = make_float2(z1, z2);                                // the exact state at the block's first sample
                if (!TSB) sp[1] = make_float2(0.0f, 0.0f);
            }
            sp += 2;
            for (int pr = 0; pr < pairs_per_block; pr++) {
                take(gi + 1, xb);                                           // (a block is a whole number of pairs)
                run(xa);
                if (gi + 2 < ngroups) take(gi + 2, xa);
                run(xb);
                gi += 2;
            }
        }
        if (live) {
            if (comp) { state[v].pad0 = z1; state[v].pad1 = z2; }
            else { state[v].rf_z1 = z1; state[v].rf_z2 = z2; }
        }
    }
}

// ---------------------------------------------------------------- pass 2: IIR2 from the recorded state, second mix, AGC
// One lane per (stream, channel, block).  True sideband: Weaver second mix Re{(yI + j yQ) exp(+/- j theta2)}, doubled;
// otherwise the reference's demodSSB on {y, y} (USB = y + y, LSB = y - y).  AGC in sweeps (ssb_common.cuh).
// SP > 0: the 255-tap decimator runs on the lane as well (LaneDecim<SP>, the fused pass of ssb.cu) and the audio-rate
// samples go straight to a48 [unit][block][m]; SP = 0 (decimation factors the fused pass does not take): the
// post-AGC samples go to xagc [unit][block][n] through a shared-memory tile and k_ssb_fir (ssb.cu) decimates them.
template <bool TSB, bool CHECK, int SP>
__global__ void __launch_bounds__(SSBX_WARPS * 32)
k_ssbx_agc(const float* __restrict__ M, const __grid_constant__ SsbParams P, uint32_t inc2,
           const float4* __restrict__ start, float* __restrict__ xagc, const float* __restrict__ tapsE, int rows,
           float* __restrict__ a48, size_t n_frames, int n_blocks, unsigned long long base_samples) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int E = TSB ? 2 : 1, NS = 16 / E;       // samples per 64-byte chunk of M
    constexpr int OP = NS + 4;
    __shared__ const unsigned char* rows_all[SSBX_WARPS][32];
    __shared__ __align__(16) float s_out[SSBX_WARPS][32 * OP];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t vf0 = ((size_t)blockIdx.x * SSBX_WARPS + warp) * 32;
    const size_t vf = vf0 + lane;
    const bool live = vf < n_frames;
    const size_t vfc = live ? vf : n_frames - 1;
    const size_t b = vfc % n_blocks;
    const size_t block_bytes = (size_t)P.n * E * sizeof(float);   // input bytes per (unit, block)
    // true sideband: M is planar per unit (I row, then Q row, each n_blocks * n floats); the lane's I block starts at
    // row 2 u, block b = element (2 vf - b) n; its Q block lies one row (n_blocks * n floats) further
    rows_all[warp][lane] = reinterpret_cast<const unsigned char*>(M) + (TSB ? 2 * vfc - b : vfc) * (size_t)P.n * sizeof(float);
    float* hT = reinterpret_cast<float*>(smem + SSBX_WARPS * SSB_MAIN_STAGES * SSB_TILE_BYTES);
    if (SP > 0)
        for (int i = threadIdx.x; i < SP * rows * SP; i += blockDim.x) hT[i] = tapsE[i];
    __syncthreads();
    if (vf0 >= n_frames) return;
    TileStreamX<SSB_MAIN_STAGES> ts = {rows_all[warp], smem + warp * SSB_MAIN_STAGES * SSB_TILE_BYTES, lane};
    ts.bind();
    if (TSB) {   // a chunk = 32 bytes of the I row (tile bytes 0..31) + 32 bytes of the Q row (32..63)
        const int sub = lane & 3, r0 = lane >> 2;
        const size_t qoff = (size_t)n_blocks * P.n * sizeof(float);
#pragma unroll
        for (int k = 0; k < 4; k++) ts.rp[k] = rows_all[warp][r0 + 8 * k] + (sub >> 1) * qoff + (sub & 1) * 16;
        ts.stride = SSB_ROW_BYTES / 2;
    }
    const int nchunks = (int)(block_bytes / SSB_ROW_BYTES);
    const float4 s0 = start[vfc];
    float iz1 = s0.x, iz2 = s0.y, qz1 = s0.z, qz2 = s0.w;
    uint32_t ph2 = (uint32_t)((unsigned long long)inc2 * (base_samples + (unsigned long long)b * (unsigned long long)P.n));
    float gain = 1.0f;
    const float omFast = __fsub_rn(1.0f, P.agcFast), omSlow = __fsub_rn(1.0f, P.agcSlow);
    float* tile = s_out[warp];
    constexpr int LPR = NS / 4, RPI = 32 / LPR;
    LaneDecim<(SP > 0 ? SP : 1)> dec;
    dec.init(P);
    float* out = a48 + vfc * (size_t)P.m;
    ts.prologue(nchunks);
    for (int c = 0; c < nchunks; c++) {
        ts.advance(c, nchunks);
        __syncwarp();
        const float4* row = reinterpret_cast<const float4*>(ts.cur(c));
        float4 dd[4];
#pragma unroll
        for (int q = 0; q < 4; q++) dd[q] = lds128_now(row + q);
        float a[NS];
        if (TSB) {
            const float xi[8] = {dd[0].x, dd[0].y, dd[0].z, dd[0].w, dd[1].x, dd[1].y, dd[1].z, dd[1].w};
            const float xq[8] = {dd[2].x, dd[2].y, dd[2].z, dd[2].w, dd[3].x, dd[3].y, dd[3].z, dd[3].w};
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const float yi = ssbx_iir(P.rf, xi[j], iz1, iz2);
                const float yq = ssbx_iir(P.rf, xq[j], qz1, qz2);
                float c2, s2;
                nco_cs(ph2, c2, s2);
                ph2 += inc2;
                const float p = __fmul_rn(yi, c2), w = __fmul_rn(yq, s2);
                const float u = P.upper ? __fsub_rn(p, w) : __fadd_rn(p, w);
                a[j < NS ? j : 0] = __fadd_rn(u, u);
            }
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const float4 d = dd[q];
            if (TSB) {
            } else {
                const float m4[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                for (int h = 0; h < 4; h++) {
                    const float y = ssbx_iir(P.rf, m4[h], iz1, iz2);
                    a[4 * q + h] = P.upper ? __fadd_rn(y, y) : __fsub_rn(y, y);
                }
            }
        }
        ssb_agc_sweeps<NS, CHECK>(a, P, omFast, omSlow, gain);
        if (SP > 0) {
            float v;
            int j;
            if (dec.template feed<NS>(a, c, hT, rows, P, v, j) && live && j < P.m) out[j] = v;
            __syncwarp();
            continue;
        }
        float4* mine = reinterpret_cast<float4*>(tile + lane * OP);
#pragma unroll
        for (int k = 0; k < LPR; k++) mine[k] = make_float4(a[4 * k], a[4 * k + 1], a[4 * k + 2], a[4 * k + 3]);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < LPR; i++) {
            const int r = i * RPI + lane / LPR, col = lane % LPR;
            const size_t fr = vf0 + r;
            if (fr < n_frames)
                *reinterpret_cast<float4*>(xagc + fr * (size_t)P.n + (size_t)c * NS + 4 * col) =
                    *reinterpret_cast<const float4*>(tile + r * OP + 4 * col);
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------- launch
// Scratch (all in `d_work`, floats): X = S * B * n * 2, then M = S * C * B * n * (1 or 2); d_agc = S * C * B * n.
size_t ssb_ext_work_floats(int n, int S, int C, int B, int true_sideband) {
    return (size_t)S * B * n * 2 + (size_t)S * C * B * n * (true_sideband ? 2 : 1);
}

template <int FMT, bool TSB>
static cudaError_t launch_ssbx_f(const void* iq, const SsbParams& P, const float* d_taps, const float* d_tapsE, int sp,
                                 const uint32_t* d_inc1, uint32_t inc2, float4* d_start, SsbStreamState* d_state,
                                 float* d_a48, float* d_agc, float* d_work, int S, int C, int B, unsigned long long base,
                                 cudaStream_t st) {
    const size_t nv = (size_t)S * C, nvf = nv * B, span = (size_t)B * P.n;
    float2* X = reinterpret_cast<float2*>(d_work);
    float* M = d_work + (size_t)S * span * 2;
    cudaError_t e;
    const size_t tsmem = (size_t)SSBX_WARPS * SSB_MAIN_STAGES * SSB_TILE_BYTES;
    if ((e = cudaFuncSetAttribute(k_ssbx_dc<FMT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem)) != cudaSuccess) return e;
    const int frames = S * B;
    k_ssbx_dc<FMT><<<(frames + SSBX_WARPS * 32 - 1) / (SSBX_WARPS * 32), SSBX_WARPS * 32, tsmem, st>>>(iq, P, X, frames);
    const int groups = (C + SSBX_MIX_CPT - 1) / SSBX_MIX_CPT;
    const size_t kblocks = (span + 1023) / 1024;
    const dim3 mgrid((unsigned)S * (unsigned)groups, (unsigned)(kblocks < 65535 ? kblocks : 65535));
    k_ssbx_mix<TSB><<<mgrid, 256, 0, st>>>(X, d_inc1, M, C, span, groups, base);
    const size_t chains = nv * (TSB ? 2 : 1);
    static const bool one_warp = getenv("SDRB_SSBX_STATE") && getenv("SDRB_SSBX_STATE")[0] == '1';
    const long long chunks_per_block = (long long)((size_t)P.n * sizeof(float) / SSB_ROW_BYTES);   // of one chain's row
    if (chunks_per_block % 8 == 0 && !one_warp) {   // a block = whole pairs of groups of K = 4 chunks
        constexpr int K = 4;
        k_ssbx_state2<TSB, K><<<(unsigned)((chains + 31) / 32), 64, (size_t)SSBX_RING_G * K * SSB_TILE_BYTES, st>>>(
            M, P, d_start, d_state, nv, B);
    } else {
        const size_t ssmem = (size_t)SSBX_STATE_STAGES * SSB_TILE_BYTES;
        if ((e = cudaFuncSetAttribute(k_ssbx_state<TSB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssmem)) != cudaSuccess) return e;
        k_ssbx_state<TSB><<<(unsigned)((chains + 31) / 32), 32, ssmem, st>>>(M, P, d_start, d_state, nv, B);
    }
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (P.m <= 0) return cudaSuccess;
    constexpr bool CHECK = FMT == FMT_CF32;
    const unsigned agrid = (unsigned)((nvf + SSBX_WARPS * 32 - 1) / (SSBX_WARPS * 32));
    const int rows = ssb_fused_rows(P);
#define SDRB_AGC(SPV)                                                                                                  \
    {                                                                                                                  \
        const size_t sm = tsmem + (size_t)(SPV) * rows * (SPV) * sizeof(float);                                        \
        if ((e = cudaFuncSetAttribute(k_ssbx_agc<TSB, CHECK, SPV>, cudaFuncAttributeMaxDynamicSharedMemorySize,        \
                                      (int)sm)) != cudaSuccess)                                                        \
            return e;                                                                                                  \
        k_ssbx_agc<TSB, CHECK, SPV><<<agrid, SSBX_WARPS * 32, sm, st>>>(M, P, inc2, d_start, d_agc, d_tapsE, rows, d_a48,  \
                                                                        nvf, B, base);                                 \
    }
    switch (sp) {
        case 0: SDRB_AGC(0) break;
        case 1: SDRB_AGC(1) break;
        case 2: SDRB_AGC(2) break;
        case 4: SDRB_AGC(4) break;
        default: return cudaErrorInvalidValue;   // ssb_ext_fused_slots() stops at 4
    }
#undef SDRB_AGC
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (sp > 0) return cudaSuccess;
    if (!d_agc) return cudaErrorMemoryAllocation;
    return launch_ssb_fir(d_agc, d_taps, d_a48, P, (int)nvf, st);
}

// d_work: ssb_ext_work_floats() floats; sp > 0 (ssb_ext_fused_slots): d_tapsE is the rotated tap table and d_agc is
// not used; sp = 0: d_agc [S * C * B][n] floats is required
cudaError_t launch_ssb_ext(int fmt, const void* iq, const SsbParams& P, const float* d_taps, const float* d_tapsE, int sp,
                           const uint32_t* d_inc1, uint32_t inc2, int true_sideband, float4* d_start,
                           SsbStreamState* d_state, float* d_a48, int16_t* d_pcm, float* d_agc, float* d_work,
                           int n_streams, int channels, int n_blocks, unsigned long long base_samples, cudaStream_t st,
                           int* launches) {
    if (n_streams <= 0 || channels <= 0 || n_blocks <= 0) return cudaSuccess;
    if (!d_work || (sp == 0 && !d_agc) || (sp > 0 && !d_tapsE)) return cudaErrorMemoryAllocation;
    if ((size_t)n_streams * channels * n_blocks >= (1ull << 31)) return cudaErrorInvalidValue;
    cudaError_t e;
#define SDRB_X(F)                                                                                                      \
    (true_sideband ? launch_ssbx_f<F, true>(iq, P, d_taps, d_tapsE, sp, d_inc1, inc2, d_start, d_state, d_a48, d_agc,  \
                                            d_work, n_streams, channels, n_blocks, base_samples, st)                   \
                   : launch_ssbx_f<F, false>(iq, P, d_taps, d_tapsE, sp, d_inc1, inc2, d_start, d_state, d_a48, d_agc, \
                                             d_work, n_streams, channels, n_blocks, base_samples, st))
    switch (fmt) {
        case FMT_CU8: e = SDRB_X(FMT_CU8); break;
        case FMT_CS16: e = SDRB_X(FMT_CS16); break;
        case FMT_CF32: e = SDRB_X(FMT_CF32); break;
        default: return cudaErrorInvalidValue;
    }
#undef SDRB_X
    if (e != cudaSuccess) return e;
    if (launches) *launches += P.m > 0 ? (sp > 0 ? 4 : 5) : 3;
    if (P.m > 0) {
        e = launch_ssb_audio(d_a48, P, d_state, d_pcm, n_streams * channels, n_blocks, st);
        if (launches) (*launches)++;
    }
    return e;
}

}  // namespace sdrb
