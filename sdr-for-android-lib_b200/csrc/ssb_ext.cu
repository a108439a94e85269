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
// Parallel structure is the one of ssb.cu with (stream, channel) as the unit that carries state:
//   k_ssbx_state : one lane per (stream, channel), blocks in order: removeDC + NCO + IIR2(I[,Q]),
//                  records the exact filter state at every block start
//   k_ssbx_main  : one lane per (stream, channel, block): full chain to ~48 kHz float audio
//   k_ssb_audio  : (ssb.cu) one lane per (stream, channel): HP/BP, boost, PCM
// PCM layout: [stream][channel][block][m].
#include "internal.h"
#include "nco.cuh"
#include "ssb_common.cuh"

namespace sdrb {

// Warp-level streamer whose 32 rows are arbitrary byte ranges (one pointer per lane's row).
template <int NST>
struct TileStreamX {
    const unsigned char* const* rowptr;   // this warp's 32 row pointers (shared memory)
    unsigned char* tile;                  // this warp's two tiles
    int lane;
    __device__ __forceinline__ void issue(int chunk, int stage) {
        const int sub = lane & 3, r0 = lane >> 2;
        unsigned char* t = tile + stage * SSB_TILE_BYTES;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int r = r0 + 8 * k;
            cp_async16(t + r * SSB_ROW_PITCH + sub * 16, rowptr[r] + (size_t)chunk * SSB_ROW_BYTES + sub * 16);
        }
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

struct SsbxLane {
    float dcr, dci;            // removeDC (complex), reset per block
    float iz1, iz2, qz1, qz2;  // IIR2 previous outputs, I and Q
    uint32_t ph1, ph2;         // NCO phases
};

// removeDC + NCO 1 + IIR2 on I (and Q): one sample.  Returns yI, yQ.
template <bool TSB>
__device__ __forceinline__ void ssbx_front(SsbxLane& L, float I, float Q, const SsbParams& P, float oma,
                                           uint32_t inc1, float& yi, float& yq) {
    L.dcr = __fadd_rn(__fmul_rn(P.dcAlpha, L.dcr), __fmul_rn(oma, I));
    L.dci = __fadd_rn(__fmul_rn(P.dcAlpha, L.dci), __fmul_rn(oma, Q));
    const float xr = __fsub_rn(I, L.dcr), xi = __fsub_rn(Q, L.dci);
    float c1, s1;
    nco_cs(L.ph1, c1, s1);
    L.ph1 += inc1;
    const float mr = __fadd_rn(__fmul_rn(xr, c1), __fmul_rn(xi, s1));
    float y = __fmul_rn(P.rf.a0, mr);
    y = __fadd_rn(y, __fmul_rn(P.rf.a1, L.iz1));
    y = __fadd_rn(y, __fmul_rn(P.rf.a2, L.iz2));
    y = __fsub_rn(y, __fmul_rn(P.rf.b1, L.iz1));
    y = __fsub_rn(y, __fmul_rn(P.rf.b2, L.iz2));
    L.iz2 = L.iz1; L.iz1 = y;
    yi = y;
    yq = 0.0f;
    if (TSB) {
        const float mi = __fsub_rn(__fmul_rn(xi, c1), __fmul_rn(xr, s1));
        float w = __fmul_rn(P.rf.a0, mi);
        w = __fadd_rn(w, __fmul_rn(P.rf.a1, L.qz1));
        w = __fadd_rn(w, __fmul_rn(P.rf.a2, L.qz2));
        w = __fsub_rn(w, __fmul_rn(P.rf.b1, L.qz1));
        w = __fsub_rn(w, __fmul_rn(P.rf.b2, L.qz2));
        L.qz2 = L.qz1; L.qz1 = w;
        yq = w;
    }
}

constexpr int SSBX_WARPS = 4;

// ---------------------------------------------------------------- pass 1
template <int FMT, bool TSB>
__global__ void __launch_bounds__(32)
k_ssbx_state(const void* __restrict__ iq, const __grid_constant__ SsbParams P, const uint32_t* __restrict__ inc1_tab,
             float4* __restrict__ start, SsbStreamState* __restrict__ state, int n_streams, int channels,
             int n_blocks, unsigned long long base_samples) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ const unsigned char* rows[32];
    const int lane = threadIdx.x;
    const size_t nv = (size_t)n_streams * channels;
    const size_t v0 = (size_t)blockIdx.x * 32;
    const size_t v = v0 + lane;
    const bool live = v < nv;
    const size_t vc = live ? v : nv - 1;
    const size_t block_bytes = (size_t)P.n * sizeof(typename Ingest<FMT>::raw_t);
    rows[lane] = static_cast<const unsigned char*>(iq) + (vc / channels) * block_bytes * n_blocks;
    __syncwarp();
    TileStreamX<SSB_STAGES> ts = {rows, smem, lane};
    const int chunks_per_block = (int)(block_bytes / SSB_ROW_BYTES);
    const int nchunks = chunks_per_block * n_blocks;
    const float oma = __fsub_rn(1.0f, P.dcAlpha);
    const uint32_t inc1 = inc1_tab[vc % channels];
    SsbxLane L = {};
    L.iz1 = state[vc].rf_z1; L.iz2 = state[vc].rf_z2; L.qz1 = state[vc].pad0; L.qz2 = state[vc].pad1;
    L.ph1 = (uint32_t)((unsigned long long)inc1 * base_samples);
    ts.prologue(nchunks);
    int cb = 0, b = 0;
    for (int c = 0; c < nchunks; c++) {
        ts.advance(c, nchunks);
        __syncwarp();
        if (cb == 0) {
            L.dcr = L.dci = 0.0f;
            if (live) start[v * n_blocks + b] = make_float4(L.iz1, L.iz2, L.qz1, L.qz2);
        }
        const uint4* row = ts.cur(c);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint4 d = row[q];
#pragma unroll
            for (int j = 0; j < SsbIngest<FMT>::kPerVec; j++) {
                float I, Q, yi, yq;
                SsbIngest<FMT>::get(d, j, I, Q);
                ssbx_front<TSB>(L, I, Q, P, oma, inc1, yi, yq);
            }
        }
        __syncwarp();
        if (++cb == chunks_per_block) { cb = 0; b++; }
    }
    if (live) { state[v].rf_z1 = L.iz1; state[v].rf_z2 = L.iz2; state[v].pad0 = L.qz1; state[v].pad1 = L.qz2; }
}

// ---------------------------------------------------------------- pass 2
template <int FMT, int SMAX, bool TSB>
__global__ void __launch_bounds__(SSBX_WARPS * 32)
k_ssbx_main(const void* __restrict__ iq, const __grid_constant__ SsbParams P, const float* __restrict__ taps,
            const uint32_t* __restrict__ inc1_tab, uint32_t inc2, const float4* __restrict__ start,
            float* __restrict__ a48, int n_streams, int channels, int n_blocks, unsigned long long base_samples) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ const unsigned char* rows_all[SSBX_WARPS][32];
    float* hT = reinterpret_cast<float*>(smem + SSBX_WARPS * SSB_MAIN_STAGES * SSB_TILE_BYTES);
    for (int i = threadIdx.x; i < P.decim * SMAX; i += SSBX_WARPS * 32) hT[i] = taps[i];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t nvf = (size_t)n_streams * channels * n_blocks;
    const size_t vf0 = ((size_t)blockIdx.x * SSBX_WARPS + warp) * 32;
    const size_t vf = vf0 + lane;
    const bool live = vf < nvf;
    const size_t vfc = live ? vf : nvf - 1;
    const size_t v = vfc / n_blocks, b = vfc % n_blocks;
    const size_t block_bytes = (size_t)P.n * sizeof(typename Ingest<FMT>::raw_t);
    rows_all[warp][lane] = static_cast<const unsigned char*>(iq) + ((v / channels) * n_blocks + b) * block_bytes;
    __syncthreads();
    if (vf0 >= nvf) return;
    TileStreamX<SSB_MAIN_STAGES> ts = {rows_all[warp], smem + warp * SSB_MAIN_STAGES * SSB_TILE_BYTES, lane};
    const int nchunks = (int)(block_bytes / SSB_ROW_BYTES);
    const float oma = __fsub_rn(1.0f, P.dcAlpha);
    const uint32_t inc1 = inc1_tab[v % channels];
    const unsigned long long k0 = base_samples + (unsigned long long)b * (unsigned long long)P.n;
    const float4 s0 = start[vfc];
    SsbxLane L = {};
    L.iz1 = s0.x; L.iz2 = s0.y; L.qz1 = s0.z; L.qz2 = s0.w;
    L.ph1 = (uint32_t)((unsigned long long)inc1 * k0);
    L.ph2 = (uint32_t)((unsigned long long)inc2 * k0);
    float gain = 1.0f;
    const float omFast = __fsub_rn(1.0f, P.agcFast), omSlow = __fsub_rn(1.0f, P.agcSlow);
    float acc[SMAX];
#pragma unroll(SMAX <= 16 ? SMAX : 1)
    for (int s = 0; s < SMAX; s++) acc[s] = 0.0f;
    const int sLast = (P.nTaps - 1) / P.decim, rLast = (P.nTaps - 1) % P.decim;
    int r = 0, q = 0;
    float* out = a48 + vfc * (size_t)P.m;

    ts.prologue(nchunks);
    for (int c = 0; c < nchunks; c++) {
        ts.advance(c, nchunks);
        __syncwarp();
        const uint4* row = ts.cur(c);
#pragma unroll
        for (int vv = 0; vv < 4; vv++) {
            const uint4 d = row[vv];
#pragma unroll
            for (int j = 0; j < SsbIngest<FMT>::kPerVec; j++) {
                float I, Q, yi, yq;
                SsbIngest<FMT>::get(d, j, I, Q);
                ssbx_front<TSB>(L, I, Q, P, oma, inc1, yi, yq);
                float a;
                if (TSB) {   // Weaver second mix: Re{(yI + j yQ) exp(+/- j theta2)}, doubled
                    float c2, s2;
                    nco_cs(L.ph2, c2, s2);
                    L.ph2 += inc2;
                    const float p = __fmul_rn(yi, c2), w = __fmul_rn(yq, s2);
                    const float u = P.upper ? __fsub_rn(p, w) : __fadd_rn(p, w);
                    a = __fadd_rn(u, u);
                } else {
                    a = P.upper ? __fadd_rn(yi, yi) : __fsub_rn(yi, yi);
                }
                const float mag = __fadd_rn(fabsf(a), 1e-8f);
                const float desired = __fdiv_rn(P.agcTarget, __fadd_rn(__fsqrt_rn(mag), 1e-6f));
                const bool fast = desired < gain;
                gain = __fadd_rn(__fmul_rn(gain, fast ? omFast : omSlow),
                                 __fmul_rn(desired, fast ? P.agcFast : P.agcSlow));
                float x = __fmul_rn(a, gain);
                x = x < -1.0f ? -1.0f : (1.0f < x ? 1.0f : x);
                const float* h = hT + r * SMAX;
#pragma unroll(SMAX <= 16 ? SMAX : 1)
                for (int s = 0; s < SMAX; s++) acc[s] = __fadd_rn(acc[s], __fmul_rn(x, h[s]));
                if (r == rLast) {
                    const int jo = q - sLast;
                    if (jo >= 0 && jo < P.m && live) out[jo] = acc[SMAX - 1];
                }
                if (++r == P.decim) {
                    r = 0;
                    q++;
#pragma unroll(SMAX <= 16 ? SMAX : 1)
                    for (int s = SMAX - 1; s > 0; s--) acc[s] = acc[s - 1];
                    acc[0] = 0.0f;
                }
            }
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------- pass 2, split (see k_ssb_agc / k_ssb_fir in ssb.cu)
// The main pass in sweeps over a 64-byte chunk, without the decimator: post-AGC samples go to xagc [vf][n].
template <int FMT, bool TSB>
__global__ void __launch_bounds__(SSBX_WARPS * 32)
k_ssbx_agc(const void* __restrict__ iq, const __grid_constant__ SsbParams P, const uint32_t* __restrict__ inc1_tab,
           uint32_t inc2, const float4* __restrict__ start, float* __restrict__ xagc, int n_streams, int channels,
           int n_blocks, unsigned long long base_samples) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int NS = 4 * SsbIngest<FMT>::kPerVec;
    constexpr int OP = NS + 4;
    __shared__ const unsigned char* rows_all[SSBX_WARPS][32];
    __shared__ __align__(16) float s_out[SSBX_WARPS][32 * OP];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t nvf = (size_t)n_streams * channels * n_blocks;
    const size_t vf0 = ((size_t)blockIdx.x * SSBX_WARPS + warp) * 32;
    const size_t vf = vf0 + lane;
    const bool live = vf < nvf;
    const size_t vfc = live ? vf : nvf - 1;
    const size_t v = vfc / n_blocks, b = vfc % n_blocks;
    const size_t block_bytes = (size_t)P.n * sizeof(typename Ingest<FMT>::raw_t);
    rows_all[warp][lane] = static_cast<const unsigned char*>(iq) + ((v / channels) * n_blocks + b) * block_bytes;
    __syncthreads();
    if (vf0 >= nvf) return;
    TileStreamX<SSB_MAIN_STAGES> ts = {rows_all[warp], smem + warp * SSB_MAIN_STAGES * SSB_TILE_BYTES, lane};
    const int nchunks = (int)(block_bytes / SSB_ROW_BYTES);
    const float oma = __fsub_rn(1.0f, P.dcAlpha);
    const uint32_t inc1 = inc1_tab[v % channels];
    const unsigned long long k0 = base_samples + (unsigned long long)b * (unsigned long long)P.n;
    const float4 s0 = start[vfc];
    SsbxLane L = {};
    L.iz1 = s0.x; L.iz2 = s0.y; L.qz1 = s0.z; L.qz2 = s0.w;
    L.ph1 = (uint32_t)((unsigned long long)inc1 * k0);
    L.ph2 = (uint32_t)((unsigned long long)inc2 * k0);
    float gain = 1.0f;
    const float omFast = __fsub_rn(1.0f, P.agcFast), omSlow = __fsub_rn(1.0f, P.agcSlow);
    const bool tgt_ok = ssb_moderate(P.agcTarget);
    float* tile = s_out[warp];
    constexpr int LPR = NS / 4, RPI = 32 / LPR;

    ts.prologue(nchunks);
    for (int c = 0; c < nchunks; c++) {
        ts.advance(c, nchunks);
        __syncwarp();
        const uint4* row = ts.cur(c);
        float a[NS], des[NS];
#pragma unroll
        for (int vv = 0; vv < 4; vv++) {
            const uint4 d = row[vv];
#pragma unroll
            for (int j = 0; j < SsbIngest<FMT>::kPerVec; j++) {
                float I, Q, yi, yq;
                SsbIngest<FMT>::get(d, j, I, Q);
                ssbx_front<TSB>(L, I, Q, P, oma, inc1, yi, yq);
                float av;
                if (TSB) {   // Weaver second mix: Re{(yI + j yQ) exp(+/- j theta2)}, doubled
                    float c2, s2;
                    nco_cs(L.ph2, c2, s2);
                    L.ph2 += inc2;
                    const float p = __fmul_rn(yi, c2), w = __fmul_rn(yq, s2);
                    const float u = P.upper ? __fsub_rn(p, w) : __fadd_rn(p, w);
                    av = __fadd_rn(u, u);
                } else {
                    av = P.upper ? __fadd_rn(yi, yi) : __fsub_rn(yi, yi);
                }
                a[vv * SsbIngest<FMT>::kPerVec + j] = av;
            }
        }
        bool ok = tgt_ok;
#pragma unroll
        for (int i = 0; i < NS; i++) {
            const float mag = __fadd_rn(fabsf(a[i]), 1e-8f);
            ok = ok && ssb_moderate(mag);
            des[i] = ssb_div_rn(P.agcTarget, __fadd_rn(ssb_sqrt_rn(mag), 1e-6f));
        }
        if (!ok) {
#pragma unroll 1
            for (int i = 0; i < NS; i++) {
                float ai = 0.0f;
#pragma unroll
                for (int k = 0; k < NS; k++) if (k == i) ai = a[k];
                const float mag = __fadd_rn(fabsf(ai), 1e-8f);
                const float dsr = __fdiv_rn(P.agcTarget, __fadd_rn(__fsqrt_rn(mag), 1e-6f));
#pragma unroll
                for (int k = 0; k < NS; k++) if (k == i) des[k] = dsr;
            }
        }
#pragma unroll
        for (int i = 0; i < NS; i++) {
            const bool fast = des[i] < gain;
            gain = __fadd_rn(__fmul_rn(gain, fast ? omFast : omSlow), __fmul_rn(des[i], fast ? P.agcFast : P.agcSlow));
            const float x = __fmul_rn(a[i], gain);
            a[i] = x < -1.0f ? -1.0f : (1.0f < x ? 1.0f : x);
        }
        float4* mine = reinterpret_cast<float4*>(tile + lane * OP);
#pragma unroll
        for (int k = 0; k < LPR; k++) mine[k] = make_float4(a[4 * k], a[4 * k + 1], a[4 * k + 2], a[4 * k + 3]);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < LPR; i++) {
            const int r = i * RPI + lane / LPR, col = lane % LPR;
            const size_t fr = vf0 + r;
            if (fr < nvf)
                *reinterpret_cast<float4*>(xagc + fr * (size_t)P.n + (size_t)c * NS + 4 * col) =
                    *reinterpret_cast<const float4*>(tile + r * OP + 4 * col);
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------- launch
template <int FMT, int SMAX, bool TSB>
static cudaError_t launch_mainx_t(const void* iq, const SsbParams& P, const float* taps, const uint32_t* inc1,
                                  uint32_t inc2, const float4* start, float* a48, int S, int C, int B,
                                  unsigned long long base, cudaStream_t st) {
    const size_t smem = (size_t)SSBX_WARPS * SSB_MAIN_STAGES * SSB_TILE_BYTES + (size_t)P.decim * SMAX * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(k_ssbx_main<FMT, SMAX, TSB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const size_t nvf = (size_t)S * C * B;
    const int grid = (int)((nvf + SSBX_WARPS * 32 - 1) / (SSBX_WARPS * 32));
    k_ssbx_main<FMT, SMAX, TSB><<<grid, SSBX_WARPS * 32, smem, st>>>(iq, P, taps, inc1, inc2, start, a48, S, C, B, base);
    return cudaGetLastError();
}

template <int FMT, bool TSB>
static cudaError_t launch_mainx_f(int smax, const void* iq, const SsbParams& P, const float* taps, const uint32_t* inc1,
                                  uint32_t inc2, const float4* start, float* a48, int S, int C, int B,
                                  unsigned long long base, cudaStream_t st) {
    switch (smax) {
        case 1: return launch_mainx_t<FMT, 1, TSB>(iq, P, taps, inc1, inc2, start, a48, S, C, B, base, st);
        case 2: return launch_mainx_t<FMT, 2, TSB>(iq, P, taps, inc1, inc2, start, a48, S, C, B, base, st);
        case 4: return launch_mainx_t<FMT, 4, TSB>(iq, P, taps, inc1, inc2, start, a48, S, C, B, base, st);
        case 8: return launch_mainx_t<FMT, 8, TSB>(iq, P, taps, inc1, inc2, start, a48, S, C, B, base, st);
        case 16: return launch_mainx_t<FMT, 16, TSB>(iq, P, taps, inc1, inc2, start, a48, S, C, B, base, st);
        default: return launch_mainx_t<FMT, 256, TSB>(iq, P, taps, inc1, inc2, start, a48, S, C, B, base, st);
    }
}

template <int FMT, bool TSB>
static cudaError_t launch_ssbx_f(const void* iq, const SsbParams& P, const float* d_taps, const uint32_t* d_inc1,
                                 uint32_t inc2, float4* d_start, SsbStreamState* d_state, float* d_a48, float* d_agc,
                                 int S, int C, int B, unsigned long long base, cudaStream_t st) {
    const size_t nv = (size_t)S * C;
    k_ssbx_state<FMT, TSB><<<(unsigned)((nv + 31) / 32), 32, SSB_STAGES * SSB_TILE_BYTES, st>>>(iq, P, d_inc1, d_start, d_state,
                                                                                      S, C, B, base);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const size_t nvf = nv * B;
    if (d_agc && nvf < (1ull << 31)) {
        const int grid = (int)((nvf + SSBX_WARPS * 32 - 1) / (SSBX_WARPS * 32));
        const size_t asmem = (size_t)SSBX_WARPS * SSB_MAIN_STAGES * SSB_TILE_BYTES;
        if ((e = cudaFuncSetAttribute(k_ssbx_agc<FMT, TSB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)asmem)) != cudaSuccess)
            return e;
        k_ssbx_agc<FMT, TSB><<<grid, SSBX_WARPS * 32, asmem, st>>>(
            iq, P, d_inc1, inc2, d_start, d_agc, S, C, B, base);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        return launch_ssb_fir(d_agc, d_taps, d_a48, P, (int)nvf, st);
    }
    return launch_mainx_f<FMT, TSB>(ssb_smax(P), iq, P, d_taps, d_inc1, inc2, d_start, d_a48, S, C, B, base, st);
}

cudaError_t launch_ssb_ext(int fmt, const void* iq, const SsbParams& P, const float* d_taps, const uint32_t* d_inc1,
                           uint32_t inc2, int true_sideband, float4* d_start, SsbStreamState* d_state, float* d_a48,
                           int16_t* d_pcm, float* d_agc, int n_streams, int channels, int n_blocks,
                           unsigned long long base_samples, cudaStream_t st, int* launches) {
    if (n_streams <= 0 || channels <= 0 || n_blocks <= 0) return cudaSuccess;
    cudaError_t e;
#define SDRB_X(F)                                                                                                      \
    (true_sideband ? launch_ssbx_f<F, true>(iq, P, d_taps, d_inc1, inc2, d_start, d_state, d_a48, d_agc, n_streams, channels, \
                                            n_blocks, base_samples, st)                                                \
                   : launch_ssbx_f<F, false>(iq, P, d_taps, d_inc1, inc2, d_start, d_state, d_a48, d_agc, n_streams, channels, \
                                             n_blocks, base_samples, st))
    switch (fmt) {
        case FMT_CU8: e = SDRB_X(FMT_CU8); break;
        case FMT_CS16: e = SDRB_X(FMT_CS16); break;
        case FMT_CF32: e = SDRB_X(FMT_CF32); break;
        default: return cudaErrorInvalidValue;
    }
#undef SDRB_X
    if (e != cudaSuccess) return e;
    if (launches) *launches += (d_agc && P.m > 0) ? 3 : 2;
    if (P.m > 0) {
        e = launch_ssb_audio(d_a48, P, d_state, d_pcm, n_streams * channels, n_blocks, st);
        if (launches) (*launches)++;
    }
    return e;
}

}  // namespace sdrb
