// estimator.cu — evaluateSignalStrength on the GPU (src/dsp/fft_process.cpp:122-378).
//
// Three kernels:
//   k_band_scan : one THREAD per (frame, band); band 0 = focus band, 1..nWin = reference windows.
//                 The float sums the reference does sequentially (signalPowerSum :148, window sums
//                 :199, the sliding run-sum of best1kHzMean :163-180 / :310-318) are done
//                 sequentially here too, un-fused, so that the arg-max bin, the best-1 kHz start
//                 bin and every band sum round exactly like the reference.  32 bands per warp.
//   k_estimator : one WARP per frame: dB figures of the bands (double log10 rounded once; glibc's
//                 log10f is not reproducible on the GPU, difference <= 1 ulp), stable rank sort of
//                 the windows by mean dB (:228-229), bottom-40 % statistics (:235-247, :292-300),
//                 pooled per-bin noise (:252-269: dB of every bin of the quiet windows, mean,
//                 median absolute deviation by a 4-pass radix select), Gumbel / best-1 kHz SNR.
//   k_tracker   : one thread per stream, the sequential part that carries state from frame to
//                 frame (frequency tracking :333-361, detection flag :365-378, hold-over of
//                 outputs on skipped / invalid frames) and writes the sdrb_frame_result records.
// The pooled per-bin mean (:261-262) is a tree sum here (the reference adds sequentially); it
// only feeds tolerance-checked scalars (0.01 dB), not an index or a sort decision.
#include <algorithm>
#include <cstdlib>

#include <cuda_fp16.h>

#include "internal.h"

namespace sdrb {

constexpr int EST_WARPS = 4;                 // frames per CTA in k_estimator (one warp each)
constexpr int EST_SMEM_POOL_FLOATS = 8192;   // pooled bins staged in shared memory per warp (32 KiB x 4 warps at most); above: global scratch
constexpr int EST_MAX_BANDS = 11;

__device__ __forceinline__ float db10(float p) {
    // 10.0f * log10f(p / refPower + 1e-20f), refPower = 1  (src/dsp/fft_process.cpp:149)
    return __fmul_rn(10.0f, (float)log10((double)__fadd_rn(p, 1e-20f)));
}
__device__ __forceinline__ float db10_fast(float p) {   // per-bin noise statistics only
    return __fmul_rn(10.0f, log10f(__fadd_rn(p, 1e-20f)));
}

struct BandStat {
    float sum;       // sequential sum of the band
    float maxP;      // max power, strict >, first index
    int maxIdx;      // offset inside the band
    float bestRun;   // max sliding run-sum of winBins1k bins (or the plain sum when shorter)
    int bestStart;   // offset of the first window that reaches bestRun
};

// One band in order: sum + arg-max (:146-155 focus, :198-202 windows) and the best run of W bins (:163-180 / :310-318:
// run-sum with +new -old updates, strict >).  Shared by k_band_scan and by the estimator's inline scan of short bands.
__device__ __forceinline__ BandStat scan_band(const float* __restrict__ p, int len, int W, bool focus) {
    float sum = 0.0f, maxP = focus ? -1.0f : 0.0f;
    int maxIdx = 0;
#pragma unroll 32
    for (int i = 0; i < len; i++) {   // 32 loads in flight: the scan is bound by memory latency, not by the adds
        const float v = p[i];
        sum = __fadd_rn(sum, v);
        if (v > maxP) { maxP = v; maxIdx = i; }
    }
    BandStat r;
    r.sum = sum; r.maxP = maxP; r.maxIdx = maxIdx;
    if (len < W) {
        r.bestRun = sum; r.bestStart = 0;
    } else {
        float run = 0.0f;
        for (int i = 0; i < W; i++) run = __fadd_rn(run, p[i]);
        float best = run;
        int bestStart = 0;
#pragma unroll 16
        for (int s = 1; s + W - 1 < len; s++) {
            run = __fadd_rn(run, __fsub_rn(p[s + W - 1], p[s - 1]));
            if (run > best) { best = run; bestStart = s; }
        }
        r.bestRun = best; r.bestStart = bestStart;
    }
    return r;
}

// ---------------------------------------------------------------- k_band_scan
// Mirrors :146-155 (sum + arg-max), :198-202 (sum + max) and best1kHzMean :163-180 / :310-318
// (run-sum with +new -old updates, strict >).
// STAGED: the CTA first brings the bands of its `fpc` frames into shared memory with 4-byte cp.async
// copies (one warp per band slot, lanes along the band: coalesced, nothing waits on a register, all copies of
// the CTA are in flight together), then every thread scans its own slot.  Slots are `pitch` floats apart,
// pitch odd, so the 32 lanes of a warp (32 consecutive slots, same position i) fall on 32 different banks.
// One thread per band walking global memory instead touches 32 different lines per warp load and is bound by
// L1 tag lookups and miss latency.
template <bool STAGED>
__global__ void __launch_bounds__(128)
k_band_scan(const float* __restrict__ power, const EstGeometry* __restrict__ geo_p, BandStat* __restrict__ stats,
            int fpc, int n_frames, int pitch) {
    extern __shared__ float band_smem[];
    __shared__ EstGeometry g;
    {
        const int* src = reinterpret_cast<const int*>(geo_p);
        int* dst = reinterpret_cast<int*>(&g);
        for (int i = threadIdx.x; i < (int)(sizeof(EstGeometry) / 4); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    if (g.focusLen <= 0) return;
    const int nSeg = g.nWin + 1;
    const int frame0 = blockIdx.x * fpc;
    const int nf = min(fpc, n_frames - frame0);
    if (STAGED) {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        for (int slot = warp; slot < nf * nSeg; slot += 4) {
            const int fl = slot / nSeg, seg = slot - fl * nSeg;
            const int lo = seg == 0 ? g.focusLo : g.winLo[seg - 1];
            const int len = g.segOff[seg + 1] - g.segOff[seg];
            const float* src = power + (size_t)(frame0 + fl) * g.n + lo;
            const unsigned dst = (unsigned)__cvta_generic_to_shared(band_smem + (size_t)slot * pitch);
            for (int i = lane; i < len; i += 32)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + 4u * i), "l"(src + i) : "memory");
        }
        asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
        __syncthreads();
    }
    const int id = threadIdx.x;
    if (id >= nf * nSeg) return;
    const int fl = id / nSeg, seg = id % nSeg;
    const int frame = frame0 + fl;
    const int lo = seg == 0 ? g.focusLo : g.winLo[seg - 1];
    const int len = g.segOff[seg + 1] - g.segOff[seg];
    const int W = g.winBins1k;
    const float* __restrict__ p = STAGED ? band_smem + (size_t)id * pitch : power + (size_t)frame * g.n + lo;
    stats[(size_t)frame * EST_MAX_BANDS + seg] = scan_band(p, len, W, seg == 0);
}

// ---------------------------------------------------------------- k_estimator
// Median-of-gaps for <= 4 values (:238-242, :296-300): sorted |x - mean| [nb / 2].
__device__ __forceinline__ void cswap(float& a, float& b) { const float lo = fminf(a, b), hi = fmaxf(a, b); a = lo; b = hi; }
__device__ __forceinline__ float mad4(const float* x, int nb, float mean) {
    float g[4];
#pragma unroll
    for (int i = 0; i < 4; i++) g[i] = i < nb ? fabsf(__fsub_rn(x[i], mean)) : 3.0e38f;
    cswap(g[0], g[1]); cswap(g[2], g[3]); cswap(g[0], g[2]); cswap(g[1], g[3]); cswap(g[1], g[2]);
    const int m = nb / 2;
    return m == 0 ? g[0] : (m == 1 ? g[1] : g[2]);
}

// U: unroll factor of the pooled-bin loops (8 when many bins are pooled per frame, 1 otherwise; see the launcher)
// INLINE: the bands are short (<= 32 bins): lanes 0 .. nWin scan their own band here, no k_band_scan launch, no round
// trip of the band statistics (short frames: the band scan was a fifth of the whole FFT stage at 1024 points)
template <bool USE_SMEM, int U, bool INLINE>
__global__ void __launch_bounds__(EST_WARPS * 32)
k_estimator(const float* __restrict__ power, const EstGeometry* __restrict__ geo_p, const BandStat* __restrict__ stats,
            FrameLocal* __restrict__ local, float* __restrict__ gscratch, int pooled_cap, int n_frames) {
    extern __shared__ float stage_smem[];
    __shared__ EstGeometry g;
    __shared__ int hist_all[EST_WARPS][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    {
        const int* src = reinterpret_cast<const int*>(geo_p);
        int* dst = reinterpret_cast<int*>(&g);
        for (int i = threadIdx.x; i < (int)(sizeof(EstGeometry) / 4); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const int frame = blockIdx.x * EST_WARPS + warp;
    if (frame >= n_frames) return;
    if (g.focusLen <= 0) {  // :139 early return, the frame leaves every output untouched
        if (lane == 0) {
            FrameLocal z = {};
            z.peakBin = -1;
            z.best1kStart = -1;
            local[frame] = z;
        }
        return;
    }
    const unsigned FULL = 0xffffffffu;
    const int nWin = g.nWin, W = g.winBins1k;
    const bool valid = nWin >= 2;  // :218
    const bool isWin = lane >= 1 && lane <= nWin;
    BandStat bs = {};
    const int len = lane <= nWin ? g.segOff[lane + 1] - g.segOff[lane] : 1;
    if (lane <= nWin) {
        if (INLINE) bs = scan_band(power + (size_t)frame * g.n + (lane == 0 ? g.focusLo : g.winLo[lane - 1]), len, W, lane == 0);
        else bs = stats[(size_t)frame * EST_MAX_BANDS + lane];
    }
    // dB figures of the bands, one lane per band (:149-155, :203-213)
    float dbA = 0.0f, dbB = 0.0f;   // focus: (absPeakDb candidate, signalPowerDb); window: (meanDb, best1kDb)
    const float focusSum = __shfl_sync(FULL, bs.sum, 0), focusBestRun = __shfl_sync(FULL, bs.bestRun, 0);
    // the focus band's best-1 kHz mean (:310-318), whose dB figure is needed at the very end: lane 31 (never a window)
    // takes it through the same pair of log10 calls instead of a third, divergent one
    const float fb = g.focusLen < W ? __fdiv_rn(focusSum, (float)g.focusLen) : __fdiv_rn(focusBestRun, (float)W);
    float dbFb;
    {   // arguments first, then ONE pair of log10 calls for the whole warp (the double log10 is ~130 instructions;
        // under a lane-0 / window branch it would run four times)
        const bool want = lane == 0 || (isWin && valid);
        const float meanP = __fdiv_rn(bs.sum, (float)(lane == 0 ? g.focusLen : len));
        const float argA = lane == 31 ? fb : (lane == 0 ? bs.maxP : meanP);
        const float argB = (lane == 0 || len < W) ? meanP : __fdiv_rn(bs.bestRun, (float)W);
        const float a = db10((want || lane == 31) ? argA : 1.0f), b = db10(want ? argB : 1.0f);
        if (want) { dbA = a; dbB = b; }
        dbFb = __shfl_sync(FULL, a, 31);
    }
    FrameLocal o = {};
    {   // ---- :142-155 focus band
        const float dbMax = __shfl_sync(FULL, dbA, 0);
        const int maxIdx = __shfl_sync(FULL, bs.maxIdx, 0);
        if (dbMax > -130.0f) { o.absPeakDb = dbMax; o.peakBin = g.focusLo + maxIdx; }
        else { o.absPeakDb = -130.0f; o.peakBin = g.focusLo; }
        o.signalPowerDb = __shfl_sync(FULL, dbB, 0);
        o.valid = valid ? 1 : 0;
        o.best1kStart = -1;
    }
    const int focusBestStart = __shfl_sync(FULL, bs.bestStart, 0);
    if (!valid) {
        if (lane == 0) local[frame] = o;
        return;
    }
    // ---- :228-229 sort by meanDb, as a stable rank (one window per lane)
    int rank = 0;
    for (int v = 1; v <= nWin; v++) {
        const float other = __shfl_sync(FULL, dbA, v);
        if (isWin && (other < dbA || (other == dbA && v < lane))) rank++;
    }
    const int nb = g.nBottom;   // <= 4
    float x[4], x1k[4];
    int plo[4], cum[5];
    cum[0] = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        x[i] = x1k[i] = 0.0f;
        plo[i] = 0;
        cum[i + 1] = cum[i];
        if (i < nb) {
            const int src = __ffs(__ballot_sync(FULL, isWin && rank == i)) - 1;   // window src - 1, sorted position i
            x[i] = __shfl_sync(FULL, dbA, src);
            x1k[i] = __shfl_sync(FULL, dbB, src);
            plo[i] = g.winLo[src - 1];
            cum[i + 1] = cum[i] + (g.segOff[src + 1] - g.segOff[src]);
        }
    }
    // ---- :235-247 mean-energy SNR (sequential sums over <= 4 values, uniform across lanes)
    float mean = 0.0f, mean1k = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; i++)
        if (i < nb) { mean = __fadd_rn(mean, x[i]); mean1k = __fadd_rn(mean1k, x1k[i]); }
    mean = __fdiv_rn(mean, (float)nb);
    mean1k = __fdiv_rn(mean1k, (float)nb);
    const float sigma = fmaxf(__fmul_rn(1.4816f, mad4(x, nb, mean)), 0.5f);
    o.meanSnrDb = __fsub_rn(o.signalPowerDb, mean);
    o.meanSnrSigma = __fdiv_rn(o.meanSnrDb, sigma);
    const float mad1k = __fmul_rn(1.4816f, mad4(x1k, nb, mean1k));   // :296-300, floor applied below
    // ---- :252-262 pooled per-bin dB of the quiet windows and their mean
    const int pooledN = cum[nb < 4 ? nb : 4];
    float* stage = USE_SMEM ? stage_smem + (size_t)warp * pooled_cap : gscratch + (size_t)frame * pooled_cap;
    const float* P = power + (size_t)frame * g.n;
    float part = 0.0f;
#pragma unroll
    for (int j = 0; j < 4; j++) {   // quiet window by quiet window: lanes along the window, coalesced
        if (j < nb) {
            const float* Pw = P + plo[j];
            float* sw = stage + cum[j];
            const int lenj = cum[j + 1] - cum[j];
            // (U = 8: with one warp per frame and long pooled stretches every iteration's load + log10f latency was
            // exposed; eight independent iterations in flight hide it.  The per-lane order of the sum is unchanged.)
#pragma unroll U
            for (int i = lane; i < lenj; i += 32) {
                const float d = db10_fast(__ldg(Pw + i));
                sw[i] = d;
                part += d;
            }
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(FULL, part, d);
    const float perBinMean = __fdiv_rn(part, (float)pooledN);
    // the staging loop above indexes lanes relative to each window's start, the loop below absolutely: a lane reads
    // what another lane wrote, so the writes have to be ordered before the reads (shared or global scratch alike)
    __syncwarp();
#pragma unroll U
    for (int i = lane; i < pooledN; i += 32) stage[i] = fabsf(__fsub_rn(stage[i], perBinMean));   // :265-266
    __syncwarp();
    // ---- :267-268 median of the gaps: 4-pass radix select on the bit pattern; each lane owns 8 of
    // the 256 histogram bins and the crossing bin is found with a warp scan
    int* hist = hist_all[warp];
    unsigned prefix = 0, mask = 0;
    int k = pooledN / 2;
    for (int shift = 24; shift >= 0; shift -= 8) {
#pragma unroll
        for (int j = 0; j < 8; j++) hist[lane + 32 * j] = 0;
        __syncwarp();
#pragma unroll U
        for (int i = lane; i < pooledN; i += 32) {
            const unsigned bits = __float_as_uint(stage[i]);
            if ((bits & mask) == prefix) atomicAdd(&hist[(bits >> shift) & 255u], 1);
        }
        __syncwarp();
        int h8[8], mine = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) { h8[j] = hist[8 * lane + j]; mine += h8[j]; }
        int incl = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int v = __shfl_up_sync(FULL, incl, d);
            if (lane >= d) incl += v;
        }
        const unsigned cross = __ballot_sync(FULL, incl > k);
        const int L = cross ? __ffs(cross) - 1 : 31;
        // inside the crossing lane's 8 bins: the first bin whose running count passes k (the last bin if none does)
        int digit = 8 * lane, acc = incl - mine;
        bool open = true;
#pragma unroll
        for (int j = 0; j < 7; j++) {
            open = open && (acc + h8[j] <= k);
            if (open) { acc += h8[j]; digit++; }
        }
        digit = __shfl_sync(FULL, digit, L);
        acc = __shfl_sync(FULL, acc, L);
        prefix |= (unsigned)digit << shift;
        k -= acc;
        mask |= 255u << shift;
        __syncwarp();
    }
    if (lane == 0) {
        const float sigmaBin = fmaxf(__fmul_rn(1.4816f, __uint_as_float(prefix)), 1.0f);
        o.perBinMean = perBinMean;
        o.peakAboveNoiseMeanDb = __fsub_rn(o.absPeakDb, perBinMean);                     // :274
        // :281-288 Gumbel-corrected max-bin SNR
        const float gumbelLoc = __fadd_rn(perBinMean, __fmul_rn(sigmaBin, g.sqrt2logN));
        const float gumbelSig = fmaxf(__fdiv_rn(__fmul_rn(sigmaBin, 3.14159f), g.gumbelDen), 0.5f);
        o.maxBinSnrDb = __fsub_rn(o.absPeakDb, gumbelLoc);
        o.maxBinSnrSigma = __fdiv_rn(o.maxBinSnrDb, gumbelSig);
        // :299-327 best-1 kHz SNR
        const float sigmaFloor1k = __fdiv_rn(sigmaBin, g.sqrtWin1k);
        const float sigma1k = fmaxf(fmaxf(mad1k, sigmaFloor1k), 0.5f);
        if (fb > 0.0f) {
            o.best1kHzSnrDb = __fsub_rn(dbFb, mean1k);
            o.best1kHzSnrSigma = __fdiv_rn(o.best1kHzSnrDb, sigma1k);
            o.best1kStart = g.focusLo + focusBestStart;
        } else {
            o.best1kHzSnrDb = 0.0f;
            o.best1kHzSnrSigma = 0.0f;
            o.best1kStart = -1;
        }
        local[frame] = o;
    }
}

// Frame clock (SURVEY.md 8a note T), integer arithmetic: with q = (frame + 1) * n,
//   frame_clock_ns = (q / fs) * 1e9 + ((q % fs) * 1e9) / fs

// One thread per stream walks its frames in order (:333-378) and emits the result records.
__global__ void __launch_bounds__(128)
k_tracker(const FrameLocal* __restrict__ local, const EstGeometry* __restrict__ geo_p,
          FftStreamState* __restrict__ state, sdrb_frame_result* __restrict__ results, int n_streams,
          int n_blocks) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_streams) return;
    const EstGeometry& g = *geo_p;
    FftStreamState st = state[s];
    const unsigned fs = (unsigned)g.sampleRateF;
    // The frame clock t = floor-ns of (frameCounter + 1) * n / fs is carried as whole seconds + a remainder
    // below fs, advanced by n per frame: the three 64-bit divisions per frame (hundreds of
    // dependent instructions each) of that formula become one division before the loop and an FP64 quotient with an exact
    // integer correction per frame.  Same integers.
    const unsigned long long q0 = (unsigned long long)st.frameCounter * (unsigned long long)g.n;
    unsigned long long sec = q0 / fs, rem = q0 % fs;
    const unsigned long long n_div = (unsigned)g.n / fs, n_mod = (unsigned)g.n % fs;
    const double inv_fs = 1.0 / (double)fs;
    const FrameLocal* lp = local + (size_t)s * n_blocks;
    FrameLocal f_next = lp[0];
    for (int b = 0; b < n_blocks; b++) {
        const FrameLocal f = f_next;
        if (b + 1 < n_blocks) f_next = lp[b + 1];   // in flight while this frame's chain runs
        sdrb_frame_result r = {};
        sec += n_div;
        rem += n_mod;
        if (rem >= fs) { rem -= fs; sec++; }
        const unsigned long long num = rem * 1000000000ULL;             // < 2^62
        unsigned long long frac = (unsigned long long)((double)num * inv_fs);   // within 1 of the floor
        const long long left = (long long)(num - frac * fs);
        if (left < 0) frac--;
        else if (left >= (long long)fs) frac++;
        const long long now_ns = (long long)(sec * 1000000000ULL + frac);
        st.frameCounter++;
        r.peakBin = f.peakBin;
        r.nRef = g.nWin;
        if (f.peakBin >= 0) {
            const float fcF = (float)st.centerFrequency;
            // uint32 - float promotes the integer to float first (:347-348)
            const float base = __fsub_rn(fcF, g.nyquist);
            const float peakHz = __fadd_rn(__fmul_rn((float)f.peakBin, g.freqPerBin), base);
            r.absPeakDb = f.absPeakDb;
            r.signalPowerDb = f.signalPowerDb;
            r.peakFrequencyHz = peakHz;
            r.valid = f.valid;
            if (!f.valid) {  // :220-225
                st.meanSnrDb = st.meanSnrSigma = 0.0f;
                st.peakAboveNoiseMeanDb = st.maxBinSnrDb = st.maxBinSnrSigma = 0.0f;
                st.best1kHzSnrDb = st.best1kHzSnrSigma = 0.0f;
            } else {
                st.meanSnrDb = f.meanSnrDb;
                st.meanSnrSigma = f.meanSnrSigma;
                st.perBinMean = f.perBinMean;
                st.peakAboveNoiseMeanDb = f.peakAboveNoiseMeanDb;
                st.maxBinSnrDb = f.maxBinSnrDb;
                st.maxBinSnrSigma = f.maxBinSnrSigma;
                st.best1kHzSnrDb = f.best1kHzSnrDb;
                st.best1kHzSnrSigma = f.best1kHzSnrSigma;
                if (f.best1kStart >= 0)  // :321-322
                    st.best1kHzCenterFreqHz =
                        __fadd_rn(__fmul_rn((float)(f.best1kStart + g.winBins1k / 2), g.freqPerBin), base);
            }
            // ---- :333-361 frequency tracking
            if (st.trackingFrequency == 0.0f) st.trackingFrequency = fcF;
            if (st.centerChanged) { st.trackingFrequency = fcF; st.centerChanged = 0; }
            if (f.valid && f.absPeakDb > st.maxPeakDb) {
                st.maxPeakDb = f.absPeakDb;
                st.maxPeakHz = peakHz;
                st.timeOfLastMaxPeak_ns = now_ns;
            }
            const long long msSincePeak = (now_ns - st.timeOfLastMaxPeak_ns) / 1000000LL;
            if (st.timeOfLastMaxPeakUpdate_ns < st.timeOfLastMaxPeak_ns && msSincePeak > 300) {
                st.trackingFrequency = st.maxPeakHz;
                st.timeOfLastMaxPeakUpdate_ns = now_ns;
                st.maxPeakDb = -130.0f;
            }
            // ---- :365-378 detection flag
            const bool above = f.valid && (st.meanSnrSigma >= 4.0f);
            if (above) { if (st.peakConfirmed < 1) st.peakConfirmed++; }
            else st.peakConfirmed = 0;
            const int cur = (above && st.peakConfirmed >= 1) ? 3 : 0;
            if (st.flagIndex == 0) st.flagRing[0] = cur;
            else if (st.flagIndex == 1) st.flagRing[1] = cur;
            else st.flagRing[2] = cur;
            st.flagIndex = (st.flagIndex + 1) % 3;
            st.detectionFlag = max(st.flagRing[0], max(st.flagRing[1], st.flagRing[2]));
        }
        r.meanSnrDb = st.meanSnrDb;
        r.meanSnrSigma = st.meanSnrSigma;
        r.peakAboveNoiseMeanDb = st.peakAboveNoiseMeanDb;
        r.maxBinSnrDb = st.maxBinSnrDb;
        r.maxBinSnrSigma = st.maxBinSnrSigma;
        r.best1kHzSnrDb = st.best1kHzSnrDb;
        r.best1kHzSnrSigma = st.best1kHzSnrSigma;
        r.best1kHzCenterFreqHz = st.best1kHzCenterFreqHz;
        r.perBinMean = st.perBinMean;
        r.trackingFrequencyHz = st.trackingFrequency;
        r.trackingFrequency = (int64_t)roundf(st.trackingFrequency);  // getTrackingFrequency(), fft_process.h:41
        r.detectionFlag = st.detectionFlag;
        results[(size_t)s * n_blocks + b] = r;
    }
    state[s] = st;
}

static int pooled_cap_of(const EstGeometry& g) {
    // the nBottom (<= 4) longest reference windows bound the pooled bin count
    int lens[10];
    for (int w = 0; w < g.nWin; w++) lens[w] = g.segOff[w + 2] - g.segOff[w + 1];
    for (int i = 0; i < g.nWin; i++)
        for (int j = i + 1; j < g.nWin; j++)
            if (lens[j] > lens[i]) { const int t = lens[i]; lens[i] = lens[j]; lens[j] = t; }
    int cap = 0;
    for (int i = 0; i < g.nBottom && i < g.nWin && i < 4; i++) cap += lens[i];
    return cap > 0 ? cap : 1;
}

cudaError_t estimator_prepare() {
    cudaError_t e = cudaFuncSetAttribute(k_band_scan<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    if (e != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_estimator<true, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  EST_WARPS * EST_SMEM_POOL_FLOATS * (int)sizeof(float))) != cudaSuccess)
        return e;
    if ((e = cudaFuncSetAttribute(k_estimator<true, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  EST_WARPS * EST_SMEM_POOL_FLOATS * (int)sizeof(float))) != cudaSuccess)
        return e;
    return cudaFuncSetAttribute(k_estimator<true, 8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                EST_WARPS * EST_SMEM_POOL_FLOATS * (int)sizeof(float));
}

// Scratch (floats): band statistics for every frame, then the pooled-bin staging area when it does
// not fit in shared memory.
static size_t band_floats(int n_frames) { return (size_t)n_frames * EST_MAX_BANDS * (sizeof(BandStat) / sizeof(float)); }
size_t estimator_scratch_floats(const EstGeometry& g, int n_frames) {
    const int cap = pooled_cap_of(g);
    return band_floats(n_frames) + (cap > EST_SMEM_POOL_FLOATS ? (size_t)cap * (size_t)n_frames : 0);
}

// Bands of at most 32 bins (short frames, narrow focus) are scanned inside the estimator kernel: one launch, not two.
bool estimator_inline_scan(const EstGeometry& g) {
    static const bool off = getenv("SDRB_EST_INLINE") && getenv("SDRB_EST_INLINE")[0] == '0';   // test hook
    if (g.focusLen <= 0 || off) return false;
    int maxLen = 0;
    for (int i = 0; i <= g.nWin; i++) maxLen = std::max(maxLen, g.segOff[i + 1] - g.segOff[i]);
    return maxLen <= 32 && pooled_cap_of(g) <= EST_SMEM_POOL_FLOATS;
}

// scratch = base of the context's scratch (sized for cap_frames); frame0 = index of this call's
// first frame inside the context (power / d_local are already offset by the caller).
cudaError_t launch_estimator(const float* power, const EstGeometry* d_geo, const EstGeometry& h_geo,
                             FrameLocal* d_local, float* scratch, size_t frame0, int cap_frames, int n_frames,
                             cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    if (!scratch) return cudaErrorInvalidValue;
    BandStat* stats = reinterpret_cast<BandStat*>(scratch) + frame0 * EST_MAX_BANDS;
    const int nSeg = h_geo.nWin + 1;
    const bool inline_scan = estimator_inline_scan(h_geo);
    if (h_geo.focusLen > 0 && !inline_scan) {
        // frames per CTA: one thread per band (<= 128 threads); staged when one frame's slots fit 96 KiB of
        // shared memory (measured on B200, 37888 frames of the bench geometry: 143 us walking global memory)
        int fpc = 128 / nSeg;
        int maxLen = 0;
        for (int sgi = 0; sgi < nSeg; sgi++) maxLen = std::max(maxLen, h_geo.segOff[sgi + 1] - h_geo.segOff[sgi]);
        const int pitch = maxLen | 1;
        const size_t per_frame = (size_t)nSeg * pitch * sizeof(float);
        static const bool no_stage = getenv("SDRB_BAND_STAGE") && getenv("SDRB_BAND_STAGE")[0] == '0';
        const int by_smem = (int)((96 * 1024) / per_frame);
        if (by_smem >= 1 && !no_stage) {
            // about 40 KiB per CTA keeps five CTAs on an SM: the scan is a dependent chain, it needs warps in flight
            const int by_occ = std::max(1, (int)((40 * 1024) / per_frame));
            fpc = std::min(fpc, std::min(by_smem, by_occ));
            const size_t smem = (size_t)fpc * per_frame;
            k_band_scan<true><<<(n_frames + fpc - 1) / fpc, 128, smem, st>>>(power, d_geo, stats, fpc, n_frames, pitch);
        } else {
            k_band_scan<false><<<(n_frames + fpc - 1) / fpc, 128, 0, st>>>(power, d_geo, stats, fpc, n_frames, 0);
        }
    }
    const int cap = pooled_cap_of(h_geo);
    const int grid = (n_frames + EST_WARPS - 1) / EST_WARPS;
    // Unrolled pooled-bin loops pay when a lane walks many bins (1 M points, 25 kHz focus: ~250 per lane, C5 +11 %) and
    // cost instructions when it walks a handful (1024-point frames: -8 %): measured, the switch is around 32 per lane.
    const bool deep = cap >= 1024;
    if (cap <= EST_SMEM_POOL_FLOATS) {
        const size_t sm = (size_t)EST_WARPS * cap * sizeof(float);
        if (inline_scan) k_estimator<true, 1, true><<<grid, EST_WARPS * 32, sm, st>>>(power, d_geo, stats, d_local, nullptr, cap, n_frames);
        else if (deep) k_estimator<true, 8, false><<<grid, EST_WARPS * 32, sm, st>>>(power, d_geo, stats, d_local, nullptr, cap, n_frames);
        else k_estimator<true, 1, false><<<grid, EST_WARPS * 32, sm, st>>>(power, d_geo, stats, d_local, nullptr, cap, n_frames);
    } else {
        float* gs = scratch + band_floats(cap_frames) + frame0 * (size_t)cap;
        k_estimator<false, 8, false><<<grid, EST_WARPS * 32, 0, st>>>(power, d_geo, stats, d_local, gs, cap, n_frames);
    }
    return cudaGetLastError();
}

cudaError_t launch_tracker(const FrameLocal* d_local, const EstGeometry* d_geo, FftStreamState* d_state,
                           sdrb_frame_result* d_results, int n_streams, int n_blocks, cudaStream_t st) {
    if (n_streams <= 0 || n_blocks <= 0) return cudaSuccess;
    k_tracker<<<(n_streams + 127) / 128, 128, 0, st>>>(d_local, d_geo, d_state, d_results, n_streams, n_blocks);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- dB view of the spectrum
// 10*log10f(p / refPower + 1e-20f), refPower = 1 (fft_process.cpp:142,149), for every bin: an extra output next to
// the linear payload.  Streaming: 16-byte loads, 16- or 8-byte stores.
template <bool HALF>
__global__ void __launch_bounds__(256)
k_power_view(const float4* __restrict__ p, void* __restrict__ out, size_t n4) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 v = p[i];
        const float a = 10.0f * log10f(v.x + 1e-20f), b = 10.0f * log10f(v.y + 1e-20f);
        const float c = 10.0f * log10f(v.z + 1e-20f), d = 10.0f * log10f(v.w + 1e-20f);
        if (HALF) {
            const __half2 lo = __floats2half2_rn(a, b), hi = __floats2half2_rn(c, d);
            uint2 w;
            w.x = *reinterpret_cast<const unsigned*>(&lo);
            w.y = *reinterpret_cast<const unsigned*>(&hi);
            static_cast<uint2*>(out)[i] = w;
        } else {
            static_cast<float4*>(out)[i] = make_float4(a, b, c, d);
        }
    }
}
__global__ void k_power_view_tail(const float* __restrict__ p, void* __restrict__ out, size_t from, size_t count, int half) {
    const size_t i = from + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const float a = 10.0f * log10f(p[i] + 1e-20f);
    if (half) static_cast<__half*>(out)[i] = __float2half_rn(a);
    else static_cast<float*>(out)[i] = a;
}

cudaError_t launch_power_view(const float* power, void* out, size_t count, int fmt, cudaStream_t st) {
    if (!count) return cudaSuccess;
    if (fmt != 1 && fmt != 2) return cudaErrorInvalidValue;
    const size_t n4 = count / 4;
    if (n4) {
        const int grid = (int)((n4 + 255) / 256 < 148 * 16 ? (n4 + 255) / 256 : 148 * 16);
        if (fmt == 2) k_power_view<true><<<grid, 256, 0, st>>>(reinterpret_cast<const float4*>(power), out, n4);
        else k_power_view<false><<<grid, 256, 0, st>>>(reinterpret_cast<const float4*>(power), out, n4);
    }
    if (count & 3) k_power_view_tail<<<1, 32, 0, st>>>(power, out, n4 * 4, count, fmt == 2);
    return cudaGetLastError();
}

}  // namespace sdrb

// ---------------------------------------------------------------- allocation-free estimator (sdrb_estimate_noalloc)
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The frame-local part of
// evaluateSignalStrength (k_band_scan + k_estimator above, unchanged) over caller-owned device spectra, with the
// scratch those kernels need -- the bin geometry and the per-frame band statistics that pass from one kernel to the
// other -- in a file-scope device array instead of a cudaMalloc'ed buffer.  Every GPU that loads the library holds the
// array (about 0.9 MB) until exit.  Every call shares it: nothing in it has to start at zero (the geometry and every
// band statistic a call reads are written by that call first), but two calls that overlap on the GPU race on it.
#include <cstdio>
#include <mutex>

#include "host_design.h"

namespace {
constexpr size_t kNoAllocGeoBytes = 1024;   // EstGeometry at offset 0, band statistics behind it
constexpr size_t kNoAllocBandBytes = (size_t)SDRB_NOALLOC_MAX_FRAMES * sdrb::EST_MAX_BANDS * sizeof(sdrb::BandStat);
static_assert(sizeof(sdrb::EstGeometry) <= kNoAllocGeoBytes, "geometry does not fit its slot");
static_assert(sizeof(sdrb_frame_estimate) == sizeof(sdrb::FrameLocal), "sdrb_frame_estimate mirrors FrameLocal");
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocGeoBytes + kNoAllocBandBytes];

// the geometry arrives as a kernel parameter and is written by device code, which can name the array directly:
// no host-to-device copy from pageable memory, nothing allocated
__global__ void k_noalloc_put_geometry(const __grid_constant__ sdrb::EstGeometry g) {
    const int* src = reinterpret_cast<const int*>(&g);
    int* dst = reinterpret_cast<int*>(g_noalloc_scratch);
    for (int i = threadIdx.x; i < (int)(sizeof(sdrb::EstGeometry) / 4); i += blockDim.x) dst[i] = src[i];
}

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address every kernel is given
// comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr int kNoAllocMaxDevices = 64;
unsigned char* g_noalloc_ptr[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mu;
}  // namespace

extern "C" int sdrb_estimate_noalloc(int device, uint32_t sample_rate, int samples_per_reading, int focus_khz,
                                     const float* d_power, int n_frames, sdrb_frame_estimate* d_out, void* cuda_stream) {
    using namespace sdrb;
    char msg[256];
    auto bad = [&](int code, const char* text) { set_global_error(text); return code; };
    if (device < 0 || device >= kNoAllocMaxDevices) return bad(SDRB_ERR_ARG, "sdrb_estimate_noalloc: device out of range [0, 64)");
    if (!d_power || !d_out) return bad(SDRB_ERR_ARG, "sdrb_estimate_noalloc: d_power / d_out is NULL");
    if (sample_rate == 0 || samples_per_reading < 1) return bad(SDRB_ERR_ARG, "sdrb_estimate_noalloc: sample_rate and samples_per_reading must be positive");
    if (n_frames < 1 || n_frames > SDRB_NOALLOC_MAX_FRAMES) {
        snprintf(msg, sizeof msg, "sdrb_estimate_noalloc: n_frames %d outside [1, %d] (SDRB_NOALLOC_MAX_FRAMES: the static scratch "
                                  "holds that many frames; split the batch or use sdrb_process)", n_frames, SDRB_NOALLOC_MAX_FRAMES);
        return bad(SDRB_ERR_ARG, msg);
    }
    EstGeometry g;
    design_geometry(sample_rate, samples_per_reading, focus_khz, g);
    if (pooled_cap_of(g) > EST_SMEM_POOL_FLOATS) {
        snprintf(msg, sizeof msg, "sdrb_estimate_noalloc: focus range %d kHz pools %d bins per frame; above %d the estimator needs a "
                                  "per-frame global scratch that the static array does not hold (use sdrb_process)",
                 focus_khz, pooled_cap_of(g), EST_SMEM_POOL_FLOATS);
        return bad(SDRB_ERR_UNSUPPORTED, msg);
    }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return bad(SDRB_ERR_CUDA, cudaGetErrorString(e));
    unsigned char* base;
    {
        std::lock_guard<std::mutex> lock(g_noalloc_mu);
        if (!g_noalloc_ptr[device]) {   // set-up, once per device, with that device current
            unsigned char* ptr = nullptr;
            if ((e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch)) != cudaSuccess || !ptr ||
                (e = estimator_prepare()) != cudaSuccess) {
                snprintf(msg, sizeof msg, "sdrb_estimate_noalloc: set-up on device %d failed: %s", device, cudaGetErrorString(e));
                return bad(SDRB_ERR_CUDA, msg);
            }
            g_noalloc_ptr[device] = ptr;
        }
        base = g_noalloc_ptr[device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    k_noalloc_put_geometry<<<1, 64, 0, st>>>(g);
    if ((e = cudaGetLastError()) == cudaSuccess)
        e = launch_estimator(d_power, reinterpret_cast<const EstGeometry*>(base), g, reinterpret_cast<FrameLocal*>(d_out),
                             reinterpret_cast<float*>(base + kNoAllocGeoBytes), 0, SDRB_NOALLOC_MAX_FRAMES, n_frames, st);
    if (e != cudaSuccess) return bad(SDRB_ERR_CUDA, cudaGetErrorString(e));
    return SDRB_OK;
}

