// estimator.cu — evaluateSignalStrength on the GPU (src/dsp/fft_process.cpp:122-378).
//
// Two kernels:
//   k_estimator : one WARP per frame, everything that is frame-local (focus band, reference
//                 windows, sort, MAD sigmas, pooled per-bin noise, Gumbel / best-1 kHz SNR).
//   k_tracker   : one thread per stream, the sequential part that carries state from frame to
//                 frame (frequency tracking :333-361, detection flag :365-378, hold-over of
//                 outputs on skipped / invalid frames) and writes the sdrb_frame_result records.
//
// Float sums that the reference does sequentially (signalPowerSum :148, window sums :199,
// the sliding run-sum :175, perBinMean :261) are done sequentially here too, one lane per
// band, with un-fused adds, so that they round exactly like the reference; only log10 differs
// (double-precision log10 rounded once vs glibc log10f).
#include "internal.h"

namespace sdrb {

constexpr int EST_MAX_WARPS = 4;   // frames per CTA (one warp each)
constexpr int EST_MAX_SMEM_FLOATS = 48 * 1024;  // 192 KiB of staged bins; above that: global scratch

__device__ __forceinline__ float db10(float p) {
    // 10.0f * log10f(p / refPower + 1e-20f), refPower = 1  (src/dsp/fft_process.cpp:149)
    return __fmul_rn(10.0f, (float)log10((double)__fadd_rn(p, 1e-20f)));
}

struct BandStat {
    float sum;       // sequential sum of the band
    float maxP;      // max power, strict >, first index
    int maxIdx;      // offset inside the band
    float bestRun;   // max sliding run-sum of winBins1k bins (or the plain sum when shorter)
    int bestStart;   // offset of the first window that reaches bestRun
};

// Sequential pass over one band held at p[0..len).  Mirrors :146-155 (sum + arg-max),
// :198-202 (sum + max) and best1kHzMean :163-180 / :310-318 (run-sum, strict >).
__device__ void band_scan(const float* p, int len, int W, float maxInit, BandStat& r) {
    float sum = 0.0f, maxP = maxInit;
    int maxIdx = 0;
#pragma unroll 8
    for (int i = 0; i < len; i++) {
        const float v = p[i];
        sum = __fadd_rn(sum, v);
        if (v > maxP) { maxP = v; maxIdx = i; }
    }
    r.sum = sum; r.maxP = maxP; r.maxIdx = maxIdx;
    if (len < W) { r.bestRun = sum; r.bestStart = 0; return; }
    float run = 0.0f;
    for (int i = 0; i < W; i++) run = __fadd_rn(run, p[i]);
    float best = run;
    int bestStart = 0;
#pragma unroll 4
    for (int s = 1; s + W - 1 < len; s++) {
        run = __fadd_rn(run, __fsub_rn(p[s + W - 1], p[s - 1]));
        if (run > best) { best = run; bestStart = s; }
    }
    r.bestRun = best; r.bestStart = bestStart;
}

// Median-of-gaps helper for <= 10 values (:238-242, :296-300): sorted |x - mean| [nb/2].
__device__ float small_mad(const float* x, int nb, float mean) {
    float g[10];
    for (int i = 0; i < nb; i++) g[i] = fabsf(__fsub_rn(x[i], mean));
    for (int i = 1; i < nb; i++) {
        const float t = g[i];
        int j = i - 1;
        while (j >= 0 && g[j] > t) { g[j + 1] = g[j]; j--; }
        g[j + 1] = t;
    }
    return g[nb / 2];
}

struct PoolMap {
    int nSeg;
    int base[4];   // stage offset of each pooled segment (sorted-window order)
    int cum[5];    // cumulative lengths, cum[nSeg] = pooledN
};

__device__ __forceinline__ int pool_addr(const PoolMap& pm, int i) {
    int j = 0;
    while (j + 1 < pm.nSeg && i >= pm.cum[j + 1]) j++;
    return pm.base[j] + (i - pm.cum[j]);
}

// Per-warp working set: one warp owns one frame, so the only block-wide barrier is the one that
// publishes the geometry; everything else is __syncwarp.
struct EstWarp {
    BandStat st[11];
    PoolMap pm;
    int hist[256];
    float winMeanDb[10], winBest1kDb[10];
    float focusDb[2];          // absPeakDb candidate, signalPowerDb
    FrameLocal out;
    float perBinMean;
};

template <bool USE_SMEM>
__global__ void __launch_bounds__(EST_MAX_WARPS * 32)
k_estimator(const float* __restrict__ power, const EstGeometry* __restrict__ geo_p, FrameLocal* __restrict__ local,
            float* __restrict__ gscratch, int n_frames) {
    extern __shared__ float stage_smem[];
    __shared__ EstGeometry g;
    __shared__ EstWarp ws[EST_MAX_WARPS];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    {
        const int* src = reinterpret_cast<const int*>(geo_p);
        int* dst = reinterpret_cast<int*>(&g);
        for (int i = threadIdx.x; i < (int)(sizeof(EstGeometry) / 4); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const size_t frame = (size_t)blockIdx.x * nwarps + warp;
    if (frame >= (size_t)n_frames) return;
    EstWarp& W = ws[warp];
    if (g.focusLen <= 0) {  // :139 early return, the frame leaves every output untouched
        if (lane == 0) {
            FrameLocal z = {};
            z.peakBin = -1;
            z.best1kStart = -1;
            local[frame] = z;
        }
        return;
    }
    const float* P = power + frame * (size_t)g.n;
    float* stage = USE_SMEM ? stage_smem + (size_t)warp * g.totalBins : gscratch + frame * (size_t)g.totalBins;

    for (int seg = 0; seg <= g.nWin; seg++) {
        const int lo = seg == 0 ? g.focusLo : g.winLo[seg - 1];
        const int len = g.segOff[seg + 1] - g.segOff[seg];
        float* d = stage + g.segOff[seg];
        for (int i = lane; i < len; i += 32) d[i] = P[lo + i];
    }
    __syncwarp();
    const bool valid = g.nWin >= 2;  // :218
    if (lane <= g.nWin) {
        const int len = g.segOff[lane + 1] - g.segOff[lane];
        BandStat bs;
        band_scan(stage + g.segOff[lane], len, g.winBins1k, lane == 0 ? -1.0f : 0.0f, bs);
        W.st[lane] = bs;
        // the dB figures of this band, one lane per band (:149-155, :203-213)
        if (lane == 0) {
            W.focusDb[0] = db10(bs.maxP);
            W.focusDb[1] = db10(__fdiv_rn(bs.sum, (float)g.focusLen));
        } else if (valid) {
            W.winMeanDb[lane - 1] = db10(__fdiv_rn(bs.sum, (float)len));
            const float b1k = len < g.winBins1k ? __fdiv_rn(bs.sum, (float)len)
                                                : __fdiv_rn(bs.bestRun, (float)g.winBins1k);
            W.winBest1kDb[lane - 1] = db10(b1k);
        }
    }
    __syncwarp();

    if (lane == 0) {
        FrameLocal o = {};
        // ---- :142-155 focus band
        const float dbMax = W.focusDb[0];
        if (dbMax > -130.0f) { o.absPeakDb = dbMax; o.peakBin = g.focusLo + W.st[0].maxIdx; }
        else { o.absPeakDb = -130.0f; o.peakBin = g.focusLo; }
        o.signalPowerDb = W.focusDb[1];
        o.valid = valid ? 1 : 0;
        o.best1kStart = -1;
        W.pm.nSeg = 0;
        W.pm.cum[0] = 0;
        if (valid) {
            // ---- :228-229 std::sort by meanDb (<= 10 elements: insertion sort, stable)
            int order[10];
            for (int w = 0; w < g.nWin; w++) order[w] = w;
            for (int i = 1; i < g.nWin; i++) {
                const int w = order[i];
                int j = i - 1;
                while (j >= 0 && W.winMeanDb[w] < W.winMeanDb[order[j]]) { order[j + 1] = order[j]; j--; }
                order[j + 1] = w;
            }
            const int nb = g.nBottom;
            // ---- :235-247 mean-energy SNR
            float x[10];
            float mean = 0.0f;
            for (int i = 0; i < nb; i++) { x[i] = W.winMeanDb[order[i]]; mean = __fadd_rn(mean, x[i]); }
            mean = __fdiv_rn(mean, (float)nb);
            const float sigma = fmaxf(__fmul_rn(1.4816f, small_mad(x, nb, mean)), 0.5f);
            o.meanSnrDb = __fsub_rn(o.signalPowerDb, mean);
            o.meanSnrSigma = __fdiv_rn(o.meanSnrDb, sigma);
            // ---- :292-300 best-1 kHz reference statistics (sigma1k needs sigmaBin, finished below)
            float mean1k = 0.0f;
            for (int i = 0; i < nb; i++) { x[i] = W.winBest1kDb[order[i]]; mean1k = __fadd_rn(mean1k, x[i]); }
            mean1k = __fdiv_rn(mean1k, (float)nb);
            o.best1kHzSnrDb = mean1k;                                      // parked until the final step
            o.best1kHzSnrSigma = __fmul_rn(1.4816f, small_mad(x, nb, mean1k));
            // pooled-bin map (:252-257)
            W.pm.nSeg = nb < 4 ? nb : 4;
            for (int j = 0; j < W.pm.nSeg; j++) {
                const int w = order[j];
                W.pm.base[j] = g.segOff[w + 1];
                W.pm.cum[j + 1] = W.pm.cum[j] + (g.segOff[w + 2] - g.segOff[w + 1]);
            }
        }
        W.out = o;
    }
    __syncwarp();

    if (valid) {
        const PoolMap pm = W.pm;
        const int pooledN = pm.cum[pm.nSeg];
        // per-bin dB of the quiet windows, in place (:255-256)
        for (int i = lane; i < pooledN; i += 32) {
            const int a = pool_addr(pm, i);
            stage[a] = db10(stage[a]);
        }
        __syncwarp();
        if (lane == 0) {  // :261-262 sequential mean
            float s = 0.0f;
            for (int j = 0; j < pm.nSeg; j++) {
                const float* d = stage + pm.base[j];
                const int len = pm.cum[j + 1] - pm.cum[j];
#pragma unroll 8
                for (int i = 0; i < len; i++) s = __fadd_rn(s, d[i]);
            }
            W.perBinMean = __fdiv_rn(s, (float)pooledN);
        }
        __syncwarp();
        const float perBinMean = W.perBinMean;
        for (int i = lane; i < pooledN; i += 32) {  // :265-266
            const int a = pool_addr(pm, i);
            stage[a] = fabsf(__fsub_rn(stage[a], perBinMean));
        }
        __syncwarp();
        // :267-268 median of the gaps: 4-pass radix select on the bit pattern; each lane owns 8 of
        // the 256 histogram bins and the crossing bin is found with a warp scan
        unsigned prefix = 0, mask = 0;
        int k = pooledN / 2;
        for (int shift = 24; shift >= 0; shift -= 8) {
            for (int i = lane; i < 256; i += 32) W.hist[i] = 0;
            __syncwarp();
            for (int i = lane; i < pooledN; i += 32) {
                const unsigned b = __float_as_uint(stage[pool_addr(pm, i)]);
                if ((b & mask) == prefix) atomicAdd(&W.hist[(b >> shift) & 255u], 1);
            }
            __syncwarp();
            int mine = 0;
#pragma unroll
            for (int j = 0; j < 8; j++) mine += W.hist[8 * lane + j];
            int incl = mine;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += v;
            }
            const unsigned cross = __ballot_sync(0xffffffffu, incl > k);
            const int L = cross ? __ffs(cross) - 1 : 31;
            int digit = 0, acc = 0;
            if (lane == L) {
                acc = incl - mine;
                digit = 8 * L;
                for (int j = 0; j < 7; j++) {
                    const int h = W.hist[8 * L + j];
                    if (acc + h > k) break;
                    acc += h;
                    digit++;
                }
            }
            digit = __shfl_sync(0xffffffffu, digit, L);
            acc = __shfl_sync(0xffffffffu, acc, L);
            prefix |= (unsigned)digit << shift;
            k -= acc;
            mask |= 255u << shift;
            __syncwarp();
        }
        if (lane == 0) {
            FrameLocal o = W.out;
            const float sigmaBin = fmaxf(__fmul_rn(1.4816f, __uint_as_float(prefix)), 1.0f);
            o.perBinMean = perBinMean;
            o.peakAboveNoiseMeanDb = __fsub_rn(o.absPeakDb, perBinMean);                     // :274
            // :281-288 Gumbel-corrected max-bin SNR
            const float gumbelLoc = __fadd_rn(perBinMean, __fmul_rn(sigmaBin, g.sqrt2logN));
            const float gumbelSig = fmaxf(__fdiv_rn(__fmul_rn(sigmaBin, 3.14159f), g.gumbelDen), 0.5f);
            o.maxBinSnrDb = __fsub_rn(o.absPeakDb, gumbelLoc);
            o.maxBinSnrSigma = __fdiv_rn(o.maxBinSnrDb, gumbelSig);
            // :299-327 best-1 kHz SNR
            const float mean1k = o.best1kHzSnrDb;
            const float sigmaFloor1k = __fdiv_rn(sigmaBin, g.sqrtWin1k);
            const float sigma1k = fmaxf(fmaxf(o.best1kHzSnrSigma, sigmaFloor1k), 0.5f);
            const float fb = g.focusLen < g.winBins1k ? __fdiv_rn(W.st[0].sum, (float)g.focusLen)
                                                      : __fdiv_rn(W.st[0].bestRun, (float)g.winBins1k);
            if (fb > 0.0f) {
                o.best1kHzSnrDb = __fsub_rn(db10(fb), mean1k);
                o.best1kHzSnrSigma = __fdiv_rn(o.best1kHzSnrDb, sigma1k);
                o.best1kStart = g.focusLo + W.st[0].bestStart;
            } else {
                o.best1kHzSnrDb = 0.0f;
                o.best1kHzSnrSigma = 0.0f;
                o.best1kStart = -1;
            }
            W.out = o;
        }
    }
    __syncwarp();
    if (lane == 0) local[frame] = W.out;
}

// Frame clock (SURVEY.md 8a note T): t = (frame + 1) * n / fs in ns, integer arithmetic.
__device__ __forceinline__ long long frame_clock_ns(long long frame_index, int n, unsigned fs) {
    const long long q = (frame_index + 1) * (long long)n;
    return (q / fs) * 1000000000LL + ((q % fs) * 1000000000LL) / fs;
}

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
    for (int b = 0; b < n_blocks; b++) {
        const FrameLocal f = local[(size_t)s * n_blocks + b];
        sdrb_frame_result r = {};
        const long long now_ns = frame_clock_ns(st.frameCounter, g.n, fs);
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

cudaError_t estimator_prepare() {
    return cudaFuncSetAttribute(k_estimator<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                EST_MAX_SMEM_FLOATS * (int)sizeof(float));
}

size_t estimator_scratch_floats(const EstGeometry& g, int n_frames) {
    return g.totalBins > EST_MAX_SMEM_FLOATS ? (size_t)g.totalBins * (size_t)n_frames : 0;
}

cudaError_t launch_estimator(const float* power, const EstGeometry* d_geo, const EstGeometry& h_geo,
                             FrameLocal* d_local, float* d_stage_scratch, int n_frames, cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    const int bins = h_geo.totalBins > 0 ? h_geo.totalBins : 1;
    const int use_smem = bins <= EST_MAX_SMEM_FLOATS ? 1 : 0;
    if (!use_smem && !d_stage_scratch) return cudaErrorInvalidValue;
    // frames (warps) per CTA: as many as fit in ~48 KiB of staging, 1..EST_MAX_WARPS
    int warps = use_smem ? (12 * 1024) / bins : EST_MAX_WARPS;
    warps = warps < 1 ? 1 : (warps > EST_MAX_WARPS ? EST_MAX_WARPS : warps);
    const size_t smem = use_smem ? (size_t)bins * warps * sizeof(float) : 0;
    const int grid = (n_frames + warps - 1) / warps;
    if (use_smem)
        k_estimator<true><<<grid, warps * 32, smem, st>>>(power, d_geo, d_local, d_stage_scratch, n_frames);
    else
        k_estimator<false><<<grid, warps * 32, 0, st>>>(power, d_geo, d_local, d_stage_scratch, n_frames);
    return cudaGetLastError();
}

cudaError_t launch_tracker(const FrameLocal* d_local, const EstGeometry* d_geo, FftStreamState* d_state,
                           sdrb_frame_result* d_results, int n_streams, int n_blocks, cudaStream_t st) {
    if (n_streams <= 0 || n_blocks <= 0) return cudaSuccess;
    k_tracker<<<(n_streams + 127) / 128, 128, 0, st>>>(d_local, d_geo, d_state, d_results, n_streams, n_blocks);
    return cudaGetLastError();
}

}  // namespace sdrb
