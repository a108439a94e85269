// estimator.cu — evaluateSignalStrength on the GPU (src/dsp/fft_process.cpp:122-378).
//
// Two kernels:
//   k_estimator : one CTA per frame, everything that is frame-local (focus band, reference
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

constexpr int EST_THREADS = 128;
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

__global__ void __launch_bounds__(EST_THREADS)
k_estimator(const float* __restrict__ power, const EstGeometry* __restrict__ geo_p, FrameLocal* __restrict__ local,
            float* __restrict__ gscratch, int use_smem) {
    extern __shared__ float stage_smem[];
    __shared__ EstGeometry g;
    __shared__ BandStat st[11];
    __shared__ PoolMap pm;
    __shared__ int hist[256];
    __shared__ unsigned sh_prefix;
    __shared__ int sh_k;
    __shared__ float sh_perBinMean;
    __shared__ FrameLocal out;
    __shared__ float winMeanDb[10], winBest1kDb[10];

    const int tid = threadIdx.x;
    const size_t frame = blockIdx.x;
    {
        const int* src = reinterpret_cast<const int*>(geo_p);
        int* dst = reinterpret_cast<int*>(&g);
        for (int i = tid; i < (int)(sizeof(EstGeometry) / 4); i += EST_THREADS) dst[i] = src[i];
    }
    __syncthreads();
    if (g.focusLen <= 0) {  // :139 early return, the frame leaves every output untouched
        if (tid == 0) {
            FrameLocal z = {};
            z.peakBin = -1;
            z.best1kStart = -1;
            local[frame] = z;
        }
        return;
    }
    const float* P = power + frame * (size_t)g.n;
    float* stage = use_smem ? stage_smem : gscratch + frame * (size_t)g.totalBins;

    for (int seg = 0; seg <= g.nWin; seg++) {
        const int lo = seg == 0 ? g.focusLo : g.winLo[seg - 1];
        const int len = g.segOff[seg + 1] - g.segOff[seg];
        float* d = stage + g.segOff[seg];
        for (int i = tid; i < len; i += EST_THREADS) d[i] = P[lo + i];
    }
    __syncthreads();
    if (tid <= g.nWin) {
        const int len = g.segOff[tid + 1] - g.segOff[tid];
        band_scan(stage + g.segOff[tid], len, g.winBins1k, tid == 0 ? -1.0f : 0.0f, st[tid]);
    }
    __syncthreads();

    const bool valid = g.nWin >= 2;  // :218
    if (tid == 0) {
        FrameLocal o = {};
        // ---- :142-155 focus band
        const float dbMax = db10(st[0].maxP);
        if (dbMax > -130.0f) { o.absPeakDb = dbMax; o.peakBin = g.focusLo + st[0].maxIdx; }
        else { o.absPeakDb = -130.0f; o.peakBin = g.focusLo; }
        o.signalPowerDb = db10(__fdiv_rn(st[0].sum, (float)g.focusLen));
        o.valid = valid ? 1 : 0;
        o.best1kStart = -1;
        pm.nSeg = 0;
        pm.cum[0] = 0;
        if (valid) {
            // ---- :203-213 per-window dB figures
            int order[10];
            for (int w = 0; w < g.nWin; w++) {
                const int len = g.segOff[w + 2] - g.segOff[w + 1];
                winMeanDb[w] = db10(__fdiv_rn(st[w + 1].sum, (float)len));
                const float b1k = len < g.winBins1k ? __fdiv_rn(st[w + 1].sum, (float)len)
                                                    : __fdiv_rn(st[w + 1].bestRun, (float)g.winBins1k);
                winBest1kDb[w] = db10(b1k);
                order[w] = w;
            }
            // ---- :228-229 std::sort by meanDb (<= 10 elements: insertion sort, stable)
            for (int i = 1; i < g.nWin; i++) {
                const int w = order[i];
                int j = i - 1;
                while (j >= 0 && winMeanDb[w] < winMeanDb[order[j]]) { order[j + 1] = order[j]; j--; }
                order[j + 1] = w;
            }
            const int nb = g.nBottom;
            // ---- :235-247 mean-energy SNR
            float x[10];
            float mean = 0.0f;
            for (int i = 0; i < nb; i++) { x[i] = winMeanDb[order[i]]; mean = __fadd_rn(mean, x[i]); }
            mean = __fdiv_rn(mean, (float)nb);
            const float sigma = fmaxf(__fmul_rn(1.4816f, small_mad(x, nb, mean)), 0.5f);
            o.meanSnrDb = __fsub_rn(o.signalPowerDb, mean);
            o.meanSnrSigma = __fdiv_rn(o.meanSnrDb, sigma);
            // ---- :292-300 best-1 kHz reference statistics (sigma1k needs sigmaBin, finished below)
            float mean1k = 0.0f;
            for (int i = 0; i < nb; i++) { x[i] = winBest1kDb[order[i]]; mean1k = __fadd_rn(mean1k, x[i]); }
            mean1k = __fdiv_rn(mean1k, (float)nb);
            o.best1kHzSnrDb = mean1k;                                      // parked until the final step
            o.best1kHzSnrSigma = __fmul_rn(1.4816f, small_mad(x, nb, mean1k));
            // pooled-bin map (:252-257)
            pm.nSeg = nb < 4 ? nb : 4;
            for (int j = 0; j < pm.nSeg; j++) {
                const int w = order[j];
                pm.base[j] = g.segOff[w + 1];
                pm.cum[j + 1] = pm.cum[j] + (g.segOff[w + 2] - g.segOff[w + 1]);
            }
        }
        out = o;
    }
    __syncthreads();

    if (valid) {
        const int pooledN = pm.cum[pm.nSeg];
        // per-bin dB of the quiet windows, in place (:255-256)
        for (int i = tid; i < pooledN; i += EST_THREADS) {
            const int a = pool_addr(pm, i);
            stage[a] = db10(stage[a]);
        }
        __syncthreads();
        if (tid == 0) {  // :261-262 sequential mean
            float s = 0.0f;
            for (int j = 0; j < pm.nSeg; j++) {
                const float* d = stage + pm.base[j];
                const int len = pm.cum[j + 1] - pm.cum[j];
                for (int i = 0; i < len; i++) s = __fadd_rn(s, d[i]);
            }
            sh_perBinMean = __fdiv_rn(s, (float)pooledN);
        }
        __syncthreads();
        const float perBinMean = sh_perBinMean;
        for (int i = tid; i < pooledN; i += EST_THREADS) {  // :265-266
            const int a = pool_addr(pm, i);
            stage[a] = fabsf(__fsub_rn(stage[a], perBinMean));
        }
        __syncthreads();
        // :267-268 median of the gaps: radix select over the mapped array
        unsigned prefix = 0, mask = 0;
        int k = pooledN / 2;
        for (int shift = 24; shift >= 0; shift -= 8) {
            for (int i = tid; i < 256; i += EST_THREADS) hist[i] = 0;
            __syncthreads();
            for (int i = tid; i < pooledN; i += EST_THREADS) {
                const unsigned b = __float_as_uint(stage[pool_addr(pm, i)]);
                if ((b & mask) == prefix) atomicAdd(&hist[(b >> shift) & 255u], 1);
            }
            __syncthreads();
            if (tid == 0) {
                int acc = 0, d = 0;
                for (; d < 255; d++) {
                    if (acc + hist[d] > k) break;
                    acc += hist[d];
                }
                sh_prefix = prefix | ((unsigned)d << shift);
                sh_k = k - acc;
            }
            __syncthreads();
            prefix = sh_prefix;
            k = sh_k;
            mask |= 255u << shift;
        }
        if (tid == 0) {
            FrameLocal o = out;
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
            const float fb = g.focusLen < g.winBins1k ? __fdiv_rn(st[0].sum, (float)g.focusLen)
                                                      : __fdiv_rn(st[0].bestRun, (float)g.winBins1k);
            if (fb > 0.0f) {
                o.best1kHzSnrDb = __fsub_rn(db10(fb), mean1k);
                o.best1kHzSnrSigma = __fdiv_rn(o.best1kHzSnrDb, sigma1k);
                o.best1kStart = g.focusLo + st[0].bestStart;
            } else {
                o.best1kHzSnrDb = 0.0f;
                o.best1kHzSnrSigma = 0.0f;
                o.best1kStart = -1;
            }
            out = o;
        }
    }
    __syncthreads();
    if (tid == 0) local[frame] = out;
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
    return cudaFuncSetAttribute(k_estimator, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                EST_MAX_SMEM_FLOATS * (int)sizeof(float));
}

size_t estimator_scratch_floats(const EstGeometry& g, int n_frames) {
    return g.totalBins > EST_MAX_SMEM_FLOATS ? (size_t)g.totalBins * (size_t)n_frames : 0;
}

cudaError_t launch_estimator(const float* power, const EstGeometry* d_geo, const EstGeometry& h_geo,
                             FrameLocal* d_local, float* d_stage_scratch, int n_frames, cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    const int use_smem = h_geo.totalBins <= EST_MAX_SMEM_FLOATS ? 1 : 0;
    if (!use_smem && !d_stage_scratch) return cudaErrorInvalidValue;
    const size_t smem = use_smem ? (size_t)(h_geo.totalBins > 0 ? h_geo.totalBins : 1) * sizeof(float) : 0;
    k_estimator<<<n_frames, EST_THREADS, smem, st>>>(power, d_geo, d_local, d_stage_scratch, use_smem);
    return cudaGetLastError();
}

cudaError_t launch_tracker(const FrameLocal* d_local, const EstGeometry* d_geo, FftStreamState* d_state,
                           sdrb_frame_result* d_results, int n_streams, int n_blocks, cudaStream_t st) {
    if (n_streams <= 0 || n_blocks <= 0) return cudaSuccess;
    k_tracker<<<(n_streams + 127) / 128, 128, 0, st>>>(d_local, d_geo, d_state, d_results, n_streams, n_blocks);
    return cudaGetLastError();
}

}  // namespace sdrb
