/* oracle/sdr_oracle.c — CPU restatement of the reference hot path.  See sdr_oracle.h.
 * TEST INFRASTRUCTURE ONLY; parity pinned against oracle/_ref (the reference itself).
 * Build: gcc -O2 -ffp-contract=off (no -march: x86-64 baseline has no FMA, as the
 * reference build), so float expressions round exactly like the reference's. */
#include "sdr_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ===================================================================== ingest */

void sdr_oracle_convert_u8(const uint8_t* in, int n_complex, float* iq_out) {
    /* ssb_demod_opt.cpp:37-41 */
    const float offset = 127.4f;
    const float scale = 1.0f / 128.0f;
    for (int i = 0; i < n_complex; ++i) {
        float I = (in[2 * i] - offset) * scale;
        float Q = (in[2 * i + 1] - offset) * scale;
        iq_out[2 * i] = I;
        iq_out[2 * i + 1] = Q;
    }
}

void sdr_oracle_convert_s16(const int16_t* in, int n_complex, float* iq_out) {
    for (int i = 0; i < 2 * n_complex; ++i) iq_out[i] = (float)in[i] * (1.0f / 32768.0f);
}

/* ================================================================ FFT + power */

static int is_pow2(int n) { return n > 0 && (n & (n - 1)) == 0; }

/* Forward DFT in double precision, X[k] = sum x[n] exp(-2 pi i k n / N)
 * (fft_process.cpp:77-78: FFTW_FORWARD, unnormalised). */
/* In-place radix-2 transform in double, sign -1 (m a power of two). */
static void fft_pow2_double(double* re, double* im, int m) {
    int bits = 0;
    while ((1 << bits) < m) bits++;
    for (int i = 0; i < m; i++) {
        unsigned r = 0;
        for (int b = 0; b < bits; b++) if (i & (1 << b)) r |= 1u << (bits - 1 - b);
        if ((int)r > i) { double t = re[i]; re[i] = re[r]; re[r] = t; t = im[i]; im[i] = im[r]; im[r] = t; }
    }
    for (int len = 2; len <= m; len <<= 1) {
        int half = len >> 1;
        for (int j = 0; j < half; j++) {
            double a = -2.0 * M_PI * (double)j / (double)len, wr = cos(a), wi = sin(a);
            for (int base = 0; base < m; base += len) {
                int u = base + j, v = u + half;
                double tr = re[v] * wr - im[v] * wi, ti = re[v] * wi + im[v] * wr;
                re[v] = re[u] - tr; im[v] = im[u] - ti;
                re[u] += tr; im[u] += ti;
            }
        }
    }
}

/* The DFT of any length in double (what fftwf_plan_dft_1d(sampCount, ...) computes, fft_process.cpp:77-79), for
 * lengths too long for the O(n^2) sum: chirp-z identity n k = (n^2 + k^2 - (k - n)^2) / 2 over a power-of-two
 * circular convolution.  Checker arithmetic only (double throughout). */
static int dft_bluestein_double(const float* iq, int n, double* re, double* im) {
    int m = 1;
    while (m < 2 * n - 1) m <<= 1;
    double* ar = (double*)calloc((size_t)m, sizeof(double));
    double* ai = (double*)calloc((size_t)m, sizeof(double));
    double* br = (double*)calloc((size_t)m, sizeof(double));
    double* bi = (double*)calloc((size_t)m, sizeof(double));
    double* cr = (double*)malloc(sizeof(double) * (size_t)n);
    double* ci = (double*)malloc(sizeof(double) * (size_t)n);
    if (!ar || !ai || !br || !bi || !cr || !ci) { free(ar); free(ai); free(br); free(bi); free(cr); free(ci); return -1; }
    for (long long k = 0; k < n; k++) {
        double a = -M_PI * (double)((k * k) % (2LL * n)) / (double)n;
        cr[k] = cos(a); ci[k] = sin(a);
        ar[k] = iq[2 * k] * cr[k] - iq[2 * k + 1] * ci[k];
        ai[k] = iq[2 * k] * ci[k] + iq[2 * k + 1] * cr[k];
        br[k] = cr[k]; bi[k] = -ci[k];
        if (k) { br[m - k] = cr[k]; bi[m - k] = -ci[k]; }
    }
    fft_pow2_double(ar, ai, m);
    fft_pow2_double(br, bi, m);
    for (int i = 0; i < m; i++) {   /* conj(A * B): the next forward transform is then the inverse one */
        double tr = ar[i] * br[i] - ai[i] * bi[i], ti = ar[i] * bi[i] + ai[i] * br[i];
        ar[i] = tr; ai[i] = -ti;
    }
    fft_pow2_double(ar, ai, m);
    for (int k = 0; k < n; k++) {
        double vr = ar[k] / (double)m, vi = -ai[k] / (double)m;
        re[k] = vr * cr[k] - vi * ci[k];
        im[k] = vr * ci[k] + vi * cr[k];
    }
    free(ar); free(ai); free(br); free(bi); free(cr); free(ci);
    return 0;
}

static int dft_double(const float* iq, int n, double* re, double* im) {
    if (is_pow2(n) && n >= 2) {
        int bits = 0;
        while ((1 << bits) < n) bits++;
        for (int i = 0; i < n; i++) {
            unsigned r = 0;
            for (int b = 0; b < bits; b++) if (i & (1 << b)) r |= 1u << (bits - 1 - b);
            re[r] = iq[2 * i];
            im[r] = iq[2 * i + 1];
        }
        double* wr = (double*)malloc(sizeof(double) * (n / 2 + 1));
        double* wi = (double*)malloc(sizeof(double) * (n / 2 + 1));
        if (!wr || !wi) { free(wr); free(wi); return -1; }
        for (int k = 0; k < n / 2; k++) {
            double a = -2.0 * M_PI * (double)k / (double)n;
            wr[k] = cos(a);
            wi[k] = sin(a);
        }
        for (int len = 2; len <= n; len <<= 1) {
            int half = len >> 1, step = n / len;
            for (int base = 0; base < n; base += len) {
                for (int j = 0; j < half; j++) {
                    double cr = wr[j * step], ci = wi[j * step];
                    double xr = re[base + j + half], xi = im[base + j + half];
                    double tr = xr * cr - xi * ci, ti = xr * ci + xi * cr;
                    re[base + j + half] = re[base + j] - tr;
                    im[base + j + half] = im[base + j] - ti;
                    re[base + j] += tr;
                    im[base + j] += ti;
                }
            }
        }
        free(wr);
        free(wi);
        return 0;
    }
    if (n < 1) return -2;
    if (n > 4096) return dft_bluestein_double(iq, n, re, im);
    for (int k = 0; k < n; k++) {
        double sr = 0.0, si = 0.0;
        for (int j = 0; j < n; j++) {
            double a = -2.0 * M_PI * (double)(((long long)k * j) % n) / (double)n;
            double c = cos(a), s = sin(a);
            sr += iq[2 * j] * c - iq[2 * j + 1] * s;
            si += iq[2 * j] * s + iq[2 * j + 1] * c;
        }
        re[k] = sr;
        im[k] = si;
    }
    return 0;
}

int sdr_oracle_power_shifted(const float* iq, int n, float* power_shifted) {
    double* re = (double*)malloc(sizeof(double) * n);
    double* im = (double*)malloc(sizeof(double) * n);
    if (!re || !im) { free(re); free(im); return -1; }
    int rc = dft_double(iq, n, re, im);
    if (rc == 0) {
        int half = n / 2;
        /* fft_process.cpp:83-97: power then swap halves (loops to half only). */
        for (int i = 0; i < half; i++) {
            float ar = (float)re[i + half], ai = (float)im[i + half];
            float br = (float)re[i], bi = (float)im[i];
            power_shifted[i] = ar * ar + ai * ai;
            power_shifted[i + half] = br * br + bi * bi;
        }
        if (n & 1) power_shifted[n - 1] = 0.0f; /* resize() value-initialises; odd n leaves the last bin 0 */
    }
    free(re);
    free(im);
    return rc;
}

/* ================================================================== estimator */

int64_t sdr_oracle_frame_clock_ns(int64_t frame_index, int n, uint32_t fs) {
    int64_t q = (frame_index + 1) * (int64_t)n;
    return (q / fs) * 1000000000LL + ((q % fs) * 1000000000LL) / fs;
}

void sdr_oracle_fft_init(sdr_oracle_fft_state* s, uint32_t fc, uint32_t fs, int n, int focus_khz) {
    memset(s, 0, sizeof(*s));
    s->centerFrequency = fc;
    s->sampleRate = fs;
    s->samplesPerReading = n;
    s->freqFocusRangeKhz = focus_khz;
    /* fft_process.cpp:33-35 */
    s->maxPeakDb = -130.0f;
    s->maxPeakHz = (float)fc;
    s->peakBin = -1;
}

int64_t sdr_oracle_tracking_long(const sdr_oracle_fft_state* s) {
    return (int64_t)roundf(s->trackingFrequency); /* fft_process.h:41 (std::round(float)) */
}

typedef struct { float meanDb, maxBinDb, best1kDb; int lo, hi; } ref_window;

static int cmp_float(const void* a, const void* b) {
    float x = *(const float*)a, y = *(const float*)b;
    return (x > y) - (x < y);
}

/* fft_process.cpp:163-180 */
static float best_1khz_mean(const float* p, int lo, int hi, int winBins1k) {
    const int len = hi - lo + 1;
    if (len <= 0) return 0.0f;
    if (len < winBins1k) {
        float s = 0.0f;
        for (int i = lo; i <= hi; i++) s += p[i];
        return s / len;
    }
    float runSum = 0.0f;
    for (int i = lo; i < lo + winBins1k; i++) runSum += p[i];
    float best = runSum / winBins1k;
    for (int start = lo + 1; start + winBins1k - 1 <= hi; start++) {
        runSum += p[start + winBins1k - 1] - p[start - 1];
        const float m = runSum / winBins1k;
        if (m > best) best = m;
    }
    return best;
}

void sdr_oracle_evaluate(sdr_oracle_fft_state* st, int n, const float* power_shifted, int64_t now_ns) {
    const uint32_t sampleRate = st->sampleRate, centerFrequency = st->centerFrequency;
    const float refPower = 1.0f;                                      /* fft_process.h:74-75 */
    const float freqPerBin = (float)sampleRate / (float)n;            /* :125 */
    const float X_hz = st->freqFocusRangeKhz * 1000.0f;               /* :126 */
    const float nyquist = sampleRate / 2.0f;                          /* :127 */
#define OFF_TO_BIN(off) ((int)(((off) + nyquist) / freqPerBin))       /* :131-133 */

    st->peakBin = -1;
    /* :136-139 */
    int focusLo = OFF_TO_BIN(-X_hz); if (focusLo < 0) focusLo = 0;
    int focusHi = OFF_TO_BIN(+X_hz) - 1; if (focusHi > n - 1) focusHi = n - 1;
    const int focusLen = focusHi - focusLo + 1;
    if (focusLen <= 0) return;

    /* :142-155 */
    float absPeakDb = -130.0f;
    int peakBinInFocus = 0;
    float signalPowerSum = 0.0f;
    for (int i = focusLo; i <= focusHi; i++) {
        float p = power_shifted[i];
        signalPowerSum += p;
        float dB = 10.0f * log10f(p / refPower + 1e-20f);
        if (dB > absPeakDb) { absPeakDb = dB; peakBinInFocus = i - focusLo; }
    }
    const float signalPowerDb = 10.0f * log10f((signalPowerSum / focusLen) / refPower + 1e-20f);
    st->absPeakDb = absPeakDb;
    st->signalPowerDb = signalPowerDb;
    st->peakBin = focusLo + peakBinInFocus;

    /* :160 */
    int winBins1k = (int)ceilf(1000.0f / freqPerBin); if (winBins1k < 1) winBins1k = 1;

    /* :182-216 */
    ref_window win[10];
    int nRef = 0;
    for (int k = 1; k <= 5; k++) {
        const float nearX = (4 * k - 2) * X_hz;
        const float farX = 4 * k * X_hz;
        if (farX >= nyquist) break;
        for (int side = 0; side < 2; side++) {
            int lo, hi;
            if (side == 0) { lo = OFF_TO_BIN(+nearX); hi = OFF_TO_BIN(+farX) - 1; }
            else           { lo = OFF_TO_BIN(-farX);  hi = OFF_TO_BIN(-nearX) - 1; }
            if (lo < 0) lo = 0;
            if (hi > n - 1) hi = n - 1;
            if (hi <= lo) continue;
            const int cnt = hi - lo + 1;
            float sum = 0.0f, maxP = 0.0f;
            for (int i = lo; i <= hi; i++) {
                sum += power_shifted[i];
                if (power_shifted[i] > maxP) maxP = power_shifted[i];
            }
            win[nRef].meanDb = 10.0f * log10f((sum / cnt) / refPower + 1e-20f);
            win[nRef].maxBinDb = 10.0f * log10f(maxP / refPower + 1e-20f);
            win[nRef].best1kDb = 10.0f * log10f(best_1khz_mean(power_shifted, lo, hi, winBins1k) / refPower + 1e-20f);
            win[nRef].lo = lo;
            win[nRef].hi = hi;
            nRef++;
        }
    }

    const int valid = (nRef >= 2);   /* :218-219 */
    st->valid = valid;
    if (!valid) {
        /* :220-225 */
        st->meanSnrDb = st->meanSnrSigma = 0.0f;
        st->peakAboveNoiseMeanDb = st->maxBinSnrDb = st->maxBinSnrSigma = 0.0f;
        st->best1kHzSnrDb = st->best1kHzSnrSigma = 0.0f;
    } else {
        /* :228-229 std::sort of <= 10 elements == insertion sort (stable) in libstdc++ */
        for (int i = 1; i < nRef; i++) {
            ref_window w = win[i];
            int j = i - 1;
            while (j >= 0 && w.meanDb < win[j].meanDb) { win[j + 1] = win[j]; j--; }
            win[j + 1] = w;
        }
        int nBottom = (int)(nRef * 0.4f); if (nBottom < 1) nBottom = 1;   /* :232 */

        /* :235-247 */
        {
            float mean = 0.0f;
            for (int i = 0; i < nBottom; i++) mean += win[i].meanDb;
            mean /= nBottom;
            float gaps[10];
            for (int i = 0; i < nBottom; i++) gaps[i] = fabsf(win[i].meanDb - mean);
            qsort(gaps, nBottom, sizeof(float), cmp_float);
            float sigma = 1.4816f * gaps[nBottom / 2]; if (sigma < 0.5f) sigma = 0.5f;
            const float snrDb = signalPowerDb - mean;
            st->meanSnrDb = snrDb;
            st->meanSnrSigma = snrDb / sigma;
        }

        /* :252-269 */
        size_t pooledN = 0;
        for (int j = 0; j < nBottom; j++) pooledN += (size_t)(win[j].hi - win[j].lo + 1);
        float* pooled = (float*)malloc(sizeof(float) * (pooledN ? pooledN : 1));
        size_t q = 0;
        for (int j = 0; j < nBottom; j++)
            for (int i = win[j].lo; i <= win[j].hi; i++)
                pooled[q++] = 10.0f * log10f(power_shifted[i] / refPower + 1e-20f);
        float sigmaBin = 1.0f;
        float perBinMean = 0.0f;
        if (pooledN) {
            for (size_t i = 0; i < pooledN; i++) perBinMean += pooled[i];
            perBinMean /= (float)pooledN;
            st->perBinMean = perBinMean;
            for (size_t i = 0; i < pooledN; i++) pooled[i] = fabsf(pooled[i] - perBinMean);
            qsort(pooled, pooledN, sizeof(float), cmp_float);
            sigmaBin = 1.4816f * pooled[pooledN / 2]; if (sigmaBin < 1.0f) sigmaBin = 1.0f;
        }
        free(pooled);

        st->peakAboveNoiseMeanDb = absPeakDb - perBinMean;   /* :274 */

        /* :281-288 */
        {
            const float logN = logf((float)focusLen);
            const float sqrt2logN = sqrtf(2.0f * logN);
            const float gumbelLoc = perBinMean + sigmaBin * sqrt2logN;
            float gumbelSig = sigmaBin * 3.14159f / (sqrtf(6.0f) * sqrt2logN); if (gumbelSig < 0.5f) gumbelSig = 0.5f;
            st->maxBinSnrDb = absPeakDb - gumbelLoc;
            st->maxBinSnrSigma = st->maxBinSnrDb / gumbelSig;
        }

        /* :292-327 */
        {
            float mean1k = 0.0f;
            for (int i = 0; i < nBottom; i++) mean1k += win[i].best1kDb;
            mean1k /= nBottom;
            float gaps[10];
            for (int i = 0; i < nBottom; i++) gaps[i] = fabsf(win[i].best1kDb - mean1k);
            qsort(gaps, nBottom, sizeof(float), cmp_float);
            const float sigmaFloor1k = sigmaBin / sqrtf((float)winBins1k);
            float sigma1k = 1.4816f * gaps[nBottom / 2];
            if (sigma1k < sigmaFloor1k) sigma1k = sigmaFloor1k;
            if (sigma1k < 0.5f) sigma1k = 0.5f;

            const float focusBest1kLinear = best_1khz_mean(power_shifted, focusLo, focusHi, winBins1k);
            if (focusBest1kLinear > 0.0f) {
                const float focusBest1kDb = 10.0f * log10f(focusBest1kLinear / refPower + 1e-20f);
                st->best1kHzSnrDb = focusBest1kDb - mean1k;
                st->best1kHzSnrSigma = st->best1kHzSnrDb / sigma1k;
                const int len = focusHi - focusLo + 1;
                int bestStart = focusLo;
                if (len >= winBins1k) {
                    float rs = 0.f;
                    for (int i = focusLo; i < focusLo + winBins1k; i++) rs += power_shifted[i];
                    float bv = rs;
                    for (int s = focusLo + 1; s + winBins1k - 1 <= focusHi; s++) {
                        rs += power_shifted[s + winBins1k - 1] - power_shifted[s - 1];
                        if (rs > bv) { bv = rs; bestStart = s; }
                    }
                }
                st->best1kHzCenterFreqHz = (bestStart + winBins1k / 2) * freqPerBin
                                           + ((float)centerFrequency - nyquist);
            } else {
                st->best1kHzSnrDb = st->best1kHzSnrSigma = 0.0f;
            }
        }
    }
#undef OFF_TO_BIN

    /* ---- :333-361 frequency tracking, frame clock instead of steady_clock ---- */
    if (st->trackingFrequency == 0.0f) st->trackingFrequency = (float)centerFrequency;
    if (st->centerChanged) {
        st->trackingFrequency = (float)centerFrequency;
        st->centerChanged = 0;
    }
    if (valid && absPeakDb > st->maxPeakDb) {
        st->maxPeakDb = absPeakDb;
        st->maxPeakHz = (float)((focusLo + peakBinInFocus) * freqPerBin + (centerFrequency - nyquist));
        st->timeOfLastMaxPeak_ns = now_ns;
    }
    {
        /* duration_cast<milliseconds> truncates toward zero */
        const int64_t msSincePeak = (now_ns - st->timeOfLastMaxPeak_ns) / 1000000LL;
        if (st->timeOfLastMaxPeakUpdate_ns < st->timeOfLastMaxPeak_ns && msSincePeak > 300) {
            st->trackingFrequency = st->maxPeakHz;
            st->timeOfLastMaxPeakUpdate_ns = now_ns;
            st->maxPeakDb = -130.0f;
        }
    }

    /* ---- :365-378 detection flag ---- */
    const int aboveThreshold = valid && (st->meanSnrSigma >= 4.0f);
    if (aboveThreshold) { if (st->peakConfirmed < 1) st->peakConfirmed++; }
    else st->peakConfirmed = 0;
    const int currentFlag = (aboveThreshold && st->peakConfirmed >= 1) ? 3 : 0;
    st->flagRing[st->flagIndex] = currentFlag;
    st->flagIndex = (st->flagIndex + 1) % 3;
    int mx = st->flagRing[0];
    if (st->flagRing[1] > mx) mx = st->flagRing[1];
    if (st->flagRing[2] > mx) mx = st->flagRing[2];
    st->detectionFlag = mx;
}

/* ======================================================================= SSB */

void sdr_oracle_remove_dc(float* iq, int n, float alpha) {
    /* ssb_demod_opt.cpp:49-55.  std::complex<float> scalar*complex is component-wise. */
    float dcr = 0.0f, dci = 0.0f;
    for (int i = 0; i < n; i++) {
        dcr = alpha * dcr + (1.0f - alpha) * iq[2 * i];
        dci = alpha * dci + (1.0f - alpha) * iq[2 * i + 1];
        iq[2 * i] -= dcr;
        iq[2 * i + 1] -= dci;
    }
}

void sdr_oracle_iir2_init_lowpass(sdr_oracle_biquad* f, float fs, float fc, float Q) {
    /* ssb_demod_opt.cpp:60-73; 2.0f * M_PI promotes to double, result rounds to float. */
    float w0 = 2.0f * M_PI * fc / fs;
    float cosw0 = cosf(w0);
    float sinw0 = sinf(w0);
    float alpha = sinw0 / (2.0f * Q);
    float norm = 1.0f / (1.0f + alpha);
    f->a0 = (1.0f - cosw0) / 2.0f * norm;
    f->a1 = (1.0f - cosw0) * norm;
    f->a2 = f->a0;
    f->b1 = -2.0f * cosw0 * norm;
    f->b2 = (1.0f - alpha) * norm;
    f->z1 = f->z2 = 0.0f;
}

void sdr_oracle_iir2_process(sdr_oracle_biquad* f, float* iq, int n) {
    /* ssb_demod_opt.cpp:75-84: real part only, z1/z2 are previous OUTPUTS, writes {y,y}. */
    float z1r = f->z1, z2r = f->z2;
    for (int i = 0; i < n; ++i) {
        float x = iq[2 * i];
        float y = f->a0 * x + f->a1 * z1r + f->a2 * z2r - f->b1 * z1r - f->b2 * z2r;
        z2r = z1r; z1r = y;
        iq[2 * i] = y;
        iq[2 * i + 1] = y;
    }
    f->z1 = z1r; f->z2 = z2r;
}

void sdr_oracle_demod_ssb(const float* iq, int n, int upper, float* audio) {
    /* ssb_demod_opt.cpp:89-96 */
    for (int i = 0; i < n; ++i)
        audio[i] = upper ? (iq[2 * i] + iq[2 * i + 1]) : (iq[2 * i] - iq[2 * i + 1]);
}

void sdr_oracle_adaptive_agc(float* audio, int n, float target, float fastAttack, float slowDecay) {
    /* ssb_demod_opt.cpp:101-115 */
    float gain = 1.0f;
    for (int i = 0; i < n; ++i) {
        float x = audio[i];
        float mag = fabsf(x) + 1e-8f;
        float desired = target / (sqrtf(mag) + 1e-6f);
        float rate = (desired < gain) ? fastAttack : slowDecay;
        gain = gain * (1.0f - rate) + desired * rate;
        float v = x * gain;
        audio[i] = v < -1.0f ? -1.0f : (v > 1.0f ? 1.0f : v);
    }
}

int sdr_oracle_fir_taps(int n_in, int decim, float cutoffRel, float* h) {
    /* ssb_demod_opt.cpp:122-134 */
    int N = 255;
    if (N > n_in) N = n_in | 1;
    int M = N - 1;
    float fc = cutoffRel / decim;
    for (int n = 0; n < N; n++) {
        int k = n - M / 2;
        /* the ?: has a double and a float arm -> double, then narrows to float */
        float sinc = (k == 0) ? 2.0f * M_PI * fc : (double)(sinf(2.0f * M_PI * fc * k) / (float)k);
        float w = 0.5f - 0.5f * cosf(2.0f * M_PI * n / M);
        h[n] = (sinc / M_PI) * w;
    }
    float sum = 0.0f;
    for (int n = 0; n < N; n++) sum += h[n];
    if (sum != 0.0f) for (int n = 0; n < N; n++) h[n] /= sum;
    return N;
}

int sdr_oracle_fir_decimate(const float* in, int n_in, int decim, float cutoffRel, float* out, int cap) {
    /* ssb_demod_opt.cpp:121-143 */
    float h[256];
    if (n_in <= 0) return 0;
    int N = sdr_oracle_fir_taps(n_in, decim, cutoffRel, h);
    int m = 0;
    for (size_t i = 0; i + (size_t)N <= (size_t)n_in; i += (size_t)decim) {
        float acc = 0.0f;
        for (int k = 0; k < N; k++) acc += in[i + k] * h[k];
        if (m < cap) out[m] = acc;
        m++;
    }
    return m;
}

void sdr_oracle_biquad_init_highpass(sdr_oracle_biquad* f, float fs, float f0, float Q) {
    /* ssb_demod_opt.cpp:148-164 */
    float w0 = 2.0f * M_PI * f0 / fs;
    float cosw0 = cosf(w0);
    float sinw0 = sinf(w0);
    float alpha = sinw0 / (2.0f * Q);
    float b0 = (1 + cosw0) / 2.0f;
    float b1 = -(1 + cosw0);
    float b2 = (1 + cosw0) / 2.0f;
    float a0 = 1 + alpha;
    float a1 = -2 * cosw0;
    float a2 = 1 - alpha;
    f->a0 = b0 / a0; f->a1 = b1 / a0; f->a2 = b2 / a0;
    f->b1 = a1 / a0; f->b2 = a2 / a0;
    f->z1 = f->z2 = 0.0f;
}

void sdr_oracle_biquad_init_bandpass(sdr_oracle_biquad* f, float fs, float f0, float Q) {
    /* ssb_demod_opt.cpp:166-175 */
    float w0 = 2.0f * M_PI * f0 / fs;
    float alpha = sinf(w0) / (2.0f * Q);
    float cosw0 = cosf(w0);
    float b0 = alpha, b1 = 0.0f, b2 = -alpha;
    float a0 = 1.0f + alpha, a1 = -2.0f * cosw0, a2 = 1.0f - alpha;
    f->a0 = b0 / a0; f->a1 = b1 / a0; f->a2 = b2 / a0;
    f->b1 = a1 / a0; f->b2 = a2 / a0;
    f->z1 = f->z2 = 0.0f;
}

void sdr_oracle_biquad_process(sdr_oracle_biquad* f, float* x, int n) {
    /* ssb_demod_opt.cpp:177-186 */
    float z1 = f->z1, z2 = f->z2;
    for (int i = 0; i < n; ++i) {
        float in = x[i];
        float y = f->a0 * in + f->a1 * z1 + f->a2 * z2 - f->b1 * z1 - f->b2 * z2;
        z2 = z1; z1 = y;
        x[i] = y;
    }
    f->z1 = z1; f->z2 = z2;
}

void sdr_oracle_transient_boost(float* x, int n, float coeff) {
    /* ssb_demod_opt.cpp:191-198 */
    float prev = 0.0f;
    for (int i = 0; i < n; ++i) {
        float diff = x[i] - prev;
        prev = x[i];
        x[i] = x[i] + coeff * diff;
    }
}

void sdr_oracle_float_to_pcm(const float* in, int n, float gain, int16_t* out) {
    /* ssb_demod_opt.cpp:203-210 */
    for (int i = 0; i < n; ++i) {
        float v = in[i] * gain;
        v = v < -1.0f ? -1.0f : (v > 1.0f ? 1.0f : v);
        out[i] = (int16_t)(v * 32767.0f);
    }
}

void sdr_oracle_ssb_init(sdr_oracle_ssb_state* s) {
    memset(s, 0, sizeof(*s));
    /* file-static defaults, ssb_demod_opt.cpp:17-26 */
    s->agcTarget = 0.35f; s->agcFast = 0.006f; s->agcSlow = 0.00035f; s->gain = 0.5f;
    s->lowpassBd = 3200.0f; s->lowpassQ = 0.9f; s->transientCoeff = 0.55f;
}

int sdr_oracle_process_ssb(sdr_oracle_ssb_state* s, const float* iq_in, int n_in, uint32_t fs,
                           int upper, int mode, int16_t* pcm_out, int cap) {
    /* ssb_demod_opt.cpp:224-227: block length latched on the first call; iq.resize() pads with zeros */
    if (!s->initialised) s->sampCount = n_in;
    const int n = s->sampCount;
    float* iq = (float*)calloc((size_t)(n > 0 ? n : 1) * 2, sizeof(float));
    float* audio = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
    memcpy(iq, iq_in, sizeof(float) * 2 * (size_t)(n_in < n ? n_in : n));

    /* :230-255 */
    if (mode == 2) {
        s->agcTarget = 0.45f; s->agcFast = 0.008f; s->gain = 4.5f;
        s->lowpassBd = 2200.0f; s->lowpassQ = 1.2f; s->transientCoeff = 0.7f;
    } else if (mode == 0) {
        s->agcTarget = 0.45f; s->agcFast = 0.008f; s->gain = 10.0f;
        s->lowpassBd = 2200.0f; s->lowpassQ = 1.2f; s->transientCoeff = 0.7f;
    } else if (mode == 1) {
        s->agcTarget = 0.35f; s->agcFast = 0.006f; s->agcSlow = 0.00035f; s->gain = 0.5f;
        s->lowpassBd = 3200.0f; s->lowpassQ = 0.9f; s->transientCoeff = 0.55f;
    }

    sdr_oracle_remove_dc(iq, n, 0.9995f);                                   /* :258 */
    if (!s->initialised) {
        sdr_oracle_iir2_init_lowpass(&s->rf, (float)fs, s->lowpassBd, s->lowpassQ);   /* :261-263 */
        sdr_oracle_biquad_init_highpass(&s->hp, 48000.0f, 1200.0f, 0.7f);   /* :277-282 */
        sdr_oracle_biquad_init_bandpass(&s->bp, 48000.0f, 2400.0f, 0.6f);
        s->initialised = 1;
    }
    sdr_oracle_iir2_process(&s->rf, iq, n);                                 /* :264 */
    sdr_oracle_demod_ssb(iq, n, upper, audio);                              /* :267 */
    sdr_oracle_adaptive_agc(audio, n, s->agcTarget, s->agcFast, 0.00035f);  /* :270 */

    int decim = (int)(fs / 48000.0f); if (decim < 1) decim = 1;             /* :273 */
    int mcap = n / decim + 4;
    float* a48 = (float*)malloc(sizeof(float) * (size_t)mcap);
    int m = sdr_oracle_fir_decimate(audio, n, decim, 0.45f, a48, mcap);     /* :274 */
    if (m > 0) {                                                            /* :283-289 */
        sdr_oracle_biquad_process(&s->hp, a48, m);
        sdr_oracle_biquad_process(&s->bp, a48, m);
        sdr_oracle_transient_boost(a48, m, s->transientCoeff);
    }
    int16_t* pcm = (int16_t*)malloc(sizeof(int16_t) * (size_t)(m > 0 ? m : 1));
    sdr_oracle_float_to_pcm(a48, m, s->gain, pcm);                          /* :295 */
    memcpy(pcm_out, pcm, sizeof(int16_t) * (size_t)(m < cap ? m : cap));
    free(pcm); free(a48); free(audio); free(iq);
    return m;
}
