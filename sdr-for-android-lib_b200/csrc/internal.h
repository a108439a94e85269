// internal.h — structures shared between the C-ABI (api.cu) and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "../../include/sdr_b200.h"
#include "fft_core.cuh"

namespace sdrb {

// ---------------------------------------------------------------- estimator
// Bin geometry of evaluateSignalStrength (src/dsp/fft_process.cpp:125-139,160,191-216),
// evaluated ONCE on the host in the reference's float expressions (no FMA contraction on
// the x86-64 baseline) so that every bin bound is bit-identical to the reference's.
struct EstGeometry {
    int n;                 // samplesPerReading
    int focusLo, focusHi, focusLen;   // focusLen <= 0: the reference returns early (:139)
    int winBins1k;
    int nWin;              // reference windows in push order (+k then -k), <= 10
    int winLo[10], winHi[10];
    int nBottom;           // max(1, int(nWin*0.4f))
    float freqPerBin;
    float nyquist;
    float sampleRateF;
    // scalars of :281-285,299 that depend on the geometry only, evaluated on the host with
    // the reference's float expressions (glibc logf/sqrtf)
    float sqrt2logN;       // sqrt(2 ln focusLen)
    float gumbelDen;       // sqrt(6.f) * sqrt2logN
    float sqrtWin1k;       // sqrt(float(winBins1k))
    // staging layout of the estimator kernel: segment 0 = focus band, 1..nWin = windows
    int segOff[12];        // float offsets, segOff[nWin+1] = total
    int totalBins;
};

// Frame-local result of the estimator kernel, before the per-stream sequential part.
struct FrameLocal {
    float absPeakDb, signalPowerDb;
    float meanSnrDb, meanSnrSigma;
    float perBinMean, peakAboveNoiseMeanDb;
    float maxBinSnrDb, maxBinSnrSigma;
    float best1kHzSnrDb, best1kHzSnrSigma;
    int best1kStart;       // bestStart bin, -1 when focusBest1kLinear <= 0 (:303,324-326)
    int peakBin;           // focusLo + peakBinInFocus
    int valid;             // nRef >= 2
    int pad;
};

// Carried per-stream state of FFTProcessor (src/dsp/fft_process.h:62-109) + hold-over outputs.
struct FftStreamState {
    uint32_t centerFrequency;
    int32_t centerChanged;           // sdr_bridge_internal::isCenterFrequencyChanged
    float trackingFrequency;
    float maxPeakDb, maxPeakHz;      // maxPeakAndFrequency
    int64_t timeOfLastMaxPeak_ns, timeOfLastMaxPeakUpdate_ns;
    int64_t frameCounter;            // frames processed so far (frame clock)
    int32_t peakConfirmed;
    int32_t flagRing[3];
    int32_t flagIndex;
    // outputs that persist when a frame does not overwrite them
    float meanSnrDb, meanSnrSigma, peakAboveNoiseMeanDb, maxBinSnrDb, maxBinSnrSigma;
    float best1kHzSnrDb, best1kHzSnrSigma, best1kHzCenterFreqHz, perBinMean;
    int32_t detectionFlag;
    int32_t pad;
};

// ---------------------------------------------------------------- SSB
// Per-stream carried filter state (the reference's function statics,
// src/ssb/ssb_demod_opt.cpp:261,277): previous two OUTPUTS of each section.
struct SsbStreamState {
    float rf_z1, rf_z2;
    float hp_z1, hp_z2;
    float bp_z1, bp_z2;
    float pad0, pad1;
};

struct BiquadCoef {
    float a0, a1, a2, b1, b2;
};

constexpr int SSB_MAX_TAPS = 255;

struct SsbParams {
    int n;                 // samples per block
    int decim;             // max(1, int(fs/48000.f))
    int nTaps;             // 255 (or n|1 when n < 255)
    int m;                 // PCM samples per block
    int upper;             // sideband flag passed to demodSSB
    float dcAlpha;         // 0.9995
    float agcTarget, agcFast, agcSlow;
    float gain, transientCoeff;
    BiquadCoef rf, hp, bp;
};

// ---------------------------------------------------------------- Airspy real->IQ front end
// Carried converter state (iqconverter_float_t: avg, fir_queue, delay_line) as the last 48 translated
// samples: 24 even ones for the half-band FIR (23 needed), 24 odd ones for the 12-sample delay.
constexpr int AIRSPY_TAIL = 48;
struct AirspyStreamState {
    float avg;
    float tail[AIRSPY_TAIL];   // oldest first
    uint32_t redone;           // chunks of the DC pass walked a second time (guessed start did not meet the true one)
    float pad[14];
};

// ---------------------------------------------------------------- timeline (SDRB_TRACE=1)
// Completion time of every launch group on its stream, relative to the start of the call: the launchers drop a
// CUDA event after each kernel when a trace is armed (api.cu prints it at the next sdrb_sync).  Diagnostic only.
struct TraceSink {
    virtual void mark(const char* name, cudaStream_t st) = 0;
    virtual ~TraceSink() {}
};
extern thread_local TraceSink* g_trace;
inline void trace_mark(const char* name, cudaStream_t st) { if (g_trace) g_trace->mark(name, st); }

// text for sdrb_last_error(NULL) from entry points that have no context (api.cu)
void set_global_error(const char* text);

// ---------------------------------------------------------------- launchers
// airspy.cu : raw 12-bit real samples (uint16, 2 per complex output) -> cf32 [streams][n_real/2]
cudaError_t launch_airspy_front(const uint16_t* d_raw, AirspyStreamState* d_state, cpx* d_iq, float* d_work,
                                int chunk, int warm, int n_streams, long long n_real, cudaStream_t st, int* launches);
void airspy_chunking(int* chunk, int* warm);
size_t airspy_work_floats(int n_streams, long long n_real, int chunk);
// fft16k2.cu : second-generation 16K kernel (512 threads x 32 points, TMA-staged input)
cudaError_t launch_fft16k2(int fmt, const void* iq, float* power, const cpx* tables, int n_frames,
                           int sm_count, cudaStream_t st, int front = 0, const float* win = nullptr,
                           int* ticket = nullptr);
cudaError_t fft16k2_prepare(const cpx* host_tables);   // make_f2_tables() layout; T2 goes to the constant bank
// fft_any.cu : generic power-of-two path (N != 16384)
// fft_small.cu : one-CTA shared-memory kernel for 256 .. 8192 points (tables: make_fs_tables, fft_tables.h)
cudaError_t fft_small_prepare();
bool fft_small_supports(int n);
int fft_small_table_elems(int n);
// front / win: the extension front end (bit 1: subtract the block mean, bit 0: window table of n floats), fused into phase 1
cudaError_t launch_fft_small(int fmt, const void* iq, float* power, const cpx* tables, int n, int n_frames, int sm_count,
                             cudaStream_t st, int front = 0, const float* win = nullptr);
// two-pass scheme for 32768 / 131072 / 262144 points: radix-A columns pass + the one-CTA kernel on the rows
// (tables: make_tp_tables(n, B, log2 A) with B = 4096 (8192 for 262144); Y: n_frames * n complex of scratch)
bool fft_twopass_supports(int n);
int fft_twopass_table_elems(int n);
cudaError_t launch_fft_twopass(int fmt, const void* iq, float* power, cpx* Y, const cpx* tables, int n, int n_frames,
                               int sm_count, cudaStream_t st);
cudaError_t launch_fft_generic(int fmt, const void* iq, float* power, cpx* scratch_a, cpx* scratch_b,
                               const cpx* twiddles, int n, int n_frames, cudaStream_t st, int* launches);
int fft_generic_passes(int n);
// any other length (Bluestein over the power-of-two passes): M = bluestein_m(n) >= 2n - 1
int bluestein_m(int n);
cudaError_t launch_fft_bluestein(int fmt, const void* iq, float* power, cpx* sa, cpx* sb, const cpx* chirp,
                                 const cpx* bf, const cpx* tw_m, int n, int m, int n_frames, cudaStream_t st,
                                 int* launches);
// extension front end (window / block-mean removal) for the multi-pass paths: raw IQ -> cf32
cudaError_t launch_fft_front(int fmt, const void* iq, cpx* out, const float* win, int n, int n_frames, int front,
                             cudaStream_t st);
// fft_big.cu : multi-CTA four-step FFT (N = 65536, 1048576), bulk-TMA staged row tiles
cudaError_t fft_big_prepare();
bool fft_big_supports(int n);
int fft_big_table_elems(int n);
int fft_big_best_group(int n, int max_group);   // frames per L2-resident group that fill whole waves of the rows kernel
cudaError_t launch_fft_big(int fmt, const void* iq, float* power, cpx* A, const cpx* tables, int n, int n_frames,
                           cudaStream_t st);
// fused persistent variant (one kernel, intermediate in an L2-resident ring of `slots` frames; ctl: 1 + 2 n_frames words)
int fft_big_ring_slots(int n);
cudaError_t launch_fft_big_fused(int fmt, const void* iq, float* power, cpx* ring, int slots, unsigned* ctl,
                                 const cpx* tables, int n, int n_frames, cudaStream_t st);
// estimator.cu
bool estimator_inline_scan(const EstGeometry& g);   // true: launch_estimator is ONE kernel (short bands scanned inline)
cudaError_t launch_estimator(const float* power, const EstGeometry* d_geo, const EstGeometry& h_geo,
                             FrameLocal* d_local, float* scratch, size_t frame0, int cap_frames, int n_frames,
                             cudaStream_t st);
cudaError_t estimator_prepare();
size_t estimator_scratch_floats(const EstGeometry& g, int n_frames);
cudaError_t launch_tracker(const FrameLocal* d_local, const EstGeometry* d_geo, FftStreamState* d_state,
                           sdrb_frame_result* d_results, int n_streams, int n_blocks, cudaStream_t st);
// estimator.cu : dB view of linear power, out = float (fmt 1) or __half (fmt 2); 10*log10f(p + 1e-20f)
cudaError_t launch_power_view(const float* power, void* out, size_t count, int fmt, cudaStream_t st);
// ssb.cu.  d_taps is the [decim][ssb_smax] slot-major tap table (split decimator, extended chain);
// d_tapsE the rotated table of the fused pass (ssb_fused_tap_table), sp = ssb_fused_slots (0: split pair).
int ssb_smax(const SsbParams& P);
int ssb_fused_slots(const SsbParams& P, int fmt);
int ssb_ext_fused_slots(const SsbParams& P);   // the same for the extended chain's pass (0: split pair)
int ssb_fused_rows(const SsbParams& P);
std::vector<float> ssb_fused_tap_table(const float* h, const SsbParams& P, int sp);
cudaError_t ssb_prepare();
cudaError_t launch_ssb(int fmt, const void* iq, const SsbParams& P, const float* d_taps, const float* d_tapsE, int sp,
                       float2* d_start, SsbStreamState* d_state, float* d_a48, int16_t* d_pcm, float* d_agc,
                       int n_streams, int n_blocks, int sm_count, cudaStream_t st, int* launches);
// any block length (bytes not a multiple of 64): one thread per stream, scalar loads; scratch [n_streams][P.n] floats
cudaError_t launch_ssb_any(int fmt, const void* iq, const SsbParams& P, const float* taps_lin, SsbStreamState* d_state,
                           float* scratch, int16_t* d_pcm, int n_streams, int n_blocks, cudaStream_t st);
cudaError_t launch_ssb_audio(const float* d_a48, const SsbParams& P, SsbStreamState* d_state, int16_t* d_pcm,
                             int n_streams, int n_blocks, cudaStream_t st);
// ssb_ext.cu : NCO channel bank + Weaver sideband selection (extension stages)
size_t ssb_ext_work_floats(int n, int S, int C, int B, int true_sideband);   // X (removeDC output) + M (mixed samples)
cudaError_t launch_ssb_ext(int fmt, const void* iq, const SsbParams& P, const float* d_taps, const float* d_tapsE, int sp,
                           const uint32_t* d_inc1, uint32_t inc2, int true_sideband, float4* d_start,
                           SsbStreamState* d_state, float* d_a48, int16_t* d_pcm, float* d_agc, float* d_work,
                           int n_streams, int channels, int n_blocks, unsigned long long base_samples, cudaStream_t st,
                           int* launches);
// ssb.cu : 255-tap decimator over post-AGC samples [n_frames][P.n] -> [n_frames][P.m]
cudaError_t launch_ssb_fir(const float* xagc, const float* taps, float* a48, const SsbParams& P, int n_frames,
                           cudaStream_t st);
// synth.cu
cudaError_t launch_synth(int fmt, uint32_t fs, int first_stream, int n_streams, int64_t first_sample,
                         int64_t n_samples, void* d_out, const int16_t* d_sine, cudaStream_t st);
void synth_host(int fmt, uint32_t fs, int first_stream, int n_streams, int64_t first_sample,
                int64_t n_samples, void* out);
const int16_t* synth_sine_table();  // 4096-entry Q15 sine, host memory

}  // namespace sdrb
