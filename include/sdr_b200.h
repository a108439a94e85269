/* include/sdr_b200.h — C-ABI of the B200-native IQ hot path.
 *
 * This is the drop-in boundary for the reference's per-reading DSP path.  The
 * reference (alexandreGellibert/SDR-for-Android-lib) has no FFI table for this
 * path: its JNI bridge owns three C++ globals and calls them directly
 *     FFTProcessor fftProcessor; SSBProcessor ssbProcessor;        src/sdr-bridge-java-soapy.cpp:101-108
 *     fftProcessor.process(buf,len) / ssbProcessor.enqueueData()    src/sdr-bridge-java-soapy.cpp:441-444
 * so the entry points below are what a cgo/JNI/ctypes binding of those objects
 * would bind, batch-first: S independent streams x B readings per call.
 * Header-compatible C++ front ends (FFTProcessor / SSBProcessor / processSSB_opt,
 * batch of one) live in sdr-for-android-lib_b200/host/ and call only this ABI.
 *
 * Conventions: plain pointers and sizes, no C++/torch types.  Every function
 * returns SDRB_OK (0) or a negative sdrb_status; sdrb_last_error() gives text.
 * The reference's DSP has no error channel at all (all void), so
 * status codes are additive.  There is NO CPU fallback: without a CUDA device or
 * with an unsupported shape the call fails loudly.
 *
 * Paths cited as file:line are relative to the reference repository root. */
#ifndef SDR_B200_H
#define SDR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SDRB_ABI_VERSION 2

typedef enum {
    SDRB_OK = 0,
    SDRB_ERR_ARG = -1,        /* bad argument / shape */
    SDRB_ERR_CUDA = -2,       /* CUDA runtime error (text in sdrb_last_error) */
    SDRB_ERR_UNSUPPORTED = -3,/* valid in the reference, not built here (e.g. non power-of-two N) */
    SDRB_ERR_STATE = -4,      /* call order (not configured, nothing processed, ...) */
    SDRB_ERR_NOMEM = -5
} sdrb_status;

/* Wire formats of one complex sample (interleaved I,Q).
 *  CU8 : RTL-SDR offset binary; f = (b - 127.4f) * (1/128.f)     src/ssb/ssb_demod_opt.cpp:33-44
 *  CS16: LimeSDR-style; f = s * 2^-15  (LimeSuite is a missing submodule: convention, parity unpinned)
 *  CF32: what SoapySDR hands the reference today ("CF32")        src/sdr-bridge-java-soapy.cpp:262-263
*  RU16: Airspy raw ADC stream, SURVEY.md 8(f) row 1: REAL 12-bit samples in uint16 at twice the complex
 *        rate (two words per complex output sample), turned into IQ by the libairspy converter chain
 *        (x-2048)/2048, DC removal, fs/4 translation, 47-tap half-band, delay
 *        third_party/libairspy/libairspy/src/airspy.c:311-321, iqconverter_float.c:418-475
 *        with its state (DC average, filter history) carried per stream across calls. */
typedef enum { SDRB_CU8 = 0, SDRB_CS16 = 1, SDRB_CF32 = 2, SDRB_RU16 = 3 } sdrb_format;

#define SDRB_STAGE_FFT 1u   /* FFTProcessor::process       src/dsp/fft_process.cpp:42-105 */
#define SDRB_STAGE_SSB 2u   /* processSSB_opt              src/ssb/ssb_demod_opt.cpp:221-296 */
/* the two beacon-pulse detectors that consume the path's outputs. */
#define SDRB_STAGE_SPECTRAL_PULSE 4u /* SpectralPulseDetector::process per reading (needs SDRB_STAGE_FFT)
                                        src/dsp/spectral_pulse_detector.cpp:10-13, call site
                                        src/sdr-bridge-java-soapy.cpp:477-487 */
#define SDRB_STAGE_AUDIO_PULSE 8u    /* AudioPulseDetector::process(pcm) per reading (needs SDRB_STAGE_SSB)
                                        src/ssb/audio_pulse_detector.cpp:112-125, call site
                                        src/ssb/ssb_processor.cpp:109-113 */

/* Per-context configuration, shared by all streams of the context.
 * First block = FftProcessorConfig (src/dsp/fft_process.h:20-26) + the BridgeConfig
 * fields the path reads (src/bridge-config.h:41-49).  Per-stream centre
 * frequencies can be changed afterwards with sdrb_set_center_frequency. */
typedef struct {
    uint32_t centerFrequency;     /* Hz */
    uint32_t sampleRate;          /* Hz */
    int32_t  samplesPerReading;   /* N = block length = FFT length */
    int32_t  freqFocusRangeKhz;
    int32_t  soundMode;           /* 0,1,2: src/ssb/ssb_demod_opt.cpp:230-255 */
    int32_t  inputFormat;         /* sdrb_format */
    int32_t  upperSideband;       /* the reference always passes true: src/ssb/ssb_processor.cpp:103 */
    /* Extension stages the reference does not have; 0 = identity = reference behaviour. */
    int32_t  fftWindow;           /* 0 rectangular (reference), 1 Hann */
    int32_t  fftRemoveDc;         /* 0 off (reference), 1 subtract the block mean before the FFT */
    int32_t  ssbTrueSideband;     /* 0 reference quirk (USB=2y, LSB=0), 1 phasing-method USB/LSB */
    int32_t  ssbChannels;         /* 0/1 = one channel at the tuned centre (reference); >1 = NCO bank */
    float    ssbChannelOffsetHz;  /* NCO offset of channel 0 from centre */
    float    ssbChannelSpacingHz; /* raster between channels */
    int32_t  reserved[3];
} sdrb_config;

/* One record per (stream, reading): everything soapyCallback reads back through the
 * getters (src/dsp/fft_process.h:38-55; use at src/sdr-bridge-java-soapy.cpp:446-455)
 * plus the frame-local internals the parity tests compare. */
typedef struct {
    float   meanSnrDb;              /* (F)V "signal strength" */
    float   meanSnrSigma;
    float   peakAboveNoiseMeanDb;   /* (F)V */
    float   maxBinSnrDb;            /* (FF)V */
    float   maxBinSnrSigma;
    float   best1kHzSnrDb;          /* (FF)V */
    float   best1kHzSnrSigma;
    float   best1kHzCenterFreqHz;
    float   perBinMean;             /* noise level (F)V */
    float   trackingFrequencyHz;    /* float state behind getTrackingFrequency() */
    int64_t trackingFrequency;      /* std::round of the above, (J)V */
    int32_t detectionFlag;          /* 0 or 3, (I)V */
    /* frame-local internals (src/dsp/fft_process.cpp:142-155,344-348) */
    float   absPeakDb;
    float   signalPowerDb;
    float   peakFrequencyHz;        /* candidate computed for this frame's peak bin */
    int32_t peakBin;                /* focusLo + peakBinInFocus in the shifted spectrum; -1 if skipped */
    int32_t valid;                  /* nRef >= 2 */
    int32_t nRef;
    int32_t reserved;
} sdrb_frame_result;

/* AudioPulseDetector::Config (src/ssb/audio_pulse_detector.h:20-38), field for field. */
typedef struct {
    float   sampleRate, fMin, fMax, fsEnergy, smoothCutoff, zDefaultS, tTargetInit, dtTolS;
    float   snrMin, snrRhythm, snrStrong, dispersionMax;
    int32_t sumNMax;
    float   liveWindowT, liveDivisor;
    int32_t noiseRefFar, noiseRefNear;
} sdrb_audio_pulse_config;

/* One record per (stream, reading) [x channel for the NCO bank: (stream, channel, reading)]:
 * the state both detectors expose after that reading's process() call. */
typedef struct {
    /* SpectralPulseDetector -> (FIJ)V payload (src/sdr-bridge-java-soapy.cpp:480-486) */
    float   spectralSnrSigma;      /* the frame's best1kHzSnrSigma as handed to process() */
    int32_t spectralLiveEtat;      /* liveEtat() 0..5 */
    int64_t spectralEstFreqHz;     /* llround(estimatedFreqHz()) */
    float   spectralEstFreqHzF;    /* estimatedFreqHz() (src/dsp/spectral_pulse_detector.cpp:173-196) */
    int32_t spectralLevel;         /* PulseLevel 0..3 */
    int32_t spectralLocked;
    float   spectralLastSnr;       /* lastPulseStrength() */
    float   spectralTTarget;       /* lockedPeriodS() */
    /* AudioPulseDetector -> (FI)V payload (src/ssb/ssb_processor.cpp:110-113) */
    float   audioLastSnr;          /* lastPulseStrength() */
    int32_t audioLiveEtat;
    int32_t audioLevel;
    int32_t audioLocked;
    float   audioTTarget;
    int32_t overflow;              /* ROI store overflow count (never in the reference: deques grow) */
    int32_t reserved;
} sdrb_pulse_result;

/* What the configuration resolves to: bin geometry of evaluateSignalStrength
 * (src/dsp/fft_process.cpp:125-139,160,191-216,232) and the SSB chain's parameters
 * (src/ssb/ssb_demod_opt.cpp:60-73,122-134,148-175,230-255,273).  Pure host computation. */
typedef struct {
    int32_t n, focusLo, focusHi, focusLen, winBins1k, nWin, nBottom;
    int32_t winLo[10], winHi[10];
    float   freqPerBin, nyquist;
    int32_t decim, nTaps, pcmPerBlock;
    int32_t fftPath;               /* 0 generic multi-pass, 1 single-CTA 16K kernel, 2 four-step, 3 chirp-z, 4 single-CTA 256..8192, 5 two-pass 32768/131072/262144 */
    float   rf[5], hp[5], bp[5];   /* a0,a1,a2,b1,b2 of struct IIR2 / Biquad (src/ssb/ssb_demod_opt.h:14-26) */
    float   agcTarget, agcFast, agcSlow, gain, transientCoeff, dcAlpha;
    float   firTaps[255];
} sdrb_plan;

typedef struct sdrb_ctx sdrb_ctx;

/* ---- lifecycle ---------------------------------------------------------- */
int  sdrb_abi_version(void);
/* Creates a context on CUDA device `device` (one context per GPU; one process per GPU when sharded). */
int  sdrb_create(int device, sdrb_ctx** out);
int  sdrb_destroy(sdrb_ctx* ctx);
/* Text of the last error on this context (ctx may be NULL for create-time errors). */
const char* sdrb_last_error(const sdrb_ctx* ctx);
/* Run all work of this context on a caller-owned cudaStream_t (NULL = the context's own stream). */
int  sdrb_set_cuda_stream(sdrb_ctx* ctx, void* cuda_stream);

/* ---- configuration (FFTProcessor::configure, BridgeConfig) -------------- */
/* src/dsp/fft_process.cpp:20-39, src/sdr-bridge-java-soapy.cpp:1123-1128.
 * Allocates device buffers for n_streams x max_blocks readings and resets all per-stream state. */
int  sdrb_configure(sdrb_ctx* ctx, const sdrb_config* cfg, int n_streams, int max_blocks);
/* A new geometry (sample rate, block length, focus, mode, centre) for the SAME streams, keeping every carried
 * state: the reference's FFTProcessor::configure keeps its tracking / detection members (src/dsp/fft_process.cpp:20-39),
 * the SSB filters are function statics (src/ssb/ssb_demod_opt.cpp:261,277) and the audio pulse detector is a member
 * of SSBProcessor (src/ssb/ssb_processor.h:52): none of them restarts when the bridge pushes a new configuration.
 * The spectral pulse detector is rebuilt (its frame rate follows the geometry).  A new centre frequency raises
 * the isCenterFrequencyChanged side channel.  The channel count cannot change. */
int  sdrb_replan(sdrb_ctx* ctx, const sdrb_config* cfg, int max_blocks);
/* setFrequency: new centre for one stream (or all with stream = -1) and raise the
 * isCenterFrequencyChanged side channel (src/sdr-bridge-java-soapy.cpp:897-905, src/sdr-bridge-internal.h:9). */
int  sdrb_set_center_frequency(sdrb_ctx* ctx, int stream, uint32_t center_hz);
/* setFrequencyFocusRange (src/sdr-bridge-java-soapy.cpp:1031-1036) */
int  sdrb_set_focus_range(sdrb_ctx* ctx, int focus_khz);
/* BridgeConfig::setSoundMode (src/bridge-config.h:61); only gains/AGC/boost follow the new mode,
 * the RF low-pass keeps the first mode like the reference's static rfFilter (src/ssb/ssb_demod_opt.cpp:261-263). */
int  sdrb_set_sound_mode(sdrb_ctx* ctx, int mode);
/* Number of int16 PCM samples produced per reading: floor((N-255)/decim)+1 (src/ssb/ssb_demod_opt.cpp:137). */
int  sdrb_pcm_per_block(const sdrb_ctx* ctx);

/* Resolve a configuration without touching a GPU (used by host-side tests and tools). */
int  sdrb_plan_config(const sdrb_config* cfg, sdrb_plan* out);

/* ---- the hot path -------------------------------------------------------- */
/* One batched submit = soapyCallback for every (stream, reading)
 * (src/sdr-bridge-java-soapy.cpp:424-493): iq is [n_streams][n_blocks][N] samples of
 * cfg.inputFormat, in device memory (on_device=1) or host memory (on_device=0; copied
 * inside the call).  Borrowed until the call returns (host) / until sdrb_sync (device).
 * stages: SDRB_STAGE_FFT | SDRB_STAGE_SSB.  Asynchronous on the context's stream. */
int  sdrb_process(sdrb_ctx* ctx, const void* iq, int on_device, int n_blocks, unsigned stages);
int  sdrb_sync(sdrb_ctx* ctx);
/* The same call host-to-host, pipelined: iq_host in, every payload of soapyCallback back out
 * (records = the scalar callbacks, power = ([F)V, pcm = ([S)V; any output may be NULL and then
 * stays on the device).  Streams are cut into chunks whose H2D copy, kernels and D2H copies
 * overlap on three CUDA streams; pass pinned host memory for full overlap.  Synchronous. */
int  sdrb_process_host(sdrb_ctx* ctx, const void* iq_host, int n_blocks, unsigned stages,
                       sdrb_frame_result* results_out, float* power_out, int16_t* pcm_out);

/* Results of the last sdrb_process, [n_streams][n_blocks] records (synchronises). */
int  sdrb_get_results(sdrb_ctx* ctx, sdrb_frame_result* host_out, size_t capacity_records);
/* Device-resident records (same layout), valid until the next sdrb_process. */
int  sdrb_results_device(sdrb_ctx* ctx, const sdrb_frame_result** dev_ptr);
/* Linear, fft-shifted, un-normalised power spectrum = the ([F)V payload
 * (src/dsp/fft_process.cpp:83-97).  Device layout [n_streams][n_blocks][N]. */
int  sdrb_power_device(sdrb_ctx* ctx, const float** dev_ptr);
int  sdrb_get_power(sdrb_ctx* ctx, int stream, int block, float* host_out);
/* All spectra of the last call, [n_streams][n_blocks][N] floats, into host memory (pinned or pageable). */
int  sdrb_get_power_all(sdrb_ctx* ctx, float* host_out, size_t capacity_floats);
/* dB view of the spectrum (extra output; the callback payload stays the linear spectrum above):
 * 10*log10f(p/refPower + 1e-20f) with refPower = 1, the expression evaluateSignalStrength applies to the bins it
 * looks at (src/dsp/fft_process.cpp:142,149; src/dsp/fft_process.h:74-75), for every bin.  Within 1e-3 dB of
 * 10*log10f of the linear output. */
enum sdrb_power_format {
    SDRB_POWER_LINEAR_F32 = 0,   /* float32 linear power: the reference's ([F)V payload, the default everywhere */
    SDRB_POWER_DB_F32 = 1,       /* float32 dB */
    SDRB_POWER_DB_F16 = 2        /* IEEE half dB (0.0625 dB steps at 64..128 dB): half the device-to-host bytes */
};
int  sdrb_get_power_db(sdrb_ctx* ctx, int stream, int block, float* host_out);
/* sdrb_process_host with the spectra delivered in `power_format` (power_out then holds float32 or half values,
 * [n_streams][n_blocks][N]).  The host side of the call is bound by the device-to-host path (4 B per sample of
 * linear spectra against 2 B of u8 IQ in: profiles/r02_host_path.json); the half-precision dB view is the opt-in
 * payload that halves it for consumers that only draw the spectrum. */
int  sdrb_process_host_ex(sdrb_ctx* ctx, const void* iq_host, int n_blocks, unsigned stages,
                          sdrb_frame_result* results_out, void* power_out, int power_format, int16_t* pcm_out);
/* int16 PCM = the ([S)V payload; layout [n_streams][n_blocks][pcm_per_block]
 * (x ssbChannels when the NCO bank is on: [stream][channel][block][m]). */
int  sdrb_pcm_device(sdrb_ctx* ctx, const int16_t** dev_ptr);
int  sdrb_get_pcm(sdrb_ctx* ctx, int16_t* host_out, size_t capacity_samples);

/* ---- pulse detectors -------------------------------------------------- */
/* Defaults of AudioPulseDetector::Config. */
void sdrb_audio_pulse_config_default(sdrb_audio_pulse_config* cfg);
/* SSBProcessor::setPulseConfig (src/ssb/ssb_processor.cpp:69-74,90-96): every stream's audio detector
 * is rebuilt from cfg (state reset) before the next reading. */
int  sdrb_set_pulse_config(sdrb_ctx* ctx, const sdrb_audio_pulse_config* cfg);
/* Records of the last sdrb_process that carried a *_PULSE stage.  Spectral fields are indexed
 * [stream][reading]; with an NCO bank the audio fields are [stream][channel][reading] and the
 * spectral fields are valid in channel 0's records. */
int  sdrb_get_pulse_results(sdrb_ctx* ctx, sdrb_pulse_result* host_out, size_t capacity_records);
int  sdrb_pulse_results_device(sdrb_ctx* ctx, const sdrb_pulse_result** dev_ptr);

/* Diagnostic of the SDRB_RU16 front end (libairspy iqconverter_float.c:418-431, remove_dc): the DC tracker is
 * walked in chunks from guessed starts and every chunk is checked against its predecessor's end value; a
 * chunk whose guess did not land on the sequential value is walked again in order.  Sum over streams since
 * sdrb_configure / sdrb_reset_state of the chunks that needed the second walk (0 on ordinary noisy ADC data;
 * the output is bit-exact either way). */
int  sdrb_airspy_redone_chunks(sdrb_ctx* ctx, uint64_t* out);
/* The converter's output of the last SDRB_RU16 call: [n_streams][n_blocks][N] interleaved (I, Q) floats, what
 * iqconverter_float_process leaves in the sample buffer (iqconverter_float.c:471-475) and SoapyAirspy hands
 * to readStream as CF32.  Synchronises. */
int  sdrb_get_airspy_iq(sdrb_ctx* ctx, float* host_out, size_t capacity_floats);

/* ---- allocation-free estimator ------------------------------------------ */
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The frame-local part of
 * evaluateSignalStrength -- what sdrb_process(SDRB_STAGE_FFT) computes per reading before the per-stream tracker --
 * over n_frames caller-owned DEVICE spectra d_power[n_frames][samples_per_reading] (linear, fft-shifted, e.g. from
 * sdrb_power_device), into caller-owned DEVICE records d_out[n_frames], on `cuda_stream` (NULL = the default stream)
 * of `device`.  Needs no context and calls no cudaMalloc / cudaFree: the scratch the kernels need lives in a static
 * device array of about 0.9 MB that every GPU loading the library holds until exit.
 * Limits: n_frames <= SDRB_NOALLOC_MAX_FRAMES, and a focus range whose pooled noise bins fit the estimator's
 * shared-memory pool (8192 bins per frame; the defaults pool 328) -- beyond either the call fails with
 * SDRB_ERR_ARG / SDRB_ERR_UNSUPPORTED and a message in sdrb_last_error(NULL).
 * Every call shares the array: calls that overlap on the GPU (two streams, two threads) race; serialise them.
 * Asynchronous: d_out is valid once the stream has reached the end of the call's work. */
#define SDRB_NOALLOC_MAX_FRAMES 4096
typedef struct sdrb_frame_estimate {
    float absPeakDb, signalPowerDb;
    float meanSnrDb, meanSnrSigma;
    float perBinMean, peakAboveNoiseMeanDb;
    float maxBinSnrDb, maxBinSnrSigma;
    float best1kHzSnrDb, best1kHzSnrSigma;
    int32_t best1kStart;   /* first bin of the best 1 kHz window, -1 when there is none */
    int32_t peakBin;       /* arg-max bin of the focus band (spectrum index), -1 when the focus band is empty */
    int32_t valid;         /* at least two reference windows: the SNR fields are meaningful */
    int32_t reserved;
} sdrb_frame_estimate;
int  sdrb_estimate_noalloc(int device, uint32_t sample_rate, int samples_per_reading, int focus_khz,
                           const float* d_power, int n_frames, sdrb_frame_estimate* d_out, void* cuda_stream);

/* ---- per-stream state (checkpoint / resume) ----------------------------- */
/* Opaque blob: three biquad states, tracking memory, detection ring, frame counter.  The pulse detectors'
 * histories and the Airspy converter state are NOT part of the blob: they restart with sdrb_configure /
 * sdrb_reset_state (the reference's detectors restart with setPulseConfig / reset() in the same way). */
size_t sdrb_state_bytes(const sdrb_ctx* ctx);
int  sdrb_save_state(sdrb_ctx* ctx, void* host_blob, size_t bytes);
int  sdrb_load_state(sdrb_ctx* ctx, const void* host_blob, size_t bytes);
int  sdrb_reset_state(sdrb_ctx* ctx);

/* ---- deterministic synthetic IQ source (replaces Soapy/USB capture on the box) ---- */
/* Integer-only generator, bit-identical on host and device.  Stream s (global id) uses
 * seed 0x5D120000+s: keyed CW tone inside the default focus band, a static tone in a
 * reference window, approximately Gaussian noise.  Fills [n_streams][n_samples]. */
int  sdrb_synth_host(int format, uint32_t sample_rate, int first_stream, int n_streams,
                     int64_t first_sample, int64_t n_samples, void* host_out);
int  sdrb_synth_device(sdrb_ctx* ctx, int format, uint32_t sample_rate, int first_stream, int n_streams,
                       int64_t first_sample, int64_t n_samples, void* dev_out);
size_t sdrb_format_bytes(int format); /* bytes per complex sample */

/* ---- introspection -------------------------------------------------------- */
/* Kernels launched by this context since creation (bench.py's gpu_launches). */
uint64_t sdrb_launch_count(const sdrb_ctx* ctx);
/* Device time in ms of the dominant (FFT) kernel launches of the last sdrb_process, measured
 * with CUDA events on the launching stream; returns count of launches timed. */
int  sdrb_last_fft_kernel_ms(sdrb_ctx* ctx, float* ms_total);
/* Readings covered by the launch(es) timed above (the 16K path times its first chunk only). */
int  sdrb_last_fft_kernel_frames(const sdrb_ctx* ctx);
int  sdrb_enable_kernel_timing(sdrb_ctx* ctx, int on);

#ifdef __cplusplus
}
#endif
#endif /* SDR_B200_H */
