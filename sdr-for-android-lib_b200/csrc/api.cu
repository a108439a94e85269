// api.cu — the C-ABI of include/sdr_b200.h over the CUDA kernels.  No CPU fallback: every
// compute entry point needs a CUDA device and fails with SDRB_ERR_CUDA otherwise.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "fft_tables.h"
#include "host_design.h"
#include "internal.h"
#include "pulse.h"
#include "nco.cuh"
#include "synth.cuh"

using namespace sdrb;

namespace sdrb { thread_local TraceSink* g_trace = nullptr; }

namespace {
thread_local std::string g_create_error;

// SDRB_TRACE=1: events after every launch group, printed (ms since the call's first mark) by sdrb_sync
struct EventTrace : sdrb::TraceSink {
    std::vector<std::pair<std::string, cudaEvent_t>> marks;
    std::vector<cudaEvent_t> pool;
    void mark(const char* name, cudaStream_t st) override {
        cudaEvent_t e;
        if (!pool.empty()) { e = pool.back(); pool.pop_back(); }
        else if (cudaEventCreate(&e) != cudaSuccess) return;
        cudaEventRecord(e, st);
        marks.emplace_back(name, e);
    }
    void dump() {
        if (marks.empty()) return;
        cudaEvent_t t0 = marks[0].second;
        for (auto& m : marks) {
            float ms = 0.f;
            if (cudaEventSynchronize(m.second) == cudaSuccess && cudaEventElapsedTime(&ms, t0, m.second) == cudaSuccess)
                fprintf(stderr, "[sdrb trace] %-22s %8.3f ms\n", m.first.c_str(), ms);
        }
        for (auto& m : marks) pool.push_back(m.second);
        marks.clear();
    }
    ~EventTrace() override { for (auto& m : marks) cudaEventDestroy(m.second); for (auto e : pool) cudaEventDestroy(e); }
};

constexpr size_t kGenericScratchBytes = 256u << 20;  // per ping-pong buffer of the generic FFT
constexpr size_t kTwoPassScratchBytes = 1024u << 20; // intermediate of the two-pass scheme (32768 .. 262144 points)
// Four-step intermediate per launch pair.  Round 1 kept it within 96 MB "so that it stays in L2"; ncu showed it did not
// (kernel B read 97 MB per group from DRAM, 16 % L2 hits) and a sweep on the B200 is monotonic in the other direction:
// 24 MB 0.99 ms, 96 MB 0.665 ms, 384 MB 0.603 ms, whole batch 0.600 ms (256 x 8 x 65536 cs16); 1M points 0.608 -> 0.491 ms.
// Fewer, longer launches win (tails and launch gaps), so the cap is only there to bound the scratch.  SDRB_FB_L2_MB overrides.
constexpr size_t kFourStepL2Bytes = 1024u << 20;

struct StateBlobHeader {
    uint32_t magic, version;
    int32_t n_streams, rfMode;
    uint64_t ssb_samples;     // NCO phase origin of the extended SSB chain
    uint32_t rfFs, pad;
};
constexpr int kMaxChunks = 32;
constexpr uint32_t kBlobMagic = 0x53445242u;  // "SDRB"
}  // namespace

struct sdrb_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    std::string err;
    bool configured = false;
    sdrb_config cfg = {};
    int n_streams = 0, max_blocks = 0;
    int rfMode = -1;            // latched by the first SSB call (static rfFilter, ssb_demod_opt.cpp:261-263)
    uint32_t rfFs = 0;          // sample rate latched with it (iir2InitLowpass runs once, with the first call's rate)
    int effMode = 1;            // last sound mode in 0..2: any other value leaves the reference's file statics as the
                                // previous call set them (ssb_demod_opt.cpp:17-26,230-255); they start as mode 1
    int fftPath = 0;
    EstGeometry geo = {};
    SsbParams ssb = {};
    float taps255[256] = {};
    int smax = 1;
    // extension stages (identity = reference behaviour)
    int channels = 1;           // SSB channels per stream (NCO bank)
    bool ssb_ext = false;       // any of: NCO offset, channel bank, true sideband
    uint32_t inc2 = 0;          // Weaver second-mix increment
    unsigned long long ssb_samples = 0;  // samples per stream since reset (NCO phase origin)
    int front = 0;              // bit 0 window, bit 1 block-mean removal
    // device buffers
    EstGeometry* d_geo = nullptr;
    void* d_in = nullptr;
    float* d_power = nullptr;
    void* d_view = nullptr;     // dB view of the spectra (float or half), allocated with the first request for one
    FrameLocal* d_local = nullptr;
    sdrb_frame_result* d_results = nullptr;
    FftStreamState* d_fft_state = nullptr;
    SsbStreamState* d_ssb_state = nullptr;
    void* d_start = nullptr;    // float2 per frame (reference-exact path) or float4 per (frame, channel)
    float* d_a48 = nullptr;
    int16_t* d_pcm = nullptr;
    float* d_taps = nullptr;
    float* d_taps_lin = nullptr;   // the 255 taps in linear order (any-length pass)
    bool ssb_any = false;          // block bytes not a multiple of 64: the scalar any-length pass
    float* d_tapsE = nullptr;   // rotated tap table of the fused AGC + decimator pass (ssb.cu:k_ssb_main)
    int ssb_sp = 0;             // its accumulator count; 0 = the decimation factor is outside its range (split pair)
    cpx* d_tables16k = nullptr;   // 16K kernel tables (TP | T2)
    int* d_ticket = nullptr;      // transform counter of the 16K kernel when it shares the SMs with the SSB chain
    cpx* d_tw = nullptr;
    cpx *d_chirp = nullptr, *d_bf = nullptr;   // Bluestein: chirp c[k], transform of the wrapped conjugate chirp
    int blue_m = 0;
    cpx *d_sa = nullptr, *d_sb = nullptr;
    cpx* d_fb_tables = nullptr;   // four-step twiddle tables
    cpx* d_fs_tables = nullptr;   // small one-CTA kernel's twiddle tables (fft_small.cu)
    cpx* d_fbA = nullptr;         // four-step intermediate A' (fb_group frames; fused variant: a ring of fb_slots frames)
    int fb_group = 0;
    unsigned* d_fb_ctl = nullptr; // fused four-step: ticket + per-frame counters
    int fb_slots = 0;             // > 0: the fused one-kernel variant is in use
    float* d_est_scratch = nullptr;
    int16_t* d_sine = nullptr;
    uint32_t* d_inc1 = nullptr;
    float* d_win = nullptr;
    cpx* d_front = nullptr;
    // pulse detectors
    sdrb_audio_pulse_config au_user = {};
    bool au_user_set = false;
    PulseCfg sp_cfg = {}, au_cfg = {};
    PulseState *d_sp_state = nullptr, *d_au_state = nullptr;
    float *d_sp_e = nullptr, *d_sp_f = nullptr, *d_au_e = nullptr;
    sdrb_pulse_result* d_pulse = nullptr;
    // Airspy real->IQ front end (inputFormat SDRB_RU16)
    AirspyStreamState* d_air_state = nullptr;
    cpx* d_air_iq = nullptr;
    float* d_ext_work = nullptr;   // extended chain: removeDC output X and mixed samples M (ssb_ext.cu)
    float* d_agc = nullptr;        // post-AGC samples [streams][blocks][N] between k_ssb_agc and k_ssb_fir
    float* d_air_work = nullptr;   // per-chunk start / end values of the DC pass
    size_t air_work_per_stream = 0;
    int air_chunk = 0, air_warm = 0;
    // last call
    int last_blocks = 0;
    unsigned last_stages = 0;
    uint64_t launches = 0;
    bool timing = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool ev_valid = false;
    int last_fft_launches = 0;
    int last_fft_frames = 0;    // frames covered by the timed FFT launch(es)
    int est_chunk = 0;          // frames per FFT->estimator chunk on the 16K path (0 = whole call)
    // SDRB_EST_OVERLAP=1: estimator / tracker of chunk i on a second stream UNDER the FFT kernel of chunk i + 1
    // (the FFT CTA leaves a quarter of the register file and 42 KiB of shared memory per SM free).  Measured on
    // the B200: 0.509 ms per step vs 0.411 ms (four 2048-frame FFT launches cost 0.36 ms against 0.315 ms for one,
    // and the co-resident estimator CTAs slow them further) -> off by default.
    bool est_overlap = false;
    cudaStream_t s_aux = nullptr;
    cudaEvent_t ev_chunk[4] = {}, ev_aux = nullptr;
    // FFT + SSB in one call: the SSB chain runs on its own stream beside the FFT chain (both read the same raw IQ,
    // src/sdr-bridge-java-soapy.cpp:441-444 hands one buffer to both processors).  The per-stream state pass is a
    // latency-bound walk of one warp per 32 streams; under the FFT kernel it costs issue slots only.  SDRB_SSB_FORK=0
    // serialises the two chains on the caller's stream.
    EventTrace* trace = nullptr;
    bool ssb_fork = true;
    bool ssb_first = false;     // SDRB_SSB_FIRST=1: launch the SSB chain before the FFT chain (measured slower)
    cudaStream_t s_ssb = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    // host-to-host pipeline (sdrb_process_host)
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t ev_in[32] = {}, ev_done[32] = {};
    int host_chunks = 8;        // SDRB_HOST_CHUNKS: pipeline depth of sdrb_process_host
};

namespace sdrb { void set_global_error(const char* text) { g_create_error = text ? text : ""; } }

static int fail(sdrb_ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf; else g_create_error = buf;
    return code;
}
#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(ctx, e_ == cudaErrorMemoryAllocation ? SDRB_ERR_NOMEM : SDRB_ERR_CUDA,     \
                        "%s: %s", #call, cudaGetErrorString(e_));                                  \
    } while (0)

static void free_buffers(sdrb_ctx* c) {
    void** ptrs[] = {(void**)&c->d_geo, &c->d_in, (void**)&c->d_power, &c->d_view, (void**)&c->d_local, (void**)&c->d_results,
                     (void**)&c->d_fft_state, (void**)&c->d_ssb_state, &c->d_start,
                     (void**)&c->d_a48, (void**)&c->d_pcm, (void**)&c->d_taps, (void**)&c->d_taps_lin, (void**)&c->d_tapsE, (void**)&c->d_tw, (void**)&c->d_chirp, (void**)&c->d_bf, (void**)&c->d_sa,
                     (void**)&c->d_sb, (void**)&c->d_est_scratch, (void**)&c->d_inc1, (void**)&c->d_win,
                     (void**)&c->d_front, (void**)&c->d_fb_tables, (void**)&c->d_fs_tables, (void**)&c->d_fbA, (void**)&c->d_fb_ctl,
                     (void**)&c->d_sp_state, (void**)&c->d_au_state, (void**)&c->d_sp_e, (void**)&c->d_sp_f,
                     (void**)&c->d_au_e, (void**)&c->d_pulse, (void**)&c->d_air_state,
                     (void**)&c->d_air_iq, (void**)&c->d_air_work, (void**)&c->d_agc, (void**)&c->d_ext_work};
    for (void** p : ptrs) {
        if (*p) cudaFree(*p);
        *p = nullptr;
    }
}

extern "C" {

int sdrb_abi_version(void) { return SDRB_ABI_VERSION; }

size_t sdrb_format_bytes(int format) {
    return format == SDRB_CU8 ? 2 : (format == SDRB_CS16 || format == SDRB_RU16 ? 4 : (format == SDRB_CF32 ? 8 : 0));
}

const char* sdrb_last_error(const sdrb_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int sdrb_create(int device, sdrb_ctx** out) {
    sdrb_ctx* ctx = nullptr;
    if (!out) return fail(nullptr, SDRB_ERR_ARG, "sdrb_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0)
        return fail(nullptr, SDRB_ERR_CUDA, "no CUDA device (%s); this library has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "count = 0");
    if (device < 0 || device >= count) return fail(nullptr, SDRB_ERR_ARG, "device %d out of range [0,%d)", device, count);
    CK(cudaSetDevice(device));
    ctx = new sdrb_ctx();
    ctx->device = device;
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    if ((e = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreate(&ctx->ev0)) != cudaSuccess || (e = cudaEventCreate(&ctx->ev1)) != cudaSuccess ||
        (e = fft_big_prepare()) != cudaSuccess || (e = fft_small_prepare()) != cudaSuccess ||
        (e = estimator_prepare()) != cudaSuccess ||
        (e = ssb_prepare()) != cudaSuccess) {
        fail(nullptr, SDRB_ERR_CUDA, "context setup: %s", cudaGetErrorString(e));
        delete ctx;
        return SDRB_ERR_CUDA;
    }
    ctx->stream = ctx->own_stream;
    // constant tables: 16K twiddles and the synthetic source's sine table
    const char* chunk_env = getenv("SDRB_EST_CHUNK");
    ctx->est_chunk = chunk_env ? atoi(chunk_env) : 0;   // measured: chunking for L2 residency loses (0.67 vs 0.45 ms per step)
    const char* hc_env = getenv("SDRB_HOST_CHUNKS");
    if (hc_env) ctx->host_chunks = atoi(hc_env);
    const char* ov_env = getenv("SDRB_EST_OVERLAP");
    ctx->est_overlap = ov_env && ov_env[0] == '1';
    const char* tr_env = getenv("SDRB_TRACE");
    if (tr_env && tr_env[0] == '1') ctx->trace = new EventTrace();
    const char* fk_env = getenv("SDRB_SSB_FORK");
    ctx->ssb_fork = !(fk_env && fk_env[0] == '0');
    const char* sf_env = getenv("SDRB_SSB_FIRST");
    ctx->ssb_first = sf_env && sf_env[0] == '1';
    const std::vector<cpx> t16 = make_f2_tables();
    if ((e = cudaMalloc(&ctx->d_tables16k, t16.size() * sizeof(cpx))) != cudaSuccess ||
        (e = cudaMemcpy(ctx->d_tables16k, t16.data(), t16.size() * sizeof(cpx), cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = fft16k2_prepare(t16.data())) != cudaSuccess ||
        (e = cudaMalloc(&ctx->d_ticket, sizeof(int))) != cudaSuccess ||
        (e = cudaMalloc(&ctx->d_sine, SYNTH_SINE_N * sizeof(int16_t))) != cudaSuccess ||
        (e = cudaMemcpy(ctx->d_sine, synth_sine_table(), SYNTH_SINE_N * sizeof(int16_t), cudaMemcpyHostToDevice)) != cudaSuccess) {
        fail(nullptr, SDRB_ERR_CUDA, "table upload: %s", cudaGetErrorString(e));
        sdrb_destroy(ctx);
        return SDRB_ERR_CUDA;
    }
    *out = ctx;
    return SDRB_OK;
}

int sdrb_destroy(sdrb_ctx* ctx) {
    if (!ctx) return SDRB_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    free_buffers(ctx);
    if (ctx->d_tables16k) cudaFree(ctx->d_tables16k);
    if (ctx->d_ticket) cudaFree(ctx->d_ticket);
    if (ctx->d_sine) cudaFree(ctx->d_sine);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->s_aux) cudaStreamDestroy(ctx->s_aux);
    if (ctx->s_ssb) cudaStreamDestroy(ctx->s_ssb);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    for (auto& e : ctx->ev_chunk) if (e) cudaEventDestroy(e);
    if (ctx->ev_aux) cudaEventDestroy(ctx->ev_aux);
    if (ctx->s_in) cudaStreamDestroy(ctx->s_in);
    if (ctx->s_out) cudaStreamDestroy(ctx->s_out);
    for (int i = 0; i < kMaxChunks; i++) {
        if (ctx->ev_in[i]) cudaEventDestroy(ctx->ev_in[i]);
        if (ctx->ev_done[i]) cudaEventDestroy(ctx->ev_done[i]);
    }
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx->trace;
    delete ctx;
    return SDRB_OK;
}

int sdrb_set_cuda_stream(sdrb_ctx* ctx, void* cuda_stream) {
    if (!ctx) return SDRB_ERR_ARG;
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
    return SDRB_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ configuration
static bool is_pow2(int n) { return n > 0 && (n & (n - 1)) == 0; }

static int check_config(sdrb_ctx* ctx, const sdrb_config* cfg) {
    if (!cfg) return fail(ctx, SDRB_ERR_ARG, "config is NULL");
    if (cfg->samplesPerReading < 1) return fail(ctx, SDRB_ERR_ARG, "samplesPerReading must be >= 1");
    if (cfg->sampleRate == 0) return fail(ctx, SDRB_ERR_ARG, "sampleRate must be > 0");
    if (cfg->inputFormat < SDRB_CU8 || cfg->inputFormat > SDRB_RU16)
        return fail(ctx, SDRB_ERR_ARG, "inputFormat %d is not a sdrb_format", cfg->inputFormat);
    if (cfg->fftWindow < 0 || cfg->fftWindow > 1) return fail(ctx, SDRB_ERR_ARG, "fftWindow must be 0 (rectangular) or 1 (Hann)");
    if (cfg->ssbChannels < 0 || cfg->ssbChannels > 4096) return fail(ctx, SDRB_ERR_ARG, "ssbChannels must be in [0, 4096]");
    if (!(cfg->ssbChannelOffsetHz == cfg->ssbChannelOffsetHz) || !(cfg->ssbChannelSpacingHz == cfg->ssbChannelSpacingHz))
        return fail(ctx, SDRB_ERR_ARG, "ssbChannelOffsetHz / ssbChannelSpacingHz must be numbers");
    return SDRB_OK;
}

// 1: one-CTA 16K kernel, 4: one-CTA kernel for 256 .. 8192 points, 2: multi-CTA four-step (64K, 1M), 0: multi-pass power
// of two, 3: any other length (Bluestein)
static int fft_path_for(int n) {
    static const bool no_small = getenv("SDRB_FFT_SMALL") && getenv("SDRB_FFT_SMALL")[0] == '0';   // test hook: Stockham passes instead
    if (n == F2_N) return 1;
    if (fft_small_supports(n) && !no_small) return 4;
    if (fft_twopass_supports(n) && !no_small) return 5;   // radix-A columns + the one-CTA kernel on 4096 / 8192-point rows
    return fft_big_supports(n) ? 2 : (is_pow2(n) && n >= 2 ? 0 : 3);
}
constexpr int kMaxBluesteinN = 1 << 22;   // convolution length 2^23: 64 MiB per transform and scratch buffer

extern "C" int sdrb_plan_config(const sdrb_config* cfg, sdrb_plan* out) {
    if (!out) return fail(nullptr, SDRB_ERR_ARG, "plan out is NULL");
    int rc = check_config(nullptr, cfg);
    if (rc != SDRB_OK) return rc;
    memset(out, 0, sizeof(*out));
    EstGeometry g;
    design_geometry(cfg->sampleRate, cfg->samplesPerReading, cfg->freqFocusRangeKhz, g);
    out->n = g.n; out->focusLo = g.focusLo; out->focusHi = g.focusHi; out->focusLen = g.focusLen;
    out->winBins1k = g.winBins1k; out->nWin = g.nWin; out->nBottom = g.nBottom;
    for (int i = 0; i < 10; i++) { out->winLo[i] = g.winLo[i]; out->winHi[i] = g.winHi[i]; }
    out->freqPerBin = g.freqPerBin; out->nyquist = g.nyquist;
    SsbParams P;
    float taps[256] = {};
    design_ssb(cfg->sampleRate, cfg->samplesPerReading, cfg->soundMode, cfg->sampleRate, cfg->soundMode,
               cfg->upperSideband, cfg->ssbTrueSideband != 0, P, taps);
    out->decim = P.decim; out->nTaps = P.nTaps; out->pcmPerBlock = P.m;
    out->fftPath = fft_path_for(cfg->samplesPerReading);
    const BiquadCoef* cs[3] = {&P.rf, &P.hp, &P.bp};
    float* dst[3] = {out->rf, out->hp, out->bp};
    for (int i = 0; i < 3; i++) {
        dst[i][0] = cs[i]->a0; dst[i][1] = cs[i]->a1; dst[i][2] = cs[i]->a2; dst[i][3] = cs[i]->b1; dst[i][4] = cs[i]->b2;
    }
    out->agcTarget = P.agcTarget; out->agcFast = P.agcFast; out->agcSlow = P.agcSlow;
    out->gain = P.gain; out->transientCoeff = P.transientCoeff; out->dcAlpha = P.dcAlpha;
    for (int i = 0; i < P.nTaps && i < 255; i++) out->firTaps[i] = taps[i];
    return SDRB_OK;
}

static int upload_geometry(sdrb_ctx* ctx) {
    design_geometry(ctx->cfg.sampleRate, ctx->cfg.samplesPerReading, ctx->cfg.freqFocusRangeKhz, ctx->geo);
    CK(cudaMemcpyAsync(ctx->d_geo, &ctx->geo, sizeof(EstGeometry), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));  // ctx->geo is pageable host memory that may change again
    const size_t need = estimator_scratch_floats(ctx->geo, ctx->n_streams * ctx->max_blocks);
    if (ctx->d_est_scratch) { cudaFree(ctx->d_est_scratch); ctx->d_est_scratch = nullptr; }
    if (need) CK(cudaMalloc(&ctx->d_est_scratch, need * sizeof(float)));
    return SDRB_OK;
}

static int upload_ssb(sdrb_ctx* ctx) {
    if (ctx->cfg.soundMode >= 0 && ctx->cfg.soundMode <= 2) ctx->effMode = ctx->cfg.soundMode;
    const int rf = ctx->rfMode >= 0 ? ctx->rfMode : ctx->effMode;
    const bool tsb = ctx->cfg.ssbTrueSideband != 0;
    design_ssb(ctx->cfg.sampleRate, ctx->cfg.samplesPerReading, rf, ctx->rfFs, ctx->effMode, ctx->cfg.upperSideband,
               tsb, ctx->ssb, ctx->taps255);
    // NCO increments of the channel bank (extension; oracle/sdr_oracle_ext.c): channel c is tuned to
    // offset + c * spacing; with true sideband the first mix goes to the sideband centre (+/- B/2)
    // and the second mix brings it back, B = the low-pass bandwidth latched with the first block.
    const double halfBand = tsb ? (double)ssb_mode(rf).lowpassBd * 0.5 : 0.0;
    const double sgn = ctx->cfg.upperSideband ? 1.0 : -1.0;
    std::vector<uint32_t> inc(ctx->channels);
    for (int c = 0; c < ctx->channels; c++)
        inc[c] = nco_inc((double)ctx->cfg.ssbChannelOffsetHz + (double)c * (double)ctx->cfg.ssbChannelSpacingHz +
                             sgn * halfBand, ctx->cfg.sampleRate);
    ctx->inc2 = tsb ? nco_inc(halfBand, ctx->cfg.sampleRate) : 0u;
    if (ctx->d_inc1) { cudaFree(ctx->d_inc1); ctx->d_inc1 = nullptr; }
    CK(cudaMalloc(&ctx->d_inc1, inc.size() * sizeof(uint32_t)));
    CK(cudaMemcpy(ctx->d_inc1, inc.data(), inc.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    ctx->smax = ssb_smax(ctx->ssb);
    const std::vector<float> t = slot_major_taps(ctx->taps255, ctx->ssb.nTaps, ctx->ssb.decim, ctx->smax);
    if (ctx->d_taps) { cudaFree(ctx->d_taps); ctx->d_taps = nullptr; }
    CK(cudaMalloc(&ctx->d_taps, t.size() * sizeof(float)));
    CK(cudaMemcpy(ctx->d_taps, t.data(), t.size() * sizeof(float), cudaMemcpyHostToDevice));
    if (ctx->d_taps_lin) { cudaFree(ctx->d_taps_lin); ctx->d_taps_lin = nullptr; }
    CK(cudaMalloc(&ctx->d_taps_lin, 256 * sizeof(float)));
    CK(cudaMemcpy(ctx->d_taps_lin, ctx->taps255, 256 * sizeof(float), cudaMemcpyHostToDevice));
    ctx->ssb_any = ((size_t)ctx->cfg.samplesPerReading * sdrb_format_bytes(ctx->cfg.inputFormat == SDRB_RU16 ? (int)SDRB_CF32 : ctx->cfg.inputFormat)) % 64 != 0;
    // fused AGC + decimator pass (both chains): accumulator count and rotated tap table
    if (ctx->d_tapsE) { cudaFree(ctx->d_tapsE); ctx->d_tapsE = nullptr; }
    const int kfmt = ctx->cfg.inputFormat == SDRB_RU16 ? (int)FMT_CF32 : ctx->cfg.inputFormat;
    const char* fe = getenv("SDRB_SSB_FUSED");
    ctx->ssb_sp = (ctx->ssb_any || (fe && fe[0] == '0')) ? 0
                  : (ctx->ssb_ext ? ssb_ext_fused_slots(ctx->ssb) : ssb_fused_slots(ctx->ssb, kfmt));
    if (ctx->ssb_sp > 0) {
        const std::vector<float> te = ssb_fused_tap_table(ctx->taps255, ctx->ssb, ctx->ssb_sp);
        CK(cudaMalloc(&ctx->d_tapsE, te.size() * sizeof(float)));
        CK(cudaMemcpy(ctx->d_tapsE, te.data(), te.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    return SDRB_OK;
}

// (Re)build the detectors' configuration and, once their buffers exist, reset their state
// (SpectralPulseDetector::reset / AudioPulseDetector construction, ssb_processor.cpp:90-96).
static int reset_pulse(sdrb_ctx* ctx, bool spectral, bool audio) {
    if (spectral) pulse_cfg_spectral(&ctx->sp_cfg, ctx->cfg.sampleRate, ctx->cfg.samplesPerReading);
    if (audio) {
        if (!ctx->au_user_set) sdrb_audio_pulse_config_default(&ctx->au_user);
        pulse_cfg_audio(&ctx->au_cfg, &ctx->au_user);
    }
    if (spectral && ctx->d_sp_state) {
        PulseState init;
        pulse_state_init(&init, ctx->sp_cfg);
        std::vector<PulseState> v((size_t)ctx->n_streams, init);
        CK(cudaMemcpy(ctx->d_sp_state, v.data(), v.size() * sizeof(PulseState), cudaMemcpyHostToDevice));
    }
    if (audio && ctx->d_au_state) {
        // a new audio configuration may change the history length: the ring is sized for it
        if (ctx->d_au_e) { cudaFree(ctx->d_au_e); ctx->d_au_e = nullptr; }
        const size_t units = (size_t)ctx->n_streams * ctx->channels;
        CK(cudaMalloc(&ctx->d_au_e, units * ctx->au_cfg.ringCap * sizeof(float)));
        PulseState init;
        pulse_state_init(&init, ctx->au_cfg);
        std::vector<PulseState> v(units, init);
        CK(cudaMemcpy(ctx->d_au_state, v.data(), v.size() * sizeof(PulseState), cudaMemcpyHostToDevice));
    }
    return SDRB_OK;
}

static int ensure_pulse_buffers(sdrb_ctx* ctx, unsigned stages) {
    const size_t S = (size_t)ctx->n_streams, units = S * ctx->channels;
    if (!ctx->d_pulse) {
        CK(cudaMalloc(&ctx->d_pulse, units * ctx->max_blocks * sizeof(sdrb_pulse_result)));
        CK(cudaMemset(ctx->d_pulse, 0, units * ctx->max_blocks * sizeof(sdrb_pulse_result)));
    }
    if ((stages & SDRB_STAGE_SPECTRAL_PULSE) && !ctx->d_sp_state) {
        CK(cudaMalloc(&ctx->d_sp_state, S * sizeof(PulseState)));
        CK(cudaMalloc(&ctx->d_sp_e, S * ctx->sp_cfg.ringCap * sizeof(float)));
        CK(cudaMalloc(&ctx->d_sp_f, S * ctx->sp_cfg.ringCap * sizeof(float)));
        int rc = reset_pulse(ctx, true, false);
        if (rc != SDRB_OK) return rc;
    }
    if ((stages & SDRB_STAGE_AUDIO_PULSE) && !ctx->d_au_state) {
        CK(cudaMalloc(&ctx->d_au_state, units * sizeof(PulseState)));
        int rc = reset_pulse(ctx, false, true);
        if (rc != SDRB_OK) return rc;
    }
    return SDRB_OK;
}

static int reset_state(sdrb_ctx* ctx) {
    std::vector<FftStreamState> fs(ctx->n_streams);
    for (auto& s : fs) {
        memset(&s, 0, sizeof(s));
        s.centerFrequency = ctx->cfg.centerFrequency;
        s.maxPeakDb = -130.0f;                                  // src/dsp/fft_process.cpp:33-35
        s.maxPeakHz = static_cast<float>(ctx->cfg.centerFrequency);
    }
    CK(cudaMemcpy(ctx->d_fft_state, fs.data(), fs.size() * sizeof(FftStreamState), cudaMemcpyHostToDevice));
    CK(cudaMemset(ctx->d_ssb_state, 0, (size_t)ctx->n_streams * ctx->channels * sizeof(SsbStreamState)));
    ctx->rfMode = -1;
    ctx->rfFs = 0;
    ctx->ssb_samples = 0;
    ctx->last_blocks = 0;
    ctx->last_stages = 0;
    if (ctx->d_air_state) CK(cudaMemset(ctx->d_air_state, 0, (size_t)ctx->n_streams * sizeof(AirspyStreamState)));
    return reset_pulse(ctx, true, true);
}

extern "C" {

int sdrb_configure(sdrb_ctx* ctx, const sdrb_config* cfg, int n_streams, int max_blocks) {
    if (!ctx) return SDRB_ERR_ARG;
    int rc = check_config(ctx, cfg);
    if (rc != SDRB_OK) return rc;
    if (n_streams < 1 || max_blocks < 1) return fail(ctx, SDRB_ERR_ARG, "n_streams and max_blocks must be >= 1");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    free_buffers(ctx);
    ctx->configured = false;
    ctx->cfg = *cfg;
    ctx->n_streams = n_streams;
    ctx->max_blocks = max_blocks;
    const int n = cfg->samplesPerReading;
    const size_t frames = (size_t)n_streams * max_blocks;
    ctx->fftPath = fft_path_for(n);
    ctx->channels = cfg->ssbChannels > 1 ? cfg->ssbChannels : 1;
    ctx->ssb_ext = ctx->channels > 1 || cfg->ssbTrueSideband != 0 || cfg->ssbChannelOffsetHz != 0.0f;
    ctx->front = (cfg->fftWindow ? 1 : 0) | (cfg->fftRemoveDc ? 2 : 0);
    if (ctx->front & 1) {   // symmetric Hann in double, rounded once (oracle/sdr_oracle_ext.c:sdr_oracle_window)
        std::vector<float> w(n);
        for (int i = 0; i < n; i++)
            w[i] = n > 1 ? (float)(0.5 - 0.5 * std::cos(2.0 * 3.14159265358979323846 * (double)i / (double)(n - 1))) : 1.0f;
        CK(cudaMalloc(&ctx->d_win, (size_t)n * sizeof(float)));
        CK(cudaMemcpy(ctx->d_win, w.data(), (size_t)n * sizeof(float), cudaMemcpyHostToDevice));
    }
    CK(cudaMalloc(&ctx->d_geo, sizeof(EstGeometry)));
    CK(cudaMalloc(&ctx->d_fft_state, n_streams * sizeof(FftStreamState)));
    CK(cudaMalloc(&ctx->d_ssb_state, (size_t)n_streams * ctx->channels * sizeof(SsbStreamState)));
    CK(cudaMalloc(&ctx->d_local, frames * sizeof(FrameLocal)));
    CK(cudaMalloc(&ctx->d_results, frames * sizeof(sdrb_frame_result)));
    CK(cudaMemset(ctx->d_results, 0, frames * sizeof(sdrb_frame_result)));
    CK(cudaMalloc(&ctx->d_start, frames * ctx->channels * sizeof(float4)));
    if ((rc = upload_geometry(ctx)) != SDRB_OK) return rc;
    if ((rc = upload_ssb(ctx)) != SDRB_OK) return rc;
    // power spectrum, audio and PCM are allocated on first use of the stage (sdrb_process)
    if ((rc = reset_state(ctx)) != SDRB_OK) return rc;
    ctx->configured = true;
    return SDRB_OK;
}

// A new geometry for the SAME streams with every carried state kept: what the reference does when the bridge
// pushes a new configuration (FFTProcessor::configure keeps its tracking / detection members, fft_process.cpp:20-39;
// the SSB filters are function statics and the audio pulse detector a member of SSBProcessor, so they survive any
// change of sample rate or block length: ssb_demod_opt.cpp:261,277, ssb_processor.h:52).
int sdrb_replan(sdrb_ctx* ctx, const sdrb_config* cfg, int max_blocks) {
    if (!ctx || !ctx->configured) return fail(ctx, SDRB_ERR_STATE, "sdrb_replan before sdrb_configure");
    int rc = check_config(ctx, cfg);
    if (rc != SDRB_OK) return rc;
    const int channels = cfg->ssbChannels > 1 ? cfg->ssbChannels : 1;
    if (channels != ctx->channels) return fail(ctx, SDRB_ERR_ARG, "sdrb_replan cannot change the channel count");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    const int S = ctx->n_streams;
    std::vector<FftStreamState> fs((size_t)S);
    std::vector<SsbStreamState> ss((size_t)S * channels);
    CK(cudaMemcpy(fs.data(), ctx->d_fft_state, fs.size() * sizeof(FftStreamState), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(ss.data(), ctx->d_ssb_state, ss.size() * sizeof(SsbStreamState), cudaMemcpyDeviceToHost));
    const int rfMode = ctx->rfMode;
    const uint32_t rfFs = ctx->rfFs;
    const unsigned long long ssb_samples = ctx->ssb_samples;
    // the audio detectors do not depend on the geometry (they run at the nominal 48 kHz): detach them from the
    // context so that sdrb_configure neither frees nor resets them
    PulseState* au_state = ctx->d_au_state;
    float* au_e = ctx->d_au_e;
    ctx->d_au_state = nullptr;
    ctx->d_au_e = nullptr;
    rc = sdrb_configure(ctx, cfg, S, max_blocks);
    if (rc != SDRB_OK) {
        if (au_state) cudaFree(au_state);
        if (au_e) cudaFree(au_e);
        return rc;
    }
    ctx->d_au_state = au_state;
    ctx->d_au_e = au_e;
    for (auto& f : fs)
        if (f.centerFrequency != cfg->centerFrequency) { f.centerFrequency = cfg->centerFrequency; f.centerChanged = 1; }
    CK(cudaMemcpy(ctx->d_fft_state, fs.data(), fs.size() * sizeof(FftStreamState), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->d_ssb_state, ss.data(), ss.size() * sizeof(SsbStreamState), cudaMemcpyHostToDevice));
    ctx->rfMode = rfMode;
    ctx->rfFs = rfFs;
    ctx->ssb_samples = ssb_samples;
    return upload_ssb(ctx);
}

int sdrb_set_center_frequency(sdrb_ctx* ctx, int stream, uint32_t center_hz) {
    if (!ctx || !ctx->configured) return fail(ctx, SDRB_ERR_STATE, "not configured");
    if (stream < -1 || stream >= ctx->n_streams) return fail(ctx, SDRB_ERR_ARG, "stream %d out of range", stream);
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    std::vector<FftStreamState> fs(ctx->n_streams);
    CK(cudaMemcpy(fs.data(), ctx->d_fft_state, fs.size() * sizeof(FftStreamState), cudaMemcpyDeviceToHost));
    for (int s = 0; s < ctx->n_streams; s++)
        if (stream < 0 || s == stream) { fs[s].centerFrequency = center_hz; fs[s].centerChanged = 1; }
    CK(cudaMemcpy(ctx->d_fft_state, fs.data(), fs.size() * sizeof(FftStreamState), cudaMemcpyHostToDevice));
    if (stream < 0) ctx->cfg.centerFrequency = center_hz;
    return SDRB_OK;
}

int sdrb_set_focus_range(sdrb_ctx* ctx, int focus_khz) {
    if (!ctx || !ctx->configured) return fail(ctx, SDRB_ERR_STATE, "not configured");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->cfg.freqFocusRangeKhz = focus_khz;
    return upload_geometry(ctx);
}

int sdrb_set_sound_mode(sdrb_ctx* ctx, int mode) {
    if (!ctx || !ctx->configured) return fail(ctx, SDRB_ERR_STATE, "not configured");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->cfg.soundMode = mode;
    return upload_ssb(ctx);
}

int sdrb_pcm_per_block(const sdrb_ctx* ctx) { return (ctx && ctx->configured) ? ctx->ssb.m : SDRB_ERR_STATE; }

}  // extern "C"

// ------------------------------------------------------------------ the hot path
static int ensure_fft_buffers(sdrb_ctx* ctx) {
    const int n = ctx->cfg.samplesPerReading;
    const size_t frames = (size_t)ctx->n_streams * ctx->max_blocks;
    if (!ctx->d_power) CK(cudaMalloc(&ctx->d_power, frames * (size_t)n * sizeof(float)));
    if (ctx->fftPath == 2 && !ctx->d_fbA) {
        const std::vector<cpx> tb = n == 65536 ? make_fb_tables<16>() : make_fb_tables<32>();
        CK(cudaMalloc(&ctx->d_fb_tables, tb.size() * sizeof(cpx)));
        CK(cudaMemcpy(ctx->d_fb_tables, tb.data(), tb.size() * sizeof(cpx), cudaMemcpyHostToDevice));
        size_t l2_bytes = kFourStepL2Bytes;
        if (const char* e = getenv("SDRB_FB_L2_MB")) { const long mb = atol(e); if (mb > 0) l2_bytes = (size_t)mb << 20; }
        size_t group = l2_bytes / ((size_t)n * sizeof(cpx));
        if (group < 1) group = 1;
        if (group > frames) group = frames;
        if (group < frames) group = (size_t)fft_big_best_group(n, (int)group);
        ctx->fb_group = (int)group;
        bool fused = false;   // SDRB_FB_FUSED=1: one persistent kernel, intermediate in an L2-resident ring
        if (const char* e = getenv("SDRB_FB_FUSED")) fused = atoi(e) != 0;
        size_t a_frames = group;
        if (fused) {
            ctx->fb_slots = fft_big_ring_slots(n);
            a_frames = (size_t)ctx->fb_slots;
            CK(cudaMalloc(&ctx->d_fb_ctl, (1 + 2 * group) * sizeof(unsigned)));
        }
        CK(cudaMalloc(&ctx->d_fbA, a_frames * (size_t)n * sizeof(cpx)));
        if (ctx->front) CK(cudaMalloc(&ctx->d_front, group * (size_t)n * sizeof(cpx)));
    }
    if (ctx->fftPath == 3 && !ctx->d_chirp) {
        const int m = bluestein_m(n);
        ctx->blue_m = m;
        std::vector<cpx> chirp, bf, tw((size_t)m);
        make_bluestein_tables(n, m, chirp, bf);
        for (int k = 0; k < m; k++) tw[(size_t)k] = twiddle(k, m);
        CK(cudaMalloc(&ctx->d_chirp, chirp.size() * sizeof(cpx)));
        CK(cudaMemcpy(ctx->d_chirp, chirp.data(), chirp.size() * sizeof(cpx), cudaMemcpyHostToDevice));
        CK(cudaMalloc(&ctx->d_bf, bf.size() * sizeof(cpx)));
        CK(cudaMemcpy(ctx->d_bf, bf.data(), bf.size() * sizeof(cpx), cudaMemcpyHostToDevice));
        CK(cudaMalloc(&ctx->d_tw, tw.size() * sizeof(cpx)));
        CK(cudaMemcpy(ctx->d_tw, tw.data(), tw.size() * sizeof(cpx), cudaMemcpyHostToDevice));
        size_t chunk = kGenericScratchBytes / ((size_t)m * sizeof(cpx));
        if (chunk < 1) chunk = 1;
        if (chunk > frames) chunk = frames;
        CK(cudaMalloc(&ctx->d_sa, chunk * (size_t)m * sizeof(cpx)));
        CK(cudaMalloc(&ctx->d_sb, chunk * (size_t)m * sizeof(cpx)));
        if (ctx->front) CK(cudaMalloc(&ctx->d_front, chunk * (size_t)n * sizeof(cpx)));
    }
    if (ctx->fftPath == 5 && !ctx->d_fs_tables) {
        const int B = n == 262144 ? 8192 : 4096, loga = n == 32768 ? 3 : 5;
        const std::vector<cpx> tb = make_tp_tables(n, B, loga);
        CK(cudaMalloc(&ctx->d_fs_tables, tb.size() * sizeof(cpx)));
        CK(cudaMemcpy(ctx->d_fs_tables, tb.data(), tb.size() * sizeof(cpx), cudaMemcpyHostToDevice));
        size_t chunk = kTwoPassScratchBytes / ((size_t)n * sizeof(cpx));
        if (chunk < 1) chunk = 1;
        if (chunk > frames) chunk = frames;
        CK(cudaMalloc(&ctx->d_sa, chunk * (size_t)n * sizeof(cpx)));
    }
    if (ctx->fftPath == 4 && !ctx->d_fs_tables) {
        const std::vector<cpx> tb = make_fs_tables(n);
        CK(cudaMalloc(&ctx->d_fs_tables, tb.size() * sizeof(cpx)));
        CK(cudaMemcpy(ctx->d_fs_tables, tb.data(), tb.size() * sizeof(cpx), cudaMemcpyHostToDevice));
    }
    if ((ctx->fftPath == 0 || ctx->fftPath == 5) && ctx->front && !ctx->d_front) {   // (path 4 fuses the front end)
        size_t chunk = kGenericScratchBytes / ((size_t)n * sizeof(cpx));
        if (chunk < 1) chunk = 1;
        if (chunk > frames) chunk = frames;
        CK(cudaMalloc(&ctx->d_front, chunk * (size_t)n * sizeof(cpx)));
    }
    if (ctx->fftPath == 0 && !ctx->d_tw) {
        std::vector<cpx> tw(n);
        for (int k = 0; k < n; k++) tw[k] = twiddle(k, n);
        CK(cudaMalloc(&ctx->d_tw, (size_t)n * sizeof(cpx)));
        CK(cudaMemcpy(ctx->d_tw, tw.data(), (size_t)n * sizeof(cpx), cudaMemcpyHostToDevice));
        if (fft_generic_passes(n) > 1) {
            size_t chunk = kGenericScratchBytes / ((size_t)n * sizeof(cpx));
            if (chunk < 1) chunk = 1;
            if (chunk > frames) chunk = frames;
            CK(cudaMalloc(&ctx->d_sa, chunk * (size_t)n * sizeof(cpx)));
            if (fft_generic_passes(n) > 2) CK(cudaMalloc(&ctx->d_sb, chunk * (size_t)n * sizeof(cpx)));
        }
    }
    return SDRB_OK;
}

static int ensure_ssb_buffers(sdrb_ctx* ctx) {
    const size_t frames = (size_t)ctx->n_streams * ctx->max_blocks * ctx->channels;
    const size_t m = ctx->ssb.m > 0 ? ctx->ssb.m : 1;
    if (!ctx->d_a48) CK(cudaMalloc(&ctx->d_a48, frames * m * sizeof(float)));
    if (!ctx->d_pcm) CK(cudaMalloc(&ctx->d_pcm, frames * m * sizeof(int16_t)));
    // scratch of the split AGC / decimator pair: one float per input sample (and per channel of an NCO bank), needed
    // only when the fused pass does not take the decimation factor.  The extended chain also has its pre-pass
    // buffers (removeDC output and mixed samples).
    if (!ctx->d_agc && ctx->ssb_sp == 0) {
        const size_t bytes = frames * (size_t)ctx->cfg.samplesPerReading * sizeof(float);   // frames includes the channels
        if (cudaMalloc(&ctx->d_agc, bytes) != cudaSuccess) {
            cudaGetLastError();
            ctx->d_agc = nullptr;
            return fail(ctx, SDRB_ERR_NOMEM, "no memory for the %zu-byte AGC scratch of the SSB chain", bytes);
        }
    }
    if (ctx->ssb_ext && !ctx->d_ext_work) {
        const size_t fl = ssb_ext_work_floats(ctx->cfg.samplesPerReading, ctx->n_streams, ctx->channels, ctx->max_blocks,
                                              ctx->cfg.ssbTrueSideband != 0);
        if (cudaMalloc(&ctx->d_ext_work, fl * sizeof(float)) != cudaSuccess) {
            cudaGetLastError();
            ctx->d_ext_work = nullptr;
            return fail(ctx, SDRB_ERR_NOMEM, "no memory for the %zu-byte pre-pass buffers of the extended SSB chain", fl * sizeof(float));
        }
    }
    return SDRB_OK;
}

extern "C" {

static int validate_process(sdrb_ctx* ctx, const void* iq, int n_blocks, unsigned stages) {
    if (!ctx || !ctx->configured) return fail(ctx, SDRB_ERR_STATE, "sdrb_process before sdrb_configure");
    if (!iq) return fail(ctx, SDRB_ERR_ARG, "iq is NULL");
    if (n_blocks < 1 || n_blocks > ctx->max_blocks)
        return fail(ctx, SDRB_ERR_ARG, "n_blocks %d outside [1, %d]", n_blocks, ctx->max_blocks);
    if ((stages & ~(SDRB_STAGE_FFT | SDRB_STAGE_SSB | SDRB_STAGE_SPECTRAL_PULSE | SDRB_STAGE_AUDIO_PULSE)) || stages == 0)
        return fail(ctx, SDRB_ERR_ARG, "stages must be a combination of SDRB_STAGE_FFT | SDRB_STAGE_SSB | SDRB_STAGE_*_PULSE");
    if ((stages & SDRB_STAGE_SPECTRAL_PULSE) && !(stages & SDRB_STAGE_FFT))
        return fail(ctx, SDRB_ERR_ARG, "SDRB_STAGE_SPECTRAL_PULSE consumes the FFT records: add SDRB_STAGE_FFT");
    if ((stages & SDRB_STAGE_AUDIO_PULSE) && !(stages & SDRB_STAGE_SSB))
        return fail(ctx, SDRB_ERR_ARG, "SDRB_STAGE_AUDIO_PULSE consumes the PCM: add SDRB_STAGE_SSB");
    const int n = ctx->cfg.samplesPerReading;
    const size_t bps = sdrb_format_bytes(ctx->cfg.inputFormat);
    if ((stages & SDRB_STAGE_FFT) && (is_pow2(n) ? n > (1 << 24) : n > kMaxBluesteinN))
        return fail(ctx, SDRB_ERR_UNSUPPORTED, "FFT length %d: powers of two up to 2^24 and any other length up to 2^22 are built", n);
    if ((stages & SDRB_STAGE_SSB) && ctx->ssb_ext && n % 32 != 0)
        return fail(ctx, SDRB_ERR_UNSUPPORTED, "extended SSB chain with a block of %d samples: the block length must be a "
                                               "multiple of 32 samples", n);
    int rc;
    CK(cudaSetDevice(ctx->device));
    if ((stages & SDRB_STAGE_FFT) && (rc = ensure_fft_buffers(ctx)) != SDRB_OK) return rc;
    if (stages & SDRB_STAGE_SSB) {
        if (ctx->rfMode < 0) {  // first SSB block latches the RF low-pass design
            if (ctx->cfg.soundMode >= 0 && ctx->cfg.soundMode <= 2) ctx->effMode = ctx->cfg.soundMode;
            ctx->rfMode = ctx->effMode;
            ctx->rfFs = ctx->cfg.sampleRate;
            if ((rc = upload_ssb(ctx)) != SDRB_OK) return rc;
        }
        if ((rc = ensure_ssb_buffers(ctx)) != SDRB_OK) return rc;
    }
    if (ctx->cfg.inputFormat == SDRB_RU16 && !ctx->d_air_state) {
        if (n % 4) return fail(ctx, SDRB_ERR_UNSUPPORTED, "SDRB_RU16 needs samplesPerReading %% 4 == 0");
        const size_t S = (size_t)ctx->n_streams, real = 2 * (size_t)ctx->max_blocks * n;
        CK(cudaMalloc(&ctx->d_air_state, S * sizeof(AirspyStreamState)));
        CK(cudaMemset(ctx->d_air_state, 0, S * sizeof(AirspyStreamState)));
        CK(cudaMalloc(&ctx->d_air_iq, S * (real / 2) * sizeof(cpx)));
        airspy_chunking(&ctx->air_chunk, &ctx->air_warm);
        ctx->air_work_per_stream = airspy_work_floats(1, (long long)real, ctx->air_chunk);
        CK(cudaMalloc(&ctx->d_air_work, S * ctx->air_work_per_stream * sizeof(float)));
    }
    if ((stages & (SDRB_STAGE_SPECTRAL_PULSE | SDRB_STAGE_AUDIO_PULSE)) && (rc = ensure_pulse_buffers(ctx, stages)) != SDRB_OK)
        return rc;
    return SDRB_OK;
}

// Every kernel of the path for streams [s0, s0 + ns) of a call of n_blocks readings, on stream st.
// d_iq points at stream s0's first sample; device outputs are laid out [n_streams][n_blocks][...].
static int run_streams(sdrb_ctx* ctx, const void* d_iq, int s0, int ns, int n_blocks, unsigned stages,
                       cudaStream_t st, bool time_fft) {
    const int n = ctx->cfg.samplesPerReading;
    int fmt = ctx->cfg.inputFormat;
    const int frames = ns * n_blocks;
    const size_t f0s = (size_t)s0 * n_blocks;   // first frame of the range
    int launches = 0;
    struct TraceArm {
        TraceArm(TraceSink* t) { g_trace = t; }
        ~TraceArm() { g_trace = nullptr; }
    } arm(ctx->trace);
    trace_mark("call start", st);
    if (fmt == SDRB_RU16) {
        // Airspy: raw real ADC words -> cf32 through the libairspy converter chain (state carried per stream),
        // then the CF32 kernels.  The intermediate is laid out for THIS call's n_blocks.
        const long long real = 2LL * n_blocks * n;
        cpx* iq_out = ctx->d_air_iq + f0s * n;
        CK(launch_airspy_front(static_cast<const uint16_t*>(d_iq), ctx->d_air_state + s0, iq_out,
                               ctx->d_air_work + (size_t)s0 * ctx->air_work_per_stream, ctx->air_chunk, ctx->air_warm, ns,
                               real, st, &launches));
        d_iq = iq_out;
        fmt = FMT_CF32;
    }
    const size_t bps = sdrb_format_bytes(fmt);
    // the SSB chain on `ss` (the caller's stream, or the side stream when it runs beside the FFT chain)
    auto run_ssb = [&](cudaStream_t ss) -> int {
        const size_t m = ctx->ssb.m > 0 ? ctx->ssb.m : 0;
        if (!ctx->ssb_ext && ctx->ssb_any) {
            CK(launch_ssb_any(fmt, d_iq, ctx->ssb, ctx->d_taps_lin, ctx->d_ssb_state + s0, ctx->d_agc + (size_t)s0 * n,
                              ctx->d_pcm + f0s * m, ns, n_blocks, ss));
            launches += 1;
            trace_mark("ssb any-length", ss);
        } else if (!ctx->ssb_ext) {
            CK(launch_ssb(fmt, d_iq, ctx->ssb, ctx->d_taps, ctx->d_tapsE, ctx->ssb_sp, static_cast<float2*>(ctx->d_start) + f0s,
                          ctx->d_ssb_state + s0, ctx->d_a48 + f0s * m, ctx->d_pcm + f0s * m,
                          ctx->d_agc ? ctx->d_agc + f0s * (size_t)n : nullptr, ns, n_blocks, ctx->sm_count, ss, &launches));
        } else {
            const size_t C = (size_t)ctx->channels;
            CK(launch_ssb_ext(fmt, d_iq, ctx->ssb, ctx->d_taps, ctx->d_tapsE, ctx->ssb_sp, ctx->d_inc1, ctx->inc2,
                              ctx->cfg.ssbTrueSideband != 0,
                              static_cast<float4*>(ctx->d_start) + f0s * C, ctx->d_ssb_state + (size_t)s0 * C,
                              ctx->d_a48 + f0s * C * m, ctx->d_pcm + f0s * C * m,
                              ctx->d_agc ? ctx->d_agc + f0s * C * (size_t)n : nullptr,
                              ctx->d_ext_work + ssb_ext_work_floats(n, s0, ctx->channels, n_blocks, ctx->cfg.ssbTrueSideband != 0),
                              ns, ctx->channels, n_blocks, ctx->ssb_samples, ss, &launches));
        }
        return SDRB_OK;
    };
    const bool fork = (stages & SDRB_STAGE_FFT) && (stages & SDRB_STAGE_SSB) && ctx->ssb_fork;
    if (fork) {
        if (!ctx->s_ssb) {
            CK(cudaStreamCreateWithFlags(&ctx->s_ssb, cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
        }
        CK(cudaEventRecord(ctx->ev_fork, st));
        CK(cudaStreamWaitEvent(ctx->s_ssb, ctx->ev_fork, 0));
        if (ctx->ssb_first) {
            int rc = run_ssb(ctx->s_ssb);
            if (rc != SDRB_OK) return rc;
            CK(cudaEventRecord(ctx->ev_join, ctx->s_ssb));
        }
    }
    if (stages & SDRB_STAGE_FFT) {
        float* power = ctx->d_power + f0s * n;
        if (time_fft && ctx->fftPath != 1) CK(cudaEventRecord(ctx->ev0, st));
        int fft_launches = 0;
        bool estimator_done = false, tracker_done = false;
        if (ctx->fftPath == 1) {
            // window / block-mean removal (extension) are fused into phase 1 of the 16K kernel.
            // Frames go through FFT -> estimator in chunks whose spectra (64 KiB each) are still in
            // the 126 MB L2 when the estimator reads its bands back.
            int chunk = ctx->est_chunk > 0 ? ctx->est_chunk : frames;
            bool overlap = false;
            if (ctx->est_chunk <= 0 && ctx->est_overlap && ns >= 4 && frames >= 8 * ctx->sm_count) {
                chunk = ((ns + 3) / 4) * n_blocks;       // four chunks, cut on stream boundaries (the tracker walks streams)
                overlap = true;
                if (!ctx->s_aux) {
                    CK(cudaStreamCreateWithFlags(&ctx->s_aux, cudaStreamNonBlocking));
                    for (auto& e : ctx->ev_chunk) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                    CK(cudaEventCreateWithFlags(&ctx->ev_aux, cudaEventDisableTiming));
                }
            }
            int ci = 0;
            for (int f0 = 0; f0 < frames; f0 += chunk, ci++) {
                const int nf = frames - f0 < chunk ? frames - f0 : chunk;
                const void* src = static_cast<const unsigned char*>(d_iq) + (size_t)f0 * n * bps;
                if (time_fft && f0 == 0) CK(cudaEventRecord(ctx->ev0, st));
                CK(launch_fft16k2(fmt, src, power + (size_t)f0 * n, ctx->d_tables16k, nf, ctx->sm_count, st, ctx->front,
                                  ctx->d_win, fork ? ctx->d_ticket : nullptr));
                if (time_fft && f0 == 0) { CK(cudaEventRecord(ctx->ev1, st)); ctx->ev_valid = true; ctx->last_fft_frames = nf; }
                trace_mark("fft16k", st);
                fft_launches++;
                if (chunk < frames) {
                    cudaStream_t es = st;
                    if (overlap) {
                        CK(cudaEventRecord(ctx->ev_chunk[ci & 3], st));
                        CK(cudaStreamWaitEvent(ctx->s_aux, ctx->ev_chunk[ci & 3], 0));
                        es = ctx->s_aux;
                    }
                    CK(launch_estimator(power + (size_t)f0 * n, ctx->d_geo, ctx->geo, ctx->d_local + f0s + f0,
                                        ctx->d_est_scratch, f0s + f0, ctx->n_streams * ctx->max_blocks, nf, es));
                    launches += (ctx->geo.focusLen > 0 && !estimator_inline_scan(ctx->geo)) ? 2 : 1;
                    if (overlap) {
                        CK(launch_tracker(ctx->d_local + f0s + f0, ctx->d_geo, ctx->d_fft_state + s0 + f0 / n_blocks,
                                          ctx->d_results + f0s + f0, nf / n_blocks, n_blocks, es));
                        launches += 1;
                    }
                }
            }
            estimator_done = chunk < frames;
            if (overlap) {
                CK(cudaEventRecord(ctx->ev_aux, ctx->s_aux));
                CK(cudaStreamWaitEvent(st, ctx->ev_aux, 0));
                tracker_done = true;
            }
        } else if (ctx->fftPath == 2) {
            // four-step: frames in groups whose intermediate stays L2-resident between the two kernels
            for (int f0 = 0; f0 < frames; f0 += ctx->fb_group) {
                const int nf = frames - f0 < ctx->fb_group ? frames - f0 : ctx->fb_group;
                const void* src = static_cast<const unsigned char*>(d_iq) + (size_t)f0 * n * bps;
                int src_fmt = fmt;
                if (ctx->front) {
                    CK(launch_fft_front(fmt, src, ctx->d_front, ctx->d_win, n, nf, ctx->front, st));
                    fft_launches++;
                    src = ctx->d_front;
                    src_fmt = FMT_CF32;
                }
                if (ctx->fb_slots) {
                    CK(launch_fft_big_fused(src_fmt, src, power + (size_t)f0 * n, ctx->d_fbA, ctx->fb_slots, ctx->d_fb_ctl,
                                            ctx->d_fb_tables, n, nf, st));
                    fft_launches += 1;
                } else {
                    CK(launch_fft_big(src_fmt, src, power + (size_t)f0 * n, ctx->d_fbA, ctx->d_fb_tables, n, nf, st));
                    fft_launches += 2;
                }
            }
        } else if (ctx->fftPath == 3) {
            size_t chunk = kGenericScratchBytes / ((size_t)ctx->blue_m * sizeof(cpx));
            if (chunk < 1) chunk = 1;
            for (size_t f0 = 0; f0 < (size_t)frames; f0 += chunk) {
                const int nf = (int)((size_t)frames - f0 < chunk ? (size_t)frames - f0 : chunk);
                const void* src = static_cast<const unsigned char*>(d_iq) + f0 * n * bps;
                int src_fmt = fmt;
                if (ctx->front) {
                    CK(launch_fft_front(fmt, src, ctx->d_front, ctx->d_win, n, nf, ctx->front, st));
                    fft_launches++;
                    src = ctx->d_front;
                    src_fmt = FMT_CF32;
                }
                CK(launch_fft_bluestein(src_fmt, src, power + f0 * n, ctx->d_sa, ctx->d_sb, ctx->d_chirp, ctx->d_bf,
                                        ctx->d_tw, n, ctx->blue_m, nf, st, &fft_launches));
            }
        } else {
            // multi-pass Stockham (scratch-sized chunks), or the small one-CTA kernel (4): whole call unless the
            // extension front end needs its cf32 pre-pass buffer
            size_t chunk = kGenericScratchBytes / ((size_t)n * sizeof(cpx));
            if (chunk < 1) chunk = 1;
            if (ctx->fftPath == 4) chunk = (size_t)frames;
            if (ctx->fftPath == 5 && !ctx->front) chunk = kTwoPassScratchBytes / ((size_t)n * sizeof(cpx));
            for (size_t f0 = 0; f0 < (size_t)frames; f0 += chunk) {
                const int nf = (int)((size_t)frames - f0 < chunk ? (size_t)frames - f0 : chunk);
                const void* src = static_cast<const unsigned char*>(d_iq) + f0 * n * bps;
                int src_fmt = fmt;
                if (ctx->front && ctx->fftPath != 4) {   // extension front end as a pre-pass: raw IQ -> (x - mean) * w in cf32
                    CK(launch_fft_front(fmt, src, ctx->d_front, ctx->d_win, n, nf, ctx->front, st));
                    fft_launches++;
                    src = ctx->d_front;
                    src_fmt = FMT_CF32;
                }
                if (ctx->fftPath == 4) {
                    CK(launch_fft_small(src_fmt, src, power + f0 * n, ctx->d_fs_tables, n, nf, ctx->sm_count, st, ctx->front,
                                        ctx->d_win));
                    fft_launches++;
                } else if (ctx->fftPath == 5) {
                    CK(launch_fft_twopass(src_fmt, src, power + f0 * n, ctx->d_sa, ctx->d_fs_tables, n, nf, ctx->sm_count, st));
                    fft_launches += 2;
                } else {
                    CK(launch_fft_generic(src_fmt, src, power + f0 * n, ctx->d_sa, ctx->d_sb, ctx->d_tw, n, nf, st,
                                          &fft_launches));
                }
            }
        }
        if (time_fft && ctx->fftPath != 1) { CK(cudaEventRecord(ctx->ev1, st)); ctx->ev_valid = true; ctx->last_fft_frames = frames; }
        ctx->last_fft_launches = ctx->fftPath == 1 ? 1 : fft_launches;
        launches += fft_launches;
        if (!estimator_done) {
            CK(launch_estimator(power, ctx->d_geo, ctx->geo, ctx->d_local + f0s, ctx->d_est_scratch, f0s,
                                ctx->n_streams * ctx->max_blocks, frames, st));
            launches += (ctx->geo.focusLen > 0 && !estimator_inline_scan(ctx->geo)) ? 2 : 1;
        }
        if (!estimator_done) trace_mark("band scan + estimator", st);
        if (!tracker_done) {
            CK(launch_tracker(ctx->d_local + f0s, ctx->d_geo, ctx->d_fft_state + s0, ctx->d_results + f0s, ns, n_blocks, st));
            launches += 1;
            trace_mark("tracker", st);
        }
    }
    if ((stages & SDRB_STAGE_SSB) && !fork) {
        int rc = run_ssb(st);
        if (rc != SDRB_OK) return rc;
    }
    if (fork) {
        // The FFT chain is launched first: its persistent CTAs (one per SM, 181 KiB of shared memory) are resident
        // before the state pass's one-warp CTAs arrive, which then fit beside them two to an SM; the other way round
        // an SM that collected three of those has no room for its FFT CTA until they end.
        if (!ctx->ssb_first) {
            int rc = run_ssb(ctx->s_ssb);
            if (rc != SDRB_OK) return rc;
            CK(cudaEventRecord(ctx->ev_join, ctx->s_ssb));
        }
        CK(cudaStreamWaitEvent(st, ctx->ev_join, 0));
        trace_mark("join", st);
    }
    if (stages & (SDRB_STAGE_SPECTRAL_PULSE | SDRB_STAGE_AUDIO_PULSE)) {
        const size_t C = (size_t)ctx->channels;
        sdrb_pulse_result* out = ctx->d_pulse + f0s * C;
        CK(cudaMemsetAsync(out, 0, (size_t)frames * C * sizeof(sdrb_pulse_result), st));
        if (stages & SDRB_STAGE_SPECTRAL_PULSE) {
            CK(launch_pulse_spectral(ctx->d_results + f0s, ctx->d_sp_state + s0, ctx->d_sp_e + (size_t)s0 * ctx->sp_cfg.ringCap,
                                     ctx->d_sp_f + (size_t)s0 * ctx->sp_cfg.ringCap, ctx->sp_cfg, out, ns, n_blocks,
                                     ctx->channels, st));
            launches += 1;
        }
        if ((stages & SDRB_STAGE_AUDIO_PULSE) && ctx->ssb.m > 0) {
            CK(launch_pulse_audio(ctx->d_pcm + f0s * C * ctx->ssb.m, ctx->d_au_state + (size_t)s0 * C,
                                  ctx->d_au_e + (size_t)s0 * C * ctx->au_cfg.ringCap, ctx->au_cfg, out, ns * ctx->channels,
                                  n_blocks, ctx->ssb.m, st));
            launches += 1;
        }
    }
    ctx->launches += (uint64_t)launches;
    return SDRB_OK;
}

int sdrb_process(sdrb_ctx* ctx, const void* iq, int on_device, int n_blocks, unsigned stages) {
    int rc = validate_process(ctx, iq, n_blocks, stages);
    if (rc != SDRB_OK) return rc;
    const size_t in_bytes = (size_t)ctx->n_streams * n_blocks * ctx->cfg.samplesPerReading *
                            sdrb_format_bytes(ctx->cfg.inputFormat);
    cudaStream_t st = ctx->stream;
    const void* d_iq = iq;
    if (!on_device) {
        if (!ctx->d_in) CK(cudaMalloc(&ctx->d_in, (size_t)ctx->n_streams * ctx->max_blocks *
                                                   ctx->cfg.samplesPerReading * sdrb_format_bytes(ctx->cfg.inputFormat)));
        CK(cudaMemcpyAsync(ctx->d_in, iq, in_bytes, cudaMemcpyHostToDevice, st));
        d_iq = ctx->d_in;
    }
    ctx->ev_valid = false;
    if ((rc = run_streams(ctx, d_iq, 0, ctx->n_streams, n_blocks, stages, st, ctx->timing)) != SDRB_OK) return rc;
    ctx->last_blocks = n_blocks;
    ctx->last_stages |= stages;
    if (stages & SDRB_STAGE_SSB) ctx->ssb_samples += (unsigned long long)n_blocks * ctx->cfg.samplesPerReading;
    if (!on_device) CK(cudaStreamSynchronize(st));  // the host buffer is only borrowed for the call
    return SDRB_OK;
}

// Host-to-host call: what soapyCallback does for every reading (copy in, both processors, all
// payloads back out: src/sdr-bridge-java-soapy.cpp:441-466, ssb_processor.cpp:103-107), for a
// whole batch.  Streams are cut into chunks; the H2D copy of chunk c+1, the kernels of chunk c
// and the D2H copies of chunk c-1 run concurrently on three CUDA streams (full-duplex PCIe).
// Any of results_out / power_out / pcm_out may be NULL (that payload then stays on the device).
int sdrb_process_host(sdrb_ctx* ctx, const void* iq_host, int n_blocks, unsigned stages,
                      sdrb_frame_result* results_out, float* power_out, int16_t* pcm_out) {
    return sdrb_process_host_ex(ctx, iq_host, n_blocks, stages, results_out, power_out, SDRB_POWER_LINEAR_F32, pcm_out);
}

static int ensure_view(sdrb_ctx* ctx) {
    if (!ctx->d_view)
        CK(cudaMalloc(&ctx->d_view, (size_t)ctx->n_streams * ctx->max_blocks * ctx->cfg.samplesPerReading * sizeof(float)));
    return SDRB_OK;
}

int sdrb_process_host_ex(sdrb_ctx* ctx, const void* iq_host, int n_blocks, unsigned stages,
                         sdrb_frame_result* results_out, void* power_out, int power_format, int16_t* pcm_out) {
    int rc = validate_process(ctx, iq_host, n_blocks, stages);
    if (rc != SDRB_OK) return rc;
    if (power_format < SDRB_POWER_LINEAR_F32 || power_format > SDRB_POWER_DB_F16)
        return fail(ctx, SDRB_ERR_ARG, "power_format %d is not a sdrb_power_format", power_format);
    const bool view = power_out && power_format != SDRB_POWER_LINEAR_F32 && (stages & SDRB_STAGE_FFT);
    if (view && (rc = ensure_view(ctx)) != SDRB_OK) return rc;
    const size_t view_elem = power_format == SDRB_POWER_DB_F16 ? 2 : 4;
    const int n = ctx->cfg.samplesPerReading, S = ctx->n_streams;
    const size_t bps = sdrb_format_bytes(ctx->cfg.inputFormat);
    if (!ctx->d_in) CK(cudaMalloc(&ctx->d_in, (size_t)S * ctx->max_blocks * n * bps));
    if (!ctx->s_in) {
        CK(cudaStreamCreateWithFlags(&ctx->s_in, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking));
        for (int i = 0; i < kMaxChunks; i++) {
            CK(cudaEventCreateWithFlags(&ctx->ev_in[i], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming));
        }
    }
    const int want = ctx->host_chunks < 1 ? 1 : (ctx->host_chunks > kMaxChunks ? kMaxChunks : ctx->host_chunks);
    const int chunks = S < want ? S : want;
    const size_t stream_in = (size_t)n_blocks * n * bps;          // bytes of one stream's IQ
    const size_t m = ctx->ssb.m > 0 ? ctx->ssb.m : 0;
    cudaStream_t st = ctx->stream;
    CK(cudaStreamSynchronize(st));                                  // earlier work on the compute stream
    ctx->ev_valid = false;
    // An SSB chain with a carried-state pass (several readings per call, or the extended chain) is not cut into
    // chunks: that pass is one latency-bound warp per 32 chains and takes the same time for 32 streams as for 256, so
    // every chunk would pay it in full (C3: 8 chunks, 315 ms per call against 41 ms of kernels).  The chunks carry the
    // copies and the FFT chain; the SSB chain runs once, over all streams, when the last chunk is in, and its PCM goes
    // out behind it.  One reading per call has no state pass (the fused kernel carries the state) and stays chunked.
    const unsigned whole = (ctx->ssb_ext || n_blocks > 1) ? (stages & (SDRB_STAGE_SSB | SDRB_STAGE_AUDIO_PULSE)) : 0u;
    const unsigned chunk_stages = stages & ~whole;
    int last_in = -1;
    for (int c = 0; c < chunks; c++) {
        const int s0 = (int)((long long)S * c / chunks), s1 = (int)((long long)S * (c + 1) / chunks);
        const int ns = s1 - s0;
        if (ns <= 0) continue;
        const size_t f0 = (size_t)s0 * n_blocks, nf = (size_t)ns * n_blocks;
        unsigned char* d_chunk = static_cast<unsigned char*>(ctx->d_in) + s0 * stream_in;
        CK(cudaMemcpyAsync(d_chunk, static_cast<const unsigned char*>(iq_host) + s0 * stream_in, ns * stream_in,
                           cudaMemcpyHostToDevice, ctx->s_in));
        CK(cudaEventRecord(ctx->ev_in[c], ctx->s_in));
        last_in = c;
        if (!chunk_stages) continue;
        CK(cudaStreamWaitEvent(st, ctx->ev_in[c], 0));
        if ((rc = run_streams(ctx, d_chunk, s0, ns, n_blocks, chunk_stages, st, false)) != SDRB_OK) return rc;
        CK(cudaEventRecord(ctx->ev_done[c], st));
        CK(cudaStreamWaitEvent(ctx->s_out, ctx->ev_done[c], 0));
        if ((stages & SDRB_STAGE_FFT) && results_out)
            CK(cudaMemcpyAsync(results_out + f0, ctx->d_results + f0, nf * sizeof(sdrb_frame_result),
                               cudaMemcpyDeviceToHost, ctx->s_out));
        if (view) {   // dB view of this chunk's spectra on the compute stream, behind its kernels
            CK(launch_power_view(ctx->d_power + f0 * n, static_cast<unsigned char*>(ctx->d_view) + f0 * n * view_elem,
                                 nf * n, power_format, st));
            ctx->launches += 1;
            CK(cudaEventRecord(ctx->ev_done[c], st));
            CK(cudaStreamWaitEvent(ctx->s_out, ctx->ev_done[c], 0));
            CK(cudaMemcpyAsync(static_cast<unsigned char*>(power_out) + f0 * n * view_elem,
                               static_cast<unsigned char*>(ctx->d_view) + f0 * n * view_elem, nf * n * view_elem,
                               cudaMemcpyDeviceToHost, ctx->s_out));
        } else if ((stages & SDRB_STAGE_FFT) && power_out)
            CK(cudaMemcpyAsync(static_cast<float*>(power_out) + f0 * n, ctx->d_power + f0 * n, nf * n * sizeof(float),
                               cudaMemcpyDeviceToHost, ctx->s_out));
        if ((chunk_stages & SDRB_STAGE_SSB) && pcm_out && m)
            CK(cudaMemcpyAsync(pcm_out + f0 * m * ctx->channels, ctx->d_pcm + f0 * m * ctx->channels,
                               nf * m * ctx->channels * sizeof(int16_t), cudaMemcpyDeviceToHost, ctx->s_out));
    }
    if (whole && last_in >= 0) {
        CK(cudaStreamWaitEvent(st, ctx->ev_in[last_in], 0));        // the copies are in order on s_in: all chunks are in
        if ((rc = run_streams(ctx, ctx->d_in, 0, S, n_blocks, whole, st, false)) != SDRB_OK) return rc;
        CK(cudaEventRecord(ctx->ev_done[0], st));
        CK(cudaStreamWaitEvent(ctx->s_out, ctx->ev_done[0], 0));
        if (pcm_out && m)
            CK(cudaMemcpyAsync(pcm_out, ctx->d_pcm, (size_t)S * n_blocks * m * ctx->channels * sizeof(int16_t),
                               cudaMemcpyDeviceToHost, ctx->s_out));
    }
    ctx->last_blocks = n_blocks;
    ctx->last_stages |= stages;
    if (stages & SDRB_STAGE_SSB) ctx->ssb_samples += (unsigned long long)n_blocks * n;
    CK(cudaStreamSynchronize(ctx->s_out));
    CK(cudaStreamSynchronize(st));
    return SDRB_OK;
}

int sdrb_sync(sdrb_ctx* ctx) {
    if (!ctx) return SDRB_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->trace) ctx->trace->dump();
    return SDRB_OK;
}

int sdrb_get_results(sdrb_ctx* ctx, sdrb_frame_result* host_out, size_t capacity_records) {
    if (!ctx || !ctx->configured || !(ctx->last_stages & SDRB_STAGE_FFT))
        return fail(ctx, SDRB_ERR_STATE, "no FFT results: call sdrb_process with SDRB_STAGE_FFT first");
    const size_t nrec = (size_t)ctx->n_streams * ctx->last_blocks;
    if (!host_out || capacity_records < nrec) return fail(ctx, SDRB_ERR_ARG, "results buffer holds %zu < %zu records", capacity_records, nrec);
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(host_out, ctx->d_results, nrec * sizeof(sdrb_frame_result), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SDRB_OK;
}

void sdrb_audio_pulse_config_default(sdrb_audio_pulse_config* c) {
    if (!c) return;
    // AudioPulseDetector::Config defaults (src/ssb/audio_pulse_detector.h:20-38)
    c->sampleRate = 48000.f; c->fMin = 1500.f; c->fMax = 4000.f; c->fsEnergy = 100.f; c->smoothCutoff = 5.f;
    c->zDefaultS = 0.666f; c->tTargetInit = 1.75f; c->dtTolS = 0.150f;
    c->snrMin = 1.0f; c->snrRhythm = 1.1f; c->snrStrong = 2.0f; c->dispersionMax = 1.3f; c->sumNMax = 7;
    c->liveWindowT = 4.0f; c->liveDivisor = 3.0f; c->noiseRefFar = 80; c->noiseRefNear = 40;
}

int sdrb_set_pulse_config(sdrb_ctx* ctx, const sdrb_audio_pulse_config* cfg) {
    if (!ctx || !cfg) return SDRB_ERR_ARG;
    if (!ctx->configured) return fail(ctx, SDRB_ERR_STATE, "sdrb_set_pulse_config before sdrb_configure");
    if (!(cfg->fsEnergy > 0.f) || !(cfg->sampleRate > 0.f) || cfg->fsEnergy > 1.0e5f)
        return fail(ctx, SDRB_ERR_ARG, "audio pulse config: sampleRate and fsEnergy must be positive (fsEnergy <= 1e5)");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->au_user = *cfg;
    ctx->au_user_set = true;
    return reset_pulse(ctx, false, true);
}

int sdrb_get_pulse_results(sdrb_ctx* ctx, sdrb_pulse_result* host_out, size_t capacity_records) {
    if (!ctx || !ctx->configured || !(ctx->last_stages & (SDRB_STAGE_SPECTRAL_PULSE | SDRB_STAGE_AUDIO_PULSE)))
        return fail(ctx, SDRB_ERR_STATE, "no pulse results: call sdrb_process with a SDRB_STAGE_*_PULSE stage first");
    const size_t nrec = (size_t)ctx->n_streams * ctx->channels * ctx->last_blocks;
    if (!host_out || capacity_records < nrec) return fail(ctx, SDRB_ERR_ARG, "pulse buffer holds %zu < %zu records", capacity_records, nrec);
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(host_out, ctx->d_pulse, nrec * sizeof(sdrb_pulse_result), cudaMemcpyDeviceToHost));
    return SDRB_OK;
}

int sdrb_pulse_results_device(sdrb_ctx* ctx, const sdrb_pulse_result** dev_ptr) {
    if (!ctx || !dev_ptr || !ctx->d_pulse) return fail(ctx, SDRB_ERR_STATE, "no pulse results");
    *dev_ptr = ctx->d_pulse;
    return SDRB_OK;
}

int sdrb_get_airspy_iq(sdrb_ctx* ctx, float* host_out, size_t capacity_floats) {
    if (!ctx || !host_out) return fail(ctx, SDRB_ERR_ARG, "null argument");
    if (!ctx->d_air_iq || ctx->last_blocks <= 0) return fail(ctx, SDRB_ERR_STATE, "no SDRB_RU16 call yet");
    const size_t need = 2 * (size_t)ctx->n_streams * ctx->last_blocks * ctx->cfg.samplesPerReading;
    if (capacity_floats < need) return fail(ctx, SDRB_ERR_ARG, "capacity %zu < %zu floats", capacity_floats, need);
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(host_out, ctx->d_air_iq, need * sizeof(float), cudaMemcpyDeviceToHost));
    return SDRB_OK;
}

int sdrb_airspy_redone_chunks(sdrb_ctx* ctx, uint64_t* out) {
    if (!ctx || !out) return fail(ctx, SDRB_ERR_ARG, "null argument");
    *out = 0;
    if (!ctx->d_air_state) return SDRB_OK;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    std::vector<AirspyStreamState> h((size_t)ctx->n_streams);
    CK(cudaMemcpy(h.data(), ctx->d_air_state, h.size() * sizeof(AirspyStreamState), cudaMemcpyDeviceToHost));
    for (const auto& a : h) *out += a.redone;
    return SDRB_OK;
}

int sdrb_results_device(sdrb_ctx* ctx, const sdrb_frame_result** dev_ptr) {
    if (!ctx || !ctx->configured || !dev_ptr) return fail(ctx, SDRB_ERR_STATE, "not configured");
    *dev_ptr = ctx->d_results;
    return SDRB_OK;
}

int sdrb_power_device(sdrb_ctx* ctx, const float** dev_ptr) {
    if (!ctx || !ctx->configured || !dev_ptr) return fail(ctx, SDRB_ERR_STATE, "not configured");
    if (!ctx->d_power) return fail(ctx, SDRB_ERR_STATE, "no power spectrum yet");
    *dev_ptr = ctx->d_power;
    return SDRB_OK;
}

int sdrb_get_power(sdrb_ctx* ctx, int stream, int block, float* host_out) {
    if (!ctx || !ctx->configured || !ctx->d_power) return fail(ctx, SDRB_ERR_STATE, "no power spectrum yet");
    if (stream < 0 || stream >= ctx->n_streams || block < 0 || block >= ctx->last_blocks || !host_out)
        return fail(ctx, SDRB_ERR_ARG, "stream/block out of range");
    const size_t n = ctx->cfg.samplesPerReading;
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(host_out, ctx->d_power + ((size_t)stream * ctx->last_blocks + block) * n, n * sizeof(float),
                       cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SDRB_OK;
}

int sdrb_get_power_db(sdrb_ctx* ctx, int stream, int block, float* host_out) {
    if (!ctx || !ctx->configured || !ctx->d_power) return fail(ctx, SDRB_ERR_STATE, "no power spectrum yet");
    if (stream < 0 || stream >= ctx->n_streams || block < 0 || block >= ctx->last_blocks || !host_out)
        return fail(ctx, SDRB_ERR_ARG, "stream/block out of range");
    const size_t n = ctx->cfg.samplesPerReading;
    CK(cudaSetDevice(ctx->device));
    int rc = ensure_view(ctx);
    if (rc != SDRB_OK) return rc;
    const size_t off = ((size_t)stream * ctx->last_blocks + block) * n;
    CK(launch_power_view(ctx->d_power + off, static_cast<float*>(ctx->d_view) + off, n, SDRB_POWER_DB_F32, ctx->stream));
    ctx->launches += 1;
    CK(cudaMemcpyAsync(host_out, static_cast<float*>(ctx->d_view) + off, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SDRB_OK;
}

int sdrb_get_power_all(sdrb_ctx* ctx, float* host_out, size_t capacity_floats) {
    if (!ctx || !ctx->configured || !ctx->d_power || !(ctx->last_stages & SDRB_STAGE_FFT))
        return fail(ctx, SDRB_ERR_STATE, "no power spectrum yet");
    const size_t nf = (size_t)ctx->n_streams * ctx->last_blocks * ctx->cfg.samplesPerReading;
    if (!host_out || capacity_floats < nf) return fail(ctx, SDRB_ERR_ARG, "power buffer holds %zu < %zu floats", capacity_floats, nf);
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(host_out, ctx->d_power, nf * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SDRB_OK;
}

int sdrb_pcm_device(sdrb_ctx* ctx, const int16_t** dev_ptr) {
    if (!ctx || !ctx->configured || !dev_ptr) return fail(ctx, SDRB_ERR_STATE, "not configured");
    if (!ctx->d_pcm) return fail(ctx, SDRB_ERR_STATE, "no PCM yet");
    *dev_ptr = ctx->d_pcm;
    return SDRB_OK;
}

int sdrb_get_pcm(sdrb_ctx* ctx, int16_t* host_out, size_t capacity_samples) {
    if (!ctx || !ctx->configured || !ctx->d_pcm || !(ctx->last_stages & SDRB_STAGE_SSB))
        return fail(ctx, SDRB_ERR_STATE, "no PCM: call sdrb_process with SDRB_STAGE_SSB first");
    const size_t ns = (size_t)ctx->n_streams * ctx->channels * ctx->last_blocks * ctx->ssb.m;
    if (capacity_samples < ns || (!host_out && ns)) return fail(ctx, SDRB_ERR_ARG, "PCM buffer holds %zu < %zu samples", capacity_samples, ns);
    CK(cudaSetDevice(ctx->device));
    if (ns) CK(cudaMemcpyAsync(host_out, ctx->d_pcm, ns * sizeof(int16_t), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SDRB_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ state, synth, introspection
extern "C" {

size_t sdrb_state_bytes(const sdrb_ctx* ctx) {
    if (!ctx || !ctx->configured) return 0;
    return sizeof(StateBlobHeader) +
           (size_t)ctx->n_streams * (sizeof(FftStreamState) + (size_t)ctx->channels * sizeof(SsbStreamState));
}

int sdrb_save_state(sdrb_ctx* ctx, void* host_blob, size_t bytes) {
    if (!ctx || !ctx->configured) return fail(ctx, SDRB_ERR_STATE, "not configured");
    if (!host_blob || bytes < sdrb_state_bytes(ctx)) return fail(ctx, SDRB_ERR_ARG, "state blob too small");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    unsigned char* p = static_cast<unsigned char*>(host_blob);
    StateBlobHeader h = {kBlobMagic, SDRB_ABI_VERSION, ctx->n_streams, ctx->rfMode, ctx->ssb_samples, ctx->rfFs, 0u};
    memcpy(p, &h, sizeof h);
    p += sizeof h;
    CK(cudaMemcpy(p, ctx->d_fft_state, (size_t)ctx->n_streams * sizeof(FftStreamState), cudaMemcpyDeviceToHost));
    p += (size_t)ctx->n_streams * sizeof(FftStreamState);
    CK(cudaMemcpy(p, ctx->d_ssb_state, (size_t)ctx->n_streams * ctx->channels * sizeof(SsbStreamState),
                  cudaMemcpyDeviceToHost));
    return SDRB_OK;
}

int sdrb_load_state(sdrb_ctx* ctx, const void* host_blob, size_t bytes) {
    if (!ctx || !ctx->configured) return fail(ctx, SDRB_ERR_STATE, "not configured");
    if (!host_blob || bytes < sdrb_state_bytes(ctx)) return fail(ctx, SDRB_ERR_ARG, "state blob too small");
    const unsigned char* p = static_cast<const unsigned char*>(host_blob);
    StateBlobHeader h;
    memcpy(&h, p, sizeof h);
    if (h.magic != kBlobMagic || h.version != SDRB_ABI_VERSION || h.n_streams != ctx->n_streams)
        return fail(ctx, SDRB_ERR_ARG, "state blob does not match this context (magic/version/stream count)");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    p += sizeof h;
    CK(cudaMemcpy(ctx->d_fft_state, p, (size_t)ctx->n_streams * sizeof(FftStreamState), cudaMemcpyHostToDevice));
    p += (size_t)ctx->n_streams * sizeof(FftStreamState);
    CK(cudaMemcpy(ctx->d_ssb_state, p, (size_t)ctx->n_streams * ctx->channels * sizeof(SsbStreamState),
                  cudaMemcpyHostToDevice));
    ctx->rfMode = h.rfMode;
    ctx->rfFs = h.rfFs;
    ctx->ssb_samples = h.ssb_samples;
    return upload_ssb(ctx);
}

int sdrb_reset_state(sdrb_ctx* ctx) {
    if (!ctx || !ctx->configured) return fail(ctx, SDRB_ERR_STATE, "not configured");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    int rc = reset_state(ctx);
    if (rc != SDRB_OK) return rc;
    return upload_ssb(ctx);
}

int sdrb_synth_host(int format, uint32_t sample_rate, int first_stream, int n_streams, int64_t first_sample,
                    int64_t n_samples, void* host_out) {
    if (sdrb_format_bytes(format) == 0 || sample_rate == 0 || n_streams < 0 || n_samples < 0 || first_sample < 0 ||
        !host_out)
        return fail(nullptr, SDRB_ERR_ARG, "sdrb_synth_host: bad argument");
    synth_host(format, sample_rate, first_stream, n_streams, first_sample, n_samples, host_out);
    return SDRB_OK;
}

int sdrb_synth_device(sdrb_ctx* ctx, int format, uint32_t sample_rate, int first_stream, int n_streams,
                      int64_t first_sample, int64_t n_samples, void* dev_out) {
    if (!ctx) return SDRB_ERR_ARG;
    if (sdrb_format_bytes(format) == 0 || sample_rate == 0 || n_streams < 0 || n_samples < 0 || first_sample < 0 ||
        !dev_out)
        return fail(ctx, SDRB_ERR_ARG, "sdrb_synth_device: bad argument");
    CK(cudaSetDevice(ctx->device));
    CK(launch_synth(format, sample_rate, first_stream, n_streams, first_sample, n_samples, dev_out, ctx->d_sine,
                    ctx->stream));
    ctx->launches += (uint64_t)((n_streams + 32767) / 32768);
    return SDRB_OK;
}

uint64_t sdrb_launch_count(const sdrb_ctx* ctx) { return ctx ? ctx->launches : 0; }

int sdrb_enable_kernel_timing(sdrb_ctx* ctx, int on) {
    if (!ctx) return SDRB_ERR_ARG;
    ctx->timing = on != 0;
    ctx->ev_valid = false;
    return SDRB_OK;
}

int sdrb_last_fft_kernel_frames(const sdrb_ctx* ctx) { return ctx ? ctx->last_fft_frames : 0; }

int sdrb_last_fft_kernel_ms(sdrb_ctx* ctx, float* ms_total) {
    if (!ctx || !ms_total) return SDRB_ERR_ARG;
    if (!ctx->ev_valid) return fail(ctx, SDRB_ERR_STATE, "kernel timing is off or no FFT stage ran");
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventSynchronize(ctx->ev1));
    CK(cudaEventElapsedTime(ms_total, ctx->ev0, ctx->ev1));
    return ctx->last_fft_launches;
}

}  // extern "C"
