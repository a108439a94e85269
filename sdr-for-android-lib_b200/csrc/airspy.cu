// airspy.cu — Airspy real->IQ front end on the GPU, sm_100a.
// Reference (third_party/libairspy/libairspy/src): convert_samples_float airspy.c:311-321,
// iqconverter_float_process iqconverter_float.c:471-475 = remove_dc (:418-431) + translate_fs_4 (:433-469)
// [fs/4 shift, 47-tap half-band fir_interleaved_24 :287-326 on the even samples, delay_interleaved :391-414
// on the odd samples].  Checker: oracle/sdr_oracle_airspy.c, pinned against the reference's own file.
//
// Two kernels (round 2: the half-band FIR moved INTO the walk; the translated scratch, its 8 B written and 8 B
// re-read per complex sample, and the third kernel are gone):
//   k_airspy_front  one lane per (stream, CHUNK of real samples).  The DC tracker avg += 0.01f * (x - avg) is a
//                 rounded float recurrence: its bits are only reproduced by walking the samples in order (three
//                 dependent operations per sample), which with one thread per stream leaves the GPU empty
//                 (256 streams x 1 M samples: 36 ms).  The recurrence forgets its start, though: two runs over
//                 the same samples from different starts come within an ulp after ~1700 samples (0.99 per
//                 sample) and then land on the SAME float (measured: 1650-2150 samples on noisy ADC data) - and
//                 from then on they are the same run.  So every chunk but a stream's first starts from a GUESS
//                 (avg = 0, 4096 samples before the chunk start) and walks its run-in + its own samples.  The
//                 lane keeps the last 24 even / 12 odd translated samples in registers and, for its own samples,
//                 turns every piece of 32 translated samples into 16 complex outputs on the spot: twelve
//                 un-fused products summed in the reference's left-to-right order (I), the 12-sample delay (Q).
//                 The dependent chain of the walk leaves most issue slots of a lane empty; the FIR fills them.
//                 A chunk records the avg it held 48 samples BEFORE its own first sample and 48 samples before
//                 its end.
//   k_airspy_fix  one thread per stream: chunk c is exact if the avg it held 48 samples before its start is
//                 bit-identical to the one chunk c-1 held at the same sample (then both lanes walked the same 48
//                 samples from the same value: same FIR history, same avg at the boundary; chunk 0 starts from
//                 the carried state, so by induction every chunk that passes is the sequential run).  A chunk
//                 that fails (it does on noise-free constant input, where the recurrence stalls inside a
//                 ~100-ulp dead band) is walked again from the right value, in order, and counted.  Worst case =
//                 one thread per stream; the result is bit-exact in every case.  Also carries avg and the
//                 48-sample tail into the next call.
// Output is cf32 [streams][blocks][N], which the CF32 FFT / SSB kernels consume unchanged.
#include <cstdlib>

#include "internal.h"

namespace sdrb {

namespace {

constexpr unsigned AIR_NO_CHECK = 0x7fc00001u;   // "no avg recorded 48 samples before my start": never equals an avg

__device__ __forceinline__ float air_step(unsigned raw, float& avg, int phase) {
    // (raw - 2048) * (1 / 2048) (airspy.c:316, SAMPLE_SCALE :58-62) without the int -> float conversion: the word
    // dropped into the mantissa of 2^23 is the float 2^23 + raw, and (2^23 + raw) / 2048 - 4097 = (raw - 2048) / 2048
    // is exact at every step of one fused multiply-add, so the bits are those of the convert-and-multiply.
    float x = __fmaf_rn(__uint_as_float(0x4B000000u | raw), 1.0f / 2048.0f, -4097.0f);
    x = __fsub_rn(x, avg);                                           // iqconverter_float.c:425
    avg = __fadd_rn(avg, __fmul_rn(0.01f, x));                       // :426
    // translate_fs_4 (:459-464): -s, -s*hbc, s, s*hbc with hbc = 0.5 (exact)
    return phase == 0 ? -x : phase == 1 ? -0.5f * x : phase == 2 ? x : 0.5f * x;
}

// eight samples (one 16-byte word; the first is a multiple of 4 into the stream) -> t[0..7]
__device__ __forceinline__ void air_walk8(const uint4 v, float& avg, float* t) {
    t[0] = air_step(v.x & 0xffffu, avg, 0); t[1] = air_step(v.x >> 16, avg, 1);
    t[2] = air_step(v.y & 0xffffu, avg, 2); t[3] = air_step(v.y >> 16, avg, 3);
    t[4] = air_step(v.z & 0xffffu, avg, 0); t[5] = air_step(v.z >> 16, avg, 1);
    t[6] = air_step(v.w & 0xffffu, avg, 2); t[7] = air_step(v.w >> 16, avg, 3);
}

// The FIR history of a lane: with j0 the first complex output of the coming piece, e[i] = even translated sample
// j0 - 24 + i (sample 2 (j0 - 24 + i) of the stream), o[i] = odd translated sample j0 - 12 + i.
struct AirHist {
    float e[24], o[12];
};

// One piece: 32 translated samples t[] (even E[j0 + i] = t[2 i], odd O[j0 + i] = t[2 i + 1]) -> 16 complex outputs
// out[2 m], out[2 m + 1].  Output j = { sum_k h[k] * (E[j - k] + E[j - 23 + k]), O[j - 12] }: fir_interleaved_24
// (iqconverter_float.c:287-326: twelve products, each of a tap with the SUM of its two mirrored samples, added
// left to right) and delay_interleaved (:391-414).  All indices are literals after unrolling, the taps immediates.
__device__ __forceinline__ void air_fir16(const AirHist& hs, const float* t, float* out) {
    constexpr float h[12] = {   // even taps 0..22 of HB_KERNEL_FLOAT (filters.h:30-52); taps 24..46 mirror them
        -0.000998606272947510f, 0.001695637278417295f, -0.003054430179754289f, 0.005055504379767936f,
        -0.007901319195893647f, 0.011873357051047719f, -0.017411159379930066f, 0.025304817427568772f,
        -0.037225225204559217f, 0.057533286997004301f, -0.102327462004259350f, 0.317034472508947400f};
    float ee[40];
#pragma unroll
    for (int i = 0; i < 24; i++) ee[i] = hs.e[i];
#pragma unroll
    for (int i = 0; i < 16; i++) ee[24 + i] = t[2 * i];
#pragma unroll
    for (int m = 0; m < 16; m++) {
        float acc = __fmul_rn(h[0], __fadd_rn(ee[24 + m], ee[1 + m]));
#pragma unroll
        for (int k = 1; k < 12; k++) acc = __fadd_rn(acc, __fmul_rn(h[k], __fadd_rn(ee[24 + m - k], ee[1 + m + k])));
        out[2 * m] = acc;
        out[2 * m + 1] = m < 12 ? hs.o[m] : t[2 * (m - 12) + 1];
    }
}
// the history 16 outputs later
__device__ __forceinline__ void air_shift(AirHist& hs, const float* t) {
#pragma unroll
    for (int i = 0; i < 8; i++) hs.e[i] = hs.e[16 + i];
#pragma unroll
    for (int i = 0; i < 16; i++) hs.e[8 + i] = t[2 * i];
#pragma unroll
    for (int i = 0; i < 12; i++) hs.o[i] = t[2 * (4 + i) + 1];
}
// from the 48 translated samples before an output (oldest first): a carried tail, or a walk of 48 samples
__device__ __forceinline__ void air_hist_from(AirHist& hs, const float* t48) {
#pragma unroll
    for (int i = 0; i < 24; i++) hs.e[i] = t48[2 * i];
#pragma unroll
    for (int i = 0; i < 12; i++) hs.o[i] = t48[2 * (12 + i) + 1];
}

// The 32 lanes of a warp walk the SAME chunk of 32 different streams, so everything about the walk but the data is
// warp-uniform (run-in, length, piece count, the marks) and every address is affine in the row.  Lanes sit a whole
// stream apart in the input and in the output: direct 16-byte loads and stores would touch 32 different lines per
// instruction.  Instead the warp moves pieces per row: 128 bytes (one line, two pieces of 32 samples) of every row
// come in with cp.async (8 lanes x 16 B per row, three loads in flight), a lane walks its own row from shared memory
// and leaves the 16 complex outputs (128 bytes) of every piece in an output tile, which the warp writes out 4 rows x
// 128 contiguous bytes per instruction.  The row pitch (144 B in and out) keeps the 16-byte accesses of a
// quarter-warp on distinct banks.
// Lanes past the last stream shadow it (reads only).
constexpr int AIRDC_WARPS = 2;
#ifndef AIRDC_SUPER
#define AIRDC_SUPER 2
#endif
constexpr int AIRDC_SUP = AIRDC_SUPER;                        // 32-sample pieces per load: 2 = whole 128-byte lines
constexpr int AIRDC_IN_PITCH = 64 * AIRDC_SUP + 16, AIRDC_OUT_PITCH = 144;
constexpr int AIRDC_IN_TILE = 32 * AIRDC_IN_PITCH, AIRDC_OUT_TILE = 32 * AIRDC_OUT_PITCH;
constexpr int AIRDC_STAGES = AIRDC_SUP == 1 ? 4 : 3;          // loads in flight per warp
constexpr int AIRDC_LPR = 4 * AIRDC_SUP;                      // lanes (16 bytes each) along a row of a load

// chunk c of a stream = real samples [c * chunk, (c + 1) * chunk), the last one through the end of the call
// (so every chunk but a lone one is at least `chunk` >= 64 samples long).
// from48[s][c] = the avg the lane held 48 samples before its own first sample (AIR_NO_CHECK if its run-in is
// shorter than that), upto48[s][c] = the avg it held 48 samples before its end.
#ifndef AIRDC_MINB
#define AIRDC_MINB 6   // 6 CTAs per SM: what the shared memory allows
#endif
__global__ void __launch_bounds__(AIRDC_WARPS * 32, AIRDC_MINB)
k_airspy_front(const uint16_t* __restrict__ raw, const AirspyStreamState* __restrict__ state, float* __restrict__ iq,
               float* __restrict__ from48, float* __restrict__ upto48, int n_streams, long long n_real, int n_chunks,
               int chunk, int warm) {
    __shared__ __align__(16) unsigned char s_in[AIRDC_WARPS][AIRDC_STAGES * AIRDC_IN_TILE];
    __shared__ __align__(16) unsigned char s_out[AIRDC_WARPS][AIRDC_OUT_TILE];
    // read once: under register pressure the compiler otherwise re-reads the thread id
    // (S2R, tens of cycles on the short scoreboard) inside the piece loop for every lane-derived address
    const unsigned tid = __shfl_sync(0xffffffffu, threadIdx.x, threadIdx.x & 31);   // (a shuffle is not re-derived)
    const int lane = tid & 31, warp = tid >> 5;
    const int groups = (n_streams + 31) >> 5;
    const long long task = (long long)blockIdx.x * AIRDC_WARPS + warp;     // (stream group, chunk), chunk fastest
    if (task >= (long long)groups * n_chunks) return;
    const int g = (int)(task / n_chunks), c = (int)(task - (long long)g * n_chunks);
    const int rows_live = min(32, n_streams - 32 * g);
    const int s = 32 * g + min(lane, rows_live - 1);
    // ---- warp-uniform geometry
    const long long pos = (long long)c * chunk;
    const int len = (int)(c == n_chunks - 1 ? n_real - pos : chunk);
    const int w = c == 0 ? 0 : (int)(pos < warm ? pos : warm);
    const int total = w + len, P = (total + 31) >> 5, p_own = w >> 5;
    const bool from_state = pos == w;   // a walk that starts on the call's first sample starts from the carried state
    const int g_from = (c > 0 && w >= AIRSPY_TAIL) ? (w - AIRSPY_TAIL) >> 3 : -1;        // in groups of 8 samples
    const int g_upto = total >= AIRSPY_TAIL ? (total - AIRSPY_TAIL) >> 3 : -1;
    const size_t rec = (size_t)s * n_chunks + c;
    if (c > 0 && g_from < 0) from48[rec] = __uint_as_float(AIR_NO_CHECK);

    float avg = from_state ? state[s].avg : 0.0f;
    AirHist hs;
    if (from_state) {
        float t48[AIRSPY_TAIL];
#pragma unroll
        for (int k = 0; k < AIRSPY_TAIL; k++) t48[k] = state[s].tail[k];
        air_hist_from(hs, t48);
    } else {
#pragma unroll
        for (int k = 0; k < 24; k++) hs.e[k] = 0.0f;
#pragma unroll
        for (int k = 0; k < 12; k++) hs.o[k] = 0.0f;
    }
    unsigned char* tin = s_in[warp];
    unsigned char* tout = s_out[warp];
    // loads: lane = (row r0 + (32 / LPR) k, 16-byte part `sub`) of the load piece
    const int sub = lane % AIRDC_LPR, r0 = lane / AIRDC_LPR;
    const uint16_t* const src0 = raw + (size_t)(32 * g) * n_real + (pos - w);
    const unsigned dst0 = (unsigned)__cvta_generic_to_shared(tin + r0 * AIRDC_IN_PITCH + sub * 16);
    const int L = (total + 32 * AIRDC_SUP - 1) / (32 * AIRDC_SUP);   // load pieces
    auto issue = [&](int l, int stage) {
        int off = 32 * AIRDC_SUP * l + 8 * sub;              // samples from the row's first
        off = min(off, total - 8);                           // past the row's end: re-read its last 16 bytes
#pragma unroll
        for (int k = 0; k < AIRDC_LPR; k++) {
            const int r = min(r0 + (32 / AIRDC_LPR) * k, rows_live - 1);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst0 + stage * AIRDC_IN_TILE +
                                                                            (32 / AIRDC_LPR) * k * AIRDC_IN_PITCH),
                         "l"(src0 + (size_t)r * n_real + off) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    // stores: instruction i covers rows 4i .. 4i+3, 8 lanes x 16 B along each row
    const int orow = lane >> 3, col = lane & 7;
    float* const out0 = iq + (size_t)(32 * g + orow) * n_real + pos + 4 * col;
    const size_t out_step = 4 * (size_t)n_real;              // floats between the rows of consecutive instructions
    const unsigned char* const tsrc = tout + orow * AIRDC_OUT_PITCH + col * 16;

    for (int k = 0; k < AIRDC_STAGES - 1; k++) {   // one commit group per load, empty past the end
        if (k < L) issue(k, k);
        else asm volatile("cp.async.commit_group;" ::: "memory");
    }
    for (int p = 0; p < P; p++) {
        const int l = p / AIRDC_SUP, h = p % AIRDC_SUP;
        if (h == 0) {
            if (l + AIRDC_STAGES - 1 < L) issue(l + AIRDC_STAGES - 1, (l + AIRDC_STAGES - 1) % AIRDC_STAGES);
            else asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group %0;" ::"n"(AIRDC_STAGES - 1) : "memory");
            __syncwarp();
        }
        const uint4* row = reinterpret_cast<const uint4*>(tin + (l % AIRDC_STAGES) * AIRDC_IN_TILE + lane * AIRDC_IN_PITCH + 64 * h);
        // past the row's end the walk goes on over re-read bytes: nothing it produces there is kept
        float t[32], aq[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            aq[q] = avg;
            air_walk8(row[q], avg, t + 8 * q);
        }
        // (lanes past the last stream walk its data and write its records: same values to the same addresses)
        if (p == (g_from >> 2)) from48[rec] = aq[2];           // g_from = 4 (w / 32) - 6: always the third group
        if (p == (g_upto >> 2)) {
            const int q = g_upto & 3;
            upto48[rec] = q == 0 ? aq[0] : q == 1 ? aq[1] : q == 2 ? aq[2] : aq[3];
        }
        if (p >= p_own) {                                      // the chunk's own samples: 16 complex outputs
            float o[32];
            air_fir16(hs, t, o);
            float4* mine = reinterpret_cast<float4*>(tout + lane * AIRDC_OUT_PITCH);
#pragma unroll
            for (int q = 0; q < 8; q++) mine[q] = make_float4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
        }
        if (p + 2 >= p_own) air_shift(hs, t);                  // the history matters from 48 samples before the own ones
        if (p >= p_own) {
            __syncwarp();
            float* dst = out0 + 32 * (p - p_own);
            if (32 * p + 4 * col < total) {
                if (rows_live == 32) {
#pragma unroll
                    for (int i = 0; i < 8; i++)
                        *reinterpret_cast<float4*>(dst + i * out_step) =
                            *reinterpret_cast<const float4*>(tsrc + 4 * i * AIRDC_OUT_PITCH);
                } else {
#pragma unroll 1
                    for (int i = 0; 4 * i + orow < rows_live; i++)
                        *reinterpret_cast<float4*>(dst + i * out_step) =
                            *reinterpret_cast<const float4*>(tsrc + 4 * i * AIRDC_OUT_PITCH);
                }
            }
        }
        __syncwarp();
    }
}

// 48 samples walked in order from `in` (16-byte aligned, a multiple of 4 into the stream) -> t48[]
__device__ __forceinline__ void air_walk48(const uint16_t* in, float& avg, float* t48) {
#pragma unroll
    for (int q = 0; q < 6; q++) air_walk8(__ldg(reinterpret_cast<const uint4*>(in) + q), avg, t48 + 8 * q);
}

__global__ void __launch_bounds__(32)
k_airspy_fix(const uint16_t* __restrict__ raw, AirspyStreamState* __restrict__ state, float* __restrict__ iq,
             const float* __restrict__ from48, float* __restrict__ upto48, int n_streams, long long n_real, int n_chunks,
             int chunk) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_streams) return;
    const uint16_t* x = raw + (size_t)s * n_real;
    float* orow = iq + (size_t)s * n_real;
    const float* fr = from48 + (size_t)s * n_chunks;
    float* up = upto48 + (size_t)s * n_chunks;
    AirspyStreamState* st = state + s;
    unsigned redone = 0;
    for (int c = 1; c < n_chunks; c++) {
        float avg = up[c - 1];                                   // the avg at sample pos - 48, exact by induction
        if (__float_as_uint(fr[c]) == __float_as_uint(avg)) continue;
        const long long pos = (long long)c * chunk;
        const long long len = c == n_chunks - 1 ? n_real - pos : chunk;
        AirHist hs;
        {
            float t48[AIRSPY_TAIL];
            air_walk48(x + pos - AIRSPY_TAIL, avg, t48);
            air_hist_from(hs, t48);
        }
        for (long long i = 0; i < len; i += 32) {
            float t[32], o[32];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const long long at = i + 8 * q;
                if (at < len) {
                    if (at == len - AIRSPY_TAIL) up[c] = avg;
                    air_walk8(__ldg(reinterpret_cast<const uint4*>(x + pos + at)), avg, t + 8 * q);
                } else {
#pragma unroll
                    for (int k = 0; k < 8; k++) t[8 * q + k] = 0.0f;
                }
            }
            air_fir16(hs, t, o);
            air_shift(hs, t);
#pragma unroll
            for (int q = 0; q < 8; q++)
                if (i + 4 * q < len)
                    *reinterpret_cast<float4*>(orow + pos + i + 4 * q) = make_float4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
        }
        redone++;
    }
    // carry: the avg after the call's last sample and the last AIRSPY_TAIL translated samples
    if (n_real >= AIRSPY_TAIL) {
        float avg = up[n_chunks - 1];
        float t48[AIRSPY_TAIL];
        air_walk48(x + n_real - AIRSPY_TAIL, avg, t48);
        st->avg = avg;
#pragma unroll
        for (int k = 0; k < AIRSPY_TAIL; k++) st->tail[k] = t48[k];
    } else {
        // a call shorter than the tail (a lone chunk, walked from the carried state): the old tail slides
        float avg = st->avg;
        const int n = (int)n_real;
        for (int k = 0; k + n < AIRSPY_TAIL; k++) st->tail[k] = st->tail[k + n];
        for (int i = 0; i < n; i += 8) {
            float t[8];
            air_walk8(__ldg(reinterpret_cast<const uint4*>(x + i)), avg, t);
            for (int k = 0; k < 8; k++) st->tail[AIRSPY_TAIL - n + i + k] = t[k];
        }
        st->avg = avg;
    }
    st->redone += redone;
}

}  // namespace

// Chunking of the walk: `chunk` real samples per lane, `warm` samples of run-in before a guessed start.
// Two environment variables are test hooks (a tiny run-in forces the second walk, a huge chunk gives the plain
// sequential walk); they are read when a context allocates its Airspy buffers.  Without them the chunk follows
// the call: the smallest power of two in [2048, 16384] whose warps (one per 32 streams and chunk) are all resident
// at once on a B200 (148 SMs x 6 CTAs x 2 warps) - every lane pays the 4096-sample run-in, so a smaller chunk only
// pays while it fills empty SMs (measured, 256 streams x 1 M samples: 8192 -> 112 GS/s, 4096 -> 103, 16384 -> 96).
constexpr int AIR_CHUNK_MIN = 2048, AIR_CHUNK_MAX = 16384, AIR_WARM = 4096;
constexpr long long AIR_RESIDENT_WARPS = 148 * AIRDC_MINB * AIRDC_WARPS;
void airspy_chunking(int* chunk, int* warm) {
    const char* ce = getenv("SDRB_AIRSPY_CHUNK");
    const char* we = getenv("SDRB_AIRSPY_WARM");
    int c = ce ? atoi(ce) : 0, w = we ? atoi(we) : AIR_WARM;
    if (ce && c < 64) c = 64;   // a chunk holds the 48 samples its successor is checked against
    if (w < 0) w = 0;
    *chunk = c & ~31;   // pieces of 32 samples: chunk starts and the end of the run-in fall on piece boundaries
    *warm = w & ~31;    // chunk 0 = follow the call
}
static int airspy_chunk_for(int fixed, int n_streams, long long n_real) {
    if (fixed > 0) return fixed;
    const long long groups = (n_streams + 31) / 32;
    int c = AIR_CHUNK_MIN;
    while (c < AIR_CHUNK_MAX && groups * (n_real / c) > AIR_RESIDENT_WARPS) c *= 2;
    return c;
}
size_t airspy_work_floats(int n_streams, long long n_real, int chunk) {
    const int c = chunk > 0 ? chunk : AIR_CHUNK_MIN;
    return 2 * (size_t)n_streams * (size_t)((n_real + c - 1) / c);
}

// raw: [n_streams][n_real] uint16 (n_real = 2 * blocks * N, a multiple of 8); iq: [n_streams][n_real / 2] cf32;
// d_work: airspy_work_floats(n_streams, n_real, chunk) floats
cudaError_t launch_airspy_front(const uint16_t* d_raw, AirspyStreamState* d_state, cpx* d_iq, float* d_work,
                                int chunk, int warm, int n_streams, long long n_real, cudaStream_t st, int* launches) {
    if (n_streams <= 0 || n_real <= 0) return cudaSuccess;
    chunk = airspy_chunk_for(chunk, n_streams, n_real);
    if (n_real % 8 || chunk < 64 || chunk % 32 || warm % 32) return cudaErrorInvalidValue;
    const long long whole = n_real / chunk;
    const int n_chunks = (int)(whole > 0 ? whole : 1);   // the last chunk takes the remainder
    float* from48 = d_work;
    float* upto48 = d_work + (size_t)n_streams * n_chunks;
    const long long tasks = (long long)((n_streams + 31) / 32) * n_chunks;   // one warp per (32 streams, chunk)
    float* out = reinterpret_cast<float*>(d_iq);
    k_airspy_front<<<(unsigned)((tasks + AIRDC_WARPS - 1) / AIRDC_WARPS), 32 * AIRDC_WARPS, 0, st>>>(
        d_raw, d_state, out, from48, upto48, n_streams, n_real, n_chunks, chunk, warm);
    k_airspy_fix<<<(n_streams + 31) / 32, 32, 0, st>>>(d_raw, d_state, out, from48, upto48, n_streams, n_real, n_chunks, chunk);
    if (launches) *launches += 2;
    return cudaGetLastError();
}

}  // namespace sdrb
