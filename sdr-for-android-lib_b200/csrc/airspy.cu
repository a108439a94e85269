// airspy.cu — Airspy real->IQ front end on the GPU, sm_100a.
// Reference (third_party/libairspy/libairspy/src): convert_samples_float airspy.c:311-321,
// iqconverter_float_process iqconverter_float.c:471-475 = remove_dc (:418-431) + translate_fs_4 (:433-469)
// [fs/4 shift, 47-tap half-band fir_interleaved_24 :287-326 on the even samples, delay_interleaved :391-414
// on the odd samples].  Checker: oracle/sdr_oracle_airspy.c, pinned against the reference's own file.
//
// Three kernels:
//   k_airspy_dc   one thread per (stream, CHUNK of 16384 real samples).  The DC tracker
//                 avg += 0.01f * (x - avg) is a rounded float recurrence: its bits are only reproduced by
//                 walking the samples in order (three dependent operations per sample), which with one thread
//                 per stream leaves the GPU empty (256 streams x 1 M samples: 36 ms).  The recurrence forgets
//                 its start, though: two runs over the same samples from different starts come within an ulp
//                 after ~1700 samples (0.99 per sample) and then land on the SAME float (measured: 1650-2150
//                 samples on noisy ADC data) - and from then on they are the same run.  So every chunk but a
//                 stream's first starts from a GUESS (avg = 0, 4096 samples before the chunk start) and walks
//                 its warm-up + its own samples; the chunk also records the avg it started its own samples
//                 from and the avg it ended with.
//   k_airspy_fix  one thread per stream: chunk c is exact if the avg it started from is bit-identical to the
//                 avg chunk c-1 ended with (chunk 0 starts from the carried state, so by induction every chunk
//                 that passes is the sequential run).  A chunk that fails (it does on noise-free constant
//                 input, where the recurrence stalls inside a ~100-ulp dead band) is walked again from the
//                 right value, in order, and counted.  Worst case = the old one-thread-per-stream cost;
//                 the result is bit-exact in every case.  Also carries avg and the 48-sample tail.
//   k_airspy_fir  fully parallel: one thread per complex output, tile of translated samples staged in shared
//                 memory; twelve un-fused multiply-adds in the reference's left-to-right order (I), a 12-sample
//                 delay (Q).  The 48 translated samples before the call come from the carried state.
// Output is cf32 [streams][blocks][N], which the CF32 FFT / SSB kernels consume unchanged.
#include <cstdlib>

#include "internal.h"

namespace sdrb {

__constant__ float c_hb[12] = {   // even taps 0..22 of HB_KERNEL_FLOAT (filters.h:30-52); taps 24..46 mirror them
    -0.000998606272947510f, 0.001695637278417295f, -0.003054430179754289f, 0.005055504379767936f,
    -0.007901319195893647f, 0.011873357051047719f, -0.017411159379930066f, 0.025304817427568772f,
    -0.037225225204559217f, 0.057533286997004301f, -0.102327462004259350f, 0.317034472508947400f};

namespace {

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

// n samples (a multiple of 8) from `in` (16-byte aligned), in order; translated samples to `out` when WRITE.
// The recurrence is latency-bound on its own (three dependent operations per sample); the loads must not add
// to that chain, so eight 16-byte loads (64 samples) are in flight ahead of the arithmetic.
template <bool WRITE>
__device__ __forceinline__ void air_walk(const uint4* __restrict__ in, float4* __restrict__ out, long long n8, float& avg) {
    constexpr int PF = 8;
    uint4 w[PF];
#pragma unroll
    for (int k = 0; k < PF; k++) w[k] = k < n8 ? __ldg(in + k) : make_uint4(0, 0, 0, 0);
    for (long long i = 0; i < n8; i += PF) {
#pragma unroll
        for (int k = 0; k < PF; k++) {
            if (i + k >= n8) break;
            const uint4 c = w[k];
            if (i + k + PF < n8) w[k] = __ldg(in + i + k + PF);
            float4 a, b;
            a.x = air_step(c.x & 0xffffu, avg, 0); a.y = air_step(c.x >> 16, avg, 1);
            a.z = air_step(c.y & 0xffffu, avg, 2); a.w = air_step(c.y >> 16, avg, 3);
            b.x = air_step(c.z & 0xffffu, avg, 0); b.y = air_step(c.z >> 16, avg, 1);
            b.z = air_step(c.w & 0xffffu, avg, 2); b.w = air_step(c.w >> 16, avg, 3);
            if (WRITE) {
                out[2 * (i + k)] = a;
                out[2 * (i + k) + 1] = b;
            }
        }
    }
}

// T is [n_streams][AIRSPY_TAIL + n_real]: the carried tail, then this call's translated samples.
// from[s][c] = the avg chunk c started its own samples with, upto[s][c] = the avg it ended with.
//
// A lane walks one chunk, so the 32 lanes of a warp sit 32 KiB (input) / 64 KiB (output) apart: direct 16-byte
// loads and stores would touch 32 different lines per instruction.  Instead the warp moves PIECES of 32 samples
// per row: 64 bytes of every row come in with cp.async (4 lanes x 16 B per row, four pieces in flight), a lane walks
// its own row from shared memory, leaves its 32 translated samples in an output tile, and the warp writes the
// tile out 4 rows x 128 contiguous bytes per instruction.  Row pitches (80 B in, 144 B out) keep the 16-byte
// accesses of a quarter-warp on distinct banks.
constexpr int AIRDC_WARPS = 2;
constexpr int AIRDC_IN_PITCH = 80, AIRDC_OUT_PITCH = 144;
constexpr int AIRDC_IN_TILE = 32 * AIRDC_IN_PITCH, AIRDC_OUT_TILE = 32 * AIRDC_OUT_PITCH;
constexpr int AIRDC_STAGES = 4;   // pieces in flight per warp: a 32-sample piece is shorter than a memory round trip
struct __align__(8) AirRow {
    const uint16_t* in;   // first sample the lane walks (run-in included)
    float* out;           // where the chunk's own first translated sample goes
    int total, w;         // samples walked, of which run-in (not written); both multiples of 8, w of 32
};

__global__ void __launch_bounds__(AIRDC_WARPS * 32)
k_airspy_dc(const uint16_t* __restrict__ raw, const AirspyStreamState* __restrict__ state, float* __restrict__ T,
            float* __restrict__ from, float* __restrict__ upto, int n_streams, long long n_real, int n_chunks,
            int chunk, int warm) {
    __shared__ __align__(16) unsigned char s_in[AIRDC_WARPS][AIRDC_STAGES * AIRDC_IN_TILE];
    __shared__ __align__(16) unsigned char s_out[AIRDC_WARPS][AIRDC_OUT_TILE];
    __shared__ AirRow s_rows[AIRDC_WARPS][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long lanes = (long long)n_streams * n_chunks;
    const long long id0 = ((long long)blockIdx.x * AIRDC_WARPS + warp) * 32;
    if (id0 >= lanes) return;
    const bool live = id0 + lane < lanes;
    const long long id = live ? id0 + lane : id0;   // idle lanes shadow lane 0's row (reads only)
    const int s = (int)(id / n_chunks), c = (int)(id - (long long)s * n_chunks);
    float* t = T + (size_t)s * (AIRSPY_TAIL + n_real);
    const uint16_t* x = raw + (size_t)s * n_real;
    const long long pos = (long long)c * chunk;
    const int len = (int)(n_real - pos < chunk ? n_real - pos : chunk);
    const int w = c == 0 ? 0 : (int)(pos < warm ? pos : warm);
    // a run-in that reaches back to the call's first sample starts from the carried avg: not a guess
    float avg = (live && pos <= warm) ? state[s].avg : 0.0f;
    if (c == 0 && live) {
#pragma unroll 4
        for (int k = 0; k < AIRSPY_TAIL; k++) t[k] = state[s].tail[k];
    }
    AirRow me;
    me.in = x + pos - w;
    me.out = t + AIRSPY_TAIL + pos;
    me.total = live ? w + len : 0;
    me.w = w;
    AirRow* rows = s_rows[warp];
    rows[lane] = me;
    int pieces = (me.total + 31) >> 5, P = pieces;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) P = max(P, __shfl_xor_sync(0xffffffffu, P, d));
    __syncwarp();
    unsigned char* tin = s_in[warp];
    unsigned char* tout = s_out[warp];
    const int sub = lane & 3, r0 = lane >> 2;
    auto issue = [&](int p, int stage) {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int r = r0 + 8 * k;
            const AirRow ri = rows[r];
            int off = 32 * p + 8 * sub;                      // samples from ri.in
            if (off > ri.total - 8) off = ri.total - 8;      // past the row's end: re-read its last 16 bytes
            if (off < 0) off = 0;
            const unsigned d = (unsigned)__cvta_generic_to_shared(tin + stage * AIRDC_IN_TILE + r * AIRDC_IN_PITCH + sub * 16);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(ri.in + off) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int k = 0; k < AIRDC_STAGES - 1; k++) {   // one commit group per piece, empty past the end
        if (k < P) issue(k, k);
        else asm volatile("cp.async.commit_group;" ::: "memory");
    }
    for (int p = 0; p < P; p++) {
        if (p + AIRDC_STAGES - 1 < P) issue(p + AIRDC_STAGES - 1, (p + AIRDC_STAGES - 1) % AIRDC_STAGES);
        else asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group %0;" ::"n"(AIRDC_STAGES - 1) : "memory");
        __syncwarp();
        if (32 * p == me.w && live) from[id] = avg;          // the run-in ends on a piece boundary
        const uint4* row = reinterpret_cast<const uint4*>(tin + (p % AIRDC_STAGES) * AIRDC_IN_TILE + lane * AIRDC_IN_PITCH);
        float4* mine = reinterpret_cast<float4*>(tout + lane * AIRDC_OUT_PITCH);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            if (32 * p + 8 * q < me.total) {
                const uint4 v = row[q];
                float4 a, b;
                a.x = air_step(v.x & 0xffffu, avg, 0); a.y = air_step(v.x >> 16, avg, 1);
                a.z = air_step(v.y & 0xffffu, avg, 2); a.w = air_step(v.y >> 16, avg, 3);
                b.x = air_step(v.z & 0xffffu, avg, 0); b.y = air_step(v.z >> 16, avg, 1);
                b.z = air_step(v.w & 0xffffu, avg, 2); b.w = air_step(v.w >> 16, avg, 3);
                mine[2 * q] = a;
                mine[2 * q + 1] = b;
            }
        }
        __syncwarp();
        // the tile out: instruction i covers rows 4i .. 4i+3, 8 lanes x 16 B along each row
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int r = 4 * i + (lane >> 3), col = lane & 7;
            const AirRow ri = rows[r];
            const int at = 32 * p + 4 * col;                 // samples from ri.in
            if (at >= ri.w && at < ri.total)
                *reinterpret_cast<float4*>(ri.out + (at - ri.w)) =
                    *reinterpret_cast<const float4*>(tout + r * AIRDC_OUT_PITCH + col * 16);
        }
        __syncwarp();
    }
    if (live) upto[id] = avg;
}

__global__ void __launch_bounds__(32)
k_airspy_fix(const uint16_t* __restrict__ raw, AirspyStreamState* __restrict__ state, float* __restrict__ T,
             const float* __restrict__ from, float* __restrict__ upto, int n_streams, long long n_real, int n_chunks,
             int chunk) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_streams) return;
    float* t = T + (size_t)s * (AIRSPY_TAIL + n_real);
    const uint16_t* x = raw + (size_t)s * n_real;
    const float* fr = from + (size_t)s * n_chunks;
    float* up = upto + (size_t)s * n_chunks;
    unsigned redone = 0;
    for (int c = 1; c < n_chunks; c++) {
        float avg = up[c - 1];
        if (__float_as_uint(fr[c]) == __float_as_uint(avg)) continue;
        const long long pos = (long long)c * chunk;
        const long long len = n_real - pos < chunk ? n_real - pos : chunk;
        air_walk<true>(reinterpret_cast<const uint4*>(x + pos), reinterpret_cast<float4*>(t + AIRSPY_TAIL + pos), len >> 3, avg);
        up[c] = avg;
        redone++;
    }
    AirspyStreamState* st = state + s;
    st->avg = up[n_chunks - 1];
    st->redone += redone;
    // the new tail: the last AIRSPY_TAIL translated samples (n_real >= AIRSPY_TAIL, or the tail straddles the
    // old one: both are in t[])
#pragma unroll 4
    for (int k = 0; k < AIRSPY_TAIL; k++) st->tail[k] = t[n_real + k];
}

constexpr int AIR_TILE = 1024;      // complex outputs per tile: 256 threads x 4 consecutive outputs
constexpr int AIR_THREADS = 256;
constexpr int AIR_TILES_PER_CTA = 4;
constexpr int AIR_HALF = AIR_TILE + AIRSPY_TAIL / 2;   // even (or odd) translated samples a tile needs: 24 of history

// Output j of a stream = { sum_k h[k] * (E[j-k] + E[j-23+k]), O[j-12] } with E / O the even / odd translated
// samples (E[i] = sample 2i).  A tile de-interleaves its stretch of T into E and O in shared memory on the way
// in (16-byte loads, 8-byte conflict-free stores); a thread then owns 4 consecutive outputs and reads its 28
// even samples as seven 16-byte vectors (a quarter-warp reads 128 contiguous bytes: no bank conflicts) instead
// of 24 scalar loads per output.
__global__ void __launch_bounds__(AIR_THREADS)
k_airspy_fir(const float* __restrict__ T, cpx* __restrict__ iq, long long n_real) {
    __shared__ __align__(16) float E[AIR_HALF];
    __shared__ __align__(16) float O[AIR_HALF];
    const int s = blockIdx.y;
    const long long n_out = n_real >> 1;
    const float* trow = T + (size_t)s * (AIRSPY_TAIL + n_real);   // trow[0] = translated sample -48 = E[-24]
    cpx* orow = iq + (size_t)s * n_out;
    float h[12];
#pragma unroll
    for (int k = 0; k < 12; k++) h[k] = c_hb[k];
    for (int tile = 0; tile < AIR_TILES_PER_CTA; tile++) {
        const long long j0 = ((long long)blockIdx.x * AIR_TILES_PER_CTA + tile) * AIR_TILE;
        if (j0 >= n_out) break;
        // pairs available from trow + 2*j0: 24 of history + the outputs left in the row
        const long long pairs = (n_out - j0 < AIR_TILE ? n_out - j0 : AIR_TILE) + AIRSPY_TAIL / 2;
        const float4* src = reinterpret_cast<const float4*>(trow + 2 * j0);
        if (tile) __syncthreads();   // the previous tile's readers are done
        for (int i = threadIdx.x; 2 * i < AIR_HALF; i += AIR_THREADS) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (2 * i + 1 < pairs) v = __ldg(src + i);                     // pairs is even (n_out - j0 and 24 are)
            *reinterpret_cast<float2*>(E + 2 * i) = make_float2(v.x, v.z);
            *reinterpret_cast<float2*>(O + 2 * i) = make_float2(v.y, v.w);
        }
        __syncthreads();
        const int l = 4 * threadIdx.x;            // local output index of this thread's first output
        if (j0 + l < n_out) {
            // E[l .. l+27] in tile coordinates = even samples j-24 .. j+3 of output j = j0 + l
            float e[28];
#pragma unroll
            for (int q = 0; q < 7; q++) {
                const float4 v = *reinterpret_cast<const float4*>(E + l + 4 * q);
                e[4 * q] = v.x; e[4 * q + 1] = v.y; e[4 * q + 2] = v.z; e[4 * q + 3] = v.w;
            }
            const float4 o = *reinterpret_cast<const float4*>(O + l + 12);   // odd samples j-12 .. j-9
            const float qv[4] = {o.x, o.y, o.z, o.w};
            float2 out[4];
#pragma unroll
            for (int m = 0; m < 4; m++) {
                // output j0+l+m: E[j-k] = e[24+m-k], E[j-23+k] = e[1+m+k]; the reference's left-to-right order
                float acc = __fmul_rn(h[0], __fadd_rn(e[24 + m], e[1 + m]));
#pragma unroll
                for (int k = 1; k < 12; k++) acc = __fadd_rn(acc, __fmul_rn(h[k], __fadd_rn(e[24 + m - k], e[1 + m + k])));
                out[m] = make_float2(acc, qv[m]);
            }
            // n_out is even but not always a multiple of 4: the last thread of a row may own 2 outputs
            float4* dst = reinterpret_cast<float4*>(orow + j0 + l);
            dst[0] = make_float4(out[0].x, out[0].y, out[1].x, out[1].y);
            if (j0 + l + 2 < n_out) dst[1] = make_float4(out[2].x, out[2].y, out[3].x, out[3].y);
        }
    }
}

}  // namespace

// Chunking of the DC pass: `chunk` real samples per lane, `warm` samples of run-in before a guessed start.
// Two environment variables are test hooks (a tiny run-in forces the second walk, a huge chunk gives the plain
// sequential walk); they are read when a context allocates its Airspy buffers.  Without them the chunk follows
// the call: a lane's walk is a latency-bound chain, so the pass wants ~50 k lanes in flight (a dozen warps per
// SM), but every lane also pays the 4096-sample run-in, so chunks stay within [2048, 16384].
constexpr int AIR_CHUNK_MIN = 2048, AIR_CHUNK_MAX = 16384, AIR_WARM = 4096;
void airspy_chunking(int* chunk, int* warm) {
    const char* ce = getenv("SDRB_AIRSPY_CHUNK");
    const char* we = getenv("SDRB_AIRSPY_WARM");
    int c = ce ? atoi(ce) : 0, w = we ? atoi(we) : AIR_WARM;
    if (ce && c < 32) c = 32;
    if (w < 0) w = 0;
    *chunk = c & ~31;   // pieces of 32 samples: chunk starts and the end of the run-in fall on piece boundaries
    *warm = w & ~31;    // chunk 0 = follow the call
}
static int airspy_chunk_for(int fixed, int n_streams, long long n_real) {
    if (fixed > 0) return fixed;
    const long long want = (long long)n_streams * n_real / 49152;
    int c = AIR_CHUNK_MIN;
    while (c < AIR_CHUNK_MAX && 2LL * c <= want) c *= 2;
    return c;
}
size_t airspy_work_floats(int n_streams, long long n_real, int chunk) {
    const int c = chunk > 0 ? chunk : AIR_CHUNK_MIN;
    return 2 * (size_t)n_streams * (size_t)((n_real + c - 1) / c);
}

// raw: [n_streams][n_real] uint16 (n_real = 2 * blocks * N, a multiple of 8); iq: [n_streams][n_real / 2] cf32;
// d_work: airspy_work_floats(n_streams, n_real, chunk) floats
cudaError_t launch_airspy_front(const uint16_t* d_raw, AirspyStreamState* d_state, float* d_T, cpx* d_iq, float* d_work,
                                int chunk, int warm, int n_streams, long long n_real, cudaStream_t st, int* launches) {
    if (n_streams <= 0 || n_real <= 0) return cudaSuccess;
    chunk = airspy_chunk_for(chunk, n_streams, n_real);
    if (n_real % 8 || chunk < 32 || chunk % 32 || warm % 32) return cudaErrorInvalidValue;
    const int n_chunks = (int)((n_real + chunk - 1) / chunk);
    float* from = d_work;
    float* upto = d_work + (size_t)n_streams * n_chunks;
    const long long lanes = (long long)n_streams * n_chunks;
    k_airspy_dc<<<(unsigned)((lanes + 32 * AIRDC_WARPS - 1) / (32 * AIRDC_WARPS)), 32 * AIRDC_WARPS, 0, st>>>(d_raw, d_state, d_T, from, upto, n_streams, n_real, n_chunks,
                                                             chunk, warm);
    k_airspy_fix<<<(n_streams + 31) / 32, 32, 0, st>>>(d_raw, d_state, d_T, from, upto, n_streams, n_real, n_chunks, chunk);
    const long long n_out = n_real / 2;
    const long long per_cta = (long long)AIR_TILE * AIR_TILES_PER_CTA;
    dim3 grid((unsigned)((n_out + per_cta - 1) / per_cta), (unsigned)n_streams);
    k_airspy_fir<<<grid, AIR_THREADS, 0, st>>>(d_T, d_iq, n_real);
    if (launches) *launches += 3;
    return cudaGetLastError();
}

}  // namespace sdrb
