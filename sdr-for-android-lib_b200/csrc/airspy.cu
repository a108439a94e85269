// airspy.cu — Airspy real->IQ front end on the GPU, sm_100a.
// Reference (third_party/libairspy/libairspy/src): convert_samples_float airspy.c:311-321,
// iqconverter_float_process iqconverter_float.c:471-475 = remove_dc (:418-431) + translate_fs_4 (:433-469)
// [fs/4 shift, 47-tap half-band fir_interleaved_24 :287-326 on the even samples, delay_interleaved :391-414
// on the odd samples].  Checker: oracle/sdr_oracle_airspy.c, pinned against the reference's own file.
//
// Two kernels:
//   k_airspy_dc   one thread per stream.  The DC tracker avg += 0.01f * (x - avg) is a rounded float
//                 recurrence; its bits are only reproduced by walking the samples in order (three dependent
//                 operations per sample), so this pass does nothing else: convert, subtract, update, apply
//                 the exact fs/4 signs and halvings, write the translated sample.  16-byte loads (8 samples)
//                 and stores keep every lane on whole sectors.
//   k_airspy_fir  fully parallel: one thread per complex output, tile of translated samples staged in shared
//                 memory; twelve un-fused multiply-adds in the reference's left-to-right order (I), a 12-sample
//                 delay (Q).  The 48 translated samples before the call come from the carried state.
// Output is cf32 [streams][blocks][N], which the CF32 FFT / SSB kernels consume unchanged.
#include "internal.h"

namespace sdrb {

__constant__ float c_hb[12] = {   // even taps 0..22 of HB_KERNEL_FLOAT (filters.h:30-52); taps 24..46 mirror them
    -0.000998606272947510f, 0.001695637278417295f, -0.003054430179754289f, 0.005055504379767936f,
    -0.007901319195893647f, 0.011873357051047719f, -0.017411159379930066f, 0.025304817427568772f,
    -0.037225225204559217f, 0.057533286997004301f, -0.102327462004259350f, 0.317034472508947400f};

namespace {

__device__ __forceinline__ float air_step(unsigned raw, float& avg, int phase) {
    float x = __fmul_rn((float)((int)raw - 2048), 1.0f / 2048.0f);   // airspy.c:316, SAMPLE_SCALE :58-62
    x = __fsub_rn(x, avg);                                           // iqconverter_float.c:425
    avg = __fadd_rn(avg, __fmul_rn(0.01f, x));                       // :426
    // translate_fs_4 (:459-464): -s, -s*hbc, s, s*hbc with hbc = 0.5 (exact)
    return phase == 0 ? -x : phase == 1 ? -0.5f * x : phase == 2 ? x : 0.5f * x;
}

// T is [n_streams][AIRSPY_TAIL + n_real]: the carried tail, then this call's translated samples.
__global__ void __launch_bounds__(32)
k_airspy_dc(const uint16_t* __restrict__ raw, AirspyStreamState* __restrict__ state, float* __restrict__ T,
            int n_streams, long long n_real) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_streams) return;
    AirspyStreamState* st = state + s;
    float* t = T + (size_t)s * (AIRSPY_TAIL + n_real);
#pragma unroll 4
    for (int k = 0; k < AIRSPY_TAIL; k++) t[k] = st->tail[k];
    float avg = st->avg;
    const uint4* in = reinterpret_cast<const uint4*>(raw + (size_t)s * n_real);
    float4* out = reinterpret_cast<float4*>(t + AIRSPY_TAIL);
    const long long n8 = n_real >> 3;
    // the recurrence is latency-bound on its own (three dependent operations per sample); the loads must not add
    // to that chain, so eight 16-byte loads (64 samples) are in flight ahead of the arithmetic
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
            out[2 * (i + k)] = a;
            out[2 * (i + k) + 1] = b;
        }
    }
    st->avg = avg;
    // the new tail: the last AIRSPY_TAIL translated samples (this thread wrote them; n_real >= AIRSPY_TAIL or
    // the tail straddles the old one, both are in t[])
#pragma unroll 4
    for (int k = 0; k < AIRSPY_TAIL; k++) st->tail[k] = t[n_real + k];
}

constexpr int AIR_TILE = 256;   // complex outputs per CTA

__global__ void __launch_bounds__(AIR_TILE)
k_airspy_fir(const float* __restrict__ T, cpx* __restrict__ iq, long long n_real) {
    __shared__ float sh[2 * AIR_TILE + AIRSPY_TAIL];
    const int s = blockIdx.y;
    const long long n_out = n_real >> 1;
    const long long j0 = (long long)blockIdx.x * AIR_TILE;
    const float* t = T + (size_t)s * (AIRSPY_TAIL + n_real) + 2 * j0;   // t[0] = translated sample 2*j0 - 48
    const long long avail = 2 * (n_out - j0) + AIRSPY_TAIL;
    for (int i = threadIdx.x; i < 2 * AIR_TILE + AIRSPY_TAIL; i += AIR_TILE) sh[i] = i < avail ? t[i] : 0.f;
    __syncthreads();
    const long long j = j0 + threadIdx.x;
    if (j >= n_out) return;
    // even sample e[j - k] sits at sh[48 + 2*(tid - k)]
    const float* e = sh + AIRSPY_TAIL + 2 * threadIdx.x;
    float acc = __fmul_rn(c_hb[0], __fadd_rn(e[0], e[-46]));
#pragma unroll
    for (int k = 1; k < 12; k++) acc = __fadd_rn(acc, __fmul_rn(c_hb[k], __fadd_rn(e[-2 * k], e[-46 + 2 * k])));
    // Q: the odd sample delayed by 12 (iqconverter_float.c:399-412): o[j - 12] = sample 2*(j - 12) + 1
    iq[(size_t)s * n_out + j] = mk(acc, e[-23]);
}

}  // namespace

// raw: [n_streams][n_real] uint16 (n_real = 2 * blocks * N, a multiple of 8); iq: [n_streams][n_real / 2] cf32
cudaError_t launch_airspy_front(const uint16_t* d_raw, AirspyStreamState* d_state, float* d_T, cpx* d_iq,
                                int n_streams, long long n_real, cudaStream_t st, int* launches) {
    if (n_streams <= 0 || n_real <= 0) return cudaSuccess;
    if (n_real % 8) return cudaErrorInvalidValue;
    k_airspy_dc<<<(n_streams + 31) / 32, 32, 0, st>>>(d_raw, d_state, d_T, n_streams, n_real);
    const long long n_out = n_real / 2;
    dim3 grid((unsigned)((n_out + AIR_TILE - 1) / AIR_TILE), (unsigned)n_streams);
    k_airspy_fir<<<grid, AIR_TILE, 0, st>>>(d_T, d_iq, n_real);
    if (launches) *launches += 2;
    return cudaGetLastError();
}

}  // namespace sdrb
