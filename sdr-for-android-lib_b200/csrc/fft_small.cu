// fft_small.cu — one-CTA-per-transform shared-memory FFT kernels for 256 .. 8192 points (sm_100a).
// See fft_small.cuh for the decomposition and the phase functions.  Ingest convert, |X|^2 and fftshift are fused as
// in the 16K kernel (reference: src/dsp/fft_process.cpp:62-97); before this kernel these lengths went through the
// radix-16 Stockham passes over global memory (16 B per sample and pass: 59-74 GS/s against 372 at 16384 points).
//
//   k_fft_small<FMT, C> (256 .. 4096 points), k_fft_8k<FMT> (8192): persistent, 256 threads, 256 / (16 C) transforms per CTA
//   and round;
//   raw IQ straight into registers (a warp reads 32 consecutive samples per load), two in-place exchanges through
//   shared memory (16 C + 1 complex per plane), twiddle tables in shared memory, spectra out 128 bytes per warp store.
#include <cstdlib>

#include "fft_small.cuh"
#include "internal.h"

namespace sdrb {

// Extension front end between load and transform (no reference code: FFTProcessor::process feeds the raw block to FFTW,
// src/dsp/fft_process.cpp:62-79; definition: oracle/sdr_oracle_ext.c).  front bit 1: subtract the block mean (per-thread
// sum, shuffle tree over the transform's lanes of a warp, then over its warps through shared memory); bit 0: multiply by
// the window table.  P points per thread at stride T; a transform owns T consecutive threads (T = 16 shares a warp with
// another transform: the shuffle tree stays inside its 16 lanes).  Every thread of the CTA calls it (barrier inside).
template <int P, int T>
__device__ __forceinline__ void fs_front(int front, int t, int sub, cpx* v, const float* __restrict__ win, cpx* red) {
    if (front & 2) {
        cpx s = mk(0.f, 0.f);
#pragma unroll
        for (int a = 0; a < P; a++) s = cadd(s, v[a]);
#pragma unroll
        for (int d = (T < 32 ? T : 32) / 2; d > 0; d >>= 1) {
            s.x += __shfl_xor_sync(0xffffffffu, s.x, d);
            s.y += __shfl_xor_sync(0xffffffffu, s.y, d);
        }
        constexpr int W = T > 32 ? T / 32 : 1;               // warps per transform
        if (T > 32) {
            if ((t & 31) == 0) red[sub * W + (t >> 5)] = s;
            __syncthreads();
            s = mk(0.f, 0.f);
#pragma unroll
            for (int w = 0; w < W; w++) s = cadd(s, red[sub * W + w]);
            __syncthreads();                                 // red is reused by the next round
        }
        const cpx m = mk(s.x * (1.0f / (float)(P * T)), s.y * (1.0f / (float)(P * T)));
#pragma unroll
        for (int a = 0; a < P; a++) v[a] = csub(v[a], m);
    }
    if (front & 1) {
#pragma unroll
        for (int a = 0; a < P; a++) {
            const float w = __ldg(win + T * a + t);
            v[a].x *= w;
            v[a].y *= w;
        }
    }
}

// ROWS = 0: n_frames independent N-point frames.  ROWS = A > 0 (the rows pass of the two-pass scheme): n_frames counts
// ROWS, row r = (big frame r / A, ka = r % A), spectra interleaved into the big frame.
template <int FMT, int C, bool STRIDED, bool FRONT>
__global__ void __launch_bounds__(FsGeom<C>::kThreads)
k_fft_small(const typename Ingest2<FMT>::raw_t* __restrict__ in, float* __restrict__ out, const cpx* __restrict__ tables,
            int n_frames, int rows_per_frame, int front, const float* __restrict__ win) {
    typedef FsGeom<C> G;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ cpx red[8];
    cpx* Eall = reinterpret_cast<cpx*>(smem_raw);
    cpx* TP = Eall + G::kPerCta * G::kE;
    cpx* T2 = TP + G::kTP;
    for (int i = threadIdx.x; i < G::kTables; i += G::kThreads) TP[i] = tables[i];
    const int sub = threadIdx.x / G::T, t = threadIdx.x % G::T;
    cpx* E = Eall + sub * G::kE;
    __syncthreads();
    for (long long f0 = (long long)blockIdx.x * G::kPerCta; f0 < n_frames; f0 += (long long)gridDim.x * G::kPerCta) {
        const long long f = f0 + sub;
        const bool live = f < n_frames;
        if (FRONT) {
            cpx v[16];
            if (live) fs_phase1_load<FMT, C>(t, in + (size_t)f * G::N, v);
            else {
#pragma unroll
                for (int a = 0; a < 16; a++) v[a] = mk(0.f, 0.f);
            }
            fs_front<16, G::T>(front, t, sub, v, win, red);
            if (live) fs_phase1_finish<C>(t, v, TP, E);
        } else if (live) {
            fs_phase1<FMT, C>(t, in + (size_t)f * G::N, TP, E);
        }
        __syncthreads();
        if (live) fs_phase2<C>(t, E);
        __syncthreads();
        if (live) {
            if (STRIDED) {
                const long long big = f / rows_per_frame;
                const FsOutStrided where = {rows_per_frame, (int)(f - big * rows_per_frame)};
                fs_phase3<C>(t, E, T2, out + (size_t)big * rows_per_frame * G::N, where);
            } else {
                fs_phase3<C>(t, E, T2, out + (size_t)f * G::N, FsOutPlain());
            }
        }
        __syncthreads();   // the next round's phase 1 overwrites E
    }
}

// 8192 points: the 32 * 16 * 16 shape (fft_small.cuh: Fs8k), 256 threads x 32 points, every thread busy in every phase
template <int FMT, bool STRIDED, bool FRONT>
__global__ void __launch_bounds__(256, 2)
k_fft_8k(const typename Ingest2<FMT>::raw_t* __restrict__ in, float* __restrict__ out, const cpx* __restrict__ tables,
         int n_frames, int rows_per_frame, int front, const float* __restrict__ win) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ cpx red[8];
    cpx* E = reinterpret_cast<cpx*>(smem_raw);
    cpx* TP = E + Fs8k::kE;
    cpx* T2 = TP + Fs8k::kTP;
    const int t = threadIdx.x;
    for (int i = t; i < Fs8k::kTables; i += 256) TP[i] = tables[i];
    __syncthreads();
    for (long long f = blockIdx.x; f < n_frames; f += gridDim.x) {
        if (FRONT) {
            cpx v[32];
            fs8k_phase1_load<FMT>(t, in + (size_t)f * Fs8k::N, v);
            fs_front<32, 256>(front, t, 0, v, win, red);
            fs8k_phase1_finish(t, v, TP, E);
        } else {
            fs8k_phase1<FMT>(t, in + (size_t)f * Fs8k::N, TP, E);
        }
        __syncthreads();
        fs8k_phase2(t, E);
        __syncthreads();
        if (STRIDED) {
            const long long big = f / rows_per_frame;
            const FsOutStrided where = {rows_per_frame, (int)(f - big * rows_per_frame)};
            fs8k_phase3(t, E, T2, out + (size_t)big * rows_per_frame * Fs8k::N, where);
        } else {
            fs8k_phase3(t, E, T2, out + (size_t)f * Fs8k::N, FsOutPlain());
        }
        __syncthreads();
    }
}
constexpr int kSmem8k = (Fs8k::kE + Fs8k::kTables) * (int)sizeof(cpx);
template <int FMT, bool STRIDED>
static cudaError_t launch_8k(const void* iq, float* power, const cpx* tables, int n_frames, int rows_per_frame, int sm_count,
                             cudaStream_t st, int front = 0, const float* win = nullptr) {
    static int per_sm = 0;
    if (!per_sm && (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_fft_8k<FMT, STRIDED, false>, 256, kSmem8k) != cudaSuccess || per_sm < 1))
        per_sm = 1;
    const long long slots = (long long)per_sm * (sm_count > 0 ? sm_count : 148);
    const int grid = (int)(n_frames < slots ? n_frames : slots);
    const typename Ingest2<FMT>::raw_t* src = static_cast<const typename Ingest2<FMT>::raw_t*>(iq);
    if (front && !STRIDED) k_fft_8k<FMT, false, true><<<grid, 256, kSmem8k, st>>>(src, power, tables, n_frames, rows_per_frame, front, win);
    else k_fft_8k<FMT, STRIDED, false><<<grid, 256, kSmem8k, st>>>(src, power, tables, n_frames, rows_per_frame, 0, nullptr);
    return cudaGetLastError();
}

// Rows pass with FOUR rows of a big frame per CTA (4096-point rows, A a multiple of 4): the powers of the four
// spectra are gathered in shared memory and bin kb of rows 4g .. 4g + 3 goes out as ONE 16-byte store at 4g + A kb.
// With one row per CTA (k_fft_small<.., true>) every float of a warp store lands in its own sector.
__global__ void __launch_bounds__(1024, 1)
k_fft_rows4(const cpx* __restrict__ Y, float* __restrict__ out, const cpx* __restrict__ tables, int n_groups, int A) {
    typedef FsGeom<16> G;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cpx* Eall = reinterpret_cast<cpx*>(smem_raw);
    cpx* TP = Eall + 4 * G::kE;
    cpx* T2 = TP + G::kTP;
    float* S = reinterpret_cast<float*>(smem_raw);          // [4][4096] powers, over the exchange buffers
    for (int i = threadIdx.x; i < G::kTables; i += 1024) TP[i] = tables[i];
    const int sub = threadIdx.x >> 8, t = threadIdx.x & 255;
    cpx* E = Eall + sub * G::kE;
    const int N = A * G::N;
    __syncthreads();
    for (int grp = blockIdx.x; grp < n_groups; grp += gridDim.x) {
        fs_phase1<FMT_CF32, 16>(t, Y + ((size_t)4 * grp + sub) * G::N, TP, E);
        __syncthreads();
        fs_phase2<16>(t, E);
        __syncthreads();
        float pw[16];
        fs_phase3<16>(t, E, T2, pw, FsOutKc());
        __syncthreads();                                    // every thread holds its bins: the buffers are free
#pragma unroll
        for (int kc = 0; kc < 16; kc++) S[sub * G::N + t + 256 * kc] = pw[kc];
        __syncthreads();
        const long long row0 = (long long)4 * grp;
        const long long big = row0 / A;
        const int ka0 = (int)(row0 - big * A);
        float* o = out + (size_t)big * N;
#pragma unroll
        for (int it = 0; it < 4; it++) {
            const int k = it * 1024 + threadIdx.x;
            const float4 v = make_float4(S[k], S[G::N + k], S[2 * G::N + k], S[3 * G::N + k]);
            *reinterpret_cast<float4*>(o + ((ka0 + A * k + N / 2) & (N - 1))) = v;
        }
        __syncthreads();                                    // the next group's phase 1 overwrites S
    }
}
constexpr int kSmemRows4 = (4 * FsGeom<16>::kE + FsGeom<16>::kTables) * (int)sizeof(cpx);

template <int C>
static constexpr int fs_smem() { return (FsGeom<C>::kPerCta * FsGeom<C>::kE + FsGeom<C>::kTables) * (int)sizeof(cpx); }

template <int C>
static cudaError_t fs_prepare_c() {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_fft_small<FMT_CU8, C, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fs_smem<C>())) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fft_small<FMT_CS16, C, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fs_smem<C>())) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fft_small<FMT_CF32, C, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fs_smem<C>())) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fft_small<FMT_CU8, C, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, fs_smem<C>())) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fft_small<FMT_CS16, C, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, fs_smem<C>())) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(k_fft_small<FMT_CF32, C, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, fs_smem<C>())) != cudaSuccess) return e;
    if (C == 16) return cudaFuncSetAttribute(k_fft_small<FMT_CF32, C, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fs_smem<C>());
    return cudaSuccess;
}
cudaError_t fft_small_prepare() {
    cudaError_t e;
    if ((e = fs_prepare_c<1>()) != cudaSuccess) return e;
    if ((e = fs_prepare_c<2>()) != cudaSuccess) return e;
    if ((e = fs_prepare_c<4>()) != cudaSuccess) return e;
    if ((e = fs_prepare_c<8>()) != cudaSuccess) return e;
    if ((e = fs_prepare_c<16>()) != cudaSuccess) return e;
#define SDRB_8K_ATTR(F, S, R)                                                                                          \
    if ((e = cudaFuncSetAttribute(k_fft_8k<F, S, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem8k)) != cudaSuccess) return e;
    SDRB_8K_ATTR(FMT_CU8, false, false) SDRB_8K_ATTR(FMT_CS16, false, false) SDRB_8K_ATTR(FMT_CF32, false, false)
    SDRB_8K_ATTR(FMT_CU8, false, true) SDRB_8K_ATTR(FMT_CS16, false, true) SDRB_8K_ATTR(FMT_CF32, false, true)
    SDRB_8K_ATTR(FMT_CF32, true, false)
#undef SDRB_8K_ATTR
    return cudaFuncSetAttribute(k_fft_rows4, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemRows4);
}

bool fft_small_supports(int n) { return n >= 256 && n <= 8192 && (n & (n - 1)) == 0; }
int fft_small_table_elems(int n) { return !fft_small_supports(n) ? 0 : (n == 8192 ? Fs8k::kTables : 80 * (n / 256)); }

template <int FMT, int C>
static cudaError_t fs_launch_fc(const void* iq, float* power, const cpx* tables, int n_frames, int sm_count, cudaStream_t st,
                                int front, const float* win) {
    typedef FsGeom<C> G;
    static int per_sm = 0;
    if (!per_sm && (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_fft_small<FMT, C, false, false>, G::kThreads, fs_smem<C>()) != cudaSuccess || per_sm < 1))
        per_sm = 1;
    const long long rounds = ((long long)n_frames + G::kPerCta - 1) / G::kPerCta;
    const long long slots = (long long)per_sm * (sm_count > 0 ? sm_count : 148);
    const int grid = (int)(rounds < slots ? rounds : slots);
    const typename Ingest2<FMT>::raw_t* src = static_cast<const typename Ingest2<FMT>::raw_t*>(iq);
    if (front) k_fft_small<FMT, C, false, true><<<grid, G::kThreads, fs_smem<C>(), st>>>(src, power, tables, n_frames, 1, front, win);
    else k_fft_small<FMT, C, false, false><<<grid, G::kThreads, fs_smem<C>(), st>>>(src, power, tables, n_frames, 1, 0, nullptr);
    return cudaGetLastError();
}
template <int C>
static cudaError_t fs_launch_c(int fmt, const void* iq, float* power, const cpx* tables, int n_frames, int sm_count,
                               cudaStream_t st, int front, const float* win) {
    switch (fmt) {
        case FMT_CU8: return fs_launch_fc<FMT_CU8, C>(iq, power, tables, n_frames, sm_count, st, front, win);
        case FMT_CS16: return fs_launch_fc<FMT_CS16, C>(iq, power, tables, n_frames, sm_count, st, front, win);
        case FMT_CF32: return fs_launch_fc<FMT_CF32, C>(iq, power, tables, n_frames, sm_count, st, front, win);
        default: return cudaErrorInvalidValue;
    }
}
cudaError_t launch_fft_small(int fmt, const void* iq, float* power, const cpx* tables, int n, int n_frames, int sm_count,
                             cudaStream_t st, int front, const float* win) {
    if (n_frames <= 0) return cudaSuccess;
    if ((front & 1) && !win) return cudaErrorInvalidValue;
    switch (n) {
        case 256: return fs_launch_c<1>(fmt, iq, power, tables, n_frames, sm_count, st, front, win);
        case 512: return fs_launch_c<2>(fmt, iq, power, tables, n_frames, sm_count, st, front, win);
        case 1024: return fs_launch_c<4>(fmt, iq, power, tables, n_frames, sm_count, st, front, win);
        case 2048: return fs_launch_c<8>(fmt, iq, power, tables, n_frames, sm_count, st, front, win);
        case 4096: return fs_launch_c<16>(fmt, iq, power, tables, n_frames, sm_count, st, front, win);
        case 8192:
            switch (fmt) {
                case FMT_CU8: return launch_8k<FMT_CU8, false>(iq, power, tables, n_frames, 1, sm_count, st, front, win);
                case FMT_CS16: return launch_8k<FMT_CS16, false>(iq, power, tables, n_frames, 1, sm_count, st, front, win);
                case FMT_CF32: return launch_8k<FMT_CF32, false>(iq, power, tables, n_frames, 1, sm_count, st, front, win);
                default: return cudaErrorInvalidValue;
            }
        default: return cudaErrorInvalidValue;
    }
}

// ---------------------------------------------------------------- two-pass scheme, 32768 .. 262144 points
template <int FMT, int LOGA>
__global__ void __launch_bounds__(256)
k_fs_cols(const typename Ingest2<FMT>::raw_t* __restrict__ in, cpx* __restrict__ Y, const cpx* __restrict__ TPc, int B) {
    const int b = blockIdx.x * 256 + threadIdx.x;           // B is a multiple of 256
    const size_t frame = blockIdx.y, N = (size_t)B << LOGA;
    fs_cols<FMT, LOGA>(b, B, in + frame * N, TPc, Y + frame * N);
}

bool fft_twopass_supports(int n) { return n == 32768 || n == 131072 || n == 262144; }
static int twopass_b(int n) { return n == 262144 ? 8192 : 4096; }
static int twopass_loga(int n) { return n == 32768 ? 3 : 5; }
// tables: TPc[j][b] = W_N^(b 2^j), j < log2 A, then the small kernel's tables for B points (make_tp_tables)
int fft_twopass_table_elems(int n) { return fft_twopass_supports(n) ? twopass_loga(n) * twopass_b(n) + fft_small_table_elems(twopass_b(n)) : 0; }

template <int FMT, int LOGA>
static void tp_cols(const void* iq, cpx* Y, const cpx* TPc, int B, int n_frames, cudaStream_t st) {
    const dim3 grid((unsigned)(B / 256), (unsigned)n_frames);
    k_fs_cols<FMT, LOGA><<<grid, 256, 0, st>>>(static_cast<const typename Ingest2<FMT>::raw_t*>(iq), Y, TPc, B);
}
template <int C>
static cudaError_t tp_rows(const cpx* Y, float* power, const cpx* tables, int n_rows, int A, int sm_count, cudaStream_t st) {
    typedef FsGeom<C> G;
    static int per_sm = 0;
    if (!per_sm && (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_fft_small<FMT_CF32, C, true, false>, G::kThreads, fs_smem<C>()) != cudaSuccess || per_sm < 1))
        per_sm = 1;
    const long long slots = (long long)per_sm * (sm_count > 0 ? sm_count : 148);
    const int grid = (int)(n_rows < slots ? n_rows : slots);
    k_fft_small<FMT_CF32, C, true, false><<<grid, G::kThreads, fs_smem<C>(), st>>>(Y, power, tables, n_rows, A, 0, nullptr);
    return cudaGetLastError();
}
// Y: n_frames * n complex of scratch.  n_frames <= 65535 per call (the caller's chunks are far smaller).
cudaError_t launch_fft_twopass(int fmt, const void* iq, float* power, cpx* Y, const cpx* tables, int n, int n_frames,
                               int sm_count, cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    if (!fft_twopass_supports(n) || n_frames > 65535) return cudaErrorInvalidValue;
    const int B = twopass_b(n), loga = twopass_loga(n), A = 1 << loga;
    const cpx* TPc = tables;
    const cpx* small = tables + (size_t)loga * B;
#define SDRB_COLS(F)                                                                   \
    if (loga == 3) tp_cols<F, 3>(iq, Y, TPc, B, n_frames, st);                         \
    else tp_cols<F, 5>(iq, Y, TPc, B, n_frames, st);
    switch (fmt) {
        case FMT_CU8: SDRB_COLS(FMT_CU8) break;
        case FMT_CS16: SDRB_COLS(FMT_CS16) break;
        case FMT_CF32: SDRB_COLS(FMT_CF32) break;
        default: return cudaErrorInvalidValue;
    }
#undef SDRB_COLS
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (B == 4096) {
        static const bool one_row = getenv("SDRB_FFT_ROWS1") && getenv("SDRB_FFT_ROWS1")[0] == '1';   // test hook
        if (one_row) return tp_rows<16>(Y, power, small, n_frames * A, A, sm_count, st);
        const int n_groups = n_frames * A / 4;
        const int slots = sm_count > 0 ? sm_count : 148;
        k_fft_rows4<<<n_groups < slots ? n_groups : slots, 1024, kSmemRows4, st>>>(Y, power, small, n_groups, A);
        return cudaGetLastError();
    }
    return launch_8k<FMT_CF32, true>(Y, power, small, n_frames * A, A, sm_count, st);
}

}  // namespace sdrb
