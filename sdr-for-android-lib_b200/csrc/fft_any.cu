// fft_any.cu — generic power-of-two FFT for block lengths without a dedicated kernel.
//
// Replaces fftwf_plan_dft_1d/fftwf_execute (src/dsp/fft_process.cpp:77-79) for any
// N = 2^p, 1 <= p <= 24, as a sequence of Stockham autosort passes of radix 16 (and one
// pass of radix 8/4/2 for the remainder) over global memory.  The first pass fuses the
// ingest convert (K1), the last pass fuses |X|^2, the wire-format scale and the fftshift
// (K3/K4, src/dsp/fft_process.cpp:83-97).  This path moves 16 B per sample per pass, so it
// is the correctness path; N = 16384 has its own single-pass kernel (fft16k.cu) and the
// 64K / 1M sizes their four-step kernels (fft_big.cu).
#include "fft16k2.cuh"
#include "internal.h"

namespace sdrb {

SDRB_HD void fft8(cpx* v) {
    // X[k] is left in v[(k & 1) * 4 + (k >> 1)]
    fft2(v[0], v[4]);
    fft2(v[1], v[5]);
    fft2(v[2], v[6]);
    fft2(v[3], v[7]);
    v[5] = mul_w8_1(v[5]);
    v[6] = mul_mi(v[6]);
    v[7] = mul_w8_3(v[7]);
    fft4(v[0], v[1], v[2], v[3]);
    fft4(v[4], v[5], v[6], v[7]);
}

template <int R>
struct Radix;
template <>
struct Radix<16> {
    SDRB_HD static void run(cpx* v) { fft16(v); }
    SDRB_HD static int slot(int k) { return rev16(k); }
};
template <>
struct Radix<8> {
    SDRB_HD static void run(cpx* v) { fft8(v); }
    SDRB_HD static int slot(int k) { return (k & 1) * 4 + (k >> 1); }
};
template <>
struct Radix<4> {
    SDRB_HD static void run(cpx* v) { fft4(v[0], v[1], v[2], v[3]); }
    SDRB_HD static int slot(int k) { return k; }
};
template <>
struct Radix<2> {
    SDRB_HD static void run(cpx* v) { fft2(v[0], v[1]); }
    SDRB_HD static int slot(int k) { return k; }
};

// One Stockham pass.  Ns = product of the radices of the passes already done.
// FMT >= 0: input is the raw wire format (first pass); FMT < 0: input is cpx.
// LAST: write scaled, shifted power instead of cpx.
template <int R, int FMT, bool LAST>
__global__ void __launch_bounds__(256)
k_stockham(const void* __restrict__ in_, void* __restrict__ out_, const cpx* __restrict__ tw, int n, int Ns,
           float power_scale) {
    const int nb = n / R;  // butterflies per frame
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nb) return;
    const size_t frame = blockIdx.y;
    cpx v[R];
    if (FMT >= 0) {
        typedef typename Ingest<(FMT >= 0 ? FMT : 0)>::raw_t raw_t;
        const raw_t* in = static_cast<const raw_t*>(in_) + frame * (size_t)n;
#pragma unroll
        for (int r = 0; r < R; r++) v[r] = Ingest<(FMT >= 0 ? FMT : 0)>::cvt(in[j + r * nb]);
    } else {
        const cpx* in = static_cast<const cpx*>(in_) + frame * (size_t)n;
#pragma unroll
        for (int r = 0; r < R; r++) v[r] = in[j + r * nb];
    }
    const int k = j & (Ns - 1);
    if (Ns > 1) {
        const int tstep = n / (Ns * R);  // W_{Ns R}^(k r) = W_n^(k r tstep)
#pragma unroll
        for (int r = 1; r < R; r++) v[r] = cmul(v[r], tw[(size_t)k * r * tstep]);
    }
    Radix<R>::run(v);
    const int j0 = (j - k) * R + k;
    if (LAST) {
        float* out = static_cast<float*>(out_) + frame * (size_t)n;
        const int half = n >> 1;
#pragma unroll
        for (int r = 0; r < R; r++) {
            const cpx x = v[Radix<R>::slot(r)];
            const int bin = j0 + r * Ns;
            // re*re + im*im, then fftshift: bin k lands at (k + N/2) mod N
            const float p = __fadd_rn(__fmul_rn(x.x, x.x), __fmul_rn(x.y, x.y));
            out[(bin + half) & (n - 1)] = p * power_scale;
        }
    } else {
        cpx* out = static_cast<cpx*>(out_) + frame * (size_t)n;
#pragma unroll
        for (int r = 0; r < R; r++) out[j0 + r * Ns] = v[Radix<R>::slot(r)];
    }
}

template <int R, int FMT>
static void launch_pass(bool last, const void* in, void* out, const cpx* tw, int n, int Ns, float ps,
                        int n_frames, cudaStream_t st) {
    const int nb = n / R;
    const int threads = nb < 256 ? (nb < 32 ? 32 : nb) : 256;
    dim3 grid((unsigned)((nb + threads - 1) / threads), (unsigned)n_frames);
    if (last)
        k_stockham<R, FMT, true><<<grid, threads, 0, st>>>(in, out, tw, n, Ns, ps);
    else
        k_stockham<R, FMT, false><<<grid, threads, 0, st>>>(in, out, tw, n, Ns, ps);
}

template <int FMT>
static void launch_pass_r(int R, bool last, const void* in, void* out, const cpx* tw, int n, int Ns, float ps,
                          int n_frames, cudaStream_t st) {
    switch (R) {
        case 16: launch_pass<16, FMT>(last, in, out, tw, n, Ns, ps, n_frames, st); break;
        case 8: launch_pass<8, FMT>(last, in, out, tw, n, Ns, ps, n_frames, st); break;
        case 4: launch_pass<4, FMT>(last, in, out, tw, n, Ns, ps, n_frames, st); break;
        default: launch_pass<2, FMT>(last, in, out, tw, n, Ns, ps, n_frames, st); break;
    }
}

// Extension front end for the multi-pass paths (no reference code; definition in
// oracle/sdr_oracle_ext.c:sdr_oracle_fft_front): out[i] = (x[i] - mean) * w[i] as cf32, one CTA per
// frame.  front bit 0 = window, bit 1 = subtract the block mean.  The 16K kernel fuses this instead.
template <int FMT>
__global__ void __launch_bounds__(256)
k_fft_front(const typename Ingest2<FMT>::raw_t* __restrict__ in, cpx* __restrict__ out, const float* __restrict__ win,
            int n, int front) {
    __shared__ cpx red[8];
    const size_t frame = blockIdx.x;
    const typename Ingest2<FMT>::raw_t* src = in + frame * (size_t)n;
    cpx* dst = out + frame * (size_t)n;
    cpx m = mk(0.f, 0.f);
    if (front & 2) {
        cpx s = mk(0.f, 0.f);
        for (int i = threadIdx.x; i < n; i += 256) s = cadd(s, Ingest2<FMT>::cvt(src[i]));
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            s.x += __shfl_xor_sync(0xffffffffu, s.x, d);
            s.y += __shfl_xor_sync(0xffffffffu, s.y, d);
        }
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
        __syncthreads();
#pragma unroll
        for (int w = 0; w < 8; w++) m = cadd(m, red[w]);
        m.x = m.x / (float)n;
        m.y = m.y / (float)n;
    }
    for (int i = threadIdx.x; i < n; i += 256) {
        cpx v = csub(Ingest2<FMT>::cvt(src[i]), m);
        if (front & 1) { const float w = win[i]; v.x *= w; v.y *= w; }
        dst[i] = v;
    }
}

cudaError_t launch_fft_front(int fmt, const void* iq, cpx* out, const float* win, int n, int n_frames, int front,
                             cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    if ((front & 1) && !win) return cudaErrorInvalidValue;
    switch (fmt) {
        case FMT_CU8: k_fft_front<FMT_CU8><<<n_frames, 256, 0, st>>>(static_cast<const uint16_t*>(iq), out, win, n, front); break;
        case FMT_CS16: k_fft_front<FMT_CS16><<<n_frames, 256, 0, st>>>(static_cast<const uint32_t*>(iq), out, win, n, front); break;
        case FMT_CF32: k_fft_front<FMT_CF32><<<n_frames, 256, 0, st>>>(static_cast<const cpx*>(iq), out, win, n, front); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

int fft_generic_passes(int n) {
    int p = 0;
    while ((1 << p) < n) p++;
    return (p + 3) / 4;
}

cudaError_t launch_fft_generic(int fmt, const void* iq, float* power, cpx* scratch_a, cpx* scratch_b,
                               const cpx* twiddles, int n, int n_frames, cudaStream_t st, int* launches) {
    if (n_frames <= 0) return cudaSuccess;
    if (n < 2 || (n & (n - 1)) != 0) return cudaErrorInvalidValue;
    int bits = 0;
    while ((1 << bits) < n) bits++;
    float ps = fmt == FMT_CU8 ? Ingest<FMT_CU8>::kPowerScale
                              : (fmt == FMT_CS16 ? Ingest<FMT_CS16>::kPowerScale : 1.0f);
    // gridDim.y limit
    for (int f0 = 0; f0 < n_frames; f0 += 65535) {
        const int nf = n_frames - f0 < 65535 ? n_frames - f0 : 65535;
        const size_t in_bytes = fmt == FMT_CU8 ? 2 : (fmt == FMT_CS16 ? 4 : 8);
        const void* src = static_cast<const unsigned char*>(iq) + (size_t)f0 * n * in_bytes;
        float* pw = power + (size_t)f0 * n;
        cpx* sa = scratch_a;  // scratch holds one launch group at a time (stream order serialises reuse)
        cpx* sb = scratch_b;
        int left = bits, Ns = 1, pass = 0;
        const void* cur = src;
        while (left > 0) {
            const int rb = left >= 4 ? 4 : left;
            const int R = 1 << rb;
            const bool last = (left == rb);
            void* dst = last ? static_cast<void*>(pw) : static_cast<void*>((pass & 1) ? sb : sa);
            if (pass == 0) {
                switch (fmt) {
                    case FMT_CU8: launch_pass_r<FMT_CU8>(R, last, cur, dst, twiddles, n, Ns, ps, nf, st); break;
                    case FMT_CS16: launch_pass_r<FMT_CS16>(R, last, cur, dst, twiddles, n, Ns, ps, nf, st); break;
                    case FMT_CF32: launch_pass_r<FMT_CF32>(R, last, cur, dst, twiddles, n, Ns, ps, nf, st); break;
                    default: return cudaErrorInvalidValue;
                }
            } else {
                launch_pass_r<-1>(R, last, cur, dst, twiddles, n, Ns, ps, nf, st);
            }
            if (launches) (*launches)++;
            cur = dst;
            Ns *= R;
            left -= rb;
            pass++;
        }
    }
    return cudaGetLastError();
}


// ---------------------------------------------------------------- any other length: Bluestein
// fftwf_plan_dft_1d takes every length (src/dsp/fft_process.cpp:77) and BridgeConfig::setSamplesPerReading does not
// restrict it (src/bridge-config.h:56).  For N that is not a power of two the transform is the chirp-z identity
//   n k = (n^2 + k^2 - (k - n)^2) / 2   ->   X[k] = c[k] * sum_n (x[n] c[n]) * conj(c[k - n]),   c[n] = exp(-i pi n^2 / N)
// i.e. a circular convolution of length M = 2^p >= 2N - 1, done with the power-of-two passes above:
//   a = x * c, zero padded  ->  FFT_M  ->  * FFT_M(b) (host table, b[n] = b[M - n] = conj(c[n]) for n < N)
//   ->  inverse FFT_M as conj(FFT_M(conj(.))) / M  ->  * c[k]  ->  |.|^2, halves swapped.
// The swap is the reference's loop (src/dsp/fft_process.cpp:92-97): i runs to N/2 only, so an odd N leaves its last
// output bin at the value resize() gave it, zero.  Chirp phases come from n^2 mod 2N in integers (host, double).
template <int FMT>
__global__ void __launch_bounds__(256)
k_blue_pre(const typename Ingest2<FMT>::raw_t* __restrict__ in, const cpx* __restrict__ chirp, cpx* __restrict__ a,
           int n, int m) {
    const size_t frame = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    cpx v = mk(0.f, 0.f);
    if (i < n) v = cmul(Ingest2<FMT>::cvt(in[frame * (size_t)n + i]), chirp[i]);
    a[frame * (size_t)m + i] = v;
}
// a <- conj(a * bf): the conjugate turns the following forward transform into the inverse one
__global__ void __launch_bounds__(256)
k_blue_mul(cpx* __restrict__ a, const cpx* __restrict__ bf, int m) {
    const size_t frame = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const cpx v = cmul(a[frame * (size_t)m + i], bf[i]);
    a[frame * (size_t)m + i] = mk(v.x, -v.y);
}
__global__ void __launch_bounds__(256)
k_blue_post(const cpx* __restrict__ a, const cpx* __restrict__ chirp, float* __restrict__ power, int n, int m, float inv_m) {
    const size_t frame = blockIdx.y;
    const int o = blockIdx.x * blockDim.x + threadIdx.x;   // output (shifted) bin
    if (o >= n) return;
    const int half = n >> 1;
    float p = 0.0f;                                         // odd n: the last bin is never written by the reference
    if (o < 2 * half) {
        const int k = o < half ? o + half : o - half;
        const cpx r = a[frame * (size_t)m + k];
        const cpx x = cmul(mk(r.x * inv_m, -r.y * inv_m), chirp[k]);
        p = __fadd_rn(__fmul_rn(x.x, x.x), __fmul_rn(x.y, x.y));   // re*re + im*im, :85
    }
    power[frame * (size_t)n + o] = p;
}

int bluestein_m(int n) {
    int m = 1;
    while (m < 2 * n - 1) m <<= 1;
    return m;
}

// power-of-two cpx -> cpx transform of every frame: `buf` in, result in the returned buffer (buf or other)
static cpx* blue_fft(cpx* buf, cpx* other, const cpx* tw, int m, int n_frames, cudaStream_t st, int* launches) {
    int bits = 0;
    while ((1 << bits) < m) bits++;
    int left = bits, Ns = 1;
    cpx *cur = buf, *nxt = other;
    while (left > 0) {
        const int rb = left >= 4 ? 4 : left;
        launch_pass_r<-1>(1 << rb, false, cur, nxt, tw, m, Ns, 1.0f, n_frames, st);
        if (launches) (*launches)++;
        cpx* t = cur; cur = nxt; nxt = t;
        Ns <<= rb;
        left -= rb;
    }
    return cur;
}

// chirp: [n] c[k]; bf: [m] FFT_M of the wrapped conjugate chirp; tw_m: [m] W_m^k; sa, sb: [n_frames][m] each
cudaError_t launch_fft_bluestein(int fmt, const void* iq, float* power, cpx* sa, cpx* sb, const cpx* chirp,
                                 const cpx* bf, const cpx* tw_m, int n, int m, int n_frames, cudaStream_t st,
                                 int* launches) {
    if (n_frames <= 0) return cudaSuccess;
    const size_t in_bytes = fmt == FMT_CU8 ? 2 : (fmt == FMT_CS16 ? 4 : 8);
    for (int f0 = 0; f0 < n_frames; f0 += 65535) {   // gridDim.y limit
        const int nf = n_frames - f0 < 65535 ? n_frames - f0 : 65535;
        const void* src = static_cast<const unsigned char*>(iq) + (size_t)f0 * n * in_bytes;
        dim3 gm((unsigned)((m + 255) / 256), (unsigned)nf), gn((unsigned)((n + 255) / 256), (unsigned)nf);
        switch (fmt) {
            case FMT_CU8: k_blue_pre<FMT_CU8><<<gm, 256, 0, st>>>(static_cast<const uint16_t*>(src), chirp, sa, n, m); break;
            case FMT_CS16: k_blue_pre<FMT_CS16><<<gm, 256, 0, st>>>(static_cast<const uint32_t*>(src), chirp, sa, n, m); break;
            case FMT_CF32: k_blue_pre<FMT_CF32><<<gm, 256, 0, st>>>(static_cast<const cpx*>(src), chirp, sa, n, m); break;
            default: return cudaErrorInvalidValue;
        }
        cpx* a = blue_fft(sa, sb, tw_m, m, nf, st, launches);
        cpx* o = a == sa ? sb : sa;
        k_blue_mul<<<gm, 256, 0, st>>>(a, bf, m);
        a = blue_fft(a, o, tw_m, m, nf, st, launches);
        k_blue_post<<<gn, 256, 0, st>>>(a, chirp, power + (size_t)f0 * n, n, m, 1.0f / (float)m);
        if (launches) (*launches) += 3;
    }
    return cudaGetLastError();
}

}  // namespace sdrb
