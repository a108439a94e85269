// nco.cuh — numerically controlled oscillator shared by the extended SSB kernels.
// Extension stage (the reference has no NCO: it demodulates at the tuned centre only,
// src/ssb/ssb_demod_opt.cpp:221-296).  Definition = oracle/sdr_oracle_ext.c:sdr_oracle_nco;
// written with explicit fused multiply-adds so host, oracle and device agree bit for bit.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "fft_core.cuh"

namespace sdrb {

// x with its sign bit XORed by `sign` (0 or 0x80000000): exact negation
SDRB_HD float nco_flip(float x, uint32_t sign) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(__float_as_uint(x) ^ sign);
#else
    uint32_t u;
    memcpy(&u, &x, 4);
    u ^= sign;
    memcpy(&x, &u, 4);
    return x;
#endif
}

// cos/sin of a 32-bit phase (turns * 2^32): quadrant reduction, Cephes minimax polynomials.
SDRB_HD void nco_cs(uint32_t phase, float& c_out, float& s_out) {
    const uint32_t q = (phase + 0x20000000u) >> 30;
    const int32_t r = (int32_t)(phase - (q << 30));
#if defined(__CUDA_ARCH__)
    const float x = __fmul_rn((float)r, 1.4629180792671596e-9f);
    const float z = __fmul_rn(x, x);
    const float xz = __fmul_rn(x, z);
#else
    const float x = (float)r * 1.4629180792671596e-9f;
    const float z = x * x;
    const float xz = x * z;
#endif
    float sp = fmaf(z, -1.9515295891e-4f, 8.3321608736e-3f);
    sp = fmaf(z, sp, -1.6666654611e-1f);
    const float s = fmaf(xz, sp, x);
    float cp = fmaf(z, 2.443315711809948e-5f, -1.388731625493765e-3f);
    cp = fmaf(z, cp, 4.166664568298827e-2f);
    cp = fmaf(z, cp, -0.5f);
    const float c = fmaf(z, cp, 1.0f);
    // quadrant: (c, s), (-s, c), (-c, -s), (s, -c) for q = 0..3, as one swap and two sign flips (no branches:
    // a 4-way ?: chain compiled to branches cost the one-warp state pass a cycle per instruction)
    const uint32_t qq = q & 3u;
    const bool swap = (qq & 1u) != 0;
    const float cc = swap ? s : c, ss = swap ? c : s;
    c_out = nco_flip(cc, ((qq + 1u) & 2u) << 30);
    s_out = nco_flip(ss, (qq & 2u) << 30);
}

// Phase increment of f_hz at sample rate fs (two's complement for negative offsets).
inline uint32_t nco_inc(double f_hz, uint32_t fs) {
    return (uint32_t)(int64_t)llround(f_hz / (double)fs * 4294967296.0);
}

}  // namespace sdrb
