"""bench_synth.py — numpy restatement of the deterministic synthetic IQ source
(sdr-for-android-lib_b200/csrc/synth.cuh: integer-only, seed 0x5D120000 + stream id).

The CPU legs of bench.py (`--impl reference`, `cpu_baseline`) generate their input with this file so
that they never map the product library; tests/test_cpu_host.py checks it byte for byte against
`sdrb_synth_host`.  Stands in for SoapySDR readStream (src/sdr-bridge-java-soapy.cpp:532-539).
"""
import numpy as np

CU8, CS16, CF32 = 0, 1, 2
_M32 = np.uint64(0xFFFFFFFF)


def _u32(x):
    return np.asarray(x, dtype=np.uint64) & _M32


def _mix32(x):
    x = _u32(x)
    x ^= x >> np.uint64(16)
    x = _u32(x * np.uint64(0x7FEB352D))
    x ^= x >> np.uint64(15)
    x = _u32(x * np.uint64(0x846CA68B))
    x ^= x >> np.uint64(16)
    return x


def _sine_table():
    i = np.arange(4096, dtype=np.float64)
    return np.rint(32767.0 * np.sin(2.0 * np.pi * i / 4096.0)).astype(np.int64)


_TAB = _sine_table()


def _cos(ph):
    return _TAB[((ph >> np.uint64(20)) + np.uint64(1024)) & np.uint64(4095)]


def _sin(ph):
    return _TAB[(ph >> np.uint64(20)) & np.uint64(4095)]


def _tdiv(a, b):
    """C integer division (truncation toward zero)."""
    q = abs(a) // abs(b)
    return q if (a >= 0) == (b >= 0) else -q


def stream_params(stream_id, fs):
    seed = (0x5D120000 + stream_id) & 0xFFFFFFFF
    h1 = int(_mix32(seed ^ 0xA5A5A5A5))
    h2 = int(_mix32((h1 + 0x9E3779B9) & 0xFFFFFFFF))
    h3 = int(_mix32((h2 + 0x9E3779B9) & 0xFFFFFFFF))
    tone1 = -4000 + h1 % 8001
    tone2 = (12500 + 1000 * (h2 % 5)) * (-1 if (h2 >> 16) & 1 else 1)
    step1 = _tdiv(tone1 * 4294967296, fs) & 0xFFFFFFFF
    step2 = _tdiv(tone2 * 4294967296, fs) & 0xFFFFFFFF
    per = (fs * 7) // 4 or 1
    key_on = per // 10
    key_off = 0 if stream_id % 4 == 0 else h3 % per
    return dict(seed=seed, step1=step1, step2=step2, keyPeriod=per, keyOn=key_on, keyOffset=key_off)


def synth(fmt, fs, stream_id, first_sample, n_samples):
    """[n_samples, 2] in the wire dtype (u8 / i16 / f32) of one stream, identical to sdrb_synth_host."""
    p = stream_params(stream_id, fs)
    i = np.arange(first_sample, first_sample + n_samples, dtype=np.uint64)
    lo, hi = i & _M32, i >> np.uint64(32)
    seed = np.uint64(p["seed"])
    hI = _mix32(_u32(lo * np.uint64(0x9E3779B1)) ^ _u32(seed * np.uint64(0x85EBCA77)) ^ _u32(hi * np.uint64(0xC2B2AE3D)))
    hQ = _mix32(hI ^ np.uint64(0x68E31DA4))

    def bytesum(h):
        return ((h & np.uint64(255)) + ((h >> np.uint64(8)) & np.uint64(255)) + ((h >> np.uint64(16)) & np.uint64(255)) +
                (h >> np.uint64(24))).astype(np.int64) - 510

    nI, nQ = bytesum(hI), bytesum(hQ)
    ph1 = _u32(lo * np.uint64(p["step1"]))
    ph2 = _u32(lo * np.uint64(p["step2"]))
    on = ((i + np.uint64(p["keyOffset"])) % np.uint64(p["keyPeriod"])) < np.uint64(p["keyOn"])
    if fmt == CU8:
        hD = _mix32(_u32(hQ + np.uint64(0x9E3779B9)))
        vi = 127 + ((hD & np.uint64(255)) < 102).astype(np.int64) + ((nI * 7 + 128) >> 8) + ((6 * _cos(ph2) + 16384) >> 15)
        vq = 127 + (((hD >> np.uint64(8)) & np.uint64(255)) < 102).astype(np.int64) + ((nQ * 7 + 128) >> 8) + \
            ((6 * _sin(ph2) + 16384) >> 15)
        vi = vi + np.where(on, (38 * _cos(ph1) + 16384) >> 15, 0)
        vq = vq + np.where(on, (38 * _sin(ph1) + 16384) >> 15, 0)
        return np.stack([np.clip(vi, 0, 255), np.clip(vq, 0, 255)], axis=1).astype(np.uint8)
    vi = nI * 2 + ((1638 * _cos(ph2) + 16384) >> 15)
    vq = nQ * 2 + ((1638 * _sin(ph2) + 16384) >> 15)
    vi = vi + np.where(on, (9830 * _cos(ph1) + 16384) >> 15, 0)
    vq = vq + np.where(on, (9830 * _sin(ph1) + 16384) >> 15, 0)
    out = np.stack([np.clip(vi, -32767, 32767), np.clip(vq, -32767, 32767)], axis=1)
    if fmt == CS16:
        return out.astype(np.int16)
    return (out.astype(np.float32) * np.float32(1.0 / 32768.0)).astype(np.float32)


def to_cf32(raw, fmt):
    """The wire-format conversion in front of the reference's CF32 interface: u8 follows convertIQ
    (src/ssb/ssb_demod_opt.cpp:37-41), cs16 is x * 2^-15 (stated convention, LimeSuite source absent)."""
    if fmt == CU8:
        return ((raw.astype(np.float32) - np.float32(127.4)) * np.float32(1.0 / 128.0)).astype(np.float32)
    if fmt == CS16:
        return (raw.astype(np.float32) * np.float32(1.0 / 32768.0)).astype(np.float32)
    return np.ascontiguousarray(raw, dtype=np.float32)
