"""CPU tests of the extension-stage definitions (oracle/sdr_oracle_ext.c): with every knob at
identity they reduce to the pinned restatement, and the Weaver chain really selects a sideband."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol


def test_nco_accuracy_and_symmetry():
    o = ol.oracle()
    rng = np.random.default_rng(1)
    phases = np.concatenate([rng.integers(0, 2**32, 5000, dtype=np.uint64),
                             np.array([0, 2**30, 2**31, 3 * 2**30, 2**29, 2**29 - 1, 2**32 - 1], dtype=np.uint64)])
    c, s = C.c_float(), C.c_float()
    err = 0.0
    for p in phases:
        o.sdr_oracle_nco(int(p), C.byref(c), C.byref(s))
        a = 2 * np.pi * float(p) / 2**32
        err = max(err, abs(c.value - np.cos(a)), abs(s.value - np.sin(a)))
    assert err < 3e-7
    o.sdr_oracle_nco(0, C.byref(c), C.byref(s))
    assert c.value == 1.0 and s.value == 0.0
    assert o.sdr_oracle_nco_inc(-1000.0, 2_000_000) == (2**32 - o.sdr_oracle_nco_inc(1000.0, 2_000_000))


@pytest.mark.parametrize("upper,mode", [(1, 1), (0, 1), (1, 0)])
def test_identity_settings_reduce_to_pinned_chain(sdrb, upper, mode):
    fs, n, B = 2_000_000, 16384, 4
    iq = ol.to_cf32(sdrb.synth_host(sdrb.CU8, fs, 5, 1, 0, B * n)[0].reshape(B, n, 2), sdrb.CU8)
    a, b = ol.OracleSsb(), ol.OracleSsbx(0.0, 0)
    for k in range(B):
        assert np.array_equal(a.process(iq[k], fs, upper, mode), b.process(iq[k], fs, upper, mode))


def test_front_identity_and_properties():
    rng = np.random.default_rng(2)
    iq = rng.standard_normal((1024, 2)).astype(np.float32) + np.float32(0.25)
    assert np.array_equal(ol.oracle_fft_front(iq, 0, 0), iq)
    d = ol.oracle_fft_front(iq, 0, 1)
    assert abs(d[:, 0].astype(np.float64).mean()) < 1e-7 and abs(d[:, 1].astype(np.float64).mean()) < 1e-7
    w = ol.oracle_fft_front(np.ones((1024, 2), np.float32), 1, 0)[:, 0]
    assert w[0] == 0.0 and w[-1] == 0.0 and abs(w[511] - 1.0) < 1e-5 and np.allclose(w, w[::-1], atol=1e-7)


def _tone(fs, n_total, f_hz, amp=0.2):
    t = np.arange(n_total, dtype=np.float64) / fs
    z = amp * np.exp(2j * np.pi * f_hz * t)
    return np.stack([z.real, z.imag], axis=-1).astype(np.float32)


def test_weaver_selects_the_sideband():
    """A tone 1 kHz above the channel centre is audible in USB and attenuated in LSB (and vice
    versa); the reference's own switch cannot do this (USB = 2y, LSB = 0)."""
    fs, n, B, off = 2_000_000, 16384, 24, 20_000.0

    def audio_rms(f_tone, upper):
        iq = _tone(fs, B * n, off + f_tone).reshape(B, n, 2)
        ch = ol.OracleSsbx(off, 1)
        out = [ch.process(iq[k], fs, upper, 1) for k in range(B)]
        x = np.concatenate(out[B // 2:]).astype(np.float64)
        return np.sqrt(np.mean(x * x))

    usb_in, usb_out = audio_rms(+1000.0, 1), audio_rms(-1000.0, 1)
    lsb_in, lsb_out = audio_rms(-1000.0, 0), audio_rms(+1000.0, 0)
    assert usb_in > 2.0 * usb_out and lsb_in > 2.0 * lsb_out
    assert usb_in > 500 and lsb_in > 500
