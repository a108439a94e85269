"""Pins the C restatement (oracle/sdr_oracle.c) against the reference itself
(oracle/_ref/libsdr_ref.so = the reference's fft_process.cpp + ssb_demod_opt.cpp + vendored FFTW,
compiled unmodified by oracle/Makefile).  The reference has no tests or golden vectors of its own
(SURVEY.md section 4), so this is the pin.  CPU only."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref/libsdr_ref.so not built")
fp = ol._fp


def synth_block(sdrb, fmt, fs, stream, block, n):
    raw = sdrb.synth_host(fmt, fs, stream, 1, block * n, n)[0]
    return ol.to_cf32(raw, fmt)


@pytest.mark.parametrize("n,fs,focus", [(16384, 2_000_000, 5), (16384, 2_000_000, 200), (4096, 2_500_000, 5),
                                        (65536, 30_720_000, 5), (1024, 250_000, 5), (16384, 2_000_000, 300)])
def test_fft_path_matches_reference(sdrb, n, fs, focus):
    fc = 430_000_000
    rf = ol.RefFft(fc, fs, n, focus)
    of_same_power = ol.OracleFft(fc, fs, n, focus)   # estimator pinned bit-exactly on identical power
    of_own_fft = ol.OracleFft(fc, fs, n, focus)      # whole path, own double-precision DFT
    for b in range(6):
        iq = synth_block(sdrb, 0, fs, 3, b, n)
        r = rf.process(iq)
        p_ref = rf.power()
        e = of_same_power.evaluate(p_ref)
        for k in ol.RESULT_FIELDS + ("trackingFrequencyHz", "maxPeakDb", "maxPeakHz"):
            assert np.float32(e[k]).tobytes() == np.float32(r[k]).tobytes(), (b, k, e[k], r[k])
        assert e["trackingFrequency"] == r["trackingFrequency"]
        assert e["detectionFlag"] == r["detectionFlag"]
        # power: float FFTW vs double DFT, 1e-5 relative of the spectrum's scale
        p_or = ol.oracle_power(iq)
        scale = max(float(p_ref.max()), 1e-30)
        assert np.all(np.abs(p_or - p_ref) <= 2e-5 * p_ref + 1e-7 * scale)
        o = of_own_fft.evaluate(p_or)
        for k in ("meanSnrDb", "peakAboveNoiseMeanDb", "maxBinSnrDb", "best1kHzSnrDb", "perBinMean"):
            assert abs(o[k] - r[k]) < 0.01, (b, k, o[k], r[k])
    rf.close()


def test_tracking_and_flag_sequence(sdrb):
    """Long enough for the 300 ms tracking gate and the detection ring to act (fft_process.cpp:333-378)."""
    n, fs, fc = 16384, 2_000_000, 430_000_000
    rf = ol.RefFft(fc, fs, n, 5)
    of = ol.OracleFft(fc, fs, n, 5)
    seen_track, seen_flag = set(), set()
    for b in range(60):
        iq = synth_block(sdrb, 0, fs, 0, b, n)
        if b == 40:
            ol.ref().sdr_ref_set_center_frequency_changed(1)
            of.st.centerChanged = 1
        r = rf.process(iq)
        e = of.evaluate(rf.power())
        assert e["trackingFrequency"] == r["trackingFrequency"]
        assert e["detectionFlag"] == r["detectionFlag"]
        seen_track.add(r["trackingFrequency"])
        seen_flag.add(r["detectionFlag"])
    assert len(seen_track) >= 2 and seen_flag == {0, 3}
    rf.close()


def test_ssb_stages_bit_exact(sdrb):
    o, r = ol.oracle(), ol.ref()
    n, fs = 16384, 2_000_000
    iq = synth_block(sdrb, 0, fs, 1, 0, n)
    a, b = iq.copy(), iq.copy()
    o.sdr_oracle_remove_dc(fp(a), n, 0.9995)
    r.sdr_ref_removeDC(fp(b), n, 0.9995)
    assert a.tobytes() == b.tobytes()
    for (fc_, q) in ((3200.0, 0.9), (2200.0, 1.2)):
        bq = ol.Biquad()
        o.sdr_oracle_iir2_init_lowpass(C.byref(bq), fs, fc_, q)
        cs = np.zeros(7, np.float32)
        r.sdr_ref_iir2_init_lowpass(fp(cs), fs, fc_, q)
        assert np.array([bq.a0, bq.a1, bq.a2, bq.b1, bq.b2], np.float32).tobytes() == cs[:5].tobytes()
    a2, b2 = a.copy(), b.copy()
    o.sdr_oracle_iir2_process(C.byref(bq), fp(a2), n)
    r.sdr_ref_iir2_process(fp(cs), fp(b2), n)
    assert a2.tobytes() == b2.tobytes() and (bq.z1, bq.z2) == (cs[5], cs[6])
    for upper in (1, 0):
        au, bu = np.zeros(n, np.float32), np.zeros(n, np.float32)
        o.sdr_oracle_demod_ssb(fp(a2), n, upper, fp(au))
        r.sdr_ref_demodSSB(fp(b2), n, upper, fp(bu))
        assert au.tobytes() == bu.tobytes()
        if not upper:
            assert not au.any()          # the reference's LSB quirk: y - y = 0
    au, bu = np.zeros(n, np.float32), np.zeros(n, np.float32)
    o.sdr_oracle_demod_ssb(fp(a2), n, 1, fp(au))
    bu[:] = au
    o.sdr_oracle_adaptive_agc(fp(au), n, 0.35, 0.006, 0.00035)
    r.sdr_ref_adaptiveAGC(fp(bu), n, 0.35, 0.006, 0.00035)
    assert au.tobytes() == bu.tobytes()
    for decim in (41, 1, 5, 640, 52):
        oa, ob = np.zeros(n, np.float32), np.zeros(n, np.float32)
        ma = o.sdr_oracle_fir_decimate(fp(au), n, decim, 0.45, fp(oa), n)
        mb = r.sdr_ref_fir_decimate(fp(bu), n, decim, 0.45, fp(ob), n)
        assert ma == mb == (n - 255) // decim + 1 and oa.tobytes() == ob.tobytes()
    for name in ("highpass", "bandpass"):
        bq = ol.Biquad()
        getattr(o, f"sdr_oracle_biquad_init_{name}")(C.byref(bq), 48000.0, 1200.0, 0.7)
        cs = np.zeros(7, np.float32)
        getattr(r, f"sdr_ref_biquad_init_{name}")(fp(cs), 48000.0, 1200.0, 0.7)
        assert np.array([bq.a0, bq.a1, bq.a2, bq.b1, bq.b2], np.float32).tobytes() == cs[:5].tobytes()
        xa, xb = oa[:394].copy(), ob[:394].copy()
        o.sdr_oracle_biquad_process(C.byref(bq), fp(xa), 394)
        r.sdr_ref_biquad_process(fp(cs), fp(xb), 394)
        assert xa.tobytes() == xb.tobytes()
    xa, xb = oa[:394].copy(), ob[:394].copy()
    o.sdr_oracle_transient_boost(fp(xa), 394, 0.55)
    r.sdr_ref_transient_boost(fp(xb), 394, 0.55)
    assert xa.tobytes() == xb.tobytes()
    pa, pb = np.zeros(394, np.int16), np.zeros(394, np.int16)
    o.sdr_oracle_float_to_pcm(fp(xa), 394, 10.0, pa.ctypes.data_as(ol.i16p))
    r.sdr_ref_float_to_pcm(fp(xb), 394, 10.0, pb.ctypes.data_as(ol.i16p))
    assert pa.tobytes() == pb.tobytes() and pa.any()


def test_convert_u8_matches_reference():
    raw = np.arange(512, dtype=np.uint8).reshape(256, 2)
    a = ol.to_cf32(raw, 0)
    b = np.zeros_like(a)
    ol.ref().sdr_ref_convertIQ(raw.ctypes.data_as(C.c_void_p), raw.size, fp(b))
    assert a.tobytes() == b.tobytes()


_SSB_CHILD = r"""
import sys, numpy as np, importlib
sys.path.insert(0, {tests!r}); sys.path.insert(0, {root!r})
import oracle_lib as ol
sdrb = importlib.import_module("sdr-for-android-lib_b200")
n, fs, mode, upper, stream, blocks = {n}, {fs}, {mode}, {upper}, {stream}, {blocks}
os_ = ol.OracleSsb()
for b in range(blocks):
    iq = ol.to_cf32(sdrb.synth_host(0, fs, stream, 1, b * n, n)[0], 0)
    out = np.zeros(n + 8, np.int16)
    m = ol.ref().sdr_ref_process_ssb(ol._fp(iq), n, fs, upper, mode, out.ctypes.data_as(ol.i16p), n + 8)
    p = os_.process(iq, fs, upper, mode)
    assert m == len(p), (m, len(p))
    assert out[:m].tobytes() == p.tobytes(), (b, np.abs(out[:m].astype(int) - p).max())
    if upper: assert p.any()
    else: assert not p.any()
print("OK", m)
"""


@pytest.mark.parametrize("n,fs,mode,upper", [(16384, 2_000_000, 1, 1), (16384, 2_000_000, 0, 1),
                                             (16384, 2_500_000, 2, 1), (16384, 2_000_000, 1, 0),
                                             (65536, 30_720_000, 1, 1), (4096, 250_000, 1, 1)])
def test_process_ssb_bit_exact(n, fs, mode, upper):
    """processSSB_opt keeps its state in function statics, so each case runs in its own process."""
    import subprocess
    import sys
    code = _SSB_CHILD.format(tests=ol.ROOT + "/tests", root=ol.ROOT, n=n, fs=fs, mode=mode, upper=upper, stream=2,
                             blocks=5)
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and "OK" in res.stdout, res.stdout + res.stderr
