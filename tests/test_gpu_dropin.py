"""GPU test of the header-compatible C++ front end (sdr-for-android-lib_b200/host): FFTProcessor,
processSSB_opt and SSBProcessor driven by host/dropin_check.cpp the way soapyCallback drives the
reference objects (src/sdr-bridge-java-soapy.cpp:424-493), compared with the oracle."""
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "sdr-for-android-lib_b200", "dropin_check.bin")
FC = 430_000_000


@pytest.mark.parametrize("n,fs,B,focus,mode", [(16384, 2_000_000, 40, 5, 1), (4096, 250_000, 12, 5, 2),
                                                (65536, 2_500_000, 6, 200, 0)])
def test_dropin_front_end(sdrb, tmp_path, n, fs, B, focus, mode):
    assert os.path.exists(BIN), "dropin_check.bin missing: run __graft_entry__.build()"
    out = tmp_path / "dropin.bin"
    subprocess.run([BIN, str(out), str(n), str(fs), str(B), str(focus), str(mode)], check=True, timeout=300)
    blob = out.read_bytes()
    hdr = np.frombuffer(blob, dtype=np.int32, count=4)
    assert hdr[0] == n and hdr[1] == B and hdr[3] == 0
    m = int(hdr[2])
    iq = sdrb.synth_host(sdrb.CF32, fs, 3, 1, 0, B * n)[0].reshape(B, n, 2)
    fft, ssb = ol.OracleFft(FC, fs, n, focus), ol.OracleSsb()
    off = 16
    for b in range(B):
        g = np.frombuffer(blob, dtype=np.float32, count=9, offset=off); off += 36
        tr = int(np.frombuffer(blob, dtype=np.int64, count=1, offset=off)[0]); off += 8
        flag, peak = np.frombuffer(blob, dtype=np.int32, count=2, offset=off); off += 8
        power = np.frombuffer(blob, dtype=np.float32, count=n, offset=off); off += 4 * n
        pcm = np.frombuffer(blob, dtype=np.int16, count=m, offset=off); off += 2 * m
        if b == B // 2:
            fft.set_center(FC + 25_000)
        e = fft.evaluate(power)                       # the estimator on the spectrum the getter returned
        for i, k in enumerate(ol.RESULT_FIELDS):
            tol = 0.01 if not k.endswith("Sigma") else 0.02 + 1e-3 * abs(e[k])
            if k == "best1kHzCenterFreqHz":
                assert np.float32(g[i]) == np.float32(e[k]), (b, k)
            else:
                assert abs(float(g[i]) - e[k]) < tol, (b, k, float(g[i]), e[k])
        assert tr == e["trackingFrequency"] and int(flag) == e["detectionFlag"] and int(peak) == e["peakBin"], b
        ref_p = ol.oracle_power(iq[b])
        total = float(ref_p.astype(np.float64).sum())
        tol = 1e-5 * ref_p + 6e-7 * np.sqrt(ref_p) * np.sqrt(total) + 1e-9 * total / n
        assert np.all(np.abs(power - ref_p) <= tol), b
        exp = ssb.process(iq[b], fs, 1, mode)
        assert len(exp) == m
        assert np.abs(pcm.astype(np.int32) - exp.astype(np.int32)).max(initial=0) <= 1, b
    same = np.frombuffer(blob, dtype=np.int32, count=1, offset=off)[0]
    assert same == 1, "SSBProcessor (queue + worker) PCM differs from direct processSSB_opt calls"
