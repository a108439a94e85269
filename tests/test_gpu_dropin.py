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
    off += 4
    # pass 3: the first call freezes the block length (ssb_demod_opt.cpp:225-227): a half block is zero-extended, a
    # double block cut; the last call passes sound mode 7, which leaves the previous mode's parameters in place
    sizes = np.frombuffer(blob, dtype=np.int32, count=3, offset=off); off += 12
    assert tuple(sizes) == (m, m, m)
    fresh = ol.OracleSsb()
    feeds = ((iq[0], mode), (iq[1][:n // 2], mode), (np.concatenate([iq[2], iq[3]]), 7))
    for k, (block, md) in enumerate(feeds):
        got = np.frombuffer(blob, dtype=np.int16, count=m, offset=off); off += 2 * m
        exp = fresh.process(block, fs, 1, md)        # the oracle latches the first length itself (sdr_oracle.c)
        assert len(exp) == m and np.abs(got.astype(np.int32) - exp.astype(np.int32)).max(initial=0) <= 1, k
        assert (got == exp).mean() >= 0.999
    assert off == len(blob)


def test_ragged_reads_through_the_batcher_equal_a_direct_call(sdrb, gpu_ctx, tmp_path):
    """host/iq_chunker.h in front of sdrb_process_host on the GPU: per-stream ragged reads ->
    exact blocks -> [S][B][N] batches -> one pipelined host-to-host call per batch, two batches so that the carried
    state crosses a batch boundary.  Records, spectra and PCM must equal a direct sdrb_process on the same samples."""
    exe = os.path.join(ROOT, "sdr-for-android-lib_b200", "batcher_check.bin")
    assert os.path.exists(exe), "batcher_check.bin missing: run __graft_entry__.build()"
    S, B, n, batches, fs = 5, 3, 16384, 2, 2_000_000
    out = tmp_path / "batcher.bin"
    subprocess.run([exe, str(out), str(S), str(B), str(n), str(batches), str(fs)], check=True, timeout=300)
    blob = out.read_bytes()
    hdr = np.frombuffer(blob, dtype=np.int32, count=4)
    m = int(hdr[3])
    assert tuple(hdr[:3]) == (S, B, n) and m > 0
    gpu_ctx.configure(sdrb.Config(FC, fs, n, 5, inputFormat=sdrb.CU8), S, B)
    off = 16
    for k in range(batches):
        rec = np.frombuffer(blob, dtype=sdrb.RESULT_DTYPE, count=S * B, offset=off).reshape(S, B); off += rec.nbytes
        power = np.frombuffer(blob, dtype=np.float32, count=S * B * n, offset=off).reshape(S, B, n); off += power.nbytes
        pcm = np.frombuffer(blob, dtype=np.int16, count=S * B * m, offset=off).reshape(S, B, m); off += pcm.nbytes
        iq = np.ascontiguousarray(sdrb.synth_host(sdrb.CU8, fs, 0, S, k * B * n, B * n).reshape(S, B, n, 2))
        gpu_ctx.process(iq, B, sdrb.STAGE_FFT | sdrb.STAGE_SSB)
        assert np.array_equal(gpu_ctx.results(), rec), k
        assert np.array_equal(gpu_ctx.power_all(), power), k
        assert np.array_equal(gpu_ctx.pcm(), pcm), k
    assert off == len(blob)
