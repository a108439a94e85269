"""Golden vectors generated from the reference itself (tests/golden/make_golden.py).
CPU: the oracle restatement against them.  GPU: the CUDA path through the C-ABI against them."""
import glob
import os

import numpy as np
import pytest

import oracle_lib as ol

GOLD = os.path.join(ol.ROOT, "tests", "golden")
FFT_FILES = sorted(glob.glob(os.path.join(GOLD, "fft_*.npz")))
SSB_FILES = sorted(glob.glob(os.path.join(GOLD, "ssb_*.npz")))
DB_KEYS = ("meanSnrDb", "peakAboveNoiseMeanDb", "maxBinSnrDb", "best1kHzSnrDb", "perBinMean")


def blocks_of(sdrb, fmt, fs, stream, blocks, n):
    return np.ascontiguousarray(sdrb.synth_host(fmt, fs, stream, 1, 0, blocks * n).reshape(1, blocks, n, 2))


def spectrum_close(got, g, n, b):
    sub, cen = got[::16], got[n // 2 - 128:n // 2 + 128]
    scale = float(g["power_every16"][b].max())
    for a, e in ((sub, g["power_every16"][b]), (cen, g["power_center"][b])):
        assert np.all(np.abs(a - e) <= 3e-5 * e + 2e-6 * scale)


def test_fixture_inventory():
    assert len(FFT_FILES) == 4 and len(SSB_FILES) == 4


@pytest.mark.parametrize("path", FFT_FILES, ids=os.path.basename)
def test_oracle_fft_vs_golden(sdrb, path):
    g = np.load(path)
    fmt, fs, n, focus, stream, blocks = (int(v) for v in g["meta"])
    iq = blocks_of(sdrb, fmt, fs, stream, blocks, n)[0]
    of = ol.OracleFft(430_000_000, fs, n, focus)
    for b in range(blocks):
        p = ol.oracle_power(ol.to_cf32(iq[b], fmt))
        e = of.evaluate(p)
        if b < 2:
            spectrum_close(p, g, n, b)
        for k in DB_KEYS:
            assert abs(e[k] - float(g[k][b])) < 0.01, (b, k)
        assert e["detectionFlag"] == int(g["detectionFlag"][b])
        assert e["trackingFrequency"] == int(g["trackingFrequency"][b])
        assert np.float32(e["maxPeakHz"]) == g["maxPeakHz"][b]
        assert np.float32(e["best1kHzCenterFreqHz"]) == g["best1kHzCenterFreqHz"][b]


@pytest.mark.parametrize("path", SSB_FILES, ids=os.path.basename)
def test_oracle_ssb_vs_golden(sdrb, path):
    g = np.load(path)
    fmt, fs, n, mode, upper, stream, blocks = (int(v) for v in g["meta"])
    iq = blocks_of(sdrb, fmt, fs, stream, blocks, n)[0]
    o = ol.OracleSsb()
    for b in range(blocks):
        assert np.array_equal(o.process(ol.to_cf32(iq[b], fmt), fs, upper, mode), g["pcm"][b])


@pytest.mark.gpu
@pytest.mark.parametrize("path", FFT_FILES, ids=os.path.basename)
def test_gpu_fft_vs_golden(sdrb, gpu_ctx, path):
    g = np.load(path)
    fmt, fs, n, focus, stream, blocks = (int(v) for v in g["meta"])
    iq = blocks_of(sdrb, fmt, fs, stream, blocks, n)
    gpu_ctx.configure(sdrb.Config(430_000_000, fs, n, focus, inputFormat=fmt), 1, blocks)
    gpu_ctx.process(iq, blocks, sdrb.STAGE_FFT)
    res = gpu_ctx.results()[0]
    for b in range(blocks):
        if b < 2:
            spectrum_close(gpu_ctx.power(0, b), g, n, b)
        for k in DB_KEYS:
            assert abs(float(res[b][k]) - float(g[k][b])) < 0.01, (b, k)
        assert int(res[b]["detectionFlag"]) == int(g["detectionFlag"][b])
        assert int(res[b]["trackingFrequency"]) == int(g["trackingFrequency"][b])
        assert np.float32(res[b]["best1kHzCenterFreqHz"]) == g["best1kHzCenterFreqHz"][b]


@pytest.mark.gpu
@pytest.mark.parametrize("path", SSB_FILES, ids=os.path.basename)
def test_gpu_ssb_vs_golden(sdrb, gpu_ctx, path):
    g = np.load(path)
    fmt, fs, n, mode, upper, stream, blocks = (int(v) for v in g["meta"])
    iq = blocks_of(sdrb, fmt, fs, stream, blocks, n)
    gpu_ctx.configure(sdrb.Config(430_000_000, fs, n, 5, soundMode=mode, inputFormat=fmt, upperSideband=upper), 1,
                      blocks)
    gpu_ctx.process(iq, blocks, sdrb.STAGE_SSB)
    pcm = gpu_ctx.pcm()[0]
    assert pcm.shape == g["pcm"].shape
    assert np.abs(pcm.astype(int) - g["pcm"].astype(int)).max(initial=0) <= 1
    assert (pcm == g["pcm"]).mean() >= 0.999 if pcm.size else True
