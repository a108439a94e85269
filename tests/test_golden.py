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


SIGMA_KEYS = ("meanSnrSigma", "maxBinSnrSigma", "best1kHzSnrSigma")


@pytest.mark.gpu
@pytest.mark.parametrize("path", FFT_FILES, ids=os.path.basename)
def test_gpu_fft_vs_golden(sdrb, gpu_ctx, path):
    """Every field the reference-generated fixture holds, against the CUDA path through the C-ABI."""
    g = np.load(path)
    fmt, fs, n, focus, stream, blocks = (int(v) for v in g["meta"])
    iq = blocks_of(sdrb, fmt, fs, stream, blocks, n)
    cfg = sdrb.Config(430_000_000, fs, n, focus, inputFormat=fmt)
    plan = sdrb.plan_config(cfg)
    gpu_ctx.configure(cfg, 1, blocks)
    gpu_ctx.process(iq, blocks, sdrb.STAGE_FFT)
    res = gpu_ctx.results()[0]
    candidates = 0
    for b in range(blocks):
        power = gpu_ctx.power(0, b)
        if b < 2:
            spectrum_close(power, g, n, b)
        for k in DB_KEYS:
            assert abs(float(res[b][k]) - float(g[k][b])) < 0.01, (b, k)
        for k in SIGMA_KEYS:   # dB / sigma with sigma floors >= 0.5 (fft_process.cpp:242,268,285,300)
            assert abs(float(res[b][k]) - float(g[k][b])) < 0.02 + 1e-3 * abs(float(g[k][b])), (b, k)
        assert int(res[b]["detectionFlag"]) == int(g["detectionFlag"][b])
        assert int(res[b]["trackingFrequency"]) == int(g["trackingFrequency"][b])
        assert np.float32(res[b]["trackingFrequencyHz"]) == g["trackingFrequencyHz"][b]
        assert np.float32(res[b]["best1kHzCenterFreqHz"]) == g["best1kHzCenterFreqHz"][b]
        # maxPeakAndFrequency (fft_process.cpp:343-348) is the tracker's candidate: it moves only in a frame whose
        # peak beats it.  In those frames the reference's candidate IS this frame's peak: its float32 frequency
        # must come out bit-identical and its level within 0.01 dB (away from exact dB ties).
        moved = b == 0 or g["maxPeakHz"][b] != g["maxPeakHz"][b - 1] or g["maxPeakDb"][b] != g["maxPeakDb"][b - 1]
        f = np.sort(power[plan.focusLo:plan.focusHi + 1])[::-1]
        tie = len(f) >= 2 and (f[0] - f[1]) <= 4e-6 * f[0]
        if moved and int(res[b]["valid"]) and g["maxPeakDb"][b] > -130.0 and not tie:
            assert abs(float(res[b]["absPeakDb"]) - float(g["maxPeakDb"][b])) < 0.01, b
            assert np.float32(res[b]["peakFrequencyHz"]) == g["maxPeakHz"][b], b
            candidates += 1
    assert candidates >= 1


@pytest.mark.gpu
def test_gpu_spectrum_error_vs_fftw(sdrb, gpu_ctx):
    """Whole spectra of the reference build (FFTW float plan) against the GPU transform, bin class by bin class.
    north_star asks for 1e-5 relative power; a float32 transform cannot give that on WEAK bins (its error is
    ~eps * sqrt(total energy) in amplitude whatever the bin holds, FFTW's own included), so the bound that is
    asserted is: 1e-5 relative for every bin within 40 dB of the strongest, and the error floor below that.
    The measured worst cases are printed (pytest -s) and quoted in DESIGN.md section 3."""
    g = np.load(os.path.join(GOLD, "fft_u8_16k.npz"))
    fmt, fs, n, focus, stream, blocks = (int(v) for v in g["meta"])
    iq = blocks_of(sdrb, fmt, fs, stream, 2, n)
    gpu_ctx.configure(sdrb.Config(430_000_000, fs, n, focus, inputFormat=fmt), 1, 2)
    gpu_ctx.process(iq, 2, sdrb.STAGE_FFT)
    for b in range(2):
        ref, got = g["power_full"][b].astype(np.float64), gpu_ctx.power(0, b).astype(np.float64)
        rel = np.abs(got - ref) / np.maximum(ref, 1e-300)
        strong = ref >= 1e-4 * ref.max()
        total = ref.sum()
        floor = 2.0 * np.sqrt(ref) * 3e-7 * np.sqrt(total) + 1e-9 * total / n
        print(f"block {b}: {strong.sum()} bins within 40 dB of the peak, worst relative error {rel[strong].max():.2e}; "
              f"{(~strong).sum()} weaker bins, worst relative error {rel[~strong].max():.2e}, "
              f"worst error / float-FFT floor {(np.abs(got - ref)[~strong] / floor[~strong]).max():.3f}")
        assert rel[strong].max() <= 1e-5
        assert np.all(np.abs(got - ref)[~strong] <= 1e-5 * ref[~strong] + floor[~strong])
        d = np.abs(10 * np.log10(got[strong] + 1e-20) - 10 * np.log10(ref[strong] + 1e-20))
        assert d.max() < 1e-3


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
