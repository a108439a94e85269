"""GPU parity tests: the CUDA path, called through the C-ABI (include/sdr_b200.h), against the oracle
on identical seeded synthetic IQ.  Tolerances are BASELINE.json's: peak bin / frequencies bit-exact
away from ties, power 1e-5 relative (on top of the float-FFT error floor, stated below), scalars
0.01 dB, PCM +-1 LSB."""
import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
FC = 430_000_000


def make_iq(sdrb, fmt, fs, first_stream, n_streams, n_blocks, n, first_block=0):
    """[S, B, n, 2] wire-format blocks, contiguous in time per stream."""
    raw = sdrb.synth_host(fmt, fs, first_stream, n_streams, first_block * n, n_blocks * n)
    return np.ascontiguousarray(raw.reshape(n_streams, n_blocks, n, 2))


def power_tolerance(p_ref):
    # 1e-5 relative power + the float32 FFT error floor (amplitude error ~ eps * sqrt(total energy)),
    # the same bound tests/cpp/fft_sim.cpp uses against a double-precision DFT.
    total = float(p_ref.astype(np.float64).sum())
    return 1e-5 * p_ref + 2.0 * np.sqrt(p_ref) * 3e-7 * np.sqrt(total) + 1e-9 * total / p_ref.size


def test_synth_device_matches_host(sdrb, gpu_ctx):
    import torch
    for fmt in (sdrb.CU8, sdrb.CS16, sdrb.CF32):
        S, n, first = 3, 5000, 123_456_789_012
        host = sdrb.synth_host(fmt, 2_000_000, 7, S, first, n)
        dev = torch.empty(host.shape, dtype={0: torch.uint8, 1: torch.int16, 2: torch.float32}[fmt], device="cuda")
        gpu_ctx.synth_device(fmt, 2_000_000, 7, S, first, n, dev.data_ptr())
        gpu_ctx.sync()
        assert np.array_equal(dev.cpu().numpy(), host)


@pytest.mark.parametrize("fmt", [0, 1, 2])
@pytest.mark.parametrize("n,fs", [(16384, 2_000_000), (4096, 2_000_000), (65536, 30_720_000), (256, 250_000),
                                  (32768, 2_500_000), (2, 250_000), (8, 250_000), (512, 250_000), (1024, 1_000_000),
                                  (2048, 2_000_000), (8192, 2_000_000), (128, 250_000), (131072, 10_000_000),
                                  (262144, 10_000_000), (524288, 10_000_000)])
def test_power_spectrum_parity(sdrb, gpu_ctx, fmt, n, fs):
    S, B = 3, 2
    iq = make_iq(sdrb, fmt, fs, 11, S, B, n)
    gpu_ctx.configure(sdrb.Config(FC, fs, n, 5, inputFormat=fmt), S, B)
    gpu_ctx.process(iq, B, sdrb.STAGE_FFT)
    for s in range(S):
        for b in range(B):
            ref = ol.oracle_power(ol.to_cf32(iq[s, b], fmt))
            got = gpu_ctx.power(s, b)
            assert np.all(np.isfinite(got))
            assert np.all(np.abs(got - ref) <= power_tolerance(ref)), (s, b, np.abs(got - ref).max())
            # dB view (10*log10f) within 1e-3 dB for every bin above the float-FFT floor
            strong = ref > 1e-4 * ref.max()
            d = np.abs(10 * np.log10(got[strong] + 1e-20) - 10 * np.log10(ref[strong] + 1e-20))
            assert d.max() < 1e-3


def check_results(got, exp, n_frames_checked, where):
    """got: numpy record, exp: oracle dict.  Scalars within 0.01 dB; indices/frequencies exact."""
    for k in ("meanSnrDb", "peakAboveNoiseMeanDb", "maxBinSnrDb", "best1kHzSnrDb", "perBinMean", "absPeakDb",
              "signalPowerDb"):
        assert abs(float(got[k]) - exp[k]) < 0.01, (where, k, float(got[k]), exp[k])
    for k in ("meanSnrSigma", "maxBinSnrSigma", "best1kHzSnrSigma"):   # dB / sigma, sigma floors >= 0.5
        assert abs(float(got[k]) - exp[k]) < 0.02 + 1e-3 * abs(exp[k]), (where, k, float(got[k]), exp[k])
    assert int(got["peakBin"]) == exp["peakBin"], where
    assert int(got["valid"]) == exp["valid"], where
    assert int(got["detectionFlag"]) == exp["detectionFlag"], where
    assert np.float32(got["best1kHzCenterFreqHz"]) == np.float32(exp["best1kHzCenterFreqHz"]), where
    assert np.float32(got["trackingFrequencyHz"]) == np.float32(exp["trackingFrequencyHz"]), where
    assert int(got["trackingFrequency"]) == exp["trackingFrequency"], where


def tie_free(power, plan, margin=4e-6):
    """True when the focus-band maximum is unique beyond dB-float resolution."""
    f = np.sort(power[plan.focusLo:plan.focusHi + 1])[::-1]
    return len(f) < 2 or (f[0] - f[1]) > margin * f[0]


@pytest.mark.parametrize("fmt,n,fs,focus", [(0, 16384, 2_000_000, 5), (0, 16384, 2_000_000, 200),
                                            (1, 65536, 30_720_000, 5), (2, 4096, 2_500_000, 5),
                                            (0, 16384, 2_000_000, 300), (2, 16384, 2_000_000, 0),
                                            (0, 1024, 250_000, 5)])
def test_estimator_parity(sdrb, gpu_ctx, fmt, n, fs, focus):
    """Estimator + tracker against the oracle fed (a) the GPU's own power spectrum — isolates
    evaluateSignalStrength — and (b) the oracle's own DFT — the whole FFT path."""
    S, B = 4, 12
    cfg = sdrb.Config(FC, fs, n, focus, inputFormat=fmt)
    plan = sdrb.plan_config(cfg)
    iq = make_iq(sdrb, fmt, fs, 0, S, B, n)
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(iq, B, sdrb.STAGE_FFT)
    res = gpu_ctx.results()
    checked = 0
    for s in range(S):
        same, own = ol.OracleFft(FC, fs, n, focus), ol.OracleFft(FC, fs, n, focus)
        for b in range(B):
            p_gpu = gpu_ctx.power(s, b)
            e1 = same.evaluate(p_gpu)
            e2 = own.process(ol.to_cf32(iq[s, b], fmt))
            if plan.focusLen <= 0:
                assert int(res[s, b]["peakBin"]) == -1 and e1["peakBin"] == -1
                continue
            check_results(res[s, b], e1, checked, ("same-power", s, b))
            if tie_free(p_gpu, plan) and e1["peakBin"] == e2["peakBin"]:
                for k in ("meanSnrDb", "peakAboveNoiseMeanDb", "maxBinSnrDb", "best1kHzSnrDb", "perBinMean"):
                    assert abs(float(res[s, b][k]) - e2[k]) < 0.01, ("own-fft", s, b, k)
                checked += 1
            assert np.float32(res[s, b]["peakFrequencyHz"]) == np.float32(
                np.float32(np.float32(e1["peakBin"]) * np.float32(plan.freqPerBin))
                + np.float32(np.float32(FC) - np.float32(plan.nyquist)))
    if plan.focusLen > 0:
        assert checked >= S * B // 2


def test_tracking_follows_center_change(sdrb, gpu_ctx):
    n, fs, S, B = 16384, 2_000_000, 2, 50
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=0)
    iq = make_iq(sdrb, 0, fs, 0, S, 2 * B, n)
    gpu_ctx.configure(cfg, S, B)
    orc = [ol.OracleFft(FC, fs, n, 5) for _ in range(S)]
    tracked = set()
    for call in range(2):
        if call == 1:
            gpu_ctx.set_center_frequency(1, FC + 25_000)
            orc[1].set_center(FC + 25_000)
        blk = np.ascontiguousarray(iq[:, call * B:(call + 1) * B])
        gpu_ctx.process(blk, B, sdrb.STAGE_FFT)
        res = gpu_ctx.results()
        for s in range(S):
            for b in range(B):
                e = orc[s].evaluate(gpu_ctx.power(s, b))
                assert int(res[s, b]["trackingFrequency"]) == e["trackingFrequency"], (call, s, b)
                assert int(res[s, b]["detectionFlag"]) == e["detectionFlag"]
                tracked.add(e["trackingFrequency"])
    assert len(tracked) >= 3


@pytest.mark.parametrize("fmt,n,fs,mode,upper,B", [(0, 16384, 2_000_000, 1, 1, 9), (0, 16384, 2_000_000, 0, 1, 4),
                                                   (0, 16384, 2_500_000, 2, 1, 4), (0, 16384, 2_000_000, 1, 0, 3),
                                                   (1, 65536, 30_720_000, 1, 1, 5), (2, 16384, 2_000_000, 1, 1, 4),
                                                   (2, 4096, 250_000, 1, 1, 6), (1, 16384, 1_024_000, 1, 1, 3),
                                                   (0, 128, 2_000_000, 1, 1, 3),
                                                   (0, 160, 2_000_000, 1, 1, 5),    # not a multiple of 128: per-chunk state pass
                                                   (1, 4000, 2_000_000, 1, 1, 4)])
def test_ssb_pcm_parity(sdrb, gpu_ctx, fmt, n, fs, mode, upper, B):
    S = 5
    cfg = sdrb.Config(FC, fs, n, 5, soundMode=mode, inputFormat=fmt, upperSideband=upper)
    iq = make_iq(sdrb, fmt, fs, 20, S, B, n)
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(iq, B, sdrb.STAGE_SSB)
    pcm = gpu_ctx.pcm()
    m = gpu_ctx.pcm_per_block
    assert pcm.shape == (S, B, m)
    exact = total = 0
    for s in range(S):
        o = ol.OracleSsb()
        for b in range(B):
            exp = o.process(ol.to_cf32(iq[s, b], fmt), fs, upper, mode)
            assert len(exp) == m
            d = np.abs(pcm[s, b].astype(np.int32) - exp.astype(np.int32))
            assert d.max(initial=0) <= 1, (s, b, int(d.max()), int(np.argmax(d)))
            exact += int((d == 0).sum())
            total += m
            if not upper:
                assert not pcm[s, b].any()
    if total:
        assert exact >= 0.999 * total, (exact, total)   # the state hand-off is exact: expect bit-identical PCM
        if upper and m > 10:
            assert np.abs(pcm).max() > 50


def test_ssb_pcm_parity_on_extreme_cf32_input(sdrb, gpu_ctx):
    """cf32 input with bursts far outside any SDR's range (1e33, 1e-30, exact zeros): the AGC's sqrt / divide leave
    the range the branch-free sequences of k_ssb_main are used for and the library-sequence path takes over; the
    PCM must stay the oracle's (finite arithmetic throughout, so the reference's behaviour is defined)."""
    n, fs, S, B = 16384, 2_000_000, 3, 3
    cfg = sdrb.Config(FC, fs, n, 5, soundMode=1, inputFormat=2, upperSideband=1)
    iq = make_iq(sdrb, 2, fs, 40, S, B, n).copy()
    f = iq.view(np.float32).reshape(S, B, n, 2)
    rng = np.random.default_rng(3)
    for s in range(S):
        for b in range(B):
            at = rng.integers(0, n - 64, 6)
            f[s, b, at[0]:at[0] + 40] *= np.float32(1e33)      # |y| beyond 2^80 for a few hundred samples
            f[s, b, at[1]:at[1] + 9, 0] = np.float32(-3e33)
            f[s, b, at[2]:at[2] + 64] *= np.float32(1e-30)
            f[s, b, at[3]:at[3] + 64] = 0.0
            f[s, b, at[4], 0] = np.float32(3e37)
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(iq, B, sdrb.STAGE_SSB)
    pcm = gpu_ctx.pcm()
    worst = 0
    for s in range(S):
        o = ol.OracleSsb()
        for b in range(B):
            exp = o.process(ol.to_cf32(iq[s, b], 2), fs, 1, 1)
            d = np.abs(pcm[s, b].astype(np.int32) - exp.astype(np.int32))
            worst = max(worst, int(d.max(initial=0)))
    assert worst <= 1, worst


def test_calls_are_continuous_and_batch_independent(sdrb, gpu_ctx):
    """Blocks split over several calls == one call; stream i of a batch == the same stream alone;
    state save/load reproduces the continuation."""
    n, fs, S, B = 16384, 2_000_000, 3, 8
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=0)
    iq = make_iq(sdrb, 0, fs, 40, S, B, n)
    both = sdrb.STAGE_FFT | sdrb.STAGE_SSB
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(iq, B, both)
    res_all, pcm_all = gpu_ctx.results().copy(), gpu_ctx.pcm().copy()
    # split 3 + 5 with a checkpoint in between
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(np.ascontiguousarray(iq[:, :3]), 3, both)
    r1, p1 = gpu_ctx.results().copy(), gpu_ctx.pcm().copy()
    blob = gpu_ctx.save_state()
    gpu_ctx.process(np.ascontiguousarray(iq[:, 3:]), 5, both)
    r2, p2 = gpu_ctx.results().copy(), gpu_ctx.pcm().copy()
    assert np.array_equal(np.concatenate([r1, r2], 1), res_all)
    assert np.array_equal(np.concatenate([p1, p2], 1), pcm_all)
    gpu_ctx.reset_state()
    gpu_ctx.load_state(blob)
    gpu_ctx.process(np.ascontiguousarray(iq[:, 3:]), 5, both)
    assert np.array_equal(gpu_ctx.results(), r2) and np.array_equal(gpu_ctx.pcm(), p2)
    # stream 1 alone
    gpu_ctx.configure(cfg, 1, B)
    gpu_ctx.process(np.ascontiguousarray(iq[1:2]), B, both)
    assert np.array_equal(gpu_ctx.results()[0], res_all[1])
    assert np.array_equal(gpu_ctx.pcm()[0], pcm_all[1])


def test_device_input_equals_host_input(sdrb, gpu_ctx):
    import torch
    n, fs, S, B = 16384, 2_000_000, 2, 2
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=1)
    iq = make_iq(sdrb, 1, fs, 0, S, B, n)
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(iq, B)
    r_h, p_h = gpu_ctx.results().copy(), gpu_ctx.pcm().copy()
    gpu_ctx.reset_state()
    d = torch.from_numpy(iq).cuda()
    gpu_ctx.process(d, B)
    gpu_ctx.sync()
    assert np.array_equal(gpu_ctx.results(), r_h) and np.array_equal(gpu_ctx.pcm(), p_h)


def test_pipelined_host_call_equals_plain_call(sdrb, gpu_ctx):
    """sdrb_process_host (chunked H2D / kernels / D2H overlap) returns exactly what sdrb_process does."""
    n, fs, S, B = 16384, 2_000_000, 11, 3
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=0)
    iq = make_iq(sdrb, 0, fs, 60, S, B, n)
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(iq, B)
    res, pcm = gpu_ctx.results().copy(), gpu_ctx.pcm().copy()
    pw = gpu_ctx.power_all().copy()
    gpu_ctx.reset_state()
    r2 = np.zeros((S, B), dtype=sdrb.RESULT_DTYPE)
    p2 = np.zeros((S, B, n), dtype=np.float32)
    c2 = np.zeros((S, B, gpu_ctx.pcm_per_block), dtype=np.int16)
    gpu_ctx.process_host(iq, B, results=r2, power=p2, pcm=c2)
    assert np.array_equal(r2, res) and np.array_equal(p2, pw) and np.array_equal(c2, pcm)
    gpu_ctx.reset_state()
    gpu_ctx.process_host(iq, B, sdrb.STAGE_FFT, results=r2)      # outputs are optional
    assert np.array_equal(r2, res)


def test_error_behaviour(sdrb, gpu_ctx):
    cfg = sdrb.Config(FC, 2_000_000, 12000, 5, inputFormat=2)
    gpu_ctx.configure(cfg, 1, 1)
    iq = np.zeros((1, 1, 12000, 2), np.float32)
    with pytest.raises(sdrb.SdrbError):
        gpu_ctx.process(iq, 2, sdrb.STAGE_SSB)          # more blocks than configured
    with pytest.raises(sdrb.SdrbError):
        gpu_ctx.configure(sdrb.Config(FC, 0, 1024, 5), 1, 1)
    big = sdrb.Config(FC, 2_000_000, (1 << 22) + 2, 5, inputFormat=2)   # beyond the any-length transform's range
    gpu_ctx.configure(big, 1, 1)
    with pytest.raises(sdrb.SdrbError) as e:
        gpu_ctx.process(np.zeros((1, 1, (1 << 22) + 2, 2), np.float32), 1, sdrb.STAGE_FFT)
    assert e.value.code == -3
    fresh = sdrb.Context(0)
    with pytest.raises(sdrb.SdrbError) as e:
        fresh.process(iq, 1)
    assert e.value.code == -4
    fresh.close()


@pytest.mark.parametrize("fmt,n,fs", [(2, 12000, 2_000_000), (0, 10000, 2_000_000), (1, 5000, 2_500_000),
                                      (2, 12001, 2_000_000), (0, 4099, 1_000_000), (2, 1000, 250_000), (2, 3, 250_000),
                                      (2, 1, 250_000)])
def test_any_block_length(sdrb, gpu_ctx, fmt, n, fs):
    """fftwf_plan_dft_1d takes every length (fft_process.cpp:77) and processSSB_opt every block size: lengths that
    are not powers of two run the chirp-z transform over the power-of-two passes, SSB blocks whose bytes are not a
    multiple of 64 the scalar pass.  Same checks as for the fast paths (an odd length leaves its last bin at zero,
    as the reference's half-swap loop does); the float transform's error floor is that of the 2N-point convolution."""
    S, B = 3, 3
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=fmt)
    plan = sdrb.plan_config(cfg)
    iq = make_iq(sdrb, fmt, fs, 5, S, B, n)
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(iq, B, sdrb.STAGE_FFT | sdrb.STAGE_SSB)
    res, pcm = gpu_ctx.results(), gpu_ctx.pcm()
    assert pcm.shape == (S, B, plan.pcmPerBlock)
    for s in range(S):
        of, os_ = ol.OracleFft(FC, fs, n, 5), ol.OracleSsb()
        for b in range(B):
            f32 = ol.to_cf32(iq[s, b], fmt)
            ref, got = ol.oracle_power(f32), gpu_ctx.power(s, b)
            total = float(ref.astype(np.float64).sum())
            tol = 1e-5 * ref + 3.0 * (2.0 * np.sqrt(ref) * 3e-7 * np.sqrt(total) + 1e-9 * total / max(n, 1))
            assert np.all(np.abs(got - ref) <= tol), (s, b, float(np.abs(got - ref).max()))
            if n & 1:
                assert got[n - 1] == 0.0
            if plan.focusLen > 0:
                check_results(res[s, b], of.evaluate(got), 0, ("any-n", s, b))
            exp = os_.process(f32, fs, 1, 1)
            assert exp.size == pcm.shape[2]
            if exp.size:
                assert np.abs(pcm[s, b].astype(int) - exp.astype(int)).max() <= 1, (s, b)
                assert (pcm[s, b] == exp).mean() >= 0.999


def test_full_size_properties_c2(sdrb, gpu_ctx):
    """BASELINE config 2 at full size (1024 u8 streams x 16384): Parseval per frame, peak bin at the
    keyed tone of the streams that are keyed on, and a checksum that is stable across two runs."""
    import torch
    S, n, fs = 1024, 16384, 2_000_000
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=0)
    gpu_ctx.configure(cfg, S, 1)
    d = torch.empty((S, 1, n, 2), dtype=torch.uint8, device="cuda")
    gpu_ctx.synth_device(0, fs, 0, S, 0, n, d.data_ptr())
    gpu_ctx.process(d, 1, sdrb.STAGE_FFT)
    gpu_ctx.sync()
    res = gpu_ctx.results()
    x = (d.to(torch.float64) - 127.4) / 128.0
    energy = (x * x).sum(dim=(1, 2, 3)).cpu().numpy() * n
    got = np.array([gpu_ctx.power(s, 0).astype(np.float64).sum() for s in range(0, S, 37)])
    assert np.allclose(got, energy[::37], rtol=2e-5)
    plan = sdrb.plan_config(cfg)
    keyed = [s for s in range(0, S, 4)]          # streams with s % 4 == 0 start keyed on
    host = sdrb.synth_host(0, fs, 0, 1, 0, 8)    # touch the host generator too
    assert np.array_equal(host[0], d[0, 0, :8].cpu().numpy())
    for s in keyed[:64]:
        assert plan.focusLo <= int(res[s, 0]["peakBin"]) <= plan.focusHi
        assert float(res[s, 0]["peakAboveNoiseMeanDb"]) > 30.0
    chk1 = res["peakBin"].astype(np.int64).sum()
    gpu_ctx.reset_state()
    gpu_ctx.process(d, 1, sdrb.STAGE_FFT)
    assert gpu_ctx.results()["peakBin"].astype(np.int64).sum() == chk1


@pytest.mark.parametrize("fmt", [2, 0])
def test_power_spectrum_parity_1m(sdrb, gpu_ctx, fmt):
    """BASELINE config C5: 1 048 576-point transform (multi-CTA four-step path), estimator included."""
    n, fs, S, B = 1 << 20, 10_000_000, 2, 2
    iq = make_iq(sdrb, fmt, fs, 3, S, B, n)
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=fmt)
    assert sdrb.plan_config(cfg).fftPath == 2
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(iq, B, sdrb.STAGE_FFT)
    res = gpu_ctx.results()
    for s in range(S):
        orc = ol.OracleFft(FC, fs, n, 5)
        for b in range(B):
            ref = ol.oracle_power(ol.to_cf32(iq[s, b], fmt))
            got = gpu_ctx.power(s, b)
            assert np.all(np.abs(got - ref) <= power_tolerance(ref)), (s, b, float(np.abs(got - ref).max()))
            check_results(res[s, b], orc.evaluate(got), 0, ("1M", s, b))


def test_ssb_one_reading_per_call_is_one_kernel_and_exact(sdrb, gpu_ctx):
    """One reading per stream per call: the whole of processSSB_opt is ONE kernel (no state pass, audio chain on the
    lane); the carried filter states go from call to call.  PCM against the oracle, call after call, and against the
    same readings handed over in one multi-reading call (state pass + fused pass + audio pass)."""
    n, fs, S, B = 16384, 2_000_000, 37, 6
    for fmt, mode in ((0, 1), (1, 0), (2, 2)):
        cfg = sdrb.Config(FC, fs, n, 5, soundMode=mode, inputFormat=fmt)
        iq = make_iq(sdrb, fmt, fs, 100, S, B, n)
        gpu_ctx.configure(cfg, S, B)
        gpu_ctx.process(iq, B, sdrb.STAGE_SSB)
        pcm_multi = gpu_ctx.pcm().copy()
        gpu_ctx.reset_state()
        orc = [ol.OracleSsb() for _ in range(S)]
        for b in range(B):
            l0 = gpu_ctx.launch_count
            gpu_ctx.process(np.ascontiguousarray(iq[:, b:b + 1]), 1, sdrb.STAGE_SSB)
            assert gpu_ctx.launch_count - l0 == 1, "one reading per stream must be a single kernel"
            pcm = gpu_ctx.pcm()
            assert np.array_equal(pcm[:, 0], pcm_multi[:, b]), (fmt, b)
            for s in (0, 17, 36):
                exp = orc[s].process(ol.to_cf32(iq[s, b], fmt), fs, 1, mode)
                assert np.abs(pcm[s, 0].astype(int) - exp.astype(int)).max() <= 1
                assert (pcm[s, 0] == exp).mean() >= 0.999


def test_ssb_fused_pass_equals_split_pass(sdrb, gpu_ctx, monkeypatch):
    """The fused AGC + decimator pass (rotated tap table, accumulators on the lane) against the split pair
    (post-AGC samples through a scratch, one thread per output): bit-identical PCM, several decimation factors."""
    for fmt, n, fs in ((0, 16384, 2_000_000), (0, 16384, 2_500_000), (0, 16384, 3_200_000), (1, 65536, 30_720_000),
                       (2, 16384, 1_600_000), (1, 16384, 1_024_000), (0, 16384, 1_536_000)):
        S, B = 33, 3
        cfg = sdrb.Config(FC, fs, n, 5, inputFormat=fmt)
        iq = make_iq(sdrb, fmt, fs, 7, S, B, n)
        out = []
        for fused in ("1", "0"):
            monkeypatch.setenv("SDRB_SSB_FUSED", fused)
            gpu_ctx.configure(cfg, S, B)
            gpu_ctx.process(iq, B, sdrb.STAGE_SSB)
            out.append(gpu_ctx.pcm().copy())
        assert out[0].size and np.array_equal(out[0], out[1]), (fmt, n, fs)
    monkeypatch.delenv("SDRB_SSB_FUSED")


def test_fft_and_ssb_side_by_side_equal_one_after_the_other(sdrb, gpu_ctx, monkeypatch):
    """FFT chain and SSB chain on two CUDA streams (the default) against both on one stream: same records, same
    spectra, same PCM."""
    n, fs, S, B = 16384, 2_000_000, 300, 2
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=0)
    iq = make_iq(sdrb, 0, fs, 0, S, B, n)
    out = []
    for fork in ("1", "0"):
        monkeypatch.setenv("SDRB_SSB_FORK", fork)
        ctx = sdrb.Context(0)            # the switch is read when the context is created
        ctx.configure(cfg, S, B)
        ctx.process(iq, B, sdrb.STAGE_FFT | sdrb.STAGE_SSB)
        out.append((ctx.results().copy(), ctx.pcm().copy(), ctx.power_all().copy()))
        ctx.close()
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])
    assert np.array_equal(out[0][2], out[1][2])


def test_stream_shards_equal_the_whole(sdrb, gpu_ctx):
    """SURVEY.md section 4 layer 4 (same results at 1 / 2 / 4 / 8 GPUs): the records, spectra and PCM of global stream
    s are the same whether s is processed in one context holding every stream or in the shard context of its rank
    (bench.py's sharding, two shard contexts standing in for two ranks on this one GPU)."""
    n, fs, S, B, world = 16384, 2_000_000, 10, 3, 2
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=0)
    both = sdrb.STAGE_FFT | sdrb.STAGE_SSB
    iq = make_iq(sdrb, 0, fs, 0, S, B, n)
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(iq, B, both)
    res, pcm, pw = gpu_ctx.results().copy(), gpu_ctx.pcm().copy(), gpu_ctx.power_all().copy()
    for rank in range(world):
        first, count = sdrb.shard_streams(S, world, rank)
        shard = sdrb.Context(0)
        shard.configure(cfg, count, B)
        mine = make_iq(sdrb, 0, fs, first, count, B, n)          # a rank generates / receives only its own streams
        shard.process(mine, B, both)
        assert np.array_equal(shard.results(), res[first:first + count]), rank
        assert np.array_equal(shard.pcm(), pcm[first:first + count]), rank
        assert np.array_equal(shard.power_all(), pw[first:first + count]), rank
        shard.close()


def test_replan_keeps_every_carried_state(sdrb, gpu_ctx):
    """sdrb_replan: a new sample rate for the same streams; the SSB filter states (function statics in the
    reference) and the FFT tracking state go on: the continuation equals the oracle's, which is never reset."""
    n, S = 16384, 3
    orc = [ol.OracleSsb() for _ in range(S)]
    gpu_ctx.configure(sdrb.Config(FC, 2_000_000, n, 5, inputFormat=2), S, 2)
    for step, fs in enumerate((2_000_000, 2_400_000, 2_000_000)):
        cfg = sdrb.Config(FC, fs, n, 5, inputFormat=2)
        if step:
            gpu_ctx.replan(cfg, 2)
        iq = make_iq(sdrb, 2, fs, 50, S, 2, n, first_block=2 * step)
        gpu_ctx.process(iq, 2, sdrb.STAGE_SSB)
        pcm = gpu_ctx.pcm()
        for s in range(S):
            for b in range(2):
                exp = orc[s].process(ol.to_cf32(iq[s, b], 2), fs, 1, 1)
                assert exp.size == pcm.shape[2]
                assert np.abs(pcm[s, b].astype(int) - exp.astype(int)).max() <= 1, (step, s, b)


def test_db_view_of_the_spectrum(sdrb, gpu_ctx):
    """north_star: '|X|^2 to dB'.  The dB view (10*log10f(p + 1e-20f), fft_process.cpp:149) of every bin within 1e-3 dB
    of the same expression on the linear output; the half-precision view of the host-to-host call within its own
    quantisation (0.0625 dB at 64..128 dB); the linear payload is untouched by either."""
    n, fs, S, B = 16384, 2_000_000, 6, 2
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=0)
    iq = make_iq(sdrb, 0, fs, 0, S, B, n)
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(iq, B, sdrb.STAGE_FFT)
    lin = gpu_ctx.power_all().copy()
    for s, b in ((0, 0), (5, 1)):
        want = (np.float32(10.0) * np.log10(lin[s, b] + np.float32(1e-20))).astype(np.float32)
        assert np.abs(gpu_ctx.power_db(s, b) - want).max() < 1e-3
        ref = ol.oracle_power(ol.to_cf32(iq[s, b], 0))
        strong = ref > 1e-4 * ref.max()
        assert np.abs(gpu_ctx.power_db(s, b)[strong] - 10 * np.log10(ref[strong] + 1e-20)).max() < 1e-3
    gpu_ctx.reset_state()
    r = np.zeros((S, B), dtype=sdrb.RESULT_DTYPE)
    h16 = np.zeros((S, B, n), dtype=np.float16)
    gpu_ctx.process_host(iq, B, sdrb.STAGE_FFT, results=r, power=h16, power_format=sdrb.POWER_DB_F16)
    want = 10.0 * np.log10(lin.astype(np.float64) + 1e-20)
    assert np.abs(h16.astype(np.float64) - want).max() <= 0.0625 * 0.5 + 1e-3 + np.abs(want).max() * 2.0 ** -11
    f32 = np.zeros((S, B, n), dtype=np.float32)
    gpu_ctx.reset_state()
    gpu_ctx.process_host(iq, B, sdrb.STAGE_FFT, results=r, power=f32, power_format=sdrb.POWER_DB_F32)
    assert np.abs(f32.astype(np.float64) - want).max() < 1e-3
    assert np.array_equal(gpu_ctx.power_all(), lin)


@pytest.mark.parametrize("n,fmt,S,B,hooks", [(65536, 1, 3, 5, {}), (65536, 0, 5, 4, {"SDRB_FB_DELAY": "2", "SDRB_FB_SLOTS": "4"}),
                                             (65536, 2, 2, 3, {"SDRB_FB_DELAY": "1", "SDRB_FB_SLOTS": "2"}),
                                             (1048576, 2, 2, 2, {}), (1048576, 1, 3, 2, {"SDRB_FB_DELAY": "1", "SDRB_FB_SLOTS": "2"})])
def test_four_step_fused_kernel_equals_the_pair(sdrb, monkeypatch, n, fmt, S, B, hooks):
    """SDRB_FB_FUSED=1: both four-step passes in one persistent kernel, the intermediate in a ring that stays in L2,
    work handed out by tickets with per-frame counters (fft_big.cu).  Same phase functions, so every bit of every
    spectrum must equal the two-kernel path's - with the default ring and with a ring of 2-4 frames and a delay of
    1-2 steps, where a column tile has to wait for its ring slot and a row group for its frame."""
    fs = 30_720_000 if n == 65536 else 10_000_000
    cfg = sdrb.Config(FC, fs, n, 25, inputFormat=fmt)
    iq = make_iq(sdrb, fmt, fs, 3, S, B, n)
    out = []
    for fused in ("0", "1"):
        monkeypatch.setenv("SDRB_FB_FUSED", fused)
        for k, v in hooks.items():
            monkeypatch.setenv(k, v)
        ctx = sdrb.Context(0)                # the switch is read when the FFT buffers are allocated
        try:
            ctx.configure(cfg, S, B)
            for call in range(2):            # the ticket and the counters are reset per call
                ctx.process(iq, B, sdrb.STAGE_FFT)
            out.append((ctx.power_all().copy().view(np.uint32), ctx.results().copy()))
        finally:
            ctx.close()
    assert out[0][0].size == S * B * n and np.array_equal(out[0][0], out[1][0])
    assert np.array_equal(out[0][1], out[1][1])


def test_full_size_properties_default_workload(sdrb):
    """The bench's default workload at full size (65536 u8 streams x one reading of 16384, FFT + peak + strength + USB
    demod, the fused one-kernel SSB pass beside the FFT chain): 24 streams spread over the batch against the oracle
    (records from the GPU's own spectrum, PCM bit for bit), Parseval on the same streams, the first 1000 streams
    against a 1000-stream context (batch independence at full size), and checksums stable across two runs."""
    import torch
    S, n, fs = 65536, 16384, 2_000_000
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=0)
    ctx = sdrb.Context(0)
    small = sdrb.Context(0)
    try:
        ctx.configure(cfg, S, 1)
        d = torch.empty((S, 1, n, 2), dtype=torch.uint8, device="cuda")
        ctx.synth_device(0, fs, 0, S, 0, n, d.data_ptr())
        st = sdrb.STAGE_FFT | sdrb.STAGE_SSB
        ctx.process(d, 1, st)
        ctx.sync()
        res, pcm = ctx.results().copy(), ctx.pcm().copy()
        m = ctx.pcm_per_block
        assert pcm.shape == (S, 1, m) and m > 300
        picks = [(k * 2731 + 5) % S for k in range(24)]
        for s in picks:
            raw = d[s, 0].cpu().numpy()
            x = ol.to_cf32(raw, 0)
            exp = ol.OracleSsb().process(x, fs, 1, 1)
            assert np.array_equal(pcm[s, 0], exp), s                       # first block from a zero state: exact
            p = ctx.power(s, 0)
            check_results(res[s, 0], ol.OracleFft(FC, fs, n, 5).evaluate(p), 0, ("default", s))
            e = float((x.astype(np.float64) ** 2).sum()) * n
            assert abs(float(p.astype(np.float64).sum()) - e) <= 2e-5 * e
        small.configure(cfg, 1000, 1)
        small.process(d[:1000], 1, st)
        small.sync()
        assert np.array_equal(small.results(), res[:1000]) and np.array_equal(small.pcm(), pcm[:1000])
        chk = (int(res["peakBin"].astype(np.int64).sum()), int(pcm.astype(np.int64).sum()),
               int(np.bitwise_xor.reduce(pcm.reshape(-1).view(np.uint16).astype(np.uint32))))
        ctx.reset_state()
        ctx.process(d, 1, st)
        ctx.sync()
        res2, pcm2 = ctx.results(), ctx.pcm()
        assert chk == (int(res2["peakBin"].astype(np.int64).sum()), int(pcm2.astype(np.int64).sum()),
                       int(np.bitwise_xor.reduce(pcm2.reshape(-1).view(np.uint16).astype(np.uint32))))
    finally:
        small.close()
        ctx.close()


@pytest.mark.parametrize("n,S,B", [(512, 4001, 3), (4096, 2500, 2), (8192, 1300, 1), (32768, 150, 2)])
def test_small_fft_kernels_over_many_rounds(sdrb, gpu_ctx, n, S, B):
    """The persistent one-CTA kernels (256 .. 8192 points) and the two-pass scheme on batches that take every CTA
    through many rounds, with a frame count that is not a multiple of the transforms per CTA: 24 frames spread over
    the batch against the oracle, and every frame against the Parseval sum."""
    fs = 2_000_000
    iq = make_iq(sdrb, 0, fs, 5, S, B, n)
    gpu_ctx.configure(sdrb.Config(FC, fs, n, 5, inputFormat=0), S, B)
    gpu_ctx.process(iq, B, sdrb.STAGE_FFT)
    pw = gpu_ctx.power_all().reshape(S, B, n)
    x = (iq.astype(np.float64) - 127.4) / 128.0
    energy = (x * x).sum(axis=(2, 3)) * n
    assert np.allclose(pw.astype(np.float64).sum(axis=2), energy, rtol=2e-5)
    for k in range(24):
        s, b = (k * 977 + 3) % S, k % B
        ref = ol.oracle_power(ol.to_cf32(iq[s, b], 0))
        assert np.all(np.abs(pw[s, b] - ref) <= power_tolerance(ref)), (s, b)
