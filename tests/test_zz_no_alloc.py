"""sdrb_estimate_noalloc: the allocation-free entry point (no counterpart in the reference).
Its kernel scratch is a static device array, so the checks are: the records equal what the
existing entry point (sdrb_process + sdrb_get_results) gives for the same spectra, on two calls in a row that share the
array, and sizes beyond the stated maximum are rejected.  Collected last on purpose;
runs alone in a fresh process or after the other tests, with a context of its own."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
FC, FS, N, FOCUS = 430_000_000, 2_000_000, 16384, 5
SAME = ("meanSnrDb", "meanSnrSigma", "peakAboveNoiseMeanDb", "maxBinSnrDb", "maxBinSnrSigma", "best1kHzSnrDb",
        "best1kHzSnrSigma", "perBinMean", "absPeakDb", "signalPowerDb", "peakBin", "valid")


def test_estimate_noalloc_equals_existing_entry_point(sdrb):
    import torch
    S, B = 40, 3
    ctx = sdrb.Context(0)
    try:
        ctx.configure(sdrb.Config(FC, FS, N, FOCUS, inputFormat=sdrb.CU8), S, B)
        out = torch.zeros((S * B * sdrb.ESTIMATE_DTYPE.itemsize,), dtype=torch.uint8, device="cuda")   # caller-owned
        for call in range(2):                       # two calls in a row share the static scratch
            iq = np.ascontiguousarray(sdrb.synth_host(sdrb.CU8, FS, 100 * call, S, call * B * N, B * N).reshape(S, B, N, 2))
            ctx.process(iq, B, sdrb.STAGE_FFT)
            ctx.sync()
            want = ctx.results().reshape(-1)
            out.fill_(0xA5 if call else 0)          # whatever the last call left must not matter
            torch.cuda.synchronize()
            sdrb.estimate_noalloc(0, FS, N, FOCUS, ctx.power_device_ptr(), S * B, out.data_ptr(), 0)
            torch.cuda.synchronize()
            got = out.cpu().numpy().view(sdrb.ESTIMATE_DTYPE)
            assert int(want["valid"].min()) == 1    # default geometry: ten reference windows
            for k in SAME:                          # same kernels on the same spectra: bit-identical
                assert np.array_equal(got[k], want[k]), (call, k)
            assert (got["best1kStart"] >= 0).all()
    finally:
        ctx.close()


def test_estimate_noalloc_rejects_what_the_static_scratch_cannot_hold(sdrb):
    import torch
    p = torch.zeros((8, N), dtype=torch.float32, device="cuda")
    o = torch.zeros((8 * sdrb.ESTIMATE_DTYPE.itemsize,), dtype=torch.uint8, device="cuda")
    with pytest.raises(sdrb.SdrbError) as e:
        sdrb.estimate_noalloc(0, FS, N, FOCUS, p.data_ptr(), sdrb.NOALLOC_MAX_FRAMES + 1, o.data_ptr(), 0)
    assert e.value.code == -1 and "SDRB_NOALLOC_MAX_FRAMES" in str(e.value)
    with pytest.raises(sdrb.SdrbError) as e:     # 200 kHz focus pools far more bins than the shared-memory pool
        sdrb.estimate_noalloc(0, FS, 1 << 20, 200, p.data_ptr(), 1, o.data_ptr(), 0)
    assert e.value.code == -3
    with pytest.raises(sdrb.SdrbError):
        sdrb.estimate_noalloc(0, FS, N, FOCUS, 0, 8, o.data_ptr(), 0)
    # a legal call afterwards still works (all-zero spectra: finite records, peak at the first focus bin)
    sdrb.estimate_noalloc(0, FS, N, FOCUS, p.data_ptr(), 8, o.data_ptr(), 0)
    torch.cuda.synchronize()
    got = o.cpu().numpy().view(sdrb.ESTIMATE_DTYPE)
    assert np.isfinite(got["meanSnrDb"]).all() and (got["peakBin"] == sdrb.plan_config(sdrb.Config(FC, FS, N, FOCUS)).focusLo).all()
