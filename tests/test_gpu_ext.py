"""GPU parity of the EXTENSION stages (sdrb_config.fftWindow / fftRemoveDc / ssbTrueSideband /
ssbChannels / ssbChannelOffsetHz / ssbChannelSpacingHz) against their CPU definition in
oracle/sdr_oracle_ext.c.  The reference has no code for these stages, so this
parity is "unpinned" by construction; identity settings are covered by test_gpu_parity.py."""
import numpy as np
import pytest

import oracle_lib as ol
from test_gpu_parity import FC, make_iq, power_tolerance

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("fmt", [0, 1, 2])
@pytest.mark.parametrize("n,fs", [(16384, 2_000_000), (4096, 2_000_000), (65536, 30_720_000), (256, 250_000), (1024, 1_000_000),
                                  (8192, 2_000_000), (32768, 2_500_000)])
@pytest.mark.parametrize("window,remove_dc", [(1, 0), (0, 1), (1, 1)])
def test_fft_front_end(sdrb, gpu_ctx, fmt, n, fs, window, remove_dc):
    S, B = 2, 3
    iq = make_iq(sdrb, fmt, fs, 31, S, B, n)
    gpu_ctx.configure(sdrb.Config(FC, fs, n, 5, inputFormat=fmt, fftWindow=window, fftRemoveDc=remove_dc), S, B)
    gpu_ctx.process(iq, B, sdrb.STAGE_FFT)
    for s in range(S):
        for b in range(B):
            x = ol.to_cf32(iq[s, b], fmt)
            ref = ol.oracle_power(ol.oracle_fft_front(x, window, remove_dc))
            got = gpu_ctx.power(s, b)
            # the float tree-summed block mean differs from the double one by a few 1e-6 relative
            # (measured 3e-6 at n = 65536): an extra n * |mean| * 5e-6 of amplitude in the bins the
            # window spreads DC into
            plain = ol.oracle_power(x)
            dc_amp = np.sqrt(plain[n // 2]) * 5e-6 if remove_dc else 0.0
            tol = power_tolerance(ref) + 2.0 * np.sqrt(ref) * dc_amp + dc_amp * dc_amp
            assert np.all(np.abs(got - ref) <= tol), (s, b, float(np.abs(got - ref).max()))
            if remove_dc and not window:
                assert got[n // 2] <= 1e-6 * plain[n // 2] + tol[n // 2]


@pytest.mark.parametrize("fmt,n,fs,mode,upper,tsb,C,off,spacing,B", [
    (0, 16384, 2_000_000, 1, 1, 0, 1, 12_500.0, 0.0, 5),          # NCO offset only
    (0, 16384, 2_000_000, 1, 1, 0, 3, -25_000.0, 25_000.0, 4),    # channel bank, reference demod
    (0, 16384, 2_000_000, 1, 1, 1, 1, 0.0, 0.0, 6),               # true USB at the centre
    (1, 65536, 30_720_000, 1, 0, 1, 2, 100_000.0, 50_000.0, 3),   # C3-like: cs16, 64K blocks, true LSB
    (2, 16384, 2_000_000, 2, 0, 1, 5, -50_000.0, 25_000.0, 33),   # cf32, true LSB, 5 channels, > 32 blocks
    (0, 16384, 4_800_000, 1, 1, 1, 2, 30_000.0, 40_000.0, 3),     # decim 100: fused decimator, four accumulators
    (2, 16384, 9_600_000, 1, 1, 0, 2, -60_000.0, 80_000.0, 3),    # decim 200: two accumulators, reference demod
    (1, 16384, 2_000_000, 1, 1, 1, 1, 5_000.0, 0.0, 3),           # cs16 at decim 41: the split AGC / decimator pair
])
def test_ssb_extended_chain(sdrb, gpu_ctx, fmt, n, fs, mode, upper, tsb, C, off, spacing, B):
    S = 3
    cfg = sdrb.Config(FC, fs, n, 5, soundMode=mode, inputFormat=fmt, upperSideband=upper, ssbTrueSideband=tsb,
                      ssbChannels=C, ssbChannelOffsetHz=off, ssbChannelSpacingHz=spacing)
    iq = make_iq(sdrb, fmt, fs, 40, S, 2 * B, n)
    gpu_ctx.configure(cfg, S, B)
    orc = [[ol.OracleSsbx(np.float32(off).item() + c * np.float32(spacing).item(), tsb) for c in range(C)]
           for _ in range(S)]
    m = gpu_ctx.pcm_per_block
    worst = 0
    for call in range(2):   # two calls: filter state and NCO phase carry across calls
        blk = np.ascontiguousarray(iq[:, call * B:(call + 1) * B])
        gpu_ctx.process(blk, B, sdrb.STAGE_SSB)
        pcm = gpu_ctx.pcm().reshape(S, C, B, m)
        for s in range(S):
            for c in range(C):
                for b in range(B):
                    exp = orc[s][c].process(ol.to_cf32(blk[s, b], fmt), fs, upper, mode)
                    assert len(exp) == m
                    d = np.abs(pcm[s, c, b].astype(np.int32) - exp.astype(np.int32)).max(initial=0)
                    worst = max(worst, int(d))
                    assert d <= 1, (call, s, c, b, int(d))
    assert np.abs(pcm).max() > 100   # audio is really there


def test_ssb_extended_state_round_trip(sdrb, gpu_ctx):
    fs, n, S, B, C = 2_000_000, 16384, 2, 3, 2
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=0, ssbTrueSideband=1, ssbChannels=C, ssbChannelOffsetHz=10_000.0,
                      ssbChannelSpacingHz=20_000.0)
    iq = make_iq(sdrb, 0, fs, 50, S, 2 * B, n)
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.process(np.ascontiguousarray(iq[:, :B]), B, sdrb.STAGE_SSB)
    blob = gpu_ctx.save_state()
    gpu_ctx.process(np.ascontiguousarray(iq[:, B:]), B, sdrb.STAGE_SSB)
    want = gpu_ctx.pcm().copy()
    gpu_ctx.configure(cfg, S, B)
    gpu_ctx.load_state(blob)
    gpu_ctx.process(np.ascontiguousarray(iq[:, B:]), B, sdrb.STAGE_SSB)
    assert np.array_equal(gpu_ctx.pcm(), want)


@pytest.mark.parametrize("stages", ["both", "ssb"])
def test_ssb_extended_chain_host_call_equals_plain_call(sdrb, gpu_ctx, stages):
    """sdrb_process_host with the extended chain: the chunks carry the copies and the FFT chain, the SSB chain runs
    once over all streams behind the last chunk (its state pass costs the same for 3 streams as for 300) - the
    payloads must be those of the plain call, over two calls (carried filter state and NCO phase)."""
    fs, n, S, B, C = 2_000_000, 16384, 11, 3, 3
    cfg = sdrb.Config(FC, fs, n, 5, inputFormat=0, ssbTrueSideband=1, ssbChannels=C, ssbChannelOffsetHz=-25_000.0,
                      ssbChannelSpacingHz=25_000.0)
    st = (sdrb.STAGE_FFT | sdrb.STAGE_SSB) if stages == "both" else sdrb.STAGE_SSB
    iq = make_iq(sdrb, 0, fs, 70, S, 2 * B, n)
    gpu_ctx.configure(cfg, S, B)
    m = gpu_ctx.pcm_per_block
    want = []
    for call in range(2):
        gpu_ctx.process(np.ascontiguousarray(iq[:, call * B:(call + 1) * B]), B, st)
        want.append((gpu_ctx.pcm().copy(), gpu_ctx.results().copy() if stages == "both" else None,
                     gpu_ctx.power_all().copy() if stages == "both" else None))
    gpu_ctx.configure(cfg, S, B)            # fresh state
    for call in range(2):
        r2 = np.zeros((S, B), dtype=sdrb.RESULT_DTYPE)
        p2 = np.zeros((S, B, n), dtype=np.float32)
        c2 = np.zeros((S, C, B, m), dtype=np.int16)
        blk = np.ascontiguousarray(iq[:, call * B:(call + 1) * B])
        if stages == "both":
            gpu_ctx.process_host(blk, B, st, results=r2, power=p2, pcm=c2)
            assert np.array_equal(r2, want[call][1]) and np.array_equal(p2.reshape(-1), want[call][2].reshape(-1))
        else:
            gpu_ctx.process_host(blk, B, st, pcm=c2)
        assert np.array_equal(c2.reshape(-1), want[call][0].reshape(-1)), call
    assert np.abs(c2).max() > 100
