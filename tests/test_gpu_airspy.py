"""GPU parity of the Airspy real->IQ front end (inputFormat SDRB_RU16) through
the C-ABI: raw 12-bit ADC words in, the FFT path's spectrum / records and the SSB path's PCM out, against
oracle/sdr_oracle_airspy.c (pinned bit-exactly against the reference's iqconverter_float.c) followed by
the pinned FFT / SSB oracle.  The converter is bit-exact, so downstream tolerances are the CF32 path's."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from test_oracle_airspy import AirspyState, oracle_airspy, synth_adc

pytestmark = pytest.mark.gpu


def _oracle_iq(raw_blocks):
    """raw_blocks: [B, 2n] uint16 of ONE stream -> [B, n, 2] float32, state carried over the blocks."""
    o = oracle_airspy()
    st = AirspyState()
    o.sdr_oracle_airspy_init(C.byref(st))
    out = np.zeros(raw_blocks.shape, np.float32)
    for b in range(raw_blocks.shape[0]):
        chunk = np.ascontiguousarray(raw_blocks[b])
        o.sdr_oracle_airspy_process(C.byref(st), chunk.ctypes.data, chunk.size, out[b].ctypes.data)
    return out.reshape(raw_blocks.shape[0], -1, 2)


@pytest.mark.parametrize("n,fs,S,B,calls", [(16384, 10_000_000, 3, 4, 3), (65536, 10_000_000, 2, 2, 2), (4096, 2_500_000, 2, 8, 2)])
def test_airspy_fft_path_matches_oracle(sdrb, gpu_ctx, n, fs, S, B, calls):
    fc = 430_000_000
    cfg = sdrb.Config(fc, fs, n, 25, inputFormat=sdrb.RU16)
    gpu_ctx.configure(cfg, S, B)
    raw = np.stack([synth_adc(2 * n * B * calls, 100 + s, dc=20 + 7 * s) for s in range(S)]).reshape(S, calls, B, 2 * n)
    iq_ref = np.stack([_oracle_iq(raw[s].reshape(calls * B, 2 * n)) for s in range(S)]).reshape(S, calls, B, n, 2)
    for c in range(calls):
        gpu_ctx.process(np.ascontiguousarray(raw[:, c].reshape(S, B, n, 2)), B, sdrb.STAGE_FFT)
        res = gpu_ctx.results()
        for s in range(S):
            of = ol.OracleFft(fc, fs, n, 25) if c == 0 else None
            for b in range(B):
                ref_p = ol.oracle_power(np.ascontiguousarray(iq_ref[s, c, b]))
                got_p = gpu_ctx.power(s, b)
                total = float(ref_p.astype(np.float64).sum())
                tol = 1e-5 * ref_p + 6e-7 * np.sqrt(ref_p) * np.sqrt(total) + 1e-9 * total / n
                assert np.all(np.abs(got_p - ref_p) <= tol), (c, s, b)
                if of is not None:
                    e = of.evaluate(got_p)
                    assert int(res[s, b]["peakBin"]) == e["peakBin"]
                    assert abs(float(res[s, b]["meanSnrDb"]) - e["meanSnrDb"]) < 0.01


def test_airspy_ssb_pcm_matches_oracle(sdrb, gpu_ctx):
    n, fs, S, B = 16384, 2_500_000, 2, 4
    cfg = sdrb.Config(430_000_000, fs, n, 5, inputFormat=sdrb.RU16)
    gpu_ctx.configure(cfg, S, B)
    raw = np.stack([synth_adc(2 * n * B, 300 + s) for s in range(S)]).reshape(S, B, 2 * n)
    gpu_ctx.process(np.ascontiguousarray(raw.reshape(S, B, n, 2)), B, sdrb.STAGE_SSB)
    pcm = gpu_ctx.pcm()
    for s in range(S):
        iq = _oracle_iq(raw[s])
        os_ = ol.OracleSsb()
        for b in range(B):
            exp = os_.process(np.ascontiguousarray(iq[b]), fs, 1, 1)
            assert np.abs(pcm[s, b].astype(int) - exp.astype(int)).max() <= 1, (s, b)


def _run_walk(sdrb, ctx, monkeypatch, raw, n, fs, chunk, warm):
    """All spectra of every call (bit patterns) + the count of chunks walked twice, for one chunking."""
    S, calls, B = raw.shape[:3]
    if chunk is None:
        monkeypatch.delenv("SDRB_AIRSPY_CHUNK", raising=False)
        monkeypatch.delenv("SDRB_AIRSPY_WARM", raising=False)
    else:
        monkeypatch.setenv("SDRB_AIRSPY_CHUNK", str(chunk))
        monkeypatch.setenv("SDRB_AIRSPY_WARM", str(warm))
    ctx.configure(sdrb.Config(430_000_000, fs, n, 25, inputFormat=sdrb.RU16), S, B)   # buffers + chunking are set up here
    out = []
    for c in range(calls):
        ctx.process(np.ascontiguousarray(raw[:, c].reshape(S, B, n, 2)), B, sdrb.STAGE_FFT)
        out.append(np.stack([[ctx.power(s, b).view(np.uint32) for b in range(B)] for s in range(S)]))
    return np.stack(out), ctx.airspy_redone_chunks()


@pytest.mark.parametrize("kind", ["noisy", "constant", "quiet"])
def test_airspy_chunked_dc_walk_is_the_sequential_walk(sdrb, gpu_ctx, monkeypatch, kind):
    """The DC tracker is walked in chunks from guessed starts (airspy.cu); whatever the chunking, and whether or
    not the guesses land, the converter output - hence every bit of the spectra - must be that of the one-chunk
    (purely sequential) walk, which the tests above hold against the oracle."""
    n, fs, S, B, calls = 16384, 10_000_000, 3, 4, 2
    total = 2 * n * B * calls
    if kind == "noisy":
        raw = np.stack([synth_adc(total, 500 + s, dc=11 * s - 9) for s in range(S)])
    elif kind == "constant":   # noise-free: the recurrence stalls in its dead band, guesses cannot land
        raw = np.stack([np.full(total, 2048 + 31 * s, np.uint16) for s in range(S)])
    else:                      # 1 LSB of dither around a large offset
        rng = np.random.default_rng(9)
        raw = np.stack([(2500 + 100 * s + rng.integers(0, 2, total)).astype(np.uint16) for s in range(S)])
    raw = raw.reshape(S, calls, B, 2 * n)
    seq, redone_seq = _run_walk(sdrb, gpu_ctx, monkeypatch, raw, n, fs, 1 << 30, 0)
    assert redone_seq == 0
    dflt, redone_dflt = _run_walk(sdrb, gpu_ctx, monkeypatch, raw, n, fs, None, None)
    assert np.array_equal(seq, dflt)
    if kind == "noisy":
        assert redone_dflt == 0          # 4096 samples of run-in: every guess lands
    forced, redone_forced = _run_walk(sdrb, gpu_ctx, monkeypatch, raw, n, fs, 1024, 8)
    assert np.array_equal(seq, forced)
    if kind == "noisy":
        assert redone_forced > 0         # 8 samples of run-in: the second walk did the work
    odd, _ = _run_walk(sdrb, gpu_ctx, monkeypatch, raw, n, fs, 5000, 2048)   # chunk not dividing the call
    assert np.array_equal(seq, odd)


@pytest.mark.parametrize("n,S,B,calls,chunking", [(16384, 3, 4, 2, None), (4096, 5, 3, 3, None), (65536, 2, 2, 2, None),
                                                  (16384, 2, 4, 2, (1024, 8)), (2048, 33, 1, 3, (1000, 64)),
                                                  (4, 2, 2, 4, None)])
def test_airspy_converter_output_is_bit_exact(sdrb, gpu_ctx, monkeypatch, n, S, B, calls, chunking):
    """The cf32 the converter hands on, bit for bit against the oracle (itself pinned bit-exactly against the
    reference's iqconverter_float.c), over several calls (carried avg + 48-sample tail), with the default
    chunking of the DC walk and with chunkings that force the second walk."""
    if chunking is None:
        monkeypatch.delenv("SDRB_AIRSPY_CHUNK", raising=False)
        monkeypatch.delenv("SDRB_AIRSPY_WARM", raising=False)
    else:
        monkeypatch.setenv("SDRB_AIRSPY_CHUNK", str(chunking[0]))
        monkeypatch.setenv("SDRB_AIRSPY_WARM", str(chunking[1]))
    gpu_ctx.configure(sdrb.Config(430_000_000, 10_000_000, n, 25, inputFormat=sdrb.RU16), S, B)
    raw = np.stack([synth_adc(2 * n * B * calls, 700 + s, dc=5 * s - 20) for s in range(S)]).reshape(S, calls, B, 2 * n)
    ref = np.stack([_oracle_iq(raw[s].reshape(calls * B, 2 * n)) for s in range(S)]).reshape(S, calls, B, n, 2)
    for c in range(calls):
        gpu_ctx.process(np.ascontiguousarray(raw[:, c].reshape(S, B, n, 2)), B, sdrb.STAGE_FFT)
        got = gpu_ctx.airspy_iq()
        assert np.array_equal(got.view(np.uint32), np.ascontiguousarray(ref[:, c]).view(np.uint32)), c
