"""GPU parity of the batched pulse detectors through the C-ABI:
SDRB_STAGE_SPECTRAL_PULSE / SDRB_STAGE_AUDIO_PULSE against oracle/sdr_oracle_pulse.c (itself pinned
bit-exactly against the reference's SpectralPulseDetector / AudioPulseDetector).  The detectors are
fed what the reference feeds them — the path's own per-reading best-1 kHz sigma/frequency and PCM — so
every field is compared bit for bit, with the detector state carried across several sdrb_process calls."""
import numpy as np
import pytest

import pulse_lib as pl

pytestmark = pytest.mark.gpu

SP_FIELDS = ("spectralLiveEtat", "spectralLevel", "spectralLocked", "spectralLastSnr", "spectralTTarget",
             "spectralEstFreqHzF")
AU_FIELDS = ("audioLiveEtat", "audioLevel", "audioLocked", "audioLastSnr", "audioTTarget")


def _run(sdrb, ctx, cfg, S, B, calls, stages, fmt):
    n, fs = cfg.samplesPerReading, cfg.sampleRate
    ctx.configure(cfg, S, B)
    res, pul, pcm = [], [], []
    for c in range(calls):
        iq = sdrb.synth_host(fmt, fs, 0, S, c * B * n, B * n).reshape(S, B, n, 2)
        ctx.process(np.ascontiguousarray(iq), B, stages)
        res.append(ctx.results())
        pul.append(ctx.pulse_results())
        if stages & sdrb.STAGE_SSB:
            pcm.append(ctx.pcm())
    return (np.concatenate(res, axis=1), np.concatenate(pul, axis=1),
            np.concatenate(pcm, axis=1) if pcm else None)


@pytest.mark.parametrize("n,S,B,calls", [(16384, 6, 256, 5), (4096, 3, 512, 5)])
def test_spectral_pulse_matches_oracle(sdrb, gpu_ctx, n, S, B, calls):
    fs = 2_000_000
    cfg = sdrb.Config(430_000_000, fs, n, 5, inputFormat=sdrb.CU8)
    res, pul, _ = _run(sdrb, gpu_ctx, cfg, S, B, calls, sdrb.STAGE_FFT | sdrb.STAGE_SPECTRAL_PULSE, sdrb.CU8)
    fs_energy = float(np.float32(fs) / np.float32(n))
    top = 0
    for s in range(S):
        o = pl.OraclePulse(False, fs_energy)
        for b in range(B * calls):
            sigma, freq = float(res[s, b]["best1kHzSnrSigma"]), float(res[s, b]["best1kHzCenterFreqHz"])
            exp = o.spectral(sigma, freq)
            got = np.array([pul[s, b][k] for k in SP_FIELDS], dtype=np.float32)
            assert got.tobytes() == exp.tobytes(), (s, b, got, exp)
            assert pul[s, b]["spectralSnrSigma"] == res[s, b]["best1kHzSnrSigma"]
            assert int(pul[s, b]["spectralEstFreqHz"]) == int(np.round(np.float64(exp[5]))) or \
                abs(int(pul[s, b]["spectralEstFreqHz"]) - float(exp[5])) <= 0.5
            assert pul[s, b]["overflow"] == 0
            top = max(top, int(exp[0]))
    assert top >= 1, "the keyed synthetic tone should raise the live state on some stream"


@pytest.mark.parametrize("overrides", [None, {"snrStrong": 1.4, "fsEnergy": 50.0, "noiseRefFar": 30, "noiseRefNear": 10}])
def test_audio_pulse_matches_oracle(sdrb, gpu_ctx, overrides):
    fs, n, S, B, calls = 2_000_000, 16384, 4, 256, 5
    cfg = sdrb.Config(430_000_000, fs, n, 5, inputFormat=sdrb.CU8)
    gpu_ctx.configure(cfg, S, B)
    if overrides:
        ac = sdrb.audio_pulse_config_default()
        for k, v in overrides.items():
            setattr(ac, k, v)
        gpu_ctx.set_pulse_config(ac)
    else:
        gpu_ctx.set_pulse_config(None)
    stages = sdrb.STAGE_FFT | sdrb.STAGE_SSB | sdrb.STAGE_AUDIO_PULSE | sdrb.STAGE_SPECTRAL_PULSE
    pul, pcm = [], []
    for c in range(calls):
        iq = sdrb.synth_host(sdrb.CU8, fs, 0, S, c * B * n, B * n).reshape(S, B, n, 2)
        gpu_ctx.process(np.ascontiguousarray(iq), B, stages)
        pul.append(gpu_ctx.pulse_results())
        pcm.append(gpu_ctx.pcm())
    pul, pcm = np.concatenate(pul, axis=1), np.concatenate(pcm, axis=1)
    seen = set()
    for s in range(S):
        o = pl.OraclePulse(True, overrides=overrides)
        for b in range(B * calls):
            exp = o.pcm(pcm[s, b])
            got = np.array([pul[s, b][k] for k in AU_FIELDS], dtype=np.float32)
            assert got.tobytes() == exp.tobytes(), (s, b, got, exp)
            seen.add(int(exp[0]))
    assert len(seen) >= 1


def test_pulse_stage_argument_checks(sdrb, gpu_ctx):
    cfg = sdrb.Config(430_000_000, 2_000_000, 16384, 5, inputFormat=sdrb.CU8)
    gpu_ctx.configure(cfg, 1, 1)
    iq = sdrb.synth_host(sdrb.CU8, 2_000_000, 0, 1, 0, 16384).reshape(1, 1, 16384, 2)
    with pytest.raises(sdrb.SdrbError):
        gpu_ctx.process(iq, 1, sdrb.STAGE_SPECTRAL_PULSE)          # needs the FFT stage
    with pytest.raises(sdrb.SdrbError):
        gpu_ctx.process(iq, 1, sdrb.STAGE_FFT | sdrb.STAGE_AUDIO_PULSE)  # needs the SSB stage
    gpu_ctx.process(iq, 1, sdrb.STAGE_FFT)
    with pytest.raises(sdrb.SdrbError):
        gpu_ctx.pulse_results()                                     # no pulse stage has run
