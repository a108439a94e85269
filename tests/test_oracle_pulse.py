"""Pins oracle/sdr_oracle_pulse.c against the reference's own SpectralPulseDetector and
AudioPulseDetector (oracle/_ref/libsdr_ref.so, built from src/dsp/spectral_pulse_detector.cpp and
src/ssb/audio_pulse_detector.cpp): every output after every frame / PCM block, bit for bit."""
import os

import numpy as np
import pytest

import oracle_lib as ol
import pulse_lib as pl

needs_ref = pytest.mark.skipif(not os.path.exists(ol.REF_SO), reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("fs_energy,n_frames,strength,jitter,seed", [
    (2e6 / 16384, 4000, 6.0, 0.0, 1),      # config C1/C2 frame rate, clean beacon -> locks
    (2e6 / 16384, 4000, 3.2, 0.05, 2),     # weak, jittered: rhythm admissions and penalties
    (30.72e6 / 65536, 6000, 5.0, 0.3, 3),  # C3 frame rate, strong jitter: confusion penalty
    (10e6 / 1048576, 400, 6.0, 0.0, 4),    # C5 frame rate (9.5 Hz): tiny buffers
    (20.0, 1500, 0.0, 0.0, 5),             # noise only
])
def test_spectral_restatement_is_bit_exact(fs_energy, n_frames, strength, jitter, seed):
    fs_energy = float(np.float32(fs_energy))
    s, f = pl.synth_sigma(n_frames, fs_energy, seed, strength=strength, jitter=jitter)
    o, r = pl.OraclePulse(False, fs_energy), pl.RefPulse(False, fs_energy)
    seen = set()
    for i in range(n_frames):
        a, b = o.spectral(float(s[i]), float(f[i])), r.spectral(float(s[i]), float(f[i]))
        assert a.tobytes() == b.tobytes(), (i, a, b)
        seen.add(int(a[0]))
    if strength >= 5.0 and jitter == 0.0:
        assert max(seen) >= 3, "the clean beacon should reach a MEDIUM/STRONG live state"


@needs_ref
@pytest.mark.parametrize("seed,amp,block", [(1, 6000.0, 394), (2, 2500.0, 394), (3, 0.0, 103), (4, 9000.0, 5041)])
def test_audio_restatement_is_bit_exact(seed, amp, block):
    pcm = pl.synth_pcm(48000 * 14, seed, amp=amp)
    o, r = pl.OraclePulse(True), pl.RefPulse(True)
    seen = set()
    for i in range(0, pcm.size - block + 1, block):
        a, b = o.pcm(pcm[i:i + block]), r.pcm(pcm[i:i + block])
        assert a.tobytes() == b.tobytes(), (i, a, b)
        seen.add(int(a[0]))
    if amp >= 6000.0:
        assert max(seen) >= 1, "tone bursts should be detected"
