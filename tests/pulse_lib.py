"""ctypes access to the pulse-detector checkers (TEST INFRASTRUCTURE ONLY): the C restatement
oracle/sdr_oracle_pulse.c and the reference's own SpectralPulseDetector / AudioPulseDetector inside
oracle/_ref/libsdr_ref.so.  Also the synthetic detector inputs shared by the CPU and GPU tests."""
import ctypes as C

import numpy as np

import oracle_lib as ol


class PulseCfg(C.Structure):
    _fields_ = [(k, C.c_float) for k in ("fsEnergy", "zDefaultS", "tTargetInit", "dtTolS", "snrMin", "snrRhythm",
                                         "snrStrong", "dispersionMax")] + [("sumNMax", C.c_int32)] + \
               [(k, C.c_float) for k in ("liveWindowT", "liveDivisor", "sampleRate", "fMin", "fMax", "smoothCutoff")] + \
               [("noiseRefFar", C.c_int32), ("noiseRefNear", C.c_int32)]


_o = None


def lib():
    global _o
    if _o is None:
        o = C.CDLL(ol.build_oracle())
        o.sdr_oracle_pulse_cfg_spectral.argtypes = [C.POINTER(PulseCfg), C.c_float]
        o.sdr_oracle_pulse_cfg_audio.argtypes = [C.POINTER(PulseCfg)]
        o.sdr_oracle_pulse_create.restype = C.c_void_p
        o.sdr_oracle_pulse_create.argtypes = [C.POINTER(PulseCfg), C.c_int]
        o.sdr_oracle_pulse_destroy.argtypes = [C.c_void_p]
        o.sdr_oracle_pulse_spectral.argtypes = [C.c_void_p, C.c_float, C.c_float]
        o.sdr_oracle_pulse_audio_pcm.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        for f in ("live_etat", "level", "locked", "n_rois"):
            getattr(o, "sdr_oracle_pulse_" + f).restype = C.c_int
            getattr(o, "sdr_oracle_pulse_" + f).argtypes = [C.c_void_p]
        for f in ("last_snr", "t_target", "est_freq"):
            getattr(o, "sdr_oracle_pulse_" + f).restype = C.c_float
            getattr(o, "sdr_oracle_pulse_" + f).argtypes = [C.c_void_p]
        _o = o
    return _o


class OraclePulse:
    """audio=False: SpectralPulseDetector(fs_energy); audio=True: AudioPulseDetector()."""

    def __init__(self, audio, fs_energy=None, overrides=None):
        o = lib()
        cfg = PulseCfg()
        if audio:
            o.sdr_oracle_pulse_cfg_audio(C.byref(cfg))
        else:
            o.sdr_oracle_pulse_cfg_spectral(C.byref(cfg), fs_energy)
        for k, v in (overrides or {}).items():
            setattr(cfg, k, v)
        self.h = o.sdr_oracle_pulse_create(C.byref(cfg), int(audio))
        self.audio = audio

    def out(self):
        o = lib()
        r = [o.sdr_oracle_pulse_live_etat(self.h), o.sdr_oracle_pulse_level(self.h), o.sdr_oracle_pulse_locked(self.h),
             o.sdr_oracle_pulse_last_snr(self.h), o.sdr_oracle_pulse_t_target(self.h)]
        if not self.audio:
            r.append(o.sdr_oracle_pulse_est_freq(self.h))
        return np.array(r, dtype=np.float32)

    def spectral(self, snr, freq):
        lib().sdr_oracle_pulse_spectral(self.h, snr, freq)
        return self.out()

    def pcm(self, pcm):
        pcm = np.ascontiguousarray(pcm, dtype=np.int16)
        lib().sdr_oracle_pulse_audio_pcm(self.h, pcm.ctypes.data, pcm.size)
        return self.out()

    def __del__(self):
        if lib and self.h:
            lib().sdr_oracle_pulse_destroy(self.h)
            self.h = None


class RefPulse:
    """The reference's own detector classes (oracle/_ref/libsdr_ref.so)."""

    def __init__(self, audio, fs_energy=None):
        r = C.CDLL(ol.REF_SO)
        r.sdr_ref_spd_create.restype = C.c_void_p
        r.sdr_ref_spd_create.argtypes = [C.c_float]
        r.sdr_ref_spd_process.argtypes = [C.c_void_p, C.c_float, C.c_float, ol.f32p]
        r.sdr_ref_spd_destroy.argtypes = [C.c_void_p]
        r.sdr_ref_apd_create.restype = C.c_void_p
        r.sdr_ref_apd_process.argtypes = [C.c_void_p, C.c_void_p, C.c_int, ol.f32p]
        r.sdr_ref_apd_destroy.argtypes = [C.c_void_p]
        self.r, self.audio = r, audio
        self.h = r.sdr_ref_apd_create() if audio else r.sdr_ref_spd_create(fs_energy)

    def spectral(self, snr, freq):
        out = np.zeros(6, np.float32)
        self.r.sdr_ref_spd_process(self.h, snr, freq, out.ctypes.data_as(ol.f32p))
        return out

    def pcm(self, pcm):
        pcm = np.ascontiguousarray(pcm, dtype=np.int16)
        out = np.zeros(5, np.float32)
        self.r.sdr_ref_apd_process(self.h, pcm.ctypes.data, pcm.size, out.ctypes.data_as(ol.f32p))
        return out


def synth_sigma(n_frames, fs_energy, seed, period=1.75, strength=6.0, jitter=0.0, fc=430e6):
    """Per-frame (best1kHzSnrSigma, best1kHzCenterFreqHz) of a keyed beacon: noise ~ N(0.8, 0.6),
    a bump of `strength` sigma every `period` seconds, a slowly drifting centre frequency."""
    rng = np.random.default_rng(seed)
    t = np.arange(n_frames) / fs_energy
    s = (0.8 + 0.6 * rng.standard_normal(n_frames)).astype(np.float32)
    k = 0
    tp = 0.9
    while tp < t[-1]:
        i = int(round(tp * fs_energy))
        w = max(1, int(0.1 * fs_energy))
        for j in range(max(0, i - w), min(n_frames, i + w + 1)):
            s[j] += np.float32(strength * np.exp(-0.5 * ((j - i) / (0.4 * w + 1e-9)) ** 2))
        k += 1
        tp += period + jitter * rng.standard_normal()
    f = (fc + 1500.0 + 2.0 * t + 30.0 * rng.standard_normal(n_frames)).astype(np.float32)
    return s, f


def synth_pcm(n_samples, seed, period=1.75, tone=2500.0, amp=6000.0, noise=900.0, fs=48000.0):
    """int16 audio: noise plus a 0.2 s tone burst every `period` seconds."""
    rng = np.random.default_rng(seed)
    t = np.arange(n_samples) / fs
    x = noise * rng.standard_normal(n_samples)
    gate = ((t - 1.0) % period) < 0.2
    x += gate * amp * np.sin(2 * np.pi * tone * t)
    return np.clip(np.round(x), -32768, 32767).astype(np.int16)
