"""ctypes binding of include/sdr_b200.h (the drop-in C-ABI of the reference's DSP objects,
src/sdr-bridge-java-soapy.cpp:101-108,441-455).  One-to-one with the header; no arithmetic here."""
import ctypes as C
import os

import numpy as np

CU8, CS16, CF32, RU16 = 0, 1, 2, 3
STAGE_FFT, STAGE_SSB = 1, 2
STAGE_SPECTRAL_PULSE, STAGE_AUDIO_PULSE = 4, 8
POWER_LINEAR_F32, POWER_DB_F32, POWER_DB_F16 = 0, 1, 2
_HERE = os.path.dirname(os.path.abspath(__file__))
lib_path = os.environ.get("SDRB_LIB") or os.path.join(_HERE, "libsdr_b200.so")   # SDRB_LIB: A/B builds of the same ABI


class SdrbError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"sdrb error {code}: {text}")
        self.code = code


class Config(C.Structure):
    _fields_ = [("centerFrequency", C.c_uint32), ("sampleRate", C.c_uint32),
                ("samplesPerReading", C.c_int32), ("freqFocusRangeKhz", C.c_int32),
                ("soundMode", C.c_int32), ("inputFormat", C.c_int32), ("upperSideband", C.c_int32),
                ("fftWindow", C.c_int32), ("fftRemoveDc", C.c_int32), ("ssbTrueSideband", C.c_int32),
                ("ssbChannels", C.c_int32), ("ssbChannelOffsetHz", C.c_float),
                ("ssbChannelSpacingHz", C.c_float), ("reserved", C.c_int32 * 3)]

    def __init__(self, centerFrequency=430_000_000, sampleRate=2_000_000, samplesPerReading=16384,
                 freqFocusRangeKhz=5, soundMode=1, inputFormat=CU8, upperSideband=1, **kw):
        super().__init__(centerFrequency=centerFrequency, sampleRate=sampleRate,
                         samplesPerReading=samplesPerReading, freqFocusRangeKhz=freqFocusRangeKhz,
                         soundMode=soundMode, inputFormat=inputFormat, upperSideband=upperSideband, **kw)


class FrameResult(C.Structure):
    _fields_ = [("meanSnrDb", C.c_float), ("meanSnrSigma", C.c_float), ("peakAboveNoiseMeanDb", C.c_float),
                ("maxBinSnrDb", C.c_float), ("maxBinSnrSigma", C.c_float), ("best1kHzSnrDb", C.c_float),
                ("best1kHzSnrSigma", C.c_float), ("best1kHzCenterFreqHz", C.c_float),
                ("perBinMean", C.c_float), ("trackingFrequencyHz", C.c_float),
                ("trackingFrequency", C.c_int64), ("detectionFlag", C.c_int32),
                ("absPeakDb", C.c_float), ("signalPowerDb", C.c_float), ("peakFrequencyHz", C.c_float),
                ("peakBin", C.c_int32), ("valid", C.c_int32), ("nRef", C.c_int32), ("reserved", C.c_int32)]


RESULT_DTYPE = np.dtype(FrameResult)


class AudioPulseConfig(C.Structure):
    """sdrb_audio_pulse_config = AudioPulseDetector::Config (src/ssb/audio_pulse_detector.h:20-38)."""
    _fields_ = [(k, C.c_float) for k in ("sampleRate", "fMin", "fMax", "fsEnergy", "smoothCutoff", "zDefaultS",
                                         "tTargetInit", "dtTolS", "snrMin", "snrRhythm", "snrStrong",
                                         "dispersionMax")] + [("sumNMax", C.c_int32)] + \
               [("liveWindowT", C.c_float), ("liveDivisor", C.c_float), ("noiseRefFar", C.c_int32),
                ("noiseRefNear", C.c_int32)]


class PulseResult(C.Structure):
    _fields_ = [("spectralSnrSigma", C.c_float), ("spectralLiveEtat", C.c_int32), ("spectralEstFreqHz", C.c_int64),
                ("spectralEstFreqHzF", C.c_float), ("spectralLevel", C.c_int32), ("spectralLocked", C.c_int32),
                ("spectralLastSnr", C.c_float), ("spectralTTarget", C.c_float),
                ("audioLastSnr", C.c_float), ("audioLiveEtat", C.c_int32), ("audioLevel", C.c_int32),
                ("audioLocked", C.c_int32), ("audioTTarget", C.c_float), ("overflow", C.c_int32),
                ("reserved", C.c_int32)]


PULSE_DTYPE = np.dtype(PulseResult)


class FrameEstimate(C.Structure):
    """sdrb_frame_estimate: the frame-local estimator record of sdrb_estimate_noalloc."""
    _fields_ = [(k, C.c_float) for k in ("absPeakDb", "signalPowerDb", "meanSnrDb", "meanSnrSigma", "perBinMean",
                                         "peakAboveNoiseMeanDb", "maxBinSnrDb", "maxBinSnrSigma", "best1kHzSnrDb",
                                         "best1kHzSnrSigma")] + \
               [(k, C.c_int32) for k in ("best1kStart", "peakBin", "valid", "reserved")]


ESTIMATE_DTYPE = np.dtype(FrameEstimate)
NOALLOC_MAX_FRAMES = 4096


def estimate_noalloc(device, sample_rate, n, focus_khz, d_power_ptr, n_frames, d_out_ptr, stream_ptr=0):
    """sdrb_estimate_noalloc: device pointers in, nothing allocated; raises SdrbError on a rejected call."""
    _check(lib.sdrb_estimate_noalloc(device, sample_rate, n, focus_khz, C.c_void_p(d_power_ptr), n_frames,
                                     C.c_void_p(d_out_ptr), C.c_void_p(stream_ptr)))


class Plan(C.Structure):
    _fields_ = [("n", C.c_int32), ("focusLo", C.c_int32), ("focusHi", C.c_int32), ("focusLen", C.c_int32),
                ("winBins1k", C.c_int32), ("nWin", C.c_int32), ("nBottom", C.c_int32),
                ("winLo", C.c_int32 * 10), ("winHi", C.c_int32 * 10),
                ("freqPerBin", C.c_float), ("nyquist", C.c_float),
                ("decim", C.c_int32), ("nTaps", C.c_int32), ("pcmPerBlock", C.c_int32), ("fftPath", C.c_int32),
                ("rf", C.c_float * 5), ("hp", C.c_float * 5), ("bp", C.c_float * 5),
                ("agcTarget", C.c_float), ("agcFast", C.c_float), ("agcSlow", C.c_float), ("gain", C.c_float),
                ("transientCoeff", C.c_float), ("dcAlpha", C.c_float), ("firTaps", C.c_float * 255)]


EXPORTS = {
    "sdrb_abi_version": (C.c_int, []),
    "sdrb_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "sdrb_destroy": (C.c_int, [C.c_void_p]),
    "sdrb_last_error": (C.c_char_p, [C.c_void_p]),
    "sdrb_set_cuda_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "sdrb_configure": (C.c_int, [C.c_void_p, C.POINTER(Config), C.c_int, C.c_int]),
    "sdrb_replan": (C.c_int, [C.c_void_p, C.POINTER(Config), C.c_int]),
    "sdrb_set_center_frequency": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32]),
    "sdrb_set_focus_range": (C.c_int, [C.c_void_p, C.c_int]),
    "sdrb_set_sound_mode": (C.c_int, [C.c_void_p, C.c_int]),
    "sdrb_pcm_per_block": (C.c_int, [C.c_void_p]),
    "sdrb_plan_config": (C.c_int, [C.POINTER(Config), C.POINTER(Plan)]),
    "sdrb_process": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint]),
    "sdrb_sync": (C.c_int, [C.c_void_p]),
    "sdrb_process_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sdrb_process_host_ex": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint, C.c_void_p, C.c_void_p, C.c_int,
                                       C.c_void_p]),
    "sdrb_get_power_db": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "sdrb_get_results": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "sdrb_results_device": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "sdrb_audio_pulse_config_default": (None, [C.POINTER(AudioPulseConfig)]),
    "sdrb_set_pulse_config": (C.c_int, [C.c_void_p, C.POINTER(AudioPulseConfig)]),
    "sdrb_get_pulse_results": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "sdrb_pulse_results_device": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "sdrb_airspy_redone_chunks": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "sdrb_get_airspy_iq": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "sdrb_power_device": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "sdrb_get_power": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "sdrb_get_power_all": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "sdrb_pcm_device": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "sdrb_get_pcm": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "sdrb_estimate_noalloc": (C.c_int, [C.c_int, C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "sdrb_state_bytes": (C.c_size_t, [C.c_void_p]),
    "sdrb_save_state": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "sdrb_load_state": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "sdrb_reset_state": (C.c_int, [C.c_void_p]),
    "sdrb_synth_host": (C.c_int, [C.c_int, C.c_uint32, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p]),
    "sdrb_synth_device": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_int, C.c_int, C.c_int64, C.c_int64,
                                    C.c_void_p]),
    "sdrb_format_bytes": (C.c_size_t, [C.c_int]),
    "sdrb_launch_count": (C.c_uint64, [C.c_void_p]),
    "sdrb_last_fft_kernel_ms": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "sdrb_enable_kernel_timing": (C.c_int, [C.c_void_p, C.c_int]),
    "sdrb_last_fft_kernel_frames": (C.c_int, [C.c_void_p]),
}

if not os.path.exists(lib_path):
    raise ImportError(f"{lib_path} is missing: build it with `make -C sdr-for-android-lib_b200/csrc` "
                      "(or __graft_entry__.build()); there is no CPU fallback")
lib = C.CDLL(lib_path)
for _name, (_res, _args) in EXPORTS.items():
    _f = getattr(lib, _name)
    _f.restype, _f.argtypes = _res, _args


def format_bytes(fmt):
    return int(lib.sdrb_format_bytes(fmt))


def format_dtype(fmt):
    """numpy dtype and per-sample width of a wire format: [.., n, 2] of u8 / i16 / f32."""
    return {CU8: np.uint8, CS16: np.int16, CF32: np.float32, RU16: np.uint16}[fmt]


def _check(rc, ctx=None):
    if rc < 0:
        raise SdrbError(rc, (lib.sdrb_last_error(ctx) or b"").decode())
    return rc


def audio_pulse_config_default():
    c = AudioPulseConfig()
    lib.sdrb_audio_pulse_config_default(C.byref(c))
    return c


def plan_config(cfg):
    p = Plan()
    _check(lib.sdrb_plan_config(C.byref(cfg), C.byref(p)))
    return p


def synth_host(fmt, sample_rate, first_stream, n_streams, first_sample, n_samples):
    """Deterministic synthetic IQ, [n_streams, n_samples, 2] in the wire dtype (host generator)."""
    out = np.empty((n_streams, n_samples, 2), dtype=format_dtype(fmt))
    _check(lib.sdrb_synth_host(fmt, sample_rate, first_stream, n_streams, first_sample, n_samples,
                               out.ctypes.data_as(C.c_void_p)))
    return out


def _ptr(x):
    """Device or host address of a numpy array / torch tensor / int."""
    if isinstance(x, (int, np.integer)):
        return int(x), None
    if isinstance(x, np.ndarray):
        return x.ctypes.data, False
    return x.data_ptr(), bool(x.is_cuda)


class Context:
    """One GPU context = the reference's FFTProcessor + SSBProcessor globals for S streams."""

    def __init__(self, device=0):
        h = C.c_void_p()
        _check(lib.sdrb_create(device, C.byref(h)))
        self._h = h
        self.cfg = None
        self.n_streams = 0
        self.last_blocks = 0

    def close(self):
        if self._h:
            lib.sdrb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def configure(self, cfg, n_streams, max_blocks):
        _check(lib.sdrb_configure(self._h, C.byref(cfg), n_streams, max_blocks), self._h)
        self.cfg, self.n_streams, self.max_blocks = cfg, n_streams, max_blocks

    def replan(self, cfg, max_blocks):
        """New geometry for the same streams, every carried state kept (sdrb_replan)."""
        _check(lib.sdrb_replan(self._h, C.byref(cfg), max_blocks), self._h)
        self.cfg, self.max_blocks = cfg, max_blocks

    def set_cuda_stream(self, stream_ptr):
        _check(lib.sdrb_set_cuda_stream(self._h, C.c_void_p(stream_ptr)), self._h)

    def set_center_frequency(self, stream, hz):
        _check(lib.sdrb_set_center_frequency(self._h, stream, hz), self._h)

    def set_focus_range(self, khz):
        _check(lib.sdrb_set_focus_range(self._h, khz), self._h)

    def set_sound_mode(self, mode):
        _check(lib.sdrb_set_sound_mode(self._h, mode), self._h)

    @property
    def pcm_per_block(self):
        return _check(lib.sdrb_pcm_per_block(self._h), self._h)

    def process(self, iq, n_blocks, stages=STAGE_FFT | STAGE_SSB, on_device=None):
        addr, is_cuda = _ptr(iq)
        if on_device is None:
            on_device = bool(is_cuda)
        _check(lib.sdrb_process(self._h, C.c_void_p(addr), 1 if on_device else 0, n_blocks, stages), self._h)
        self.last_blocks = n_blocks

    def process_host(self, iq, n_blocks, stages=STAGE_FFT | STAGE_SSB, results=None, power=None, pcm=None,
                     power_format=POWER_LINEAR_F32):
        """Pipelined host-to-host call; outputs are preallocated numpy arrays / pinned tensors or None.
        power_format: POWER_LINEAR_F32 (the reference's payload), POWER_DB_F32 or POWER_DB_F16 (dB view)."""
        ptr = lambda x: C.c_void_p(_ptr(x)[0]) if x is not None else C.c_void_p(0)
        _check(lib.sdrb_process_host_ex(self._h, ptr(iq), n_blocks, stages, ptr(results), ptr(power), power_format,
                                        ptr(pcm)), self._h)
        self.last_blocks = n_blocks

    def power_db(self, stream, block):
        """dB view of one spectrum: 10*log10f(p + 1e-20f) (src/dsp/fft_process.cpp:149)."""
        out = np.empty(self.cfg.samplesPerReading, dtype=np.float32)
        _check(lib.sdrb_get_power_db(self._h, stream, block, out.ctypes.data_as(C.c_void_p)), self._h)
        return out

    def sync(self):
        _check(lib.sdrb_sync(self._h), self._h)

    def results(self):
        out = np.zeros((self.n_streams, self.last_blocks), dtype=RESULT_DTYPE)
        _check(lib.sdrb_get_results(self._h, out.ctypes.data_as(C.c_void_p), out.size), self._h)
        return out

    def pulse_results(self):
        """[n_streams, last_blocks] records ([n_streams, channels, last_blocks] with an NCO bank)."""
        ch = max(1, int(self.cfg.ssbChannels))
        shape = (self.n_streams, self.last_blocks) if ch == 1 else (self.n_streams, ch, self.last_blocks)
        out = np.zeros(shape, dtype=PULSE_DTYPE)
        _check(lib.sdrb_get_pulse_results(self._h, out.ctypes.data_as(C.c_void_p), out.size), self._h)
        return out

    def airspy_iq(self):
        """[n_streams, last_blocks, N, 2] float32: the real->IQ converter's output of the last RU16 call."""
        out = np.zeros((self.n_streams, self.last_blocks, int(self.cfg.samplesPerReading), 2), dtype=np.float32)
        _check(lib.sdrb_get_airspy_iq(self._h, out.ctypes.data_as(C.c_void_p), out.size), self._h)
        return out

    def airspy_redone_chunks(self):
        """Chunks of the RU16 DC pass that were walked a second time (diagnostic; 0 on noisy ADC data)."""
        n = C.c_uint64(0)
        _check(lib.sdrb_airspy_redone_chunks(self._h, C.byref(n)), self._h)
        return int(n.value)

    def set_pulse_config(self, cfg=None):
        """SSBProcessor::setPulseConfig: rebuild every audio detector (None = the reference defaults)."""
        if cfg is None:
            cfg = audio_pulse_config_default()
        _check(lib.sdrb_set_pulse_config(self._h, C.byref(cfg)), self._h)

    def power(self, stream, block):
        out = np.empty(self.cfg.samplesPerReading, dtype=np.float32)
        _check(lib.sdrb_get_power(self._h, stream, block, out.ctypes.data_as(C.c_void_p)), self._h)
        return out

    def power_all(self, out=None):
        """All spectra of the last call into `out` (numpy array or pinned torch tensor) or a new array."""
        n = self.n_streams * self.last_blocks * self.cfg.samplesPerReading
        if out is None:
            out = np.empty((self.n_streams, self.last_blocks, self.cfg.samplesPerReading), dtype=np.float32)
        addr, _ = _ptr(out)
        _check(lib.sdrb_get_power_all(self._h, C.c_void_p(addr), n), self._h)
        return out

    def results_into(self, out):
        """Records of the last call into a preallocated buffer (numpy RESULT_DTYPE array or pinned bytes)."""
        addr, _ = _ptr(out)
        _check(lib.sdrb_get_results(self._h, C.c_void_p(addr), self.n_streams * self.last_blocks), self._h)
        return out

    def power_device_ptr(self):
        p = C.c_void_p()
        _check(lib.sdrb_power_device(self._h, C.byref(p)), self._h)
        return p.value

    def results_device_ptr(self):
        p = C.c_void_p()
        _check(lib.sdrb_results_device(self._h, C.byref(p)), self._h)
        return p.value

    def pcm_device_ptr(self):
        p = C.c_void_p()
        _check(lib.sdrb_pcm_device(self._h, C.byref(p)), self._h)
        return p.value

    def pcm(self):
        """[streams, blocks, m]; with the NCO channel bank on: [streams, channels, blocks, m]."""
        m = self.pcm_per_block
        ch = int(self.cfg.ssbChannels)
        shape = (self.n_streams, ch, self.last_blocks, m) if ch > 1 else (self.n_streams, self.last_blocks, m)
        out = np.zeros(shape, dtype=np.int16)
        _check(lib.sdrb_get_pcm(self._h, out.ctypes.data_as(C.c_void_p), out.size), self._h)
        return out

    def save_state(self):
        n = lib.sdrb_state_bytes(self._h)
        buf = np.zeros(n, dtype=np.uint8)
        _check(lib.sdrb_save_state(self._h, buf.ctypes.data_as(C.c_void_p), n), self._h)
        return buf

    def load_state(self, buf):
        _check(lib.sdrb_load_state(self._h, buf.ctypes.data_as(C.c_void_p), buf.size), self._h)

    def reset_state(self):
        _check(lib.sdrb_reset_state(self._h), self._h)

    def synth_device(self, fmt, sample_rate, first_stream, n_streams, first_sample, n_samples, dev_ptr):
        _check(lib.sdrb_synth_device(self._h, fmt, sample_rate, first_stream, n_streams, first_sample, n_samples,
                                     C.c_void_p(dev_ptr)), self._h)

    @property
    def launch_count(self):
        return int(lib.sdrb_launch_count(self._h))

    def enable_kernel_timing(self, on=True):
        _check(lib.sdrb_enable_kernel_timing(self._h, 1 if on else 0), self._h)

    @property
    def last_fft_kernel_frames(self):
        return int(lib.sdrb_last_fft_kernel_frames(self._h))

    def last_fft_kernel_ms(self):
        ms = C.c_float()
        n = _check(lib.sdrb_last_fft_kernel_ms(self._h, C.byref(ms)), self._h)
        return ms.value, n
