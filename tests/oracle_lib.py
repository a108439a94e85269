"""ctypes access to the two CPU checkers (TEST INFRASTRUCTURE ONLY):

* ``oracle/libsdr_oracle.so``  — the plain-C restatement (oracle/sdr_oracle.c), always buildable;
* ``oracle/_ref/libsdr_ref.so`` — the reference's own sources compiled in place (oracle/Makefile),
  prebuilt in the authoring container and shipped to the GPU box with the snapshot.

Nothing under sdr-for-android-lib_b200/ imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libsdr_oracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libsdr_ref.so")

f32p = C.POINTER(C.c_float)
i16p = C.POINTER(C.c_int16)


def _fp(a):
    return a.ctypes.data_as(f32p)


class FftState(C.Structure):
    _fields_ = [("centerFrequency", C.c_uint32), ("sampleRate", C.c_uint32),
                ("samplesPerReading", C.c_int32), ("freqFocusRangeKhz", C.c_int32),
                ("trackingFrequency", C.c_float), ("maxPeakDb", C.c_float), ("maxPeakHz", C.c_float),
                ("timeOfLastMaxPeak_ns", C.c_int64), ("timeOfLastMaxPeakUpdate_ns", C.c_int64),
                ("peakConfirmed", C.c_int32), ("flagRing", C.c_int32 * 3), ("flagIndex", C.c_int32),
                ("centerChanged", C.c_int32),
                ("meanSnrDb", C.c_float), ("meanSnrSigma", C.c_float), ("detectionFlag", C.c_int32),
                ("peakAboveNoiseMeanDb", C.c_float), ("maxBinSnrDb", C.c_float), ("maxBinSnrSigma", C.c_float),
                ("best1kHzSnrDb", C.c_float), ("best1kHzSnrSigma", C.c_float),
                ("best1kHzCenterFreqHz", C.c_float), ("perBinMean", C.c_float),
                ("absPeakDb", C.c_float), ("signalPowerDb", C.c_float), ("peakBin", C.c_int32),
                ("valid", C.c_int32)]


class Biquad(C.Structure):
    _fields_ = [(k, C.c_float) for k in ("a0", "a1", "a2", "b1", "b2", "z1", "z2")]


class SsbState(C.Structure):
    _fields_ = [("initialised", C.c_int32), ("sampCount", C.c_int32),
                ("rf", Biquad), ("hp", Biquad), ("bp", Biquad),
                ("agcTarget", C.c_float), ("agcFast", C.c_float), ("agcSlow", C.c_float), ("gain", C.c_float),
                ("lowpassBd", C.c_float), ("lowpassQ", C.c_float), ("transientCoeff", C.c_float)]


class SsbxState(C.Structure):
    _fields_ = [("base", SsbState), ("rfq", Biquad), ("samples", C.c_uint64), ("halfBand", C.c_double)]


def build_oracle():
    """(Re)build the C restatement; cheap, so done on demand."""
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("sdr_oracle.c", "sdr_oracle_ext.c", "sdr_oracle_pulse.c", "sdr_oracle_airspy.c", "sdr_oracle.h",
                                                "sdr_oracle_pulse.h")]
    if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle"], stdout=subprocess.DEVNULL)
    return ORACLE_SO


_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        o = C.CDLL(build_oracle())
        o.sdr_oracle_power_shifted.restype = C.c_int
        o.sdr_oracle_power_shifted.argtypes = [f32p, C.c_int, f32p]
        o.sdr_oracle_fft_init.argtypes = [C.POINTER(FftState), C.c_uint32, C.c_uint32, C.c_int, C.c_int]
        o.sdr_oracle_evaluate.argtypes = [C.POINTER(FftState), C.c_int, f32p, C.c_int64]
        o.sdr_oracle_tracking_long.restype = C.c_int64
        o.sdr_oracle_tracking_long.argtypes = [C.POINTER(FftState)]
        o.sdr_oracle_frame_clock_ns.restype = C.c_int64
        o.sdr_oracle_frame_clock_ns.argtypes = [C.c_int64, C.c_int, C.c_uint32]
        o.sdr_oracle_convert_u8.argtypes = [C.c_void_p, C.c_int, f32p]
        o.sdr_oracle_convert_s16.argtypes = [C.c_void_p, C.c_int, f32p]
        o.sdr_oracle_ssb_init.argtypes = [C.POINTER(SsbState)]
        o.sdr_oracle_process_ssb.restype = C.c_int
        o.sdr_oracle_process_ssb.argtypes = [C.POINTER(SsbState), f32p, C.c_int, C.c_uint32, C.c_int, C.c_int,
                                             i16p, C.c_int]
        o.sdr_oracle_fir_taps.restype = C.c_int
        o.sdr_oracle_fir_taps.argtypes = [C.c_int, C.c_int, C.c_float, f32p]
        o.sdr_oracle_fir_decimate.restype = C.c_int
        o.sdr_oracle_fir_decimate.argtypes = [f32p, C.c_int, C.c_int, C.c_float, f32p, C.c_int]
        o.sdr_oracle_remove_dc.argtypes = [f32p, C.c_int, C.c_float]
        o.sdr_oracle_iir2_init_lowpass.argtypes = [C.POINTER(Biquad), C.c_float, C.c_float, C.c_float]
        o.sdr_oracle_iir2_process.argtypes = [C.POINTER(Biquad), f32p, C.c_int]
        o.sdr_oracle_demod_ssb.argtypes = [f32p, C.c_int, C.c_int, f32p]
        o.sdr_oracle_adaptive_agc.argtypes = [f32p, C.c_int, C.c_float, C.c_float, C.c_float]
        o.sdr_oracle_biquad_init_highpass.argtypes = [C.POINTER(Biquad), C.c_float, C.c_float, C.c_float]
        o.sdr_oracle_biquad_init_bandpass.argtypes = [C.POINTER(Biquad), C.c_float, C.c_float, C.c_float]
        o.sdr_oracle_biquad_process.argtypes = [C.POINTER(Biquad), f32p, C.c_int]
        o.sdr_oracle_transient_boost.argtypes = [f32p, C.c_int, C.c_float]
        o.sdr_oracle_float_to_pcm.argtypes = [f32p, C.c_int, C.c_float, i16p]
        # extension stages (oracle/sdr_oracle_ext.c): definitions without reference code
        o.sdr_oracle_window.argtypes = [C.c_int, C.c_int, f32p]
        o.sdr_oracle_fft_front.argtypes = [f32p, C.c_int, C.c_int, C.c_int, f32p]
        o.sdr_oracle_nco.argtypes = [C.c_uint32, f32p, f32p]
        o.sdr_oracle_nco_inc.restype = C.c_uint32
        o.sdr_oracle_nco_inc.argtypes = [C.c_double, C.c_uint32]
        o.sdr_oracle_ssbx_init.argtypes = [C.POINTER(SsbxState)]
        o.sdr_oracle_process_ssbx.restype = C.c_int
        o.sdr_oracle_process_ssbx.argtypes = [C.POINTER(SsbxState), f32p, C.c_int, C.c_uint32, C.c_int, C.c_int,
                                              C.c_int, C.c_double, i16p, C.c_int]
        _oracle = o
    return _oracle


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    """The reference itself (FFTProcessor + ssb_demod_opt.cpp + vendored FFTW) behind a C harness."""
    global _ref
    if _ref is None:
        r = C.CDLL(REF_SO)
        r.sdr_ref_fft_create.restype = C.c_void_p
        r.sdr_ref_fft_destroy.argtypes = [C.c_void_p]
        r.sdr_ref_fft_configure.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int]
        r.sdr_ref_set_center_frequency_changed.argtypes = [C.c_int]
        r.sdr_ref_fft_process.argtypes = [C.c_void_p, f32p, C.c_uint32, C.c_int64]
        r.sdr_ref_fft_results.argtypes = [C.c_void_p, f32p, C.POINTER(C.c_int64), C.POINTER(C.c_int)]
        r.sdr_ref_fft_power.restype = C.c_int
        r.sdr_ref_fft_power.argtypes = [C.c_void_p, f32p, C.c_int]
        r.sdr_ref_convertIQ.argtypes = [C.c_void_p, C.c_int, f32p]
        r.sdr_ref_removeDC.argtypes = [f32p, C.c_int, C.c_float]
        r.sdr_ref_iir2_init_lowpass.argtypes = [f32p, C.c_float, C.c_float, C.c_float]
        r.sdr_ref_iir2_process.argtypes = [f32p, f32p, C.c_int]
        r.sdr_ref_demodSSB.argtypes = [f32p, C.c_int, C.c_int, f32p]
        r.sdr_ref_adaptiveAGC.argtypes = [f32p, C.c_int, C.c_float, C.c_float, C.c_float]
        r.sdr_ref_fir_decimate.restype = C.c_int
        r.sdr_ref_fir_decimate.argtypes = [f32p, C.c_int, C.c_int, C.c_float, f32p, C.c_int]
        r.sdr_ref_biquad_init_highpass.argtypes = [f32p, C.c_float, C.c_float, C.c_float]
        r.sdr_ref_biquad_init_bandpass.argtypes = [f32p, C.c_float, C.c_float, C.c_float]
        r.sdr_ref_biquad_process.argtypes = [f32p, f32p, C.c_int]
        r.sdr_ref_transient_boost.argtypes = [f32p, C.c_int, C.c_float]
        r.sdr_ref_float_to_pcm.argtypes = [f32p, C.c_int, C.c_float, i16p]
        r.sdr_ref_process_ssb.restype = C.c_int
        r.sdr_ref_process_ssb.argtypes = [f32p, C.c_int, C.c_uint32, C.c_int, C.c_int, i16p, C.c_int]
        r.sdr_ref_run_frames.restype = C.c_double
        r.sdr_ref_run_frames.argtypes = [C.c_void_p, f32p, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_int, C.c_int,
                                         C.c_int]
        _ref = r
    return _ref


# ----------------------------------------------------------------------------- helpers
def to_cf32(raw, fmt):
    """Wire format [.., n, 2] -> interleaved float32 [.., n, 2] exactly like the ingest stage."""
    o = oracle()
    raw = np.ascontiguousarray(raw)
    out = np.empty(raw.shape, dtype=np.float32)
    n = raw.size // 2
    if fmt == 0:
        o.sdr_oracle_convert_u8(raw.ctypes.data_as(C.c_void_p), n, _fp(out))
    elif fmt == 1:
        o.sdr_oracle_convert_s16(raw.ctypes.data_as(C.c_void_p), n, _fp(out))
    else:
        out[...] = raw
    return out


def oracle_power(iq_f32):
    """Shifted linear power of one block (double-precision DFT rounded once)."""
    iq = np.ascontiguousarray(iq_f32, dtype=np.float32)
    n = iq.shape[0]
    out = np.empty(n, dtype=np.float32)
    rc = oracle().sdr_oracle_power_shifted(_fp(iq), n, _fp(out))
    assert rc == 0
    return out


RESULT_FIELDS = ("meanSnrDb", "meanSnrSigma", "peakAboveNoiseMeanDb", "maxBinSnrDb", "maxBinSnrSigma",
                 "best1kHzSnrDb", "best1kHzSnrSigma", "best1kHzCenterFreqHz", "perBinMean")


class OracleFft:
    """Per-stream FFTProcessor restatement driven by the frame clock."""

    def __init__(self, fc, fs, n, focus_khz):
        self.st = FftState()
        self.n, self.fs = n, fs
        self.frame = 0
        oracle().sdr_oracle_fft_init(C.byref(self.st), fc, fs, n, focus_khz)

    def set_center(self, fc):
        self.st.centerFrequency = fc
        self.st.centerChanged = 1

    def evaluate(self, power_shifted):
        o = oracle()
        p = np.ascontiguousarray(power_shifted, dtype=np.float32)
        now = o.sdr_oracle_frame_clock_ns(self.frame, self.n, self.fs)
        self.frame += 1
        o.sdr_oracle_evaluate(C.byref(self.st), self.n, _fp(p), now)
        d = {k: getattr(self.st, k) for k in RESULT_FIELDS}
        d.update(trackingFrequencyHz=self.st.trackingFrequency,
                 trackingFrequency=o.sdr_oracle_tracking_long(C.byref(self.st)),
                 detectionFlag=self.st.detectionFlag, absPeakDb=self.st.absPeakDb,
                 signalPowerDb=self.st.signalPowerDb, peakBin=self.st.peakBin, valid=self.st.valid,
                 maxPeakDb=self.st.maxPeakDb, maxPeakHz=self.st.maxPeakHz)
        return d

    def process(self, iq_f32):
        return self.evaluate(oracle_power(iq_f32))


class OracleSsb:
    """Per-stream processSSB_opt restatement (explicit state instead of function statics)."""

    def __init__(self):
        self.st = SsbState()
        oracle().sdr_oracle_ssb_init(C.byref(self.st))

    def process(self, iq_f32, fs, upper=1, mode=1):
        iq = np.ascontiguousarray(iq_f32, dtype=np.float32)
        n = iq.shape[0]
        cap = n + 8
        out = np.zeros(cap, dtype=np.int16)
        m = oracle().sdr_oracle_process_ssb(C.byref(self.st), _fp(iq), n, fs, upper, mode,
                                            out.ctypes.data_as(i16p), cap)
        return out[:m].copy()


class RefFft:
    """The reference's FFTProcessor (FFTW) under the frame clock."""

    def __init__(self, fc, fs, n, focus_khz):
        r = ref()
        self.h = r.sdr_ref_fft_create()
        self.n, self.fs, self.frame = n, fs, 0
        r.sdr_ref_fft_configure(self.h, fc, fs, n, focus_khz)

    def close(self):
        if self.h:
            ref().sdr_ref_fft_destroy(self.h)
            self.h = None

    def process(self, iq_f32):
        r = ref()
        iq = np.ascontiguousarray(iq_f32, dtype=np.float32)
        now = oracle().sdr_oracle_frame_clock_ns(self.frame, self.n, self.fs)
        self.frame += 1
        r.sdr_ref_fft_process(self.h, _fp(iq), self.n, now)
        out = np.zeros(13, dtype=np.float32)
        tl, fl = C.c_int64(), C.c_int()
        r.sdr_ref_fft_results(self.h, _fp(out), C.byref(tl), C.byref(fl))
        d = {k: float(out[i]) for i, k in enumerate(RESULT_FIELDS)}
        d.update(trackingFrequencyHz=float(out[9]), maxPeakDb=float(out[10]), maxPeakHz=float(out[11]),
                 trackingFrequency=tl.value, detectionFlag=fl.value)
        return d

    def power(self):
        out = np.empty(self.n, dtype=np.float32)
        ref().sdr_ref_fft_power(self.h, _fp(out), self.n)
        return out


def oracle_fft_front(iq_f32, window_kind, remove_dc):
    """Extension front end of the FFT path: (x - block mean) * window (oracle/sdr_oracle_ext.c)."""
    iq = np.ascontiguousarray(iq_f32, dtype=np.float32)
    out = np.empty_like(iq)
    oracle().sdr_oracle_fft_front(_fp(iq), iq.shape[0], window_kind, remove_dc, _fp(out))
    return out


class OracleSsbx:
    """One channel of the extended SSB chain (NCO mix + optional Weaver sideband selection)."""

    def __init__(self, offset_hz=0.0, true_sideband=0):
        self.st = SsbxState()
        self.offset, self.tsb = float(offset_hz), int(true_sideband)
        oracle().sdr_oracle_ssbx_init(C.byref(self.st))

    def process(self, iq_f32, fs, upper=1, mode=1):
        iq = np.ascontiguousarray(iq_f32, dtype=np.float32)
        n = iq.shape[0]
        out = np.zeros(n + 8, dtype=np.int16)
        m = oracle().sdr_oracle_process_ssbx(C.byref(self.st), _fp(iq), n, fs, upper, mode, self.tsb, self.offset,
                                             out.ctypes.data_as(i16p), n + 8)
        return out[:m].copy()
