"""Generates tests/golden/*.npz from the REFERENCE ITSELF (oracle/_ref/libsdr_ref.so = the reference's
own fft_process.cpp / ssb_demod_opt.cpp + vendored FFTW, built by oracle/Makefile in the authoring
container where /root/reference exists).  The inputs are not stored: they are the deterministic
synthetic source (sdrb_synth_host), identified by (format, fs, stream, first block, n).

    python tests/golden/make_golden.py        # rewrites the fixtures

Each SSB case runs in a fresh process because processSSB_opt keeps its state in function statics."""
import importlib
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

FFT_CASES = [  # name, fmt, fs, n, focus_khz, stream, blocks
    ("fft_u8_16k", 0, 2_000_000, 16384, 5, 0, 24),
    ("fft_u8_16k_wide", 0, 2_000_000, 16384, 200, 4, 6),
    ("fft_s16_64k", 1, 30_720_000, 65536, 5, 1, 4),
    ("fft_f32_1k", 2, 250_000, 1024, 5, 2, 8),
]
SSB_CASES = [  # name, fmt, fs, n, mode, upper, stream, blocks
    ("ssb_u8_usb_m1", 0, 2_000_000, 16384, 1, 1, 0, 6),
    ("ssb_u8_usb_m0", 0, 2_000_000, 16384, 0, 1, 3, 3),
    ("ssb_u8_lsb", 0, 2_000_000, 16384, 1, 0, 0, 2),
    ("ssb_s16_64k", 1, 30_720_000, 65536, 2, 1, 1, 3),
]


def gen_fft(case):
    import oracle_lib as ol
    sdrb = importlib.import_module("sdr-for-android-lib_b200")
    name, fmt, fs, n, focus, stream, blocks = case
    rf = ol.RefFft(430_000_000, fs, n, focus)
    keys = ol.RESULT_FIELDS + ("trackingFrequencyHz", "maxPeakDb", "maxPeakHz")
    rec = {k: [] for k in keys}
    rec["trackingFrequency"], rec["detectionFlag"] = [], []
    spectra = []
    for b in range(blocks):
        iq = ol.to_cf32(sdrb.synth_host(fmt, fs, stream, 1, b * n, n)[0], fmt)
        r = rf.process(iq)
        for k in rec:
            rec[k].append(r[k])
        if b < 2:
            spectra.append(rf.power())
    out = {k: np.array(v, dtype=np.int64 if k in ("trackingFrequency", "detectionFlag") else np.float32)
           for k, v in rec.items()}
    # the 16K / 64K spectra are stored decimated to keep the fixture small: every 16th bin + the focus band
    sp = np.stack(spectra)
    out["power_every16"] = sp[:, ::16].copy()
    out["power_center"] = sp[:, n // 2 - 128:n // 2 + 128].copy()
    if name == "fft_u8_16k":
        out["power_full"] = sp.copy()   # two whole FFTW spectra: the GPU-vs-FFTW error per bin class is measured on them
    out["meta"] = np.array([fmt, fs, n, focus, stream, blocks], dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)


def gen_ssb(case):
    import oracle_lib as ol
    sdrb = importlib.import_module("sdr-for-android-lib_b200")
    name, fmt, fs, n, mode, upper, stream, blocks = case
    pcm = []
    for b in range(blocks):
        iq = np.ascontiguousarray(ol.to_cf32(sdrb.synth_host(fmt, fs, stream, 1, b * n, n)[0], fmt))
        out = np.zeros(n + 8, np.int16)
        m = ol.ref().sdr_ref_process_ssb(ol._fp(iq), n, fs, upper, mode, out.ctypes.data_as(ol.i16p), n + 8)
        pcm.append(out[:m].copy())
    np.savez_compressed(os.path.join(HERE, name + ".npz"), pcm=np.stack(pcm),
                        meta=np.array([fmt, fs, n, mode, upper, stream, blocks], dtype=np.int64))


if __name__ == "__main__":
    if len(sys.argv) == 3:
        kind, idx = sys.argv[1], int(sys.argv[2])
        (gen_fft if kind == "fft" else gen_ssb)((FFT_CASES if kind == "fft" else SSB_CASES)[idx])
    else:
        for i in range(len(FFT_CASES)):
            subprocess.check_call([sys.executable, __file__, "fft", str(i)])
        for i in range(len(SSB_CASES)):
            subprocess.check_call([sys.executable, __file__, "ssb", str(i)])
        print("golden fixtures written to", HERE)
