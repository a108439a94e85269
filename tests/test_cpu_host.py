"""CPU-side tests (no GPU): the C-ABI library loads and exports every symbol of include/sdr_b200.h,
fails loudly without a device, and its host-side parameter design (geometry, coefficients, FIR taps,
synthetic source) agrees with the oracle."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle_lib as ol

ROOT = ol.ROOT


def test_every_declared_symbol_is_exported(sdrb):
    hdr = open(os.path.join(ROOT, "include", "sdr_b200.h")).read()
    declared = set(re.findall(r"\b(sdrb_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 28
    for name in sorted(declared):
        assert hasattr(sdrb.lib, name), f"{name} declared in include/sdr_b200.h but not exported"
    assert declared == set(sdrb.binding.EXPORTS), declared ^ set(sdrb.binding.EXPORTS)
    assert sdrb.lib.sdrb_abi_version() == 2
    assert [sdrb.format_bytes(f) for f in (0, 1, 2, 3, 4)] == [2, 4, 8, 4, 0]


def test_struct_layouts_match_header(sdrb):
    # sizes the C side was compiled with (sdrb_frame_result has an int64 at offset 40)
    assert C.sizeof(sdrb.Config) == 64
    assert C.sizeof(sdrb.FrameResult) == 80 and sdrb.FrameResult.trackingFrequency.offset == 40
    assert sdrb.RESULT_DTYPE.itemsize == 80


def test_no_cpu_fallback(sdrb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    rc = sdrb.lib.sdrb_create(0, C.byref(h))
    assert rc == -2 and not h.value
    assert b"no CPU fallback" in sdrb.lib.sdrb_last_error(None)
    with pytest.raises(sdrb.SdrbError):
        sdrb.Context(0)


def ref_geometry(fs, n, focus_khz):
    """evaluateSignalStrength's bin bounds in float32 numpy (src/dsp/fft_process.cpp:125-139,191-216)."""
    f32 = np.float32
    fpb = f32(fs) / f32(n)
    X = f32(focus_khz) * f32(1000.0)
    nyq = f32(fs) / f32(2.0)
    o2b = lambda off: int((f32(off) + nyq) / fpb)
    lo, hi = max(0, o2b(-X)), min(n - 1, o2b(X) - 1)
    wins = []
    for k in range(1, 6):
        near, far = f32(4 * k - 2) * X, f32(4 * k) * X
        if far >= nyq:
            break
        for a, b in ((max(0, o2b(near)), min(n - 1, o2b(far) - 1)), (max(0, o2b(-far)), min(n - 1, o2b(-near) - 1))):
            if b > a:
                wins.append((a, b))
    return lo, hi, wins


@pytest.mark.parametrize("fs,n,focus", [(2_000_000, 16384, 5), (2_000_000, 16384, 200), (2_500_000, 16384, 5),
                                        (30_720_000, 65536, 5), (10_000_000, 1 << 20, 5), (250_000, 1024, 5),
                                        (2_000_000, 16384, 0), (2_000_000, 16384, 300), (61_440_000, 65536, 25)])
def test_plan_geometry(sdrb, fs, n, focus):
    p = sdrb.plan_config(sdrb.Config(430_000_000, fs, n, focus))
    lo, hi, wins = ref_geometry(fs, n, focus)
    assert (p.focusLo, p.focusHi, p.focusLen) == (lo, hi, hi - lo + 1)
    if hi - lo + 1 > 0:
        assert p.nWin == len(wins)
        assert [(p.winLo[i], p.winHi[i]) for i in range(p.nWin)] == wins
        assert p.nBottom == max(1, int(np.float32(len(wins)) * np.float32(0.4)))
        assert p.winBins1k == max(1, int(np.ceil(np.float32(1000.0) / (np.float32(fs) / np.float32(n)))))


def test_plan_defaults_match_survey(sdrb):
    p = sdrb.plan_config(sdrb.Config())         # fs 2 MS/s, N 16384, focus 5 kHz
    assert (p.focusLo, p.focusHi, p.focusLen, p.winBins1k, p.nWin, p.nBottom) == (8151, 8231, 81, 9, 10, 4)
    assert (p.decim, p.nTaps, p.pcmPerBlock, p.fftPath) == (41, 255, 394, 1)


@pytest.mark.parametrize("fs,n,mode", [(2_000_000, 16384, 1), (2_500_000, 16384, 0), (30_720_000, 65536, 2),
                                       (250_000, 4096, 1), (61_440_000, 65536, 1), (2_000_000, 128, 1)])
def test_plan_ssb_parameters_bit_exact(sdrb, fs, n, mode):
    o = ol.oracle()
    p = sdrb.plan_config(sdrb.Config(430_000_000, fs, n, 5, soundMode=mode))
    decim = max(1, int(np.float32(fs) / np.float32(48000.0)))
    assert p.decim == decim
    h = np.zeros(256, np.float32)
    nt = o.sdr_oracle_fir_taps(n, decim, 0.45, ol._fp(h))
    assert p.nTaps == nt and np.array(p.firTaps[:nt], np.float32).tobytes() == h[:nt].tobytes()
    assert p.pcmPerBlock == ((n - nt) // decim + 1 if n >= nt else 0)
    st = ol.SsbState()
    o.sdr_oracle_ssb_init(C.byref(st))
    pcm = np.zeros(n + 8, np.int16)
    o.sdr_oracle_process_ssb(C.byref(st), ol._fp(np.zeros((n, 2), np.float32)), n, fs, 1, mode,
                             pcm.ctypes.data_as(ol.i16p), n + 8)
    for mine, theirs in ((p.rf, st.rf), (p.hp, st.hp), (p.bp, st.bp)):
        a = np.array(list(mine), np.float32)
        b = np.array([theirs.a0, theirs.a1, theirs.a2, theirs.b1, theirs.b2], np.float32)
        assert a.tobytes() == b.tobytes()
    assert (np.float32(p.agcTarget), np.float32(p.agcFast), np.float32(p.gain), np.float32(p.transientCoeff)) == \
           (np.float32(st.agcTarget), np.float32(st.agcFast), np.float32(st.gain), np.float32(st.transientCoeff))


def test_config_errors(sdrb):
    for bad in (sdrb.Config(sampleRate=0), sdrb.Config(samplesPerReading=0), sdrb.Config(inputFormat=7)):
        with pytest.raises(sdrb.SdrbError) as e:
            sdrb.plan_config(bad)
        assert e.value.code == -1
    for bad in (sdrb.Config(fftWindow=2), sdrb.Config(ssbChannels=-1), sdrb.Config(ssbChannels=5000),
                sdrb.Config(ssbChannelOffsetHz=float("nan"))):
        with pytest.raises(sdrb.SdrbError) as e:
            sdrb.plan_config(bad)
        assert e.value.code == -1
    # extension stages plan like the reference chain; the true-sideband low-pass is designed at B/2
    a, b = sdrb.plan_config(sdrb.Config()), sdrb.plan_config(sdrb.Config(ssbTrueSideband=1, fftWindow=1))
    assert a.pcmPerBlock == b.pcmPerBlock and a.rf[0] > b.rf[0] > 0


def test_synth_source_properties(sdrb):
    fs, n = 2_000_000, 1 << 16
    a = sdrb.synth_host(0, fs, 5, 2, 1000, n)
    b = sdrb.synth_host(0, fs, 5, 2, 1000, n)
    assert np.array_equal(a, b)                                   # deterministic
    c = sdrb.synth_host(0, fs, 5, 1, 1000 + 777, 100)             # random access == sequential
    assert np.array_equal(c[0], a[0, 777:877])
    assert not np.array_equal(a[0], a[1])                         # streams differ
    x = a.astype(np.float64) - 127.4
    assert abs(x.mean()) < 0.05                                   # no DC spike after the 127.4 offset
    s16 = sdrb.synth_host(1, fs, 5, 1, 0, 4096)
    f32 = sdrb.synth_host(2, fs, 5, 1, 0, 4096)
    assert np.array_equal(f32, s16.astype(np.float32) / 32768.0)  # CF32 = CS16 * 2^-15 exactly
    # the keyed tone of a keyed-on stream (s % 4 == 0) is inside the +-5 kHz focus band
    k = sdrb.synth_host(2, fs, 8, 1, 0, 16384)[0]
    spec = np.abs(np.fft.fftshift(np.fft.fft(k[:, 0] + 1j * k[:, 1]))) ** 2
    p = sdrb.plan_config(sdrb.Config())
    assert p.focusLo <= int(np.argmax(spec)) <= p.focusHi


def test_dropin_front_end_exports_reference_interface():
    """libsdr_b200_dropin.so defines the reference's C++ entry points (src/dsp/fft_process.h:29-35,
    src/ssb/ssb_demod_opt.h:64-65, src/ssb/ssb_processor.h:22-31) and binds only the C-ABI."""
    import subprocess
    lib = os.path.join(ROOT, "sdr-for-android-lib_b200", "libsdr_b200_dropin.so")
    assert os.path.exists(lib), "run __graft_entry__.build()"
    syms = subprocess.run(["nm", "-D", "-C", lib], capture_output=True, text=True, check=True).stdout
    defined = [l for l in syms.splitlines() if " T " in l or " W " in l]
    for want in ("FFTProcessor::FFTProcessor()", "FFTProcessor::configure(FftProcessorConfig const&)",
                 "FFTProcessor::process(std::complex<float> const*, unsigned int)",
                 "processSSB_opt(std::vector<std::complex<float>", "lastBlockContainsPulse()",
                 "SSBProcessor::startProcessing(", "SSBProcessor::stopProcessing()",
                 "SSBProcessor::enqueueData(std::vector<std::complex<float>"):
        assert any(want in l for l in defined), want
    undefined = [l.split()[-1] for l in syms.splitlines() if " U " in l]
    assert any(u.startswith("sdrb_process_host") for u in undefined)
    assert not any(u.startswith("cuda") or u.startswith("cu") and u[2:3].isupper() for u in undefined)


def test_bridge_chunker_and_replay(tmp_path):
    """host/iq_chunker.h: exact blocks from ragged reads, drop-oldest at 20, 60 ms
    consumer wait, file replay and batch assembly — the reader thread of src/sdr-bridge-java-soapy.cpp:496-575."""
    import subprocess
    exe = str(tmp_path / "chunker_test")
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", "-I", os.path.join(ROOT, "sdr-for-android-lib_b200", "host"),
                    "-o", exe, os.path.join(ROOT, "tests", "cpp", "chunker_test.cpp")], check=True)
    r = subprocess.run([exe, str(tmp_path / "cap.iq")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "chunker OK" in r.stdout, r.stdout + r.stderr


def test_bench_numpy_source_equals_the_library_generator(sdrb):
    """bench_synth.py (numpy) feeds the CPU legs of bench.py so that they never map the product library; it must be the
    same integer generator as sdrb_synth_host (csrc/synth.cuh), byte for byte, for every wire format."""
    import bench_synth as bs
    for fmt in (sdrb.CU8, sdrb.CS16, sdrb.CF32):
        for fs, stream, first in ((2_000_000, 0, 0), (30_720_000, 5, 3_400_000), (2_500_000, 70_000, 5_000_000_000)):
            a = sdrb.synth_host(fmt, fs, stream, 1, first, 4000)[0]
            b = bs.synth(fmt, fs, stream, first, 4000)
            assert a.dtype == b.dtype and np.array_equal(a, b), (fmt, fs, stream, first)
    raw = bs.synth(sdrb.CU8, 2_000_000, 1, 0, 64)
    assert np.array_equal(bs.to_cf32(raw, sdrb.CU8), ol.to_cf32(raw, sdrb.CU8))


def test_reference_arm_never_loads_the_product(tmp_path):
    """`bench.py --impl reference` times the reference's CPU code only: it must run with the product library out of
    reach (SDRB_LIB pointing nowhere makes any import of the package raise) and print the contract's JSON line with
    the same `config` keys as our arm."""
    import json
    import subprocess
    import sys
    env = dict(os.environ, SDRB_LIB=str(tmp_path / "absent.so"))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and d["unit"] == "MS/s" and d["value"] > 0 and d["gpu_launches"] == 0
    assert "ssb" in d["config"]["stages"] and "fft" in d["config"]["stages"]
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    import bench
    ours = bench.config_dict(dict(bench.WORKLOADS["default"], name="default"), 1)
    assert set(ours) == set(d["config"]) and ours == d["config"]
