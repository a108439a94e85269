#!/usr/bin/env python
"""bench.py — IQ mega-samples/s through the batched hot path on N B200s (one process per GPU).

Default workload = the pipeline the metric names, FFT + peak + SSB over ONE buffer of raw IQ
(src/sdr-bridge-java-soapy.cpp:441-444 hands each reading to both processors): the per-stream work of
BASELINE.json configs[0] (RTL-SDR u8 at 2 MS/s, 16384-pt FFT + power/fftshift + peak + signal strength +
USB demod to ~48 kHz int16 PCM), batched over thousands of independent streams per GPU like configs[1].
One step = one pass over one batch of `--blocks` readings per stream; streams are sharded by rank with
no data-path collective ("scaling": "weak": per-GPU work is fixed as N grows).

    python bench.py [--gpus N --steps K --warmup W]            # our arm, default workload
    python bench.py --config c2|c1|c3|c3ref|c3fft|c4|c5|c5ru16  # the other BASELINE configurations
    python bench.py --config n1024|n4096|n8192|n32768|n4096ssb  # other block lengths
    python bench.py --impl reference [...]                     # the reference's CPU path, host cores

Prints ONE JSON line (rank 0).  `value` is device-timed (CUDA events on the launching stream) with the
inputs resident in HBM; `e2e` is the same metric through the C-ABI with HOST buffers (H2D of the IQ and
D2H of records + spectra + PCM inside the timed region); `roofline` is the dominant kernel against the
measured HBM copy peak, with the whole step beside it; `cpu_baseline` is the reference's own code
(oracle/_ref) on this box's host cores; `sustained` repeats the step for seconds with the clocks sampled.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FC = 430_000_000
FOCUS_KHZ = 5
CU8, CS16, CF32, RU16 = 0, 1, 2, 3
F_FFT, F_SSB = 1, 2
FMT_NAME = {CU8: "cu8", CS16: "cs16", CF32: "cf32", RU16: "ru16"}
FMT_BYTES = {CU8: 2, CS16: 4, CF32: 8, RU16: 4}
METRIC = "IQ MS/s through FFT+peak+SSB at 1/2/4/8 B200, % HBM roofline, vs host CPU"

# name -> workload.  S = streams per GPU, B = readings per stream per step.  The C3 shapes take 32 readings per call
# (68 ms of signal): with 256 streams the per-(stream, block) passes of the SSB chain only have S x B lanes to fill the GPU.
WORKLOADS = {
    "default": dict(
        what="65536 independent RTL-SDR u8 streams per GPU, one reading of each per step (the reference's own call shape, one "
             "soapyCallback per reading), 2 MS/s: 16384-pt FFT + power/fftshift + peak + signal strength + USB demod to "
             "int16 PCM (the per-stream work of BASELINE configs[0], batched like configs[1])",
        S=65536, B=1, n=16384, fs=2_000_000, fmt=CU8, stages=F_FFT | F_SSB, extra={}),
    "c2": dict(what="1024 independent RTL-SDR u8 streams per GPU, 16384-pt FFT + power/fftshift + peak + signal strength "
                    "only (BASELINE configs[1])",
               S=1024, B=37, n=16384, fs=2_000_000, fmt=CU8, stages=F_FFT, extra={}),
    "c1": dict(what="ONE RTL-SDR u8 stream, 2 MS/s, 16384-pt FFT + peak + strength + USB demod, 64 readings per call "
                    "(BASELINE configs[0], the CPU reference's own case)",
               S=1, B=64, n=16384, fs=2_000_000, fmt=CU8, stages=F_FFT | F_SSB, extra={}),
    "c3": dict(what="256 LimeSDR-style cs16 streams, 30.72 MS/s, 65536-pt FFT + true-LSB demod (BASELINE configs[2]; "
                    "extension stage: the reference's LSB output is all zeros)",
               S=256, B=32, n=65536, fs=30_720_000, fmt=CS16, stages=F_FFT | F_SSB,
               extra={"upperSideband": 0, "ssbTrueSideband": 1}),
    "c3ref": dict(what="256 cs16 streams, 30.72 MS/s, 65536-pt FFT + USB demod in reference mode (configs[2] without the "
                       "extension stage)",
                  S=256, B=32, n=65536, fs=30_720_000, fmt=CS16, stages=F_FFT | F_SSB, extra={}),
    "c3fft": dict(what="256 cs16 streams, 65536-pt FFT + peak + strength only",
                  S=256, B=32, n=65536, fs=30_720_000, fmt=CS16, stages=F_FFT, extra={}),
    "c4": dict(what="wideband cf32 61.44 MS/s, 256 SSB channels per stream on a 25 kHz raster, SSB only (BASELINE "
                    "configs[3]; extension stage, no reference code)",
               S=32, B=32, n=16384, fs=61_440_000, fmt=CF32, stages=F_SSB,
               extra={"ssbChannels": 256, "ssbChannelOffsetHz": -3.2e6, "ssbChannelSpacingHz": 25e3}),
    "c5": dict(what="cf32 10 MS/s, 1048576-pt FFT (four-step) + peak tracking (BASELINE configs[4])",
               S=16, B=4, n=1048576, fs=10_000_000, fmt=CF32, stages=F_FFT, extra={}),
    "c5ru16": dict(what="raw Airspy 12-bit ADC words (real->IQ on the GPU), 65536-pt FFT + peak tracking",
                   S=256, B=8, n=65536, fs=10_000_000, fmt=RU16, stages=F_FFT, extra={}),
    # other block lengths (samplesPerReading is a run-time setting of the reference): the same number of samples per call
    "n1024": dict(what="u8 streams, 1024-pt FFT + peak + strength (one-CTA shared-memory kernel)",
                  S=16384, B=16, n=1024, fs=2_000_000, fmt=CU8, stages=F_FFT, extra={}),
    "n4096": dict(what="u8 streams, 4096-pt FFT + peak + strength (one-CTA shared-memory kernel)",
                  S=8192, B=8, n=4096, fs=2_000_000, fmt=CU8, stages=F_FFT, extra={}),
    "n8192": dict(what="u8 streams, 8192-pt FFT + peak + strength (one-CTA shared-memory kernel, 32 x 16 x 16)",
                  S=4096, B=8, n=8192, fs=2_000_000, fmt=CU8, stages=F_FFT, extra={}),
    "n32768": dict(what="u8 streams, 32768-pt FFT + peak + strength (two passes: radix-8 columns + 4096-point rows)",
                   S=1024, B=8, n=32768, fs=2_000_000, fmt=CU8, stages=F_FFT, extra={}),
    "n4096ssb": dict(what="u8 streams, 2 MS/s, 4096-pt FFT + peak + strength + USB demod, one reading per call",
                     S=65536, B=1, n=4096, fs=2_000_000, fmt=CU8, stages=F_FFT | F_SSB, extra={}),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="default", choices=sorted(WORKLOADS))
    ap.add_argument("--streams", type=int, default=0, help="streams per GPU (0 = the workload's own)")
    ap.add_argument("--blocks", type=int, default=0, help="readings per stream per step (0 = the workload's own)")
    ap.add_argument("--e2e-steps", type=int, default=8)
    ap.add_argument("--e2e-frames", type=int, default=16384,
                    help="readings per host-to-host call of the e2e leg (bounds the pinned host memory per rank: 0.5 GiB of "
                         "u8 IQ in, 1 GiB of spectra out at 16384 points)")
    ap.add_argument("--sustain-seconds", type=float, default=3.0)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def workload(args):
    w = dict(WORKLOADS[args.config])
    if args.streams:
        w["S"] = args.streams
    if args.blocks:
        w["B"] = args.blocks
    w["name"] = args.config
    return w


def decim_of(fs):
    return max(1, int(np.float32(fs) / np.float32(48000.0)))     # src/ssb/ssb_demod_opt.cpp:273


def pcm_per_block(n, fs):
    d, taps = decim_of(fs), 255 if n >= 255 else (n | 1)
    return (n - taps) // d + 1 if n >= taps else 0                 # src/ssb/ssb_demod_opt.cpp:135


def stage_names(w):
    parts = []
    if w["stages"] & F_FFT:
        parts.append("fft+power+fftshift+estimator+tracker")
    if w["stages"] & F_SSB:
        parts.append("ssb(dc+iir2+usb/lsb+agc+fir-decimate+hp/bp+boost+pcm)")
    return "+".join(parts)


def config_dict(w, world):
    """Identical in both arms (the reference arm times a bounded sample of this same workload)."""
    S, B, n = w["S"], w["B"], w["n"]
    bi = FMT_BYTES[w["fmt"]]
    return {"workload": w["what"], "name": w["name"], "streams_per_gpu": S, "readings_per_stream_per_step": B,
            "fft": n, "format": FMT_NAME[w["fmt"]], "sample_rate": w["fs"], "focus_khz": FOCUS_KHZ,
            "stages": stage_names(w), "extra": {k: v for k, v in w["extra"].items()},
            "parallelism": f"streams sharded x{world}",
            "l2": f"2 alternating batches of {S * B * n * bi >> 20} MiB in per step (> 126 MB L2)"}


def algorithmic_bytes_per_sample(w):
    """one read of the raw sample, one f32 power bin, 2/decim B of PCM (per channel)."""
    b = float(FMT_BYTES[w["fmt"]])      # RU16: two 16-bit real ADC words per complex sample
    if w["stages"] & F_FFT:
        b += 4.0
    if w["stages"] & F_SSB:
        ch = max(1, int(w["extra"].get("ssbChannels", 1)))
        b += 2.0 * ch * pcm_per_block(w["n"], w["fs"]) / w["n"]
    return b


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def mark(self):
        return time.perf_counter()

    def summary(self, t0=None, t1=None):
        rows = [r for (t, r) in self.rows if (t0 is None or t >= t0) and (t1 is None or t <= t1)]
        sm = [float(r[0]) for r in rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        pw = [float(r[6]) for r in rows if len(r) > 6 and r[6].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_min_mhz": min(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "power_w_max": max(pw) if pw else None,
                "reasons": reasons, "samples": len(sm)}

    def stop(self):
        if not self.proc:
            return
        time.sleep(0.12)
        self.proc.terminate()
        self.t.join(timeout=2)


# ----------------------------------------------------------------------------- CPU arm
# The only place besides tests/ and smoke() that executes oracle/: as the thing TIMED for the CPU
# baseline (never on our arm's product path).  Input comes from bench_synth.py (numpy), so these
# processes never map the product library.
_W = {}


def _cpu_worker_init(tests_dir, root, stream_id, n, fs, fmt, focus_khz):
    sys.path.insert(0, tests_dir)
    sys.path.insert(0, root)
    import bench_synth as bs
    import oracle_lib as ol
    nbuf = 8 if n <= 65536 else 2
    src_fmt = CS16 if fmt == RU16 else fmt        # the reference is handed cf32 either way
    raw = bs.synth(src_fmt, fs, stream_id, 0, nbuf * n)
    _W.update(iq=np.ascontiguousarray(bs.to_cf32(raw, src_fmt)), nbuf=nbuf, ol=ol, n=n, fs=fs)
    if ol.have_ref():
        r = ol.ref()
        _W["h"] = r.sdr_ref_fft_create()
        r.sdr_ref_fft_configure(_W["h"], FC, fs, n, focus_khz)
        _W["kind"] = "reference"
    else:
        _W["of"], _W["os"] = ol.OracleFft(FC, fs, n, focus_khz), ol.OracleSsb()
        _W["kind"] = "port"
    return _W["kind"]


def _cpu_worker_run(frames, do_fft, do_ssb):
    """FFTProcessor::process and/or processSSB_opt on `frames` readings, exactly the two calls soapyCallback
    makes per reading (src/sdr-bridge-java-soapy.cpp:441-444); returns seconds."""
    ol, n, fs = _W["ol"], _W["n"], _W["fs"]
    t0 = time.perf_counter()
    if _W["kind"] == "reference":
        ol.ref().sdr_ref_run_frames(_W["h"], ol._fp(_W["iq"]), n, _W["nbuf"], frames, fs, int(do_fft), int(do_ssb), 1)
    else:
        iq = _W["iq"].reshape(_W["nbuf"], n, 2)
        for f in range(frames):
            if do_ssb:
                _W["os"].process(iq[f % _W["nbuf"]], fs, 1, 1)
            if do_fft:
                _W["of"].process(iq[f % _W["nbuf"]])
    return time.perf_counter() - t0


def _stage_label(do_fft, do_ssb):
    return "+".join(x for x, on in (("FFTProcessor::process", do_fft), ("processSSB_opt", do_ssb)) if on)


def cpu_single_core(w, seconds):
    """Rank 0, N = 1: the reference on ONE host core, a bounded sample of this workload; rows for FFT only and SSB
    only beside the full pipeline (the reference runs them on two threads: SURVEY.md 8(d))."""
    n, fs = w["n"], w["fs"]
    # extension configurations have no reference code: c3 is timed in reference mode (USB), c4 has no CPU arm
    do_fft = bool(w["stages"] & F_FFT)
    do_ssb = bool(w["stages"] & F_SSB) and (not w["extra"] or w["name"] == "c3")
    if not do_fft and not do_ssb:
        return None
    kind = _cpu_worker_init(os.path.join(ROOT, "tests"), ROOT, 0, n, fs, w["fmt"], FOCUS_KHZ)
    rows = {}

    def one(df, ds, secs):
        probe = 4 if n <= 65536 else 1
        t = _cpu_worker_run(probe, df, ds)
        frames = max(probe, int(secs / max(t / probe, 1e-6)))
        t = _cpu_worker_run(frames, df, ds)
        return frames, t

    frames, t = one(do_fft, do_ssb, seconds)
    if do_fft and do_ssb:
        f1, t1 = one(True, False, seconds / 4)
        f2, t2 = one(False, True, seconds / 4)
        rows = {"fft_only": {"value": f1 * n / t1 / 1e6, "unit": "MS/s", "readings": f1},
                "ssb_only": {"value": f2 * n / t2 / 1e6, "unit": "MS/s", "readings": f2}}
    out = {"value": frames * n / t / 1e6, "unit": "MS/s", "cores": 1, "kind": kind,
           "sample": f"{frames} readings of {n} {FMT_NAME[w['fmt']]}-derived cf32 samples of one stream through "
                     f"{_stage_label(do_fft, do_ssb)} on one host core, {t:.1f} s"}
    if rows:
        out["rows"] = rows
    return out


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation on all host cores, multi-process because the FFTW
    planner is not thread-safe, process() plans on every call and processSSB_opt keeps its filter state in
    function statics: one stream per process."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = workload(args)
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    if w["extra"] and w["name"] != "c3":
        print(json.dumps({"impl": "reference", "unavailable": f"config {w['name']} is an extension stage: the reference "
                                                             "has no code for it (SURVEY.md section 0 rows 6, 8)"}))
        return
    import multiprocessing as mp
    n, fs = w["n"], w["fs"]
    do_fft, do_ssb = bool(w["stages"] & F_FFT), bool(w["stages"] & F_SSB)
    cores = min(len(os.sched_getaffinity(0)) or 1, 128)
    ctx = mp.get_context("spawn")
    pools = [ctx.Pool(1, initializer=_cpu_worker_init,
                      initargs=(os.path.join(ROOT, "tests"), ROOT, i, n, fs, w["fmt"], FOCUS_KHZ)) for i in range(cores)]
    probe_n = 8 if n <= 65536 else 1
    probe = [p.apply_async(_cpu_worker_run, (probe_n, do_fft, do_ssb)) for p in pools]
    per_frame = max(r.get() for r in probe) / probe_n
    frames = max(2, int(1.5 / per_frame))          # each step ~1.5 s of wall clock
    kind = "reference" if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libsdr_ref.so")) else "port"

    def step():
        rs = [p.apply_async(_cpu_worker_run, (frames, do_fft, do_ssb)) for p in pools]
        for r in rs:
            r.get()

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    for p in pools:
        p.close()
    val = args.steps * cores * frames * n / dt / 1e6
    note = " (USB: the reference's LSB output is identically zero)" if w["name"] == "c3" else ""
    sample = (f"{cores} processes x {frames} readings of {n} samples per step through "
              f"{_stage_label(do_fft, do_ssb)} ({kind}), one stream per process{note}")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "MS/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(w, world),
        "cpu_baseline": {"value": val, "unit": "MS/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": "MS/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ----------------------------------------------------------------------------- our arm
def ncu_traffic(name):
    """Per-launch dram__bytes_read.sum + dram__bytes_write.sum from the committed `ncu --set full` captures
    (profiles/r02_traffic.json, keyed by kernel name), None when not captured."""
    for fn in ("r02_traffic.json", "r01_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", fn)) as f:
                v = json.load(f).get(name)
                if v is not None:
                    return v
        except (OSError, ValueError):
            pass
    return None


def fill_input(sdrb, torch, ctx, w, S, B, first_stream, batch_index):
    n, fmt = w["n"], w["fmt"]
    t = torch.empty((S * B * n * FMT_BYTES[fmt],), dtype=torch.uint8, device="cuda")
    if fmt == RU16:
        # 12-bit ADC words: tone near fs/4 + noise (the integer generator has no real-sample format)
        g = torch.Generator(device="cuda").manual_seed(1 + batch_index)
        i = torch.arange(S * B * n * 2, device="cuda", dtype=torch.float32)
        x = 2048 + 600 * torch.cos(2 * np.pi * 0.263 * i) + 40 * torch.randn(i.shape, device="cuda", generator=g)
        t.view(torch.int16).copy_(x.clamp(0, 4095).round().to(torch.int16))
    else:
        ctx.synth_device(fmt, w["fs"], first_stream, S, batch_index * B * n, B * n, t.data_ptr())
    return t


def run_ours(args):
    import torch
    import torch.distributed as dist
    sdrb = importlib.import_module("sdr-for-android-lib_b200")   # raises if libsdr_b200.so is missing

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    w = workload(args)
    S, B, n, fs, fmt, stages = w["S"], w["B"], w["n"], w["fs"], w["fmt"], w["stages"]
    first_stream, _ = sdrb.shard_streams(S * world, world, rank)      # weak scaling: S streams on every rank
    samples_per_step = S * B * n
    stream = torch.cuda.Stream()
    ctx = sdrb.Context(local)
    ctx.set_cuda_stream(stream.cuda_stream)
    cfg = sdrb.Config(FC, fs, n, FOCUS_KHZ, inputFormat=fmt, **w["extra"])
    ctx.configure(cfg, S, B)
    has_fft = bool(stages & F_FFT)
    ctx.enable_kernel_timing(has_fft)

    # synthetic IQ generated on the device, resident in HBM before the timed region.  Two alternating
    # batches, each larger than L2, so successive steps never find their input in L2.
    n_batches = 2
    d_iq = [fill_input(sdrb, torch, ctx, w, S, B, first_stream, i) for i in range(n_batches)]
    ctx.sync()
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(i, st=stages):
        ctx.process(d_iq[i % n_batches], B, st, on_device=True)

    def timed(k, st=stages):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record()
            for i in range(k):
                step(i, st)
            e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for i in range(max(args.warmup, 3)):
        step(i)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.1)
    launches0 = ctx.launch_count
    t_a = sampler.mark()
    ms_total = timed(args.steps)
    launches = ctx.launch_count - launches0
    value = world * samples_per_step * args.steps / (ms_total * 1e-3) / 1e6
    fft_ms = [ctx.last_fft_kernel_ms()[0]] if has_fft else []
    fft_frames = ctx.last_fft_kernel_frames if has_fft else 0

    # sustained: the same step back to back for seconds, clocks sampled throughout
    sustained = None
    if args.sustain_seconds > 0:
        per = ms_total / args.steps * 1e-3
        k = max(args.steps, int(args.sustain_seconds / max(per, 1e-6)))
        t_s0 = sampler.mark()
        ms_s = timed(k)
        t_s1 = sampler.mark()
        sustained = {"steps": k, "seconds": ms_s * 1e-3, "value": world * samples_per_step * k / (ms_s * 1e-3) / 1e6,
                     "unit": "MS/s", "clocks": sampler.summary(t_s0, t_s1) if rank == 0 else None}
    t_b = sampler.mark()
    if rank == 0:
        sampler.stop()
    clocks = sampler.summary(t_a, t_b) if rank == 0 else None

    # the dominant kernel, launch by launch (every launch bracketed by the library's events on the launching stream)
    roof = None
    peak, peak_src = measured_peak_gbs()
    step_bytes = algorithmic_bytes_per_sample(w) * samples_per_step
    step_achieved = step_bytes / (ms_total / args.steps * 1e-3) / 1e9
    if has_fft:
        for i in range(5):
            step(i)
            ctx.sync()
            fft_ms.append(ctx.last_fft_kernel_ms()[0])
        k_ms = float(np.median(fft_ms))
        k_name = {16384: f"k_fft16k2<{FMT_NAME[CF32 if fmt == RU16 else fmt].upper()}>", 65536: "k_fbA<16>+k_fbB<16>",
                  1048576: "k_fbA<32>+k_fbB<32>", 8192: "k_fft_8k", 32768: "k_fs_cols<3>+k_fft_rows4",
                  131072: "k_fs_cols<5>+k_fft_rows4", 262144: "k_fs_cols<5>+k_fft_8k"}.get(
                      n, "k_fft_small" if 256 <= n <= 4096 and n & (n - 1) == 0 else "k_stockham passes")
        k_bytes = (FMT_BYTES[CF32 if fmt == RU16 else fmt] + 4) * fft_frames * n
        k_ach = k_bytes / (k_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": k_ach, "peak": peak, "unit": "GB/s", "frac": k_ach / peak,
                "traffic": ncu_traffic(k_name), "kernel": k_name, "kernel_ms": k_ms, "readings_per_launch": fft_frames,
                "algorithmic_bytes_per_launch": k_bytes, "peak_source": peak_src,
                "note": "timed live inside the step; with SSB in the step the per-stream SSB state pass runs on a second "
                        "stream under this kernel"}
    else:
        roof = {"bound": "hbm", "achieved": step_achieved, "peak": peak, "unit": "GB/s", "frac": step_achieved / peak,
                "traffic": None, "kernel": "whole SSB step (no FFT stage in this configuration)",
                "algorithmic_bytes_per_launch": step_bytes, "peak_source": peak_src}
    roof["step"] = {"algorithmic_bytes_per_sample": algorithmic_bytes_per_sample(w), "algorithmic_bytes": step_bytes,
                    "achieved": step_achieved, "frac": step_achieved / peak, "traffic": ncu_traffic("step:" + w["name"])}

    # the FFT-only step of the same batch (what round 1 reported as `value`), for context
    fft_only = None
    if has_fft and (stages & F_SSB):
        for i in range(3):
            step(i, F_FFT)
        barrier()
        ms_f = timed(args.steps, F_FFT)
        kf = []
        for i in range(3):
            step(i, F_FFT)
            ctx.sync()
            kf.append(ctx.last_fft_kernel_ms()[0])
        kf_ms = float(np.median(kf))
        fft_only = {"value": world * samples_per_step * args.steps / (ms_f * 1e-3) / 1e6, "unit": "MS/s",
                    "ms_per_step": ms_f / args.steps, "kernel_ms": kf_ms,
                    "kernel_frac": (FMT_BYTES[fmt] + 4) * ctx.last_fft_kernel_frames * n / (kf_ms * 1e-3) / 1e9 / peak}

    # e2e: host buffers through the C-ABI; H2D of the IQ and D2H of records + spectra + PCM in the region
    e2e = None
    peak_rec = None
    if not args.no_e2e:
        Se = max(1, min(S, args.e2e_frames // B))
        ch = max(1, int(w["extra"].get("ssbChannels", 1)))
        m = pcm_per_block(n, fs) if stages & F_SSB else 0
        ectx = sdrb.Context(local)
        ectx.configure(cfg, Se, B)
        bi = FMT_BYTES[fmt]
        h_iq = torch.empty((Se * B * n * bi,), dtype=torch.uint8).pin_memory()
        h_iq.copy_(d_iq[0][: Se * B * n * bi].cpu())
        h_pow = torch.empty((Se, B, n), dtype=torch.float32).pin_memory() if has_fft else None
        h_rec_t = torch.empty((Se * B * sdrb.RESULT_DTYPE.itemsize,), dtype=torch.uint8).pin_memory() if has_fft else None
        h_pcm = torch.empty((max(1, Se * ch * B * m),), dtype=torch.int16).pin_memory() if m else None

        h_pow16 = torch.empty((Se, B, n), dtype=torch.float16).pin_memory() if has_fft else None

        def e2e_step(power=True, half=False):
            # one C-ABI call: H2D of the IQ, all kernels, D2H of every payload, chunk-pipelined over CUDA streams
            if half:
                ectx.process_host(h_iq, B, stages, results=h_rec_t, power=h_pow16, pcm=h_pcm, power_format=sdrb.POWER_DB_F16)
            else:
                ectx.process_host(h_iq, B, stages, results=h_rec_t, power=h_pow if power else None, pcm=h_pcm)

        def e2e_time(**kw):
            e2e_step(**kw)
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.e2e_steps):
                e2e_step(**kw)
            barrier()
            t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return world * Se * B * n * args.e2e_steps / float(t.item()) / 1e6

        e2e_val = e2e_time()
        e2e_small = e2e_time(power=False)          # records + PCM only, for context
        e2e_half = e2e_time(half=True) if has_fft else None   # opt-in payload: half-precision dB view of the spectra
        d2h = (Se * B * n * 4 + Se * B * sdrb.RESULT_DTYPE.itemsize if has_fft else 0) + Se * ch * B * m * 2
        e2e = {"value": e2e_val, "unit": "MS/s", "h2d_bytes_per_step": Se * B * n * bi, "d2h_bytes_per_step": d2h,
               "streams_per_call": Se, "without_spectra_value": e2e_small, "db_f16_view_value": e2e_half,
               "d2h_gbs_per_gpu": d2h * 1e-9 / (Se * B * n / (e2e_val / world * 1e6)),
               "host_ceiling": "profiles/r02_host_path.json: one GPU alone moves 48-50 GB/s each way with both directions "
                               "busy; eight GPUs together 94 GB/s out (11.8 GB/s per GPU)",
               "note": "one sdrb_process_host call per step: pinned host IQ in; records, float32 linear spectra and PCM out; "
                       "chunk-pipelined over 3 CUDA streams (H2D, kernels, D2H).  db_f16_view_value: the same call with "
                       "the opt-in half-precision dB view instead of the float32 spectra (sdrb_process_host_ex)"}
        if has_fft:
            h_rec = h_rec_t.numpy().view(sdrb.RESULT_DTYPE).reshape(Se, B)
            peak_rec = sdrb.strongest_peak(h_rec, first_stream, dist if world > 1 else None)
        ectx.close()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_single_core(w, args.cpu_seconds)
    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "MS/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(w, world), "roofline": roof, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": int(launches), "clocks": clocks, "sustained": sustained, "fft_only": fft_only}
        if peak_rec:
            out["strongest_peak"] = {"absPeakDb": peak_rec[0], "stream": peak_rec[1], "block": peak_rec[2],
                                     "bin": peak_rec[3]}
        print(json.dumps(out))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_ours(a)
