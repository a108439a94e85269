#!/usr/bin/env python
"""bench.py — IQ mega-samples/s through the batched hot path on N B200s (one process per GPU).

Workload (BASELINE.json configs[1], the configuration the metric is quoted on): 1024 independent
RTL-SDR u8 streams per GPU, 16384-point FFT + |X|^2/fftshift + peak / signal-strength estimator.
One step = one pass over one batch of `--blocks` readings per stream; streams are sharded by rank
with no data-path collective ("scaling": "weak": per-GPU work is fixed as N grows).

    python bench.py [--gpus N --steps K --warmup W]            # our arm
    python bench.py --impl reference [...]                     # the reference's CPU path, host cores

Prints ONE JSON line (rank 0).  `value` is device-timed with the inputs resident in HBM; `e2e` is
the same metric through the C-ABI with HOST buffers (H2D of the IQ and D2H of records + spectra
inside the timed region); `roofline` is the 16K FFT kernel against the measured HBM copy peak;
`cpu_baseline` is the reference's own code (oracle/_ref) on this box's host cores.
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

N_FFT = 16384
FS = 2_000_000
FC = 430_000_000
FOCUS_KHZ = 5
FMT_CU8 = 0
BYTES_IN = 2          # u8 I + u8 Q
ALGO_BYTES = BYTES_IN + 4   # one read of the raw sample + one f32 power bin
METRIC = "IQ MS/s through FFT+peak+SSB at 1/2/4/8 B200, % HBM roofline, vs host CPU"
WORKLOAD = ("1024 independent RTL-SDR u8 streams per GPU, 16384-pt FFT + power/fftshift + peak + "
            "signal strength (BASELINE configs[1])")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--streams", type=int, default=1024, help="streams per GPU")
    ap.add_argument("--blocks", type=int, default=37,
                    help="readings per stream per step (1024 x 37 = 37888 transforms = 256 whole rounds of 148 SMs)")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU arm
# The only place besides tests/ and smoke() that executes oracle/: as the thing TIMED for the CPU
# baseline (never on our arm's product path).
_W = {}


def _cpu_worker_init(tests_dir, root, first_stream):
    sys.path.insert(0, tests_dir)
    sys.path.insert(0, root)
    import oracle_lib as ol
    sdrb = importlib.import_module("sdr-for-android-lib_b200")
    nbuf = 8
    raw = sdrb.synth_host(FMT_CU8, FS, first_stream, 1, 0, nbuf * N_FFT)[0]
    _W["iq"] = np.ascontiguousarray(ol.to_cf32(raw, FMT_CU8))
    _W["nbuf"] = nbuf
    _W["ol"] = ol
    if ol.have_ref():
        r = ol.ref()
        _W["h"] = r.sdr_ref_fft_create()
        r.sdr_ref_fft_configure(_W["h"], FC, FS, N_FFT, FOCUS_KHZ)
        _W["kind"] = "reference"
    else:
        _W["of"] = ol.OracleFft(FC, FS, N_FFT, FOCUS_KHZ)
        _W["kind"] = "port"
    return _W["kind"]


def _cpu_worker_run(frames):
    """FFTProcessor::process on `frames` readings (src/sdr-bridge-java-soapy.cpp:444); returns seconds."""
    ol = _W["ol"]
    t0 = time.perf_counter()
    if _W["kind"] == "reference":
        ol.ref().sdr_ref_run_frames(_W["h"], ol._fp(_W["iq"]), N_FFT, _W["nbuf"], frames, FS, 1, 0, 1)
    else:
        iq = _W["iq"].reshape(_W["nbuf"], N_FFT, 2)
        for f in range(frames):
            _W["of"].process(iq[f % _W["nbuf"]])
    return time.perf_counter() - t0


def cpu_single_core(seconds):
    kind = _cpu_worker_init(os.path.join(ROOT, "tests"), ROOT, 0)
    t = _cpu_worker_run(20)
    frames = max(20, int(seconds / max(t / 20, 1e-6)))
    t = _cpu_worker_run(frames)
    return {"value": frames * N_FFT / t / 1e6, "unit": "MS/s", "cores": 1, "kind": kind,
            "sample": f"{frames} readings of {N_FFT} u8-derived cf32 samples through FFTProcessor::process "
                      f"(FFTW plan+execute, power, fftshift, estimator) on one host core, {t:.1f} s"}


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation on all host cores, multi-process because
    the FFTW planner is not thread-safe and process() plans on every call."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    cores = min(len(os.sched_getaffinity(0)) or 1, 128)
    ctx = mp.get_context("spawn")
    pools = [ctx.Pool(1, initializer=_cpu_worker_init, initargs=(os.path.join(ROOT, "tests"), ROOT, i))
             for i in range(cores)]
    probe = [p.apply_async(_cpu_worker_run, (20,)) for p in pools]
    per_frame = max(r.get() for r in probe) / 20
    frames = max(10, int(1.5 / per_frame))          # each step ~1.5 s of wall clock
    kind = "reference" if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libsdr_ref.so")) else "port"

    def step():
        rs = [p.apply_async(_cpu_worker_run, (frames,)) for p in pools]
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
    samples = args.steps * cores * frames * N_FFT
    val = samples / dt / 1e6
    sample = (f"{cores} processes x {frames} readings of {N_FFT} samples per step through "
              f"FFTProcessor::process ({kind}), one stream per process")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "MS/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "fft": N_FFT, "sample_rate": FS, "format": "cu8->cf32 on host",
                   "bounded_sample": sample},
        "cpu_baseline": {"value": val, "unit": "MS/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": "MS/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ----------------------------------------------------------------------------- our arm
FFT_KERNEL = {"3": "k_fft16k3<CU8>", "4": "k_fft16k4<CU8>", "5": "k_fft16k5<CU8>"}.get(
    os.environ.get("SDRB_FFT16K", "2")[:1], "k_fft16k2<CU8>")


def ncu_traffic_bytes():
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel, from the committed
    `ncu --set full` capture of this same command (profiles/r01_traffic.json); None if not captured."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            return json.load(f).get(FFT_KERNEL)
    except (OSError, ValueError):
        return None


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

    S, B, n = args.streams, args.blocks, N_FFT
    first_stream, _ = sdrb.shard_streams(S * world, world, rank)      # weak scaling: S streams on every rank
    samples_per_step = S * B * n
    stream = torch.cuda.Stream()
    ctx = sdrb.Context(local)
    ctx.set_cuda_stream(stream.cuda_stream)
    cfg = sdrb.Config(FC, FS, n, FOCUS_KHZ, inputFormat=sdrb.CU8)
    ctx.configure(cfg, S, B)
    ctx.enable_kernel_timing(True)

    # synthetic IQ generated on the device, resident in HBM before the timed region.  Two batches
    # (2 x 256 MiB in, 512 MiB of spectra out per step) so that successive steps never hit in L2.
    n_batches = 2
    d_iq = [torch.empty((S, B, n, 2), dtype=torch.uint8, device="cuda") for _ in range(n_batches)]
    for i, t in enumerate(d_iq):
        ctx.synth_device(sdrb.CU8, FS, first_stream, S, i * B * n, B * n, t.data_ptr())
    ctx.sync()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(i):
        ctx.process(d_iq[i % n_batches], B, sdrb.STAGE_FFT, on_device=True)

    for i in range(args.warmup):
        step(i)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fft_ms = []
    with torch.cuda.stream(stream):
        e0.record()
        for i in range(args.steps):
            step(i)
        e1.record()
    # per-launch duration of the dominant kernel: CUDA events recorded by the library around the
    # FFT launch on the launching stream; read for the last step after the region is closed
    barrier()
    fft_ms.append(ctx.last_fft_kernel_ms()[0])
    fft_frames = ctx.last_fft_kernel_frames      # readings covered by one timed launch of the FFT kernel
    ms_total = e0.elapsed_time(e1)
    launches = ctx.launch_count - launches0
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * samples_per_step * args.steps / (ms_total * 1e-3) / 1e6

    # FFT kernel alone, timed launch by launch (every launch bracketed by the library's events)
    for i in range(5):
        step(i)
        ctx.sync()
        fft_ms.append(ctx.last_fft_kernel_ms()[0])
    k_ms = float(np.median(fft_ms))
    peak, peak_src = measured_peak_gbs()
    achieved = ALGO_BYTES * fft_frames * n / (k_ms * 1e-3) / 1e9

    # e2e: host buffers through the C-ABI; H2D of the IQ and D2H of records + spectra in the region
    h_iq = torch.empty((S, B, n, 2), dtype=torch.uint8).pin_memory()
    h_iq.copy_(d_iq[0].cpu())
    h_pow = torch.empty((S, B, n), dtype=torch.float32).pin_memory()
    h_rec_t = torch.empty((S * B * sdrb.RESULT_DTYPE.itemsize,), dtype=torch.uint8).pin_memory()
    h_rec = h_rec_t.numpy().view(sdrb.RESULT_DTYPE).reshape(S, B)

    def e2e_step():
        # one C-ABI call: H2D of the IQ, all kernels, D2H of records + spectra, chunk-pipelined
        ctx.process_host(h_iq, B, sdrb.STAGE_FFT, results=h_rec_t, power=h_pow)

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_val = world * samples_per_step * args.e2e_steps / float(t.item()) / 1e6
    # the same without hauling the spectra back (records only), for context
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        ctx.process_host(h_iq, B, sdrb.STAGE_FFT, results=h_rec_t)
    torch.cuda.synchronize()
    e2e_rec_val = samples_per_step * args.e2e_steps / (time.perf_counter() - t0) / 1e6
    peak_rec = sdrb.strongest_peak(h_rec, first_stream, dist if world > 1 else None)

    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        cpu = cpu_single_core(args.cpu_seconds)
    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "MS/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "streams_per_gpu": S, "readings_per_stream_per_step": B,
                       "fft": n, "format": "cu8", "sample_rate": FS, "focus_khz": FOCUS_KHZ,
                       "stages": "fft+power+fftshift+estimator+tracker", "parallelism": f"streams sharded x{world}",
                       "l2": f"{n_batches} alternating batches of {S * B * n * BYTES_IN >> 20} MiB in / "
                             f"{S * B * n * 4 >> 20} MiB out per step (> 126 MB L2)"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": ncu_traffic_bytes(), "kernel": FFT_KERNEL,
                         "kernel_ms": k_ms, "readings_per_launch": fft_frames,
                         "algorithmic_bytes_per_launch": ALGO_BYTES * fft_frames * n,
                         "peak_source": peak_src},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_val, "unit": "MS/s", "h2d_bytes_per_step": S * B * n * BYTES_IN,
                    "d2h_bytes_per_step": S * B * n * 4 + h_rec.nbytes,
                    "records_only_value": e2e_rec_val,
                    "note": "one sdrb_process_host call per step: pinned host IQ in, records + spectra out, chunk-pipelined over 3 CUDA streams"},
            "gpu_launches": int(launches), "clocks": clocks,
            "strongest_peak": {"absPeakDb": peak_rec[0], "stream": peak_rec[1], "block": peak_rec[2],
                               "bin": peak_rec[3]}}))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_ours(a)
