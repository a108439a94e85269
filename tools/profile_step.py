"""tools/profile_step.py — the bare step of one bench.py workload, for ncu launch lists and captures.

    python tools/profile_step.py [--config default] [--streams S] [--blocks B] [--steps 3] [--stages fft|ssb|both]

Prints the wall-clock ms per step (informational only: numbers under a profiler are never bench values)."""
import argparse
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="default")
    ap.add_argument("--streams", type=int, default=0)
    ap.add_argument("--blocks", type=int, default=0)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--stages", default="")
    a = ap.parse_args()
    import torch
    sdrb = importlib.import_module("sdr-for-android-lib_b200")
    w = dict(bench.WORKLOADS[a.config], name=a.config)
    if a.streams:
        w["S"] = a.streams
    if a.blocks:
        w["B"] = a.blocks
    stages = {"": w["stages"], "fft": 1, "ssb": 2, "both": 3}[a.stages]
    ctx = sdrb.Context(0)
    ctx.configure(sdrb.Config(bench.FC, w["fs"], w["n"], bench.FOCUS_KHZ, inputFormat=w["fmt"], **w["extra"]), w["S"], w["B"])
    d = [bench.fill_input(sdrb, torch, ctx, w, w["S"], w["B"], 0, i) for i in range(2)]
    ctx.sync()
    torch.cuda.synchronize()
    for i in range(2):
        ctx.process(d[i % 2], w["B"], stages, on_device=True)
    ctx.sync()
    t0 = time.perf_counter()
    for i in range(a.steps):
        ctx.process(d[i % 2], w["B"], stages, on_device=True)
    ctx.sync()
    dt = (time.perf_counter() - t0) / a.steps
    print(f"{a.config} S={w['S']} B={w['B']} stages={stages}: {dt * 1e3:.3f} ms per step, "
          f"{w['S'] * w['B'] * w['n'] / dt / 1e9:.1f} GS/s (wall clock)")
    ctx.close()


if __name__ == "__main__":
    main()
