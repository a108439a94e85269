"""tools/block_lengths.py — FFT stage at every power-of-two block length that has a kernel of its own, 268 M u8
samples per call: wall-clock ms of the whole FFT stage and the device-timed FFT kernel(s) (informational; the
clock-stamped lines are bench.py --config n1024 ... n32768).

    python tools/block_lengths.py"""
import importlib
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
sdrb = importlib.import_module("sdr-for-android-lib_b200")
for n in (256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 131072, 262144):
    S = (1 << 28) // n     # 268 M samples per call
    ctx = sdrb.Context(0)
    ctx.configure(sdrb.Config(430_000_000, 2_000_000, n, 5, inputFormat=0), S, 1)
    ctx.enable_kernel_timing(True)
    d = torch.empty((S, 1, n, 2), dtype=torch.uint8, device="cuda")
    ctx.synth_device(0, 2_000_000, 0, S, 0, n, d.data_ptr())
    for _ in range(2): ctx.process(d, 1, sdrb.STAGE_FFT, on_device=True)
    ctx.sync(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5): ctx.process(d, 1, sdrb.STAGE_FFT, on_device=True)
    ctx.sync(); dt = (time.perf_counter() - t0) / 5
    km = ctx.last_fft_kernel_ms()[0]
    print(f"n={n:7d} S={S:6d}: {dt*1e3:8.3f} ms  {S*n/dt/1e9:7.1f} GS/s   fft kernels {km:7.3f} ms = {S*n/km/1e6:7.1f} GS/s, {S*n*6/km/1e6/6537:5.2f} of HBM", flush=True)
    ctx.close(); del d
