"""Throughput of the other BASELINE.json configurations (C1, C3, C4, C5) on one B200, device-resident
synthetic IQ, one JSON line per configuration: IQ MS/s, algorithmic GB/s (bytes_in + 4 per
sample for the FFT path, + 2/decim for PCM) and the fraction of the measured HBM peak.  bench.py stays the
contract line (C2); this script is the measurement of the parity-test configurations at their full sizes.

    python bench_configs.py [--iters 10] [--only c3]
"""
import argparse
import importlib
import json
import os
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))


def peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except (OSError, KeyError, ValueError):
        return 6500.0, "fallback"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    import torch
    sdrb = importlib.import_module("sdr-for-android-lib_b200")
    if not torch.cuda.is_available():
        raise SystemExit("needs a CUDA device: the product has no CPU fallback")
    F, S_ = sdrb.STAGE_FFT, sdrb.STAGE_SSB
    # name, streams, blocks, N, fs, format, stages, extra config
    cases = [
        ("c1", "1 RTL-SDR stream, u8 2 MS/s, 16384-pt FFT + peak + strength + USB demod (64 readings per call)",
         1, 64, 16384, 2_000_000, sdrb.CU8, F | S_, {}),
        ("c1x1024", "the same per-stream work, 1024 streams x 8 readings", 1024, 8, 16384, 2_000_000, sdrb.CU8, F | S_, {}),
        ("c1x1024-b64", "the same, 64 readings per call (65536 AGC lanes instead of 8192)", 1024, 64, 16384, 2_000_000, sdrb.CU8,
         F | S_, {}),
        ("c3-fft", "256 cs16 streams, 65536-pt FFT + peak + strength only", 256, 8, 65536, 30_720_000, sdrb.CS16, F, {}),
        ("c3", "256 LimeSDR-style cs16 streams, 30.72 MS/s, 65536-pt FFT + true-LSB demod", 256, 8, 65536,
         30_720_000, sdrb.CS16, F | S_, {"upperSideband": 0, "ssbTrueSideband": 1}),
        ("c4", "wideband cf32 61.44 MS/s, 256 SSB channels per stream on a 25 kHz raster (SSB only)", 8, 32, 16384,
         61_440_000, sdrb.CF32, S_, {"ssbChannels": 256, "ssbChannelOffsetHz": -3.2e6, "ssbChannelSpacingHz": 25e3}),
        ("c5", "cf32 10 MS/s, 1048576-pt FFT + peak tracking", 16, 4, 1048576, 10_000_000, sdrb.CF32, F, {}),
        ("c5-ru16", "raw Airspy ADC words (SDRB_RU16) -> IQ on the GPU, 65536-pt FFT + peak tracking", 256, 8, 65536,
         10_000_000, sdrb.RU16, F, {}),
    ]
    peak, peak_src = peak_gbs()
    ctx = sdrb.Context(0)
    for name, what, S, B, n, fs, fmt, stages, extra in cases:
        if args.only and args.only != name:
            continue
        cfg = sdrb.Config(430_000_000, fs, n, 5, inputFormat=fmt, **extra)
        ctx.configure(cfg, S, B)
        bps = sdrb.format_bytes(fmt)
        d_iq = torch.empty((S * B * n * bps,), dtype=torch.uint8, device="cuda")
        if fmt == sdrb.RU16:
            # 12-bit ADC words: tone near fs/4 + noise (no device generator for this format)
            g = torch.Generator(device="cuda").manual_seed(1)
            i = torch.arange(S * B * n * 2, device="cuda", dtype=torch.float32)
            x = 2048 + 600 * torch.cos(2 * np.pi * 0.263 * i) + 40 * torch.randn(i.shape, device="cuda", generator=g)
            d_iq.view(torch.int16).copy_(x.clamp(0, 4095).round().to(torch.int16))
        else:
            ctx.synth_device(fmt, fs, 0, S, 0, B * n, d_iq.data_ptr())
        ctx.sync()
        for _ in range(3):
            ctx.process(d_iq, B, stages, on_device=True)
        ctx.sync()
        t0 = time.perf_counter()
        for _ in range(args.iters):
            ctx.process(d_iq, B, stages, on_device=True)
        ctx.sync()
        dt = (time.perf_counter() - t0) / args.iters
        samples = S * B * n
        m = ctx.pcm_per_block if stages & S_ else 0
        ch = max(1, int(extra.get("ssbChannels", 1)))
        algo = samples * bps + (samples * 4 if stages & F else 0) + S * ch * B * m * 2
        print(json.dumps({"config": name, "workload": what, "streams": S, "readings_per_call": B, "fft": n,
                          "format": {0: "cu8", 1: "cs16", 2: "cf32", 3: "ru16"}[fmt],
                          "stages": "+".join(k for k, b in (("fft", F), ("ssb", S_)) if stages & b),
                          "ms_per_call": dt * 1e3, "value": samples / dt / 1e6, "unit": "MS/s",
                          "pcm_per_block": m, "algorithmic_gbs": algo / dt / 1e9,
                          "frac_of_hbm_peak": algo / dt / 1e9 / peak, "peak_gbs": peak, "peak_source": peak_src,
                          "timing": "wall clock over %d calls bracketed by stream syncs, inputs resident in HBM" % args.iters}),
              flush=True)
        del d_iq
    ctx.close()


if __name__ == "__main__":
    main()
