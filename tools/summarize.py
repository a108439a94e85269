"""tools/summarize.py — one line per bench JSON, or per-kernel totals of an ncu launch list (csv)."""
import collections
import csv
import json
import sys


def bench(fn):
    try:
        d = json.loads(open(fn).read().strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        print(fn, "ERR", e, open(fn).read()[-300:])
        return
    r, c = d.get("roofline") or {}, d.get("config", {})
    fo = d.get("fft_only") or {}
    print(f"{fn}: {c.get('name')} S={c.get('streams_per_gpu')} B={c.get('readings_per_stream_per_step')} "
          f"value {d['value'] / 1e3:.1f} GS/s  {d['ms_per_step']:.3f} ms/step  kernel {r.get('kernel_ms', 0):.3f} ms "
          f"frac {r.get('frac', 0):.3f}  step frac {r.get('step', {}).get('frac', 0):.3f}  "
          f"fft_only {fo.get('value', 0) / 1e3:.1f}  launches {d.get('gpu_launches')}"
          + (f"  e2e {d['e2e']['value'] / 1e3:.2f}" if d.get("e2e") else ""))


def launches(fn, skip=0):
    rows = list(csv.reader(open(fn)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    H = rows[hdr]
    ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[hdr + 1:]:
        if len(r) <= vi or not r[vi].replace(",", "").replace(".", "").isdigit():
            continue
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] in ("ns", "nsecond") else (v * 1e3 if r[ui] in ("ms", "msecond") else v)
        agg.setdefault(r[ki][:70], []).append(v)
    tot = 0.0
    for k, v in agg.items():
        print(f"{k:70s} n={len(v):3d}  mean {sum(v) / len(v):9.1f} us  last {v[-1]:9.1f}")
        if "synth" not in k:
            tot += v[-1]
    print(f"sum of last launches (without synth): {tot:.1f} us")


if __name__ == "__main__":
    for fn in sys.argv[1:]:
        (launches if fn.endswith(".csv") else bench)(fn)
