"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum --csv`) or a `--page raw --csv`
export into a per-kernel markdown table.  usage:
    python tools/summarize_ncu.py launches gpurun_out/launches_bf16.csv > profiles/r1_launches_bf16.md
    python tools/summarize_ncu.py raw gpurun_out/r1_sections_raw.csv > profiles/r1_sections.md
"""
import collections
import csv
import re
import sys


def short(name):
    name = re.sub(r"\(.*", "", name)
    return name.replace("void ", "").replace("stts::<unnamed>::", "stts::")[:80]


def launches(path):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr, data = rows[hi], rows[hi + 1:]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = nat = 0.0
    n_nat = 0
    for r in data:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1e-3)
        k = short(r[ki])
        agg[k][0] += 1
        agg[k][1] += v
        tot += v
        if "stts::" in r[ki]:
            nat += v
            n_nat += 1
    print(f"total {tot / 1e3:.2f} ms over {sum(a[0] for a in agg.values())} launches; "
          f"native (stts::) {nat / 1e3:.2f} ms over {n_nat} launches; "
          f"torch token-rate ops {(tot - nat) / 1e3:.2f} ms\n")
    print("| kernel | launches | total us | share |")
    print("|---|---|---|---|")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
        print(f"| `{k}` | {n} | {t:.1f} | {100 * t / tot:.1f} % |")


def raw(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    H = {h: i for i, h in enumerate(hdr)}

    def g(r, k):
        try:
            return float(r[H[k]].replace(",", ""))
        except Exception:
            return float("nan")

    cols = [("gpu__time_duration.sum", "ms"),
            ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor %"),
            ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %"),
            ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %"),
            ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
            ("dram__bytes.sum.per_second", "DRAM TB/s")]
    print("| # | kernel | grid | " + " | ".join(c[1] for c in cols) + " |")
    print("|---|---|---|" + "---|" * len(cols))
    seen = set()
    for i, r in enumerate(data):
        key = (short(r[H["Kernel Name"]]), r[H["Grid Size"]])
        if key in seen:
            continue
        seen.add(key)
        print(f"| {i} | `{key[0]}` | {key[1]} | " + " | ".join(f"{g(r, c[0]):.3f}" for c in cols) + " |")


if __name__ == "__main__":
    {"launches": launches, "raw": raw}[sys.argv[1]](sys.argv[2])
