"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list into kernel time shares (markdown)."""
import collections
import csv
import re
import sys


def main(path, per_step_launch_groups=4):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = r["Kernel Name"]
        val = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ns = val * {"ns": 1, "us": 1e3, "ms": 1e6, "nsecond": 1, "usecond": 1e3, "msecond": 1e6}.get(unit, 1)
        rows.append((name, ns))
    agg = collections.OrderedDict()
    for name, ns in rows:
        short = re.sub(r"\(.*$", "", name)
        short = re.sub(r"^void ", "", short)
        if not short.startswith("jm::"):
            short = "(torch) " + short[:60]
        t, n = agg.get(short, (0.0, 0))
        agg[short] = (t + ns, n + 1)
    jm_total = sum(t for k, (t, n) in agg.items() if k.startswith("jm::"))
    print(f"jm:: kernels: {jm_total / 1e6:.1f} ms over {sum(n for k, (t, n) in agg.items() if k.startswith('jm::'))} launches")
    print()
    print("| kernel | time (ms) | share | launches | avg (us) |")
    print("|---|---|---|---|---|")
    for k, (t, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        if not k.startswith("jm::"):
            continue
        print(f"| `{k}` | {t / 1e6:.1f} | {100 * t / jm_total:.1f}% | {n} | {t / n / 1e3:.1f} |")
    groups = collections.Counter()
    for k, (t, n) in agg.items():
        if not k.startswith("jm::"):
            continue
        g = ("GEMM" if "gemm" in k else "attention" if "attn" in k else "LayerNorm" if ("layernorm" in k or "ln_param" in k)
             else "other")
        groups[g] += t
    print()
    for g, t in groups.most_common():
        print(f"- {g}: {t / 1e6:.1f} ms ({100 * t / jm_total:.1f}%)")


if __name__ == "__main__":
    main(sys.argv[1])
