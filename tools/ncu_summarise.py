"""Summarise the ncu launch list of tools/ncu_job.py (gpurun_out/launches_c3.csv) into per-kernel shares of one
prefill step and one decode step.  Steps are delimited by embed_kernel launches (the first kernel of every step).
  python tools/ncu_summarise.py gpurun_out/launches_c3.csv"""
import collections
import csv
import re
import sys


def short(name):
    m = re.match(r"(?:void )?(?:\w+::)*(\w+)(<[^>]*>)?", name)      # any namespace prefix (kge::b16::atc::...)
    base = m.group(1)
    t = m.group(2) or ""
    t = t.replace("(bool)", "").replace("(int)", "").replace(" ", "")
    return base + t


def main(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = [(short(r["Kernel Name"]), r["Grid Size"], float(r["Metric Value"]) / 1e3) for r in csv.DictReader(lines)]
    steps, cur = [], None
    for k, g, us in rows:
        if k.startswith("embed_kernel") and (cur is None or len(cur) > 8):
            cur = []
            steps.append(cur)
        if cur is not None:
            cur.append((k, g, us))
    prefill = [s for s in steps if any(k.startswith("attn_tc") for k, _, _ in s)]
    decode = [s for s in steps if any(k.startswith("attn_decode") for k, _, _ in s) and not any(k.startswith("attn_tc") for k, _, _ in s)]
    print(f"{len(rows)} launches, {len(steps)} steps: {len(prefill)} prefill, {len(decode)} decode")
    for title, group in (("prefill step (4 x 2048 tokens), mean over steps", prefill), ("decode step (64 requests, ctx 2049..), mean over steps", decode)):
        if not group:
            continue
        agg = collections.OrderedDict()
        for s in group:
            for k, g, us in s:
                a = agg.setdefault(k, [0, 0.0])
                a[0] += 1
                a[1] += us
        n = len(group)
        tot = sum(v[1] for v in agg.values()) / n
        print(f"\n### {title}: {sum(v[0] for v in agg.values()) // n} launches, {tot / 1e3:.2f} ms of kernel time\n")
        print("| kernel | launches/step | µs/launch | µs/step | share |")
        print("|---|---|---|---|---|")
        for k, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            print(f"| {k} | {c / n:.0f} | {us / c:.1f} | {us / n:.0f} | {100 * us / n / tot:.1f} % |")


if __name__ == "__main__":
    main(sys.argv[1])
