#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals and the last search step.

usage: python tools/ncu_launch_summary.py profiles/r1_ncu_launches_c2.csv > profiles/r1_ncu_launches_c2_summary.txt
"""
import csv
import io
import sys
from collections import OrderedDict


def short(name: str) -> str:
    return name.split("(")[0][:70]


def main(path: str) -> None:
    lines = [l for l in open(path, errors="replace") if l.startswith('"')]
    rows = list(csv.DictReader(io.StringIO("".join(lines))))
    rows = [r for r in rows if r["Metric Name"] == "gpu__time_duration.sum"]
    unit = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
    launches = [(int(r["ID"]), short(r["Kernel Name"]), float(r["Metric Value"].replace(",", "")) * unit.get(r["Metric Unit"], 1e-6))
                for r in rows]
    tot = OrderedDict()
    for _, n, ms in launches:
        c, t = tot.get(n, (0, 0.0))
        tot[n] = (c + 1, t + ms)
    total = sum(t for _, t in tot.values())
    print("ncu --metrics gpu__time_duration.sum --clock-control none -c 400, python bench.py --steps 2 --warmup 3 --no-cpu-baseline")
    print("(whole process: index build, warm-up, timed steps, e2e steps, validity check; per-launch times are cold-cache and "
          "serialised: compare SHARES)")
    print("kernel | launches | total ms | share")
    for n, (c, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print(f"{n} | {c} | {t:.3f} | {100 * t / total:.2f}%")
    # the last search step: from the last merge_arena launch back to the convert launches that open it
    last = max((i for i, (_, n, _) in enumerate(launches) if "merge_arena" in n), default=None)
    if last is None:
        return
    first = last
    while first > 0 and ("nxf::" in launches[first - 1][1]) and "merge_arena" not in launches[first - 1][1]:
        first -= 1
    print("\none search step (the last one in the list), launch by launch:")
    step = launches[first:last + 1]
    for i, n, ms in step:
        print(f"  {i} {n} {ms:.4f} ms")
    st = sum(ms for _, _, ms in step)
    k1 = [ms for _, n, ms in step if "k1_search" in n]
    if k1:
        print(f"  step total {st:.3f} ms; k1_search_kernel ({len(k1)} launches: estimate, sample 1/256, sample 1/16, main, "
              f"fall-back no-op) {sum(k1):.3f} ms = {100 * sum(k1) / st:.1f} % of the step")


if __name__ == "__main__":
    main(sys.argv[1])
