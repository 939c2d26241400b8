"""Summarise an ncu CSV launch list of one train step (gpu__time_duration + DRAM bytes per launch).

    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
        -s 2500 -c 3000 --csv --log-file gpurun_out/step_launches.csv \
        python bench.py --steps 1 --warmup 3 --no-graph --no-prof --no-cpu-baseline
    (any window of >= 2 steps after the first: the step's launch count is found as the period of the kernel-name sequence)
    python tools/ncu_step_summary.py gpurun_out/step_launches.csv <batch_per_gpu> profiles/r1_step_ncu_summary

Writes <out>.txt (per-kernel table: share of the step, launches, average time, DRAM bytes per launch) and <out>.json
(what bench.py reads for roofline.traffic).  Per-launch times under ncu are cold-cache and serialised: compare SHARES.
"""
import csv
import json
import re
import sys
from collections import defaultdict


def short(name):
    name = re.sub(r"\(.*", "", name)
    name = name.replace("void ", "").replace("<unnamed>::", "")
    return name.strip()


def main():
    src, batch, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    rows = []
    with open(src, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    rd = csv.DictReader(lines)
    per = defaultdict(dict)   # launch id -> {metric: value}
    names = {}
    for r in rd:
        i = int(r["ID"])
        names[i] = r["Kernel Name"]
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "")
        m = r["Metric Name"]
        if m == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)
        else:
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
        per[i][m] = v
    # keep ONE step: the capture may span several; the launch sequence is periodic with the step's launch count
    ids = sorted(per)
    seq = [short(names[i]) for i in ids]
    period = None
    for L in range(200, len(seq) // 2 + 1):
        if all(seq[j] == seq[j + L] for j in range(len(seq) - L)):
            period = L
            break
    if period is not None:
        ids = ids[-period:]
        per = {i: per[i] for i in ids}
    agg = defaultdict(lambda: [0, 0.0, 0.0])
    for i, d in per.items():
        k = short(names[i])
        a = agg[k]
        a[0] += 1
        a[1] += d.get("gpu__time_duration.sum", 0.0)
        a[2] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    total_us = sum(a[1] for a in agg.values())
    total_n = sum(a[0] for a in agg.values())
    with open(out + ".txt", "w") as f:
        f.write("# %s  (batch %d / GPU, %d launches, %.1f us summed; ncu per-launch times are cold-cache and serialised)\n"
                % (src, batch, total_n, total_us))
        f.write("%10s %6s %6s %10s %14s  kernel\n" % ("total_us", "share", "n", "avg_us", "dram_B/launch"))
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("%10.1f %5.1f%% %6d %10.1f %14.0f  %s\n" % (a[1], 100 * a[1] / total_us, a[0], a[1] / a[0], a[2] / a[0], k))
    g = [a for k, a in agg.items() if k.startswith("gemm_bf16_tcgen05")]
    gn, gus, gb = sum(a[0] for a in g), sum(a[1] for a in g), sum(a[2] for a in g)
    js = {"source": src, "batch_per_gpu": batch, "launches_per_step": total_n, "period_detected": period is not None, "sum_us": round(total_us, 1),
          "gemm": {"launches": gn, "share_of_step": round(gus / total_us, 4), "avg_us": round(gus / max(gn, 1), 2),
                   "dram_bytes_per_launch": round(gb / max(gn, 1), 1)}}
    json.dump(js, open(out + ".json", "w"), indent=1)
    print(json.dumps(js))


if __name__ == "__main__":
    main()
