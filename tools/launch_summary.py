"""Summarise an ncu launch list (CSV of gpu__time_duration.sum, dram__bytes_read/write.sum and the tensor-pipe metric, one
row per launch and metric) into the table kept under profiles/ and the DRAM-traffic JSON bench.py reads.

    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,\
sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -c 800 --csv \
        --log-file gpurun_out/launches.csv python tools/ncu_target.py
    python tools/launch_summary.py gpurun_out/launches.csv profiles/r01_launch_list_v5 "<one-line description>"
"""
import collections
import csv
import json
import re
import sys

TENSOR = "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"
GEMM_FAMILY = ("conv_gemm_umma2_kernel", "resblock_pair_kernel", "ff_fused_kernel", "resqkv_fused_kernel")   # attention_umma_kernel is reported on its own line


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"^mb::", "", name)
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("(int)", "")
    if name.startswith("at::native::"):
        name = name.split("<")[0].split("::")[-1]
    return name


def main():
    src, out, desc = sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ""
    rows = list(csv.reader(open(src)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    cols = rows[h]
    launches = collections.OrderedDict()
    for r in rows[h + 1:]:
        if len(r) < len(cols):
            continue
        rec = dict(zip(cols, r))
        d = launches.setdefault(int(rec["ID"]), {"name": short(rec["Kernel Name"])})
        d[rec["Metric Name"]] = float(rec["Metric Value"].replace(",", ""))
    agg = collections.OrderedDict()
    for d in launches.values():
        a = agg.setdefault(d["name"], {"n": 0, "ns": 0.0, "dram": 0.0, "tens_w": 0.0})
        ns = d.get("gpu__time_duration.sum", 0.0)
        a["n"] += 1
        a["ns"] += ns
        a["dram"] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
        a["tens_w"] += d.get(TENSOR, 0.0) * ns
    total = sum(a["ns"] for a in agg.values())
    lines = [desc, "Per-launch times are serialised and cold-cache; the SHARE per kernel is what agrees with bench.py's event-timed breakdown.", "",
             "%-58s %8s %9s %6s %8s %15s %s" % ("kernel", "launches", "total ms", "share", "avg us", "dram MB/launch", "tensor-pipe active (time-weighted)")]
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1]["ns"]):
        lines.append("%-58s %8d %9.3f %5.1f%% %8.1f %15.2f %10.1f%%" % (name[:58], a["n"], a["ns"] / 1e6, 100 * a["ns"] / total, a["ns"] / a["n"] / 1e3,
                                                                     a["dram"] / a["n"] / 1e6, a["tens_w"] / max(a["ns"], 1)))
    lines.append("%-58s %8d %9.3f" % ("TOTAL", len(launches), total / 1e6))
    fam = [a for n, a in agg.items() if n.startswith(GEMM_FAMILY)]
    n = sum(a["n"] for a in fam)
    ns = sum(a["ns"] for a in fam)
    dram = sum(a["dram"] for a in fam)
    lines += ["", "tcgen05 conv-GEMM family (conv_gemm_umma2_kernel + resblock_pair_kernel + ff_fused_kernel + resqkv_fused_kernel, all instantiations): %d launches, %.3f ms = %.1f%% of "
              "the step, DRAM traffic %.2f MB per launch (%.2f GB per step)" % (n, ns / 1e6, 100 * ns / total, dram / n / 1e6, dram / 1e9)]
    open(out + "_summary.txt", "w").write("\n".join(lines) + "\n")
    json.dump({"source": out + "_summary.txt (ncu dram__bytes_read.sum + dram__bytes_write.sum, one eager cfg2 pass)",
               "kernel": "conv_gemm_umma2_kernel + resblock_pair_kernel + ff_fused_kernel + resqkv_fused_kernel (all instantiations)", "launches_per_step": n,
               "dram_bytes_per_launch": dram / n, "dram_bytes_per_step": dram, "share_of_step_time": ns / total},
              open(sys.argv[4] if len(sys.argv) > 4 else "profiles/r02_dram_traffic.json", "w"), indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
