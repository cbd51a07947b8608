"""Text summary of one `ncu --set full --import-source on` report (key raw metrics, stall-reason totals, hottest SASS lines).

    python tools/ncu_full_summary.py gpurun_out/pair_k11.ncu-rep "description" >> profiles/r01_ncu_full_resblock_pair.txt
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size"]


def run(args):
    return subprocess.run(["ncu", "-i", sys.argv[1]] + args, capture_output=True, text=True).stdout


raw = list(csv.reader(io.StringIO(run(["--page", "raw", "--csv"]))))
hdr, units, vals = raw[0], raw[1], raw[2]
col = {n: i for i, n in enumerate(hdr)}
print("=== %s" % vals[col["Kernel Name"]][:110])
print("    " + (sys.argv[2] if len(sys.argv) > 2 else ""))
for k in KEYS:
    if k in col:
        print("    %-70s %18s %s" % (k, vals[col[k]], units[col[k]]))
src = list(csv.reader(io.StringIO(run(["--page", "source", "--csv"]))))
h = next(i for i, r in enumerate(src) if r and r[0] == "Address")
cols = {n: i for i, n in enumerate(src[h])}
stalls = [c for c in src[h] if c.startswith("stall_") and "Not Issued" not in c]
tot = collections.Counter()
lines = []
for r in src[h + 1:]:
    if len(r) < len(cols) or not r[cols["# Samples"]].isdigit():   # a report with several launches repeats the header row
        continue
    n = int(r[cols["# Samples"]])
    lines.append((n, r[cols["Source"]].strip()))
    for s in stalls:
        tot[s[6:]] += int(r[cols[s]])
print("    warp-stall samples: %d; by reason: %s" % (sum(n for n, _ in lines), ", ".join("%s %d" % kv for kv in tot.most_common(9))))
print("    hottest SASS lines:")
for n, s in sorted(lines, reverse=True)[:10]:
    print("      %6d  %s" % (n, s[:90]))
print()
