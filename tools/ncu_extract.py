"""Turn an `ncu --set full` report of one tc2_fwd_kernel launch into the two text summaries kept under profiles/:
usage: ncu_extract.py report.ncu-rep out_prefix "description of the run"  (needs the ncu CLI; no GPU)."""
import csv
import subprocess
import sys

rep, prefix, desc = sys.argv[1], sys.argv[2], sys.argv[3]


def page(name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader([l for l in out.splitlines() if not l.startswith("==")]))


rows = page("raw")
hdr, units, vals = rows[0], rows[1], rows[2]
keys = ["gpu__time_duration.sum", "launch__grid_size", "launch__cluster_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max", "sm__cycles_active.avg",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed.sum"]
out = [desc, "kernel: " + vals[hdr.index("Kernel Name")], ""]
for k in keys:
    if k in hdr:
        i = hdr.index(k)
        out.append(f"{k:75s} {vals[i]:>18s} {units[i]}")
open(prefix + "_ncu_metrics.txt", "w").write("\n".join(out) + "\n")

rows = page("source")
hdr, data = rows[1], rows[2:]
isrc, ins = hdr.index("Source"), hdr.index("# Samples")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[ins] or 0) for r in data)
agg = {hdr[i]: sum(int(r[i] or 0) for r in data) for i in stall}
out = [desc, f"{tot} warp samples over {len(data)} SASS instructions", "", "stall reason totals (all warps; the 16 producer warps dominate the sample count):"]
for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:10]:
    out.append(f"  {k:24s} {v:8d}  {100 * v / max(tot, 1):5.1f} %")
out += ["", "top 25 instructions by samples:"]
for r in sorted(data, key=lambda r: -int(r[ins] or 0))[:25]:
    top = sorted(((int(r[i] or 0), hdr[i]) for i in stall), reverse=True)[:2]
    out.append(f"  {int(r[ins]):7d}  {r[isrc][:70]:70s}  " + ", ".join(f"{n}={v}" for v, n in top if v))
open(prefix + "_top_stalls.txt", "w").write("\n".join(out) + "\n")
print(open(prefix + "_ncu_metrics.txt").read())
