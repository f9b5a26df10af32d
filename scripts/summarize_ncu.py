"""Turns the ncu launch list of scripts/profile_round.sh into the tracked files under profiles/:
  r01_launches.csv          one line per launch (id, kernel, grid, block, ns, DRAM read/write bytes)
  r01_launch_summary.md     per-kernel totals and shares
  r01_conv_tc_traffic.json  mean measured DRAM traffic of the conv_tc_kernel launches
usage: python scripts/summarize_ncu.py gpurun_out/r01_launches_full.csv [round tag]
"""
import collections
import csv
import json
import re
import sys

src = sys.argv[1]
tag = sys.argv[2] if len(sys.argv) > 2 else "r01"
rows = []
with open(src, newline="") as f:
    lines = [ln for ln in f if not ln.startswith("==")]
rd = csv.DictReader(lines)
per = collections.OrderedDict()
for r in rd:
    key = int(r["ID"])
    e = per.setdefault(key, {"kernel": r["Kernel Name"], "grid": r["Grid Size"], "block": r["Block Size"]})
    val = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    name = r["Metric Name"]
    if name == "gpu__time_duration.sum":
        e["ns"] = val * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1)
    else:
        mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
        e["rd" if "read" in name else "wr"] = val * mul


def short(k):
    k = re.sub(r"^void ", "", k)
    k = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", k)
    k = re.sub(r"\(.*$", "", k)
    return k[:90]


with open(f"profiles/{tag}_launches.csv", "w") as f:
    f.write("id,kernel,grid,block,time_ns,dram_read_bytes,dram_write_bytes\n")
    for i, e in per.items():
        f.write(f'{i},"{short(e["kernel"])}","{e["grid"]}","{e["block"]}",{int(e.get("ns", 0))},'
                f'{int(e.get("rd", 0))},{int(e.get("wr", 0))}\n')

agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for e in per.values():
    a = agg[short(e["kernel"])]
    a[0] += 1
    a[1] += e.get("ns", 0)
    a[2] += e.get("rd", 0)
    a[3] += e.get("wr", 0)
total = sum(a[1] for a in agg.values())
cmd = "python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --pipeline 0"
with open(f"profiles/{tag}_launch_summary.md", "w") as f:
    f.write(f"# {tag}: ncu launch list of `{cmd}`\n\n")
    f.write("`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum "
            "--clock-control none -c 9000` (the command exited 0 without ncu first). "
            f"{len(per)} launches: one I frame + two P frames of the timed loop and the P frame of the "
            "per-kernel profile pass, sequential schedule. Per-launch times are cold-cache and "
            "serialised, so only the kernels' SHARES are comparable with bench.py's live CUDA-event "
            f"table (`{tag}_pframe_kernel_table.json`). Compact per-launch list: `{tag}_launches.csv`.\n\n")
    f.write("| kernel | launches | total us | share | DRAM bytes / launch |\n|---|---:|---:|---:|---:|\n")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
        f.write(f"| `{k}` | {a[0]} | {a[1] / 1e3:.0f} | {a[1] / total:.3f} | "
                f"{(a[2] + a[3]) / a[0] / 1e6:.1f} MB |\n")
    f.write(f"\ntotal {total / 1e6:.1f} ms over {len(per)} launches\n")

a = agg.get("conv_tc_kernel")
if a:
    json.dump({"kernel": "conv_tc_kernel", "launches": a[0],
               "dram_bytes_per_launch": (a[2] + a[3]) / a[0],
               "dram_read_bytes_per_launch": a[2] / a[0],
               "dram_write_bytes_per_launch": a[3] / a[0],
               "avg_launch_us": a[1] / a[0] / 1e3,
               "share_of_listed_time": a[1] / total,
               "source": f"profiles/{tag}_launches.csv (ncu --metrics dram__bytes_read.sum,"
                         "dram__bytes_write.sum,gpu__time_duration.sum; 1 I + 3 P frames, "
                         "sequential schedule)"},
              open(f"profiles/{tag}_conv_tc_traffic.json", "w"), indent=1)
print(f"{len(per)} launches, {total / 1e6:.1f} ms")
