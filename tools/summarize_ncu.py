"""Summarise an `ncu --set full` report: one block per captured launch with the metrics DESIGN.md quotes.
usage: ncu -i X.ncu-rep --page raw --csv > raw.csv ; python tools/summarize_ncu.py raw.csv > profiles/rNN_ncu_top_kernels.txt"""
import csv, re, sys

WANT = [
    ("time", "gpu__time_duration.sum"),
    ("dram_rd", "dram__bytes_read.sum"),
    ("dram_wr", "dram__bytes_write.sum"),
    ("grid", "launch__grid_size"),
    ("block", "launch__block_size"),
    ("regs", "launch__registers_per_thread"),
    ("smem_dyn", "launch__shared_mem_per_block_dynamic"),
    ("warps_active%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("issue_active%", "sm__inst_issued.avg.pct_of_peak_sustained_active"),
    ("warp_inst", "smsp__inst_executed.sum"),
    ("l1_hit%", "l1tex__t_sector_hit_rate.pct"),
    ("l2_hit%", "lts__t_sector_hit_rate.pct"),
    ("l2_thr%", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("sm_thr%", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("dram_thr%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0,
         "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}

rows = list(csv.reader(l for l in open(sys.argv[1], newline="") if not l.startswith("==")))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
print("# ncu --set full --clock-control none; one block per captured launch; bytes and times are per launch")
for r in data:
    name = re.sub(r"\(.*", "", r[col["Kernel Name"]])
    print("\n" + name)
    vals = {}
    for label, metric in WANT:
        if metric not in col:
            continue
        v, u = r[col[metric]].replace(",", ""), units[col[metric]]
        try:
            vals[label] = float(v) * SCALE.get(u, 1.0)
        except ValueError:
            continue
        print("  %-16s %s %s" % (label, v, u))
    if "time" in vals and "dram_rd" in vals:
        print("  %-16s %.0f GB/s (dram bytes / duration under ncu)"
              % ("dram_bw", (vals["dram_rd"] + vals.get("dram_wr", 0.0)) / vals["time"] / 1e9))
