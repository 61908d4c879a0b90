"""summarise an .ncu-rep (raw page) per kernel launch: duration, DRAM traffic, throughput, stalls"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
def g(r, k):
    try: return float(r[col[k]].replace(",", ""))
    except Exception: return float("nan")
want = [("gpu__time_duration.sum", "dur"), ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%")]
print("| kernel | grid | " + " | ".join(n + (" [" + units[col[k]] + "]" if k in col else "") for k, n in want) + " |")
print("|---|---|" + "---|" * len(want))
for r in data:
    name = re.sub(r"\(.*", "", r[col["Kernel Name"]])
    grid = r[col["Grid Size"]] if "Grid Size" in col else ""
    print(f"| {name} | {grid} | " + " | ".join(f"{g(r, k):.4g}" if k in col else "-" for k, _ in want) + " |")
if len(sys.argv) > 2:  # stall reasons for launch index argv[2]
    r = data[int(sys.argv[2])]
    st = [(h, g(r, h)) for h in hdr if "smsp__average_warp" in h and "issue_stalled" in h and h.endswith("_per_issue_active.ratio") and "not_issued" not in h]
    for h, v in sorted(st, key=lambda x: -x[1])[:8]:
        print(f"  {h.split('issue_stalled_')[1].split('_per_')[0]:28s} {v:.2f}")
