"""Summarise an .ncu-rep (raw page) into a markdown table:  python tools/ncu_summary.py rep.ncu-rep out.md "title" """
import csv, subprocess, sys
rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
keep = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "smsp__inst_executed.sum", "sm__cycles_elapsed.avg"]
keep = [k for k in keep if k in idx]
with open(out, "w") as f:
    f.write(f"# {title}\n\n| kernel | " + " | ".join(keep) + " |\n|---|" + "---|" * len(keep) + "\n")
    for r in rows[2:]:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").replace("b200moe::", "")
        f.write("| " + name + " | " + " | ".join(f"{r[idx[k]]} {units[idx[k]]}".strip() for k in keep) + " |\n")
print(open(out).read())
