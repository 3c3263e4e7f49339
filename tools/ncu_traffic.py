"""DRAM traffic of the grouped-GEMM launches of ONE profiled step, from an ncu --set full report, merged into
profiles/r02_traffic.json under the key bench.py looks up (workload + tokens, e.g. A3200, C65533):

    python tools/ncu_traffic.py gpurun_out/ncu_A.ncu-rep A3200 "tools/profile_step.py A 8 400 1" [out.json]
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, key, how = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def nbytes(r, col):
    return float(r[idx[col]]) * SCALE[units[idx[col]]]


gemms = [r for r in rows[2:] if "grouped_gemm_kernel" in r[idx["Kernel Name"]]]
total = sum(nbytes(r, "dram__bytes_read.sum") + nbytes(r, "dram__bytes_write.sum") for r in gemms)
path = sys.argv[4] if len(sys.argv) > 4 else os.path.join(ROOT, "profiles", "r02_traffic.json")
j = json.load(open(path)) if os.path.exists(path) else {}
j[key] = {"gemm_family_dram_bytes_per_step": total, "gemm_launches_per_step": len(gemms),
          "source": f"ncu --set full --clock-control none, {how}: dram__bytes_read.sum + dram__bytes_write.sum of the "
                    f"{len(gemms)} grouped_gemm_kernel launches"}
json.dump(j, open(path, "w"), indent=1)
print(key, j[key])
