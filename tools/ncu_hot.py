"""Top stalled SASS instructions of one kernel in an .ncu-rep:  python tools/ncu_hot.py rep.ncu-rep <kernel regex> [N]"""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
raw = subprocess.run(["ncu", "-i", rep, "--kernel-name-base", "demangled", "-k", "regex:" + pat, "--page", "source",
                      "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}
        blocks.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None and r:
        cur["rows"].append(r)
for b in blocks:
    h = {k: i for i, k in enumerate(b["hdr"])}
    si = h["Warp Stall Sampling (All Samples)"]
    stall_cols = [k for k in b["hdr"] if k.startswith("stall_") and "Not Issued" not in k]
    tot = sum(int(r[si] or 0) for r in b["rows"])
    print("==", b["name"][:110], "samples", tot)
    agg = {k: sum(int(r[h[k]] or 0) for r in b["rows"]) for k in stall_cols}
    print("   stall mix:", ", ".join(f"{k[6:]}={v}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    top = sorted(range(len(b["rows"])), key=lambda i: -int(b["rows"][i][si] or 0))[:n]
    for i in sorted(top):
        r = b["rows"][i]
        why = max(stall_cols, key=lambda k: int(r[h[k]] or 0))
        print(f"   {i:5d} {int(r[si] or 0):6d} {why[6:]:14s} {r[h['Source']].strip()[:100]}")
