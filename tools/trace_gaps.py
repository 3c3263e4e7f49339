"""GPU timeline of a chrome trace (torch.profiler): kernels in time order with the idle gap before each.
python tools/trace_gaps.py trace.json [min_gap_us]"""
import json, sys
tr = json.load(open(sys.argv[1]))
ming = float(sys.argv[2]) if len(sys.argv) > 2 else 20.0
ev = [e for e in tr["traceEvents"] if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
ev.sort(key=lambda e: e["ts"])
t0 = ev[0]["ts"]
end = t0
busy = 0.0
print(f"{len(ev)} gpu events, span {(max(e['ts']+e['dur'] for e in ev)-t0)/1e3:.3f} ms")
for e in ev:
    gap = e["ts"] - end
    if gap > ming:
        print(f"  t={(e['ts']-t0)/1e3:8.3f} ms  gap {gap:7.1f} us before  {e['name'][:70]}  (stream {e['args'].get('stream')})")
    busy += max(0.0, e["ts"] + e["dur"] - max(end, e["ts"]))
    end = max(end, e["ts"] + e["dur"])
print(f"busy {busy/1e3:.3f} ms  idle {(end-t0-busy)/1e3:.3f} ms")
# longest kernels
agg = {}
for e in ev:
    k = e["name"][:60]
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += e["dur"]
for k, (n, d) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:22]:
    print(f"  {d/1e3:8.3f} ms  x{n:<4d} {k}")
