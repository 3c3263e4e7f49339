"""Pretty-print a bench.py JSON line: python tools/show_bench.py gpurun_out/bench.json"""
import json, sys
j = json.load(open(sys.argv[1]))
print("value %.4g %s  ms/step %.4f  graph=%s eager_ms=%.3f  launches/step=%s" % (
    j["value"], j["unit"], j["ms_per_step"], j.get("cuda_graph"), j.get("eager_ms_per_step", 0), j.get("gpu_launches_per_step")))
if "graph_error" in j: print("graph_error", j["graph_error"][:200])
print("e2e", j.get("e2e")); print("roofline", j.get("roofline")); print("cpu", j.get("cpu_baseline")); print("clocks", j.get("clocks"))
print("per_call", j.get("per_call_ms"))
for s in j.get("sweep", []):
    if "error" in s: print(s); continue
    print(s["workload"], s["tokens"], "ms=%.3f" % s["ms_per_step"], "graph", s["cuda_graph"], "tok/s=%.3g" % s["tokens_per_s"],
          "gemm_ms=%.3f" % s["gemm_ms"], "TF=%.0f" % s["gemm_tflops"], "frac=%.3f" % s["gemm_frac_of_peak"])
    print("    ", {k: round(v, 3) for k, v in s["per_call_ms"].items()})
    if "hbm_kernels" in s:
        print("     hbm:", {k: (v["gbs"], v["frac"]) for k, v in s["hbm_kernels"].items()})
