"""Opcode evidence for the built library: per kernel, how many tcgen05 / TMEM / TMA / reduction instructions its SASS
holds (B200_PROFILING.md "What proves a Blackwell-native kernel").  Runs on the CPU box:

    python tools/sass_histogram.py > profiles/r02_sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "multimodal-moe_b200", "b200moe", "libb200moe.so")
# SASS mnemonic prefix -> what was written in the source
WATCH = [("UTCHMMA", "tcgen05.mma kind::f16"), ("UTCBAR", "tcgen05.commit"), ("LDTM", "tcgen05.ld"), ("STTM", "tcgen05.st"),
         ("UTCATOMSWS", "tcgen05.alloc/dealloc"), ("UTMALDG", "TMA tensor load"), ("UTMASTG", "TMA tensor store"),
         ("UTMAREDG", "TMA tensor reduce-add"), ("UBLKCP", "cp.async.bulk"), ("UTMAPF", "TMA prefetch"),
         ("REDG", "red.global"), ("ATOMG", "atom.global"), ("HMMA", "legacy mma.sync"), ("LDGSTS", "cp.async"),
         ("SYNCS", "mbarrier"), ("MUFU.TANH", "tanh.approx"), ("MUFU.EX2", "ex2.approx"), ("ACQBULK", "cp.async.bulk wait")]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in sass.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur is not None:
            op = m.group(1)
            cur["_total"] += 1
            for pre, _ in WATCH:
                if op.startswith(pre):
                    cur[pre] += 1
                    if ".2CTA" in op:
                        cur[pre + ".2CTA"] += 1
    names = demangle(list(kernels))
    print(f"# SASS opcode histogram of {os.path.relpath(LIB, ROOT)} (cuobjdump -sass, sm_100a), {len(kernels)} kernels")
    print("# legend: " + "; ".join(f"{a} = {b}" for a, b in WATCH))
    total = collections.Counter()
    for k, c in kernels.items():
        short = re.sub(r"\(.*", "", names.get(k, k)).replace("b200moe::", "").replace("void ", "")
        parts = [f"{op}={n}" for op, n in sorted(c.items()) if op != "_total"]
        print(f"{short}: instructions={c['_total']} " + " ".join(parts))
        total.update(c)
    print("TOTAL: " + " ".join(f"{op}={n}" for op, n in sorted(total.items())))
    if total["HMMA"]:
        print("WARNING: legacy HMMA present", file=sys.stderr)


if __name__ == "__main__":
    main()
