"""Static evidence from the built library (no GPU needed): per kernel, registers / shared memory / spills (cuobjdump -res-usage)
and the SASS mnemonics that show what the hot kernels are made of — UBLKCP = cp.async.bulk (TMA 1-D bulk copy), SYNCS.* = mbarrier,
MEMBAR.*.SYS / .SYS loads and stores = system-scope ordering of the peer transport, DADD/DMUL/DFMA = fp64 arithmetic.
Usage: python tools/sass_summary.py > profiles/rNN_sass_resource_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "conjugategradient_b200", "lib", "libMgcgGpu.so")
PAT = re.compile(r"\b(UBLKCP|UTMALDG|UTMASTG|SYNCS|MEMBAR\.[A-Z.]+|DFMA|DADD|DMUL|REDUX|ATOMG|RED|LDG|STG|LDS|STS|SHFL|BAR)\b")


def demangle(names):
    return subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.split("\n")


def main():
    res = subprocess.run(["cuobjdump", "-res-usage", SO], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for line in res.split("\n"):
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
        elif cur and "REG:" in line:
            usage[cur] = " ".join(re.findall(r"(?:REG|STACK|SHARED|LOCAL):\d+", line))
            cur = None
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    funcs = re.split(r"\n\s*Function : ", sass)[1:]
    names = [f.split("\n", 1)[0].strip() for f in funcs]
    dem = demangle(names)
    print(f"# {os.path.relpath(SO, ROOT)}: {len(funcs)} kernels, sm_100a SASS (cuobjdump -sass / -res-usage)")
    want = sys.argv[1:] or ["k_spmv_tma", "k_spmv_vector", "k_spmv_stream", "k_update_xr", "k_update_p", "k_peer_", "k_spgemm_wmerge",
                            "k_spgemm_merge", "k_loop_condition", "k_mg_tail", "k_gen_poisson", "k_ell_"]
    for n, d, f in zip(names, dem, funcs):
        if not any(w in d for w in want):
            continue
        c = collections.Counter(PAT.findall(f))
        ninstr = len(re.findall(r"^\s+/\*[0-9a-f]{4}\*/", f, flags=re.M))
        short = re.sub(r"\(.*", "", d)
        print(f"{short:<58s} {usage.get(n, ''):<40s} instr={ninstr:<5d} " + " ".join(f"{k}={v}" for k, v in sorted(c.items())))


if __name__ == "__main__":
    main()
