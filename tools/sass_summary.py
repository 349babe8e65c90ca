"""Static listing of the kernels in libpynama_b200.so for profiles/: per kernel the registers, shared and
local (spill) bytes from `cuobjdump -res-usage`, and for the kernels of the hot loop the SASS instruction mix
(`cuobjdump -sass`): loads by width, DFMA / DADD / DMUL, shuffles, barriers, atomics, and whether any of
TMA (UBLKCP / UTMALDG), tcgen05 (UTC*MMA) or cluster instructions appear.  Runs without a GPU.

    python tools/sass_summary.py > profiles/r2_sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "pynama_b200", "libpynama_b200.so")
# the instantiations the 3-D bench workload (configs[1], ngl = 4) launches
HOT = ("spmv_kernel<3, 3, 32,", "spmv_kernel<3, 6,", "spmv_kernel<6, 3,", "cg_update_kernel", "cg_direction_kernel",
       "cg_scalars_p2p", "asm_mult_kernel<3>", "asm_sum_kernel<3>", "asm_coarse_fused_kernel<3>",
       "numeric_kle_kernel<3, 3, 2", "numeric_kernel<3, 3, 16>", "numeric_kernel<3, 6, 16>", "elem_kle_kernel<3, 4, 256>",
       "mf_apply_kernel<3, 4>", "mf_gather_kernel<3", "vtensv", "stress_combine", "maxpy_kernel", "noalloc", "fp64_peak")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def short(d):
    d = re.sub(r"^void ", "", d)
    d = re.sub(r"\(.*$", "", d)
    return d.replace("(anonymous namespace)::", "")


def main():
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    rows = []
    fn = None
    for line in res.split("\n"):
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            fn = m.group(1)
            continue
        m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+) CONSTANT\[0\]:(\d+)", line)
        if m and fn:
            rows.append((fn,) + tuple(int(v) for v in m.groups()))
            fn = None
    dm = demangle([r[0] for r in rows])
    print(f"# {os.path.relpath(LIB)}: {len(rows)} kernels, all sm_100a (cuobjdump -res-usage)")
    print(f"# {'kernel':<78} {'regs':>4} {'stack':>5} {'smem':>6} {'local':>5}")
    spilled = 0
    for fn, reg, stack, sh, loc, cst in sorted(rows, key=lambda r: short(dm[r[0]])):
        spilled += loc > 0
        print(f"{short(dm[fn])[:78]:<78} {reg:>4} {stack:>5} {sh:>6} {loc:>5}")
    print(f"# kernels with local-memory (spill) bytes: {spilled}")
    print()
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    cur, mix = None, {}
    for line in sass.split("\n"):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            mix[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_]*(?:\.[A-Z0-9_.]+)?)", line)
        if m and cur:
            mix[cur][m.group(1)] += 1
    dm2 = demangle(list(mix))
    print("# SASS instruction mix of the hot-loop kernels (cuobjdump -sass): total | LDG by width | DFMA DADD DMUL | SHFL | BAR |")
    print("# ATOM/RED | system-scope ld/st (peer flags) | TMA / tcgen05 / cluster instructions")
    for fn in sorted(mix, key=lambda f: short(dm2[f])):
        name = short(dm2[fn])
        if not any(h in name for h in HOT):
            continue
        c = mix[fn]
        tot = sum(c.values())
        grp = lambda pred: sum(v for k, v in c.items() if pred(k))      # noqa: E731
        ldg = {w: grp(lambda k, w=w: k.startswith("LDG") and f".{w}" in k) for w in ("128", "64")}
        ldg["32/other"] = grp(lambda k: k.startswith("LDG")) - sum(ldg.values())
        sysscope = grp(lambda k: (k.startswith("LD") or k.startswith("ST")) and ".SYS" in k)
        special = grp(lambda k: k.startswith(("UBLKCP", "UTMA", "UTC", "UCGABAR", "SYNCS")))
        print(f"{name[:70]:<70} {tot:>5} | LDG.128 {ldg['128']:>3} .64 {ldg['64']:>3} other {ldg['32/other']:>3} | "
              f"DFMA {c['DFMA']:>4} DADD {c['DADD']:>3} DMUL {c['DMUL']:>3} | SHFL {grp(lambda k: k.startswith('SHFL')):>3} | "
              f"BAR {grp(lambda k: k.startswith('BAR')):>2} | ATOM/RED {grp(lambda k: k.startswith(('ATOM', 'RED'))):>2} | "
              f"sys {sysscope:>2} | tma/tcgen05/cluster {special}")


if __name__ == "__main__":
    sys.exit(main())
