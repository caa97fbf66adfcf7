#!/usr/bin/env python
"""SASS evidence for the shipped library: per-kernel counts of the mnemonics that prove the Blackwell paths
(tcgen05.mma = UTCHMMA, TMA = UTMALDG/UTMASTG/UBLKCP, TMEM = LDTM/STTM, tcgen05.commit = UTCBAR) and of the
peer-memory traffic (ld.global.nc on peer pointers = LDG.*.CONSTANT, system-scope LDG/ST, RED/ATOM), plus full
listings of the fused compute+collective kernels. Runs without a GPU.
    python scripts/sass_summary.py            -> profiles/sass_summary_r2.txt, profiles/sass_*_r2.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "dlti_b200", "lib", "libdlti_b200.so")
TAG = sys.argv[1] if len(sys.argv) > 1 else "r2"
BASES = ("UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UBLKCP", "LDTM", "STTM", "UTCBAR", "UTCCP", "HMMA", "LDGSTS", "RED",
         "REDG", "ATOM", "ATOMG", "MEMBAR", "ACQBULK", "PREEXIT", "UCGABAR_ARV", "UCGABAR_WAIT")
FULL = {"gemm2_kernelILb0": f"sass_gemm2_nt_k1_k10_k11_{TAG}.txt", "k9_step_kernel": f"sass_k9_step_{TAG}.txt",
        "rs_wait_accumulate": f"sass_k9_rs_wait_accumulate_{TAG}.txt",
        "tp_own_reduce_norm_kernelI13__nv_bfloat16Li4": f"sass_k10_own_reduce_norm_{TAG}.txt",
        "embedding_peer_kernelI13": f"sass_k15_embedding_peer_{TAG}.txt",
        "tp_reduce_add_norm_kernelI13__nv_bfloat16Li4ELb1": f"sass_k10_decode_barrier_reduce_norm_{TAG}.txt"}


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"\n\s*Function : ", txt)
    rows, bodies = [], {}
    for p in parts[1:]:
        name = p.split("\n", 1)[0].strip()
        mn = collections.Counter()
        n = 0
        for m in re.finditer(r"/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", p):
            n += 1
            op = m.group(1)
            base = op.split(".")[0]
            if base in BASES:
                mn[base] += 1
            if base == "LDG" and "CONSTANT" in op:
                mn["LDG.CONSTANT"] += 1
            if base in ("LDG", "LD") and ".SYS" in op:
                mn["LD*.SYS"] += 1
            if base in ("STG", "ST") and ".SYS" in op:
                mn["ST*.SYS"] += 1
            if ".2CTA" in op and base in ("UTCHMMA", "UTMALDG", "UTCBAR"):
                mn[base + ".2CTA"] += 1
        rows.append((name, n, mn))
        bodies[name] = p
    dem = subprocess.run(["c++filt"], input="\n".join(r[0] for r in rows), capture_output=True, text=True).stdout.splitlines()
    out = [f"SASS mnemonic summary of dlti_b200/lib/libdlti_b200.so (cuobjdump -sass, sm_100a; scripts/sass_summary.py)",
           "tcgen05.mma = UTCHMMA, TMA = UTMALDG/UTMASTG/UBLKCP, TMEM = LDTM/STTM, tcgen05.commit = UTCBAR; peer memory:",
           "LDG.CONSTANT = ld.global.nc (weight / activation pulls), LD*/ST*.SYS = system-scope flags and data, RED/ATOM.", ""]
    for (name, n, mn), d in sorted(zip(rows, dem), key=lambda x: -x[0][1]):
        short = d.replace("(anonymous namespace)::", "")[:120]
        out.append(f"{short}\n    {n} instructions | " + ", ".join(f"{k} {v}" for k, v in sorted(mn.items())))
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    open(os.path.join(ROOT, "profiles", f"sass_summary_{TAG}.txt"), "w").write("\n".join(out) + "\n")
    for key, fn in FULL.items():
        for name, body in bodies.items():
            if key in name:
                open(os.path.join(ROOT, "profiles", fn), "w").write("Function : " + body)
                break
    print(len(rows), "kernels summarised")


if __name__ == "__main__":
    main()
