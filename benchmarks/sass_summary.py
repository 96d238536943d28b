"""Per-kernel SASS summary of the built libraries (sm_100a): instruction count, registers, and the mnemonics that show
what the kernels are made of -- TMA bulk copies (UBLKCP), mbarrier operations (SYNCS), 128-bit global loads / stores,
warp votes / matches / shuffles, reductions.  No tensor-core instruction is expected: neither path is a contraction.

    python benchmarks/sass_summary.py > profiles/r2_sass_summary.txt
"""
import collections
import os
import re
import subprocess

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["UBLKCP", "SYNCS", "LDG.E.128", "LDG.E.64", "STG.E.128", "LDS", "STS", "LDL", "STL", "RED", "ATOM", "MATCH", "SHFL", "VOTE",
        "BAR", "FFMA", "MUFU", "PRMT", "HMMA", "UTCMMA", "NANOSLEEP"]


def main():
    for lib in ("libb200sc.so", "libb200pt.so"):
        path = os.path.join(REPO, "cuda_pathtracer_b200", lib)
        sass = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
        res = subprocess.run(["cuobjdump", "-res-usage", path], capture_output=True, text=True).stdout
        regs = dict(re.findall(r"Function (\S+):\s*\n\s*REG:(\d+)", res))
        print(f"== {lib} ({os.path.getsize(path)} bytes) ==")
        cur, counts, total = None, None, 0
        out = []

        def flush():
            if cur:
                demangled = subprocess.run(["c++filt", cur], capture_output=True, text=True).stdout.strip().split("(")[0]
                keys = "  ".join(f"{k} {counts[k]}" for k in KEYS if counts[k])
                out.append(f"{demangled[:70]:70s} {total:6d} instr  {regs.get(cur, '?'):>3s} regs  | {keys}")

        for ln in sass.splitlines():
            m = re.match(r"\s*Function : (\S+)", ln)
            if m:
                flush()
                cur, counts, total = m.group(1), collections.Counter(), 0
                continue
            m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
            if m and cur:
                total += 1
                op = m.group(1)
                for k in KEYS:
                    if k in ("LDG.E.128", "LDG.E.64", "STG.E.128"):        # (cache-hint suffixes sit between the opcode and the width)
                        if op.startswith(k[:3]) and op.endswith("." + k.rsplit(".", 1)[1]) or op.startswith(k[:3]) and ("." + k.rsplit(".", 1)[1] + ".") in op:
                            counts[k] += 1
                    elif op.startswith(k):
                        counts[k] += 1
        flush()
        print("\n".join(sorted(out)))
        print()


if __name__ == "__main__":
    main()
