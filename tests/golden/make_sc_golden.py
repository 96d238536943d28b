"""Generates tests/golden/sc_golden.npz from the REFERENCE's own CPU scan/compact
(stream_compaction/stream_compaction/cpu.cu, compiled unmodified into oracle/_ref/libref_sc_cpu.so
by oracle/Makefile).  Run in the build container (needs /root/reference):

    make oracle && python tests/golden/make_sc_golden.py

Vectors: the harness's own generator (glibc rand, seed 565, last POT element 0) at n = 2^12 and
2^12-3 for scan (maxval 50) and compact (maxval 4); an int32-overflow case; tiny/ragged sizes.
Stored compactly (inputs + reference outputs).
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from oracle_bindings import ScOracle, gen_array  # noqa: E402


def main():
    ref = ScOracle(reference=True)
    out = {}
    pot = 1 << 12
    a50 = gen_array(pot, 50)
    a4 = gen_array(pot, 4)
    out["scan_in"] = a50
    out["scan_pot"] = ref.scan(a50)
    out["scan_npot"] = ref.scan(a50[:pot - 3])
    out["compact_in"] = a4
    out["compact_pot"] = ref.compact_without_scan(a4)
    out["compact_npot"] = ref.compact_without_scan(a4[:pot - 3])
    out["compact_ws_pot"] = ref.compact_with_scan(a4)
    out["compact_ws_npot"] = ref.compact_with_scan(a4[:pot - 3])
    rng = np.random.RandomState(565)
    big = rng.randint(-2**31, 2**31 - 1, size=1000, dtype=np.int64).astype(np.int32)
    out["wrap_in"] = big
    out["wrap_scan"] = ref.scan(big)
    for n in (1, 2, 3, 31, 32, 33, 63, 64, 65):
        x = rng.randint(0, 4, size=n).astype(np.int32)
        out[f"tiny{n}_in"] = x
        out[f"tiny{n}_scan"] = ref.scan(x)
        out[f"tiny{n}_compact"] = ref.compact_without_scan(x)
    path = os.path.join(os.path.dirname(__file__), "sc_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
