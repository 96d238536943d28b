"""Two compactions of 2^28 ints (75 % kept) through the library's device entry point: the target of a single-launch ncu capture
(benchmarks/run_tma.sh: `ncu --set full -k regex:k_compact_tma -s 1 -c 1 python benchmarks/compact_once.py`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cuda_pathtracer_b200 import stream_compaction as sc  # noqa: E402

n = 1 << 28
a = torch.randint(0, 4, (n,), dtype=torch.int32, device="cuda")
out = torch.empty_like(a)
cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
ctx = sc.Context()
for _ in range(2):
    ctx.compact(n, out, a, cnt)
torch.cuda.synchronize()
print("kept", int(cnt.item()), "of", n)
