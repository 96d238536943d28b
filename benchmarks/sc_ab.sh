#!/bin/bash
# A/B of prebuilt scan-library variants (variants/<name>/libb200sc.so): bit-exact tests, then scan / compact at 2^28 and 2^24
mkdir -p gpurun_out
cp cuda_pathtracer_b200/libb200sc.so /tmp/keep_sc.so
for v in main "$@"; do
  [ "$v" != main ] && cp variants/$v/libb200sc.so cuda_pathtracer_b200/libb200sc.so
  timeout 600 python -m pytest tests/test_sc_gpu.py -m gpu -x -q > gpurun_out/sc_ab_${v}_pytest.log 2>&1; echo "$v pytest rc=$? $(tail -1 gpurun_out/sc_ab_${v}_pytest.log)"
  python - "$v" <<'PY'
import sys, torch
sys.path.insert(0, '.')
from cuda_pathtracer_b200 import stream_compaction as sc
ctx = sc.Context(); s = torch.cuda.current_stream()
for lg in (28, 26, 24):
    n = 1 << lg
    a = torch.randint(0, 50, (n,), dtype=torch.int32, device='cuda'); b = a % 4; out = torch.empty_like(a)
    cnt = torch.zeros(1, dtype=torch.int32, device='cuda')
    def timed(fn, reps=20):
        for _ in range(3): fn()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s); fn(); e1.record(s); e1.synchronize(); ts.append(e0.elapsed_time(e1))
        ts.sort(); return ts[len(ts) // 2]
    ms_s = timed(lambda: ctx.scan(n, out, a, sc.LOOKBACK, stream=s.cuda_stream))
    ms_c = timed(lambda: ctx.compact(n, out, b, cnt, stream=s.cuda_stream)); k = int(cnt.item())
    print(sys.argv[1], 'lg', lg, 'scan %.4f ms %.3f  compact %.4f ms %.3f' % (ms_s, 8.0 * n / ms_s / 1e6 / 6552.3, ms_c, (4.0 * n + 4.0 * k) / ms_c / 1e6 / 6552.3))
    del a, b, out
PY
done
cp /tmp/keep_sc.so cuda_pathtracer_b200/libb200sc.so
