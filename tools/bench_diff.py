#!/usr/bin/env python3
"""Ad-hoc: K4 manifest diff on device-resident manifests (rows of 16 bytes), against HBM bandwidth."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tpi_b200 as tpi

with tpi.Context([0]) as c:
    st = torch.cuda.current_stream().cuda_stream
    for n in (1 << 20, 1 << 24, 1 << 27):
        a = torch.randint(0, 256, (n, 16), dtype=torch.uint8, device="cuda")
        b = a.clone()
        b[:: 20, 7] ^= 1                                        # 5 % changed
        idx = torch.empty(n, dtype=torch.int32, device="cuda")
        cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
        for _ in range(3):
            c.dev_diff(a, b, n, idx, cnt, stream=st)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            c.dev_diff(a, b, n, idx, cnt, stream=st)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        assert cnt.item() == (n + 19) // 20
        alg = 2 * 16 * n + 4 * cnt.item()                       # each manifest read once (32 B per row) + indices written
        print(json.dumps({"rows": n, "ms": round(ms, 4), "rows_per_s": round(n / ms * 1e3), "algorithmic_GBps": round(alg / ms / 1e6, 1)}))
