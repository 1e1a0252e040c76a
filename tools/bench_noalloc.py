#!/usr/bin/env python3
"""The allocation-free S3 ETag entry point next to the ordinary one, device-resident data, CUDA events."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import tpi_b200 as tpi  # noqa: E402

torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream
with tpi.Context([0]) as c:
    for size, part in (((8 << 30), 1 << 20), ((8 << 30) + 12345, (1 << 20) + 16), (10000 * (100 << 10), 100 << 10), ((2 << 30) + 3, 8 << 20)):
        d = torch.empty((size + 31) // 16 * 16, dtype=torch.uint8, device="cuda")
        c.dev_synth_fill(d, 12, 0, 16, d.numel() // 16, stream=st)
        a, b = torch.zeros(16, dtype=torch.uint8, device="cuda"), torch.zeros(16, dtype=torch.uint8, device="cuda")
        res = {}
        for name, fn in (("dev_s3_etag", lambda: c.dev_s3_etag(d, size, part, a, stream=st)),
                         ("dev_s3_etag_noalloc", lambda: tpi.dev_s3_etag_noalloc(0, d, size, part, b, stream=st))):
            for _ in range(2):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3):
                fn()
            e1.record()
            torch.cuda.synchronize()
            res[name] = e0.elapsed_time(e1) / 3
        assert torch.equal(a, b)
        print(json.dumps({"bytes": size, "part": part, "parts": tpi.s3_nparts(size, part),
                          "dev_s3_etag_ms": round(res["dev_s3_etag"], 3), "noalloc_ms": round(res["dev_s3_etag_noalloc"], 3),
                          "noalloc_GBps": round(size / res["dev_s3_etag_noalloc"] / 1e6, 1)}))
        del d
