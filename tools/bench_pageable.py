#!/usr/bin/env python3
"""Ad-hoc: tpimd5_hash_buffers from pageable host memory (what a Go slice is): packed through pinned staging
by the worker pool (default) against a direct cudaMemcpy from pageable memory (TPIMD5_PAGEABLE_DIRECT=1)."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tpi_b200 as tpi
n, flen = 1 << 19, 4096
buf = np.random.default_rng(1).integers(0, 256, n * flen, dtype=np.uint8)
off = np.arange(n, dtype=np.uint64) * flen
ln = np.full(n, flen, dtype=np.uint64)
out = np.empty((n, 16), np.uint8)
with tpi.Context([0]) as c:
    for _ in range(2):
        c.hash_buffers(buf, off, ln, out=out)
    t0 = time.perf_counter()
    for _ in range(5):
        c.hash_buffers(buf, off, ln, out=out)
    dt = (time.perf_counter() - t0) / 5
print(json.dumps({"mode": "direct" if os.environ.get("TPIMD5_PAGEABLE_DIRECT") == "1" else "packed", "GiB": n * flen / 2**30, "ms": round(dt * 1e3, 1), "GBps": round(n * flen / dt / 1e9, 2)}))
