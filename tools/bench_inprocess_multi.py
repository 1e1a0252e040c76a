#!/usr/bin/env python3
"""ONE process, one context over N GPUs (what a cgo caller has): C2-shaped units in one pinned host buffer,
tpimd5_hash_buffers end to end, 4 GiB per device (the same per-GPU work as a torchrun rank of bench.py), next to the
H2D-only ceiling of the same buffer with all devices copying at once."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import oracle
import tpi_b200 as tpi
ndev = tpi.device_count()
n_max, flen = (1 << 20) * ndev, 4096
host = torch.empty(n_max * flen, dtype=torch.uint8, pin_memory=True)
hn = host.numpy()
hn[:] = 0
hn[: 1 << 26] = np.random.default_rng(1).integers(0, 256, 1 << 26, dtype=np.uint8)
for devs in sorted({1, 2, 4, ndev}):
    if devs > ndev:
        continue
    n = (1 << 20) * devs
    off = np.arange(n, dtype=np.uint64) * flen
    ln = np.full(n, flen, dtype=np.uint64)
    out = np.empty((n, 16), np.uint8)
    with tpi.Context(list(range(devs))) as c:
        for _ in range(2):
            c.hash_buffers(hn.ctypes.data, off, ln, out=out)
        t0 = time.perf_counter()
        for _ in range(3):
            c.hash_buffers(hn.ctypes.data, off, ln, out=out)
        dt = (time.perf_counter() - t0) / 3
    # ceiling of this leg: the same pinned buffer cut into `devs` slices, every device copying its slice at once
    per = (n // devs) * flen
    dst = [torch.empty(per, dtype=torch.uint8, device=f"cuda:{d}") for d in range(devs)]
    streams = [torch.cuda.Stream(device=f"cuda:{d}") for d in range(devs)]

    def copy_all():
        for d in range(devs):
            with torch.cuda.stream(streams[d]):
                dst[d].copy_(host[d * per:(d + 1) * per], non_blocking=True)
        for d in range(devs):
            streams[d].synchronize()
    copy_all()
    t0 = time.perf_counter()
    for _ in range(3):
        copy_all()
    h2d = devs * per * 3 / (time.perf_counter() - t0) / 1e9
    del dst
    sel = [0, 1, n // 2, n - 1]
    want = oracle.md5_batch(hn, off[sel], ln[sel])
    assert np.array_equal(out[sel], want)
    print(json.dumps({"devices_in_one_context": devs, "GiB": n * flen / 2**30, "ms": round(dt * 1e3, 1), "GBps": round(n * flen / dt / 1e9, 1),
                      "h2d_only_ceiling_GBps": round(h2d, 1), "frac_of_ceiling": round(n * flen / dt / 1e9 / h2d, 3)}))
