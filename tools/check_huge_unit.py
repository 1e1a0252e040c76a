#!/usr/bin/env python3
"""One unit of 4 GiB + 100 bytes (byte count above 2^32, bit count above 2^35) through the device API
(one lane of the general kernel, ~35 s) and, chained across staging buffers, through the host API."""
import hashlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tpi_b200 as tpi
L = (4 << 30) + 100
with tpi.Context([0]) as c:
    d = torch.empty((L + 15) // 16 * 16, dtype=torch.uint8, device="cuda")
    c.dev_synth_fill(d, 11, 0, 16, d.numel() // 16)
    dig = torch.zeros((1, 16), dtype=torch.uint8, device="cuda")
    t0 = time.time()
    c.dev_hash_buffers(d, L, 1, dig, stride=(L + 15) // 16 * 16)
    torch.cuda.synchronize()
    t1 = time.time()
    host = d.cpu().numpy()[:L]
    want = hashlib.md5(host).digest()
    assert dig.cpu().numpy().tobytes() == want, "device API digest differs"
    t2 = time.time()
    got = c.hash_buffers(host, [0], [L])
    assert got[0].tobytes() == want, "host API digest differs"
    print(f"huge unit ok: device {t1 - t0:.1f} s ({L / (t1 - t0) / 1e6:.0f} MB/s), host path {time.time() - t2:.1f} s")
