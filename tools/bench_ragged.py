#!/usr/bin/env python3
"""Ad-hoc: kernel-only throughput of ragged device-resident units (uniform 0..8 KiB and a Zipf mix)."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tpi_b200 as tpi

def run(name, lengths):
    n = len(lengths)
    off = np.concatenate([[0], np.cumsum((lengths + 15) // 16 * 16)[:-1]]).astype(np.int64)
    total = int(off[-1] + lengths[-1])
    with tpi.Context([0]) as c:
        d = torch.empty((total + 15) // 16 * 16, dtype=torch.uint8, device="cuda")
        c.dev_synth_fill(d, 9, 0, 16, d.numel() // 16)
        d_off, d_len = torch.from_numpy(off).cuda(), torch.from_numpy(lengths.astype(np.int64)).cuda()
        dig = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(3):
            c.dev_hash_buffers(d, total, n, dig, d_offsets=d_off, d_lengths=d_len, stream=st)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            c.dev_hash_buffers(d, total, n, dig, d_offsets=d_off, d_lengths=d_len, stream=st)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(json.dumps({"case": name, "units": n, "GB": round(lengths.sum() / 1e9, 3), "ms": round(ms, 3), "GBps": round(lengths.sum() / ms / 1e6, 1)}))

rng = np.random.default_rng(1)
run("uniform 0..8KiB x 1Mi", rng.integers(0, 8192, 1 << 20))
k = np.arange(1, 400001)
z = np.maximum(1024, ((64 << 20) * k ** -1.111).astype(np.int64)); rng.shuffle(z)
run("zipf 1KiB..64MiB x 400k", z)
run("all 4096 (offset arrays) x 1Mi", np.full(1 << 20, 4096))
