#!/usr/bin/env python3
"""Differential fuzz of the device-pointer API: strided units (tensor-map kernel when the stride is a
multiple of 16, general kernel otherwise), ragged last unit, offset/length arrays below and above the
bucketing threshold, S3 ETags with and without library scratch and through the allocation-free entry.
    python tools/fuzz_device_api.py [seconds] [seed]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import oracle
import tpi_b200 as tpi

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 4321
rng = np.random.default_rng(seed)
ENG = oracle.ENGINE_OPENSSL if oracle.lib().orc_has_openssl() else 0
t_end, rounds = time.time() + budget, 0
with tpi.Context([0]) as c:
    while time.time() < t_end:
        kind = rounds % 3
        if kind == 0:                                                     # strided
            stride = int(rng.choice([16, 32, 48, 64, 80, 128, 144, 250, 256, 1000, 4096, 4100, 65536, 1 << 20]))
            n = int(rng.integers(1, max(2, min(30000, (256 << 20) // stride))))
            shift = int(rng.choice([0, 0, 16, 3]))
            cut = int(rng.integers(0, stride)) if rng.random() < 0.5 else 0
            d = torch.empty(stride * n + 64, dtype=torch.uint8, device="cuda")
            c.dev_synth_fill(d, seed + rounds, 0, 16, d.numel() // 16)
            view = d[shift:]
            nbytes = stride * n - cut
            dig = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
            c.dev_hash_buffers(view, nbytes, n, dig, stride=stride)
            torch.cuda.synchronize()
            host = view.cpu().numpy()
            off = np.arange(n, dtype=np.uint64) * stride
            ln = np.minimum(np.uint64(stride), np.uint64(nbytes) - off).astype(np.uint64)
            want = oracle.md5_batch(host, off, ln, threads=8, engine=ENG)
            bad = np.nonzero((dig.cpu().numpy() != want).any(axis=1))[0]
            assert bad.size == 0, f"seed {seed} round {rounds}: strided stride {stride} n {n} shift {shift} cut {cut}: units {bad[:5]}"
        elif kind == 1:                                                   # offset arrays, ragged
            n = int(rng.choice([1, 100, 4095, 4096, 20000]))
            ln = (rng.pareto(1.2, n) * 400).astype(np.int64) % (1 << 20)
            align = int(rng.choice([1, 16]))
            off = np.concatenate([[0], np.cumsum((ln + align - 1) // align * align)[:-1]]).astype(np.int64)
            total = int(off[-1] + ln[-1]) + 16
            d = torch.empty((total + 15) // 16 * 16, dtype=torch.uint8, device="cuda")
            c.dev_synth_fill(d, seed + rounds, 0, 16, d.numel() // 16)
            dig = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
            c.dev_hash_buffers(d, total, n, dig, d_offsets=torch.from_numpy(off).cuda(), d_lengths=torch.from_numpy(ln).cuda())
            torch.cuda.synchronize()
            want = oracle.md5_batch(d.cpu().numpy(), off.astype(np.uint64), ln.astype(np.uint64), threads=8, engine=ENG)
            bad = np.nonzero((dig.cpu().numpy() != want).any(axis=1))[0]
            assert bad.size == 0, f"seed {seed} round {rounds}: ragged n {n} align {align}: units {bad[:5]} len {ln[bad[:5]]}"
        else:                                                             # S3 ETag
            part = int(rng.choice([64, 4096, 65536, (1 << 20) + 16, 5 << 20]))
            nparts = int(rng.integers(1, min(9000, max(2, (128 << 20) // part))))
            size = part * nparts - int(rng.integers(0, part))
            d = torch.empty((size + 31) // 16 * 16, dtype=torch.uint8, device="cuda")
            c.dev_synth_fill(d, seed + rounds, 0, 16, d.numel() // 16)
            e = [torch.zeros(16, dtype=torch.uint8, device="cuda") for _ in range(3)]
            parts = torch.zeros((tpi.s3_nparts(size, part), 16), dtype=torch.uint8, device="cuda")
            c.dev_s3_etag(d, size, part, e[0], d_part_digests=parts)
            c.dev_s3_etag(d, size, part, e[1])
            tpi.dev_s3_etag_noalloc(0, d, size, part, e[2])
            torch.cuda.synchronize()
            o_etag, o_n, o_parts = oracle.s3_etag(d.cpu().numpy()[:size], part, threads=8, engine=ENG)
            assert np.array_equal(parts.cpu().numpy(), o_parts), f"seed {seed} round {rounds}: parts differ (part {part} size {size})"
            for x in e:
                assert x.cpu().numpy().tobytes() == o_etag, f"seed {seed} round {rounds}: etag differs (part {part} size {size})"
        rounds += 1
print(f"fuzz ok: {rounds} rounds, seed {seed}")
