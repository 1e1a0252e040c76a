#!/usr/bin/env python3
"""Differential fuzz of the host-pointer API against the oracle (OpenSSL engine): random unit layouts
(dense / gapped / overlapping / shuffled), lengths around every block, stage and batch boundary, random
alignment, small staging buffers (many batches, chained units), pinned and pageable sources, files.
    python tools/fuzz_host_api.py [seconds] [seed]"""
import os, sys, tempfile, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import oracle
import tpi_b200 as tpi

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 12345
rng = np.random.default_rng(seed)
ENG = oracle.ENGINE_OPENSSL if oracle.lib().orc_has_openssl() else 0
EDGES = [0, 1, 55, 56, 63, 64, 65, 119, 120, 127, 128, 129, 191, 192, 255, 256, 257, 511, 512, 4095, 4096, 4097]


def lengths(n):
    kind = rng.integers(0, 5)
    if kind == 0:
        return rng.choice(EDGES, n)
    if kind == 1:
        return rng.integers(0, 300, n)
    if kind == 2:
        return (rng.pareto(1.1, n) * 500).astype(np.int64) % (3 << 20)
    if kind == 3:
        return np.full(n, int(rng.choice(EDGES + [8192, 70000])))
    base = rng.choice([64, 128, 256, 1 << 20])
    return (rng.integers(0, 6, n) * base + rng.choice([-1, 0, 1], n)).clip(0)


t_end, rounds, units = time.time() + budget, 0, 0
tmp = tempfile.mkdtemp(prefix="tpi_fuzz_")
while time.time() < t_end:
    n = int(rng.choice([1, 2, 31, 32, 33, 100, 1000, 3000]))
    ln = lengths(n).astype(np.int64)
    if rng.random() < 0.15:
        ln[int(rng.integers(0, n))] = int(rng.integers(1 << 20, 12 << 20))          # one unit larger than staging
    align = int(rng.choice([1, 4, 16, 64]))
    layout = rng.integers(0, 4)
    pos = np.concatenate([[0], np.cumsum((ln + align - 1) // align * align)[:-1]])
    if layout == 1:
        pos = pos + np.cumsum(rng.integers(0, 100, n)) * align                       # gaps
    total = int((pos + ln).max()) + 1
    if layout == 2:
        pos = rng.integers(0, max(1, total - ln.max()), n)                           # overlapping, unordered
    elif layout == 3:
        perm = rng.permutation(n)
        pos, ln = pos[perm], ln[perm]                                                # shuffled
    buf = rng.integers(0, 256, total + 64, dtype=np.uint8)
    off, ln64 = pos.astype(np.uint64), ln.astype(np.uint64)
    want = oracle.md5_batch(buf, off, ln64, threads=8, engine=ENG)
    staging = int(rng.choice([1, 2, 8])) << 20
    with tpi.Context([0], staging_bytes=staging) as c:
        src = buf
        if rng.random() < 0.5:
            pinned = torch.empty(buf.size, dtype=torch.uint8, pin_memory=True)
            pinned.numpy()[:] = buf
            src = pinned.numpy()
        got = c.hash_buffers(src, off, ln64)
        bad = np.nonzero((got != want).any(axis=1))[0]
        assert bad.size == 0, f"seed {seed} round {rounds}: units {bad[:5]} len {ln[bad[:5]]} off {pos[bad[:5]]} align {align} layout {layout} staging {staging}"
        if rounds % 5 == 0 and n <= 1000:
            paths = []
            for i in range(min(n, 200)):
                p = os.path.join(tmp, f"f{i}")
                with open(p, "wb") as f:
                    f.write(buf[int(pos[i]):int(pos[i] + ln[i])].tobytes())
                paths.append(p)
            gotf = c.hash_files(paths, ln[: len(paths)].astype(np.uint64))
            assert np.array_equal(gotf, want[: len(paths)]), f"seed {seed} round {rounds}: files differ"
    rounds += 1
    units += n
print(f"fuzz ok: {rounds} rounds, {units} units, seed {seed}")
