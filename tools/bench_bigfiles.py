#!/usr/bin/env python3
"""A workdir with several files larger than a staging buffer next to small ones, through tpimd5_hash_files: every
long file is one serial chain, so the pass cannot end before (longest file / ~0.16 GB/s); checked against the oracle.
    python tools/bench_bigfiles.py [big files] [bytes each]"""
import os, sys, time, shutil, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import tpi_b200 as tpi, oracle
root = tempfile.mkdtemp(prefix="tpi_big_", dir="/dev/shm")
try:
    nbig, big = int(sys.argv[1]) if len(sys.argv) > 1 else 6, int(sys.argv[2]) if len(sys.argv) > 2 else (1 << 30)
    chunk = oracle.synth_files(9, 0, 1 << 20, 64, threads=16).tobytes()
    paths = []
    for i in range(nbig):
        p = os.path.join(root, f"big{i}.bin"); paths.append(p)
        with open(p, "wb") as f:
            left = big + i * 4097
            while left:
                k = min(left, len(chunk)); f.write(chunk[i:k] if k > i else chunk[:k]); left -= len(chunk[i:k]) if k > i else k
    for i in range(3000):
        p = os.path.join(root, f"s{i}.bin"); paths.append(p)
        with open(p, "wb") as f: f.write(chunk[i: i + (i * 733) % 300000])
    sizes = [os.path.getsize(p) for p in paths]
    ctx = tpi.Context([0])
    ctx.hash_files(paths[nbig:])     # warm
    t0 = time.perf_counter(); got = ctx.hash_files(paths); t = time.perf_counter() - t0
    t0 = time.perf_counter(); want = oracle.md5_files(paths, threads=16); tc = time.perf_counter() - t0
    print("files", len(paths), "bytes GiB", round(sum(sizes) / 2**30, 2), "gpu s", round(t, 2), "cpu(16 thr) s", round(tc, 2), "equal", bool(np.array_equal(got, want)),
          "longest chain floor s", round(max(sizes) / 0.161e9, 2))
finally:
    shutil.rmtree(root, ignore_errors=True)
