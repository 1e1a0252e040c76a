#!/usr/bin/env python3
"""Ad-hoc: BASELINE config 2 through real files (a slice of it): N x 4 KiB files in a 1,024-way
directory fan-out on /dev/shm, tpimd5_hash_files against the CPU (OpenSSL, one file per thread)."""
import json
import os
import shutil
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
import tpi_b200 as tpi  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
root = tempfile.mkdtemp(prefix="tpi_c2f_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
try:
    buf = oracle.synth_bytes(2, 0, 4096).tobytes()
    paths = []
    for d in range(1024):
        os.makedirs(os.path.join(root, f"{d:04d}"))
    per_dir = (n + 1023) // 1024
    for i in range(n):
        p = os.path.join(root, f"{i // per_dir:04d}", f"{i:07d}")       # listing order: directory by directory
        with open(p, "wb") as f:
            f.write(buf[: 4096 - (i % 3)] + bytes(i % 3))
        paths.append(p)
    sizes = [4096] * n
    with tpi.Context([0]) as c:
        c.hash_files(paths[:1000], sizes[:1000])
        t = []
        for _ in range(3):
            t0 = time.perf_counter()
            got = c.hash_files(paths, sizes)
            t.append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        m = tpi.Manifest.scan(c, root, tpi.Filter([]))
        t_scan = time.perf_counter() - t0
        assert len(m) == n
    eng = oracle.ENGINE_OPENSSL if oracle.lib().orc_has_openssl() else 0
    oracle.md5_files(paths[:1000], 4, eng)
    t0 = time.perf_counter()
    want = oracle.md5_files(paths, os.cpu_count(), eng)
    tc = time.perf_counter() - t0
    assert np.array_equal(got, want)
    print(json.dumps({"files": n, "gpu_path_s": round(min(t), 3), "gpu_files_per_s": round(n / min(t)), "gpu_GBps": round(n * 4096 / min(t) / 1e9, 3),
                      "manifest_scan_s": round(t_scan, 3), "cpu_s": round(tc, 3), "cpu_files_per_s": round(n / tc), "cpu_GBps": round(n * 4096 / tc / 1e9, 3), "cpu_threads": os.cpu_count()}))
finally:
    shutil.rmtree(root, ignore_errors=True)
