#!/usr/bin/env python3
"""Ad-hoc: C3 end to end for one file on tmpfs: tpimd5_s3_etag_file (8 MiB parts) against the CPU
(OpenSSL, one part per thread), for two staging sizes."""
import json, os, sys, tempfile, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle
import tpi_b200 as tpi
size, part = (8 << 30) + 12345, 8 << 20
d = tempfile.mkdtemp(prefix="tpi_s3_", dir="/dev/shm")
path = os.path.join(d, "ckpt.bin")
try:
    chunk = np.random.default_rng(1).integers(0, 256, 64 << 20, dtype=np.uint8).tobytes()
    with open(path, "wb") as f:
        left = size
        while left:
            k = min(left, len(chunk)); f.write(chunk[:k]); left -= k
    res = {}
    for staging in (256 << 20, 512 << 20):
        with tpi.Context([0], staging_bytes=staging) as c:
            c.s3_etag_file(path, size, part)
            t0 = time.perf_counter()
            etag, n, parts = c.s3_etag_file(path, size, part)
            res[f"gpu_staging_{staging >> 20}MiB_GBps"] = round(size / (time.perf_counter() - t0) / 1e9, 2)
    buf = np.fromfile(path, dtype=np.uint8)
    eng = oracle.ENGINE_OPENSSL if oracle.lib().orc_has_openssl() else 0
    t0 = time.perf_counter()
    o_etag, o_n, o_parts = oracle.s3_etag(buf, part, threads=os.cpu_count(), engine=eng)
    res["cpu_in_memory_GBps"] = round(size / (time.perf_counter() - t0) / 1e9, 2)
    assert (etag, n) == (o_etag, o_n) and np.array_equal(parts, o_parts)
    res["etag"] = tpi.s3_etag_text(etag, n)
    print(json.dumps(res))
finally:
    import shutil; shutil.rmtree(d, ignore_errors=True)
