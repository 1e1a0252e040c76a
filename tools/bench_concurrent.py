#!/usr/bin/env python3
"""Several callers on one context and one GPU at once (Terraform runs ~10 resource operations in one plugin
process): K threads each hash their own workdir (files 1 KiB..8 MiB, ~240 MiB) through tpimd5_hash_buffers.
One such job is bound by its longest file's MD5 chain (8 MiB: ~60 ms) and uses a few percent of the SMs, so
K jobs should take about as long as one.
    python tools/bench_concurrent.py [threads]"""
import json
import os
import sys
import threading
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
import tpi_b200 as tpi  # noqa: E402


def workdir(seed, n=200, total=232 << 20):
    rng = np.random.default_rng(seed)
    sizes = np.exp(rng.uniform(np.log(1 << 10), np.log(8 << 20), n))
    sizes = np.maximum(1, (sizes * (total / sizes.sum())).astype(np.int64))
    sizes[0] = 8 << 20
    off = np.concatenate([[0], np.cumsum((sizes + 15) // 16 * 16)[:-1]]).astype(np.uint64)
    buf = oracle.synth_files(seed, 0, 1 << 20, int(off[-1] + sizes[-1] + (1 << 20)) >> 20, threads=os.cpu_count() or 4)
    return buf, off, sizes.astype(np.uint64)


def main():
    k = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    jobs = [workdir(100 + i) for i in range(k)]
    eng = oracle.ENGINE_OPENSSL if oracle.lib().orc_has_openssl() else 0
    want = [oracle.md5_batch(b, o, l, threads=os.cpu_count() or 4, engine=eng) for b, o, l in jobs]
    with tpi.Context([0]) as c:
        got = [None] * k

        def run(i):
            got[i] = c.hash_buffers(*jobs[i])
        times = []
        for rep in range(5):                              # the first rounds allocate staging buffers, tables and streams
            th = [threading.Thread(target=run, args=(i,)) for i in range(k)]
            t0 = time.perf_counter()
            for t in th:
                t.start()
            for t in th:
                t.join()
            times.append(time.perf_counter() - t0)
        allk = sorted(times[2:])[1]                       # median of the last three rounds
        ok = all(np.array_equal(g, w) for g, w in zip(got, want))
        t0 = time.perf_counter()
        for _ in range(3):
            c.hash_buffers(*jobs[0])
        one = (time.perf_counter() - t0) / 3
    print(json.dumps({"threads": k, "one_job_ms": round(1e3 * one, 1), "all_jobs_ms": round(1e3 * allk, 1),
                      "ratio": round(allk / one, 2), "serial_would_be_ms": round(1e3 * one * k, 1), "bit_exact": ok,
                      "job": "200 files, 1 KiB..8 MiB, ~240 MiB, pageable host memory, tpimd5_hash_buffers"}))
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
