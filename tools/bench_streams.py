#!/usr/bin/env python3
"""Ad-hoc: K3 -- many hash.Hash streams advanced together (tpimd5_stream_write_many)."""
import hashlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tpi_b200 as tpi
rng = np.random.default_rng(3)
for nstreams, chunk, rounds in ((4096, 1 << 18, 4), (65536, 1 << 14, 4), (256, 4 << 20, 4)):
    with tpi.Context([0]) as c:
        streams = [c.stream() for _ in range(nstreams)]
        data = rng.integers(0, 256, nstreams * chunk, dtype=np.uint8)
        chunks = [data[i * chunk:(i + 1) * chunk - (i % 7)] for i in range(nstreams)]       # ragged ends: remainders carry over
        c.stream_write_many(streams, chunks)
        t0 = time.perf_counter()
        for _ in range(rounds):
            c.stream_write_many(streams, chunks)
        dt = time.perf_counter() - t0
        total = sum(x.size for x in chunks) * rounds
        h = hashlib.md5()
        for _ in range(rounds + 1):
            h.update(chunks[5].tobytes())
        assert streams[5].sum() == h.digest()
        print(json.dumps({"streams": nstreams, "chunk": chunk, "GBps": round(total / dt / 1e9, 2), "ms_per_round": round(dt / rounds * 1e3, 1)}))
