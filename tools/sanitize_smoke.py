#!/usr/bin/env python3
"""Small run of every kernel for compute-sanitizer (memcheck / racecheck / synccheck), e.g.

    compute-sanitizer --tool racecheck python tools/sanitize_smoke.py

Sizes are tiny on purpose (the tools slow kernels down by one to two orders of magnitude); results are
still checked against the oracle so a silent corruption under the tool is caught too."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
import tpi_b200 as tpi  # noqa: E402
import torch  # noqa: E402


def main():
    with tpi.Context([0], staging_bytes=1 << 20) as c:
        # general kernel: aligned + unaligned, ragged, units chained across launches (1 MiB staging)
        for align in (16, 1):
            lengths = list(range(0, 140)) + [4096] * 40 + [3 << 20, 70001]
            buf, off, ln = oracle.synth_packed(81, lengths, align=align)
            assert np.array_equal(c.hash_buffers(buf, off, ln), oracle.md5_batch(buf, off, ln, threads=4))
        # tensor-map kernel: several strides incl. partial stage and ragged tail
        for stride, n in ((16, 70), (112, 40), (128, 33), (4096, 200), (4112, 37)):
            d = torch.empty(stride * n, dtype=torch.uint8, device="cuda")
            c.dev_synth_fill(d, 82, 0, 16, stride * n // 16)
            dig = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
            nbytes = stride * n - 5
            c.dev_hash_buffers(d, nbytes, n, dig, stride=stride)
            torch.cuda.synchronize()
            o = np.arange(n, dtype=np.uint64) * stride
            l = np.minimum(np.uint64(stride), np.uint64(nbytes) - o).astype(np.uint64)
            assert np.array_equal(dig.cpu().numpy(), oracle.md5_batch(d.cpu().numpy(), o, l))
        # S3 ETag (K2 + K2b), with library scratch and with the built-in array
        d = torch.empty(3 << 20, dtype=torch.uint8, device="cuda")
        c.dev_synth_fill(d, 83, 0, 1 << 20, 3)
        e1 = torch.zeros(16, dtype=torch.uint8, device="cuda")
        e2 = torch.zeros(16, dtype=torch.uint8, device="cuda")
        c.dev_s3_etag(d, (3 << 20) - 7, 65536, e1)
        tpi.dev_s3_etag_noalloc(0, d, (3 << 20) - 7, 65536, e2)
        torch.cuda.synchronize()
        want = oracle.s3_etag(d.cpu().numpy()[: (3 << 20) - 7], 65536)[0]
        assert e1.cpu().numpy().tobytes() == want == e2.cpu().numpy().tobytes()
        # K4
        rng = np.random.default_rng(84)
        old = rng.integers(0, 256, (5000, 16), dtype=np.uint8)
        new = old.copy()
        new[::97, 5] ^= 1
        assert np.array_equal(c.diff(old, new), oracle.diff(old, new))
        # streams
        s = c.stream()
        data = oracle.synth_bytes(85, 0, 200000).tobytes()
        s.write(data[:77]); s.write(data[77:])
        assert s.sum() == oracle.md5(data)
    print("sanitize_smoke ok")


if __name__ == "__main__":
    main()
