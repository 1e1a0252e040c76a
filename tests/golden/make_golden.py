#!/usr/bin/env python3
"""Regenerates tests/golden/vectors.json.

The reference (Go + un-vendored rclone v1.57.0) cannot run on this image, so the digests here
come from an implementation independent of both the oracle and the CUDA path: CPython's hashlib
(OpenSSL).  Inputs are either literal strings (RFC 1321 A.5), the `i mod 251` pattern of
SURVEY.md appendix B, or the counter-PRNG workdir bytes, whose generator is itself pinned by
the `synth_prefix` entries.  The file-set cases restate the expectations of
/root/reference/task/common/machine/storage_test.go:55-101 (TestTransferExcludes).

Run from the repo root:  python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import oracle  # noqa: E402  (only used for the synthetic byte generator)


def pat(n):
    return bytes(i % 251 for i in range(n))


def main():
    v = {}
    v["rfc1321"] = [
        ["", "d41d8cd98f00b204e9800998ecf8427e"],
        ["a", "0cc175b9c0f1b6a831c399e269772661"],
        ["abc", "900150983cd24fb0d6963f7d28e17f72"],
        ["message digest", "f96b697d7cb7938d525a2f31aaf161d0"],
        ["abcdefghijklmnopqrstuvwxyz", "c3fcd3d76192e4007dfb496cca67e13b"],
        ["ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789", "d174ab98d277d9f5a5611c2c9f419d9f"],
        ["1234567890" * 8, "57edf4a22be3c955ac49da2e2107b67a"],
    ]
    for s, h in v["rfc1321"]:
        assert hashlib.md5(s.encode()).hexdigest() == h
    v["pattern251"] = {str(n): hashlib.md5(pat(n)).hexdigest()
                       for n in (0, 1, 55, 56, 57, 63, 64, 65, 119, 120, 121, 127, 128, 129, 4095, 4096, 4097)}
    # SURVEY appendix B multipart example
    data = pat(5 << 20) * 2 + b"tail"
    parts = [data[i:i + (5 << 20)] for i in range(0, len(data), 5 << 20)]
    dd = b"".join(hashlib.md5(p).digest() for p in parts)
    v["s3_example"] = {"part_size": 5 << 20, "etag": hashlib.md5(dd).hexdigest() + "-%d" % len(parts),
                       "whole": hashlib.md5(data).hexdigest()}
    assert v["s3_example"]["etag"] == "63a84bdfe2824bf190d348ef62e9ce73-3"
    assert v["s3_example"]["whole"] == "28761385735a63bdc3670a0661af302d"
    # synthetic generator pin + digests at edge lengths
    v["synth_prefix"] = [[seed, idx, off, oracle.synth_bytes(seed, idx, 40, off).tobytes().hex()]
                         for seed, idx, off in ((0, 0, 0), (2, 7, 0), (2, 7, 5), (5, 123456789, 4093))]
    edge = [0, 1, 3, 4, 15, 16, 17, 54, 55, 56, 57, 62, 63, 64, 65, 111, 119, 120, 127, 128, 129, 183, 184, 191,
            192, 255, 256, 257, 511, 512, 1000, 1023, 1024, 4095, 4096, 4097, 65535, 65536, 65537, 1 << 20]
    v["synth_md5"] = [[2, i, n, hashlib.md5(oracle.synth_bytes(2, i, n).tobytes()).hexdigest()]
                      for i, n in enumerate(edge)]
    # S3 ETags of synthetic files: ragged last part, exact multiple, single part, empty
    s3 = []
    for idx, (n, ps) in enumerate(((3 * 65536 + 17, 65536), (4 * 65536, 65536), (100, 65536), (0, 65536),
                                   ((5 << 20) + 1, 5 << 20))):
        b = oracle.synth_bytes(3, idx, n).tobytes()
        pp = [b[i:i + ps] for i in range(0, n, ps)] or [b""]
        d = b"".join(hashlib.md5(p).digest() for p in pp)
        s3.append([3, idx, n, ps, hashlib.md5(d).hexdigest() + "-%d" % len(pp)])
    v["synth_s3"] = s3
    # rclone part-size rule: [size, chunk, max_parts, expected]
    MiB, GiB = 1 << 20, 1 << 30
    v["s3_part_size"] = [
        [200 * MiB, 5 * MiB, 10000, 5 * MiB],
        [64 * GiB, 5 * MiB, 10000, 7 * MiB],
        [64 * GiB, 8 * MiB, 10000, 8 * MiB],
        [5 * MiB * 10000, 5 * MiB, 10000, 6 * MiB],
        [5 * MiB * 10000 - 1, 5 * MiB, 10000, 5 * MiB],
        [5 * 1024 * GiB, 5 * MiB, 10000, 525 * MiB],
    ]
    # storage_test.go:55-101 -- listing of testdata/transferTest and expected transferred FILES
    listing = ["a.txt", "main.tf", "temp/a.txt", "temp/b.txt"]
    v["transfer_excludes"] = {"listing": listing, "cases": [
        {"exclude": [], "expect": ["a.txt", "temp/a.txt", "temp/b.txt"]},
        {"exclude": ["**.txt"], "expect": []},
        {"exclude": ["/a.txt"], "expect": ["temp/a.txt", "temp/b.txt"]},
        {"exclude": ["a.txt"], "expect": ["temp/a.txt", "temp/b.txt"]},
    ]}
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "vectors.json")
    with open(out, "w") as f:
        json.dump(v, f, indent=1, sort_keys=True)
    print("wrote", out)


if __name__ == "__main__":
    main()
