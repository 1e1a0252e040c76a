"""CPU-side checks of the drop-in boundary: the built library exports exactly what
include/tpimd5.h declares, pure host entry points behave, and compute calls fail loudly
(never silently fall back) when there is no GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

import oracle
import tpi_b200 as tpi
from conftest import ROOT

binding = __import__("importlib").import_module("terraform-provider-iterative_b200.binding")


def header_functions():
    src = open(os.path.join(ROOT, "include", "tpimd5.h")).read()
    return sorted(set(re.findall(r"TPIMD5_API[^;(]*?\b(tpimd5_\w+)\s*\(", src)))


def test_header_and_binding_list_the_same_symbols():
    assert header_functions() == sorted(binding.EXPORTS)


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(tpi.lib_path())
    for name in header_functions():
        assert hasattr(lib, name), name


def test_header_is_plain_c(tmp_path):
    # the header must compile as C (cgo includes it), with no C++ or CUDA types in signatures
    c = tmp_path / "t.c"
    c.write_text('#include "tpimd5.h"\nint main(void){return (int)sizeof(tpimd5_ctx*) * 0;}\n')
    import subprocess
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(c),
                           "-o", str(tmp_path / "t.o")])


def test_part_size_rule_matches_oracle(golden):
    for size, chunk, maxp, want in golden["s3_part_size"]:
        assert tpi.s3_part_size(size, chunk, maxp) == want == oracle.s3_part_size(size, chunk, maxp)
    rng = np.random.default_rng(3)
    for _ in range(200):
        size = int(rng.integers(0, 1 << 44))
        chunk = int(rng.integers(1, 64)) << 20
        assert tpi.s3_part_size(size, chunk, 10000) == oracle.s3_part_size(size, chunk, 10000)
        ps = tpi.s3_part_size(size, chunk, 10000)
        assert tpi.s3_nparts(size, ps) == oracle.s3_nparts(size, ps) <= 10000


def test_shard_lpt_is_balanced_and_total():
    rng = np.random.default_rng(8)
    lens = (rng.pareto(1.2, 5000) * 4096).astype(np.uint64) + 1
    for shards in (1, 2, 4, 8):
        assign = tpi.shard_lpt(lens, shards)
        assert assign.min() >= 0 and assign.max() < shards
        loads = np.array([lens[assign == s].sum() for s in range(shards)], dtype=np.float64)
        # greedy LPT: no shard exceeds the mean by more than the largest unit
        assert loads.max() - loads.mean() <= float(lens.max()) + 72 * len(lens)
    assert tpi.shard_lpt([], 4).size == 0


@pytest.mark.skipif(tpi.device_count() > 0, reason="only meaningful without a GPU")
def test_no_cpu_fallback_without_gpu():
    with pytest.raises(tpi.TpiMd5Error) as e:
        tpi.Context()
    assert "no CPU fallback" in str(e.value)


def test_digest_string_forms():
    import base64
    import hashlib
    for msg in (b"", b"abc", bytes(range(256))):
        d = hashlib.md5(msg).digest()
        assert tpi.md5_hex(d) == d.hex()
        assert tpi.md5_base64(d) == base64.b64encode(d).decode()
        assert tpi.md5_from_etag(d.hex()) == d and tpi.md5_from_etag('"' + d.hex() + '"') == d
        # rclone lowercases the ETag and trims quotes before matching [0-9a-f]{32} (restated from memory of v1.57.0)
        assert tpi.md5_from_etag(d.hex().upper()) == d and tpi.md5_from_etag('"' + d.hex().upper() + '"') == d
        assert tpi.md5_from_etag(d.hex()[:-1] + 'g') is None and tpi.md5_from_etag('') is None
        assert tpi.md5_from_etag(d.hex() + "-3") is None and tpi.md5_from_etag(d.hex()[:-1]) is None
        assert tpi.s3_etag_text(d, 8192) == d.hex() + "-8192"


def test_s3_plan_decision_matches_the_restated_rule():
    """tpimd5_s3_plan_parts is pure host arithmetic (no GPU): single PutObject below the cutoff, multipart at and
    above it, part size by rclone's growth rule -- against the oracle's restatement, defaults included."""
    import oracle
    cases = [(0, 0, 0, 0), (1, 0, 0, 0), ((200 << 20) - 1, 0, 0, 0), (200 << 20, 0, 0, 0), (64 << 30, 0, 0, 0), (64 << 30, 8 << 20, 0, 0),
             (5 << 40, 0, 0, 0), (1000, 64, 1, 10), (1000, 64, 1001, 10), (10000 * (5 << 20), 0, 0, 0), (10000 * (5 << 20) - 1, 0, 0, 0)]
    for size, chunk, cutoff, maxp in cases:
        multi, part, n = tpi.s3_plan_parts(size, chunk, cutoff, maxp)
        c, k, m = chunk or (5 << 20), cutoff or (200 << 20), maxp or 10000
        if size >= k:
            want_part = oracle.s3_part_size(size, c, m)
            assert (multi, part, n) == (True, want_part, oracle.s3_nparts(size, want_part)), (size, chunk, cutoff, maxp)
            assert n <= m or size // c < m
        else:
            assert (multi, part, n) == (False, size, 1)


def test_etag_rule_agrees_with_the_oracle():
    import hashlib
    import oracle
    d = hashlib.md5(b"x").hexdigest()
    for etag in (d, d.upper(), '"' + d + '"', '"' + d.upper() + '"', d + "-2", d[:-1], d[:-1] + "g", "", '""', '"' + d, d + '"'):
        assert tpi.md5_from_etag(etag) == oracle.md5_from_etag(etag), etag


def test_pathlist_keeps_its_strings_alive():
    pl = tpi.PathList(["/tmp/a", "/tmp/b/c"])
    assert len(pl) == 2 and pl.array[1] == b"/tmp/b/c"
