"""Pins the CPU oracle (oracle/) before anything is compared against it.

Sources of truth, all independent of the oracle's own C code: the committed golden vectors
(RFC 1321 A.5, SURVEY.md appendix B, hashlib-derived digests of synthetic files), hashlib at
random lengths, OpenSSL's MD5 through libcrypto, coreutils md5sum, and for the file-set rules
the expectations of /root/reference/task/common/machine/storage_test.go:55-101.
"""
import hashlib
import os
import shutil
import subprocess

import numpy as np
import pytest

import oracle


def pat(n):
    return bytes(i % 251 for i in range(n))


def test_rfc1321_vectors(golden):
    for s, h in golden["rfc1321"]:
        assert oracle.md5_hex(s.encode()) == h


def test_pattern_vectors(golden):
    for n, h in golden["pattern251"].items():
        assert oracle.md5_hex(pat(int(n))) == h, n


def test_synth_generator_pinned(golden):
    for seed, idx, off, hx in golden["synth_prefix"]:
        assert oracle.synth_bytes(seed, idx, 40, off).tobytes().hex() == hx
    # any window of a file equals the same window of the whole file
    whole = oracle.synth_bytes(9, 3, 1000)
    for off, n in ((0, 1), (1, 7), (7, 9), (8, 8), (13, 500), (999, 1)):
        assert np.array_equal(oracle.synth_bytes(9, 3, n, off), whole[off:off + n])


def test_synth_md5_vectors(golden):
    for seed, idx, n, h in golden["synth_md5"]:
        assert oracle.md5_hex(oracle.synth_bytes(seed, idx, n)) == h, n


def test_against_hashlib_random_lengths():
    rng = np.random.default_rng(1321)
    for n in list(range(0, 200)) + [int(x) for x in rng.integers(200, 300000, 40)]:
        b = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert oracle.md5(b) == hashlib.md5(b).digest(), n


def test_streaming_matches_one_shot_and_sum_is_non_destructive():
    rng = np.random.default_rng(7)
    data = rng.integers(0, 256, 10_000, dtype=np.uint8).tobytes()
    s = oracle.MD5Stream()
    pos = 0
    for step in (1, 63, 64, 65, 7, 128, 1000, 3, 5000, 10_000):
        s.write(data[pos:pos + step])
        pos = min(len(data), pos + step)
        assert s.sum() == hashlib.md5(data[:pos]).digest()
        assert s.sum() == hashlib.md5(data[:pos]).digest()      # Sum twice: unchanged
    s.reset()
    assert s.sum().hex() == "d41d8cd98f00b204e9800998ecf8427e"


def test_bit_length_beyond_32_bits():
    # 512 MiB + 3 bytes: the bit count needs 33 bits.  Streamed through a 1 MiB zero buffer.
    n = (512 << 20) + 3
    chunk = bytes(1 << 20)
    s, h = oracle.MD5Stream(), hashlib.md5()
    left = n
    while left:
        k = min(left, len(chunk))
        s.write(chunk[:k]); h.update(chunk[:k]); left -= k
    assert s.sum() == h.digest()


def test_batch_and_openssl_engine_agree():
    lengths = [0, 1, 55, 56, 64, 100, 4096, 4096, 70000, 3]
    buf, off, ln = oracle.synth_packed(11, lengths, align=16)
    d0 = oracle.md5_batch(buf, off, ln, threads=1)
    d3 = oracle.md5_batch(buf, off, ln, threads=3)
    assert np.array_equal(d0, d3)
    for i, (o, l) in enumerate(zip(off.tolist(), ln.tolist())):
        assert d0[i].tobytes() == hashlib.md5(buf[o:o + l].tobytes()).digest()
    if oracle.lib().orc_has_openssl():
        d1 = oracle.md5_batch(buf, off, ln, threads=2, engine=oracle.ENGINE_OPENSSL)
        assert np.array_equal(d0, d1)


@pytest.mark.skipif(shutil.which("md5sum") is None, reason="coreutils md5sum absent")
def test_against_md5sum_and_manifest_line(tmp_path):
    names = ["a.bin", "dir with space.bin", "x"]
    lines = []
    for i, nm in enumerate(names):
        b = oracle.synth_bytes(4, i, 1000 * i + 5).tobytes()
        (tmp_path / nm).write_bytes(b)
        lines.append(oracle.md5sum_line(oracle.md5(b), nm))
    got = subprocess.check_output(["md5sum"] + names, cwd=tmp_path).decode()
    assert got == "".join(lines)


def test_s3_etag_vectors(golden):
    ex = golden["s3_example"]
    data = pat(5 << 20) * 2 + b"tail"
    etag, n, parts = oracle.s3_etag(data, ex["part_size"])
    assert oracle.s3_etag_text(etag, n) == ex["etag"]
    assert oracle.md5_hex(data) == ex["whole"]
    assert parts.shape == (3, 16)
    for seed, idx, size, ps, want in golden["synth_s3"]:
        etag, n, _ = oracle.s3_etag(oracle.synth_bytes(seed, idx, size), ps, threads=2)
        assert oracle.s3_etag_text(etag, n) == want


def test_s3_part_size_rule(golden):
    for size, chunk, maxp, want in golden["s3_part_size"]:
        assert oracle.s3_part_size(size, chunk, maxp) == want
    # exact multiple => no empty trailing part
    assert oracle.s3_nparts(4 * 65536, 65536) == 4
    assert oracle.s3_nparts(4 * 65536 + 1, 65536) == 5
    assert oracle.s3_nparts(0, 65536) == 1


def test_diff():
    rng = np.random.default_rng(5)
    old = rng.integers(0, 256, (1000, 16), dtype=np.uint8)
    new = old.copy()
    changed = sorted(rng.choice(1000, 50, replace=False).tolist())
    for i in changed:
        new[i, rng.integers(0, 16)] ^= 1 << int(rng.integers(0, 8))
    assert oracle.diff(old, new).tolist() == changed
    assert oracle.diff(old, old).size == 0


def test_equal_decision_table():
    a, b = bytes(range(16)), bytes(range(1, 17))
    E = oracle.equal
    assert E(1, 2, 0, 0, a, a) == (False, False)                                   # size differs
    assert E(1, 1, 5, 5, a, b) == (True, False)                                    # mtime equal wins (default)
    assert E(1, 1, 5, 9, a, a) == (True, True)                                     # hash equal: fix mtime
    assert E(1, 1, 5, 9, a, b) == (False, False)
    assert E(1, 1, 5, 9, a, None) == (False, False)                                # no hash to vouch
    assert E(1, 1, 5, 9, a, b, mode=oracle.MODE_CHECKSUM) == (False, False)
    assert E(1, 1, 5, 5, a, b, mode=oracle.MODE_CHECKSUM) == (False, False)        # mtime ignored
    assert E(1, 1, 5, 9, a, None, mode=oracle.MODE_CHECKSUM) == (True, False)      # unavailable => equal by size
    assert E(1, 1, 5, 9, a, b, mode=oracle.MODE_SIZE_ONLY) == (True, False)


def test_transfer_excludes_match_reference_test(golden):
    te = golden["transfer_excludes"]
    for case in te["cases"]:
        assert oracle.list_transfer(te["listing"], case["exclude"]) == sorted(case["expect"]), case


def test_filter_details():
    fs = oracle.FilterSet(oracle.transfer_rules(None))
    assert not fs.include_file("main.tf")
    assert fs.include_file("sub/main.tf")                      # anchored at the root only
    assert not fs.include_file("terraform.tfstate.backup")
    assert not fs.include_file(".terraform/providers/x")
    assert not fs.include_file(".terraform.lock.hcl")          # "/.terraform**"
    lim = oracle.limit_transfer("out", oracle.transfer_rules(["cache"]))
    assert lim[-3:] == ["+ /out", "+ /out/**", "- /**"]
    fs = oracle.FilterSet(lim)
    assert fs.include_file("out/file") and fs.include_file("out/deep/er")
    assert not fs.include_file("cache/file") and not fs.include_file("other")
    assert oracle.limit_transfer(".", ["- /x"]) == ["- /x"]


def test_dir_globs_derivation():
    """The directory globs rclone derives from a file glob (fs/filter globToDirGlobs), on worked examples."""
    g = oracle.glob_to_dir_globs
    assert g("/a/b/**") == ["/a/b/**/", "/a/b/", "/a/", "/"]
    assert g("/out") == ["/"]
    assert g("/a/b/c.txt") == ["/a/b/", "/a/", "/"]
    assert g("x.txt") == ["/**"]                       # could be anywhere: every directory stays in
    assert g("a/b.txt") == ["a/"]
    fs = oracle.FilterSet(oracle.limit_transfer("a/b", oracle.transfer_rules([])))
    assert fs.include_dir("a") and fs.include_dir("a/b") and fs.include_dir("a/b/c") and not fs.include_dir("z")
    assert fs.include_path("a/b/f") and not fs.include_path("a/f") and not fs.include_path("z/a/b/f")
