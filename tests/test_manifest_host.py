"""Host logic either side of the hash (csrc/manifest.cc), checked on CPU: rule assembly and glob
semantics against the oracle's restatement and the reference's own TestTransferExcludes
expectations (storage_test.go:55-101); manifest persistence; the digest-free branches of the
transfer plan."""
import os
import random
import subprocess

import numpy as np
import pytest

import oracle
import tpi_b200 as tpi


def test_rule_assembly_matches_go_logic():
    assert tpi.Filter().rules() == oracle.transfer_rules(None)
    ex = ["**.txt", "/a.txt", "a.txt", "dir/../b", "./c/", "+ /keep/**", "- *.tmp"]
    assert tpi.Filter(ex).rules() == oracle.transfer_rules(ex)
    assert tpi.Filter(["cache"], "out").rules() == oracle.limit_transfer("out", oracle.transfer_rules(["cache"]))
    assert tpi.Filter(["cache"], ".").rules() == oracle.transfer_rules(["cache"])
    assert tpi.Filter([], "a/b/").rules()[-3:] == ["+ /a/b", "+ /a/b/**", "- /**"]


def test_reference_transfer_excludes_cases(golden):
    te = golden["transfer_excludes"]
    for case in te["cases"]:
        f = tpi.Filter(case["exclude"])
        assert sorted(p for p in te["listing"] if f.include(p)) == sorted(case["expect"]), case


def test_glob_semantics_against_oracle_randomised():
    rnd = random.Random(57)
    segs = ["a", "b", "main.tf", "x.txt", "data", ".terraform", "terraform.tfstate", "out", "a.b", "c-d", "1", "ab"]
    paths = ["/".join(rnd.choice(segs) for _ in range(rnd.randint(1, 4))) for _ in range(400)]
    globs = ["*.txt", "**.txt", "/a/**", "a/*", "/*/b", "data/", "?", "/??", "[ab]", "/[!a]*", "{a,b}/x.txt", "/{a,data}/**",
             "*.{txt,tf}", "a?b", "**/out/**", "out", "/out", "c-d", "*a*", "/**/1", "[a-c]*", "x.t?t", "**"]
    for _ in range(60):
        rules = []
        for _ in range(rnd.randint(0, 4)):
            g = rnd.choice(globs)
            rules.append(rnd.choice(["+ ", "- ", ""]) + g)
        limit = rnd.choice([None, None, "out", "data/a"])
        f = tpi.Filter(rules, limit)
        o_rules = oracle.transfer_rules(rules)
        if limit:
            o_rules = oracle.limit_transfer(limit, o_rules)
        o = oracle.FilterSet(o_rules)
        assert f.rules() == o_rules
        for p in paths:
            assert f.include(p) == o.include_file(p), (rules, limit, p)


def test_bad_globs_are_rejected():
    for bad in ["- [abc", "- {a,b", "- a}b", "- "]:
        with pytest.raises(tpi.TpiMd5Error):
            tpi.Filter([bad])


def test_manifest_roundtrip_and_md5sum_lines(tmp_path):
    paths = ["b/z.bin", "a.txt", "dir with space/x", "b/a"]
    sizes = [10, 0, 5, 7]
    mt = [111, 222, 333, 444]
    md = np.stack([np.frombuffer(oracle.md5(p.encode()), dtype=np.uint8) for p in paths])
    m = tpi.Manifest.from_arrays(paths, sizes, mt, md)
    assert m.paths() == sorted(paths)                              # kept sorted by path
    want = {p: (s, t, d.tobytes()) for p, s, t, d in zip(paths, sizes, mt, md)}
    for p, s, t, d in m.entries():
        assert want[p] == (s, t, d)
    f = str(tmp_path / "m.tpim")
    m.save(f)
    assert tpi.Manifest.load(f).entries() == m.entries()
    txt = str(tmp_path / "m.md5")
    m.write_md5sum(txt)
    assert open(txt).read() == "".join(oracle.md5sum_line(want[p][2], p) for p in sorted(paths))
    with pytest.raises(tpi.TpiMd5Error):
        tpi.Manifest.from_arrays(["x", "/x"], [1, 1])              # duplicate after normalisation
    (tmp_path / "junk").write_bytes(b"not a manifest")
    with pytest.raises(tpi.TpiMd5Error):
        tpi.Manifest.load(str(tmp_path / "junk"))


def test_md5sum_text_is_what_coreutils_prints(tmp_path):
    names = ["plain", "with space", "back\\slash"]
    digs = []
    for i, n in enumerate(names):
        b = oracle.synth_bytes(8, i, 100 + i).tobytes()
        (tmp_path / n).write_bytes(b)
        digs.append(np.frombuffer(oracle.md5(b), dtype=np.uint8))
    m = tpi.Manifest.from_arrays(names, [100, 101, 102], None, np.stack(digs))
    out = str(tmp_path / "sum.txt")
    m.write_md5sum(out)
    got = subprocess.check_output(["md5sum"] + sorted(names), cwd=tmp_path).decode()
    assert open(out).read() == got


def test_plan_branches_that_need_no_digest_comparison():
    src = tpi.Manifest.from_arrays(["a", "b", "c", "d"], [1, 2, 3, 4], [10, 10, 10, 10])
    dst = tpi.Manifest.from_arrays(["a", "b", "c", "zz"], [1, 9, 3, 4], [10, 10, 99, 10])
    act, nt, nd = tpi.plan_transfer(None, src, dst, tpi.manifest.MODE_SIZE_ONLY)
    assert act.tolist() == [0, 2, 0, 1] and nt == 2 and nd == 1
    # default mode without hashes: equal mtime -> skip, different mtime -> transfer ("nothing vouches")
    act, nt, nd = tpi.plan_transfer(None, src, dst, tpi.manifest.MODE_MODTIME)
    assert act.tolist() == [0, 2, 2, 1]
    # --checksum with no hashes on either side: equal by size
    act, _, _ = tpi.plan_transfer(None, src, dst, tpi.manifest.MODE_CHECKSUM)
    assert act.tolist() == [0, 2, 0, 1]
    # the same decisions from the oracle's equal()
    for mode in (0, 1, 2):
        act, _, _ = tpi.plan_transfer(None, src, dst, mode)
        d = {p: (s, t) for p, s, t, _ in dst.entries()}
        for (p, s, t, _), a in zip(src.entries(), act.tolist()):
            if p not in d:
                assert a == 1
            else:
                eq, fix = oracle.equal(s, d[p][0], t, d[p][1], None, None, mode=mode)
                assert (a == 0) == eq and not fix


def test_plan_refuses_to_compare_digests_without_a_gpu():
    h = np.zeros((1, 16), np.uint8)
    src = tpi.Manifest.from_arrays(["a"], [1], [1], h)
    dst = tpi.Manifest.from_arrays(["a"], [1], [2], h)
    with pytest.raises(tpi.TpiMd5Error) as e:
        tpi.plan_transfer(None, src, dst, tpi.manifest.MODE_CHECKSUM)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_directory_rules_against_oracle_randomised():
    """rclone decides directories by rules derived from the file rules (globs with "**", trailing "/",
    parents of include rules).  The C++ filter and the oracle's restatement must agree on every
    directory, and a nested DirectoryOut must stay reachable."""
    rnd = random.Random(58)
    segs = ["a", "b", "data", "out", "deep", ".terraform", "x.d", "1"]
    dirs = ["/".join(rnd.choice(segs) for _ in range(rnd.randint(1, 4))) for _ in range(300)]
    globs = ["*.txt", "**.txt", "/a/**", "a/*", "/*/b", "data/", "/out/", "out/deep/**", "/a/b/c.bin", "*", "**", "/**", "x.d/**",
             "{a,b}/**", "/data/deep/", "b/1"]
    for _ in range(80):
        rules = [rnd.choice(["+ ", "- ", ""]) + rnd.choice(globs) for _ in range(rnd.randint(0, 4))]
        limit = rnd.choice([None, "out", "data/deep", "a/b/1"])
        f = tpi.Filter(rules, limit)
        o_rules = oracle.transfer_rules(rules)
        if limit:
            o_rules = oracle.limit_transfer(limit, o_rules)
        o = oracle.FilterSet(o_rules)
        for d in dirs:
            assert f.include_dir(d) == o.include_dir(d), (rules, limit, d)
    f = tpi.Filter(["cache"], "data/deep")                    # Pull with a nested output directory
    assert f.include_dir("data") and f.include_dir("data/deep") and f.include_dir("data/deep/er")
    assert not f.include_dir("cache") and not f.include_dir("data/other") and not f.include_dir("out")
    assert f.include("data/deep/file") and not f.include("data/file")
    f = tpi.Filter([])                                        # built-ins: "/.terraform**" prunes the directory itself
    assert not f.include_dir(".terraform") and not f.include_dir(".terraform/providers") and f.include_dir("src/.terraform")
