"""GPU tests of the callers either side of the hash: manifest scan, transfer plan, and
machine.Transfer for local directories -- the latter written like the reference's own
TestTransferExcludes (/root/reference/task/common/machine/storage_test.go:55-101)."""
import hashlib
import os

import numpy as np
import pytest

import oracle
import tpi_b200 as tpi
from tpi_b200 import machine

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = tpi.Context([0], staging_bytes=16 << 20)
    yield c
    c.close()


def list_dir(root):
    """The reference test's listDir: every file and directory below root, as "/relative/path"."""
    out = []
    for d, dirs, files in os.walk(root):
        for n in dirs + files:
            out.append("/" + os.path.relpath(os.path.join(d, n), root).replace(os.sep, "/"))
    return sorted(out)


def make_transfer_test_tree(root):
    """testdata/transferTest of the reference: four empty files."""
    os.makedirs(os.path.join(root, "temp"))
    for p in ("a.txt", "main.tf", "temp/a.txt", "temp/b.txt"):
        open(os.path.join(root, p), "wb").close()


@pytest.mark.parametrize("exclude,expect", [
    (None, ["/a.txt", "/temp", "/temp/a.txt", "/temp/b.txt"]),           # built-in rules drop main.tf
    (["**.txt"], ["/temp"]),                                             # directory still gets transferred
    (["/a.txt"], ["/temp", "/temp/a.txt", "/temp/b.txt"]),               # explicitly anchored
    (["a.txt"], ["/temp", "/temp/a.txt", "/temp/b.txt"]),                # implicitly anchored
])
def test_transfer_excludes(ctx, tmp_path, exclude, expect):
    src, dst = str(tmp_path / "transferTest"), str(tmp_path / "dst")
    make_transfer_test_tree(src)
    machine.Transfer(ctx, src, dst, exclude)
    assert list_dir(dst) == sorted(expect)


def build_workdir(root, seed, n=60):
    rng = np.random.default_rng(seed)
    files = {}
    for i in range(n):
        d = ["", "data", "data/deep", "out", "cache", ".terraform/providers"][int(rng.integers(0, 6))]
        name = f"f{i}" + [".bin", ".txt", ".ckpt"][int(rng.integers(0, 3))]
        rel = (d + "/" if d else "") + name
        size = int(rng.choice([0, 1, 55, 64, 4096, 70000, 3 << 20]))
        os.makedirs(os.path.join(root, d), exist_ok=True)
        b = oracle.synth_bytes(seed, i, size).tobytes()
        with open(os.path.join(root, rel), "wb") as f:
            f.write(b)
        files[rel] = b
    for rel in ("main.tf", "terraform.tfstate", "terraform.tfstate.backup"):
        with open(os.path.join(root, rel), "wb") as f:
            f.write(b"x")
    os.symlink("main.tf", os.path.join(root, "link"))                       # rclone's local backend skips symlinks
    return files


def test_manifest_scan_is_the_filtered_listing_with_oracle_digests(ctx, tmp_path):
    root = str(tmp_path / "w")
    files = build_workdir(root, 61)
    flt = tpi.Filter(["cache", "**.txt"])
    m = tpi.Manifest.scan(ctx, root, flt)
    o = oracle.FilterSet(oracle.transfer_rules(["cache", "**.txt"]))
    want = sorted(p for p in files if o.include_path(p))
    assert m.paths() == want and "link" not in m.paths()
    for p, size, mtime, md5 in m.entries():
        assert size == len(files[p]) and md5 == oracle.md5(files[p])
        assert mtime == os.stat(os.path.join(root, p)).st_mtime_ns
    out = str(tmp_path / "manifest.md5")
    m.write_md5sum(out)
    assert open(out).read() == "".join(oracle.md5sum_line(oracle.md5(files[p]), p) for p in want)
    m.save(str(tmp_path / "m.tpim"))
    assert tpi.Manifest.load(str(tmp_path / "m.tpim")).entries() == m.entries()


def test_plan_transfer_follows_rclone_equal(ctx):
    rng = np.random.default_rng(62)
    n = 3000
    paths = [f"d{i % 17}/f{i}" for i in range(n)]
    sizes = rng.integers(0, 1000, n)
    mt = rng.integers(0, 5, n)
    md = rng.integers(0, 256, (n, 16), dtype=np.uint8)
    have_s = rng.random(n) < 0.9
    dsizes, dmt, dmd = sizes.copy(), mt.copy(), md.copy()
    have_d = rng.random(n) < 0.9
    for arr, frac in ((dsizes, 0.1), (dmt, 0.4)):
        k = rng.random(n) < frac
        arr[k] += 1
    flip = rng.random(n) < 0.2
    dmd[flip, 0] ^= 0xFF
    present = rng.random(n) < 0.85
    src = tpi.Manifest.from_arrays(paths, sizes, mt, md, have_s.astype(np.uint8))
    keep = np.nonzero(present)[0]
    dst = tpi.Manifest.from_arrays([paths[i] for i in keep] + ["zz/only-at-dst"], np.append(dsizes[keep], 5),
                                   np.append(dmt[keep], 0), np.vstack([dmd[keep], np.zeros((1, 16), np.uint8)]),
                                   np.append(have_d[keep], 1).astype(np.uint8))
    order = {p: i for i, p in enumerate(paths)}
    for mode in (tpi.manifest.MODE_MODTIME, tpi.manifest.MODE_CHECKSUM, tpi.manifest.MODE_SIZE_ONLY):
        act, nt, nd = tpi.plan_transfer(ctx, src, dst, mode)
        assert nd == 1
        for (p, *_), a in zip(src.entries(), act.tolist()):
            i = order[p]
            if not present[i]:
                assert a == tpi.manifest.ACT_TRANSFER_NEW
                continue
            eq, fix = oracle.equal(int(sizes[i]), int(dsizes[i]), int(mt[i]), int(dmt[i]),
                                   md[i].tobytes() if have_s[i] else None, dmd[i].tobytes() if have_d[i] else None, mode=mode)
            want = tpi.manifest.ACT_FIX_MTIME if (eq and fix) else (tpi.manifest.ACT_SKIP if eq else tpi.manifest.ACT_TRANSFER)
            assert a == want, (mode, p)
        assert nt == int(((act == 1) | (act == 2)).sum())


def test_transfer_then_resync_with_checksum_finds_exactly_the_changed_files(ctx, tmp_path):
    """BASELINE config 1 in miniature: sync a workdir to a local "remote", then mutate files without
    touching size or mtime and resync with --checksum: only those move; a default-mode sync misses
    them (as rclone's would), a resync after nothing changed moves nothing."""
    src, dst = str(tmp_path / "src"), str(tmp_path / "dst")
    files = build_workdir(src, 63, n=80)
    st = machine.Transfer(ctx, src, dst, ["cache"])
    o = oracle.FilterSet(oracle.transfer_rules(["cache"]))
    moved = sorted(p for p in files if o.include_path(p))
    assert st["transferred"] == len(moved) == st["files_listed"] and st["bytes_listed"] == sum(len(files[p]) for p in moved)
    for p in moved:
        assert open(os.path.join(dst, p), "rb").read() == files[p]
        assert os.stat(os.path.join(dst, p)).st_mtime_ns == os.stat(os.path.join(src, p)).st_mtime_ns
    # "cache" becomes "- /cache": that is a FILE named cache at the root (rclone glob semantics), so the
    # directory cache/ still travels -- the oracle's filter agrees (it produced `moved`)
    assert not os.path.exists(os.path.join(dst, "main.tf")) and not os.path.exists(os.path.join(dst, "terraform.tfstate"))
    assert machine.Transfer(ctx, src, dst, ["cache"])["transferred"] == 0                      # unchanged: nothing moves
    victims = [p for p in moved if len(files[p]) > 0][::7]
    for p in victims:
        full = os.path.join(src, p)
        stt = os.stat(full)
        b = bytearray(files[p]); b[len(b) // 2] ^= 1
        with open(full, "wb") as f:
            f.write(bytes(b))
        os.utime(full, ns=(stt.st_atime_ns, stt.st_mtime_ns))                                   # same size, same mtime
        files[p] = bytes(b)
    assert machine.Transfer(ctx, src, dst, ["cache"])["transferred"] == 0                      # size+mtime cannot see it
    st = machine.Transfer(ctx, src, dst, ["cache"], checksum=True)
    assert st["transferred"] == len(victims) and st["skipped"] == len(moved) - len(victims)
    for p in moved:
        assert open(os.path.join(dst, p), "rb").read() == files[p]
    # mtime-only change: default mode hashes the pair and fixes the destination mtime instead of copying
    p = moved[0]
    os.utime(os.path.join(src, p), ns=(1_000_000_000, 1_234_567_890_000_000_000))
    st = machine.Transfer(ctx, src, dst, ["cache"])
    assert st["transferred"] == 0 and st["mtime_fixed"] == 1
    assert os.stat(os.path.join(dst, p)).st_mtime_ns == 1_234_567_890_000_000_000


def test_pull_limited_to_the_output_directory(ctx, tmp_path):
    """Pull: Transfer(remote/data, Directory, LimitTransfer(DirectoryOut, ExcludeList))
    (/root/reference/task/aws/task.go:306-313): only out/ comes back, as the smoke test asserts."""
    src, dst = str(tmp_path / "bucket-data"), str(tmp_path / "workdir")
    files = build_workdir(src, 64, n=80)
    machine.Transfer(ctx, src, dst, machine.LimitTransfer("out", ["cache"]))
    got = [p[1:] for p in list_dir(dst) if os.path.isfile(os.path.join(dst, p[1:]))]
    assert got == sorted(p for p in files if p.startswith("out/")) and got
    assert not os.path.exists(os.path.join(dst, "cache")) and not os.path.exists(os.path.join(dst, "data"))


def test_pull_limited_to_a_nested_output_directory(ctx, tmp_path):
    """DirectoryOut may be nested: "+ /data/deep", "+ /data/deep/**", "- /**".  rclone still descends through
    data/ (directory rules derived from the include rules) but brings back nothing else from it."""
    src, dst = str(tmp_path / "bucket-data"), str(tmp_path / "workdir")
    files = build_workdir(src, 66, n=120)
    machine.Transfer(ctx, src, dst, machine.LimitTransfer("data/deep", []))
    got = [p[1:] for p in list_dir(dst) if os.path.isfile(os.path.join(dst, p[1:]))]
    want = sorted(p for p in files if p.startswith("data/deep/"))
    assert got == want and got
    assert sorted(os.listdir(dst)) == ["data"] and sorted(os.listdir(os.path.join(dst, "data"))) == ["deep"]
    o = oracle.FilterSet(oracle.limit_transfer("data/deep", oracle.transfer_rules([])))
    assert want == sorted(p for p in files if o.include_path(p))


def test_transfer_refuses_cloud_remotes(ctx, tmp_path):
    with pytest.raises(tpi.TpiMd5Error):
        machine.Transfer(ctx, str(tmp_path), ":s3,provider=AWS:bucket/data")


CLI = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "terraform-provider-iterative_b200", "tpimd5sum")


@pytest.mark.skipif(not os.path.exists(CLI), reason="tpimd5sum not built")
def test_cli_prints_what_md5sum_prints_and_detects_changes(tmp_path):
    """The VM script's change detectors (machine-script.sh.tpl:111,119) answered from the GPU."""
    import subprocess
    root = str(tmp_path / "logs")
    files = build_workdir(root, 65, n=40)
    names = sorted(p for p in files if "/" not in p)                        # like `md5sum "$DIR"/*`
    got = subprocess.run([CLI] + names, cwd=root, capture_output=True, text=True)
    assert got.returncode == 0, got.stderr
    assert got.stdout == subprocess.check_output(["md5sum"] + names, cwd=root, text=True)
    man = str(tmp_path / "old.tpim")
    out = subprocess.run([CLI, "--manifest", root, "-x", "**.txt", "-o", man], capture_output=True, text=True)
    o = oracle.FilterSet(oracle.transfer_rules(["**.txt"]))
    keep = sorted(p for p in files if o.include_path(p))
    assert out.returncode == 0 and out.stdout == "".join(oracle.md5sum_line(oracle.md5(files[p]), p) for p in keep)
    # nothing changed -> exit 0 and no lines
    r = subprocess.run([CLI, "--changed", man, root, "-x", "**.txt"], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout == ""
    # one byte flipped with size and mtime restored, one new file -> exactly those two, exit 1
    victim = next(p for p in keep if len(files[p]) > 10)
    full = os.path.join(root, victim)
    st = os.stat(full)
    b = bytearray(files[victim]); b[3] ^= 0x80
    open(full, "wb").write(bytes(b))
    os.utime(full, ns=(st.st_atime_ns, st.st_mtime_ns))
    open(os.path.join(root, "zz_new.bin"), "wb").write(b"new")
    r = subprocess.run([CLI, "--changed", man, root, "-x", "**.txt"], capture_output=True, text=True)
    assert r.returncode == 1 and sorted(r.stdout.splitlines()) == sorted([f"* {victim}", "+ zz_new.bin"])


@pytest.mark.skipif(not os.path.exists(CLI), reason="tpimd5sum not built")
def test_cli_changed_without_an_old_manifest_is_the_loops_first_turn(tmp_path):
    """The patched VM loop (integration/machine-script.sh.tpl.patch) calls `tpimd5sum --changed OLD DIR -o OLD`
    every 10 s; on the first turn OLD does not exist: everything is new (exit 1, sync runs) and the manifest is
    written, so the second turn over an unchanged tree exits 0."""
    import subprocess
    root = str(tmp_path / "data")
    files = build_workdir(root, 66, n=12)
    man = str(tmp_path / "state.tpim")
    r = subprocess.run([CLI, "--changed", man, root, "-o", man], capture_output=True, text=True)
    assert r.returncode == 1 and os.path.exists(man), r.stderr
    o = oracle.FilterSet(oracle.transfer_rules([]))
    assert sorted(r.stdout.splitlines()) == sorted("+ " + p for p in files if o.include_path(p))
    r = subprocess.run([CLI, "--changed", man, root, "-o", man], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout == ""


def test_s3_upload_plan_matches_the_restated_rclone_rule(ctx, tmp_path):
    """tpimd5_s3_upload_plan against the oracle's restatement of rclone v1.57.0 backend/s3 Object.Update: single
    PutObject below the cutoff (Content-MD5 = base64 MD5, ETag = hex MD5), multipart at and above it (part size
    by the growth rule, per-part Content-MD5, Md5chksum, ETag hex(md5(md5s))-N), and --s3-disable-checksum."""
    import base64
    cases = [  # (size, chunk_size, upload_cutoff, max_parts)
        (0, 0, 0, 0), (1, 0, 0, 0), (3 << 20, 1 << 20, 3 << 20, 0), ((3 << 20) - 1, 1 << 20, 3 << 20, 0),
        ((5 << 20) + 77, 1 << 20, 1 << 20, 3), (40 * 4096 + 5, 4096, 8192, 10000), (1 << 20, 64, 1, 100),
    ]
    for k, (size, chunk, cutoff, maxp) in enumerate(cases):
        data = oracle.synth_bytes(70 + k, 0, size)
        path = str(tmp_path / f"obj{k}.bin")
        data.tofile(path)
        got = ctx.s3_upload_plan(path, size, chunk, cutoff, maxp)
        want = oracle.s3_upload_plan(data, chunk or (5 << 20), cutoff or (200 << 20), maxp or 10000)
        assert tpi.s3_plan_parts(size, chunk, cutoff, maxp) == (want["multipart"], want["part_size"], want["nparts"]), (k, got)
        for key in ("multipart", "part_size", "nparts", "md5", "content_md5", "etag", "part_content_md5"):
            assert got[key] == want[key], (k, key, got[key], want[key])
        assert np.array_equal(got["part_md5"], want["part_md5"])
        assert got["content_md5"] == base64.b64encode(hashlib.md5(data.tobytes()).digest()).decode()
        if want["multipart"]:
            lean = ctx.s3_upload_plan(path, size, chunk, cutoff, maxp, no_whole_md5=True)
            assert lean["md5"] is None and lean["content_md5"] is None and lean["etag"] == want["etag"]
    # a file that is not the listed size is an error, as everywhere
    with pytest.raises(tpi.TpiMd5Error):
        ctx.s3_upload_plan(str(tmp_path / "obj1.bin"), 2)


def test_progress_counters_advance_with_the_calls(ctx):
    before = ctx.progress()
    buf, off, ln = oracle.synth_packed(71, [100, 5000, 0, 70000], align=16)
    ctx.hash_buffers(buf, off, ln)
    after = ctx.progress()
    assert after["calls_started"] - before["calls_started"] == 1 == after["calls_finished"] - before["calls_finished"]
    assert after["units_done"] - before["units_done"] == 4
    assert after["bytes_done"] - before["bytes_done"] == 75100 == after["bytes_staged"] - before["bytes_staged"]
