"""GPU parity: every result of the CUDA path, taken through the C-ABI, against the CPU oracle
(bit-exact -- this is integer/byte work) and against the committed golden vectors."""
import hashlib
import os

import numpy as np
import pytest

import oracle
import tpi_b200 as tpi

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = tpi.Context([0], staging_bytes=8 << 20)   # small staging: exercises multi-batch + chunked units
    yield c
    c.close()


def pat(n):
    return bytes(i % 251 for i in range(n))


def check_units(ctx, buf, off, ln):
    got = ctx.hash_buffers(buf, off, ln)
    want = oracle.md5_batch(buf, off, ln, threads=4)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert bad.size == 0, f"{bad.size} digests differ, first unit {bad[:5]} len {np.asarray(ln)[bad[:5]]}"
    return got


def test_rfc_and_pattern_vectors(ctx, golden):
    msgs = [s.encode() for s, _ in golden["rfc1321"]] + [pat(int(n)) for n in golden["pattern251"]]
    want = [h for _, h in golden["rfc1321"]] + list(golden["pattern251"].values())
    buf = np.frombuffer(b"".join(msgs) + b"\0", dtype=np.uint8)
    ln = np.array([len(m) for m in msgs], dtype=np.uint64)
    off = np.concatenate([[0], np.cumsum(ln)[:-1]]).astype(np.uint64)
    got = ctx.hash_buffers(buf, off, ln)
    assert [g.tobytes().hex() for g in got] == want


def test_golden_synth_vectors(ctx, golden):
    for seed, idx, n, h in golden["synth_md5"]:
        b = oracle.synth_bytes(seed, idx, n)
        got = ctx.hash_buffers(b if n else np.zeros(1, np.uint8), [0], [n])
        assert got[0].tobytes().hex() == h, n


@pytest.mark.parametrize("align", [16, 1, 4])
def test_every_padding_length(ctx, align):
    # all lengths 0..300: covers len mod 64 in {0, 55, 56, 63} several times; packed at the given alignment
    lengths = list(range(0, 301))
    buf, off, ln = oracle.synth_packed(21, lengths, align=align)
    check_units(ctx, buf, off, ln)


def test_ragged_batch_sorted_and_unsorted(ctx):
    rng = np.random.default_rng(42)
    lengths = np.concatenate([rng.integers(0, 20000, 3000), rng.integers(0, 70, 1000), [0, 0, 1 << 20, 3 << 20]])
    rng.shuffle(lengths)
    buf, off, ln = oracle.synth_packed(22, lengths.tolist(), align=1)
    check_units(ctx, buf, off, ln)


def test_scattered_and_overlapping_units_use_the_pack_path(ctx):
    rng = np.random.default_rng(43)
    buf = oracle.synth_bytes(23, 0, 1 << 20)
    off = rng.integers(0, (1 << 20) - 5000, 2000).astype(np.uint64)      # unordered, overlapping
    ln = rng.integers(0, 5000, 2000).astype(np.uint64)
    check_units(ctx, buf, off, ln)


def test_unit_larger_than_staging_is_chained_across_launches(ctx):
    # 8 MiB staging; a 30 MiB unit + neighbours => several launches with carried chaining values
    lengths = [1000, (30 << 20) + 77, 64, 0, (9 << 20), 55]
    buf, off, ln = oracle.synth_packed(24, lengths, align=1)
    got = check_units(ctx, buf, off, ln)
    assert got[1].tobytes() == hashlib.md5(buf[int(off[1]):int(off[1] + ln[1])].tobytes()).digest()


def test_config2_shape_small(ctx):
    # BASELINE config 2 in miniature: many 4 KiB files, contiguous, through the host API
    n = 20000
    buf, off, ln = oracle.synth_packed(2, [4096] * n, align=1)
    check_units(ctx, buf, off, ln)


def test_s3_etag_golden_and_oracle(ctx, golden):
    ex = golden["s3_example"]
    data = np.frombuffer(pat(5 << 20) * 2 + b"tail", dtype=np.uint8)
    etag, n, parts = ctx.s3_etag_buffer(data, ex["part_size"])
    assert tpi.s3_etag_text(etag, n) == ex["etag"]
    for seed, idx, size, ps, want in golden["synth_s3"]:
        b = oracle.synth_bytes(seed, idx, size)
        etag, n, parts = ctx.s3_etag_buffer(b if size else np.zeros(1, np.uint8)[:0], ps)
        assert tpi.s3_etag_text(etag, n) == want
        o_etag, o_n, o_parts = oracle.s3_etag(b, ps)
        assert n == o_n and np.array_equal(parts, o_parts)


def test_s3_etag_file_and_hash_files(ctx, tmp_path):
    sizes = [0, 1, 4096, 100000, (9 << 20) + 5, 63, 64]
    paths = []
    for i, n in enumerate(sizes):
        p = tmp_path / f"f{i}.bin"
        p.write_bytes(oracle.synth_bytes(25, i, n).tobytes())
        paths.append(str(p))
    got = ctx.hash_files(paths)                                  # sizes by fstat
    got2 = ctx.hash_files(paths, sizes)
    for i, n in enumerate(sizes):
        assert got[i].tobytes() == oracle.md5(oracle.synth_bytes(25, i, n)) == got2[i].tobytes()
    etag, n, parts = ctx.s3_etag_file(paths[4], sizes[4], 1 << 20)
    o = oracle.s3_etag(oracle.synth_bytes(25, 4, sizes[4]), 1 << 20)
    assert (etag, n) == (o[0], o[1]) and np.array_equal(parts, o[2])
    # a file that changed size since it was listed is an error, not a digest
    with pytest.raises(tpi.TpiMd5Error) as e:
        ctx.hash_files(paths, [s + (1 if i == 2 else 0) for i, s in enumerate(sizes)])
    assert e.value.code == -4 and e.value.failed_index == 2
    with pytest.raises(tpi.TpiMd5Error) as e:
        ctx.hash_files(paths, [s - (1 if i == 3 else 0) for i, s in enumerate(sizes)])          # listed shorter than it is
    assert e.value.code == -4 and e.value.failed_index == 3
    with pytest.raises(tpi.TpiMd5Error) as e:
        ctx.hash_files(paths, [1 if i == 0 else s for i, s in enumerate(sizes)])                # empty file listed as 1 byte
    assert e.value.failed_index == 0
    open(paths[0], "wb").write(b"x")                                                            # listed empty, has grown
    with pytest.raises(tpi.TpiMd5Error) as e:
        ctx.hash_files(paths, sizes)
    assert e.value.failed_index == 0
    with pytest.raises(tpi.TpiMd5Error):
        ctx.hash_files(paths + [str(tmp_path / "missing")])


@pytest.mark.parametrize("staging", [1 << 20, 4 << 20])
def test_long_units_are_hashed_in_slices_side_by_side(tmp_path, staging):
    """The packed path takes long units in slices, up to 384 of them side by side, each batch's slices continuing
    from the previous batch's chaining values while whole short units ride in the same buffers as a second launch
    (run_device_impl, csrc/tpimd5.cu).  Small staging buffers put every case of that planner within a few MiB:
    units above half a buffer (always sliced), many equal long units (the parts of one object: sliced because
    whole units in order would end later), more long units than are active at once, lengths that end exactly on a
    slice, on a slice + 1 and on a 64-byte multiple, files that grew or shrank in a sliced unit."""
    rng = np.random.default_rng(staging)
    sizes = [0, 1, 63, 64, 65, 4096, staging // 2 - 1, staging // 2, staging // 2 + 1, staging, staging + 1, 3 * staging + 77,
             7 * staging, staging // 4, staging // 4 + 64, 2 * staging + 64]
    sizes += [int(x) for x in rng.integers(0, 3000, 300)] + [int(x) for x in rng.integers(staging // 4, staging, 40)]
    blob = oracle.synth_bytes(31, 0, max(sizes) + len(sizes))
    paths = []
    for i, n in enumerate(sizes):
        p = tmp_path / f"u{i}.bin"
        p.write_bytes(blob[i:i + n].tobytes())
        paths.append(str(p))
    want = np.stack([np.frombuffer(oracle.md5(blob[i:i + n]), dtype=np.uint8) for i, n in enumerate(sizes)])
    with tpi.Context([0], staging_bytes=staging) as c:
        assert np.array_equal(c.hash_files(paths, sizes), want)
        assert np.array_equal(c.hash_files(paths), want)
        # the same units out of pageable memory (packed), in caller order rather than longest first
        off = np.cumsum([0] + [n + 3 for n in sizes[:-1]]).astype(np.uint64)
        buf = np.zeros(int(off[-1]) + sizes[-1], dtype=np.uint8)
        for i, n in enumerate(sizes):
            buf[int(off[i]):int(off[i]) + n] = blob[i:i + n]
        assert np.array_equal(c.hash_buffers(buf, off, np.array(sizes, dtype=np.uint64)), want)
        # the parts of one object: 500 equal long units and a ragged last one, more than are active at once
        part = staging // 4 + 128
        size = 500 * part + 12345
        obj = tmp_path / "object.bin"
        data = oracle.synth_files(32, 0, size, 1, threads=8)
        obj.write_bytes(data.tobytes())
        etag, n, parts = c.s3_etag_file(str(obj), size, part)
        o = oracle.s3_etag(data, part, threads=8)
        assert (etag, n) == (o[0], o[1]) and np.array_equal(parts, o[2])
        # a sliced file that is longer / shorter than listed is an error that names it
        for delta in (1, -1):
            listed = list(sizes)
            listed[12] += delta
            with pytest.raises(tpi.TpiMd5Error) as e:
                c.hash_files(paths, listed)
            assert e.value.code == -4 and e.value.failed_index == 12


def test_ingest_probe_reads_what_hash_files_reads(ctx, tmp_path):
    """tpimd5_ingest_probe (the host-ingest ceiling bench.py quotes): same planning and reads as hash_files, no
    digests; a file that is missing or shorter than listed is the same error, with its index."""
    sizes = [0, 5, 70000, (3 << 20) + 1]
    paths = []
    for i, n in enumerate(sizes):
        p = tmp_path / f"p{i}.bin"
        p.write_bytes(oracle.synth_bytes(41, i, n).tobytes())
        paths.append(str(p))
    before = ctx.launch_count()
    ctx.ingest_probe(paths, sizes)
    ctx.ingest_probe([paths[3]] * 3, [1 << 20, 1 << 20, (1 << 20) + 1], file_offsets=[0, 1 << 20, 2 << 20])
    assert ctx.launch_count() == before                       # nothing was hashed
    with pytest.raises(tpi.TpiMd5Error) as e:
        ctx.ingest_probe(paths, [0, 5, 70001, sizes[3]])
    assert e.value.code == -4
    with pytest.raises(tpi.TpiMd5Error):
        ctx.ingest_probe(paths + [str(tmp_path / "missing")], sizes + [1])
    got = ctx.hash_files(paths, sizes)                        # and the context still hashes afterwards
    for i, n in enumerate(sizes):
        assert got[i].tobytes() == oracle.md5(oracle.synth_bytes(41, i, n))


def test_stream_is_a_go_hash(ctx):
    data = oracle.synth_bytes(26, 0, (9 << 20) + 13).tobytes()
    s, o = ctx.stream(), oracle.MD5Stream()
    assert s.sum().hex() == "d41d8cd98f00b204e9800998ecf8427e"
    pos = 0
    for step in (1, 62, 1, 64, 65, 4096, 5 << 20, 3, 4 << 20, 1 << 30):
        s.write(data[pos:pos + step]); o.write(data[pos:pos + step])
        pos = min(len(data), pos + step)
        assert s.sum() == o.sum() == hashlib.md5(data[:pos]).digest()
        assert s.sum() == o.sum()                                # Sum leaves the stream as it was
    s.reset()
    s.write(b"abc")
    assert s.sum().hex() == "900150983cd24fb0d6963f7d28e17f72"
    s.close()


def test_many_streams_advance_in_one_launch(ctx):
    rng = np.random.default_rng(44)
    n = 300
    streams = [ctx.stream() for _ in range(n)]
    ref = [hashlib.md5() for _ in range(n)]
    for _ in range(4):
        chunks = [rng.integers(0, 256, int(rng.integers(0, 3000)), dtype=np.uint8) for _ in range(n)]
        ctx.stream_write_many(streams, chunks)
        for r, c in zip(ref, chunks):
            r.update(c.tobytes())
    for s, r in zip(streams, ref):
        assert s.sum() == r.digest()


def test_diff_against_oracle(ctx):
    rng = np.random.default_rng(45)
    for n in (1, 31, 32, 1023, 1024, 1025, 100_000):
        old = rng.integers(0, 256, (n, 16), dtype=np.uint8)
        new = old.copy()
        k = max(1, n // 20)
        changed = np.sort(rng.choice(n, k, replace=False))
        new[changed, rng.integers(0, 16, k)] ^= 0x10
        assert np.array_equal(ctx.diff(old, new), oracle.diff(old, new))
        assert np.array_equal(ctx.diff(old, new), changed.astype(np.uint32))
        assert ctx.diff(old, old).size == 0


def test_device_generator_matches_host_generator(ctx):
    import torch
    file_len, n = 4096, 257
    d = torch.empty(file_len * n, dtype=torch.uint8, device="cuda")
    ctx.dev_synth_fill(d, seed=2, first_file_index=5, file_len=file_len, n_files=n)
    torch.cuda.synchronize()
    host = d.cpu().numpy().reshape(n, file_len)
    for i in (0, 1, 100, 256):
        assert np.array_equal(host[i], oracle.synth_bytes(2, 5 + i, file_len))


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4, 7, 9, 41, 44, 64, 67])
def test_every_kernel_variant_agrees(variant):
    with tpi.Context([0], staging_bytes=4 << 20) as c:
        c.set_variant(variant)
        lengths = list(range(0, 200)) + [4096] * 300 + [70000, 5 << 20]
        for align in (16, 1):
            buf, off, ln = oracle.synth_packed(27, lengths, align=align)
            check_units(c, buf, off, ln)


def test_device_resident_entry_points():
    import torch
    with tpi.Context([0]) as c:
        n, flen = 5000, 4096
        d = torch.empty(n * flen, dtype=torch.uint8, device="cuda")
        c.dev_synth_fill(d, 2, 0, flen, n)
        dig = torch.empty((n, 16), dtype=torch.uint8, device="cuda")
        c.dev_hash_buffers(d, n * flen, n, dig, stride=flen, stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        want = oracle.md5_batch(d.cpu().numpy(), np.arange(n, dtype=np.uint64) * flen, np.full(n, flen, np.uint64), 4)
        assert np.array_equal(dig.cpu().numpy(), want)
        # explicit offsets on the device, unaligned
        off = torch.tensor([1, 4099, 9000, 9001], dtype=torch.int64, device="cuda")
        ln = torch.tensor([4000, 4901, 1, 0], dtype=torch.int64, device="cuda")
        dig2 = torch.empty((4, 16), dtype=torch.uint8, device="cuda")
        c.dev_hash_buffers(d, n * flen, 4, dig2, d_offsets=off, d_lengths=ln)
        torch.cuda.synchronize()
        want2 = oracle.md5_batch(d.cpu().numpy(), off.cpu().numpy().astype(np.uint64), ln.cpu().numpy().astype(np.uint64))
        assert np.array_equal(dig2.cpu().numpy(), want2)
        # S3 ETag on device data, library scratch for the part digests
        etag = torch.empty(16, dtype=torch.uint8, device="cuda")
        size, ps = n * flen - 100, 1 << 20
        c.dev_s3_etag(d, size, ps, etag)
        torch.cuda.synchronize()
        o = oracle.s3_etag(d.cpu().numpy()[:size], ps, threads=4)
        assert etag.cpu().numpy().tobytes() == o[0]
        # diff on device
        a = torch.randint(0, 256, (3000, 16), dtype=torch.uint8, device="cuda")
        b = a.clone()
        b[[5, 77, 2999], 3] ^= 1
        idx = torch.empty(3000, dtype=torch.int32, device="cuda")
        cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
        c.dev_diff(a, b, 3000, idx, cnt)
        torch.cuda.synchronize()
        assert cnt.item() == 3 and idx[:3].tolist() == [5, 77, 2999]


def test_full_size_properties_config2():
    """BASELINE config 2 at full size (1 Mi files x 4 KiB, generated in HBM): too big for the
    oracle in seconds, so check size-independent properties -- a sample against the oracle, a
    repeat run is identical, one flipped byte changes exactly one digest (found by K4)."""
    import torch
    n, flen = 1 << 20, 4096
    with tpi.Context([0]) as c:
        d = torch.empty(n * flen, dtype=torch.uint8, device="cuda")
        c.dev_synth_fill(d, 2, 0, flen, n)
        dig = torch.empty((n, 16), dtype=torch.uint8, device="cuda")
        c.dev_hash_buffers(d, n * flen, n, dig, stride=flen)
        torch.cuda.synchronize()
        host = dig.cpu().numpy()
        rng = np.random.default_rng(2)
        for i in [0, 1, n - 1] + rng.integers(0, n, 200).tolist():
            assert host[i].tobytes() == oracle.md5(oracle.synth_bytes(2, i, flen)), i
        assert len({r.tobytes() for r in host[:100000]}) == 100000        # no collisions / stuck lanes
        victim = 777_777
        d[victim * flen + 1234] ^= 0x40
        dig2 = torch.empty_like(dig)
        c.dev_hash_buffers(d, n * flen, n, dig2, stride=flen)
        idx = torch.empty(n, dtype=torch.int32, device="cuda")
        cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
        c.dev_diff(dig, dig2, n, idx, cnt)
        torch.cuda.synchronize()
        assert cnt.item() == 1 and idx[0].item() == victim


@pytest.mark.parametrize("variant", [0, 201, 202, 203, 204, 241, 244, 262, 245, 246, 247, 265,
                                     341, 342, 343, 344, 345, 346, 347, 348, 349, 301, 363,
                                     441, 442, 443, 446, 448, 449, 461, 365, 364, 262])
def test_tensor_map_kernel_for_strided_units(variant):
    """Equal-sized units at a fixed stride through the TMA kernel: every stride class (multiples of
    16 around the 64/128-byte block and stage boundaries), unit counts that do not fill the last
    warp or CTA, and a buffer that ends inside its last unit."""
    import torch
    with tpi.Context([0]) as c:
        c.set_variant(variant)
        cases = [(16, 1000), (48, 333), (64, 129), (112, 70), (128, 1), (144, 31), (192, 500), (256, 4097),
                 (4096, 2051), (4112, 300), (65536 + 16, 40), (1 << 20, 9)]
        for stride, n in cases:
            for ragged in (0, 5):
                nbytes = stride * n - (stride // 2 + 3 if ragged else 0)
                d = torch.empty(stride * n, dtype=torch.uint8, device="cuda")
                flen = 1 << 12 if (stride * n) % (1 << 12) == 0 else 16
                c.dev_synth_fill(d, 30 + ragged, 0, flen, (stride * n) // flen)
                dig = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
                c.dev_hash_buffers(d, nbytes, n, dig, stride=stride)
                torch.cuda.synchronize()
                host = d.cpu().numpy()
                off = np.arange(n, dtype=np.uint64) * stride
                ln = np.minimum(np.uint64(stride), np.uint64(nbytes) - off).astype(np.uint64)
                want = oracle.md5_batch(host, off, ln, threads=4)
                got = dig.cpu().numpy()
                bad = np.nonzero((got != want).any(axis=1))[0]
                assert bad.size == 0, (variant, stride, n, ragged, bad[:5])


def _free_gib():
    import torch
    free, _ = torch.cuda.mem_get_info()
    return free / (1 << 30)


def test_full_size_properties_config3_s3_etag():
    """BASELINE config 3, one of its 64 GiB files with 8 MiB parts, generated in HBM.  The oracle
    cannot hash 64 GiB in seconds, so: a sample of part digests against the oracle, the
    digest-of-digests against the oracle's MD5 of the GPU's own part digests (checks K2b and the
    part count 8192), and the allocation-free/explicit-scratch variants agree."""
    import torch
    size, part = 64 << 30, 8 << 20
    if _free_gib() < 70:
        pytest.skip("needs 64 GiB of free HBM")
    with tpi.Context([0]) as c:
        d = torch.empty(size, dtype=torch.uint8, device="cuda")
        c.dev_synth_fill(d, 3, 0, size, 1)                       # file 0 of workdir seed 3
        n = tpi.s3_nparts(size, part)
        assert n == 8192
        parts = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
        etag = torch.zeros(16, dtype=torch.uint8, device="cuda")
        c.dev_s3_etag(d, size, part, etag, d_part_digests=parts)
        torch.cuda.synchronize()
        ph = parts.cpu().numpy()
        for i in (0, 1, 4095, 8191):
            assert ph[i].tobytes() == oracle.md5(oracle.synth_bytes(3, 0, part, i * part)), i
        assert etag.cpu().numpy().tobytes() == oracle.md5(ph.tobytes())
        etag2 = torch.zeros(16, dtype=torch.uint8, device="cuda")
        c.dev_s3_etag(d, size, part, etag2)                      # library scratch
        torch.cuda.synchronize()
        assert torch.equal(etag, etag2)
        # ragged: drop 5 bytes -> last part short, same part count; grow by one byte -> 8193 parts
        c.dev_s3_etag(d, size - 5, part, etag2, d_part_digests=parts)
        torch.cuda.synchronize()
        ph2 = parts.cpu().numpy()
        assert np.array_equal(ph2[:-1], ph[:-1]) and not np.array_equal(ph2[-1], ph[-1])
        assert ph2[-1].tobytes() == oracle.md5(oracle.synth_bytes(3, 0, part - 5, 8191 * part))
        assert etag2.cpu().numpy().tobytes() == oracle.md5(ph2.tobytes())


def test_full_size_properties_config5_manifest_rebuild_and_diff():
    """BASELINE config 5: 25,600 files x 4 MiB = 100 GiB, 1,280 files mutated by one byte (same size,
    so only a checksum sees it).  Manifest rebuild + K4 diff must return exactly those 1,280 indices."""
    import torch
    n, flen = 25600, 4 << 20
    if _free_gib() < 105:
        pytest.skip("needs 100 GiB of free HBM")
    with tpi.Context([0]) as c:
        d = torch.empty(n * flen, dtype=torch.uint8, device="cuda")
        c.dev_synth_fill(d, 5, 0, flen, n)
        old = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
        c.dev_hash_buffers(d, n * flen, n, old, stride=flen)
        rng = np.random.default_rng(5)
        victims = np.sort(rng.choice(n, 1280, replace=False))
        offs = rng.integers(0, flen, 1280)
        pos = torch.from_numpy(victims.astype(np.int64) * flen + offs).cuda()
        d[pos] ^= 0x01
        new = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
        c.dev_hash_buffers(d, n * flen, n, new, stride=flen)
        idx = torch.zeros(n, dtype=torch.int32, device="cuda")
        cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
        c.dev_diff(old, new, n, idx, cnt)
        torch.cuda.synchronize()
        assert cnt.item() == 1280
        assert np.array_equal(idx[:1280].cpu().numpy().astype(np.int64), victims)
        oh = old.cpu().numpy()
        for i in (0, 12345, n - 1):
            assert oh[i].tobytes() == oracle.md5(oracle.synth_bytes(5, i, flen)), i
        v = int(victims[7])
        b = oracle.synth_bytes(5, v, flen).copy()
        b[int(offs[7])] ^= 0x01
        assert new[v].cpu().numpy().tobytes() == oracle.md5(b)


def test_zipf_sizes_config4_shape_scaled():
    """BASELINE config 4 scaled down (Zipf sizes 1 KiB .. 256 MiB, ~2 GiB): whole-file content-MD5 of a
    skewed workdir through the host API; every digest against the oracle."""
    k = np.arange(1, 4000)
    sizes = np.maximum(1024, ((256 << 20) * k ** -1.111).astype(np.int64))
    rng = np.random.default_rng(4)
    rng.shuffle(sizes)
    buf, off, ln = oracle.synth_packed(4, sizes.tolist(), align=1)
    with tpi.Context([0], staging_bytes=64 << 20) as c:
        got = c.hash_buffers(buf, off, ln)
    want = oracle.md5_batch(buf, off, ln, threads=8, engine=oracle.ENGINE_OPENSSL if oracle.lib().orc_has_openssl() else 0)
    assert np.array_equal(got, want)


def test_sliced_calls_from_many_threads_share_the_staging_buffers():
    """Eight threads, one context with 1 MiB staging buffers, every call with units of several buffers' length: the
    slices of a call continue from chaining values written by kernels in buffers that are meanwhile leased, handed
    back and taken by other calls; each call's digests must still be its own."""
    import threading
    stage = 1 << 20
    jobs = []
    for k in range(8):
        rng = np.random.default_rng(500 + k)
        lengths = [int(x) for x in rng.integers(0, 4000, 150)] + [int(x) for x in rng.integers(stage // 4, 5 * stage, 6)] + [3 * stage + k]
        rng.shuffle(lengths)
        buf, off, ln = oracle.synth_packed(500 + k, lengths, align=1)
        jobs.append((buf, off, ln, oracle.md5_batch(buf, off, ln, threads=2)))
    errs = []
    with tpi.Context([0], staging_bytes=stage) as c:
        def worker(k):
            try:
                buf, off, ln, want = jobs[k]
                for _ in range(2):
                    got = c.hash_buffers(buf, off, ln)
                    bad = np.nonzero((got != want).any(axis=1))[0]
                    assert bad.size == 0, (k, bad[:4], ln[bad[:4]])
            except Exception as e:      # noqa: BLE001
                errs.append(e)

        th = [threading.Thread(target=worker, args=(k,)) for k in range(8)]
        [t.start() for t in th]
        [t.join() for t in th]
    assert not errs, errs


def test_one_context_many_threads(ctx):
    """Terraform runs resource operations concurrently in one plugin process and rclone hashes from
    8 checker + 4 transfer goroutines: a context must take calls from many threads at once."""
    import threading
    lengths = [int(x) for x in np.random.default_rng(71).integers(0, 50000, 300)]
    buf, off, ln = oracle.synth_packed(71, lengths, align=1)
    want = oracle.md5_batch(buf, off, ln, threads=4)
    errs = []

    def worker(k):
        try:
            for _ in range(3):
                sel = np.arange(k, len(lengths), 4)
                got = ctx.hash_buffers(buf, off[sel], ln[sel])
                assert np.array_equal(got, want[sel])
                assert ctx.diff(want, want).size == 0
        except Exception as e:      # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs


def test_mixed_calls_from_many_threads_and_an_error_in_one_of_them(ctx, tmp_path):
    """Twelve threads on one context, every kind of host call at once -- buffers, files, S3 ETags, streams, diffs --
    while one thread keeps asking for a file that has shrunk since its listing.  That call must fail with TPIMD5_EIO every
    time, hand back everything it leased, and leave the other threads' results bit-exact."""
    import threading
    rng = np.random.default_rng(72)
    lengths = [int(x) for x in rng.integers(0, 200000, 120)] + [3 << 20, (5 << 20) + 7]
    buf, off, ln = oracle.synth_packed(72, lengths, align=16)
    want = oracle.md5_batch(buf, off, ln, threads=4)
    paths = []
    for i in range(40):
        p = tmp_path / f"f{i:02d}.bin"
        buf[int(off[i]): int(off[i] + ln[i])].tofile(p)
        paths.append(str(p))
    short = tmp_path / "shrunk.bin"
    short.write_bytes(b"x" * 100)
    big = buf[int(off[-1]): int(off[-1] + ln[-1])]
    etag_want = oracle.s3_etag(big, 1 << 20)
    errs, eio = [], []

    def worker(k):
        try:
            for it in range(4):
                kind = (k + it) % 6
                if kind == 0:
                    sel = np.arange(k % 5, len(lengths), 5)
                    assert np.array_equal(ctx.hash_buffers(buf, off[sel], ln[sel]), want[sel])
                elif kind == 1:
                    assert np.array_equal(ctx.hash_files(paths, ln[:40]), want[:40])
                elif kind == 2:
                    e, n, parts = ctx.s3_etag_buffer(big, 1 << 20)
                    assert (e, n) == etag_want[:2] and np.array_equal(parts, etag_want[2])
                elif kind == 3:
                    h = ctx.stream()
                    for lo in range(0, int(ln[-2]), 700001):
                        h.write(buf[int(off[-2]) + lo: int(off[-2]) + min(int(ln[-2]), lo + 700001)])
                    assert h.sum() == want[-2].tobytes()
                    h.close()
                elif kind == 4:
                    other = want.copy()
                    other[[5, 77], 3] ^= 1
                    assert ctx.diff(want, other).tolist() == [5, 77]
                else:
                    try:
                        ctx.hash_files(paths[:10] + [str(short)], list(ln[:10]) + [5000])
                        errs.append("a shrunk file produced digests")
                    except tpi.TpiMd5Error as e:
                        assert e.code == -4 and e.failed_index == 10, (e.code, e.failed_index)
                        eio.append(1)
        except Exception as e:      # noqa: BLE001
            errs.append(repr(e))

    th = [threading.Thread(target=worker, args=(k,)) for k in range(12)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    assert len(eio) == 8                                           # 12 threads x 4 rounds, one kind in six
    p = ctx.progress()
    assert p["calls_started"] == p["calls_finished"]
    assert np.array_equal(ctx.hash_buffers(buf, off, ln), want)    # and the context is as good as new


def test_units_extent_and_the_binding_bounds_check():
    """tpimd5_units_extent (the helper a binding checks its slice length with): the furthest unit end, over more
    units than one pool item, overflow refused; the Python binding refuses a unit past the end of its buffer."""
    import ctypes as C
    rng = np.random.default_rng(12)
    n = 300000
    off = rng.integers(0, 1 << 40, n).astype(np.uint64)
    ln = rng.integers(0, 1 << 30, n).astype(np.uint64)
    with tpi.Context([0]) as c:
        end = C.c_uint64(7)
        L = tpi.load_library()
        assert L.tpimd5_units_extent(c._h, off.ctypes.data, ln.ctypes.data, n, C.byref(end)) == 0
        assert end.value == int((off + ln).max())
        assert L.tpimd5_units_extent(c._h, off.ctypes.data, ln.ctypes.data, 0, C.byref(end)) == 0 and end.value == 0
        off[123456] = np.uint64(2**64 - 5)
        ln[123456] = np.uint64(10)
        assert L.tpimd5_units_extent(c._h, off.ctypes.data, ln.ctypes.data, n, C.byref(end)) == -6   # TPIMD5_ERANGE
        buf = np.zeros(1000, dtype=np.uint8)
        with pytest.raises(ValueError):
            c.hash_buffers(buf, np.array([0, 990], dtype=np.uint64), np.array([10, 11], dtype=np.uint64))
        got = c.hash_buffers(buf, np.array([0, 990], dtype=np.uint64), np.array([10, 10], dtype=np.uint64))
        assert np.array_equal(got, oracle.md5_batch(buf, np.array([0, 990]), np.array([10, 10])))


@pytest.mark.parametrize("variant", [0, 501, 502, 503])
def test_ragged_latency_kernel_shapes(variant):
    """Ragged host batches through md5_ragged_chain_kernel in each of its shapes (512-byte stages; 256-byte stages in step
    modes 4 and 6, which the library picks from 593 and 889 warps of units): every length class -- empty, sub-block,
    55/56/63/64 bytes mod 64, stage multiples +- 1, lengths that end in every block of an iteration -- sorted and
    unsorted, plus units longer than a staging buffer (chained across launches) and a batch of more than 592 warps."""
    rng = np.random.default_rng(90 + variant)
    lengths = list(range(0, 200)) + [256 * k + d for k in range(1, 40) for d in (-1, 0, 1, 55, 56, 63)]
    lengths += [int(x) for x in rng.integers(0, 300000, 600)] + [(9 << 20) + 5, 5 << 20, 0, 0]
    rng.shuffle(lengths)
    buf, off, ln = oracle.synth_packed(73, lengths, align=1)
    want = oracle.md5_batch(buf, off, ln, threads=8)
    with tpi.Context([0], staging_bytes=8 << 20) as c:
        if variant:
            c.set_variant(variant)
        got = c.hash_buffers(buf, off, ln)
        bad = np.nonzero((got != want).any(axis=1))[0]
        assert bad.size == 0, (variant, bad[:5], ln[bad[:5]])
    # many small ragged units in one batch: more than 592 warps of them
    lengths = [int(x) for x in rng.integers(0, 3000, 24000)]
    buf, off, ln = oracle.synth_packed(74, lengths, align=1)
    want = oracle.md5_batch(buf, off, ln, threads=8)
    with tpi.Context([0]) as c:
        if variant:
            c.set_variant(variant)
        got = c.hash_buffers(buf, off, ln)
        bad = np.nonzero((got != want).any(axis=1))[0]
        assert bad.size == 0, (variant, bad[:5], ln[bad[:5]])


def test_concurrent_callers_overlap_on_one_gpu():
    """Eight threads, one context, one GPU, each hashing its own workdir whose longest file is 8 MiB (one job is
    bound by that file's serial chain and leaves the GPU almost idle): the calls lease staging buffers, tables
    and streams instead of queueing behind one lock, so all eight take about twice as long as one (packing and
    copying eight workdirs is 80 ms of host and PCIe work by itself; one job alone is 60 ms) instead of eight
    times -- and every digest is the oracle's."""
    import subprocess
    import sys
    import json
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "bench_concurrent.py"), "8"], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    rec = json.loads(r.stdout.strip().splitlines()[-1])
    assert rec["bit_exact"] and rec["ratio"] <= 2.5 and rec["all_jobs_ms"] <= 0.4 * rec["serial_would_be_ms"], rec


@pytest.mark.skipif(tpi.device_count() < 2, reason="needs two GPUs in one process")
def test_one_context_two_devices_shards_and_gathers():
    """In-process multi-GPU: units split byte-balanced over the context's devices, digests gathered
    through pinned host memory; results must not depend on the split."""
    rng = np.random.default_rng(72)
    lengths = [int(x) for x in (rng.pareto(1.2, 3000) * 3000).astype(np.int64)] + [0, 40 << 20, 7]
    buf, off, ln = oracle.synth_packed(72, lengths, align=16)
    want = oracle.md5_batch(buf, off, ln, threads=8, engine=oracle.ENGINE_OPENSSL if oracle.lib().orc_has_openssl() else 0)
    with tpi.Context(list(range(min(tpi.device_count(), 8))), staging_bytes=16 << 20) as c:
        assert np.array_equal(c.hash_buffers(buf, off, ln), want)
        etag, n, parts = c.s3_etag_buffer(buf[: 50 << 20], 1 << 20)
        o = oracle.s3_etag(buf[: 50 << 20], 1 << 20, threads=4)
        assert (etag, n) == (o[0], o[1]) and np.array_equal(parts, o[2])


def test_device_ragged_units_are_bucketed_by_length():
    """Device-resident units of arbitrary length and alignment (>= 4096 of them take the radix-sort
    bucketing path): digests must land at the caller's indices."""
    import torch
    rng = np.random.default_rng(73)
    n = 30000
    lengths = np.concatenate([rng.integers(0, 9000, n - 3), [0, 1 << 20, 63]]).astype(np.int64)
    rng.shuffle(lengths)
    buf, off, ln = oracle.synth_packed(73, lengths.tolist(), align=1)
    want = oracle.md5_batch(buf, off, ln, threads=8, engine=oracle.ENGINE_OPENSSL if oracle.lib().orc_has_openssl() else 0)
    with tpi.Context([0]) as c:
        d = torch.from_numpy(buf).cuda()
        d_off = torch.from_numpy(off.astype(np.int64)).cuda()
        d_len = torch.from_numpy(ln.astype(np.int64)).cuda()
        dig = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
        for _ in range(2):                                           # twice: the scratch is reused
            dig.zero_()
            c.dev_hash_buffers(d, buf.size, n, dig, d_offsets=d_off, d_lengths=d_len)
            torch.cuda.synchronize()
            assert np.array_equal(dig.cpu().numpy(), want)


def test_repeated_launches_are_deterministic():
    """compute-sanitizer is closed on this pool, so shared-memory races are hunted the blunt way: the
    same launch 30 times over, every kernel family, all digests identical every time."""
    import torch
    with tpi.Context([0]) as c:
        n, flen = 40000, 4096 + 48
        d = torch.empty(n * flen, dtype=torch.uint8, device="cuda")
        c.dev_synth_fill(d, 74, 0, 16, n * flen // 16)
        ref_s = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
        c.dev_hash_buffers(d, n * flen, n, ref_s, stride=flen)                      # tensor-map kernel
        off = torch.arange(n, dtype=torch.int64, device="cuda") * flen + 3          # general kernel, unaligned
        ln = torch.full((n,), flen - 7, dtype=torch.int64, device="cuda")
        ref_g = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
        c.dev_hash_buffers(d, n * flen, n, ref_g, d_offsets=off, d_lengths=ln)
        torch.cuda.synchronize()
        out = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
        for _ in range(30):
            c.dev_hash_buffers(d, n * flen, n, out, stride=flen)
            assert torch.equal(out, ref_s)
            c.dev_hash_buffers(d, n * flen, n, out, d_offsets=off, d_lengths=ln)
            assert torch.equal(out, ref_g)


def test_more_units_than_one_batch_of_descriptors(ctx):
    """1.3 M tiny units: a batch closes at 1 Mi descriptor rows even though the staging buffer is nearly
    empty, so this goes through several batches; plus empty units in bulk."""
    n = 1_300_000
    rng = np.random.default_rng(75)
    ln = rng.integers(0, 9, n).astype(np.uint64)
    off = np.concatenate([[0], np.cumsum(ln)[:-1]]).astype(np.uint64)
    buf = oracle.synth_bytes(75, 0, int(ln.sum()) + 1)
    got = ctx.hash_buffers(buf, off, ln)
    want = oracle.md5_batch(buf, off, ln, threads=8, engine=oracle.ENGINE_OPENSSL if oracle.lib().orc_has_openssl() else 0)
    assert np.array_equal(got, want)


def test_bad_arguments_are_refused_with_a_message(ctx):
    import torch
    d = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    dig = torch.zeros((4, 16), dtype=torch.uint8, device="cuda")
    cases = [
        lambda: ctx.dev_hash_buffers(d, 4096, 4, dig, stride=0),                           # no stride, no offsets
        lambda: ctx.dev_hash_buffers(d, 4096, 4, dig, stride=2048),                        # 4 units do not fit
        lambda: ctx.dev_hash_buffers(d, 4096, 4, dig.view(-1)[1:], stride=1024),           # digests not 16-byte aligned
        lambda: ctx.dev_hash_buffers(d, 4096, 4, dig, d_offsets=d),                        # offsets without lengths
        lambda: ctx.dev_s3_etag(d, 4096, 0, dig),                                          # part size 0
        lambda: ctx.dev_synth_fill(d, 1, 0, 24, 10),                                       # file_len not a multiple of 16
        lambda: ctx.set_variant(1234),
        lambda: tpi.Context([99]),                                                         # no such device
        lambda: tpi.Context([0], staging_bytes=1000),                                      # staging below 1 MiB
    ]
    for i, f in enumerate(cases):
        with pytest.raises(tpi.TpiMd5Error) as e:
            f()
        assert e.value.code in (-1, -2, -6) and len(str(e.value)) > 20, i
    # and the context is still usable afterwards
    assert ctx.hash_buffers(b"abc", [0], [3])[0].tobytes().hex() == "900150983cd24fb0d6963f7d28e17f72"


def test_pinned_source_takes_the_direct_copy_path(ctx):
    """Pageable caller memory is packed through pinned staging; a pinned buffer (what bench.py's e2e leg
    and a cgo caller with C-allocated pinned memory pass) is copied straight from where it lies, for any
    alignment of its units.  Same digests either way."""
    import torch
    lengths = [int(x) for x in np.random.default_rng(76).integers(0, 30000, 2000)] + [(20 << 20) + 5, 0, 64]
    for align in (16, 1):
        buf, off, ln = oracle.synth_packed(76, lengths, align=align)
        pinned = torch.empty(buf.size, dtype=torch.uint8, pin_memory=True)
        pinned.numpy()[:] = buf
        want = oracle.md5_batch(buf, off, ln, threads=4)
        assert np.array_equal(ctx.hash_buffers(pinned.numpy(), off, ln), want)        # direct
        assert np.array_equal(ctx.hash_buffers(buf, off, ln), want)                   # packed


def test_streams_mixed_write_paths_randomised(ctx):
    """Small buffered writes, large zero-copy writes and grouped writes on the same streams, in random
    order, with Sum taken in between: always what hashlib gives for the bytes written so far."""
    rng = np.random.default_rng(77)
    n = 12
    streams = [ctx.stream() for _ in range(n)]
    ref = [hashlib.md5() for _ in range(n)]
    sizes = [0, 1, 63, 64, 65, 1000, 40000, (4 << 20) - 1, 4 << 20, (5 << 20) + 3]
    pool = rng.integers(0, 256, 6 << 20, dtype=np.uint8)
    for step in range(40):
        if rng.random() < 0.5:
            i = int(rng.integers(0, n))
            k = int(rng.choice(sizes))
            o = int(rng.integers(0, pool.size - k))
            streams[i].write(pool[o:o + k]); ref[i].update(pool[o:o + k].tobytes())
        else:
            who = sorted(rng.choice(n, int(rng.integers(1, n + 1)), replace=False).tolist())
            chunks = []
            for i in who:
                k = int(rng.choice(sizes[:8]))
                o = int(rng.integers(0, pool.size - k))
                chunks.append(pool[o:o + k]); ref[i].update(pool[o:o + k].tobytes())
            ctx.stream_write_many([streams[i] for i in who], chunks)
        if step % 5 == 4:
            j = int(rng.integers(0, n))
            assert streams[j].sum() == ref[j].digest(), (step, j)
    for s, r in zip(streams, ref):
        assert s.sum() == r.digest()
        s.close()


def test_fuzz_host_api_short():
    """Ten seconds of tools/fuzz_host_api.py (random layouts, alignments, staging sizes, pinned/pageable,
    files) against the oracle.  Longer runs: `python tools/fuzz_host_api.py 300 <seed>`."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "fuzz_host_api.py"), "10", "777"], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0 and "fuzz ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
