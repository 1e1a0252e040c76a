"""Large units and BASELINE config 4 at a size that holds multi-GiB files, through the C-ABI, against the oracle.

These are the cases round 1 left to tools: units above 4 GiB (byte counts past 2^32) and above the longest
tensor-map row, the Zipf workdir of config 4 scaled so that its largest file is 8 GiB, a slice of the
device-API fuzz, and every kernel instantiation the library picks by launch size run thirty times over.
"""
import hashlib
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

import oracle
import tpi_b200 as tpi

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENG = oracle.ENGINE_OPENSSL if oracle.lib().orc_has_openssl() else oracle.ENGINE_RESTATED


def _free_gib():
    import torch
    free, _ = torch.cuda.mem_get_info()
    return free / (1 << 30)


def test_long_units_continue_from_launch_to_launch(monkeypatch):
    """A unit longer than one tensor-map row is hashed as a run of launches that hand the chaining value on
    through the digest table.  The row limit (2 GiB) and the segment size (1 GiB) are shrunk through the
    environment so that every boundary case runs on kilobytes: unit = exactly k segments, one byte more or
    less, padding block spilling into a segment of its own, unit counts that leave warps half empty, and a
    buffer that ends inside its last unit."""
    import torch
    monkeypatch.setenv("TPIMD5_ROW_MAX", "4096")
    monkeypatch.setenv("TPIMD5_CHAIN_SEG", "1024")
    with tpi.Context([0]) as c:
        cases = [(4096 + 16, 70), (5 * 1024, 33), (5 * 1024 + 16, 1), (8192 - 16, 31), (8192 + 64, 3), (65536 + 48, 40),
                 (1 << 20, 2), (3 * 1024 + 1024 - 16, 100)]
        for stride, n in cases:
            for cut in (0, 7, stride - 9, 1024 + 55, 1024 + 56):
                nbytes = stride * n - cut
                d = torch.empty(stride * n, dtype=torch.uint8, device="cuda")
                c.dev_synth_fill(d, 91, 0, 16, d.numel() // 16)
                dig = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
                c.dev_hash_buffers(d, nbytes, n, dig, stride=stride)
                torch.cuda.synchronize()
                host = d.cpu().numpy()
                off = np.arange(n, dtype=np.uint64) * stride
                ln = np.minimum(np.uint64(stride), np.uint64(nbytes) - off).astype(np.uint64)
                want = oracle.md5_batch(host, off, ln, threads=4)
                bad = np.nonzero((dig.cpu().numpy() != want).any(axis=1))[0]
                assert bad.size == 0, (stride, n, cut, bad[:5])
        # S3 parts longer than a row, and the digest of digests behind them
        d = torch.empty(40 * 5008 + 300, dtype=torch.uint8, device="cuda")
        c.dev_synth_fill(d[: d.numel() // 16 * 16], 92, 0, 16, d.numel() // 16)
        etag = torch.zeros(16, dtype=torch.uint8, device="cuda")
        c.dev_s3_etag(d, d.numel(), 5008, etag)
        torch.cuda.synchronize()
        assert etag.cpu().numpy().tobytes() == oracle.s3_etag(d.cpu().numpy(), 5008)[0]


def _zipf_sizes(top, s_exp=1.111, seed=4):
    n = int(np.ceil((top / 1024) ** (1 / s_exp)))
    k = np.arange(1, n + 1, dtype=np.float64)
    sizes = np.maximum(1024, np.floor(top * k ** (-s_exp))).astype(np.int64)
    np.random.default_rng(seed).shuffle(sizes)
    return sizes


def test_config4_zipf_with_multi_gib_files_both_apis():
    """BASELINE config 4 at a quarter of its size: rank-size Zipf from 8 GiB down to 1 KiB (about 1.7 M files,
    64 GiB), whole-file content-MD5, resident in HBM.  Device API: the files of 256 MiB and more go in as
    single strided units on streams of their own (the 8 GiB and 3.7 GiB ones are longer than a tensor-map
    row: segments), everything else as one ragged batch.  Host API, at the same time from another thread:
    the 8 GiB file out of pageable host memory (byte count above 2^32, chained across dozens of staging
    buffers) and a sample of small files.  Every digest is checked against the oracle's OpenSSL engine."""
    import torch
    if _free_gib() < 90:
        pytest.skip("needs 90 GiB of free HBM")
    sizes = _zipf_sizes(8 << 30)
    assert sizes.max() == 8 << 30 and (sizes > (4 << 30)).sum() == 1 and sizes.sum() > (60 << 30)
    off = np.concatenate([[0], np.cumsum((sizes + 15) // 16 * 16)[:-1]]).astype(np.int64)
    total = int(off[-1] + (sizes[-1] + 15) // 16 * 16)
    with tpi.Context([0]) as c:
        data = torch.empty(total, dtype=torch.uint8, device="cuda")
        c.dev_synth_fill(data, 4, 0, 16, total // 16)
        torch.cuda.synchronize()
        big = np.nonzero(sizes >= (256 << 20))[0]
        small = np.nonzero(sizes < (256 << 20))[0]
        dig = torch.zeros((sizes.size, 16), dtype=torch.uint8, device="cuda")
        # host side of the 8 GiB file, fetched before the timed work so both APIs hash at the same time
        top = int(np.argmax(sizes))
        top_host = data[int(off[top]): int(off[top]) + int(sizes[top])].cpu().numpy()
        host_result = {}

        def host_api():
            try:
                host_result["top"] = c.hash_buffers(top_host, [0], [top_host.size])[0].tobytes()
            except Exception as e:                                   # surfaced by the assert below
                host_result["error"] = e
        th = threading.Thread(target=host_api)
        th.start()
        streams = [torch.cuda.Stream() for _ in big]
        for st, i in zip(streams, big.tolist()):
            ln = int(sizes[i])
            c.dev_hash_buffers(data[int(off[i]):], ln, 1, dig[i:i + 1], stride=(ln + 15) // 16 * 16, stream=st.cuda_stream)
        d_off, d_len = torch.from_numpy(off[small]).cuda(), torch.from_numpy(sizes[small]).cuda()
        dig_small = torch.zeros((small.size, 16), dtype=torch.uint8, device="cuda")
        c.dev_hash_buffers(data, total, small.size, dig_small, d_offsets=d_off, d_lengths=d_len)
        torch.cuda.synchronize()
        th.join()
        assert "error" not in host_result, host_result.get("error")
        got_small = dig_small.cpu().numpy()
        got_big = dig.cpu().numpy()
        # small files against the oracle, 4 GiB of the buffer at a time
        order = small[np.argsort(off[small])]
        pos = {int(i): k for k, i in enumerate(small.tolist())}
        lo = 0
        while lo < order.size:
            hi = lo
            base = int(off[order[lo]])
            while hi < order.size and int(off[order[hi]]) + int(sizes[order[hi]]) - base <= (4 << 30):
                hi += 1
            chunk = data[base: int(off[order[hi - 1]]) + int(sizes[order[hi - 1]])].cpu().numpy()
            want = oracle.md5_batch(chunk, (off[order[lo:hi]] - base).astype(np.uint64), sizes[order[lo:hi]].astype(np.uint64),
                                    threads=os.cpu_count() or 4, engine=ENG)
            got = got_small[[pos[int(i)] for i in order[lo:hi]]]
            bad = np.nonzero((got != want).any(axis=1))[0]
            assert bad.size == 0, (lo, bad[:5], sizes[order[lo:hi]][bad[:5]])
            lo = hi
        # the big ones (hashlib releases the GIL: all of them at once, bounded by the 8 GiB file)
        from concurrent.futures import ThreadPoolExecutor

        def cpu_md5(i):
            host = top_host if i == top else data[int(off[i]): int(off[i]) + int(sizes[i])].cpu().numpy()
            return hashlib.md5(host).digest()
        with ThreadPoolExecutor(8) as ex:
            wants = list(ex.map(cpu_md5, big.tolist()))
        for i, want in zip(big.tolist(), wants):
            assert got_big[i].tobytes() == want, ("device API", i, int(sizes[i]))
            if i == top:
                assert host_result["top"] == want, "host API, 8 GiB unit"
        # host API over a sample of the small files, out of pageable memory
        sample = order[:: max(1, order.size // 100000)]
        base = int(off[sample[0]])
        span = data[base: int(off[sample[-1]]) + int(sizes[sample[-1]])].cpu().numpy() if \
            int(off[sample[-1]]) + int(sizes[sample[-1]]) - base <= (8 << 30) else None
        if span is not None:
            got = c.hash_buffers(span, (off[sample] - base).astype(np.uint64), sizes[sample].astype(np.uint64))
            assert np.array_equal(got, got_small[[pos[int(i)] for i in sample]])


def test_fuzz_device_api_short():
    """Ten seconds of tools/fuzz_device_api.py (strided units through the TMA kernels, ragged batches below and
    above the bucketing threshold, S3 ETags through all three device entry points) against the oracle."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_device_api.py"), "10", "909"], capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0 and "fuzz ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("n", [3000, 9000, 25000, 40000, 200000])
def test_every_kernel_the_library_picks_is_deterministic(n):
    """compute-sanitizer is closed on this pool, so shared-memory and barrier races are hunted the blunt way:
    thirty identical launches per kernel the library chooses by launch size, all digests identical to the
    oracle's every time.  3,000 / 9,000 / 25,000 / 40,000 units take the latency kernel in its four shapes
    (2, 4 and 3 hashing warps per CTA in step mode 4; 4 warps in mode 6), 200,000 the throughput kernel."""
    import torch
    with tpi.Context([0]) as c:
        flen = 4096 + 48 if n > 50000 else 16384 + 48
        d = torch.empty(n * flen, dtype=torch.uint8, device="cuda")
        c.dev_synth_fill(d, 74, 0, 16, n * flen // 16)
        host = d.cpu().numpy()
        off = np.arange(n, dtype=np.uint64) * flen
        want = torch.from_numpy(oracle.md5_batch(host, off, np.full(n, flen, dtype=np.uint64), threads=os.cpu_count() or 4,
                                                 engine=ENG)).cuda()
        out = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
        for rep in range(30):
            out.zero_()
            c.dev_hash_buffers(d, n * flen, n, out, stride=flen)
            assert torch.equal(out, want), (n, rep)
