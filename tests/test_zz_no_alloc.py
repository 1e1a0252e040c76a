"""The allocation-free S3 ETag entry point (tpimd5_dev_s3_etag_noalloc), through the C-ABI.

Collected last on purpose (file name): with `-x` every other test has already run if this fails.

The entry point is for callers that cannot allocate device memory -- a checkpoint hook inside a
training process whose framework pool owns the GPU, or a host program held to a fixed device
memory budget.  It must give exactly the ETag the ordinary entry point gives (bit-exact, this is
byte work) and the CPU oracle gives, on back-to-back calls that share its one built-in scratch
array, and it must refuse sizes that do not fit that array.
"""
import numpy as np
import pytest

import oracle
import tpi_b200 as tpi

pytestmark = pytest.mark.gpu


def _etag(t):
    return bytes(t.cpu().numpy().tobytes())


def test_noalloc_etag_matches_the_ordinary_entry_point_twice_in_a_row():
    import torch
    dev = 0
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    with tpi.Context([dev]) as ctx:
        n_bytes = (48 << 20) + 12345
        data = torch.empty(n_bytes + 64, dtype=torch.uint8, device="cuda")
        ctx.dev_synth_fill(data, 9, 0, 1 << 20, 48, stream=stream)
        data[48 << 20:] = 7
        # two calls in a row on the same stream: first many small parts (fills a long stretch of
        # the shared array), then fewer, larger, ragged parts of an unaligned slice -- stale
        # digests from the first call must not leak into the second
        cases = [(data, n_bytes, 8192), (data[3:], (20 << 20) + 5, (1 << 20) + 16)]
        want, got = [], []
        for buf, size, part in cases:
            ref = torch.zeros(16, dtype=torch.uint8, device="cuda")
            ctx.dev_s3_etag(buf, size, part, ref, stream=stream)
            want.append(ref)
        for buf, size, part in cases:
            out = torch.zeros(16, dtype=torch.uint8, device="cuda")
            tpi.dev_s3_etag_noalloc(dev, buf, size, part, out, stream=stream)
            got.append(out)
        torch.cuda.synchronize()
        host = data.cpu().numpy()
        for (buf, size, part), w, g in zip(cases, want, got):
            assert _etag(w) == _etag(g)
            lo = 3 if buf is not data else 0
            o_etag, _, _ = oracle.s3_etag(host[lo: lo + size], part, threads=4, engine=oracle.ENGINE_OPENSSL
                                          if oracle.lib().orc_has_openssl() else oracle.ENGINE_RESTATED)
            assert _etag(g) == o_etag                     # not only CUDA against CUDA: the CPU oracle agrees
        # and once more, the first case again after the second
        again = torch.zeros(16, dtype=torch.uint8, device="cuda")
        tpi.dev_s3_etag_noalloc(dev, cases[0][0], cases[0][1], cases[0][2], again, stream=stream)
        torch.cuda.synchronize()
        assert _etag(again) == _etag(want[0])


def test_noalloc_edges_empty_object_and_exactly_the_maximum_number_of_parts():
    import torch
    data = torch.arange(0, 1 << 20, dtype=torch.int32, device="cuda").view(torch.uint8)[: 10000 * 100 + 37].contiguous()
    host = data.cpu().numpy()
    out = torch.zeros(16, dtype=torch.uint8, device="cuda")
    # size 0: S3 (and rclone) still upload one empty part
    tpi.dev_s3_etag_noalloc(0, data, 0, 5 << 20, out)
    torch.cuda.synchronize()
    assert _etag(out) == oracle.s3_etag(host[:0], 5 << 20)[0]
    # exactly 10,000 parts (the last one ragged), the most the built-in array holds ...
    size, part = 10000 * 100 - 63, 100
    assert tpi.s3_nparts(size, part) == tpi.NOALLOC_MAX_PARTS
    tpi.dev_s3_etag_noalloc(0, data, size, part, out)
    torch.cuda.synchronize()
    assert _etag(out) == oracle.s3_etag(host[:size], part)[0]
    # ... and one more is refused before anything is launched
    with pytest.raises(tpi.TpiMd5Error) as e:
        tpi.dev_s3_etag_noalloc(0, data, 10000 * 100 + 1, 100, out)
    assert e.value.code == -6


def test_noalloc_refuses_more_parts_than_its_array_holds():
    import torch
    data = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
    out = torch.zeros(16, dtype=torch.uint8, device="cuda")
    with pytest.raises(tpi.TpiMd5Error) as e:
        tpi.dev_s3_etag_noalloc(0, data, 1 << 20, 64, out)          # 16,384 parts > 10,000
    assert e.value.code == -6 and str(tpi.NOALLOC_MAX_PARTS) in str(e.value)
    tpi.dev_s3_etag_noalloc(0, data, 1 << 20, 128, out)             # 8,192 parts: fine
    torch.cuda.synchronize()


def test_noalloc_on_another_device_leaves_the_callers_device_current():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    torch.cuda.set_device(0)
    with torch.cuda.device(1):
        data = torch.full((3 << 20,), 5, dtype=torch.uint8, device="cuda")
        out = torch.zeros(16, dtype=torch.uint8, device="cuda")
    tpi.dev_s3_etag_noalloc(1, data, data.numel(), 1 << 20, out)      # called with device 0 current
    assert torch.cuda.current_device() == 0                            # cudaGetDevice: the thread is back on its own GPU
    torch.cuda.synchronize(1)
    assert _etag(out) == oracle.s3_etag(np.full(3 << 20, 5, dtype=np.uint8), 1 << 20)[0]
