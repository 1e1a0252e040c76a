"""The allocation-free S3 ETag entry point (tpimd5_dev_s3_etag_noalloc), through the C-ABI.

Collected last on purpose (file name): with `-x` every other test has already run if this fails.

The entry point is for callers that cannot allocate device memory -- a checkpoint hook inside a
training process whose framework pool owns the GPU, or a host program held to a fixed device
memory budget.  It must give exactly the ETag the ordinary entry point gives (bit-exact, this is
byte work), on back-to-back calls that share its one built-in scratch array, and it must refuse
sizes that do not fit that array.
"""
import numpy as np
import pytest

import tpi_b200 as tpi

pytestmark = pytest.mark.gpu


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
        for w, g in zip(want, got):
            assert np.array_equal(w.cpu().numpy(), g.cpu().numpy())
            assert w.cpu().numpy().any()
        # and once more, the first case again after the second
        again = torch.zeros(16, dtype=torch.uint8, device="cuda")
        tpi.dev_s3_etag_noalloc(dev, cases[0][0], cases[0][1], cases[0][2], again, stream=stream)
        torch.cuda.synchronize()
        assert np.array_equal(again.cpu().numpy(), want[0].cpu().numpy())


def test_noalloc_refuses_more_parts_than_its_array_holds():
    import torch
    data = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
    out = torch.zeros(16, dtype=torch.uint8, device="cuda")
    with pytest.raises(tpi.TpiMd5Error) as e:
        tpi.dev_s3_etag_noalloc(0, data, 1 << 20, 64, out)          # 16,384 parts > 10,000
    assert e.value.code == -6 and str(tpi.NOALLOC_MAX_PARTS) in str(e.value)
    tpi.dev_s3_etag_noalloc(0, data, 1 << 20, 128, out)             # 8,192 parts: fine
    torch.cuda.synchronize()
