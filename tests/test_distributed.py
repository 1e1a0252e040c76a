"""world_size-2 run of the one-process-per-GPU layout on CPU (gloo): sharding + digest gather.

The hash itself needs a B200, so each rank's `hasher` here is a test double answering from the
oracle; what is under test is the product's host logic in sharding.py (who hashes what, every unit
exactly once, manifest reassembled in listing order on every rank)."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

import oracle


class OracleHasher:
    def hash_buffers(self, base, off, ln):
        return oracle.md5_batch(base, off, ln)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, lengths, out_dir):
    import importlib
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sharding = importlib.import_module("terraform-provider-iterative_b200.sharding")
        buf, off, ln = oracle.synth_packed(31, lengths, align=16)
        units = sharding.my_units(ln, rank, world)
        manifest = sharding.hash_sharded(OracleHasher(), buf, off, ln, rank, world)
        np.save(os.path.join(out_dir, f"manifest{rank}.npy"), manifest)
        np.save(os.path.join(out_dir, f"units{rank}.npy"), units)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_two_ranks_shard_and_gather(tmp_path):
    rng = np.random.default_rng(6)
    lengths = [int(x) for x in (rng.pareto(1.1, 400) * 2000).astype(np.int64)] + [0, 0, 1 << 20, 5]
    port = _free_port()
    mp.spawn(_worker, args=(2, port, lengths, str(tmp_path)), nprocs=2, join=True)
    buf, off, ln = oracle.synth_packed(31, lengths, align=16)
    want = oracle.md5_batch(buf, off, ln)
    m0, m1 = np.load(tmp_path / "manifest0.npy"), np.load(tmp_path / "manifest1.npy")
    assert np.array_equal(m0, want) and np.array_equal(m1, want)
    u0, u1 = np.load(tmp_path / "units0.npy"), np.load(tmp_path / "units1.npy")
    assert np.array_equal(np.sort(np.concatenate([u0, u1])), np.arange(len(lengths)))
    b0, b1 = float(ln[u0].sum()), float(ln[u1].sum())
    assert abs(b0 - b1) <= float(ln.max()) + 72 * len(lengths)          # byte balanced up to one unit
