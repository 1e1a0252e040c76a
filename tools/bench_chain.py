#!/usr/bin/env python3
"""Chain-latency sweep: launches with too few MD5 chains to fill the SMs (S3 parts, large files), per kernel variant.

For every (units, unit bytes) shape and kernel variant: ms per launch, GB/s, and SM cycles per 64-byte block of
one chain (ms x SM clock / blocks per unit; 768 = three dependent ops x 4 cycles x 64 steps is the floor).
    python tools/bench_chain.py [--variants 0,341,...] [--shapes 8192x1048576,...]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import tpi_b200 as tpi  # noqa: E402
from bench import ClockSampler  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants", default="0,241,245,246,341,342,343,344,345,346,347,348,361,363")
    ap.add_argument("--shapes", default="8192x1048576,16384x1048576,25600x524288,1x131072,32x1048576,40000x262144")
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    variants = [int(v) for v in args.variants.split(",")]
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    ctx = tpi.Context([0])
    stream = torch.cuda.current_stream().cuda_stream
    for shape in args.shapes.split(","):
        n, ulen = (int(x) for x in shape.split("x"))
        data = torch.empty(n * ulen, dtype=torch.uint8, device=dev)
        ctx.dev_synth_fill(data, 11, 0, 1 << 16, n * ulen >> 16, stream=stream)
        ref = None
        for v in variants:
            ctx.set_variant(v)
            dig = torch.zeros((n, 16), dtype=torch.uint8, device=dev)
            for _ in range(2):
                ctx.dev_hash_buffers(data, n * ulen, n, dig, stride=ulen, stream=stream)
            torch.cuda.synchronize()
            if ref is None:
                ref = dig.clone()
            same = bool(torch.equal(ref, dig))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with ClockSampler(0, 0.002) as cs:
                e0.record()
                for _ in range(args.reps):
                    ctx.dev_hash_buffers(data, n * ulen, n, dig, stride=ulen, stream=stream)
                e1.record()
                torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.reps
            mhz = cs.summary()["sm_mhz"] or 0.0
            blocks = ((ulen + 8) >> 6) + 1
            print(json.dumps({"units": n, "unit_bytes": ulen, "variant": v, "ms": round(ms, 4),
                              "GBps": round(n * ulen / ms / 1e6, 1), "sm_mhz": mhz,
                              "cycles_per_block": round(ms * 1e-3 * mhz * 1e6 / blocks, 1) if mhz else None,
                              "agrees_with_first_variant": same}), flush=True)
        del data
    ctx.close()


if __name__ == "__main__":
    main()
