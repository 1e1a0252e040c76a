#!/usr/bin/env python3
"""Register-only MD5 probe (tpimd5_dev_int_probe): cycles per MD5 step per scheduler, per step formulation and occupancy.

    python tools/probe_once.py                 # sweep: modes x CTAs/SM (4 warps each), JSON lines
    python tools/probe_once.py MODE CTAS ITERS # one launch (for ncu)
ctas_per_sm = 1 is one warp per SM sub-partition: the pure dependent-chain latency (12 cycles per step = three
dependent 4-cycle ops is the floor); 8 = 32 warps/SM: the integer-issue ceiling."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import tpi_b200 as tpi  # noqa: E402
from bench import ClockSampler  # noqa: E402

torch.cuda.set_device(0)
ctx = tpi.Context([0])
out = torch.empty(148 * 16 * 128 * 16, dtype=torch.uint8, device="cuda")
stream = torch.cuda.current_stream().cuda_stream
if len(sys.argv) >= 4:
    mode, ctas, iters = (int(x) for x in sys.argv[1:4])
    ctx.dev_int_probe(mode, ctas, iters, out, stream=stream)
    torch.cuda.synchronize()
    sys.exit(0)
modes = [int(x) for x in os.environ.get("PROBE_MODES", "0,4,5,6,7,8").split(",")]
for mode in modes:
    for ctas in (1, 2, 3, 4, 8):
        ctx.dev_int_probe(mode, ctas, 256, out, stream=stream)
        iters = 16384 if ctas <= 2 else 8192
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        with ClockSampler(0, 0.005) as cs:
            e0.record()
            blocks = 0
            for _ in range(6):
                blocks += ctx.dev_int_probe(mode, ctas, iters, out, stream=stream)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        mhz = cs.summary()["sm_mhz"] or 0.0
        # every warp runs 6*iters blocks of 64 steps back to back
        cyc_step_warp = ms * 1e-3 * mhz * 1e6 / (6 * iters * 64) if mhz else None
        print(json.dumps({"probe_mode": mode, "warps_per_smsp": ctas, "ms": round(ms, 3), "sm_mhz": mhz,
                          "GBps_equiv": round(blocks * 64 / ms / 1e6, 1),
                          "cycles_per_step_per_warp": None if cyc_step_warp is None else round(cyc_step_warp, 3),
                          "cycles_per_step_per_smsp": None if cyc_step_warp is None else round(cyc_step_warp / ctas, 3)}), flush=True)
ctx.close()
