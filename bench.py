#!/usr/bin/env python3
"""bench.py -- workdir checksum throughput (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path over one batch of synthetic input.  The headline workload is
BASELINE.json configs[1] ("C2"): a workdir of 1,048,576 files x 4 KiB per GPU, one MD5 per file
(the manifest rclone's checkers would build one file at a time).  Units are independent, so at
N GPUs every rank hashes its own shard of N x 1 Mi files with no data-path collective (weak
scaling); torch.distributed only provides the barrier and the max-over-ranks of the timings.

The one JSON line carries, next to the C2 headline:
  value      C2 kernel-only GB/s (10^9 bytes of file content per second), inputs resident in HBM,
             timed with CUDA events on the launching stream
  e2e        the same metric through the host-pointer C-ABI call a cgo caller makes
             (tpimd5_hash_buffers): pinned host buffer -> H2D -> kernel -> D2H of the digests,
             all inside the timed region
  roofline   the C2 kernel against measured HBM copy bandwidth (MEASURED_PEAKS.json), the
             register-only integer-issue ceiling measured in the same run, and `sustained`: the
             same launch repeated for >= 2 s (the power-capped regime a 100 GiB workdir runs in)
  cpu_baseline  OpenSSL MD5 (stand-in for Go crypto/md5; Go is not installed on this image)
             driven like rclone's per-file hashing on the box's host cores, bounded sample
  c3         BASELINE configs[2]: S3 multipart ETag of 8 x 64 GiB files, 8 MiB parts, the 8 files
             spread over the N ranks (STRONG scaling: total work fixed), kernel-only
  c5         configs[4]: 100 GiB manifest rebuild + digest diff, 5 % mutated, spread over the ranks
  c4         configs[3] scaled 1/64: Zipf sizes, whole-file MD5, byte-balanced LPT sharding
  c1         (N=1) configs[0]: 1,000 files / 1 GiB on tmpfs through machine.Transfer --checksum
  c2_files   (N=1) C2 from REAL files: 200,000 x 4 KiB files through tpimd5_hash_files
  c3_file    (N=1) C3 from a REAL file: one 8 GiB checkpoint on tmpfs through tpimd5_s3_etag_file
each with its own value, ms_per_step, roofline (for chain-starved workloads also the chain-latency
bound resident_chains x 64 B x f_clk / 768 cycles) and, at N=1, a CPU figure beside it.
`--impl reference` times the CPU path alone over the full C2 file set, all host threads, and
prints the same line shape with the same `config`.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Workdir checksum GB/s (MD5 + S3 ETag) at 1/2/4/8 B200 vs Go crypto/md5"
FILE_BYTES = 4096
FILES_PER_GPU = 1 << 20
SEED = 2
WORKLOAD = "C2: 1,048,576 files x 4 KiB per GPU, one-lane-per-file MD5 manifest"
CHAIN_FLOOR_CYCLES = 768          # 64 steps x three dependent 4-cycle integer ops (LOP3 -> IADD3 -> LEA.HI)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--variant", type=int, default=int(os.environ.get("TPIMD5_VARIANT", "0")))
    ap.add_argument("--files", type=int, default=FILES_PER_GPU, help="files per GPU (default: the C2 size)")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--sub-steps", type=int, default=2, help="timed steps of the c3/c5 sub-records")
    ap.add_argument("--sustained-s", type=float, default=2.5, help="length of the sustained C2 loop (0 = skip)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--only-c2", action="store_true", help="skip the c1/c3/c4/c5/c2_files sub-records")
    ap.add_argument("--sweep", action="store_true", help="print kernel-only GB/s for every kernel variant, then exit")
    ap.add_argument("--workload", default="all", choices=["all", "c1", "c2", "c2_files", "c3", "c3_file", "c4", "c5"],
                    help="all = the driver's line (C2 headline + every sub-record); c2 = headline only; "
                         "any other = that sub-record alone, printed as its own line")
    return ap.parse_args()


def c2_config(args):
    """The `config` of the headline line; --impl reference prints the identical dict."""
    return {"workload": WORKLOAD, "file_bytes": FILE_BYTES, "files_per_gpu": args.files,
            "bytes_per_gpu_per_step": args.files * FILE_BYTES,
            "l2": "4 GiB of input per step, far larger than the 126 MB L2 (no flush needed)",
            "kernel_variant": args.variant, "sharding": "independent shard per rank, no collective"}


# ------------------------------------------------------------------------------------------------
# clocks during the timed region (NVML; falls back to nvidia-smi)
# ------------------------------------------------------------------------------------------------
REASON_BITS = {
    0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
    0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
    0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
}


class ClockSampler:
    def __init__(self, index: int, period_s: float = 0.05):
        self.index, self.period = index, period_s
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nvml = None

    def _sample(self):
        if self._nvml is not None:
            p = self._nvml
            try:
                self.samples.append(float(p.nvmlDeviceGetClockInfo(self._h, p.NVML_CLOCK_SM)))
                try:
                    mask = p.nvmlDeviceGetCurrentClocksEventReasons(self._h)
                except Exception:
                    mask = p.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                for bit, name in REASON_BITS.items():
                    if mask & bit and name != "gpu_idle":
                        self.reasons.add(name)
                return
            except Exception:
                pass
        try:
            import subprocess
            out = subprocess.check_output(
                ["nvidia-smi", f"--id={self.index}", "--query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.active",
                 "--format=csv,noheader,nounits"], timeout=5).decode().strip().split(",")
            self.samples.append(float(out[0]))
            self.max_mhz = float(out[1])
            mask = int(out[2].strip(), 16)
            for bit, name in REASON_BITS.items():
                if mask & bit and name != "gpu_idle":
                    self.reasons.add(name)
        except Exception:
            pass

    def __enter__(self):
        def run():
            while not self._stop.is_set():
                self._sample()
                self._stop.wait(self.period)
        self._thread = threading.Thread(target=run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._thread.join()

    def summary(self):
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle's OpenSSL engine on all host threads
# ------------------------------------------------------------------------------------------------
def cpu_engine():
    import oracle
    if oracle.lib().orc_has_openssl():
        return oracle.ENGINE_OPENSSL, "OpenSSL MD5 (asm)"
    return oracle.ENGINE_RESTATED, "restated C MD5"


def cpu_pass(sample, off, ln, threads):
    import oracle
    eng, _ = cpu_engine()
    t0 = time.perf_counter()
    oracle.md5_batch(sample, off, ln, threads=threads, engine=eng)
    return time.perf_counter() - t0


def cpu_baseline(budget_s=12.0, sample_files=1 << 16, host_buf=None):
    import oracle
    threads = os.cpu_count() or 1
    if host_buf is not None:
        sample = host_buf[: sample_files * FILE_BYTES]
    else:
        sample = oracle.synth_files(SEED, 0, FILE_BYTES, sample_files, threads=threads)
    off = np.arange(sample_files, dtype=np.uint64) * FILE_BYTES
    ln = np.full(sample_files, FILE_BYTES, dtype=np.uint64)
    cpu_pass(sample, off, ln, threads)                      # warm (page in, spin up)
    passes, spent = 0, 0.0
    while spent < budget_s and passes < 10000:
        spent += cpu_pass(sample, off, ln, threads)
        passes += 1
    gbs = passes * sample.size / spent / 1e9
    return {"value": round(gbs, 3), "unit": "GB/s", "cores": threads, "kind": "port",
            "sample": f"{sample_files} files x {FILE_BYTES} B of the C2 workdir ({sample.size >> 20} MiB), {passes} passes "
                      f"in {spent:.1f} s, {cpu_engine()[1]}, one file per task on {threads} threads; Go crypto/md5 + rclone "
                      f"not runnable here (no Go toolchain)"}


def run_reference(args, rank, world):
    """--impl reference: the CPU implementation of the path over the FULL C2 file set (the same
    1,048,576 x 4 KiB workdir the GPU arm hashes per GPU), all host threads, rank 0 only."""
    if rank != 0:
        return
    import oracle
    threads = os.cpu_count() or 1
    n = args.files
    data = oracle.synth_files(SEED, 0, FILE_BYTES, n, threads=threads)
    off = np.arange(n, dtype=np.uint64) * FILE_BYTES
    ln = np.full(n, FILE_BYTES, dtype=np.uint64)
    for _ in range(max(1, args.warmup)):
        cpu_pass(data, off, ln, threads)
    t_total = 0.0
    for _ in range(args.steps):
        t_total += cpu_pass(data, off, ln, threads)
    gbs = args.steps * data.size / t_total / 1e9
    sample_txt = (f"every step one full pass over the {n:,} x {FILE_BYTES} B C2 workdir ({data.size >> 20} MiB) held in host "
                  f"memory, {cpu_engine()[1]}, one file per task on {threads} host threads; stand-in for Go crypto/md5 + rclone, "
                  f"which cannot run here (no Go toolchain)")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbs, 3), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": max(1, args.warmup), "ms_per_step": round(1e3 * t_total / args.steps, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": c2_config(args),
        "cpu_baseline": {"value": round(gbs, 3), "unit": "GB/s", "cores": threads, "kind": "port", "sample": sample_txt},
        "e2e": {"value": round(gbs, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def bind_near_gpu(index: int):
    """Run this rank's host threads (and first-touch its pinned buffers) on the CPUs next to its GPU:
    at 8 ranks the H2D copies otherwise cross the socket interconnect.  Best effort; returns the CPU list."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [64 * w + b for w, mask in enumerate(words) for b in range(64) if (mask >> b) & 1 and 64 * w + b < ncpu]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if len(allowed) >= 2:
            os.sched_setaffinity(0, allowed)
            return allowed
    except Exception:
        pass
    return None


class Env:
    """What every section needs: the device, the context, the ranks and how to time across them."""

    def __init__(self, args, torch, dist, tpi, ctx, dev, rank, local_rank, world):
        self.args, self.torch, self.dist, self.tpi, self.ctx, self.dev = args, torch, dist, tpi, ctx, dev
        self.rank, self.local_rank, self.world = rank, local_rank, world
        self.stream = torch.cuda.current_stream().cuda_stream
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            self.peak, self.peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
        else:
            self.peak, self.peak_src = 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, step, steps, warmup, period=0.002):
        """W warm-up steps, then K steps between a barrier + synchronize on both sides, CUDA events on
        the launching stream; -> (ms per step as the max over ranks, this rank's own, clocks, launches)."""
        torch = self.torch
        for _ in range(warmup):
            step()
        self.barrier()
        l0 = self.ctx.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ClockSampler(self.local_rank, period) as clocks:
            e0.record()
            for _ in range(steps):
                step()
            e1.record()
            self.barrier()
        local = e0.elapsed_time(e1) / steps
        return self.max_over_ranks(local), local, clocks.summary(), int(self.ctx.launch_count() - l0)

    def free(self):
        self.torch.cuda.synchronize()
        self.torch.cuda.empty_cache()


def chain_roofline(env, per_gpu_gbps, resident_chains, blocks_per_chain_per_step, ms_local, clocks):
    """Roofline record of a chain-starved workload: against HBM as always, and against the bound that
    actually binds -- resident chains x 64 B per CHAIN_FLOOR_CYCLES at the clock seen."""
    mhz = clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0
    chain_bound = resident_chains * 64.0 * mhz * 1e6 / CHAIN_FLOOR_CYCLES / 1e9
    cyc = ms_local * 1e-3 * mhz * 1e6 / blocks_per_chain_per_step
    return {"bound": "hbm", "achieved": round(per_gpu_gbps, 1), "peak": env.peak, "unit": "GB/s",
            "frac": round(per_gpu_gbps / env.peak, 4), "traffic": None, "peak_source": env.peak_src,
            "binding": "chain latency: too few independent MD5 chains to fill the SMs",
            "resident_chains_per_gpu": int(resident_chains),
            "chain_latency_bound_GBps": round(chain_bound, 1), "frac_of_chain_latency_bound": round(per_gpu_gbps / chain_bound, 4),
            "cycles_per_block": round(cyc, 1), "cycles_per_block_floor": CHAIN_FLOOR_CYCLES}


def run_c2(env):
    """The headline: kernel-only C2, the sustained loop, end to end through the host-pointer call."""
    args, torch, ctx, dev = env.args, env.torch, env.ctx, env.dev
    n, flen = args.files, FILE_BYTES
    nbytes = n * flen
    data = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    digests = torch.empty((n, 16), dtype=torch.uint8, device=dev)
    ctx.dev_synth_fill(data, SEED, env.rank * n, flen, n, stream=env.stream)       # this rank's shard of the workdir
    torch.cuda.synchronize()

    def step():
        ctx.dev_hash_buffers(data, nbytes, n, digests, stride=flen, stream=env.stream)

    # ---- integer-issue ceiling: the same 64-step body, registers only ----------------------------
    # best over the step formulations at full occupancy = what the SM's integer pipes can do at most;
    # taken before the timed steps, on a GPU that is not yet held back by the power cap
    probe_out = torch.empty(148 * 16 * 128 * 16, dtype=torch.uint8, device=dev)
    int_peak, int_peak_mode = 0.0, None
    for mode in (4, 5, 6):
        ctx.dev_int_probe(mode, 8, 64, probe_out, stream=env.stream)
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        probe_blocks = ctx.dev_int_probe(mode, 8, 2048, probe_out, stream=env.stream)
        p1.record()
        torch.cuda.synchronize()
        g = probe_blocks * 64 / (p0.elapsed_time(p1) * 1e-3) / 1e9
        if g > int_peak:
            int_peak, int_peak_mode = g, mode

    # ---- kernel-only: W warm-up + K timed steps; 4 GiB of input per step >> 126 MB of L2 ---------
    args.warmup = max(args.warmup, 3)          # timing rule: at least 3 warm-up steps
    ms_step, ms_local, clocks, launches = env.timed(step, args.steps, args.warmup)
    value = env.world * nbytes / (ms_step * 1e-3) / 1e9
    achieved = nbytes / (ms_local * 1e-3) / 1e9

    # ---- sustained: the same launch back to back for seconds (board power cap regime) ------------
    sustained = None
    if args.sustained_s > 0:
        k = max(args.steps, int(args.sustained_s * 1e3 / ms_local) + 1)
        s_ms, s_local, s_clocks, _ = env.timed(step, k, 0, period=0.02)
        s_ach = nbytes / (s_local * 1e-3) / 1e9
        sustained = {"achieved": round(s_ach, 1), "frac": round(s_ach / env.peak, 4), "unit": "GB/s", "steps": k,
                     "seconds": round(k * s_local * 1e-3, 2), "ms_per_step": round(s_ms, 4),
                     "value_all_gpus": round(env.world * nbytes / (s_ms * 1e-3) / 1e9, 1), "clocks": s_clocks}

    # ---- end to end through the host-pointer entry point -----------------------------------------
    e2e, host_np = None, None
    near = getattr(env, "near", None)
    if not args.no_e2e:
        host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        host.copy_(data)
        torch.cuda.synchronize()
        host_np = host.numpy()
        off = np.arange(n, dtype=np.uint64) * flen
        ln = np.full(n, flen, dtype=np.uint64)
        out = np.empty((n, 16), dtype=np.uint8)            # reused across calls, as a caller's digest slice would be
        for _ in range(2):
            ctx.hash_buffers(host_np, off, ln, out=out)
        if not np.array_equal(out, digests.cpu().numpy()):
            raise SystemExit("bench.py: end-to-end digests differ from the kernel-only digests")
        # ceiling of this leg: the same pinned buffer copied to the device and nothing else (all ranks
        # at once, as in the timed region: GPUs may share PCIe uplinks and the host's memory bandwidth)
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        scratch = torch.empty_like(data)
        scratch.copy_(host, non_blocking=True)
        env.barrier()
        c0.record()
        for _ in range(3):
            scratch.copy_(host, non_blocking=True)
        c1.record()
        torch.cuda.synchronize()
        h2d_gbps = 3 * nbytes / (c0.elapsed_time(c1) * 1e-3) / 1e9
        del scratch
        env.barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            ctx.hash_buffers(host_np, off, ln, out=out)
        env.barrier()
        dt = env.max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": round(env.world * nbytes * args.e2e_steps / dt / 1e9, 3), "unit": "GB/s",
               "h2d_bytes_per_step": int(nbytes + n * 32), "d2h_bytes_per_step": int(n * 16),
               "steps": args.e2e_steps, "ms_per_step": round(1e3 * dt / args.e2e_steps, 3),
               "h2d_ceiling_GBps_per_gpu": round(h2d_gbps, 2), "host_cpus_bound_near_gpu": (len(near) if near else None),
               "frac_of_h2d_ceiling": round(nbytes * args.e2e_steps / dt / 1e9 / h2d_gbps, 3),
               "api": "tpimd5_hash_buffers (host pointers, pinned source, H2D + kernel + D2H inside the timed region)"}

    # ---- CPU baseline beside it (rank 0, N=1 only) ------------------------------------------------
    cpu = None
    if env.rank == 0 and env.world == 1 and not args.no_cpu:
        cpu = cpu_baseline(host_buf=host_np)

    # DRAM traffic per launch is an ncu number (one --set full capture); it is only quoted when the
    # capture on file is of this kernel at this size, otherwise null
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            t = json.load(open(tpath))
            if t.get("kernel") == "md5_chain_tma_kernel<4,6,2,1,6,2>" and t.get("algorithmic_bytes_per_launch") == nbytes and \
                    t.get("kernel_variant", 0) == args.variant:
                traffic = t.get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": "GB/s", "n_gpus": env.world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": c2_config(args),
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": env.peak, "unit": "GB/s",
                     "frac": round(achieved / env.peak, 4), "traffic": traffic, "peak_source": env.peak_src,
                     "kernel": "md5_chain_tma_kernel<4,6,2,1,6,2> (K1, one lane per file, producer warp + TMA)", "algorithmic_bytes_per_launch": nbytes,
                     "int_issue_peak_GBps": round(int_peak, 1), "frac_of_int_issue": round(achieved / int_peak, 4),
                     "int_issue_source": f"register-only probe of the same 64-step MD5 body (best step formulation: mode {int_peak_mode}), 32 warps/SM, same run",
                     "sustained": sustained},
        "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    del data, digests
    env.free()
    return line


def run_c3(env):
    """BASELINE config 3: 8 files x 64 GiB, S3 multipart ETag with 8 MiB parts, kernel-only.  The 8
    files are dealt to the ranks (strong scaling).  One file is only 8,192 chains, so a rank with two
    or more files keeps two resident (128 of the 180 GB) and hashes them concurrently on two streams."""
    args, torch, ctx, dev, tpi = env.args, env.torch, env.ctx, env.dev, env.tpi
    size, part, files = 64 << 30, 8 << 20, 8
    mine = [f for f in range(files) if f % env.world == env.rank]
    resident = 2 if len(mine) >= 2 else 1
    nparts = tpi.s3_nparts(size, part)
    bufs, parts, etags, streams = [], [], [], []
    for r in range(resident):
        b = torch.empty(size, dtype=torch.uint8, device=dev)
        ctx.dev_synth_fill(b, 3, env.rank * 2 + r, size, 1, stream=env.stream)
        bufs.append(b)
        parts.append(torch.empty((nparts, 16), dtype=torch.uint8, device=dev))
        etags.append(torch.zeros(16, dtype=torch.uint8, device=dev))
        streams.append(torch.cuda.Stream(device=dev))
    torch.cuda.synchronize()

    def step():
        cur = torch.cuda.current_stream()
        for st_ in streams:
            st_.wait_stream(cur)
        for i, _ in enumerate(mine):            # the resident buffers stand in, in turn, for this rank's files
            r = i % resident
            ctx.dev_s3_etag(bufs[r], size, part, etags[r], d_part_digests=parts[r], stream=streams[r].cuda_stream)
        for st_ in streams:
            cur.wait_stream(st_)

    ms, ms_local, clocks, launches = env.timed(step, args.sub_steps, 1) if mine else (env.timed(lambda: None, 1, 0))
    first = [bytes(e.cpu().numpy().tobytes()) for e in etags]
    if mine:
        step()
        torch.cuda.synchronize()
        assert first == [bytes(e.cpu().numpy().tobytes()) for e in etags] and any(first[0]), "c3: ETag not reproducible"
    value = files * size / (ms * 1e-3) / 1e9
    rec = {"workload": f"C3: 8 files x 64 GiB, S3 multipart ETag, 8 MiB parts (8,192 chains per file + digest of digests), "
                       f"files dealt to the ranks, {resident} file(s) resident and in flight per GPU",
           "value": round(value, 1), "unit": "GB/s", "scaling": "strong", "steps": args.sub_steps, "warmup": 1,
           "ms_per_step": round(ms, 3), "files_per_rank": len(mine), "gpu_launches": launches, "clocks": clocks,
           "etag_of_first_resident_file": tpi.s3_etag_text(first[0], nparts) if mine else None}
    if mine:
        rounds = (len(mine) + resident - 1) // resident            # a chain hashes `rounds` parts in a row per step
        per_gpu = len(mine) * size / (ms_local * 1e-3) / 1e9
        rec["roofline"] = chain_roofline(env, per_gpu, resident * nparts, rounds * (((part + 8) >> 6) + 1), ms_local, clocks)
    del bufs, parts, etags
    env.free()
    if env.rank == 0 and env.world == 1 and not args.no_cpu:
        import oracle
        threads = os.cpu_count() or 1
        sample = oracle.synth_files(3, 0, 1 << 30, 1, threads=threads)
        eng, kind = cpu_engine()
        oracle.s3_etag(sample, part, threads=threads, engine=eng)
        t0, passes = time.perf_counter(), 0
        while time.perf_counter() - t0 < 3.0:
            oracle.s3_etag(sample, part, threads=threads, engine=eng)
            passes += 1
        dt = time.perf_counter() - t0
        rec["cpu_baseline"] = {"value": round(passes * sample.size / dt / 1e9, 3), "unit": "GB/s", "cores": threads, "kind": "port",
                               "sample": f"1 GiB of one file, 8 MiB parts + digest of digests, {passes} passes in {dt:.1f} s, {kind}, "
                                         f"one part per task on {threads} threads (rclone itself hashes 4 parts at a time)"}
    return rec


def run_c5(env):
    """BASELINE config 5: 25,600 files x 4 MiB = 100 GiB, 5 % mutated (one flipped byte, size and mtime
    unchanged): full manifest rebuild + digest diff against the old manifest, files spread over the ranks."""
    args, torch, ctx, dev = env.args, env.torch, env.ctx, env.dev
    n, flen = 25600, 4 << 20
    n_local = n // env.world
    data = torch.empty(n_local * flen, dtype=torch.uint8, device=dev)
    ctx.dev_synth_fill(data, 5, env.rank * n_local, flen, n_local, stream=env.stream)
    old = torch.empty((n_local, 16), dtype=torch.uint8, device=dev)
    new = torch.empty((n_local, 16), dtype=torch.uint8, device=dev)
    ctx.dev_hash_buffers(data, n_local * flen, n_local, old, stride=flen, stream=env.stream)
    torch.cuda.synchronize()
    g = torch.Generator(device="cpu").manual_seed(5 + env.rank)
    victims = torch.randperm(n_local, generator=g)[: n_local // 20].to(dev)
    data[victims * flen + 4321] ^= 1
    idx = torch.empty(n_local, dtype=torch.int32, device=dev)
    cnt = torch.zeros(1, dtype=torch.int64, device=dev)

    def step():
        ctx.dev_hash_buffers(data, n_local * flen, n_local, new, stride=flen, stream=env.stream)
        ctx.dev_diff(old, new, n_local, idx, cnt, stream=env.stream)

    ms, ms_local, clocks, launches = env.timed(step, args.sub_steps, 1)
    found = int(cnt.item())
    assert found == n_local // 20, "c5: the diff did not return exactly the mutated files"
    assert torch.equal(torch.sort(idx[:found].long()).values, torch.sort(victims).values), "c5: wrong files reported"
    value = n * flen / (ms * 1e-3) / 1e9
    per_gpu = n_local * flen / (ms_local * 1e-3) / 1e9
    rec = {"workload": "C5: 25,600 files x 4 MiB = 100 GiB, 5 % mutated: manifest rebuild (K1) + digest diff (K4), files spread over the ranks",
           "value": round(value, 1), "unit": "GB/s", "scaling": "strong", "steps": args.sub_steps, "warmup": 1,
           "ms_per_step": round(ms, 3), "changed_found": found * env.world if env.world == 1 else None,
           "changed_found_this_rank": found, "changed_expected_this_rank": n_local // 20,
           "gpu_launches": launches, "clocks": clocks,
           "roofline": chain_roofline(env, per_gpu, n_local, ((flen + 8) >> 6) + 1, ms_local, clocks)}
    del data, old, new, idx
    env.free()
    if env.rank == 0 and env.world == 1 and not args.no_cpu:
        import oracle
        threads = os.cpu_count() or 1
        k = 256
        sample = oracle.synth_files(5, 0, flen, k, threads=threads)
        off = np.arange(k, dtype=np.uint64) * flen
        ln = np.full(k, flen, dtype=np.uint64)
        eng, kind = cpu_engine()
        was = oracle.md5_batch(sample, off, ln, threads=threads, engine=eng)
        t0, passes = time.perf_counter(), 0
        while time.perf_counter() - t0 < 3.0:
            now = oracle.md5_batch(sample, off, ln, threads=threads, engine=eng)
            oracle.diff(was, now)
            passes += 1
        dt = time.perf_counter() - t0
        rec["cpu_baseline"] = {"value": round(passes * sample.size / dt / 1e9, 3), "unit": "GB/s", "cores": threads, "kind": "port",
                               "sample": f"{k} of the 4 MiB files (1 GiB) rehashed + digest compare, {passes} passes in {dt:.1f} s, {kind}, "
                                         f"one file per task on {threads} threads"}
    return rec


def run_c4(env):
    """BASELINE config 4 scaled by 1/64 so that a step takes seconds: rank-size Zipf floor(512 MiB * k^-1.111),
    smallest 1 KiB, whole-file content-MD5 (Azure/GCS); files spread over the ranks by byte-balanced LPT.
    One file is one serial chain, so the largest file bounds the pass: the tail bound is reported beside it."""
    args, torch, ctx, dev, tpi = env.args, env.torch, env.ctx, env.dev, env.tpi
    s_exp, top = 1.111, 512 << 20
    nfiles = int(np.ceil((top / 1024) ** (1 / s_exp)))
    k = np.arange(1, nfiles + 1, dtype=np.float64)
    sizes = np.maximum(1024, np.floor(top * k ** (-s_exp))).astype(np.int64)
    np.random.default_rng(4).shuffle(sizes)
    assign = tpi.shard_lpt(sizes, env.world)
    mine = sizes[assign == env.rank]
    off = np.concatenate([[0], np.cumsum((mine + 15) // 16 * 16)[:-1]]).astype(np.int64)
    total = int(off[-1] + mine[-1])
    data = torch.empty((total + 15) // 16 * 16, dtype=torch.uint8, device=dev)
    ctx.dev_synth_fill(data, 4, env.rank, 16, data.numel() // 16, stream=env.stream)
    # the caller knows the listing: files of 8 MiB and more go in as single strided units on streams of their
    # own (the latency kernel, fed by TMA; a unit longer than a tensor-map row continues across launches),
    # everything else as one ragged batch
    big = np.nonzero(mine >= (8 << 20))[0]
    small = np.nonzero(mine < (8 << 20))[0]
    d_off, d_len = torch.from_numpy(off[small]).to(dev), torch.from_numpy(mine[small]).to(dev)
    dig = torch.empty((mine.size, 16), dtype=torch.uint8, device=dev)
    dig_small = torch.empty((small.size, 16), dtype=torch.uint8, device=dev)
    side = [torch.cuda.Stream(device=dev) for _ in big]

    def step():
        cur = torch.cuda.current_stream()
        for st_, i in zip(side, big.tolist()):
            st_.wait_stream(cur)
            ln = int(mine[i])
            ctx.dev_hash_buffers(data[int(off[i]):], ln, 1, dig[i:i + 1], stride=(ln + 15) // 16 * 16, stream=st_.cuda_stream)
        ctx.dev_hash_buffers(data, total, small.size, dig_small, d_offsets=d_off, d_lengths=d_len, stream=env.stream)
        for st_ in side:
            cur.wait_stream(st_)

    ms, ms_local, clocks, launches = env.timed(step, 1, 1)
    work_bytes = int(sizes.sum())
    value = work_bytes / (ms * 1e-3) / 1e9
    lane_rate = top / (ms * 1e-3) / 1e9                     # the pass IS the largest file's chain
    rec = {"workload": f"C4 scaled 1/64: {nfiles:,} files, Zipf 1 KiB..512 MiB ({work_bytes / 2**30:.1f} GiB), whole-file "
                       f"content-MD5, byte-balanced LPT sharding over the ranks",
           "value": round(value, 3), "unit": "GB/s", "scaling": "strong", "steps": 1, "warmup": 1, "ms_per_step": round(ms, 1),
           "largest_file_bytes": top, "single_chain_rate_GBps": round(lane_rate, 4),
           "tail_bound": "one file is one serial MD5 chain: a pass cannot end before largest file / single-chain rate",
           "tail_bound_this_scale_s": round(top / (lane_rate * 1e9), 2),
           "tail_bound_full_size_s": round((32 << 30) / (lane_rate * 1e9), 1),
           "full_size_note": "at full size (256 GiB, largest file 32 GiB) the pass is that one chain; a CPU core hashes such a "
                             "file at 0.6-0.75 GB/s, so very large single files are where this design loses",
           "gpu_launches": launches, "clocks": clocks}
    del data, dig
    env.free()
    return rec


def _tmp_root(prefix):
    import tempfile
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    return tempfile.mkdtemp(prefix=prefix, dir=base)


def run_c1(env):
    """BASELINE config 1: 1,000 files / 1 GiB synced to a local-filesystem remote with --checksum.
    Pass 1 copies, pass 2 (timed) finds nothing changed: both sides are hashed, 2 GiB through the GPU.
    Beside it the same pass on the CPU, rclone-shaped (8 checkers x 2 sides, 32 KiB reads, OpenSSL MD5)."""
    import shutil
    import oracle
    from tpi_b200 import machine
    args, ctx = env.args, env.ctx
    root = _tmp_root("tpi_c1_")
    try:
        rng = np.random.default_rng(1)
        n = 1000
        sizes = np.exp(rng.uniform(np.log(1 << 10), np.log(8 << 20), n))
        sizes = np.maximum(1, (sizes * ((1 << 30) / sizes.sum())).astype(np.int64))
        sizes[-1] += (1 << 30) - int(sizes.sum())
        src, dst = os.path.join(root, "src"), os.path.join(root, "dst")
        os.makedirs(src)
        paths = []
        for i, sz in enumerate(sizes.tolist()):
            d = os.path.join(src, f"d{i % 16:02d}")
            os.makedirs(d, exist_ok=True)
            p = os.path.join(d, f"f{i:04d}.bin")
            with open(p, "wb") as f:
                f.write(oracle.synth_bytes(1, i, sz).tobytes())
            paths.append(p)
        l0 = ctx.launch_count()
        t0 = time.perf_counter()
        st1 = machine.Transfer(ctx, src, dst, None, checksum=True)
        t_copy = time.perf_counter() - t0
        assert st1["transferred"] == n
        machine.Transfer(ctx, src, dst, None, checksum=True)          # warm
        reps, t_gpu = max(3, args.e2e_steps), []
        for _ in range(reps):
            t0 = time.perf_counter()
            st2 = machine.Transfer(ctx, src, dst, None, checksum=True)
            t_gpu.append(time.perf_counter() - t0)
            assert st2["transferred"] == 0 and st2["skipped"] == n and st2["bytes_hashed"] == 2 << 30
        both = paths + [p.replace(src, dst, 1) for p in paths]
        eng, kind = cpu_engine()
        cpu = {}
        for label, th in (("rclone_shape_16_threads", 16), ("all_cores", os.cpu_count() or 1)):
            oracle.md5_files(both, threads=th, engine=eng)
            t0 = time.perf_counter()
            for _ in range(reps):
                oracle.md5_files(both, threads=th, engine=eng)
            cpu[label] = round(reps * (2 << 30) / (time.perf_counter() - t0) / 1e9, 3)
        mean = sum(t_gpu) / reps
        both_sizes = np.concatenate([sizes, sizes]).astype(np.uint64)
        ceiling = ingest_ceiling(ctx, env.tpi.PathList(both), both_sizes, 2 << 30, (2 << 30) / mean / 1e9)
        longest = int(sizes.max())
        ceiling["longest_file_chain_ms"] = round(1e3 * longest / 0.161e9, 1)       # one file = one serial chain (~0.161 GB/s)
        return {"ceiling": ceiling,
                "workload": "C1: 1,000 files / 1 GiB on tmpfs, machine.Transfer local->local with --checksum, unchanged second "
                            "pass (2 GiB hashed: both sides); end to end: listing, file reads, H2D, kernels, D2H, plan",
                "value": round((2 << 30) / mean / 1e9, 3), "unit": "GB/s", "steps": reps, "warmup": 1,
                "ms_per_step": round(1e3 * mean, 2), "best_ms": round(1e3 * min(t_gpu), 2),
                "first_pass_copy_verify_s": round(t_copy, 3), "gpu_launches": int(ctx.launch_count() - l0),
                "cpu_baseline": {"value": cpu["rclone_shape_16_threads"], "unit": "GB/s", "cores": 16, "kind": "port",
                                 "sample": f"the same 2,000 files read through 32 KiB buffers into {kind}, one file per "
                                           "thread, 16 threads (rclone: 8 checkers x 2 sides); all cores: "
                                           f"{cpu['all_cores']} GB/s on {os.cpu_count()} threads"}}
    finally:
        shutil.rmtree(root, ignore_errors=True)


def ingest_ceiling(ctx, listing, sizes, nbytes, value_gbps, file_offsets=None, reps=3):
    """SURVEY 8d "end-to-end ceiling (a)": the same listing read into the same pinned staging buffers by the same
    reader threads with nothing copied to the device and nothing hashed (tpimd5_ingest_probe), in the same run."""
    ctx.ingest_probe(listing, sizes, file_offsets)
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        ctx.ingest_probe(listing, sizes, file_offsets)
        t.append(time.perf_counter() - t0)
    mean = sum(t) / reps
    ceil = nbytes / mean / 1e9
    return {"host_ingest_GBps": round(ceil, 3), "ms": round(1e3 * mean, 1), "frac_of_host_ingest": round(value_gbps / ceil, 3),
            "how": "tpimd5_ingest_probe: open/read/close of the same listing into pinned staging, no H2D, no kernels"}


def run_c2_files(env, n=200_000):
    """C2 end to end from REAL files: n x 4 KiB files in a 1,024-way directory fan-out on
    tmpfs through tpimd5_hash_files (open/read/close on the host, H2D, kernel, D2H), the CPU path
    (rclone-shaped: one file per thread, all cores) over the same files beside it."""
    import shutil
    from concurrent.futures import ThreadPoolExecutor
    import oracle
    ctx = env.ctx
    root = _tmp_root("tpi_c2f_")
    try:
        content = oracle.synth_files(SEED, 0, FILE_BYTES, 4096, threads=os.cpu_count() or 1)
        for d in range(1024):
            os.makedirs(os.path.join(root, f"{d:04d}"))
        per_dir = (n + 1023) // 1024
        paths = [os.path.join(root, f"{i // per_dir:04d}", f"{i:07d}") for i in range(n)]   # listing order: directory by directory

        def write_range(lo, hi):
            for i in range(lo, hi):
                fd = os.open(paths[i], os.O_WRONLY | os.O_CREAT | os.O_TRUNC, 0o644)
                k = i % 4096
                os.write(fd, content[k * FILE_BYTES:(k + 1) * FILE_BYTES].tobytes())
                os.close(fd)
        with ThreadPoolExecutor(8) as ex:
            list(ex.map(lambda lo: write_range(lo, min(n, lo + 4096)), range(0, n, 4096)))
        sizes = np.full(n, FILE_BYTES, dtype=np.uint64)
        listing = env.tpi.PathList(paths)          # the listing as C strings, encoded once (both arms use it)
        l0 = ctx.launch_count()
        ctx.hash_files(listing, sizes)             # warm: staging buffers are allocated on first use
        t = []
        for _ in range(3):
            t0 = time.perf_counter()
            got = ctx.hash_files(listing, sizes)
            t.append(time.perf_counter() - t0)
        launches = int(ctx.launch_count() - l0)
        eng, kind = cpu_engine()
        threads = os.cpu_count() or 1
        oracle.md5_files(paths[:2000], threads, eng)
        tc = []
        for _ in range(3):
            t0 = time.perf_counter()
            want = oracle.md5_files(listing, threads, eng)
            tc.append(time.perf_counter() - t0)
        if not np.array_equal(got, want):
            raise SystemExit("bench.py: c2_files digests differ from the CPU path's")
        mean, cmean = sum(t) / 3, sum(tc) / 3
        ceiling = ingest_ceiling(ctx, listing, sizes, n * FILE_BYTES, n * FILE_BYTES / mean / 1e9)
        return {"ceiling": ceiling,
                "workload": f"C2 from real files: {n:,} x 4 KiB files, 1,024-way directory fan-out on tmpfs, tpimd5_hash_files "
                            "(open/read/close + H2D + kernel + D2H)",
                "value": round(n * FILE_BYTES / mean / 1e9, 3), "unit": "GB/s", "files_per_s": round(n / mean), "steps": 3,
                "warmup": 1, "ms_per_step": round(1e3 * mean, 1), "best_ms": round(1e3 * min(t), 1), "gpu_launches": launches,
                "cpu_baseline": {"value": round(n * FILE_BYTES / cmean / 1e9, 3), "unit": "GB/s", "files_per_s": round(n / cmean),
                                 "cores": threads, "kind": "port",
                                 "sample": f"the same {n:,} files, open + 32 KiB reads + close into {kind}, one file per thread "
                                           f"on {threads} threads, 3 passes"}}
    finally:
        shutil.rmtree(root, ignore_errors=True)


def run_c3_file(env, size=(8 << 30) + 12345, part=8 << 20):
    """C3 end to end for one checkpoint file: 8 GiB (+ a ragged last part) on tmpfs through tpimd5_s3_etag_file --
    file reads into pinned staging, H2D, one lane per 8 MiB part, digest of digests, D2H -- against the CPU doing
    the same from the file already in memory (one part per thread, all cores)."""
    import shutil
    import oracle
    ctx = env.ctx
    root = _tmp_root("tpi_c3f_")
    path = os.path.join(root, "ckpt.bin")
    try:
        chunk = oracle.synth_files(3, 0, 1 << 20, 64, threads=os.cpu_count() or 1).tobytes()
        with open(path, "wb") as f:
            left = size
            while left:
                k = min(left, len(chunk))
                f.write(chunk[:k])
                left -= k
        l0 = ctx.launch_count()
        etag, n, parts = ctx.s3_etag_file(path, size, part)            # warm: staging buffers
        t = []
        for _ in range(3):
            t0 = time.perf_counter()
            etag, n, parts = ctx.s3_etag_file(path, size, part)
            t.append(time.perf_counter() - t0)
        launches = int(ctx.launch_count() - l0)
        buf = np.fromfile(path, dtype=np.uint8)
        eng, kind = cpu_engine()
        threads = os.cpu_count() or 1
        t0 = time.perf_counter()
        o_etag, o_n, o_parts = oracle.s3_etag(buf, part, threads=threads, engine=eng)
        tc = time.perf_counter() - t0
        if (etag, n) != (o_etag, o_n) or not np.array_equal(parts, o_parts):
            raise SystemExit("bench.py: c3_file ETag differs from the CPU path's")
        mean = sum(t) / 3
        offs = np.arange(n, dtype=np.uint64) * np.uint64(part)
        lens = np.minimum(np.uint64(size) - offs, np.uint64(part))
        ceiling = ingest_ceiling(ctx, env.tpi.PathList([path] * n), lens, size, size / mean / 1e9, file_offsets=offs)
        return {"ceiling": ceiling,
                "workload": f"C3 from a real file: one {size / 2**30:.1f} GiB checkpoint on tmpfs, S3 multipart ETag with 8 MiB parts "
                            f"({n} parts, the last one ragged), tpimd5_s3_etag_file (reads + H2D + kernels + D2H)",
                "value": round(size / mean / 1e9, 3), "unit": "GB/s", "steps": 3, "warmup": 1, "ms_per_step": round(1e3 * mean, 1),
                "best_ms": round(1e3 * min(t), 1), "etag": env.tpi.s3_etag_text(etag, n), "gpu_launches": launches,
                "cpu_baseline": {"value": round(size / tc / 1e9, 3), "unit": "GB/s", "cores": threads, "kind": "port",
                                 "sample": f"the same file already read into memory, {kind}, one part per task on {threads} threads, one pass"}}
    finally:
        shutil.rmtree(root, ignore_errors=True)


def run_sweep(env):
    args, torch, ctx, dev = env.args, env.torch, env.ctx, env.dev
    n, flen = args.files, FILE_BYTES
    nbytes = n * flen
    data = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    digests = torch.empty((n, 16), dtype=torch.uint8, device=dev)
    ctx.dev_synth_fill(data, SEED, env.rank * n, flen, n, stream=env.stream)
    for variant in [0, 4, 44, 64, 202, 242, 262, 0]:
        ctx.set_variant(variant)
        ms, _, _, _ = env.timed(lambda: ctx.dev_hash_buffers(data, nbytes, n, digests, stride=flen, stream=env.stream), 10, 3)
        print(json.dumps({"variant": variant, "ms": round(ms, 4), "GBps": round(nbytes / ms / 1e6, 1)}), flush=True)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return 0

    if not os.path.exists(os.path.join(ROOT, "terraform-provider-iterative_b200", "libtpimd5.so")) and rank == 0:
        import __graft_entry__
        __graft_entry__.build()          # normally prebuilt; never a fallback to anything but this library
    import torch
    import torch.distributed as dist
    import tpi_b200 as tpi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the checksum path has no CPU fallback (use --impl reference for the CPU arm)")
    near = bind_near_gpu(local_rank) if world > 1 else None
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    ctx = tpi.Context([local_rank])
    if args.variant:
        ctx.set_variant(args.variant)
    env = Env(args, torch, dist, tpi, ctx, dev, rank, local_rank, world)
    env.near = near

    sections = {"c1": run_c1, "c2_files": run_c2_files, "c3": run_c3, "c3_file": run_c3_file, "c4": run_c4, "c5": run_c5}
    if args.sweep:
        run_sweep(env)
    elif args.workload in sections:
        # one sub-record alone, as its own line (side runs, profiling)
        rec = sections[args.workload](env)
        if rank == 0:
            rec = dict(rec)
            rec.update({"metric": METRIC, "n_gpus": world, "higher_is_better": True, "vs_baseline": None, "dtype": "u32",
                        "data": "synthetic", "config": {"workload": rec.pop("workload"), "kernel_variant": args.variant}})
            print(json.dumps(rec), flush=True)
    else:
        line = run_c2(env)
        if args.workload == "all" and not args.only_c2:
            # every other BASELINE config, inside the same line.  Order: device-resident ones first
            # (they need the HBM the C2 buffers just released), then the file-based ones at N=1.
            subs = {"c3": run_c3(env), "c5": run_c5(env), "c4": run_c4(env)}
            if world == 1:
                subs["c1"] = run_c1(env)
                subs["c2_files"] = run_c2_files(env)
                subs["c3_file"] = run_c3_file(env)
            line.update(subs)
        if rank == 0:
            print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
