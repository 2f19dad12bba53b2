#!/usr/bin/env python
"""bench.py — LZ4E hot path benchmark (driver contract in the task statement).

A "step" is one pass of the hot path over one batch of synthetic blocks:
LZ4E_compress_default on every block, then the LZ4E decompressor on every
compressed block (device-level C ABI, inputs resident in HBM).

  N=1 workload  = BASELINE.json configs[1]: 1 GiB of independent 4 KiB blocks, one
                  bvec each (byU16), content class = block index mod 4.
  N>1           = weak scaling: every rank owns 1 GiB of the N GiB batch, sharded by
                  block index, no collective on the data path.
  extra.cfg3_64k = a 2 GiB subset of configs[2] (64 KiB blocks over 16 odd,
                  non-contiguous segments, mid-bvec iterator start) on rank 0.

  --impl reference  times the reference's own C code (oracle/_ref, built from
                  /root/reference's sources unmodified; else the oracle port) on the
                  host cores, on a bounded sample of the same workload.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools", "synth"))

METRIC = "compress & decompress input GB/s at 4/64 KiB blocks, 1/2/4/8 B200, byte-identical"
# ONE unit string for `value`, `e2e` and `cpu_baseline` in both arms (the driver only forms
# ratios between equal units)
UNIT = "GB/s of uncompressed input through compress+decompress (one step = both passes)"
BLOCK = 4096
BLOCKS_PER_GPU = 262144          # 1 GiB
DST_STRIDE = 4224                # >= LZ4E_COMPRESSBOUND(4096) = 4128, multiple of 128
CFG3_BLOCKS = 32768              # 2 GiB subset of config 3
CFG3_SIZE = 65536


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML polled from a thread
    every few ms (the timed region of a short run is well below nvidia-smi's start-up time);
    `nvidia-smi --query-gpu` in a loop is the fallback when NVML cannot be loaded."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu, uuid=None):
        self.gpu, self.uuid = gpu, uuid
        self.sm, self.reasons, self.power = [], set(), []
        self.sm_max = None
        self.nvml = self.handle = self.thread = self.proc = None
        self.stop_flag = threading.Event()
        self.lines = []
        self.how = None
        try:
            import pynvml
            pynvml.nvmlInit()
            h = None
            if uuid:
                try:
                    h = pynvml.nvmlDeviceGetHandleByUUID(uuid if uuid.startswith("GPU-") else "GPU-" + uuid)
                except Exception:
                    h = None
            if h is None:
                h = pynvml.nvmlDeviceGetHandleByIndex(gpu)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self.nvml, self.handle, self.how = pynvml, h, "nvml"
        except Exception:
            self.nvml = None

    def _poll(self):
        nv, h = self.nvml, self.handle
        names = (("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown),
                 ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                 ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown),
                 ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap))
        while not self.stop_flag.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                for name, bit in names:
                    if r & bit:
                        self.reasons.add(name)
                self.power.append(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)
            except Exception:
                pass
            self.stop_flag.wait(0.004)

    def start(self):
        if self.nvml is not None:
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
            self.how = "nvidia-smi"
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.thread is not None:
            self.stop_flag.set()
            self.thread.join(timeout=1.0)
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None,
                    "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons),
                    "samples": len(self.sm), "how": self.how,
                    "power_w_max": round(max(self.power), 1) if self.power else None}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml and nvidia-smi unavailable"],
                    "samples": 0}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm), "how": self.how}


# ----------------------------------------------------------------------------- workloads
def cfg2_tables(pkg, src_base, dst_base, out_base, nblocks):
    """4 KiB blocks, one bvec each; destination = one (multi-page) bvec of COMPRESSBOUND."""
    idx = np.arange(nblocks, dtype=np.uint64)
    bound = pkg.compress_bound(BLOCK)

    def tab(base, stride, length):
        bv = np.zeros(nblocks, dtype=pkg.DEV_BVEC)
        bv["addr"] = np.uint64(base) + idx * np.uint64(stride)
        bv["len"] = length
        it = np.zeros(nblocks, dtype=pkg.DEV_ITER)
        it["first"] = idx.astype(np.uint32)
        it["bi_size"] = length
        return bv, it
    return tab(src_base, BLOCK, BLOCK), tab(dst_base, DST_STRIDE, bound), tab(out_base, BLOCK, BLOCK)


def cfg3_tables(pkg, base, nblocks, size, capacity, salt):
    """`size`-byte blocks over 16 non-contiguous segments with odd offsets and lengths and a
    mid-bvec iterator start; each segment owns a private run of 3 pages."""
    nseg = 16
    b = np.arange(nblocks, dtype=np.int64)[:, None]
    i = np.arange(nseg, dtype=np.int64)[None, :]
    per = capacity // nseg
    delta = ((b * 31 + i * 17 + salt) % 129) - 64
    delta = np.where(i % 2 == 1, -delta, delta)
    lens = per + delta
    skip = 1 + ((b[:, 0] * 2654435761 + salt) >> 7) % 4095
    lens[:, 0] += skip
    lens[:, -1] += capacity - (lens.sum(axis=1) - skip)
    offs = (53 * (i + b) + 7 + salt) % 509
    assert (lens > 0).all() and (offs + lens <= 3 * 4096).all()
    seg_index = (b * nseg + i).astype(np.uint64)
    bv = np.zeros(nblocks * nseg, dtype=pkg.DEV_BVEC)
    bv["addr"] = (np.uint64(base) + seg_index * np.uint64(3 * 4096) + offs.astype(np.uint64)).reshape(-1)
    bv["len"] = lens.reshape(-1).astype(np.uint32)
    it = np.zeros(nblocks, dtype=pkg.DEV_ITER)
    it["first"] = (np.arange(nblocks) * nseg).astype(np.uint32)
    it["bi_size"] = size
    it["bi_bvec_done"] = skip.astype(np.uint32)
    return bv, it, nblocks * nseg * 3 * 4096


def event_ms(torch, fn, stream):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    fn()
    e1.record(stream)
    return e0, e1


# ----------------------------------------------------------------------------- product arm
def run_ours(args):
    import torch
    import lz4e_synth
    pkg = importlib.import_module("lz4-sgori_b200")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    ctx = pkg.Context(local)          # raises without a B200: no CPU path
    stream = torch.cuda.Stream(device=local)
    sp = stream.cuda_stream

    total_blocks = BLOCKS_PER_GPU * world
    first, count = pkg.shard_range(total_blocks, rank, world)
    nb = int(count)
    in_bytes = nb * BLOCK

    # ---- device-resident batch (config 2 shard) ----
    d_src = ctx.dev_alloc(nb * BLOCK)
    d_dst = ctx.dev_alloc(nb * DST_STRIDE)
    d_out = ctx.dev_alloc(nb * BLOCK)
    (sbv, sit), (dbv, dit), (obv, oit) = cfg2_tables(pkg, d_src, d_dst, d_out, nb)
    t = [ctx.upload(x) for x in (sbv, sit, dbv, dit, obv, oit)]
    d_rets, d_rets2 = ctx.dev_alloc(4 * nb), ctx.dev_alloc(4 * nb)
    lz4e_synth.dev(local, t[0], t[1], nb, int(first), BLOCK)
    ctx.compress_dev(t[0], t[1], t[2], t[3], nb, BLOCK, d_rets, flags=pkg.DEV_DST_LINEAR, stream=sp)
    ctx.sync(sp)
    rets = ctx.download(d_rets, np.int32, nb)
    assert (rets > 0).all()
    comp_bytes = int(rets.sum())
    cit = dit.copy()
    cit["bi_size"] = rets.astype(np.uint32)
    d_cit = ctx.upload(cit)
    max_c = int(rets.max())

    def step(evs=None):
        if evs is not None:
            evs.append(event_ms(torch, lambda: ctx.compress_dev(
                t[0], t[1], t[2], t[3], nb, BLOCK, d_rets, flags=pkg.DEV_DST_LINEAR, stream=sp),
                stream))
            evs.append(event_ms(torch, lambda: ctx.decompress_dev(
                t[2], d_cit, t[4], t[5], nb, max_c, BLOCK, d_rets2,
                flags=pkg.DEV_DST_LINEAR | pkg.DEV_SRC_LINEAR, stream=sp), stream))
        else:
            ctx.compress_dev(t[0], t[1], t[2], t[3], nb, BLOCK, d_rets, flags=pkg.DEV_DST_LINEAR, stream=sp)
            ctx.decompress_dev(t[2], d_cit, t[4], t[5], nb, max_c, BLOCK, d_rets2,
                               flags=pkg.DEV_DST_LINEAR | pkg.DEV_SRC_LINEAR, stream=sp)

    for _ in range(max(args.warmup, 3)):
        step()
    ctx.sync(sp)
    # round-trip check on the warm-up result (size-independent property at full size)
    r2 = ctx.download(d_rets2, np.int32, nb)
    assert (r2 == BLOCK).all(), "round trip sizes"
    src_t = ctx.download(d_src, np.uint8, min(in_bytes, 64 << 20))
    out_t = ctx.download(d_out, np.uint8, min(in_bytes, 64 << 20))
    assert np.array_equal(src_t, out_t), "round trip bytes"
    del src_t, out_t
    # every 1024th block's source and compressed bytes, kept for the oracle byte-compare of
    # the cpu_baseline leg (rank 0, N=1)
    parity = None
    if rank == 0 and world == 1 and not args.no_cpu:
        parity = [(b, ctx.download(d_src + b * BLOCK, np.uint8, BLOCK),
                   ctx.download(d_dst + b * DST_STRIDE, np.uint8, int(rets[b])))
                  for b in range(0, nb, 1024)]

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local, uuid=str(torch.cuda.get_device_properties(local).uuid))
    l0 = ctx.launches()
    evs = []
    barrier()
    sampler.start()
    e_begin, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e_begin.record(stream)
    for _ in range(args.steps):
        step(evs)
    e_end.record(stream)
    ctx.sync(sp)
    barrier()
    clocks = sampler.stop()
    launches = ctx.launches() - l0
    total_ms = e_begin.elapsed_time(e_end)
    c_ms = float(np.mean([a.elapsed_time(b) for a, b in evs[0::2]]))
    d_ms = float(np.mean([a.elapsed_time(b) for a, b in evs[1::2]]))
    if dist is not None:
        tt = torch.tensor([total_ms, c_ms, d_ms], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms, c_ms, d_ms = [float(x) for x in tt.tolist()]
        cb = torch.tensor([comp_bytes], device=f"cuda:{local}", dtype=torch.int64)
        dist.all_reduce(cb)
        comp_total = int(cb.item())
    else:
        comp_total = comp_bytes
    ms_per_step = total_ms / args.steps
    job_in = total_blocks * BLOCK

    result = None
    if rank == 0:
        peak, peak_src = peaks()
        c_gbs = job_in / (c_ms * 1e-3) / 1e9
        d_gbs = job_in / (d_ms * 1e-3) / 1e9
        alg = (job_in + comp_total) / world    # N + C per launch (per GPU)
        dom = "lz4e_compress_kernel" if c_ms >= d_ms else "lz4e_decompress_kernel"
        dom_ms = max(c_ms, d_ms)
        ach = alg / (dom_ms * 1e-3) / 1e9
        result = {
            "metric": METRIC,
            "value": round(job_in / (ms_per_step * 1e-3) / 1e9, 3),
            "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "configs[1]: %d x 4 KiB blocks per GPU (1 GiB/GPU), 1 bvec each "
                                   "(byU16), class = block index mod 4 (text/random/sparse/mixed), "
                                   "dst capacity LZ4E_COMPRESSBOUND" % BLOCKS_PER_GPU,
                       "l2": "inputs (1 GiB per GPU) exceed the 126 MB L2; no explicit flush",
                       "sharding": "contiguous block ranges per rank, no collective"},
            "compress_gbs": round(c_gbs, 3), "decompress_gbs": round(d_gbs, 3),
            "decompress_compressed_in_gbs": round(comp_total / (d_ms * 1e-3) / 1e9, 3),
            "ratio": round(job_in / comp_total, 4),
            "kernel_ms": {"compress": round(c_ms, 4), "decompress": round(d_ms, 4)},
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": round(ach, 2), "peak": peak,
                         "peak_source": peak_src, "unit": "GB/s", "frac": round(ach / peak, 5),
                         "traffic": profile_traffic(dom),
                         "algorithmic_bytes_per_launch": int(alg)},
            "roofline_other": {
                "kernel": "lz4e_decompress_lane_kernel" if dom.startswith("lz4e_c") else "lz4e_compress_kernel",
                "achieved": round(alg / (min(c_ms, d_ms) * 1e-3) / 1e9, 2),
                "frac": round(alg / (min(c_ms, d_ms) * 1e-3) / 1e9 / peak, 5),
                "traffic": profile_traffic("lz4e_decompress_lane_kernel" if dom.startswith("lz4e_c")
                                           else "lz4e_compress_kernel")},
            "gpu_launches": int(launches), "clocks": clocks,
        }
    for p in (d_src, d_dst, d_out, d_rets, d_rets2, d_cit, *t):
        ctx.dev_free(p)

    if not args.no_extra:
        # configs[2]/[3]: every rank runs its 2 GiB share of the 64 KiB workload
        cfg3 = run_cfg3(torch, pkg, ctx, stream, args, dist, rank, world, local)
        if rank == 0:
            result["extra"] = {"cfg3_64k": cfg3}
    if not args.no_extra and not args.no_replay:
        # configs[4]: one replay harness per GPU (its own process, its own request stream),
        # started together; GB/s add up, latencies are the worst rank's
        if dist is not None:
            dist.barrier()
        rep = run_replay(device=local, seed=1 + rank)
        if dist is not None:
            allrep = [None] * world
            dist.all_gather_object(allrep, rep)
        else:
            allrep = [rep]
        if rank == 0:
            result["extra"]["cfg5_replay"] = merge_replay(allrep)
    if not args.no_e2e:
        e2e = run_e2e(torch, pkg, ctx, args, dist, rank, world, local)
        if rank == 0:
            result["e2e"] = e2e
    elif rank == 0:
        result["e2e"] = None
    if rank == 0 and world == 1 and not args.no_cpu:
        result["cpu_baseline"] = cpu_baseline(pkg, budget_s=args.cpu_seconds, parity=parity)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(result))
    ctx.close()


def profile_traffic(kernel):
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of ONE launch of `kernel` at
    this bench's size, from the committed `ncu --set full` capture (profiles/traffic.json,
    written by tools/ncu_traffic.py); None when no capture of that kernel is committed."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        e = t.get(kernel)
        if e and int(e.get("blocks", 0)) == BLOCKS_PER_GPU:
            return int(e["dram_bytes_per_launch"])
    except (OSError, ValueError, KeyError):
        pass
    return None


def merge_replay(reps):
    """Aggregate of the per-GPU replay lines: throughput and counters add up, batch latencies
    are the slowest rank's, every rank must have matched the oracle."""
    bad = [r for r in reps if "error" in r]
    if bad:
        return {"error": bad[0]["error"], "n_gpus": len(reps)}
    out = dict(reps[0])
    out["n_gpus"] = len(reps)
    if len(reps) == 1:
        return out
    for k in ("gbs", "write_gbs"):
        out[k] = round(sum(r[k] for r in reps), 3)
    for k in ("batch_ms_p50", "batch_ms_p99", "batch_ms_max"):
        out[k] = max(r[k] for r in reps)
    for k in ("write_reqs", "read_reqs", "write_bytes", "read_bytes", "compressed_bytes",
              "oracle_compressed_bytes", "failed", "mismatches", "gpu_launches"):
        out[k] = sum(r[k] for r in reps)
    out["per_bio_digest_equal"] = all(r.get("per_bio_digest_equal") for r in reps)
    out["per_gpu_gbs"] = [r["gbs"] for r in reps]
    out["workload"] += " (one harness process per GPU, different request streams)"
    return out


def run_replay(device=0, seed=1):
    """BASELINE.json configs[4]: the C replay harness of the chunk layer / request path
    (tools/replay/lz4e_replay.c), QD256 random 4-128 KiB writes/reads on this GPU; --oracle
    makes it compare every bio's compressed bytes with the reference (checker, untimed)."""
    import subprocess
    exe = os.path.join(ROOT, "tools", "replay", "lz4e_replay")
    try:
        if not os.path.exists(exe):
            subprocess.run(["make", "-C", os.path.dirname(exe)], check=True, capture_output=True)
        # two warm-up batches (run and checked like the others, not timed): the first batch on a
        # fresh context pays its pinned staging, pool and kernel-attribute set-up (about 100 ms)
        cmd = [exe, "--arena-bios", "--batches", "48", "--warmup", "2", "--device", str(device),
               "--seed", str(seed)]
        for so in (os.path.join(ROOT, "oracle", "_ref", "liblz4e_ref.so"),
                   os.path.join(ROOT, "oracle", "_build", "liblz4e_port.so")):
            if os.path.exists(so):
                cmd += ["--oracle", so]
                break
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=180)
        if out.returncode != 0:
            return {"error": "rc=%d %s" % (out.returncode, out.stderr[-300:])}
        return json.loads(out.stdout.strip().splitlines()[-1])
    except Exception as e:
        return {"error": repr(e)}


def run_cfg3(torch, pkg, ctx, stream, args, dist, rank, world, local):
    """configs[2] + configs[3]: 64 KiB blocks over 16 odd, non-contiguous segments on both
    sides, mid-bvec iterator start (byU32).  EVERY rank runs CFG3_BLOCKS blocks (2 GiB; 8 ranks
    = the full 16 GiB), sharded by block index, bracketed by barriers; the slowest rank's
    kernel times count and the GB/s are whole-job aggregates."""
    import lz4e_synth
    sp = stream.cuda_stream
    nb = CFG3_BLOCKS
    first = rank * nb
    bound = pkg.compress_bound(CFG3_SIZE)
    region = nb * 16 * 3 * 4096
    err = None
    try:
        d_src, d_dst, d_out = ctx.dev_alloc(region), ctx.dev_alloc(region), ctx.dev_alloc(region)
        sbv, sit, _ = cfg3_tables(pkg, d_src, nb, CFG3_SIZE, CFG3_SIZE, 0)
        dbv, dit, _ = cfg3_tables(pkg, d_dst, nb, bound, bound, 11)
        obv, oit, _ = cfg3_tables(pkg, d_out, nb, CFG3_SIZE, CFG3_SIZE, 5)
        t = [ctx.upload(x) for x in (sbv, sit, dbv, dit, obv, oit)]
        d_rets, d_rets2 = ctx.dev_alloc(4 * nb), ctx.dev_alloc(4 * nb)
        lz4e_synth.dev(local, t[0], t[1], nb, first, CFG3_SIZE)
        ctx.compress_dev(t[0], t[1], t[2], t[3], nb, CFG3_SIZE, d_rets, flags=pkg.DEV_NO_BYU16, stream=sp)
        ctx.sync(sp)
        rets = ctx.download(d_rets, np.int32, nb)
        assert (rets > 0).all()
        cit = dit.copy()
        cit["bi_size"] = rets.astype(np.uint32)
        d_cit = ctx.upload(cit)
        cms, dms = [], []
        warm, reps = 2, 4
        for it in range(warm + reps):
            if it == warm:
                if dist is not None:
                    dist.barrier()
                torch.cuda.synchronize()
            ec = event_ms(torch, lambda: ctx.compress_dev(t[0], t[1], t[2], t[3], nb, CFG3_SIZE, d_rets,
                                                          flags=pkg.DEV_NO_BYU16, stream=sp), stream)
            ed = event_ms(torch, lambda: ctx.decompress_dev(t[2], d_cit, t[4], t[5], nb, int(rets.max()),
                                                            CFG3_SIZE, d_rets2, stream=sp), stream)
            ctx.sync(sp)
            if it >= warm:
                cms.append(ec[0].elapsed_time(ec[1]))
                dms.append(ed[0].elapsed_time(ed[1]))
        r2 = ctx.download(d_rets2, np.int32, nb)
        assert (r2 == CFG3_SIZE).all(), "64k round trip sizes"
        # bit-exact round trip of a sample of blocks (gathered through the segment tables)
        for b in range(0, nb, 4096):
            assert np.array_equal(_gather_dev(ctx, sbv, sit, b, CFG3_SIZE),
                                  _gather_dev(ctx, obv, oit, b, CFG3_SIZE)), "64k round trip bytes"
        for p in (d_src, d_dst, d_out, d_rets, d_rets2, d_cit, *t):
            ctx.dev_free(p)
        c_ms, d_ms, cbytes = float(np.mean(cms)), float(np.mean(dms)), int(rets.sum())
    except Exception as e:  # the headline line must still print
        err, c_ms, d_ms, cbytes = repr(e), 0.0, 0.0, 0
    if dist is not None:
        tt = torch.tensor([c_ms, d_ms, 1.0 if err else 0.0], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        c_ms, d_ms, bad = [float(x) for x in tt.tolist()]
        cb = torch.tensor([cbytes], device=f"cuda:{local}", dtype=torch.int64)
        dist.all_reduce(cb)
        cbytes = int(cb.item())
        if bad and not err:
            err = "another rank failed"
    if err:
        return {"error": err}
    nbytes = nb * CFG3_SIZE * world
    peak, _ = peaks()
    return {"workload": "configs[2]/[3]: %d x 64 KiB blocks per GPU (2 GiB/GPU, %d GiB in all), 16 odd "
                        "non-contiguous segments per side, mid-bvec start (byU32); every rank runs "
                        "its share, barrier-bracketed, slowest rank counts" % (nb, 2 * world),
            "n_gpus": world,
            "compress_gbs": round(nbytes / (c_ms * 1e-3) / 1e9, 3),
            "decompress_gbs": round(nbytes / (d_ms * 1e-3) / 1e9, 3),
            "decompress_compressed_in_gbs": round(cbytes / (d_ms * 1e-3) / 1e9, 3),
            "ratio": round(nbytes / cbytes, 4),
            "kernel_ms": {"compress": round(c_ms, 3), "decompress": round(d_ms, 3)},
            "algorithmic_bytes_per_launch": int((nbytes + cbytes) // world),
            "decompress_roofline_frac": round((nbytes + cbytes) / (d_ms * 1e-3) / 1e9 / (peak * world), 5),
            "compress_roofline_frac": round((nbytes + cbytes) / (c_ms * 1e-3) / 1e9 / (peak * world), 5)}


def _gather_dev(ctx, bv, it, b, nbytes):
    """Linear bytes of block b from device memory, walking its dev_bvec list."""
    out, i, done = [], int(it["first"][b]) + int(it["bi_idx"][b]), int(it["bi_bvec_done"][b])
    left = nbytes
    while left > 0:
        take = min(left, int(bv["len"][i]) - done)
        out.append(ctx.download(int(bv["addr"][i]) + done, np.uint8, take))
        left -= take
        done = 0
        i += 1
    return np.concatenate(out) if out else np.zeros(0, np.uint8)


def run_e2e(torch, pkg, ctx, args, dist=None, rank=0, world=1, local=0):
    """The same step through the reference-facing host API (lz4e_b200_compress_batch /
    lz4e_b200_decompress_batch = LZ4E_compress_default per block) with HOST buffers in pinned
    page arenas: host<->device copies of pages, descriptor tables and results are inside the
    timed region.  N > 1: every rank runs its own shard through its own GPU (no collective);
    the timed loop is bracketed by barriers and the slowest rank's time counts."""
    sg = pkg.sg
    nb = args.e2e_blocks
    # How compressed blocks (40 % of their LZ4E_COMPRESSBOUND slots) go up for decompression:
    # alone on the host a GPU is fastest with one copy-engine transfer of the spanned range
    # (29.5 vs 33 ms per GiB); when several GPUs share the host's memory system the bytes count
    # and the exact-byte transfer kernel wins (8 GPUs: 45.1 vs 39.8 GB/s, 4 GPUs: 33.4 vs 30.8;
    # 2 GPUs still prefer the range: 33.6 vs 29.1; profiles/r2ab_e2e_n8.md).  The library's
    # default is the former; set per deployment.
    if world >= 4:
        os.environ.setdefault("LZ4E_B200_SPARSE_SRC", "kernel")
    src_arena = pkg.Arena(ctx, nb + 1)
    dst_pages = (nb * DST_STRIDE + 4095) // 4096 + 2
    dst_arena = pkg.Arena(ctx, dst_pages)
    out_arena = pkg.Arena(ctx, nb + 1)
    bound = pkg.compress_bound(BLOCK)
    idx = np.arange(nb, dtype=np.uint64)

    def lay(arena, stride, length):
        bv = np.zeros(nb, dtype=pkg.BIO_VEC)
        a = np.uint64(arena.base) + idx * np.uint64(stride)
        bv["bv_page"] = a & ~np.uint64(4095)
        bv["bv_offset"] = (a & np.uint64(4095)).astype(np.uint32)
        bv["bv_len"] = length
        it = np.zeros(nb, dtype=pkg.BVEC_ITER)
        it["bi_size"] = length
        return sg.Layout(arena, bv, idx.copy(), it, arena.npages)
    src, dst, out = lay(src_arena, BLOCK, BLOCK), lay(dst_arena, DST_STRIDE, bound), \
        lay(out_arena, BLOCK, BLOCK)
    import lz4e_synth
    lz4e_synth.host_blocks(src_arena.mem, BLOCK, BLOCK, nb, first=rank * nb)  # global index = class, seed
    creqs, keep1 = pkg.make_reqs(src, dst)
    # the decompress requests read the compressed blocks where compress left them; only the
    # compressed sizes change from step to step
    csrc = sg.Layout(dst_arena, dst.bvecs, dst.first, dst.iters.copy(), dst.npages_used)
    dreqs, keep2 = pkg.make_reqs(csrc, out)
    d_src_iters = keep2[0]

    def once():
        creqs["ret"] = 0
        rc = ctx.compress_batch(creqs, flags=pkg.F_KEEP_ITERS)
        assert rc == 0, ctx.err()
        sizes = creqs["ret"]
        d_src_iters["bi_size"] = sizes
        rc = ctx.decompress_batch(dreqs, flags=0)
        assert rc == 0, ctx.err()
        return sizes.copy(), dreqs["ret"].copy()

    for _ in range(2):
        sizes, r2 = once()
    assert (sizes > 0).all() and (r2 == BLOCK).all()
    assert np.array_equal(out_arena.mem[:nb * BLOCK], src_arena.mem[:nb * BLOCK]), "e2e round trip"
    reps = 3
    ctx.sync()
    h0, d0 = ctx.traffic()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        once()
    dt = (time.perf_counter() - t0) / reps
    ctx.sync()
    h1, d1 = ctx.traffic()
    # bytes that crossed PCIe per step, counted by the library where it enqueues the copies
    # (copy engines) and by the transfer kernel on the device (exact bytes per block)
    h2d, d2h = (h1 - h0) // reps, (d1 - d0) // reps
    if dist is not None:
        tt = torch.tensor([dt], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
    serial_dt = dt

    # ---- the same steps with TWO in flight: one thread calls lz4e_b200_compress_batch for
    # step s + 1 while another calls lz4e_b200_decompress_batch for step s (what a block device
    # with many bios in flight does; the context runs two batches side by side).  A compress
    # batch is upload-heavy, a decompress batch download-heavy, and PCIe is full duplex.  Every
    # step still uploads its own inputs and downloads its own results; compressed blocks of
    # consecutive steps alternate between two destination arenas. ----
    import threading
    dst_arena2 = pkg.Arena(ctx, dst_pages)
    dst2 = lay(dst_arena2, DST_STRIDE, bound)
    creqs2, keep3 = pkg.make_reqs(src, dst2)
    csrc2 = sg.Layout(dst_arena2, dst2.bvecs, dst2.first, dst2.iters.copy(), dst2.npages_used)
    dreqs2, keep4 = pkg.make_reqs(csrc2, out)
    CR, DR, DI = (creqs, creqs2), (dreqs, dreqs2), (d_src_iters, keep4[0])
    warm, timed = 2, 6
    free = [threading.Semaphore(1), threading.Semaphore(1)]
    ready = [threading.Semaphore(0), threading.Semaphore(0)]
    stamps, errs = {}, []

    def comp_thread():
        try:
            for st in range(warm + timed):
                b = st & 1
                free[b].acquire()
                CR[b]["ret"] = 0
                assert ctx.compress_batch(CR[b], flags=pkg.F_KEEP_ITERS) == 0, ctx.err()
                ready[b].release()
        except Exception as e:  # noqa: BLE001
            errs.append(repr(e))
            for sem in ready:
                sem.release()

    def dec_thread():
        try:
            for st in range(warm + timed):
                b = st & 1
                ready[b].acquire()
                if errs:
                    return
                DI[b]["bi_size"] = CR[b]["ret"]
                assert ctx.decompress_batch(DR[b], flags=0) == 0, ctx.err()
                assert (DR[b]["ret"] == BLOCK).all()
                stamps[st] = time.perf_counter()
                free[b].release()
        except Exception as e:  # noqa: BLE001
            errs.append(repr(e))
            for sem in free:
                sem.release()

    out_arena.mem[:nb * BLOCK] = 0
    ctx.sync()
    h0, d0 = ctx.traffic()
    if dist is not None:
        dist.barrier()
    ths = [threading.Thread(target=comp_thread), threading.Thread(target=dec_thread)]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    assert not errs, errs
    ctx.sync()
    h1, d1 = ctx.traffic()
    assert np.array_equal(out_arena.mem[:nb * BLOCK], src_arena.mem[:nb * BLOCK]), "pipelined e2e round trip"
    dt = (stamps[warm + timed - 1] - stamps[warm - 1]) / timed
    if dist is not None:
        tt = torch.tensor([dt], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
    dst_arena2.close()
    for a in (src_arena, dst_arena, out_arena):
        a.close()
    pipe_dt, dt = dt, serial_dt
    return {"value": round(world * nb * BLOCK / dt / 1e9, 3),
            "how": "one step at a time: compress batch, then decompress batch of what it produced",
            # measured as well, not the headline: on this box both PCIe directions together
            # deliver about 70 GB/s, so two steps in flight gain nothing
            "two_steps_in_flight": {
                "value": round(world * nb * BLOCK / pipe_dt / 1e9, 3),
                "ms_per_step": round(pipe_dt * 1e3, 3),
                "how": "two host threads on one context: compress batch of step s+1 next to the "
                       "decompress batch of step s; %d timed steps after %d" % (timed, warm)},
            "unit": UNIT, "buffers": "host (pinned page arenas); copies inside the timed region",
            "sparse_src": os.environ.get("LZ4E_B200_SPARSE_SRC", "range when >= 25 % full (default)"),
            "algorithmic_pcie_bytes_per_step": int(2 * (nb * BLOCK + int(sizes.sum()))) * world,
            "blocks": world * nb, "ms_per_step": round(dt * 1e3, 3),
            "h2d_bytes_per_step": int(h2d) * world, "d2h_bytes_per_step": int(d2h) * world,
            "api": "lz4e_b200_compress_batch + lz4e_b200_decompress_batch (pinned page arenas)"}


# ----------------------------------------------------------------------------- CPU arms
# Nothing below loads the product library: the reference's C code (oracle/_ref, else the
# oracle port) is what runs, and the synthetic contents come from the oracle driver's own
# copy of the generator (tools/synth/lz4e_synth.h).
def cpu_sample(sg, lib, nblocks, threads, first_index=0):
    """Host copy of `nblocks` blocks of the config-2 workload + bio_vec tables."""
    from oracle import oracle
    arena = sg.HostArena(nblocks + 1)
    specs = [sg.segs_single(BLOCK)] * nblocks
    src = sg.layout_from_segments(arena, [dict(s, gap=0) for s in specs])
    oracle.synth_blocks(lib, arena.mem, BLOCK, BLOCK, nblocks, first=first_index, nthreads=threads)
    bound = sg.compress_bound(BLOCK)
    darena = sg.HostArena(nblocks * 2 + 1)
    dst = sg.layout_from_segments(darena, [dict(sg.segs_single(bound), gap=0)] * nblocks)
    return src, dst


def cpu_run(lib, src, dst, threads):
    """One CPU step on a sample: compress every block, then decompress every block."""
    from oracle import oracle
    n = src.nblocks()
    rets, t_c = oracle.compress_batch(lib, src.bvecs, src.first, src.iters, dst.bvecs, dst.first,
                                      dst.iters, nthreads=threads)
    assert (rets > 0).all()
    doff = (dst.bvecs["bv_page"] - np.uint64(dst.arena.base)).astype(np.uint64)
    outbuf = np.zeros(n * BLOCK + 64, dtype=np.uint8)
    r2, t_d = oracle.decompress_batch(lib, dst.arena.mem, doff, rets, outbuf,
                                      np.arange(n, dtype=np.uint64) * np.uint64(BLOCK),
                                      np.full(n, BLOCK, dtype=np.int32), nthreads=threads)
    assert (r2 == BLOCK).all()
    return t_c, t_d


def cpu_cfg3(sg, lib, threads, nblocks):
    """The reference C path on a subset of configs[2]: 64 KiB blocks over 16 odd segments
    (compress: SG source and destination as on the GPU; decompress: the reference decoder only
    takes contiguous buffers, so it runs on the linearised compressed blocks)."""
    from oracle import oracle

    class _Np:  # cfg3_tables only needs the dtypes
        DEV_BVEC = np.dtype([("addr", "<u8"), ("len", "<u4"), ("rsvd", "<u4")])
        DEV_ITER = np.dtype([("first", "<u4"), ("bi_size", "<u4"), ("bi_idx", "<u4"),
                             ("bi_bvec_done", "<u4")])
    bound = sg.compress_bound(CFG3_SIZE)
    region = nblocks * 16 * 3 * 4096
    smem = np.zeros(region + 4096, dtype=np.uint8)
    dmem = np.zeros(region + 4096, dtype=np.uint8)
    sbase = (smem.ctypes.data + 4095) & ~4095
    dbase = (dmem.ctypes.data + 4095) & ~4095

    def host_tables(base, size, cap, salt):
        bv, it, _ = cfg3_tables(_Np, base, nblocks, size, cap, salt)
        hv = np.zeros(len(bv), dtype=oracle.BIO_VEC)
        hv["bv_page"] = bv["addr"] & ~np.uint64(4095)
        hv["bv_offset"] = (bv["addr"] & np.uint64(4095)).astype(np.uint32)
        hv["bv_len"] = bv["len"]
        hi = np.zeros(nblocks, dtype=oracle.BVEC_ITER)
        hi["bi_size"] = it["bi_size"]
        hi["bi_bvec_done"] = it["bi_bvec_done"]
        return bv, it, hv, hi, it["first"].astype(np.uint64)
    sbv, sit, shv, shi, sfirst = host_tables(sbase, CFG3_SIZE, CFG3_SIZE, 0)
    dbv, dit, dhv, dhi, dfirst = host_tables(dbase, bound, bound, 11)
    # contents: generate linear, scatter into the segments
    lin = np.zeros(nblocks * CFG3_SIZE, dtype=np.uint8)
    oracle.synth_blocks(lib, lin, CFG3_SIZE, CFG3_SIZE, nblocks, first=0, nthreads=threads)

    def walk(bv, it, base_addr, mem, b, nbytes, data=None):
        out, i, done, left, pos = [], int(it["first"][b]), int(it["bi_bvec_done"][b]), nbytes, 0
        off0 = mem.ctypes.data
        while left > 0:
            take = min(left, int(bv["len"][i]) - done)
            o = int(bv["addr"][i]) + done - off0
            if data is None:
                out.append(mem[o:o + take])
            else:
                mem[o:o + take] = data[pos:pos + take]
            left, pos, done, i = left - take, pos + take, 0, i + 1
        return np.concatenate(out) if data is None else None
    for b in range(nblocks):
        walk(sbv, sit, sbase, smem, b, CFG3_SIZE, lin[b * CFG3_SIZE:(b + 1) * CFG3_SIZE])
    rets, t_c = oracle.compress_batch(lib, shv, sfirst, shi, dhv, dfirst, dhi, nthreads=threads)
    assert (rets > 0).all()
    comp = [walk(dbv, dit, dbase, dmem, b, int(rets[b])) for b in range(nblocks)]
    cbuf = np.concatenate(comp + [np.zeros(64, np.uint8)])
    coff = np.concatenate([[0], np.cumsum(rets)[:-1]]).astype(np.uint64)
    out = np.zeros(nblocks * CFG3_SIZE + 64, dtype=np.uint8)
    r2, t_d = oracle.decompress_batch(lib, cbuf, coff, rets, out,
                                      np.arange(nblocks, dtype=np.uint64) * np.uint64(CFG3_SIZE),
                                      np.full(nblocks, CFG3_SIZE, dtype=np.int32), nthreads=threads)
    assert (r2 == CFG3_SIZE).all() and np.array_equal(out[:nblocks * CFG3_SIZE], lin)
    nbytes = nblocks * CFG3_SIZE
    return {"value": round(nbytes / (t_c + t_d) / 1e9, 4), "unit": UNIT,
            "compress_gbs": round(nbytes / t_c / 1e9, 4), "decompress_gbs": round(nbytes / t_d / 1e9, 4),
            "sample": "first %d blocks (%d MiB) of the configs[2] workload; decompress on the "
                      "linearised compressed blocks (the reference decoder is contiguous-only)"
                      % (nblocks, nbytes >> 20)}


def cpu_baseline(pkg, budget_s=12.0, parity=None):
    from oracle import oracle
    sg = pkg.sg
    lib = oracle.load("auto")
    threads = os.cpu_count() or 1
    kind = "reference" if lib.kind == "reference" else "port"
    src, dst = cpu_sample(sg, lib, 1024, threads)
    t_c, t_d = cpu_run(lib, src, dst, threads)
    per_block = (t_c + t_d) / 1024
    n = int(min(max(budget_s / max(per_block, 1e-9), 2048), 65536))
    src, dst = cpu_sample(sg, lib, n, threads)
    t_c, t_d = cpu_run(lib, src, dst, threads)
    nbytes = n * BLOCK
    res = {"value": round(nbytes / (t_c + t_d) / 1e9, 4), "unit": UNIT,
           "cores": threads, "kind": kind,
           "sample": "first %d blocks (%d MiB) of the configs[1] workload, one block per task, "
                     "one thread per core with its own wrkmem" % (n, nbytes >> 20),
           "compress_gbs": round(nbytes / t_c / 1e9, 4), "decompress_gbs": round(nbytes / t_d / 1e9, 4)}
    try:
        res["cfg3_64k"] = cpu_cfg3(sg, lib, threads, 1024)
    except Exception as e:
        res["cfg3_64k"] = {"error": repr(e)}
    if parity:
        # the oracle as the checker: the GPU's compressed bytes of every 1024th block of the
        # timed workload against the reference's, byte for byte
        k = len(parity)
        arena = sg.HostArena(k + 1)
        ps = sg.layout_from_segments(arena, [dict(sg.segs_single(BLOCK), gap=0)] * k)
        for i, (_, data, _) in enumerate(parity):
            arena.mem[i * BLOCK:(i + 1) * BLOCK] = data
        darena = sg.HostArena(2 * k + 1)
        pd = sg.layout_from_segments(darena, [dict(sg.segs_single(sg.compress_bound(BLOCK)), gap=0)] * k)
        r, _ = oracle.compress_batch(lib, ps.bvecs, ps.first, ps.iters, pd.bvecs, pd.first, pd.iters,
                                     nthreads=threads)
        same = all(int(r[i]) == len(c) and np.array_equal(pd.read_block(i, int(r[i])), c)
                   for i, (_, _, c) in enumerate(parity))
        res["parity_sample"] = {"blocks": k, "byte_identical": bool(same), "checker": kind}
        assert same, "GPU output differs from the oracle on the parity sample"
    return res


def run_reference(args):
    """--impl reference: the reference's own C implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pkg = importlib.import_module("lz4-sgori_b200")  # pure-Python layout helpers only (pkg.sg);
    sg = pkg.sg                                       # the product .so is never loaded here
    from oracle import oracle
    lib = oracle.load("auto")
    threads = os.cpu_count() or 1
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    src, dst = cpu_sample(sg, lib, 512, threads)
    t_c, t_d = cpu_run(lib, src, dst, threads)
    per_block = (t_c + t_d) / 512
    total_steps = args.steps + args.warmup
    n = int(min(max(60.0 / total_steps / max(per_block, 1e-9), 1024), 32768))
    src, dst = cpu_sample(sg, lib, n, threads)
    for _ in range(args.warmup):
        cpu_run(lib, src, dst, threads)
    tc = td = 0.0
    for _ in range(args.steps):
        a, b = cpu_run(lib, src, dst, threads)
        tc, td = tc + a, td + b
    nbytes = n * BLOCK * args.steps
    val = round(nbytes / (tc + td) / 1e9, 4)
    kind = "reference" if lib.kind == "reference" else "port"
    sample = ("each step = first %d blocks (%d MiB) of the configs[1] workload; whole job is %d GiB"
              % (n, (n * BLOCK) >> 20, world))
    cb = {"value": val, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample}
    try:
        cb["cfg3_64k"] = cpu_cfg3(sg, lib, threads, 512)
    except Exception as e:
        cb["cfg3_64k"] = {"error": repr(e)}
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round((tc + td) / args.steps * 1e3, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "configs[1]: %d x 4 KiB blocks per GPU (1 GiB/GPU), 1 bvec each "
                               "(byU16), class = block index mod 4 (text/random/sparse/mixed), "
                               "dst capacity LZ4E_COMPRESSBOUND" % BLOCKS_PER_GPU,
                   "sample": sample},
        "compress_gbs": round(nbytes / tc / 1e9, 4), "decompress_gbs": round(nbytes / td / 1e9, 4),
        "cpu_baseline": cb,
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--no-replay", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-blocks", type=int, default=BLOCKS_PER_GPU)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
