#!/usr/bin/env python
"""tools/class_bench.py — per content class timing (4 KiB blocks, one bvec):
usage: class_bench.py [nblocks] [reps]"""
import importlib
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "synth"))
import lz4e_synth  # noqa: E402  (tools/synth: benchmark data generator)
import bench  # noqa: E402

pkg = importlib.import_module("lz4-sgori_b200")


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    ctx = pkg.Context(0)
    S = int(os.environ.get("CLASS_SIZE", "4096"))  # block size (one bvec per block)
    if S != 4096:
        return main_sized(ctx, n, reps, S)
    d_src, d_dst, d_out = ctx.dev_alloc(n * 4096), ctx.dev_alloc(n * bench.DST_STRIDE), ctx.dev_alloc(n * 4096)
    (sbv, sit), (dbv, dit), (obv, oit) = bench.cfg2_tables(pkg, d_src, d_dst, d_out, n)
    cflags = pkg.DEV_DST_LINEAR
    if os.environ.get("CLASS_U32"):  # occupancy probe: raw bv_len > 4096 selects byU32 (8 KiB table)
        sbv["len"] = 8192
        cflags |= pkg.DEV_NO_BYU16
    t = [ctx.upload(x) for x in (sbv, sit, dbv, dit, obv, oit)]
    r1, r2 = ctx.dev_alloc(4 * n), ctx.dev_alloc(4 * n)
    host = np.zeros((n, 4096), dtype=np.uint8)
    only = os.environ.get("CLASS_ONLY")
    for cls in ([int(only)] if only else range(4)):
        for b in range(n):
            lz4e_synth.lib().lz4e_synth_fill_host(host[b].ctypes.data, 4096, cls, 4 * b + cls)
        ctx.check(pkg.lib().lz4e_b200_copy_h2d(ctx.h, None, d_src, host.ctypes.data, n * 4096), "h2d")
        ctx.sync()
        ctx.compress_dev(t[0], t[1], t[2], t[3], n, 4096, r1, flags=cflags)
        ctx.sync()
        t0 = time.perf_counter()
        for _ in range(reps):
            ctx.compress_dev(t[0], t[1], t[2], t[3], n, 4096, r1, flags=cflags)
        ctx.sync()
        tc = (time.perf_counter() - t0) / reps
        rets = ctx.download(r1, np.int32, n)
        cit = dit.copy()
        cit["bi_size"] = rets.astype(np.uint32)
        d_cit = ctx.upload(cit)
        ctx.decompress_dev(t[2], d_cit, t[4], t[5], n, int(rets.max()), 4096, r2, flags=pkg.DEV_DST_LINEAR)
        ctx.sync()
        t0 = time.perf_counter()
        for _ in range(reps):
            ctx.decompress_dev(t[2], d_cit, t[4], t[5], n, int(rets.max()), 4096, r2, flags=pkg.DEV_DST_LINEAR)
        ctx.sync()
        td = (time.perf_counter() - t0) / reps
        assert (ctx.download(r2, np.int32, n) == 4096).all()
        print("class %d: ratio %.2f  compress %.2f GB/s (%.1f us/block/warp-slot)  decompress %.2f GB/s" %
              (cls, n * 4096 / rets.sum(), n * 4096 / tc / 1e9, tc * 1e6, n * 4096 / td / 1e9))
    ctx.close()


def main_sized(ctx, n, reps, S):
    """Same per-class timing for blocks of S bytes (one multi-page bvec each: byU32)."""
    bound = pkg.compress_bound(S)
    stride = (bound + 255) & ~255
    d_src, d_dst, d_out = ctx.dev_alloc(n * S), ctx.dev_alloc(n * stride), ctx.dev_alloc(n * S)
    idx = np.arange(n, dtype=np.uint64)

    def tab(base, st, length):
        bv = np.zeros(n, dtype=pkg.DEV_BVEC)
        bv["addr"] = np.uint64(base) + idx * np.uint64(st)
        bv["len"] = length
        it = np.zeros(n, dtype=pkg.DEV_ITER)
        it["first"] = idx.astype(np.uint32)
        it["bi_size"] = length
        return bv, it
    (sbv, sit), (dbv, dit), (obv, oit) = tab(d_src, S, S), tab(d_dst, stride, bound), tab(d_out, S, S)
    t = [ctx.upload(x) for x in (sbv, sit, dbv, dit, obv, oit)]
    r1, r2 = ctx.dev_alloc(4 * n), ctx.dev_alloc(4 * n)
    host = np.zeros((n, S), dtype=np.uint8)
    only = os.environ.get("CLASS_ONLY")
    cf = pkg.DEV_DST_LINEAR | pkg.DEV_NO_BYU16
    df = pkg.DEV_DST_LINEAR | pkg.DEV_SRC_LINEAR
    for cls in ([int(only)] if only else range(4)):
        for b in range(n):
            lz4e_synth.lib().lz4e_synth_fill_host(host[b].ctypes.data, S, cls, 4 * b + cls)
        ctx.check(pkg.lib().lz4e_b200_copy_h2d(ctx.h, None, d_src, host.ctypes.data, n * S), "h2d")
        ctx.sync()
        res = []
        for what in ("c", "d"):
            if what == "c":
                fn = lambda: ctx.compress_dev(t[0], t[1], t[2], t[3], n, S, r1, flags=cf)  # noqa: E731
            else:
                rets = ctx.download(r1, np.int32, n)
                cit = dit.copy()
                cit["bi_size"] = rets.astype(np.uint32)
                d_cit = ctx.upload(cit)
                fn = lambda: ctx.decompress_dev(t[2], d_cit, t[4], t[5], n, int(rets.max()), S, r2, flags=df)  # noqa: E731
            fn()
            ctx.sync()
            t0 = time.perf_counter()
            for _ in range(reps):
                fn()
            ctx.sync()
            res.append((time.perf_counter() - t0) / reps)
        assert (ctx.download(r2, np.int32, n) == S).all()
        print("class %d (%d B blocks): ratio %.2f  compress %.2f GB/s  decompress %.2f GB/s" %
              (cls, S, n * S / rets.sum(), n * S / res[0] / 1e9, n * S / res[1] / 1e9))
    ctx.close()


if __name__ == "__main__":
    main()
