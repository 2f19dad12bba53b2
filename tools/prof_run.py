#!/usr/bin/env python
"""tools/prof_run.py — a short, fixed sequence of launches for ncu:
   [synth x2] compress(4 KiB) decompress(4 KiB) [synth x2] compress(64 KiB) decompress(64 KiB)
Sizes are small so `ncu --set full` (about 40 replays per kernel) stays cheap."""
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
    n4 = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
    n64 = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    ctx = pkg.Context(0)
    # ---- 4 KiB, one bvec ----
    d_src, d_dst, d_out = ctx.dev_alloc(n4 * 4096), ctx.dev_alloc(n4 * bench.DST_STRIDE), ctx.dev_alloc(n4 * 4096)
    (sbv, sit), (dbv, dit), (obv, oit) = bench.cfg2_tables(pkg, d_src, d_dst, d_out, n4)
    t = [ctx.upload(x) for x in (sbv, sit, dbv, dit, obv, oit)]
    r1, r2 = ctx.dev_alloc(4 * n4), ctx.dev_alloc(4 * n4)
    ctx.sync(); lz4e_synth.dev(0, t[0], t[1], n4, 0, 4096)
    ctx.compress_dev(t[0], t[1], t[2], t[3], n4, 4096, r1, flags=pkg.DEV_DST_LINEAR)
    ctx.sync()
    t0 = time.perf_counter()
    for _ in range(reps):
        ctx.compress_dev(t[0], t[1], t[2], t[3], n4, 4096, r1, flags=pkg.DEV_DST_LINEAR)
    ctx.sync()
    tc = (time.perf_counter() - t0) / reps
    rets = ctx.download(r1, np.int32, n4)
    cit = dit.copy()
    cit["bi_size"] = rets.astype(np.uint32)
    d_cit = ctx.upload(cit)
    ctx.decompress_dev(t[2], d_cit, t[4], t[5], n4, int(rets.max()), 4096, r2, flags=pkg.DEV_DST_LINEAR)
    ctx.sync()
    t0 = time.perf_counter()
    for _ in range(reps):
        ctx.decompress_dev(t[2], d_cit, t[4], t[5], n4, int(rets.max()), 4096, r2, flags=pkg.DEV_DST_LINEAR)
    ctx.sync()
    td = (time.perf_counter() - t0) / reps
    assert (ctx.download(r2, np.int32, n4) == 4096).all()
    print("4k ok", n4, int(rets.sum()), "compress %.2f GB/s  decompress %.2f GB/s (wall, incl. launch)" %
          (n4 * 4096 / tc / 1e9, n4 * 4096 / td / 1e9))
    # ---- 64 KiB, 16 odd segments ----
    if n64:
        bound = pkg.compress_bound(65536)
        region = n64 * 16 * 3 * 4096
        a, b, c = ctx.dev_alloc(region), ctx.dev_alloc(region), ctx.dev_alloc(region)
        sbv, sit, _ = bench.cfg3_tables(pkg, a, n64, 65536, 65536, 0)
        dbv, dit, _ = bench.cfg3_tables(pkg, b, n64, bound, bound, 11)
        obv, oit, _ = bench.cfg3_tables(pkg, c, n64, 65536, 65536, 5)
        t = [ctx.upload(x) for x in (sbv, sit, dbv, dit, obv, oit)]
        r1, r2 = ctx.dev_alloc(4 * n64), ctx.dev_alloc(4 * n64)
        ctx.sync(); lz4e_synth.dev(0, t[0], t[1], n64, 0, 65536)
        ctx.compress_dev(t[0], t[1], t[2], t[3], n64, 65536, r1, flags=pkg.DEV_NO_BYU16)
        ctx.sync()
        t0 = time.perf_counter()
        for _ in range(reps):
            ctx.compress_dev(t[0], t[1], t[2], t[3], n64, 65536, r1, flags=pkg.DEV_NO_BYU16)
        ctx.sync()
        tc = (time.perf_counter() - t0) / reps
        rets = ctx.download(r1, np.int32, n64)
        cit = dit.copy()
        cit["bi_size"] = rets.astype(np.uint32)
        d_cit = ctx.upload(cit)
        ctx.decompress_dev(t[2], d_cit, t[4], t[5], n64, int(rets.max()), 65536, r2)
        ctx.sync()
        t0 = time.perf_counter()
        for _ in range(reps):
            ctx.decompress_dev(t[2], d_cit, t[4], t[5], n64, int(rets.max()), 65536, r2)
        ctx.sync()
        td = (time.perf_counter() - t0) / reps
        assert (ctx.download(r2, np.int32, n64) == 65536).all()
        print("64k ok", n64, int(rets.sum()), "compress %.2f GB/s  decompress %.2f GB/s (wall, incl. launch)" %
              (n64 * 65536 / tc / 1e9, n64 * 65536 / td / 1e9))
    ctx.close()


if __name__ == "__main__":
    main()
