#!/usr/bin/env python
"""tools/latency_bench.py — latency of ONE block (and of 148 equal blocks, one per SM) per
content class and block size: the serial time of a block is what bounds small batches
(BASELINE.json configs[4]).  usage: latency_bench.py [size ...]   (default 4096 65536 131072)"""
import importlib
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "synth"))
import lz4e_synth  # noqa: E402  (tools/synth: benchmark data generator)
pkg = importlib.import_module("lz4-sgori_b200")
NAMES = ["text", "random", "sparse", "mixed"]


def run(ctx, size, cls, n, reps=5):
    bound = pkg.compress_bound(size)
    stride_d = (bound + 4095) & ~4095
    host = np.zeros((n, size), dtype=np.uint8)
    for b in range(n):
        lz4e_synth.lib().lz4e_synth_fill_host(host[b].ctypes.data, size, cls, 4 * b + cls)
    d_src, d_dst, d_out = ctx.upload(host), ctx.dev_alloc(n * stride_d), ctx.dev_alloc(n * size)

    def tab(base, stride, length):
        bv = np.zeros(n, dtype=pkg.DEV_BVEC)
        bv["addr"] = np.uint64(base) + np.arange(n, dtype=np.uint64) * np.uint64(stride)
        bv["len"] = length
        it = np.zeros(n, dtype=pkg.DEV_ITER)
        it["first"] = np.arange(n, dtype=np.uint32)
        it["bi_size"] = length
        return bv, it
    (sbv, sit), (dbv, dit), (obv, oit) = tab(d_src, size, size), tab(d_dst, stride_d, bound), \
        tab(d_out, size, size)
    t = [ctx.upload(x) for x in (sbv, sit, dbv, dit, obv, oit)]
    r1, r2 = ctx.dev_alloc(4 * n), ctx.dev_alloc(4 * n)

    def timed(fn):
        fn()
        ctx.sync()
        best = 1e9
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            ctx.sync()
            best = min(best, time.perf_counter() - t0)
        return best * 1e3
    tc = timed(lambda: ctx.compress_dev(t[0], t[1], t[2], t[3], n, size, r1, flags=pkg.DEV_DST_LINEAR))
    rets = ctx.download(r1, np.int32, n)
    cit = dit.copy()
    cit["bi_size"] = rets.astype(np.uint32)
    d_cit = ctx.upload(cit)
    td = timed(lambda: ctx.decompress_dev(t[2], d_cit, t[4], t[5], n, int(rets.max()), size, r2,
                                          flags=pkg.DEV_DST_LINEAR))
    assert (ctx.download(r2, np.int32, n) == size).all()
    assert np.array_equal(ctx.download(d_out, np.uint8, n * size).reshape(n, size), host)
    for p in (d_src, d_dst, d_out, r1, r2, d_cit, *t):
        ctx.dev_free(p)
    return tc, td, float(rets.mean())


def main():
    sizes = [int(x) for x in sys.argv[1:]] or [4096, 65536, 131072]
    ctx = pkg.Context(0)
    print("| size | class | blocks | compress ms | decompress ms | compressed bytes | compress MB/s per block | decompress MB/s per block |")
    print("|---|---|---|---|---|---|---|---|")
    for size in sizes:
        for cls in range(4):
            for n in (1, 148):
                tc, td, c = run(ctx, size, cls, n)
                print(f"| {size} | {NAMES[cls]} | {n} | {tc:.3f} | {td:.3f} | {c:.0f} | "
                      f"{size / tc / 1e3:.1f} | {size / td / 1e3:.1f} |")
    ctx.close()


if __name__ == "__main__":
    main()
