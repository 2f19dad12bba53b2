#!/usr/bin/env python
"""tools/dbench.py — decoder variants side by side on one B200: the 4 KiB config (linear
source/destination) and the 64 KiB config (16 odd segments per side), device-resident, CUDA
events on the launching stream, round trip checked for every variant.
usage: dbench.py [n4k [n64k [reps]]]"""
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools", "synth"))
import bench  # noqa: E402
import lz4e_synth  # noqa: E402

pkg = importlib.import_module("lz4-sgori_b200")


def main():
    import torch
    n4 = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
    n64 = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    ctx = pkg.Context(0)
    stream = torch.cuda.Stream(device=0)
    sp = stream.cuda_stream
    peak, _ = bench.peaks()

    def timed(fn):
        ms = []
        for _ in range(2):
            fn()
        ctx.sync(sp)
        for _ in range(reps):
            e = bench.event_ms(torch, fn, stream)
            ctx.sync(sp)
            ms.append(e[0].elapsed_time(e[1]))
        return float(np.mean(ms)), float(np.min(ms))

    only = os.environ.get("DBENCH_ONLY")  # "variant,group,stage": just that decoder

    def run(name, n, size, t, cit_dev, rets, flags, src_ptr, out_check, variants):
        nbytes, cbytes = n * size, int(rets.sum())
        if only:
            variants = [tuple(int(x) for x in (only + ",1").split(",")[:4])]
        r2 = ctx.dev_alloc(4 * n)
        for var, grp, stg, order in variants:
            ctx.tune("dvariant", var)
            ctx.tune("dgroup", grp)
            ctx.tune("dstage", stg)
            ctx.tune("dorder", 2048 if order else 0)
            try:
                mean, best = timed(lambda: ctx.decompress_dev(t[2], cit_dev, t[4], t[5], n, int(rets.max()),
                                                              size, r2, flags=flags, stream=sp))
                ok = bool((ctx.download(r2, np.int32, n) == size).all()) and out_check()
            except Exception as e:
                print(json.dumps({"cfg": name, "variant": var, "group": grp, "stage": stg, "error": repr(e)}))
                continue
            print(json.dumps({"cfg": name, "variant": var, "group": grp, "stage": stg, "order": order,
                              "ms": round(mean, 3), "ms_best": round(best, 3),
                              "out_gbs": round(nbytes / mean / 1e6, 1),
                              "frac": round((nbytes + cbytes) / mean / 1e6 / peak, 4), "ok": ok}), flush=True)
        ctx.dev_free(r2)

    if n4:
        d_src, d_dst, d_out = ctx.dev_alloc(n4 * 4096), ctx.dev_alloc(n4 * bench.DST_STRIDE), ctx.dev_alloc(n4 * 4096)
        (sbv, sit), (dbv, dit), (obv, oit) = bench.cfg2_tables(pkg, d_src, d_dst, d_out, n4)
        t = [ctx.upload(x) for x in (sbv, sit, dbv, dit, obv, oit)]
        r1 = ctx.dev_alloc(4 * n4)
        lz4e_synth.dev(0, t[0], t[1], n4, 0, 4096)
        mean, _ = timed(lambda: ctx.compress_dev(t[0], t[1], t[2], t[3], n4, 4096, r1,
                                                 flags=pkg.DEV_DST_LINEAR, stream=sp))
        rets = ctx.download(r1, np.int32, n4)
        print(json.dumps({"cfg": "4k", "compress_ms": round(mean, 3),
                          "compress_gbs": round(n4 * 4096 / mean / 1e6, 1)}), flush=True)
        cit = dit.copy()
        cit["bi_size"] = rets.astype(np.uint32)
        d_cit = ctx.upload(cit)
        src_h = ctx.download(d_src, np.uint8, min(n4 * 4096, 256 << 20))

        def chk():
            return bool(np.array_equal(src_h, ctx.download(d_out, np.uint8, len(src_h))))
        lin = pkg.DEV_DST_LINEAR | pkg.DEV_SRC_LINEAR
        run("4k-linear", n4, 4096, t, d_cit, rets, lin, d_src, chk,
            [(3, 32, 0, 0), (3, 32, 0, 1), (0, 1, -1, 0), (0, 1, -1, 1), (0, 8, 0, 0), (2, 8, 0, 0), (2, 8, 0, 1), (2, 4, 0, 1),
             (2, 16, 0, 1), (1, 8, 0, 1), (1, 4, 0, 1), (2, 8, 1, 1), (2, 4, 1, 1)])
        for p in (d_src, d_dst, d_out, r1, d_cit, *t):
            ctx.dev_free(p)
    if n64:
        S = 65536
        bound = pkg.compress_bound(S)
        region = n64 * 16 * 3 * 4096
        a, b, c = ctx.dev_alloc(region), ctx.dev_alloc(region), ctx.dev_alloc(region)
        sbv, sit, _ = bench.cfg3_tables(pkg, a, n64, S, S, 0)
        dbv, dit, _ = bench.cfg3_tables(pkg, b, n64, bound, bound, 11)
        obv, oit, _ = bench.cfg3_tables(pkg, c, n64, S, S, 5)
        t = [ctx.upload(x) for x in (sbv, sit, dbv, dit, obv, oit)]
        r1 = ctx.dev_alloc(4 * n64)
        lz4e_synth.dev(0, t[0], t[1], n64, 0, S)
        mean, _ = timed(lambda: ctx.compress_dev(t[0], t[1], t[2], t[3], n64, S, r1,
                                                 flags=pkg.DEV_NO_BYU16, stream=sp))
        rets = ctx.download(r1, np.int32, n64)
        print(json.dumps({"cfg": "64k", "compress_ms": round(mean, 3),
                          "compress_gbs": round(n64 * S / mean / 1e6, 1)}), flush=True)
        cit = dit.copy()
        cit["bi_size"] = rets.astype(np.uint32)
        d_cit = ctx.upload(cit)
        sample = list(range(0, n64, max(1, n64 // 16)))
        want = [bench._gather_dev(ctx, sbv, sit, i, S) for i in sample]

        def chk64():
            return all(np.array_equal(w, bench._gather_dev(ctx, obv, oit, i, S)) for w, i in zip(want, sample))
        run("64k-sg16", n64, S, t, d_cit, rets, pkg.DEV_DST_SEGS32, a, chk64,
            [(3, 32, 0, 0), (3, 32, 0, 1), (0, 8, 0, 0), (2, 8, 0, 0), (2, 8, 0, 1), (2, 4, 0, 1), (2, 16, 0, 1), (1, 8, 0, 1),
             (1, 4, 0, 1)])
    ctx.close()


if __name__ == "__main__":
    main()
