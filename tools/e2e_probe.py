#!/usr/bin/env python
"""tools/e2e_probe.py — where the host-level batch time goes: raw PCIe copy rates, then
compress_batch / decompress_batch separately (set LZ4E_B200_TRACE=1 for the C-side split).
usage: e2e_probe.py [nblocks]"""
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
sg = pkg.sg
BLOCK, DST_STRIDE = bench.BLOCK, bench.DST_STRIDE


def main():
    nb = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
    ctx = pkg.Context(0)
    src_arena = pkg.Arena(ctx, nb + 1)
    dst_arena = pkg.Arena(ctx, (nb * DST_STRIDE + 4095) // 4096 + 2)
    out_arena = pkg.Arena(ctx, nb + 1)
    for name, fn in (("h2d", src_arena.upload), ("d2h", src_arena.download)):
        fn()
        t0 = time.perf_counter()
        fn()
        dt = time.perf_counter() - t0
        print("%s %.1f MiB: %.2f ms = %.1f GB/s" % (name, src_arena.npages * 4096 / 2**20, dt * 1e3,
                                                   src_arena.npages * 4096 / dt / 1e9))
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
    lz4e_synth.host_blocks(src_arena.mem, BLOCK, BLOCK, nb, first=0)
    creqs, keep1 = pkg.make_reqs(src, dst)
    for rep in range(3):
        t0 = time.perf_counter()
        assert ctx.compress_batch(creqs, flags=pkg.F_KEEP_ITERS) == 0, ctx.err()
        t1 = time.perf_counter()
        sizes = creqs["ret"].copy()
        csrc = sg.Layout(dst_arena, dst.bvecs, dst.first, dst.iters.copy(), dst.npages_used)
        csrc.iters["bi_size"] = sizes.astype(np.uint32)
        dreqs, keep2 = pkg.make_reqs(csrc, out)
        t2 = time.perf_counter()
        assert ctx.decompress_batch(dreqs) == 0, ctx.err()
        t3 = time.perf_counter()
        print("rep %d: compress_batch %.2f ms, python glue %.2f ms, decompress_batch %.2f ms"
              % (rep, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3))
    assert np.array_equal(out_arena.mem[:nb * BLOCK], src_arena.mem[:nb * BLOCK])
    ctx.close()


if __name__ == "__main__":
    main()
