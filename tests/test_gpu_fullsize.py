"""BASELINE.json configs at FULL size on one GPU, through size-independent properties (the
oracle cannot run 1 GiB in seconds): bit-exact round trip of every block, the known size of
every incompressible block, and byte-identity with the oracle on a deterministic sample
(every 64th / 512th block).  Data is generated on the device by the same generator the host
uses (lz4e_synth.h), which the sample also checks."""
import os
import sys

import numpy as np
import pytest

import common
from common import pkg, sg

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402  (table builders of the benchmark workloads)

pytestmark = pytest.mark.gpu


def _gather(ctx, bv, it, b, nbytes):
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


def test_config2_full_size(ctx, ref_lib):
    """configs[1]: 262144 x 4 KiB blocks (1 GiB), class = block index mod 4."""
    n, B = bench.BLOCKS_PER_GPU, bench.BLOCK
    d_src, d_dst, d_out = (ctx.dev_alloc(n * B), ctx.dev_alloc(n * bench.DST_STRIDE),
                           ctx.dev_alloc(n * B))
    (sbv, sit), (dbv, dit), (obv, oit) = bench.cfg2_tables(pkg, d_src, d_dst, d_out, n)
    t = [ctx.upload(x) for x in (sbv, sit, dbv, dit, obv, oit)]
    r1, r2 = ctx.dev_alloc(4 * n), ctx.dev_alloc(4 * n)
    common.synth_fill_dev(ctx, t[0], t[1], n, B)
    ctx.compress_dev(t[0], t[1], t[2], t[3], n, B, r1, flags=pkg.DEV_DST_LINEAR)
    ctx.sync()
    rets = ctx.download(r1, np.int32, n)
    assert (rets > 0).all()
    # incompressible blocks expand to 4114 unless 4 bytes repeat by chance
    assert (rets[1::4] <= 4114).all() and (rets[1::4] == 4114).mean() > 0.99
    cit = dit.copy()
    cit["bi_size"] = rets.astype(np.uint32)
    d_cit = ctx.upload(cit)
    ctx.decompress_dev(t[2], d_cit, t[4], t[5], n, int(rets.max()), B, r2,
                       flags=pkg.DEV_DST_LINEAR)
    ctx.sync()
    assert (ctx.download(r2, np.int32, n) == B).all()
    src = ctx.download(d_src, np.uint8, n * B)
    assert np.array_equal(src, ctx.download(d_out, np.uint8, n * B)), "round trip"
    # sample against the oracle
    sample = list(range(0, n, 64)) + [1, 2, 3, n - 1]
    datas = [common.synth(B, b & 3, b) for b in sample]
    for b, d in zip(sample, datas):
        assert np.array_equal(src[b * B:(b + 1) * B], d), f"device generator, block {b}"
    hsrc = common.build_layout(sg.HostArena, [sg.segs_single(B)] * len(sample), datas)
    ref_dst, (ref_rets, _) = common.oracle_compress(
        ref_lib, hsrc, [sg.segs_single(sg.compress_bound(B))] * len(sample))
    assert np.array_equal(rets[sample], ref_rets)
    for k, b in enumerate(sample):
        got = ctx.download(d_dst + b * bench.DST_STRIDE, np.uint8, int(rets[b]))
        assert np.array_equal(got, ref_dst.read_block(k, ref_rets[k])), b
    for p in (d_src, d_dst, d_out, r1, r2, d_cit, *t):
        ctx.dev_free(p)


def test_config3_one_gpu_share(ctx, ref_lib):
    """configs[2] as one of 8 GPUs sees it: 32768 x 64 KiB blocks (2 GiB), every block over 16
    odd non-contiguous segments with a mid-bvec start on both sides (byU32)."""
    n, S = 32768, 65536
    bound = pkg.compress_bound(S)
    region = n * 16 * 3 * 4096
    a, b_, c = ctx.dev_alloc(region), ctx.dev_alloc(region), ctx.dev_alloc(region)
    sbv, sit, _ = bench.cfg3_tables(pkg, a, n, S, S, 0)
    dbv, dit, _ = bench.cfg3_tables(pkg, b_, n, bound, bound, 11)
    obv, oit, _ = bench.cfg3_tables(pkg, c, n, S, S, 5)
    t = [ctx.upload(x) for x in (sbv, sit, dbv, dit, obv, oit)]
    r1, r2 = ctx.dev_alloc(4 * n), ctx.dev_alloc(4 * n)
    common.synth_fill_dev(ctx, t[0], t[1], n, S)
    ctx.compress_dev(t[0], t[1], t[2], t[3], n, S, r1, flags=pkg.DEV_NO_BYU16)
    ctx.sync()
    rets = ctx.download(r1, np.int32, n)
    assert (rets > 0).all()
    # random 64 KiB -> 65794 unless 4 bytes repeat by chance
    assert (rets[1::4] <= 65794).all() and (rets[1::4] == 65794).mean() > 0.9
    cit = dit.copy()
    cit["bi_size"] = rets.astype(np.uint32)
    d_cit = ctx.upload(cit)
    ctx.decompress_dev(t[2], d_cit, t[4], t[5], n, int(rets.max()), S, r2)
    ctx.sync()
    assert (ctx.download(r2, np.int32, n) == S).all()
    sample = list(range(0, n, 512)) + [1, 2, 3, n - 1]
    datas = [common.synth(S, b & 3, b) for b in sample]
    hsrc = common.build_layout(sg.HostArena, [sg.segs_odd(S, 16, b) for b in sample], datas)
    ref_dst, (ref_rets, _) = common.oracle_compress(
        ref_lib, hsrc, [sg.segs_single(bound)] * len(sample))
    assert np.array_equal(rets[sample], ref_rets)
    for k, b in enumerate(sample):
        assert np.array_equal(_gather(ctx, sbv, sit, b, S), datas[k]), f"generator, block {b}"
        assert np.array_equal(_gather(ctx, obv, oit, b, S), datas[k]), f"round trip, block {b}"
        assert np.array_equal(_gather(ctx, dbv, dit, b, int(rets[b])),
                              ref_dst.read_block(k, ref_rets[k])), f"compressed bytes, block {b}"
    for p in (a, b_, c, r1, r2, d_cit, *t):
        ctx.dev_free(p)
