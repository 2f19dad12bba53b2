"""lz4e_b200_compress_dev_static: the compress entry point that allocates nothing (its scratch is a
fixed array in the library's device image, include/lz4e_b200.h).  It has no counterpart in the
reference; the check is against lz4e_b200_compress_dev on the same inputs, bit for bit, on two
calls in a row that share the array.  Collected last on purpose
(file name)."""
import numpy as np
import pytest

import common
from common import pkg, sg

pytestmark = pytest.mark.gpu


def _arena(ctx):
    return lambda n: pkg.Arena(ctx, n)


class _Batch:
    """One batch resident on the device: source and destination arenas, descriptor tables,
    result arrays."""

    def __init__(self, ctx, datas, specs, dspecs, flags):
        self.ctx, self.flags, self.n = ctx, flags, len(datas)
        self.src = common.build_layout(_arena(ctx), specs, datas)
        self.dst = common.build_layout(_arena(ctx), dspecs)
        self.dst.arena.mem[:] = 0xEE
        self.src.arena.upload()
        self.dst.arena.upload()
        self.s, self.d = common.DevTables(ctx, self.src), common.DevTables(ctx, self.dst)
        self.d_rets = ctx.dev_alloc(4 * self.n)
        self.d_ends = ctx.dev_alloc(8 * self.n)
        self.max_in = max(len(d) for d in datas)

    def launch(self, fn):
        fn(self.s.bv, self.s.it, self.d.bv, self.d.it, self.n, self.max_in, self.d_rets,
           self.d_ends, flags=self.flags)

    def fetch(self):
        rets = self.ctx.download(self.d_rets, np.int32, self.n)
        ends = self.ctx.download(self.d_ends, pkg.DEV_END, self.n)
        self.dst.arena.download()
        return rets, ends, self.dst.arena.mem.copy()

    def close(self):
        self.s.close()
        self.d.close()
        self.ctx.dev_free(self.d_rets)
        self.ctx.dev_free(self.d_ends)
        self.src.arena.close()
        self.dst.arena.close()


def _batches(seed):
    """(datas, src specs, dst specs, flags): 64 KiB blocks over 16 odd segments into SG
    destinations — more blocks than the array has 64 KiB slots, so slots are reused — and ragged
    blocks up to 4 KiB written in place, some with too little room (0 returns)."""
    rng = np.random.default_rng(seed)
    big = [common.synth(65536, b & 3, seed * 1000 + b) for b in range(300)]
    bspecs = [sg.segs_odd(65536, 16, b) for b in range(300)]
    bdst = [sg.segs_odd(sg.compress_bound(65536), 16, b + 7) for b in range(300)]
    small, sspecs, sdst = [], [], []
    for b in range(1500):
        n = int(rng.choice([0, 1, 12, 13, 100, 4095, 4096])) if rng.random() < 0.3 \
            else int(rng.integers(0, 4097))
        small.append(common.synth(n, int(rng.integers(0, 4)), seed * 7777 + b))
        sspecs.append(sg.segs_random(n, rng, max_segs=6))
        cap = sg.compress_bound(n) if b % 9 else max(n // 3, 1)
        sdst.append(dict(sg.segs_single(cap), size=cap))
    return [(big, bspecs, bdst, 0), (small, sspecs, sdst, pkg.DEV_DST_LINEAR)]


def _run_pair(ctx, fn_name, seed):
    """Both batches through one entry point, launched back to back with no synchronisation in
    between (same stream: they run one after the other and share the array)."""
    bs = [_Batch(ctx, *b) for b in _batches(seed)]
    fn = getattr(ctx, fn_name)
    for b in bs:
        b.launch(fn)
    ctx.sync()
    out = [b.fetch() for b in bs]
    for b in bs:
        b.close()
    return out


@pytest.mark.parametrize("seed", [1, 2])
def test_static_scratch_equals_compress_dev(ctx, seed):
    got = _run_pair(ctx, "compress_dev_static", seed)
    want = _run_pair(ctx, "compress_dev", seed)
    for (r0, e0, m0), (r1, e1, m1) in zip(got, want):
        assert np.array_equal(r0, r1)
        assert np.array_equal(e0["src_pos"], e1["src_pos"])
        assert np.array_equal(e0["dst_pos"], e1["dst_pos"])
        assert np.array_equal(m0, m1)  # every destination byte, written or not
    assert (got[0][0] > 0).all() and (got[1][0] == 0).any()


def test_static_scratch_rejects_large_blocks(ctx):
    with pytest.raises(pkg.Lz4eError, match="max_in_bytes 65537 is above"):
        ctx.compress_dev_static(None, None, None, None, 1, 65537, None)
