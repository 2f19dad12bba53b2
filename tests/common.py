"""Shared helpers for the tests: workload construction and oracle runs."""
import importlib
import os
import sys

import numpy as np

from oracle import oracle

pkg = importlib.import_module("lz4-sgori_b200")
sg = pkg.sg
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                "tools", "synth"))
import lz4e_synth  # noqa: E402  (benchmark data generator, tools/synth)


def synth(n, cls, seed):
    return lz4e_synth.host(n, cls, seed)


def synth_fill_dev(ctx, d_bv, d_it, nblocks, max_bytes, first_index=0):
    """Device-resident blocks filled by the generator's own kernels (tools/synth)."""
    ctx.sync()
    lz4e_synth.dev(ctx.device, d_bv, d_it, nblocks, first_index, max_bytes)


def lcg_text(n, seed=12345):
    """SURVEY.md section 8(c) generator: LCG-picked words + ' '."""
    words = ("lorem ipsum dolor sit amet consectetur adipiscing elit sed do eiusmod "
             "tempor").split()
    s, out = seed, bytearray()
    while len(out) < n:
        s = (s * 1664525 + 1013904223) & 0xFFFFFFFF
        out += words[(s >> 16) % 12].encode() + b" "
    return np.frombuffer(bytes(out[:n]), dtype=np.uint8)


def build_layout(arena_factory, specs, datas=None):
    arena = arena_factory(sg.pages_needed(specs) + 2)
    lay = sg.layout_from_segments(arena, specs)
    if datas is not None:
        for b, d in enumerate(datas):
            lay.write_block(b, d)
    return lay


def oracle_compress(lib, src, dst_specs, nthreads=4, want_iters=False):
    """Run the oracle over `src` into a fresh host arena laid out by dst_specs."""
    dst = build_layout(sg.HostArena, dst_specs)
    dst.arena.mem[:] = 0xEE
    res = oracle.compress_batch(lib, src.bvecs, src.first, src.iters, dst.bvecs, dst.first,
                                dst.iters, nthreads=nthreads, want_iters=want_iters)
    return dst, res


def oracle_decode_linear(lib, comps, caps, nthreads=4):
    """Reference decoder on contiguous buffers: returns (rets, list of outputs)."""
    lens = [len(c) for c in comps]
    src = np.concatenate([np.asarray(c, dtype=np.uint8) for c in comps] +
                         [np.zeros(64, dtype=np.uint8)])
    soff = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint64)
    slots = [c + 64 for c in caps]
    doff = np.concatenate([[0], np.cumsum(slots)[:-1]]).astype(np.uint64)
    dst = np.full(int(sum(slots)) + 64, 0xEE, dtype=np.uint8)
    rets, _ = oracle.decompress_batch(lib, src, soff, lens, dst, doff, caps, nthreads=nthreads)
    outs = [dst[int(o):int(o) + max(int(r), 0)].copy() for o, r in zip(doff, rets)]
    return rets, outs


class DevTables:
    """Upload the descriptor tables of a Layout; frees on close()."""

    def __init__(self, ctx, layout):
        self.ctx = ctx
        self.bv_np, self.it_np = pkg.to_dev(layout)
        self.bv = ctx.upload(self.bv_np)
        self.it = ctx.upload(self.it_np)
        self.n = layout.nblocks()

    def with_sizes(self, sizes):
        it = self.it_np.copy()
        it["bi_size"] = np.asarray(sizes, dtype=np.uint32)
        return self.ctx.upload(it)

    def close(self):
        self.ctx.dev_free(self.bv)
        self.ctx.dev_free(self.it)


def gpu_compress(ctx, src, dst, max_in=None, flags=0, want_ends=False):
    """Device-level compress of Layout src -> Layout dst (both over pkg.Arena).  Returns rets
    (and ends).  Uploads the source arena, downloads the destination arena."""
    n = src.nblocks()
    src.arena.upload()
    dst.arena.upload()
    s, d = DevTables(ctx, src), DevTables(ctx, dst)
    d_rets = ctx.dev_alloc(4 * n)
    d_ends = ctx.dev_alloc(8 * n) if want_ends else None
    if max_in is None:
        max_in = int(src.iters["bi_size"].max()) if n else 0
    ctx.compress_dev(s.bv, s.it, d.bv, d.it, n, max_in, d_rets, d_ends, flags=flags)
    ctx.sync()
    rets = ctx.download(d_rets, np.int32, n)
    ends = ctx.download(d_ends, pkg.DEV_END, n) if want_ends else None
    dst.arena.download()
    s.close()
    d.close()
    ctx.dev_free(d_rets)
    if want_ends:
        ctx.dev_free(d_ends)
        return rets, ends
    return rets


def gpu_decompress(ctx, src, sizes, dst, flags=0):
    """Device-level decompress: block b reads sizes[b] bytes of Layout src, writes Layout dst."""
    n = src.nblocks()
    src.arena.upload()
    dst.arena.upload()
    s, d = DevTables(ctx, src), DevTables(ctx, dst)
    sit = s.with_sizes(sizes)
    d_rets = ctx.dev_alloc(4 * n)
    ctx.decompress_dev(s.bv, sit, d.bv, d.it, n, int(max(sizes)) if n else 0,
                       int(dst.iters["bi_size"].max()) if n else 0, d_rets, flags=flags)
    ctx.sync()
    rets = ctx.download(d_rets, np.int32, n)
    dst.arena.download()
    s.close()
    d.close()
    ctx.dev_free(sit)
    ctx.dev_free(d_rets)
    return rets
