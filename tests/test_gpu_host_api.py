"""GPU tests of the HOST-level API: the reference's own entry points (LZ4E_compress_default,
LZ4E_decompress_safe; lz4e/include/lz4e.h:47-51) and the batched form of the same argument
tuple (lz4e_b200_compress_batch / lz4e_b200_decompress_batch), against the oracle.  Covers
pages inside pinned arenas (bulk copies, chunks pipelined over several streams), pages
anywhere in host memory (bounced), iterator end states, and compacted output."""
import ctypes
import os

import numpy as np
import pytest

import common
from common import pkg, sg

pytestmark = pytest.mark.gpu


def _arena(ctx):
    return lambda npages: pkg.Arena(ctx, npages)


def _vp(a):
    return ctypes.c_void_p(a.ctypes.data)


def _batch_roundtrip(ctx, ref_lib, datas, specs, dspecs, ospecs, src_fac, dst_fac, out_fac,
                     flags=0):
    src = common.build_layout(src_fac, specs, datas)
    dst = common.build_layout(dst_fac, dspecs)
    dst.arena.mem[:] = 0xEE
    creqs, keep = pkg.make_reqs(src, dst)
    assert ctx.compress_batch(creqs, flags=flags) == 0, ctx.err()
    rets = creqs["ret"].copy()
    ref_dst, (ref_rets, _, so, do) = common.oracle_compress(ref_lib, src, dspecs, want_iters=True)
    assert np.array_equal(rets, ref_rets)
    for b in range(len(datas)):
        assert np.array_equal(dst.read_block(b, rets[b]), ref_dst.read_block(b, ref_rets[b])), b
    if not (flags & pkg.F_KEEP_ITERS):
        # *srcIter / *dstIter end where the reference leaves them (lz4e_compress.c:304-305,
        # :528-529)
        si, di = keep
        for f in ("bi_size", "bi_idx", "bi_bvec_done"):
            assert np.array_equal(si[f], so[f]), f
            assert np.array_equal(di[f], do[f]), f
    out = common.build_layout(out_fac, ospecs)
    csrc = sg.Layout(dst.arena, dst.bvecs, dst.first, dst.iters.copy(), dst.npages_used)
    csrc.iters["bi_size"] = rets.astype(np.uint32)
    dreqs, keep2 = pkg.make_reqs(csrc, out)
    assert ctx.decompress_batch(dreqs) == 0, ctx.err()
    for b, d in enumerate(datas):
        assert dreqs["ret"][b] == len(d), b
        assert np.array_equal(out.read_block(b), d), b
    return rets


@pytest.mark.parametrize("chunk_mb", ["1", "32"])
def test_batch_arena_pipelined(ctx, ref_lib, chunk_mb, monkeypatch):
    """4 KiB blocks in three separate arenas; with 1 MiB chunks the batch runs as many chunks
    overlapped on several streams and must give the same bytes."""
    monkeypatch.setenv("LZ4E_B200_CHUNK_MB", chunk_mb)
    n = 1500
    datas = [common.synth(4096, b & 3, 77 + b) for b in range(n)]
    specs = [dict(sg.segs_single(4096), gap=0)] * n
    dspecs = [dict(sg.segs_single(sg.compress_bound(4096)), gap=0)] * n
    _batch_roundtrip(ctx, ref_lib, datas, specs, dspecs, specs, _arena(ctx), _arena(ctx),
                     _arena(ctx))


def test_batch_mixed_sizes_odd_segments(ctx, ref_lib, monkeypatch):
    monkeypatch.setenv("LZ4E_B200_CHUNK_MB", "1")
    rng = np.random.default_rng(21)
    sizes = [int(rng.choice([0, 5, 12, 13, 100, 4096, 8192, 65536])) if rng.random() < 0.5
             else int(rng.integers(0, 65537)) for _ in range(160)]
    datas = [common.synth(n, int(rng.integers(0, 4)), i) for i, n in enumerate(sizes)]
    specs = [sg.segs_random(n, rng) for n in sizes]
    dspecs = [sg.segs_random(sg.compress_bound(n), rng) for n in sizes]
    ospecs = [sg.segs_random(n, rng) for n in sizes]
    _batch_roundtrip(ctx, ref_lib, datas, specs, dspecs, ospecs, _arena(ctx), _arena(ctx),
                     _arena(ctx))


def test_batch_interleaved_single_arena(ctx, ref_lib, monkeypatch):
    """Source and destination pages interleaved inside ONE arena: the bulk ranges cross, so the
    batch must fall back to a single chunk and still be exact."""
    monkeypatch.setenv("LZ4E_B200_CHUNK_MB", "1")
    n = 600
    datas = [common.synth(4096, b & 3, 5 + b) for b in range(n)]
    arena = pkg.Arena(ctx, n * 8 + 8)
    specs, dspecs = [], []
    src = sg.layout_from_segments(arena, [dict(sg.segs_single(4096), gap=3)] * n, first_page=0)
    dst = sg.layout_from_segments(arena, [dict(sg.segs_single(sg.compress_bound(4096)), gap=2)] * n,
                                  first_page=1)
    for b, d in enumerate(datas):
        src.write_block(b, d)
    before = arena.mem.copy()
    creqs, keep = pkg.make_reqs(src, dst)
    assert ctx.compress_batch(creqs, flags=pkg.F_DST_PRESERVE) == 0, ctx.err()
    rets = creqs["ret"].copy()
    ref_dst, (ref_rets, _) = common.oracle_compress(
        ref_lib, src, [sg.segs_single(sg.compress_bound(4096))] * n)
    assert np.array_equal(rets, ref_rets)
    for b in range(n):
        assert np.array_equal(dst.read_block(b, rets[b]), ref_dst.read_block(b, ref_rets[b])), b
        assert np.array_equal(src.read_block(b), datas[b]), b  # sources untouched
    # DST_PRESERVE: bytes nobody wrote keep their value
    mask = np.ones(len(before), dtype=bool)
    for b in range(n):
        for off, take in dst._pieces(b):
            mask[off:off + min(take, int(rets[b]))] = False
            break
    assert np.array_equal(arena.mem[mask], before[mask])


def test_batch_pageable_memory_bounced(ctx, ref_lib):
    """Pages that are in no arena (plain host memory) go through device bounce buffers."""
    rng = np.random.default_rng(4)
    sizes = [4096, 13, 65536, 1000, 0, 20000]
    datas = [common.synth(n, i & 3, 300 + i) for i, n in enumerate(sizes)]
    specs = [sg.segs_random(n, rng) for n in sizes]
    dspecs = [sg.segs_random(sg.compress_bound(n), rng) for n in sizes]
    ospecs = [sg.segs_random(n, rng) for n in sizes]
    _batch_roundtrip(ctx, ref_lib, datas, specs, dspecs, ospecs, sg.HostArena, sg.HostArena,
                     sg.HostArena)


def test_reference_entry_points(ctx, ref_lib):
    """LZ4E_compress_default / LZ4E_decompress_safe exactly as the reference declares them
    (lz4e.h:47-51): one call, caller-owned wrkmem, iterators moved, bytes written or 0."""
    lib = pkg.lib()
    rng = np.random.default_rng(8)
    wrk = np.full(pkg.MEM_COMPRESS if hasattr(pkg, "MEM_COMPRESS") else 17440, 0x5A, np.uint8)
    for i, n in enumerate([4096, 65536, 12, 13, 0, 30000]):
        d = common.synth(n, i & 3, 900 + i)
        spec, dspec = sg.segs_random(n, rng), sg.segs_random(sg.compress_bound(n), rng)
        src = common.build_layout(sg.HostArena, [spec], [d])
        dst = common.build_layout(sg.HostArena, [dspec])
        dst.arena.mem[:] = 0xEE
        si, di = src.iters.copy(), dst.iters.copy()
        ret = lib.LZ4E_compress_default(_vp(src.bvecs), _vp(dst.bvecs), _vp(si), _vp(di), _vp(wrk))
        ref_dst, (ref_rets, _, so, do) = common.oracle_compress(ref_lib, src, [dspec],
                                                                want_iters=True)
        assert ret == ref_rets[0], (n, ret, ref_rets[0])
        assert np.array_equal(dst.arena.mem, ref_dst.arena.mem[:len(dst.arena.mem)]), n
        for f in ("bi_size", "bi_idx", "bi_bvec_done"):
            assert si[f][0] == so[f][0] and di[f][0] == do[f][0], (n, f)
        assert not wrk.any()  # cleared like lz4e_compress.c:548
        wrk[:] = 0x5A
        if ret > 0:
            comp = np.ascontiguousarray(dst.read_block(0, ret))
            out = np.full(n + 64, 0xEE, np.uint8)
            r2 = lib.LZ4E_decompress_safe(_vp(comp), _vp(out), int(ret), int(n))
            assert r2 == n
            assert np.array_equal(out[:n], d) and (out[n:] == 0xEE).all()
            if n > 40:  # a truncated stream fails with the reference's negative code
                ref_r, _ = common.oracle_decode_linear(ref_lib, [comp[:ret // 2]], [n])
                assert lib.LZ4E_decompress_safe(_vp(comp), _vp(out), int(ret // 2), int(n)) == ref_r[0]


def test_compacted_output(ctx, ref_lib):
    """lz4e_b200_compress_compact_dev: all blocks back to back in ONE destination SG stream,
    offsets = exclusive scan of the reference's return values."""
    rng = np.random.default_rng(6)
    n = 700
    sizes = [4096 if rng.random() < 0.7 else int(rng.integers(0, 9000)) for _ in range(n)]
    datas = [common.synth(s, int(rng.integers(0, 4)), i) for i, s in enumerate(sizes)]
    specs = [sg.segs_random(s, rng, max_segs=6) for s in sizes]
    src = common.build_layout(_arena(ctx), specs, datas)
    ref_dst, (ref_rets, _) = common.oracle_compress(
        ref_lib, src, [sg.segs_single(sg.compress_bound(s)) for s in sizes])
    total = int(ref_rets.sum())
    skip = 777
    dspec = sg.segs_random(total + 100, rng, max_segs=20)
    dspec["skip"], dspec["lead"] = 0, 0
    dspec["segs"][0] = (dspec["segs"][0][0], dspec["segs"][0][1] + skip)
    dst = common.build_layout(_arena(ctx), [dspec])
    dst.arena.mem[:] = 0xEE
    src.arena.upload()
    dst.arena.upload()
    s, d = common.DevTables(ctx, src), common.DevTables(ctx, dst)
    d_rets, d_offs = ctx.dev_alloc(4 * n), ctx.dev_alloc(8 * (n + 1))
    ctx.compress_compact_dev(s.bv, s.it, d.bv, len(dst.bvecs), skip, total + 100, n,
                             max(sizes), d_rets, d_offs)
    ctx.sync()
    rets = ctx.download(d_rets, np.int32, n)
    offs = ctx.download(d_offs, np.uint64, n + 1)
    dst.arena.download()
    assert np.array_equal(rets, ref_rets)
    assert np.array_equal(offs, np.concatenate([[0], np.cumsum(ref_rets)]).astype(np.uint64))
    dst.iters["bi_bvec_done"] = skip
    dst.iters["bi_size"] = total
    packed = dst.read_block(0, total)
    for b in range(n):
        o = int(offs[b])
        assert np.array_equal(packed[o:o + int(rets[b])], ref_dst.read_block(b, ref_rets[b])), b


@pytest.mark.parametrize("mode", ["1", "2"])
def test_batch_moves_exact_bytes(ctx, ref_lib, mode, monkeypatch):
    """The transfer kernel brings back rets[b] bytes per block and nothing else: without
    LZ4E_B200_F_DST_PRESERVE every host byte the codec did not produce keeps its value, and the
    PCIe byte counters show produced bytes, not destination slots."""
    monkeypatch.setenv("LZ4E_B200_XFER", mode)
    monkeypatch.setenv("LZ4E_B200_SPARSE_SRC", "kernel")  # default: the spanned range when >= 25 % full
    monkeypatch.setenv("LZ4E_B200_CHUNK_MB", "2")
    n = 2000
    datas = [common.synth(4096, b & 3, 11 + b) for b in range(n)]
    specs = [dict(sg.segs_single(4096), gap=0)] * n
    dspecs = [dict(sg.segs_single(sg.compress_bound(4096)), gap=0)] * n
    src = common.build_layout(_arena(ctx), specs, datas)
    dst = common.build_layout(_arena(ctx), dspecs)
    dst.arena.mem[:] = 0xEE
    creqs, keep = pkg.make_reqs(src, dst)
    ctx.sync()
    h0, d0 = ctx.traffic()
    assert ctx.compress_batch(creqs, flags=pkg.F_KEEP_ITERS) == 0, ctx.err()
    ctx.sync()
    h1, d1 = ctx.traffic()
    rets = creqs["ret"].copy()
    ref_dst, (ref_rets, _) = common.oracle_compress(ref_lib, src, dspecs)
    assert np.array_equal(rets, ref_rets)
    assert np.array_equal(dst.arena.mem, ref_dst.arena.mem[:len(dst.arena.mem)])  # 0xEE elsewhere
    tables = n * 64 + 8 * 256 * 16
    assert n * 4096 <= h1 - h0 <= n * 4096 + tables + 4096 * 16
    assert int(rets.sum()) <= d1 - d0 <= int(rets.sum()) + n * 12 + 8 * 256 * 16
    # decompress: only the compressed bytes go up
    out = common.build_layout(_arena(ctx), specs)
    csrc = sg.Layout(dst.arena, dst.bvecs, dst.first, dst.iters.copy(), dst.npages_used)
    csrc.iters["bi_size"] = rets.astype(np.uint32)
    dreqs, keep2 = pkg.make_reqs(csrc, out)
    assert ctx.decompress_batch(dreqs) == 0, ctx.err()
    ctx.sync()
    h2, d2 = ctx.traffic()
    assert int(rets.sum()) <= h2 - h1 <= int(rets.sum()) + tables
    assert (dreqs["ret"] == 4096).all()
    assert np.array_equal(out.arena.mem[:n * 4096], src.arena.mem[:n * 4096])


@pytest.mark.parametrize("sparse", ["kernel", "hull", "pack", "rows"])
def test_batch_sparse_source_modes(ctx, ref_lib, sparse, monkeypatch):
    """The three ways a sparse source side can go up give the same results."""
    monkeypatch.setenv("LZ4E_B200_SPARSE_SRC", sparse)
    monkeypatch.setenv("LZ4E_B200_CHUNK_MB", "1")
    rng = np.random.default_rng(33)
    sizes = [4096] * 700 + [int(rng.integers(0, 30000)) for _ in range(60)]
    datas = [common.synth(n, i & 3, 40 + i) for i, n in enumerate(sizes)]
    specs = [sg.segs_random(n, rng, max_segs=5) for n in sizes]
    dspecs = [dict(sg.segs_single(sg.compress_bound(n)), gap=0) for n in sizes]
    ospecs = [sg.segs_random(n, rng, max_segs=5) for n in sizes]
    _batch_roundtrip(ctx, ref_lib, datas, specs, dspecs, ospecs, _arena(ctx), _arena(ctx),
                     _arena(ctx))


def test_batch_sparse_sources_kernel_path(ctx, ref_lib, monkeypatch):
    """64 KiB blocks over 16 odd segments with guard pages in between: neither side is dense,
    so sources go up and results come down through the transfer kernel only."""
    monkeypatch.setenv("LZ4E_B200_CHUNK_MB", "1")
    monkeypatch.setenv("LZ4E_B200_SPARSE_SRC", "kernel")
    n = 40
    datas = [common.synth(65536, b & 3, 500 + b) for b in range(n)]
    specs = [sg.segs_odd(65536, 16, b) for b in range(n)]
    dspecs = [sg.segs_odd(sg.compress_bound(65536), 16, b + 7) for b in range(n)]
    ospecs = [sg.segs_odd(65536, 16, b + 3) for b in range(n)]
    ctx.sync()
    h0, d0 = ctx.traffic()
    rets = _batch_roundtrip(ctx, ref_lib, datas, specs, dspecs, ospecs, _arena(ctx), _arena(ctx),
                            _arena(ctx))
    ctx.sync()
    h1, d1 = ctx.traffic()
    moved = n * 65536 + int(rets.sum())
    assert moved <= h1 - h0 <= moved + (1 << 20)
    assert moved <= d1 - d0 <= moved + (1 << 20)


def test_contexts_in_threads_and_destroy(ref_lib):
    """Per-context workspaces: two contexts on the same device run batches from two threads at
    once; destroying one (with its arenas) leaves the other usable, and a context created
    afterwards starts from a clean workspace."""
    import threading
    n = 800
    datas = [common.synth(4096, b & 3, 900 + b) for b in range(n)]
    specs = [dict(sg.segs_single(4096), gap=0)] * n
    dspecs = [dict(sg.segs_single(sg.compress_bound(4096)), gap=0)] * n
    want = None
    errs = []

    def work(c, reps):
        try:
            for _ in range(reps):
                src = common.build_layout(_arena(c), specs, datas)
                dst = common.build_layout(_arena(c), dspecs)
                creqs, keep = pkg.make_reqs(src, dst)
                assert c.compress_batch(creqs, flags=pkg.F_KEEP_ITERS) == 0, c.err()
                got = [bytes(dst.read_block(b, creqs["ret"][b])) for b in range(0, n, 37)]
                assert got == want
        except Exception as e:  # noqa: BLE001
            errs.append(repr(e))

    c0 = pkg.Context(0)
    src = common.build_layout(_arena(c0), specs, datas)
    ref_dst, (ref_rets, _) = common.oracle_compress(ref_lib, src, dspecs)
    want = [bytes(ref_dst.read_block(b, ref_rets[b])) for b in range(0, n, 37)]
    c1 = pkg.Context(0)
    ts = [threading.Thread(target=work, args=(c, 3)) for c in (c0, c1)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    # two threads on ONE context: the context hands each batch its own workspace
    ts = [threading.Thread(target=work, args=(c1, 3)) for _ in range(3)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    c0.close()  # its arenas and workspace go with it
    work(c1, 1)
    c2 = pkg.Context(0)
    work(c2, 1)
    assert not errs, errs
    c1.close()
    c2.close()
