"""CPU tests of the CUDA kernel BODIES (lz4e_compress_body.cuh / lz4e_decompress_body.cuh)
compiled for the host against the 32-lane warp emulator in tests/simt — the very source
nvcc builds — compared with the oracle.  Small sizes: the emulator is slow."""
import numpy as np
import pytest

import common
import emu
from common import sg


@pytest.fixture(scope="module")
def simt():
    return emu.load()


def _host_layout(specs, datas=None):
    return common.build_layout(sg.HostArena, specs, datas)


def _cases():
    rng = np.random.default_rng(7)
    datas, specs = [], []
    for b in range(8):  # config-2 shape
        datas.append(common.synth(4096, b & 3, b))
        specs.append(sg.segs_single(4096))
    for n in (0, 1, 12, 13, 14, 64, 300, 1500):  # short inputs, lz4e_compress.c:268-271
        datas.append(common.synth(n, 3, 50 + n) if n else np.zeros(0, np.uint8))
        specs.append(sg.segs_single(n))
    datas.append(common.lcg_text(9000))  # > 4096 in one bvec: byU32
    specs.append(sg.segs_single(9000))
    datas.append(np.zeros(5000, np.uint8))  # long RLE match
    specs.append(sg.segs_single(5000))
    for b in range(4):  # odd segments, mid-bvec start
        n = int(rng.integers(2000, 12000))
        datas.append(common.synth(n, (b + 1) & 3, 900 + b))
        specs.append(sg.segs_odd(n, int(rng.integers(2, 9)), b))
    return datas, specs


# streamed output / whole block staged / written in place (linear destination)
@pytest.mark.parametrize("out_cap,gin", [(1024, False), (16384, False), (64, False),
                                         (1024, True), (64, True)])
def test_compress_body_matches_oracle(simt, ref_lib, out_cap, gin):
    datas, specs = _cases()
    src = _host_layout(specs, datas)
    dspecs = [sg.segs_single(sg.compress_bound(len(d))) for d in datas]
    dst = _host_layout(dspecs)
    dst.arena.mem[:] = 0xEE
    rets, ends, faults = emu.compress(simt, src, dst, in_cap=12288, out_cap=out_cap, gin=gin)
    ref_dst, (ref_rets, _, so, do) = common.oracle_compress(ref_lib, src, dspecs, want_iters=True)
    assert faults == 0
    assert np.array_equal(rets, ref_rets)
    for b in range(len(datas)):
        assert np.array_equal(dst.read_block(b, rets[b]), ref_dst.read_block(b, ref_rets[b])), b
        if rets[b] > 0:
            assert int(ends[b]["src_pos"]) == len(datas[b]) - int(so[b]["bi_size"]), b
            assert int(ends[b]["dst_pos"]) == int(dspecs[b]["size"]) - int(do[b]["bi_size"]), b


@pytest.mark.parametrize("out_cap", [1024, 64])
@pytest.mark.parametrize("sg_dst", [False, True])
def test_compress_body_limited_output(simt, ref_lib, out_cap, sg_dst):
    """limitedOutput (lz4e_compress.c:358-363, :425-430, :509): return values, and after a 0
    return the iterator positions and the destination prefix up to *dstIter the reference leaves
    (the token byte of the sequence in progress is garbage on both sides: <= 1 byte differs)."""
    if sg_dst and out_cap == 64:
        pytest.skip("out_cap 64 selects the in-place kernels: linear destinations only")
    datas = [common.synth(4096, c, 77 + c) for c in range(4)] * 6
    specs = [sg.segs_single(4096)] * len(datas)
    src = _host_layout(specs, datas)
    full = _host_layout([sg.segs_single(sg.compress_bound(4096))] * len(datas))
    sizes, _ = common.oracle.compress_batch(ref_lib, src.bvecs, src.first, src.iters,
                                            full.bvecs, full.first, full.iters)
    deltas = [0, -1, -7, 5] * 3 + [-2, -3, -5, -6, -9, -12, -40, -100, -300, -1000, -1500, -16]
    caps = [max(int(s) + d if d > -1000 else int(s) // (2 if d == -1000 else 3), 1)
            for s, d in zip(sizes, deltas)]
    dspecs = [sg.segs_odd(c, 5, i) if sg_dst and c > 64 else sg.segs_single(c)
              for i, c in enumerate(caps)]
    dst = _host_layout(dspecs)
    dst.arena.mem[:] = 0xEE
    rets, ends, faults = emu.compress(simt, src, dst, in_cap=4096, out_cap=out_cap)
    ref_dst, (ref_rets, _, so, do) = common.oracle_compress(ref_lib, src, dspecs, want_iters=True)
    assert faults == 0
    assert np.array_equal(rets, ref_rets)
    assert (rets == 0).sum() >= 12
    for b in range(len(datas)):
        assert np.array_equal(dst.read_block(b, rets[b]), ref_dst.read_block(b, ref_rets[b])), b
        dp = caps[b] - int(do[b]["bi_size"])
        assert int(ends[b]["src_pos"]) == 4096 - int(so[b]["bi_size"]), b
        assert int(ends[b]["dst_pos"]) == dp, b
        if rets[b] == 0 and dp:
            got, want = dst.read_block(b, dp), ref_dst.read_block(b, dp)
            assert int((got != want).sum()) <= 1, b


def test_compress_body_long_length_runs(simt, ref_lib):
    """Length runs of hundreds of bytes (one length byte per 255 bytes of a literal run or a
    match): last literals of an incompressible 64 KiB block, a long literal run inside the loop,
    a 64 KiB match — into linear and scattered destinations through the 1 KiB window.  A window
    that could not hold such a run by itself is refused (fault flag), never overrun: the launch
    sizes it from the largest block (lz4e_b200_compress_dev)."""
    rnd = common.synth(65536, 1, 5)
    inloop = np.concatenate([rnd[:40000], np.tile(common.lcg_text(500), 60)])[:65536]
    zeros = np.zeros(65536, np.uint8)
    both = np.concatenate([rnd[:33000], np.zeros(32536, np.uint8)])
    datas = [rnd, inloop, zeros, both] * 2
    specs = [sg.segs_single(65536)] * 4 + [sg.segs_odd(65536, 16, i) for i in range(4)]
    dspecs = [sg.segs_single(sg.compress_bound(65536))] * 4 + \
             [sg.segs_odd(sg.compress_bound(65536), 7, i) for i in range(4)]
    src = _host_layout(specs, datas)
    dst = _host_layout(dspecs)
    dst.arena.mem[:] = 0xEE
    rets, ends, faults = emu.compress(simt, src, dst, in_cap=65536, out_cap=1024)
    ref_dst, (ref_rets, _) = common.oracle_compress(ref_lib, src, dspecs)
    assert faults == 0
    assert np.array_equal(rets, ref_rets)
    for b in range(len(datas)):
        assert np.array_equal(dst.read_block(b, rets[b]), ref_dst.read_block(b, ref_rets[b])), b
    assert np.array_equal(dst.arena.mem, ref_dst.arena.mem[:len(dst.arena.mem)])
    rets, _, faults = emu.compress(simt, src, dst, in_cap=65536, out_cap=256)
    assert faults == 1 and (rets == 0).all()


@pytest.mark.parametrize("out_cap", [1024, 4352, 64])
def test_compress_body_packed_table(simt, ref_lib, out_cap):
    """Blocks of at most 4 KiB with the packed 12-bit table (12 KiB for the 8192 byU16 entries):
    same bytes as the reference on every class, on sizes around the limits, on page-split and
    odd segmentations (byU16) and on a multi-page bvec (byU32 inside the same launch)."""
    rng = np.random.default_rng(12)
    datas, specs = [], []
    for c in range(4):
        for n in (4096, 4095, 2500, 300, 13, 12, 0):
            datas.append(common.synth(n, c, 31 * c + n))
            specs.append(sg.segs_single(n))
    for b in range(8):  # several small segments: still byU16
        n = int(rng.integers(500, 4097))
        datas.append(common.synth(n, b & 3, 500 + b))
        specs.append(sg.segs_random(n, rng, max_segs=8, allow_big=False))
    datas.append(common.lcg_text(4096))  # raw bv_len above 4096: byU32, plain 8 KiB table
    specs.append({"size": 4096, "segs": [(100, 6000)], "skip": 700})
    datas.append(np.zeros(4096, np.uint8))
    specs.append(sg.segs_single(4096))
    src = _host_layout(specs, datas)
    dspecs = [sg.segs_single(sg.compress_bound(len(d))) for d in datas]
    dst = _host_layout(dspecs)
    dst.arena.mem[:] = 0xEE
    rets, ends, faults = emu.compress(simt, src, dst, in_cap=4096, out_cap=out_cap, table_bytes=12288)
    ref_dst, (ref_rets, _) = common.oracle_compress(ref_lib, src, dspecs)
    assert faults == 0
    assert np.array_equal(rets, ref_rets)
    for b in range(len(datas)):
        assert np.array_equal(dst.read_block(b, rets[b]), ref_dst.read_block(b, ref_rets[b])), b


@pytest.mark.parametrize("group", [1, 2, 4, 8, 16, 32])
def test_decompress_body_roundtrip(simt, ref_lib, group):
    datas, specs = _cases()
    keep = [i for i, d in enumerate(datas) if len(d) > 0]
    datas, specs = [datas[i] for i in keep], [specs[i] for i in keep]
    src = _host_layout(specs, datas)
    dspecs = [sg.segs_single(sg.compress_bound(len(d))) for d in datas]
    cdst, (sizes, _) = common.oracle_compress(ref_lib, src, dspecs)
    out = _host_layout([sg.segs_single(len(d)) for d in datas])
    csrc = sg.Layout(cdst.arena, cdst.bvecs, cdst.first, cdst.iters.copy(), cdst.npages_used)
    csrc.iters["bi_size"] = sizes.astype(np.uint32)
    rets, faults = emu.decompress(simt, csrc, out, group=group, stage_cap=0, scratch_stride=16384)
    assert faults == 0
    assert np.array_equal(rets, [len(d) for d in datas])
    for b, d in enumerate(datas):
        assert np.array_equal(out.read_block(b), d), b


def test_lane_decoder_scattered_source_and_destination(simt, ref_lib):
    """Lane-per-block decoder (lz4e_decompress_lane.cuh): compressed bytes and output both over
    odd, non-contiguous segments (gather into a slot, decode, scatter), 40 blocks so lanes
    take a second block."""
    rng = np.random.default_rng(11)
    datas = [common.synth(int(rng.integers(1, 9000)), b & 3, 300 + b) for b in range(40)]
    src = _host_layout([sg.segs_single(len(d)) for d in datas], datas)
    cdst, (sizes, _) = common.oracle_compress(ref_lib, src, [sg.segs_single(sg.compress_bound(len(d)))
                                                             for d in datas])
    comps = [cdst.read_block(b, sizes[b]) for b in range(len(datas))]
    csrc = _host_layout([sg.segs_random(len(c), rng) for c in comps], comps)
    out = _host_layout([sg.segs_random(len(d), rng) for d in datas])
    rets, faults = emu.decompress(simt, csrc, out, group=1, stage_cap=0, scratch_stride=16384)
    assert faults == 0
    assert np.array_equal(rets, [len(d) for d in datas])
    for b, d in enumerate(datas):
        assert np.array_equal(out.read_block(b), d), b


def test_lane_decoder_malformed_streams(simt, ref_lib):
    """Mutated, truncated and random streams: return values (incl. the -(offset)-1 codes,
    lz4e_decompress.c:458-459) and the bytes of accepted streams equal the reference decoder's."""
    rng = np.random.default_rng(5)
    base = [common.synth(int(rng.integers(20, 3000)), int(rng.integers(0, 4)), i) for i in range(96)]
    src = _host_layout([sg.segs_single(len(d)) for d in base], base)
    cdst, (sizes, _) = common.oracle_compress(ref_lib, src, [sg.segs_single(sg.compress_bound(len(d)))
                                                             for d in base])
    comps, caps = [], []
    for b, d in enumerate(base):
        c = bytearray(cdst.read_block(b, sizes[b]).tobytes())
        r = rng.random()
        if r < 0.3 and len(c) > 2:
            c = c[:int(rng.integers(1, len(c)))]
        elif r < 0.6:
            for _ in range(int(rng.integers(1, 4))):
                c[int(rng.integers(0, len(c)))] = int(rng.integers(0, 256))
        elif r < 0.75:
            c += bytes(rng.integers(0, 256, int(rng.integers(1, 20)), dtype=np.uint8))
        elif r < 0.85:
            c = bytearray(rng.integers(0, 256, int(rng.integers(1, 200)), dtype=np.uint8).tobytes())
        comps.append(np.frombuffer(bytes(c), dtype=np.uint8))
        caps.append(max(0, len(d) + int(rng.integers(-20, 20))) if rng.random() < 0.4 else len(d))
    ref_rets, ref_outs = common.oracle_decode_linear(ref_lib, comps, caps)
    assert (ref_rets < 0).sum() > 10 and (ref_rets >= 0).sum() > 10
    for seg_mode in (0, 1):
        mk = (lambda n: sg.segs_random(n, rng)) if seg_mode else \
            (lambda n: dict(sg.segs_single(max(n, 1)), size=n))
        csrc = _host_layout([mk(len(c)) for c in comps], comps)
        out = _host_layout([mk(c) for c in caps])
        rets, faults = emu.decompress(simt, csrc, out, group=1, stage_cap=0, scratch_stride=16384)
        assert faults == 0
        assert np.array_equal(rets, ref_rets)
        for b in range(len(comps)):
            if rets[b] > 0:
                assert np.array_equal(out.read_block(b, rets[b]), ref_outs[b]), b


def test_lane_decoder_medium_runs(simt, ref_lib):
    """Runs of 64..511 bytes on every lane at once (unit lists longer than the shared-memory
    table: several passes), periods below and above the unit size, literal runs in between."""
    rng = np.random.default_rng(21)
    datas = []
    for b in range(64):
        parts = []
        while sum(len(p) for p in parts) < 3000:
            period = int(rng.choice([1, 2, 3, 5, 7, 8, 9, 12, 16, 31, 40, 100]))
            pat = rng.integers(0, 256, period, dtype=np.uint8)
            n = int(rng.integers(64, 520))
            parts.append(np.tile(pat, n // period + 1)[:n])
            parts.append(rng.integers(0, 256, int(rng.integers(1, 90)), dtype=np.uint8))
        datas.append(np.concatenate(parts))
    src = _host_layout([sg.segs_single(len(d)) for d in datas], datas)
    cdst, (sizes, _) = common.oracle_compress(ref_lib, src, [sg.segs_single(sg.compress_bound(len(d)))
                                                             for d in datas])
    csrc = sg.Layout(cdst.arena, cdst.bvecs, cdst.first, cdst.iters.copy(), cdst.npages_used)
    csrc.iters["bi_size"] = sizes.astype(np.uint32)
    out = _host_layout([sg.segs_single(len(d)) for d in datas])
    rets, faults = emu.decompress(simt, csrc, out, group=1, stage_cap=0, scratch_stride=16384)
    assert faults == 0
    assert np.array_equal(rets, [len(d) for d in datas])
    for b, d in enumerate(datas):
        assert np.array_equal(out.read_block(b), d), b


@pytest.mark.parametrize("gin", [True, False])
def test_compress_body_blocks_above_64k(simt, ref_lib, gin):
    """above 64 KiB the table holds 32-bit positions (input in a global slot, or
    in shared memory for batches small enough to be resident at once);
    byU32 by construction (a bvec > 4096 bytes or more than 16 bvecs)."""
    datas = [common.lcg_text(70000), common.synth(98304, 3, 5), np.zeros(150000, np.uint8),
             common.synth(131072, 0, 9)]
    specs = [sg.segs_single(70000), sg.segs_odd(98304, 24, 1), sg.segs_single(150000),
             sg.segs_odd(131072, 32, 2)]
    src = _host_layout(specs, datas)
    dspecs = [sg.segs_single(sg.compress_bound(len(d))) for d in datas]
    dst = _host_layout(dspecs)
    dst.arena.mem[:] = 0xEE
    rets, ends, faults = emu.compress(simt, src, dst, in_cap=150016, out_cap=2048, gin=gin)
    ref_dst, (ref_rets, _, so, do) = common.oracle_compress(ref_lib, src, dspecs, want_iters=True)
    assert faults == 0
    assert np.array_equal(rets, ref_rets)
    for b in range(len(datas)):
        assert np.array_equal(dst.read_block(b, rets[b]), ref_dst.read_block(b, ref_rets[b])), b
        assert int(ends[b]["src_pos"]) == len(datas[b]) - int(so[b]["bi_size"]), b


# ---------------------------------------------------------------------------------------------
# lz4e_decompress_group.cuh: in-place scattered output through the segment map, parse window,
# pipelined short matches.  (group, variant, stage): variant 1 = synchronous, 2 = pipelined.
# (32, 3, False) = the wave decoder (lz4e_decompress_wave.cuh): one warp per block.
GROUP_CFGS = [(32, 3, False), (8, 2, False), (4, 2, False), (16, 2, False), (8, 1, False),
              (4, 1, False), (8, 2, True), (4, 1, True), (16, 2, True)]


def _compress_all(ref_lib, datas):
    src = _host_layout([sg.segs_single(len(d)) for d in datas], datas)
    cdst, (sizes, _) = common.oracle_compress(ref_lib, src, [sg.segs_single(sg.compress_bound(len(d)))
                                                             for d in datas])
    return [cdst.read_block(b, sizes[b]) for b in range(len(datas))]


@pytest.mark.parametrize("group,variant,stage", GROUP_CFGS)
def test_group_decoder_roundtrip(simt, ref_lib, group, variant, stage):
    """Valid streams of every class over linear, page-wise, odd and random segmentations on
    both sides (more than 32 destination segments included: the scratch-slot fallback)."""
    rng = np.random.default_rng(31 + group + variant)
    datas = []
    for b in range(44):
        n = int(rng.integers(1, 4096 if stage else 20000))
        datas.append(common.synth(n, b & 3, 700 + b))
    datas.append(np.zeros(4000 if stage else 30000, np.uint8))
    datas.append(common.lcg_text(4096 if stage else 24000))
    comps = _compress_all(ref_lib, datas)

    def dst_spec(i, n):
        k = i % 5
        if k == 0:
            return sg.segs_single(n)
        if k == 1:
            return sg.segs_pages(n)
        if k == 2:
            return sg.segs_odd(n, 16, i)
        if k == 3:
            return sg.segs_random(n, rng, max_segs=24)
        return sg.segs_random(n, rng, max_segs=80)   # > LZ4E_G_MAXSEG now and then
    csrc = _host_layout([sg.segs_random(len(c), rng) if i % 3 else sg.segs_single(len(c))
                         for i, c in enumerate(comps)], comps)
    out = _host_layout([dst_spec(i, len(d)) for i, d in enumerate(datas)])
    out.arena.mem[:] = 0xEE
    rets, faults = emu.decompress(simt, csrc, out, group=group, stage_cap=4112 if stage else 0,
                                  scratch_stride=30208, variant=variant)
    assert faults == 0
    assert np.array_equal(rets, [len(d) for d in datas])
    for b, d in enumerate(datas):
        assert np.array_equal(out.read_block(b), d), b


@pytest.mark.parametrize("group,variant,stage", GROUP_CFGS[:7])
def test_group_decoder_malformed_streams(simt, ref_lib, group, variant, stage):
    """Mutated, truncated and random streams: return values (incl. the -(offset)-1 codes,
    lz4e_decompress.c:458-459) and the bytes of accepted streams equal the reference decoder's,
    with linear and with scattered source/destination."""
    rng = np.random.default_rng(5 + group)
    base = [common.synth(int(rng.integers(20, 3000)), int(rng.integers(0, 4)), i) for i in range(96)]
    good = _compress_all(ref_lib, base)
    comps, caps = [], []
    for b, d in enumerate(base):
        c = bytearray(good[b].tobytes())
        r = rng.random()
        if r < 0.3 and len(c) > 2:
            c = c[:int(rng.integers(1, len(c)))]
        elif r < 0.6:
            for _ in range(int(rng.integers(1, 4))):
                c[int(rng.integers(0, len(c)))] = int(rng.integers(0, 256))
        elif r < 0.75:
            c += bytes(rng.integers(0, 256, int(rng.integers(1, 20)), dtype=np.uint8))
        elif r < 0.85:
            c = bytearray(rng.integers(0, 256, int(rng.integers(1, 200)), dtype=np.uint8).tobytes())
        comps.append(np.frombuffer(bytes(c), dtype=np.uint8))
        caps.append(max(0, len(d) + int(rng.integers(-20, 20))) if rng.random() < 0.4 else len(d))
    ref_rets, ref_outs = common.oracle_decode_linear(ref_lib, comps, caps)
    assert (ref_rets < 0).sum() > 10 and (ref_rets >= 0).sum() > 10
    for seg_mode in (0, 1):
        mk = (lambda n: sg.segs_random(n, rng)) if seg_mode else \
            (lambda n: dict(sg.segs_single(max(n, 1)), size=n))
        csrc = _host_layout([mk(len(c)) for c in comps], comps)
        out = _host_layout([mk(c) for c in caps])
        rets, faults = emu.decompress(simt, csrc, out, group=group, stage_cap=3200 if stage else 0,
                                      scratch_stride=16384, variant=variant)
        assert faults == 0
        assert np.array_equal(rets, ref_rets)
        for b in range(len(comps)):
            if rets[b] > 0:
                assert np.array_equal(out.read_block(b, rets[b]), ref_outs[b]), b


@pytest.mark.parametrize("group,variant", [(32, 3), (8, 2), (4, 2), (16, 1)])
def test_group_decoder_overlaps_across_segments(simt, ref_lib, group, variant):
    """Runs and patterns of every period crossing destination segment boundaries (short and
    long matches whose source lies in the segment before or further back), plus matches at
    offsets of several segments."""
    rng = np.random.default_rng(77 + group)
    datas = []
    for b in range(24):
        parts = []
        while sum(len(p) for p in parts) < 6000:
            period = int(rng.choice([1, 2, 3, 4, 5, 7, 8, 9, 12, 15, 16, 17, 31, 40, 100]))
            pat = rng.integers(0, 256, period, dtype=np.uint8)
            n = int(rng.integers(5, 700))
            parts.append(np.tile(pat, n // period + 1)[:n])
            parts.append(rng.integers(0, 256, int(rng.integers(0, 40)), dtype=np.uint8))
            if len(parts) > 6 and rng.random() < 0.5:   # far back-reference
                allb = np.concatenate(parts)
                o = int(rng.integers(0, len(allb) - 4))
                parts.append(allb[o:o + int(rng.integers(4, 300))].copy())
        datas.append(np.concatenate(parts))
    comps = _compress_all(ref_lib, datas)
    csrc = _host_layout([sg.segs_single(len(c)) for c in comps], comps)

    def tiny_segs(n, b):
        # many small segments (17..300 bytes) so that nearly every match meets a boundary
        segs, left = [], n
        while left > 0:
            ln = min(left, int(rng.integers(17, 300)))
            segs.append((int(rng.integers(0, 3000)), ln))
            left -= ln
        return {"size": n, "segs": segs[:LZ4E_MAXSEG_TEST] if len(segs) <= LZ4E_MAXSEG_TEST else
                _merge_tail(segs, LZ4E_MAXSEG_TEST), "skip": 0}
    out = _host_layout([tiny_segs(len(d), b) for b, d in enumerate(datas)])
    rets, faults = emu.decompress(simt, csrc, out, group=group, stage_cap=0, scratch_stride=0,
                                  variant=variant)
    assert faults == 0
    assert np.array_equal(rets, [len(d) for d in datas])
    for b, d in enumerate(datas):
        assert np.array_equal(out.read_block(b), d), b


LZ4E_MAXSEG_TEST = 32


def _merge_tail(segs, k):
    """Keep the first k-1 segments, fold the rest into one (stays within the decoder's map)."""
    tail = sum(ln for _, ln in segs[k - 1:])
    return segs[:k - 1] + [(0, tail)]
