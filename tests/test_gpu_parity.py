"""GPU parity tests proper: the CUDA path, called through the C ABI, against the oracle on the
same seeded inputs.  Bit-exact: compressed bytes, return values, round trips."""
import numpy as np
import pytest

import common
from common import pkg, sg

pytestmark = pytest.mark.gpu


def _arena(ctx):
    return lambda npages: pkg.Arena(ctx, npages)


def _check_compress(ctx, ref_lib, datas, specs, dst_specs=None, flags=0, check_untouched=True):
    src = common.build_layout(_arena(ctx), specs, datas)
    if dst_specs is None:
        dst_specs = [sg.segs_single(sg.compress_bound(len(d))) for d in datas]
    dst = common.build_layout(_arena(ctx), dst_specs)
    dst.arena.mem[:] = 0xEE
    rets, ends = common.gpu_compress(ctx, src, dst, flags=flags, want_ends=True)
    ref_dst, (ref_rets, _, so, do) = common.oracle_compress(ref_lib, src, dst_specs, want_iters=True)
    assert np.array_equal(rets, ref_rets)
    for b in range(len(datas)):
        assert np.array_equal(dst.read_block(b, rets[b]), ref_dst.read_block(b, ref_rets[b])), b
        # iterator end state (lz4e_compress.c:304-305, :528-529), also after a 0 return
        # (:363, :430, :509 return from inside the loop)
        dp = int(dst_specs[b]["size"]) - int(do[b]["bi_size"])
        assert int(ends[b]["src_pos"]) == len(datas[b]) - int(so[b]["bi_size"]), b
        assert int(ends[b]["dst_pos"]) == dp, b
        if rets[b] == 0 and dp:
            # the prefix the reference leaves; the token byte of the sequence in progress is
            # garbage on both sides (include/lz4e.h)
            got, want = dst.read_block(b, dp), ref_dst.read_block(b, dp)
            assert int((got != want).sum()) <= 1, b
    if check_untouched:  # no byte outside [dst start, +ret) is touched
        assert np.array_equal(dst.arena.mem, ref_dst.arena.mem[:len(dst.arena.mem)])
    return src, dst, rets


def test_config2_4k_mixed(ctx, ref_lib):
    """BASELINE config 2 shape at reduced count: 4 KiB blocks, 1 bvec, class = index mod 4."""
    n = 512
    datas = [common.synth(4096, b & 3, b) for b in range(n)]
    src, dst, rets = _check_compress(ctx, ref_lib, datas, [sg.segs_single(4096)] * n)
    assert rets[1] == 4114  # incompressible expands
    # linear destinations written in place (no staging) give the same bytes
    _check_compress(ctx, ref_lib, datas, [sg.segs_single(4096)] * n, flags=pkg.DEV_DST_LINEAR)
    # config 4: decompress what was produced, from the (multi-page) destination bvecs
    out = common.build_layout(_arena(ctx), [sg.segs_single(4096)] * n)
    r2 = common.gpu_decompress(ctx, dst, rets, out)
    assert (r2 == 4096).all()
    for b in range(n):
        assert np.array_equal(out.read_block(b), datas[b])


def test_config3_64k_odd_segments(ctx, ref_lib):
    """BASELINE config 3 shape: 64 KiB blocks over 16 odd segments, mid-bvec start, SG dst."""
    n = 48
    datas = [common.synth(65536, b & 3, 1000 + b) for b in range(n)]
    specs = [sg.segs_odd(65536, 16, b) for b in range(n)]
    dspecs = [sg.segs_odd(sg.compress_bound(65536), 16, b + 7) for b in range(n)]
    src, dst, rets = _check_compress(ctx, ref_lib, datas, specs, dspecs,
                                     flags=pkg.DEV_NO_BYU16)
    out = common.build_layout(_arena(ctx), [sg.segs_odd(65536, 16, b + 3) for b in range(n)])
    r2 = common.gpu_decompress(ctx, dst, rets, out)
    assert (r2 == 65536).all()
    for b in range(n):
        assert np.array_equal(out.read_block(b), datas[b])
    # linear destination takes the in-place path
    out2 = common.build_layout(_arena(ctx), [sg.segs_single(65536)] * n)
    r3 = common.gpu_decompress(ctx, dst, rets, out2, flags=pkg.DEV_DST_LINEAR)
    assert (r3 == 65536).all()
    for b in range(n):
        assert np.array_equal(out2.read_block(b), datas[b])


def test_64k_page_bvecs_byu16(ctx, ref_lib):
    """16 x 4 KiB page bvecs select byU16 with a 64 KiB window."""
    datas = [common.lcg_text(65536), common.synth(65536, 3, 5), np.zeros(65536, np.uint8)]
    _, _, rets = _check_compress(ctx, ref_lib, datas, [sg.segs_pages(65536)] * 3)
    assert rets[0] == 26412 and rets[2] == 267  # SURVEY.md 8(c) sanity values


def test_survey_sanity_values(ctx, ref_lib):
    t = common.lcg_text(65536)
    datas = [t[:4096], t, np.zeros(4096, np.uint8), t[:12]]
    specs = [sg.segs_single(4096), sg.segs_single(65536), sg.segs_single(4096), sg.segs_single(12)]
    _, _, rets = _check_compress(ctx, ref_lib, datas, specs)
    assert list(rets) == [1688, 23373, 26, 13]


def test_fuzz_segmentations_and_sizes(ctx, ref_lib):
    rng = np.random.default_rng(11)
    datas, specs, dspecs = [], [], []
    edge = [0, 1, 5, 11, 12, 13, 14, 20, 100, 1000, 4095, 4096, 4097, 8192, 20000, 65535, 65536]
    for i in range(300):
        n = int(rng.choice(edge)) if rng.random() < 0.5 else int(rng.integers(0, 65537))
        datas.append(common.synth(n, int(rng.integers(0, 4)), i))
        specs.append(sg.segs_random(n, rng))
        dspecs.append(sg.segs_random(sg.compress_bound(n), rng))
    src, dst, rets = _check_compress(ctx, ref_lib, datas, specs, dspecs)
    out = common.build_layout(_arena(ctx), [sg.segs_random(max(len(d), 0), rng) for d in datas])
    r2 = common.gpu_decompress(ctx, dst, rets, out)
    for b, d in enumerate(datas):
        assert r2[b] == len(d), (b, r2[b], len(d))
        assert np.array_equal(out.read_block(b), d)


def test_limited_output(ctx, ref_lib):
    """Capacity below LZ4E_COMPRESSBOUND: exact fit succeeds, one byte less returns 0."""
    rng = np.random.default_rng(5)
    base = [common.synth(int(rng.integers(13, 9000)), int(rng.integers(0, 4)), i) for i in range(40)]
    src = common.build_layout(_arena(ctx), [sg.segs_single(len(d)) for d in base], base)
    full = common.build_layout(_arena(ctx), [sg.segs_single(sg.compress_bound(len(d))) for d in base])
    sizes = common.gpu_compress(ctx, src, full)
    datas, specs, dspecs = [], [], []
    for d, c in zip(base, sizes):
        for cap in (int(c), int(c) - 1, max(int(c) - 9, 1), int(c) // 2, 1):  # cap 0: the reference BUG_ONs
            datas.append(d)
            specs.append(sg.segs_single(len(d)))
            dspecs.append(dict(sg.segs_single(max(cap, 1)), size=cap))
    # a failed call leaves the reference's iterator positions and its prefix up to *dstIter
    # (compared in _check_compress); bytes behind *dstIter are unspecified on both sides
    for flags in (0, pkg.DEV_DST_LINEAR):
        _, _, rets = _check_compress(ctx, ref_lib, datas, specs, dspecs, flags=flags,
                                     check_untouched=False)
        assert (rets[0::5] == sizes).all() and (rets[1::5] == 0).all()


def test_runs_longer_than_the_window(ctx, ref_lib):
    """Blocks above 64 KiB whose length runs alone exceed the 1 KiB staging window (one length
    byte per 255 bytes, lz4e_compress.c:365-377, :432-446, :511-522): 1 MiB of noise (last
    literals, 4112 length bytes), 1 MiB of zeros (one match), 700 KB of noise in front of text
    (literal run inside the loop), next to ordinary blocks in the same launch whose shared
    memory such a run must not touch."""
    noise = np.concatenate([common.synth(65536, 1, 900 + b) for b in range(16)])
    mixed = np.concatenate([noise[:700000], np.tile(common.lcg_text(3000), 120)])[:1048576]
    tail = np.concatenate([noise[:300000], np.zeros(748576, np.uint8)])
    datas = [noise, np.zeros(1048576, np.uint8), mixed, tail] + \
            [common.synth(65536, b & 3, 50 + b) for b in range(12)]
    specs = [sg.segs_pages(1048576), sg.segs_single(1048576), sg.segs_odd(1048576, 40, 3),
             sg.segs_single(1048576)] + [sg.segs_odd(65536, 16, b) for b in range(12)]
    dspecs = [sg.segs_odd(sg.compress_bound(len(d)), 9, 3 + i) for i, d in enumerate(datas)]
    src, dst, rets = _check_compress(ctx, ref_lib, datas, specs, dspecs)
    out = common.build_layout(_arena(ctx), [sg.segs_odd(len(d), 7, i) for i, d in enumerate(datas)])
    r2 = common.gpu_decompress(ctx, dst, rets, out)
    for b, d in enumerate(datas):
        assert r2[b] == len(d)
        assert np.array_equal(out.read_block(b), d)


def test_too_many_bvecs(ctx, ref_lib):
    d = common.synth(300, 0, 1)
    spec257 = {"size": 257, "segs": [(i % 4000, 1) for i in range(257)], "skip": 0, "gap": 0}
    spec256 = {"size": 256, "segs": [(i % 4000, 1) for i in range(256)], "skip": 0, "gap": 0}
    _, _, rets = _check_compress(ctx, ref_lib, [d[:257], d[:256]], [spec257, spec256])
    assert rets[0] == 0 and rets[1] > 0


def test_decompress_malformed(ctx, ref_lib):
    """Mutated / truncated / random streams: return values equal the reference decoder's,
    including the -(offset)-1 error codes (lz4e_decompress.c:458-459)."""
    rng = np.random.default_rng(3)
    base = [common.synth(int(rng.integers(20, 3000)), int(rng.integers(0, 4)), i) for i in range(200)]
    src = common.build_layout(_arena(ctx), [sg.segs_single(len(d)) for d in base], base)
    full = common.build_layout(_arena(ctx), [sg.segs_single(sg.compress_bound(len(d))) for d in base])
    sizes = common.gpu_compress(ctx, src, full)
    comps, caps = [], []
    for b, d in enumerate(base):
        c = bytearray(full.read_block(b, sizes[b]).tobytes())
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
    assert (ref_rets < 0).sum() > 20 and (ref_rets >= 0).sum() > 20
    for seg_mode in (0, 1):
        mk = (lambda n: sg.segs_random(n, rng)) if seg_mode else \
            (lambda n: dict(sg.segs_single(max(n, 1)), size=n))
        csrc = common.build_layout(_arena(ctx), [mk(len(c)) for c in comps], comps)
        out = common.build_layout(_arena(ctx), [mk(c) for c in caps])
        rets = common.gpu_decompress(ctx, csrc, [len(c) for c in comps], out)
        assert np.array_equal(rets, ref_rets)
        for b in range(len(comps)):
            if rets[b] > 0:
                assert np.array_equal(out.read_block(b, rets[b]), ref_outs[b])


def test_decompress_stress_overlap_and_literals(ctx, ref_lib):
    """Config 4 stress sets: RLE-like data (offsets 1..31, long overlapping matches) and long
    literal runs, at 4 KiB and 64 KiB."""
    rng = np.random.default_rng(9)
    datas = []
    for size in (4096, 65536):
        for period in (1, 2, 3, 4, 5, 7, 8, 13, 16, 17, 31, 32, 33, 64):
            pat = rng.integers(0, 256, period, dtype=np.uint8)
            d = np.tile(pat, size // period + 1)[:size].copy()
            d[rng.integers(0, size, 6)] ^= 0x55  # break the run a few times
            datas.append(d)
        datas.append(rng.integers(0, 256, size, dtype=np.uint8))
    specs = [sg.segs_single(len(d)) for d in datas]
    src, dst, rets = _check_compress(ctx, ref_lib, datas, specs)
    for big in (False, True):
        sel = [i for i, d in enumerate(datas) if (len(d) > 4096) == big]
        sub = common.build_layout(_arena(ctx), [dict(sg.segs_single(sg.compress_bound(len(datas[i]))),
                                                     size=int(rets[i])) for i in sel],
                                  [dst.read_block(i, rets[i]) for i in sel])
        out = common.build_layout(_arena(ctx), [sg.segs_single(len(datas[i])) for i in sel])
        r2 = common.gpu_decompress(ctx, sub, [rets[i] for i in sel], out)
        for k, i in enumerate(sel):
            assert r2[k] == len(datas[i])
            assert np.array_equal(out.read_block(k), datas[i])


def test_blocks_above_64k(ctx, ref_lib):
    """bios of 68 KiB ... 1 MiB (multi-page bvecs, more than 16 bvecs): 32-bit
    table positions, byte-identical to the reference, round trip through the SG decoder."""
    sizes = [70000, 98304, 131072, 150000, 262144, 1048576]
    datas = [common.lcg_text(sizes[0]), common.synth(sizes[1], 3, 5), common.synth(sizes[2], 0, 9),
             np.zeros(sizes[3], np.uint8), common.synth(sizes[4], 2, 11),
             np.concatenate([common.synth(65536, b & 3, 70 + b) for b in range(16)])]
    specs = [sg.segs_single(sizes[0]), sg.segs_odd(sizes[1], 24, 1), sg.segs_pages(sizes[2]),
             sg.segs_single(sizes[3]), sg.segs_odd(sizes[4], 64, 2), sg.segs_pages(sizes[5])]
    dspecs = [sg.segs_odd(sg.compress_bound(n), 9, 3 + i) for i, n in enumerate(sizes)]
    src, dst, rets = _check_compress(ctx, ref_lib, datas, specs, dspecs)
    out = common.build_layout(_arena(ctx), [sg.segs_odd(n, 7, i) for i, n in enumerate(sizes)])
    r2 = common.gpu_decompress(ctx, dst, rets, out)
    assert np.array_equal(r2, np.array(sizes, np.int32))
    for b in range(len(sizes)):
        assert np.array_equal(out.read_block(b), datas[b])
    # a mixed batch: small blocks next to large ones in one launch
    datas2 = [common.synth(4096, 3, 1), datas[1], common.synth(65536, 0, 2), datas[0]]
    specs2 = [sg.segs_single(4096), specs[1], sg.segs_odd(65536, 16, 4), specs[0]]
    _check_compress(ctx, ref_lib, datas2, specs2)
