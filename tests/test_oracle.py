"""CPU tests of the oracle itself: the linear-position port (oracle/lz4e_port.c) and, when it
is available, the reference's own object code (oracle/_ref) against the committed golden
vectors, the reference's fixtures and each other."""
import os

import numpy as np
import pytest

import common
import golden_util
from common import sg
from oracle import oracle

G = golden_util.load()
LIBS = ["port"] + (["reference"] if oracle.have_ref() or os.path.isdir("/root/reference") else [])


@pytest.fixture(scope="module", params=LIBS)
def lib(request):
    return oracle.load(request.param)


def test_golden_compress(lib):
    cases = G["compress"]
    datas, specs = golden_util.compress_inputs(cases)
    src = common.build_layout(sg.HostArena, specs, datas)
    dspecs = [sg.segs_single(sg.compress_bound(c["n"])) for c in cases]
    dst, (rets, _, so, do) = common.oracle_compress(lib, src, dspecs, want_iters=True)
    for b, c in enumerate(cases):
        assert rets[b] == c["ret"], c
        assert golden_util.sha(dst.read_block(b, rets[b])) == c["sha256"], c
        if "hex" in c:
            assert dst.read_block(b, rets[b]).tobytes().hex() == c["hex"]
        assert int(so[b]["bi_size"]) == c["src_left"] and int(do[b]["bi_size"]) == c["dst_left"], c


def test_golden_limited_output(lib):
    cases = G["limited"]
    datas = [common.synth(c["n"], c["cls"], c["seed"]) for c in cases]
    src = common.build_layout(sg.HostArena, [sg.segs_single(c["n"]) for c in cases], datas)
    dspecs = [dict(sg.segs_single(max(c["cap"], 1)), size=c["cap"]) for c in cases]
    dst, (rets, _) = common.oracle_compress(lib, src, dspecs)
    for b, c in enumerate(cases):
        assert rets[b] == c["ret"], c
        assert golden_util.sha(dst.read_block(b, rets[b])) == c["sha256"]


def test_golden_decompress(lib):
    cases = G["decompress"]
    comps = [np.frombuffer(bytes.fromhex(c["stream"]), dtype=np.uint8) for c in cases]
    rets, outs = common.oracle_decode_linear(lib, comps, [c["cap"] for c in cases])
    for b, c in enumerate(cases):
        assert rets[b] == c["ret"], (b, c["ret"], rets[b])
        assert golden_util.sha(outs[b]) == c["sha256"]


@pytest.mark.skipif(not os.path.isdir("/root/reference/test/test_files"),
                    reason="the reference's fixtures only exist in the build container")
def test_reference_fixtures_roundtrip(lib):
    """test/bash_tests/test_proxy.sh:22-38 of the reference: its three fixtures survive
    compress -> decompress; compressed sizes as recorded from the reference build."""
    for c in G["fixtures"]:
        d = np.fromfile(os.path.join("/root/reference/test/test_files", c["name"]), dtype=np.uint8)
        assert golden_util.sha(d) == c["in_sha256"]
        src = common.build_layout(sg.HostArena, [golden_util.seg_spec(c["seg"], len(d))], [d])
        dst, (rets, _) = common.oracle_compress(lib, src, [sg.segs_single(sg.compress_bound(len(d)))])
        assert rets[0] == c["ret"] and golden_util.sha(dst.read_block(0, rets[0])) == c["sha256"]
        r2, outs = common.oracle_decode_linear(lib, [dst.read_block(0, rets[0])], [len(d)])
        assert r2[0] == len(d) and np.array_equal(outs[0], d)


def test_config1_1mib_text_roundtrip(lib):
    """BASELINE.json configs[0]: 1 MiB synthetic text in 256 x 4 KiB bio_vec pages (byU32);
    SURVEY.md 8(c) records 372813 compressed bytes for the reference."""
    d = common.lcg_text(1 << 20)
    src = common.build_layout(sg.HostArena, [sg.segs_pages(1 << 20)], [d])
    dst, (rets, _) = common.oracle_compress(lib, src, [sg.segs_single(sg.compress_bound(1 << 20))])
    assert rets[0] == 372813
    r2, outs = common.oracle_decode_linear(lib, [dst.read_block(0, rets[0])], [1 << 20])
    assert r2[0] == (1 << 20) and np.array_equal(outs[0], d)


def test_survey_sanity_sizes(lib):
    t = common.lcg_text(65536)
    rnd = np.zeros(0)
    cases = [(t[:4096], sg.segs_single(4096), 1688), (t, sg.segs_pages(65536), 26412),
             (t, sg.segs_single(65536), 23373), (np.zeros(4096, np.uint8), sg.segs_single(4096), 26),
             (np.zeros(65536, np.uint8), sg.segs_pages(65536), 267), (t[:12], sg.segs_single(12), 13)]
    src = common.build_layout(sg.HostArena, [c[1] for c in cases], [c[0] for c in cases])
    _, (rets, _) = common.oracle_compress(lib, src, [sg.segs_single(sg.compress_bound(len(c[0]))) for c in cases])
    assert list(rets) == [c[2] for c in cases]
    del rnd


def test_257_bvecs_fail_256_ok(lib):
    d = common.synth(300, 0, 1)
    s257 = {"size": 257, "segs": [(i % 4000, 1) for i in range(257)], "skip": 0, "gap": 0}
    s256 = {"size": 256, "segs": [(i % 4000, 1) for i in range(256)], "skip": 0, "gap": 0}
    src = common.build_layout(sg.HostArena, [s257, s256], [d[:257], d[:256]])
    _, (rets, _) = common.oracle_compress(lib, src, [sg.segs_single(sg.compress_bound(300))] * 2)
    assert rets[0] == 0 and rets[1] > 0


@pytest.mark.skipif("reference" not in LIBS, reason="needs oracle/_ref")
def test_port_equals_reference_fuzz():
    """The structural facts the CUDA design rests on: output is a function of the
    linear bytes and the table type; destination segmentation is irrelevant; iterator end states;
    decoder return codes on malformed input."""
    ref, port = oracle.load("reference"), oracle.load("port")
    rng = np.random.default_rng(1)
    edge = [0, 1, 12, 13, 14, 100, 4095, 4096, 4097, 65535, 65536, 100000]
    datas, specs, dspecs = [], [], []
    for i in range(400):
        n = int(rng.choice(edge)) if rng.random() < 0.4 else int(rng.integers(0, 70000))
        datas.append(common.synth(n, int(rng.integers(0, 4)), i))
        specs.append(sg.segs_random(n, rng, max_segs=int(rng.choice([3, 16, 40, 300]))))
        cap = sg.compress_bound(n) if rng.random() < 0.7 else int(rng.integers(1, n + 20))
        dspecs.append(sg.segs_random(cap, rng))
    src = common.build_layout(sg.HostArena, specs, datas)
    a, (ra, _, sa, da) = common.oracle_compress(ref, src, dspecs, want_iters=True)
    b, (rb, _, sb, db) = common.oracle_compress(port, src, dspecs, want_iters=True)
    assert np.array_equal(ra, rb)
    ok = ra > 0
    assert np.array_equal(sa[ok], sb[ok]) and np.array_equal(da[ok], db[ok])
    for k in range(len(datas)):
        assert np.array_equal(a.read_block(k, ra[k]), b.read_block(k, rb[k])), k
    noLimit = np.array([int(d["size"]) >= sg.compress_bound(len(x)) and len(s["segs"]) <= 256
                        for d, x, s in zip(dspecs, datas, specs)])
    assert (ra[noLimit] > 0).all()  # noLimit mode cannot fail below 257 bvecs
    # decoder on valid + mutated streams
    comps, caps = [], []
    for k in np.nonzero(ok)[0][:250]:
        c = bytearray(a.read_block(k, ra[k]).tobytes())
        r = rng.random()
        if r < 0.3 and len(c) > 2:
            c = c[:int(rng.integers(1, len(c)))]
        elif r < 0.6:
            c[int(rng.integers(0, len(c)))] = int(rng.integers(0, 256))
        comps.append(np.frombuffer(bytes(c), dtype=np.uint8))
        caps.append(len(datas[k]) + (int(rng.integers(-9, 9)) if rng.random() < 0.3 else 0))
    caps = [max(c, 0) for c in caps]
    r1, o1 = common.oracle_decode_linear(ref, comps, caps)
    r2, o2 = common.oracle_decode_linear(port, comps, caps)
    assert np.array_equal(r1, r2)
    for x, y, r in zip(o1, o2, r1):
        if r >= 0:
            assert np.array_equal(x, y)
