"""Hardware parity of the kernels the size heuristics rarely pick in small tests: every decoder
(lane-per-block, wave, group, pipelined group) forced through lz4e_b200_ctx_tune and run over
the malformed-stream, fuzz and scatter-gather cases against the reference decoder; byU64 table
mode on the GPU; limitedOutput through the global-input (GIN) compress kernels."""
import numpy as np
import pytest

import common
from common import pkg, sg

import test_gpu_parity as P

pytestmark = pytest.mark.gpu

# (dvariant, dgroup): 0/1 = lane-per-block decoder, 3 = wave decoder (one warp per block),
# 2 = pipelined group decoder, 1 = group decoder, 0/8 = round-1 group decoder
DECODERS = [(0, 1), (3, 32), (2, 8), (2, 16), (1, 8), (0, 8)]
DEFAULTS = {"dvariant": -1, "dgroup": 0, "dstage": -1, "dorder": 2048, "gin_small": 2,
            "gin_min": 12288, "tma": 0, "wave_serial": 32, "p12": 0}


@pytest.fixture
def tuned(ctx):
    def apply(**kw):
        for k, v in kw.items():
            ctx.tune(k, v)
    yield apply
    for k, v in DEFAULTS.items():
        ctx.tune(k, v)


@pytest.mark.parametrize("variant,group", DECODERS)
def test_forced_decoder_malformed(ctx, ref_lib, tuned, variant, group):
    """Error codes -(ip)-1 and partial outputs of every decoder, linear and SG layouts."""
    tuned(dvariant=variant, dgroup=group, dorder=0)
    P.test_decompress_malformed(ctx, ref_lib)
    tuned(dorder=1)  # size-ordered work queue on even for this small batch
    P.test_decompress_malformed(ctx, ref_lib)


@pytest.mark.parametrize("variant,group", DECODERS)
def test_forced_decoder_fuzz_sg(ctx, ref_lib, tuned, variant, group):
    """Fuzzed sizes and segmentations: SG source (gather slot) and SG destination (scatter
    slot / segment map) of every decoder, plus the overlap / long-literal stress set."""
    tuned(dvariant=variant, dgroup=group)
    P.test_fuzz_segmentations_and_sizes(ctx, ref_lib)
    P.test_decompress_stress_overlap_and_literals(ctx, ref_lib)
    P.test_config3_64k_odd_segments(ctx, ref_lib)


def test_byu64_on_gpu(ctx, ref_lib):
    """Table mode byU64 (hash4, hashLog 11; lz4e_compress.c:50-51, :204-205): selected by a RAW
    bv_len above 16 MiB, whatever the iterator covers.  One 16 MiB + 4 KiB bvec per block,
    bi_size 64 KiB / 4 KiB / 70 000, mid-bvec start for one of them."""
    big = (1 << 24) + 4096
    datas = [common.lcg_text(65536), common.synth(65536, 3, 5), common.synth(4096, 0, 2),
             common.synth(70000, 2, 3), np.zeros(65536, np.uint8)]
    specs = [{"size": len(d), "segs": [(0, big)], "skip": 0} for d in datas]
    specs[1] = {"size": 65536, "segs": [(123, big)], "skip": 4321}
    # a byU32 twin of block 0 in the same launch must still differ from the byU64 one
    datas.append(datas[0])
    specs.append(sg.segs_single(65536))
    src, dst, rets = P._check_compress(ctx, ref_lib, datas, specs)
    assert rets[0] != rets[5]  # different table modes give different streams
    out = common.build_layout(lambda n: pkg.Arena(ctx, n), [sg.segs_single(len(d)) for d in datas])
    r2 = common.gpu_decompress(ctx, dst, rets, out)
    for b, d in enumerate(datas):
        assert r2[b] == len(d)
        assert np.array_equal(out.read_block(b), d)


def test_byu64_host_api(ctx, ref_lib):
    """The same rule through lz4e_b200_compress_batch (the host walk builds the tables)."""
    big = (1 << 24) + 4096
    d = common.synth(65536, 3, 17)
    spec = {"size": 65536, "segs": [(0, big)], "skip": 0}
    arena = lambda n: pkg.Arena(ctx, n)  # noqa: E731
    src = common.build_layout(arena, [spec], [d])
    dst = common.build_layout(arena, [sg.segs_single(sg.compress_bound(65536))])
    reqs, keep = pkg.make_reqs(src, dst)
    assert ctx.compress_batch(reqs, flags=0) == 0, ctx.err()
    ref_dst, (ref_rets, _) = common.oracle_compress(ref_lib, src, [sg.segs_single(sg.compress_bound(65536))])
    assert int(reqs["ret"][0]) == int(ref_rets[0])
    assert np.array_equal(dst.read_block(0, ref_rets[0]), ref_dst.read_block(0, ref_rets[0]))


@pytest.mark.parametrize("gin_small", [0, 2])
def test_limited_output_64k(ctx, ref_lib, tuned, gin_small):
    """limitedOutput (lz4e_compress.c:358-363, :425-430, :509) on 64 KiB blocks: exact fit
    succeeds, one byte less returns 0 — through the GIN kernels (gin_small=0 forces them for a
    small batch) and the shared-memory kernels, linear and SG destinations."""
    tuned(gin_small=gin_small)
    base = [common.synth(65536, c, 40 + c) for c in range(4)] + [common.lcg_text(50000)]
    specs0 = [sg.segs_odd(len(d), 16, i) for i, d in enumerate(base)]
    src = common.build_layout(lambda n: pkg.Arena(ctx, n), specs0, base)
    full = common.build_layout(lambda n: pkg.Arena(ctx, n),
                               [sg.segs_single(sg.compress_bound(len(d))) for d in base])
    sizes = common.gpu_compress(ctx, src, full)
    datas, specs, dlin, dsg = [], [], [], []
    for i, (d, c) in enumerate(zip(base, sizes)):
        for k, cap in enumerate((int(c), int(c) - 1, int(c) - 9, int(c) // 2, 100)):
            datas.append(d)
            specs.append(specs0[i])
            dlin.append(dict(sg.segs_single(cap), size=cap))
            dsg.append(sg.segs_odd(cap, 5, i + k))
    for dspecs, flags in ((dlin, 0), (dlin, pkg.DEV_DST_LINEAR), (dsg, 0)):
        _, _, rets = P._check_compress(ctx, ref_lib, datas, specs, dspecs, flags=flags,
                                       check_untouched=False)
        assert (rets[0::5] == sizes).all() and (rets[1::5] == 0).all()


def test_tma_staged_input_parity(ctx, ref_lib, tuned):
    """north_star's "per-page 1-D TMA bulk staging": with tma=1 the shared-memory compress
    kernels bring every 16-byte-congruent segment in with cp.async.bulk + mbarrier (heads, tails
    and incongruent segments by the lanes).  Same bytes as the reference on aligned pages, page
    bvecs, odd segments and fuzzed layouts."""
    tuned(tma=1)
    P.test_config2_4k_mixed(ctx, ref_lib)
    P.test_64k_page_bvecs_byu16(ctx, ref_lib)
    P.test_config3_64k_odd_segments(ctx, ref_lib)
    P.test_fuzz_segmentations_and_sizes(ctx, ref_lib)
    P.test_limited_output(ctx, ref_lib)
    P.test_survey_sanity_values(ctx, ref_lib)


@pytest.mark.parametrize("p12", [0, 1])
def test_packed_table_parity(ctx, ref_lib, tuned, p12):
    """Blocks of at most 4 KiB: the packed 12-bit hash table (an option) and the plain 16-bit one
    produce the reference's bytes; larger blocks in other launches are unaffected."""
    tuned(p12=p12)
    P.test_config2_4k_mixed(ctx, ref_lib)
    P.test_limited_output(ctx, ref_lib)
    rng = np.random.default_rng(77)
    datas, specs = [], []
    for i in range(200):
        n = int(rng.choice([0, 1, 12, 13, 14, 100, 4095, 4096])) if rng.random() < 0.4 \
            else int(rng.integers(0, 4097))
        datas.append(common.synth(n, int(rng.integers(0, 4)), i))
        specs.append(sg.segs_random(n, rng, max_segs=10))
    P._check_compress(ctx, ref_lib, datas, specs)
