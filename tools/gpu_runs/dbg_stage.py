import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, common
from common import pkg, sg
from oracle import oracle
ref = oracle.load("auto")
ctx = pkg.Context(0)
n = 64
datas = [common.synth(4096, b & 3, b) for b in range(n)]
src = common.build_layout(sg.HostArena, [sg.segs_single(4096)] * n, datas)
cdst, (sizes, _) = common.oracle_compress(ref, src, [sg.segs_single(sg.compress_bound(4096))] * n)
comps = [cdst.read_block(b, sizes[b]) for b in range(n)]
mk = lambda: pkg.Arena(ctx, 4096)
csrc = common.build_layout(lambda k: pkg.Arena(ctx, k), [sg.segs_single(len(c)) for c in comps], comps)
for var, grp, stg in [(2, 8, 1), (1, 8, 1), (2, 8, 0), (2, 4, 1)]:
    ctx.tune("dvariant", var); ctx.tune("dgroup", grp); ctx.tune("dstage", stg)
    out = common.build_layout(lambda k: pkg.Arena(ctx, k), [sg.segs_single(4096)] * n)
    out.arena.mem[:] = 0xEE
    try:
        r = common.gpu_decompress(ctx, csrc, sizes, out)
        good = sum(np.array_equal(out.read_block(b), datas[b]) for b in range(n))
        print(var, grp, stg, "rets", r[:8], "good", good, flush=True)
    except Exception as e:
        print(var, grp, stg, "EXC", e, flush=True)
