#!/usr/bin/env python
"""tools/seq_stats.py — sequence statistics of the bench content classes (CPU, oracle):
sequences per block, literal/match length and offset distributions."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import common  # noqa: E402
from common import sg  # noqa: E402
from oracle import oracle  # noqa: E402


def parse(c):
    ip, seqs = 0, []
    n = len(c)
    while ip < n:
        tok = c[ip]; ip += 1
        lit = tok >> 4
        if lit == 15:
            while True:
                s = c[ip]; ip += 1; lit += s
                if s != 255:
                    break
        ip += lit
        if ip >= n:
            seqs.append((lit, 0, 0))
            break
        off = c[ip] | (c[ip + 1] << 8); ip += 2
        ml = tok & 15
        if ml == 15:
            while True:
                s = c[ip]; ip += 1; ml += s
                if s != 255:
                    break
        seqs.append((lit, ml + 4, off))
    return seqs


def main():
    size = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    lib = oracle.load("auto")
    for cls in range(4):
        datas = [common.synth(size, cls, 4 * b + cls) for b in range(8)]
        src = common.build_layout(sg.HostArena, [sg.segs_single(size)] * 8, datas)
        cdst, (sizes, _) = common.oracle_compress(lib, src, [sg.segs_single(sg.compress_bound(size))] * 8)
        S = []
        for b in range(8):
            S += parse(bytes(cdst.read_block(b, sizes[b])))
        a = np.array(S)
        lit, ml, off = a[:, 0], a[:, 1], a[:, 2]
        print("class %d: %.0f seq/block, %.1f B/seq; lit mean %.1f p50 %d p90 %d max %d (zero %.0f%%); "
              "match mean %.1f p50 %d p90 %d max %d; off p10 %d p50 %d p90 %d; overlap %.0f%%; units8 lit %.2f match %.2f" %
              (cls, len(S) / 8, size * 8 / len(S), lit.mean(), np.median(lit), np.percentile(lit, 90), lit.max(),
               100 * (lit == 0).mean(), ml.mean(), np.median(ml), np.percentile(ml, 90), ml.max(),
               np.percentile(off, 10), np.median(off), np.percentile(off, 90), 100 * ((off < ml) & (ml > 0)).mean(),
               np.ceil(lit / 8).mean(), np.ceil(ml / 8).mean()))


if __name__ == "__main__":
    main()
