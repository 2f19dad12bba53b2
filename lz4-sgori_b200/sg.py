"""Scatter-gather helpers: numpy mirrors of struct bio_vec / struct bvec_iter and
builders that lay blocks out over page arenas the way the reference's callers do
(doc/API.md:46-82; lz4e_bdev/lz4e_req.c:115-142 builds bvecs page by page).

Pure numpy, no CUDA: used by tests/ and bench.py to describe workloads.
"""
import numpy as np

PAGE_SIZE = 4096
BIO_MAX_VECS = 256

# include/lz4e_bvec.h
BIO_VEC = np.dtype([("bv_page", "<u8"), ("bv_len", "<u4"), ("bv_offset", "<u4")])
BVEC_ITER = np.dtype([("bi_sector", "<u8"), ("bi_size", "<u4"), ("bi_idx", "<u4"),
                      ("bi_bvec_done", "<u4")])


def compress_bound(n):
    """LZ4E_COMPRESSBOUND, lz4e/include/lz4e.h:25-28."""
    n = int(n)
    return 0 if n > 0x7E000000 else n + n // 255 + 16


class HostArena:
    """npages 4 KiB pages of page-aligned host memory; page i is `struct page *`
    base + i*4096."""

    def __init__(self, npages, buf=None, base_addr=None):
        self.npages = int(npages)
        if buf is None:
            raw = np.zeros(self.npages * PAGE_SIZE + PAGE_SIZE, dtype=np.uint8)
            shift = (-raw.ctypes.data) % PAGE_SIZE
            buf = raw[shift:shift + self.npages * PAGE_SIZE]
            self._raw = raw
        self.mem = buf
        self.base = int(base_addr if base_addr is not None else buf.ctypes.data)
        assert self.base % PAGE_SIZE == 0

    def page_addr(self, i):
        return self.base + int(i) * PAGE_SIZE


class Layout:
    """A batch of blocks laid over an arena: shared bvec table, per-block index of
    bvec[0] (`first`), per-block iterator, plus where every byte lives."""

    def __init__(self, arena, bvecs, first, iters, npages_used):
        self.arena = arena
        self.bvecs = bvecs
        self.first = first
        self.iters = iters
        self.npages_used = npages_used

    def nblocks(self):
        return len(self.iters)

    def _pieces(self, b):
        """Yield (arena byte offset, length) for block b in stream order."""
        it = self.iters[b]
        idx = int(self.first[b]) + int(it["bi_idx"])
        done = int(it["bi_bvec_done"])
        left = int(it["bi_size"])
        while left:
            v = self.bvecs[idx]
            take = min(left, int(v["bv_len"]) - done)
            off = int(v["bv_page"]) - self.arena.base + int(v["bv_offset"]) + done
            yield off, take
            left -= take
            idx += 1
            done = 0

    def write_block(self, b, data):
        data = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        pos = 0
        for off, take in self._pieces(b):
            self.arena.mem[off:off + take] = data[pos:pos + take]
            pos += take
        assert pos == len(data), (pos, len(data))

    def read_block(self, b, nbytes=None):
        out = []
        left = None if nbytes is None else int(nbytes)
        for off, take in self._pieces(b):
            if left is not None:
                take = min(take, left)
                left -= take
            out.append(self.arena.mem[off:off + take])
            if left == 0:
                break
        return np.concatenate(out) if out else np.zeros(0, dtype=np.uint8)


def layout_from_segments(arena, blocks, first_page=0):
    """blocks: list of dicts {size, segs: [(bv_offset, bv_len), ...], skip, lead}.
    `segs` are the RAW bvecs (bv_offset may be anywhere in its first page, bv_len may
    span pages = multi-page bvec), `lead` extra bvecs placed BEFORE the iterator's
    bi_idx (so bi_idx != 0), `skip` = bi_bvec_done of the first covered bvec.
    Every bvec gets its own run of pages, one guard page apart, so segments are
    non-contiguous in memory."""
    bvecs, first, iters = [], [], []
    page = int(first_page)
    for blk in blocks:
        segs = list(blk["segs"])
        lead = int(blk.get("lead", 0))
        skip = int(blk.get("skip", 0))
        size = int(blk["size"])
        first.append(len(bvecs))
        for _ in range(lead):
            bvecs.append((arena.page_addr(page), 512, 0))
            page += 2
        covered = -skip
        for (off, ln) in segs:
            npg = (off + ln + PAGE_SIZE - 1) // PAGE_SIZE
            bvecs.append((arena.page_addr(page), ln, off))
            page += npg + int(blk.get("gap", 1))
            covered += ln
        assert covered >= size, "segments do not cover the block"
        iters.append((0, size, lead, skip))
    assert page <= arena.npages, (page, arena.npages)
    return Layout(arena, np.array(bvecs, dtype=BIO_VEC).reshape(-1),
                  np.array(first, dtype=np.uint64), np.array(iters, dtype=BVEC_ITER).reshape(-1), page)


def pages_needed(blocks):
    tot = 0
    for blk in blocks:
        tot += 2 * int(blk.get("lead", 0))
        for (off, ln) in blk["segs"]:
            tot += (off + ln + PAGE_SIZE - 1) // PAGE_SIZE + int(blk.get("gap", 1))
    return tot


def segs_single(size):
    """One bvec at offset 0 (multi-page when size > 4096)."""
    return {"size": size, "segs": [(0, max(size, 1))], "skip": 0}


def segs_pages(size):
    """4 KiB page bvecs at offset 0 — how lz4e_add_buf_to_bio (lz4e_req.c:115-142)
    wraps a buffer."""
    segs, left = [], size
    while left > 0:
        segs.append((0, min(left, PAGE_SIZE)))
        left -= PAGE_SIZE
    return {"size": size, "segs": segs or [(0, 1)], "skip": 0}


def segs_odd(size, nseg=16, block_index=0):
    """Config 3 of BASELINE.json: `nseg` non-contiguous segments with odd bv_offset
    ((53*i+7) mod 509 style), odd lengths around size/nseg and a mid-bvec iterator
    start."""
    base = size // nseg
    skip = 1 + (block_index * 2654435761 >> 7) % min(4095, max(base - 1, 1))
    lens = []
    for i in range(nseg):
        delta = ((block_index * 31 + i * 17) % 129) - 64
        if i % 2:
            delta = -delta
        lens.append(max(1, base + delta))
    lens[0] += skip
    tot = sum(lens) - skip
    lens[-1] += size - tot
    if lens[-1] < 1:  # tiny blocks: fold into one segment
        return {"size": size, "segs": [(7, size + skip)], "skip": skip}
    segs = [(((53 * (i + block_index) + 7) % 509), ln) for i, ln in enumerate(lens)]
    return {"size": size, "segs": segs, "skip": skip}


def segs_random(size, rng, max_segs=24, allow_big=True):
    """Random segmentation for fuzzing: random count, lengths, offsets, mid-bvec
    start and leading bvecs."""
    if size == 0:
        return {"size": 0, "segs": [(int(rng.integers(0, 4000)), 8)], "skip": 0,
                "lead": int(rng.integers(0, 3))}
    nseg = int(rng.integers(1, max(2, min(max_segs, size) + 1)))
    cuts = np.sort(rng.choice(np.arange(1, size), size=min(nseg - 1, max(size - 1, 0)), replace=False)) \
        if size > 1 and nseg > 1 else np.zeros(0, dtype=np.int64)
    edges = [0] + [int(c) for c in cuts] + [size]
    lens = [edges[i + 1] - edges[i] for i in range(len(edges) - 1)]
    skip = int(rng.integers(0, 3000)) if rng.random() < 0.6 else 0
    lens[0] += skip
    if rng.random() < 0.4:  # slack after the data in the last bvec
        lens[-1] += int(rng.integers(1, 700))
    segs = []
    for ln in lens:
        off = int(rng.integers(0, PAGE_SIZE)) if (allow_big or ln <= PAGE_SIZE) else \
            int(rng.integers(0, PAGE_SIZE - ln + 1))
        segs.append((off, ln))
    return {"size": size, "segs": segs, "skip": skip, "lead": int(rng.integers(0, 3))}
