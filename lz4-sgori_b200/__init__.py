"""lz4-sgori_b200 — B200-native LZ4E scatter-gather compression hot path.

This package is a thin ctypes view of liblz4e_b200.so (CUDA kernels for sm_100a
behind the C ABI of include/lz4e_b200.h plus the C host library with the
reference's LZ4E API).  Python is only used by tests/ and bench.py; the product
is the shared library.  There is no CPU fallback: importing works anywhere (so
the exported symbols can be checked), but every compute call needs a B200 and
raises otherwise.
"""
import ctypes
import os
import subprocess

import numpy as np

from . import sg  # noqa: F401  (re-exported helpers)
from .sg import BIO_VEC, BVEC_ITER, PAGE_SIZE, compress_bound  # noqa: F401

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LZ4E_B200_LIB") or os.path.join(HERE, "liblz4e_b200.so")

DEV_BVEC = np.dtype([("addr", "<u8"), ("len", "<u4"), ("rsvd", "<u4")])
DEV_ITER = np.dtype([("first", "<u4"), ("bi_size", "<u4"), ("bi_idx", "<u4"),
                     ("bi_bvec_done", "<u4")])
DEV_END = np.dtype([("src_pos", "<u4"), ("dst_pos", "<u4")])
REQ = np.dtype([("src", "<u8"), ("dst", "<u8"), ("src_iter", "<u8"), ("dst_iter", "<u8"),
                ("ret", "<i4"), ("rsvd", "<i4")])

F_SRC_RESIDENT, F_DST_ON_DEVICE, F_KEEP_ITERS, F_DST_PRESERVE = 1, 2, 4, 8
DEV_NO_BYU16, DEV_DST_LINEAR, DEV_SRC_LINEAR, DEV_DST_SEGS32 = 1, 2, 4, 8


class Lz4eError(RuntimeError):
    pass


def build(verbose=False):
    """Compile liblz4e_b200.so in-tree (nvcc -gencode arch=compute_100a,code=sm_100a)."""
    out = subprocess.run(["make", "-C", HERE], capture_output=True, text=True)
    if out.returncode != 0:
        raise Lz4eError("building liblz4e_b200.so failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)


_lib = None


def lib():
    """The loaded shared library.  Raises if it has not been built: the product has no
    other implementation to fall back to."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Lz4eError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; "
                        "g.build()'` (needs nvcc); there is no CPU fallback")
    L = ctypes.CDLL(LIB_PATH)
    vp, u32, u64, i32, sz = (ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_int,
                             ctypes.c_size_t)
    sig = {
        "lz4e_b200_device_count": (i32, []),
        "lz4e_b200_ctx_create": (i32, [i32, ctypes.POINTER(vp)]),
        "lz4e_b200_ctx_destroy": (None, [vp]),
        "lz4e_b200_default_ctx": (vp, []),
        "lz4e_b200_strerror": (ctypes.c_char_p, [vp]),
        "lz4e_b200_ctx_tune": (i32, [vp, ctypes.c_char_p, i32]),
        "lz4e_b200_ctx_device": (i32, [vp]),
        "lz4e_b200_ctx_sm_count": (i32, [vp]),
        "lz4e_b200_kernel_launches": (u64, [vp]),
        "lz4e_b200_sync": (i32, [vp, vp]),
        "lz4e_b200_compress_dev": (i32, [vp, vp, vp, vp, vp, vp, u32, u32, u32, vp, vp]),
        "lz4e_b200_compress_dev_static": (i32, [vp, vp, vp, vp, vp, vp, u32, u32, u32, vp, vp]),
        "lz4e_b200_decompress_dev": (i32, [vp, vp, vp, vp, vp, vp, u32, u32, u32, u32, vp]),
        "lz4e_b200_compress_compact_dev": (i32, [vp, vp, vp, vp, vp, u32, u32, u64, u32, u32,
                                                 u32, vp, vp]),
        "lz4e_b200_dev_alloc": (vp, [vp, sz]),
        "lz4e_b200_dev_free": (None, [vp, vp]),
        "lz4e_b200_host_alloc": (vp, [vp, sz]),
        "lz4e_b200_host_free": (None, [vp, vp]),
        "lz4e_b200_copy_h2d": (i32, [vp, vp, vp, vp, sz]),
        "lz4e_b200_copy_d2h": (i32, [vp, vp, vp, vp, sz]),
        "lz4e_b200_xfer_dev": (i32, [vp, vp, vp, vp, vp, u32, i32]),
        "lz4e_b200_xfer_skip_dev": (i32, [vp, vp, vp, vp, vp, u32, i32, u32]),
        "lz4e_b200_copy2d_h2d": (i32, [vp, vp, vp, sz, vp, sz, sz, sz]),
        "lz4e_b200_traffic": (i32, [vp, ctypes.POINTER(u64), ctypes.POINTER(u64)]),
        "lz4e_b200_arena_create": (i32, [vp, sz, ctypes.POINTER(vp)]),
        "lz4e_b200_arena_destroy": (None, [vp]),
        "lz4e_b200_arena_pages": (vp, [vp]),
        "lz4e_b200_arena_dev_base": (vp, [vp]),
        "lz4e_b200_arena_npages": (sz, [vp]),
        "lz4e_b200_arena_upload": (i32, [vp, sz, sz]),
        "lz4e_b200_arena_download": (i32, [vp, sz, sz]),
        "lz4e_b200_arena_translate": (u64, [vp, vp, ctypes.POINTER(vp)]),
        "lz4e_b200_compress_batch": (i32, [vp, vp, sz, ctypes.c_uint]),
        "lz4e_b200_decompress_batch": (i32, [vp, vp, sz, ctypes.c_uint]),
        "lz4e_b200_shard_range": (None, [u64, u32, u32, ctypes.POINTER(u64),
                                         ctypes.POINTER(u64)]),
        "LZ4E_compress_default": (i32, [vp, vp, vp, vp, vp]),
        "LZ4E_decompress_safe": (i32, [vp, vp, i32, i32]),
        "LZ4E_decompress_safe_sg": (i32, [vp, vp, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def shard_range(nblocks, rank, world):
    """Block range [first, first+count) of `rank`."""
    first, count = ctypes.c_uint64(), ctypes.c_uint64()
    lib().lz4e_b200_shard_range(nblocks, rank, world, ctypes.byref(first), ctypes.byref(count))
    return first.value, count.value


def _p(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return ctypes.c_void_p(a.ctypes.data)
    return ctypes.c_void_p(int(a))


class Context:
    """One GPU.  Raises Lz4eError when no B200 is usable (no CPU path)."""

    def __init__(self, device=0):
        h = ctypes.c_void_p()
        rc = lib().lz4e_b200_ctx_create(int(device), ctypes.byref(h))
        if rc != 0 or not h.value:
            raise Lz4eError(f"lz4e_b200_ctx_create(device={device}) failed with {rc}: "
                            "a CUDA sm_100 device is required, there is no CPU fallback")
        self.h = h
        self.device = device

    def close(self):
        if self.h:
            lib().lz4e_b200_ctx_destroy(self.h)
            self.h = None

    def err(self):
        return lib().lz4e_b200_strerror(self.h).decode()

    def check(self, rc, what):
        if rc != 0:
            raise Lz4eError(f"{what} failed with {rc}: {self.err()}")

    def tune(self, key, value):
        self.check(lib().lz4e_b200_ctx_tune(self.h, key.encode(), int(value)), "tune " + key)

    def sync(self, stream=None):
        self.check(lib().lz4e_b200_sync(self.h, _p(stream)), "sync")

    def launches(self):
        return int(lib().lz4e_b200_kernel_launches(self.h))

    def sm_count(self):
        return int(lib().lz4e_b200_ctx_sm_count(self.h))

    # ---- raw device memory (numpy <-> device) ----
    def dev_alloc(self, nbytes):
        p = lib().lz4e_b200_dev_alloc(self.h, int(nbytes))
        if not p:
            raise Lz4eError("device allocation failed: " + self.err())
        return int(p)

    def dev_free(self, p):
        lib().lz4e_b200_dev_free(self.h, _p(p))

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        p = self.dev_alloc(max(arr.nbytes, 1))
        self.check(lib().lz4e_b200_copy_h2d(self.h, None, _p(p), _p(arr), arr.nbytes), "h2d")
        self.sync()
        return p

    def download(self, p, dtype, count):
        out = np.zeros(count, dtype=dtype)
        self.check(lib().lz4e_b200_copy_d2h(self.h, None, _p(out), _p(p), out.nbytes), "d2h")
        self.sync()
        return out

    # ---- device-level calls ----
    def compress_dev(self, sbv, sit, dbv, dit, nblocks, max_in, rets, ends=None, flags=0,
                     stream=None):
        self.check(lib().lz4e_b200_compress_dev(self.h, _p(stream), _p(sbv), _p(sit), _p(dbv),
                                                _p(dit), nblocks, max_in, flags, _p(rets),
                                                _p(ends)), "compress_dev")

    def compress_dev_static(self, sbv, sit, dbv, dit, nblocks, max_in, rets, ends=None, flags=0,
                            stream=None):
        """compress_dev with the kernel's scratch in the library's fixed device array: no
        allocation inside the call (include/lz4e_b200.h)."""
        self.check(lib().lz4e_b200_compress_dev_static(self.h, _p(stream), _p(sbv), _p(sit),
                                                       _p(dbv), _p(dit), nblocks, max_in, flags,
                                                       _p(rets), _p(ends)), "compress_dev_static")

    def decompress_dev(self, sbv, sit, dbv, dit, nblocks, max_in, max_out, rets, flags=0,
                       stream=None):
        self.check(lib().lz4e_b200_decompress_dev(self.h, _p(stream), _p(sbv), _p(sit), _p(dbv),
                                                  _p(dit), nblocks, max_in, max_out, flags,
                                                  _p(rets)), "decompress_dev")

    def compress_compact_dev(self, sbv, sit, dbv, dst_nbvecs, dst_done, dst_cap, nblocks, max_in,
                             rets, offsets, flags=0, stream=None):
        self.check(lib().lz4e_b200_compress_compact_dev(
            self.h, _p(stream), _p(sbv), _p(sit), _p(dbv), dst_nbvecs, dst_done, dst_cap, nblocks,
            max_in, flags, _p(rets), _p(offsets)), "compress_compact_dev")

    def xfer_dev(self, bv, it, lens, nblocks, to_host, stream=None):
        self.check(lib().lz4e_b200_xfer_dev(self.h, _p(stream), _p(bv), _p(it), _p(lens), nblocks,
                                            1 if to_host else 0), "xfer_dev")

    def traffic(self):
        """(h2d bytes, d2h bytes) moved through this context so far."""
        a, b = ctypes.c_uint64(), ctypes.c_uint64()
        self.check(lib().lz4e_b200_traffic(self.h, ctypes.byref(a), ctypes.byref(b)), "traffic")
        return a.value, b.value

    # ---- host-level calls ----
    def compress_batch(self, reqs, flags=0):
        assert reqs.dtype == REQ
        return lib().lz4e_b200_compress_batch(self.h, _p(reqs), len(reqs), flags)

    def decompress_batch(self, reqs, flags=0):
        assert reqs.dtype == REQ
        return lib().lz4e_b200_decompress_batch(self.h, _p(reqs), len(reqs), flags)


class Arena(sg.HostArena):
    """Pinned host pages with a device mirror (include/lz4e_b200.h, layer 2)."""

    def __init__(self, ctx, npages):
        h = ctypes.c_void_p()
        ctx.check(lib().lz4e_b200_arena_create(ctx.h, int(npages), ctypes.byref(h)),
                  "arena_create")
        self.h = h
        self.ctx = ctx
        host = lib().lz4e_b200_arena_pages(h)
        buf = np.ctypeslib.as_array(ctypes.cast(host, ctypes.POINTER(ctypes.c_uint8)),
                                    shape=(int(npages) * PAGE_SIZE,))
        super().__init__(npages, buf=buf, base_addr=host)
        self.dev_base = int(lib().lz4e_b200_arena_dev_base(h))

    def upload(self, first=0, npages=None):
        npages = self.npages - first if npages is None else npages
        self.ctx.check(lib().lz4e_b200_arena_upload(self.h, first, npages), "arena_upload")
        self.ctx.sync()

    def download(self, first=0, npages=None):
        npages = self.npages - first if npages is None else npages
        self.ctx.check(lib().lz4e_b200_arena_download(self.h, first, npages), "arena_download")
        self.ctx.sync()

    def close(self):
        if self.h:
            lib().lz4e_b200_arena_destroy(self.h)
            self.h = None


def to_dev(layout, dev_base=None):
    """bio_vec/bvec_iter tables of a Layout -> device descriptor tables (numpy), with page
    addresses moved from the arena's host base to `dev_base` (default: the arena's mirror)."""
    if dev_base is None:
        dev_base = layout.arena.dev_base
    delta = np.uint64((int(dev_base) - layout.arena.base) % (1 << 64))
    bv = np.zeros(len(layout.bvecs), dtype=DEV_BVEC)
    bv["addr"] = layout.bvecs["bv_page"] + layout.bvecs["bv_offset"].astype(np.uint64) + delta
    bv["len"] = layout.bvecs["bv_len"]
    it = np.zeros(len(layout.iters), dtype=DEV_ITER)
    it["first"] = layout.first.astype(np.uint32)
    it["bi_size"] = layout.iters["bi_size"]
    it["bi_idx"] = layout.iters["bi_idx"]
    it["bi_bvec_done"] = layout.iters["bi_bvec_done"]
    return bv, it


def make_reqs(src_layout, dst_layout):
    """struct lz4e_b200_req array over two Layouts (host pointers into their tables).
    Returns (reqs, keepalive) — keep `keepalive` referenced while reqs is in use."""
    n = src_layout.nblocks()
    reqs = np.zeros(n, dtype=REQ)
    si = src_layout.iters.copy()
    di = dst_layout.iters.copy()
    reqs["src"] = src_layout.bvecs.ctypes.data + src_layout.first * np.uint64(16)
    reqs["dst"] = dst_layout.bvecs.ctypes.data + dst_layout.first * np.uint64(16)
    reqs["src_iter"] = si.ctypes.data + np.arange(n, dtype=np.uint64) * np.uint64(20)
    reqs["dst_iter"] = di.ctypes.data + np.arange(n, dtype=np.uint64) * np.uint64(20)
    return reqs, (si, di)
