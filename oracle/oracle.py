"""oracle/oracle.py — TEST INFRASTRUCTURE, not product code.

ctypes loader for the two CPU checkers:
  * oracle/_ref/liblz4e_ref.so   — the reference's own lz4e_compress.c /
    lz4e_decompress.c compiled unmodified against oracle/shim ("reference")
  * oracle/_build/liblz4e_port.so — oracle/lz4e_port.c, the linear-position
    restatement ("port")

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may
import this module.  The product (lz4-sgori_b200/) never does.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "liblz4e_ref.so")
PORT_SO = os.path.join(HERE, "_build", "liblz4e_port.so")

# struct bio_vec / struct bvec_iter exactly as the reference's callers lay them
# out (doc/API.md:46-82; Linux include/linux/bvec.h)
BIO_VEC = np.dtype([("bv_page", "<u8"), ("bv_len", "<u4"), ("bv_offset", "<u4")])
BVEC_ITER = np.dtype([("bi_sector", "<u8"), ("bi_size", "<u4"), ("bi_idx", "<u4"),
                      ("bi_bvec_done", "<u4")])
assert BIO_VEC.itemsize == 16 and BVEC_ITER.itemsize == 20


def build(verbose=False):
    """Build the port always, and the reference oracle when /root/reference exists."""
    out = subprocess.run(["make", "-C", HERE, "all"], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)


def _bind(path):
    lib = ctypes.CDLL(path)
    vp = ctypes.c_void_p
    lib.oracle_compress_batch.restype = ctypes.c_double
    lib.oracle_compress_batch.argtypes = [vp, vp, vp, vp, vp, vp, ctypes.c_int64, vp, vp, vp,
                                          ctypes.c_int]
    lib.oracle_decompress_batch.restype = ctypes.c_double
    lib.oracle_decompress_batch.argtypes = [vp, vp, vp, vp, vp, vp, ctypes.c_int64, vp,
                                            ctypes.c_int]
    lib.oracle_synth_blocks.restype = None
    lib.oracle_synth_blocks.argtypes = [vp, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint64,
                                        ctypes.c_uint64, ctypes.c_int]
    lib.LZ4E_decompress_safe.restype = ctypes.c_int
    lib.LZ4E_decompress_safe.argtypes = [vp, vp, ctypes.c_int, ctypes.c_int]
    assert lib.oracle_sizeof_bio_vec() == 16
    assert lib.oracle_sizeof_bvec_iter() == 20
    assert lib.oracle_sizeof_page() == 4096
    return lib


def have_ref():
    return os.path.exists(REF_SO)


def load(kind="auto"):
    """kind: 'reference' | 'port' | 'auto' (reference when present, else port)."""
    if kind == "auto":
        kind = "reference" if have_ref() else "port"
    path = REF_SO if kind == "reference" else PORT_SO
    if not os.path.exists(path):
        build()
    if not os.path.exists(path):
        raise RuntimeError(f"oracle library {path} is not available")
    lib = _bind(path)
    lib.kind = kind
    return lib


def _p(a):
    return ctypes.c_void_p(a.ctypes.data) if a is not None else None


def compress_batch(lib, src_bvecs, src_first, src_iters, dst_bvecs, dst_first, dst_iters,
                   nthreads=1, want_iters=False):
    """Run LZ4E_compress_default once per block.  All arrays are numpy arrays in
    host memory whose bv_page fields hold HOST addresses.  Returns (rets, seconds)
    or (rets, seconds, src_iters_out, dst_iters_out)."""
    n = len(src_iters)
    assert src_bvecs.dtype == BIO_VEC and dst_bvecs.dtype == BIO_VEC
    assert src_iters.dtype == BVEC_ITER and dst_iters.dtype == BVEC_ITER
    src_first = np.ascontiguousarray(src_first, dtype=np.uint64)
    dst_first = np.ascontiguousarray(dst_first, dtype=np.uint64)
    rets = np.zeros(n, dtype=np.int32)
    so = np.zeros(n, dtype=BVEC_ITER) if want_iters else None
    do = np.zeros(n, dtype=BVEC_ITER) if want_iters else None
    secs = lib.oracle_compress_batch(_p(src_bvecs), _p(src_first), _p(src_iters), _p(dst_bvecs),
                                     _p(dst_first), _p(dst_iters), n, _p(rets), _p(so), _p(do),
                                     int(nthreads))
    if want_iters:
        return rets, secs, so, do
    return rets, secs


def decompress_batch(lib, src, src_off, src_len, dst, dst_off, dst_cap, nthreads=1):
    """Contiguous LZ4E_decompress_safe per block; src/dst are uint8 numpy arrays."""
    n = len(src_len)
    src_off = np.ascontiguousarray(src_off, dtype=np.uint64)
    dst_off = np.ascontiguousarray(dst_off, dtype=np.uint64)
    src_len = np.ascontiguousarray(src_len, dtype=np.int32)
    dst_cap = np.ascontiguousarray(dst_cap, dtype=np.int32)
    rets = np.zeros(n, dtype=np.int32)
    secs = lib.oracle_decompress_batch(_p(src), _p(src_off), _p(src_len), _p(dst), _p(dst_off),
                                       _p(dst_cap), n, _p(rets), int(nthreads))
    return rets, secs


def synth_blocks(lib, mem, stride, n, nblocks, first=0, nthreads=1):
    """Benchmark contents (tools/synth/lz4e_synth.h) for nblocks blocks of n bytes at
    mem[b * stride:]; class and seed from first + b."""
    assert mem.dtype == np.uint8 and mem.nbytes >= (nblocks - 1) * stride + n
    lib.oracle_synth_blocks(_p(mem), int(stride), int(n), int(nblocks), int(first), int(nthreads))
