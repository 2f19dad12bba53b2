"""tests/emu.py — TEST INFRASTRUCTURE: ctypes front end of tests/simt (the CUDA
kernel bodies compiled for the host and run by a 32-lane coroutine emulator)."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SIMT_DIR = os.path.join(HERE, "simt")
SIMT_SO = os.path.join(SIMT_DIR, "_build", "liblz4e_simt.so")

DEV_BVEC = np.dtype([("addr", "<u8"), ("len", "<u4"), ("rsvd", "<u4")])
DEV_ITER = np.dtype([("first", "<u4"), ("bi_size", "<u4"), ("bi_idx", "<u4"),
                     ("bi_bvec_done", "<u4")])
DEV_END = np.dtype([("src_pos", "<u4"), ("dst_pos", "<u4")])


def load():
    out = subprocess.run(["make", "-C", SIMT_DIR], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("simt build failed:\n" + out.stdout + out.stderr)
    lib = ctypes.CDLL(SIMT_SO)
    vp, u32 = ctypes.c_void_p, ctypes.c_uint32
    lib.simt_compress_batch.restype = ctypes.c_int
    lib.simt_compress_batch.argtypes = [vp, vp, vp, vp, u32, u32, u32, u32, vp, vp, vp, u32, u32]
    lib.simt_decompress_batch.restype = ctypes.c_int
    lib.simt_decompress_batch.argtypes = [vp, vp, vp, vp, u32, u32, u32, vp, vp, vp, u32, u32, u32]
    return lib


def to_dev(layout, addr_delta=0):
    """bio_vec/bvec_iter layout -> device descriptor tables (addresses shifted by
    addr_delta = device base - host base)."""
    bv = np.zeros(len(layout.bvecs), dtype=DEV_BVEC)
    bv["addr"] = layout.bvecs["bv_page"] + layout.bvecs["bv_offset"] + np.uint64(addr_delta)
    bv["len"] = layout.bvecs["bv_len"]
    it = np.zeros(len(layout.iters), dtype=DEV_ITER)
    it["first"] = layout.first.astype(np.uint32)
    it["bi_size"] = layout.iters["bi_size"]
    it["bi_idx"] = layout.iters["bi_idx"]
    it["bi_bvec_done"] = layout.iters["bi_bvec_done"]
    return bv, it


def _p(a):
    return ctypes.c_void_p(a.ctypes.data) if a is not None else None


def compress(lib, src_layout, dst_layout, in_cap=4096, out_cap=4352, seed=1, table_bytes=16384,
             gin=False):
    sbv, sit = to_dev(src_layout)
    dbv, dit = to_dev(dst_layout)
    n = len(sit)
    rets = np.full(n, -777, dtype=np.int32)
    ends = np.zeros(n, dtype=DEV_END)
    faults = np.zeros(1, dtype=np.uint32)
    rc = lib.simt_compress_batch(_p(sbv), _p(sit), _p(dbv), _p(dit), n, table_bytes, in_cap, out_cap,
                                 _p(rets), _p(ends), _p(faults), seed, 1 if gin else 0)
    assert rc == 0
    return rets, ends, int(faults[0])


def decompress(lib, src_layout, dst_layout, group=8, stage_cap=4096, scratch_stride=0, seed=1,
               variant=0):
    sbv, sit = to_dev(src_layout)
    dbv, dit = to_dev(dst_layout)
    n = len(sit)
    rets = np.full(n, -777, dtype=np.int32)
    faults = np.zeros(1, dtype=np.uint32)
    scratch = np.zeros(max(1, (32 // group) * scratch_stride + 64), dtype=np.uint8)
    rc = lib.simt_decompress_batch(_p(sbv), _p(sit), _p(dbv), _p(dit), n, group, stage_cap,
                                   _p(rets), _p(faults), _p(scratch), scratch_stride, seed, variant)
    assert rc == 0
    return rets, int(faults[0])
