"""tools/synth/lz4e_synth.py — ctypes view of tools/synth/liblz4e_synth.so, the synthetic
block generator of the benchmark configs (BENCHMARK/TEST DATA, not product code and not the
oracle).  Used by tests/, bench.py (both arms) and the profiling tools."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liblz4e_synth.so")
_lib = None


def build():
    out = subprocess.run(["make", "-C", HERE], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("building liblz4e_synth.so failed:\n" + out.stdout + out.stderr)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = ctypes.CDLL(LIB_PATH)
        vp, u32, u64 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64
        L.lz4e_synth_fill_host.restype = None
        L.lz4e_synth_fill_host.argtypes = [vp, u32, u32, u32]
        L.lz4e_synth_fill_host_blocks.restype = None
        L.lz4e_synth_fill_host_blocks.argtypes = [vp, u64, u32, u64, u64, ctypes.c_int]
        L.lz4e_synth_fill_dev.restype = ctypes.c_int
        L.lz4e_synth_fill_dev.argtypes = [ctypes.c_int, vp, vp, u32, u32, u32]
        _lib = L
    return _lib


def host(n, cls, seed):
    """One block of n bytes of class cls (0 text, 1 random, 2 sparse, 3 mixed)."""
    buf = np.zeros(max(int(n), 1), dtype=np.uint8)
    lib().lz4e_synth_fill_host(buf.ctypes.data, int(n), int(cls), int(seed))
    return buf[:n]


def host_blocks(mem, stride, n, nblocks, first=0, threads=None):
    """Fill nblocks blocks of n bytes at mem[b * stride:], class/seed from first + b."""
    assert mem.dtype == np.uint8 and mem.nbytes >= (nblocks - 1) * stride + n
    lib().lz4e_synth_fill_host_blocks(mem.ctypes.data, int(stride), int(n), int(nblocks),
                                      int(first), int(threads or os.cpu_count() or 1))


def dev(device, d_bvecs, d_iters, nblocks, first_index, max_bytes):
    """Fill device-resident SG blocks (device pointers to the descriptor tables)."""
    rc = lib().lz4e_synth_fill_dev(int(device), ctypes.c_void_p(int(d_bvecs)),
                                   ctypes.c_void_p(int(d_iters)), int(nblocks),
                                   int(first_index), int(max_bytes))
    if rc != 0:
        raise RuntimeError("lz4e_synth_fill_dev failed")
