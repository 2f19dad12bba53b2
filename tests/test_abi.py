"""CPU tests of the boundary: the C-ABI library loads, exports every function that
include/*.h declares, fails loudly without a device (no CPU fallback), and the
host-side sharding helper agrees across ranks (gloo, world_size 2)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INCLUDE = os.path.join(ROOT, "include")


def _declared():
    names = set()
    for h in ("lz4e.h", "lz4e_b200.h"):
        text = open(os.path.join(INCLUDE, h)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        text = re.sub(r"^\s*#.*$", "", text, flags=re.M)
        for m in re.finditer(r"\b(LZ4E_[A-Za-z0-9_]+|lz4e_b200_[a-z0-9_]+)\s*\(", text):
            names.add(m.group(1))
    return sorted(n for n in names if not n.isupper())


def test_library_exports_every_declared_symbol(pkg):
    lib = ctypes.CDLL(pkg.LIB_PATH)
    names = _declared()
    assert "LZ4E_compress_default" in names and "lz4e_b200_compress_dev" in names
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_sizing_macros_match_reference():
    """lz4e.h:11-28 of the reference: LZ4E_MEM_COMPRESS 17440, COMPRESSBOUND."""
    src = ('#include <stdio.h>\n#include "lz4e.h"\nint main(void){printf("%zu %d %d %d %u %u\\n",'
           '(size_t)LZ4E_MEM_COMPRESS, LZ4E_COMPRESSBOUND(4096), LZ4E_COMPRESSBOUND(65536),'
           'LZ4E_COMPRESSBOUND(0x7E000001), (unsigned)sizeof(struct bio_vec),'
           '(unsigned)sizeof(struct bvec_iter));return 0;}')
    exe = "/tmp/lz4e_macro_probe"
    subprocess.run(["gcc", "-x", "c", "-", "-I", INCLUDE, "-o", exe], input=src.encode(), check=True)
    out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()
    assert out == ["17440", "4128", "65809", "0", "16", "20"]


def test_stream_typedefs_match_reference():
    """lz4e.h:30-45 of the reference: LZ4E_stream_t / LZ4E_stream_t_internal exist, the union is
    LZ4E_MEM_COMPRESS bytes and the fields sit where the reference puts them."""
    src = ('#include <stdio.h>\n#include <stddef.h>\n#include "lz4e.h"\nint main(void){LZ4E_stream_t s;'
           'printf("%zu %zu %zu %zu %zu\\n", sizeof(s), sizeof(s.table)/sizeof(s.table[0]),'
           'offsetof(LZ4E_stream_t_internal, hashTable), offsetof(LZ4E_stream_t_internal, currentOffset),'
           'sizeof(s.internal_donotuse.hashTable));return 0;}')
    exe = "/tmp/lz4e_stream_probe"
    subprocess.run(["gcc", "-x", "c", "-", "-I", INCLUDE, "-o", exe], input=src.encode(), check=True)
    out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()
    assert out == ["17440", "2180", "1024", "17408", "16384"]


def test_no_cpu_fallback_without_device(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    assert pkg.lib().lz4e_b200_device_count() <= 0
    with pytest.raises(Exception):
        pkg.Context(0)
    # the reference-shaped entry point fails the reference's way: 0 = failure
    code = ("import importlib,ctypes,numpy as np;p=importlib.import_module('lz4-sgori_b200');"
            "l=p.lib();bv=np.zeros(1,dtype=p.BIO_VEC);it=np.zeros(1,dtype=p.BVEC_ITER);"
            "w=np.zeros(17440,dtype=np.uint8);vp=ctypes.c_void_p;"
            "l.LZ4E_compress_default.restype=ctypes.c_int;"
            "print(l.LZ4E_compress_default(vp(bv.ctypes.data),vp(bv.ctypes.data),"
            "vp(it.ctypes.data),vp(it.ctypes.data),vp(w.ctypes.data)))")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT)
    assert out.stdout.strip().splitlines()[-1] == "0", out.stdout + out.stderr


def test_shard_range_partitions(pkg):
    for n in (0, 1, 7, 262144, 262145, 1000003):
        for world in (1, 2, 3, 8):
            got = [pkg.shard_range(n, r, world) for r in range(world)]
            assert got[0][0] == 0 and sum(c for _, c in got) == n
            for (f0, c0), (f1, _) in zip(got, got[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in got) - min(c for _, c in got) <= 1


_GLOO_WORKER = r"""
import importlib, os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.environ["LZ4E_ROOT"])
pkg = importlib.import_module("lz4-sgori_b200")
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 1000
first, count = pkg.shard_range(n, rank, world)
# every rank "processes" its contiguous block range; host-side aggregation only (SURVEY 8e)
owned = torch.zeros(n, dtype=torch.int64)
owned[first:first + count] = 1
sizes = torch.tensor([float(count * 4096)])
dist.all_reduce(owned)
dist.all_reduce(sizes)
t = torch.tensor([1.0 + rank])
dist.all_reduce(t, op=dist.ReduceOp.MAX)
assert bool((owned == 1).all()), "ranges overlap or leave gaps"
assert sizes.item() == n * 4096 and t.item() == float(world)
if rank == 0:
    print("gloo-ok", world)
dist.destroy_process_group()
"""


def test_sharding_world2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    env = dict(os.environ, LZ4E_ROOT=ROOT, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                          "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                          "--master-port", "29571", str(script)],
                         capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "gloo-ok 2" in out.stdout
