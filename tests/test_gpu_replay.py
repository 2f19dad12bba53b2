"""SURVEY.md rows f1/f2 and BASELINE.json configs[4]: the userspace chunk layer
(lz4-sgori_b200/host/lz4e_chunk.c, mirroring lz4e_bdev/lz4e_chunk.c:19-159) and the request
path (lz4e_bdev/lz4e_req.c:53-75, :144-213) driven by the C replay harness.  The harness checks
every read against a shadow device and the statistics counters itself; with --oracle it also
compresses every write bio with the reference's own C code and compares the total compressed
size with what the GPU path produced."""
import json
import os
import subprocess

import pytest

import common  # noqa: F401  (puts the repo root on sys.path)
from oracle import oracle

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tools", "replay", "lz4e_replay")


def _run(*args):
    if not os.path.exists(EXE):
        subprocess.run(["make", "-C", os.path.dirname(EXE)], check=True, capture_output=True)
    so = oracle.REF_SO if oracle.have_ref() else oracle.PORT_SO
    if not os.path.exists(so):
        oracle.build()
    out = subprocess.run([EXE, "--oracle", so, *args], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    return json.loads(out.stdout.strip().splitlines()[-1])


@pytest.mark.parametrize("bios", [(), ("--arena-bios",)])
def test_replay_4k_to_128k(bios):
    """fio-like QD64 random 4-128 KiB writes/reads; bios above 64 KiB take the 32-bit table."""
    r = _run("--qd", "64", "--batches", "6", "--capacity-mib", "64", "--seed", "3", *bios)
    assert r["failed"] == 0 and r["mismatches"] == 0
    assert r["compressed_bytes"] == r["oracle_compressed_bytes"] > 0
    assert r["per_bio_digest_equal"] is True
    assert r["per_bio_digest_equal"] is True  # the BYTES of every bio, not just the total
    assert r["gpu_launches"] >= 12  # a compress and a decompress launch per write batch
    assert r["read_reqs"] > 0 and r["write_reqs"] > 0


def test_replay_chunk_size_only():
    """the bdev's own chunk size: 4 KiB bios only, overlapping rewrites of a small device"""
    r = _run("--qd", "256", "--batches", "4", "--min-kib", "4", "--max-kib", "4",
             "--capacity-mib", "1", "--read-pct", "50", "--seed", "9")
    assert r["failed"] == 0 and r["mismatches"] == 0
    assert r["compressed_bytes"] == r["oracle_compressed_bytes"] > 0
    assert r["per_bio_digest_equal"] is True


def test_replay_two_submitters_one_process():
    """One submitter thread per GPU in ONE process (two contexts, two chunk pools, thread-bound
    with lz4e_chunk_bind; on a one-GPU box both share the device): same checks, counters add
    up."""
    r = _run("--qd", "64", "--batches", "4", "--capacity-mib", "64", "--seed", "5", "--devices", "2",
             "--arena-bios")
    assert r["submitters"] == 2
    assert r["failed"] == 0 and r["mismatches"] == 0
    assert r["compressed_bytes"] == r["oracle_compressed_bytes"] > 0
    assert r["per_bio_digest_equal"] is True
