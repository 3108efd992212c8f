"""World-size-2 test of the multi-process plumbing (rs-vgaligner_b200/shards.py) over gloo on CPU.

The N>1 path has no data-path collective: rank 0 parses/simulates the reads and scatters contiguous
shards, the index image is replicated by ONE broadcast, every rank maps its shard, and rank-ordered
concatenation of the GAF text must equal the single-process output (output order = input order,
reference src/map.rs:123-133).  There is no GPU here, so the per-rank mapper in this test is the
oracle (the checker standing in for vga_map_batch); on the GPU box bench.py drives the same helpers
with NCCL and the CUDA library.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r)
import numpy as np
import torch
import torch.distributed as dist
import vgaligner_b200 as vg
from oracle import binding as orc
import importlib
shards = sys.modules["rs_vgaligner_b200"].shards if hasattr(sys.modules["rs_vgaligner_b200"], "shards") else importlib.import_module("rs_vgaligner_b200.shards")

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
gfa = os.path.join(%(root)r, "tests", "golden", "graphs", "19-MICB-4227-smooth.gfa")
ix = vg.Index.build(gfa, 11)                 # every rank can build the host index (the oracle needs it)
reads = vg.simulate_reads(ix, 101, 150, 0.01, 0.0005, 0.0005, 0.1, 5) if rank == 0 else None

# one broadcast replicates the "image" (here: a byte pattern standing for the device blob)
img = torch.arange(0, 100003, dtype=torch.int64).to(torch.uint8) if rank == 0 else None
img = shards.broadcast_image(img, rank, dist)
assert img.numel() == 100003 and int(img[777]) == 777 %% 256 and int(img[-1]) == 100002 %% 256

mine = shards.scatter_reads(reads, rank, world, dist)
again = shards.scatter_reads(reads, rank, world, dist, with_names=False)
assert again.n == mine.n and np.array_equal(again.seqs, mine.seqs) and np.array_equal(again.offsets, mine.offsets)
# packed result bytes of every rank arrive on rank 0 in rank order (sizes differ, one is empty)
mineb = np.arange(1000 * rank + 7 * (1 - rank), dtype=np.int64).astype(np.uint8)
got = shards.gather_bytes(mineb, rank, world, dist)
if rank == 0:
    assert [g.size for g in got] == [7, 1000] and np.array_equal(got[1], np.arange(1000).astype(np.uint8)) and np.array_equal(got[0], mineb)
else:
    assert got is None
assert shards.gather_bytes(np.zeros(0, np.uint8), rank, world, dist) in (None,) or rank == 0
b = shards.shard_bounds(101, world)
assert mine.n == b[rank + 1] - b[rank], (mine.n, b)
oi = orc.OracleIndex(ix)
res = oi.map_reads(mine)
# names stay on rank 0: the shard carries positional names, fix them up to the global ones for the text
mine.names = ["read%%d" %% i for i in range(b[rank], b[rank + 1])]
text = res.gaf(oi, mine, 1) + "\x00" + res.gaf(oi, mine, 0)
aln = shards.gather_text(text.split("\x00")[0], rank, world, dist)
chn = shards.gather_text(text.split("\x00")[1], rank, world, dist)
if rank == 0:
    reads.names = ["read%%d" %% i for i in range(reads.n)]
    whole = oi.map_reads(reads)
    assert aln == whole.gaf(oi, reads, 1), "alignments GAF differs from the single-process run"
    assert chn == whole.gaf(oi, reads, 0), "chains GAF differs from the single-process run"
    assert aln.count("\n") == 101
    print("GLOO_OK")
dist.barrier()
dist.destroy_process_group()
'''


def test_shard_bounds_cover_everything():
    sys.path.insert(0, ROOT)
    import vgaligner_b200  # noqa: F401  (registers rs_vgaligner_b200)
    import importlib
    sh = importlib.import_module("rs_vgaligner_b200.shards")
    for n in (0, 1, 7, 100, 101):
        for w in (1, 2, 3, 8):
            b = sh.shard_bounds(n, w)
            assert b[0] == 0 and b[-1] == n and len(b) == w + 1
            assert all(b[i] <= b[i + 1] for i in range(w))
            assert max(b[i + 1] - b[i] for i in range(w)) - min(b[i + 1] - b[i] for i in range(w)) <= 1


def test_shard_reads_rebases_offsets():
    sys.path.insert(0, ROOT)
    import vgaligner_b200  # noqa: F401
    import importlib
    sh = importlib.import_module("rs_vgaligner_b200.shards")
    whole = sh.ReadShard(np.frombuffer(b"AAACCCCGGT", dtype=np.uint8), np.array([0, 3, 7, 9, 10], dtype=np.uint64))
    parts = [sh.shard_reads(whole, r, 3) for r in range(3)]
    assert [p.n for p in parts] == [1, 1, 2]
    assert [p.seq(0) for p in parts] == ["AAA", "CCCC", "GG"] and parts[2].seq(1) == "T"
    assert all(int(p.offsets[0]) == 0 for p in parts)


@pytest.mark.timeout(300)
def test_two_ranks_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29541", WORLD_SIZE="2", OMP_NUM_THREADS="1")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            p.kill()
            out, _ = p.communicate()
        outs.append(out)
    assert all(p.returncode == 0 for p in procs), "\n----\n".join(outs)
    assert "GLOO_OK" in outs[0]
