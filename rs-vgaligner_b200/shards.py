"""Multi-GPU plumbing for the map path: one process per GPU, reads sharded, index replicated.

Reads are independent units (the reference's per-read closures share only the immutable `&Index`,
reference src/map.rs:56-111 and 162-167), so the path shards with NO data-path collective:

* the index image is ONE allocation, replicated by a single broadcast (NCCL over NVLink on the GPU
  box; the same code runs over gloo with host tensors in the CPU tests);
* rank r maps the contiguous read range [r*n/W, (r+1)*n/W) — output order must equal input order
  (src/map.rs:123-133), so rank-ordered concatenation of the per-rank GAF text is the whole output;
* the only other traffic is start-up scatter of the read shards from the rank that parsed the
  FASTA/FASTQ file and the final gather of GAF text to it.

torch.distributed is plumbing only; nothing here touches oracle/.
"""
import numpy as np


def shard_bounds(n_reads, world):
    """Contiguous read ranges: shard g is [g*n/W, (g+1)*n/W)."""
    return [(g * n_reads) // world for g in range(world + 1)]


class ReadShard:
    """A slice of a read batch (duck-types vgaligner_b200.Reads for map_reads)."""

    def __init__(self, seqs, offsets, names=None):
        self.seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        self.n = len(self.offsets) - 1
        self.names = names if names is not None else ["read%d" % i for i in range(self.n)]
        self._h = None

    def seq(self, i):
        return self.seqs[int(self.offsets[i]):int(self.offsets[i + 1])].tobytes().decode()


def shard_reads(reads, rank, world, with_names=True):
    """The reads rank `rank` of `world` maps, rebased to offset 0.  with_names=False leaves the (lazily built) name list of a
    simulated set of millions of reads alone."""
    b = shard_bounds(reads.n, world)
    lo, hi = b[rank], b[rank + 1]
    o = np.asarray(reads.offsets[lo:hi + 1], dtype=np.uint64)
    base = int(o[0]) if len(o) else 0
    seqs = np.asarray(reads.seqs[base:int(o[-1])] if len(o) else np.zeros(0, np.uint8))
    return ReadShard(seqs, o - np.uint64(base), list(reads.names[lo:hi]) if with_names and reads.names else None)


def scatter_reads(reads, rank, world, dist, device=None, src=0, with_names=True):
    """Rank `src` holds `reads` (others pass None); every rank returns its shard.

    The shard sizes travel by one small broadcast, the payloads (bytes, offsets) as ONE batch of point-to-point operations, so
    the sends to different ranks are in flight together; names stay on `src` (only the GAF writer needs them and it runs where
    the names are)."""
    import torch
    dev = device if device is not None else torch.device("cpu")
    if world == 1:
        return shard_reads(reads, 0, 1, with_names)
    meta = torch.zeros(world, 2, dtype=torch.int64, device=dev)
    parts = None
    if rank == src:
        parts = [shard_reads(reads, r, world, with_names) for r in range(world)]
        meta.copy_(torch.tensor([[s.n, int(s.offsets[-1]) if s.n else 0] for s in parts], dtype=torch.int64))
    dist.broadcast(meta, src)
    ops, keep = [], []
    if rank == src:
        for r in range(world):
            if r == src:
                continue
            s = parts[r]
            seqs_t = torch.from_numpy(s.seqs.copy()).to(dev) if s.seqs.size else torch.zeros(1, dtype=torch.uint8, device=dev)
            offs_t = torch.from_numpy(s.offsets.astype(np.int64)).to(dev)
            keep += [seqs_t, offs_t]
            ops += [dist.P2POp(dist.isend, seqs_t, r), dist.P2POp(dist.isend, offs_t, r)]
    else:
        n, nb = int(meta[rank, 0]), int(meta[rank, 1])
        seqs_t = torch.empty(max(nb, 1), dtype=torch.uint8, device=dev)
        offs_t = torch.empty(n + 1, dtype=torch.int64, device=dev)
        ops = [dist.P2POp(dist.irecv, seqs_t, src), dist.P2POp(dist.irecv, offs_t, src)]
    if ops:
        for q in dist.batch_isend_irecv(ops):
            q.wait()
    if rank == src:
        return parts[src]
    return ReadShard(seqs_t.cpu().numpy()[:nb], offs_t.cpu().numpy().astype(np.uint64))


def broadcast_image(image, rank, dist, device=None, src=0):
    """Replicate the index image (a uint8 tensor on `src`, None elsewhere) with ONE broadcast
    (plus an 8-byte size broadcast so the receivers can allocate)."""
    import torch
    dev = device if device is not None else torch.device("cpu")
    sz = torch.tensor([image.numel() if rank == src else 0], dtype=torch.int64, device=dev)
    dist.broadcast(sz, src)
    if rank != src:
        image = torch.empty(int(sz.item()), dtype=torch.uint8, device=dev)
    dist.broadcast(image, src)
    return image


def gather_text(text, rank, world, dist, dst=0):
    """Rank-ordered concatenation of per-rank GAF text on `dst` (None elsewhere)."""
    if world == 1:
        return text
    parts = [None] * world if rank == dst else None
    dist.gather_object(text, parts, dst=dst)
    return "".join(parts) if rank == dst else None


def gather_bytes(buf, rank, world, dist, device=None, dst=0):
    """Every rank contributes a uint8 numpy array (its packed result arrays); `dst` returns all of them in rank order,
    the other ranks None.  Sizes travel by one all_gather, payloads as ONE batch of point-to-point operations."""
    import torch
    if world == 1:
        return [buf]
    dev = device if device is not None else torch.device("cpu")
    sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([buf.size], dtype=torch.int64, device=dev))
    if rank != dst:
        t = torch.from_numpy(buf).to(dev) if buf.size else torch.zeros(1, dtype=torch.uint8, device=dev)
        for q in dist.batch_isend_irecv([dist.P2POp(dist.isend, t, dst)]):
            q.wait()
        return None
    bufs = {r: torch.empty(max(int(sizes[r].item()), 1), dtype=torch.uint8, device=dev) for r in range(world) if r != dst}
    for q in dist.batch_isend_irecv([dist.P2POp(dist.irecv, bufs[r], r) for r in bufs]):
        q.wait()
    return [buf if r == dst else bufs[r].cpu().numpy()[:int(sizes[r].item())] for r in range(world)]
