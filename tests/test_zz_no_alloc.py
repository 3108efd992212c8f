"""vga_seed_count_noalloc (the allocation-free K1 counting entry point, include/vga.h) against the existing entry point:
anchors per read must equal VgaResultsView.n_anchors of vga_map_batch, bit for bit, on two calls in a row that share the
library's fixed scratch array.  Collected last on purpose.

Run on the B200 box: python -m pytest tests/test_zz_no_alloc.py -m gpu
"""
import ctypes as C

import numpy as np
import pytest

import vgaligner_b200 as vg
from conftest import fixture_path

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 64 << 20          # VGA_NOALLOC_SCRATCH_BYTES


def count_noalloc(L, gix, d_seqs, d_offs, n, n_bases, d_out):
    return L.vga_seed_count_noalloc(gix, C.c_void_p(d_seqs.data_ptr()), C.c_void_p(d_offs.data_ptr()), n, n_bases,
                                    C.c_void_p(d_out.data_ptr()))


def test_no_alloc_seed_count_matches_map_batch_twice():
    import torch
    L = vg.gpu_lib()
    ix = vg.Index.build(fixture_path("graphs/2-DRB1-3123.gfa"), 11)
    gix = ix.upload()
    dev = torch.device("cuda:0")
    # two different read sets: the second call must not see anything the first one left in the scratch array
    sets = [vg.simulate_reads(ix, 3000, 150, seed=5), vg.simulate_reads(ix, 1200, 400, 0.02, 0.002, 0.002, seed=9)]
    for reads in sets:
        want = vg.map_reads(ix, reads, params=vg.default_params(flags=vg.VGA_FLAG_NO_ALIGN)).n_anchors
        seqs = np.ascontiguousarray(reads.seqs)
        offs = np.ascontiguousarray(reads.offsets).view(np.int64)
        d_seqs = torch.from_numpy(seqs).to(dev)
        d_offs = torch.from_numpy(offs).to(dev)
        d_out = torch.full((reads.n,), -1, dtype=torch.int32, device=dev)        # the caller's output buffer
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc = count_noalloc(L, gix, d_seqs, d_offs, reads.n, int(offs[-1]), d_out)
        assert rc == 0, L.vga_last_error().decode()
        if reads is not sets[0]:          # the first call may load the kernels' module; after that the device's free memory stays put
            assert torch.cuda.mem_get_info()[0] >= free_before
        got = d_out.cpu().numpy().view(np.uint32)
        assert got.shape == want.shape
        assert np.array_equal(got, want), "anchors per read differ at %s" % np.flatnonzero(got != want)[:5]
        assert int(got.sum()) > 0
    # first set again after the second: same answer from the shared array a third time
    reads = sets[0]
    want = vg.map_reads(ix, reads, params=vg.default_params(flags=vg.VGA_FLAG_NO_ALIGN)).n_anchors
    d_seqs = torch.from_numpy(np.ascontiguousarray(reads.seqs)).to(dev)
    offs = np.ascontiguousarray(reads.offsets).view(np.int64)
    d_offs = torch.from_numpy(offs).to(dev)
    d_out = torch.zeros(reads.n, dtype=torch.int32, device=dev)
    assert count_noalloc(L, gix, d_seqs, d_offs, reads.n, int(offs[-1]), d_out) == 0, L.vga_last_error().decode()
    assert np.array_equal(d_out.cpu().numpy().view(np.uint32), want)


def test_no_alloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    L = vg.gpu_lib()
    ix = vg.Index.build(fixture_path("test.gfa"), 4)
    gix = ix.upload()
    dev = torch.device("cuda:0")
    d = torch.zeros(16, dtype=torch.int64, device=dev)
    # too many bases for 8 bytes each, then too many reads for 8 bytes each: refused before anything is launched
    for n, n_bases in ((1, SCRATCH_BYTES // 8 + 1), (SCRATCH_BYTES // 8, 0)):
        rc = count_noalloc(L, gix, d, d, n, n_bases, d)
        assert rc == -1, rc                                  # VGA_ERR_INVALID
        assert "fixed scratch" in L.vga_last_error().decode()
    # offsets that describe more k-mers than n_bases admits are caught after the scan, before any slot is written
    seqs = np.frombuffer(b"ACGTTGCAACGT" * 8, dtype=np.uint8).copy()
    offs = np.arange(9, dtype=np.int64) * 12
    offs[-1] = SCRATCH_BYTES                                 # a last read of 64 M bases (never dereferenced)
    d_offs = torch.from_numpy(offs).to(dev)
    d_seqs = torch.from_numpy(seqs).to(dev)
    d_out = torch.zeros(8, dtype=torch.int32, device=dev)
    rc = count_noalloc(L, gix, d_seqs, d_offs, 8, 96, d_out)
    assert rc == -1 and "more than n_bases allows" in L.vga_last_error().decode()
