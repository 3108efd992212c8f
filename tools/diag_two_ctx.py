"""How much of a step is recoverable by keeping TWO calls in flight on one GPU?  Two host threads, two uploads of the same index
(a workspace each), each thread maps half of the reads with vga_map_batch_device; compared with one thread mapping all of them.
usage: python tools/diag_two_ctx.py [C4|C5|C2|C3a] [reads] [batch]"""
import ctypes as C
import os
import sys
import threading
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import vgaligner_b200 as vg  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "C4"
wl = dict(bench.WORKLOADS[name])
n_total = int(sys.argv[2]) if len(sys.argv) > 2 else 2000000
batch = int(sys.argv[3]) if len(sys.argv) > 3 else 500000
L = vg.gpu_lib()
ix, _ = bench.build_index(vg, wl)
reads = vg.simulate_reads_mt(ix, n_total, wl["length"], *wl["err"], 0.1, wl["seed"])
dev = torch.device("cuda:0")
d_seqs = torch.from_numpy(np.ascontiguousarray(reads.seqs)).to(dev)
offs = np.ascontiguousarray(reads.offsets).view(np.int64)
d_offs = torch.from_numpy(offs).to(dev)
n_bases = int(offs[-1])
params = vg.default_params()


def upload():
    out = C.c_void_p()
    assert L.vga_index_upload_ex(C.byref(ix.view), 0, C.byref(out)) == 0
    return out


def run(gix, lo, hi):
    n_aln = 0
    for b in range(lo, hi, batch):
        e = min(hi, b + batch)
        out = C.c_void_p()
        rc = L.vga_map_batch_device(gix, C.c_void_p(d_seqs.data_ptr()), C.c_void_p(d_offs.data_ptr() + 8 * b), e - b, n_bases,
                                    C.byref(params), C.byref(out))
        assert rc == 0, L.vga_last_error().decode()
        st = vg.VgaStats()
        L.vga_results_stats(out, C.byref(st))
        n_aln += st.n_alignments
        L.vga_results_free(out)
    return n_aln


g = [upload(), upload()]
for mode in ("one", "two", "one", "two"):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    if mode == "one":
        run(g[0], 0, n_total)
    else:
        th = [threading.Thread(target=run, args=(g[i], n_total * i // 2, n_total * (i + 1) // 2)) for i in range(2)]
        for t in th:
            t.start()
        for t in th:
            t.join()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("%s %s: %d reads in calls of %d, %s in flight: %.1f ms, %.2f M reads/s" % (name, mode, n_total, batch, mode, dt * 1e3, n_total / dt / 1e6),
          flush=True)
