"""Debug helper: repeated vga_map_batch_device calls as bench.py issues them."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import vgaligner_b200 as vg
import bench
wl = bench.WORKLOADS[sys.argv[1]]; n = int(sys.argv[2]); reps = int(sys.argv[3])
ix, _ = bench.build_index(vg, wl)
reads = vg.simulate_reads(ix, n, wl["length"], *wl["err"], 0.1, wl["seed"])
L = vg.gpu_lib(); gix = ix.upload()
d_seqs = torch.from_numpy(reads.seqs).cuda(); d_offs = torch.from_numpy(reads.offsets.view(np.int64)).cuda()
p = vg.default_params()
for i in range(reps):
    out = C.c_void_p()
    rc = L.vga_map_batch_device(gix, C.c_void_p(d_seqs.data_ptr()), C.c_void_p(d_offs.data_ptr()), reads.n, int(reads.offsets[-1]), C.byref(p), C.byref(out))
    print(i, rc, L.vga_last_error().decode() if rc else "ok", flush=True)
    if rc: break
    L.vga_results_free(out)
