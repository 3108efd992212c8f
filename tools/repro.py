"""Debug helper: map N simulated reads of a workload on the GPU and print the stage stats."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgaligner_b200 as vg
import bench

wl = bench.WORKLOADS[sys.argv[1]]
n = int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
ix, _ = bench.build_index(vg, wl)
reads = vg.simulate_reads(ix, n, wl["length"], *wl["err"], 0.1, wl["seed"])
for _ in range(reps):
    r = vg.map_reads(ix, reads)
    print(n, "ok", int((r.aln_status == 1).sum()), json.dumps({k: (round(v, 2) if isinstance(v, float) else v) for k, v in r.stats.items()}))
    r.free()
