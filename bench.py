#!/usr/bin/env python
"""bench.py — `vgaligner map --also-align` hot path on N B200s (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C2|C3a|C3b|C4|C5] [--impl reference]

A step = one pass of the whole hot path (seeding -> chaining -> subgraphs -> POA -> traceback ->
result compaction) over one batch of simulated reads.  `value` is reads/s with the reads already
resident in HBM (vga_map_batch_device, CUDA-event timed on the library's stream); `e2e` is the same
through the reference-facing C-ABI call vga_map_batch with pinned HOST buffers, host<->device
copies inside the timed region.  Reads shard across ranks with no data-path collective; the index
is replicated by ONE NCCL broadcast (weak scaling: fixed reads per GPU).

`--impl reference` times the CPU implementation of the same path (the oracle port, worker threads over
reads on all host cores; the Rust binary cannot be built in this image) on a bounded
sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GRAPHS = os.path.join(ROOT, "tests", "golden", "graphs")
# per-GPU workloads (BASELINE.json configs; reads per GPU are fixed -> weak scaling)
WORKLOADS = {
    "C2": dict(graph="2-DRB1-3123.gfa", k=11, reads=100000, length=150, err=(0.01, 0.0005, 0.0005), seed=77,
               desc="HLA-zoo DRB1-3123 graph, k=11, 100k simulated 150bp reads per GPU"),
    "C3a": dict(graph="17-DRB1-3123-smooth.gfa", k=11, reads=125000, length=150, err=(0.01, 0.0005, 0.0005), seed=77,
                desc="HLA-zoo 17-DRB1-3123-smooth, k=11, 1M simulated 150bp reads / 8 GPUs = 125k per GPU"),
    "C3b": dict(graph="19-MICB-4227-smooth.gfa", k=11, reads=125000, length=150, err=(0.01, 0.0005, 0.0005), seed=77,
                desc="HLA-zoo 19-MICB-4227-smooth, k=11, 1M simulated 150bp reads / 8 GPUs = 125k per GPU"),
    "C4": dict(graph=None, synth=5000000, k=15, reads=1250000, length=150, err=(0.01, 0.0005, 0.0005), seed=2,
               desc="synthetic 5 Mbp variation graph, k=15, 10M simulated 150bp reads / 8 GPUs = 1.25M per GPU"),
    "C5": dict(graph=None, synth=5000000, k=15, reads=12500, length=10000, err=(0.04, 0.03, 0.03), seed=3,
               desc="synthetic 5 Mbp variation graph, k=15, 100k simulated 10kb reads at 10% error / 8 GPUs = 12.5k per GPU"),
}


def build_index(vg, wl):
    t0 = time.time()
    if wl.get("graph"):
        ix = vg.Index.build(os.path.join(GRAPHS, wl["graph"]), wl["k"])
    else:
        cache = "/tmp/vga_synth_%d_k%d.vgidx" % (wl["synth"], wl["k"])
        if os.path.exists(cache):
            ix = vg.Index.load_from_file(cache)
        else:
            ix = vg.Index.build_from_text(vg.synth_gfa(wl["synth"], 100, 1), wl["k"])
            try:
                ix.store_in_file(cache)
            except Exception:
                pass
    return ix, time.time() - t0


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons DURING the timed region.  NVML in-process (cheap); falls back to
    spawning nvidia-smi when pynvml is not importable."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.rows, self.stop_flag = gpu, [], False
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # LOCAL_RANK indexes the visible devices; honour CUDA_VISIBLE_DEVICES when it lists ordinals
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = gpu
            if vis:
                parts = [x.strip() for x in vis.split(",") if x.strip()]
                if gpu < len(parts) and parts[gpu].isdigit():
                    idx = int(parts[gpu])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def sample(self):
        if self.nvml is not None:
            n = self.nvml
            sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
            mx = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
            pw = n.nvmlDeviceGetPowerUsage(self.handle) / 1000.0
            r = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
                else n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
            bits = [getattr(n, "nvmlClocksThrottleReasonHwSlowdown", 0x8), getattr(n, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                    getattr(n, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20), getattr(n, "nvmlClocksThrottleReasonSwPowerCap", 0x4)]
            return [str(sm), str(mx), "%.1f" % pw] + ["Active" if (r & b) else "Not Active" for b in bits]
        out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout.strip()
        return [x.strip() for x in out.split(",")] if out else None

    def run(self):
        while not self.stop_flag:
            try:
                row = self.sample()
                if row:
                    self.rows.append(row)
            except Exception:
                pass
            time.sleep(0.02 if self.nvml is not None else 0.2)

    def summary(self):
        self.stop_flag = True
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        reasons = sorted({self.NAMES[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        pw = [float(r[2]) for r in self.rows if len(r) > 2]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows), "power_w_max": max(pw) if pw else None,
                "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def cpu_run(vg, orc, ix, reads, n_sample, threads, steps):
    """Oracle port on the host cores over the first n_sample reads; returns reads/s (best of steps)."""
    import numpy as np

    class Sub:
        pass

    s = Sub()
    s.n = n_sample
    s.names = reads.names[:n_sample]
    s.offsets = reads.offsets[:n_sample + 1].copy()
    s.seqs = reads.seqs[:int(s.offsets[-1])]
    oi = orc.OracleIndex(ix)
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        r = oi.map_reads(s, orc.default_params(), n_threads=threads)
        times.append(time.perf_counter() - t0)
        aligned = int((r.aln_status == 1).sum())
        cells = int(r.aln_cells.sum())
        del r
    return n_sample / (sum(times) / len(times)), sum(times) / len(times), aligned, cells


def pick_cpu_sample(vg, orc, ix, reads, threads, budget_s):
    n0 = max(8, min(reads.n, 64 if reads.offsets[1] > 2000 else 2000))
    rate, _, _, _ = cpu_run(vg, orc, ix, reads, n0, threads, 1)
    return int(max(n0, min(reads.n, rate * budget_s)))


# stdout carries exactly ONE JSON line: file descriptor 1 is pointed at stderr for the whole run (NCCL's banner, any
# library chatter), and the line is written to the saved descriptor at the end.
_JSON_FD = None


def capture_stdout():
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    os.write(_JSON_FD if _JSON_FD is not None else 1, data)


def main():
    capture_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("VGA_WORKLOAD", "C2"), choices=sorted(WORKLOADS))
    ap.add_argument("--reads", type=int, default=0, help="override reads per GPU")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU work per baseline pass")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = dict(WORKLOADS[args.workload])
    if args.reads:
        wl["reads"] = args.reads
    threads = os.cpu_count() or 1

    import vgaligner_b200 as vg

    base = {"metric": "aligned reads/sec", "unit": "reads/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": args.workload + ": " + wl["desc"], "reads_per_gpu": wl["reads"], "read_length": wl["length"],
                       "k": wl["k"], "also_align": True, "params": "bandwidth=50 max_gap=1000 min_anchors=3 best_n=1 abPOA defaults",
                       "l2": "flushed between timed steps (256 MiB write); intermediates per step exceed L2"}}

    # ------------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return 0
        from oracle import binding as orc
        ix, _ = build_index(vg, wl)
        n_gen = min(wl["reads"], 200000 if wl["length"] <= 1000 else 2000)
        reads = vg.simulate_reads(ix, n_gen, wl["length"], *wl["err"], 0.1, wl["seed"])
        n_sample = pick_cpu_sample(vg, orc, ix, reads, threads, args.cpu_seconds)
        for _ in range(min(args.warmup, 1)):
            cpu_run(vg, orc, ix, reads, max(8, n_sample // 8), threads, 1)
        rate, sec, aligned, cells = cpu_run(vg, orc, ix, reads, n_sample, threads, args.steps)
        line = dict(base)
        line.update({"impl": "reference", "value": rate, "ms_per_step": sec * 1e3,
                     "cpu_baseline": {"value": rate, "unit": "reads/s", "cores": threads, "kind": "port",
                                      "sample": "%d reads of the workload per step, oracle port, threads over reads" % n_sample},
                     "e2e": {"value": rate, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                     "gpu_launches": 0, "aligned_fraction": aligned / n_sample, "cpu_gcups": cells / sec / 1e9})
        emit(line)
        return 0

    # ------------------------------------------------------------------ our arm (GPU)
    import numpy as np
    import torch
    import torch.distributed as dist
    import ctypes as C

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the map path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL prints its version banner on stdout (NCCL_DEBUG=VERSION/WARN/INFO); emit() keeps it off the JSON stream
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    L = vg.gpu_lib()
    rc = L.vga_init(local_rank)
    if rc != 0:
        raise SystemExit("vga_init: " + L.vga_last_error().decode())

    # index: rank 0 builds + uploads; ONE NCCL broadcast of the device image replicates it
    t_index = 0.0
    if rank == 0:
        ix, t_index = build_index(vg, wl)
        ix.upload()
        nbytes = ix.image_size()
    if world > 1:
        sz = torch.tensor([nbytes if rank == 0 else 0], dtype=torch.int64, device=dev)
        dist.broadcast(sz, 0)
        nbytes = int(sz.item())
        img = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        if rank == 0:
            ix.copy_image_to(img.data_ptr())
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        dist.broadcast(img, 0)
        torch.cuda.synchronize()
        t_bcast = time.perf_counter() - t0
        if rank != 0:
            ix, _ = (None, None)
            gix = C.c_void_p()
            rc = L.vga_index_import_device(C.c_void_p(img.data_ptr()), nbytes, C.byref(gix))
            if rc != 0:
                raise SystemExit("vga_index_import_device: " + L.vga_last_error().decode())
        del img
    else:
        t_bcast = 0.0
    if rank == 0:
        gix = ix.upload()

    # reads: every rank simulates its own shard (rank 0 has the host index; other ranks rebuild it only
    # for the simulator when the graph is a fixture, else receive the reads from rank 0)
    if rank == 0:
        shards = [vg.simulate_reads(ix, wl["reads"], wl["length"], *wl["err"], 0.1, wl["seed"] + 1000 * r) for r in range(world)]
        reads = shards[0]
    if world > 1:
        meta = torch.zeros(2, dtype=torch.int64, device=dev)
        for r in range(1, world):
            if rank == 0:
                s = shards[r]
                meta[0], meta[1] = s.n, int(s.offsets[-1])
                dist.send(meta, r)
                dist.send(torch.from_numpy(s.seqs).to(dev), r)
                dist.send(torch.from_numpy(s.offsets.astype(np.int64)).to(dev), r)
            elif rank == r:
                dist.recv(meta, 0)
                seqs_t = torch.empty(int(meta[1]), dtype=torch.uint8, device=dev)
                offs_t = torch.empty(int(meta[0]) + 1, dtype=torch.int64, device=dev)
                dist.recv(seqs_t, 0)
                dist.recv(offs_t, 0)

                class Shard:
                    pass
                reads = Shard()
                reads.n = int(meta[0])
                reads.seqs = seqs_t.cpu().numpy()
                reads.offsets = offs_t.cpu().numpy().astype(np.uint64)
        if rank == 0:
            del shards
    n_reads = reads.n
    n_bases = int(reads.offsets[-1])

    # pinned host copies (e2e leg) and device-resident copies (device leg)
    h_seqs = torch.from_numpy(np.ascontiguousarray(reads.seqs)).pin_memory()
    h_offs = torch.from_numpy(np.ascontiguousarray(reads.offsets).view(np.int64)).pin_memory()
    d_seqs = h_seqs.to(dev)
    d_offs = h_offs.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    params = vg.default_params()

    def step_device():
        out = C.c_void_p()
        rc = L.vga_map_batch_device(gix, C.c_void_p(d_seqs.data_ptr()), C.c_void_p(d_offs.data_ptr()), n_reads, n_bases,
                                    C.byref(params), C.byref(out))
        if rc != 0:
            raise SystemExit("vga_map_batch_device: " + L.vga_last_error().decode())
        return out

    def step_host():
        out = C.c_void_p()
        rc = L.vga_map_batch(gix, C.c_void_p(h_seqs.data_ptr()), C.c_void_p(h_offs.data_ptr()), n_reads, C.byref(params), C.byref(out))
        if rc != 0:
            raise SystemExit("vga_map_batch: " + L.vga_last_error().decode())
        return out

    def stats_of(out):
        st = vg.VgaStats()
        L.vga_results_stats(out, C.byref(st))
        return {f: getattr(st, f) for f, _ in vg.VgaStats._fields_}

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        acc, last = [], None
        for _ in range(steps):
            flush.fill_(1)
            torch.cuda.synchronize()
            out = fn()
            st = stats_of(out)
            acc.append(st)
            if last is not None:
                L.vga_results_free(last)
            last = out
        return acc, last

    # warm-up in the same pattern as the timed loop (the previous step's results stay alive while the next
    # step runs, so both sets of pinned result blocks exist before timing starts)
    _, wlast = timed(step_device, args.warmup)
    if wlast is not None:
        L.vga_results_free(wlast)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    t_wall0 = time.perf_counter()
    acc, last = timed(step_device, args.steps)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    ms_dev = sum(s["ms_total"] for s in acc)

    # aligned fraction from the last step's results
    view = vg.VgaResultsView()
    L.vga_results_view(last, C.byref(view))
    status = np.ctypeslib.as_array(view.aln_status, shape=(n_reads,)).copy()
    aligned = int((status == vg.ALN_OK).sum())
    L.vga_results_free(last)

    e2e = None
    if not args.no_e2e:
        _, wlast = timed(step_host, max(1, min(args.warmup, 2)))
        L.vga_results_free(wlast)
        barrier()
        acc_h, last_h = timed(step_host, args.steps)
        barrier()
        L.vga_results_free(last_h)
        ms_e2e = sum(s["ms_total"] for s in acc_h)
        e2e = (ms_e2e, acc_h[-1]["h2d_bytes"], acc_h[-1]["d2h_bytes"],
               {k: sum(x[k] for x in acc_h) / len(acc_h) for k in ("ms_h2d", "ms_seed", "ms_sort", "ms_chain", "ms_subgraph", "ms_poa", "ms_trace", "ms_d2h", "ms_total")})

    clocks = sampler.summary()      # sampled over both timed legs

    # max over ranks
    red = torch.tensor([ms_dev, e2e[0] if e2e else 0.0], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(n_reads), float(aligned)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_dev_max, ms_e2e_max = float(red[0]), float(red[1])
    total_reads, total_aligned = float(tot[0]), float(tot[1])

    if rank == 0:
        s = acc[-1]
        K = args.steps
        mean = lambda k: sum(x[k] for x in acc) / K
        int32_peak = vg.measure_int32_peak()
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak, hbm_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
        traffic = {}
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload, {})
        except Exception:
            pass
        poa_ms, poa_launches = mean("ms_poa"), max(1, s["poa_launches"])
        gcups = s["poa_cells"] / (poa_ms * 1e-3) / 1e9 if poa_ms > 0 else 0.0
        gcups_peak = int32_peak / 17.0 / 1e9
        roof_poa = {"kernel": "poa_fill_pair_kernel (short reads) / poa_fill_kernel<SLIDE> (long reads): all fill launches of a step", "bound": "int32_alu", "achieved": gcups, "peak": gcups_peak, "unit": "GCUPS",
                    "frac": gcups / gcups_peak if gcups_peak else None, "traffic": traffic.get("poa_fill_bytes_per_launch"),
                    "peak_source": "measured here: %.1f T int32 op/s (IADD+IMNMX streams) / 17 ops per cell" % (int32_peak / 1e12),
                    "launches_per_step": poa_launches, "ms_per_launch": poa_ms / poa_launches,
                    "cells_per_launch": s["poa_cells"] / poa_launches}
        seed_ms = mean("ms_seed")
        gbs = s["lookup_bytes"] / (seed_ms * 1e-3) / 1e9 if seed_ms > 0 else 0.0
        roof_seed = {"kernel": "seed_count_kernel+seed_emit_kernel (K1 stage)", "bound": "hbm", "achieved": gbs, "peak": hbm_peak,
                     "unit": "GB/s", "frac": gbs / hbm_peak, "traffic": traffic.get("seed_bytes_per_launch"),
                     "peak_source": hbm_src, "ms_per_launch": seed_ms, "bytes_per_launch": s["lookup_bytes"]}
        stage = {k: mean(k) for k in ("ms_h2d", "ms_seed", "ms_sort", "ms_chain", "ms_subgraph", "ms_poa", "ms_trace", "ms_d2h", "ms_total")}
        dominant = roof_poa if stage["ms_poa"] >= stage["ms_seed"] else roof_seed
        line = dict(base)
        line.update({"value": total_reads * K / (ms_dev_max * 1e-3), "ms_per_step": ms_dev_max / K,
                     "ms_steps": [round(x["ms_total"], 3) for x in acc], "clocks": clocks, "gpu_launches": int(sum(x["kernel_launches"] for x in acc)),
                     "aligned_fraction": total_aligned / total_reads, "stage_ms": stage,
                     "roofline": dominant, "roofline_poa": roof_poa, "roofline_lookup": roof_seed,
                     "setup": {"index_build_s": t_index, "index_image_bytes": nbytes if world > 1 else ix.image_size(),
                               "index_broadcast_s": t_bcast, "wall_s_timed_region": t_wall},
                     "counts": {k: s[k] for k in ("n_probes", "n_anchors", "n_chains", "n_alignments", "poa_cells", "lookup_bytes")}})
        if e2e:
            line["e2e"] = {"value": total_reads * K / (ms_e2e_max * 1e-3), "unit": "reads/s",
                           "h2d_bytes_per_step": int(e2e[1]), "d2h_bytes_per_step": int(e2e[2]), "ms_per_step": ms_e2e_max / K, "stage_ms": e2e[3],
                           "ms_steps": [round(x["ms_total"], 3) for x in acc_h]}
        if not args.no_cpu_baseline:
            from oracle import binding as orc       # CPU baseline leg: the one place bench.py executes oracle/
            n_sample = pick_cpu_sample(vg, orc, ix, reads, threads, args.cpu_seconds)
            rate, sec, _, cells = cpu_run(vg, orc, ix, reads, n_sample, threads, 1)
            line["cpu_baseline"] = {"value": rate, "unit": "reads/s", "cores": threads, "kind": "port",
                                    "sample": "first %d reads of rank 0's shard, oracle port with threads over reads, %.1f s" % (n_sample, sec),
                                    "cpu_gcups": cells / sec / 1e9}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
