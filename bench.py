#!/usr/bin/env python
"""bench.py — `vgaligner map --also-align` hot path on N B200s (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C4+C5|C2|C3a|C3b|C4|C5] [--impl reference]

BASELINE.json's metric is "aligned reads/sec (150bp & 10kb) at 1/2/4/8 B200", quoted on C4 (synthetic 5 Mbp graph, k=15,
10 M x 150 bp) and C5 (same graph, 100 k x 10 kb at 10 % error).  The default run measures BOTH: the JSON line is the C4
record (`value`, `e2e`, `roofline`, `cpu_baseline`, `parity_sample`) and carries the C5 record under `long_reads`.
The TOTAL read count of a workload is fixed and sharded over the ranks (strong scaling: 10 M / N and 100 k / N reads per GPU).

A step = one pass of the whole hot path (seeding -> chaining -> subgraphs -> POA -> traceback -> result compaction) over the
rank's shard, in batches of --batch-reads reads per C-ABI call (the library itself cuts at 2 M reads / 320 MB of bases).
  value   reads/s with the shard resident in HBM (vga_map_batch_device), CUDA-event time of the calls on the library's
          stream, max over ranks
  e2e     the same through the reference-facing C-ABI call vga_map_batch with HOST buffers, timed by the CALLER's clock
          (perf_counter around the loop of calls, max over ranks): pinned and pageable caller buffers
  e2e_gaf call + both GAF texts (what reference src/map.rs:123-184 produces), on one batch
Reads shard across ranks with no data-path collective; the index is replicated by ONE NCCL broadcast.

`--impl reference` times the CPU implementation of the same path (the oracle port, worker threads over reads on all host
cores; the Rust binary cannot be built in this image) on a bounded sample of the same workloads.
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
# BASELINE.json configs; `reads` is the TOTAL over all GPUs
WORKLOADS = {
    "C2": dict(graph="2-DRB1-3123.gfa", k=11, reads=100000, length=150, err=(0.01, 0.0005, 0.0005), seed=77, batch=100000,
               desc="HLA-zoo DRB1-3123 graph, k=11, 100k simulated 150bp reads"),
    "C3a": dict(graph="17-DRB1-3123-smooth.gfa", k=11, reads=1000000, length=150, err=(0.01, 0.0005, 0.0005), seed=77, batch=1000000,
                desc="HLA-zoo 17-DRB1-3123-smooth, k=11, 1M simulated 150bp reads"),
    "C3b": dict(graph="19-MICB-4227-smooth.gfa", k=11, reads=1000000, length=150, err=(0.01, 0.0005, 0.0005), seed=77, batch=250000,
                desc="HLA-zoo 19-MICB-4227-smooth, k=11, 1M simulated 150bp reads"),
    "C4": dict(graph=None, synth=5000000, k=15, reads=10000000, length=150, err=(0.01, 0.0005, 0.0005), seed=2, batch=1000000,
               desc="synthetic 5 Mbp variation graph (~1% SNP/indel bubbles), k=15, 10M simulated 150bp reads"),
    "C5": dict(graph=None, synth=5000000, k=15, reads=100000, length=10000, err=(0.04, 0.03, 0.03), seed=3, batch=12500,
               desc="synthetic 5 Mbp variation graph, k=15, 100k simulated 10kb reads at 10% error"),
}
PARITY_FIELDS = ("n_anchors", "chain_off", "chain_anchor_off", "ca_qb", "ca_tb", "ca_te", "aln_status", "aln_chain", "aln_score",
                 "aln_cells", "aln_path_len", "aln_start_off", "aln_end_off", "aln_n_aligned", "aln_path_off", "aln_path",
                 "aln_run_off", "aln_runs")


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def build_index(vg, wl, rank=0, world=1, barrier=None):
    """rank 0 builds (or loads a cached) host index; the other ranks of the box load the file it wrote."""
    t0 = time.time()
    if wl.get("graph"):
        return vg.Index.build(os.path.join(GRAPHS, wl["graph"]), wl["k"]), time.time() - t0
    cache = "/tmp/vga_synth_%d_k%d.vgidx" % (wl["synth"], wl["k"])
    if rank == 0 and not os.path.exists(cache):
        ix = vg.Index.build_from_text(vg.synth_gfa(wl["synth"], 100, 1), wl["k"])
        try:
            ix.store_in_file(cache + ".tmp")
            os.replace(cache + ".tmp", cache)
        except Exception:
            pass
        t = time.time() - t0
        if barrier:
            barrier()
        return ix, t
    if barrier:
        barrier()
    return vg.Index.load_from_file(cache), time.time() - t0


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons DURING the timed region.  NVML in-process (cheap); falls back to
    spawning nvidia-smi when pynvml is not importable."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, gpu, n_gpus=1):
        """gpu: this rank's local ordinal; n_gpus > 1 (rank 0 of a multi-rank run): sample ALL local GPUs round-robin from this one
        thread — eight ranks each polling NVML at 50 Hz contend with every rank's CUDA calls for the driver's locks."""
        super().__init__(daemon=True)
        self.gpu, self.rows, self.stop_flag = gpu, [], False
        self.nvml, self.handles, self.turn, self.max_mhz = None, [], 0, {}
        try:
            import pynvml
            pynvml.nvmlInit()
            # local ordinals index the visible devices; honour CUDA_VISIBLE_DEVICES when it lists ordinals
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            parts = [x.strip() for x in vis.split(",") if x.strip()] if vis else []
            for g in (range(n_gpus) if n_gpus > 1 else [gpu]):
                idx = int(parts[g]) if g < len(parts) and parts[g].isdigit() else g
                self.handles.append(pynvml.nvmlDeviceGetHandleByIndex(idx))
            self.handle = self.handles[0]
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def sample(self):
        if self.nvml is not None:
            n = self.nvml
            self.turn = (self.turn + 1) % len(self.handles)
            self.handle = self.handles[self.turn]
            sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
            if self.turn not in self.max_mhz:
                self.max_mhz[self.turn] = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
            mx = self.max_mhz[self.turn]
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
                "samples": len(self.rows), "power_w_max": max(pw) if pw else None, "gpus_sampled": max(1, len(self.handles)),
                "sm_min_mhz": sm[0] if sm else None, "source": "nvml" if self.nvml is not None else "nvidia-smi"}


class Sub:
    """the first n reads of a read set (duck-types vgaligner_b200.Reads)"""

    def __init__(self, reads, n):
        self.n = n
        self.offsets = reads.offsets[:n + 1].copy()
        self.seqs = reads.seqs[:int(self.offsets[-1])]


def cpu_run(orc, oi, reads, n_sample, threads, steps=1):
    """Oracle port on the host cores over the first n_sample reads -> (reads/s, seconds, aligned, cells, results)."""
    s = Sub(reads, n_sample)
    times, r = [], None
    for _ in range(steps):
        t0 = time.perf_counter()
        r = oi.map_reads(s, orc.default_params(), n_threads=threads)
        times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    return n_sample / sec, sec, int((r.aln_status == 1).sum()), int(r.aln_cells.sum()), r


def pick_cpu_sample(orc, oi, reads, threads, budget_s):
    n0 = max(8, min(reads.n, 64 if reads.offsets[1] > 2000 else 2000))
    rate, _, _, _, _ = cpu_run(orc, oi, reads, n0, threads)
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


def config_of(name, wl, world):
    return {"workload": name + ": " + wl["desc"], "total_reads": wl["reads"], "reads_per_gpu": wl["reads"] // world,
            "batch_reads_per_call": wl["batch"], "read_length": wl["length"], "k": wl["k"], "also_align": True,
            "params": "bandwidth=50 max_gap=1000 min_anchors=3 best_n=1 abPOA defaults",
            "l2": "inputs and intermediates of a step exceed L2 many times over (%.1f GB of read bytes per GPU and step); no flush needed"
                  % (wl["reads"] // world * wl["length"] / 1e9) if wl["reads"] // world * wl["length"] > (512 << 20) else
                  "flushed between timed steps (256 MiB write)"}


# ---------------------------------------------------------------------------------------------------------------
# reference arm (CPU)
# ---------------------------------------------------------------------------------------------------------------
def reference_record(vg, orc, name, wl, threads, args):
    ix, _ = build_index(vg, wl)
    n_gen = min(wl["reads"], 400000 if wl["length"] <= 1000 else 2000)
    reads = vg.simulate_reads_mt(ix, n_gen, wl["length"], *wl["err"], 0.1, wl["seed"])
    oi = orc.OracleIndex(ix)
    n_sample = pick_cpu_sample(orc, oi, reads, threads, args.cpu_seconds)
    for _ in range(min(args.warmup, 1)):
        cpu_run(orc, oi, reads, max(8, n_sample // 8), threads)
    rate, sec, aligned, cells, _ = cpu_run(orc, oi, reads, n_sample, threads, max(1, min(args.steps, 3)))
    return {"value": rate, "ms_per_step": sec * 1e3, "config": config_of(name, wl, 1),
            "cpu_baseline": {"value": rate, "unit": "reads/s", "cores": threads, "kind": "port",
                             "sample": "first %d reads of the workload per step, oracle port, threads over reads" % n_sample},
            "e2e": {"value": rate, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "aligned_fraction": aligned / n_sample, "cpu_gcups": cells / sec / 1e9}


# ---------------------------------------------------------------------------------------------------------------
# our arm (GPU): one workload
# ---------------------------------------------------------------------------------------------------------------
class Gpu:
    pass


def gpu_record(G, name, wl, args):
    """Measures one workload on this rank's shard; returns the record on rank 0 (None elsewhere)."""
    import ctypes as C

    import numpy as np
    import torch
    import torch.distributed as dist
    vg, L, dev, rank, world = G.vg, G.L, G.dev, G.rank, G.world

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- index: rank 0 builds + uploads; ONE NCCL broadcast of the device image replicates it
    ix, t_index = build_index(vg, wl, rank, world, barrier if world > 1 else None)
    t_bcast, nbytes = 0.0, 0
    gix = ix.upload() if rank == 0 else None
    if world > 1:
        # the library replicates the image itself (vga_index_broadcast: ONE ncclBroadcast over NVLink, NCCL bound at run time);
        # torch.distributed only carries the 128-byte communicator id to the other ranks
        cid = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            cid.copy_(torch.frombuffer(bytearray(vg.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(cid, 0)
        id_bytes = bytes(cid.cpu().numpy().tobytes())
        barrier()
        t0 = time.perf_counter()
        gix = vg.index_broadcast(gix, id_bytes, rank, world)
        torch.cuda.synchronize()
        t_bcast = time.perf_counter() - t0
        ix._gpu = gix
    nbytes = ix.image_size()
    # GPU index build of the same graph (vga_index_build_device; the tests compare its arrays with the host build's)
    t_build_dev = None
    if rank == 0 and not args.no_index_build:
        try:
            t0 = time.perf_counter()
            ms_dev = vg.index_build_device(ix)[4]
            t_build_dev = {"call_s": time.perf_counter() - t0, "device_ms": ms_dev}
        except vg.VgaError as e:
            t_build_dev = {"unsupported": str(e)}

    # ---- reads: contiguous shard [rank*n/world, (rank+1)*n/world) of ONE seeded read set (one RNG stream per read)
    total = wl["reads"]
    lo, hi = total * rank // world, total * (rank + 1) // world
    t0 = time.time()
    reads = vg.simulate_reads_mt(ix, hi - lo, wl["length"], *wl["err"], 0.1, wl["seed"], first_id=lo)
    t_sim = time.time() - t0
    n_reads, n_bases = reads.n, int(reads.offsets[-1])
    batch = max(1, min(args.batch_reads or wl["batch"], n_reads))
    n_calls = (n_reads + batch - 1) // batch            # equal calls: a short last call would run the device half empty
    cuts = [n_reads * i // n_calls for i in range(n_calls + 1)]
    log("%s rank %d: %d reads [%d, %d), %d calls per step, index %.1f s, reads %.1f s" % (name, rank, n_reads, lo, hi, len(cuts) - 1, t_index, t_sim))

    h_seqs_pageable = np.ascontiguousarray(reads.seqs)
    h_offs_np = np.ascontiguousarray(reads.offsets)
    h_seqs = torch.from_numpy(h_seqs_pageable).pin_memory()
    h_offs = torch.from_numpy(h_offs_np.view(np.int64)).pin_memory()
    d_seqs = h_seqs.to(dev)
    d_offs = h_offs.to(dev)
    small = n_bases <= (512 << 20)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if small else None
    params = vg.default_params()

    def check(rc, what):
        if rc != 0:
            raise SystemExit("%s: %s" % (what, L.vga_last_error().decode()))

    def call_device(b, e):
        out = C.c_void_p()
        check(L.vga_map_batch_device(gix, C.c_void_p(d_seqs.data_ptr()), C.c_void_p(d_offs.data_ptr() + 8 * b), e - b, n_bases,
                                     C.byref(params), C.byref(out)), "vga_map_batch_device")
        return out

    def call_host(seq_ptr, off_ptr):
        def f(b, e):
            out = C.c_void_p()
            check(L.vga_map_batch(gix, C.c_void_p(seq_ptr), C.c_void_p(off_ptr + 8 * b), e - b, C.byref(params), C.byref(out)), "vga_map_batch")
            return out
        return f

    def stats_of(out):
        st = vg.VgaStats()
        L.vga_results_stats(out, C.byref(st))
        return {f: getattr(st, f) for f, _ in vg.VgaStats._fields_}

    def one_step(fn, keep_first=None):
        """one pass over the shard; the previous call's results stay alive while the next call runs (as a caller that
        formats batch i while batch i+1 is mapped would hold them)"""
        acc, last = [], None
        t0 = time.perf_counter()
        for i in range(len(cuts) - 1):
            out = fn(cuts[i], cuts[i + 1])
            acc.append(stats_of(out))
            if last is not None:
                L.vga_results_free(last)
            last = out
        wall = time.perf_counter() - t0
        if last is not None:
            L.vga_results_free(last)
        return acc, wall

    def leg(fn, steps, warmup):
        for _ in range(warmup):
            one_step(fn)
        barrier()
        per_step, walls = [], []
        for _ in range(steps):
            if flush is not None:
                flush.fill_(1)
                torch.cuda.synchronize()
            acc, wall = one_step(fn)
            per_step.append(acc)
            walls.append(wall)
        barrier()
        return per_step, walls

    # ---- device-resident leg (value)
    sampler = ClockSampler(G.local_rank, world) if rank == 0 and not os.environ.get("VGA_BENCH_NO_SAMPLER") else None     # rank 0 watches every GPU of the box
    if sampler:
        sampler.start()
    t_wall0 = time.perf_counter()
    dev_steps, dev_walls = leg(call_device, args.steps, args.warmup)
    t_wall = time.perf_counter() - t_wall0
    ms_dev = sum(s["ms_total"] for acc in dev_steps for s in acc)

    # ---- e2e legs: host buffers, caller's clock
    e2e = None
    if not args.no_e2e:
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        pin_steps, pin_walls = leg(call_host(h_seqs.data_ptr(), h_offs.data_ptr()), e2e_steps, max(1, min(args.warmup, 2)))
        pg_steps, pg_walls = leg(call_host(h_seqs_pageable.ctypes.data, h_offs_np.ctypes.data), e2e_steps, 1)
        e2e = (sum(pin_walls), sum(pg_walls), e2e_steps, pin_steps)
    clocks = sampler.summary() if sampler else None

    # ---- one batch end to end INCLUDING both GAF texts (bounded: one batch)
    gaf = None
    if not args.no_gaf:
        nb = min(cuts[1], args.gaf_reads if wl["length"] <= 1000 else max(1, args.gaf_reads // 64))
        f = call_host(h_seqs.data_ptr(), h_offs.data_ptr())
        L.vga_results_free(f(0, nb))
        t0 = time.perf_counter()
        out = f(0, nb)
        t_map = time.perf_counter() - t0
        view = vg.VgaResultsView()
        L.vga_results_view(out, C.byref(view))
        t0 = time.perf_counter()
        nbytes_gaf, failed = 0, 0
        for which in (0, 1):
            n_, nf_ = C.c_uint64(), C.c_uint64()
            p = vg.host_lib().vgah_gaf_raw(ix._h, C.byref(view), None, lo, C.c_void_p(h_seqs.data_ptr()), C.c_void_p(h_offs.data_ptr()),
                                           which, 0, C.byref(n_), C.byref(nf_))
            if not p:
                raise SystemExit("vgah_gaf_raw: " + vg.host_lib().vgah_last_error().decode())
            vg.host_lib().vgah_free_str(p)
            nbytes_gaf += n_.value
            failed += nf_.value
        t_gaf = time.perf_counter() - t0
        L.vga_results_free(out)
        gaf = {"value": nb / (t_map + t_gaf), "unit": "reads/s", "reads": nb, "map_s": t_map, "gaf_s": t_gaf, "gaf_bytes": nbytes_gaf,
               "gaf_mb_per_s": nbytes_gaf / t_gaf / 1e6, "host_threads": os.cpu_count(), "failed_alignments": failed,
               "what": "vga_map_batch + chains GAF + alignments GAF text (parallel host formatter) for one batch, caller's clock, rank 0"}

    # ---- aligned fraction + parity sample (rank 0): oracle on the first reads of the shard vs the CUDA path on the same reads
    status_ok = 0
    view = vg.VgaResultsView()
    out = call_device(0, cuts[1])
    L.vga_results_view(out, C.byref(view))
    status_ok = int((np.ctypeslib.as_array(view.aln_status, shape=(cuts[1],)) == vg.ALN_OK).sum())
    frac_local = status_ok / cuts[1]
    L.vga_results_free(out)

    red = torch.tensor([ms_dev, e2e[0] if e2e else 0.0, e2e[1] if e2e else 0.0, t_sim], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(n_reads), frac_local * n_reads], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_dev_max, e2e_pin_max, e2e_pg_max = float(red[0]), float(red[1]), float(red[2])
    total_reads, total_aligned = float(tot[0]), float(tot[1])
    if rank != 0:
        ix.free_device()                                        # with its workspace: the next workload sizes its trace pool by the free memory
        if not args.no_sharded:
            G.keep.append((None, name, wl, ix, batch))          # this rank's part of the sharded leg (after every record is complete)
        return None

    K = args.steps
    last = dev_steps[-1]
    sumf = lambda acc, k: sum(x[k] for x in acc)
    mean = lambda k: sum(sumf(acc, k) for acc in dev_steps) / K           # per step (all calls of the step)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, hbm_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
    traffic = {}
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = dict(tj.get(name, {}))
        traffic["source"] = tj.get("source")
    except Exception:
        pass
    def per_unit(key):
        """DRAM bytes per algorithmic unit (cell / lookup byte) of the profiled launch (profiles/traffic.json, regenerated by
        tools/make_profiles.py from the ncu captures of the current build); scaled to this run's launches below"""
        b, u = traffic.get(key + "_bytes_per_launch"), traffic.get(key + "_units_in_launch")
        return b / u if b and u else None

    cells = sumf(last, "poa_cells")
    poa_ms, poa_launches = mean("ms_poa"), max(1, sumf(last, "poa_launches"))
    gcups = cells / (poa_ms * 1e-3) / 1e9 if poa_ms > 0 else 0.0
    gcups_peak = G.int32_peak / 17.0 / 1e9
    kern = ("poa_fill_pair_kernel (16x2 fill, short reads)" if wl["length"] < 246 else
            "poa_fill_long_kernel (fused fill + traceback per warp) + poa_fill_wide_kernel (bands beyond 1024 columns)")
    roof_poa = {"kernel": kern + ": all fill launches of a step", "bound": "int32_alu", "achieved": gcups, "peak": gcups_peak,
                "unit": "GCUPS", "frac": gcups / gcups_peak if gcups_peak else None,
                "traffic": per_unit("poa_fill") * (cells / poa_launches) if per_unit("poa_fill") else None, "traffic_source": traffic.get("source"),
                "traffic_note": traffic.get("poa_fill_note"),
                "peak_source": "measured here: %.1f T int32 op/s (IADD+IMNMX streams) / 17 ops per cell" % (G.int32_peak / 1e12),
                "launches_per_step": poa_launches, "ms_per_launch": poa_ms / poa_launches, "cells_per_launch": cells / poa_launches}
    seed_ms = mean("ms_seed")
    lbytes = sumf(last, "lookup_bytes")
    gbs = lbytes / (seed_ms * 1e-3) / 1e9 if seed_ms > 0 else 0.0
    roof_seed = {"kernel": "seed_count_kernel+seed_emit_kernel (K1 stage)", "bound": "hbm", "achieved": gbs, "peak": hbm_peak,
                 "unit": "GB/s", "frac": gbs / hbm_peak, "traffic": per_unit("seed") * lbytes if per_unit("seed") else None,
                 "traffic_source": traffic.get("source"), "traffic_note": traffic.get("seed_note"), "peak_source": hbm_src, "ms_per_step": seed_ms, "bytes_per_step": lbytes}
    stage_keys = ("ms_h2d", "ms_seed", "ms_sort", "ms_chain", "ms_subgraph", "ms_poa", "ms_trace", "ms_d2h", "ms_total")
    stage = {k: mean(k) for k in stage_keys}
    rec = {"value": total_reads * K / (ms_dev_max * 1e-3), "ms_per_step": ms_dev_max / K,
           "ms_steps": [round(sumf(acc, "ms_total"), 3) for acc in dev_steps], "wall_s_steps": [round(x, 4) for x in dev_walls],
           "config": config_of(name, wl, world), "clocks": clocks,
           "gpu_launches": int(sum(sumf(acc, "kernel_launches") for acc in dev_steps)),
           "aligned_fraction": total_aligned / total_reads, "stage_ms": stage,
           "roofline": roof_poa if stage["ms_poa"] >= stage["ms_seed"] else roof_seed, "roofline_poa": roof_poa, "roofline_lookup": roof_seed,
           "setup": {"index_build_or_load_s": t_index, "index_image_bytes": nbytes, "index_broadcast_s": t_bcast, "index_build_on_gpu": t_build_dev,
                     "read_simulation_s": float(red[3]), "wall_s_timed_region": t_wall},
           "counts": {k: sumf(last, k) for k in ("n_probes", "n_anchors", "n_chains", "n_alignments", "poa_cells", "lookup_bytes")}}
    if e2e:
        pin_last = e2e[3][-1]
        rec["e2e"] = {"value": total_reads * e2e[2] / e2e_pin_max, "unit": "reads/s", "clock": "caller (perf_counter around the calls), max over ranks",
                      "caller_buffers": "pinned", "pageable_value": total_reads * e2e[2] / e2e_pg_max,
                      "h2d_bytes_per_step": int(sumf(pin_last, "h2d_bytes")), "d2h_bytes_per_step": int(sumf(pin_last, "d2h_bytes")),
                      "ms_per_step": e2e_pin_max / e2e[2] * 1e3, "ms_per_step_pageable": e2e_pg_max / e2e[2] * 1e3, "steps": e2e[2],
                      "library_event_ms_per_step": sum(sumf(acc, "ms_total") for acc in e2e[3]) / e2e[2],
                      "stage_ms": {k: sum(sumf(acc, k) for acc in e2e[3]) / e2e[2] for k in stage_keys}}
    if gaf:
        rec["e2e_gaf"] = gaf
    if not args.no_cpu_baseline:
        from oracle import binding as orc       # CPU baseline leg: the one place bench.py executes oracle/ (as the checker and the timed CPU arm)
        threads = os.cpu_count() or 1
        oi = orc.OracleIndex(ix)
        n_sample = min(pick_cpu_sample(orc, oi, reads, threads, args.cpu_seconds), cuts[1])
        rate, sec, _, ccells, o = cpu_run(orc, oi, reads, n_sample, threads)
        n1 = max(4, int(n_sample * min(1.0, 3.0 / max(sec, 1e-3)) / threads))
        rate1, sec1, _, _, _ = cpu_run(orc, oi, reads, n1, 1)
        rec["cpu_baseline"] = {"value": rate, "unit": "reads/s", "cores": threads, "kind": "port",
                               "sample": "first %d reads of rank 0's shard, oracle port with threads over reads, %.1f s" % (n_sample, sec),
                               "cpu_gcups": ccells / sec / 1e9, "t1_value": rate1, "t1_sample": "first %d reads, 1 thread, %.1f s" % (n1, sec1),
                               "note": "scalar C++ restatement of the reference path (reference binary: unavailable on this box, no cargo); "
                                       "it allocates per DP row, so the GPU/CPU ratio says nothing about kernel quality"}
        # parity sample: the CUDA path on exactly these reads, every result array against the oracle's
        sub = Sub(reads, n_sample)
        g = vg.map_reads(ix, sub)
        bad = [f for f in PARITY_FIELDS if not np.array_equal(getattr(g, f), getattr(o, f))]
        rec["parity_sample"] = {"n": n_sample, "equal": not bad, "fields": len(PARITY_FIELDS), "differing": bad}
        g.free()
        if bad:
            log("PARITY FAILURE on %s: %s differ from the oracle" % (name, bad))
    ix.free_device()                                            # with its workspace: the next workload sizes its trace pool by the free memory
    if world > 1 and not args.no_sharded:
        G.keep.append((rec, name, wl, ix, batch))               # the sharded leg runs after every record is complete
    return rec


def pack_results(vg, L, outs):
    """Every result array of the calls of one shard as ONE byte buffer (what a rank ships to the process that owns the output)."""
    import ctypes as C

    import numpy as np
    parts = []
    u8 = C.POINTER(C.c_uint8)

    def raw(ptr, count, dt):
        nb = int(count) * np.dtype(dt).itemsize
        return np.ctypeslib.as_array(C.cast(ptr, u8), shape=(nb,)) if nb and ptr else np.zeros(0, np.uint8)

    for out in outs:
        v = vg.VgaResultsView()
        L.vga_results_view(out, C.byref(v))
        n, c, m = v.n_reads, v.n_chains, v.n_chain_anchors
        size = {"n": n, "n+1": n + 1, "c+1": c + 1, "m": m}
        n_path = int(v.aln_path_off[n]) if n else 0
        n_runs = int(v.aln_run_off[n]) if n else 0
        parts.append(np.array([n, c, m, n_path, n_runs], dtype=np.uint64).view(np.uint8))
        for fname, (sz, dt) in vg._FIELDS.items():
            parts.append(raw(getattr(v, fname), size[sz], dt))
        parts.append(raw(v.aln_path, n_path, np.uint64))
        parts.append(raw(v.aln_runs, n_runs, np.uint32))
    return np.concatenate(parts) if parts else np.zeros(0, np.uint8)


def sharded_leg(G, name, wl, ix, batch, args):
    """ONE process owns the reads and wants every result array back (what `vgaligner map` is): rank 0 holds the whole read set, scatters
    contiguous shards (host -> device -> NCCL -> device -> host), every rank maps its shard from host buffers through vga_map_batch,
    and the packed result arrays are gathered on rank 0.  One pass, caller's clock, barriers between the phases."""
    import ctypes as C
    import importlib

    import torch
    import torch.distributed as dist
    shards = importlib.import_module("rs_vgaligner_b200.shards")
    vg, L, dev, rank, world = G.vg, G.L, G.dev, G.rank, G.world
    total = wl["reads"]

    def barrier():
        torch.cuda.synchronize()
        dist.barrier()

    # the device index again (the records above released theirs together with the workspaces): upload on rank 0 + ONE broadcast
    cid = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        cid.copy_(torch.frombuffer(bytearray(vg.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(cid, 0)
    gix = vg.index_broadcast(ix.upload() if rank == 0 else None, bytes(cid.cpu().numpy().tobytes()), rank, world)
    ix._gpu = gix
    t0 = time.perf_counter()
    whole = vg.simulate_reads_mt(ix, total, wl["length"], *wl["err"], 0.1, wl["seed"], first_id=0) if rank == 0 else None
    t_sim = time.perf_counter() - t0
    params = vg.default_params()
    barrier()
    t0 = time.perf_counter()
    mine = shards.scatter_reads(whole, rank, world, dist, device=dev, with_names=False)
    barrier()
    t1 = time.perf_counter()
    n_calls = max(1, (mine.n + batch - 1) // batch)
    cuts = [mine.n * i // n_calls for i in range(n_calls + 1)]
    outs = []
    for i in range(n_calls):
        out = C.c_void_p()
        rc = L.vga_map_batch(gix, C.c_void_p(mine.seqs.ctypes.data), C.c_void_p(mine.offsets.ctypes.data + 8 * cuts[i]), cuts[i + 1] - cuts[i],
                             C.byref(params), C.byref(out))
        if rc != 0:
            raise RuntimeError("vga_map_batch: " + L.vga_last_error().decode())
        outs.append(out)
    barrier()
    t2 = time.perf_counter()
    packed = pack_results(vg, L, outs)
    t_pack = time.perf_counter() - t2
    got = shards.gather_bytes(packed, rank, world, dist, device=dev)
    barrier()
    t3 = time.perf_counter()
    for out in outs:
        L.vga_results_free(out)
    if rank != 0:
        return None
    return {"value": total / (t3 - t0), "unit": "reads/s", "scatter_s": t1 - t0, "map_s": t2 - t1, "gather_s": t3 - t2, "pack_s_rank0": t_pack,
            "scatter_bytes": int(whole.offsets[-1]) + 8 * (total + world), "gather_bytes": int(sum(g.size for g in got)),
            "read_simulation_s_rank0": t_sim, "passes": 1,
            "what": "rank 0 owns all reads: scatter of the shards + vga_map_batch on every rank (pageable host buffers) + gather of every result "
                    "array on rank 0, caller's clock; the limiter is the phase with the largest share"}


def main():
    capture_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("VGA_WORKLOAD", "C4+C5"), choices=sorted(WORKLOADS) + ["C4+C5"])
    ap.add_argument("--reads", type=int, default=0, help="override the TOTAL number of reads of the (first) workload")
    ap.add_argument("--batch-reads", type=int, default=0, help="reads per C-ABI call (default: per workload)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU work per baseline pass")
    ap.add_argument("--e2e-steps", type=int, default=2, help="timed steps of each host-buffer leg (<= --steps)")
    ap.add_argument("--gaf-reads", type=int, default=250000, help="reads of the batch that is also formatted as GAF text")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-gaf", action="store_true")
    ap.add_argument("--no-sharded", action="store_true", help="skip the scatter + map + gather leg of a multi-rank run (e2e_sharded)")
    ap.add_argument("--sharded-timeout", type=float, default=240.0, help="seconds after which the sharded leg is abandoned and the line printed without it")
    ap.add_argument("--no-index-build", action="store_true", help="skip the GPU index build measurement of the setup section")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    names = args.workload.split("+")
    wls = [dict(WORKLOADS[n]) for n in names]
    if args.reads:
        wls[0]["reads"] = args.reads

    import vgaligner_b200 as vg

    base = {"metric": "aligned reads/sec", "unit": "reads/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32", "data": "synthetic"}

    if args.impl == "reference":
        if rank != 0:
            return 0
        from oracle import binding as orc
        threads = os.cpu_count() or 1
        recs = [reference_record(vg, orc, n, w, threads, args) for n, w in zip(names, wls)]
        line = dict(base)
        line.update(recs[0])
        line.update({"impl": "reference", "gpu_launches": 0})
        if len(recs) > 1:
            line["long_reads"] = recs[1]
        emit(line)
        return 0

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the map path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    G = Gpu()
    G.vg, G.rank, G.local_rank, G.world = vg, rank, local_rank, world
    G.dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL prints its version banner on stdout (NCCL_DEBUG=VERSION/WARN/INFO); emit() keeps it off the JSON stream
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=G.dev)
    G.L = vg.gpu_lib()
    if G.L.vga_init(local_rank) != 0:
        raise SystemExit("vga_init: " + G.L.vga_last_error().decode())
    G.int32_peak = vg.measure_int32_peak() if rank == 0 else 0.0

    G.keep = []
    recs = [gpu_record(G, n, w, args) for n, w in zip(names, wls)]

    def build_line():
        line = dict(base)
        line.update(recs[0])
        if len(recs) > 1:
            line["long_reads"] = recs[1]
            line["config"]["also_measured"] = recs[1]["config"]["workload"] + " (record under long_reads)"
        return line

    if G.keep:
        # every record is complete: the sharded leg may not take the line with it if a rank falls out of step
        def give_up():
            log("sharded leg: no end after %.0f s, printing the line without it" % args.sharded_timeout)
            if rank == 0:
                for r in recs:
                    r.setdefault("e2e_sharded", {"error": "abandoned after %.0f s" % args.sharded_timeout})
                emit(build_line())
            os._exit(0)
        timer = threading.Timer(args.sharded_timeout, give_up)
        timer.daemon = True
        timer.start()
        for rec, n, w, ix, batch in G.keep:
            try:
                res = sharded_leg(G, n, w, ix, batch, args)
            except Exception as e:           # a rank that fails here leaves the others waiting: the timer ends the run
                log("sharded leg of %s failed on rank %d: %r" % (n, rank, e))
                res = {"error": repr(e)}
            if rank == 0:
                rec["e2e_sharded"] = res
            ix.free_device()
        timer.cancel()
    if rank == 0:
        line = build_line()
        emit(line)
        bad = [r["config"]["workload"] for r in recs if r.get("parity_sample") and not r["parity_sample"]["equal"]]
        if bad:
            log("parity sample differs from the oracle on", bad)
            if world > 1:
                dist.destroy_process_group()
            return 3
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
