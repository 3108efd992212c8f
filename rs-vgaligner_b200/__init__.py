"""rs-vgaligner_b200 — B200-native `vgaligner map --also-align` hot path.

Python mirror of the reference's operator surface for this path, over the C ABI in
``include/vga.h`` (libvgaligner_gpu.so) and the host library (libvga_host.so):

* :class:`Index`      — ``Index::build`` / ``load_from_file`` (reference src/index.rs:109-305)
* :func:`read_seqs_from_file` — reference src/io.rs:74-131
* :func:`map_reads`   — reference src/map.rs:27-216 (the GPU path; fails loudly without CUDA)

Nothing in this package imports, links or executes anything under ``oracle/``.
The directory name carries a hyphen (the repo layout asks for it); import it through the
``vgaligner_b200`` shim at the repository root.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)

u8p, u32p, u64p, i32p, f64p = (C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64),
                               C.POINTER(C.c_int32), C.POINTER(C.c_double))


class VgaIndexView(C.Structure):
    _fields_ = [("k", C.c_uint32), ("seq_len", C.c_uint64), ("n_nodes", C.c_uint64), ("n_edges", C.c_uint64),
                ("n_keys", C.c_uint64), ("n_table", C.c_uint64), ("seq_fwd", u8p), ("node_ref", u64p),
                ("edges", u64p), ("key_kmer2bit", u64p), ("key_offset", u64p), ("tab_start", u64p),
                ("tab_end", u64p)]


class VgaParams(C.Structure):
    _fields_ = [("bandwidth", C.c_uint32), ("max_gap", C.c_uint32), ("min_anchors", C.c_uint32),
                ("best_n", C.c_uint32), ("match", C.c_int32), ("mismatch", C.c_int32), ("o1", C.c_int32),
                ("e1", C.c_int32), ("o2", C.c_int32), ("e2", C.c_int32), ("wb", C.c_int32), ("wf", C.c_float),
                ("flags", C.c_uint32)]


class VgaResultsView(C.Structure):
    _fields_ = [("n_reads", C.c_uint32), ("n_anchors", u32p), ("chain_off", u64p), ("n_chains", C.c_uint64),
                ("chain_anchor_off", u64p), ("n_chain_anchors", C.c_uint64), ("ca_id", u32p), ("ca_qb", u32p),
                ("ca_tb", u32p), ("ca_te", u32p), ("ca_score", f64p), ("aln_status", u8p), ("aln_chain", u32p),
                ("aln_path_off", u64p), ("aln_path", u64p), ("aln_run_off", u64p), ("aln_runs", u32p),
                ("aln_path_len", u32p), ("aln_start_off", u32p), ("aln_end_off", u32p), ("aln_n_aligned", u32p),
                ("aln_score", i32p), ("aln_cells", u64p), ("dbg_anchor_off", u64p), ("dbg_qb", u32p),
                ("dbg_tb", u32p), ("dbg_te", u32p), ("dbg_sorted", u32p), ("dbg_f", f64p), ("dbg_pred", i32p),
                ("dbg_sub_off", u64p), ("dbg_sub_handles", u64p), ("dbg_edge_off", u64p), ("dbg_edges", u32p)]


class VgaStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("n_reads", "n_bases", "n_probes", "n_positions_read", "n_anchors",
                                          "n_chains", "n_alignments", "poa_cells", "lookup_bytes")] + \
               [(n, C.c_float) for n in ("ms_h2d", "ms_seed", "ms_sort", "ms_chain", "ms_subgraph", "ms_poa",
                                         "ms_trace", "ms_d2h", "ms_total", "ms_wall")] + \
               [("kernel_launches", C.c_uint32), ("poa_launches", C.c_uint32), ("h2d_bytes", C.c_uint64),
                ("d2h_bytes", C.c_uint64)]


VGA_FLAG_NO_ALIGN, VGA_FLAG_KEEP_DEBUG, VGA_FLAG_CHAIN_SCORES = 1, 2, 4
VGA_FLAG_BAND_GRAN8, VGA_FLAG_BAND_GRAN16, VGA_FLAG_ALL_ORIENTATIONS = 0x100, 0x200, 0x1000
VGA_INDEX_ALL_ORIENTATIONS = 1
ALN_PLACEHOLDER, ALN_OK, ALN_EMPTY, ALN_TRACE_ERR, ALN_BAND_ERR, ALN_UNSUPPORTED, ALN_TOO_WIDE = range(7)

GPU_SYMBOLS = ["vga_init", "vga_device_count", "vga_index_upload", "vga_index_free", "vga_index_image_size",
               "vga_index_image_copy_to_device", "vga_index_import_device", "vga_default_params", "vga_map_batch",
               "vga_map_batch_device", "vga_results_view", "vga_results_stats", "vga_results_free",
               "vga_last_error", "vga_measure_int32_peak", "vga_comm_unique_id", "vga_index_broadcast", "vga_index_build_device",
               "vga_index_build_view", "vga_index_build_free", "vga_index_upload_ex", "vga_seed_count_noalloc"]

HASH_FN = C.CFUNCTYPE(C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p)     # vgah_hash_fn: (kmer bytes, k, user) -> key hash

_host = None
_gpu = None


class VgaError(RuntimeError):
    pass


def host_lib():
    global _host
    if _host is None:
        path = os.path.join(_HERE, "libvga_host.so")
        if not os.path.exists(path):
            raise VgaError("libvga_host.so is not built: run `make host` (or __graft_entry__.build())")
        L = C.CDLL(path)
        L.vgah_last_error.restype = C.c_char_p
        for f in ("vgah_index_build_file", "vgah_index_build_text", "vgah_index_load", "vgah_reads_from_file",
                  "vgah_reads_simulate"):
            getattr(L, f).restype = C.c_void_p
        L.vgah_index_build_file.argtypes = [C.c_char_p, C.c_uint32, C.c_uint64, C.c_uint64]
        L.vgah_index_build_text.argtypes = [C.c_char_p, C.c_uint32, C.c_uint64, C.c_uint64]
        L.vgah_index_load.argtypes = [C.c_char_p]
        L.vgah_index_save.argtypes = [C.c_void_p, C.c_char_p]
        L.vgah_index_save_idx.argtypes = [C.c_void_p, C.c_char_p]
        L.vgah_index_free.argtypes = [C.c_void_p]
        L.vgah_index_view.argtypes = [C.c_void_p, C.POINTER(VgaIndexView)]
        L.vgah_index_keys.argtypes = [C.c_void_p, u64p]
        L.vgah_index_keys.restype = C.c_void_p
        L.vgah_index_seq_rev.argtypes = [C.c_void_p]
        L.vgah_index_seq_rev.restype = C.c_char_p
        L.vgah_reads_from_file.argtypes = [C.c_char_p]
        L.vgah_reads_simulate.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_double, C.c_double, C.c_double,
                                          C.c_double, C.c_uint64]
        L.vgah_reads_simulate_mt.restype = C.c_void_p
        L.vgah_reads_simulate_mt.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_double, C.c_double, C.c_double,
                                             C.c_double, C.c_uint64, C.c_int]
        L.vgah_reads_free.argtypes = [C.c_void_p]
        L.vgah_reads_count.argtypes = [C.c_void_p]
        L.vgah_reads_count.restype = C.c_uint32
        L.vgah_reads_seqs.argtypes = [C.c_void_p]
        L.vgah_reads_seqs.restype = C.c_void_p
        L.vgah_reads_offsets.argtypes = [C.c_void_p]
        L.vgah_reads_offsets.restype = u64p
        L.vgah_reads_names.argtypes = [C.c_void_p]
        L.vgah_reads_names.restype = C.c_char_p
        L.vgah_reads_write_fasta.argtypes = [C.c_void_p, C.c_char_p]
        L.vgah_synth_gfa.argtypes = [C.c_uint64, C.c_uint32, C.c_uint64, u64p]
        L.vgah_synth_gfa.restype = C.c_void_p
        L.vgah_free_str.argtypes = [C.c_void_p]
        L.vgah_gaf.argtypes = [C.c_void_p, C.POINTER(VgaResultsView), C.c_void_p, C.c_int, u64p]
        L.vgah_gaf.restype = C.c_void_p
        L.vgah_gaf_raw.argtypes = [C.c_void_p, C.POINTER(VgaResultsView), C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int,
                                   C.c_int, u64p, u64p]
        L.vgah_gaf_raw.restype = C.c_void_p
        L.vgah_derive_keys.argtypes = [C.POINTER(VgaIndexView), u64p, HASH_FN, C.c_void_p, u64p, u64p, u64p]
        _host = L
    return _host


def gpu_lib():
    """The CUDA library.  There is no fallback: a missing library is an error."""
    global _gpu
    if _gpu is None:
        path = os.environ.get("VGA_GPU_LIB") or os.path.join(_HERE, "libvgaligner_gpu.so")
        if not os.path.exists(path):
            raise VgaError("libvgaligner_gpu.so is not built: run `make gpu` (or __graft_entry__.build()); "
                           "the map path has no CPU fallback")
        L = C.CDLL(path)
        L.vga_last_error.restype = C.c_char_p
        L.vga_init.argtypes = [C.c_int]
        L.vga_index_upload.argtypes = [C.POINTER(VgaIndexView), C.POINTER(C.c_void_p)]
        L.vga_index_upload_ex.argtypes = [C.POINTER(VgaIndexView), C.c_uint32, C.POINTER(C.c_void_p)]
        L.vga_index_free.argtypes = [C.c_void_p]
        L.vga_index_image_size.argtypes = [C.c_void_p, u64p]
        L.vga_index_image_copy_to_device.argtypes = [C.c_void_p, C.c_void_p]
        L.vga_index_import_device.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p)]
        L.vga_default_params.argtypes = [C.POINTER(VgaParams)]
        L.vga_map_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(VgaParams),
                                    C.POINTER(C.c_void_p)]
        L.vga_map_batch_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64,
                                           C.POINTER(VgaParams), C.POINTER(C.c_void_p)]
        L.vga_seed_count_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p]
        L.vga_results_view.argtypes = [C.c_void_p, C.POINTER(VgaResultsView)]
        L.vga_results_stats.argtypes = [C.c_void_p, C.POINTER(VgaStats)]
        L.vga_results_free.argtypes = [C.c_void_p]
        L.vga_measure_int32_peak.argtypes = [C.POINTER(C.c_double)]
        _gpu = L
    return _gpu


def _check(rc, what):
    if rc != 0:
        raise VgaError("%s failed (%d): %s" % (what, rc, gpu_lib().vga_last_error().decode()))


def default_params(**kw):
    """map_reads' defaults (reference src/subcommands/map_main.rs:30-52,103; abPOA defaults)."""
    p = VgaParams(50, 1000, 3, 1, 2, 4, 4, 2, 24, 1, 10, 0.01, 0)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def _arr(ptr, n, dtype):
    if n == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(int(n),)).view(dtype).copy()


class Index:
    """Host-side index (the reference's `Index`)."""

    def __init__(self, handle):
        if not handle:
            raise VgaError("index: " + host_lib().vgah_last_error().decode())
        self._h = C.c_void_p(handle)
        self.view = VgaIndexView()
        host_lib().vgah_index_view(self._h, C.byref(self.view))
        self._gpu = None

    @classmethod
    def build(cls, gfa_path, kmer_length, max_furcations=100, max_degree=100):
        return cls(host_lib().vgah_index_build_file(os.fsencode(gfa_path), kmer_length, max_furcations, max_degree))

    @classmethod
    def build_from_text(cls, gfa_text, kmer_length, max_furcations=100, max_degree=100):
        return cls(host_lib().vgah_index_build_text(gfa_text.encode(), kmer_length, max_furcations, max_degree))

    @classmethod
    def load_from_file(cls, path):
        return cls(host_lib().vgah_index_load(os.fsencode(path)))

    def store_as_idx(self, path):
        """the reference's own index file (bincode of `Index`, src/index.rs:284-305; layout in host/idx_file.cpp)"""
        if host_lib().vgah_index_save_idx(self._h, os.fsencode(path)) != 0:
            raise VgaError(host_lib().vgah_last_error().decode())

    def store_in_file(self, path):
        if host_lib().vgah_index_save(self._h, os.fsencode(path)) != 0:
            raise VgaError(host_lib().vgah_last_error().decode())

    # -- plain-data accessors
    @property
    def kmer_length(self):
        return self.view.k

    @property
    def seq_fwd(self):
        return C.string_at(self.view.seq_fwd, self.view.seq_len).decode()

    @property
    def seq_rev(self):
        return host_lib().vgah_index_seq_rev(self._h).decode()

    @property
    def node_ref(self):
        return _arr(self.view.node_ref, 3 * (self.view.n_nodes + 1), np.uint64).reshape(-1, 3)

    @property
    def edges(self):
        return _arr(self.view.edges, self.view.n_edges, np.uint64)

    @property
    def n_kmers(self):
        return int(self.view.n_keys)

    @property
    def n_kmer_pos(self):
        return int(self.view.n_table)

    def keys_bytes(self):
        n = C.c_uint64()
        p = host_lib().vgah_index_keys(self._h, C.byref(n))
        return C.string_at(p, n.value * self.view.k), n.value

    def table(self):
        return (_arr(self.view.tab_start, self.view.n_table, np.uint64), _arr(self.view.tab_end, self.view.n_table, np.uint64),
                _arr(self.view.key_offset, self.view.n_keys, np.uint64))

    def derive_keys(self, key_hash, hash_fn, n_keys=None):
        """`Index::walk_kmer` for every key: the 2-bit packed k-mers rebuilt from the table positions, the
        linearisation and the edge lists only, candidates told apart by `hash_fn(kmer_bytes) == key_hash[i]`.
        Returns (key_kmer2bit, n_ambiguous, n_hopping)."""
        v = VgaIndexView()
        C.memmove(C.byref(v), C.byref(self.view), C.sizeof(VgaIndexView))
        v.key_kmer2bit = None                  # not available to the walk
        if n_keys is not None:
            v.n_keys = min(int(n_keys), self.view.n_keys)
        kh = np.ascontiguousarray(key_hash if key_hash is not None else np.zeros(1), dtype=np.uint64)
        out = np.zeros(v.n_keys, dtype=np.uint64)
        amb, hop = C.c_uint64(), C.c_uint64()
        cb = HASH_FN(lambda p, k, _u: int(hash_fn(C.string_at(p, k))) & 0xFFFFFFFFFFFFFFFF) if hash_fn is not None else HASH_FN(0)
        rc = host_lib().vgah_derive_keys(C.byref(v), kh.ctypes.data_as(u64p) if hash_fn is not None else None, cb, None,
                                         out.ctypes.data_as(u64p), C.byref(amb), C.byref(hop))
        if rc != 0:
            raise VgaError("vgah_derive_keys failed (%d)" % rc)
        return out, amb.value, hop.value

    # -- device side
    def upload(self, device=None):
        """vga_index_upload: device-resident k-mer table + graph arrays."""
        if self._gpu is None:
            L = gpu_lib()
            if device is not None:
                _check(L.vga_init(int(device)), "vga_init")
            out = C.c_void_p()
            _check(L.vga_index_upload_ex(C.byref(self.view), int(getattr(self, "upload_flags", 0)), C.byref(out)), "vga_index_upload_ex")
            self._gpu = out
        return self._gpu

    def adopt_device_image(self, dev_ptr, nbytes):
        """Adopt an index image that an NCCL broadcast delivered to this rank's GPU."""
        out = C.c_void_p()
        _check(gpu_lib().vga_index_import_device(C.c_void_p(dev_ptr), nbytes, C.byref(out)), "vga_index_import_device")
        self._gpu = out
        return out

    def image_size(self):
        n = C.c_uint64()
        _check(gpu_lib().vga_index_image_size(self.upload(), C.byref(n)), "vga_index_image_size")
        return n.value

    def copy_image_to(self, dev_ptr):
        _check(gpu_lib().vga_index_image_copy_to_device(self.upload(), C.c_void_p(dev_ptr)), "vga_index_image_copy_to_device")

    def free_device(self):
        if self._gpu is not None:
            gpu_lib().vga_index_free(self._gpu)
            self._gpu = None

    def __del__(self):
        try:
            self.free_device()
            if self._h:
                host_lib().vgah_index_free(self._h)
                self._h = None
        except Exception:
            pass


class Reads:
    """A batch of query sequences (the reference's Vec<QuerySequence>)."""

    def __init__(self, handle):
        if not handle:
            raise VgaError("reads: " + host_lib().vgah_last_error().decode())
        self._h = C.c_void_p(handle)
        L = host_lib()
        self.n = L.vgah_reads_count(self._h)
        self.offsets = _arr(L.vgah_reads_offsets(self._h), self.n + 1, np.uint64)
        nb = int(self.offsets[-1]) if self.n else 0
        if nb:
            self.seqs = np.ctypeslib.as_array(C.cast(L.vgah_reads_seqs(self._h), u8p), shape=(nb,)).copy()
        else:
            self.seqs = np.zeros(0, np.uint8)
        self._names = None

    @property
    def names(self):
        """read names (materialised on first use: simulated sets of millions of reads never need them as a list)"""
        if self._names is None:
            self._names = host_lib().vgah_reads_names(self._h).decode().split("\n") if self.n else []
        return self._names

    @names.setter
    def names(self, value):
        self._names = list(value)

    def seq(self, i):
        return self.seqs[int(self.offsets[i]):int(self.offsets[i + 1])].tobytes().decode()

    def write_fasta(self, path):
        host_lib().vgah_reads_write_fasta(self._h, os.fsencode(path))

    def __del__(self):
        try:
            if self._h:
                host_lib().vgah_reads_free(self._h)
                self._h = None
        except Exception:
            pass


def read_seqs_from_file(path):
    return Reads(host_lib().vgah_reads_from_file(os.fsencode(path)))


def simulate_reads(index, n, length, sub=0.01, ins=0.0005, dele=0.0005, rev_frac=0.1, seed=77):
    return Reads(host_lib().vgah_reads_simulate(index._h, n, length, sub, ins, dele, rev_frac, seed))


def simulate_reads_mt(index, n, length, sub=0.01, ins=0.0005, dele=0.0005, rev_frac=0.1, seed=77, first_id=0, threads=0):
    """Same generator with one RNG stream per read (any sub-range reproducible on any rank), on all host cores."""
    return Reads(host_lib().vgah_reads_simulate_mt(index._h, first_id, n, length, sub, ins, dele, rev_frac, seed, threads))


def synth_gfa(backbone_len, site_every=100, seed=1):
    n = C.c_uint64()
    p = host_lib().vgah_synth_gfa(backbone_len, site_every, seed, C.byref(n))
    s = C.string_at(p, n.value).decode()
    host_lib().vgah_free_str(p)
    return s


_FIELDS = {
    "n_anchors": ("n", np.uint32), "chain_off": ("n+1", np.uint64), "chain_anchor_off": ("c+1", np.uint64),
    "ca_id": ("m", np.uint32), "ca_qb": ("m", np.uint32), "ca_tb": ("m", np.uint32), "ca_te": ("m", np.uint32),
    "ca_score": ("m", np.float64), "aln_status": ("n", np.uint8), "aln_chain": ("n", np.uint32),
    "aln_path_off": ("n+1", np.uint64), "aln_run_off": ("n+1", np.uint64), "aln_path_len": ("n", np.uint32),
    "aln_start_off": ("n", np.uint32), "aln_end_off": ("n", np.uint32), "aln_n_aligned": ("n", np.uint32),
    "aln_score": ("n", np.int32), "aln_cells": ("n", np.uint64),
}


_VIEW_PTR = {"n_anchors": u32p, "chain_off": u64p, "chain_anchor_off": u64p, "ca_id": u32p, "ca_qb": u32p, "ca_tb": u32p,
             "ca_te": u32p, "ca_score": f64p, "aln_status": u8p, "aln_chain": u32p, "aln_path_off": u64p, "aln_path": u64p,
             "aln_run_off": u64p, "aln_runs": u32p, "aln_path_len": u32p, "aln_start_off": u32p, "aln_end_off": u32p,
             "aln_n_aligned": u32p, "aln_score": i32p, "aln_cells": u64p}


def view_from_arrays(res, n_reads):
    """A VgaResultsView over numpy arrays carrying the C ABI's field names (e.g. results gathered from several ranks).
    The arrays must stay alive while the view is used."""
    v = VgaResultsView()
    v.n_reads = n_reads
    keep = []
    for name, pt in _VIEW_PTR.items():
        a = getattr(res, name, None)
        if a is None or a.size == 0:
            continue
        a = np.ascontiguousarray(a)
        keep.append(a)
        setattr(v, name, a.ctypes.data_as(pt))
    v.n_chains = max(0, len(res.chain_anchor_off) - 1)
    v.n_chain_anchors = len(res.ca_qb)
    v._keep = keep
    return v


def gaf_text(index, view, which, seqs, offsets, names=None, name_base=0, threads=0):
    """GAF text of a result view (which=0 chains file, 1 alignments file) -> (text bytes, reads whose alignment failed)."""
    n, nf = C.c_uint64(), C.c_uint64()
    seqs = np.ascontiguousarray(seqs)
    offsets = np.ascontiguousarray(offsets)
    joined = None if names is None else "\n".join(names).encode()
    p = host_lib().vgah_gaf_raw(index._h, C.byref(view), joined, name_base, seqs.ctypes.data_as(C.c_void_p),
                                offsets.ctypes.data_as(C.c_void_p), which, threads, C.byref(n), C.byref(nf))
    if not p:
        raise VgaError("gaf: " + host_lib().vgah_last_error().decode())
    s = C.string_at(p, n.value)
    host_lib().vgah_free_str(p)
    return s, nf.value


class MapResults:
    """Flat result arrays of one vga_map_batch call (numpy copies) + the GAF writers."""

    def __init__(self, handle, index, reads):
        L = gpu_lib()
        self._h = handle
        self.index, self.reads = index, reads
        v = VgaResultsView()
        _check(L.vga_results_view(handle, C.byref(v)), "vga_results_view")
        self.view = v
        n, c, m = v.n_reads, v.n_chains, v.n_chain_anchors
        size = {"n": n, "n+1": n + 1, "c+1": c + 1, "m": m}
        for name, (sz, dt) in _FIELDS.items():
            setattr(self, name, _arr(getattr(v, name), size[sz], dt))
        self.aln_path = _arr(v.aln_path, int(self.aln_path_off[-1]), np.uint64)
        self.aln_runs = _arr(v.aln_runs, int(self.aln_run_off[-1]), np.uint32)
        if v.dbg_anchor_off:
            self.dbg_anchor_off = _arr(v.dbg_anchor_off, n + 1, np.uint64)
            a = int(self.dbg_anchor_off[-1])
            for name, dt in (("dbg_qb", np.uint32), ("dbg_tb", np.uint32), ("dbg_te", np.uint32), ("dbg_sorted", np.uint32),
                             ("dbg_f", np.float64), ("dbg_pred", np.int32)):
                setattr(self, name, _arr(getattr(v, name), a, dt))
            self.dbg_sub_off = _arr(v.dbg_sub_off, n + 1, np.uint64)
            self.dbg_sub_handles = _arr(v.dbg_sub_handles, int(self.dbg_sub_off[-1]), np.uint64)
            self.dbg_edge_off = _arr(v.dbg_edge_off, n + 1, np.uint64)
            self.dbg_edges = _arr(v.dbg_edges, 2 * int(self.dbg_edge_off[-1]), np.uint32).reshape(-1, 2)
        st = VgaStats()
        _check(L.vga_results_stats(handle, C.byref(st)), "vga_results_stats")
        self.stats = {f: getattr(st, f) for f, _ in VgaStats._fields_}

    def gaf(self, which):
        """GAF text: which=0 `<prefix>-chains.gaf`, which=1 `<prefix>-alignments.gaf` (src/map.rs:135-184)."""
        s, self.n_failed = gaf_text(self.index, self.view, which, self.reads.seqs, self.reads.offsets, self.reads.names)
        return s.decode()

    def free(self):
        if self._h is not None:
            gpu_lib().vga_results_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def map_reads(index, inputs, bandwidth=50, max_gap=1000, chain_min_n_anchors=3, also_align=True, align_best_n=1,
              out_prefix=None, keep_debug=False, chain_scores=False, params=None):
    """map_reads (reference src/map.rs:27-216) on the GPU.

    Returns :class:`MapResults`; when ``out_prefix`` is given also writes ``<prefix>-chains.gaf`` and
    (with ``also_align``) ``<prefix>-alignments.gaf`` exactly like the reference (a prefix ending in
    ``.gaf`` is used as the file name for both, the alignments overwriting the chains, src/map.rs:136-139,175-178).
    """
    L = gpu_lib()
    p = params or default_params()
    if params is None:
        p.bandwidth, p.max_gap, p.min_anchors, p.best_n = bandwidth, max_gap, chain_min_n_anchors, align_best_n
        p.flags = ((0 if also_align else VGA_FLAG_NO_ALIGN) | (VGA_FLAG_KEEP_DEBUG if keep_debug else 0) |
                   (VGA_FLAG_CHAIN_SCORES if chain_scores else 0))
    gix = index.upload()
    out = C.c_void_p()
    seqs = np.ascontiguousarray(inputs.seqs)
    offs = np.ascontiguousarray(inputs.offsets)
    _check(L.vga_map_batch(gix, seqs.ctypes.data_as(C.c_void_p), offs.ctypes.data_as(C.c_void_p), inputs.n, C.byref(p),
                           C.byref(out)), "vga_map_batch")
    res = MapResults(out, index, inputs)
    if out_prefix is not None:
        gaf_name = out_prefix if out_prefix.endswith(".gaf") else out_prefix + "-chains.gaf"
        with open(gaf_name, "w") as f:
            f.write(res.gaf(0))
        if not (p.flags & VGA_FLAG_NO_ALIGN):
            gaf_name = out_prefix if out_prefix.endswith(".gaf") else out_prefix + "-alignments.gaf"
            with open(gaf_name, "w") as f:
                f.write(res.gaf(1))
    return res


class VgaGraphView(C.Structure):
    _fields_ = [("seq_len", C.c_uint64), ("n_nodes", C.c_uint64), ("n_edges", C.c_uint64), ("seq_fwd", C.c_void_p),
                ("node_ref", u64p), ("edges", u64p)]


def index_build_device(index, k=None):
    """vga_index_build_device: the k-mer table of `Index::build` on the GPU from the linearised graph of `index` (seq_fwd, NodeRef,
    edges).  Returns (key_kmer2bit, key_offset, tab_start, tab_end, device milliseconds); raises VgaError(-4) for graphs the
    device build does not admit (cyclic / high-degree: build those on the host)."""
    L = gpu_lib()
    v = index.view
    g = VgaGraphView(v.seq_len, v.n_nodes, v.n_edges, C.cast(v.seq_fwd, C.c_void_p), v.node_ref, v.edges)
    out = C.c_void_p()
    L.vga_index_build_device.argtypes = [C.POINTER(VgaGraphView), C.c_uint32, C.POINTER(C.c_void_p)]
    L.vga_index_build_view.argtypes = [C.c_void_p, C.POINTER(VgaGraphView), C.c_uint32, C.POINTER(VgaIndexView), C.POINTER(C.c_float)]
    L.vga_index_build_free.argtypes = [C.c_void_p]
    kk = int(k or v.k)
    _check(L.vga_index_build_device(C.byref(g), kk, C.byref(out)), "vga_index_build_device")
    bv, ms = VgaIndexView(), C.c_float()
    _check(L.vga_index_build_view(out, C.byref(g), kk, C.byref(bv), C.byref(ms)), "vga_index_build_view")
    res = (_arr(bv.key_kmer2bit, bv.n_keys, np.uint64), _arr(bv.key_offset, bv.n_keys, np.uint64),
           _arr(bv.tab_start, bv.n_table, np.uint64), _arr(bv.tab_end, bv.n_table, np.uint64), ms.value)
    L.vga_index_build_free(out)
    return res


def comm_unique_id():
    """vga_comm_unique_id: the 128 bytes rank 0 hands to the other ranks before vga_index_broadcast."""
    buf = (C.c_ubyte * 128)()
    _check(gpu_lib().vga_comm_unique_id(buf), "vga_comm_unique_id")
    return bytes(buf)


def index_broadcast(gix, id_bytes, rank, world):
    """vga_index_broadcast on the calling thread's device: rank 0 passes its uploaded handle, the others None; returns the
    handle of this rank's replica (ONE ncclBroadcast of the index image inside the library)."""
    h = C.c_void_p(gix.value if (rank == 0 and gix is not None) else None)
    _check(gpu_lib().vga_index_broadcast(C.byref(h), id_bytes, rank, world), "vga_index_broadcast")
    return h


class _Slice:
    """reads [lo, hi) of a read set, sharing the sequence bytes (offsets stay absolute, as vga_map_batch accepts them)"""

    def __init__(self, reads, lo, hi):
        self.n, self.seqs, self.offsets = hi - lo, reads.seqs, reads.offsets[lo:hi + 1]
        self._reads, self._lo, self._hi = reads, lo, hi

    @property
    def names(self):
        return self._reads.names[self._lo:self._hi]


def map_reads_multi(index, inputs, n_gpus, params=None):
    """map_reads sharded over `n_gpus` devices of this box through the C ABI alone (no torch): one host thread per GPU,
    index replicated by vga_index_broadcast, rank r maps the contiguous range [r*n/N, (r+1)*n/N).  Returns the per-rank
    MapResults in rank order (their concatenation is the single-GPU result: output order = input order, src/map.rs:123-133)."""
    import threading
    L = gpu_lib()
    p = params or default_params()
    _check(L.vga_init(0), "vga_init")
    g0 = index.upload()
    cid = comm_unique_id() if n_gpus > 1 else bytes(128)
    out, errs = [None] * n_gpus, [None] * n_gpus
    seqs = np.ascontiguousarray(inputs.seqs)
    offs = np.ascontiguousarray(inputs.offsets)

    def worker(r):
        try:
            _check(L.vga_init(r), "vga_init")
            gix = index_broadcast(g0, cid, r, n_gpus)
            lo, hi = inputs.n * r // n_gpus, inputs.n * (r + 1) // n_gpus
            res = C.c_void_p()
            _check(L.vga_map_batch(gix, seqs.ctypes.data_as(C.c_void_p), C.c_void_p(offs.ctypes.data + 8 * lo), hi - lo, C.byref(p),
                                   C.byref(res)), "vga_map_batch")
            out[r] = MapResults(res, index, _Slice(inputs, lo, hi))
            if r != 0:
                L.vga_index_free(gix)
        except Exception as e:      # noqa: BLE001 - reported to the caller below
            errs[r] = e

    th = [threading.Thread(target=worker, args=(r,)) for r in range(1, n_gpus)]
    for t in th:
        t.start()
    worker(0)
    for t in th:
        t.join()
    L.vga_init(0)
    for e in errs:
        if e is not None:
            raise e
    return out


def measure_int32_peak():
    v = C.c_double()
    _check(gpu_lib().vga_measure_int32_peak(C.byref(v)), "vga_measure_int32_peak")
    return v.value
