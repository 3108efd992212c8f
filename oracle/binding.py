"""TEST INFRASTRUCTURE — ctypes binding of oracle/_build/liboracle.so (see oracle/oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
u32p, u64p = C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)


class OrcParams(C.Structure):
    _fields_ = [("bandwidth", C.c_uint32), ("max_gap", C.c_uint32), ("min_anchors", C.c_uint32),
                ("best_n", C.c_uint32), ("match", C.c_int32), ("mismatch", C.c_int32), ("o1", C.c_int32),
                ("e1", C.c_int32), ("o2", C.c_int32), ("e2", C.c_int32), ("wb", C.c_int32), ("wf", C.c_float),
                ("flags", C.c_uint32)]


FLAG_NO_ALIGN, FLAG_KEEP_DEBUG, FLAG_LITERAL_EXTEND = 1, 2, 4
FLAG_BAND_GRAN8, FLAG_BAND_GRAN16, FLAG_INT16_SCORES, FLAG_NO_CS_PREFIX, FLAG_ALL_ORIENTATIONS = 0x100, 0x200, 0x400, 0x800, 0x1000
REV_BIT = 1 << 63
_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "_build", "liboracle.so")
        if not os.path.exists(path):
            raise RuntimeError("oracle not built: run `make oracle`")
        L = C.CDLL(path)
        L.orc_index_create.restype = C.c_void_p
        L.orc_index_create.argtypes = [C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64, C.c_char_p, C.c_void_p, C.c_void_p,
                                       C.c_uint64, C.c_char_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
        L.orc_index_free.argtypes = [C.c_void_p]
        for f in ("orc_bv_rank", "orc_bv_inverse_rank", "orc_bv_select"):
            getattr(L, f).restype = C.c_uint64
            getattr(L, f).argtypes = [C.c_void_p, C.c_uint64]
        L.orc_handle_from_seqpos.restype = C.c_uint64
        L.orc_handle_from_seqpos.argtypes = [C.c_void_p, C.c_int, C.c_uint64]
        for f in ("orc_incoming", "orc_outgoing"):
            getattr(L, f).restype = C.c_uint64
            getattr(L, f).argtypes = [C.c_void_p, C.c_uint64, u64p, C.c_uint64]
        L.orc_seq_from_handle.restype = C.c_uint64
        L.orc_seq_from_handle.argtypes = [C.c_void_p, C.c_uint64, C.c_char_p, C.c_uint64]
        L.orc_positions.restype = C.c_uint64
        L.orc_positions.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, u64p, u64p, C.c_uint64]
        L.orc_anchors.restype = C.c_uint64
        L.orc_anchors.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_int, u64p, u64p, u64p, u64p, C.c_uint64]
        L.orc_score_anchor.restype = C.c_double
        L.orc_score_anchor.argtypes = [C.c_uint64] * 4 + [C.c_double] + [C.c_uint64] * 6
        L.orc_map_reads.restype = C.c_void_p
        L.orc_map_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(OrcParams), C.c_int]
        L.orc_results_free.argtypes = [C.c_void_p]
        L.orc_results_field.restype = C.c_void_p
        L.orc_results_field.argtypes = [C.c_void_p, C.c_char_p, u64p]
        L.orc_results_gaf.restype = C.c_void_p
        L.orc_results_gaf.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, C.c_void_p, C.c_void_p, C.c_int, u64p]
        L.orc_poa_align.restype = C.c_void_p
        L.orc_poa_align.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_char_p, C.c_uint32,
                                    C.POINTER(OrcParams)]
        L.orc_default_params.argtypes = [C.POINTER(OrcParams)]
        _lib = L
    return _lib


def default_params(**kw):
    p = OrcParams()
    lib().orc_default_params(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


_DT = {"n_anchors": np.uint32, "chain_off": np.uint64, "chain_anchor_off": np.uint64, "ca_id": np.uint32,
       "ca_qb": np.uint32, "ca_tb": np.uint32, "ca_te": np.uint32, "ca_score": np.float64, "aln_status": np.uint8,
       "aln_chain": np.uint32, "aln_path_off": np.uint64, "aln_path": np.uint64, "aln_run_off": np.uint64,
       "aln_runs": np.uint32, "aln_path_len": np.uint32, "aln_start_off": np.uint32, "aln_end_off": np.uint32,
       "aln_n_aligned": np.uint32, "aln_score": np.int32, "aln_cells": np.uint64, "dbg_anchor_off": np.uint64,
       "dbg_qb": np.uint32, "dbg_tb": np.uint32, "dbg_te": np.uint32, "dbg_sorted": np.uint32, "dbg_f": np.float64,
       "dbg_pred": np.int32, "dbg_sub_off": np.uint64, "dbg_sub_handles": np.uint64, "dbg_edge_off": np.uint64,
       "dbg_edges": np.uint32}


class OracleResults:
    def __init__(self, handle):
        self._h = C.c_void_p(handle)
        L = lib()
        for name, dt in _DT.items():
            n = C.c_uint64()
            p = L.orc_results_field(self._h, name.encode(), C.byref(n))
            if n.value == 0 or not p:
                a = np.zeros(0, dtype=dt)
            else:
                a = np.frombuffer(C.string_at(p, n.value * np.dtype(dt).itemsize), dtype=dt).copy()
            setattr(self, name, a)
        self.dbg_edges = self.dbg_edges.reshape(-1, 2)

    def gaf(self, oracle_index, reads, which):
        n = C.c_uint64()
        seqs = np.ascontiguousarray(reads.seqs)
        offs = np.ascontiguousarray(reads.offsets)
        p = lib().orc_results_gaf(self._h, oracle_index._h, "\n".join(reads.names).encode(), seqs.ctypes.data_as(C.c_void_p),
                                  offs.ctypes.data_as(C.c_void_p), which, C.byref(n))
        return C.string_at(p, n.value).decode()

    def __del__(self):
        try:
            if self._h:
                lib().orc_results_free(self._h)
                self._h = None
        except Exception:
            pass


class OracleIndex:
    """Oracle-side copy of an index, built from the host index's flat arrays."""

    def __init__(self, host_index):
        v = host_index.view
        keys, n_keys = host_index.keys_bytes()
        self.k = v.k
        self._h = C.c_void_p(lib().orc_index_create(v.k, v.seq_len, v.n_nodes, v.n_edges,
                                                    C.cast(v.seq_fwd, C.c_char_p), C.cast(v.node_ref, C.c_void_p),
                                                    C.cast(v.edges, C.c_void_p), n_keys, keys,
                                                    C.cast(v.key_offset, C.c_void_p), v.n_table,
                                                    C.cast(v.tab_start, C.c_void_p), C.cast(v.tab_end, C.c_void_p)))

    def rank(self, pos): return lib().orc_bv_rank(self._h, pos)
    def inverse_rank(self, pos): return lib().orc_bv_inverse_rank(self._h, pos)
    def select(self, n): return lib().orc_bv_select(self._h, n)
    def handle_from_seqpos(self, reverse, pos): return lib().orc_handle_from_seqpos(self._h, int(reverse), pos)

    def _handles(self, fn, h):
        buf = (C.c_uint64 * 4096)()
        n = fn(self._h, h, buf, 4096)
        return [int(buf[i]) for i in range(n)]

    def incoming(self, h): return self._handles(lib().orc_incoming, h)
    def outgoing(self, h): return self._handles(lib().orc_outgoing, h)

    def seq_from_handle(self, h):
        buf = C.create_string_buffer(1 << 20)
        n = lib().orc_seq_from_handle(self._h, h, buf, 1 << 20)
        return buf.raw[:n].decode()

    def positions(self, kmer):
        s = (C.c_uint64 * 65536)(); e = (C.c_uint64 * 65536)()
        n = lib().orc_positions(self._h, kmer.encode(), len(kmer), s, e, 65536)
        f = lambda x: ("R" if x & REV_BIT else "F", x & ~REV_BIT)
        return [(f(int(s[i])), f(int(e[i]))) for i in range(n)]

    def anchors(self, read, only_forward=True):
        cap = 1 << 20
        a = [(C.c_uint64 * cap)() for _ in range(4)]
        n = lib().orc_anchors(self._h, read.encode(), len(read), int(only_forward), a[0], a[1], a[2], a[3], cap)
        f = lambda x: ("R" if x & REV_BIT else "F", x & ~REV_BIT)
        return [(int(a[0][i]), int(a[1][i]), f(int(a[2][i])), f(int(a[3][i]))) for i in range(n)]

    def map_reads(self, reads, params=None, n_threads=1):
        p = params or default_params()
        seqs = np.ascontiguousarray(reads.seqs)
        offs = np.ascontiguousarray(reads.offsets)
        return OracleResults(lib().orc_map_reads(self._h, seqs.ctypes.data_as(C.c_void_p), offs.ctypes.data_as(C.c_void_p),
                                                 reads.n, C.byref(p), n_threads))

    def __del__(self):
        try:
            if self._h:
                lib().orc_index_free(self._h)
                self._h = None
        except Exception:
            pass


def score_anchor(a, b, seed_length, max_gap):
    """a, b = (qb, qe, tb, te, f); positions may carry REV_BIT."""
    return lib().orc_score_anchor(a[0], a[1], a[2], a[3], a[4], b[0], b[1], b[2], b[3], seed_length, max_gap)


def poa_align(labels, edges, query, params=None):
    p = params or default_params()
    cat = "".join(labels).encode()
    off = np.zeros(len(labels) + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(x) for x in labels])
    ed = np.ascontiguousarray(np.array(edges, dtype=np.uint32).reshape(-1, 2))
    return OracleResults(lib().orc_poa_align(cat, off.ctypes.data_as(C.c_void_p), len(labels), ed.ctypes.data_as(C.c_void_p),
                                             len(edges), query.encode(), len(query), C.byref(p)))
