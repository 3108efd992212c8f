// K5 — banded sequence-to-graph POA, replacing the abPOA call at the reference's FFI seam
// (src/align.rs:190-202, AbpoaAligner::create_align_safe(nodes, edges, read, Global)).
// Definition of the alignment (recurrences, adaptive band, tie-breaks): DESIGN.md "POA definition",
// identical to oracle/oracle.cpp poa_align().
//
// Fill kernel: one warp per (read, chain).  Rows (graph bases) are visited in subgraph node order,
// which is a topological order because only edges s<e are kept (src/align.rs:717-718).  Each lane
// owns C consecutive query columns.  FIXED instances (32*C > qlen) pin column j to lane j/C for the
// whole alignment; SLIDE instances (long reads) use a cyclic block layout (column j -> block j/C ->
// lane block%32) so the adaptive band can slide without moving data.  The previous row's H/E1'/E2'
// stay in registers for all rows inside a node label.  The first row of a node rebuilds them from
// the trace words of its predecessors' last rows.
//
// Trace: one 32-bit word per DP cell that holds the cell's complete state relative to H:
//   bits  0..15  H(i,j) - ref(i), clamped to int16; ref(i) is 0 in the FIXED instances (scores are
//                bounded by 6*(rows+qlen) there) and the previous row's maximum in the SLIDE instances,
//                which also verify that no reachable cell was clamped
//   bits 16..18  E1'(i,j) - (H(i,j) - o1 - e1)            in [0, o1]
//   bits 19..23  E2'(i,j) - (H(i,j) - o2 - e2)            in [0, o2]
//   bits 24..26  max(F1(i,j) - (H(i,j-1) - o1 - e1), -1) + 1   in [0, o1+1]
//   bits 27..31  max(F2(i,j) - (H(i,j-1) - o2 - e2), -1) + 1   in [0, o2+1]
// (bounds: E <= H, F(i,j) <= H(i,j-1) - e; proofs in DESIGN.md).  The traceback kernel re-derives
// H/E1'/E2'/F1/F2 from these words and runs abPOA's value-driven backtrack exactly as the oracle
// does, so no per-cell decision flags have to be computed in the fill loop.
// The horizontal gap recurrences (F1,F2) are prefix-max scans: serial inside a lane, Kogge-Stone
// across lanes with the constant decay e*C per lane.  Row maxima use REDUX.
// Roofline: INT32 ALU (no tensor cores: nothing here is a contraction).  Algorithmic work =
// band cells; see DESIGN.md for ops/cell.
#include <climits>
#include <cstdlib>

#include "common.cuh"
#include "internal.h"
#include "k_poa_trace.cuh"

namespace vga {

static const int POA_WARPS = 4;
static const u32 FULL = 0xffffffffu;
static const u64 POOL_CHUNK = 8192;   // trace words taken from the pool per refill

__device__ __forceinline__ i32 imax(i32 a, i32 b) { return a > b ? a : b; }
__device__ __forceinline__ i32 imin(i32 a, i32 b) { return a < b ? a : b; }
__device__ __forceinline__ i32 sext16(u32 w) { return (i32)(short)(w & 0xffffu); }

template <int C, bool SLIDE>
__global__ void __launch_bounds__(POA_WARPS * 32) poa_fill_kernel(DevIndex ix, PoaBuffers b) {
    const u32 lane = threadIdx.x & 31;
    const i32 e1 = b.e1, e2 = b.e2, oe1 = b.o1 + b.e1, oe2 = b.o2 + b.e2;
    const i32 smatch = b.match, smis = -b.mismatch;
    u64 pool_cur = 0, pool_end = 0;      // this warp's current chunk of the trace pool (words)
    for (;;) {
        u32 wi = 0;
        if (lane == 0) wi = atomicAdd(b.work_ctr, 1u);
        wi = __shfl_sync(FULL, wi, 0);
        if (wi >= b.n_items) break;
        const u32 w = b.work[wi];
        const AlignItem it = b.items[w];
        const i32 qlen = (i32)it.qlen;
        const u8* __restrict__ q = (const u8*)b.seqs + it.q_off;
        const i32 wband = b.wb + (i32)(b.wf * (float)qlen);
        const u32 nb = it.node_base, S = it.S;
        const u32* __restrict__ rs = b.nd_row_start + nb;
        const u32* __restrict__ poff = b.nd_pred_off + nb;
        const u32* __restrict__ pl = b.nd_pred + it.pred_base;
        NodeState* __restrict__ nst = b.nd_state + nb;
        RowInfo* __restrict__ rows = b.rows + it.row_base;
        const i32 remain_src = b.nd_remain_last[nb] + (i32)(rs[1] - rs[0]) - 1 + 1;
        const i32 end0 = band_round_end0(imin(qlen, imax(0, qlen - remain_src) + wband), b.bgran_mask, qlen);
        u32 status = VGA_ALN_OK;
        // FIXED instances keep absolute H as int16: every DP step loses at most max(mismatch, o1+e1), so
        // reachable scores are >= -that*(rows+qlen); longer alignments go to a SLIDE instance
        if (!SLIDE && (qlen >= 32 * C || (i64)imax(b.mismatch, oe1) * ((i64)it.nrows + qlen) > 32000)) status = IT_RETRY_WIDER;
        i32 mxrow = 0;                       // maximum of the previous row
        u64 cells = 0;
        i32 Hp[C], E1p[C], E2p[C];
        u32 qv[C];
        i32 qblk = INT_MIN, pbb = 0, ppb = 0, ppe = -1, lpos = 0, rpos = 0;
#pragma unroll
        for (int t = 0; t < C; t++) {
            Hp[t] = E1p[t] = E2p[t] = POA_NEG_INF;
            qv[t] = 0;
            if (!SLIDE) { const i32 j = (i32)lane * C + t; if (j >= 1 && j <= qlen) qv[t] = q[j - 1]; }
        }

        for (u32 s = 0; s < S && status == VGA_ALN_OK; s++) {
            const u32 len = rs[s + 1] - rs[s];
            const u32 src = b.nd_src[nb + s];
            const u32 po = poff[s], pn = poff[s + 1] - po;
            const i32 remain_last = b.nd_remain_last[nb + s];
            // ---- predecessor summary (lane x holds predecessor x of the current chunk of 32)
            i32 my_pb = INT_MAX, my_pe = -1, my_l = INT_MAX - 1, my_r = -1, my_mx = INT_MIN, my_ref = 0;
            u32 my_toff = 0, my_prow = 0, prow0 = 0;
            auto load_preds = [&](u32 x0) {
                my_pb = INT_MAX; my_pe = -1; my_l = INT_MAX - 1; my_r = -1; my_mx = INT_MIN; my_ref = 0; my_toff = 0;
                if (x0 + lane < pn) {
                    const u32 p = pl[po + x0 + lane];
                    const NodeState ns = nst[p];
                    my_pb = ns.beg; my_pe = ns.end; my_l = ns.lpos; my_r = ns.rpos; my_mx = ns.mx;
                    my_prow = rs[p + 1] - 1;
                    const RowInfo pr = rows[my_prow];
                    my_toff = pr.toff; my_ref = pr.ref;
                }
            };
            i32 f_mpl = 1, f_mpr = 1, f_min_pb = 0, f_max_pe = end0, f_mx = 0;   // virtual source row
            if (pn > 0) {
                f_mpl = INT_MAX; f_mpr = 0; f_min_pb = INT_MAX; f_max_pe = -1; f_mx = INT_MIN;
                for (u32 x0 = 0; x0 < pn; x0 += 32) {
                    load_preds(x0);
                    if (x0 == 0) prow0 = my_prow;           // lane 0: row of the first predecessor
                    f_mpl = imin(f_mpl, __reduce_min_sync(FULL, my_l + 1));
                    f_mpr = imax(f_mpr, __reduce_max_sync(FULL, my_r + 1));
                    f_min_pb = imin(f_min_pb, __reduce_min_sync(FULL, my_pb));
                    f_max_pe = imax(f_max_pe, __reduce_max_sync(FULL, my_pe));
                    if (SLIDE) f_mx = imax(f_mx, __reduce_max_sync(FULL, my_mx));
                }
            }
            u32 bases = 0;
            i32 beg = 0, end = 0;
            for (u32 o = 0; o < len; o++) {
                const bool first = o == 0;
                if ((o & 31) == 0) bases = (o + lane < len) ? (u32)ix.seq_both[src + o + lane] : 0u;
                const u32 base = __shfl_sync(FULL, bases, o & 31);
                // ---- adaptive band (abPOA GET_AD_DP_BEGIN/END + clipping to the predecessors' bands)
                i32 mpl, mpr, min_pb, max_pe;
                if (!first) { mpl = lpos + 1; mpr = rpos + 1; min_pb = ppb; max_pe = ppe; }
                else { mpl = f_mpl; mpr = f_mpr; min_pb = f_min_pb; max_pe = f_max_pe; }
                const i32 r = qlen - (remain_last + (i32)(len - 1 - o));
                beg = imax(0, imin(mpl, r) - wband);
                end = imin(qlen, imax(mpr, r) + wband);
                beg = imax(beg, min_pb);
                end = imin(end, max_pe + 1);
                band_round(beg, end, b.bgran_mask, qlen, max_pe);
                if (beg > end) status = VGA_ALN_BAND_ERR;
                const i32 W = end - beg + 1;
                const i32 cbb = SLIDE ? beg / C : 0;
                // the 32 lanes cover columns [cbb*C, (cbb+32)*C): the band must fit from its block-aligned start
                if (SLIDE && status == VGA_ALN_OK && end - cbb * C + 1 > 32 * C) status = IT_RETRY_WIDER;
                // ---- pool space: one trace word per band cell, rows start on 16-byte boundaries
                const u64 need = (u64)((W + (beg & 7) + 3) & ~3);          // word 0 of a row is column beg & ~7 (trace_col0)
                if (status == VGA_ALN_OK && pool_cur + need > pool_end) {
                    unsigned long long got = 0;
                    const u64 take = need > POOL_CHUNK ? need : POOL_CHUNK;
                    if (lane == 0) got = atomicAdd(b.pool_ctr, (unsigned long long)take);
                    got = __shfl_sync(FULL, got, 0);
                    if (got + take > b.pool_words) status = IT_RETRY_POOL;
                    else { pool_cur = got; pool_end = got + take; }
                }
                if (status != VGA_ALN_OK) break;
                u32* __restrict__ tr = b.pool + pool_cur;
                const i32 ref = SLIDE ? (first ? f_mx : mxrow) : 0;
                if (lane == 0) {
                    RowInfo ri;
                    ri.beg = beg; ri.end = end; ri.toff = (u32)(pool_cur >> 2); ri.ref = ref;
                    ri.prow = first ? prow0 : rs[s] + o - 1;
                    ri.meta = (first ? (pn < 0xffffffu ? pn : 0xffffffu) : 1u) | (base << 24);
                    ri.node = s; ri.off = o;
                    rows[rs[s] + o] = ri;
                }
                pool_cur += need;
                cells += (u64)W;
                // ---- column mapping of this row
                const i32 cbl = SLIDE ? cbb + (i32)((lane - (u32)cbb) & 31u) : (i32)lane;
                const i32 ll = cbl - cbb;                 // logical lane
                const i32 j0 = cbl * C;
                if (SLIDE && cbl != qblk) {
#pragma unroll
                    for (int t = 0; t < C; t++) { const i32 j = j0 + t; qv[t] = (j >= 1 && j <= qlen) ? q[j - 1] : 0u; }
                    qblk = cbl;
                }
                i32 hleft_old;
                const u32 src_lane = (lane + 31) & 31;
                if (first) {
                    // combined predecessor row: H at columns j0-1..j0+C-1, E1'/E2' at j0..j0+C-1
                    hleft_old = POA_NEG_INF;
#pragma unroll
                    for (int t = 0; t < C; t++) { Hp[t] = E1p[t] = E2p[t] = POA_NEG_INF; }
                    if (pn == 0) {
#pragma unroll
                        for (int t = -1; t < C; t++) {
                            const i32 j = j0 + t;
                            if (j >= 0 && j <= end0) {
                                const i32 hv = j == 0 ? 0 : imax(-(b.o1 + b.e1 * j), -(b.o2 + b.e2 * j));
                                if (t < 0) hleft_old = hv;
                                else { Hp[t < 0 ? 0 : t] = hv; if (j == 0) { E1p[t < 0 ? 0 : t] = -oe1; E2p[t < 0 ? 0 : t] = -oe2; } }
                            }
                        }
                    } else {
                        __syncwarp();       // trace words of earlier rows were written by other lanes
                        for (u32 x0 = 0; x0 < pn; x0 += 32) {
                            if (pn > 32) load_preds(x0);
                            const u32 nh = imin(32, (i32)(pn - x0));
                            for (u32 x = 0; x < nh; x++) {
                                const i32 pb = __shfl_sync(FULL, my_pb, x), pe = __shfl_sync(FULL, my_pe, x);
                                const i32 pref = SLIDE ? __shfl_sync(FULL, my_ref, x) : 0;
                                const u32* __restrict__ tw = b.pool + (u64)__shfl_sync(FULL, my_toff, x) * 4;
                                u32 wv[C + 1];
#pragma unroll
                                for (int t = -1; t < C; t++) {
                                    const i32 j = j0 + t;
                                    wv[t + 1] = (j >= pb && j <= pe) ? tw[j - trace_col0(pb)] : 0xffff8000u;
                                }
#pragma unroll
                                for (int t = -1; t < C; t++) {
                                    const i32 j = j0 + t;
                                    if (j >= pb && j <= pe) {
                                        const u32 v = wv[t + 1];
                                        const i32 hv = pref + sext16(v);
                                        if (t < 0) hleft_old = imax(hleft_old, hv);
                                        else {
                                            Hp[t < 0 ? 0 : t] = imax(Hp[t < 0 ? 0 : t], hv);
                                            E1p[t < 0 ? 0 : t] = imax(E1p[t < 0 ? 0 : t], hv - oe1 + (i32)((v >> 16) & 7u));
                                            E2p[t < 0 ? 0 : t] = imax(E2p[t < 0 ? 0 : t], hv - oe2 + (i32)((v >> 19) & 31u));
                                        }
                                    }
                                }
                            }
                        }
                    }
                    ppb = min_pb; ppe = max_pe;
                } else if (SLIDE) {
                    const i32 lv = __shfl_sync(FULL, Hp[C - 1], src_lane);
                    const i32 src_pblk = pbb + (i32)((src_lane - (u32)pbb) & 31u);
                    hleft_old = src_pblk == cbl - 1 ? lv : POA_NEG_INF;
                    const i32 pbl = pbb + (i32)((lane - (u32)pbb) & 31u);
                    if (pbl != cbl) {
#pragma unroll
                        for (int t = 0; t < C; t++) { Hp[t] = E1p[t] = E2p[t] = POA_NEG_INF; }
                    }
                } else {
                    hleft_old = __shfl_sync(FULL, Hp[C - 1], src_lane);
                }
                if (j0 == 0) hleft_old = POA_DEEP;       // column -1: M(i,0) does not exist
                const i32 tlo = beg - j0, thi = end - j0;     // in band iff tlo <= t <= thi
                // ---- sweep 1: Hhat = max(M, E1, E2) and the lane-local part of the F scans
                i32 hh[C], f1l[C], f2l[C];
                i32 run1 = POA_DEEP, run2 = POA_DEEP;
#pragma unroll
                for (int t = 0; t < C; t++) {
                    const i32 hd = t == 0 ? hleft_old : Hp[t - 1];
                    const i32 sc = qv[t] == base ? smatch : smis;
                    hh[t] = imax(imax(hd + sc, E1p[t]), E2p[t]);
                    const i32 hg = (t >= tlo && t <= thi) ? hh[t] : POA_DEEP;
                    f1l[t] = run1; f2l[t] = run2;
                    run1 = imax(run1 - e1, hg - oe1);
                    run2 = imax(run2 - e2, hg - oe2);
                }
                // ---- cross-lane carry of F1/F2 (Kogge-Stone in logical lane order)
                i32 v1 = run1, v2 = run2;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const i32 t1 = __shfl_sync(FULL, v1, (lane - d) & 31) - e1 * C * d;
                    const i32 t2 = __shfl_sync(FULL, v2, (lane - d) & 31) - e2 * C * d;
                    if (ll >= d) { v1 = imax(v1, t1); v2 = imax(v2, t2); }
                }
                i32 cin1 = __shfl_sync(FULL, v1, src_lane), cin2 = __shfl_sync(FULL, v2, src_lane);
                if (ll == 0) { cin1 = POA_DEEP; cin2 = POA_DEEP; }
                // ---- sweep 2: H, E', trace words, masking, row maximum
                u32 wd[C];
                i32 b1p = 0, b2p = 0, f1_0 = 0, f2_0 = 0;
                i32 lmax = INT_MIN, lfirst = INT_MAX, llast = -1, lmin = INT_MAX;
#pragma unroll
                for (int t = 0; t < C; t++) {
                    const i32 j = j0 + t;
                    const bool inb = t >= tlo && t <= thi;
                    i32 f1 = imax(f1l[t], cin1 - e1 * t), f2 = imax(f2l[t], cin2 - e2 * t);
                    if (t == tlo) { f1 = POA_NEG_INF; f2 = POA_NEG_INF; }       // F(i,beg) does not exist
                    const i32 h = imax(hh[t], imax(f1, f2));
                    const i32 b1 = h - oe1, b2 = h - oe2;
                    const i32 n1 = imax(E1p[t] - e1, b1), n2 = imax(E2p[t] - e2, b2);
                    u32 wv = ((u32)imax(h - ref, -32768) & 0xffffu) + ((u32)(n1 - b1) << 16) + ((u32)(n2 - b2) << 19);
                    if (SLIDE && inb && h > -(1 << 27)) lmin = imin(lmin, h);   // reachable cells only
                    if (t > 0) wv += ((u32)imax(f1 - b1p + 1, 0) << 24) + ((u32)imax(f2 - b2p + 1, 0) << 27);
                    else { f1_0 = f1; f2_0 = f2; }
                    wd[t] = wv;
                    b1p = b1; b2p = b2;
                    Hp[t] = inb ? h : POA_NEG_INF;
                    E1p[t] = inb ? n1 : POA_NEG_INF;
                    E2p[t] = inb ? n2 : POA_NEG_INF;
                    if (inb) {
                        if (h > lmax) { lmax = h; lfirst = j; }
                        if (h >= lmax) llast = j;
                    }
                }
                {   // F fields of the lane's first column need the left lane's new last H
                    const i32 hl = __shfl_sync(FULL, Hp[C - 1], src_lane);
                    if (ll > 0) wd[0] += ((u32)imax(f1_0 - (hl - oe1) + 1, 0) << 24) + ((u32)imax(f2_0 - (hl - oe2) + 1, 0) << 27);
                }
#pragma unroll
                for (int t = 0; t < C; t++) {
                    if (t >= tlo && t <= thi) tr[j0 + t - trace_col0(beg)] = wd[t];
                }
                // ---- row maximum and its leftmost / rightmost column (abPOA simd_abpoa_ada_max_i)
                const i32 mx = __reduce_max_sync(FULL, lmax);
                lpos = __reduce_min_sync(FULL, lmax == mx ? lfirst : INT_MAX);
                rpos = __reduce_max_sync(FULL, lmax == mx ? llast : -1);
                pbb = cbb; ppb = beg; ppe = end; mxrow = mx;
                // a reachable cell further than int16 below the row reference cannot be encoded
                if (SLIDE && __any_sync(FULL, lmin != INT_MAX && lmin - ref < -32000)) status = VGA_ALN_UNSUPPORTED;
            }
            if (status != VGA_ALN_OK) break;
            // ---- the node's last row may be a predecessor of later nodes' first rows
            if (lane == 0) {
                NodeState ns;
                ns.beg = beg; ns.end = end; ns.lpos = lpos; ns.rpos = rpos; ns.mx = mxrow; ns.pad0 = ns.pad1 = ns.pad2 = 0;
                nst[s] = ns;
            }
            __syncwarp();       // NodeState / RowInfo are read by other lanes when a later node loads its predecessors
        }
        // ---- global end: sink predecessors (nodes without successors, ascending) that reach column qlen
        i32 best = INT_MIN;
        u32 bs = 0xffffffffu;
        if (status == VGA_ALN_OK) {
            __syncwarp();
            for (u32 s0 = 0; s0 < S; s0 += 32) {
                const u32 s = s0 + lane;
                i32 h = INT_MIN;
                if (s < S && b.nd_first_succ[nb + s] < 0) {
                    const RowInfo ri = rows[rs[s + 1] - 1];
                    if (ri.end >= qlen) h = ri.ref + sext16(b.pool[(u64)ri.toff * 4 + (u32)(qlen - trace_col0(ri.beg))]);
                }
                const i32 mx = __reduce_max_sync(FULL, h);
                if (mx > best) {                             // strict: the first node wins ties
                    best = mx;
                    bs = __reduce_min_sync(FULL, h == mx ? s : 0xffffffffu);
                }
            }
            if (bs == 0xffffffffu) status = VGA_ALN_EMPTY;
        }
        if (lane == 0) {
            AlignItem* o = b.items + w;
            o->status = status;
            o->score = status == VGA_ALN_OK ? best : 0;
            o->end_node = bs;
            o->cells = cells;
            o->n_runs = 0; o->n_path = 0; o->path_len = 0; o->start_off = 0; o->end_off = 0; o->n_aligned = 0;
        }
    }
}

// `spread` = threads per alignment (a power of two <= 32): only the first of them works.  With few, long
// alignments a warp of 32 different tracebacks would serialise its divergent steps; spreading them over
// warps trades idle lanes (there are plenty) for independent progress.  The traceback itself: k_poa_trace.cuh.
template <bool FAST>
__global__ void __launch_bounds__(128) poa_trace_kernel(DevIndex ix, TraceBuffers b, u32 spread) {
    const u32 tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid & (spread - 1)) return;
    const u32 idx = tid / spread;
    if (idx >= b.n_items) return;
    poa_trace_one<FAST>(b, b.work[idx]);
}

// classes: 0 = short reads on the 16x2 kernel (k_poa_pair.cu: qlen < 246, default gap model, int16 absolute scores bounded
// by the row count), 1 = the same kernel with a per-row range check for longer subgraphs, 2 = short reads with 32 lanes
// (k_poa_short.cu, any gap model, reads < 256), 3..7 = SLIDE C = 8, 12, 16, 24, 32 (bands up to 256 / 384 / 512 / 768 / 1024 columns),
// 8 = bands of any width (k_poa_wide.cu).
// An item that does not fit its class moves to the next.
// VGA_POA_PAIR=0 restores the 8 / 16 / 32-lane kernels of k_poa_short.cu on classes 0..2 (A/B measurements).
static int g_pair = getenv("VGA_POA_PAIR") ? atoi(getenv("VGA_POA_PAIR")) : 1;
// Class 1, the range-checked 16x2 instance, re-centres its int16 planes on the row maximum (RowInfo.ref), so subgraphs thousands
// of rows longer than their read (17-DRB1-3123-smooth: 1.5 k items per 125 k reads) stay inside int16 — but it is off by default:
// measured 125 ms against 74 ms on the SLIDE kernel for those items (gpurun_out/r02l).  Few items mean two or three warps per SM,
// so what counts is the latency of ONE row, and an 8-lane step of ~550 dependent instructions is longer than a 32-lane SLIDE row.
// VGA_POA_CHK=1 enables it (read per batch: the tests run both).
static int g_pair_chk = 0, g_force_wide = 0;
void poa_begin_batch() {
    const char* e = getenv("VGA_POA_CHK");
    g_pair_chk = e ? atoi(e) : 0;
    e = getenv("VGA_POA_WIDE");
    g_force_wide = e ? atoi(e) == 1 : 0;
}
// long reads (and everything else that is not a short-read class) run on the shared-memory kernel of k_poa_long.cu as
// class 3; VGA_POA_LONG=0 restores the register-resident SLIDE instances on classes 3..7 (A/B measurements)
// The shared-memory kernel wins on throughput (work proportional to the band, 64 registers), the register-resident
// SLIDE instances on the latency of one alignment: a batch with only a few long alignments (the very long subgraphs of
// 17-DRB1-3123-smooth: 1.5 k items, ~100 ms tail) stays on SLIDE.  Decided once per batch before classification and
// carried in the call's PoaBuffers (no process-wide state: two host threads may map on two indexes at once).
int poa_long_mode(u64 n_long_items, int n_sms) {
    const char* e = getenv("VGA_POA_LONG");            // read per batch: tests force either family
    return e ? atoi(e) : (n_long_items >= (u64)n_sms * 24 ? 1 : 0);
}
static int g_short_w8 = getenv("VGA_POA_W8") ? atoi(getenv("VGA_POA_W8")) : 0;
int poa_class_count() { return 9; }
bool poa_pair_active() { return g_pair != 0; }
int poa_class_for(i64 qlen, i64 nrows, i64 west, const VgaParams& P, int long_mode) {
    if (g_force_wide) return 8;                                                      // tests (VGA_POA_WIDE=1): every alignment on the CTA-per-alignment kernel
    if (g_pair) {
        if (poa_pair_ok(qlen, nrows, P)) return 0;
        if (g_pair_chk && poa_pair_checked_ok(qlen, nrows, P)) return 1;
        if (poa_short_ok(qlen, nrows, P)) return 2;
    } else if (poa_short_ok(qlen, nrows, P)) return west <= (i64)g_short_w8 ? 0 : west <= 160 ? 1 : 2;
    if (long_mode) return 3;
    // SLIDE instances compute all 32*C columns of every row, so an item goes to the narrowest one its estimate allows
    if (west <= 256 - 8) return 3;
    if (west <= 384 - 12) return 4;
    if (west <= 512 - 16) return 5;
    if (west <= 768 - 24) return 6;
    return 7;
}
// class 8 = bands wider than 1024 columns, one CTA per alignment (k_poa_wide.cu)
int poa_class_wider(int cls, int long_mode) { return cls >= 8 ? -1 : (long_mode && cls >= 3) ? 8 : cls + 1; }

template <int C, bool SLIDE>
static void launch_fill(const DevIndex& ix, const PoaBuffers& b, int n_sms, cudaStream_t st) {
    // persistent warps pulling work from a counter; enough CTAs to fill every SM
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, poa_fill_kernel<C, SLIDE>, POA_WARPS * 32, 0);
    if (per_sm < 1) per_sm = 1;
    long long ctas = (long long)n_sms * per_sm;
    long long need = ((long long)b.n_items + POA_WARPS - 1) / POA_WARPS;
    if (ctas > need) ctas = need;
    if (ctas < 1) ctas = 1;
    poa_fill_kernel<C, SLIDE><<<(unsigned)ctas, POA_WARPS * 32, 0, st>>>(ix, b);
}

int launch_poa_fill(const DevIndex& ix, const PoaBuffers& b, int cls, int n_sms, cudaStream_t st, const TraceBuffers* fused_tb) {
    if (b.n_items == 0) return 0;
    switch (cls) {
        case 0: case 1: if (g_pair) return launch_poa_fill_pair(ix, b, cls, n_sms, st);
        // fall through
        case 2: return launch_poa_fill_short(ix, b, cls, n_sms, st);
        case 3: if (b.long_mode) return launch_poa_fill_long(ix, b, n_sms, st, fused_tb);
                launch_fill<8, true>(ix, b, n_sms, st); break;
        case 4: launch_fill<12, true>(ix, b, n_sms, st); break;
        case 5: launch_fill<16, true>(ix, b, n_sms, st); break;
        case 6: launch_fill<24, true>(ix, b, n_sms, st); break;
        case 7: launch_fill<32, true>(ix, b, n_sms, st); break;
        default: return launch_poa_fill_wide(ix, b, n_sms, st);
    }
    return 1;
}

int launch_poa_trace(const DevIndex& ix, const TraceBuffers& b, cudaStream_t st) {
    if (b.n_items == 0) return 0;
    // enough tracebacks to fill the machine: one per thread; otherwise spread them out (up to one per warp)
    u32 spread = 1;
    while (spread < 32 && (u64)b.n_items * spread * 2 <= 148ull * 2048) spread <<= 1;
    if (const char* e = getenv("VGA_TRACE_SPREAD")) { const int v = atoi(e); if (v >= 1 && v <= 32 && (v & (v - 1)) == 0) spread = (u32)v; }
    const u64 threads = (u64)b.n_items * spread;
    // Tried and dropped: L2 prefetch of the diagonal trace word 2-8 rows ahead (consecutive rows lie back to back in a pool
    // chunk, so the stride is predictable): C2 3.44 -> 3.55-3.70 ms, C4 14.0 -> 18.6-19.8 ms, C5 unchanged (gpurun_out/ab13).
    // Tried and dropped (round 2): storing the fill kernel's int16 planes (H, E1', E2'; 6 B per cell) as the trace of short reads
    // and re-deriving F in the traceback.  28 % fewer instructions in the fill (2.5 -> 1.8 warp-instructions per cell) but no
    // time gained — the fill is bound by dependency latency at 16 warps per SM, not by issue — and the traceback of the wider
    // rows took 36 instead of 23 ms per 2 M reads (gpurun_out/r02h, r02i).
    if (spread >= 4) poa_trace_kernel<true><<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(ix, b, spread);
    else poa_trace_kernel<false><<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(ix, b, spread);
    return 1;
}

// ---------------------------------------------------------------------------------------------
// INT32 ALU peak: independent IADD / VIMNMX streams, the instruction mix of the DP inner loop.
// ---------------------------------------------------------------------------------------------
__global__ void int32_peak_kernel(int iters, int* out) {
    int a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const int c = blockIdx.x + 3;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            a0 = max(a0 + c, a1); a1 = max(a1 + c, a2); a2 = max(a2 + c, a3); a3 = max(a3 + c, a4);
            a4 = max(a4 + c, a5); a5 = max(a5 + c, a6); a6 = max(a6 + c, a7); a7 = max(a7 + c, a0);
        }
    }
    if (a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 == 0x7fffffff) *out = a0;
}

double measure_int32_peak() {
    int* d = nullptr;
    VGA_CUDA(cudaMalloc(&d, 4));
    const int iters = 4096, blocks = 148 * 8, threads = 256;
    cudaEvent_t e0, e1;
    VGA_CUDA(cudaEventCreate(&e0)); VGA_CUDA(cudaEventCreate(&e1));
    int32_peak_kernel<<<blocks, threads>>>(64, d);
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        VGA_CUDA(cudaEventRecord(e0));
        int32_peak_kernel<<<blocks, threads>>>(iters, d);
        VGA_CUDA(cudaEventRecord(e1));
        VGA_CUDA(cudaEventSynchronize(e1));
        float ms; VGA_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    // per inner step: 8 adds + 8 max = 16 INT32 ops per thread (ptxas fuses each pair into one VIADDMNMX)
    double ops = (double)iters * 16.0 * 16.0 * (double)blocks * threads;
    return ops / (best * 1e-3);
}

}  // namespace vga
