// K5 — banded sequence-to-graph POA, replacing the abPOA call at the reference's FFI seam
// (src/align.rs:190-202, AbpoaAligner::create_align_safe(nodes, edges, read, Global)).
// Definition of the alignment (recurrences, adaptive band, tie-breaks): DESIGN.md "POA definition",
// identical to oracle/oracle.cpp poa_align().
//
// Fill kernel: one warp per (read, chain).  Rows (graph bases) are visited in subgraph node order,
// which is a topological order because only edges s<e are kept (src/align.rs:717-718).  Each lane
// owns C consecutive query columns of a cyclic block layout (column j -> block j/C -> lane
// block%32), so the adaptive band can slide without moving data: the previous row's H/E1/E2 stay
// in registers for all rows inside a node label; only the last row of every node is saved to the
// pool, because only those rows can be predecessors of another node's first row.
// The horizontal gap recurrences (F1,F2) are prefix-max scans: serial inside a lane, Kogge-Stone
// across lanes with the constant decay e*C per lane.  Row maxima use REDUX.  Traceback state is a
// 12-bit flag word per cell (u16) computed from the same equalities abPOA's value-driven
// backtrack evaluates, so the traceback kernel reproduces its decisions without the H/E/F matrices.
// Roofline: INT32 ALU (no tensor cores: nothing here is a contraction).  Algorithmic work =
// band cells; see DESIGN.md for ops/cell.
#include <climits>

#include "common.cuh"
#include "internal.h"

namespace vga {

static const int POA_WARPS = 4;
static const u32 FULL = 0xffffffffu;
static const u64 POOL_CHUNK = 16384;   // cells taken from the pool per refill

enum { M_OP = 1, E1_OP = 2, E2_OP = 4, F1_OP = 8, F2_OP = 16, ALL_OP = 31 };

__device__ __forceinline__ i32 imax(i32 a, i32 b) { return a > b ? a : b; }
__device__ __forceinline__ i32 imin(i32 a, i32 b) { return a < b ? a : b; }

template <int C>
__global__ void __launch_bounds__(POA_WARPS * 32) poa_fill_kernel(DevIndex ix, PoaBuffers b) {
    const u32 lane = threadIdx.x & 31;
    const i32 e1 = b.e1, e2 = b.e2, oe1 = b.o1 + b.e1, oe2 = b.o2 + b.e2;
    u64 pool_cur = 0, pool_end = 0;      // this warp's current chunk of the trace pool (cells)
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
        const i32 remain_src = b.nd_remain_last[nb] + (i32)(rs[1] - rs[0]) - 1 + 1;
        const i32 end0 = imin(qlen, imax(0, qlen - remain_src) + wband);
        u32 status = VGA_ALN_OK;
        u64 cells = 0;
        i32 Hp[C], E1p[C], E2p[C];
        u32 qv[C];
        i32 qblk = INT_MIN, pbb = 0, ppb = 0, ppe = -1, lpos = 0, rpos = 0;
#pragma unroll
        for (int t = 0; t < C; t++) { Hp[t] = E1p[t] = E2p[t] = POA_NEG_INF; qv[t] = 0; }

        // value of the combined predecessor row of node s at column j (first rows only)
        auto pred_vals = [&](u32 po, u32 pn, i32 j, i32& hv, i32& x1, i32& x2) {
            hv = POA_NEG_INF; x1 = POA_NEG_INF; x2 = POA_NEG_INF;
            if (pn == 0) {                       // virtual source row
                if (j >= 0 && j <= end0) {
                    hv = j == 0 ? 0 : imax(-(b.o1 + b.e1 * j), -(b.o2 + b.e2 * j));
                    if (j == 0) { x1 = -oe1; x2 = -oe2; }
                }
                return;
            }
            for (u32 x = 0; x < pn; x++) {
                const u32 p = nb + pl[po + x];
                const i32 pb = b.nd_beg[p], pe = b.nd_end[p];
                if (j >= pb && j <= pe) {
                    const i32* base = (const i32*)(b.pool + (u64)b.nd_store[p] * 4);
                    const i32 Wp = pe - pb + 1;
                    hv = imax(hv, base[j - pb]);
                    x1 = imax(x1, base[Wp + j - pb]);
                    x2 = imax(x2, base[2 * Wp + j - pb]);
                }
            }
        };

        for (u32 s = 0; s < S && status == VGA_ALN_OK; s++) {
            const u32 len = rs[s + 1] - rs[s];
            const u32 src = b.nd_src[nb + s];
            const u32 po = poff[s], pn = poff[s + 1] - po;
            const i32 remain_last = b.nd_remain_last[nb + s];
            i32 beg = 0, end = 0;
            for (u32 o = 0; o < len; o++) {
                const bool first = o == 0;
                const u32 base = ix.seq_both[src + o];
                // ---- adaptive band (abPOA GET_AD_DP_BEGIN/END + clipping to the predecessors' bands)
                i32 mpl, mpr, min_pb, max_pe;
                if (!first) { mpl = lpos + 1; mpr = rpos + 1; min_pb = ppb; max_pe = ppe; }
                else if (pn == 0) { mpl = 1; mpr = 1; min_pb = 0; max_pe = end0; }
                else {
                    mpl = INT_MAX; mpr = 0; min_pb = INT_MAX; max_pe = -1;
                    for (u32 x = 0; x < pn; x++) {
                        const u32 p = nb + pl[po + x];
                        mpl = imin(mpl, b.nd_lpos[p] + 1); mpr = imax(mpr, b.nd_rpos[p] + 1);
                        min_pb = imin(min_pb, b.nd_beg[p]); max_pe = imax(max_pe, b.nd_end[p]);
                    }
                    if (pn > 254) status = VGA_ALN_UNSUPPORTED;
                }
                const i32 r = qlen - (remain_last + (i32)(len - 1 - o));
                beg = imax(0, imin(mpl, r) - wband);
                end = imin(qlen, imax(mpr, r) + wband);
                beg = imax(beg, min_pb);
                end = imin(end, max_pe + 1);
                if (beg > end) status = VGA_ALN_BAND_ERR;
                const i32 W = end - beg + 1;
                // the 32 lanes cover columns [cbb*C, (cbb+32)*C): the band must fit from its block-aligned start
                if (status == VGA_ALN_OK && end - (beg / C) * C + 1 > 32 * C) status = IT_RETRY_WIDER;
                // ---- pool space: W trace cells (+2W for the per-predecessor bytes of a first row)
                const u64 Wr = (u64)((W + 3) & ~3);
                const u64 need = first ? 3 * Wr : Wr;
                if (status == VGA_ALN_OK && pool_cur + need > pool_end) {
                    unsigned long long got = 0;
                    const u64 take = need > POOL_CHUNK ? need : POOL_CHUNK;
                    if (lane == 0) got = atomicAdd(b.pool_ctr, (unsigned long long)take);
                    got = __shfl_sync(FULL, got, 0);
                    if (got + take > b.pool_cells) status = IT_RETRY_POOL;
                    else { pool_cur = got; pool_end = got + take; }
                }
                if (status != VGA_ALN_OK) break;
                u16* __restrict__ tr = b.pool + pool_cur;
                u32* __restrict__ ptr = (u32*)(b.pool + pool_cur + Wr);
                if (lane == 0) {
                    b.row_beg[it.row_base + rs[s] + o] = beg;
                    b.row_toff[it.row_base + rs[s] + o] = (u32)(pool_cur >> 2);
                    if (first) b.nd_ptoff[nb + s] = (u32)((pool_cur + Wr) >> 2);
                }
                pool_cur += need;
                cells += (u64)W;
                // ---- column mapping of this row
                const i32 cbb = beg / C;
                const i32 cbl = cbb + (i32)((lane - (u32)cbb) & 31u);
                const i32 ll = cbl - cbb;                 // logical lane
                const i32 j0 = cbl * C;
                if (cbl != qblk) {
#pragma unroll
                    for (int t = 0; t < C; t++) { const i32 j = j0 + t; qv[t] = (j >= 1 && j <= qlen) ? q[j - 1] : 0u; }
                    qblk = cbl;
                }
                i32 hleft_old;
                const u32 src_lane = (lane + 31) & 31;
                if (first) {
                    i32 d1, d2;
                    pred_vals(po, pn, j0 - 1, hleft_old, d1, d2);
#pragma unroll
                    for (int t = 0; t < C; t++) pred_vals(po, pn, j0 + t, Hp[t], E1p[t], E2p[t]);
                    ppb = min_pb; ppe = max_pe;
                } else {
                    const i32 lv = __shfl_sync(FULL, Hp[C - 1], src_lane);
                    const i32 src_pblk = pbb + (i32)((src_lane - (u32)pbb) & 31u);
                    hleft_old = src_pblk == cbl - 1 ? lv : POA_NEG_INF;
                    const i32 pbl = pbb + (i32)((lane - (u32)pbb) & 31u);
                    if (pbl != cbl) {
#pragma unroll
                        for (int t = 0; t < C; t++) { Hp[t] = E1p[t] = E2p[t] = POA_NEG_INF; }
                    }
                }
                // ---- sweep 1: Hhat = max(M, E1, E2) and the lane-local part of the F scans
                i32 hh[C], f1l[C], f2l[C];
                i32 run1 = POA_DEEP, run2 = POA_DEEP;
#pragma unroll
                for (int t = 0; t < C; t++) {
                    const i32 j = j0 + t;
                    const i32 hd = t == 0 ? hleft_old : Hp[t - 1];
                    const i32 sc = qv[t] == base ? b.match : -b.mismatch;
                    const i32 m = j >= 1 ? hd + sc : POA_NEG_INF;
                    hh[t] = imax(m, imax(E1p[t], E2p[t]));
                    const i32 hg = (j >= beg && j <= end) ? hh[t] : POA_DEEP;
                    f1l[t] = run1; f2l[t] = run2;
                    run1 = imax(hg - oe1, run1 - e1);
                    run2 = imax(hg - oe2, run2 - e2);
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
                // ---- sweep 2: H, E', flags, masking, row maximum
                u32 fl[C];
                i32 oldh_prev = hleft_old, newh_prev = POA_NEG_INF, f1_0 = 0, f2_0 = 0;
                i32 lmax = INT_MIN, lfirst = INT_MAX, llast = -1;
#pragma unroll
                for (int t = 0; t < C; t++) {
                    const i32 j = j0 + t;
                    const bool inb = j >= beg && j <= end;
                    i32 f1 = imax(f1l[t], cin1 - e1 * t), f2 = imax(f2l[t], cin2 - e2 * t);
                    if (j == beg) { f1 = POA_NEG_INF; f2 = POA_NEG_INF; }
                    const i32 h = imax(hh[t], imax(f1, f2));
                    const i32 x1 = E1p[t], x2 = E2p[t];
                    const i32 n1 = imax(x1 - e1, h - oe1), n2 = imax(x2 - e2, h - oe2);
                    const bool ismatch = qv[t] == base;
                    const i32 m = j >= 1 ? oldh_prev + (ismatch ? b.match : -b.mismatch) : POA_NEG_INF;
                    const bool pin_m = j - 1 >= ppb && j - 1 <= ppe && j >= 1;
                    const bool pin_e = j >= ppb && j <= ppe;
                    u32 f = 0;
                    if (pin_m && m == h) f |= TB_M;
                    if (pin_e && x1 == h) f |= TB_E1;
                    if (pin_e && x2 == h) f |= TB_E2;
                    if (h == f1) f |= TB_HF1;
                    if (h == f2) f |= TB_HF2;
                    if (h - oe1 == n1) f |= TB_E1OPEN;
                    if (h - oe2 == n2) f |= TB_E2OPEN;
                    if (ismatch) f |= TB_MATCH;
                    if (pin_e && n1 == x1 - e1) f |= TB_E1EXT;
                    if (pin_e && n2 == x2 - e2) f |= TB_E2EXT;
                    if (t > 0) {
                        if (newh_prev - oe1 == f1) f |= TB_F1OPEN;
                        if (newh_prev - oe2 == f2) f |= TB_F2OPEN;
                    } else { f1_0 = f1; f2_0 = f2; }
                    fl[t] = f;
                    oldh_prev = Hp[t];
                    Hp[t] = inb ? h : POA_NEG_INF;
                    E1p[t] = inb ? n1 : POA_NEG_INF;
                    E2p[t] = inb ? n2 : POA_NEG_INF;
                    newh_prev = Hp[t];
                    if (inb) {
                        if (h > lmax) { lmax = h; lfirst = j; }
                        if (h >= lmax) llast = j;
                    }
                }
                {   // F-open flags of the lane's first column need the left lane's new last H
                    const i32 hl = __shfl_sync(FULL, Hp[C - 1], src_lane);
                    if (ll > 0) {
                        if (hl - oe1 == f1_0) fl[0] |= TB_F1OPEN;
                        if (hl - oe2 == f2_0) fl[0] |= TB_F2OPEN;
                    }
                }
                if (first) {
                    // exact per-predecessor decisions in list order (backtrack tries predecessors in order)
#pragma unroll
                    for (int t = 0; t < C; t++) {
                        const i32 j = j0 + t;
                        if (j < beg || j > end) continue;
                        const i32 h = Hp[t], n1 = E1p[t], n2 = E2p[t];
                        const i32 sc = qv[t] == base ? b.match : -b.mismatch;
                        u32 mp = 0xff, ep = 0xff, ex = 0, e1p = 0xff, e2p = 0xff;
                        const u32 np = pn == 0 ? 1u : pn;
                        for (u32 x = 0; x < np; x++) {
                            i32 pb, pe, a1 = POA_NEG_INF, a2 = POA_NEG_INF, hm = POA_NEG_INF;
                            if (pn == 0) {
                                pb = 0; pe = end0;
                                if (j - 1 >= 0 && j - 1 <= end0) hm = j - 1 == 0 ? 0 : imax(-(b.o1 + b.e1 * (j - 1)), -(b.o2 + b.e2 * (j - 1)));
                                if (j == 0) { a1 = -oe1; a2 = -oe2; }
                            } else {
                                const u32 p = nb + pl[po + x];
                                pb = b.nd_beg[p]; pe = b.nd_end[p];
                                const i32* bs = (const i32*)(b.pool + (u64)b.nd_store[p] * 4);
                                const i32 Wp = pe - pb + 1;
                                if (j - 1 >= pb && j - 1 <= pe) hm = bs[j - 1 - pb];
                                if (j >= pb && j <= pe) { a1 = bs[Wp + j - pb]; a2 = bs[2 * Wp + j - pb]; }
                            }
                                                        const bool in_m = j >= 1 && j - 1 >= pb && j - 1 <= pe;
                            const bool in_e = j >= pb && j <= pe;
                            if (mp == 0xff && in_m && hm + sc == h) mp = x;
                            if (ep == 0xff && in_e) { if (a1 == h) { ep = x; ex = 1; } else if (a2 == h) { ep = x; ex = 2; } }
                            if (e1p == 0xff && in_e && n1 == a1 - e1) e1p = x;
                            if (e2p == 0xff && in_e && n2 == a2 - e2) e2p = x;
                        }
                        u32 f = fl[t] & ~(u32)(TB_M | TB_E1 | TB_E2 | TB_E1EXT | TB_E2EXT);
                        if (mp != 0xff) f |= TB_M;
                        if (ex == 1) f |= TB_E1;
                        if (ex == 2) f |= TB_E2;
                        if (e1p != 0xff) f |= TB_E1EXT;
                        if (e2p != 0xff) f |= TB_E2EXT;
                        fl[t] = f;
                        ptr[j - beg] = mp | (ep << 8) | (e1p << 16) | (e2p << 24);
                    }
                }
#pragma unroll
                for (int t = 0; t < C; t++) {
                    const i32 j = j0 + t;
                    if (j >= beg && j <= end) tr[j - beg] = (u16)fl[t];
                }
                // ---- row maximum and its leftmost / rightmost column (abPOA simd_abpoa_ada_max_i)
                const i32 mx = __reduce_max_sync(FULL, lmax);
                lpos = __reduce_min_sync(FULL, lmax == mx ? lfirst : INT_MAX);
                rpos = __reduce_max_sync(FULL, lmax == mx ? llast : -1);
                pbb = cbb; ppb = beg; ppe = end;
            }
            if (status != VGA_ALN_OK) break;
            // ---- save the node's last row: it may be a predecessor of later nodes' first rows
            {
                const i32 W = end - beg + 1;
                const u64 Wr = (u64)((W + 1) & ~1);
                const u64 need = ((6 * Wr) + 3) & ~3ull;       // 3 ints per column, in u16 cells
                if (pool_cur + need > pool_end) {
                    unsigned long long got = 0;
                    const u64 take = need > POOL_CHUNK ? need : POOL_CHUNK;
                    if (lane == 0) got = atomicAdd(b.pool_ctr, (unsigned long long)take);
                    got = __shfl_sync(FULL, got, 0);
                    if (got + take > b.pool_cells) { status = IT_RETRY_POOL; break; }
                    pool_cur = got; pool_end = got + take;
                }
                i32* st = (i32*)(b.pool + pool_cur);
                const i32 cbl = pbb + (i32)((lane - (u32)pbb) & 31u);
#pragma unroll
                for (int t = 0; t < C; t++) {
                    const i32 j = cbl * C + t;
                    if (j >= beg && j <= end) { st[j - beg] = Hp[t]; st[W + j - beg] = E1p[t]; st[2 * W + j - beg] = E2p[t]; }
                }
                if (lane == 0) {
                    b.nd_beg[nb + s] = beg; b.nd_end[nb + s] = end; b.nd_lpos[nb + s] = lpos; b.nd_rpos[nb + s] = rpos;
                    b.nd_store[nb + s] = (u32)(pool_cur >> 2);
                }
                pool_cur += need;
                __syncwarp();
            }
        }
        // ---- global end: sink predecessors (nodes without successors, ascending) that reach column qlen
        i32 best = INT_MIN;
        u32 bs = 0xffffffffu;
        if (status == VGA_ALN_OK) {
            for (u32 s0 = 0; s0 < S; s0 += 32) {
                const u32 s = s0 + lane;
                i32 h = INT_MIN;
                if (s < S && b.nd_first_succ[nb + s] < 0 && b.nd_end[nb + s] >= qlen) {
                    const i32* st = (const i32*)(b.pool + (u64)b.nd_store[nb + s] * 4);
                    h = st[qlen - b.nd_beg[nb + s]];
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

// ---------------------------------------------------------------------------------------------
// Traceback: one thread per alignment, abPOA's backtrack state machine driven by the flag words.
// Runs and path nodes are written back to front, so they end up in alignment order.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) poa_trace_kernel(TraceBuffers b) {
    const u32 idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= b.n_items) return;
    const u32 w = b.work[idx];
    AlignItem* itp = b.items + w;
    if (itp->status != VGA_ALN_OK) return;
    const AlignItem it = *itp;
    const u32 nb = it.node_base;
    const u32* __restrict__ rs = b.nd_row_start + nb;
    const u32* __restrict__ poff = b.nd_pred_off + nb;
    const u32* __restrict__ pl = b.nd_pred + it.pred_base;
    const u32 cap = it.nrows + it.qlen + 2;
    u32* runs = b.out_runs + it.out_base;
    u32* path = b.out_path + nb;
    u32 wpos = cap, ppos = it.S;
    u32 cur_op = 0xffffffffu, cur_len = 0, n_aligned = 0, path_len = 0;
    u32 s = it.end_node, o = rs[s + 1] - rs[s] - 1;
    i32 j = (i32)it.qlen;
    i32 cur = ALL_OP;
    bool at_source = false, err = false, have_row = false;
    u32 last_node = 0xffffffffu, start_off = 0, end_off = 0;
    auto push = [&](u32 op) {
        if (op == cur_op) { cur_len++; return; }
        if (cur_len) runs[--wpos] = (cur_op << 28) | cur_len;
        cur_op = op; cur_len = 1;
    };
    auto visit_row = [&]() {          // an M/X/D op consumes graph row (s,o)
        path_len++;
        if (s != last_node) { path[--ppos] = s; last_node = s; }
        if (!have_row) { end_off = o; have_row = true; }
        start_off = o;
    };
    while (!at_source && j > 0) {
        const u32 row = it.row_base + rs[s] + o;
        const i32 rb = b.row_beg[row];
        const u32 cell = b.pool[(u64)b.row_toff[row] * 4 + (u32)(j - rb)];
        const i32 jc = j;
        const bool first = o == 0;
        bool hit = false, fresh = false;
        int kind = 0, which = 0;      // kind 1 = M move, 2 = D move
        for (int attempt = 0; attempt < 2 && !hit; attempt++) {
            if ((cur & M_OP) && (cell & TB_M)) {
                push((cell & TB_MATCH) ? 0u : 1u);
                visit_row();
                n_aligned++;
                j--;
                cur = ALL_OP;
                kind = 1; hit = true;
            }
            if (!hit && (cur & (E1_OP | E2_OP))) {
                fresh = (cur & M_OP) != 0;
                which = 0;
                if (cur & E1_OP) { if (fresh ? (cell & TB_E1) : (cell & TB_E1EXT)) which = 1; }
                if (!which && (cur & E2_OP)) { if (fresh ? (cell & TB_E2) : (cell & TB_E2EXT)) which = 2; }
                if (which) {
                    push(2u);
                    visit_row();
                    cur = which == 1 ? E1_OP : E2_OP;     // refined below from the predecessor cell
                    kind = 2; hit = true;
                }
            }
            if (!hit && (cur & (F1_OP | F2_OP)) && j > rb) {
                if (cur & F1_OP) {
                    if (!(cur & M_OP) || (cell & TB_HF1)) { cur = (cell & TB_F1OPEN) ? (M_OP | E1_OP | E2_OP) : F1_OP; hit = true; }
                }
                if (!hit && (cur & F2_OP)) {
                    if (!(cur & M_OP) || (cell & TB_HF2)) { cur = (cell & TB_F2OPEN) ? (M_OP | E1_OP | E2_OP) : F2_OP; hit = true; }
                }
                if (hit) { push(3u); j--; }
            }
            if (!hit) {
                if (cur == ALL_OP) break;
                cur = ALL_OP;         // definition: an exhausted restricted state falls back to ALL once
            }
        }
        if (!hit) { err = true; break; }
        if (kind) {
            const u32 po = poff[s], pn = poff[s + 1] - po;
            if (!first) {
                o--;
            } else if (pn == 0) {
                at_source = true;
            } else {
                const u32 pt = ((const u32*)(b.pool + (u64)b.nd_ptoff[nb + s] * 4))[jc - rb];
                u32 k = kind == 1 ? (pt & 0xff) : fresh ? ((pt >> 8) & 0xff) : which == 1 ? ((pt >> 16) & 0xff) : ((pt >> 24) & 0xff);
                if (k >= pn) { err = true; break; }      // inconsistent trace word: report, never fault
                s = pl[po + k];
                o = rs[s + 1] - rs[s] - 1;
            }
            if (kind == 2 && !at_source) {
                const u32 rowp = it.row_base + rs[s] + o;
                if (jc < b.row_beg[rowp]) { err = true; break; }
                const u32 cellp = b.pool[(u64)b.row_toff[rowp] * 4 + (u32)(jc - b.row_beg[rowp])];
                const bool opened = cellp & (which == 1 ? TB_E1OPEN : TB_E2OPEN);
                cur = opened ? (M_OP | F1_OP | F2_OP) : (which == 1 ? E1_OP : E2_OP);
            }
        }
    }
    if (err) {
        itp->status = VGA_ALN_TRACE_ERR;
        itp->score = 0;
        return;
    }
    if (at_source) for (; j > 0; j--) push(3u);          // leading insertions; no leading deletions
    if (cur_len) runs[--wpos] = (cur_op << 28) | cur_len;
    itp->run_pos = wpos; itp->n_runs = cap - wpos;
    itp->path_pos = ppos; itp->n_path = it.S - ppos;
    itp->path_len = path_len; itp->n_aligned = n_aligned;
    itp->start_off = start_off; itp->end_off = end_off;
}

int launch_poa_fill(const DevIndex& ix, const PoaBuffers& b, int C, int n_sms, cudaStream_t st) {
    if (b.n_items == 0) return 0;
    // persistent warps pulling work from a counter; enough CTAs to fill every SM
    int per_sm = C <= 2 ? 12 : C <= 4 ? 10 : C <= 8 ? 6 : 3;
    long long ctas = (long long)n_sms * per_sm;
    long long need = ((long long)b.n_items + POA_WARPS - 1) / POA_WARPS;
    if (ctas > need) ctas = need;
    if (ctas < 1) ctas = 1;
    dim3 g((unsigned)ctas), t(POA_WARPS * 32);
    switch (C) {
        case 1: poa_fill_kernel<1><<<g, t, 0, st>>>(ix, b); break;
        case 2: poa_fill_kernel<2><<<g, t, 0, st>>>(ix, b); break;
        case 4: poa_fill_kernel<4><<<g, t, 0, st>>>(ix, b); break;
        case 8: poa_fill_kernel<8><<<g, t, 0, st>>>(ix, b); break;
        case 16: poa_fill_kernel<16><<<g, t, 0, st>>>(ix, b); break;
        default: poa_fill_kernel<32><<<g, t, 0, st>>>(ix, b); break;
    }
    return 1;
}

int launch_poa_trace(const TraceBuffers& b, cudaStream_t st) {
    if (b.n_items == 0) return 0;
    poa_trace_kernel<<<(b.n_items + 127) / 128, 128, 0, st>>>(b);
    return 1;
}

// ---------------------------------------------------------------------------------------------
// INT32 ALU peak: independent IADD3 / VIMNMX streams, the instruction mix of the DP inner loop.
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
    // per inner step: 8 adds + 8 max = 16 INT32 ops per thread
    double ops = (double)iters * 16.0 * 16.0 * (double)blocks * threads;
    return ops / (best * 1e-3);
}

}  // namespace vga
