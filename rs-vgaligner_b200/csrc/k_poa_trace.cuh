// Traceback of one alignment (K5b), shared by the traceback kernel (k_poa.cu) and by the fused long-read kernel (k_poa_long.cu),
// which traces an alignment right after filling it.
#pragma once
#include "common.cuh"
#include "internal.h"

namespace vga {

enum { M_OP = 1, E1_OP = 2, E2_OP = 4, F1_OP = 8, F2_OP = 16, ALL_OP = 31 };
namespace trace_detail {
__device__ __forceinline__ i32 imax(i32 a, i32 b) { return a > b ? a : b; }
__device__ __forceinline__ i32 imin(i32 a, i32 b) { return a < b ? a : b; }
__device__ __forceinline__ i32 sext16(u32 w) { return (i32)(short)(w & 0xffffu); }
}  // namespace trace_detail

// ---------------------------------------------------------------------------------------------
// Traceback: one thread per alignment, abPOA's value-driven backtrack on the values re-derived
// from the trace words (same order of tests as oracle/oracle.cpp poa_align()).
// Runs and path nodes are written back to front, so they end up in alignment order.
// ---------------------------------------------------------------------------------------------
struct TCell { i32 h, e1, e2; bool ok; };

// One traceback, by the calling thread: item w of b.items (its fill must be complete and visible).  FAST: see below.
template <bool FAST>
__device__ __forceinline__ void poa_trace_one(const TraceBuffers& b, u32 w) {
    using namespace trace_detail;
    AlignItem* itp = b.items + w;
    if (itp->status != VGA_ALN_OK) return;
    const AlignItem it = *itp;
    const u32 nb = it.node_base;
    const u32* __restrict__ rs = b.nd_row_start + nb;
    const u32* __restrict__ poff = b.nd_pred_off + nb;
    const u32* __restrict__ pl = b.nd_pred + it.pred_base;
    const RowInfo* __restrict__ rows = b.rows + it.row_base;
    const u32* __restrict__ pool = b.pool;
    const u8* __restrict__ q = (const u8*)b.seqs + it.q_off;
    const i32 e1 = b.e1, e2 = b.e2, oe1 = b.o1 + b.e1, oe2 = b.o2 + b.e2;
    const i32 qlen = (i32)it.qlen;
    const i32 wband = b.wb + (i32)(b.wf * (float)qlen);
    const i32 remain_src = b.nd_remain_last[nb] + (i32)(rs[1] - rs[0]) - 1 + 1;
    const i32 end0 = band_round_end0(imin(qlen, imax(0, qlen - remain_src) + wband), b.bgran_mask, qlen);
    const u32 cap = it.nrows + it.qlen + 2;
    u32* runs = b.out_runs + it.out_base;
    u32* path = b.out_path + nb;
    u32 wpos = cap, ppos = it.S;
    u32 cur_op = 0xffffffffu, cur_len = 0, n_aligned = 0, path_len = 0;
    i32 j = qlen;
    i32 cur = ALL_OP;
    bool at_source = false, err = false, have_row = false;
    u32 last_node = 0xffffffffu, start_off = 0, end_off = 0;
    auto push = [&](u32 op) {
        if (op == cur_op) { cur_len++; return; }
        if (cur_len) runs[--wpos] = (cur_op << 28) | cur_len;
        cur_op = op; cur_len = 1;
    };
    auto word_at = [&](const RowInfo& r, i32 jj) -> u32 { return pool[(u64)r.toff * 4 + (u32)(jj - trace_col0(r.beg))]; };
    auto decode = [&](const RowInfo& r, i32 jj, u32 v) {
        TCell c; c.h = c.e1 = c.e2 = POA_NEG_INF; c.ok = false;
        if (jj >= r.beg && jj <= r.end) {
            c.h = r.ref + sext16(v); c.e1 = c.h - oe1 + (i32)((v >> 16) & 7u); c.e2 = c.h - oe2 + (i32)((v >> 19) & 31u);
            c.ok = true;
        }
        return c;
    };
    auto source_cell = [&](i32 jj) {  // virtual source row
        TCell c; c.h = c.e1 = c.e2 = POA_NEG_INF; c.ok = false;
        if (jj >= 0 && jj <= end0) {
            c.ok = true;
            c.h = jj == 0 ? 0 : imax(-(b.o1 + b.e1 * jj), -(b.o2 + b.e2 * jj));
            if (jj == 0) { c.e1 = -oe1; c.e2 = -oe2; }
        }
        return c;
    };
    // current row, its word at column j, and (single-predecessor rows) the predecessor's RowInfo, which
    // is always fetched one step ahead so that a step costs one memory round trip
    RowInfo ri = rows[rs[it.end_node + 1] - 1];
    if (j < ri.beg || j > ri.end) { itp->status = VGA_ALN_TRACE_ERR; itp->score = 0; return; }
    u32 vij = word_at(ri, j);
    RowInfo rip = ri;
    bool have_p = false;
    while (!at_source && j > 0) {
        const u32 np = ri.meta & 0xffffffu;
        const i32 hij = ri.ref + sext16(vij);
        const i32 e1ij = hij - oe1 + (i32)((vij >> 16) & 7u), e2ij = hij - oe2 + (i32)((vij >> 19) & 31u);
        const bool ismatch = q[j - 1] == (u8)(ri.meta >> 24);
        const i32 sc = ismatch ? b.match : -b.mismatch;
        const i32 jc = j;
        // ---- the common step, on its own short path: a row with one predecessor, unrestricted state, and the diagonal cell
        // explains H (an M op — the first test of the general path below, which stays the definition).  Its two loads, the
        // diagonal word and the RowInfo one row further up, depend only on what the previous step left in registers, so
        // the step costs ONE memory round trip; lanes that miss fall through to the general path.  Only when a warp holds few
        // tracebacks (FAST, spread >= 4): with 16 or 32 per warp some lane misses on almost every step, and the warp then
        // pays for both paths (measured: C5 62 -> 50 ms with it, C2 3.4 -> 4.2 ms when it was unconditional).
        if (FAST && np == 1 && cur == ALL_OP && have_p && jc - 1 >= rip.beg && jc - 1 <= rip.end) {
            const u32 wdf = word_at(rip, jc - 1);
            const bool p1 = (rip.meta & 0xffffffu) == 1u;
            RowInfo rnx = rip;
            if (p1) rnx = rows[rip.prow];
            if (rip.ref + sext16(wdf) + sc == hij) {
                push(ismatch ? 0u : 1u);
                n_aligned++; path_len++; j--;
                if (ri.node != last_node) { path[--ppos] = ri.node; last_node = ri.node; }
                if (!have_row) { end_off = ri.off; have_row = true; }
                start_off = ri.off;
                ri = rip; rip = rnx; have_p = p1; vij = wdf;
                continue;
            }
        }
        // ---- issue every load of this step together
        if (np == 1 && !have_p) { rip = rows[ri.prow]; have_p = true; }
        u32 wl = 0, wd = 0, wv = 0;
        if (jc > ri.beg) wl = word_at(ri, jc - 1);
        RowInfo rpp = rip;
        bool have_pp = false;
        if (np == 1) {
            if (jc - 1 >= rip.beg && jc - 1 <= rip.end) wd = word_at(rip, jc - 1);
            if (jc >= rip.beg && jc <= rip.end) wv = word_at(rip, jc);
            if ((rip.meta & 0xffffffu) == 1u) { rpp = rows[rip.prow]; have_pp = true; }
        }
        const u32 po = np > 1 ? poff[ri.node] : 0;
        RowInfo rsel = rip;               // row moved to by an M / D op
        u32 wsel = 0;
        auto pred_cell = [&](u32 kx, i32 jj, bool diag) {
            if (np == 0) return source_cell(jj);
            if (np == 1) { wsel = diag ? wd : wv; return decode(rip, jj, wsel); }
            const u32 p = pl[po + kx];
            rsel = rows[rs[p + 1] - 1];
            wsel = (jj >= rsel.beg && jj <= rsel.end) ? word_at(rsel, jj) : 0u;
            return decode(rsel, jj, wsel);
        };
        const u32 npl = np == 0 ? 1u : np;
        bool hit = false, moved = false;
        for (int attempt = 0; attempt < 2 && !hit; attempt++) {
            if (cur & M_OP) {
                for (u32 kx = 0; kx < npl; kx++) {
                    const TCell c = pred_cell(kx, jc - 1, true);
                    if (!c.ok) continue;
                    if (c.h + sc == hij) {
                        push(ismatch ? 0u : 1u);
                        n_aligned++;
                        moved = true; j--; cur = ALL_OP; hit = true;
                        break;
                    }
                }
            }
            if (!hit && (cur & (E1_OP | E2_OP))) {
                for (u32 kx = 0; kx < npl && !hit; kx++) {
                    const TCell c = pred_cell(kx, jc, false);
                    if (!c.ok) continue;
                    if (cur & E1_OP) {
                        const bool t = (cur & M_OP) ? (hij == c.e1) : (e1ij == c.e1 - e1);
                        if (t) { cur = (c.h - oe1 == c.e1) ? (M_OP | F1_OP | F2_OP) : E1_OP; hit = true; }
                    }
                    if (!hit && (cur & E2_OP)) {
                        const bool t = (cur & M_OP) ? (hij == c.e2) : (e2ij == c.e2 - e2);
                        if (t) { cur = (c.h - oe2 == c.e2) ? (M_OP | F1_OP | F2_OP) : E2_OP; hit = true; }
                    }
                    if (hit) { push(2u); moved = true; }
                }
            }
            if (!hit && (cur & (F1_OP | F2_OP)) && jc > ri.beg) {
                const i32 hl = ri.ref + sext16(wl);
                if (cur & F1_OP) {
                    const i32 f = hl - oe1 + (i32)((vij >> 24) & 7u) - 1;
                    if (!(cur & M_OP) || hij == f) { cur = (hl - oe1 == f) ? (M_OP | E1_OP | E2_OP) : F1_OP; hit = true; }
                }
                if (!hit && (cur & F2_OP)) {
                    const i32 f = hl - oe2 + (i32)((vij >> 27) & 31u) - 1;
                    if (!(cur & M_OP) || hij == f) { cur = (hl - oe2 == f) ? (M_OP | E1_OP | E2_OP) : F2_OP; hit = true; }
                }
                if (hit) { push(3u); j--; vij = wl; }
            }
            if (!hit) {
                if (cur == ALL_OP) break;
                cur = ALL_OP;         // definition: an exhausted restricted state falls back to ALL once
            }
        }
        if (!hit) { err = true; break; }
        if (moved) {                  // an M/X/D op consumed graph row ri and moves to a predecessor row
            path_len++;
            if (ri.node != last_node) { path[--ppos] = ri.node; last_node = ri.node; }
            if (!have_row) { end_off = ri.off; have_row = true; }
            start_off = ri.off;
            if (np == 0) at_source = true;
            else if (np == 1) { ri = rip; rip = rpp; have_p = have_pp; vij = wsel; }
            else { ri = rsel; have_p = false; vij = wsel; }
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


}  // namespace vga
