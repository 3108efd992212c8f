// K5b (short reads) — traceback over the PLANE trace format of the 16x2 fill kernel (k_poa_pair.cu): the fill stores,
// per DP row and 8-column block, the three int16 planes it keeps in shared memory anyway (H, E1', E2'; 48 bytes per
// block, AlignItem.fmt = 1) instead of composing a 32-bit word per cell with the E / F deltas — a third fewer
// instructions per cell in the dominant kernel.  What the word format carried and the planes do not, F1 / F2, is
// re-derived here on demand: F_x(i,j) is a left-to-right scan of Hhat(i,.) = max(M, E1, E2) from the row's band begin,
// with Hhat recomputed from the predecessor rows' planes.  It is needed only where the backtrack tests an insertion
// (M and both E tests failed, or the state is restricted to F): a fraction of a percent of the steps of a short read.
// Same value-driven state machine, order of tests and outputs as poa_trace_kernel (k_poa.cu) and the oracle's
// poa_align() (abPOA's backtrack; replaces the result of the call at reference src/align.rs:190-202).
#include <climits>

#include "common.cuh"
#include "internal.h"

namespace vga {

namespace {
enum { M_OP = 1, E1_OP = 2, E2_OP = 4, F1_OP = 8, F2_OP = 16, ALL_OP = 31 };
struct PCell { i32 h, e1, e2; bool ok; };
__device__ __forceinline__ i32 imax(i32 a, i32 b) { return a > b ? a : b; }
__device__ __forceinline__ i32 imin(i32 a, i32 b) { return a < b ? a : b; }
}  // namespace

__global__ void __launch_bounds__(128) poa_trace_planes_kernel(DevIndex ix, TraceBuffers b, u32 spread) {
    const u32 tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid & (spread - 1)) return;
    const u32 idx = tid / spread;
    if (idx >= b.n_items) return;
    const u32 w = b.work[idx];
    AlignItem* itp = b.items + w;
    if (itp->status != VGA_ALN_OK || itp->fmt != 1u) return;
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
    const i32 end0 = imin(qlen, imax(0, qlen - remain_src) + wband);
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
    auto cell = [&](const RowInfo& r, i32 jj) {
        PCell c; c.h = c.e1 = c.e2 = POA_NEG_INF; c.ok = false;
        if (jj >= r.beg && jj <= r.end) {
            const short* p = (const short*)(pool + (u64)r.toff * 4) + planes_cell(r.beg, jj);
            c.h = p[0]; c.e1 = p[8]; c.e2 = p[16];
            c.ok = true;
        }
        return c;
    };
    auto source_cell = [&](i32 jj) {  // virtual source row
        PCell c; c.h = c.e1 = c.e2 = POA_NEG_INF; c.ok = false;
        if (jj >= 0 && jj <= end0) {
            c.ok = true;
            c.h = jj == 0 ? 0 : imax(-(b.o1 + b.e1 * jj), -(b.o2 + b.e2 * jj));
            if (jj == 0) { c.e1 = -oe1; c.e2 = -oe2; }
        }
        return c;
    };
    // current row and, for rows with a single predecessor, that predecessor's record fetched one step ahead (rip) and the one
    // after it in flight (rpp): the loads of a common step — own cell, the predecessor's two cells, the next record — are
    // independent of each other, one memory round trip
    RowInfo ri = rows[rs[it.end_node + 1] - 1];
    if (j < ri.beg || j > ri.end) { itp->status = VGA_ALN_TRACE_ERR; itp->score = 0; return; }
    RowInfo rip = ri;
    bool have_p = false;
    while (!at_source && j > 0) {
        const u32 np = ri.meta & 0xffffffu;
        if (np == 1 && !have_p) { rip = rows[ri.prow]; have_p = true; }
        const PCell me = cell(ri, j);
        const i32 hij = me.h, e1ij = me.e1, e2ij = me.e2;
        const u8 rbase = (u8)(ri.meta >> 24);
        const bool ismatch = q[j - 1] == rbase;
        const i32 sc = ismatch ? b.match : -b.mismatch;
        const i32 jc = j;
        PCell pd = me, pv = me;               // predecessor cells of a single-predecessor row: diagonal and vertical
        RowInfo rpp = rip;
        bool have_pp = false;
        if (np == 1) {
            pd = cell(rip, jc - 1); pv = cell(rip, jc);
            if ((rip.meta & 0xffffffu) == 1u) { rpp = rows[rip.prow]; have_pp = true; }
        }
        const u32 po = np > 1 ? poff[ri.node] : 0;
        RowInfo rsel = rip;               // row moved to by an M / D op
        auto pred_cell = [&](u32 kx, i32 jj, bool diag) {
            if (np == 0) return source_cell(jj);
            if (np == 1) return diag ? pd : pv;
            rsel = rows[rs[pl[po + kx] + 1] - 1];
            return cell(rsel, jj);
        };
        const u32 npl = np == 0 ? 1u : np;
        // F1(i,jc), F2(i,jc): the recurrences of the fill (F(beg) = -inf, F(j) = max(Hhat(j-1) - oe, F(j-1) - e)) from the band begin
        auto f_values = [&](i32& f1, i32& f2) {
            f1 = POA_NEG_INF; f2 = POA_NEG_INF;
            for (i32 jj = ri.beg; jj < jc; jj++) {
                i32 hh = POA_NEG_INF;
                const i32 s2 = (jj >= 1 && q[jj - 1] == rbase) ? b.match : -b.mismatch;
                for (u32 kx = 0; kx < npl; kx++) {
                    PCell d, v;
                    if (np == 0) { d = source_cell(jj - 1); v = source_cell(jj); }
                    else {
                        const RowInfo rr = np == 1 ? rip : rows[rs[pl[po + kx] + 1] - 1];
                        d = cell(rr, jj - 1); v = cell(rr, jj);
                    }
                    if (d.ok && jj >= 1) hh = imax(hh, d.h + s2);
                    if (v.ok) hh = imax(hh, imax(v.e1, v.e2));
                }
                f1 = imax(hh - oe1, f1 - e1);
                f2 = imax(hh - oe2, f2 - e2);
            }
        };
        bool hit = false, moved = false;
        for (int attempt = 0; attempt < 2 && !hit; attempt++) {
            if (cur & M_OP) {
                for (u32 kx = 0; kx < npl; kx++) {
                    const PCell c = pred_cell(kx, jc - 1, true);
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
                    const PCell c = pred_cell(kx, jc, false);
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
                const i32 hl = cell(ri, jc - 1).h;
                i32 f1v, f2v;
                f_values(f1v, f2v);
                if (cur & F1_OP) {
                    if (!(cur & M_OP) || hij == f1v) { cur = (hl - oe1 == f1v) ? (M_OP | E1_OP | E2_OP) : F1_OP; hit = true; }
                }
                if (!hit && (cur & F2_OP)) {
                    if (!(cur & M_OP) || hij == f2v) { cur = (hl - oe2 == f2v) ? (M_OP | E1_OP | E2_OP) : F2_OP; hit = true; }
                }
                if (hit) { push(3u); j--; }
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
            else if (np == 1) { ri = rip; rip = rpp; have_p = have_pp; }
            else { ri = rsel; have_p = false; }
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

int launch_poa_trace_planes(const DevIndex& ix, const TraceBuffers& b, u32 spread, cudaStream_t st) {
    const u64 threads = (u64)b.n_items * spread;
    poa_trace_planes_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(ix, b, spread);
    return 1;
}

}  // namespace vga
