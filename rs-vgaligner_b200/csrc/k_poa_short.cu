// K5 (short reads) — banded sequence-to-graph POA fill for reads shorter than 256 bases, replacing the
// abPOA call at the reference's FFI seam (src/align.rs:190-202).  Same alignment definition, trace
// word format, RowInfo / NodeState records and work-queue protocol as k_poa.cu, so the
// traceback kernel is shared.
//
// Two things make the general kernel slow on short reads: it computes every query column of a row
// although the adaptive band covers a fraction of them, and one warp spends ~120 bookkeeping
// instructions per row (band, trace pool, RowInfo, scans, reductions) on a few dozen cells.  Here
//   * a GROUP of G lanes (G = 8, 16 or 32) owns one (read, chain), so a warp advances 32/G
//     alignments at once and the per-row bookkeeping is SIMD across them.  Every group runs the
//     same flattened state machine (fetch item -> node prologue -> one DP row per iteration), all
//     its state is group-uniform registers, and all exchanges use the group's lane mask;
//   * the work per row is proportional to the band: a row uses the smallest block size CR such
//     that G lanes x CR columns cover [beg,end] from its CR-aligned start (the groups of a warp
//     that are converged agree on the largest CR among them, so they stay converged), and lane l
//     computes columns [(beg/CR + l)*CR, +CR);
//   * since that mapping moves from row to row, the previous row (H, E1', E2') lives in shared
//     memory as int16 (absolute scores fit: same criterion as the trace words), indexed by query
//     column; out-of-band cells are written as -inf plus one guard cell per side, so readers need
//     no band test.  Rows inside a label, and first rows whose only predecessor is the previous
//     node, read it as is; other first rows merge their predecessors' last rows from the trace words;
//   * substitution scores come from a per-alignment query profile (int8 per column for A,C,G,T; any
//     other graph byte gets a profile row built on demand — raw byte equality, like the oracle).
// F1/F2 are prefix-max scans: serial inside the lane block, Kogge-Stone across the group.
// Roofline: INT32 ALU issue.  Algorithmic work = band cells.
#include <climits>
#include <cstdlib>

#include "common.cuh"
#include "internal.h"

namespace vga {

static const int PG_WARPS = 4;           // warps per CTA
#ifndef PG_MIN_CTAS
#define PG_MIN_CTAS 5
#endif
static const u64 GROUP_POOL_CHUNK = 4096;   // trace words a group takes from the pool per refill
static const int PAD = 4;                // smem index of query column 0 (column -1 is the DEEP guard at PAD-1)
static const int CR_MAX = 12;
static const int DUMMY = 272;            // lane blocks entirely right of the band are parked on columns DUMMY..DUMMY+CR_MAX-1
static const int SW = PAD + DUMMY + CR_MAX;
static const i32 S16 = -32000;           // -inf of the int16 planes; reachable scores are >= -30000 in this class
static const i32 SHORT_SCORE_LIMIT = 30000;

__device__ __forceinline__ i32 imax(i32 a, i32 b) { return a > b ? a : b; }
__device__ __forceinline__ i32 imin(i32 a, i32 b) { return a < b ? a : b; }
__device__ __forceinline__ i32 sext16(u32 w) { return (i32)(short)(w & 0xffffu); }

struct GroupSmem {
    short h[SW], e1[SW], e2[SW];          // previous row; updated in place (loads precede the group's scan, stores follow it)
    signed char prof[5][SW];              // substitution score of column j against A, C, T, G (2-bit code order) and the custom base
};

struct RowConst { i32 e1, e2, oe1, oe2, o1, o2; };
struct RowOut { i32 mx, lpos, rpos; };

template <int G> __device__ __forceinline__ u32 group_mask(u32 lane) {
    return G == 32 ? 0xffffffffu : (((1u << (G & 31)) - 1u) << (lane & ~(u32)(G - 1)));
}
template <int G> __device__ __forceinline__ i32 group_max(u32 gm, i32 v) {
    if constexpr (G == 32) return __reduce_max_sync(0xffffffffu, v);
    else {
#pragma unroll
    for (int d = G / 2; d > 0; d >>= 1) v = imax(v, __shfl_xor_sync(gm, v, d, G));
    return v;
    }
}
template <int G> __device__ __forceinline__ i32 group_min(u32 gm, i32 v) {
    if constexpr (G == 32) return __reduce_min_sync(0xffffffffu, v);
    else {
#pragma unroll
    for (int d = G / 2; d > 0; d >>= 1) v = imin(v, __shfl_xor_sync(gm, v, d, G));
    return v;
    }
}

// One DP row of one group.  `tr` is the row's trace words (word t = column beg + t).  Lane l computes the CR
// columns from beg + l*CR.  Only in-band cells are stored (previous row in shared memory, trace words); the
// generator of F is masked outside the band, so an out-of-band H is either -inf (left of beg) or an F that decays
// from the last band cell (right of end) and never reaches the row maximum (e1, e2 >= 1).
template <int G, int CR>
__device__ __forceinline__ RowOut row_body(const RowConst& K, u32 gm, u32 gl, i32 beg, i32 end, GroupSmem& sm,
                                           const signed char* __restrict__ prof, u32* __restrict__ tr) {
    const i32 j0 = beg + (i32)gl * CR;
    const i32 idx = imin(j0, DUMMY) + PAD;                 // lanes right of the band read the dummy block (all their cells are masked)
    const i32 thi = end - j0;                              // column t of the lane is in the band iff t <= thi
    i32 x1[CR], x2[CR], hh[CR], f1l[CR], f2l[CR];
    // ---- sweep 1: Hhat = max(M, E1, E2) and the lane-local part of the F scans
    {
        i32 hd = sm.h[idx - 1];
        i32 run1 = POA_DEEP, run2 = POA_DEEP;
#pragma unroll
        for (int t = 0; t < CR; t++) {
            x1[t] = sm.e1[idx + t]; x2[t] = sm.e2[idx + t];
            const i32 m = hd + (i32)prof[idx + t];
            hd = sm.h[idx + t];
            hh[t] = t <= thi ? __vimax3_s32(m, x1[t], x2[t]) : POA_DEEP;
            f1l[t] = run1; f2l[t] = run2;
            run1 = __viaddmax_s32(run1, -K.e1, hh[t] - K.oe1);
            run2 = __viaddmax_s32(run2, -K.e2, hh[t] - K.oe2);
        }
        // carry of F1/F2 across the group (Kogge-Stone; shfl_up leaves lanes < d unchanged, and max(v, v - c) = v)
#pragma unroll
        for (int d = 1; d < G; d <<= 1) {
            const i32 t1 = __shfl_up_sync(gm, run1, d, G), t2 = __shfl_up_sync(gm, run2, d, G);
            run1 = __viaddmax_s32(t1, -K.e1 * CR * d, run1);
            run2 = __viaddmax_s32(t2, -K.e2 * CR * d, run2);
        }
        run1 = __shfl_up_sync(gm, run1, 1, G); run2 = __shfl_up_sync(gm, run2, 1, G);
        if (gl == 0) { run1 = POA_DEEP; run2 = POA_DEEP; }
        // F arriving at the lane's first column; F(i,beg) = -inf falls out of the masked generators
#pragma unroll
        for (int t = 0; t < CR; t++) {
            f1l[t] = __viaddmax_s32(run1, -K.e1 * t, f1l[t]);
            f2l[t] = __viaddmax_s32(run2, -K.e2 * t, f2l[t]);
        }
    }
    // ---- sweep 2: H, E', trace words, row maximum keyed by column
    i32 h[CR];
    u32 wd[CR];
    i32 kl = INT_MIN, kr = INT_MIN;          // max of (H, -column) and (H, column): row maximum with its leftmost / rightmost column
    short* const ph = sm.h + idx; short* const pe1 = sm.e1 + idx; short* const pe2 = sm.e2 + idx;
#pragma unroll
    for (int t = 0; t < CR; t++) {
        h[t] = imax(__vimax3_s32(hh[t], f1l[t], f2l[t]), S16);
        const i32 d1 = __viaddmax_s32(x1[t], K.o1 - h[t], 0);          // E1' - (H - oe1) = max(E1 - e1 - H + oe1, 0)
        const i32 d2 = __viaddmax_s32(x2[t], K.o2 - h[t], 0);
        u32 hi = (u32)d1 + ((u32)d2 << 3);
        if (t > 0) hi += ((u32)__viaddmax_s32(f1l[t], 1 + K.oe1 - h[t - 1], 0) << 8) + ((u32)__viaddmax_s32(f2l[t], 1 + K.oe2 - h[t - 1], 0) << 11);
        wd[t] = __byte_perm((u32)h[t], hi, 0x5410);
        if (t <= thi) {
            ph[t] = (short)h[t];
            pe1[t] = (short)(d1 + h[t] - K.oe1);
            pe2[t] = (short)(d2 + h[t] - K.oe2);
        }
        kl = imax(kl, h[t] * 512 + (511 - t));
        kr = imax(kr, h[t] * 512 + t);
    }
    {   // F fields of the lane's first column: relative to the left lane's last H
        const i32 hl = __shfl_up_sync(gm, h[CR - 1], 1, G);
        if (gl > 0) wd[0] += ((u32)imax(f1l[0] + 1 + K.oe1 - hl, 0) << 24) + ((u32)imax(f2l[0] + 1 + K.oe2 - hl, 0) << 27);
    }
    if (gl == 0 && beg > 0) sm.h[beg - 1 + PAD] = (short)S16;                              // left guard (diagonal of column beg)
    if (gl == (1 % G)) { sm.e1[end + 1 + PAD] = (short)S16; sm.e2[end + 1 + PAD] = (short)S16; }   // right guard (vertical of column end+1)
    u32* const trl = tr + gl * CR;
#pragma unroll
    for (int t = 0; t < CR; t++)
        if (t <= thi) trl[t] = wd[t];
    // ---- row maximum and its leftmost / rightmost column (abPOA simd_abpoa_ada_max_i); keys carry the column
    kl = group_max<G>(gm, kl - j0);          // (H, -(j0 + t)) as H*512 + 511 - t - j0
    kr = group_max<G>(gm, kr + j0);
    RowOut o;
    o.mx = kl >> 9;
    o.lpos = 511 - (kl & 511);
    o.rpos = kr & 511;
    return o;
}

// CR choice: the smallest block size of the instance's set with G * CR >= W; 0 = the band does not fit this instance.
template <int G> __device__ __forceinline__ int pick_cr(i32 W) {
    if (G == 8) return W <= 16 ? 2 : W <= 32 ? 4 : W <= 64 ? 8 : W <= 96 ? 12 : 0;
    if (G == 16) return W <= 32 ? 2 : W <= 64 ? 4 : W <= 96 ? 6 : W <= 160 ? 10 : 0;
    return W <= 32 ? 1 : W <= 64 ? 2 : W <= 96 ? 3 : W <= 160 ? 5 : 8;      // 32 lanes x 8 covers every band of a read < 256
}

template <int G>
__global__ void __launch_bounds__(PG_WARPS * 32, PG_MIN_CTAS) poa_fill_group_kernel(DevIndex ix, PoaBuffers b) {
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    GroupSmem& sm = ((GroupSmem*)dyn_smem)[threadIdx.x / G];
    const u32 lane = threadIdx.x & 31;
    const u32 gl = lane & (G - 1);
    const u32 gm = group_mask<G>(lane);
    RowConst K;
    K.e1 = b.e1; K.e2 = b.e2; K.oe1 = b.o1 + b.e1; K.oe2 = b.o2 + b.e2; K.o1 = b.o1; K.o2 = b.o2;
    const i32 oe1 = K.oe1, oe2 = K.oe2;
    const i32 smatch = b.match, smis = -b.mismatch;
    if (gl == 0) sm.h[PAD - 1] = (short)S16;        // column -1: M(i,0) does not exist
    // ---- group-uniform state of the alignment in flight
    bool active = false;
    u32 w = 0, nb = 0, S = 0, pred_base = 0, row_base = 0, nrows = 0;
    i32 qlen = 0, wband = 0, end0 = 0;
    const u8* __restrict__ q = nullptr;
    u32 row = 0, s = 0, o = 0, metas = 0;
    i32 beg = 0, end = 0, ppb = 0, ppe = -1, lpos = 0, rpos = 0, mxrow = 0;
    u64 cells = 0;
    u32 status = VGA_ALN_OK, custom_base = 0xffffffffu;
    u64 pool_cur = 0, pool_end = 0;                 // this group's current chunk of the trace pool (words)

    // result record of the item in flight; the group goes back to fetching
    auto finish = [&]() {
        i32 best = INT_MIN;
        u32 bs = 0xffffffffu;
        if (status == VGA_ALN_OK) {
            // global end: sink predecessors (nodes without successors, ascending) that reach column qlen
            const u32* __restrict__ rs = b.nd_row_start + nb;
            const RowInfo* __restrict__ rows = b.rows + row_base;
            __syncwarp(gm);
            for (u32 s0 = 0; s0 < S; s0 += G) {
                const u32 sx = s0 + gl;
                i32 h = INT_MIN;
                if (sx < S && b.nd_first_succ[nb + sx] < 0) {
                    const RowInfo ri = rows[rs[sx + 1] - 1];
                    if (ri.end >= qlen) h = ri.ref + sext16(b.pool[(u64)ri.toff * 4 + (u32)(qlen - trace_col0(ri.beg))]);
                }
                const i32 mx = group_max<G>(gm, h);
                if (mx > best) {                             // strict: the first node wins ties
                    best = mx;
                    bs = (u32)group_min<G>(gm, h == mx ? (i32)sx : INT_MAX);
                }
            }
            if (bs == 0xffffffffu) status = VGA_ALN_EMPTY;
        }
        if (gl == 0) {
            AlignItem* out = b.items + w;
            out->status = status;
            out->score = status == VGA_ALN_OK ? best : 0;
            out->end_node = bs;
            out->cells = cells;
            out->n_runs = 0; out->n_path = 0; out->path_len = 0; out->start_off = 0; out->end_off = 0; out->n_aligned = 0;
        }
        active = false;
    };

    for (;;) {
        if (!active) {
            // ---- next (read, chain) of the work list
            u32 wi = 0;
            if (gl == 0) wi = atomicAdd(b.work_ctr, 1u);
            wi = __shfl_sync(gm, wi, 0, G);
            if (wi >= b.n_items) break;
            w = b.work[wi];
            const AlignItem it = b.items[w];
            qlen = (i32)it.qlen;
            q = (const u8*)b.seqs + it.q_off;
            wband = b.wb + (i32)(b.wf * (float)qlen);
            nb = it.node_base; S = it.S; row_base = it.row_base; pred_base = it.pred_base; nrows = it.nrows;
            status = VGA_ALN_OK;
            cells = 0; row = 0; s = 0; o = 0; ppb = 0; ppe = -1; lpos = 0; rpos = 0; mxrow = 0;
            custom_base = 0xffffffffu;
            active = true;
            if (S == 0 || nrows == 0) { status = it.status == VGA_ALN_UNSUPPORTED ? VGA_ALN_UNSUPPORTED : VGA_ALN_EMPTY; finish(); continue; }
            // absolute H is kept as int16 (trace words and shared memory): every DP step loses at most
            // max(mismatch, o1+e1), so reachable scores are >= -that*(rows+qlen); other alignments go to the general kernel
            if (qlen >= 256 || (i64)imax(b.mismatch, oe1) * ((i64)nrows + qlen) > SHORT_SCORE_LIMIT) {
                status = IT_RETRY_WIDER;
                finish();
                continue;
            }
            {   // band end of the virtual source row: max_remain of the first row + 1
                const i32 remain_src = (i32)(b.row_meta[row_base] >> 11) + 1;
                end0 = band_round_end0(imin(qlen, imax(0, qlen - remain_src) + wband), b.bgran_mask, qlen);
            }
            // query profile: score of column j (query base j-1) against A, C, T, G (order of the 2-bit code (c >> 1) & 3)
            __syncwarp(gm);
            for (i32 j = (i32)gl; j < SW - PAD; j += G) {
                const u32 c = (j >= 1 && j <= qlen) ? q[j - 1] : 0u;
                sm.prof[0][j + PAD] = (signed char)(c == 'A' ? smatch : smis);
                sm.prof[1][j + PAD] = (signed char)(c == 'C' ? smatch : smis);
                sm.prof[2][j + PAD] = (signed char)(c == 'T' ? smatch : smis);
                sm.prof[3][j + PAD] = (signed char)(c == 'G' ? smatch : smis);
            }
            __syncwarp(gm);
        }
        // ---- one DP row; its record (sub_rowmeta_kernel) is fetched G rows at a time
        if ((row & (G - 1)) == 0) metas = (row + gl < nrows) ? b.row_meta[row_base + row + gl] : 0u;
        const u32 meta = __shfl_sync(gm, metas, row & (G - 1), G);
        const u32 base = meta & 0xffu, kind = (meta >> 8) & 3u;
        // predecessor row: the previous row unless this is a first row with other predecessors
        i32 mpl = lpos + 1, mpr = rpos + 1, min_pb = ppb, max_pe = ppe;
        u32 prow = row - 1, np = 1, po = 0;
        if (kind == 1) { mpl = 1; mpr = 1; min_pb = 0; max_pe = end0; np = 0; prow = 0; }      // virtual source row
        else if (kind == 2) {
            __syncwarp(gm);       // NodeState / RowInfo / trace words of earlier rows were written by other lanes of the group
            po = b.nd_pred_off[nb + s];
            np = b.nd_pred_off[nb + s + 1] - po;
            const u32* __restrict__ pl = b.nd_pred + pred_base + po;
            mpl = INT_MAX; mpr = 0; min_pb = INT_MAX; max_pe = -1;
            for (u32 x = 0; x < np; x++) {
                const NodeState ns = b.nd_state[nb + pl[x]];
                mpl = imin(mpl, ns.lpos + 1); mpr = imax(mpr, ns.rpos + 1);
                min_pb = imin(min_pb, ns.beg); max_pe = imax(max_pe, ns.end);
            }
            prow = b.nd_row_start[nb + pl[0] + 1] - 1;
        }
        // adaptive band (abPOA GET_AD_DP_BEGIN/END + clipping to the predecessors' bands)
        const i32 r = qlen - (i32)(meta >> 11);
        beg = imax(imax(0, imin(mpl, r) - wband), min_pb);
        end = imin(imin(qlen, imax(mpr, r) + wband), max_pe + 1);
        band_round(beg, end, b.bgran_mask, qlen, max_pe);
        const i32 W = end - beg + 1;
        int cr = pick_cr<G>(W);
        if (beg > end) status = VGA_ALN_BAND_ERR;
        else if (cr == 0) status = IT_RETRY_WIDER;
        // pool space: one trace word per band cell, rows start on 16-byte boundaries
        const u64 need = (u64)((W + (beg & 7) + 3) & ~3);          // word 0 of a row is column beg & ~7 (trace_col0)
        if (status == VGA_ALN_OK && pool_cur + need > pool_end) {
            unsigned long long got = 0;
            const u64 take = need > GROUP_POOL_CHUNK ? need : GROUP_POOL_CHUNK;
            if (gl == 0) got = atomicAdd(b.pool_ctr, (unsigned long long)take);
            got = __shfl_sync(gm, got, 0, G);
            if (got + take > b.pool_words) status = IT_RETRY_POOL;
            else { pool_cur = got; pool_end = got + take; }
        }
        if (status != VGA_ALN_OK) { finish(); continue; }
        u32* __restrict__ tr = b.pool + pool_cur + (beg & 7);
        if (gl == 0) {
            uint4* dst = (uint4*)(b.rows + row_base + row);
            dst[0] = make_uint4((u32)beg, (u32)end, (u32)(pool_cur >> 2), 0u);
            dst[1] = make_uint4(prow, (np < 0xffffffu ? np : 0xffffffu) | (base << 24), s, o);
        }
        pool_cur += need;
        cells += (u64)W;
        if (kind != 0) {
            // the predecessor row is not the one in shared memory: build it there
            const i32 jlo = imax(beg - 1, 0);
            if (kind == 1) {          // virtual source row
                __syncwarp(gm);
                for (i32 j = jlo + (i32)gl; j <= end + 1; j += G) {
                    const bool in0 = j <= end0;
                    sm.h[j + PAD] = (short)(!in0 ? S16 : j == 0 ? 0 : imax(-(b.o1 + b.e1 * j), -(b.o2 + b.e2 * j)));
                    sm.e1[j + PAD] = (short)(j == 0 ? -oe1 : S16);
                    sm.e2[j + PAD] = (short)(j == 0 ? -oe2 : S16);
                }
            } else {                  // merge of the predecessors' last rows, decoded from their trace words
                const u32* __restrict__ pl = b.nd_pred + pred_base + po;
                for (u32 x = 0; x < np; x++) {
                    // one predecessor at a time: its band and trace row are read once, its words stream through the lanes
                    const u32 p = pl[x];
                    const int2 pbe = *(const int2*)(b.nd_state + nb + p);
                    const u32* __restrict__ tw = b.pool + (u64)b.rows[row_base + b.nd_row_start[nb + p + 1] - 1].toff * 4;
#pragma unroll 2
                    for (i32 j = jlo + (i32)gl; j <= end + 1; j += G) {
                        i32 hv = S16, e1v = S16, e2v = S16;
                        if (j >= pbe.x && j <= pbe.y) {
                            const u32 v = tw[j - trace_col0(pbe.x)];
                            hv = sext16(v);
                            e1v = hv - oe1 + (i32)((v >> 16) & 7u);
                            e2v = hv - oe2 + (i32)((v >> 19) & 31u);
                        }
                        if (x > 0) {
                            hv = imax(hv, (i32)sm.h[j + PAD]); e1v = imax(e1v, (i32)sm.e1[j + PAD]); e2v = imax(e2v, (i32)sm.e2[j + PAD]);
                        }
                        sm.h[j + PAD] = (short)hv; sm.e1[j + PAD] = (short)e1v; sm.e2[j + PAD] = (short)e2v;
                    }
                }
            }
        }
        // substitution profile row of this graph base
        u32 code = (base >> 1) & 3u;
        if (base != ((0x47544341u >> (8 * code)) & 0xffu)) {          // not A, C, T, G: custom row
            code = 4;
            if (base != custom_base) {
                __syncwarp(gm);
                for (i32 j = (i32)gl; j < SW - PAD; j += G)
                    sm.prof[4][j + PAD] = (signed char)((j >= 1 && j <= qlen && q[j - 1] == base) ? smatch : smis);
                custom_base = base;
            }
        }
        __syncwarp(gm);
        const signed char* pr = sm.prof[code];
        // the groups of the warp that are here together agree on the largest block size among them (stay converged)
        cr = __reduce_max_sync(__activemask(), cr);
        RowOut ro;
        if (G == 8) {
            switch (cr) {
                case 2: ro = row_body<G, 2>(K, gm, gl, beg, end, sm, pr, tr); break;
                case 4: ro = row_body<G, 4>(K, gm, gl, beg, end, sm, pr, tr); break;
                case 8: ro = row_body<G, 8>(K, gm, gl, beg, end, sm, pr, tr); break;
                default: ro = row_body<G, 12>(K, gm, gl, beg, end, sm, pr, tr); break;
            }
        } else if (G == 16) {
            switch (cr) {
                case 2: ro = row_body<G, 2>(K, gm, gl, beg, end, sm, pr, tr); break;
                case 4: ro = row_body<G, 4>(K, gm, gl, beg, end, sm, pr, tr); break;
                case 6: ro = row_body<G, 6>(K, gm, gl, beg, end, sm, pr, tr); break;
                default: ro = row_body<G, 10>(K, gm, gl, beg, end, sm, pr, tr); break;
            }
        } else {
            switch (cr) {
                case 1: ro = row_body<G, 1>(K, gm, gl, beg, end, sm, pr, tr); break;
                case 2: ro = row_body<G, 2>(K, gm, gl, beg, end, sm, pr, tr); break;
                case 3: ro = row_body<G, 3>(K, gm, gl, beg, end, sm, pr, tr); break;
                case 5: ro = row_body<G, 5>(K, gm, gl, beg, end, sm, pr, tr); break;
                default: ro = row_body<G, 8>(K, gm, gl, beg, end, sm, pr, tr); break;
            }
        }
        lpos = ro.lpos; rpos = ro.rpos; mxrow = ro.mx;
        ppb = beg; ppe = end;
        if (meta & 0x400u) {
            // last row of its node: it may be a predecessor of later nodes' first rows
            if (gl == 0) {
                int4* dst = (int4*)(b.nd_state + nb + s);
                dst[0] = make_int4(beg, end, lpos, rpos);
                dst[1] = make_int4(mxrow, 0, 0, 0);
            }
            s++; o = 0;
        } else o++;
        row++;
        if (row == nrows) finish();
    }
}

template <int G>
static void launch_group(const DevIndex& ix, const PoaBuffers& b, int n_sms, cudaStream_t st) {
    int per_sm = 0;
    const int groups_per_cta = PG_WARPS * 32 / G;
    const size_t smem_bytes = (size_t)groups_per_cta * sizeof(GroupSmem);
    static bool attr_set = false;
    if (!attr_set) {
        VGA_CUDA(cudaFuncSetAttribute(poa_fill_group_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
        attr_set = true;
    }
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, poa_fill_group_kernel<G>, PG_WARPS * 32, smem_bytes);
    if (per_sm < 1) per_sm = 1;
    long long ctas = (long long)n_sms * per_sm;
    const long long need = ((long long)b.n_items + groups_per_cta - 1) / groups_per_cta;
    if (ctas > need) ctas = need;
    if (ctas < 1) ctas = 1;
    poa_fill_group_kernel<G><<<(unsigned)ctas, PG_WARPS * 32, smem_bytes, st>>>(ix, b);
}

// short-read classes: 0 = 8 lanes per alignment (bands up to 96 columns), 1 = 16 lanes (up to 160), 2 = 32 lanes (any read < 256)
int launch_poa_fill_short(const DevIndex& ix, const PoaBuffers& b, int cls, int n_sms, cudaStream_t st) {
    if (b.n_items == 0) return 0;
    if (cls == 0) launch_group<8>(ix, b, n_sms, st);
    else if (cls == 1) launch_group<16>(ix, b, n_sms, st);
    else launch_group<32>(ix, b, n_sms, st);
    return 1;
}

bool poa_short_ok(i64 qlen, i64 nrows, const VgaParams& P) {
    const i64 step = P.mismatch > P.o1 + P.e1 ? P.mismatch : P.o1 + P.e1;
    return qlen < 256 && step * (nrows + qlen) <= SHORT_SCORE_LIMIT;
}

}  // namespace vga
