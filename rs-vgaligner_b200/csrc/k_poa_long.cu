// K5 (long reads) — banded sequence-to-graph POA fill for reads of any length whose band stays within 1024 columns,
// replacing the abPOA call at the reference's FFI seam (src/align.rs:190-202).  Same alignment definition, trace word
// format (scores relative to ref(row) = maximum of the previous row), RowInfo / NodeState records and
// work-queue protocol as k_poa.cu, so the traceback kernel is shared.
//
// The SLIDE instances of k_poa.cu keep the previous row in registers, C columns per lane, C fixed per kernel: every
// row costs 32*C columns although the adaptive band of a 10 kb read covers a third of that on average, and an
// alignment has to run on the instance its WIDEST row needs.  Here (the design of k_poa_short.cu, 32-bit):
//   * one warp per (read, chain); the work of a row is proportional to ITS band: the row uses the smallest block size
//     CR of {5,9,13,17,25,33} with 32 lanes x CR columns >= band, lane l computes columns [beg + l*CR, +CR) — CR is odd
//     so that the 32 lanes, CR int32 words apart, hit 32 different shared-memory banks;
//   * since that mapping moves from row to row, the previous row (H as int32, E1'/E2' as one byte of deltas to H) lives in shared memory in a
//     window of WIN columns that follows the band: buffer index = column - cbase + PAD.  The window is re-anchored
//     (data moved to the front) when the band reaches its end, and re-built on rows whose predecessor is not the
//     previous row.  Out-of-band cells hold -inf plus one guard cell per side, so readers need no band test;
//   * the query bytes of the window are staged next to it (the band advances a column or two per row);
//   * F1/F2 are prefix-max scans: serial inside the lane block, Kogge-Stone across the warp.
// Roofline: INT32 ALU issue.  Algorithmic work = band cells.
#include <climits>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "internal.h"
#include "k_poa_trace.cuh"

namespace vga {

static const int PL_WARPS = 2;              // warps per CTA (7.8 KB of shared memory each)
static const u64 LONG_POOL_CHUNK = 8192;    // trace words a warp takes from the pool per refill
static const u64 LONG_OVF_CHUNK = 1u << 18;  // fused kernel: words an alignment takes from the overflow region at a time
static const u64 LONG_ARENA_MIN = 4u << 20;  // fused kernel: smallest useful arena (words): a 10 kb alignment writes 3-4 M
static const int LPAD = 8;                  // buffer index of column cbase (column cbase-1 .. are guards)
static const int WIN = 1280;                // columns of the window (bands are at most 1024 + the odd block size slack)
static const int LBUF = LPAD + WIN + 8;     // index of column beg + 32*CR + 1 stays below LPAD + WIN (re-anchoring rule below)
static const u32 FULLW = 0xffffffffu;

__device__ __forceinline__ i32 imax(i32 a, i32 b) { return a > b ? a : b; }
__device__ __forceinline__ i32 imin(i32 a, i32 b) { return a < b ? a : b; }
__device__ __forceinline__ i32 sext16(u32 w) { return (i32)(short)(w & 0xffffu); }

struct LongSmem {
    i32 h[LBUF];                          // H of the previous row; updated in place (loads precede the warp's scan, stores follow it)
    u8 dd[LBUF];                          // E1' - (H - oe1) in bits 0..2 and E2' - (H - oe2) in bits 3..7 of the same cell: both
                                          // are in [0, o] for every reachable cell, also after a predecessor merge
    u8 q[LBUF];                           // query byte of column j (base j-1) at the same index
};

struct LongConst { i32 e1, e2, oe1, oe2, o1, o2, smatch, smis; };
struct LongRowOut { i32 mx, lpos, rpos, lmin; };

// One DP row.  `x0` is the buffer index of column beg; lane l computes the CR columns from beg + l*CR.  Only in-band
// cells are stored; the generator of F is masked outside the band, so an out-of-band H is an F that decays from the
// last band cell and never reaches the row maximum (e1, e2 >= 1).  CR is a run-time value and the column loops are
// NOT unrolled per block size: sweep 2 recomputes Hhat and the lane-local F scan of sweep 1 instead of keeping them in
// per-column registers, so there is ONE small loop body for every band width (a version with an unrolled body per
// block size spent 60 % of its cycles waiting for instruction fetch) and 6 bytes of shared memory per window column.
// SRC: the previous row is the virtual source row, whose E is -inf except in column 0 (not expressible as H + delta).
template <bool SRC>
__device__ __forceinline__ LongRowOut long_row(const LongConst& K, u32 lane, i32 CR, i32 beg, i32 end, i32 x0, u32 base, i32 ref,
                                               LongSmem& sm, u32* __restrict__ tr) {
    const i32 j0 = beg + (i32)lane * CR;
    const i32 idx = x0 + (i32)lane * CR;
    const i32 thi = end - j0;                              // column t of the lane is in the band iff t <= thi
    const i32 n = imin(CR, thi + 1);                       // the lane's band columns (<= 0: none); only those are computed
    i32* const sh = sm.h + idx;
    u8* const sd = sm.dd + idx;
    const u8* const pq = sm.q + idx;
    const i32 hleft = sh[-1];
    // E1, E2 arriving from the previous row at column t: H_old - oe + delta
    auto e_of = [&](i32 ho, u32 dv, i32 t, i32& x1, i32& x2) {
        if (SRC) {
            const bool c0 = j0 + t == 0;
            x1 = c0 ? -K.oe1 : POA_NEG_INF; x2 = c0 ? -K.oe2 : POA_NEG_INF;
        } else {
            x1 = ho - K.oe1 + (i32)(dv & 7u); x2 = ho - K.oe2 + (i32)(dv >> 3);
        }
    };
    // ---- sweep 1: lane-local part of the F scans over Hhat = max(M, E1, E2)
    i32 run1 = POA_DEEP, run2 = POA_DEEP;
    {
        i32 hd = hleft;
#pragma unroll 4
        for (i32 t = 0; t < n; t++) {
            const i32 ho = sh[t];
            i32 x1, x2;
            e_of(ho, sd[t], t, x1, x2);
            const i32 m = hd + ((u32)pq[t] == base ? K.smatch : K.smis);
            const i32 hh = __vimax3_s32(m, x1, x2);
            hd = ho;
            run1 = __viaddmax_s32(run1, -K.e1, hh - K.oe1);
            run2 = __viaddmax_s32(run2, -K.e2, hh - K.oe2);
        }
    }
    // carry of F1/F2 across the warp (Kogge-Stone; shfl_up leaves lanes < d unchanged, and max(v, v - c) = v)
    i32 cin1 = run1, cin2 = run2;
    {
        const i32 s1 = K.e1 * CR, s2 = K.e2 * CR;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const i32 t1 = __shfl_up_sync(FULLW, cin1, d), t2 = __shfl_up_sync(FULLW, cin2, d);
            cin1 = __viaddmax_s32(t1, -s1 * d, cin1);
            cin2 = __viaddmax_s32(t2, -s2 * d, cin2);
        }
    }
    cin1 = __shfl_up_sync(FULLW, cin1, 1); cin2 = __shfl_up_sync(FULLW, cin2, 1);
    if (lane == 0) { cin1 = POA_DEEP; cin2 = POA_DEEP; }
    // ---- sweep 2: H, E', trace words, row maximum keyed by column, smallest reachable H
    i32 kl = INT_MIN, kr = INT_MIN;          // max of (H, -t) and (H, t): lane maximum with its leftmost / rightmost column
    u32 lminu = 0xffffffffu;                 // min of H + 2^27 as unsigned: unreachable cells (H <= -2^27) wrap to huge values
    i32 hprev_new = 1 << 29;                 // new H of column t-1; column 0 of the lane: so large that both F fields come out as 0 —
                                             // they are relative to the LEFT lane's last new H and are added after the loop
    u32 w0 = 0;
    u32* const trl = tr + lane * CR;
    {
        i32 hd = hleft;
        run1 = POA_DEEP; run2 = POA_DEEP;
#pragma unroll 2
        for (i32 t = 0; t < n; t++) {
            const i32 ho = sh[t];
            i32 x1, x2;
            e_of(ho, sd[t], t, x1, x2);
            const i32 m = hd + ((u32)pq[t] == base ? K.smatch : K.smis);
            const i32 hh = __vimax3_s32(m, x1, x2);
            hd = ho;
            const i32 f1 = __viaddmax_s32(cin1, -K.e1 * t, run1), f2 = __viaddmax_s32(cin2, -K.e2 * t, run2);
            const i32 h = imax(__vimax3_s32(hh, f1, f2), POA_DEEP);
            run1 = __viaddmax_s32(run1, -K.e1, hh - K.oe1);
            run2 = __viaddmax_s32(run2, -K.e2, hh - K.oe2);
            const i32 d1 = __viaddmax_s32(x1, K.o1 - h, 0);          // E1' - (H - oe1) = max(E1 - e1 - H + oe1, 0)
            const i32 d2 = __viaddmax_s32(x2, K.o2 - h, 0);
            u32 hi = (u32)d2 * 8u + (u32)d1;
            const u32 dnew = hi;
            hi = (u32)__viaddmax_s32(f1, 1 + K.oe1 - hprev_new, 0) * 256u + hi;
            hi = (u32)__viaddmax_s32(f2, 1 + K.oe2 - hprev_new, 0) * 2048u + hi;
            const i32 hrel = imax(h - ref, -32768);
            const u32 wd = __byte_perm((u32)hrel, hi, 0x5410);
            if (t == 0) w0 = wd;
            hprev_new = h;
            sh[t] = h;
            sd[t] = (u8)dnew;
            trl[t] = wd;                                             // column 0 is stored again below, with its F fields
            const i32 hk = hrel * 64;                                // the row maximum is far above ref - 32768
            kl = imax(kl, hk + 63 - t);
            kr = imax(kr, hk + t);
            lminu = min(lminu, (u32)(h + (1 << 27)));
        }
    }
    {   // F fields of the lane's first column: relative to the left lane's last H
        const i32 hl = __shfl_up_sync(FULLW, hprev_new, 1);
        const i32 f1_0 = imax(cin1, POA_DEEP), f2_0 = imax(cin2, POA_DEEP);     // F of column 0: the carry, no lane-local part yet
        if (lane > 0) w0 += ((u32)imax(f1_0 + 1 + K.oe1 - hl, 0) << 24) + ((u32)imax(f2_0 + 1 + K.oe2 - hl, 0) << 27);
        if (n > 0) trl[0] = w0;
    }
    // guards: H(beg-1) = -inf (diagonal of column beg in the next row) and H(end+1) = -inf (E'(end+1) = H - oe + delta, the
    // vertical of column end+1; as a diagonal it only serves column end+2, which is outside the next band)
    if (lane == 0) sm.h[x0 - 1] = POA_NEG_INF;
    if (lane == 1) sm.h[x0 + (end - beg) + 1] = POA_NEG_INF;
    // ---- row maximum and its leftmost / rightmost column (abPOA simd_abpoa_ada_max_i)
    LongRowOut o;
    const i32 lmax = kl == INT_MIN ? INT_MIN : (kl >> 6);                // relative to ref
    o.mx = __reduce_max_sync(FULLW, lmax) + ref;
    o.lpos = __reduce_min_sync(FULLW, lmax == o.mx - ref ? j0 + 63 - (kl & 63) : INT_MAX);
    o.rpos = __reduce_max_sync(FULLW, lmax == o.mx - ref ? j0 + (kr & 63) : -1);
    const u32 mnu = __reduce_min_sync(FULLW, lminu);
    o.lmin = mnu >= 0x80000000u ? INT_MAX : (i32)mnu - (1 << 27);       // no reachable cell in the row: nothing to check
    return o;
}

__device__ __forceinline__ int long_pick_cr(i32 W) {
    // columns per lane: the smallest odd count that covers the band (odd: lane stride of the int32 plane without bank conflicts)
    return W > 1024 ? 0 : (((W + 31) >> 5) | 1) < 3 ? 3 : (((W + 31) >> 5) | 1);
}

// FUSED: every warp owns a private arena of the trace pool (arena_words from arena 0 of warp 0 on), fills an alignment into it and
// traces it right away (lane 0; k_poa_trace.cuh) before it takes the next one — no waves, no traceback launches that leave the
// machine to a few thousand threads, no host round trip between fill and traceback.  An alignment that outgrows its arena goes on
// in a shared overflow region (a quarter of the pool, bump-allocated); only when that is exhausted it ends as IT_RETRY_POOL and goes
// through the wave path afterwards.
template <bool FUSED, int MINB>
__global__ void __launch_bounds__(PL_WARPS * 32, MINB) poa_fill_long_kernel(DevIndex ix, PoaBuffers b, TraceBuffers tb, u64 arena_words) {
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    LongSmem& sm = ((LongSmem*)dyn_smem)[threadIdx.x >> 5];
    const u32 lane = threadIdx.x & 31;
    LongConst K;
    K.e1 = b.e1; K.e2 = b.e2; K.oe1 = b.o1 + b.e1; K.oe2 = b.o2 + b.e2; K.o1 = b.o1; K.o2 = b.o2;
    K.smatch = b.match; K.smis = -b.mismatch;
    u64 pool_cur = 0, pool_end = 0;                 // this warp's current chunk of the trace pool (words)
    const u64 arena_base = FUSED ? ((u64)blockIdx.x * PL_WARPS + (threadIdx.x >> 5)) * arena_words : 0;
    for (;;) {
        // ---- next (read, chain) of the work list
        u32 wi = 0;
        if (lane == 0) wi = atomicAdd(b.work_ctr, 1u);
        wi = __shfl_sync(FULLW, wi, 0);
        if (wi >= b.n_items) break;
        if (FUSED) { pool_cur = arena_base; pool_end = arena_base + arena_words; }     // the arena starts over with every alignment
        const u32 w = b.work[wi];
        const AlignItem it = b.items[w];
        const i32 qlen = (i32)it.qlen;
        const u8* __restrict__ q = (const u8*)b.seqs + it.q_off;
        const i32 wband = b.wb + (i32)(b.wf * (float)qlen);
        const u32 nb = it.node_base, S = it.S, row_base = it.row_base, pred_base = it.pred_base, nrows = it.nrows;
        u32 status = VGA_ALN_OK;
        u64 cells = 0;
        u32 s = 0, o = 0, metas = 0;
        i32 beg = 0, end = 0, ppb = 0, ppe = -1, lpos = 0, rpos = 0, mxrow = 0;
        i32 cbase = 0;                               // column at buffer index LPAD
        i32 qfill = -1;                              // query bytes of columns cbase .. qfill are staged
        i32 end0 = 0;
        if (S == 0 || nrows == 0) status = it.status == VGA_ALN_UNSUPPORTED ? VGA_ALN_UNSUPPORTED : VGA_ALN_EMPTY;   // K4 refused the chain (reverse-strand anchors)
        else {
            const i32 remain_src = (i32)(b.row_meta[row_base] >> 11) + 1;
            end0 = band_round_end0(imin(qlen, imax(0, qlen - remain_src) + wband), b.bgran_mask, qlen);
        }
        for (u32 row = 0; row < nrows && status == VGA_ALN_OK; row++) {
            if ((row & 31) == 0) metas = (row + lane < nrows) ? b.row_meta[row_base + row + lane] : 0u;
            const u32 meta = __shfl_sync(FULLW, metas, row & 31);
            const u32 base = meta & 0xffu, kind = (meta >> 8) & 3u;
            // predecessor row: the previous row unless this is a first row with other predecessors
            i32 mpl = lpos + 1, mpr = rpos + 1, min_pb = ppb, max_pe = ppe, ref = mxrow;
            u32 prow = row - 1, np = 1, po = 0;
            if (kind == 1) { mpl = 1; mpr = 1; min_pb = 0; max_pe = end0; np = 0; prow = 0; ref = 0; }      // virtual source row
            else if (kind == 2) {
                __syncwarp();       // NodeState / RowInfo / trace words of earlier rows were written by other lanes
                po = b.nd_pred_off[nb + s];
                np = b.nd_pred_off[nb + s + 1] - po;
                const u32* __restrict__ pl = b.nd_pred + pred_base + po;
                mpl = INT_MAX; mpr = 0; min_pb = INT_MAX; max_pe = -1; ref = INT_MIN;
                for (u32 x = 0; x < np; x++) {
                    const NodeState ns = b.nd_state[nb + pl[x]];
                    mpl = imin(mpl, ns.lpos + 1); mpr = imax(mpr, ns.rpos + 1);
                    min_pb = imin(min_pb, ns.beg); max_pe = imax(max_pe, ns.end);
                    ref = imax(ref, ns.mx);
                }
                prow = b.nd_row_start[nb + pl[0] + 1] - 1;
            }
            // adaptive band (abPOA GET_AD_DP_BEGIN/END + clipping to the predecessors' bands)
            const i32 r = qlen - (i32)(meta >> 11);
            beg = imax(imax(0, imin(mpl, r) - wband), min_pb);
            end = imin(imin(qlen, imax(mpr, r) + wband), max_pe + 1);
            band_round(beg, end, b.bgran_mask, qlen, max_pe);
            const i32 W = end - beg + 1;
            const int cr = long_pick_cr(W);
            if (beg > end) status = VGA_ALN_BAND_ERR;
            else if (cr == 0) status = IT_RETRY_WIDER;       // wider than 1024 columns: the CTA-per-alignment kernel (k_poa_wide.cu)
            // pool space: one trace word per band cell, rows stored from an 8-column boundary (trace_col0)
            const u64 need = (u64)((W + (beg & 7) + 3) & ~3);
            if (status == VGA_ALN_OK && pool_cur + need > pool_end) {
                // FUSED: the arena is full — the alignment goes on in chunks of the shared overflow region behind the arenas (the
                // counter starts there); when that is exhausted too, the wave path takes the alignment
                unsigned long long got = 0;
                const u64 chunk = FUSED ? LONG_OVF_CHUNK : LONG_POOL_CHUNK;
                const u64 take = need > chunk ? need : chunk;
                if (lane == 0) got = atomicAdd(b.pool_ctr, (unsigned long long)take);
                got = __shfl_sync(FULLW, got, 0);
                if (got + take > b.pool_words) status = IT_RETRY_POOL;
                else { pool_cur = got; pool_end = got + take; }
            }
            if (status != VGA_ALN_OK) break;
            u32* __restrict__ tr = b.pool + pool_cur + (beg & 7);
            if (lane == 0) {
                uint4* dst = (uint4*)(b.rows + row_base + row);
                dst[0] = make_uint4((u32)beg, (u32)end, (u32)(pool_cur >> 2), (u32)ref);
                dst[1] = make_uint4(prow, (np < 0xffffffu ? np : 0xffffffu) | (base << 24), s, o);
            }
            pool_cur += need;
            cells += (u64)W;
            // ---- the window: columns beg-1 .. end+1 (+ the last lane's block) must lie inside the buffer
            __syncwarp();
            if (kind != 0) {
                // the predecessor row is not the one in shared memory: anchor the window at this band and build it there
                cbase = imax(beg - 1, 0) & ~3;
                qfill = cbase - 1;
                const i32 jlo = imax(beg - 1, 0);
                if (lane == 0 && cbase == 0) sm.h[LPAD - 1] = POA_NEG_INF;           // column -1: M(i,0) does not exist
                if (kind == 1) {          // virtual source row: H only, its E is handled by the SRC instance of the row body
                    for (i32 j = jlo + (i32)lane; j <= end + 1; j += 32) {
                        const bool in0 = j <= end0;
                        const int x = j - cbase + LPAD;
                        sm.h[x] = !in0 ? POA_NEG_INF : j == 0 ? 0 : imax(-(b.o1 + b.e1 * j), -(b.o2 + b.e2 * j));
                        sm.dd[x] = 0;
                    }
                } else {                  // merge of the predecessors' last rows, decoded from their trace words
                    const u32* __restrict__ pl = b.nd_pred + pred_base + po;
                    for (u32 x = 0; x < np; x++) {
                        const u32 p = pl[x];
                        const int2 pbe = *(const int2*)(b.nd_state + nb + p);
                        const RowInfo pri = b.rows[row_base + b.nd_row_start[nb + p + 1] - 1];
                        const u32* __restrict__ tw = b.pool + (u64)pri.toff * 4;
                        const i32 pc0 = trace_col0(pbe.x);
                        for (i32 j = jlo + (i32)lane; j <= end + 1; j += 32) {
                            // H = max over predecessors, E' = max over predecessors: kept as H and E' - (H - oe), which stays in
                            // [0, o] (every predecessor has E' >= its H - oe, and E' - (H_max - oe) <= E' - (its H - oe) <= o)
                            i32 hv = POA_NEG_INF, a1 = 0, a2 = 0;
                            if (j >= pbe.x && j <= pbe.y) {
                                const u32 v = tw[j - pc0];
                                hv = pri.ref + sext16(v);
                                a1 = (i32)((v >> 16) & 7u); a2 = (i32)((v >> 19) & 31u);
                            }
                            const int xx = j - cbase + LPAD;
                            if (x > 0) {
                                const i32 ho = sm.h[xx];
                                const u32 dv = sm.dd[xx];
                                const i32 hm = imax(ho, hv);
                                // absolute E' of both sides relative to hm - oe (the oe cancels)
                                const i32 n1 = imax(ho + (i32)(dv & 7u), hv + a1) - hm, n2 = imax(ho + (i32)(dv >> 3), hv + a2) - hm;
                                hv = hm; a1 = imax(n1, 0); a2 = imax(n2, 0);
                            }
                            sm.h[xx] = hv; sm.dd[xx] = (u8)((u32)a1 | ((u32)a2 << 3));
                        }
                    }
                }
                __syncwarp();
            } else if (beg + 32 * cr + 2 - cbase > WIN) {        // lane blocks reach up to column beg + 32*CR - 1
                // the band reached the end of the window: move columns beg-1 .. ppe+1 (the previous row and its guards) to the front
                const i32 nbase = imax(beg - 1, 0) & ~3;
                const i32 shift = nbase - cbase;                 // > 0
                const i32 lo = imax(beg - 1, 0), hiq = ppe + 1;
                const i32 hi_all = imax(hiq, qfill);
                for (i32 j0 = lo; j0 <= hi_all; j0 += 32) {
                    const i32 j = j0 + (i32)lane;
                    const int xs = j - cbase + LPAD;
                    i32 a = 0; u32 c = 0, qq = 0;
                    if (j <= hiq) { a = sm.h[xs]; c = sm.dd[xs]; }
                    if (j <= qfill) qq = sm.q[xs];
                    __syncwarp();
                    if (j <= hiq) { sm.h[xs - shift] = a; sm.dd[xs - shift] = (u8)c; }
                    if (j <= qfill) sm.q[xs - shift] = (u8)qq;
                    __syncwarp();
                }
                cbase = nbase;
                if (qfill < lo - 1) qfill = lo - 1;
            }
            // query bytes of the columns this row can touch (lane blocks may reach past end; those cells are masked).  Staged
            // 128 columns ahead, so the global loads are off the critical path of all but one row in ~100
            {
                const i32 need_q = imin(end + 1, qlen);
                if (need_q > qfill) {
                    const i32 upto = imin(imin(need_q + 128, qlen), cbase + WIN - 1);
                    for (i32 j = qfill + 1 + (i32)lane; j <= upto; j += 32) sm.q[j - cbase + LPAD] = j >= 1 ? q[j - 1] : (u8)0;
                    qfill = upto;
                }
            }
            __syncwarp();
            const i32 x0 = beg - cbase + LPAD;
            LongRowOut ro;
            if (kind == 1) ro = long_row<true>(K, lane, cr, beg, end, x0, base, ref, sm, tr);
            else ro = long_row<false>(K, lane, cr, beg, end, x0, base, ref, sm, tr);
            // a reachable cell further than int16 below the row reference cannot be encoded
            if (ro.lmin != INT_MAX && ro.lmin - ref < -32000) { status = VGA_ALN_UNSUPPORTED; break; }
            lpos = ro.lpos; rpos = ro.rpos; mxrow = ro.mx;
            ppb = beg; ppe = end;
            if (meta & 0x400u) {
                // last row of its node: it may be a predecessor of later nodes' first rows
                if (lane == 0) {
                    int4* dst = (int4*)(b.nd_state + nb + s);
                    dst[0] = make_int4(beg, end, lpos, rpos);
                    dst[1] = make_int4(mxrow, 0, 0, 0);
                }
                s++; o = 0;
            } else o++;
        }
        // ---- result record
        i32 best = INT_MIN;
        u32 bs = 0xffffffffu;
        if (status == VGA_ALN_OK) {
            // global end: sink predecessors (nodes without successors, ascending) that reach column qlen
            const u32* __restrict__ rs = b.nd_row_start + nb;
            const RowInfo* __restrict__ rows = b.rows + row_base;
            __syncwarp();
            for (u32 s0 = 0; s0 < S; s0 += 32) {
                const u32 sx = s0 + lane;
                i32 h = INT_MIN;
                if (sx < S && b.nd_first_succ[nb + sx] < 0) {
                    const RowInfo ri = rows[rs[sx + 1] - 1];
                    if (ri.end >= qlen) h = ri.ref + sext16(b.pool[(u64)ri.toff * 4 + (u32)(qlen - trace_col0(ri.beg))]);
                }
                const i32 mx = __reduce_max_sync(FULLW, h);
                if (mx > best) {                             // strict: the first node wins ties
                    best = mx;
                    bs = (u32)__reduce_min_sync(FULLW, h == mx ? (i32)sx : INT_MAX);
                }
            }
            if (bs == 0xffffffffu) status = VGA_ALN_EMPTY;
        }
        if (lane == 0) {
            AlignItem* out = b.items + w;
            out->status = status;
            out->score = status == VGA_ALN_OK ? best : 0;
            out->end_node = bs;
            out->cells = cells;
            out->n_runs = 0; out->n_path = 0; out->path_len = 0; out->start_off = 0; out->end_off = 0; out->n_aligned = 0;
        }
        __syncwarp();
        if (FUSED && status == VGA_ALN_OK) {
            if (lane == 0) poa_trace_one<false>(tb, w);        // rows, trace words and the record above are visible after the barrier
            __syncwarp();
        }
    }
}

template <bool FUSED, int MINB>
static int launch_long(const DevIndex& ix, const PoaBuffers& b, const TraceBuffers& tb, int n_sms, cudaStream_t st) {
    int per_sm = 0;
    const size_t smem_bytes = (size_t)PL_WARPS * sizeof(LongSmem);
    static bool attr_set = false;
    if (!attr_set) {
        VGA_CUDA(cudaFuncSetAttribute(poa_fill_long_kernel<FUSED, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
        attr_set = true;
    }
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, poa_fill_long_kernel<FUSED, MINB>, PL_WARPS * 32, smem_bytes);
    if (per_sm < 1) per_sm = 1;
    long long ctas = (long long)n_sms * per_sm;
    const long long need = ((long long)b.n_items + PL_WARPS - 1) / PL_WARPS;
    if (ctas > need) ctas = need;
    if (ctas < 1) ctas = 1;
    u64 arena = 0;
    if (FUSED) {
        // one arena per resident warp; an alignment larger than its arena is left to the wave path (IT_RETRY_POOL).  Fewer, larger
        // arenas when the pool is small for the machine: at least LONG_ARENA_MIN words each
        const u64 arena_words_total = b.pool_words / 4 * 3;          // a quarter of the pool is the shared overflow region
        arena = (arena_words_total / (u64)(ctas * PL_WARPS)) & ~(u64)3;
        while (arena < LONG_ARENA_MIN && ctas > 1) { ctas = (ctas + 1) / 2; arena = (arena_words_total / (u64)(ctas * PL_WARPS)) & ~(u64)3; }
        const unsigned long long ovf_base = (unsigned long long)arena * (u64)(ctas * PL_WARPS);
        VGA_CUDA(cudaMemcpyAsync(b.pool_ctr, &ovf_base, 8, cudaMemcpyHostToDevice, st));
    }
    poa_fill_long_kernel<FUSED, MINB><<<(unsigned)ctas, PL_WARPS * 32, smem_bytes, st>>>(ix, b, tb, arena);
    return 1;
}

// fused_tb != NULL: fill + traceback per warp in private arenas of the pool (the pool must hold nothing that still awaits a traceback)
int launch_poa_fill_long(const DevIndex& ix, const PoaBuffers& b, int n_sms, cudaStream_t st, const TraceBuffers* fused_tb) {
    if (b.n_items == 0) return 0;
    // cut for 8 CTAs (16 warps) per SM: 115 registers, no spills.  Measured on C5 (25 k reads): 218 ms, against 261 ms when cut for 12
    // CTAs (80 registers, 48 B of spills) and 307 ms for 16 (64 registers, 112 B) — gpurun_out/r02y
    if (fused_tb) return launch_long<true, 8>(ix, b, *fused_tb, n_sms, st);
    TraceBuffers none;
    memset(&none, 0, sizeof(none));
    return launch_long<false, 8>(ix, b, none, n_sms, st);
}

}  // namespace vga
