// K5 (bands of any width) — sequence-to-graph POA fill for the alignments whose adaptive band grows beyond the 1024
// columns of the warp-per-alignment kernels (k_poa_long.cu, k_poa.cu), replacing the abPOA call at the reference's FFI
// seam (src/align.rs:190-202).  abPOA itself has no band limit (its band is clipped to the read, App. E of SURVEY.md):
// a chain through a few random k-mer hits of a reverse-strand 10 kb read still gets its (meaningless, score < 0) global
// alignment with rows thousands of columns wide, and parity wants exactly that alignment.  Same alignment definition,
// trace word format (scores relative to ref(row)), RowInfo / NodeState records and work-queue protocol as
// k_poa_long.cu, so the traceback kernel is shared.
//
// One CTA per (read, chain).  The whole previous row lives in shared memory at ABSOLUTE column indices (H as int32,
// E1'/E2' as one byte of deltas to H, the query bytes beside it: 6 B per query column, 60 KB for a 10 kb read), so
// there is no window to move.  A row is split into PW_THREADS contiguous lane blocks of CR columns; F1/F2 are
// prefix-max scans (serial in the lane block, Kogge-Stone in the warp, the warp totals combined through shared
// memory).  Two CTA barriers per row.
// Roofline: INT32 ALU issue.  Algorithmic work = band cells.
#include <climits>
#include <cstdlib>

#include "common.cuh"
#include "internal.h"

namespace vga {

static const int PW_THREADS = 128;
static const int PW_WARPS = PW_THREADS / 32;
static const int WPAD = 8;                  // buffer index of column 0 (column -1 is a guard)
static const u64 WIDE_POOL_CHUNK = 1u << 16;
static const u32 FULLW = 0xffffffffu;

__device__ __forceinline__ i32 imax(i32 a, i32 b) { return a > b ? a : b; }
__device__ __forceinline__ i32 imin(i32 a, i32 b) { return a < b ? a : b; }
__device__ __forceinline__ i32 sext16(u32 w) { return (i32)(short)(w & 0xffffu); }

struct WideConst { i32 e1, e2, oe1, oe2, o1, o2, smatch, smis; };
struct WideShared {                          // fixed part; the row planes follow
    i32 tot1[PW_WARPS], tot2[PW_WARPS];      // F1 / F2 leaving each warp (inclusive over the warp's lanes)
    i32 rmx[PW_WARPS], rl[PW_WARPS], rr[PW_WARPS];
    u32 rmin[PW_WARPS];
    unsigned long long bcast;                // work index / pool offset from thread 0
    unsigned long long key[PW_WARPS];
};
struct WideRowOut { i32 mx, lpos, rpos, lmin; };

// One DP row; thread `tid` computes the CR columns from beg + tid*CR (only those inside the band).
// SRC: the previous row is the virtual source row, whose E is -inf except in column 0.
template <bool SRC>
__device__ __forceinline__ WideRowOut wide_row(const WideConst& K, u32 tid, i32 CR, i32 beg, i32 end, u32 base, i32 ref,
                                               WideShared& ws, i32* __restrict__ smh, u8* __restrict__ smd,
                                               const u8* __restrict__ smq, u32* __restrict__ tr) {
    const u32 lane = tid & 31, wv = tid >> 5;
    const i32 j0 = beg + (i32)tid * CR;
    const i32 idx = j0 + WPAD;
    const i32 thi = end - j0;
    const i32 n = imin(CR, thi + 1);                       // the thread's band columns (<= 0: none)
    i32* const sh = smh + idx;
    u8* const sd = smd + idx;
    const u8* const pq = smq + idx;
    const i32 hleft = n > 0 ? sh[-1] : 0;
    auto e_of = [&](i32 ho, u32 dv, i32 t, i32& x1, i32& x2) {
        if (SRC) {
            const bool c0 = j0 + t == 0;
            x1 = c0 ? -K.oe1 : POA_NEG_INF; x2 = c0 ? -K.oe2 : POA_NEG_INF;
        } else {
            x1 = ho - K.oe1 + (i32)(dv & 7u); x2 = ho - K.oe2 + (i32)(dv >> 3);
        }
    };
    // ---- sweep 1: thread-local part of the F scans over Hhat = max(M, E1, E2)
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
    // ---- carry of F1 / F2 across the CTA: inclusive decayed max inside the warp, then the warps in order
    const i32 s1 = K.e1 * CR, s2 = K.e2 * CR;
    i32 in1 = run1, in2 = run2;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const i32 t1 = __shfl_up_sync(FULLW, in1, d), t2 = __shfl_up_sync(FULLW, in2, d);
        in1 = __viaddmax_s32(t1, -s1 * d, in1);          // shfl_up leaves lanes < d unchanged, and max(v, v - c) = v
        in2 = __viaddmax_s32(t2, -s2 * d, in2);
    }
    if (lane == 31) { ws.tot1[wv] = in1; ws.tot2[wv] = in2; }
    i32 cin1 = __shfl_up_sync(FULLW, in1, 1), cin2 = __shfl_up_sync(FULLW, in2, 1);
    if (lane == 0) { cin1 = POA_DEEP; cin2 = POA_DEEP; }
    __syncthreads();                                       // barrier 1: warp totals visible; every hleft is in a register
    {
        i32 p1 = POA_DEEP, p2 = POA_DEEP;                  // F leaving the last thread of the previous warp (all warps before it folded in)
        for (u32 u = 0; u < wv; u++) {
            p1 = __viaddmax_s32(p1, -s1 * 32, ws.tot1[u]);
            p2 = __viaddmax_s32(p2, -s2 * 32, ws.tot2[u]);
        }
        if (wv > 0) {
            cin1 = __viaddmax_s32(p1, -s1 * (i32)lane, cin1);
            cin2 = __viaddmax_s32(p2, -s2 * (i32)lane, cin2);
        }
        cin1 = imax(cin1, POA_DEEP); cin2 = imax(cin2, POA_DEEP);
    }
    // ---- sweep 2: H, E', trace words, row maximum keyed by column, smallest reachable H
    i32 kl = INT_MIN, kr = INT_MIN;          // max of (H, -t) and (H, t): thread maximum with its leftmost / rightmost column
    u32 lminu = 0xffffffffu;
    i32 hprev_new = 0;
    u32 w0 = 0; i32 f1_0 = 0, f2_0 = 0;
    u32* const trl = tr + (size_t)tid * CR;
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
            if (t > 0) {
                hi = (u32)__viaddmax_s32(f1, 1 + K.oe1 - hprev_new, 0) * 256u + hi;
                hi = (u32)__viaddmax_s32(f2, 1 + K.oe2 - hprev_new, 0) * 2048u + hi;
            }
            const i32 hrel = imax(h - ref, -32768);
            const u32 wd = __byte_perm((u32)hrel, hi, 0x5410);
            if (t == 0) { w0 = wd; f1_0 = f1; f2_0 = f2; }
            hprev_new = h;
            sh[t] = h;
            sd[t] = (u8)dnew;
            if (t > 0) trl[t] = wd;
            const i32 hk = hrel * 1024;                              // the row maximum is far above ref - 32768; t < 1024
            kl = imax(kl, hk + 1023 - t);
            kr = imax(kr, hk + t);
            lminu = min(lminu, (u32)(h + (1 << 27)));
        }
    }
    // guards for the next row: H(beg-1) = -inf (diagonal of column beg) and H(end+1) = -inf (E'(end+1) = H - oe + delta).
    // Nobody reads them from shared memory during sweep 2 (hleft is in a register, out-of-band cells are not computed).
    if (tid == 0) smh[beg - 1 + WPAD] = POA_NEG_INF;
    if (tid == 1) smh[end + 1 + WPAD] = POA_NEG_INF;
    // ---- row maximum with its leftmost / rightmost column: per warp, then over the warps
    {
        const i32 lmax = kl == INT_MIN ? INT_MIN : (kl >> 10);               // relative to ref
        const i32 wmx = __reduce_max_sync(FULLW, lmax);
        const i32 wl = __reduce_min_sync(FULLW, (lmax == wmx && kl != INT_MIN) ? j0 + 1023 - (kl & 1023) : INT_MAX);
        const i32 wr = __reduce_max_sync(FULLW, (lmax == wmx && kl != INT_MIN) ? j0 + (kr & 1023) : -1);
        const u32 wmn = __reduce_min_sync(FULLW, lminu);
        if (lane == 0) { ws.rmx[wv] = wmx; ws.rl[wv] = wl; ws.rr[wv] = wr; ws.rmin[wv] = wmn; }
    }
    __syncthreads();                                       // barrier 2: the new row, its guards and the warp maxima are visible
    {   // F fields of the thread's first column: relative to the left thread's last new H (now in the row buffer)
        if (n > 0) {
            if (tid > 0) {
                const i32 hl = sh[-1];
                w0 += ((u32)imax(f1_0 + 1 + K.oe1 - hl, 0) << 24) + ((u32)imax(f2_0 + 1 + K.oe2 - hl, 0) << 27);
            }
            trl[0] = w0;
        }
    }
    WideRowOut o;
    i32 mx = INT_MIN;
#pragma unroll
    for (int u = 0; u < PW_WARPS; u++) mx = imax(mx, ws.rmx[u]);
    i32 lp = INT_MAX, rp = -1;
    u32 mnu = 0xffffffffu;
#pragma unroll
    for (int u = 0; u < PW_WARPS; u++) {
        if (ws.rmx[u] == mx) { lp = imin(lp, ws.rl[u]); rp = imax(rp, ws.rr[u]); }
        mnu = min(mnu, ws.rmin[u]);
    }
    o.mx = mx + ref; o.lpos = lp; o.rpos = rp;
    o.lmin = mnu >= 0x80000000u ? INT_MAX : (i32)mnu - (1 << 27);
    return o;
}

__global__ void __launch_bounds__(PW_THREADS) poa_fill_wide_kernel(DevIndex ix, PoaBuffers b, int lbuf) {
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    WideShared& ws = *(WideShared*)dyn_smem;
    i32* const smh = (i32*)(dyn_smem + ((sizeof(WideShared) + 15) & ~(size_t)15));
    u8* const smd = (u8*)(smh + lbuf);
    u8* const smq = smd + lbuf;
    const u32 tid = threadIdx.x, lane = tid & 31, wv = tid >> 5;
    WideConst K;
    K.e1 = b.e1; K.e2 = b.e2; K.oe1 = b.o1 + b.e1; K.oe2 = b.o2 + b.e2; K.o1 = b.o1; K.o2 = b.o2;
    K.smatch = b.match; K.smis = -b.mismatch;
    u64 pool_cur = 0, pool_end = 0;                 // this CTA's current chunk of the trace pool (words); uniform over the CTA
    for (;;) {
        // ---- next (read, chain) of the work list
        __syncthreads();
        if (tid == 0) ws.bcast = atomicAdd(b.work_ctr, 1u);
        __syncthreads();
        const u32 wi = (u32)ws.bcast;
        if (wi >= b.n_items) break;
        const u32 w = b.work[wi];
        const AlignItem it = b.items[w];
        const i32 qlen = (i32)it.qlen;
        const u8* __restrict__ q = (const u8*)b.seqs + it.q_off;
        const i32 wband = b.wb + (i32)(b.wf * (float)qlen);
        const u32 nb = it.node_base, S = it.S, row_base = it.row_base, pred_base = it.pred_base, nrows = it.nrows;
        u32 status = VGA_ALN_OK;
        u64 cells = 0;
        u32 s = 0, o = 0;
        i32 beg = 0, end = 0, ppb = 0, ppe = -1, lpos = 0, rpos = 0, mxrow = 0;
        i32 end0 = 0;
        if (S == 0 || nrows == 0) status = it.status == VGA_ALN_UNSUPPORTED ? VGA_ALN_UNSUPPORTED : VGA_ALN_EMPTY;
        else if (qlen + 2 * WPAD + 8 > lbuf) status = VGA_ALN_TOO_WIDE;     // the read does not fit the shared-memory row
        else {
            const i32 remain_src = (i32)(b.row_meta[row_base] >> 11) + 1;
            end0 = band_round_end0(imin(qlen, imax(0, qlen - remain_src) + wband), b.bgran_mask, qlen);
            for (i32 j = (i32)tid; j <= qlen + 1; j += PW_THREADS) smq[j + WPAD] = (j >= 1 && j <= qlen) ? q[j - 1] : (u8)0;
            if (tid == 0) smh[WPAD - 1] = POA_NEG_INF;                        // column -1: M(i,0) does not exist
        }
        u32 meta_next = (status == VGA_ALN_OK) ? b.row_meta[row_base] : 0u;
        for (u32 row = 0; row < nrows && status == VGA_ALN_OK; row++) {
            const u32 meta = meta_next;
            if (row + 1 < nrows) meta_next = b.row_meta[row_base + row + 1];  // the next row's record is in flight during this row
            const u32 base = meta & 0xffu, kind = (meta >> 8) & 3u;
            i32 mpl = lpos + 1, mpr = rpos + 1, min_pb = ppb, max_pe = ppe, ref = mxrow;
            u32 prow = row - 1, np = 1, po = 0;
            if (kind == 1) { mpl = 1; mpr = 1; min_pb = 0; max_pe = end0; np = 0; prow = 0; ref = 0; }      // virtual source row
            else if (kind == 2) {
                // NodeState / RowInfo / trace words of earlier rows: written before an earlier CTA barrier
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
            const i32 CR = imax(3, (W + PW_THREADS - 1) / PW_THREADS) | 1;
            if (beg > end) status = VGA_ALN_BAND_ERR;
            else if (CR >= 1024) status = VGA_ALN_TOO_WIDE;
            const u64 need = (u64)((W + (beg & 7) + 3) & ~3);
            if (status == VGA_ALN_OK && pool_cur + need > pool_end) {      // uniform over the CTA
                const u64 take = need > WIDE_POOL_CHUNK ? need : WIDE_POOL_CHUNK;
                __syncthreads();
                if (tid == 0) ws.bcast = atomicAdd(b.pool_ctr, (unsigned long long)take);
                __syncthreads();
                const u64 got = ws.bcast;
                if (got + take > b.pool_words) status = IT_RETRY_POOL;
                else { pool_cur = got; pool_end = got + take; }
            }
            if (status != VGA_ALN_OK) break;
            u32* __restrict__ tr = b.pool + pool_cur + (beg & 7);
            if (tid == 0) {
                uint4* dst = (uint4*)(b.rows + row_base + row);
                dst[0] = make_uint4((u32)beg, (u32)end, (u32)(pool_cur >> 2), (u32)ref);
                dst[1] = make_uint4(prow, (np < 0xffffffu ? np : 0xffffffu) | (base << 24), s, o);
            }
            pool_cur += need;
            cells += (u64)W;
            if (kind != 0) {
                // the predecessor row is not the one in shared memory: build it there (columns beg-1 .. end+1)
                __syncthreads();          // the previous row's first-column fix-up still reads the buffer
                const i32 jlo = imax(beg - 1, 0);
                if (kind == 1) {          // virtual source row: H only, its E is handled by the SRC instance of the row body
                    for (i32 j = jlo + (i32)tid; j <= end + 1; j += PW_THREADS) {
                        const bool in0 = j <= end0;
                        smh[j + WPAD] = !in0 ? POA_NEG_INF : j == 0 ? 0 : imax(-(b.o1 + b.e1 * j), -(b.o2 + b.e2 * j));
                        smd[j + WPAD] = 0;
                    }
                } else {                  // merge of the predecessors' last rows, decoded from their trace words
                    const u32* __restrict__ pl = b.nd_pred + pred_base + po;
                    for (u32 x = 0; x < np; x++) {
                        const u32 p = pl[x];
                        const int2 pbe = *(const int2*)(b.nd_state + nb + p);
                        const RowInfo pri = b.rows[row_base + b.nd_row_start[nb + p + 1] - 1];
                        const u32* __restrict__ tw = b.pool + (u64)pri.toff * 4;
                        const i32 pc0 = trace_col0(pbe.x);
                        for (i32 j = jlo + (i32)tid; j <= end + 1; j += PW_THREADS) {     // a thread revisits only its own columns
                            i32 hv = POA_NEG_INF, a1 = 0, a2 = 0;
                            if (j >= pbe.x && j <= pbe.y) {
                                const u32 v = tw[j - pc0];
                                hv = pri.ref + sext16(v);
                                a1 = (i32)((v >> 16) & 7u); a2 = (i32)((v >> 19) & 31u);
                            }
                            const int xx = j + WPAD;
                            if (x > 0) {
                                const i32 ho = smh[xx];
                                const u32 dv = smd[xx];
                                const i32 hm = imax(ho, hv);
                                const i32 n1 = imax(ho + (i32)(dv & 7u), hv + a1) - hm, n2 = imax(ho + (i32)(dv >> 3), hv + a2) - hm;
                                hv = hm; a1 = imax(n1, 0); a2 = imax(n2, 0);
                            }
                            smh[xx] = hv; smd[xx] = (u8)((u32)a1 | ((u32)a2 << 3));
                        }
                    }
                }
                __syncthreads();
            }
            WideRowOut ro;
            if (kind == 1) ro = wide_row<true>(K, tid, CR, beg, end, base, ref, ws, smh, smd, smq, tr);
            else ro = wide_row<false>(K, tid, CR, beg, end, base, ref, ws, smh, smd, smq, tr);
            // a reachable cell further than int16 below the row reference cannot be encoded
            if (ro.lmin != INT_MAX && ro.lmin - ref < -32000) { status = VGA_ALN_UNSUPPORTED; break; }
            lpos = ro.lpos; rpos = ro.rpos; mxrow = ro.mx;
            ppb = beg; ppe = end;
            if (meta & 0x400u) {
                // last row of its node: it may be a predecessor of later nodes' first rows
                if (tid == 0) {
                    int4* dst = (int4*)(b.nd_state + nb + s);
                    dst[0] = make_int4(beg, end, lpos, rpos);
                    dst[1] = make_int4(mxrow, 0, 0, 0);
                }
                s++; o = 0;
            } else o++;
        }
        // ---- result record
        __syncthreads();            // the last rows' trace words, RowInfo and NodeState are visible to every thread
        i32 best = INT_MIN;
        u32 bs = 0xffffffffu;
        if (status == VGA_ALN_OK) {
            // global end: sink predecessors (nodes without successors, ascending) that reach column qlen; first wins ties
            const u32* __restrict__ rs = b.nd_row_start + nb;
            const RowInfo* __restrict__ rows = b.rows + row_base;
            unsigned long long key = 0;          // (H + 2^31) << 32 | ~node : maximum = best H, smallest node
            for (u32 sx = tid; sx < S; sx += PW_THREADS) {
                if (b.nd_first_succ[nb + sx] < 0) {
                    const RowInfo ri = rows[rs[sx + 1] - 1];
                    if (ri.end >= qlen) {
                        const i32 h = ri.ref + sext16(b.pool[(u64)ri.toff * 4 + (u32)(qlen - trace_col0(ri.beg))]);
                        const unsigned long long k2 = ((unsigned long long)((u32)h ^ 0x80000000u) << 32) | (u32)~sx;
                        if (k2 > key) key = k2;
                    }
                }
            }
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                const unsigned long long o2 = __shfl_xor_sync(FULLW, key, d);
                if (o2 > key) key = o2;
            }
            if (lane == 0) ws.key[wv] = key;
            __syncthreads();
            key = 0;
#pragma unroll
            for (int u = 0; u < PW_WARPS; u++) if (ws.key[u] > key) key = ws.key[u];
            if (key == 0) status = VGA_ALN_EMPTY;
            else { best = (i32)((u32)(key >> 32) ^ 0x80000000u); bs = ~(u32)key; }
        }
        if (tid == 0) {
            AlignItem* out = b.items + w;
            out->status = status;
            out->score = status == VGA_ALN_OK ? best : 0;
            out->end_node = bs;
            out->cells = cells;
            out->n_runs = 0; out->n_path = 0; out->path_len = 0; out->start_off = 0; out->end_off = 0; out->n_aligned = 0;
        }
    }
}

// shared memory the wide class needs for reads up to qmax bases; 0 = does not fit one SM
size_t poa_wide_smem_bytes(int qmax, int* lbuf_out) {
    const int lbuf = (qmax + 2 * WPAD + 8 + 15) & ~15;
    const size_t bytes = ((sizeof(WideShared) + 15) & ~(size_t)15) + (size_t)lbuf * 6;
    if (lbuf_out) *lbuf_out = lbuf;
    return bytes;
}

int launch_poa_fill_wide(const DevIndex& ix, const PoaBuffers& b, int n_sms, cudaStream_t st) {
    if (b.n_items == 0) return 0;
    int dev = 0, max_smem = 0;
    VGA_CUDA(cudaGetDevice(&dev));
    VGA_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    int lbuf = 0;
    size_t smem_bytes = poa_wide_smem_bytes(b.qmax_wide, &lbuf);
    if (smem_bytes > (size_t)max_smem) {
        // reads too long for a shared-memory row end as VGA_ALN_TOO_WIDE inside the kernel; the others still run
        lbuf = (int)(((size_t)max_smem - ((sizeof(WideShared) + 15) & ~(size_t)15)) / 6) & ~15;
        smem_bytes = ((sizeof(WideShared) + 15) & ~(size_t)15) + (size_t)lbuf * 6;
    }
    VGA_CUDA(cudaFuncSetAttribute(poa_fill_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, poa_fill_wide_kernel, PW_THREADS, smem_bytes);
    if (per_sm < 1) per_sm = 1;
    long long ctas = (long long)n_sms * per_sm;
    if (ctas > (long long)b.n_items) ctas = b.n_items;
    poa_fill_wide_kernel<<<(unsigned)ctas, PW_THREADS, smem_bytes, st>>>(ix, b, lbuf);
    return 1;
}

}  // namespace vga
