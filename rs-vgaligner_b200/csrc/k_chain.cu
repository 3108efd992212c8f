// K3 — chaining: chain_anchors / score_anchor (reference src/chain.rs:274-655) with a warp per
// read.  The predecessor window (bandwidth, 50 in map_main.rs:103) is scored across lanes and
// reduced by shuffle max; ties keep the nearest predecessor exactly like the reference's
// descending-j loop with a strict `>` (src/chain.rs:417-443).
//
// f64 determinism: this file is compiled with -fmad=false, every arithmetic step
// uses an explicit round-to-nearest intrinsic in the reference's evaluation order, Rust's
// f64::round is CUDA's round() (half away from zero), and 0.01*k*g + 0.5*log2(g) comes from a
// host-computed table so no device transcendental is involved.  Scores are bit-identical to the
// CPU evaluation, not merely within a tolerance.
#include <cfloat>
#include <climits>
#include <cstdlib>

#include "common.cuh"
#include "internal.h"

namespace vga {

static const int CHAIN_WARPS = 4;
static const int RING = 128;           // supports bandwidth <= 96 (checked by the caller)
static const u32 FULLM = 0xffffffffu;

// score_anchor (src/chain.rs:274-368) for forward/forward anchors of equal length k, returned as the
// integer round(x * 1000) the reference divides by 1000.0 (:360-363); INT_MIN stands for -f64::MAX.
// The f64 score is a strictly increasing function of that integer, so every comparison of the
// chaining DP (`>` at :430-447, `==` at :469) can be made on the integers, and only the winning
// score of an anchor is turned back into the f64 the reference stores (one division per anchor).
__device__ __forceinline__ i32 score_pair(u32 qb_a, u32 tb_a, u32 te_a, double f_a, u32 qb_b, u32 tb_b, u32 te_b,
                                          u32 k, u32 max_gap, const double* __restrict__ gapcost) {
    // all four positions must share their orientation (bit 31), :280-283; equal orientations compare like plain positions
    if (qb_a >= qb_b || te_a >= te_b || (((tb_a ^ te_a) | (tb_a ^ tb_b) | (te_a ^ te_b)) >> 31)) return INT_MIN;   // :277-311
    const u32 ql = qb_b - qb_a;                                       // :313 (both terms are equal)
    const u32 tbd = tb_b > tb_a ? tb_b - tb_a : tb_a - tb_b;          // :324-327
    const u32 ted = te_b - te_a;                                      // :328-331
    const u32 tl = min(tbd, ted);
    const u32 g = ql > tl ? ql - tl : tl - ql;                        // :338-341
    if (g > max_gap) return INT_MIN;                                  // :343-344
    const double gc = __ldg(gapcost + g);                             // :348-354
    const u32 ml = min(min(ql, tl), k);                               // :357
    double x = __dadd_rn(f_a, (double)ml);
    x = __dsub_rn(x, gc);
    x = __dmul_rn(x, 1000.0);
    return __double2int_rn(round(x));                                 // f64::round: half away from zero
}

// The same function without branches, for the two-pass body below: an invalid pair (or a lane without a predecessor,
// `live` false) still runs the arithmetic on whatever it loaded — the gap is clamped into the table, F2I saturates — and
// is replaced by INT_MIN at the end, so two evaluations can be interleaved by the scheduler.
__device__ __forceinline__ i32 score_pair_nb(bool live, u32 qb_a, u32 tb_a, u32 te_a, double f_a, u32 qb_b, u32 tb_b, u32 te_b,
                                             u32 k, u32 max_gap, const double* __restrict__ gapcost) {
    live = live && qb_a < qb_b && te_a < te_b && !(((tb_a ^ te_a) | (tb_a ^ tb_b) | (te_a ^ te_b)) >> 31);
    const u32 ql = qb_b - qb_a;
    const u32 tbd = tb_b > tb_a ? tb_b - tb_a : tb_a - tb_b;
    const u32 ted = te_b - te_a;
    const u32 tl = min(tbd, ted);
    const u32 g = ql > tl ? ql - tl : tl - ql;
    live = live && g <= max_gap;
    const double gc = __ldg(gapcost + min(g, max_gap));
    const u32 ml = min(min(ql, tl), k);
    double x = __dadd_rn(f_a, (double)ml);
    x = __dsub_rn(x, gc);
    x = __dmul_rn(x, 1000.0);
    const i32 r = __double2int_rn(round(x));
    return live ? r : INT_MIN;
}

// The all-integer form: with A = 1000 * f_a (the integer the reference's stored score is the quotient of) the f64 evaluation
// above equals A + 1000 * m - K[g], K[g] = round(1000 * gap_cost(g)), for every A — provided no 1000 * gap_cost(g) lies within
// the f64 rounding error of a rounding boundary, which the host checks when it builds K (vga_api.cu upload_gapcost).
__device__ __forceinline__ i32 score_pair_int(bool live, u32 qb_a, u32 tb_a, u32 te_a, i32 r_a, u32 qb_b, u32 tb_b, u32 te_b,
                                              u32 k, u32 max_gap, const i32* __restrict__ sK) {
    live = live && qb_a < qb_b && te_a < te_b && !(((tb_a ^ te_a) | (tb_a ^ tb_b) | (te_a ^ te_b)) >> 31);
    const u32 ql = qb_b - qb_a;
    const u32 tbd = tb_b > tb_a ? tb_b - tb_a : tb_a - tb_b;
    const u32 ted = te_b - te_a;
    const u32 tl = min(tbd, ted);
    const u32 g = ql > tl ? ql - tl : tl - ql;
    live = live && g <= max_gap;
    const i32 kg = sK[min(g, max_gap)];
    const u32 ml = min(min(ql, tl), k);
    const i32 r = r_a + (i32)(ml * 1000u) - kg;
    return live ? r : INT_MIN;
}

// r / 1000.0 correctly rounded, for an integer score r, without the division subroutine on the DP's critical path:
// q0 = RN(r * RN(1/1000)), one FMA for the exact remainder, one FMA for the correction.  Equal to the IEEE quotient for
// EVERY r in [-1e6, 2^31) — checked exhaustively on the host (tests/test_host.py::test_div1000_identity runs a stride of it).
__device__ __forceinline__ double div1000(i32 r) {
    const double a = (double)r;
    const double q0 = __dmul_rn(a, 0.001);
    const double rem = __fma_rn(-1000.0, q0, a);
    return __fma_rn(rem, 0.001, q0);
}

// One warp per read.  FUSED: reads with at most CHAIN_SMAX anchors — the stable sort by
// target_end.position (src/chain.rs:386-389) is a bitonic sort of (te, id) keys in shared memory,
// and the integer scores / predecessor distances of the whole read stay in shared memory for the
// backtracking.  !FUSED: the reads of big_list (more anchors than that), already sorted by CUB.
// INTDP: the all-integer DP (scores as round(x * 1000), gap costs from the integer table in shared memory; one packed
// reduction per anchor when the scores fit 25 bits, which they always do in the fused kernel).
template <bool FUSED, bool INTDP>
__global__ void __launch_bounds__(CHAIN_WARPS * 32) chain_kernel(ChainBuffers b, bool two_pass_on) {
    __shared__ u32 s_q[INTDP ? 1 : CHAIN_WARPS][INTDP ? 1 : RING], s_tb[INTDP ? 1 : CHAIN_WARPS][INTDP ? 1 : RING], s_te[INTDP ? 1 : CHAIN_WARPS][INTDP ? 1 : RING];
    __shared__ uint4 s_ring[INTDP ? CHAIN_WARPS : 1][INTDP ? RING : 1];      // INTDP: {query_begin, target_begin, target_end, score} of a staged anchor: one 16-byte load per predecessor
    __shared__ double s_f[INTDP ? 1 : CHAIN_WARPS][INTDP ? 1 : RING];

    __shared__ unsigned long long s_key[FUSED ? CHAIN_WARPS : 1][FUSED ? CHAIN_SMAX : 1];
    __shared__ u8 s_pd[FUSED ? CHAIN_WARPS : 1][FUSED ? CHAIN_SMAX : 1];
    const u32 lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 n_warps = (gridDim.x * blockDim.x) >> 5;
    u32* rq = s_q[INTDP ? 0 : wl]; u32* rtb = s_tb[INTDP ? 0 : wl]; u32* rte = s_te[INTDP ? 0 : wl]; double* rf = s_f[INTDP ? 0 : wl];
    uint4* ring = s_ring[INTDP ? wl : 0];
    unsigned long long* keys = s_key[FUSED ? wl : 0];
    // integer score of sorted anchor i: kept in the HIGH word of keys[i] (its te) once the anchor has been staged into the
    // ring (a block of 32 is staged before any score of that block is written), the low word (anchor id) stays
    i32* rr2 = (i32*)keys;
    u8* pd = s_pd[FUSED ? wl : 0];
    const i32 rk = (i32)b.k * 1000;                                       // Anchor::new: max_chain_score = k
    const u32 n_work = FUSED ? b.n_reads : b.n_big;
    const bool two_pass = two_pass_on && b.bandwidth > 32 && b.bandwidth <= 64;
    for (u32 x = warp; x < n_work; x += n_warps) {
        const u32 r = FUSED ? x : b.big_list[x];
        const u32 base = b.read_aoff[r];
        const u32 n = b.read_aoff[r + 1] - base;
        if (n == 0) {
            if (lane == 0) { b.n_chains[r] = 0; b.n_members[r] = 0; }
            continue;
        }
        if (FUSED && n > CHAIN_SMAX) continue;                               // handled by the !FUSED pass
        if (FUSED) {
            // ---- stable sort by te: keys (te, id) are unique, so any sort is the stable one
            u32 np2 = 32;
            while (np2 < n) np2 <<= 1;
            for (u32 i = lane; i < np2; i += 32)
                keys[i] = i < n ? (((unsigned long long)(b.a_te[base + i] ^ 0x80000000u) << 32) | i) : ~0ull;     // Reverse first
            __syncwarp();
            for (u32 k2 = 2; k2 <= np2; k2 <<= 1) {
                for (u32 j = k2 >> 1; j > 0; j >>= 1) {
                    for (u32 t = lane; t < (np2 >> 1); t += 32) {
                        const u32 i = ((t & ~(j - 1)) << 1) | (t & (j - 1));   // the lower index of pair t
                        const u32 p = i | j;
                        const unsigned long long a = keys[i], c = keys[p];
                        const bool up = (i & k2) == 0;
                        if ((a > c) == up) { keys[i] = c; keys[p] = a; }
                    }
                    __syncwarp();
                }
            }
            for (u32 i = lane; i < n; i += 32) b.sidx[base + i] = (u32)keys[i];
        }
        i32 r_max = 0;                                                   // curr_max = 0.0 (:398)
        if (INTDP) {
            // All-integer DP, software-pipelined (bandwidth <= 64; the host gates this instance).  Of the up to 64 predecessors of
            // anchor i only i-1 was scored in the previous iteration; its score is warp-uniform and stays in a register.  Everything
            // else anchor i needs (coordinates of its window, gap costs, the older scores from the ring) was loaded and reduced to
            // "valid, T = 1000 m - K[g], predecessor score" during iteration i-1, before that iteration's reduction — so the serial
            // dependency between anchors is one add, one packed REDUX and the select, not a round trip through shared memory.
            auto stage = [&](u32 first) {                                // anchors first .. first+31 of the sorted order into the ring
                const u32 ii = first + lane;
                if (ii < n) {
                    const u32 id = FUSED ? (u32)keys[ii] : b.sidx[base + ii];
                    const u32 te = FUSED ? ((u32)(keys[ii] >> 32) ^ 0x80000000u) : b.a_te[base + id];
                    ring[ii % RING] = make_uint4(b.a_qb[base + id], b.a_tb[base + id], te, 0u);
                }
            };
            stage(0); stage(32);
            __syncwarp();
            const u32 d1 = lane + 1, d2 = lane + 33;
            bool v1 = false, v2 = false;                                 // operands of the anchor about to be scored
            i32 t1 = 0, t2 = 0, ra1 = 0, ra2 = 0;
            i32 far_ub = INT_MIN;                                        // warp-uniform: best predecessor score at distances 33 .. bandwidth
            auto one = [&](u32 ia, bool live, u32 sj, bool& v, i32& t, i32& ra) {
                const uint4 eb = ring[ia % RING];
                const u32 qb_b = eb.x, tb_b = eb.y, te_b = eb.z;
                const uint4 ea = ring[sj];
                const u32 qb_a = ea.x, tb_a = ea.y, te_a = ea.z;
                live = live && qb_a < qb_b && te_a < te_b && !(((tb_a ^ te_a) | (tb_a ^ tb_b) | (te_a ^ te_b)) >> 31);
                const u32 ql = qb_b - qb_a;
                const u32 tbd = tb_b > tb_a ? tb_b - tb_a : tb_a - tb_b;
                const u32 tl = min(tbd, te_b - te_a);
                const u32 g = ql > tl ? ql - tl : tl - ql;
                v = live && g <= b.max_gap;
                t = (i32)(min(min(ql, tl), b.k) * 1000u) - __ldg(b.gapcost_i + min(g, b.max_gap));
                ra = (i32)ea.w;
            };
            // The far half of the window (distances 33 .. bandwidth) is only BOUNDED ahead of time: a pair scores at most its
            // predecessor's score + 1000 k (gap costs are >= 0), and a far predecessor replaces a near one only with a strictly
            // higher score (ties keep the nearest).  In a collinear run of anchors the near half always wins, so the far pairs are
            // evaluated only when the bound says they could matter (chain starts, gaps in the run).
            auto prefetch = [&](u32 ia) {                                // operands of anchor ia (its predecessors ia-2 .. are in the ring)
                one(ia, d1 <= b.bandwidth && d1 <= ia, (ia - d1) % RING, v1, t1, ra1);
                const bool live2 = d2 <= b.bandwidth && d2 <= ia;
                far_ub = __reduce_max_sync(FULLM, live2 ? ((const i32*)&ring[(ia - d2) % RING])[3] : INT_MIN);
            };
            i32 r_prev = rk;
            for (u32 i = 0; i < n; i++) {
                if (lane == 0) ra1 = r_prev;                             // predecessor i-1: just computed
                const i32 a1 = v1 ? ra1 + t1 : INT_MIN;
                i32 m;
                u32 dmin;
                if (FUSED) {
                    // |score| < 2^25 here (at most CHAIN_SMAX anchors of at most 32 bases, gap costs below 2^21): the score and the
                    // nearest-predecessor rule (ties keep the smallest distance) travel as ONE key through one reduction
                    const i32 k1 = v1 ? a1 * 64 + (64 - (i32)d1) : INT_MIN;
                    i32 km = __reduce_max_sync(FULLM, k1);
                    if (far_ub != INT_MIN && (km == INT_MIN || far_ub + rk > (km >> 6))) {          // warp-uniform
                        one(i, d2 <= b.bandwidth && d2 <= i, (i - d2) % RING, v2, t2, ra2);
                        const i32 k2 = v2 ? (ra2 + t2) * 64 + (64 - (i32)d2) : INT_MIN;
                        km = max(km, __reduce_max_sync(FULLM, k2));
                    }
                    m = km == INT_MIN ? INT_MIN : (km >> 6);
                    dmin = 64u - (u32)(km & 63);
                } else {
                    i32 a2 = INT_MIN;
                    const i32 m1 = __reduce_max_sync(FULLM, a1);
                    if (far_ub != INT_MIN && (m1 == INT_MIN || (i64)far_ub + rk > (i64)m1)) {         // warp-uniform
                        one(i, d2 <= b.bandwidth && d2 <= i, (i - d2) % RING, v2, t2, ra2);
                        a2 = v2 ? ra2 + t2 : INT_MIN;
                    }
                    const i32 best = max(a1, a2);
                    m = max(m1, __reduce_max_sync(FULLM, a2));
                    dmin = __reduce_min_sync(FULLM, best == m && best != INT_MIN ? (a1 == m && v1 ? d1 : d2) : 0x7fffffffu);
                }
                // operands of anchor i+1 while the reduction is in flight (its own coordinates were staged a block ahead)
                if (((i + 1) & 31) == 0) { stage(i + 33); __syncwarp(); }
                if (i + 1 < n) prefetch(i + 1);
                i32 r_i = rk;
                i32 p = -1;
                if (m > rk) { r_i = m; p = (i32)(i - dmin); }            // :430-443 strict >
                if (m > r_max) r_max = m;                                // :445-447 (all proposals)
                r_prev = r_i;
                if (lane == 0) {
                    ((i32*)&ring[i % RING])[3] = r_i;
                    if (FUSED) { rr2[2 * i + 1] = r_i; pd[i] = p < 0 ? 0 : (u8)dmin; }       // the f64 scores follow after the loop, 32 at a time
                    else { b.f[base + i] = div1000(r_i); b.pred[base + i] = p; }            // :360-363
                    if (b.pred_dbg) b.pred_dbg[base + i] = p < 0 ? -1 : (i32)(FUSED ? (u32)keys[p] : b.sidx[base + p]);
                }
                __syncwarp();
            }
            if (FUSED)
                for (u32 i = lane; i < n; i += 32) b.f[base + i] = div1000(rr2[2 * i + 1]);     // :360-363
        }
        if (!INTDP)
        for (u32 i = 0; i < n; i++) {
            if ((i & 31) == 0) {           // stage the next 32 anchors of the sorted order
                const u32 ii = i + lane;
                if (ii < n) {
                    const u32 id = FUSED ? (u32)keys[ii] : b.sidx[base + ii];
                    rq[ii % RING] = b.a_qb[base + id];
                    rtb[ii % RING] = b.a_tb[base + id];
                    rte[ii % RING] = FUSED ? ((u32)(keys[ii] >> 32) ^ 0x80000000u) : b.a_te[base + id];
                }
                __syncwarp();
            }
            const u32 si = i % RING;
            const u32 qb_i = rq[si], tb_i = rtb[si], te_i = rte[si];
            i32 best = INT_MIN;
            u32 bestd = 0x7fffffffu;
            i32 m;
            u32 dmin;
            {
            if (two_pass) {
                // bandwidth in (32, 64] (map_main.rs:103 passes 50): both halves of the window as straight-line code
                const u32 d1 = lane + 1, d2 = lane + 33;
                const u32 s1 = (i - d1) % RING, s2 = (i - d2) % RING;
                const i32 a1 = score_pair_nb(d1 <= i, rq[s1], rtb[s1], rte[s1], rf[s1], qb_i, tb_i, te_i, b.k, b.max_gap, b.gapcost);
                const i32 a2 = score_pair_nb(d2 <= b.bandwidth && d2 <= i, rq[s2], rtb[s2], rte[s2], rf[s2], qb_i, tb_i, te_i, b.k,
                                             b.max_gap, b.gapcost);
                if (a1 > best) { best = a1; bestd = d1; }
                if (a2 > best) { best = a2; bestd = d2; }
            } else
            for (u32 d0 = 0; d0 < b.bandwidth; d0 += 32) {               // j = i-1 down to max(0, i-bandwidth) :404-417
                const u32 d = d0 + lane + 1;
                if (d <= b.bandwidth && d <= i) {
                    const u32 sj = (i - d) % RING;
                    const i32 s = score_pair(rq[sj], rtb[sj], rte[sj], rf[sj], qb_i, tb_i, te_i, b.k, b.max_gap, b.gapcost);
                    if (s > best) { best = s; bestd = d; }
                }
            }
            m = __reduce_max_sync(FULLM, best);
            dmin = __reduce_min_sync(FULLM, best == m ? bestd : 0x7fffffffu);   // ties keep the nearest j
            }
            i32 r_i = rk;
            i32 p = -1;
            if (m > rk) { r_i = m; p = (i32)(i - dmin); }                // :430-443 strict >
            if (m > r_max) r_max = m;                                    // :445-447 (all proposals)
            const double f_i = div1000(r_i);                                 // :360-363
            if (lane == 0) {
                rf[si] = f_i;
                b.f[base + i] = f_i;
                if (FUSED) { rr2[2 * i + 1] = r_i; pd[i] = p < 0 ? 0 : (u8)dmin; }
                else b.pred[base + i] = p;
                if (b.pred_dbg) b.pred_dbg[base + i] = p < 0 ? -1 : (i32)(FUSED ? (u32)keys[p] : b.sidx[base + p]);
            }
            __syncwarp();
        }
        // ---- backtracking (:455-558).  The reference visits i = n-1 .. 0 and starts a chain at every anchor that still has
        // a predecessor and carries the maximum score; candidates are found 32 anchors at a time by ballot (a chain consumes
        // only anchors below its start, so the block is re-examined after every extraction), the walk itself is lane 0's.
        {
            u32* creg = b.cmem + 2ull * base + 2ull * r;
            u32* cl = b.clen + base + r;
            u32 n_ch = 0, tail = 0;                                      // lane 0's
            const double curr_max = div1000(r_max);
            for (i32 c0 = (i32)((n - 1) & ~31u); c0 >= 0; c0 -= 32) {
                for (;;) {
                    const u32 ii = (u32)c0 + lane;
                    bool cand = false;
                    if (ii < n) cand = FUSED ? (pd[ii] != 0 && rr2[2 * ii + 1] == r_max)
                                             : (b.pred[base + ii] >= 0 && b.f[base + ii] == curr_max);     // :469
                    const u32 cm = __ballot_sync(FULLM, cand);
                    if (cm == 0) break;
                    if (lane == 0) {
                        const u32 i = (u32)c0 + 31u - (u32)__clz((int)cm);
                        const u32 start = tail;
                        u32 cur = i;
                        for (;;) {
                            i32 p;
                            if (FUSED) { p = pd[cur] ? (i32)(cur - pd[cur]) : -1; pd[cur] = 0; }
                            else { p = b.pred[base + cur]; b.pred[base + cur] = -1; }       // :478 consume
                            creg[tail++] = cur;                              // :498 the terminal is pushed too
                            if (p < 0) break;
                            cur = (u32)p;
                        }
                        const u32 len = tail - start;
                        if (len >= b.min_anchors) {                          // :545-548
                            for (u32 y = 0; y < len / 2; y++) {
                                u32 t = creg[start + y];
                                creg[start + y] = creg[tail - 1 - y];
                                creg[tail - 1 - y] = t;
                            }
                            cl[n_ch++] = len;
                        } else {
                            tail = start;
                        }
                    }
                    __syncwarp();
                }
            }
            if (lane == 0) {
                b.n_chains[r] = n_ch;
                b.n_members[r] = tail;
            }
        }
        __syncwarp();
    }
}

// anchors per read -> u32 offsets, plus the list of reads the fused kernel leaves to the CUB path
__global__ void chain_prep_kernel(const u64* __restrict__ aoff64, u32 n, u32* __restrict__ aoff32, u32* big_ctr,
                                  u32* __restrict__ big_list, u32* __restrict__ big_beg, u32* __restrict__ big_end) {
    const u32 r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r > n) return;
    aoff32[r] = (u32)aoff64[r];
    if (r == n) return;
    const u64 cnt = aoff64[r + 1] - aoff64[r];
    if (cnt > CHAIN_SMAX) {
        const u32 x = atomicAdd(big_ctr, 1u);
        big_list[x] = r; big_beg[x] = (u32)aoff64[r]; big_end[x] = (u32)aoff64[r + 1];
    }
}

__global__ void __launch_bounds__(256) chain_gather_kernel(ChainGather g) {
    const u32 lane = threadIdx.x & 31;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 n_warps = (gridDim.x * blockDim.x) >> 5;
    for (u32 r = warp; r < g.n_reads; r += n_warps) {
        const u32 base = g.read_aoff[r];
        const u32 nc = g.n_chains[r];
        const u64 c0 = g.chain_off[r];
        u64 m0 = g.member_off[r];
        const u32* creg = g.cmem + 2ull * base + 2ull * r;
        const u32* cl = g.clen + base + r;
        u32 local = 0;
        for (u32 c = 0; c < nc; c++) {
            const u32 len = cl[c];
            if (lane == 0) g.chain_anchor_off[c0 + c] = m0 + local;
            for (u32 x = lane; x < len; x += 32) {
                const u32 p = creg[local + x];
                const u32 id = g.sidx[base + p];
                const u64 o = m0 + local + x;
                g.ca_id[o] = id;
                g.ca_qb[o] = g.a_qb[base + id];
                g.ca_tb[o] = g.a_tb[base + id];
                g.ca_te[o] = g.a_te[base + id];
                g.ca_score[o] = g.f[base + p];
            }
            local += len;
        }
        if (r == g.n_reads - 1 && lane == 0) g.chain_anchor_off[g.chain_off[g.n_reads]] = g.member_off[g.n_reads];
    }
}

static int grid_for_warps(u32 n_items, int warps_per_cta) {
    long long ctas = ((long long)n_items + warps_per_cta - 1) / warps_per_cta;
    const long long cap = 148 * 16;
    if (ctas > cap) ctas = cap;
    if (ctas < 1) ctas = 1;
    return (int)ctas;
}

int launch_chain_prep(const u64* aoff64, u32 n, u32* aoff32, u32* big_ctr, u32* big_list, u32* big_beg, u32* big_end,
                      cudaStream_t st) {
    chain_prep_kernel<<<(n + 1 + 255) / 256, 256, 0, st>>>(aoff64, n, aoff32, big_ctr, big_list, big_beg, big_end);
    return 1;
}

int launch_chain(const ChainBuffers& b, bool fused, cudaStream_t st) {
    // the straight-line two-pass window body for bandwidths in (32, 64]: C2 chain 4.66 -> 4.08 ms, C4 35.0 -> 31.3 ms, parity
    // green (gpurun_out/ab16); VGA_CHAIN_TWO_PASS=0 restores the generic loop (A/B)
    const char* e = getenv("VGA_CHAIN_TWO_PASS");
    const bool tp = e ? atoi(e) != 0 : true;
    // all-integer DP when the host proved it exact for this (k, max_gap) (VGA_CHAIN_INT=0: the f64 body, A/B and tests)
    const char* ei = getenv("VGA_CHAIN_INT");
    const bool idp = b.gapcost_i != nullptr && b.bandwidth <= 64 && (ei ? atoi(ei) != 0 : true);
    if (fused) {
        if (idp) chain_kernel<true, true><<<grid_for_warps(b.n_reads, CHAIN_WARPS), CHAIN_WARPS * 32, 0, st>>>(b, tp);
        else chain_kernel<true, false><<<grid_for_warps(b.n_reads, CHAIN_WARPS), CHAIN_WARPS * 32, 0, st>>>(b, tp);
    } else if (b.n_big) {
        if (idp) chain_kernel<false, true><<<grid_for_warps(b.n_big, CHAIN_WARPS), CHAIN_WARPS * 32, 0, st>>>(b, tp);
        else chain_kernel<false, false><<<grid_for_warps(b.n_big, CHAIN_WARPS), CHAIN_WARPS * 32, 0, st>>>(b, tp);
    }
    else return 0;
    return 1;
}

int launch_chain_gather(const ChainGather& g, cudaStream_t st) {
    chain_gather_kernel<<<grid_for_warps(g.n_reads, 8), 256, 0, st>>>(g);
    return 1;
}

}  // namespace vga
