// K5 (short reads, 16x2 arithmetic) — banded sequence-to-graph POA fill for reads shorter than 246 bases, replacing
// the abPOA call at the reference's FFI seam (src/align.rs:190-202).  Same alignment definition, trace word
// format, RowInfo / NodeState records and work-queue protocol as k_poa.cu / k_poa_short.cu; the
// traceback kernel is shared.  This instance is compiled for abPOA's default gap model (o1=4, e1=2, o2=24, e2=1 —
// the only one the reference ever uses, src/align.rs:190-202), so every gap constant is an immediate; other gap
// models run on the 32-bit kernels.
//
// What is different from k_poa_short.cu: the DP arithmetic runs two query columns per 32-bit register on the
// sm_100a 16x2 integer instructions (VIADD.16x2, VIMNMX3.S16x2, VIADDMNMX.S16x2[.RELU]); absolute scores fit
// int16 in this class (same criterion as before).
//   * Columns are grouped in aligned BLOCKS of 8 (block b = columns 8b-8 .. 8b-1, so column -1 exists).  Inside a
//     block, 32-bit word t (t = 0..3) holds column t in its low half and column t+4 in its high half.  A block is
//     thus two independent 4-column chains (A = columns 0..3 in the low halves, B = columns 4..7 in the high
//     halves): the left neighbour of pair t is pair t-1 for BOTH halves, so the diagonal operand and the serial
//     part of the F1/F2 prefix-max scans need no per-pair shuffling.  Chains are stitched per lane in scalar code
//     (2 chains per block), lanes by Kogge-Stone over the group.
//   * The previous row (H, E1', E2') lives in shared memory in exactly this layout: one LDS.128 / STS.128 per plane
//     and block.  The substitution scores are NOT stored per base: the read is kept as one plane of 16-bit PRMT
//     selectors per column pair (nibbles: code of column t, code | 8, code of column t+4, code | 8; code = A0 C1 T2 G3,
//     4 = anything else), and a row turns a selector into the pair of scores with ONE byte permute of an 8-byte
//     look-up register pair (byte c = score of the row's base against code c; a selector nibble with bit 3 set
//     replicates the sign, which makes the int8 score an int16).  8 bytes of shared memory per block instead of 80:
//     1.3 KB per alignment instead of 2.7, so twice the groups fit an SM.
//   * A GROUP of G lanes owns one (read, chain).  A row is computed in SPANS of G*NB blocks (lane l: NB consecutive
//     blocks); a warp step advances every group of the warp by one span, so a group with a narrow band moves on to
//     its next row while a group with a wide band works on the next span of the same row — nobody computes columns
//     it does not need.  F, the old H of the boundary column and 1-H of the boundary column are carried between
//     spans in registers.  All 32 lanes execute every step (groups without work run it with stores disabled), so
//     the shuffles of the step use the full mask.
//   * Band limits inside the first / last block are applied with validity words from a small shared table
//     (masked cells get Hhat = -inf exactly); trace rows are stored as whole blocks with 16-byte vectors, which is
//     why word 0 of a trace row is column beg & ~7 (trace_col0, common.cuh).
//   * Row maximum: 16x2 running max, one group reduction, then VIMNMX.S16x2 predicate outputs mark the columns
//     equal to it; leftmost / rightmost column from the lane's bit mask and one packed group reduction.
// Roofline: INT32 ALU issue.  Algorithmic work = band cells.
#include <climits>
#include <cstdlib>

#include "common.cuh"
#include "internal.h"

namespace vga {

static const int PP_WARPS = 4;                 // warps per CTA
static const u64 PAIR_POOL_CHUNK = 4096;       // trace words a group takes from the pool per refill (multiple of 8)
static const i32 NEG16 = -32000;               // -inf of the int16 planes; reachable scores are >= -30000 in this class
static const u32 NEGP = 0x83008300u;           // NEG16 in both halves
static const i32 PAIR_SCORE_LIMIT = 30000;
static const i32 PAIR_REBASE_AT = 8000;        // CHK instances: a row maximum this far below ref re-centres the planes
static const i32 PAIR_UNREACH = 31000;         // values below -this are -inf (NEG16 plus what an unreachable cell can gain)
static const int VT_WORDS = 8 * 9;             // validity table entries (uint4 each): [a = 0..7][z = 0..8]
// abPOA's default convex gap model (rs-abPOA defaults)
static const i32 PO1 = 4, PE1 = 2, PO2 = 24, PE2 = 1, POE1 = PO1 + PE1, POE2 = PO2 + PE2;
#define P16(v) ((u32)(((u32)(v) & 0xffffu) * 0x10001u))           // v in both halves (compile-time)
#define P16LH(lo, hi) ((u32)(((u32)(lo) & 0xffffu) | (((u32)(hi) & 0xffffu) << 16)))

__device__ __forceinline__ i32 imax(i32 a, i32 b) { return a > b ? a : b; }
__device__ __forceinline__ i32 imin(i32 a, i32 b) { return a < b ? a : b; }
__device__ __forceinline__ i32 sext16(u32 w) { return (i32)(short)(w & 0xffffu); }
__device__ __forceinline__ u32 pair16(i32 lo, i32 hi) { return __byte_perm((u32)lo, (u32)hi, 0x5410); }
__device__ __forceinline__ u32 both16(i32 v) { return __byte_perm((u32)v, (u32)v, 0x1010); }
// generic byte permute (selector nibbles with bit 3 set replicate the sign of the selected byte; only bits 0..15 of c are used)
__device__ __forceinline__ u32 prmt(u32 a, u32 b, u32 c) {
    u32 d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// selector code of a read byte: A 0, C 1, T 2, G 3 (bits 1..2 of the ASCII code), anything else 4
__device__ __forceinline__ u32 base_code(u32 c) {
    const u32 cc = (c >> 1) & 3u;
    return c == ((0x47544341u >> (8 * cc)) & 0xffu) ? cc : 4u;
}
__device__ __forceinline__ u32 sel16_of(u32 clo, u32 chi) { return clo | ((clo | 8u) << 4) | (chi << 8) | ((chi | 8u) << 12); }
// half-word index of query column j (j >= -8) inside a plane
__device__ __forceinline__ int hw_index(i32 j) {
    const i32 c = j + 8;
    return (int)(((c >> 3) << 3) | ((c & 3) << 1) | ((c >> 2) & 1));
}

struct SpanOut { i32 mx, lpos, rpos, mn; };   // mn: smallest Hhat of a band cell (CHK instances only)
// carried from one span of a row to the next (group-uniform)
struct SpanCarry {
    i32 fc1, fc2;      // F1', F2' arriving at the first column of the span (F' = F + o + e)
    u32 hq;            // low half: old H (previous row) of the column left of the span; high half: 1 - new H of that column
};

template <int G> __device__ __forceinline__ u32 pair_group_mask(u32 lane) {
    return G == 32 ? 0xffffffffu : (((1u << (G & 31)) - 1u) << (lane & ~(u32)(G - 1)));
}

// One span (G*NB blocks) of one DP row, executed by all 32 lanes of the warp (one group each G lanes).
// planes = H, E1', E2' (nblk blocks each) of the group; prof = profile plane of the row's graph base; vtab = validity
// table; tr = trace word of column beg & ~7.  live = false: a group without work (no stores).
// CHK: also return the smallest in-band Hhat of the span (Hhat <= H, so it bounds H from below) — for alignments whose
// row count does not prove a priori that every reachable score fits the int16 planes.
template <int G, int NB, bool CHK>
__device__ __forceinline__ SpanOut span_step(u32 gl, i32 beg, i32 end, i32 bdone, bool live, bool last, SpanCarry& cy,
                                             uint4* __restrict__ planes, int nblk, const uint2* __restrict__ sel, u32 lutA, u32 lutB,
                                             const uint4* __restrict__ vtab, u32* __restrict__ tr) {
    const u32 FULLM = 0xffffffffu;
    const i32 b0 = (beg + 8) >> 3, bend = (end + 8) >> 3;
    const i32 bsp = b0 + bdone;                         // first block of the span (bdone blocks of the row are computed)
    const i32 bme = bsp + (i32)gl * NB;                 // the lane's first block
    const i32 cfirst = bme * 8 - 8;                     // and its first column
    const i32 D = beg - cfirst, Eo = end - cfirst + 1;
    uint4* const Hpl = planes;
    uint4* const E1pl = planes + nblk;
    uint4* const E2pl = planes + 2 * nblk;
    u32 hh[NB][4], fl1[NB][4], fl2[NB][4], x1[NB][4], x2[NB][4];
    u32 R1[NB], R2[NB];
    u32 mnacc = 0x7fff7fffu;
    // ---- sweep 1: Hhat = max(M, E1, E2) masked to the band, chain-local part of the F scans (F' = F + o + e)
    u32 hprev = ((const u32*)(Hpl + (imin(bme, nblk - 1) - 1)))[3];
    if (gl == 0 && bdone > 0) hprev = cy.hq << 16;         // the previous span has already overwritten that block
#pragma unroll
    for (int k = 0; k < NB; k++) {
        const int bk = imin(bme + k, nblk - 1);         // lanes right of the band read the spare block (all their cells are masked)
        const uint4 hv = Hpl[bk], e1v = E1pl[bk], e2v = E2pl[bk];
        const uint2 sv = sel[bk];
        const i32 a = __viaddmax_s32(D, -8 * k, 0);                    // invalid cells at the block's left  (0..7)
        const i32 z = imin(__viaddmax_s32(Eo, -8 * k, 0), 8);          // cells r < z are left of end+1      (0..8)
        const uint4 V = vtab[a * 9 + z];
        const u32 hw[4] = {hv.x, hv.y, hv.z, hv.w};
        const u32 pw[4] = {prmt(lutA, lutB, sv.x), prmt(lutA, lutB, sv.x >> 16), prmt(lutA, lutB, sv.y), prmt(lutA, lutB, sv.y >> 16)};
        const u32 vw[4] = {V.x, V.y, V.z, V.w};
        x1[k][0] = e1v.x; x1[k][1] = e1v.y; x1[k][2] = e1v.z; x1[k][3] = e1v.w;
        x2[k][0] = e2v.x; x2[k][1] = e2v.y; x2[k][2] = e2v.z; x2[k][3] = e2v.w;
        u32 hd = __byte_perm(hprev, hv.w, 0x5432);      // (column 8k-1, column 8k+3): diagonal operands of pair 0
        u32 r1 = NEGP, r2 = NEGP;
#pragma unroll
        for (int t = 0; t < 4; t++) {
            const u32 m = __vadd2(hd, pw[t]);
            u32 h = __vimax3_s16x2(m, x1[k][t], x2[k][t]);
            if (CHK) mnacc = __vmins2(mnacc, (h & vw[t]) | (0x7fff7fffu & ~vw[t]));
            h = (h & vw[t]) | (NEGP & ~vw[t]);
            hh[k][t] = h;
            fl1[k][t] = r1; fl2[k][t] = r2;
            r1 = __viaddmax_s16x2(r1, P16(-PE1), h);
            r2 = __viaddmax_s16x2(r2, P16(-PE2), h);
            hd = hw[t];
        }
        hprev = hv.w;
        R1[k] = r1; R2[k] = r2;
    }
    const u32 hold_last = hprev;            // old H of the lane's last block, word 3 (its high half is the column left of the next lane)
    // ---- stitch the chains of the lane (A0, B0, A1, B1, ...), then the lanes of the group
    u32 in1[NB], in2[NB];
    i32 c1 = NEG16, c2 = NEG16;
#pragma unroll
    for (int k = 0; k < NB; k++) {
        const i32 a1 = sext16(R1[k]), b1 = (i32)R1[k] >> 16, a2 = sext16(R2[k]), b2 = (i32)R2[k] >> 16;
        const i32 ia1 = c1, ia2 = c2;
        c1 = __viaddmax_s32(c1, -4 * PE1, a1); c2 = __viaddmax_s32(c2, -4 * PE2, a2);
        in1[k] = pair16(ia1, c1); in2[k] = pair16(ia2, c2);
        c1 = __viaddmax_s32(c1, -4 * PE1, b1); c2 = __viaddmax_s32(c2, -4 * PE2, b2);
    }
    // what arrives from the previous spans passes through lane 0's columns
    if (gl == 0) { c1 = __viaddmax_s32(cy.fc1, -PE1 * 8 * NB, c1); c2 = __viaddmax_s32(cy.fc2, -PE2 * 8 * NB, c2); }
#pragma unroll
    for (int d = 1; d < G; d <<= 1) {       // shfl_up leaves lanes < d unchanged, and max(v, v - c) = v
        const i32 t1 = __shfl_up_sync(FULLM, c1, d, G), t2 = __shfl_up_sync(FULLM, c2, d, G);
        c1 = __viaddmax_s32(t1, -PE1 * 8 * NB * d, c1);
        c2 = __viaddmax_s32(t2, -PE2 * 8 * NB * d, c2);
    }
    const i32 tot1 = c1, tot2 = c2;         // inclusive: F' arriving right of the lane's last column
    c1 = __shfl_up_sync(FULLM, c1, 1, G); c2 = __shfl_up_sync(FULLM, c2, 1, G);
    if (gl == 0) { c1 = cy.fc1; c2 = cy.fc2; }
    {
        const u32 c1p = both16(c1), c2p = both16(c2);
#pragma unroll
        for (int k = 0; k < NB; k++) {
            in1[k] = __viaddmax_s16x2(c1p, P16LH(-PE1 * 8 * k, -PE1 * (8 * k + 4)), in1[k]);
            in2[k] = __viaddmax_s16x2(c2p, P16LH(-PE2 * 8 * k, -PE2 * (8 * k + 4)), in2[k]);
        }
    }
    // ---- sweep 2a: H and the running maximum
    u32 H[NB][4], f1[NB][4], f2[NB][4];
    u32 mxacc = NEGP;
#pragma unroll
    for (int k = 0; k < NB; k++) {
#pragma unroll
        for (int t = 0; t < 4; t++) {
            f1[k][t] = __viaddmax_s16x2(in1[k], P16(-PE1 * t), fl1[k][t]);
            f2[k][t] = __viaddmax_s16x2(in2[k], P16(-PE2 * t), fl2[k][t]);
            const u32 hp = __viaddmax_s16x2(f1[k][t], P16(-POE1), hh[k][t]);
            H[k][t] = __viaddmax_s16x2(f2[k][t], P16(-POE2), hp);
            mxacc = __vmaxs2(mxacc, H[k][t]);
        }
    }
    // 1 - H of the column left of the lane's first one (F fields of pair 0 of block 0); the first column of a row has none (don't care)
    const u32 qlast = __vadd2(~H[NB - 1][3], 0x00020002u);
    u32 qleft = __shfl_up_sync(FULLM, qlast, 1, G);
    if (gl == 0) qleft = cy.hq;
    // carries for the next span: everything comes from the last lane of the group
    cy.fc1 = __shfl_sync(FULLM, tot1, G - 1, G);
    cy.fc2 = __shfl_sync(FULLM, tot2, G - 1, G);
    cy.hq = __shfl_sync(FULLM, __byte_perm(hold_last, qlast, 0x7632), G - 1, G);
    // ---- sweep 2b: E', trace words, stores
#pragma unroll
    for (int k = 0; k < NB; k++) {
        u32 q[4], wlo[4], whi[4], e1n[4], e2n[4];
#pragma unroll
        for (int t = 0; t < 4; t++) q[t] = __vadd2(~H[k][t], 0x00020002u);         // 1 - H
        const u32 ql0 = __byte_perm(qleft, q[3], 0x5432);                          // (column 8k-1, column 8k+3)
        qleft = q[3];
#pragma unroll
        for (int t = 0; t < 4; t++) {
            const u32 nq = ~H[k][t];
            const u32 u1 = __vadd2(nq, P16(PO1 + 1)), u2 = __vadd2(nq, P16(PO2 + 1));   // o - H
            const u32 d1 = __viaddmin_s16x2_relu(x1[k][t], u1, P16(PO1));           // E1' - (H - oe1) in [0, o1]
            const u32 d2 = __viaddmin_s16x2_relu(x2[k][t], u2, P16(PO2));
            const u32 ql = t == 0 ? ql0 : q[t - 1];
            const u32 g1 = __viaddmin_s16x2_relu(f1[k][t], ql, P16(PO1 + 1));       // max(F1 - (Hleft - oe1), -1) + 1
            const u32 g2 = __viaddmin_s16x2_relu(f2[k][t], ql, P16(PO2 + 1));
            const u32 hi = d1 + d2 * 8u + g1 * 256u + g2 * 2048u;                   // fields are bounded: no carry between halves
            wlo[t] = __byte_perm(H[k][t], hi, 0x5410);
            whi[t] = __byte_perm(H[k][t], hi, 0x7632);
            e1n[t] = __viaddmax_s16x2(x1[k][t], P16(-PE1), __vadd2(H[k][t], P16(-POE1)));
            e2n[t] = __viaddmax_s16x2(x2[k][t], P16(-PE2), __vadd2(H[k][t], P16(-POE2)));
        }
        if (live && bme + k <= bend) {
            uint4* const t4 = (uint4*)(tr + 8 * (bme + k - b0));
            t4[0] = make_uint4(wlo[0], wlo[1], wlo[2], wlo[3]);
            t4[1] = make_uint4(whi[0], whi[1], whi[2], whi[3]);
            Hpl[bme + k] = make_uint4(H[k][0], H[k][1], H[k][2], H[k][3]);
            E1pl[bme + k] = make_uint4(e1n[0], e1n[1], e1n[2], e1n[3]);
            E2pl[bme + k] = make_uint4(e2n[0], e2n[1], e2n[2], e2n[3]);
        }
    }
    // guards: H(beg-1) = -inf (diagonal of column beg in the next row) and E'(end+1) = -inf (vertical of column end+1).
    // A guard inside a block stored this row is written by the block's owner after its stores; other blocks are not stored.
    if (live && bdone == 0 && gl == 0 && (beg & 7) == 0) ((short*)Hpl)[hw_index(beg - 1)] = (short)NEG16;
    if (live && last) {
        const i32 ow = imin((((end + 9) >> 3) - bsp) / NB, G - 1);
        if ((i32)gl == ow) {
            const int x = hw_index(end + 1);
            ((short*)E1pl)[x] = (short)NEG16;
            ((short*)E2pl)[x] = (short)NEG16;
        }
    }
    // ---- maximum of the span and its leftmost / rightmost column (abPOA simd_abpoa_ada_max_i).  Masked cells are below
    // every in-band value of their block: left of beg they are -inf, right of end they hold an F that decays from H(end).
    // Every lane marks the columns that attain ITS maximum; (max, leftmost) and (max, rightmost) are then reduced as two
    // independent keys in the same three shuffle rounds (a lane whose maximum is below the span's loses both).  Columns are
    // biased by 8 (a lane left of the band starts at column -8); they stay below 512 in this class.
    const i32 mxl = imax(sext16(mxacc), (i32)mxacc >> 16);
    const u32 mxp = both16(mxl);
    u32 eq = 0;
#pragma unroll
    for (int k = 0; k < NB; k++) {
#pragma unroll
        for (int t = 0; t < 4; t++) {
            bool ph, pl;
            (void)__vibmax_s16x2(H[k][t], mxp, &ph, &pl);          // H >= lane max  <=>  H == lane max
            if (pl) eq |= 1u << (8 * k + t);
            if (ph) eq |= 1u << (8 * k + t + 4);
        }
    }
    const u32 mxb = (u32)(mxl + 32768) << 9;
    u32 kL = mxb | (u32)(511 - (cfirst + 8 + __ffs((int)eq) - 1));
    u32 kR = mxb | (u32)(cfirst + 8 + 31 - __clz((int)eq));
#pragma unroll
    for (int d = G / 2; d > 0; d >>= 1) {
        const u32 oL = __shfl_xor_sync(FULLM, kL, d, G), oR = __shfl_xor_sync(FULLM, kR, d, G);
        kL = max(kL, oL); kR = max(kR, oR);
    }
    const i32 mx = (i32)(kL >> 9) - 32768;
    SpanOut o;
    o.mn = 0;
    if (CHK) {
        i32 mn = imin(sext16(mnacc), (i32)mnacc >> 16);
#pragma unroll
        for (int d = G / 2; d > 0; d >>= 1) mn = imin(mn, __shfl_xor_sync(FULLM, mn, d, G));
        o.mn = mn;
    }
    o.mx = mx;
    o.lpos = 511 - 8 - (i32)(kL & 511u);
    o.rpos = (i32)(kR & 511u) - 8;
    return o;
}

enum { PH_FETCH = 0, PH_ROW = 1, PH_SPAN = 2, PH_DONE = 3 };

// NBMAX: a step computes up to G*NBMAX blocks of a row; the width of a step (1..NBMAX blocks per lane) is the largest
// any group of the warp still needs, so a warp of narrow rows runs the short body.
template <int G, int NBMAX, int MINB, bool CHK>
__global__ void __launch_bounds__(PP_WARPS * 32, MINB) poa_fill_pair_kernel(DevIndex ix, PoaBuffers b, int nblk) {
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    const u32 FULLM = 0xffffffffu;
    uint4* const vtab = (uint4*)dyn_smem;
    uint4* const planes = vtab + VT_WORDS + (size_t)(threadIdx.x / G) * 4 * nblk;     // H, E1', E2' (nblk uint4 each), read selectors and custom-base selectors (nblk uint2 each)
    uint2* const selq = (uint2*)(planes + 3 * nblk);
    uint2* const selc = selq + nblk;
    short* const Hs = (short*)planes;
    short* const E1s = (short*)(planes + nblk);
    short* const E2s = (short*)(planes + 2 * nblk);
    const u32 lane = threadIdx.x & 31;
    const u32 gl = lane & (G - 1);
    const u32 gm = pair_group_mask<G>(lane);
    // validity table: entry (a, z), word t: low half = column t, high half = column t+4; valid iff a <= column < z
    for (int i = threadIdx.x; i < VT_WORDS; i += blockDim.x) {
        const int a = i / 9, z = i % 9;
        u32 w[4];
        for (int t = 0; t < 4; t++) w[t] = ((a <= t && t < z) ? 0xffffu : 0u) | ((a <= t + 4 && t + 4 < z) ? 0xffff0000u : 0u);
        vtab[i] = make_uint4(w[0], w[1], w[2], w[3]);
    }
    __syncthreads();
    const i32 smatch = b.match, smis = -b.mismatch;
    const u32 mis4 = ((u32)smis & 0xffu) * 0x01010101u;                 // look-up bytes: every code mismatches ...
    const u32 lutB = ((u32)smis & 0xffu) | (((u32)smatch & 0xffu) << 8);   // ... code 4 = other, code 5 = equal to a custom base
    if (gl == 0) Hs[hw_index(-1)] = (short)NEG16;       // column -1: M(i,0) does not exist
    const int ncols = (nblk - 1) * 8;                   // columns -8 .. ncols-9 are backed by real blocks
    // ---- group-uniform state of the alignment in flight
    int phase = PH_FETCH;
    u32 w = 0, nb = 0, S = 0, pred_base = 0, row_base = 0, nrows = 0;
    i32 qlen = 0, wband = 0, end0 = 0;
    const u8* __restrict__ q = nullptr;
    u32 row = 0, s = 0, o = 0, meta = 0, meta_next = 0;
    i32 beg = 0, end = -1, ppb = 0, ppe = -1, lpos = 0, rpos = 0, mxrow = 0;
    i32 ref = 0;                                    // CHK instances: the planes and trace words of a row hold score - ref (RowInfo.ref)
    i32 bdone = 0, nblocks_row = 1, rl = 0, rr = 0;
    u64 cells = 0;
    u32 status = VGA_ALN_OK, custom_base = 0xffffffffu;
    u64 pool_cur = 0, pool_end = 0;                 // this group's current chunk of the trace pool (words)
    u32* tr = b.pool;
    const uint2* pr = selq;
    u32 lutA = mis4;
    SpanCarry cy;
    cy.fc1 = NEG16; cy.fc2 = NEG16; cy.hq = 0;

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
                i32 mx = h;
#pragma unroll
                for (int d = G / 2; d > 0; d >>= 1) mx = imax(mx, __shfl_xor_sync(gm, mx, d, G));
                if (mx > best) {                             // strict: the first node wins ties
                    best = mx;
                    i32 mn = h == mx ? (i32)sx : INT_MAX;
#pragma unroll
                    for (int d = G / 2; d > 0; d >>= 1) mn = imin(mn, __shfl_xor_sync(gm, mn, d, G));
                    bs = (u32)mn;
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
        phase = PH_FETCH;
    };

    for (;;) {
        // ---- A: next (read, chain) of the work list (groups without an item)
        if (phase == PH_FETCH) {
            u32 wi = 0;
            if (gl == 0) wi = atomicAdd(b.work_ctr, 1u);
            wi = __shfl_sync(gm, wi, 0, G);
            if (wi >= b.n_items) phase = PH_DONE;
            else {
                w = b.work[wi];
                const AlignItem it = b.items[w];
                qlen = (i32)it.qlen;
                q = (const u8*)b.seqs + it.q_off;
                wband = b.wb + (i32)(b.wf * (float)qlen);
                nb = it.node_base; S = it.S; row_base = it.row_base; pred_base = it.pred_base; nrows = it.nrows;
                status = VGA_ALN_OK;
                cells = 0; row = 0; s = 0; o = 0; ppb = 0; ppe = -1; lpos = 0; rpos = 0; mxrow = 0; ref = 0;
                custom_base = 0xffffffffu;
                phase = PH_ROW;
                if (S == 0 || nrows == 0) { status = it.status == VGA_ALN_UNSUPPORTED ? VGA_ALN_UNSUPPORTED : VGA_ALN_EMPTY; finish(); }   // K4 refused the chain (reverse-strand anchors)
                // absolute H is kept as int16 (trace words and shared memory): every DP step loses at most
                // max(mismatch, o1+e1), so reachable scores are >= -that*(rows+qlen); other alignments go to the general kernel
                else if (qlen + 2 > ncols - 8 || (!CHK && (i64)imax(b.mismatch, POE1) * ((i64)nrows + qlen) > PAIR_SCORE_LIMIT)) {
                    status = IT_RETRY_WIDER;
                    finish();
                } else {
                    {   // band end of the virtual source row: max_remain of the first row + 1
                        meta_next = b.row_meta[row_base];
                        const i32 remain_src = (i32)(meta_next >> 11) + 1;
                        end0 = band_round_end0(imin(qlen, imax(0, qlen - remain_src) + wband), b.bgran_mask, qlen);
                    }
                    // selectors of the read: pair p = (block p / 4, word p % 4) holds columns 8*blk - 8 + word and + 4 (query base j-1)
                    __syncwarp(gm);
                    for (int p = (int)gl; p < 4 * nblk; p += G) {
                        const i32 j = 8 * (p >> 2) - 8 + (p & 3);
                        const u32 clo = (j >= 1 && j <= qlen) ? base_code(q[j - 1]) : 4u;
                        const u32 chi = (j + 4 >= 1 && j + 4 <= qlen) ? base_code(q[j + 3]) : 4u;
                        ((u16*)selq)[p] = (u16)sel16_of(clo, chi);
                    }
                }
            }
        }
        // ---- B: row prologue (groups that start a DP row).  The row record (sub_rowmeta_kernel) comes by a cached load that every
        // lane of the group issues for the same address, one row ahead; the lines it will need are pulled into L1 a few dozen
        // rows ahead by prefetches, which have no destination register.  (Round 1 loaded G rows at a time and shuffled; the
        // shuffle then waited on the warp-wide scoreboard for whatever load ANOTHER group of the warp had just issued: 10 % of
        // all stall samples.  A load consumed in the same prologue waits just as long: profiles/r02_pair_stalls.txt, L426.)
        if (phase == PH_ROW) {
            if ((row & 31u) == 0 && gl == 0) {
                const u32* nx = b.row_meta + row_base + row + 32;
                if (row == 0) asm volatile("prefetch.global.L1 [%0];" ::"l"(nx - 32 + 31));     // rest of the first 128 bytes
                if (row + 32 < nrows) asm volatile("prefetch.global.L1 [%0];" ::"l"(nx));
                if (row + 64 < nrows) asm volatile("prefetch.global.L1 [%0];" ::"l"(nx + 32));
            }
            // one row ahead: the record of row + 1 is requested now and consumed by the next prologue of this group, at least one
            // step body later (an L1 / L2 hit by then: 32 records per 128-byte line)
            meta = meta_next;
            if (row + 1 < nrows) meta_next = __ldca(b.row_meta + row_base + row + 1);
        }
        if (phase == PH_ROW) {
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
                i32 mxp = INT_MIN;
                for (u32 x = 0; x < np; x++) {
                    const NodeState ns = b.nd_state[nb + pl[x]];
                    mpl = imin(mpl, ns.lpos + 1); mpr = imax(mpr, ns.rpos + 1);
                    min_pb = imin(min_pb, ns.beg); max_pe = imax(max_pe, ns.end);
                    mxp = imax(mxp, ns.mx);
                }
                prow = b.nd_row_start[nb + pl[0] + 1] - 1;
                if (CHK) ref = imin(mxp, 0);            // the merged row is rebuilt below, relative to the best predecessor maximum
            }
            if (CHK && kind == 1) ref = 0;
            if (CHK && kind == 0 && mxrow < -PAIR_REBASE_AT) {
                // Scores have drifted down (a subgraph thousands of rows longer than its read is mostly deletions): re-centre the
                // previous row on its maximum.  Cells already at -inf stay there; nothing else of this alignment is relative to ref.
                const i32 delta = -mxrow;
                __syncwarp(gm);
                for (int x = (int)gl; x < 24 * nblk; x += G) {
                    const i32 v = ((short*)planes)[x];
                    ((short*)planes)[x] = (short)(v < -PAIR_UNREACH ? NEG16 : v + delta);
                }
                __syncwarp(gm);
                ref -= delta; mxrow = 0;
            }
            // adaptive band (abPOA GET_AD_DP_BEGIN/END + clipping to the predecessors' bands)
            const i32 r = qlen - (i32)(meta >> 11);
            beg = imax(imax(0, imin(mpl, r) - wband), min_pb);
            end = imin(imin(qlen, imax(mpr, r) + wband), max_pe + 1);
            band_round(beg, end, b.bgran_mask, qlen, max_pe);
            const i32 W = end - beg + 1;
            const i32 nblocks = ((end + 8) >> 3) - ((beg + 8) >> 3) + 1;
            if (beg > end) status = VGA_ALN_BAND_ERR;
            // pool space: whole blocks, one trace word per cell
            const u64 need = (u64)nblocks * 8;
            if (status == VGA_ALN_OK && pool_cur + need > pool_end) {
                unsigned long long got = 0;
                const u64 take = need > PAIR_POOL_CHUNK ? need : PAIR_POOL_CHUNK;
                if (gl == 0) got = atomicAdd(b.pool_ctr, (unsigned long long)take);
                got = __shfl_sync(gm, got, 0, G);
                if (got + take > b.pool_words) status = IT_RETRY_POOL;
                else { pool_cur = got; pool_end = got + take; }
            }
            if (status != VGA_ALN_OK) finish();
            else {
                tr = b.pool + pool_cur;
                if (gl == 0) {
                    uint4* dst = (uint4*)(b.rows + row_base + row);
                    dst[0] = make_uint4((u32)beg, (u32)end, (u32)(pool_cur >> 2), (u32)ref);
                    dst[1] = make_uint4(prow, (np < 0xffffffu ? np : 0xffffffu) | (base << 24), s, o);
                }
                pool_cur += need;
                cells += (u64)W;
                if (kind != 0) {
                    // the predecessor row is not the one in shared memory: build it there (columns beg-1 .. end+1)
                    const i32 jlo = imax(beg - 1, 0);
                    if (kind == 1) {          // virtual source row
                        __syncwarp(gm);
                        for (i32 j = jlo + (i32)gl; j <= end + 1; j += G) {
                            const bool in0 = j <= end0;
                            const int x = hw_index(j);
                            Hs[x] = (short)(!in0 ? NEG16 : j == 0 ? 0 : imax(-(PO1 + PE1 * j), -(PO2 + PE2 * j)));
                            E1s[x] = (short)(j == 0 ? -POE1 : NEG16);
                            E2s[x] = (short)(j == 0 ? -POE2 : NEG16);
                        }
                    } else {                  // merge of the predecessors' last rows, decoded from their trace words
                        bool lost = false;
                        const u32* __restrict__ pl = b.nd_pred + pred_base + po;
                        for (u32 x = 0; x < np; x++) {
                            // one predecessor at a time: its band and trace row are read once, its words stream through the lanes
                            const u32 p = pl[x];
                            const int2 pbe = *(const int2*)(b.nd_state + nb + p);
                            // this kernel keeps the trace offset of a node's last row in NodeState.pad0 (one dependent load less)
                            const u32* __restrict__ tw = b.pool + (u64)(u32)b.nd_state[nb + p].pad0 * 4;
                            const i32 pc0 = trace_col0(pbe.x);
                            const i32 adj = CHK ? b.nd_state[nb + p].pad1 - ref : 0;      // the predecessor row's ref (pad1) against this row's
#pragma unroll 2
                            for (i32 j = jlo + (i32)gl; j <= end + 1; j += G) {
                                i32 hv = NEG16, e1v = NEG16, e2v = NEG16;
                                if (j >= pbe.x && j <= pbe.y) {
                                    const u32 v = tw[j - pc0];
                                    hv = sext16(v);
                                    if (CHK) {
                                        // a reachable cell that cannot be expressed against this row's ref: the alignment goes to the 32-bit kernel
                                        if (hv >= -PAIR_UNREACH && hv + adj < -PAIR_SCORE_LIMIT) lost = true;
                                        hv = hv < -PAIR_UNREACH ? NEG16 : hv + adj;
                                    }
                                    e1v = hv - POE1 + (i32)((v >> 16) & 7u);
                                    e2v = hv - POE2 + (i32)((v >> 19) & 31u);
                                }
                                const int xx = hw_index(j);
                                if (x > 0) {
                                    hv = imax(hv, (i32)Hs[xx]); e1v = imax(e1v, (i32)E1s[xx]); e2v = imax(e2v, (i32)E2s[xx]);
                                }
                                Hs[xx] = (short)hv; E1s[xx] = (short)e1v; E2s[xx] = (short)e2v;
                            }
                        }
                        if (CHK && __any_sync(gm, lost)) status = IT_RETRY_WIDER;
                    }
                }
                // substitution scores of this graph base: look-up bytes for the read's selectors
                const u32 code = (base >> 1) & 3u;
                if (base == ((0x47544341u >> (8 * code)) & 0xffu)) {
                    lutA = mis4 ^ ((((u32)smis ^ (u32)smatch) & 0xffu) << (8 * code));
                    pr = selq;
                } else {                                                     // not A, C, T, G: selectors 5 = equal to this base, 4 = not
                    lutA = mis4;
                    if (base != custom_base) {
                        __syncwarp(gm);
                        for (int p = (int)gl; p < 4 * nblk; p += G) {
                            const i32 j = 8 * (p >> 2) - 8 + (p & 3);
                            const u32 clo = (j >= 1 && j <= qlen && q[j - 1] == base) ? 5u : 4u;
                            const u32 chi = (j + 4 >= 1 && j + 4 <= qlen && q[j + 3] == base) ? 5u : 4u;
                            ((u16*)selc)[p] = (u16)sel16_of(clo, chi);
                        }
                        custom_base = base;
                    }
                    pr = selc;
                }
                bdone = 0;
                nblocks_row = nblocks;
                cy.fc1 = NEG16; cy.fc2 = NEG16; cy.hq = 0;
                phase = PH_SPAN;
                if (CHK && status != VGA_ALN_OK) finish();
            }
        }
        // ---- C: one span of every group's row; all 32 lanes take part (groups without a row run it with stores disabled)
        __syncwarp();
        if (__all_sync(FULLM, phase == PH_DONE)) break;
        const bool live = phase == PH_SPAN;
        const i32 left = live ? nblocks_row - bdone : 0;
        const int nbw = __reduce_max_sync(FULLM, (u32)imin((left + G - 1) / G, NBMAX));     // blocks per lane of this step
        SpanOut so;
        if (NBMAX >= 3 && nbw == 3)
            so = span_step<G, (NBMAX >= 3 ? 3 : 1), CHK>(gl, live ? beg : 0, live ? end : -1, live ? bdone : 0, live, live && left <= G * 3, cy, planes, nblk, pr, lutA, lutB, vtab, tr);
        else if (NBMAX >= 2 && nbw == 2)
            so = span_step<G, (NBMAX >= 2 ? 2 : 1), CHK>(gl, live ? beg : 0, live ? end : -1, live ? bdone : 0, live, live && left <= G * 2, cy, planes, nblk, pr, lutA, lutB, vtab, tr);
        else
            so = span_step<G, 1, CHK>(gl, live ? beg : 0, live ? end : -1, live ? bdone : 0, live, live && left <= G, cy, planes, nblk, pr, lutA, lutB, vtab, tr);
        // ---- D: row epilogue
        if (live) {
            // a band cell below the int16 range of this class: the alignment is redone by the 32-bit kernel
            if (CHK && so.mn < -PAIR_SCORE_LIMIT) { status = IT_RETRY_WIDER; finish(); }
        }
        if (phase == PH_SPAN) {
            if (bdone == 0 || so.mx > mxrow) { mxrow = so.mx; rl = so.lpos; rr = so.rpos; }
            else if (so.mx == mxrow) rr = so.rpos;
            bdone += G * nbw;
            if (bdone >= nblocks_row) {
                lpos = rl; rpos = rr;
                ppb = beg; ppe = end;
                if (meta & 0x400u) {
                    // last row of its node: it may be a predecessor of later nodes' first rows
                    if (gl == 0) {
                        int4* dst = (int4*)(b.nd_state + nb + s);
                        dst[0] = make_int4(beg, end, lpos, rpos);
                        dst[1] = make_int4(ref + mxrow, (i32)(u32)((u64)(tr - b.pool) >> 2), ref, 0);
                    }
                    s++; o = 0;
                } else o++;
                row++;
                phase = PH_ROW;
                if (row == nrows) finish();
            }
        }
    }
}

template <int G, int NBMAX, int MINB, bool CHK>
static void launch_pair(const DevIndex& ix, const PoaBuffers& b, int n_sms, cudaStream_t st) {
    int per_sm = 0;
    const int groups_per_cta = PP_WARPS * 32 / G;
    int qmax = b.qmax_short < 16 ? 16 : b.qmax_short > 255 ? 255 : b.qmax_short;
    const int nblk = ((qmax + 9) >> 3) + 2;         // columns -8 .. qmax+1, plus the spare block
    const size_t smem_bytes = ((size_t)VT_WORDS + (size_t)groups_per_cta * 4 * nblk) * sizeof(uint4);
    static size_t attr_bytes = 0;
    if (smem_bytes > attr_bytes) {
        VGA_CUDA(cudaFuncSetAttribute(poa_fill_pair_kernel<G, NBMAX, MINB, CHK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
        attr_bytes = smem_bytes;
    }
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, poa_fill_pair_kernel<G, NBMAX, MINB, CHK>, PP_WARPS * 32, smem_bytes);
    if (per_sm < 1) per_sm = 1;
    long long ctas = (long long)n_sms * per_sm;
    const long long need = ((long long)b.n_items + groups_per_cta - 1) / groups_per_cta;
    if (ctas > need) ctas = need;
    if (ctas < 1) ctas = 1;
    poa_fill_pair_kernel<G, NBMAX, MINB, CHK><<<(unsigned)ctas, PP_WARPS * 32, smem_bytes, st>>>(ix, b, nblk);
}

// pair classes: 0 = alignments whose row count bounds every reachable score inside int16 (poa_pair_ok), 1 = longer
// subgraphs, same kernel with the per-row range check (poa_pair_checked_ok).  8 lanes per alignment, steps of up to
// nb_hint blocks per lane.
int launch_poa_fill_pair(const DevIndex& ix, const PoaBuffers& b, int cls, int n_sms, cudaStream_t st) {
    if (b.n_items == 0) return 0;
    static const int nbenv = getenv("VGA_POA_NBMAX") ? atoi(getenv("VGA_POA_NBMAX")) : 0;     // A/B measurements
    static const int genv = getenv("VGA_POA_G") ? atoi(getenv("VGA_POA_G")) : 0;
    const int nbmax = nbenv ? nbenv : b.nb_hint;
    // subgraphs much longer than their reads (nb_hint 3: most rows at full band width) keep 8 lanes x 3 blocks; otherwise a
    // row has ~9 blocks and 4 lanes x 2 blocks waste fewer block slots (C4: fill 170 -> 141 ms per 2 M reads, gpurun_out/r02d)
    const int g = genv ? genv : (nbmax >= 3 ? 8 : 4);
    if (cls == 1) launch_pair<8, 3, 4, true>(ix, b, n_sms, st);      // subgraphs far longer than the read: most rows at full band width
    // (tried: the same instance cut to 102 / 85 registers for 5 / 6 CTAs per SM, with and without reloading E' in sweep 2b:
    // 137-144 ms and 160-197 ms against 139 ms per 2 M reads of C4 — spills cost more than the extra warps hide; gpurun_out/r02i, r02r)
    else if (g == 4) { if (nbmax <= 2) launch_pair<4, 2, 4, false>(ix, b, n_sms, st); else launch_pair<4, 3, 4, false>(ix, b, n_sms, st); }
    else if (g == 2) { if (nbmax <= 2) launch_pair<2, 2, 4, false>(ix, b, n_sms, st); else launch_pair<2, 3, 4, false>(ix, b, n_sms, st); }
    else if (nbmax == 1) launch_pair<8, 1, 4, false>(ix, b, n_sms, st);
    else if (nbmax == 2) launch_pair<8, 2, 4, false>(ix, b, n_sms, st);
    else launch_pair<8, 3, 4, false>(ix, b, n_sms, st);
    return 1;
}

// the 16x2 kernel is compiled for abPOA's default gap model; match / mismatch come in through the query profile
bool poa_pair_checked_ok(i64 qlen, i64 nrows, const VgaParams& P) {
    (void)nrows;
    return qlen < 246 && P.o1 == PO1 && P.e1 == PE1 && P.o2 == PO2 && P.e2 == PE2 &&
           P.match >= 0 && P.match <= 64 && P.mismatch >= 0 && P.mismatch <= 64;
}
bool poa_pair_ok(i64 qlen, i64 nrows, const VgaParams& P) {
    const i64 step = P.mismatch > POE1 ? P.mismatch : POE1;
    return qlen < 246 && step * (nrows + qlen) <= PAIR_SCORE_LIMIT && P.o1 == PO1 && P.e1 == PE1 && P.o2 == PO2 && P.e2 == PE2 &&
           P.match >= 0 && P.match <= 64 && P.mismatch >= 0 && P.mismatch <= 64;
}

}  // namespace vga
