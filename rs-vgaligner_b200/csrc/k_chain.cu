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

#include "common.cuh"
#include "internal.h"

namespace vga {

static const int CHAIN_WARPS = 4;
static const int RING = 128;           // supports bandwidth <= 96 (checked by the caller)

// score_anchor (src/chain.rs:274-368) for forward/forward anchors of equal length k
__device__ __forceinline__ double score_pair(u32 qb_a, u32 tb_a, u32 te_a, double f_a, u32 qb_b, u32 tb_b,
                                             u32 te_b, u32 k, u32 max_gap, const double* __restrict__ gapcost) {
    if (qb_a >= qb_b || te_a >= te_b) return -DBL_MAX;               // :277-311
    const u32 ql = qb_b - qb_a;                                       // :313 (both terms are equal)
    const u32 tbd = tb_b > tb_a ? tb_b - tb_a : tb_a - tb_b;          // :324-327
    const u32 ted = te_b - te_a;                                      // :328-331
    const u32 tl = min(tbd, ted);
    const u32 g = ql > tl ? ql - tl : tl - ql;                        // :338-341
    if (g > max_gap) return -DBL_MAX;                                 // :343-344
    const double gc = __ldg(gapcost + g);                             // :348-354
    const u32 ml = min(min(ql, tl), k);                               // :357
    double x = __dadd_rn(f_a, (double)ml);
    x = __dsub_rn(x, gc);
    x = __dmul_rn(x, 1000.0);
    x = round(x);
    x = __ddiv_rn(x, 1000.0);
    return __dadd_rn(x, 0.0);                                         // + query_overlap (always 0) :360-363
}

__global__ void __launch_bounds__(CHAIN_WARPS * 32) chain_kernel(ChainBuffers b) {
    __shared__ u32 s_q[CHAIN_WARPS][RING], s_tb[CHAIN_WARPS][RING], s_te[CHAIN_WARPS][RING];
    __shared__ double s_f[CHAIN_WARPS][RING];
    const u32 lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 n_warps = (gridDim.x * blockDim.x) >> 5;
    u32* rq = s_q[wl]; u32* rtb = s_tb[wl]; u32* rte = s_te[wl]; double* rf = s_f[wl];
    const double kf = (double)b.k;
    for (u32 r = warp; r < b.n_reads; r += n_warps) {
        const u32 base = b.read_aoff[r];
        const u32 n = b.read_aoff[r + 1] - base;
        if (n == 0) {
            if (lane == 0) { b.n_chains[r] = 0; b.n_members[r] = 0; }
            continue;
        }
        double curr_max = 0.0;                                           // :398
        for (u32 i = 0; i < n; i++) {
            if ((i & 31) == 0) {           // stage the next 32 anchors of the sorted order (:386-389)
                const u32 ii = i + lane;
                if (ii < n) {
                    const u32 id = b.sidx[base + ii];
                    rq[ii % RING] = b.a_qb[base + id];
                    rtb[ii % RING] = b.a_tb[base + id];
                    rte[ii % RING] = b.a_te[base + id];
                }
                __syncwarp();
            }
            const u32 si = i % RING;
            const u32 qb_i = rq[si], tb_i = rtb[si], te_i = rte[si];
            double best = -DBL_MAX;
            u32 bestd = 0x7fffffffu;
            for (u32 d0 = 0; d0 < b.bandwidth; d0 += 32) {               // j = i-1 down to max(0, i-bandwidth) :404-417
                const u32 d = d0 + lane + 1;
                if (d <= b.bandwidth && d <= i) {
                    const u32 sj = (i - d) % RING;
                    const double s = score_pair(rq[sj], rtb[sj], rte[sj], rf[sj], qb_i, tb_i, te_i, b.k, b.max_gap, b.gapcost);
                    if (s > best) { best = s; bestd = d; }
                }
            }
            double m = best;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
            const u32 dmin = __reduce_min_sync(0xffffffffu, best == m ? bestd : 0x7fffffffu);
            double f_i = kf;                                             // Anchor::new: max_chain_score = k
            i32 p = -1;
            if (m > kf) { f_i = m; p = (i32)(i - dmin); }                // :430-443 strict >
            if (m > curr_max) curr_max = m;                              // :445-447 (all proposals)
            if (lane == 0) {
                rf[si] = f_i;
                b.f[base + i] = f_i;
                b.pred[base + i] = p;
                if (b.pred_dbg) b.pred_dbg[base + i] = p < 0 ? -1 : (i32)b.sidx[base + p];
            }
            __syncwarp();
        }
        // ---- backtracking (:455-558), sequential by construction
        if (lane == 0) {
            u32* creg = b.cmem + 2ull * base + 2ull * r;
            u32* cl = b.clen + base + r;
            u32 n_ch = 0, tail = 0;
            for (u32 i = n; i-- > 0;) {
                if (b.pred[base + i] >= 0 && b.f[base + i] == curr_max) {     // :469 exact f64 equality
                    const u32 start = tail;
                    u32 cur = i;
                    i32 p;
                    while ((p = b.pred[base + cur]) >= 0) {
                        b.pred[base + cur] = -1;                         // :478 consume
                        creg[tail++] = cur;
                        cur = (u32)p;
                    }
                    creg[tail++] = cur;                                  // :498 terminal (root or already consumed)
                    const u32 len = tail - start;
                    if (len >= b.min_anchors) {                          // :545-548
                        for (u32 x = 0; x < len / 2; x++) {
                            u32 t = creg[start + x];
                            creg[start + x] = creg[tail - 1 - x];
                            creg[tail - 1 - x] = t;
                        }
                        cl[n_ch++] = len;
                    } else {
                        tail = start;
                    }
                }
            }
            b.n_chains[r] = n_ch;
            b.n_members[r] = tail;
        }
        __syncwarp();
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

int launch_chain(const ChainBuffers& b, cudaStream_t st) {
    chain_kernel<<<grid_for_warps(b.n_reads, CHAIN_WARPS), CHAIN_WARPS * 32, 0, st>>>(b);
    return 1;
}

int launch_chain_gather(const ChainGather& g, cudaStream_t st) {
    chain_gather_kernel<<<grid_for_warps(g.n_reads, 8), 256, 0, st>>>(g);
    return 1;
}

}  // namespace vga
