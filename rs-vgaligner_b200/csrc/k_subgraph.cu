// K4 — chain -> node range -> extension -> aligner subgraph, a warp per (read, chain):
//   find_range_chain            reference src/align.rs:267-402   (Forward/Forward case; the others are
//                               unreachable because map_reads only emits forward anchors, src/map.rs:62)
//   extend_range_chain_2        src/align.rs:523-665
//   find_nodes_edges_for_abpoa  src/align.rs:670-724
//   Index graph accessors       src/index.rs:489-606
// The frontier walk of extend_range_chain_2 has no visited set in the reference (exponential on
// bubble chains); here a node is re-expanded only when reached with a larger remaining length,
// which yields the same handle SET — and the set (sort + dedup, :658-659) is all the function
// returns.  Latency-bound stage (pointer chasing over the graph); reported, no roofline claim.
#include "common.cuh"
#include "internal.h"

namespace vga {

static const int SUB_WARPS = 4;

struct NbrList { const u32* p; u32 n; u32 flip; };

// incoming_edges_from_handle / outgoing_edges_from_handle as SETS (src/index.rs:559-606): for a
// reverse handle the opposite list of the flipped handle, each entry flipped (order is irrelevant
// for the callers here except first-successor selection, which only sees forward lists below).
__device__ __forceinline__ NbrList neighbours(const DevIndex& ix, u32 h, bool left) {
    const u32 id = h >> 1, rev = h & 1;
    const u32 b = ix.edge_idx[id - 1], nin = ix.edges_in[id - 1], e = ix.edge_idx[id];
    NbrList l;
    if (left != (rev != 0)) { l.p = ix.edges + b; l.n = nin; }
    else { l.p = ix.edges + b + nin; l.n = e - b - nin; }
    l.flip = rev;
    return l;
}

__device__ __forceinline__ u32 node_len(const DevIndex& ix, u32 id) { return ix.seq_idx[id] - ix.seq_idx[id - 1]; }

__device__ __forceinline__ u32 lower_bound_u32(const u32* a, u32 n, u32 x) {
    u32 lo = 0, hi = n;
    while (lo < hi) { u32 mid = (lo + hi) >> 1; if (a[mid] < x) lo = mid + 1; else hi = mid; }
    return lo;
}

struct SubView {      // membership / position helpers over (range ∪ sorted ext)
    u32 mn, mx, n_range, n_ext;
    const u32* ext;
    __device__ __forceinline__ bool in_range(u32 h) const { return !(h & 1) && (h >> 1) >= mn && (h >> 1) <= mx; }
    __device__ __forceinline__ u32 range_below(u32 x) const {      // # range handles < x
        if (x <= 2 * mn) return 0;
        u32 c = (x - 2 * mn + 1) >> 1;
        return c < n_range ? c : n_range;
    }
    // position in the sorted handle list or 0xffffffff
    __device__ __forceinline__ u32 position(u32 h) const {
        if (in_range(h)) return ((h >> 1) - mn) + lower_bound_u32(ext, n_ext, h);
        u32 j = lower_bound_u32(ext, n_ext, h);
        if (j < n_ext && ext[j] == h) return j + range_below(h);
        return 0xffffffffu;
    }
    __device__ __forceinline__ u32 handle_at_range(u32 i) const { return 2 * (mn + i); }
};

// ---- BFS scratch: a generation-tagged hash (handle -> best remaining length) + a ring queue ----
struct Scratch {
    u32* hkey; u32* hval; u32* htag; u32 hmask;
    u32* q; u32 qcap;          // entries of 2 words
    u32* ext; u32 ecap;
    u32* gen;
};

__device__ __forceinline__ Scratch carve(u32* base, u32 words) {
    Scratch s;
    u32 hs = 1;
    while (hs * 2 * 3 <= words / 4) hs <<= 1;       // hash gets <= words/4
    s.hmask = hs - 1;
    s.hkey = base; s.hval = base + hs; s.htag = base + 2 * hs;
    u32 used = 3 * hs;
    s.gen = base + used; used += 1;
    s.qcap = (words - used) / 3 / 2;                // queue gets 2/3 of the rest
    s.q = base + used; used += s.qcap * 2;
    s.ext = base + used; s.ecap = words - used;
    return s;
}

// one directional walk of extend_range_chain_2 (:551-591 left, :616-656 right); lane 0 only
__device__ bool walk(const DevIndex& ix, Scratch& sc, u32 start, u32 budget, bool left, u32 mn, u32 mx,
                     u32& n_ext) {
    const u32 gen = ++(*sc.gen);
    u32 head = 0, tail = 0;      // monotone counters, ring indexed modulo qcap
    {
        NbrList l = neighbours(ix, start, left);
        for (u32 x = 0; x < l.n; x++) {
            if (tail - head >= sc.qcap) return false;
            sc.q[2 * (tail % sc.qcap)] = l.p[x] ^ l.flip;
            sc.q[2 * (tail % sc.qcap) + 1] = budget;
            tail++;
        }
    }
    u32 n_hash = 0;
    while (head != tail) {
        const u32 h = sc.q[2 * (head % sc.qcap)], rem = sc.q[2 * (head % sc.qcap) + 1];
        head++;
        u32 slot = (h * 2654435761u) & sc.hmask;
        bool found = false;
        while (sc.htag[slot] == gen) {
            if (sc.hkey[slot] == h) { found = true; break; }
            slot = (slot + 1) & sc.hmask;
        }
        if (found) {
            if (sc.hval[slot] >= rem) continue;
            sc.hval[slot] = rem;
        } else {
            if (++n_hash > (sc.hmask + 1) / 2) return false;
            sc.hkey[slot] = h; sc.hval[slot] = rem; sc.htag[slot] = gen;
            const bool inr = !(h & 1) && (h >> 1) >= mn && (h >> 1) <= mx;
            if (!inr) {
                if (n_ext >= sc.ecap) return false;
                sc.ext[n_ext++] = h;                            // :568 / :633 push
            }
        }
        const u32 len = node_len(ix, h >> 1);
        if (len < rem) {                                        // :572 / :637
            NbrList l = neighbours(ix, h, left);
            for (u32 x = 0; x < l.n; x++) {
                if (tail - head >= sc.qcap) return false;
                sc.q[2 * (tail % sc.qcap)] = l.p[x] ^ l.flip;
                sc.q[2 * (tail % sc.qcap) + 1] = rem - len;
                tail++;
            }
        }
    }
    return true;
}

__global__ void __launch_bounds__(SUB_WARPS * 32) sub_count_kernel(DevIndex ix, SubBuffers b) {
    const u32 lane = threadIdx.x & 31;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 n_warps = (gridDim.x * blockDim.x) >> 5;
    if (warp >= b.n_warps) return;
    Scratch sc = carve(b.scratch + (u64)warp * b.scratch_per_warp, b.scratch_per_warp);
    for (u32 w = warp; w < b.n_items; w += n_warps) {
        AlignItem it = b.items[w];
        const u64 cb = b.chain_anchor_off[it.chain], ce = b.chain_anchor_off[it.chain + 1];
        // ---- find_range_chain (:286-308): min/max handle over anchor starts and inclusive ends
        u32 mn = 0xffffffffu, mx = 0, rev_any = 0;
        for (u64 m = cb + lane; m < ce; m += 32) {
            const u32 tb = b.ca_tb[m], te = b.ca_te[m];
            rev_any |= (tb | te) >> 31;
            if ((tb | te) >> 31) continue;
            const u32 n1 = ix.pos2node[tb], n2 = ix.pos2node[te - 1];
            mn = min(mn, min(n1, n2));
            mx = max(mx, max(n1, n2));
        }
        if (__any_sync(0xffffffffu, rev_any)) {
            // RangeOrient::Reverse / Both (src/align.rs:365-387): only reachable with VGA_FLAG_ALL_ORIENTATIONS; not aligned (vga.h)
            if (lane == 0) {
                AlignItem* o = b.items + w;
                o->S = 0; o->nrows = 0; o->E = 0; o->qlen = (u32)(b.seq_off[it.read + 1] - b.seq_off[it.read]); o->q_off = b.seq_off[it.read];
                o->ext_off = 0; o->ext_cnt = 0; o->min_node = 1; o->max_node = 0;
                o->status = VGA_ALN_UNSUPPORTED;
            }
            continue;
        }
        mn = __reduce_min_sync(0xffffffffu, mn);
        mx = __reduce_max_sync(0xffffffffu, mx);
        const u32 qlen = (u32)(b.seq_off[it.read + 1] - b.seq_off[it.read]);
        // ---- extend_range_chain_2 budgets (:536-547, :593-612)
        u32 prefix = b.ca_qb[cb];
        const u32 slack = b.ca_tb[cb] - ix.seq_idx[mn - 1];
        prefix = slack < prefix ? prefix - slack : 0;
        u32 suffix = qlen - (b.ca_qb[ce - 1] + ix.k);
        const u32 eslack = ix.seq_idx[mx] - b.ca_te[ce - 1];      // select(last+1) - 1 - (te - 1)
        suffix = eslack > suffix ? 0 : suffix - eslack;
        u32 n_ext = 0, ok = 1;
        if (lane == 0) {
            if (prefix > 0) ok = walk(ix, sc, 2 * mn, prefix, true, mn, mx, n_ext);
            if (ok && suffix > 0) ok = walk(ix, sc, 2 * mx, suffix, false, mn, mx, n_ext);
            if (ok) {                                             // sort + dedup (:658-659)
                for (u32 i = 1; i < n_ext; i++) {
                    u32 v = sc.ext[i], j = i;
                    while (j > 0 && sc.ext[j - 1] > v) { sc.ext[j] = sc.ext[j - 1]; j--; }
                    sc.ext[j] = v;
                }
                u32 u = 0;
                for (u32 i = 0; i < n_ext; i++)
                    if (i == 0 || sc.ext[i] != sc.ext[i - 1]) sc.ext[u++] = sc.ext[i];
                n_ext = u;
            }
        }
        ok = __shfl_sync(0xffffffffu, ok, 0);
        n_ext = __shfl_sync(0xffffffffu, n_ext, 0);
        u32 eoff = 0;
        if (ok && lane == 0) {
            eoff = atomicAdd(b.ext_ctr, n_ext);
            if (eoff + n_ext > b.ext_cap) ok = 0;
        }
        ok = __shfl_sync(0xffffffffu, ok, 0);
        eoff = __shfl_sync(0xffffffffu, eoff, 0);
        if (!ok) {
            if (lane == 0) { atomicExch(b.overflow, 1u); }
            continue;
        }
        __syncwarp();
        for (u32 i = lane; i < n_ext; i += 32) b.ext_pool[eoff + i] = sc.ext[i];
        __syncwarp();
        // ---- sizes: nodes, rows, predecessor entries
        SubView sv{mn, mx, mx - mn + 1, n_ext, b.ext_pool + eoff};
        const u32 S = sv.n_range + n_ext;
        u32 rows = 0, E = 0;
        for (u32 p = lane; p < S; p += 32) {
            const u32 t = p < sv.n_range ? sv.handle_at_range(p) : sv.ext[p - sv.n_range];
            if (p >= sv.n_range) rows += node_len(ix, t >> 1);
            NbrList l = neighbours(ix, t, true);
            for (u32 x = 0; x < l.n; x++) {
                const u32 s = l.p[x] ^ l.flip;
                if (s < t && sv.position(s) != 0xffffffffu) E++;   // keep s<e (:717-718)
            }
        }
        rows = __reduce_add_sync(0xffffffffu, rows) + (ix.seq_idx[mx] - ix.seq_idx[mn - 1]);
        E = __reduce_add_sync(0xffffffffu, E);
        if (lane == 0) {
            AlignItem* o = b.items + w;
            o->S = S; o->nrows = rows; o->E = E; o->qlen = qlen; o->q_off = b.seq_off[it.read];
            o->ext_off = eoff; o->ext_cnt = n_ext; o->min_node = mn; o->max_node = mx;
            o->status = IT_PENDING;
            // row records keep max_remain in 21 bits (sub_rowmeta_kernel): a subgraph of 2^21 rows or more is refused, not clamped
            if (rows >= (1u << 21)) { o->S = 0; o->nrows = 0; o->E = 0; o->ext_cnt = 0; o->min_node = 1; o->max_node = 0; o->status = VGA_ALN_UNSUPPORTED; }
        }
    }
}

__global__ void __launch_bounds__(SUB_WARPS * 32) sub_fill_kernel(DevIndex ix, SubBuffers b) {
    const u32 lane = threadIdx.x & 31;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 n_warps = (gridDim.x * blockDim.x) >> 5;
    for (u32 w = warp; w < b.n_items; w += n_warps) {
        const AlignItem it = b.items[w];
        SubView sv{it.min_node, it.max_node, it.max_node - it.min_node + 1, it.ext_cnt, b.ext_pool + it.ext_off};
        const u32 S = it.S, nb = it.node_base;
        u32* H = b.nd_handle + nb;
        // handles in sorted order = merge of the contiguous range and the sorted extension list
        for (u32 i = lane; i < sv.n_range; i += 32) {
            const u32 h = sv.handle_at_range(i);
            H[i + lower_bound_u32(sv.ext, sv.n_ext, h)] = h;
        }
        for (u32 j = lane; j < sv.n_ext; j += 32) H[j + sv.range_below(sv.ext[j])] = sv.ext[j];
        __syncwarp();
        // row starts (labels concatenated in handle order) and label sources (seq_from_handle, src/index.rs:503-533)
        u32 carry = 0, pcarry = 0;
        for (u32 p0 = 0; p0 < S; p0 += 32) {
            const u32 p = p0 + lane;
            u32 len = 0, np = 0, t = 0;
            if (p < S) {
                t = H[p];
                const u32 id = t >> 1;
                len = node_len(ix, id);
                b.nd_src[nb + p] = (t & 1) ? ix.L + (ix.L - ix.seq_idx[id]) : ix.seq_idx[id - 1];
                NbrList l = neighbours(ix, t, true);
                for (u32 x = 0; x < l.n; x++) {
                    const u32 s = l.p[x] ^ l.flip;
                    if (s < t && sv.position(s) != 0xffffffffu) np++;
                }
            }
            u32 il = len, ip = np;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                u32 a = __shfl_up_sync(0xffffffffu, il, d), c = __shfl_up_sync(0xffffffffu, ip, d);
                if (lane >= (u32)d) { il += a; ip += c; }
            }
            if (p < S) {
                b.nd_row_start[nb + p] = carry + il - len;
                const u32 po = pcarry + ip - np;
                b.nd_pred_off[nb + p] = po;
                // predecessor node positions, ascending = the order find_nodes_edges_for_abpoa emits the
                // edges that end in t (sources ascending, :689-711)
                u32* dst = b.nd_pred + it.pred_base + po;
                u32 cnt = 0;
                NbrList l = neighbours(ix, t, true);
                for (u32 x = 0; x < l.n; x++) {
                    const u32 s = l.p[x] ^ l.flip;
                    if (!(s < t)) continue;
                    const u32 ps = sv.position(s);
                    if (ps == 0xffffffffu) continue;
                    u32 j = cnt++;
                    while (j > 0 && dst[j - 1] > ps) { dst[j] = dst[j - 1]; j--; }
                    dst[j] = ps;
                }
                // first successor in Index order (decides abPOA's max_remain path)
                i32 fs = -1;
                NbrList r = neighbours(ix, t, false);
                for (u32 x = 0; x < r.n; x++) {
                    const u32 e = r.p[r.flip ? r.n - 1 - x : x] ^ r.flip;    // reverse handle: flipped AND reversed list
                    if (e > t) { const u32 pe = sv.position(e); if (pe != 0xffffffffu) { fs = (i32)pe; break; } }
                }
                b.nd_first_succ[nb + p] = fs;
            }
            carry += __shfl_sync(0xffffffffu, il, 31);
            pcarry += __shfl_sync(0xffffffffu, ip, 31);
        }
        if (lane == 0) { b.nd_row_start[nb + S] = carry; b.nd_pred_off[nb + S] = pcarry; }
        __syncwarp();
        // max_remain along the first out-edge (abPOA abpoa_BFS_set_node_remain with equal weights)
        if (lane == 0) {
            for (u32 p = S; p-- > 0;) {
                const i32 fs = b.nd_first_succ[nb + p];
                i32 rl = 0;
                if (fs >= 0) {
                    const i32 flen = (i32)(b.nd_row_start[nb + fs + 1] - b.nd_row_start[nb + fs]);
                    rl = b.nd_remain_last[nb + fs] + flen - 1 + 1;
                }
                b.nd_remain_last[nb + p] = rl;
            }
        }
        __syncwarp();
    }
}

// Per-row record for the short-read POA kernel: graph base, how the row gets its predecessor row, whether it is the
// last row of its node, and max_remain of the row (abPOA's remaining graph length along the first out-edge):
//   bits 0..7 base | 8..9 kind | 10 last | 11..31 remain
// kind 0: the predecessor row is the previous row (inside a label, or a first row whose only predecessor is the
// previous node); 1: virtual source; 2: any other first row (predecessors merged from their trace words).
__global__ void __launch_bounds__(SUB_WARPS * 32) sub_rowmeta_kernel(DevIndex ix, SubBuffers b, u32* __restrict__ row_meta) {
    const u32 lane = threadIdx.x & 31;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 n_warps = (gridDim.x * blockDim.x) >> 5;
    for (u32 w = warp; w < b.n_items; w += n_warps) {
        const AlignItem it = b.items[w];
        const u32 nb = it.node_base;
        const u32* __restrict__ rs = b.nd_row_start + nb;
        u32* __restrict__ out = row_meta + it.row_base;
        // a lane per row; its node by binary search over the row starts (labels are short, so a loop over nodes would idle lanes)
        for (u32 row = lane; row < it.nrows; row += 32) {
            u32 lo = 0, hi = it.S;                       // rs[lo] <= row < rs[hi]
            while (hi - lo > 1) { const u32 mid = (lo + hi) >> 1; if (rs[mid] <= row) lo = mid; else hi = mid; }
            const u32 s = lo, r0 = rs[s], len = rs[s + 1] - r0, o = row - r0;
            u32 kind = 0;
            if (o == 0) {
                const u32 po = b.nd_pred_off[nb + s], pn = b.nd_pred_off[nb + s + 1] - po;
                kind = pn == 0 ? 1u : (pn == 1 && s > 0 && b.nd_pred[it.pred_base + po] == s - 1) ? 0u : 2u;
            }
            const u32 remain = (u32)b.nd_remain_last[nb + s] + (len - 1 - o);
            out[row] = (u32)ix.seq_both[b.nd_src[nb + s] + o] | (kind << 8) | ((o == len - 1 ? 1u : 0u) << 10) |
                       ((remain < (1u << 21) ? remain : (1u << 21) - 1) << 11);
        }
    }
}

static int grid_for_warps(u32 n_items, int warps_per_cta, u32 max_warps) {
    long long ctas = ((long long)n_items + warps_per_cta - 1) / warps_per_cta;
    long long cap = max_warps / warps_per_cta;
    if (ctas > cap) ctas = cap;
    if (ctas < 1) ctas = 1;
    return (int)ctas;
}

int launch_sub_count(const DevIndex& ix, const SubBuffers& b, cudaStream_t st) {
    sub_count_kernel<<<grid_for_warps(b.n_items, SUB_WARPS, b.n_warps), SUB_WARPS * 32, 0, st>>>(ix, b);
    return 1;
}

int launch_sub_fill(const DevIndex& ix, const SubBuffers& b, cudaStream_t st) {
    sub_fill_kernel<<<grid_for_warps(b.n_items, SUB_WARPS, 148 * 64), SUB_WARPS * 32, 0, st>>>(ix, b);
    return 1;
}

int launch_sub_rowmeta(const DevIndex& ix, const SubBuffers& b, u32* row_meta, cudaStream_t st) {
    sub_rowmeta_kernel<<<grid_for_warps(b.n_items, SUB_WARPS, 148 * 64), SUB_WARPS * 32, 0, st>>>(ix, b, row_meta);
    return 1;
}

}  // namespace vga
