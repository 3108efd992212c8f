// GPU index build — the k-mer table of `Index::build` (reference src/index.rs:109-281) on the device:
//   generate_kmers_parallel / generate_kmer_with_handle_orient   src/kmer.rs:277-505   (k-mer walks from every handle and offset)
//   stable sort by k-mer + consecutive dedup                       src/kmer.rs:298-301
//   generate_pos_on_ref_2                                          src/kmer.rs:816-928   (positions per k-mer, sorted, delimiter)
// The host keeps what is sequential and cheap (GFA parsing, linearisation: seq_fwd, NodeRef, edges — src/utils.rs:81-146) and
// hands those arrays in; out come exactly the arrays of VgaIndexView (key_kmer2bit, key_offset, tab_start, tab_end), identical to
// what the host build (host/index_build.cpp) produces — the tests compare them on every fixture graph.
//
// One thread per START (handle orientation, offset) walks the graph depth-first with the reference's explicit stack (neighbours
// pushed in stored order, popped LIFO), so every start visits its k-mers in the reference's order.  The reference's result order
// matters twice: its sort is STABLE on the k-mer alone and its dedup removes only ADJACENT identical entries.  Both are kept by
// laying the emitted entries out in generation order — per handle, forward orientation then reverse; inside a handle orientation
// first the k-mers that end inside the handle (offset ascending), then the walks (offset DESCENDING, because all offsets push
// onto one stack before the first pop; pop order inside an offset) — and sorting them with a stable radix sort.  A handle
// orientation that meets an 'N' anywhere contributes nothing (src/kmer.rs:401-403, 459-461).
// Limits (VGA_ERR_UNSUPPORTED, the host build stays the reference for such graphs): a walk stack deeper than WALK_STACK entries,
// more than WALK_BUDGET steps from one start (the microsatellite loops of 5-B3106: 17^10 walks), more than 2^31 k-mers.
// Roofline: the walks are latency-bound pointer chasing (report only); the sorts are CUB radix sorts (HBM).
#include <cub/cub.cuh>

#include <vector>

#include "common.cuh"
#include "internal.h"

namespace vga {

namespace {

static const int WALK_STACK = 96;
static const u32 WALK_BUDGET = 1u << 16;
enum : u32 { BF_N = 1u, BF_OVERFLOW = 2u, BF_NOT_DNA = 4u };

struct GraphDev {
    u32 k, n_nodes, L, n_edges;
    const u8* seq_fwd;      // L
    const u8* seq_rev;      // L (reverse complement)
    const u32* seq_idx;     // n_nodes + 1
    const u32* edge_idx;    // n_nodes + 1
    const u32* edges_in;    // n_nodes + 1
    const u32* edges;       // packed handles
    const u32* ho_base;     // 2 * n_nodes + 1: first start index of every handle orientation (fwd of node 1, rev of node 1, ...)
};

__device__ __forceinline__ u32 node_len(const GraphDev& g, u32 id) { return g.seq_idx[id] - g.seq_idx[id - 1]; }
// first position of a handle on its own strand (get_seq_pos, src/kmer.rs:752-770)
__device__ __forceinline__ u32 seqpos(const GraphDev& g, u32 id, bool rev) {
    return rev ? g.L - g.seq_idx[id - 1] - node_len(g, id) : g.seq_idx[id - 1];
}
__device__ __forceinline__ const u8* bases(const GraphDev& g, u32 id, bool rev) {
    return (rev ? g.seq_rev : g.seq_fwd) + seqpos(g, id, rev);
}
// handle_edges_iter(h, Right) of the HashGraph (handlegraph 0.5.0): right edges of a forward handle in stored order; for a reverse
// handle the LEFT edges of the node, flipped, in stored order (host/index_build.cpp HashGraph::edges_right)
__device__ __forceinline__ u32 out_degree(const GraphDev& g, u32 id, bool rev) {
    const u32 nin = g.edges_in[id - 1];
    return rev ? nin : g.edge_idx[id] - g.edge_idx[id - 1] - nin;
}
__device__ __forceinline__ u32 out_edge(const GraphDev& g, u32 id, bool rev, u32 x) {
    const u32 b = g.edge_idx[id - 1], nin = g.edges_in[id - 1];
    return rev ? (g.edges[b + x] ^ 1u) : g.edges[b + nin + x];
}
__device__ __forceinline__ u32 code_of(u8 c, u32& bad) {
    switch (c) {
        case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3;
        case 'N': bad |= BF_N; return 0;
        default: bad |= BF_NOT_DNA; return 0;
    }
}

// One walk of the reference's explicit stack from (handle orientation, offset): SINK receives every k-mer that passes through
// later handles, in pop order, as (packed k-mer, end position word, forks).  Returns 1 when the k-mer ends inside the handle
// (then *key0 is that k-mer and nothing is walked).  flags collects N / overflow / non-DNA.
template <class SINK>
__device__ __forceinline__ u32 walk_start(const GraphDev& g, u32 id, bool rev, u32 off, u32& flags, u64* key0, SINK&& sink) {
    const u32 k = g.k;
    const u32 len = node_len(g, id);
    const u8* s = bases(g, id, rev);
    u32 bad = 0;
    u64 key = 0;
    const u32 take = min(k, len - off);
    for (u32 t = 0; t < take; t++) key = (key << 2) | code_of(s[off + t], bad);
    if (bad) { flags |= bad; return 0; }                  // an N in the partial k-mer: the whole handle orientation is dropped
    if (take == k) { *key0 = key; return 1; }
    u32 st_h[WALK_STACK];
    u64 st_key[WALK_STACK];
    u16 st_len[WALK_STACK];
    u16 st_forks[WALK_STACK];
    int sp = 0;
    {
        const u32 deg = out_degree(g, id, rev);
        for (u32 x = 0; x < deg; x++) {
            if (sp >= WALK_STACK) { flags |= BF_OVERFLOW; return 0; }
            st_h[sp] = out_edge(g, id, rev, x); st_key[sp] = key; st_len[sp] = (u16)take; st_forks[sp] = (u16)(deg > 1 ? 1 : 0);
            sp++;
        }
    }
    u32 steps = 0;
    while (sp > 0) {
        if (++steps > WALK_BUDGET) { flags |= BF_OVERFLOW; return 0; }
        sp--;
        const u32 h = st_h[sp];
        u64 kk = st_key[sp];
        const u32 have = st_len[sp], forks = st_forks[sp];
        const u32 hid = h >> 1;
        const bool hrev = h & 1u;
        const u32 add = min(k - have, node_len(g, hid));
        const u8* hs = bases(g, hid, hrev);
        u32 bad2 = 0;
        for (u32 t = 0; t < add; t++) kk = (kk << 2) | code_of(hs[t], bad2);
        if (bad2) { flags |= bad2; return 0; }
        if (have + add == k) sink(kk, (seqpos(g, hid, hrev) + add) | (hrev ? 0x80000000u : 0u), forks);
        else {
            const u32 deg = out_degree(g, hid, hrev);
            for (u32 x = 0; x < deg; x++) {
                if (sp >= WALK_STACK) { flags |= BF_OVERFLOW; return 0; }
                st_h[sp] = out_edge(g, hid, hrev, x); st_key[sp] = kk; st_len[sp] = (u16)(have + add);
                st_forks[sp] = (u16)(forks + (deg > 1 ? 1 : 0));
                sp++;
            }
        }
    }
    return 0;
}

// start index s = ho_base[ho] + offset; ho = 2 * (id - 1) + (reverse ? 1 : 0)
__device__ __forceinline__ bool start_of(const GraphDev& g, u32 s, u32& ho, u32& off) {
    // binary search over ho_base (2 * n_nodes + 1 entries)
    u32 lo = 0, hi = 2 * g.n_nodes;
    while (hi - lo > 1) { const u32 mid = (lo + hi) >> 1; if (g.ho_base[mid] <= s) lo = mid; else hi = mid; }
    ho = lo; off = s - g.ho_base[lo];
    return true;
}

__global__ void ib_count_kernel(GraphDev g, u32 n_starts, u32* __restrict__ n0, u32* __restrict__ n1r, u32* __restrict__ ho_flags) {
    const u32 s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_starts) return;
    u32 ho, off;
    start_of(g, s, ho, off);
    const u32 id = (ho >> 1) + 1;
    const bool rev = ho & 1u;
    u32 b = 0, flags = 0;
    u64 key0 = 0;
    const u32 a = walk_start(g, id, rev, off, flags, &key0, [&](u64, u32, u32) { b++; });
    if (flags) atomicOr(ho_flags + ho, flags);
    n0[s] = flags ? 0u : a;
    // walks are laid out in offset-descending order inside a handle orientation: store the count at the mirrored index
    n1r[g.ho_base[ho] + (node_len(g, id) - 1 - off)] = flags ? 0u : b;
}

// starts of a dropped handle orientation contribute nothing
__global__ void ib_mask_kernel(GraphDev g, u32 n_starts, const u32* __restrict__ ho_flags, u32* __restrict__ n0, u32* __restrict__ n1r) {
    const u32 s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_starts) return;
    u32 ho, off;
    start_of(g, s, ho, off);
    if (ho_flags[ho] & BF_N) { n0[s] = 0; n1r[s] = 0; }
}

__global__ void ib_emit_kernel(GraphDev g, u32 n_starts, const u32* __restrict__ ho_flags, const u64* __restrict__ e0,
                               const u64* __restrict__ e1, u64* __restrict__ keys, u64* __restrict__ vals, u8* __restrict__ forks) {
    const u32 s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_starts) return;
    u32 ho, off;
    start_of(g, s, ho, off);
    if (ho_flags[ho] & BF_N) return;
    const u32 id = (ho >> 1) + 1;
    const bool rev = ho & 1u;
    const u32 hb = g.ho_base[ho], hb_next = g.ho_base[ho + 1];
    // output position of this handle orientation's first entry, its in-handle k-mers, its walks (see the file header)
    const u64 out_base = e0[hb] + e1[hb];
    const u64 t0 = e0[hb_next] - e0[hb];
    const u64 pos0 = out_base + (e0[s] - e0[hb]);
    const u64 pos1 = out_base + t0 + (e1[hb + (node_len(g, id) - 1 - off)] - e1[hb]);
    const u32 start_word = (seqpos(g, id, rev) + off) | (rev ? 0x80000000u : 0u);
    u32 flags = 0;
    u64 key0 = 0, w = pos1;
    const u32 a = walk_start(g, id, rev, off, flags, &key0, [&](u64 kk, u32 end_word, u32 fk) {
        keys[w] = kk;
        vals[w] = ((u64)start_word << 32) | end_word;
        forks[w] = (u8)fk;
        w++;
    });
    if (a) {
        keys[pos0] = key0;
        vals[pos0] = ((u64)start_word << 32) | (u64)((seqpos(g, id, rev) + off + g.k) | (rev ? 0x80000000u : 0u));
        forks[pos0] = 0;
    }
}

__global__ void ib_iota_kernel(u32 n, u32* __restrict__ idx) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) idx[i] = i;
}
__global__ void ib_gather_kernel(u32 n, const u32* __restrict__ idx, const u64* __restrict__ vals, const u8* __restrict__ forks,
                                 u64* __restrict__ vals_o, u8* __restrict__ forks_o) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    vals_o[i] = vals[idx[i]];
    forks_o[i] = forks[idx[i]];
}
// consecutive dedup on all fields (src/kmer.rs:301 with the derived PartialEq of GraphKmer, src/kmer.rs:47)
__global__ void ib_uniq_flag_kernel(u32 n, const u64* __restrict__ keys, const u64* __restrict__ vals, const u8* __restrict__ forks,
                                    u8* __restrict__ keep) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    keep[i] = (i == 0 || keys[i] != keys[i - 1] || vals[i] != vals[i - 1] || forks[i] != forks[i - 1]) ? 1 : 0;
}
__global__ void ib_head_flag_kernel(u32 n, const u64* __restrict__ keys, u32* __restrict__ head) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u;
}
// table with one delimiter after every group (KMER_POS_DELIMITER, src/kmer.rs:740-749): entry i of group gi lands at i + gi
__global__ void ib_table_kernel(u32 n, const u64* __restrict__ keys, const u64* __restrict__ vals, const u32* __restrict__ head,
                                const u32* __restrict__ gincl, u64* __restrict__ key_out, u64* __restrict__ key_off,
                                u64* __restrict__ tab_start, u64* __restrict__ tab_end) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 gi = gincl[i] - 1;
    const u64 at = (u64)i + gi;
    const u64 v = vals[i];
    const u32 sw = (u32)(v >> 32), ew = (u32)v;
    tab_start[at] = (u64)(sw & 0x7fffffffu) | ((sw >> 31) ? (1ull << 63) : 0ull);
    tab_end[at] = (u64)(ew & 0x7fffffffu) | ((ew >> 31) ? (1ull << 63) : 0ull);
    if (head[i]) { key_out[gi] = keys[i]; key_off[gi] = at; }
    if (i + 1 == n || head[i + 1]) { tab_start[at + 1] = ~0ull; tab_end[at + 1] = ~0ull; }
}

template <class T> struct DArr {
    T* p = nullptr;
    ~DArr() { if (p) cudaFree(p); }
    T* alloc(size_t n) { if (p) cudaFree(p); p = nullptr; VGA_CUDA(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T))); return p; }
};

}  // namespace
}  // namespace vga

using namespace vga;

struct vga_index_build {
    std::vector<u64> key_kmer2bit, key_offset, tab_start, tab_end;
    float ms_device = 0;
};

extern "C" {

int vga_index_build_device(const VgaGraphView* gv, uint32_t k, vga_index_build_t** out) {
    if (!gv || !out || !gv->seq_fwd || !gv->node_ref || (gv->n_edges && !gv->edges)) { set_error("vga_index_build_device: null argument"); return VGA_ERR_INVALID; }
    *out = nullptr;
    if (k == 0 || k > 32) { set_error("kmer length outside 1..32"); return VGA_ERR_UNSUPPORTED; }
    const u64 L = gv->seq_len, N = gv->n_nodes;
    if (L >= (1ull << 31) || N >= (1ull << 30) || gv->n_edges >= (1ull << 32)) { set_error("graph too large for 32-bit device coordinates"); return VGA_ERR_UNSUPPORTED; }
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); set_error("no CUDA device available (there is no CPU fallback)"); return VGA_ERR_NO_DEVICE; }
    vga_index_build* B = new vga_index_build();
    try {
        // ---- graph arrays to the device (u32 forms, reverse complement, start index of every handle orientation)
        std::vector<u32> sidx(N + 1), eidx(N + 1), ein(N + 1), edges(gv->n_edges), hob(2 * N + 1);
        for (u64 i = 0; i <= N; i++) { sidx[i] = (u32)gv->node_ref[3 * i]; eidx[i] = (u32)gv->node_ref[3 * i + 1]; ein[i] = (u32)gv->node_ref[3 * i + 2]; }
        if (sidx[0] != 0 || sidx[N] != L || eidx[N] != gv->n_edges) { set_error("node_ref does not match seq_len / n_edges"); throw CudaFail{VGA_ERR_INVALID}; }
        for (u64 i = 0; i < gv->n_edges; i++) {
            if ((gv->edges[i] >> 1) == 0 || (gv->edges[i] >> 1) > N) { set_error("edge handle outside 1..n_nodes"); throw CudaFail{VGA_ERR_INVALID}; }
            edges[i] = (u32)gv->edges[i];
        }
        std::vector<u8> rev(L);
        for (u64 i = 0; i < L; i++) {
            const u8 c = gv->seq_fwd[L - 1 - i];
            rev[i] = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : c == 'a' ? 't' : c == 'c' ? 'g' : c == 'g' ? 'c' : c == 't' ? 'a' :
                     (c == 'U' || c == 'u') ? (c == 'U' ? 'A' : 'a') : 'N';
        }
        u64 run = 0;
        for (u64 id = 1; id <= N; id++) {
            if (sidx[id] <= sidx[id - 1]) { set_error("node_ref not increasing"); throw CudaFail{VGA_ERR_INVALID}; }
            const u32 len = sidx[id] - sidx[id - 1];
            hob[2 * (id - 1)] = (u32)run; run += len;
            hob[2 * (id - 1) + 1] = (u32)run; run += len;
        }
        hob[2 * N] = (u32)run;
        for (u64 id = 1; id <= N; id++)      // max_degree / max_furcations of the reference default to 100 and never bind below that
            if (eidx[id] - eidx[id - 1] > 100) { set_error("GPU index build: a node with more than 100 edges (the reference's degree limits would bind); use the host build"); throw CudaFail{VGA_ERR_UNSUPPORTED}; }
        const u32 n_starts = (u32)run;                     // 2 L
        DArr<u8> d_fwd, d_rev, d_forks, d_forks2, d_keep;
        DArr<u32> d_sidx, d_eidx, d_ein, d_edges, d_hob, d_n0, d_n1r, d_flags, d_idx, d_idx2, d_head, d_gincl;
        DArr<u64> d_e0, d_e1, d_keys, d_keys2, d_vals, d_vals2, d_kout, d_koff, d_ts, d_te;
        DArr<u8> d_tmp;
        cudaStream_t st = nullptr;                          // default stream: this is an offline, blocking call
        VGA_CUDA(cudaMemcpy(d_fwd.alloc(L), gv->seq_fwd, L, cudaMemcpyHostToDevice));
        VGA_CUDA(cudaMemcpy(d_rev.alloc(L), rev.data(), L, cudaMemcpyHostToDevice));
        VGA_CUDA(cudaMemcpy(d_sidx.alloc(N + 1), sidx.data(), 4 * (N + 1), cudaMemcpyHostToDevice));
        VGA_CUDA(cudaMemcpy(d_eidx.alloc(N + 1), eidx.data(), 4 * (N + 1), cudaMemcpyHostToDevice));
        VGA_CUDA(cudaMemcpy(d_ein.alloc(N + 1), ein.data(), 4 * (N + 1), cudaMemcpyHostToDevice));
        VGA_CUDA(cudaMemcpy(d_edges.alloc(edges.size()), edges.data(), 4 * edges.size(), cudaMemcpyHostToDevice));
        VGA_CUDA(cudaMemcpy(d_hob.alloc(hob.size()), hob.data(), 4 * hob.size(), cudaMemcpyHostToDevice));
        GraphDev g;
        g.k = k; g.n_nodes = (u32)N; g.L = (u32)L; g.n_edges = (u32)gv->n_edges;
        g.seq_fwd = d_fwd.p; g.seq_rev = d_rev.p; g.seq_idx = d_sidx.p; g.edge_idx = d_eidx.p; g.edges_in = d_ein.p; g.edges = d_edges.p;
        g.ho_base = d_hob.p;
        cudaEvent_t ev0, ev1;
        VGA_CUDA(cudaEventCreate(&ev0)); VGA_CUDA(cudaEventCreate(&ev1));
        VGA_CUDA(cudaEventRecord(ev0, st));
        // ---- count, mask the dropped handle orientations, scan
        d_n0.alloc(n_starts); d_n1r.alloc(n_starts); d_flags.alloc(2 * N);
        VGA_CUDA(cudaMemsetAsync(d_flags.p, 0, 4 * 2 * N, st));
        const u32 TB = 128, GB = (n_starts + TB - 1) / TB;
        ib_count_kernel<<<GB, TB, 0, st>>>(g, n_starts, d_n0.p, d_n1r.p, d_flags.p);
        {
            // any overflow / non-DNA flag ends the build (reduced on the host: 2 N words)
            std::vector<u32> hf(2 * N);
            VGA_CUDA(cudaMemcpy(hf.data(), d_flags.p, 4 * 2 * N, cudaMemcpyDeviceToHost));
            u32 any = 0;
            for (u32 f : hf) any |= f;
            if (any & BF_NOT_DNA) { set_error("Input sequence base is not DNA (src/dna.rs:10-13)"); throw CudaFail{VGA_ERR_INVALID}; }
            if (any & BF_OVERFLOW) {
                set_error("GPU index build: a start has more walks than the device build admits (cyclic / high-degree graph); use the host build");
                throw CudaFail{VGA_ERR_UNSUPPORTED};
            }
        }
        ib_mask_kernel<<<GB, TB, 0, st>>>(g, n_starts, d_flags.p, d_n0.p, d_n1r.p);
        d_e0.alloc((size_t)n_starts + 1); d_e1.alloc((size_t)n_starts + 1);
        // exclusive sums with the total at index n_starts: a zero in front, the inclusive sum behind it
        size_t tb = 0;
        cub::DeviceScan::InclusiveSum(nullptr, tb, d_n0.p, d_e0.p + 1, (int)n_starts, st);
        d_tmp.alloc(tb);
        VGA_CUDA(cudaMemsetAsync(d_e0.p, 0, 8, st));
        VGA_CUDA(cudaMemsetAsync(d_e1.p, 0, 8, st));
        {
            size_t sb = tb;
            VGA_CUDA(cub::DeviceScan::InclusiveSum(d_tmp.p, sb, d_n0.p, d_e0.p + 1, (int)n_starts, st));
            sb = tb;
            VGA_CUDA(cub::DeviceScan::InclusiveSum(d_tmp.p, sb, d_n1r.p, d_e1.p + 1, (int)n_starts, st));
        }
        u64 tot0 = 0, tot1 = 0;
        VGA_CUDA(cudaMemcpy(&tot0, d_e0.p + n_starts, 8, cudaMemcpyDeviceToHost));
        VGA_CUDA(cudaMemcpy(&tot1, d_e1.p + n_starts, 8, cudaMemcpyDeviceToHost));
        const u64 M = tot0 + tot1;
        if (M >= (1ull << 31)) { set_error("GPU index build: more than 2^31 graph k-mers"); throw CudaFail{VGA_ERR_UNSUPPORTED}; }
        if (M == 0) { *out = B; cudaEventDestroy(ev0); cudaEventDestroy(ev1); return VGA_OK; }
        const u32 m = (u32)M;
        // ---- emit in generation order
        d_keys.alloc(m); d_vals.alloc(m); d_forks.alloc(m);
        ib_emit_kernel<<<GB, TB, 0, st>>>(g, n_starts, d_flags.p, d_e0.p, d_e1.p, d_keys.p, d_vals.p, d_forks.p);
        // ---- stable sort by k-mer (payload = generation index), gather, consecutive dedup
        d_keys2.alloc(m); d_idx.alloc(m); d_idx2.alloc(m); d_vals2.alloc(m); d_forks2.alloc(m); d_keep.alloc(m);
        {
            ib_iota_kernel<<<(m + 255) / 256, 256, 0, st>>>(m, d_idx.p);
            size_t sb = 0;
            cub::DeviceRadixSort::SortPairs(nullptr, sb, d_keys.p, d_keys2.p, d_idx.p, d_idx2.p, (int)m, 0, (int)(2 * k), st);
            if (sb > tb) { d_tmp.alloc(sb); tb = sb; }
            VGA_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp.p, sb, d_keys.p, d_keys2.p, d_idx.p, d_idx2.p, (int)m, 0, (int)(2 * k), st));
            VGA_CUDA(cudaStreamSynchronize(st));
        }
        const u32 GM = (m + 255) / 256;
        ib_gather_kernel<<<GM, 256, 0, st>>>(m, d_idx2.p, d_vals.p, d_forks.p, d_vals2.p, d_forks2.p);
        ib_uniq_flag_kernel<<<GM, 256, 0, st>>>(m, d_keys2.p, d_vals2.p, d_forks2.p, d_keep.p);
        // compact keys and positions (forks are not needed after the dedup)
        DArr<u32> d_nsel;
        d_nsel.alloc(1);
        {
            size_t sb = 0;
            cub::DeviceSelect::Flagged(nullptr, sb, d_keys2.p, d_keep.p, d_keys.p, d_nsel.p, (int)m, st);
            if (sb > tb) { d_tmp.alloc(sb); tb = sb; }
            VGA_CUDA(cub::DeviceSelect::Flagged(d_tmp.p, sb, d_keys2.p, d_keep.p, d_keys.p, d_nsel.p, (int)m, st));
            VGA_CUDA(cub::DeviceSelect::Flagged(d_tmp.p, sb, d_vals2.p, d_keep.p, d_vals.p, d_nsel.p, (int)m, st));
        }
        u32 n = 0;
        VGA_CUDA(cudaMemcpy(&n, d_nsel.p, 4, cudaMemcpyDeviceToHost));
        // ---- positions of a k-mer in the derived order of KmerPos (start orient, start, end orient, end): sort by the position
        // word, then stable by k-mer
        {
            size_t sb = 0;
            cub::DeviceRadixSort::SortPairs(nullptr, sb, d_vals.p, d_vals2.p, d_keys.p, d_keys2.p, (int)n, 0, 64, st);
            if (sb > tb) { d_tmp.alloc(sb); tb = sb; }
            VGA_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp.p, sb, d_vals.p, d_vals2.p, d_keys.p, d_keys2.p, (int)n, 0, 64, st));
            VGA_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp.p, sb, d_keys2.p, d_keys.p, d_vals2.p, d_vals.p, (int)n, 0, (int)(2 * k), st));
        }
        // ---- groups, key arrays, table with delimiters
        const u32 GN = (n + 255) / 256;
        d_head.alloc(n); d_gincl.alloc(n);
        ib_head_flag_kernel<<<GN, 256, 0, st>>>(n, d_keys.p, d_head.p);
        {
            size_t sb = 0;
            cub::DeviceScan::InclusiveSum(nullptr, sb, d_head.p, d_gincl.p, (int)n, st);
            if (sb > tb) { d_tmp.alloc(sb); tb = sb; }
            VGA_CUDA(cub::DeviceScan::InclusiveSum(d_tmp.p, sb, d_head.p, d_gincl.p, (int)n, st));
        }
        u32 n_keys = 0;
        VGA_CUDA(cudaMemcpy(&n_keys, d_gincl.p + (n - 1), 4, cudaMemcpyDeviceToHost));
        const u64 n_table = (u64)n + n_keys;
        d_kout.alloc(n_keys); d_koff.alloc(n_keys); d_ts.alloc(n_table); d_te.alloc(n_table);
        ib_table_kernel<<<GN, 256, 0, st>>>(n, d_keys.p, d_vals.p, d_head.p, d_gincl.p, d_kout.p, d_koff.p, d_ts.p, d_te.p);
        VGA_CUDA(cudaEventRecord(ev1, st));
        VGA_CUDA(cudaEventSynchronize(ev1));
        cudaEventElapsedTime(&B->ms_device, ev0, ev1);
        cudaEventDestroy(ev0); cudaEventDestroy(ev1);
        B->key_kmer2bit.resize(n_keys); B->key_offset.resize(n_keys); B->tab_start.resize(n_table); B->tab_end.resize(n_table);
        VGA_CUDA(cudaMemcpy(B->key_kmer2bit.data(), d_kout.p, 8ull * n_keys, cudaMemcpyDeviceToHost));
        VGA_CUDA(cudaMemcpy(B->key_offset.data(), d_koff.p, 8ull * n_keys, cudaMemcpyDeviceToHost));
        VGA_CUDA(cudaMemcpy(B->tab_start.data(), d_ts.p, 8ull * n_table, cudaMemcpyDeviceToHost));
        VGA_CUDA(cudaMemcpy(B->tab_end.data(), d_te.p, 8ull * n_table, cudaMemcpyDeviceToHost));
        // The reference's final-k-mer flush (src/kmer.rs:885-889) fires on every element equal to the last one: a duplicate of the
        // last GraphKmer inside the last group would hand boomphf a duplicate key.  The host build refuses such a graph too.
        if (n >= 2) {
            const size_t last = B->tab_start.size() - 2;      // last position entry (the delimiter follows)
            const size_t first = (size_t)B->key_offset.back();
            for (size_t x = first; x < last; x++)
                if (B->tab_start[x] == B->tab_start[last] && B->tab_end[x] == B->tab_end[last]) {
                    set_error("index build: the reference's final-kmer flush (src/kmer.rs:885) would fire twice");
                    throw CudaFail{VGA_ERR_UNSUPPORTED};
                }
        }
    } catch (CudaFail& f) {
        delete B;
        return f.code;
    }
    *out = B;
    return VGA_OK;
}

int vga_index_build_view(const vga_index_build_t* b, const VgaGraphView* g, uint32_t k, VgaIndexView* view, float* ms_device) {
    if (!b || !g || !view) { set_error("vga_index_build_view: null argument"); return VGA_ERR_INVALID; }
    view->k = k; view->seq_len = g->seq_len; view->n_nodes = g->n_nodes; view->n_edges = g->n_edges;
    view->n_keys = b->key_kmer2bit.size(); view->n_table = b->tab_start.size();
    view->seq_fwd = g->seq_fwd; view->node_ref = g->node_ref; view->edges = g->edges;
    view->key_kmer2bit = b->key_kmer2bit.data(); view->key_offset = b->key_offset.data();
    view->tab_start = b->tab_start.data(); view->tab_end = b->tab_end.data();
    if (ms_device) *ms_device = b->ms_device;
    return VGA_OK;
}

int vga_index_build_free(vga_index_build_t* b) { delete b; return VGA_OK; }

}  // extern "C"
