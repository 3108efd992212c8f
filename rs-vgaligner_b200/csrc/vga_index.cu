// Index upload: flattens the reference's `Index` (src/index.rs:36-90) into one device image.
// The k-mer lookup table stands for `kmer_pos_ref.contains(hash)` + `bhpf.get(hash)`
// (src/index.rs:319-320): open addressing keyed by the 2-bit packed k-mer, exact membership,
// identical positions in identical (table) order.  Only forward/forward positions are kept because
// map_reads hard-codes only_forward=true (src/map.rs:62, src/chain.rs:154-156).
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "internal.h"

namespace vga {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
const char* get_error() { return g_err.c_str(); }

// vga_init selects the device for the calling host thread (like cudaSetDevice) and records it as the process default for
// threads that never called it.
thread_local int g_device = -1;
static std::atomic<int> g_default_device{-1};

static inline u64 align_up(u64 x, u64 a) { return (x + a - 1) / a * a; }

static char ref_complement(char c, bool* ok) {    // src/dna.rs:19-40
    switch (c) {
        case 'a': return 't'; case 'c': return 'g'; case 't': return 'a'; case 'g': return 'c';
        case 'u': return 'a'; case 'A': return 'T'; case 'C': return 'G'; case 'T': return 'A';
        case 'G': return 'C'; case 'U': return 'A'; case 'N': return 'N';
        default: *ok = false; return 'N';
    }
}

void resolve_pointers(vga_index* ix) {
    const IndexHeader& h = ix->hdr;
    const u8* b = (const u8*)ix->d_blob;
    DevIndex& d = ix->dv;
    d.k = h.k; d.n_nodes = h.n_nodes; d.n_edges = h.n_edges; d.L = h.L;
    d.slot_mask = (u32)(h.n_slots - 1);
    d.seq_both = b + h.off_seq_both;
    d.seq_idx = (const u32*)(b + h.off_seq_idx);
    d.edge_idx = (const u32*)(b + h.off_edge_idx);
    d.edges_in = (const u32*)(b + h.off_edges_in);
    d.edges = (const u32*)(b + h.off_edges);
    d.pos2node = (const u32*)(b + h.off_pos2node);
    d.slots = (const Slot*)(b + h.off_slots);
    d.pos_tb = (const u32*)(b + h.off_pos_tb);
    d.pos_te = (const u32*)(b + h.off_pos_te);
    d.n_buckets = (u32)h.n_buckets;
    d.buckets = (const uint4*)(b + h.off_buckets);
    d.pos2 = (const uint2*)(b + h.off_pos2);
}

static int build_image(const VgaIndexView* v, bool keep_all, std::vector<u8>& img, IndexHeader& hdr) {
    if (!v || !v->seq_fwd || !v->node_ref || (v->n_edges && !v->edges)) {
        set_error("vga_index_upload: null pointer in view");
        return VGA_ERR_INVALID;
    }
    if (v->k == 0 || v->k > 32) { set_error("kmer length outside 1..32"); return VGA_ERR_UNSUPPORTED; }
    if (v->seq_len >= (1ull << 31) || v->n_edges >= (1ull << 32) || v->n_nodes >= (1ull << 31)) {
        set_error("graph too large for 32-bit device coordinates (positions must be < 2^31)");
        return VGA_ERR_UNSUPPORTED;
    }
    const u64 L = v->seq_len, N = v->n_nodes;
    // ---- k-mer table: walk each key's table run up to the delimiter, keep fwd/fwd entries
    std::vector<u32> ptb, pte;
    struct Ent { u64 key; u32 off, cnt; };
    std::vector<Ent> ents;
    ents.reserve(v->n_keys);
    for (u64 i = 0; i < v->n_keys; i++) {
        u64 p = v->key_offset[i];
        u32 off = (u32)ptb.size(), cnt = 0;
        while (true) {
            if (p >= v->n_table) { set_error("kmer_pos_table run without delimiter"); return VGA_ERR_INVALID; }
            u64 s = v->tab_start[p], e = v->tab_end[p];
            if (s == VGA_POS_DELIM && e == VGA_POS_DELIM) break;
            if (keep_all || (!(s & VGA_POS_REV) && !(e & VGA_POS_REV))) {
                // start > end is legal: a forward/forward k-mer may cross reverse handles (17-DRB1-3123-smooth has such links)
                const u64 sp = s & ~VGA_POS_REV, ep = e & ~VGA_POS_REV;
                if (sp >= L || ep > L || ep == 0) { set_error("k-mer position outside the linearisation"); return VGA_ERR_INVALID; }
                ptb.push_back((u32)sp | ((s & VGA_POS_REV) ? 0x80000000u : 0u));
                pte.push_back((u32)ep | ((e & VGA_POS_REV) ? 0x80000000u : 0u));
                cnt++;
            }
            p++;
        }
        if (v->k < 32 && (v->key_kmer2bit[i] >> (2 * v->k)) != 0) {
            set_error("key_kmer2bit has bits above 2k"); return VGA_ERR_INVALID;
        }
        if (cnt) ents.push_back({kmer_planes_from_2bit(v->key_kmer2bit[i], v->k), off, cnt});
    }
    if (ptb.size() >= (1ull << 32)) { set_error("too many k-mer positions"); return VGA_ERR_UNSUPPORTED; }
    u64 n_slots = 16;
    const char* ct_env = getenv("VGA_TABLE_COMPACT");
    // by default only for tables that would not stay in L2 as 16-byte slots (measured: C4 seed stage 7.8 -> 6.0 ms per 2 M reads,
    // C5 6.8 -> 5.6 ms; on the L2-resident 2.8 MB table of C2 the longer probe body costs 0.39 -> 0.44 ms; gpurun_out/r02g)
    const bool compact = !keep_all && v->k <= 16 && L < (1ull << 26) && ptb.size() + ents.size() < (1ull << 26) && !ents.empty() &&
                         (ct_env ? atoi(ct_env) != 0 : ents.size() >= (1u << 20));
    if (!compact) while (n_slots < 2 * ents.size()) n_slots <<= 1;            // the compact table replaces the slots
    if (n_slots > (1ull << 32)) { set_error("k-mer table too large"); return VGA_ERR_UNSUPPORTED; }
    // ---- compact bucket table (common.cuh) when the key fits 32 bits and every field its bits; VGA_TABLE_COMPACT=0: the
    // open-addressed 16-byte slots (A/B measurements and tests; also what k > 16 uses)
    std::vector<u64> bkt;
    std::vector<u32> pos2;                  // interleaved {start, end}
    u64 n_buckets = 0;
    {
        if (compact) {
            n_buckets = (ents.size() * 5 + 23) / 24 + 1;            // load 0.6: one bucket per 4.8 keys
            bkt.assign(n_buckets * 8, 0);
            pos2.reserve(2 * ptb.size() + 64);
            for (const Ent& en : ents) {
                const u32 key32 = (u32)(en.key & 0xffffu) | ((u32)(en.key >> 32) << 16);
                u32 payload;
                const u64 dlt = (u64)pte[en.off] - ptb[en.off] - v->k;      // unsigned: start > end wraps to a huge value
                if (en.cnt == 1 && dlt <= CT_MAX_INLINE_DLT) payload = CT_INLINE | ptb[en.off] | ((u32)dlt << 26);
                else {
                    const u32 first = (u32)(pos2.size() / 2);
                    if (en.cnt > CT_MAX_CNT) { pos2.push_back(en.cnt); pos2.push_back(0); }
                    for (u32 t = 0; t < en.cnt; t++) { pos2.push_back(ptb[en.off + t]); pos2.push_back(pte[en.off + t]); }
                    payload = first | ((en.cnt > CT_MAX_CNT ? CT_MAX_CNT + 1 : en.cnt) << 26);
                }
                u64 bi = ct_bucket(key32, (u32)n_buckets);
                for (;;) {
                    u64* slot = bkt.data() + bi * 8;
                    int x = 0;
                    while (x < 8 && slot[x] != 0) {
                        if ((u32)slot[x] == key32) { set_error("duplicate k-mer key in index"); return VGA_ERR_INVALID; }
                        x++;
                    }
                    if (x < 8) { slot[x] = (u64)key32 | ((u64)payload << 32); break; }
                    bi = bi + 1 == n_buckets ? 0 : bi + 1;
                }
            }
        }
    }
    // ---- layout
    memset(&hdr, 0, sizeof(hdr));
    hdr.magic = INDEX_MAGIC;
    hdr.k = v->k; hdr.n_nodes = (u32)N; hdr.n_edges = (u32)v->n_edges; hdr.L = (u32)L;
    hdr.n_keys = ents.size(); hdr.n_pos = ptb.size(); hdr.n_slots = n_slots;
    u64 o = align_up(sizeof(IndexHeader), 256);
    auto place = [&](u64& field, u64 bytes) { field = o; o = align_up(o + bytes, 256); };
    place(hdr.off_seq_both, 2 * L);
    place(hdr.off_seq_idx, 4 * (N + 1));
    place(hdr.off_edge_idx, 4 * (N + 1));
    place(hdr.off_edges_in, 4 * (N + 1));
    place(hdr.off_edges, 4 * v->n_edges);
    place(hdr.off_pos2node, 4 * L);
    place(hdr.off_slots, sizeof(Slot) * n_slots);
    place(hdr.off_pos_tb, compact ? 0 : 4 * ptb.size());
    place(hdr.off_pos_te, compact ? 0 : 4 * pte.size());
    hdr.n_buckets = n_buckets; hdr.n_pos2 = pos2.size() / 2;
    hdr.all_orient = keep_all ? 1 : 0;
    place(hdr.off_buckets, 64 * n_buckets);
    place(hdr.off_pos2, 4 * pos2.size());
    hdr.bytes = o;
    img.assign(o, 0);
    u8* b = img.data();
    // ---- fill
    memcpy(b + hdr.off_seq_both, v->seq_fwd, L);
    bool ok = true;
    for (u64 i = 0; i < L; i++) b[hdr.off_seq_both + L + i] = (u8)ref_complement((char)v->seq_fwd[L - 1 - i], &ok);
    if (!ok) { set_error("Input sequence base is not DNA (src/dna.rs:10-13)"); return VGA_ERR_INVALID; }
    u32* sidx = (u32*)(b + hdr.off_seq_idx);
    u32* eidx = (u32*)(b + hdr.off_edge_idx);
    u32* ein = (u32*)(b + hdr.off_edges_in);
    for (u64 i = 0; i <= N; i++) {
        sidx[i] = (u32)v->node_ref[3 * i];
        eidx[i] = (u32)v->node_ref[3 * i + 1];
        ein[i] = (u32)v->node_ref[3 * i + 2];
        if (i > 0 && (sidx[i] <= sidx[i - 1] || eidx[i] < eidx[i - 1])) { set_error("node_ref not increasing"); return VGA_ERR_INVALID; }
        // incoming edges are the first edges_to_node entries of the node's run (src/index.rs:559-606)
        if (i > 0 && ein[i - 1] > eidx[i] - eidx[i - 1]) { set_error("node_ref: edges_to_node exceeds the node's edge run"); return VGA_ERR_INVALID; }
    }
    if (sidx[0] != 0 || eidx[0] != 0) { set_error("node_ref[0] must start at sequence / edge offset 0"); return VGA_ERR_INVALID; }
    if (sidx[N] != L || eidx[N] != v->n_edges) { set_error("node_ref end marker mismatch"); return VGA_ERR_INVALID; }
    u32* ed = (u32*)(b + hdr.off_edges);
    for (u64 i = 0; i < v->n_edges; i++) {
        u64 h = v->edges[i];
        if ((h >> 1) == 0 || (h >> 1) > N) { set_error("edge handle outside 1..n_nodes"); return VGA_ERR_INVALID; }
        ed[i] = (u32)h;
    }
    u32* p2n = (u32*)(b + hdr.off_pos2node);
    for (u64 id = 1; id <= N; id++)
        for (u32 p = sidx[id - 1]; p < sidx[id]; p++) p2n[p] = (u32)id;   // get_bv_rank (src/index.rs:427-439)
    Slot* slots = (Slot*)(b + hdr.off_slots);
    if (!compact)
    for (const Ent& e : ents) {
        u64 h = kmer_slot_hash(e.key) & (n_slots - 1);
        while (slots[h].cnt != 0) {
            if (slots[h].key == e.key) { set_error("duplicate k-mer key in index"); return VGA_ERR_INVALID; }
            h = (h + 1) & (n_slots - 1);
        }
        slots[h].key = e.key; slots[h].off = e.off; slots[h].cnt = e.cnt;
    }
    if (!ptb.empty() && !compact) {
        memcpy(b + hdr.off_pos_tb, ptb.data(), 4 * ptb.size());
        memcpy(b + hdr.off_pos_te, pte.data(), 4 * pte.size());
    }
    if (n_buckets) {
        memcpy(b + hdr.off_buckets, bkt.data(), 64 * n_buckets);
        if (!pos2.empty()) memcpy(b + hdr.off_pos2, pos2.data(), 4 * pos2.size());
    }
    memcpy(b, &hdr, sizeof(hdr));
    return VGA_OK;
}

}  // namespace vga

using namespace vga;

extern "C" {

const char* vga_last_error(void) { return get_error(); }

int vga_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int vga_init(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        set_error("no CUDA device available (there is no CPU fallback)");
        return VGA_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= n) { set_error("vga_init: device ordinal out of range"); return VGA_ERR_INVALID; }
    if (cudaSetDevice(device) != cudaSuccess) { set_error("cudaSetDevice failed"); return VGA_ERR_CUDA; }
    g_device = device;
    g_default_device = device;
    return VGA_OK;
}

static int require_device() {
    if (g_device < 0) {
        const int d = g_default_device.load();
        int rc = vga_init(d < 0 ? 0 : d);
        if (rc != VGA_OK) return rc;
    }
    if (cudaSetDevice(g_device) != cudaSuccess) { set_error("cudaSetDevice failed"); return VGA_ERR_CUDA; }
    return VGA_OK;
}

int vga_index_upload(const VgaIndexView* view, vga_index_t** out) { return vga_index_upload_ex(view, 0, out); }

int vga_index_upload_ex(const VgaIndexView* view, uint32_t flags, vga_index_t** out) {
    if (!out) { set_error("vga_index_upload: out is null"); return VGA_ERR_INVALID; }
    if (flags & ~VGA_INDEX_ALL_ORIENTATIONS) { set_error("vga_index_upload_ex: unknown flag"); return VGA_ERR_INVALID; }
    *out = nullptr;
    int rc = require_device();
    if (rc != VGA_OK) return rc;
    std::vector<u8> img;
    IndexHeader hdr;
    rc = build_image(view, (flags & VGA_INDEX_ALL_ORIENTATIONS) != 0, img, hdr);
    if (rc != VGA_OK) return rc;
    vga_index* ix = new vga_index();
    try {
        ix->device = g_device;
        ix->hdr = hdr;
        VGA_CUDA(cudaMalloc(&ix->d_blob, hdr.bytes));
        VGA_CUDA(cudaMemcpy(ix->d_blob, img.data(), hdr.bytes, cudaMemcpyHostToDevice));
        resolve_pointers(ix);
    } catch (CudaFail& f) {
        if (ix->d_blob) cudaFree(ix->d_blob);
        delete ix;
        return f.code;
    }
    *out = ix;
    return VGA_OK;
}

int vga_index_image_size(const vga_index_t* ix, uint64_t* bytes) {
    if (!ix || !bytes) { set_error("null argument"); return VGA_ERR_INVALID; }
    *bytes = ix->hdr.bytes;
    return VGA_OK;
}

int vga_index_image_copy_to_device(const vga_index_t* ix, void* dev_dst) {
    if (!ix || !dev_dst) { set_error("null argument"); return VGA_ERR_INVALID; }
    try {
        VGA_CUDA(cudaSetDevice(ix->device));
        VGA_CUDA(cudaMemcpy(dev_dst, ix->d_blob, ix->hdr.bytes, cudaMemcpyDeviceToDevice));
    } catch (CudaFail& f) { return f.code; }
    return VGA_OK;
}

int vga_index_import_device(const void* dev_blob, uint64_t bytes, vga_index_t** out) {
    if (!dev_blob || !out || bytes < sizeof(IndexHeader)) { set_error("bad index image"); return VGA_ERR_INVALID; }
    *out = nullptr;
    int rc = require_device();
    if (rc != VGA_OK) return rc;
    vga_index* ix = new vga_index();
    try {
        ix->device = g_device;
        VGA_CUDA(cudaMemcpy(&ix->hdr, dev_blob, sizeof(IndexHeader), cudaMemcpyDeviceToHost));
        if (ix->hdr.magic != INDEX_MAGIC || ix->hdr.bytes != bytes) {
            set_error("index image header mismatch");
            delete ix;
            return VGA_ERR_INVALID;
        }
        VGA_CUDA(cudaMalloc(&ix->d_blob, bytes));
        VGA_CUDA(cudaMemcpy(ix->d_blob, dev_blob, bytes, cudaMemcpyDeviceToDevice));
        resolve_pointers(ix);
    } catch (CudaFail& f) {
        if (ix->d_blob) cudaFree(ix->d_blob);
        delete ix;
        return f.code;
    }
    *out = ix;
    return VGA_OK;
}

int vga_index_free(vga_index_t* ix) {
    if (!ix) return VGA_OK;
    cudaSetDevice(ix->device);
    free_workspace(ix);
    if (ix->d_blob) cudaFree(ix->d_blob);
    delete ix;
    return VGA_OK;
}

void vga_default_params(VgaParams* p) {
    if (!p) return;
    p->bandwidth = 50; p->max_gap = 1000; p->min_anchors = 3; p->best_n = 1;
    p->match = 2; p->mismatch = 4; p->o1 = 4; p->e1 = 2; p->o2 = 24; p->e2 = 1;
    p->wb = 10; p->wf = 0.01f; p->flags = 0;
}

}  // extern "C"
