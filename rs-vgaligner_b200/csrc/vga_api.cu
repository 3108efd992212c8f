// vga_map_batch: the batched replacement of map_reads' per-read loops (reference src/map.rs:56-111
// and 162-167).  Stages per chunk of reads, all on one stream:
//   H2D reads -> K1 seeding -> K2 segmented stable sort (CUB) -> K3 chaining -> K4 subgraphs ->
//   K5 POA fill + traceback (in waves that fit the trace pool) -> selection/compaction -> D2H.
// No CPU fallback: every stage is a CUDA kernel of this library (or CUB scan/sort).
#include <cub/cub.cuh>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"
#include "internal.h"

using namespace vga;

// ---------------------------------------------------------------------------------------------
// workspace: grow-only device buffers reused across calls
// ---------------------------------------------------------------------------------------------
struct DBuf {
    void* p = nullptr;
    size_t cap = 0;
    void* ensure(size_t bytes) {
        if (bytes > cap) {
            if (p) VGA_CUDA(cudaFree(p));
            p = nullptr; cap = 0;
            size_t want = bytes + bytes / 8 + 256;
            VGA_CUDA(cudaMalloc(&p, want));
            cap = want;
        }
        return p;
    }
    template <class T> T* get(size_t n) { return (T*)ensure(std::max<size_t>(n, 1) * sizeof(T)); }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

enum {
    B_SEQS, B_SEQOFF, B_KOFF, B_KCNT, B_KSLOT, B_ACNT, B_AOFF, B_AQB, B_ATB, B_ATE, B_KEYIN, B_VALIN, B_KEYOUT, B_VALOUT,
    B_CUB, B_BIG, B_F, B_PRED, B_PREDDBG, B_NCH, B_NMEM, B_CLEN, B_CMEM, B_CHOFF, B_MEMOFF, B_CAOFF, B_CAID, B_CAQB, B_CATB,
    B_CATE, B_CASC, B_GAP, B_GAPI, B_ITEMCNT, B_ITEMOFF, B_ITEMS, B_EXT, B_CTR, B_SCRATCH, B_SZ, B_SZOFF, B_NDH, B_NDRS,
    B_NDSRC, B_NDPO, B_NDPRED, B_NDFS, B_NDRL, B_NDSTATE, B_ROWS, B_ROWMETA,
    B_POOL, B_WORK, B_WORK2, B_STATUS2, B_TWORK, B_OUTRUNS, B_OUTPATH, B_SEL, B_RCNT, B_ROFF, B_PCNT, B_POFF2, B_RUNS, B_PATH, B_SCAL, B_STATUS,
    B_DBGSUBCNT, B_DBGSUBOFF, B_DBGSUB, B_DBGEDGECNT, B_DBGEDGEOFF, B_DBGEDGE, B_CELLS, B_COUNT
};

struct vga_workspace {
    cudaStream_t st = nullptr, st2 = nullptr;   // st2: result copies overlapped with the POA stage
    cudaStream_t st3 = nullptr;                 // the few long alignments of a short-read batch, beside the 16x2 kernel
    cudaEvent_t ev[12];
    cudaEvent_t ev_e0 = nullptr, ev_e1 = nullptr;   // span of the early wave on st3 (timed)
    cudaEvent_t ev_stf = nullptr;                   // end of the latest fill wave on st
    cudaEvent_t ev_copy = nullptr, ev_fork = nullptr, ev_join = nullptr;
    DBuf b[B_COUNT];
    int n_sms = 148;
    u32 scratch_words = 0;      // BFS scratch per warp currently allocated
    u32 sub_warps = 0;
    bool scratch_zeroed = false;
    u64 pool_capped_words = 0;  // size the trace pool was cut to by free memory (0: never cut)
    double gap_k = -1, gap_max = -1;
    bool gap_int_ok = false;
    // host-side scratch of the item classification, kept across calls (fresh vectors of this size cost milliseconds of page faults)
    std::vector<u64> h_est;
    std::vector<u32> h_key;
    std::vector<std::vector<u32>> h_cls;
    std::vector<u64> h_off;      // read offsets of the call on the host (chunking); kept for the same reason
    std::vector<u32> h_trace;    // alignments that wait for their traceback
};

namespace vga {
void free_workspace(vga_index* ix) {
    if (!ix->ws) return;
    for (auto& d : ix->ws->b) d.release();
    for (auto& e : ix->ws->ev) cudaEventDestroy(e);
    if (ix->ws->ev_copy) cudaEventDestroy(ix->ws->ev_copy);
    if (ix->ws->ev_fork) cudaEventDestroy(ix->ws->ev_fork);
    if (ix->ws->ev_join) cudaEventDestroy(ix->ws->ev_join);
    if (ix->ws->st2) cudaStreamDestroy(ix->ws->st2);
    if (ix->ws->st3) cudaStreamDestroy(ix->ws->st3);
    if (ix->ws->ev_e0) cudaEventDestroy(ix->ws->ev_e0);
    if (ix->ws->ev_e1) cudaEventDestroy(ix->ws->ev_e1);
    if (ix->ws->ev_stf) cudaEventDestroy(ix->ws->ev_stf);
    if (ix->ws->st) cudaStreamDestroy(ix->ws->st);
    delete ix->ws;
    ix->ws = nullptr;
}
}  // namespace vga

static vga_workspace* workspace(vga_index* ix) {
    if (ix->ws) return ix->ws;
    vga_workspace* w = new vga_workspace();
    VGA_CUDA(cudaStreamCreate(&w->st));
    VGA_CUDA(cudaStreamCreateWithFlags(&w->st2, cudaStreamNonBlocking));
    VGA_CUDA(cudaStreamCreateWithFlags(&w->st3, cudaStreamNonBlocking));
    VGA_CUDA(cudaEventCreate(&w->ev_e0));
    VGA_CUDA(cudaEventCreate(&w->ev_e1));
    VGA_CUDA(cudaEventCreate(&w->ev_stf));
    VGA_CUDA(cudaEventCreateWithFlags(&w->ev_copy, cudaEventDisableTiming));
    VGA_CUDA(cudaEventCreateWithFlags(&w->ev_fork, cudaEventDisableTiming));
    VGA_CUDA(cudaEventCreateWithFlags(&w->ev_join, cudaEventDisableTiming));
    for (auto& e : w->ev) VGA_CUDA(cudaEventCreate(&e));
    cudaDeviceProp prop;
    VGA_CUDA(cudaGetDeviceProperties(&prop, ix->device));
    w->n_sms = prop.multiProcessorCount;
    ix->ws = w;
    return w;
}

// ---------------------------------------------------------------------------------------------
// results: pinned host arrays (process-wide cache of pinned blocks, so repeated calls do not pay
// cudaHostAlloc again); one OutSet per chunk, merged when a call needed several chunks
// ---------------------------------------------------------------------------------------------
struct HBlock { void* p; size_t cap; };
static std::mutex g_pin_mu;
static std::vector<HBlock> g_pin_free;

// Requests are rounded up to quarter-octave size classes and matched exactly, so a steady stream of
// calls with similar batch sizes re-uses the same blocks (cudaHostAlloc of ~100 MB costs tens of ms).
static size_t pin_class(size_t bytes) {
    if (bytes < 4096) return 4096;
    size_t p = 4096;
    while (p * 2 <= bytes) p *= 2;            // p <= bytes < 2p
    const size_t q = p / 4;
    return p + (bytes - p + q - 1) / q * q;     // p, 1.25p, 1.5p, 1.75p, 2p
}
static size_t g_pin_free_bytes = 0;
static HBlock pin_get(size_t bytes) {
    const size_t cls = pin_class(bytes);
    {
        std::lock_guard<std::mutex> lk(g_pin_mu);
        for (int i = (int)g_pin_free.size() - 1; i >= 0; i--)
            if (g_pin_free[i].cap == cls) {
                HBlock b = g_pin_free[i];
                g_pin_free.erase(g_pin_free.begin() + i);
                g_pin_free_bytes -= b.cap;
                return b;
            }
    }
    HBlock b{nullptr, cls};
    VGA_CUDA(cudaHostAlloc(&b.p, b.cap, cudaHostAllocDefault));
    return b;
}
static void pin_put(HBlock b) {
    if (!b.p) return;
    std::lock_guard<std::mutex> lk(g_pin_mu);
    if (g_pin_free.size() >= 1024 || g_pin_free_bytes + b.cap > (8ull << 30)) { cudaFreeHost(b.p); return; }
    g_pin_free.push_back(b);
    g_pin_free_bytes += b.cap;
}

template <class T> struct HArr {
    HBlock blk{nullptr, 0};
    size_t n = 0;
    T* data() const { return (T*)blk.p; }
    void alloc(size_t count) { release(); blk = pin_get(std::max<size_t>(count, 1) * sizeof(T)); n = count; }
    void release() { pin_put(blk); blk = HBlock{nullptr, 0}; n = 0; }
    T& operator[](size_t i) { return data()[i]; }
};

struct OutSet {
    u32 n = 0;
    u64 NC = 0, NM = 0;
    HArr<u32> n_anchors, aln_chain, aln_path_len, aln_start_off, aln_end_off, aln_n_aligned;
    HArr<u64> chain_off, chain_anchor_off, aln_path_off, aln_path, aln_run_off, aln_cells;
    HArr<u32> ca_id, ca_qb, ca_tb, ca_te, aln_runs;
    HArr<double> ca_score;
    HArr<u8> aln_status;
    HArr<i32> aln_score;
    HArr<u64> dbg_anchor_off, dbg_sub_off, dbg_sub_handles, dbg_edge_off;
    HArr<u32> dbg_qb, dbg_tb, dbg_te, dbg_sorted, dbg_edges;
    HArr<double> dbg_f;
    HArr<i32> dbg_pred;
    void release_all() {
        n_anchors.release(); aln_chain.release(); aln_path_len.release(); aln_start_off.release(); aln_end_off.release();
        aln_n_aligned.release(); chain_off.release(); chain_anchor_off.release(); aln_path_off.release(); aln_path.release();
        aln_run_off.release(); aln_cells.release(); ca_id.release(); ca_qb.release(); ca_tb.release(); ca_te.release();
        aln_runs.release(); ca_score.release(); aln_status.release(); aln_score.release(); dbg_anchor_off.release();
        dbg_sub_off.release(); dbg_sub_handles.release(); dbg_edge_off.release(); dbg_qb.release(); dbg_tb.release();
        dbg_te.release(); dbg_sorted.release(); dbg_edges.release(); dbg_f.release(); dbg_pred.release();
    }
};

struct vga_results {
    u32 n = 0;
    bool debug = false, scores = false;
    OutSet o;
    std::vector<OutSet*> parts;
    VgaStats stats;
    ~vga_results() {
        o.release_all();
        for (OutSet* p : parts) { p->release_all(); delete p; }
    }
};

// ---------------------------------------------------------------------------------------------
// small glue kernels
// ---------------------------------------------------------------------------------------------
__global__ void item_count_kernel(const u32* n_chains, u32 n, u32 best_n, u32* cnt) {
    u32 r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n) cnt[r] = min(n_chains[r], best_n);      // first min(best_n, len) chains (src/align.rs:43-45)
}

__global__ void item_build_kernel(const u32* cnt, const u64* item_off, const u64* chain_off, u32 n, AlignItem* items) {
    u32 r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    for (u32 c = 0; c < cnt[r]; c++) {
        AlignItem it;
        memset(&it, 0, sizeof(it));
        it.read = r;
        it.chain = (u32)(chain_off[r] + c);
        it.status = IT_PENDING;
        items[item_off[r] + c] = it;
    }
}

// per item: sizes[0..3] = S+1, E, nrows, nrows+qlen+2
__global__ void item_sizes_kernel(const AlignItem* items, u32 n, u64* s0, u64* s1, u64* s2, u64* s3) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const AlignItem& it = items[i];
    s0[i] = it.S + 1; s1[i] = it.E; s2[i] = it.nrows; s3[i] = (u64)it.nrows + it.qlen + 2;
}

// s_out bit 31: the chain's anchors cover less than an eighth of the read — on long reads the mark of a chain through random k-mer
// hits (reverse-strand reads), whose band will leave the 1024 columns of the warp kernels: such an alignment is started on the wide
// kernel right away (a routing hint only: every class computes the same alignment)
__global__ void item_bases_kernel(AlignItem* items, u32 n, const u64* o0, const u64* o1, const u64* o2, const u64* o3,
                                  const u64* chain_anchor_off, u32 k, u32* nrows_out, u32* qlen_out, u32* s_out) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    AlignItem& it = items[i];
    it.node_base = (u32)o0[i]; it.pred_base = (u32)o1[i]; it.row_base = (u32)o2[i]; it.out_base = (u32)o3[i];
    const u64 n_anc = chain_anchor_off[it.chain + 1] - chain_anchor_off[it.chain];
    const u32 sparse = n_anc * k * 8 < (u64)it.qlen ? 0x80000000u : 0u;
    nrows_out[i] = it.nrows; qlen_out[i] = it.qlen; s_out[i] = it.S | sparse;
}

__global__ void item_status_kernel(const AlignItem* items, const u32* work, u32 n, u32* status) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) status[i] = items[work[i]].status;
}

// total DP cells over all alignment items of a chunk (the GCUPS numerator), one atomic per warp
__global__ void cells_sum_kernel(const AlignItem* items, u32 n, unsigned long long* total) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long v = i < n ? items[i].cells : 0ull;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(total, v);
}

// best_alignment_for_query (src/align.rs:34-55): stable sort by path_length descending, first wins.
// sc: SoA block of 6 arrays of n (chain, path_len, start_off, end_off, n_aligned, score)
__global__ void select_kernel(const AlignItem* items, const u32* cnt, const u64* item_off, u32 n, u32 max_c_status,
                              i32* sel, u8* status, u32* sc, u64* cells, u64* run_cnt, u64* path_cnt,
                              u64* sub_cnt, u64* edge_cnt) {
    u32 r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    i32 best = -1;
    u32 best_len = 0;
    for (u32 c = 0; c < cnt[r]; c++) {
        const AlignItem& it = items[item_off[r] + c];
        const u32 pl = it.status == VGA_ALN_OK ? it.path_len : 0;
        if (best < 0 || pl > best_len) { best = (i32)(item_off[r] + c); best_len = pl; }
    }
    sel[r] = best;
    u32 v[6] = {0, 0, 0, 0, 0, 0};
    u8 st = VGA_ALN_PLACEHOLDER;
    u64 rc = 0, pc = 0, ce = 0, sn = 0, en = 0;
    if (best >= 0) {
        const AlignItem& it = items[best];
        st = (u8)(it.status >= IT_PENDING ? max_c_status : it.status);
        v[0] = (u32)(best - (i32)item_off[r]);
        if (it.status == VGA_ALN_OK) {
            v[1] = it.path_len; v[2] = it.start_off; v[3] = it.end_off; v[4] = it.n_aligned; v[5] = (u32)it.score;
            rc = it.n_runs; pc = it.n_path;
        }
        ce = it.cells; sn = it.S; en = it.E;
    }
    status[r] = st; cells[r] = ce; run_cnt[r] = rc; path_cnt[r] = pc;
#pragma unroll
    for (int q = 0; q < 6; q++) sc[(u64)q * n + r] = v[q];
    if (sub_cnt) { sub_cnt[r] = sn; edge_cnt[r] = en; }
}

__global__ void gather_kernel(const AlignItem* items, const i32* sel, u32 n, const u64* run_off, const u64* path_off,
                              const u32* out_runs, const u32* out_path, const u32* nd_handle, u32* runs, u64* path) {
    const u32 lane = threadIdx.x & 31;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 n_warps = (gridDim.x * blockDim.x) >> 5;
    for (u32 r = warp; r < n; r += n_warps) {
        if (sel[r] < 0) continue;
        const AlignItem& it = items[sel[r]];
        if (it.status != VGA_ALN_OK) continue;
        for (u32 x = lane; x < it.n_runs; x += 32) runs[run_off[r] + x] = out_runs[it.out_base + it.run_pos + x];
        for (u32 x = lane; x < it.n_path; x += 32)
            path[path_off[r] + x] = nd_handle[it.node_base + out_path[it.node_base + it.path_pos + x]];
    }
}

__global__ void dbg_sub_kernel(const AlignItem* items, const i32* sel, u32 n, const u64* sub_off, const u64* edge_off,
                               const u32* nd_handle, const u32* nd_pred_off, const u32* nd_pred, u64* sub, u32* edges) {
    const u32 lane = threadIdx.x & 31;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 n_warps = (gridDim.x * blockDim.x) >> 5;
    for (u32 r = warp; r < n; r += n_warps) {
        if (sel[r] < 0) continue;
        const AlignItem& it = items[sel[r]];
        for (u32 x = lane; x < it.S; x += 32) {
            sub[sub_off[r] + x] = nd_handle[it.node_base + x];
            const u32 pb = nd_pred_off[it.node_base + x], pe = nd_pred_off[it.node_base + x + 1];
            for (u32 y = pb; y < pe; y++) {
                edges[2 * (edge_off[r] + y)] = nd_pred[it.pred_base + y];
                edges[2 * (edge_off[r] + y) + 1] = x;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// helpers
// ---------------------------------------------------------------------------------------------
struct Ctx {
    vga_index* ix;
    vga_workspace* ws;
    cudaStream_t st;
    const VgaParams* P;
    vga_results* R;
    u32 launches = 0;
};

// VGA_HOST_TIMELINE=1: print host wall-clock marks of every call (where the host thread waits);
// VGA_HOST_TIMELINE=<ms> (> 1): only for calls that took longer than that many milliseconds
static int host_timeline() { static int v = -1; if (v < 0) { const char* e = getenv("VGA_HOST_TIMELINE"); v = e ? atoi(e) : 0; } return v; }
static std::chrono::steady_clock::time_point g_tl0;
static std::vector<std::pair<double, const char*>> g_tl;
static void tl_mark(const char* what) {
    if (!host_timeline()) return;
    g_tl.push_back({std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - g_tl0).count(), what});
}
static void tl_flush(double wall_ms) {
    if (host_timeline() && (host_timeline() == 1 || wall_ms > host_timeline()))
        for (auto& m : g_tl) fprintf(stderr, "[vga-tl] %9.3f ms  %s\n", m.first, m.second);
    g_tl.clear();
}

// VGA_DEBUG_SYNC=1: synchronise and check after every stage so a device fault names its stage
static bool debug_sync() { static int v = -1; if (v < 0) { const char* e = getenv("VGA_DEBUG_SYNC"); v = e && *e == '1'; } return v; }
static void stage_check(cudaStream_t st, const char* name) {
    if (!debug_sync()) return;
    cudaError_t e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { set_error(std::string("stage ") + name + ": " + cudaGetErrorString(e)); throw CudaFail{VGA_ERR_CUDA}; }
    fprintf(stderr, "[vga] stage %s ok\n", name);
}

template <class T> static T d2h_one(const T* dptr, cudaStream_t st) {
    T v;
    VGA_CUDA(cudaMemcpyAsync(&v, dptr, sizeof(T), cudaMemcpyDeviceToHost, st));
    VGA_CUDA(cudaStreamSynchronize(st));
    return v;
}

static thread_local u64 g_d2h_bytes = 0;
// device -> pinned result array (async on `st`)
template <class T> static void d2h_into(HArr<T>& dst, const T* dptr, size_t n, cudaStream_t st) {
    dst.alloc(n);
    g_d2h_bytes += n * sizeof(T);
    if (n) VGA_CUDA(cudaMemcpyAsync(dst.data(), dptr, n * sizeof(T), cudaMemcpyDeviceToHost, st));
}

// inclusive scan of in[0..n) into out[1..n], out[0]=0  (u64 output)
template <class InT> static void scan_to_offsets(Ctx& c, const InT* in, u64* out, u32 n) {
    VGA_CUDA(cudaMemsetAsync(out, 0, sizeof(u64), c.st));
    if (n == 0) return;
    size_t tb = 0;
    cub::DeviceScan::InclusiveSum(nullptr, tb, in, out + 1, (int)n, c.st);
    void* tmp = c.ws->b[B_CUB].ensure(tb);
    VGA_CUDA(cub::DeviceScan::InclusiveSum(tmp, tb, in, out + 1, (int)n, c.st));
    c.launches += 2;
}

static void upload_gapcost(Ctx& c, u32 k, u32 max_gap) {
    vga_workspace* w = c.ws;
    if (w->gap_k == k && w->gap_max == max_gap) return;
    // gamma_c of the reference, evaluated on the host exactly as written (src/chain.rs:348-354):
    // 0.01 * seed_length * gap_length + 0.5 * log2(gap_length); entry 0 is the `gap_length == 0` arm.
    std::vector<double> g(max_gap + 1);
    g[0] = 0.0;
    for (u32 x = 1; x <= max_gap; x++) {
        volatile double a = 0.01 * (double)k;
        volatile double b = a * (double)x;
        volatile double l = 0.5 * std::log2((double)x);
        g[x] = b + l;
    }
    // Integer form of the same table for the all-integer DP of k_chain.cu: the reference's score is
    // round((f_a + m - gap_cost) * 1000) / 1000 with f_a = A / 1000 for an integer A, i.e. round(A + 1000 m - 1000 gap_cost)
    // up to the rounding errors of the three f64 operations (< 2e-7 for |score| < 2^18).  1000 * gap_cost(g) is a fixed
    // real number per g; when its fractional part is further than 1e-4 from 1/2 for EVERY g, no rounding error can move the
    // result across a rounding boundary and the f64 evaluation equals A + 1000 m - K[g] with K[g] = nearest integer —
    // exactly, for every A.  (k = 11, 15, max_gap = 1000: the closest is g = 881 at 9e-4.)  Otherwise the f64 body runs.
    std::vector<i32> gi(max_gap + 1);
    bool int_ok = max_gap < CHAIN_KMAX;
    gi[0] = 0;
    for (u32 x = 1; x <= max_gap && int_ok; x++) {
        const long double t = (long double)g[x] * 1000.0L;       // exact: a double times 1000 fits the 64-bit mantissa
        const long double fl = floorl(t);
        const long double fr = t - fl;
        if (fabsl(fr - 0.5L) < 1e-4L || t > 2.0e9L) int_ok = false;
        gi[x] = (i32)(fl + (fr > 0.5L ? 1 : 0));
    }
    double* d = w->b[B_GAP].get<double>(max_gap + 1);
    i32* di = w->b[B_GAPI].get<i32>(max_gap + 1);
    VGA_CUDA(cudaMemcpyAsync(d, g.data(), g.size() * sizeof(double), cudaMemcpyHostToDevice, c.st));
    VGA_CUDA(cudaMemcpyAsync(di, gi.data(), gi.size() * sizeof(i32), cudaMemcpyHostToDevice, c.st));
    VGA_CUDA(cudaStreamSynchronize(c.st));
    w->gap_k = k; w->gap_max = max_gap; w->gap_int_ok = int_ok;
}

// ---------------------------------------------------------------------------------------------
// one chunk of reads already resident on the device
// returns false when the chunk must be split (too many anchors for 32-bit offsets)
// ---------------------------------------------------------------------------------------------
static bool process_chunk(Ctx& c, const char* d_seqs, const u64* d_seq_off, u32 n, u64 n_bases, OutSet& O) {
    vga_workspace* w = c.ws;
    cudaStream_t st = c.st;
    const DevIndex& dv = c.ix->dv;
    const VgaParams& P = *c.P;
    vga_results& R = *c.R;
    const bool dbg = P.flags & VGA_FLAG_KEEP_DEBUG;
    const bool want_scores = dbg || (P.flags & VGA_FLAG_CHAIN_SCORES);
    cudaStream_t st2 = w->st2;
    VGA_CUDA(cudaEventRecord(w->ev[0], st));
    u32 A_total = 0;
    // ---- K1 seeding
    SeedBuffers sb;
    sb.seqs = d_seqs; sb.seq_off = d_seq_off; sb.n_reads = n;
    sb.read_koff = w->b[B_KOFF].get<u64>(n + 1);
    c.launches += launch_read_prep(dv, sb, st);
    stage_check(st, "read_prep");
    {
        size_t tb = 0;
        cub::DeviceScan::InclusiveSum(nullptr, tb, sb.read_koff + 1, sb.read_koff + 1, (int)n, st);
        void* tmp = w->b[B_CUB].ensure(tb);
        VGA_CUDA(cub::DeviceScan::InclusiveSum(tmp, tb, sb.read_koff + 1, sb.read_koff + 1, (int)n, st));
        c.launches += 2;
    }
    const u64 total_k = d2h_one(sb.read_koff + n, st);
    sb.kcnt = w->b[B_KCNT].get<u32>(total_k);
    sb.koff = w->b[B_KSLOT].get<u32>(total_k);
    sb.read_acnt = w->b[B_ACNT].get<u32>(n);
    sb.read_aoff = w->b[B_AOFF].get<u32>(n + 1);
    c.launches += launch_seed_count(dv, sb, st);
    stage_check(st, "seed_count");
    // anchors per read -> offsets; reads with more than CHAIN_SMAX anchors are listed for the CUB sort path.
    // Anchors may not exceed 2^31 per chunk (32-bit anchor offsets, CUB takes int counts).
    u32* big_ctr = w->b[B_CTR].get<u32>(16) + 8;
    u32* big_list = w->b[B_BIG].get<u32>(3ull * n);
    u32 n_big = 0;
    {
        u64* tmp64 = w->b[B_SZOFF].get<u64>(n + 1);
        scan_to_offsets(c, sb.read_acnt, tmp64, n);
        VGA_CUDA(cudaMemsetAsync(big_ctr, 0, 4, st));
        c.launches += launch_chain_prep(tmp64, n, sb.read_aoff, big_ctr, big_list, big_list + n, big_list + 2ull * n, st);
        u64 total64 = 0;
        VGA_CUDA(cudaMemcpyAsync(&total64, tmp64 + n, 8, cudaMemcpyDeviceToHost, st));
        VGA_CUDA(cudaMemcpyAsync(&n_big, big_ctr, 4, cudaMemcpyDeviceToHost, st));
        VGA_CUDA(cudaStreamSynchronize(st));
        if (total64 >= (1ull << 31) - 1) return false;
        A_total = (u32)total64;
    }
    const u32 A = A_total;
    sb.a_qb = w->b[B_AQB].get<u32>(A); sb.a_tb = w->b[B_ATB].get<u32>(A); sb.a_te = w->b[B_ATE].get<u32>(A);
    sb.sort_key_in = w->b[B_KEYIN].get<u32>(A); sb.sort_val_in = w->b[B_VALIN].get<u32>(A);
    u32* key_out = w->b[B_KEYOUT].get<u32>(A);
    u32* val_out = w->b[B_VALOUT].get<u32>(A);
    c.launches += launch_seed_emit(dv, sb, st);
    stage_check(st, "seed_emit");
    VGA_CUDA(cudaEventRecord(w->ev[1], st));
    tl_mark("seed done (host)");
    // ---- K2 + K3: per-read stable sort by target_end.position (src/chain.rs:386-389) and chaining.
    // Reads with up to CHAIN_SMAX anchors: one fused kernel (shared-memory bitonic sort + DP + backtrack).
    // The others: CUB segmented sort over just their segments, then the same DP on the sorted order.
    upload_gapcost(c, dv.k, P.max_gap);
    ChainBuffers cb;
    cb.n_reads = n; cb.read_aoff = sb.read_aoff; cb.a_qb = sb.a_qb; cb.a_tb = sb.a_tb; cb.a_te = sb.a_te;
    cb.sidx = val_out; cb.gapcost = (const double*)w->b[B_GAP].p;
    cb.gapcost_i = w->gap_int_ok ? (const i32*)w->b[B_GAPI].p : nullptr;
    cb.big_list = big_list; cb.n_big = n_big;
    cb.f = w->b[B_F].get<double>(A); cb.pred = w->b[B_PRED].get<i32>(A);
    cb.pred_dbg = dbg ? w->b[B_PREDDBG].get<i32>(A) : nullptr;
    cb.n_chains = w->b[B_NCH].get<u32>(n); cb.n_members = w->b[B_NMEM].get<u32>(n);
    cb.clen = w->b[B_CLEN].get<u32>((u64)A + n); cb.cmem = w->b[B_CMEM].get<u32>(2ull * A + 2ull * n);
    cb.k = dv.k; cb.bandwidth = P.bandwidth; cb.max_gap = P.max_gap; cb.min_anchors = P.min_anchors;
    // The few reads with more anchors than the fused kernel takes are a long serial tail on a handful of warps: their CUB
    // sort + DP go to the second stream FIRST (it is idle here) and run beside the fused kernel, which skips those reads.
    if (n_big > 0) {
        VGA_CUDA(cudaEventRecord(w->ev_fork, st));
        VGA_CUDA(cudaStreamWaitEvent(st2, w->ev_fork, 0));
        size_t tb = 0;
        cub::DeviceSegmentedSort::StableSortPairs(nullptr, tb, sb.sort_key_in, key_out, sb.sort_val_in, val_out, (int)A,
                                                  (int)n_big, big_list + n, big_list + 2ull * n, st2);
        void* tmp = w->b[B_CUB].ensure(tb);
        VGA_CUDA(cub::DeviceSegmentedSort::StableSortPairs(tmp, tb, sb.sort_key_in, key_out, sb.sort_val_in, val_out,
                                                           (int)A, (int)n_big, big_list + n, big_list + 2ull * n, st2));
        c.launches += 3;
        c.launches += launch_chain(cb, false, st2);
        stage_check(st2, "chain (big reads)");
        VGA_CUDA(cudaEventRecord(w->ev_join, st2));
    }
    if (n_big < n) c.launches += launch_chain(cb, true, st);
    stage_check(st, "chain (fused)");
    if (n_big > 0) VGA_CUDA(cudaStreamWaitEvent(st, w->ev_join, 0));
    VGA_CUDA(cudaEventRecord(w->ev[2], st));
    tl_mark("sort+chain queued");
    u64* chain_off = w->b[B_CHOFF].get<u64>(n + 1);
    u64* member_off = w->b[B_MEMOFF].get<u64>(n + 1);
    scan_to_offsets(c, cb.n_chains, chain_off, n);
    scan_to_offsets(c, cb.n_members, member_off, n);
    u64 NC = 0, NM = 0;
    VGA_CUDA(cudaMemcpyAsync(&NC, chain_off + n, 8, cudaMemcpyDeviceToHost, st));
    VGA_CUDA(cudaMemcpyAsync(&NM, member_off + n, 8, cudaMemcpyDeviceToHost, st));
    VGA_CUDA(cudaStreamSynchronize(st));
    ChainGather cg;
    cg.n_reads = n; cg.read_aoff = sb.read_aoff; cg.a_qb = sb.a_qb; cg.a_tb = sb.a_tb; cg.a_te = sb.a_te; cg.sidx = val_out;
    cg.f = cb.f; cg.n_chains = cb.n_chains; cg.clen = cb.clen; cg.cmem = cb.cmem; cg.chain_off = chain_off;
    cg.member_off = member_off;
    cg.chain_anchor_off = w->b[B_CAOFF].get<u64>(NC + 1);
    cg.ca_id = w->b[B_CAID].get<u32>(NM); cg.ca_qb = w->b[B_CAQB].get<u32>(NM); cg.ca_tb = w->b[B_CATB].get<u32>(NM);
    cg.ca_te = w->b[B_CATE].get<u32>(NM); cg.ca_score = w->b[B_CASC].get<double>(NM);
    c.launches += launch_chain_gather(cg, st);
    stage_check(st, "chain_gather");
    VGA_CUDA(cudaEventRecord(w->ev[3], st));

    tl_mark("chain_gather queued");
    // ---- chain results leave on the copy stream while the alignment stages run.  They are queued
    // only after the first POA fill launch (or right away when nothing gets aligned): the small
    // blocking D2H reads of the stages in between would otherwise wait behind ~100 MB on the copy engine.
    O.n = n; O.NC = NC; O.NM = NM;
    bool chain_d2h_queued = false;
    auto queue_chain_d2h = [&]() {
        if (chain_d2h_queued) return;
        chain_d2h_queued = true;
        VGA_CUDA(cudaStreamWaitEvent(st2, w->ev_copy, 0));
        d2h_into(O.n_anchors, (const u32*)sb.read_acnt, n, st2);
        d2h_into(O.chain_off, (const u64*)chain_off, (size_t)n + 1, st2);
        d2h_into(O.chain_anchor_off, (const u64*)cg.chain_anchor_off, NC + 1, st2);
        d2h_into(O.ca_qb, (const u32*)cg.ca_qb, NM, st2); d2h_into(O.ca_tb, (const u32*)cg.ca_tb, NM, st2);
        d2h_into(O.ca_te, (const u32*)cg.ca_te, NM, st2);
        if (want_scores) { d2h_into(O.ca_id, (const u32*)cg.ca_id, NM, st2); d2h_into(O.ca_score, (const double*)cg.ca_score, NM, st2); }
        if (dbg) {
            d2h_into(O.dbg_anchor_off, (const u64*)w->b[B_DBGSUBCNT].p, (size_t)n + 1, st2);
            d2h_into(O.dbg_qb, (const u32*)sb.a_qb, A, st2); d2h_into(O.dbg_tb, (const u32*)sb.a_tb, A, st2);
            d2h_into(O.dbg_te, (const u32*)sb.a_te, A, st2); d2h_into(O.dbg_sorted, (const u32*)val_out, A, st2);
            d2h_into(O.dbg_f, (const double*)cb.f, A, st2); d2h_into(O.dbg_pred, (const i32*)cb.pred_dbg, A, st2);
        }
    };
    if (dbg) {
        u64* aoff64 = w->b[B_DBGSUBCNT].get<u64>(n + 1);
        scan_to_offsets(c, sb.read_acnt, aoff64, n);
    }
    VGA_CUDA(cudaEventRecord(w->ev_copy, st));
    R.stats.n_probes += total_k; R.stats.n_anchors += A; R.stats.n_positions_read += A; R.stats.n_chains += NC;
    R.stats.n_bases += n_bases;
    R.stats.lookup_bytes += n_bases + 32ull * total_k + 8ull * A + 16ull * A;

    tl_mark("chain d2h queued");
    // ---- alignment
    u32 NI = 0;
    u32* item_cnt = w->b[B_ITEMCNT].get<u32>(n);
    u64* item_off = w->b[B_ITEMOFF].get<u64>(n + 1);
    AlignItem* items = nullptr;
    if (!(P.flags & VGA_FLAG_NO_ALIGN)) {
        item_count_kernel<<<(n + 255) / 256, 256, 0, st>>>(cb.n_chains, n, P.best_n, item_cnt);
        c.launches++;
        scan_to_offsets(c, item_cnt, item_off, n);
        NI = (u32)d2h_one(item_off + n, st);
    }
    if (NI == 0) {
        // every read is a placeholder (or alignment was not requested)
        queue_chain_d2h();
        O.aln_status.alloc(n); O.aln_chain.alloc(n); O.aln_path_len.alloc(n); O.aln_start_off.alloc(n); O.aln_end_off.alloc(n);
        O.aln_n_aligned.alloc(n); O.aln_score.alloc(n); O.aln_cells.alloc(n); O.aln_path_off.alloc((size_t)n + 1);
        O.aln_run_off.alloc((size_t)n + 1); O.aln_path.alloc(0); O.aln_runs.alloc(0);
        memset(O.aln_status.data(), VGA_ALN_PLACEHOLDER, n); memset(O.aln_chain.data(), 0, 4ull * n);
        memset(O.aln_path_len.data(), 0, 4ull * n); memset(O.aln_start_off.data(), 0, 4ull * n);
        memset(O.aln_end_off.data(), 0, 4ull * n); memset(O.aln_n_aligned.data(), 0, 4ull * n);
        memset(O.aln_score.data(), 0, 4ull * n); memset(O.aln_cells.data(), 0, 8ull * n);
        memset(O.aln_path_off.data(), 0, 8ull * (n + 1)); memset(O.aln_run_off.data(), 0, 8ull * (n + 1));
        if (dbg) {
            O.dbg_sub_off.alloc((size_t)n + 1); O.dbg_edge_off.alloc((size_t)n + 1); O.dbg_sub_handles.alloc(0); O.dbg_edges.alloc(0);
            memset(O.dbg_sub_off.data(), 0, 8ull * (n + 1)); memset(O.dbg_edge_off.data(), 0, 8ull * (n + 1));
        }
        VGA_CUDA(cudaEventRecord(w->ev[4], st)); VGA_CUDA(cudaEventRecord(w->ev[5], st)); VGA_CUDA(cudaEventRecord(w->ev[6], st));
        VGA_CUDA(cudaEventRecord(w->ev[7], st));
        VGA_CUDA(cudaStreamSynchronize(st2));
        VGA_CUDA(cudaStreamSynchronize(st));
        return true;
    }
    tl_mark("NI known");
    items = w->b[B_ITEMS].get<AlignItem>(NI);
    item_build_kernel<<<(n + 255) / 256, 256, 0, st>>>(item_cnt, item_off, chain_off, n, items);
    c.launches++;
    // ---- K4 subgraphs
    SubBuffers sub;
    memset(&sub, 0, sizeof(sub));
    sub.n_items = NI; sub.items = items; sub.seqs = d_seqs; sub.seq_off = d_seq_off;
    sub.chain_anchor_off = cg.chain_anchor_off; sub.ca_qb = cg.ca_qb; sub.ca_tb = cg.ca_tb; sub.ca_te = cg.ca_te;
    u32* ctr = w->b[B_CTR].get<u32>(16);     // [0] ext ctr, [1] overflow, [2] work ctr, [4..5] pool ctr (u64)
    u32 ext_cap = std::max<u32>(1u << 16, NI * 8);
    for (int attempt = 0;; attempt++) {
        if (w->scratch_words == 0) { w->scratch_words = 1u << 15; w->scratch_zeroed = false; }
        const u32 want_warps = std::min<u32>((u32)w->n_sms * 32, std::max<u32>(NI, 4));
        if (want_warps != w->sub_warps) { w->sub_warps = want_warps; w->scratch_zeroed = false; }
        const u64 sbytes = (u64)w->sub_warps * w->scratch_words * 4;
        void* prevp = w->b[B_SCRATCH].p;
        u32* scratch = (u32*)w->b[B_SCRATCH].ensure(sbytes);
        if (scratch != prevp) w->scratch_zeroed = false;
        if (!w->scratch_zeroed) { VGA_CUDA(cudaMemsetAsync(scratch, 0, sbytes, st)); w->scratch_zeroed = true; }
        sub.ext_pool = w->b[B_EXT].get<u32>(ext_cap); sub.ext_cap = ext_cap;
        sub.ext_ctr = ctr; sub.overflow = ctr + 1;
        sub.scratch = scratch; sub.scratch_per_warp = w->scratch_words; sub.n_warps = w->sub_warps;
        VGA_CUDA(cudaMemsetAsync(ctr, 0, 64, st));
        c.launches += launch_sub_count(dv, sub, st);
    stage_check(st, "sub_count");
        u32 h[2];
        VGA_CUDA(cudaMemcpyAsync(h, ctr, 8, cudaMemcpyDeviceToHost, st));
        VGA_CUDA(cudaStreamSynchronize(st));
        if (!h[1]) break;
        if (attempt >= 12) { set_error("subgraph extension exceeded the scratch limits"); throw CudaFail{VGA_ERR_NOMEM}; }
        if (h[0] > ext_cap) ext_cap = h[0] + h[0] / 4 + 1024;      // the counter kept counting: exact need is known
        else { w->scratch_words *= 2; w->scratch_zeroed = false; }
    }
    tl_mark("sub_count done");
    u64* sz = w->b[B_SZ].get<u64>(4ull * NI);
    u64* szo = w->b[B_SZOFF].get<u64>(4ull * (NI + 1));
    item_sizes_kernel<<<(NI + 255) / 256, 256, 0, st>>>(items, NI, sz, sz + NI, sz + 2ull * NI, sz + 3ull * NI);
    c.launches++;
    u64 tot[4];
    for (int q = 0; q < 4; q++) scan_to_offsets(c, sz + (u64)q * NI, szo + (u64)q * (NI + 1), NI);
    for (int q = 0; q < 4; q++) VGA_CUDA(cudaMemcpyAsync(&tot[q], szo + (u64)q * (NI + 1) + NI, 8, cudaMemcpyDeviceToHost, st));
    VGA_CUDA(cudaStreamSynchronize(st));
    for (int q = 0; q < 4; q++)
        if (tot[q] >= (1ull << 32)) { set_error("alignment batch exceeds 32-bit offsets; use smaller batches"); throw CudaFail{VGA_ERR_UNSUPPORTED}; }
    tl_mark("sizes scanned");
    u32* d_nrows = w->b[B_RCNT].get<u32>(3ull * NI);
    item_bases_kernel<<<(NI + 255) / 256, 256, 0, st>>>(items, NI, szo, szo + (NI + 1), szo + 2ull * (NI + 1), szo + 3ull * (NI + 1),
                                                        cg.chain_anchor_off, dv.k, d_nrows, d_nrows + NI, d_nrows + 2ull * NI);
    c.launches++;
    // rows / read length / size class of every item go to the host now: the classification below runs on the host while
    // sub_fill and sub_rowmeta run on the device
    HArr<u32> h_meta_pin;
    h_meta_pin.alloc(3ull * NI);
    struct PinGuard { HArr<u32>& a; ~PinGuard() { a.release(); } } h_meta_guard{h_meta_pin};
    const u32* h_meta = h_meta_pin.data();
    VGA_CUDA(cudaMemcpyAsync(h_meta_pin.data(), d_nrows, 12ull * NI, cudaMemcpyDeviceToHost, st));
    sub.nd_handle = w->b[B_NDH].get<u32>(tot[0]); sub.nd_row_start = w->b[B_NDRS].get<u32>(tot[0]);
    sub.nd_src = w->b[B_NDSRC].get<u32>(tot[0]); sub.nd_pred_off = w->b[B_NDPO].get<u32>(tot[0]);
    sub.nd_pred = w->b[B_NDPRED].get<u32>(tot[1]); sub.nd_first_succ = w->b[B_NDFS].get<i32>(tot[0]);
    sub.nd_remain_last = w->b[B_NDRL].get<i32>(tot[0]);
    u32* d_row_meta = w->b[B_ROWMETA].get<u32>(tot[2]);
    VGA_CUDA(cudaStreamSynchronize(st));
    tl_mark("meta on host");
    c.launches += launch_sub_fill(dv, sub, st);
    c.launches += launch_sub_rowmeta(dv, sub, d_row_meta, st);
    stage_check(st, "sub_fill");
    VGA_CUDA(cudaEventRecord(w->ev[4], st));
    tl_mark("sub_fill queued");
    // ---- K5 POA in waves
    PoaBuffers pb;
    memset(&pb, 0, sizeof(pb));
    pb.items = items; pb.seqs = d_seqs;
    pb.nd_row_start = sub.nd_row_start; pb.nd_src = sub.nd_src; pb.nd_pred_off = sub.nd_pred_off; pb.nd_pred = sub.nd_pred;
    pb.nd_first_succ = sub.nd_first_succ; pb.nd_remain_last = sub.nd_remain_last;
    pb.nd_state = w->b[B_NDSTATE].get<NodeState>(tot[0]);
    pb.rows = w->b[B_ROWS].get<RowInfo>(tot[2]);
    pb.row_meta = d_row_meta;
    pb.match = P.match; pb.mismatch = P.mismatch; pb.o1 = P.o1; pb.e1 = P.e1; pb.o2 = P.o2; pb.e2 = P.e2; pb.wb = P.wb; pb.wf = P.wf;
    pb.qmax_short = 0; pb.nb_hint = 3;
    pb.bgran_mask = (P.flags & VGA_FLAG_BAND_GRAN16) ? 15 : (P.flags & VGA_FLAG_BAND_GRAN8) ? 7 : 0;
    pb.work_ctr = ctr + 2; pb.pool_ctr = (unsigned long long*)(ctr + 4);
    TraceBuffers tbuf;
    memset(&tbuf, 0, sizeof(tbuf));
    tbuf.items = items; tbuf.seqs = d_seqs;
    tbuf.nd_row_start = sub.nd_row_start; tbuf.nd_src = sub.nd_src; tbuf.nd_pred_off = sub.nd_pred_off; tbuf.nd_pred = sub.nd_pred;
    tbuf.nd_remain_last = sub.nd_remain_last; tbuf.rows = pb.rows;
    tbuf.out_runs = w->b[B_OUTRUNS].get<u32>(tot[3]); tbuf.out_path = w->b[B_OUTPATH].get<u32>(tot[0]);
    tbuf.bgran_mask = (P.flags & VGA_FLAG_BAND_GRAN16) ? 15 : (P.flags & VGA_FLAG_BAND_GRAN8) ? 7 : 0;
    tbuf.match = P.match; tbuf.mismatch = P.mismatch; tbuf.o1 = P.o1; tbuf.e1 = P.e1; tbuf.o2 = P.o2; tbuf.e2 = P.e2; tbuf.wb = P.wb; tbuf.wf = P.wf;
    // estimated trace words per item and its kernel class; inside a class the large items go first
    // (persistent warps pull work in list order, so the long alignments do not end up as a tail)
    poa_begin_batch();
    const int NCLS = poa_class_count();
    std::vector<u64>& est = w->h_est;
    est.resize(NI);
    std::vector<std::vector<u32>>& cls = w->h_cls;     // [NCLS]: chains K4 refused (no subgraph: VGA_ALN_UNSUPPORTED stays), never launched
    cls.resize(NCLS + 1);
    for (auto& l : cls) l.clear();
    u64 est_total = 0;
    u64 sum_rows = 0, sum_q = 0;
    // Host side of the classification, spread over a few threads for big batches (1.2 M items took 11 ms on one thread —
    // longer than the sub_fill kernels it hides behind).  Every thread owns a contiguous item range; per-thread
    // histograms turn into per-thread write positions, so the lists come out exactly as the serial loop built them.
    {
        const int NB = 48;
        const size_t NK = (size_t)(NCLS + 1) * NB;
        u32 NT = NI >= 200000 ? std::min<u32>(8, std::max<u32>(1, std::thread::hardware_concurrency())) : 1;
        if (const char* e = getenv("VGA_CLASSIFY_THREADS")) { const int v = atoi(e); if (v >= 1 && v <= 64) NT = (u32)v; }   // tests
        if (NT > NI) NT = 1;
        auto par = [&](auto&& fn) {
            if (NT == 1) { fn(0u, 0u, NI); return; }
            std::vector<std::thread> th;
            for (u32 t = 0; t < NT; t++)
                th.emplace_back([&fn, t, NT, NI]() { fn(t, (u32)((u64)NI * t / NT), (u32)((u64)NI * (t + 1) / NT)); });
            for (auto& x : th) x.join();
        };
        std::vector<u64> part((size_t)NT * 8, 0);               // per thread: n_long, est_total, sum_rows, sum_q, qmax
        par([&](u32 t, u32 i0, u32 i1) {
            u64 nl = 0;
            for (u32 i = i0; i < i1; i++) nl += poa_short_ok(h_meta[NI + i], h_meta[i], P) ? 0 : 1;
            part[(size_t)t * 8] = nl;
        });
        tl_mark("classify: long count");
        u64 n_long = 0;
        for (u32 t = 0; t < NT; t++) n_long += part[(size_t)t * 8];
        pb.long_mode = poa_long_mode(n_long, w->n_sms);
        std::vector<u32>& key = w->h_key;
        key.resize(NI);
        std::vector<u32> hist((size_t)NT * NK, 0);
        par([&](u32 t, u32 i0, u32 i1) {
            u32* h = hist.data() + (size_t)t * NK;
            u64 et = 0, sr = 0, sq = 0, qm = 0;
            for (u32 i = i0; i < i1; i++) {
                const i64 nrows = h_meta[i], qlen = h_meta[NI + i];
                if ((h_meta[2ull * NI + i] & 0x7fffffffu) == 0) { est[i] = 0; key[i] = (u32)(NCLS * NB); h[key[i]]++; continue; }
                const i64 wband = P.wb + (i64)(P.wf * (float)qlen);
                const i64 west = std::min<i64>(qlen + 1, 2 * wband + 1 + std::llabs(nrows - qlen) + 8);
                // rows are stored from an 8-column boundary (trace_col0); no kernel computes a band wider than 1024 columns
                est[i] = (u64)(nrows * (std::min<i64>(west, 1024) + 12));
                if (qlen < 256) { if ((u64)qlen > qm) qm = (u64)qlen; sr += (u64)nrows; sq += (u64)qlen; }
                int ci = poa_class_for(qlen, nrows, west, P, pb.long_mode);
                if (ci >= 3 && ci < 8 && qlen >= 2048 && (h_meta[2ull * NI + i] >> 31)) {
                    ci = 8;                                                  // sparse chain on a long read: the wide kernel from the start
                    est[i] = std::max<u64>(est[i], (u64)nrows * ((u64)qlen / 8 + 256));
                }
                et += est[i];
                const int lb = 63 - __builtin_clzll(est[i] | 1);            // floor(log2)
                key[i] = (u32)(ci * NB + (NB - 1 - std::min(lb, NB - 1)));
                h[key[i]]++;
            }
            u64* pt = part.data() + (size_t)t * 8;
            pt[1] = et; pt[2] = sr; pt[3] = sq; pt[4] = qm;
        });
        for (u32 t = 0; t < NT; t++) {
            const u64* pt = part.data() + (size_t)t * 8;
            est_total += pt[1]; sum_rows += pt[2]; sum_q += pt[3];
            if ((i64)pt[4] > (i64)pb.qmax_short) pb.qmax_short = (i32)pt[4];
        }
        tl_mark("classify: keys");
        // position of thread t's first item of key k inside its class list: items of smaller keys of the class, then the
        // same key in the threads before t
        std::vector<u32> start((size_t)NT * NK);
        for (int ci = 0; ci <= NCLS; ci++) {
            u32 run = 0;
            for (int kb = 0; kb < NB; kb++)
                for (u32 t = 0; t < NT; t++) {
                    const size_t k = (size_t)ci * NB + kb;
                    start[(size_t)t * NK + k] = run;
                    run += hist[(size_t)t * NK + k];
                }
            cls[ci].resize(run);
        }
        par([&](u32 t, u32 i0, u32 i1) {
            u32* st_t = start.data() + (size_t)t * NK;
            for (u32 i = i0; i < i1; i++) cls[key[i] / NB][st_t[key[i]]++] = i;
        });
    }
    tl_mark("classify: lists");
    // subgraphs much longer than their reads spend most rows at full band width: worth the 3-block step body (more registers)
    pb.nb_hint = sum_rows >= 3 * sum_q ? 3 : 2;
    const u64 warp_slack = (u64)w->n_sms * 64 * 8192;
    u64 want_words = est_total + est_total / 4 + warp_slack;
    // The free-memory query only when the pool has to grow: cudaMemGetInfo takes a driver lock that NVML queries of
    // another thread (a clock sampler, a monitoring agent) also take — measured as sporadic 10-90 ms stalls of this call.
    // A pool that was already cut to what the device can spare keeps that size.
    if (w->pool_capped_words && want_words >= w->pool_capped_words) want_words = w->pool_capped_words;
    else if (want_words * 4 > w->b[B_POOL].cap) {
        size_t free_b = 0, total_b = 0;
        VGA_CUDA(cudaMemGetInfo(&free_b, &total_b));
        const u64 max_words = (u64)((free_b + w->b[B_POOL].cap) * 0.6) / 4;
        if (want_words > max_words) { want_words = max_words; w->pool_capped_words = max_words; }
    }
    if (want_words > (1ull << 34) - (1u << 20)) want_words = (1ull << 34) - (1u << 20);   // RowInfo.toff: u32 in 4-word units
    if (want_words < (1ull << 22)) want_words = 1ull << 22;
    // VGA_POOL_MAX_WORDS caps the pool (tests: forces several waves and the pool-overflow path on small inputs)
    u64 pool_cap_env = ~0ull;
    if (const char* e = getenv("VGA_POOL_MAX_WORDS")) { const long long v = atoll(e); if (v > 0) pool_cap_env = (u64)v; }
    if (want_words > pool_cap_env) want_words = pool_cap_env;
    pb.pool = w->b[B_POOL].get<u32>(want_words);
    pb.pool_words = std::min<u64>(std::min<u64>(w->b[B_POOL].cap / 4, (1ull << 34) - 16), pool_cap_env);
    tbuf.pool = pb.pool;
    if (host_timeline()) fprintf(stderr, "[vga] trace pool: %.2f G words wanted, %.2f G in use (capped at %.2f G by the free memory seen when it last grew)\n",
                                 (double)(est_total + est_total / 4 + warp_slack) / 1e9, (double)pb.pool_words / 1e9, (double)w->pool_capped_words / 1e9);
    tl_mark("pool ready");
    u32* d_work = w->b[B_WORK].get<u32>(NI);
    u32* d_twork = w->b[B_TWORK].get<u32>(NI);
    u32* d_status = w->b[B_STATUS].get<u32>(NI);
    HArr<u32> h_status;                          // pinned: the per-wave status read is on the critical path between fill and traceback
    h_status.alloc(NI);
    PinGuard h_status_guard{h_status};
    float ms_fill = 0, ms_trace = 0;
    // Fill launches of all classes share the trace pool; the traceback of everything filled so far runs as ONE
    // launch when the pool is about to run out (or at the end), after which the pool starts over.
    std::vector<u32>& trace_list = w->h_trace;
    trace_list.clear();
    trace_list.reserve(NI);
    u64 pool_acc = warp_slack;                  // estimated words in use since the last traceback
    u64 used_prev = 0;                          // pool counter after the previous wave (timeline print only)
    VGA_CUDA(cudaMemsetAsync(ctr + 4, 0, 8, st));
    std::function<void()> collect_early_fn;      // set below; a traceback that empties the pool takes the early wave's rows with it
    auto flush_trace = [&]() {
        if (collect_early_fn) collect_early_fn();
        if (!trace_list.empty()) {
            VGA_CUDA(cudaMemcpyAsync(d_twork, trace_list.data(), 4ull * trace_list.size(), cudaMemcpyHostToDevice, st));
            tbuf.n_items = (u32)trace_list.size(); tbuf.work = d_twork;
            VGA_CUDA(cudaEventRecord(w->ev[9], st));
            c.launches += launch_poa_trace(dv, tbuf, st);
            stage_check(st, "poa_trace");
            VGA_CUDA(cudaEventRecord(w->ev[10], st));
            VGA_CUDA(cudaEventSynchronize(w->ev[10]));
            float t = 0;
            cudaEventElapsedTime(&t, w->ev[9], w->ev[10]);
            ms_trace += t;
            if (host_timeline()) fprintf(stderr, "[vga] poa trace: %zu items, %.3f ms\n", trace_list.size(), t);
            trace_list.clear();
        }
        VGA_CUDA(cudaMemsetAsync(ctr + 4, 0, 8, st));
        pool_acc = warp_slack;
        used_prev = 0;
    };
    // a traceback launched right behind its fill wave (below) is timed when the stream has passed it
    bool trace_time_pending = false;
    auto collect_trace_time = [&]() {
        if (!trace_time_pending) return;
        VGA_CUDA(cudaEventSynchronize(w->ev_e1));
        float t = 0;
        cudaEventElapsedTime(&t, w->ev_e0, w->ev_e1);
        ms_trace += t;
        if (host_timeline()) fprintf(stderr, "[vga] poa trace (behind the fill): %.3f ms\n", t);
        trace_time_pending = false;
    };
    // what a wave's statuses mean for the lists: retries move on (a wider class, or the same class again after a traceback
    // has emptied the pool), completed alignments wait for their traceback
    auto absorb = [&](int ci, const u32* wave, u32 nw, const u32* hs, bool traced, std::vector<u32>& again, u32& done) {
        for (u32 x = 0; x < nw; x++) {
            const u32 it = wave[x];
            if (hs[x] == IT_RETRY_WIDER) {
                const int nc = poa_class_wider(ci, pb.long_mode);
                if (nc == 8) {
                    // the band left the 1024 columns of the warp kernels: nothing bounds it but the read.  Measured on C5's
                    // reverse-strand reads (random k-mer hits chained): 0.07-0.1 of rows x read length (0.19 of rows x (read / 4 + 1024),
                    // which cut the 1.1 k such alignments of a 12.5 k-read call into three waves)
                    const u64 nr = h_meta[it], ql = h_meta[NI + it];
                    est[it] = std::max<u64>(est[it], nr * (ql / 8 + 256));
                }
                if (nc >= 0) cls[nc].push_back(it); else done++;
            }
            else if (hs[x] == IT_RETRY_POOL) again.push_back(it);
            else { done++; if (hs[x] == VGA_ALN_OK && !traced) trace_list.push_back(it); }
        }
    };
    // Early wave.  A short-read batch usually holds a few alignments outside the short-read classes (17-DRB1-3123-smooth: 1.5 k
    // subgraphs of thousands of rows per 125 k reads).  One warp each, they are two or three warps per SM and run for as long as
    // their longest row sequence takes (74 ms) — after the 24 ms of the 16x2 kernel that is a tail of idle SMs.  When both fit the
    // pool together, the long class is launched FIRST on its own stream, with its own work list and counter, and the 16x2 kernel's
    // CTAs fill what the few long-class CTAs leave of every SM.  Its statuses are absorbed when the class loop reaches it, or
    // before any traceback empties the pool.
    int early_ci = -1;
    u32 early_n = 0;
    HArr<u32> h_status2;
    PinGuard h_status2_guard{h_status2};
    float ms_early_span = 0;
    bool early_timed = false, st_fill_after_fork = false;
    auto collect_early = [&]() {
        if (early_ci < 0 || early_n == 0) return;
        VGA_CUDA(cudaEventSynchronize(w->ev_e1));
        cudaEventElapsedTime(&ms_early_span, w->ev_e0, w->ev_e1);
        early_timed = true;
        // the POA stage lasts until both streams are done: what the early wave ran beyond the fill waves of the main stream
        float t_st = 0;
        if (st_fill_after_fork) { VGA_CUDA(cudaEventSynchronize(w->ev_stf)); cudaEventElapsedTime(&t_st, w->ev_e0, w->ev_stf); }
        if (ms_early_span > t_st) ms_fill += ms_early_span - t_st;
        std::vector<u32> again;
        u32 done = 0;
        std::vector<u32>& el = cls[early_ci];
        absorb(early_ci, el.data(), early_n, h_status2.data(), false, again, done);
        if (host_timeline()) fprintf(stderr, "[vga] poa fill class %d (early wave, beside the short-read classes): %u items, done %.3f ms after the fork\n",
                                     early_ci, early_n, ms_early_span);
        // the wave's items leave the list; pool retries come back at its end with a doubled estimate
        el.erase(el.begin(), el.begin() + early_n);
        for (u32 it : again) est[it] *= 2;
        el.insert(el.end(), again.begin(), again.end());
        early_n = 0;
    };
    collect_early_fn = collect_early;
    {
        const char* ee = getenv("VGA_POA_EARLY");
        const bool on = ee ? atoi(ee) != 0 : true;
        int cj = -1;
        for (int x = 2; x < NCLS && cj < 0; x++) if (!cls[x].empty()) cj = x;
        const size_t n_short = cls[0].size() + cls[1].size();
        if (on && cj >= 3 && cj < 8 && n_short > 0 && cls[cj].size() <= (size_t)w->n_sms * 32) {
            u64 need = warp_slack;
            for (int x = 0; x < NCLS; x++)
                for (u32 it : cls[x]) need += est[it] + est[it] / 4 + 64;
            if (need <= pb.pool_words) {
                early_ci = cj;
                early_n = (u32)cls[cj].size();
                u32* d_work2 = w->b[B_WORK2].get<u32>(early_n);
                u32* d_status2 = w->b[B_STATUS2].get<u32>(early_n);
                h_status2.alloc(early_n);
                PoaBuffers pb2 = pb;
                pb2.n_items = early_n; pb2.work = d_work2; pb2.work_ctr = ctr + 3;
                VGA_CUDA(cudaMemcpyAsync(d_work2, cls[cj].data(), 4ull * early_n, cudaMemcpyHostToDevice, st));
                VGA_CUDA(cudaMemsetAsync(ctr + 3, 0, 4, st));
                VGA_CUDA(cudaEventRecord(w->ev_e0, st));
                VGA_CUDA(cudaStreamWaitEvent(w->st3, w->ev_e0, 0));
                c.launches += launch_poa_fill(dv, pb2, cj, w->n_sms, w->st3);
                stage_check(w->st3, "poa_fill (early wave)");
                R.stats.poa_launches++;
                item_status_kernel<<<(early_n + 255) / 256, 256, 0, w->st3>>>(items, d_work2, early_n, d_status2);
                c.launches++;
                VGA_CUDA(cudaMemcpyAsync(h_status2.data(), d_status2, 4ull * early_n, cudaMemcpyDeviceToHost, w->st3));
                VGA_CUDA(cudaEventRecord(w->ev_e1, w->st3));
                for (u32 it : cls[cj]) pool_acc += est[it] + est[it] / 4 + 64;
            }
        }
    }
    for (int ci = 0; ci < NCLS; ci++) {
        if (ci == early_ci) collect_early();
        std::vector<u32>& list = cls[ci];
        size_t pos = 0;
        int stall = 0;
        // Waves are admitted on the a-priori estimate (widest band on every row).  Long-read classes: measured pool use is
        // 0.43-0.52 of that estimate wave after wave (C5), so they are admitted at 0.62 of it — fuller waves, fewer
        // tracebacks; an overflow is handled like any other (IT_RETRY_POOL: trace, start the pool over, rerun).  Tried and
        // dropped: learning the factor from the previous wave (overflows and reruns cost more than fuller waves gain) and
        // accounting the real use instead of the estimate (leaves room for tiny follow-up waves that cannot fill the machine).
        // Short-read classes: 0.34-0.43 of the estimate on C2 / C3 / C4 (the estimate assumes the widest band on every row, most
        // rows are narrower), admitted at 0.55: a 1 M-read call of C4 needs 2 waves instead of 4.
        const double scale = ci >= 3 ? 0.62 : 0.55;
        auto cost = [&](u32 it) { return (u64)((double)(est[it] + est[it] / 4) * scale) + 64; };
        // Long reads on the shared-memory kernel: ONE fused launch first — every warp fills an alignment into its private arena of
        // the pool and traces it right away (k_poa_long.cu).  No waves, no traceback launches of a few thousand threads, no host
        // round trips in between; alignments larger than an arena come back as pool retries and take the wave path below.
        {
            const char* fe = getenv("VGA_POA_FUSED");
            if (ci == 3 && pb.long_mode && !list.empty() && (fe ? atoi(fe) != 0 : true)) {
                flush_trace();                                   // the arenas take the whole pool: nothing may wait in it
                const u32 nw = (u32)list.size();
                VGA_CUDA(cudaMemcpyAsync(d_work, list.data(), 4ull * nw, cudaMemcpyHostToDevice, st));
                VGA_CUDA(cudaMemsetAsync(ctr + 2, 0, 4, st));
                pb.n_items = nw; pb.work = d_work;
                VGA_CUDA(cudaEventRecord(w->ev[8], st));
                c.launches += launch_poa_fill(dv, pb, ci, w->n_sms, st, &tbuf);
                stage_check(st, "poa_fill (fused long reads)");
                R.stats.poa_launches++;
                VGA_CUDA(cudaEventRecord(w->ev[9], st));
                item_status_kernel<<<(nw + 255) / 256, 256, 0, st>>>(items, d_work, nw, d_status);
                c.launches++;
                queue_chain_d2h();
                VGA_CUDA(cudaMemcpyAsync(h_status.data(), d_status, 4ull * nw, cudaMemcpyDeviceToHost, st));
                VGA_CUDA(cudaStreamSynchronize(st));
                float a = 0;
                cudaEventElapsedTime(&a, w->ev[8], w->ev[9]);
                ms_fill += a;
                std::vector<u32> again;
                u32 done = 0;
                absorb(ci, list.data(), nw, h_status.data(), true, again, done);
                if (host_timeline()) fprintf(stderr, "[vga] poa fill+trace class 3 (fused, per-warp arenas): %u items, %.3f ms, %zu left to the wave path\n", nw, a, again.size());
                list.swap(again);
                VGA_CUDA(cudaMemsetAsync(ctr + 4, 0, 8, st));
                pool_acc = warp_slack;
                used_prev = 0;
            }
        }
        while (pos < list.size()) {
            // take a wave whose estimate fits what is left of the pool
            const u64 first = cost(list[pos]);
            if (pool_acc + first > pb.pool_words && pool_acc > warp_slack) flush_trace();
            const bool fresh_pool = pool_acc == warp_slack;
            u64 acc = pool_acc, wave_est = 0;
            size_t end = pos;
            while (end < list.size() && (end == pos || acc + cost(list[end]) <= pb.pool_words)) {
                acc += cost(list[end]);
                wave_est += est[list[end]] + est[list[end]] / 4;
                end++;
            }
            const u32 nw = (u32)(end - pos);
            if (ci == 8) {
                pb.qmax_wide = 0;
                for (size_t x = pos; x < end; x++) pb.qmax_wide = std::max<i32>(pb.qmax_wide, (i32)h_meta[NI + list[x]]);
            }
            VGA_CUDA(cudaMemcpyAsync(d_work, list.data() + pos, 4ull * nw, cudaMemcpyHostToDevice, st));
            VGA_CUDA(cudaMemsetAsync(ctr + 2, 0, 4, st));
            pb.n_items = nw; pb.work = d_work;
            VGA_CUDA(cudaEventRecord(w->ev[8], st));
            c.launches += launch_poa_fill(dv, pb, ci, w->n_sms, st);
            stage_check(st, "poa_fill");
            R.stats.poa_launches++;
            VGA_CUDA(cudaEventRecord(w->ev[9], st));
            if (early_n) { VGA_CUDA(cudaEventRecord(w->ev_stf, st)); st_fill_after_fork = true; }
            // A wave that starts on an empty pool (nothing else waits for a traceback) is traced right behind its fill, from the
            // same work list, without the host round trip in between: the traceback skips every item whose status is not OK
            // (items a retry brings back run and are traced again later), and the pool starts over for the next wave — the host
            // reads the statuses while the traceback runs.
            const bool traced_here = trace_list.empty() && fresh_pool && early_n == 0;
            item_status_kernel<<<(nw + 255) / 256, 256, 0, st>>>(items, d_work, nw, d_status);
            c.launches++;
            queue_chain_d2h();          // host-side queueing (and any pinned allocation) hides behind the fill kernel
            VGA_CUDA(cudaMemcpyAsync(h_status.data(), d_status, 4ull * nw, cudaMemcpyDeviceToHost, st));
            unsigned long long used_now = 0;
            VGA_CUDA(cudaMemcpyAsync(&used_now, ctr + 4, 8, cudaMemcpyDeviceToHost, st));
            VGA_CUDA(cudaEventRecord(w->ev_fork, st));            // fill + statuses of this wave
            if (traced_here) {
                collect_trace_time();                             // the previous wave's traceback ended before this fill began
                tbuf.n_items = nw; tbuf.work = d_work;
                VGA_CUDA(cudaEventRecord(w->ev_e0, st));
                c.launches += launch_poa_trace(dv, tbuf, st);
                stage_check(st, "poa_trace");
                VGA_CUDA(cudaEventRecord(w->ev_e1, st));
                trace_time_pending = true;
            }
            VGA_CUDA(cudaEventSynchronize(w->ev_fork));
            float a = 0;
            cudaEventElapsedTime(&a, w->ev[8], w->ev[9]);
            ms_fill += a;
            if (host_timeline()) fprintf(stderr, "[vga] poa fill class %d: %u items, %.3f ms (pool: %.2f of the estimate)\n", ci, nw, a,
                                         wave_est && used_now >= used_prev ? (double)(used_now - used_prev) / (double)wave_est : 0.0);
            used_prev = used_now;
            pool_acc = acc;
            if (traced_here) {                       // everything complete in this wave has its traceback: the pool is free again
                VGA_CUDA(cudaMemsetAsync(ctr + 4, 0, 8, st));
                pool_acc = warp_slack;
                used_prev = 0;
            }
            std::vector<u32> again;
            u32 done = 0;
            absorb(ci, list.data() + pos, nw, h_status.data(), traced_here, again, done);
            pos = end;
            if (!again.empty()) {
                // the pool ran out: trace what is complete, start the pool over and run the rest again
                flush_trace();
                if (fresh_pool && done == 0 && again.size() == nw) {
                    // nothing fitted: the pool is too small for even one alignment of this wave
                    if (++stall > 2) { set_error("POA trace pool too small for a single alignment"); throw CudaFail{VGA_ERR_NOMEM}; }
                }
                for (u32 it : again) est[it] *= 2;
                list.insert(list.end(), again.begin(), again.end());
            }
        }
    }
    collect_early();
    collect_trace_time();
    flush_trace();
    VGA_CUDA(cudaEventRecord(w->ev[5], st));
    tl_mark("poa waves done");
    // ---- selection + compaction
    i32* sel = w->b[B_SEL].get<i32>(n);
    u8* d_st8 = (u8*)w->b[B_STATUS].ensure(std::max<size_t>(4ull * NI, n));
    u32* d_sc = w->b[B_SCAL].get<u32>(6ull * n);
    u64* d_cells = w->b[B_CELLS].get<u64>(n);
    u64* cnts = w->b[B_PCNT].get<u64>(4ull * n);
    u64* offs = w->b[B_POFF2].get<u64>(4ull * (n + 1));
    select_kernel<<<(n + 255) / 256, 256, 0, st>>>(items, item_cnt, item_off, n, (u32)VGA_ALN_TOO_WIDE, sel, d_st8, d_sc, d_cells, cnts,
                                                   cnts + n, dbg ? cnts + 2ull * n : nullptr, cnts + 3ull * n);
    c.launches++;
    const int nscan = dbg ? 4 : 2;
    u64 tots[4] = {0, 0, 0, 0};
    for (int q = 0; q < nscan; q++) scan_to_offsets(c, cnts + (u64)q * n, offs + (u64)q * (n + 1), n);
    for (int q = 0; q < nscan; q++) VGA_CUDA(cudaMemcpyAsync(&tots[q], offs + (u64)q * (n + 1) + n, 8, cudaMemcpyDeviceToHost, st));
    VGA_CUDA(cudaStreamSynchronize(st));
    u32* d_runs = w->b[B_RUNS].get<u32>(tots[0]);
    u64* d_path = w->b[B_PATH].get<u64>(tots[1]);
    gather_kernel<<<std::min<u32>((n + 7) / 8, 148 * 16), 256, 0, st>>>(items, sel, n, offs, offs + (n + 1), tbuf.out_runs,
                                                                         tbuf.out_path, sub.nd_handle, d_runs, d_path);
    c.launches++;
    VGA_CUDA(cudaEventRecord(w->ev[6], st));
    // ---- D2H of the alignment results
    d2h_into(O.aln_status, (const u8*)d_st8, n, st);
    d2h_into(O.aln_cells, (const u64*)d_cells, n, st);
    d2h_into(O.aln_chain, (const u32*)d_sc, n, st); d2h_into(O.aln_path_len, (const u32*)(d_sc + n), n, st);
    d2h_into(O.aln_start_off, (const u32*)(d_sc + 2ull * n), n, st); d2h_into(O.aln_end_off, (const u32*)(d_sc + 3ull * n), n, st);
    d2h_into(O.aln_n_aligned, (const u32*)(d_sc + 4ull * n), n, st); d2h_into(O.aln_score, (const i32*)(d_sc + 5ull * n), n, st);
    d2h_into(O.aln_run_off, (const u64*)offs, (size_t)n + 1, st); d2h_into(O.aln_path_off, (const u64*)(offs + (n + 1)), (size_t)n + 1, st);
    d2h_into(O.aln_runs, (const u32*)d_runs, tots[0], st);
    d2h_into(O.aln_path, (const u64*)d_path, tots[1], st);
    if (dbg) {
        u64* d_sub = w->b[B_DBGSUB].get<u64>(tots[2]);
        u32* d_edges = w->b[B_DBGEDGE].get<u32>(2 * tots[3]);
        dbg_sub_kernel<<<std::min<u32>((n + 7) / 8, 148 * 16), 256, 0, st>>>(items, sel, n, offs + 2ull * (n + 1), offs + 3ull * (n + 1),
                                                                              sub.nd_handle, sub.nd_pred_off, sub.nd_pred, d_sub, d_edges);
        c.launches++;
        d2h_into(O.dbg_sub_off, (const u64*)(offs + 2ull * (n + 1)), (size_t)n + 1, st);
        d2h_into(O.dbg_edge_off, (const u64*)(offs + 3ull * (n + 1)), (size_t)n + 1, st);
        d2h_into(O.dbg_sub_handles, (const u64*)d_sub, tots[2], st);
        d2h_into(O.dbg_edges, (const u32*)d_edges, 2 * tots[3], st);
    }
    tl_mark("aln d2h queued");
    // total cells over all items (GCUPS numerator): reduced on the device, 8 bytes come back
    unsigned long long h_cells_total = 0;
    {
        unsigned long long* d_total = (unsigned long long*)(ctr + 6);
        VGA_CUDA(cudaMemsetAsync(d_total, 0, 8, st));
        cells_sum_kernel<<<(NI + 255) / 256, 256, 0, st>>>(items, NI, d_total);
        c.launches++;
        VGA_CUDA(cudaMemcpyAsync(&h_cells_total, d_total, 8, cudaMemcpyDeviceToHost, st));
    }
    VGA_CUDA(cudaEventRecord(w->ev[7], st));
    VGA_CUDA(cudaStreamSynchronize(st2));
    VGA_CUDA(cudaStreamSynchronize(st));
    tl_mark("chunk done");
    R.stats.poa_cells += h_cells_total;
    R.stats.n_alignments += NI;
    R.stats.ms_poa += ms_fill; R.stats.ms_trace += ms_trace;
    return true;
}

static void add_stage_times(Ctx& c) {
    vga_workspace* w = c.ws;
    float t;
    VgaStats& s = c.R->stats;
    if (cudaEventElapsedTime(&t, w->ev[0], w->ev[1]) == cudaSuccess) s.ms_seed += t;
    if (cudaEventElapsedTime(&t, w->ev[1], w->ev[2]) == cudaSuccess) s.ms_sort += t;
    if (cudaEventElapsedTime(&t, w->ev[2], w->ev[3]) == cudaSuccess) s.ms_chain += t;
    if (cudaEventElapsedTime(&t, w->ev[3], w->ev[4]) == cudaSuccess) s.ms_subgraph += t;
    if (cudaEventElapsedTime(&t, w->ev[6], w->ev[7]) == cudaSuccess) s.ms_d2h += t;
    cudaGetLastError();
}

// chunk driver: reads [b,e) whose bytes are resident at d_seqs (offsets absolute into d_seqs)
static void run_range(Ctx& c, const char* d_seqs, const u64* d_seq_off, const u64* h_seq_off, u32 b, u32 e) {
    if (b >= e) return;
    const u64 bases = h_seq_off[e] - h_seq_off[b];
    const u64 MAX_BASES = 320ull << 20;
    const u32 MAX_READS = 1u << 21;
    bool split = (e - b > 1) && (bases > MAX_BASES || e - b > MAX_READS);
    if (!split) {
        OutSet* O = new OutSet();
        c.R->parts.push_back(O);
        if (process_chunk(c, d_seqs, d_seq_off + b, e - b, bases, *O)) { add_stage_times(c); return; }
        c.R->parts.pop_back();
        O->release_all();
        delete O;
        if (e - b == 1) { set_error("a single read produced more than 2^31 anchors"); throw CudaFail{VGA_ERR_UNSUPPORTED}; }
    }
    const u32 mid = b + (e - b) / 2;
    run_range(c, d_seqs, d_seq_off, h_seq_off, b, mid);
    run_range(c, d_seqs, d_seq_off, h_seq_off, mid, e);
}

// concatenate per-chunk outputs (only when a call needed more than one chunk)
template <class T> static void cat_plain(HArr<T>& dst, std::vector<OutSet*>& parts, HArr<T> OutSet::*f) {
    size_t tot = 0;
    for (OutSet* p : parts) tot += (p->*f).n;
    dst.alloc(tot);
    size_t o = 0;
    for (OutSet* p : parts) { memcpy(dst.data() + o, (p->*f).data(), (p->*f).n * sizeof(T)); o += (p->*f).n; }
}
// offsets arrays: each part has count+1 entries starting at 0; rebase by the running total of `payload`
static void cat_offsets(HArr<u64>& dst, std::vector<OutSet*>& parts, HArr<u64> OutSet::*f) {
    size_t tot = 1;
    for (OutSet* p : parts) tot += (p->*f).n ? (p->*f).n - 1 : 0;
    dst.alloc(tot);
    dst[0] = 0;
    size_t o = 1;
    u64 base = 0;
    for (OutSet* p : parts) {
        const HArr<u64>& a = p->*f;
        if (a.n == 0) continue;
        for (size_t i = 1; i < a.n; i++) dst[o++] = base + a.data()[i];
        base += a.data()[a.n - 1];
    }
}

static void merge_parts(vga_results* R) {
    if (R->parts.size() == 1) {
        R->o = *R->parts[0];           // move the blocks (shallow copy, then forget the part)
        delete R->parts[0];
        R->parts.clear();
        return;
    }
    OutSet& o = R->o;
    auto& ps = R->parts;
    for (OutSet* p : ps) { o.n += p->n; o.NC += p->NC; o.NM += p->NM; }
    cat_plain(o.n_anchors, ps, &OutSet::n_anchors); cat_offsets(o.chain_off, ps, &OutSet::chain_off);
    cat_offsets(o.chain_anchor_off, ps, &OutSet::chain_anchor_off);
    cat_plain(o.ca_qb, ps, &OutSet::ca_qb); cat_plain(o.ca_tb, ps, &OutSet::ca_tb); cat_plain(o.ca_te, ps, &OutSet::ca_te);
    if (R->scores) { cat_plain(o.ca_id, ps, &OutSet::ca_id); cat_plain(o.ca_score, ps, &OutSet::ca_score); }
    cat_plain(o.aln_status, ps, &OutSet::aln_status); cat_plain(o.aln_chain, ps, &OutSet::aln_chain);
    cat_plain(o.aln_path_len, ps, &OutSet::aln_path_len); cat_plain(o.aln_start_off, ps, &OutSet::aln_start_off);
    cat_plain(o.aln_end_off, ps, &OutSet::aln_end_off); cat_plain(o.aln_n_aligned, ps, &OutSet::aln_n_aligned);
    cat_plain(o.aln_score, ps, &OutSet::aln_score); cat_plain(o.aln_cells, ps, &OutSet::aln_cells);
    cat_offsets(o.aln_path_off, ps, &OutSet::aln_path_off); cat_offsets(o.aln_run_off, ps, &OutSet::aln_run_off);
    cat_plain(o.aln_path, ps, &OutSet::aln_path); cat_plain(o.aln_runs, ps, &OutSet::aln_runs);
    if (R->debug) {
        cat_offsets(o.dbg_anchor_off, ps, &OutSet::dbg_anchor_off); cat_plain(o.dbg_qb, ps, &OutSet::dbg_qb);
        cat_plain(o.dbg_tb, ps, &OutSet::dbg_tb); cat_plain(o.dbg_te, ps, &OutSet::dbg_te);
        cat_plain(o.dbg_sorted, ps, &OutSet::dbg_sorted); cat_plain(o.dbg_f, ps, &OutSet::dbg_f);
        cat_plain(o.dbg_pred, ps, &OutSet::dbg_pred); cat_offsets(o.dbg_sub_off, ps, &OutSet::dbg_sub_off);
        cat_plain(o.dbg_sub_handles, ps, &OutSet::dbg_sub_handles); cat_offsets(o.dbg_edge_off, ps, &OutSet::dbg_edge_off);
        cat_plain(o.dbg_edges, ps, &OutSet::dbg_edges);
    }
    for (OutSet* p : ps) { p->release_all(); delete p; }
    ps.clear();
}

static int map_common(vga_index_t* ix, const char* seqs, const u64* seq_off, const char* d_seqs_in, const u64* d_off_in,
                      u32 n_reads, u64 n_bases_in, const VgaParams* params, vga_results_t** out) {
    if (!ix || !out || !params || (n_reads && (!seq_off && !d_off_in))) { set_error("vga_map_batch: null argument"); return VGA_ERR_INVALID; }
    *out = nullptr;
    if (params->bandwidth == 0 || params->bandwidth > 96) { set_error("chaining bandwidth must be in 1..96"); return VGA_ERR_UNSUPPORTED; }
    if (params->best_n == 0 || params->min_anchors == 0) { set_error("best_n and min_anchors must be >= 1"); return VGA_ERR_INVALID; }
    if (params->max_gap > (1u << 24)) { set_error("max_gap too large"); return VGA_ERR_UNSUPPORTED; }
    if (((params->flags & VGA_FLAG_ALL_ORIENTATIONS) != 0) != (ix->hdr.all_orient != 0)) {
        set_error("VGA_FLAG_ALL_ORIENTATIONS needs an index uploaded with VGA_INDEX_ALL_ORIENTATIONS (and such an index serves nothing else)");
        return VGA_ERR_INVALID;
    }
    {   // the 32-bit trace word keeps E'/F relative to H in 3 + 5 bit fields (k_poa.cu): bounds on the gap model
        const VgaParams& q = *params;
        if (q.match < 0 || q.mismatch < 0 || q.e1 < 1 || q.e2 < 1 || q.o1 < 0 || q.o2 < 0 || q.o1 > 6 || q.o2 > 30 ||
            q.e2 > q.o1 + q.e1 || q.e1 > q.o2 + q.e2 || q.match > 64 || q.mismatch > 64 || q.e1 > 64 || q.e2 > 64) {
            set_error("POA gap model outside the supported range (e1, e2 >= 1, o1 <= 6, o2 <= 30, e2 <= o1+e1, e1 <= o2+e2)");
            return VGA_ERR_UNSUPPORTED;
        }
    }
    vga_results* R = new vga_results();
    std::lock_guard<std::mutex> index_lock(ix->mu);      // calls on the same index are serialised (vga.h)
    try {
        VGA_CUDA(cudaSetDevice(ix->device));
        Ctx c;
        c.ix = ix; c.ws = workspace(ix); c.st = c.ws->st; c.P = params; c.R = R;
        R->n = n_reads;
        R->debug = params->flags & VGA_FLAG_KEEP_DEBUG;
        R->scores = R->debug || (params->flags & VGA_FLAG_CHAIN_SCORES);
        memset(&R->stats, 0, sizeof(R->stats));
        cudaEvent_t t0 = c.ws->ev[11];
        g_d2h_bytes = 0;
        std::vector<u64>& h_off = c.ws->h_off;
        const char* d_seqs = d_seqs_in;
        const u64* d_off = d_off_in;
        VGA_CUDA(cudaEventRecord(t0, c.st));
        auto wall0 = std::chrono::steady_clock::now();
        g_tl0 = wall0;
        if (n_reads > 0) {
            if (!d_seqs_in) {
                for (u32 i = 0; i < n_reads; i++)
                    if (seq_off[i + 1] < seq_off[i] || seq_off[i + 1] - seq_off[i] >= (1ull << 32)) {
                        set_error("seq_off not monotone (or a read of 2^32 bases or more)"); delete R; return VGA_ERR_INVALID;
                    }
                const u64 nb = seq_off[n_reads] - seq_off[0];
                char* ds = c.ws->b[B_SEQS].get<char>(nb + 16);
                u64* doff = c.ws->b[B_SEQOFF].get<u64>(n_reads + 1);
                h_off.resize(n_reads + 1);
                for (u32 i = 0; i <= n_reads; i++) h_off[i] = seq_off[i] - seq_off[0];
                VGA_CUDA(cudaMemcpyAsync(ds, seqs + seq_off[0], nb, cudaMemcpyHostToDevice, c.st));
                VGA_CUDA(cudaMemcpyAsync(doff, h_off.data(), (n_reads + 1) * 8, cudaMemcpyHostToDevice, c.st));
                d_seqs = ds; d_off = doff;
                R->stats.h2d_bytes += nb + (n_reads + 1) * 8ull;
            } else {
                h_off.resize(n_reads + 1);
                VGA_CUDA(cudaMemcpyAsync(h_off.data(), d_off_in, (n_reads + 1) * 8, cudaMemcpyDeviceToHost, c.st));
                VGA_CUDA(cudaStreamSynchronize(c.st));
                for (u32 i = 0; i < n_reads; i++)
                    if (h_off[i + 1] < h_off[i] || h_off[i + 1] - h_off[i] >= (1ull << 32)) {
                        set_error("dev_seq_off not monotone (or a read of 2^32 bases or more)"); delete R; return VGA_ERR_INVALID;
                    }
                if (h_off[n_reads] > n_bases_in) { set_error("dev_seq_off runs past n_bases"); delete R; return VGA_ERR_INVALID; }
            }
            VGA_CUDA(cudaEventRecord(c.ws->ev[0], c.st));
            float th2d = 0;
            VGA_CUDA(cudaStreamSynchronize(c.st));
            cudaEventElapsedTime(&th2d, t0, c.ws->ev[0]);
            R->stats.ms_h2d = th2d;
            run_range(c, d_seqs, d_off, h_off.data(), 0, n_reads);
            merge_parts(R);
        } else {
            R->o.chain_off.alloc(1); R->o.chain_off[0] = 0; R->o.chain_anchor_off.alloc(1); R->o.chain_anchor_off[0] = 0;
            R->o.aln_path_off.alloc(1); R->o.aln_path_off[0] = 0; R->o.aln_run_off.alloc(1); R->o.aln_run_off[0] = 0;
            if (R->debug) {
                R->o.dbg_anchor_off.alloc(1); R->o.dbg_anchor_off[0] = 0; R->o.dbg_sub_off.alloc(1); R->o.dbg_sub_off[0] = 0;
                R->o.dbg_edge_off.alloc(1); R->o.dbg_edge_off[0] = 0;
            }
        }
        VGA_CUDA(cudaEventRecord(c.ws->ev[9], c.st));
        VGA_CUDA(cudaStreamSynchronize(c.st));
        VGA_CUDA(cudaEventElapsedTime(&R->stats.ms_total, t0, c.ws->ev[9]));
        R->stats.ms_wall = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - wall0).count();
        tl_mark("call done");
        tl_flush(R->stats.ms_wall);
        R->stats.d2h_bytes = g_d2h_bytes;
        R->stats.n_reads = n_reads;
        R->stats.kernel_launches = c.launches;
    } catch (CudaFail& f) {
        delete R;
        return f.code;
    } catch (std::bad_alloc&) {
        delete R;
        set_error("host allocation failed");
        return VGA_ERR_NOMEM;
    }
    *out = R;
    return VGA_OK;
}

extern "C" {

int vga_map_batch(vga_index_t* ix, const char* seqs, const uint64_t* seq_off, uint32_t n_reads, const VgaParams* params,
                  vga_results_t** out) {
    if (n_reads && !seqs) { set_error("vga_map_batch: seqs is null"); return VGA_ERR_INVALID; }
    return map_common(ix, seqs, seq_off, nullptr, nullptr, n_reads, 0, params, out);
}

int vga_map_batch_device(vga_index_t* ix, const char* dev_seqs, const uint64_t* dev_seq_off, uint32_t n_reads, uint64_t n_bases,
                         const VgaParams* params, vga_results_t** out) {
    if (n_reads && (!dev_seqs || !dev_seq_off)) { set_error("vga_map_batch_device: null device pointer"); return VGA_ERR_INVALID; }
    return map_common(ix, nullptr, nullptr, dev_seqs, dev_seq_off, n_reads, n_bases, params, out);
}

int vga_results_view(const vga_results_t* r, VgaResultsView* v) {
    if (!r || !v) { set_error("null argument"); return VGA_ERR_INVALID; }
    memset(v, 0, sizeof(*v));
    const OutSet& o = r->o;
    v->n_reads = r->n;
    v->n_anchors = o.n_anchors.data(); v->chain_off = o.chain_off.data();
    v->n_chains = o.chain_anchor_off.n ? o.chain_anchor_off.n - 1 : 0; v->chain_anchor_off = o.chain_anchor_off.data();
    v->n_chain_anchors = o.ca_qb.n;
    v->ca_qb = o.ca_qb.data(); v->ca_tb = o.ca_tb.data(); v->ca_te = o.ca_te.data();
    if (r->scores) { v->ca_id = o.ca_id.data(); v->ca_score = o.ca_score.data(); }
    v->aln_status = o.aln_status.data(); v->aln_chain = o.aln_chain.data(); v->aln_path_off = o.aln_path_off.data();
    v->aln_path = o.aln_path.data(); v->aln_run_off = o.aln_run_off.data(); v->aln_runs = o.aln_runs.data();
    v->aln_path_len = o.aln_path_len.data(); v->aln_start_off = o.aln_start_off.data(); v->aln_end_off = o.aln_end_off.data();
    v->aln_n_aligned = o.aln_n_aligned.data(); v->aln_score = o.aln_score.data(); v->aln_cells = o.aln_cells.data();
    if (r->debug) {
        v->dbg_anchor_off = o.dbg_anchor_off.data(); v->dbg_qb = o.dbg_qb.data(); v->dbg_tb = o.dbg_tb.data();
        v->dbg_te = o.dbg_te.data(); v->dbg_sorted = o.dbg_sorted.data(); v->dbg_f = o.dbg_f.data();
        v->dbg_pred = o.dbg_pred.data(); v->dbg_sub_off = o.dbg_sub_off.data(); v->dbg_sub_handles = o.dbg_sub_handles.data();
        v->dbg_edge_off = o.dbg_edge_off.data(); v->dbg_edges = o.dbg_edges.data();
    }
    return VGA_OK;
}

int vga_results_stats(const vga_results_t* r, VgaStats* s) {
    if (!r || !s) { set_error("null argument"); return VGA_ERR_INVALID; }
    *s = r->stats;
    return VGA_OK;
}

int vga_results_free(vga_results_t* r) {
    delete r;
    return VGA_OK;
}

int vga_measure_int32_peak(double* ops_per_s) {
    if (!ops_per_s) { set_error("null argument"); return VGA_ERR_INVALID; }
    try {
        if (g_device < 0) { int rc = vga_init(0); if (rc != VGA_OK) return rc; }
        VGA_CUDA(cudaSetDevice(g_device));
        *ops_per_s = measure_int32_peak();
    } catch (CudaFail& f) { return f.code; }
    return VGA_OK;
}

}  // extern "C"
