// vga_seed_count_noalloc: the K1 counting stage (anchors per read) for callers that may not allocate device memory —
// code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
// Nothing here calls cudaMalloc / cudaFree / an allocating library routine.
//
// The stage's scratch (per-read k-mer ranges, per-k-mer position counts and table offsets — what vga_map_batch keeps in its
// per-index workspace) lives in ONE file-scope device array of VGA_NOALLOC_SCRATCH_BYTES.  Every GPU that loads the library
// holds that array until the process exits.  All calls share it and it keeps the last call's contents: the stage writes every
// word it later reads (the scan kernel below writes read_koff[0] = 0 itself), so nothing has to be cleared between calls —
// but two calls that overlap on the same GPU race on it.  Calls of one process are serialised by a mutex.
//
// The array's name is a host-side stand-in in host code, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress, taken once per device with that device current.
#include <mutex>

#include "../../include/vga.h"
#include "common.cuh"
#include "internal.h"

namespace vga {

static __device__ __align__(256) unsigned char g_noalloc_scratch[VGA_NOALLOC_SCRATCH_BYTES];

static const int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES];     // device address of the array, per device ordinal

// inclusive sum of read_koff[1..n] in place (read_prep wrote the per-read k-mer counts there); ONE CTA of 1024 threads walks
// the array in tiles, so no temporary storage is needed.  read_koff[0] = 0 is written here: the array holds stale data.
__global__ void __launch_bounds__(1024) noalloc_scan_kernel(u64* __restrict__ read_koff, u32 n) {
    __shared__ u64 s_warp[32];
    __shared__ u64 s_carry;
    const u32 lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) { s_carry = 0; read_koff[0] = 0; }
    __syncthreads();
    for (u32 t0 = 0; t0 < n; t0 += 1024) {
        const u32 i = t0 + threadIdx.x;
        u64 v = i < n ? read_koff[i + 1] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const u64 t = __shfl_up_sync(0xffffffffu, v, d);
            if (lane >= (u32)d) v += t;
        }
        if (lane == 31) s_warp[wid] = v;
        __syncthreads();
        if (wid == 0) {
            u64 x = s_warp[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const u64 t = __shfl_up_sync(0xffffffffu, x, d);
                if (lane >= (u32)d) x += t;
            }
            s_warp[lane] = x;
        }
        __syncthreads();
        const u64 carry = s_carry;
        v += carry + (wid ? s_warp[wid - 1] : 0);
        if (i < n) read_koff[i + 1] = v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = v;
        __syncthreads();
    }
}

static int noalloc_scratch_for(int device, unsigned char** out) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) { set_error("vga_seed_count_noalloc: device ordinal beyond 63"); return VGA_ERR_INVALID; }
    if (!g_noalloc_ptr[device]) {
        unsigned char* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);     // the current device's copy of the array
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            set_error(std::string("vga_seed_count_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
            return VGA_ERR_CUDA;
        }
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return VGA_OK;
}

}  // namespace vga

using namespace vga;

extern "C" int vga_seed_count_noalloc(vga_index_t* ix, const char* dev_seqs, const uint64_t* dev_seq_off, uint32_t n_reads,
                                      uint64_t n_bases, uint32_t* dev_n_anchors) {
    if (!ix || !dev_seq_off || !dev_n_anchors || (!dev_seqs && n_bases)) { set_error("vga_seed_count_noalloc: null argument"); return VGA_ERR_INVALID; }
    if (n_reads == 0) return VGA_OK;
    // layout: read_koff u64[n_reads + 1] | kcnt u32[n_kmers] | koff u32[n_kmers]; a read of L bases has at most L k-mers
    const u64 koff_bytes = (8ull * ((u64)n_reads + 1) + 255) & ~255ull;
    const u64 kmer_cap = koff_bytes < VGA_NOALLOC_SCRATCH_BYTES ? (VGA_NOALLOC_SCRATCH_BYTES - koff_bytes) / 8 : 0;
    if (koff_bytes >= VGA_NOALLOC_SCRATCH_BYTES || n_bases > kmer_cap) {
        set_error("vga_seed_count_noalloc: " + std::to_string(n_reads) + " reads / " + std::to_string(n_bases) +
                  " bases need more than the fixed scratch of " + std::to_string((u64)VGA_NOALLOC_SCRATCH_BYTES) +
                  " bytes (8 bytes per read + 8 bytes per base); use vga_map_batch_device or split the call");
        return VGA_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    try {
        VGA_CUDA(cudaSetDevice(ix->device));
        unsigned char* scratch = nullptr;
        const int rc = noalloc_scratch_for(ix->device, &scratch);
        if (rc != VGA_OK) return rc;
        cudaStream_t st = 0;
        SeedBuffers sb{};
        sb.seqs = dev_seqs; sb.seq_off = dev_seq_off; sb.n_reads = n_reads;
        sb.read_koff = (u64*)scratch;
        sb.kcnt = (u32*)(scratch + koff_bytes);
        sb.koff = sb.kcnt + kmer_cap;
        sb.read_acnt = dev_n_anchors;
        launch_read_prep(ix->dv, sb, st);
        noalloc_scan_kernel<<<1, 1024, 0, st>>>(sb.read_koff, n_reads);
        // the offsets are the caller's: what they add up to is checked before any k-mer slot is written
        u64 total_k = 0;
        VGA_CUDA(cudaMemcpyAsync(&total_k, sb.read_koff + n_reads, 8, cudaMemcpyDeviceToHost, st));
        VGA_CUDA(cudaStreamSynchronize(st));
        if (total_k > kmer_cap) {
            set_error("vga_seed_count_noalloc: the read offsets describe " + std::to_string(total_k) + " k-mers, more than n_bases allows (" +
                      std::to_string(kmer_cap) + " fit the fixed scratch)");
            return VGA_ERR_INVALID;
        }
        launch_seed_count(ix->dv, sb, st);
        VGA_CUDA(cudaGetLastError());
        VGA_CUDA(cudaStreamSynchronize(st));
    } catch (const CudaFail& f) {
        return f.code;
    }
    return VGA_OK;
}
