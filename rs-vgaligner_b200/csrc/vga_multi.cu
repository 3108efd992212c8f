// Multi-GPU plumbing behind the C ABI: reads are independent units (reference src/map.rs:56-111, 162-167 share
// only the immutable &Index), so the path shards with no data-path collective; the one collective is the replication of the
// index image, ONE ncclBroadcast over NVLink.  A Rust host has no torch: it gets the replication from this library.
//   one process (or host thread) per GPU:   vga_init(rank's device); rank 0: vga_index_upload + vga_comm_unique_id, hands the
//   128 id bytes to the others by whatever it has (pipe, file, MPI, a channel between threads); every rank:
//   vga_index_broadcast(&ix, id, rank, world); then vga_map_batch on its contiguous read range; outputs concatenated in
//   rank order (output order = input order, src/map.rs:123-133).
// NCCL is bound at run time (dlopen "libnccl.so.2"): the library carries no link-time dependency on it, a process that maps
// on one GPU never loads it, and inside a process that already holds an NCCL (torch) the same copy is used.
#include <dlfcn.h>
#include <nccl.h>

#include <mutex>

#include "common.cuh"
#include "internal.h"

namespace vga {
void resolve_pointers(vga_index* ix);

namespace {
struct NcclApi {
    void* h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

bool load_nccl() {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.h) return true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { set_error(std::string("NCCL not found (dlopen libnccl.so.2): ") + dlerror()); return false; }
    NcclApi a;
    a.h = h;
    a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    a.CommInitRank = (decltype(a.CommInitRank))dlsym(h, "ncclCommInitRank");
    a.Broadcast = (decltype(a.Broadcast))dlsym(h, "ncclBroadcast");
    a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
    if (!a.GetUniqueId || !a.CommInitRank || !a.Broadcast || !a.CommDestroy || !a.GetErrorString) {
        set_error("libnccl.so.2 lacks a required entry point");
        return false;
    }
    g_nccl = a;
    return true;
}
}  // namespace
}  // namespace vga

using namespace vga;

#define VGA_NCCL(call)                                                                                       \
    do {                                                                                                     \
        ncclResult_t r__ = (call);                                                                           \
        if (r__ != ncclSuccess) {                                                                            \
            set_error(std::string(#call) + ": " + g_nccl.GetErrorString(r__));                               \
            throw CudaFail{VGA_ERR_CUDA};                                                                    \
        }                                                                                                    \
    } while (0)

extern "C" {

int vga_comm_unique_id(void* id128) {
    if (!id128) { set_error("vga_comm_unique_id: null argument"); return VGA_ERR_INVALID; }
    if (!load_nccl()) return VGA_ERR_UNSUPPORTED;
    ncclUniqueId id;
    ncclResult_t r = g_nccl.GetUniqueId(&id);
    if (r != ncclSuccess) { set_error(std::string("ncclGetUniqueId: ") + g_nccl.GetErrorString(r)); return VGA_ERR_CUDA; }
    static_assert(sizeof(id) == VGA_COMM_ID_BYTES, "ncclUniqueId size");
    memcpy(id128, &id, sizeof(id));
    return VGA_OK;
}

int vga_index_broadcast(vga_index_t** ix, const void* id128, int rank, int world) {
    if (!ix || !id128 || world < 1 || rank < 0 || rank >= world) { set_error("vga_index_broadcast: bad argument"); return VGA_ERR_INVALID; }
    if (rank == 0 && !*ix) { set_error("vga_index_broadcast: rank 0 passes the uploaded index"); return VGA_ERR_INVALID; }
    if (rank != 0 && *ix) { set_error("vga_index_broadcast: the other ranks pass *ix = NULL"); return VGA_ERR_INVALID; }
    if (world == 1) return VGA_OK;
    if (!load_nccl()) return VGA_ERR_UNSUPPORTED;
    if (g_device < 0) { set_error("vga_index_broadcast: call vga_init(device) on this thread first"); return VGA_ERR_INVALID; }
    ncclComm_t comm = nullptr;
    cudaStream_t st = nullptr;
    vga_index* rep = nullptr;
    unsigned long long* d_size = nullptr;
    int rc = VGA_OK;
    try {
        VGA_CUDA(cudaSetDevice(g_device));
        ncclUniqueId id;
        memcpy(&id, id128, sizeof(id));
        VGA_NCCL(g_nccl.CommInitRank(&comm, world, id, rank));
        VGA_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        // the image size first (8 bytes), then the image itself: the ONE collective of the path
        VGA_CUDA(cudaMalloc(&d_size, 8));
        unsigned long long bytes = rank == 0 ? (*ix)->hdr.bytes : 0ull;
        VGA_CUDA(cudaMemcpyAsync(d_size, &bytes, 8, cudaMemcpyHostToDevice, st));
        VGA_NCCL(g_nccl.Broadcast(d_size, d_size, 8, ncclUint8, 0, comm, st));
        VGA_CUDA(cudaMemcpyAsync(&bytes, d_size, 8, cudaMemcpyDeviceToHost, st));
        VGA_CUDA(cudaStreamSynchronize(st));
        if (bytes < sizeof(IndexHeader)) { set_error("vga_index_broadcast: bad image size from rank 0"); throw CudaFail{VGA_ERR_INVALID}; }
        void* blob = nullptr;
        if (rank == 0) blob = (*ix)->d_blob;
        else {
            rep = new vga_index();
            rep->device = g_device;
            VGA_CUDA(cudaMalloc(&rep->d_blob, bytes));
            blob = rep->d_blob;
        }
        VGA_NCCL(g_nccl.Broadcast(blob, blob, (size_t)bytes, ncclUint8, 0, comm, st));
        VGA_CUDA(cudaStreamSynchronize(st));
        if (rank != 0) {
            VGA_CUDA(cudaMemcpy(&rep->hdr, rep->d_blob, sizeof(IndexHeader), cudaMemcpyDeviceToHost));
            if (rep->hdr.magic != INDEX_MAGIC || rep->hdr.bytes != bytes) { set_error("index image header mismatch after the broadcast"); throw CudaFail{VGA_ERR_INVALID}; }
            resolve_pointers(rep);
            *ix = rep;
            rep = nullptr;
        }
    } catch (CudaFail& f) { rc = f.code; }
    if (rep) { if (rep->d_blob) cudaFree(rep->d_blob); delete rep; }
    if (d_size) cudaFree(d_size);
    if (st) cudaStreamDestroy(st);
    if (comm) g_nccl.CommDestroy(comm);
    return rc;
}

}  // extern "C"
