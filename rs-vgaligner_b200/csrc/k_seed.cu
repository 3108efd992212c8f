// K1 — seeding: anchors_for_query (reference src/chain.rs:134-173) over a batch of reads.
//   split_into_kmers            src/io.rs:41-56      every overlapping k-mer, raw bytes
//   find_start_position_in_index src/index.rs:309-325 membership + start offset  -> slot probe
//   find_positions_for_query_kmer src/index.rs:353-382 positions up to the delimiter -> (off,cnt)
// A warp owns a read.  Pass 1 (seed_count) probes the open-addressed table once per k-mer and keeps
// (off,cnt); pass 2 (seed_emit) turns them into anchors in (query position asc, table order) —
// the emission order of the reference, which also defines Anchor.id.
// Roofline: HBM.  Algorithmic bytes per k-mer: 1 (read byte) + 32 (one DRAM sector per probe)
// + 8 per position read + 16 per anchor written.
#include "common.cuh"
#include "internal.h"

namespace vga {

static const int SEED_WARPS = 8;   // warps per CTA

__global__ void read_prep_kernel(const u64* __restrict__ seq_off, u32 n, u32 k, u64* __restrict__ read_koff) {
    u32 r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r == 0) read_koff[0] = 0;
    if (r >= n) return;
    u64 len = seq_off[r + 1] - seq_off[r];
    read_koff[r + 1] = len >= k ? len - k + 1 : 0;     // src/io.rs:47
}

// 2-bit code of an uppercase ACGT byte (A0 C1 G2 T3) or 4 for anything else; a k-mer with any
// other byte can never equal an index k-mer (index k-mers are ACGT only, src/kmer.rs:401,459).
__device__ __forceinline__ u32 base_code(u32 c) {
    u32 x = (c >> 1) & 3;            // A0 C1 T2 G3
    x ^= x >> 1;                     // A0 C1 T3 G2
    bool ok = (c == 'A') | (c == 'C') | (c == 'G') | (c == 'T');
    return ok ? x : 4u;
}

// A warp owns a read.  Each lane classifies ONE read byte per 32-position chunk; three ballots give the bit planes of
// the 2-bit codes (and of the non-ACGT / past-the-end flag) for 32 positions, and the k-mer starting at position
// i0+lane is a funnel shift over the planes of this chunk and the next one — the slot keys are stored in the same
// bit-plane form (kmer_planes_from_2bit), so no per-base packing happens here.
// COMPACT: the bucket table of common.cuh (k <= 16) — one 64-byte bucket per probe; koff receives the entry's payload
// (an inline anchor, or the first element of the key's run in pos2), kcnt the number of positions.
template <bool COMPACT>
__global__ void __launch_bounds__(SEED_WARPS * 32)
seed_count_kernel(DevIndex ix, const char* __restrict__ seqs, const u64* __restrict__ seq_off, u32 n,
                  const u64* __restrict__ read_koff, u32* __restrict__ kcnt, u32* __restrict__ koff,
                  u32* __restrict__ read_acnt) {
    const u32 FULL = 0xffffffffu;
    const u32 lane = threadIdx.x & 31;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 n_warps = (gridDim.x * blockDim.x) >> 5;
    const u32 k = ix.k;
    const u32 kmask = k >= 32 ? 0xffffffffu : ((1u << k) - 1u);
    for (u32 r = warp; r < n; r += n_warps) {
        const u64 sb = seq_off[r];
        const u32 len = (u32)(seq_off[r + 1] - sb);
        const u64 kb = read_koff[r];
        const u32 nk = (u32)(read_koff[r + 1] - kb);
        const u8* s = (const u8*)seqs + sb;
        u32 total = 0;
        // planes of the current chunk
        u32 c = lane < len ? base_code(s[lane]) : 4u;
        u32 p0 = __ballot_sync(FULL, c & 1u), p1 = __ballot_sync(FULL, c & 2u), pb = __ballot_sync(FULL, c & 4u);
        for (u32 i0 = 0; i0 < nk; i0 += 32) {
            const u32 jn = i0 + 32 + lane;
            c = jn < len ? base_code(s[jn]) : 4u;
            const u32 n0 = __ballot_sync(FULL, c & 1u), n1 = __ballot_sync(FULL, c & 2u), nbad = __ballot_sync(FULL, c & 4u);
            const u32 i = i0 + lane;
            u32 cnt = 0, off = 0;
            if (i < nk) {
                const u32 bad = __funnelshift_r(pb, nbad, lane) & kmask;
                if (!bad) {
                    const u32 lo = __funnelshift_r(p0, n0, lane) & kmask, hi = __funnelshift_r(p1, n1, lane) & kmask;
                    if (COMPACT) {
                        const u32 key32 = lo | (hi << 16);
                        u32 bi = ct_bucket(key32, ix.n_buckets);
                        while (true) {
                            const uint4* bp = ix.buckets + 4ull * bi;
                            const uint4 e0 = __ldg(bp), e1 = __ldg(bp + 1), e2 = __ldg(bp + 2), e3 = __ldg(bp + 3);
                            u32 pay = 0;
                            bool hit = false;
                            if (e0.x == key32 && (e0.x | e0.y)) { pay = e0.y; hit = true; }
                            if (e0.z == key32 && (e0.z | e0.w)) { pay = e0.w; hit = true; }
                            if (e1.x == key32 && (e1.x | e1.y)) { pay = e1.y; hit = true; }
                            if (e1.z == key32 && (e1.z | e1.w)) { pay = e1.w; hit = true; }
                            if (e2.x == key32 && (e2.x | e2.y)) { pay = e2.y; hit = true; }
                            if (e2.z == key32 && (e2.z | e2.w)) { pay = e2.w; hit = true; }
                            if (e3.x == key32 && (e3.x | e3.y)) { pay = e3.y; hit = true; }
                            if (e3.z == key32 && (e3.z | e3.w)) { pay = e3.w; hit = true; }
                            if (hit) {
                                if (pay & CT_INLINE) { cnt = 1; off = pay; }
                                else {
                                    cnt = (pay >> 26) & 31u; off = pay & CT_LOW26;
                                    if (cnt > CT_MAX_CNT) { cnt = __ldg(ix.pos2 + off).x; off++; }      // header element: the real count
                                }
                                break;
                            }
                            if ((e3.z | e3.w) == 0) break;           // room in this bucket: the key is not in the index
                            bi = bi + 1 == ix.n_buckets ? 0 : bi + 1;
                        }
                    } else {
                    const u64 key = ((u64)hi << 32) | lo;
                    u32 h = (u32)kmer_slot_hash(key) & ix.slot_mask;
                    while (true) {
                        const uint4 raw = __ldg((const uint4*)(ix.slots + h));
                        if (raw.w == 0) break;                      // empty slot: not in index
                        if (raw.x == lo && raw.y == hi) { off = raw.z; cnt = raw.w; break; }
                        h = (h + 1) & ix.slot_mask;
                    }
                    }
                }
                kcnt[kb + i] = cnt;
                koff[kb + i] = off;
            }
            total += __reduce_add_sync(FULL, cnt);
            p0 = n0; p1 = n1; pb = nbad;
        }
        if (lane == 0) read_acnt[r] = total;
    }
}

template <bool COMPACT>
__global__ void __launch_bounds__(SEED_WARPS * 32)
seed_emit_kernel(DevIndex ix, u32 n, const u64* __restrict__ read_koff, const u32* __restrict__ kcnt,
                 const u32* __restrict__ koff, const u32* __restrict__ read_aoff, u32* __restrict__ a_qb,
                 u32* __restrict__ a_tb, u32* __restrict__ a_te, u32* __restrict__ sort_key,
                 u32* __restrict__ sort_val) {
    const u32 lane = threadIdx.x & 31;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 n_warps = (gridDim.x * blockDim.x) >> 5;
    for (u32 r = warp; r < n; r += n_warps) {
        const u64 kb = read_koff[r];
        const u32 nk = (u32)(read_koff[r + 1] - kb);
        const u32 abase = read_aoff[r];
        if (read_aoff[r + 1] == abase) continue;
        // (te, id) sort pairs only for the reads the CUB segmented sort handles; the fused chain kernel builds its own keys
        const bool big = read_aoff[r + 1] - abase > CHAIN_SMAX;
        u32 carry = 0;
        for (u32 i0 = 0; i0 < nk; i0 += 32) {
            const u32 i = i0 + lane;
            u32 cnt = 0, off = 0;
            if (i < nk) { cnt = kcnt[kb + i]; off = koff[kb + i]; }
            u32 incl = cnt;                                // warp inclusive scan
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                u32 t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= (u32)d) incl += t;
            }
            const u32 first = carry + incl - cnt;          // Anchor.id of this k-mer's first anchor
            for (u32 t = 0; t < cnt; t++) {                // table order (src/index.rs:372-378)
                const u32 id = first + t;
                u32 tb, te;
                if (COMPACT) {
                    if (off & CT_INLINE) { tb = off & CT_LOW26; te = tb + ix.k + ((off >> 26) & 31u); }
                    else { const uint2 pp = __ldg(ix.pos2 + off + t); tb = pp.x; te = pp.y; }
                } else { tb = __ldg(ix.pos_tb + off + t); te = __ldg(ix.pos_te + off + t); }
                a_qb[abase + id] = i;                      // query_begin = i, query_end = i + k
                a_tb[abase + id] = tb;
                a_te[abase + id] = te;
                // sort key: Reverse before Forward, then the position (src/chain.rs:386-389); bit 31 of te is the orientation
                if (big) { sort_key[abase + id] = te ^ 0x80000000u; sort_val[abase + id] = id; }
            }
            carry += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
}

static int grid_for_warps(u32 n_items, int warps_per_cta) {
    long long ctas = ((long long)n_items + warps_per_cta - 1) / warps_per_cta;
    const long long cap = 148 * 16;        // grid-stride beyond this
    if (ctas > cap) ctas = cap;
    if (ctas < 1) ctas = 1;
    return (int)ctas;
}

int launch_read_prep(const DevIndex& ix, const SeedBuffers& b, cudaStream_t st) {
    u32 threads = 256, blocks = (b.n_reads + threads - 1) / threads;
    if (blocks == 0) blocks = 1;
    read_prep_kernel<<<blocks, threads, 0, st>>>(b.seq_off, b.n_reads, ix.k, b.read_koff);
    return 1;
}

int launch_seed_count(const DevIndex& ix, const SeedBuffers& b, cudaStream_t st) {
    if (ix.n_buckets)
        seed_count_kernel<true><<<grid_for_warps(b.n_reads, SEED_WARPS), SEED_WARPS * 32, 0, st>>>(
            ix, b.seqs, b.seq_off, b.n_reads, b.read_koff, b.kcnt, b.koff, b.read_acnt);
    else
    seed_count_kernel<false><<<grid_for_warps(b.n_reads, SEED_WARPS), SEED_WARPS * 32, 0, st>>>(
        ix, b.seqs, b.seq_off, b.n_reads, b.read_koff, b.kcnt, b.koff, b.read_acnt);
    return 1;
}

int launch_seed_emit(const DevIndex& ix, const SeedBuffers& b, cudaStream_t st) {
    if (ix.n_buckets)
        seed_emit_kernel<true><<<grid_for_warps(b.n_reads, SEED_WARPS), SEED_WARPS * 32, 0, st>>>(
            ix, b.n_reads, b.read_koff, b.kcnt, b.koff, b.read_aoff, b.a_qb, b.a_tb, b.a_te, b.sort_key_in,
            b.sort_val_in);
    else
    seed_emit_kernel<false><<<grid_for_warps(b.n_reads, SEED_WARPS), SEED_WARPS * 32, 0, st>>>(
        ix, b.n_reads, b.read_koff, b.kcnt, b.koff, b.read_aoff, b.a_qb, b.a_tb, b.a_te, b.sort_key_in,
        b.sort_val_in);
    return 1;
}

}  // namespace vga
