// K3 -- checkpoint snapshot kernels for sm_100a: PR-SUM64 checksum, pack (gather + checksum)
// and unpack (scatter + checksum), plus the pinned host arena, chunked async copies and the
// native atomic file writer.
//
// Replaces the snapshot half of the reference's save_ckpt_vanilla / save_ckpt_distributed
// (pyrecover/checkpoint.py:60-84, 254-274: state_dict -> per-storage synchronous pageable
// D2H -> file -> re-read for md5) and the restore half of load_ckpt_* (:155-199, 352-356).
//
// Roofline: HBM.  Algorithmic bytes per packed byte: 2 (read tensor + write image) for
// pack/unpack into device memory, 1 for checksum.  The D2H leg that follows is PCIe-bound
// and is done by the copy engines (pr_copy_d2h_async), not by SMs.
//
// Work decomposition: the image is cut into tiles of kTileBytes that never straddle two
// tensors; a device table maps tile -> tensor (built once per model layout, this is the
// plan's scratch memory).  A CTA streams one tile per iteration with 16-byte vectors,
// kUnroll loads in flight per thread; the checksum rides on the registers holding the data.
#include <errno.h>
#include <fcntl.h>
#include <stdlib.h>
#include <unistd.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;
constexpr uint64_t kTileBytes = 64 * 1024;

// ---- PR-SUM64 on one 16-byte vector whose first word has global word index g0 ---------
__device__ __forceinline__ void sum_vec(const uint4 &w, uint64_t g0, uint64_t &A, uint64_t &B) {
    uint64_t s = (uint64_t)w.x + w.y + w.z + w.w;
    uint64_t t = (uint64_t)w.y + 2ull * w.z + 3ull * w.w;
    A += s;
    B += (g0 + 1) * s + t;
}

// Reduce (A, B) over the CTA and add into sum[0..1].  Integer addition mod 2^64 commutes,
// so the atomics give a deterministic result regardless of CTA scheduling.
__device__ __forceinline__ void block_sum_commit(uint64_t A, uint64_t B, uint64_t *sum) {
    __shared__ unsigned long long sA[kThreads / 32], sB[kThreads / 32];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        A += __shfl_down_sync(0xffffffffu, (unsigned long long)A, off);
        B += __shfl_down_sync(0xffffffffu, (unsigned long long)B, off);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { sA[warp] = A; sB[warp] = B; }
    __syncthreads();
    if (warp == 0) {
        A = lane < kThreads / 32 ? sA[lane] : 0ull;
        B = lane < kThreads / 32 ? sB[lane] : 0ull;
#pragma unroll
        for (int off = 4; off > 0; off >>= 1) {
            A += __shfl_down_sync(0xffffffffu, (unsigned long long)A, off);
            B += __shfl_down_sync(0xffffffffu, (unsigned long long)B, off);
        }
        if (lane == 0 && (A | B)) {
            atomicAdd((unsigned long long *)&sum[0], (unsigned long long)A);
            atomicAdd((unsigned long long *)&sum[1], (unsigned long long)B);
        }
    }
}

// Assemble up to 16 bytes starting at src (any alignment) into a zero-padded vector.
__device__ __forceinline__ uint4 load_bytes(const uint8_t *src, int nvalid) {
    uint32_t w[4] = {0u, 0u, 0u, 0u};
    for (int k = 0; k < nvalid; ++k) w[k >> 2] |= (uint32_t)src[k] << (8 * (k & 3));
    return make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __forceinline__ void store_bytes(uint8_t *dst, const uint4 &v, int nvalid) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    for (int k = 0; k < nvalid; ++k) dst[k] = (uint8_t)(w[k >> 2] >> (8 * (k & 3)));
}

// ---- checksum of a flat byte range ------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
checksum_kernel(const uint8_t *__restrict__ data, uint64_t nbytes, uint64_t word_base,
                uint64_t *sum) {
    uint64_t A = 0, B = 0;
    const uint64_t nvec = nbytes / 16;
    const uint64_t stride = (uint64_t)gridDim.x * kThreads;
    const uint4 *d4 = reinterpret_cast<const uint4 *>(data);
    uint64_t i = (uint64_t)blockIdx.x * kThreads + threadIdx.x;
    for (; i + (kUnroll - 1) * stride < nvec; i += kUnroll * stride) {
        uint4 w[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) w[u] = ld_stream_nc(d4 + i + u * stride);
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) sum_vec(w[u], word_base + 4 * (i + u * stride), A, B);
    }
    for (; i < nvec; i += stride) sum_vec(ld_stream_nc(d4 + i), word_base + 4 * i, A, B);
    if (blockIdx.x == 0 && threadIdx.x == 0 && (nbytes & 15)) {
        uint4 w = load_bytes(data + 16 * nvec, (int)(nbytes & 15));
        sum_vec(w, word_base + 4 * nvec, A, B);
    }
    block_sum_commit(A, B, sum);
}

// Same for a start address that is not 16-byte aligned (slow path, byte loads).
__global__ void __launch_bounds__(kThreads)
checksum_unaligned_kernel(const uint8_t *__restrict__ data, uint64_t nbytes, uint64_t word_base,
                          uint64_t *sum) {
    uint64_t A = 0, B = 0;
    const uint64_t nvec = (nbytes + 15) / 16;
    const uint64_t stride = (uint64_t)gridDim.x * kThreads;
    for (uint64_t i = (uint64_t)blockIdx.x * kThreads + threadIdx.x; i < nvec; i += stride) {
        uint64_t left = nbytes - 16 * i;
        uint4 w = load_bytes(data + 16 * i, left < 16 ? (int)left : 16);
        sum_vec(w, word_base + 4 * i, A, B);
    }
    block_sum_commit(A, B, sum);
}

// ---- pack / unpack ----------------------------------------------------------------------
struct PlanDev {
    const uint64_t *seg_ptr;      // tensor base address
    const uint64_t *seg_off;      // byte offset of the tensor in the image (multiple of 16)
    const uint64_t *seg_bytes;
    const uint64_t *tile_first;   // index of the tensor's first tile
    const uint32_t *tile_seg;     // tile -> tensor
    uint64_t ntiles;
};

template <bool UNPACK, bool WITH_SUM>
__global__ void __launch_bounds__(kThreads)
pack_kernel(PlanDev pd, uint8_t *image, uint64_t *sum) {
    uint64_t A = 0, B = 0;
    for (uint64_t t = blockIdx.x; t < pd.ntiles; t += gridDim.x) {
        const uint32_t s = pd.tile_seg[t];
        const uint64_t toff = (t - pd.tile_first[s]) * kTileBytes;
        const uint64_t sbytes = pd.seg_bytes[s];
        uint8_t *tensor = reinterpret_cast<uint8_t *>(pd.seg_ptr[s]) + toff;
        const uint64_t ioff = pd.seg_off[s] + toff;       // multiple of 16
        uint8_t *img = image + ioff;
        const uint64_t len = sbytes - toff < kTileBytes ? sbytes - toff : kTileBytes;
        const uint32_t nfull = (uint32_t)(len / 16);
        const uint64_t w0 = ioff / 4;
        if ((reinterpret_cast<uintptr_t>(tensor) & 15u) == 0) {
            const uint4 *src = reinterpret_cast<const uint4 *>(UNPACK ? img : tensor);
            uint4 *dst = reinterpret_cast<uint4 *>(UNPACK ? tensor : img);
            uint32_t k = threadIdx.x;
            for (; k + (kUnroll - 1) * kThreads < nfull; k += kUnroll * kThreads) {
                uint4 w[kUnroll];
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) w[u] = ld_stream(src + k + u * kThreads);
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) {
                    if (WITH_SUM) sum_vec(w[u], w0 + 4ull * (k + u * kThreads), A, B);
                    st_stream(dst + k + u * kThreads, w[u]);
                }
            }
            for (; k < nfull; k += kThreads) {
                uint4 w = ld_stream(src + k);
                if (WITH_SUM) sum_vec(w, w0 + 4ull * k, A, B);
                st_stream(dst + k, w);
            }
        } else {
            for (uint32_t k = threadIdx.x; k < nfull; k += kThreads) {
                uint4 w;
                if (UNPACK) {
                    w = ld_stream(reinterpret_cast<const uint4 *>(img) + k);
                    store_bytes(tensor + 16ull * k, w, 16);
                } else {
                    w = load_bytes(tensor + 16ull * k, 16);
                    st_stream(reinterpret_cast<uint4 *>(img) + k, w);
                }
                if (WITH_SUM) sum_vec(w, w0 + 4ull * k, A, B);
            }
        }
        const int rem = (int)(len & 15);
        if (rem && threadIdx.x == 0) {   // zero-padded last vector of the tensor
            uint4 w;
            if (UNPACK) {
                w = ld_stream(reinterpret_cast<const uint4 *>(img) + nfull);
                store_bytes(tensor + 16ull * nfull, w, rem);
            } else {
                w = load_bytes(tensor + 16ull * nfull, rem);
                st_stream(reinterpret_cast<uint4 *>(img) + nfull, w);
            }
            if (WITH_SUM) sum_vec(w, w0 + 4ull * nfull, A, B);
        }
    }
    if (WITH_SUM) block_sum_commit(A, B, sum);
}

// ---- bulk-copy pack / unpack: TMA-staged through shared memory -------------------------------
// The same gather / scatter + PR-SUM64, but the bytes move as bulk asynchronous copies
// (cp.async.bulk: global -> shared -> global, UBLKCP in SASS) issued by ONE elected thread per CTA and
// tracked by mbarriers; no data passes through registers except for the checksum, which the other
// threads take from the shared-memory tile while the copy engine of the SM already writes it out.
// A handful of CTAs keep enough bytes in flight (kBulkStages x kBulkChunk = 192 KiB of shared memory
// per CTA), which leaves the remaining SMs, their registers and issue slots to the forward pass this
// kernel runs beside (north_star: "pack kernel ... via TMA-staged async copies on a side stream").
//
// Pipeline of one CTA over its chunks q = 0, 1, ... (a chunk = <= kBulkChunk bytes of one tile):
//   load(q)  : mbarrier.arrive.expect_tx(full[q % S]) + cp.async.bulk g->s        (thread 0)
//   wait(q)  : mbarrier.try_wait.parity(full[q % S])                              (all threads)
//   store(q) : cp.async.bulk s->g + commit_group                                  (thread 0)
//   sum(q)   : PR-SUM64 over the tile in shared memory                            (all threads)
//   refill   : at the top of iteration q, stage (q-2) % S is reloaded with chunk q+S-2 once
//              cp.async.bulk.wait_group.read 1 says store(q-2) has drained it and the CTA barrier says
//              every thread has finished summing it.
// Requires every tensor base to be 16-byte aligned (the plan checks); a tensor's last < 16 bytes
// are moved by thread 0 with byte accesses, zero-padded in the image like the LDG kernel does.
constexpr int kBulkThreads = 256;
constexpr uint32_t kBulkChunk = 32 * 1024;
constexpr int kBulkStages = 6;
constexpr size_t kBulkSmemBytes = (size_t)kBulkStages * kBulkChunk + 8 * kBulkStages;
static_assert(kTileBytes % kBulkChunk == 0, "a tile is a whole number of chunks");

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" :: "r"(smem_addr(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar, uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(smem_addr(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_addr(bar)), "l"(pol) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *gdst, const void *smem_src, uint32_t bytes, uint64_t pol) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
                 :: "l"(gdst), "r"(smem_addr(smem_src)), "r"(bytes), "l"(pol) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

struct BulkChunk {
    uint8_t *tensor;       // first byte of the chunk in the tensor
    uint64_t ioff;         // its offset in the image (multiple of 16)
    uint32_t len;          // bytes (<= kBulkChunk); only a tensor's last chunk may not be a multiple of 16
};

// Walks the chunks of tiles blockIdx.x, blockIdx.x + gridDim.x, ...  The tile -> tensor table is only
// consulted when a tile leaves the tensor the previous one was in (tensors are arenas of gigabytes).
struct BulkCursor {
    uint64_t t, first = 0, ntiles_seg = 0, ptr = 0, off = 0, bytes = 0;
    uint32_t h = 0;
    bool cached = false;
    __device__ __forceinline__ explicit BulkCursor(uint64_t t0) : t(t0) {}
    __device__ __forceinline__ bool next(const PlanDev &pd, BulkChunk &c) {
        while (t < pd.ntiles) {
            if (!cached || t < first || t >= first + ntiles_seg) {
                const uint32_t s = pd.tile_seg[t];
                first = pd.tile_first[s];
                ptr = pd.seg_ptr[s];
                off = pd.seg_off[s];
                bytes = pd.seg_bytes[s];
                ntiles_seg = (bytes + kTileBytes - 1) / kTileBytes;
                cached = true;
            }
            const uint64_t toff = (t - first) * kTileBytes + (uint64_t)h * kBulkChunk;
            const bool exists = toff < bytes;
            if (exists) {
                const uint64_t left = bytes - toff;
                c.tensor = reinterpret_cast<uint8_t *>(ptr) + toff;
                c.ioff = off + toff;
                c.len = left < kBulkChunk ? (uint32_t)left : kBulkChunk;
            }
            if (exists && h + 1 < kTileBytes / kBulkChunk && toff + kBulkChunk < bytes) {
                ++h;
            } else {
                h = 0;
                t += gridDim.x;
            }
            if (exists) return true;
        }
        return false;
    }
};

template <bool UNPACK, bool WITH_SUM>
__global__ void __launch_bounds__(kBulkThreads, 1)
pack_bulk_kernel(PlanDev pd, uint8_t *image, uint64_t *sum) {
    extern __shared__ __align__(128) uint8_t bulk_smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(bulk_smem + (size_t)kBulkStages * kBulkChunk);
    const bool leader = threadIdx.x == 0;
    if (!WITH_SUM && !leader) return;            // without a checksum the elected thread does everything
    if (leader) {
        for (int s = 0; s < kBulkStages; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (WITH_SUM) __syncthreads();
    const uint64_t pol = l2_evict_first_policy();
    uint64_t A = 0, B = 0;
    BulkCursor prod(blockIdx.x), cons(blockIdx.x);
    BulkChunk pc, cc;
    uint32_t issued = 0;
    auto issue = [&](const BulkChunk &c) {        // leader only: chunk number `issued` -> stage issued % S
        uint64_t *bar = &full[issued % kBulkStages];
        const uint32_t nb = c.len & ~15u;
        mbar_arrive_expect_tx(bar, nb);           // nb == 0 (a tensor's < 16-byte tail chunk): the phase completes at once
        if (nb) bulk_g2s(bulk_smem + (size_t)(issued % kBulkStages) * kBulkChunk, UNPACK ? image + c.ioff : c.tensor, nb, bar, pol);
        ++issued;
    };
    if (leader)
        for (int i = 0; i < kBulkStages - 2; ++i)
            if (prod.next(pd, pc)) issue(pc);
    for (uint32_t q = 0; cons.next(pd, cc); ++q) {
        if (WITH_SUM) __syncthreads();            // every thread is done with the tiles of chunks < q
        if (leader && prod.next(pd, pc)) {
            bulk_wait_read<1>();                  // store(q-2) has drained the stage about to be refilled
            issue(pc);
        }
        const uint32_t stage = q % kBulkStages;
        uint8_t *tile = bulk_smem + (size_t)stage * kBulkChunk;
        const uint32_t nb = cc.len & ~15u;
        mbar_wait(&full[stage], (q / kBulkStages) & 1u);
        if (leader) {
            if (nb) bulk_s2g(UNPACK ? cc.tensor : image + cc.ioff, tile, nb, pol);
            bulk_commit();                        // one group per chunk, empty or not: keeps wait_group counts uniform
        }
        const uint64_t w0 = cc.ioff / 4;
        if (WITH_SUM) {
            const uint4 *tv = reinterpret_cast<const uint4 *>(tile);
            const uint32_t nvec = nb / 16;
#pragma unroll 4
            for (uint32_t k = threadIdx.x; k < nvec; k += kBulkThreads) sum_vec(tv[k], w0 + 4ull * k, A, B);
        }
        const int rem = (int)(cc.len & 15u);
        if (rem && leader) {                      // zero-padded last vector of the tensor, as in pack_kernel
            uint4 w;
            uint8_t *img = image + cc.ioff + nb;
            if (UNPACK) {
                w = ld_stream(img);
                store_bytes(cc.tensor + nb, w, rem);
            } else {
                w = load_bytes(cc.tensor + nb, rem);
                st_stream(img, w);
            }
            if (WITH_SUM) sum_vec(w, w0 + nb / 4, A, B);
        }
    }
    if (leader) bulk_wait_all();                  // shared memory must outlive the stores reading it
    if (WITH_SUM) block_sum_commit(A, B, sum);
}

// Resident CTAs per SM, measured on B200 over 2 GiB tensors (round 1): the copy-shaped pack / unpack
// kernels peak with FEW concurrent streams (3 CTAs/SM: 6300 GB/s = 0.98 of the measured copy peak;
// 8: 0.93), the read-only checksum with many (12: 6855 GB/s; 3: 5401).  PR_CKPT_BLOCKS_PER_SM
// overrides both.
int ckpt_blocks_per_sm(bool copy_shaped) {
    static int env = [] {
        const char *e = getenv("PR_CKPT_BLOCKS_PER_SM");
        int b = e ? atoi(e) : 0;
        return b > 0 ? b : 0;
    }();
    if (env) return env;
    return copy_shaped ? 3 : 12;
}

}  // namespace

// Validation shared by pr_pack_plan_create and pr_ckpt_pack_noalloc: every slot inside the image (no
// uint64 wrap-around) and no two non-empty slots overlapping.  `order` is caller-provided scratch of n
// ints (the no-allocation entry point must not touch the heap).  Empty tensors own no bytes and are
// left out of the sweep, so one cannot hide an overlap between its neighbours.
static int validate_slots(const char *who, const void *const *tensor, const uint64_t *nbytes, const uint64_t *dst_off,
                          int n, uint64_t image_bytes, int *order) {
    int m = 0;
    for (int i = 0; i < n; ++i) {
        PR_REQUIRE(dst_off[i] % 16 == 0, "%s: dst_off[%d]=%llu is not a multiple of 16", who, i,
                   (unsigned long long)dst_off[i]);
        PR_REQUIRE(nbytes[i] <= UINT64_MAX - 15, "%s: tensor %d has an impossible size", who, i);
        const uint64_t padded = (nbytes[i] + 15) / 16 * 16;
        PR_REQUIRE(dst_off[i] <= image_bytes && padded <= image_bytes - dst_off[i],
                   "%s: tensor %d (%llu B at %llu) overruns the image (%llu B)", who, i,
                   (unsigned long long)nbytes[i], (unsigned long long)dst_off[i], (unsigned long long)image_bytes);
        PR_REQUIRE(nbytes[i] == 0 || tensor[i] != nullptr, "%s: tensor %d is NULL", who, i);
        if (nbytes[i]) order[m++] = i;
    }
    std::sort(order, order + m, [&](int a, int b) { return dst_off[a] < dst_off[b]; });
    for (int k = 1; k < m; ++k) {
        const int a = order[k - 1], b = order[k];
        PR_REQUIRE(dst_off[a] + (nbytes[a] + 15) / 16 * 16 <= dst_off[b], "%s: tensors %d and %d overlap in the image",
                   who, a, b);
    }
    return PR_OK;
}

struct pr_pack_plan {
    int device = -1;
    int n = 0;
    uint64_t image_bytes = 0;
    uint64_t payload_bytes = 0;
    void *table = nullptr;   // one device allocation holding all five arrays
    PlanDev dev{};
    bool bulk_ok = true;     // every tensor base is 16-byte aligned: the bulk-copy kernels apply
};

extern "C" int pr_checksum(const void *data, uint64_t nbytes, uint64_t word_base, uint64_t *sum,
                           void *stream_) {
    PR_REQUIRE(sum != nullptr, "pr_checksum: sum_dev is NULL");
    if (nbytes == 0) return PR_OK;
    PR_REQUIRE(data != nullptr, "pr_checksum: data is NULL");
    cudaStream_t stream = (cudaStream_t)stream_;
    uint64_t nvec = (nbytes + 15) / 16;
    uint64_t need = (nvec + (uint64_t)kThreads * kUnroll - 1) / ((uint64_t)kThreads * kUnroll);
    uint64_t cap = (uint64_t)pr_sm_count() * ckpt_blocks_per_sm(false);
    int grid = (int)(need < cap ? need : cap);
    if (pr_aligned16(data))
        checksum_kernel<<<grid, kThreads, 0, stream>>>((const uint8_t *)data, nbytes, word_base, sum);
    else
        checksum_unaligned_kernel<<<grid, kThreads, 0, stream>>>((const uint8_t *)data, nbytes,
                                                                 word_base, sum);
    PR_CUDA(cudaGetLastError());
    pr_count_launch();
    return PR_OK;
}

extern "C" int pr_pack_plan_create(const void *const *tensor, const uint64_t *nbytes,
                                   const uint64_t *dst_off, int n, uint64_t image_bytes,
                                   pr_pack_plan **out) {
    PR_REQUIRE(out != nullptr, "pr_pack_plan_create: plan_out is NULL");
    *out = nullptr;
    PR_REQUIRE(n >= 0 && (n == 0 || (tensor && nbytes && dst_off)), "pr_pack_plan_create: bad table");
    std::vector<uint64_t> ptrs(n), offs(n), lens(n), first(n);
    std::vector<uint32_t> tile_seg;
    std::vector<int> order(n);
    int vrc = validate_slots("pr_pack_plan_create", tensor, nbytes, dst_off, n, image_bytes, order.data());
    if (vrc) return vrc;
    uint64_t payload = 0;
    bool bulk_ok = true;
    for (int i = 0; i < n; ++i) {
        ptrs[i] = (uint64_t)(uintptr_t)tensor[i];
        offs[i] = dst_off[i];
        lens[i] = nbytes[i];
        first[i] = tile_seg.size();
        uint64_t nt = (nbytes[i] + kTileBytes - 1) / kTileBytes;
        tile_seg.insert(tile_seg.end(), nt, (uint32_t)i);
        payload += nbytes[i];
        if (nbytes[i] && !pr_aligned16(tensor[i])) bulk_ok = false;
    }
    int device = 0;
    PR_CUDA(cudaGetDevice(&device));
    auto *plan = new pr_pack_plan();
    plan->device = device;
    plan->n = n;
    plan->image_bytes = image_bytes;
    plan->payload_bytes = payload;
    plan->bulk_ok = bulk_ok;
    const uint64_t ntiles = tile_seg.size();
    const size_t seg_bytes = sizeof(uint64_t) * (size_t)n;
    const size_t total = 4 * seg_bytes + sizeof(uint32_t) * ntiles;
    if (total) {
        cudaError_t e = cudaMalloc(&plan->table, total);
        if (e != cudaSuccess) {
            delete plan;
            pr_set_error("pr_pack_plan_create: cudaMalloc(%zu) failed: %s", total, cudaGetErrorString(e));
            return PR_ERR_CUDA;
        }
        std::vector<uint8_t> host(total);
        memcpy(host.data() + 0 * seg_bytes, ptrs.data(), seg_bytes);
        memcpy(host.data() + 1 * seg_bytes, offs.data(), seg_bytes);
        memcpy(host.data() + 2 * seg_bytes, lens.data(), seg_bytes);
        memcpy(host.data() + 3 * seg_bytes, first.data(), seg_bytes);
        memcpy(host.data() + 4 * seg_bytes, tile_seg.data(), sizeof(uint32_t) * ntiles);
        e = cudaMemcpy(plan->table, host.data(), total, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            cudaFree(plan->table);
            delete plan;
            pr_set_error("pr_pack_plan_create: table upload failed: %s", cudaGetErrorString(e));
            return PR_ERR_CUDA;
        }
        uint8_t *base = (uint8_t *)plan->table;
        plan->dev.seg_ptr = (const uint64_t *)(base + 0 * seg_bytes);
        plan->dev.seg_off = (const uint64_t *)(base + 1 * seg_bytes);
        plan->dev.seg_bytes = (const uint64_t *)(base + 2 * seg_bytes);
        plan->dev.tile_first = (const uint64_t *)(base + 3 * seg_bytes);
        plan->dev.tile_seg = (const uint32_t *)(base + 4 * seg_bytes);
    }
    plan->dev.ntiles = ntiles;
    *out = plan;
    return PR_OK;
}

extern "C" int pr_pack_plan_destroy(pr_pack_plan *plan) {
    if (!plan) return PR_OK;
    if (plan->table) cudaFree(plan->table);
    delete plan;
    return PR_OK;
}

// Default grid of the bulk-copy kernels.  One CTA keeps kBulkStages - 2 chunks of loads in flight, so a few
// dozen CTAs already move data at a large fraction of the HBM rate while leaving most SMs to whatever runs
// beside the snapshot (the next forward pass).  PR_CKPT_BULK_BLOCKS overrides; per call: max_blocks.
constexpr int kBulkDefaultBlocks = 64;

static int bulk_default_blocks() {
    static int env = [] {
        const char *e = getenv("PR_CKPT_BULK_BLOCKS");
        int b = e ? atoi(e) : 0;
        return b > 0 ? b : 0;
    }();
    return env ? env : kBulkDefaultBlocks;
}

template <bool UNPACK, bool WITH_SUM>
static int launch_bulk(const PlanDev &pd, uint8_t *image, uint64_t *sum, int grid, cudaStream_t stream) {
    auto kernel = pack_bulk_kernel<UNPACK, WITH_SUM>;
    static bool configured[64] = {};
    int dev = 0;
    PR_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        PR_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBulkSmemBytes));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    kernel<<<grid, kBulkThreads, kBulkSmemBytes, stream>>>(pd, image, sum);
    PR_CUDA(cudaGetLastError());
    pr_count_launch();
    return PR_OK;
}

template <bool UNPACK>
static int launch_pack(const pr_pack_plan *plan, void *image, uint64_t *sum, int algo, int max_blocks, void *stream_) {
    const char *who = UNPACK ? "pr_ckpt_unpack" : "pr_ckpt_pack";
    PR_REQUIRE(plan != nullptr, "%s: plan is NULL", who);
    PR_REQUIRE(algo == PR_PACK_AUTO || algo == PR_PACK_LDG || algo == PR_PACK_BULK, "%s: algo %d unknown", who, algo);
    PR_REQUIRE(max_blocks >= 0, "%s: max_blocks must be >= 0", who);
    if (plan->dev.ntiles == 0) return PR_OK;
    PR_REQUIRE(image != nullptr && pr_aligned16(image), "%s: image must be non-NULL and 16-B aligned", who);
    cudaStream_t stream = (cudaStream_t)stream_;
    bool bulk = algo == PR_PACK_BULK;
    if (algo != PR_PACK_LDG) {
        // bulk copies address device memory; an image in mapped pinned host memory takes the LDG/STG kernel
        cudaPointerAttributes attr;
        const bool on_device = cudaPointerGetAttributes(&attr, image) == cudaSuccess && attr.type == cudaMemoryTypeDevice;
        cudaGetLastError();
        if (algo == PR_PACK_BULK) {
            PR_REQUIRE(plan->bulk_ok, "%s: PR_PACK_BULK needs every tensor base 16-byte aligned", who);
            PR_REQUIRE(on_device, "%s: PR_PACK_BULK needs the image in device memory", who);
        } else {
            bulk = plan->bulk_ok && on_device;
        }
    }
    if (bulk) {
        uint64_t cap = (uint64_t)(max_blocks > 0 ? max_blocks : bulk_default_blocks());
        if (cap > (uint64_t)pr_sm_count()) cap = (uint64_t)pr_sm_count();     // 192 KiB of shared memory: one CTA per SM
        int grid = (int)(plan->dev.ntiles < cap ? plan->dev.ntiles : cap);
        return sum ? launch_bulk<UNPACK, true>(plan->dev, (uint8_t *)image, sum, grid, stream)
                   : launch_bulk<UNPACK, false>(plan->dev, (uint8_t *)image, nullptr, grid, stream);
    }
    uint64_t cap = (uint64_t)pr_sm_count() * ckpt_blocks_per_sm(true);
    if (max_blocks > 0 && (uint64_t)max_blocks < cap) cap = (uint64_t)max_blocks;
    int grid = (int)(plan->dev.ntiles < cap ? plan->dev.ntiles : cap);
    if (sum)
        pack_kernel<UNPACK, true><<<grid, kThreads, 0, stream>>>(plan->dev, (uint8_t *)image, sum);
    else
        pack_kernel<UNPACK, false><<<grid, kThreads, 0, stream>>>(plan->dev, (uint8_t *)image, nullptr);
    PR_CUDA(cudaGetLastError());
    pr_count_launch();
    return PR_OK;
}

extern "C" int pr_ckpt_pack(const pr_pack_plan *plan, void *image, uint64_t *sum, void *stream) {
    return launch_pack<false>(plan, image, sum, PR_PACK_AUTO, 0, stream);
}

extern "C" int pr_ckpt_unpack(const pr_pack_plan *plan, const void *image, uint64_t *sum, void *stream) {
    return launch_pack<true>(plan, const_cast<void *>(image), sum, PR_PACK_AUTO, 0, stream);
}

extern "C" int pr_ckpt_pack_ex(const pr_pack_plan *plan, void *image, uint64_t *sum, int algo, int max_blocks,
                               void *stream) {
    return launch_pack<false>(plan, image, sum, algo, max_blocks, stream);
}

extern "C" int pr_ckpt_unpack_ex(const pr_pack_plan *plan, const void *image, uint64_t *sum, int algo, int max_blocks,
                                 void *stream) {
    return launch_pack<true>(plan, const_cast<void *>(image), sum, algo, max_blocks, stream);
}

// ---- host arena + copies + file writer --------------------------------------------------
extern "C" int pr_host_arena_alloc(uint64_t nbytes, void **out) {
    PR_REQUIRE(out != nullptr, "pr_host_arena_alloc: out is NULL");
    *out = nullptr;
    if (nbytes == 0) return PR_OK;
    void *p = nullptr;
    PR_CUDA(cudaHostAlloc(&p, nbytes, cudaHostAllocPortable | cudaHostAllocMapped));
    memset(p, 0, nbytes);
    *out = p;
    return PR_OK;
}

extern "C" int pr_host_arena_free(void *p) {
    if (p) PR_CUDA(cudaFreeHost(p));
    return PR_OK;
}

extern "C" int pr_host_arena_device_ptr(void *host_ptr, void **dev_out) {
    PR_REQUIRE(host_ptr && dev_out, "pr_host_arena_device_ptr: NULL argument");
    PR_CUDA(cudaHostGetDevicePointer(dev_out, host_ptr, 0));
    return PR_OK;
}

static int copy_chunked(void *dst, const void *src, uint64_t nbytes, uint64_t chunk,
                        cudaMemcpyKind kind, void *stream_) {
    if (nbytes == 0) return PR_OK;
    PR_REQUIRE(dst && src, "pr_copy: NULL pointer");
    cudaStream_t stream = (cudaStream_t)stream_;
    if (chunk == 0) chunk = nbytes;
    for (uint64_t o = 0; o < nbytes; o += chunk) {
        uint64_t len = nbytes - o < chunk ? nbytes - o : chunk;
        PR_CUDA(cudaMemcpyAsync((uint8_t *)dst + o, (const uint8_t *)src + o, len, kind, stream));
    }
    return PR_OK;
}

extern "C" int pr_copy_d2h_async(void *dst_host, const void *src_dev, uint64_t nbytes,
                                 uint64_t chunk, void *stream) {
    return copy_chunked(dst_host, src_dev, nbytes, chunk, cudaMemcpyDeviceToHost, stream);
}

extern "C" int pr_copy_h2d_async(void *dst_dev, const void *src_host, uint64_t nbytes,
                                 uint64_t chunk, void *stream) {
    return copy_chunked(dst_dev, src_host, nbytes, chunk, cudaMemcpyHostToDevice, stream);
}

extern "C" int pr_file_write_atomic(const char *path, const void *const *pieces,
                                    const uint64_t *piece_bytes, int n, int do_fsync) {
    PR_REQUIRE(path && (n == 0 || (pieces && piece_bytes)), "pr_file_write_atomic: NULL argument");
    std::string tmp = std::string(path) + ".tmp";
    int fd = open(tmp.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) {
        pr_set_error("pr_file_write_atomic: open(%s): %s", tmp.c_str(), strerror(errno));
        return PR_ERR_IO;
    }
    const uint64_t kMaxWrite = 64ull << 20;
    for (int i = 0; i < n; ++i) {
        const uint8_t *p = (const uint8_t *)pieces[i];
        uint64_t left = piece_bytes[i];
        while (left) {
            ssize_t w = write(fd, p, left < kMaxWrite ? left : kMaxWrite);
            if (w < 0) {
                if (errno == EINTR) continue;
                pr_set_error("pr_file_write_atomic: write(%s): %s", tmp.c_str(), strerror(errno));
                close(fd);
                unlink(tmp.c_str());
                return PR_ERR_IO;
            }
            p += w;
            left -= (uint64_t)w;
        }
    }
    if (do_fsync && fsync(fd) != 0) {
        pr_set_error("pr_file_write_atomic: fsync(%s): %s", tmp.c_str(), strerror(errno));
        close(fd);
        unlink(tmp.c_str());
        return PR_ERR_IO;
    }
    close(fd);
    if (rename(tmp.c_str(), path) != 0) {
        pr_set_error("pr_file_write_atomic: rename(%s): %s", path, strerror(errno));
        unlink(tmp.c_str());
        return PR_ERR_IO;
    }
    return PR_OK;
}

// ---- pack without any allocation ------------------------------------------------------------
// Who needs this: a checkpoint is most valuable exactly when it is hardest to take -- the
// walltime / preemption handler of a job that has already filled HBM, or a process whose memory
// manager forbids cudaMalloc after start-up.  pr_pack_plan_create needs a device table (one
// cudaMalloc, which can fail with the device full and which synchronises the whole device);
// those callers get the same gather + PR-SUM64 here with the table kept in a fixed array that
// is part of the library image.
//
// The array below is that table.  It costs kNoAllocScratchBytes of global memory on every GPU
// that loads the library, for the life of the process.  In host code its NAME is only a host-side
// stand-in -- the device address comes from cudaGetSymbolAddress, looked up once per device with
// that device current, return code checked, and handed to the kernel as an ordinary pointer.
namespace {

constexpr int kNoAllocMaxTensors = 1024;
constexpr uint64_t kNoAllocMaxTiles = 65536;                       // 65536 x 64 KiB = 4 GiB of tiles
constexpr size_t kNoAllocSegBytes = sizeof(uint64_t) * kNoAllocMaxTensors;
constexpr size_t kNoAllocScratchBytes = 4 * kNoAllocSegBytes + sizeof(uint32_t) * kNoAllocMaxTiles;   // 288 KiB

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mutex;
void *g_noalloc_dev_ptr[kNoAllocMaxDevices] = {};                  // per-device address of the array
alignas(16) unsigned char g_noalloc_host_table[kNoAllocScratchBytes];   // host image of the table (no malloc)
int g_noalloc_order[kNoAllocMaxTensors];

}  // namespace

extern "C" int pr_ckpt_pack_noalloc(const void *const *tensor, const uint64_t *nbytes,
                                    const uint64_t *dst_off, int n, void *image, uint64_t image_bytes,
                                    uint64_t *sum, void *stream_) {
    PR_REQUIRE(n >= 0 && (n == 0 || (tensor && nbytes && dst_off)), "pr_ckpt_pack_noalloc: bad table");
    PR_REQUIRE(n <= kNoAllocMaxTensors,
               "pr_ckpt_pack_noalloc: %d tensors exceed the built-in table (max %d); use pr_pack_plan_create",
               n, kNoAllocMaxTensors);
    if (n == 0) return PR_OK;
    PR_REQUIRE(image != nullptr && pr_aligned16(image), "pr_ckpt_pack_noalloc: image must be non-NULL and 16-B aligned");
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);     // the host table and the device array are shared

    uint64_t *ptrs = reinterpret_cast<uint64_t *>(g_noalloc_host_table + 0 * kNoAllocSegBytes);
    uint64_t *offs = reinterpret_cast<uint64_t *>(g_noalloc_host_table + 1 * kNoAllocSegBytes);
    uint64_t *lens = reinterpret_cast<uint64_t *>(g_noalloc_host_table + 2 * kNoAllocSegBytes);
    uint64_t *first = reinterpret_cast<uint64_t *>(g_noalloc_host_table + 3 * kNoAllocSegBytes);
    uint32_t *tile_seg = reinterpret_cast<uint32_t *>(g_noalloc_host_table + 4 * kNoAllocSegBytes);
    int vrc = validate_slots("pr_ckpt_pack_noalloc", tensor, nbytes, dst_off, n, image_bytes, g_noalloc_order);
    if (vrc) return vrc;
    uint64_t ntiles = 0;
    for (int i = 0; i < n; ++i) {
        uint64_t nt = (nbytes[i] + kTileBytes - 1) / kTileBytes;
        PR_REQUIRE(nt <= kNoAllocMaxTiles && ntiles + nt <= kNoAllocMaxTiles,
                   "pr_ckpt_pack_noalloc: the tensors need more than %llu tiles of %llu KiB (about %llu GiB of payload); "
                   "use pr_pack_plan_create for larger images",
                   (unsigned long long)kNoAllocMaxTiles, (unsigned long long)(kTileBytes >> 10),
                   (unsigned long long)((kNoAllocMaxTiles * kTileBytes) >> 30));
        ptrs[i] = (uint64_t)(uintptr_t)tensor[i];
        offs[i] = dst_off[i];
        lens[i] = nbytes[i];
        first[i] = ntiles;
        for (uint64_t t = 0; t < nt; ++t) tile_seg[ntiles + t] = (uint32_t)i;
        ntiles += nt;
    }
    if (ntiles == 0) return PR_OK;

    // Set-up: resolve the device address of the array with this device current.  The first resolution on
    // a device is what loads the module and materialises the array there (CUDA loads modules lazily); the
    // address is remembered per device, and looked up again on every call so that a pointer from before a
    // cudaDeviceReset is never handed to the kernel (after the first time this is a table lookup).
    int dev = 0;
    PR_CUDA(cudaGetDevice(&dev));
    PR_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, "pr_ckpt_pack_noalloc: device ordinal %d unsupported", dev);
    {
        void *ptr = nullptr;
        PR_CUDA(cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch));
        PR_REQUIRE(ptr != nullptr, "pr_ckpt_pack_noalloc: cudaGetSymbolAddress returned NULL");
        g_noalloc_dev_ptr[dev] = ptr;
    }
    unsigned char *base = static_cast<unsigned char *>(g_noalloc_dev_ptr[dev]);

    cudaStream_t stream = (cudaStream_t)stream_;
    // Every entry the kernel reads ([0, n) of the four tensor arrays, [0, ntiles) of the tile map)
    // is rewritten here, so nothing in the array has to start at zero and whatever the previous
    // call left behind its own range is never looked at.  Pageable source: the call returns once
    // the host table has been consumed, so the next call may overwrite it.
    for (int part = 0; part < 4; ++part)
        PR_CUDA(cudaMemcpyAsync(base + part * kNoAllocSegBytes, g_noalloc_host_table + part * kNoAllocSegBytes,
                                sizeof(uint64_t) * (size_t)n, cudaMemcpyHostToDevice, stream));
    PR_CUDA(cudaMemcpyAsync(base + 4 * kNoAllocSegBytes, g_noalloc_host_table + 4 * kNoAllocSegBytes,
                            sizeof(uint32_t) * (size_t)ntiles, cudaMemcpyHostToDevice, stream));
    PlanDev pd;
    pd.seg_ptr = reinterpret_cast<const uint64_t *>(base + 0 * kNoAllocSegBytes);
    pd.seg_off = reinterpret_cast<const uint64_t *>(base + 1 * kNoAllocSegBytes);
    pd.seg_bytes = reinterpret_cast<const uint64_t *>(base + 2 * kNoAllocSegBytes);
    pd.tile_first = reinterpret_cast<const uint64_t *>(base + 3 * kNoAllocSegBytes);
    pd.tile_seg = reinterpret_cast<const uint32_t *>(base + 4 * kNoAllocSegBytes);
    pd.ntiles = ntiles;
    uint64_t cap = (uint64_t)pr_sm_count() * ckpt_blocks_per_sm(true);
    int grid = (int)(ntiles < cap ? ntiles : cap);
    if (sum)
        pack_kernel<false, true><<<grid, kThreads, 0, stream>>>(pd, (uint8_t *)image, sum);
    else
        pack_kernel<false, false><<<grid, kThreads, 0, stream>>>(pd, (uint8_t *)image, nullptr);
    PR_CUDA(cudaGetLastError());
    pr_count_launch();
    return PR_OK;
}
