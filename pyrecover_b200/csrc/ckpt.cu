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

struct pr_pack_plan {
    int device = -1;
    int n = 0;
    uint64_t image_bytes = 0;
    uint64_t payload_bytes = 0;
    void *table = nullptr;   // one device allocation holding all five arrays
    PlanDev dev{};
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
    uint64_t payload = 0;
    for (int i = 0; i < n; ++i) {
        PR_REQUIRE(dst_off[i] % 16 == 0, "pr_pack_plan_create: dst_off[%d]=%llu is not a multiple of 16",
                   i, (unsigned long long)dst_off[i]);
        uint64_t padded = (nbytes[i] + 15) / 16 * 16;
        PR_REQUIRE(dst_off[i] + padded <= image_bytes,
                   "pr_pack_plan_create: tensor %d (%llu B at %llu) overruns the image (%llu B)", i,
                   (unsigned long long)nbytes[i], (unsigned long long)dst_off[i],
                   (unsigned long long)image_bytes);
        PR_REQUIRE(nbytes[i] == 0 || tensor[i] != nullptr, "pr_pack_plan_create: tensor %d is NULL", i);
        ptrs[i] = (uint64_t)(uintptr_t)tensor[i];
        offs[i] = dst_off[i];
        lens[i] = nbytes[i];
        first[i] = tile_seg.size();
        uint64_t nt = (nbytes[i] + kTileBytes - 1) / kTileBytes;
        tile_seg.insert(tile_seg.end(), nt, (uint32_t)i);
        payload += nbytes[i];
    }
    // reject overlapping slots (sorted sweep)
    {
        std::vector<int> order(n);
        for (int i = 0; i < n; ++i) order[i] = i;
        std::sort(order.begin(), order.end(), [&](int a, int b) { return offs[a] < offs[b]; });
        for (int k = 1; k < n; ++k) {
            int a = order[k - 1], b = order[k];
            PR_REQUIRE(offs[a] + (lens[a] + 15) / 16 * 16 <= offs[b] || lens[a] == 0 || lens[b] == 0,
                       "pr_pack_plan_create: tensors %d and %d overlap in the image", a, b);
        }
    }
    auto *plan = new pr_pack_plan();
    plan->n = n;
    plan->image_bytes = image_bytes;
    plan->payload_bytes = payload;
    PR_CUDA(cudaGetDevice(&plan->device));
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

template <bool UNPACK>
static int launch_pack(const pr_pack_plan *plan, void *image, uint64_t *sum, void *stream_) {
    PR_REQUIRE(plan != nullptr, "pr_ckpt_%s: plan is NULL", UNPACK ? "unpack" : "pack");
    if (plan->dev.ntiles == 0) return PR_OK;
    PR_REQUIRE(image != nullptr && pr_aligned16(image), "pr_ckpt_%s: image must be non-NULL and 16-B aligned",
               UNPACK ? "unpack" : "pack");
    cudaStream_t stream = (cudaStream_t)stream_;
    uint64_t cap = (uint64_t)pr_sm_count() * ckpt_blocks_per_sm(true);
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
    return launch_pack<false>(plan, image, sum, stream);
}

extern "C" int pr_ckpt_unpack(const pr_pack_plan *plan, const void *image, uint64_t *sum, void *stream) {
    return launch_pack<true>(plan, const_cast<void *>(image), sum, stream);
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
    uint64_t ntiles = 0;
    for (int i = 0; i < n; ++i) {
        PR_REQUIRE(dst_off[i] % 16 == 0, "pr_ckpt_pack_noalloc: dst_off[%d]=%llu is not a multiple of 16", i,
                   (unsigned long long)dst_off[i]);
        uint64_t padded = (nbytes[i] + 15) / 16 * 16;
        PR_REQUIRE(dst_off[i] + padded <= image_bytes,
                   "pr_ckpt_pack_noalloc: tensor %d (%llu B at %llu) overruns the image (%llu B)", i,
                   (unsigned long long)nbytes[i], (unsigned long long)dst_off[i], (unsigned long long)image_bytes);
        PR_REQUIRE(nbytes[i] == 0 || tensor[i] != nullptr, "pr_ckpt_pack_noalloc: tensor %d is NULL", i);
        uint64_t nt = (nbytes[i] + kTileBytes - 1) / kTileBytes;
        PR_REQUIRE(ntiles + nt <= kNoAllocMaxTiles,
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
        g_noalloc_order[i] = i;
    }
    std::sort(g_noalloc_order, g_noalloc_order + n, [&](int a, int b) { return offs[a] < offs[b]; });
    for (int k = 1; k < n; ++k) {
        int a = g_noalloc_order[k - 1], b = g_noalloc_order[k];
        PR_REQUIRE(offs[a] + (lens[a] + 15) / 16 * 16 <= offs[b] || lens[a] == 0 || lens[b] == 0,
                   "pr_ckpt_pack_noalloc: tensors %d and %d overlap in the image", a, b);
    }
    if (ntiles == 0) return PR_OK;

    // Set-up, once per device: resolve the device address of the array with this device current.
    int dev = 0;
    PR_CUDA(cudaGetDevice(&dev));
    PR_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, "pr_ckpt_pack_noalloc: device ordinal %d unsupported", dev);
    if (g_noalloc_dev_ptr[dev] == nullptr) {
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
