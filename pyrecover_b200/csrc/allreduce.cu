// K1 -- peer-to-peer gradient allreduce over NVLink / NVSwitch for sm_100a, and its fusion
// with AdamW (K1+K2).
//
// Replaces the DDP Reducer's bucketed allreduce of the reference (train.py:110-115 builds
// DDP; the collective fires inside loss.backward(), train.py:269):
//     g <- sum_r g_r * (1/W)          (torch/csrc/distributed/c10d/reducer.hpp:500)
// and, fused, optimizer.step() (train.py:274).
//
// Design.  One process per GPU.  Every rank cudaMalloc's its symmetric regions and a block
// of flag words, exports CUDA-IPC handles (the caller ships the blobs with one all_gather on
// NCCL/gloo -- bootstrap only) and maps every peer.  The kernels then work on plain global
// pointers: 16-byte coalesced loads pull the same vector from all W ranks (W independent
// NVLink requests in flight per thread per unrolled vector), the sum is taken in fp32 in
// rank order 0..W-1 so every rank computes bit-identical results, and the rounded result is
// pushed to all W ranks with 16-byte stores.  Reads and pushes run in the same phase, so
// both NVLink directions carry (W-1)/W * N bytes each: 2(W-1)/W * N per direction in total,
// the bandwidth-optimal volume (roofline: measured 770 GB/s peer copy / 900 GB/s nominal
// per direction per GPU; bus GB/s = N/t * 2(W-1)/W).
//
// Cross-GPU ordering: CTA b of every rank meets CTA b of every other rank at a flag barrier:
// st.release.sys of the launch sequence number into the peer's flag word, ld.acquire.sys spin
// on the own word.  bar.sync before/after extends the release/acquire to the whole CTA.
// Slots 0 (begin) and 1 (end) alternate so a fast rank can never overwrite a flag its peer
// has not consumed.  Grids never exceed what is co-resident (<= 1 CTA per SM).
//
// A communicator can also be "emulated": all W ranks' regions live on one device and ONE
// cooperative launch runs the W ranks as blockIdx.y, with the same flag protocol.  That is
// how the kernels are tested on a single GPU (separate spinning launches on one GPU are not
// guaranteed to run concurrently).
//
// NVSwitch multicast (NVLS).  When the driver offers it, the symmetric regions are CUDA VMM
// allocations (shared as POSIX file descriptors instead of CUDA-IPC handles -- fabric.cu) bound
// to a multicast object, and two more kernels exist: rank r fetches slice r ALREADY SUMMED by
// the switch (multimem.ld_reduce, fp32 accumulation) and broadcasts the result with one
// multimem.st.  Per direction a GPU then moves N(1 + 1/W) bytes instead of 2N(W-1)/W: less
// than the P2P volume from W = 4 up (1.125 N vs 1.75 N at W = 8), more at W = 2 -- so the
// automatic choice is multicast for W >= 4 and P2P below.  The order in which the switch adds
// the W contributions is not specified: results are identical on every rank (one rank
// computes each slice) but equal to the rank-order fp32 sum only within summation-order
// tolerance, and the bf16 sum is rounded by the switch BEFORE the 1/W scale (exact for W a
// power of two).
#include <stdlib.h>
#include <unistd.h>

#include "adamw_math.cuh"
#include "fabric.cuh"

namespace {

constexpr int kThreads = 512;
constexpr int kMaxBlocks = 512;          // flag words are provisioned for this many CTAs
constexpr int kSlots = 2;
// default grid of the multicast kernels (tunable per call: max_blocks).  The switch does the adding, so
// a GPU only has to keep N/W of requests in flight: measured on 8 x B200, 16 CTAs are fastest up
// to 512 MiB and 32 above (845.8 GB/s bus at 2 GiB); 148 CTAs are SLOWER (714 GB/s)
constexpr int kMulticastBlocksSmall = 16, kMulticastBlocksLarge = 32;
constexpr uint64_t kMulticastLargeBytes = 1ull << 30;
constexpr int kOneShotVecPerThread = 4;  // register budget of the one-shot kernel
constexpr size_t kBarrierWords = (size_t)kSlots * kMaxBlocks * PR_MAX_WORLD;
constexpr size_t kAbortWord = kBarrierWords;          // set by a peer whose barrier timed out: everybody gives up
constexpr size_t kSignalWord = kBarrierWords + 1;     // a rank's posted signal (pr_comm_signal_post), read by every peer
constexpr size_t kFlagWords = kBarrierWords + 64;
constexpr uint32_t kBlobMagic = 0x50523142u;  // "PR1B"

struct CommDev {
    int world;
    int rank;                        // real communicator; emulated uses blockIdx.y
    int emulated;
    uint32_t seq;
    unsigned long long timeout_ns;
    int *err;                        // mapped host word: set to 1 when a barrier timed out
    int absent_rank;                 // emulated communicators only (tests): this rank never shows up; -1 = none
    unsigned long long *stats;       // device: [0] ns CTA 0 waited at the begin barrier, [1] launches, [2] ns CTA 0 ran
    unsigned long long *signal_host; // mapped host word: (launch sequence << 32) | rank 0's posted signal
    uint8_t *grad[PR_MAX_WORLD];     // region base on rank i as mapped here
    uint8_t *param[PR_MAX_WORLD];    // second region (fused kernel pushes parameters)
    uint8_t *mom1[PR_MAX_WORLD];     // exp_avg region (fused kernel; only the own rank's is touched)
    uint8_t *mom2[PR_MAX_WORLD];     // exp_avg_sq region
    uint32_t *flags[PR_MAX_WORLD];
    uint8_t *mc_grad;                // multicast mapping of the gradient region (NULL without NVLS)
    uint8_t *mc_param;               // multicast mapping of the parameter region
};

__device__ __forceinline__ unsigned long long globaltimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ int my_rank(const CommDev &c) { return c.emulated ? (int)blockIdx.y : c.rank; }

// CTA blockIdx.x of all ranks rendezvous.  Returns false when a peer did not show up within the
// timeout (or a peer already gave up): the error word is raised, every peer's abort word is set so
// that nobody else waits out its own timeout, and the caller must return WITHOUT touching any data
// -- a rank that reads gradients a peer never finished writing, updates its parameters from them and
// pushes the result into every peer's arena has destroyed the job's state by the time the host finds
// out.  The verdict is uniform over the CTA.
__device__ __forceinline__ bool xbarrier(const CommDev &c, int rank, int slot) {
    __syncthreads();
    int failed = 0;
    if ((int)threadIdx.x < c.world) {
        const int peer = threadIdx.x;
        const size_t idx = ((size_t)slot * kMaxBlocks + blockIdx.x) * PR_MAX_WORLD;
        __threadfence_system();
        st_release_sys(c.flags[peer] + idx + rank, c.seq);
        const uint32_t *mine = c.flags[rank] + idx + peer;
        const uint32_t *abort_word = c.flags[rank] + kAbortWord;
        unsigned long long t0 = 0;
        uint32_t spins = 0;
        while ((int32_t)(ld_acquire_sys(mine) - c.seq) < 0) {
            if ((++spins & 0x3ffu) == 0) {
                unsigned long long now = globaltimer();
                if (t0 == 0) t0 = now;
                else if (now - t0 > c.timeout_ns || ld_acquire_sys(abort_word) != 0) { failed = 1; break; }
            }
        }
    }
    failed = __syncthreads_or(failed);
    if (failed && (int)threadIdx.x < c.world) {
        if (threadIdx.x == 0) *c.err = 1;
        st_release_sys(c.flags[threadIdx.x] + kAbortWord, 1u);
    }
    return failed == 0;
}

// Begin barrier of every K1 kernel.  Besides the rendezvous: an emulated rank marked absent leaves
// at once (tests of the timeout path); CTA 0 adds the time it waited for the slowest peer to the
// communicator's statistics and forwards rank 0's signal word to this rank's host.
struct KernelClock { unsigned long long t_start; };

__device__ __forceinline__ bool k1_begin(const CommDev &c, int rank, KernelClock &clk) {
    if (c.emulated && c.absent_rank == rank) return false;
    const bool lead = blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0;   // one clock per launch (emulated: rank 0's CTA 0)
    if (lead) clk.t_start = globaltimer();
    int dead = 0;
    if (threadIdx.x == 0) dead = ld_acquire_sys(c.flags[rank] + kAbortWord) != 0;      // a peer gave up earlier
    if (__syncthreads_or(dead)) {
        if (threadIdx.x == 0) *c.err = 1;
        return false;
    }
    const bool ok = xbarrier(c, rank, 0);
    if (lead) {
        atomicAdd(c.stats + 0, globaltimer() - clk.t_start);
        atomicAdd(c.stats + 1, 1ull);
        if (ok && c.signal_host != nullptr) {
            const uint32_t sig = ld_acquire_sys(c.flags[0] + kSignalWord);
            *reinterpret_cast<volatile unsigned long long *>(c.signal_host) = ((unsigned long long)c.seq << 32) | sig;
        }
    }
    return ok;
}

__device__ __forceinline__ void k1_end(const CommDev &c, int rank, const KernelClock &clk) {
    xbarrier(c, rank, 1);
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) atomicAdd(c.stats + 2, globaltimer() - clk.t_start);
}

__global__ void signal_post_kernel(uint32_t *word, uint32_t value) { st_release_sys(word, value); }

// ---- element helpers ------------------------------------------------------------------
template <bool F32> struct Acc;
template <> struct Acc<false> {   // 8 bf16 per vector
    float a[8];
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = 0.0f;
    }
    __device__ __forceinline__ void add(const uint4 &v) {
        const uint32_t *w = &v.x;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            a[2 * k] = __fadd_rn(a[2 * k], bf_lo(w[k]));
            a[2 * k + 1] = __fadd_rn(a[2 * k + 1], bf_hi(w[k]));
        }
    }
    __device__ __forceinline__ uint4 finish(float scale) const {
        uint4 o;
        uint32_t *w = &o.x;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            w[k] = pack_bf2(__fmul_rn(a[2 * k], scale), __fmul_rn(a[2 * k + 1], scale));
        return o;
    }
};
template <> struct Acc<true> {    // 4 fp32 per vector
    float a[4];
    __device__ __forceinline__ void zero() { a[0] = a[1] = a[2] = a[3] = 0.0f; }
    __device__ __forceinline__ void add(const uint4 &v) {
        a[0] = __fadd_rn(a[0], __uint_as_float(v.x));
        a[1] = __fadd_rn(a[1], __uint_as_float(v.y));
        a[2] = __fadd_rn(a[2], __uint_as_float(v.z));
        a[3] = __fadd_rn(a[3], __uint_as_float(v.w));
    }
    __device__ __forceinline__ uint4 finish(float scale) const {
        return make_uint4(__float_as_uint(__fmul_rn(a[0], scale)), __float_as_uint(__fmul_rn(a[1], scale)),
                          __float_as_uint(__fmul_rn(a[2], scale)), __float_as_uint(__fmul_rn(a[3], scale)));
    }
};

__device__ __forceinline__ void slice_bounds(uint64_t nvec, int world, int rank, uint64_t &lo,
                                             uint64_t &hi) {
    uint64_t per = (nvec + world - 1) / world;
    lo = per * rank;
    hi = lo + per;
    if (lo > nvec) lo = nvec;
    if (hi > nvec) hi = nvec;
}

// ---- two-shot (reduce-scatter + push all-gather in one sweep) ----------------------------
template <int W, bool F32>
__global__ void __launch_bounds__(kThreads, 1)
allreduce_twoshot_kernel(CommDev c, uint64_t off, uint64_t nvec, float scale) {
    const int rank = my_rank(c);
    KernelClock clk;
    if (!k1_begin(c, rank, clk)) return;
    uint64_t lo, hi;
    slice_bounds(nvec, W, rank, lo, hi);
    const uint64_t stride = (uint64_t)gridDim.x * kThreads;
    constexpr int U = 2;
    for (uint64_t i = lo + (uint64_t)blockIdx.x * kThreads + threadIdx.x; i < hi; i += U * stride) {
        uint4 v[U][W];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t j = i + u * stride;
            if (j < hi) {
#pragma unroll
                for (int r = 0; r < W; ++r) v[u][r] = ld_stream(c.grad[r] + off + 16 * j);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t j = i + u * stride;
            if (j < hi) {
                Acc<F32> acc;
                acc.zero();
#pragma unroll
                for (int r = 0; r < W; ++r) acc.add(v[u][r]);
                const uint4 o = acc.finish(scale);
#pragma unroll
                for (int r = 0; r < W; ++r) st_stream(c.grad[r] + off + 16 * j, o);
            }
        }
    }
    k1_end(c, rank, clk);
}

// ---- one-shot (every rank reads everything; small messages) -------------------------------
template <int W, bool F32>
__global__ void __launch_bounds__(kThreads, 1)
allreduce_oneshot_kernel(CommDev c, uint64_t off, uint64_t nvec, float scale) {
    const int rank = my_rank(c);
    KernelClock clk;
    if (!k1_begin(c, rank, clk)) return;
    const uint64_t stride = (uint64_t)gridDim.x * kThreads;
    const uint64_t i0 = (uint64_t)blockIdx.x * kThreads + threadIdx.x;
    uint4 res[kOneShotVecPerThread];
#pragma unroll
    for (int k = 0; k < kOneShotVecPerThread; ++k) {
        const uint64_t j = i0 + k * stride;
        if (j < nvec) {
            uint4 v[W];
#pragma unroll
            for (int r = 0; r < W; ++r) v[r] = ld_stream(c.grad[r] + off + 16 * j);
            Acc<F32> acc;
            acc.zero();
#pragma unroll
            for (int r = 0; r < W; ++r) acc.add(v[r]);
            res[k] = acc.finish(scale);
        }
    }
    if (!xbarrier(c, rank, 1)) return;   // every rank has finished reading: the buffers may be overwritten
#pragma unroll
    for (int k = 0; k < kOneShotVecPerThread; ++k) {
        const uint64_t j = i0 + k * stride;
        if (j < nvec) st_stream(c.grad[rank] + off + 16 * j, res[k]);
    }
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) atomicAdd(c.stats + 2, globaltimer() - clk.t_start);
}

// ---- fused reduce-scatter -> AdamW -> parameter all-gather (bf16) ----------------------------
template <int W>
__global__ void __launch_bounds__(kThreads, 1)
allreduce_adamw_kernel(CommDev c, uint64_t goff, uint64_t nvec, AdamWConsts ac_) {
    const AdamWConsts ac = resolve_gscale(ac_);
    if (ac.skip) return;      // uniform over CTAs and ranks: nobody reaches the barrier
    const int rank = my_rank(c);
    const uint64_t poff = goff;
    uint4 *__restrict__ m = reinterpret_cast<uint4 *>(c.mom1[rank] + goff);
    uint4 *__restrict__ v = reinterpret_cast<uint4 *>(c.mom2[rank] + goff);
    const AdamWFast fast = make_adamw_fast(ac);
    KernelClock clk;
    if (!k1_begin(c, rank, clk)) return;
    uint64_t lo, hi;
    slice_bounds(nvec, W, rank, lo, hi);
    const uint64_t stride = (uint64_t)gridDim.x * kThreads;
    for (uint64_t i = lo + (uint64_t)blockIdx.x * kThreads + threadIdx.x; i < hi; i += stride) {
        uint4 G[W];
#pragma unroll
        for (int r = 0; r < W; ++r) G[r] = ld_stream(c.grad[r] + goff + 16 * i);
        uint4 P = ld_stream(c.param[rank] + poff + 16 * i);
        uint4 M = ld_stream(m + i), V = ld_stream(v + i);
        Acc<false> acc;
        acc.zero();
#pragma unroll
        for (int r = 0; r < W; ++r) acc.add(G[r]);
        adamw_vec8_fast(P, acc.a, M, V, fast);  // 1/W rides in ac.gscale
        st_stream(m + i, M);
        st_stream(v + i, V);
#pragma unroll
        for (int r = 0; r < W; ++r) st_stream(c.param[r] + poff + 16 * i, P);
    }
    k1_end(c, rank, clk);
}

// ---- all-gather of per-rank slices (push) ----------------------------------------------------
template <int W>
__global__ void __launch_bounds__(kThreads, 1)
allgather_slices_kernel(CommDev c, uint64_t off, uint64_t nvec) {
    const int rank = my_rank(c);
    KernelClock clk;
    if (!k1_begin(c, rank, clk)) return;
    uint64_t lo, hi;
    slice_bounds(nvec, W, rank, lo, hi);
    const uint64_t stride = (uint64_t)gridDim.x * kThreads;
    for (uint64_t i = lo + (uint64_t)blockIdx.x * kThreads + threadIdx.x; i < hi; i += stride) {
        const uint4 x = ld_stream(c.grad[rank] + off + 16 * i);
#pragma unroll
        for (int r = 0; r < W; ++r)
            if (r != rank) st_stream(c.grad[r] + off + 16 * i, x);
    }
    k1_end(c, rank, clk);
}

// ---- NVSwitch multicast kernels ------------------------------------------------------------------
// multimem.ld_reduce: one request, the switch reads the vector from every bound GPU and returns the
// sum (bf16x2 inputs accumulated in fp32).  multimem.st: one store, the switch writes every GPU.
__device__ __forceinline__ uint4 mm_ld_reduce_bf16(const void *mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
    return r;
}
__device__ __forceinline__ uint4 mm_ld_reduce_f32(const void *mc) {
    float a, b, c, d;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(a), "=f"(b), "=f"(c), "=f"(d) : "l"(mc) : "memory");
    return make_uint4(__float_as_uint(a), __float_as_uint(b), __float_as_uint(c), __float_as_uint(d));
}
__device__ __forceinline__ void mm_st(void *mc, const uint4 &v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(mc), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
                    "f"(__uint_as_float(v.w)) : "memory");
}
template <bool F32>
__device__ __forceinline__ uint4 scale_vec(const uint4 &v, float scale) {
    uint4 o;
    const uint32_t *w = &v.x;
    uint32_t *q = &o.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (F32) q[k] = __float_as_uint(__fmul_rn(__uint_as_float(w[k]), scale));
        else q[k] = pack_bf2(__fmul_rn(bf_lo(w[k]), scale), __fmul_rn(bf_hi(w[k]), scale));
    }
    return o;
}

template <bool F32>
__global__ void __launch_bounds__(kThreads, 1)
allreduce_mc_kernel(CommDev c, uint64_t off, uint64_t nvec, float scale) {
    const int rank = c.rank;
    KernelClock clk;
    if (!k1_begin(c, rank, clk)) return;
    uint64_t lo, hi;
    slice_bounds(nvec, c.world, rank, lo, hi);
    const uint64_t stride = (uint64_t)gridDim.x * kThreads;
    uint8_t *base = c.mc_grad + off;
    constexpr int U = 4;
    for (uint64_t i = lo + (uint64_t)blockIdx.x * kThreads + threadIdx.x; i < hi; i += U * stride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t j = i + u * stride;
            if (j < hi) v[u] = F32 ? mm_ld_reduce_f32(base + 16 * j) : mm_ld_reduce_bf16(base + 16 * j);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t j = i + u * stride;
            if (j < hi) mm_st(base + 16 * j, scale == 1.0f ? v[u] : scale_vec<F32>(v[u], scale));
        }
    }
    k1_end(c, rank, clk);
}

// fused: switch-reduced gradient slice -> AdamW on the local (p, m, v) slice -> multicast store of
// the updated parameters into every rank's parameter region
__global__ void __launch_bounds__(kThreads, 1)
allreduce_adamw_mc_kernel(CommDev c, uint64_t goff, uint64_t nvec, AdamWConsts ac_) {
    const AdamWConsts ac = resolve_gscale(ac_);
    if (ac.skip) return;      // uniform over CTAs and ranks: nobody reaches the barrier
    const int rank = c.rank;
    const uint8_t *__restrict__ p = c.param[rank] + goff;
    uint4 *__restrict__ m = reinterpret_cast<uint4 *>(c.mom1[rank] + goff);
    uint4 *__restrict__ v = reinterpret_cast<uint4 *>(c.mom2[rank] + goff);
    const uint8_t *mg = c.mc_grad + goff;
    uint8_t *mp = c.mc_param + goff;
    const AdamWFast fast = make_adamw_fast(ac);
    KernelClock clk;
    if (!k1_begin(c, rank, clk)) return;
    uint64_t lo, hi;
    slice_bounds(nvec, c.world, rank, lo, hi);
    const uint64_t stride = (uint64_t)gridDim.x * kThreads;
    constexpr int U = 2;
    for (uint64_t i = lo + (uint64_t)blockIdx.x * kThreads + threadIdx.x; i < hi; i += U * stride) {
        uint4 G[U], P[U], M[U], V[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t j = i + u * stride;
            if (j < hi) {
                G[u] = mm_ld_reduce_bf16(mg + 16 * j);
                P[u] = ld_stream(p + 16 * j);
                M[u] = ld_stream(m + j);
                V[u] = ld_stream(v + j);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t j = i + u * stride;
            if (j < hi) {
                float g[8];
                const uint32_t *w = &G[u].x;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    g[2 * k] = bf_lo(w[k]);
                    g[2 * k + 1] = bf_hi(w[k]);
                }
                adamw_vec8_fast(P[u], g, M[u], V[u], fast);   // 1/W rides in ac.gscale
                st_stream(m + j, M[u]);
                st_stream(v + j, V[u]);
                mm_st(mp + 16 * j, P[u]);
            }
        }
    }
    k1_end(c, rank, clk);
}

// all-gather of per-rank slices through the switch: one multicast store instead of W - 1 pushes
__global__ void __launch_bounds__(kThreads, 1)
allgather_slices_mc_kernel(CommDev c, uint64_t off, uint64_t nvec) {
    const int rank = c.rank;
    KernelClock clk;
    if (!k1_begin(c, rank, clk)) return;
    uint64_t lo, hi;
    slice_bounds(nvec, c.world, rank, lo, hi);
    const uint64_t stride = (uint64_t)gridDim.x * kThreads;
    for (uint64_t i = lo + (uint64_t)blockIdx.x * kThreads + threadIdx.x; i < hi; i += stride)
        mm_st(c.mc_grad + off + 16 * i, ld_stream(c.grad[rank] + off + 16 * i));
    k1_end(c, rank, clk);
}

struct Blob {
    uint32_t magic;
    int32_t rank;
    int32_t world;
    int32_t pid;
    uint64_t region_bytes[PR_MAX_REGIONS];
    cudaIpcMemHandle_t flags;
    cudaIpcMemHandle_t region[PR_MAX_REGIONS];   // unused for VMM regions (those travel as descriptors)
    int32_t vmm;                                 // regions are VMM allocations shared as file descriptors
    int32_t want_mc;                             // this rank can and wants to join multicast objects
    char sock[48];                               // abstract unix socket the descriptors are sent to
};
static_assert(sizeof(Blob) <= PR_EXPORT_BLOB_BYTES, "export blob too large");

struct FdMsg {                                   // one datagram per (sender, receiver) pair
    uint32_t magic;
    int32_t rank;
    int32_t n_region_fds;                        // descriptors of the sender's regions, in region order
    int32_t n_mc_fds;                            // rank 0 only: multicast objects, in region order (after the regions)
};

}  // namespace

struct pr_comm {
    int rank = 0, world = 1, device = 0;
    bool emulated = false, imported = false;
    uint32_t seq = 0;
    int *err_host = nullptr, *err_dev = nullptr;
    unsigned long long *signal_host = nullptr, *signal_dev = nullptr;   // mapped host word written by K1's begin barrier
    unsigned long long *stats_dev = nullptr;                             // 4 x u64 device counters (CommDev::stats)
    int absent_rank = -1;                                                // emulated only: test hook
    unsigned long long timeout_ns = 20ull * 1000000000ull;
    // [rank][region]; for a real comm only [self] is owned, the rest are IPC / VMM mappings
    uint8_t *region[PR_MAX_WORLD][PR_MAX_REGIONS] = {};
    uint64_t region_bytes[PR_MAX_REGIONS] = {};
    uint32_t *flags[PR_MAX_WORLD] = {};
    // NVSwitch multicast state (fabric.cu)
    bool vmm = false, want_mc = false, mc_ready = false;
    int mc_policy = -1;                          // -1 auto (W >= 4), 0 never, 1 whenever available
    PrFabricCaps caps;
    PrVmmBlock block[PR_MAX_WORLD][PR_MAX_REGIONS];
    CUmemGenericAllocationHandle mc_handle[PR_MAX_REGIONS] = {};
    CUdeviceptr mc_va[PR_MAX_REGIONS] = {};
    bool mc_bound[PR_MAX_REGIONS] = {};
    PrSock sock;
};

static bool use_multicast(const pr_comm *c) {
    if (!c->mc_ready || c->mc_policy == 0) return false;
    return c->mc_policy == 1 || c->world >= 4;
}

static int fill_dev(pr_comm *c, int grad_region, int param_region, CommDev &d, int m_region = -1,
                    int v_region = -1) {
    PR_REQUIRE(c->emulated || c->imported || c->world == 1, "pr_comm: peers not imported yet");
    d.world = c->world;
    d.rank = c->rank;
    d.emulated = c->emulated ? 1 : 0;
    d.seq = ++c->seq;
    d.timeout_ns = c->timeout_ns;
    d.err = c->err_dev;
    d.absent_rank = c->absent_rank;
    d.stats = c->stats_dev;
    d.signal_host = c->signal_dev;
    for (int r = 0; r < PR_MAX_WORLD; ++r) {
        d.grad[r] = r < c->world ? c->region[r][grad_region] : nullptr;
        d.param[r] = (r < c->world && param_region >= 0) ? c->region[r][param_region] : nullptr;
        d.mom1[r] = (r < c->world && m_region >= 0) ? c->region[r][m_region] : nullptr;
        d.mom2[r] = (r < c->world && v_region >= 0) ? c->region[r][v_region] : nullptr;
        d.flags[r] = r < c->world ? c->flags[r] : nullptr;
    }
    d.mc_grad = c->mc_ready ? (uint8_t *)c->mc_va[grad_region] : nullptr;
    d.mc_param = (c->mc_ready && param_region >= 0) ? (uint8_t *)c->mc_va[param_region] : nullptr;
    return PR_OK;
}

static int pick_grid(pr_comm *c, int max_blocks, const void *kernel, uint64_t work_vecs) {
    int cap = pr_sm_count();
    if (cap > kMaxBlocks) cap = kMaxBlocks;
    if (c->emulated) {
        int per_sm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kThreads, 0) != cudaSuccess || per_sm < 1)
            per_sm = 1;
        cap = (pr_sm_count() * per_sm) / c->world;
        if (cap > kMaxBlocks) cap = kMaxBlocks;
    }
    if (max_blocks > 0 && max_blocks < cap) cap = max_blocks;
    uint64_t need = (work_vecs + kThreads - 1) / kThreads;
    if (need < 1) need = 1;
    int grid = (int)(need < (uint64_t)cap ? need : (uint64_t)cap);
    return grid < 1 ? 1 : grid;
}

template <typename... Args>
static int launch(pr_comm *c, const void *kernel, int grid, cudaStream_t stream, Args... args) {
    void *argv[] = {(void *)&args...};
    dim3 g(grid, c->emulated ? c->world : 1, 1), b(kThreads, 1, 1);
    if (c->emulated) PR_CUDA(cudaLaunchCooperativeKernel(kernel, g, b, argv, 0, stream));
    else PR_CUDA(cudaLaunchKernel(kernel, g, b, argv, 0, stream));
    pr_count_launch();
    return PR_OK;
}

#define PR_DISPATCH_W(world, ...)                                    \
    switch (world) {                                                 \
        case 1: { constexpr int W_ = 1; __VA_ARGS__; } break;               \
        case 2: { constexpr int W_ = 2; __VA_ARGS__; } break;               \
        case 3: { constexpr int W_ = 3; __VA_ARGS__; } break;               \
        case 4: { constexpr int W_ = 4; __VA_ARGS__; } break;               \
        case 5: { constexpr int W_ = 5; __VA_ARGS__; } break;               \
        case 6: { constexpr int W_ = 6; __VA_ARGS__; } break;               \
        case 7: { constexpr int W_ = 7; __VA_ARGS__; } break;               \
        case 8: { constexpr int W_ = 8; __VA_ARGS__; } break;               \
        default: pr_set_error("world %d unsupported", world); return PR_ERR_INVALID; \
    }

static void comm_timeout_from_env(pr_comm *c) {
    if (const char *e = getenv("PR_COMM_TIMEOUT_S")) {
        double s = atof(e);
        if (s > 0) c->timeout_ns = (unsigned long long)(s * 1e9);
    }
}

// Host words, counters and flag blocks of a communicator.  Any failure is returned to the caller,
// which destroys the half-built object (pr_comm_destroy copes with whatever is still NULL).
static int comm_init_common(pr_comm *c) {
    PR_CUDA(cudaGetDevice(&c->device));
    PR_CUDA(cudaHostAlloc((void **)&c->err_host, sizeof(int), cudaHostAllocMapped));
    *c->err_host = 0;
    PR_CUDA(cudaHostGetDevicePointer((void **)&c->err_dev, c->err_host, 0));
    PR_CUDA(cudaHostAlloc((void **)&c->signal_host, sizeof(unsigned long long), cudaHostAllocMapped));
    *c->signal_host = 0;
    PR_CUDA(cudaHostGetDevicePointer((void **)&c->signal_dev, c->signal_host, 0));
    PR_CUDA(cudaMalloc((void **)&c->stats_dev, 4 * sizeof(unsigned long long)));
    PR_CUDA(cudaMemset(c->stats_dev, 0, 4 * sizeof(unsigned long long)));
    for (int r = 0; r < c->world; ++r) {
        if (!c->emulated && r != c->rank) continue;
        PR_CUDA(cudaMalloc((void **)&c->flags[r], kFlagWords * sizeof(uint32_t)));
        PR_CUDA(cudaMemset(c->flags[r], 0, kFlagWords * sizeof(uint32_t)));
    }
    return PR_OK;
}

extern "C" int pr_comm_create(int rank, int world, pr_comm **out) {
    PR_REQUIRE(out != nullptr, "pr_comm_create: comm_out is NULL");
    *out = nullptr;
    PR_REQUIRE(world >= 1 && world <= PR_MAX_WORLD, "pr_comm_create: world %d not in [1, %d]", world, PR_MAX_WORLD);
    PR_REQUIRE(rank >= 0 && rank < world, "pr_comm_create: rank %d not in [0, %d)", rank, world);
    auto *c = new pr_comm();
    c->rank = rank;
    c->world = world;
    comm_timeout_from_env(c);
    if (int rc = comm_init_common(c)) {
        pr_comm_destroy(c);
        return rc;
    }
    // NVSwitch multicast needs the regions to be VMM allocations; PR_COMM_MULTICAST=0 keeps the
    // cudaMalloc + CUDA-IPC regions (P2P kernels only), =1 uses multicast whenever it is available,
    // unset = multicast from W = 4 up
    const char *mc_env = getenv("PR_COMM_MULTICAST");
    if (mc_env && *mc_env) c->mc_policy = atoi(mc_env) != 0 ? 1 : 0;
    if (world > 1 && c->mc_policy != 0) {
        c->caps = pr_fabric_caps(c->device, world);
        c->vmm = c->caps.driver && c->caps.posix_fd && c->caps.multicast;
        c->want_mc = c->vmm;
        if (c->vmm && pr_sock_open(&c->sock, rank) != PR_OK) c->vmm = c->want_mc = false;
        if (getenv("PR_COMM_DEBUG"))
            fprintf(stderr, "[pr_comm rank %d] driver %d posix_fd %d multicast %d alloc_gran %zu mc_gran min %zu rec %zu -> vmm %d\n",
                    rank, (int)c->caps.driver, (int)c->caps.posix_fd, (int)c->caps.multicast, c->caps.alloc_gran,
                    c->caps.mc_gran_min, c->caps.mc_gran_rec, (int)c->vmm);
    }
    *out = c;
    return PR_OK;
}

extern "C" int pr_comm_create_emulated(int world, pr_comm **out) {
    PR_REQUIRE(out != nullptr, "pr_comm_create_emulated: comm_out is NULL");
    *out = nullptr;
    PR_REQUIRE(world >= 1 && world <= PR_MAX_WORLD, "pr_comm_create_emulated: world %d not in [1, %d]", world, PR_MAX_WORLD);
    auto *c = new pr_comm();
    c->world = world;
    c->emulated = true;
    c->timeout_ns = 5ull * 1000000000ull;
    comm_timeout_from_env(c);
    if (int rc = comm_init_common(c)) {
        pr_comm_destroy(c);
        return rc;
    }
    *out = c;
    return PR_OK;
}

extern "C" int pr_comm_destroy(pr_comm *c) {
    if (!c) return PR_OK;
    cudaDeviceSynchronize();
    for (int k = 0; k < PR_MAX_REGIONS; ++k)
        if (c->mc_handle[k] || c->mc_va[k])
            pr_mc_detach(c->mc_handle[k], c->device, c->mc_va[k], c->block[c->rank][k].size, c->mc_bound[k]);
    for (int r = 0; r < c->world; ++r) {
        const bool own = c->emulated || r == c->rank;
        for (int k = 0; k < PR_MAX_REGIONS; ++k) {
            if (!c->region[r][k]) continue;
            if (c->vmm) pr_vmm_free(&c->block[r][k]);
            else if (own) cudaFree(c->region[r][k]);
            else cudaIpcCloseMemHandle(c->region[r][k]);
        }
        if (c->flags[r]) {
            if (own) cudaFree(c->flags[r]);
            else cudaIpcCloseMemHandle(c->flags[r]);
        }
    }
    pr_sock_close(&c->sock);
    if (c->err_host) cudaFreeHost(c->err_host);
    if (c->signal_host) cudaFreeHost(c->signal_host);
    if (c->stats_dev) cudaFree(c->stats_dev);
    delete c;
    return PR_OK;
}

extern "C" int pr_comm_region_alloc(pr_comm *c, int region, uint64_t nbytes) {
    PR_REQUIRE(c != nullptr, "pr_comm_region_alloc: comm is NULL");
    PR_REQUIRE(region >= 0 && region < PR_MAX_REGIONS, "pr_comm_region_alloc: region %d out of range", region);
    PR_REQUIRE(!c->imported, "pr_comm_region_alloc: regions must be allocated before pr_comm_import");
    PR_REQUIRE(c->region_bytes[region] == 0, "pr_comm_region_alloc: region %d already allocated", region);
    PR_REQUIRE(nbytes > 0 && nbytes % 16 == 0, "pr_comm_region_alloc: nbytes must be a positive multiple of 16");
    for (int r = 0; r < c->world; ++r) {
        if (!c->emulated && r != c->rank) continue;
        if (c->vmm) {
            const size_t rounded = pr_fabric_round(c->caps, nbytes);
            int rc = pr_vmm_alloc(c->device, rounded, &c->block[r][region]);
            if (rc) return rc;
            c->region[r][region] = (uint8_t *)c->block[r][region].va;
            PR_CUDA(cudaMemset(c->region[r][region], 0, rounded));
        } else {
            PR_CUDA(cudaMalloc((void **)&c->region[r][region], nbytes));
            PR_CUDA(cudaMemset(c->region[r][region], 0, nbytes));
        }
    }
    c->region_bytes[region] = nbytes;
    return PR_OK;
}

extern "C" int pr_comm_region_ptr(pr_comm *c, int region, int rank, void **out) {
    PR_REQUIRE(c && out, "pr_comm_region_ptr: NULL argument");
    PR_REQUIRE(region >= 0 && region < PR_MAX_REGIONS && c->region_bytes[region], "pr_comm_region_ptr: region %d not allocated", region);
    int r = c->emulated ? rank : c->rank;
    PR_REQUIRE(r >= 0 && r < c->world, "pr_comm_region_ptr: rank %d out of range", r);
    *out = c->region[r][region];
    return PR_OK;
}

extern "C" int pr_comm_export(pr_comm *c, void *blob_out) {
    PR_REQUIRE(c && blob_out, "pr_comm_export: NULL argument");
    PR_REQUIRE(!c->emulated, "pr_comm_export: emulated communicators have nothing to export");
    Blob b;
    memset(&b, 0, sizeof(b));
    b.magic = kBlobMagic;
    b.rank = c->rank;
    b.world = c->world;
    b.pid = (int32_t)getpid();
    PR_CUDA(cudaIpcGetMemHandle(&b.flags, c->flags[c->rank]));
    b.vmm = c->vmm ? 1 : 0;
    b.want_mc = c->want_mc ? 1 : 0;
    memcpy(b.sock, c->sock.name, sizeof(b.sock));
    for (int k = 0; k < PR_MAX_REGIONS; ++k) {
        b.region_bytes[k] = c->region_bytes[k];
        if (c->region_bytes[k] && !c->vmm) PR_CUDA(cudaIpcGetMemHandle(&b.region[k], c->region[c->rank][k]));
    }
    memset(blob_out, 0, PR_EXPORT_BLOB_BYTES);
    memcpy(blob_out, &b, sizeof(b));
    return PR_OK;
}

// VMM regions: every rank sends the descriptors of its regions to every peer over the unix
// sockets named in the blobs (rank 0 adds the multicast objects it created), then maps what it
// received and joins the multicast team.  cuMulticastBindMem blocks until all W devices have been
// added, so returning from here means the whole team is bound.
static int import_vmm(pr_comm *c, const Blob *blobs) {
    int nreg = 0, reg_ids[PR_MAX_REGIONS];
    for (int k = 0; k < PR_MAX_REGIONS; ++k)
        if (c->region_bytes[k]) reg_ids[nreg++] = k;
    if (nreg == 0) return PR_OK;                 // nothing to share (every rank sees the same region set)
    bool mc = true;
    for (int r = 0; r < c->world; ++r) mc = mc && blobs[r].want_mc;
    int fds[2 * PR_MAX_REGIONS], nfds = 0;
    for (int i = 0; i < nreg; ++i) {
        int rc = pr_vmm_export_fd(c->block[c->rank][reg_ids[i]], &fds[nfds]);
        if (rc) return rc;
        ++nfds;
    }
    FdMsg out;
    out.magic = kBlobMagic;
    out.rank = c->rank;
    out.n_region_fds = nreg;
    out.n_mc_fds = 0;
    if (mc && c->rank == 0) {
        for (int i = 0; i < nreg; ++i) {
            const int k = reg_ids[i];
            int fd = -1;
            if (pr_mc_create(c->world, c->block[0][k].size, &c->mc_handle[k], &fd) != PR_OK) {
                for (int j = 0; j < i; ++j) {            // multicast stays off for the whole team
                    close(fds[nreg + j]);
                    pr_mc_detach(c->mc_handle[reg_ids[j]], c->device, 0, 0, false);
                    c->mc_handle[reg_ids[j]] = 0;
                }
                nfds = nreg;
                out.n_mc_fds = 0;
                break;
            }
            fds[nfds++] = fd;
            out.n_mc_fds = i + 1;
        }
    }
    int rc = PR_OK;
    for (int r = 0; r < c->world && rc == PR_OK; ++r)
        if (r != c->rank) rc = pr_sock_send(c->sock, blobs[r].sock, &out, sizeof(out), fds, nfds);
    for (int i = 0; i < nfds; ++i) close(fds[i]);
    if (rc) return rc;
    int n_mc_seen = c->rank == 0 ? out.n_mc_fds : -1;
    for (int got = 0; got < c->world - 1; ++got) {
        FdMsg in;
        int rfds[16], nr = 0;
        if ((rc = pr_sock_recv(c->sock, &in, sizeof(in), rfds, 16, &nr, 120 * 1000))) return rc;
        bool ok = in.magic == kBlobMagic && in.rank >= 0 && in.rank < c->world && in.rank != c->rank &&
                  in.n_region_fds == nreg && nr == in.n_region_fds + in.n_mc_fds && !c->region[in.rank][reg_ids[0]];
        for (int i = 0; ok && i < nreg; ++i) {
            const int k = reg_ids[i];
            rc = pr_vmm_import_fd(c->device, rfds[i], c->block[c->rank][k].size, &c->block[in.rank][k]);
            if (rc) ok = false;
            else c->region[in.rank][k] = (uint8_t *)c->block[in.rank][k].va;
        }
        if (ok && in.rank == 0) {
            n_mc_seen = in.n_mc_fds;
            for (int i = 0; ok && i < in.n_mc_fds && in.n_mc_fds == nreg; ++i)
                if ((rc = pr_mc_import(rfds[nreg + i], &c->mc_handle[reg_ids[i]]))) ok = false;
        }
        for (int i = 0; i < nr; ++i) close(rfds[i]);
        if (!ok) {
            if (rc == PR_OK) {
                pr_set_error("pr_comm_import: malformed descriptor message from rank %d", in.rank);
                rc = PR_ERR_STATE;
            }
            return rc;
        }
    }
    if (mc && n_mc_seen == nreg && nreg > 0) {
        for (int i = 0; i < nreg; ++i) {
            const int k = reg_ids[i];
            if ((rc = pr_mc_attach(c->mc_handle[k], c->device, c->block[c->rank][k], &c->mc_va[k]))) return rc;
            c->mc_bound[k] = true;
        }
        c->mc_ready = true;
    }
    if (getenv("PR_COMM_DEBUG"))
        fprintf(stderr, "[pr_comm rank %d] %d VMM regions mapped from %d peers, multicast %s\n", c->rank, nreg, c->world - 1,
                c->mc_ready ? "bound" : "off");
    return PR_OK;
}

extern "C" int pr_comm_import(pr_comm *c, const void *blobs) {
    PR_REQUIRE(c && blobs, "pr_comm_import: NULL argument");
    PR_REQUIRE(!c->emulated, "pr_comm_import: emulated communicators need no import");
    PR_REQUIRE(!c->imported, "pr_comm_import: already imported");
    Blob all[PR_MAX_WORLD];
    for (int r = 0; r < c->world; ++r) {
        Blob &b = all[r];
        memcpy(&b, (const uint8_t *)blobs + (size_t)r * PR_EXPORT_BLOB_BYTES, sizeof(b));
        PR_REQUIRE(b.magic == kBlobMagic && b.rank == r && b.world == c->world,
                   "pr_comm_import: blob %d is malformed (magic %08x rank %d world %d)", r, b.magic, b.rank, b.world);
        for (int k = 0; k < PR_MAX_REGIONS; ++k)
            PR_REQUIRE(b.region_bytes[k] == c->region_bytes[k],
                       "pr_comm_import: rank %d region %d is %llu B here but %llu B there", r, k,
                       (unsigned long long)c->region_bytes[k], (unsigned long long)b.region_bytes[k]);
        PR_REQUIRE((b.vmm != 0) == c->vmm, "pr_comm_import: rank %d %s VMM regions but this rank %s (set PR_COMM_MULTICAST "
                   "identically on every rank)", r, b.vmm ? "uses" : "does not use", c->vmm ? "does" : "does not");
        b.sock[sizeof(b.sock) - 1] = 0;
    }
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank) continue;
        PR_CUDA(cudaIpcOpenMemHandle((void **)&c->flags[r], all[r].flags, cudaIpcMemLazyEnablePeerAccess));
        if (c->vmm) continue;
        for (int k = 0; k < PR_MAX_REGIONS; ++k)
            if (c->region_bytes[k])
                PR_CUDA(cudaIpcOpenMemHandle((void **)&c->region[r][k], all[r].region[k], cudaIpcMemLazyEnablePeerAccess));
    }
    if (c->vmm && c->world > 1) {
        int rc = import_vmm(c, all);
        if (rc) return rc;
    }
    c->imported = true;
    return PR_OK;
}

extern "C" int pr_comm_multicast(pr_comm *c, int policy, int *available, int *in_use) {
    PR_REQUIRE(c != nullptr, "pr_comm_multicast: comm is NULL");
    PR_REQUIRE(policy >= -2 && policy <= 1, "pr_comm_multicast: policy must be -2 (query), -1 (auto), 0 (off) or 1 (on)");
    if (policy != -2) c->mc_policy = policy;
    if (available) *available = c->mc_ready ? 1 : 0;
    if (in_use) *in_use = use_multicast(c) ? 1 : 0;
    return PR_OK;
}

static int check_range(pr_comm *c, int region, uint64_t offset, uint64_t nbytes, const char *who) {
    PR_REQUIRE(c != nullptr, "%s: comm is NULL", who);
    PR_REQUIRE(region >= 0 && region < PR_MAX_REGIONS && c->region_bytes[region], "%s: region %d not allocated", who, region);
    PR_REQUIRE(offset % 16 == 0 && nbytes % 16 == 0, "%s: offset and nbytes must be multiples of 16", who);
    PR_REQUIRE(offset + nbytes <= c->region_bytes[region], "%s: range exceeds region %d", who, region);
    if (*c->err_host) {
        pr_set_error("%s: a previous collective timed out waiting for a peer", who);
        return PR_ERR_STATE;
    }
    return PR_OK;
}

extern "C" int pr_allreduce(pr_comm *c, int region, uint64_t offset, uint64_t nbytes, int dtype,
                            float scale, int algo, int max_blocks, void *stream_) {
    int rc = check_range(c, region, offset, nbytes, "pr_allreduce");
    if (rc) return rc;
    PR_REQUIRE(dtype == PR_BF16 || dtype == PR_F32, "pr_allreduce: dtype %d unsupported", dtype);
    if (nbytes == 0) return PR_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    const uint64_t nvec = nbytes / 16;
    const bool f32 = dtype == PR_F32;
    bool one_shot = algo == PR_AR_ONE_SHOT;
    if (algo == PR_AR_AUTO) one_shot = nbytes <= (512u << 10);
    CommDev d;
    if ((rc = fill_dev(c, region, -1, d))) return rc;
    if (one_shot) {
        const void *k = nullptr;
        PR_DISPATCH_W(c->world, k = f32 ? (const void *)allreduce_oneshot_kernel<W_, true>
                                        : (const void *)allreduce_oneshot_kernel<W_, false>);
        int grid = pick_grid(c, max_blocks, k, nvec);
        if ((uint64_t)grid * kThreads * kOneShotVecPerThread < nvec) {
            PR_REQUIRE(algo == PR_AR_AUTO, "pr_allreduce: %llu B is too large for the one-shot kernel with %d CTAs",
                       (unsigned long long)nbytes, grid);
            one_shot = false;
        } else {
            return launch(c, k, grid, stream, d, offset, nvec, scale);
        }
    }
    const void *k = nullptr;
    if (algo == PR_AR_MULTIMEM || (algo == PR_AR_AUTO && use_multicast(c))) {
        if (!c->mc_ready) {
            pr_set_error("pr_allreduce: PR_AR_MULTIMEM asked for, but this communicator has no NVSwitch multicast mapping");
            return PR_ERR_UNSUPPORTED;
        }
        k = f32 ? (const void *)allreduce_mc_kernel<true> : (const void *)allreduce_mc_kernel<false>;
        // the switch does the summing: a few dozen CTAs keep the links full (4 vectors in flight per thread)
        const int def_blocks = nbytes >= kMulticastLargeBytes ? kMulticastBlocksLarge : kMulticastBlocksSmall;
        int grid = pick_grid(c, max_blocks > 0 ? max_blocks : def_blocks, k, (nvec + c->world - 1) / c->world / 4 + 1);
        return launch(c, k, grid, stream, d, offset, nvec, scale);
    }
    PR_DISPATCH_W(c->world, k = f32 ? (const void *)allreduce_twoshot_kernel<W_, true>
                                    : (const void *)allreduce_twoshot_kernel<W_, false>);
    int grid = pick_grid(c, max_blocks, k, (nvec + c->world - 1) / c->world);
    return launch(c, k, grid, stream, d, offset, nvec, scale);
}

extern "C" int pr_allreduce_adamw(pr_comm *c, int grad_region, int param_region, int m_region,
                                  int v_region, uint64_t elem_offset, uint64_t n, int dtype,
                                  const pr_adamw_args *a, int max_blocks, void *stream_) {
    PR_REQUIRE(dtype == PR_BF16, "pr_allreduce_adamw: only bf16 arenas are supported");
    PR_REQUIRE(a != nullptr && a->numerics == 0 && a->step >= 1, "pr_allreduce_adamw: bad args (numerics must be 0, step >= 1)");
    PR_REQUIRE(elem_offset % 8 == 0 && n % 8 == 0, "pr_allreduce_adamw: elem_offset and n must be multiples of 8");
    const int regs[4] = {grad_region, param_region, m_region, v_region};
    for (int i = 0; i < 4; ++i) {
        int rc = check_range(c, regs[i], elem_offset * 2, n * 2, "pr_allreduce_adamw");
        if (rc) return rc;
        for (int j = 0; j < i; ++j)
            PR_REQUIRE(regs[i] != regs[j], "pr_allreduce_adamw: the four regions must be distinct");
    }
    if (n == 0) return PR_OK;
    CommDev d;
    int rc = fill_dev(c, grad_region, param_region, d, m_region, v_region);
    if (rc) return rc;
    AdamWConsts ac = make_adamw_consts(a);
    const uint64_t nvec = n / 8;
    const void *k = nullptr;
    uint64_t goff = elem_offset * 2;
    if (use_multicast(c)) {
        k = (const void *)allreduce_adamw_mc_kernel;
        int grid = pick_grid(c, max_blocks, k, (nvec + c->world - 1) / c->world);
        return launch(c, k, grid, (cudaStream_t)stream_, d, goff, nvec, ac);
    }
    PR_DISPATCH_W(c->world, k = (const void *)allreduce_adamw_kernel<W_>);
    int grid = pick_grid(c, max_blocks, k, (nvec + c->world - 1) / c->world);
    return launch(c, k, grid, (cudaStream_t)stream_, d, goff, nvec, ac);
}

extern "C" int pr_allgather_slices(pr_comm *c, int region, uint64_t offset, uint64_t nbytes,
                                   int max_blocks, void *stream_) {
    int rc = check_range(c, region, offset, nbytes, "pr_allgather_slices");
    if (rc) return rc;
    if (nbytes == 0) return PR_OK;
    CommDev d;
    if ((rc = fill_dev(c, region, -1, d))) return rc;
    const uint64_t nvec = nbytes / 16;
    const void *k = nullptr;
    if (use_multicast(c)) {
        k = (const void *)allgather_slices_mc_kernel;
        int grid = pick_grid(c, max_blocks > 0 ? max_blocks : kMulticastBlocksLarge, k, (nvec + c->world - 1) / c->world);
        return launch(c, k, grid, (cudaStream_t)stream_, d, offset, nvec);
    }
    PR_DISPATCH_W(c->world, k = (const void *)allgather_slices_kernel<W_>);
    int grid = pick_grid(c, max_blocks, k, (nvec + c->world - 1) / c->world);
    return launch(c, k, grid, (cudaStream_t)stream_, d, offset, nvec);
}

extern "C" int pr_comm_error(pr_comm *c, int *timed_out) {
    PR_REQUIRE(c && timed_out, "pr_comm_error: NULL argument");
    *timed_out = *c->err_host;
    return PR_OK;
}

// ---- statistics, signal word, test hook ------------------------------------------------------------
extern "C" int pr_comm_stats(pr_comm *c, uint64_t *out3, int reset) {
    PR_REQUIRE(c && out3, "pr_comm_stats: NULL argument");
    unsigned long long h[4] = {0, 0, 0, 0};
    PR_CUDA(cudaMemcpy(h, c->stats_dev, sizeof(h), cudaMemcpyDeviceToHost));      // synchronises: call it off the hot loop
    out3[0] = h[0];
    out3[1] = h[1];
    out3[2] = h[2];
    if (reset) PR_CUDA(cudaMemset(c->stats_dev, 0, sizeof(h)));
    return PR_OK;
}

extern "C" int pr_comm_signal_post(pr_comm *c, uint32_t value, void *stream_) {
    PR_REQUIRE(c != nullptr, "pr_comm_signal_post: comm is NULL");
    const int r = c->emulated ? 0 : c->rank;
    PR_REQUIRE(c->flags[r] != nullptr, "pr_comm_signal_post: communicator has no flag block");
    signal_post_kernel<<<1, 1, 0, (cudaStream_t)stream_>>>(c->flags[r] + kSignalWord, value);
    PR_CUDA(cudaGetLastError());
    pr_count_launch();
    return PR_OK;
}

extern "C" int pr_comm_signal_read(pr_comm *c, uint32_t *seq_out, uint32_t *value_out) {
    PR_REQUIRE(c && seq_out && value_out, "pr_comm_signal_read: NULL argument");
    const unsigned long long w = *reinterpret_cast<volatile unsigned long long *>(c->signal_host);
    *seq_out = (uint32_t)(w >> 32);
    *value_out = (uint32_t)(w & 0xffffffffull);
    return PR_OK;
}

extern "C" int pr_comm_seq(pr_comm *c, uint32_t *seq_out) {
    PR_REQUIRE(c && seq_out, "pr_comm_seq: NULL argument");
    *seq_out = c->seq;
    return PR_OK;
}

extern "C" int pr_comm_test_absent_rank(pr_comm *c, int rank) {
    PR_REQUIRE(c != nullptr && c->emulated, "pr_comm_test_absent_rank: only emulated communicators (a test aid)");
    PR_REQUIRE(rank >= -1 && rank < c->world, "pr_comm_test_absent_rank: rank %d out of range", rank);
    c->absent_rank = rank;
    return PR_OK;
}
