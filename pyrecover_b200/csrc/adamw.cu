// K2 -- flat-arena AdamW for sm_100a.
//
// Replaces optimizer.step() of the reference (train.py:274; ctor train.py:120-122), i.e.
// torch/optim/adam.py:347-551 and torch._fused_adamw_.  The whole model is ONE launch over
// four contiguous arenas (p, g, m, v): no per-tensor launches, no multi-tensor-apply
// metadata, no bf16 temporaries.
//
// Roofline: HBM.  Algorithmic bytes = 14 B per bf16 element (read p,g,m,v; write p,m,v),
// 28 B per fp32 element.  Each thread owns 16-byte vectors (8 bf16 / 4 fp32 elements) of
// the four arenas, UNROLL vectors per grid-stride iteration, so 4*UNROLL independent 16-B
// loads are in flight per thread; the grid is a whole number of CTAs per SM.
#include "adamw_math.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 2;
constexpr int kMinBlocks = 4;    // <= 64 registers per thread: 32 resident warps per SM

template <bool GRAD_F32>
__device__ __forceinline__ void load_grad8(const void *g_, uint64_t j, float (&g)[8]) {
    if (GRAD_F32) {
        const uint4 *g4 = reinterpret_cast<const uint4 *>(g_) + 2 * j;
        uint4 a = ld_stream_nc(g4), b = ld_stream_nc(g4 + 1);
        g[0] = __uint_as_float(a.x); g[1] = __uint_as_float(a.y);
        g[2] = __uint_as_float(a.z); g[3] = __uint_as_float(a.w);
        g[4] = __uint_as_float(b.x); g[5] = __uint_as_float(b.y);
        g[6] = __uint_as_float(b.z); g[7] = __uint_as_float(b.w);
    } else {
        unpack_bf8(ld_stream_nc(reinterpret_cast<const uint4 *>(g_) + j), g);
    }
}

template <int NUMERICS>
__device__ __forceinline__ void update_vec8(uint4 &P, const float (&g)[8], uint4 &M, uint4 &V,
                                            const AdamWConsts &c, const AdamWFast &f) {
    if (NUMERICS == 0) adamw_vec8_fast(P, g, M, V, f);
    else adamw_vec8<NUMERICS>(P, g, M, V, c);
}

template <int NUMERICS, bool GRAD_F32>
__global__ void __launch_bounds__(kThreads, kMinBlocks)
adamw_bf16_kernel(uint4 *__restrict__ p, const void *__restrict__ g_, uint4 *__restrict__ m,
                  uint4 *__restrict__ v, uint64_t nvec, AdamWConsts c) {
    const AdamWFast f = make_adamw_fast(c);
    const uint64_t stride = (uint64_t)gridDim.x * kThreads;
    uint64_t i = (uint64_t)blockIdx.x * kThreads + threadIdx.x;
    // main loop: kUnroll full vectors per thread, all 4*kUnroll loads issued before any math
    for (; i + (kUnroll - 1) * stride < nvec; i += stride * kUnroll) {
        uint4 P[kUnroll], M[kUnroll], V[kUnroll];
        float g[kUnroll][8];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const uint64_t j = i + u * stride;
            P[u] = ld_stream(p + j);
            M[u] = ld_stream(m + j);
            V[u] = ld_stream(v + j);
            load_grad8<GRAD_F32>(g_, j, g[u]);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const uint64_t j = i + u * stride;
            update_vec8<NUMERICS>(P[u], g[u], M[u], V[u], c, f);
            st_stream(p + j, P[u]);
            st_stream(m + j, M[u]);
            st_stream(v + j, V[u]);
        }
    }
    for (; i < nvec; i += stride) {
        uint4 P = ld_stream(p + i), M = ld_stream(m + i), V = ld_stream(v + i);
        float g[8];
        load_grad8<GRAD_F32>(g_, i, g);
        update_vec8<NUMERICS>(P, g, M, V, c, f);
        st_stream(p + i, P);
        st_stream(m + i, M);
        st_stream(v + i, V);
    }
}

__global__ void __launch_bounds__(kThreads)
adamw_f32_kernel(float4 *__restrict__ p, const float4 *__restrict__ g, float4 *__restrict__ m,
                 float4 *__restrict__ v, uint64_t nvec, AdamWConsts c) {
    const uint64_t stride = (uint64_t)gridDim.x * kThreads;
    for (uint64_t i = (uint64_t)blockIdx.x * kThreads + threadIdx.x; i < nvec; i += stride) {
        float4 P = p[i], G = __ldg(g + i), M = m[i], V = v[i];
        adamw_fp32(P.x, G.x, M.x, V.x, c);
        adamw_fp32(P.y, G.y, M.y, V.y, c);
        adamw_fp32(P.z, G.z, M.z, V.z, c);
        adamw_fp32(P.w, G.w, M.w, V.w, c);
        p[i] = P; m[i] = M; v[i] = V;
    }
}

// Element-wise path for the < 1 vector tail and for arenas that are not 16-B aligned.
template <int NUMERICS>
__global__ void adamw_scalar_kernel(void *p_, const void *g_, void *m_, void *v_, uint64_t first,
                                    uint64_t n, int dtype, int grad_dtype, AdamWConsts c) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = first + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        if (dtype == PR_F32) {
            float *p = (float *)p_, *m = (float *)m_, *v = (float *)v_;
            adamw_fp32(p[i], ((const float *)g_)[i], m[i], v[i], c);
        } else {
            uint16_t *p = (uint16_t *)p_, *m = (uint16_t *)m_, *v = (uint16_t *)v_;
            float g = grad_dtype == PR_F32 ? ((const float *)g_)[i]
                                           : __uint_as_float((uint32_t)((const uint16_t *)g_)[i] << 16);
            float pf = __uint_as_float((uint32_t)p[i] << 16);
            float mf = __uint_as_float((uint32_t)m[i] << 16);
            float vf = __uint_as_float((uint32_t)v[i] << 16);
            if (NUMERICS == 1) adamw_stepwise_bf16(pf, g, mf, vf, c);
            else adamw_fp32(pf, g, mf, vf, c);
            p[i] = (uint16_t)f2bf_bits(pf);
            m[i] = (uint16_t)f2bf_bits(mf);
            v[i] = (uint16_t)f2bf_bits(vf);
        }
    }
}

int grid_for(uint64_t work_items, int per_block, int blocks_per_sm) {
    uint64_t need = (work_items + per_block - 1) / per_block;
    uint64_t cap = (uint64_t)pr_sm_count() * blocks_per_sm;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

}  // namespace

extern "C" int pr_adamw(void *p, const void *g, void *m, void *v, uint64_t n, int dtype,
                        const pr_adamw_args *a, void *stream_) {
    PR_REQUIRE(a != nullptr, "pr_adamw: args is NULL");
    if (n == 0) return PR_OK;
    PR_REQUIRE(p && g && m && v, "pr_adamw: NULL arena pointer");
    PR_REQUIRE(dtype == PR_BF16 || dtype == PR_F32, "pr_adamw: dtype %d unsupported", dtype);
    PR_REQUIRE(a->grad_dtype == PR_BF16 || a->grad_dtype == PR_F32, "pr_adamw: grad_dtype %d", a->grad_dtype);
    PR_REQUIRE(!(dtype == PR_F32 && a->grad_dtype != PR_F32), "pr_adamw: fp32 params need fp32 grads");
    PR_REQUIRE(a->numerics == 0 || (a->numerics == 1 && dtype == PR_BF16),
               "pr_adamw: numerics %d unsupported for dtype %d", a->numerics, dtype);
    PR_REQUIRE(!(a->grad_dtype == PR_F32 && a->numerics != 0), "pr_adamw: fp32 grads support numerics 0 only");
    PR_REQUIRE(a->step >= 1, "pr_adamw: step must be >= 1 (got %lld)", (long long)a->step);
    cudaStream_t stream = (cudaStream_t)stream_;
    AdamWConsts c = make_adamw_consts(a);

    const bool aligned = pr_aligned16(p) && pr_aligned16(g) && pr_aligned16(m) && pr_aligned16(v);
    const int per_vec = dtype == PR_F32 ? 4 : 8;
    uint64_t nvec = aligned ? n / per_vec : 0;
    static int blocks_per_sm = [] {
        const char *e = getenv("PR_ADAMW_BLOCKS_PER_SM");
        int b = e ? atoi(e) : 4;
        return b > 0 ? b : 4;
    }();
    if (nvec) {
        int grid = grid_for(nvec, kThreads * (dtype == PR_F32 ? 1 : kUnroll), blocks_per_sm);
        if (dtype == PR_F32) {
            adamw_f32_kernel<<<grid, kThreads, 0, stream>>>((float4 *)p, (const float4 *)g,
                                                           (float4 *)m, (float4 *)v, nvec, c);
        } else if (a->grad_dtype == PR_F32) {
            adamw_bf16_kernel<0, true><<<grid, kThreads, 0, stream>>>((uint4 *)p, g, (uint4 *)m,
                                                                       (uint4 *)v, nvec, c);
        } else if (a->numerics == 1) {
            adamw_bf16_kernel<1, false><<<grid, kThreads, 0, stream>>>((uint4 *)p, g, (uint4 *)m,
                                                                        (uint4 *)v, nvec, c);
        } else {
            adamw_bf16_kernel<0, false><<<grid, kThreads, 0, stream>>>((uint4 *)p, g, (uint4 *)m,
                                                                        (uint4 *)v, nvec, c);
        }
        PR_CUDA(cudaGetLastError());
        pr_count_launch();
    }
    uint64_t done = nvec * per_vec;
    if (done < n) {
        int grid = grid_for(n - done, 256, 4);
        if (a->numerics == 1)
            adamw_scalar_kernel<1><<<grid, 256, 0, stream>>>(p, g, m, v, done, n, dtype, a->grad_dtype, c);
        else
            adamw_scalar_kernel<0><<<grid, 256, 0, stream>>>(p, g, m, v, done, n, dtype, a->grad_dtype, c);
        PR_CUDA(cudaGetLastError());
        pr_count_launch();
    }
    return PR_OK;
}
