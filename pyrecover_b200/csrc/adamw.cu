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
#include <vector>

#include "adamw_math.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 2;
constexpr int kMinBlocks = 3;    // <= 85 registers per thread, 24 resident warps per SM (the grid is sized for 3 CTAs/SM anyway)
// CTAs per SM the grid is sized for.  Measured on B200, 128 MiB - 8 GiB arenas (round 1): 3 CTAs/SM give
// 0.95 - 0.98 of the measured HBM copy peak (6310 GB/s at 2 GiB), 4 give 0.87 - 0.92, 5 give 0.78: with four
// streams per thread fewer concurrent CTAs keep more DRAM pages open.  PR_ADAMW_BLOCKS_PER_SM overrides.
constexpr int kDefaultBlocksPerSM = 3;
#ifndef PR_STEPWISE_FAST
#define PR_STEPWISE_FAST 1
#endif
constexpr bool kStepwiseFast = PR_STEPWISE_FAST != 0;   // numerics 1 through the branch-free sequences (0: IEEE intrinsics)

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
    else if (kStepwiseFast) adamw_vec8_stepwise_fast(P, g, M, V, f);
    else adamw_vec8<NUMERICS>(P, g, M, V, c);
}

template <int NUMERICS, bool GRAD_F32>
__global__ void __launch_bounds__(kThreads, kMinBlocks)
adamw_bf16_kernel(uint4 *__restrict__ p, const void *__restrict__ g_, uint4 *__restrict__ m,
                  uint4 *__restrict__ v, uint64_t nvec, AdamWConsts c_) {
    const AdamWConsts c = resolve_gscale(c_);
    if (c.skip) return;
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
                 float4 *__restrict__ v, uint64_t nvec, AdamWConsts c_) {
    const AdamWConsts c = resolve_gscale(c_);
    if (c.skip) return;
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
                                    uint64_t n, int dtype, int grad_dtype, AdamWConsts c_) {
    const AdamWConsts c = resolve_gscale(c_);
    if (c.skip) return;
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

// ---- K2 with gradients left where autograd allocated them -----------------------------------
// At world size 1 nothing needs the gradients in one buffer, so copying 147-291 tensors into the
// g arena first (2 B read + 2 B write per parameter) is pure overhead: this variant reads tensor
// s's gradient from its own pointer.  p / m / v stay flat.  Work is cut into tiles of kTileElems
// that never straddle tensors (device table built once per layout); the per-launch gradient
// pointers travel as kernel parameters, so there is no per-step upload to synchronise.
constexpr int kMaxGradPtrs = 384;
#ifndef PR_ADAMW_TILE_ELEMS
#define PR_ADAMW_TILE_ELEMS 16384
#endif
constexpr uint32_t kTileElems = PR_ADAMW_TILE_ELEMS;    // measured in the 1 B step: 4096 -> 2.44 ms, 16384 -> 2.39, 65536 -> 2.38

struct GradPtrs { const void *ptr[kMaxGradPtrs]; };

struct AdamWPlanDev {
    const uint64_t *elem_off;     // tensor -> first element in the arenas
    const uint64_t *numel;
    const uint64_t *tile_first;   // tensor -> its first tile
    const uint32_t *tile_seg;     // tile -> tensor
};

// The tensor a tile belongs to, kept across a CTA's tiles: the tile -> tensor table costs two dependent loads
// (~1 us) when consulted, and a CTA's next tile (gridDim.x tiles further on) mostly lies in the same tensor.
struct SegCursor {
    uint64_t first = 0, tiles = 0, numel = 0, base = 0;
    const uint16_t *grad = nullptr;
    __device__ __forceinline__ void seek(const AdamWPlanDev &pd, const GradPtrs &gp, uint32_t seg_lo, uint64_t t) {
        if (t - first >= tiles) {                              // also true on the first call (tiles == 0)
            const uint32_t s = pd.tile_seg[t];
            first = pd.tile_first[s];
            numel = pd.numel[s];
            tiles = (numel + kTileElems - 1) / kTileElems;
            base = pd.elem_off[s];
            grad = reinterpret_cast<const uint16_t *>(gp.ptr[s - seg_lo]);
        }
    }
};

// numerics 0 fits 64 registers without spilling and measured 1 % faster that way inside the training step
// (2.38 vs 2.41 ms); the stepwise numerics needs the 85 that 3 CTAs/SM allow
template <int NUMERICS>
__global__ void __launch_bounds__(kThreads, NUMERICS == 0 ? 4 : kMinBlocks)
adamw_bf16_scattered_kernel(uint16_t *__restrict__ p, uint16_t *__restrict__ m, uint16_t *__restrict__ v,
                            AdamWPlanDev pd, uint64_t tile_lo, uint64_t tile_hi, uint32_t seg_lo,
                            const __grid_constant__ GradPtrs gp, AdamWConsts c_) {
    const AdamWConsts c = resolve_gscale(c_);
    if (c.skip) return;
    const AdamWFast f = make_adamw_fast(c);
    SegCursor sc;
    for (uint64_t t = tile_lo + blockIdx.x; t < tile_hi; t += gridDim.x) {
        sc.seek(pd, gp, seg_lo, t);
        const uint64_t seg_first = sc.first, seg_numel = sc.numel, seg_base = sc.base;
        const uint16_t *seg_grad = sc.grad;
        const uint64_t off = (t - seg_first) * kTileElems;
        const uint64_t left = seg_numel - off;
        const uint32_t n = left < kTileElems ? (uint32_t)left : kTileElems;
        const uint64_t base = seg_base + off;                  // multiple of 8: slots are 256-B aligned
        uint4 *P = reinterpret_cast<uint4 *>(p + base);
        uint4 *M = reinterpret_cast<uint4 *>(m + base);
        uint4 *V = reinterpret_cast<uint4 *>(v + base);
        const uint16_t *gs = seg_grad + off;
        const uint4 *G = reinterpret_cast<const uint4 *>(gs);
        const uint32_t nvec = n / 8;
        uint32_t k = threadIdx.x;
        for (; k + kThreads < nvec; k += 2 * kThreads) {
            uint4 Pa = ld_stream(P + k), Ma = ld_stream(M + k), Va = ld_stream(V + k), Ga = ld_stream_nc(G + k);
            uint4 Pb = ld_stream(P + k + kThreads), Mb = ld_stream(M + k + kThreads), Vb = ld_stream(V + k + kThreads),
                  Gb = ld_stream_nc(G + k + kThreads);
            float ga[8], gb[8];
            unpack_bf8(Ga, ga);
            unpack_bf8(Gb, gb);
            update_vec8<NUMERICS>(Pa, ga, Ma, Va, c, f);
            st_stream(P + k, Pa); st_stream(M + k, Ma); st_stream(V + k, Va);
            update_vec8<NUMERICS>(Pb, gb, Mb, Vb, c, f);
            st_stream(P + k + kThreads, Pb); st_stream(M + k + kThreads, Mb); st_stream(V + k + kThreads, Vb);
        }
        for (; k < nvec; k += kThreads) {
            uint4 Pa = ld_stream(P + k), Ma = ld_stream(M + k), Va = ld_stream(V + k), Ga = ld_stream_nc(G + k);
            float ga[8];
            unpack_bf8(Ga, ga);
            update_vec8<NUMERICS>(Pa, ga, Ma, Va, c, f);
            st_stream(P + k, Pa); st_stream(M + k, Ma); st_stream(V + k, Va);
        }
        if (threadIdx.x < (n & 7u)) {       // ragged tail of a tensor whose numel is not a multiple of 8
            const uint64_t i = base + 8ull * nvec + threadIdx.x;
            float pf = __uint_as_float((uint32_t)p[i] << 16), mf = __uint_as_float((uint32_t)m[i] << 16);
            float vf = __uint_as_float((uint32_t)v[i] << 16);
            float g = __uint_as_float((uint32_t)gs[8u * nvec + threadIdx.x] << 16);
            if (NUMERICS == 1) adamw_stepwise_bf16(pf, g, mf, vf, c);
            else adamw_fp32(pf, g, mf, vf, c);
            p[i] = (uint16_t)f2bf_bits(pf); m[i] = (uint16_t)f2bf_bits(mf); v[i] = (uint16_t)f2bf_bits(vf);
        }
    }
}

// ---- gather: gradients from where autograd left them into the flat g arena ----------------------
// With W > 1 the gradient exchange needs the gradients in the symmetric (peer-mapped, multicast-bound)
// arena.  This is that copy, as ONE launch per bucket of tensors [seg_lo, seg_hi) over the same tile
// table as the scattered K2 -- instead of a library multi-tensor copy of all 147-291 tensors after
// backward has finished (reference: the Reducer's grad -> bucket copy, torch/csrc/distributed/c10d/
// reducer.cpp mark_variable_ready_dense; call site train.py:269).  HBM: 2 B read + 2 B write per element.
__global__ void __launch_bounds__(kThreads)
gather_grads_kernel(uint16_t *__restrict__ g, AdamWPlanDev pd, uint64_t tile_lo, uint64_t tile_hi, uint32_t seg_lo,
                    const __grid_constant__ GradPtrs gp) {
    SegCursor sc;
    for (uint64_t t = tile_lo + blockIdx.x; t < tile_hi; t += gridDim.x) {
        sc.seek(pd, gp, seg_lo, t);
        const uint64_t off = (t - sc.first) * kTileElems;
        const uint64_t left = sc.numel - off;
        const uint32_t n = left < kTileElems ? (uint32_t)left : kTileElems;
        const uint64_t base = sc.base + off;
        const uint16_t *src = sc.grad + off;
        const uint4 *S = reinterpret_cast<const uint4 *>(src);
        uint4 *D = reinterpret_cast<uint4 *>(g + base);
        const uint32_t nvec = n / 8;
        uint32_t k = threadIdx.x;
        for (; k + 3 * kThreads < nvec; k += 4 * kThreads) {
            uint4 a = ld_stream_nc(S + k), b = ld_stream_nc(S + k + kThreads), c = ld_stream_nc(S + k + 2 * kThreads),
                  d = ld_stream_nc(S + k + 3 * kThreads);
            st_stream(D + k, a); st_stream(D + k + kThreads, b); st_stream(D + k + 2 * kThreads, c);
            st_stream(D + k + 3 * kThreads, d);
        }
        for (; k < nvec; k += kThreads) st_stream(D + k, ld_stream_nc(S + k));
        if (threadIdx.x < (n & 7u)) g[base + 8ull * nvec + threadIdx.x] = src[8u * nvec + threadIdx.x];
    }
}

// ---- global gradient norm (clip_grad_norm_, reference utils.py:84-89) -------------------------
// Roofline: HBM, 2 B per bf16 gradient element read once.  Stage 1: fp32 sum of squares per
// thread, double from the warp shuffle on; one double per CTA.  Stage 2: one CTA adds the CTA
// partials in a fixed order and writes norm / clip coefficient / non-finite flag.
constexpr int kNormThreads = 256;

__device__ __forceinline__ float sumsq_vec(const uint4 &G, float s) {
    const uint32_t *w = &G.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float a = bf_lo(w[k]), b = bf_hi(w[k]);
        s = __fmaf_rn(a, a, s);
        s = __fmaf_rn(b, b, s);
    }
    return s;
}

__device__ __forceinline__ void block_partial(float s, double *partial) {
    __shared__ double warp_sum[kNormThreads / 32];
    double d = (double)s;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = d;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < kNormThreads / 32; ++w) t += warp_sum[w];
        partial[blockIdx.x] = t;
    }
}

template <bool F32>
__global__ void __launch_bounds__(kNormThreads)
sumsq_flat_kernel(const void *__restrict__ g_, uint64_t n, double *__restrict__ partial) {
    const uint64_t stride = (uint64_t)gridDim.x * kNormThreads;
    const uint64_t i0 = (uint64_t)blockIdx.x * kNormThreads + threadIdx.x;
    float s = 0.0f;
    const int per_vec = F32 ? 4 : 8;
    const uint64_t nvec = n / per_vec;
    const uint4 *G = reinterpret_cast<const uint4 *>(g_);
    uint64_t i = i0;
    for (; i + 3 * stride < nvec; i += 4 * stride) {          // 4 independent 16-B loads in flight
        uint4 a = ld_stream_nc(G + i), b = ld_stream_nc(G + i + stride), c = ld_stream_nc(G + i + 2 * stride),
              d = ld_stream_nc(G + i + 3 * stride);
        if (F32) {
            const uint4 q[4] = {a, b, c, d};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float x = __uint_as_float(q[k].x), y = __uint_as_float(q[k].y), z = __uint_as_float(q[k].z),
                      w = __uint_as_float(q[k].w);
                s = __fmaf_rn(x, x, s); s = __fmaf_rn(y, y, s); s = __fmaf_rn(z, z, s); s = __fmaf_rn(w, w, s);
            }
        } else {
            s = sumsq_vec(a, s); s = sumsq_vec(b, s); s = sumsq_vec(c, s); s = sumsq_vec(d, s);
        }
    }
    for (; i < nvec; i += stride) {
        uint4 a = ld_stream_nc(G + i);
        if (F32) {
            float x = __uint_as_float(a.x), y = __uint_as_float(a.y), z = __uint_as_float(a.z), w = __uint_as_float(a.w);
            s = __fmaf_rn(x, x, s); s = __fmaf_rn(y, y, s); s = __fmaf_rn(z, z, s); s = __fmaf_rn(w, w, s);
        } else {
            s = sumsq_vec(a, s);
        }
    }
    for (uint64_t e = nvec * per_vec + i0; e < n; e += stride) {     // < 1 vector of tail elements
        const float x = F32 ? reinterpret_cast<const float *>(g_)[e]
                            : __uint_as_float((uint32_t)reinterpret_cast<const uint16_t *>(g_)[e] << 16);
        s = __fmaf_rn(x, x, s);
    }
    block_partial(s, partial);
}

__global__ void __launch_bounds__(kNormThreads)
sumsq_scattered_kernel(AdamWPlanDev pd, uint64_t tile_lo, uint64_t tile_hi, uint32_t seg_lo,
                       const __grid_constant__ GradPtrs gp, double *__restrict__ partial) {
    float s = 0.0f;
    SegCursor sc;
    for (uint64_t t = tile_lo + blockIdx.x; t < tile_hi; t += gridDim.x) {
        sc.seek(pd, gp, seg_lo, t);
        const uint64_t off = (t - sc.first) * kTileElems;
        const uint64_t left = sc.numel - off;
        const uint32_t n = left < kTileElems ? (uint32_t)left : kTileElems;
        const uint16_t *gs = sc.grad + off;
        const uint4 *G = reinterpret_cast<const uint4 *>(gs);
        const uint32_t nvec = n / 8;
        uint32_t k = threadIdx.x;
        for (; k + kNormThreads < nvec; k += 2 * kNormThreads) {
            uint4 a = ld_stream_nc(G + k), b = ld_stream_nc(G + k + kNormThreads);
            s = sumsq_vec(a, s);
            s = sumsq_vec(b, s);
        }
        for (; k < nvec; k += kNormThreads) s = sumsq_vec(ld_stream_nc(G + k), s);
        if (threadIdx.x < (n & 7u)) {
            const float x = __uint_as_float((uint32_t)gs[8u * nvec + threadIdx.x] << 16);
            s = __fmaf_rn(x, x, s);
        }
    }
    block_partial(s, partial);
}

// partial[first .. first + count) -> out (norm, coefficient, non-finite flag).  `accumulate` adds to
// the running sum kept in partial[PR_GRAD_NORM_SCRATCH - 1] instead of finishing (the scattered
// variant launches once per group of kMaxGradPtrs tensors).
__global__ void __launch_bounds__(kNormThreads)
norm_finish_kernel(double *__restrict__ partial, int count, float prescale, float max_norm,
                   float *__restrict__ out, int first_group, int last_group) {
    __shared__ double sh[kNormThreads];
    double t = 0.0;
    for (int i = threadIdx.x; i < count; i += kNormThreads) t += partial[i];
    sh[threadIdx.x] = t;
    __syncthreads();
    for (int o = kNormThreads / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        double total = sh[0] + (first_group ? 0.0 : partial[PR_GRAD_NORM_SCRATCH - 1]);
        partial[PR_GRAD_NORM_SCRATCH - 1] = total;
        if (last_group) {
            const float norm = __fmul_rn(prescale, (float)sqrt(total));
            const bool finite = isfinite(norm);
            float coef = 1.0f;
            if (max_norm > 0.0f && finite) {
                coef = __fdiv_rn(max_norm, __fadd_rn(norm, 1e-6f));
                if (coef > 1.0f) coef = 1.0f;
            }
            out[0] = norm;
            out[1] = finite ? coef : -1.0f;      // negative: K2 leaves p, m, v untouched (resolve_gscale)
            out[2] = finite ? 0.0f : 1.0f;
            if (!finite) out[3] = out[3] + 1.0f; // sticky count of non-finite norms (the host checks it lazily)
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
    static int blocks_env = [] {
        const char *e = getenv("PR_ADAMW_BLOCKS_PER_SM");
        int b = e ? atoi(e) : 0;
        return b > 0 ? b : 0;
    }();
    // the stepwise numerics (a rounding after every operation) is issue-bound, not DRAM-page-bound: it keeps 4
    const int blocks_per_sm = blocks_env ? blocks_env : (a->numerics == 1 ? kMinBlocks : kDefaultBlocksPerSM);
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

struct pr_adamw_plan {
    int n = 0;
    uint64_t arena_elems = 0, ntiles = 0;
    void *table = nullptr;
    AdamWPlanDev dev{};
    std::vector<uint64_t> tile_first;     // host copy, to split launches by tensor range
};

extern "C" int pr_adamw_plan_create(const uint64_t *elem_off, const uint64_t *numel, int n, uint64_t arena_elems,
                                    pr_adamw_plan **out) {
    PR_REQUIRE(out != nullptr, "pr_adamw_plan_create: plan_out is NULL");
    *out = nullptr;
    PR_REQUIRE(n > 0 && elem_off && numel, "pr_adamw_plan_create: empty layout");
    std::vector<uint64_t> first(n + 1);
    std::vector<uint32_t> tile_seg;
    for (int i = 0; i < n; ++i) {
        PR_REQUIRE(elem_off[i] % 8 == 0, "pr_adamw_plan_create: tensor %d starts at element %llu (not 16-B aligned)", i,
                   (unsigned long long)elem_off[i]);
        PR_REQUIRE(elem_off[i] + numel[i] <= arena_elems, "pr_adamw_plan_create: tensor %d overruns the arena", i);
        first[i] = tile_seg.size();
        tile_seg.insert(tile_seg.end(), (numel[i] + kTileElems - 1) / kTileElems, (uint32_t)i);
    }
    first[n] = tile_seg.size();
    auto *plan = new pr_adamw_plan();
    plan->n = n;
    plan->arena_elems = arena_elems;
    plan->ntiles = tile_seg.size();
    plan->tile_first = first;
    const size_t seg = sizeof(uint64_t) * (size_t)n;
    const size_t total = 3 * seg + sizeof(uint32_t) * tile_seg.size();
    cudaError_t e = cudaMalloc(&plan->table, total);
    if (e != cudaSuccess) {
        delete plan;
        pr_set_error("pr_adamw_plan_create: cudaMalloc(%zu) failed: %s", total, cudaGetErrorString(e));
        return PR_ERR_CUDA;
    }
    std::vector<uint8_t> host(total);
    memcpy(host.data(), elem_off, seg);
    memcpy(host.data() + seg, numel, seg);
    memcpy(host.data() + 2 * seg, first.data(), seg);
    memcpy(host.data() + 3 * seg, tile_seg.data(), sizeof(uint32_t) * tile_seg.size());
    e = cudaMemcpy(plan->table, host.data(), total, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(plan->table);
        delete plan;
        pr_set_error("pr_adamw_plan_create: table upload failed: %s", cudaGetErrorString(e));
        return PR_ERR_CUDA;
    }
    uint8_t *b = (uint8_t *)plan->table;
    plan->dev.elem_off = (const uint64_t *)b;
    plan->dev.numel = (const uint64_t *)(b + seg);
    plan->dev.tile_first = (const uint64_t *)(b + 2 * seg);
    plan->dev.tile_seg = (const uint32_t *)(b + 3 * seg);
    *out = plan;
    return PR_OK;
}

extern "C" int pr_adamw_plan_destroy(pr_adamw_plan *plan) {
    if (!plan) return PR_OK;
    if (plan->table) cudaFree(plan->table);
    delete plan;
    return PR_OK;
}

extern "C" int pr_adamw_scattered(const pr_adamw_plan *plan, void *p, const void *const *grad_dev, void *m, void *v,
                                  const pr_adamw_args *a, void *stream_) {
    PR_REQUIRE(plan && p && grad_dev && m && v && a, "pr_adamw_scattered: NULL argument");
    PR_REQUIRE(a->grad_dtype == PR_BF16 && (a->numerics == 0 || a->numerics == 1) && a->step >= 1,
               "pr_adamw_scattered: bf16 gradients, numerics 0/1 and step >= 1 are required");
    PR_REQUIRE(pr_aligned16(p) && pr_aligned16(m) && pr_aligned16(v), "pr_adamw_scattered: arenas must be 16-B aligned");
    for (int i = 0; i < plan->n; ++i)
        PR_REQUIRE(grad_dev[i] != nullptr && pr_aligned16(grad_dev[i]),
                   "pr_adamw_scattered: gradient %d is NULL or not 16-B aligned", i);
    cudaStream_t stream = (cudaStream_t)stream_;
    AdamWConsts c = make_adamw_consts(a);
    static int blocks_env = [] {
        const char *e = getenv("PR_ADAMW_BLOCKS_PER_SM");
        int b = e ? atoi(e) : 0;
        return b > 0 ? b : 0;
    }();
    // the stepwise numerics (a rounding after every operation) is issue-bound, not DRAM-page-bound: it keeps 4
    const int blocks_per_sm = blocks_env ? blocks_env : (a->numerics == 1 ? kMinBlocks : kDefaultBlocksPerSM);
    for (int s0 = 0; s0 < plan->n; s0 += kMaxGradPtrs) {
        const int s1 = s0 + kMaxGradPtrs < plan->n ? s0 + kMaxGradPtrs : plan->n;
        GradPtrs gp;
        memset(&gp, 0, sizeof(gp));
        for (int s = s0; s < s1; ++s) gp.ptr[s - s0] = grad_dev[s];
        const uint64_t t0 = plan->tile_first[s0], t1 = plan->tile_first[s1];
        if (t1 == t0) continue;
        uint64_t cap = (uint64_t)pr_sm_count() * blocks_per_sm;
        int grid = (int)((t1 - t0) < cap ? (t1 - t0) : cap);
        if (a->numerics == 1)
            adamw_bf16_scattered_kernel<1><<<grid, kThreads, 0, stream>>>((uint16_t *)p, (uint16_t *)m, (uint16_t *)v,
                                                                         plan->dev, t0, t1, (uint32_t)s0, gp, c);
        else
            adamw_bf16_scattered_kernel<0><<<grid, kThreads, 0, stream>>>((uint16_t *)p, (uint16_t *)m, (uint16_t *)v,
                                                                         plan->dev, t0, t1, (uint32_t)s0, gp, c);
        PR_CUDA(cudaGetLastError());
        pr_count_launch();
    }
    return PR_OK;
}

extern "C" int pr_grad_norm(const void *g, uint64_t n, int dtype, float prescale, float max_norm, float *out,
                            double *scratch, void *stream_) {
    PR_REQUIRE(out && scratch, "pr_grad_norm: NULL output or scratch");
    PR_REQUIRE(dtype == PR_BF16 || dtype == PR_F32, "pr_grad_norm: dtype %d unsupported", dtype);
    PR_REQUIRE(n == 0 || (g != nullptr && pr_aligned16(g)), "pr_grad_norm: gradient buffer is NULL or not 16-B aligned");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int per_vec = dtype == PR_F32 ? 4 : 8;
    int grid = grid_for(n / per_vec + 1, kNormThreads * 4, 4);
    if (grid > PR_GRAD_NORM_SCRATCH - 1) grid = PR_GRAD_NORM_SCRATCH - 1;
    if (dtype == PR_F32) sumsq_flat_kernel<true><<<grid, kNormThreads, 0, stream>>>(g, n, scratch);
    else sumsq_flat_kernel<false><<<grid, kNormThreads, 0, stream>>>(g, n, scratch);
    PR_CUDA(cudaGetLastError());
    norm_finish_kernel<<<1, kNormThreads, 0, stream>>>(scratch, grid, prescale, max_norm, out, 1, 1);
    PR_CUDA(cudaGetLastError());
    pr_count_launch(2);
    return PR_OK;
}

extern "C" int pr_grad_norm_scattered(const pr_adamw_plan *plan, const void *const *grad_dev, float prescale,
                                      float max_norm, float *out, double *scratch, void *stream_) {
    PR_REQUIRE(plan && grad_dev && out && scratch, "pr_grad_norm_scattered: NULL argument");
    for (int i = 0; i < plan->n; ++i)
        PR_REQUIRE(grad_dev[i] != nullptr && pr_aligned16(grad_dev[i]),
                   "pr_grad_norm_scattered: gradient %d is NULL or not 16-B aligned", i);
    cudaStream_t stream = (cudaStream_t)stream_;
    for (int lo = 0; lo < plan->n; lo += kMaxGradPtrs) {
        const int hi = lo + kMaxGradPtrs < plan->n ? lo + kMaxGradPtrs : plan->n;
        GradPtrs gp;
        memset(&gp, 0, sizeof(gp));
        for (int i = lo; i < hi; ++i) gp.ptr[i - lo] = grad_dev[i];
        const uint64_t t_lo = plan->tile_first[lo], t_hi = plan->tile_first[hi];
        int grid = grid_for(t_hi - t_lo, 1, 4);
        if (grid > PR_GRAD_NORM_SCRATCH - 1) grid = PR_GRAD_NORM_SCRATCH - 1;
        sumsq_scattered_kernel<<<grid, kNormThreads, 0, stream>>>(plan->dev, t_lo, t_hi, (uint32_t)lo, gp, scratch);
        PR_CUDA(cudaGetLastError());
        norm_finish_kernel<<<1, kNormThreads, 0, stream>>>(scratch, grid, prescale, max_norm, out, lo == 0 ? 1 : 0,
                                                           hi == plan->n ? 1 : 0);
        PR_CUDA(cudaGetLastError());
        pr_count_launch(2);
    }
    return PR_OK;
}

extern "C" int pr_gather_grads(const pr_adamw_plan *plan, const void *const *grad_dev, int seg_lo, int seg_hi,
                               void *g_dev, int max_blocks, void *stream_) {
    PR_REQUIRE(plan && grad_dev && g_dev, "pr_gather_grads: NULL argument");
    PR_REQUIRE(seg_lo >= 0 && seg_lo <= seg_hi && seg_hi <= plan->n, "pr_gather_grads: tensor range [%d, %d) not in [0, %d]",
               seg_lo, seg_hi, plan->n);
    PR_REQUIRE(pr_aligned16(g_dev), "pr_gather_grads: the arena must be 16-B aligned");
    for (int i = seg_lo; i < seg_hi; ++i)
        PR_REQUIRE(grad_dev[i] != nullptr && pr_aligned16(grad_dev[i]), "pr_gather_grads: gradient %d is NULL or not 16-B aligned", i);
    cudaStream_t stream = (cudaStream_t)stream_;
    for (int s0 = seg_lo; s0 < seg_hi; s0 += kMaxGradPtrs) {
        const int s1 = s0 + kMaxGradPtrs < seg_hi ? s0 + kMaxGradPtrs : seg_hi;
        GradPtrs gp;
        memset(&gp, 0, sizeof(gp));
        for (int s = s0; s < s1; ++s) gp.ptr[s - s0] = grad_dev[s];
        const uint64_t t0 = plan->tile_first[s0], t1 = plan->tile_first[s1];
        if (t1 == t0) continue;
        uint64_t cap = (uint64_t)pr_sm_count() * kDefaultBlocksPerSM;
        if (max_blocks > 0 && (uint64_t)max_blocks < cap) cap = (uint64_t)max_blocks;
        int grid = (int)((t1 - t0) < cap ? (t1 - t0) : cap);
        gather_grads_kernel<<<grid, kThreads, 0, stream>>>((uint16_t *)g_dev, plan->dev, t0, t1, (uint32_t)s0, gp);
        PR_CUDA(cudaGetLastError());
        pr_count_launch();
    }
    return PR_OK;
}
