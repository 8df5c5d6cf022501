/*
 * pyrecover_b200.h -- C ABI of libpyrecover_b200.so (sm_100a only).
 *
 * The reference (Shaswat-G/PyRecover) is pure Python and has no FFI of its own; the three
 * places where its training step hands work to native code are library calls made from
 * train.py and pyrecover/checkpoint.py.  Each entry point below names the reference call
 * site it stands in for.  The Python binding a maintainer adds is in INTEGRATION.md.
 *
 * Conventions
 *   - every function returns 0 (PR_OK) or a negative pr_status; pr_last_error() gives the
 *     text of the last failure on the calling thread.  No exception crosses the boundary.
 *   - pointers named *_dev are device pointers, *_host are host pointers; `stream` is a
 *     cudaStream_t passed as void* (NULL = the legacy default stream).
 *   - the caller owns every buffer it passes; the library borrows it until the work queued
 *     on `stream` has completed.
 *   - there is no CPU fallback: without a Blackwell GPU the calls fail with PR_ERR_CUDA.
 */
#ifndef PYRECOVER_B200_H
#define PYRECOVER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PR_ABI_VERSION 3
#define PR_MAX_WORLD 8            /* one NVSwitch domain: 8 x B200                          */
#define PR_MAX_REGIONS 4          /* symmetric regions per communicator                     */
#define PR_EXPORT_BLOB_BYTES 512  /* bytes pr_comm_export writes per rank                   */

typedef enum pr_status {
    PR_OK = 0,
    PR_ERR_INVALID = -1,   /* bad argument (null, misaligned, out of range)                 */
    PR_ERR_CUDA = -2,      /* a CUDA runtime call failed; text in pr_last_error()           */
    PR_ERR_UNSUPPORTED = -3,
    PR_ERR_STATE = -4,     /* call made in the wrong state (e.g. allreduce before import)   */
    PR_ERR_IO = -5
} pr_status;

typedef enum pr_dtype { PR_BF16 = 0, PR_F32 = 1 } pr_dtype;

/* Last error text for this thread ("" if none).  Never NULL. */
const char *pr_last_error(void);
int pr_abi_version(void);
/* SM count / compute capability of the current device. */
int pr_device_info(int *sm_count, int *cc_major, int *cc_minor);
/* Number of kernels this library has launched in this process (all entry points). */
uint64_t pr_launch_count(void);

/* ---------------------------------------------------------------- K2: AdamW ----------
 * Replaces optimizer.step() -- reference train.py:274, constructed at train.py:120-122 as
 * torch.optim.AdamW(model.parameters(), lr, fused=args.fused_optimizer); the arithmetic is
 * torch/optim/adam.py:347-551 (_single_tensor_adam) / :803-900 (_fused_adamw_).
 * One launch updates a flat arena of n elements: p, m, v in place, g read-only.
 *   numerics = 0: fp32 op-math, one rounding at the store  (class of fused=True)
 *   numerics = 1: result of every torch op rounded to the storage dtype (class of the
 *                 reference's default fused=False path on bf16 tensors)
 * lr is a runtime scalar because the reference's LambdaLR (utils.py:59-81, train.py:275)
 * changes it every warm-up step.  `step` is the 1-based step count t of this update.
 * grad_scale multiplies g first (1/W of the DDP mean, or a loss-scale inverse); 1.0 = off.
 * grad_dtype may be PR_F32 with dtype PR_BF16 (fp32 sums out of the reduce-scatter).
 */
typedef struct pr_adamw_args {
    double lr, beta1, beta2, eps, weight_decay, grad_scale;
    int64_t step;
    int32_t numerics;
    int32_t grad_dtype;  /* pr_dtype of g */
    const float *grad_scale_dev;  /* optional DEVICE scalar multiplied (in binary32) into grad_scale when the
                                     kernel starts -- the clip coefficient pr_grad_norm leaves on the device,
                                     so clipping costs no host synchronisation.  NULL = none. */
} pr_adamw_args;

int pr_adamw(void *p_dev, const void *g_dev, void *m_dev, void *v_dev, uint64_t n,
             int dtype, const pr_adamw_args *args, void *stream);

/* K2 reading every tensor's gradient from where autograd left it (world size 1: no exchange
 * needs the gradients in one buffer, so the copy into a flat g arena is skipped -- the reference's
 * optimizer.step() reads p.grad in place too, train.py:274).  The plan fixes the layout: tensor i
 * occupies elements [elem_off[i], elem_off[i] + numel[i]) of the p / m / v arenas (elem_off a
 * multiple of 8).  grad_dev[i] is a HOST array entry holding the DEVICE pointer of tensor i's
 * contiguous bf16 gradient (16-B aligned).  Same arithmetic as pr_adamw. */
typedef struct pr_adamw_plan pr_adamw_plan;
int pr_adamw_plan_create(const uint64_t *elem_off, const uint64_t *numel, int n_tensors,
                         uint64_t arena_elems, pr_adamw_plan **plan_out);
int pr_adamw_plan_destroy(pr_adamw_plan *plan);
int pr_adamw_scattered(const pr_adamw_plan *plan, void *p_dev, const void *const *grad_dev,
                       void *m_dev, void *v_dev, const pr_adamw_args *args, void *stream);

/* Copy the gradients of tensors [seg_lo, seg_hi) of the plan from where autograd left them into the flat
 * gradient arena g_dev (element offsets as in the plan) -- the Reducer's gradient -> bucket copy
 * (torch/csrc/distributed/c10d/reducer.cpp, fired from loss.backward(), reference train.py:269), as one
 * launch per bucket.  grad_dev has one entry per tensor of the plan; only [seg_lo, seg_hi) are read.
 * max_blocks bounds the grid (0 = 3 CTAs per SM). */
int pr_gather_grads(const pr_adamw_plan *plan, const void *const *grad_dev, int seg_lo, int seg_hi,
                    void *g_dev, int max_blocks, void *stream);

/* Global-norm gradient clipping -- the reference's clip_grad_norm_ (utils.py:84-89: total =
 * get_total_norm(grads), then clip_grads_with_norm_(params, max_norm, total)), whose call is
 * commented out at train.py:271-272 with the flag --grad-max-norm (utils.py:171-175) left behind.
 *   out_dev[0] = prescale * sqrt(sum g_i^2)                    (the total L2 norm)
 *   out_dev[1] = min(1, max_norm / (out_dev[0] + 1e-6))        (1 when max_norm <= 0)
 *                and -1 when the norm is not finite: a negative (or NaN) device scale makes pr_adamw,
 *                pr_adamw_scattered and pr_allreduce_adamw return without touching p, m or v
 *   out_dev[2] = 1 if the norm is not finite, else 0
 *   out_dev[3] += 1 when the norm is not finite (sticky; zero it once)
 * The reference raises on a non-finite norm (utils.py:87, error_if_nonfinite=True) before the optimizer
 * runs.  Here the host learns of it later (it reads out_dev[3] when it logs or saves), so the update is
 * withheld on the device instead: inf/NaN gradients never reach the parameters, the moments or a checkpoint.
 * out_dev holds 4 floats.  Pass out_dev + 1 as pr_adamw_args.grad_scale_dev and the update applies the coefficient in
 * fp32 inside K2 -- no extra pass over the gradients and no rounding of the clipped gradients to
 * bf16 as clip_grads_with_norm_ does.  Two launches; squares are summed in fp32 per thread, in
 * double across threads and CTAs, in a fixed order: the result is deterministic for a given n
 * and GPU (bit-exact resume depends on it).  scratch_dev: >= PR_GRAD_NORM_SCRATCH doubles owned
 * by the caller (no allocation happens here).
 * pr_grad_norm_scattered reads every tensor's gradient where autograd left it (see
 * pr_adamw_scattered). */
#define PR_GRAD_NORM_SCRATCH 2048
int pr_grad_norm(const void *g_dev, uint64_t n, int dtype, float prescale, float max_norm,
                 float *out_dev, double *scratch_dev, void *stream);
int pr_grad_norm_scattered(const pr_adamw_plan *plan, const void *const *grad_dev, float prescale,
                           float max_norm, float *out_dev, double *scratch_dev, void *stream);

/* ---------------------------------------------------------------- K3: checkpoint -----
 * PR-SUM64, the on-GPU checksum that stands in for hashlib.md5 over the file -- reference
 * pyrecover/checkpoint.py:79-80 (save) and :162-163 (load).  The byte stream is read as
 * little-endian u32 words w_i (tail zero-padded); mod 2^64:
 *     A = sum w_i          B = sum (word_base + i + 1) * w_i
 * sum_dev[0] += A, sum_dev[1] += B  (the caller zeroes sum_dev; partial sums over chunks
 * taken with their own word_base add up to the sum of the whole).
 */
int pr_checksum(const void *data_dev, uint64_t nbytes, uint64_t word_base, uint64_t *sum_dev,
                void *stream);

/* A pack plan describes the static layout "n tensors -> one packed image": the role of
 * model.state_dict() / optimizer.state_dict() feeding torch.save -- reference
 * pyrecover/checkpoint.py:60-74 (vanilla) and :254-274 (torch.distributed.checkpoint).
 * dst_off[i] must be a multiple of 16; bytes between segments are not written (keep the
 * image zero-initialised).  The plan owns a small device table (scratch memory). */
typedef struct pr_pack_plan pr_pack_plan;

int pr_pack_plan_create(const void *const *tensor_dev, const uint64_t *nbytes,
                        const uint64_t *dst_off, int n, uint64_t image_bytes,
                        pr_pack_plan **plan_out);
int pr_pack_plan_destroy(pr_pack_plan *plan);

/* Gather every tensor of the plan into image (device memory, or mapped pinned host memory)
 * and add PR-SUM64 of the packed image into sum_dev (2 x u64, zeroed by the caller; may be
 * NULL to skip the checksum).  Replaces the per-storage synchronous D2H of torch.save --
 * reference pyrecover/checkpoint.py:74. */
int pr_ckpt_pack(const pr_pack_plan *plan, void *image, uint64_t *sum_dev, void *stream);

/* Scatter the packed image back into the plan's tensors and add PR-SUM64 of the image
 * into sum_dev: the role of torch.load + load_state_dict -- reference
 * pyrecover/checkpoint.py:182-199 -- with the verification of :155-178 folded in. */
int pr_ckpt_unpack(const pr_pack_plan *plan, const void *image, uint64_t *sum_dev,
                   void *stream);

/* The same two calls with the kernel and its grid chosen by the caller.
 *   PR_PACK_BULK: the bytes move as bulk asynchronous copies staged through shared memory
 *     (cp.async.bulk global -> shared -> global; one elected thread per CTA issues them, mbarriers track
 *     them, the checksum is taken from the shared-memory tile) -- the north_star's "TMA-staged" pack.  A few
 *     dozen CTAs carry it, so the snapshot that runs on a side stream beside the next forward pass
 *     (reference: the save the training loop blocks on at train.py:310-340) takes few SMs, no registers
 *     and no issue slots from it.  Needs every tensor base 16-byte aligned and the image in device memory.
 *   PR_PACK_LDG: 16-byte loads / stores through registers on all SMs; any alignment, image may be mapped
 *     pinned host memory.
 *   PR_PACK_AUTO (what pr_ckpt_pack / pr_ckpt_unpack use): BULK when its two conditions hold, else LDG.
 * max_blocks bounds the grid (0 = the kernel's default: PR_CKPT_BULK_BLOCKS or 64 CTAs for BULK, 3 CTAs
 * per SM for LDG).  Results (image bytes, tensors, PR-SUM64) are identical whichever kernel runs. */
typedef enum pr_pack_algo { PR_PACK_AUTO = 0, PR_PACK_LDG = 1, PR_PACK_BULK = 2 } pr_pack_algo;
int pr_ckpt_pack_ex(const pr_pack_plan *plan, void *image, uint64_t *sum_dev, int algo, int max_blocks,
                    void *stream);
int pr_ckpt_unpack_ex(const pr_pack_plan *plan, const void *image, uint64_t *sum_dev, int algo,
                      int max_blocks, void *stream);

/* pr_ckpt_pack for callers that must not allocate.  (No counterpart in the reference: its
 * save path allocates freely.)  A checkpoint matters most when it is hardest to take: in the
 * walltime / preemption handler of a job that has filled HBM, where the cudaMalloc behind
 * pr_pack_plan_create can fail or stall the device, or in a process whose memory manager forbids
 * device allocation after start-up.  Here the tile table lives in a fixed 288 KiB array inside the
 * library (held on every GPU that loads it, until exit) and the call performs no cudaMalloc,
 * cudaFree or other allocating call; the caller still supplies the tensors, the image and sum_dev.
 *   Limits: n <= 1024 tensors and at most 65536 tiles of 64 KiB (4 GiB of payload when tensors
 *   are tile multiples); beyond either the call fails with PR_ERR_INVALID and says so.
 *   Every call resolves the array's device address for the current device.  The FIRST call on a device
 *   is also what loads the kernel image and materialises the 288 KiB array there (CUDA loads modules
 *   lazily): make one small warm-up call per device at start-up -- before memory fills, and before any
 *   stream capture begins -- so that the call that matters later finds everything in place.
 *   All calls share the one array and it keeps the previous call's table; each call rewrites every
 *   entry it reads, so nothing needs clearing -- but two calls whose GPU work overlaps (different
 *   streams, or threads racing on one device) would overwrite each other's table: serialise them.
 * Same results as pr_pack_plan_create + pr_ckpt_pack: image bytes and PR-SUM64 added into sum_dev
 * (zeroed by the caller; NULL skips the checksum). */
int pr_ckpt_pack_noalloc(const void *const *tensor_dev, const uint64_t *nbytes,
                         const uint64_t *dst_off, int n, void *image, uint64_t image_bytes,
                         uint64_t *sum_dev, void *stream);

/* Pinned, device-mapped host arena the snapshot lands in (replaces the pageable staging
 * torch.save does implicitly -- reference pyrecover/checkpoint.py:74).  Zero-filled. */
int pr_host_arena_alloc(uint64_t nbytes, void **host_ptr_out);
int pr_host_arena_free(void *host_ptr);
/* Device-visible alias of an arena pointer (for pr_ckpt_pack's direct-to-host mode). */
int pr_host_arena_device_ptr(void *host_ptr, void **dev_ptr_out);
/* Async copies in <= chunk_bytes pieces on `stream` (0 = one piece). */
int pr_copy_d2h_async(void *dst_host, const void *src_dev, uint64_t nbytes,
                      uint64_t chunk_bytes, void *stream);
int pr_copy_h2d_async(void *dst_dev, const void *src_host, uint64_t nbytes,
                      uint64_t chunk_bytes, void *stream);

/* Native file writer: streams `n` (host pointer, length) pieces to `path + ".tmp"` with
 * large write()s, fsyncs and renames to `path` -- the durable half of
 * pyrecover/checkpoint.py:74 without the non-atomic in-place write.  Blocking; call it
 * from a helper thread (it never touches the Python GIL). */
int pr_file_write_atomic(const char *path, const void *const *pieces_host,
                         const uint64_t *piece_bytes, int n, int do_fsync);

/* ---------------------------------------------------------------- K1: allreduce ------
 * Replaces the DDP Reducer's bucket allreduce -- reference train.py:110-115 (DDP ctor),
 * firing inside loss.backward() at train.py:269: g <- sum_r g_r * (1/W) per flat bucket
 * (torch/csrc/distributed/c10d/reducer.hpp:500).  One process per GPU; every rank maps
 * every peer's symmetric regions and flag words over NVLink (CUDA IPC); the kernels pull
 * 16-byte vectors from all peers, accumulate in fp32 in rank order 0..W-1 (every rank gets
 * bit-identical results), scale, round once and push the result to every rank.
 * NCCL/gloo are used only by the caller, to exchange the export blobs.
 * With NVSwitch multicast (pr_comm_multicast) the sum is taken by the switch instead: identical on
 * every rank, but in an unspecified order (summation-order tolerance vs the rank-order sum) and,
 * for bf16, rounded to bf16 before the scale is applied (exact when the scale is a power of two).
 */
typedef struct pr_comm pr_comm;

int pr_comm_create(int rank, int world, pr_comm **comm_out);
/* Emulated communicator: all `world` ranks live on the current device and one cooperative
 * launch runs them side by side (single-GPU tests of the same kernels and flag protocol). */
int pr_comm_create_emulated(int world, pr_comm **comm_out);
int pr_comm_destroy(pr_comm *comm);
/* Allocate symmetric region `region` (cudaMalloc, 256-B aligned, zero-filled). */
int pr_comm_region_alloc(pr_comm *comm, int region, uint64_t nbytes);
/* Local base pointer of a region (rank = the emulated rank, ignored for a real comm). */
int pr_comm_region_ptr(pr_comm *comm, int region, int rank, void **dev_ptr_out);
/* Write this rank's PR_EXPORT_BLOB_BYTES handle blob (after all regions are allocated). */
int pr_comm_export(pr_comm *comm, void *blob_out);
/* Import all ranks' blobs (world * PR_EXPORT_BLOB_BYTES, rank-major) and map the peers. */
int pr_comm_import(pr_comm *comm, const void *blobs);
/* 1 in *timed_out if a kernel of this communicator gave up waiting for a peer. */
int pr_comm_error(pr_comm *comm, int *timed_out);
/* NVSwitch multicast (NVLS).  Where the driver supports it (and PR_COMM_MULTICAST is not "0" in the
 * environment at pr_comm_create), regions are CUDA VMM allocations: pr_comm_import then ships them as
 * POSIX file descriptors over abstract unix sockets instead of CUDA-IPC handles, binds each region
 * to a multicast object and maps it.  The summing then happens inside the switch -- the job NCCL's
 * NVLS algorithm does for the reference's DDP allreduce (train.py:110-115 / :269).
 * policy: -2 = only query, -1 = automatic (use multicast for world >= 4, where it moves fewer bytes
 * per link than P2P), 0 = never, 1 = whenever available.  *available: the mappings exist;
 * *in_use: PR_AR_AUTO, pr_allreduce_adamw and pr_allgather_slices take the multicast kernels.
 * Every rank must choose the same policy.  Either out pointer may be NULL. */
int pr_comm_multicast(pr_comm *comm, int policy, int *available, int *in_use);

/* A barrier that times out (PR_COMM_TIMEOUT_S, default 20 s: a peer died or is stuck) makes the kernel
 * return BEFORE it reads or writes any gradient, parameter or moment, raises the error word that
 * pr_comm_error reports and tells every peer to give up as well; the communicator refuses further work
 * (PR_ERR_STATE).  State that a late peer might have half-written is never consumed.
 *
 * Statistics of the kernels launched on this communicator so far (read with a synchronising copy: call
 * it outside the training loop): out3[0] = nanoseconds CTA 0 spent in the begin barrier waiting for the
 * slowest rank, out3[1] = launches, out3[2] = nanoseconds CTA 0 ran in total.  Separates rank skew from
 * transfer time.  reset != 0 zeroes the counters. */
int pr_comm_stats(pr_comm *comm, uint64_t *out3, int reset);

/* One word from rank 0 to every rank without a collective of its own and without a host synchronisation:
 * the job of the per-step `dist.broadcast(should_stop)` + `.item()` of reference train.py:342-346.
 * pr_comm_signal_post (rank 0; other ranks' posts are ignored) enqueues a store of `value` into rank 0's
 * signal word on `stream`.  The begin barrier of every later collective of this communicator reads it on
 * every rank and drops (launch sequence number << 32 | value) into a host-mapped word;
 * pr_comm_signal_read returns that pair without touching the GPU.  pr_comm_seq is the sequence number of
 * the most recent launch: a value read with seq_out >= the pr_comm_seq noted after some step's exchange
 * was posted no later than that step. */
int pr_comm_signal_post(pr_comm *comm, uint32_t value, void *stream);
int pr_comm_signal_read(pr_comm *comm, uint32_t *seq_out, uint32_t *value_out);
int pr_comm_seq(pr_comm *comm, uint32_t *seq_out);
/* Test aid (emulated communicators only): rank `rank` never reaches a barrier, as if its process had
 * died; -1 = everybody present. */
int pr_comm_test_absent_rank(pr_comm *comm, int rank);

typedef enum pr_ar_algo {
    PR_AR_AUTO = 0,       /* one-shot <= 512 KiB, else multimem when pr_comm_multicast says in_use, else two-shot */
    PR_AR_ONE_SHOT = 1,
    PR_AR_TWO_SHOT = 2,
    PR_AR_MULTIMEM = 3    /* NVSwitch in-switch reduction (multimem.ld_reduce / multimem.st); PR_ERR_UNSUPPORTED
                             when the communicator has no multicast mapping */
} pr_ar_algo;

/* In-place allreduce of region[offset, offset+nbytes) with result = round(sum * scale).
 * offset and nbytes must be multiples of 16.  max_blocks bounds the grid (0 = one CTA per
 * SM); every rank must pass identical arguments in the same order. */
int pr_allreduce(pr_comm *comm, int region, uint64_t offset, uint64_t nbytes, int dtype,
                 float scale, int algo, int max_blocks, void *stream);

/* Fused reduce-scatter -> AdamW -> all-gather: rank r reduces slice r of the gradient
 * region in fp32 registers, applies AdamW to slice r of (p, m, v) and pushes the updated
 * parameters to every rank's parameter region.  Replaces train.py:269 + train.py:274 in
 * one launch.  The moments live in two more symmetric regions of which a rank only ever
 * touches its own slice (sharded optimizer state; pr_allgather_slices re-replicates them
 * when a rank-0 checkpoint needs the whole).  n = elements, a multiple of 8; the gradient
 * mean's 1/W goes in args->grad_scale; args->numerics must be 0. */
int pr_allreduce_adamw(pr_comm *comm, int grad_region, int param_region, int m_region,
                       int v_region, uint64_t elem_offset, uint64_t n, int dtype,
                       const pr_adamw_args *args, int max_blocks, void *stream);

/* All-gather of per-rank slices of a region (slice r of rank r -> every rank): used to
 * re-replicate sharded moments before a rank-0 vanilla checkpoint (the reference's
 * save_ckpt_vanilla writes the whole optimizer state from rank 0, checkpoint.py:58-74). */
int pr_allgather_slices(pr_comm *comm, int region, uint64_t offset, uint64_t nbytes,
                        int max_blocks, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PYRECOVER_B200_H */
