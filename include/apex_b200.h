/*
 * apex_b200.h — C ABI of libapex_b200.so
 *
 * The B200-native (sm_100a) replacement for the native layer that the reference
 * script reaches through NVIDIA/apex (pinned at f3a960f8 by /root/reference/README.md:16-19,
 * /root/reference/test_apex_distributed_spawn.py:13-16): the pybind modules `amp_C`
 * (multi_tensor_scale / axpby / sgd / adam) and `apex_C` (flatten / unflatten), plus the
 * host-side LossScaler state machine that apex keeps in Python.
 *
 * The reference call sites that end up here:
 *   test_apex_distributed_spawn.py:160 (`with amp.scale_loss(...)`, exit)  -> unscale + overflow check + scale update
 *   test_apex_distributed_spawn.py:161 (`scaled_loss.backward()`, DDP hooks) -> bucket flatten + allreduce epilogue
 *   test_apex_distributed_spawn.py:164 (`optimizer.step()`)                 -> fused Adam / SGD + master->model copy
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / pybind types.
 *   - every pointer named dev_* is a device address, host_* a host address.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - kernels never allocate, free or synchronise; they launch on `stream` and return.
 *   - return value: 0 on success, a positive cudaError_t, or a negative B200_E* code.
 *     b200_last_error() returns a thread-local message for the last non-zero return.
 *   - numerical overflow of gradients is NOT an error: it raises the device flag and the
 *     step is skipped on the device (same contract as apex's noop_flag / torch's found_inf).
 */
#ifndef APEX_B200_H_
#define APEX_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 1

/* element types */
#define B200_F32  0
#define B200_F16  1
#define B200_BF16 2
#define B200_NONE (-1)          /* "column absent" (e.g. no fp16 model copy) */

/* error codes (negative; positive values are cudaError_t) */
#define B200_EINVAL   (-1)      /* bad argument */
#define B200_ESIZE    (-2)      /* buffer too small / size beyond a stated maximum */
#define B200_EDTYPE   (-3)      /* dtype combination not instantiated */
#define B200_ENODEV   (-4)      /* no usable sm_100 device */

#define B200_MAX_DEPTH   5      /* pointer columns per tensor row */
#define B200_MAX_GROUPS  8      /* hyper-parameter groups one launch can carry */
#define B200_TILE_ELEMS_DEFAULT 4096

int32_t     b200_abi_version(void);
const char* b200_last_error(void);

/* ------------------------------------------------------------------------------------------
 * S1 — chunk table ("multi_tensor_apply" scheduler).
 * Replaces apex csrc/multi_tensor_apply.cuh (K1 in SURVEY.md §2.4), which re-packs tensor
 * addresses into a <4 KB kernel-argument struct on every call and needs ceil(chunks/320)
 * launches.  Here the list is described ONCE by a table that lives in device memory; every
 * kernel below takes the table's device address and covers the whole list in one launch.
 *
 * Image layout (identical bytes on host and device):
 *   b200_table_header | b200_table_row[n_tensors] | b200_table_tile[n_tiles]
 * A row holds up to B200_MAX_DEPTH column pointers for one tensor (what each column means is
 * fixed by the kernel that consumes the table, see below), its element count and its
 * hyper-parameter group.  A tile is (tensor index, tile index inside the tensor); tile k of a
 * tensor covers elements [k*tile_elems, min(numel,(k+1)*tile_elems)).
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_table_header {
    uint32_t magic;             /* 0x42325442 "B2TB" */
    uint32_t version;           /* B200_ABI_VERSION */
    uint32_t n_tensors;
    uint32_t depth;
    uint32_t tile_elems;
    uint32_t n_tiles;
    uint64_t rows_off;          /* byte offset of the row array from the image start */
    uint64_t tiles_off;         /* byte offset of the tile array */
    uint64_t total_elems;
    uint64_t image_bytes;
    uint64_t reserved;
} b200_table_header;            /* 64 bytes */

typedef struct b200_table_row {
    void*    ptr[B200_MAX_DEPTH];
    int64_t  numel;
    uint32_t group;
    uint32_t flags;             /* bit 0: every column is aligned for the 128/256-bit path */
    uint64_t reserved;          /* the row's index in the list (set by b200_table_build) */
} b200_table_row;               /* 64 bytes */

typedef struct b200_table_tile {
    uint32_t tensor;
    uint32_t tile;
} b200_table_tile;              /* 8 bytes */

/* Bytes of the image for a list with these element counts (zero-sized tensors get no tile).
 * Returns <0 on error. */
int64_t b200_table_bytes(int32_t n_tensors, const int64_t* numel, int32_t tile_elems);

/* Fill `host_image` (at least b200_table_bytes() bytes).  `ptrs` is column-major:
 * ptrs[d * n_tensors + t] is column d of tensor t (device addresses; may be NULL for an
 * absent column).  elem_size[d] is the byte size of column d's element type (2 or 4) and
 * decides the row's alignment flag.  `group` may be NULL (all zero).  Pure host code: usable
 * without a GPU.  Returns the number of tiles (>=0) or a negative error. */
int64_t b200_table_build(void* host_image, int64_t image_bytes,
                         int32_t n_tensors, int32_t depth,
                         const void* const* ptrs, const int32_t* elem_size,
                         const int64_t* numel, const int32_t* group,
                         int32_t tile_elems);

/* Convenience for non-torch callers: cudaMemcpyAsync(host image -> device) on `stream`. */
int32_t b200_table_upload(void* dev_image, const void* host_image, int64_t image_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * S4 — loss-scaler state, resident on the device.
 * Replaces apex/amp/scaler.py LossScaler.update_scale (a6 in SURVEY.md §8a), which reads the
 * overflow flag with .item() (a full stream sync per step).  Here the flag, the scale and the
 * window counter stay in HBM; the optimizer kernels read `found_inf` / `inv_scale` directly.
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_scaler_state {
    float   loss_scale;         /* scale the NEXT backward is multiplied with */
    float   inv_scale;          /* 1 / (scale of the backward whose grads await the step) */
    int32_t found_inf;          /* 1: skip the pending step */
    int32_t unskipped;          /* clean steps since the last change of scale */
    int32_t overflow_acc;       /* raised by check / scale / epilogue kernels; cleared by update */
    int32_t seq;                /* number of updates so far (lets the host see fresh mirrors) */
    int32_t had_overflow;       /* overflow_acc as seen by the last update (also set when static) */
    float   inv_loss_scale;     /* 1 / loss_scale: what an unscale BEFORE the next update multiplies by */
    int32_t skipped_total;      /* updates so far that ended in a skip (lets a late host count them) */
    int32_t reserved[3];
} b200_scaler_state;            /* 48 bytes */

/* Initialise *dev_state (loss_scale = init_scale, inv_scale = inv_loss_scale = 1/init_scale, rest 0). */
int32_t b200_scaler_init(b200_scaler_state* dev_state, float init_scale, void* stream);

/* `dynamic` argument of b200_scaler_update: 0 = static scale, 1 = dynamic (the two values ABI version 1
 * has always taken).  OR-ing in B200_SCALER_KEEP_SKIP says "no optimizer step has run since the previous
 * update" (a further backward of a gradient-accumulation cycle): a pending found_inf then stays set, the
 * way apex's one-shot skip_step stays patched in until step() is called. */
#define B200_SCALER_DYNAMIC   1
#define B200_SCALER_KEEP_SKIP 2

/* One LossScaler.update_scale() on the device:
 *   flag = overflow_acc != 0;  inv_scale = 1/loss_scale;  found_inf = dynamic && flag  (| pending one, see above)
 *   dynamic: flag ? (loss_scale = max(min_scale, loss_scale/2), unskipped = 0) : ++unskipped;
 *            unskipped == scale_window ? (loss_scale = min(max_scale, loss_scale*scale_factor), unskipped = 0)
 *   overflow_acc = 0; ++seq; inv_loss_scale = 1/loss_scale
 * min_scale <= 0 means "no minimum".  If `mirror` is non-NULL the new state is also stored
 * there (a device-accessible host address, e.g. pinned memory) so the host can print apex's
 * "Gradient overflow. Skipping step" line later without synchronising. */
int32_t b200_scaler_update(b200_scaler_state* dev_state, int32_t dynamic,
                           int32_t scale_window, float scale_factor,
                           float min_scale, float max_scale,
                           b200_scaler_state* mirror, void* stream);

/* ------------------------------------------------------------------------------------------
 * amp_C.multi_tensor_scale (K2): column 0 = in, column 1 = out.  out = in * scale, and
 * *dev_flag = 1 if any INPUT element is inf/NaN.  Used for the generic unscale
 * (fp16 model grad -> fp32 master grad), the master->model weight copy (scale 1) and the DDP
 * flatten (grads -> bucket views).  If dev_scale is non-NULL the factor is read from the
 * device (e.g. &state->inv_scale) and `scale` is ignored.
 * ------------------------------------------------------------------------------------------ */
int32_t b200_multi_tensor_scale(const void* dev_table, int64_t n_tiles,
                                int32_t in_dtype, int32_t out_dtype,
                                float scale, const float* dev_scale,
                                int32_t* dev_flag, void* stream);

/* amp_C.multi_tensor_axpby (K3): columns x, y, out.  out = a*x + b*y; the flag is raised on
 * non-finite x (arg_to_check 0), y (1) or either (-1). */
int32_t b200_multi_tensor_axpby(const void* dev_table, int64_t n_tiles,
                                int32_t x_dtype, int32_t y_dtype, int32_t out_dtype,
                                float a, float b, int32_t arg_to_check,
                                int32_t* dev_flag, void* stream);

/* Overflow check only (read-only pass over column `col`): *dev_flag = 1 if any element is
 * inf/NaN.  This is the "check" half of apex's unscale when the arithmetic half is fused into
 * the optimizer kernel below. */
int32_t b200_multi_tensor_check(const void* dev_table, int64_t n_tiles,
                                int32_t col, int32_t dtype,
                                int32_t* dev_flag, void* stream);

/* amp_C.multi_tensor_l2norm fused with the overflow check: ONE read-only
 * pass over column `col` adds the list's sum of squares to *dev_sumsq and raises *dev_flag on
 * inf/NaN.  Block partials go through `dev_scratch` (>= B200_NORM_SCRATCH_BYTES, zeroed once by
 * the caller) and are summed in block order by the last block to finish — deterministic for a given
 * grid, no floating-point atomics.  Several lists (dtypes) accumulate into the same *dev_sumsq.
 *
 * b200_clip_scale turns the accumulated sum of squares into the factor the step kernels take as
 * `dev_inv_scale`:  norm = sqrt(sumsq) * inv;  scale_out = inv * min(1, max_norm / (norm + 1e-6))
 * (torch.nn.utils.clip_grad_norm_'s rule), with inv = dev_inv_scale ? *dev_inv_scale : inv_scale.
 * It zeroes *dev_sumsq for the next step.  Clipping therefore costs no extra pass over the
 * gradients and no fp32 master gradients: the unscale, the clip and the update share one kernel. */
#define B200_NORM_SCRATCH_BYTES 16384
int32_t b200_multi_tensor_l2norm(const void* dev_table, int64_t n_tiles, int32_t col, int32_t dtype,
                                 float* dev_sumsq, void* dev_scratch, int64_t scratch_bytes,
                                 int32_t* dev_flag, void* stream);
int32_t b200_clip_scale(float* dev_sumsq, const float* dev_inv_scale, float inv_scale, float max_norm,
                        float* dev_scale_out, float* dev_norm_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * S2 — fused optimizer step: unscale + Adam/SGD + master->model cast in ONE pass.
 * Replaces the three passes apex makes per step: multi_tensor_scale
 * (unscale, a5), multi_tensor_adam / multi_tensor_sgd (K4/K5, a7) and multi_tensor_scale
 * again (master->model copy, a8).
 *
 * Table columns: 0 = grad (read), 1 = param / fp32 master (read+write), 2 = exp_avg or
 * momentum (rw), 3 = exp_avg_sq (rw, Adam only), last = model copy (write, optional):
 *   Adam: depth 4 (g,p,m,v) or 5 (g,p,m,v,copy);  SGD: depth 3 (g,p,mom) or 4 (g,p,mom,copy).
 * Grad handling: g = float(grad) * s, s = dev_inv_scale ? *dev_inv_scale : inv_scale.
 * Skip: if dev_found_inf is non-NULL and *dev_found_inf != 0 the kernel writes nothing.
 * Arithmetic is fp32 with IEEE sqrt/div and no fused multiply-add, in torch.optim's order
 * (TORCH/optim/adam.py:414-547, TORCH/optim/sgd.py:340-380), so results are bit-identical to
 * oracle/amp_step_oracle.c.
 * ------------------------------------------------------------------------------------------ */
/* Exact step numbers without a host sync.
 * apex decides "skip" on the host, so its step counter (Adam's bias correction, SGD's first_run)
 * only advances on steps that really ran.  Here the decision is taken on the device, after the
 * host has already enqueued the step kernel, so the host cannot know the exact step number when
 * it computes the hyper-parameters.  b200_step_sync closes the gap: the kernel keeps
 * {calls, skips} in `dev_counters` (and mirrors them to `mirror`); the host passes the skips it
 * had observed (`known_skips`) and derives every per-step constant for the B200_STEP_WINDOW
 * candidates step, step-1, ... ; the kernel picks candidate (skips_now - known_skips).
 * The caller must not let more than B200_STEP_WINDOW-1 launches be unobserved (it synchronises
 * if the mirror lags that far; in steady state it never does).  NULL = the host's `step` is exact. */
#define B200_STEP_WINDOW 4
typedef struct b200_step_sync {
    int32_t* dev_counters;      /* int32[2] {calls, skips}, zeroed by the caller before first use */
    int32_t* mirror;            /* device-accessible host int32[2] (pinned), or NULL */
    int32_t  known_skips;       /* skips already reflected in the groups' `step` */
    int32_t  read_only;         /* 1: use the counters but do not advance them (2nd, 3rd ... launch
                                   of one optimizer step that is split over several dtype sets) */
} b200_step_sync;

typedef struct b200_adam_group {
    double  lr, beta1, beta2, eps, weight_decay;
    int64_t step;               /* 1-based number of this step if no unobserved skip happened */
    int32_t bias_correction;    /* 0: bc1 = bc2 = 1 */
    int32_t adam_w_mode;        /* 1: decoupled decay (AdamW, apex default); 0: L2 */
} b200_adam_group;

int32_t b200_fused_adam(const void* dev_table, int64_t n_tiles,
                        int32_t grad_dtype, int32_t state_dtype, int32_t copy_dtype,
                        const b200_adam_group* groups, int32_t n_groups,
                        float inv_scale, const float* dev_inv_scale,
                        const int32_t* dev_found_inf, const b200_step_sync* sync,
                        void* stream);

typedef struct b200_sgd_group {
    double  lr, momentum, dampening, weight_decay;
    int64_t step;               /* as above; step 1 is apex's first_run (momentum buffer = grad) */
    int32_t nesterov;
    int32_t wd_after_momentum;
} b200_sgd_group;

int32_t b200_fused_sgd(const void* dev_table, int64_t n_tiles,
                       int32_t grad_dtype, int32_t state_dtype, int32_t copy_dtype,
                       const b200_sgd_group* groups, int32_t n_groups,
                       float inv_scale, const float* dev_inv_scale,
                       const int32_t* dev_found_inf, const b200_step_sync* sync,
                       void* stream);

/* ------------------------------------------------------------------------------------------
 * S3 — DDP bucket epilogue.  Replaces apex DistributedDataParallel.allreduce_bucket's
 * `mul_(1/world_size)` + apex_C.unflatten + multi_tensor_scale copy-back (a4) and folds the
 * overflow check of a5 into the same pass: in place on the flat bucket,
 *   x = dtype(float(x) * factor);  *dev_flag = 1 if any result is inf/NaN.
 * factor == 1 makes it a read-only check.  The bucket is one contiguous buffer, so no table.
 * ------------------------------------------------------------------------------------------ */
int32_t b200_bucket_epilogue(void* dev_bucket, int64_t numel, int32_t dtype,
                             float factor, int32_t* dev_flag, void* stream);

/* ------------------------------------------------------------------------------------------
 * S3' — gradient mean over NVLink peer memory, fused with its neighbours (no counterpart in
 * apex, which calls ncclAllReduce on a flattened copy and then multiplies by 1/world).
 * Every rank keeps its gradients in one flat buffer G_r that all peers can address (CUDA IPC;
 * NVSwitch gives every pair full bandwidth).  The all-reduce is split where the data flow splits:
 *
 *   b200_peer_reduce_shard  (reduce-scatter + mean + overflow check, ONE kernel):
 *       rank r sums its shard [begin,end) of all `world` buffers — fp32 accumulation in rank
 *       order 0..world-1, one rounding to the gradient dtype, then x factor with a second
 *       rounding (apex: SUM in the gradient dtype, then mul_(1/world)) — writes the result at
 *       the same offsets of its own `reduced` buffer and raises *dev_flag on inf/NaN.
 *   b200_fused_adam_peer / b200_fused_sgd_peer  (all-gather fused INTO the optimizer):
 *       the step kernels read each tile's gradient straight from the `reduced` buffer of the rank
 *       that owns it (local HBM for the own shard, NVLink loads for the others), so the
 *       "gather" overlaps the HBM-bound update tile by tile and is never materialised.
 *
 * Between the two the caller needs a cross-rank barrier that also ORs the overflow flags (a
 * 4-byte NCCL all-reduce(MAX) on the stream does both), and one more before the next reduce.
 * Shard bounds must not split a tile of the optimizer's table (tables.tile_aligned_shards()).
 * ------------------------------------------------------------------------------------------ */
#define B200_MAX_PEERS 8
typedef struct b200_peer_pull {
    const void* local_base;                 /* this rank's flat buffer: column 0 of the table points into it */
    int64_t     bounds[B200_MAX_PEERS + 1]; /* rank r owns elements [bounds[r], bounds[r+1]) */
    int64_t     delta[B200_MAX_PEERS];      /* bytes from local_base to rank r's reduced buffer (this process's view) */
    int32_t     world;
    int32_t     elem_size;                  /* bytes per gradient element (2 or 4) */
    uint32_t    first_tile;                 /* index (in the table's tile list) of this rank's first own tile: the
                                               kernel starts its sweep there, so that at any moment the ranks pull
                                               from DIFFERENT owners instead of all from the same one */
    uint32_t    reserved;
} b200_peer_pull;

/* Set-up / tear-down of the peer-addressable buffers (the only entry points that allocate; call
 * them once, outside the step).  b200_peer_alloc: cudaMalloc + cudaIpcGetMemHandle (64-byte
 * handle to ship to the other ranks).  b200_peer_open: map a peer's buffer into this process with
 * peer access enabled.  Close every mapping (all ranks, then a barrier) before the owner frees. */
#define B200_IPC_HANDLE_BYTES 64
int32_t b200_peer_alloc(int64_t bytes, void** dev_ptr, void* handle_out);
int32_t b200_peer_open(const void* handle, void** dev_ptr);
int32_t b200_peer_close(void* dev_ptr);
int32_t b200_peer_free(void* dev_ptr);

/* In-kernel cross-GPU hand-shake (optional; replaces the two NCCL barrier all-reduces).
 * Every rank owns one 128-byte signal pad in peer-addressable memory, zeroed at set-up.  With a
 * b200_peer_sync, b200_peer_reduce_shard (a) tells every rank "my gradients of `epoch` are in
 * place" and waits until every rank has said so before it reads their buffers, and (b) has its
 * last block publish "my shard of `epoch` is written" plus this rank's overflow bit to every pad;
 * b200_peer_wait_flags (one tiny kernel, next on the stream) waits for all ranks' (b) and ORs the
 * overflow bits into *dev_flag.  `epoch` starts at 1 and grows by one per step on every rank.
 * Waits are bounded (~30 s): a rank that never arrives sets pad->error instead of hanging the GPU, and the
 * kernel that gave up raises the overflow flag without reading any peer buffer, so the step that follows is
 * SKIPPED (nothing is written from incomplete sums); pad->error stays set, every later step is skipped too,
 * and the host reports it (apex_example_b200.parallel.peer.PeerBuffers.check_signals / close). */
typedef struct b200_peer_signal_pad {
    int32_t ready[B200_MAX_PEERS];          /* ready[r]: highest epoch rank r has announced */
    int32_t done[B200_MAX_PEERS];           /* done[r]:  (epoch << 1) | overflow bit of rank r */
    int32_t error;                          /* set when a bounded wait gave up */
    int32_t local_flag, ticket;             /* kernel-internal */
    int32_t reserved[13];
} b200_peer_signal_pad;                     /* 128 bytes */

typedef struct b200_peer_sync {
    b200_peer_signal_pad* const* pads;      /* [world]: rank r's pad as addressable from this process */
    int32_t world, rank, epoch, reserved;
} b200_peer_sync;

int32_t b200_peer_reduce_shard(const void* const* peer_bufs, int32_t world, void* reduced,
                               int64_t begin, int64_t end, int32_t dtype, float factor,
                               int32_t* dev_flag, const b200_peer_sync* sync, void* stream);
int32_t b200_peer_wait_flags(const b200_peer_sync* sync, int32_t* dev_flag, void* stream);

/* The all-gather half as a kernel of its own, for callers that need the averaged gradient
 * materialised (apex.parallel.DistributedDataParallel keeps apex's contract that param.grad
 * holds the mean after backward): waits for every rank's shard (if `sync`), ORs the overflow bits
 * into *dev_flag and copies each element of [0, bounds[world]) from its owner's `reduced` buffer
 * into `local_out`.  bounds[1..world-1] must be multiples of 8 or equal to bounds[world] (empty trailing
 * shards of a small bucket).  b200_peer_reduce_shard +
 * b200_peer_gather is a complete all-reduce over peer memory in two launches, no NCCL. */
int32_t b200_peer_gather(const void* const* peer_reduced, int32_t world, void* local_out,
                         const int64_t* bounds, int32_t dtype, int32_t* dev_flag,
                         const b200_peer_sync* sync, void* stream);

/* Cap the grid of the peer kernels (0 = fill the GPU).  Under DDP they overlap with backward and
 * spin while they wait for slower ranks, so they are run as a small "communication" grid. */
int32_t b200_set_peer_grid(int32_t max_ctas);

int32_t b200_fused_adam_peer(const void* dev_table, int64_t n_tiles,
                             int32_t grad_dtype, int32_t state_dtype, int32_t copy_dtype,
                             const b200_adam_group* groups, int32_t n_groups,
                             float inv_scale, const float* dev_inv_scale,
                             const int32_t* dev_found_inf, const b200_step_sync* sync,
                             const b200_peer_pull* pull, void* stream);

int32_t b200_fused_sgd_peer(const void* dev_table, int64_t n_tiles,
                            int32_t grad_dtype, int32_t state_dtype, int32_t copy_dtype,
                            const b200_sgd_group* groups, int32_t n_groups,
                            float inv_scale, const float* dev_inv_scale,
                            const int32_t* dev_found_inf, const b200_step_sync* sync,
                            const b200_peer_pull* pull, void* stream);

/* ------------------------------------------------------------------------------------------
 * S3''' — sharded step: reduce-scatter + speculative Adam on the own shard + all-gather of the
 * new 16-bit weights, ONE kernel per rank.  (Counterpart in apex: apex.contrib.optimizers.
 * DistributedFusedAdam — ZeRO-1 style sharding of the optimizer state over the data-parallel ranks; not used
 * by the reference script, whose ranks keep full replicas, test_apex_distributed_spawn.py:119-122.)
 *
 * The table lists ONLY this rank's shard (tensor segments clipped to the shard):
 *   column 0 = the segment inside this rank's flat gradient buffer   (fan->grads[rank])
 *   columns 1-3 = fp32 master / m / v of the segment, double-buffered as for b200_fused_adam_spec
 *   column 4 = the segment inside this rank's flat model-weight buffer (fan->models[rank])
 * For every element the kernel reads the gradient of EVERY rank at the same offset (fan->grads[r]), forms
 * the mean exactly like b200_peer_reduce_shard (fp32 sum in rank order, rounded to the gradient type,
 * times `factor`, rounded), inspects it, writes it back over this rank's own gradient (column 0: after the step
 * the shard of the local gradient buffer holds the mean), runs the speculative update, and stores the new
 * weights at the same offset of EVERY rank's model buffer (fan->models[r]).  With `psync`: the kernel first announces /
 * awaits "gradients in place, model weights no longer read" and its last block publishes done[rank] =
 * (epoch << 1) | overflow to every pad; then call b200_peer_wait_flags(psync, &state->overflow_acc) and
 * b200_spec_commit, and b200_peer_shard_repair with epoch + 1 (so `epoch` grows by TWO per step).  Without
 * `psync` (all buffers on one GPU: tests) the overflow is raised in *dev_flag.
 * b200_peer_shard_repair exits at once unless *dev_found_inf != 0 (the same on every rank after the
 * commit); otherwise it re-sends cast(current master) of the shard to every rank and waits until every rank
 * has done so.  Bounded waits; a rank that never arrives poisons the step (it is skipped).
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_peer_fan {
    const void* grads[B200_MAX_PEERS];      /* rank r's flat gradient buffer as addressable from this process */
    void*       models[B200_MAX_PEERS];     /* rank r's flat model-weight buffer */
    int32_t     world, rank;
    float       factor;                     /* 1/world for the mean, 1 for the sum */
    int32_t     reserved;
    void*       model_mc;                   /* optional: MULTICAST address of the model buffers (an NVSwitch multicast
                                               object every rank's buffer is bound to at the same offset).  With it a
                                               16-byte group of new weights leaves this GPU ONCE (multimem.st) and the
                                               switch writes it into every rank's buffer; models[] still serves the
                                               ragged elements.  NULL: one store per rank.  Only stores use it. */
    const void* grads_mc;                   /* optional: MULTICAST address of the gradient buffers.  With it a 16-byte group
                                               of gradients is summed INSIDE the switch (multimem.ld_reduce, fp32
                                               accumulation, rounded once; the order of the addends is not specified, so
                                               the mean can differ from the rank-ordered one by a few 16-bit ulps; 4 seen at 4 ranks) and only
                                               the sum crosses the link.  NULL: one load per rank, rank order, bit-exact. */
} b200_peer_fan;

int32_t b200_peer_shard_step(const void* dev_table, int64_t n_tiles,
                             int32_t grad_dtype, int32_t state_dtype, int32_t copy_dtype,
                             const b200_adam_group* groups, int32_t n_groups,
                             float inv_scale, const float* dev_inv_scale,
                             int32_t* dev_flag, const int32_t* dev_parity, const b200_step_sync* sync,
                             const b200_peer_fan* fan, const b200_peer_sync* psync, void* stream);
int32_t b200_peer_shard_repair(const void* dev_table, int64_t n_tiles, int32_t state_dtype, int32_t copy_dtype,
                               const int32_t* dev_found_inf, const int32_t* dev_parity,
                               const b200_peer_fan* fan, const b200_peer_sync* psync, void* stream);

/* ------------------------------------------------------------------------------------------
 * S2'' — per-tensor L2 norms and FusedLAMB.
 * Replaces amp_C.multi_tensor_l2norm(..., per_tensor=True) (apex csrc/multi_tensor_l2norm_kernel.cu)
 * and amp_C.multi_tensor_lamb (apex csrc/multi_tensor_lamb.cu: stage 1 = moments + update,
 * stage 2 = trust-ratio apply) as apex/optimizers/fused_lamb.py calls them.  apex itself is not
 * vendored in the reference (it is a pip dependency of test_apex_distributed_spawn.py:14-16);
 * the rule restated in oracle/amp_step_oracle.c (lamb_*) follows the LAMB paper and apex's
 * published kernel; parity for this row is UNPINNED against apex (no apex build here).
 *
 * b200_l2norm_per_tensor: sum of squares of column `col` of every tensor, deterministic (one CTA
 * per tile in fixed order -> one warp per tensor over its tiles -> one warp over the tensors; no
 * floating-point atomics).  dev_first_tile[n_tensors + 1] is the exclusive prefix sum of the
 * tensors' tile counts (ceil(numel / tile_elems)); dev_tile_ss holds n_tiles floats of scratch.
 * Writes dev_sumsq_per_tensor[t], dev_norm_per_tensor[t] = sqrt (may be NULL), ADDS the list's
 * total to *dev_total_sumsq (may be NULL; several dtype lists accumulate) and raises *dev_flag
 * on inf/NaN (may be NULL).
 *
 * b200_lamb_stage1, columns 0 g, 1 p (read only), 2 m, 3 v, 4 u (fp32 out):
 *   g' = g * inv_scale / max(1, ||g * inv_scale||_all / max_grad_norm)      (max_grad_norm <= 0: no clip)
 *   L2 mode: g' += wd * p;   m = b1*m + b3*g';  v = b2*v + (1-b2)*g'^2;  b3 = grad_averaging ? 1-b1 : 1
 *   u = (m/bc1) / (sqrt(v/bc2) + eps);   AdamW mode: u += wd * p
 * b200_lamb_stage2, columns 0 u (fp32), 1 p, 2 model copy (optional):
 *   ratio = lr * ||p|| / ||u|| if (use_nvlamb or wd != 0) and both norms non-zero, else lr;  p -= ratio * u
 * dev_p_norm / dev_u_norm are indexed by the row's index in the list.  Both stages skip when
 * *dev_found_inf != 0. */
typedef struct b200_lamb_group {
    double  lr, beta1, beta2, eps, weight_decay;
    int64_t step;
    int32_t bias_correction;
    int32_t adam_w_mode;
    int32_t grad_averaging;
    int32_t use_nvlamb;
} b200_lamb_group;

int32_t b200_l2norm_per_tensor(const void* dev_table, int64_t n_tiles, int32_t col, int32_t dtype,
                               const uint32_t* dev_first_tile, int32_t n_tensors, float* dev_tile_ss,
                               float* dev_sumsq_per_tensor, float* dev_norm_per_tensor,
                               float* dev_total_sumsq, int32_t* dev_flag, void* stream);
int32_t b200_lamb_stage1(const void* dev_table, int64_t n_tiles, int32_t grad_dtype, int32_t state_dtype,
                         const b200_lamb_group* groups, int32_t n_groups,
                         float inv_scale, const float* dev_inv_scale, const int32_t* dev_found_inf,
                         const float* dev_grad_sumsq, float max_grad_norm, void* stream);
int32_t b200_lamb_stage2(const void* dev_table, int64_t n_tiles, int32_t state_dtype, int32_t copy_dtype,
                         const b200_lamb_group* groups, int32_t n_groups, const int32_t* dev_found_inf,
                         const float* dev_p_norm, const float* dev_u_norm, void* stream);

/* NovoGrad (amp_C.multi_tensor_novograd as apex.optimizers.FusedNovoGrad calls it; apex is not vendored in the
 * reference: parity UNPINNED, rule restated in oracle/amp_step_oracle.c).  The second moment is ONE float per tensor:
 *   b200_novograd_blend:  n = sqrt(grad_sumsq[t]) * inv_scale;  v[t] = sqrt(b2 * v[t]^2 + (1-b2) * n^2), with v[t] starting
 *                         from n on step 1 unless init_zero  (grad_sumsq from b200_l2norm_per_tensor, indexed by row)
 *   b200_fused_novograd, columns 0 g, 1 p, 2 m, 3 model copy (optional):  denom = v[t]/bc2 + eps, bc2 = sqrt(1 - b2^step)
 *     reg_inside_moment:  g' = g/denom + wd*p;  m = b1*m + b3*g';  p -= lr * m/bc1
 *     otherwise (default): m = b1*m + b3*g;  p -= lr * ((m/bc1)/denom + wd*p)          b3 = grad_averaging ? 1-b1 : 1
 * Both skip when *dev_found_inf != 0. */
typedef struct b200_novograd_group {
    double  lr, beta1, beta2, eps, weight_decay;
    int64_t step;
    int32_t bias_correction;
    int32_t reg_inside_moment;
    int32_t grad_averaging;
    int32_t init_zero;
} b200_novograd_group;

int32_t b200_novograd_blend(const void* dev_table, int32_t n_tensors, const b200_novograd_group* groups, int32_t n_groups,
                            float inv_scale, const float* dev_inv_scale, const int32_t* dev_found_inf,
                            float* dev_v, const float* dev_grad_sumsq, void* stream);
int32_t b200_fused_novograd(const void* dev_table, int64_t n_tiles, int32_t grad_dtype, int32_t state_dtype,
                            int32_t copy_dtype, const b200_novograd_group* groups, int32_t n_groups,
                            float inv_scale, const float* dev_inv_scale, const int32_t* dev_found_inf,
                            const float* dev_v, void* stream);

/* ------------------------------------------------------------------------------------------
 * S2' — speculative single-pass step: the overflow check rides on the update pass.
 * (No counterpart in apex, whose unscale pass inspects every gradient before multi_tensor_adam
 * starts — SURVEY.md §7.3 "overflow flag vs all-or-nothing skip".)
 *
 * The all-or-nothing skip forces "inspect every gradient before the first state element is
 * overwritten", i.e. a 2 B/param pass in front of the 28 B/param update (b200_multi_tensor_check,
 * or the peer reduce-scatter, which cannot overlap the update for the same reason).  This variant
 * double-buffers p, m and v instead: every state tensor of the table (columns 1..3) is followed by
 * a SHADOW copy of itself that starts b200_spec_shadow_delta(numel, elem_size) bytes after the
 * primary one (the next 256-byte boundary).  A device word `parity` says which copy is current.
 *
 *   b200_fused_adam_spec : one pass — reads grad + the current p/m/v, raises *dev_flag on a
 *                          non-finite gradient, writes the new p/m/v into the OTHER copy and the
 *                          new model weights into column 4.  Never skipped, never destructive.
 *   b200_spec_commit     : one thread — LossScaler.update_scale (exactly b200_scaler_update), then
 *                          parity ^= 1 unless found_inf; advances the step counters of `sync`.
 *   b200_spec_repair     : after a skipped step the model copy holds weights derived from the
 *                          poisoned gradient: rewrite it from the current masters.  Exits at once
 *                          when *dev_found_inf == 0 (every normal step), so it is always enqueued.
 * The unscale factor must be the one of the backward that produced the gradients; because the
 * scaler update now runs AFTER the pass, pass &state->inv_loss_scale as dev_inv_scale (not
 * &state->inv_scale).  Same arithmetic and results as b200_fused_adam, 28 B/param of traffic, at
 * the price of 12 B/param more resident memory.
 * ------------------------------------------------------------------------------------------ */
int64_t b200_spec_shadow_delta(int64_t numel, int32_t elem_size);

int32_t b200_fused_adam_spec(const void* dev_table, int64_t n_tiles,
                             int32_t grad_dtype, int32_t state_dtype, int32_t copy_dtype,
                             const b200_adam_group* groups, int32_t n_groups,
                             float inv_scale, const float* dev_inv_scale,
                             int32_t* dev_flag, const int32_t* dev_parity,
                             const b200_step_sync* sync, const b200_peer_pull* pull, void* stream);

int32_t b200_spec_commit(b200_scaler_state* dev_state, int32_t mode,
                         int32_t scale_window, float scale_factor, float min_scale, float max_scale,
                         b200_scaler_state* mirror, int32_t* dev_parity,
                         const b200_step_sync* sync, void* stream);

int32_t b200_spec_repair(const void* dev_table, int64_t n_tiles, int32_t state_dtype, int32_t copy_dtype,
                         const int32_t* dev_found_inf, const int32_t* dev_parity, void* stream);

/* ------------------------------------------------------------------------------------------
 * S3'' — gradient mean + speculative step as ONE kernel per rank (no counterpart in apex).
 * b200_peer_reduce_shard followed by b200_fused_adam_peer costs NVLink time PLUS HBM time: the
 * in-place update may not start before every rank has reduced (and inspected) its whole shard.
 * With the speculative update (b200_fused_adam_spec) nothing is destroyed, so both halves run at
 * the same time inside one kernel: the first `n_reducers` CTAs REDUCE this rank's shard — runs of
 * 64 pieces of 2048 elements, every rank's slice of a piece staged in shared memory by cp.async,
 * several pieces deep, arithmetic exactly as b200_peer_reduce_shard — and when the last piece of
 * a `chunk_elems`-element chunk is folded the chunk is announced to every rank: each rank's signal
 * pad is FOLLOWED by a progress board of B200_MAX_PEERS x B200_PEER_STEP_MAX_CHUNKS int32 epoch
 * flags (board[owner][chunk]), i.e. the pad allocation must be B200_PEER_STEP_PAD_BYTES long and
 * zeroed.  All other CTAs UPDATE: they stride over the tile list, wait until the owner of a tile's
 * gradient has announced the chunk(s) it lies in, pull the reduced gradient over NVLink and run the
 * speculative Adam pass.  Every rank inspects every reduced gradient it pulls, so *dev_flag is
 * complete without a flag exchange; follow with b200_spec_commit and b200_spec_repair exactly as
 * after b200_fused_adam_spec.
 *
 * Requirements: shards that never cut a tile (tables.tile_aligned_shards); the table's tile list
 * ordered by b200_table_order_for_peer_step (chunk-major, own shard first) with the same
 * chunk_elems (a multiple of 131072, at most B200_PEER_STEP_MAX_CHUNKS chunks per shard);
 * `dev_scratch` >= (8 + chunks per shard) * 4 bytes, zeroed once; `psync->epoch` growing by one per
 * step on every rank.
 * ------------------------------------------------------------------------------------------ */
#define B200_PEER_STEP_MAX_CHUNKS 1024
#define B200_PEER_STEP_PAD_BYTES  (128 + B200_MAX_PEERS * B200_PEER_STEP_MAX_CHUNKS * 4)
int32_t b200_table_order_for_peer_step(void* host_image, const void* local_base, int32_t grad_elem_size,
                                       const int64_t* bounds, int32_t world, int32_t rank, int64_t chunk_elems);

int32_t b200_peer_spec_step(const void* dev_table, int64_t n_tiles,
                            int32_t grad_dtype, int32_t state_dtype, int32_t copy_dtype,
                            const b200_adam_group* groups, int32_t n_groups,
                            float inv_scale, const float* dev_inv_scale,
                            int32_t* dev_flag, const int32_t* dev_parity, const b200_step_sync* sync,
                            const b200_peer_pull* pull, const void* const* peer_grads, void* reduced,
                            float factor, const b200_peer_sync* psync,
                            int32_t n_reducers, int64_t chunk_elems,
                            void* dev_scratch, int64_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * b200_fused_adam_noalloc — the fused Adam step for callers that must not allocate.
 * (No counterpart in the reference: apex always goes through the torch allocator.)
 *
 * Who this is for: code that has no way to allocate device memory at the point where it needs an
 * optimizer step — a plugin that its host application handed a fixed set of buffers and a rule
 * "no further allocation", a callback that runs in a context where cudaMalloc and the torch
 * caching allocator must not be entered (they may take locks or synchronise the device), a C
 * caller that brings its own p/m/v/grad buffers and has nowhere to put a tensor-list table.
 * b200_fused_adam needs such a table in device memory and leaves allocating it to the caller;
 * this entry point keeps the table in a B200_NOALLOC_TABLE_BYTES static device array that is
 * part of the library image instead (every GPU that loads the library carries those bytes until
 * the process exits).
 *
 * Arguments are b200_table_build's description of the list (`ptrs` column-major over the Adam
 * columns g, p, m, v [, copy]; depth 4 when copy_dtype is B200_NONE, else 5), one
 * hyper-parameter group, and b200_fused_adam's scale / skip arguments.  The caller still owns
 * every input and output buffer; a NULL entry in a column the kernel uses is refused with
 * B200_EINVAL.  Same arithmetic, same results as b200_fused_adam.
 *
 * Limits and rules:
 *   - the list must fit: b200_table_bytes(n_tensors, numel, 4096) <= B200_NOALLOC_TABLE_BYTES
 *     (256 KiB = 64 B + 64 B per tensor + 8 B per 4096-element tile: e.g. the whole 148-tensor,
 *     124M-parameter GPT-2-small list fits).  Larger lists are refused with B200_ESIZE and a
 *     message that says so.
 *   - it calls no cudaMalloc / cudaFree and nothing that allocates; the table image is assembled
 *     in a static (pageable) host buffer and sent with one cudaMemcpyAsync on `stream` at call time.
 *   - the first call on each device looks up the array's device address (a one-time set-up with
 *     that device current).  Make that first call before starting any stream capture.
 *   - all calls share the one array, and each call overwrites the previous table in full
 *     (nothing in it needs to start out zero).  Two calls whose GPU work can overlap — different
 *     streams, or different host threads without ordering between their streams — race on it:
 *     order them yourself.  Host-side the call is serialised by a mutex.
 * ------------------------------------------------------------------------------------------ */
#define B200_NOALLOC_TABLE_BYTES (256 * 1024)
int32_t b200_fused_adam_noalloc(int32_t n_tensors, const void* const* ptrs, const int64_t* numel,
                                int32_t grad_dtype, int32_t state_dtype, int32_t copy_dtype,
                                const b200_adam_group* group,
                                float inv_scale, const float* dev_inv_scale,
                                const int32_t* dev_found_inf, void* stream);

/* Tuning knobs (process-wide; defaults chosen by measurement): CTAs per SM of
 * the persistent grids (0 = per-kernel default), threads per CTA (256 or 512) and how many
 * 8-element groups each thread keeps in flight (1 or 2). */
int32_t b200_set_launch_config(int32_t ctas_per_sm, int32_t threads, int32_t unroll);
/* Number of kernels this library has launched since load (for bench.py's gpu_launches). */
int64_t b200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* APEX_B200_H_ */
