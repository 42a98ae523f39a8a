/*
 * gradavg.h -- C ABI of the B200 gradient-averaging library (libgradavg.so).
 *
 * One hot path, rebuilt from scratch for sm_100a: the post-backward gradient averaging of
 * dblakely/pytorch-distributed-example,
 *
 *     for p in self.net.parameters():                                  mnist/main.py:125
 *         distributed.all_reduce(p.grad.data, op=reduce_op.SUM)        mnist/main.py:126
 *         p.grad.data /= float(world_size)                             mnist/main.py:127
 *
 * and its toy analogue (one int32 all_reduce SUM, toy/main.py:19-20).  Everything below is
 * plain C: pointers, sizes, ints.  No torch types cross this boundary; a stream is passed as
 * the `cudaStream_t` value cast to `void*` (NULL = the legacy default stream).
 *
 * Conventions
 *   - every function returns 0 on success and a negative GA_E* code on failure; the message
 *     for the calling thread's last failure is `ga_last_error()`.  Nothing throws.
 *   - a context hosts `n_local` consecutive ranks of a `world`-rank job on ONE device.  The
 *     normal case is one process per GPU: n_local = 1 (`ga_create`).  `ga_create_local` hosts
 *     all `world` ranks on one device in one process (single launch, blockIdx.y = rank); it is
 *     the same kernels and flag protocol and is how the parity tests run on a one-GPU box.
 *   - every entry point makes the context's device current while it runs and restores the caller's
 *     device before returning.
 *   - a launch carries no host-side counter (the flag epoch lives in device memory), so calls can be
 *     captured in a CUDA graph and replayed, provided the pointer table fits the kernel parameters
 *     (world-hosted ranks * n_tensors <= 128; larger tables are staged through a pinned ring, which
 *     synchronises on reuse and therefore cannot be captured) and nothing has to be allocated
 *     (create the context, and grow its bucket, before capture).
 *   - calls are collective: every rank calls the same function with the same sizes, in the
 *     same order.  Calls on one context must come from one thread at a time.
 *   - launches of one context must be ORDERED with respect to each other on the GPU: the flag epoch,
 *     the staging regions and the bucket belong to the context, so two of its kernels running at
 *     once would share them and give wrong gradients.  Calls on one stream are ordered.  When a call
 *     arrives on a different stream than the previous one, the library makes the new stream wait for
 *     the previous launch (one event record + wait) -- except while either stream is being captured,
 *     where the caller's own graph edges must provide the order.  Use one context per concurrent
 *     stream if launches are meant to overlap.
 *   - sums are taken in rank order, ((g0 + g1) + g2) + ..., each add one IEEE binary32
 *     operation, then ONE IEEE division by (float)world.  Results are therefore identical on
 *     every rank and bit-identical to oracle/fold.c.
 */
#ifndef GRADAVG_H_
#define GRADAVG_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GA_MAX_WORLD        8      /* ranks behind one NVSwitch domain */
#define GA_IPC_HANDLE_BYTES 64     /* sizeof(cudaIpcMemHandle_t) */

/* error codes */
#define GA_OK            0
#define GA_EINVAL       -1   /* bad argument (message says which) */
#define GA_ECUDA        -2   /* a CUDA runtime call failed (message carries cudaGetErrorString) */
#define GA_ECAPACITY    -3   /* gradient set / byte count does not fit the bucket capacity */
#define GA_ESTATE       -4   /* wrong call order (peers not opened, context broken, ...) */
#define GA_ETIMEOUT     -5   /* a kernel gave up waiting for a peer's flag (see ga_status) */

/* variant of the reduction */
#define GA_VARIANT_AUTO     0   /* chosen from world size and byte count by a table read off measured sweeps (ga_auto_variant) */
#define GA_VARIANT_ONESHOT  1   /* every rank reads all `world` buckets: 1 barrier, (world-1)*S ingress */
#define GA_VARIANT_TWOSHOT  2   /* reduce-scatter + all-gather by peer reads: 2 barriers, 2(world-1)/world*S */
#define GA_VARIANT_ONESHOT_PUSH 3 /* every rank STORES its set into every peer's staging slot, local fold   */
#define GA_VARIANT_TWOSHOT_PUSH 4 /* scatter to shard owners by remote stores, owner folds, result stored to
                                     every peer's gather region; pack folds into the scatter (tensor calls) */
#define GA_VARIANT_HOSTED_DIRECT 6 /* ga_create_local only: every rank's tensors are addressable from one launch, so
                                     each tile is read from all ranks, folded and written back: 2*W*S bytes, no bucket */
#define GA_VARIANT_TWOSHOT_PIPE 5 /* two-shot push, software-pipelined in rounds: scatter of round q+1 and
                                     unpack of round q-1 overlap the fold of round q (large tensor calls)   */
#define GA_VARIANT_VIEW     7   /* ga_average_view only: in place on the view region, scatter blocks and fold blocks
                                     running at once with one flag per chunk; bit-exact like every variant above */
#define GA_VARIANT_NVLS     8   /* ga_average_view only, OPT-IN, one rank per GPU with NVSwitch multicast: the switch
                                     adds the replicas (multimem.ld_reduce) and replicates the result (multimem.st).
                                     Half the link bytes.  NOT the rank-order sum: the switch returns the correctly
                                     rounded exact sum of the `world` values, so results can differ from every other
                                     variant in the last bit (oracle: oracle/fold.c fold_exact_*); never picked by AUTO */

/* element type + operation */
#define GA_OP_F32_AVG   0   /* fp32: rank-order sum, then / (float)world      mnist/main.py:126-127 */
#define GA_OP_I32_SUM   1   /* int32: sum, no division                          toy/main.py:19-20   */
#define GA_OP_F32_SUM   2   /* fp32: rank-order sum only (all_reduce alone)    mnist/main.py:126   */

/* options for ga_set_option */
#define GA_OPT_ONESHOT_MAX_BYTES 1   /* AUTO uses one-shot when bytes <= value */
#define GA_OPT_TIMEOUT_MS        2   /* flag-wait timeout inside kernels; 0 = wait forever */
#define GA_OPT_MAX_BLOCKS        3   /* cap on thread blocks per rank (0 = fill the device) */
#define GA_OPT_TRACE             6   /* 1: the push kernels record per-block phase timestamps (globaltimer ns) */
#define GA_OPT_HOST_CHUNK_BYTES  5   /* ga_average_host: chunk size of the copy-in / reduce / copy-out pipeline (0 = default: 1/8 of the set, 1-8 MiB) */
#define GA_OPT_PIPE_GROUP        4   /* TWOSHOT_PIPE: tile slots per block per round (0 = automatic) */
#define GA_OPT_VIEW_CHUNK_TILES  7   /* GA_VARIANT_VIEW: 2 KiB tiles per chunk (0 = automatic) */
#define GA_OPT_VIEW_ONESHOT_MAX_BYTES 8 /* GA_VARIANT_VIEW: sets up to this size take the one-shot form (every rank receives
                                        every tile, folds all of it locally); -1 = automatic, 0 = never */
#define GA_OPT_VIEW_MIN_BYTES    9   /* ga_average_view with GA_VARIANT_AUTO: smaller sets go through the bucket kernels
                                        (region = one tensor), larger ones through the in-place kernel (default 32 MiB) */

/* flags for ga_view_create */
#define GA_VIEW_MULTICAST        1   /* also map the regions through an NVSwitch multicast object when the GPUs
                                        support it (one rank per GPU only); enables GA_VARIANT_NVLS */
#define GA_VIEW_HANDLE_BYTES   128   /* what ga_view_export writes per rank */

typedef struct ga_ctx ga_ctx;

/* Last error message of the calling thread ("" if none). */
const char *ga_last_error(void);

/*
 * Create the per-process context of rank `rank` in a `world`-rank job on CUDA device `device`.
 * Replaces the state `distributed.init_process_group` sets up for this path
 * (mnist/main.py:198-204, toy/main.py:28-33) plus `distributed.get_world_size()`
 * (mnist/main.py:123), which becomes a field of the context.
 * cudaMallocs, owned by the context and freed by ga_destroy:
 *   - the persistent flat bucket of `capacity_bytes` (rounded up to 2 KiB tiles) preceded by
 *     the flag pad (one IPC-exportable allocation),
 *   - a same-size local result region used by the bucket-direct calls.
 */
int ga_create(int rank, int world, int device, size_t capacity_bytes, ga_ctx **out);

/*
 * Same, but all `world` ranks live in this process on this one device; no IPC, peers are
 * opened implicitly.  Tensor-list arguments then carry world * n_tensors pointers,
 * rank-major.  No counterpart call in the reference: its READMEs start the ranks as separate
 * processes on one host (mnist/README.md:18-33); this is that, for one GPU.
 */
int ga_create_local(int world, int device, size_t capacity_bytes, ga_ctx **out);

/* CUDA-IPC handle of this rank's bucket allocation; exchange the 64 bytes once over the
 * process group the script already created (mnist/main.py:199-203). */
int ga_export_handle(ga_ctx *ctx, unsigned char out[GA_IPC_HANDLE_BYTES]);

/* `handles` = world * 64 bytes, rank order (own entry ignored).  Opens every peer's bucket
 * (cudaIpcOpenMemHandle, peer access enabled lazily).  Call once, after every rank exported. */
int ga_open_peers(ga_ctx *ctx, const unsigned char *handles);

/*
 * The hot path: Trainer.average_gradients (mnist/main.py:122-127) for one parameter list.
 * grad_ptrs[i] / numel[i]: device pointer and element count of the i-th contiguous fp32
 * gradient, in parameters() order (n_local * n_tensors pointers for a local world).  Pointers
 * are re-read every call because autograd re-allocates `.grad` each step.
 * Enqueues ONE kernel on `stream`: pack into the bucket -> peer reduce in rank order (+ the
 * division) -> unpack straight into the same gradient tensors.  Asynchronous; the gradients
 * are borrowed until the enqueued work completes.
 */
int ga_average(ga_ctx *ctx, const void *const *grad_ptrs, const int64_t *numel, int n_tensors,
               void *stream, int variant);

/* Generalisation of ga_average over the element op (GA_OP_I32_SUM covers toy/main.py:19-20,
 * where the tensor list is one 1-element int32 tensor). */
int ga_allreduce_tensors(ga_ctx *ctx, const void *const *ptrs, const int64_t *numel, int n_tensors,
                         int op, void *stream, int variant);

/* Same with a per-call cap on thread blocks per rank (0 = none): what the overlap mode uses for a
 * bucket that is averaged while loss.backward() (mnist/main.py:92) still runs, so that the remaining
 * backward kernels keep their SMs -- without touching the context's GA_OPT_MAX_BLOCKS. */
int ga_allreduce_tensors_ex(ga_ctx *ctx, const void *const *ptrs, const int64_t *numel, int n_tensors,
                            int op, void *stream, int variant, int max_blocks);

/*
 * Replicates rank `root`'s tensors into the same-shaped tensors of every other rank (32-bit
 * elements, any type).  The reference builds Net() on every rank and relies on identical initial
 * parameters that it never arranges (no seed, no broadcast: mnist/main.py:185); this is the missing
 * step, over the peer mappings and flag protocol already in place: the root stores its tensors
 * straight into every peer's gather region, the peers unpack.  One kernel on `stream`; collective.
 */
int ga_broadcast_tensors(ga_ctx *ctx, const void *const *ptrs, const int64_t *numel, int n_tensors,
                         int root, void *stream);

/*
 * Bucket-direct form, for the size sweep and the same-buffer comparison with NCCL: the caller
 * has written `nbytes` into ga_send_ptr() (stream-ordered before this call); the averaged
 * result is left in ga_recv_ptr().  Returns (stream-wise) only after every peer has finished
 * reading this rank's bucket, so the caller may refill it at once.
 */
int ga_allreduce_bucket(ga_ctx *ctx, size_t nbytes, int op, void *stream, int variant);

/*
 * Host-buffer form of ga_average, the call a CPU-side user of the reference makes (its
 * default gloo path keeps gradients in host memory, mnist/main.py:181-182 with --no-cuda):
 * copies each host tensor into its bucket slot (H2D), runs the bucket-direct reduction,
 * copies the results back over the same host tensors (D2H), all on `stream`.  Host memory
 * should be pinned for the copies to be asynchronous.  n_local must be 1 unless the context is
 * a local world, in which case host_ptrs is rank-major like grad_ptrs.
 */
int ga_average_host(ga_ctx *ctx, void *const *host_ptrs, const int64_t *numel, int n_tensors,
                    void *stream, int variant);

/*
 * Gradients as views.  The steps either side of the hot path only touch p.grad:
 * zero_grad() + backward() write it (mnist/main.py:91-92), optimizer.step() reads it (:96).  If every
 * p.grad IS a slice of one flat region that the peers have mapped, averaging (:122-127) needs neither
 * pack nor unpack: ga_average_view works in place on that region.
 *
 *   ga_view_create   allocates `nbytes` (<= the context's capacity: the bucket serves as staging) per
 *                    hosted rank.  With GA_VIEW_MULTICAST, on a one-rank-per-GPU world whose GPUs report
 *                    CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, the region is a cuMem allocation that is
 *                    also bound to an NVSwitch multicast object; otherwise a plain cudaMalloc shared
 *                    through CUDA-IPC like the bucket.  Collective: same arguments on every rank.
 *   ga_view_export   GA_VIEW_HANDLE_BYTES to all-gather over the script's own process group
 *                    (mnist/main.py:199-203), like ga_export_handle.
 *   ga_view_open     maps every peer's region (and joins the multicast object).  `handles` = world *
 *                    GA_VIEW_HANDLE_BYTES in rank order.  For the cuMem kind the descriptors are fetched
 *                    from the exporting processes with pidfd_getfd (same user, Linux >= 5.6), so every
 *                    exporter must still be alive and must not have called ga_view_bind yet.
 *   -- barrier --
 *   ga_view_bind     binds this rank's memory to the multicast object and maps it (a no-op for the
 *                    CUDA-IPC kind).  Barrier again before the first ga_average_view.
 *   ga_view_ptr      device pointer of hosted rank `local_rank`'s region: lay the gradients out in it
 *                    with ga_bucket_offset (2 KiB-aligned tensors) or any layout every rank shares --
 *                    the kernels treat it as one flat array.
 * A local world (ga_create_local) needs only ga_view_create.  The regions are unmapped by
 * ga_close_peers and freed by ga_destroy.
 */
int ga_view_create(ga_ctx *ctx, size_t nbytes, int flags);
int ga_view_export(ga_ctx *ctx, unsigned char out[GA_VIEW_HANDLE_BYTES]);
int ga_view_open(ga_ctx *ctx, const unsigned char *handles);
int ga_view_bind(ga_ctx *ctx);
void *ga_view_ptr(ga_ctx *ctx, int local_rank);
/* Unmaps and frees the view region again (collective: barrier first, no peer may still be using it), e.g. to
 * create it anew without GA_VIEW_MULTICAST after ga_view_open failed to fetch the peers' descriptors. */
int ga_view_destroy(ga_ctx *ctx);
size_t ga_view_bytes(const ga_ctx *ctx);
int ga_view_has_multicast(const ga_ctx *ctx);

/* 1 if `device` supports NVSwitch multicast objects with POSIX descriptors (what GA_VIEW_MULTICAST needs). */
int ga_multicast_supported(int device);

/*
 * Trainer.average_gradients (mnist/main.py:122-127) IN PLACE on the first `nbytes` of every rank's
 * view region: afterwards every rank's region holds the rank-order sum divided by (float)world
 * (GA_OP_F32_AVG; the other ops as in ga_allreduce_tensors).  One kernel on `stream`, asynchronous.
 * variant: GA_VARIANT_VIEW (the in-place kernel), GA_VARIANT_AUTO (that kernel from GA_OPT_VIEW_MIN_BYTES on,
 * below it the bucket kernels with the region as their one tensor: same bits, quicker for small sets)
 * or GA_VARIANT_NVLS (opt-in, see above).
 * This is also the same-buffer form to time against NCCL all_reduce, which is in place too.
 */
int ga_average_view(ga_ctx *ctx, size_t nbytes, int op, void *stream, int variant);

/*
 * Call on `stream` before writing into ga_send_ptr() if the previous call on this context was a
 * tensor-list call (ga_average / ga_allreduce_tensors): those let their peers finish reading this
 * rank's bucket lazily, and this enqueues the wait for them.  A no-op after bucket-direct calls.
 */
int ga_bucket_acquire(ga_ctx *ctx, void *stream);

/* Device pointers of local rank `local_rank`'s bucket (what peers read) and result region. */
void *ga_send_ptr(ga_ctx *ctx, int local_rank);
void *ga_recv_ptr(ga_ctx *ctx, int local_rank);

/* Byte offset of tensor `i` inside the bucket for the given element counts (segments are
 * padded to 2 KiB tiles); with i == n_tensors returns the total bytes used. */
int64_t ga_bucket_offset(const int64_t *numel, int n_tensors, int i);

/* world size / first hosted rank / number of hosted ranks / capacity in bytes */
int ga_world(const ga_ctx *ctx);
int ga_rank(const ga_ctx *ctx);
int ga_n_local(const ga_ctx *ctx);
size_t ga_capacity(const ga_ctx *ctx);

/* What AUTO resolves to for a tensor-list call of `nbytes` on this context (one of GA_VARIANT_ONESHOT ...
 * GA_VARIANT_HOSTED_DIRECT; never VIEW or NVLS, which belong to ga_average_view). */
int ga_pick_variant(const ga_ctx *ctx, size_t nbytes);

/* The same rule without a context (pure host arithmetic, no GPU needed): what AUTO would pick for a
 * job of `world` ranks of which `n_local` are hosted per process, with `capacity_bytes` buckets. */
int ga_auto_variant(int world, int n_local, size_t capacity_bytes, size_t nbytes);

/* Kernels launched by this context so far (for bench.py's gpu_launches). */
uint64_t ga_launch_count(const ga_ctx *ctx);

/* Grid the last launch used: blocks per rank and threads per block. */
int ga_last_grid(const ga_ctx *ctx, int *blocks, int *threads);

int ga_set_option(ga_ctx *ctx, int option, int64_t value);

/*
 * Phase timestamps of the LAST launch (needs GA_OPT_TRACE): blocks * hosted ranks * slots words of
 * %globaltimer nanoseconds, slot k of block b at out[(rank * blocks + b) * slots + k].  Slots of
 * ga_push_kernel: 0 start, 1 previous call's readers gone, 2 scatter issued, 3 flags raised (stores
 * acknowledged), 4 peers' flags seen, 5 fold + result stores issued, 6 flags raised, 7 peers' flags
 * seen, 8 unpack done, 9 end.  ncu cannot profile a multi-rank run; this is the multi-GPU substitute.
 */
int ga_read_trace(ga_ctx *ctx, uint64_t *out, size_t max_words, int *blocks, int *slots);

/* 0, or GA_ETIMEOUT once a kernel of this context gave up waiting for a peer (the context is
 * then unusable).  Reads a host-mapped word; call after synchronising the stream. */
int ga_status(ga_ctx *ctx);

/*
 * Gradient average for callers that must not allocate.  No counterpart in the reference.
 *
 * Who needs it: code that runs where a cudaMalloc is off limits -- inside a host application
 * whose own pool has claimed the GPU's memory and treats foreign allocations as a bug, inside a
 * latency-critical region that cannot afford the device-wide synchronisation cudaMalloc/cudaFree
 * imply, or under a plug-in contract that forbids allocation after start-up.  Every other entry
 * point needs a context, and ga_create / ga_create_local allocate (bucket, flag pad, pinned status
 * word).  This one needs no context and calls no cudaMalloc, cudaFree, cudaHostAlloc or any
 * library routine that allocates: bucket, flags and status live in a fixed array inside the
 * library image.
 *
 * What it does: exactly ga_average on a ga_create_local(world) context -- all `world` ranks hosted
 * on `device`, grad_ptrs = world * n_tensors device pointers (rank-major), fp32, averaged in place,
 * rank-order sum then one division, one kernel on `stream` -- and its results are bit-identical
 * to that call.
 *
 * Limits: world * (bucket bytes of the set, each tensor padded to 2 KiB) <= 8 MiB, and
 * world * n_tensors <= 128 pointers; beyond either it returns GA_ECAPACITY and launches nothing.
 * The array is 8 945 664 bytes (8.53 MiB: 8 MiB of bucket + 8 flag pads) and stays resident on
 * every GPU that has used the library, until the process exits.
 *
 * First call on a device: resolves the array's address on that device (cudaGetSymbolAddress, with
 * the device made current) and caches it.  With CUDA's default lazy module loading that is also the
 * moment the driver loads the library's device code and reserves the array on that GPU -- the one
 * device-memory reservation this entry point ever causes.  So make that first call (or
 * ga_static_status) at start-up: before any stream capture, before the no-allocation region begins
 * and before the host application's pool claims the rest of the memory; or run with
 * CUDA_MODULE_LOADING=EAGER, which moves the reservation to context creation.
 *
 * All calls on one device share the one array.  It keeps whatever the previous call left, so each
 * call first clears, on `stream`, the words its flag protocol counts up from zero (three small
 * strided memsets).  Two calls whose GPU work can overlap (different streams without an event
 * between them, or two host threads) race on that array and give wrong gradients; order them
 * yourself.
 *
 * Errors inside the kernel: a flag wait that expires (GA_TIMEOUT_MS in the environment, default 30 s)
 * makes the blocks leave early with the gradients partly averaged.  The call itself has long returned
 * GA_OK by then; ga_static_status reports it.
 */
int ga_average_static(int world, int device, const void *const *grad_ptrs, const int64_t *numel,
                      int n_tensors, void *stream, int variant);

/* Waits for `stream`, then reads the status word the ga_average_static kernels leave in the array
 * (a 4-byte copy into a stack variable: no allocation here either).  GA_OK, or GA_ETIMEOUT if a kernel
 * of the last call gave up on a flag.  No counterpart in the reference. */
int ga_static_status(int device, void *stream);

/* Waits for everything enqueued on `stream` (for hosts without a CUDA runtime of their own, e.g. a
 * plain C program using ga_average_host), then returns ga_status(). */
int ga_synchronize(ga_ctx *ctx, void *stream);

/*
 * Tear-down is two collective steps, because an exporter must not free memory that a peer still
 * has mapped: every rank calls ga_close_peers (unmaps the peers' buckets; the context can no longer
 * launch), barrier, then every rank calls ga_destroy.  Replaces destroy_process_group for this path.
 */
int ga_close_peers(ga_ctx *ctx);

/* Frees everything.  Every rank must have closed its peers first (see ga_close_peers). */
int ga_destroy(ga_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* GRADAVG_H_ */
