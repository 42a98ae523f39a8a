// gradavg.cu -- sm_100a kernels + C ABI of libgradavg.so (declared in include/gradavg.h).
//
// The one hot path of dblakely/pytorch-distributed-example, /root/reference/mnist/main.py:122-127
// (all_reduce(SUM) of every p.grad, then /= float(world_size)) and toy/main.py:19-20, as ONE
// fused kernel per call.  Two families:
//
//   pull  ga_fused_kernel        pack the gradients into this rank's flat bucket (16-byte vectors,
//                                every tensor padded to a 2 KiB tile so a warp tile never straddles
//                                two tensors) -> per-block flags -> LOAD the peers' buckets over
//                                NVLink (plain ld.global on CUDA-IPC mapped pointers), add in rank
//                                order, divide by (float)world, write straight into the gradients.
//                                One-shot: every rank reads everything; two-shot: rank r reduces
//                                shard r in place, then every rank pulls the other shards.
//   push  ga_push1_kernel        one-shot: STORE every gradient tile into every rank's staging slot
//         ga_push_kernel         two-shot: scatter tiles to the shard owners, owner folds and stores
//         ga_push_pipe_kernel    the result into every peer's gather region; the pipelined form
//                                overlaps scatter / fold / unpack of consecutive rounds.
//         (stores are posted and pack folds into the transfer; AUTO uses these across GPUs)
//   local ga_hosted_direct_kernel  all ranks hosted in one process: no bucket, no flags.
//
// Cross-GPU ordering is by per-block flag rows written into the peers' memory (system-scope
// fence + relaxed stores, one polling thread + one fence on the reader's side), with the flag epoch
// kept in device memory so that a launch can be replayed from a CUDA graph.  Nothing here is a
// contraction, so no tensor cores; the kernels are HBM-bound on one GPU and NVLink-bound across
// GPUs.  Build: see __graft_entry__.build().
#include "gradavg_internal.cuh"

namespace {

// ------------------------------------------------------------------------------------------
// tile movers (one warp moves one 2 KiB tile; lane l owns vectors l, l+32, l+64, l+96)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int find_seg(const Seg *segs, int n, uint32_t tile) {
    int lo = 0, hi = n - 1;          // largest s with segs[s].start_tile <= tile
    while (lo < hi) {
        int m = (lo + hi + 1) >> 1;
        if (segs[m].start_tile <= tile) lo = m; else hi = m - 1;
    }
    return lo;
}

// gradient tensor -> registers (zero fill past the end of the tensor)
__device__ __forceinline__ void load_grad_tile(const Seg &sg, uint32_t tile, int lane, Vec (&v)[kLaneVecs]) {
    const uint32_t *src = (const uint32_t *)sg.ptr;
    const size_t n = sg.numel;
    const size_t base = (size_t)(tile - sg.start_tile) * kTileElems;
    const bool aligned = (((uintptr_t)src) & 15) == 0;
#pragma unroll
    for (int u = 0; u < kLaneVecs; ++u) {
        const size_t e = base + (size_t)(u * 32 + lane) * 4;
        if (aligned && e + 4 <= n) {
            v[u] = __ldcs((const Vec *)(src + e));
        } else {
            v[u].x = e + 0 < n ? __ldcs(src + e + 0) : 0u;
            v[u].y = e + 1 < n ? __ldcs(src + e + 1) : 0u;
            v[u].z = e + 2 < n ? __ldcs(src + e + 2) : 0u;
            v[u].w = e + 3 < n ? __ldcs(src + e + 3) : 0u;
        }
    }
}

// registers -> gradient tensor (elements past the end are dropped)
template <int NV>
__device__ __forceinline__ void store_grad_vecs(const Seg &sg, uint32_t tile, int lane, int u0, const Vec (&v)[NV]) {
    uint32_t *dst = (uint32_t *)sg.ptr;
    const size_t n = sg.numel;
    const size_t base = (size_t)(tile - sg.start_tile) * kTileElems;
    const bool aligned = (((uintptr_t)dst) & 15) == 0;
#pragma unroll
    for (int u = 0; u < NV; ++u) {
        const size_t e = base + (size_t)((u0 + u) * 32 + lane) * 4;
        // __stcg = st.global.cg: the pointer comes out of the kernel parameters as a generic one; naming the
        // state space keeps these STG instead of generic ST (and the gradients' next reader is another kernel: L2)
        if (aligned && e + 4 <= n) {
            __stcg((Vec *)(dst + e), v[u]);
        } else {
            if (e + 0 < n) __stcg(dst + e + 0, v[u].x);
            if (e + 1 < n) __stcg(dst + e + 1, v[u].y);
            if (e + 2 < n) __stcg(dst + e + 2, v[u].z);
            if (e + 3 < n) __stcg(dst + e + 3, v[u].w);
        }
    }
}

// Where a finished tile goes: the gradient tensors (fused unpack) or the result region.
struct Sink {
    const Seg *segs;   // this rank's pointer table (nullptr in bucket-direct mode)
    int n_seg;
    Vec *recv;         // bucket-direct destination
};

template <int NV>
__device__ __forceinline__ void sink_store(const Sink &s, int si, uint32_t tile, int lane, int u0, const Vec (&v)[NV]) {
    if (s.segs) {      // a tensor call: entry `si` of the pointer table is the tile's tensor (an index, not a Seg or a
                       // pointer-or-null to one: those cost four registers or a local-memory copy -- ptxas -v)
        store_grad_vecs<NV>(s.segs[si], tile, lane, u0, v);
    } else {
        Vec *d = s.recv + (size_t)tile * kTileVecs + u0 * 32 + lane;
#pragma unroll
        for (int u = 0; u < NV; ++u) __stcg(d + u * 32, v[u]);
    }
}

// Vectors per lane fetched from EVERY rank before the adds start: W * kBatch loads in flight
// per thread, capped so that the kernel stays within 64 registers (2 blocks of 512 per SM).
template <int W> struct Batch { static constexpr int value = W <= 2 ? 4 : (W <= 4 ? 2 : 1); };

// Reduce one tile from all W buckets in rank order, divide, store to `own` (two-shot in-place
// shard, may be nullptr) and to the sink.
template <int W, int OP>
__device__ __forceinline__ void reduce_tile(const Params &p, uint32_t tile, int lane, float wf,
                                            Vec *own, const Sink &sink, int si) {
    constexpr int NB = Batch<W>::value;
    const size_t off = (size_t)tile * kTileVecs + lane;
#pragma unroll(W > 4 ? 1 : kLaneVecs)
    for (int u0 = 0; u0 < kLaneVecs; u0 += NB) {
        Vec v[W][NB];
#pragma unroll
        for (int r = 0; r < W; ++r)
#pragma unroll
            for (int u = 0; u < NB; ++u)
                v[r][u] = __ldcg((const Vec *)p.send[r] + off + (u0 + u) * 32);
        Vec acc[NB];
#pragma unroll
        for (int u = 0; u < NB; ++u) {
            acc[u] = v[0][u];
#pragma unroll
            for (int r = 1; r < W; ++r) {          // ((g0 + g1) + g2) + ...
                acc[u].x = ElemOp<OP>::add(acc[u].x, v[r][u].x);
                acc[u].y = ElemOp<OP>::add(acc[u].y, v[r][u].y);
                acc[u].z = ElemOp<OP>::add(acc[u].z, v[r][u].z);
                acc[u].w = ElemOp<OP>::add(acc[u].w, v[r][u].w);
            }
            acc[u].x = ElemOp<OP>::fin(acc[u].x, wf);
            acc[u].y = ElemOp<OP>::fin(acc[u].y, wf);
            acc[u].z = ElemOp<OP>::fin(acc[u].z, wf);
            acc[u].w = ElemOp<OP>::fin(acc[u].w, wf);
        }
        if (own) {
#pragma unroll
            for (int u = 0; u < NB; ++u) __stcg(own + off + (u0 + u) * 32, acc[u]);
        }
        sink_store<NB>(sink, si, tile, lane, u0, acc);
    }
}

// ------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------
// grid = (blocks per rank, hosted ranks).  Block b of rank r only ever waits for block b of
// the other ranks, and only touches bucket tiles of its own "column": tile slot
// i = (k * gridDim.x + b) * warps + warp, for k = 0, 1, ...  In the one-shot variant slot i is
// bucket tile i; in the two-shot variant it is tile i of EVERY shard (j * shard_tiles + i).
// That makes the producer of every byte a block of the same index b, so a row of per-block
// flags is the only synchronisation.
template <int W, int VARIANT, int OP>
__global__ void __launch_bounds__(kThreads, 2) ga_fused_kernel(const __grid_constant__ Params p) {
    __shared__ int s_abort;
    __shared__ int s_last;
    const int ly = blockIdx.y;
    const int r = p.rank_base + ly;
    const int b = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int wpb = blockDim.x >> 5;
    const uint32_t stride = gridDim.x * wpb;
    const uint32_t first = b * wpb + warp;
    const uint32_t T = p.total_tiles;
    const float wf = (float)W;

    Pad *mypad = p.pad[r];
    const uint32_t epoch = ld_relaxed_sys(&mypad->epoch) + 1;     // same value in every block of this launch
    Vec *mysend = (Vec *)p.send[r];
    Sink sink;
    sink.n_seg = (int)p.n_seg;
    sink.segs = p.direct ? nullptr : (p.dev_segs ? p.dev_segs : p.inline_segs) + (size_t)ly * p.n_seg;
    sink.recv = (Vec *)p.recv[ly];

    if (threadIdx.x == 0) { s_abort = 0; s_last = 0; }
    __syncthreads();

    // (0) The previous call's readers must be done with my bucket before pack overwrites it.
    //     (Bucket-direct calls end with this wait instead, see the bottom.)
    if (!p.direct) {
        if (!wait_flags<W>(mypad->done, epoch - 1, p, &s_abort)) return;
    }

    // (1) pack: gradient tensors -> my bucket
    if (!p.direct) {
        if (VARIANT == GA_VARIANT_ONESHOT) {
            for (uint32_t t = first; t < T; t += stride) {
                const Seg sg = sink.segs[find_seg(sink.segs, sink.n_seg, t)];
                Vec v[kLaneVecs];
                load_grad_tile(sg, t, lane, v);
#pragma unroll
                for (int u = 0; u < kLaneVecs; ++u) __stcg(mysend + (size_t)t * kTileVecs + u * 32 + lane, v[u]);
            }
        } else {
            for (uint32_t i = first; i < p.shard_tiles; i += stride) {
                for (int j = 0; j < W; ++j) {
                    const uint32_t t = j * p.shard_tiles + i;
                    if (t >= T) break;
                    const Seg sg = sink.segs[find_seg(sink.segs, sink.n_seg, t)];
                    Vec v[kLaneVecs];
                    load_grad_tile(sg, t, lane, v);
#pragma unroll
                    for (int u = 0; u < kLaneVecs; ++u) __stcg(mysend + (size_t)t * kTileVecs + u * 32 + lane, v[u]);
                }
            }
        }
    }

    // (2) barrier 1: my column is packed / wait for the same column of every rank
    const size_t ready_off = offsetof(Pad, ready) / 4 + (size_t)b * kMaxWorld + r;
    raise_flags<W>(p, ready_off, epoch);
    if (!wait_flags<W>(mypad->ready[b], epoch, p, &s_abort)) return;

    if (VARIANT == GA_VARIANT_ONESHOT) {
        // (3) every rank reduces every tile and writes its own copy of the result
        for (uint32_t t = first; t < T; t += stride) {
            const int si = sink.segs ? find_seg(sink.segs, sink.n_seg, t) : 0;
            reduce_tile<W, OP>(p, t, lane, wf, nullptr, sink, si);
        }
    } else {
        // (3) reduce-scatter: rank r reduces shard r in place in its bucket (+ its own copy out)
        for (uint32_t i = first; i < p.shard_tiles; i += stride) {
            const uint32_t t = r * p.shard_tiles + i;
            if (t >= T) break;
            const int si = sink.segs ? find_seg(sink.segs, sink.n_seg, t) : 0;
            reduce_tile<W, OP>(p, t, lane, wf, mysend, sink, si);
        }
        // (4) barrier 2: shard column reduced everywhere
        const size_t mid_off = offsetof(Pad, mid) / 4 + (size_t)b * kMaxWorld + r;
        raise_flags<W>(p, mid_off, epoch);
        if (!wait_flags<W>(mypad->mid[b], epoch, p, &s_abort)) return;
        // (5) all-gather: pull the other ranks' reduced shards, starting with the next rank so
        //     that at any moment the W ranks read from W different sources
        for (uint32_t i = first; i < p.shard_tiles; i += stride) {
#pragma unroll 1
            for (int jj = 1; jj < W; ++jj) {
                const int j = (r + jj) % W;
                const uint32_t t = j * p.shard_tiles + i;
                if (t >= T) continue;
                const int si = sink.segs ? find_seg(sink.segs, sink.n_seg, t) : 0;
                const Vec *src = (const Vec *)p.send[j] + (size_t)t * kTileVecs + lane;
                Vec v[kLaneVecs];
#pragma unroll
                for (int u = 0; u < kLaneVecs; ++u) v[u] = __ldcg(src + u * 32);
                sink_store<kLaneVecs>(sink, si, t, lane, 0, v);
            }
        }
    }

    // (6) tell every rank that this rank no longer reads their buckets (last block only)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&mypad->finished_blocks, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        if (threadIdx.x == 0) { mypad->finished_blocks = 0; mypad->epoch = epoch; mypad->last_kind = 0; }
        const size_t done_off = offsetof(Pad, done) / 4 + r;
        if (threadIdx.x == 0) {
            asm volatile("fence.acq_rel.sys;" ::: "memory");
#pragma unroll
            for (int j = 0; j < W; ++j) st_relaxed_sys((uint32_t *)p.pad[j] + done_off, epoch);
        }
        if (p.direct) wait_flags<W>(mypad->done, epoch, p, &s_abort);
    }
}

// ------------------------------------------------------------------------------------------
// push variants: the same averaging with remote STORES instead of remote loads
// ------------------------------------------------------------------------------------------
// NVLink writes are posted (no request/response round trip), and pack folds into the transfer:
// a gradient tile is read once from the tensor and stored straight into the owner's staging slot.
//   two-shot push: (1) scatter: tile i of shard j -> rank j's staging slot [me]      (remote stores)
//                  (2) flags    (3) rank j folds its W slots in rank order, divides, writes its own
//                  gradients and stores the result into every peer's gather region   (remote stores)
//                  (4) flags    (5) unpack the other shards from my gather region    (local)
//   one-shot push: (1) every tile -> every rank's staging slot [me]  (2) flags  (3) local fold
// The staging area is the bucket (`send`), slot s at s * slot_tiles; the gather region sits behind
// it in the same IPC allocation.  Tensor-list calls only.
template <int W, int OP, typename SrcFn, typename OutFn>
__device__ __forceinline__ void fold_tile(SrcFn src, float wf, int lane, OutFn out) {
    constexpr int NB = Batch<W>::value;
    // worlds above 4 keep W vectors per row in flight: unrolling the four rows on top of that spilled (ptxas -v)
#pragma unroll(W > 4 ? 1 : kLaneVecs)
    for (int u0 = 0; u0 < kLaneVecs; u0 += NB) {
        Vec v[W][NB];
#pragma unroll
        for (int r = 0; r < W; ++r) {
            const Vec *s = src(r) + lane;
#pragma unroll
            for (int u = 0; u < NB; ++u) v[r][u] = __ldcg(s + (u0 + u) * 32);
        }
        Vec acc[NB];
#pragma unroll
        for (int u = 0; u < NB; ++u) {
            acc[u] = v[0][u];
#pragma unroll
            for (int r = 1; r < W; ++r) {          // ((g0 + g1) + g2) + ...
                acc[u].x = ElemOp<OP>::add(acc[u].x, v[r][u].x);
                acc[u].y = ElemOp<OP>::add(acc[u].y, v[r][u].y);
                acc[u].z = ElemOp<OP>::add(acc[u].z, v[r][u].z);
                acc[u].w = ElemOp<OP>::add(acc[u].w, v[r][u].w);
            }
            acc[u].x = ElemOp<OP>::fin(acc[u].x, wf);
            acc[u].y = ElemOp<OP>::fin(acc[u].y, wf);
            acc[u].z = ElemOp<OP>::fin(acc[u].z, wf);
            acc[u].w = ElemOp<OP>::fin(acc[u].w, wf);
        }
        out(u0, acc);
    }
}

template <int W, int OP>
__global__ void __launch_bounds__(kThreads, 2) ga_push_kernel(const __grid_constant__ Params p) {
    __shared__ int s_abort;
    __shared__ int s_last;
    constexpr int NB = Batch<W>::value;
    const int ly = blockIdx.y;
    const int r = p.rank_base + ly;
    const int b = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int wpb = blockDim.x >> 5;
    const uint32_t stride = gridDim.x * wpb;
    const uint32_t first = b * wpb + warp;
    const uint32_t T = p.total_tiles;
    const uint32_t Ts = p.shard_tiles;            // tiles per staging slot
    const float wf = (float)W;

    Pad *mypad = p.pad[r];
    const uint32_t epoch = ld_relaxed_sys(&mypad->epoch) + 1;     // same value in every block of this launch
    Vec *mystage = (Vec *)p.send[r];
    Vec *mygather = (Vec *)p.gather[r];
    Sink sink;
    sink.n_seg = (int)p.n_seg;
    sink.segs = (p.dev_segs ? p.dev_segs : p.inline_segs) + (size_t)ly * p.n_seg;
    sink.recv = nullptr;

    if (threadIdx.x == 0) { s_abort = 0; s_last = 0; }
    __syncthreads();
    GA_TRACE(p, 0);

    // (0) every rank must have consumed the staging data of the previous call before anyone
    //     stores into it again (needed after a one-shot call; free when the ranks are in step)
    if (!wait_flags<W>(mypad->done, epoch - 1, p, &s_abort)) return;
    GA_TRACE(p, 1);

    // (1) scatter my gradient tiles into the owners' staging slots, nearest rank first
    {
        for (uint32_t i = first; i < Ts; i += stride) {
#pragma unroll 1
            for (int jj = 0; jj < W; ++jj) {
                const int j = (r + jj) % W;
                const uint32_t t = j * Ts + i;
                if (t >= T) continue;
                const Seg sg = sink.segs[find_seg(sink.segs, sink.n_seg, t)];
                Vec v[kLaneVecs];
                load_grad_tile(sg, t, lane, v);
                Vec *dst = (Vec *)p.send[j] + ((size_t)r * Ts + i) * kTileVecs + lane;
#pragma unroll
                for (int u = 0; u < kLaneVecs; ++u) __stcg(dst + u * 32, v[u]);
            }
        }
    }

    // (2) my column has landed everywhere / wait for everyone's column
    GA_TRACE(p, 2);
    const size_t ready_off = offsetof(Pad, ready) / 4 + (size_t)b * kMaxWorld + r;
    raise_flags<W>(p, ready_off, epoch);
    GA_TRACE(p, 3);
    if (!wait_flags<W>(mypad->ready[b], epoch, p, &s_abort)) return;
    GA_TRACE(p, 4);

    {
        // (3) fold my shard from the W local slots; result -> my gradients and every peer's gather region
        for (uint32_t i = first; i < Ts; i += stride) {
            const uint32_t t = r * Ts + i;
            if (t >= T) break;
            const Seg sg = sink.segs[find_seg(sink.segs, sink.n_seg, t)];
            fold_tile<W, OP>([&](int s) { return mystage + ((size_t)s * Ts + i) * kTileVecs; }, wf, lane,
                             [&](int u0, const Vec(&acc)[NB]) {
                                 store_grad_vecs<NB>(sg, t, lane, u0, acc);
                                 auto to_peer = [&](int jj) {
                                     const int j = (r + jj) % W;
                                     Vec *dst = (Vec *)p.gather[j] + (size_t)t * kTileVecs + u0 * 32 + lane;
#pragma unroll
                                     for (int u = 0; u < NB; ++u) __stcg(dst + u * 32, acc[u]);
                                 };
                                 if constexpr (W > 4) {          // seven hoisted peer addresses spilled (ptxas -v)
#pragma unroll 1
                                     for (int jj = 1; jj < W; ++jj) to_peer(jj);
                                 } else {
#pragma unroll
                                     for (int jj = 1; jj < W; ++jj) to_peer(jj);
                                 }
                             });
        }
        // (4) my reduced column has landed everywhere / wait for everyone's
        GA_TRACE(p, 5);
        const size_t mid_off = offsetof(Pad, mid) / 4 + (size_t)b * kMaxWorld + r;
        raise_flags<W>(p, mid_off, epoch);
        GA_TRACE(p, 6);
        if (!wait_flags<W>(mypad->mid[b], epoch, p, &s_abort)) return;
        GA_TRACE(p, 7);
        // (5) unpack the other ranks' shards from my gather region (local traffic only)
        for (uint32_t i = first; i < Ts; i += stride) {
#pragma unroll 1
            for (int jj = 1; jj < W; ++jj) {
                const int j = (r + jj) % W;
                const uint32_t t = j * Ts + i;
                if (t >= T) continue;
                const Seg sg = sink.segs[find_seg(sink.segs, sink.n_seg, t)];
                const Vec *src = mygather + (size_t)t * kTileVecs + lane;
                Vec v[kLaneVecs];
#pragma unroll
                for (int u = 0; u < kLaneVecs; ++u) v[u] = __ldcg(src + u * 32);
                store_grad_vecs<kLaneVecs>(sg, t, lane, 0, v);
            }
        }
    }

    // (6) this rank is done with its staging and gather data
    __syncthreads();
    GA_TRACE(p, 8);
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&mypad->finished_blocks, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        if (threadIdx.x == 0) { mypad->finished_blocks = 0; mypad->epoch = epoch; mypad->last_kind = 0; }
        const size_t done_off = offsetof(Pad, done) / 4 + r;
        if (threadIdx.x == 0) {
            asm volatile("fence.acq_rel.sys;" ::: "memory");
#pragma unroll
            for (int j = 0; j < W; ++j) st_relaxed_sys((uint32_t *)p.pad[j] + done_off, epoch);
        }
    }
    GA_TRACE(p, 9);
}

// One-shot push, latency form.  A block owns `group` consecutive tiles; the scatter is spread over
// the warps as (tile, destination) items and the fold as (tile, 512-byte row) items, so a small set
// still engages many SMs.  With room for two staging halves (parity_stride != 0) consecutive
// launches alternate halves by epoch parity: a peer can then never be overwritten while it is still
// folding the previous call, so such a launch skips the wait for the peers' `done` words whenever
// this rank's previous launch was of the same kind (recorded in the pad, so it also holds under
// graph replay).  Any other predecessor (a pull kernel's readers, a two-shot's staging) is waited for.
template <int W, int OP>
__global__ void __launch_bounds__(kThreads, 2) ga_push1_kernel(const __grid_constant__ Params p) {
    __shared__ int s_abort;
    __shared__ int s_last;
    const int ly = blockIdx.y;
    const int r = p.rank_base + ly;
    const int b = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t wpb = blockDim.x >> 5;
    const uint32_t T = p.total_tiles, G = p.group;
    const float wf = (float)W;
    const uint32_t t0 = b * G;
    const uint32_t nt = t0 >= T ? 0u : (T - t0 < G ? T - t0 : G);

    Pad *mypad = p.pad[r];
    const uint32_t epoch = ld_relaxed_sys(&mypad->epoch) + 1;
    const bool dbuf = p.parity_stride != 0;
    const size_t half = dbuf && (epoch & 1u) ? (size_t)p.parity_stride : 0;
    const Vec *mystage = (const Vec *)((const char *)p.send[r] + half);
    const Seg *segs = (p.dev_segs ? p.dev_segs : p.inline_segs) + (size_t)ly * p.n_seg;
    const int n_seg = (int)p.n_seg;

    if (threadIdx.x == 0) { s_abort = 0; s_last = 0; }
    __syncthreads();
    GA_TRACE(p, 0);
    if (!(dbuf && ld_relaxed_sys(&mypad->last_kind) == 1u)) {
        if (!wait_flags<W>(mypad->done, epoch - 1, p, &s_abort)) return;
    }
    GA_TRACE(p, 1);

    for (uint32_t it = warp; it < nt * W; it += wpb) {
        const uint32_t t = t0 + it / W;
        const int j = (r + (int)(it % W)) % W;
        const Seg sg = segs[find_seg(segs, n_seg, t)];
        Vec v[kLaneVecs];
        load_grad_tile(sg, t, lane, v);
        Vec *dst = (Vec *)((char *)p.send[j] + half) + ((size_t)r * T + t) * kTileVecs + lane;
#pragma unroll
        for (int u = 0; u < kLaneVecs; ++u) __stcg(dst + u * 32, v[u]);
    }
    GA_TRACE(p, 2);
    const size_t ready_off = offsetof(Pad, ready) / 4 + (size_t)b * kMaxWorld + r;
    raise_flags<W>(p, ready_off, epoch);
    GA_TRACE(p, 3);
    if (!wait_flags<W>(mypad->ready[b], epoch, p, &s_abort)) return;
    GA_TRACE(p, 4);

    for (uint32_t it = warp; it < nt * kLaneVecs; it += wpb) {
        const uint32_t t = t0 + it / kLaneVecs;
        const int u = (int)(it % kLaneVecs);
        const Seg sg = segs[find_seg(segs, n_seg, t)];
        Vec v[W];
#pragma unroll
        for (int s = 0; s < W; ++s) v[s] = __ldcg(mystage + ((size_t)s * T + t) * kTileVecs + u * 32 + lane);
        Vec acc[1] = {v[0]};
#pragma unroll
        for (int s = 1; s < W; ++s) {              // ((g0 + g1) + g2) + ...
            acc[0].x = ElemOp<OP>::add(acc[0].x, v[s].x);
            acc[0].y = ElemOp<OP>::add(acc[0].y, v[s].y);
            acc[0].z = ElemOp<OP>::add(acc[0].z, v[s].z);
            acc[0].w = ElemOp<OP>::add(acc[0].w, v[s].w);
        }
        acc[0].x = ElemOp<OP>::fin(acc[0].x, wf);
        acc[0].y = ElemOp<OP>::fin(acc[0].y, wf);
        acc[0].z = ElemOp<OP>::fin(acc[0].z, wf);
        acc[0].w = ElemOp<OP>::fin(acc[0].w, wf);
        store_grad_vecs<1>(sg, t, lane, u, acc);
    }

    __syncthreads();
    GA_TRACE(p, 8);
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&mypad->finished_blocks, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        if (threadIdx.x == 0) { mypad->finished_blocks = 0; mypad->epoch = epoch; mypad->last_kind = dbuf ? 1u : 0u; }
        const size_t done_off = offsetof(Pad, done) / 4 + r;
        if (threadIdx.x == 0) {
            asm volatile("fence.acq_rel.sys;" ::: "memory");
#pragma unroll
            for (int j = 0; j < W; ++j) st_relaxed_sys((uint32_t *)p.pad[j] + done_off, epoch);
        }
    }
    GA_TRACE(p, 9);
}

// Two-shot push, software-pipelined.  A block owns `group` tile slots per round; round q+1 is
// scattered before round q is folded, and round q-1 is unpacked after round q's results were
// pushed, so the links carry data while flags are in flight and while the local unpack runs:
//
//   scatter(0) | scatter(1) fold(0) | scatter(2) fold(1) unpack(0) | ... | unpack(R-1)
//
// Work inside a phase is spread over the 16 warps as (slot, destination) tiles for scatter and
// unpack and as (slot, 512-byte row) items for the fold, whose W inputs must meet in one thread.
template <int W, int OP>
__global__ void __launch_bounds__(kThreads, 2) ga_push_pipe_kernel(const __grid_constant__ Params p) {
    __shared__ int s_abort;
    __shared__ int s_last;
    const int ly = blockIdx.y;
    const int r = p.rank_base + ly;
    const int b = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t wpb = blockDim.x >> 5;
    const uint32_t T = p.total_tiles, Ts = p.shard_tiles, G = p.group;
    const int R = (int)p.rounds;
    const float wf = (float)W;

    Pad *mypad = p.pad[r];
    const uint32_t epoch = ld_relaxed_sys(&mypad->epoch) + 1;     // first round's flag value
    const Vec *mystage = (const Vec *)p.send[r];
    const Vec *mygather = (const Vec *)p.gather[r];
    const Seg *segs = (p.dev_segs ? p.dev_segs : p.inline_segs) + (size_t)ly * p.n_seg;
    const int n_seg = (int)p.n_seg;
    const size_t ready_off = offsetof(Pad, ready) / 4 + (size_t)b * kMaxWorld + r;
    const size_t mid_off = offsetof(Pad, mid) / 4 + (size_t)b * kMaxWorld + r;

    if (threadIdx.x == 0) { s_abort = 0; s_last = 0; }
    __syncthreads();
    if (!wait_flags<W>(mypad->done, epoch - 1, p, &s_abort)) return;

    auto slot0 = [&](int q) { return ((uint32_t)q * gridDim.x + b) * G; };

    auto scatter = [&](int q) {
        const uint32_t s0 = slot0(q);
        for (uint32_t it = warp; it < G * W; it += wpb) {
            const uint32_t i = s0 + it / W;
            const int j = (r + (int)(it % W)) % W;
            const uint32_t t = j * Ts + i;
            if (i >= Ts || t >= T) continue;
            const Seg sg = segs[find_seg(segs, n_seg, t)];
            Vec v[kLaneVecs];
            load_grad_tile(sg, t, lane, v);
            Vec *dst = (Vec *)p.send[j] + ((size_t)r * Ts + i) * kTileVecs + lane;
#pragma unroll
            for (int u = 0; u < kLaneVecs; ++u) __stcg(dst + u * 32, v[u]);
        }
    };

    auto fold = [&](int q) {
        const uint32_t s0 = slot0(q);
        for (uint32_t it = warp; it < G * kLaneVecs; it += wpb) {
            const uint32_t i = s0 + it / kLaneVecs;
            const int u = (int)(it % kLaneVecs);
            const uint32_t t = r * Ts + i;
            if (i >= Ts || t >= T) continue;
            const Seg sg = segs[find_seg(segs, n_seg, t)];
            Vec v[W];
#pragma unroll
            for (int s = 0; s < W; ++s) v[s] = __ldcg(mystage + ((size_t)s * Ts + i) * kTileVecs + u * 32 + lane);
            Vec acc[1] = {v[0]};
#pragma unroll
            for (int s = 1; s < W; ++s) {          // ((g0 + g1) + g2) + ...
                acc[0].x = ElemOp<OP>::add(acc[0].x, v[s].x);
                acc[0].y = ElemOp<OP>::add(acc[0].y, v[s].y);
                acc[0].z = ElemOp<OP>::add(acc[0].z, v[s].z);
                acc[0].w = ElemOp<OP>::add(acc[0].w, v[s].w);
            }
            acc[0].x = ElemOp<OP>::fin(acc[0].x, wf);
            acc[0].y = ElemOp<OP>::fin(acc[0].y, wf);
            acc[0].z = ElemOp<OP>::fin(acc[0].z, wf);
            acc[0].w = ElemOp<OP>::fin(acc[0].w, wf);
#pragma unroll
            for (int jj = 1; jj < W; ++jj) {
                const int j = (r + jj) % W;
                __stcg((Vec *)p.gather[j] + (size_t)t * kTileVecs + u * 32 + lane, acc[0]);
            }
            store_grad_vecs<1>(sg, t, lane, u, acc);
        }
    };

    auto unpack = [&](int q) {
        if (W == 1) return;
        const uint32_t s0 = slot0(q);
        constexpr uint32_t kPeers = W > 1 ? W - 1 : 1;
        for (uint32_t it = warp; it < G * kPeers; it += wpb) {
            const uint32_t i = s0 + it / kPeers;
            const int j = (r + 1 + (int)(it % kPeers)) % W;
            const uint32_t t = j * Ts + i;
            if (i >= Ts || t >= T) continue;
            const Seg sg = segs[find_seg(segs, n_seg, t)];
            const Vec *src = mygather + (size_t)t * kTileVecs + lane;
            Vec v[kLaneVecs];
#pragma unroll
            for (int u = 0; u < kLaneVecs; ++u) v[u] = __ldcg(src + u * 32);
            store_grad_vecs<kLaneVecs>(sg, t, lane, 0, v);
        }
    };

    scatter(0);
    raise_flags<W>(p, ready_off, epoch);
    for (int q = 0; q < R; ++q) {
        if (q + 1 < R) {
            scatter(q + 1);
            raise_flags<W>(p, ready_off, epoch + q + 1);
        }
        if (!wait_flags<W>(mypad->ready[b], epoch + q, p, &s_abort)) return;
        fold(q);
        raise_flags<W>(p, mid_off, epoch + q);
        if (q > 0) {
            if (!wait_flags<W>(mypad->mid[b], epoch + q - 1, p, &s_abort)) return;
            unpack(q - 1);
        }
    }
    if (!wait_flags<W>(mypad->mid[b], epoch + R - 1, p, &s_abort)) return;
    unpack(R - 1);

    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&mypad->finished_blocks, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        if (threadIdx.x == 0) { mypad->finished_blocks = 0; mypad->epoch = epoch + R - 1; mypad->last_kind = 0; }
        const size_t done_off = offsetof(Pad, done) / 4 + r;
        if (threadIdx.x == 0) {
            asm volatile("fence.acq_rel.sys;" ::: "memory");
#pragma unroll
            for (int j = 0; j < W; ++j) st_relaxed_sys((uint32_t *)p.pad[j] + done_off, epoch + R - 1);
        }
    }
}

// Hosted world, no exchange at all: when every rank of the job lives in this process, their
// gradient tensors are all addressable from one launch, so a tile is read once from each rank's
// tensor, folded in rank order, divided and written back to every rank's tensor.  2*W*S bytes in
// total, no bucket, no flags (each tile is complete in registers before anything is stored).
__device__ __forceinline__ Vec load_grad_vec(const Seg &sg, uint32_t tile, int lane, int u) {
    const uint32_t *src = (const uint32_t *)sg.ptr;
    const size_t n = sg.numel;
    const size_t e = (size_t)(tile - sg.start_tile) * kTileElems + (size_t)(u * 32 + lane) * 4;
    Vec v;
    if ((((uintptr_t)src) & 15) == 0 && e + 4 <= n) {
        v = __ldcs((const Vec *)(src + e));
    } else {
        v.x = e + 0 < n ? __ldcs(src + e + 0) : 0u;
        v.y = e + 1 < n ? __ldcs(src + e + 1) : 0u;
        v.z = e + 2 < n ? __ldcs(src + e + 2) : 0u;
        v.w = e + 3 < n ? __ldcs(src + e + 3) : 0u;
    }
    return v;
}

template <int W, int OP>
__global__ void __launch_bounds__(kThreads, 2) ga_hosted_direct_kernel(const __grid_constant__ Params p) {
    const Seg *segs = p.dev_segs ? p.dev_segs : p.inline_segs;      // [rank][n_seg], same layout for every rank
    const int n_seg = (int)p.n_seg;
    const int lane = threadIdx.x & 31;
    const uint32_t wpb = blockDim.x >> 5;
    const uint32_t stride = gridDim.x * wpb;
    const float wf = (float)W;
    constexpr int NU = W <= 4 ? 2 : 1;                               // vectors per rank in flight per thread
    for (uint32_t t = blockIdx.x * wpb + (threadIdx.x >> 5); t < p.total_tiles; t += stride) {
        const int si = find_seg(segs, n_seg, t);
#pragma unroll
        for (int u0 = 0; u0 < kLaneVecs; u0 += NU) {
            Vec v[W][NU];
#pragma unroll
            for (int r = 0; r < W; ++r)
#pragma unroll
                for (int u = 0; u < NU; ++u) v[r][u] = load_grad_vec(segs[(size_t)r * n_seg + si], t, lane, u0 + u);
            Vec acc[NU];
#pragma unroll
            for (int u = 0; u < NU; ++u) {
                acc[u] = v[0][u];
#pragma unroll
                for (int r = 1; r < W; ++r) {      // ((g0 + g1) + g2) + ...
                    acc[u].x = ElemOp<OP>::add(acc[u].x, v[r][u].x);
                    acc[u].y = ElemOp<OP>::add(acc[u].y, v[r][u].y);
                    acc[u].z = ElemOp<OP>::add(acc[u].z, v[r][u].z);
                    acc[u].w = ElemOp<OP>::add(acc[u].w, v[r][u].w);
                }
                acc[u].x = ElemOp<OP>::fin(acc[u].x, wf);
                acc[u].y = ElemOp<OP>::fin(acc[u].y, wf);
                acc[u].z = ElemOp<OP>::fin(acc[u].z, wf);
                acc[u].w = ElemOp<OP>::fin(acc[u].w, wf);
            }
#pragma unroll
            for (int r = 0; r < W; ++r) store_grad_vecs<NU>(segs[(size_t)r * n_seg + si], t, lane, u0, acc);
        }
    }
}

// Parameter broadcast: the reference builds Net() on every rank and relies on
// identical seeds it never sets (/root/reference/mnist/main.py:185).  The root packs its tensors
// straight into every peer's gather region (remote stores, pack folded in), flags per block column;
// the peers unpack their column into their own tensors.  Same flag protocol and epoch as the averaging
// kernels, so calls can be mixed freely on one context.
template <int W>
__global__ void __launch_bounds__(kThreads, 2) ga_bcast_kernel(const __grid_constant__ Params p) {
    __shared__ int s_abort;
    __shared__ int s_last;
    const int ly = blockIdx.y;
    const int r = p.rank_base + ly;
    const int b = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int wpb = blockDim.x >> 5;
    const uint32_t stride = gridDim.x * wpb;
    const uint32_t first = b * wpb + warp;
    const uint32_t T = p.total_tiles;
    const int root = p.root;

    Pad *mypad = p.pad[r];
    const uint32_t epoch = ld_relaxed_sys(&mypad->epoch) + 1;
    const Seg *segs = (p.dev_segs ? p.dev_segs : p.inline_segs) + (size_t)ly * p.n_seg;
    const int n_seg = (int)p.n_seg;
    if (threadIdx.x == 0) { s_abort = 0; s_last = 0; }
    __syncthreads();
    // every rank has left its previous launch: the gather regions are free to be overwritten
    if (!wait_flags<W>(mypad->done, epoch - 1, p, &s_abort)) return;

    if (r == root) {
        for (uint32_t t = first; t < T; t += stride) {
            const Seg sg = segs[find_seg(segs, n_seg, t)];
            Vec v[kLaneVecs];
            load_grad_tile(sg, t, lane, v);
#pragma unroll 1
            for (int jj = 1; jj < W; ++jj) {
                Vec *dst = (Vec *)p.gather[(r + jj) % W] + (size_t)t * kTileVecs + lane;
#pragma unroll
                for (int u = 0; u < kLaneVecs; ++u) __stcg(dst + u * 32, v[u]);
            }
        }
        raise_flags<W>(p, offsetof(Pad, ready) / 4 + (size_t)b * kMaxWorld + root, epoch);
    } else {
        if (threadIdx.x == 0) {
            if (!spin_until(&mypad->ready[b][root], epoch, p.timeout_ns)) {
                s_abort = 1;
                *(volatile int *)p.status = GA_ETIMEOUT;
            }
            asm volatile("fence.acq_rel.sys;" ::: "memory");
        }
        __syncthreads();
        if (s_abort) return;
        const Vec *mygather = (const Vec *)p.gather[r];
        for (uint32_t t = first; t < T; t += stride) {
            const Seg sg = segs[find_seg(segs, n_seg, t)];
            const Vec *src = mygather + (size_t)t * kTileVecs + lane;
            Vec v[kLaneVecs];
#pragma unroll
            for (int u = 0; u < kLaneVecs; ++u) v[u] = __ldcg(src + u * 32);
            store_grad_vecs<kLaneVecs>(sg, t, lane, 0, v);
        }
    }

    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&mypad->finished_blocks, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
        mypad->finished_blocks = 0; mypad->epoch = epoch; mypad->last_kind = 0;
        asm volatile("fence.acq_rel.sys;" ::: "memory");
#pragma unroll
        for (int j = 0; j < W; ++j) st_relaxed_sys((uint32_t *)p.pad[j] + offsetof(Pad, done) / 4 + r, epoch);
    }
}

// Launched only when a bucket-direct call follows a tensor-list call: the caller is about to
// overwrite the bucket from the host side of the stream, so every peer must have finished
// reading it (tensor-list calls check this lazily, at the start of the NEXT kernel).
__global__ void ga_drain_kernel(const Pad *const *pads, int n_local, int world,
                                unsigned long long timeout_ns, int *status) {
    const int l = threadIdx.x / kMaxWorld, src = threadIdx.x % kMaxWorld;
    if (l < n_local && src < world) {
        const uint32_t epoch = ld_relaxed_sys(&pads[l]->epoch);     // last value this rank's kernels used
        if (!spin_until(&pads[l]->done[src], epoch, timeout_ns)) *(volatile int *)status = GA_ETIMEOUT;
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

int ga_fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

// Launches of one context share its flag epoch and staging regions, so they must be ordered with
// respect to each other.  On one stream they are; when the caller comes back on ANOTHER stream the
// new stream is made to wait for the previous launch here.  (While either stream is being captured
// nothing is inserted: inside a graph the caller's own edges are the ordering.)
int ga_order_stream(ga_ctx *c, void *stream) {
    if (c->have_last_stream && c->last_stream != stream) {
        cudaStreamCaptureStatus a = cudaStreamCaptureStatusNone, b = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing((cudaStream_t)c->last_stream, &a) != cudaSuccess) { cudaGetLastError(); a = cudaStreamCaptureStatusInvalidated; }
        if (cudaStreamIsCapturing((cudaStream_t)stream, &b) != cudaSuccess) { cudaGetLastError(); b = cudaStreamCaptureStatusInvalidated; }
        if (a == cudaStreamCaptureStatusNone && b == cudaStreamCaptureStatusNone) {
            if (!c->order_event) CUDA_TRY(cudaEventCreateWithFlags(&c->order_event, cudaEventDisableTiming));
            if (cudaEventRecord(c->order_event, (cudaStream_t)c->last_stream) == cudaSuccess)
                CUDA_TRY(cudaStreamWaitEvent((cudaStream_t)stream, c->order_event, 0));
            else
                cudaGetLastError();      // the previous stream no longer exists: nothing left to order against
        }
    }
    c->last_stream = stream;
    c->have_last_stream = true;
    return GA_OK;
}

namespace {

typedef void (*KernelFn)(Params);

template <int W, int VARIANT> KernelFn pick_op(int op) {
    return op == GA_OP_I32_SUM   ? (KernelFn)ga_fused_kernel<W, VARIANT, GA_OP_I32_SUM>
           : op == GA_OP_F32_SUM ? (KernelFn)ga_fused_kernel<W, VARIANT, GA_OP_F32_SUM>
                                 : (KernelFn)ga_fused_kernel<W, VARIANT, GA_OP_F32_AVG>;
}
template <int W> KernelFn pick_push_op(int op) {
    return op == GA_OP_I32_SUM   ? (KernelFn)ga_push_kernel<W, GA_OP_I32_SUM>
           : op == GA_OP_F32_SUM ? (KernelFn)ga_push_kernel<W, GA_OP_F32_SUM>
                                 : (KernelFn)ga_push_kernel<W, GA_OP_F32_AVG>;
}
template <int W> KernelFn pick_push1_op(int op) {
    return op == GA_OP_I32_SUM   ? (KernelFn)ga_push1_kernel<W, GA_OP_I32_SUM>
           : op == GA_OP_F32_SUM ? (KernelFn)ga_push1_kernel<W, GA_OP_F32_SUM>
                                 : (KernelFn)ga_push1_kernel<W, GA_OP_F32_AVG>;
}
template <int W> KernelFn pick_pipe_op(int op) {
    return op == GA_OP_I32_SUM   ? (KernelFn)ga_push_pipe_kernel<W, GA_OP_I32_SUM>
           : op == GA_OP_F32_SUM ? (KernelFn)ga_push_pipe_kernel<W, GA_OP_F32_SUM>
                                 : (KernelFn)ga_push_pipe_kernel<W, GA_OP_F32_AVG>;
}
template <int W> KernelFn pick_hosted_op(int op) {
    return op == GA_OP_I32_SUM   ? (KernelFn)ga_hosted_direct_kernel<W, GA_OP_I32_SUM>
           : op == GA_OP_F32_SUM ? (KernelFn)ga_hosted_direct_kernel<W, GA_OP_F32_SUM>
                                 : (KernelFn)ga_hosted_direct_kernel<W, GA_OP_F32_AVG>;
}
template <int W> KernelFn pick_variant(int variant, int op) {
    switch (variant) {
        case GA_VARIANT_HOSTED_DIRECT: return pick_hosted_op<W>(op);
        case GA_VARIANT_TWOSHOT_PIPE: return pick_pipe_op<W>(op);
        case GA_VARIANT_TWOSHOT: return pick_op<W, GA_VARIANT_TWOSHOT>(op);
        case GA_VARIANT_ONESHOT_PUSH: return pick_push1_op<W>(op);
        case GA_VARIANT_TWOSHOT_PUSH: return pick_push_op<W>(op);
    }
    return pick_op<W, GA_VARIANT_ONESHOT>(op);
}
constexpr int kVariantBcast = 15;    // internal: ga_bcast_kernel through the common launch path
KernelFn pick_bcast(int world) {
    switch (world) {
        case 1: return (KernelFn)ga_bcast_kernel<1>;
        case 2: return (KernelFn)ga_bcast_kernel<2>;
        case 3: return (KernelFn)ga_bcast_kernel<3>;
        case 4: return (KernelFn)ga_bcast_kernel<4>;
        case 5: return (KernelFn)ga_bcast_kernel<5>;
        case 6: return (KernelFn)ga_bcast_kernel<6>;
        case 7: return (KernelFn)ga_bcast_kernel<7>;
        case 8: return (KernelFn)ga_bcast_kernel<8>;
    }
    return nullptr;
}
KernelFn pick_kernel(int world, int variant, int op) {
    if (variant == kVariantBcast) return pick_bcast(world);
    switch (world) {
        case 1: return pick_variant<1>(variant, op);
        case 2: return pick_variant<2>(variant, op);
        case 3: return pick_variant<3>(variant, op);
        case 4: return pick_variant<4>(variant, op);
        case 5: return pick_variant<5>(variant, op);
        case 6: return pick_variant<6>(variant, op);
        case 7: return pick_variant<7>(variant, op);
        case 8: return pick_variant<8>(variant, op);
    }
    return nullptr;
}

}  // namespace

namespace {

size_t tiles_for_elems(int64_t numel) { return (size_t)((numel + kTileElems - 1) / kTileElems); }

// AUTO, one rank per GPU: which kernel is quickest for a set of `bytes`, per world-size class, read
// off the committed sweeps of this pool (profiles/r02_bench_n{2,4,8}_sweep.json; tests/test_abi.py
// re-checks every row of those files: AUTO's pick must be within 3 % of the best fused variant).
//   world 2      one-shot push up to 32 MiB (same link bytes as two-shot, one barrier less), two-shot push beyond
//   world 3..8   one-shot push while its (world-1)*S egress stays below 8 MiB, then the pipelined kernel in
//                its one-round form up to 32 MiB, two-shot push beyond (its rounds no longer pay: the round-1
//                gain of 3-10 % between 32 and 128 MiB is gone since the barriers poll with one thread)
struct AutoRow { int world_max; size_t bytes_max_num, bytes_max_div_world_minus_1; int variant; };
constexpr AutoRow kAutoTable[] = {
    {2, (size_t)32 << 20, 0, GA_VARIANT_ONESHOT_PUSH},
    {2, (size_t)-1, 0, GA_VARIANT_TWOSHOT_PUSH},
    {8, (size_t)8 << 20, 1, GA_VARIANT_ONESHOT_PUSH},
    {8, (size_t)32 << 20, 0, GA_VARIANT_TWOSHOT_PIPE},
    {8, (size_t)-1, 0, GA_VARIANT_TWOSHOT_PUSH},
};

int auto_from_table(int world, size_t bytes) {
    for (const AutoRow &r : kAutoTable) {
        if (world > r.world_max) continue;
        const size_t lim = r.bytes_max_div_world_minus_1 ? r.bytes_max_num / (size_t)(world > 1 ? world - 1 : 1) : r.bytes_max_num;
        if (bytes <= lim) return r.variant;
    }
    return GA_VARIANT_TWOSHOT_PUSH;
}

// Thresholds of the remaining cases (hosted sub-worlds move everything through HBM and prefer the
// pull kernels) and of GA_OPT_ONESHOT_MAX_BYTES, which overrides the table's first row when set.
void set_auto_thresholds(ga_ctx *c) {
    const int world = c->world, n_local = c->n_local;
    if (world <= 2) c->oneshot_max_bytes = (size_t)32 << 20;
    else if (n_local > 1) c->oneshot_max_bytes = (size_t)512 << 10;
    else c->oneshot_max_bytes = ((size_t)8 << 20) / (size_t)(world - 1);
    c->pipe_min_bytes = (size_t)32 << 20;
    c->pipe_max_bytes = (size_t)32 << 20;
}

int create_body(ga_ctx *c, int device);

int create_common(int rank_base, int n_local, int world, int device, size_t capacity_bytes, ga_ctx **out) {
    if (!out) return fail(GA_EINVAL, "out is NULL");
    *out = nullptr;
    if (world < 1 || world > kMaxWorld) return fail(GA_EINVAL, "world=%d outside 1..%d", world, kMaxWorld);
    if (n_local < 1 || rank_base < 0 || rank_base + n_local > world)
        return fail(GA_EINVAL, "rank=%d (+%d hosted) outside world=%d", rank_base, n_local, world);
    if (capacity_bytes == 0) return fail(GA_EINVAL, "capacity_bytes is 0");
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(GA_EINVAL, "device=%d but %d CUDA devices visible", device, ndev);
    ENTER_DEVICE(device);

    ga_ctx *c = (ga_ctx *)calloc(1, sizeof(ga_ctx));
    if (!c) return fail(GA_EINVAL, "out of host memory");
    c->world = world; c->rank_base = rank_base; c->n_local = n_local; c->device = device;
    c->capacity_bytes = (capacity_bytes + kTileBytes - 1) / kTileBytes * kTileBytes;
    if (c->capacity_bytes / kTileBytes > 0xffffffffull) { free(c); return fail(GA_ECAPACITY, "capacity too large"); }
    c->region_bytes = c->capacity_bytes + (size_t)kMaxWorld * kTileBytes;
    set_auto_thresholds(c);
    c->timeout_ns = 30ull * 1000000000ull;
    c->host_chunk_bytes = (size_t)8 << 20;
    c->view_oneshot_max_opt = -1;
    // world 2: two-shot push on the region is as quick as the in-place kernel at every size (r02 N=2 sweep)
    c->view_min_bytes = world <= 2 ? (size_t)-1 : (size_t)32 << 20;
    const char *env = getenv("GA_TIMEOUT_MS");
    if (env) c->timeout_ns = strtoull(env, nullptr, 10) * 1000000ull;
    const int rc = create_body(c, device);
    if (rc != GA_OK) {
        // whatever was allocated before the failing call is released (ga_destroy tolerates the gaps);
        // the message of the failure is kept
        char keep[sizeof(g_err)];
        memcpy(keep, g_err, sizeof keep);
        ga_destroy(c);
        memcpy(g_err, keep, sizeof keep);
        return rc;
    }
    c->opened = (n_local == world);
    *out = c;
    return GA_OK;
}

int create_body(ga_ctx *c, int device) {
    const int n_local = c->n_local, rank_base = c->rank_base;
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;

    for (int l = 0; l < n_local; ++l) {
        CUDA_TRY(cudaMalloc(&c->base[l], kPadBytes + 2 * c->region_bytes));
        CUDA_TRY(cudaMemset(c->base[l], 0, kPadBytes));
        CUDA_TRY(cudaMalloc(&c->recv[l], c->capacity_bytes));
        c->pad_all[rank_base + l] = (Pad *)c->base[l];
        c->send_all[rank_base + l] = (char *)c->base[l] + kPadBytes;
        c->gather_all[rank_base + l] = (char *)c->base[l] + kPadBytes + c->region_bytes;
    }
    CUDA_TRY(cudaHostAlloc((void **)&c->status_host, sizeof(int), cudaHostAllocMapped));
    *c->status_host = 0;
    CUDA_TRY(cudaHostGetDevicePointer((void **)&c->status_dev, c->status_host, 0));
    {
        const Pad *own[kMaxWorld] = {};
        for (int l = 0; l < n_local; ++l) own[l] = (const Pad *)c->base[l];
        CUDA_TRY(cudaMalloc((void **)&c->own_pads_dev, sizeof(own)));
        CUDA_TRY(cudaMemcpy(c->own_pads_dev, own, sizeof(own), cudaMemcpyHostToDevice));
    }
    CUDA_TRY(cudaDeviceSynchronize());
    return GA_OK;
}

int ensure_seg_ring(ga_ctx *c, size_t entries) {
    if (entries <= c->seg_cap) return GA_OK;
    size_t cap = 1024;
    while (cap < entries) cap *= 2;
    for (int i = 0; i < kSegRing; ++i) {
        if (c->seg_used[i]) CUDA_TRY(cudaEventSynchronize(c->seg_event[i]));
        if (c->seg_host[i]) CUDA_TRY(cudaFreeHost(c->seg_host[i]));
        if (c->seg_dev[i]) CUDA_TRY(cudaFree(c->seg_dev[i]));
        c->seg_host[i] = nullptr; c->seg_dev[i] = nullptr;
        CUDA_TRY(cudaHostAlloc((void **)&c->seg_host[i], cap * sizeof(Seg), cudaHostAllocDefault));
        CUDA_TRY(cudaMalloc((void **)&c->seg_dev[i], cap * sizeof(Seg)));
        if (!c->seg_event[i]) CUDA_TRY(cudaEventCreateWithFlags(&c->seg_event[i], cudaEventDisableTiming));
        c->seg_used[i] = false;
    }
    c->seg_cap = cap;
    return GA_OK;
}

int blocks_per_sm(ga_ctx *c, KernelFn fn, int variant) {
    int &cached = c->blocks_per_sm[variant];
    if (cached == 0) {
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, (const void *)fn, kThreads, 0) != cudaSuccess || n < 1) n = 1;
        cached = n;
    }
    return cached;
}

// Common launch path.  `segs` (n_local * n_seg entries, rank-major) is nullptr for bucket-direct.
int launch(ga_ctx *c, const Seg *segs, uint32_t n_seg, uint32_t total_tiles, int op, void *stream, int variant, bool direct,
           size_t byte_offset = 0, int root = 0, int max_blocks = 0) {
    if (c->broken) return fail(GA_ESTATE, "context is unusable after an earlier timeout");
    if (*(volatile int *)c->status_host != 0) {
        // a kernel of this context already gave up on a peer: fail now instead of queueing more
        // launches that would each spin until the timeout
        c->broken = true;
        return fail(GA_ETIMEOUT, "an earlier kernel timed out waiting for a peer rank's flag; the context is unusable");
    }
    if (!c->opened) return fail(GA_ESTATE, "ga_open_peers has not been called");
    if (op != GA_OP_F32_AVG && op != GA_OP_I32_SUM && op != GA_OP_F32_SUM) return fail(GA_EINVAL, "unknown op %d", op);
    if (variant == GA_VARIANT_AUTO) {
        variant = ga_pick_variant(c, (size_t)total_tiles * kTileBytes);
        if (direct && variant == GA_VARIANT_HOSTED_DIRECT)
            variant = (size_t)total_tiles * kTileBytes <= c->oneshot_max_bytes ? GA_VARIANT_ONESHOT : GA_VARIANT_TWOSHOT;
    }
    if ((variant < GA_VARIANT_ONESHOT || variant > GA_VARIANT_HOSTED_DIRECT) && variant != kVariantBcast)
        return fail(GA_EINVAL, "unknown variant %d (GA_VARIANT_VIEW / GA_VARIANT_NVLS belong to ga_average_view)", variant);
    if (variant == GA_VARIANT_HOSTED_DIRECT && c->n_local != c->world)
        return fail(GA_EINVAL, "GA_VARIANT_HOSTED_DIRECT needs every rank hosted in this process (ga_create_local)");
    if (variant == GA_VARIANT_HOSTED_DIRECT && direct)
        variant = (size_t)total_tiles * kTileBytes <= c->oneshot_max_bytes ? GA_VARIANT_ONESHOT : GA_VARIANT_TWOSHOT;
    if (total_tiles == 0) return GA_OK;    // nothing to average: no launch, epoch unchanged on every rank
    // Bucket-direct calls exist in the pull form only (their input already sits in the bucket).
    if (direct && variant == GA_VARIANT_ONESHOT_PUSH) variant = GA_VARIANT_ONESHOT;
    if (direct && (variant == GA_VARIANT_TWOSHOT_PUSH || variant == GA_VARIANT_TWOSHOT_PIPE)) variant = GA_VARIANT_TWOSHOT;
    // One-shot push stages W copies of the set in every rank's bucket; beyond that, two-shot push.
    if (variant == GA_VARIANT_ONESHOT_PUSH && (size_t)total_tiles * kTileBytes * c->world > c->capacity_bytes)
        variant = GA_VARIANT_TWOSHOT_PUSH;
    ENTER_DEVICE(c->device);
    cudaStream_t st = (cudaStream_t)stream;

    Params p;
    memset(&p, 0, sizeof(p));
    // byte_offset: the call works on the sub-bucket [byte_offset, byte_offset + tiles) of every rank
    for (int r = 0; r < c->world; ++r) {
        p.send[r] = (char *)c->send_all[r] + byte_offset;
        p.gather[r] = c->gather_all[r];
        p.pad[r] = c->pad_all[r];
    }
    for (int l = 0; l < c->n_local; ++l) p.recv[l] = (char *)c->recv[l] + byte_offset;
    p.status = c->status_dev;
    p.timeout_ns = c->timeout_ns;
    p.n_seg = n_seg;
    p.total_tiles = total_tiles;
    p.shard_tiles = (total_tiles + c->world - 1) / c->world;
    p.rank_base = c->rank_base;
    p.direct = direct ? 1 : 0;
    p.trace = c->trace_dev;
    p.root = root;

    if (segs) {
        const size_t entries = (size_t)n_seg * c->n_local;
        if (entries <= (size_t)kInlineSegs) {
            memcpy(p.inline_segs, segs, entries * sizeof(Seg));
        } else {
            int rc = ensure_seg_ring(c, entries);
            if (rc) return rc;
            const int s = c->seg_pos;
            c->seg_pos = (s + 1) % kSegRing;
            if (c->seg_used[s]) CUDA_TRY(cudaEventSynchronize(c->seg_event[s]));
            memcpy(c->seg_host[s], segs, entries * sizeof(Seg));
            CUDA_TRY(cudaMemcpyAsync(c->seg_dev[s], c->seg_host[s], entries * sizeof(Seg), cudaMemcpyHostToDevice, st));
            CUDA_TRY(cudaEventRecord(c->seg_event[s], st));
            c->seg_used[s] = true;
            p.dev_segs = c->seg_dev[s];
        }
    }

    KernelFn fn = pick_kernel(c->world, variant, op);
    // Grid: one warp per tile slot, at most what is co-resident (blocks spin on peers' flags).
    const bool two_shot = variant == GA_VARIANT_TWOSHOT || variant == GA_VARIANT_TWOSHOT_PUSH || variant == GA_VARIANT_TWOSHOT_PIPE;
    const uint32_t slots = two_shot ? p.shard_tiles : total_tiles;
    const int wpb = kThreads / 32;
    int resident = blocks_per_sm(c, fn, variant) * c->sm_count / c->n_local;
    if (resident > kMaxBlocks) resident = kMaxBlocks;
    if (c->max_blocks_opt > 0 && resident > c->max_blocks_opt) resident = c->max_blocks_opt;
    if (max_blocks > 0 && resident > max_blocks) resident = max_blocks;       // per-call cap (overlap mode)
    if (resident < 1) resident = 1;
    int blocks = (int)((slots + wpb - 1) / wpb);
    uint32_t rounds = 1;
    if (variant == GA_VARIANT_TWOSHOT_PIPE) {
        // Two regimes, both measured at 8 GPUs (profiles/r01b_explore_n8.md, r01d_explore_mid_n8.md):
        //  * below pipe_min_bytes one round, and the slots spread over ~64 blocks (2..16 per block) so
        //    that scatter, fold and unpack are each one or two items per warp instead of a serial
        //    chain of W load->store steps;
        //  * from pipe_min_bytes on, rounds of >= 32 MiB per GPU on one block per SM: a round costs a
        //    flag round trip of a few tens of microseconds and must move enough bytes to hide it.
        uint32_t group = (uint32_t)c->pipe_group_opt;
        if ((size_t)total_tiles * kTileBytes < c->pipe_min_bytes) {
            if (group == 0) {
                group = (slots + 63) / 64;
                if (group < 2) group = 2;
                if (group > 16) group = 16;
            }
        } else {
            if (c->max_blocks_opt == 0 && max_blocks == 0 && resident > c->sm_count / c->n_local) resident = c->sm_count / c->n_local;
            if (resident < 1) resident = 1;
            if (group == 0) {
                const size_t per_slot = (size_t)c->world * kTileBytes * (size_t)resident;
                group = (uint32_t)((((size_t)32 << 20) + per_slot - 1) / per_slot);
                if (group < 16) group = 16;
                if (group > 64) group = 64;
            }
        }
        blocks = (int)((slots + group - 1) / group);
        if (blocks > resident) blocks = resident;
        if (blocks < 1) blocks = 1;
        rounds = (slots + (uint32_t)blocks * group - 1) / ((uint32_t)blocks * group);
        if (rounds < 1) rounds = 1;
        p.group = group;
    }
    if (variant == GA_VARIANT_ONESHOT_PUSH) {
        // one block per `group` tiles, as many blocks as are resident; two staging halves if they fit
        // at least 16 (tile, destination) items per block so that every warp scatters
        uint32_t group = (total_tiles + (uint32_t)resident - 1) / (uint32_t)resident;
        // (one item per warp up to 2 MiB, two beyond: N=2 sweeps r01k/r01l)
        const uint32_t per_block = total_tiles <= 1024 ? kThreads / 32 : 2 * (kThreads / 32);
        const uint32_t min_group = (per_block + (uint32_t)c->world - 1) / (uint32_t)c->world;
        if (group < min_group) group = min_group;
        blocks = (int)((total_tiles + group - 1) / group);
        p.group = group;
        // The odd half starts at a FIXED offset (middle of the bucket), whatever the set size: with a
        // size-dependent offset a smaller set's odd half could land inside a larger predecessor's even half.
        const size_t one = (size_t)total_tiles * kTileBytes * c->world;
        const size_t mid = c->capacity_bytes / 2 / kTileBytes * kTileBytes;
        p.parity_stride = one <= mid ? mid : 0;
    }
    p.rounds = rounds;
    if (blocks > resident) blocks = resident;
    if (blocks < 1) blocks = 1;

    if (variant == GA_VARIANT_HOSTED_DIRECT) {
        // one launch over the tiles, every rank's tensors addressed directly: no flags, no epoch
        int all = blocks_per_sm(c, fn, variant) * c->sm_count;
        if (c->max_blocks_opt > 0 && all > c->max_blocks_opt) all = c->max_blocks_opt;
        if (max_blocks > 0 && all > max_blocks) all = max_blocks;
        blocks = (int)((total_tiles + wpb - 1) / wpb);
        if (blocks > all) blocks = all;
        if (blocks < 1) blocks = 1;
        void *hargs[] = {&p};
        CUDA_TRY(cudaLaunchKernel((const void *)fn, dim3(blocks), dim3(kThreads), hargs, 0, st));
        c->launches += 1;
        c->last_blocks = blocks;
        c->last_threads = kThreads;
        return GA_OK;
    }
    dim3 grid(blocks, c->n_local), block(kThreads);
    void *args[] = {&p};
    if (c->n_local > 1) {
        // hosted ranks wait on one another inside one launch: co-residency must be guaranteed
        CUDA_TRY(cudaLaunchCooperativeKernel((const void *)fn, grid, block, args, 0, st));
    } else {
        CUDA_TRY(cudaLaunchKernel((const void *)fn, grid, block, args, 0, st));
    }
    c->launches += 1;
    c->last_was_tensor_call = !direct;
    c->last_blocks = blocks;
    c->last_threads = kThreads;
    return GA_OK;
}

// Called before the host side of the stream overwrites the bucket (bucket-direct callers do
// that themselves, ga_average_host does it with its H2D copies).
int drain_if_needed(ga_ctx *c, void *stream) {
    if (!c->last_was_tensor_call) return GA_OK;
    ENTER_DEVICE(c->device);
    ga_drain_kernel<<<1, kMaxWorld * kMaxWorld, 0, (cudaStream_t)stream>>>(
        c->own_pads_dev, c->n_local, c->world, c->timeout_ns, c->status_dev);
    CUDA_TRY(cudaGetLastError());
    c->launches += 1;
    c->last_was_tensor_call = false;
    return GA_OK;
}

int ensure_host_pipeline(ga_ctx *c) {
    if (c->s_in) return GA_OK;
    CUDA_TRY(cudaStreamCreateWithFlags(&c->s_in, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&c->s_k, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&c->s_out, cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreateWithFlags(&c->ev_start, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&c->ev_done, cudaEventDisableTiming));
    for (int i = 0; i < kHostChunks; ++i) {
        CUDA_TRY(cudaEventCreateWithFlags(&c->ev_in[i], cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&c->ev_k[i], cudaEventDisableTiming));
    }
    return GA_OK;
}

// Copies the parts of every hosted rank's tensors that fall into bucket bytes [lo, hi) between the
// host tensors and `dev_base[l] + bucket offset` (to_device) or back (!to_device).
int copy_chunk(ga_ctx *c, void *const *host_ptrs, const int64_t *numel, int n_tensors, size_t lo, size_t hi,
               bool to_device, cudaStream_t st) {
    for (int l = 0; l < c->n_local; ++l) {
        char *dev = to_device ? (char *)c->send_all[c->rank_base + l] : (char *)c->recv[l];
        size_t off = 0;                                       // bucket offset of tensor i
        for (int i = 0; i < n_tensors; ++i) {
            const size_t bytes = (size_t)numel[i] * 4;
            const size_t a = off > lo ? off : lo;
            const size_t b = off + bytes < hi ? off + bytes : hi;
            if (a < b) {
                char *h = (char *)host_ptrs[(size_t)l * n_tensors + i] + (a - off);
                if (to_device) CUDA_TRY(cudaMemcpyAsync(dev + a, h, b - a, cudaMemcpyHostToDevice, st));
                else CUDA_TRY(cudaMemcpyAsync(h, dev + a, b - a, cudaMemcpyDeviceToHost, st));
            }
            off += tiles_for_elems(numel[i]) * kTileBytes;
            if (off >= hi) break;
        }
    }
    return GA_OK;
}

// Fills rank-major Seg entries + returns total tiles.  ptrs may be nullptr (layout only).
int build_layout(const int64_t *numel, int n_tensors, uint64_t *total_tiles) {
    uint64_t t = 0;
    for (int i = 0; i < n_tensors; ++i) {
        if (numel[i] < 0) return fail(GA_EINVAL, "numel[%d] is negative", i);
        if (numel[i] > 0xffffffffll) return fail(GA_EINVAL, "numel[%d] exceeds 2^32-1 elements", i);
        t += tiles_for_elems(numel[i]);
    }
    *total_tiles = t;
    return GA_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------
extern "C" {

int ga_status(ga_ctx *c);

const char *ga_last_error(void) { return g_err; }

int ga_create(int rank, int world, int device, size_t capacity_bytes, ga_ctx **out) {
    return create_common(rank, 1, world, device, capacity_bytes, out);
}

int ga_create_local(int world, int device, size_t capacity_bytes, ga_ctx **out) {
    return create_common(0, world, world, device, capacity_bytes, out);
}

int ga_export_handle(ga_ctx *c, unsigned char out[GA_IPC_HANDLE_BYTES]) {
    if (!c || !out) return fail(GA_EINVAL, "NULL argument");
    if (c->n_local != 1) return fail(GA_ESTATE, "a local world has no handles to exchange");
    static_assert(sizeof(cudaIpcMemHandle_t) == GA_IPC_HANDLE_BYTES, "IPC handle size");
    ENTER_DEVICE(c->device);
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, c->base[0]));
    memcpy(out, &h, sizeof(h));
    return GA_OK;
}

int ga_open_peers(ga_ctx *c, const unsigned char *handles) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (c->opened) return c->n_local == c->world ? GA_OK : fail(GA_ESTATE, "peers already opened");
    if (!handles) return fail(GA_EINVAL, "NULL handles");
    ENTER_DEVICE(c->device);
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank_base) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + (size_t)r * GA_IPC_HANDLE_BYTES, sizeof(h));
        void *ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            // leave nothing half-opened: a retry would map the earlier peers a second time
            for (int q = 0; q < c->world; ++q) {
                if (!c->peer_base[q]) continue;
                cudaIpcCloseMemHandle(c->peer_base[q]);
                c->peer_base[q] = nullptr; c->pad_all[q] = nullptr; c->send_all[q] = nullptr; c->gather_all[q] = nullptr;
            }
            return fail(GA_ECUDA, "cudaIpcOpenMemHandle(rank %d) failed: %s", r, cudaGetErrorString(e));
        }
        c->peer_base[r] = ptr;
        c->pad_all[r] = (Pad *)ptr;
        c->send_all[r] = (char *)ptr + kPadBytes;
        c->gather_all[r] = (char *)ptr + kPadBytes + c->region_bytes;
    }
    c->opened = true;
    return GA_OK;
}

namespace {
int tensors_call(ga_ctx *c, const void *const *ptrs, const int64_t *numel, int n_tensors,
                 int op, void *stream, int variant, int root, int max_blocks) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (n_tensors < 0 || (n_tensors > 0 && (!ptrs || !numel))) return fail(GA_EINVAL, "bad tensor list");
    uint64_t total = 0;
    int rc = build_layout(numel, n_tensors, &total);
    if (rc) return rc;
    if (total * kTileBytes > c->capacity_bytes)
        return fail(GA_ECAPACITY, "gradient set needs %llu bucket bytes, capacity is %zu",
                    (unsigned long long)(total * kTileBytes), c->capacity_bytes);
    if (total == 0) return GA_OK;
    const size_t entries = (size_t)n_tensors * c->n_local;
    Seg stack_segs[kInlineSegs];
    Seg *segs = entries <= (size_t)kInlineSegs ? stack_segs : (Seg *)malloc(entries * sizeof(Seg));
    if (!segs) return fail(GA_EINVAL, "out of host memory");
    for (int l = 0; l < c->n_local; ++l) {
        uint32_t start = 0;
        for (int i = 0; i < n_tensors; ++i) {
            Seg &s = segs[(size_t)l * n_tensors + i];
            s.ptr = ptrs[(size_t)l * n_tensors + i];
            s.start_tile = start;
            s.numel = (uint32_t)numel[i];
            if (numel[i] > 0 && !s.ptr) {
                if (segs != stack_segs) free(segs);
                return fail(GA_EINVAL, "pointer %d of hosted rank %d is NULL", i, l);
            }
            if (((uintptr_t)s.ptr & 3) != 0) {
                if (segs != stack_segs) free(segs);
                return fail(GA_EINVAL, "pointer %d of hosted rank %d is not 4-byte aligned", i, l);
            }
            start += (uint32_t)tiles_for_elems(numel[i]);
        }
    }
    rc = ga_order_stream(c, stream);
    if (rc == GA_OK) rc = launch(c, segs, (uint32_t)n_tensors, (uint32_t)total, op, stream, variant, false, 0, root, max_blocks);
    if (segs != stack_segs) free(segs);
    return rc;
}
}  // namespace

int ga_allreduce_tensors(ga_ctx *c, const void *const *ptrs, const int64_t *numel, int n_tensors,
                         int op, void *stream, int variant) {
    return tensors_call(c, ptrs, numel, n_tensors, op, stream, variant, 0, 0);
}

int ga_allreduce_tensors_ex(ga_ctx *c, const void *const *ptrs, const int64_t *numel, int n_tensors,
                            int op, void *stream, int variant, int max_blocks) {
    if (max_blocks < 0) return fail(GA_EINVAL, "max_blocks is negative");
    return tensors_call(c, ptrs, numel, n_tensors, op, stream, variant, 0, max_blocks);
}

int ga_broadcast_tensors(ga_ctx *c, const void *const *ptrs, const int64_t *numel, int n_tensors, int root, void *stream) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (root < 0 || root >= c->world) return fail(GA_EINVAL, "root=%d outside world=%d", root, c->world);
    return tensors_call(c, ptrs, numel, n_tensors, GA_OP_F32_SUM, stream, kVariantBcast, root, 0);
}


int ga_average(ga_ctx *c, const void *const *grad_ptrs, const int64_t *numel, int n_tensors,
               void *stream, int variant) {
    return ga_allreduce_tensors(c, grad_ptrs, numel, n_tensors, GA_OP_F32_AVG, stream, variant);
}

int ga_allreduce_bucket(ga_ctx *c, size_t nbytes, int op, void *stream, int variant) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (nbytes > c->capacity_bytes) return fail(GA_ECAPACITY, "%zu bytes exceed the bucket capacity %zu", nbytes, c->capacity_bytes);
    if (nbytes % 4) return fail(GA_EINVAL, "nbytes must be a multiple of 4");
    const uint32_t tiles = (uint32_t)((nbytes + kTileBytes - 1) / kTileBytes);
    const int rc = ga_order_stream(c, stream);
    if (rc) return rc;
    return launch(c, nullptr, 0, tiles, op, stream, variant, true);
}

int ga_average_host(ga_ctx *c, void *const *host_ptrs, const int64_t *numel, int n_tensors,
                    void *stream, int variant) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (n_tensors < 0 || (n_tensors > 0 && (!host_ptrs || !numel))) return fail(GA_EINVAL, "bad tensor list");
    uint64_t total = 0;
    int rc = build_layout(numel, n_tensors, &total);
    if (rc) return rc;
    if (total * kTileBytes > c->capacity_bytes)
        return fail(GA_ECAPACITY, "gradient set needs %llu bucket bytes, capacity is %zu",
                    (unsigned long long)(total * kTileBytes), c->capacity_bytes);
    if (total == 0) return GA_OK;
    if (c->broken) return fail(GA_ESTATE, "context is unusable after an earlier timeout");
    ENTER_DEVICE(c->device);
    cudaStream_t st = (cudaStream_t)stream;
    rc = ga_order_stream(c, stream);
    if (rc) return rc;
    rc = drain_if_needed(c, stream);
    if (rc) return rc;
    const size_t bucket_bytes = (size_t)total * kTileBytes;
    // Chunk = 1/8 of the set, between 1 MiB and host_chunk_bytes (8 MiB): the pipeline's fill and drain
    // (first copy-in, last copy-out run alone) cost one chunk each, and every chunk costs a handful of
    // API calls.  Round-2 sweep at 64 MiB per rank (tools/e2e_chunk_sweep.py): 8 MiB chunks take 6.4 / 2.9 /
    // 3.9 ms at N = 1 (4 hosted) / 2 / 4 where the former 2 MiB chunks took 7.6 / 3.6 / 4.5.
    size_t chunk = (bucket_bytes / 8 + kTileBytes - 1) / kTileBytes * kTileBytes;
    if (chunk < ((size_t)1 << 20)) chunk = (size_t)1 << 20;
    if (chunk > c->host_chunk_bytes) chunk = c->host_chunk_bytes;
    if (c->host_chunk_set) chunk = c->host_chunk_bytes;        // GA_OPT_HOST_CHUNK_BYTES given: exactly that
    if ((bucket_bytes + chunk - 1) / chunk > (size_t)kHostChunks)
        chunk = ((bucket_bytes + kHostChunks - 1) / kHostChunks + kTileBytes - 1) / kTileBytes * kTileBytes;
    const int n_chunks = (int)((bucket_bytes + chunk - 1) / chunk);
    if (n_chunks <= 1) {
        // small set: one H2D pass, one kernel, one D2H pass, all on the caller's stream
        rc = copy_chunk(c, host_ptrs, numel, n_tensors, 0, bucket_bytes, true, st);
        if (rc) return rc;
        rc = launch(c, nullptr, 0, (uint32_t)total, GA_OP_F32_AVG, stream, variant, true);
        if (rc) return rc;
        return copy_chunk(c, host_ptrs, numel, n_tensors, 0, bucket_bytes, false, st);
    }
    // Large set: cut the bucket into chunks and run copy-in, reduction and copy-out of different
    // chunks concurrently (PCIe is full duplex), fork/join around the caller's stream.  AUTO is
    // resolved on the whole set so that every chunk uses the same variant.
    if (variant == GA_VARIANT_AUTO) variant = ga_pick_variant(c, bucket_bytes);
    rc = ensure_host_pipeline(c);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(c->ev_start, st));
    CUDA_TRY(cudaStreamWaitEvent(c->s_in, c->ev_start, 0));
    CUDA_TRY(cudaStreamWaitEvent(c->s_k, c->ev_start, 0));
    CUDA_TRY(cudaStreamWaitEvent(c->s_out, c->ev_start, 0));
    // (A tapered schedule -- 1, 2, 4 MiB chunks at both ends to shorten the pipeline's fill and drain -- was
    // measured and lost to uniform chunks, 6.9 against 6.5 ms: six more chunks of API calls and small copies.)
    for (int k = 0; k < n_chunks; ++k) {
        const size_t lo = (size_t)k * chunk, hi = lo + chunk < bucket_bytes ? lo + chunk : bucket_bytes;
        rc = copy_chunk(c, host_ptrs, numel, n_tensors, lo, hi, true, c->s_in);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(c->ev_in[k], c->s_in));
        CUDA_TRY(cudaStreamWaitEvent(c->s_k, c->ev_in[k], 0));
        rc = launch(c, nullptr, 0, (uint32_t)((hi - lo) / kTileBytes), GA_OP_F32_AVG, (void *)c->s_k, variant, true, lo);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(c->ev_k[k], c->s_k));
        CUDA_TRY(cudaStreamWaitEvent(c->s_out, c->ev_k[k], 0));
        rc = copy_chunk(c, host_ptrs, numel, n_tensors, lo, hi, false, c->s_out);
        if (rc) return rc;
    }
    CUDA_TRY(cudaEventRecord(c->ev_done, c->s_out));
    CUDA_TRY(cudaStreamWaitEvent(st, c->ev_done, 0));
    return GA_OK;
}

int ga_bucket_acquire(ga_ctx *c, void *stream) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (c->broken) return fail(GA_ESTATE, "context is unusable after an earlier timeout");
    const int rc = ga_order_stream(c, stream);
    if (rc) return rc;
    return drain_if_needed(c, stream);
}

void *ga_send_ptr(ga_ctx *c, int local_rank) {
    if (!c || local_rank < 0 || local_rank >= c->n_local) return nullptr;
    return c->send_all[c->rank_base + local_rank];
}

void *ga_recv_ptr(ga_ctx *c, int local_rank) {
    if (!c || local_rank < 0 || local_rank >= c->n_local) return nullptr;
    return c->recv[local_rank];
}

int64_t ga_bucket_offset(const int64_t *numel, int n_tensors, int i) {
    if (!numel || i < 0 || i > n_tensors) return -1;
    int64_t off = 0;
    for (int k = 0; k < i; ++k) off += (int64_t)tiles_for_elems(numel[k]) * kTileBytes;
    return off;
}

int ga_world(const ga_ctx *c) { return c ? c->world : 0; }
int ga_rank(const ga_ctx *c) { return c ? c->rank_base : -1; }
int ga_n_local(const ga_ctx *c) { return c ? c->n_local : 0; }
size_t ga_capacity(const ga_ctx *c) { return c ? c->capacity_bytes : 0; }

int ga_pick_variant(const ga_ctx *c, size_t nbytes) {
    if (!c) return GA_VARIANT_ONESHOT;
    const bool one_shot = nbytes <= c->oneshot_max_bytes;
    if (c->n_local > 1 && c->n_local == c->world) return GA_VARIANT_HOSTED_DIRECT;
    if (c->n_local > 1 || c->world == 1) return one_shot ? GA_VARIANT_ONESHOT : GA_VARIANT_TWOSHOT;
    int v;
    if (c->oneshot_max_set)      // the caller's threshold replaces the table's one-shot rows
        v = one_shot ? GA_VARIANT_ONESHOT_PUSH : (nbytes <= c->pipe_max_bytes ? GA_VARIANT_TWOSHOT_PIPE : GA_VARIANT_TWOSHOT_PUSH);
    else
        v = auto_from_table(c->world, nbytes);
    // one-shot push stages `world` copies of the set in every rank's bucket
    if (v == GA_VARIANT_ONESHOT_PUSH && nbytes * c->world > c->capacity_bytes) v = c->world <= 2 ? GA_VARIANT_TWOSHOT_PUSH : GA_VARIANT_TWOSHOT_PIPE;
    if (v == GA_VARIANT_TWOSHOT_PIPE && (nbytes > c->pipe_max_bytes || c->world <= 2)) v = GA_VARIANT_TWOSHOT_PUSH;
    return v;
}

int ga_auto_variant(int world, int n_local, size_t capacity_bytes, size_t nbytes) {
    if (world < 1 || world > kMaxWorld || n_local < 1 || n_local > world) return fail(GA_EINVAL, "bad world / n_local");
    ga_ctx c;
    memset(&c, 0, sizeof(c));
    c.world = world; c.n_local = n_local; c.capacity_bytes = capacity_bytes;
    set_auto_thresholds(&c);
    return ga_pick_variant(&c, nbytes);
}

uint64_t ga_launch_count(const ga_ctx *c) { return c ? c->launches : 0; }

int ga_last_grid(const ga_ctx *c, int *blocks, int *threads) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (blocks) *blocks = c->last_blocks;
    if (threads) *threads = c->last_threads;
    return GA_OK;
}

int ga_set_option(ga_ctx *c, int option, int64_t value) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    switch (option) {
        case GA_OPT_ONESHOT_MAX_BYTES: c->oneshot_max_bytes = value < 0 ? 0 : (size_t)value; c->oneshot_max_set = true; return GA_OK;
        case GA_OPT_TIMEOUT_MS: c->timeout_ns = value <= 0 ? 0 : (unsigned long long)value * 1000000ull; return GA_OK;
        case GA_OPT_MAX_BLOCKS: c->max_blocks_opt = value < 0 ? 0 : (int)value; return GA_OK;
        case GA_OPT_TRACE: {
            const size_t bytes = (size_t)kMaxBlocks * c->n_local * kTraceSlots * sizeof(unsigned long long);
            if (value && !c->trace_dev) {
                ENTER_DEVICE(c->device);
                CUDA_TRY(cudaMalloc((void **)&c->trace_dev, bytes));
                CUDA_TRY(cudaMemset(c->trace_dev, 0, bytes));
            } else if (!value && c->trace_dev) {
                ENTER_DEVICE(c->device);
                CUDA_TRY(cudaDeviceSynchronize());
                CUDA_TRY(cudaFree(c->trace_dev));
                c->trace_dev = nullptr;
            }
            return GA_OK;
        }
        case GA_OPT_HOST_CHUNK_BYTES:
            c->host_chunk_bytes = value < (int64_t)kTileBytes ? kTileBytes : ((size_t)value + kTileBytes - 1) / kTileBytes * kTileBytes;
            c->host_chunk_set = value > 0;
            if (value <= 0) c->host_chunk_bytes = (size_t)8 << 20;
            return GA_OK;
        case GA_OPT_PIPE_GROUP: c->pipe_group_opt = value < 0 ? 0 : (value > 64 ? 64 : (int)value); return GA_OK;
        case GA_OPT_VIEW_CHUNK_TILES: c->view_chunk_opt = value < 0 ? 0 : (value > 65536 ? 65536 : (int)value); return GA_OK;
        case GA_OPT_VIEW_ONESHOT_MAX_BYTES: c->view_oneshot_max_opt = value < 0 ? -1 : (long long)value; return GA_OK;
        case GA_OPT_VIEW_MIN_BYTES: c->view_min_bytes = value < 0 ? 0 : (size_t)value; return GA_OK;
    }
    return fail(GA_EINVAL, "unknown option %d", option);
}

int ga_synchronize(ga_ctx *c, void *stream) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    ENTER_DEVICE(c->device);
    CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    return ga_status(c);
}

int ga_read_trace(ga_ctx *c, uint64_t *out, size_t max_words, int *blocks, int *slots) {
    if (!c || !out) return fail(GA_EINVAL, "NULL argument");
    if (!c->trace_dev) return fail(GA_ESTATE, "tracing is off (ga_set_option(ctx, GA_OPT_TRACE, 1))");
    const size_t words = (size_t)c->last_blocks * c->n_local * kTraceSlots;
    if (words > max_words) return fail(GA_ECAPACITY, "trace needs %zu words, buffer has %zu", words, max_words);
    ENTER_DEVICE(c->device);
    CUDA_TRY(cudaDeviceSynchronize());
    CUDA_TRY(cudaMemcpy(out, c->trace_dev, words * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    if (blocks) *blocks = c->last_blocks;
    if (slots) *slots = kTraceSlots;
    return GA_OK;
}

int ga_status(ga_ctx *c) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    int s = *(volatile int *)c->status_host;
    if (s != 0) {
        c->broken = true;
        return fail(s, "a kernel timed out waiting for a peer rank's flag; the context is unusable");
    }
    return GA_OK;
}

int ga_close_peers(ga_ctx *c) {
    if (!c) return GA_OK;
    ENTER_DEVICE(c->device);
    CUDA_TRY(cudaDeviceSynchronize());
    for (int r = 0; r < c->world; ++r) {
        if (c->peer_base[r]) {
            CUDA_TRY(cudaIpcCloseMemHandle(c->peer_base[r]));
            c->peer_base[r] = nullptr;
            c->pad_all[r] = nullptr;
            c->send_all[r] = nullptr;
            c->gather_all[r] = nullptr;
        }
    }
    ga_view_release(c, false);
    if (c->n_local != c->world) c->opened = false;
    return GA_OK;
}

int ga_destroy(ga_ctx *c) {
    if (!c) return GA_OK;
    DeviceGuard device_guard_(c->device);
    cudaDeviceSynchronize();
    ga_view_release(c, true);
    if (c->order_event) cudaEventDestroy(c->order_event);
    for (int r = 0; r < c->world; ++r)
        if (c->peer_base[r]) cudaIpcCloseMemHandle(c->peer_base[r]);
    for (int l = 0; l < c->n_local; ++l) {
        if (c->base[l]) cudaFree(c->base[l]);
        if (c->recv[l]) cudaFree(c->recv[l]);
    }
    for (int i = 0; i < kSegRing; ++i) {
        if (c->seg_host[i]) cudaFreeHost(c->seg_host[i]);
        if (c->seg_dev[i]) cudaFree(c->seg_dev[i]);
        if (c->seg_event[i]) cudaEventDestroy(c->seg_event[i]);
    }
    if (c->status_host) cudaFreeHost(c->status_host);
    if (c->own_pads_dev) cudaFree(c->own_pads_dev);
    if (c->trace_dev) cudaFree(c->trace_dev);
    if (c->s_in) {
        cudaStreamDestroy(c->s_in); cudaStreamDestroy(c->s_k); cudaStreamDestroy(c->s_out);
        cudaEventDestroy(c->ev_start); cudaEventDestroy(c->ev_done);
        for (int i = 0; i < kHostChunks; ++i) { cudaEventDestroy(c->ev_in[i]); cudaEventDestroy(c->ev_k[i]); }
    }
    free(c);
    return GA_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// ga_average_static: the hosted-world average with NO allocation anywhere
// ------------------------------------------------------------------------------------------
// Some callers of this library are not allowed to allocate device memory at the moment they
// need a gradient average: code that runs while another component owns the allocator (a
// framework that has reserved "all" of HBM for its own pool and treats any foreign cudaMalloc as
// an out-of-memory risk), code inside a latency-critical section where the implicit device
// synchronisation of cudaMalloc/cudaFree is unacceptable, and plug-in hosts that forbid
// allocation after start-up.  ga_create cannot serve them: it cudaMallocs the bucket, the flag
// pad and a pinned status word.  This entry point keeps all of that in one file-scope device
// array instead.  The array is part of the library's device image: with CUDA's default lazy module
// loading the driver reserves its 8.53 MiB on a GPU the first time the library touches that GPU
// (the first call below), and the GPU then keeps it until the process exits.
namespace {

constexpr size_t kStaticBucketBytes = (size_t)8 << 20;    // bucket space shared by ALL hosted ranks
constexpr size_t kStaticPadsBytes = kMaxWorld * kPadBytes;
constexpr size_t kStaticBytes = kStaticPadsBytes + kStaticBucketBytes;      // 8 945 664 bytes
constexpr int kStaticMaxDevices = 64;
static_assert(sizeof(((Pad *)0)->reserved) >= sizeof(int), "the status word lives in the pad's spare words");

// In host code this name is only a handle for cudaGetSymbolAddress, never an address to pass on.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

struct StaticDevice {
    bool ready;
    unsigned char *scratch;      // device address of g_static_scratch on this device
    int sm_count;
    int blocks_per_sm[2][kMaxWorld + 1];
};
StaticDevice g_static_dev[kStaticMaxDevices];

unsigned long long static_timeout_ns() {
    static unsigned long long ns = [] {
        const char *env = getenv("GA_TIMEOUT_MS");           // the same knob the contexts read
        return env ? strtoull(env, nullptr, 10) * 1000000ull : 30ull * 1000000000ull;
    }();
    return ns;
}

// Once per device, with that device current: the array's address ON THIS DEVICE.
int static_setup(int device, StaticDevice **out) {
    StaticDevice &sd = g_static_dev[device];
    if (!sd.ready) {
        void *addr = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&addr, g_static_scratch));
        if (!addr) return fail(GA_ECUDA, "cudaGetSymbolAddress returned a null address");
        CUDA_TRY(cudaDeviceGetAttribute(&sd.sm_count, cudaDevAttrMultiProcessorCount, device));
        sd.scratch = (unsigned char *)addr;
        sd.ready = true;
    }
    *out = &sd;
    return GA_OK;
}

}  // namespace

extern "C" int ga_average_static(int world, int device, const void *const *grad_ptrs, const int64_t *numel,
                                 int n_tensors, void *stream, int variant) {
    if (world < 1 || world > kMaxWorld) return fail(GA_EINVAL, "world=%d outside 1..%d", world, kMaxWorld);
    if (device < 0 || device >= kStaticMaxDevices) return fail(GA_EINVAL, "device=%d outside 0..%d", device, kStaticMaxDevices - 1);
    if (n_tensors < 0 || (n_tensors > 0 && (!grad_ptrs || !numel))) return fail(GA_EINVAL, "bad tensor list");
    if ((size_t)n_tensors * world > (size_t)kInlineSegs)
        return fail(GA_ECAPACITY, "ga_average_static carries at most %d pointers (world * n_tensors = %zu): "
                    "a larger table would need a device allocation", kInlineSegs, (size_t)n_tensors * world);
    uint64_t total = 0;
    int rc = build_layout(numel, n_tensors, &total);
    if (rc) return rc;
    if (total * kTileBytes * world > kStaticBucketBytes)
        return fail(GA_ECAPACITY, "ga_average_static: %d ranks x %llu bucket bytes exceed the built-in %zu-byte scratch; "
                    "use ga_create_local + ga_average for larger gradient sets",
                    world, (unsigned long long)(total * kTileBytes), kStaticBucketBytes);
    if (total == 0) return GA_OK;
    if (variant == GA_VARIANT_AUTO)
        variant = (world <= 2 || total * kTileBytes <= ((size_t)512 << 10)) ? GA_VARIANT_ONESHOT : GA_VARIANT_TWOSHOT;
    if (variant != GA_VARIANT_ONESHOT && variant != GA_VARIANT_TWOSHOT) return fail(GA_EINVAL, "unknown variant %d", variant);

    ENTER_DEVICE(device);
    StaticDevice *sdp = nullptr;
    rc = static_setup(device, &sdp);
    if (rc) return rc;
    StaticDevice &sd = *sdp;
    cudaStream_t st = (cudaStream_t)stream;

    Params p;
    memset(&p, 0, sizeof(p));
    const size_t per_rank = (size_t)total * kTileBytes;
    for (int r = 0; r < world; ++r) {
        p.pad[r] = (Pad *)(sd.scratch + (size_t)r * kPadBytes);
        p.send[r] = sd.scratch + kStaticPadsBytes + (size_t)r * per_rank;
    }
    p.status = (int *)(sd.scratch + offsetof(Pad, reserved));      // rank 0's spare pad words: read back by ga_static_status
    p.timeout_ns = static_timeout_ns();
    p.n_seg = (uint32_t)n_tensors;
    p.total_tiles = (uint32_t)total;
    p.shard_tiles = ((uint32_t)total + world - 1) / world;
    p.rank_base = 0;
    p.direct = 0;
    for (int l = 0; l < world; ++l) {
        uint32_t start = 0;
        for (int i = 0; i < n_tensors; ++i) {
            Seg &s = p.inline_segs[(size_t)l * n_tensors + i];
            s.ptr = grad_ptrs[(size_t)l * n_tensors + i];
            s.start_tile = start;
            s.numel = (uint32_t)numel[i];
            if (numel[i] > 0 && !s.ptr) return fail(GA_EINVAL, "pointer %d of hosted rank %d is NULL", i, l);
            if (((uintptr_t)s.ptr & 3) != 0) return fail(GA_EINVAL, "pointer %d of hosted rank %d is not 4-byte aligned", i, l);
            start += (uint32_t)tiles_for_elems(numel[i]);
        }
    }

    KernelFn fn = pick_kernel(world, variant, GA_OP_F32_AVG);
    int &bps = sd.blocks_per_sm[variant == GA_VARIANT_TWOSHOT][world];
    if (bps == 0) {
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, (const void *)fn, kThreads, 0) != cudaSuccess || n < 1) n = 1;
        bps = n;
    }
    const uint32_t slots = variant == GA_VARIANT_TWOSHOT ? p.shard_tiles : p.total_tiles;
    const int wpb = kThreads / 32;
    int resident = bps * sd.sm_count / world;
    if (resident > kMaxBlocks) resident = kMaxBlocks;
    if (resident < 1) resident = 1;
    int blocks = (int)((slots + wpb - 1) / wpb);
    if (blocks > resident) blocks = resident;

    // The array still holds the previous call's flags.  What the protocol counts up from zero is
    // cleared on the stream before the launch, and only that: the flag rows of the blocks this launch
    // uses and the tail of each pad (done words, block counter, epoch, status) -- a few KB per rank in
    // three strided memsets, whatever the world size.
    unsigned char *pad0 = sd.scratch;
    const size_t rows = (size_t)blocks * kMaxWorld * sizeof(uint32_t);
    CUDA_TRY(cudaMemset2DAsync(pad0 + offsetof(Pad, ready), kPadBytes, 0, rows, (size_t)world, st));
    if (variant == GA_VARIANT_TWOSHOT)
        CUDA_TRY(cudaMemset2DAsync(pad0 + offsetof(Pad, mid), kPadBytes, 0, rows, (size_t)world, st));
    CUDA_TRY(cudaMemset2DAsync(pad0 + offsetof(Pad, done), kPadBytes, 0, sizeof(Pad) - offsetof(Pad, done), (size_t)world, st));

    dim3 grid(blocks, world), block(kThreads);
    void *args[] = {&p};
    if (world > 1) CUDA_TRY(cudaLaunchCooperativeKernel((const void *)fn, grid, block, args, 0, st));
    else CUDA_TRY(cudaLaunchKernel((const void *)fn, grid, block, args, 0, st));
    return GA_OK;
}

// Outcome of the ga_average_static calls enqueued on `stream` so far, for callers that cannot
// allocate either: waits for the stream, then copies the one status word out of the device array
// into a stack variable.  GA_OK, or GA_ETIMEOUT if a block of the last call gave up waiting for a flag
// (its gradients are then only partly averaged).
extern "C" int ga_static_status(int device, void *stream) {
    if (device < 0 || device >= kStaticMaxDevices) return fail(GA_EINVAL, "device=%d outside 0..%d", device, kStaticMaxDevices - 1);
    ENTER_DEVICE(device);
    StaticDevice *sd = nullptr;
    int rc = static_setup(device, &sd);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    int word = 0;
    CUDA_TRY(cudaMemcpy(&word, sd->scratch + offsetof(Pad, reserved), sizeof word, cudaMemcpyDeviceToHost));
    if (word != 0) return fail(word, "a ga_average_static kernel timed out waiting for a flag: its gradients are only partly averaged");
    return GA_OK;
}
