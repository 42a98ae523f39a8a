// gradavg_internal.cuh -- declarations shared by the translation units of libgradavg.so
// (gradavg.cu: bucket kernels + C ABI; gradavg_view.cu: gradient views, NVSwitch multicast,
// parameter broadcast).  Not installed, not part of the ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/gradavg.h"

#define GA_HIDDEN __attribute__((visibility("hidden")))

namespace {


// ------------------------------------------------------------------------------------------
// constants and device-visible structures
// ------------------------------------------------------------------------------------------
constexpr int kMaxWorld = GA_MAX_WORLD;
constexpr int kTileVecs = 128;                 // 16-byte vectors per warp tile
constexpr int kTileBytes = kTileVecs * 16;     // 2 KiB
constexpr int kTileElems = kTileVecs * 4;      // 512 32-bit elements
constexpr int kLaneVecs = kTileVecs / 32;      // 4 vectors per lane per tile
constexpr int kThreads = 512;                  // 16 warps per block
constexpr int kMaxBlocks = 1024;               // flag rows per rank
constexpr int kInlineSegs = 128;               // pointer-table entries carried in kernel params
constexpr int kSegRing = 4;                    // staging slots for larger pointer tables
constexpr int kTraceSlots = 16;                // phase timestamps per block when tracing is on
constexpr int kHostChunks = 64;                // most chunks one ga_average_host call is cut into

typedef uint4 Vec;

struct Seg {               // one gradient tensor of one rank
    const void *ptr;       // device pointer (re-read every call)
    uint32_t start_tile;   // first bucket tile of this tensor
    uint32_t numel;        // 32-bit elements
};

struct Pad {                                    // head of every rank's IPC allocation
    uint32_t ready[kMaxBlocks][kMaxWorld];      // [block][src rank]: src packed its tiles (epoch)
    uint32_t mid[kMaxBlocks][kMaxWorld];        // [block][src rank]: src reduced its shard (epoch)
    uint32_t done[kMaxWorld];                   // [src rank]: src finished reading my bucket (epoch)
    uint32_t finished_blocks;                   // local block counter of the running launch
    uint32_t epoch;                             // last flag value this rank used; read at kernel start, advanced
                                                // by the last block, so a launch carries no host-side counter
                                                // and can be replayed from a CUDA graph
    uint32_t last_kind;                         // 1 if this rank's previous launch was a double-buffered one-shot
                                                // push (its successor of the same kind need not wait for `done`)
    uint32_t reserved[21];
};
constexpr size_t kPadBytes = (sizeof(Pad) + 4095) / 4096 * 4096;

struct Params {
    void *send[kMaxWorld];      // every rank's bucket, by global rank (peer-mapped)
    Pad *pad[kMaxWorld];        // every rank's flag pad, by global rank (peer-mapped)
    void *recv[kMaxWorld];      // result regions of the hosted ranks, by local index
    void *gather[kMaxWorld];    // every rank's gather region, by global rank (push variants write into it)
    const Seg *dev_segs;        // [n_local][n_seg], or nullptr -> inline_segs
    int *status;                // host-mapped error word
    unsigned long long timeout_ns;
    uint32_t n_seg;
    uint32_t total_tiles;
    uint32_t shard_tiles;       // two-shot: tiles per rank shard
    int rank_base;
    int direct;                 // 1: bucket-direct (no pack, result -> recv)
    unsigned long long parity_stride;   // one-shot push: byte offset of the odd-epoch staging half (0 = single buffer)
    unsigned long long *trace;  // optional phase timestamps, kTraceSlots per block (GA_OPT_TRACE)
    uint32_t rounds;            // pipelined two-shot push: rounds per launch (flags use epoch .. epoch+rounds-1)
    uint32_t group;             //   and tile slots per block per round
    int root;                   // ga_bcast_kernel: the rank whose tensors are replicated
    Seg inline_segs[kInlineSegs];
};

// ------------------------------------------------------------------------------------------
// element operations
// ------------------------------------------------------------------------------------------
template <int OP> struct ElemOp;
template <> struct ElemOp<GA_OP_F32_AVG> {
    // __fadd_rn / __fdiv_rn: single IEEE binary32 operations, never contracted or replaced by
    // a reciprocal multiply (x/3.0f != x*(1/3.0f)).
    static __device__ __forceinline__ uint32_t add(uint32_t a, uint32_t b) {
        return __float_as_uint(__fadd_rn(__uint_as_float(a), __uint_as_float(b)));
    }
    static __device__ __forceinline__ uint32_t fin(uint32_t a, float w) {
        return __float_as_uint(__fdiv_rn(__uint_as_float(a), w));
    }
};
template <> struct ElemOp<GA_OP_F32_SUM> {
    static __device__ __forceinline__ uint32_t add(uint32_t a, uint32_t b) {
        return __float_as_uint(__fadd_rn(__uint_as_float(a), __uint_as_float(b)));
    }
    static __device__ __forceinline__ uint32_t fin(uint32_t a, float) { return a; }
};
template <> struct ElemOp<GA_OP_I32_SUM> {
    static __device__ __forceinline__ uint32_t add(uint32_t a, uint32_t b) { return a + b; }
    static __device__ __forceinline__ uint32_t fin(uint32_t a, float) { return a; }
};

// ------------------------------------------------------------------------------------------
// flags
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Phase timestamp of this block (thread 0), only when the context has tracing switched on.
#define GA_TRACE(p, k)                                                                                   \
    do {                                                                                                 \
        if ((p).trace && threadIdx.x == 0)                                                               \
            (p).trace[((size_t)blockIdx.y * gridDim.x + blockIdx.x) * kTraceSlots + (k)] = global_ns(); \
    } while (0)

// Spin until *flag has reached `target` (wrap-safe).  Returns false on timeout.
__device__ __forceinline__ bool spin_until(const uint32_t *flag, uint32_t target, unsigned long long timeout_ns) {
    unsigned long long t0 = 0;
    uint32_t spins = 0;
    while ((int32_t)(ld_relaxed_sys(flag) - target) < 0) {
        if ((++spins & 1023u) == 0 && timeout_ns) {
            unsigned long long now = global_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > timeout_ns) return false;
        }
    }
    return true;
}

// Block-wide wait for one flag per source rank.  ONE thread polls the W words in turn (they turn
// up within a round trip of each other) and issues ONE system-scope fence: with eight polling
// threads each needed its own fence, and system fences are what these barriers cost (more blocks
// made every variant slower until this was changed).  Returns false for the whole block on timeout.
template <int W, typename P>
__device__ __forceinline__ bool wait_flags(const uint32_t *row, uint32_t target, const P &p, int *s_abort) {
    if (threadIdx.x == 0) {
#pragma unroll 1
        for (int s = 0; s < W; ++s) {
            if (!spin_until(row + s, target, p.timeout_ns)) {
                *s_abort = 1;
                *(volatile int *)p.status = GA_ETIMEOUT;
                break;
            }
        }
        asm volatile("fence.acq_rel.sys;" ::: "memory");   // acquire: peers' data before the block's loads of it
    }
    __syncthreads();
    return *s_abort == 0;
}

__device__ __forceinline__ void st_relaxed_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Block-wide: everything this block wrote so far becomes visible (one system-scope fence by one
// thread, after the block barrier), then that thread posts the flag word into every rank's pad.
template <int W, typename P>
__device__ __forceinline__ void raise_flags(const P &p, size_t word_off, uint32_t value) {
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("fence.acq_rel.sys;" ::: "memory");
#pragma unroll
        for (int j = 0; j < W; ++j) st_relaxed_sys((uint32_t *)p.pad[j] + word_off, value);
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// host side shared by the translation units
// ------------------------------------------------------------------------------------------
GA_HIDDEN int ga_fail(int code, const char *fmt, ...);
#define fail ga_fail

#define CUDA_TRY(expr)                                                                         \
    do {                                                                                       \
        cudaError_t e_ = (expr);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            return fail(GA_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// Makes `device` current for the duration of an entry point and puts the caller's device back on
// the way out, so that a process driving several GPUs is not left on another device by a call.
struct DeviceGuard {
    int prev = -1;
    cudaError_t err;
    explicit DeviceGuard(int device) {
        err = cudaGetDevice(&prev);
        if (err == cudaSuccess && prev != device) err = cudaSetDevice(device);
        else if (err != cudaSuccess) { prev = -1; err = cudaSetDevice(device); }
        if (prev == device) prev = -1;          // nothing to restore
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
#define ENTER_DEVICE(dev)                                                                      \
    DeviceGuard device_guard_(dev);                                                            \
    if (device_guard_.err != cudaSuccess)                                                      \
        return fail(GA_ECUDA, "cudaSetDevice(%d) failed: %s", (dev), cudaGetErrorString(device_guard_.err))

struct ga_ctx {
    int world, rank_base, n_local, device;
    size_t capacity_bytes;          // per rank, multiple of kTileBytes
    size_t region_bytes;            // capacity + slack: two-shot push stages world * ceil(T / world) >= T tiles
    void *base[kMaxWorld];          // own allocations [Pad | bucket], by local index
    void *recv[kMaxWorld];          // own result regions, by local index
    void *peer_base[kMaxWorld];     // IPC-opened peer allocations, by global rank (nullptr = not opened)
    void *send_all[kMaxWorld];      // by global rank
    void *gather_all[kMaxWorld];    // by global rank (second half of the IPC allocation)
    Pad *pad_all[kMaxWorld];        // by global rank
    bool opened, broken;
    int sm_count, blocks_per_sm[16];                 // by variant; 0 = not queried yet
    int max_blocks_opt;
    unsigned long long *trace_dev;  // GA_OPT_TRACE: kMaxBlocks * n_local * kTraceSlots words, or nullptr
    int pipe_group_opt;             // GA_OPT_PIPE_GROUP: tile slots per block per round (0 = automatic)
    size_t oneshot_max_bytes;
    bool oneshot_max_set;           // GA_OPT_ONESHOT_MAX_BYTES was given: it replaces the table's one-shot rows
    size_t pipe_min_bytes;          // pipelined kernel: one round below min, 32 MiB rounds from min on; AUTO stops
    size_t pipe_max_bytes;          //   using it above max (N=4/8 sweeps: 3-10 % gain up to 128 MiB, a loss beyond)
    unsigned long long timeout_ns;
    int *status_host, *status_dev;
    Seg *seg_host[kSegRing], *seg_dev[kSegRing];
    size_t seg_cap;                 // entries per ring slot
    cudaEvent_t seg_event[kSegRing];
    bool seg_used[kSegRing];
    int seg_pos;
    uint64_t launches;
    int last_blocks, last_threads;
    bool last_was_tensor_call;      // previous launch checks its readers lazily (see ga_drain_kernel)
    const Pad **own_pads_dev;       // device array of the hosted ranks' pads, for ga_drain_kernel
    // ga_average_host pipeline: copy-in / kernel / copy-out streams and per-chunk events
    cudaStream_t s_in, s_k, s_out;
    cudaEvent_t ev_start, ev_done, ev_in[kHostChunks], ev_k[kHostChunks];
    size_t host_chunk_bytes;
    bool host_chunk_set;            // GA_OPT_HOST_CHUNK_BYTES was given: use exactly that chunk size
    // launches of one context must be ordered with respect to each other (the flag epoch and the
    // staging regions are per context): when the caller's stream changes, the new stream is made to
    // wait for the previous launch through this event (ga_order_stream)
    void *last_stream;
    bool have_last_stream;
    cudaEvent_t order_event;
    // gradient views (gradavg_view.cu): the gradients live in a peer-mapped region of their own
    int view_kind;                          // 0 none, 1 cudaMalloc + CUDA-IPC, 2 cuMem (VMM) + POSIX descriptors (+ multicast)
    size_t view_bytes;                      // usable bytes per rank (tile multiple)
    size_t view_map_bytes;                  // mapped bytes per rank (allocation granularity multiple)
    void *view_own[kMaxWorld];              // own view regions, by local index
    void *view_all[kMaxWorld];              // every rank's view region, by global rank (peer-mapped)
    void *view_peer_ipc[kMaxWorld];         // kind 1: IPC-opened peer mappings, by global rank
    unsigned long long view_mem[kMaxWorld]; // kind 2: allocation handles (own + imported), by global rank
    unsigned long long view_mc_handle;      // kind 2: the NVSwitch multicast object (0 = none)
    void *view_mc;                          // multicast mapping of all ranks' view regions (multimem.* address), or nullptr
    int view_fd, view_mc_fd;                // descriptors exported to the peers, closed by ga_view_bind
    bool view_opened, view_mc_added, view_mc_bound;
    int view_chunk_opt;                     // GA_OPT_VIEW_CHUNK_TILES (0 = automatic)
    long long view_oneshot_max_opt;         // GA_OPT_VIEW_ONESHOT_MAX_BYTES (-1 = automatic)
    size_t view_min_bytes;                  // GA_OPT_VIEW_MIN_BYTES: AUTO uses the in-place kernel from this size on
};

// stream ordering between consecutive launches of one context (gradavg.cu)
GA_HIDDEN int ga_order_stream(ga_ctx *c, void *stream);
// frees the view regions (gradavg_view.cu); called by ga_close_peers (peer mappings) and ga_destroy (everything)
GA_HIDDEN void ga_view_release(ga_ctx *c, bool own_too);
