// gradavg_view.cu -- gradients as views of a peer-mapped region, the NVSwitch
// multicast variant (8f-4), both in place on that region.
//
// The step after the hot path, optimizer.step() (/root/reference/mnist/main.py:96), reads p.grad;
// the step before it, zero_grad() + backward() (:91-92), writes it.  When every p.grad IS a slice of
// one flat region that the peers have mapped, the averaging of :122-127 needs neither pack nor
// unpack: ga_view_kernel scatters the region's shards straight into the owners' staging slots,
// the owner folds them in rank order and stores the result straight into every rank's region.
//
//   scatter blocks (even blockIdx.x)    chunk c of shard j: my region -> rank j's staging slot [me]   remote stores
//   fold blocks    (odd  blockIdx.x)    wait for chunk c from every rank, fold in rank order, divide,
//                                       store the result into EVERY rank's region                     remote stores
// Both kinds run at once, chunk flags instead of grid-wide phases, so the links always have stores
// queued in both directions (profiles/r02_nvlink_probe.md: stores are the fastest SM transport,
// 630-650 GB/s all-to-all; loads + stores mixed in one thread 557).
//
// ga_nvls_kernel is the opt-in in-switch form: multimem.ld_reduce of the owner's shard (the switch
// adds the W replicas), divide, multimem.st to all replicas.  Half the link bytes; the switch returns
// the correctly rounded exact sum, which is NOT the rank-order fold, so it has its own oracle and is
// never what AUTO picks.
#include "gradavg_internal.cuh"

#include <cuda.h>          // driver TYPES only: the functions are resolved through cudaGetDriverEntryPoint,
                           // so libgradavg.so has no link dependency on libcuda (it loads on a CPU-only box)
#include <errno.h>
#include <sys/syscall.h>
#include <unistd.h>

namespace {

struct ViewParams {
    void *view[kMaxWorld];      // every rank's gradient region, by global rank (peer-mapped)
    void *stage[kMaxWorld];     // every rank's staging area (the context's bucket), by global rank
    Pad *pad[kMaxWorld];
    void *mc;                   // multicast mapping of the regions (ga_nvls_kernel), or nullptr
    int *status;
    unsigned long long timeout_ns;
    unsigned long long *trace;
    uint32_t total_tiles;
    uint32_t shard_tiles;       // tiles per rank shard
    uint32_t chunk_tiles;       // tiles per chunk (one flag per chunk and source rank)
    uint32_t n_chunks;          // chunks per shard, <= kMaxBlocks
    int rank_base;
    int world;                  // ga_nvls_kernel: runtime world size
    int oneshot;                // ga_view_kernel: 1 = every rank receives every tile and folds everything locally
};

template <int OP, int N>
__device__ __forceinline__ Vec fold_vecs(const Vec (&v)[N], float wf) {
    Vec acc = v[0];
#pragma unroll
    for (int s = 1; s < N; ++s) {              // ((g0 + g1) + g2) + ...
        acc.x = ElemOp<OP>::add(acc.x, v[s].x);
        acc.y = ElemOp<OP>::add(acc.y, v[s].y);
        acc.z = ElemOp<OP>::add(acc.z, v[s].z);
        acc.w = ElemOp<OP>::add(acc.w, v[s].w);
    }
    acc.x = ElemOp<OP>::fin(acc.x, wf);
    acc.y = ElemOp<OP>::fin(acc.y, wf);
    acc.z = ElemOp<OP>::fin(acc.z, wf);
    acc.w = ElemOp<OP>::fin(acc.w, wf);
    return acc;
}

// phase timestamp written by whichever single thread calls it (GA_TRACE is thread 0 only)
template <typename P> __device__ __forceinline__ void trace_at(const P &p, int slot) {
    if (p.trace) p.trace[((size_t)blockIdx.y * gridDim.x + blockIdx.x) * kTraceSlots + slot] = global_ns();
}

constexpr uint32_t kRing = 16;     // chunk ordinals a block's data warps may run ahead of its control warp

__device__ __forceinline__ uint32_t ld_vol(const uint32_t *p) { return *(const volatile uint32_t *)p; }
__device__ __forceinline__ void st_vol(uint32_t *p, uint32_t v) { *(volatile uint32_t *)p = v; }

// Two warps of every block are CONTROL warps, one thread each: warp 1 is the PUBLISHER (waits until the
// data warps have issued a chunk, pays the system-scope fence, raises the chunk's flags in every
// rank's pad); in fold blocks warp 0 is the POLLER (waits for a chunk's flags from every rank, then
// releases it to the data warps).  The other warps are DATA warps that stream through the block's
// chunks and never wait for a fence: they tick a shared-memory counter per finished chunk.  Measured
// (profiles/r02_*): with the whole block behind one barrier per chunk, or with one control thread
// doing both jobs, a fence -- 40-50 us when the links are busy, it waits for the queue in front of it
// -- stood between a block and its next chunk, and more than one round per block made the kernel
// slower.  A fence by one thread does not slow the other warps' stores (nvlink_probe st16_fence).
//
// Trace slots (even blocks scatter, odd blocks fold): 0 start; 1 staging released by the previous
// launch's owners (scatter); 3 first chunk's flags raised (scatter) / first chunk arrived (fold);
// 5 a data warp has issued its last chunk; 6 last chunk arrived (fold); 2 last flags raised;
// 4 every peer's results seen (fold); 9 end.
template <int W, int OP>
__global__ void __launch_bounds__(kThreads, 2) ga_view_kernel(const __grid_constant__ ViewParams p) {
    __shared__ int s_abort;
    __shared__ int s_last;
    __shared__ uint32_t s_cnt[kRing];      // data warps that have finished chunk ordinal q (slot q % kRing)
    __shared__ uint32_t s_sig;             // ordinals whose flags the publisher has raised (frees ring slots)
    __shared__ uint32_t s_ready;           // fold blocks: ordinals whose inputs have arrived from every rank
    const int ly = blockIdx.y;
    const int r = p.rank_base + ly;
    const int lane = threadIdx.x & 31;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t T = p.total_tiles, Ts = p.shard_tiles, C = p.chunk_tiles, NC = p.n_chunks;
    const uint32_t k = blockIdx.x >> 1, nb = gridDim.x >> 1;     // index / count within my role
    const uint32_t nq = NC > k ? (NC - k + nb - 1) / nb : 0;      // my chunks: c = k + q * nb, q = 0 .. nq-1
    const bool scatter_role = (blockIdx.x & 1u) == 0;
    // scatter blocks: warp 1 publishes, warps 0, 2..15 move data; fold blocks: warp 0 polls, warp 1 publishes, 2..15 fold
    const uint32_t ndata = (blockDim.x >> 5) - (scatter_role ? 1u : 2u);
    const uint32_t dw = scatter_role ? (warp == 0 ? 0u : warp - 1) : warp - 2;   // data-warp index (control warps never use it)
    const float wf = (float)W;

    Pad *mypad = p.pad[r];
    const uint32_t epoch = ld_relaxed_sys(&mypad->epoch) + 1;    // same value in every block of this launch
    if (threadIdx.x < kRing) s_cnt[threadIdx.x] = 0;
    if (threadIdx.x == 0) { s_abort = 0; s_last = 0; s_sig = 0; s_ready = 0; }
    __syncthreads();
    GA_TRACE(p, 0);

    // data warp: this warp's part of chunk ordinal q is issued
    auto finish = [&](uint32_t q) {
        __syncwarp();
        if (lane == 0) { __threadfence_block(); atomicAdd(&s_cnt[q % kRing], 1u); }
    };
    // data warp: ring slot of ordinal q is free again
    auto ring_wait = [&](uint32_t q) {
        if (q >= kRing) {
            if (lane == 0) while (ld_vol(&s_sig) + kRing <= q && *(volatile int *)&s_abort == 0) {}
            __syncwarp();
        }
    };
    // publisher thread: as the data warps finish chunk ordinal q -> fence, raise `row`[c][r] in every pad
    auto publisher = [&](size_t row_off, bool raise) {
        for (uint32_t q = 0; q < nq; ++q) {
            while (ld_vol(&s_cnt[q % kRing]) < ndata) { if (*(volatile int *)&s_abort) return; }
            st_vol(&s_cnt[q % kRing], 0);
            if (raise) {
                asm volatile("fence.acq_rel.sys;" ::: "memory");
                const size_t off = row_off + (size_t)(k + q * nb) * kMaxWorld + r;
#pragma unroll
                for (int j = 0; j < W; ++j) st_relaxed_sys((uint32_t *)p.pad[j] + off, epoch);
            }
            st_vol(&s_sig, q + 1);
            if (q == 0 && scatter_role) trace_at(p, 3);
        }
        trace_at(p, 2);
    };

    if (scatter_role) {
        // the owners must have folded the previous launch's staging data before it is overwritten
        if (!wait_flags<W>(mypad->done, epoch - 1, p, &s_abort)) return;
        GA_TRACE(p, 1);
        if (warp == 1) {
            if (lane == 0) publisher(offsetof(Pad, ready) / 4, true);
        } else {
            constexpr uint32_t kPeers = W > 1 ? W - 1 : 1;
            for (uint32_t q = 0; q < nq; ++q) {
                ring_wait(q);
                const uint32_t i0 = (k + q * nb) * C;
                if (W > 1 && p.oneshot) {
                    // one-shot: tile t goes to EVERY peer, into the slot that peer keeps for me (T tiles per slot)
                    for (uint32_t it = dw; it < C; it += ndata) {
                        const uint32_t t = i0 + it;
                        if (t >= T) continue;
                        const Vec *src = (const Vec *)p.view[r] + (size_t)t * kTileVecs + lane;
                        Vec v[kLaneVecs];
#pragma unroll
                        for (int u = 0; u < kLaneVecs; ++u) v[u] = __ldcs(src + u * 32);
#pragma unroll 1
                        for (int jj = 1; jj < W; ++jj) {
                            const int j = (r + jj) % W;
                            Vec *dst = (Vec *)p.stage[j] + ((size_t)(r < j ? r : r - 1) * T + t) * kTileVecs + lane;
#pragma unroll
                            for (int u = 0; u < kLaneVecs; ++u) __stcg(dst + u * 32, v[u]);
                        }
                    }
                } else if (W > 1) {
                    for (uint32_t it = dw; it < C * kPeers; it += ndata) {
                        const uint32_t i = i0 + it / kPeers;
                        const int j = (r + 1 + (int)(it % kPeers)) % W;       // nearest rank first: W ranks, W different destinations
                        const uint32_t t = j * Ts + i;
                        if (i >= Ts || t >= T) continue;
                        const Vec *src = (const Vec *)p.view[r] + (size_t)t * kTileVecs + lane;
                        Vec v[kLaneVecs];
#pragma unroll
                        for (int u = 0; u < kLaneVecs; ++u) v[u] = __ldcs(src + u * 32);
                        Vec *dst = (Vec *)p.stage[j] + ((size_t)r * Ts + i) * kTileVecs + lane;
#pragma unroll
                        for (int u = 0; u < kLaneVecs; ++u) __stcg(dst + u * 32, v[u]);
                    }
                }
                finish(q);
            }
            if (warp == 0 && lane == 0) trace_at(p, 5);
        }
    } else {
        if (warp == 0) {
            if (lane == 0) {
                bool ok = true;
                for (uint32_t q = 0; q < nq && ok; ++q) {
                    const uint32_t *row = mypad->ready[k + q * nb];
#pragma unroll 1
                    for (int s = 0; s < W && ok; ++s) ok = spin_until(row + s, epoch, p.timeout_ns);
                    if (!ok) break;
                    asm volatile("fence.acq_rel.sys;" ::: "memory");     // acquire: the ranks' chunk before the data warps' loads
                    st_vol(&s_ready, q + 1);
                    if (q == 0) trace_at(p, 3);
                    if (q == nq - 1) trace_at(p, 6);
                }
                // every rank's results for my chunks have landed in my region (block k of every rank owns
                // the same chunks, so when all my blocks are through, all peers' blocks are); one-shot
                // results are local, nothing to wait for
                for (uint32_t q = 0; q < nq && ok && !p.oneshot; ++q) {
                    const uint32_t *row = mypad->mid[k + q * nb];
#pragma unroll 1
                    for (int s = 0; s < W && ok; ++s) ok = spin_until(row + s, epoch, p.timeout_ns);
                }
                if (!ok) {
                    *(volatile int *)p.status = GA_ETIMEOUT;
                    *(volatile int *)&s_abort = 1;
                } else {
                    asm volatile("fence.acq_rel.sys;" ::: "memory");
                }
                trace_at(p, 4);
            }
        } else if (warp == 1) {
            if (lane == 0) publisher(offsetof(Pad, mid) / 4, !p.oneshot);
        } else {
            const Vec *mystage = (const Vec *)p.stage[r];
            const bool one = p.oneshot != 0;
            for (uint32_t q = 0; q < nq; ++q) {
                if (lane == 0) while (ld_vol(&s_ready) <= q && *(volatile int *)&s_abort == 0) {}
                __syncwarp();
                if (*(volatile int *)&s_abort) break;
                __threadfence_block();
                ring_wait(q);
                const uint32_t i0 = (k + q * nb) * C;
                // items = (tile, group of NB 512-byte rows): W * NB vectors in flight per thread.  Small worlds
                // take more rows per item, or the fold is latency-bound on its local loads (2 x 16 B per thread
                // at world 2 could not feed the link)
                constexpr int NB = W <= 2 ? 4 : (W <= 4 ? 2 : 1);
                constexpr uint32_t kGroups = kLaneVecs / NB;
                for (uint32_t it = dw; it < C * kGroups; it += ndata) {
                    const uint32_t i = i0 + it / kGroups;
                    const uint32_t t = one ? i : r * Ts + i;                   // one-shot: chunks run over all T tiles
                    if ((!one && i >= Ts) || t >= T) continue;
                    const size_t row = (size_t)(it % kGroups) * NB * 32 + lane;
                    Vec v[NB][W];
#pragma unroll
                    for (int s = 0; s < W; ++s) {
                        const Vec *src = s == r ? (const Vec *)p.view[r] + (size_t)t * kTileVecs
                                         : one  ? mystage + ((size_t)(s < r ? s : s - 1) * T + t) * kTileVecs
                                                : mystage + ((size_t)s * Ts + i) * kTileVecs;
#pragma unroll
                        for (int u = 0; u < NB; ++u) v[u][s] = __ldcg(src + row + u * 32);
                    }
#pragma unroll
                    for (int u = 0; u < NB; ++u) {
                        const Vec acc = fold_vecs<OP, W>(v[u], wf);
                        if (one) {
                            __stcg((Vec *)p.view[r] + (size_t)t * kTileVecs + row + u * 32, acc);
                        } else {
#pragma unroll
                            for (int jj = 0; jj < W; ++jj) {
                                const int j = (r + jj) % W;
                                __stcg((Vec *)p.view[j] + (size_t)t * kTileVecs + row + u * 32, acc);
                            }
                        }
                    }
                }
                finish(q);
            }
            if (warp == 2 && lane == 0) trace_at(p, 5);
        }
    }

    __syncthreads();
    if (s_abort) return;
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&mypad->finished_blocks, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        if (threadIdx.x == 0) {
            mypad->finished_blocks = 0; mypad->epoch = epoch; mypad->last_kind = 0;
            asm volatile("fence.acq_rel.sys;" ::: "memory");
#pragma unroll
            for (int j = 0; j < W; ++j) st_relaxed_sys((uint32_t *)p.pad[j] + offsetof(Pad, done) / 4 + r, epoch);
        }
    }
    GA_TRACE(p, 9);
}

// ---- NVSwitch multicast ------------------------------------------------------------------------
template <int OP> struct McOp;
template <> struct McOp<GA_OP_F32_AVG> {
    static __device__ __forceinline__ Vec ld(const Vec *a) {
        float x, y, z, w;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "l"(a) : "memory");
        return make_uint4(__float_as_uint(x), __float_as_uint(y), __float_as_uint(z), __float_as_uint(w));
    }
};
template <> struct McOp<GA_OP_F32_SUM> : McOp<GA_OP_F32_AVG> {};
template <> struct McOp<GA_OP_I32_SUM> {
    static __device__ __forceinline__ Vec ld(const Vec *a) {
        Vec v;
        const uint32_t *q = (const uint32_t *)a;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];" : "=r"(v.x) : "l"(q + 0) : "memory");
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];" : "=r"(v.y) : "l"(q + 1) : "memory");
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];" : "=r"(v.z) : "l"(q + 2) : "memory");
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];" : "=r"(v.w) : "l"(q + 3) : "memory");
        return v;
    }
};
__device__ __forceinline__ void mc_st(Vec *a, const Vec &v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%4], {%0,%1,%2,%3};"
                 ::"f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w)), "l"(a) : "memory");
}

// runtime-W forms of the block barriers (the NVLS kernel's only use of W is the divisor)
__device__ __forceinline__ bool wait_flags_n(const uint32_t *row, int W, uint32_t target, const ViewParams &p, int *s_abort) {
    if (threadIdx.x == 0) {
        for (int s = 0; s < W; ++s) {
            if (!spin_until(row + s, target, p.timeout_ns)) {
                *s_abort = 1;
                *(volatile int *)p.status = GA_ETIMEOUT;
                break;
            }
        }
        asm volatile("fence.acq_rel.sys;" ::: "memory");
    }
    __syncthreads();
    return *s_abort == 0;
}
__device__ __forceinline__ void raise_flags_n(const ViewParams &p, int W, size_t word_off, uint32_t value) {
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("fence.acq_rel.sys;" ::: "memory");
        for (int j = 0; j < W; ++j) st_relaxed_sys((uint32_t *)p.pad[j] + word_off, value);
    }
}

// One rank per GPU only (a multicast object has one member per device).  Block b of every rank:
// barrier "my region is final" -> in-switch reduce of my part of my shard -> divide -> multicast
// store -> barrier "my stores have landed everywhere".
template <int OP>
__global__ void __launch_bounds__(kThreads, 2) ga_nvls_kernel(const __grid_constant__ ViewParams p) {
    __shared__ int s_abort;
    __shared__ int s_last;
    const int r = p.rank_base, W = p.world;
    const int b = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const uint32_t warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const uint32_t T = p.total_tiles, Ts = p.shard_tiles;
    const float wf = (float)W;
    Pad *mypad = p.pad[r];
    const uint32_t epoch = ld_relaxed_sys(&mypad->epoch) + 1;
    if (threadIdx.x == 0) { s_abort = 0; s_last = 0; }
    __syncthreads();
    GA_TRACE(p, 0);
    raise_flags_n(p, W, offsetof(Pad, ready) / 4 + (size_t)b * kMaxWorld + r, epoch);
    if (!wait_flags_n(mypad->ready[b], W, epoch, p, &s_abort)) return;
    GA_TRACE(p, 1);
    Vec *mc = (Vec *)p.mc;
    for (uint32_t i = b * wpb + warp; i < Ts; i += gridDim.x * wpb) {
        const uint32_t t = r * Ts + i;
        if (t >= T) break;
        Vec *a = mc + (size_t)t * kTileVecs + lane;
        Vec v[kLaneVecs];
#pragma unroll
        for (int u = 0; u < kLaneVecs; ++u) v[u] = McOp<OP>::ld(a + u * 32);
#pragma unroll
        for (int u = 0; u < kLaneVecs; ++u) {
            v[u].x = ElemOp<OP>::fin(v[u].x, wf);
            v[u].y = ElemOp<OP>::fin(v[u].y, wf);
            v[u].z = ElemOp<OP>::fin(v[u].z, wf);
            v[u].w = ElemOp<OP>::fin(v[u].w, wf);
            mc_st(a + u * 32, v[u]);
        }
    }
    GA_TRACE(p, 2);
    raise_flags_n(p, W, offsetof(Pad, mid) / 4 + (size_t)b * kMaxWorld + r, epoch);
    if (!wait_flags_n(mypad->mid[b], W, epoch, p, &s_abort)) return;
    GA_TRACE(p, 4);
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&mypad->finished_blocks, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
        mypad->finished_blocks = 0; mypad->epoch = epoch; mypad->last_kind = 0;
        asm volatile("fence.acq_rel.sys;" ::: "memory");
        for (int j = 0; j < W; ++j) st_relaxed_sys((uint32_t *)p.pad[j] + offsetof(Pad, done) / 4 + r, epoch);
    }
    GA_TRACE(p, 9);
}

typedef void (*ViewFn)(ViewParams);
template <int W> ViewFn pick_view_op(int op) {
    return op == GA_OP_I32_SUM   ? (ViewFn)ga_view_kernel<W, GA_OP_I32_SUM>
           : op == GA_OP_F32_SUM ? (ViewFn)ga_view_kernel<W, GA_OP_F32_SUM>
                                 : (ViewFn)ga_view_kernel<W, GA_OP_F32_AVG>;
}
ViewFn pick_view(int world, int op) {
    switch (world) {
        case 1: return pick_view_op<1>(op);
        case 2: return pick_view_op<2>(op);
        case 3: return pick_view_op<3>(op);
        case 4: return pick_view_op<4>(op);
        case 5: return pick_view_op<5>(op);
        case 6: return pick_view_op<6>(op);
        case 7: return pick_view_op<7>(op);
        case 8: return pick_view_op<8>(op);
    }
    return nullptr;
}
ViewFn pick_nvls(int op) {
    return op == GA_OP_I32_SUM   ? (ViewFn)ga_nvls_kernel<GA_OP_I32_SUM>
           : op == GA_OP_F32_SUM ? (ViewFn)ga_nvls_kernel<GA_OP_F32_SUM>
                                 : (ViewFn)ga_nvls_kernel<GA_OP_F32_AVG>;
}

// ---- driver entry points (cuMem* / cuMulticast*), resolved at first use ---------------------------
struct Driver {
    bool tried, ok;
    decltype(&cuDeviceGet) DeviceGet;
    decltype(&cuDeviceGetAttribute) DeviceGetAttribute;
    decltype(&cuGetErrorString) GetErrorString;
    decltype(&cuMemGetAllocationGranularity) MemGetAllocationGranularity;
    decltype(&cuMemCreate) MemCreate;
    decltype(&cuMemRelease) MemRelease;
    decltype(&cuMemAddressReserve) MemAddressReserve;
    decltype(&cuMemAddressFree) MemAddressFree;
    decltype(&cuMemMap) MemMap;
    decltype(&cuMemUnmap) MemUnmap;
    decltype(&cuMemSetAccess) MemSetAccess;
    decltype(&cuMemExportToShareableHandle) MemExportToShareableHandle;
    decltype(&cuMemImportFromShareableHandle) MemImportFromShareableHandle;
    decltype(&cuMulticastCreate) MulticastCreate;
    decltype(&cuMulticastAddDevice) MulticastAddDevice;
    decltype(&cuMulticastBindMem) MulticastBindMem;
    decltype(&cuMulticastUnbind) MulticastUnbind;
    decltype(&cuMulticastGetGranularity) MulticastGetGranularity;
};
Driver g_drv;

template <typename F> bool resolve(const char *name, F *out) {
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
        cudaGetLastError();
        return false;
    }
    *out = (F)fn;
    return true;
}

bool driver_ready() {
    Driver &d = g_drv;
    if (d.tried) return d.ok;
    d.tried = true;
    d.ok = resolve("cuDeviceGet", &d.DeviceGet) && resolve("cuDeviceGetAttribute", &d.DeviceGetAttribute) &&
           resolve("cuGetErrorString", &d.GetErrorString) &&
           resolve("cuMemGetAllocationGranularity", &d.MemGetAllocationGranularity) && resolve("cuMemCreate", &d.MemCreate) &&
           resolve("cuMemRelease", &d.MemRelease) && resolve("cuMemAddressReserve", &d.MemAddressReserve) &&
           resolve("cuMemAddressFree", &d.MemAddressFree) && resolve("cuMemMap", &d.MemMap) && resolve("cuMemUnmap", &d.MemUnmap) &&
           resolve("cuMemSetAccess", &d.MemSetAccess) && resolve("cuMemExportToShareableHandle", &d.MemExportToShareableHandle) &&
           resolve("cuMemImportFromShareableHandle", &d.MemImportFromShareableHandle) &&
           resolve("cuMulticastCreate", &d.MulticastCreate) && resolve("cuMulticastAddDevice", &d.MulticastAddDevice) &&
           resolve("cuMulticastBindMem", &d.MulticastBindMem) && resolve("cuMulticastUnbind", &d.MulticastUnbind) &&
           resolve("cuMulticastGetGranularity", &d.MulticastGetGranularity);
    return d.ok;
}

#define CU_TRY(expr)                                                                                    \
    do {                                                                                                \
        CUresult e_ = (expr);                                                                           \
        if (e_ != CUDA_SUCCESS) {                                                                       \
            const char *s_ = nullptr;                                                                   \
            g_drv.GetErrorString(e_, &s_);                                                              \
            return fail(GA_ECUDA, "%s failed: %s (%s:%d)", #expr, s_ ? s_ : "?", __FILE__, __LINE__);   \
        }                                                                                               \
    } while (0)

// What travels between the ranks for a view region (GA_VIEW_HANDLE_BYTES = 128 per rank).
struct ViewHandle {
    unsigned char ipc[64];      // kind 1: cudaIpcMemHandle_t
    int32_t kind;               // 1 or 2
    int32_t pid;                // kind 2: exporter's process id
    int32_t mem_fd;             // kind 2: descriptor of the exporter's allocation, in the exporter's table
    int32_t mc_fd;              // kind 2, rank 0 only: descriptor of the multicast object (-1 = none)
    uint64_t map_bytes;
    uint64_t view_bytes;
    unsigned char pad[32];
};
static_assert(sizeof(ViewHandle) == GA_VIEW_HANDLE_BYTES, "view handle size");

// Duplicates descriptor `fd` of process `pid` into this process (Linux >= 5.6; same user).
int fetch_fd(int pid, int fd) {
    if (getenv("GA_TEST_FAIL_FETCH_FD")) { errno = EPERM; return -1; }     // test hook: a box where pidfd_getfd is not permitted
    if (pid == (int)getpid()) return dup(fd);
    const int pidfd = (int)syscall(SYS_pidfd_open, pid, 0);
    if (pidfd < 0) return -1;
    const int got = (int)syscall(SYS_pidfd_getfd, pidfd, fd, 0);
    const int e = errno;
    close(pidfd);
    errno = e;
    return got;
}

size_t round_up(size_t x, size_t g) { return (x + g - 1) / g * g; }

int multicast_supported(int device) {
    if (!driver_ready()) return 0;
    CUdevice dev;
    int sup = 0, posix = 0;
    if (g_drv.DeviceGet(&dev, device) != CUDA_SUCCESS) return 0;
    g_drv.DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
    g_drv.DeviceGetAttribute(&posix, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
    return sup && posix;
}

int map_handle(ga_ctx *c, CUmemGenericAllocationHandle h, size_t bytes, size_t gran, void **out) {
    CUdeviceptr va = 0;
    CU_TRY(g_drv.MemAddressReserve(&va, bytes, gran, 0, 0));
    CUresult e = g_drv.MemMap(va, bytes, 0, h, 0);
    if (e == CUDA_SUCCESS) {
        CUmemAccessDesc ad;
        memset(&ad, 0, sizeof ad);
        ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ad.location.id = c->device;
        ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        e = g_drv.MemSetAccess(va, bytes, &ad, 1);
        if (e != CUDA_SUCCESS) g_drv.MemUnmap(va, bytes);
    }
    if (e != CUDA_SUCCESS) {
        g_drv.MemAddressFree(va, bytes);
        const char *s = nullptr;
        g_drv.GetErrorString(e, &s);
        return fail(GA_ECUDA, "mapping a view region failed: %s", s ? s : "?");
    }
    *out = (void *)va;
    return GA_OK;
}

void unmap_va(void *va, size_t bytes) {
    if (!va) return;
    g_drv.MemUnmap((CUdeviceptr)va, bytes);
    g_drv.MemAddressFree((CUdeviceptr)va, bytes);
}

constexpr size_t kViewGranularity = (size_t)2 << 20;    // cuMem + multicast minimum on B200 (checked at create)

}  // namespace

void ga_view_release(ga_ctx *c, bool own_too) {
    if (!c || c->view_kind == 0) return;
    if (c->view_kind == 1) {
        for (int r = 0; r < c->world; ++r)
            if (c->view_peer_ipc[r]) { cudaIpcCloseMemHandle(c->view_peer_ipc[r]); c->view_peer_ipc[r] = nullptr; c->view_all[r] = nullptr; }
        if (own_too)
            for (int l = 0; l < c->n_local; ++l)
                if (c->view_own[l]) { cudaFree(c->view_own[l]); c->view_own[l] = nullptr; c->view_all[c->rank_base + l] = nullptr; }
    } else {
        const int me = c->rank_base;
        if (c->view_mc) { unmap_va(c->view_mc, c->view_map_bytes); c->view_mc = nullptr; }
        if (c->view_mc_handle) {
            if (c->view_mc_bound) {
                CUdevice dev;
                if (g_drv.DeviceGet(&dev, c->device) == CUDA_SUCCESS) g_drv.MulticastUnbind(c->view_mc_handle, dev, 0, c->view_map_bytes);
                c->view_mc_bound = false;
            }
            g_drv.MemRelease(c->view_mc_handle);
            c->view_mc_handle = 0;
            c->view_mc_added = false;
        }
        for (int r = 0; r < c->world; ++r) {
            if (r == me) continue;
            if (c->view_all[r]) { unmap_va(c->view_all[r], c->view_map_bytes); c->view_all[r] = nullptr; }
            if (c->view_mem[r]) { g_drv.MemRelease(c->view_mem[r]); c->view_mem[r] = 0; }
        }
        if (c->view_fd > 0) { close(c->view_fd); c->view_fd = 0; }
        if (c->view_mc_fd > 0) { close(c->view_mc_fd); c->view_mc_fd = 0; }
        if (own_too) {
            if (c->view_own[0]) { unmap_va(c->view_own[0], c->view_map_bytes); c->view_own[0] = nullptr; c->view_all[me] = nullptr; }
            if (c->view_mem[me]) { g_drv.MemRelease(c->view_mem[me]); c->view_mem[me] = 0; }
        }
    }
    c->view_opened = false;
    if (own_too) { c->view_kind = 0; c->view_bytes = 0; c->view_map_bytes = 0; }
}

extern "C" {

int ga_multicast_supported(int device) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) { cudaGetLastError(); return 0; }
    return multicast_supported(device);
}

int ga_view_create(ga_ctx *c, size_t nbytes, int flags) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (c->view_kind != 0) return fail(GA_ESTATE, "this context already has a view region");
    if (nbytes == 0) return fail(GA_EINVAL, "nbytes is 0");
    const size_t want = round_up(nbytes, kTileBytes);
    if (want > c->capacity_bytes)
        return fail(GA_ECAPACITY, "a %zu-byte view region needs a context of at least that capacity (the bucket is its "
                    "staging area), this one has %zu", want, c->capacity_bytes);
    ENTER_DEVICE(c->device);
    const bool vmm = (flags & GA_VIEW_MULTICAST) && c->n_local == 1 && c->world > 1 && multicast_supported(c->device);
    if (!vmm) {
        for (int l = 0; l < c->n_local; ++l) {
            cudaError_t e = cudaMalloc(&c->view_own[l], want);
            if (e == cudaSuccess) e = cudaMemset(c->view_own[l], 0, want);
            if (e != cudaSuccess) {
                for (int k = 0; k <= l; ++k) { if (c->view_own[k]) cudaFree(c->view_own[k]); c->view_own[k] = nullptr; }
                return fail(GA_ECUDA, "allocating the view region failed: %s", cudaGetErrorString(e));
            }
            c->view_all[c->rank_base + l] = c->view_own[l];
        }
        c->view_kind = 1;
        c->view_bytes = want;
        c->view_map_bytes = want;
        c->view_opened = (c->n_local == c->world);
        return GA_OK;
    }
    // cuMem allocation that can be exported as a POSIX descriptor and bound to a multicast object
    const int me = c->rank_base;
    CUmemAllocationProp ap;
    memset(&ap, 0, sizeof ap);
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = c->device;
    ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t gran = 0;
    CU_TRY(g_drv.MemGetAllocationGranularity(&gran, &ap, CU_MEM_ALLOC_GRANULARITY_MINIMUM));
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof mp);
    mp.numDevices = (unsigned)c->world;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    mp.size = round_up(want, kViewGranularity);
    size_t mgran = 0;
    CU_TRY(g_drv.MulticastGetGranularity(&mgran, &mp, CU_MULTICAST_GRANULARITY_MINIMUM));
    if (mgran > gran) gran = mgran;
    const size_t map_bytes = round_up(want, gran);
    CUmemGenericAllocationHandle h = 0;
    CU_TRY(g_drv.MemCreate(&h, map_bytes, &ap, 0));
    c->view_mem[me] = h;
    c->view_kind = 2;
    c->view_bytes = want;
    c->view_map_bytes = map_bytes;
    int rc = map_handle(c, h, map_bytes, gran, &c->view_own[0]);
    if (rc) { ga_view_release(c, true); return rc; }
    c->view_all[me] = c->view_own[0];
    cudaError_t e = cudaMemset(c->view_own[0], 0, map_bytes);
    if (e != cudaSuccess) { ga_view_release(c, true); return fail(GA_ECUDA, "clearing the view region failed: %s", cudaGetErrorString(e)); }
    int fd = -1;
    CUresult ce = g_drv.MemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (ce != CUDA_SUCCESS || fd <= 0) { ga_view_release(c, true); return fail(GA_ECUDA, "cuMemExportToShareableHandle failed (%d)", (int)ce); }
    c->view_fd = fd;
    if (me == 0) {
        mp.size = map_bytes;
        CUmemGenericAllocationHandle mc = 0;
        ce = g_drv.MulticastCreate(&mc, &mp);
        if (ce != CUDA_SUCCESS) { ga_view_release(c, true); return fail(GA_ECUDA, "cuMulticastCreate failed (%d)", (int)ce); }
        c->view_mc_handle = mc;
        int mfd = -1;
        ce = g_drv.MemExportToShareableHandle(&mfd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
        if (ce != CUDA_SUCCESS || mfd <= 0) { ga_view_release(c, true); return fail(GA_ECUDA, "exporting the multicast object failed (%d)", (int)ce); }
        c->view_mc_fd = mfd;
    }
    return GA_OK;
}

int ga_view_export(ga_ctx *c, unsigned char out[GA_VIEW_HANDLE_BYTES]) {
    if (!c || !out) return fail(GA_EINVAL, "NULL argument");
    if (c->view_kind == 0) return fail(GA_ESTATE, "no view region (ga_view_create)");
    if (c->n_local != 1) return fail(GA_ESTATE, "a local world has no handles to exchange");
    ViewHandle h;
    memset(&h, 0, sizeof h);
    h.kind = c->view_kind;
    h.pid = (int32_t)getpid();
    h.mem_fd = -1;
    h.mc_fd = -1;
    h.map_bytes = c->view_map_bytes;
    h.view_bytes = c->view_bytes;
    if (c->view_kind == 1) {
        ENTER_DEVICE(c->device);
        cudaIpcMemHandle_t ih;
        CUDA_TRY(cudaIpcGetMemHandle(&ih, c->view_own[0]));
        memcpy(h.ipc, &ih, sizeof ih);
    } else {
        h.mem_fd = c->view_fd;
        h.mc_fd = c->view_mc_fd > 0 ? c->view_mc_fd : -1;
    }
    memcpy(out, &h, sizeof h);
    return GA_OK;
}

int ga_view_open(ga_ctx *c, const unsigned char *handles) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (c->view_kind == 0) return fail(GA_ESTATE, "no view region (ga_view_create)");
    if (c->view_opened) return c->n_local == c->world ? GA_OK : fail(GA_ESTATE, "view regions already opened");
    if (!handles) return fail(GA_EINVAL, "NULL handles");
    ENTER_DEVICE(c->device);
    const int me = c->rank_base;
    const ViewHandle *hs = (const ViewHandle *)handles;
    for (int r = 0; r < c->world; ++r) {
        ViewHandle h;
        memcpy(&h, hs + r, sizeof h);
        if (h.kind != c->view_kind || h.view_bytes != c->view_bytes || h.map_bytes != c->view_map_bytes)
            return fail(GA_ESTATE, "rank %d created a different view region (kind %d, %llu bytes; mine: kind %d, %zu bytes): "
                        "every rank must call ga_view_create with the same arguments on the same hardware",
                        r, h.kind, (unsigned long long)h.view_bytes, c->view_kind, c->view_bytes);
    }
    int rc = GA_OK;
    if (c->view_kind == 1) {
        for (int r = 0; r < c->world && rc == GA_OK; ++r) {
            if (r == me) continue;
            cudaIpcMemHandle_t ih;
            memcpy(&ih, hs[r].ipc, sizeof ih);
            void *ptr = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&ptr, ih, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) { rc = fail(GA_ECUDA, "cudaIpcOpenMemHandle(view of rank %d) failed: %s", r, cudaGetErrorString(e)); break; }
            c->view_peer_ipc[r] = ptr;
            c->view_all[r] = ptr;
        }
    } else {
        for (int r = 0; r < c->world && rc == GA_OK; ++r) {
            if (r == me) continue;
            const int fd = fetch_fd(hs[r].pid, hs[r].mem_fd);
            if (fd < 0) { rc = fail(GA_ECUDA, "pidfd_getfd(rank %d, pid %d, fd %d) failed: %s", r, hs[r].pid, hs[r].mem_fd, strerror(errno)); break; }
            CUmemGenericAllocationHandle h = 0;
            CUresult ce = g_drv.MemImportFromShareableHandle(&h, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
            close(fd);
            if (ce != CUDA_SUCCESS) { rc = fail(GA_ECUDA, "cuMemImportFromShareableHandle(rank %d) failed (%d)", r, (int)ce); break; }
            c->view_mem[r] = h;
            rc = map_handle(c, h, c->view_map_bytes, kViewGranularity, &c->view_all[r]);
        }
        if (rc == GA_OK && hs[0].mc_fd > 0) {
            if (me != 0) {
                const int fd = fetch_fd(hs[0].pid, hs[0].mc_fd);
                if (fd < 0) rc = fail(GA_ECUDA, "pidfd_getfd(multicast object of rank 0) failed: %s", strerror(errno));
                else {
                    CUmemGenericAllocationHandle mc = 0;
                    CUresult ce = g_drv.MemImportFromShareableHandle(&mc, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
                    close(fd);
                    if (ce != CUDA_SUCCESS) rc = fail(GA_ECUDA, "importing the multicast object failed (%d)", (int)ce);
                    else c->view_mc_handle = mc;
                }
            }
            if (rc == GA_OK) {
                CUdevice dev;
                CUresult ce = g_drv.DeviceGet(&dev, c->device);
                if (ce == CUDA_SUCCESS) ce = g_drv.MulticastAddDevice(c->view_mc_handle, dev);
                if (ce != CUDA_SUCCESS) rc = fail(GA_ECUDA, "cuMulticastAddDevice failed (%d)", (int)ce);
                else c->view_mc_added = true;
            }
        }
    }
    if (rc != GA_OK) { ga_view_release(c, false); return rc; }
    c->view_opened = true;
    return GA_OK;
}

int ga_view_bind(ga_ctx *c) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (c->view_kind != 2) return GA_OK;                      // nothing to bind without a multicast object
    if (!c->view_opened) return fail(GA_ESTATE, "ga_view_open has not been called");
    ENTER_DEVICE(c->device);
    // every peer has fetched my descriptors by now (the caller's barrier between open and bind)
    if (c->view_fd > 0) { close(c->view_fd); c->view_fd = 0; }
    if (c->view_mc_fd > 0) { close(c->view_mc_fd); c->view_mc_fd = 0; }
    if (!c->view_mc_handle || !c->view_mc_added || c->view_mc) return GA_OK;
    CU_TRY(g_drv.MulticastBindMem(c->view_mc_handle, 0, c->view_mem[c->rank_base], 0, c->view_map_bytes, 0));
    c->view_mc_bound = true;
    return map_handle(c, c->view_mc_handle, c->view_map_bytes, kViewGranularity, &c->view_mc);
}

int ga_view_destroy(ga_ctx *c) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (c->view_kind == 0) return GA_OK;
    ENTER_DEVICE(c->device);
    CUDA_TRY(cudaDeviceSynchronize());
    ga_view_release(c, true);
    return GA_OK;
}

void *ga_view_ptr(ga_ctx *c, int local_rank) {
    if (!c || local_rank < 0 || local_rank >= c->n_local) return nullptr;
    return c->view_own[local_rank];
}

size_t ga_view_bytes(const ga_ctx *c) { return c ? c->view_bytes : 0; }

int ga_view_has_multicast(const ga_ctx *c) { return c && c->view_mc ? 1 : 0; }

int ga_average_view(ga_ctx *c, size_t nbytes, int op, void *stream, int variant) {
    if (!c) return fail(GA_EINVAL, "NULL context");
    if (c->view_kind == 0 || !c->view_opened) return fail(GA_ESTATE, "no opened view region (ga_view_create / ga_view_open)");
    if (c->broken) return fail(GA_ESTATE, "context is unusable after an earlier timeout");
    if (*(volatile int *)c->status_host != 0) {
        c->broken = true;
        return fail(GA_ETIMEOUT, "an earlier kernel timed out waiting for a peer rank's flag; the context is unusable");
    }
    if (!c->opened) return fail(GA_ESTATE, "ga_open_peers has not been called");
    if (op != GA_OP_F32_AVG && op != GA_OP_I32_SUM && op != GA_OP_F32_SUM) return fail(GA_EINVAL, "unknown op %d", op);
    if (nbytes % 4) return fail(GA_EINVAL, "nbytes must be a multiple of 4");
    if (nbytes > c->view_bytes) return fail(GA_ECAPACITY, "%zu bytes exceed the view region (%zu)", nbytes, c->view_bytes);
    if (variant == GA_VARIANT_AUTO) {
        // Below ~32 MiB the bucket kernels are quicker even with their pack and unpack (their one-shot
        // push skips the start-of-launch wait, profiles/r02_view_sweep_*): AUTO hands such sets to the
        // tensor-list path with the region as its one tensor, and takes the in-place kernel from there on.
        if (nbytes < c->view_min_bytes && c->world > 1) {
            const void *ptrs[kMaxWorld];
            for (int l = 0; l < c->n_local; ++l) ptrs[l] = c->view_own[l];
            const int64_t numel = (int64_t)(nbytes / 4);
            return ga_allreduce_tensors(c, ptrs, &numel, 1, op, stream, GA_VARIANT_AUTO);
        }
        variant = GA_VARIANT_VIEW;
    }
    if (variant != GA_VARIANT_VIEW && variant != GA_VARIANT_NVLS) return fail(GA_EINVAL, "ga_average_view takes GA_VARIANT_VIEW or GA_VARIANT_NVLS, not %d", variant);
    if (variant == GA_VARIANT_NVLS && !c->view_mc)
        return fail(GA_ESTATE, "GA_VARIANT_NVLS needs a multicast mapping: create the view region with GA_VIEW_MULTICAST on a "
                    "one-rank-per-GPU world whose GPUs support NVSwitch multicast, then ga_view_open, barrier, ga_view_bind");
    const uint32_t tiles = (uint32_t)((nbytes + kTileBytes - 1) / kTileBytes);
    if (tiles == 0) return GA_OK;
    ENTER_DEVICE(c->device);
    int rc = ga_order_stream(c, stream);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;

    ViewParams p;
    memset(&p, 0, sizeof p);
    for (int r = 0; r < c->world; ++r) {
        p.view[r] = c->view_all[r];
        p.stage[r] = c->send_all[r];
        p.pad[r] = c->pad_all[r];
    }
    p.mc = c->view_mc;
    p.status = c->status_dev;
    p.timeout_ns = c->timeout_ns;
    p.trace = c->trace_dev;
    p.total_tiles = tiles;
    p.shard_tiles = (tiles + c->world - 1) / c->world;
    p.rank_base = c->rank_base;
    p.world = c->world;

    ViewFn fn = variant == GA_VARIANT_NVLS ? pick_nvls(op) : pick_view(c->world, op);
    int &bps = c->blocks_per_sm[variant];
    if (bps == 0) {
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, (const void *)fn, kThreads, 0) != cudaSuccess || n < 1) n = 1;
        bps = n;
    }
    int resident = bps * c->sm_count / c->n_local;
    if (resident > kMaxBlocks) resident = kMaxBlocks;
    if (c->max_blocks_opt > 0 && resident > c->max_blocks_opt) resident = c->max_blocks_opt;
    const int wpb = kThreads / 32;
    int blocks;
    if (variant == GA_VARIANT_NVLS) {
        if (resident < 1) resident = 1;
        blocks = (int)((p.shard_tiles + wpb - 1) / wpb);
        if (blocks > resident) blocks = resident;
        if (blocks < 1) blocks = 1;
    } else {
        // One-shot (every rank sends its whole region to every peer, folds everything locally, one
        // flag round): same link bytes as two-shot at world 2, and one barrier less for small sets.
        const size_t set_bytes = (size_t)tiles * kTileBytes;
        // (N = 2 / 4 sweeps: one-shot wins up to 4 MiB / 1 MiB; beyond, its local fold no longer hides behind the transfer)
        const size_t one_max = c->view_oneshot_max_opt >= 0 ? (size_t)c->view_oneshot_max_opt
                                                            : ((size_t)4 << 20) / (size_t)(c->world > 1 ? c->world - 1 : 1);
        p.oneshot = c->world > 1 && set_bytes <= one_max && set_bytes * (size_t)(c->world - 1) <= c->capacity_bytes;
        // Scatter and fold blocks in pairs.  Chunk policy from the N = 2/4/8 sweeps (profiles/r02_view_sweep_*):
        // the links are the limit and a chunk's flag only arrives after the queue in front of it has
        // drained, so finer chunks do not start the fold any earlier -- large sets take one round of up
        // to 64 tiles per block on every SM, mid sizes run on half the SMs with fuller blocks, small
        // sets on at most 32 blocks of >= 4 tiles (more blocks only add flag traffic).
        if (resident < 2) resident = 2;
        const uint32_t nb_max = (uint32_t)resident / 2;
        const uint32_t Ts = p.oneshot ? tiles : p.shard_tiles;     // tiles the chunks of one rank run over
        uint32_t chunk = (uint32_t)c->view_chunk_opt;
        if (chunk == 0) {
            const uint32_t nb_half = nb_max > 1 ? nb_max / 2 : 1;
            if (Ts <= 1024) {
                chunk = (Ts + 31) / 32;
                if (chunk < 4) chunk = 4;
            } else if (Ts < nb_max * 32) {
                chunk = (Ts + nb_half - 1) / nb_half;
            } else {
                const uint32_t rounds = (Ts + nb_max * 64 - 1) / (nb_max * 64);
                chunk = (Ts + nb_max * rounds - 1) / (nb_max * rounds);
            }
        }
        while ((Ts + chunk - 1) / chunk > (uint32_t)kMaxBlocks) ++chunk;     // one flag row per chunk
        p.chunk_tiles = chunk;
        p.n_chunks = (Ts + chunk - 1) / chunk;
        uint32_t nb = p.n_chunks < nb_max ? p.n_chunks : nb_max;
        if (nb < 1) nb = 1;
        blocks = (int)(2 * nb);
    }
    dim3 grid(blocks, c->n_local), block(kThreads);
    void *args[] = {&p};
    if (c->n_local > 1) CUDA_TRY(cudaLaunchCooperativeKernel((const void *)fn, grid, block, args, 0, st));
    else CUDA_TRY(cudaLaunchKernel((const void *)fn, grid, block, args, 0, st));
    c->launches += 1;
    c->last_was_tensor_call = true;        // staging readers are checked lazily, like the tensor-list calls
    c->last_blocks = blocks;
    c->last_threads = kThreads;
    return GA_OK;
}

}  // extern "C"
