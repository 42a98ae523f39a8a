// nvlink_probe.cu -- what the NVLink fabric gives SM-issued and engine-issued traffic on this pool
// (build-time tool, not part of the library).  One process drives G GPUs with peer access; every
// GPU talks to ALL its peers at once (the averaging kernels' pattern), so the figures are per GPU
// per direction with both directions busy.  Transports:
//   ce        cudaMemcpyPeerAsync, one stream per (src, dst) pair
//   st16      local ld.v4 -> remote st.v4        (what round 1's push kernels use)
//   st32      local ld.v8 -> remote st.v8        (256-bit vectors, new on sm_100)
//   ld16      remote ld.v4 -> local st.v4        (pull)
//   ld32      remote ld.v8 -> local st.v8
//   tma_st    local cp.async.bulk global->shared, then cp.async.bulk shared->REMOTE global
//   tma_ld    cp.async.bulk REMOTE global->shared, then cp.async.bulk shared->local global
//   duplex16  remote ld.v4 from peer p AND remote st.v4 to peer p in the same thread (in-place exchange)
//   mc_st     multimem.st to a multicast mapping: one store lands on every GPU (needs NVSwitch multicast)
//   mc_ldred  multimem.ld_reduce.add.f32 from the multicast mapping (in-switch reduction)
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o nvlink_probe nvlink_probe.cu -lcuda
// Run:   ./nvlink_probe [MiB per GPU = 256] [G = all]
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s (line %d)\n", #x, cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)
#define CU(x) do { CUresult e_ = (x); if (e_ != CUDA_SUCCESS) { const char *s_ = 0; cuGetErrorString(e_, &s_); printf("%s: %s (line %d)\n", #x, s_ ? s_ : "?", __LINE__); return false; } } while (0)

constexpr int kMaxG = 8;
struct Ptrs { char *p[kMaxG]; };

struct V8 { float f[8]; };
__device__ __forceinline__ V8 ld8(const void *p) {
    V8 v;
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v.f[0]), "=f"(v.f[1]), "=f"(v.f[2]), "=f"(v.f[3]), "=f"(v.f[4]), "=f"(v.f[5]), "=f"(v.f[6]), "=f"(v.f[7]) : "l"(p));
    return v;
}
__device__ __forceinline__ void st8(void *p, const V8 &v) {
    asm volatile("st.global.v8.f32 [%8], {%0,%1,%2,%3,%4,%5,%6,%7};"
                 :: "f"(v.f[0]), "f"(v.f[1]), "f"(v.f[2]), "f"(v.f[3]), "f"(v.f[4]), "f"(v.f[5]), "f"(v.f[6]), "f"(v.f[7]), "l"(p) : "memory");
}

// GPU `me` moves chunk (me -> peer) for every peer: bytes_per_peer each.  src/dst selection by mode.
//   push: src = own[me] + peer * B, dst = other[peer] + me * B
//   pull: src = own-side buffer of the peer, dst local
template <int VEC, bool PULL>
__global__ void __launch_bounds__(512, 2) simt_kernel(Ptrs in, Ptrs out, int me, int G, size_t bytes_per_peer) {
    const size_t nvec = bytes_per_peer / VEC;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
    constexpr int U = 4;
#pragma unroll 1
    for (int jj = 1; jj < G; ++jj) {
        const int peer = (me + jj) % G;
        const char *src = PULL ? in.p[peer] + (size_t)me * bytes_per_peer : in.p[me] + (size_t)peer * bytes_per_peer;
        char *dst = PULL ? out.p[me] + (size_t)peer * bytes_per_peer : out.p[peer] + (size_t)me * bytes_per_peer;
        size_t i = tid;
        for (; i + (U - 1) * nthr < nvec; i += U * nthr) {
            if (VEC == 16) {
                uint4 v[U];
#pragma unroll
                for (int u = 0; u < U; ++u) v[u] = __ldcg((const uint4 *)src + i + u * nthr);
#pragma unroll
                for (int u = 0; u < U; ++u) __stcg((uint4 *)dst + i + u * nthr, v[u]);
            } else {
                V8 v[U];
#pragma unroll
                for (int u = 0; u < U; ++u) v[u] = ld8(src + (i + u * nthr) * 32);
#pragma unroll
                for (int u = 0; u < U; ++u) st8(dst + (i + u * nthr) * 32, v[u]);
            }
        }
        for (; i < nvec; i += nthr) {
            if (VEC == 16) __stcg((uint4 *)dst + i, __ldcg((const uint4 *)src + i));
            else st8(dst + i * 32, ld8(src + i * 32));
        }
    }
}

// In-place exchange pattern: every thread loads a vector from EVERY peer and stores a vector to EVERY peer.
__global__ void __launch_bounds__(512, 2) duplex_kernel(Ptrs in, Ptrs out, int me, int G, size_t bytes_per_peer) {
    const size_t nvec = bytes_per_peer / 16;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
    for (size_t i = tid; i < nvec; i += nthr) {
        uint4 v[kMaxG];
#pragma unroll
        for (int jj = 1; jj < kMaxG; ++jj)
            if (jj < G) v[jj] = __ldcg((const uint4 *)(in.p[(me + jj) % G] + (size_t)me * bytes_per_peer) + i);
        uint4 acc = v[1];
#pragma unroll
        for (int jj = 2; jj < kMaxG; ++jj)
            if (jj < G) { acc.x += v[jj].x; acc.y ^= v[jj].y; acc.z += v[jj].z; acc.w ^= v[jj].w; }
#pragma unroll
        for (int jj = 1; jj < kMaxG; ++jj)
            if (jj < G) __stcg((uint4 *)(out.p[(me + jj) % G] + (size_t)me * bytes_per_peer) + i, acc);
    }
}

// ---- TMA bulk (1-D cp.async.bulk) --------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint64_t *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem, const void *g, uint32_t bytes, uint64_t *b) {
    asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem)), "l"(g), "r"(bytes), "r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *g, const void *smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(smem_u32(smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }

// One thread per block drives a ring of STAGES shared-memory chunks: load chunk (local or remote) with
// mbarrier completion, store it (remote or local) as a bulk group.  PULL: the global->shared side is remote.
template <bool PULL>
__global__ void tma_kernel(Ptrs in, Ptrs out, int me, int G, size_t bytes_per_peer, int chunk, int stages) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bar[8];
    if (threadIdx.x != 0) return;
    for (int s = 0; s < stages; ++s) mbar_init(&bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    const size_t chunks_per_peer = bytes_per_peer / chunk;
    const size_t total = chunks_per_peer * (G - 1);
    // chunk c of this block: global chunk index k = c * gridDim.x + blockIdx.x; peer rotates fastest
    auto addr = [&](size_t k, const char *&src, char *&dst) {
        const int jj = 1 + (int)(k % (G - 1));
        const size_t off = (k / (G - 1)) * (size_t)chunk;
        const int peer = (me + jj) % G;
        src = PULL ? in.p[peer] + (size_t)me * bytes_per_peer + off : in.p[me] + (size_t)peer * bytes_per_peer + off;
        dst = PULL ? out.p[me] + (size_t)peer * bytes_per_peer + off : out.p[peer] + (size_t)me * bytes_per_peer + off;
    };
    const size_t mine = total > blockIdx.x ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    uint32_t phase[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    size_t issued = 0;
    for (; issued < mine && issued < (size_t)stages; ++issued) {
        const char *src; char *dst; addr(issued * gridDim.x + blockIdx.x, src, dst);
        mbar_expect(&bar[issued], chunk);
        bulk_g2s(smem + issued * chunk, src, chunk, &bar[issued]);
    }
    for (size_t c = 0; c < mine; ++c) {
        const int s = (int)(c % stages);
        const char *src; char *dst; addr(c * gridDim.x + blockIdx.x, src, dst);
        mbar_wait(&bar[s], phase[s]); phase[s] ^= 1;
        bulk_s2g(dst, smem + s * chunk, chunk);
        bulk_commit();
        if (c >= 1 && issued < mine) {
            // all stores but the newest have read their shared-memory source: the slot of chunk c-1 is free
            bulk_wait_read<1>();
            const int s2 = (int)((c - 1) % stages);
            const char *src2; char *dst2; addr(issued * gridDim.x + blockIdx.x, src2, dst2);
            mbar_expect(&bar[s2], chunk);
            bulk_g2s(smem + s2 * chunk, src2, chunk, &bar[s2]);
            ++issued;
        }
    }
    bulk_wait<0>();
}

// multicast: store own slice to the multicast address (lands on all G GPUs) / load-reduce a slice
__global__ void mc_st_kernel(const char *src, char *mc_dst, size_t bytes) {
    const size_t nvec = bytes / 16;
    const size_t nthr = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += nthr) {
        float4 v = __ldcg((const float4 *)src + i);
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%4], {%0,%1,%2,%3};" ::"f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"((float4 *)mc_dst + i) : "memory");
    }
}
__global__ void mc_ldred_kernel(const char *mc_src, char *dst, size_t bytes) {
    const size_t nvec = bytes / 16;
    const size_t nthr = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += nthr) {
        float4 v;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"((const float4 *)mc_src + i) : "memory");
        __stcg((float4 *)dst + i, v);
    }
}

// system-scope fence cost after a burst of remote stores, and a flag ping-pong round trip
__global__ void fence_kernel(char *remote, size_t bytes_per_block, unsigned long long *out_ns) {
    uint4 *dst = (uint4 *)(remote + (size_t)blockIdx.x * bytes_per_block);
    const size_t n = bytes_per_block / 16;
    unsigned long long t0, t1, t2;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (size_t i = threadIdx.x; i < n; i += blockDim.x) __stcg(dst + i, make_uint4(1, 2, 3, 4));
    __syncthreads();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (threadIdx.x == 0) asm volatile("fence.acq_rel.sys;" ::: "memory");
    __syncthreads();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t2));
    if (threadIdx.x == 0) { out_ns[blockIdx.x * 2] = t1 - t0; out_ns[blockIdx.x * 2 + 1] = t2 - t1; }
}
__global__ void pingpong_kernel(volatile uint32_t *mine, volatile uint32_t *theirs, int first, int iters, unsigned long long *out_ns) {
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (int i = 1; i <= iters; ++i) {
        if (first) { *theirs = i; while (*mine < (uint32_t)i) {} }
        else { while (*mine < (uint32_t)i) {} *theirs = i; }
    }
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    *out_ns = t1 - t0;
}


// Do system-scope fences issued by ONE thread slow down the other warps' remote stores?  Warps 1..15
// stream the all-to-all copy; warp 0 lane 0 issues fence.acq_rel.sys back to back (mode 1), once per
// `period` microseconds (mode 2), or never (mode 0), until the data warps are done.
__global__ void __launch_bounds__(512, 2) fence_stream_kernel(Ptrs in, Ptrs out, int me, int G, size_t bytes_per_peer, int mode, int period_us,
                                                              unsigned long long *stats) {
    __shared__ unsigned int s_done;
    if (threadIdx.x == 0) s_done = 0;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        if (lane == 0) {
            unsigned long long n = 0, t_in = 0;
            while (*(volatile unsigned int *)&s_done < 15u) {
                if (mode == 0) continue;
                unsigned long long a, b;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(a));
                asm volatile("fence.acq_rel.sys;" ::: "memory");
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(b));
                ++n; t_in += b - a;
                if (mode == 2) { unsigned long long c = b; while (c - b < (unsigned long long)period_us * 1000ull && *(volatile unsigned int *)&s_done < 15u) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(c)); }
            }
            if (stats) { atomicAdd(stats, n); atomicAdd(stats + 1, t_in); }
        }
        return;
    }
    const size_t nvec = bytes_per_peer / 16;
    const size_t tid = (size_t)blockIdx.x * 480 + (threadIdx.x - 32), nthr = (size_t)gridDim.x * 480;
#pragma unroll 1
    for (int jj = 1; jj < G; ++jj) {
        const int peer = (me + jj) % G;
        const uint4 *src = (const uint4 *)(in.p[me] + (size_t)peer * bytes_per_peer);
        uint4 *dst = (uint4 *)(out.p[peer] + (size_t)me * bytes_per_peer);
        size_t i = tid;
        for (; i + 3 * nthr < nvec; i += 4 * nthr) {
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = __ldcg(src + i + u * nthr);
#pragma unroll
            for (int u = 0; u < 4; ++u) __stcg(dst + i + u * nthr, v[u]);
        }
        for (; i < nvec; i += nthr) __stcg(dst + i, __ldcg(src + i));
    }
    __syncwarp();
    if (lane == 0) atomicAdd(&s_done, 1u);
}

// cost of polling W flags one after the other (what a block barrier's poller does): W dependent
// ld.relaxed.sys of local memory + one fence.acq_rel.sys, nothing outstanding
__global__ void poll_cost_kernel(const uint32_t *flags, int W, unsigned long long *out_ns) {
    unsigned long long t0, t1, t2;
    uint32_t acc = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (int s = 0; s < W; ++s) { uint32_t v; asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + s * 32) : "memory"); acc += v; }
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    asm volatile("fence.acq_rel.sys;" ::: "memory");
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t2));
    out_ns[0] = t1 - t0; out_ns[1] = t2 - t1; out_ns[2] = acc;
}

int G = 0;
size_t bytes = 0;
char *bufA[kMaxG], *bufB[kMaxG];
cudaStream_t st[kMaxG], pair_st[kMaxG][kMaxG];
cudaEvent_t e0[kMaxG], e1[kMaxG];

template <typename F> void timed(const char *name, const char *cfg, double bytes_per_gpu_dir, F launch) {
    float best = 1e30f, worst_of_best = 0;
    float bestd[kMaxG];
    for (int d = 0; d < G; ++d) bestd[d] = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        for (int d = 0; d < G; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
        for (int d = 0; d < G; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(e0[d], st[d])); }
        for (int d = 0; d < G; ++d) { CK(cudaSetDevice(d)); launch(d); }
        for (int d = 0; d < G; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(e1[d], st[d])); }
        for (int d = 0; d < G; ++d) {
            CK(cudaSetDevice(d)); CK(cudaEventSynchronize(e1[d]));
            float ms; CK(cudaEventElapsedTime(&ms, e0[d], e1[d]));
            if (ms < bestd[d]) bestd[d] = ms;
        }
    }
    for (int d = 0; d < G; ++d) { if (bestd[d] < best) best = bestd[d]; if (bestd[d] > worst_of_best) worst_of_best = bestd[d]; }
    printf("{\"transport\": \"%s\", \"cfg\": \"%s\", \"gpus\": %d, \"gbps_per_gpu_dir_slowest\": %.1f, \"gbps_fastest\": %.1f, \"ms_slowest\": %.4f}\n",
           name, cfg, G, bytes_per_gpu_dir / worst_of_best / 1e6, bytes_per_gpu_dir / best / 1e6, worst_of_best);
    fflush(stdout);
}

bool multicast_probe(size_t mc_bytes);

int main(int argc, char **argv) {
    bytes = (argc > 1 ? (size_t)atol(argv[1]) : 256) << 20;
    CK(cudaGetDeviceCount(&G));
    if (argc > 2 && atoi(argv[2]) < G) G = atoi(argv[2]);
    if (G > kMaxG) G = kMaxG;
    if (G < 2) { printf("{\"error\": \"need 2 GPUs\"}\n"); return 0; }
    for (int d = 0; d < G; ++d) {
        CK(cudaSetDevice(d));
        for (int p = 0; p < G; ++p) if (p != d) CK(cudaDeviceEnablePeerAccess(p, 0));
        CK(cudaMalloc(&bufA[d], bytes)); CK(cudaMalloc(&bufB[d], bytes));
        CK(cudaMemset(bufA[d], d + 1, bytes)); CK(cudaMemset(bufB[d], 0, bytes));
        CK(cudaStreamCreateWithFlags(&st[d], cudaStreamNonBlocking)); CK(cudaEventCreate(&e0[d])); CK(cudaEventCreate(&e1[d]));
        for (int p = 0; p < G; ++p) CK(cudaStreamCreateWithFlags(&pair_st[d][p], cudaStreamNonBlocking));
    }
    Ptrs A, B;
    for (int d = 0; d < kMaxG; ++d) { A.p[d] = d < G ? bufA[d] : nullptr; B.p[d] = d < G ? bufB[d] : nullptr; }
    // each GPU's buffer is cut into G slices; slice p is what it exchanges with peer p
    const size_t per_peer = bytes / G / 65536 * 65536;
    const double dir_bytes = (double)per_peer * (G - 1);          // per GPU per direction
    printf("{\"probe\": \"all-to-all\", \"gpus\": %d, \"bytes_per_gpu\": %zu, \"bytes_per_peer\": %zu}\n", G, bytes, per_peer);

    // copy engines: for the event on st[d] to cover the copies, fork/join through events
    {
        cudaEvent_t fork[kMaxG], join[kMaxG][kMaxG];
        for (int d = 0; d < G; ++d) { CK(cudaSetDevice(d)); CK(cudaEventCreateWithFlags(&fork[d], cudaEventDisableTiming)); for (int p = 0; p < G; ++p) CK(cudaEventCreateWithFlags(&join[d][p], cudaEventDisableTiming)); }
        timed("ce", "cudaMemcpyPeerAsync per pair", dir_bytes, [&](int d) {
            CK(cudaEventRecord(fork[d], st[d]));
            for (int jj = 1; jj < G; ++jj) {
                const int p = (d + jj) % G;
                CK(cudaStreamWaitEvent(pair_st[d][p], fork[d], 0));
                CK(cudaMemcpyPeerAsync(bufB[p] + (size_t)d * per_peer, p, bufA[d] + (size_t)p * per_peer, d, per_peer, pair_st[d][p]));
                CK(cudaEventRecord(join[d][p], pair_st[d][p]));
                CK(cudaStreamWaitEvent(st[d], join[d][p], 0));
            }
        });
    }
    const int grids[][2] = {{148, 512}, {296, 512}, {592, 256}, {74, 512}, {32, 512}};
    for (auto &g : grids) {
        char cfg[64];
        snprintf(cfg, sizeof cfg, "%d blocks x %d thr", g[0], g[1]);
        timed("st16", cfg, dir_bytes, [&](int d) { simt_kernel<16, false><<<g[0], g[1], 0, st[d]>>>(A, B, d, G, per_peer); CK(cudaGetLastError()); });
        timed("st32", cfg, dir_bytes, [&](int d) { simt_kernel<32, false><<<g[0], g[1], 0, st[d]>>>(A, B, d, G, per_peer); });
        timed("ld16", cfg, dir_bytes, [&](int d) { simt_kernel<16, true><<<g[0], g[1], 0, st[d]>>>(A, B, d, G, per_peer); });
        timed("ld32", cfg, dir_bytes, [&](int d) { simt_kernel<32, true><<<g[0], g[1], 0, st[d]>>>(A, B, d, G, per_peer); });
        timed("duplex16", cfg, dir_bytes, [&](int d) { duplex_kernel<<<g[0], g[1], 0, st[d]>>>(A, B, d, G, per_peer / 2 / 16 * 16); CK(cudaGetLastError()); });
    }
    {
        unsigned long long *stats[kMaxG];
        for (int d = 0; d < G; ++d) { CK(cudaSetDevice(d)); CK(cudaMalloc(&stats[d], 16)); }
        const int modes[][2] = {{0, 0}, {1, 0}, {2, 20}, {2, 5}};
        for (auto &m : modes) {
            char cfg[96];
            snprintf(cfg, sizeof cfg, "296 x 512 (15 data warps), fences: %s", m[0] == 0 ? "none" : m[0] == 1 ? "back to back by 1 thread per block" : m[1] == 20 ? "every 20 us" : "every 5 us");
            for (int d = 0; d < G; ++d) { CK(cudaSetDevice(d)); CK(cudaMemset(stats[d], 0, 16)); }
            timed("st16_fence", cfg, dir_bytes, [&](int d) { fence_stream_kernel<<<296, 512, 0, st[d]>>>(A, B, d, G, per_peer, m[0], m[1], stats[d]); CK(cudaGetLastError()); });
            unsigned long long h[2]; CK(cudaSetDevice(0)); CK(cudaMemcpy(h, stats[0], 16, cudaMemcpyDeviceToHost));
            printf("{\"fences_issued_gpu0_over_5_reps\": %llu, \"mean_fence_us\": %.2f}\n", h[0], h[0] ? (double)h[1] / h[0] / 1e3 : 0.0);
        }
        uint32_t *fl; unsigned long long *pc; CK(cudaSetDevice(0)); CK(cudaMalloc(&fl, 4096)); CK(cudaMemset(fl, 0, 4096)); CK(cudaMallocHost(&pc, 32));
        for (int rep = 0; rep < 3; ++rep) { poll_cost_kernel<<<1, 1, 0, st[0]>>>(fl, 8, pc); CK(cudaStreamSynchronize(st[0])); }
        printf("{\"poll_8_flags_sequentially_us\": %.2f, \"idle_fence_us\": %.2f}\n", pc[0] / 1e3, pc[1] / 1e3);
    }
    // note: duplex moves per_peer/2 bytes per peer pair as load data AND the same as store data, i.e. per_peer per pair per direction
    const int tma_cfg[][3] = {{148, 16384, 4}, {148, 32768, 4}, {148, 8192, 8}, {296, 16384, 4}, {74, 32768, 6}, {32, 32768, 6}};
    for (auto &c : tma_cfg) {
        char cfg[64];
        snprintf(cfg, sizeof cfg, "%d blocks, %d B chunks x %d stages", c[0], c[1], c[2]);
        const size_t sm = (size_t)c[1] * c[2];
        for (int d = 0; d < G; ++d) {
            CK(cudaSetDevice(d));
            CK(cudaFuncSetAttribute(tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
            CK(cudaFuncSetAttribute(tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        }
        timed("tma_st", cfg, dir_bytes, [&](int d) { tma_kernel<false><<<c[0], 32, sm, st[d]>>>(A, B, d, G, per_peer, c[1], c[2]); });
        timed("tma_ld", cfg, dir_bytes, [&](int d) { tma_kernel<true><<<c[0], 32, sm, st[d]>>>(A, B, d, G, per_peer, c[1], c[2]); });
        CK(cudaGetLastError());
    }
    // fence cost after remote stores, GPU 0 -> GPU 1, alone on the link
    {
        unsigned long long *ns; CK(cudaSetDevice(0)); CK(cudaMallocHost(&ns, 148 * 2 * 8));
        const size_t per_block[] = {2048, 16384, 131072, 1048576};
        for (size_t pb : per_block) {
            for (int rep = 0; rep < 2; ++rep) { fence_kernel<<<148, 512, 0, st[0]>>>(bufB[1], pb, ns); CK(cudaStreamSynchronize(st[0])); }
            double a = 0, b = 0, bm = 0;
            for (int i = 0; i < 148; ++i) { a += ns[2 * i]; b += ns[2 * i + 1]; if (ns[2 * i + 1] > bm) bm = ns[2 * i + 1]; }
            printf("{\"fence_after_remote_stores\": {\"bytes_per_block\": %zu, \"blocks\": 148, \"issue_us_mean\": %.2f, \"fence_us_mean\": %.2f, \"fence_us_max\": %.2f}}\n",
                   pb, a / 148 / 1e3, b / 148 / 1e3, bm / 1e3);
        }
        // flag ping-pong 0 <-> 1
        uint32_t *f0, *f1; CK(cudaSetDevice(0)); CK(cudaMalloc(&f0, 256)); CK(cudaMemset(f0, 0, 256));
        CK(cudaSetDevice(1)); CK(cudaMalloc(&f1, 256)); CK(cudaMemset(f1, 0, 256));
        unsigned long long *pn; CK(cudaMallocHost(&pn, 16));
        CK(cudaDeviceSynchronize()); CK(cudaSetDevice(0)); CK(cudaDeviceSynchronize());
        const int iters = 2000;
        CK(cudaSetDevice(0)); pingpong_kernel<<<1, 1, 0, st[0]>>>(f0, f1, 1, iters, ns);
        CK(cudaSetDevice(1)); pingpong_kernel<<<1, 1, 0, st[1]>>>(f1, f0, 0, iters, pn);
        CK(cudaStreamSynchronize(st[1])); CK(cudaSetDevice(0)); CK(cudaStreamSynchronize(st[0]));
        printf("{\"flag_round_trip_us\": %.3f}\n", (double)ns[0] / iters / 1e3);
    }
    size_t mcb = bytes < ((size_t)256 << 20) ? bytes : ((size_t)256 << 20);
    if (!multicast_probe(mcb)) printf("{\"multicast\": \"unavailable or failed (see message above)\"}\n");
    return 0;
}

bool multicast_probe(size_t mc_bytes) {
    CU(cuInit(0));
    CUdevice dev[kMaxG];
    int all = 1;
    for (int d = 0; d < G; ++d) {
        CU(cuDeviceGet(&dev[d], d));
        int sup = 0;
        CU(cuDeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev[d]));
        int fabric = 0;
        cuDeviceGetAttribute(&fabric, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, dev[d]);
        int posix = 0;
        cuDeviceGetAttribute(&posix, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev[d]);
        if (d == 0) printf("{\"multicast_supported\": %d, \"fabric_handles\": %d, \"posix_fd_handles\": %d}\n", sup, fabric, posix);
        all &= sup;
    }
    if (!all) return false;
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof mp);
    mp.numDevices = G;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    mp.size = mc_bytes;
    size_t gran = 0, gran_min = 0;
    CU(cuMulticastGetGranularity(&gran_min, &mp, CU_MULTICAST_GRANULARITY_MINIMUM));
    CU(cuMulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    printf("{\"multicast_granularity_min\": %zu, \"recommended\": %zu}\n", gran_min, gran);
    mp.size = (mc_bytes + gran - 1) / gran * gran;
    const size_t size = mp.size;
    CUmemGenericAllocationHandle mc;
    CU(cuMulticastCreate(&mc, &mp));
    for (int d = 0; d < G; ++d) CU(cuMulticastAddDevice(mc, dev[d]));
    CUmemGenericAllocationHandle mem[kMaxG];
    CUdeviceptr uc[kMaxG], mcva[kMaxG];
    for (int d = 0; d < G; ++d) {
        CK(cudaSetDevice(d));
        CUmemAllocationProp ap;
        memset(&ap, 0, sizeof ap);
        ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ap.location.id = d;
        ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        CU(cuMemCreate(&mem[d], size, &ap, 0));
        CU(cuMulticastBindMem(mc, 0, mem[d], 0, size, 0));
        CUmemAccessDesc ad;
        memset(&ad, 0, sizeof ad);
        ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ad.location.id = d;
        ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        CU(cuMemAddressReserve(&uc[d], size, gran, 0, 0));
        CU(cuMemMap(uc[d], size, 0, mem[d], 0));
        CU(cuMemSetAccess(uc[d], size, &ad, 1));
        CU(cuMemAddressReserve(&mcva[d], size, gran, 0, 0));
        CU(cuMemMap(mcva[d], size, 0, mc, 0));
        CU(cuMemSetAccess(mcva[d], size, &ad, 1));
        CK(cudaMemset((void *)uc[d], 0, size));
    }
    // correctness: GPU 0 multicasts a pattern, every GPU must see it in its unicast mapping; then ld_reduce
    const size_t slice = mc_bytes / G / 65536 * 65536;
    for (int d = 0; d < G; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
    for (int d = 0; d < G; ++d) {
        CK(cudaSetDevice(d));
        mc_st_kernel<<<148, 512, 0, st[d]>>>(bufA[d], (char *)mcva[d] + (size_t)d * slice, slice);
        CK(cudaStreamSynchronize(st[d]));
    }
    int ok = 1;
    for (int d = 0; d < G; ++d) {
        unsigned char h[kMaxG];
        for (int s = 0; s < G; ++s) {
            CK(cudaSetDevice(d));
            CK(cudaMemcpy(&h[s], (char *)uc[d] + (size_t)s * slice + 12345, 1, cudaMemcpyDeviceToHost));
            if (h[s] != (unsigned char)(s + 1)) ok = 0;
        }
    }
    printf("{\"multicast_store_lands_everywhere\": %d}\n", ok);
    const double st_bytes = (double)slice;                 // egress per GPU; ingress per GPU = G * slice
    timed("mc_st", "148 x 512, each GPU multicasts its 1/G slice (egress S/G, ingress S)", st_bytes * G, [&](int d) {
        mc_st_kernel<<<148, 512, 0, st[d]>>>(bufA[d], (char *)mcva[d] + (size_t)d * slice, slice);
    });
    timed("mc_st", "296 x 512", st_bytes * G, [&](int d) {
        mc_st_kernel<<<296, 512, 0, st[d]>>>(bufA[d], (char *)mcva[d] + (size_t)d * slice, slice);
    });
    timed("mc_ldred", "148 x 512, each GPU reduces its 1/G slice in the switch (ingress S/G, egress (G-1)S/G)", st_bytes * G, [&](int d) {
        mc_ldred_kernel<<<148, 512, 0, st[d]>>>((const char *)mcva[d] + (size_t)d * slice, bufB[d], slice);
    });
    timed("mc_ldred", "296 x 512", st_bytes * G, [&](int d) {
        mc_ldred_kernel<<<296, 512, 0, st[d]>>>((const char *)mcva[d] + (size_t)d * slice, bufB[d], slice);
    });
    // is the in-switch sum the rank-order sum?  fill each GPU's slice 0 with values whose fp32 sum depends on the order
    {
        const int n = 1 << 16;
        float *h = (float *)malloc(n * sizeof(float) * kMaxG);
        for (int d = 0; d < G; ++d) {
            unsigned s = 12345u + 977u * d;
            for (int i = 0; i < n; ++i) { s = s * 1664525u + 1013904223u; h[d * n + i] = ((int)(s >> 8) - (1 << 23)) * (1.0f / (1 << 20)) * (1 + (s & 7)); }
            CK(cudaSetDevice(d)); CK(cudaMemcpy((void *)uc[d], h + d * n, n * sizeof(float), cudaMemcpyHostToDevice));
        }
        for (int d = 0; d < G; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
        CK(cudaSetDevice(0));
        mc_ldred_kernel<<<32, 256, 0, st[0]>>>((const char *)mcva[0], bufB[0], n * sizeof(float));
        CK(cudaStreamSynchronize(st[0]));
        float *got = (float *)malloc(n * sizeof(float));
        CK(cudaMemcpy(got, bufB[0], n * sizeof(float), cudaMemcpyDeviceToHost));
        int same_rank_order = 0, same_reverse = 0, same_double = 0;
        for (int i = 0; i < n; ++i) {
            float a = h[i]; for (int d = 1; d < G; ++d) a = a + h[d * n + i];
            float b = h[(G - 1) * n + i]; for (int d = G - 2; d >= 0; --d) b = b + h[d * n + i];
            double c = 0; for (int d = 0; d < G; ++d) c += h[d * n + i];
            same_rank_order += (got[i] == a); same_reverse += (got[i] == b); same_double += (got[i] == (float)c);
        }
        printf("{\"ld_reduce_vs_host\": {\"n\": %d, \"equals_rank_order_fold\": %d, \"equals_reverse_fold\": %d, \"equals_rounded_exact_sum\": %d}}\n",
               n, same_rank_order, same_reverse, same_double);
    }
    return true;
}
