// nvlink_probe.cu -- raw peer bandwidth between GPU 0 and GPU 1 of one process (build-time tool, not
// part of the library): SM-issued 16-byte stores (push) and loads (pull), one direction and both
// directions at once, against cudaMemcpyPeerAsync.  Tells how close the averaging kernels are to
// what the links give.   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o nvlink_probe nvlink_probe.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void copy_kernel(const uint4 *__restrict__ src, uint4 *__restrict__ dst, size_t n, int unroll4) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    if (unroll4) {
        for (; i + 3 * stride < n; i += 4 * stride) {
            uint4 a = __ldcg(src + i), b = __ldcg(src + i + stride), c = __ldcg(src + i + 2 * stride), d = __ldcg(src + i + 3 * stride);
            __stcg(dst + i, a); __stcg(dst + i + stride, b); __stcg(dst + i + 2 * stride, c); __stcg(dst + i + 3 * stride, d);
        }
    }
    for (; i < n; i += stride) __stcg(dst + i, __ldcg(src + i));
}

int main(int argc, char **argv) {
    size_t bytes = argc > 1 ? (size_t)atol(argv[1]) << 20 : (size_t)256 << 20;
    int ndev = 0; CK(cudaGetDeviceCount(&ndev));
    if (ndev < 2) { printf("need 2 GPUs\n"); return 0; }
    uint4 *buf[2][2]; cudaStream_t st[2]; cudaEvent_t e0[2], e1[2];
    for (int d = 0; d < 2; ++d) {
        CK(cudaSetDevice(d)); CK(cudaDeviceEnablePeerAccess(1 - d, 0));
        for (int k = 0; k < 2; ++k) { CK(cudaMalloc(&buf[d][k], bytes)); CK(cudaMemset(buf[d][k], d + k, bytes)); }
        CK(cudaStreamCreate(&st[d])); CK(cudaEventCreate(&e0[d])); CK(cudaEventCreate(&e1[d]));
    }
    const size_t n = bytes / 16;
    auto run = [&](const char *name, int mode, int both, int blocks, int threads, int unroll) {
        float best[2] = {1e30f, 1e30f};
        for (int rep = 0; rep < 6; ++rep) {
            for (int d = 0; d < (both ? 2 : 1); ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
            for (int d = 0; d < (both ? 2 : 1); ++d) {
                CK(cudaSetDevice(d)); CK(cudaEventRecord(e0[d], st[d]));
                if (mode == 0) copy_kernel<<<blocks, threads, 0, st[d]>>>(buf[d][0], buf[1 - d][1], n, unroll);        // push: local load, remote store
                else if (mode == 1) copy_kernel<<<blocks, threads, 0, st[d]>>>(buf[1 - d][0], buf[d][1], n, unroll);   // pull: remote load, local store
                else CK(cudaMemcpyPeerAsync(buf[1 - d][1], 1 - d, buf[d][0], d, bytes, st[d]));
                CK(cudaEventRecord(e1[d], st[d]));
            }
            for (int d = 0; d < (both ? 2 : 1); ++d) {
                CK(cudaSetDevice(d)); CK(cudaEventSynchronize(e1[d]));
                float ms; CK(cudaEventElapsedTime(&ms, e0[d], e1[d])); if (ms < best[d]) best[d] = ms;
            }
        }
        printf("%-28s %s blocks=%4d thr=%4d unroll=%d : %7.1f GB/s", name, both ? "both dirs" : "one dir  ", blocks, threads, unroll, bytes / best[0] / 1e6);
        if (both) printf("  | %7.1f GB/s", bytes / best[1] / 1e6);
        printf("\n");
    };
    for (int both = 0; both < 2; ++both) {
        run("memcpyPeerAsync", 2, both, 0, 0, 0);
        for (int unroll = 0; unroll < 2; ++unroll) {
            run("SM push (remote st.v4)", 0, both, 148, 1024, unroll);
            run("SM push (remote st.v4)", 0, both, 296, 512, unroll);
            run("SM push (remote st.v4)", 0, both, 74, 512, unroll);
            run("SM push (remote st.v4)", 0, both, 32, 512, unroll);
            run("SM pull (remote ld.v4)", 1, both, 148, 1024, unroll);
            run("SM pull (remote ld.v4)", 1, both, 296, 512, unroll);
            run("SM pull (remote ld.v4)", 1, both, 74, 512, unroll);
        }
    }
    return 0;
}
