// Micro-benchmark: rate of mma.sync m16n8k256 b1 AND+POPC (emulated through IMMA on sm_100a) versus the
// CUDA-core AND + POPC + IADD path, in "word-ops" (32 bit-pairs ANDed and counted) per second.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void k_mma(const uint32_t* src, int* out, int iters) {
    uint32_t a[4], b[4][2]; int c[4][4] = {};
    for (int i = 0; i < 4; ++i) a[i] = src[threadIdx.x + 32 * i];
    for (int j = 0; j < 4; ++j) { b[j][0] = src[threadIdx.x + 7 * j]; b[j][1] = src[threadIdx.x + 3 * j + 1]; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            asm volatile("mma.sync.aligned.m16n8k256.row.col.s32.b1.b1.s32.and.popc {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                : "+r"(c[j][0]), "+r"(c[j][1]), "+r"(c[j][2]), "+r"(c[j][3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[j][0]), "r"(b[j][1]));
        a[0] ^= it;
    }
    int s = 0; for (int j = 0; j < 4; ++j) for (int i = 0; i < 4; ++i) s += c[j][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_popc(const uint32_t* src, int* out, int iters) {
    uint32_t a[4], b[4]; int c[4][4] = {};
    for (int i = 0; i < 4; ++i) { a[i] = src[threadIdx.x + 32 * i]; b[i] = src[threadIdx.x + 5 * i]; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) c[i][j] += __popc(a[i] & b[j]);
        a[0] ^= it; b[1] += it;
    }
    int s = 0; for (int j = 0; j < 4; ++j) for (int i = 0; i < 4; ++i) s += c[j][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    uint32_t* src; int* out; cudaMalloc(&src, 1 << 20); cudaMalloc(&out, 148 * 8 * 256 * 4); cudaMemset(src, 0x5a, 1 << 20);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1); float ms;
    const int iters = 20000, grid = 148 * 4, block = 256;
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0); k_mma<<<grid, block>>>(src, out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
        double wops = (double)grid * (block / 32) * iters * 4.0 * (16 * 8 * 8);
        printf("b1 mma : %.3f ms  %.3e word-ops/s\n", ms, wops / (ms * 1e-3));
        cudaEventRecord(e0); k_popc<<<grid, block>>>(src, out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
        wops = (double)grid * block * iters * 16.0;
        printf("popc   : %.3f ms  %.3e word-ops/s\n", ms, wops / (ms * 1e-3));
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
