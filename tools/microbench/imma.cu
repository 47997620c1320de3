// Microbenchmark: throughput of the legacy warp-level int8 tensor-core path (mma.sync.m16n8k32.s8) and of POPC / LOP3 on
// B200, to decide whether a dot-product formulation of the 256-bit Hamming distance (hamming = |q| + |d| - 2 q.d on
// 0/1 bytes) could beat the POPC pipe.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o imma imma.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void k_imma(int iters, int* out)
{
    int c[8][4] = {};
    uint32_t a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[j][0]), "+r"(c[j][1]), "+r"(c[j][2]), "+r"(c[j][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0 + j), "r"(b1));
    }
    int s = 0;
    for (int j = 0; j < 8; j++) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_popc(int iters, int* out)
{
    uint32_t x[8];
    for (int j = 0; j < 8; j++) x[j] = threadIdx.x * (2 * j + 1) + 12345;
    int acc = 0;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) { acc += __popc(x[j]); x[j] += acc; }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

int main()
{
    int* d;
    cudaMalloc(&d, 4 << 20);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = 148 * 8, block = 256, iters = 4096;
    float ms;
    k_imma<<<grid, block>>>(64, d);
    cudaEventRecord(e0);
    k_imma<<<grid, block>>>(iters, d);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    double ops = (double)grid * (block / 32) * iters * 8 * (16.0 * 8 * 32 * 2);
    printf("mma.sync m16n8k32 s8: %.3f ms, %.1f TOP/s dense int8\n", ms, ops / ms / 1e9);
    k_popc<<<grid, block>>>(64, d);
    cudaEventRecord(e0);
    k_popc<<<grid, block>>>(iters, d);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    double pops = (double)grid * block * iters * 8;
    printf("POPC (+2 IADD each): %.3f ms, %.2f T POPC/s = %.1f per clk per SM at 1965 MHz\n", ms, pops / ms / 1e9, pops / (ms * 1e-3) / (148 * 1.965e9));
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
