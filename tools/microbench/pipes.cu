// Microbenchmark: issue rates (thread-instructions per clock per SM) of the integer instructions the extractor kernels
// are made of, alone and in pairs, to tell which pipe each one occupies on B200 (sm_100a):
//   IDP.4A / IDP.2A (dp4a / dp2a), IMAD, IMAD.HI, PRMT, SHF, LOP3, IADD3, VABSDIFF4, VIMNMX3.S16x2, LDS.
// 8 independent dependency chains per thread, 1024 threads per SM resident.  clock64() around the loop, SM clock rate
// therefore cancels.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pipes pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

enum Op { IDP4A, IDP2A, IMAD_, IMADHI, PRMT_, SHF_, LOP3_, IADD3_, VABS4, VMNMX3, MIX_IDP_PRMT, MIX_IDP_IMAD, MIX_IMAD_PRMT, MIX_VABS_PRMT,
          MIX_VMNMX_PRMT, MIX_VMNMX_IMAD, MIX_IDP_VABS, N_OPS };
static const char* NAMES[N_OPS] = {"IDP.4A", "IDP.2A", "IMAD", "IMAD.HI", "PRMT", "SHF", "LOP3", "IADD3", "VABSDIFF4", "VIMNMX3.S16x2",
                                   "IDP.4A+PRMT (1:1)", "IDP.4A+IMAD (1:1)", "IMAD+PRMT (1:1)", "VABSDIFF4+PRMT (1:1)", "VIMNMX3+PRMT (1:1)",
                                   "VIMNMX3+IMAD (1:1)", "IDP.4A+VABSDIFF4 (1:1)"};

template <int OP>
__device__ __forceinline__ void body(uint32_t& x, uint32_t& y, uint32_t k)
{
    if (OP == IDP4A) x = __dp4a(x, k, x);
    if (OP == IDP2A) x = __dp2a_lo(x, k, x);
    if (OP == IMAD_) x = x * k + x;
    if (OP == IMADHI) x = __umulhi(x, k) + x;
    if (OP == PRMT_) x = __byte_perm(x, k, 0x6130);
    if (OP == SHF_) x = __funnelshift_r(x, k, 8);
    if (OP == LOP3_) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(k), "r"(y));
    if (OP == IADD3_) asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(k));
    if (OP == VABS4) x = __vabsdiffu4(x, k);
    if (OP == VMNMX3) x = __vimin3_s16x2(x, k, y);
    if (OP == MIX_IDP_PRMT) { x = __dp4a(x, k, x); y = __byte_perm(y, k, 0x6130); }
    if (OP == MIX_IDP_IMAD) { x = __dp4a(x, k, x); y = y * k + y; }
    if (OP == MIX_IMAD_PRMT) { x = x * k + x; y = __byte_perm(y, k, 0x6130); }
    if (OP == MIX_VABS_PRMT) { x = __vabsdiffu4(x, k); y = __byte_perm(y, k, 0x6130); }
    if (OP == MIX_VMNMX_PRMT) { x = __vimin3_s16x2(x, k, k + 1); y = __byte_perm(y, k, 0x6130); }
    if (OP == MIX_VMNMX_IMAD) { x = __vimin3_s16x2(x, k, k + 1); y = y * k + y; }
    if (OP == MIX_IDP_VABS) { x = __dp4a(x, k, x); y = __vabsdiffu4(y, k); }
}

template <int OP>
__global__ void __launch_bounds__(256) k_rate(int iters, uint32_t seed, long long* clocks, uint32_t* sink)
{
    uint32_t x[8], y[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { x[j] = threadIdx.x * (2 * j + 1) + seed; y[j] = x[j] ^ 0x5a5a5a5au; }
    const uint32_t k = seed * 2654435761u | 1u;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
            for (int j = 0; j < 8; j++) body<OP>(x[j], y[j], k);
    }
    const long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s ^= x[j] ^ y[j];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clocks[blockIdx.x] = t1 - t0;
}

template <int OP>
static void run(long long* dClk, uint32_t* dSink)
{
    const int grid = 148 * 4, block = 256, iters = 2048;  // 4 CTAs x 256 threads = 32 warps per SM
    k_rate<OP><<<grid, block>>>(16, 1, dClk, dSink);
    k_rate<OP><<<grid, block>>>(iters, 3, dClk, dSink);
    cudaDeviceSynchronize();
    static long long h[148 * 4];
    cudaMemcpy(h, dClk, sizeof h, cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < grid; i++) avg += (double)h[i];
    avg /= grid;
    const int perBody = OP >= MIX_IDP_PRMT ? 2 : 1;
    const double threadInsts = 4.0 * block * (double)iters * 32 * perBody;  // per SM (4 resident CTAs run concurrently)
    printf("%-24s %7.1f thread-insts/clk/SM (%.2f warp-insts/clk/SMSP)\n", NAMES[OP], threadInsts / avg, threadInsts / avg / 128.0);
}

int main()
{
    long long* dClk;
    uint32_t* dSink;
    cudaMalloc(&dClk, 148 * 4 * sizeof(long long));
    cudaMalloc(&dSink, 148 * 4 * 256 * 4);
    run<IDP4A>(dClk, dSink); run<IDP2A>(dClk, dSink); run<IMAD_>(dClk, dSink); run<IMADHI>(dClk, dSink);
    run<PRMT_>(dClk, dSink); run<SHF_>(dClk, dSink); run<LOP3_>(dClk, dSink); run<IADD3_>(dClk, dSink);
    run<VABS4>(dClk, dSink); run<VMNMX3>(dClk, dSink);
    run<MIX_IDP_PRMT>(dClk, dSink); run<MIX_IDP_IMAD>(dClk, dSink); run<MIX_IMAD_PRMT>(dClk, dSink); run<MIX_VABS_PRMT>(dClk, dSink);
    run<MIX_VMNMX_PRMT>(dClk, dSink); run<MIX_VMNMX_IMAD>(dClk, dSink); run<MIX_IDP_VABS>(dClk, dSink);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
