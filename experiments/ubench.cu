// ubench.cu — issue/pipe throughput microbenchmarks on B200 (not product code).
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench ubench.cu && ./ubench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;
constexpr int ILP = 8;

template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, float a, float b, uint32_t ua)
{
    float acc[ILP];
    float2 acc2[ILP];
    uint32_t u[ILP];
#pragma unroll
    for (int j = 0; j < ILP; ++j) {
        acc[j] = threadIdx.x * 1e-3f + j;
        acc2[j] = make_float2(acc[j], acc[j] + 0.5f);
        u[j] = threadIdx.x + j * 77u;
    }
    const float2 a2 = make_float2(a, a * 1.0001f), b2 = make_float2(b, b * 0.999f);
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int j = 0; j < ILP; ++j) {
            if (MODE == 0) acc[j] = __fmaf_rn(acc[j], a, b);
            if (MODE == 1) acc2[j] = __ffma2_rn(acc2[j], a2, b2);
            if (MODE == 2) u[j] = u[j] * ua + 12345u;
            if (MODE == 3) { acc[j] = __fmaf_rn(acc[j], a, b); u[j] = (u[j] >> 3) ^ (u[j] + ua); }
            if (MODE == 4) acc2[j] = __fmul2_rn(acc2[j], a2);
            if (MODE == 5) acc2[j] = __fadd2_rn(acc2[j], b2);
            if (MODE == 6) { acc2[j] = __ffma2_rn(acc2[j], a2, b2); u[j] = (u[j] >> 3) ^ (u[j] + ua); }
            if (MODE == 7) acc[j] = __fmul_rn(acc[j], a);
            if (MODE == 8) acc[j] = __fadd_rn(acc[j], b);
            if (MODE == 9) u[j] = (u[j] >> 3) ^ (u[j] + ua);      // SHF + IADD + LOP
            if (MODE == 10) acc[j] = __fmaf_rn(acc[j], acc[(j + 1) % ILP], acc[(j + 3) % ILP]);  // 3 distinct regs
            if (MODE == 11) acc2[j] = __ffma2_rn(acc2[j], acc2[(j + 1) % ILP], acc2[(j + 3) % ILP]);
            if (MODE == 12) { acc[j] = __fmaf_rn(acc[j], a, b); u[j] = u[j] * ua + 12345u; }  // FFMA + IMAD
        }
    }
    float s = 0;
#pragma unroll
    for (int j = 0; j < ILP; ++j) s += acc[j] + acc2[j].x + acc2[j].y + (float)u[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char *name, double ops_per_inner, float *out)
{
    const int blocks = 148 * 8, threads = 256;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<MODE><<<blocks, threads>>>(out, 1.0001f, 0.5f, 747796405u);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MODE><<<blocks, threads>>>(out, 1.0001f, 0.5f, 747796405u);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double warp_instr = (double)blocks * threads / 32 * ITERS * ILP * ops_per_inner;
    int clk_khz;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    // per SM per clock at max clock (upper bound on clocks => lower bound on IPC)
    const double ipc = warp_instr / (ms * 1e-3) / 148 / (clk_khz * 1e3);
    printf("%-34s %8.3f ms  %.3e warp-instr/s  >= %.2f warp-instr/clk/SM (at %d MHz)\n", name, ms,
           warp_instr / (ms * 1e-3), ipc, clk_khz / 1000);
}

int main()
{
    float *out;
    cudaMalloc(&out, 148 * 8 * 256 * sizeof(float));
    run<0>("FFMA (reg,const,const)", 1, out);
    run<10>("FFMA (3 distinct regs)", 1, out);
    run<1>("FFMA2 (reg,const,const)", 1, out);
    run<11>("FFMA2 (3 distinct regs)", 1, out);
    run<7>("FMUL", 1, out);
    run<8>("FADD", 1, out);
    run<4>("FMUL2", 1, out);
    run<5>("FADD2", 1, out);
    run<2>("IMAD", 1, out);
    run<9>("SHF+IADD+LOP (3 alu)", 3, out);
    run<3>("FFMA + 3 alu", 4, out);
    run<6>("FFMA2 + 3 alu", 4, out);
    run<12>("FFMA + IMAD", 2, out);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
