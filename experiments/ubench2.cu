// ubench2.cu — FMA-port cost of FFMA2/FMUL2/FADD2 operand forms vs scalar (not product code).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int ITERS = 4096, ILP = 8;
template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, float2 u1, float2 u2, const float2 *tab)
{
    float2 acc[ILP];
    float s[2 * ILP];
    const float2 B = tab[threadIdx.x & 7], C = tab[8 + (threadIdx.x & 7)];   // run-time register pairs
#pragma unroll
    for (int j = 0; j < ILP; ++j) { acc[j] = make_float2(threadIdx.x * 1e-3f + j, j + 0.5f); s[2*j] = acc[j].x; s[2*j+1] = acc[j].y; }
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int j = 0; j < ILP; ++j) {
            if (MODE == 0) acc[j] = __ffma2_rn(acc[j], u1, u2);
            if (MODE == 1) acc[j] = __ffma2_rn(acc[j], acc[j], u2);
            if (MODE == 2) acc[j] = __ffma2_rn(acc[j], B, u2);
            if (MODE == 3) acc[j] = __ffma2_rn(acc[j], B, C);
            if (MODE == 4) acc[j] = __fadd2_rn(acc[j], B);
            if (MODE == 5) acc[j] = __fmul2_rn(acc[j], B);
            if (MODE == 6) { s[2*j] = __fmaf_rn(s[2*j], B.x, C.x); s[2*j+1] = __fmaf_rn(s[2*j+1], B.y, C.y); }
            if (MODE == 7) { s[2*j] = __fadd_rn(s[2*j], B.x); s[2*j+1] = __fadd_rn(s[2*j+1], B.y); }
            if (MODE == 8) { s[2*j] = __fmul_rn(s[2*j], B.x); s[2*j+1] = __fmul_rn(s[2*j+1], B.y); }
            if (MODE == 9) acc[j] = __ffma2_rn(B, C, acc[j]);          // accumulate form
            if (MODE == 10) { s[2*j] = __fmaf_rn(B.x, C.x, s[2*j]); s[2*j+1] = __fmaf_rn(B.y, C.y, s[2*j+1]); }
            if (MODE == 11) { s[2*j] = __fmaf_rn(s[2*j], s[2*j+1], C.x); s[2*j+1] = __fmaf_rn(s[2*j+1], B.y, s[2*j]); }
        }
    }
    float r = 0;
#pragma unroll
    for (int j = 0; j < ILP; ++j) r += acc[j].x + acc[j].y + s[2*j] + s[2*j+1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE>
void run(const char *name, double lane_ops_per_inner, double instr_per_inner, float *out, float2 *tab)
{
    const int blocks = 148 * 8, threads = 256;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, threads>>>(out, make_float2(1.0001f, 0.9999f), make_float2(0.5f, 0.25f), tab);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MODE><<<blocks, threads>>>(out, make_float2(1.0001f, 0.9999f), make_float2(0.5f, 0.25f), tab);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double inner = (double)blocks * threads / 32 * ITERS * ILP;
    const double cyc = ms * 1e-3 * 1965e6 * 148 * 4;   // SMSP-cycles available
    printf("%-40s %7.3f ms  %.2f SMSP-cycles per instr, %.2f per scalar-equivalent op\n", name, ms,
           cyc / (inner * instr_per_inner), cyc / (inner * lane_ops_per_inner));
}
int main()
{
    float *out; float2 *tab; float2 h[16];
    for (int i = 0; i < 16; ++i) h[i] = make_float2(1.0f + i * 1e-4f, 0.5f - i * 1e-4f);
    cudaMalloc(&out, 148 * 8 * 256 * 4); cudaMalloc(&tab, sizeof h); cudaMemcpy(tab, h, sizeof h, cudaMemcpyHostToDevice);
    run<0>("FFMA2 acc*U+U", 2, 1, out, tab);
    run<1>("FFMA2 acc*acc+U", 2, 1, out, tab);
    run<2>("FFMA2 acc*B+U", 2, 1, out, tab);
    run<3>("FFMA2 acc*B+C", 2, 1, out, tab);
    run<9>("FFMA2 B*C+acc", 2, 1, out, tab);
    run<4>("FADD2 acc+B", 2, 1, out, tab);
    run<5>("FMUL2 acc*B", 2, 1, out, tab);
    run<6>("2x FFMA s*b+c", 2, 2, out, tab);
    run<10>("2x FFMA b*c+s", 2, 2, out, tab);
    run<11>("2x FFMA dependent pairs 3 var regs", 2, 2, out, tab);
    run<7>("2x FADD s+b", 2, 2, out, tab);
    run<8>("2x FMUL s*b", 2, 2, out, tab);
    printf("status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
