// div3_exhaustive.cu — is the ONE-correction division  q = fma(fma(-q0,y,x), yh, q0), q0 = RN(x*yh),
// yh = RN(1/y)  correctly rounded for EVERY pair of binary32 significands?  (Scale invariance: away
// from under/overflow the rounding of x/y depends only on the two 24-bit significands, so 2^23 x 2^23
// pairs with x,y in [1,2) cover all normal operands.)  Not product code.
//   ./div3_exhaustive <y_mantissa_lo> <y_mantissa_hi>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

struct Res { unsigned long long bad, bad_unfaithful_q0, n; uint32_t ex_x[8], ex_y[8]; unsigned int nex; };

__global__ void __launch_bounds__(256) kern(uint32_t ylo, uint32_t yhi, Res *res)
{
    unsigned long long bad = 0, unf = 0, n = 0;
    // grid-stride over y significands; every thread of a block walks all x significands for its y's cooperatively
    for (uint32_t ym = ylo + blockIdx.x; ym < yhi; ym += gridDim.x) {
        const float y = __uint_as_float(0x3F800000u | ym);
        const float q = rcp_approx(y);
        const float yh = __fmaf_rn(__fmaf_rn(-y, q, 1.0f), q, q);     // == __frcp_rn(y) (proven by exhaustion elsewhere)
        for (uint32_t xm = threadIdx.x; xm < (1u << 23); xm += blockDim.x) {
            const float x = __uint_as_float(0x3F800000u | xm);
            const float q0 = __fmul_rn(x, yh);
            const float q1 = __fmaf_rn(__fmaf_rn(-q0, y, x), yh, q0);
            const float ref = __fdiv_rn(x, y);
            if (__float_as_uint(q1) != __float_as_uint(ref)) {
                ++bad;
                const unsigned int slot = atomicAdd(&res->nex, 1u);
                if (slot < 8) { res->ex_x[slot] = __float_as_uint(x); res->ex_y[slot] = __float_as_uint(y); }
            }
            const int d = (int)__float_as_uint(q0) - (int)__float_as_uint(ref);
            unf += (d > 1 || d < -1);
            ++n;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        bad += __shfl_down_sync(0xFFFFFFFFu, bad, o);
        unf += __shfl_down_sync(0xFFFFFFFFu, unf, o);
        n += __shfl_down_sync(0xFFFFFFFFu, n, o);
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&res->bad, bad); atomicAdd(&res->bad_unfaithful_q0, unf); atomicAdd(&res->n, n); }
}

int main(int argc, char **argv)
{
    const uint32_t lo = argc > 1 ? strtoul(argv[1], 0, 0) : 0, hi = argc > 2 ? strtoul(argv[2], 0, 0) : (1u << 23);
    Res *r; cudaMallocManaged(&r, sizeof *r); *r = Res{};
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    // chunk launches so no single kernel runs for more than ~a second
    const uint32_t chunk = 1u << 16;
    for (uint32_t a = lo; a < hi; a += chunk) {
        const uint32_t b = (a + chunk < hi) ? a + chunk : hi;
        kern<<<148 * 8, 256>>>(a, b, r);
    }
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("y significands [%u,%u): pairs=%llu  one-correction_bad=%llu  q0_more_than_1ulp_off=%llu  (%.1f s)  %s\n", lo, hi, r->n,
           r->bad, r->bad_unfaithful_q0, ms * 1e-3, cudaGetErrorString(e));
    for (unsigned i = 0; i < r->nex && i < 8; ++i) printf("  example x=0x%08X y=0x%08X\n", r->ex_x[i], r->ex_y[i]);
    return e != cudaSuccess;
}
