// exactmath.cu — exhaustive / randomised verification of short exactly-rounded sequences
// against the IEEE intrinsics (__fsqrt_rn, __frcp_rn, __fdiv_rn).  Not product code.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ float rsq_approx(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

struct Counts { unsigned long long sqrt_bad, rcp1_bad, rcp2_bad, rcpB_bad, n; };

// all floats in [lo_bits, hi_bits)
__global__ void sqrt_rcp_exhaustive(uint32_t lo_bits, uint32_t hi_bits, Counts *c)
{
    unsigned long long sb = 0, r1b = 0, r2b = 0, rBb = 0, n = 0;
    for (uint64_t b = (uint64_t)lo_bits + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < hi_bits;
         b += (uint64_t)gridDim.x * blockDim.x) {
        const float x = __uint_as_float((uint32_t)b);
        const float r = rsq_approx(x);
        const float s0 = __fmul_rn(x, r), h = __fmul_rn(r, 0.5f);
        const float e = __fmaf_rn(-s0, s0, x);
        const float s = __fmaf_rn(e, h, s0);
        const float sref = __fsqrt_rn(x);
        sb += (__float_as_uint(s) != __float_as_uint(sref));
        // reciprocal of the exact sqrt, seeded by the rsqrt approximation
        const float e1 = __fmaf_rn(-sref, r, 1.0f);
        const float r1 = __fmaf_rn(e1, r, r);
        const float e2 = __fmaf_rn(-sref, r1, 1.0f);
        const float r2 = __fmaf_rn(e2, r1, r1);
        const float rref = __frcp_rn(sref);
        r1b += (__float_as_uint(r1) != __float_as_uint(rref));
        r2b += (__float_as_uint(r2) != __float_as_uint(rref));
        // seeded by MUFU.RCP instead
        const float q = rcp_approx(sref);
        const float f1 = __fmaf_rn(-sref, q, 1.0f);
        const float q1 = __fmaf_rn(f1, q, q);
        rBb += (__float_as_uint(q1) != __float_as_uint(rref));
        ++n;
    }
    atomicAdd(&c->sqrt_bad, sb); atomicAdd(&c->rcp1_bad, r1b); atomicAdd(&c->rcp2_bad, r2b);
    atomicAdd(&c->rcpB_bad, rBb); atomicAdd(&c->n, n);
}

// rcp over all floats y in [lo,hi): variants seeded by MUFU.RCP
__global__ void rcp_exhaustive(uint32_t lo_bits, uint32_t hi_bits, Counts *c)
{
    unsigned long long b1 = 0, b2 = 0, n = 0;
    for (uint64_t b = (uint64_t)lo_bits + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < hi_bits;
         b += (uint64_t)gridDim.x * blockDim.x) {
        const float y = __uint_as_float((uint32_t)b);
        const float q = rcp_approx(y);
        const float f1 = __fmaf_rn(-y, q, 1.0f);
        const float q1 = __fmaf_rn(f1, q, q);
        const float f2 = __fmaf_rn(-y, q1, 1.0f);
        const float q2 = __fmaf_rn(f2, q1, q1);
        const float ref = __frcp_rn(y);
        b1 += (__float_as_uint(q1) != __float_as_uint(ref));
        b2 += (__float_as_uint(q2) != __float_as_uint(ref));
        ++n;
    }
    atomicAdd(&c->rcp1_bad, b1); atomicAdd(&c->rcp2_bad, b2); atomicAdd(&c->n, n);
}

__device__ __forceinline__ uint32_t hash(uint32_t v)
{
    uint32_t s = v * 747796405u + 2891336453u;
    uint32_t w = ((s >> ((s >> 28) + 4u)) ^ s) * 277803737u;
    return (w >> 22) ^ w;
}

struct DivCounts { unsigned long long q0_bad, q1_bad, q2_bad, q0_unfaithful, n; };

// random (v, y): v in (-4,4) any mantissa, y in [1,4) any mantissa, yhat = __frcp_rn(y) (exact reciprocal)
__global__ void div_random(uint64_t per_thread, uint32_t seed, DivCounts *c, int exact_rcp)
{
    unsigned long long b0 = 0, b1 = 0, b2 = 0, unf = 0;
    uint32_t st = hash(seed + blockIdx.x * blockDim.x + threadIdx.x);
    for (uint64_t i = 0; i < per_thread; ++i) {
        st = st * 747796405u + 2891336453u; const uint32_t a = hash(st);
        st = st * 747796405u + 2891336453u; const uint32_t bb = hash(st);
        // v: sign + exponent in {2^-8 .. 2^1} + random mantissa ; y: exponent {0,1}
        const uint32_t ve = 119u + (a >> 24) % 10u;
        const float v = __uint_as_float((a & 0x807FFFFFu) | (ve << 23));
        const float y = __uint_as_float((bb & 0x007FFFFFu) | ((127u + ((bb >> 23) & 1u)) << 23));
        float yh;
        if (exact_rcp) yh = __frcp_rn(y);
        else { const float q = rcp_approx(y); yh = __fmaf_rn(__fmaf_rn(-y, q, 1.0f), q, q); }
        const float q0 = __fmul_rn(v, yh);
        const float r0 = __fmaf_rn(-q0, y, v);
        const float q1 = __fmaf_rn(r0, yh, q0);
        const float r1 = __fmaf_rn(-q1, y, v);
        const float q2 = __fmaf_rn(r1, yh, q1);
        const float ref = __fdiv_rn(v, y);
        b0 += (__float_as_uint(q0) != __float_as_uint(ref));
        b1 += (__float_as_uint(q1) != __float_as_uint(ref));
        b2 += (__float_as_uint(q2) != __float_as_uint(ref));
        const int d = abs((int)(__float_as_uint(q0) & 0x7FFFFFFFu) - (int)(__float_as_uint(ref) & 0x7FFFFFFFu));
        unf += (d > 1);
    }
    atomicAdd(&c->q0_bad, b0); atomicAdd(&c->q1_bad, b1); atomicAdd(&c->q2_bad, b2);
    atomicAdd(&c->q0_unfaithful, unf); atomicAdd(&c->n, (unsigned long long)per_thread);
}

// x / D for ALL floats x in [0, xmax]: 3-op and 5-op sequences with yhat = __frcp_rn(D)
__global__ void div_const_exhaustive(float D, uint32_t hi_bits, DivCounts *c)
{
    unsigned long long b0 = 0, b1 = 0, b2 = 0, n = 0;
    const float yh = __frcp_rn(D);
    for (uint64_t b = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b <= hi_bits; b += (uint64_t)gridDim.x * blockDim.x) {
        const float v = __uint_as_float((uint32_t)b);
        const float q0 = __fmul_rn(v, yh);
        const float r0 = __fmaf_rn(-q0, D, v);
        const float q1 = __fmaf_rn(r0, yh, q0);
        const float r1 = __fmaf_rn(-q1, D, v);
        const float q2 = __fmaf_rn(r1, yh, q1);
        const float ref = __fdiv_rn(v, D);
        b0 += (__float_as_uint(q0) != __float_as_uint(ref));
        b1 += (__float_as_uint(q1) != __float_as_uint(ref));
        b2 += (__float_as_uint(q2) != __float_as_uint(ref));
        ++n;
    }
    atomicAdd(&c->q0_bad, b0); atomicAdd(&c->q1_bad, b1); atomicAdd(&c->q2_bad, b2); atomicAdd(&c->n, n);
}

int main()
{
    Counts *c; DivCounts *d;
    cudaMallocManaged(&c, sizeof *c); cudaMallocManaged(&d, sizeof *d);
    auto f2u = [](float f) { uint32_t u; memcpy(&u, &f, 4); return u; };

    *c = Counts{};
    sqrt_rcp_exhaustive<<<148 * 16, 256>>>(f2u(1.0f), f2u(16777216.0f), c);
    cudaDeviceSynchronize();
    printf("sqrt/rcp over all floats in [1,2^24): n=%llu sqrt_bad=%llu rcp(rsq-seed,1it)_bad=%llu rcp(rsq-seed,2it)_bad=%llu rcp(mufu.rcp,1it)_bad=%llu\n",
           c->n, c->sqrt_bad, c->rcp1_bad, c->rcp2_bad, c->rcpB_bad);
    *c = Counts{};
    sqrt_rcp_exhaustive<<<148 * 16, 256>>>(f2u(1.0f / 1048576.0f), f2u(1.0f), c);
    cudaDeviceSynchronize();
    printf("sqrt/rcp over all floats in [2^-20,1): n=%llu sqrt_bad=%llu rcp1_bad=%llu rcp2_bad=%llu rcpB_bad=%llu\n",
           c->n, c->sqrt_bad, c->rcp1_bad, c->rcp2_bad, c->rcpB_bad);
    *c = Counts{};
    rcp_exhaustive<<<148 * 16, 256>>>(f2u(1.0f), f2u(4.0f), c);
    cudaDeviceSynchronize();
    printf("rcp over all floats in [1,4): n=%llu 1it_bad=%llu 2it_bad=%llu\n", c->n, c->rcp1_bad, c->rcp2_bad);

    for (int exact = 1; exact >= 0; --exact) {
        *d = DivCounts{};
        div_random<<<148 * 16, 256>>>(1u << 16, 1234u + exact, d, exact);
        cudaDeviceSynchronize();
        printf("div random (yhat %s): n=%llu q0_bad=%llu q0_unfaithful=%llu q1_bad(3-op)=%llu q2_bad(5-op)=%llu\n",
               exact ? "= __frcp_rn" : "= rcp.approx+1 NR", d->n, d->q0_bad, d->q0_unfaithful, d->q1_bad, d->q2_bad);
    }
    const float divs[] = {1080.f, 180.f, 2160.f, 360.f, 720.f, 77.f, 36.f, 3.f, 1.f, 1079.f, 1081.f, 768.f, 1440.f, 4320.f, 1234.567f};
    for (float D : divs) {
        *d = DivCounts{};
        div_const_exhaustive<<<148 * 16, 256>>>(D, f2u(65536.0f), d);
        cudaDeviceSynchronize();
        printf("x/%g for all floats x in [0,65536]: n=%llu q0_bad=%llu 3-op_bad=%llu 5-op_bad=%llu\n", D, d->n, d->q0_bad,
               d->q1_bad, d->q2_bad);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
