// Micro-benchmark of the integer SIMD / reduction instructions the SGM kernels are built
// from (sm_100a).  Prints warp-instructions per SM per clock for each.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#define ITERS 4096
template <int OP> __global__ void k(uint32_t* out, uint32_t seed, long long* clk)
{
    uint32_t a[8];
    for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 7 + i * 13;
    uint32_t b = seed * 3 + 1, c = seed ^ 0x12345;
    __shared__ uint32_t sm[1024 + 64];
    sm[threadIdx.x] = threadIdx.x;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (OP == 0) a[i] = __viaddmin_u16x2(a[i], b, c);
            if (OP == 1) a[i] = __vminu2(a[i], b);
            if (OP == 2) a[i] = __vadd2(a[i], b);
            if (OP == 3) a[i] = __byte_perm(a[i], b, 0x5432);
            if (OP == 4) a[i] = __reduce_min_sync(0xffffffffu, a[i]) + i;
            if (OP == 5) a[i] = __shfl_up_sync(0xffffffffu, a[i], 1) + 1;
            if (OP == 6) a[i] = sm[(a[i] + threadIdx.x) & 1023];
            if (OP == 7) a[i] = a[i] + b;
            if (OP == 8) a[i] = __vimin3_u16x2(a[i], b, c);
            if (OP == 9) a[i] = __viaddmax_s16x2_relu(a[i], b, c);
            if (OP == 10) a[i] = a[i] * b + c;
            if (OP == 11) a[i] = min(a[i], b) ;
            if (OP == 12) { a[i] = (a[i] & 0xffffu) * 0x10001u; }
            if (OP == 13) { if (i & 1) a[i] = __vminu2(a[i], b); else a[i] = a[i] * 1u + c; }     // ALU + FMA pipe mix
            if (OP == 14) { if (i & 1) a[i] = __viaddmin_u16x2(a[i], b, c); else asm volatile("mad.lo.u32 %0, %1, 1, %2;" : "=r"(a[i]) : "r"(a[i]), "r"(b)); }
            if (OP == 15) { uint4 v = *reinterpret_cast<const uint4*>(&sm[((a[i] & 63) * 4 + threadIdx.x * 4) & 1020]); a[i] = v.x + v.y + v.z + v.w; }
            if (OP == 16) a[i] = __vsub2(a[i], b);
            if (OP == 17) a[i] = __vimax3_s16x2_relu(a[i], b, c);
            if (OP == 18) a[i] = __vaddus2(a[i], b);
            if (OP == 19) { if (i & 1) a[i] = __vminu2(a[i], b); else a[i] = __shfl_up_sync(0xffffffffu, a[i], 1); }
            if (OP == 20) a[i] = a[i] + b + c;
        }
        b += 1;
    }
    long long t1 = clock64();
    uint32_t s = 0;
    for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}
template <int OP> void run(const char* name, int threads)
{
    uint32_t* out; long long* clk; long long h;
    cudaMalloc(&out, 148 * 8 * 1024 * 4); cudaMalloc(&clk, 8);
    k<OP><<<148, threads>>>(out, 1, clk);
    k<OP><<<148, threads>>>(out, 2, clk);
    cudaDeviceSynchronize();
    cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    double winstr = (double)ITERS * 8 * (threads / 32);
    printf("%-28s threads/SM %4d  clk %9lld  warp-instr/clk/SM %.3f  (dep-chain latency at 1 warp: %.1f clk)\n", name, threads, h,
           winstr / h, (double)h / (ITERS * 8));
    cudaFree(out); cudaFree(clk);
}
int main()
{
    const int T[3] = {32, 512, 1024};
    for (int t = 0; t < 3; ++t) {
        run<0>("VIADDMNMX.U16x2", T[t]); run<1>("VIMNMX.U16x2", T[t]); run<2>("VIADD.16x2", T[t]);
        run<3>("PRMT", T[t]); run<4>("CREDUX.MIN(+add)", T[t]); run<5>("SHFL.UP(+add)", T[t]);
        run<6>("LDS.32", T[t]); run<7>("IADD", T[t]); run<8>("VIMNMX3.U16x2", T[t]);
        run<9>("VIADDMNMX.S16x2.RELU", T[t]); run<10>("IMAD", T[t]); run<11>("IMNMX.U32", T[t]);
        run<12>("pk_dup (and+imad)", T[t]);
        run<13>("mix VIMNMX.U16x2 + IMAD(add)", T[t]); run<14>("mix VIADDMNMX + mad.lo x1", T[t]);
        run<15>("LDS.128(+3 add)", T[t]); run<16>("VSUB2", T[t]); run<17>("VIMNMX3.S16x2.RELU", T[t]);
        run<18>("VADDUS2 (saturating)", T[t]); run<19>("mix VIMNMX + SHFL", T[t]); run<20>("IADD3", T[t]);
    }
    return 0;
}
