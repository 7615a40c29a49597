// probe2.cu - issue/pipe throughput of packed fp32 (FADD2/FMUL2/FFMA2) against scalar fp32 and the
// ALU-pipe instructions the DCT kernels mix with them, timed with CUDA events (wall clock) so the
// figure does not depend on what clock64() counts.  ops/clk/SM is quoted at the 1.965 GHz max clock;
// the nvidia-smi clock during the run is printed by the caller.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe2 tools/probe2.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define OP2(name, ins) __device__ __forceinline__ float2 name(float2 a, float2 b){ float2 d; \
 asm volatile("{.reg .b64 ra, rb, rd; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; " ins " rd, ra, rb; mov.b64 {%0,%1}, rd;}" : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y)); return d;}
OP2(add2, "add.rn.f32x2") OP2(mul2, "mul.rn.f32x2")
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c){ float2 d;
 asm volatile("{.reg .b64 ra, rb, rc, rd; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mov.b64 rc, {%6,%7}; fma.rn.f32x2 rd, ra, rb, rc; mov.b64 {%0,%1}, rd;}" : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y)); return d;}

enum { P_FADD, P_FADD2, P_FMUL2C, P_FFMA2C, P_FFMA2R, P_FFMAC, P_PRMT, P_IADD, P_I2IP, P_FADD2_PRMT, P_FADD2_IADD, P_FADD_PRMT, P_FADD2_FADD, P_FADD2x2_PRMT, P_VIADDMNMX, P_COUNT };
static const char *names[] = {"FADD", "FADD2", "FMUL2 r,const", "FFMA2 r,const,r", "FFMA2 r,r,r", "FFMA r,const,r", "PRMT", "IADD3", "I2IP",
  "1 FADD2 + 1 PRMT", "1 FADD2 + 1 IADD3", "1 FADD + 1 PRMT", "1 FADD2 + 1 FADD", "2 FADD2 + 1 PRMT", "VIADDMNMX (add+max)"};
static const int insts[] = {1,1,1,1,1,1,1,1,1,2,2,2,2,3,1};
static const int flops[] = {1,2,2,2,2,1,0,0,0,2,2,1,3,4,0};

template <int OP>
__global__ void __launch_bounds__(256) tput(uint32_t *out, int iters, float fa, uint32_t ua)
{
    float2 f[8]; uint32_t u[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { f[k] = make_float2(fa + k + threadIdx.x, fa * k + 1.f); u[k] = ua + k * 77 + threadIdx.x; }
    const float2 c = make_float2(1.0000001f, 1.0000001f);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const int k1 = (k + 1) & 7, k2 = (k + 2) & 7;
                if (OP == P_FADD) f[k].x = __fadd_rn(f[k].x, f[k1].x);
                if (OP == P_FADD2) f[k] = add2(f[k], f[k1]);
                if (OP == P_FMUL2C) f[k] = mul2(f[k], c);
                if (OP == P_FFMA2C) f[k] = fma2(f[k], c, f[k1]);
                if (OP == P_FFMA2R) f[k] = fma2(f[k], f[k1], f[k2]);
                if (OP == P_FFMAC) f[k].x = __fmaf_rn(f[k].x, 1.0000001f, f[k1].x);
                if (OP == P_PRMT) u[k] = __byte_perm(u[k], u[k1], 0x5140);
                if (OP == P_IADD) u[k] = u[k] + u[k1];
                if (OP == P_I2IP) asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(u[k]) : "r"(u[k]), "r"(u[k1]), "r"(u[k2]));
                if (OP == P_FADD2_PRMT) { f[k] = add2(f[k], f[k1]); u[k] = __byte_perm(u[k], u[k1], 0x5140); }
                if (OP == P_FADD2_IADD) { f[k] = add2(f[k], f[k1]); u[k] = u[k] + u[k1]; }
                if (OP == P_FADD_PRMT) { f[k].x = __fadd_rn(f[k].x, f[k1].x); u[k] = __byte_perm(u[k], u[k1], 0x5140); }
                if (OP == P_FADD2_FADD) { f[k] = add2(f[k], f[k1]); u[k] = __float_as_uint(__fadd_rn(__uint_as_float(u[k]), __uint_as_float(u[k1]))); }
                if (OP == P_FADD2x2_PRMT) { f[k] = add2(f[k], f[k1]); f[k2] = mul2(f[k2], c); u[k] = __byte_perm(u[k], u[k1], 0x5140); }
                if (OP == P_VIADDMNMX) u[k] = (uint32_t)max((int)u[k] + (int)u[k1], 3);
            }
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) acc += __float_as_uint(f[k].x) ^ __float_as_uint(f[k].y) ^ u[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int OP>
void run(uint32_t *out, int sms)
{
    const int iters = 4000, threads = 256, ctas = sms * 8;  // 2048 threads per SM resident
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    tput<OP><<<ctas, threads>>>(out, 200, 1.0f, 3u);
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        tput<OP><<<ctas, threads>>>(out, iters, 1.0f, 3u);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double groups = 8.0 * threads * (double)iters * 32;   // per SM: op-groups executed by all its threads
    const double clk = best * 1e-3 * 1.965e9;
    printf("%-22s %7.1f thread-instr/clk/SM  %7.1f flop-lanes/clk/SM   (%.3f ms)\n", names[OP], groups * insts[OP] / clk, groups * flops[OP] / clk, best);
}

// plain u8 stream copy, 16 B per thread per access: what a pure pass-through kernel reaches
__global__ void __launch_bounds__(256) copy16(const uint4 *__restrict__ in, uint4 *__restrict__ out, size_t n)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i + 3 * stride < n; i += 4 * stride) {
        uint4 a = in[i], b = in[i + stride], c = in[i + 2 * stride], d = in[i + 3 * stride];
        out[i] = a; out[i + stride] = b; out[i + 2 * stride] = c; out[i + 3 * stride] = d;
    }
    for (; i < n; i += stride) out[i] = in[i];
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    printf("device %s sm_%d%d SMs %d\n", p.name, p.major, p.minor, p.multiProcessorCount);
    uint32_t *out;
    cudaMalloc(&out, 4096 * 512 * 4);
    const int sms = p.multiProcessorCount;
    run<P_FADD>(out, sms); run<P_FADD2>(out, sms); run<P_FMUL2C>(out, sms); run<P_FFMA2C>(out, sms); run<P_FFMA2R>(out, sms);
    run<P_FFMAC>(out, sms); run<P_PRMT>(out, sms); run<P_IADD>(out, sms); run<P_I2IP>(out, sms); run<P_FADD2_PRMT>(out, sms);
    run<P_FADD2_IADD>(out, sms); run<P_FADD_PRMT>(out, sms); run<P_FADD2_FADD>(out, sms); run<P_FADD2x2_PRMT>(out, sms); run<P_VIADDMNMX>(out, sms);
    for (size_t mb : {128, 256, 1024}) {
        const size_t bytes = mb << 20, n = bytes / 16;
        uint4 *a, *b; cudaMalloc(&a, bytes); cudaMalloc(&b, bytes); cudaMemset(a, 1, bytes);
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int ctas_per_sm : {4, 8, 16}) {
            float best = 1e30f;
            for (int rep = 0; rep < 6; rep++) {
                cudaEventRecord(e0);
                copy16<<<sms * ctas_per_sm, 256>>>(a, b, n);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep && ms < best) best = ms;
            }
            printf("copy16 %4zu MiB grid %2d/SM: %.1f GB/s (read+write)\n", mb, ctas_per_sm, 2.0 * bytes / (best * 1e-3) / 1e9);
        }
        cudaFree(a); cudaFree(b);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
