// probe.cu - answers micro-questions about sm_100a that the kernels depend on:
//   (1) byte order of cvt.pack.sat.u8.s32.b32 (I2IP)
//   (2) issue throughput (lane-ops / clk / SM) of the instruction kinds on the hot path
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe tools/probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void pack_probe(uint32_t *out)
{
    int a = 0x11, b = 0x22; uint32_t c = 0xAABBCCDDu, d;
    asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    out[0] = d;
    int a2 = 300, b2 = -5;
    asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a2), "r"(b2), "r"(0u));
    out[1] = d;
}

enum { OP_FADD, OP_FMULC, OP_FFMA, OP_FFMAC, OP_PRMT, OP_I2F8, OP_F2I, OP_I2IP, OP_FMNMX, OP_IADD, OP_LEA, OP_VIMNMX, OP_MIX, OP_COUNT };
static const char *names[] = {"FADD r,r", "FMUL r,const", "FFMA r,r,r", "FFMA r,const,r", "PRMT", "I2F.U8 (cvt.rn.f32.u8)", "F2I (cvt.rni.s32.f32)",
                              "I2IP (cvt.pack.sat)", "FMNMX", "IADD3", "LEA", "VIMNMX", "MIX 2xFADD+FMUL+PRMT"};

template <int OP>
__global__ void __launch_bounds__(512) tput(uint32_t *out, int iters, float fa, uint32_t ua, long long *clk)
{
    float f[8]; uint32_t u[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { f[k] = fa + k + threadIdx.x; u[k] = ua + k * 77 + threadIdx.x; }
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (OP == OP_FADD) f[k] = __fadd_rn(f[k], f[(k + 1) & 7]);
                if (OP == OP_FMULC) f[k] = __fmul_rn(f[k], 1.0000001f);
                if (OP == OP_FFMA) f[k] = __fmaf_rn(f[k], f[(k + 1) & 7], f[(k + 2) & 7]);
                if (OP == OP_FFMAC) f[k] = __fmaf_rn(f[k], 1.0000001f, f[(k + 1) & 7]);
                if (OP == OP_PRMT) u[k] = __byte_perm(u[k], u[(k + 1) & 7], 0x5140);
                if (OP == OP_I2F8) { float t; asm volatile("cvt.rn.f32.u8 %0, %1;" : "=f"(t) : "r"(u[k])); u[k] = __float_as_uint(t) + 1; }
                if (OP == OP_F2I) { int t; asm volatile("cvt.rni.s32.f32 %0, %1;" : "=r"(t) : "f"(f[k])); f[k] = __int_as_float(t | 0x3f800000); }
                if (OP == OP_I2IP) asm volatile("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(u[k]) : "r"(u[k]), "r"(u[(k + 1) & 7]), "r"(u[(k + 2) & 7]));
                if (OP == OP_FMNMX) f[k] = fmaxf(f[k], f[(k + 1) & 7]);
                if (OP == OP_IADD) u[k] = u[k] + u[(k + 1) & 7];
                if (OP == OP_LEA) u[k] = (u[k] << 8) + u[(k + 1) & 7];
                if (OP == OP_VIMNMX) u[k] = (uint32_t)max((int)u[k], (int)u[(k + 1) & 7]);
                if (OP == OP_MIX) {
                    f[k] = __fadd_rn(f[k], f[(k + 1) & 7]);
                    f[(k + 3) & 7] = __fsub_rn(f[(k + 3) & 7], f[(k + 5) & 7]);
                    f[(k + 2) & 7] = __fmul_rn(f[(k + 2) & 7], 1.0000001f);
                    u[k] = __byte_perm(u[k], u[(k + 1) & 7], 0x5140);
                }
            }
        }
    }
    long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) acc += __float_as_uint(f[k]) ^ u[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(uint32_t *out, long long *clk, int sms)
{
    const int iters = 2000, threads = 512, ctas = sms * 4;  // 2048 threads per SM resident
    tput<OP><<<ctas, threads>>>(out, 10, 1.0f, 3u, clk);
    cudaDeviceSynchronize();
    tput<OP><<<ctas, threads>>>(out, iters, 1.0f, 3u, clk);
    cudaDeviceSynchronize();
    long long h[4096];
    cudaMemcpy(h, clk, sizeof(long long) * ctas, cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < ctas; i++) avg += (double)h[i];
    avg /= ctas;
    const int per = (OP == OP_MIX) ? 4 : 1;
    // each SM ran 4 CTAs concurrently for ~avg cycles
    const double lane_ops = 4.0 * threads * (double)iters * 32 * per;
    printf("%-28s %7.1f lane-ops/clk/SM   (cycles %.0f)\n", names[OP], lane_ops / avg, avg);
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    printf("device %s sm_%d%d SMs %d clock %d kHz\n", p.name, p.major, p.minor, p.multiProcessorCount, p.clockRate);
    uint32_t *out; long long *clk;
    cudaMalloc(&out, 4096 * 512 * 4); cudaMalloc(&clk, 4096 * 8);
    pack_probe<<<1, 1>>>(out);
    uint32_t h[2];
    cudaMemcpy(h, out, 8, cudaMemcpyDeviceToHost);
    printf("cvt.pack.sat.u8.s32 d(a=0x11,b=0x22,c=0xAABBCCDD) = 0x%08x ; d(a=300,b=-5,c=0) = 0x%08x\n", h[0], h[1]);
    const int sms = p.multiProcessorCount;
    run<OP_FADD>(out, clk, sms); run<OP_FMULC>(out, clk, sms); run<OP_FFMA>(out, clk, sms); run<OP_FFMAC>(out, clk, sms);
    run<OP_PRMT>(out, clk, sms); run<OP_I2F8>(out, clk, sms); run<OP_F2I>(out, clk, sms); run<OP_I2IP>(out, clk, sms);
    run<OP_FMNMX>(out, clk, sms); run<OP_IADD>(out, clk, sms); run<OP_LEA>(out, clk, sms); run<OP_VIMNMX>(out, clk, sms);
    run<OP_MIX>(out, clk, sms);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
