// sdct_noalloc.cu - sdct_roundtrip_noalloc_device (include/sdct.h): the two-kernel round trip (forward kernel -> quantised levels
// in scratch memory -> inverse kernel) for callers that cannot allocate device memory at call time.  sdct_roundtrip_device takes
// that scratch from a stream-ordered pool (cudaMallocFromPoolAsync, SDCT_TENSOR jobs outside the fused kernel's reach); here it
// is a file-scope __device__ array of a fixed size, so the call path contains no cudaMalloc / cudaFree / pool routine at all.
//
// The arrays below exist once per GPU that loads this library, from module load to process exit, and are shared by EVERY call on
// that GPU: two calls whose kernels overlap on the device (different streams, different host threads, different contexts) race
// on them.  They keep the previous call's contents; nothing in them has to start at zero (the forward kernel writes every level
// the inverse kernel reads, and the operand array is rewritten whenever its (layout, LUT) key changes), so nothing is cleared.
//
// In HOST code the arrays' names are host-side stand-ins, not device addresses: the kernels live in other translation units and
// get the address that cudaGetSymbolAddress returns - looked up once per device, with that device current, return code checked.
#include "sdct_internal.h"

#include <mutex>

namespace {

constexpr int kMaxDevices = 64;

__device__ __align__(256) unsigned char g_noalloc_coefs[SDCT_NOALLOC_MAX_COEF_BYTES];
__device__ __align__(256) unsigned char g_noalloc_inv_operand[128 * 64 * 2];   // fp16 digits, UMMA canonical layout (16 KiB)

struct PerDevice {
    std::once_flag once;
    cudaError_t err = cudaErrorUnknown;
    SdctStaticScratch ss;
};
PerDevice g_dev[kMaxDevices];

// set-up, once per device: the device addresses of the two arrays
SdctStaticScratch *static_scratch(int device, cudaError_t *err)
{
    PerDevice &d = g_dev[device];
    std::call_once(d.once, [&] {
        DeviceScope scope(device);   // the lookup answers for the CURRENT device
        d.err = scope.err;
        void *p = nullptr;
        if (d.err == cudaSuccess) d.err = cudaGetSymbolAddress(&p, g_noalloc_coefs);
        d.ss.coefs = (uint8_t *)p;
        if (d.err == cudaSuccess) d.err = cudaGetSymbolAddress(&d.ss.inv_operand, g_noalloc_inv_operand);
        if (d.err != cudaSuccess) cudaGetLastError();
    });
    *err = d.err;
    return d.err == cudaSuccess ? &d.ss : nullptr;
}

}  // namespace

extern "C" int sdct_roundtrip_noalloc_device(sdct_context *ctx, const sdct_job *job, const void *d_pixels_in, const float *lut,
                                             void *d_pixels_out, void *stream)
{
    if (!ctx) return fail(SDCT_INVALID_PARAMETER, "context is null");
    if (ctx->device < 0 || ctx->device >= kMaxDevices) return fail(SDCT_NOT_SUPPORTED, "device %d: the allocation-free entry point serves devices 0..%d", ctx->device, kMaxDevices - 1);
    cudaError_t e;
    SdctStaticScratch *ss = static_scratch(ctx->device, &e);
    if (!ss) return fail(SDCT_CUDA_ERROR, "cudaGetSymbolAddress (static scratch of device %d) failed: %s", ctx->device, cudaGetErrorString(e));
    return sdct_roundtrip_two_pass_static(ctx, job, d_pixels_in, lut, d_pixels_out, stream, ss, SDCT_NOALLOC_MAX_COEF_BYTES);
}
