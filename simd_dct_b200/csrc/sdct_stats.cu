// sdct_stats.cu - per-plane statistics of layout-C coefficient sets (the step after the hot path).
//
// Layout C exists so that each of the 64 coefficient planes is one homogeneous byte stream (src/simd_dct.cpp:1061-1069,
// :256-264): plane i of a set holds coefficient i of every block in block order.  What an entropy coder wants to
// know about such a stream before it picks a code - how many levels are zero (byte 127: the quantiser's bias), in how
// many runs they come, and how large the others are (their sum, and the OR of their magnitudes = the bit width) - is one streaming pass over the bytes: HBM-bound, 64 B read per
// block, nothing written but 64 small records per set.
//
// A CTA takes 16 KiB pieces of one plane; a thread reads 16-byte vectors, keeps its counts in registers, the warp
// reduces them with shuffles and one lane adds them to the plane's record with 64-bit atomics.  A zero run is
// counted where a zero byte follows a non-zero byte (or starts the plane's window), so pieces need only the one
// byte before them.
#include "sdct_kernels.h"

namespace sdct {

constexpr int STAT_THREADS = 256;
constexpr uint32_t STAT_PIECE = 16384;   // bytes of a plane per CTA step

struct StatParams {
    const uint8_t *coefs;
    unsigned long long *stats;   // [set][plane][4]: zeros, zero_runs, sum_abs, or_abs
    uint64_t coef_frame_stride, coef_comp_stride, coef_plane_stride;
    uint64_t first, length;      // the window of every plane: bytes [first, first + length)
    uint32_t comps, sets, pieces_per_plane;
    uint32_t aligned16;
};

__device__ __forceinline__ void count_byte(uint32_t v, uint32_t prev, uint32_t &zeros, uint32_t &runs, uint32_t &sum, uint32_t &orr)
{
    const uint32_t a = v >= 127u ? v - 127u : 127u - v;
    zeros += a == 0;
    runs += (a == 0) & (prev != 127u);
    sum += a;
    orr |= a;
}

__global__ void __launch_bounds__(STAT_THREADS) plane_stats_kernel(const StatParams p)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t total = (uint64_t)p.sets * 64 * p.pieces_per_plane;
    for (uint64_t job = blockIdx.x; job < total; job += gridDim.x) {
        const uint32_t piece = (uint32_t)(job % p.pieces_per_plane);
        const uint64_t sp = job / p.pieces_per_plane;
        const uint32_t plane = (uint32_t)(sp & 63), set = (uint32_t)(sp >> 6);
        const uint32_t frame = set / p.comps, comp = set - frame * p.comps;
        const uint8_t *base = p.coefs + (uint64_t)frame * p.coef_frame_stride + (uint64_t)comp * p.coef_comp_stride +
                              (uint64_t)plane * p.coef_plane_stride + p.first;
        const uint64_t b0 = (uint64_t)piece * STAT_PIECE;
        const uint32_t n = (uint32_t)min((uint64_t)STAT_PIECE, p.length - b0);
        const uint8_t *src = base + b0;
        uint32_t zeros = 0, runs = 0, sum = 0, orr = 0;
        if (p.aligned16) {
            // whole 16-byte vectors: four bytes per step, SIMD within the word.  x = bytes ^ 127 is zero where the level is
            // zero; ~(((x & 0x7f..) + 0x7f..) | x) has a byte's top bit set iff the byte is zero.  "The byte before this
            // vector is zero" comes from the lane below (its vector precedes this one); lane 0 reads that byte.
            static_assert(STAT_PIECE % (16 * STAT_THREADS) == 0 && STAT_PIECE / 16 / STAT_THREADS <= 4, "vectors per thread");
            const uint32_t vectors = n >> 4;
            constexpr int U = STAT_PIECE / 16 / STAT_THREADS;   // vectors per thread and piece: all loads issued up front
            uint4 q[U];
#pragma unroll
            for (int i = 0; i < U; i++) {
                const uint32_t v = i * STAT_THREADS + threadIdx.x;
                q[i] = v < vectors ? __ldg(reinterpret_cast<const uint4 *>(src) + v) : make_uint4(0, 0, 0, 0);
            }
            uint32_t edge = 0;   // lane 0: the byte before each of its vectors
            if (lane == 0) {
#pragma unroll
                for (int i = 0; i < U; i++) {
                    const uint32_t v = i * STAT_THREADS + threadIdx.x;
                    if (v < vectors && (b0 + 16ull * v) && __ldg(src + 16u * v - 1) == 127u) edge |= 0x80u << (8 * i);
                }
            }
#pragma unroll
            for (int i = 0; i < U; i++) {
                const uint32_t v = i * STAT_THREADS + threadIdx.x;
                const bool live = v < vectors;
                const uint32_t w[4] = {q[i].x, q[i].y, q[i].z, q[i].w};
                uint32_t zm[4];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t x = w[k] ^ 0x7f7f7f7fu;
                    zm[k] = ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u;   // 0x80 per zero byte
                }
                uint32_t carry = __shfl_up_sync(0xffffffffu, zm[3] >> 24, 1);
                if (lane == 0) carry = (edge >> (8 * i)) & 0x80u;
                if (live) {
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const uint32_t before = (zm[k] << 8) | carry;   // 0x80 per byte whose predecessor is zero
                        zeros += __popc(zm[k]);
                        runs += __popc(zm[k] & ~before);
                        carry = zm[k] >> 24;
                        sum = __vsadu4(w[k], 0x7f7f7f7fu) + sum;
                        orr |= __vabsdiffu4(w[k], 0x7f7f7f7fu);
                    }
                }
            }
            orr = (orr | (orr >> 8) | (orr >> 16) | (orr >> 24)) & 0xffu;
            if (threadIdx.x == 0) {   // the last piece's tail
                for (uint32_t k = vectors << 4; k < n; k++) {
                    const uint32_t prev = (b0 + k) ? (uint32_t)__ldg(src + k - 1) : 0u;
                    count_byte(__ldg(src + k), prev, zeros, runs, sum, orr);
                }
            }
        } else {
            for (uint32_t k = threadIdx.x; k < n; k += STAT_THREADS) {
                const uint32_t prev = (b0 + k) ? (uint32_t)__ldg(src + k - 1) : 0u;
                count_byte(__ldg(src + k), prev, zeros, runs, sum, orr);
            }
        }
#pragma unroll
        for (int d = 16; d; d >>= 1) {
            zeros += __shfl_xor_sync(0xffffffffu, zeros, d);
            runs += __shfl_xor_sync(0xffffffffu, runs, d);
            sum += __shfl_xor_sync(0xffffffffu, sum, d);
            orr |= __shfl_xor_sync(0xffffffffu, orr, d);
        }
        if (lane == 0) {
            unsigned long long *rec = p.stats + ((uint64_t)set * 64 + plane) * 4;
            if (zeros) atomicAdd(rec, (unsigned long long)zeros);
            if (runs) atomicAdd(rec + 1, (unsigned long long)runs);
            if (sum) atomicAdd(rec + 2, (unsigned long long)sum);
            if (orr) atomicOr(rec + 3, (unsigned long long)orr);
        }
    }
}

}  // namespace sdct

extern "C" cudaError_t sdct_k_plane_stats(const KParams *kp, void *d_stats, cudaStream_t s)
{
    using namespace sdct;
    StatParams p;
    p.coefs = kp->coefs;
    p.stats = static_cast<unsigned long long *>(d_stats);
    p.coef_frame_stride = kp->coef_frame_stride;
    p.coef_comp_stride = kp->coef_comp_stride;
    p.coef_plane_stride = kp->coef_plane_stride;
    p.first = (uint64_t)kp->row0 * kp->bpr;
    p.length = (uint64_t)kp->rows * kp->bpr;
    p.comps = kp->comps;
    p.sets = kp->comps * kp->frames;
    cudaError_t e = cudaMemsetAsync(d_stats, 0, (size_t)p.sets * 64 * 4 * sizeof(unsigned long long), s);
    if (e != cudaSuccess || p.length == 0 || p.sets == 0) return e;
    p.pieces_per_plane = (uint32_t)((p.length + STAT_PIECE - 1) / STAT_PIECE);
    p.aligned16 = kp->coef_aligned16;
    const unsigned long long total = (unsigned long long)p.sets * 64 * p.pieces_per_plane;
    const unsigned resident = 8u * (kp->sm_count ? kp->sm_count : 148u);
    plane_stats_kernel<<<(unsigned)(total < resident ? total : resident), STAT_THREADS, 0, s>>>(p);
    return cudaGetLastError();
}
