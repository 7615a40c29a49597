// sdct_math.cuh - per-thread fp32 arithmetic of the 8x8 block-DCT hot path (sm_100a).
//
// Three arithmetic definitions live here:
//   EXACT / scalar : the reference's scalar flow graph, src/simd_dct.cpp:138-172, every
//                    operation an individually rounded IEEE binary32 op (no FMA contraction)
//   EXACT / paired : the reference's AVX2 lane arithmetic, src/simd_dct.cpp:2160-2183
//   FAST / inverse : this repo's FMA-contracted forms, defined in oracle/sdct_oracle.c
// A thread owns a whole 8x8 block in registers (float b[64], statically indexed), so the
// "transposes" between the row and the column pass of the reference are index renamings.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace sdct {

// src/simd_dct.cpp:140-146
#define SDCT_KA 1.3870398453221474618216191915664f
#define SDCT_KB 1.3065629648763765278566431734272f
#define SDCT_KC 1.1758756024193587169744671046113f
#define SDCT_KD 0.78569495838710218127789736765722f
#define SDCT_KE 0.54119610014619698439972320536639f
#define SDCT_KF 0.27589937928294301233595756366937f
#define SDCT_KN 0.35355339059327376220042218105242f

#define SDCT_DEV __device__ __forceinline__

SDCT_DEV float fadd(float a, float b) { return __fadd_rn(a, b); }
SDCT_DEV float fsub(float a, float b) { return __fsub_rn(a, b); }
SDCT_DEV float fmul(float a, float b) { return __fmul_rn(a, b); }

// ---- 8-point forward, reference scalar order (:148-171); v[0], v[S], ... v[7S] ----------
template <int S>
SDCT_DEV void dct8_scalar(float *v)
{
    const float s07 = fadd(v[0], v[7 * S]), s16 = fadd(v[S], v[6 * S]);
    const float s25 = fadd(v[2 * S], v[5 * S]), s34 = fadd(v[3 * S], v[4 * S]);
    const float d07 = fsub(v[0], v[7 * S]), d61 = fsub(v[6 * S], v[S]);
    const float d25 = fsub(v[2 * S], v[5 * S]), d43 = fsub(v[4 * S], v[3 * S]);
    const float pp = fadd(s07, s34), pm = fsub(s07, s34), qq = fadd(s16, s25), qm = fsub(s16, s25);

    v[0] = fmul(SDCT_KN, fadd(pp, qq));
    v[2 * S] = fmul(SDCT_KN, fadd(fmul(SDCT_KB, pm), fmul(SDCT_KE, qm)));
    v[4 * S] = fmul(SDCT_KN, fsub(pp, qq));
    v[6 * S] = fmul(SDCT_KN, fsub(fmul(SDCT_KE, pm), fmul(SDCT_KB, qm)));
    v[S] = fmul(SDCT_KN, fsub(fadd(fsub(fmul(SDCT_KA, d07), fmul(SDCT_KC, d61)), fmul(SDCT_KD, d25)), fmul(SDCT_KF, d43)));
    v[3 * S] = fmul(SDCT_KN, fadd(fsub(fadd(fmul(SDCT_KC, d07), fmul(SDCT_KF, d61)), fmul(SDCT_KA, d25)), fmul(SDCT_KD, d43)));
    v[5 * S] = fmul(SDCT_KN, fsub(fadd(fadd(fmul(SDCT_KD, d07), fmul(SDCT_KA, d61)), fmul(SDCT_KF, d25)), fmul(SDCT_KC, d43)));
    v[7 * S] = fmul(SDCT_KN, fadd(fadd(fadd(fmul(SDCT_KF, d07), fmul(SDCT_KD, d61)), fmul(SDCT_KC, d25)), fmul(SDCT_KA, d43)));
}

// ---- 8-point forward, reference AVX2 lane order (:2160-2183); X3 = true keeps :2181's sign
template <int S, bool X3>
SDCT_DEV void dct8_paired(float *v)
{
    const float s07 = fadd(v[0], v[7 * S]), s16 = fadd(v[S], v[6 * S]);
    const float s25 = fadd(v[2 * S], v[5 * S]), s34 = fadd(v[3 * S], v[4 * S]);
    const float d07 = fsub(v[0], v[7 * S]), d61 = fsub(v[6 * S], v[S]);
    const float d25 = fsub(v[2 * S], v[5 * S]), d43 = fsub(v[4 * S], v[3 * S]);
    const float pp = fadd(s07, s34), pm = fsub(s07, s34), qq = fadd(s16, s25), qm = fsub(s16, s25);

    v[0] = fmul(SDCT_KN, fadd(pp, qq));
    v[4 * S] = fmul(SDCT_KN, fsub(pp, qq));
    v[2 * S] = fmul(SDCT_KN, fadd(fmul(SDCT_KB, pm), fmul(SDCT_KE, qm)));
    v[6 * S] = fmul(SDCT_KN, fsub(fmul(SDCT_KE, pm), fmul(SDCT_KB, qm)));
    v[S] = fmul(SDCT_KN, fadd(fsub(fmul(SDCT_KA, d07), fmul(SDCT_KC, d61)), fsub(fmul(SDCT_KD, d25), fmul(SDCT_KF, d43))));
    const float x3r = X3 ? fadd(fmul(SDCT_KA, d25), fmul(SDCT_KD, d43)) : fsub(fmul(SDCT_KA, d25), fmul(SDCT_KD, d43));
    v[3 * S] = fmul(SDCT_KN, fsub(fadd(fmul(SDCT_KC, d07), fmul(SDCT_KF, d61)), x3r));
    v[5 * S] = fmul(SDCT_KN, fadd(fadd(fmul(SDCT_KD, d07), fmul(SDCT_KA, d61)), fsub(fmul(SDCT_KF, d25), fmul(SDCT_KC, d43))));
    v[7 * S] = fmul(SDCT_KN, fadd(fadd(fmul(SDCT_KF, d07), fmul(SDCT_KD, d61)), fadd(fmul(SDCT_KC, d25), fmul(SDCT_KA, d43))));
}

// ---- 8-point forward, FAST (oracle/sdct_oracle.c: sdct_oracle_dct8_fast), unnormalised -----
template <int S>
SDCT_DEV void dct8_fast(float *v)
{
    const float s07 = fadd(v[0], v[7 * S]), s16 = fadd(v[S], v[6 * S]);
    const float s25 = fadd(v[2 * S], v[5 * S]), s34 = fadd(v[3 * S], v[4 * S]);
    const float d07 = fsub(v[0], v[7 * S]), d61 = fsub(v[6 * S], v[S]);
    const float d25 = fsub(v[2 * S], v[5 * S]), d43 = fsub(v[4 * S], v[3 * S]);
    const float pp = fadd(s07, s34), pm = fsub(s07, s34), qq = fadd(s16, s25), qm = fsub(s16, s25);

    v[0] = fadd(pp, qq);
    v[4 * S] = fsub(pp, qq);
    v[2 * S] = __fmaf_rn(SDCT_KB, pm, fmul(SDCT_KE, qm));
    v[6 * S] = __fmaf_rn(SDCT_KE, pm, -fmul(SDCT_KB, qm));
    v[S] = __fmaf_rn(-SDCT_KF, d43, __fmaf_rn(SDCT_KD, d25, __fmaf_rn(-SDCT_KC, d61, fmul(SDCT_KA, d07))));
    v[3 * S] = __fmaf_rn(SDCT_KD, d43, __fmaf_rn(-SDCT_KA, d25, __fmaf_rn(SDCT_KF, d61, fmul(SDCT_KC, d07))));
    v[5 * S] = __fmaf_rn(-SDCT_KC, d43, __fmaf_rn(SDCT_KF, d25, __fmaf_rn(SDCT_KA, d61, fmul(SDCT_KD, d07))));
    v[7 * S] = __fmaf_rn(SDCT_KA, d43, __fmaf_rn(SDCT_KC, d25, __fmaf_rn(SDCT_KD, d61, fmul(SDCT_KF, d07))));
}

// ---- 8-point inverse (oracle/sdct_oracle.c: sdct_oracle_idct8), unnormalised ----------------
// SAT = true saturates the eight outputs to [0,1] in the final add (the pixel clamp).
template <int S, bool SAT>
SDCT_DEV void idct8(float *v)
{
    const float e0 = fadd(v[0], v[4 * S]), e1 = fsub(v[0], v[4 * S]);
    const float e2 = __fmaf_rn(SDCT_KB, v[2 * S], fmul(SDCT_KE, v[6 * S]));
    const float e3 = __fmaf_rn(SDCT_KE, v[2 * S], -fmul(SDCT_KB, v[6 * S]));
    const float a0 = fadd(e0, e2), a3 = fsub(e0, e2), a1 = fadd(e1, e3), a2 = fsub(e1, e3);
    const float o0 = __fmaf_rn(SDCT_KF, v[7 * S], __fmaf_rn(SDCT_KD, v[5 * S], __fmaf_rn(SDCT_KC, v[3 * S], fmul(SDCT_KA, v[S]))));
    const float o1 = __fmaf_rn(SDCT_KD, v[7 * S], __fmaf_rn(SDCT_KA, v[5 * S], __fmaf_rn(SDCT_KF, v[3 * S], -fmul(SDCT_KC, v[S]))));
    const float o2 = __fmaf_rn(SDCT_KC, v[7 * S], __fmaf_rn(SDCT_KF, v[5 * S], __fmaf_rn(-SDCT_KA, v[3 * S], fmul(SDCT_KD, v[S]))));
    const float o3 = __fmaf_rn(SDCT_KA, v[7 * S], __fmaf_rn(-SDCT_KC, v[5 * S], __fmaf_rn(SDCT_KD, v[3 * S], -fmul(SDCT_KF, v[S]))));
    if (SAT) {
        v[0] = __saturatef(fadd(a0, o0)); v[7 * S] = __saturatef(fsub(a0, o0));
        v[S] = __saturatef(fsub(a1, o1)); v[6 * S] = __saturatef(fadd(a1, o1));
        v[2 * S] = __saturatef(fadd(a2, o2)); v[5 * S] = __saturatef(fsub(a2, o2));
        v[3 * S] = __saturatef(fsub(a3, o3)); v[4 * S] = __saturatef(fadd(a3, o3));
    } else {
        v[0] = fadd(a0, o0); v[7 * S] = fsub(a0, o0);
        v[S] = fsub(a1, o1); v[6 * S] = fadd(a1, o1);
        v[2 * S] = fadd(a2, o2); v[5 * S] = fsub(a2, o2);
        v[3 * S] = fsub(a3, o3); v[4 * S] = fadd(a3, o3);
    }
}

// ---- u8 -> f32 --------------------------------------------------------------------------
// byte k of w as an exact float: PRMT the byte into the mantissa of 2^23, subtract 2^23.
SDCT_DEV float byte_to_float(uint32_t w, int k)
{
    return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7650 + k)) - 8388608.0f;
}

// x / 255.f correctly rounded (the reference's `u8 / 255.f`, :217 / :340) without a divide:
// q0 = x*r, residual by FMA, one correction.  Verified for all 256 inputs against IEEE
// division in tests/test_math_identities.py (CPU) and tests/test_gpu_parity.py (GPU).
SDCT_DEV float div255(float x)
{
    const float r = 1.0f / 255.0f;
    const float q0 = fmul(x, r);
    const float e = __fmaf_rn(-q0, 255.0f, x);
    return __fmaf_rn(e, r, q0);
}

// ---- quantisers -> integer in [0,255] -----------------------------------------------------
// reference scalar path (:245 / :362): (u8)roundf(clamp(c*rq + 127/255, 0, 1) * 255)
SDCT_DEV uint32_t quant_scalar(float c, float rq)
{
    const float y = fmul(__saturatef(fadd(fmul(c, rq), 127.0f / 255.0f)), 255.0f);
    // roundf = half away from zero.  t = rne(y) via the 2^23 trick (0 <= y <= 255); the only
    // case where rne differs from half-away is an exact tie that rne resolved downwards.
    const float t = fadd(y, 8388608.0f);
    const float rn = fsub(t, 8388608.0f);
    uint32_t n = __float_as_uint(t) & 0x1ffu;
    if (fsub(y, rn) == 0.5f) n += 1;
    return n;
}

// reference AVX2 path (:2224): max(min(127 + cvtps_epi32(c*q), 255), 0), cvtps = nearest even
SDCT_DEV uint32_t quant_avx(float c, float q)
{
    // bits(v + 1.5*2^23) - bits(1.5*2^23) == rne(v) for |v| < 2^22 and stays monotone in v
    // beyond that, so the integer clamp below is exact for every finite v < 2^31.
    const float t = fadd(fmul(c, q), 12582912.0f);
    const int r = (int)__float_as_uint(t) - (0x4B400000 - 127);
    return (uint32_t)min(max(r, 0), 255);
}

// FAST forward quantiser and inverse pixel rounding: rne(255 * sat(.)) through 1.5*2^23
SDCT_DEV uint32_t unit_to_u8(float y01) { return __float_as_uint(__fmaf_rn(y01, 255.0f, 12582912.0f)) & 0xffu; }
SDCT_DEV uint32_t quant_fast(float c, float qs) { return unit_to_u8(__saturatef(__fmaf_rn(c, qs, 127.0f / 255.0f))); }

SDCT_DEV uint32_t pack4(uint32_t b0, uint32_t b1, uint32_t b2, uint32_t b3)
{
    return b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
}

}  // namespace sdct
