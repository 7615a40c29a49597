// sdct_math.cuh - per-thread fp32 arithmetic of the 8x8 block-DCT hot path (sm_100a).
//
// A thread owns TWO 8x8 blocks, P and Q, held as 64 packed pairs (F2 = {P value, Q value}), and
// every operation of the two blocks is issued once as a Blackwell packed-fp32 instruction
// (add/sub/mul/fma.f32x2 -> FADD2 / FMUL2 / FFMA2).  The packed forms do not raise the fp32 rate
// (the FMA pipe still retires 128 lanes/clk/SM, profiles/r01_probe2_issue_rates.txt) but they
// halve the issue slots the arithmetic needs, which is what bounds this kernel: the conversion
// (XU pipe), packing (ALU pipe) and staging (LSU) instructions then issue in the shadow of the
// FMA pipe.  The "transposes" between the row and the column pass are index renamings.
//
// Arithmetic definitions:
//   EXACT / scalar : the reference's scalar flow graph, src/simd_dct.cpp:138-172
//   EXACT / paired : the reference's AVX2 lane arithmetic, src/simd_dct.cpp:2160-2183
//     every operation an individually rounded IEEE binary32 op, in the reference's order.
//   FAST / inverse : this repo's FMA-contracted forms, defined in oracle/sdct_oracle.c.
//
// ptxas 12.9 contracts `mul.rn.f32x2` + `add.rn.f32x2` into FFMA2 even though both carry .rn and
// --fmad=false is given (it does not do that to the scalar forms).  In the EXACT graphs every add
// that consumes a product is therefore protected (mulx/sadd/ssub below: either the product becomes an
// unfusable fma(a, k, -0.0), or the add is issued as two scalar FADDs on the halves);
// tests/test_gpu_parity.py::test_dct8_bitwise checks the result bit for bit against the oracle.
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

typedef float2 F2;

#define SDCT_OP2(name, ins)                                                                          \
    SDCT_DEV F2 name(F2 a, F2 b)                                                                     \
    {                                                                                                \
        F2 d;                                                                                        \
        asm("{.reg .b64 ra, rb, rd;\n\tmov.b64 ra, {%2,%3};\n\tmov.b64 rb, {%4,%5};\n\t" ins         \
            " rd, ra, rb;\n\tmov.b64 {%0,%1}, rd;}"                                                  \
            : "=f"(d.x), "=f"(d.y)                                                                   \
            : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));                                               \
        return d;                                                                                    \
    }
SDCT_OP2(add2, "add.rn.f32x2")
SDCT_OP2(sub2, "sub.rn.f32x2")
SDCT_OP2(mul2, "mul.rn.f32x2")
SDCT_OP2(mul2_rd, "mul.rm.f32x2")
#undef SDCT_OP2

SDCT_DEV F2 fma2(F2 a, F2 b, F2 c)
{
    F2 d;
    asm("{.reg .b64 ra, rb, rc, rd;\n\tmov.b64 ra, {%2,%3};\n\tmov.b64 rb, {%4,%5};\n\tmov.b64 rc, {%6,%7};\n\t"
        "fma.rn.f32x2 rd, ra, rb, rc;\n\tmov.b64 {%0,%1}, rd;}"
        : "=f"(d.x), "=f"(d.y)
        : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return d;
}

SDCT_DEV F2 splat(float k) { return make_float2(k, k); }
SDCT_DEV F2 mulk(F2 a, float k) { return mul2(a, splat(k)); }
SDCT_DEV F2 fmak(float k, F2 a, F2 c) { return fma2(a, splat(k), c); }   // k*a + c
// EXACT graphs, two ways to keep ptxas from contracting a packed product into the add that consumes it:
//   FP = 1    : the product is issued as fma(a, k, -0.0), which is fl(a * k) bit for bit (x + -0.0 == x for
//               every x, signed zeros included) and cannot be fused any further, so the add stays a packed
//               FADD2.  The -0.0 comes from constant memory: ptxas must not know its value, or it would
//               simplify the FMA back into the multiply it is allowed to fuse.  (scalar-order graph: -12 %)
//   FP = 0    : packed multiply, the add as two scalar FADDs on the halves.  (AVX-order graph: measured
//               3 % faster than FP = 1, its outer adds see two sums and are packed anyway)
//   FP = 2    : packed multiply; the add is issued as fma(b, 1.0, a) (a - b: fma(b, -1.0, a)) - fl(b * 1 + a) is
//               fl(a + b) bit for bit, b's multiply cannot be contracted into an FMA that already has a multiplier,
//               and the +-1.0 pairs live in constant memory, so ptxas neither knows their value nor spends a
//               register on them (FFMA2 r, c[], r runs at the full packed rate).  No scalar instruction is left:
//               half the issue slots of FP = false for these adds, and no packed/scalar switching on the FMA pipe
//               (profiles/r01_probe2_issue_rates.txt: alternating FADD2 / FADD reaches 102 of 127 lanes/clk/SM).
static __constant__ float c_neg_zero = -0.0f;
static __constant__ float2 c_one2 = {1.0f, 1.0f}, c_neg_one2 = {-1.0f, -1.0f};
template <int FP> SDCT_DEV F2 mulx(F2 a, float k) { return FP == 1 ? fma2(a, splat(k), splat(c_neg_zero)) : mulk(a, k); }
template <int FP> SDCT_DEV F2 sadd(F2 a, F2 b)
{
    return FP == 1 ? add2(a, b) : FP == 2 ? fma2(b, c_one2, a) : make_float2(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y));
}
template <int FP> SDCT_DEV F2 ssub(F2 a, F2 b)
{
    return FP == 1 ? sub2(a, b) : FP == 2 ? fma2(b, c_neg_one2, a) : make_float2(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y));
}

// first butterfly stage: packed when the inputs are not products (converted pixels), scalar when
// they are (the second pass reads the first pass's K_N * (...) results)
template <bool PROD_IN, int FP> SDCT_DEV F2 add_in(F2 a, F2 b) { return PROD_IN ? sadd<FP>(a, b) : add2(a, b); }
template <bool PROD_IN, int FP> SDCT_DEV F2 sub_in(F2 a, F2 b) { return PROD_IN ? ssub<FP>(a, b) : sub2(a, b); }

// ---- 8-point forward, reference scalar order (:148-171); v[0], v[S], ... v[7S] ----------------
// how the scalar-order graph keeps its products out of FFMA2 contraction (see mulx): measured per build
#ifndef SDCT_SCALAR_FP
#define SDCT_SCALAR_FP 2
#endif
constexpr int SFP = SDCT_SCALAR_FP;
template <int S, bool PROD_IN>
SDCT_DEV void dct8_scalar(F2 *v)
{
    const F2 s07 = add_in<PROD_IN, SFP>(v[0], v[7 * S]), s16 = add_in<PROD_IN, SFP>(v[S], v[6 * S]);
    const F2 s25 = add_in<PROD_IN, SFP>(v[2 * S], v[5 * S]), s34 = add_in<PROD_IN, SFP>(v[3 * S], v[4 * S]);
    const F2 d07 = sub_in<PROD_IN, SFP>(v[0], v[7 * S]), d61 = sub_in<PROD_IN, SFP>(v[6 * S], v[S]);
    const F2 d25 = sub_in<PROD_IN, SFP>(v[2 * S], v[5 * S]), d43 = sub_in<PROD_IN, SFP>(v[4 * S], v[3 * S]);
    const F2 pp = add2(s07, s34), pm = sub2(s07, s34), qq = add2(s16, s25), qm = sub2(s16, s25);

    v[0] = mulx<SFP>(add2(pp, qq), SDCT_KN);
    v[2 * S] = mulx<SFP>(sadd<SFP>(mulx<SFP>(pm, SDCT_KB), mulx<SFP>(qm, SDCT_KE)), SDCT_KN);
    v[4 * S] = mulx<SFP>(sub2(pp, qq), SDCT_KN);
    v[6 * S] = mulx<SFP>(ssub<SFP>(mulx<SFP>(pm, SDCT_KE), mulx<SFP>(qm, SDCT_KB)), SDCT_KN);
    // odd outputs, left to right: ((t0 op t1) op t2) op t3 - every add has a product operand
    v[S] = mulx<SFP>(ssub<SFP>(sadd<SFP>(ssub<SFP>(mulx<SFP>(d07, SDCT_KA), mulx<SFP>(d61, SDCT_KC)), mulx<SFP>(d25, SDCT_KD)), mulx<SFP>(d43, SDCT_KF)), SDCT_KN);
    v[3 * S] = mulx<SFP>(sadd<SFP>(ssub<SFP>(sadd<SFP>(mulx<SFP>(d07, SDCT_KC), mulx<SFP>(d61, SDCT_KF)), mulx<SFP>(d25, SDCT_KA)), mulx<SFP>(d43, SDCT_KD)), SDCT_KN);
    v[5 * S] = mulx<SFP>(ssub<SFP>(sadd<SFP>(sadd<SFP>(mulx<SFP>(d07, SDCT_KD), mulx<SFP>(d61, SDCT_KA)), mulx<SFP>(d25, SDCT_KF)), mulx<SFP>(d43, SDCT_KC)), SDCT_KN);
    v[7 * S] = mulx<SFP>(sadd<SFP>(sadd<SFP>(sadd<SFP>(mulx<SFP>(d07, SDCT_KF), mulx<SFP>(d61, SDCT_KD)), mulx<SFP>(d25, SDCT_KC)), mulx<SFP>(d43, SDCT_KA)), SDCT_KN);
}

// first butterfly stage of every forward graph (:148-155 / :2160-2167): exact whenever the inputs are integers
struct Stage1 { F2 s07, s16, s25, s34, d07, d61, d25, d43; };

template <int S, bool PROD_IN, int FP>
SDCT_DEV Stage1 stage1_of(const F2 *v)
{
    Stage1 a;
    a.s07 = add_in<PROD_IN, FP>(v[0], v[7 * S]); a.s16 = add_in<PROD_IN, FP>(v[S], v[6 * S]);
    a.s25 = add_in<PROD_IN, FP>(v[2 * S], v[5 * S]); a.s34 = add_in<PROD_IN, FP>(v[3 * S], v[4 * S]);
    a.d07 = sub_in<PROD_IN, FP>(v[0], v[7 * S]); a.d61 = sub_in<PROD_IN, FP>(v[6 * S], v[S]);
    a.d25 = sub_in<PROD_IN, FP>(v[2 * S], v[5 * S]); a.d43 = sub_in<PROD_IN, FP>(v[4 * S], v[3 * S]);
    return a;
}

// The same stage straight from the pixel bytes of one row of the blocks P and Q (words p0 = x0..x3, p1 = x4..x7),
// with the u8 -> f32 conversion folded into it: PRMT builds the float 2^23 + x from a byte (ALU pipe), an integer
// subtract turns its bits into those of x - 2^23 (bits = 0xCB000000 - 2x), and
//     (2^23 + a) + (b - 2^23) = a + b,      (2^23 + a) - (2^23 + b) = a - b
// are exact fp32 operations with exactly representable results - the sums and differences the reference computes
// from (float)x (:2107-2151, :2160-2167), bit for bit, with no conversion instruction (XU pipe: 1 I2F per 8 clk per
// scheduler) and no extra FMA-pipe operation.
SDCT_DEV F2 magic_plus(uint32_t wp, uint32_t wq, int k)
{
    const uint32_t sel = 0x7540u | (uint32_t)k;
    return make_float2(__uint_as_float(__byte_perm(wp, 0x4B000000u, sel)), __uint_as_float(__byte_perm(wq, 0x4B000000u, sel)));
}
SDCT_DEV F2 magic_minus(F2 m)   // bits of (x - 2^23) from the bits of (2^23 + x)
{
    const uint32_t a = __float_as_uint(m.x), b = __float_as_uint(m.y);
    return make_float2(__uint_as_float((0x61000000u - a) - a), __uint_as_float((0x61000000u - b) - b));
}
SDCT_DEV Stage1 stage1_pixels(uint32_t p0, uint32_t p1, uint32_t q0, uint32_t q1)
{
    const F2 m0 = magic_plus(p0, q0, 0), m1 = magic_plus(p0, q0, 1), m2 = magic_plus(p0, q0, 2), m3 = magic_plus(p0, q0, 3);
    const F2 m4 = magic_plus(p1, q1, 0), m5 = magic_plus(p1, q1, 1), m6 = magic_plus(p1, q1, 2), m7 = magic_plus(p1, q1, 3);
    Stage1 a;
    a.s07 = add2(m0, magic_minus(m7)); a.s16 = add2(m1, magic_minus(m6));
    a.s25 = add2(m2, magic_minus(m5)); a.s34 = add2(m3, magic_minus(m4));
    a.d07 = sub2(m0, m7); a.d61 = sub2(m6, m1); a.d25 = sub2(m2, m5); a.d43 = sub2(m4, m3);
    return a;
}

// ---- 8-point forward, reference AVX2 lane order (:2160-2183); X3 = true keeps :2181's sign -----
// how the AVX-order graph keeps its products out of FFMA2 contraction (see mulx): measured per build
#ifndef SDCT_PAIRED_FP
#define SDCT_PAIRED_FP 2
#endif
constexpr int PFP = SDCT_PAIRED_FP;
// even outputs X0, X2, X4, X6 from the four sums, odd outputs X1, X3, X5, X7 from the four differences: the two halves
// share nothing
SDCT_DEV void paired_even(F2 s07, F2 s16, F2 s25, F2 s34, F2 &x0, F2 &x2, F2 &x4, F2 &x6)
{
    const F2 pp = add2(s07, s34), pm = sub2(s07, s34), qq = add2(s16, s25), qm = sub2(s16, s25);
    x0 = mulx<PFP>(add2(pp, qq), SDCT_KN);
    x4 = mulx<PFP>(sub2(pp, qq), SDCT_KN);
    x2 = mulx<PFP>(sadd<PFP>(mulx<PFP>(pm, SDCT_KB), mulx<PFP>(qm, SDCT_KE)), SDCT_KN);
    x6 = mulx<PFP>(ssub<PFP>(mulx<PFP>(pm, SDCT_KE), mulx<PFP>(qm, SDCT_KB)), SDCT_KN);
}
// Q: the reference variant's sign defect - 0 none (the DCT), 1 = AVX2 / AVX-512VL X3 (:2181), 2 = SSE X1 (:788, :550; the
// SSE forms' other sign flips and negative factors are exact negations, so their sums are these sums bit for bit)
template <int Q>
SDCT_DEV void paired_odd(F2 d07, F2 d61, F2 d25, F2 d43, F2 &x1, F2 &x3, F2 &x5, F2 &x7)
{
    constexpr bool X3 = Q == 1, X1 = Q == 2;
    // pairwise: (t0 op t1) op (t2 op t3); the outer add sees two sums -> packed
    const F2 x1r = X1 ? sadd<PFP>(mulx<PFP>(d25, SDCT_KD), mulx<PFP>(d43, SDCT_KF)) : ssub<PFP>(mulx<PFP>(d25, SDCT_KD), mulx<PFP>(d43, SDCT_KF));
    x1 = mulx<PFP>(add2(ssub<PFP>(mulx<PFP>(d07, SDCT_KA), mulx<PFP>(d61, SDCT_KC)), x1r), SDCT_KN);
    const F2 x3r = X3 ? sadd<PFP>(mulx<PFP>(d25, SDCT_KA), mulx<PFP>(d43, SDCT_KD)) : ssub<PFP>(mulx<PFP>(d25, SDCT_KA), mulx<PFP>(d43, SDCT_KD));
    x3 = mulx<PFP>(sub2(sadd<PFP>(mulx<PFP>(d07, SDCT_KC), mulx<PFP>(d61, SDCT_KF)), x3r), SDCT_KN);
    x5 = mulx<PFP>(add2(sadd<PFP>(mulx<PFP>(d07, SDCT_KD), mulx<PFP>(d61, SDCT_KA)), ssub<PFP>(mulx<PFP>(d25, SDCT_KF), mulx<PFP>(d43, SDCT_KC))), SDCT_KN);
    x7 = mulx<PFP>(add2(sadd<PFP>(mulx<PFP>(d07, SDCT_KF), mulx<PFP>(d61, SDCT_KD)), sadd<PFP>(mulx<PFP>(d25, SDCT_KC), mulx<PFP>(d43, SDCT_KA))), SDCT_KN);
}
template <int S, int Q>
SDCT_DEV void dct8_paired_tail(const Stage1 &a, F2 *v)
{
    paired_even(a.s07, a.s16, a.s25, a.s34, v[0], v[2 * S], v[4 * S], v[6 * S]);
    paired_odd<Q>(a.d07, a.d61, a.d25, a.d43, v[S], v[3 * S], v[5 * S], v[7 * S]);
}
template <int S, bool PROD_IN, int Q>
SDCT_DEV void dct8_paired(F2 *v) { dct8_paired_tail<S, Q>(stage1_of<S, PROD_IN, PFP>(v), v); }

// ---- 8-point forward, FAST (oracle/sdct_oracle.c: sdct_oracle_dct8_fast), unnormalised ---------
// explicit FMAs; no multiply feeds a plain add, so there is nothing left for ptxas to contract
SDCT_DEV void fast_even(F2 s07, F2 s16, F2 s25, F2 s34, F2 &x0, F2 &x2, F2 &x4, F2 &x6)
{
    const F2 pp = add2(s07, s34), pm = sub2(s07, s34), qq = add2(s16, s25), qm = sub2(s16, s25);
    x0 = add2(pp, qq);
    x4 = sub2(pp, qq);
    x2 = fmak(SDCT_KB, pm, mulk(qm, SDCT_KE));
    x6 = fmak(SDCT_KE, pm, mulk(qm, -SDCT_KB));
}
SDCT_DEV void fast_odd(F2 d07, F2 d61, F2 d25, F2 d43, F2 &x1, F2 &x3, F2 &x5, F2 &x7)
{
    x1 = fmak(-SDCT_KF, d43, fmak(SDCT_KD, d25, fmak(-SDCT_KC, d61, mulk(d07, SDCT_KA))));
    x3 = fmak(SDCT_KD, d43, fmak(-SDCT_KA, d25, fmak(SDCT_KF, d61, mulk(d07, SDCT_KC))));
    x5 = fmak(-SDCT_KC, d43, fmak(SDCT_KF, d25, fmak(SDCT_KA, d61, mulk(d07, SDCT_KD))));
    x7 = fmak(SDCT_KA, d43, fmak(SDCT_KC, d25, fmak(SDCT_KD, d61, mulk(d07, SDCT_KF))));
}
template <int S>
SDCT_DEV void dct8_fast_tail(const Stage1 &a, F2 *v)
{
    fast_even(a.s07, a.s16, a.s25, a.s34, v[0], v[2 * S], v[4 * S], v[6 * S]);
    fast_odd(a.d07, a.d61, a.d25, a.d43, v[S], v[3 * S], v[5 * S], v[7 * S]);
}
template <int S>
SDCT_DEV void dct8_fast(F2 *v) { dct8_fast_tail<S>(stage1_of<S, false, 1>(v), v); }

// ---- 8-point inverse (oracle/sdct_oracle.c: sdct_oracle_idct8), unnormalised -------------------
template <int S>
SDCT_DEV void idct8(F2 *v)
{
    const F2 e0 = add2(v[0], v[4 * S]), e1 = sub2(v[0], v[4 * S]);
    const F2 e2 = fmak(SDCT_KB, v[2 * S], mulk(v[6 * S], SDCT_KE));
    const F2 e3 = fmak(SDCT_KE, v[2 * S], mulk(v[6 * S], -SDCT_KB));
    const F2 a0 = add2(e0, e2), a3 = sub2(e0, e2), a1 = add2(e1, e3), a2 = sub2(e1, e3);
    const F2 o0 = fmak(SDCT_KF, v[7 * S], fmak(SDCT_KD, v[5 * S], fmak(SDCT_KC, v[3 * S], mulk(v[S], SDCT_KA))));
    const F2 o1 = fmak(SDCT_KD, v[7 * S], fmak(SDCT_KA, v[5 * S], fmak(SDCT_KF, v[3 * S], mulk(v[S], -SDCT_KC))));
    const F2 o2 = fmak(SDCT_KC, v[7 * S], fmak(SDCT_KF, v[5 * S], fmak(-SDCT_KA, v[3 * S], mulk(v[S], SDCT_KD))));
    const F2 o3 = fmak(SDCT_KA, v[7 * S], fmak(-SDCT_KC, v[5 * S], fmak(SDCT_KD, v[3 * S], mulk(v[S], -SDCT_KF))));
    v[0] = add2(a0, o0); v[7 * S] = sub2(a0, o0);
    v[S] = sub2(a1, o1); v[6 * S] = add2(a1, o1);
    v[2 * S] = add2(a2, o2); v[5 * S] = sub2(a2, o2);
    v[3 * S] = sub2(a3, o3); v[4 * S] = add2(a3, o3);
}

// ---- u8 -> f32 -----------------------------------------------------------------------------------
// I2F.U8 with a byte selector: one XU-pipe instruction per sample, none on the FMA or ALU pipe
SDCT_DEV float byte_to_float(uint32_t w, int k) { return (float)((w >> (8 * k)) & 0xffu); }

// x / 255.f correctly rounded (the reference's `u8 / 255.f`, :217 / :340) without a divide:
// fma(x, r, x * rl) with r = fl(1/255), rl = fl(1/255 - r).  Equal to IEEE division for all 256
// inputs (tests/test_math_identities.py on the CPU, tests/test_gpu_parity.py on the GPU).
#define SDCT_R255_HI 0.003921568859368562698364257812500f
#define SDCT_R255_LO (-2.319175823606300543e-10f)
SDCT_DEV F2 div255(F2 x) { return fma2(x, splat(SDCT_R255_HI), mulk(x, SDCT_R255_LO)); }

// ---- quantisers ------------------------------------------------------------------------------------
// All of them return the quantised level as a *bit pattern*: multiplying by 2^-149 (the smallest
// denormal) makes the FMA pipe's own rounding the round-to-integer step, and the result's bits
// are the integer itself (sign-magnitude), ready for the saturating pack.  COMPLEMENTED levels
// (255 - q) are produced where the reference's "+127 after rounding" would otherwise cost an
// integer add per sample: sat_u8(128 - rne(v)) == 255 - clamp(rne(v) + 127, 0, 255).
#define SDCT_DENORM_MIN __int_as_float(0x00000001)        /*  2^-149 */
#define SDCT_NEG_DENORM_MIN __int_as_float(0x80000001u)   /* -2^-149 */
#define SDCT_128_DENORM __int_as_float(0x00000080)        /*  128 * 2^-149 */

// reference AVX2 path (:2224): max(min(127 + cvtps_epi32(c*q), 255), 0), cvtps = nearest even.
// bits = 128 - rne(fl(c*q)) (complemented level before saturation) while |c*q| < 2^23 (the result is a denormal whose
// bit pattern IS the integer); beyond that the pattern is a normal float's, same sign and larger magnitude, which the
// saturating pack clamps to the same byte.  From 2^31 the reference's cvtps_epi32 returns INT_MIN (:2224): sdct_api.cpp
// (build_tables) rejects tables under which |c*q| could get there.
SDCT_DEV F2 quant_rne_c(F2 c, F2 q) { return fma2(mul2(c, q), splat(SDCT_NEG_DENORM_MIN), splat(SDCT_128_DENORM)); }

// reference SSE stereo encoders (:1018-1024): min(255, max(0, cvtps_epi32(c * q + 127.f))) - product and sum rounded
// separately (the sum is issued as fma(p, 1.0, 127): fl(p + 127) bit for bit, and p's multiply cannot be contracted into
// it), then nearest even.  bits = the level itself; negative sums carry the sign bit and saturate to 0 in the pack.
SDCT_DEV F2 quant_sse(F2 c, F2 q) { return mulk(sadd<2>(splat(127.0f), mul2(c, q)), SDCT_DENORM_MIN); }

// reference scalar path (:245 / :362): (u8)roundf(clamp(c*rq + 127/255, 0, 1) * 255).
// roundf (half away from zero) of y in [0,255] = floor(y + 0.5) with the sum rounded DOWN, which
// cannot cross an integer.  bits = the level itself (not complemented).
SDCT_DEV F2 quant_scalar(F2 c, F2 rq)
{
    const F2 p = mul2(c, rq);
    const F2 s = make_float2(__saturatef(__fadd_rn(p.x, 127.0f / 255.0f)), __saturatef(__fadd_rn(p.y, 127.0f / 255.0f)));
    const F2 y = mulk(s, 255.0f);
    const F2 u = make_float2(__fadd_rd(y.x, 0.5f), __fadd_rd(y.y, 0.5f));
    return mul2_rd(u, splat(SDCT_DENORM_MIN));
}

// inverse: pixel level = clamp(rne(x), 0, 255): bits = rne(x) for x >= 0, sign bit set otherwise
SDCT_DEV F2 pixel_bits(F2 x) { return mulk(x, SDCT_DENORM_MIN); }

// ---- saturating pack (ALU pipe): d = (c & 0xffff) << 16 | sat_u8(a) << 8 | sat_u8(b) ---------------
SDCT_DEV uint32_t pack_sat(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// low 16 bits = [sat(lo), sat(hi)] = bytes (P, Q)
SDCT_DEV uint32_t pack_pq(F2 bits) { return pack_sat(__float_as_uint(bits.y), __float_as_uint(bits.x), 0u); }
// bytes [b0, b1, b2, b3] in memory order
SDCT_DEV uint32_t pack4_sat(float b0, float b1, float b2, float b3)
{
    return pack_sat(__float_as_uint(b1), __float_as_uint(b0), pack_sat(__float_as_uint(b3), __float_as_uint(b2), 0u));
}

}  // namespace sdct
