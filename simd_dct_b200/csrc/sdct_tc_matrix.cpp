// sdct_tc_matrix.cpp - host-side construction of the tensor-core B operand (sdct_tc.cu).
//
// K[c][p] = round(2^22 * C[u][y] * C[v][x]) with C the orthonormal 8-point DCT-II matrix,
// c = u*8+v (u vertical, v horizontal frequency), p = y*8+x.  The entries of C are the
// reference's own constants (src/simd_dct.cpp:140-146) taken as double literals, so the integer
// matrix is the same on every IEEE-754 host (two double multiplies, one scaling by 2^22, one
// rint).  oracle/sdct_oracle.c restates this function independently; tests compare the two.
#include "sdct_internal.h"

#include <cmath>
#include <cstring>
#include <vector>

namespace {

// sqrt(2) * cos(k * pi / 16), k = 0..8
const double kV[9] = {1.4142135623730950488016887242097, 1.3870398453221474618216191915664, 1.3065629648763765278566431734272,
                      1.1758756024193587169744671046113, 1.0, 0.78569495838710218127789736765722,
                      0.54119610014619698439972320536639, 0.27589937928294301233595756366937, 0.0};
const double kN = 0.35355339059327376220042218105242;  // 1/sqrt(8)

double dct_entry(int u, int y)
{
    if (u == 0) return kN;
    int m = ((2 * y + 1) * u) % 32;   // angle m * pi / 16
    if (m > 16) m = 32 - m;           // cos(2 pi - a) = cos(a)
    double sign = 1.0;
    if (m > 8) { m = 16 - m; sign = -1.0; }   // cos(pi - a) = -cos(a)
    return kN * (sign * kV[m]);
}

// IEEE binary16 bits of a small integer (|v| <= 2048: exactly representable)
uint16_t half_bits_of_int(int v)
{
    if (v == 0) return 0;
    const uint16_t sign = v < 0 ? 0x8000u : 0u;
    unsigned a = (unsigned)(v < 0 ? -v : v);
    int e = 0;
    while ((a >> (e + 1)) != 0) e++;          // a in [2^e, 2^(e+1))
    const unsigned mant = (a << (10 - e)) & 0x3ffu;   // e <= 10 here; wider values are rejected by the caller
    return (uint16_t)(sign | (unsigned)((e + 15) << 10) | mant);
}

}  // namespace

void sdct_tc_integer_matrix(int32_t K[64 * 64])
{
    for (int u = 0; u < 8; u++)
        for (int v = 0; v < 8; v++)
            for (int y = 0; y < 8; y++)
                for (int x = 0; x < 8; x++)
                    K[(u * 8 + v) * 64 + y * 8 + x] = (int32_t)std::llrint(std::ldexp(dct_entry(u, y) * dct_entry(v, x), 22));
}

void sdct_tc_operand(int layout, uint16_t out[128 * 64], bool positional)
{
    std::vector<int32_t> Kv(64 * 64);   // (no static scratch: contexts may be created concurrently)
    int32_t *K = Kv.data();
    sdct_tc_integer_matrix(K);
    memset(out, 0, 128 * 64 * sizeof(uint16_t));
    for (int n = 0; n < 128; n++) {
        const int digit = n >> 6, j = n & 63;                  // rows 0..63: K1 (hi), 64..127: K0 (lo)
        const int pos = positional ? j : sdct_tc_column_position(layout, j);
        // layout A stores (and quantises) coefficient (u, v) at position v*8+u; B and C at u*8+v
        const int u = layout == SDCT_K_LAYOUT_A ? pos & 7 : pos >> 3, v = layout == SDCT_K_LAYOUT_A ? pos >> 3 : pos & 7;
        for (int k = 0; k < 64; k++) {
            const int32_t full = K[(u * 8 + v) * 64 + k];
            const int32_t hi = (full + 512) >> 10;             // floor((full + 512) / 1024)
            const int32_t lo = full - hi * 1024;               // in [-512, 511]
            const int val = digit == 0 ? hi : lo;
            // canonical K-major, no swizzle: 8-row x 16-byte core matrices, contiguous along N, then along K
            const size_t off = (size_t)(k >> 3) * (128 * 16) + (size_t)(n >> 3) * 128 + (size_t)(n & 7) * 16 + (size_t)(k & 7) * 2;
            out[off / 2] = half_bits_of_int(val);
        }
    }
}

// ---- inverse (sdct_tc.cu: inverse_tc_kernel) ---------------------------------------------------------------------
// The dequantiser scale is folded into the matrix, so the operand depends on the LUT: E[p][i] = W[p][i] * tab[i]
// (W = unnormalised inverse basis = 8 * C[u][y] * C[v][x], tab = the inverse scale table indexed by the layout's own
// coefficient index i), scaled by the largest 2^S that keeps |round(2^S E)| < 2^20 - 512, split into two fp16 digits.
// Row n of the operand = digit n / 64 of D column j = n % 64 = pixel sdct_tc_inverse_column_pixel(j, rows).
// oracle/sdct_oracle.c (inverse_tensor_matrix) restates this independently.
int sdct_tc_inverse_operand(int layout, const float tab[64], uint16_t out[128 * 64], float off_cols[64], bool rows)
{
    std::vector<double> Ev(64 * 64);
    double *E = Ev.data();
    double emax = 0.0;
    for (int y = 0; y < 8; y++)
        for (int x = 0; x < 8; x++)
            for (int u = 0; u < 8; u++)
                for (int v = 0; v < 8; v++) {
                    const int i = layout == SDCT_K_LAYOUT_A ? v * 8 + u : u * 8 + v;
                    const double w = (dct_entry(u, y) / kN) * (dct_entry(v, x) / kN);
                    const double e = w * (double)tab[i];
                    E[(y * 8 + x) * 64 + i] = e;
                    if (std::fabs(e) > emax) emax = std::fabs(e);
                }
    int S = 0;
    const double lim = 1048576.0 - 512.0;
    if (emax > 0.0) {
        S = (int)std::floor(std::log2(lim / emax));
        while (std::ldexp(emax, S) >= lim) S--;
        while (std::ldexp(emax, S + 1) < lim) S++;
    }
    memset(out, 0, 128 * 64 * sizeof(uint16_t));
    for (int j = 0; j < 64; j++) {
        const int px = sdct_tc_inverse_column_pixel(j, rows);
        int64_t sum = 0;   // of the rounded entries: the level offset -127 must go through the same matrix as the levels
        for (int k = 0; k < 64; k++) {
            const double e = E[px * 64 + k];
            const int32_t full = (int32_t)std::llrint(std::ldexp(e, S));
            sum += full;
            const int32_t hi = (full + 512) >> 10, lo = full - hi * 1024;
            for (int digit = 0; digit < 2; digit++) {
                const int n = digit * 64 + j;
                const size_t off = (size_t)(k >> 3) * (128 * 16) + (size_t)(n >> 3) * 128 + (size_t)(n & 7) * 16 + (size_t)(k & 7) * 2;
                out[off / 2] = half_bits_of_int(digit == 0 ? hi : lo);
            }
        }
        off_cols[j] = (float)(-127.0 * std::ldexp((double)sum, -S));
    }
    return S;
}
