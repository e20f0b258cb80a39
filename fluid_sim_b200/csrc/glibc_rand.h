// Host-side model of the generator behind the reference's rand() calls
// (/root/reference/src/fluid_sim.cpp:378-379; the reference never calls srand, so the
// stream is glibc's default-seed TYPE_3 additive-feedback sequence).
//
// The raw word sequence obeys a[q+31] = a[q+28] + a[q] (mod 2^32) and output number q is
// a[q+31] >> 1.  A "state" is the window (a[q] .. a[q+30]).  Because the recurrence is
// linear over Z/2^32, advancing a window by n positions is a 31x31 matrix whose row m
// holds the coefficients of x^(n+m) mod (x^31 - x^28 - 1); the device noise kernel uses
// such matrices to start every warp / lane at its own stream position.
#pragma once
#include <cstdint>
#include <cstring>

namespace fsb {

constexpr int kRandDeg = 31;

struct RandPoly { uint32_t c[kRandDeg]; };                 // sum c[j] x^j
struct RandJump { uint32_t m[kRandDeg][kRandDeg]; };        // new[r] = sum_j m[r][j] * old[j]

// window at stream position 0 for srand(seed)
inline void rand_seed_window(uint32_t seed, uint32_t win[kRandDeg])
{
    int32_t r[344];
    if (seed == 0) seed = 1;
    r[0] = (int32_t)seed;
    for (int i = 1; i < 31; i++) {                            // Lehmer step by Schrage's method
        int32_t hi = r[i - 1] / 127773, lo = r[i - 1] % 127773;
        int32_t w = 16807 * lo - 2836 * hi;
        if (w < 0) w += 2147483647;
        r[i] = w;
    }
    for (int i = 31; i < 34; i++) r[i] = r[i - 31];
    for (int i = 34; i < 344; i++) r[i] = (int32_t)((uint32_t)r[i - 31] + (uint32_t)r[i - 3]);
    for (int j = 0; j < kRandDeg; j++) win[j] = (uint32_t)r[313 + j];
}

inline RandPoly poly_mul(const RandPoly &a, const RandPoly &b)
{
    uint32_t t[2 * kRandDeg - 1];
    std::memset(t, 0, sizeof(t));
    for (int i = 0; i < kRandDeg; i++)
        for (int j = 0; j < kRandDeg; j++) t[i + j] += a.c[i] * b.c[j];
    for (int d = 2 * kRandDeg - 2; d >= kRandDeg; d--) {      // x^d = x^(d-3) + x^(d-31)
        t[d - 3] += t[d];
        t[d - 31] += t[d];
    }
    RandPoly r;
    for (int i = 0; i < kRandDeg; i++) r.c[i] = t[i];
    return r;
}

inline RandPoly poly_x_pow(uint64_t n)                        // x^n mod f
{
    RandPoly result, base;
    std::memset(&result, 0, sizeof(result));
    std::memset(&base, 0, sizeof(base));
    result.c[0] = 1;
    base.c[1] = 1;
    while (n) {
        if (n & 1) result = poly_mul(result, base);
        base = poly_mul(base, base);
        n >>= 1;
    }
    return result;
}

inline RandJump jump_from_poly(RandPoly p)                    // rows: p, p*x, p*x^2, ...
{
    RandJump J;
    for (int m = 0; m < kRandDeg; m++) {
        for (int j = 0; j < kRandDeg; j++) J.m[m][j] = p.c[j];
        uint32_t top = p.c[kRandDeg - 1];                      // multiply by x and reduce
        for (int j = kRandDeg - 1; j > 0; j--) p.c[j] = p.c[j - 1];
        p.c[0] = top;
        p.c[28] += top;
    }
    return J;
}

inline RandJump jump_matrix(uint64_t n) { return jump_from_poly(poly_x_pow(n)); }

inline void apply_jump(const RandJump &J, uint32_t win[kRandDeg])
{
    uint32_t out[kRandDeg];
    for (int m = 0; m < kRandDeg; m++) {
        uint32_t acc = 0;
        for (int j = 0; j < kRandDeg; j++) acc += J.m[m][j] * win[j];
        out[m] = acc;
    }
    std::memcpy(win, out, sizeof(out));
}

inline uint32_t window_next(uint32_t win[kRandDeg])            // one sequential output (host checks)
{
    uint32_t w = win[0] + win[28];
    for (int j = 0; j < kRandDeg - 1; j++) win[j] = win[j + 1];
    win[kRandDeg - 1] = w;
    return w >> 1;
}

}  // namespace fsb
