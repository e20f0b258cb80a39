// Shared by noise.cu (the stand-alone noise kernels), chain.cu (the noise as a phase of the chained launch) and the CPU
// check of the phase's stream arithmetic (tests/noise_phase_check.cpp, compiled with g++).
#pragma once
#include <cstdint>
#include "glibc_rand.h"

#if defined(__CUDACC__)
#define FSB_HD __host__ __device__ __forceinline__
#else
#define FSB_HD inline
#endif

namespace fsb {

#if defined(__CUDACC__)
// one rand() output -> the value velocityStep adds to a velocity component (src/fluid_sim.cpp:378-379):
// ((float)rand() / RAND_MAX - 0.5f) * NOISE_SCALE, with (float)RAND_MAX == 2^31
__device__ __forceinline__ float noise_of(uint32_t word)
{
    const int r = (int)(word >> 1);                       // rand() output
    const float f = __fmul_rn((float)r, 4.656612873077393e-10f);
    return __fmul_rn(__fsub_rn(f, 0.5f), 1e-4f);
}
#endif

// ---- the noise as a phase of the chained launch: stream positions of a block of cw columns ---------------------------
// Cell (i, j) uses the outputs 2 (i H + j) and 2 (i H + j) + 1 of the step.  A warp of the CTA takes 32 consecutive
// rows of the block, lane = row; each lane reaches the window of (its row, first column of the block) from the step's base
// window through the bits of the block index, of the warp index and of the lane, generates the row's 2 cw outputs
// sequentially, and jumps on to the same columns 32 * kNzWarps rows further down.  Jump polynomials, 32 words each:
constexpr int kNzLanePoly = 0;               // [0..4]   x^(2 H 2^b)                    one row, two rows, ...
constexpr int kNzWarpPoly = 5;               // [5..8]   x^(64 H 2^b)                   32 rows, 64 rows, ...
constexpr int kNzStridePoly = 9;             // [9]      x^(64 H kNzWarps - 2 cw)       end of a row piece -> the same piece, the warp's next rows
constexpr int kNzBlockPoly = 10;             // [10..19] x^(2 cw 2^b)                   one block, two blocks, ...
constexpr int kNzPolys = 20;
constexpr int kNzWarps = 10;                 // warps of a CTA of the chained launch
static_assert(kNzWarps <= 16, "four warp-index bits");

// w[]: window a[q .. q+30] in natural order -> (take ? the window k outputs further on : unchanged), p = coefficients of
// x^k mod (x^31 - x^28 - 1).  new[m] = sum_j p[j] a[q+m+j]: a correlation with the window extended by 30 sequential
// outputs, on the thread's own registers (961 multiply-adds; the 31 coefficients are the only memory operands).
FSB_HD void nz_jump(uint32_t (&w)[kRandDeg], const uint32_t *p, const bool take)
{
    uint32_t ext[2 * kRandDeg - 1];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int j = 0; j < kRandDeg; ++j) ext[j] = w[j];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int k = 0; k < kRandDeg - 1; ++k) ext[kRandDeg + k] = ext[k] + ext[k + 28];   // a[q+31+k] = a[q+k] + a[q+k+28]
    uint32_t nw[kRandDeg];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int m = 0; m < kRandDeg; ++m) nw[m] = 0u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int j = 0; j < kRandDeg; ++j) {
        const uint32_t c = p[j];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int m = 0; m < kRandDeg; ++m) nw[m] += c * ext[m + j];
    }
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int m = 0; m < kRandDeg; ++m) w[m] = take ? nw[m] : w[m];
}

// the table of jump polynomials for a grid of H columns cut into blocks of cw: host[kNzPolys * 32]
inline void nz_fill_polys(int H, int cw, uint32_t *host)
{
    for (int i = 0; i < kNzPolys * 32; ++i) host[i] = 0u;
    auto put = [&](int slot, unsigned long long k) {
        const RandPoly q = poly_x_pow(k);
        for (int j = 0; j < kRandDeg; ++j) host[(size_t)slot * 32 + j] = q.c[j];
    };
    const unsigned long long row = 2ull * (unsigned long long)H;          // outputs per grid row
    for (int bit = 0; bit < 5; ++bit) put(kNzLanePoly + bit, row << bit);
    for (int bit = 0; bit < 4; ++bit) put(kNzWarpPoly + bit, (32ull * row) << bit);
    put(kNzStridePoly, 32ull * kNzWarps * row - 2ull * (unsigned long long)cw);
    for (int bit = 0; bit < 10; ++bit) put(kNzBlockPoly + bit, (2ull * (unsigned long long)cw) << bit);
}

}  // namespace fsb
