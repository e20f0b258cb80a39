// Velocity noise of velocityStep (src/fluid_sim.cpp:374-383) on the device.
//
// The reference draws 2 rand() numbers per cell, X then Y, i-outer / j-inner over the FULL grid,
// from glibc's never-seeded generator.  That stream is a linear recurrence over Z/2^32
// (glibc_rand.h), so any position can be reached with a 31x31 matrix product.  Work split:
//   warp  -> 8192 consecutive cells; its window is reached from the step's base window by the
//            binary decomposition of the warp index (<= 24 cooperative mat-vecs, one word per lane)
//   lane  -> 256 consecutive cells of that run; lane windows are derived one after the other with
//            the fixed 512-output jump, then each lane keeps its 31 words in registers and
//            generates sequentially with static indexing (31 outputs per rotation)
//   store -> noise is staged in shared memory and applied with the warp walking one lane's run at a
//            time, so u / v are read and written in 124-byte contiguous pieces.
#include "fsb_internal.h"
#include "noise.cuh"

namespace fsb {

namespace {

constexpr int kWarpsPerBlock = 4;
constexpr int kMatStride = kRandDeg * 32;   // transposed 31x31 matrix, rows padded to 32 words
constexpr int kStagePitch = 63;             // 62 noise floats per lane per round, odd pitch: no bank conflicts
constexpr int kWinPitch = 33;

__device__ __forceinline__ uint32_t matvec(const uint32_t *__restrict__ mt, uint32_t win, int lane)
{
    uint32_t acc = 0;
#pragma unroll
    for (int j = 0; j < kRandDeg; ++j) acc += __ldg(mt + j * 32 + lane) * __shfl_sync(0xffffffffu, win, j);
    return acc;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32) noise_kernel(float *__restrict__ u, float *__restrict__ v,
                                                                   size_t ncells, const NoiseWindow base,
                                                                   const uint32_t *__restrict__ tables)
{
    __shared__ uint32_t s_win[kWarpsPerBlock][32 * kWinPitch];
    __shared__ float s_stage[kWarpsPerBlock][32 * kStagePitch];

    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const size_t warp_global = (size_t)blockIdx.x * kWarpsPerBlock + wib;
    const size_t cell_base = warp_global * kNoiseCellsPerWarp;
    if (cell_base >= ncells) return;

    // 1. window of this warp's first cell, one word per lane
    uint32_t win = (lane < kRandDeg) ? base.w[lane] : 0u;
    for (int b = 0; b < kNoiseMaxLevels; ++b)
        if ((warp_global >> b) & 1) win = matvec(tables + (size_t)(1 + b) * kMatStride, win, lane);

    // 2. windows of the 32 lanes, 512 outputs apart
    uint32_t *sw = s_win[wib];
    for (int r = 0; r < 32; ++r) {
        if (lane < kRandDeg) sw[r * kWinPitch + lane] = win;
        win = matvec(tables, win, lane);
    }
    __syncwarp();
    uint32_t w[kRandDeg];
#pragma unroll
    for (int m = 0; m < kRandDeg; ++m) w[m] = sw[lane * kWinPitch + m];

    // 3. generate and apply: 8 rounds of 31 cells, then a tail of 8 cells
    float *stg = s_stage[wib];
    const size_t lane_cell0 = cell_base + (size_t)lane * kNoiseCellsPerLane;
    for (int round = 0; round <= 8; ++round) {
        const int ncell_round = (round < 8) ? 31 : 8;
        __syncwarp();
        if (round < 8) {
#pragma unroll
            for (int n = 0; n < 62; ++n) {
                const uint32_t nw = w[n % kRandDeg] + w[(n + 28) % kRandDeg];
                w[n % kRandDeg] = nw;
                stg[lane * kStagePitch + n] = noise_of(nw);
            }
        } else {
#pragma unroll
            for (int n = 0; n < 16; ++n) {
                const uint32_t nw = w[n % kRandDeg] + w[(n + 28) % kRandDeg];
                w[n % kRandDeg] = nw;
                stg[lane * kStagePitch + n] = noise_of(nw);
            }
        }
        __syncwarp();
        if (lane < ncell_round) {
            // eight runs at a time: all loads are issued before the first add / store
#pragma unroll 1
            for (int r0 = 0; r0 < 32; r0 += 8) {
                float uu[8], vv[8];
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    const size_t c = cell_base + (size_t)(r0 + r) * kNoiseCellsPerLane + (size_t)round * 31 + lane;
                    uu[r] = (c < ncells) ? u[c] : 0.0f;
                    vv[r] = (c < ncells) ? v[c] : 0.0f;
                }
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    const size_t c = cell_base + (size_t)(r0 + r) * kNoiseCellsPerLane + (size_t)round * 31 + lane;
                    if (c < ncells) {
                        u[c] = __fadd_rn(uu[r], stg[(r0 + r) * kStagePitch + 2 * lane]);
                        v[c] = __fadd_rn(vv[r], stg[(r0 + r) * kStagePitch + 2 * lane + 1]);
                    }
                }
            }
        }
    }
    (void)lane_cell0;
}

// ---- column-block variant: any rank's block of the grid, addressed by stream position ---------------------
// One warp per (row, 1024-cell piece of the owned columns).  The piece starts at stream position
// 2 * (i * H + j): the warp reaches it from the step's base window through the binary decomposition of
// that position (matrices x^(2^b)), lane windows are 64 outputs apart.
constexpr int kNbLevels = 40;
constexpr int kNbPitch = 65;

// consecutive rows of one piece a warp handles (one full position jump, then one fixed jump per row): fewer for long
// pieces, so that the number of warps - the only parallelism of these dependent mat-vec chains - stays the same
constexpr int nb_rows(int cpl) { return 256 / cpl; }

// CPL: cells per lane (32, 64 or 128): a warp's piece is 32 * CPL consecutive cells of one row.  The 32 lane windows of
// a piece cost 32 dependent mat-vecs whatever its length, so wide blocks use long pieces.
template <int CPL>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) noise_block_kernel(float *__restrict__ u, float *__restrict__ v,
                                                                         const GridView G, const NoiseWindow base,
                                                                         const uint32_t *__restrict__ pow2)
{
    constexpr int kPiece = 32 * CPL;
    constexpr int kNbRows = 256 / CPL;      // = nb_rows(CPL)
    constexpr int kLaneJump = (CPL == 32) ? 6 : ((CPL == 64) ? 7 : 8);   // x^(2 * CPL)
    __shared__ float s_stage[kWarpsPerBlock][32 * kNbPitch];   // also holds the lane windows before generation
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const int nown = G.j1 - G.j0;
    const int pieces = (nown + kPiece - 1) / kPiece;
    const int row_groups = (G.W + kNbRows - 1) / kNbRows;
    const long long wg = (long long)blockIdx.x * kWarpsPerBlock + wib;
    if (wg >= (long long)row_groups * pieces) return;
    const int i_first = (int)(wg / pieces) * kNbRows;
    const int jstart = G.j0 + (int)(wg % pieces) * kPiece;
    const unsigned long long pos = 2ull * ((unsigned long long)i_first * G.H + jstart);

    // window at the piece's first cell of the first row: binary decomposition of the stream position
    uint32_t row_win = (lane < kRandDeg) ? base.w[lane] : 0u;
    for (int b = 0; b < kNbLevels; ++b)
        if ((pos >> b) & 1ull) row_win = matvec(pow2 + (size_t)b * kMatStride, row_win, lane);
    uint32_t *sw = reinterpret_cast<uint32_t *>(s_stage[wib]);
    float *stg = s_stage[wib];
    for (int ri = 0; ri < kNbRows; ++ri) {
        const int i = i_first + ri;
        if (i >= G.W) break;
        uint32_t win = row_win;
        for (int r = 0; r < 32; ++r) {
            if (lane < kRandDeg) sw[r * kWinPitch + lane] = win;
            win = matvec(pow2 + (size_t)kLaneJump * kMatStride, win, lane);   // the next lane's CPL cells
        }
        __syncwarp();
        uint32_t w[kRandDeg];
#pragma unroll
        for (int m = 0; m < kRandDeg; ++m) w[m] = sw[lane * kWinPitch + m];
        __syncwarp();
        // lane l generates the cells [l * CPL, (l + 1) * CPL) of the piece, 32 cells (64 outputs) per round
#pragma unroll
        for (int round = 0; round < CPL / 32; ++round) {
            // w[] is a circular buffer: output number N overwrites index N mod 31, and 64 mod 31 = 2, so round r starts
            // at index 2 r (the rounds are unrolled: every index is a compile-time constant, w[] stays in registers)
            const int rot = (2 * round) % kRandDeg;
#pragma unroll
            for (int n = 0; n < 64; ++n) {
                const uint32_t nw = w[(n + rot) % kRandDeg] + w[(n + rot + 28) % kRandDeg];
                w[(n + rot) % kRandDeg] = nw;
                stg[lane * kNbPitch + n] = noise_of(nw);
            }
            __syncwarp();
#pragma unroll 4
            for (int r = 0; r < 32; ++r) {
                const int j = jstart + r * CPL + 32 * round + lane;
                if (j < G.j1) {
                    const size_t c = (size_t)i * G.pitch + (j - G.jbase);
                    u[c] = __fadd_rn(u[c], stg[r * kNbPitch + 2 * lane]);
                    v[c] = __fadd_rn(v[c], stg[r * kNbPitch + 2 * lane + 1]);
                }
            }
            __syncwarp();
        }
        // the same piece one row down is 2 * H outputs further on: one fixed jump (the matrix after the powers of two)
        row_win = matvec(pow2 + (size_t)kNbLevels * kMatStride, row_win, lane);
    }
}

}  // namespace

cudaError_t noise_tables_create(NoiseTables &t)
{
    const size_t words = (size_t)(1 + kNoiseMaxLevels) * kMatStride;
    std::vector<uint32_t> host(words, 0u);
    auto put = [&](int slot, const RandJump &J) {
        for (int m = 0; m < kRandDeg; ++m)
            for (int j = 0; j < kRandDeg; ++j) host[(size_t)slot * kMatStride + j * 32 + m] = J.m[m][j];
    };
    put(0, jump_matrix(2ull * kNoiseCellsPerLane));
    RandPoly p = poly_x_pow(2ull * kNoiseCellsPerWarp);
    for (int b = 0; b < kNoiseMaxLevels; ++b) {
        put(1 + b, jump_from_poly(p));
        p = poly_mul(p, p);
    }
    cudaError_t e = cudaMalloc(&t.dev, words * sizeof(uint32_t));
    if (e != cudaSuccess) return e;
    return cudaMemcpy(t.dev, host.data(), words * sizeof(uint32_t), cudaMemcpyHostToDevice);
}

cudaError_t noise_tables_create_pow2(NoiseTables &t, int H)
{
    // x^(2^b), b = 0 .. kNbLevels - 1, then x^(2 H): one grid row further on in the stream
    const size_t words = (size_t)(kNbLevels + 1) * kMatStride;
    std::vector<uint32_t> host(words, 0u);
    auto put = [&](int slot, const RandJump &J) {
        for (int m = 0; m < kRandDeg; ++m)
            for (int j = 0; j < kRandDeg; ++j) host[(size_t)slot * kMatStride + j * 32 + m] = J.m[m][j];
    };
    RandPoly p = poly_x_pow(1);
    for (int b = 0; b < kNbLevels; ++b) {
        put(b, jump_from_poly(p));
        p = poly_mul(p, p);
    }
    put(kNbLevels, jump_matrix(2ull * (unsigned long long)H));
    cudaError_t e = cudaMalloc(&t.dev_pow2, words * sizeof(uint32_t));
    if (e != cudaSuccess) return e;
    return cudaMemcpy(t.dev_pow2, host.data(), words * sizeof(uint32_t), cudaMemcpyHostToDevice);
}

void noise_tables_destroy(NoiseTables &t)
{
    if (t.dev) cudaFree(t.dev);
    if (t.dev_pow2) cudaFree(t.dev_pow2);
    t.dev = nullptr;
    t.dev_pow2 = nullptr;
}

cudaError_t launch_noise_block(float *u, float *v, const GridView &G, const NoiseWindow &base, const NoiseTables &t,
                               int zero_mode, cudaStream_t st)
{
    if (zero_mode) {
        // like the whole-grid path: the reference still adds +0.0f everywhere (halo columns are refreshed later)
        cudaError_t e = launch_add_zero(u, (size_t)G.W * G.pitch, st);
        if (e != cudaSuccess) return e;
        return launch_add_zero(v, (size_t)G.W * G.pitch, st);
    }
    if (!t.dev_pow2) return cudaErrorInvalidValue;
    const int nown = G.j1 - G.j0;
    const int cpl = nown >= 4096 ? 128 : (nown >= 2048 ? 64 : 32);
    const long long pieces = (nown + 32 * cpl - 1) / (32 * cpl);
    const int rows = nb_rows(cpl);
    const long long warps = (long long)((G.W + rows - 1) / rows) * pieces;
    const unsigned blocks = (unsigned)((warps + kWarpsPerBlock - 1) / kWarpsPerBlock);
    if (cpl == 128) noise_block_kernel<128><<<blocks, kWarpsPerBlock * 32, 0, st>>>(u, v, G, base, t.dev_pow2);
    else if (cpl == 64) noise_block_kernel<64><<<blocks, kWarpsPerBlock * 32, 0, st>>>(u, v, G, base, t.dev_pow2);
    else noise_block_kernel<32><<<blocks, kWarpsPerBlock * 32, 0, st>>>(u, v, G, base, t.dev_pow2);
    return cudaGetLastError();
}

cudaError_t launch_noise(float *u, float *v, size_t ncells, const NoiseWindow &base, const NoiseTables &t,
                         int zero_mode, cudaStream_t st)
{
    if (zero_mode) {
        // rand() == 0x3FFFFFFF: ((float)r / RAND_MAX - 0.5f) * 1e-4f == +0.0f, and the reference still adds it.
        cudaError_t e = launch_add_zero(u, ncells, st);
        if (e != cudaSuccess) return e;
        return launch_add_zero(v, ncells, st);
    }
    const size_t warps = (ncells + kNoiseCellsPerWarp - 1) / kNoiseCellsPerWarp;
    if (warps >> kNoiseMaxLevels) return cudaErrorInvalidValue;
    const unsigned blocks = (unsigned)((warps + kWarpsPerBlock - 1) / kWarpsPerBlock);
    noise_kernel<<<blocks, kWarpsPerBlock * 32, 0, st>>>(u, v, ncells, base, t.dev);
    return cudaGetLastError();
}

}  // namespace fsb
