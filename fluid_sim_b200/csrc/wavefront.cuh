// Exact-order Gauss-Seidel / SOR sweeps as a systolic strip pipeline (sm_100a).
//
// What it replaces: the lexicographic, in-place relaxation loops of the reference,
//   diffuse()  src/fluid_sim.cpp:50-61   (SOR, omega 1.5, `k < 5`)
//   project()  src/fluid_sim.cpp:319-331 (Gauss-Seidel, `k < 20`)
// including the setBoundary() call that follows every sweep (:60, :330).  The update order
// of those loops decides the result, so it is reproduced exactly:
//
//   * a warp owns a strip of 32 "global lanes" and S consecutive sweeps ("levels").  Lane g keeps
//     x[0..S] in registers, x[k] being the level-k value of cell (i, j) = (t - 32 - k - g, g - k) at
//     time step t (in the warp's own frame).  Each step every lane moves one row down; level k
//     trails level k-1 by one row and one column.  With that skew all four neighbours a sweep
//     needs are values of step t-1: own registers (up / right) or the registers of lane g-1
//     (left / down), read as float4s from the warp's state buffer in shared memory.  Hence S sweeps
//     are in flight per warp.
//   * the right-hand side (divergence or diffusion source) travels with the cell through the same
//     buffer (two steps per level).
//   * stages: the sweeps of a pass are split over `nstg` warps per strip.  Stage g is the same
//     program shifted by g*S columns and 2*g*S time steps; it takes its level-0 stream (x and
//     right-hand side, lane-aligned) from the shared-memory ring stage g-1 writes its level-S
//     stream to.  Stage 0 is fed from HBM through cp.async row tiles, the last stage writes the
//     result rows back.  All sweeps of a pass therefore read and write each cell of HBM once.
//   * strips: lane 0 takes its left neighbours from the previous strip of the same stage: through a
//     shared-memory ring inside a CTA, through a ring in global memory between CTAs (units of 8 steps,
//     moved by two transfer warps; every 16-byte word carries its own tag, so there are no flags and no
//     fences on that path), and through the same kind of ring in the neighbouring GPU's memory between
//     ranks.  The launch is cooperative so that every CTA of a chain is resident while its right
//     neighbour waits for it.
//   * there is no CTA-wide barrier in the main loop: every compute warp publishes the last step it
//     has completed in shared memory and, before each group of 4 steps, checks that its producers
//     (left strip, previous stage) are far enough ahead and its consumers (right strip, next stage)
//     have freed the ring slots it is about to overwrite.  Two extra warps per CTA move the CTA-edge
//     rings to and from global memory, so compute warps never wait on L2 latency.
//   * wall cells are evaluated in the pipeline with the value setBoundary() would have left in the
//     array at the moment the reference reads it (warp-uniform fix-ups, only near walls).
//
// All arithmetic uses explicit round-to-nearest intrinsics: no FMA contraction (F7).
#pragma once
#include "fsb_internal.h"
#include <cstdlib>

#include <cstdio>

namespace fsb {

namespace {

constexpr int kToff = 32;        // time offset: lane g sits on row t - kToff - g at level 0
constexpr int kTileRows = 64;    // rows of an HBM staging ring (input tiles, result tile)
constexpr int kTileFloats = kTileRows * 32;
// Compile-time switches (defaults are what ships; the others are kept as measured experiments):
//   WF_GROUP (4)          steps a warp runs between two progress checks
//   WF_XFER (8)           steps per ring hand-over between CTAs: 4 saturates the transfer warps, 16 adds latency
//   WF_FLUSH_OFFLOAD (0)  result write-back by the export warp instead of the last-stage warp: 8 % slower
//   WF_CLUSTER_MODE (0)   2 = hand-over through distributed shared memory inside thread-block clusters: slower
//   WF_PROFILE            in-kernel cycle counters printed per warp (development builds only)
//   WF_RELAXED_FENCES     volatile accesses instead of acquire/release instructions on the shared-memory counters
// Result rows are written back by the last-stage warp itself (0) or by the export helper warp (1).
// Measured on B200 at 8192^2: the helper variant loses about 8 % (it competes for issue slots with the
// compute warps of its scheduler), so the default is 0.
#ifndef WF_FLUSH_OFFLOAD
#define WF_FLUSH_OFFLOAD 0
#endif
#ifndef WF_GROUP
#define WF_GROUP 4
#endif
#ifndef WF_XFER
#define WF_XFER 8           // steps per ring hand-over between CTAs (4, 8 or 16)
#endif
#ifndef WF_CLUSTER_MODE
#define WF_CLUSTER_MODE 0   // 0: no cluster code, 1: cluster barriers only, 2: hand-over through distributed shared memory
#endif
constexpr int kGroup = WF_GROUP;  // steps a warp runs between two progress checks (4, 8 or 16)
constexpr int kEdgeDepth = 64;   // steps held by the CTA-edge rings (four batches)
constexpr int kMidDepth = 16;    // steps held by the rings between strips of a CTA
constexpr int kBatchBias = 64;   // keeps published batch numbers positive (later stages start at t < 0)
constexpr int kMaxStages = 16;

template <int S>
struct WfLayout {
    static constexpr int XW = (S + 1 + 3) & ~3;   // x[0..S], padded to float4
    static constexpr int DW = (S + 3) & ~3;       // rhs levels 0..S-1
    static constexpr int EXW = XW + DW;           // floats one lane hands to the next strip per step
    static constexpr int kBoundaryFloats = kWfSlots * kWfBatch * EXW * 2;   // x2: every float travels with a tag
    // per-lane strides of a warp's own state buffers: an odd number of 16-byte chunks keeps the
    // 128-bit shared-memory accesses of a quarter warp on distinct banks
    static constexpr int XP = ((XW / 4) & 1) ? XW : XW + 4;
    static constexpr int DP = ((DW / 4) & 1) ? DW : DW + 4;
    static constexpr int kStateFloats = 2 * 32 * XP + 4 * 32 * DP;   // x of 2 steps, rhs of 4 steps
    // stage-to-stage ring: the writer runs 2S steps (plus up to two groups) ahead of the reader
    static constexpr int kHandRows = (2 * S + 2 * kGroup + 2 <= 16) ? 16 : ((2 * S + 2 * kGroup + 2 <= 32) ? 32 : 64);
    static constexpr int kHandFloats = kHandRows * 32;
    static constexpr int kMaxStagesS = (S <= 2) ? 16 : 8;
    // compute warps per CTA: bounded by the register file (S >= 10 needs ~168 registers per thread)
    static constexpr int kMaxComputeWarps = (S <= 5) ? 16 : 8;
    static constexpr int kMaxThreads = (kMaxComputeWarps + 2) * 32;   // + the two comm warps
    static constexpr int kMinBlocks = 1;
};

struct WfKernelParams {
    SweepField fld[kMaxSweepFields];
    int W, H;               // global grid
    int pitch, jbase;       // local arrays: floats per row, global column of local column 0
    int nstr, nstg, ncta;   // strips per CTA, stages per strip, CTAs of this launch (this rank's part of the chain)
    int c_base;             // global index of this launch's first CTA
    float a, c_recip, omega;
    float *xchg;
    unsigned long long *flags;
    unsigned long long seq_base;
    SweepLink link;         // chain hand-over to neighbouring ranks
    unsigned long long link_seq_base;
};

// Per-warp wiring of the pipeline.
struct WfWarp {
    const float *imp;   // ring written by the strip on the left (same stage)
    float *exp;         // ring read by the strip on the right
    int imp_mask, exp_mask;
    const float *in_x;  // level-0 stream of the iterate (HBM tile or previous stage), may be unused
    const float *in_d;  // level-0 stream of the right-hand side
    int in_mask, in_off;
    bool has_in_x;
    float *out_x;       // level-S stream (next stage, or the result tile of the last stage)
    float *out_d;       // level-S right-hand side for the next stage (nullptr on the last stage)
    int out_mask;
    float *bx;          // this warp's x[0..S] of the last 2 steps, [2][32 lanes][XP]
    float *bd;          // this warp's rhs levels 0..S-1 of the last 4 steps, [4][32 lanes][DP]
    bool first;         // stage 0: level 1 is the first sweep of the pass (walls read from the array)
    int W, H;
    float si, sj, a, c_recip, omega;
};

enum { kFixLeft = 1, kFixRight = 2, kFixTop = 4, kFixBottom = 8 };

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_release_u64(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ void st_release_sys_u64(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Rank-to-rank rings carry the flag inside the data: every 16-byte store holds two floats and two copies of
// a tag unique to (pass, batch), and the reader retries until all tags of a batch match.  A 16-byte access is
// performed as a unit (8-byte halves at worst, each with its own tag), so neither side needs a fence and the
// writer never waits for NVLink acknowledgements.
__device__ __forceinline__ void st_volatile_v4(uint4 *p, uint4 v)
{
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ uint4 ld_volatile_v4(const uint4 *p)
{
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_volatile_u64(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- thread-block cluster helpers: neighbouring CTAs of a cluster hand the strip chain over through
// distributed shared memory instead of global rings ----
__device__ __forceinline__ unsigned cluster_ctarank()
{
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}

__device__ __forceinline__ unsigned cluster_nctarank()
{
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}

__device__ __forceinline__ void cluster_barrier()
{
#if WF_CLUSTER_MODE >= 1
    asm volatile("barrier.cluster.arrive.release;\n\tbarrier.cluster.wait.acquire;" ::: "memory");
#endif
}

// shared::cluster address of `p` (a pointer into this CTA's shared memory) in CTA `rank` of the cluster
__device__ __forceinline__ unsigned cluster_map(const void *p, unsigned rank)
{
    unsigned a = (unsigned)__cvta_generic_to_shared(p), r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
    return r;
}

__device__ __forceinline__ void st_cluster_f32x4(unsigned addr, float4 v)
{
    asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__device__ __forceinline__ void st_release_cluster_s32(unsigned addr, int v)
{
    asm volatile("st.release.cluster.shared::cluster.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

__device__ __forceinline__ int ld_acquire_cluster_s32(const int *p)
{
    int v;
    asm volatile("ld.acquire.cluster.shared::cta.s32 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
    return v;
}

// Progress words in shared memory: acquire / release at CTA scope, as the PTX memory model asks for a flag that
// publishes data written by another warp.  -DWF_RELAXED_FENCES swaps them for volatile accesses (relying on an SM
// performing shared-memory accesses in issue order, which is not a PTX guarantee): measured 0.9 % faster on the
// whole step at 8192^2, P = 40 (profiles/), bit-identical in every parity test - not worth leaving the model.
__device__ __forceinline__ int ld_acquire_cta_s32(const int *p)
{
    int v;
#ifndef WF_RELAXED_FENCES
    asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
#else
    asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
#endif
    return v;
}

__device__ __forceinline__ void st_release_cta_s32(int *p, int v)
{
#ifndef WF_RELAXED_FENCES
    asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
#else
    asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
#endif
}

__device__ __forceinline__ void cp_async_f32(float *smem_dst, const float *gsrc, bool valid)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    const int n = valid ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(gsrc), "r"(n) : "memory");
}

__device__ __forceinline__ void cp_async_f32x4(float *smem_dst, const float *gsrc, bool valid)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    const int n = valid ? 16 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(n) : "memory");
}

__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }


// One time step of one warp.  Every lane publishes its x[0..S] and right-hand-side levels in the warp's
// state buffer (lane 31: in the ring read by the next strip instead) and reads lane g-1's state of the
// previous steps from there (lane 0: from the ring written by the previous strip).  U receives lane
// g-1's x of step t-1; Up is the U of the step before (centre value of SOR, first-sweep walls).
// FIX: kStepFast no wall in reach; kStepLeft / kStepRight the strip contains the j == 0 / j == H-1 wall and all
// rows are interior (the common case for the first and last strips: one compare + select per level);
// kStepFull anything (first / last rows of the grid, tiny grids).
enum { kStepFast = 0, kStepLeft = 1, kStepRight = 2, kStepFull = 3, kStepTop = 4 };

__device__ __forceinline__ float bit_select(float keep, float take, unsigned mask)
{
    return __uint_as_float((__float_as_uint(keep) & ~mask) | (__float_as_uint(take) & mask));
}

template <int S, int MODE, int FIX>
__device__ __forceinline__ void wf_step(float (&x)[S + 1], float (&U)[S + 1], const float (&Up)[S + 1], const int t,
                                        const int lane, const int gp, const int fix, const unsigned (&wmask)[S + 1],
                                        const WfWarp &cx)
{
    using L = WfLayout<S>;
    float Dc[S + 1];
    {
        const float *px = (lane == 0) ? cx.imp + ((t - 1) & cx.imp_mask) * L::EXW
                                      : cx.bx + (((t - 1) & 1) * 32 + (lane - 1)) * L::XP;
        const float *pd = (lane == 0) ? cx.imp + ((t - 2) & cx.imp_mask) * L::EXW + L::XW
                                      : cx.bd + (((t - 2) & 3) * 32 + (lane - 1)) * L::DP;
        const float4 *ix = reinterpret_cast<const float4 *>(px);
        const float4 *id = reinterpret_cast<const float4 *>(pd);
#pragma unroll
        for (int q = 0; q < L::XW / 4; ++q) {
            const float4 v = ix[q];
            if (4 * q + 0 <= S) U[(4 * q + 0 <= S) ? 4 * q + 0 : 0] = v.x;
            if (4 * q + 1 <= S) U[(4 * q + 1 <= S) ? 4 * q + 1 : 0] = v.y;
            if (4 * q + 2 <= S) U[(4 * q + 2 <= S) ? 4 * q + 2 : 0] = v.z;
            if (4 * q + 3 <= S) U[(4 * q + 3 <= S) ? 4 * q + 3 : 0] = v.w;
        }
#pragma unroll
        for (int q = 0; q < L::DW / 4; ++q) {
            const float4 v = id[q];
            if (4 * q + 0 < S) Dc[(4 * q + 0 < S) ? 4 * q + 1 : 0] = v.x;
            if (4 * q + 1 < S) Dc[(4 * q + 1 < S) ? 4 * q + 2 : 0] = v.y;
            if (4 * q + 2 < S) Dc[(4 * q + 2 < S) ? 4 * q + 3 : 0] = v.z;
            if (4 * q + 3 < S) Dc[(4 * q + 3 < S) ? 4 * q + 4 : 0] = v.w;
        }
    }
    const int islot = ((t + cx.in_off - lane) & cx.in_mask) * 32 + lane;
    const float nx0 = cx.has_in_x ? cx.in_x[islot] : 0.0f;
    Dc[0] = cx.in_d[islot];
    const float cold1 = Up[0];
    // The S updates of a step are independent of each other (they read only values of step t-1), so
    // they are written stage by stage across all levels: consecutive instructions belong to different
    // dependency chains.
    //   down = U[k-1] (i+1, j) sweep k-1      up    = x[k]   (i-1, j) sweep k
    //   left = U[k]   (i, j-1) sweep k        right = x[k-1] (i, j+1) sweep k-1     centre = Up[k-1]
    float acc[S + 1];
    acc[0] = 0.0f;
    if (MODE == kGaussSeidel) {
        // src/fluid_sim.cpp:325-327: ((((div + down) + up) + right) + left) / 4
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fadd_rn(Dc[k], U[k - 1]);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fadd_rn(acc[k], x[k]);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fadd_rn(acc[k], x[k - 1]);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fadd_rn(acc[k], U[k]);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fmul_rn(acc[k], 0.25f);
    } else {
        // src/fluid_sim.cpp:56-57: nv = (src + a*(((down + up) + right) + left)) * cRecip; x += omega*(nv - x)
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fadd_rn(U[k - 1], x[k]);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fadd_rn(acc[k], x[k - 1]);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fadd_rn(acc[k], U[k]);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fmul_rn(cx.a, acc[k]);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fadd_rn(Dc[k], acc[k]);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fmul_rn(acc[k], cx.c_recip);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fsub_rn(acc[k], Up[k - 1]);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fmul_rn(cx.omega, acc[k]);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = __fadd_rn(Up[k - 1], acc[k]);
    }
    // Cells on a wall carry what setBoundary() (src/fluid_sim.cpp:347-361) left in the array at the moment
    // the reference reads them during sweep k: the array's own value before the first sweep of the pass,
    // +-(the adjacent interior cell's sweep k-1 value) afterwards.  `fix` is warp-uniform and non-zero only
    // for the few batches / strips that touch a wall.  Cells outside the grid hold garbage that no cell
    // inside the grid ever reads.
    if (FIX == kStepLeft) {
        // wmask[k] is all ones in the one lane that sits on column 0 at level k (loop-invariant registers:
        // one LOP3 per level, no predicates).  Gauss-Seidel runs on the pressure only, whose wall sign is +1.
#pragma unroll
        for (int k = 1; k <= S; ++k) {
            const float wr = (MODE == kGaussSeidel) ? x[k - 1] : __fmul_rn(cx.sj, x[k - 1]);
            const float wv = (k == 1 && cx.first) ? cold1 : wr;
            acc[k] = bit_select(acc[k], wv, wmask[k]);
        }
    } else if (FIX == kStepRight) {
        // wmask[k] is all ones in the one lane that sits on column H-1 at level k
#pragma unroll
        for (int k = 1; k <= S; ++k) {
            const float wv = (MODE == kGaussSeidel) ? U[k] : __fmul_rn(cx.sj, U[k]);
            acc[k] = bit_select(acc[k], wv, wmask[k]);
        }
    } else if (FIX == kStepTop) {
        // only the top wall row is in reach (the start of every strip, which the hand-over chain runs through):
        // the wall cell of level k carries +-(row 1, sweep k-1)
        const int ib = t - kToff - gp;   // row of level 0
#pragma unroll
        for (int k = 1; k <= S; ++k) {
            const float wr = (MODE == kGaussSeidel) ? U[k - 1] : __fmul_rn(cx.si, U[k - 1]);
            const float wv = (k == 1 && cx.first) ? cold1 : wr;
            acc[k] = (ib == k) ? wv : acc[k];
        }
    } else if (FIX == kStepFull) {
        if (fix & kFixLeft) {
#pragma unroll
            for (int k = 1; k <= S; ++k) {
                const float wv = (k == 1 && cx.first) ? cold1 : __fmul_rn(cx.sj, x[k - 1]);
                acc[k] = (gp == k) ? wv : acc[k];
            }
        }
        if (fix & kFixRight) {
            const int gw = gp - (cx.H - 1);
#pragma unroll
            for (int k = 1; k <= S; ++k) acc[k] = (gw == k) ? __fmul_rn(cx.sj, U[k]) : acc[k];
        }
        if (fix & kFixTop) {
            const int ib = t - kToff - gp;   // row of level 0
#pragma unroll
            for (int k = 1; k <= S; ++k) {
                const float wv = (k == 1 && cx.first) ? cold1 : __fmul_rn(cx.si, U[k - 1]);
                acc[k] = (ib == k) ? wv : acc[k];
            }
        }
        if (fix & kFixBottom) {
            const int ibw = t - kToff - gp - (cx.W - 1);
#pragma unroll
            for (int k = 1; k <= S; ++k) acc[k] = (ibw == k) ? __fmul_rn(cx.si, x[k]) : acc[k];
        }
    }
#pragma unroll
    for (int k = 1; k <= S; ++k) x[k] = acc[k];
    x[0] = nx0;
    {
        float *wx = (lane == 31) ? cx.exp + (t & cx.exp_mask) * L::EXW : cx.bx + ((t & 1) * 32 + lane) * L::XP;
        float *wd = (lane == 31) ? cx.exp + (t & cx.exp_mask) * L::EXW + L::XW : cx.bd + ((t & 3) * 32 + lane) * L::DP;
        float4 *ex = reinterpret_cast<float4 *>(wx);
#pragma unroll
        for (int q = 0; q < L::XW / 4; ++q) {
            float4 v;
            v.x = x[(4 * q + 0 <= S) ? 4 * q + 0 : 0];
            v.y = x[(4 * q + 1 <= S) ? 4 * q + 1 : 0];
            v.z = x[(4 * q + 2 <= S) ? 4 * q + 2 : 0];
            v.w = x[(4 * q + 3 <= S) ? 4 * q + 3 : 0];
            ex[q] = v;
        }
        float4 *ed = reinterpret_cast<float4 *>(wd);
#pragma unroll
        for (int q = 0; q < L::DW / 4; ++q) {
            float4 v;
            v.x = Dc[(4 * q + 0 < S) ? 4 * q + 0 : 0];
            v.y = Dc[(4 * q + 1 < S) ? 4 * q + 1 : 0];
            v.z = Dc[(4 * q + 2 < S) ? 4 * q + 2 : 0];
            v.w = Dc[(4 * q + 3 < S) ? 4 * q + 3 : 0];
            ed[q] = v;
        }
    }
    const int oslot = ((t - lane) & cx.out_mask) * 32 + lane;
    cx.out_x[oslot] = x[S];
    if (cx.out_d) cx.out_d[oslot] = Dc[S];
    __syncwarp();
}

// One pass of the sweep pipeline for chain CTA `cl` (of P.ncta) and sweep field `fi`, run by all threads of the
// calling CTA: warps [0, nstr*nstg) compute, the next two move the CTA-edge rings, any further warps only take
// part in the two start-up barriers.  `smem` is the CTA's dynamic shared memory (wf_smem_bytes<S>(nstr, nstg)).
// Called by wf_kernel (one pass per launch) and by the phases of chain_kernel (chain.cu).
template <int S, int MODE, bool HAS_X0>
__device__ __forceinline__ void wf_pass(const WfKernelParams &P, const int cl, const int fi, float *smem)
{
    using L = WfLayout<S>;

    const int NSTR = P.nstr, NSTG = P.nstg;
    const int NW = NSTR * NSTG;                  // compute warps; warps NW and NW+1 move the CTA-edge rings
    const int nthreads = blockDim.x;
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int w = tid >> 5;
    const int c = P.c_base + cl;                 // CTA index inside the global chain (geometry); cl: inside this launch
    const SweepField F = P.fld[fi];
    const int W = P.W, H = P.H;
    const int pitch = P.pitch, jbase = P.jbase;
    const bool link_l = (cl == 0) && P.link.has_left;             // left neighbours come from another rank
    const bool link_r = (cl == P.ncta - 1) && P.link.has_right;   // right edge goes to another rank
    const bool has_left = (cl > 0) || link_l;
    const bool has_right = (cl + 1 < P.ncta) || link_r;
    // CTAs of one cluster are consecutive in the chain: inside a cluster the right edge of a CTA goes straight
    // into the next CTA's edge_in ring (distributed shared memory), without the comm warps and global rings
#if WF_CLUSTER_MODE >= 2
    const unsigned crank = cluster_ctarank(), csize = cluster_nctarank();
    const bool dsm_l = (crank > 0) && (cl > 0) && (cl < P.ncta);
    const bool dsm_r = (crank + 1 < csize) && (cl + 1 < P.ncta);
#else
    const unsigned crank = 0;
    const bool dsm_l = false, dsm_r = false;
#endif
    const int j0_first = 32 * NSTR * c;

    // ---- shared memory carve-up -------------------------------------------------------------
    int *prog = reinterpret_cast<int *>(smem);                          // [NW] last step completed by a compute warp
    int *imp_steps = prog + 64;                                         // [NSTG] last step present in edge_in
    int *exp_steps = imp_steps + 16;                                    // [NSTG] last step moved out of edge_out
    int *flushed_steps = exp_steps + 16;                                // [NSTR] last result slot written back to HBM
    float *edge_in = smem + 112;                                         // [NSTG][kEdgeDepth * EXW]
    float *edge_out = edge_in + NSTG * kEdgeDepth * L::EXW;             // [NSTG][kEdgeDepth * EXW]
    float *mid = edge_out + NSTG * kEdgeDepth * L::EXW;                 // [NSTR-1][NSTG][kMidDepth * EXW]
    float *tiles = mid + (NSTR - 1) * NSTG * kMidDepth * L::EXW;        // [NSTR][rhs | x0 | result]
    float *hand = tiles + NSTR * 3 * kTileFloats;                       // [NSTR][NSTG-1][x | rhs]
    float *state = hand + NSTR * (NSTG - 1) * 2 * L::kHandFloats;       // [NW][kStateFloats]
    {
        const int total = 112 + 2 * NSTG * kEdgeDepth * L::EXW + (NSTR - 1) * NSTG * kMidDepth * L::EXW +
                          NSTR * 3 * kTileFloats + NSTR * (NSTG - 1) * 2 * L::kHandFloats + NW * L::kStateFloats;
        for (int i = tid; i < total; i += nthreads) smem[i] = 0.0f;
    }
    __syncthreads();
    // geometry helpers (a warp's own time / column frame): first step, last step
    auto warp_j0w = [&](int sidx_, int g_) { return j0_first + 32 * sidx_ - g_ * S; };
    auto first_step = [&](int j0w_) { return (j0w_ >> 4) << 4; };                   // batch aligned, may be < 0
    auto last_needed = [&](int j0w_) { return (W - 1) + S + kToff + j0w_ + 31; };   // level S reaches the bottom wall row
    auto last_step = [&](int j0w_) { return ((last_needed(j0w_) >> 4) << 4) + 15; };
    if (tid < NW) prog[tid] = first_step(warp_j0w(tid / NSTG, tid % NSTG)) - 1;
    if (tid < NSTG) {
        imp_steps[tid] = first_step(warp_j0w(0, tid)) - 1;
        exp_steps[tid] = first_step(warp_j0w(NSTR - 1, tid)) - 1;
    }
    if (tid < NSTR) flushed_steps[tid] = first_step(warp_j0w(tid, NSTG - 1)) - 48 - 1;
    __syncthreads();
    cluster_barrier();     // every CTA of the cluster has initialised its rings and counters
    if (cl >= P.ncta) {    // padding CTA (the grid is a whole number of clusters)
        cluster_barrier();
        return;
    }
    if (w >= NW + 2) return;   // spare warps of a CTA that is sized for another pass of the same kernel

    // =============================== comm warps ==================================================
    // Warp NW moves the left neighbours in (global ring written by the CTA on the left -> edge_in), warp
    // NW+1 moves the right edge out (edge_out -> global ring read by the CTA on the right); lane gi serves
    // stage gi.  Every round handles all transfers that are ready; the import loads of several stages are in
    // flight together, so the cost of a round hardly grows with the number of stages.
    if (w >= NW) {
        const bool is_exp = (w == NW + 1);
        // lanes 0..15: one ring-transfer machine per stage; lanes 16..31 of the export warp: one result
        // write-back machine per strip of the CTA (rows the last stage has finished, staged in its tile)
        const bool is_flush = WF_FLUSH_OFFLOAD && is_exp && lane >= 16;
        const int gi = lane & 15;
        const bool valid = is_flush ? (gi < NSTR)
                                    : ((lane < 16) && (gi < NSTG) && (is_exp ? (has_right && !dsm_r) : (has_left && !dsm_l)));
        // Ring hand-overs move units of kXfer steps.  Every 16-byte word of a ring (global memory on this GPU, or
        // the neighbouring rank's memory) holds two floats and two copies of a tag unique to (launch, unit): the
        // reader polls the data itself and retries until all tags of a unit match, so neither side executes a
        // fence or waits for a flag round trip.  Only the credit (units consumed) travels back as a counter.
        constexpr int kXfer = WF_XFER;
        constexpr int kXferLog = (kXfer == 4) ? 2 : ((kXfer == 8) ? 3 : 4);
        static_assert(kXfer % kGroup == 0 && kXfer <= 16, "hand-over unit: a whole number of groups, at most a batch");
        constexpr int kRingUnits = kWfSlots * kWfBatch / kXfer;
        constexpr int kUnitFloats = kXfer * L::EXW;
        constexpr int kWords = kUnitFloats / 2;
        constexpr int kWordsPerLane = (kWords + 31) / 32;
        int n = 0, n_end = -1, n_first_right = 0, fl_tau_e = 0;
        if (valid) {
            if (is_flush) {
                const int jl = warp_j0w(gi, NSTG - 1);
                n = (first_step(jl) - 48) >> 4;        // block n = result slots [16n, 16n+16)
                fl_tau_e = last_step(jl);
                n_end = fl_tau_e >> 4;
            } else if (!is_exp) {
                n = first_step(warp_j0w(0, gi)) >> kXferLog;
                n_end = last_needed(warp_j0w(0, gi) - 32) >> kXferLog;    // last unit of the strip on the left
            } else {
                n = first_step(warp_j0w(NSTR - 1, gi)) >> kXferLog;
                n_end = last_needed(warp_j0w(NSTR - 1, gi)) >> kXferLog;
                n_first_right = first_step(warp_j0w(NSTR, gi)) >> kXferLog;   // first unit the reader takes
            }
        }
        unsigned long long *fl = P.flags + ((size_t)fi * kMaxStages + (gi < kMaxStages ? gi : 0)) * kWfMaxCtas * 2;
        const int lk = fi * kLinkStages + (gi < kLinkStages ? gi : 0);   // slot of (field, stage) in the rank links
        const bool via_link = is_exp ? link_r : link_l;
        const unsigned long long flag_base = via_link ? P.link_seq_base : P.seq_base;
        const unsigned tag_seq = (unsigned)(flag_base >> 32) << 18;
        auto unit_tag = [tag_seq](int unit) { return tag_seq + (unsigned)(unit + kBatchBias + 1); };
        // ring of (stage st) this warp reads (import) or writes (export), as floats
        auto ring_of = [&](int st) -> float * {
            if (via_link)
                return (is_exp ? P.link.right_rings : const_cast<float *>(P.link.in_rings)) +
                       (size_t)(fi * kLinkStages + st) * kLinkRingFloats;
            return P.xchg + (((size_t)fi * NSTG + st) * P.ncta + (is_exp ? cl : cl - 1)) * L::kBoundaryFloats;
        };
        const float *my_ring = ring_of(gi < NSTG ? gi : 0);
        // credits: units the reader has consumed
        const volatile unsigned long long *credit_in = link_r ? P.link.cons + lk : fl + (size_t)cl * 2 + 1;
        unsigned long long *credit_out = link_l ? P.link.left_cons + lk : fl + (size_t)(cl > 0 ? cl - 1 : 0) * 2 + 1;
        const int *poll_prog = is_flush ? prog + (gi < NSTR ? gi : 0) * NSTG + (NSTG - 1)
                                        : (is_exp ? prog + (NSTR - 1) * NSTG + (gi < NSTG ? gi : 0) : prog + (gi < NSTG ? gi : 0));
        bool done = !valid || n > n_end;
        while (!__all_sync(0xffffffffu, done)) {
            bool ready = false;
            if (!done) {
                if (is_flush) {
                    // every lane of the strip has written slot q once the warp has finished step q + 31
                    ready = ld_acquire_cta_s32(poll_prog) >= min(16 * n + 15 + 32, fl_tau_e);
                } else if (!is_exp) {
                    // unit n replaces unit n - kEdgeDepth/kXfer in edge_in, whose last slot is read one and two steps later
                    ready = ld_acquire_cta_s32(poll_prog) >= kXfer * (n - kEdgeDepth / kXfer + 1) + 1;
                    if (ready)   // hint only: the tag of the unit's last word (the copy below checks all of them)
                        ready = *reinterpret_cast<const volatile unsigned *>(
                                    my_ring + 2 * (size_t)(n & (kRingUnits - 1)) * kUnitFloats + 2 * kUnitFloats - 1) == unit_tag(n);
                } else {
                    ready = ld_acquire_cta_s32(poll_prog) >= kXfer * n + kXfer - 1 &&
                            (n - kRingUnits < n_first_right ||
                             *credit_in >= flag_base + (unsigned long long)(n - kRingUnits + kBatchBias + 1));
                }
            }
            unsigned m = __ballot_sync(0xffffffffu, ready);
            if (m == 0) {
                __nanosleep(32);
                continue;
            }
            if (!is_exp) {
                // the loads of several stages' units are in flight together (one L2 round trip per burst, not per stage)
                constexpr int kBurst = (kWordsPerLane <= 1) ? 8 : ((kWordsPerLane == 2) ? 4 : 2);
                unsigned mm = m & 0xffffu;
                while (mm) {
                    int srcs[kBurst], bns[kBurst];
                    uint4 v[kBurst][kWordsPerLane];
#pragma unroll
                    for (int b = 0; b < kBurst; ++b) {
                        srcs[b] = -1;
                        bns[b] = 0;
                        if (mm) {
                            srcs[b] = __ffs(mm) - 1;
                            mm &= mm - 1;
                            bns[b] = __shfl_sync(0xffffffffu, n, srcs[b]);
                            const uint4 *gring = reinterpret_cast<const uint4 *>(
                                ring_of(srcs[b]) + 2 * (size_t)(bns[b] & (kRingUnits - 1)) * kUnitFloats);
#pragma unroll
                            for (int k = 0; k < kWordsPerLane; ++k) {
                                const int i = lane + 32 * k;
                                if (i < kWords) v[b][k] = ld_volatile_v4(gring + i);
                            }
                        }
                    }
#pragma unroll
                    for (int b = 0; b < kBurst; ++b) {
                        if (srcs[b] < 0) continue;
                        const unsigned tag = unit_tag(bns[b]);
                        bool ok = true;
#pragma unroll
                        for (int k = 0; k < kWordsPerLane; ++k) {
                            const int i = lane + 32 * k;
                            if (i < kWords) ok = ok && v[b][k].y == tag && v[b][k].w == tag;
                        }
                        if (__all_sync(0xffffffffu, ok)) {
                            float2 *sring = reinterpret_cast<float2 *>(edge_in + srcs[b] * kEdgeDepth * L::EXW +
                                                                       ((kXfer * bns[b]) & (kEdgeDepth - 1)) * L::EXW);
#pragma unroll
                            for (int k = 0; k < kWordsPerLane; ++k) {
                                const int i = lane + 32 * k;
                                if (i < kWords) sring[i] = make_float2(__uint_as_float(v[b][k].x), __uint_as_float(v[b][k].z));
                            }
                        } else if (lane == srcs[b]) {
                            ready = false;    // part of the unit is still in flight: try again next round
                        }
                    }
                }
            } else {
                for (unsigned mm = m & 0xffffu; mm; mm &= mm - 1) {
                    const int src = __ffs(mm) - 1;
                    const int bn = __shfl_sync(0xffffffffu, n, src);
                    uint4 *gring = reinterpret_cast<uint4 *>(ring_of(src) + 2 * (size_t)(bn & (kRingUnits - 1)) * kUnitFloats);
                    const float2 *sring = reinterpret_cast<const float2 *>(edge_out + src * kEdgeDepth * L::EXW +
                                                                           ((kXfer * bn) & (kEdgeDepth - 1)) * L::EXW);
                    const unsigned tag = unit_tag(bn);
#pragma unroll
                    for (int k = 0; k < kWordsPerLane; ++k) {
                        const int i = lane + 32 * k;
                        if (i < kWords) {
                            const float2 d = sring[i];
                            st_volatile_v4(gring + i, make_uint4(__float_as_uint(d.x), tag, __float_as_uint(d.y), tag));
                        }
                    }
                }
            }
            for (unsigned mm = m >> 16; mm; mm &= mm - 1) {
                // write back result block bn of strip `strip` (lane = column inside the strip)
                const int strip = __ffs(mm) - 1;
                const int bn = __shfl_sync(0xffffffffu, n, 16 + strip);
                const int jl = warp_j0w(strip, NSTG - 1);
                const float *tile = tiles + strip * 3 * kTileFloats + 2 * kTileFloats;
                const int col = jl + lane - S;
                const bool colok = (col >= 1) && (col <= H - 2);
                const int row0 = 16 * bn - kToff - S - jl;
                float vals[kWfBatch];
#pragma unroll
                for (int it = 0; it < kWfBatch; ++it) vals[it] = tile[((16 * bn + it) & (kTileRows - 1)) * 32 + lane];
                if (colok && row0 + kWfBatch - 1 >= 1 && row0 <= W - 2) {
                    float *dst = F.out + ((long long)row0 * pitch + (col - jbase));
#pragma unroll
                    for (int it = 0; it < kWfBatch; ++it)
                        if (row0 + it >= 1 && row0 + it <= W - 2) dst[(long long)it * pitch] = vals[it];
                }
            }
            __syncwarp();
            if (ready) {
                if (is_flush) {
                    st_release_cta_s32(flushed_steps + gi, 16 * n + 15);
                } else if (!is_exp) {
                    st_release_cta_s32(imp_steps + gi, kXfer * n + kXfer - 1);
                    // the unit has been read (its tags were checked), so the credit needs no ordering
                    st_volatile_u64(credit_out, flag_base + (unsigned long long)(n + kBatchBias + 1));
                } else {
                    st_release_cta_s32(exp_steps + gi, kXfer * n + kXfer - 1);
                }
                n += 1;
                done = n > n_end;
            }
        }
        cluster_barrier();   // nobody leaves while a neighbour may still write into this CTA's shared memory
        return;
    }

    // =============================== compute warps =================================================
    const int sidx = w / NSTG;
    const int g = w - sidx * NSTG;
    float *t_rhs = tiles + sidx * 3 * kTileFloats;
    float *t_x0 = t_rhs + kTileFloats;
    float *t_out = t_x0 + kTileFloats;

    WfWarp cx;
    cx.imp = (sidx == 0) ? edge_in + g * kEdgeDepth * L::EXW : mid + ((sidx - 1) * NSTG + g) * kMidDepth * L::EXW;
    cx.imp_mask = (sidx == 0) ? (kEdgeDepth - 1) : (kMidDepth - 1);
    cx.exp = (sidx == NSTR - 1) ? edge_out + g * kEdgeDepth * L::EXW : mid + (sidx * NSTG + g) * kMidDepth * L::EXW;
    cx.exp_mask = (sidx == NSTR - 1) ? (kEdgeDepth - 1) : (kMidDepth - 1);
    if (g == 0) {
        cx.in_x = t_x0;
        cx.in_d = t_rhs;
        cx.in_mask = kTileRows - 1;
        cx.in_off = 0;
        cx.has_in_x = HAS_X0;
    } else {
        cx.in_x = hand + (sidx * (NSTG - 1) + (g - 1)) * 2 * L::kHandFloats;
        cx.in_d = cx.in_x + L::kHandFloats;
        cx.in_mask = L::kHandRows - 1;
        cx.in_off = 2 * S;
        cx.has_in_x = true;
    }
    if (g == NSTG - 1) {
        cx.out_x = t_out;
        cx.out_d = nullptr;
        cx.out_mask = kTileRows - 1;
    } else {
        cx.out_x = hand + (sidx * (NSTG - 1) + g) * 2 * L::kHandFloats;
        cx.out_d = cx.out_x + L::kHandFloats;
        cx.out_mask = L::kHandRows - 1;
    }
    cx.bx = state + w * L::kStateFloats;
    cx.bd = cx.bx + 2 * 32 * L::XP;
    cx.first = (g == 0);
    cx.W = W;
    cx.H = H;
    cx.si = F.si;
    cx.sj = F.sj;
    cx.a = P.a;
    cx.c_recip = P.c_recip;
    cx.omega = P.omega;

    const int j0 = j0_first + 32 * sidx;         // strip base column in stage 0's frame
    const int j0w = warp_j0w(sidx, g);           // ... in this stage's frame
    const int gp = j0w + lane;
    const int tau_b = first_step(j0w);
    const int tau_e = last_step(j0w);
    // last stage of a strip whose result columns reach into the block of the rank on the left
    const bool to_peer = (g == NSTG - 1) && (F.out_peer != nullptr) && (j0w - S < P.link.peer_hi);

    // ---- what this warp waits for before a group of steps starting at t: counter >= min(t + add, clamp) ----
    //   lane 0: left neighbours present      lane 1: previous stage far enough ahead
    //   lane 2: right strip / export freed the slots we overwrite      lane 3: next stage freed its ring slots
    //   lane 4: (last stage) the result tile slots we overwrite have been written back
    // hand-over inside a cluster: this warp pushes its right-edge rows into the next CTA's edge_in ring and
    // publishes its progress there (push_r); the first strip tells the previous CTA how far it has read (cred_l)
    const bool push_r = dsm_r && (sidx == NSTR - 1);
    const bool cred_l = dsm_l && (sidx == 0);
    const unsigned r_edge = push_r ? cluster_map(edge_in + g * kEdgeDepth * L::EXW, crank + 1) : 0u;
    const unsigned r_imp = push_r ? cluster_map(imp_steps + g, crank + 1) : 0u;
    const unsigned l_exp = cred_l ? cluster_map(exp_steps + g, crank - 1) : 0u;
    const bool cluster_waits = push_r || cred_l;   // some counter this warp polls is written by another CTA of the cluster
    const int *wait_ptr = nullptr;
    int wait_add = 0, wait_clamp = 0x7fffffff;
    if (lane == 0) {
        if (sidx > 0) {
            wait_ptr = prog + (w - NSTG);
            wait_clamp = last_step(j0w - 32);
        } else if (has_left) {
            wait_ptr = imp_steps + g;
            wait_clamp = WF_XFER * (last_needed(j0w - 32) / WF_XFER) + WF_XFER - 1;   // end of the last unit handed over
        }
        wait_add = kGroup - 2;
    } else if (lane == 1) {
        if (g > 0) {
            wait_ptr = prog + (w - 1);
            wait_clamp = last_step(j0w + S);
        }
        wait_add = kGroup - 1 + 2 * S;
    } else if (lane == 2) {
        if (sidx < NSTR - 1) {
            wait_ptr = prog + (w + NSTG);
            wait_add = kGroup - 1 - (kMidDepth - 2);
        } else if (has_right) {
            wait_ptr = exp_steps + g;
            wait_add = kGroup - 1 - kEdgeDepth;
            if (dsm_r) {   // exp_steps[g] is the progress of the reader in the next CTA; its ring has kEdgeDepth rows
                wait_add = kGroup - 1 - (kEdgeDepth - 2);
            }
        }
    } else if (lane == 3) {
        if (g < NSTG - 1) wait_ptr = prog + (w + 1);
        wait_add = kGroup - 1 - L::kHandRows - 2 * S;
    } else if (lane == 4) {
        // last stage: the result tile slot we are about to overwrite has been written back
        if (WF_FLUSH_OFFLOAD && g == NSTG - 1) wait_ptr = flushed_steps + sidx;
        wait_add = kGroup - 1 - kTileRows;
    }

    float x[S + 1], Ua[S + 1], Ub[S + 1];
#pragma unroll
    for (int k = 0; k <= S; ++k) {
        x[k] = 0.0f; Ua[k] = 0.0f; Ub[k] = 0.0f;
    }

    const bool vec_ok = ((H & 3) == 0) && ((pitch & 3) == 0) && ((jbase & 3) == 0);
    int fix_cols = 0;
    if (j0w - S <= 0 && j0w + 30 >= 0) fix_cols |= kFixLeft;
    if (j0w - S <= H - 1 && j0w + 30 >= H - 1) fix_cols |= kFixRight;
    int fix = fix_cols;
    // lane masks of the column wall this strip contains (if exactly one); the empty asm keeps them in
    // registers instead of letting the compiler recompute them inside the loop
    unsigned wmask[S + 1];
#pragma unroll
    for (int k = 0; k <= S; ++k) {
        const int col = gp - k;
        wmask[k] = ((fix_cols == kFixLeft && col == 0) || (fix_cols == kFixRight && col == H - 1)) ? 0xffffffffu : 0u;
        asm volatile("" : "+r"(wmask[k]));
    }
    // stage 0: copy the 16 array rows the strip reaches during steps [q0, q0+16) into the input tiles (cp.async)
    auto stage_tile_rows = [&](const int q0) {
        if (vec_ok) {
            const int rr = lane >> 3, cg = lane & 7;
#pragma unroll
            for (int it = 0; it < kWfBatch / 4; ++it) {
                const int q = q0 + it * 4 + rr;
                const int row = q - kToff - j0;
                const int col = j0 + cg * 4;
                const bool ok = (row >= 0) && (row < W) && (col < H);
                const size_t off = ok ? ((size_t)row * pitch + (col - jbase)) : 0;
                const int so = (q & (kTileRows - 1)) * 32 + cg * 4;
                cp_async_f32x4(t_rhs + so, F.rhs + off, ok);
                if (HAS_X0) cp_async_f32x4(t_x0 + so, F.x0 + off, ok);
            }
        } else {
#pragma unroll 4
            for (int it = 0; it < kWfBatch; ++it) {
                const int q = q0 + it;
                const int row = q - kToff - j0;
                const int col = j0 + lane;
                const bool ok = (row >= 0) && (row < W) && (col < H);
                const size_t off = ok ? ((size_t)row * pitch + (col - jbase)) : 0;
                const int so = (q & (kTileRows - 1)) * 32 + lane;
                cp_async_f32(t_rhs + so, F.rhs + off, ok);
                if (HAS_X0) cp_async_f32(t_x0 + so, F.x0 + off, ok);
            }
        }
    };
#ifdef WF_PROFILE
    long long pf_wait = 0, pf_batch = 0, pf_step = 0, pf_first = 0;
    int pf_groups = 0;
    long long pf_spins = 0;
    long long pf_mine = 0;   // spins in which this lane's own condition was the unmet one
    long long pf_top = 0, pf_early = 0;
    int pf_top_n = 0;
#endif
#pragma unroll 1
    for (int t = tau_b; t <= tau_e; t += kGroup) {
#ifdef WF_PROFILE
        const long long pf0 = clock64();
#endif
        // ---- wait for producers / consumers ------------------------------------------------------
        {
            const int need = min(t + wait_add, wait_clamp);
            for (;;) {
                bool ok = true;
                if (wait_ptr) ok = ld_acquire_cta_s32(wait_ptr) >= need;
#ifdef WF_PROFILE
                if (!ok) pf_mine++;
#endif
                if (__all_sync(0xffffffffu, ok)) break;
#ifdef WF_PROFILE
                pf_spins++;
#endif
            }
            // counters written by another CTA of the cluster were polled with plain loads: order the data after them
            if (cluster_waits) asm volatile("fence.acq_rel.cluster;" ::: "memory");
        }
#ifdef WF_PROFILE
        const long long pf1 = clock64();
        if (pf_groups == 0) pf_first = pf1 - pf0; else pf_wait += pf1 - pf0;
#endif
        // ---- once per batch of 16 steps ----------------------------------------------------------
        if ((t & (kWfBatch - 1)) == 0) {
            if (g == 0) {
                // stage the rows this strip reaches during the next batch
                cp_async_wait_all();
                stage_tile_rows(t + kWfBatch);
                cp_async_commit();
            }
            if (!WF_FLUSH_OFFLOAD && g == NSTG - 1) {
                // write back the result rows that every lane has finished (all loads first, then all stores)
                const int col = gp - S;
                const bool colok = (col >= 1) && (col <= H - 2);
                const int row0 = (t - 48) - kToff - S - j0w;
                float vals[kWfBatch];
#pragma unroll
                for (int it = 0; it < kWfBatch; ++it) vals[it] = t_out[((t - 48 + it) & (kTileRows - 1)) * 32 + lane];
                if (colok && row0 + kWfBatch - 1 >= 1 && row0 <= W - 2) {
                    float *dst = F.out + ((long long)row0 * pitch + (col - jbase));
#pragma unroll
                    for (int it = 0; it < kWfBatch; ++it) {
                        if (row0 + it >= 1 && row0 + it <= W - 2) dst[(long long)it * pitch] = vals[it];
                    }
                    if (to_peer && col < P.link.peer_hi) {
                        // these columns belong to (or border) the block of the rank on the left: store them there too
                        float *pd = F.out_peer + ((long long)row0 * pitch + (col - P.link.peer_jbase));
#pragma unroll
                        for (int it = 0; it < kWfBatch; ++it) {
                            if (row0 + it >= 1 && row0 + it <= W - 2) pd[(long long)it * pitch] = vals[it];
                        }
                    }
                }
            }
            // which walls the (lane, level) cells of this batch can touch
            const int i_min = t - kToff - j0w - 31 - S;
            const int i_max = t + kWfBatch - 1 - kToff - j0w - 1;
            fix = fix_cols;
            if (i_min <= 0 && i_max >= 0) fix |= kFixTop;
            if (i_min <= W - 1 && i_max >= W - 1) fix |= kFixBottom;
            __syncwarp();
        }
        // ---- four time steps ----------------------------------------------------------------------
#ifdef WF_PROFILE
        const long long pf2 = clock64();
        pf_batch += pf2 - pf1;
#endif
        // two step bodies (the U / Up roles alternate), kept small for the instruction cache; the wall
        // fix-ups live in a second copy that only strips / batches near a wall execute
        if (fix == 0) {
#pragma unroll 1
            for (int u2 = 0; u2 < kGroup; u2 += 2) {
                wf_step<S, MODE, kStepFast>(x, Ua, Ub, t + u2, lane, gp, 0, wmask, cx);
                wf_step<S, MODE, kStepFast>(x, Ub, Ua, t + u2 + 1, lane, gp, 0, wmask, cx);
            }
        } else if (fix == kFixLeft) {
#pragma unroll 1
            for (int u2 = 0; u2 < kGroup; u2 += 2) {
                wf_step<S, MODE, kStepLeft>(x, Ua, Ub, t + u2, lane, gp, fix, wmask, cx);
                wf_step<S, MODE, kStepLeft>(x, Ub, Ua, t + u2 + 1, lane, gp, fix, wmask, cx);
            }
        } else if (fix == kFixRight) {
#pragma unroll 1
            for (int u2 = 0; u2 < kGroup; u2 += 2) {
                wf_step<S, MODE, kStepRight>(x, Ua, Ub, t + u2, lane, gp, fix, wmask, cx);
                wf_step<S, MODE, kStepRight>(x, Ub, Ua, t + u2 + 1, lane, gp, fix, wmask, cx);
            }
        } else if (fix == kFixTop) {
#pragma unroll 1
            for (int u2 = 0; u2 < kGroup; u2 += 2) {
                wf_step<S, MODE, kStepTop>(x, Ua, Ub, t + u2, lane, gp, fix, wmask, cx);
                wf_step<S, MODE, kStepTop>(x, Ub, Ua, t + u2 + 1, lane, gp, fix, wmask, cx);
            }
        } else {
#pragma unroll 1
            for (int u2 = 0; u2 < kGroup; u2 += 2) {
                wf_step<S, MODE, kStepFull>(x, Ua, Ub, t + u2, lane, gp, fix, wmask, cx);
                wf_step<S, MODE, kStepFull>(x, Ub, Ua, t + u2 + 1, lane, gp, fix, wmask, cx);
            }
        }
        if (lane == 0) st_release_cta_s32(prog + w, t + kGroup - 1);
        if (push_r) {
            // rows t .. t+kGroup-1 of the local right-edge ring -> the same rows of the next CTA's edge_in
            const int row_off = (t & (kEdgeDepth - 1)) * L::EXW;
            const float4 *src = reinterpret_cast<const float4 *>(cx.exp + row_off);
#pragma unroll
            for (int i = lane; i < kGroup * L::EXW / 4; i += 32) st_cluster_f32x4(r_edge + 4u * (unsigned)row_off + 16u * (unsigned)i, src[i]);
            __syncwarp();
            if (lane == 0) st_release_cluster_s32(r_imp, t + kGroup - 1);
        }
        if (cred_l && lane == 0) st_release_cluster_s32(l_exp, t + kGroup - 1);
#ifdef WF_PROFILE
        pf_step += clock64() - pf2;
        if (fix & kFixTop) { pf_top += clock64() - pf0; pf_top_n++; }
        if (pf_groups >= 1 && pf_groups <= 9) pf_early += clock64() - pf0;
        pf_groups++;
#endif
    }
#ifdef WF_PROFILE
    {
        const long long m0 = __shfl_sync(0xffffffffu, pf_mine, 0), m1 = __shfl_sync(0xffffffffu, pf_mine, 1);
        const long long m2 = __shfl_sync(0xffffffffu, pf_mine, 2), m3 = __shfl_sync(0xffffffffu, pf_mine, 3);
        if (lane == 0 && fi == 0 && (c == P.ncta / 2 || c == 0 || c == P.ncta - 1 || c == P.ncta - 2))
            printf("cta %d warp %d: unmet per group: left %.1f prev-stage %.1f right/export %.1f next-stage %.1f\n", c, w,
                   (double)m0 / pf_groups, (double)m1 / pf_groups, (double)m2 / pf_groups, (double)m3 / pf_groups);
    }
    if (lane == 0 && fi == 0 && (c == P.ncta / 2 || c == 0 || c == P.ncta - 1 || c == P.ncta - 2))
        printf("cta %d warp %d (strip %d stage %d): groups %d  first-wait %lld  per group: wait %lld batch %lld steps %lld  spins/group %.1f\n",
               c, w, sidx, g, pf_groups, pf_first, pf_wait / pf_groups, pf_batch / pf_groups, pf_step / pf_groups,
               (double)pf_spins / pf_groups);
    if (lane == 0 && fi == 0 && (c == P.ncta / 2 || c == 0))
        printf("cta %d warp %d: top-wall groups %d, %lld cycles each; groups 1..9 after the first wait: %lld cycles each\n", c, w,
               pf_top_n, pf_top / max(pf_top_n, 1), pf_early / 9);
#endif
    cp_async_wait_all();
    if (!WF_FLUSH_OFFLOAD && g == NSTG - 1) {
        // drain: rows finished during the last three batches
        __syncwarp();
        const int col = gp - S;
        const bool colok = (col >= 1) && (col <= H - 2);
        for (int q = tau_e + 1 - 48; q <= tau_e; ++q) {
            const int row = q - kToff - S - j0w;
            if (colok && row >= 1 && row <= W - 2) {
                const float val = t_out[(q & (kTileRows - 1)) * 32 + lane];
                F.out[(size_t)row * pitch + (col - jbase)] = val;
                if (to_peer && col < P.link.peer_hi) F.out_peer[(size_t)row * pitch + (col - P.link.peer_jbase)] = val;
            }
        }
        if (to_peer) {
            // all result columns this strip owes the left rank are stored: tell it
            __threadfence_system();
            __syncwarp();
            if (lane == 0) atomicAdd_system(P.link.left_ret + fi, 1ull);
        }
    }
    cluster_barrier();
}

template <int S>
size_t wf_smem_bytes(int nstr, int nstg)
{
    using L = WfLayout<S>;
    return sizeof(float) * ((size_t)112 + (size_t)2 * nstg * kEdgeDepth * L::EXW +
                            (size_t)(nstr - 1) * nstg * kMidDepth * L::EXW + (size_t)nstr * 3 * kTileFloats +
                            (size_t)nstr * (nstg - 1) * 2 * L::kHandFloats + (size_t)nstr * nstg * L::kStateFloats);
}

}  // namespace

}  // namespace fsb
