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
//     (left / down), fetched with one shuffle per level.  Hence S sweeps are in flight per warp.
//   * the right-hand side (divergence or diffusion source) travels with the cell through a second
//     shuffle pipeline (two steps per level).
//   * stages: the sweeps of a pass are split over `nstg` warps per strip.  Stage g is the same
//     program shifted by g*S columns and 2*g*S time steps; it takes its level-0 stream (x and
//     right-hand side, lane-aligned) from the shared-memory ring stage g-1 writes its level-S
//     stream to.  Stage 0 is fed from HBM through cp.async row tiles, the last stage writes the
//     result rows back.  All sweeps of a pass therefore read and write each cell of HBM once.
//   * strips: lane 0 takes its left neighbours from the previous strip of the same stage: through a
//     shared-memory ring inside a CTA, through a global ring with acquire/release progress flags
//     between CTAs (batches of 16 steps).  The launch is cooperative so that every CTA of a chain
//     is resident while its right neighbour waits for it.
//   * the warps of a CTA advance in lockstep, one __syncthreads per time step; stage g runs
//     g*stage_off steps behind stage 0 (a multiple of 16, so batch boundaries coincide).
//   * wall cells are evaluated in the pipeline with the value setBoundary() would have left in the
//     array at the moment the reference reads it (warp-uniform fix-ups, only near walls).
//
// All arithmetic uses explicit round-to-nearest intrinsics: no FMA contraction (F7).
#include "fsb_internal.h"

namespace fsb {

namespace {

constexpr int kToff = 32;        // time offset: lane g sits on row t - kToff - g at level 0
constexpr int kTileRows = 64;    // rows of an HBM staging ring (input tiles, result tile)
constexpr int kTileFloats = kTileRows * 32;
constexpr int kHandRows = 32;    // steps held by a stage-to-stage ring
constexpr int kHandFloats = kHandRows * 32;
constexpr int kEdgeDepth = 32;   // steps held by the CTA-edge rings (two batches)
constexpr int kMidDepth = 4;     // steps held by the rings between strips of a CTA
constexpr int kBatchBias = 64;   // keeps published batch numbers positive (later stages start at t < 0)
constexpr int kMaxStages = 32;

template <int S>
struct WfLayout {
    static constexpr int XW = (S + 1 + 3) & ~3;   // x[0..S], padded to float4
    static constexpr int DW = (S + 3) & ~3;       // rhs levels 0..S-1
    static constexpr int EXW = XW + DW;           // floats one lane hands to the next strip per step
    static constexpr int kBoundaryFloats = kWfSlots * kWfBatch * EXW;
    static constexpr int kMaxThreads = (S <= 5) ? 512 : 256;
};

struct WfKernelParams {
    SweepField fld[kMaxSweepFields];
    int W, H;
    int nstr, nstg, ncta;   // strips per CTA, stages per strip, CTAs per chain
    int stage_off;          // wall-clock steps between consecutive stages (multiple of 16, >= 2S+2)
    float a, c_recip, omega;
    float *xchg;
    unsigned long long *flags;
    unsigned long long seq_base;
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


// One time step of one warp.  U receives lane g-1's x of the previous step, Up is the U of the step
// before (needed for the centre value of SOR and for first-sweep walls).
template <int S, int MODE>
__device__ __forceinline__ void wf_step(float (&x)[S + 1], float (&U)[S + 1], const float (&Up)[S + 1],
                                        float (&Dc)[S + 1], const int t, const int lane, const int gp,
                                        const int fix, const WfWarp &cx)
{
    using L = WfLayout<S>;
#pragma unroll
    for (int k = 0; k <= S; ++k) U[k] = __shfl_up_sync(0xffffffffu, x[k], 1);
#pragma unroll
    for (int k = S; k >= 1; --k) Dc[k] = __shfl_up_sync(0xffffffffu, Dc[k - 1], 1);
    if (lane == 0) {
        const float4 *ix = reinterpret_cast<const float4 *>(cx.imp + ((t - 1) & cx.imp_mask) * L::EXW);
        const float4 *id = reinterpret_cast<const float4 *>(cx.imp + ((t - 2) & cx.imp_mask) * L::EXW + L::XW);
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
    if (fix) {
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
    if (lane == 31) {
        float4 *ex = reinterpret_cast<float4 *>(cx.exp + (t & cx.exp_mask) * L::EXW);
#pragma unroll
        for (int q = 0; q < L::XW / 4; ++q) {
            float4 v;
            v.x = x[(4 * q + 0 <= S) ? 4 * q + 0 : 0];
            v.y = x[(4 * q + 1 <= S) ? 4 * q + 1 : 0];
            v.z = x[(4 * q + 2 <= S) ? 4 * q + 2 : 0];
            v.w = x[(4 * q + 3 <= S) ? 4 * q + 3 : 0];
            ex[q] = v;
        }
        float4 *ed = ex + L::XW / 4;
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
}

template <int S, int MODE, bool HAS_X0>
__global__ void __launch_bounds__(WfLayout<S>::kMaxThreads, 1) wf_kernel(const WfKernelParams P)
{
    using L = WfLayout<S>;
    extern __shared__ float4 smem4[];
    float *smem = reinterpret_cast<float *>(smem4);

    const int NSTR = P.nstr, NSTG = P.nstg;
    const int nthreads = NSTR * NSTG * 32;
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int w = tid >> 5;
    const int sidx = w / NSTG;
    const int g = w - sidx * NSTG;
    const int c = blockIdx.x;
    const int fi = blockIdx.y;
    const SweepField F = P.fld[fi];
    const int W = P.W, H = P.H;

    // ---- shared memory carve-up -------------------------------------------------------------
    float *edge_in = smem;                                              // [NSTG][kEdgeDepth * EXW]
    float *edge_out = edge_in + NSTG * kEdgeDepth * L::EXW;             // [NSTG][kEdgeDepth * EXW]
    float *mid = edge_out + NSTG * kEdgeDepth * L::EXW;                 // [NSTR-1][NSTG][kMidDepth * EXW]
    float *tiles = mid + (NSTR - 1) * NSTG * kMidDepth * L::EXW;        // [NSTR][rhs | x0 | result]
    float *hand = tiles + NSTR * 3 * kTileFloats;                       // [NSTR][NSTG-1][x | rhs]
    {
        const int total = 2 * NSTG * kEdgeDepth * L::EXW + (NSTR - 1) * NSTG * kMidDepth * L::EXW +
                          NSTR * 3 * kTileFloats + NSTR * (NSTG - 1) * 2 * kHandFloats;
        for (int i = tid; i < total; i += nthreads) smem[i] = 0.0f;
    }
    __syncthreads();

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
        cx.in_x = hand + (sidx * (NSTG - 1) + (g - 1)) * 2 * kHandFloats;
        cx.in_d = cx.in_x + kHandFloats;
        cx.in_mask = kHandRows - 1;
        cx.in_off = 2 * S;
        cx.has_in_x = true;
    }
    if (g == NSTG - 1) {
        cx.out_x = t_out;
        cx.out_d = nullptr;
        cx.out_mask = kTileRows - 1;
    } else {
        cx.out_x = hand + (sidx * (NSTG - 1) + g) * 2 * kHandFloats;
        cx.out_d = cx.out_x + kHandFloats;
        cx.out_mask = kHandRows - 1;
    }
    cx.first = (g == 0);
    cx.W = W;
    cx.H = H;
    cx.si = F.si;
    cx.sj = F.sj;
    cx.a = P.a;
    cx.c_recip = P.c_recip;
    cx.omega = P.omega;

    // ---- geometry of this warp (its own time / column frame) ------------------------------------
    const int koff = g * S;                      // sweeps done by earlier stages
    const int j0_first = 32 * NSTR * c;
    const int j0 = j0_first + 32 * sidx;         // strip base column in stage 0's frame
    const int j0w = j0 - koff;                   // ... in this stage's frame
    const int gp = j0w + lane;
    const int off_w = g * P.stage_off;           // wall-clock lag behind stage 0
    const int tau_b = (j0w >> 4) << 4;           // first step (batch aligned; arithmetic shift: may be < 0)
    const int tau_last = (W - 1) + S + kToff + j0w + 31;   // level S reaches the bottom wall row (the next stage reads it)
    const int tau_e = ((tau_last >> 4) << 4) + 15;
    const int n_last_left = ((W - 1) + S + kToff + (j0_first - 32 - koff) + 31) >> 4;
    const int n_first_right = (j0_first + 32 * NSTR - koff) >> 4;
    const int T0 = j0_first;
    int T1;
    {
        const int kl = (NSTG - 1) * S;
        const int jl = j0_first + 32 * (NSTR - 1) - kl;
        const int tl = (W - 1) + S + kToff + jl + 31;
        T1 = (((tl >> 4) << 4) + 15) + (NSTG - 1) * P.stage_off;
    }

    unsigned long long *flags_g = P.flags + ((size_t)fi * kMaxStages + g) * kWfMaxCtas * 2;
    float *xchg_g = P.xchg + ((size_t)fi * NSTG + g) * P.ncta * L::kBoundaryFloats;

    float x[S + 1], Ua[S + 1], Ub[S + 1], Da[S + 1], Db[S + 1];
#pragma unroll
    for (int k = 0; k <= S; ++k) {
        x[k] = 0.0f; Ua[k] = 0.0f; Ub[k] = 0.0f; Da[k] = 0.0f; Db[k] = 0.0f;
    }

    const bool vec_ok = ((H & 3) == 0);
    int fix_cols = 0;
    if (j0w - S <= 0 && j0w + 30 >= 0) fix_cols |= kFixLeft;
    if (j0w - S <= H - 1 && j0w + 30 >= H - 1) fix_cols |= kFixRight;
    int fix = fix_cols;

    // Work of a warp at the first step of a batch.
    auto batch_begin = [&](const int t0) {
        const int n = t0 >> 4;
        // left neighbours of this batch from the CTA on the left (same stage)
        if (sidx == 0 && c > 0 && n <= n_last_left) {
            if (lane == 0) {
                const unsigned long long want = P.seq_base + (unsigned long long)(n + kBatchBias + 1);
                const unsigned long long *pf = flags_g + (size_t)(c - 1) * 2;
                while (ld_acquire_u64(pf) < want) __nanosleep(32);
            }
            __syncwarp();
            const float4 *src = reinterpret_cast<const float4 *>(
                xchg_g + (size_t)(c - 1) * L::kBoundaryFloats + (size_t)(n & (kWfSlots - 1)) * kWfBatch * L::EXW);
            float4 *dst = reinterpret_cast<float4 *>(const_cast<float *>(cx.imp) + (t0 & (kEdgeDepth - 1)) * L::EXW);
            for (int i = lane; i < kWfBatch * L::EXW / 4; i += 32) dst[i] = __ldcg(src + i);
            __syncwarp();
            if (lane == 0) st_release_u64(flags_g + (size_t)(c - 1) * 2 + 1, P.seq_base + (unsigned long long)(n + kBatchBias + 1));
        }
        // stage 0: stage the rows this strip reaches during the next batch
        if (g == 0) {
            cp_async_wait_all();
            const int q0 = t0 + kWfBatch;
            if (vec_ok) {
                const int rr = lane >> 3, cg = lane & 7;
#pragma unroll
                for (int it = 0; it < kWfBatch / 4; ++it) {
                    const int q = q0 + it * 4 + rr;
                    const int row = q - kToff - j0;
                    const int col = j0 + cg * 4;
                    const bool ok = (row >= 0) && (row < W) && (col < H);
                    const size_t off = ok ? ((size_t)row * H + col) : 0;
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
                    const size_t off = ok ? ((size_t)row * H + col) : 0;
                    const int so = (q & (kTileRows - 1)) * 32 + lane;
                    cp_async_f32(t_rhs + so, F.rhs + off, ok);
                    if (HAS_X0) cp_async_f32(t_x0 + so, F.x0 + off, ok);
                }
            }
            cp_async_commit();
        }
        // last stage: write back the result rows that every lane has finished
        if (g == NSTG - 1) {
            const int col = gp - S;
            const bool colok = (col >= 1) && (col <= H - 2);
#pragma unroll 4
            for (int it = 0; it < kWfBatch; ++it) {
                const int q = t0 - 48 + it;
                const int row = q - kToff - S - j0w;
                if (colok && row >= 1 && row <= W - 2)
                    F.out[(size_t)row * H + col] = t_out[(q & (kTileRows - 1)) * 32 + lane];
            }
        }
        // which walls the (lane, level) cells of this batch can touch
        const int i_min = t0 - kToff - j0w - 31 - S;
        const int i_max = t0 + kWfBatch - 1 - kToff - j0w - 1;
        fix = fix_cols;
        if (i_min <= 0 && i_max >= 0) fix |= kFixTop;
        if (i_min <= W - 1 && i_max >= W - 1) fix |= kFixBottom;
    };

    // Work of a warp after the last step of a batch.
    auto batch_end = [&](const int t0) {
        const int n = t0 >> 4;
        if (sidx == NSTR - 1 && c + 1 < P.ncta) {
            __syncwarp();
            if (n - kWfSlots >= n_first_right) {
                if (lane == 0) {
                    const unsigned long long want = P.seq_base + (unsigned long long)(n - kWfSlots + kBatchBias + 1);
                    const unsigned long long *pf = flags_g + (size_t)c * 2 + 1;
                    while (ld_acquire_u64(pf) < want) __nanosleep(32);
                }
                __syncwarp();
            }
            float4 *dst = reinterpret_cast<float4 *>(
                xchg_g + (size_t)c * L::kBoundaryFloats + (size_t)(n & (kWfSlots - 1)) * kWfBatch * L::EXW);
            const float4 *src = reinterpret_cast<const float4 *>(cx.exp + (t0 & (kEdgeDepth - 1)) * L::EXW);
            for (int i = lane; i < kWfBatch * L::EXW / 4; i += 32) dst[i] = src[i];
            __threadfence();
            __syncwarp();
            if (lane == 0) st_release_u64(flags_g + (size_t)c * 2, P.seq_base + (unsigned long long)(n + kBatchBias + 1));
        }
        if (g == NSTG - 1 && t0 + kWfBatch - 1 == tau_e) {
            // drain: rows finished during the last three batches
            __syncwarp();
            const int col = gp - S;
            const bool colok = (col >= 1) && (col <= H - 2);
            for (int q = tau_e + 1 - 48; q <= tau_e; ++q) {
                const int row = q - kToff - S - j0w;
                if (colok && row >= 1 && row <= W - 2)
                    F.out[(size_t)row * H + col] = t_out[(q & (kTileRows - 1)) * 32 + lane];
            }
        }
    };

#pragma unroll 1
    for (int T = T0; T <= T1; T += 2) {
        const int t = T - off_w;   // even
        const bool active = (t >= tau_b) && (t <= tau_e);
        if (active) {
            if ((t & (kWfBatch - 1)) == 0) batch_begin(t);
            wf_step<S, MODE>(x, Ua, Ub, Da, t, lane, gp, fix, cx);
        }
        __syncthreads();
        if (active) {
            wf_step<S, MODE>(x, Ub, Ua, Db, t + 1, lane, gp, fix, cx);
            if (((t + 1) & (kWfBatch - 1)) == kWfBatch - 1) batch_end(t + 1 - (kWfBatch - 1));
        }
        __syncthreads();
    }
    cp_async_wait_all();
}

template <int S>
size_t wf_smem_bytes(int nstr, int nstg)
{
    using L = WfLayout<S>;
    return sizeof(float) * ((size_t)2 * nstg * kEdgeDepth * L::EXW + (size_t)(nstr - 1) * nstg * kMidDepth * L::EXW +
                            (size_t)nstr * 3 * kTileFloats + (size_t)nstr * (nstg - 1) * 2 * kHandFloats);
}

template <int S, int MODE, bool HAS_X0>
cudaError_t wf_launch_inst(const SweepLaunch &Lc, SweepScratch &scr, cudaStream_t st)
{
    using L = WfLayout<S>;
    auto kern = wf_kernel<S, MODE, HAS_X0>;
    const int H = Lc.H;
    const int nstg = Lc.stages;
    const int total_levels = S * nstg;
    const int NS = (H + total_levels + 31) / 32;     // strips: the last stage's frame is shifted left by all earlier sweeps
    int dev = 0, nsm = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    if (nstg < 1 || nstg > kMaxStages || nstg * 32 > L::kMaxThreads) return cudaErrorInvalidValue;

    int nstr = Lc.strips_per_cta > 0 ? Lc.strips_per_cta : 1;
    while (nstr > 1 && nstr * nstg * 32 > L::kMaxThreads) nstr >>= 1;
    if (nstr > NS) nstr = NS;
    int ncta = 0;
    size_t smem = 0;
    for (;;) {
        ncta = (NS + nstr - 1) / nstr;
        smem = wf_smem_bytes<S>(nstr, nstg);
        bool fits = smem <= 227 * 1024 && ncta <= kWfMaxCtas;
        if (fits) {
            e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
            int per_sm = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, nstr * nstg * 32, smem);
            if (e != cudaSuccess) return e;
            fits = (long long)ncta * Lc.nfields <= (long long)per_sm * nsm;
        }
        if (fits) break;
        if ((nstr * 2) * nstg * 32 > L::kMaxThreads) return cudaErrorCooperativeLaunchTooLarge;
        nstr *= 2;
    }
    const size_t need = (size_t)Lc.nfields * nstg * ncta * L::kBoundaryFloats;
    if (need > scr.xchg_floats) {
        // grow-only scratch for the global boundary rings (first use of a larger configuration)
        e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) return e;
        if (scr.xchg) cudaFree(scr.xchg);
        scr.xchg = nullptr;
        scr.xchg_floats = 0;
        e = cudaMalloc(&scr.xchg, need * sizeof(float));
        if (e != cudaSuccess) return e;
        scr.xchg_floats = need;
    }

    scr.seq += 1;
    WfKernelParams P;
    for (int i = 0; i < kMaxSweepFields; ++i) P.fld[i] = Lc.fld[i < Lc.nfields ? i : 0];
    P.W = Lc.W;
    P.H = Lc.H;
    P.nstr = nstr;
    P.nstg = nstg;
    P.ncta = ncta;
    P.stage_off = ((2 * S + 2) + 15) / 16 * 16;
    P.a = Lc.a;
    P.c_recip = Lc.c_recip;
    P.omega = Lc.omega;
    P.xchg = scr.xchg;
    P.flags = scr.flags;
    P.seq_base = scr.seq << 32;
    void *args[] = {&P};
    return cudaLaunchCooperativeKernel((const void *)kern, dim3(ncta, Lc.nfields), dim3(nstr * nstg * 32), args, smem, st);
}

template <int S>
cudaError_t wf_launch_levels(const SweepLaunch &Lc, SweepScratch &scr, cudaStream_t st)
{
    if (Lc.mode == kSor) return wf_launch_inst<S, kSor, true>(Lc, scr, st);
    bool has_x0 = Lc.fld[0].x0 != nullptr;
    for (int i = 1; i < Lc.nfields; ++i)
        if ((Lc.fld[i].x0 != nullptr) != has_x0) return cudaErrorInvalidValue;
    if (has_x0) return wf_launch_inst<S, kGaussSeidel, true>(Lc, scr, st);
    return wf_launch_inst<S, kGaussSeidel, false>(Lc, scr, st);
}

}  // namespace

size_t sweep_flag_words() { return (size_t)kMaxSweepFields * kMaxStages * kWfMaxCtas * 2; }

int sweep_max_stages(int levels)
{
    switch (levels) {
    case 1: case 2: case 5: return 512 / 32;
    case 10: case 20: return 256 / 32;
    default: return 0;
    }
}

cudaError_t launch_sweep_pass(const SweepLaunch &L, SweepScratch &scr, cudaStream_t st, int *launches)
{
    if (L.nfields < 1 || L.nfields > kMaxSweepFields || L.W < 3 || L.H < 3) return cudaErrorInvalidValue;
    if (launches) *launches += 1;
    switch (L.levels) {
    case 1: return wf_launch_levels<1>(L, scr, st);
    case 2: return wf_launch_levels<2>(L, scr, st);
    case 5: return wf_launch_levels<5>(L, scr, st);
    case 10: return wf_launch_levels<10>(L, scr, st);
    case 20: return wf_launch_levels<20>(L, scr, st);
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace fsb
