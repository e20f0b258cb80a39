// Exact-order Gauss-Seidel / SOR sweeps as a strip-pipelined wavefront (sm_100a).
//
// What it replaces: the lexicographic, in-place relaxation loops of the reference,
//   diffuse()  src/fluid_sim.cpp:50-61   (SOR, omega 1.5, `k < 5`)
//   project()  src/fluid_sim.cpp:319-331 (Gauss-Seidel, `k < 20`)
// including the setBoundary() call that follows every sweep (:60, :330).  The update order
// of those loops decides the result, so it is reproduced exactly:
//
//   * a warp owns a strip of 32 "global lanes"; lane g keeps S+1 values x[0..S] in
//     registers, x[k] being the sweep-k value of cell (i, j) = (t - 32 - k - g, g - k) at
//     global time step t.  Each step every lane moves one row down; level k trails level
//     k-1 by one row and one column.  With that skew all four neighbours a sweep needs
//     are values of step t-1: own registers (up / right) or the registers of lane g-1
//     (left / down), fetched with one shuffle per level.  Hence S sweeps are in flight at
//     once and every cell of the pass is read from and written to HBM exactly once.
//   * the right-hand side (divergence or diffusion source) travels with the cell through
//     a second shuffle pipeline (two steps per level).
//   * wall cells are evaluated in the pipeline with the value setBoundary() would have
//     left in the array at the moment the reference reads it.
//   * lane 0 of a strip takes its left neighbours from the previous strip: through a
//     shared-memory ring inside a CTA (one __syncthreads per step), through a global ring
//     with acquire/release progress flags between CTAs (batches of 16 steps).  CTAs of a
//     chain therefore run a few batches apart; the launch is cooperative so that every
//     CTA of a chain is resident while its right neighbour waits for it.
//
// All arithmetic uses explicit round-to-nearest intrinsics: no FMA contraction (F7).
#include "fsb_internal.h"

namespace fsb {

namespace {

constexpr int kToff = 32;        // time offset: lane g sits on row t - kToff - g at level 0
constexpr int kTileRows = 64;    // rows of a per-warp staging ring
constexpr int kTileFloats = kTileRows * 32;
constexpr int kEdgeDepth = 32;   // steps held by the CTA-edge rings (two batches)
constexpr int kMidDepth = 4;     // steps held by the rings between warps of a CTA

template <int S>
struct WfLayout {
    static constexpr int XW = (S + 1 + 3) & ~3;   // x[0..S], padded to float4
    static constexpr int DW = (S + 3) & ~3;       // rhs levels 0..S-1
    static constexpr int EXW = XW + DW;           // floats one lane hands to the next strip per step
};
constexpr int kExwMax = WfLayout<kWfMaxLevels>::EXW;
constexpr int kBoundaryFloats = kWfSlots * kWfBatch * kExwMax;

struct WfKernelParams {
    SweepField fld[kMaxSweepFields];
    int W, H, NW, ncta;
    float a, c_recip, omega;
    float *xchg;
    unsigned long long *flags;
    unsigned long long seq_base;
};

struct WfCtx {
    const float *imp;
    float *exp;
    int imp_mask, exp_mask;
    const float *t_rhs;
    const float *t_x0;
    float *t_out;
    int W, H;
    float si, sj, a, c_recip, omega;
};

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

// Which boundary logic a step needs (warp-uniform, chosen per batch of 16 steps):
//   kPathFast   every (lane, level) of the strip sits on an interior cell
//   kPathLeft   interior rows only; the strip contains the j == 0 wall (one level per lane)
//   kPathRight  interior rows only; the strip contains the j == H-1 wall
//   kPathFull   anything (first / last rows of the grid, tiny grids): all walls, branch-free
enum { kPathFast = 0, kPathLeft = 1, kPathRight = 2, kPathFull = 3 };

// One global time step of one strip.  U receives lane g-1's x of the previous step, Up is
// the U of the step before (needed for the centre value of SOR and for first-level walls).
template <int S, int MODE, bool HAS_X0, int PATH>
__device__ __forceinline__ void wf_step(float (&x)[S + 1], float (&U)[S + 1], const float (&Up)[S + 1],
                                        float (&Dc)[S + 1], const int t, const int lane, const int g,
                                        const WfCtx &cx)
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
    const int slot = ((t - lane) & (kTileRows - 1)) * 32 + lane;
    const float nx0 = HAS_X0 ? cx.t_x0[slot] : 0.0f;
    Dc[0] = cx.t_rhs[slot];
    const int ib = t - kToff - g;
    const float cold1 = Up[0];
    // The S updates of a step are independent of each other (they read only values of step t-1), so
    // they are written stage by stage across all levels: consecutive instructions belong to different
    // dependency chains, which lets a single warp keep its FP pipe busy.
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
    // Cells on a wall carry what setBoundary() (src/fluid_sim.cpp:347-361) left in the array at the
    // moment the reference reads them during sweep k: the array's own value before the first sweep,
    // +-(the adjacent interior cell's sweep k-1 value) afterwards.  Cells outside the grid are never
    // consumed by a cell inside it.
    if (PATH == kPathLeft) {
#pragma unroll
        for (int k = 1; k <= S; ++k) {
            const float wv = (k == 1) ? cold1 : __fmul_rn(cx.sj, x[k - 1]);
            acc[k] = (g == k) ? wv : acc[k];
        }
    } else if (PATH == kPathRight) {
        const int gw = g - (cx.H - 1);
#pragma unroll
        for (int k = 1; k <= S; ++k) acc[k] = (gw == k) ? __fmul_rn(cx.sj, U[k]) : acc[k];
    } else if (PATH == kPathFull) {
#pragma unroll
        for (int k = 1; k <= S; ++k) {
            const int i = ib - k;
            const int j = g - k;
            const bool inrow = (unsigned)(i - 1) <= (unsigned)(cx.W - 3);
            const bool incol = (unsigned)(j - 1) <= (unsigned)(cx.H - 3);
            const float wtop = (k == 1) ? cold1 : __fmul_rn(cx.si, U[k - 1]);
            const float wi = (i == 0) ? wtop : __fmul_rn(cx.si, x[k]);
            const bool on_i = (i == 0) | (i == cx.W - 1);
            const float wleft = (k == 1) ? cold1 : __fmul_rn(cx.sj, x[k - 1]);
            const float wj = (j == 0) ? wleft : __fmul_rn(cx.sj, U[k]);
            const bool on_j = (j == 0) | (j == cx.H - 1);
            const float row_in = incol ? acc[k] : (on_j ? wj : 0.0f);
            const float row_out = (on_i & incol) ? wi : 0.0f;
            acc[k] = inrow ? row_in : row_out;
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
    cx.t_out[slot] = x[S];
}

// Sixteen steps of one path (kept out of line per path so the batch loop stays small).
template <int S, int MODE, bool HAS_X0, int PATH>
__device__ __forceinline__ void wf_batch(float (&x)[S + 1], float (&Ua)[S + 1], float (&Ub)[S + 1], float (&Da)[S + 1],
                                         float (&Db)[S + 1], const int t0, const int lane, const int g,
                                         const WfCtx &cx)
{
#pragma unroll 1
    for (int s2 = 0; s2 < kWfBatch / 2; ++s2) {
        const int t = t0 + 2 * s2;
        wf_step<S, MODE, HAS_X0, PATH>(x, Ua, Ub, Da, t, lane, g, cx);
        __syncthreads();
        wf_step<S, MODE, HAS_X0, PATH>(x, Ub, Ua, Db, t + 1, lane, g, cx);
        __syncthreads();
    }
}

template <int S, int MODE, bool HAS_X0>
__global__ void __launch_bounds__(256, 1) wf_kernel(const WfKernelParams P)
{
    using L = WfLayout<S>;
    extern __shared__ float4 smem4[];
    float *smem = reinterpret_cast<float *>(smem4);

    const int NW = P.NW;
    const int nthreads = NW * 32;
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int w = tid >> 5;
    const int c = blockIdx.x;
    const int fi = blockIdx.y;
    const SweepField F = P.fld[fi];
    const int W = P.W, H = P.H;

    float *ring_in = smem;
    float *ring_out = ring_in + kEdgeDepth * L::EXW;
    float *ring_mid = ring_out + kEdgeDepth * L::EXW;
    float *tiles = ring_mid + (NW - 1) * kMidDepth * L::EXW;
    float *t_rhs = tiles + w * 3 * kTileFloats;
    float *t_x0 = t_rhs + kTileFloats;
    float *t_out = t_x0 + kTileFloats;

    {
        const int total = 2 * kEdgeDepth * L::EXW + (NW - 1) * kMidDepth * L::EXW + NW * 3 * kTileFloats;
        for (int i = tid; i < total; i += nthreads) smem[i] = 0.0f;
    }
    __syncthreads();

    WfCtx cx;
    cx.imp = (w == 0) ? ring_in : ring_mid + (w - 1) * kMidDepth * L::EXW;
    cx.imp_mask = (w == 0) ? (kEdgeDepth - 1) : (kMidDepth - 1);
    cx.exp = (w == NW - 1) ? ring_out : ring_mid + w * kMidDepth * L::EXW;
    cx.exp_mask = (w == NW - 1) ? (kEdgeDepth - 1) : (kMidDepth - 1);
    cx.t_rhs = t_rhs;
    cx.t_x0 = t_x0;
    cx.t_out = t_out;
    cx.W = W;
    cx.H = H;
    cx.si = F.si;
    cx.sj = F.sj;
    cx.a = P.a;
    cx.c_recip = P.c_recip;
    cx.omega = P.omega;

    const int j0_first = 32 * NW * c;
    const int j0 = j0_first + 32 * w;
    const int g = j0 + lane;
    const int t_last = (W - 2) + S + kToff + j0_first + 32 * NW - 1;
    const int n_first = j0_first / kWfBatch;
    const int n_last = t_last / kWfBatch;
    const int n_last_left = ((W - 2) + S + kToff + j0_first - 1) / kWfBatch;
    const int n_first_right = (j0_first + 32 * NW) / kWfBatch;

    unsigned long long *flags_f = P.flags + (size_t)fi * kWfMaxCtas * 2;
    float *xchg_f = P.xchg + (size_t)fi * P.ncta * kBoundaryFloats;

    float x[S + 1], Ua[S + 1], Ub[S + 1], Da[S + 1], Db[S + 1];
#pragma unroll
    for (int k = 0; k <= S; ++k) {
        x[k] = 0.0f; Ua[k] = 0.0f; Ub[k] = 0.0f; Da[k] = 0.0f; Db[k] = 0.0f;
    }

    const bool vec_ok = ((H & 3) == 0);
    const bool has_left = (j0 - S <= 0);            // some (lane, level) of the strip has column <= 0
    const bool has_right = (j0 + 30 >= H - 1);      // ... or column >= H-1

    for (int n = n_first; n <= n_last; ++n) {
        const int t0 = n * kWfBatch;

        // (a) left neighbours of this batch from the CTA on the left
        if (c > 0 && n <= n_last_left) {
            if (tid == 0) {
                const unsigned long long want = P.seq_base + (unsigned long long)(n + 1);
                const unsigned long long *pf = flags_f + (size_t)(c - 1) * 2;
                while (ld_acquire_u64(pf) < want) __nanosleep(40);
            }
            __syncthreads();
            const float4 *src = reinterpret_cast<const float4 *>(
                xchg_f + (size_t)(c - 1) * kBoundaryFloats + (size_t)(n % kWfSlots) * kWfBatch * kExwMax);
            float4 *dst = reinterpret_cast<float4 *>(ring_in + (t0 & (kEdgeDepth - 1)) * L::EXW);
            for (int i = tid; i < kWfBatch * L::EXW / 4; i += nthreads) dst[i] = __ldcg(src + i);
            __syncthreads();
            if (tid == 0) st_release_u64(flags_f + (size_t)(c - 1) * 2 + 1, P.seq_base + (unsigned long long)(n + 1));
        }

        // (b) stage the rows this strip reaches during the next batch
        cp_async_wait_all();
        {
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

        // (c) write back the result rows that every lane has finished
        {
            const int col = g - S;
            const bool colok = (col >= 1) && (col <= H - 2);
#pragma unroll 4
            for (int it = 0; it < kWfBatch; ++it) {
                const int q = t0 - 48 + it;
                const int row = q - kToff - S - j0;
                if (colok && row >= 1 && row <= W - 2)
                    F.out[(size_t)row * H + col] = t_out[(q & (kTileRows - 1)) * 32 + lane];
            }
        }
        __syncthreads();

        // (e) sixteen time steps; every warp executes the same number of barriers on every path
        const bool rows_interior = (t0 - kToff - j0 - 31 - S >= 1) && (t0 + kWfBatch - 1 - kToff - j0 - 1 <= W - 2);
        int path = kPathFull;
        if (rows_interior) {
            if (!has_left && !has_right) path = kPathFast;
            else if (has_left && !has_right && S <= 31 && j0 == 0) path = kPathLeft;
            else if (has_right && !has_left) path = kPathRight;
        }
        if (path == kPathFast) wf_batch<S, MODE, HAS_X0, kPathFast>(x, Ua, Ub, Da, Db, t0, lane, g, cx);
        else if (path == kPathLeft) wf_batch<S, MODE, HAS_X0, kPathLeft>(x, Ua, Ub, Da, Db, t0, lane, g, cx);
        else if (path == kPathRight) wf_batch<S, MODE, HAS_X0, kPathRight>(x, Ua, Ub, Da, Db, t0, lane, g, cx);
        else wf_batch<S, MODE, HAS_X0, kPathFull>(x, Ua, Ub, Da, Db, t0, lane, g, cx);

        // (f) hand this batch's right edge to the CTA on the right
        if (c + 1 < P.ncta) {
            if (n - kWfSlots >= n_first_right) {
                if (tid == 0) {
                    const unsigned long long want = P.seq_base + (unsigned long long)(n - kWfSlots + 1);
                    const unsigned long long *pf = flags_f + (size_t)c * 2 + 1;
                    while (ld_acquire_u64(pf) < want) __nanosleep(40);
                }
                __syncthreads();
            }
            float4 *dst = reinterpret_cast<float4 *>(
                xchg_f + (size_t)c * kBoundaryFloats + (size_t)(n % kWfSlots) * kWfBatch * kExwMax);
            const float4 *src = reinterpret_cast<const float4 *>(ring_out + (t0 & (kEdgeDepth - 1)) * L::EXW);
            for (int i = tid; i < kWfBatch * L::EXW / 4; i += nthreads) dst[i] = src[i];
            __threadfence();
            __syncthreads();
            if (tid == 0) st_release_u64(flags_f + (size_t)c * 2, P.seq_base + (unsigned long long)(n + 1));
        }
    }

    // drain: rows finished during the last three batches
    cp_async_wait_all();
    __syncthreads();
    {
        const int t_end = (n_last + 1) * kWfBatch;
        const int col = g - S;
        const bool colok = (col >= 1) && (col <= H - 2);
        for (int q = t_end - 48; q < t_end; ++q) {
            const int row = q - kToff - S - j0;
            if (colok && row >= 1 && row <= W - 2)
                F.out[(size_t)row * H + col] = t_out[(q & (kTileRows - 1)) * 32 + lane];
        }
    }
}

template <int S>
size_t wf_smem_bytes(int NW)
{
    using L = WfLayout<S>;
    return sizeof(float) * ((size_t)2 * kEdgeDepth * L::EXW + (size_t)(NW - 1) * kMidDepth * L::EXW +
                            (size_t)NW * 3 * kTileFloats);
}

template <int S, int MODE, bool HAS_X0>
cudaError_t wf_launch_inst(const SweepLaunch &Lc, SweepScratch &scr, cudaStream_t st)
{
    auto kern = wf_kernel<S, MODE, HAS_X0>;
    const int H = Lc.H;
    const int NS = (H + S + 31) / 32;
    int dev = 0, nsm = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;

    int NW = Lc.warps_per_cta > 0 ? Lc.warps_per_cta : 4;
    if (NW > 8) NW = 8;
    if (NW > NS) NW = NS;
    int ncta = 0;
    size_t smem = 0;
    for (;;) {
        ncta = (NS + NW - 1) / NW;
        smem = wf_smem_bytes<S>(NW);
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        int per_sm = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, NW * 32, smem);
        if (e != cudaSuccess) return e;
        if ((long long)ncta * Lc.nfields <= (long long)per_sm * nsm && ncta <= kWfMaxCtas) break;
        if (NW >= 8) return cudaErrorCooperativeLaunchTooLarge;
        NW *= 2;
    }
    if ((size_t)Lc.nfields * ncta * kBoundaryFloats > scr.xchg_floats) return cudaErrorInvalidValue;

    scr.seq += 1;
    WfKernelParams P;
    for (int i = 0; i < kMaxSweepFields; ++i) P.fld[i] = Lc.fld[i < Lc.nfields ? i : 0];
    P.W = Lc.W;
    P.H = Lc.H;
    P.NW = NW;
    P.ncta = ncta;
    P.a = Lc.a;
    P.c_recip = Lc.c_recip;
    P.omega = Lc.omega;
    P.xchg = scr.xchg;
    P.flags = scr.flags;
    P.seq_base = scr.seq << 32;
    void *args[] = {&P};
    return cudaLaunchCooperativeKernel((const void *)kern, dim3(ncta, Lc.nfields), dim3(NW * 32), args, smem, st);
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

size_t sweep_xchg_floats(int H)
{
    // worst case: one warp per CTA, shortest pass
    const size_t ns = (size_t)(H + kWfMaxLevels + 31) / 32 + 1;
    return (size_t)kMaxSweepFields * ns * kBoundaryFloats;
}

size_t sweep_flag_words() { return (size_t)kMaxSweepFields * kWfMaxCtas * 2; }

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
