// Several passes of FluidSim::step chained inside ONE cooperative launch (sm_100a).
//
// Why: an exact-order sweep pass (wavefront.cuh) is a wave that crosses the grid's columns from left to right; the
// CTA that owns the last columns starts when the wave arrives and finishes one full column walk later, so a
// stand-alone launch spends more time filling and draining the pipeline than running it full.  Between two passes
// of velocityStep (src/fluid_sim.cpp:370-406) nothing forces that drain: diffuse -> divergence -> pressure sweeps ->
// gradient only ever need the previous pass's results a few columns to the right of the cell being worked on.  So
// this kernel keeps every CTA resident across the passes ("phases") and lets each CTA go on to the next phase as
// soon as the columns it needs are done, instead of waiting for the whole grid.  The stagger between CTAs is paid
// once per launch, not once per pass.
//
//   * a phase is a list of work items, one per block of columns (all rows); item `i` of every phase is run by
//     CTA `i mod gridDim.x`, phases in order.  Sweep phases (wf_pass) have at most gridDim.x items, which therefore
//     run concurrently and hand the strip chain over exactly as the CTAs of a stand-alone launch do;
//   * done[phase][item] holds the launch number once the item's results are in global memory (release store
//     after a CTA barrier).  Before an item starts, its CTA waits (acquire loads) for the items of earlier phases
//     whose results it reads (RAW) or whose inputs it is about to overwrite (WAR), found from column ranges;
//   * the pointwise phases in between (divergence, gradient + boundaries) evaluate wall values on the fly from the
//     adjacent interior cell, as setBoundary() (src/fluid_sim.cpp:347-361) defines them, so no separate boundary
//     pass sits between two phases.  The final walls and corners of u and v are written by the gradient phase.
//
// Every wait loop carries a watchdog (global timer): a protocol error traps instead of hanging the GPU.
// All arithmetic uses round-to-nearest intrinsics in the reference's association order (no FMA).
#include "wavefront.cuh"
#include "gather.cuh"

#include <algorithm>

namespace fsb {

namespace {

constexpr int kChainThreads = 320;      // 8 compute warps + 2 transfer warps: the widest sweep pass
constexpr int kChainMaxPhases = 24;
constexpr int kChainMaxItems = 1024;    // items per phase the done[] array is sized for
constexpr int kChainMaxDeps = 4;
constexpr int kChainMaxSweeps = 7;      // sweep phases per launch: diffusion + up to three passes per projection
constexpr int kChainMaxGsPasses = 3;
constexpr int kChainCols = 64;          // columns of a pointwise work item

enum ChainKind { kPhSor = 1, kPhGs = 2, kPhDiv = 3, kPhGrad = 4, kPhAdvect = 5, kPhGsX0 = 6, kPhBndP = 7 };

// "wait until phase `phase` has finished every item whose result columns meet [b0 - lo, b1 + hi]", where [b0, b1]
// are the waiting item's own columns: the ones it reads (from_out = 0) or the ones it writes (from_out = 1).
struct ChainDep { int phase, from_out, lo, hi; };

struct ChainPhase {
    int kind;
    int nitems;      // work items = ncl * nf
    int ncl;         // column ranges
    int nf;          // fields per column range (sweep passes over several fields: one item per field)
    int cpv;         // columns per range
    int shift;       // range c reads columns [c*cpv, (c+1)*cpv) and writes [c*cpv - shift, (c+1)*cpv - shift)
    int wf;          // sweep phases: index into ChainParams::wf
    int ndeps;
    ChainDep dep[kChainMaxDeps];
    // pointwise phases
    //   divergence: u, v -> q                      gradient: q = pressure, u / v updated in place, then boundaries;
    //   advection : dst_a (dst_b) = src_a (src_b) traced back through u, v, then boundaries ba (bb) on dst
    float *u, *v;
    float *q;
    float *dst_a, *dst_b;
    const float *src_a, *src_b;
    float *bnd_extra;   // gradient: one more field whose boundaries (b = 0, interior read as v + 0.0f) are set here
    float c0;           // -0.5f * h (divergence) / h (gradient) / -dt0 (advection)
    int reach_phase;    // advection: the gradient phase that published u, v (its per-block max |v| bounds the trace)
    int ba, bb;         // advection: setBoundary's b of dst_a / dst_b
    int pz;             // advection: interior taps of src_a are read as (v + 0.0f) (diffuse with a == 0)
};

struct ChainParams {
    int W, H, nphases;
    int smem_floats;                // dynamic shared memory of the launch, in floats
    unsigned epoch;                 // number of this launch (done[] words hold the epoch of their last completion)
    unsigned *done;                 // [kChainMaxPhases][kChainMaxItems]
    unsigned long long *err;        // watchdog report: (phase << 32 | item) + 1 of the first wait that timed out
    unsigned long long timeout_ns;
    unsigned long long *trace;      // [kChainMaxPhases][kChainMaxItems][3] global-timer ns: wait begins, work begins, work ends
    float *vmax;                    // [2][kChainMaxPhases][kChainMaxItems] max |v|, max |u| of a block after its gradient phase
    ChainPhase ph[kChainMaxPhases];
    WfKernelParams wf[kChainMaxSweeps];
};
static_assert(sizeof(ChainParams) <= 16000, "kernel parameter space (CUDA 12.1+: up to 32764 bytes)");

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_release_u32(unsigned *p, unsigned v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ unsigned long long global_timer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ int floor_div(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

// Spins until done-word `f` carries this launch's epoch; traps if that takes longer than the watchdog allows.
__device__ __forceinline__ void chain_poll(const ChainParams &C, const unsigned *f, const int p, const int cl,
                                           unsigned long long &t0)
{
    unsigned spins = 0;
    while ((int)(ld_acquire_u32(f) - C.epoch) < 0) {
        __nanosleep(64);
        if ((++spins & 0x3ffu) == 0) {
            const unsigned long long now = global_timer_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > C.timeout_ns) {
                atomicCAS(C.err, 0ull, (((unsigned long long)p << 32) | (unsigned)cl) + 1ull);
                __threadfence_system();
                __trap();
            }
        }
    }
}

// Blocks the CTA until the dependencies of item (phase p, column range cl) are met.  Thread d polls dependency d.
__device__ __forceinline__ void chain_wait(const ChainParams &C, const int p, const int cl)
{
    const ChainPhase &ph = C.ph[p];
    const int d = threadIdx.x;
    if (d < ph.ndeps) {
        const ChainDep dep = ph.dep[d];
        const ChainPhase &q = C.ph[dep.phase];
        const int b0 = cl * ph.cpv - (dep.from_out ? ph.shift : 0);
        const int b1 = b0 + ph.cpv - 1;
        int a = floor_div(b0 - dep.lo + q.shift, q.cpv), b = floor_div(b1 + dep.hi + q.shift, q.cpv);
        a = max(a, 0);
        b = min(b, q.ncl - 1);
        const unsigned *flags = C.done + (size_t)dep.phase * kChainMaxItems;
        unsigned long long t0 = 0;
        for (int x = a * q.nf; x < (b + 1) * q.nf; ++x) chain_poll(C, flags + x, p, cl, t0);
    }
    __syncthreads();
}

// setBoundary(b, x) (src/fluid_sim.cpp:347-367) on the columns [c0, c0 + 64) of x, whose interior cells this CTA has
// just written (the caller placed a CTA barrier after them).  Walls, then the corners the block contains, each
// rebuilt from the interior so that no ordering between wall and corner threads is needed.  pz: interior values are
// read as (v + 0.0f).
__device__ __forceinline__ void chain_boundary(float *x, const int b, const int pz, const int W, const int H, const int c0)
{
    const int tid = threadIdx.x, nt = blockDim.x;
    const float si = (b == 1) ? -1.0f : 1.0f, sj = (b == 2) ? -1.0f : 1.0f;
    auto in = [&](int i, int j) {
        const float v = __ldcg(x + (size_t)i * H + j);
        return pz ? __fadd_rn(v, 0.0f) : v;
    };
    for (int t = tid; t < 2 * kChainCols; t += nt) {
        const int j = c0 + (t % kChainCols);
        if (j < 1 || j > H - 2) continue;
        const int i = (t < kChainCols) ? 0 : W - 1, iin = (t < kChainCols) ? 1 : W - 2;
        const float v = in(iin, j);
        x[(size_t)i * H + j] = (b == 1) ? -v : v;                  // :356-360
    }
    for (int side = 0; side < 2; ++side) {
        // a wall column is written by the block that owns the interior column it mirrors (the CTA that has just
        // written it), which need not be the block the wall column itself falls into
        const int j = side ? H - 1 : 0, jin = side ? H - 2 : 1;
        if (jin < c0 || jin >= c0 + kChainCols) continue;
        for (int i = 1 + tid; i <= W - 2; i += nt) {
            const float v = in(i, jin);
            x[(size_t)i * H + j] = (b == 2) ? -v : v;              // :350-354
        }
        if (tid < 2) {
            // corner = 0.5f * (x[iin][j] + x[i][jin]): a j-wall cell and an i-wall cell (:363-366)
            const int i = tid ? W - 1 : 0, iin = tid ? W - 2 : 1;
            const float v = in(iin, jin);
            x[(size_t)i * H + j] = __fmul_rn(0.5f, __fadd_rn(__fmul_rn(sj, v), __fmul_rn(si, v)));
        }
    }
}

// ---- divergence (src/fluid_sim.cpp:307-314) on the columns [c0, c0 + 64) ---------------------------------------------
// A warp walks a run of rows with the rows above / below kept in registers; every lane owns two columns (c0 + lane,
// c0 + 32 + lane), the j-neighbours come from the neighbouring lanes (two edge loads per row for the columns outside
// the block).  Wall cells are read as setBoundary(1, u) / setBoundary(2, v) would have left them:
//   u[0][j] = -u[1][j], u[W-1][j] = -u[W-2][j], v[i][0] = -v[i][1], v[i][H-1] = -v[i][H-2]   (:350-360)
constexpr int kPwRun = 64;   // rows per warp run
constexpr int kPwBatch = 4;  // rows whose loads are in flight together

__device__ __forceinline__ void chain_divergence(const ChainPhase &ph, const int W, const int H, const int cl)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int c0 = cl * kChainCols;
    const float *u = ph.u, *v = ph.v;
    float *div = ph.q;
    const float mhh = ph.c0;
    const int ja = c0 + lane, jb = c0 + 32 + lane;
    const bool ina = ja >= 1 && ja <= H - 2, inb = jb >= 1 && jb <= H - 2;
    const bool edge_l = (lane == 0) && (c0 - 1 >= 1), edge_r = (lane == 31) && (c0 + 64 <= H - 2);
    const int je = (lane == 0) ? c0 - 1 : c0 + 64;
    for (int r0 = 1 + kPwRun * w; r0 <= W - 2; r0 += kPwRun * nw) {
        const int r1 = min(r0 + kPwRun, W - 1);   // rows [r0, r1)
        const size_t row0 = (size_t)r0 * H;
        float uca = ina ? __ldcg(u + row0 + ja) : 0.0f, ucb = inb ? __ldcg(u + row0 + jb) : 0.0f;
        float uma, umb;
        if (r0 == 1) {
            uma = -uca;
            umb = -ucb;
        } else {
            uma = ina ? __ldcg(u + row0 - H + ja) : 0.0f;
            umb = inb ? __ldcg(u + row0 - H + jb) : 0.0f;
        }
        for (int i0 = r0; i0 < r1; i0 += kPwBatch) {
            float upa[kPwBatch], upb[kPwBatch], vca[kPwBatch], vcb[kPwBatch], ve[kPwBatch];
#pragma unroll
            for (int k = 0; k < kPwBatch; ++k) {
                const int i = i0 + k;
                const size_t row = (size_t)i * H;
                const bool rv = i < r1;
                const bool nx = rv && (i + 1 <= W - 2);
                upa[k] = (nx && ina) ? __ldcg(u + row + H + ja) : 0.0f;
                upb[k] = (nx && inb) ? __ldcg(u + row + H + jb) : 0.0f;
                vca[k] = (rv && ina) ? __ldcg(v + row + ja) : 0.0f;
                vcb[k] = (rv && inb) ? __ldcg(v + row + jb) : 0.0f;
                ve[k] = (rv && (edge_l || edge_r)) ? __ldcg(v + row + je) : 0.0f;
            }
#pragma unroll
            for (int k = 0; k < kPwBatch; ++k) {
                const int i = i0 + k;
                if (i >= r1) break;                       // warp-uniform
                const size_t row = (size_t)i * H;
                const float nua = (i + 1 == W - 1) ? -uca : upa[k];
                const float nub = (i + 1 == W - 1) ? -ucb : upb[k];
                // j-neighbours from the neighbouring lanes
                float la = __shfl_up_sync(0xffffffffu, vca[k], 1), ra = __shfl_down_sync(0xffffffffu, vca[k], 1);
                float lb = __shfl_up_sync(0xffffffffu, vcb[k], 1), rb = __shfl_down_sync(0xffffffffu, vcb[k], 1);
                const float b_first = __shfl_sync(0xffffffffu, vcb[k], 0), a_last = __shfl_sync(0xffffffffu, vca[k], 31);
                if (lane == 0) { la = ve[k]; lb = a_last; }
                if (lane == 31) { ra = b_first; rb = ve[k]; }
                if (ja - 1 == 0) la = -vca[k];
                if (ja + 1 == H - 1) ra = -vca[k];
                if (jb - 1 == 0) lb = -vcb[k];
                if (jb + 1 == H - 1) rb = -vcb[k];
                // (-0.5f*h) * (((u[i+1][j] - u[i-1][j]) + v[i][j+1]) - v[i][j-1])
                if (ina) div[row + ja] = __fmul_rn(mhh, __fsub_rn(__fadd_rn(__fsub_rn(nua, uma), ra), la));
                if (inb) div[row + jb] = __fmul_rn(mhh, __fsub_rn(__fadd_rn(__fsub_rn(nub, umb), rb), lb));
                uma = uca; umb = ucb;
                uca = nua; ucb = nub;
            }
        }
    }
}

// ---- pressure gradient (src/fluid_sim.cpp:334-341) on the columns [c0, c0 + 64), then setBoundary(1, u) and
// setBoundary(2, v) (:342-343) for those columns.  Pressure walls are read as setBoundary(0, p) defines them
// (p[0][j] = p[1][j], p[i][0] = p[i][1], ...), so the pressure array's own wall cells are never touched.
__device__ __forceinline__ void chain_gradient(const ChainPhase &ph, const int W, const int H, const int cl, float *vmax_out,
                                               float *umax_out)
{
    __shared__ float s_vmax[kChainThreads / 32], s_umax[kChainThreads / 32];
    float vmax = 0.0f, umax = 0.0f;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int c0 = cl * kChainCols;
    float *u = ph.u, *v = ph.v;
    const float *p = ph.q;
    const float h = ph.c0;
    const int ja = c0 + lane, jb = c0 + 32 + lane;
    const bool ina = ja >= 1 && ja <= H - 2, inb = jb >= 1 && jb <= H - 2;
    const bool edge_l = (lane == 0) && (c0 - 1 >= 1), edge_r = (lane == 31) && (c0 + 64 <= H - 2);
    const int je = (lane == 0) ? c0 - 1 : c0 + 64;
    for (int r0 = 1 + kPwRun * w; r0 <= W - 2; r0 += kPwRun * nw) {
        const int r1 = min(r0 + kPwRun, W - 1);
        const size_t row0 = (size_t)r0 * H;
        float pca = ina ? __ldcg(p + row0 + ja) : 0.0f, pcb = inb ? __ldcg(p + row0 + jb) : 0.0f;
        float pma, pmb;
        if (r0 == 1) {
            pma = pca;
            pmb = pcb;
        } else {
            pma = ina ? __ldcg(p + row0 - H + ja) : 0.0f;
            pmb = inb ? __ldcg(p + row0 - H + jb) : 0.0f;
        }
        float pe = (edge_l || edge_r) ? __ldcg(p + row0 + je) : 0.0f;   // this row's pressure just outside the block
        for (int i0 = r0; i0 < r1; i0 += kPwBatch) {
            float ppa[kPwBatch], ppb[kPwBatch], pen[kPwBatch], ua[kPwBatch], ub[kPwBatch], va[kPwBatch], vb[kPwBatch];
#pragma unroll
            for (int k = 0; k < kPwBatch; ++k) {
                const int i = i0 + k;
                const size_t row = (size_t)i * H;
                const bool rv = i < r1;
                const bool nx = rv && (i + 1 <= W - 2);
                ppa[k] = (nx && ina) ? __ldcg(p + row + H + ja) : 0.0f;
                ppb[k] = (nx && inb) ? __ldcg(p + row + H + jb) : 0.0f;
                pen[k] = (nx && (edge_l || edge_r)) ? __ldcg(p + row + H + je) : 0.0f;
                ua[k] = (rv && ina) ? __ldcg(u + row + ja) : 0.0f;
                ub[k] = (rv && inb) ? __ldcg(u + row + jb) : 0.0f;
                va[k] = (rv && ina) ? __ldcg(v + row + ja) : 0.0f;
                vb[k] = (rv && inb) ? __ldcg(v + row + jb) : 0.0f;
            }
#pragma unroll
            for (int k = 0; k < kPwBatch; ++k) {
                const int i = i0 + k;
                if (i >= r1) break;
                const size_t row = (size_t)i * H;
                const float npa = (i + 1 == W - 1) ? pca : ppa[k];
                const float npb = (i + 1 == W - 1) ? pcb : ppb[k];
                float la = __shfl_up_sync(0xffffffffu, pca, 1), ra = __shfl_down_sync(0xffffffffu, pca, 1);
                float lb = __shfl_up_sync(0xffffffffu, pcb, 1), rb = __shfl_down_sync(0xffffffffu, pcb, 1);
                const float b_first = __shfl_sync(0xffffffffu, pcb, 0), a_last = __shfl_sync(0xffffffffu, pca, 31);
                if (lane == 0) { la = pe; lb = a_last; }
                if (lane == 31) { ra = b_first; rb = pe; }
                if (ja - 1 == 0) la = pca;
                if (ja + 1 == H - 1) ra = pca;
                if (jb - 1 == 0) lb = pcb;
                if (jb + 1 == H - 1) rb = pcb;
                // u -= (0.5f*(p[i+1][j] - p[i-1][j])) / h ; v -= (0.5f*(p[i][j+1] - p[i][j-1])) / h
                if (ina) {
                    const float nv = __fsub_rn(va[k], __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(ra, la)), h));
                    const float nu = __fsub_rn(ua[k], __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(npa, pma)), h));
                    u[row + ja] = nu;
                    v[row + ja] = nv;
                    vmax = fmaxf(vmax, fabsf(nv));
                    umax = fmaxf(umax, fabsf(nu));
                }
                if (inb) {
                    const float nv = __fsub_rn(vb[k], __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(rb, lb)), h));
                    const float nu = __fsub_rn(ub[k], __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(npb, pmb)), h));
                    u[row + jb] = nu;
                    v[row + jb] = nv;
                    vmax = fmaxf(vmax, fabsf(nv));
                    umax = fmaxf(umax, fabsf(nu));
                }
                pma = pca; pmb = pcb;
                pca = npa; pcb = npb;
                pe = pen[k];
            }
        }
    }
    // largest |v| of the block (walls repeat interior magnitudes): bounds how many columns an advection trace that
    // starts in or passes through this block can move
    for (int o = 16; o > 0; o >>= 1) {
        vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
        umax = fmaxf(umax, __shfl_xor_sync(0xffffffffu, umax, o));
    }
    if (lane == 0) {
        s_vmax[w] = vmax;
        s_umax[w] = umax;
    }
    __syncthreads();   // every interior cell of the block carries its new velocity
    if (threadIdx.x == 0) {
        float m = 0.0f, mu = 0.0f;
        for (int k = 0; k < nw; ++k) {
            m = fmaxf(m, s_vmax[k]);
            mu = fmaxf(mu, s_umax[k]);
        }
        *vmax_out = m;
        *umax_out = mu;     // bounds the rows a trace can move: sizes the shared-memory tiles of the advection phase
    }
    chain_boundary(u, 1, 0, W, H, c0);    // setBoundary(1, u), :342
    chain_boundary(v, 2, 0, W, H, c0);    // setBoundary(2, v), :343
    // density "diffusion" with a == 0 leaves d + 0.0f in the interior and runs setBoundary(0, d): the
    // walls are set here, the + 0.0f is folded into the taps of the density advection
    if (ph.bnd_extra) chain_boundary(ph.bnd_extra, 0, 1, W, H, c0);
}

// ---- RK4 advection (src/fluid_sim.cpp:261-298) of the columns [c0, c0 + 64), then setBoundary on the result -------
// The back-trace may leave the block: before it starts, the CTA waits until the gradient phase has published every
// block the traces can reach.  A trace moves at most dt0 * max|v| columns per stage (bilinear samples never exceed
// the largest tap), so with m = max |v| over the blocks [b - k, b + k] the reach is ceil(dt0 * m) + 2 columns; k grows
// until that fits inside those blocks (one block for the 4-cell displacement of the bench input, the whole grid
// in the worst case).
__device__ __forceinline__ void chain_advect_wait(const ChainParams &C, const ChainPhase &ph, const int p, const int b,
                                                  const float dt0, int *reach_rows, int *reach_cols)
{
    __shared__ int s_k, s_ri, s_rj;
    const ChainPhase &g = C.ph[ph.reach_phase];
    const unsigned *flags = C.done + (size_t)ph.reach_phase * kChainMaxItems;
    const float *vm = C.vmax + (size_t)ph.reach_phase * kChainMaxItems;
    const int nb = g.ncl;
    int k = 1, waited_lo = b + 1, waited_hi = b;   // blocks [waited_lo, waited_hi] are known to be done
    for (;;) {
        const int lo = max(b - k, 0), hi = min(b + k, nb - 1);
        unsigned long long t0 = 0;
        for (int x = lo + (int)threadIdx.x; x <= hi; x += blockDim.x)
            if (x < waited_lo || x > waited_hi) chain_poll(C, flags + x, p, b, t0);
        __syncthreads();
        waited_lo = lo;
        waited_hi = hi;
        if (threadIdx.x == 0) {
            float m = 0.0f, mu = 0.0f;
            for (int x = lo; x <= hi; ++x) {
                m = fmaxf(m, __ldcg(vm + x));
                mu = fmaxf(mu, __ldcg(vm + (size_t)kChainMaxPhases * kChainMaxItems + x));
            }
            const float r = dt0 * m, ru = dt0 * mu;
            int need = nb;                                      // non-finite velocities: wait for everything
            if (r < 1.0e8f) need = ((int)ceilf(r) + 2 + kChainCols - 1) / kChainCols;
            s_k = max(need, 1);
            s_rj = (r < 1.0e8f) ? (int)ceilf(r) + 2 : (1 << 30);
            s_ri = (ru < 1.0e8f) ? (int)ceilf(ru) + 2 : (1 << 30);
        }
        __syncthreads();
        const int need = s_k;
        *reach_rows = s_ri;
        *reach_cols = s_rj;
        __syncthreads();
        if (need <= k || (lo == 0 && hi == nb - 1)) break;
        k = need;
    }
}

// ---- the same advection with the fields staged in shared memory -----------------------------------------------------
// A CTA is alone on its SM and nearly all of the SM's memory is shared memory (the sweep phases need it), so the L1
// that would serve the 36 taps of a cell is tiny and the plain version waits on L2 for most of them (57 % of its
// stall samples are long-scoreboard).  When the traces of the block stay within `ri` rows and `rj` columns (from the
// per-block velocity maxima, see chain_advect_wait), the rows a stage of 40 output rows can reach are copied into shared
// memory with cp.async (double buffered) and every tap is a shared-memory load.  Same taps, same arithmetic.
constexpr int kAdvRows = 40;

template <bool PZ>
__device__ __forceinline__ float sample_tile(const float *T, const Taps &t, const int row0, const int cb, const int TC,
                                             const int W, const int H)
{
    const float *r0 = T + (t.i0 - row0) * TC + (t.j0 - cb);
    float f00 = r0[0], f01 = r0[1], f10 = r0[TC], f11 = r0[TC + 1];
    if (PZ) {   // density only: interior cells are read as d + 0.0f (diffuse with a == 0, src/fluid_sim.cpp:56-57)
        const bool i0in = t.i0 >= 1 && t.i0 <= W - 2, i1in = t.i0 + 1 >= 1 && t.i0 + 1 <= W - 2;
        const bool j0in = t.j0 >= 1 && t.j0 <= H - 2, j1in = t.j0 + 1 >= 1 && t.j0 + 1 <= H - 2;
        if (i0in && j0in) f00 = __fadd_rn(f00, 0.0f);
        if (i0in && j1in) f01 = __fadd_rn(f01, 0.0f);
        if (i1in && j0in) f10 = __fadd_rn(f10, 0.0f);
        if (i1in && j1in) f11 = __fadd_rn(f11, 0.0f);
    }
    const float a = __fadd_rn(__fmul_rn(t.t0, f00), __fmul_rn(t.t1, f01));
    const float b = __fadd_rn(__fmul_rn(t.t0, f10), __fmul_rn(t.t1, f11));
    return __fadd_rn(__fmul_rn(t.s0, a), __fmul_rn(t.s1, b));
}

// returns false (nothing done) when the tiles for this reach do not fit the shared memory of the launch
template <bool TWO, bool PZ>
__device__ __forceinline__ bool chain_advect_tiled(const ChainPhase &ph, const int W, const int H, const int cl, float *smem,
                                                   const int smem_floats, const int ri, const int rj)
{
    constexpr int NARR = TWO ? 2 : 3;            // velocity: u, v are sources as well; density: u, v and the density
    if ((H & 3) != 0 || ri > 64 || rj > 60) return false;
    const int halo_c = (rj + 1 + 3) & ~3;        // columns staged on each side (taps j0, j0 + 1), 16-byte aligned
    const int halo_r = ri + 1;
    const int TC = kChainCols + 2 * halo_c, TR = kAdvRows + 2 * halo_r;
    const int stage_floats = NARR * TR * TC;
    if (2 * stage_floats > smem_floats) return false;
    const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, w = tid >> 5;
    const int c0 = cl * kChainCols, cb = c0 - halo_c;
    const float *arr[3] = {ph.u, ph.v, ph.src_a};
    float *dst_a = ph.dst_a, *dst_b = ph.dst_b;
    const float ndt0 = ph.c0;
    const float xm = (float)W - 1.5f, ym = (float)H - 1.5f;
    const int nq = TC >> 2;                      // 16-byte pieces per tile row
    const int my_q = tid % nq, my_r = tid / nq, r_step = nt / nq;
    const int nstages = (W - 2 + kAdvRows - 1) / kAdvRows;
    auto issue = [&](int k) {
        if (k < nstages && my_r < r_step) {
            float *slot = smem + (k & 1) * stage_floats;
            const int row_lo = 1 + k * kAdvRows - halo_r;
            const int col = cb + 4 * my_q;
            if (col >= 0 && col < H) {
#pragma unroll
                for (int a = 0; a < NARR; ++a)
                    for (int r = my_r; r < TR; r += r_step) {
                        const int row = row_lo + r;
                        if (row < 0 || row > W - 1) continue;
                        cp_async_f32x4(slot + (a * TR + r) * TC + 4 * my_q, arr[a] + (size_t)row * H + col, true);
                    }
            }
        }
        cp_async_commit();
    };
    issue(0);
    for (int k = 0; k < nstages; ++k) {
        issue(k + 1);                            // the other buffer: consumed in iteration k - 1 (barrier at its end)
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();                         // stage k has landed for every thread
        const float *slot = smem + (k & 1) * stage_floats;
        const float *Tu = slot, *Tv = slot + TR * TC, *Ts = slot + 2 * TR * TC;
        const int row0 = 1 + k * kAdvRows - halo_r;
        constexpr int N = 4;                     // independent traces per thread: two columns x two rows
#pragma unroll 1
        for (int it = 0; it < 2; ++it) {
            int ci[N], cj[N];
            bool ok[N];
            float x[N], y[N], k1x[N], k1y[N], k2x[N], k2y[N], k3x[N], k3y[N], k4x[N], k4y[N];
            Taps t[N];
#pragma unroll
            for (int n = 0; n < N; ++n) {
                const int r = 4 * w + 2 * it + (n >> 1);
                ci[n] = 1 + k * kAdvRows + r;
                cj[n] = c0 + 32 * (n & 1) + lane;
                ok[n] = r < kAdvRows && ci[n] <= W - 2 && cj[n] >= 1 && cj[n] <= H - 2;
                // cells outside the interior trace from the first cell of the stage's block (inside the tile), store nothing
                x[n] = ok[n] ? (float)ci[n] : (float)(1 + k * kAdvRows);
                y[n] = ok[n] ? (float)cj[n] : (float)max(c0, 1);
                t[n] = make_taps(x[n], y[n], xm, ym);
            }
#pragma unroll
            for (int n = 0; n < N; ++n) {
                k1x[n] = __fmul_rn(sample_tile<false>(Tu, t[n], row0, cb, TC, W, H), ndt0);
                k1y[n] = __fmul_rn(sample_tile<false>(Tv, t[n], row0, cb, TC, W, H), ndt0);
            }
#pragma unroll
            for (int n = 0; n < N; ++n)
                t[n] = make_taps(__fadd_rn(x[n], __fmul_rn(k1x[n], 0.5f)), __fadd_rn(y[n], __fmul_rn(k1y[n], 0.5f)), xm, ym);
#pragma unroll
            for (int n = 0; n < N; ++n) {
                k2x[n] = __fmul_rn(sample_tile<false>(Tu, t[n], row0, cb, TC, W, H), ndt0);
                k2y[n] = __fmul_rn(sample_tile<false>(Tv, t[n], row0, cb, TC, W, H), ndt0);
            }
#pragma unroll
            for (int n = 0; n < N; ++n)
                t[n] = make_taps(__fadd_rn(x[n], __fmul_rn(k2x[n], 0.5f)), __fadd_rn(y[n], __fmul_rn(k2y[n], 0.5f)), xm, ym);
#pragma unroll
            for (int n = 0; n < N; ++n) {
                k3x[n] = __fmul_rn(sample_tile<false>(Tu, t[n], row0, cb, TC, W, H), ndt0);
                k3y[n] = __fmul_rn(sample_tile<false>(Tv, t[n], row0, cb, TC, W, H), ndt0);
            }
#pragma unroll
            for (int n = 0; n < N; ++n) t[n] = make_taps(__fadd_rn(x[n], k3x[n]), __fadd_rn(y[n], k3y[n]), xm, ym);
#pragma unroll
            for (int n = 0; n < N; ++n) {
                k4x[n] = __fmul_rn(sample_tile<false>(Tu, t[n], row0, cb, TC, W, H), ndt0);
                k4y[n] = __fmul_rn(sample_tile<false>(Tv, t[n], row0, cb, TC, W, H), ndt0);
            }
#pragma unroll
            for (int n = 0; n < N; ++n) {
                const float dx = __fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(k1x[n], __fmul_rn(2.0f, k2x[n])), __fmul_rn(2.0f, k3x[n])), k4x[n]), 6.0f);
                const float dy = __fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(k1y[n], __fmul_rn(2.0f, k2y[n])), __fmul_rn(2.0f, k3y[n])), k4y[n]), 6.0f);
                t[n] = make_taps(__fadd_rn(x[n], dx), __fadd_rn(y[n], dy), xm, ym);
            }
#pragma unroll
            for (int n = 0; n < N; ++n) {
                const float ra = sample_tile<PZ>(TWO ? Tu : Ts, t[n], row0, cb, TC, W, H);
                const float rb = TWO ? sample_tile<false>(Tv, t[n], row0, cb, TC, W, H) : 0.0f;
                if (ok[n]) {
                    const size_t c = (size_t)ci[n] * H + cj[n];
                    dst_a[c] = ra;
                    if (TWO) dst_b[c] = rb;
                }
            }
        }
        __syncthreads();                         // stage k is consumed: its buffer may be refilled
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    chain_boundary(dst_a, ph.ba, 0, W, H, c0);
    if (TWO) chain_boundary(dst_b, ph.bb, 0, W, H, c0);
    return true;
}

template <bool TWO, bool PZ>
__device__ __forceinline__ void chain_advect(const ChainPhase &ph, const int W, const int H, const int cl)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int c0 = cl * kChainCols;
    const GridView G{W, H, H, 0, 0, H};
    const float *u = ph.u, *v = ph.v;
    const float *src_a = ph.src_a, *src_b = ph.src_b;
    float *dst_a = ph.dst_a, *dst_b = ph.dst_b;
    const float ndt0 = ph.c0;
    const float xm = (float)W - 1.5f, ym = (float)H - 1.5f;
    constexpr int N = 4;     // independent traces per thread: two columns x two rows
    for (int i0 = 1 + 2 * w; i0 <= W - 2; i0 += 2 * nw) {
        int ci[N], cj[N];
        bool ok[N];
        float x[N], y[N], k1x[N], k1y[N], k2x[N], k2y[N], k3x[N], k3y[N], k4x[N], k4y[N];
        Taps t[N];
#pragma unroll
        for (int n = 0; n < N; ++n) {
            ci[n] = i0 + (n >> 1);
            cj[n] = c0 + 32 * (n & 1) + lane;
            ok[n] = ci[n] <= W - 2 && cj[n] >= 1 && cj[n] <= H - 2;
            // cells outside the interior trace from a harmless in-grid position and store nothing
            x[n] = ok[n] ? (float)ci[n] : 1.0f;
            y[n] = ok[n] ? (float)cj[n] : 1.0f;
            t[n] = make_taps(x[n], y[n], xm, ym);
        }
#pragma unroll
        for (int n = 0; n < N; ++n) {
            k1x[n] = __fmul_rn(sample<false, int, false>(u, t[n], G), ndt0);
            k1y[n] = __fmul_rn(sample<false, int, false>(v, t[n], G), ndt0);
        }
#pragma unroll
        for (int n = 0; n < N; ++n)
            t[n] = make_taps(__fadd_rn(x[n], __fmul_rn(k1x[n], 0.5f)), __fadd_rn(y[n], __fmul_rn(k1y[n], 0.5f)), xm, ym);
#pragma unroll
        for (int n = 0; n < N; ++n) {
            k2x[n] = __fmul_rn(sample<false, int, false>(u, t[n], G), ndt0);
            k2y[n] = __fmul_rn(sample<false, int, false>(v, t[n], G), ndt0);
        }
#pragma unroll
        for (int n = 0; n < N; ++n)
            t[n] = make_taps(__fadd_rn(x[n], __fmul_rn(k2x[n], 0.5f)), __fadd_rn(y[n], __fmul_rn(k2y[n], 0.5f)), xm, ym);
#pragma unroll
        for (int n = 0; n < N; ++n) {
            k3x[n] = __fmul_rn(sample<false, int, false>(u, t[n], G), ndt0);
            k3y[n] = __fmul_rn(sample<false, int, false>(v, t[n], G), ndt0);
        }
#pragma unroll
        for (int n = 0; n < N; ++n) t[n] = make_taps(__fadd_rn(x[n], k3x[n]), __fadd_rn(y[n], k3y[n]), xm, ym);
#pragma unroll
        for (int n = 0; n < N; ++n) {
            k4x[n] = __fmul_rn(sample<false, int, false>(u, t[n], G), ndt0);
            k4y[n] = __fmul_rn(sample<false, int, false>(v, t[n], G), ndt0);
        }
#pragma unroll
        for (int n = 0; n < N; ++n) {
            // ((k1 + 2*k2) + 2*k3 + k4) / 6, componentwise (:290)
            const float dx = __fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(k1x[n], __fmul_rn(2.0f, k2x[n])), __fmul_rn(2.0f, k3x[n])), k4x[n]), 6.0f);
            const float dy = __fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(k1y[n], __fmul_rn(2.0f, k2y[n])), __fmul_rn(2.0f, k3y[n])), k4y[n]), 6.0f);
            t[n] = make_taps(__fadd_rn(x[n], dx), __fadd_rn(y[n], dy), xm, ym);
        }
#pragma unroll
        for (int n = 0; n < N; ++n) {
            const float ra = sample<PZ, int, false>(src_a, t[n], G);
            const float rb = TWO ? sample<PZ, int, false>(src_b, t[n], G) : 0.0f;
            if (ok[n]) {
                const size_t c = (size_t)ci[n] * H + cj[n];
                dst_a[c] = ra;
                if (TWO) dst_b[c] = rb;
            }
        }
    }
    __syncthreads();
    chain_boundary(dst_a, ph.ba, 0, W, H, c0);
    if (TWO) chain_boundary(dst_b, ph.bb, 0, W, H, c0);
}

template <int SD, int SP>
__global__ void __launch_bounds__(kChainThreads, 1) chain_kernel(const __grid_constant__ ChainParams C)
{
    extern __shared__ float4 smem4[];
    float *smem = reinterpret_cast<float *>(smem4);
    for (int p = 0; p < C.nphases; ++p) {
        const ChainPhase &ph = C.ph[p];
        for (int item = blockIdx.x; item < ph.nitems; item += gridDim.x) {
            const int cl = item / ph.nf, fi = item - cl * ph.nf;
            unsigned long long *tr = C.trace + ((size_t)p * kChainMaxItems + item) * 3;
            if (threadIdx.x == 0) tr[0] = global_timer_ns();
            chain_wait(C, p, cl);
            int reach_rows = 1 << 30, reach_cols = 1 << 30;
            if (ph.kind == kPhAdvect) chain_advect_wait(C, ph, p, cl, -ph.c0, &reach_rows, &reach_cols);
            if (threadIdx.x == 0) tr[1] = global_timer_ns();
            if (ph.kind == kPhSor) {
                if constexpr (SD > 0) wf_pass<SD, kSor, true>(C.wf[ph.wf], cl, fi, smem);
            } else if (ph.kind == kPhGs) {
                wf_pass<SP, kGaussSeidel, false>(C.wf[ph.wf], cl, fi, smem);
            } else if (ph.kind == kPhGsX0) {
                wf_pass<SP, kGaussSeidel, true>(C.wf[ph.wf], cl, fi, smem);     // a later pass of the same projection
            } else if (ph.kind == kPhBndP) {
                chain_boundary(ph.q, 0, 0, C.W, C.H, cl * kChainCols);           // setBoundary(0, p) between two passes (:330)
            } else if (ph.kind == kPhDiv) {
                chain_divergence(ph, C.W, C.H, cl);
            } else if (ph.kind == kPhGrad) {
                chain_gradient(ph, C.W, C.H, cl, C.vmax + (size_t)p * kChainMaxItems + item,
                               C.vmax + (size_t)(kChainMaxPhases + p) * kChainMaxItems + item);
            } else if (ph.kind == kPhAdvect) {
                if (ph.dst_b) {
                    if (!chain_advect_tiled<true, false>(ph, C.W, C.H, cl, smem, C.smem_floats, reach_rows, reach_cols))
                        chain_advect<true, false>(ph, C.W, C.H, cl);
                } else if (ph.pz) {
                    if (!chain_advect_tiled<false, true>(ph, C.W, C.H, cl, smem, C.smem_floats, reach_rows, reach_cols))
                        chain_advect<false, true>(ph, C.W, C.H, cl);
                } else {
                    if (!chain_advect_tiled<false, false>(ph, C.W, C.H, cl, smem, C.smem_floats, reach_rows, reach_cols))
                        chain_advect<false, false>(ph, C.W, C.H, cl);
                }
            }
            __syncthreads();   // the item's results are written, its shared memory is free
            if (threadIdx.x == 0) {
                __threadfence();
                st_release_u32(C.done + (size_t)p * kChainMaxItems + item, C.epoch);
                tr[2] = global_timer_ns();
            }
        }
    }
}

// Strips per CTA of a sweep phase inside the chain kernel: as many as the 8 compute warps, shared memory and
// the grid allow (a hand-over inside a CTA is far cheaper than one through the global rings).
template <int S>
bool chain_pick(int H, int nstg, int nfields, int max_ctas, int *nstr_out, int *ncta_out, size_t *smem_out)
{
    using L = WfLayout<S>;
    if (nstg < 1 || nstg > L::kMaxStagesS) return false;
    const int NS = (H + S * nstg + 31) / 32;
    const int max_warps = kChainThreads / 32 - 2;
    int nstr = 4;
    while (nstr > 1 && (nstr * nstg > max_warps || wf_smem_bytes<S>(nstr, nstg) > 227 * 1024 || nstr > NS)) nstr >>= 1;
    if (nstr * nstg > max_warps || wf_smem_bytes<S>(nstr, nstg) > 227 * 1024) return false;
    const int ncta = (NS + nstr - 1) / nstr;
    if ((long long)ncta * nfields > max_ctas || ncta > kWfMaxCtas) return false;
    *nstr_out = nstr;
    *ncta_out = ncta;
    *smem_out = wf_smem_bytes<S>(nstr, nstg);
    return true;
}

constexpr int kChainSD = 5, kChainSP = 10;   // sweeps per warp of the diffusion / pressure phases (the instantiated pair)

struct ChainPlan {
    bool ok = false;
    int W = 0, H = 0, D = 0, P = 0;
    int sor_nstr = 0, sor_ncta = 0, gs_nstr = 0, gs_ncta = 0, grid = 0;
    int gs_npass = 0, gs_stages[kChainMaxGsPasses] = {0, 0, 0};   // passes of one projection (gs_ncta: of the widest pass)
    size_t smem = 0;
    size_t sor_ring_floats = 0, gs_ring_floats = 0;
};

}  // namespace

struct ChainState {
    ChainPlan plan[2];                    // [0] with the diffusion phase, [1] without
    unsigned char *buf = nullptr;         // [done | err | vmax | trace x 2 | flags x 3 | rings sor | rings gs x 2]
    size_t bytes = 0;
    float *extra[2] = {nullptr, nullptr}; // pressure / divergence of the second projection of a whole-step launch
    size_t extra_floats = 0;
    unsigned epoch = 0;
    int last_nphases[2] = {0, 0};         // of the most recent launch of each kind (for chain_read_trace)
};

namespace {

// Can (D diffusion sweeps of u and v +) P pressure sweeps run as phases of one launch on this grid?
ChainPlan make_plan(int W, int H, int D, int P)
{
    ChainPlan pl;
    pl.W = W; pl.H = H; pl.D = D; pl.P = P;
    if (P < kChainSP || P % kChainSP != 0) return pl;
    if (D != 0 && D % kChainSD != 0) return pl;
    int dev = 0, nsm = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return pl;
    if (cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return pl;
    size_t smem_gs = 0, smem_sor = 0;
    {
        // Passes of one projection: the fewest strips per CTA for which a pass fits the SMs (more SMs at work, fewer
        // warps sharing one), as many chained warps per strip as the 8 compute warps and the shared memory allow, the
        // P / 10 stages spread evenly over the passes.
        using L = WfLayout<kChainSP>;
        const int T = P / kChainSP;
        bool found = false;
        for (int nstr = 1; nstr <= 4 && !found; nstr *= 2) {
            int maxst = std::min((kChainThreads / 32 - 2) / nstr, L::kMaxStagesS);
            while (maxst >= 1 && wf_smem_bytes<kChainSP>(nstr, maxst) > 227 * 1024) --maxst;
            if (maxst < 1) continue;
            const int npass = (T + maxst - 1) / maxst;
            if (npass > kChainMaxGsPasses) continue;
            const int widest = (T + npass - 1) / npass;
            const int NS = (H + kChainSP * widest + 31) / 32;
            const int ncta = (NS + std::min(nstr, NS) - 1) / std::min(nstr, NS);
            if (ncta > nsm || ncta > kWfMaxCtas) continue;
            pl.gs_nstr = std::min(nstr, NS);
            pl.gs_ncta = ncta;
            pl.gs_npass = npass;
            for (int k = 0; k < npass; ++k) pl.gs_stages[k] = T / npass + (k < T % npass ? 1 : 0);
            smem_gs = wf_smem_bytes<kChainSP>(pl.gs_nstr, widest);
            found = true;
        }
        if (!found) return pl;
    }
    if (D > 0 && !chain_pick<kChainSD>(H, D / kChainSD, 2, nsm, &pl.sor_nstr, &pl.sor_ncta, &smem_sor)) return pl;
    // the advection phases stage their fields in the same shared memory: give them room on small grids too
    pl.smem = std::max(std::max(smem_gs, smem_sor), (size_t)160 * 1024);
    const int nblk = (H + kChainCols - 1) / kChainCols;
    pl.grid = std::max(std::max(pl.gs_ncta, 2 * pl.sor_ncta), std::min(nblk, nsm));
    if (pl.grid > nsm || nblk > kChainMaxItems || pl.grid > kChainMaxItems) return pl;
    auto kern = D > 0 ? chain_kernel<kChainSD, kChainSP> : chain_kernel<0, kChainSP>;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem) != cudaSuccess) {
        (void)cudaGetLastError();
        return pl;
    }
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kChainThreads, pl.smem) != cudaSuccess) {
        (void)cudaGetLastError();
        return pl;
    }
    if ((long long)per_sm * nsm < pl.grid) return pl;
    pl.gs_ring_floats = (size_t)pl.gs_stages[0] * pl.gs_ncta * WfLayout<kChainSP>::kBoundaryFloats;   // of the widest pass
    pl.sor_ring_floats = D > 0 ? (size_t)2 * (D / kChainSD) * pl.sor_ncta * WfLayout<kChainSD>::kBoundaryFloats : 0;
    pl.ok = true;
    return pl;
}

void fill_wf(WfKernelParams &P, int W, int H, int nstr, int nstg, int ncta, float a, float c_recip, float omega,
             float *rings, unsigned long long *flags, unsigned long long seq)
{
    P.W = W;
    P.H = H;
    P.pitch = H;
    P.jbase = 0;
    P.nstr = nstr;
    P.nstg = nstg;
    P.ncta = ncta;
    P.c_base = 0;
    P.a = a;
    P.c_recip = c_recip;
    P.omega = omega;
    P.xchg = rings;
    P.flags = flags;
    P.seq_base = seq << 32;
    P.link = SweepLink{};
    P.link_seq_base = 0;
}

constexpr size_t kDoneBytes = (size_t)kChainMaxPhases * kChainMaxItems * sizeof(unsigned);
constexpr size_t kVmaxBytes = (size_t)2 * kChainMaxPhases * kChainMaxItems * sizeof(float);   // max |v|, then max |u|
constexpr size_t kTraceOne = (size_t)kChainMaxPhases * kChainMaxItems * 3 * sizeof(unsigned long long);

}  // namespace

void chain_state_destroy(ChainState *cs)
{
    if (!cs) return;
    if (cs->buf) cudaFree(cs->buf);
    for (int k = 0; k < 2; ++k)
        if (cs->extra[k]) cudaFree(cs->extra[k]);
    delete cs;
}

// Passes of FluidSim::step as phases of one launch (see the head of this file):
//   diffuse(u), diffuse(v)                               when A.sor_sweeps > 0          (src/fluid_sim.cpp:390-391)
//   project(u, v, p, div)                                always                         (:394)
//   advect(u), advect(v) into (p, div), project again    when A.with_advect             (:401-405)
//   densityStep with DIFFUSION == 0                      when A.with_density            (:409-422)
// Returns cudaErrorNotSupported when this grid / these sweep counts have no chained configuration (the caller then
// runs the passes as separate launches).
cudaError_t launch_chain(const ChainArgs &A, ChainState *&cs, SweepScratch &scr, cudaStream_t st, bool query_only)
{
    if (!cs) cs = new ChainState();
    const int kind = A.sor_sweeps > 0 ? 0 : 1;
    ChainPlan &pl = cs->plan[kind];
    if (pl.W != A.W || pl.H != A.H || pl.D != A.sor_sweeps || pl.P != A.gs_sweeps) pl = make_plan(A.W, A.H, A.sor_sweeps, A.gs_sweeps);
    if (!pl.ok) return cudaErrorNotSupported;
    if (A.with_density && !A.with_advect) return cudaErrorInvalidValue;
    if (query_only) return cudaSuccess;

    cudaError_t e;
    if (A.with_advect) {
        const size_t nfl = (size_t)A.W * A.H;
        if (nfl > cs->extra_floats) {
            e = cudaStreamSynchronize(st);
            if (e != cudaSuccess) return e;
            for (int k = 0; k < 2; ++k) {
                if (cs->extra[k]) cudaFree(cs->extra[k]);
                cs->extra[k] = nullptr;
            }
            cs->extra_floats = 0;
            for (int k = 0; k < 2; ++k) {
                e = cudaMalloc(&cs->extra[k], nfl * sizeof(float));
                if (e != cudaSuccess) return e;
            }
            cs->extra_floats = nfl;
        }
    }
    // scratch: done words, watchdog word, per-block maxima, timelines, progress words and boundary rings of the sweep phases
    const size_t flag_bytes = sweep_flag_words() * sizeof(unsigned long long);
    const size_t sor_rings = std::max(cs->plan[0].sor_ring_floats, pl.sor_ring_floats) * sizeof(float);
    const size_t gs_rings = std::max(std::max(cs->plan[0].gs_ring_floats, cs->plan[1].gs_ring_floats), pl.gs_ring_floats) * sizeof(float);
    const size_t need = kDoneBytes + 256 + kVmaxBytes + 2 * kTraceOne + kChainMaxSweeps * flag_bytes + sor_rings +
                        2 * kChainMaxGsPasses * gs_rings;
    if (need > cs->bytes) {
        e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) return e;
        if (cs->buf) cudaFree(cs->buf);
        cs->buf = nullptr;
        cs->bytes = 0;
        e = cudaMalloc(&cs->buf, need);
        if (e != cudaSuccess) return e;
        e = cudaMemsetAsync(cs->buf, 0, need, st);   // done epoch 0 and ring tag 0 are never produced
        if (e != cudaSuccess) return e;
        cs->bytes = need;
    }
    unsigned char *b = cs->buf;
    unsigned *done = reinterpret_cast<unsigned *>(b);
    unsigned long long *err = reinterpret_cast<unsigned long long *>(b + kDoneBytes);
    float *vmax = reinterpret_cast<float *>(b + kDoneBytes + 256);
    unsigned long long *trace = reinterpret_cast<unsigned long long *>(b + kDoneBytes + 256 + kVmaxBytes + kind * kTraceOne);
    unsigned char *b2 = b + kDoneBytes + 256 + kVmaxBytes + 2 * kTraceOne;
    unsigned long long *flags[kChainMaxSweeps];
    for (int k = 0; k < kChainMaxSweeps; ++k) flags[k] = reinterpret_cast<unsigned long long *>(b2 + k * flag_bytes);
    float *rings_sor = reinterpret_cast<float *>(b2 + kChainMaxSweeps * flag_bytes);
    float *rings_gs[2 * kChainMaxGsPasses];
    for (int k = 0; k < 2 * kChainMaxGsPasses; ++k)
        rings_gs[k] = reinterpret_cast<float *>(b2 + kChainMaxSweeps * flag_bytes + sor_rings + k * gs_rings);

    ChainParams C{};
    C.W = A.W;
    C.H = A.H;
    C.smem_floats = (int)(pl.smem / sizeof(float));
    C.epoch = ++cs->epoch;
    C.done = done;
    C.err = err;
    C.timeout_ns = 20ull * 1000 * 1000 * 1000;
    C.trace = trace;
    C.vmax = vmax;
    const int nblk = (A.H + kChainCols - 1) / kChainCols;
    const float h = 1.0f / (float)A.W;            // src/fluid_sim.cpp:304
    int np = 0, nwf = 0;

    auto pointwise_phase = [&](int kind_) -> ChainPhase & {
        ChainPhase &ph = C.ph[np++];
        ph.kind = kind_;
        ph.ncl = ph.nitems = nblk;
        ph.nf = 1;
        ph.cpv = kChainCols;
        ph.shift = 0;
        ph.ndeps = 0;
        return ph;
    };
    // divergence -> pressure sweeps -> gradient (+ boundaries) of (u, v) with scratch (p, div); `after` = phase that
    // produced u, v (or -1); returns the index of the gradient phase
    auto projection = [&](float *u, float *v, float *p, float *div, int after, int gs_slot, float *bnd_extra) -> int {
        const int ph_div = np;
        {
            ChainPhase &ph = pointwise_phase(kPhDiv);
            ph.u = u;
            ph.v = v;
            ph.q = div;
            ph.c0 = -0.5f * h;                    // evaluated first (:311)
            if (after >= 0) ph.dep[ph.ndeps++] = ChainDep{after, 0, 1, 1};    // u, v of the block and one column either side
        }
        int ph_gs = -1, ph_prev = ph_div;
        for (int k = 0; k < pl.gs_npass; ++k) {
            if (k > 0) {
                // setBoundary(0, p) after the previous pass's last sweep (:330): the next pass reads the walls of its
                // first sweep from the array
                ChainPhase &pb = pointwise_phase(kPhBndP);
                pb.q = p;
                pb.dep[pb.ndeps++] = ChainDep{ph_gs, 0, 0, 0};
                ph_prev = np - 1;
            }
            const int ph_last = ph_gs;
            ph_gs = np;
            ChainPhase &ph = C.ph[np++];
            ph.kind = k == 0 ? kPhGs : kPhGsX0;
            const int nstrips = (A.H + kChainSP * pl.gs_stages[k] + 31) / 32;     // the last stage's frame is shifted left by the pass's sweeps
            const int ncta = (nstrips + pl.gs_nstr - 1) / pl.gs_nstr;
            ph.ncl = ph.nitems = ncta;
            ph.nf = 1;
            ph.cpv = 32 * pl.gs_nstr;
            ph.shift = kChainSP * pl.gs_stages[k];
            ph.wf = nwf;
            ph.ndeps = 0;
            ph.dep[ph.ndeps++] = ChainDep{ph_prev, 0, 0, 0};                  // divergence (first pass) / walls of p (later passes) of the columns it loads
            if (k > 0) ph.dep[ph.ndeps++] = ChainDep{ph_last, 1, 0, 0};       // the previous pass has produced the columns this one overwrites
            WfKernelParams &P = C.wf[nwf++];
            const int slot = gs_slot * kChainMaxGsPasses + k;
            fill_wf(P, A.W, A.H, pl.gs_nstr, pl.gs_stages[k], ncta, 0.0f, 1.0f, 1.0f, rings_gs[slot], flags[1 + slot], ++scr.seq);
            P.fld[0] = SweepField{k == 0 ? nullptr : p, div, p, 1.0f, 1.0f};  // p starts at 0 (:313, :316)
            P.fld[1] = P.fld[2] = P.fld[0];
        }
        ChainPhase &ph = pointwise_phase(kPhGrad);
        ph.u = u;
        ph.v = v;
        ph.q = p;
        ph.c0 = h;
        ph.bnd_extra = bnd_extra;
        ph.dep[ph.ndeps++] = ChainDep{ph_gs, 0, 1, 1};    // pressure of the block and one column either side
        ph.dep[ph.ndeps++] = ChainDep{ph_div, 1, 1, 1};   // the divergence blocks that read the u, v about to change
        return np - 1;
    };

    int ph_sor = -1;
    if (A.sor_sweeps > 0) {
        ChainPhase &ph = C.ph[np];
        ph.kind = kPhSor;
        ph.ncl = pl.sor_ncta;
        ph.nf = 2;
        ph.nitems = 2 * pl.sor_ncta;
        ph.cpv = 32 * pl.sor_nstr;
        ph.shift = A.sor_sweeps;
        ph.wf = nwf;
        ph.ndeps = 0;
        WfKernelParams &P = C.wf[nwf++];
        fill_wf(P, A.W, A.H, pl.sor_nstr, A.sor_sweeps / kChainSD, pl.sor_ncta, A.a, A.c_recip, A.omega, rings_sor, flags[0], ++scr.seq);
        P.fld[0] = SweepField{A.u, A.u, A.u, -1.0f, 1.0f};    // setBoundary(1, u): negated on the i-walls
        P.fld[1] = SweepField{A.v, A.v, A.v, 1.0f, -1.0f};    // setBoundary(2, v): negated on the j-walls
        P.fld[2] = P.fld[0];
        ph_sor = np++;
    }
    const int ph_grad1 = projection(A.u, A.v, A.p, A.div, ph_sor, 0, nullptr);
    if (A.with_advect) {
        // advect(1, u), advect(2, v) (:401-402) into the buffers the first projection used as pressure / divergence
        const int ph_gs1 = ph_grad1 - 1;
        const int ph_adv = np;
        {
            ChainPhase &ph = pointwise_phase(kPhAdvect);
            ph.u = A.u;
            ph.v = A.v;
            ph.src_a = A.u;
            ph.src_b = A.v;
            ph.dst_a = A.p;
            ph.dst_b = A.div;
            ph.ba = 1;
            ph.bb = 2;
            ph.pz = 0;
            ph.c0 = -(A.dt * (float)A.W);              // -dt0, dt0 = dt * width (:264)
            ph.reach_phase = ph_grad1;
            ph.dep[ph.ndeps++] = ChainDep{ph_grad1, 1, 1, 1};                                // gradient blocks still reading p here
            const int sh = C.ph[ph_gs1].shift;                                             // of the projection's last pass
            ph.dep[ph.ndeps++] = ChainDep{ph_gs1, 1, sh, -sh};                             // sweep items that load div here
        }
        // the second projection works in scratch of its own: u, v stay intact for traces that start far away
        const float *dens = A.with_density ? A.d : nullptr;
        const int ph_grad2 = projection(A.p, A.div, cs->extra[0], cs->extra[1], ph_adv, 1,
                                        (A.with_density && A.density_pz) ? A.d : nullptr);
        if (dens) {
            ChainPhase &ph = pointwise_phase(kPhAdvect);
            ph.u = A.p;
            ph.v = A.div;
            ph.src_a = A.d;
            ph.dst_a = A.nd;
            ph.ba = 0;
            ph.pz = A.density_pz;
            ph.c0 = -(A.dt * (float)A.W);
            ph.reach_phase = ph_grad2;
        }
    }
    C.nphases = np;
    cs->last_nphases[kind] = np;

    void *args[] = {&C};
    const void *kern = A.sor_sweeps > 0 ? (const void *)chain_kernel<kChainSD, kChainSP> : (const void *)chain_kernel<0, kChainSP>;
    return cudaLaunchCooperativeKernel(kern, dim3(pl.grid), dim3(kChainThreads), args, pl.smem, st);
}

// Start / end times of every work item of the most recent chained launch of one kind (which = 0: launches that begin
// with the diffusion phase, 1: the others): out[(phase * max_items + item) * 3 + k] in nanoseconds of the GPU's global
// timer (k = 0 wait begins, 1 work begins, 2 results published); 0 = item did not run.  The caller synchronises first.
cudaError_t chain_read_trace(ChainState *cs, int which, unsigned long long *out, int max_phases, int max_items, int *nphases)
{
    *nphases = 0;
    if (!cs || !cs->buf || which < 0 || which > 1) return cudaSuccess;
    const unsigned long long *trace =
        reinterpret_cast<const unsigned long long *>(cs->buf + kDoneBytes + 256 + kVmaxBytes + which * kTraceOne);
    const int np = std::min(cs->last_nphases[which], max_phases), ni = std::min(max_items, kChainMaxItems);
    for (int p = 0; p < np; ++p) {
        cudaError_t e = cudaMemcpy(out + (size_t)p * max_items * 3, trace + (size_t)p * kChainMaxItems * 3,
                                   (size_t)ni * 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) return e;
    }
    *nphases = np;
    return cudaSuccess;
}

}  // namespace fsb
