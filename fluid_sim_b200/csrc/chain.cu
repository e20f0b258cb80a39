// Several passes of FluidSim::step -- on one GPU all of them -- chained inside ONE launch (sm_100a).
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
//   * on one GPU the whole step is such a launch, the noise included (chain_noise), and the launches of consecutive
//     steps overlap: launch n + 1 is given the SMs launch n's CTAs leave one by one (programmatic stream serialization)
//     and orders itself against launch n block by block through the same done-words (ChainPhase::xdep); the fields
//     rotate through enough buffers that what a launch overwrites was last read two launches ago (launch_chain).
//
// Every wait loop carries a watchdog (global timer): a protocol error traps instead of hanging the GPU.
// All arithmetic uses round-to-nearest intrinsics in the reference's association order (no FMA).
#include "wavefront.cuh"
#include "gather.cuh"
#include "noise.cuh"

#include <algorithm>

namespace fsb {

namespace {

constexpr int kChainThreads = 320;      // 8 compute warps + 2 transfer warps: the widest sweep pass
constexpr int kChainMaxPhases = 24;
constexpr int kChainMaxItems = 1088;    // items per phase (numbered along the whole grid) the done[] array is sized for:
                                        // 1024 column blocks, and the few more items a sweep phase's result shift adds
constexpr int kChainMaxDeps = 4;
constexpr int kChainMaxXDeps = 3;     // dependencies on the previous launch (overlapped steps)
constexpr int kChainMaxSweeps = 7;      // sweep phases per launch: diffusion + up to three passes per projection
constexpr int kChainMaxGsPasses = 3;
constexpr int kChainCols = 64;          // columns of a pointwise work item
constexpr int kChainMaxSources = kChainSources;   // addDensity / addVelocity calls a launch can apply itself (kernel parameters)

enum ChainKind { kPhSor = 1, kPhGs = 2, kPhDiv = 3, kPhGrad = 4, kPhAdvect = 5, kPhGsX0 = 6, kPhBndP = 7, kPhNoise = 8 };

// "wait until phase `phase` has finished every item whose result columns meet [b0 - lo, b1 + hi]", where [b0, b1]
// are the waiting item's own columns: the ones it reads (from_out = 0) or the ones it writes (from_out = 1).
struct ChainDep { int phase, from_out, lo, hi; };

// This rank's view of the grid (one GPU: the whole grid) and its neighbours in a multi-rank run.
struct ChainGrid {
    int W, H;                  // global grid
    int pitch, jbase;          // local arrays: floats per row, global column of local column 0
    int j0, j1;                // owned columns
    int cw;                    // columns of a pointwise work item: 64, or 32 when the owned range has few blocks
    int halo;                  // columns stored (and kept current by the neighbours' pushes) beyond the owned range
    int has_left, has_right;
    int jbase_l, jbase_r;      // the neighbours' jbase (same pitch everywhere)
};

struct ChainPhase {
    int kind;
    int item0;       // global index of this rank's first work item (items are numbered along the whole grid)
    int nitems;      // this rank's work items
    int ncl;         // column ranges of the whole grid (nitems of all ranks / nf)
    int nf;          // fields per column range (sweep passes over several fields: one item per field)
    int cpv;         // columns per range
    int shift;       // range c reads columns [c*cpv, (c+1)*cpv) and writes [c*cpv - shift, (c+1)*cpv - shift)
    int wf;          // sweep phases: index into ChainParams::wf
    int ndeps;
    ChainDep dep[kChainMaxDeps];
    // the same against the PREVIOUS launch (same phase numbering), used when two steps overlap: the items of the step
    // before whose results this item reads, or which still read what this item overwrites
    int nxdeps;
    ChainDep xdep[kChainMaxXDeps];
    // pointwise phases
    //   divergence: u, v -> q                      gradient: q = pressure, u / v updated in place, then boundaries;
    //   advection : dst_a (dst_b) = src_a (src_b) traced back through u, v, then boundaries ba (bb) on dst
    //   noise     : u, v = src_a, src_b + the step's noise, every cell of the block (walls included)
    float *u, *v;
    float *q;
    float *dst_a, *dst_b;
    const float *src_a, *src_b;
    float *bnd_extra;   // gradient: one more field whose boundaries (b = 0, interior read as v + 0.0f) are set here
    float c0;           // -0.5f * h (divergence) / h (gradient) / -dt0 (advection)
    // multi-rank: the same fields in the neighbours' memory (peer mappings).  Blocks within push_cols columns of a rank
    // boundary write their results there as well, so every rank reads only its own memory.
    float *peer_l[3], *peer_r[3];   // gradient: u, v, bnd_extra; advection: dst_a, dst_b
    int push_cols;
    int reach_phase;    // advection: the gradient phase that published u, v (its per-block max |v| bounds the trace)
    int ba, bb;         // advection: setBoundary's b of dst_a / dst_b
    float *snap;        // advection of one field: a second copy of the result (walls included), or nullptr
    int pz;             // advection: interior taps of src_a are read as (v + 0.0f) (diffuse with a == 0)
};

// One pending addDensity (field 0) / addVelocity (fields 1, 2: x and y component) call, src/fluid_sim.cpp:425-438
struct ChainSrc { int row, col, field; float amount; };

struct ChainParams {
    int W, H, nphases;
    ChainGrid g;
    int smem_floats;                // dynamic shared memory of the launch, in floats
    // multi-rank: the neighbours' done[] / vmax[] arrays (items near a rank boundary publish there too), the words in
    // which the neighbours announce the last step they have finished, and the words this rank announces its own in
    unsigned *done_l, *done_r;
    float *vmax_l, *vmax_r;
    unsigned *step_in;              // [0]: written by the left neighbour, [1]: by the right one
    unsigned *step_out_l, *step_out_r;
    unsigned *cta_count;            // [2], by launch parity: CTAs of the launch that have finished
    unsigned *completed;            // [2], by launch parity: number of the last launch all of whose CTAs have finished
    // steps that overlap (one GPU): launch n + 1 starts on the SMs the first CTAs of launch n have left, see launch_chain
    int xstep;                      // the previous launch was the step before this one: ChainPhase::xdep applies
    int pdl;                        // let the next launch of the stream start before this one has drained
    // noise phase (chain_noise)
    int nz_zero;                    // FSB_NOISE_ZERO: every cell gets + 0.0f
    uint32_t nz_base[kRandDeg];     // window of the rand() stream at the step's first output
    const uint32_t *nz_polys;       // jump polynomials
    // sources the noise phase applies before it adds the noise (velocity) / to the density the step starts from, in call
    // order per cell; `density`: that field
    int nsrc, nsrc_density;
    float *density;
    ChainSrc src[kChainMaxSources];
    unsigned epoch;                 // number of this launch (done[] words hold the epoch of their last completion)
    unsigned *done;                 // [kChainMaxPhases][kChainMaxItems]
    unsigned long long *err;        // watchdog report: (phase << 32 | item) + 1 of the first wait that timed out
    unsigned long long timeout_ns;
    unsigned long long *trace;      // [kChainMaxPhases][kChainMaxItems][3] global-timer ns: wait begins, work begins, work ends
    float *vmax;                    // [2][kChainMaxPhases][kChainMaxItems] max |v|, max |u| of a block after its gradient phase
                                    // (overlapping launches alternate between two such arrays)
    ChainPhase ph[kChainMaxPhases];
    WfKernelParams wf[kChainMaxSweeps];
};
static_assert(sizeof(ChainParams) <= 24000, "kernel parameter space (CUDA 12.1+: up to 32764 bytes)");

// system scope on a multi-rank launch: the word may have been written by a neighbouring GPU through peer memory
template <bool MR>
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p)
{
    unsigned v;
    if constexpr (MR) asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    else asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_release_sys_u32(unsigned *p, unsigned v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
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

// Spins until word `f` has reached `want` (words only grow); traps if that takes longer than the watchdog allows.
template <bool MR>
__device__ __forceinline__ void chain_poll(const ChainParams &C, const unsigned *f, const unsigned want, const int p, const int cl,
                                           unsigned long long &t0)
{
    unsigned spins = 0;
    while ((int)(ld_acquire_u32<MR>(f) - want) < 0) {
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

// does the item with result columns [lo, hi] lie within `dist` columns of the boundary to the left / right rank?
__device__ __forceinline__ bool near_left(const ChainGrid &g, int lo, int dist) { return g.has_left && lo < g.j0 + dist; }
__device__ __forceinline__ bool near_right(const ChainGrid &g, int hi, int dist) { return g.has_right && hi >= g.j1 - dist; }

// Blocks the CTA until the dependencies of item (phase p, column range cl, numbered along the whole grid) are met.
// Thread d polls dependency d; `wait_l` / `wait_r`: the item is about to write into the left / right neighbour's
// memory, which that rank may still be reading in its previous step: wait until it has announced the end of that step.
template <bool MR>
__device__ __forceinline__ void chain_wait(const ChainParams &C, const int p, const int cl, const bool wait_l, const bool wait_r)
{
    const ChainPhase &ph = C.ph[p];
    const int d = threadIdx.x;
    unsigned long long t0 = 0;
    auto wait_for = [&](const ChainDep dep, const unsigned want) {
        const ChainPhase &q = C.ph[dep.phase];
        const int b0 = cl * ph.cpv - (dep.from_out ? ph.shift : 0);
        const int b1 = b0 + ph.cpv - 1;
        int a = floor_div(b0 - dep.lo + q.shift, q.cpv), b = floor_div(b1 + dep.hi + q.shift, q.cpv);
        a = max(a, 0);
        b = min(b, q.ncl - 1);
        const unsigned *flags = C.done + (size_t)dep.phase * kChainMaxItems;
        for (int x = a * q.nf; x < (b + 1) * q.nf; ++x) chain_poll<MR>(C, flags + x, want, p, cl, t0);
    };
    if (d < ph.ndeps) {
        wait_for(ph.dep[d], C.epoch);
    } else if (!MR && d >= 8 && d - 8 < ph.nxdeps && C.xstep) {
        wait_for(ph.xdep[d - 8], C.epoch - 1);      // the step before, still running
    } else if (d == 32 && wait_l) {
        chain_poll<MR>(C, C.step_in + 0, C.epoch - 1, p, cl, t0);
    } else if (d == 33 && wait_r) {
        chain_poll<MR>(C, C.step_in + 1, C.epoch - 1, p, cl, t0);
    }
    __syncthreads();
}

// Where a pointwise item's results go besides this rank's arrays: the neighbours' halo copies of the same fields.
struct ChainPush {
    bool l, r;
    long long off_l, off_r;     // add to a local element index to get the element index in the neighbour's array
};
__device__ __forceinline__ ChainPush make_push(const ChainGrid &g, const ChainPhase &ph, const int c0)
{
    ChainPush P;
    P.l = ph.push_cols > 0 && near_left(g, c0, ph.push_cols);
    P.r = ph.push_cols > 0 && near_right(g, c0 + g.cw - 1, ph.push_cols);
    P.off_l = (long long)g.jbase - g.jbase_l;     // local column j - jbase  ->  j - jbase_l
    P.off_r = (long long)g.jbase - g.jbase_r;
    return P;
}
// MR: multi-rank launch (on one GPU the pushes are compiled out)
template <bool MR>
__device__ __forceinline__ void store_pushed(float *x, float *xl, float *xr, const ChainPush &P, const long long idx, const float v)
{
    x[idx] = v;
    if constexpr (MR) {
        if (P.l) xl[idx + P.off_l] = v;
        if (P.r) xr[idx + P.off_r] = v;
    }
}

#define CAT(g, i, j) ((long long)(i) * (g).pitch + ((j) - (g).jbase))

// setBoundary(b, x) (src/fluid_sim.cpp:347-367) on the columns [c0, c0 + cw) of x, whose interior cells this CTA has
// just written (the caller placed a CTA barrier after them).  Walls, then the corners the block contains, each
// rebuilt from the interior so that no ordering between wall and corner threads is needed.  pz: interior values are
// read as (v + 0.0f).
template <bool MR, int CW>
__device__ __forceinline__ void chain_boundary(float *x, float *xl, float *xr, const ChainPush &P, const int b, const int pz,
                                               const ChainGrid &g, const int c0)
{
    const int W = g.W, H = g.H;
    constexpr int cw = CW;
    const int tid = threadIdx.x, nt = blockDim.x;
    const float si = (b == 1) ? -1.0f : 1.0f, sj = (b == 2) ? -1.0f : 1.0f;
    auto in = [&](int i, int j) {
        const float v = __ldcg(x + CAT(g, i, j));
        return pz ? __fadd_rn(v, 0.0f) : v;
    };
    for (int t = tid; t < 2 * cw; t += nt) {
        const int j = c0 + (t % cw);
        if (j < 1 || j > H - 2) continue;
        const int i = (t < cw) ? 0 : W - 1, iin = (t < cw) ? 1 : W - 2;
        const float v = in(iin, j);
        store_pushed<MR>(x, xl, xr, P, CAT(g, i, j), (b == 1) ? -v : v);                  // :356-360
    }
    for (int side = 0; side < 2; ++side) {
        // a wall column is written by the block that owns the interior column it mirrors (the CTA that has just
        // written it), which need not be the block the wall column itself falls into
        const int j = side ? H - 1 : 0, jin = side ? H - 2 : 1;
        if (jin < c0 || jin >= c0 + cw) continue;
        for (int i = 1 + tid; i <= W - 2; i += nt) {
            const float v = in(i, jin);
            store_pushed<MR>(x, xl, xr, P, CAT(g, i, j), (b == 2) ? -v : v);              // :350-354
        }
        if (tid < 2) {
            // corner = 0.5f * (x[iin][j] + x[i][jin]): a j-wall cell and an i-wall cell (:363-366)
            const int i = tid ? W - 1 : 0, iin = tid ? W - 2 : 1;
            const float v = in(iin, jin);
            store_pushed<MR>(x, xl, xr, P, CAT(g, i, j), __fmul_rn(0.5f, __fadd_rn(__fmul_rn(sj, v), __fmul_rn(si, v))));
        }
    }
}

// ---- divergence (src/fluid_sim.cpp:307-314) on the columns [c0, c0 + cw) ---------------------------------------------
// A warp walks a run of rows with the rows above / below kept in registers; every lane owns one column (cw = 32) or two
// (c0 + lane, c0 + 32 + lane), the j-neighbours come from the neighbouring lanes (two edge loads per row for the columns
// outside the block).  Wall cells are read as setBoundary(1, u) / setBoundary(2, v) would have left them:
//   u[0][j] = -u[1][j], u[W-1][j] = -u[W-2][j], v[i][0] = -v[i][1], v[i][H-1] = -v[i][H-2]   (:350-360)
constexpr int kPwRun = 64;   // rows per warp run
constexpr int kPwBatch = 8;  // rows whose loads are in flight together

template <int CW>
__device__ __forceinline__ void chain_divergence(const ChainPhase &ph, const ChainGrid &g, const int blk)
{
    const int W = g.W, H = g.H;
    constexpr int cw = CW;
    const long long pitch = g.pitch;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int c0 = blk * cw;
    const float *u = ph.u, *v = ph.v;
    float *div = ph.q;
    const float mhh = ph.c0;
    constexpr bool two = cw == 64;
    const int ja = c0 + lane, jb = c0 + 32 + lane;
    const bool ina = ja >= 1 && ja <= H - 2, inb = two && jb >= 1 && jb <= H - 2;
    const bool edge_l = (lane == 0) && (c0 - 1 >= 1), edge_r = (lane == 31) && (c0 + cw <= H - 2);
    const int je = (lane == 0) ? c0 - 1 : c0 + cw;
    for (int r0 = 1 + kPwRun * w; r0 <= W - 2; r0 += kPwRun * nw) {
        const int r1 = min(r0 + kPwRun, W - 1);   // rows [r0, r1)
        const long long row0 = (long long)r0 * pitch - g.jbase;
        float uca = ina ? __ldcg(u + row0 + ja) : 0.0f, ucb = inb ? __ldcg(u + row0 + jb) : 0.0f;
        float uma, umb;
        if (r0 == 1) {
            uma = -uca;
            umb = -ucb;
        } else {
            uma = ina ? __ldcg(u + row0 - pitch + ja) : 0.0f;
            umb = inb ? __ldcg(u + row0 - pitch + jb) : 0.0f;
        }
        for (int i0 = r0; i0 < r1; i0 += kPwBatch) {
            float upa[kPwBatch], upb[kPwBatch], vca[kPwBatch], vcb[kPwBatch], ve[kPwBatch];
#pragma unroll
            for (int k = 0; k < kPwBatch; ++k) {
                const int i = i0 + k;
                const long long row = (long long)i * pitch - g.jbase;
                const bool rv = i < r1;
                const bool nx = rv && (i + 1 <= W - 2);
                upa[k] = (nx && ina) ? __ldcg(u + row + pitch + ja) : 0.0f;
                upb[k] = (nx && inb) ? __ldcg(u + row + pitch + jb) : 0.0f;
                vca[k] = (rv && ina) ? __ldcg(v + row + ja) : 0.0f;
                vcb[k] = (rv && inb) ? __ldcg(v + row + jb) : 0.0f;
                ve[k] = (rv && (edge_l || edge_r)) ? __ldcg(v + row + je) : 0.0f;
            }
#pragma unroll
            for (int k = 0; k < kPwBatch; ++k) {
                const int i = i0 + k;
                if (i >= r1) break;                       // warp-uniform
                const long long row = (long long)i * pitch - g.jbase;
                const float nua = (i + 1 == W - 1) ? -uca : upa[k];
                const float nub = (i + 1 == W - 1) ? -ucb : upb[k];
                // j-neighbours from the neighbouring lanes
                float la = __shfl_up_sync(0xffffffffu, vca[k], 1), ra = __shfl_down_sync(0xffffffffu, vca[k], 1);
                float lb = __shfl_up_sync(0xffffffffu, vcb[k], 1), rb = __shfl_down_sync(0xffffffffu, vcb[k], 1);
                const float b_first = __shfl_sync(0xffffffffu, vcb[k], 0), a_last = __shfl_sync(0xffffffffu, vca[k], 31);
                if (lane == 0) { la = ve[k]; lb = a_last; }
                if (lane == 31) { ra = two ? b_first : ve[k]; rb = ve[k]; }
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

// ---- pressure gradient (src/fluid_sim.cpp:334-341) on the columns [c0, c0 + cw), then setBoundary(1, u) and
// setBoundary(2, v) (:342-343) for those columns.  Pressure walls are read as setBoundary(0, p) defines them
// (p[0][j] = p[1][j], p[i][0] = p[i][1], ...), so the pressure array's own wall cells are never touched.
template <bool MR, int CW>
__device__ __forceinline__ void chain_gradient(const ChainPhase &ph, const ChainGrid &g, const int blk, float *vmax_out,
                                               float *umax_out)
{
    __shared__ float s_vmax[kChainThreads / 32], s_umax[kChainThreads / 32];
    float vmax = 0.0f, umax = 0.0f;
    const int W = g.W, H = g.H;
    constexpr int cw = CW;
    const long long pitch = g.pitch;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int c0 = blk * cw;
    const ChainPush P = make_push(g, ph, c0);
    float *u = ph.u, *v = ph.v;
    const float *p = ph.q;
    const float h = ph.c0;
    constexpr bool two = cw == 64;
    const int ja = c0 + lane, jb = c0 + 32 + lane;
    const bool ina = ja >= 1 && ja <= H - 2, inb = two && jb >= 1 && jb <= H - 2;
    const bool edge_l = (lane == 0) && (c0 - 1 >= 1), edge_r = (lane == 31) && (c0 + cw <= H - 2);
    const int je = (lane == 0) ? c0 - 1 : c0 + cw;
    for (int r0 = 1 + kPwRun * w; r0 <= W - 2; r0 += kPwRun * nw) {
        const int r1 = min(r0 + kPwRun, W - 1);
        const long long row0 = (long long)r0 * pitch - g.jbase;
        float pca = ina ? __ldcg(p + row0 + ja) : 0.0f, pcb = inb ? __ldcg(p + row0 + jb) : 0.0f;
        float pma, pmb;
        if (r0 == 1) {
            pma = pca;
            pmb = pcb;
        } else {
            pma = ina ? __ldcg(p + row0 - pitch + ja) : 0.0f;
            pmb = inb ? __ldcg(p + row0 - pitch + jb) : 0.0f;
        }
        float pe = (edge_l || edge_r) ? __ldcg(p + row0 + je) : 0.0f;   // this row's pressure just outside the block
        for (int i0 = r0; i0 < r1; i0 += kPwBatch) {
            float ppa[kPwBatch], ppb[kPwBatch], pen[kPwBatch], ua[kPwBatch], ub[kPwBatch], va[kPwBatch], vb[kPwBatch];
#pragma unroll
            for (int k = 0; k < kPwBatch; ++k) {
                const int i = i0 + k;
                const long long row = (long long)i * pitch - g.jbase;
                const bool rv = i < r1;
                const bool nx = rv && (i + 1 <= W - 2);
                ppa[k] = (nx && ina) ? __ldcg(p + row + pitch + ja) : 0.0f;
                ppb[k] = (nx && inb) ? __ldcg(p + row + pitch + jb) : 0.0f;
                pen[k] = (nx && (edge_l || edge_r)) ? __ldcg(p + row + pitch + je) : 0.0f;
                ua[k] = (rv && ina) ? __ldcg(u + row + ja) : 0.0f;
                ub[k] = (rv && inb) ? __ldcg(u + row + jb) : 0.0f;
                va[k] = (rv && ina) ? __ldcg(v + row + ja) : 0.0f;
                vb[k] = (rv && inb) ? __ldcg(v + row + jb) : 0.0f;
            }
#pragma unroll
            for (int k = 0; k < kPwBatch; ++k) {
                const int i = i0 + k;
                if (i >= r1) break;
                const long long row = (long long)i * pitch - g.jbase;
                const float npa = (i + 1 == W - 1) ? pca : ppa[k];
                const float npb = (i + 1 == W - 1) ? pcb : ppb[k];
                float la = __shfl_up_sync(0xffffffffu, pca, 1), ra = __shfl_down_sync(0xffffffffu, pca, 1);
                float lb = __shfl_up_sync(0xffffffffu, pcb, 1), rb = __shfl_down_sync(0xffffffffu, pcb, 1);
                const float b_first = __shfl_sync(0xffffffffu, pcb, 0), a_last = __shfl_sync(0xffffffffu, pca, 31);
                if (lane == 0) { la = pe; lb = a_last; }
                if (lane == 31) { ra = two ? b_first : pe; rb = pe; }
                if (ja - 1 == 0) la = pca;
                if (ja + 1 == H - 1) ra = pca;
                if (jb - 1 == 0) lb = pcb;
                if (jb + 1 == H - 1) rb = pcb;
                // u -= (0.5f*(p[i+1][j] - p[i-1][j])) / h ; v -= (0.5f*(p[i][j+1] - p[i][j-1])) / h
                if (ina) {
                    const float nv = __fsub_rn(va[k], __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(ra, la)), h));
                    const float nu = __fsub_rn(ua[k], __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(npa, pma)), h));
                    store_pushed<MR>(u, ph.peer_l[0], ph.peer_r[0], P, row + ja, nu);
                    store_pushed<MR>(v, ph.peer_l[1], ph.peer_r[1], P, row + ja, nv);
                    vmax = fmaxf(vmax, fabsf(nv));
                    umax = fmaxf(umax, fabsf(nu));
                }
                if (inb) {
                    const float nv = __fsub_rn(vb[k], __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(rb, lb)), h));
                    const float nu = __fsub_rn(ub[k], __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(npb, pmb)), h));
                    store_pushed<MR>(u, ph.peer_l[0], ph.peer_r[0], P, row + jb, nu);
                    store_pushed<MR>(v, ph.peer_l[1], ph.peer_r[1], P, row + jb, nv);
                    vmax = fmaxf(vmax, fabsf(nv));
                    umax = fmaxf(umax, fabsf(nu));
                }
                pma = pca; pmb = pcb;
                pca = npa; pcb = npb;
                pe = pen[k];
            }
        }
    }
    // largest |v| and |u| of the block (walls repeat interior magnitudes): they bound how many columns / rows an
    // advection trace that starts in or passes through this block can move
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
    chain_boundary<MR, CW>(u, ph.peer_l[0], ph.peer_r[0], P, 1, 0, g, c0);    // setBoundary(1, u), :342
    chain_boundary<MR, CW>(v, ph.peer_l[1], ph.peer_r[1], P, 2, 0, g, c0);    // setBoundary(2, v), :343
    // density "diffusion" with a == 0 leaves d + 0.0f in the interior and runs setBoundary(0, d): the
    // walls are set here, the + 0.0f is folded into the taps of the density advection
    if (ph.bnd_extra) chain_boundary<MR, CW>(ph.bnd_extra, ph.peer_l[2], ph.peer_r[2], P, 0, 1, g, c0);
}

// ---- RK4 advection (src/fluid_sim.cpp:261-298) of the columns [c0, c0 + cw), then setBoundary on the result -------
// The back-trace may leave the block: before it starts, the CTA waits until the gradient phase has published every
// block the traces can reach.  A trace moves at most dt0 * max|v| columns per stage (bilinear samples never exceed
// the largest tap), so with m = max |v| over the blocks [b - k, b + k] the reach is ceil(dt0 * m) + 2 columns; k grows
// until that fits inside those blocks (one block for the 4-cell displacement of the bench input, the whole grid
// in the worst case).  On a column block of a multi-rank run the traces must stay inside the halo the neighbours keep
// current: a trace that would leave it is reported (err word) instead of waited for.
template <bool MR>
__device__ __forceinline__ void chain_advect_wait(const ChainParams &C, const ChainPhase &ph, const int p, const int b,
                                                  const float dt0, int *reach_rows, int *reach_cols)
{
    __shared__ int s_k, s_ri, s_rj;
    const ChainGrid &g = C.g;
    const ChainPhase &gr = C.ph[ph.reach_phase];
    const unsigned *flags = C.done + (size_t)ph.reach_phase * kChainMaxItems;
    const float *vm = C.vmax + (size_t)ph.reach_phase * kChainMaxItems;
    // blocks whose gradient results this rank can see: the whole grid on one GPU, own blocks + halo on a column block
    const int vis_lo = g.has_left ? max((g.j0 - g.halo) / g.cw, 0) : 0;
    const int vis_hi = g.has_right ? min((g.j1 + g.halo) / g.cw - 1, gr.ncl - 1) : gr.ncl - 1;
    int k = 1, waited_lo = b + 1, waited_hi = b;   // blocks [waited_lo, waited_hi] are known to be done
    for (;;) {
        const int lo = max(b - k, vis_lo), hi = min(b + k, vis_hi);
        unsigned long long t0 = 0;
        for (int x = lo + (int)threadIdx.x; x <= hi; x += blockDim.x)
            if (x < waited_lo || x > waited_hi) chain_poll<MR>(C, flags + x, C.epoch, p, b, t0);
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
            int need = gr.ncl;                                  // non-finite velocities: wait for everything
            if (r < 1.0e8f) need = ((int)ceilf(r) + 2 + g.cw - 1) / g.cw;
            s_k = max(need, 1);
            s_rj = (r < 1.0e8f) ? (int)ceilf(r) + 2 : (1 << 30);
            s_ri = (ru < 1.0e8f) ? (int)ceilf(ru) + 2 : (1 << 30);
        }
        __syncthreads();
        const int need = s_k;
        *reach_rows = s_ri;
        *reach_cols = s_rj;
        __syncthreads();
        if (need <= k || (lo == vis_lo && hi == vis_hi)) {
            // the traces may want blocks this rank cannot see: the step's result is invalid, say so
            if (threadIdx.x == 0 && need > k && (b - need < vis_lo && vis_lo > 0 || b + need > vis_hi && vis_hi < gr.ncl - 1))
                atomicCAS(C.err, 0ull, 0x8000000000000000ull | (unsigned)b);
            break;
        }
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

// The four independent traces of a thread: two columns x two rows on 64-column items, one column x four rows on
// 32-column items.  `r` = row inside the stage / run, `c` = column offset inside the block.
__device__ __forceinline__ void trace_cell(const bool two, const int n, const int lane, int *r, int *c)
{
    *r = two ? (n >> 1) : n;
    *c = (two ? 32 * (n & 1) : 0) + lane;
}

// returns false (nothing done) when the tiles for this reach do not fit the shared memory of the launch
template <bool TWO, bool PZ, bool MR, int CW>
__device__ __forceinline__ bool chain_advect_tiled(const ChainPhase &ph, const ChainGrid &g, const int blk, float *smem,
                                                   const int smem_floats, const int ri, const int rj)
{
    constexpr int NARR = TWO ? 2 : 3;            // velocity: u, v are sources as well; density: u, v and the density
    const int W = g.W, H = g.H;
    constexpr int cw = CW;
    if ((g.pitch & 3) != 0 || (g.jbase & 3) != 0 || (H & 3) != 0 || ri > 64 || rj > 60) return false;
    const int halo_c = (rj + 1 + 3) & ~3;        // columns staged on each side (taps j0, j0 + 1), 16-byte aligned
    const int halo_r = ri + 1;
    const int TC = cw + 2 * halo_c, TR = kAdvRows + 2 * halo_r;
    const int stage_floats = NARR * TR * TC;
    if (2 * stage_floats > smem_floats) return false;
    const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, w = tid >> 5;
    const int c0 = blk * cw, cb = c0 - halo_c;
    const ChainPush P = make_push(g, ph, c0);
    const float *arr[3] = {ph.u, ph.v, ph.src_a};
    float *dst_a = ph.dst_a, *dst_b = ph.dst_b;
    const float ndt0 = ph.c0;
    const float xm = (float)W - 1.5f, ym = (float)H - 1.5f;
    constexpr bool two = cw == 64;
    constexpr int rows_per_pass = two ? 2 : 4, npass = two ? 2 : 1;     // 4 rows per warp and stage either way
    const int nq = TC >> 2;                      // 16-byte pieces per tile row
    const int my_q = tid % nq, my_r = tid / nq, r_step = nt / nq;
    const int nstages = (W - 2 + kAdvRows - 1) / kAdvRows;
    // columns the local arrays hold: [jbase, jbase + pitch) within the grid
    const int col_lo = max(g.jbase, 0), col_hi = min(g.jbase + g.pitch, H);
    auto issue = [&](int k) {
        if (k < nstages && my_r < r_step) {
            float *slot = smem + (k & 1) * stage_floats;
            const int row_lo = 1 + k * kAdvRows - halo_r;
            const int col = cb + 4 * my_q;
            if (col >= col_lo && col + 3 < col_hi) {
#pragma unroll
                for (int a = 0; a < NARR; ++a)
                    for (int r = my_r; r < TR; r += r_step) {
                        const int row = row_lo + r;
                        if (row < 0 || row > W - 1) continue;
                        cp_async_f32x4(slot + (a * TR + r) * TC + 4 * my_q, arr[a] + CAT(g, row, col), true);
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
        constexpr int N = 4;
#pragma unroll 1
        for (int it = 0; it < npass; ++it) {
            int ci[N], cj[N];
            bool ok[N];
            float x[N], y[N], k1x[N], k1y[N], k2x[N], k2y[N], k3x[N], k3y[N], k4x[N], k4y[N];
            Taps t[N];
#pragma unroll
            for (int n = 0; n < N; ++n) {
                int rr, cc;
                trace_cell(two, n, lane, &rr, &cc);
                const int r = 4 * w + rows_per_pass * it + rr;
                ci[n] = 1 + k * kAdvRows + r;
                cj[n] = c0 + cc;
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
                    const long long c = CAT(g, ci[n], cj[n]);
                    store_pushed<MR>(dst_a, ph.peer_l[0], ph.peer_r[0], P, c, ra);
                    if (TWO) store_pushed<MR>(dst_b, ph.peer_l[1], ph.peer_r[1], P, c, rb);
                    if (!TWO && !MR && ph.snap) ph.snap[c] = ra;
                }
            }
        }
        __syncthreads();                         // stage k is consumed: its buffer may be refilled
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    chain_boundary<MR, CW>(dst_a, ph.peer_l[0], ph.peer_r[0], P, ph.ba, 0, g, c0);
    if (TWO) chain_boundary<MR, CW>(dst_b, ph.peer_l[1], ph.peer_r[1], P, ph.bb, 0, g, c0);
    if (!TWO && !MR && ph.snap) chain_boundary<MR, CW>(ph.snap, nullptr, nullptr, P, ph.ba, 0, g, c0);
    return true;
}

template <bool TWO, bool PZ, bool MR, int CW>
__device__ __forceinline__ void chain_advect(const ChainPhase &ph, const ChainGrid &g, const int blk)
{
    const int W = g.W, H = g.H;
    constexpr int cw = CW;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int c0 = blk * cw;
    const ChainPush P = make_push(g, ph, c0);
    const GridView G{W, H, g.pitch, g.jbase, g.j0, g.j1};
    const float *u = ph.u, *v = ph.v;
    const float *src_a = ph.src_a, *src_b = ph.src_b;
    float *dst_a = ph.dst_a, *dst_b = ph.dst_b;
    const float ndt0 = ph.c0;
    const float xm = (float)W - 1.5f, ym = (float)H - 1.5f;
    constexpr bool two = cw == 64;
    constexpr int rows_per_pass = two ? 2 : 4;
    constexpr int N = 4;     // independent traces per thread
    for (int i0 = 1 + rows_per_pass * w; i0 <= W - 2; i0 += rows_per_pass * nw) {
        int ci[N], cj[N];
        bool ok[N];
        float x[N], y[N], k1x[N], k1y[N], k2x[N], k2y[N], k3x[N], k3y[N], k4x[N], k4y[N];
        Taps t[N];
#pragma unroll
        for (int n = 0; n < N; ++n) {
            int rr, cc;
            trace_cell(two, n, lane, &rr, &cc);
            ci[n] = i0 + rr;
            cj[n] = c0 + cc;
            ok[n] = ci[n] <= W - 2 && cj[n] >= 1 && cj[n] <= H - 2;
            // cells outside the interior trace from a harmless position this rank holds and store nothing
            x[n] = ok[n] ? (float)ci[n] : 1.0f;
            y[n] = ok[n] ? (float)cj[n] : (float)max(g.j0, 1);
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
                const long long c = CAT(g, ci[n], cj[n]);
                store_pushed<MR>(dst_a, ph.peer_l[0], ph.peer_r[0], P, c, ra);
                if (TWO) store_pushed<MR>(dst_b, ph.peer_l[1], ph.peer_r[1], P, c, rb);
                if (!TWO && !MR && ph.snap) ph.snap[c] = ra;
            }
        }
    }
    __syncthreads();
    chain_boundary<MR, CW>(dst_a, ph.peer_l[0], ph.peer_r[0], P, ph.ba, 0, g, c0);
    if (TWO) chain_boundary<MR, CW>(dst_b, ph.peer_l[1], ph.peer_r[1], P, ph.bb, 0, g, c0);
    if (!TWO && !MR && ph.snap) chain_boundary<MR, CW>(ph.snap, nullptr, nullptr, P, ph.ba, 0, g, c0);
}


// ---- velocity noise (src/fluid_sim.cpp:374-383) on the columns [c0, c0 + cw), all rows: u, v = src + noise ----------
// The reference draws two rand() numbers per cell, X then Y, rows outer / columns inner over the full grid, so cell
// (i, j) uses the outputs 2 (i H + j) and 2 (i H + j) + 1 of the step.  rand() is a linear recurrence over Z/2^32
// (glibc_rand.h): the window k outputs further on is new[m] = sum_j p[j] a[m + j], p = x^k mod (x^31 - x^28 - 1) and a[]
// the window extended by 30 sequential outputs -- a correlation every thread can do on its own registers (961 multiply-adds,
// 31 coefficients from shared memory).  Work split: a warp takes 32 consecutive rows of the block, lane = row.  Each lane
// reaches the window of (its row, c0) from the step's base window through the bits of the block index (x^(2 cw 2^b),
// skipped warp-wide where the bit is clear), of the warp index (x^(64 H 2^b)) and of the lane (x^(2 H 2^b)), generates the
// row's 2 cw outputs sequentially, and jumps on to the same columns 32 * (warps of the CTA) rows further down with one
// fixed polynomial.  The values are staged in shared memory so that the warp applies them one row at a time (coalesced).
constexpr int kNzPitch = 65;                 // 64 outputs per lane and round, odd pitch
static_assert(kNzWarps == kChainThreads / 32, "noise.cuh: warps of a CTA");

template <int CW>
__device__ __forceinline__ void chain_noise(const ChainParams &C, const ChainPhase &ph, const ChainGrid &g, const int blk, float *smem)
{
    const int W = g.W, H = g.H;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int c0 = blk * CW;
    const float *su = ph.src_a, *sv = ph.src_b;
    float *du = ph.u, *dv = ph.v;
    // Pending addDensity / addVelocity calls (src/fluid_sim.cpp:425-438) that fall into this block, in call order: the
    // density ones are applied to the density the step starts from right here (that block is final: the launch waits
    // for the previous step's density advection of it), the velocity ones to the value read below, before the noise.
    int *s_nvs = reinterpret_cast<int *>(smem) + kNzPolys * 32;
    int *s_list = s_nvs + 4;
    int nvs = 0;
    if (C.nsrc > 0) {
        if (tid == 0) {
            int m = 0;
            for (int k = 0; k < C.nsrc; ++k) {
                const ChainSrc e = C.src[k];
                if (e.col < c0 || e.col >= c0 + CW) continue;
                if (e.field == 0) {
                    float *q = C.density + CAT(g, e.row, e.col);
                    *q = __fadd_rn(__ldcg(q), e.amount);
                } else {
                    s_list[m++] = k;
                }
            }
            s_nvs[0] = m;
        }
        __syncthreads();
        nvs = s_nvs[0];
    }
    // is one of the block's velocity sources in the rows [lo, hi)?  (warp-uniform)
    auto rows_have_sources = [&](const int lo, const int hi) {
        bool any = false;
        for (int k0 = 0; k0 < nvs; k0 += 32) {
            const int k = k0 + lane;
            const int r = k < nvs ? C.src[s_list[k]].row : -1;
            any = any || __any_sync(0xffffffffu, r >= lo && r < hi);
        }
        return any;
    };
    auto add_sources = [&](const int i, const int j, float &uu, float &vv) {
        for (int k = 0; k < nvs; ++k) {
            const ChainSrc e = C.src[s_list[k]];
            if (e.row == i && e.col == j) {
                if (e.field == 1) uu = __fadd_rn(uu, e.amount);
                else vv = __fadd_rn(vv, e.amount);
            }
        }
    };
    if (C.nz_zero) {
        // rand() replaced by the value that makes the noise + 0.0f (FSB_NOISE_ZERO): the reference still adds it
        for (int i = wid; i < W; i += kNzWarps) {
            const bool srcs = nvs > 0 && rows_have_sources(i, i + 1);
#pragma unroll
            for (int k = 0; k < CW / 32; ++k) {
                const int j = c0 + 32 * k + lane;
                if (j < H) {
                    const long long c = CAT(g, i, j);
                    float uu = __ldcg(su + c), vv = __ldcg(sv + c);
                    if (srcs) add_sources(i, j, uu, vv);
                    du[c] = __fadd_rn(uu, 0.0f);
                    dv[c] = __fadd_rn(vv, 0.0f);
                }
            }
        }
        return;
    }
    uint32_t *s_poly = reinterpret_cast<uint32_t *>(smem);
    float *stg = smem + kNzPolys * 32 + 4 + kChainMaxSources + wid * (32 * kNzPitch);
    for (int i = tid; i < kNzPolys * 32; i += blockDim.x) s_poly[i] = C.nz_polys[i];
    __syncthreads();
    uint32_t w[kRandDeg];
#pragma unroll
    for (int m = 0; m < kRandDeg; ++m) w[m] = C.nz_base[m];
    int R0 = 32 * wid;                                   // first row of the warp's current group of 32
#pragma unroll 1
    for (int it = 0;; ++it) {
        const uint32_t *p;
        bool take = true;
        if (it < 10) {                                   // columns: 2 c0 = 2 cw blk outputs
            if (!((blk >> it) & 1)) continue;
            p = s_poly + (kNzBlockPoly + it) * 32;
        } else if (it < 14) {                            // rows: 32 wid ...
            if (!((wid >> (it - 10)) & 1)) continue;
            p = s_poly + (kNzWarpPoly + it - 10) * 32;
        } else if (it < 19) {                            // ... + lane
            take = (lane >> (it - 14)) & 1;
            p = s_poly + (kNzLanePoly + it - 14) * 32;
        } else {
            if (R0 >= W) break;
            const bool srcs = nvs > 0 && rows_have_sources(R0, R0 + 32);
#pragma unroll
            for (int round = 0; round < CW / 32; ++round) {
                // w[] is a circular buffer: output number n overwrites index n mod 31; 64 mod 31 = 2, so round r starts
                // at index 2 r (everything unrolled: the indices are constants, w[] stays in registers)
                const int rot = (64 * round) % kRandDeg;
#pragma unroll
                for (int n = 0; n < 64; ++n) {
                    const uint32_t x = w[(n + rot) % kRandDeg] + w[(n + rot + 28) % kRandDeg];
                    w[(n + rot) % kRandDeg] = x;
                    stg[lane * kNzPitch + n] = noise_of(x);
                }
                __syncwarp();
                const int j = c0 + 32 * round + lane;
#pragma unroll 1
                for (int r0 = 0; r0 < 32; r0 += 8) {
                    float uu[8], vv[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        const int i = R0 + r0 + r;
                        const bool ok = i < W && j < H;
                        const long long c = CAT(g, ok ? i : 0, ok ? j : g.jbase);
                        uu[r] = ok ? __ldcg(su + c) : 0.0f;
                        vv[r] = ok ? __ldcg(sv + c) : 0.0f;
                    }
                    if (srcs) {
#pragma unroll
                        for (int r = 0; r < 8; ++r) add_sources(R0 + r0 + r, j, uu[r], vv[r]);
                    }
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        const int i = R0 + r0 + r;
                        if (i < W && j < H) {
                            const long long c = CAT(g, i, j);
                            du[c] = __fadd_rn(uu[r], stg[(r0 + r) * kNzPitch + 2 * lane]);       // X component first
                            dv[c] = __fadd_rn(vv[r], stg[(r0 + r) * kNzPitch + 2 * lane + 1]);
                        }
                    }
                }
                __syncwarp();
            }
            R0 += 32 * kNzWarps;
            if (R0 >= W) break;
            // back to natural order: 2 cw outputs later the window starts at index (2 cw) mod 31
            constexpr int rot_end = (2 * CW) % kRandDeg;
            uint32_t t[kRandDeg];
#pragma unroll
            for (int m = 0; m < kRandDeg; ++m) t[m] = w[(m + rot_end) % kRandDeg];
#pragma unroll
            for (int m = 0; m < kRandDeg; ++m) w[m] = t[m];
            p = s_poly + kNzStridePoly * 32;
        }
        nz_jump(w, p, take);
    }
}

// What an item does besides its arithmetic, recomputed from (phase, local item) wherever it is needed so that nothing
// but those two integers stays live across the sweep pipeline's loops (the kernel runs at the 168-register cap of a
// 10-warp CTA, and anything more spills inside those loops).
struct ChainItem {
    int item, cl, fi, lo, hi;
    bool sweep, data_l, data_r;
};
__device__ __forceinline__ ChainItem chain_item(const ChainParams &C, const int p, const int li)
{
    const ChainPhase &ph = C.ph[p];
    const ChainGrid &g = C.g;
    ChainItem I;
    I.sweep = ph.kind == kPhSor || ph.kind == kPhGs || ph.kind == kPhGsX0;
    I.item = ph.item0 + li;                       // numbered along the whole grid
    I.cl = I.item / ph.nf;
    I.fi = I.item - I.cl * ph.nf;
    I.lo = I.cl * ph.cpv - ph.shift;              // result columns
    I.hi = I.lo + ph.cpv - 1;
    // does this item write into a neighbour's memory?  (sweep phases: the result columns owed to the left rank)
    I.data_l = I.sweep ? (g.has_left && I.lo < g.j0 + 1) : (ph.push_cols > 0 && near_left(g, I.lo, ph.push_cols));
    I.data_r = I.sweep ? false : (ph.push_cols > 0 && near_right(g, I.hi, ph.push_cols));
    return I;
}

template <int SD, int SP, bool MR, int CW>
__global__ void __launch_bounds__(kChainThreads, 1) chain_kernel(const __grid_constant__ ChainParams C)
{
    extern __shared__ float4 smem4[];
    float *smem = reinterpret_cast<float *>(smem4);
    const ChainGrid &g = C.g;
    if constexpr (!MR) {
        // Overlapping steps.  Every CTA of this launch is resident by the time all of them have executed the instruction
        // below; from then on the next launch of the stream (if it was launched for programmatic stream serialization)
        // may take the SMs this launch's CTAs leave one by one.  It orders itself against this launch through the
        // done-words (ChainPhase::xdep); what it overwrites that the launch before this one was still reading far from
        // the writing block (traced velocities, densities, block maxima) it protects by waiting, once, for that launch
        // to have finished altogether: the buffers alternate, so nothing closer in time is at risk.
        if (C.pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        if (C.xstep) {
            if (threadIdx.x == 0) {
                unsigned long long t0 = 0;
                chain_poll<MR>(C, C.completed + (C.epoch & 1u), C.epoch - 2u, kChainMaxPhases, 0, t0);
            }
            __syncthreads();
        }
    }
    for (int p = 0; p < C.nphases; ++p) {
        const ChainPhase &ph = C.ph[p];
        for (int li = blockIdx.x; li < ph.nitems; li += gridDim.x) {
            {
                const ChainItem I = chain_item(C, p, li);
                if (threadIdx.x == 0) C.trace[((size_t)p * kChainMaxItems + li) * 3] = global_timer_ns();
                chain_wait<MR>(C, p, I.cl, MR && I.data_l, MR && I.data_r);
            }
            if (ph.kind == kPhSor) {
                if (threadIdx.x == 0) C.trace[((size_t)p * kChainMaxItems + li) * 3 + 1] = global_timer_ns();
                if constexpr (SD > 0) wf_pass<SD, kSor, true>(C.wf[ph.wf], li / ph.nf, li % ph.nf, smem);
            } else if (ph.kind == kPhGs) {
                if (threadIdx.x == 0) C.trace[((size_t)p * kChainMaxItems + li) * 3 + 1] = global_timer_ns();
                wf_pass<SP, kGaussSeidel, false>(C.wf[ph.wf], li, 0, smem);
            } else if (ph.kind == kPhGsX0) {
                if (threadIdx.x == 0) C.trace[((size_t)p * kChainMaxItems + li) * 3 + 1] = global_timer_ns();
                wf_pass<SP, kGaussSeidel, true>(C.wf[ph.wf], li, 0, smem);     // a later pass of the same projection
            } else {
                const int cl = ph.item0 + li;     // pointwise phases: one field, item = block
                int reach_rows = 1 << 30, reach_cols = 1 << 30;
                if (ph.kind == kPhAdvect) chain_advect_wait<MR>(C, ph, p, cl, -ph.c0, &reach_rows, &reach_cols);
                if (threadIdx.x == 0) C.trace[((size_t)p * kChainMaxItems + li) * 3 + 1] = global_timer_ns();
                if (ph.kind == kPhBndP) {
                    ChainPush none{false, false, 0, 0};
                    chain_boundary<MR, CW>(ph.q, nullptr, nullptr, none, 0, 0, g, cl * CW);       // setBoundary(0, p) between two passes (:330)
                } else if (ph.kind == kPhNoise) {
                    if constexpr (!MR) chain_noise<CW>(C, ph, g, cl, smem);
                } else if (ph.kind == kPhDiv) {
                    chain_divergence<CW>(ph, g, cl);
                } else if (ph.kind == kPhGrad) {
                    chain_gradient<MR, CW>(ph, g, cl, C.vmax + (size_t)p * kChainMaxItems + cl,
                                   C.vmax + (size_t)(kChainMaxPhases + p) * kChainMaxItems + cl);
                } else if (ph.kind == kPhAdvect) {
                    if (ph.dst_b) {
                        if (!chain_advect_tiled<true, false, MR, CW>(ph, g, cl, smem, C.smem_floats, reach_rows, reach_cols))
                            chain_advect<true, false, MR, CW>(ph, g, cl);
                    } else if (ph.pz) {
                        if (!chain_advect_tiled<false, true, MR, CW>(ph, g, cl, smem, C.smem_floats, reach_rows, reach_cols))
                            chain_advect<false, true, MR, CW>(ph, g, cl);
                    } else {
                        if (!chain_advect_tiled<false, false, MR, CW>(ph, g, cl, smem, C.smem_floats, reach_rows, reach_cols))
                            chain_advect<false, false, MR, CW>(ph, g, cl);
                    }
                }
            }
            asm volatile("" : "+r"(li), "+r"(p));        // everything below is recomputed, not kept
            const ChainItem I = chain_item(C, p, li);
            if constexpr (MR) {
                if (I.data_l || I.data_r) __threadfence_system();   // this thread's stores into the neighbour's memory
            }
            __syncthreads();   // the item's results are written, its shared memory is free
            if (threadIdx.x == 0) {
                const size_t vi = (size_t)p * kChainMaxItems + I.item, ui = (size_t)(kChainMaxPhases + p) * kChainMaxItems + I.item;
                bool pushed = false;
                if constexpr (MR) {
                    // Items near a rank boundary publish in the neighbour's arrays as well (results, where the neighbour
                    // needs them, were stored there by this item): velocity maxima first, then the done-word.
                    const int reach = g.halo + 2 * ph.cpv;
                    const bool flag_l = near_left(g, I.lo, reach), flag_r = near_right(g, I.hi, reach);
                    if (flag_l || flag_r) {
                        pushed = true;
                        __threadfence_system();
                        if (ph.kind == kPhGrad) {
                            const float m = C.vmax[vi], mu = C.vmax[ui];
                            if (flag_l) { C.vmax_l[vi] = m; C.vmax_l[ui] = mu; }
                            if (flag_r) { C.vmax_r[vi] = m; C.vmax_r[ui] = mu; }
                            __threadfence_system();
                        }
                        if (flag_l) st_release_sys_u32(C.done_l + vi, C.epoch);
                        if (flag_r) st_release_sys_u32(C.done_r + vi, C.epoch);
                    }
                }
                if (!pushed) __threadfence();
                st_release_u32(C.done + vi, C.epoch);
                C.trace[((size_t)p * kChainMaxItems + li) * 3 + 2] = global_timer_ns();
            }
        }
    }
    // the last CTA to finish announces that this launch no longer reads anything: to the neighbouring ranks (which write
    // into this rank's arrays in their next step) and to the launch after next of this GPU (overlapping steps)
    __syncthreads();
    if (threadIdx.x == 0) {
        if constexpr (MR) __threadfence_system();
        else __threadfence();
        unsigned *count = C.cta_count + (C.epoch & 1u);
        const unsigned old = atomicAdd(count, 1u);
        if (old == gridDim.x - 1) {
            __threadfence();
            *count = 0;
            if constexpr (MR) {
                if (g.has_left) st_release_sys_u32(C.step_out_l, C.epoch);
                if (g.has_right) st_release_sys_u32(C.step_out_r, C.epoch);
            }
            st_release_u32(C.completed + (C.epoch & 1u), C.epoch);
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

const void *chain_kernel_ptr(bool with_sor, bool multi_rank, int cw)
{
#define FSB_CK(SD_, MR_, CW_) (const void *)chain_kernel<SD_, kChainSP, MR_, CW_>
    if (with_sor) {
        if (multi_rank) return cw == 64 ? FSB_CK(kChainSD, true, 64) : FSB_CK(kChainSD, true, 32);
        return cw == 64 ? FSB_CK(kChainSD, false, 64) : FSB_CK(kChainSD, false, 32);
    }
    if (multi_rank) return cw == 64 ? FSB_CK(0, true, 64) : FSB_CK(0, true, 32);
    return cw == 64 ? FSB_CK(0, false, 64) : FSB_CK(0, false, 32);
#undef FSB_CK
}

struct ChainPlan {
    bool ok = false;
    int W = 0, H = 0, D = 0, P = 0, ncols = 0;
    bool multi = false;
    int sor_nstr = 0, sor_ncta = 0, gs_nstr = 0, gs_ncta = 0, grid = 0;   // ncta: upper bounds for this rank
    int cw = kChainCols;                                          // columns of a pointwise item
    int gs_npass = 0, gs_stages[kChainMaxGsPasses] = {0, 0, 0};   // passes of one projection
    size_t smem = 0;
    size_t sor_ring_floats = 0, gs_ring_floats = 0;
};

}  // namespace

// What the most recent launch was: the next one may overlap with it only if it is the step after it
struct ChainLast {
    bool valid = false;
    unsigned epoch = 0;
    int kind = 0, nphases = 0, with_density = 0, density_pz = 0, noise_zero = 0;
    const float *vin_u = nullptr, *vin_v = nullptr;
    float *u = nullptr, *v = nullptr, *p = nullptr, *div = nullptr, *yu = nullptr, *yv = nullptr, *d = nullptr, *nd = nullptr;
};

struct ChainState {
    ChainPlan plan[2];                    // [0] with the diffusion phase, [1] without
    unsigned char *buf = nullptr;         // [done | misc | vmax x 2 | trace x 4 | link regions | flags | rings sor | rings gs]
    size_t bytes = 0;
    float *extra[2] = {nullptr, nullptr}; // pressure / divergence of the second projection of a whole-step launch
    size_t extra_floats = 0;
    unsigned epoch = 0;
    int last_nphases[2] = {0, 0};         // of the most recent launch of each kind (for chain_read_trace)
    int last_trace[2] = {0, 1};           // ... and the timeline buffer it wrote
    uint32_t *nz_polys = nullptr;         // jump polynomials of the noise phase for (nz_H, nz_cw)
    int nz_H = 0, nz_cw = 0;
    ChainLast last;
};

namespace {

// Can (D diffusion sweeps of u and v +) P pressure sweeps run as phases of one launch on this grid (or on a column
// block of `ncols` columns of it: every rank must arrive at the same strips per CTA, so only W, H, ncols enter)?
// dry_nsm > 0: no device is asked anything; the plan is made for a GPU of that many SMs (chain_describe)
ChainPlan make_plan(int W, int H, int D, int P, int ncols, bool multi, int dry_nsm = 0)
{
    ChainPlan pl;
    pl.W = W; pl.H = H; pl.D = D; pl.P = P; pl.ncols = ncols; pl.multi = multi;
    if (P < kChainSP || P % kChainSP != 0) return pl;
    if (D != 0 && D % kChainSD != 0) return pl;
    if (multi && D == 0) return pl;
    int dev = 0, nsm = dry_nsm;
    if (dry_nsm <= 0) {
        if (cudaGetDevice(&dev) != cudaSuccess) return pl;
        if (cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return pl;
    }
    size_t smem_gs = 0, smem_sor = 0;
    // strips a rank may have to run: its own columns, plus (last rank) the strips the result shift adds
    auto strips_of = [&](int shift) { return multi ? ncols / 32 + (shift + 31) / 32 + 1 : (H + shift + 31) / 32; };
    {
        // Passes of one projection: the fewest strips per CTA for which a pass fits the SMs (more SMs at work, fewer
        // warps sharing one), as many chained warps per strip as the 8 compute warps and the shared memory allow, the
        // P / 10 stages spread evenly over the passes.
        using L = WfLayout<kChainSP>;
        const int T = P / kChainSP;
        bool found = false;
        for (int nstr = 1; nstr <= 4 && !found; nstr *= 2) {
            if (multi && ncols % (32 * nstr) != 0) continue;
            int maxst = std::min((kChainThreads / 32 - 2) / nstr, L::kMaxStagesS);
            if (multi) maxst = std::min(maxst, kLinkStages);
            while (maxst >= 1 && wf_smem_bytes<kChainSP>(nstr, maxst) > 227 * 1024) --maxst;
            if (maxst < 1) continue;
            const int npass = (T + maxst - 1) / maxst;
            if (npass > kChainMaxGsPasses) continue;
            const int widest = (T + npass - 1) / npass;
            const int NS = strips_of(kChainSP * widest);
            const int ncta = (NS + std::min(nstr, NS) - 1) / std::min(nstr, NS);
            if (ncta > nsm || ncta > kWfMaxCtas) continue;
            // items are numbered along the whole grid (every rank's done[] array is indexed that way)
            const int NS_all = (H + kChainSP * widest + 31) / 32;
            if ((NS_all + std::min(nstr, NS) - 1) / std::min(nstr, NS) > kChainMaxItems) continue;
            pl.gs_nstr = std::min(nstr, NS);
            pl.gs_ncta = ncta;
            pl.gs_npass = npass;
            for (int k = 0; k < npass; ++k) pl.gs_stages[k] = T / npass + (k < T % npass ? 1 : 0);
            smem_gs = wf_smem_bytes<kChainSP>(pl.gs_nstr, widest);
            found = true;
        }
        if (!found) return pl;
    }
    if (D > 0) {
        // diffusion: as many strips per CTA as fit (a hand-over inside a CTA is far cheaper than one through the rings)
        using L = WfLayout<kChainSD>;
        const int nstg = D / kChainSD;
        if (nstg < 1 || nstg > L::kMaxStagesS) return pl;
        const int NS = strips_of(D);
        int nstr = 4;
        auto bad = [&](int n) {
            return n * nstg > kChainThreads / 32 - 2 || wf_smem_bytes<kChainSD>(n, nstg) > 227 * 1024 || n > NS ||
                   (multi && ncols % (32 * n) != 0);
        };
        while (nstr > 1 && bad(nstr)) nstr >>= 1;
        if (bad(nstr)) return pl;
        pl.sor_nstr = nstr;
        pl.sor_ncta = (NS + nstr - 1) / nstr;
        if (2 * pl.sor_ncta > nsm || pl.sor_ncta > kWfMaxCtas) return pl;
        if (2 * (((H + D + 31) / 32 + nstr - 1) / nstr) > kChainMaxItems) return pl;     // numbered along the whole grid
        smem_sor = wf_smem_bytes<kChainSD>(nstr, nstg);
    }
    // the advection phases stage their fields in the same shared memory: give them room on small grids too
    pl.smem = std::max(std::max(smem_gs, smem_sor), (size_t)160 * 1024);
    // pointwise items: 64 columns, or 32 when that is what it takes to give most SMs a block
    {
        static const int forced = [] { const char *e = std::getenv("FSB_CHAIN_CW"); return e ? std::atoi(e) : 0; }();
        pl.cw = (forced == 32 || forced == 64) ? forced : (((ncols + 63) / 64 < (3 * nsm) / 4) ? 32 : 64);
        if (multi && ncols % pl.cw != 0) return pl;
    }
    const int nblk = (ncols + pl.cw - 1) / pl.cw;
    pl.grid = std::max(std::max(pl.gs_ncta, 2 * pl.sor_ncta), std::min(nblk, nsm));
    if (pl.grid > nsm || (H + pl.cw - 1) / pl.cw > kChainMaxItems || pl.grid > kChainMaxItems) return pl;
    if (dry_nsm <= 0) {
        const void *kern = chain_kernel_ptr(D > 0, multi, pl.cw);
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
    }
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
constexpr size_t kMiscBytes = 256;        // +0 error word, +64 step words written by the neighbours, +128 finished-CTA counts, +192 completed launches
constexpr size_t kVmaxOne = (size_t)2 * kChainMaxPhases * kChainMaxItems * sizeof(float);     // max |v|, then max |u|
constexpr size_t kVmaxBytes = 2 * kVmaxOne;                                                  // overlapping launches alternate
constexpr size_t kTraceOne = (size_t)kChainMaxPhases * kChainMaxItems * 3 * sizeof(unsigned long long);
// everything up to here sits at the same offsets on every rank (the neighbours address it through peer mappings)
constexpr size_t kVmaxOff = kDoneBytes + kMiscBytes;
constexpr size_t kTraceOff = kVmaxOff + kVmaxBytes;
constexpr size_t kLinksOff = kTraceOff + 4 * kTraceOne;            // one link region per sweep phase
constexpr size_t kFixedBytes = kLinksOff + kChainMaxSweeps * kLinkBytes;

// hand-over of the strip chain to / from the neighbouring ranks for sweep phase `slot`
SweepLink chain_link(const ChainBlock &B, unsigned char *buf, int slot, unsigned long long seq)
{
    SweepLink L;
    const size_t off = kLinksOff + (size_t)slot * kLinkBytes;
    L.has_left = B.has_left;
    L.has_right = B.has_right;
    L.in_rings = reinterpret_cast<const float *>(buf + off);
    L.in_prod = reinterpret_cast<const unsigned long long *>(buf + off + kLinkRingBytes);
    L.cons = reinterpret_cast<const unsigned long long *>(buf + off + kLinkRingBytes + kLinkFlagBytes);
    if (B.has_left) {
        L.left_cons = reinterpret_cast<unsigned long long *>(B.left.chain_buf + off + kLinkRingBytes + kLinkFlagBytes);
        L.left_ret = reinterpret_cast<unsigned long long *>(B.left.chain_buf + off + kLinkRingBytes + 2 * kLinkFlagBytes);
        L.peer_jbase = B.G.jbase - B.ncols;        // same halo width and pitch on every rank
        L.peer_hi = B.G.j0 + 1;                    // the left rank's last owned columns and its first halo column
    }
    if (B.has_right) {
        L.right_rings = reinterpret_cast<float *>(B.right.chain_buf + off);
        L.right_prod = reinterpret_cast<unsigned long long *>(B.right.chain_buf + off + kLinkRingBytes);
    }
    L.seq = seq;
    return L;
}

}  // namespace

void chain_state_destroy(ChainState *cs)
{
    if (!cs) return;
    if (cs->buf) cudaFree(cs->buf);
    for (int k = 0; k < 2; ++k)
        if (cs->extra[k]) cudaFree(cs->extra[k]);
    if (cs->nz_polys) cudaFree(cs->nz_polys);
    delete cs;
}

void chain_shared_buffers(ChainState *cs, void *out[3])
{
    out[0] = cs ? cs->buf : nullptr;
    out[1] = cs ? cs->extra[0] : nullptr;
    out[2] = cs ? cs->extra[1] : nullptr;
}

cudaError_t chain_read_error(ChainState *cs, unsigned long long *word)
{
    *word = 0;
    if (!cs || !cs->buf) return cudaSuccess;
    return cudaMemcpy(word, cs->buf + kDoneBytes, sizeof(*word), cudaMemcpyDeviceToHost);
}

// Passes of FluidSim::step as phases of one launch (see the head of this file):
//   diffuse(u), diffuse(v)                               when A.sor_sweeps > 0          (src/fluid_sim.cpp:390-391)
//   project(u, v, p, div)                                always                         (:394)
//   advect(u), advect(v) into (p, div), project again    when A.with_advect             (:401-405)
//   densityStep with DIFFUSION == 0                      when A.with_density            (:409-422)
// Returns cudaErrorNotSupported when this grid / these sweep counts have no chained configuration (the caller then
// runs the passes as separate launches).
cudaError_t launch_chain(const ChainArgs &A, ChainState *&cs, SweepScratch &scr, cudaStream_t st, int mode)
{
    if (!cs) cs = new ChainState();
    const ChainBlock *B = A.blk;
    const bool multi = B != nullptr;
    const int kind = A.sor_sweeps > 0 ? 0 : 1;
    const int ncols = multi ? B->ncols : A.H;
    ChainPlan &pl = cs->plan[kind];
    const bool dry = mode == 3;                // describe only: no device call, no allocation, no launch
    if (dry && (!A.describe || A.describe_nsm <= 0)) return cudaErrorInvalidValue;
    if (dry || pl.W != A.W || pl.H != A.H || pl.D != A.sor_sweeps || pl.P != A.gs_sweeps || pl.ncols != ncols || pl.multi != multi)
    {
        pl = make_plan(A.W, A.H, A.sor_sweeps, A.gs_sweeps, ncols, multi, dry ? A.describe_nsm : 0);
        cs->last.valid = false;
    }
    if (!pl.ok) return cudaErrorNotSupported;
    if (A.with_density && !A.with_advect) return cudaErrorInvalidValue;
    if (multi && (!A.with_advect || B->halo < pl.cw || B->halo < A.gs_sweeps + 2)) return cudaErrorNotSupported;
    const bool with_noise = A.vin_u != nullptr;
    const bool ovl = A.overlap && !multi;      // launches of consecutive steps may overlap
    if (with_noise && (multi || !A.vin_v || !A.with_advect || (!A.noise_zero && !A.noise_base))) return cudaErrorInvalidValue;
    if ((A.yu == nullptr) != (A.yv == nullptr) || (A.yu && !A.with_advect)) return cudaErrorInvalidValue;
    if (A.nsources < 0 || A.nsources > kChainMaxSources || (A.nsources > 0 && (!with_noise || !A.sources))) return cudaErrorInvalidValue;
    if (A.snap && (multi || !A.with_density)) return cudaErrorInvalidValue;
    if (mode == 1) return cudaSuccess;

    const GridView G = multi ? B->G : whole_grid(A.W, A.H);
    cudaError_t e;
    if (A.with_advect && !dry) {
        const size_t nfl = (size_t)A.W * G.pitch;
        if (nfl > cs->extra_floats) {
            if (multi && cs->extra[0]) return cudaErrorInvalidValue;   // the neighbours have these buffers mapped
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
                e = cudaMemsetAsync(cs->extra[k], 0, nfl * sizeof(float), st);
                if (e != cudaSuccess) return e;
            }
            cs->extra_floats = nfl;
        }
    }
    // scratch: done words, watchdog / step words, per-block maxima, timelines, link regions, then progress words and
    // boundary rings of the sweep phases
    const size_t flag_bytes = sweep_flag_words() * sizeof(unsigned long long);
    const size_t sor_rings = std::max(cs->plan[0].sor_ring_floats, pl.sor_ring_floats) * sizeof(float);
    const size_t gs_rings = std::max(std::max(cs->plan[0].gs_ring_floats, cs->plan[1].gs_ring_floats), pl.gs_ring_floats) * sizeof(float);
    const size_t need = kFixedBytes + kChainMaxSweeps * flag_bytes + sor_rings + 2 * kChainMaxGsPasses * gs_rings;
    if (need > cs->bytes && !dry) {
        if (multi && cs->buf) return cudaErrorInvalidValue;            // mapped by the neighbours: must not move
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
    if (mode == 2) return cudaStreamSynchronize(st);
    if (!dry && with_noise && !A.noise_zero && (cs->nz_H != A.H || cs->nz_cw != pl.cw || !cs->nz_polys)) {
        // jump polynomials of the noise phase (chain_noise): x^k mod (x^31 - x^28 - 1), 32 words each
        std::vector<uint32_t> host((size_t)kNzPolys * 32, 0u);
        nz_fill_polys(A.H, pl.cw, host.data());
        e = cudaStreamSynchronize(st);         // a launch in flight may still read the old table
        if (e != cudaSuccess) return e;
        if (!cs->nz_polys) {
            e = cudaMalloc(&cs->nz_polys, host.size() * sizeof(uint32_t));
            if (e != cudaSuccess) return e;
        }
        e = cudaMemcpy(cs->nz_polys, host.data(), host.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) return e;
        cs->nz_H = A.H;
        cs->nz_cw = pl.cw;
    }
    unsigned char *b = dry ? reinterpret_cast<unsigned char *>(0x100000000ull) : cs->buf;   // (dry: addresses are never used)
    unsigned *done = reinterpret_cast<unsigned *>(b);
    unsigned long long *err = reinterpret_cast<unsigned long long *>(b + kDoneBytes);
    // (committed to cs->epoch only when the launch has been queued: the launch after next waits for THIS number to
    // appear in `completed`, so a number that never ran must not exist)
    const unsigned epoch = cs->epoch + 1;
    const int par = ovl ? (int)(epoch & 1u) : 0;          // overlapping launches alternate between two sets of these
    float *vmax = reinterpret_cast<float *>(b + kVmaxOff + par * kVmaxOne);
    const int trace_idx = kind + 2 * par;
    unsigned long long *trace = reinterpret_cast<unsigned long long *>(b + kTraceOff + trace_idx * kTraceOne);
    unsigned char *b2 = b + kFixedBytes;
    unsigned long long *flags[kChainMaxSweeps];
    for (int k = 0; k < kChainMaxSweeps; ++k) flags[k] = reinterpret_cast<unsigned long long *>(b2 + k * flag_bytes);
    float *rings_sor = reinterpret_cast<float *>(b2 + kChainMaxSweeps * flag_bytes);
    float *rings_gs[2 * kChainMaxGsPasses];
    for (int k = 0; k < 2 * kChainMaxGsPasses; ++k)
        rings_gs[k] = reinterpret_cast<float *>(b2 + kChainMaxSweeps * flag_bytes + sor_rings + k * gs_rings);

    const int cw = pl.cw;
    ChainParams C{};
    C.W = A.W;
    C.H = A.H;
    C.g = ChainGrid{A.W, A.H, G.pitch, G.jbase, G.j0, G.j1, cw, multi ? B->halo : 0, multi && B->has_left, multi && B->has_right,
                    multi ? G.jbase - B->ncols : 0, multi ? G.jbase + B->ncols : 0};
    C.smem_floats = (int)(pl.smem / sizeof(float));
    C.epoch = epoch;
    C.done = done;
    C.err = err;
    C.timeout_ns = 20ull * 1000 * 1000 * 1000;
    C.trace = trace;
    C.vmax = vmax;
    C.step_in = reinterpret_cast<unsigned *>(b + kDoneBytes + 64);
    C.cta_count = reinterpret_cast<unsigned *>(b + kDoneBytes + 128);
    C.completed = reinterpret_cast<unsigned *>(b + kDoneBytes + 192);
    C.pdl = ovl ? 1 : 0;
    C.nz_zero = A.noise_zero;
    C.nz_polys = cs->nz_polys;
    if (with_noise && !A.noise_zero) std::memcpy(C.nz_base, A.noise_base, sizeof(C.nz_base));
    C.nsrc = A.nsources;
    C.nsrc_density = 0;
    C.density = A.d;
    for (int k = 0; k < A.nsources; ++k) {
        const SourceEntry &q = A.sources[k];
        C.src[k] = ChainSrc{q.cell / G.pitch, q.cell % G.pitch + G.jbase, q.field, q.amount};
        if (q.field == 0) {
            if (!A.d) return cudaErrorInvalidValue;
            C.nsrc_density += 1;
        }
    }
    // Is this the step after the one the previous launch ran, in buffers that rotate as the overlap needs?  Then it may
    // start while that launch drains.  What this launch writes must not be anything the previous one reads far from
    // the writing block: its two velocity sets (traced by its advection phases), the density it advects from; its
    // results must be this step's inputs.  Otherwise: an ordinary launch, which starts when the stream is idle.
    {
        const ChainLast &L = cs->last;
        auto differs = [](const void *x, std::initializer_list<const void *> ys) {
            for (const void *y : ys)
                if (x == y) return false;
            return true;
        };
        C.xstep = ovl && with_noise && A.yu && A.with_density && L.valid && L.epoch + 1 == epoch && L.kind == kind &&
                  L.with_density == A.with_density && L.density_pz == A.density_pz && L.noise_zero == A.noise_zero &&
                  L.yu && A.vin_u == L.yu && A.vin_v == L.yv && A.d == L.nd && A.p == L.p && A.div == L.div &&
                  differs(A.u, {L.u, L.v, L.yu, L.yv}) && differs(A.v, {L.u, L.v, L.yu, L.yv}) &&
                  differs(A.yu, {L.u, L.v, L.yu, L.yv, L.vin_u, L.vin_v}) && differs(A.yv, {L.u, L.v, L.yu, L.yv, L.vin_u, L.vin_v}) &&
                  differs(A.nd, {L.d, L.nd});
    }
    const ChainPeer none;
    const ChainPeer &PL = (multi && B->has_left) ? B->left : none, &PR = (multi && B->has_right) ? B->right : none;
    if (PL.chain_buf) {
        C.done_l = reinterpret_cast<unsigned *>(PL.chain_buf);
        C.vmax_l = reinterpret_cast<float *>(PL.chain_buf + kVmaxOff);
        C.step_out_l = reinterpret_cast<unsigned *>(PL.chain_buf + kDoneBytes + 64) + 1;    // this rank is its right neighbour
    }
    if (PR.chain_buf) {
        C.done_r = reinterpret_cast<unsigned *>(PR.chain_buf);
        C.vmax_r = reinterpret_cast<float *>(PR.chain_buf + kVmaxOff);
        C.step_out_r = reinterpret_cast<unsigned *>(PR.chain_buf + kDoneBytes + 64) + 0;
    }
    // this rank's pointwise items: blocks of cw columns, numbered along the whole grid
    const int nblk_all = (A.H + cw - 1) / cw;
    const int blk0 = G.j0 / cw;
    const int nblk = (G.j1 >= A.H ? nblk_all : G.j1 / cw) - blk0;
    const float h = 1.0f / (float)A.W;            // src/fluid_sim.cpp:304
    int np = 0, nwf = 0;

    auto pointwise_phase = [&](int kind_) -> ChainPhase & {
        ChainPhase &ph = C.ph[np++];
        ph.kind = kind_;
        ph.item0 = blk0;
        ph.nitems = nblk;
        ph.ncl = nblk_all;
        ph.nf = 1;
        ph.cpv = cw;
        ph.shift = 0;
        ph.ndeps = 0;
        ph.nxdeps = 0;
        return ph;
    };
    // a sweep phase of nstg stages with nstr strips per CTA over nf fields: this rank's CTAs of the global chain
    auto sweep_phase = [&](int kind_, int nstr, int nstg, int S, int nf, float a, float c_recip, float omega, float *rings,
                           int slot) -> ChainPhase & {
        ChainPhase &ph = C.ph[np++];
        const int shift = S * nstg;
        const int NS_all = (A.H + shift + 31) / 32;       // the last stage's frame is shifted left by the pass's sweeps
        const int s_begin = multi ? B->strip_begin : 0;
        const int s_end = (!multi || B->strip_end < 0 || B->strip_end > NS_all) ? NS_all : B->strip_end;
        const int ncta = (s_end - s_begin + nstr - 1) / nstr;
        ph.kind = kind_;
        ph.item0 = (s_begin / nstr) * nf;
        ph.nitems = ncta * nf;
        ph.ncl = (NS_all + nstr - 1) / nstr;
        ph.nf = nf;
        ph.cpv = 32 * nstr;
        ph.shift = shift;
        ph.wf = nwf;
        ph.ndeps = 0;
        ph.nxdeps = 0;
        // overlapping steps: the rings between this item and the next are reused by the next launch; the item on the
        // right must have taken everything the step before sent it (items cl and cl + 1 of the same phase)
        ph.xdep[ph.nxdeps++] = ChainDep{np - 1, 1, 0, 32 * nstr};
        WfKernelParams &P = C.wf[nwf++];
        fill_wf(P, A.W, A.H, nstr, nstg, ncta, a, c_recip, omega, rings, flags[slot], ++scr.seq);
        P.pitch = G.pitch;
        P.jbase = G.jbase;
        P.c_base = s_begin / nstr;
        if (multi) {
            P.link = chain_link(*B, b, slot, B->step * 8 + (unsigned long long)slot + 1);
            P.link_seq_base = P.link.seq << 32;
        }
        return ph;
    };
    // divergence -> pressure sweeps -> gradient (+ boundaries) of (u, v) with scratch (p, div); `after` = phase that
    // produced u, v (or -1); returns the index of the gradient phase.  lu, lv, lp / ru, rv: the neighbours' copies.
    auto projection = [&](float *u, float *v, float *p, float *div, int after, int gs_slot, float *bnd_extra, float *lu, float *lv,
                          float *lp, float *lx, float *ru, float *rv, float *rx) -> int {
        const int ph_div = np;
        {
            ChainPhase &ph = pointwise_phase(kPhDiv);
            ph.u = u;
            ph.v = v;
            ph.q = div;
            ph.c0 = -0.5f * h;                    // evaluated first (:311)
            if (after >= 0) ph.dep[ph.ndeps++] = ChainDep{after, 0, 1, 1};    // u, v of the block and one column either side
        }
        int ph_gs = -1, ph_prev = ph_div, ph_gs_first = -1;
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
            if (k == 0) ph_gs_first = np;
            const int slot = 1 + gs_slot * kChainMaxGsPasses + k;
            ChainPhase &ph = sweep_phase(k == 0 ? kPhGs : kPhGsX0, pl.gs_nstr, pl.gs_stages[k], kChainSP, 1, 0.0f, 1.0f, 1.0f,
                                         rings_gs[slot - 1], slot);
            ph.dep[ph.ndeps++] = ChainDep{ph_prev, 0, 0, 0};                  // divergence (first pass) / walls of p (later passes) of the columns it loads
            if (k > 0) ph.dep[ph.ndeps++] = ChainDep{ph_last, 1, 0, 0};       // the previous pass has produced the columns this one overwrites
            WfKernelParams &P = C.wf[ph.wf];
            P.fld[0] = SweepField{k == 0 ? nullptr : p, div, p, 1.0f, 1.0f};  // p starts at 0 (:313, :316)
            P.fld[0].out_peer = lp;                                           // result columns owed to the left rank
            P.fld[1] = P.fld[2] = P.fld[0];
        }
        ChainPhase &ph = pointwise_phase(kPhGrad);
        ph.u = u;
        ph.v = v;
        ph.q = p;
        ph.c0 = h;
        ph.bnd_extra = bnd_extra;
        ph.peer_l[0] = lu; ph.peer_l[1] = lv; ph.peer_l[2] = lx;
        ph.peer_r[0] = ru; ph.peer_r[1] = rv; ph.peer_r[2] = rx;
        ph.push_cols = multi ? B->halo : 0;               // the neighbours' traces reach this far into these columns
        ph.dep[ph.ndeps++] = ChainDep{ph_gs, 0, 1, 1};    // pressure of the block and one column either side
        ph.dep[ph.ndeps++] = ChainDep{ph_div, 1, 1, 1};   // the divergence blocks that read the u, v about to change
        // Overlapping steps: pressure and divergence are scratch of this projection alone, but the step before used the
        // same arrays.  Its sweep items must have loaded the divergence this step overwrites (all passes: the last
        // one's items, a block wider for the passes' different shifts), its gradient blocks must have read the
        // pressure this step's sweeps (and wall updates) overwrite.
        const int ph_grad = np - 1;
        {
            const int sh = C.ph[ph_gs].shift;
            ChainPhase &dv = C.ph[ph_div];
            dv.xdep[dv.nxdeps++] = ChainDep{ph_gs, 1, sh + kChainCols, -sh + kChainCols};
            for (int q = ph_gs_first; q < ph_grad; ++q) {
                ChainPhase &g = C.ph[q];
                g.xdep[g.nxdeps++] = ChainDep{ph_grad, 1, 1, 1};
            }
        }
        return ph_grad;
    };

    int ph_noise = -1;
    if (with_noise) {
        // u, v = (velocity the step starts from) + noise, every cell (src/fluid_sim.cpp:374-383)
        ph_noise = np;
        ChainPhase &ph = pointwise_phase(kPhNoise);
        ph.u = A.u;
        ph.v = A.v;
        ph.src_a = A.vin_u;
        ph.src_b = A.vin_v;
    }
    int ph_sor = -1;
    if (A.sor_sweeps > 0) {
        ph_sor = np;
        ChainPhase &ph = sweep_phase(kPhSor, pl.sor_nstr, A.sor_sweeps / kChainSD, kChainSD, 2, A.a, A.c_recip, A.omega, rings_sor, 0);
        if (ph_noise >= 0) ph.dep[ph.ndeps++] = ChainDep{ph_noise, 0, 0, 0};        // the columns the item loads
        WfKernelParams &P = C.wf[ph.wf];
        P.fld[0] = SweepField{A.u, A.u, A.u, -1.0f, 1.0f};    // setBoundary(1, u): negated on the i-walls
        P.fld[1] = SweepField{A.v, A.v, A.v, 1.0f, -1.0f};    // setBoundary(2, v): negated on the j-walls
        P.fld[0].out_peer = PL.u;
        P.fld[1].out_peer = PL.v;
        P.fld[2] = P.fld[0];
    }
    const bool fold = A.with_density && A.density_pz;
    const int ph_grad1 = projection(A.u, A.v, A.p, A.div, ph_sor >= 0 ? ph_sor : ph_noise, 0, nullptr, PL.u, PL.v, PL.p, nullptr,
                                    PR.u, PR.v, nullptr);
    float *yu = A.yu ? A.yu : A.p, *yv = A.yu ? A.yv : A.div;     // the advected velocity
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
            ph.dst_a = yu;
            ph.dst_b = yv;
            ph.ba = 1;
            ph.bb = 2;
            ph.pz = 0;
            ph.c0 = -(A.dt * (float)A.W);              // -dt0, dt0 = dt * width (:264)
            ph.reach_phase = ph_grad1;
            ph.peer_l[0] = PL.p; ph.peer_l[1] = PL.div;
            ph.peer_r[0] = PR.p; ph.peer_r[1] = PR.div;
            ph.push_cols = multi ? cw : 0;             // the neighbour's divergence reads one column across
            if (!A.yu) {
                ph.dep[ph.ndeps++] = ChainDep{ph_grad1, 1, 1, 1};                            // gradient blocks still reading p here
                const int sh = C.ph[ph_gs1].shift;                                         // of the projection's last pass
                ph.dep[ph.ndeps++] = ChainDep{ph_gs1, 1, sh, -sh};                         // sweep items that load div here
            }
        }
        // the second projection works in scratch of its own: u, v stay intact for traces that start far away
        const int ph_grad2 = projection(yu, yv, cs->extra[0], cs->extra[1], ph_adv, 1, fold ? A.d : nullptr, PL.p, PL.div,
                                        PL.extra[0], fold ? PL.d : nullptr, PR.p, PR.div, fold ? PR.d : nullptr);
        if (A.with_density) {
            const int ph_advd = np;
            ChainPhase &ph = pointwise_phase(kPhAdvect);
            ph.u = yu;
            ph.v = yv;
            ph.src_a = A.d;
            ph.dst_a = A.nd;
            ph.ba = 0;
            ph.pz = A.density_pz;
            ph.c0 = -(A.dt * (float)A.W);
            ph.reach_phase = ph_grad2;
            ph.snap = A.snap;
            ph.peer_l[0] = PL.nd;
            ph.peer_r[0] = PR.nd;
            ph.push_cols = multi ? B->halo : 0;        // the next step's density traces reach into these columns
            // Overlapping steps: the density this step starts from is the result of the step before.  Its gradient
            // phase (which sets that density's walls, and which the density advection waits for over its whole reach)
            // takes the block from the previous launch's density advection ...
            ChainPhase &g2 = C.ph[ph_grad2];
            g2.xdep[g2.nxdeps++] = ChainDep{ph_advd, 0, 1, 1};
        }
        if (ph_noise >= 0) {
            // ... and the noise phase takes the velocity from its second gradient phase; it overwrites what the step
            // before read as ITS starting velocity, block by block
            ChainPhase &nz = C.ph[ph_noise];
            nz.xdep[nz.nxdeps++] = ChainDep{ph_grad2, 0, 1, 1};
            nz.xdep[nz.nxdeps++] = ChainDep{ph_noise, 0, 0, 0};
            // density sources are added to the previous step's result where it lies: that block must be final
            if (C.nsrc_density > 0 && A.with_density) nz.xdep[nz.nxdeps++] = ChainDep{np - 1, 0, 0, 0};
        }
    }
    C.nphases = np;
    if (dry) {
        ChainDesc &D = *A.describe;
        D = ChainDesc{};
        D.nphases = np;
        D.grid = pl.grid;
        D.cw = pl.cw;
        for (int q = 0; q < np; ++q) {
            const ChainPhase &ph = C.ph[q];
            ChainPhaseDesc &d = D.ph[q];
            d.kind = ph.kind; d.nitems = ph.nitems; d.ncl = ph.ncl; d.nf = ph.nf; d.cpv = ph.cpv; d.shift = ph.shift;
            d.ndeps = ph.ndeps; d.nxdeps = ph.nxdeps;
            for (int k = 0; k < ph.ndeps; ++k) { d.dep[k][0] = ph.dep[k].phase; d.dep[k][1] = ph.dep[k].from_out; d.dep[k][2] = ph.dep[k].lo; d.dep[k][3] = ph.dep[k].hi; }
            for (int k = 0; k < ph.nxdeps; ++k) { d.xdep[k][0] = ph.xdep[k].phase; d.xdep[k][1] = ph.xdep[k].from_out; d.xdep[k][2] = ph.xdep[k].lo; d.xdep[k][3] = ph.xdep[k].hi; }
            const void *ptrs[8] = {ph.u, ph.v, ph.q, ph.dst_a, ph.dst_b, ph.src_a, ph.src_b, ph.bnd_extra};
            for (int k = 0; k < 8; ++k) d.ptr[k] = ptrs[k];
            d.reach_phase = ph.reach_phase;
            d.item0 = ph.item0;
            d.push_cols = ph.push_cols;
            if (ph.kind == kPhSor || ph.kind == kPhGs || ph.kind == kPhGsX0)
                for (int f = 0; f < 2; ++f) {
                    const SweepField &F = C.wf[ph.wf].fld[f];
                    d.wf[f][0] = F.x0; d.wf[f][1] = F.rhs; d.wf[f][2] = F.out;
                }
        }
        return cudaSuccess;
    }
    C.xstep = C.xstep && cs->last.nphases == np;
    cs->last_nphases[kind] = np;
    cs->last_trace[kind] = trace_idx;
    {
        ChainLast &L = cs->last;
        L.valid = true;
        L.epoch = epoch;
        L.kind = kind;
        L.nphases = np;
        L.with_density = A.with_density;
        L.density_pz = A.density_pz;
        L.noise_zero = A.noise_zero;
        L.vin_u = A.vin_u; L.vin_v = A.vin_v;
        L.u = A.u; L.v = A.v; L.p = A.p; L.div = A.div;
        L.yu = A.yu; L.yv = A.yv;
        L.d = A.d; L.nd = A.nd;
    }

    void *args[] = {&C};
    const void *kern = chain_kernel_ptr(A.sor_sweeps > 0, multi, pl.cw);
    // Cooperative (all CTAs guaranteed co-resident) unless steps may overlap: a cooperative launch never starts before
    // the launch before it has drained, a launch with programmatic stream serialization does, and is given its SMs
    // only once every CTA of that launch is resident (they trigger at their first instruction), so the CTAs of both
    // are co-resident by construction (the grid fits the GPU: make_plan).  It still starts after everything else the
    // stream holds between the two.
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3(pl.grid);
    lc.blockDim = dim3(kChainThreads);
    lc.dynamicSmemBytes = pl.smem;
    lc.stream = st;
    cudaLaunchAttribute at[1];
    lc.attrs = at;
    lc.numAttrs = 0;
    if (!ovl) {
        at[0].id = cudaLaunchAttributeCooperative;
        at[0].val.cooperative = 1;
        lc.numAttrs = 1;
    } else if (C.xstep) {
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        lc.numAttrs = 1;
    }
    e = cudaLaunchKernelExC(&lc, kern, args);
    if (e == cudaSuccess) cs->epoch = epoch;
    else cs->last.valid = false;
    return e;
}

// The phase table launch_chain would build for these arguments on a GPU of `nsm` SMs, without touching a device: the
// items, their dependencies inside the launch and on the previous launch, and the arrays they read and write (the field
// pointers of `A` may be any distinct addresses).  For the CPU check of the overlap protocol (tests/test_chain_deps.py).
cudaError_t chain_describe(const ChainArgs &A0, int nsm, float *extra0, float *extra1, ChainDesc *out)
{
    ChainArgs A = A0;
    A.describe = out;
    A.describe_nsm = nsm;
    ChainState *cs = new ChainState();
    cs->extra[0] = extra0;      // stand-ins for the second projection's scratch
    cs->extra[1] = extra1;
    SweepScratch scr{};
    const cudaError_t e = launch_chain(A, cs, scr, nullptr, 3);
    cs->extra[0] = cs->extra[1] = nullptr;
    chain_state_destroy(cs);
    return e;
}

// Start / end times of every work item of the most recent chained launch of one kind (which = 0: launches that begin
// with the diffusion phase, 1: the others): out[(phase * max_items + item) * 3 + k] in nanoseconds of the GPU's global
// timer (k = 0 wait begins, 1 work begins, 2 results published); 0 = item did not run.  The caller synchronises first.
cudaError_t chain_read_trace(ChainState *cs, int which, unsigned long long *out, int max_phases, int max_items, int *nphases)
{
    *nphases = 0;
    if (!cs || !cs->buf || which < 0 || which > 1) return cudaSuccess;
    const unsigned long long *trace =
        reinterpret_cast<const unsigned long long *>(cs->buf + kTraceOff + cs->last_trace[which] * kTraceOne);
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
