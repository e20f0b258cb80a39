// Pointwise, stencil and gather passes of FluidSim::step (sm_100a).
// One thread per cell, j (contiguous) across the lanes of a warp.  Every expression keeps the
// reference's association order and uses round-to-nearest intrinsics, so nothing is
// contracted into an FMA (the reference's x86-64 -O2 build has none).
#include "fsb_internal.h"
#include "gather.cuh"

namespace fsb {

namespace {

constexpr int kBx = 32, kBy = 8;

__device__ __forceinline__ float plusz(float v, int plus_zero) { return plus_zero ? __fadd_rn(v, 0.0f) : v; }

// ---- setBoundary (src/fluid_sim.cpp:347-367) ------------------------------------------------
// One thread per perimeter cell; corner threads rebuild the two wall values they average from
// the interior, so there is no ordering between wall and corner threads.
struct BndJobs { BndJob j[4]; int n; };


__device__ __forceinline__ float wall_j(const float *x, const GridView &G, int i, int jin, int b, int pz)
{   // value setBoundary gives to (i, 0) or (i, H-1); jin = 1 or H-2
    const float v = plusz(x[AT(G, i, jin)], pz);
    return b == 2 ? -v : v;
}
__device__ __forceinline__ float wall_i(const float *x, const GridView &G, int iin, int j, int b, int pz)
{   // value setBoundary gives to (0, j) or (W-1, j); iin = 1 or W-2
    const float v = plusz(x[AT(G, iin, j)], pz);
    return b == 1 ? -v : v;
}

// Only perimeter cells in owned columns [j0, j1) are written; perimeter cells in halo columns arrive with the
// halo exchange from the rank that owns them.
__global__ void set_boundary_kernel(const BndJobs jobs, const GridView G)
{
    const BndJob J = jobs.j[blockIdx.y];
    float *x = J.x;
    const int b = J.b, pz = J.plus_zero;
    const int W = G.W, H = G.H;
    const int nown = G.j1 - G.j0;
    const int per = 2 * nown + 2 * W;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < per; t += gridDim.x * blockDim.x) {
        if (t < 2 * nown) {
            // rows i = 0 and i = W-1, owned columns
            const int top = t < nown;
            const int j = G.j0 + (top ? t : t - nown);
            const int i = top ? 0 : W - 1;
            const int iin = top ? 1 : W - 2;
            float v;
            if (j >= 1 && j <= H - 2) {
                v = wall_i(x, G, iin, j, b, pz);
            } else {
                // corner: 0.5f * (x[iin][jc] + x[i][jin]) with jc = j, jin = 1 or H-2 (:363-366)
                const int jin = (j == 0) ? 1 : H - 2;
                const float a = wall_j(x, G, iin, jin, b, pz);   // x[iin][j]  (a j-wall cell)
                const float c = wall_i(x, G, iin, jin, b, pz);   // x[i][jin]  (an i-wall cell)
                v = __fmul_rn(0.5f, __fadd_rn(a, c));
            }
            x[AT(G, i, j)] = v;
        } else {
            const int s = t - 2 * nown;
            const int left = s < W;
            const int i = left ? s : s - W;
            const int j = left ? 0 : H - 1;
            const int jin = left ? 1 : H - 2;
            if (j >= G.j0 && j < G.j1 && i >= 1 && i <= W - 2) x[AT(G, i, j)] = wall_j(x, G, i, jin, b, pz);
        }
    }
}

// ---- divergence (src/fluid_sim.cpp:307-314) -------------------------------------------------------
__global__ void divergence_kernel(const float *__restrict__ u, const float *__restrict__ v, float *__restrict__ div,
                                  const GridView G, float mhh)
{
    const int W = G.W, H = G.H;
    const int j = G.j0 + blockIdx.x * kBx + threadIdx.x;
    const int i = blockIdx.y * kBy + threadIdx.y;
    if (i < 1 || i > W - 2 || j < 1 || j > H - 2 || j >= G.j1) return;
    const size_t c = AT(G, i, j);
    const float du = __fsub_rn(u[c + G.pitch], u[c - G.pitch]);
    const float s = __fsub_rn(__fadd_rn(du, v[c + 1]), v[c - 1]);
    div[c] = __fmul_rn(mhh, s);
}

// ---- pressure gradient (src/fluid_sim.cpp:334-341) ---------------------------------------------------
// vmax_bits (may be null): receives the largest |v| written, as the bit pattern of a non-negative float (atomicMax on
// unsigned keeps the float order).  It bounds how many columns the next advection's back-traces can move.
__device__ __forceinline__ void publish_vmax(float m, unsigned *vmax_bits)
{
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (((threadIdx.y * blockDim.x + threadIdx.x) & 31) == 0 && m > 0.0f) atomicMax(vmax_bits, __float_as_uint(m));
}

__global__ void gradient_kernel(float *__restrict__ u, float *__restrict__ v, const float *__restrict__ p,
                                const GridView G, float h, unsigned *vmax_bits)
{
    const int W = G.W, H = G.H;
    const int j = G.j0 + blockIdx.x * kBx + threadIdx.x;
    const int i = blockIdx.y * kBy + threadIdx.y;
    float m = 0.0f;
    if (!(i < 1 || i > W - 2 || j < 1 || j > H - 2 || j >= G.j1)) {
        const size_t c = AT(G, i, j);
        const float gx = __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(p[c + G.pitch], p[c - G.pitch])), h);
        const float gy = __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(p[c + 1], p[c - 1])), h);
        u[c] = __fsub_rn(u[c], gx);
        const float nv = __fsub_rn(v[c], gy);
        v[c] = nv;
        m = fabsf(nv);
    }
    if (vmax_bits) publish_vmax(m, vmax_bits);
}

// ---- the same two stencils, four columns and kRows rows per thread -------------------------------------------------
// Used when every row starts 16-byte aligned (pitch, jbase and j0 multiples of 4).  A thread walks down kRows rows
// with the rows above / below kept in registers, so each value of the i-stencil is loaded once per thread; the two
// j-neighbours outside the quad are scalar loads that hit the lines the neighbouring lanes load anyway.  Quads that
// contain a wall column (or the end of the owned block) store per component.
constexpr int kRows = 4;

__device__ __forceinline__ float4 ld4(const float *p) { return *reinterpret_cast<const float4 *>(p); }

__device__ __forceinline__ void store_quad(float *dst, float4 r, int j, int jlo, int jhi)
{   // dst = &x[i][j]; only columns in [jlo, jhi] are written
    if (j >= jlo && j + 3 <= jhi) {
        *reinterpret_cast<float4 *>(dst) = r;
    } else {
        if (j >= jlo && j <= jhi) dst[0] = r.x;
        if (j + 1 >= jlo && j + 1 <= jhi) dst[1] = r.y;
        if (j + 2 >= jlo && j + 2 <= jhi) dst[2] = r.z;
        if (j + 3 >= jlo && j + 3 <= jhi) dst[3] = r.w;
    }
}

__global__ void __launch_bounds__(256) divergence_vec_kernel(const float *__restrict__ u, const float *__restrict__ v,
                                                             float *__restrict__ div, const GridView G, float mhh)
{
    const int W = G.W, H = G.H;
    const int j = G.j0 + 4 * (blockIdx.x * kBx + threadIdx.x);
    const int i0 = 1 + kRows * (blockIdx.y * kBy + threadIdx.y);
    if (j >= G.j1 || i0 > W - 2) return;
    const int jlo = max(1, G.j0), jhi = min(H - 2, G.j1 - 1);
    const int jmax = G.jbase + G.pitch - 1;                 // last column the local array holds
    size_t c = AT(G, i0, j);
    float4 um = ld4(u + c - G.pitch), uc = ld4(u + c);
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
        const int i = i0 + r;
        if (i > W - 2) break;
        const float4 up = ld4(u + c + G.pitch);
        const float4 vc = ld4(v + c);
        const float vl = (j - 1 >= G.jbase) ? v[c - 1] : 0.0f;
        const float vr = (j + 4 <= jmax) ? v[c + 4] : 0.0f;
        float4 d;
        d.x = __fmul_rn(mhh, __fsub_rn(__fadd_rn(__fsub_rn(up.x, um.x), vc.y), vl));
        d.y = __fmul_rn(mhh, __fsub_rn(__fadd_rn(__fsub_rn(up.y, um.y), vc.z), vc.x));
        d.z = __fmul_rn(mhh, __fsub_rn(__fadd_rn(__fsub_rn(up.z, um.z), vc.w), vc.y));
        d.w = __fmul_rn(mhh, __fsub_rn(__fadd_rn(__fsub_rn(up.w, um.w), vr), vc.z));
        store_quad(div + c, d, j, jlo, jhi);
        um = uc;
        uc = up;
        c += G.pitch;
    }
}

__global__ void __launch_bounds__(256) gradient_vec_kernel(float *__restrict__ u, float *__restrict__ v,
                                                           const float *__restrict__ p, const GridView G, float h,
                                                           unsigned *vmax_bits)
{
    const int W = G.W, H = G.H;
    const int j = G.j0 + 4 * (blockIdx.x * kBx + threadIdx.x);
    const int i0 = 1 + kRows * (blockIdx.y * kBy + threadIdx.y);
    float m = 0.0f;
    if (j >= G.j1 || i0 > W - 2) {
        if (vmax_bits) publish_vmax(m, vmax_bits);
        return;
    }
    const int jlo = max(1, G.j0), jhi = min(H - 2, G.j1 - 1);
    const int jmax = G.jbase + G.pitch - 1;
    size_t c = AT(G, i0, j);
    float4 pm = ld4(p + c - G.pitch), pc = ld4(p + c);
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
        const int i = i0 + r;
        if (i > W - 2) break;
        const float4 pp = ld4(p + c + G.pitch);
        const float pl = (j - 1 >= G.jbase) ? p[c - 1] : 0.0f;
        const float pr = (j + 4 <= jmax) ? p[c + 4] : 0.0f;
        float4 uu = ld4(u + c), vv = ld4(v + c);
        uu.x = __fsub_rn(uu.x, __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(pp.x, pm.x)), h));
        uu.y = __fsub_rn(uu.y, __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(pp.y, pm.y)), h));
        uu.z = __fsub_rn(uu.z, __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(pp.z, pm.z)), h));
        uu.w = __fsub_rn(uu.w, __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(pp.w, pm.w)), h));
        vv.x = __fsub_rn(vv.x, __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(pc.y, pl)), h));
        vv.y = __fsub_rn(vv.y, __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(pc.z, pc.x)), h));
        vv.z = __fsub_rn(vv.z, __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(pc.w, pc.y)), h));
        vv.w = __fsub_rn(vv.w, __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(pr, pc.z)), h));
        store_quad(u + c, uu, j, jlo, jhi);
        store_quad(v + c, vv, j, jlo, jhi);
        if (j >= jlo && j <= jhi) m = fmaxf(m, fabsf(vv.x));
        if (j + 1 >= jlo && j + 1 <= jhi) m = fmaxf(m, fabsf(vv.y));
        if (j + 2 >= jlo && j + 2 <= jhi) m = fmaxf(m, fabsf(vv.z));
        if (j + 3 >= jlo && j + 3 <= jhi) m = fmaxf(m, fabsf(vv.w));
        pm = pc;
        pc = pp;
        c += G.pitch;
    }
    if (vmax_bits) publish_vmax(m, vmax_bits);
}

inline bool quad_aligned(const GridView &G) { return (G.pitch % 4 == 0) && (G.jbase % 4 == 0) && (G.j0 % 4 == 0); }
inline dim3 quad_grid(const GridView &G)
{
    const int quads = (G.j1 - G.j0 + 3) / 4, rows = (G.W - 2 + kRows - 1) / kRows;
    return dim3((quads + kBx - 1) / kBx, (rows + kBy - 1) / kBy);
}

// ---- RK4 back-trace (src/fluid_sim.cpp:261-298) ------------------------------------------------------------
// dst_a = src_a sampled at the traced position; dst_b likewise (the two velocity components share
// one trace because both advect() calls of velocityStep use prevVelocityX/Y, :401-402).
template <bool TWO, bool PZ, typename Idx>
__global__ void advect_rk4_kernel(float *__restrict__ dst_a, const float *__restrict__ src_a,
                                  float *__restrict__ dst_b, const float *__restrict__ src_b,
                                  const float *__restrict__ u, const float *__restrict__ v, const GridView G, float ndt0)
{
    const int W = G.W, H = G.H;
    const int j = G.j0 + blockIdx.x * kBx + threadIdx.x;
    const int i = blockIdx.y * kBy + threadIdx.y;
    if (i < 1 || i > W - 2 || j < 1 || j > H - 2 || j >= G.j1) return;
    const float xm = (float)W - 1.5f, ym = (float)H - 1.5f;
    const float x = (float)i, y = (float)j;

    Taps t = make_taps(x, y, xm, ym);
    const float k1x = __fmul_rn(sample<false, Idx>(u, t, G), ndt0);
    const float k1y = __fmul_rn(sample<false, Idx>(v, t, G), ndt0);

    t = make_taps(__fadd_rn(x, __fmul_rn(k1x, 0.5f)), __fadd_rn(y, __fmul_rn(k1y, 0.5f)), xm, ym);
    const float k2x = __fmul_rn(sample<false, Idx>(u, t, G), ndt0);
    const float k2y = __fmul_rn(sample<false, Idx>(v, t, G), ndt0);

    t = make_taps(__fadd_rn(x, __fmul_rn(k2x, 0.5f)), __fadd_rn(y, __fmul_rn(k2y, 0.5f)), xm, ym);
    const float k3x = __fmul_rn(sample<false, Idx>(u, t, G), ndt0);
    const float k3y = __fmul_rn(sample<false, Idx>(v, t, G), ndt0);

    t = make_taps(__fadd_rn(x, k3x), __fadd_rn(y, k3y), xm, ym);
    const float k4x = __fmul_rn(sample<false, Idx>(u, t, G), ndt0);
    const float k4y = __fmul_rn(sample<false, Idx>(v, t, G), ndt0);

    const float dx = __fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(k1x, __fmul_rn(2.0f, k2x)), __fmul_rn(2.0f, k3x)), k4x), 6.0f);
    const float dy = __fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(k1y, __fmul_rn(2.0f, k2y)), __fmul_rn(2.0f, k3y)), k4y), 6.0f);

    t = make_taps(__fadd_rn(x, dx), __fadd_rn(y, dy), xm, ym);
    const size_t c = AT(G, i, j);
    dst_a[c] = sample<PZ, Idx>(src_a, t, G);
    if (TWO) dst_b[c] = sample<PZ, Idx>(src_b, t, G);
}

// ---- single semi-Lagrangian trace (src/fluid_sim.cpp:65-100 and the two halves of :103-168) ---------------------
template <bool PZ, typename Idx>
__global__ void advect_sl_kernel(float *__restrict__ dst, const float *__restrict__ src, const float *__restrict__ u,
                                 const float *__restrict__ v, const GridView G, float dt0, int forward)
{
    const int W = G.W, H = G.H;
    const int j = G.j0 + blockIdx.x * kBx + threadIdx.x;
    const int i = blockIdx.y * kBy + threadIdx.y;
    if (i < 1 || i > W - 2 || j < 1 || j > H - 2 || j >= G.j1) return;
    const size_t c = AT(G, i, j);
    const float du = __fmul_rn(dt0, u[c]), dv = __fmul_rn(dt0, v[c]);
    const float x = forward ? __fadd_rn((float)i, du) : __fsub_rn((float)i, du);
    const float y = forward ? __fadd_rn((float)j, dv) : __fsub_rn((float)j, dv);
    const Taps t = make_taps(x, y, (float)W - 1.5f, (float)H - 1.5f);
    dst[c] = sample<PZ, Idx>(src, t, G);
}

// ---- MacCormack correction and clamp (src/fluid_sim.cpp:170-207) --------------------------------------------------
template <bool PZ>
__global__ void maccormack_combine_kernel(float *__restrict__ dst, const float *__restrict__ src,
                                          const float *__restrict__ fwd, const float *__restrict__ back, const GridView G)
{
    const int W = G.W, H = G.H;
    const int j = G.j0 + blockIdx.x * kBx + threadIdx.x;
    const int i = blockIdx.y * kBy + threadIdx.y;
    if (i < 1 || i > W - 2 || j < 1 || j > H - 2 || j >= G.j1) return;
    const size_t c = AT(G, i, j);
    const float s = tap_val<PZ>(src, i, j, G);
    const float err = __fsub_rn(s, back[c]);
    float r = __fadd_rn(fwd[c], __fmul_rn(0.5f, err));
    float lo = s, hi = s;
#pragma unroll
    for (int di = -1; di <= 1; ++di)
#pragma unroll
        for (int dj = -1; dj <= 1; ++dj) {
            const float q = tap_val<PZ>(src, i + di, j + dj, G);
            lo = (q < lo) ? q : lo;
            hi = (hi < q) ? q : hi;
        }
    r = (r < hi) ? r : hi;
    r = (lo < r) ? r : lo;
    dst[c] = r;
}

__global__ void add_zero_kernel(float *x, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        x[i] = __fadd_rn(x[i], 0.0f);
}

// Holds the stream until a counter in this GPU's memory, incremented by kernels of the neighbouring GPU through
// peer memory, reaches `expected`.
__global__ void wait_counter_kernel(const unsigned long long *counter, unsigned long long expected)
{
    const volatile unsigned long long *c = counter;
    while (*c < expected) __nanosleep(200);
    __threadfence_system();
}

__global__ void axpy_field_kernel(float *x, const float *y, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        x[i] = __fadd_rn(x[i], y[i]);
}

// ---- addDensity / addVelocity (src/fluid_sim.cpp:424-439), applied in call order per cell -----------------------------
__global__ void apply_sources_kernel(float *d, float *u, float *v, const SourceEntry *e, const int *seg_start, int nseg)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nseg) return;
    for (int k = seg_start[s]; k < seg_start[s + 1]; ++k) {
        float *f = e[k].field == 0 ? d : (e[k].field == 1 ? u : v);
        f[e[k].cell] = __fadd_rn(f[e[k].cell], e[k].amount);
    }
}

inline dim3 cell_grid(const GridView &G) { return dim3((G.j1 - G.j0 + kBx - 1) / kBx, (G.W + kBy - 1) / kBy); }

}  // namespace

cudaError_t launch_set_boundary(const BndJob *jobs, int njobs, const GridView &G, cudaStream_t st)
{
    if (njobs < 1 || njobs > 4) return cudaErrorInvalidValue;
    BndJobs J;
    J.n = njobs;
    for (int i = 0; i < 4; ++i) J.j[i] = jobs[i < njobs ? i : 0];
    const int per = 2 * G.W + 2 * (G.j1 - G.j0);
    int blocks = (per + 255) / 256;
    if (blocks > 1024) blocks = 1024;
    set_boundary_kernel<<<dim3(blocks, njobs), 256, 0, st>>>(J, G);
    return cudaGetLastError();
}

cudaError_t launch_divergence(const float *u, const float *v, float *div, const GridView &G, cudaStream_t st)
{
    const float h = 1.0f / (float)G.W;      // src/fluid_sim.cpp:304
    const float mhh = -0.5f * h;            // -0.5f * h, evaluated first (:311)
    if (quad_aligned(G)) divergence_vec_kernel<<<quad_grid(G), dim3(kBx, kBy), 0, st>>>(u, v, div, G, mhh);
    else divergence_kernel<<<cell_grid(G), dim3(kBx, kBy), 0, st>>>(u, v, div, G, mhh);
    return cudaGetLastError();
}

cudaError_t launch_gradient(float *u, float *v, const float *p, const GridView &G, cudaStream_t st, unsigned *vmax_bits)
{
    const float h = 1.0f / (float)G.W;
    if (quad_aligned(G)) gradient_vec_kernel<<<quad_grid(G), dim3(kBx, kBy), 0, st>>>(u, v, p, G, h, vmax_bits);
    else gradient_kernel<<<cell_grid(G), dim3(kBx, kBy), 0, st>>>(u, v, p, G, h, vmax_bits);
    return cudaGetLastError();
}

cudaError_t launch_advect_rk4(float *dst_a, const float *src_a, float *dst_b, const float *src_b, const float *u,
                              const float *v, int src_plus_zero, const GridView &G, float dt, cudaStream_t st)
{
    const float dt0 = dt * (float)G.W;      // src/fluid_sim.cpp:264
    const float ndt0 = -dt0;
    const dim3 g = cell_grid(G), b(kBx, kBy);
    const bool small = (size_t)G.W * (size_t)G.pitch < ((size_t)1 << 31);
#define FSB_RK4(TWO_, PZ_)                                                                                         \
    do {                                                                                                           \
        if (small) advect_rk4_kernel<TWO_, PZ_, int><<<g, b, 0, st>>>(dst_a, src_a, dst_b, src_b, u, v, G, ndt0);  \
        else advect_rk4_kernel<TWO_, PZ_, long long><<<g, b, 0, st>>>(dst_a, src_a, dst_b, src_b, u, v, G, ndt0);  \
    } while (0)
    if (dst_b) {
        if (src_plus_zero) FSB_RK4(true, true);
        else FSB_RK4(true, false);
    } else {
        if (src_plus_zero) FSB_RK4(false, true);
        else FSB_RK4(false, false);
    }
#undef FSB_RK4
    return cudaGetLastError();
}

cudaError_t launch_advect_sl(float *dst, const float *src, const float *u, const float *v, int src_plus_zero,
                             const GridView &G, float dt, float sign, cudaStream_t st)
{
    const float dt0 = dt * (float)G.W;
    const dim3 g = cell_grid(G), b(kBx, kBy);
    const int fwd = sign > 0.0f;
    const bool small = (size_t)G.W * (size_t)G.pitch < ((size_t)1 << 31);
    if (src_plus_zero) {
        if (small) advect_sl_kernel<true, int><<<g, b, 0, st>>>(dst, src, u, v, G, dt0, fwd);
        else advect_sl_kernel<true, long long><<<g, b, 0, st>>>(dst, src, u, v, G, dt0, fwd);
    } else {
        if (small) advect_sl_kernel<false, int><<<g, b, 0, st>>>(dst, src, u, v, G, dt0, fwd);
        else advect_sl_kernel<false, long long><<<g, b, 0, st>>>(dst, src, u, v, G, dt0, fwd);
    }
    return cudaGetLastError();
}

cudaError_t launch_maccormack_combine(float *dst, const float *src, const float *fwd, const float *back,
                                      int src_plus_zero, const GridView &G, cudaStream_t st)
{
    const dim3 g = cell_grid(G), b(kBx, kBy);
    if (src_plus_zero) maccormack_combine_kernel<true><<<g, b, 0, st>>>(dst, src, fwd, back, G);
    else maccormack_combine_kernel<false><<<g, b, 0, st>>>(dst, src, fwd, back, G);
    return cudaGetLastError();
}

cudaError_t launch_add_zero(float *x, size_t n, cudaStream_t st)
{
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    add_zero_kernel<<<(unsigned)blocks, 256, 0, st>>>(x, n);
    return cudaGetLastError();
}

cudaError_t launch_axpy_field(float *x, const float *y, size_t n, cudaStream_t st)
{
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    axpy_field_kernel<<<(unsigned)blocks, 256, 0, st>>>(x, y, n);
    return cudaGetLastError();
}

cudaError_t launch_wait_counter(const unsigned long long *counter, unsigned long long expected, cudaStream_t st)
{
    wait_counter_kernel<<<1, 1, 0, st>>>(counter, expected);
    return cudaGetLastError();
}

cudaError_t launch_apply_sources(float *d, float *u, float *v, const SourceEntry *entries, const int *seg_start,
                                 int nseg, cudaStream_t st)
{
    if (nseg <= 0) return cudaSuccess;
    apply_sources_kernel<<<(nseg + 127) / 128, 128, 0, st>>>(d, u, v, entries, seg_start, nseg);
    return cudaGetLastError();
}

}  // namespace fsb
