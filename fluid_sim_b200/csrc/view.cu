// Headless views: the reference viewer's per-cell density quads with its density shader,
// and its velocity glyph lines, evaluated per pixel by CUDA kernels into an RGBA8 device buffer.  No graphics API is
// involved.
//
// What the reference draws (src/main.cpp):
//   * clear colour (0, 0, 0, 1)                                                        :345
//   * blending SRC_ALPHA / ONE_MINUS_SRC_ALPHA                                          :376-377
//   * one quad per cell (i, j) with density >= 0.01, corners
//       x = (float)i / width * 2 - 1,  y = (float)j / height * 2 - 1,  + 2/width, + 2/height        :388-427
//   * fragment colour mix((0,0,1), (1,0,0), d), alpha min(2 d, 0.8)                                  :32-41
// Quads of neighbouring cells tile the viewport, so a pixel centre is covered by one quad; where float rounding lets
// two neighbouring quads overlap by an ulp and a pixel centre falls exactly there, both are blended in the reference's
// draw order over the 8-bit framebuffer value.  Rows of the output are window rows: row 0 is the bottom of the window
// (y = -1), as in GL.
#include "fsb_internal.h"

namespace fsb {

namespace {

// NDC coordinate of the left / lower edge of cell k out of n (the reference's float expression, :397-398)
__device__ __forceinline__ float cell_edge(int k, float n) { return __fsub_rn(__fmul_rn(__fdiv_rn((float)k, n), 2.0f), 1.0f); }

// Cells whose quads cover NDC coordinate c along one axis: `k` (the last cell whose lower edge is <= c, or -1 if c lies
// in a rounding gap above its upper edge) and, where float rounding lets two neighbouring quads overlap by an ulp,
// also k - 1 (`prev`).  Quads are half-open: [edge, edge + size).
struct Cover { int k; bool cur, prev; };

__device__ __forceinline__ Cover covering_cells(float c, int n)
{
    const float fn = (float)n;
    const float size = __fdiv_rn(2.0f, fn);                     // cellWidth / cellHeight (:399-400)
    int k = __float2int_rd(__fmul_rn(__fmul_rn(__fadd_rn(c, 1.0f), 0.5f), fn));
    k = max(0, min(n - 1, k));
    while (k > 0 && c < cell_edge(k, fn)) --k;
    while (k < n - 1 && c >= cell_edge(k + 1, fn)) ++k;
    Cover r;
    r.k = k;
    const float lo = cell_edge(k, fn);
    r.cur = (c >= lo) && (c < __fadd_rn(lo, size));
    r.prev = (k > 0) && (c < __fadd_rn(cell_edge(k - 1, fn), size));
    return r;
}

__device__ __forceinline__ float clamp01(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }
__device__ __forceinline__ int to_unorm8(float v) { return __float2int_rd(__fadd_rn(__fmul_rn(clamp01(v), 255.0f), 0.5f)); }

// One quad's fragment blended over the 8-bit framebuffer value `px` (SRC_ALPHA, ONE_MINUS_SRC_ALPHA).
__device__ __forceinline__ void blend_cell(float den, int (&px)[4])
{
    if (den < 0.01f) return;                                    // :404 skips these cells
    // mix(x, y, t) = x * (1 - t) + y * t per component; fragment outputs clamp to [0, 1] in a fixed-point framebuffer
    const float sr = clamp01(__fadd_rn(__fmul_rn(0.0f, __fsub_rn(1.0f, den)), __fmul_rn(1.0f, den)));
    const float sb = clamp01(__fadd_rn(__fmul_rn(1.0f, __fsub_rn(1.0f, den)), __fmul_rn(0.0f, den)));
    const float sa = clamp01(fminf(__fmul_rn(den, 2.0f), 0.8f));
    const float ia = __fsub_rn(1.0f, sa);
    const float src[4] = {sr, 0.0f, sb, sa};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const float dst = __fdiv_rn((float)px[c], 255.0f);
        px[c] = to_unorm8(__fadd_rn(__fmul_rn(src[c], sa), __fmul_rn(dst, ia)));
    }
}

__global__ void view_density_kernel(const float *__restrict__ d, const GridView G, int pw, int ph, uchar4 *__restrict__ out)
{
    const int px = blockIdx.x * blockDim.x + threadIdx.x;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= pw || py >= ph) return;
    // pixel centre in NDC (viewport = the whole window, :341)
    const float xc = __fsub_rn(__fmul_rn(__fdiv_rn(__fadd_rn((float)px, 0.5f), (float)pw), 2.0f), 1.0f);
    const float yc = __fsub_rn(__fmul_rn(__fdiv_rn(__fadd_rn((float)py, 0.5f), (float)ph), 2.0f), 1.0f);
    const Cover ci = covering_cells(xc, G.W), cj = covering_cells(yc, G.H);
    int v[4] = {0, 0, 0, 255};                                  // the cleared framebuffer (:345)
    // draw order of the reference: i outer, j inner (:392-395)
#pragma unroll
    for (int a = 1; a >= 0; --a) {
        if (!(a ? ci.prev : ci.cur)) continue;
        const int i = ci.k - a;
#pragma unroll
        for (int b = 1; b >= 0; --b) {
            if (!(b ? cj.prev : cj.cur)) continue;
            const int j = cj.k - b;
            blend_cell(d[(size_t)i * G.pitch + (j - G.jbase)], v);
        }
    }
    out[(size_t)py * pw + px] = make_uchar4((unsigned char)v[0], (unsigned char)v[1], (unsigned char)v[2], (unsigned char)v[3]);
}

// ---- velocity glyphs (src/main.cpp:452-514, colour ramp src/fluid_sim.cpp:482-519) ---------------
// Every second cell in i and j with |v| >= 0.1 draws a line from the cell centre, 0.05 NDC long, in the direction of the
// velocity, ramp colour, alpha 0.8, glLineWidth(2).  GL leaves the exact pixels of a wide non-antialiased line to the
// implementation; here a pixel belongs to a line when its centre lies in the half-open span [start, end) along the
// major axis and within [-1, 1) pixels of the line along the minor axis (two pixels per column / row).  Lines are
// blended in the reference's draw order (i outer, j inner) over the density image.
struct Glyph { float ax, ay, bx, by, r, g, b; bool on; };

__device__ __forceinline__ Glyph make_glyph(float vx, float vy, int i, int j, float fw, float fh, float pw, float ph)
{
    Glyph q;
    const float mag = sqrtf(__fadd_rn(__fmul_rn(vx, vx), __fmul_rn(vy, vy)));      // glm::length
    q.on = !(mag < 0.1f);
    float nx = 0.0f, ny = 0.0f;
    if (mag > 0.001f) {
        nx = __fdiv_rn(vx, mag);
        ny = __fdiv_rn(vy, mag);
    }
    const float ns = fminf(__fdiv_rn(mag, 10.0f), 1.0f);
    if (ns < 0.25f) {
        q.r = 0.0f; q.g = __fdiv_rn(ns, 0.25f); q.b = 1.0f;
    } else if (ns < 0.5f) {
        q.r = 0.0f; q.g = 1.0f; q.b = __fsub_rn(1.0f, __fdiv_rn(__fsub_rn(ns, 0.25f), 0.25f));
    } else if (ns < 0.75f) {
        q.r = __fdiv_rn(__fsub_rn(ns, 0.5f), 0.25f); q.g = 1.0f; q.b = 0.0f;
    } else {
        q.r = 1.0f; q.g = __fsub_rn(1.0f, __fdiv_rn(__fsub_rn(ns, 0.75f), 0.25f)); q.b = 0.0f;
    }
    const float cx = __fadd_rn(__fsub_rn(__fmul_rn(__fdiv_rn((float)i, fw), 2.0f), 1.0f), __fdiv_rn(1.0f, fw));
    const float cy = __fadd_rn(__fsub_rn(__fmul_rn(__fdiv_rn((float)j, fh), 2.0f), 1.0f), __fdiv_rn(1.0f, fh));
    const float ex = __fadd_rn(cx, __fmul_rn(nx, 0.05f));
    const float ey = __fadd_rn(cy, __fmul_rn(ny, 0.05f));
    // viewport transform to window coordinates
    q.ax = __fmul_rn(__fmul_rn(__fadd_rn(cx, 1.0f), 0.5f), pw);
    q.ay = __fmul_rn(__fmul_rn(__fadd_rn(cy, 1.0f), 0.5f), ph);
    q.bx = __fmul_rn(__fmul_rn(__fadd_rn(ex, 1.0f), 0.5f), pw);
    q.by = __fmul_rn(__fmul_rn(__fadd_rn(ey, 1.0f), 0.5f), ph);
    return q;
}

__device__ __forceinline__ bool glyph_covers(const Glyph &q, float pcx, float pcy)
{
    const float dx = __fsub_rn(q.bx, q.ax), dy = __fsub_rn(q.by, q.ay);
    const bool xmajor = fabsf(dx) >= fabsf(dy);
    const float a0 = xmajor ? q.ax : q.ay, a1 = xmajor ? q.bx : q.by;      // major axis
    const float m0 = xmajor ? q.ay : q.ax;                                  // minor axis at the start
    const float dmaj = xmajor ? dx : dy, dmin = xmajor ? dy : dx;
    const float pc = xmajor ? pcx : pcy, pm = xmajor ? pcy : pcx;
    if (dmaj == 0.0f) return false;
    const float lo = fminf(a0, a1), hi = fmaxf(a0, a1);
    if (!(pc >= lo && pc < hi)) return false;
    const float at = __fadd_rn(m0, __fmul_rn(__fsub_rn(pc, a0), __fdiv_rn(dmin, dmaj)));
    const float off = __fsub_rn(pm, at);
    return off >= -1.0f && off < 1.0f;
}

// One record per sampled cell, built once per image: everything a pixel needs to test and blend the cell's line without
// redoing the square root, the six divisions and the colour ramp for every (pixel, cell) pair.
struct GlyphRec {
    float a0, a1;        // start / end along the major axis (window coordinates)
    float m0, slope;     // minor-axis coordinate at the start, d(minor) / d(major)
    float cr, cg, cb;    // source colour, already clamped and multiplied by the source alpha
    unsigned flags;      // bit 0: drawn (|v| >= 0.1 and non-degenerate), bit 1: x is the major axis
};
static_assert(sizeof(GlyphRec) == 32, "two 16-byte loads per record");

__global__ void glyph_build_kernel(const float *__restrict__ u, const float *__restrict__ v, const GridView G, int pw, int ph,
                                   int ni, int nj, GlyphRec *__restrict__ rec)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x, a = blockIdx.y * blockDim.y + threadIdx.y;   // a: i / 2, b: j / 2
    if (a >= ni || b >= nj) return;
    const int i = 2 * a, j = 2 * b;
    const size_t o = (size_t)i * G.pitch + (j - G.jbase);
    const Glyph q = make_glyph(u[o], v[o], i, j, (float)G.W, (float)G.H, (float)pw, (float)ph);
    const float dx = __fsub_rn(q.bx, q.ax), dy = __fsub_rn(q.by, q.ay);
    const bool xmajor = fabsf(dx) >= fabsf(dy);
    const float dmaj = xmajor ? dx : dy, dmin = xmajor ? dy : dx;
    GlyphRec r;
    r.a0 = xmajor ? q.ax : q.ay;
    r.a1 = xmajor ? q.bx : q.by;
    r.m0 = xmajor ? q.ay : q.ax;
    r.slope = (dmaj == 0.0f) ? 0.0f : __fdiv_rn(dmin, dmaj);
    r.cr = __fmul_rn(clamp01(q.r), 0.8f);
    r.cg = __fmul_rn(clamp01(q.g), 0.8f);
    r.cb = __fmul_rn(clamp01(q.b), 0.8f);
    r.flags = ((q.on && dmaj != 0.0f) ? 1u : 0u) | (xmajor ? 2u : 0u);
    rec[(size_t)a * nj + b] = r;
}

__global__ void view_glyphs_kernel(const GlyphRec *__restrict__ rec, int ni, int nj, const GridView G, int pw, int ph,
                                   uchar4 *__restrict__ img)
{
    const int px = blockIdx.x * blockDim.x + threadIdx.x;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= pw || py >= ph) return;
    const float fw = (float)G.W, fh = (float)G.H, fpw = (float)pw, fph = (float)ph;
    const float pcx = __fadd_rn((float)px, 0.5f), pcy = __fadd_rn((float)py, 0.5f);
    // sampled cells whose line can reach this pixel: centre within 0.05 NDC (+ 2 pixels) of it, generously rounded.
    // (Their number grows with the grid: the reference draws one line per sampled cell, and all of them within reach
    // are blended in draw order.)
    const float rx = 0.05f * 0.5f * fw + 2.0f * fw / fpw + 2.0f, ry = 0.05f * 0.5f * fh + 2.0f * fh / fph + 2.0f;
    const float ci = pcx / fpw * fw, cj = pcy / fph * fh;          // the pixel in cell units
    int i0 = max(0, (int)floorf(ci - rx)), i1 = min(G.W - 1, (int)ceilf(ci + rx));
    int j0 = max(0, (int)floorf(cj - ry)), j1 = min(G.H - 1, (int)ceilf(cj + ry));
    i0 += i0 & 1;
    j0 += j0 & 1;
    const uchar4 in = img[(size_t)py * pw + px];
    int c[4] = {in.x, in.y, in.z, in.w};
    const float sa = 0.8f, ia = __fsub_rn(1.0f, 0.8f);
    const float ca = __fmul_rn(sa, sa);
    for (int i = i0; i <= i1; i += 2)
        for (int j = j0; j <= j1; j += 2) {
            const float4 *rp = reinterpret_cast<const float4 *>(rec + (size_t)(i >> 1) * nj + (j >> 1));
            const float4 g0 = __ldg(rp);                                   // a0, a1, m0, slope
            const float lo = fminf(g0.x, g0.y), hi = fmaxf(g0.x, g0.y);
            const float4 g1 = __ldg(rp + 1);                               // colour, flags
            const unsigned fl = __float_as_uint(g1.w);
            const float pc = (fl & 2u) ? pcx : pcy, pm = (fl & 2u) ? pcy : pcx;
            if (!(fl & 1u) || !(pc >= lo && pc < hi)) continue;
            const float at = __fadd_rn(g0.z, __fmul_rn(__fsub_rn(pc, g0.x), g0.w));
            const float off = __fsub_rn(pm, at);
            if (!(off >= -1.0f && off < 1.0f)) continue;
            const float cs[4] = {g1.x, g1.y, g1.z, ca};
#pragma unroll
            for (int k = 0; k < 4; ++k)
                c[k] = to_unorm8(__fadd_rn(cs[k], __fmul_rn(__fdiv_rn((float)c[k], 255.0f), ia)));
        }
    img[(size_t)py * pw + px] = make_uchar4((unsigned char)c[0], (unsigned char)c[1], (unsigned char)c[2], (unsigned char)c[3]);
}

}  // namespace

size_t view_glyph_scratch_bytes(const GridView &G)
{
    return (size_t)((G.W + 1) / 2) * ((G.H + 1) / 2) * sizeof(GlyphRec);
}

// scratch: view_glyph_scratch_bytes(G) bytes of device memory for the per-cell line records
cudaError_t launch_view_glyphs(const float *u, const float *v, const GridView &G, int pw, int ph, unsigned char *rgba,
                               void *scratch, cudaStream_t st)
{
    if (pw < 1 || ph < 1 || !scratch) return cudaErrorInvalidValue;
    const int ni = (G.W + 1) / 2, nj = (G.H + 1) / 2;
    GlyphRec *rec = static_cast<GlyphRec *>(scratch);
    const dim3 bb(32, 8), gb((nj + 31) / 32, (ni + 7) / 8);
    glyph_build_kernel<<<gb, bb, 0, st>>>(u, v, G, pw, ph, ni, nj, rec);
    const dim3 b(32, 8), g((pw + 31) / 32, (ph + 7) / 8);
    view_glyphs_kernel<<<g, b, 0, st>>>(rec, ni, nj, G, pw, ph, reinterpret_cast<uchar4 *>(rgba));
    return cudaGetLastError();
}

cudaError_t launch_view_density(const float *d, const GridView &G, int pw, int ph, unsigned char *rgba, cudaStream_t st)
{
    if (pw < 1 || ph < 1) return cudaErrorInvalidValue;
    const dim3 b(32, 8), g((pw + 31) / 32, (ph + 7) / 8);
    view_density_kernel<<<g, b, 0, st>>>(d, G, pw, ph, reinterpret_cast<uchar4 *>(rgba));
    return cudaGetLastError();
}

}  // namespace fsb
