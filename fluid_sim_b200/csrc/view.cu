// Headless density view: the reference viewer's per-cell quads and its density shader,
// evaluated per pixel by a CUDA kernel into an RGBA8 device buffer.  No graphics API is involved.
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

}  // namespace

cudaError_t launch_view_density(const float *d, const GridView &G, int pw, int ph, unsigned char *rgba, cudaStream_t st)
{
    if (pw < 1 || ph < 1) return cudaErrorInvalidValue;
    const dim3 b(32, 8), g((pw + 31) / 32, (ph + 7) / 8);
    view_density_kernel<<<g, b, 0, st>>>(d, G, pw, ph, reinterpret_cast<uchar4 *>(rgba));
    return cudaGetLastError();
}

}  // namespace fsb
