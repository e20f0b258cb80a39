/*
 * TEST INFRASTRUCTURE ONLY.  CPU restatement of what the reference VIEWER draws for the density field
 * (SURVEY.md 8(f) rank 1), used solely as the checker of fsb_view_density in tests/.
 *
 * Parity status: UNPINNED.  The reference renders through GLFW/OpenGL (src/main.cpp:307-533); no graphics stack
 * exists on this box, so its framebuffer cannot be produced here.  This file restates the reference's own
 * arithmetic: quad corners (src/main.cpp:388-427), density shader (:32-41), blending (:376-377), clear colour
 * (:345), window 800 x 600 (:9-10).  What GL leaves implementation-defined is fixed here: a pixel belongs to a quad
 * when its centre lies in [x, x + cellWidth) x [y, y + cellHeight); the framebuffer is RGBA8 with round-to-nearest.
 *
 * Unlike the CUDA kernel (one thread per pixel looks its cell up), this follows the reference literally: it walks the
 * cells in draw order and blends every quad over the pixels it covers.
 */
#include <math.h>
#include <stdlib.h>

static float clamp01(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }
static int to_unorm8(float v) { return (int)floorf(clamp01(v) * 255.0f + 0.5f); }

/* first pixel whose centre is >= edge, searched around the analytic guess */
static int first_pixel_at_or_after(float edge, int npx)
{
    int p = (int)floorf((edge + 1.0f) * 0.5f * (float)npx - 0.5f);
    if (p < 0) p = 0;
    if (p > npx) p = npx;
    while (p > 0 && (((float)(p - 1) + 0.5f) / (float)npx * 2.0f - 1.0f) >= edge) --p;
    while (p < npx && (((float)p + 0.5f) / (float)npx * 2.0f - 1.0f) < edge) ++p;
    return p;
}

/* d: W x H floats, d[i*H + j]; out: ph x pw x 4 bytes, row 0 = bottom window row */
void ora_view_density(const float *d, int W, int H, int pw, int ph, unsigned char *out)
{
    for (long n = 0; n < (long)pw * ph; ++n) {   /* glClear with (0, 0, 0, 1) */
        out[4 * n + 0] = 0; out[4 * n + 1] = 0; out[4 * n + 2] = 0; out[4 * n + 3] = 255;
    }
    for (int i = 0; i < W; ++i) {
        for (int j = 0; j < H; ++j) {
            const float x = (float)i / (float)W * 2.0f - 1.0f;
            const float y = (float)j / (float)H * 2.0f - 1.0f;
            const float cellWidth = 2.0f / (float)W;
            const float cellHeight = 2.0f / (float)H;
            const float density = d[(long)i * H + j];
            if (density < 0.01f) continue;
            /* fragment shader */
            const float sr = clamp01(0.0f * (1.0f - density) + 1.0f * density);
            const float sg = 0.0f;
            const float sb = clamp01(1.0f * (1.0f - density) + 0.0f * density);
            const float sa = clamp01(fminf(density * 2.0f, 0.8f));
            const float src[4] = {sr, sg, sb, sa};
            const int px0 = first_pixel_at_or_after(x, pw), px1 = first_pixel_at_or_after(x + cellWidth, pw);
            const int py0 = first_pixel_at_or_after(y, ph), py1 = first_pixel_at_or_after(y + cellHeight, ph);
            for (int py = py0; py < py1; ++py)
                for (int px = px0; px < px1; ++px) {
                    unsigned char *q = out + 4 * ((long)py * pw + px);
                    for (int c = 0; c < 4; ++c) {
                        const float dst = (float)q[c] / 255.0f;
                        q[c] = (unsigned char)to_unorm8(src[c] * sa + dst * (1.0f - sa));
                    }
                }
        }
    }
}

/* ---- velocity glyphs: src/main.cpp:452-514, colour ramp src/fluid_sim.cpp:482-519 ----
 * Literal: walk the sampled cells in draw order, build the line, blend it over every pixel it covers.  The
 * coverage rule (GL leaves wide non-antialiased lines to the implementation) is the one stated in
 * include/fluid_sim_b200.h: pixel centre in [start, end) along the major axis, within [-1, 1) pixels of the line
 * along the minor axis. */
static int line_covers(float ax, float ay, float bx, float by, float pcx, float pcy)
{
    const float dx = bx - ax, dy = by - ay;
    const int xmajor = fabsf(dx) >= fabsf(dy);
    const float a0 = xmajor ? ax : ay, a1 = xmajor ? bx : by;
    const float m0 = xmajor ? ay : ax;
    const float dmaj = xmajor ? dx : dy, dmin = xmajor ? dy : dx;
    const float pc = xmajor ? pcx : pcy, pm = xmajor ? pcy : pcx;
    if (dmaj == 0.0f) return 0;
    const float lo = fminf(a0, a1), hi = fmaxf(a0, a1);
    if (!(pc >= lo && pc < hi)) return 0;
    const float at = m0 + (pc - a0) * (dmin / dmaj);
    const float off = pm - at;
    return off >= -1.0f && off < 1.0f;
}

/* u, v: W x H floats; img: ph x pw x 4 bytes already holding the density image (ora_view_density) */
void ora_view_glyphs(const float *u, const float *v, int W, int H, int pw, int ph, unsigned char *img)
{
    const float fw = (float)W, fh = (float)H, fpw = (float)pw, fph = (float)ph;
    for (int i = 0; i < W; i += 2) {
        for (int j = 0; j < H; j += 2) {
            const float vx = u[(long)i * H + j], vy = v[(long)i * H + j];
            const float magnitude = sqrtf(vx * vx + vy * vy);          /* glm::length */
            float nx = 0.0f, ny = 0.0f;
            if (magnitude > 0.001f) { nx = vx / magnitude; ny = vy / magnitude; }
            if (magnitude < 0.1f) continue;
            /* getVelocityColor */
            const float ns = fminf(magnitude / 10.0f, 1.0f);
            float r, g, b;
            if (ns < 0.25f) { r = 0.0f; g = ns / 0.25f; b = 1.0f; }
            else if (ns < 0.5f) { r = 0.0f; g = 1.0f; b = 1.0f - (ns - 0.25f) / 0.25f; }
            else if (ns < 0.75f) { r = (ns - 0.5f) / 0.25f; g = 1.0f; b = 0.0f; }
            else { r = 1.0f; g = 1.0f - (ns - 0.75f) / 0.25f; b = 0.0f; }
            const float centerX = (float)i / fw * 2.0f - 1.0f + (1.0f / fw);
            const float centerY = (float)j / fh * 2.0f - 1.0f + (1.0f / fh);
            const float endX = centerX + nx * 0.05f, endY = centerY + ny * 0.05f;
            const float ax = (centerX + 1.0f) * 0.5f * fpw, ay = (centerY + 1.0f) * 0.5f * fph;
            const float bx = (endX + 1.0f) * 0.5f * fpw, by = (endY + 1.0f) * 0.5f * fph;
            int x0 = (int)floorf(fminf(ax, bx)) - 2, x1 = (int)ceilf(fmaxf(ax, bx)) + 2;
            int y0 = (int)floorf(fminf(ay, by)) - 2, y1 = (int)ceilf(fmaxf(ay, by)) + 2;
            if (x0 < 0) x0 = 0;
            if (y0 < 0) y0 = 0;
            if (x1 > pw - 1) x1 = pw - 1;
            if (y1 > ph - 1) y1 = ph - 1;
            const float sa = 0.8f, ia = 1.0f - 0.8f;
            const float src[4] = {clamp01(r), clamp01(g), clamp01(b), sa};
            for (int py = y0; py <= y1; ++py)
                for (int px = x0; px <= x1; ++px) {
                    if (!line_covers(ax, ay, bx, by, (float)px + 0.5f, (float)py + 0.5f)) continue;
                    unsigned char *q = img + 4 * ((long)py * pw + px);
                    for (int c = 0; c < 4; ++c)
                        q[c] = (unsigned char)to_unorm8(src[c] * sa + ((float)q[c] / 255.0f) * ia);
                }
        }
    }
}
