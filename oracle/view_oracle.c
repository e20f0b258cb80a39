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
