// Bilinear sampling shared by the advection kernels (pointwise.cu) and the advection phases of the chained launch
// (chain.cu).  Same association order and round-to-nearest intrinsics as the reference (src/fluid_sim.cpp:231-252).
#pragma once
#include "fsb_internal.h"

namespace fsb {

namespace {

#define AT(G, i, j) ((size_t)(i) * (G).pitch + ((j) - (G).jbase))

// ---- bilinear sample (src/fluid_sim.cpp:231-252) -------------------------------------------------------
struct Taps { int i0, j0; float s0, s1, t0, t1; };

__device__ __forceinline__ Taps make_taps(float x, float y, float xm, float ym)
{
    x = (x < xm) ? x : xm;        // std::min(width - 1.5f, x)
    x = (0.5f < x) ? x : 0.5f;    // std::max(0.5f, .)
    y = (y < ym) ? y : ym;
    y = (0.5f < y) ? y : 0.5f;
    const int i0 = __float2int_rz(x);
    const int j0 = __float2int_rz(y);
    Taps t;
    t.s1 = __fsub_rn(x, (float)i0);
    t.s0 = __fsub_rn(1.0f, t.s1);
    t.t1 = __fsub_rn(y, (float)j0);
    t.t0 = __fsub_rn(1.0f, t.t1);
    t.i0 = i0;
    t.j0 = j0;
    return t;
}

template <bool PZ>
__device__ __forceinline__ float tap_val(const float *__restrict__ f, int i, int j, const GridView &G)
{
    const int W = G.W, H = G.H;
    float v = __ldg(f + AT(G, i, j));
    if (PZ) {
        // density only: the reference's diffuse pass with a == 0 rewrites every interior cell as
        // d + 1.5f * ((d + 0*nb) - d), i.e. d + 0.0f (src/fluid_sim.cpp:56-57); folded in here.
        if (i >= 1 && i <= W - 2 && j >= 1 && j <= H - 2) v = __fadd_rn(v, 0.0f);
    }
    return v;
}

// Idx: int when the local array has fewer than 2^31 elements (one 32-bit multiply-add per tap row instead of 64-bit
// address arithmetic per tap), long long otherwise.
// NC: the field is read-only for the whole kernel (ld.global.nc); false inside kernels that also write it.
template <bool PZ, typename Idx, bool NC = true>
__device__ __forceinline__ float sample(const float *f, const Taps &t, const GridView &G)
{
    const float *r0 = f + ((Idx)t.i0 * (Idx)G.pitch + (Idx)(t.j0 - G.jbase));
    const float *r1 = r0 + G.pitch;
    float f00, f01, f10, f11;
    if (NC) {
        f00 = __ldg(r0); f01 = __ldg(r0 + 1); f10 = __ldg(r1); f11 = __ldg(r1 + 1);
    } else {
        f00 = r0[0]; f01 = r0[1]; f10 = r1[0]; f11 = r1[1];
    }
    if (PZ) {
        // density only: the reference's diffuse pass with a == 0 rewrites every interior cell as
        // d + 1.5f * ((d + 0*nb) - d), i.e. d + 0.0f (src/fluid_sim.cpp:56-57); folded in here.
        const bool i0in = t.i0 >= 1 && t.i0 <= G.W - 2, i1in = t.i0 + 1 >= 1 && t.i0 + 1 <= G.W - 2;
        const bool j0in = t.j0 >= 1 && t.j0 <= G.H - 2, j1in = t.j0 + 1 >= 1 && t.j0 + 1 <= G.H - 2;
        if (i0in && j0in) f00 = __fadd_rn(f00, 0.0f);
        if (i0in && j1in) f01 = __fadd_rn(f01, 0.0f);
        if (i1in && j0in) f10 = __fadd_rn(f10, 0.0f);
        if (i1in && j1in) f11 = __fadd_rn(f11, 0.0f);
    }
    const float a = __fadd_rn(__fmul_rn(t.t0, f00), __fmul_rn(t.t1, f01));
    const float b = __fadd_rn(__fmul_rn(t.t0, f10), __fmul_rn(t.t1, f11));
    return __fadd_rn(__fmul_rn(t.s0, a), __fmul_rn(t.s1, b));
}

}  // namespace

}  // namespace fsb
