// Launch side of the exact-order sweep pipeline (device code: wavefront.cuh): one pass per launch (wf_kernel),
// configuration of strips per CTA for a cooperative launch, and the allocation-free pressure sweeps.
#include "wavefront.cuh"

namespace fsb {

namespace {

template <int S, int MODE, bool HAS_X0>
__global__ void __launch_bounds__(WfLayout<S>::kMaxThreads, WfLayout<S>::kMinBlocks) wf_kernel(const WfKernelParams P)
{
    extern __shared__ float4 smem4[];
    wf_pass<S, MODE, HAS_X0>(P, blockIdx.x, blockIdx.y, reinterpret_cast<float *>(smem4));
}

struct WfConfig { int nstr, ncta, c_base; size_t smem; int cluster; };

// Chooses strips per CTA so that every CTA of the launch is co-resident (cooperative launch).
template <int S, int MODE, bool HAS_X0>
cudaError_t wf_configure(const SweepLaunch &Lc, WfConfig &cfg)
{
    using L = WfLayout<S>;
    auto kern = wf_kernel<S, MODE, HAS_X0>;
    const int nstg = Lc.stages;
    const int NS_all = (Lc.H + S * nstg + 31) / 32;   // strips: the last stage's frame is shifted left by all earlier sweeps
    const int s_begin = Lc.strip_begin;
    const int s_end = (Lc.strip_end < 0 || Lc.strip_end > NS_all) ? NS_all : Lc.strip_end;
    const int NS = s_end - s_begin;                   // strips this rank runs
    if (NS < 1 || s_begin < 0) return cudaErrorInvalidValue;
    int dev = 0, nsm = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    if (nstg < 1 || nstg > L::kMaxStagesS) return cudaErrorInvalidValue;
    const int max_compute_warps = L::kMaxThreads / 32 - 2;
    // default: as many strips per CTA as fit (up to 4) -- a hand-over inside a CTA costs one group of steps, one
    // through the global rings about five, and the chain of hand-overs along the top row is what a launch waits for
    int nstr = Lc.strips_per_cta > 0 ? Lc.strips_per_cta : 4;
    while (nstr > 1 && nstr * nstg > max_compute_warps) nstr >>= 1;
    if (Lc.strips_per_cta <= 0)
        while (nstr > 1 && (wf_smem_bytes<S>(nstr, nstg) > 227 * 1024 || nstr > NS || s_begin % nstr != 0)) nstr >>= 1;
    if (nstr > NS) nstr = NS;
    const bool multi_rank = Lc.link.has_left || Lc.link.has_right || s_begin > 0;
    for (;;) {
        if (s_begin % nstr != 0) return cudaErrorInvalidValue;   // rank boundaries must fall on CTA boundaries
        cfg.nstr = nstr;
        cfg.c_base = s_begin / nstr;
        cfg.ncta = (NS + nstr - 1) / nstr;
        cfg.smem = wf_smem_bytes<S>(nstr, nstg);
        bool fits = cfg.smem <= 227 * 1024 && cfg.ncta <= kWfMaxCtas;
        if (fits) {
            e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem);
            if (e != cudaSuccess) return e;
            int per_sm = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, (nstr * nstg + 2) * 32, cfg.smem);
            if (e != cudaSuccess) return e;
            fits = (long long)cfg.ncta * Lc.nfields <= (long long)per_sm * nsm;
        }
        if (fits) {
            // Largest cluster size (CTAs that hand over through distributed shared memory) for which the whole
            // grid, padded to full clusters, is still co-resident.  FSB_CLUSTER caps it (1 = no clusters).
            static const int cap = [] {
                const char *e = std::getenv("FSB_CLUSTER");
                const int v = e ? std::atoi(e) : 8;
                return (WF_CLUSTER_MODE < 2 || v < 1) ? 1 : (v > 8 ? 8 : v);
            }();
            cfg.cluster = 1;
            for (int cs = 8; cs >= 2; cs >>= 1) {
                if (cs > cap || cfg.ncta < 2) continue;
                const int gx = (cfg.ncta + cs - 1) / cs * cs;
                cudaLaunchConfig_t lc{};
                lc.gridDim = dim3(gx, Lc.nfields);
                lc.blockDim = dim3((nstr * nstg + 2) * 32);
                lc.dynamicSmemBytes = cfg.smem;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = cs;
                at[0].val.clusterDim.y = 1;
                at[0].val.clusterDim.z = 1;
                lc.attrs = at;
                lc.numAttrs = 1;
                int nclusters = 0;
                if (cudaOccupancyMaxActiveClusters(&nclusters, kern, &lc) != cudaSuccess) {
                    (void)cudaGetLastError();
                    continue;
                }
                if ((long long)nclusters >= (long long)(gx / cs) * Lc.nfields) {
                    cfg.cluster = cs;
                    break;
                }
            }
            return cudaSuccess;
        }
        // every rank of a multi-rank chain must use the same strips-per-CTA: no silent adjustment there
        if (multi_rank || (nstr * 2) * nstg > max_compute_warps || wf_smem_bytes<S>(nstr * 2, nstg) > 227 * 1024 || nstr >= NS)
            return cudaErrorCooperativeLaunchTooLarge;
        nstr *= 2;
    }
}

template <int S, int MODE, bool HAS_X0>
cudaError_t wf_launch_inst(const SweepLaunch &Lc, SweepScratch &scr, cudaStream_t st, bool query_only)
{
    using L = WfLayout<S>;
    auto kern = wf_kernel<S, MODE, HAS_X0>;
    WfConfig cfg;
    cudaError_t e = wf_configure<S, MODE, HAS_X0>(Lc, cfg);
    if (e != cudaSuccess || query_only) return e;
    const int nstg = Lc.stages, nstr = cfg.nstr, ncta = cfg.ncta;
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
        e = cudaMemsetAsync(scr.xchg, 0, need * sizeof(float), st);   // tag 0 is never used
        if (e != cudaSuccess) return e;
        scr.xchg_floats = need;
    }

    scr.seq += 1;
    WfKernelParams P;
    for (int i = 0; i < kMaxSweepFields; ++i) P.fld[i] = Lc.fld[i < Lc.nfields ? i : 0];
    P.W = Lc.W;
    P.H = Lc.H;
    P.pitch = Lc.pitch > 0 ? Lc.pitch : Lc.H;
    P.jbase = Lc.jbase;
    P.nstr = nstr;
    P.nstg = nstg;
    P.ncta = ncta;
    P.c_base = cfg.c_base;
    P.a = Lc.a;
    P.c_recip = Lc.c_recip;
    P.omega = Lc.omega;
    P.xchg = scr.xchg;
    P.flags = scr.flags;
    P.seq_base = scr.seq << 32;
    P.link = Lc.link;
    P.link_seq_base = Lc.link.seq << 32;
    if ((Lc.link.has_left || Lc.link.has_right) && (nstg > kLinkStages || L::EXW > 44)) return cudaErrorInvalidValue;
    void *args[] = {&P};
    const int cs = cfg.cluster;
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3((ncta + cs - 1) / cs * cs, Lc.nfields);   // padding CTAs leave at once
    lc.blockDim = dim3((nstr * nstg + 2) * 32);
    lc.dynamicSmemBytes = cfg.smem;
    lc.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeCooperative;      // every CTA resident: the chain spins on its neighbours
    at[0].val.cooperative = 1;
    at[1].id = cudaLaunchAttributeClusterDimension;
    at[1].val.clusterDim.x = cs;
    at[1].val.clusterDim.y = 1;
    at[1].val.clusterDim.z = 1;
    lc.attrs = at;
    lc.numAttrs = cs > 1 ? 2 : 1;
    return cudaLaunchKernelExC(&lc, (const void *)kern, args);
}

template <int S>
cudaError_t wf_launch_levels(const SweepLaunch &Lc, SweepScratch &scr, cudaStream_t st, bool query_only)
{
    if (Lc.mode == kSor) return wf_launch_inst<S, kSor, true>(Lc, scr, st, query_only);
    bool has_x0 = Lc.fld[0].x0 != nullptr;
    for (int i = 1; i < Lc.nfields; ++i)
        if ((Lc.fld[i].x0 != nullptr) != has_x0) return cudaErrorInvalidValue;
    if (has_x0) return wf_launch_inst<S, kGaussSeidel, true>(Lc, scr, st, query_only);
    return wf_launch_inst<S, kGaussSeidel, false>(Lc, scr, st, query_only);
}

}  // namespace

size_t sweep_flag_words() { return (size_t)kMaxSweepFields * kMaxStages * kWfMaxCtas * 2; }

int sweep_max_stages(int levels)
{
    switch (levels) {
    case 1: return WfLayout<1>::kMaxStagesS;
    case 2: return WfLayout<2>::kMaxStagesS;
    case 5: return WfLayout<5>::kMaxStagesS;
    case 10: return WfLayout<10>::kMaxStagesS;
    case 20: return WfLayout<20>::kMaxStagesS;
    default: return 0;
    }
}

// ------------------------------------------------------------------------------------------------
// Pressure sweeps for hosts that must not allocate (extern "C" fsb_pressure_sweeps_noalloc, declared in
// include/fluid_sim_b200.h).  Same kernels as the projection of fsb_step, but the boundary rings and
// progress words the sweep pipeline needs between CTAs live in one fixed array inside the library image.
// ------------------------------------------------------------------------------------------------
namespace {

constexpr size_t kNaFlagBytes = (size_t)kMaxStages * kWfMaxCtas * 2 * sizeof(unsigned long long);   // 256 KiB
constexpr size_t kNaRingBytes = (size_t)16 << 20;                                                     // 16 MiB (the rings carry a tag with every float)
constexpr size_t kNaScratchBytes = kNaFlagBytes + kNaRingBytes;

// Resident on every GPU that loads the library, until the process exits.  The name below is only a
// host-side handle; the device address is obtained once per device with cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaScratchBytes];

constexpr int kNaMaxDevices = 64;
void *g_na_addr[kNaMaxDevices];                 // device address of g_noalloc_scratch, per device ordinal
unsigned long long g_na_seq[kNaMaxDevices];     // launch sequence number carried by the ring tags
// (width, height, sweeps) of the previous call per device: the ring layout inside the array depends on them
struct NaShape { int W, H, sweeps; };
NaShape g_na_shape[kNaMaxDevices];

template <int S>
cudaError_t na_launch(const SweepLaunch &Lc, float *rings, size_t ring_floats, unsigned long long *flags,
                      unsigned long long seq, cudaStream_t st, bool *too_big)
{
    using L = WfLayout<S>;
    *too_big = false;
    WfConfig cfg;
    const bool has_x0 = Lc.fld[0].x0 != nullptr;
    cudaError_t e = has_x0 ? wf_configure<S, kGaussSeidel, true>(Lc, cfg) : wf_configure<S, kGaussSeidel, false>(Lc, cfg);
    if (e != cudaSuccess) return e;
    const size_t need = (size_t)Lc.stages * cfg.ncta * L::kBoundaryFloats;
    if (need > ring_floats) {
        *too_big = true;
        return cudaSuccess;
    }
    WfKernelParams P;
    for (int i = 0; i < kMaxSweepFields; ++i) P.fld[i] = Lc.fld[0];
    P.W = Lc.W;
    P.H = Lc.H;
    P.pitch = Lc.H;
    P.jbase = 0;
    P.nstr = cfg.nstr;
    P.nstg = Lc.stages;
    P.ncta = cfg.ncta;
    P.c_base = 0;
    P.link = SweepLink{};
    P.link_seq_base = 0;
    P.a = 0.0f;
    P.c_recip = 1.0f;
    P.omega = 1.0f;
    P.xchg = rings;
    P.flags = flags;
    P.seq_base = seq << 32;
    void *args[] = {&P};
    const void *kern = has_x0 ? (const void *)wf_kernel<S, kGaussSeidel, true> : (const void *)wf_kernel<S, kGaussSeidel, false>;
    return cudaLaunchCooperativeKernel(kern, dim3(cfg.ncta, 1), dim3((cfg.nstr * Lc.stages + 2) * 32), args, cfg.smem, st);
}

cudaError_t na_launch_levels(const SweepLaunch &Lc, float *rings, size_t ring_floats, unsigned long long *flags,
                             unsigned long long seq, cudaStream_t st, bool *too_big)
{
    switch (Lc.levels) {
    case 1: return na_launch<1>(Lc, rings, ring_floats, flags, seq, st, too_big);
    case 2: return na_launch<2>(Lc, rings, ring_floats, flags, seq, st, too_big);
    case 5: return na_launch<5>(Lc, rings, ring_floats, flags, seq, st, too_big);
    case 10: return na_launch<10>(Lc, rings, ring_floats, flags, seq, st, too_big);
    case 20: return na_launch<20>(Lc, rings, ring_floats, flags, seq, st, too_big);
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace

// Returns 0 on success, 1 if the request does not fit the built-in scratch, 2 for a bad argument, otherwise
// -(cudaError_t).  `msg` receives a short description on failure.
int pressure_sweeps_noalloc(const float *d_div, float *d_p, int W, int H, int sweeps, cudaStream_t st, std::string &msg)
{
    if (!d_div || !d_p || W < 3 || H < 3 || sweeps < 1) {
        msg = "fsb_pressure_sweeps_noalloc: null pointer, grid smaller than 3x3 or sweeps < 1";
        return 2;
    }
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kNaMaxDevices) {
        msg = "fsb_pressure_sweeps_noalloc: no usable current CUDA device";
        return e != cudaSuccess ? -(int)e : 2;
    }
    if (!g_na_addr[dev]) {
        // first call on this device: translate the host-side symbol into this device's address
        void *ptr = nullptr;
        e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr) {
            msg = std::string("fsb_pressure_sweeps_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e);
            return -(int)e;
        }
        g_na_addr[dev] = ptr;
        g_na_seq[dev] = 0;
    }
    unsigned long long *flags = reinterpret_cast<unsigned long long *>(g_na_addr[dev]);
    float *rings = reinterpret_cast<float *>(static_cast<unsigned char *>(g_na_addr[dev]) + kNaFlagBytes);
    const size_t ring_floats = kNaRingBytes / sizeof(float);

    // The array keeps whatever the previous call left behind.  The progress words must not look "ahead" of
    // this call: they are cleared here.  Ring words are recognised by a tag made of (launch number mod 2^14,
    // unit); a call with the same shape as the previous one rewrites every word it later reads, but another
    // shape lays the rings out differently and could, 16384 launches later, meet a word of an older layout
    // whose tag matches by accident.  So a change of shape clears the ring region as well (a memset on the
    // caller's stream: still no allocation).  Tag 0 is never produced.
    const bool reshaped = g_na_shape[dev].W != W || g_na_shape[dev].H != H || g_na_shape[dev].sweeps != sweeps;
    e = cudaMemsetAsync(flags, 0, reshaped ? kNaScratchBytes : kNaFlagBytes, st);
    if (e != cudaSuccess) {
        msg = std::string("fsb_pressure_sweeps_noalloc: clearing the built-in scratch failed: ") + cudaGetErrorString(e);
        return -(int)e;
    }
    g_na_shape[dev] = NaShape{W, H, sweeps};

    int left = sweeps;
    bool first = true;
    while (left > 0) {
        bool placed = false;
        for (int lv : {10, 20, 5, 2, 1}) {
            if (lv > left) continue;
            int stg = left / lv;
            if (stg > sweep_max_stages(lv)) stg = sweep_max_stages(lv);
            for (; stg >= 1 && !placed; --stg) {
                SweepLaunch L{};
                L.nfields = 1;
                L.fld[0] = SweepField{first ? nullptr : d_p, d_div, d_p, 1.0f, 1.0f};
                L.W = W;
                L.H = H;
                L.mode = kGaussSeidel;
                L.levels = lv;
                L.stages = stg;
                L.strips_per_cta = 0;
                bool too_big = false;
                g_na_seq[dev] += 1;
                e = na_launch_levels(L, rings, ring_floats, flags, g_na_seq[dev], st, &too_big);
                if (e == cudaErrorCooperativeLaunchTooLarge) {
                    (void)cudaGetLastError();
                    continue;   // fewer stages need fewer resident CTAs
                }
                if (e != cudaSuccess) {
                    msg = std::string("fsb_pressure_sweeps_noalloc: launch failed: ") + cudaGetErrorString(e);
                    return -(int)e;
                }
                if (too_big) {
                    char buf[256];
                    snprintf(buf, sizeof(buf),
                             "fsb_pressure_sweeps_noalloc: a %d x %d grid with %d sweeps needs more than the %zu-byte "
                             "built-in scratch (about 8000 columns at 40 sweeps fit); use fsb_step / fsb_op_project instead",
                             W, H, sweeps, kNaRingBytes);
                    msg = buf;
                    return 1;
                }
                placed = true;
                left -= lv * stg;
            }
            if (placed) break;
        }
        if (!placed) {
            msg = "fsb_pressure_sweeps_noalloc: no launchable pipeline configuration for this grid";
            return 1;
        }
        // walls and corners of p after the pass (setBoundary(0, p), src/fluid_sim.cpp:330)
        BndJob job{d_p, 0, 0};
        e = launch_set_boundary(&job, 1, whole_grid(W, H), st);
        if (e != cudaSuccess) {
            msg = std::string("fsb_pressure_sweeps_noalloc: boundary kernel failed: ") + cudaGetErrorString(e);
            return -(int)e;
        }
        first = false;
    }
    return 0;
}

static cudaError_t sweep_dispatch(const SweepLaunch &L, SweepScratch &scr, cudaStream_t st, bool query_only)
{
    if (L.nfields < 1 || L.nfields > kMaxSweepFields || L.W < 3 || L.H < 3) return cudaErrorInvalidValue;
    switch (L.levels) {
    case 1: return wf_launch_levels<1>(L, scr, st, query_only);
    case 2: return wf_launch_levels<2>(L, scr, st, query_only);
    case 5: return wf_launch_levels<5>(L, scr, st, query_only);
    case 10: return wf_launch_levels<10>(L, scr, st, query_only);
    case 20: return wf_launch_levels<20>(L, scr, st, query_only);
    default: return cudaErrorInvalidValue;
    }
}

bool sweep_config_fits(const SweepLaunch &L)
{
    SweepScratch none;
    const cudaError_t e = sweep_dispatch(L, none, nullptr, true);
    if (e != cudaSuccess) (void)cudaGetLastError();
    return e == cudaSuccess;
}

cudaError_t launch_sweep_pass(const SweepLaunch &L, SweepScratch &scr, cudaStream_t st, int *launches)
{
    if (launches) *launches += 1;
    return sweep_dispatch(L, scr, st, false);
}

}  // namespace fsb
