// C ABI of libfluidsim_b200.so (include/fluid_sim_b200.h): simulation handle, host-side
// orchestration of FluidSim::step, lazy host mirror for the per-cell getters.
//
// Pass order follows the reference exactly (velocityStep src/fluid_sim.cpp:370-406, densityStep
// :409-422); what changes is where data lives and how passes are fused:
//   - six float32 fields resident in HBM (the reference's six live arrays);
//     the std::copy calls to prev* (:386-387, :397-398, :412, :418) become pointer swaps or
//     vanish, because every device pass writes to a different buffer than it gathers from or is
//     safe in place;
//   - the two velocity advects share one RK4 back-trace (:401-402 trace through the same fields);
//   - density "diffusion" with DIFFUSION == 0 (include/fluid_sim.h:14) is the identity d + 0.0f plus
//     the boundary update and is folded into the advect gather.
#include "../../include/fluid_sim_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>

#include "fsb_internal.h"

using namespace fsb;

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string &msg)
{
    g_last_error = msg;
    return code;
}

int fail_cuda(cudaError_t e, const char *what)
{
    g_last_error = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    return e == cudaErrorMemoryAllocation ? FSB_ENOMEM : FSB_ECUDA;
}

#define FSB_CUDA(call)                                         \
    do {                                                       \
        cudaError_t e__ = (call);                              \
        if (e__ != cudaSuccess) return fail_cuda(e__, #call);  \
    } while (0)

#define FSB_TRY(call)                  \
    do {                               \
        int r__ = (call);              \
        if (r__ != FSB_OK) return r__; \
    } while (0)

constexpr int kNumFields = 6;

struct SweepPlan { int levels, stages; };   // one launch of the sweep pipeline: sweeps per warp x chained warps per strip

// Position of the process-wide rand() stream (fsb_params.noise_process_stream): outputs consumed so far by all
// simulations that share it.  The reference's generator is libc's, one per process (src/fluid_sim.cpp:378-379).
std::mutex g_proc_noise_mu;
uint64_t g_proc_noise_pos = 0;

// Live simulations of this process (whole-grid ones).  Overlapping steps give up the co-residency guarantee of a
// cooperative launch: a launch is resident as a whole because its predecessor in the stream was.  Two simulations that
// step concurrently on one device could each hold part of the SMs and wait for the rest, so the overlap is used only
// by a simulation that is the only one of its device that has it enabled.
// (never destroyed: a namespace-scope FluidSim, src/main.cpp:7, is torn down after this file's own statics)
std::mutex &g_sims_mu = *new std::mutex;
std::vector<struct fsb_sim *> &g_sims = *new std::vector<struct fsb_sim *>;

}  // namespace

struct fsb_sim {
    fsb_params prm;
    int W = 0, H = 0;      // global grid
    GridView G{};          // this process's view of it (the whole grid unless the simulation is a column block)
    size_t ncells = 0;     // floats per local field (W * pitch)
    int device = 0;
    cudaStream_t stream = nullptr;
    float *f[kNumFields] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    float *temp[2] = {nullptr, nullptr};   // MacCormack scratch (src/fluid_sim.cpp:52-53 tempField1/2), lazy
    SweepScratch sweep;
    ChainState *chain = nullptr;           // passes chained inside one launch (chain.cu), lazily created
    int chain_level = 2;                   // 0: every pass its own launch, 1: diffuse + project chained, 2: the whole step
    int chain_overlap = 1;                 // the whole step INCLUDING the noise as one launch, consecutive steps overlapping
    // Buffers the overlapping steps rotate through besides f[]: [0..1], [2..3] two more velocity sets (a step must not
    // write what the step before still traces through), [4..5] pressure / divergence of the first projection (f[]'s prev
    // arrays hold a velocity set then), [6] a third density.  Allocated on the first such step.
    float *spare[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    // Asynchronous density reads of a caller that reads after every step (fsb_read_field_async): the step's launch
    // writes a second copy of its density (two buffers, alternating) which the copy stream then transfers, instead of
    // a device-to-device snapshot between two launches (which would keep them from overlapping).
    float *snap[2] = {nullptr, nullptr};
    cudaEvent_t ev_snap_copied[2] = {nullptr, nullptr};
    bool snap_busy[2] = {false, false};    // a transfer out of snap[k] has been queued and not yet waited for
    int snap_next = 0, snap_idx = -1;      // buffer the next step writes; buffer that holds the current density (-1: none)
    bool want_snap = false;                // the caller has been reading the density after its steps
    int steps_since_read = 0;
    cudaEvent_t last_copy = nullptr;       // completion of the most recent asynchronous read
    NoiseTables noise_tables;
    uint32_t noise_win[kRandDeg];          // window of the emulated rand() stream at noise_pos
    uint64_t noise_pos = 0;
    RandJump noise_step_jump;              // advance by 2*W*H outputs
    std::vector<SourceEntry> pending;
    SourceEntry *d_entries = nullptr;
    int *d_segs = nullptr;
    size_t d_entries_cap = 0, d_segs_cap = 0;
    struct SourceStage { unsigned char *host = nullptr; size_t bytes = 0; cudaEvent_t done = nullptr; bool used = false; };
    std::vector<std::pair<uint64_t, std::vector<SweepPlan>>> plan_cache;   // (sweeps, fields, mode) -> launches
    SourceStage src_stage[2];              // page-locked staging of pending sources, used alternately
    int src_next = 0;
    std::vector<float> mirror[3];
    bool mirror_valid[3] = {false, false, false};
    int probes[3] = {0, 0, 0};             // single-cell reads served from the device since the fields last changed
    uint64_t launches = 0;
    int tune_strips = 0;   // strips per CTA (0 = automatic)
    int tune_levels = 0;   // sweeps per warp (0 = automatic)
    int tune_stages = 0;   // cap on chained warps per strip (0 = automatic)
    // column block of a multi-rank run (is_block == false: the whole grid)
    bool is_block = false;
    int strip_begin = 0, strip_end = -1;       // 32-column strips of the sweep chain this rank runs
    // fsb_read_field_async: a device-to-host copy on its own stream, ordered against later writers of that buffer
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_ready = nullptr, ev_copied = nullptr;
    bool copy_pending = false;
    unsigned *vmax_word = nullptr;              // column blocks: largest |v| written by the gradient passes (float bits)
    float *snapshot = nullptr;                  // device-side copy of the field an asynchronous read is transferring
    unsigned char *view_buf = nullptr;          // RGBA8 image of the last fsb_view_density* call
    size_t view_bytes = 0;
    void *glyph_buf = nullptr;                  // per-cell line records of the velocity-glyph view
    size_t glyph_bytes = 0;
    unsigned char *link_buf = nullptr;          // [rings | in_prod | cons], shared with both neighbours through CUDA IPC
    void *peer_left = nullptr, *peer_right = nullptr;   // the neighbours' link_buf mapped into this process
    float *peer_left_f[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // left rank's fields (same roles)
    void *peer_left_base[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr}; // as opened (for closing)
    int left_ncols = 0;                                  // columns owned by the left rank
    // chained launch on column blocks: the neighbours' fields (role order, swapped with this rank's), second-projection
    // scratch and chain buffers, mapped through CUDA IPC
    int halo = 0;
    bool blk_chain_ready = false;
    float *peer_l_f[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    float *peer_r_f[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    void *peer_chain[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};   // [left / right][buf, extra0, extra1]
    void *peer_r_base[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    unsigned long long blk_steps = 0;
    // optional per-kernel-class timing (CUDA events on the simulation's stream)
    bool profiling = false;
    struct ProfRec { int cat; cudaEvent_t a, b; };
    std::vector<ProfRec> prof_open;                 // recorded, not yet read back
    std::vector<cudaEvent_t> prof_free;             // recycled events
    double prof_ms[FSB_PROF_COUNT] = {0};
    uint64_t prof_n[FSB_PROF_COUNT] = {0};
};

namespace {

struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) {
            cudaSetDevice(dev);
            switched = true;
        }
    }
    ~DeviceGuard()
    {
        if (switched) cudaSetDevice(prev);
    }
};

cudaEvent_t prof_event(fsb_sim *s)
{
    cudaEvent_t e = nullptr;
    if (!s->prof_free.empty()) {
        e = s->prof_free.back();
        s->prof_free.pop_back();
    } else if (cudaEventCreate(&e) != cudaSuccess) {
        e = nullptr;
    }
    return e;
}

// Brackets the kernels launched inside its scope with two events when profiling is on.
struct ProfScope {
    fsb_sim *s;
    int cat;
    cudaEvent_t a = nullptr;
    ProfScope(fsb_sim *sim, int c) : s(sim), cat(c)
    {
        if (s->profiling && (a = prof_event(s))) cudaEventRecord(a, s->stream);
    }
    ~ProfScope()
    {
        if (!a) return;
        cudaEvent_t b = prof_event(s);
        if (!b) { s->prof_free.push_back(a); return; }
        cudaEventRecord(b, s->stream);
        s->prof_open.push_back(fsb_sim::ProfRec{cat, a, b});
    }
};

inline float sign_i(int b) { return b == 1 ? -1.0f : 1.0f; }
inline float sign_j(int b) { return b == 2 ? -1.0f : 1.0f; }


// Splits n sweeps into launches of (sweeps per warp) x (stages): one launch when n factors into an
// instantiated size that can be launched with every CTA resident, otherwise greedily.
// pref / max_stages are tuning overrides (0 = automatic).
std::vector<SweepPlan> plan_sweeps_uncached(const fsb_sim *s, int n, int nfields, int mode);

// The plan depends on (sweeps, fields, mode) and the tuning only; probing a configuration costs several runtime calls
// (function attributes, occupancy), so plans are kept until fsb_set_tuning changes the tuning.
const std::vector<SweepPlan> &plan_sweeps(fsb_sim *s, int n, int nfields, int mode)
{
    const uint64_t key = ((uint64_t)(uint32_t)n << 8) | ((uint64_t)nfields << 4) | (uint64_t)mode;
    for (auto &c : s->plan_cache)
        if (c.first == key) return c.second;
    s->plan_cache.emplace_back(key, plan_sweeps_uncached(s, n, nfields, mode));
    return s->plan_cache.back().second;
}

std::vector<SweepPlan> plan_sweeps_uncached(const fsb_sim *s, int n, int nfields, int mode)
{
    const int pref = s->tune_levels, max_stages = s->tune_stages;
    std::vector<int> sizes;
    if (pref > 0 && sweep_max_stages(pref) > 0) sizes.push_back(pref);
    for (int lv : {10, 20, 5, 2, 1})
        if (lv != pref) sizes.push_back(lv);
    auto cap = [&](int lv) {
        int m = sweep_max_stages(lv);
        return (max_stages > 0 && max_stages < m) ? max_stages : m;
    };
    auto fits = [&](int lv, int st) {
        SweepLaunch L{};
        L.nfields = nfields;
        L.W = s->W;
        L.H = s->H;
        L.mode = mode;
        L.levels = lv;
        L.stages = st;
        L.strips_per_cta = s->tune_strips;
        for (int i = 0; i < nfields; ++i) L.fld[i] = SweepField{s->f[0], s->f[0], s->f[0], 1.0f, 1.0f};
        return sweep_config_fits(L);
    };
    std::vector<SweepPlan> out;
    for (int lv : sizes)
        if (n > 0 && n % lv == 0 && n / lv <= cap(lv) && fits(lv, n / lv)) {
            out.push_back(SweepPlan{lv, n / lv});
            return out;
        }
    while (n > 0) {
        bool placed = false;
        for (int lv : sizes) {
            if (lv > n) continue;
            int st = std::min(n / lv, cap(lv));
            while (st > 1 && !fits(lv, st)) --st;
            if (!fits(lv, st)) continue;
            out.push_back(SweepPlan{lv, st});
            n -= lv * st;
            placed = true;
            break;
        }
        if (!placed) {   // nothing launchable: let the launch report the error
            out.push_back(SweepPlan{1, 1});
            n -= 1;
        }
    }
    return out;
}

int set_noise_window(fsb_sim *s, uint32_t seed, uint64_t pos)
{
    rand_seed_window(seed, s->noise_win);
    if (pos) {
        RandJump J = jump_matrix(pos);
        apply_jump(J, s->noise_win);
    }
    s->noise_pos = pos;
    s->prm.noise_seed = seed;
    return FSB_OK;
}

int set_boundary(fsb_sim *s, const BndJob *jobs, int n)
{
    ProfScope ps(s, FSB_PROF_BOUNDARY);
    FSB_CUDA(launch_set_boundary(jobs, n, s->G, s->stream));
    s->launches += 1;
    return FSB_OK;
}

int set_boundary1(fsb_sim *s, float *x, int b, int plus_zero = 0)
{
    BndJob j{x, b, plus_zero};
    return set_boundary(s, &j, 1);
}

// An asynchronous read transfers a device-side snapshot (fsb_read_field_async), so later work may overwrite the field.
int order_after_copy(fsb_sim *) { return FSB_OK; }

enum SourceSel { kAllSources = 0, kVelocitySources = 1, kDensitySources = 2 };

int flush_sources(fsb_sim *s, SourceSel sel = kAllSources)
{
    if (sel == kAllSources) FSB_TRY(order_after_copy(s));      // every entry point but fsb_step comes through here
    if (s->pending.empty()) return FSB_OK;
    s->snap_idx = -1;                                           // the density is about to change
    std::vector<SourceEntry> chosen, rest;
    if (sel != kAllSources) {
        for (const SourceEntry &q : s->pending) ((sel == kDensitySources) == (q.field == 0) ? chosen : rest).push_back(q);
        s->pending.swap(chosen);                                  // flushed now; `rest` is restored below
        if (s->pending.empty()) {
            s->pending.swap(rest);
            return FSB_OK;
        }
    }
    struct Restore {
        std::vector<SourceEntry> &dst, &src;
        ~Restore() { dst.insert(dst.end(), src.begin(), src.end()); }
    } restore{s->pending, rest};
    // group by (field, cell) keeping call order inside a group: `+=` is applied in the order of the calls
    std::vector<SourceEntry> &e = s->pending;
    std::stable_sort(e.begin(), e.end(), [](const SourceEntry &a, const SourceEntry &b) {
        return a.field != b.field ? a.field < b.field : a.cell < b.cell;
    });
    std::vector<int> segs;
    for (size_t k = 0; k < e.size(); ++k)
        if (k == 0 || e[k].field != e[k - 1].field || e[k].cell != e[k - 1].cell) segs.push_back((int)k);
    const int nseg = (int)segs.size();
    segs.push_back((int)e.size());
    // Staging: the entries and segment starts go to the device from one of two page-locked buffers; a buffer is reused
    // only after the copy that read it last has finished (event), so fsb_step never waits for the GPU here (the
    // reference's mouse handler adds sources every frame, src/main.cpp:216-252).  Buffers grow geometrically.
    const size_t ebytes = e.size() * sizeof(SourceEntry), sbytes = segs.size() * sizeof(int);
    const size_t need = ebytes + sbytes;
    fsb_sim::SourceStage &stg = s->src_stage[s->src_next];
    s->src_next ^= 1;
    if (stg.used) FSB_CUDA(cudaEventSynchronize(stg.done));      // two flushes ago: long finished
    if (need > stg.bytes) {
        if (stg.host) cudaFreeHost(stg.host);
        stg.host = nullptr;
        stg.bytes = 0;
        const size_t cap = std::max<size_t>(need * 2, 1 << 16);
        FSB_CUDA(cudaHostAlloc(&stg.host, cap, cudaHostAllocDefault));
        stg.bytes = cap;
    }
    if (!stg.done) FSB_CUDA(cudaEventCreateWithFlags(&stg.done, cudaEventDisableTiming));
    if (e.size() > s->d_entries_cap) {
        FSB_CUDA(cudaStreamSynchronize(s->stream));                // rare: the device buffer grows
        if (s->d_entries) cudaFree(s->d_entries);
        s->d_entries = nullptr;
        s->d_entries_cap = std::max<size_t>(e.size() * 2, 4096);
        FSB_CUDA(cudaMalloc(&s->d_entries, s->d_entries_cap * sizeof(SourceEntry)));
    }
    if (segs.size() > s->d_segs_cap) {
        FSB_CUDA(cudaStreamSynchronize(s->stream));
        if (s->d_segs) cudaFree(s->d_segs);
        s->d_segs = nullptr;
        s->d_segs_cap = std::max<size_t>(segs.size() * 2, 4096);
        FSB_CUDA(cudaMalloc(&s->d_segs, s->d_segs_cap * sizeof(int)));
    }
    std::memcpy(stg.host, e.data(), ebytes);
    std::memcpy(stg.host + ebytes, segs.data(), sbytes);
    FSB_CUDA(cudaMemcpyAsync(s->d_entries, stg.host, ebytes, cudaMemcpyHostToDevice, s->stream));
    FSB_CUDA(cudaMemcpyAsync(s->d_segs, stg.host + ebytes, sbytes, cudaMemcpyHostToDevice, s->stream));
    FSB_CUDA(launch_apply_sources(s->f[FSB_DENSITY], s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y], s->d_entries,
                                  s->d_segs, nseg, s->stream));
    FSB_CUDA(cudaEventRecord(stg.done, s->stream));
    stg.used = true;
    s->launches += 1;
    e.clear();
    return FSB_OK;
}

// diffuse() of src/fluid_sim.cpp:43-62 on up to two fields sharing the coefficient `diff`.
// fold_identity: with a == 0 only the boundary is updated and *folded is set, the caller then reads
// interior values as (v + 0.0f).
int diffuse_fields(fsb_sim *s, int nf, float *const *xs, const int *bs, float diff, float dt, bool fold_identity,
                   int *folded)
{
    if (folded) *folded = 0;
    const int D = s->prm.diff_iters;
    if (D <= 0) return FSB_OK;
    const float a = dt * diff * (float)s->W * (float)s->H;   // ((dt*diff)*W)*H, :45
    const float c_recip = 1.0f / (1.0f + 4.0f * a);          // :46
    if (a == 0.0f) {
        BndJob jobs[2];
        for (int i = 0; i < nf; ++i) {
            if (!fold_identity) {
                FSB_CUDA(launch_add_zero(xs[i], s->ncells, s->stream));
                s->launches += 1;
            }
            jobs[i] = BndJob{xs[i], bs[i], fold_identity ? 1 : 0};
        }
        FSB_TRY(set_boundary(s, jobs, nf));
        if (folded && fold_identity) *folded = 1;
        return FSB_OK;
    }
    const std::vector<SweepPlan> &passes = plan_sweeps(s, D, nf, kSor);
    const bool multi = passes.size() > 1;
    float *rhs[2] = {xs[0], nf > 1 ? xs[1] : nullptr};
    if (multi) {
        // the source must stay the pre-diffusion field for every pass: keep a copy (prev* in the reference)
        float *scratch[2] = {s->f[FSB_PREV_VEL_X], s->f[FSB_PREV_VEL_Y]};
        for (int i = 0; i < nf; ++i) {
            if (scratch[i] == xs[i]) scratch[i] = s->f[FSB_PREV_DENSITY];
            FSB_CUDA(cudaMemcpyAsync(scratch[i], xs[i], s->ncells * sizeof(float), cudaMemcpyDeviceToDevice, s->stream));
            rhs[i] = scratch[i];
        }
    }
    for (size_t pi = 0; pi < passes.size(); ++pi) {
        SweepLaunch L{};
        L.nfields = nf;
        for (int i = 0; i < nf; ++i) L.fld[i] = SweepField{xs[i], rhs[i], xs[i], sign_i(bs[i]), sign_j(bs[i])};
        L.W = s->W;
        L.H = s->H;
        L.mode = kSor;
        L.levels = passes[pi].levels;
        L.stages = passes[pi].stages;
        L.a = a;
        L.c_recip = c_recip;
        L.omega = s->prm.omega;
        L.strips_per_cta = s->tune_strips;
        int nl = 0;
        {
            ProfScope ps(s, FSB_PROF_SOR);
            FSB_CUDA(launch_sweep_pass(L, s->sweep, s->stream, &nl));
        }
        s->launches += nl;
        BndJob jobs[2];
        for (int i = 0; i < nf; ++i) jobs[i] = BndJob{xs[i], bs[i], 0};
        FSB_TRY(set_boundary(s, jobs, nf));
    }
    return FSB_OK;
}

// project() of src/fluid_sim.cpp:301-344
int project(fsb_sim *s, float *u, float *v, float *p, float *div)
{
    {
        ProfScope ps(s, FSB_PROF_DIVERGENCE);
        FSB_CUDA(launch_divergence(u, v, div, s->G, s->stream));
    }
    s->launches += 1;
    const std::vector<SweepPlan> &passes = plan_sweeps(s, s->prm.proj_iters, 1, kGaussSeidel);
    if (passes.empty()) FSB_CUDA(cudaMemsetAsync(p, 0, s->ncells * sizeof(float), s->stream));
    for (size_t pi = 0; pi < passes.size(); ++pi) {
        SweepLaunch L{};
        L.nfields = 1;
        L.fld[0] = SweepField{pi == 0 ? nullptr : p, div, p, 1.0f, 1.0f};   // p starts at 0 (:313, :316)
        L.W = s->W;
        L.H = s->H;
        L.mode = kGaussSeidel;
        L.levels = passes[pi].levels;
        L.stages = passes[pi].stages;
        L.a = 0.0f;
        L.c_recip = 1.0f;
        L.omega = 1.0f;
        L.strips_per_cta = s->tune_strips;
        int nl = 0;
        {
            ProfScope ps(s, FSB_PROF_GS);
            FSB_CUDA(launch_sweep_pass(L, s->sweep, s->stream, &nl));
        }
        s->launches += nl;
        FSB_TRY(set_boundary1(s, p, 0));
    }
    {
        ProfScope ps(s, FSB_PROF_GRADIENT);
        FSB_CUDA(launch_gradient(u, v, p, s->G, s->stream));
    }
    s->launches += 1;
    BndJob jobs[2] = {BndJob{u, 1, 0}, BndJob{v, 2, 0}};
    return set_boundary(s, jobs, 2);
}

int ensure_temps(fsb_sim *s)
{
    for (int i = 0; i < 2; ++i)
        if (!s->temp[i]) {
            FSB_CUDA(cudaMalloc(&s->temp[i], s->ncells * sizeof(float)));
            FSB_CUDA(cudaMemsetAsync(s->temp[i], 0, s->ncells * sizeof(float), s->stream));
        }
    return FSB_OK;
}

// advect() of src/fluid_sim.cpp:212-228 for one field (dst must not alias src, u or v)
int advect_one(fsb_sim *s, int b, float *dst, const float *src, const float *u, const float *v, float dt, int pz)
{
    const int scheme = s->prm.advect_scheme;
    if (scheme == FSB_ADVECT_RK4) {
        ProfScope ps(s, FSB_PROF_ADVECT_SCALAR);
        FSB_CUDA(launch_advect_rk4(dst, src, nullptr, nullptr, u, v, pz, s->G, dt, s->stream));
        s->launches += 1;
    } else if (scheme == FSB_ADVECT_SEMI_LAGRANGIAN) {
        FSB_CUDA(launch_advect_sl(dst, src, u, v, pz, s->G, dt, -1.0f, s->stream));
        s->launches += 1;
    } else {
        FSB_TRY(ensure_temps(s));
        FSB_CUDA(launch_advect_sl(s->temp[0], src, u, v, pz, s->G, dt, -1.0f, s->stream));
        FSB_TRY(set_boundary1(s, s->temp[0], b));
        FSB_CUDA(launch_advect_sl(s->temp[1], s->temp[0], u, v, 0, s->G, dt, +1.0f, s->stream));
        FSB_TRY(set_boundary1(s, s->temp[1], b));
        FSB_CUDA(launch_maccormack_combine(dst, src, s->temp[0], s->temp[1], pz, s->G, s->stream));
        s->launches += 3;
    }
    return set_boundary1(s, dst, b);
}

// The window of the rand() stream at the first output of this step's noise; moves the stream on by the 2*W*H outputs
// the step consumes (src/fluid_sim.cpp:374-383).
void take_noise_window(fsb_sim *s, NoiseWindow *nw)
{
    const bool shared = s->prm.noise_process_stream && s->prm.noise_mode == FSB_NOISE_GLIBC && !s->is_block;
    std::unique_lock<std::mutex> lock(g_proc_noise_mu, std::defer_lock);
    if (shared) {
        // another simulation of the process may have drawn from the stream since this one's last step
        lock.lock();
        if (s->noise_pos != g_proc_noise_pos) set_noise_window(s, s->prm.noise_seed, g_proc_noise_pos);
    }
    std::memcpy(nw->w, s->noise_win, sizeof(nw->w));
    // the reference consumes 2*W*H outputs per step in either mode (the zero mode replaces rand() itself)
    if (s->prm.noise_mode != FSB_NOISE_ZERO) {
        apply_jump(s->noise_step_jump, s->noise_win);
        s->noise_pos += 2ull * (uint64_t)s->W * (uint64_t)s->H;
    }
    if (shared) g_proc_noise_pos = s->noise_pos;
}

int add_noise(fsb_sim *s)
{
    NoiseWindow nw;
    take_noise_window(s, &nw);
    const int zero = s->prm.noise_mode == FSB_NOISE_ZERO;
    ProfScope ps(s, FSB_PROF_NOISE);
    if (s->is_block)
        FSB_CUDA(launch_noise_block(s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y], s->G, nw, s->noise_tables, zero, s->stream));
    else
        FSB_CUDA(launch_noise(s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y], s->ncells, nw, s->noise_tables, zero, s->stream));
    s->launches += zero ? 2 : 1;
    return FSB_OK;
}

// Passes of the step as phases of ONE launch whose CTAs move from pass to pass as soon as the columns they need are
// done (chain.cu).  with_diffuse: diffuse(u), diffuse(v) + project; otherwise project alone; whole_step: everything
// after the noise (diffuse, project, advect, project, and the density step when its diffusion coefficient is 0).
// *done = false: no chained configuration for this grid / these sweep counts / this tuning; the caller runs the passes
// one launch at a time.  *density_done tells whether the density step was part of the launch.
int chained_launch(fsb_sim *s, bool with_diffuse, bool whole_step, float dt, bool *done, bool *density_done)
{
    *done = false;
    if (density_done) *density_done = false;
    if (s->chain_level < 1 || s->is_block || s->tune_strips || s->tune_levels || s->tune_stages) return FSB_OK;
    // Measured (tools/chain_ab.py, ms per step, one launch per pass -> chained): 8192^2 P=40 20.9 -> 16.4, 1024 x 8192
    // 12.7 -> 6.5, 4096^2 10.0 -> 7.8, 2048^2 4.33 -> 3.69, 200^2 0.64 -> 0.62; grids much taller than wide lose
    // with few columns lose (16384 x 2048: 14.8 -> 18.7: every block is a long serial walk for one CTA and there are not
    // enough blocks for the SMs) and keep one launch per pass; 32768 x 4096: 68.4 -> 66.2.
    // FSB_CHAIN=3 / fsb_set_chain_level(3) forces the chained launch wherever a configuration exists.
    if (s->chain_level < 3 && s->W > 2 * s->H && s->H < 4096) return FSB_OK;
    // the advection phases of the chained launch are RK4 (the live scheme, src/fluid_sim.cpp:219)
    if (whole_step && (s->chain_level < 2 || s->prm.advect_scheme != FSB_ADVECT_RK4)) return FSB_OK;
    ChainArgs A{};
    A.W = s->W;
    A.H = s->H;
    A.u = s->f[FSB_VELOCITY_X];
    A.v = s->f[FSB_VELOCITY_Y];
    A.p = s->f[FSB_PREV_VEL_X];
    A.div = s->f[FSB_PREV_VEL_Y];
    A.gs_sweeps = s->prm.proj_iters;
    A.sor_sweeps = 0;
    A.a = 0.0f;
    A.c_recip = 1.0f;
    A.omega = s->prm.omega;
    A.dt = dt;
    if (with_diffuse) {
        A.a = dt * s->prm.viscosity * (float)s->W * (float)s->H;   // ((dt*diff)*W)*H, src/fluid_sim.cpp:45
        A.c_recip = 1.0f / (1.0f + 4.0f * A.a);                    // :46
        A.sor_sweeps = s->prm.diff_iters;
        if (A.sor_sweeps <= 0 || A.a == 0.0f) return FSB_OK;
    }
    if (whole_step) {
        A.with_advect = 1;
        const float ad = dt * s->prm.diffusion * (float)s->W * (float)s->H;
        if (ad == 0.0f) {
            // density diffusion with a == 0 (DIFFUSION, include/fluid_sim.h:14): identity d + 0.0f plus the boundary
            // update when the loop runs at all; the advection is then a phase of the same launch
            A.with_density = 1;
            A.density_pz = s->prm.diff_iters > 0 ? 1 : 0;
            A.d = s->f[FSB_DENSITY];
            A.nd = s->f[FSB_PREV_DENSITY];
        }
    }
    cudaError_t e = launch_chain(A, s->chain, s->sweep, s->stream, 1);
    if (e == cudaErrorNotSupported) return FSB_OK;
    if (e == cudaSuccess) {
        ProfScope ps(s, FSB_PROF_CHAIN);
        e = launch_chain(A, s->chain, s->sweep, s->stream, 0);
    }
    if (e != cudaSuccess) return fail_cuda(e, "chained launch");
    s->launches += 1;
    *done = true;
    if (whole_step) {
        // the advected velocity sits in the former prev buffers (src/fluid_sim.cpp:397-402 copies, here a swap of roles)
        std::swap(s->f[FSB_VELOCITY_X], s->f[FSB_PREV_VEL_X]);
        std::swap(s->f[FSB_VELOCITY_Y], s->f[FSB_PREV_VEL_Y]);
        if (A.with_density) {
            std::swap(s->f[FSB_DENSITY], s->f[FSB_PREV_DENSITY]);
            if (density_done) *density_done = true;
        }
    }
    return FSB_OK;
}

// FluidSim::step (src/fluid_sim.cpp:25-28) as ONE launch, noise included, in buffers that rotate so that the launch of
// the next step may start on the SMs this one's first CTAs have left (chain.cu, launch_chain).  Step n reads the velocity
// V(n-1) = f[velocity], writes V(n-1) + noise, diffuses and projects in X(n), advects into Y(n) and projects again; it
// advects the density f[density] into D(n).  Afterwards f[velocity] = Y(n), f[prev velocity] = X(n), f[density] = D(n),
// f[prev density] = the density the step started from, and the two velocity sets / the density that were current and
// previous BEFORE the step become the spares the next step writes: nothing a step writes is read by the step before it
// anywhere but in the block being written.
// *done = false: this grid / these parameters have no such configuration; nothing has been changed.
bool only_overlapping_sim_of_its_device(const fsb_sim *s);

int chained_step(fsb_sim *s, float dt, bool *done)
{
    *done = false;
    if (s->chain_level < 2 || !s->chain_overlap || s->is_block || s->tune_strips || s->tune_levels || s->tune_stages) return FSB_OK;
    if (!only_overlapping_sim_of_its_device(s)) return FSB_OK;
    if (s->chain_level < 3 && s->W > 2 * s->H && s->H < 4096) return FSB_OK;      // see chained_launch
    if (s->prm.advect_scheme != FSB_ADVECT_RK4) return FSB_OK;
    if (s->prm.noise_mode != FSB_NOISE_GLIBC && s->prm.noise_mode != FSB_NOISE_ZERO) return FSB_OK;
    const float ad = dt * s->prm.diffusion * (float)s->W * (float)s->H;
    if (ad != 0.0f) return FSB_OK;               // the density diffusion would run between the launches
    ChainArgs A{};
    A.W = s->W;
    A.H = s->H;
    A.gs_sweeps = s->prm.proj_iters;
    A.omega = s->prm.omega;
    A.dt = dt;
    A.a = dt * s->prm.viscosity * (float)s->W * (float)s->H;       // ((dt*diff)*W)*H, src/fluid_sim.cpp:45
    A.c_recip = 1.0f / (1.0f + 4.0f * A.a);                        // :46
    A.sor_sweeps = s->prm.diff_iters;
    if (A.sor_sweeps <= 0 || A.a == 0.0f) return FSB_OK;
    A.with_advect = 1;
    A.with_density = 1;
    A.density_pz = s->prm.diff_iters > 0 ? 1 : 0;
    A.overlap = 1;
    A.noise_zero = s->prm.noise_mode == FSB_NOISE_ZERO;
    // any non-null pointers do for the query
    A.vin_u = s->f[FSB_VELOCITY_X];
    A.vin_v = s->f[FSB_VELOCITY_Y];
    NoiseWindow nw{};
    A.noise_base = nw.w;
    cudaError_t e = launch_chain(A, s->chain, s->sweep, s->stream, 1);
    if (e == cudaErrorNotSupported) return FSB_OK;
    if (e != cudaSuccess) return fail_cuda(e, "chained step");
    for (float *&b : s->spare)
        if (!b) {
            if (cudaMalloc(&b, s->ncells * sizeof(float)) != cudaSuccess) {
                // not enough device memory for the rotation: this simulation keeps the noise as its own launch and its
                // steps one after the other (seven fields fewer)
                (void)cudaGetLastError();
                b = nullptr;
                FSB_CUDA(cudaStreamSynchronize(s->stream));
                for (float *&q : s->spare) {
                    if (q) cudaFree(q);
                    q = nullptr;
                }
                s->chain_overlap = 0;
                return FSB_OK;
            }
            FSB_CUDA(cudaMemsetAsync(b, 0, s->ncells * sizeof(float), s->stream));
        }
    // pending sources: a few (the reference's mouse handler, src/main.cpp:216-252) travel as parameters of the launch
    // and are applied by its noise phase; more than that go through the staged copy + kernel
    if ((int)s->pending.size() > kChainSources) FSB_TRY(flush_sources(s));
    std::stable_sort(s->pending.begin(), s->pending.end(), [](const SourceEntry &a, const SourceEntry &b) {
        return a.field != b.field ? a.field < b.field : a.cell < b.cell;
    });
    A.sources = s->pending.data();
    A.nsources = (int)s->pending.size();
    int snap_k = -1;
    if (s->want_snap) {
        snap_k = s->snap_next;
        if (!s->snap[snap_k] && cudaMalloc(&s->snap[snap_k], s->ncells * sizeof(float)) != cudaSuccess) {
            (void)cudaGetLastError();          // no memory for the second copy: reads go through the device-side snapshot
            s->snap[snap_k] = nullptr;
            s->want_snap = false;
            snap_k = -1;
        }
    }
    if (snap_k >= 0) {
        if (!s->ev_snap_copied[snap_k]) FSB_CUDA(cudaEventCreateWithFlags(&s->ev_snap_copied[snap_k], cudaEventDisableTiming));
        // the transfer out of this buffer was queued two steps ago
        if (s->snap_busy[snap_k]) FSB_CUDA(cudaStreamWaitEvent(s->stream, s->ev_snap_copied[snap_k], 0));
        A.snap = s->snap[snap_k];
    }
    take_noise_window(s, &nw);
    A.u = s->spare[0];
    A.v = s->spare[1];
    A.yu = s->spare[2];
    A.yv = s->spare[3];
    A.p = s->spare[4];
    A.div = s->spare[5];
    A.d = s->f[FSB_DENSITY];
    A.nd = s->spare[6];
    {
        ProfScope ps(s, FSB_PROF_CHAIN);
        e = launch_chain(A, s->chain, s->sweep, s->stream, 0);
    }
    if (e != cudaSuccess) return fail_cuda(e, "chained step");
    s->launches += 1;
    float *const vin_u = s->f[FSB_VELOCITY_X], *const vin_v = s->f[FSB_VELOCITY_Y];
    float *const prev_u = s->f[FSB_PREV_VEL_X], *const prev_v = s->f[FSB_PREV_VEL_Y];
    s->f[FSB_VELOCITY_X] = A.yu;
    s->f[FSB_VELOCITY_Y] = A.yv;
    s->f[FSB_PREV_VEL_X] = A.u;
    s->f[FSB_PREV_VEL_Y] = A.v;
    s->spare[0] = vin_u;
    s->spare[1] = vin_v;
    s->spare[2] = prev_u;
    s->spare[3] = prev_v;
    float *const prev_d = s->f[FSB_PREV_DENSITY];
    s->f[FSB_PREV_DENSITY] = s->f[FSB_DENSITY];
    s->f[FSB_DENSITY] = A.nd;
    s->spare[6] = prev_d;
    s->pending.clear();
    s->snap_idx = snap_k;
    if (snap_k >= 0) {
        s->snap_next ^= 1;
        if (++s->steps_since_read > 2) s->want_snap = false;       // nobody reads: stop writing the second copy
    }
    *done = true;
    return FSB_OK;
}

bool only_overlapping_sim_of_its_device(const fsb_sim *s)
{
    std::lock_guard<std::mutex> lock(g_sims_mu);
    for (const fsb_sim *o : g_sims)
        if (o != s && o->device == s->device && o->chain_overlap && o->chain_level >= 2) return false;
    return true;
}

int chained_project(fsb_sim *s, bool with_diffuse, float dt, bool *done)
{
    return chained_launch(s, with_diffuse, false, dt, done, nullptr);
}

int velocity_step_after_noise(fsb_sim *s, float dt)
{
    bool chained = false;
    FSB_TRY(chained_project(s, true, dt, &chained));
    if (!chained) {
        float *xs[2] = {s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y]};
        const int bs[2] = {1, 2};
        FSB_TRY(diffuse_fields(s, 2, xs, bs, s->prm.viscosity, dt, false, nullptr));
        FSB_TRY(project(s, s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y], s->f[FSB_PREV_VEL_X], s->f[FSB_PREV_VEL_Y]));
    }
    // advect both components out of (u, v) into the free prev buffers, then swap roles
    float *u = s->f[FSB_VELOCITY_X], *v = s->f[FSB_VELOCITY_Y];
    float *nu = s->f[FSB_PREV_VEL_X], *nv = s->f[FSB_PREV_VEL_Y];
    if (s->prm.advect_scheme == FSB_ADVECT_RK4) {
        {
            ProfScope ps(s, FSB_PROF_ADVECT_VELOCITY);
            FSB_CUDA(launch_advect_rk4(nu, u, nv, v, u, v, 0, s->G, dt, s->stream));
        }
        s->launches += 1;
        BndJob jobs[2] = {BndJob{nu, 1, 0}, BndJob{nv, 2, 0}};
        FSB_TRY(set_boundary(s, jobs, 2));
    } else {
        FSB_TRY(advect_one(s, 1, nu, u, u, v, dt, 0));
        FSB_TRY(advect_one(s, 2, nv, v, u, v, dt, 0));
    }
    std::swap(s->f[FSB_VELOCITY_X], s->f[FSB_PREV_VEL_X]);
    std::swap(s->f[FSB_VELOCITY_Y], s->f[FSB_PREV_VEL_Y]);
    FSB_TRY(chained_project(s, false, dt, &chained));
    if (chained) return FSB_OK;
    return project(s, s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y], s->f[FSB_PREV_VEL_X], s->f[FSB_PREV_VEL_Y]);
}

int velocity_step(fsb_sim *s, float dt)
{
    FSB_TRY(add_noise(s));
    return velocity_step_after_noise(s, dt);
}

int density_step(fsb_sim *s, float dt)
{
    s->snap_idx = -1;
    float *xs[1] = {s->f[FSB_DENSITY]};
    const int bs[1] = {0};
    int folded = 0;
    FSB_TRY(diffuse_fields(s, 1, xs, bs, s->prm.diffusion, dt, true, &folded));
    FSB_TRY(advect_one(s, 0, s->f[FSB_PREV_DENSITY], s->f[FSB_DENSITY], s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y], dt,
                       folded));
    std::swap(s->f[FSB_DENSITY], s->f[FSB_PREV_DENSITY]);
    return FSB_OK;
}

void invalidate_mirror(fsb_sim *s)
{
    s->mirror_valid[0] = s->mirror_valid[1] = s->mirror_valid[2] = false;
    s->probes[0] = s->probes[1] = s->probes[2] = 0;
}

// A caller that probes a few cells after a step (not the viewer's full sweep over the grid, src/main.cpp:388-396) should
// not pay for a whole-field download: the first kProbeCells reads after a change fetch single cells; a caller that keeps
// reading gets the host mirror of the whole field, as the reference's per-cell read loop wants.
constexpr int kProbeCells = 32;

int read_cell(fsb_sim *s, int field, int x, int y, float *out, bool *served)
{
    *served = false;
    if (s->is_block) return fail(FSB_EINVAL, "per-cell getters are not available on a column block; read the field instead");
    if (s->mirror_valid[field] || s->probes[field] >= kProbeCells) return FSB_OK;
    FSB_TRY(flush_sources(s));
    s->probes[field] += 1;
    FSB_CUDA(cudaMemcpyAsync(out, s->f[field] + (size_t)x * s->H + y, sizeof(float), cudaMemcpyDeviceToHost, s->stream));
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    *served = true;
    return FSB_OK;
}

int refresh_mirror(fsb_sim *s, int field)
{
    if (s->is_block) return fail(FSB_EINVAL, "per-cell getters are not available on a column block; read the field instead");
    if (s->mirror_valid[field]) return FSB_OK;
    FSB_TRY(flush_sources(s));
    s->mirror[field].resize(s->ncells);
    FSB_CUDA(cudaMemcpyAsync(s->mirror[field].data(), s->f[field], s->ncells * sizeof(float), cudaMemcpyDeviceToHost,
                             s->stream));
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    s->mirror_valid[field] = true;
    return FSB_OK;
}

inline bool in_range(const fsb_sim *s, int x, int y) { return x >= 0 && x < s->W && y >= 0 && y < s->H; }

bool valid_field(int f) { return f >= 0 && f < kNumFields; }

}  // namespace

// ===================================== C ABI ============================================

extern "C" {

void fsb_default_params(fsb_params *p)
{
    if (!p) return;
    std::memset(p, 0, sizeof(*p));
    p->width = 200;
    p->height = 200;
    p->proj_iters = 20;
    p->diff_iters = 5;
    p->viscosity = 0.0001f;
    p->diffusion = 0.0f;
    p->omega = 1.5f;
    p->advect_scheme = FSB_ADVECT_RK4;
    p->noise_mode = FSB_NOISE_GLIBC;
    p->noise_seed = 1;
    p->noise_pos = 0;
    p->device = -1;
}

static int create_common(const fsb_params *p, const GridView *block, fsb_sim **out);

int fsb_create(const fsb_params *p, fsb_sim **out) { return create_common(p, nullptr, out); }

int fsb_create_block(const fsb_params *p, int col_begin, int col_end, int halo, fsb_sim **out)
{
    if (!p || !out) return fail(FSB_EINVAL, "fsb_create_block: null argument");
    if (col_begin < 0 || col_end > p->height || col_begin >= col_end || halo < 0 || (halo & 31) || (col_begin & 31))
        return fail(FSB_EINVAL, "fsb_create_block: columns must lie in the grid, col_begin and halo must be multiples of 32");
    GridView G;
    G.W = p->width;
    G.H = p->height;
    G.pitch = ((col_end - col_begin) + 2 * halo + 3) & ~3;
    G.jbase = col_begin - halo;
    G.j0 = col_begin;
    G.j1 = col_end;
    const int rc = create_common(p, &G, out);
    if (rc == FSB_OK) (*out)->halo = halo;
    return rc;
}

static int create_common(const fsb_params *p, const GridView *block, fsb_sim **out)
{
    if (!out) return fail(FSB_EINVAL, "fsb_create: out is null");
    *out = nullptr;
    fsb_params prm;
    if (p) prm = *p;
    else fsb_default_params(&prm);
    if (prm.width < 3 || prm.height < 3) return fail(FSB_EINVAL, "fsb_create: grid must be at least 3x3");
    if (prm.proj_iters < 0 || prm.diff_iters < 0) return fail(FSB_EINVAL, "fsb_create: negative sweep count");
    if (prm.advect_scheme < 0 || prm.advect_scheme > 2) return fail(FSB_EINVAL, "fsb_create: unknown advection scheme");
    if ((long long)prm.width * prm.height > (1ll << 31) - 1) return fail(FSB_ELIMIT, "fsb_create: more than 2^31-1 cells");
    if (block && (long long)prm.width * block->pitch > (1ll << 31) - 1)
        return fail(FSB_ELIMIT, "fsb_create_block: more than 2^31-1 cells in the local arrays (cell indices are 32-bit)");

    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return fail(FSB_ECUDA, "fsb_create: no CUDA device (this library has no CPU fallback)");
    int dev = prm.device;
    if (dev < 0) FSB_CUDA(cudaGetDevice(&dev));
    if (dev >= ndev) return fail(FSB_EINVAL, "fsb_create: device ordinal out of range");

    fsb_sim *s = new (std::nothrow) fsb_sim();
    if (!s) return fail(FSB_ENOMEM, "fsb_create: host allocation failed");
    s->prm = prm;
    s->W = prm.width;
    s->H = prm.height;
    s->G = block ? *block : whole_grid(prm.width, prm.height);
    s->is_block = block != nullptr;
    s->ncells = (size_t)prm.width * s->G.pitch;
    if (block) {
        s->strip_begin = block->j0 / 32;
        s->strip_end = (block->j1 >= prm.height) ? -1 : block->j1 / 32;
    }
    s->device = dev;
    DeviceGuard guard(dev);

    auto bail = [&](int code) {
        fsb_destroy(s);
        return code;
    };
#define FSB_CREATE_CUDA(call)                                           \
    do {                                                                \
        cudaError_t e__ = (call);                                       \
        if (e__ != cudaSuccess) return bail(fail_cuda(e__, #call));     \
    } while (0)

    FSB_CREATE_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    for (int k = 0; k < kNumFields; ++k) {
        FSB_CREATE_CUDA(cudaMalloc(&s->f[k], s->ncells * sizeof(float)));
        FSB_CREATE_CUDA(cudaMemsetAsync(s->f[k], 0, s->ncells * sizeof(float), s->stream));
    }
    FSB_CREATE_CUDA(cudaMalloc(&s->sweep.flags, sweep_flag_words() * sizeof(unsigned long long)));
    FSB_CREATE_CUDA(cudaMemsetAsync(s->sweep.flags, 0, sweep_flag_words() * sizeof(unsigned long long), s->stream));
    FSB_CREATE_CUDA(noise_tables_create(s->noise_tables));
    if (block) {
        FSB_CREATE_CUDA(noise_tables_create_pow2(s->noise_tables, prm.height));
        FSB_CREATE_CUDA(cudaMalloc(&s->link_buf, kLinkBytes));
        FSB_CREATE_CUDA(cudaMemsetAsync(s->link_buf, 0, kLinkBytes, s->stream));
    }
    {
        // FSB_CHAIN = 0: every pass as its own launch, 1: diffuse + project / project as chained launches, 2 (default):
        // the whole step after the noise as one chained launch where that is faster, 3: wherever it is possible
        const char *e = std::getenv("FSB_CHAIN");
        if (e && e[0] >= '0' && e[0] <= '3') s->chain_level = e[0] - '0';
        // FSB_CHAIN_OVERLAP = 0: the noise as its own launch and every step's launch after the one before has drained
        const char *o = std::getenv("FSB_CHAIN_OVERLAP");
        if (o && (o[0] == '0' || o[0] == '1')) s->chain_overlap = o[0] - '0';
    }
    s->noise_step_jump = jump_matrix(2ull * (uint64_t)prm.width * (uint64_t)prm.height);
    set_noise_window(s, prm.noise_seed, prm.noise_pos);
    FSB_CREATE_CUDA(cudaStreamSynchronize(s->stream));
#undef FSB_CREATE_CUDA
    if (!s->is_block) {
        std::lock_guard<std::mutex> lock(g_sims_mu);
        g_sims.push_back(s);
    }
    *out = s;
    return FSB_OK;
}

int fsb_destroy(fsb_sim *s)
{
    if (!s) return FSB_OK;
    {
        std::lock_guard<std::mutex> lock(g_sims_mu);
        g_sims.erase(std::remove(g_sims.begin(), g_sims.end(), s), g_sims.end());
    }
    {
        DeviceGuard guard(s->device);
        // tolerate a runtime that is already shutting down (a namespace-scope FluidSim, src/main.cpp:7,
        // is destroyed after main returns): errors are ignored here on purpose
        if (s->stream) cudaStreamSynchronize(s->stream);
        for (int k = 0; k < kNumFields; ++k)
            if (s->f[k]) cudaFree(s->f[k]);
        for (int k = 0; k < 2; ++k)
            if (s->temp[k]) cudaFree(s->temp[k]);
        for (float *b : s->spare)
            if (b) cudaFree(b);
        if (s->sweep.xchg) cudaFree(s->sweep.xchg);
        if (s->sweep.flags) cudaFree(s->sweep.flags);
        chain_state_destroy(s->chain);
        s->chain = nullptr;
        if (s->d_entries) cudaFree(s->d_entries);
        if (s->d_segs) cudaFree(s->d_segs);
        for (auto &stg : s->src_stage) {
            if (stg.host) cudaFreeHost(stg.host);
            if (stg.done) cudaEventDestroy(stg.done);
        }
        noise_tables_destroy(s->noise_tables);
        for (int k = 0; k < 6; ++k)
            if (s->peer_left_base[k]) cudaIpcCloseMemHandle(s->peer_left_base[k]);
        for (int k = 0; k < 6; ++k)
            if (s->peer_r_base[k]) cudaIpcCloseMemHandle(s->peer_r_base[k]);
        for (int side = 0; side < 2; ++side)
            for (int k = 0; k < 3; ++k)
                if (s->peer_chain[side][k]) cudaIpcCloseMemHandle(s->peer_chain[side][k]);
        if (s->peer_left) cudaIpcCloseMemHandle(s->peer_left);
        if (s->peer_right) cudaIpcCloseMemHandle(s->peer_right);
        if (s->link_buf) cudaFree(s->link_buf);
        if (s->view_buf) cudaFree(s->view_buf);
        if (s->glyph_buf) cudaFree(s->glyph_buf);
        if (s->snapshot) cudaFree(s->snapshot);
        for (int k = 0; k < 2; ++k) {
            if (s->snap[k]) cudaFree(s->snap[k]);
            if (s->ev_snap_copied[k]) cudaEventDestroy(s->ev_snap_copied[k]);
        }
        if (s->vmax_word) cudaFree(s->vmax_word);
        for (auto &r : s->prof_open) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
        for (auto &e : s->prof_free) cudaEventDestroy(e);
        if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
        if (s->ev_ready) cudaEventDestroy(s->ev_ready);
        if (s->ev_copied) cudaEventDestroy(s->ev_copied);
        if (s->stream) cudaStreamDestroy(s->stream);
        (void)cudaGetLastError();
    }
    delete s;
    return FSB_OK;
}

int fsb_reset(fsb_sim *s)
{
    if (!s) return fail(FSB_EINVAL, "fsb_reset: null handle");
    DeviceGuard guard(s->device);
    s->pending.clear();
    s->snap_idx = -1;
    FSB_TRY(order_after_copy(s));
    for (int k = 0; k < kNumFields; ++k) FSB_CUDA(cudaMemsetAsync(s->f[k], 0, s->ncells * sizeof(float), s->stream));
    for (int k = 0; k < 2; ++k)
        if (s->temp[k]) FSB_CUDA(cudaMemsetAsync(s->temp[k], 0, s->ncells * sizeof(float), s->stream));
    invalidate_mirror(s);
    return FSB_OK;
}

int fsb_step(fsb_sim *s, float dt)
{
    if (!s) return fail(FSB_EINVAL, "fsb_step: null handle");
    DeviceGuard guard(s->device);
    invalidate_mirror(s);
    s->snap_idx = -1;
    {
        // the whole step as one launch (which also applies the pending sources), or everything after the noise as one
        // launch, when this grid / these parameters have a chained configuration
        bool query = false, dens = false;
        if (s->chain_level >= 2 && !s->is_block) {
            FSB_TRY(chained_step(s, dt, &query));
            if (query) return FSB_OK;
            FSB_TRY(flush_sources(s));       // density sources too: the velocity step never reads the density
            FSB_TRY(add_noise(s));
            FSB_TRY(chained_launch(s, true, true, dt, &query, &dens));
            if (query) {
                if (!dens) FSB_TRY(density_step(s, dt));
                return FSB_OK;
            }
            FSB_TRY(velocity_step_after_noise(s, dt));
            return density_step(s, dt);
        }
    }
    FSB_TRY(flush_sources(s));
    FSB_TRY(velocity_step(s, dt));   // src/fluid_sim.cpp:26
    FSB_TRY(density_step(s, dt));    // src/fluid_sim.cpp:27
    return FSB_OK;
}

int fsb_sync(fsb_sim *s)
{
    if (!s) return fail(FSB_EINVAL, "fsb_sync: null handle");
    DeviceGuard guard(s->device);
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    return FSB_OK;
}

int fsb_add_density(fsb_sim *s, int x, int y, float amount)
{
    if (!s) return fail(FSB_EINVAL, "fsb_add_density: null handle");
    if (!in_range(s, x, y)) return FSB_OK;   // silently ignored, src/fluid_sim.cpp:426
    // a column block also keeps the columns of its halo current (the neighbour that owns them applies the same call)
    if (y < s->G.jbase || y >= s->G.jbase + s->G.pitch) return FSB_OK;   // a column this rank does not store
    const int cell = x * s->G.pitch + (y - s->G.jbase);
    s->pending.push_back(SourceEntry{cell, 0, amount});
    if (s->mirror_valid[0]) s->mirror[0][cell] += amount;
    return FSB_OK;
}

int fsb_add_velocity(fsb_sim *s, int x, int y, float ax, float ay)
{
    if (!s) return fail(FSB_EINVAL, "fsb_add_velocity: null handle");
    if (!in_range(s, x, y)) return FSB_OK;   // src/fluid_sim.cpp:434
    if (y < s->G.jbase || y >= s->G.jbase + s->G.pitch) return FSB_OK;   // a column this rank does not store
    const int cell = x * s->G.pitch + (y - s->G.jbase);
    s->pending.push_back(SourceEntry{cell, 1, ax});
    s->pending.push_back(SourceEntry{cell, 2, ay});
    if (s->mirror_valid[1]) s->mirror[1][cell] += ax;
    if (s->mirror_valid[2]) s->mirror[2][cell] += ay;
    return FSB_OK;
}

int fsb_add_density_n(fsb_sim *s, int n, const int *xs, const int *ys, const float *amounts)
{
    if (!s || n < 0 || (n > 0 && (!xs || !ys || !amounts))) return fail(FSB_EINVAL, "fsb_add_density_n: bad argument");
    for (int k = 0; k < n; ++k) FSB_TRY(fsb_add_density(s, xs[k], ys[k], amounts[k]));
    return FSB_OK;
}

int fsb_add_velocity_n(fsb_sim *s, int n, const int *xs, const int *ys, const float *axs, const float *ays)
{
    if (!s || n < 0 || (n > 0 && (!xs || !ys || !axs || !ays))) return fail(FSB_EINVAL, "fsb_add_velocity_n: bad argument");
    for (int k = 0; k < n; ++k) FSB_TRY(fsb_add_velocity(s, xs[k], ys[k], axs[k], ays[k]));
    return FSB_OK;
}

int fsb_get_density(fsb_sim *s, int x, int y, float *out)
{
    if (!s || !out) return fail(FSB_EINVAL, "fsb_get_density: bad argument");
    *out = 0.0f;
    if (!in_range(s, x, y)) return FSB_OK;   // src/fluid_sim.cpp:447
    DeviceGuard guard(s->device);
    bool served = false;
    FSB_TRY(read_cell(s, 0, x, y, out, &served));
    if (served) return FSB_OK;
    FSB_TRY(refresh_mirror(s, 0));
    *out = s->mirror[0][(size_t)x * s->H + y];
    return FSB_OK;
}

int fsb_get_velocity(fsb_sim *s, int x, int y, float *out_xy)
{
    if (!s || !out_xy) return fail(FSB_EINVAL, "fsb_get_velocity: bad argument");
    out_xy[0] = out_xy[1] = 0.0f;
    if (!in_range(s, x, y)) return FSB_OK;   // src/fluid_sim.cpp:456
    DeviceGuard guard(s->device);
    for (int k = 0; k < 2; ++k) {
        bool served = false;
        FSB_TRY(read_cell(s, 1 + k, x, y, out_xy + k, &served));
        if (served) continue;
        FSB_TRY(refresh_mirror(s, 1 + k));
        out_xy[k] = s->mirror[1 + k][(size_t)x * s->H + y];
    }
    return FSB_OK;
}

int fsb_get_velocity_magnitude(fsb_sim *s, int x, int y, float *out)
{
    float v[2];
    if (!out) return fail(FSB_EINVAL, "fsb_get_velocity_magnitude: bad argument");
    FSB_TRY(fsb_get_velocity(s, x, y, v));
    *out = in_range(s, x, y) ? std::sqrt(v[0] * v[0] + v[1] * v[1]) : 0.0f;   // glm::length, src/fluid_sim.cpp:477
    return FSB_OK;
}

int fsb_get_normalized_velocity(fsb_sim *s, int x, int y, float *out_xy)
{
    float v[2];
    if (!out_xy) return fail(FSB_EINVAL, "fsb_get_normalized_velocity: bad argument");
    FSB_TRY(fsb_get_velocity(s, x, y, v));
    out_xy[0] = out_xy[1] = 0.0f;
    if (in_range(s, x, y)) {
        const float m = std::sqrt(v[0] * v[0] + v[1] * v[1]);
        if (m > 0.001f) {   // src/fluid_sim.cpp:465
            out_xy[0] = v[0] / m;
            out_xy[1] = v[1] / m;
        }
    }
    return FSB_OK;
}

int fsb_get_velocity_color(fsb_sim *s, int x, int y, float *rgb)
{
    float speed = 0.0f;
    if (!rgb) return fail(FSB_EINVAL, "fsb_get_velocity_color: bad argument");
    FSB_TRY(fsb_get_velocity_magnitude(s, x, y, &speed));
    // four-segment blue -> cyan -> green -> yellow -> red ramp, src/fluid_sim.cpp:482-519
    const float q = speed / 10.0f;
    const float n = (q < 1.0f) ? q : 1.0f;
    if (n < 0.25f) {
        rgb[0] = 0.0f; rgb[1] = n / 0.25f; rgb[2] = 1.0f;
    } else if (n < 0.5f) {
        rgb[0] = 0.0f; rgb[1] = 1.0f; rgb[2] = 1.0f - (n - 0.25f) / 0.25f;
    } else if (n < 0.75f) {
        rgb[0] = (n - 0.5f) / 0.25f; rgb[1] = 1.0f; rgb[2] = 0.0f;
    } else {
        rgb[0] = 1.0f; rgb[1] = 1.0f - (n - 0.75f) / 0.25f; rgb[2] = 0.0f;
    }
    return FSB_OK;
}

int fsb_get_width(const fsb_sim *s) { return s ? s->W : 0; }
int fsb_get_height(const fsb_sim *s) { return s ? s->H : 0; }

int fsb_read_field(fsb_sim *s, int field, float *host_dst)
{
    if (!s || !host_dst || !valid_field(field)) return fail(FSB_EINVAL, "fsb_read_field: bad argument");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    FSB_CUDA(cudaMemcpyAsync(host_dst, s->f[field], s->ncells * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    return FSB_OK;
}

int fsb_write_field(fsb_sim *s, int field, const float *host_src)
{
    if (!s || !host_src || !valid_field(field)) return fail(FSB_EINVAL, "fsb_write_field: bad argument");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    s->snap_idx = -1;
    FSB_CUDA(cudaMemcpyAsync(s->f[field], host_src, s->ncells * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    if (field < 3) s->mirror_valid[field] = false;
    return FSB_OK;
}

/* Queues a device-to-host copy of a field as it is after all work queued so far: a device-side snapshot on the
 * simulation's stream (so later steps and sources may overwrite the field at once), then the transfer to the host on a
 * second stream, overlapping with later fsb_step calls.  One read at a time. */
int fsb_read_field_async(fsb_sim *s, int field, float *host_dst)
{
    if (!s || !host_dst || !valid_field(field)) return fail(FSB_EINVAL, "fsb_read_field_async: bad argument");
    if (s->is_block) return fail(FSB_EINVAL, "fsb_read_field_async: not available on a column block");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    if (!s->copy_stream) {
        FSB_CUDA(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
        FSB_CUDA(cudaEventCreateWithFlags(&s->ev_ready, cudaEventDisableTiming));
        FSB_CUDA(cudaEventCreateWithFlags(&s->ev_copied, cudaEventDisableTiming));
    }
    if (field == FSB_DENSITY) {
        // a caller that reads the density after its steps: from now on the steps' launches leave a second copy of it
        s->want_snap = true;
        s->steps_since_read = 0;
        if (s->snap_idx >= 0) {
            // the last step did: transfer that copy once the step has finished (an event, nothing else, goes between
            // this step's launch and the next one's, so they still overlap)
            const int k = s->snap_idx;
            FSB_CUDA(cudaEventRecord(s->ev_ready, s->stream));
            FSB_CUDA(cudaStreamWaitEvent(s->copy_stream, s->ev_ready, 0));
            FSB_CUDA(cudaMemcpyAsync(host_dst, s->snap[k], s->ncells * sizeof(float), cudaMemcpyDeviceToHost, s->copy_stream));
            FSB_CUDA(cudaEventRecord(s->ev_snap_copied[k], s->copy_stream));
            s->snap_busy[k] = true;
            s->last_copy = s->ev_snap_copied[k];
            s->copy_pending = true;
            return FSB_OK;
        }
    }
    if (!s->snapshot) FSB_CUDA(cudaMalloc(&s->snapshot, s->ncells * sizeof(float)));
    if (s->copy_pending) FSB_CUDA(cudaStreamWaitEvent(s->stream, s->ev_copied, 0));   // the snapshot buffer is free again
    FSB_CUDA(cudaMemcpyAsync(s->snapshot, s->f[field], s->ncells * sizeof(float), cudaMemcpyDeviceToDevice, s->stream));
    FSB_CUDA(cudaEventRecord(s->ev_ready, s->stream));
    FSB_CUDA(cudaStreamWaitEvent(s->copy_stream, s->ev_ready, 0));
    FSB_CUDA(cudaMemcpyAsync(host_dst, s->snapshot, s->ncells * sizeof(float), cudaMemcpyDeviceToHost, s->copy_stream));
    FSB_CUDA(cudaEventRecord(s->ev_copied, s->copy_stream));
    s->last_copy = s->ev_copied;
    s->copy_pending = true;
    return FSB_OK;
}

/* Blocks the host until the copy queued by fsb_read_field_async has arrived (no-op if there is none). */
int fsb_read_wait(fsb_sim *s)
{
    if (!s) return fail(FSB_EINVAL, "fsb_read_wait: null handle");
    if (!s->copy_pending) return FSB_OK;
    DeviceGuard guard(s->device);
    FSB_CUDA(cudaEventSynchronize(s->last_copy));     // the copy stream is in order: every earlier read has arrived too
    s->copy_pending = false;
    s->snap_busy[0] = s->snap_busy[1] = false;
    return FSB_OK;
}

int fsb_add_field(fsb_sim *s, int field, const float *host_src)
{
    if (!s || !host_src || field < 0 || field > 2) return fail(FSB_EINVAL, "fsb_add_field: bad argument");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    // stage through a scratch field that no pass reads between steps
    float *stage = s->f[FSB_PREV_DENSITY];
    FSB_CUDA(cudaMemcpyAsync(stage, host_src, s->ncells * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    FSB_CUDA(launch_axpy_field(s->f[field], stage, s->ncells, s->stream));
    s->launches += 1;
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    s->mirror_valid[field] = false;
    return FSB_OK;
}

int fsb_device_field(fsb_sim *s, int field, void **dptr)
{
    if (!s || !dptr || !valid_field(field)) return fail(FSB_EINVAL, "fsb_device_field: bad argument");
    *dptr = s->f[field];
    return FSB_OK;
}

int fsb_get_noise_pos(const fsb_sim *s, uint64_t *pos)
{
    if (!s || !pos) return fail(FSB_EINVAL, "fsb_get_noise_pos: bad argument");
    if (s->prm.noise_process_stream && !s->is_block) {
        std::lock_guard<std::mutex> lock(g_proc_noise_mu);
        *pos = g_proc_noise_pos;
    } else {
        *pos = s->noise_pos;
    }
    return FSB_OK;
}

int fsb_set_noise(fsb_sim *s, int mode, uint32_t seed, uint64_t pos)
{
    if (!s || (mode != FSB_NOISE_GLIBC && mode != FSB_NOISE_ZERO)) return fail(FSB_EINVAL, "fsb_set_noise: bad argument");
    s->prm.noise_mode = mode;
    if (s->prm.noise_process_stream && !s->is_block) {     // repositions the stream for every simulation that shares it
        std::lock_guard<std::mutex> lock(g_proc_noise_mu);
        g_proc_noise_pos = pos;
    }
    return set_noise_window(s, seed, pos);
}

int fsb_op_add_noise(fsb_sim *s)
{
    if (!s) return fail(FSB_EINVAL, "null handle");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    invalidate_mirror(s);
    return add_noise(s);
}

int fsb_op_set_boundary(fsb_sim *s, int b, int field)
{
    if (!s || !valid_field(field) || b < 0 || b > 2) return fail(FSB_EINVAL, "fsb_op_set_boundary: bad argument");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    invalidate_mirror(s);
    return set_boundary1(s, s->f[field], b);
}

int fsb_op_diffuse(fsb_sim *s, int b, int field, float diff, float dt)
{
    if (!s || !valid_field(field) || b < 0 || b > 2) return fail(FSB_EINVAL, "fsb_op_diffuse: bad argument");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    invalidate_mirror(s);
    float *xs[1] = {s->f[field]};
    const int bs[1] = {b};
    return diffuse_fields(s, 1, xs, bs, diff, dt, false, nullptr);
}

int fsb_op_project(fsb_sim *s)
{
    if (!s) return fail(FSB_EINVAL, "null handle");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    invalidate_mirror(s);
    return project(s, s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y], s->f[FSB_PREV_VEL_X], s->f[FSB_PREV_VEL_Y]);
}

int fsb_op_advect(fsb_sim *s, int b, int dst, int src, int u, int v, float dt)
{
    if (!s || !valid_field(dst) || !valid_field(src) || !valid_field(u) || !valid_field(v) || b < 0 || b > 2)
        return fail(FSB_EINVAL, "fsb_op_advect: bad argument");
    if (dst == src || dst == u || dst == v) return fail(FSB_EINVAL, "fsb_op_advect: dst must not alias an input");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    invalidate_mirror(s);
    return advect_one(s, b, s->f[dst], s->f[src], s->f[u], s->f[v], dt, 0);
}

int fsb_op_velocity_step(fsb_sim *s, float dt)
{
    if (!s) return fail(FSB_EINVAL, "null handle");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    invalidate_mirror(s);
    return velocity_step(s, dt);
}

int fsb_op_density_step(fsb_sim *s, float dt)
{
    if (!s) return fail(FSB_EINVAL, "null handle");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    invalidate_mirror(s);
    return density_step(s, dt);
}

int fsb_launch_count(const fsb_sim *s, uint64_t *count)
{
    if (!s || !count) return fail(FSB_EINVAL, "fsb_launch_count: bad argument");
    *count = s->launches;
    return FSB_OK;
}

int fsb_set_tuning(fsb_sim *s, int strips_per_cta, int sweeps_per_warp, int max_stages)
{
    if (!s || strips_per_cta < 0 || strips_per_cta > 16 || sweeps_per_warp < 0 || max_stages < 0)
        return fail(FSB_EINVAL, "fsb_set_tuning: bad argument");
    if (sweeps_per_warp > 0 && sweep_max_stages(sweeps_per_warp) == 0)
        return fail(FSB_EINVAL, "fsb_set_tuning: sweeps_per_warp must be 1, 2, 5, 10 or 20");
    s->tune_strips = strips_per_cta;
    s->tune_levels = sweeps_per_warp;
    s->tune_stages = max_stages;
    s->plan_cache.clear();
    return FSB_OK;
}

int fsb_set_chain_level(fsb_sim *s, int level)
{
    if (!s || level < 0 || level > 3) return fail(FSB_EINVAL, "fsb_set_chain_level: bad argument");
    s->chain_level = level;
    return FSB_OK;
}

int fsb_dbg_chain_describe(int variant, int width, int height, int diff_iters, int proj_iters, int nsm, int col_begin,
                           int col_end, int halo, long long *out, int max_words, int *nphases, int *grid, int *cw)
{
    if (!out || !nphases || !grid || !cw || width < 3 || height < 3 || nsm < 1 || variant < 0 || variant > 3)
        return fail(FSB_EINVAL, "fsb_dbg_chain_describe: bad argument");
    if (variant == 2 && (col_begin < 0 || col_end > height || col_begin >= col_end || halo < 0 || (halo & 31) || (col_begin & 31)))
        return fail(FSB_EINVAL, "fsb_dbg_chain_describe: bad column block");
    // the arguments chained_step (variant 0) / chained_launch for a whole step (variant 1) pass, with small integers
    // standing in for the arrays
    auto arr = [](int id) { return reinterpret_cast<float *>((size_t)id * 0x1000); };
    ChainArgs A{};
    A.W = width;
    A.H = height;
    A.gs_sweeps = proj_iters;
    A.sor_sweeps = diff_iters;
    A.a = 1.0f;
    A.c_recip = 0.2f;
    A.omega = 1.5f;
    A.dt = 0.01f;
    A.with_advect = 1;
    A.with_density = 1;
    A.density_pz = 1;
    A.overlap = 1;
    A.noise_zero = 1;
    const SourceEntry some_sources[2] = {SourceEntry{width + 1, 0, 1.0f}, SourceEntry{width + 1, 1, 1.0f}};
    if (variant == 0 || variant == 3) {
        A.vin_u = arr(1); A.vin_v = arr(2);
        A.yu = arr(5); A.yv = arr(6);
        if (variant == 3) {         // the same launch carrying pending sources (one addDensity, one addVelocity)
            A.sources = some_sources;
            A.nsources = 2;
        }
    } else {
        A.overlap = 0;      // the noise is its own launch; the advected velocity goes where the first projection had its scratch
    }
    ChainBlock B{};
    if (variant == 2) {
        // a column block of a multi-rank run, as fsb_create_block / blk_chain_block set it up (the neighbours' arrays:
        // the same stand-ins, 16 / 32 higher)
        B.G.W = width; B.G.H = height;
        B.G.pitch = ((col_end - col_begin) + 2 * halo + 3) & ~3;
        B.G.jbase = col_begin - halo; B.G.j0 = col_begin; B.G.j1 = col_end;
        B.halo = halo;
        B.strip_begin = col_begin / 32;
        B.strip_end = col_end >= height ? -1 : col_end / 32;
        B.has_left = col_begin > 0;
        B.has_right = col_end < height;
        B.ncols = col_end - col_begin;
        B.step = 1;
        auto peer = [&](ChainPeer &P, int base) {
            P.u = arr(base + 3); P.v = arr(base + 4); P.p = arr(base + 7); P.div = arr(base + 8); P.d = arr(base + 9); P.nd = arr(base + 10);
            P.extra[0] = arr(base + 11); P.extra[1] = arr(base + 12);
            P.chain_buf = reinterpret_cast<unsigned char *>(0x200000000ull + (size_t)base * 0x10000000ull);
        };
        if (B.has_left) peer(B.left, 16);
        if (B.has_right) peer(B.right, 32);
        A.blk = &B;
    }
    A.u = arr(3); A.v = arr(4);
    A.p = arr(7); A.div = arr(8);
    A.d = arr(9); A.nd = arr(10);
    ChainDesc D;
    const cudaError_t e = chain_describe(A, nsm, arr(11), arr(12), &D);
    if (e == cudaErrorNotSupported) return fail(FSB_ELIMIT, "fsb_dbg_chain_describe: no chained configuration for this grid");
    if (e != cudaSuccess) return fail_cuda(e, "fsb_dbg_chain_describe");
    constexpr int kWords = 53;
    if (D.nphases * kWords > max_words) return fail(FSB_EINVAL, "fsb_dbg_chain_describe: output too small");
    auto id = [](const void *q) { return (long long)((size_t)q / 0x1000); };
    for (int q = 0; q < D.nphases; ++q) {
        const ChainPhaseDesc &d = D.ph[q];
        long long *o = out + (size_t)q * kWords;
        const int head[8] = {d.kind, d.nitems, d.ncl, d.nf, d.cpv, d.shift, d.ndeps, d.nxdeps};
        for (int k = 0; k < 8; ++k) o[k] = head[k];
        for (int k = 0; k < 4; ++k)
            for (int m = 0; m < 4; ++m) o[8 + 4 * k + m] = d.dep[k][m];
        for (int k = 0; k < 3; ++k)
            for (int m = 0; m < 4; ++m) o[24 + 4 * k + m] = d.xdep[k][m];
        for (int k = 0; k < 8; ++k) o[36 + k] = id(d.ptr[k]);
        for (int f = 0; f < 2; ++f)
            for (int m = 0; m < 3; ++m) o[44 + 3 * f + m] = id(d.wf[f][m]);
        o[50] = d.reach_phase;
        o[51] = d.item0;
        o[52] = d.push_cols;
    }
    *nphases = D.nphases;
    *grid = D.grid;
    *cw = D.cw;
    return FSB_OK;
}

int fsb_set_chain_overlap(fsb_sim *s, int on)
{
    if (!s || on < 0 || on > 1) return fail(FSB_EINVAL, "fsb_set_chain_overlap: bad argument");
    s->chain_overlap = on;
    return FSB_OK;
}

int fsb_get_stream(fsb_sim *s, void **stream)
{
    if (!s || !stream) return fail(FSB_EINVAL, "fsb_get_stream: bad argument");
    *stream = (void *)s->stream;
    return FSB_OK;
}

int fsb_profile_enable(fsb_sim *s, int on)
{
    if (!s) return fail(FSB_EINVAL, "fsb_profile_enable: null handle");
    s->profiling = on != 0;
    return FSB_OK;
}

int fsb_chain_trace(fsb_sim *s, int which, uint64_t *out, int max_phases, int max_items, int *nphases)
{
    if (!s || !out || !nphases || max_phases < 1 || max_items < 1) return fail(FSB_EINVAL, "fsb_chain_trace: bad argument");
    DeviceGuard guard(s->device);
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    std::memset(out, 0, (size_t)max_phases * max_items * 3 * sizeof(uint64_t));
    FSB_CUDA(chain_read_trace(s->chain, which, reinterpret_cast<unsigned long long *>(out), max_phases, max_items, nphases));
    return FSB_OK;
}

int fsb_profile_read(fsb_sim *s, double *ms, uint64_t *counts, int reset)
{
    if (!s) return fail(FSB_EINVAL, "fsb_profile_read: null handle");
    DeviceGuard guard(s->device);
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    for (auto &r : s->prof_open) {
        float t = 0.0f;
        if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) {
            s->prof_ms[r.cat] += t;
            s->prof_n[r.cat] += 1;
        }
        s->prof_free.push_back(r.a);
        s->prof_free.push_back(r.b);
    }
    s->prof_open.clear();
    for (int c = 0; c < FSB_PROF_COUNT; ++c) {
        if (ms) ms[c] = s->prof_ms[c];
        if (counts) counts[c] = s->prof_n[c];
        if (reset) { s->prof_ms[c] = 0.0; s->prof_n[c] = 0; }
    }
    return FSB_OK;
}

/* ---------------------------- column blocks of a multi-rank run ---------------------------- */

int fsb_block_info(const fsb_sim *s, int *pitch, int *jbase, int *j0, int *j1)
{
    if (!s) return fail(FSB_EINVAL, "fsb_block_info: null handle");
    if (pitch) *pitch = s->G.pitch;
    if (jbase) *jbase = s->G.jbase;
    if (j0) *j0 = s->G.j0;
    if (j1) *j1 = s->G.j1;
    return FSB_OK;
}

int fsb_link_handle(fsb_sim *s, void *handle64)
{
    if (!s || !handle64 || !s->link_buf) return fail(FSB_EINVAL, "fsb_link_handle: not a block simulation");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    DeviceGuard guard(s->device);
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    FSB_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t *>(handle64), s->link_buf));
    return FSB_OK;
}

int fsb_link_connect(fsb_sim *s, const void *left_handle64, const void *right_handle64)
{
    if (!s || !s->link_buf) return fail(FSB_EINVAL, "fsb_link_connect: not a block simulation");
    DeviceGuard guard(s->device);
    if (left_handle64) {
        cudaIpcMemHandle_t h;
        std::memcpy(&h, left_handle64, sizeof(h));
        FSB_CUDA(cudaIpcOpenMemHandle(&s->peer_left, h, cudaIpcMemLazyEnablePeerAccess));
    }
    if (right_handle64) {
        cudaIpcMemHandle_t h;
        std::memcpy(&h, right_handle64, sizeof(h));
        FSB_CUDA(cudaIpcOpenMemHandle(&s->peer_right, h, cudaIpcMemLazyEnablePeerAccess));
    }
    return FSB_OK;
}

/* CUDA IPC handles of the six field buffers, in role order (6 x 64 bytes): the rank on the RIGHT maps them so that
 * its sweep kernels can store the result columns that belong to this rank straight into these arrays. */
int fsb_field_handles(fsb_sim *s, void *handles6x64)
{
    if (!s || !handles6x64 || !s->is_block) return fail(FSB_EINVAL, "fsb_field_handles: not a block simulation");
    DeviceGuard guard(s->device);
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    for (int k = 0; k < 6; ++k)
        FSB_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t *>(static_cast<unsigned char *>(handles6x64) + 64 * k), s->f[k]));
    return FSB_OK;
}

int fsb_link_connect_fields(fsb_sim *s, const void *left_handles6x64, int left_ncols)
{
    if (!s || !left_handles6x64 || !s->is_block) return fail(FSB_EINVAL, "fsb_link_connect_fields: bad argument");
    DeviceGuard guard(s->device);
    for (int k = 0; k < 6; ++k) {
        cudaIpcMemHandle_t h;
        std::memcpy(&h, static_cast<const unsigned char *>(left_handles6x64) + 64 * k, sizeof(h));
        FSB_CUDA(cudaIpcOpenMemHandle(&s->peer_left_base[k], h, cudaIpcMemLazyEnablePeerAccess));
        s->peer_left_f[k] = static_cast<float *>(s->peer_left_base[k]);
    }
    s->left_ncols = left_ncols;
    return FSB_OK;
}

/* Holds this simulation's stream until the rank on the right has stored `expected` strips' worth of result columns
 * of sweep-field slot `slot` into this rank's arrays (counted since creation). */
int fsb_blk_wait_returns(fsb_sim *s, int slot, unsigned long long expected)
{
    if (!s || !s->link_buf || slot < 0 || slot >= kMaxSweepFields) return fail(FSB_EINVAL, "fsb_blk_wait_returns: bad argument");
    DeviceGuard guard(s->device);
    const unsigned long long *c = reinterpret_cast<const unsigned long long *>(s->link_buf + kLinkRingBytes + 2 * kLinkFlagBytes) + slot;
    FSB_CUDA(launch_wait_counter(c, expected, s->stream));
    s->launches += 1;
    return FSB_OK;
}

/* ---- the whole step of a column block as ONE chained launch (chain.cu) -----------------------------------------------
 * Every rank runs its blocks of the wave; items near a rank boundary store their results into the neighbour's halo as
 * well and publish their done-words there, so a rank only ever reads its own memory and no collective, halo exchange
 * or host round trip sits inside a step.  Set-up: fsb_blk_chain_prepare on every rank, exchange the handles
 * (fsb_blk_chain_handles, fsb_field_handles), fsb_blk_chain_connect; then fsb_blk_step per step. */
static ChainArgs blk_chain_args(fsb_sim *s, float dt, const ChainBlock *B)
{
    ChainArgs A{};
    A.W = s->W;
    A.H = s->H;
    A.u = s->f[FSB_VELOCITY_X];
    A.v = s->f[FSB_VELOCITY_Y];
    A.p = s->f[FSB_PREV_VEL_X];
    A.div = s->f[FSB_PREV_VEL_Y];
    A.d = s->f[FSB_DENSITY];
    A.nd = s->f[FSB_PREV_DENSITY];
    A.gs_sweeps = s->prm.proj_iters;
    A.sor_sweeps = s->prm.diff_iters;
    A.a = dt * s->prm.viscosity * (float)s->W * (float)s->H;
    A.c_recip = 1.0f / (1.0f + 4.0f * A.a);
    A.omega = s->prm.omega;
    A.dt = dt;
    A.with_advect = 1;
    A.with_density = 1;
    A.density_pz = s->prm.diff_iters > 0 ? 1 : 0;
    A.blk = B;
    return A;
}

static void blk_chain_block(const fsb_sim *s, ChainBlock &B)
{
    B.G = s->G;
    B.halo = s->halo;
    B.strip_begin = s->strip_begin;
    B.strip_end = s->strip_end;
    B.has_left = s->peer_chain[0][0] != nullptr;
    B.has_right = s->peer_chain[1][0] != nullptr;
    B.ncols = s->G.j1 - s->G.j0;
    B.step = s->blk_steps;
    auto fill = [&](ChainPeer &P, float *const *f, void *const *c) {
        P.d = f[FSB_DENSITY]; P.u = f[FSB_VELOCITY_X]; P.v = f[FSB_VELOCITY_Y];
        P.nd = f[FSB_PREV_DENSITY]; P.p = f[FSB_PREV_VEL_X]; P.div = f[FSB_PREV_VEL_Y];
        P.chain_buf = static_cast<unsigned char *>(c[0]);
        P.extra[0] = static_cast<float *>(c[1]);
        P.extra[1] = static_cast<float *>(c[2]);
    };
    if (B.has_left) fill(B.left, s->peer_l_f, s->peer_chain[0]);
    if (B.has_right) fill(B.right, s->peer_r_f, s->peer_chain[1]);
}

/* Is there a chained configuration for this block (sweep counts that are multiples of 10 / 5, non-zero viscosity, zero
 * density diffusion, RK4, a halo at least as wide as the fused sweeps)?  Allocates the buffers the neighbours will map.
 * Returns FSB_OK, or FSB_ELIMIT when the block must use the pass-by-pass entry points instead. */
int fsb_blk_chain_prepare(fsb_sim *s, float dt)
{
    if (!s || !s->is_block) return fail(FSB_EINVAL, "fsb_blk_chain_prepare: not a block simulation");
    DeviceGuard guard(s->device);
    const float ad = dt * s->prm.diffusion * (float)s->W * (float)s->H;
    const float av = dt * s->prm.viscosity * (float)s->W * (float)s->H;
    if (s->prm.advect_scheme != FSB_ADVECT_RK4 || ad != 0.0f || av == 0.0f || s->prm.diff_iters <= 0 || s->chain_level < 1)
        return fail(FSB_ELIMIT, "fsb_blk_chain_prepare: these parameters have no chained configuration");
    ChainBlock B;
    blk_chain_block(s, B);
    B.has_left = B.has_right = false;        // sizes do not depend on the neighbours
    ChainArgs A = blk_chain_args(s, dt, &B);
    const cudaError_t e = launch_chain(A, s->chain, s->sweep, s->stream, 2);
    if (e == cudaErrorNotSupported) return fail(FSB_ELIMIT, "fsb_blk_chain_prepare: this grid / these sweep counts have no chained configuration");
    if (e != cudaSuccess) return fail_cuda(e, "fsb_blk_chain_prepare");
    return FSB_OK;
}

/* 3 x 64 bytes: CUDA IPC handles of the chain buffer and the two second-projection scratch fields. */
int fsb_blk_chain_handles(fsb_sim *s, void *handles3x64)
{
    if (!s || !handles3x64 || !s->chain) return fail(FSB_EINVAL, "fsb_blk_chain_handles: call fsb_blk_chain_prepare first");
    DeviceGuard guard(s->device);
    void *bufs[3];
    chain_shared_buffers(s->chain, bufs);
    for (int k = 0; k < 3; ++k) {
        if (!bufs[k]) return fail(FSB_EINVAL, "fsb_blk_chain_handles: buffers not allocated");
        FSB_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t *>(static_cast<unsigned char *>(handles3x64) + 64 * k), bufs[k]));
    }
    return FSB_OK;
}

/* left / right: the neighbours' fsb_blk_chain_handles (3 x 64) and fsb_field_handles (6 x 64) output, NULL where there
 * is no neighbour.  The left neighbour's fields may already be mapped by fsb_link_connect_fields. */
int fsb_blk_chain_connect(fsb_sim *s, const void *left_chain3x64, const void *left_fields6x64, const void *right_chain3x64,
                          const void *right_fields6x64)
{
    if (!s || !s->is_block || !s->chain) return fail(FSB_EINVAL, "fsb_blk_chain_connect: call fsb_blk_chain_prepare first");
    DeviceGuard guard(s->device);
    auto open = [&](const void *h64, void **out) -> int {
        cudaIpcMemHandle_t h;
        std::memcpy(&h, h64, sizeof(h));
        FSB_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
        return FSB_OK;
    };
    if (left_chain3x64) {
        for (int k = 0; k < 3; ++k) FSB_TRY(open(static_cast<const unsigned char *>(left_chain3x64) + 64 * k, &s->peer_chain[0][k]));
        if (!s->peer_left_f[0]) {
            if (!left_fields6x64) return fail(FSB_EINVAL, "fsb_blk_chain_connect: left field handles missing");
            for (int k = 0; k < 6; ++k) {
                FSB_TRY(open(static_cast<const unsigned char *>(left_fields6x64) + 64 * k, &s->peer_left_base[k]));
                s->peer_left_f[k] = static_cast<float *>(s->peer_left_base[k]);
            }
        }
        for (int k = 0; k < 6; ++k) s->peer_l_f[k] = s->peer_left_f[k];
    }
    if (right_chain3x64) {
        if (!right_fields6x64) return fail(FSB_EINVAL, "fsb_blk_chain_connect: right field handles missing");
        for (int k = 0; k < 3; ++k) FSB_TRY(open(static_cast<const unsigned char *>(right_chain3x64) + 64 * k, &s->peer_chain[1][k]));
        for (int k = 0; k < 6; ++k) {
            FSB_TRY(open(static_cast<const unsigned char *>(right_fields6x64) + 64 * k, &s->peer_r_base[k]));
            s->peer_r_f[k] = static_cast<float *>(s->peer_r_base[k]);
        }
    }
    s->blk_chain_ready = true;
    return FSB_OK;
}

/* One FluidSim::step of this rank's column block: the velocity noise, then everything else as one chained launch.
 * Every rank calls it once per step.  Asynchronous; fsb_blk_chain_error tells, after a sync, whether a back-trace
 * left the halo (the step's results are then invalid: create the blocks with a wider halo). */
int fsb_blk_step(fsb_sim *s, float dt)
{
    if (!s || !s->is_block || !s->blk_chain_ready) return fail(FSB_EINVAL, "fsb_blk_step: chain not connected");
    DeviceGuard guard(s->device);
    FSB_TRY(flush_sources(s));
    invalidate_mirror(s);
    FSB_TRY(add_noise(s));
    ChainBlock B;
    blk_chain_block(s, B);
    ChainArgs A = blk_chain_args(s, dt, &B);
    cudaError_t e;
    {
        ProfScope ps(s, FSB_PROF_CHAIN);
        e = launch_chain(A, s->chain, s->sweep, s->stream, 0);
    }
    if (e != cudaSuccess) return fail_cuda(e, "fsb_blk_step: chained launch");
    s->launches += 1;
    s->blk_steps += 1;
    // the advected fields sit in the former prev buffers: swap roles, here and in the neighbours' tables
    auto swap_roles = [](float **f) {
        std::swap(f[FSB_VELOCITY_X], f[FSB_PREV_VEL_X]);
        std::swap(f[FSB_VELOCITY_Y], f[FSB_PREV_VEL_Y]);
        std::swap(f[FSB_DENSITY], f[FSB_PREV_DENSITY]);
    };
    swap_roles(s->f);
    swap_roles(s->peer_l_f);
    swap_roles(s->peer_r_f);
    swap_roles(s->peer_left_f);
    return FSB_OK;
}

int fsb_blk_chain_error(fsb_sim *s, unsigned long long *word)
{
    if (!s || !word) return fail(FSB_EINVAL, "fsb_blk_chain_error: bad argument");
    DeviceGuard guard(s->device);
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    FSB_CUDA(chain_read_error(s->chain, word));
    return FSB_OK;
}

static SweepLink make_link(const fsb_sim *s, unsigned long long seq)
{
    SweepLink L;
    L.has_left = s->peer_left != nullptr;
    L.has_right = s->peer_right != nullptr;
    L.in_rings = reinterpret_cast<const float *>(s->link_buf);
    L.in_prod = reinterpret_cast<const unsigned long long *>(s->link_buf + kLinkRingBytes);
    L.cons = reinterpret_cast<const unsigned long long *>(s->link_buf + kLinkRingBytes + kLinkFlagBytes);
    if (s->peer_left)
        L.left_cons = reinterpret_cast<unsigned long long *>(static_cast<unsigned char *>(s->peer_left) + kLinkRingBytes + kLinkFlagBytes);
    if (s->peer_right) {
        L.right_rings = reinterpret_cast<float *>(s->peer_right);
        L.right_prod = reinterpret_cast<unsigned long long *>(static_cast<unsigned char *>(s->peer_right) + kLinkRingBytes);
    }
    if (s->peer_left && s->peer_left_f[0]) {
        L.left_ret = reinterpret_cast<unsigned long long *>(static_cast<unsigned char *>(s->peer_left) + kLinkRingBytes + 2 * kLinkFlagBytes);
        L.peer_jbase = s->G.jbase - s->left_ncols;     // same halo width and pitch on every rank
        L.peer_hi = s->G.j0 + 1;                        // the left rank's last owned columns and its first halo column
    }
    L.seq = seq;
    return L;
}

/* One launch of the sweep chain on this rank's strips.  mode 0: Gauss-Seidel of the pressure (rhs = divergence in
 * FSB_PREV_VEL_Y, result in FSB_PREV_VEL_X, `first` = start from p == 0); mode 1: SOR diffusion of `nfields` fields in
 * place.  `seq` must be the same, strictly increasing number on every rank.  The result columns that belong to the
 * left neighbour land in this rank's left halo; the caller returns them and then applies the boundaries. */
int fsb_blk_sweep(fsb_sim *s, int mode, int nfields, const int *fields, const int *bs, int levels, int stages, int first,
                  float diff, float dt, unsigned long long seq, const int *rhs_fields)
{
    if (!s || !s->is_block || nfields < 1 || nfields > kMaxSweepFields) return fail(FSB_EINVAL, "fsb_blk_sweep: bad argument");
    DeviceGuard guard(s->device);
    SweepLaunch L{};
    L.nfields = nfields;
    L.W = s->W;
    L.H = s->H;
    L.levels = levels;
    L.stages = stages;
    L.strips_per_cta = s->tune_strips > 0 ? s->tune_strips : 2;
    // diffusion runs one stage per strip: four strips per CTA where every rank's block allows it (same choice on all ranks)
    if (mode == 1 && stages == 1 && s->strip_begin % 4 == 0 && (s->G.j1 - s->G.j0) % 128 == 0) L.strips_per_cta = 4;
    L.pitch = s->G.pitch;
    L.jbase = s->G.jbase;
    L.strip_begin = s->strip_begin;
    L.strip_end = s->strip_end;
    L.link = make_link(s, seq);
    if (mode == 0) {
        L.mode = kGaussSeidel;
        L.fld[0] = SweepField{first ? nullptr : s->f[FSB_PREV_VEL_X], s->f[FSB_PREV_VEL_Y], s->f[FSB_PREV_VEL_X], 1.0f, 1.0f};
        if (L.link.left_ret) L.fld[0].out_peer = s->peer_left_f[FSB_PREV_VEL_X];
        L.a = 0.0f; L.c_recip = 1.0f; L.omega = 1.0f;
    } else {
        if (!fields || !bs) return fail(FSB_EINVAL, "fsb_blk_sweep: fields missing");
        L.mode = kSor;
        const float a = dt * diff * (float)s->W * (float)s->H;
        L.a = a;
        L.c_recip = 1.0f / (1.0f + 4.0f * a);
        L.omega = s->prm.omega;
        for (int i = 0; i < nfields; ++i) {
            if (!valid_field(fields[i])) return fail(FSB_EINVAL, "fsb_blk_sweep: bad field");
            float *x = s->f[fields[i]];
            // several passes over one diffusion: the right-hand side stays the field as it was before the first pass
            // (prev* in the reference, src/fluid_sim.cpp:386-391), kept by the caller in another buffer
            const float *rhs = x;
            if (rhs_fields) {
                if (!valid_field(rhs_fields[i]) || rhs_fields[i] == fields[i]) return fail(FSB_EINVAL, "fsb_blk_sweep: bad rhs field");
                rhs = s->f[rhs_fields[i]];
            }
            L.fld[i] = SweepField{x, rhs, x, sign_i(bs[i]), sign_j(bs[i])};
            if (L.link.left_ret) L.fld[i].out_peer = s->peer_left_f[fields[i]];
        }
    }
    int nl = 0;
    {
        ProfScope ps(s, mode == 0 ? FSB_PROF_GS : FSB_PROF_SOR);
        FSB_CUDA(launch_sweep_pass(L, s->sweep, s->stream, &nl));
    }
    s->launches += nl;
    return FSB_OK;
}

int fsb_blk_set_boundary(fsb_sim *s, int n, const int *fields, const int *bs, const int *plus_zero)
{
    if (!s || n < 1 || n > 4 || !fields || !bs) return fail(FSB_EINVAL, "fsb_blk_set_boundary: bad argument");
    DeviceGuard guard(s->device);
    BndJob jobs[4];
    for (int i = 0; i < n; ++i) {
        if (!valid_field(fields[i])) return fail(FSB_EINVAL, "fsb_blk_set_boundary: bad field");
        jobs[i] = BndJob{s->f[fields[i]], bs[i], plus_zero ? plus_zero[i] : 0};
    }
    return set_boundary(s, jobs, n);
}

/* x = x + 0.0f over the local arrays: what the reference's diffuse() leaves in the interior when a == 0
 * (src/fluid_sim.cpp:56-57); the caller applies the boundaries afterwards. */
int fsb_blk_add_zero(fsb_sim *s, int n, const int *fields)
{
    if (!s || n < 1 || !fields) return fail(FSB_EINVAL, "fsb_blk_add_zero: bad argument");
    DeviceGuard guard(s->device);
    for (int i = 0; i < n; ++i) {
        if (!valid_field(fields[i])) return fail(FSB_EINVAL, "fsb_blk_add_zero: bad field");
        FSB_CUDA(launch_add_zero(s->f[fields[i]], s->ncells, s->stream));
        s->launches += 1;
    }
    return FSB_OK;
}

int fsb_blk_divergence(fsb_sim *s)
{
    if (!s) return fail(FSB_EINVAL, "null handle");
    DeviceGuard guard(s->device);
    ProfScope ps(s, FSB_PROF_DIVERGENCE);
    FSB_CUDA(launch_divergence(s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y], s->f[FSB_PREV_VEL_Y], s->G, s->stream));
    s->launches += 1;
    return FSB_OK;
}

int fsb_blk_gradient(fsb_sim *s)
{
    if (!s) return fail(FSB_EINVAL, "null handle");
    DeviceGuard guard(s->device);
    if (!s->vmax_word) {
        FSB_CUDA(cudaMalloc(&s->vmax_word, 256));
        FSB_CUDA(cudaMemsetAsync(s->vmax_word, 0, 256, s->stream));
    }
    ProfScope ps(s, FSB_PROF_GRADIENT);
    FSB_CUDA(launch_gradient(s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y], s->f[FSB_PREV_VEL_X], s->G, s->stream, s->vmax_word));
    s->launches += 1;
    return FSB_OK;
}

/* Device address of one float: the largest |velocityY| the fsb_blk_gradient calls wrote into this rank's owned columns
 * since the caller last cleared it (it only grows).  velocityY moves a back-trace along the columns
 * (src/fluid_sim.cpp:275-291), so this bounds the halo width the next advection needs. */
int fsb_blk_vmax_word(fsb_sim *s, void **dptr)
{
    if (!s || !dptr) return fail(FSB_EINVAL, "fsb_blk_vmax_word: bad argument");
    DeviceGuard guard(s->device);
    if (!s->vmax_word) {
        FSB_CUDA(cudaMalloc(&s->vmax_word, 256));
        FSB_CUDA(cudaMemsetAsync(s->vmax_word, 0, 256, s->stream));
    }
    *dptr = s->vmax_word;
    return FSB_OK;
}

/* RK4 advection of both velocity components (halo columns must hold the neighbours' values out to the trace
 * radius); the fields swap roles with the prev* buffers, boundaries are left to the caller. */
int fsb_blk_advect_velocity(fsb_sim *s, float dt, int extend_halo)
{
    if (!s) return fail(FSB_EINVAL, "null handle");
    DeviceGuard guard(s->device);
    GridView G = s->G;
    if (extend_halo) {      // one more column towards every neighbouring rank (the grid's own walls stay excluded)
        if (G.j0 > 0 && G.j0 - 1 >= G.jbase) G.j0 -= 1;
        if (G.j1 < G.H && G.j1 + 1 <= G.jbase + G.pitch) G.j1 += 1;
    }
    {
        ProfScope ps(s, FSB_PROF_ADVECT_VELOCITY);
        FSB_CUDA(launch_advect_rk4(s->f[FSB_PREV_VEL_X], s->f[FSB_VELOCITY_X], s->f[FSB_PREV_VEL_Y], s->f[FSB_VELOCITY_Y],
                                   s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y], 0, G, dt, s->stream));
    }
    s->launches += 1;
    std::swap(s->f[FSB_VELOCITY_X], s->f[FSB_PREV_VEL_X]);
    std::swap(s->f[FSB_VELOCITY_Y], s->f[FSB_PREV_VEL_Y]);
    std::swap(s->peer_left_f[FSB_VELOCITY_X], s->peer_left_f[FSB_PREV_VEL_X]);   // every rank swaps the same roles
    std::swap(s->peer_left_f[FSB_VELOCITY_Y], s->peer_left_f[FSB_PREV_VEL_Y]);
    return FSB_OK;
}

int fsb_blk_advect_density(fsb_sim *s, float dt, int plus_zero)
{
    if (!s) return fail(FSB_EINVAL, "null handle");
    DeviceGuard guard(s->device);
    {
        ProfScope ps(s, FSB_PROF_ADVECT_SCALAR);
        FSB_CUDA(launch_advect_rk4(s->f[FSB_PREV_DENSITY], s->f[FSB_DENSITY], nullptr, nullptr, s->f[FSB_VELOCITY_X],
                                   s->f[FSB_VELOCITY_Y], plus_zero, s->G, dt, s->stream));
    }
    s->launches += 1;
    std::swap(s->f[FSB_DENSITY], s->f[FSB_PREV_DENSITY]);
    std::swap(s->peer_left_f[FSB_DENSITY], s->peer_left_f[FSB_PREV_DENSITY]);
    return FSB_OK;
}

/* ---------------------------- headless density view ---------------------------- */

static int render_density(fsb_sim *s, int px_w, int px_h)
{
    if (px_w < 1 || px_h < 1 || (size_t)px_w * px_h > ((size_t)1 << 28)) return fail(FSB_EINVAL, "fsb_view_density: bad image size");
    if (s->is_block) return fail(FSB_EINVAL, "fsb_view_density: not available on a column block (gather the density first)");
    FSB_TRY(flush_sources(s));
    const size_t need = (size_t)px_w * px_h * 4;
    if (need > s->view_bytes) {
        FSB_CUDA(cudaStreamSynchronize(s->stream));
        if (s->view_buf) cudaFree(s->view_buf);
        s->view_buf = nullptr;
        s->view_bytes = 0;
        FSB_CUDA(cudaMalloc(&s->view_buf, need));
        s->view_bytes = need;
    }
    FSB_CUDA(launch_view_density(s->f[FSB_DENSITY], s->G, px_w, px_h, s->view_buf, s->stream));
    s->launches += 1;
    return FSB_OK;
}

/* What the reference draws with the velocity vectors switched on (src/main.cpp:452-514): the density image, then one
 * line per sampled cell. */
int fsb_view_velocity(fsb_sim *s, int px_w, int px_h, unsigned char *host_rgba)
{
    if (!s || !host_rgba) return fail(FSB_EINVAL, "fsb_view_velocity: bad argument");
    DeviceGuard guard(s->device);
    FSB_TRY(render_density(s, px_w, px_h));
    {
        const size_t need = view_glyph_scratch_bytes(s->G);
        if (need > s->glyph_bytes) {
            FSB_CUDA(cudaStreamSynchronize(s->stream));
            if (s->glyph_buf) cudaFree(s->glyph_buf);
            s->glyph_buf = nullptr;
            s->glyph_bytes = 0;
            FSB_CUDA(cudaMalloc(&s->glyph_buf, need));
            s->glyph_bytes = need;
        }
    }
    FSB_CUDA(launch_view_glyphs(s->f[FSB_VELOCITY_X], s->f[FSB_VELOCITY_Y], s->G, px_w, px_h, s->view_buf, s->glyph_buf, s->stream));
    s->launches += 2;
    FSB_CUDA(cudaMemcpyAsync(host_rgba, s->view_buf, (size_t)px_w * px_h * 4, cudaMemcpyDeviceToHost, s->stream));
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    return FSB_OK;
}

int fsb_view_density_device(fsb_sim *s, int px_w, int px_h, void **d_rgba)
{
    if (!s || !d_rgba) return fail(FSB_EINVAL, "fsb_view_density_device: bad argument");
    DeviceGuard guard(s->device);
    FSB_TRY(render_density(s, px_w, px_h));
    *d_rgba = s->view_buf;
    return FSB_OK;
}

int fsb_view_density(fsb_sim *s, int px_w, int px_h, unsigned char *host_rgba)
{
    if (!s || !host_rgba) return fail(FSB_EINVAL, "fsb_view_density: bad argument");
    DeviceGuard guard(s->device);
    FSB_TRY(render_density(s, px_w, px_h));
    FSB_CUDA(cudaMemcpyAsync(host_rgba, s->view_buf, (size_t)px_w * px_h * 4, cudaMemcpyDeviceToHost, s->stream));
    FSB_CUDA(cudaStreamSynchronize(s->stream));
    return FSB_OK;
}

int fsb_pressure_sweeps_noalloc(const float *d_divergence, float *d_pressure, int width, int height, int sweeps,
                                void *stream)
{
    std::string msg;
    const int rc = pressure_sweeps_noalloc(d_divergence, d_pressure, width, height, sweeps, (cudaStream_t)stream, msg);
    if (rc == 0) return FSB_OK;
    g_last_error = msg;
    if (rc == 1) return FSB_ELIMIT;
    if (rc == 2) return FSB_EINVAL;
    return FSB_ECUDA;
}

void *fsb_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
        g_last_error = "fsb_host_alloc: cudaHostAlloc failed";
        return nullptr;
    }
    return p;
}

void fsb_host_free(void *p)
{
    if (p) cudaFreeHost(p);
}

const char *fsb_last_error(void) { return g_last_error.c_str(); }
const char *fsb_version(void) { return "fluid_sim_b200 0.1 (sm_100a)"; }

}  // extern "C"
