// Internal declarations shared by the CUDA translation units of libfluidsim_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <string>
#include <vector>

#include "glibc_rand.h"

namespace fsb {

// A rank's view of the global W x H grid: all rows of the global columns [j0, j1) ("owned"), stored with
// `pitch` floats per row, local column 0 holding global column `jbase` (jbase <= j0: halo columns left and
// right of the owned block).  Single GPU: pitch = H, jbase = 0, j0 = 0, j1 = H.
struct GridView {
    int W, H;
    int pitch;
    int jbase;
    int j0, j1;
};
inline GridView whole_grid(int W, int H) { return GridView{W, H, H, 0, 0, H}; }

// ------------------------------------------------------------------------------------
// Strip-pipelined exact-order sweeps (wavefront.cu)
// ------------------------------------------------------------------------------------
enum SweepMode { kGaussSeidel = 0, kSor = 1 };

constexpr int kMaxSweepFields = 3;   // fields relaxed by one launch (independent chains)
constexpr int kWfBatch = 16;         // time steps between inter-CTA hand-offs
constexpr int kWfSlots = 8;          // depth (in batches) of the global boundary ring
constexpr int kWfMaxCtas = 1024;     // CTAs per chain the flag array is sized for
constexpr int kWfMaxLevels = 20;     // largest sweeps-per-warp instantiation

struct SweepField {
    const float *x0;    // initial iterate, nullptr = all zeros
    const float *rhs;   // divergence (GS) or source field (SOR)
    float *out;         // result (interior cells only; may alias x0)
    float si, sj;       // sign applied on the i-walls / j-walls (setBoundary's b)
    float *out_peer = nullptr;   // multi-rank: the same field of the rank on the LEFT (peer memory), or nullptr
};

// Hand-over of the strip chain between two ranks that own adjacent column blocks (peer memory over
// NVLink).  All pointers are device addresses valid on THIS rank; "peer" ones map the neighbour's memory.
// Layout per field f and stage g: rings[(f * kLinkStages + g) * kLinkRingFloats], flags[(f * kLinkStages + g)].
constexpr int kLinkStages = 16;
// 44 = floats per lane-step at 20 sweeps per warp; x2: every float travels with a 4-byte tag (flag-in-data)
constexpr int kLinkRingFloats = kWfSlots * kWfBatch * 44 * 2;
struct SweepLink {
    bool has_left = false, has_right = false;
    const float *in_rings = nullptr;                 // local: written by the left rank (tagged 16-byte words)
    const unsigned long long *in_prod = nullptr;     // unused since the rings carry their own tags
    unsigned long long *left_cons = nullptr;         // peer (left rank): batches this rank has consumed
    float *right_rings = nullptr;                    // peer (right rank): its in_rings
    unsigned long long *right_prod = nullptr;        // peer (right rank): its in_prod
    const unsigned long long *cons = nullptr;        // local: batches the right rank has consumed
    unsigned long long seq = 0;                      // pass number, identical on all ranks
    // result columns that belong to the left rank (its last owned columns and its first halo column) are
    // stored straight into its arrays; every strip that wrote some adds 1 to left_ret[field] when it is done
    unsigned long long *left_ret = nullptr;          // peer (left rank): completion counters
    int peer_jbase = 0;                              // global column of the left rank's local column 0
    int peer_hi = 0;                                 // columns < peer_hi go to the left rank as well
};

struct SweepLaunch {
    SweepField fld[kMaxSweepFields];
    int nfields;
    int W, H;           // global grid
    int mode;           // SweepMode
    int levels;         // sweeps per warp (1, 2, 5, 10 or 20)
    int stages;         // warps chained per strip; the launch applies levels * stages sweeps
    float a, c_recip, omega;   // SOR coefficients (src/fluid_sim.cpp:45-47)
    int strips_per_cta; // 0 = choose
    // column block of a multi-rank run (defaults: the whole grid on one GPU)
    int pitch = 0, jbase = 0;            // local row pitch / global column of local column 0 (pitch 0: = H)
    int strip_begin = 0, strip_end = -1; // global 32-column strips this rank runs (-1: to the end)
    SweepLink link;
};

struct SweepScratch {
    float *xchg = nullptr;                 // global boundary rings
    size_t xchg_floats = 0;
    unsigned long long *flags = nullptr;   // producer / consumer progress words
    unsigned long long seq = 0;            // launch sequence number (flag epoch)
};

size_t sweep_flag_words();
int sweep_max_stages(int levels);   // 0 if `levels` is not an instantiated size
bool sweep_config_fits(const SweepLaunch &L);   // can (levels, stages) be launched with all CTAs co-resident?

cudaError_t launch_sweep_pass(const SweepLaunch &L, SweepScratch &scr, cudaStream_t st, int *launches);

// Passes of FluidSim::step chained inside one cooperative launch (chain.cu): diffuse(u), diffuse(v) when
// sor_sweeps > 0, project(u, v, p, div); with_advect: advect(u), advect(v) into (p, div) and the second projection on
// them; with_density: densityStep (DIFFUSION == 0) with the final velocity, d -> nd.
// cudaErrorNotSupported: no chained configuration for this grid / these sweep counts.
// layout of a link buffer (the rings, flags and counters through which neighbouring ranks hand the strip chain over)
constexpr size_t kLinkRingBytes = (size_t)kMaxSweepFields * kLinkStages * kLinkRingFloats * sizeof(float);
constexpr size_t kLinkFlagBytes = (size_t)kMaxSweepFields * kLinkStages * sizeof(unsigned long long);
constexpr size_t kLinkRetBytes = 64;   // completion counters of the columns stored here by the right rank
constexpr size_t kLinkBytes = kLinkRingBytes + 2 * kLinkFlagBytes + kLinkRetBytes;

struct ChainState;
// One neighbouring rank of a column block, as seen from this GPU (every pointer is a peer mapping).
struct ChainPeer {
    float *u = nullptr, *v = nullptr, *p = nullptr, *div = nullptr, *d = nullptr, *nd = nullptr;
    float *extra[2] = {nullptr, nullptr};      // its second-projection scratch
    unsigned char *chain_buf = nullptr;        // its ChainState buffer (done words, maxima, step words, link regions)
};
struct ChainBlock {
    GridView G;
    int halo;
    int strip_begin, strip_end;                // 32-column strips of the sweep chain this rank runs (-1: to the end)
    bool has_left, has_right;
    ChainPeer left, right;
    int ncols;                                 // columns every rank owns
    unsigned long long step;                   // number of this step, identical on all ranks
};
constexpr int kChainSources = 256;
// The phase table of a chained launch as plain data (chain_describe): what tests/test_chain_deps.py checks the
// dependency lists against.  dep / xdep rows: {phase, from_out, lo, hi} (see ChainDep in chain.cu).
struct ChainPhaseDesc {
    int kind, nitems, ncl, nf, cpv, shift, ndeps, nxdeps;
    int dep[4][4], xdep[3][4];
    const void *ptr[8];        // u, v, q, dst_a, dst_b, src_a, src_b, bnd_extra
    const void *wf[2][3];      // sweep phases: {initial iterate, right-hand side, result} of fields 0 and 1
    int reach_phase;           // advection: the gradient phase whose blocks it waits for
    int item0, push_cols;      // column block of a multi-rank run: global index of the rank's first item; see ChainPhase
};
struct ChainDesc {
    int nphases, grid, cw;
    ChainPhaseDesc ph[24];
};
struct ChainArgs {
    int W, H;
    float *u, *v, *p, *div;
    int sor_sweeps;            // 0: no diffusion phase
    float a, c_recip, omega;   // SOR coefficients (src/fluid_sim.cpp:45-47)
    int gs_sweeps;
    int with_advect, with_density;
    float dt;
    float *d, *nd;             // density and the buffer that receives the advected density
    int density_pz;            // diffusion sweeps ran with a == 0: boundaries of d + folded (v + 0.0f) taps
    // The whole step including its noise, in buffers that let consecutive steps overlap (one GPU, see launch_chain):
    // vin_u / vin_v: the velocity the step starts from; the noise phase writes vin + noise into u, v (nullptr: the caller
    // has added the noise to u, v itself).  yu / yv: receive the advected velocity, which the second projection then
    // works on (nullptr: p, div are reused for that, as the reference reuses its prev arrays).
    const float *vin_u = nullptr, *vin_v = nullptr;
    float *yu = nullptr, *yv = nullptr;
    const uint32_t *noise_base = nullptr;      // window of the rand() stream at the step's first output (31 words)
    int noise_zero = 0;                        // FSB_NOISE_ZERO
    int overlap = 0;                           // the next step's launch may start before this one has drained
    // pending addDensity / addVelocity calls, sorted by (field, cell) in call order: applied by the noise phase instead
    // of a separate copy + kernel between two launches (which would keep them from overlapping); at most kChainSources
    const struct SourceEntry *sources = nullptr;
    int nsources = 0;
    float *snap = nullptr;                     // a second copy of the advected density (for an asynchronous read)
    ChainDesc *describe = nullptr;             // mode 3: fill this in instead of launching (no device is touched) ...
    int describe_nsm = 0;                      // ... planning for a GPU of this many SMs
    const ChainBlock *blk = nullptr;   // column block of a multi-rank run (nullptr: the whole grid on this GPU)
};
// mode 0: launch; 1: only tell whether a configuration exists; 2: also allocate the scratch (so that its IPC handles can
// be exchanged before the first launch)
cudaError_t launch_chain(const ChainArgs &A, ChainState *&cs, SweepScratch &scr, cudaStream_t st, int mode);
// the buffers neighbouring ranks map: [0] ChainState buffer, [1], [2] second-projection scratch (after mode 2)
void chain_shared_buffers(ChainState *cs, void *out[3]);
// error word of the chained launches (0: none; bit 63: an advection trace left the columns this rank can see)
cudaError_t chain_read_error(ChainState *cs, unsigned long long *word);
void chain_state_destroy(ChainState *cs);
cudaError_t chain_read_trace(ChainState *cs, int which, unsigned long long *out, int max_phases, int max_items, int *nphases);
// the phase table of a launch as data, no device involved (see ChainDesc)
cudaError_t chain_describe(const ChainArgs &A, int nsm, float *extra0, float *extra1, ChainDesc *out);

// Allocation-free pressure sweeps on caller-owned device buffers (wavefront.cu); see fsb_pressure_sweeps_noalloc.
int pressure_sweeps_noalloc(const float *d_div, float *d_p, int W, int H, int sweeps, cudaStream_t st, std::string &msg);

// ------------------------------------------------------------------------------------
// Pointwise / gather kernels (pointwise.cu)
// ------------------------------------------------------------------------------------
struct BndJob { float *x; int b; int plus_zero; };   // plus_zero: interior values are read as (v + 0.0f)

cudaError_t launch_set_boundary(const BndJob *jobs, int njobs, const GridView &G, cudaStream_t st);
cudaError_t launch_divergence(const float *u, const float *v, float *div, const GridView &G, cudaStream_t st);
// vmax_bits (optional): atomicMax target for the largest |v| written (bit pattern of a non-negative float)
cudaError_t launch_gradient(float *u, float *v, const float *p, const GridView &G, cudaStream_t st,
                            unsigned *vmax_bits = nullptr);
// RK4 back-trace of (src_a[, src_b]) through (u, v); dst_b/src_b may be null.
cudaError_t launch_advect_rk4(float *dst_a, const float *src_a, float *dst_b, const float *src_b,
                              const float *u, const float *v, int src_plus_zero, const GridView &G, float dt,
                              cudaStream_t st);
cudaError_t launch_advect_sl(float *dst, const float *src, const float *u, const float *v, int src_plus_zero,
                             const GridView &G, float dt, float sign, cudaStream_t st);
cudaError_t launch_maccormack_combine(float *dst, const float *src, const float *fwd, const float *back,
                                      int src_plus_zero, const GridView &G, cudaStream_t st);
cudaError_t launch_add_zero(float *x, size_t n, cudaStream_t st);
cudaError_t launch_axpy_field(float *x, const float *y, size_t n, cudaStream_t st);
cudaError_t launch_wait_counter(const unsigned long long *counter, unsigned long long expected, cudaStream_t st);

struct SourceEntry { int cell; int field; float amount; };   // field: 0 d, 1 u, 2 v
cudaError_t launch_apply_sources(float *d, float *u, float *v, const SourceEntry *entries, const int *seg_start,
                                 int nseg, cudaStream_t st);

// ------------------------------------------------------------------------------------
// Headless density view (view.cu): the reference viewer's quads + density shader per pixel, RGBA8, row 0 = bottom
// ------------------------------------------------------------------------------------
cudaError_t launch_view_density(const float *d, const GridView &G, int px_w, int px_h, unsigned char *rgba,
                                cudaStream_t st);
// velocity glyph lines (every second cell, |v| >= 0.1) blended over the image already in `rgba`
// scratch: view_glyph_scratch_bytes(G) bytes of device memory (one 32-byte record per sampled cell)
size_t view_glyph_scratch_bytes(const GridView &G);
cudaError_t launch_view_glyphs(const float *u, const float *v, const GridView &G, int px_w, int px_h, unsigned char *rgba,
                               void *scratch, cudaStream_t st);

// ------------------------------------------------------------------------------------
// Noise (noise.cu): u, v += ((float)rand()/RAND_MAX - 0.5f) * 1e-4f over the full grid
// ------------------------------------------------------------------------------------
constexpr int kNoiseCellsPerLane = 256;                          // contiguous cells generated by one lane
constexpr int kNoiseCellsPerWarp = 32 * kNoiseCellsPerLane;
constexpr int kNoiseMaxLevels = 24;                              // warp-index bits covered by jump matrices

struct NoiseTables {
    uint32_t *dev = nullptr;        // [1 + kNoiseMaxLevels] transposed 31x31 matrices: lane jump, then warp jumps 2^b
    uint32_t *dev_pow2 = nullptr;   // column-block runs: x^(2^b), b = 0..39
};

struct NoiseWindow { uint32_t w[kRandDeg]; };

cudaError_t noise_tables_create(NoiseTables &t);
void noise_tables_destroy(NoiseTables &t);
cudaError_t launch_noise(float *u, float *v, size_t ncells, const NoiseWindow &base, const NoiseTables &t,
                         int zero_mode, cudaStream_t st);
cudaError_t noise_tables_create_pow2(NoiseTables &t, int H);
cudaError_t launch_noise_block(float *u, float *v, const GridView &G, const NoiseWindow &base, const NoiseTables &t,
                               int zero_mode, cudaStream_t st);

}  // namespace fsb
