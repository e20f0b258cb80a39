/*
 * fluid_sim_b200 -- C ABI of the B200-native FluidSim::step.
 *
 * This is the drop-in boundary underneath the reference's C++ class
 * (reference: include/fluid_sim.h:17-37).  The reference has no FFI layer of its
 * own; each entry point below names the class member (or private method) of the
 * reference that it replaces.  Plain pointers and sizes only: no C++ or torch types.
 *
 * Conventions
 *   - every function returns FSB_OK (0) or a negative FSB_E* code; the message of
 *     the last failure on the calling thread is available from fsb_last_error().
 *   - fields are float32, row-major [width][height] with the y index contiguous,
 *     exactly like the reference's float f[GRID_SIZE_X][GRID_SIZE_Y]
 *     (reference: include/fluid_sim.h:42-47); linear index = x * height + y.
 *   - a simulation is bound to one CUDA device and one stream and is not
 *     thread-safe, like the reference object (single-threaded main loop,
 *     reference: src/main.cpp:369-519).
 *   - there is no CPU fallback: without a CUDA device fsb_create fails.
 */
#ifndef FLUID_SIM_B200_H
#define FLUID_SIM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fsb_sim fsb_sim;

enum {
    FSB_OK = 0,
    FSB_EINVAL = -1,   /* bad argument */
    FSB_ECUDA = -2,    /* CUDA runtime / launch failure */
    FSB_ENOMEM = -3,   /* host or device allocation failed */
    FSB_ELIMIT = -4    /* size beyond what this build supports */
};

/* field ids (the reference's member arrays, include/fluid_sim.h:42-47) */
enum {
    FSB_DENSITY = 0,      /* density       */
    FSB_VELOCITY_X = 1,   /* velocityX     */
    FSB_VELOCITY_Y = 2,   /* velocityY     */
    FSB_PREV_DENSITY = 3, /* prevDensity   (scratch) */
    FSB_PREV_VEL_X = 4,   /* prevVelocityX (scratch; pressure during project) */
    FSB_PREV_VEL_Y = 5    /* prevVelocityY (scratch; divergence during project) */
};

/* advection schemes (reference: src/fluid_sim.cpp:212-228; RK4 is the live call :219) */
enum { FSB_ADVECT_RK4 = 0, FSB_ADVECT_MACCORMACK = 1, FSB_ADVECT_SEMI_LAGRANGIAN = 2 };

/* velocity-noise modes (reference: src/fluid_sim.cpp:374-383) */
enum {
    FSB_NOISE_GLIBC = 0,  /* glibc rand() stream, addressed by position (bit-exact with the reference) */
    FSB_NOISE_ZERO = 1    /* rand() == 0x3FFFFFFF for every call: the noise term is exactly 0 */
};

typedef struct fsb_params {
    int width;             /* GRID_SIZE_X, include/fluid_sim.h:11 (default 200) */
    int height;            /* GRID_SIZE_Y, include/fluid_sim.h:12 (default 200) */
    int proj_iters;        /* Gauss-Seidel sweeps per projection, src/fluid_sim.cpp:319 (20) */
    int diff_iters;        /* SOR sweeps per diffusion, src/fluid_sim.cpp:50 (5) */
    float viscosity;       /* VISCOSITY, include/fluid_sim.h:13 (1e-4) */
    float diffusion;       /* DIFFUSION, include/fluid_sim.h:14 (0) */
    float omega;           /* SOR relaxation factor, src/fluid_sim.cpp:47 (1.5) */
    int advect_scheme;     /* FSB_ADVECT_*, src/fluid_sim.cpp:219-227 (RK4) */
    int noise_mode;        /* FSB_NOISE_* */
    uint32_t noise_seed;   /* srand() seed of the emulated stream (1 = glibc default) */
    uint64_t noise_pos;    /* number of rand() outputs already consumed by the process */
    int device;            /* CUDA device ordinal, -1 = current device */
    int noise_process_stream; /* 1: draw the noise from ONE stream shared by every simulation of the process that sets
                              * this flag, like the reference, whose rand() is process-global (src/fluid_sim.cpp:378-379:
                              * two FluidSim objects stepped alternately consume alternate stretches of it, a copy goes
                              * on where the original stopped).  0 (default): the simulation owns its position. */
    int reserved[6];
} fsb_params;

/* Fill *p with the reference's literals (200x200, 20/5 sweeps, ...). */
void fsb_default_params(fsb_params *p);

/* Replaces FluidSim::FluidSim() (src/fluid_sim.cpp:4-21): zero-filled fields on the device. */
int fsb_create(const fsb_params *p, fsb_sim **out);
int fsb_destroy(fsb_sim *s);

/* Replaces `sim = FluidSim();` (src/main.cpp:289): zero all fields; like the
 * reference, the position in the rand() stream is NOT rewound. */
int fsb_reset(fsb_sim *s);

/* Replaces FluidSim::step (src/fluid_sim.cpp:23-28).  Asynchronous: returns once the
 * work is queued on the simulation's stream; readers synchronise. */
int fsb_step(fsb_sim *s, float dt);
int fsb_sync(fsb_sim *s);

/* Replace FluidSim::addDensity / addVelocity (src/fluid_sim.cpp:424-439): out-of-range
 * coordinates are silently ignored, amounts are += in call order. */
int fsb_add_density(fsb_sim *s, int x, int y, float amount);
int fsb_add_velocity(fsb_sim *s, int x, int y, float ax, float ay);
/* Batched forms of the same (one call per mouse event / injection pattern). */
int fsb_add_density_n(fsb_sim *s, int n, const int *xs, const int *ys, const float *amounts);
int fsb_add_velocity_n(fsb_sim *s, int n, const int *xs, const int *ys, const float *axs, const float *ays);

/* Replace FluidSim::getDensity / getVelocity (src/fluid_sim.cpp:441-457): out of range
 * yields 0.  Served from a host mirror that is refreshed lazily after a step. */
int fsb_get_density(fsb_sim *s, int x, int y, float *out);
int fsb_get_velocity(fsb_sim *s, int x, int y, float *out_xy);
/* Replace getNormalizedVelocity / getVelocityMagnitude / getVelocityColor
 * (src/fluid_sim.cpp:459-519). */
int fsb_get_normalized_velocity(fsb_sim *s, int x, int y, float *out_xy);
int fsb_get_velocity_magnitude(fsb_sim *s, int x, int y, float *out);
int fsb_get_velocity_color(fsb_sim *s, int x, int y, float *out_rgb);
/* Replace getWidth / getHeight (include/fluid_sim.h:36-37). */
int fsb_get_width(const fsb_sim *s);
int fsb_get_height(const fsb_sim *s);

/* Whole-field transfers with HOST buffers of width*height floats (the bulk form of the
 * per-cell getters the reference's render loop issues, src/main.cpp:388-396). */
int fsb_read_field(fsb_sim *s, int field, float *host_dst);
/* The same copy without stalling the frame loop (the reference's loop reads every cell after every step,
 * src/main.cpp:388-396): fsb_read_field_async queues the device-to-host copy of the field as it is after all work
 * queued so far on a second stream and returns; later fsb_step calls overlap with it.  The copy reads a device-side
 * snapshot, never the live field: taken on the simulation's stream at the first such read; once a caller reads the
 * density after its steps, written by the steps' own launches (two more buffers of the grid's size), so that nothing
 * but an event sits between two consecutive steps and they keep overlapping (fsb_set_chain_overlap).
 * fsb_read_wait blocks until the data is in host_dst.  One outstanding read per simulation; host_dst should be
 * page-locked (fsb_host_alloc) or the copy will not overlap.  Not available on column blocks. */
int fsb_read_field_async(fsb_sim *s, int field, float *host_dst);
int fsb_read_wait(fsb_sim *s);
int fsb_write_field(fsb_sim *s, int field, const float *host_src);
/* density += host_src (one addDensity per cell); velocity likewise */
int fsb_add_field(fsb_sim *s, int field, const float *host_src);

/* Page-locked host buffers for the whole-field transfers above (optional; any host pointer works,
 * pinned ones copy at full PCIe rate).  No counterpart in the reference. */
void *fsb_host_alloc(size_t bytes);
void fsb_host_free(void *p);

/* Device pointer of a field (valid until the next step swaps buffers); for callers
 * that keep their data on the GPU. */
int fsb_device_field(fsb_sim *s, int field, void **dptr);

/* Position bookkeeping of the emulated rand() stream (2*width*height outputs per step). */
int fsb_get_noise_pos(const fsb_sim *s, uint64_t *pos);
int fsb_set_noise(fsb_sim *s, int mode, uint32_t seed, uint64_t pos);

/* The reference's private passes, exposed for per-kernel parity tests.  `b` is the
 * boundary kind of setBoundary (0 scalar, 1 x-velocity, 2 y-velocity). */
int fsb_op_add_noise(fsb_sim *s);                                          /* src/fluid_sim.cpp:374-383 */
int fsb_op_set_boundary(fsb_sim *s, int b, int field);                     /* :347-367 */
int fsb_op_diffuse(fsb_sim *s, int b, int field, float diff, float dt);    /* :43-62 (src == field's content on entry) */
int fsb_op_project(fsb_sim *s);                                            /* :301-344 on velocityX/velocityY */
int fsb_op_advect(fsb_sim *s, int b, int dst, int src, int u, int v, float dt); /* :212-298 */
int fsb_op_velocity_step(fsb_sim *s, float dt);                            /* :370-406 */
int fsb_op_density_step(fsb_sim *s, float dt);                             /* :409-422 */

/* The CUDA stream (cudaStream_t) all work of `s` is queued on, for callers that time or order
 * their own work against it. */
int fsb_get_stream(fsb_sim *s, void **stream);

/* Optional per-kernel-class timing with CUDA events on that stream.  fsb_profile_read synchronises,
 * then fills ms[FSB_PROF_COUNT] (accumulated milliseconds) and counts[FSB_PROF_COUNT] (number of
 * bracketed launches); reset != 0 clears the accumulators. */
enum {
    FSB_PROF_NOISE = 0,            /* velocity noise                      src/fluid_sim.cpp:374-383 */
    FSB_PROF_SOR = 1,              /* diffusion sweeps (one fused pass)   :50-61   */
    FSB_PROF_GS = 2,               /* projection sweeps (one fused pass)  :319-331 */
    FSB_PROF_DIVERGENCE = 3,       /* :307-314 */
    FSB_PROF_GRADIENT = 4,         /* :334-341 */
    FSB_PROF_ADVECT_VELOCITY = 5,  /* both velocity components, :401-402  */
    FSB_PROF_ADVECT_SCALAR = 6,    /* one field, :421                     */
    FSB_PROF_BOUNDARY = 7,         /* setBoundary, :347-367               */
    FSB_PROF_CHAIN = 8,            /* diffuse + project (or project alone) chained inside one launch: the sweep passes,
                                    * divergence, gradient and their boundaries of :390-394 / :405 */
    FSB_PROF_COUNT = 9
};
int fsb_profile_enable(fsb_sim *s, int on);
int fsb_profile_read(fsb_sim *s, double *ms, uint64_t *counts, int reset);
/* Diagnostic with no counterpart in the reference: timeline of the most recent chained launch of one kind (which = 0: the kernel that runs
 * diffuse + project, 1: project alone, of velocityStep, src/fluid_sim.cpp:390-394 / :405, as phases of one launch).
 * out[(phase * max_items + item) * 3 + k], k = 0 wait begins, 1 work begins, 2 results published, in nanoseconds of
 * the GPU's global timer; items are blocks of columns, left to right.  Synchronises the simulation's stream. */
int fsb_chain_trace(fsb_sim *s, int which, uint64_t *out, int max_phases, int max_items, int *nphases);

/* ---- column blocks: one simulation per GPU, each owning all rows of a range of columns --------------------
 * No counterpart in the reference (it is single-process); this is the sharded form of the same passes.
 * A block stores global columns [col_begin - halo, col_end + halo) with `pitch` floats per row
 * (fsb_block_info); fsb_read_field / fsb_write_field move the whole local array (width * pitch floats).
 * The exact-order sweeps run as ONE chain across all ranks: neighbouring blocks hand the strip boundary over
 * through a small buffer mapped with CUDA IPC (fsb_link_handle / fsb_link_connect), inside the kernels.
 * Halo columns are exchanged by the caller between passes (fluid_sim_b200/multi.py does it with NCCL). */
int fsb_create_block(const fsb_params *p, int col_begin, int col_end, int halo, fsb_sim **out);
int fsb_block_info(const fsb_sim *s, int *pitch, int *jbase, int *j0, int *j1);
int fsb_link_handle(fsb_sim *s, void *handle64);                                   /* 64-byte cudaIpcMemHandle_t */
int fsb_link_connect(fsb_sim *s, const void *left_handle64, const void *right_handle64);   /* NULL = no neighbour */
int fsb_field_handles(fsb_sim *s, void *handles6x64);                             /* 6 IPC handles, role order */
int fsb_link_connect_fields(fsb_sim *s, const void *left_handles6x64, int left_ncols);
int fsb_blk_wait_returns(fsb_sim *s, int slot, unsigned long long expected);
/* rhs_fields (mode 1, may be NULL): the buffers holding the fields as they were before the first pass of a diffusion
 * that takes several passes; NULL = the field itself (single pass, in place). */
int fsb_blk_sweep(fsb_sim *s, int mode, int nfields, const int *fields, const int *bs, int levels, int stages,
                  int first, float diff, float dt, unsigned long long seq, const int *rhs_fields);
int fsb_blk_add_zero(fsb_sim *s, int n, const int *fields);
int fsb_blk_set_boundary(fsb_sim *s, int n, const int *fields, const int *bs, const int *plus_zero);
int fsb_blk_divergence(fsb_sim *s);
int fsb_blk_gradient(fsb_sim *s);
/* The whole step of a column block as one chained launch (no halo exchange, collective or host round trip inside a
 * step: blocks near a rank boundary store their results into the neighbour's halo through peer mappings and publish
 * their progress words there).  Set-up on every rank: fsb_blk_chain_prepare (FSB_ELIMIT: these parameters need the
 * pass-by-pass entry points above), exchange fsb_blk_chain_handles (3 x 64 bytes) and fsb_field_handles with both
 * neighbours, fsb_blk_chain_connect; then fsb_blk_step once per step on every rank.  fsb_blk_chain_error (synchronises):
 * non-zero if a back-trace left the halo (bit 63) or a wait timed out. */
int fsb_blk_chain_prepare(fsb_sim *s, float dt);
int fsb_blk_chain_handles(fsb_sim *s, void *handles3x64);
int fsb_blk_chain_connect(fsb_sim *s, const void *left_chain3x64, const void *left_fields6x64, const void *right_chain3x64,
                          const void *right_fields6x64);
int fsb_blk_step(fsb_sim *s, float dt);
int fsb_blk_chain_error(fsb_sim *s, unsigned long long *word);
/* One float in device memory: largest |velocityY| written by fsb_blk_gradient into the owned columns since the caller
 * cleared it; bounds the columns the next advection's back-traces can move (src/fluid_sim.cpp:264, :275). */
int fsb_blk_vmax_word(fsb_sim *s, void **dptr);
/* extend_halo: also advect the first halo column on every side that has a neighbour (the divergence that follows needs
 * it; computing it here saves a halo exchange -- the halo must then hold the neighbours' values one column further). */
int fsb_blk_advect_velocity(fsb_sim *s, float dt, int extend_halo);
int fsb_blk_advect_density(fsb_sim *s, float dt, int plus_zero);

/* No counterpart in the reference.
 *
 * The pressure relaxation of project() (src/fluid_sim.cpp:313-331: p = 0, then `sweeps` lexicographic
 * Gauss-Seidel sweeps, each followed by setBoundary(0, p)) on buffers the CALLER owns, without a single
 * allocation.  It exists for hosts that cannot let this library allocate: an engine that reserved the
 * whole GPU for its own pools at start-up (a later cudaMalloc simply fails there), or a frame / control
 * loop that must never run into the device-wide synchronisation and unbounded latency that cudaMalloc and
 * cudaFree imply.  fsb_create / fsb_step size their inter-CTA boundary rings with cudaMalloc on first use;
 * this entry point keeps the same rings and progress words in one fixed 16.25 MiB array that is part of
 * the library image (every GPU that loads the library holds it until the process exits) and calls no
 * cudaMalloc, cudaFree or any other allocating routine.
 *
 *   d_divergence  device pointer, width*height floats (only interior cells are read)
 *   d_pressure    device pointer, width*height floats, fully overwritten; must not alias d_divergence
 *   stream        cudaStream_t of the caller (NULL = default stream); the work is queued, not waited for
 *
 * The result is bit-identical to what fsb_op_project leaves in FSB_PREV_VEL_X for the same divergence.
 * Limits and rules:
 *   - the rings must fit the fixed array: about 8000 columns at 40 sweeps; larger requests return FSB_ELIMIT
 *     with a message and launch nothing;
 *   - the first call on a device looks up the array's device address (once per device, with that device
 *     current): make that first call outside any stream capture;
 *   - "part of the library image" does not mean "resident from dlopen": CUDA loads a module lazily by
 *     default, so the array, and the code of each sweep-kernel variant, reach the GPU the first time they are
 *     used.  The first call on a device (the array: 16.25 MiB) and the first call with a sweep count that
 *     needs a kernel variant not used before (a few hundred KB of code) therefore make the DRIVER set device
 *     memory aside and may synchronise the device, although this library still calls no allocator.  A host
 *     that hands all GPU memory to its own pools must make one call per sweep count it will use BEFORE it
 *     reserves that memory, and a latency-critical loop must make those calls before it starts; or run the
 *     process with CUDA_MODULE_LOADING=EAGER, which moves all of it to library load.  After that warm-up a
 *     call changes nothing in cudaMemGetInfo (tests/test_zz_no_alloc.py checks this);
 *   - every call on a device uses the same array: calls whose GPU work overlaps (different streams, or a
 *     second thread) race with each other - serialise them; calls on one stream are safe back to back;
 *   - the array keeps the previous call's contents.  Every call clears the progress words; a call whose
 *     (width, height, sweeps) differ from the previous call's on that device also clears the ring region
 *     (the rings of another shape sit elsewhere in the array, and their tags repeat every 16384 launches).
 *     Both are memsets on `stream`, not allocations. */
int fsb_pressure_sweeps_noalloc(const float *d_divergence, float *d_pressure, int width, int height, int sweeps,
                                void *stream);

/* Headless density view: what the reference viewer draws for the density field
 * (src/main.cpp:388-427 quads, :32-41 fragment shader, :376-377 blending, :345 clear colour), evaluated per pixel
 * by a CUDA kernel with no graphics API.  px_w x px_h RGBA8 pixels, 4 bytes each, row 0 = bottom window row (GL
 * convention); the reference's window is 800 x 600 (src/main.cpp:9-10).  Pixel = cleared black (0,0,0,255) unless
 * its centre lies in the quad of a cell with density >= 0.01: then rgb = (d, 0, 1-d) * a, a = min(2d, 0.8),
 * alpha = a*a + (1-a), everything clamped to [0,1] and rounded to 8 bits.
 *   fsb_view_density         renders and copies the image to host_rgba (px_w*px_h*4 bytes); synchronises.
 *   fsb_view_density_device  renders into a library-owned device buffer (valid until the next view call or
 *                            fsb_destroy) and returns its address; queued on the simulation's stream, not waited for.
 * Not available on column blocks (gather the density first). */
int fsb_view_density(fsb_sim *s, int px_w, int px_h, unsigned char *host_rgba);
int fsb_view_density_device(fsb_sim *s, int px_w, int px_h, void **d_rgba);
/* The viewer's image with the velocity vectors switched on (src/main.cpp:452-514, toggle :17): the density image,
 * then for every second cell in x and y with |v| >= 0.1 a line from the cell centre, 0.05 NDC long along the
 * normalised velocity (getNormalizedVelocity), in getVelocityColor's ramp colour with alpha 0.8, line width 2.
 * GL leaves the exact pixels of such a line to the implementation; this library's rule: a pixel belongs to a line
 * when its centre lies in [start, end) along the major axis and within [-1, 1) pixels of it along the minor axis. */
int fsb_view_velocity(fsb_sim *s, int px_w, int px_h, unsigned char *host_rgba);

/* Number of kernels this library has launched on behalf of `s` so far. */
int fsb_launch_count(const fsb_sim *s, uint64_t *count);

/* Tuning knobs of the relaxation pipeline, for experiments (0 = library default): column strips per
 * CTA, sweeps kept in flight per warp (1, 2, 5, 10 or 20), cap on warps chained per strip. */
int fsb_set_tuning(fsb_sim *s, int strips_per_cta, int sweeps_per_warp, int max_stages);
/* Tuning with no counterpart in the reference: how many passes of a step share one kernel launch.  0: every pass is its
 * own launch; 1: diffuse + project (and the second project) as chained launches; 2 (default): the whole step after the
 * noise as one chained launch on the grid shapes where that was measured faster; 3: wherever a chained configuration
 * exists.  Results are bit-identical at every level.  The environment variable FSB_CHAIN sets the initial level. */
int fsb_set_chain_level(fsb_sim *s, int level);
/* Also without counterpart in the reference: at chain level >= 2, with DIFFUSION == 0 and the RK4 advection, fsb_step is
 * ONE launch that contains the noise as well, and consecutive calls overlap on the GPU: the launch of step n+1 starts on
 * the SMs the first thread blocks of step n have left while the last ones are still at work (the launches order
 * themselves block by block through words in device memory).  on = 0: the noise is its own launch and every step's launch
 * starts after the one before has drained (round-2 structure before this).  Results are bit-identical either way; the
 * fields rotate through seven more buffers of the grid's size.  Used only while the simulation is the only one of its
 * device (in this process) that has it enabled: overlapping launches are not cooperative launches, and two simulations
 * stepping concurrently could each hold part of the SMs while waiting for the rest.  Default 1; the environment variable
 * FSB_CHAIN_OVERLAP sets the initial value. */
int fsb_set_chain_overlap(fsb_sim *s, int on);

/* Diagnostics, no counterpart in the reference and no device needed: the phase table of the one-launch step
 * (fsb_set_chain_overlap; variant 0, variant 3: carrying pending sources) or of the one launch after the noise (variant 1: the advected velocity is written
 * where the first projection had its pressure and divergence) for a grid and sweep counts on a GPU of `nsm` SMs, as 53
 * integers per phase (variant 2: the same launch on the column block [col_begin, col_end) with `halo` stored columns either
 * side, as a rank of a multi-GPU run builds it; items are numbered along the whole grid): kind, items,
 * column ranges, fields per range, columns per range, result shift, number of dependencies inside the launch / on the
 * previous launch, 4 x {phase, from_out, lo, hi}, 3 x the same for the previous launch, the arrays it touches as small
 * integers (u, v, q, dst_a, dst_b, src_a, src_b, bnd_extra; per sweep field: iterate, right-hand side, result; 1, 2 =
 * the velocity the step starts from, 3, 4 / 5, 6 = the two velocity sets it writes, 7, 8 / 11, 12 = pressure and
 * divergence of the two projections, 9 / 10 = density in / out), the gradient phase an advection phase waits for, the global
 * index of the rank's first item and the columns from a rank boundary within which results are also stored in the neighbour.
 * tests/test_chain_deps.py checks on the CPU that these dependencies order every pair of conflicting accesses. */
int fsb_dbg_chain_describe(int variant, int width, int height, int diff_iters, int proj_iters, int nsm, int col_begin,
                           int col_end, int halo, long long *out, int max_words, int *nphases, int *grid, int *cw);

const char *fsb_last_error(void);
const char *fsb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* FLUID_SIM_B200_H */
