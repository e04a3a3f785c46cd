/* lbm_b200.h - C ABI of the B200-native D2Q9-BGK timestep engine (liblbm_b200.so).
 *
 * This is the drop-in boundary for the reference's per-timestep hot path.  The
 * reference (georgeherbert/lattice-boltzmann, d2q9-bgk.c) has no FFI: its boundary
 * is the four calls main() makes around the time loop.  Each entry point below
 * names the reference code it replaces (file:line relative to the reference tree).
 *
 * Conventions
 *   - plain C, opaque handle, plain pointers and sizes; no CUDA/torch types.
 *   - every int-returning call returns 0 on success, non-zero on failure; the
 *     message is then available from lbm_last_error() (thread-local).  The host
 *     driver turns a failure into the reference's die() (d2q9-bgk.c:754-759).
 *   - the caller owns all host buffers; the library owns all device memory.
 *   - one host thread drives one context.  Calls are asynchronous where noted.
 *   - there is NO CPU fallback: without a usable CUDA device lbm_create fails.
 *
 * Lattice layout on the host side of the ABI: nine planes ("speeds") of ny*nx
 * floats, row-major with x fastest, direction numbering 0=rest 1=E 2=N 3=W 4=S
 * 5=NE 6=NW 7=SW 8=SE (the reference's t_speed, d2q9-bgk.c:7-13,90-100).
 */
#ifndef LBM_B200_H
#define LBM_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define LBM_NSPEEDS 9
#define LBM_ABI_VERSION 1
#define LBM_MAX_GPUS 16
#define LBM_IPC_BLOB_BYTES 256

typedef struct lbm_ctx lbm_ctx;

/* The hot-path fields of the reference's t_param (d2q9-bgk.c:66-87). */
typedef struct lbm_params {
    int nx;             /* cells in x */
    int ny;             /* cells in y */
    int max_iters;      /* capacity of the per-step av_vels record (maxIters) */
    float density;      /* density per link */
    float accel;        /* density redistribution */
    float omega;        /* relaxation parameter */
    int obstacle_lines; /* denominator rule for av_vels: the reference divides by
                           nx*ny - <number of LINES in the obstacle file>, duplicates
                           included (d2q9-bgk.c:626,640,642).  Pass that line count to
                           reproduce it; pass a negative value to use the true number
                           of blocked cells instead. */
} lbm_params;

/* Engine options (all optional; zero-initialise for defaults). */
typedef struct lbm_options {
    int strict_fp;      /* 0 = default (strict); 1 = strict: IEEE fp32 in the reference's
                           evaluation order with no FMA contraction, so the lattice is
                           bit-identical to a plain `gcc -O3` build of the reference;
                           2 = fast: FMA contraction allowed (within check.py's tolerance) */
    int ctas_per_sm;    /* 0 = auto */
    int threads_per_cta;/* 0 = auto (multiple of 32) */
    int engine;         /* 0 = auto (4 or 3 where they apply, else 1); 1 = persistent neighbour-synchronised
                           kernel; 2 = one launch per timestep (CUDA-graph replayed, single slab); 3 = dataflow
                           over tagged records: one resident warp per 128-cell task (or per 2-3 of them), slabs
                           of at most ~7000 tasks (... 512x512, 1024x768); 4 = dataflow with one resident CTA per
                           ROW, one cell per thread: nx % 32 == 0, nx <= 1024, all rows of a slab resident (up to
                           1184 rows of 128 cells, 592 of 256, 148 of 1024); auto takes it for rows of up to 256
                           cells (the check grids 128x128, 128x256, 256x256).  Engines 3 and 4 also run row slabs
                           on several GPUs (records of the edge rows are stored straight into the ring
                           neighbour's record lattice). */
    int first_device;   /* first CUDA device ordinal used (single-process multi-GPU uses
                           first_device .. first_device+ngpus-1) */
} lbm_options;

/* ---- life cycle ------------------------------------------------------------ */

/* Replaces the allocation / initialisation half of initialise() (d2q9-bgk.c:531-604,
 * 646-651) and allocate_rows() (:483-503): splits the ny rows into `ngpus` contiguous
 * slabs by the reference's rule, allocates the A/B lattices with one halo row either
 * side, uploads the uniform initial state (w0=rho*4/9, w1=rho/9, w2=rho/36) and the
 * obstacle mask (int[ny*nx], 0/1, as filled at :639) packed to one bit per cell, and
 * links neighbouring slabs over NVLink peer access.  ngpus <= 0 means all visible
 * devices (capped at ny).  opts may be NULL. */
int lbm_create(lbm_ctx** out, const lbm_params* params, const int* obstacles, int ngpus,
               const lbm_options* opts);

/* Frees everything (finalise(), d2q9-bgk.c:656-673).  NULL is allowed.  Waits for every slab's enqueued
 * work first, so it is safe straight after lbm_run.  One-process-per-GPU mode: see lbm_rank_disconnect. */
void lbm_destroy(lbm_ctx* ctx);

/* Device buffers of destroyed contexts are kept per process (by device and exact size, up to LBM_CACHE_MB
 * megabytes, default 3072, 0 = off) for the next lbm_create of the same shape: cudaMalloc / cudaFree of lattice-sized
 * buffers cost tens of milliseconds, more than a short run.  lbm_trim_cache returns them to the driver. */
int lbm_trim_cache(void);

/* ---- the hot path ------------------------------------------------------------ */

/* Replaces `iters` trips of the time loop body (d2q9-bgk.c:191-197):
 *   accelerate_flow (:235-266) + timestep (:268-339: propagate, rebound, collision,
 *   av_velocity) + pointer swap (:194-196) + halo_exchange (:341-393),
 * recording the un-normalised per-step sum of |u| for steps
 * [steps_done, steps_done+iters) on the device.  Asynchronous: enqueues the work and
 * returns; there is no host synchronisation and no cross-GPU collective per step.
 * Fails if steps_done + iters > max_iters. */
int lbm_run(lbm_ctx* ctx, int iters);

/* Blocks until all enqueued work is complete (the device sync before comp_toc,
 * d2q9-bgk.c:205-207).  Reports asynchronous kernel errors. */
int lbm_sync(lbm_ctx* ctx);

/* ---- results (collate, d2q9-bgk.c:395-447) ------------------------------------ */

/* av_vels[t] for t in [0, n): sum over slabs of the per-step partials (the one
 * MPI_Reduce, :445) times 1/free_cells (:446, denominator rule of params.obstacle_lines).
 * n <= steps_done.  Synchronises. */
int lbm_get_av_vels(lbm_ctx* ctx, float* out, int n);

/* The un-normalised per-step sums of |u| over fluid cells (what timestep() returns and the
 * reference's -DDEBUG build prints as "av velocity", d2q9-bgk.c:193,200) for steps
 * [first, first + n), summed over this context's slabs.  first + n <= steps_done.  Synchronises. */
int lbm_get_tot_u(lbm_ctx* ctx, float* out, int first, int n);

/* Gathers every slab's rows into nine global planes of ny*nx floats (:397-443; the
 * result is what the reference holds in cells_global).  Synchronises. */
int lbm_get_cells(lbm_ctx* ctx, float* const speeds[LBM_NSPEEDS]);

/* Scatters nine global planes into the slabs (incl. halo rows).  The reference has no
 * such call (its state always starts uniform); tests and checkpoint-style callers use it.
 * Does not change steps_done. */
int lbm_set_cells(lbm_ctx* ctx, const float* const speeds[LBM_NSPEEDS]);

/* Macroscopic fields of write_values (d2q9-bgk.c:710-741), computed on the device:
 * u_x, u_y, |u|, pressure per cell (ny*nx floats each; any pointer may be NULL).
 * Obstacle cells give u = 0, pressure = density/3. */
int lbm_get_final_state(lbm_ctx* ctx, float* ux, float* uy, float* u, float* pressure);

/* av_velocity() of the current state (d2q9-bgk.c:449-481): mean |u| over cells whose
 * obstacle flag is 0 (true count).  calc_reynolds (:675-678) is this times
 * reynolds_dim / viscosity, done by the host driver. */
int lbm_av_velocity(lbm_ctx* ctx, float* out);

/* total_density() (d2q9-bgk.c:680-696), the DEBUG conservation invariant. */
int lbm_total_density(lbm_ctx* ctx, float* out);

/* lbm_av_velocity and lbm_total_density in one call that ALLOCATES NOTHING (no cudaMalloc / cudaFree, no allocating
 * library routine): for callers that cannot allocate - code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it.  This entry point has no counterpart in the reference.
 * Its scratch (a pair of doubles per lattice row and the two totals) is a file-scope device array of fixed
 * size, LBM_NOALLOC_MAX_ROWS rows per GPU = 256 KB, which every GPU that loads the library holds until the process
 * exits; a context with more rows than that on one GPU is rejected with an error (use the two allocating calls, or
 * more GPUs).  *av_velocity = mean |u| over unblocked cells (un-normalised sum in one-process-per-GPU mode, as
 * lbm_av_velocity), *total_density = sum of all populations; either pointer may be NULL.  The sums are folded in a fixed
 * order (deterministic).  Synchronises.  All calls on a GPU share the one array: calls that overlap on the same GPU
 * (two contexts driven from two host threads) race - serialise them. */
#define LBM_NOALLOC_MAX_ROWS 16384
int lbm_diagnostics_noalloc(lbm_ctx* ctx, float* av_velocity, float* total_density);

/* ---- introspection ------------------------------------------------------------ */

int lbm_abi_version(void);
/* 1 if this library was built with -DLBM_CHECKED (`make checked`): every lattice / record / halo-slot access of the
 * timestep kernels is range- and alignment-checked against the allocations registered for the device, buffers start
 * poisoned with NaN bytes, and lbm_sync fails on a violation.  The stand-in for compute-sanitizer's memcheck and
 * initcheck, which this GPU pool does not allow.  Select it with LBM_LIBRARY=<path> in the Python mirror. */
int lbm_checked_build(void);
const char* lbm_last_error(void);
/* number of usable CUDA devices (0 if none / no driver). */
int lbm_device_count(void);
int lbm_steps_done(const lbm_ctx* ctx);
int lbm_num_slabs(const lbm_ctx* ctx);
/* slab r covers global rows [*row_start, *row_start + *rows) (allocate_rows, :484-496) */
int lbm_slab_rows(const lbm_ctx* ctx, int r, int* row_start, int* rows);
/* Device time of the most recent lbm_run, measured with CUDA events on the streams the
 * kernels were launched on; max over slabs.  Valid after lbm_sync. */
int lbm_last_run_ms(lbm_ctx* ctx, float* ms);
/* Number of kernel launches issued by the most recent lbm_run (all slabs). */
int lbm_last_run_launches(const lbm_ctx* ctx);
/* Bytes copied host->device and device->host through this context so far (mask and schedule
 * uploads, lattice / av_vels / final-state downloads). */
int lbm_transfer_bytes(const lbm_ctx* ctx, long long* h2d, long long* d2h);
/* Human-readable description of the engine configuration chosen (static buffer in ctx). */
const char* lbm_describe(lbm_ctx* ctx);

/* Diagnostics of the persistent engine (off unless the environment has LBM_PROFILE=1 when the
 * context is created): per CTA and warp, the SM cycles of the most recent lbm_run spent waiting
 * for neighbouring CTAs (progress flags), waiting at the CTA's end-of-step barrier, pulling
 * tagged halo slots of a neighbouring GPU (load + spin), and in total.  out receives num_ctas * warps * 4 values ([cta][warp][4]);
 * out may be NULL to query the sizes. */
int lbm_get_profile(lbm_ctx* ctx, int slab, long long* out, int capacity, int* num_ctas, int* warps);

/* The reference's row decomposition as a pure function (allocate_rows, :483-503):
 * rows/start of rank r of `size`, and which rank owns global row ny-2.  Host-only. */
int lbm_allocate_rows(int ny, int size, int rank, int* row_start, int* rows);

/* Host-only self check of the decomposition planner (slabs -> warp-tasks -> per-CTA schedules
 * and neighbour dependencies) for an nx x ny lattice on `size` GPUs with `ctas` persistent CTAs
 * each: every task scheduled exactly once, interior tasks have no foreign neighbours, every
 * neighbour owner is a dependency, mirror-flag indices agree between ring neighbours.
 * Needs no GPU.  Returns 0 if consistent. */
int lbm_plan_selftest(int nx, int ny, int size, int ctas);
/* Host set-up code (threaded obstacle count and bit-pack) against plain loops on a hashed obstacle
 * array, for every slab of a `size`-way decomposition.  No device needed. */
int lbm_host_selftest(int nx, int ny, int size, long long seed);

/* Device self check of the kernel's exact division / square-root fast paths: evaluates them on
 * 2*n hashed operand pairs (whole accepted range, and LBM-like magnitudes) and counts results that
 * differ in any bit from the IEEE-rounded intrinsics.  *mismatches must come back 0. */
int lbm_selftest_math(long long n, long long seed, long long* mismatches);

/* Device stress test of the tagged 16-byte records that carry the halos between GPUs (and the whole lattice
 * of engines 3 and 4): a kernel on `writer_device` keeps rewriting 16 K records living on `reader_device`
 * with {t,t,t,t}, t = 1, 2, ... through the product's store (one scalar st.relaxed.sys.b128) while a kernel on
 * `reader_device` reads them `sweeps` x 4 times each through the product's read path (mode 0: ld.relaxed.sys.b128; mode 1: the
 * 16-byte cp.async into shared memory used by the halo staging).  *torn = records seen with mixed words (must be
 * 0), *changes = times a record changed between two consecutive looks (> 0 proves the kernels overlapped).
 * writer_device == reader_device runs both kernels side by side on one GPU. */
int lbm_selftest_records(int writer_device, int reader_device, int sweeps, int mode, long long* torn, long long* changes);

/* ---- one process per GPU (torchrun-style launch) -------------------------------
 * Rank r of `size` owns slab r on CUDA device `device`.  After creation every rank
 * exports an opaque blob (the CUDA IPC handle of its inbound halo slots), the
 * host program exchanges blobs with the ring neighbours by whatever transport it has
 * (torch.distributed, MPI, files), and lbm_rank_connect maps them.  lbm_run/lbm_sync/
 * lbm_get_* then act on the local slab only: lbm_get_av_vels returns this rank's
 * UN-summed, un-normalised partials (to be summed once across ranks, :445) and
 * lbm_get_cells returns the local rows [row_start, row_start+rows). */
int lbm_create_rank(lbm_ctx** out, const lbm_params* params, const int* obstacles,
                    int rank, int size, int device, const lbm_options* opts);
int lbm_rank_export(lbm_ctx* ctx, void* blob /* LBM_IPC_BLOB_BYTES */);
int lbm_rank_connect(lbm_ctx* ctx, const void* blob_up, const void* blob_down);
/* Teardown in this mode: every rank calls lbm_rank_disconnect (waits for the local kernel, forgets the neighbours'
 * pointers) and then lbm_destroy; no barrier between the ranks is needed:
 *  - no kernel can still be writing into a slab whose own kernel has finished: it has consumed every halo record a
 *    neighbour will send it (the last timestep of a run pushes nothing);
 *  - buffers whose IPC handle was exported are never returned to the driver while the process lives (they wait in the
 *    buffer cache for the next context of the same shape), and the importers keep their mappings open in a per-process
 *    cache, so a neighbour's mapping never outlives the memory behind it.
 * lbm_trim_cache is the exception: in this mode call it on all ranks only after all ranks have destroyed their
 * contexts (a host barrier), because it closes the mappings and frees the exported buffers. */
int lbm_rank_disconnect(lbm_ctx* ctx);
/* 1/free_cells by the rule of params.obstacle_lines (for the caller's final scaling). */
float lbm_inv_free_cells(const lbm_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* LBM_B200_H */
