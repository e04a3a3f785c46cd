// lbm_kernels.cuh - the warp-task update and the three timestep engines built on it (1: persistent,
// progress flags; 2: one launch per step; 3: dataflow over tagged records for small lattices).
#pragma once
#include "lbm_device.cuh"

namespace lbm {

// Geometry of one warp-task for one lane.  All lattice offsets are 32-bit ELEMENT offsets from
// the buffer base (lbm_create guarantees 9*ps < 2^32), so an address costs one IMAD.WIDE.
struct TaskGeom {
    int lr, x0, xc;
    unsigned rc, rs, rn;     // element offsets of (row, xc), (south row, xc), (north row, xc) in plane 0
    unsigned wc, ws, wn;     // same rows at the west neighbour column of this lane's first cell
    unsigned ec, es, en;     // same rows at the east neighbour column of this lane's last cell
    bool active, need_w, need_e;
};

// RING: the slab has ring neighbours on other GPUs (rows beyond its edges come from halo slots);
// otherwise it is the whole lattice and y wraps inside it.
template <bool RING>
__device__ __forceinline__ TaskGeom task_geom(const SlabDev& d, int2 td, int lane)
{
    TaskGeom g;
    g.lr = td.x;
    g.x0 = td.y * 128 + lane * 4;
    g.active = g.x0 < d.nx;
    g.xc = g.active ? g.x0 : 0;            // inactive lanes (ragged last chunk) shadow x=0, never store
    const int h = g.lr + 1;                // row index in the haloed layout
    int hs = h - 1, hn = h + 1;
    if (!RING) {                           // single slab: periodic in y inside the slab
        if (g.lr == 0) hs = d.rows;
        if (g.lr == d.rows - 1) hn = 1;
    }
    g.need_w = (lane == 0);
    g.need_e = (lane == 31) || (g.xc + 4 >= d.nx);
    const unsigned nx = (unsigned)d.nx;
    const unsigned xw = (g.xc == 0) ? nx - 1u : (unsigned)g.xc - 1u;          // d2q9-bgk.c:289
    const unsigned xe = (g.xc + 4 >= d.nx) ? 0u : (unsigned)g.xc + 4u;        // d2q9-bgk.c:288
    const unsigned bc = (unsigned)h * nx, bs = (unsigned)hs * nx, bn = (unsigned)hn * nx;
    g.rc = bc + g.xc; g.rs = bs + g.xc; g.rn = bn + g.xc;
    g.wc = bc + xw;   g.ws = bs + xw;   g.wn = bn + xw;
    g.ec = bc + xe;   g.es = bs + xe;   g.en = bn + xe;
    return g;
}

__device__ __forceinline__ unsigned ld_mask(const unsigned char* p) { return chk_ptr(p, 1) ? (unsigned)__ldg(p) : 0u; }

// SM cycle counter that the compiler may not move across barriers or memory operations (profiling)
__device__ __forceinline__ long long clock_fenced()
{
    long long t;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t)::"memory");
    return t;
}

constexpr long long kSpinLimitCycles = 20000000000LL;  // ~10 s: a neighbour (CTA, GPU or process) that never arrives

// First record of (parity, side) in an inbound halo-slot array (see "halo slots" in lbm_device.cuh):
// 4 record planes of nx/4 records each.
__device__ __forceinline__ int ll_slot(const SlabDev& d, int par, int side) { return (par * 2 + side) * d.nx; }

// The record addresses one edge task reads on side SIDE: its own group in the 4 planes, the group
// holding the west element, the group holding the east element.
template <int SIDE>
__device__ __forceinline__ void ll_addresses(const SlabDev& d, const uint4* base, const TaskGeom& g, const uint4*& pq,
                                             const uint4*& pw, const uint4*& pe)
{
    const int Q = d.nx >> 2;
    const unsigned row0 = g.rc - (unsigned)g.xc;
    pq = base + (g.xc >> 2);
    pw = base + (SIDE ? 3 : 2) * Q + ((g.wc - row0) >> 2);
    pe = base + (SIDE ? 1 : 2) * Q + ((g.ec - row0) >> 2);
}

// Pulls the three crossing populations of one edge row for this lane's 4 cells (plus the single
// element west / east of the warp's range for the lanes that need it) out of the local inbound
// slots, spinning until every record carries `tag`.  SIDE 0: row below (f2,f5,f6: west element is
// f5 = b3 of the group before, east element f6 = c0 of the group after); SIDE 1: row above
// (f4,f7,f8: west element f8 = c3, east element f7 = b0).  Warp-uniform: all 32 lanes call it.
// Only taken when the staged copy (ll_take) was not there or not complete yet, so it is kept OUT OF
// LINE: the persistent kernel's steady-state code must stay inside the 32 KB instruction cache.
struct HaloRow { float4 v0, v1, v2; float w, e; };

template <int SIDE>
__device__ __noinline__ HaloRow ll_pull_slow(const uint4* pq, const uint4* pw, const uint4* pe, int Q, bool need_w, bool need_e,
                                             unsigned tag, int* status)
{
    HaloRow h;
    h.w = 0.f; h.e = 0.f;
    const long long t0 = clock64();
    for (;;) {
        const uint4 r0 = ld_record(pq), r1 = ld_record(pq + Q), r2 = ld_record(pq + 2 * Q), r3 = ld_record(pq + 3 * Q);
        bool ok = r0.w == tag && r1.w == tag && r2.w == tag && r3.w == tag;
        h.v0 = make_float4(__uint_as_float(r0.x), __uint_as_float(r0.y), __uint_as_float(r0.z), __uint_as_float(r1.x));
        h.v1 = make_float4(__uint_as_float(r1.y), __uint_as_float(r1.z), __uint_as_float(r2.x), __uint_as_float(r2.y));
        h.v2 = make_float4(__uint_as_float(r2.z), __uint_as_float(r3.x), __uint_as_float(r3.y), __uint_as_float(r3.z));
        if (need_w) { const uint4 q = ld_record(pw); ok = ok && q.w == tag; h.w = __uint_as_float(SIDE ? q.z : q.y); }
        if (need_e) { const uint4 q = ld_record(pe); ok = ok && q.w == tag; h.e = __uint_as_float(SIDE ? q.y : q.z); }
        if (__all_sync(0xffffffffu, ok)) break;
        // give up warp-uniformly (another CTA failed, or the neighbour never arrives)
        const bool late = clock64() - t0 > kSpinLimitCycles;
        if (late) atomicExch(status, 2);
        if (__any_sync(0xffffffffu, late || ld_status(status) != 0)) break;
        __nanosleep(40);
    }
    return h;
}

template <int SIDE>
__device__ __forceinline__ void ll_pull(const SlabDev& d, const uint4* base, const TaskGeom& g, unsigned tag,
                                        float4 (&v)[3], float& w, float& e)
{
    const uint4 *pq, *pw, *pe;
    ll_addresses<SIDE>(d, base, g, pq, pw, pe);
    const HaloRow h = ll_pull_slow<SIDE>(pq, pw, pe, d.nx >> 2, g.need_w, g.need_e, tag, d.status);
    v[0] = h.v0; v[1] = h.v1; v[2] = h.v2;
    if (g.need_w) w = h.w;
    if (g.need_e) e = h.e;
}

// Starts the asynchronous fetch of one edge task's halo records into a shared-memory stage buffer
// (each lane fetches exactly the records it will read itself).  The caller commits the group.
__device__ __forceinline__ void ll_prefetch(const SlabDev& d, int par, int2 td, int lane, uint4* stage)
{
    const TaskGeom g = task_geom<true>(d, td, lane);
    const uint4 *pq, *pw, *pe;
    if (g.lr == 0) ll_addresses<0>(d, d.ll_in + ll_slot(d, par, 0), g, pq, pw, pe);
    else ll_addresses<1>(d, d.ll_in + ll_slot(d, par, 1), g, pq, pw, pe);
    const int Q = d.nx >> 2;
#ifdef LBM_STAGE_BULK
    // lane 0: the four record planes of the task (contiguous: lane l's records sit l records after lane 0's) and the
    // west record; the lane holding the task's last cell: the east record.  All counted on the buffer's mbarrier.
    void* bar = stage + kStageRecords;
    const int lanes = min(32, (d.nx - td.y * 128) >> 2);           // a ragged last chunk has fewer active lanes
    const bool east = g.need_e && g.active;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the buffer was last read through the generic proxy
    if (lane == 0) {
        stage_bar_expect(bar, 4u * 16u * (unsigned)lanes + 16u + 16u);
#pragma unroll
        for (int r = 0; r < 4; ++r) bulk_copy(stage + r * 32, pq + r * Q, 16u * (unsigned)lanes, bar);
        bulk_copy(stage + 128, pw, 16u, bar);
    }
    if (east) bulk_copy(stage + 129, pe, 16u, bar);
#else
#pragma unroll
    for (int r = 0; r < 4; ++r) cp_async_record(stage + r * 32 + lane, pq + r * Q);
    if (g.need_w) cp_async_record(stage + 128, pw);
    if (g.need_e && g.active) cp_async_record(stage + 129, pe);
#endif
}

// Takes a staged edge task out of shared memory; false if any tag is not `tag` yet.
template <int SIDE>
__device__ __forceinline__ bool ll_take(const uint4* stage, const TaskGeom& g, int lane, unsigned tag, float4 (&v)[3], float& w, float& e)
{
    const uint4 r0 = stage[lane], r1 = stage[32 + lane], r2 = stage[64 + lane], r3 = stage[96 + lane];
    bool ok = r0.w == tag && r1.w == tag && r2.w == tag && r3.w == tag;
    v[0] = make_float4(__uint_as_float(r0.x), __uint_as_float(r0.y), __uint_as_float(r0.z), __uint_as_float(r1.x));
    v[1] = make_float4(__uint_as_float(r1.y), __uint_as_float(r1.z), __uint_as_float(r2.x), __uint_as_float(r2.y));
    v[2] = make_float4(__uint_as_float(r2.z), __uint_as_float(r3.x), __uint_as_float(r3.y), __uint_as_float(r3.z));
    if (g.need_w) { const uint4 q = stage[128]; ok = ok && q.w == tag; w = __uint_as_float(SIDE ? q.z : q.y); }
    if (g.need_e && g.active) { const uint4 q = stage[129]; ok = ok && q.w == tag; e = __uint_as_float(SIDE ? q.y : q.z); }
    return __all_sync(0xffffffffu, ok);
}

// Updates the 4 cells of this lane for one task: pull (propagate), rebound / collision,
// accelerate_flow epilogue for the NEXT step on row ny-2 and halo push to the ring neighbours (both only
// when a next step follows in this run: `epilogue_accel`), store, and returns the lane's sum of |u| over its fluid cells.  On a slab edge row the row
// beyond the edge is pulled from the inbound halo slots, which must carry `tag`; pushes carry tag + 1.
template <bool S, bool RING>
__device__ __forceinline__ float process_task(const SlabDev& d, int par, int2 td, int lane, bool epilogue_accel, unsigned tag,
                                            const uint4* stage = nullptr, long long* ll_cycles = nullptr, unsigned stage_parity = 0)
{
    const TaskGeom g = task_geom<RING>(d, td, lane);
    const float* __restrict__ src = d.buf[par];
    float* __restrict__ dst = d.buf[par ^ 1];
    const unsigned ps = (unsigned)d.ps;
    // slab edge rows take the row beyond the edge from the inbound halo slots (warp-uniform)
    const bool ll_s = RING && g.lr == 0;
    const bool ll_n = RING && g.lr == d.rows - 1;

    // ---- 9 aligned 128-bit loads (all issued before first use)
    const float4 c0 = ld_cg4(src + (g.rc));
    const float4 c1 = ld_cg4(src + (1u * ps + g.rc));
    const float4 c3 = ld_cg4(src + (3u * ps + g.rc));
    float4 sv[3], nv[3];      // f2,f5,f6 of the south row; f4,f7,f8 of the north row
    if (!ll_s) {
        sv[0] = ld_cg4(src + (2u * ps + g.rs));
        sv[1] = ld_cg4(src + (5u * ps + g.rs));
        sv[2] = ld_cg4(src + (6u * ps + g.rs));
    }
    if (!ll_n) {
        nv[0] = ld_cg4(src + (4u * ps + g.rn));
        nv[1] = ld_cg4(src + (7u * ps + g.rn));
        nv[2] = ld_cg4(src + (8u * ps + g.rn));
    }
    // ---- the elements that cross the warp's x range (one lane each side)
    float w1 = 0.f, w5 = 0.f, w8 = 0.f, e3 = 0.f, e6 = 0.f, e7 = 0.f;
    if (g.need_w) {
        w1 = ld_cg1(src + (1u * ps + g.wc));
        if (!ll_s) w5 = ld_cg1(src + (5u * ps + g.ws));
        if (!ll_n) w8 = ld_cg1(src + (8u * ps + g.wn));
    }
    if (g.need_e) {
        e3 = ld_cg1(src + (3u * ps + g.ec));
        if (!ll_s) e6 = ld_cg1(src + (6u * ps + g.es));
        if (!ll_n) e7 = ld_cg1(src + (7u * ps + g.en));
    }
    const unsigned mbyte = ld_mask(d.mask + ((unsigned)(g.lr + 1) * (unsigned)d.mask_pitch + ((unsigned)g.xc >> 3)));
    const unsigned nib = (mbyte >> (g.xc & 4)) & 0xFu;
    if (ll_s | ll_n) {
        const long long t0 = ll_cycles ? clock64() : 0;
        bool staged = false;       // `stage`: this task's records were prefetched into shared memory
        if (stage) {
#ifdef LBM_STAGE_BULK
            stage_bar_wait((void*)(stage + kStageRecords), stage_parity);
#else
            cp_async_wait_all_but_last();
#endif
            staged = ll_s ? ll_take<0>(stage, g, lane, tag, sv, w5, e6) : ll_take<1>(stage, g, lane, tag, nv, w8, e7);
        }
        if (!staged) {
            if (ll_s) ll_pull<0>(d, d.ll_in + ll_slot(d, par, 0), g, tag, sv, w5, e6);
            if (ll_n) ll_pull<1>(d, d.ll_in + ll_slot(d, par, 1), g, tag, nv, w8, e7);
        }
        if (ll_cycles) *ll_cycles += clock64() - t0;
    }
    const float4 s2 = sv[0], s5 = sv[1], s6 = sv[2], n4 = nv[0], n7 = nv[1], n8 = nv[2];

    // ---- lane-crossing elements by shuffle
    const unsigned full = 0xffffffffu;
    float t;
    t = __shfl_up_sync(full, c1.w, 1);   if (!g.need_w) w1 = t;
    t = __shfl_up_sync(full, s5.w, 1);   if (!g.need_w) w5 = t;
    t = __shfl_up_sync(full, n8.w, 1);   if (!g.need_w) w8 = t;
    t = __shfl_down_sync(full, c3.x, 1); if (!g.need_e) e3 = t;
    t = __shfl_down_sync(full, s6.x, 1); if (!g.need_e) e6 = t;
    t = __shfl_down_sync(full, n7.x, 1); if (!g.need_e) e7 = t;

    // ---- pulled populations of the 4 cells  (d2q9-bgk.c:315)
    float p[4][9];
    p[0][0] = c0.x; p[1][0] = c0.y; p[2][0] = c0.z; p[3][0] = c0.w;
    p[0][1] = w1;   p[1][1] = c1.x; p[2][1] = c1.y; p[3][1] = c1.z;
    p[0][2] = s2.x; p[1][2] = s2.y; p[2][2] = s2.z; p[3][2] = s2.w;
    p[0][3] = c3.y; p[1][3] = c3.z; p[2][3] = c3.w; p[3][3] = e3;
    p[0][4] = n4.x; p[1][4] = n4.y; p[2][4] = n4.z; p[3][4] = n4.w;
    p[0][5] = w5;   p[1][5] = s5.x; p[2][5] = s5.y; p[3][5] = s5.z;
    p[0][6] = s6.y; p[1][6] = s6.z; p[2][6] = s6.w; p[3][6] = e6;
    p[0][7] = n7.y; p[1][7] = n7.z; p[2][7] = n7.w; p[3][7] = e7;
    p[0][8] = w8;   p[1][8] = n8.x; p[2][8] = n8.y; p[3][8] = n8.z;

    float usum = collide4<S>(p, nib, d.omega);

    // ---- accelerate_flow for the next step, applied to what this step stores in row ny-2
    if (epilogue_accel && g.lr == d.accel_lr) {
#pragma unroll
        for (int j = 0; j < 4; ++j) accelerate_cell(p[j], (nib >> j) & 1u, d.aw1, d.aw2);
    }

    if (g.active) {
        float4 o[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) o[k] = make_float4(p[0][k], p[1][k], p[2][k], p[3][k]);
#pragma unroll
        for (int k = 0; k < 9; ++k) st_cg4(dst + ((unsigned)k * ps + g.rc), o[k]);
        // halo push over NVLink: only the three populations that cross the slab edge, as tagged
        // records into the ring neighbour's inbound slots of the other parity.  The last step of a run
        // pushes nothing (the next run's prologue republishes the edge rows): a slab whose kernel has
        // finished has therefore consumed every record a neighbour will ever send it, so nothing can
        // still be in flight towards its memory when the host frees it (lbm_destroy).
        const int Q = d.nx >> 2, q = g.x0 >> 2;
        if (ll_n && epilogue_accel)   // my top row -> the up neighbour's side 0 (row below its first row)
            st_halo(d.ll_up + ll_slot(d, par ^ 1, 0) + q, Q, o[2], o[5], o[6], tag + 1u);
        if (ll_s && epilogue_accel)   // my bottom row -> the down neighbour's side 1 (row above its last row)
            st_halo(d.ll_down + ll_slot(d, par ^ 1, 1) + q, Q, o[4], o[7], o[8], tag + 1u);
    } else {
        usum = 0.f;
    }
    return usum;
}

// accelerate_flow (d2q9-bgk.c:235-266) applied IN PLACE to one task of row ny-2 of buffer `par`.
// Used once per lbm_run, before its first step; later steps get the body force through
// process_task's epilogue.
__device__ __forceinline__ void accelerate_task(const SlabDev& d, int par, int2 td, int lane)
{
    const TaskGeom g = task_geom<true>(d, td, lane);       // only the task's own row is touched
    if (!g.active) return;
    float* b = d.buf[par];
    const long long off = (long long)g.rc;
    const unsigned mbyte = ld_mask(d.mask + ((unsigned)(g.lr + 1) * (unsigned)d.mask_pitch + ((unsigned)g.x0 >> 3)));
    const unsigned nib = (mbyte >> (g.x0 & 4)) & 0xFu;
    const int ks[6] = {1, 3, 5, 6, 7, 8};
    float4 v[9];
#pragma unroll
    for (int i = 0; i < 6; ++i) v[ks[i]] = ld_cg4(b + ks[i] * d.ps + off);
    float p[4][9];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        const int k = ks[i];
        p[0][k] = v[k].x; p[1][k] = v[k].y; p[2][k] = v[k].z; p[3][k] = v[k].w;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) accelerate_cell(p[j], (nib >> j) & 1u, d.aw1, d.aw2);
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        const int k = ks[i];
        v[k] = make_float4(p[0][k], p[1][k], p[2][k], p[3][k]);
        st_cg4(b + k * d.ps + off, v[k]);
    }
}

// Start of a run: (re)publishes one task of a slab edge row of buffer `par` into the ring
// neighbour's inbound slots with the run's first tag, so that every lbm_run starts from what the
// lattice holds (initial state, lbm_set_cells, or the previous run's result + accelerate prologue).
// Must be called by the same lanes that ran accelerate_task on this task, after it.
__device__ __forceinline__ void publish_edge_task(const SlabDev& d, int par, int2 td, int lane, unsigned tag)
{
    const TaskGeom g = task_geom<true>(d, td, lane);
    if (!g.active) return;
    const float* b = d.buf[par] + g.rc;
    const int Q = d.nx >> 2, q = g.x0 >> 2;
    if (g.lr == d.rows - 1)
        st_halo(d.ll_up + ll_slot(d, par, 0) + q, Q, ld_cg4(b + 2 * d.ps), ld_cg4(b + 5 * d.ps), ld_cg4(b + 6 * d.ps), tag);
    if (g.lr == 0)
        st_halo(d.ll_down + ll_slot(d, par, 1) + q, Q, ld_cg4(b + 4 * d.ps), ld_cg4(b + 7 * d.ps), ld_cg4(b + 8 * d.ps), tag);
}

// Sum of a double over the warp.
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------------------------------
// Engine 2: one launch per timestep (replayed from a CUDA graph).  Single slab only.
// Grid: ceil(num_tasks / warps_per_cta) CTAs, one task per warp.
// ------------------------------------------------------------------------------------------
template <bool S>
__global__ void __launch_bounds__(256) lbm_step_kernel(SlabDev d, int par, const int* step_ptr, int step_off,
                                                        int epilogue_accel)
{
    __shared__ double red[kMaxWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int task = blockIdx.x * nw + warp;
    double s = 0.0;
    if (task < d.rows * d.tpr) s = (double)process_task<S, false>(d, par, __ldg(d.task_order + task), lane, epilogue_accel != 0, 0u);
    s = warp_sum(s);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < nw; ++w) t += red[w];
        atomicAdd(d.tot + (*step_ptr + step_off), t);
    }
}

__global__ void lbm_accel_kernel(SlabDev d, int par)
{
    const int lane = threadIdx.x & 31;
    const int chunk = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (d.accel_lr < 0 || chunk >= d.tpr) return;
    accelerate_task(d, par, make_int2(d.accel_lr, chunk), lane);
}

__global__ void lbm_advance_step_kernel(int* step_ptr, int n) { *step_ptr += n; }

// ------------------------------------------------------------------------------------------
// Engine 1: persistent, neighbour-synchronised.  Launched cooperatively (for the
// co-residency guarantee only; there is NO grid-wide barrier).  Every CTA owns a fixed set of
// tasks, loops over all `iters` timesteps and synchronises only with the CTAs of its own GPU
// that own adjacent tasks (progress flags); the ring-neighbour GPUs are coupled through the
// tagged halo slots alone.
//   progress value of a CTA = base + 1 after the prologue, base + 1 + (s+1) after local step s.
//   Boundary tasks of step s wait for all dependencies to reach base + 1 + s (their step s-1
//   results are complete, and they no longer read the cells about to be overwritten); the same
//   value is the tag step s expects on its inbound halo slots.
// A CTA's step: interior tasks (need nobody), wait, slab-edge tasks (their halo pushes are then
// acknowledged long before the release below, which waits for outstanding remote writes), the
// other boundary tasks, CTA barrier, release of the progress flag.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void wait_deps(const SlabDev& d, int dep_lo, int dep_hi, unsigned target, int lane)
{
    for (int i = dep_lo + lane; i < dep_hi; i += 32) {
        const unsigned* f = d.flags + __ldg(d.deps + i);
        const long long t0 = clock64();
        for (;;) {
            const unsigned v = ld_acquire_gpu(f);
            if ((int)(v - target) >= 0) break;
            if (ld_status(d.status) != 0) break;
            if (clock64() - t0 > kSpinLimitCycles) { atomicExch(d.status, 2); break; }
            __nanosleep(20);
        }
    }
    __syncwarp();
}

template <bool S, bool RING, bool PROF>
__global__ void __launch_bounds__(512, 1) lbm_persistent_kernel(SlabDev d, int par0, int step0, int iters, unsigned base)
{
    __shared__ double red[2][kMaxWarps];
    extern __shared__ uint4 ll_stage[];          // multi-slab: [warp][2][kStageRecords]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int b = blockIdx.x;
    const int t_lo = __ldg(d.task_begin + b), t_hi = __ldg(d.task_begin + b + 1);
    const int n_int = __ldg(d.n_interior + b);
    const int edge_end = RING ? n_int + __ldg(d.n_edge + b) : n_int;
    const int dep_lo = __ldg(d.dep_begin + b), dep_hi = __ldg(d.dep_begin + b + 1);
    const int n_tasks = t_hi - t_lo;
    // edge tasks are staged one task ahead; a one-row slab needs both sides at once and just polls
    // (only the CTAs that own edge tasks carry the staging bookkeeping through their task loop)
    const bool staging = RING && d.rows > 1 && edge_end > n_int;
    uint4* const my_stage = ll_stage + warp * 2 * kStageSlot;
    unsigned stage_phase = 0;                     // (bulk staging) bit b: parity the next wait on stage buffer b expects
#ifdef LBM_STAGE_BULK
    if (RING && lane < 2) stage_bar_init(my_stage + lane * kStageSlot + kStageRecords);
    if (RING) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#endif
    constexpr bool prof = PROF;               // LBM_PROFILE=1 runs the instantiation with cycle counters
    long long c_wait = 0, c_bar = 0, c_ll = 0;
    const long long c_begin = prof ? clock64() : 0;

    // ---- prologue: accelerate_flow for the first step of this run, in place; edge rows of a
    // slab are (re)published to the ring neighbours with the run's first tag
    if (d.accel_lr >= 0 || RING) {
        for (int k = warp; k < n_tasks; k += nw) {
            const int2 td = __ldg(d.task_order + t_lo + k);
            if (td.x == d.accel_lr) accelerate_task(d, par0, td, lane);
            if (RING && (td.x == 0 || td.x == d.rows - 1)) publish_edge_task(d, par0, td, lane, base + 1u);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) st_release_gpu(d.flags + b, base + 1u);

    int par = par0;
    for (int s = 0; s < iters; ++s) {
        const bool epi = (s + 1 < iters);
        const unsigned target = base + 1u + (unsigned)s;
        double acc = 0.0;
        bool waited = false, staged = false;
        int ord = 0;                              // ordinal of the task within this warp's step
        for (int k = warp; k < n_tasks; k += nw, ++ord) {
            bool staged_next = false;
            if (staging) {
                const int kn = k + nw;           // edge tasks sit at [n_int, edge_end) of the CTA's list
                if (kn >= n_int && kn < edge_end) {
                    ll_prefetch(d, par, __ldg(d.task_order + t_lo + kn), lane, my_stage + ((ord + 1) & 1) * kStageSlot);
                    staged_next = true;
                }
#ifndef LBM_STAGE_BULK
                cp_async_commit();               // one group per task, empty or not
#endif
            }
            if (k >= n_int && !waited) {
                const long long t0 = prof ? clock64() : 0;
                wait_deps(d, dep_lo, dep_hi, target, lane);
                if (prof) c_wait += clock64() - t0;
                waited = true;
            }
            const int2 td = __ldg(d.task_order + t_lo + k);
            acc += (double)process_task<S, RING>(d, par, td, lane, epi, target, staged ? my_stage + (ord & 1) * kStageSlot : nullptr,
                                           prof ? &c_ll : nullptr, (stage_phase >> (ord & 1)) & 1u);
            if (staged) stage_phase ^= 1u << (ord & 1);
            staged = staged_next;
        }
        acc = warp_sum(acc);
        if (lane == 0) red[s & 1][warp] = acc;
        const long long t1 = prof ? clock_fenced() : 0;
        __syncthreads();
        if (prof) c_bar += clock_fenced() - t1;
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int w = 0; w < nw; ++w) t += red[s & 1][w];
            if (chk_ptr(d.partials + ((long long)(step0 + s) * d.num_ctas + b), 8))
                d.partials[(long long)(step0 + s) * d.num_ctas + b] = t;
            st_release_gpu(d.flags + b, target + 1u);
        }
        par ^= 1;
    }
    if (prof && lane == 0) {
        long long* o = d.prof + ((long long)b * kMaxWarps + warp) * 4;
        o[0] = c_wait; o[1] = c_bar; o[2] = c_ll; o[3] = clock64() - c_begin;
    }
}

// ------------------------------------------------------------------------------------------
// Engine 3: dataflow over tagged records - the small-lattice engine.
// On a lattice with no more warp-tasks than the GPU holds resident warps (128x128 ... 512x512) a
// timestep of engine 1 is ONE dependency chain: store, fence, release flag, poll flag, load, compute;
// the arithmetic is a small part of it.  Here every warp owns one task for the whole run and the lattice
// itself lives in the tagged 16-byte record format of the halo slots (lbm_device.cuh): a row's 9
// populations are kept as three record groups by DESTINATION row - northbound (f2,f5,f6: read by the row
// above), southbound (f4,f7,f8: read by the row below), horizontal (f0,f1,f3: read by the row itself) -
// each group 4 record planes of nx/4 records {v,v,v,tag}.  A warp polls the three groups it pulls from
// until all carry the step's tag, collides, and stores its own three groups with tag + 1 into the other
// parity.  No flag, no fence, no barrier, no CTA: one warp per block.  A record is rewritten two steps
// later; by then every reader has published the output of the step that read it, which the writer had
// to see before it could compute (same argument as for the halo slots).
// The SoA lattice stays the canonical state BETWEEN runs: the prologue converts the warp's own row
// segment into records (after the in-place accelerate_flow of the first step), the last step stores
// SoA.  Per step every warp leaves its own partial sum of |u| (partials[step][task], folded later in a
// fixed order).
// Row slabs on several GPUs (RING): the record lattice has one haloed row either side, exactly like the
// SoA lattice; the top row's northbound group is stored straight into the up neighbour's haloed row 0 and
// the bottom row's southbound group into the down neighbour's last haloed row, through peer pointers -
// the consumer still polls only its own memory, and the cross-GPU hop needs nothing beyond the records
// themselves (this is engine 1's halo-slot protocol applied to every row).  This is the few-rows-per-GPU
// path: a 1024x1024 lattice on 8 GPUs is 1024 resident warps per GPU.
// ------------------------------------------------------------------------------------------
constexpr int kRecNorthbound = 0, kRecSouthbound = 1, kRecHorizontal = 2;

// Record group `kind` of haloed row `hrow` (local row r = haloed row r + 1) in a record lattice of `rows` real rows.
__device__ __forceinline__ uint4* rec_group_at(uint4* base, int rows, int nx, int par, int hrow, int kind)
{
    return base + ((size_t)(par * (rows + 2) + hrow) * 3 + kind) * (size_t)nx;
}
__device__ __forceinline__ uint4* rec_group(const SlabDev& d, int par, int hrow, int kind)
{
    return rec_group_at(d.rec, d.rows, d.nx, par, hrow, kind);
}
// Where local row lr's groups are READ from / WRITTEN to.  Single lattice: y wraps inside the slab.
template <bool RING> __device__ __forceinline__ int rec_hrow_south(const SlabDev& d, int lr) { return lr == 0 ? (RING ? 0 : d.rows) : lr; }
template <bool RING> __device__ __forceinline__ int rec_hrow_north(const SlabDev& d, int lr) { return lr == d.rows - 1 ? (RING ? d.rows + 1 : 1) : lr + 2; }
template <bool RING> __device__ __forceinline__ uint4* rec_out_north(const SlabDev& d, int par, int lr)
{
    if (RING && lr == d.rows - 1) return rec_group_at(d.rec_up, d.rows_up, d.nx, par, 0, kRecNorthbound);
    return rec_group(d, par, lr + 1, kRecNorthbound);
}
template <bool RING> __device__ __forceinline__ uint4* rec_out_south(const SlabDev& d, int par, int lr)
{
    if (RING && lr == 0) return rec_group_at(d.rec_down, d.rows_down, d.nx, par, d.rows_down + 1, kRecSouthbound);
    return rec_group(d, par, lr + 1, kRecSouthbound);
}

// One record group of this lane's 4 cells (+ the west / east element on the lanes that need one);
// returns whether every tag matched.  SIDE as in ll_pull: 0 = the west element is b3 and the east
// element c0 (northbound and horizontal groups), 1 = west c3, east b0 (southbound group).
template <int SIDE>
__device__ __forceinline__ bool rec_take(const SlabDev& d, const uint4* base, const TaskGeom& g, unsigned tag, float4 (&v)[3],
                                         float& w, float& e)
{
    const int Q = d.nx >> 2;
    const uint4 *pq, *pw, *pe;
    ll_addresses<SIDE>(d, base, g, pq, pw, pe);
    const uint4 r0 = ld_record(pq), r1 = ld_record(pq + Q), r2 = ld_record(pq + 2 * Q), r3 = ld_record(pq + 3 * Q);
    bool ok = r0.w == tag && r1.w == tag && r2.w == tag && r3.w == tag;
    v[0] = make_float4(__uint_as_float(r0.x), __uint_as_float(r0.y), __uint_as_float(r0.z), __uint_as_float(r1.x));
    v[1] = make_float4(__uint_as_float(r1.y), __uint_as_float(r1.z), __uint_as_float(r2.x), __uint_as_float(r2.y));
    v[2] = make_float4(__uint_as_float(r2.z), __uint_as_float(r3.x), __uint_as_float(r3.y), __uint_as_float(r3.z));
    if (g.need_w) { const uint4 q = ld_record(pw); ok = ok && q.w == tag; w = __uint_as_float(SIDE ? q.z : q.y); }
    if (g.need_e) { const uint4 q = ld_record(pe); ok = ok && q.w == tag; e = __uint_as_float(SIDE ? q.y : q.z); }
    return ok;
}

// One dataflow task as a lane sees it.  Tasks are numbered chunk-major (id = chunk * rows + row), so a
// warp that owns several consecutive ids owns vertically adjacent rows of one chunk.
struct DfTask {
    TaskGeom g;
    int id, q, hrow_s, hrow_n;
    unsigned nib;
};

template <bool RING>
__device__ __forceinline__ DfTask df_task(const SlabDev& d, int id, int lane)
{
    DfTask t;
    t.id = id;
    t.g = task_geom<RING>(d, make_int2(id % d.rows, id / d.rows), lane);
    t.q = t.g.x0 >> 2;
    t.hrow_s = rec_hrow_south<RING>(d, t.g.lr);
    t.hrow_n = rec_hrow_north<RING>(d, t.g.lr);
    const unsigned mbyte = ld_mask(d.mask + ((unsigned)(t.g.lr + 1) * (unsigned)d.mask_pitch + ((unsigned)t.g.xc >> 3)));
    t.nib = (mbyte >> (t.g.xc & 4)) & 0xFu;
    return t;
}

// Prologue of a run: the task's row segment of the SoA lattice -> records of parity rp0 (= step0 & 1), tag base + 1,
// after the in-place accelerate_flow of the first step (d2q9-bgk.c:235-266).
template <bool RING>
__device__ __forceinline__ void df_prologue(const SlabDev& d, const DfTask& t, int par0, int rp0, unsigned base)
{
    const TaskGeom& g = t.g;
    const int Q = d.nx >> 2;
    const unsigned ps = (unsigned)d.ps;
    const float* src = d.buf[par0];
    float4 v[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) v[k] = ld_cg4(src + ((unsigned)k * ps + g.rc));
    if (g.lr == d.accel_lr) {
        float p[4][9];
#pragma unroll
        for (int k = 0; k < 9; ++k) { p[0][k] = v[k].x; p[1][k] = v[k].y; p[2][k] = v[k].z; p[3][k] = v[k].w; }
#pragma unroll
        for (int j = 0; j < 4; ++j) accelerate_cell(p[j], (t.nib >> j) & 1u, d.aw1, d.aw2);
#pragma unroll
        for (int k = 0; k < 9; ++k) v[k] = make_float4(p[0][k], p[1][k], p[2][k], p[3][k]);
    }
    if (g.active) {
        st_halo(rec_out_north<RING>(d, rp0, g.lr) + t.q, Q, v[2], v[5], v[6], base + 1u);
        st_halo(rec_out_south<RING>(d, rp0, g.lr) + t.q, Q, v[4], v[7], v[8], base + 1u);
        st_halo(rec_group(d, rp0, g.lr + 1, kRecHorizontal) + t.q, Q, v[0], v[1], v[3], base + 1u);
    }
}

// One timestep of one task; false = bail out (status set: a neighbour never arrived).
template <bool S, bool RING>
__device__ __forceinline__ bool df_step(const SlabDev& d, const DfTask& t, int lane, int s, int iters, int par0, int step0, unsigned base)
{
    const TaskGeom& g = t.g;
    const int Q = d.nx >> 2;
    const unsigned ps = (unsigned)d.ps;
    const unsigned full = 0xffffffffu;
    const unsigned tag = base + 1u + (unsigned)s;
    const int rp = (step0 + s) & 1;   // record parity continues across runs: the next run's prologue must not hit the
                                      // parity a ring neighbour's last step may still be reading
    float4 cv[3], sv[3], nv[3];      // (f0,f1,f3) of the row; (f2,f5,f6) of the south row; (f4,f7,f8) of the north row
    float w1 = 0.f, w5 = 0.f, w8 = 0.f, e3 = 0.f, e6 = 0.f, e7 = 0.f;
    long long t0 = 0;
#ifdef LBM_DF_PROBE   // dev aid (tools/probes/df_split.py): where does a dataflow step's time go?  needs LBM_PROFILE=1
    const long long pr0 = clock_fenced();
    int pr_tries = 0;
#endif
    // all loads of the three groups go out back to back and are voted on ONCE; polling the groups one by
    // one measured slower (three dependent votes on the fast path)
    for (int tries = 0;; ++tries) {
#ifdef LBM_DF_PROBE
        pr_tries = tries;
#endif
        bool ok = rec_take<0>(d, rec_group(d, rp, g.lr + 1, kRecHorizontal), g, tag, cv, w1, e3);
        ok = rec_take<0>(d, rec_group(d, rp, t.hrow_s, kRecNorthbound), g, tag, sv, w5, e6) && ok;
        ok = rec_take<1>(d, rec_group(d, rp, t.hrow_n, kRecSouthbound), g, tag, nv, w8, e7) && ok;
        if (__all_sync(full, ok)) break;
        // give up warp-uniformly (another warp failed, or a neighbour never arrives)
        if (tries == 0) t0 = clock64();
        const bool late = clock64() - t0 > kSpinLimitCycles;
        if (late) atomicExch(d.status, 2);
        if (__any_sync(full, late || ld_status(d.status) != 0)) return false;
    }
    const float4 c0 = cv[0], c1 = cv[1], c3 = cv[2], s2 = sv[0], s5 = sv[1], s6 = sv[2], n4 = nv[0], n7 = nv[1], n8 = nv[2];
#ifdef LBM_DF_PROBE
    const long long pr1 = clock_fenced();
#endif

    // ---- lane-crossing elements by shuffle, pulled populations (d2q9-bgk.c:315): as process_task
    float x;
    x = __shfl_up_sync(full, c1.w, 1);   if (!g.need_w) w1 = x;
    x = __shfl_up_sync(full, s5.w, 1);   if (!g.need_w) w5 = x;
    x = __shfl_up_sync(full, n8.w, 1);   if (!g.need_w) w8 = x;
    x = __shfl_down_sync(full, c3.x, 1); if (!g.need_e) e3 = x;
    x = __shfl_down_sync(full, s6.x, 1); if (!g.need_e) e6 = x;
    x = __shfl_down_sync(full, n7.x, 1); if (!g.need_e) e7 = x;
    float p[4][9];
    p[0][0] = c0.x; p[1][0] = c0.y; p[2][0] = c0.z; p[3][0] = c0.w;
    p[0][1] = w1;   p[1][1] = c1.x; p[2][1] = c1.y; p[3][1] = c1.z;
    p[0][2] = s2.x; p[1][2] = s2.y; p[2][2] = s2.z; p[3][2] = s2.w;
    p[0][3] = c3.y; p[1][3] = c3.z; p[2][3] = c3.w; p[3][3] = e3;
    p[0][4] = n4.x; p[1][4] = n4.y; p[2][4] = n4.z; p[3][4] = n4.w;
    p[0][5] = w5;   p[1][5] = s5.x; p[2][5] = s5.y; p[3][5] = s5.z;
    p[0][6] = s6.y; p[1][6] = s6.z; p[2][6] = s6.w; p[3][6] = e6;
    p[0][7] = n7.y; p[1][7] = n7.z; p[2][7] = n7.w; p[3][7] = e7;
    p[0][8] = w8;   p[1][8] = n8.x; p[2][8] = n8.y; p[3][8] = n8.z;

    float usum = collide4<S>(p, t.nib, d.omega);
    const bool last = (s + 1 == iters);
    if (!last && g.lr == d.accel_lr) {              // accelerate_flow for the next step
#pragma unroll
        for (int j = 0; j < 4; ++j) accelerate_cell(p[j], (t.nib >> j) & 1u, d.aw1, d.aw2);
    }
    if (g.active) {
        float4 o[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) o[k] = make_float4(p[0][k], p[1][k], p[2][k], p[3][k]);
        if (!last) {
            st_halo(rec_out_north<RING>(d, rp ^ 1, g.lr) + t.q, Q, o[2], o[5], o[6], tag + 1u);
            st_halo(rec_out_south<RING>(d, rp ^ 1, g.lr) + t.q, Q, o[4], o[7], o[8], tag + 1u);
            st_halo(rec_group(d, rp ^ 1, g.lr + 1, kRecHorizontal) + t.q, Q, o[0], o[1], o[3], tag + 1u);
        } else {                                    // the run's result goes back to the SoA lattice
            float* dst = d.buf[par0 ^ (iters & 1)];
#pragma unroll
            for (int k = 0; k < 9; ++k) st_cg4(dst + ((unsigned)k * ps + g.rc), o[k]);
        }
    } else {
        usum = 0.f;
    }
#ifdef LBM_DF_PROBE
    if (d.prof && lane == 0) {
        long long* o = d.prof + (long long)t.id * kMaxWarps * 4;
        const long long pr2 = clock_fenced();
        o[0] += pr1 - pr0; o[1] += pr_tries; o[2] += pr2 - pr1; o[3] += 1;
    }
#endif
    const double acc = warp_sum((double)usum);
    if (lane == 0 && chk_ptr(d.partials + ((long long)(step0 + s) * d.num_ctas + t.id), 8))
        d.partials[(long long)(step0 + s) * d.num_ctas + t.id] = acc;
    return true;
}

// MULTI = false: one task per warp (block b owns task b).  MULTI = true: block b owns the df_k consecutive
// tasks [b * df_k, ...), i.e. a column of vertically adjacent rows, and walks them in order every step
// (a step-s task only ever waits for step s-1 outputs, so no cycle of waits can form).
template <bool S, bool MULTI, bool RING>
__global__ void __launch_bounds__(32, 16) lbm_dataflow_kernel(SlabDev d, int par0, int step0, int iters, unsigned base)
{
    const int lane = threadIdx.x;
    if (!MULTI) {
        // block b = task (row b / tpr, chunk b % tpr): x-neighbours in adjacent blocks (measured: 256x256 1.68 us
        // per step against 1.85 with the chunk-major numbering of the multi-task mode)
        const int id = (blockIdx.x % d.tpr) * d.rows + blockIdx.x / d.tpr;
        const DfTask t = df_task<RING>(d, id, lane);
        df_prologue<RING>(d, t, par0, step0 & 1, base);
        for (int s = 0; s < iters; ++s)
            if (!df_step<S, RING>(d, t, lane, s, iters, par0, step0, base)) return;
    } else {
        const int first = blockIdx.x * d.df_k;
        const int n = min(d.df_k, d.rows * d.tpr - first);
        for (int j = 0; j < n; ++j) df_prologue<RING>(d, df_task<RING>(d, first + j, lane), par0, step0 & 1, base);
        for (int s = 0; s < iters; ++s)
            for (int j = 0; j < n; ++j)
                if (!df_step<S, RING>(d, df_task<RING>(d, first + j, lane), lane, s, iters, par0, step0, base)) return;
    }
}

// ------------------------------------------------------------------------------------------
// Engine 4: row blocks - one CTA per lattice ROW, one cell per thread (nx <= 1024, nx % 32 == 0).
// Engine 3 pays, per timestep, one producer->consumer hop through L2 (530-1300 cycles one way, measured:
// tools/probes/hop_probe.cu) PLUS the collide chain of the consuming warp (1460 cycles for four cells per lane:
// tools/probes/df_split.py).  Cutting the tasks into independent 32-cell warps (one cell per lane, a record per cell)
// was built and measured first and gained nothing (128x128: 1.50 vs 1.52 us per step, 256x256: 1.77 vs 1.65): every
// such task waits for EIGHT other warps instead of two, and each extra neighbour adds its jitter.  Here a row is the
// unit of the dataflow - it only ever waits for the row below and the row above - while its work is spread over
// nx/32 warps on all four schedulers of the SM:
//   * f0 stays in a register; f1 / f3 move one cell along the row by warp shuffle, across warp boundaries through
//     a few words of shared memory: the horizontal record group of engines 3 / 4 never touches memory;
//   * the northbound (f2,f5,f6) and southbound (f4,f7,f8) populations go to the rows above / below as ONE tagged
//     16-byte record {a,b,c,tag} per cell and group; a thread polls exactly two records per step, and
//     the elements beyond a warp's 32 cells also come through shared memory from the warp that polled them;
//   * one __syncthreads per step (shared-memory words double-buffered by step parity), no flag, no fence.
// Row slabs on several GPUs (RING): the edge rows' records go straight into the ring neighbour's record lattice,
// exactly as in engine 3.
// (A whole lattice inside one thread-block cluster's shared memory - the other B200 option - is issue-bound instead:
// at most 16 SMs for 128 tasks x 950 instructions = 1900 cycles per step at 4 IPC, no better than engine 3, and four
// times that on 256x256; DSMEM, 393-632 cycles per hop measured, only helps pairs inside one cluster.  DESIGN.md 4.6.)
// ------------------------------------------------------------------------------------------
template <bool S, bool RING>
__global__ void __launch_bounds__(1024, 1) lbm_row_kernel(SlabDev d, int par0, int step0, int iters, unsigned base)
{
    // [parity][0: f1|f3 of the neighbouring cell, 1: its northbound element (f5|f6), 2: its southbound element (f8|f7)][warp]
    __shared__ unsigned sm_w[2][3][kMaxWarps], sm_e[2][3][kMaxWarps];
    __shared__ int sm_bail;
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int wnext = (warp + 1 == nw) ? 0 : warp + 1, wprev = (warp == 0) ? nw - 1 : warp - 1;    // periodic in x (d2q9-bgk.c:288-289)
    const int lr = blockIdx.x, x = threadIdx.x;
    const bool blocked = (ld_mask(d.mask + ((unsigned)(lr + 1) * (unsigned)d.mask_pitch + ((unsigned)x >> 3))) >> (x & 7)) & 1u;
    const bool accel_row = (lr == d.accel_lr);
    const unsigned ps = (unsigned)d.ps;
    const unsigned cell = (unsigned)(lr + 1) * (unsigned)d.nx + (unsigned)x;          // element offset in a SoA plane
    const int hs = rec_hrow_south<RING>(d, lr), hn = rec_hrow_north<RING>(d, lr);
    if (threadIdx.x == 0) sm_bail = 0;

    // ---- prologue: my cell of the SoA lattice (after the first step's accelerate_flow); f0, f1, f3 stay here,
    // the crossing populations go out as records of parity step0 & 1
    float o0, o1, o3;
    {
        const float* src = d.buf[par0];
        float p[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) p[k] = ld_cg1(src + ((unsigned)k * ps + cell));
        if (accel_row) accelerate_cell(p, blocked, d.aw1, d.aw2);
        const int rp0 = step0 & 1;
        st_record(rec_out_north<RING>(d, rp0, lr) + x, p[2], p[5], p[6], base + 1u);
        st_record(rec_out_south<RING>(d, rp0, lr) + x, p[4], p[7], p[8], base + 1u);
        o0 = p[0]; o1 = p[1]; o3 = p[3];
        if (lane == 31) sm_w[0][0][wnext] = __float_as_uint(o1);
        if (lane == 0) sm_e[0][0][wprev] = __float_as_uint(o3);
    }

#ifdef LBM_DF_PROBE   // dev aid: poll cycles, extra polls, barrier wait, collide + record stores, accumulated in registers
    long long pc0 = 0, pc1 = 0, pc2 = 0, pc3 = 0;
#endif
    // (Measured and dropped, profiles/r02_small_lattice.md: issuing a step's first poll early, behind the previous step's
    // record stores - 1.31 vs 1.17 us per step on 128x128; two to four polls in flight a fraction of an L2 round trip
    // apart - 1.32 / 1.40 / 1.48 vs 1.14.)
    for (int s = 0; s < iters; ++s) {
        const unsigned tag = base + 1u + (unsigned)s;
        const int rp = (step0 + s) & 1;            // record parity continues across runs (see df_step)
        const int sp = s & 1;                      // shared-memory parity
        const uint4* gn = rec_group(d, rp, hs, kRecNorthbound) + x;      // the south row's northbound populations
        const uint4* gs = rec_group(d, rp, hn, kRecSouthbound) + x;      // the north row's southbound populations
        uint4 N, Sg;
        long long t0 = 0;
#ifdef LBM_DF_PROBE
        const long long pr0 = clock_fenced();
        int pr_tries = 0;
#endif
        for (int tries = 0;; ++tries) {
#ifdef LBM_DF_PROBE
            pr_tries = tries;
#endif
            N = ld_record(gn); Sg = ld_record(gs);
            if (__all_sync(full, N.w == tag && Sg.w == tag)) break;
            if (tries == 0) t0 = clock64();
            const bool late = clock64() - t0 > kSpinLimitCycles;
            if (late) atomicExch(d.status, 2);
            if (__any_sync(full, late || ld_status(d.status) != 0)) { if (lane == 0) sm_bail = 1; break; }
        }
        // what the neighbouring warps need of my edge cells: f5 / f8 travel east (from my last cell), f6 / f7 west
        if (lane == 31) { sm_w[sp][1][wnext] = N.y; sm_w[sp][2][wnext] = Sg.z; }
        if (lane == 0) { sm_e[sp][1][wprev] = N.z; sm_e[sp][2][wprev] = Sg.y; }
#ifdef LBM_DF_PROBE
        const long long pr1 = clock_fenced();
#endif
        __syncthreads();
#ifdef LBM_DF_PROBE
        const long long pr2 = clock_fenced();
#endif
        if (sm_bail) return;                       // a neighbour never arrived (status is set): the whole row leaves together
        // pull (d2q9-bgk.c:315): f1,f5,f8 come from the west cell, f3,f6,f7 from the east cell
        float q[1][9];
        const unsigned w1 = __shfl_up_sync(full, __float_as_uint(o1), 1), w5 = __shfl_up_sync(full, N.y, 1), w8 = __shfl_up_sync(full, Sg.z, 1);
        const unsigned e3 = __shfl_down_sync(full, __float_as_uint(o3), 1), e6 = __shfl_down_sync(full, N.z, 1), e7 = __shfl_down_sync(full, Sg.y, 1);
        q[0][0] = o0;
        q[0][2] = __uint_as_float(N.x);
        q[0][4] = __uint_as_float(Sg.x);
        q[0][1] = __uint_as_float(lane == 0 ? sm_w[sp][0][warp] : w1);
        q[0][5] = __uint_as_float(lane == 0 ? sm_w[sp][1][warp] : w5);
        q[0][8] = __uint_as_float(lane == 0 ? sm_w[sp][2][warp] : w8);
        q[0][3] = __uint_as_float(lane == 31 ? sm_e[sp][0][warp] : e3);
        q[0][6] = __uint_as_float(lane == 31 ? sm_e[sp][1][warp] : e6);
        q[0][7] = __uint_as_float(lane == 31 ? sm_e[sp][2][warp] : e7);

        const float usum = collide_n<S, 1>(q, blocked ? 1u : 0u, d.omega);
        const bool last = (s + 1 == iters);
        if (!last && accel_row) accelerate_cell(q[0], blocked, d.aw1, d.aw2);     // accelerate_flow for the next step
        if (!last) {
            st_record(rec_out_north<RING>(d, rp ^ 1, lr) + x, q[0][2], q[0][5], q[0][6], tag + 1u);
            st_record(rec_out_south<RING>(d, rp ^ 1, lr) + x, q[0][4], q[0][7], q[0][8], tag + 1u);
            o0 = q[0][0]; o1 = q[0][1]; o3 = q[0][3];
            if (lane == 31) sm_w[sp ^ 1][0][wnext] = __float_as_uint(o1);
            if (lane == 0) sm_e[sp ^ 1][0][wprev] = __float_as_uint(o3);
        } else {                                    // the run's result goes back to the SoA lattice
            float* dst = d.buf[par0 ^ (iters & 1)];
#pragma unroll
            for (int k = 0; k < 9; ++k)
                if (chk_ptr(dst + ((unsigned)k * ps + cell), 4)) dst[(unsigned)k * ps + cell] = q[0][k];
        }
#ifdef LBM_DF_PROBE
        pc0 += pr1 - pr0; pc1 += pr_tries; pc2 += pr2 - pr1; pc3 += clock_fenced() - pr2;
#endif
        const double acc = warp_sum((double)usum);
        const long long slot = (long long)(step0 + s) * d.num_ctas + (long long)blockIdx.x * nw + warp;
        if (lane == 0 && chk_ptr(d.partials + slot, 8)) d.partials[slot] = acc;
    }
#ifdef LBM_DF_PROBE
    if (d.prof && lane == 0) {
        long long* o = d.prof + ((long long)blockIdx.x * nw + warp) * kMaxWarps * 4;
        o[0] = pc0; o[1] = pc1; o[2] = pc2; o[3] = pc3;
    }
#endif
}

}  // namespace lbm
