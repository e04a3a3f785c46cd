// lbm_kernels.cuh - the warp-task update and the two timestep engines built on it.
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

__device__ __forceinline__ TaskGeom task_geom(const SlabDev& d, int2 td, int lane)
{
    TaskGeom g;
    g.lr = td.x;
    g.x0 = td.y * 128 + lane * 4;
    g.active = g.x0 < d.nx;
    g.xc = g.active ? g.x0 : 0;            // inactive lanes (ragged last chunk) shadow x=0, never store
    const int h = g.lr + 1;                // row index in the haloed layout
    int hs = h - 1, hn = h + 1;
    if (d.wrap_y) {                        // single slab: periodic in y inside the slab
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

// Updates the 4 cells of this lane for one task: pull (propagate), rebound / collision,
// optional accelerate_flow epilogue for the NEXT step on row ny-2, store, halo push to the
// ring neighbours, and returns the lane's sum of |u| over its fluid cells.
template <bool S>
__device__ __forceinline__ float process_task(const SlabDev& d, int par, int2 td, int lane, bool epilogue_accel)
{
    const TaskGeom g = task_geom(d, td, lane);
    const float* __restrict__ src = d.buf[par];
    float* __restrict__ dst = d.buf[par ^ 1];
    const unsigned ps = (unsigned)d.ps;

    // ---- 9 aligned 128-bit loads (all issued before first use)
    const float4 c0 = ld_cg4(src + (g.rc));
    const float4 c1 = ld_cg4(src + (1u * ps + g.rc));
    const float4 c3 = ld_cg4(src + (3u * ps + g.rc));
    const float4 s2 = ld_cg4(src + (2u * ps + g.rs));
    const float4 s5 = ld_cg4(src + (5u * ps + g.rs));
    const float4 s6 = ld_cg4(src + (6u * ps + g.rs));
    const float4 n4 = ld_cg4(src + (4u * ps + g.rn));
    const float4 n7 = ld_cg4(src + (7u * ps + g.rn));
    const float4 n8 = ld_cg4(src + (8u * ps + g.rn));
    // ---- the elements that cross the warp's x range (one lane each side)
    float w1 = 0.f, w5 = 0.f, w8 = 0.f, e3 = 0.f, e6 = 0.f, e7 = 0.f;
    if (g.need_w) {
        w1 = ld_cg1(src + (1u * ps + g.wc));
        w5 = ld_cg1(src + (5u * ps + g.ws));
        w8 = ld_cg1(src + (8u * ps + g.wn));
    }
    if (g.need_e) {
        e3 = ld_cg1(src + (3u * ps + g.ec));
        e6 = ld_cg1(src + (6u * ps + g.es));
        e7 = ld_cg1(src + (7u * ps + g.en));
    }
    const unsigned mbyte = __ldg(d.mask + ((unsigned)(g.lr + 1) * (unsigned)d.mask_pitch + ((unsigned)g.xc >> 3)));
    const unsigned nib = (mbyte >> (g.xc & 4)) & 0xFu;

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
        if (!d.wrap_y) {
            // halo push over NVLink: only the three populations that cross the slab edge
            if (g.lr == d.rows - 1) {          // my top row -> up neighbour's bottom halo (row 0)
                float* u = d.up_buf[par ^ 1] + g.x0;
                st_cg4(u + 2 * d.up_ps, o[2]); st_cg4(u + 5 * d.up_ps, o[5]); st_cg4(u + 6 * d.up_ps, o[6]);
            }
            if (g.lr == 0) {                   // my bottom row -> down neighbour's top halo
                float* w = d.down_buf[par ^ 1] + (long long)(d.down_rows + 1) * d.nx + g.x0;
                st_cg4(w + 4 * d.down_ps, o[4]); st_cg4(w + 7 * d.down_ps, o[7]); st_cg4(w + 8 * d.down_ps, o[8]);
            }
        }
    } else {
        usum = 0.f;
    }
    return usum;
}

// accelerate_flow (d2q9-bgk.c:235-266) applied IN PLACE to one task of row ny-2 of buffer `par`,
// with the halo push of the touched populations.  Used once per lbm_run, before its first step;
// later steps get the body force through process_task's epilogue.
__device__ __forceinline__ void accelerate_task(const SlabDev& d, int par, int2 td, int lane)
{
    const TaskGeom g = task_geom(d, td, lane);
    if (!g.active) return;
    float* b = d.buf[par];
    const long long off = (long long)g.rc;
    const unsigned mbyte = __ldg(d.mask + ((unsigned)(g.lr + 1) * (unsigned)d.mask_pitch + ((unsigned)g.x0 >> 3)));
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
    if (!d.wrap_y) {
        if (g.lr == d.rows - 1) {
            float* u = d.up_buf[par] + g.x0;
            st_cg4(u + 5 * d.up_ps, v[5]); st_cg4(u + 6 * d.up_ps, v[6]);
        }
        if (g.lr == 0) {
            float* w = d.down_buf[par] + (long long)(d.down_rows + 1) * d.nx + g.x0;
            st_cg4(w + 7 * d.down_ps, v[7]); st_cg4(w + 8 * d.down_ps, v[8]);
        }
    }
}

// Sum of a double over the CTA; result valid in thread 0.  `red` holds kMaxWarps doubles.
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
    if (task < d.rows * d.tpr) s = (double)process_task<S>(d, par, __ldg(d.task_order + task), lane, epilogue_accel != 0);
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
// tasks, loops over all `iters` timesteps and synchronises only with the CTAs (local, or on
// the ring-neighbour GPUs through mirrored flags) that own adjacent tasks.
//   progress value of a CTA = base + 1 after the accelerate prologue, base + 1 + (s+1) after
//   local step s.  Boundary tasks of step s wait for all dependencies to reach base + 1 + s.
// ------------------------------------------------------------------------------------------
constexpr long long kSpinLimitCycles = 4000000000LL;   // ~2 s: a neighbour that never arrives

__device__ __forceinline__ bool wait_deps(const SlabDev& d, int dep_lo, int dep_hi, unsigned target, int lane)
{
    bool ok = true;
    for (int i = dep_lo + lane; i < dep_hi; i += 32) {
        const int idx = __ldg(d.deps + i);
        const unsigned* f = d.flags + idx;
        const bool remote = idx >= d.num_ctas;
        const long long t0 = clock64();
        for (;;) {
            const unsigned v = remote ? ld_acquire_sys(f) : ld_acquire_gpu(f);
            if ((int)(v - target) >= 0) break;
            if (*(volatile int*)d.status != 0) { ok = false; break; }
            if (clock64() - t0 > kSpinLimitCycles) { atomicExch(d.status, 2); ok = false; break; }
            __nanosleep(20);
        }
    }
    __syncwarp();
    return ok;
}

template <bool S>
__global__ void __launch_bounds__(512, 1) lbm_persistent_kernel(SlabDev d, int par0, int step0, int iters, unsigned base)
{
    __shared__ double red[2][kMaxWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int b = blockIdx.x;
    const int t_lo = __ldg(d.task_begin + b), t_hi = __ldg(d.task_begin + b + 1);
    const int n_int = __ldg(d.n_interior + b);
    const int dep_lo = __ldg(d.dep_begin + b), dep_hi = __ldg(d.dep_begin + b + 1);
    const int pub_up = d.wrap_y ? -1 : __ldg(d.pub_up + b);
    const int pub_down = d.wrap_y ? -1 : __ldg(d.pub_down + b);
    const int n_tasks = t_hi - t_lo;

    // ---- prologue: accelerate_flow for the first step of this run, in place
    if (d.accel_lr >= 0) {
        for (int k = warp; k < n_tasks; k += nw) {
            const int2 td = __ldg(d.task_order + t_lo + k);
            if (td.x == d.accel_lr) accelerate_task(d, par0, td, lane);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned v = base + 1u;
        if (pub_up >= 0 || pub_down >= 0) {
            __threadfence_system();
            if (pub_up >= 0) st_release_sys(d.up_flags + pub_up, v);
            if (pub_down >= 0) st_release_sys(d.down_flags + pub_down, v);
        }
        st_release_gpu(d.flags + b, v);
    }

    int par = par0;
    for (int s = 0; s < iters; ++s) {
        const bool epi = (s + 1 < iters);
        const unsigned target = base + 1u + (unsigned)s;
        double acc = 0.0;
        bool waited = false;
        for (int k = warp; k < n_tasks; k += nw) {
            if (k >= n_int && !waited) {
                wait_deps(d, dep_lo, dep_hi, target, lane);
                waited = true;
            }
            const int2 td = __ldg(d.task_order + t_lo + k);
            acc += (double)process_task<S>(d, par, td, lane, epi);
        }
        acc = warp_sum(acc);
        if (lane == 0) red[s & 1][warp] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int w = 0; w < nw; ++w) t += red[s & 1][w];
            d.partials[(long long)(step0 + s) * d.num_ctas + b] = t;
            const unsigned v = target + 1u;
            if (pub_up >= 0 || pub_down >= 0) {
                __threadfence_system();
                if (pub_up >= 0) st_release_sys(d.up_flags + pub_up, v);
                if (pub_down >= 0) st_release_sys(d.down_flags + pub_down, v);
            }
            st_release_gpu(d.flags + b, v);
        }
        par ^= 1;
    }
}

}  // namespace lbm
