// lbm_api.cu - implementation of the C ABI in include/lbm_b200.h on top of the sm_100a
// kernels in lbm_kernels.cuh.  No CPU compute path exists here: every lattice update runs
// on the device, and creation fails loudly when no CUDA device is usable.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/lbm_b200.h"
#include "lbm_kernels.cuh"
#include "lbm_plan.hpp"

namespace {

thread_local char g_err[1024] = "";

int fail(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail("%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e_)); \
    } while (0)

// ---- device buffer cache --------------------------------------------------------------------------------
// cudaMalloc / cudaFree of the lattice buffers cost 10-100 ms per context on the B200 boxes (cudaFree of a 600 MB
// buffer has been seen to stall for up to a second: it synchronises the device and unmaps), which a caller that
// creates a context per job pays on its critical path (bench.py e2e: 1024x1024 ran 211 ms and spent 73 ms in
// destroy).  Freed buffers are therefore kept per process, keyed by (device, exact size), up to LBM_CACHE_MB
// (default 3072; 0 disables the cache) and handed to the next lbm_create that asks for the same size; beyond the
// limit the oldest entries are really freed.  lbm_trim_cache() empties it.  Contents are never trusted: every
// buffer whose initial value matters is (re)initialised by slab_create as before.
struct DevCache {
    struct Entry { int dev; size_t bytes; void* p; };
    std::mutex mu;
    std::vector<Entry> idle;                                          // oldest first
    std::unordered_map<void*, std::pair<int, size_t>> live;           // handed out, not yet returned
    std::unordered_map<void*, bool> exported;                         // buffers whose CUDA IPC handle went to other processes:
                                                                      // never returned to the driver before lbm_trim_cache,
                                                                      // because an importer may still hold a mapping
    std::map<std::pair<int, std::string>, void*> ipc_open;            // (device, handle bytes) -> mapping opened in THIS process
    size_t idle_bytes = 0;
    static size_t limit()
    {
        const char* e = getenv("LBM_CACHE_MB");
        return (size_t)((e && *e) ? std::max(0, atoi(e)) : 3072) << 20;
    }
};
DevCache g_cache;

cudaError_t dmalloc_bytes(void** out, size_t bytes)
{
    int dev = 0;
    cudaGetDevice(&dev);
    bytes = std::max<size_t>(bytes, 1);
    {
        std::lock_guard<std::mutex> g(g_cache.mu);
        for (size_t i = g_cache.idle.size(); i-- > 0;) {              // newest first: most likely still mapped hot
            if (g_cache.idle[i].dev == dev && g_cache.idle[i].bytes == bytes) {
                *out = g_cache.idle[i].p;
                g_cache.idle_bytes -= bytes;
                g_cache.idle.erase(g_cache.idle.begin() + (long)i);
                g_cache.live[*out] = {dev, bytes};
                return cudaSuccess;
            }
        }
    }
    const cudaError_t e = cudaMalloc(out, bytes);
    if (e == cudaSuccess) {
        std::lock_guard<std::mutex> g(g_cache.mu);
        g_cache.live[*out] = {dev, bytes};
    }
    return e;
}

// In the checked build every buffer handed out (fresh or recycled) starts as 0xFF bytes = NaN, so a read of
// never-written memory shows up in the bit-exact parity tests.
template <class T> cudaError_t dmalloc(T** out, size_t bytes)
{
    cudaError_t e = dmalloc_bytes((void**)out, bytes);
#ifdef LBM_CHECKED
    const bool poison = !(getenv("LBM_POISON") && atoi(getenv("LBM_POISON")) == 0);
#else
    const bool poison = getenv("LBM_POISON") && atoi(getenv("LBM_POISON")) != 0;   // any build can be asked to
#endif
    if (poison && e == cudaSuccess) e = cudaMemset(*out, 0xFF, std::max<size_t>(bytes, 1));
    if (poison && e == cudaSuccess) e = cudaDeviceSynchronize();
    return e;
}

// Returns a buffer to the cache (or frees it when the cache is off / full).  The caller has synchronised its streams.
void dfree(void* p)
{
    if (!p) return;
    std::vector<DevCache::Entry> evict;
    {
        std::lock_guard<std::mutex> g(g_cache.mu);
        auto it = g_cache.live.find(p);
        if (it == g_cache.live.end()) { evict.push_back({-1, 0, p}); }
        else {
            const DevCache::Entry e{it->second.first, it->second.second, p};
            g_cache.live.erase(it);
            const size_t lim = DevCache::limit();
            const bool sticky = g_cache.exported.count(p) != 0;
            if (e.bytes > lim && !sticky) evict.push_back(e);
            else {
                g_cache.idle.push_back(e);
                g_cache.idle_bytes += e.bytes;
                for (size_t i = 0; g_cache.idle_bytes > lim && i < g_cache.idle.size();) {
                    if (g_cache.exported.count(g_cache.idle[i].p)) { ++i; continue; }     // oldest first, exported ones stay
                    evict.push_back(g_cache.idle[i]);
                    g_cache.idle_bytes -= g_cache.idle[i].bytes;
                    g_cache.idle.erase(g_cache.idle.begin() + (long)i);
                }
            }
        }
    }
    if (evict.empty()) return;
    int cur = 0;
    cudaGetDevice(&cur);
    for (const auto& e : evict) {
        if (e.dev >= 0) cudaSetDevice(e.dev);
        cudaFree(e.p);
    }
    cudaSetDevice(cur);
}

// CUDA IPC mappings of the ring neighbours' buffers are opened once per (device, handle) and kept: with the buffer
// cache the same allocations come back context after context, and cudaIpcOpen/CloseMemHandle cost milliseconds each.
cudaError_t ipc_open_cached(void** out, const cudaIpcMemHandle_t& h)
{
    int dev = 0;
    cudaGetDevice(&dev);
    const std::pair<int, std::string> key(dev, std::string((const char*)&h, sizeof(h)));
    std::lock_guard<std::mutex> g(g_cache.mu);
    auto it = g_cache.ipc_open.find(key);
    if (it != g_cache.ipc_open.end()) { *out = it->second; return cudaSuccess; }
    const cudaError_t e = cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess);
    if (e == cudaSuccess) g_cache.ipc_open[key] = *out;
    return e;
}
void mark_exported(void* p)
{
    std::lock_guard<std::mutex> g(g_cache.mu);
    g_cache.exported[p] = true;
}

// ---- checked build: the per-device table of valid address ranges (see lbm_device.cuh) ---------------------
#ifdef LBM_CHECKED
struct CheckRange { int dev; unsigned long long lo, hi; const void* owner; };
std::mutex g_check_mu;
std::vector<CheckRange> g_check_ranges;

int check_upload(int dev)
{
    lbm::CheckTable t;
    memset(&t, 0, sizeof(t));
    {
        std::lock_guard<std::mutex> g(g_check_mu);
        for (const CheckRange& r : g_check_ranges)
            if (r.dev == dev && t.n < lbm::kMaxCheckRanges) { t.lo[t.n] = r.lo; t.hi[t.n] = r.hi; ++t.n; }
    }
    int cur = 0;
    cudaGetDevice(&cur);
    cudaSetDevice(dev);
    // keep the violation record, replace the ranges
    lbm::CheckTable old;
    cudaError_t e = cudaMemcpyFromSymbol(&old, lbm::g_check, sizeof(old));
    if (e == cudaSuccess) { t.bad_count = old.bad_count; t.bad_ptr = old.bad_ptr; t.bad_bytes = old.bad_bytes; }
    e = cudaMemcpyToSymbol(lbm::g_check, &t, sizeof(t));
    cudaSetDevice(cur);
    return e == cudaSuccess ? 0 : fail("checked build: cudaMemcpyToSymbol: %s", cudaGetErrorString(e));
}
void check_add(int dev, const void* owner, const void* p, size_t bytes)
{
    if (!p || !bytes) return;
    std::lock_guard<std::mutex> g(g_check_mu);
    g_check_ranges.push_back({dev, (unsigned long long)p, (unsigned long long)p + bytes, owner});
}
void check_remove(const void* owner)
{
    std::lock_guard<std::mutex> g(g_check_mu);
    g_check_ranges.erase(std::remove_if(g_check_ranges.begin(), g_check_ranges.end(), [&](const CheckRange& r) { return r.owner == owner; }),
                         g_check_ranges.end());
}
// after a synchronisation: did any kernel on `dev` touch memory outside its ranges?
int check_report(int dev)
{
    lbm::CheckTable t;
    if (cudaMemcpyFromSymbol(&t, lbm::g_check, sizeof(t)) != cudaSuccess) { cudaGetLastError(); return 0; }
    if (t.bad_count == 0) return 0;
    const unsigned long long n = t.bad_count, p = t.bad_ptr, b = t.bad_bytes;
    t.bad_count = t.bad_ptr = t.bad_bytes = 0;
    cudaMemcpyToSymbol(lbm::g_check, &t, sizeof(t));
    return fail("checked build: %llu out-of-range or misaligned device accesses on device %d; first: %llu bytes at %p "
                "(status 3)", n, dev, b, (void*)p);
}
#else
inline int check_upload(int) { return 0; }
inline void check_add(int, const void*, const void*, size_t) {}
inline void check_remove(const void*) {}
inline int check_report(int) { return 0; }
#endif

struct Slab {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
    lbm::SlabPlan plan;
    lbm::SlabDev dev{};
    // owned device allocations
    float* buf[2] = {nullptr, nullptr};
    unsigned char* mask = nullptr;
    int* sched = nullptr;          // all schedule arrays in one allocation
    unsigned* flags = nullptr;
    double* partials = nullptr;
    double* tot = nullptr;
    int* status = nullptr;
    int* step_ptr = nullptr;       // engine 2: device copy of steps_done
    float* fields = nullptr;       // scratch for final-state planes (4 * rows * nx)
    double* scal = nullptr;        // scratch for scalar reductions (2 doubles)
    long long* prof = nullptr;     // LBM_PROFILE=1: per-warp cycle counters of the persistent engine
    // engine 2
    cudaGraphExec_t graph = nullptr;
    int graph_steps = 0;
    int graph_par = -1;
    uint4* rec = nullptr;          // engines 3 / 4: the lattice as tagged records [parity 2][haloed row][group 3][nx]
    uint4* ll_in = nullptr;        // inbound halo slots [parity 2][side 2][record 4][nx/4] (multi-slab only)
    // IPC (one-process-per-GPU mode): the ring neighbours' inbound halo slots
    void* ipc_ll_up = nullptr;
    void* ipc_ll_down = nullptr;
    void* ipc_rec_up = nullptr;    // dataflow engines across processes: the neighbours' record lattices
    void* ipc_rec_down = nullptr;
    int finalized_steps = 0;       // partials already folded into tot
};

}  // namespace

struct lbm_ctx {
    lbm_params prm{};
    lbm_options opt{};
    int nslabs = 0;          // slabs driven by this process
    int rank0 = 0;           // global rank of slabs[0]
    int size = 1;            // total slabs in the ring
    bool rank_mode = false;  // one process per GPU
    bool connected = false;
    std::vector<Slab> slabs;
    int steps_done = 0;
    int par = 0;             // buffer holding the current state
    unsigned base = 0;       // progress-flag base of the next persistent launch
    int engine = 1;
    bool strict = true;
    int threads = 512;
    int df_k = 1;            // engine 3: consecutive tasks per warp
    int ctas_budget = 0;     // persistent CTAs per slab (before clamping to the task count)
    size_t stage_bytes = 0;  // dynamic shared memory per persistent CTA
    int free_cells = 0;      // denominator of av_vels
    int blocked_cells = 0;
    float last_ms = 0.f;
    int last_launches = 0;
    long long h2d_bytes = 0, d2h_bytes = 0;   // host<->device traffic issued through this context
    char desc[512] = "";
};

namespace {

using lbm::SlabDev;

// ---------------------------------------------------------------- utility kernels
__global__ void k_fill_uniform(float* buf, long long ps, float w0, float w1, float w2)
{
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= ps) return;
    buf[i] = w0;
#pragma unroll
    for (int k = 1; k < 5; ++k) buf[k * ps + i] = w1;
#pragma unroll
    for (int k = 5; k < 9; ++k) buf[k * ps + i] = w2;
}

__device__ __forceinline__ bool cell_blocked(const SlabDev& d, int h, int x)
{
    return (d.mask[(long long)h * d.mask_pitch + (x >> 3)] >> (x & 7)) & 1u;
}

// write_values' per-cell fields (d2q9-bgk.c:710-741), same summation order (rho from 0.f, k=0..8).
__global__ void k_final_state(SlabDev d, int par, float density, float* ux, float* uy, float* u, float* pr)
{
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long n = (long long)d.rows * d.nx;
    if (i >= n) return;
    const int lr = (int)(i / d.nx), x = (int)(i - (long long)lr * d.nx);
    const float* f = d.buf[par] + (long long)(lr + 1) * d.nx + x;
    float vx = 0.f, vy = 0.f, vu = 0.f, p;
    if (cell_blocked(d, lr + 1, x)) {
        p = __fmul_rn(density, 1.f / 3.f);
    } else {
        float q[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) q[k] = f[k * d.ps];
        float rho = 0.f;
#pragma unroll
        for (int k = 0; k < 9; ++k) rho = __fadd_rn(rho, q[k]);
        vx = __fdiv_rn(__fsub_rn(__fadd_rn(__fadd_rn(q[1], q[5]), q[8]), __fadd_rn(__fadd_rn(q[3], q[6]), q[7])), rho);
        vy = __fdiv_rn(__fsub_rn(__fadd_rn(__fadd_rn(q[2], q[5]), q[6]), __fadd_rn(__fadd_rn(q[4], q[7]), q[8])), rho);
        vu = __fsqrt_rn(__fadd_rn(__fmul_rn(vx, vx), __fmul_rn(vy, vy)));
        p = __fmul_rn(rho, 1.f / 3.f);
    }
    if (ux) ux[i] = vx;
    if (uy) uy[i] = vy;
    if (u) u[i] = vu;
    if (pr) pr[i] = p;
}

// out[0] += sum |u| over free cells, out[1] += sum of all populations (double accumulation).
__global__ void k_diagnostics(SlabDev d, int par, double* out)
{
    __shared__ double sh[2][32];
    const long long n = (long long)d.rows * d.nx;
    double su = 0.0, sd = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int lr = (int)(i / d.nx), x = (int)(i - (long long)lr * d.nx);
        const float* f = d.buf[par] + (long long)(lr + 1) * d.nx + x;
        float q[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) q[k] = f[k * d.ps];
        float rho = 0.f;
#pragma unroll
        for (int k = 0; k < 9; ++k) rho = __fadd_rn(rho, q[k]);
        sd += (double)rho;
        if (!cell_blocked(d, lr + 1, x)) {
            const float vx = __fdiv_rn(__fsub_rn(__fadd_rn(__fadd_rn(q[1], q[5]), q[8]), __fadd_rn(__fadd_rn(q[3], q[6]), q[7])), rho);
            const float vy = __fdiv_rn(__fsub_rn(__fadd_rn(__fadd_rn(q[2], q[5]), q[6]), __fadd_rn(__fadd_rn(q[4], q[7]), q[8])), rho);
            su += (double)__fsqrt_rn(__fadd_rn(__fmul_rn(vx, vx), __fmul_rn(vy, vy)));
        }
    }
    su = lbm::warp_sum(su);
    sd = lbm::warp_sum(sd);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { sh[0][warp] = su; sh[1][warp] = sd; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { a += sh[0][w]; b += sh[1][w]; }
        atomicAdd(out, a);
        atomicAdd(out + 1, b);
    }
}

// tot[s] = sum over CTAs of partials[s][*], fixed order (deterministic).
__global__ void k_fold_partials(const double* partials, double* tot, int num_ctas, int s_begin, int s_end)
{
    const int s = s_begin + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (s >= s_end) return;
    const int lane = threadIdx.x & 31;
    double a = 0.0;
    for (int b = lane; b < num_ctas; b += 32) a += partials[(long long)s * num_ctas + b];
    a = lbm::warp_sum(a);
    if (lane == 0) tot[s] = a;
}

// Brute-force check of the exact fast paths in lbm_device.cuh against the IEEE intrinsics.
// Operands are hashed bit patterns: mode 0 spans the whole accepted range (and a little beyond it,
// where the range tests must reject), mode 1 concentrates on LBM-like magnitudes (rho ~ 0.1).
__host__ __device__ __forceinline__ unsigned mix32(unsigned long long x)
{
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return (unsigned)x;
}

__global__ void k_selftest_math(unsigned long long n, unsigned long long seed, int mode, unsigned long long* bad)
{
    unsigned long long local = 0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned h1 = mix32(seed + 3 * i), h2 = mix32(seed + 3 * i + 1), h3 = mix32(seed + 3 * i + 2);
        unsigned eb, ea;
        if (mode == 0) { eb = 115u + h3 % 25u; ea = 26u + (h3 >> 8) % 126u; }      // straddles both range limits
        else           { eb = 121u + h3 % 5u;  ea = 90u + (h3 >> 8) % 34u; }
        const float b = __uint_as_float((eb << 23) | (h1 & 0x7fffffu));
        float a = __uint_as_float((h2 & 0x80000000u) | (ea << 23) | (h2 & 0x7fffffu));
        if ((h3 >> 20) % 61u == 0u) a = 0.f;
        if (lbm::div_den_ok(b) && lbm::div_num_ok(a)) {
            const float q = lbm::div_fast(a, b, lbm::rcp_refined(b));
            if (__float_as_uint(q) != __float_as_uint(__fdiv_rn(a, b))) ++local;
        }
        const float x = fabsf(a) * ((h3 & 1u) ? 1.f : b);
        if (lbm::sqrt_arg_ok(x)) {
            if (__float_as_uint(lbm::sqrt_fast(x)) != __float_as_uint(__fsqrt_rn(x))) ++local;
        }
    }
    if (local) atomicAdd(bad, local);
}


// ---- torn-record stress test (lbm_selftest_records) ---------------------------------------------------
// Writer: every thread rewrites its records with {t,t,t,t}, t = 1, 2, ... through st_record, the product's
// store, at full rate, until the host tells it (through a mapped flag) that the reader has finished.  Reader (other GPU over NVLink, or other SMs of the same GPU): sweeps the same records
// through the product's two read paths - ld_record (mode 0) and the cp.async 16-byte copy into shared memory used
// by the halo staging (mode 1) - and counts records whose four words are not all equal (TORN) and how many
// times a record's value changed between two looks (proof that reads and writes really overlapped).  The reader
// never waits for the writer and the writer stops on its own after ~10 s, so nothing can hang.
__global__ void k_records_writer(uint4* slots, int n, volatile int* stop /* host-mapped */, unsigned long long* rounds_done)
{
    const long long t0 = clock64();
    unsigned t = 1;
    for (;; ++t) {
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
            lbm::st_record(slots + i, __uint_as_float(t), __uint_as_float(t), __uint_as_float(t), t);
        // until the host says the reader is done; never longer than ~10 s on its own
        if ((t & 255u) == 0u && (*stop || clock64() - t0 > lbm::kSpinLimitCycles)) break;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) *rounds_done = t;
}

__global__ void k_records_reader(const uint4* slots, int n, int sweeps, int mode, unsigned long long* out /* torn, changes */)
{
    __shared__ uint4 stage[256];
    unsigned long long torn = 0, changes = 0;
    const int i0 = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    for (int sw = 0; sw < sweeps; ++sw) {
        unsigned last = 0xffffffffu;
        for (int i = i0; i < n; i += stride) {
            for (int rep = 0; rep < 4; ++rep) {       // look at the same record a few times in a row
                uint4 v;
                if (mode == 0) {
                    v = lbm::ld_record(slots + i);
                } else {
                    lbm::cp_async_record(stage + threadIdx.x, slots + i);
                    lbm::cp_async_commit();
                    asm volatile("cp.async.wait_group 0;" ::: "memory");
                    v = stage[threadIdx.x];
                }
                if (v.x != v.w || v.y != v.w || v.z != v.w) ++torn;
                if (rep && v.w != last) ++changes;
                last = v.w;
            }
        }
    }
    if (torn) atomicAdd(out, torn);
    if (changes) atomicAdd(out + 1, changes);
}

}  // namespace

// =========================================================================== slab set-up
namespace {

int env_int(const char* name, int dflt);

template <bool PROF>
const void* persistent_pick(bool strict, bool ring)
{
    if (ring) return strict ? (const void*)lbm::lbm_persistent_kernel<true, true, PROF> : (const void*)lbm::lbm_persistent_kernel<false, true, PROF>;
    return strict ? (const void*)lbm::lbm_persistent_kernel<true, false, PROF> : (const void*)lbm::lbm_persistent_kernel<false, false, PROF>;
}

template <bool RING>
const void* dataflow_pick(bool strict, bool multi)
{
    if (multi) return strict ? (const void*)lbm::lbm_dataflow_kernel<true, true, RING> : (const void*)lbm::lbm_dataflow_kernel<false, true, RING>;
    return strict ? (const void*)lbm::lbm_dataflow_kernel<true, false, RING> : (const void*)lbm::lbm_dataflow_kernel<false, false, RING>;
}
const void* dataflow_fn(bool strict, bool multi, bool ring) { return ring ? dataflow_pick<true>(strict, multi) : dataflow_pick<false>(strict, multi); }

const void* row_fn(bool strict, bool ring)
{
    if (ring) return strict ? (const void*)lbm::lbm_row_kernel<true, true> : (const void*)lbm::lbm_row_kernel<false, true>;
    return strict ? (const void*)lbm::lbm_row_kernel<true, false> : (const void*)lbm::lbm_row_kernel<false, false>;
}

// LBM_PROFILE=1 selects the instantiations that carry the per-warp cycle counters
const void* persistent_fn(bool strict, bool ring)
{
    return env_int("LBM_PROFILE", 0) ? persistent_pick<true>(strict, ring) : persistent_pick<false>(strict, ring);
}

int env_int(const char* name, int dflt)
{
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// Runs fn(begin, end) over [0, n) on a few host threads when the range is large (set-up work on the
// 16384x1024 lattice: 16 M obstacle ints to count and pack; downloads of hundreds of MB).
// In one-process-per-GPU mode the `sharers` ranks of a node do this at the same moment: each takes its share of the cores.
std::atomic<int> g_host_sharers{1};
template <class F>
void host_parallel(size_t n, size_t grain, F fn)
{
    unsigned hw = std::thread::hardware_concurrency();
    hw = std::max(1u, (hw ? hw : 1u) / (unsigned)std::max(1, g_host_sharers.load()));
    size_t nt = std::min<size_t>(std::min<size_t>(hw, 8), n / std::max<size_t>(grain, 1));
    if (const char* e = getenv("LBM_HOST_THREADS")) nt = std::min<size_t>(nt, (size_t)std::max(1, atoi(e)));
    if (nt <= 1) { fn((size_t)0, n); return; }
    std::vector<std::thread> th;
    for (size_t t = 1; t < nt; ++t) th.emplace_back([=] { fn(n * t / nt, n * (t + 1) / nt); });
    fn((size_t)0, n / nt);
    for (auto& x : th) x.join();
}

// Packs rows [row_start-1, row_start+rows] (periodic) of the 0/1 obstacle array to bits.
std::vector<unsigned char> pack_mask(const int* obst, int nx, int ny, int row_start, int rows, int pitch)
{
    std::vector<unsigned char> m((size_t)(rows + 2) * pitch, 0);
    host_parallel((size_t)(rows + 2), (size_t)std::max(1, (1 << 20) / nx), [&](size_t h0, size_t h1) {
        for (size_t h = h0; h < h1; ++h) {
            const int gy = ((row_start - 1 + (int)h) % ny + ny) % ny;
            const int* row = obst + (size_t)gy * nx;
            unsigned char* out = m.data() + h * pitch;
            int x = 0;
            for (; x + 8 <= nx; x += 8) {
                unsigned b = 0;
                for (int i = 0; i < 8; ++i) b |= (unsigned)(row[x + i] != 0) << i;
                out[x >> 3] = (unsigned char)b;
            }
            for (; x < nx; ++x)
                if (row[x]) out[x >> 3] |= (unsigned char)(1u << (x & 7));
        }
    });
    return m;
}

// ---- large device -> host downloads (collate, d2q9-bgk.c:395-447) -------------------------------
// A plain cudaMemcpy of hundreds of MB into fresh pageable memory runs at ~4 GB/s on the B200 boxes
// (first-touch page faults and a single staging thread; bench.py e2e phase "download").  Downloads of
// 128 MB and more instead go through pinned staging chunks: a few host threads each take every nt-th chunk,
// D2H it asynchronously on their own stream into one of two pinned buffers and copy the previous
// chunk out to the caller's memory meanwhile.  The pinned buffers live in a process-wide pool.
struct Piece { void* dst; const void* src; size_t bytes; };
constexpr size_t kChunkBytes = (size_t)2 << 20;   // pinning costs ~1 ms per MB (more with 8 ranks at once): keep the pool small
constexpr int kChunksPerBlock = 8;                // 4 threads x 2 chunks, one cudaHostAlloc
std::mutex g_pin_mutex;
std::vector<void*> g_pin_pool;

void* pin_acquire()
{
    std::lock_guard<std::mutex> g(g_pin_mutex);     // held across the allocation: concurrent workers share ONE block
    if (g_pin_pool.empty()) {
        void* p = nullptr;
        if (cudaHostAlloc(&p, kChunkBytes * kChunksPerBlock, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        for (int i = 0; i < kChunksPerBlock; ++i) g_pin_pool.push_back((char*)p + (size_t)i * kChunkBytes);
    }
    void* p = g_pin_pool.back();
    g_pin_pool.pop_back();
    return p;
}
void pin_release(void* p)
{
    if (!p) return;
    std::lock_guard<std::mutex> g(g_pin_mutex);
    g_pin_pool.push_back(p);
}

int download(int device, cudaStream_t stream, const std::vector<Piece>& pieces)
{
    size_t total = 0;
    std::vector<Piece> chunks;
    for (const Piece& pc : pieces) {
        total += pc.bytes;
        for (size_t o = 0; o < pc.bytes; o += kChunkBytes)
            chunks.push_back({(char*)pc.dst + o, (const char*)pc.src + o, std::min(kChunkBytes, pc.bytes - o)});
    }
    int nt = (int)std::min<size_t>(4, chunks.size() / 4);
    if (const char* e = getenv("LBM_HOST_THREADS")) nt = std::min(nt, std::max(1, atoi(e)));
    // a destination in page-locked memory (the caller's long-lived result planes) is DMA-ed into directly
    bool pinned_dst = !pieces.empty();
    for (const Piece& pc : pieces) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, pc.dst) != cudaSuccess) { cudaGetLastError(); pinned_dst = false; break; }
        if (at.type != cudaMemoryTypeHost) { pinned_dst = false; break; }
    }
    // below 128 MB the plain copy wins once the one-off pinning of the staging block is counted (8 ranks x 75 MB at
    // 8 GPUs: ~20 ms plain against ~10 ms + a first call of 60-130 ms)
    if (pinned_dst || total < ((size_t)128 << 20) || nt < 1 || env_int("LBM_PLAIN_DOWNLOAD", 0)) {
        for (const Piece& pc : pieces) CK(cudaMemcpyAsync(pc.dst, pc.src, pc.bytes, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        return 0;
    }
    std::vector<cudaError_t> err(nt, cudaSuccess);
    auto worker = [&](int t) {
        cudaError_t e = cudaSetDevice(device);
        cudaStream_t st = nullptr;
        cudaEvent_t ev[2] = {nullptr, nullptr};
        void* pin[2] = {pin_acquire(), pin_acquire()};
        if (e == cudaSuccess && (!pin[0] || !pin[1])) e = cudaErrorMemoryAllocation;
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
        for (int i = 0; i < 2 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming);
        long prev = -1;
        int k = 0;
        for (size_t i = (size_t)t; e == cudaSuccess && i < chunks.size(); i += (size_t)nt, ++k) {
            e = cudaMemcpyAsync(pin[k & 1], chunks[i].src, chunks[i].bytes, cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess) e = cudaEventRecord(ev[k & 1], st);
            if (prev >= 0 && e == cudaSuccess) {
                e = cudaEventSynchronize(ev[(k - 1) & 1]);
                if (e == cudaSuccess) memcpy(chunks[prev].dst, pin[(k - 1) & 1], chunks[prev].bytes);
            }
            prev = (long)i;
        }
        if (prev >= 0 && e == cudaSuccess) {
            e = cudaEventSynchronize(ev[(k - 1) & 1]);
            if (e == cudaSuccess) memcpy(chunks[prev].dst, pin[(k - 1) & 1], chunks[prev].bytes);
        }
        if (st) cudaStreamSynchronize(st);
        for (int i = 0; i < 2; ++i) if (ev[i]) cudaEventDestroy(ev[i]);
        if (st) cudaStreamDestroy(st);
        pin_release(pin[0]);
        pin_release(pin[1]);
        err[t] = e;
    };
    CK(cudaStreamSynchronize(stream));          // what is downloaded was produced on `stream`
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back(worker, t);
    worker(0);
    for (auto& x : th) x.join();
    for (int t = 0; t < nt; ++t)
        if (err[t] != cudaSuccess) { cudaGetLastError(); return fail("download: %s", cudaGetErrorString(err[t])); }
    return 0;
}

void slab_free(Slab& s)
{
    if (s.stream == nullptr && s.buf[0] == nullptr) return;
    const bool report = getenv("LBM_REPORT") && atoi(getenv("LBM_REPORT")) > 1;
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    cudaSetDevice(s.device);
    check_remove(&s);
    check_upload(s.device);
    if (s.graph) cudaGraphExecDestroy(s.graph);
    // (CUDA IPC mappings of the neighbours' buffers stay open in the process-wide cache, see ipc_open_cached)
    const double t1 = now();
    dfree(s.buf[0]);
    const double t2 = now();
    dfree(s.buf[1]);
    const double t3 = now();
    dfree(s.mask); dfree(s.sched); dfree(s.flags);
    dfree(s.partials); dfree(s.tot); dfree(s.status); dfree(s.step_ptr); dfree(s.fields);
    dfree(s.scal); dfree(s.prof); dfree(s.ll_in); dfree(s.rec);
    const double t4 = now();
    if (s.ev_start) cudaEventDestroy(s.ev_start);
    if (s.ev_stop) cudaEventDestroy(s.ev_stop);
    if (s.stream) cudaStreamDestroy(s.stream);
    if (report)
        fprintf(stderr, "slab_free: graph/ipc %.4f, buf0 %.4f, buf1 %.4f, small buffers %.4f, events/stream %.4f s\n", t1 - t0,
                t2 - t1, t3 - t2, t4 - t3, now() - t4);
    s = Slab();
}

// Decides engine, CTA shape and the per-slab CTA budget from the device and the options.
int configure_engine(lbm_ctx* c, int device)
{
    // attributes, not cudaGetDeviceProperties: that call alone costs 2-3 ms, as much as the rest of lbm_create
    int cc_major = 0, cc_minor = 0, sms = 0;
    CK(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, device));
    CK(cudaDeviceGetAttribute(&cc_minor, cudaDevAttrComputeCapabilityMinor, device));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    if (cc_major < 10)
        return fail("device %d (sm_%d%d) is not a Blackwell-class GPU; this library is built for sm_100a only",
                    device, cc_major, cc_minor);
    c->strict = (c->opt.strict_fp != 2);
    if (env_int("LBM_FAST_FP", 0)) c->strict = false;
    c->engine = c->opt.engine ? c->opt.engine : env_int("LBM_ENGINE", 0);
    if (c->engine < 0 || c->engine > 4) return fail("unknown engine %d", c->engine);
    if (c->engine == 2 && c->size > 1) return fail("engine 2 (launch per timestep) supports a single slab only");
    // The dataflow engines (3: one warp per 128-cell task, 4: one CTA per row) need every task of a slab resident
    // for the whole run and keep one partial sum per warp and step; auto picks them whenever the slab is small enough
    // for both.  Every rank must decide alike, so the decision uses the LARGEST slab of the ring (rank 0's:
    // ceil(ny / size) rows).
    const bool ring = c->size > 1;
    const int rows_max = (c->prm.ny + c->size - 1) / c->size;
    const long long tasks = (long long)((c->prm.nx + lbm::kTaskCells - 1) / lbm::kTaskCells) * rows_max;
    const long long warps_row = (c->prm.nx % 32 == 0 && c->prm.nx <= 1024) ? (long long)(c->prm.nx / 32) * rows_max : -1;
    // resident warps: the smaller of the two instantiations' occupancies, so that whichever is launched
    // (one task per warp, or several) passes the cooperative launch's co-residency check
    int occ3 = 0, occ3m = 0, occ4 = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ3, dataflow_fn(c->strict, false, ring), 32, 0));
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ3m, dataflow_fn(c->strict, true, ring), 32, 0));
    if (warps_row > 0) CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ4, row_fn(c->strict, ring), c->prm.nx, 0));
    const long long warps3 = (long long)std::min(occ3, occ3m) * sms;
    const long long steps = std::max(c->prm.max_iters, 1);
    // one double per resident warp and step (a per-row fold inside engine 4 - one thread per step, or warp 0 every 64 steps
    // from a shared-memory ring - was measured: +0.1 us per step on 128x128 and 256x256, so the sums stay per warp): up to
    // 8 GiB of the 180 GB (256x256 x 80 000 steps = 1.3 GB)
    const long long partial_limit = 8LL << 30;
    const bool mem3 = tasks * steps * (long long)sizeof(double) <= partial_limit;
    const bool mem4 = warps_row * steps * (long long)sizeof(double) <= partial_limit;
    const bool fits4 = occ4 > 0 && rows_max <= (long long)occ4 * sms && mem4;
    // engine 3: a warp owns 1 task, or several consecutive ones when the slab has more tasks than resident warps.
    // Measured (profiles/r01_small_lattice.md): 2 per warp beats engine 1 by 2.3x (1024x512), 3 per warp by 1.4x
    // (1024x768: 82 vs 59 GLUPS), 4 per warp loses to it (1024x896: 74 vs 85, 1024x1024: 73 vs 98) - auto stops
    // at 3; a forced engine 3 may go to 8.
    const int k_min = (int)((tasks + warps3 - 1) / std::max(warps3, 1LL));
    const bool fits3 = k_min <= 3 && mem3;
    c->df_k = k_min;
    if (c->engine == 3) {
        const int k = std::max(k_min, env_int("LBM_DF_TASKS_PER_WARP", 1));
        if (!mem3 || k > 8)
            return fail("engine 3 needs slabs of at most %lld warp-tasks (the largest here has %lld) and at most 8 GiB of "
                        "per-step partial sums", 8 * warps3, tasks);
        c->df_k = k;
    }
    if (c->engine == 4 && !fits4)
        return fail("engine 4 (one CTA per row) needs nx %% 32 == 0, nx <= 1024, slabs of at most %lld rows (the largest here "
                    "has %d) and at most 8 GiB of per-step partial sums", (long long)occ4 * sms, rows_max);
    // auto: the row engine where it measured faster than engine 3 (rows of up to 256 cells: 128x128 1.17 vs 1.58 us per
    // step, 256x256 1.52 vs 1.68; 512x256 loses, 2.34 vs 2.16 - profiles/r02_small_lattice.md), then engine 3, then 1
    if (c->engine == 0) c->engine = (fits4 && c->prm.nx <= 256) ? 4 : fits3 ? 3 : 1;
    c->threads = c->opt.threads_per_cta ? c->opt.threads_per_cta : env_int("LBM_THREADS", 0);
    int per_sm = c->opt.ctas_per_sm ? c->opt.ctas_per_sm : env_int("LBM_CTAS_PER_SM", 0);
    if (c->engine == 1 && c->threads == 0 && per_sm == 0 && c->size <= c->prm.ny) {
        // CTA shape of the persistent kernel, by the share of INTERIOR tasks the planner reaches with two CTAs of 8 warps
        // per SM on the largest slab.  Measured on 16 slab shapes (profiles/r02_cta_shape_sweep.md): below 35 % a CTA's
        // patch is so thin that the flag waits dominate, and ONE CTA of 16 warps per SM - twice the tasks per patch - is
        // faster (1024x1024 +3.5 %, 4096x256 +26 %, 8192x128 +17 %, the 8-GPU slabs of 16384x1024 +2.3 %, never slower);
        // from 36 % up the two smaller CTAs win by 3-8 % (16384x1024, its 2- and 4-GPU slabs, 2048x1024, 1024x2048).
        try {
            const lbm::SlabPlan probe = lbm::make_plan(c->prm.nx, c->prm.ny, 0, c->size, sms * 2);
            if (lbm::interior_fraction(probe) < 0.35) { c->threads = 512; per_sm = 1; }
        } catch (const std::exception& e) {
            return fail("%s", e.what());
        }
    }
    if (c->threads == 0) c->threads = 256;
    if (per_sm == 0) per_sm = 2;
    if (c->threads % 32 || c->threads < 32 || c->threads > 512) return fail("threads_per_cta must be a multiple of 32 in [32,512]");
    // multi-slab: two halo stage buffers per warp in dynamic shared memory
    c->stage_bytes = c->size > 1 ? (size_t)(c->threads / 32) * 2 * lbm::kStageSlot * sizeof(uint4) : 0;
    CK(cudaFuncSetAttribute(persistent_fn(c->strict, c->size > 1), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->stage_bytes));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, persistent_fn(c->strict, c->size > 1), c->threads, c->stage_bytes));
    if (occ < 1) return fail("persistent kernel does not fit on an SM with %d threads", c->threads);
    per_sm = std::max(1, std::min(per_sm, occ));
    c->ctas_budget = sms * per_sm;
    if (c->engine >= 3) {                        // one block of one warp per task (engine 3: or per df_k tasks), or per row
        c->threads = 32;
        c->ctas_budget = (int)tasks;             // the plan keeps one "CTA" per wide task; engine 4 sizes its own partials
        return 0;
    }
    const int cap = env_int("LBM_MAX_CTAS", 0);
    if (cap > 0) c->ctas_budget = std::min(c->ctas_budget, cap);
    return 0;
}

int slab_create(lbm_ctx* c, Slab& s, int rank, int device, const int* obst)
{
    const lbm_params& P = c->prm;
    const bool report = env_int("LBM_REPORT", 0) > 1;           // set-up phase times on stderr (host clock)
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_begin = now();
    s.device = device;
    CK(cudaSetDevice(device));
    CK(cudaFuncSetAttribute(persistent_fn(c->strict, c->size > 1), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->stage_bytes));
    CK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&s.ev_start));
    CK(cudaEventCreate(&s.ev_stop));
    try {
        s.plan = lbm::make_plan(P.nx, P.ny, rank, c->size, c->ctas_budget);
    } catch (const std::exception& e) {
        return fail("%s", e.what());
    }
    const double t_plan = now();
    const lbm::SlabPlan& pl = s.plan;
    const long long ps = (long long)(pl.rows + 2) * P.nx;
    if (9 * ps >= (1LL << 32))
        return fail("slab of %d rows x %d cells is too large for 32-bit lattice offsets (9*(rows+2)*nx must be < 2^32); "
                    "use more GPUs", pl.rows, P.nx);
    // slots of per-step partial sums: one per persistent CTA (engine 1), per task (3) or per warp of a row CTA (4)
    const int G = (c->engine >= 4) ? pl.rows * (P.nx / 32) : pl.num_ctas;
    const int pitch = ((P.nx + 7) / 8 + 15) / 16 * 16;

    for (int i = 0; i < 2; ++i) CK(dmalloc(&s.buf[i], sizeof(float) * 9 * ps));
    CK(dmalloc(&s.mask, (size_t)(pl.rows + 2) * pitch));
    CK(dmalloc(&s.flags, sizeof(unsigned) * G));
    CK(dmalloc(&s.partials, sizeof(double) * (size_t)std::max(P.max_iters, 1) * G));
    CK(dmalloc(&s.tot, sizeof(double) * std::max(P.max_iters, 1)));
    CK(dmalloc(&s.status, sizeof(int)));
    CK(dmalloc(&s.step_ptr, sizeof(int)));
    CK(dmalloc(&s.scal, sizeof(double) * 2));
    CK(cudaMemsetAsync(s.flags, 0, sizeof(unsigned) * G, s.stream));
    if (c->size > 1) {     // tag 0 = never written; the first run's tags start at 1
        CK(dmalloc(&s.ll_in, sizeof(uint4) * 4 * (size_t)P.nx));
        CK(cudaMemsetAsync(s.ll_in, 0, sizeof(uint4) * 4 * (size_t)P.nx, s.stream));
    }
    if (c->engine >= 3) {  // tag 0 = never written
        const size_t bytes = sizeof(uint4) * 2 * 3 * (size_t)(pl.rows + 2) * P.nx;
        CK(dmalloc(&s.rec, bytes));
        CK(cudaMemsetAsync(s.rec, 0, bytes, s.stream));
    }
    CK(cudaMemsetAsync(s.tot, 0, sizeof(double) * std::max(P.max_iters, 1), s.stream));
    CK(cudaMemsetAsync(s.status, 0, sizeof(int), s.stream));
    CK(cudaMemsetAsync(s.step_ptr, 0, sizeof(int), s.stream));

    const double t_alloc = now();
    // schedule arrays, one allocation: task_order as (row, chunk) pairs first (8-byte aligned), then
    // task_begin | n_interior | n_edge | dep_begin | deps
    std::vector<int> h;
    auto put = [&h](const std::vector<int>& v) { size_t o = h.size(); h.insert(h.end(), v.begin(), v.end()); return o; };
    std::vector<int> pairs;
    pairs.reserve(2 * pl.task_order.size());
    for (int t : pl.task_order) { pairs.push_back(t / pl.tpr); pairs.push_back(t % pl.tpr); }
    const size_t o_to = put(pairs);
    const size_t o_tb = put(pl.task_begin), o_ni = put(pl.n_interior), o_ne = put(pl.n_edge),
                 o_db = put(pl.dep_begin), o_dp = put(pl.deps);
    h.push_back(0);
    CK(dmalloc(&s.sched, sizeof(int) * h.size()));
    CK(cudaMemcpyAsync(s.sched, h.data(), sizeof(int) * h.size(), cudaMemcpyHostToDevice, s.stream));
    c->h2d_bytes += (long long)(sizeof(int) * h.size());

    const double t_sched = now();
    const std::vector<unsigned char> m = pack_mask(obst, P.nx, P.ny, pl.row_start, pl.rows, pitch);
    const double t_pack = now();
    CK(cudaMemcpyAsync(s.mask, m.data(), m.size(), cudaMemcpyHostToDevice, s.stream));
    c->h2d_bytes += (long long)m.size();

    // uniform initial state incl. halo rows (d2q9-bgk.c:584-604)
    const float w0 = P.density * 4.f / 9.f, w1 = P.density / 9.f, w2 = P.density / 36.f;
    k_fill_uniform<<<(unsigned)((ps + 255) / 256), 256, 0, s.stream>>>(s.buf[0], ps, w0, w1, w2);
    k_fill_uniform<<<(unsigned)((ps + 255) / 256), 256, 0, s.stream>>>(s.buf[1], ps, w0, w1, w2);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(s.stream));   // h and m go out of scope
    if (report)
        fprintf(stderr, "slab_create[%d]: stream/events+plan %.4f, allocation %.4f, schedule upload %.4f, mask pack %.4f, "
                "uploads+fill+sync %.4f s\n", rank, t_plan - t_begin, t_alloc - t_plan, t_sched - t_alloc, t_pack - t_sched, now() - t_pack);

    SlabDev& d = s.dev;
    d.buf[0] = s.buf[0]; d.buf[1] = s.buf[1];
    d.mask = s.mask; d.ps = ps; d.nx = P.nx; d.rows = pl.rows; d.tpr = pl.tpr; d.mask_pitch = pitch;
    d.accel_lr = pl.accel_local_row;
    d.omega = P.omega;
    d.aw1 = P.density * P.accel / 9.f;       // d2q9-bgk.c:237
    d.aw2 = P.density * P.accel / 36.f;      // d2q9-bgk.c:238
    d.ll_in = s.ll_in; d.ll_up = d.ll_down = nullptr;
    d.rec = s.rec; d.rec_up = d.rec_down = nullptr; d.rows_up = d.rows_down = 0;
    d.df_k = c->df_k;
    d.task_begin = s.sched + o_tb; d.n_interior = s.sched + o_ni; d.n_edge = s.sched + o_ne; d.task_order = (const int2*)(s.sched + o_to);
    d.dep_begin = s.sched + o_db; d.deps = s.sched + o_dp;
    d.flags = s.flags;
    d.num_ctas = G;
    d.partials = s.partials; d.tot = s.tot; d.status = s.status;
    d.prof = nullptr;
    if (env_int("LBM_PROFILE", 0)) {
        const size_t n = sizeof(long long) * (size_t)G * lbm::kMaxWarps * 4;
        CK(dmalloc(&s.prof, n));
        CK(cudaMemset(s.prof, 0, n));
        d.prof = s.prof;
    }
    // checked build: what this slab's kernels may touch on this device (LBM_CHECK_INJECT=1 registers lattice B two
    // rows short - the upper halo row and the last real row of its last plane - so that the checker itself can be
    // tested: the legitimate stores of f8 to that row must then be caught)
    const size_t inject = env_int("LBM_CHECK_INJECT", 0) ? 2 * sizeof(float) * (size_t)P.nx : 0;
    check_add(device, &s, s.buf[0], sizeof(float) * 9 * ps);
    check_add(device, &s, s.buf[1], sizeof(float) * 9 * ps - inject);
    check_add(device, &s, s.mask, (size_t)(pl.rows + 2) * pitch);
    check_add(device, &s, s.flags, sizeof(unsigned) * G);
    check_add(device, &s, s.partials, sizeof(double) * (size_t)std::max(P.max_iters, 1) * G);
    check_add(device, &s, s.ll_in, sizeof(uint4) * 4 * (size_t)P.nx);
    check_add(device, &s, s.rec, sizeof(uint4) * 2 * 3 * (size_t)(pl.rows + 2) * P.nx);
    return check_upload(device);
}

// The tagged halo records need a peer link on which a 16-byte store lands as one unit; that is established (stress test,
// bit-exact tests) for NVLink / NVSwitch, which is also the only kind of link that reports native peer atomics.  A
// PCIe peer-to-peer pair is refused unless LBM_ALLOW_ANY_PEER_LINK=1.
int peer_link_unsuitable(int dev, int peer)
{
    if (dev == peer || env_int("LBM_ALLOW_ANY_PEER_LINK", 0)) return 0;
    int native = 0;
    if (cudaDeviceGetP2PAttribute(&native, cudaDevP2PAttrNativeAtomicSupported, dev, peer) != cudaSuccess) { cudaGetLastError(); native = 0; }
    if (!native)
        return fail("devices %d and %d are peers over a link without native atomics (not NVLink): the 16-byte tagged halo "
                    "records are only validated over NVLink; set LBM_ALLOW_ANY_PEER_LINK=1 to run anyway", dev, peer);
    return 0;
}

int count_cells(lbm_ctx* c, const int* obst)
{
    const lbm_params& P = c->prm;
    const size_t n = (size_t)P.nx * P.ny;
    std::atomic<long long> total{0};
    host_parallel(n, (size_t)1 << 20, [&](size_t i0, size_t i1) {
        long long k = 0;
        for (size_t i = i0; i < i1; ++i) k += (obst[i] != 0);
        total += k;
    });
    const int blocked = (int)total.load();
    c->blocked_cells = blocked;
    // the reference divides by nx*ny - <obstacle file LINES> (d2q9-bgk.c:642)
    c->free_cells = P.nx * P.ny - (P.obstacle_lines >= 0 ? P.obstacle_lines : blocked);
    if (c->free_cells <= 0) return fail("no free cells (nx*ny - obstacles = %d)", c->free_cells);
    return 0;
}

int check_params(const lbm_params* p)
{
    if (!p) return fail("params is NULL");
    if (p->nx < 4 || p->ny < 3) return fail("grid too small: nx=%d ny=%d", p->nx, p->ny);
    if (p->nx % 4) return fail("nx=%d: this build requires nx to be a multiple of 4 (128-bit lattice accesses); "
                               "the reference itself assumes nx %% 16 == 0 (d2q9-bgk.c:291)", p->nx);
    if (p->max_iters < 0) return fail("max_iters < 0");
    return 0;
}

void describe(lbm_ctx* c)
{
    const Slab& s = c->slabs[0];
    snprintf(c->desc, sizeof(c->desc),
             "engine=%s fp=%s slabs=%d/%d ctas/slab=%d threads/cta=%d tasks/slab=%d tasks/row=%d tasks/warp=%d strips=%d interior=%.0f%%",
             c->engine == 1 ? "persistent-neighbour-sync" : c->engine == 3 ? "dataflow-tagged-records"
             : c->engine == 4 ? "row-block-dataflow" : "launch-per-step-graph",
             c->strict ? "strict" : "fast",
             c->nslabs, c->size, c->engine == 4 ? s.plan.rows : s.plan.num_ctas, c->engine == 4 ? c->prm.nx : c->threads, s.plan.num_tasks, s.plan.tpr, c->engine == 3 ? c->df_k : 0, s.plan.strips,
             100.0 * lbm::interior_fraction(s.plan));
}

}  // namespace

// =========================================================================== public API
extern "C" {

int lbm_abi_version(void) { return LBM_ABI_VERSION; }
int lbm_checked_build(void)
{
#ifdef LBM_CHECKED
    return 1;
#else
    return 0;
#endif
}
const char* lbm_last_error(void) { return g_err; }

int lbm_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int lbm_allocate_rows(int ny, int size, int rank, int* row_start, int* rows)
{
    if (size < 1 || rank < 0 || rank >= size || ny < 1) return fail("lbm_allocate_rows: bad arguments");
    const lbm::SlabShape s = lbm::allocate_rows(ny, size, rank);
    if (row_start) *row_start = s.row_start;
    if (rows) *rows = s.rows;
    return 0;
}

int lbm_plan_selftest(int nx, int ny, int size, int ctas)
{
    try {
        int next_row = 0;
        for (int r = 0; r < size; ++r) {
            const lbm::SlabPlan p = lbm::make_plan(nx, ny, r, size, ctas);
            const std::string e = lbm::validate_plan(p);
            if (!e.empty()) return fail("plan of rank %d/%d invalid: %s", r, size, e.c_str());
            if (p.row_start != next_row) return fail("rank %d: slabs do not tile the rows", r);
            next_row += p.rows;
        }
        if (next_row != ny) return fail("slabs cover %d of %d rows", next_row, ny);
    } catch (const std::exception& e) {
        return fail("%s", e.what());
    }
    return 0;
}

// Host set-up code without a device: the threaded obstacle count and bit-pack (count_cells, pack_mask)
// against plain loops, on a hashed 0/1 array, for every slab of a `size`-way row decomposition.
int lbm_host_selftest(int nx, int ny, int size, long long seed)
{
    if (nx < 1 || ny < 1 || size < 1 || size > ny) return fail("bad arguments");
    std::vector<int> obst((size_t)nx * ny);
    long long want = 0;
    for (size_t i = 0; i < obst.size(); ++i) {
        obst[i] = (mix32((unsigned long long)seed * 0x9E3779B97F4A7C15ULL + i) % 7u) == 0u ? 1 : 0;
        want += obst[i];
    }
    lbm_ctx c;
    c.prm.nx = nx; c.prm.ny = ny; c.prm.obstacle_lines = -1;
    if (count_cells(&c, obst.data())) return 1;
    if (c.blocked_cells != want) return fail("count_cells: %d blocked cells, expected %lld", c.blocked_cells, want);
    const int pitch = ((nx + 7) / 8 + 15) / 16 * 16;
    for (int r = 0; r < size; ++r) {
        const lbm::SlabShape sh = lbm::allocate_rows(ny, size, r);
        const std::vector<unsigned char> m = pack_mask(obst.data(), nx, ny, sh.row_start, sh.rows, pitch);
        if (m.size() != (size_t)(sh.rows + 2) * pitch) return fail("pack_mask: wrong size");
        for (int h = 0; h < sh.rows + 2; ++h) {
            const int gy = ((sh.row_start - 1 + h) % ny + ny) % ny;
            for (int x = 0; x < pitch * 8; ++x) {
                const int bit = (m[(size_t)h * pitch + (x >> 3)] >> (x & 7)) & 1;
                const int ref = x < nx ? obst[(size_t)gy * nx + x] : 0;
                if (bit != ref) return fail("pack_mask: slab %d row %d x %d: bit %d, expected %d", r, h, x, bit, ref);
            }
        }
    }
    return 0;
}

int lbm_selftest_math(long long n, long long seed, long long* mismatches)
{
    if (!mismatches || n < 0) return fail("bad arguments");
    if (lbm_device_count() < 1) return fail("no usable CUDA device");
    unsigned long long* d = nullptr;
    CK(cudaMalloc(&d, sizeof(unsigned long long)));
    CK(cudaMemset(d, 0, sizeof(unsigned long long)));
    for (int mode = 0; mode < 2; ++mode) k_selftest_math<<<1184, 256>>>((unsigned long long)n, (unsigned long long)seed, mode, d);
    CK(cudaGetLastError());
    unsigned long long h = 0;
    CK(cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost));
    CK(cudaFree(d));
    *mismatches = (long long)h;
    return 0;
}

int lbm_selftest_records(int writer_device, int reader_device, int sweeps, int mode, long long* torn, long long* changes)
{
    if (!torn || !changes || sweeps < 1 || mode < 0 || mode > 1) return fail("bad arguments");
    const int ndev = lbm_device_count();
    if (writer_device < 0 || reader_device < 0 || writer_device >= ndev || reader_device >= ndev)
        return fail("lbm_selftest_records: devices %d / %d but %d visible", writer_device, reader_device, ndev);
    const int n = 64 * 256;                              // 256 KB of records, the size of a 16384-wide slab's inbound slots
    uint4* slots = nullptr;
    unsigned long long* out = nullptr;
    int* stop = nullptr;
    cudaStream_t sw = nullptr, sr = nullptr;
    CK(cudaSetDevice(reader_device));                    // the slots live with the reader, as the halo slots do
    CK(cudaMalloc(&slots, sizeof(uint4) * n));
    CK(cudaMemset(slots, 0, sizeof(uint4) * n));
    CK(cudaMalloc(&out, 3 * sizeof(unsigned long long)));
    CK(cudaMemset(out, 0, 3 * sizeof(unsigned long long)));
    CK(cudaHostAlloc(&stop, sizeof(int), cudaHostAllocPortable | cudaHostAllocMapped));
    *stop = 0;
    CK(cudaStreamCreateWithFlags(&sr, cudaStreamNonBlocking));
    CK(cudaDeviceSynchronize());
    if (writer_device != reader_device) {
        int can = 0;
        CK(cudaDeviceCanAccessPeer(&can, writer_device, reader_device));
        if (!can) return fail("device %d cannot access peer %d", writer_device, reader_device);
        CK(cudaSetDevice(writer_device));
        cudaError_t e = cudaDeviceEnablePeerAccess(reader_device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else if (e != cudaSuccess) return fail("cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
    }
    CK(cudaSetDevice(writer_device));
    CK(cudaStreamCreateWithFlags(&sw, cudaStreamNonBlocking));
    // Both kernels run once on nothing first: with lazy module loading the FIRST launch of a kernel may have to wait
    // for the device to go idle, which would serialise reader and writer instead of overlapping them.
    *stop = 1;
    k_records_writer<<<1, 32, 0, sw>>>(slots, 0, stop, out + 2);
    CK(cudaStreamSynchronize(sw));
    CK(cudaSetDevice(reader_device));
    k_records_reader<<<1, 32, 0, sr>>>(slots, 0, 0, mode, out);
    CK(cudaStreamSynchronize(sr));
    *stop = 0;
    CK(cudaSetDevice(writer_device));
    // 64 CTAs each: on one GPU both kernels fit side by side (148 SMs), so they overlap there as well
    k_records_writer<<<64, 256, 0, sw>>>(slots, n, stop, out + 2);
    CK(cudaGetLastError());
    CK(cudaSetDevice(reader_device));
    k_records_reader<<<64, 256, 0, sr>>>(slots, n, sweeps, mode, out);
    CK(cudaGetLastError());
    const cudaError_t er = cudaStreamSynchronize(sr);
    *(volatile int*)stop = 1;                            // the writer ends within 256 rounds of seeing this
    CK(er);
    CK(cudaSetDevice(writer_device));
    CK(cudaStreamSynchronize(sw));
    unsigned long long h[3] = {0, 0, 0};
    CK(cudaSetDevice(reader_device));
    CK(cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost));
    CK(cudaFree(slots));
    CK(cudaFree(out));
    CK(cudaFreeHost(stop));
    CK(cudaStreamDestroy(sr));
    CK(cudaSetDevice(writer_device));
    CK(cudaStreamDestroy(sw));
    *torn = (long long)h[0];
    *changes = (long long)h[1];
    return 0;
}

int lbm_trim_cache(void)
{
    std::vector<DevCache::Entry> all;
    std::map<std::pair<int, std::string>, void*> maps;
    {
        std::lock_guard<std::mutex> g(g_cache.mu);
        all.swap(g_cache.idle);
        g_cache.idle_bytes = 0;
        maps.swap(g_cache.ipc_open);
        for (const auto& e : all) g_cache.exported.erase(e.p);
    }
    int cur = 0;
    cudaGetDevice(&cur);
    for (const auto& m : maps) { cudaSetDevice(m.first.first); cudaIpcCloseMemHandle(m.second); }
    for (const auto& e : all) { cudaSetDevice(e.dev); cudaFree(e.p); }
    cudaSetDevice(cur);
    cudaGetLastError();
    return 0;
}

void lbm_destroy(lbm_ctx* c)
{
    if (!c) return;
    // Every slab's kernel must have finished before ANY slab's memory goes: a ring neighbour pushes halo
    // records into this slab's inbound slots until its second-to-last step.  (Once a slab's own kernel has
    // finished it has consumed every record it will ever be sent - the last step pushes nothing - so in
    // one-process-per-GPU mode this local synchronisation is all the teardown needs.)
    for (Slab& s : c->slabs)
        if (s.stream) { cudaSetDevice(s.device); cudaStreamSynchronize(s.stream); }
    for (Slab& s : c->slabs) slab_free(s);
    delete c;
}

static int create_common(lbm_ctx** out, const lbm_params* params, const int* obstacles, const lbm_options* opts,
                         lbm_ctx** cp)
{
    if (!out) return fail("out is NULL");
    *out = nullptr;
    if (check_params(params)) return 1;
    if (!obstacles) return fail("obstacles is NULL");
    if (lbm_device_count() < 1)
        return fail("no usable CUDA device: liblbm_b200 has no CPU fallback (cudaGetDeviceCount found none)");
    lbm_ctx* c = new lbm_ctx();
    c->prm = *params;
    if (opts) c->opt = *opts;
    const auto t0 = std::chrono::steady_clock::now();
    if (count_cells(c, obstacles)) { delete c; return 1; }
    if (env_int("LBM_REPORT", 0) > 1)
        fprintf(stderr, "create: obstacle count %.4f s\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    *cp = c;
    return 0;
}

int lbm_create(lbm_ctx** out, const lbm_params* params, const int* obstacles, int ngpus, const lbm_options* opts)
{
    lbm_ctx* c = nullptr;
    if (create_common(out, params, obstacles, opts, &c)) return 1;
    const int ndev = lbm_device_count();
    const int first = c->opt.first_device;
    if (ngpus <= 0) ngpus = ndev - first;      // "all visible devices"
    ngpus = std::min(ngpus, params->ny);
    if (ngpus < 1 || first < 0 || first + ngpus > ndev || ngpus > LBM_MAX_GPUS) {
        fail("requested %d GPUs from device %d but %d are visible", ngpus, first, ndev);
        delete c;
        return 1;
    }
    c->size = c->nslabs = ngpus;
    c->rank0 = 0;
    c->slabs.resize(ngpus);
    const bool report = env_int("LBM_REPORT", 0) != 0;      // set-up phase times on stderr (host clock)
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    cudaSetDevice(first);
    cudaFree(nullptr);                                      // creates the primary context
    const double t1 = now();
    int rc = configure_engine(c, first);
    const double t2 = now();
    for (int r = 0; r < ngpus && !rc; ++r) rc = slab_create(c, c->slabs[r], r, first + r, obstacles);
    if (report)
        fprintf(stderr, "lbm_create: context %.3f s, engine set-up %.3f s, %d slab(s) (plan, allocation, uploads enqueued) %.3f s\n",
                t1 - t0, t2 - t1, ngpus, now() - t2);
    // ring links over NVLink peer access (the reference's rank_up / rank_down, d2q9-bgk.c:501-502)
    for (int r = 0; r < ngpus && !rc && ngpus > 1; ++r) {
        Slab& s = c->slabs[r];
        Slab& up = c->slabs[(r + 1) % ngpus];
        Slab& dn = c->slabs[(r - 1 + ngpus) % ngpus];
        cudaSetDevice(s.device);
        for (Slab* o : {&up, &dn}) {
            int can = 0;
            cudaDeviceCanAccessPeer(&can, s.device, o->device);
            if (!can) { rc = fail("device %d cannot access peer %d", s.device, o->device); break; }
            if (peer_link_unsuitable(s.device, o->device)) { rc = 1; break; }
            cudaError_t e = cudaDeviceEnablePeerAccess(o->device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) { rc = fail("cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e)); break; }
        }
        s.dev.ll_up = up.ll_in;
        s.dev.ll_down = dn.ll_in;
        s.dev.rec_up = up.rec;     s.dev.rows_up = up.plan.rows;
        s.dev.rec_down = dn.rec;   s.dev.rows_down = dn.plan.rows;
        for (Slab* o : {&up, &dn}) {
            check_add(s.device, &s, o->ll_in, sizeof(uint4) * 4 * (size_t)params->nx);
            check_add(s.device, &s, o->rec, sizeof(uint4) * 2 * 3 * (size_t)(o->plan.rows + 2) * params->nx);
        }
        if (!rc) rc = check_upload(s.device);
    }
    if (rc) { lbm_destroy(c); return 1; }
    c->connected = true;
    describe(c);
    *out = c;
    return 0;
}

int lbm_steps_done(const lbm_ctx* c) { return c ? c->steps_done : -1; }
int lbm_num_slabs(const lbm_ctx* c) { return c ? c->size : -1; }
const char* lbm_describe(lbm_ctx* c) { return c ? c->desc : ""; }
float lbm_inv_free_cells(const lbm_ctx* c) { return c ? 1.f / (float)c->free_cells : 0.f; }
int lbm_last_run_launches(const lbm_ctx* c) { return c ? c->last_launches : -1; }

int lbm_transfer_bytes(const lbm_ctx* c, long long* h2d, long long* d2h)
{
    if (!c) return fail("ctx is NULL");
    if (h2d) *h2d = c->h2d_bytes;
    if (d2h) *d2h = c->d2h_bytes;
    return 0;
}

int lbm_get_profile(lbm_ctx* c, int slab, long long* out, int capacity, int* num_ctas, int* warps)
{
    if (!c || slab < 0 || slab >= c->nslabs) return fail("lbm_get_profile: bad slab index");
    Slab& s = c->slabs[slab];
    if (!s.prof) return fail("lbm_get_profile: profiling is off (set LBM_PROFILE=1 before lbm_create)");
    if (lbm_sync(c)) return 1;
    const int G = s.plan.num_ctas, nw = c->threads / 32;
    if (num_ctas) *num_ctas = G;
    if (warps) *warps = nw;
    if (!out) return 0;
    if (capacity < G * nw * 4) return fail("lbm_get_profile: need room for %d values", G * nw * 4);
    std::vector<long long> h((size_t)G * lbm::kMaxWarps * 4);
    CK(cudaSetDevice(s.device));
    CK(cudaMemcpy(h.data(), s.prof, sizeof(long long) * h.size(), cudaMemcpyDeviceToHost));
    for (int b = 0; b < G; ++b)
        for (int w = 0; w < nw; ++w)
            for (int i = 0; i < 4; ++i) out[((size_t)b * nw + w) * 4 + i] = h[((size_t)b * lbm::kMaxWarps + w) * 4 + i];
    return 0;
}

int lbm_slab_rows(const lbm_ctx* c, int r, int* row_start, int* rows)
{
    if (!c || r < 0 || r >= c->size) return fail("lbm_slab_rows: bad slab index");
    return lbm_allocate_rows(c->prm.ny, c->size, r, row_start, rows);
}

// ---- engine 2: graph of `gs` timesteps (gs even so the buffer parity returns) ----
static int build_graph(lbm_ctx* c, Slab& s, int gs)
{
    CK(cudaSetDevice(s.device));
    if (s.graph) { cudaGraphExecDestroy(s.graph); s.graph = nullptr; }
    const int nw = 8, threads = nw * 32;
    const int grid = (s.plan.num_tasks + nw - 1) / nw;
    cudaGraph_t g;
    CK(cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal));
    int par = c->par;
    for (int i = 0; i < gs; ++i) {
        if (c->strict) lbm::lbm_step_kernel<true><<<grid, threads, 0, s.stream>>>(s.dev, par, s.step_ptr, i, 1);
        else lbm::lbm_step_kernel<false><<<grid, threads, 0, s.stream>>>(s.dev, par, s.step_ptr, i, 1);
        par ^= 1;
    }
    lbm::lbm_advance_step_kernel<<<1, 1, 0, s.stream>>>(s.step_ptr, gs);
    CK(cudaStreamEndCapture(s.stream, &g));
    CK(cudaGraphInstantiate(&s.graph, g, 0));
    CK(cudaGraphDestroy(g));
    s.graph_steps = gs;
    s.graph_par = c->par;
    return 0;
}

static int run_engine2(lbm_ctx* c, int iters)
{
    Slab& s = c->slabs[0];
    CK(cudaSetDevice(s.device));
    const int nw = 8, threads = nw * 32;
    const int grid = (s.plan.num_tasks + nw - 1) / nw;
    const int gs = 64;
    int launches = 0;
    CK(cudaEventRecord(s.ev_start, s.stream));
    if (s.dev.accel_lr >= 0) {
        lbm::lbm_accel_kernel<<<(s.plan.tpr + 3) / 4, 128, 0, s.stream>>>(s.dev, c->par);
        ++launches;
    }
    int remaining = iters;
    // all but the last step carry the accelerate epilogue for the following step
    while (remaining - 1 >= gs) {
        if (!s.graph || s.graph_par != c->par) { if (build_graph(c, s, gs)) return 1; }
        CK(cudaGraphLaunch(s.graph, s.stream));
        launches += gs + 1;
        remaining -= gs;
    }
    for (int i = 0; i < remaining; ++i) {
        const int epi = (i + 1 < remaining) ? 1 : 0;
        if (c->strict) lbm::lbm_step_kernel<true><<<grid, threads, 0, s.stream>>>(s.dev, c->par, s.step_ptr, i, epi);
        else lbm::lbm_step_kernel<false><<<grid, threads, 0, s.stream>>>(s.dev, c->par, s.step_ptr, i, epi);
        c->par ^= 1;
        ++launches;
    }
    if (remaining > 0) { lbm::lbm_advance_step_kernel<<<1, 1, 0, s.stream>>>(s.step_ptr, remaining); ++launches; }
    CK(cudaGetLastError());
    CK(cudaEventRecord(s.ev_stop, s.stream));
    c->last_launches = launches;
    return 0;
}

// engines 3 and 4: one cooperative launch per slab (cooperative only for the co-residency guarantee: every warp
// polls its neighbours' records)
static int run_dataflow(lbm_ctx* c, int iters)
{
    const bool ring = c->size > 1;
    int launches = 0;
    for (Slab& s : c->slabs) {
        CK(cudaSetDevice(s.device));
        CK(cudaEventRecord(s.ev_start, s.stream));
        int par0 = c->par, step0 = c->steps_done, it = iters;
        unsigned base = c->base;
        void* args[] = {(void*)&s.dev, (void*)&par0, (void*)&step0, (void*)&it, (void*)&base};
        if (c->engine == 4) {
            CK(cudaLaunchCooperativeKernel(row_fn(c->strict, ring), dim3(s.plan.rows), dim3(c->prm.nx), args, 0, s.stream));
        } else {
            const int blocks = (s.plan.num_tasks + c->df_k - 1) / c->df_k;
            CK(cudaLaunchCooperativeKernel(dataflow_fn(c->strict, c->df_k > 1, ring), dim3(blocks), dim3(32), args, 0, s.stream));
        }
        CK(cudaEventRecord(s.ev_stop, s.stream));
        ++launches;
    }
    c->base += (unsigned)iters + 1u;
    if (iters & 1) c->par ^= 1;
    c->last_launches = launches;
    return 0;
}

static int run_engine1(lbm_ctx* c, int iters)
{
    const void* fn = persistent_fn(c->strict, c->size > 1);
    int launches = 0;
    for (Slab& s : c->slabs) {
        CK(cudaSetDevice(s.device));
        CK(cudaEventRecord(s.ev_start, s.stream));
        int par0 = c->par, step0 = c->steps_done, it = iters;
        unsigned base = c->base;
        void* args[] = {(void*)&s.dev, (void*)&par0, (void*)&step0, (void*)&it, (void*)&base};
        CK(cudaLaunchCooperativeKernel(fn, dim3(s.plan.num_ctas), dim3(c->threads), args, c->stage_bytes, s.stream));
        CK(cudaEventRecord(s.ev_stop, s.stream));
        ++launches;
    }
    c->base += (unsigned)iters + 1u;
    if (iters & 1) c->par ^= 1;
    c->last_launches = launches;
    return 0;
}

int lbm_run(lbm_ctx* c, int iters)
{
    if (!c) return fail("ctx is NULL");
    if (iters < 0) return fail("iters < 0");
    if (iters == 0) { c->last_launches = 0; return 0; }
    if (c->steps_done + iters > c->prm.max_iters)
        return fail("lbm_run: %d + %d steps exceed max_iters=%d", c->steps_done, iters, c->prm.max_iters);
    if (c->size > 1 && !c->connected) return fail("lbm_run: ring neighbours not connected (call lbm_rank_connect)");
    const int rc = (c->engine == 1) ? run_engine1(c, iters) : (c->engine == 2) ? run_engine2(c, iters) : run_dataflow(c, iters);
    if (rc) return rc;
    c->steps_done += iters;
    return 0;
}

int lbm_sync(lbm_ctx* c)
{
    if (!c) return fail("ctx is NULL");
    for (Slab& s : c->slabs) {
        CK(cudaSetDevice(s.device));
        CK(cudaStreamSynchronize(s.stream));
        int st = 0;
        CK(cudaMemcpy(&st, s.status, sizeof(int), cudaMemcpyDeviceToHost));
        if (st != 0) return fail("device reported status %d on slab %d (2 = timed out waiting for a neighbouring CTA/GPU)",
                                 st, (int)(&s - &c->slabs[0]) + c->rank0);
        if (check_report(s.device)) return 1;
    }
    return 0;
}

int lbm_last_run_ms(lbm_ctx* c, float* ms)
{
    if (!c || !ms) return fail("bad arguments");
    float worst = 0.f;
    for (Slab& s : c->slabs) {
        CK(cudaSetDevice(s.device));
        CK(cudaEventSynchronize(s.ev_stop));
        float t = 0.f;
        CK(cudaEventElapsedTime(&t, s.ev_start, s.ev_stop));
        worst = std::max(worst, t);
    }
    *ms = c->last_ms = worst;
    return 0;
}

}  // extern "C"

// =========================================================================== results
namespace {

// Folds the persistent engine's per-CTA partials of steps [finalized, steps_done) into tot[].
int fold_partials(lbm_ctx* c, Slab& s)
{
    if (c->engine == 2 || s.finalized_steps >= c->steps_done) return 0;
    CK(cudaSetDevice(s.device));
    const int n = c->steps_done - s.finalized_steps;
    k_fold_partials<<<(n + 7) / 8, 256, 0, s.stream>>>(s.partials, s.tot, s.dev.num_ctas, s.finalized_steps, c->steps_done);
    CK(cudaGetLastError());
    s.finalized_steps = c->steps_done;
    return 0;
}

}  // namespace

extern "C" {

// Sum over this context's slabs of the per-step sums of |u| for steps [first, first + n), in double.
static int gather_tot_u(lbm_ctx* c, int first, int n, std::vector<double>& sum)
{
    if (lbm_sync(c)) return 1;
    sum.assign((size_t)n, 0.0);
    std::vector<double> tmp((size_t)n);
    for (Slab& s : c->slabs) {
        if (fold_partials(c, s)) return 1;
        CK(cudaMemcpyAsync(tmp.data(), s.tot + first, sizeof(double) * n, cudaMemcpyDeviceToHost, s.stream));
        c->d2h_bytes += (long long)sizeof(double) * n;
        CK(cudaStreamSynchronize(s.stream));
        for (int t = 0; t < n; ++t) sum[t] += tmp[t];           // the one reduction over slabs (:445)
    }
    return 0;
}

int lbm_get_av_vels(lbm_ctx* c, float* out, int n)
{
    if (!c || !out) return fail("bad arguments");
    if (n < 0 || n > c->steps_done) return fail("lbm_get_av_vels: n=%d but only %d steps done", n, c->steps_done);
    std::vector<double> sum;
    if (gather_tot_u(c, 0, n, sum)) return 1;
    if (c->rank_mode) {   // un-normalised local partial; the caller sums across ranks, then scales
        for (int t = 0; t < n; ++t) out[t] = (float)sum[t];
    } else {
        const float r = 1.f / (float)c->free_cells;              // :642
        for (int t = 0; t < n; ++t) out[t] = (float)sum[t] * r;  // :446
    }
    return 0;
}

int lbm_get_tot_u(lbm_ctx* c, float* out, int first, int n)
{
    if (!c || !out) return fail("bad arguments");
    if (first < 0 || n < 0 || first + n > c->steps_done)
        return fail("lbm_get_tot_u: steps [%d,%d) requested but only %d done", first, first + n, c->steps_done);
    std::vector<double> sum;
    if (gather_tot_u(c, first, n, sum)) return 1;
    for (int t = 0; t < n; ++t) out[t] = (float)sum[t];
    return 0;
}

int lbm_get_cells(lbm_ctx* c, float* const speeds[LBM_NSPEEDS])
{
    if (!c || !speeds) return fail("bad arguments");
    if (lbm_sync(c)) return 1;
    const int nx = c->prm.nx;
    for (Slab& s : c->slabs) {
        CK(cudaSetDevice(s.device));
        const size_t row0 = c->rank_mode ? 0 : (size_t)s.plan.row_start;
        std::vector<Piece> pieces;
        for (int k = 0; k < LBM_NSPEEDS; ++k)
            pieces.push_back({speeds[k] + row0 * nx, s.buf[c->par] + k * s.dev.ps + nx, sizeof(float) * (size_t)s.plan.rows * nx});
        if (download(s.device, s.stream, pieces)) return 1;
        c->d2h_bytes += (long long)LBM_NSPEEDS * sizeof(float) * s.plan.rows * nx;
    }
    return 0;
}

int lbm_set_cells(lbm_ctx* c, const float* const speeds[LBM_NSPEEDS])
{
    if (!c || !speeds) return fail("bad arguments");
    if (c->rank_mode) return fail("lbm_set_cells is not available in one-process-per-GPU mode");
    if (lbm_sync(c)) return 1;
    const int nx = c->prm.nx, ny = c->prm.ny;
    for (Slab& s : c->slabs) {
        CK(cudaSetDevice(s.device));
        for (int k = 0; k < LBM_NSPEEDS; ++k) {
            float* plane = s.buf[c->par] + k * s.dev.ps;
            const int below = (s.plan.row_start - 1 + ny) % ny, above = (s.plan.row_start + s.plan.rows) % ny;
            CK(cudaMemcpyAsync(plane, speeds[k] + (size_t)below * nx, sizeof(float) * nx, cudaMemcpyHostToDevice, s.stream));
            CK(cudaMemcpyAsync(plane + nx, speeds[k] + (size_t)s.plan.row_start * nx,
                               sizeof(float) * (size_t)s.plan.rows * nx, cudaMemcpyHostToDevice, s.stream));
            CK(cudaMemcpyAsync(plane + (size_t)(s.plan.rows + 1) * nx, speeds[k] + (size_t)above * nx, sizeof(float) * nx,
                               cudaMemcpyHostToDevice, s.stream));
            c->h2d_bytes += (long long)sizeof(float) * (s.plan.rows + 2) * nx;
        }
        CK(cudaStreamSynchronize(s.stream));
    }
    return 0;
}

int lbm_get_final_state(lbm_ctx* c, float* ux, float* uy, float* u, float* pressure)
{
    if (!c) return fail("ctx is NULL");
    if (lbm_sync(c)) return 1;
    const int nx = c->prm.nx;
    float* host[4] = {ux, uy, u, pressure};
    for (Slab& s : c->slabs) {
        CK(cudaSetDevice(s.device));
        const size_t n = (size_t)s.plan.rows * nx;
        if (!s.fields) CK(dmalloc(&s.fields, sizeof(float) * 4 * n));
        float* dv[4];
        for (int i = 0; i < 4; ++i) dv[i] = host[i] ? s.fields + i * n : nullptr;
        k_final_state<<<(unsigned)((n + 255) / 256), 256, 0, s.stream>>>(s.dev, c->par, c->prm.density, dv[0], dv[1], dv[2], dv[3]);
        CK(cudaGetLastError());
        const size_t row0 = c->rank_mode ? 0 : (size_t)s.plan.row_start;
        std::vector<Piece> pieces;
        for (int i = 0; i < 4; ++i)
            if (host[i]) {
                pieces.push_back({host[i] + row0 * nx, dv[i], sizeof(float) * n});
                c->d2h_bytes += (long long)sizeof(float) * n;
            }
        if (download(s.device, s.stream, pieces)) return 1;
    }
    return 0;
}

static int diagnostics(lbm_ctx* c, double* sum_u, double* sum_rho)
{
    if (lbm_sync(c)) return 1;
    *sum_u = *sum_rho = 0.0;
    for (Slab& s : c->slabs) {
        CK(cudaSetDevice(s.device));
        CK(cudaMemsetAsync(s.scal, 0, sizeof(double) * 2, s.stream));
        k_diagnostics<<<592, 256, 0, s.stream>>>(s.dev, c->par, s.scal);
        CK(cudaGetLastError());
        double h[2];
        CK(cudaMemcpyAsync(h, s.scal, sizeof(h), cudaMemcpyDeviceToHost, s.stream));
        CK(cudaStreamSynchronize(s.stream));
        *sum_u += h[0];
        *sum_rho += h[1];
    }
    return 0;
}

// ---- allocation-free diagnostics (lbm_diagnostics_noalloc) -------------------------------------------------------
// The same two sums as k_diagnostics (sum of |u| over free cells, sum of all populations), for callers that must not
// allocate: the scratch - one pair of doubles per lattice row, then the two totals - is a file-scope device array of
// fixed size instead of a cudaMalloc'ed buffer.  Every GPU that loads the library holds its own copy (256 KB) until exit.
// In HOST code the array's name is only a stand-in: the device address is taken once per device with
// cudaGetSymbolAddress, with that device current, and passed to the kernels.
constexpr int kNoAllocMaxRows = LBM_NOALLOC_MAX_ROWS;
static __device__ __align__(256) double g_noalloc_scratch[2 * kNoAllocMaxRows + 2];

// one block per row: scratch[2*row] = sum of |u| over the row's free cells, scratch[2*row + 1] = sum of its populations
static __global__ void k_diag_rows(SlabDev d, int par, double* scratch)
{
    __shared__ double sh[2][32];
    const int lr = blockIdx.x;
    double su = 0.0, sd = 0.0;
    for (int x = threadIdx.x; x < d.nx; x += blockDim.x) {
        const float* f = d.buf[par] + (long long)(lr + 1) * d.nx + x;
        float q[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) q[k] = f[k * d.ps];
        float rho = 0.f;
#pragma unroll
        for (int k = 0; k < 9; ++k) rho = __fadd_rn(rho, q[k]);
        sd += (double)rho;
        if (!cell_blocked(d, lr + 1, x)) {
            const float vx = __fdiv_rn(__fsub_rn(__fadd_rn(__fadd_rn(q[1], q[5]), q[8]), __fadd_rn(__fadd_rn(q[3], q[6]), q[7])), rho);
            const float vy = __fdiv_rn(__fsub_rn(__fadd_rn(__fadd_rn(q[2], q[5]), q[6]), __fadd_rn(__fadd_rn(q[4], q[7]), q[8])), rho);
            su += (double)__fsqrt_rn(__fadd_rn(__fmul_rn(vx, vx), __fmul_rn(vy, vy)));
        }
    }
    su = lbm::warp_sum(su);
    sd = lbm::warp_sum(sd);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { sh[0][warp] = su; sh[1][warp] = sd; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { a += sh[0][w]; b += sh[1][w]; }
        scratch[2 * lr] = a;
        scratch[2 * lr + 1] = b;
    }
}

// one block: folds the `rows` pairs in a fixed order (deterministic) into scratch[2*kNoAllocMaxRows], [.. + 1]
static __global__ void k_diag_fold(double* scratch, int rows)
{
    __shared__ double sh[2][256];
    double a = 0.0, b = 0.0;
    for (int r = threadIdx.x; r < rows; r += blockDim.x) { a += scratch[2 * r]; b += scratch[2 * r + 1]; }
    sh[0][threadIdx.x] = a;
    sh[1][threadIdx.x] = b;
    __syncthreads();
    for (int o = blockDim.x >> 1; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { sh[0][threadIdx.x] += sh[0][threadIdx.x + o]; sh[1][threadIdx.x] += sh[1][threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { scratch[2 * kNoAllocMaxRows] = sh[0][0]; scratch[2 * kNoAllocMaxRows + 1] = sh[1][0]; }
}

// device address of g_noalloc_scratch on `device` (which the caller has made current), looked up once per device
static int noalloc_scratch(int device, double** out)
{
    static std::mutex mu;
    static double* ptr[LBM_MAX_GPUS * 8] = {nullptr};
    if (device < 0 || device >= (int)(sizeof(ptr) / sizeof(ptr[0]))) return fail("lbm_diagnostics_noalloc: device ordinal %d out of range", device);
    std::lock_guard<std::mutex> g(mu);
    if (!ptr[device]) {
        double* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return fail("lbm_diagnostics_noalloc: cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", device, cudaGetErrorString(e));
        ptr[device] = p;
    }
    *out = ptr[device];
    return 0;
}

int lbm_diagnostics_noalloc(lbm_ctx* c, float* av_velocity, float* total_density)
{
    if (!c || (!av_velocity && !total_density)) return fail("lbm_diagnostics_noalloc: bad arguments");
    for (const Slab& s : c->slabs)
        if (s.plan.rows > kNoAllocMaxRows)
            return fail("lbm_diagnostics_noalloc: a slab of %d rows exceeds the fixed scratch of LBM_NOALLOC_MAX_ROWS = %d rows per GPU "
                        "(a static 256 KB device array; use lbm_av_velocity / lbm_total_density, which allocate, or more GPUs)",
                        s.plan.rows, kNoAllocMaxRows);
    if (lbm_sync(c)) return 1;
    double su = 0.0, sr = 0.0;
    for (Slab& s : c->slabs) {
        CK(cudaSetDevice(s.device));
        double* scratch = nullptr;
        if (noalloc_scratch(s.device, &scratch)) return 1;
        // the array is shared by every call on this device and keeps the last call's contents: the entries this call
        // reads are all rewritten by k_diag_rows first; the totals are cleared so that a failed launch cannot leave the
        // previous call's result to be read back
        CK(cudaMemsetAsync(scratch + 2 * kNoAllocMaxRows, 0, 2 * sizeof(double), s.stream));
        k_diag_rows<<<s.plan.rows, 256, 0, s.stream>>>(s.dev, c->par, scratch);
        k_diag_fold<<<1, 256, 0, s.stream>>>(scratch, s.plan.rows);
        CK(cudaGetLastError());
        double h[2];
        CK(cudaMemcpyAsync(h, scratch + 2 * kNoAllocMaxRows, sizeof(h), cudaMemcpyDeviceToHost, s.stream));
        CK(cudaStreamSynchronize(s.stream));
        su += h[0];
        sr += h[1];
    }
    const int cells = c->prm.nx * c->prm.ny - c->blocked_cells;               // as lbm_av_velocity (d2q9-bgk.c:480)
    if (av_velocity) *av_velocity = c->rank_mode ? (float)su : (float)su / (float)cells;
    if (total_density) *total_density = (float)sr;
    return 0;
}

int lbm_av_velocity(lbm_ctx* c, float* out)
{
    if (!c || !out) return fail("bad arguments");
    double su, sr;
    if (diagnostics(c, &su, &sr)) return 1;
    // d2q9-bgk.c:480: tot_u / (float)tot_cells with the TRUE number of unblocked cells
    const int cells = c->prm.nx * c->prm.ny - c->blocked_cells;
    *out = c->rank_mode ? (float)su : (float)su / (float)cells;
    return 0;
}

int lbm_total_density(lbm_ctx* c, float* out)
{
    if (!c || !out) return fail("bad arguments");
    double su, sr;
    if (diagnostics(c, &su, &sr)) return 1;
    *out = (float)sr;
    return 0;
}

// =========================================================================== rank mode
// Blob layout: magic, rank, size, nx, then the IPC handle of the slab's inbound halo slots.
struct IpcBlob {
    unsigned magic;
    int rank, size, nx, rows, engine;
    char bus_id[20];               // PCI bus id of the exporting rank's GPU (lets the importer vet the link)
    cudaIpcMemHandle_t ll_in;      // engine 1: inbound halo slots
    cudaIpcMemHandle_t rec;        // engines 3 / 4: the record lattice (its two haloed rows are written by the neighbours)
};
static_assert(sizeof(IpcBlob) <= LBM_IPC_BLOB_BYTES, "IPC blob too large");

int lbm_create_rank(lbm_ctx** out, const lbm_params* params, const int* obstacles, int rank, int size, int device,
                    const lbm_options* opts)
{
    lbm_ctx* c = nullptr;
    if (size > 1) g_host_sharers = std::min(size, std::max(1, lbm_device_count()));   // ranks of this node set up side by side
    if (create_common(out, params, obstacles, opts, &c)) return 1;
    int rc = 0;
    if (size < 1 || size > params->ny || rank < 0 || rank >= size) rc = fail("bad rank %d / size %d", rank, size);
    if (!rc && (device < 0 || device >= lbm_device_count())) rc = fail("bad device %d", device);
    if (!rc) {
        c->size = size; c->nslabs = 1; c->rank0 = rank; c->rank_mode = true;
        c->slabs.resize(1);
        const auto t0 = std::chrono::steady_clock::now();
        cudaSetDevice(device);
        rc = configure_engine(c, device);
        if (env_int("LBM_REPORT", 0) > 1)
            fprintf(stderr, "create: engine set-up %.4f s\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    }
    if (!rc) rc = slab_create(c, c->slabs[0], rank, device, obstacles);
    if (rc) { lbm_destroy(c); return 1; }
    c->connected = (size == 1);
    describe(c);
    *out = c;
    return 0;
}

int lbm_rank_export(lbm_ctx* c, void* blob)
{
    if (!c || !blob || !c->rank_mode) return fail("lbm_rank_export: needs a context made by lbm_create_rank");
    Slab& s = c->slabs[0];
    CK(cudaSetDevice(s.device));
    IpcBlob b;
    memset(&b, 0, sizeof(b));
    b.magic = 0x4C424D32u;  // "LBM2"
    b.rank = c->rank0; b.size = c->size; b.nx = c->prm.nx; b.rows = s.plan.rows; b.engine = c->engine;
    CK(cudaDeviceGetPCIBusId(b.bus_id, (int)sizeof(b.bus_id), s.device));
    if (c->size > 1) { CK(cudaIpcGetMemHandle(&b.ll_in, s.ll_in)); mark_exported(s.ll_in); }
    if (c->size > 1 && s.rec) { CK(cudaIpcGetMemHandle(&b.rec, s.rec)); mark_exported(s.rec); }
    memset(blob, 0, LBM_IPC_BLOB_BYTES);
    memcpy(blob, &b, sizeof(b));
    return 0;
}

int lbm_rank_connect(lbm_ctx* c, const void* blob_up, const void* blob_down)
{
    if (!c || !c->rank_mode) return fail("lbm_rank_connect: needs a context made by lbm_create_rank");
    if (c->size == 1) { c->connected = true; return 0; }
    if (!blob_up || !blob_down) return fail("lbm_rank_connect: NULL blob");
    Slab& s = c->slabs[0];
    CK(cudaSetDevice(s.device));
    IpcBlob up, dn;
    memcpy(&up, blob_up, sizeof(up));
    memcpy(&dn, blob_down, sizeof(dn));
    const int r_up = (c->rank0 + 1) % c->size, r_dn = (c->rank0 - 1 + c->size) % c->size;
    if (up.magic != 0x4C424D32u || dn.magic != 0x4C424D32u) return fail("lbm_rank_connect: bad blob magic");
    if (up.rank != r_up || dn.rank != r_dn || up.size != c->size || dn.size != c->size)
        return fail("lbm_rank_connect: blobs are from ranks %d/%d, expected up=%d down=%d", up.rank, dn.rank, r_up, r_dn);
    if (up.nx != c->prm.nx || dn.nx != c->prm.nx) return fail("lbm_rank_connect: neighbour lattice width differs");
    if (up.engine != c->engine || dn.engine != c->engine)
        return fail("lbm_rank_connect: neighbours run engines %d / %d, this rank engine %d", up.engine, dn.engine, c->engine);
    for (const IpcBlob* o : {&up, &dn}) {
        int peer = -1;
        char id[sizeof(o->bus_id) + 1] = {0};
        memcpy(id, o->bus_id, sizeof(o->bus_id));
        if (cudaDeviceGetByPCIBusId(&peer, id) != cudaSuccess) { cudaGetLastError(); continue; }   // not visible here: cannot vet
        if (peer_link_unsuitable(s.device, peer)) return 1;
    }
    const bool same = (r_up == r_dn);   // size == 2: both neighbours are the same process
    CK(ipc_open_cached(&s.ipc_ll_up, up.ll_in));
    if (same) s.ipc_ll_down = s.ipc_ll_up;
    else CK(ipc_open_cached(&s.ipc_ll_down, dn.ll_in));
    s.dev.ll_up = (uint4*)s.ipc_ll_up;
    s.dev.ll_down = (uint4*)s.ipc_ll_down;
    if (s.rec) {
        CK(ipc_open_cached(&s.ipc_rec_up, up.rec));
        if (same) s.ipc_rec_down = s.ipc_rec_up;
        else CK(ipc_open_cached(&s.ipc_rec_down, dn.rec));
        s.dev.rec_up = (uint4*)s.ipc_rec_up;     s.dev.rows_up = up.rows;
        s.dev.rec_down = (uint4*)s.ipc_rec_down; s.dev.rows_down = dn.rows;
        check_add(s.device, &s, s.ipc_rec_up, sizeof(uint4) * 2 * 3 * (size_t)(up.rows + 2) * c->prm.nx);
        check_add(s.device, &s, s.ipc_rec_down, sizeof(uint4) * 2 * 3 * (size_t)(dn.rows + 2) * c->prm.nx);
    }
    check_add(s.device, &s, s.ipc_ll_up, sizeof(uint4) * 4 * (size_t)c->prm.nx);
    check_add(s.device, &s, s.ipc_ll_down, sizeof(uint4) * 4 * (size_t)c->prm.nx);
    if (check_upload(s.device)) return 1;
    c->connected = true;
    return 0;
}

int lbm_rank_disconnect(lbm_ctx* c)
{
    if (!c || !c->rank_mode) return fail("lbm_rank_disconnect: needs a context made by lbm_create_rank");
    if (c->size == 1 || !c->connected) return 0;
    if (lbm_sync(c)) return 1;                       // my kernel may still be pushing into the mappings
    Slab& s = c->slabs[0];
    CK(cudaSetDevice(s.device));
    // the mappings themselves stay open in the process-wide cache (ipc_open_cached); lbm_trim_cache closes them
    s.ipc_ll_up = s.ipc_ll_down = s.ipc_rec_up = s.ipc_rec_down = nullptr;
    s.dev.ll_up = s.dev.ll_down = s.dev.rec_up = s.dev.rec_down = nullptr;
    c->connected = false;
    return 0;
}

}  // extern "C"
