// hop_probe.cu - what does ONE producer->consumer hop between two SMs cost on B200, by transport?
// Two warps (different CTAs, hence different SMs) play ping-pong with a 16-byte tagged record: A writes tag t,
// B polls until it sees t and writes t into A's slot, A polls ... `rounds` times.  Reported: SM cycles per ONE-WAY
// hop (round trip / 2), measured with clock64 on A.  Transports:
//   0  global memory, st/ld.relaxed.sys.b128   (the product's tagged records)
//   1  global memory, st/ld.relaxed.gpu.b128
//   2  global memory, st.global.cg.v4 / ld.volatile.v4 (round-1 form)
//   3  thread-block cluster of 2 CTAs: st.shared::cluster.v4 into the peer's shared memory, peer polls its OWN
//      shared memory with ld.volatile.shared.v4
//   4  cluster: st.async + mbarrier complete_tx in the peer's shared memory, peer waits on its own mbarrier (try_wait)
// Also: the pure compute chain of a D2Q9 cell update (for scale) is measured by the library itself.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hop_probe hop_probe.cu ; run: ./hop_probe
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
namespace cg = cooperative_groups;

__device__ __forceinline__ void st_sys(uint4* p, uint4 v) {
    unsigned long long lo = ((unsigned long long)v.y << 32) | v.x, hi = ((unsigned long long)v.w << 32) | v.z;
    asm volatile("{.reg .b128 _v; mov.b128 _v, {%1,%2}; st.relaxed.sys.global.b128 [%0], _v;}" ::"l"(p), "l"(lo), "l"(hi) : "memory");
}
__device__ __forceinline__ uint4 ld_sys(const uint4* p) {
    unsigned long long lo, hi;
    asm volatile("{.reg .b128 _d; ld.relaxed.sys.global.b128 _d, [%2]; mov.b128 {%0,%1}, _d;}" : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
    return make_uint4((unsigned)lo, (unsigned)(lo >> 32), (unsigned)hi, (unsigned)(hi >> 32));
}
__device__ __forceinline__ void st_gpu(uint4* p, uint4 v) {
    unsigned long long lo = ((unsigned long long)v.y << 32) | v.x, hi = ((unsigned long long)v.w << 32) | v.z;
    asm volatile("{.reg .b128 _v; mov.b128 _v, {%1,%2}; st.relaxed.gpu.global.b128 [%0], _v;}" ::"l"(p), "l"(lo), "l"(hi) : "memory");
}
__device__ __forceinline__ uint4 ld_gpu(const uint4* p) {
    unsigned long long lo, hi;
    asm volatile("{.reg .b128 _d; ld.relaxed.gpu.global.b128 _d, [%2]; mov.b128 {%0,%1}, _d;}" : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
    return make_uint4((unsigned)lo, (unsigned)(lo >> 32), (unsigned)hi, (unsigned)(hi >> 32));
}
__device__ __forceinline__ void st_cg(uint4* p, uint4 v) {
    asm volatile("st.global.cg.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_vol(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

// slots[0] is polled by block 0 (written by block 1), slots[64] by block 1 (written by block 0); every lane of the
// warp has its own record (512 B per direction, as a warp-task's record plane), `nrec` planes per message.
template <int MODE>
__global__ void k_global(uint4* slots, int rounds, int nrec, long long* out, int other)
{
    const int lane = threadIdx.x, me = blockIdx.x;
    if (me != 0 && me != other) return;
    const int role = me == 0 ? 0 : 1;
    uint4* mine = slots + role * 4096 + lane;
    uint4* theirs = slots + (role ^ 1) * 4096 + lane;
    long long t0 = 0;
    for (int r = 1; r <= rounds; ++r) {
        if (r == 9 && role == 0) t0 = clock64();
        if (role == 0) {
            for (int k = 0; k < nrec; ++k) {
                const uint4 v = make_uint4(r, r, r, r);
                if (MODE == 0) st_sys(theirs + 32 * k, v); else if (MODE == 1) st_gpu(theirs + 32 * k, v); else st_cg(theirs + 32 * k, v);
            }
        }
        for (;;) {
            bool ok = true;
            for (int k = 0; k < nrec; ++k) {
                const uint4 v = MODE == 0 ? ld_sys(mine + 32 * k) : MODE == 1 ? ld_gpu(mine + 32 * k) : ld_vol(mine + 32 * k);
                ok = ok && v.w == (unsigned)r;
            }
            if (__all_sync(0xffffffffu, ok)) break;
        }
        if (role == 1) {
            for (int k = 0; k < nrec; ++k) {
                const uint4 v = make_uint4(r, r, r, r);
                if (MODE == 0) st_sys(theirs + 32 * k, v); else if (MODE == 1) st_gpu(theirs + 32 * k, v); else st_cg(theirs + 32 * k, v);
            }
        }
    }
    if (role == 0 && lane == 0) out[0] = (clock64() - t0) / (2LL * (rounds - 8));
}

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned mapa(unsigned addr, unsigned rank) {
    unsigned r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}

// cluster of 2 CTAs, one warp each: plain remote shared-memory stores, local polling
__global__ void __cluster_dims__(2, 1, 1) k_dsmem(int rounds, int nrec, long long* out)
{
    __shared__ __align__(16) uint4 box[32 * 8];
    cg::cluster_group cl = cg::this_cluster();
    const int lane = threadIdx.x, role = cl.block_rank();
    for (int k = 0; k < 8; ++k) box[lane + 32 * k] = make_uint4(0, 0, 0, 0);
    cl.sync();
    const unsigned remote = mapa(smem_u32(box + lane), role ^ 1);
    long long t0 = 0;
    for (int r = 1; r <= rounds; ++r) {
        if (r == 9 && role == 0) t0 = clock64();
        if (role == 0)
            for (int k = 0; k < nrec; ++k)
                asm volatile("st.shared::cluster.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(remote + 512 * k), "r"(r) : "memory");
        for (;;) {
            bool ok = true;
            for (int k = 0; k < nrec; ++k) {
                uint4 v;
                asm volatile("ld.volatile.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(box + lane + 32 * k)) : "memory");
                ok = ok && v.w == (unsigned)r;
            }
            if (__all_sync(0xffffffffu, ok)) break;
        }
        if (role == 1)
            for (int k = 0; k < nrec; ++k)
                asm volatile("st.shared::cluster.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(remote + 512 * k), "r"(r) : "memory");
    }
    if (role == 0 && lane == 0) out[0] = (clock64() - t0) / (2LL * (rounds - 8));
    cl.sync();
}

// cluster of 2: st.async (data + complete_tx on the peer's mbarrier), the peer sleeps in mbarrier.try_wait
__global__ void __cluster_dims__(2, 1, 1) k_stasync(int rounds, int nrec, long long* out)
{
    __shared__ __align__(16) uint4 box[32 * 8];
    __shared__ __align__(8) unsigned long long bar;
    cg::cluster_group cl = cg::this_cluster();
    const int lane = threadIdx.x, role = cl.block_rank();
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    cl.sync();
    const unsigned remote = mapa(smem_u32(box + lane), role ^ 1), rbar = mapa(smem_u32(&bar), role ^ 1);
    const unsigned bytes = 512u * nrec;
    long long t0 = 0;
    unsigned phase = 0;
    if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
    __syncwarp();
    for (int r = 1; r <= rounds; ++r) {
        if (r == 9 && role == 0) t0 = clock64();
        if (role == 0)
            for (int k = 0; k < nrec; ++k)
                asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1,%1,%1,%1}, [%2];" ::"r"(remote + 512 * k), "r"(r), "r"(rbar) : "memory");
        // wait for my inbox
        unsigned done = 0;
        while (!done)
            asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p;}" : "=r"(done) : "r"(smem_u32(&bar)), "r"(phase) : "memory");
        phase ^= 1;
        if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
        __syncwarp();
        volatile uint4* b = box;
        unsigned s = 0;
        for (int k = 0; k < nrec; ++k) s += b[lane + 32 * k].w;
        if (s == 0xdeadbeef) out[1] = s;
        if (role == 1)
            for (int k = 0; k < nrec; ++k)
                asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1,%1,%1,%1}, [%2];" ::"r"(remote + 512 * k), "r"(r), "r"(rbar) : "memory");
    }
    if (role == 0 && lane == 0) out[0] = (clock64() - t0) / (2LL * (rounds - 8));
    cl.sync();
}

int main()
{
    uint4* slots; long long* out; long long h;
    cudaMalloc(&slots, 2 * 4096 * sizeof(uint4)); cudaMalloc(&out, 16);
    const int rounds = 20000;
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("SM clock (attr) %d kHz\n", clk);
    const char* names[] = {"global st/ld.relaxed.sys.b128", "global st/ld.relaxed.gpu.b128", "global st.cg.v4 / ld.volatile.v4"};
    for (int nrec = 1; nrec <= 4; nrec *= 4)
      for (int other : {1, 2, 37, 74, 111, 147}) {
        for (int mode = 0; mode < 3; ++mode) {
            cudaMemset(slots, 0, 2 * 4096 * sizeof(uint4));
            void* args[] = {&slots, (void*)&rounds, &nrec, &out, &other};
            const void* fn = mode == 0 ? (const void*)k_global<0> : mode == 1 ? (const void*)k_global<1> : (const void*)k_global<2>;
            cudaError_t e = cudaLaunchCooperativeKernel(fn, dim3(148), dim3(32), args, 0, 0);
            if (e != cudaSuccess) { printf("launch: %s\n", cudaGetErrorString(e)); return 1; }
            e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("sync: %s\n", cudaGetErrorString(e)); return 1; }
            cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
            printf("%-36s block 0 <-> block %3d, %d x 512 B per message: %5lld cycles per one-way hop\n", names[mode], other, nrec, h);
        }
      }
    for (int nrec = 1; nrec <= 4; nrec *= 4) {
        k_dsmem<<<2, 32>>>(rounds, nrec, out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("dsmem: %s\n", cudaGetErrorString(e)); return 1; }
        cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
        printf("%-36s cluster pair, %d x 512 B per message: %5lld cycles per one-way hop\n", "DSMEM st.shared::cluster + local poll", nrec, h);
        k_stasync<<<2, 32>>>(rounds, nrec, out);
        e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("st.async: %s\n", cudaGetErrorString(e)); return 1; }
        cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
        printf("%-36s cluster pair, %d x 512 B per message: %5lld cycles per one-way hop\n", "DSMEM st.async + mbarrier try_wait", nrec, h);
    }
    return 0;
}
