// hop_home.cu - does the one-way hop between two SMs depend on WHERE the record lives (which L2 half is its home)?
// Ping-pong as in hop_probe.cu between two blocks, both mailboxes inside the same 4 KB granule g of a buffer (512 B
// apart), for 16 granules and several pairs of blocks; next to it the plain (dependent) L2 load latency from each of
// the two SMs to that granule, which tells on which die the granule is homed relative to the SM.
#include <cuda_runtime.h>
#include <cstdio>
__device__ __forceinline__ void st_rec(uint4* p, unsigned r) {
    unsigned long long v = ((unsigned long long)r << 32) | r;
    asm volatile("{.reg .b128 _v; mov.b128 _v, {%1,%1}; st.volatile.global.b128 [%0], _v;}" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint4 ld_rec(const uint4* p) {
    unsigned long long lo, hi;
    asm volatile("{.reg .b128 _d; ld.volatile.global.b128 _d, [%2]; mov.b128 {%0,%1}, _d;}" : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
    return make_uint4((unsigned)lo, (unsigned)(lo >> 32), (unsigned)hi, (unsigned)(hi >> 32));
}
__global__ void k(uint4* buf, int rounds, int a, int b, int g, long long* out, int* smids)
{
    const int lane = threadIdx.x, me = blockIdx.x;
    if (lane == 0) { unsigned s; asm("mov.u32 %0, %%smid;" : "=r"(s)); smids[me] = (int)s; }
    if (me != a && me != b) return;
    uint4* A = buf + g * 256 + lane;           // written by a, polled by b
    uint4* B = buf + g * 256 + 32 + lane;      // written by b, polled by a
    // plain dependent-load latency to this granule first (the buffer is zero: the next address depends on the value)
    if (lane == 0) {
        const uint4* p = buf + g * 256 + 64;
        const long long t0 = clock64();
        for (int i = 0; i < 64; ++i) p += ld_rec(p).x;
        out[me == a ? 1 : 2] = (clock64() - t0) / 64 + (p - (buf + g * 256 + 64));
    }
    __syncwarp();
    long long t0 = 0;
    for (int r = 1; r <= rounds; ++r) {
        if (r == 9 && me == a) t0 = clock64();
        if (me == a) { st_rec(A, r); while (!__all_sync(0xffffffffu, ld_rec(B).w == (unsigned)r)) {} }
        else { while (!__all_sync(0xffffffffu, ld_rec(A).w == (unsigned)r)) {} st_rec(B, r); }
    }
    if (me == a && lane == 0) out[0] = (clock64() - t0) / (2LL * (rounds - 8));
}
int main()
{
    uint4* buf; long long* out; int* smids; long long h[3]; int hs[148];
    cudaMalloc(&buf, 1 << 20); cudaMalloc(&out, 24); cudaMalloc(&smids, 148 * 4);
    const int rounds = 5000, G = 16;
    const int pairs[][2] = {{0, 1}, {0, 37}, {37, 0}, {37, 38}, {1, 37}, {0, 74}, {37, 111}, {2, 39}};
    for (auto& pr : pairs) {
        int a = pr[0], b = pr[1];
        long long hop[G], la[G], lb[G];
        for (int g = 0; g < G; ++g) {
            cudaMemset(buf, 0, 1 << 20);
            void* args[] = {&buf, (void*)&rounds, &a, &b, &g, &out, &smids};
            if (cudaLaunchCooperativeKernel((const void*)k, dim3(148), dim3(32), args, 0, 0) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 1; }
            cudaMemcpy(h, out, 24, cudaMemcpyDeviceToHost);
            hop[g] = h[0]; la[g] = h[1]; lb[g] = h[2];
        }
        cudaMemcpy(hs, smids, sizeof(hs), cudaMemcpyDeviceToHost);
        printf("blocks %3d <-> %3d (SM %3d <-> %3d)\n  one-way hop per granule:", a, b, hs[a], hs[b]);
        for (int g = 0; g < G; ++g) printf(" %5lld", hop[g]);
        printf("\n  load latency from SM %3d:", hs[a]); for (int g = 0; g < G; ++g) printf(" %5lld", la[g]);
        printf("\n  load latency from SM %3d:", hs[b]); for (int g = 0; g < G; ++g) printf(" %5lld", lb[g]);
        printf("\n");
    }
    return 0;
}
