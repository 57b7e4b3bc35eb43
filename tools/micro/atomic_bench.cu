// Micro-benchmark (development aid): throughput of global reductions on B200 for the address
// patterns the backward scatter produces.  nvcc -arch=sm_100a -O3 -o atomic_bench atomic_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t hash32(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

// each lane adds to element idx = window_base(warp) + hash % window
template <int MODE>
__global__ void k_red(float *buf, uint32_t nelem, uint32_t window, int iters)
{
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t warp = tid >> 5;
    for (int it = 0; it < iters; ++it) {
        uint32_t h = hash32(tid * 9781u + it * 7919u);
        uint32_t base = hash32(warp * 31u + it) % (nelem - window);
        uint32_t idx = base + h % window;
        if (MODE == 0) atomicAdd(buf + 8 * (size_t)idx, 1.0f);
        if (MODE == 1) atomicAdd(reinterpret_cast<float2 *>(buf + 8 * (size_t)idx), make_float2(1.f, 2.f));
        if (MODE == 2) atomicAdd(reinterpret_cast<float4 *>(buf + 8 * (size_t)idx), make_float4(1.f, 2.f, 3.f, 4.f));
        if (MODE == 3) { atomicAdd(reinterpret_cast<float4 *>(buf + 8 * (size_t)idx), make_float4(1.f, 2.f, 3.f, 4.f));
                         atomicAdd(reinterpret_cast<float4 *>(buf + 8 * (size_t)idx + 4), make_float4(1.f, 2.f, 3.f, 4.f)); }
        if (MODE == 4) { float *q = buf + 8 * (size_t)idx; atomicAdd(reinterpret_cast<float4 *>(q), make_float4(1.f, 2.f, 3.f, 4.f));
                         atomicAdd(q + 4, 1.f); atomicAdd(q + 5, 1.f); atomicAdd(q + 6, 1.f); }
        if (MODE == 5) atomicMin(reinterpret_cast<unsigned long long *>(buf) + 4 * (size_t)idx, (unsigned long long)h << 20);
        if (MODE == 6) { unsigned m = __match_any_sync(0xFFFFFFFFu, idx); if (m == 0x12345u) buf[0] = 1.f; }
    }
}

template <int MODE>
void run(const char *name, float *buf, uint32_t nelem, uint32_t window, int lanes_per_op)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 64, grid = 148 * 16, block = 256;
    k_red<MODE><<<grid, block>>>(buf, nelem, window, 4);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k_red<MODE><<<grid, block>>>(buf, nelem, window, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double ops = (double)grid * block * iters;
    printf("%-28s window %8u: %7.3f ms  %8.2f Gops/s (lane-ops; x%d instr)  err=%s\n", name, window, ms, ops / ms / 1e6, lanes_per_op,
           cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    const uint32_t nelem = 4u << 20;   // 4M slots x 32 B = 128 MB
    float *buf; cudaMalloc(&buf, (size_t)nelem * 32); cudaMemset(buf, 0, (size_t)nelem * 32);
    for (uint32_t window : {64u, 4096u, 2000000u}) {
        run<0>("red.f32", buf, nelem, window, 1);
        run<1>("red.v2.f32", buf, nelem, window, 1);
        run<2>("red.v4.f32", buf, nelem, window, 1);
        run<3>("2 x red.v4.f32 (32 B slot)", buf, nelem, window, 2);
        run<4>("red.v4 + 3 x red.f32", buf, nelem, window, 4);
        run<5>("atomicMin.u64", buf, nelem, window, 1);
        run<6>("match.any", buf, nelem, window, 1);
    }
    return 0;
}
