// Micro-benchmark: cost of the warp primitives a radix-sort ranking loop can be built from (B200, sm_100a).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/match_atoms tools/ubench/match_atoms.cu && /tmp/match_atoms
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE> __global__ void k(uint32_t* out, long long* cyc, int iters, uint32_t seed)
{
    __shared__ uint32_t sh[32][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = lane; i < 256; i += 32) sh[warp][i] = 0;
    __syncthreads();
    uint32_t x = seed * 2654435761u + threadIdx.x * 40503u + blockIdx.x;
    uint32_t acc = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        x = x * 1664525u + 1013904223u;
        uint32_t d = (x >> 24) & 255u;           // random digit
        if (MODE == 1 || MODE == 11) d = (x >> 29);      // 8 distinct
        if (MODE == 2 || MODE == 12) d = it & 255u;      // uniform
        if (MODE < 10) {                          // match.any + popc
            const unsigned peers = __match_any_sync(0xffffffffu, d);
            acc += __popc(peers & ((1u << lane) - 1u));
        } else if (MODE < 20) {                   // shared atomicAdd
            acc += atomicAdd(&sh[warp][d], 1u);
        } else if (MODE == 20) {                  // 8 ballots
            unsigned peers = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 8; b++) {
                const unsigned v = __ballot_sync(0xffffffffu, (d >> b) & 1u);
                peers &= ((d >> b) & 1u) ? v : ~v;
            }
            acc += __popc(peers & ((1u << lane) - 1u));
        } else if (MODE == 21) {                  // match + leader LDS/STS (non-atomic per-warp counting)
            const unsigned peers = __match_any_sync(0xffffffffu, d);
            const uint32_t b = sh[warp][d];
            __syncwarp();
            if (lane == 31 - __clz(peers)) sh[warp][d] = b + __popc(peers);
            __syncwarp();
            acc += b + __popc(peers & ((1u << lane) - 1u));
        } else if (MODE == 22) {                  // red.shared (no return)
            asm volatile("red.shared.add.u32 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&sh[warp][d])) : "memory");
        } else if (MODE == 23) {                  // redux min/max uniform test
            acc += __reduce_min_sync(0xffffffffu, d) == __reduce_max_sync(0xffffffffu, d);
        }
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + sh[warp][lane];
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE> void run(const char* name)
{
    uint32_t* out;
    long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4);
    cudaMalloc(&cyc, 8);
    for (int threads : {32, 256, 1024}) {
        const int iters = 4096;
        k<MODE><<<148, threads>>>(out, cyc, iters, 1);
        k<MODE><<<148, threads>>>(out, cyc, iters, 2);
        cudaDeviceSynchronize();
        long long h = 0;
        cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("%-34s %2d warps/SM: %7.1f cycles per iteration per warp, %6.2f cycles per warp-op per SM\n", name, threads / 32,
               (double)h / iters, (double)h / iters / (threads / 32));
    }
    cudaFree(out);
    cudaFree(cyc);
}

int main()
{
    run<0>("match.any random digit");
    run<1>("match.any 8 distinct");
    run<2>("match.any uniform");
    run<10>("atomicAdd.shared random (ret)");
    run<11>("atomicAdd.shared 8 distinct (ret)");
    run<12>("atomicAdd.shared uniform (ret)");
    run<22>("red.shared random");
    run<20>("8 ballots");
    run<21>("match + leader LDS/STS");
    run<23>("redux min==max");
    return 0;
}
