// Micro-benchmark for the next round: what does ONE `red.global.add.v4.f32` cost on sm_100a as a function of how its
// active lanes are laid out?  The splat kernel's reductions are L2-resident; its "single" cells (few active lanes,
// no shared sector) were measured to cost as much as the two main cells of every pixel pair
// (profiles/r1_v4_notes.md), so the cost model per instruction / per lane / per 32-byte sector decides the design.
// Every warp issues ITER reductions into a 32 MB (L2-resident) float4 buffer; the pattern picks the lanes and cells.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o micro_red_patterns micro_red_patterns.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n",cudaGetErrorString(e),__LINE__); exit(1);} }while(0)

constexpr int ITER = 2048;
constexpr size_t CELLS = (size_t)2 << 20;          // 2 Mi float4 cells = 32 MB

struct Pattern { const char* name; int stride; unsigned lane_mask; int with_scalar; };

__global__ void __launch_bounds__(128) k_red(float4* acc, float* accw, int stride, unsigned lane_mask, int with_scalar) {
    const int lane = threadIdx.x & 31;
    const size_t warp = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const bool on = (lane_mask >> lane) & 1u;
    size_t base = (warp * 4099u) % CELLS;
    for (int i = 0; i < ITER; ++i) {
        const size_t cell = (base + (size_t)lane * stride) % CELLS;
        const float v = (float)(i & 7);
        if (on) {
            asm volatile("red.global.add.v4.f32 [%0], {%1, %1, %1, %1};" :: "l"(acc + cell), "f"(v) : "memory");
            if (with_scalar) asm volatile("red.global.add.f32 [%0], %1;" :: "l"(accw + cell), "f"(v) : "memory");
        }
        base = (base + 32u * stride + 977u) % CELLS;   // move on: no same-address serialisation between iterations
    }
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    float4* acc; float* accw;
    CK(cudaMalloc(&acc, CELLS * sizeof(float4))); CK(cudaMalloc(&accw, CELLS * sizeof(float)));
    CK(cudaMemset(acc, 0, CELLS * sizeof(float4))); CK(cudaMemset(accw, 0, CELLS * sizeof(float)));
    const Pattern pats[] = {
        {"32 lanes, consecutive cells (16 full sectors)", 1, 0xffffffffu, 0},
        {"32 lanes, every 2nd cell (32 half sectors)   ", 2, 0xffffffffu, 0},
        {"32 lanes, every 8th cell (32 lines)          ", 8, 0xffffffffu, 0},
        {"16 lanes (even), consecutive-by-2 cells      ", 1, 0x55555555u, 0},
        {" 8 lanes (every 4th)                         ", 1, 0x11111111u, 0},
        {" 4 lanes (every 8th)                         ", 1, 0x01010101u, 0},
        {" 1 lane                                      ", 1, 0x00000001u, 0},
        {"32 lanes consecutive + scalar f32 (the pair) ", 1, 0xffffffffu, 1},
        {" 8 lanes (every 4th) + scalar f32            ", 1, 0x11111111u, 1},
        {"32 lanes every 2nd cell + scalar f32         ", 2, 0xffffffffu, 1},
    };
    const int blocks = p.multiProcessorCount * 4, threads = 128;      // 16 warps per SM, like the splat kernel
    const double warps = (double)blocks * (threads / 32);
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    printf("device %s, %d SMs, %d CTAs x %d threads, %d reductions per warp\n", p.name, p.multiProcessorCount, blocks, threads, ITER);
    printf("%-48s %10s %14s %14s %16s\n", "pattern", "us", "ns/instr/SM", "G lanes/s", "cycles/instr/SM");
    for (const Pattern& q : pats) {
        for (int rep = 0; rep < 2; ++rep) {
            CK(cudaEventRecord(e0));
            k_red<<<blocks, threads>>>(acc, accw, q.stride, q.lane_mask, q.with_scalar);
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        }
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double instr_per_sm = warps * ITER * (1 + q.with_scalar) / p.multiProcessorCount;
        const double lanes = warps * ITER * __builtin_popcount(q.lane_mask) * (1 + q.with_scalar);
        printf("%-48s %10.1f %14.3f %14.1f %16.2f\n", q.name, ms * 1e3, ms * 1e6 / instr_per_sm, lanes / (ms * 1e-3) / 1e9,
               ms * 1e-3 * p.clockRate * 1e3 / instr_per_sm);
    }
    CK(cudaFree(acc)); CK(cudaFree(accw));
    return 0;
}
