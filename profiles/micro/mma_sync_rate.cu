// mma.sync issue rates on sm_100a: s8 m16n8k32 (IMMA) vs f16 m16n8k16 (HMMA), and I2F.  nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cuda_runtime.h>
#include <cstdint>
template <int KIND>
__global__ void k(int iters, int *out) {
    int c[8][4] = {};
    float f[8][4] = {};
    uint32_t a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (KIND == 0)
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(c[j][0]), "+r"(c[j][1]), "+r"(c[j][2]), "+r"(c[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else if (KIND == 1)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(f[j][0]), "+f"(f[j][1]), "+f"(f[j][2]), "+f"(f[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else {
#pragma unroll
                for (int q = 0; q < 4; q++) f[j][q] += (float)(c[j][q] + i);  // I2F + FADD
            }
        }
    }
    int s = 0;
    for (int j = 0; j < 8; j++) for (int q = 0; q < 4; q++) s += c[j][q] + (int)f[j][q];
    if (s == 0x12345678) out[0] = s;
}
template <int KIND> void run(const char *name, double macs_per_op) {
    int *d; cudaMalloc(&d, 4);
    const int iters = 4096, grid = 148 * 4, block = 256;
    k<KIND><<<grid, block>>>(16, d);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<KIND><<<grid, block>>>(iters, d); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double ops = (double)grid * (block / 32) * iters * 8;  // warp-level instructions (x4 for KIND 2)
    printf("%s: %.3f ms, %.1f G warp-instr/s, %.2f per clk per SM (1.965 GHz), %.1f T MAC/s\n", name, ms, ops / ms / 1e6, ops / (ms * 1e-3) / 1.965e9 / 148, ops * macs_per_op / ms / 1e9);
}
int main() { run<0>("IMMA s8 m16n8k32", 4096); run<1>("HMMA f16 m16n8k16", 2048); run<2>("I2F+FADD x4", 0); return 0; }
