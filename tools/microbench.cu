// tools/microbench.cu -- integer-pipe issue rates on this GPU (roofline denominators for the
// distance kernel: VABSDIFF4.U8.ACC, and for the tree kernels: VIMNMX.U16x2 / PRMT / IADD3).
// Prints lane-instructions per clock per SM.  nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d; asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }

template <int OP, int ILP>
__global__ void k(uint32_t* out, int iters, uint32_t seed) {
    uint32_t a[ILP], x = seed + threadIdx.x, y = seed * 3 + blockIdx.x;
#pragma unroll
    for (int i = 0; i < ILP; i++) a[i] = i + threadIdx.x;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            if (OP == 0) a[i] = sad4(x, y + i, a[i]);
            if (OP == 1) a[i] = __vmaxu2(a[i], x + i);
            if (OP == 2) a[i] = __byte_perm(a[i], x, 0x4140 + i);
            if (OP == 3) a[i] = a[i] + x + i;                       // IADD3
            if (OP == 4) a[i] = a[i] * x + y;                       // IMAD
            if (OP == 6) a[i] = __vimax3_u16x2(a[i], x + i, y);      // VIMNMX3.U16x2
            if (OP == 7) a[i] = __vimin3_u16x2(__vmaxu2(a[i], x + i), y, x);  // VIMNMX + VIMNMX3 chain
            if (OP == 8) a[i] = (a[i] & (x + i)) ^ y;                // LOP3
            if (OP == 5) { a[i] = sad4(x, y + i, a[i]); x = x * 5 + a[i]; }  // SAD + IMAD mix
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + x;
}

template <int OP>
void run(const char* name, int sms, double mhz_hint) {
    constexpr int ILP = 16;
    int iters = 4096, threads = 256, blocks = sms * 4;
    uint32_t* out; cudaMalloc(&out, sizeof(uint32_t) * blocks * threads);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<OP, ILP><<<blocks, threads>>>(out, 64, 1);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        k<OP, ILP><<<blocks, threads>>>(out, iters, r + 2);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    double lane_instr = (double)blocks * threads * iters * ILP * (OP == 5 || OP == 7 ? 2 : 1);
    double per_s = lane_instr / (best * 1e-3);
    printf("{\"op\": \"%s\", \"ms\": %.4f, \"lane_instr_per_s\": %.4e, \"lane_instr_per_clk_per_sm_at_%.0fMHz\": %.2f}\n",
           name, best, per_s, mhz_hint, per_s / (mhz_hint * 1e6) / sms);
    cudaFree(out);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount; double mhz = p.clockRate / 1000.0;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_mhz\": %.0f}\n", p.name, sms, mhz);
    run<0>("VABSDIFF4.U8.ACC", sms, mhz);
    run<1>("VIMNMX.U16x2", sms, mhz);
    run<2>("PRMT", sms, mhz);
    run<3>("IADD3", sms, mhz);
    run<4>("IMAD", sms, mhz);
    run<5>("VABSDIFF4+IMAD", sms, mhz);
    run<6>("VIMNMX3.U16x2", sms, mhz);
    run<7>("VIMNMX.U16x2+VIMNMX3.U16x2", sms, mhz);
    run<8>("LOP3", sms, mhz);
    return 0;
}
