// tools/microbench.cu -- integer-pipe issue rates on this GPU: the roofline denominators of the
// distance kernel (VABSDIFF4.U8.ACC) and of the tree kernels (VIMNMX.U16x2 / VIMNMX3 / PRMT / LOP3 /
// IADD3) and the scalar VABSDIFF of the int16 distance path (vabsdiff2 is emulated on sm_100a).  Prints one JSON line per instruction: lane-instructions per
// clock per SM at the SM clock the GPU reports, and the mnemonic count found in this binary's
// SASS is checked by tools/gpu_microbench.sh (cuobjdump) so that a folded loop cannot pass as a rate.
//
// Every operation is an `asm volatile` statement on a loop-carried register, so neither nvcc nor
// ptxas can fold, hoist or strength-reduce it (the round-1 version used plain C: its IADD3 row
// was folded to a + n * x and reported 344 lanes/clk/SM, above the 128-lane issue limit).
//     nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

enum Op { SAD4 = 0, MAX2, PRMT, IADD, IMAD, MAX3, LOP3, SAD1, MAXMIN, N_OPS };

template <int OP>
__device__ __forceinline__ uint32_t step(uint32_t a, uint32_t x, uint32_t y) {
    uint32_t d;
    if (OP == SAD4) asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(x), "r"(y), "r"(a));
    if (OP == SAD1) asm volatile("sad.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(x), "r"(y), "r"(a));  // one 32-bit |x - y| + a: the int16 distance path
    if (OP == MAX2) asm volatile("max.u16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(x));
    if (OP == PRMT) asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(x), "r"(y & 0x7777u));
    if (OP == IADD) asm volatile("add.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(x));
    if (OP == IMAD) asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(x), "r"(y));
    if (OP == MAX3) asm volatile("{.reg .b32 t; max.u16x2 t, %1, %2; max.u16x2 %0, t, %3;}" : "=r"(d) : "r"(a), "r"(x), "r"(y));
    if (OP == LOP3) asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(x), "r"(y));
    if (OP == MAXMIN) {  // the merge's dependent pair: lo = max(a, x); d = min(lo, y)
        uint32_t t;
        asm volatile("max.u16x2 %0, %1, %2;" : "=r"(t) : "r"(a), "r"(x));
        asm volatile("min.u16x2 %0, %1, %2;" : "=r"(d) : "r"(t), "r"(y));
    }
    return d;
}

template <int OP, int ILP>
__global__ void __launch_bounds__(256) k(uint32_t* out, int iters, uint32_t seed) {
    uint32_t a[ILP];
    const uint32_t x = seed * 0x9e3779b9u + threadIdx.x, y = seed * 40503u + blockIdx.x;
#pragma unroll
    for (int i = 0; i < ILP; i++) a[i] = i * 0x01010101u + threadIdx.x;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) a[i] = step<OP>(a[i], x + i, y);
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
void run(const char* name, const char* sass, int per_step, int sms, double mhz) {
    constexpr int ILP = 16;
    const int iters = 4096, threads = 256, blocks = sms * 8;
    uint32_t* out;
    cudaMalloc(&out, sizeof(uint32_t) * blocks * threads);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<OP, ILP><<<blocks, threads>>>(out, 64, 1);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 7; r++) {
        cudaEventRecord(e0);
        k<OP, ILP><<<blocks, threads>>>(out, iters, r + 2);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double lane_instr = (double)blocks * threads * iters * ILP * per_step;
    const double per_s = lane_instr / (best * 1e-3);
    printf("{\"op\": \"%s\", \"sass\": \"%s\", \"ms\": %.4f, \"lane_instr_per_s\": %.4e, \"lane_instr_per_clk_per_sm\": %.2f, \"clock_mhz\": %.0f, \"sms\": %d}\n",
           name, sass, best, per_s, per_s / (mhz * 1e6) / sms, mhz, sms);
    cudaFree(out);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const int sms = p.multiProcessorCount;
    const double mhz = khz / 1000.0;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_mhz\": %.0f, \"note\": \"asm volatile on loop-carried registers, 16 independent chains per thread, 8 CTAs of 256 threads per SM, best of 7\"}\n",
           p.name, sms, mhz);
    run<SAD4>("VABSDIFF4.U8.ACC", "VABSDIFF4", 1, sms, mhz);
    run<SAD1>("VABSDIFF.U32 (sad.u32: one position per instruction, the int16 distance path)", "VABSDIFF", 1, sms, mhz);
    run<MAX2>("VIMNMX.U16x2", "VIMNMX", 1, sms, mhz);
    run<MAX3>("VIMNMX3.U16x2", "VIMNMX3", 1, sms, mhz);
    run<MAXMIN>("VIMNMX.U16x2 max->min dependent pair", "VIMNMX", 2, sms, mhz);
    run<PRMT>("PRMT", "PRMT", 1, sms, mhz);
    run<LOP3>("LOP3", "LOP3", 1, sms, mhz);
    run<IADD>("IADD3", "IADD3", 1, sms, mhz);
    run<IMAD>("IMAD", "IMAD", 1, sms, mhz);
    return 0;
}
