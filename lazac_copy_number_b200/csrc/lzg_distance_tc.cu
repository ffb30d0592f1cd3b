// lzg_distance_tc.cu -- the all-pairs ZCNT distance as an exact integer GEMM on the 5th-generation tensor cores (sm_100a).
//
// |a - b| is not a contraction, but it becomes one under a thermometer code.  Per position p let lo_p = min over the cells and
// w_p = max - min; code the value v as the w_p bits t_k(v) = [v - lo_p > k], k = 0 .. w_p-1.  Then |a - b| = sum_k t_k(a) xor t_k(b),
// and over a whole profile
//        D[i][j] = s_i + s_j - 2 * (T T^t)[i][j],      T in {0,1}^(n x K), K = sum_p w_p, s_i = sum of row i of T,
// bit-identical to the reference's loop (operator- include/copy_number.hpp:74-83 + breakpoint_magnitude src/copy_number.cxx:100-113)
// because every quantity is an integer below 2^31: tcgen05.mma kind::i8 multiplies unsigned bytes into int32 accumulators.
// Breakpoint profiles are sparse and narrow (cfg3: K = 2.02 L), so the product is ~10x fewer tensor-core cycles than the
// VABSDIFF4 kernel has ALU cycles; lzg_distance_run dispatches here by value range (K <= 8 L, the 8-bit format, enough tiles to
// fill the GPU) -- whole matrices, shards and the multi-GPU row-block table alike -- and keeps k_zcnt_distance for everything
// else.  LZG_TC=0 switches the path off, LZG_TC=1 forces it where it can run.
//
// Kernels: k_tc_colrange (lo_p, w_p), k_tc_expand (T and s), and the product in two forms over 128 x 256 tiles of the upper
// triangle, persistent, operands by TMA (128-byte swizzle) issued by one thread, tcgen05.mma kind::i8 issued by one thread into
// two alternating accumulators of 256 TMEM columns, four epilogue warps (tcgen05.ld -> s_i + s_j - 2 g -> D[i][j] and the
// mirror D[j][i]) working on one tile while the next tile's product runs:
//   k_tc_distance2 (default): CTA pairs, cta_group::2 -- M = 256 instructions over two SMs that share the column block, six
//                             stages of 32 KB per SM; 3.74 ms at cfg3, tensor pipe 93 % active;
//   k_tc_distance:            single CTAs, M = 128, four stages of 48 KB; 4.68 ms (LZG_TC_2CTA=0, and fewer than two row blocks).
#include <cuda.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "lzg_common.cuh"

namespace {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 128, TC_STAGES = 4, TC_UMMA_K = 32;
constexpr int TC_STAGE_A = TC_BM * TC_BK, TC_STAGE_B = TC_BN * TC_BK, TC_STAGE = TC_STAGE_A + TC_STAGE_B;
constexpr int TC_SMEM = TC_STAGES * TC_STAGE + 1024 /* alignment slack */ + 256 /* barriers, TMEM pointer */;
constexpr int TC_TMEM_COLS = 256;

// ---- per-position range over the cells: lo_p (biased byte) and hi_p ------------------------------------------------
__global__ void __launch_bounds__(128) k_tc_colrange(const uint8_t* __restrict__ d, int n, int pitch, int L, unsigned* __restrict__ lo_hi) {
    // one thread per group of sixteen positions (one 16-byte load per row, eight rows in flight), rows strided over blockIdx.y;
    // lo_hi[2p] = 255 - lo, lo_hi[2p+1] = hi, both merged with atomicMax
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (16 * g >= pitch) return;  // the padding positions of a row are read too (never reported)
    uint32_t lo4[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu}, hi4[4] = {0u, 0u, 0u, 0u};  // per-byte minima / maxima
    const uint8_t* col = d + 16 * (size_t)g;
    int i = blockIdx.y;
    for (; i + 7 * (int)gridDim.y < n; i += 8 * gridDim.y) {
        uint4 w[8];
#pragma unroll
        for (int u = 0; u < 8; u++) w[u] = __ldg(reinterpret_cast<const uint4*>(col + (size_t)(i + u * gridDim.y) * pitch));
#pragma unroll
        for (int u = 0; u < 8; u++) {
            lo4[0] = __vminu4(lo4[0], w[u].x), hi4[0] = __vmaxu4(hi4[0], w[u].x);
            lo4[1] = __vminu4(lo4[1], w[u].y), hi4[1] = __vmaxu4(hi4[1], w[u].y);
            lo4[2] = __vminu4(lo4[2], w[u].z), hi4[2] = __vmaxu4(hi4[2], w[u].z);
            lo4[3] = __vminu4(lo4[3], w[u].w), hi4[3] = __vmaxu4(hi4[3], w[u].w);
        }
    }
    for (; i < n; i += gridDim.y) {
        const uint4 w = __ldg(reinterpret_cast<const uint4*>(col + (size_t)i * pitch));
        lo4[0] = __vminu4(lo4[0], w.x), hi4[0] = __vmaxu4(hi4[0], w.x);
        lo4[1] = __vminu4(lo4[1], w.y), hi4[1] = __vmaxu4(hi4[1], w.y);
        lo4[2] = __vminu4(lo4[2], w.z), hi4[2] = __vmaxu4(hi4[2], w.z);
        lo4[3] = __vminu4(lo4[3], w.w), hi4[3] = __vmaxu4(hi4[3], w.w);
    }
#pragma unroll
    for (int q = 0; q < 4; q++)
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const int pos = 16 * g + 4 * q + e;
            if (pos < L) atomicMax(lo_hi + 2 * pos, 255u - ((lo4[q] >> (8 * e)) & 255u)), atomicMax(lo_hi + 2 * pos + 1, (hi4[q] >> (8 * e)) & 255u);
        }
}

// ---- T and its row sums ------------------------------------------------------------------------------------------
// The host lays the columns out four to an output word such that the (up to four) positions of a word lie within eight bytes
// of a 4-byte-aligned base (it starts a new word, leaving never-set columns, where they do not); a word's table entry is
// {base word index, PRMT selector, four threshold bytes, 0x01 per valid column}.  T[i][4 w + e] = [byte(i, position) >= threshold]:
// two 32-bit loads of the row, one PRMT gathers the four bytes, a SWAR compare makes the four results.
// A thread owns sixteen columns (four words: its table entries stay in registers) and walks TC_XROWS rows, four at a time;
// neighbouring threads read neighbouring positions (L1 serves the overlap), every 16-byte store is final.  s[i] (zeroed by the
// caller) collects the ones of row i: a warp reduction and one atomic per warp and row.
constexpr int TC_XROWS = 64;
__global__ void __launch_bounds__(256) k_tc_expand(const uint8_t* __restrict__ d, int n, int pitch, const uint4* __restrict__ table, int k_pad,
                                                    uint8_t* __restrict__ T, int32_t* __restrict__ s) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;  // 16-byte group of a row of T
    const bool active = 16 * g < k_pad;
    uint4 ent[4];
#pragma unroll
    for (int q = 0; q < 4; q++) ent[q] = active ? __ldg(table + 4 * g + q) : make_uint4(0, 0, 0, 0);
    const int last_word = pitch / 4 - 1;
    const int i0 = blockIdx.y * TC_XROWS;  // rows n .. n_pad-1 are zero rows
    for (int r = 0; r < TC_XROWS; r += 4) {
        uint32_t lo[4][4], hi[4][4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int i = i0 + r + u;
            const uint32_t* roww = reinterpret_cast<const uint32_t*>(d + (size_t)i * pitch);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                lo[u][q] = i < n ? __ldg(roww + ent[q].x) : 0u;
                hi[u][q] = i < n ? __ldg(roww + min((int)ent[q].x + 1, last_word)) : 0u;
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int i = i0 + r + u;
            uint32_t out[4];
            int ones = 0;
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const uint32_t x = __byte_perm(lo[u][q], hi[u][q], ent[q].y), y = ent[q].z;
                const uint32_t dlt = (x | 0x80808080u) - (y & 0x7f7f7f7fu);      // bit 7 of a byte: low seven bits of x >= those of y
                const uint32_t ge = (x & ~y) | (~(x ^ y) & dlt);                  // bit 7 of a byte: x >= y
                out[q] = i < n ? (ge >> 7) & ent[q].w : 0u;
                ones += __popc(out[q]);
            }
            if (active) *reinterpret_cast<uint4*>(T + (size_t)i * k_pad + 16 * g) = make_uint4(out[0], out[1], out[2], out[3]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) ones += __shfl_xor_sync(0xffffffffu, ones, o);
            if ((threadIdx.x & 31) == 0 && ones) atomicAdd(s + i, ones);
        }
    }
}

// ---- PTX helpers ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait: a wrong descriptor or a lost arrival must end in an error, not in a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    const long long t0 = clock64();
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) return;
        if (clock64() - t0 > 2000000000ll) __trap();  // about a second
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst), "l"(map), "r"(c0), "r"(c1),
                 "r"(bar)
                 : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], unsigned bytes into int32
__device__ __forceinline__ void tc_mma_i8(uint32_t d_tmem, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// shared-memory operand descriptor: K-major tile, 128-byte swizzle, rows of 128 bytes, groups of 8 rows 1024 bytes apart
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t addr) {
    return (uint64_t)((addr & 0x3FFFFu) >> 4) | (1ull << 16) /* leading byte offset: unused with a swizzle */ | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) /* sm_100 */ |
           (2ull << 61) /* SWIZZLE_128B */;
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, "
        "%24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]),
          "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// ---- the product and the epilogue --------------------------------------------------------------------------------------
// tiles[t] = (row block of 128, column block of 256); every tile runs the whole K.  Warp 0 (one lane) is the TMA producer, warp 1
// (one lane) the MMA issuer, warps 2 .. 5 the epilogue: each reads its 32 TMEM lanes (= tile rows) and writes D.  Two
// accumulators of 256 TMEM columns alternate, so the product of the next tile runs while this one is written out, and the
// producer keeps the four stages full across tile boundaries.  Mirrors D[j][i] are written only for rows j beyond this tile's
// row block, so a finished group of row blocks never changes again (lzg_distance_matrix copies it out while later groups compute).
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }

__global__ void __launch_bounds__(192, 1)
k_tc_distance(const __grid_constant__ CUtensorMap tmap, const int2* __restrict__ tiles, int n_tiles, int k_blocks, const int32_t* __restrict__ s,
              int32_t* __restrict__ D, int32_t* const* __restrict__ row_base, int n) {
    extern __shared__ uint8_t tc_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(tc_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t bar_full = smem_base + TC_STAGES * TC_STAGE, bar_empty = bar_full + 8 * TC_STAGES, bar_tfull = bar_empty + 8 * TC_STAGES, bar_tempty = bar_tfull + 16;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + TC_STAGES * TC_STAGE + 8 * (2 * TC_STAGES + 4));
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int q = 0; q < TC_STAGES; q++) mbar_init(bar_full + 8 * q, 1), mbar_init(bar_empty + 8 * q, 1);
        for (int q = 0; q < 2; q++) mbar_init(bar_tfull + 8 * q, 1), mbar_init(bar_tempty + 8 * q, 4);  // one arrival per epilogue warp
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * TC_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    // instruction descriptor: D int32, A and B unsigned 8-bit, both K-major, N = 256, M = 128
    constexpr uint32_t idesc = (2u << 4) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
    if (warp == 0) {
        if (lane == 0) {
            uint32_t q = 0;  // k-blocks issued so far (stage = q % STAGES, phase = q / STAGES & 1)
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                const int2 tile = tiles[t];
                const int row0 = tile.x * TC_BM, col0 = tile.y * TC_BN;
                for (int kb = 0; kb < k_blocks; kb++, q++) {
                    const uint32_t st = q % TC_STAGES, ph = (q / TC_STAGES) & 1u;
                    mbar_wait(bar_empty + 8 * st, ph ^ 1u);
                    mbar_expect_tx(bar_full + 8 * st, TC_STAGE);
                    const uint32_t a = smem_base + st * TC_STAGE, b = a + TC_STAGE_A;
                    tma_load_2d(a, &tmap, kb * TC_BK, row0, bar_full + 8 * st);
                    tma_load_2d(b, &tmap, kb * TC_BK, col0, bar_full + 8 * st);
                    tma_load_2d(b + TC_STAGE_A, &tmap, kb * TC_BK, col0 + 128, bar_full + 8 * st);
                }
            }
        }
    } else if (warp == 1) {
        // The whole warp walks the loop (warp-uniform control flow keeps addresses and descriptors in uniform registers), one
        // elected lane issues.  k_blocks is a multiple of the stage count, so the stage of a k-block is a compile-time constant
        // of the unrolled loop and its descriptors are the stage-0 descriptors plus constants.
        const uint64_t desc_a0 = tc_smem_desc(smem_base), desc_b0 = tc_smem_desc(smem_base + TC_STAGE_A);
        uint32_t rounds = 0, lt = 0;  // rounds of STAGES k-blocks so far: the phase of the full barriers
        for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, lt++) {
            const uint32_t acc = lt & 1u;
            mbar_wait(bar_tempty + 8 * acc, ((lt >> 1) & 1u) ^ 1u);  // the epilogue has drained this accumulator
            tc_fence_after();
            const uint32_t d_tmem = tmem + acc * TC_TMEM_COLS;
            for (int kb = 0; kb < k_blocks; kb += TC_STAGES, rounds++) {
#pragma unroll
                for (int st = 0; st < TC_STAGES; st++) {
                    mbar_wait(bar_full + 8 * st, rounds & 1u);
                    tc_fence_after();
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < TC_BK / TC_UMMA_K; k++) {
                            const uint64_t off = (uint64_t)((st * TC_STAGE + k * TC_UMMA_K) >> 4);
                            tc_mma_i8(d_tmem, desc_a0 + off, desc_b0 + off, idesc, (kb | st | k) != 0);
                        }
                        tc_commit(bar_empty + 8 * st);  // the stage is free once these MMAs have read it
                    }
                    __syncwarp();
                }
            }
            if (elect_one()) tc_commit(bar_tfull + 8 * acc);  // the accumulator is complete
            __syncwarp();
        }
    } else {
        const int quad = warp & 3, trow = quad * 32 + lane;  // a warp reads the TMEM lanes of its quadrant
        const bool vec = (n & 3) == 0, vec8 = (n & 7) == 0;
        uint32_t lt = 0;
        for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, lt++) {
            const int2 tile = tiles[t];
            const int row0 = tile.x * TC_BM, col0 = tile.y * TC_BN;
            const uint32_t acc = lt & 1u;
            const int i = row0 + trow;
            const int si = i < n ? __ldg(s + i) : 0;
            // row_base (several GPUs, lzg_multi.cu): one pointer per block of 128 output rows instead of the one matrix
            int32_t* drow = row_base ? row_base[tile.x] + (size_t)trow * n : D + (size_t)i * n;
            const int mirror_from = row0 + TC_BM;  // rows below this tile's row block only
            mbar_wait(bar_tfull + 8 * acc, (lt >> 1) & 1u);
            tc_fence_after();
            for (int c = 0; c < TC_BN; c += 32) {
                const int j0 = col0 + c;
                if (j0 >= n) break;  // uniform over the CTA
                uint32_t g[32];
                tc_ld32(tmem + acc * TC_TMEM_COLS + ((uint32_t)(quad * 32) << 16) + c, g);
                int v[32];
#pragma unroll
                for (int e = 0; e < 32; e++) v[e] = si + (j0 + e < n ? __ldg(s + j0 + e) : 0) - 2 * (int)g[e];
                if (i < n) {
                    if (vec8 && j0 + 32 <= n) {  // 32-byte stores: one whole sector per lane and instruction
#pragma unroll
                        for (int e = 0; e < 32; e += 8)
                            asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(drow + j0 + e), "r"(v[e]), "r"(v[e + 1]), "r"(v[e + 2]), "r"(v[e + 3]),
                                         "r"(v[e + 4]), "r"(v[e + 5]), "r"(v[e + 6]), "r"(v[e + 7])
                                         : "memory");
                    } else if (vec && j0 + 32 <= n) {
#pragma unroll
                        for (int e = 0; e < 32; e += 4) *reinterpret_cast<int4*>(drow + j0 + e) = make_int4(v[e], v[e + 1], v[e + 2], v[e + 3]);
                    } else {
#pragma unroll
                        for (int e = 0; e < 32; e++)
                            if (j0 + e < n) drow[j0 + e] = v[e];
                    }
#pragma unroll
                    for (int e = 0; e < 32; e++) {
                        const int j = j0 + e;
                        if (j >= mirror_from && j < n) {  // lanes = consecutive i: coalesced
                            if (row_base) row_base[j >> 7][(size_t)(j & 127) * n + i] = v[e];
                            else D[(size_t)j * n + i] = v[e];
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty + 8 * acc);  // this warp's lanes of the accumulator are read
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(2 * TC_TMEM_COLS) : "memory");
}


// ---- the same product on CTA PAIRS (tcgen05 cta_group::2) ------------------------------------------------------------------
// What limits k_tc_distance is the operand stream into the SM: 48 KB of shared memory filled per 128 x 256 x 128 block of
// products.  Here two CTAs of a cluster share a column block: each loads its own 128 rows of A and HALF of the 256 B rows
// (32 KB per block and SM, six stages), one thread of the leader CTA issues M = 256 instructions that read both CTAs' shared
// memory and accumulate into both CTAs' TMEM, and every CTA writes out its own 128 rows.  The two row blocks of a pair need not
// be neighbours: tiles2[t] = (row block of CTA 0, row block of CTA 1 or -1, column block).
constexpr int TC2_STAGES = 6, TC2_STAGE = 2 * TC_STAGE_A;  // A 16 KB + half of B 16 KB
constexpr int TC2_SMEM = TC2_STAGES * TC2_STAGE + 1024 + 256;

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// both CTAs of a pair issue their loads; the bytes count on the LEADER's barrier (its address with the peer bit cleared)
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst), "l"(map), "r"(c0),
                 "r"(c1), "r"(bar & 0xFEFFFFFFu)
                 : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {  // arrives on the barrier at this offset in BOTH CTAs
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc_mma_i8_pair(uint32_t d_tmem, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive_cta(uint32_t bar, uint32_t cta) {  // on the barrier at this offset in CTA `cta` of the cluster
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(bar),
        "r"(cta)
        : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
    const long long t0 = clock64();
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) return;
        if (clock64() - t0 > 2000000000ll) __trap();
    }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1)
k_tc_distance2(const __grid_constant__ CUtensorMap tmap, const int4* __restrict__ tiles, int n_tiles, int k_blocks, const int32_t* __restrict__ s,
               int32_t* __restrict__ D, int32_t* const* __restrict__ row_base, int n) {
    extern __shared__ uint8_t tc_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(tc_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t bar_full = smem_base + TC2_STAGES * TC2_STAGE, bar_empty = bar_full + 8 * TC2_STAGES, bar_tfull = bar_empty + 8 * TC2_STAGES, bar_tempty = bar_tfull + 16;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + TC2_STAGES * TC2_STAGE + 8 * (2 * TC2_STAGES + 4));
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_rank();
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    if (tid == 0) {
        for (int q = 0; q < TC2_STAGES; q++) mbar_init(bar_full + 8 * q, 1), mbar_init(bar_empty + 8 * q, 1);
        for (int q = 0; q < 2; q++) mbar_init(bar_tfull + 8 * q, 1), mbar_init(bar_tempty + 8 * q, 8);  // four epilogue warps of each CTA
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * TC_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    cluster_sync();  // both CTAs' barriers exist before anything arrives on them
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    // instruction descriptor: D int32, A and B unsigned 8-bit, both K-major, N = 256, M = 256 over the pair
    constexpr uint32_t idesc = (2u << 4) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
    if (warp == 0) {
        if (lane == 0) {
            uint32_t q = 0;
            for (int t = pair; t < n_tiles; t += n_pairs) {
                const int4 tile = tiles[t];
                const int rb = rank == 0 || tile.y < 0 ? tile.x : tile.y;  // a missing second row block: its CTA loads the first one's rows (unused)
                const int row0 = rb * TC_BM, colh = tile.z * TC_BN + 128 * (int)rank;  // this CTA's half of the column block
                for (int kb = 0; kb < k_blocks; kb++, q++) {
                    const uint32_t st = q % TC2_STAGES, ph = (q / TC2_STAGES) & 1u;
                    mbar_wait_cluster(bar_empty + 8 * st, ph ^ 1u);
                    if (rank == 0) mbar_expect_tx(bar_full + 8 * st, 2 * TC2_STAGE);  // both CTAs' bytes
                    const uint32_t a = smem_base + st * TC2_STAGE, b = a + TC_STAGE_A;
                    tma_load_2d_pair(a, &tmap, kb * TC_BK, row0, bar_full + 8 * st);
                    tma_load_2d_pair(b, &tmap, kb * TC_BK, colh, bar_full + 8 * st);
                }
            }
        }
    } else if (warp == 1) {
        if (rank == 0) {
            const uint64_t desc_a0 = tc_smem_desc(smem_base), desc_b0 = tc_smem_desc(smem_base + TC_STAGE_A);
            uint32_t q = 0, lt = 0;
            for (int t = pair; t < n_tiles; t += n_pairs, lt++) {
                const uint32_t acc = lt & 1u;
                mbar_wait_cluster(bar_tempty + 8 * acc, ((lt >> 1) & 1u) ^ 1u);  // both CTAs' epilogues have drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem + acc * TC_TMEM_COLS;
                for (int kb = 0; kb < k_blocks; kb++, q++) {
                    const uint32_t st = q % TC2_STAGES, ph = (q / TC2_STAGES) & 1u;
                    mbar_wait_cluster(bar_full + 8 * st, ph);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint64_t off0 = (uint64_t)((st * TC2_STAGE) >> 4);
#pragma unroll
                        for (int k = 0; k < TC_BK / TC_UMMA_K; k++)
                            tc_mma_i8_pair(d_tmem, desc_a0 + off0 + 2 * k, desc_b0 + off0 + 2 * k, idesc, (kb | k) != 0);
                        tc_commit_pair(bar_empty + 8 * st);
                    }
                    __syncwarp();
                }
                if (elect_one()) tc_commit_pair(bar_tfull + 8 * acc);
                __syncwarp();
            }
        }
    } else {
        const int quad = warp & 3, trow = quad * 32 + lane;
        const bool vec = (n & 3) == 0, vec8 = (n & 7) == 0;
        uint32_t lt = 0;
        for (int t = pair; t < n_tiles; t += n_pairs, lt++) {
            const int4 tile = tiles[t];
            const int rb = rank == 0 ? tile.x : tile.y;  // -1: nothing to write for this CTA
            const int row0 = rb * TC_BM, col0 = tile.z * TC_BN;
            const uint32_t acc = lt & 1u;
            const int i = row0 + trow;
            const bool live = rb >= 0 && i < n;
            const int si = live ? __ldg(s + i) : 0;
            int32_t* drow = !live ? nullptr : row_base ? row_base[rb] + (size_t)trow * n : D + (size_t)i * n;
            const int mirror_from = row0 + TC_BM;
            mbar_wait_cluster(bar_tfull + 8 * acc, (lt >> 1) & 1u);
            tc_fence_after();
            if (rb >= 0) {
                for (int c = 0; c < TC_BN; c += 32) {
                    const int j0 = col0 + c;
                    if (j0 >= n) break;
                    uint32_t g[32];
                    tc_ld32(tmem + acc * TC_TMEM_COLS + ((uint32_t)(quad * 32) << 16) + c, g);
                    int v[32];
#pragma unroll
                    for (int e = 0; e < 32; e++) v[e] = si + (j0 + e < n ? __ldg(s + j0 + e) : 0) - 2 * (int)g[e];
                    if (live) {
                        if (vec8 && j0 + 32 <= n) {
#pragma unroll
                            for (int e = 0; e < 32; e += 8)
                                asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(drow + j0 + e), "r"(v[e]), "r"(v[e + 1]), "r"(v[e + 2]),
                                             "r"(v[e + 3]), "r"(v[e + 4]), "r"(v[e + 5]), "r"(v[e + 6]), "r"(v[e + 7])
                                             : "memory");
                        } else if (vec && j0 + 32 <= n) {
#pragma unroll
                            for (int e = 0; e < 32; e += 4) *reinterpret_cast<int4*>(drow + j0 + e) = make_int4(v[e], v[e + 1], v[e + 2], v[e + 3]);
                        } else {
#pragma unroll
                            for (int e = 0; e < 32; e++)
                                if (j0 + e < n) drow[j0 + e] = v[e];
                        }
#pragma unroll
                        for (int e = 0; e < 32; e++) {
                            const int j = j0 + e;
                            if (j >= mirror_from && j < n) {
                                if (row_base) row_base[j >> 7][(size_t)(j & 127) * n + i] = v[e];
                                else D[(size_t)j * n + i] = v[e];
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cta(bar_tempty + 8 * acc, 0);  // on the leader's barrier
        }
    }
    tc_fence_before();
    cluster_sync();  // no CTA leaves while its partner may still read its shared memory or signal its barriers
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(2 * TC_TMEM_COLS) : "memory");
}

}  // namespace

// ---- host side ---------------------------------------------------------------------------------------------------------
typedef CUresult (*tc_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                 CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static tc_encode_fn tc_encoder() {
    static tc_encode_fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) fn = (tc_encode_fn)sym;
        else cudaGetLastError();
    }
    return fn;
}

// Host only: the word table of k_tc_expand from the per-position ranges (lohi[2q] = 255 - lowest byte, lohi[2q+1] = highest
// byte; bytes are values + 128).  Entry {x: word index of the 4-byte-aligned base position, y: PRMT selector (nibble e = byte
// offset of column e's position from the base, 0 .. 7), z: the four threshold bytes, w: 0x01 per valid column}.
static void tc_layout(const unsigned* lohi, int L, std::vector<uint4>& table) {
    // at most one word per column: size once, write through a pointer, shrink at the end
    size_t k = 0;
    for (int q = 0; q < L; q++) k += lohi[2 * q + 1] - (255u - lohi[2 * q]);
    table.resize(k + 1);
    uint4* out = table.data();
    size_t n_words = 0;
    unsigned base = 0, sel = 0, thr4 = 0, mask = 0;
    int fill = 0;
    for (int q = 0; q < L; q++) {
        const unsigned lo = 255u - lohi[2 * q], hi = lohi[2 * q + 1];
        for (unsigned thr = lo + 1; thr <= hi; thr++) {
            if (fill == 4 || (fill > 0 && (unsigned)q >= 4 * base + 8)) {  // full, or this position is out of the word's reach
                out[n_words++] = make_uint4(base, sel, thr4, mask);
                sel = thr4 = mask = 0, fill = 0;
            }
            if (fill == 0) base = (unsigned)q >> 2;
            sel |= ((unsigned)q - 4 * base) << (4 * fill);
            thr4 |= thr << (8 * fill);
            mask |= 1u << (8 * fill);
            fill++;
        }
    }
    if (fill) out[n_words++] = make_uint4(base, sel, thr4, mask);
    table.resize(n_words);
}

// The same for tests (no GPU, no context): lo / hi = lowest / highest VALUE of every position (-128 .. 127).  Writes up to
// cap_words entries of four uint32 and returns the number of words the code needs (call again with a larger buffer if it is
// more than cap_words); *K = sum of (hi - lo).
extern "C" int64_t lzg_distance_tc_layout(const int32_t* lo, const int32_t* hi, int32_t L, uint32_t* words, int64_t cap_words, int64_t* K) {
    if (!lo || !hi || L < 0) return -1;
    std::vector<unsigned> lohi(2 * (size_t)L);
    int64_t k = 0;
    for (int q = 0; q < L; q++) {
        if (lo[q] < -128 || hi[q] > 127 || lo[q] > hi[q]) return -1;
        lohi[2 * q] = 255u - (unsigned)(lo[q] + 128), lohi[2 * q + 1] = (unsigned)(hi[q] + 128);
        k += hi[q] - lo[q];
    }
    if (K) *K = k;
    std::vector<uint4> table;
    tc_layout(lohi.data(), L, table);
    for (size_t w = 0; w < table.size() && (int64_t)w < cap_words && words; w++)
        words[4 * w] = table[w].x, words[4 * w + 1] = table[w].y, words[4 * w + 2] = table[w].z, words[4 * w + 3] = table[w].w;
    return (int64_t)table.size();
}

void lzg_distance_tc_free(lzg_ctx* ctx, lzg_profiles* p) {
    if (!p->tc) return;
    for (cudaEvent_t ev : p->tc->ev)
        if (ev) cudaEventDestroy(ev);
    lzg_dev_free(ctx, p->tc->d_T);
    lzg_dev_free(ctx, p->tc->d_s);
    lzg_dev_free(ctx, p->tc->d_tiles);
    delete p->tc;
    p->tc = nullptr;
}

// Two steps so that several GPUs (lzg_multi.cu) overlap: _begin queues the range kernel and its read-back, _finish waits for
// it, lays the code out on the host (a loop over the positions) and queues the expansion.  Returns 1: usable, 0: these profiles
// stay with k_zcnt_distance, < 0: minus an LZG_ERR_* status.
int lzg_distance_tc_prepare_begin(lzg_ctx* ctx, lzg_profiles* p) {
    if (p->tc) return p->tc->usable;
    const char* env = getenv("LZG_TC");
    const int mode = env ? atoi(env) : -1;  // 0 off, 1 forced, otherwise by value range
    lzg_tc_state* tc = new lzg_tc_state;
    p->tc = tc;
    // fewer tiles than SMs (n below ~2,700): one tile per CTA would leave most of the GPU idle, and k_zcnt_distance splits K
    const long long whole_tiles = (long long)((p->n + TC_BM - 1) / TC_BM) * ((p->n + TC_BN - 1) / TC_BN) / 2;
    if (mode == 0 || p->width != 1 || p->n < 256 || p->L < 1 || !tc_encoder()) return 0;
    if (mode != 1 && whole_tiles < (long long)ctx->sm_count * 3 / 4) return 0;
    const int n = p->n, L = p->L;
    int rc = lzg_ensure_pinned(ctx, 16 + 2 * (size_t)L);
    if (rc != LZG_OK) return -rc;
    if (getenv("LZG_TC_TIMING") && atoi(getenv("LZG_TC_TIMING"))) {
        for (cudaEvent_t& ev : tc->ev) cudaEventCreate(&ev);
        cudaEventRecord(tc->ev[0], ctx->stream);
    }
    unsigned* d_lohi = nullptr;
    cudaError_t e = lzg_dev_alloc(ctx, (void**)&d_lohi, sizeof(unsigned) * 2 * (size_t)L);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_lohi, 0, sizeof(unsigned) * 2 * (size_t)L, ctx->stream);
    if (e == cudaSuccess) {
        // about 40 KB of loads in flight per SM: groups of sixteen positions x row slabs, eight rows a thread at a time
        const int blocks_x = (p->pitch / 16 + 127) / 128;
        const int slabs = std::max(1, std::min((n + 7) / 8, 3 * ctx->sm_count / blocks_x));
        k_tc_colrange<<<dim3(blocks_x, slabs), 128, 0, ctx->stream>>>(p->d, n, p->pitch, L, d_lohi);
        ctx->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(ctx->h_pinned + 16, d_lohi, sizeof(unsigned) * 2 * (size_t)L, cudaMemcpyDeviceToHost, ctx->stream);
    lzg_dev_free(ctx, d_lohi);
    if (e != cudaSuccess) return -lzg_fail(ctx, LZG_ERR_CUDA, "tensor-core distance: value ranges", e);
    tc->begun = 1;
    return 2;  // _finish decides
}

int lzg_distance_tc_prepare_finish(lzg_ctx* ctx, lzg_profiles* p) {
    lzg_tc_state* tc = p->tc;
    if (!tc || !tc->begun) return tc ? tc->usable : 0;
    tc->begun = 0;
    const char* env = getenv("LZG_TC");
    const int mode = env ? atoi(env) : -1;
    const int n = p->n, L = p->L;
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return -lzg_fail(ctx, LZG_ERR_CUDA, "tensor-core distance: value ranges", e);
    const unsigned* lohi = reinterpret_cast<const unsigned*>(ctx->h_pinned + 16);
    int64_t K = 0;
    for (int q = 0; q < L; q++) K += (int64_t)lohi[2 * q + 1] - (255 - (int64_t)lohi[2 * q]);
    tc->K = K;
    if (mode != 1 && K > 8ll * L) return 0;  // wide values: the VABSDIFF4 kernel is the faster one
    // the word table: four columns per word, positions within eight bytes of the word's 4-byte-aligned base
    std::vector<uint4>& table = tc->table;
    tc_layout(lohi, L, table);
    const int64_t k_used = 4 * (int64_t)table.size();
    // a multiple of STAGES k-blocks per tile: the issuer's loop is unrolled over the stages with compile-time descriptors
    constexpr int64_t KQ = (int64_t)TC_BK * TC_STAGES;
    const int64_t k_pad = std::max<int64_t>(KQ, (k_used + KQ - 1) / KQ * KQ), n_pad = ((int64_t)n + 255) / 256 * 256;
    table.resize((size_t)(k_pad / 4), make_uint4(0, 0, 0, 0));
    if (k_pad >= (1ll << 30)) return 0;
    if ((uint64_t)(n_pad * k_pad) > ((uint64_t)1 << 30)) {  // beyond 1 GB: only if it leaves half of the free memory
        size_t free_b = 0, total_b = 0;
        cudaMemGetInfo(&free_b, &total_b);
        if ((uint64_t)(n_pad * k_pad) >= free_b / 2) return 0;
    }
    uint4* d_table = nullptr;
    e = lzg_dev_alloc(ctx, (void**)&d_table, sizeof(uint4) * table.size());
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&tc->d_T, (size_t)(n_pad * k_pad));
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&tc->d_s, sizeof(int32_t) * (size_t)n_pad);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_table, table.data(), sizeof(uint4) * table.size(), cudaMemcpyHostToDevice, ctx->stream);  // pageable: staged before the call returns
    if (e == cudaSuccess) e = cudaMemsetAsync(tc->d_s, 0, sizeof(int32_t) * (size_t)n_pad, ctx->stream);
    if (e == cudaSuccess) {
        k_tc_expand<<<dim3((unsigned)((k_pad / 16 + 255) / 256), (unsigned)(n_pad / TC_XROWS)), 256, 0, ctx->stream>>>(p->d, n, p->pitch, d_table, (int)k_pad, tc->d_T, tc->d_s);
        ctx->launches++;
        e = cudaGetLastError();
    }
    lzg_dev_free(ctx, d_table);
    if (e != cudaSuccess) return -lzg_fail(ctx, LZG_ERR_CUDA, "tensor-core distance: thermometer code", e);
    // the TMA descriptor of T: bytes, [n_pad][k_pad], boxes of 128 rows x 128 bytes, 128-byte swizzle
    const cuuint64_t dims[2] = {(cuuint64_t)k_pad, (cuuint64_t)n_pad}, strides[1] = {(cuuint64_t)k_pad};
    const cuuint32_t box[2] = {TC_BK, 128}, estr[2] = {1, 1};
    const CUresult cr = tc_encoder()(reinterpret_cast<CUtensorMap*>(tc->tmap), CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, tc->d_T, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) {
        lzg_fail(ctx, LZG_ERR_CUDA, "tensor-core distance: cuTensorMapEncodeTiled");
        return -LZG_ERR_CUDA;
    }
    static bool attrs_set[64] = {};  // per device, once
    if (ctx->device >= 64 || !attrs_set[ctx->device]) {
        e = cudaFuncSetAttribute(k_tc_distance, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_tc_distance2, cudaFuncAttributeMaxDynamicSharedMemorySize, TC2_SMEM);
        if (e == cudaSuccess && ctx->device < 64) attrs_set[ctx->device] = true;
    }
    if (e != cudaSuccess) return -lzg_fail(ctx, LZG_ERR_CUDA, "tensor-core distance: shared memory", e);
    tc->k_pad = (int)k_pad, tc->n_pad = (int)n_pad;
    tc->usable = 1;
    if (tc->ev[1]) cudaEventRecord(tc->ev[1], ctx->stream);
    return 1;
}

int lzg_distance_tc_prepare(lzg_ctx* ctx, lzg_profiles* p) {
    const int r = lzg_distance_tc_prepare_begin(ctx, p);
    return r == 2 ? lzg_distance_tc_prepare_finish(ctx, p) : r;
}

// Tiles of the row blocks (128 rows, = the tile rows of k_zcnt_distance) that `shard` owns in [row_lo, row_hi), into the whole
// matrix d_D or into the row blocks of d_row_base.  owner(r): the deal of lzg_distance.cu (pairs (r, R-1-r) or snake order).
int lzg_distance_tc_run(lzg_ctx* ctx, lzg_profiles* p, int shard, int n_shards, int own_mode, int row_lo, int row_hi, int32_t* d_D, int32_t* const* d_row_base,
                        int64_t* pairs_done) {
    lzg_tc_state* tc = p->tc;
    const int n = p->n, RB = (n + TC_BM - 1) / TC_BM, CB = (n + TC_BN - 1) / TC_BN;
    auto owner = [&](int r) { return own_mode == 0 ? std::min(r, RB - 1 - r) % n_shards : lzg_snake_owner(r, n_shards); };
    std::vector<int> mine;
    for (int rb = std::max(0, row_lo); rb < std::min(RB, row_hi); rb++)
        if (owner(rb) == shard) mine.push_back(rb);
    // bands of 12 of these row blocks, column blocks swept inside a band: the ~148 tiles in flight share 12 row blocks and ~12
    // column blocks, and run through K side by side -- the operands come from L2, not from DRAM
    tc->h_tiles.clear();
    int64_t pairs = 0;
    for (size_t b0 = 0; b0 < mine.size(); b0 += 12) {
        const size_t b1 = std::min(mine.size(), b0 + 12);
        for (int cb = mine[b0] / 2; cb < CB; cb++)
            for (size_t q = b0; q < b1; q++)
                if (cb >= mine[q] / 2) tc->h_tiles.push_back(make_int2(mine[q], cb));
    }
    static bool pairs_refused = false;  // a cluster launch failed once in this process: single CTAs from then on
    bool pairs_mode = !(getenv("LZG_TC_2CTA") && atoi(getenv("LZG_TC_2CTA")) == 0) && mine.size() >= 2 && !pairs_refused;
    std::vector<int4> tiles2;
    if (pairs_mode) {
        // CTA pairs: consecutive owned row blocks share their column blocks; bands of six pairs
        for (size_t b0 = 0; b0 < mine.size(); b0 += 12) {
            const size_t b1 = std::min(mine.size(), b0 + 12);
            for (int cb = mine[b0] / 2; cb < CB; cb++)
                for (size_t q = b0; q < b1; q += 2) {
                    const int r0 = mine[q], r1 = q + 1 < b1 ? mine[q + 1] : -1;
                    if (cb >= r0 / 2) tiles2.push_back(make_int4(r0, r1, cb, 0));
                }
        }
    }
    for (int rb : mine) {
        const int64_t i0 = (int64_t)rb * TC_BM, i1 = std::min<int64_t>(n, i0 + TC_BM);
        pairs += (i1 - i0) * n - (i0 + i1 - 1) * (i1 - i0) / 2;  // sum over the rows of (n - i)
    }
    if (pairs_done) *pairs_done = pairs;
    int n_tiles = pairs_mode ? (int)tiles2.size() : (int)tc->h_tiles.size();
    if (n_tiles == 0) return LZG_OK;
    if (std::max(n_tiles, (int)tc->h_tiles.size()) > tc->tiles_cap) {
        lzg_dev_free(ctx, tc->d_tiles);
        tc->d_tiles = nullptr;
        const int cap = std::max(std::max(n_tiles, (int)tc->h_tiles.size()), RB * CB);
        LZG_CUDA(ctx, lzg_dev_alloc(ctx, (void**)&tc->d_tiles, sizeof(int4) * (size_t)cap));  // int2 tiles or int4 pair tiles
        tc->tiles_cap = cap;
    }
    CUtensorMap map;
    memcpy(&map, tc->tmap, sizeof(map));
    if (pairs_mode) {
        LZG_CUDA(ctx, cudaMemcpyAsync(tc->d_tiles, tiles2.data(), sizeof(int4) * (size_t)n_tiles, cudaMemcpyHostToDevice, ctx->stream));  // pageable: staged before the call returns
        static int max_clusters_of[64];  // resident CTA pairs of this kernel, per device (0 = not asked yet)
        int& max_clusters = max_clusters_of[ctx->device & 63];
        if (max_clusters <= 0) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(ctx->sm_count & ~1), cfg.blockDim = dim3(192), cfg.dynamicSmemBytes = TC2_SMEM;
            cudaLaunchAttribute at;
            at.id = cudaLaunchAttributeClusterDimension;
            at.val.clusterDim.x = 2, at.val.clusterDim.y = 1, at.val.clusterDim.z = 1;
            cfg.attrs = &at, cfg.numAttrs = 1;
            int nc = 0;
            if (cudaOccupancyMaxActiveClusters(&nc, k_tc_distance2, &cfg) != cudaSuccess || nc < 1) nc = ctx->sm_count / 2, cudaGetLastError();
            max_clusters = std::min(nc, ctx->sm_count / 2);
        }
        if (tc->ev[2]) cudaEventRecord(tc->ev[2], ctx->stream);
        k_tc_distance2<<<2 * std::min(n_tiles, max_clusters), 192, TC2_SMEM, ctx->stream>>>(map, reinterpret_cast<const int4*>(tc->d_tiles), n_tiles, tc->k_pad / TC_BK, tc->d_s,
                                                                                          d_D, d_row_base, n);
        if (cudaGetLastError() == cudaSuccess) {
            ctx->launches++;
            if (tc->ev[3]) cudaEventRecord(tc->ev[3], ctx->stream);
            tc->launches++;
            tc->tiles_run += 2 * (int64_t)n_tiles;
            return LZG_OK;
        }
        pairs_refused = true;  // no clusters here (a partitioned GPU, a driver limit): the single-CTA kernel below
        pairs_mode = false;
        n_tiles = (int)tc->h_tiles.size();
    }
    LZG_CUDA(ctx, cudaMemcpyAsync(tc->d_tiles, tc->h_tiles.data(), sizeof(int2) * (size_t)n_tiles, cudaMemcpyHostToDevice, ctx->stream));
    if (tc->ev[2]) cudaEventRecord(tc->ev[2], ctx->stream);
    k_tc_distance<<<std::min(n_tiles, ctx->sm_count), 192, TC_SMEM, ctx->stream>>>(map, tc->d_tiles, n_tiles, tc->k_pad / TC_BK, tc->d_s, d_D, d_row_base, n);
    LZG_LAUNCH_CHECK(ctx);
    if (tc->ev[3]) cudaEventRecord(tc->ev[3], ctx->stream);
    tc->launches++;
    tc->tiles_run += n_tiles;
    return LZG_OK;
}

// info[4], info[5]: microseconds of the preparation (range kernel, host layout, expansion) and of the last k_tc_distance launch,
// measured with events on the stream when LZG_TC_TIMING=1 (the stream must be idle); info[6] = tiles run so far.
void lzg_distance_tc_timing(const lzg_profiles* p, int64_t* info) {
    info[4] = info[5] = info[6] = 0;
    const lzg_tc_state* tc = p->tc;
    if (!tc) return;
    info[6] = tc->tiles_run;
    float ms = 0.0f;
    if (tc->ev[1] && cudaEventElapsedTime(&ms, tc->ev[0], tc->ev[1]) == cudaSuccess) info[4] = (int64_t)(ms * 1e3f);
    if (tc->ev[3] && tc->launches && cudaEventElapsedTime(&ms, tc->ev[2], tc->ev[3]) == cudaSuccess) info[5] = (int64_t)(ms * 1e3f);
    cudaGetLastError();
}
