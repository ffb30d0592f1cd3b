// lzg_nj.cu -- neighbour-joining seed tree on the GPU, bit-compatible with the Clearcut code the
// reference links (/root/reference/src/clearcut.cxx: NJ_neighbor_joining :925, NJ_min_transform
// :1938, NJ_compute_r :417, NJ_collapse :744, NJ_init_r :1413).
//
// The reference scans the whole remaining upper triangle for the minimum transformed distance
// d(i,j) - (r2[i] + r2[j]) before every join -- O(n^3) float operations on one core, the wall-clock
// bottleneck of `lazac nni` once scoring runs on the GPU.  Here the scan is a bandwidth-bound
// reduction (k_nj_argmin: n^3/6 floats in total, 0.67 TB for n = 10,000) and everything that
// decides rounding and ties is reproduced operation by operation:
//   * float32 arithmetic with the reference's association, no FMA contraction (__fadd_rn/__fsub_rn),
//   * strict '<', first in row-major order: min over (value, i * n0 + j) as one 64-bit key,
//   * the row sums r: per index the same two subtractions and one addition in the same order
//     (they are independent across indices), r2 = r / (size - 3) with IEEE division,
//   * r of the joined row: a SEQUENTIAL float32 sum in the reference's order (i = a+1 .. size-1,
//     then 0 .. a-1, including the harmless term at i = b) -- one thread, O(size) per join,
//   * the index bookkeeping: the new node takes index a, index 0 moves to index b, index 0 is
//     dropped (here: a base offset into the square matrix, as host/seed_tree.cpp does).
// Distances are integers below 2^24 (checked), so their float32 images are exact.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "lzg_common.cuh"

namespace {

__device__ __forceinline__ int ld_acquire_i(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_i(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ uint32_t float_order(float f) {  // monotone map float -> uint32
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// m[i][j] = (float)D[perm[i]][perm[j]], zero diagonal (NJ_shuffle_distance_matrix + dmat parse).
__global__ void k_nj_gather(const int32_t* __restrict__ D, const int32_t* __restrict__ perm, float* __restrict__ m, int n, int ld, int* __restrict__ flag) {
    const int i = blockIdx.y;
    const size_t src = (size_t)perm[i] * n;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < ld; j += gridDim.x * blockDim.x)
    {
        int v = 0;
        if (i != j && j < n) {
            v = D[src + perm[j]];
            if (v < 0 || v >= (1 << 24)) *flag = 1;  // its float32 image would not be exact
        }
        m[(size_t)i * ld + j] = (float)v;
    }
}

// NJ_init_r: r[k] = sequential float32 sum of d(k, j) in increasing j (rows above k first, then
// row k itself); r2[k] = r[k] / (n - 2) for all but the last index, which is never initialised.
__global__ void k_nj_init_r(const float* __restrict__ m, int n, int ld, float* __restrict__ r, float* __restrict__ r2, int* __restrict__ node) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    float s = 0.0f;
    for (int j = 0; j < n; j++) s = __fadd_rn(s, m[(size_t)j * ld + k]);  // column k == row k; the zero diagonal changes nothing
    r[k] = s;
    r2[k] = k < n - 1 ? __fdiv_rn(s, (float)(n - 2)) : 0.0f;
    node[k] = k;
}

// NJ_min_transform over the active triangle (logical indices, physical = logical + base).
__global__ void __launch_bounds__(256)
k_nj_argmin(const float* __restrict__ m, int n0, int ld, int base, int size, const float* __restrict__ r2, unsigned long long* __restrict__ key) {
    unsigned long long best = ~0ull;
    for (int i = blockIdx.x; i < size - 1; i += gridDim.x) {
        const float* row = m + (size_t)(i + base) * ld + base;
        const float r2i = r2[i + base];
        for (int j = i + 1 + threadIdx.x; j < size; j += blockDim.x) {
            const float cur = __fadd_rn(__fsub_rn(__ldg(row + j), __fadd_rn(r2i, r2[j + base])), 0.0f);  // + 0: no negative zero
            const unsigned long long k = ((unsigned long long)float_order(cur) << 32) | ((unsigned)i * (unsigned)n0 + (unsigned)j);
            best = k < best ? k : best;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
        best = other < best ? other : best;
    }
    __shared__ unsigned long long s_best[8];
    if ((threadIdx.x & 31) == 0) s_best[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; w++) best = s_best[w] < best ? s_best[w] : best;
        if (best != ~0ull) atomicMin(key, best);
    }
}

// One join: NJ_compute_r + NJ_collapse + the index-0-to-b relabelling, a single CTA (O(size) work).
__global__ void __launch_bounds__(1024)
k_nj_join(float* __restrict__ m, int n0, int ld, int base, int size, float* __restrict__ r, float* __restrict__ r2, int* __restrict__ node,
          unsigned long long* __restrict__ key, int32_t* __restrict__ left, int32_t* __restrict__ right, int joined) {
    __shared__ int s_a, s_b;
    __shared__ float s_a2, s_b2;
    const int tid = threadIdx.x;
    if (tid == 0) {
        const unsigned idx = (unsigned)(*key & 0xffffffffull);
        const int a = (int)(idx / (unsigned)n0), b = (int)(idx % (unsigned)n0);
        const float dab = m[(size_t)(a + base) * ld + b + base];
        const float r2a = r2[a + base], r2b = r2[b + base];
        s_a = a;
        s_b = b;
        s_a2 = (float)((double)__fadd_rn(dab, __fsub_rn(r2a, r2b)) / 2.0);
        s_b2 = (float)((double)__fadd_rn(dab, __fsub_rn(r2b, r2a)) / 2.0);
        left[joined] = node[a + base];
        right[joined] = node[b + base];
        *key = ~0ull;
    }
    __syncthreads();
    const int a = s_a, b = s_b;
    const float a2 = s_a2, b2 = s_b2;
    const float denom = (float)(size - 3);
    float* row_a = m + (size_t)(a + base) * ld + base;
    float* row_b = m + (size_t)(b + base) * ld + base;
    // Every index but a: two subtractions (NJ_compute_r), the joined distance, one addition
    // (NJ_collapse).  Indices are visited in the order the reference sums the joined row's r --
    // i = a+1 .. size-1, then 0 .. a-1 -- in chunks of blockDim.x staged in shared memory, so
    // that thread 0 adds chunk k one float32 at a time while the others already compute chunk k+1.
    __shared__ float s_cv[2][1024];
    float sum = 0.0f;
    const int n_terms = size - 1;
    for (int p0 = 0, buf = 0; p0 < n_terms; p0 += blockDim.x, buf ^= 1) {
        const int p = p0 + tid;
        if (p < n_terms) {
            int i = a + 1 + p;
            if (i >= size) i -= size;
            const float da = row_a[i], db = row_b[i];  // d(b, b) = 0 at i == b, as the reference reads it
            const float c = (float)((double)__fadd_rn(__fsub_rn(da, a2), __fsub_rn(db, b2)) / 2.0);
            s_cv[buf][tid] = c;
            row_a[i] = c;
            m[(size_t)(i + base) * ld + a + base] = c;
            float ri = __fsub_rn(__fsub_rn(r[i + base], da), db);
            ri = __fadd_rn(ri, c);
            r[i + base] = ri;
            r2[i + base] = __fdiv_rn(ri, denom);
        }
        __syncthreads();  // chunk staged; also: thread 0 has finished adding the previous chunk of this buffer's twin
        if (tid == 0) {
            const int cnt = min((int)blockDim.x, n_terms - p0);
#pragma unroll 8
            for (int q = 0; q < cnt; q++) sum = __fadd_rn(sum, s_cv[buf][q]);
        }
    }
    if (tid == 0) {
        r[a + base] = sum;
        r2[a + base] = __fdiv_rn(sum, denom);
        node[a + base] = joined;
    }
    __syncthreads();
    // index 0 takes over index b; index 0 is dropped by the caller (base + 1)
    const float* row_0 = m + (size_t)base * ld + base;
    for (int i = tid; i < size; i += blockDim.x) {
        if (i == b) continue;
        const float v = row_0[i];
        row_b[i] = v;
        m[(size_t)(i + base) * ld + b + base] = v;
    }
    if (tid == 0) {
        row_b[b] = 0.0f;
        r[b + base] = r[base];
        r2[b + base] = r2[base];
        node[b + base] = node[base];
    }
}

// ---- all joins in ONE launch ----------------------------------------------------------------------------
// Persistent cooperative kernel (one launch instead of two per join): per join a scan phase and an update phase with a grid
// barrier behind each.
//  * scan: the active upper triangle is cut into equal contiguous element ranges, one per CTA (row-major; the first row of a
//    range from the closed form of the row offsets), eight independent loads per thread in flight;
//  * update: thread i computes the joined distance c_i, rewrites row / column a, the row sums, and -- fused -- the relabelling
//    "index 0 takes over index b" for its own column (row b := the updated row 0);
//  * the row sum of the joined node is a SEQUENTIAL float32 sum in the reference's order (i = a+1 .. size-1, 0 .. a-1): one lane
//    adds it while every other warp already scans the NEXT join's triangle, skipping the pairs of the new node; the same warp then
//    scans those pairs itself and joins the barrier.
// Every load of data another CTA may have written goes through L2 (ld.cg): L1 is not coherent across the barriers.
struct lzg_nj_state {
    unsigned bar_count, bar_gen;
    unsigned staged;  // k_nj_fused: CTAs that have read the rows the last join's deferred stores overwrite
    unsigned long long key[2];
    int err;
    // the join's operands, read by the LAST CTA to reach the barrier behind the scan (every minimum is in by then) and before it
    // lets the others go: the update overwrites d(a, b) and r2[b] while other threads would still be reading them
    int a, b;
    float dab, r2a, r2b;
    unsigned long long t_scan, t_wait1, t_update, t_wait2;  // LZG_NJ_DEBUG: ns spent per phase by CTA 1 (summed over the joins)
};

// Grid barrier number `phase` (1, 2, ... -- every CTA counts its own barriers, so nothing is read before arriving): arrivals
// accumulate in bar_count, the CTA that brings it to phase * n_ctas publishes bar_gen = phase.
// snap_m != nullptr: the barrier behind the scan -- the last arriver reads the winning pair and its operands into *st.
__device__ __forceinline__ bool nj_barrier(lzg_nj_state* st, unsigned n_ctas, unsigned phase, const float* snap_m = nullptr, const float* snap_r2 = nullptr,
                                           int n0 = 0, int ld = 0, int base = 0, int parity = 0, const float* snap_r = nullptr, float snap_denom = 1.0f) {
    __syncthreads();
    __shared__ int s_ok;
    if (threadIdx.x == 0) {
        int ok = 1;
        unsigned old;
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(&st->bar_count) : "memory");
        if (old == phase * n_ctas - 1) {
            if (snap_m) {
                const unsigned idx = (unsigned)(__ldcg(&st->key[parity]) & 0xffffffffull);
                const int a = (int)(idx / (unsigned)n0), b = (int)(idx % (unsigned)n0);
                st->a = a, st->b = b;
                st->dab = __ldcg(snap_m + (size_t)(a + base) * ld + b + base);
                if (snap_r2) {
                    st->r2a = __ldcg(snap_r2 + a + base);
                    st->r2b = __ldcg(snap_r2 + b + base);
                } else {  // k_nj_fused keeps r only: r2 = r / (size - 2)
                    st->r2a = __fdiv_rn(__ldcg(snap_r + a + base), snap_denom);
                    st->r2b = __fdiv_rn(__ldcg(snap_r + b + base), snap_denom);
                }
            }
            st_release_i(reinterpret_cast<int*>(&st->bar_gen), (int)phase);
        } else {
            long long spins = 0;
            while ((unsigned)ld_acquire_i(reinterpret_cast<const int*>(&st->bar_gen)) != phase) {
                if (++spins > (1ll << 27)) {  // seconds: a CTA is missing -- give up instead of hanging the device
                    st->err = 1;
                    ok = 0;
                    break;
                }
            }
        }
        s_ok = ok && !__ldcg(&st->err);
    }
    __syncthreads();
    return s_ok != 0;
}

__device__ __forceinline__ unsigned long long nj_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ unsigned long long nj_key(float d, float r2i, float r2j, int i, int j, int n0) {
    const float cur = __fadd_rn(__fsub_rn(d, __fadd_rn(r2i, r2j)), 0.0f);  // + 0: no negative zero
    return ((unsigned long long)float_order(cur) << 32) | ((unsigned)i * (unsigned)n0 + (unsigned)j);
}

// sum + v[0] + v[1] + ... + v[cnt-1], one float32 add after the other (the reference's order), v in shared memory and 16-byte
// aligned.  Software pipeline: the next 32 terms are loaded (8 x LDS.128) while the 32 dependent adds of the current ones run;
// volatile asm keeps ptxas from sinking each load next to its use (it did: 10 instead of 4 cycles per term).
#define NJ_LDS8(dst, addr)                                                                                                                   \
    _Pragma("unroll") for (int u = 0; u < 8; u++) asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"                                  \
                                                               : "=f"(dst[u].x), "=f"(dst[u].y), "=f"(dst[u].z), "=f"(dst[u].w)           \
                                                               : "r"((addr) + 16 * u))
#define NJ_ADD32(sum, src)                                                       \
    _Pragma("unroll") for (int u = 0; u < 8; u++) {                              \
        asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(sum) : "f"(src[u].x));      \
        asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(sum) : "f"(src[u].y));      \
        asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(sum) : "f"(src[u].z));      \
        asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(sum) : "f"(src[u].w));      \
    }
__device__ __forceinline__ float nj_sum_in_order(const float* v_smem, int cnt, float sum) {
    int q = 0;
    while (cnt > 0 && ((unsigned)__cvta_generic_to_shared(v_smem) & 15u)) sum = __fadd_rn(sum, *v_smem++), cnt--;  // up to a 16-byte boundary
    if (cnt >= 64) {
        float4 v[8], w[8];  // two register sets in turn: no copies on the chain
        const unsigned sa = (unsigned)__cvta_generic_to_shared(v_smem);
        NJ_LDS8(v, sa);
#pragma unroll 1
        for (; q + 64 <= cnt; q += 64) {
            NJ_LDS8(w, sa + 4 * (q + 32));
            NJ_ADD32(sum, v);
            NJ_LDS8(v, sa + 4 * (q + 128 <= cnt ? q + 64 : 0));  // the last round re-reads the first block (unused)
            NJ_ADD32(sum, w);
        }
    }
    for (; q < cnt; q++) sum = __fadd_rn(sum, v_smem[q]);
    return sum;
}
#undef NJ_LDS8
#undef NJ_ADD32

// NJ_min_transform over the active triangle, all CTAs but CTA 0; returns this thread's best key.
// Physical indices p = logical + base: the active block is rows / columns [base, n0) whatever the join, so a row's pairs end at
// column n0 - 1 and rows are read as aligned float4 groups (the pitch is a multiple of 4, pad columns hold zeros).  A task is one
// chunk of 128 groups (2 KB) of one row: 32 lanes x 4 independent 16-byte loads.  Rows are paired -- p1 = base + v with
// p2 = n0 - 2 - v -- so that every virtual row has (at most) C chunks and task t -> v = t / C, chunk t % C; warps take tasks round
// robin.  SMEM: r2 comes from shared memory where the pending node and the pad columns hold -inf (their pairs evaluate to +inf).
template <bool SMEM, bool STANDIN = false>
__device__ __noinline__ unsigned long long nj_scan(const float* __restrict__ m, int n0, int ld, int base, int size, int pendp, const float* __restrict__ r2,
                                                      const float* s_r2, int r2_org, int n_scan, int scan_rank, int bp = -1,
                                                      const float* __restrict__ row_standin = nullptr) {
    const float inf = __int_as_float(0x7f800000), ninf = __int_as_float(0xff800000);
    const int lane = threadIdx.x & 31;
    const unsigned CGL = (unsigned)(n0 - 1) >> 2;  // last column group
    const unsigned V = (unsigned)size / 2;         // ceil((size - 1) / 2) virtual rows
    const unsigned C = (2 * CGL + 3 - ((unsigned)(base + n0) >> 2) + 254) >> 7;
    const unsigned T = V * C;
    const unsigned n_warps = (unsigned)n_scan * 16, warp = (unsigned)scan_rank * 16 + (threadIdx.x >> 5);
    unsigned long long best = ~0ull;
    float bestv = inf;
    for (unsigned t = warp; t < T; t += n_warps) {
        const unsigned v = t / C;
        unsigned ch = t - v * C;
        int p = base + (int)v;
        unsigned f = (unsigned)(p + 1) >> 2, nch = (CGL - f + 128) >> 7;
        if (ch >= nch) {  // the partner row's chunks
            ch -= nch;
            const int p2 = n0 - 2 - (int)v;
            if (p2 == p) continue;
            p = p2;
            f = (unsigned)(p + 1) >> 2, nch = (CGL - f + 128) >> 7;
            if (ch >= nch) continue;
        }
        if (p == pendp) continue;
        const unsigned cg0 = f + ch * 128 + lane;
        // bp >= 0 (the fused kernel): physical index bp is taking over the data of the index dropped by the last join, whose row
        // (row_standin) is still in place while the copy into row / column bp is under way -- read it from there
        const float* row = STANDIN && p == bp ? row_standin : m + (size_t)p * ld;
        const float ri = SMEM ? s_r2[p - r2_org] : __ldcg(r2 + p);
        float4 d[4], rj[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const unsigned cg = cg0 + 32 * u;
            if (cg <= CGL) {
                d[u] = __ldcg(reinterpret_cast<const float4*>(row + 4 * cg));
                if (STANDIN && p < bp && (unsigned)bp >> 2 == cg) {  // column bp of this row: the stand-in row's element at column p (symmetry)
                    const float sub = __ldcg(row_standin + p);
                    const int e = bp & 3;
                    if (e == 0) d[u].x = sub;
                    if (e == 1) d[u].y = sub;
                    if (e == 2) d[u].z = sub;
                    if (e == 3) d[u].w = sub;
                }
                rj[u] = SMEM ? *reinterpret_cast<const float4*>(s_r2 + ((int)(4 * cg) - r2_org)) : __ldcg(reinterpret_cast<const float4*>(r2 + 4 * cg));
            } else {
                d[u] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                rj[u] = make_float4(ninf, ninf, ninf, ninf);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int c0 = (int)(4 * (cg0 + 32 * u));
            float c[4];
            c[0] = __fsub_rn(d[u].x, __fadd_rn(ri, rj[u].x));
            c[1] = __fsub_rn(d[u].y, __fadd_rn(ri, rj[u].y));
            c[2] = __fsub_rn(d[u].z, __fadd_rn(ri, rj[u].z));
            c[3] = __fsub_rn(d[u].w, __fadd_rn(ri, rj[u].w));
            if (u == 0 && c0 <= p) {  // the row's first group reaches back to the diagonal
#pragma unroll
                for (int e = 0; e < 4; e++)
                    if (c0 + e <= p) c[e] = inf;
            }
            if (!SMEM) {
#pragma unroll
                for (int e = 0; e < 4; e++)
                    if (c0 + e >= n0 || c0 + e == pendp) c[e] = inf;
            }
            const float g4 = fminf(fminf(c[0], c[1]), fminf(c[2], c[3]));
            if (g4 <= bestv && g4 < inf) {  // rare once a thread has seen a few groups; on a tie the index decides
                const int e4 = c[0] == g4 ? 0 : c[1] == g4 ? 1 : c[2] == g4 ? 2 : 3;  // the first column wins a tie
                const unsigned long long key =
                    ((unsigned long long)float_order(__fadd_rn(g4, 0.0f)) << 32) | ((unsigned)(p - base) * (unsigned)n0 + (unsigned)(c0 + e4 - base));
                if (key < best) best = key, bestv = g4;
            }
        }
    }
    return best;
}

// CTA 0 of k_nj_persistent (early form): the joined row c and the new r2 of every index into shared memory, from the rows a and b
// and the row sums BEFORE the update.  Aligned groups of four physical columns (rows start 16-byte aligned, pad columns hold
// zeros), two groups a thread in flight: three L2 round trips at n = 10,000.  Its own function so that its registers do not
// weigh on the scan loop.
__device__ __noinline__ void nj_stage_join(const float* __restrict__ m, const float* __restrict__ r, int n0, int ld, int base, int size, int a, int b, float dab,
                                           float a2, float b2, float denom, float* s_c, float* s_r2) {
    const int tid = threadIdx.x;
    const float4* pa = reinterpret_cast<const float4*>(m + (size_t)(a + base) * ld);
    const float4* pb = reinterpret_cast<const float4*>(m + (size_t)(b + base) * ld);
    const float4* pr = reinterpret_cast<const float4*>(r);
    const int gbeg = base >> 2, gend = (n0 + 3) >> 2;
    for (int g0 = gbeg + tid; g0 < gend; g0 += 2 * 512) {
        float4 A[2], B[2], R[2];
#pragma unroll
        for (int u = 0; u < 2; u++) {
            const int g = g0 + 512 * u;
            if (g < gend) A[u] = __ldcg(pa + g), B[u] = __ldcg(pb + g), R[u] = __ldcg(pr + g);
        }
#pragma unroll
        for (int u = 0; u < 2; u++) {
            const int g = g0 + 512 * u;
            if (g < gend) {
                const float va[4] = {A[u].x, A[u].y, A[u].z, A[u].w}, vb[4] = {B[u].x, B[u].y, B[u].z, B[u].w}, vr[4] = {R[u].x, R[u].y, R[u].z, R[u].w};
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const int i = 4 * g + e - base;
                    if (i >= 0 && i < size) {
                        const float da = i == b ? dab : va[e], db = i == b ? 0.0f : vb[e];
                        const float c = (float)((double)__fadd_rn(__fsub_rn(da, a2), __fsub_rn(db, b2)) / 2.0);
                        const float ri = __fadd_rn(__fsub_rn(__fsub_rn(vr[e], da), db), c);
                        s_c[i] = i == a ? 0.0f : c;
                        s_r2[i] = __fdiv_rn(ri, denom);
                    }
                }
            }
        }
    }
}

__global__ void __launch_bounds__(512, 2)
k_nj_persistent(float* __restrict__ m, int n0, int ld, float* __restrict__ r, float* __restrict__ r2, int* __restrict__ node, float* __restrict__ cvec,
                int32_t* __restrict__ left, int32_t* __restrict__ right, lzg_nj_state* __restrict__ st, int r2_in_smem, int early, unsigned long long* __restrict__ dbg, int* __restrict__ smid) {
    const int tid = threadIdx.x, n_ctas = gridDim.x;
    // CTA 0 owns the joined node (its row sum and its pairs); the others share the triangle -- except a CTA that landed on CTA 0's
    // SM: it only takes part in the updates, so that the lane adding the row sum (one dependent float add after the other, the
    // serial part of every join) does not compete for issue slots with sixteen scanning warps.
    int n_scan = n_ctas - 1, scan_rank = (int)blockIdx.x - 1;
    unsigned phase = 0;  // barriers passed
    bool scans = blockIdx.x != 0;
    if (smid) {  // n_ctas <= 512
        if (tid == 0) {
            unsigned id;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(id));
            smid[blockIdx.x] = (int)id;
        }
        if (!nj_barrier(st, n_ctas, ++phase)) return;
        const int sm0 = __ldcg(smid);
        const bool other = tid > 0 && tid < n_ctas && __ldcg(smid + tid) != sm0;  // a scanning CTA
        n_scan = __syncthreads_count(other);
        scan_rank = __syncthreads_count(other && tid < (int)blockIdx.x);
        scans = blockIdx.x != 0 && __ldcg(smid + blockIdx.x) != sm0;
        if (n_scan == 0) n_scan = n_ctas - 1, scan_rank = (int)blockIdx.x - 1, scans = blockIdx.x != 0;  // a single SM: nothing to separate
    }
    __shared__ unsigned long long s_best[16];
    __shared__ __align__(16) float s_sum[256];
    __shared__ float s_r2p;
    extern __shared__ __align__(16) float s_r2[];  // r2 of the active indices (logical), when n0 floats fit
    const bool timed = scans && scan_rank == 0 && tid == 0;
    int base = 0, joined = n0, pend = -1;
    int ja = 0, jb = 0;  // the pair joined last, in the numbering it was found in (early)
    float* s_c = s_r2 + ld + 4;  // early, CTA 0: the joined row by the last join's index; s_r2: r2 after that join, same index
    for (int size = n0; size > 2; size--, base++, joined++) {
        const int parity = joined & 1;
        unsigned long long t0 = 0, ts0 = 0;
        if (timed) t0 = nj_now();
        if (dbg && tid == 0) ts0 = nj_now();
        // the node joined last time: its row sum is not known yet (CTA 0 adds it up now)
        // pend: the logical index (in this join's numbering) of the node joined last time; < 0: none
        unsigned long long best = ~0ull;
        if (blockIdx.x == 0) {
            if (pend >= 0) {
                const int terms = size;                    // the join before this one ran at size + 1: size terms,
                const float pend_denom = (float)(size - 2);  // and divided by (size + 1) - 3
                // the pairs of the new node: its row of the (symmetric) matrix.  The operands of the first 2048 are fetched now,
                // before the row sum they wait for is known
                if (early) {
                    // c and r2 of the last join were staged in shared memory right behind the barrier that published the pair (below),
                    // so this chain of adds ran alongside the other CTAs' update, their barrier and their scan
                    if (tid == 0) {
                        float sum = nj_sum_in_order(s_c + ja + 1, size - ja, 0.0f);  // i = a+1 .. size (the last join's numbering), then 0 .. a-1
                        sum = nj_sum_in_order(s_c, ja, sum);
                        const float r2p = __fdiv_rn(sum, pend_denom);
                        r[pend + base] = sum;
                        r2[pend + base] = r2p;
                        s_r2p = r2p;
                        if (dbg) dbg[n_ctas] += nj_now() - ts0;
                    }
                    __syncthreads();
                    const float r2p = s_r2p;
                    const int bq = ja != 0 ? jb - 1 : -1;  // this index took over the data of the dropped index 0
                    for (int q = tid; q < size; q += 512) {
                        if (q == pend) continue;
                        const int src = q == bq ? 0 : q + 1;
                        const unsigned long long k = nj_key(s_c[src], s_r2[src], r2p, min(q, pend), max(q, pend), n0);
                        best = k < best ? k : best;
                    }
                } else {
                const float* row_p = m + (size_t)(pend + base) * ld + base;
                float d0[4], rq0[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int q = tid + 512 * u;
                    const bool ok = q < size && q != pend;
                    d0[u] = ok ? __ldcg(row_p + q) : 0.0f;
                    rq0[u] = ok ? __ldcg(r2 + q + base) : 0.0f;
                }
                if (r2_in_smem) {
                    // every thread of the CTA fetches its share of the terms (independent L2 loads), then one lane adds them in order
#pragma unroll 8
                    for (int c = tid; c < terms; c += 512) s_r2[c] = __ldcg(cvec + c);
                    __syncthreads();
                    if (tid == 0) {
                        const float sum = nj_sum_in_order(s_r2, terms, 0.0f);
                        const float r2p = __fdiv_rn(sum, pend_denom);
                        r[pend + base] = sum;
                        r2[pend + base] = r2p;
                        s_r2p = r2p;
                        if (dbg) dbg[n_ctas] += nj_now() - ts0;
                    }
                } else if (tid < 32) {
                    // the terms come through shared memory in chunks of 256: all 32 lanes fetch the NEXT chunk while lane 0 adds
                    float sum = 0.0f;
                    float nxt[8];
#pragma unroll
                    for (int q = 0; q < 8; q++) nxt[q] = tid + 32 * q < terms ? __ldcg(cvec + tid + 32 * q) : 0.0f;
                    for (int p0 = 0; p0 < terms; p0 += 256) {
#pragma unroll
                        for (int q = 0; q < 8; q++) s_sum[tid + 32 * q] = nxt[q];
                        __syncwarp();
                        const int p1 = p0 + 256;
#pragma unroll
                        for (int q = 0; q < 8; q++) nxt[q] = p1 + tid + 32 * q < terms ? __ldcg(cvec + p1 + tid + 32 * q) : 0.0f;
                        if (tid == 0) sum = nj_sum_in_order(s_sum, min(256, terms - p0), sum);
                        __syncwarp();
                    }
                    if (tid == 0) {
                        const float r2p = __fdiv_rn(sum, pend_denom);
                        r[pend + base] = sum;
                        r2[pend + base] = r2p;
                        s_r2p = r2p;
                        if (dbg) dbg[n_ctas] += nj_now() - ts0;
                    }
                }
                __syncthreads();
                const float r2p = s_r2p;
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int q = tid + 512 * u;
                    if (q < size && q != pend) {
                        const unsigned long long k = nj_key(d0[u], q < pend ? rq0[u] : r2p, q < pend ? r2p : rq0[u], min(q, pend), max(q, pend), n0);
                        best = k < best ? k : best;
                    }
                }
                for (int q0 = tid + 4 * 512; q0 < size; q0 += 4 * 512) {
                    float d[4], rq[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const int q = q0 + 512 * u;
                        const bool ok = q < size && q != pend;
                        d[u] = ok ? __ldcg(row_p + q) : 0.0f;
                        rq[u] = ok ? __ldcg(r2 + q + base) : 0.0f;
                    }
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const int q = q0 + 512 * u;
                        if (q < size && q != pend) {
                            const unsigned long long k = nj_key(d[u], q < pend ? rq[u] : r2p, q < pend ? r2p : rq[u], min(q, pend), max(q, pend), n0);
                            best = k < best ? k : best;
                        }
                    }
                }
                }
            }
        } else if (scans) {
            const int pendp = pend >= 0 ? pend + base : -1;
            const int r2_org = base & ~3;  // s_r2[c - r2_org] = r2 of physical column c
            if (r2_in_smem) {
                // minus infinity for the pending node and the pad columns: d - (r2_i + r2_j) is then +inf and never the minimum
                for (int c = r2_org + tid; c < ld; c += 512) s_r2[c - r2_org] = c >= base && c < n0 && c != pendp ? __ldcg(r2 + c) : __int_as_float(0xff800000);
                __syncthreads();
                best = nj_scan<true>(m, n0, ld, base, size, pendp, r2, s_r2, r2_org, n_scan, scan_rank);
            } else {
                best = nj_scan<false>(m, n0, ld, base, size, pendp, r2, s_r2, r2_org, n_scan, scan_rank);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other < best ? other : best;
        }
        if ((tid & 31) == 0) s_best[tid >> 5] = best;
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < (int)(blockDim.x >> 5); w++) best = s_best[w] < best ? s_best[w] : best;
            if (early && blockIdx.x == 0) {  // the barrier behind the update, which this CTA only signed: the key was reset before it
                long long spins = 0;
                while ((unsigned)ld_acquire_i(reinterpret_cast<const int*>(&st->bar_gen)) != phase)
                    if (++spins > (1ll << 27)) {
                        st->err = 1;
                        break;
                    }
            }
            if (best != ~0ull) atomicMin(&st->key[parity], best);
            if (dbg) dbg[blockIdx.x] += nj_now() - ts0;
        }
        unsigned long long t1 = 0, t2 = 0, t3 = 0;
        if (timed) t1 = nj_now();
        if (!nj_barrier(st, n_ctas, ++phase, m, r2, n0, ld, base, parity)) return;
        if (timed) t2 = nj_now();
        // ---- update: the join (NJ_compute_r + NJ_collapse) and the relabelling, one thread per index ----
        const int a = __ldcg(&st->a), b = __ldcg(&st->b);
        float* row_a = m + (size_t)(a + base) * ld + base;
        float* row_b = m + (size_t)(b + base) * ld + base;
        const float* row_0 = m + (size_t)base * ld + base;
        const float dab = __ldcg(&st->dab);
        const float r2a = __ldcg(&st->r2a), r2b = __ldcg(&st->r2b);
        const float a2 = (float)((double)__fadd_rn(dab, __fsub_rn(r2a, r2b)) / 2.0);
        const float b2 = (float)((double)__fadd_rn(dab, __fsub_rn(r2b, r2a)) / 2.0);
        const float denom = (float)(size - 3);
        // CTA 0 keeps out of the update: ugid over the other CTAs
        const int ugid = ((int)blockIdx.x - 1) * 512 + tid, ustride = (n_ctas - 1) * 512;
        if (early && blockIdx.x == 0) {
            // the joined row c and the new r2 of every index, from the rows a and b and the row sums BEFORE the update (the
            // updating threads hold their stores back until st->staged says these reads are done)
            nj_stage_join(m, r, n0, ld, base, size, a, b, dab, a2, b2, denom, s_c, s_r2);
            __syncthreads();
            ++phase;  // the barrier behind the update: signed here, waited for only before this CTA's next atomicMin
            if (tid == 0) st_release_i(reinterpret_cast<int*>(&st->staged), n0 - size + 1);
            if (tid == 32) {
                unsigned old;
                asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(&st->bar_count) : "memory");
                if (old == phase * n_ctas - 1) st_release_i(reinterpret_cast<int*>(&st->bar_gen), (int)phase);
            }
        } else if (early) {  // one index per thread (the host checked ustride >= n0): loads, the wait, stores
            const int i = ugid;
            const bool act = i < size && i != a;
            float da = 0.0f, db = 0.0f, ro = 0.0f, v0 = 0.0f;
            if (act) {
                da = i == b ? dab : __ldcg(row_a + i), db = i == b ? 0.0f : __ldcg(row_b + i);
                ro = __ldcg(r + i + base);
                v0 = a == 0 ? 0.0f : __ldcg(row_0 + i);
            }
            if (((int)blockIdx.x - 1) * 512 < size) {
                if (tid == 0) {
                    long long spins = 0;
                    while (ld_acquire_i(reinterpret_cast<const int*>(&st->staged)) < n0 - size + 1)
                        if (++spins > (1ll << 27)) {
                            st->err = 1;
                            break;
                        }
                }
                __syncthreads();
            }
            if (act && i != b) {
                const float c = (float)((double)__fadd_rn(__fsub_rn(da, a2), __fsub_rn(db, b2)) / 2.0);
                row_a[i] = c;
                m[(size_t)(i + base) * ld + a + base] = c;
                const float ri = __fadd_rn(__fsub_rn(__fsub_rn(ro, da), db), c);
                r[i + base] = ri;
                const float r2i = __fdiv_rn(ri, denom);
                r2[i + base] = r2i;
                const float v = a == 0 ? c : v0;
                row_b[i] = v;
                m[(size_t)(i + base) * ld + b + base] = v;
                if (i == 0) {
                    row_b[a] = c;
                    row_a[b] = c;
                    r[b + base] = ri;
                    r2[b + base] = r2i;
                }
            }
        } else if (blockIdx.x != 0) {
        for (int i = ugid; i < size; i += ustride) {
            if (i == a) continue;
            // i == b: d(a, b) from the snapshot (thread 0 rewrites that cell below) and d(b, b) = 0, as the reference reads them
            const float da = i == b ? dab : __ldcg(row_a + i), db = i == b ? 0.0f : __ldcg(row_b + i);
            const float c = (float)((double)__fadd_rn(__fsub_rn(da, a2), __fsub_rn(db, b2)) / 2.0);
            cvec[i > a ? i - a - 1 : size - a - 1 + i] = c;  // the order the reference adds them in
            if (i == b) continue;  // index b is overwritten by the relabelling below
            row_a[i] = c;
            m[(size_t)(i + base) * ld + a + base] = c;
            float ri = __fsub_rn(__fsub_rn(__ldcg(r + i + base), da), db);
            ri = __fadd_rn(ri, c);
            r[i + base] = ri;
            const float r2i = __fdiv_rn(ri, denom);
            r2[i + base] = r2i;
            // index 0 (after the update) takes over index b; index 0 itself is dropped by base + 1
            const float v = a == 0 ? c : __ldcg(row_0 + i);
            row_b[i] = v;
            m[(size_t)(i + base) * ld + b + base] = v;
            if (i == 0) {  // a != 0 here: the updated row 0 also holds c_0 at column a
                row_b[a] = c;
                row_a[b] = c;
                r[b + base] = ri;
                r2[b + base] = r2i;
            }
        }
        }
        if (blockIdx.x == 1 && tid == 0) {
            left[joined] = __ldcg(node + a + base);
            right[joined] = __ldcg(node + b + base);
            const int where = a == 0 ? b : a;  // where the joined node lives from now on
            if (a != 0) node[b + base] = __ldcg(node + base);
            node[where + base] = joined;
            st->key[parity ^ 1] = ~0ull;  // the next join's minimum; nobody touches it before the barrier below
        }
        ja = a, jb = b;
        pend = (a == 0 ? b : a) - 1;  // where the joined node lives, in the next join's numbering (base moves on by one)
        if (timed) t3 = nj_now();
        if (!(early && blockIdx.x == 0) && !nj_barrier(st, n_ctas, ++phase)) return;
        if (timed) {
            const unsigned long long t4 = nj_now();
            st->t_scan += t1 - t0, st->t_wait1 += t2 - t1, st->t_update += t3 - t2, st->t_wait2 += t4 - t3;
        }
    }
    if (blockIdx.x == 1 && tid == 0) {  // the CTA that kept the node table, behind the last barrier
        left[joined] = __ldcg(node + base);
        right[joined] = __ldcg(node + base + 1);
    }
}

// The last join's effect on four consecutive physical columns c0 .. c0+3 (k_nj_fused): the joined distance c and the new r2 of
// each, from the rows a and b and the row sums of the state BEFORE the update.  x * 0.5f is the reference's (float)((double)x / 2.0)
// bit for bit (halving is exact, or rounds once into the subnormals either way).
__device__ __forceinline__ void nj_virtual4(const float* __restrict__ rowA, const float* __restrict__ rowB, const float* __restrict__ rin, int c0, float a2, float b2,
                                            float denom, float (&c)[4], float (&r2n)[4]) {
    const float4 da = __ldcg(reinterpret_cast<const float4*>(rowA + c0)), db = __ldcg(reinterpret_cast<const float4*>(rowB + c0));
    const float4 r = __ldcg(reinterpret_cast<const float4*>(rin + c0));
    const float vda[4] = {da.x, da.y, da.z, da.w}, vdb[4] = {db.x, db.y, db.z, db.w}, vr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int e = 0; e < 4; e++) {
        c[e] = __fmul_rn(__fadd_rn(__fsub_rn(vda[e], a2), __fsub_rn(vdb[e], b2)), 0.5f);
        r2n[e] = __fdiv_rn(__fadd_rn(__fsub_rn(__fsub_rn(vr[e], vda[e]), vdb[e]), c[e]), denom);
    }
}

// ---- all joins in one launch, ONE grid barrier per join ------------------------------------------------------
// k_nj_persistent spends about 9 us of every join in its update phase and the barrier behind it.  Here the update of join t-1 is
// VIRTUAL while join t is scanned: once the winning pair (a, b) is known, every CTA derives what it needs from the state before
// the update --
//   * the scanning CTAs compute r2 of every index themselves (the same float operations on r, d(a, i), d(b, i)) while staging it
//     in shared memory, skip the pairs of the joined node (CTA 0 scores them), and read the relabelled index ("index 0 takes over
//     index b") from the dropped row, which nobody overwrites;
//   * CTA 0 computes the joined row c_i into shared memory straight from the old rows a and b, adds the row sum in the
//     reference's order, and scores the joined node's pairs from shared memory;
//   * the threads that own an index i keep (c_i, r_i, d(0, i)) in registers through the scan and store them afterwards -- the
//     in-place update -- once every CTA has announced (st->staged) that it no longer reads the old rows a and b; r is double
//     buffered.  The barrier that ends the scan publishes the stores together with the minimum.
// At every barrier the global state is exactly the reference's state after the join, so the last arriver snapshots the next pair.
__global__ void __launch_bounds__(512, 2)
k_nj_fused(float* __restrict__ m, int n0, int ld, float* __restrict__ rb, const float* __restrict__ r2_init, int* __restrict__ node,
           int32_t* __restrict__ left, int32_t* __restrict__ right, lzg_nj_state* __restrict__ st, unsigned long long* __restrict__ dbg, int* __restrict__ smid) {
    const int tid = threadIdx.x, n_ctas = gridDim.x;
    const float ninf = __int_as_float(0xff800000);
    int n_scan = n_ctas - 1, scan_rank = (int)blockIdx.x - 1;
    unsigned phase = 0;  // barriers passed
    bool scans = blockIdx.x != 0;
    if (smid) {  // n_ctas <= 512: the CTA sharing CTA 0's SM does not scan (see k_nj_persistent)
        if (tid == 0) {
            unsigned id;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(id));
            smid[blockIdx.x] = (int)id;
        }
        if (!nj_barrier(st, n_ctas, ++phase)) return;
        const int sm0 = __ldcg(smid);
        const bool other = tid > 0 && tid < n_ctas && __ldcg(smid + tid) != sm0;
        n_scan = __syncthreads_count(other);
        scan_rank = __syncthreads_count(other && tid < (int)blockIdx.x);
        scans = blockIdx.x != 0 && __ldcg(smid + blockIdx.x) != sm0;
        if (n_scan == 0) n_scan = n_ctas - 1, scan_rank = (int)blockIdx.x - 1, scans = blockIdx.x != 0;
    }
    __shared__ unsigned long long s_best[16];
    __shared__ float s_r2p;
    extern __shared__ __align__(16) float s_dyn[];
    float* s_r2 = s_dyn;           // scanning CTAs: r2 by physical column; CTA 0: r2 by the index numbering of the last join
    float* s_c = s_dyn + ld + 4;   // CTA 0: the joined row, same numbering
    // the last join (none before the first scan): the pair in the numbering it was found in, and its operands
    int a = -1, b = -1;
    float dab = 0.0f, r2a = 0.0f, r2b = 0.0f;
    int base = 0, joined = n0;
    for (int size = n0; size > 2; size--, base++, joined++) {
        const int parity = joined & 1, t = n0 - size;
        const bool prev = t > 0;
        const float* rin = rb + (size_t)((t + 1) & 1) * ld;  // r before the last join's update
        float* rout = rb + (size_t)(t & 1) * ld;             // r after it (t = 0: the initial sums)
        const int Bp = base - 1;                             // the base the last join ran at
        const int pend = prev ? (a == 0 ? b : a) - 1 : -1;   // the joined node and
        const int bq = prev && a != 0 ? b - 1 : -1;          // the index that takes over the dropped one, in THIS join's numbering
        const float a2 = (float)((double)__fadd_rn(dab, __fsub_rn(r2a, r2b)) / 2.0);
        const float b2 = (float)((double)__fadd_rn(dab, __fsub_rn(r2b, r2a)) / 2.0);
        const float denom = (float)(size - 2);  // (size + 1) - 3: the last join ran at size + 1
        const float* rowA = prev ? m + (size_t)(a + Bp) * ld : m;  // rows a, b and 0 of the last join's numbering, by physical column
        const float* rowB = prev ? m + (size_t)(b + Bp) * ld : m;
        const float* row0 = prev ? m + (size_t)Bp * ld : m;
        unsigned long long ts0 = 0;
        if (dbg && tid == 0) ts0 = nj_now();
        unsigned long long best = ~0ull;
        // ---- this thread's share of the in-place update: index i of the last join's numbering, held in registers until the scan is over
        const int ui = blockIdx.x * 512 + tid;
        const bool upd = prev && ui <= size && ui != a;
        float u_da = 0.0f, u_db = 0.0f, u_r = 0.0f, u_v = 0.0f;
        if (upd) {
            u_da = __ldcg(rowA + ui + Bp);  // at ui == b: d(a, b) and the zero diagonal d(b, b), as the reference reads them
            u_db = __ldcg(rowB + ui + Bp);
            u_r = __ldcg(rin + ui + Bp);
            u_v = a == 0 ? 0.0f : __ldcg(row0 + ui + Bp);
        }
        if (blockIdx.x == 0) {
            if (prev) {
                // the joined row and r2 of every index, in the last join's numbering (i = c - Bp = 0 .. size), four columns a thread
                for (int g0 = (Bp >> 2) + tid; 4 * g0 < n0; g0 += 2 * 512) {
                    float c[2][4], r2n[2][4];
#pragma unroll
                    for (int u = 0; u < 2; u++)
                        if (4 * (g0 + 512 * u) < n0) nj_virtual4(rowA, rowB, rin, 4 * (g0 + 512 * u), a2, b2, denom, c[u], r2n[u]);
#pragma unroll
                    for (int u = 0; u < 2; u++)
#pragma unroll
                        for (int e = 0; e < 4; e++) {
                            const int i = 4 * (g0 + 512 * u) + e - Bp;
                            if (i >= 0 && i <= size) s_c[i] = i == a ? 0.0f : c[u][e], s_r2[i] = r2n[u][e];
                        }
                }
                __syncthreads();
                if (tid == 0) {
                    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(&st->staged) : "memory");
                    // the reference's order: i = a+1 .. size, then 0 .. a-1 (the term at i = b included)
                    float sum = nj_sum_in_order(s_c + a + 1, size - a, 0.0f);
                    sum = nj_sum_in_order(s_c, a, sum);
                    rout[pend + base] = sum;
                    s_r2p = __fdiv_rn(sum, denom);
                    if (dbg) dbg[n_ctas] += nj_now() - ts0;
                }
                if (tid == 32) {  // the tree: children of the new node, then the relabelling
                    left[joined - 1] = node[a + Bp];
                    right[joined - 1] = node[b + Bp];
                    if (a != 0) node[b + Bp] = node[Bp];
                    node[(a == 0 ? b : a) + Bp] = joined - 1;
                }
                __syncthreads();
                const float r2p = s_r2p;
                for (int q = tid; q < size; q += 512) {  // the pairs of the joined node, from shared memory
                    if (q == pend) continue;
                    const int src = q == bq ? 0 : q + 1;
                    const unsigned long long k = nj_key(s_c[src], s_r2[src], r2p, min(q, pend), max(q, pend), n0);
                    best = k < best ? k : best;
                }
            }
        } else {
            const int r2_org = base & ~3;  // s_r2[c - r2_org] = r2 of physical column c
            if (prev) {
                const int cb = bq >= 0 ? bq + base : -1;  // this column takes the values of the dropped index (physical column Bp)
                for (int g0 = (r2_org >> 2) + tid; 4 * g0 < ld; g0 += 2 * 512) {
                    float c[2][4], r2n[2][4];
                    float sub = 0.0f;
#pragma unroll
                    for (int u = 0; u < 2; u++) {
                        const int c0 = 4 * (g0 + 512 * u);
                        if (c0 < ld) nj_virtual4(rowA, rowB, rin, c0, a2, b2, denom, c[u], r2n[u]);
                        if (cb >= c0 && cb < c0 + 4) {
                            const float da = __ldcg(rowA + Bp), db = __ldcg(rowB + Bp);
                            const float cc = __fmul_rn(__fadd_rn(__fsub_rn(da, a2), __fsub_rn(db, b2)), 0.5f);
                            sub = __fdiv_rn(__fadd_rn(__fsub_rn(__fsub_rn(__ldcg(rin + Bp), da), db), cc), denom);
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 2; u++)
#pragma unroll
                        for (int e = 0; e < 4; e++) {
                            const int col = 4 * (g0 + 512 * u) + e;
                            if (col < ld) s_r2[col - r2_org] = col < base || col >= n0 || col - base == pend ? ninf : col == cb ? sub : r2n[u][e];
                        }
                }
            } else {
                for (int c = r2_org + tid; c < ld; c += 512) s_r2[c - r2_org] = c >= base && c < n0 ? __ldcg(r2_init + c) : ninf;
            }
            __syncthreads();
            if (prev && tid == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(&st->staged) : "memory");
            if (dbg && tid == 0) dbg[n_ctas + 1 + blockIdx.x] += nj_now() - ts0;
            if (scans) best = nj_scan<true, true>(m, n0, ld, base, size, pend >= 0 ? pend + base : -1, nullptr, s_r2, r2_org, n_scan, scan_rank, bq >= 0 ? bq + base : -1, prev ? m + (size_t)Bp * ld : m);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other < best ? other : best;
        }
        if ((tid & 31) == 0) s_best[tid >> 5] = best;
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < 16; w++) best = s_best[w] < best ? s_best[w] : best;
            if (best != ~0ull) atomicMin(&st->key[parity], best);
            if (dbg) dbg[blockIdx.x] += nj_now() - ts0;
        }
        // ---- the deferred stores of the last join: NJ_compute_r + NJ_collapse + the relabelling (k_nj_persistent's update phase)
        if (prev && blockIdx.x * 512 <= size) {
            if (tid == 0) {  // every CTA has staged what it needs of the old rows a and b
                const unsigned want = (unsigned)t * (unsigned)n_ctas;
                long long spins = 0;
                while ((int)((unsigned)ld_acquire_i(reinterpret_cast<const int*>(&st->staged)) - want) < 0)
                    if (++spins > (1ll << 27)) {
                        st->err = 1;
                        break;
                    }
            }
            __syncthreads();
            if (upd) {
                const int i = ui;
                const float c = __fmul_rn(__fadd_rn(__fsub_rn(u_da, a2), __fsub_rn(u_db, b2)), 0.5f);
                if (i != b) {  // index b is overwritten by the relabelling
                    float* row_a = m + (size_t)(a + Bp) * ld + Bp;
                    float* row_b = m + (size_t)(b + Bp) * ld + Bp;
                    row_a[i] = c;
                    m[(size_t)(i + Bp) * ld + a + Bp] = c;
                    const float ri = __fadd_rn(__fsub_rn(__fsub_rn(u_r, u_da), u_db), c);
                    rout[i + Bp] = ri;
                    // index 0 (after the update) takes over index b; index 0 itself is dropped
                    const float v = a == 0 ? c : u_v;
                    row_b[i] = v;
                    m[(size_t)(i + Bp) * ld + b + Bp] = v;
                    if (i == 0) {  // a != 0 here: the updated row 0 also holds c_0 at column a
                        row_b[a] = c;
                        row_a[b] = c;
                        rout[b + Bp] = ri;
                    }
                }
            }
        }
        if (blockIdx.x == 0 && tid == 0) st->key[parity ^ 1] = ~0ull;  // the next join's minimum
        // the initial r2 leaves its last index at zero (NJ_init_r), so the first join snapshots r2 itself
        if (!nj_barrier(st, n_ctas, ++phase, m, prev ? nullptr : r2_init, n0, ld, base, parity, rout, (float)(size - 2))) return;
        a = __ldcg(&st->a), b = __ldcg(&st->b);
        dab = __ldcg(&st->dab), r2a = __ldcg(&st->r2a), r2b = __ldcg(&st->r2b);
    }
    // the last join's bookkeeping and the final edge between the two nodes left
    if (blockIdx.x == 0 && tid == 0) {
        const int Bp = base - 1;
        if (n0 > 2) {
            left[joined - 1] = node[a + Bp];
            right[joined - 1] = node[b + Bp];
            if (a != 0) node[b + Bp] = node[Bp];
            node[(a == 0 ? b : a) + Bp] = joined - 1;
        }
        left[joined] = node[base];
        right[joined] = node[base + 1];
    }
}

__global__ void k_nj_finish(const int* __restrict__ node, int base, int32_t* __restrict__ left, int32_t* __restrict__ right, int joined) {
    left[joined] = node[base];
    right[joined] = node[base + 1];
}

}  // namespace

// D_host or d_D_in (device, n x n int32, left untouched): one of the two
static int nj_seed_tree_impl(lzg_ctx* ctx, const int32_t* D_host, const int32_t* d_D_in, int32_t n, const int32_t* perm, int32_t* left, int32_t* right) {
    if (!ctx || (!D_host && !d_D_in) || !perm || !left || !right || n < 1 || n > 65535) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_nj_seed_tree: bad argument");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    const int total = 2 * n - 1;
    for (int i = 0; i < n; i++) left[i] = right[i] = -1;
    if (n == 1) return LZG_OK;
    {
        std::vector<char> seen(n, 0);
        for (int i = 0; i < n; i++) {
            if (perm[i] < 0 || perm[i] >= n || seen[perm[i]]) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_nj_seed_tree: perm is not a permutation");
            seen[perm[i]] = 1;
        }
    }
    int32_t *d_D = nullptr, *d_perm = nullptr, *d_left = nullptr, *d_right = nullptr;
    int* d_node = nullptr;
    float *d_m = nullptr, *d_r = nullptr, *d_r2 = nullptr;
    unsigned long long* d_key = nullptr;
    float* d_cvec = nullptr;
    lzg_nj_state* d_state = nullptr;
    unsigned long long* d_dbg = nullptr;
    int* d_smid = nullptr;
    const size_t nn = (size_t)n * n;
    const int ld = (n + 3) & ~3;  // row pitch of the float matrix: rows start 16-byte aligned
    auto release = [&]() {
        if (!d_D_in) lzg_dev_free(ctx, d_D);
        lzg_dev_free(ctx, d_perm);
        lzg_dev_free(ctx, d_left);
        lzg_dev_free(ctx, d_right);
        lzg_dev_free(ctx, d_node);
        lzg_dev_free(ctx, d_m);
        lzg_dev_free(ctx, d_r);
        lzg_dev_free(ctx, d_r2);
        lzg_dev_free(ctx, d_key);
        lzg_dev_free(ctx, d_cvec);
        lzg_dev_free(ctx, d_state);
        lzg_dev_free(ctx, d_dbg);
        lzg_dev_free(ctx, d_smid);
    };
    const auto t_begin = std::chrono::steady_clock::now();
    cudaError_t e = cudaSuccess;
    if (d_D_in) d_D = const_cast<int32_t*>(d_D_in);
    else e = lzg_dev_alloc(ctx, (void**)&d_D, sizeof(int32_t) * nn);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_m, sizeof(float) * (size_t)n * ld);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_perm, sizeof(int32_t) * n);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_left, sizeof(int32_t) * total);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_right, sizeof(int32_t) * total);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_node, sizeof(int) * n);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_r, sizeof(float) * 2 * ld);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_r2, sizeof(float) * ld);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_key, sizeof(unsigned long long));
    if (e == cudaSuccess && !d_D_in) e = cudaMemcpyAsync(d_D, D_host, sizeof(int32_t) * nn, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(ctx->d_flag, 0, sizeof(int), ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_perm, perm, sizeof(int32_t) * n, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_key, 0xff, sizeof(unsigned long long), ctx->stream);
    if (e != cudaSuccess) {
        release();
        return lzg_fail(ctx, LZG_ERR_CUDA, "lzg_nj_seed_tree: setup", e);
    }
    k_nj_gather<<<dim3(std::max(1, std::min(32, (n + 255) / 256)), n), 256, 0, ctx->stream>>>(d_D, d_perm, d_m, n, ld, ctx->d_flag);
    ctx->launches++;
    k_nj_init_r<<<(n + 127) / 128, 128, 0, ctx->stream>>>(d_m, n, ld, d_r, d_r2, d_node);
    ctx->launches++;
    // the range check rode along with the gather (a host loop over n^2 values cost 57 ms at n = 10,000)
    e = cudaMemcpyAsync(ctx->h_pinned, ctx->d_flag, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess || ctx->h_pinned[0]) {
        release();
        return e != cudaSuccess ? lzg_fail(ctx, LZG_ERR_CUDA, "lzg_nj_seed_tree: gather", e) : lzg_fail(ctx, LZG_ERR_RANGE, "distance does not fit float32 exactly");
    }
    if (getenv("LZG_NJ_DEBUG")) {
        cudaStreamSynchronize(ctx->stream);
        fprintf(stderr, "[lzg nj] upload + gather + row sums: %.1f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count());
    }
    // one persistent cooperative launch for all joins; LZG_NJ_PERSISTENT=0 keeps the two-launches-per-join form (the cross-check)
    // LZG_NJ_PERSISTENT: 0 = two launches per join, 1 = k_nj_persistent (two grid barriers per join), otherwise k_nj_fused (one)
    // where its shared memory fits -- the older forms stay as cross-checks and for very large n
    // (2 forces it).  Measured on B200: fused 0.031 / 0.116 / 0.209 s at n = 2,000 / 5,000 / 7,000 against 0.036 / 0.123 / 0.213 s,
    // but 0.455 s against 0.385 s at n = 10,000, where every CTA recomputing r2 costs more than the barrier it saves.
    const int mode = getenv("LZG_NJ_PERSISTENT") ? atoi(getenv("LZG_NJ_PERSISTENT")) : 3;
    const bool persistent = mode != 0;
    bool fused = mode != 1 && (mode == 2 || n <= 8000) && sizeof(float) * 2 * (size_t)(ld + 4) <= 110 * 1024;  // two CTAs per SM: n <= 14,000
    bool done = false;
    if (persistent && n >= 4) {
        int per_sm = 0;
        const size_t smem_one = sizeof(float) * (size_t)(ld + 4);  // r2 by column (scanning CTAs) / the joined row (CTA 0)
        const size_t smem_fused = 2 * smem_one;
        const bool no_smem = getenv("LZG_NJ_SMEM") && atoi(getenv("LZG_NJ_SMEM")) == 0;  // tests: the path of very large n
        if (fused) {
            if (smem_fused > 48 * 1024) cudaFuncSetAttribute(k_nj_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_fused);
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_nj_fused, 512, smem_fused) != cudaSuccess || per_sm < 1) fused = false, cudaGetLastError();
        }
        // k_nj_persistent, early form (CTA 0 stages the joined row and the new r2 itself and adds the row sum while the others
        // update): two arrays of shared memory, two CTAs per SM -> n <= 14,000; LZG_NJ_EARLY=0 keeps the form for larger n
        bool early = !no_smem && smem_fused <= 110 * 1024 && !(getenv("LZG_NJ_EARLY") && atoi(getenv("LZG_NJ_EARLY")) == 0);
        size_t smem_r2 = early ? smem_fused : smem_one;
        if (smem_r2 > 110 * 1024 || no_smem) smem_r2 = 0;  // beyond n = 28,000 r2 and the joined row's terms come from L2
        if (smem_r2 > 48 * 1024) cudaFuncSetAttribute(k_nj_persistent, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_r2);
        cudaError_t ce = fused ? cudaSuccess : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_nj_persistent, 512, smem_r2);
        if (ce == cudaSuccess && per_sm >= 1) ce = lzg_dev_alloc(ctx, (void**)&d_cvec, sizeof(float) * (size_t)n);
        if (ce == cudaSuccess && per_sm >= 1) ce = lzg_dev_alloc(ctx, (void**)&d_state, sizeof(lzg_nj_state));
        const int max_grid = ctx->sm_count * 2 + 1;
        if (ce == cudaSuccess && per_sm >= 1 && max_grid <= 512) ce = lzg_dev_alloc(ctx, (void**)&d_smid, sizeof(int) * max_grid);
        if (ce == cudaSuccess && per_sm >= 1 && getenv("LZG_NJ_DEBUG")) {
            ce = lzg_dev_alloc(ctx, (void**)&d_dbg, sizeof(unsigned long long) * (2 * max_grid + 2));
            if (ce == cudaSuccess) ce = cudaMemsetAsync(d_dbg, 0, sizeof(unsigned long long) * (2 * max_grid + 2), ctx->stream);
        }
        if (ce == cudaSuccess && per_sm >= 1) {
            lzg_nj_state init;
            memset(&init, 0, sizeof(init));
            init.key[0] = init.key[1] = ~0ull;
            ce = cudaMemcpyAsync(d_state, &init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream);
            int grid = std::min(ctx->sm_count * std::min(per_sm, 2), 1 + std::max(1, n / 2));
            grid = std::max(grid, 2);  // CTA 0 owns the joined node, the others scan
            if ((long long)(grid - 1) * 512 < n) early = false;  // one index per updating thread
            int n_arg = n, ld_arg = ld, in_smem = smem_r2 > 0, early_arg = early;
            void* args[] = {(void*)&d_m, (void*)&n_arg, (void*)&ld_arg, (void*)&d_r, (void*)&d_r2, (void*)&d_node, (void*)&d_cvec, (void*)&d_left, (void*)&d_right, (void*)&d_state, (void*)&in_smem, (void*)&early_arg, (void*)&d_dbg, (void*)&d_smid};
            void* fargs[] = {(void*)&d_m, (void*)&n_arg, (void*)&ld_arg, (void*)&d_r, (void*)&d_r2, (void*)&d_node, (void*)&d_left, (void*)&d_right, (void*)&d_state, (void*)&d_dbg, (void*)&d_smid};
            if (ce == cudaSuccess)
                ce = fused ? cudaLaunchCooperativeKernel((void*)k_nj_fused, dim3(grid), dim3(512), fargs, smem_fused, ctx->stream)
                           : cudaLaunchCooperativeKernel((void*)k_nj_persistent, dim3(grid), dim3(512), args, smem_r2, ctx->stream);
            if (ce == cudaSuccess) {
                ctx->launches++;
                lzg_nj_state back;
                ce = cudaMemcpyAsync(&back, d_state, sizeof(back), cudaMemcpyDeviceToHost, ctx->stream);
                if (ce == cudaSuccess) ce = cudaStreamSynchronize(ctx->stream);
                if (ce == cudaSuccess && getenv("LZG_NJ_DEBUG") && !fused)
                    fprintf(stderr, "[lzg nj] n %d, %d CTAs: scan %.1f ms, wait behind the scan %.1f ms, update %.1f ms, wait behind the update %.1f ms\n", n, grid,
                            back.t_scan * 1e-6, back.t_wait1 * 1e-6, back.t_update * 1e-6, back.t_wait2 * 1e-6);
                if (ce == cudaSuccess && d_dbg) {
                    fprintf(stderr, "[lzg nj] n %d: %s, %d CTAs of 512 threads (%d per SM)\n", n, fused ? "k_nj_fused" : early ? "k_nj_persistent (early)" : "k_nj_persistent", grid, per_sm);
                    std::vector<unsigned long long> h(2 * grid + 2);
                    cudaMemcpy(h.data(), d_dbg, sizeof(unsigned long long) * (2 * grid + 2), cudaMemcpyDeviceToHost);
                    if (fused) fprintf(stderr, "[lzg nj] of the scan phase, staging r2: %.1f ms (CTA 1), %.1f ms (CTA %d)\n", h[grid + 2] * 1e-6, h[2 * grid] * 1e-6, grid - 1);
                    unsigned long long lo = ~0ull, hi = 0, sum = 0;
                    int arg = 1;
                    for (int c = 1; c < grid; c++) {
                        lo = std::min(lo, h[c]), sum += h[c];
                        if (h[c] > hi) hi = h[c], arg = c;
                    }
                    fprintf(stderr, "[lzg nj] scan phase per CTA: min %.1f, mean %.1f, max %.1f ms (CTA %d); CTA 0: %.1f ms of which the row sum %.1f ms\n", lo * 1e-6,
                            sum * 1e-6 / std::max(1, grid - 1), hi * 1e-6, arg, h[0] * 1e-6, h[grid] * 1e-6);
                }
                if (ce == cudaSuccess && back.err) {
                    release();
                    return lzg_fail(ctx, LZG_ERR_STATE, "lzg_nj_seed_tree: a grid barrier timed out");
                }
                done = ce == cudaSuccess;
            }
        }
        if (!done) cudaGetLastError();  // cooperative launch refused: the per-join launches below
    }
    if (!done) {
        int base = 0, joined = n;
        for (int size = n; size > 2; size--, base++, joined++) {
            const int grid = std::max(1, std::min(size - 1, ctx->sm_count * 8));
            k_nj_argmin<<<grid, 256, 0, ctx->stream>>>(d_m, n, ld, base, size, d_r2, d_key);
            k_nj_join<<<1, 1024, 0, ctx->stream>>>(d_m, n, ld, base, size, d_r, d_r2, d_node, d_key, d_left, d_right, joined);
            ctx->launches += 2;
        }
        k_nj_finish<<<1, 1, 0, ctx->stream>>>(d_node, base, d_left, d_right, joined);
        ctx->launches++;
    }
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(left + n, d_left + n, sizeof(int32_t) * (total - n), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(right + n, d_right + n, sizeof(int32_t) * (total - n), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    release();
    if (e != cudaSuccess) return lzg_fail(ctx, LZG_ERR_CUDA, "lzg_nj_seed_tree", e);
    return LZG_OK;
}

extern "C" int lzg_nj_seed_tree(lzg_ctx* ctx, const int32_t* D_host, int32_t n, const int32_t* perm, int32_t* left, int32_t* right) {
    if (!D_host) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_nj_seed_tree: bad argument");
    return nj_seed_tree_impl(ctx, D_host, nullptr, n, perm, left, right);
}

// The matrix already in HBM (d_D: n x n int32, e.g. what lzg_distance_matrix_device_async left there): no upload.
extern "C" int lzg_nj_seed_tree_device(lzg_ctx* ctx, const int32_t* d_D, int32_t n, const int32_t* perm, int32_t* left, int32_t* right) {
    if (!d_D) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_nj_seed_tree_device: bad argument");
    return nj_seed_tree_impl(ctx, nullptr, d_D, n, perm, left, right);
}
