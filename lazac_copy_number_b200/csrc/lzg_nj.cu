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
__global__ void k_nj_gather(const int32_t* __restrict__ D, const int32_t* __restrict__ perm, float* __restrict__ m, int n) {
    const int i = blockIdx.y;
    const size_t src = (size_t)perm[i] * n;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x)
        m[(size_t)i * n + j] = i == j ? 0.0f : (float)D[src + perm[j]];
}

// NJ_init_r: r[k] = sequential float32 sum of d(k, j) in increasing j (rows above k first, then
// row k itself); r2[k] = r[k] / (n - 2) for all but the last index, which is never initialised.
__global__ void k_nj_init_r(const float* __restrict__ m, int n, float* __restrict__ r, float* __restrict__ r2, int* __restrict__ node) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    float s = 0.0f;
    for (int j = 0; j < n; j++) s = __fadd_rn(s, m[(size_t)j * n + k]);  // column k == row k; the zero diagonal changes nothing
    r[k] = s;
    r2[k] = k < n - 1 ? __fdiv_rn(s, (float)(n - 2)) : 0.0f;
    node[k] = k;
}

// NJ_min_transform over the active triangle (logical indices, physical = logical + base).
__global__ void __launch_bounds__(256)
k_nj_argmin(const float* __restrict__ m, int n0, int base, int size, const float* __restrict__ r2, unsigned long long* __restrict__ key) {
    unsigned long long best = ~0ull;
    for (int i = blockIdx.x; i < size - 1; i += gridDim.x) {
        const float* row = m + (size_t)(i + base) * n0 + base;
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
k_nj_join(float* __restrict__ m, int n0, int base, int size, float* __restrict__ r, float* __restrict__ r2, int* __restrict__ node,
          unsigned long long* __restrict__ key, int32_t* __restrict__ left, int32_t* __restrict__ right, int joined) {
    __shared__ int s_a, s_b;
    __shared__ float s_a2, s_b2;
    const int tid = threadIdx.x;
    if (tid == 0) {
        const unsigned idx = (unsigned)(*key & 0xffffffffull);
        const int a = (int)(idx / (unsigned)n0), b = (int)(idx % (unsigned)n0);
        const float dab = m[(size_t)(a + base) * n0 + b + base];
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
    float* row_a = m + (size_t)(a + base) * n0 + base;
    float* row_b = m + (size_t)(b + base) * n0 + base;
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
            m[(size_t)(i + base) * n0 + a + base] = c;
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
    const float* row_0 = m + (size_t)base * n0 + base;
    for (int i = tid; i < size; i += blockDim.x) {
        if (i == b) continue;
        const float v = row_0[i];
        row_b[i] = v;
        m[(size_t)(i + base) * n0 + b + base] = v;
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
    unsigned long long key[2];
    int pend_phys, pend_terms, err;
    float pend_denom;
    // the join's operands, read by the LAST CTA to reach the barrier behind the scan (every minimum is in by then) and before it
    // lets the others go: the update overwrites d(a, b) and r2[b] while other threads would still be reading them
    int a, b;
    float dab, r2a, r2b;
    unsigned long long t_scan, t_wait1, t_update, t_wait2;  // LZG_NJ_DEBUG: ns spent per phase by CTA 1 (summed over the joins)
};

// snap != nullptr: the barrier behind the scan -- the last arriver reads the winning pair and its operands into *st.
__device__ __forceinline__ bool nj_barrier(lzg_nj_state* st, unsigned n_ctas, const float* snap_m = nullptr, const float* snap_r2 = nullptr, int n0 = 0,
                                           int base = 0, int parity = 0) {
    __syncthreads();
    __shared__ int s_ok;
    if (threadIdx.x == 0) {
        int ok = 1;
        const unsigned gen = (unsigned)ld_acquire_i(reinterpret_cast<const int*>(&st->bar_gen));
        __threadfence();
        if (atomicAdd(&st->bar_count, 1u) == n_ctas - 1) {
            if (snap_m) {
                __threadfence();
                const unsigned idx = (unsigned)(__ldcg(&st->key[parity]) & 0xffffffffull);
                const int a = (int)(idx / (unsigned)n0), b = (int)(idx % (unsigned)n0);
                st->a = a, st->b = b;
                st->dab = __ldcg(snap_m + (size_t)(a + base) * n0 + b + base);
                st->r2a = __ldcg(snap_r2 + a + base);
                st->r2b = __ldcg(snap_r2 + b + base);
            }
            st->bar_count = 0;
            __threadfence();
            st_release_i(reinterpret_cast<int*>(&st->bar_gen), (int)(gen + 1));
        } else {
            long long spins = 0;
            while ((unsigned)ld_acquire_i(reinterpret_cast<const int*>(&st->bar_gen)) == gen) {
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

__global__ void __launch_bounds__(512, 2)
k_nj_persistent(float* __restrict__ m, int n0, float* __restrict__ r, float* __restrict__ r2, int* __restrict__ node, float* __restrict__ cvec,
                int32_t* __restrict__ left, int32_t* __restrict__ right, lzg_nj_state* __restrict__ st, int r2_in_smem) {
    const int tid = threadIdx.x, n_ctas = gridDim.x;
    const int n_scan = n_ctas - 1;  // CTA 0 owns the joined node (its row sum and its pairs); the others share the triangle
    __shared__ unsigned long long s_best[16];
    __shared__ __align__(16) float s_sum[256];
    __shared__ float s_r2p;
    extern __shared__ float s_r2[];  // r2 of the active indices (logical), when n0 floats fit
    const bool timed = blockIdx.x == 1 && tid == 0;
    int base = 0, joined = n0;
    for (int size = n0; size > 2; size--, base++, joined++) {
        const int parity = joined & 1;
        unsigned long long t0 = 0;
        if (timed) t0 = nj_now();
        // the node joined last time: its row sum is not known yet (CTA 0 adds it up now)
        const int pend = __ldcg(&st->pend_phys) - base;  // logical index now; < 0: none
        unsigned long long best = ~0ull;
        if (blockIdx.x == 0) {
            if (pend >= 0) {
                if (tid < 32) {
                    const int terms = __ldcg(&st->pend_terms);
                    float sum = 0.0f;
                    // the terms come through shared memory in chunks of 256: all 32 lanes fetch the NEXT chunk (independent L2 loads)
                    // while lane 0 adds the current one in order -- the chain of dependent float adds is then the only cost
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
                        if (tid == 0) {
                            const int cnt = min(256, terms - p0);
                            int q = 0;
                            for (; q + 4 <= cnt; q += 4) {
                                const float4 v = *reinterpret_cast<const float4*>(s_sum + q);
                                sum = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(sum, v.x), v.y), v.z), v.w);
                            }
                            for (; q < cnt; q++) sum = __fadd_rn(sum, s_sum[q]);
                        }
                        __syncwarp();
                    }
                    if (tid == 0) {
                        const float r2p = __fdiv_rn(sum, __ldcg(&st->pend_denom));
                        r[pend + base] = sum;
                        r2[pend + base] = r2p;
                        s_r2p = r2p;
                    }
                }
                __syncthreads();
                // the pairs of the new node: its row of the (symmetric) matrix, the whole CTA, every load independent
                const float r2p = s_r2p;
                const float* row_p = m + (size_t)(pend + base) * n0 + base;
                for (int q0 = tid; q0 < size; q0 += 4 * 512) {
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
        } else {
            // The triangle as V = ceil((size-1)/2) rows of exactly `size` elements: row v (size-1-v pairs) followed by row
            // size-2-v (v+1 pairs).  Element e of that rectangle -> v = e / size, k = e % size; every CTA takes an equal
            // contiguous range of e, every thread eight independent loads at a time -- short rows cost no extra round trips.
            if (r2_in_smem) {
                for (int q = tid; q < size; q += 512) s_r2[q] = __ldcg(r2 + q + base);
                __syncthreads();
            }
            const unsigned V = (unsigned)size / 2;  // ceil((size - 1) / 2)
            const unsigned long long E = (unsigned long long)V * (unsigned)size;
            const unsigned lo = (unsigned)(E * (unsigned)(blockIdx.x - 1) / (unsigned)n_scan);
            const unsigned hi = (unsigned)(E * (unsigned)blockIdx.x / (unsigned)n_scan);
            for (unsigned e0 = lo + tid; e0 < hi; e0 += 8 * 512) {
                float d[8], ri[8], rj[8];
                int ii[8], jj[8];
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const unsigned e = e0 + q * 512;
                    const unsigned v = e / (unsigned)size, k = e - v * (unsigned)size;
                    const int w = size - 2 - (int)v;
                    const bool first = (int)k < size - 1 - (int)v;
                    const int i = first ? (int)v : w, j = first ? (int)(v + 1 + k) : (int)k;
                    const bool ok = e < hi && (first || w != (int)v) && i != pend && j != pend;
                    ii[q] = ok ? i : -1;
                    jj[q] = j;
                    d[q] = ok ? __ldcg(m + (size_t)(i + base) * n0 + j + base) : 0.0f;
                    if (r2_in_smem) {
                        ri[q] = ok ? s_r2[i] : 0.0f;
                        rj[q] = ok ? s_r2[j] : 0.0f;
                    } else {
                        ri[q] = ok ? __ldcg(r2 + i + base) : 0.0f;
                        rj[q] = ok ? __ldcg(r2 + j + base) : 0.0f;
                    }
                }
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    if (ii[q] >= 0) {
                        const unsigned long long k = nj_key(d[q], ri[q], rj[q], ii[q], jj[q], n0);
                        best = k < best ? k : best;
                    }
                }
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
            if (best != ~0ull) atomicMin(&st->key[parity], best);
        }
        unsigned long long t1 = 0, t2 = 0, t3 = 0;
        if (timed) t1 = nj_now();
        if (!nj_barrier(st, n_ctas, m, r2, n0, base, parity)) return;
        if (timed) t2 = nj_now();
        // ---- update: the join (NJ_compute_r + NJ_collapse) and the relabelling, one thread per index ----
        const int a = __ldcg(&st->a), b = __ldcg(&st->b);
        float* row_a = m + (size_t)(a + base) * n0 + base;
        float* row_b = m + (size_t)(b + base) * n0 + base;
        const float* row_0 = m + (size_t)base * n0 + base;
        const float dab = __ldcg(&st->dab);
        const float r2a = __ldcg(&st->r2a), r2b = __ldcg(&st->r2b);
        const float a2 = (float)((double)__fadd_rn(dab, __fsub_rn(r2a, r2b)) / 2.0);
        const float b2 = (float)((double)__fadd_rn(dab, __fsub_rn(r2b, r2a)) / 2.0);
        const float denom = (float)(size - 3);
        const int gid = blockIdx.x * blockDim.x + tid, gstride = n_ctas * blockDim.x;
        for (int i = gid; i < size; i += gstride) {
            if (i == a) continue;
            // i == b: d(a, b) from the snapshot (thread 0 rewrites that cell below) and d(b, b) = 0, as the reference reads them
            const float da = i == b ? dab : __ldcg(row_a + i), db = i == b ? 0.0f : __ldcg(row_b + i);
            const float c = (float)((double)__fadd_rn(__fsub_rn(da, a2), __fsub_rn(db, b2)) / 2.0);
            cvec[i > a ? i - a - 1 : size - a - 1 + i] = c;  // the order the reference adds them in
            if (i == b) continue;  // index b is overwritten by the relabelling below
            row_a[i] = c;
            m[(size_t)(i + base) * n0 + a + base] = c;
            float ri = __fsub_rn(__fsub_rn(__ldcg(r + i + base), da), db);
            ri = __fadd_rn(ri, c);
            r[i + base] = ri;
            const float r2i = __fdiv_rn(ri, denom);
            r2[i + base] = r2i;
            // index 0 (after the update) takes over index b; index 0 itself is dropped by base + 1
            const float v = a == 0 ? c : __ldcg(row_0 + i);
            row_b[i] = v;
            m[(size_t)(i + base) * n0 + b + base] = v;
            if (i == 0) {  // a != 0 here: the updated row 0 also holds c_0 at column a
                row_b[a] = c;
                row_a[b] = c;
                r[b + base] = ri;
                r2[b + base] = r2i;
            }
        }
        if (gid == 0) {
            left[joined] = __ldcg(node + a + base);
            right[joined] = __ldcg(node + b + base);
            const int where = a == 0 ? b : a;  // where the joined node lives from now on
            if (a != 0) node[b + base] = __ldcg(node + base);
            node[where + base] = joined;
            st->pend_phys = where + base;
            st->pend_terms = size - 1;
            st->pend_denom = denom;
            st->key[parity ^ 1] = ~0ull;  // the next join's minimum; nobody touches it before the barrier below
        }
        if (timed) t3 = nj_now();
        if (!nj_barrier(st, n_ctas)) return;
        if (timed) {
            const unsigned long long t4 = nj_now();
            st->t_scan += t1 - t0, st->t_wait1 += t2 - t1, st->t_update += t3 - t2, st->t_wait2 += t4 - t3;
        }
    }
    if (blockIdx.x == 0 && tid == 0) {
        left[joined] = __ldcg(node + base);
        right[joined] = __ldcg(node + base + 1);
    }
}

__global__ void k_nj_finish(const int* __restrict__ node, int base, int32_t* __restrict__ left, int32_t* __restrict__ right, int joined) {
    left[joined] = node[base];
    right[joined] = node[base + 1];
}

}  // namespace

extern "C" int lzg_nj_seed_tree(lzg_ctx* ctx, const int32_t* D_host, int32_t n, const int32_t* perm, int32_t* left, int32_t* right) {
    if (!ctx || !D_host || !perm || !left || !right || n < 1 || n > 65535) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_nj_seed_tree: bad argument");
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
        const size_t nn = (size_t)n * n;
        for (size_t q = 0; q < nn; q++)
            if (D_host[q] < 0 || D_host[q] >= (1 << 24)) return lzg_fail(ctx, LZG_ERR_RANGE, "distance does not fit float32 exactly");
    }
    int32_t *d_D = nullptr, *d_perm = nullptr, *d_left = nullptr, *d_right = nullptr;
    int* d_node = nullptr;
    float *d_m = nullptr, *d_r = nullptr, *d_r2 = nullptr;
    unsigned long long* d_key = nullptr;
    float* d_cvec = nullptr;
    lzg_nj_state* d_state = nullptr;
    const size_t nn = (size_t)n * n;
    auto release = [&]() {
        lzg_dev_free(ctx, d_D);
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
    };
    cudaError_t e = lzg_dev_alloc(ctx, (void**)&d_D, sizeof(int32_t) * nn);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_m, sizeof(float) * nn);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_perm, sizeof(int32_t) * n);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_left, sizeof(int32_t) * total);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_right, sizeof(int32_t) * total);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_node, sizeof(int) * n);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_r, sizeof(float) * n);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_r2, sizeof(float) * n);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&d_key, sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_D, D_host, sizeof(int32_t) * nn, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_perm, perm, sizeof(int32_t) * n, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_key, 0xff, sizeof(unsigned long long), ctx->stream);
    if (e != cudaSuccess) {
        release();
        return lzg_fail(ctx, LZG_ERR_CUDA, "lzg_nj_seed_tree: setup", e);
    }
    k_nj_gather<<<dim3(std::max(1, std::min(32, (n + 255) / 256)), n), 256, 0, ctx->stream>>>(d_D, d_perm, d_m, n);
    ctx->launches++;
    k_nj_init_r<<<(n + 127) / 128, 128, 0, ctx->stream>>>(d_m, n, d_r, d_r2, d_node);
    ctx->launches++;
    // one persistent cooperative launch for all joins; LZG_NJ_PERSISTENT=0 keeps the two-launches-per-join form (the cross-check)
    const bool persistent = !(getenv("LZG_NJ_PERSISTENT") && atoi(getenv("LZG_NJ_PERSISTENT")) == 0);
    bool done = false;
    if (persistent && n >= 4) {
        int per_sm = 0;
        const size_t smem_r2 = n <= 12000 ? sizeof(float) * (size_t)n : 0;  // r2 staged in shared memory under the 48 KB default
        cudaError_t ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_nj_persistent, 512, smem_r2);
        if (ce == cudaSuccess && per_sm >= 1) ce = lzg_dev_alloc(ctx, (void**)&d_cvec, sizeof(float) * (size_t)n);
        if (ce == cudaSuccess && per_sm >= 1) ce = lzg_dev_alloc(ctx, (void**)&d_state, sizeof(lzg_nj_state));
        if (ce == cudaSuccess && per_sm >= 1) {
            lzg_nj_state init;
            memset(&init, 0, sizeof(init));
            init.key[0] = init.key[1] = ~0ull;
            init.pend_phys = -1;
            ce = cudaMemcpyAsync(d_state, &init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream);
            int grid = std::min(ctx->sm_count * std::min(per_sm, 2), 1 + std::max(1, n / 2));
            grid = std::max(grid, 2);  // CTA 0 owns the joined node, the others scan
            int n_arg = n, in_smem = smem_r2 > 0;
            void* args[] = {(void*)&d_m, (void*)&n_arg, (void*)&d_r, (void*)&d_r2, (void*)&d_node, (void*)&d_cvec, (void*)&d_left, (void*)&d_right, (void*)&d_state, (void*)&in_smem};
            if (ce == cudaSuccess) ce = cudaLaunchCooperativeKernel((void*)k_nj_persistent, dim3(grid), dim3(512), args, smem_r2, ctx->stream);
            if (ce == cudaSuccess) {
                ctx->launches++;
                lzg_nj_state back;
                ce = cudaMemcpyAsync(&back, d_state, sizeof(back), cudaMemcpyDeviceToHost, ctx->stream);
                if (ce == cudaSuccess) ce = cudaStreamSynchronize(ctx->stream);
                if (ce == cudaSuccess && getenv("LZG_NJ_DEBUG"))
                    fprintf(stderr, "[lzg nj] n %d, %d CTAs: scan %.1f ms, wait behind the scan %.1f ms, update %.1f ms, wait behind the update %.1f ms\n", n, grid,
                            back.t_scan * 1e-6, back.t_wait1 * 1e-6, back.t_update * 1e-6, back.t_wait2 * 1e-6);
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
            k_nj_argmin<<<grid, 256, 0, ctx->stream>>>(d_m, n, base, size, d_r2, d_key);
            k_nj_join<<<1, 1024, 0, ctx->stream>>>(d_m, n, base, size, d_r, d_r2, d_node, d_key, d_left, d_right, joined);
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
