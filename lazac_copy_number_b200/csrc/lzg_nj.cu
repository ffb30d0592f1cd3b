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
#include <vector>

#include "lzg_common.cuh"

namespace {

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
    int base = 0, joined = n;
    for (int size = n; size > 2; size--, base++, joined++) {
        const int grid = std::max(1, std::min(size - 1, ctx->sm_count * 8));
        k_nj_argmin<<<grid, 256, 0, ctx->stream>>>(d_m, n, base, size, d_r2, d_key);
        k_nj_join<<<1, 1024, 0, ctx->stream>>>(d_m, n, base, size, d_r, d_r2, d_node, d_key, d_left, d_right, joined);
        ctx->launches += 2;
    }
    k_nj_finish<<<1, 1, 0, ctx->stream>>>(d_node, base, d_left, d_right, joined);
    ctx->launches++;
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(left + n, d_left + n, sizeof(int32_t) * (total - n), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(right + n, d_right + n, sizeof(int32_t) * (total - n), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    release();
    if (e != cudaSuccess) return lzg_fail(ctx, LZG_ERR_CUDA, "lzg_nj_seed_tree", e);
    return LZG_OK;
}
