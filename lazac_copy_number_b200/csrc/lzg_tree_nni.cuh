// lzg_tree_nni.cuh -- NNI neighbourhood scoring, move application, top-down labeling.
// Included by lzg_tree.cu only (uses its device helpers).
#pragma once

namespace {

// ---- one neighbourhood pass = ONE launch ---------------------------------------------------------------
// The reference scores a candidate (u,w,v,z) by nni -> unvisit(v) -> small_rectilinear -> read
// t[0].score -> undo (/root/reference/src/copy_number.cxx:349-366): depth(v)+1 merges along the
// path v .. root.  That root score is the optimum of the small-parsimony problem on the
// candidate topology, and the interval DP is exact for it under ANY rooting (per position every
// partial cost function has integer slopes, so min-plus convolution with |x| collapses to
// "minimum + distance to the arg-min interval").  We therefore root the candidate at the edge
// above u and need three merges only.  With UP(a) = interval of the rest of the tree seen from
// the edge above a (UP(a) = DOWN(sibling) when the parent is the root, else
// merge(UP(parent), DOWN(sibling))) and y = the child v keeps:
//     I1 = merge(w, y)          (the new v)   cost c1
//     I2 = merge(I1, z)         (the new u)   cost c2
//     c3 = cost(merge(I2, UP(u)))             (not when u is the root)
//     c0 = cost(merge(DOWN(u), UP(u)))        (the current tree re-rooted above u:
//                                              score[0] = c0 + score[u] + score_of_the_rest)
//     candidate = c1 + c2 + c3 - c0 + score[w] + score[y] + score[z] + score[0] - score[u]
// Every term is a sum over positions, so chunk groups just add their parts.  The two candidates
// of an edge (u, v) -- z = either child of v, enumerated back to back by greedy_nni,
// src/copy_number.cxx:340-346 -- read the same five intervals (w, both children of v, DOWN(u),
// UP(u)), so they are scored together as ONE work item: 10 L bytes per PAIR (6 L when u is the
// root) instead of (depth(v)+2) 2 L per candidate, nothing written but 8 bytes per CTA.
//
// The kernel itself (persistent CTAs, a producer warp feeding a shared-memory ring with bulk
// async copies, decoupled consumer warps; UP intervals built by heavy-path walkers in the same
// launch) is in lzg_tree_pass.cuh.
__device__ __forceinline__ int ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 ldcg16(const uint8_t* base, size_t off) {
    return __ldcg(reinterpret_cast<const uint4*>(base + off));
}

#include "lzg_tree_pass.cuh"

// ---- accept a move = ONE launch ------------------------------------------------------------------------
// nni(t,u,w,v,z); unvisit(t,0,v); small_rectilinear(t,0) (src/copy_number.cxx:397,408-409): rewrite
// the intervals of v, u, parent(u), ..., root.  Every thread owns 16 positions and walks the path
// with its running interval in registers.  The host lists the path in the kernel arguments
// (step k writes node[k] after merging other[k]: v <- (w, y), u <- (v, z), then the old
// ancestors of u with their siblings), so the sibling rows are prefetched two steps ahead and
// the final score update loads everything at once; a path longer than LZG_APPLY_LISTED continues
// through the (parent, sibling) table, whose entries above u the move does not touch.  Step
// k's cost goes to path_cost[k]; the CTA that finishes last turns the costs into cumulative
// scores (src/copy_number.cxx:264), leaves the root score in ctl and patches the topology arrays.
#define LZG_APPLY_LISTED 64
struct lzg_apply_args {
    int n_steps, u, w, v, z, y;
    int node[LZG_APPLY_LISTED], other[LZG_APPLY_LISTED];
};

template <int W>
__global__ void __launch_bounds__(256)
k_apply_move(int32_t* child0, int32_t* child1, int2* up, uint8_t* __restrict__ start, uint8_t* __restrict__ end,
             int32_t* __restrict__ score, int32_t* __restrict__ path_cost, int* __restrict__ ctl, const lzg_apply_args P, int l_pad) {
    const int chunk = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = chunk * 16 < l_pad;
    const size_t off = (size_t)chunk * 16;
    const int n_listed = min(P.n_steps, LZG_APPLY_LISTED);
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    iv16 acc;
    uint32_t os[8], oe[8];
    uint4 s0 = z4, e0 = z4, s1 = z4, e1 = z4, s2 = z4, e2 = z4;
    if (active) {
        widen_w<W>(ldg16(start, (size_t)P.w * l_pad + off), acc.s);
        widen_w<W>(ldg16(end, (size_t)P.w * l_pad + off), acc.e);
        s0 = ldg16(start, (size_t)P.other[0] * l_pad + off), e0 = ldg16(end, (size_t)P.other[0] * l_pad + off);
        if (n_listed > 1) s1 = ldg16(start, (size_t)P.other[1] * l_pad + off), e1 = ldg16(end, (size_t)P.other[1] * l_pad + off);
    }
    int a = 0;
    for (int k = 0; k < P.n_steps; k++) {
        int oth_next2 = -1;
        if (k + 2 < n_listed) oth_next2 = P.other[k + 2];
        if (active && oth_next2 >= 0) s2 = ldg16(start, (size_t)oth_next2 * l_pad + off), e2 = ldg16(end, (size_t)oth_next2 * l_pad + off);
        if (k < n_listed) {
            a = P.node[k];
        } else {  // beyond the listed part: follow the table (a is the node written last)
            const int2 ps = up[a];
            a = ps.x;
            if (active) s0 = ldg16(start, (size_t)ps.y * l_pad + off), e0 = ldg16(end, (size_t)ps.y * l_pad + off);
        }
        uint32_t c = 0;
        if (active) {
            widen_w<W>(s0, os);
            widen_w<W>(e0, oe);
            c = merge_w<W>(acc, os, oe);
            *reinterpret_cast<uint4*>(start + (size_t)a * l_pad + off) = narrow_w<W>(acc.s);
            *reinterpret_cast<uint4*>(end + (size_t)a * l_pad + off) = narrow_w<W>(acc.e);
        }
        c = warp_sum(c);
        if ((threadIdx.x & 31) == 0 && c) atomicAdd(path_cost + k, (int32_t)c);
        s0 = s1, e0 = e1;
        s1 = s2, e1 = e2;
    }
    __shared__ int s_last;
    __shared__ int32_t s_term[LZG_APPLY_LISTED];
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(ctl + LZG_CTL_DONE, 1) == (int)gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // cumulative scores along the path: score[node[k]] = cost[k] + score[other[k]] + score[node[k-1]] (w for k = 0)
    for (int k = threadIdx.x; k < n_listed; k += blockDim.x) s_term[k] = __ldcg(path_cost + k) + score[P.other[k]];
    __syncthreads();
    if (threadIdx.x != 0) return;
    int32_t s = score[P.w];
    for (int k = 0; k < n_listed; k++) {
        s += s_term[k];
        score[P.node[k]] = s;
    }
    for (int k = n_listed, b = P.node[LZG_APPLY_LISTED - 1]; k < P.n_steps; k++) {
        const int2 q = up[b];
        s += __ldcg(path_cost + k) + score[q.y];
        score[q.x] = s;
        b = q.x;
    }
    ctl[LZG_CTL_SCORE] = s;
    // the topology patch (nni(), src/copy_number.cxx:279-284)
    const int u = P.u, w = P.w, v = P.v, z = P.z, y = P.y;
    child0[v] = min(w, y);
    child1[v] = max(w, y);
    child0[u] = min(v, z);
    child1[u] = max(v, z);
    up[w] = make_int2(v, y);
    up[y] = make_int2(v, w);
    up[v] = make_int2(u, z);
    up[z] = make_int2(u, v);
}

// ---- top-down labeling (local_labeling / ancestral_labeling, src/copy_number.cxx:165-225) ---------
// label[root] = start[root]; label[c] = clamp(label[parent], start[c], end[c]);
// branch[c] = sum |label[c] - label[parent]|.  One launch per depth level, root first.
template <int W>
__global__ void k_label_level(const int32_t* __restrict__ order, const int2* __restrict__ up, const uint8_t* __restrict__ start,
                              const uint8_t* __restrict__ end, uint8_t* __restrict__ label, int32_t* __restrict__ branch, int l_pad) {
    int u = order[blockIdx.y];
    int p = up[u].x;
    int chunk = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t c = 0;
    if (chunk * 16 < l_pad) {
        size_t ou = (size_t)u * l_pad + chunk * 16;
        uint4 s4 = ldg16(start, ou);
        if (p < 0) {
            *reinterpret_cast<uint4*>(label + ou) = s4;
        } else {
            uint4 e4 = ldg16(end, ou);
            uint4 p4 = *reinterpret_cast<const uint4*>(label + (size_t)p * l_pad + chunk * 16);
            uint32_t s[8], e[8], pl[8], out[8];
            widen_w<W>(s4, s);
            widen_w<W>(e4, e);
            widen_w<W>(p4, pl);
#pragma unroll
            for (int k = 0; k < (W == 1 ? 8 : 4); k++) {
                uint32_t x = __vminu2(__vmaxu2(pl[k], s[k]), e[k]);
                out[k] = x;
                const uint32_t d = __vmaxu2(x, pl[k]) - __vminu2(x, pl[k]);
                c += W == 1 ? d : (d & 0xffffu) + (d >> 16);  // 16-bit lanes: fold every word (differences up to 65535)
            }
            if (W == 1) c = (c & 0xffffu) + (c >> 16);
            *reinterpret_cast<uint4*>(label + ou) = narrow_w<W>(out);
        }
    }
    c = warp_sum(c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(branch + u, (int32_t)c);
}

#include "lzg_tree_climb.cuh"

// biased bytes / 16-bit words [rows][l_pad bytes] -> int32 [rows][L]
template <int W>
__global__ void k_unpack_i32(const uint8_t* __restrict__ src, int32_t* __restrict__ dst, int L, int l_pad) {
    int r = blockIdx.y;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < L; k += gridDim.x * blockDim.x) {
        if (W == 1) dst[(size_t)r * L + k] = (int32_t)src[(size_t)r * l_pad + k] - LZG_BIAS;
        else dst[(size_t)r * L + k] = (int32_t)reinterpret_cast<const uint16_t*>(src + (size_t)r * l_pad)[k] - LZG_BIAS16;
    }
}

}  // namespace
