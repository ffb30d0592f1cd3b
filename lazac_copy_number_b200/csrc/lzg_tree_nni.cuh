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
// with its running interval in registers.  The new children of v are (w, y) and the other child
// of u is z; everything above u is read from the (parent, sibling) table, whose entries for u and
// its ancestors the move does not touch, so thread 0 of CTA 0 can patch the topology arrays while
// the others walk.  Step k's cost goes to path_cost[k]; the CTA that finishes last turns the
// costs into cumulative scores (src/copy_number.cxx:264) and leaves the root score in ctl.
__global__ void __launch_bounds__(256)
k_apply_move(int32_t* child0, int32_t* child1, int2* up, uint8_t* __restrict__ start, uint8_t* __restrict__ end,
             int32_t* __restrict__ score, int32_t* __restrict__ path_cost, int* __restrict__ ctl, int u, int w, int v, int z,
             int y, int l_pad) {
    const int chunk = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = chunk * 16 < l_pad;
    const size_t off = (size_t)chunk * 16;
    iv16 acc;
    uint32_t os[8], oe[8];
    if (active) {
        widen(ldg16(start, (size_t)w * l_pad + off), acc.s);
        widen(ldg16(end, (size_t)w * l_pad + off), acc.e);
        widen(ldg16(start, (size_t)y * l_pad + off), os);
        widen(ldg16(end, (size_t)y * l_pad + off), oe);
    }
    int a = v, step = 0;
    int2 ps = make_int2(u, z);  // (parent, sibling) of v after the move
    while (true) {
        uint32_t c = 0;
        if (active) {
            c = merge16(acc, os, oe);
            *reinterpret_cast<uint4*>(start + (size_t)a * l_pad + off) = narrow(acc.s);
            *reinterpret_cast<uint4*>(end + (size_t)a * l_pad + off) = narrow(acc.e);
        }
        c = warp_sum(c);
        if ((threadIdx.x & 31) == 0 && c) atomicAdd(path_cost + step, (int32_t)c);
        if (a == 0) break;
        if (active) {
            widen(ldg16(start, (size_t)ps.y * l_pad + off), os);
            widen(ldg16(end, (size_t)ps.y * l_pad + off), oe);
        }
        a = ps.x;
        if (a != 0) ps = up[a];
        step++;
    }
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(ctl + LZG_CTL_DONE, 1) == (int)gridDim.x - 1;
    __syncthreads();
    if (!s_last || threadIdx.x != 0) return;
    __threadfence();
    // cumulative scores along the path, then the topology patch (nni(), src/copy_number.cxx:279-284)
    int32_t s = __ldcg(path_cost + 0) + score[w] + score[y];
    score[v] = s;
    s += __ldcg(path_cost + 1) + score[z];
    score[u] = s;
    step = 2;
    for (int b = u; b != 0; step++) {
        const int2 q = up[b];
        s += __ldcg(path_cost + step) + score[q.y];
        score[q.x] = s;
        b = q.x;
    }
    ctl[LZG_CTL_SCORE] = s;
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
            widen(s4, s);
            widen(e4, e);
            widen(p4, pl);
#pragma unroll
            for (int k = 0; k < 8; k++) {
                uint32_t x = __vminu2(__vmaxu2(pl[k], s[k]), e[k]);
                out[k] = x;
                c += __vmaxu2(x, pl[k]) - __vminu2(x, pl[k]);
            }
            c = (c & 0xffffu) + (c >> 16);
            *reinterpret_cast<uint4*>(label + ou) = narrow(out);
        }
    }
    c = warp_sum(c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(branch + u, (int32_t)c);
}

// biased bytes [rows][l_pad] -> int32 [rows][L]
__global__ void k_unpack_i32(const uint8_t* __restrict__ src, int32_t* __restrict__ dst, int L, int l_pad) {
    int r = blockIdx.y;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < L; k += gridDim.x * blockDim.x)
        dst[(size_t)r * L + k] = (int32_t)src[(size_t)r * l_pad + k] - LZG_BIAS;
}

}  // namespace
