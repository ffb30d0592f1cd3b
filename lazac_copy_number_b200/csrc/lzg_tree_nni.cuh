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
// Every term is a sum over positions, so chunk groups just add their parts: 10 L bytes per
// candidate instead of (depth(v)+2) 2 L, nothing written but 4 bytes per CTA.
//
// Work items are drawn from an atomic ticket counter in this order:
//   1. (heavy path, chunk group): a CTA walks one heavy path of internal nodes top-down with
//      the running UP interval in registers, storing UP(a) and releasing a per-(node, group)
//      epoch flag at every node; its head waits for the flag of its parent (on an earlier path).
//      Cross-CTA hops are light edges only (<= log2 n on any root-leaf path), however deep the
//      tree.  Skipped when the UP intervals are still valid.
//   2. (candidate, chunk group): waits for the flag of (u, group), scores, adds its part.
//   3. the CTA that finishes last reduces min (score, index) over the candidates below the
//      current root score (greedy: min index) -- the reference's strict '<', first-wins rule
//      (src/copy_number.cxx:353-362).
// A ticket is only drawn by a CTA that is already running and only waits for smaller tickets, so
// the spins cannot deadlock whatever the number of resident CTAs.  Cross-CTA data goes through
// L2: stores, bar.sync, st.release by one thread / ld.acquire by one thread, bar.sync, ld.cg.
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

struct lzg_pass_args {
    const int32_t* path_off;  // [n_paths + 1] into order
    const int32_t* order;     // internal non-root nodes, heavy paths back to back, parents-first
    const int32_t* moves;     // [n_cand][4]
    int n_paths, n_cand, groups, l_pad, epoch, greedy;
    const int32_t *child0, *child1;
    const int2* up;
    const uint8_t *start, *end;
    uint8_t *ustart, *uend;
    const int32_t* score;
    int* ready;          // [n_nodes][groups]
    int* ctl;            // [0] ticket, [1] finished candidate CTAs, [2] best index, [3] best score
    int32_t* cand_score;  // [n_cand], zeroed
};

enum { LZG_CTL_TICKET = 0, LZG_CTL_DONE = 1, LZG_CTL_BEST = 2, LZG_CTL_SCORE = 3, LZG_CTL_LAST = 4, LZG_CTL_INTS = 16 };

__global__ void __launch_bounds__(256)
k_nni_pass(const lzg_pass_args A) {
    __shared__ int s_item;
    __shared__ int32_t s_part[8];
    __shared__ unsigned long long s_key[8];
    const int tid = threadIdx.x;
    if (tid == 0) s_item = atomicAdd(A.ctl + LZG_CTL_TICKET, 1);
    __syncthreads();
    const int item = s_item;
    const int g = item % A.groups;
    const int chunk = g * blockDim.x + tid;
    const bool active = chunk * 16 < A.l_pad;
    const size_t off = (size_t)chunk * 16;
    const int l_pad = A.l_pad;
    uint32_t os[8], oe[8];

    if (item < A.n_paths * A.groups) {
        // ---- heavy-path walker -----------------------------------------------------------------
        const int p = item / A.groups;
        const int k0 = A.path_off[p], k1 = A.path_off[p + 1];
        int a = A.order[k0];
        int2 ps = A.up[a];
        uint4 ss = make_uint4(0, 0, 0, 0), se = ss;
        if (active) {
            ss = ldg16(A.start, (size_t)ps.y * l_pad + off);
            se = ldg16(A.end, (size_t)ps.y * l_pad + off);
        }
        iv16 acc;
        if (ps.x != 0) {
            if (tid == 0) {
                const int* flag = A.ready + (size_t)ps.x * A.groups + g;
                while (ld_acquire(flag) != A.epoch) {
                }
            }
            __syncthreads();
            if (active) {
                widen(ldcg16(A.ustart, (size_t)ps.x * l_pad + off), acc.s);
                widen(ldcg16(A.uend, (size_t)ps.x * l_pad + off), acc.e);
                widen(ss, os);
                widen(se, oe);
                merge16(acc, os, oe);
            }
        } else if (active) {
            widen(ss, acc.s);
            widen(se, acc.e);
        }
        for (int k = k0;; k++) {
            // prefetch the next node's sibling interval while this node's result is stored
            int a_next = 0;
            uint4 ns = make_uint4(0, 0, 0, 0), ne = ns;
            if (k + 1 < k1) {
                a_next = A.order[k + 1];
                const int sib = A.up[a_next].y;
                if (active) {
                    ns = ldg16(A.start, (size_t)sib * l_pad + off);
                    ne = ldg16(A.end, (size_t)sib * l_pad + off);
                }
            }
            if (active) {
                *reinterpret_cast<uint4*>(A.ustart + (size_t)a * l_pad + off) = narrow(acc.s);
                *reinterpret_cast<uint4*>(A.uend + (size_t)a * l_pad + off) = narrow(acc.e);
            }
            __syncthreads();
            if (tid == 0) st_release(A.ready + (size_t)a * A.groups + g, A.epoch);
            if (k + 1 >= k1) break;
            if (active) {
                widen(ns, os);
                widen(ne, oe);
                merge16(acc, os, oe);
            }
            a = a_next;
        }
        return;
    }

    // ---- candidate -------------------------------------------------------------------------------
    const int cand = (item - A.n_paths * A.groups) / A.groups;
    const int4 mv = *reinterpret_cast<const int4*>(A.moves + 4 * cand);
    const int u = mv.x, w = mv.y, v = mv.z, z = mv.w;
    const int vc0 = A.child0[v], vc1 = A.child1[v];
    const int y = vc0 == z ? vc1 : vc0;
    int32_t part = 0;
    uint4 sw, ew, sy, ey, sz, ez, sd, ed;
    if (active) {
        sw = ldg16(A.start, (size_t)w * l_pad + off), ew = ldg16(A.end, (size_t)w * l_pad + off);
        sy = ldg16(A.start, (size_t)y * l_pad + off), ey = ldg16(A.end, (size_t)y * l_pad + off);
        sz = ldg16(A.start, (size_t)z * l_pad + off), ez = ldg16(A.end, (size_t)z * l_pad + off);
        if (u != 0) sd = ldg16(A.start, (size_t)u * l_pad + off), ed = ldg16(A.end, (size_t)u * l_pad + off);
    }
    if (u != 0) {
        if (tid == 0) {
            const int* flag = A.ready + (size_t)u * A.groups + g;
            while (ld_acquire(flag) != A.epoch) {
            }
        }
        __syncthreads();
    }
    if (active) {
        iv16 acc;
        widen(sw, acc.s);
        widen(ew, acc.e);
        widen(sy, os);
        widen(ey, oe);
        uint32_t c = merge16(acc, os, oe);
        widen(sz, os);
        widen(ez, oe);
        c += merge16(acc, os, oe);
        if (u != 0) {
            const uint4 us = ldcg16(A.ustart, (size_t)u * l_pad + off), ue = ldcg16(A.uend, (size_t)u * l_pad + off);
            widen(us, os);
            widen(ue, oe);
            c += merge16(acc, os, oe);
            widen(sd, acc.s);
            widen(ed, acc.e);
            part = -(int32_t)merge16(acc, os, oe);
        }
        part += (int32_t)c;
    }
    part = (int32_t)warp_sum((uint32_t)part);  // two's complement: the wrapped sum is the signed sum
    if ((tid & 31) == 0) s_part[tid >> 5] = part;
    __syncthreads();
    __shared__ int s_last;
    if (tid == 0) {
        int32_t tot = 0;
        for (int k = 0; k < (blockDim.x + 31) / 32; k++) tot += s_part[k];
        if (g == 0) tot += A.score[w] + A.score[y] + A.score[z] + (u != 0 ? A.score[0] - A.score[u] : 0);
        if (tot) atomicAdd(A.cand_score + cand, tot);
        __threadfence();
        s_last = atomicAdd(A.ctl + LZG_CTL_DONE, 1) == A.n_cand * A.groups - 1;
    }
    __syncthreads();
    if (!s_last) return;

    // ---- the last CTA: argmin -----------------------------------------------------------------------
    __threadfence();
    const int32_t base = A.score[0];
    unsigned long long best = ~0ull;
    for (int i = tid; i < A.n_cand; i += blockDim.x) {
        const int32_t sc = __ldcg(A.cand_score + i);
        if (sc < base) {
            unsigned long long key = A.greedy ? (unsigned long long)i : (((unsigned long long)(uint32_t)sc) << 32) | (uint32_t)i;
            best = key < best ? key : best;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
        best = other < best ? other : best;
    }
    if ((tid & 31) == 0) s_key[tid >> 5] = best;
    __syncthreads();
    if (tid == 0) {
        for (int k = 1; k < (blockDim.x + 31) / 32; k++) best = s_key[k] < best ? s_key[k] : best;
        A.ctl[LZG_CTL_LAST] = __ldcg(A.cand_score + A.n_cand - 1);  // see lzg_tree_last_candidate_score
        if (best == ~0ull) {
            A.ctl[LZG_CTL_BEST] = -1;
            A.ctl[LZG_CTL_SCORE] = base;
        } else {
            const int idx = (int)(best & 0xffffffffu);
            A.ctl[LZG_CTL_BEST] = idx;
            A.ctl[LZG_CTL_SCORE] = __ldcg(A.cand_score + idx);
        }
    }
}

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
