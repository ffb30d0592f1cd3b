// lzg_tree_nni.cuh -- NNI neighbourhood scoring, move application, top-down labeling.
// Included by lzg_tree.cu only (uses its device helpers).
#pragma once

namespace {

// ---- neighbourhood scorer ----------------------------------------------------------------------
// The reference scores a candidate (u,w,v,z) by nni -> unvisit(v) -> small_rectilinear -> read
// t[0].score -> undo (/root/reference/src/copy_number.cxx:349-366): depth(v)+1 merges along the
// path v .. root.  That root score is the optimum of the small-parsimony problem on the
// candidate topology, and the interval DP is exact for it under ANY rooting (per position every
// partial cost function has integer slopes, so min-plus convolution with |x| collapses to
// "minimum + distance to the arg-min interval").  We therefore root the candidate at the edge
// above u and need three merges only:
//     I1 = merge(w, y)   (the new v; y = the child v keeps)      cost c1
//     I2 = merge(I1, z)  (the new u)                              cost c2
//     c3 = cost(merge(I2, UP(u)))                                 (skipped when u is the root)
//     score = c1 + c2 + c3 + score[w] + score[y] + score[z] + upscore[u]
// where UP(u) / upscore[u] are the interval and cumulative score of the rest of the tree seen
// from the edge above u (k_up_level below).  8 L bytes per candidate instead of
// (depth(v)+2) 2 L, nothing written but 4 bytes per CTA.  tests/ check every candidate's score
// against the reference's path algorithm.
__global__ void __launch_bounds__(256)
k_nni_scores(const int32_t* __restrict__ moves, const int32_t* __restrict__ child0, const int32_t* __restrict__ child1,
             const uint8_t* __restrict__ start, const uint8_t* __restrict__ end, const uint8_t* __restrict__ ustart,
             const uint8_t* __restrict__ uend, const int32_t* __restrict__ score, const int32_t* __restrict__ upscore,
             int32_t* __restrict__ cand_score, int l_pad) {
    const int cand = blockIdx.x;
    const int4 mv = *reinterpret_cast<const int4*>(moves + 4 * cand);
    const int u = mv.x, w = mv.y, v = mv.z, z = mv.w;
    const int vc0 = child0[v], vc1 = child1[v];
    const int y = vc0 == z ? vc1 : vc0;
    const int chunk = blockIdx.y * blockDim.x + threadIdx.x;
    const size_t off = (size_t)chunk * 16;
    uint32_t cost = 0;
    if (chunk * 16 < l_pad) {
        uint4 sw = ldg16(start, (size_t)w * l_pad + off), ew = ldg16(end, (size_t)w * l_pad + off);
        uint4 sy = ldg16(start, (size_t)y * l_pad + off), ey = ldg16(end, (size_t)y * l_pad + off);
        uint4 sz = ldg16(start, (size_t)z * l_pad + off), ez = ldg16(end, (size_t)z * l_pad + off);
        uint4 su = sz, eu = ez;
        if (u != 0) {
            su = ldg16(ustart, (size_t)u * l_pad + off);
            eu = ldg16(uend, (size_t)u * l_pad + off);
        }
        iv16 acc;
        uint32_t os[8], oe[8];
        widen(sw, acc.s);
        widen(ew, acc.e);
        widen(sy, os);
        widen(ey, oe);
        cost += merge16(acc, os, oe);
        widen(sz, os);
        widen(ez, oe);
        cost += merge16(acc, os, oe);
        if (u != 0) {
            widen(su, os);
            widen(eu, oe);
            cost += merge16(acc, os, oe);
        }
    }
    cost = warp_sum(cost);
    __shared__ uint32_t wsum[8];
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = cost;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tot = 0;
        for (int k = 0; k < (blockDim.x + 31) / 32; k++) tot += wsum[k];
        int32_t add = (int32_t)tot;
        if (blockIdx.y == 0) add += score[w] + score[y] + score[z] + (u != 0 ? upscore[u] : 0);
        if (add) atomicAdd(cand_score + cand, add);
    }
}

// ---- UP intervals: ONE dependency-driven launch ----------------------------------------------------
// For a non-root node a with parent p and sibling s:
//     p is the root:  UP(a) = DOWN(s),                 upscore[a] = score[s]
//     otherwise:      UP(a) = merge(UP(p), DOWN(s)),   upscore[a] = cost + upscore[p] + score[s]
// (the root has two children and no data of its own, so seen from one child the rest of the tree
// is just the other child).  Only internal nodes are listed: a leaf is never the u of a move.
//
// A level-by-level schedule costs one launch per tree level (27 at cfg4, hundreds on a deep NJ
// tree), each a few microseconds of mostly idle GPU.  Instead the work items (node, chunk group)
// are listed parents-first (BFS order from the host mirror) and CTAs draw them from an atomic
// ticket counter; an item waits on a per-(parent, group) epoch flag.  A ticket is only drawn by
// a CTA that is already running and only waits for smaller tickets, so the spin cannot deadlock
// whatever the number of resident CTAs.  Cross-CTA data goes through L2 (st + __threadfence +
// release flag on one side, acquire flag + ld.cg on the other).
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

__global__ void __launch_bounds__(256)
k_up_chain(const int32_t* __restrict__ order, int groups, const int2* __restrict__ up, const uint8_t* __restrict__ start,
           const uint8_t* __restrict__ end, uint8_t* __restrict__ ustart, uint8_t* __restrict__ uend,
           int32_t* __restrict__ upcost, int* __restrict__ ready, int epoch, int* __restrict__ ticket, int l_pad) {
    __shared__ int s_item;
    if (threadIdx.x == 0) s_item = atomicAdd(ticket, 1);
    __syncthreads();
    const int item = s_item;
    const int a = order[item / groups], g = item % groups;
    const int2 ps = up[a];
    const int chunk = g * blockDim.x + threadIdx.x;
    const bool active = chunk * 16 < l_pad;
    const size_t off = (size_t)chunk * 16;
    uint4 ss = make_uint4(0, 0, 0, 0), se = ss;
    if (active) {  // the sibling's DOWN interval does not depend on the wait
        ss = ldg16(start, (size_t)ps.y * l_pad + off);
        se = ldg16(end, (size_t)ps.y * l_pad + off);
    }
    uint32_t c = 0;
    if (ps.x != 0) {
        if (threadIdx.x == 0) {
            const int* flag = ready + (size_t)ps.x * groups + g;
            while (ld_acquire(flag) != epoch) __nanosleep(20);
        }
        __syncthreads();
        if (active) {
            uint4 us = ldcg16(ustart, (size_t)ps.x * l_pad + off), ue = ldcg16(uend, (size_t)ps.x * l_pad + off);
            iv16 acc;
            uint32_t os[8], oe[8];
            widen(us, acc.s);
            widen(ue, acc.e);
            widen(ss, os);
            widen(se, oe);
            c = merge16(acc, os, oe);
            ss = narrow(acc.s);
            se = narrow(acc.e);
        }
    }
    if (active) {
        *reinterpret_cast<uint4*>(ustart + (size_t)a * l_pad + off) = ss;
        *reinterpret_cast<uint4*>(uend + (size_t)a * l_pad + off) = se;
    }
    c = warp_sum(c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(upcost + a, (int32_t)c);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) st_release(ready + (size_t)a * groups + g, epoch);
}

// upscore[a] = sum over b = a, parent(a), ... (below the root's children) of upcost[b] + score[sibling(b)]:
// one thread per listed node walks up; no level structure needed.
__global__ void k_up_scores(const int32_t* __restrict__ order, int n_items, const int2* __restrict__ up,
                            const int32_t* __restrict__ upcost, const int32_t* __restrict__ score,
                            int32_t* __restrict__ upscore) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_items) return;
    const int a = order[i];
    int32_t tot = 0;
    for (int b = a; b != 0;) {
        int2 ps = up[b];
        tot += score[ps.y] + (ps.x != 0 ? upcost[b] : 0);
        b = ps.x;
    }
    upscore[a] = tot;
}

// Strict '<' with first-wins ties (src/copy_number.cxx:353-357): minimise (score, index) over the
// candidates below the current root score; greedy: minimise the index alone (:358-362).
// result[0] = index (-1 if none), result[1] = its score.
__global__ void k_nni_argmin(const int32_t* __restrict__ cand_score, int n, const int32_t* __restrict__ score,
                             int greedy, int32_t* __restrict__ result) {
    const int32_t base = score[0];
    unsigned long long best = ~0ull;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        int32_t s = cand_score[i];
        if (s < base) {
            unsigned long long key = greedy ? (unsigned long long)i
                                            : (((unsigned long long)(uint32_t)s) << 32) | (uint32_t)i;
            best = key < best ? key : best;
        }
    }
    __shared__ unsigned long long sm[32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
        best = other < best ? other : best;
    }
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < (blockDim.x + 31) / 32; k++) best = sm[k] < best ? sm[k] : best;
        if (best == ~0ull) {
            result[0] = -1;
            result[1] = base;
        } else {
            int idx = (int)(best & 0xffffffffu);
            result[0] = idx;
            result[1] = cand_score[idx];
        }
    }
}

// ---- apply a move --------------------------------------------------------------------------------
// nni(t,u,w,v,z) on the device topology (src/copy_number.cxx:279-284).
__global__ void k_apply_topology(int32_t* child0, int32_t* child1, int2* up, int u, int w, int v, int z) {
    int y = child0[v] == z ? child1[v] : child0[v];
    child0[v] = min(w, y);
    child1[v] = max(w, y);
    child0[u] = min(v, z);
    child1[u] = max(v, z);
    up[w] = make_int2(v, y);
    up[y] = make_int2(v, w);
    up[v] = make_int2(u, z);
    up[z] = make_int2(u, v);
}

// unvisit(t,0,v) + small_rectilinear(t,0) after the move (src/copy_number.cxx:408-409): rewrite
// the intervals of v, u, parent(u), ..., root.  Every thread owns 16 positions and walks the path
// with its running interval in registers; step k's cost goes to path_cost[k].
__global__ void __launch_bounds__(256)
k_rescore_path(const int32_t* __restrict__ child0, const int32_t* __restrict__ child1, const int2* __restrict__ up,
               uint8_t* __restrict__ start, uint8_t* __restrict__ end, int32_t* __restrict__ path_cost, int v, int l_pad) {
    const int chunk = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = chunk * 16 < l_pad;
    const size_t off = (size_t)chunk * 16;
    iv16 acc;
    uint32_t os[8], oe[8];
    int a = v, step = 0;
    if (active) {
        int c0 = child0[v], c1 = child1[v];
        widen(ldg16(start, (size_t)c0 * l_pad + off), acc.s);
        widen(ldg16(end, (size_t)c0 * l_pad + off), acc.e);
        widen(ldg16(start, (size_t)c1 * l_pad + off), os);
        widen(ldg16(end, (size_t)c1 * l_pad + off), oe);
    }
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
        int2 ps = up[a];
        if (active) {
            widen(ldg16(start, (size_t)ps.y * l_pad + off), os);
            widen(ldg16(end, (size_t)ps.y * l_pad + off), oe);
        }
        a = ps.x;
        step++;
    }
}

// Cumulative scores along the re-scored path (src/copy_number.cxx:264), one thread.
__global__ void k_path_scores(const int32_t* __restrict__ child0, const int32_t* __restrict__ child1, const int2* __restrict__ up,
                              const int32_t* __restrict__ path_cost, int32_t* __restrict__ score, int v) {
    int a = v, step = 0;
    while (true) {
        score[a] = path_cost[step] + score[child0[a]] + score[child1[a]];
        if (a == 0) break;
        a = up[a].x;
        step++;
    }
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
