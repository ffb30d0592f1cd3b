// lzg_tree_climb.cuh -- hill_climb (/root/reference/src/copy_number.cxx:375-417) with the host out of
// the loop.  Included by lzg_tree.cu inside its anonymous namespace, after lzg_tree_nni.cuh.
//
// One iteration of the reference's loop is greedy_nni (score the whole neighbourhood, pick) ->
// nni + unvisit + small_rectilinear (accept).  Here that is four kernels that read their operands from
// device memory only, so the host can queue iterations blindly -- or not at all: the whole climb
// runs as ONE CUDA graph whose WHILE node repeats {k_apply2, k_up_df, k_nni_items, k_nni_argmin} until
// the arg-min finds no improving candidate (cudaGraphSetConditional from the last kernel).
//
// What makes it possible: inside one hill_climb the shuffled
// edge order is fixed and an edge index stays with its child end (src/copy_number.cxx:381-388,
// 401-406), and an NNI never changes which nodes are internal.  So the candidates of the internal
// non-root node v are ALWAYS number cbase[v] (z = the child of v with the smaller id, the std::set
// order of :340-346) and cbase[v] + 1: the work items never move, an accepted move rewrites five of
// them (those of u, v, w, z and y) and nothing is packed or validated on the host between passes.
//
//  * k_up_df   UP intervals by heavy-path walkers, DATAFLOW scheduled: a path's head is pushed to a
//              per-chunk-group ready queue by the walker that publishes the UP row it hangs off
//              (atomicAdd on the tail, st.release of epoch << 32 | node); persistent CTAs draw tickets
//              (slot, group) and wait for queue entry `slot`.  An entry is always pushed while a
//              smaller ticket is processed, and tickets are drawn by running CTAs in order, so the
//              waits cannot deadlock.  The walk follows the node4 records (heavy child first) two
//              nodes ahead; there are no host-built path tables, so an accepted move only has to
//              patch two node4 records and the path count.
//  * k_apply2  accepts the winner that k_nni_argmin left in ctl[BEST] (or an explicit move): derives
//              (u, w, v, z) from the topology arrays, rewrites the intervals along v .. root following
//              the (parent, sibling) table two steps ahead, and its last CTA turns the per-node
//              costs into cumulative scores, patches topology, node4 records, path count, the five
//              item records, logs the move and applies hill_climb's stopping rule (:411-412).
#pragma once

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// node4 record of an internal node with children c0 < c1 whose records are current (the device twin of
// make_node4 in lzg_tree.cu): heavy = more internal nodes below, c0 on a tie.
__device__ __forceinline__ int4 make_node4_dev(const int4* nd, int c0, int c1, int parent) {
    const int z0 = nd[c0].z >> 2, z1 = nd[c1].z >> 2;
    const bool h0 = z0 >= z1;
    return make_int4(h0 ? c0 : c1, h0 ? c1 : c0, ((1 + z0 + z1) << 2) | (((h0 ? z1 : z0) > 0) << 1) | ((h0 ? z0 : z1) > 0), parent);
}
// path heads among the children of p: every internal child of the root; below it the light child when
// both children are internal (the heavy one continues its parent's path)
__device__ __forceinline__ int heads_below(int p, const int4& rec) {
    return p == 0 ? (rec.z & 1) + ((rec.z >> 1) & 1) : ((rec.z & 3) == 3 ? 1 : 0);
}

// ---- UP pass, dataflow -----------------------------------------------------------------------------
// One heavy path x one chunk group.  head = first node of the path; its parent's UP row is published.
__device__ __forceinline__ void up_walk_df(const lzg_pass_args& A, int head, int g, int tid, int epoch) {
    const int chunk = g * blockDim.x + tid;
    const bool active = chunk * 16 < A.l_pad;
    const size_t off = (size_t)chunk * 16, lp = (size_t)A.l_pad;
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    int4 na = __ldg(A.node4 + head);
    const int p = na.w;
    const int4 pn = __ldg(A.node4 + p);
    const int sib = pn.x == head ? pn.y : pn.x;
    iv16 acc;
    uint32_t os[8], oe[8];
    uint4 l1s = z4, l1e = z4, l2s = z4, l2e = z4;
    int4 nb = make_int4(-1, -1, 0, -1);
    if (na.z & 1) {
        nb = __ldg(A.node4 + na.x);
        if (active) l1s = ldg16(A.start, na.y * lp + off), l1e = ldg16(A.end, na.y * lp + off);
    }
    if (active) {
        const uint4 ss = ldg16(A.start, sib * lp + off), se = ldg16(A.end, sib * lp + off);
        if (p != 0) {
            widen(ldcg16(A.ustart, p * lp + off), acc.s);
            widen(ldcg16(A.uend, p * lp + off), acc.e);
            widen(ss, os);
            widen(se, oe);
            merge16(acc, os, oe);
        } else {
            widen(ss, acc.s);
            widen(se, acc.e);
        }
    }
    int a = head;
    for (;;) {
        int4 nc = make_int4(-1, -1, 0, -1);
        if ((na.z & 1) && (nb.z & 1)) {  // two nodes ahead: the record after next, the rows the next merge reads
            nc = __ldg(A.node4 + nb.x);
            if (active) l2s = ldg16(A.start, nb.y * lp + off), l2e = ldg16(A.end, nb.y * lp + off);
        }
        if (active) {
            *reinterpret_cast<uint4*>(A.ustart + a * lp + off) = narrow(acc.s);
            *reinterpret_cast<uint4*>(A.uend + a * lp + off) = narrow(acc.e);
        }
        if (na.z & 2) {  // the light child is internal: its path may start (uniform over the CTA)
            __syncthreads();
            if (tid == 0) {
                const int pos = atomicAdd(A.tail + g, 1);
                st_release_u64(A.queue + (size_t)g * A.qcap + pos, ((unsigned long long)(uint32_t)epoch << 32) | (uint32_t)na.y);
            }
        }
        if (!(na.z & 1)) break;  // the heavy child is a leaf: end of the path
        if (active) {
            widen(l1s, os);
            widen(l1e, oe);
            merge16(acc, os, oe);
        }
        a = na.x;
        na = nb;
        nb = nc;
        l1s = l2s;
        l1e = l2e;
    }
}

// Persistent CTAs.  Self-resetting: the last CTA to leave zeroes the ticket and the tails and advances
// the epoch, so the next pass needs no memset.  A wait that lasts implausibly long (a path count that
// does not match the node4 records would leave a ticket without a producer) raises STOP = 4 instead of
// hanging the GPU.
template <bool DESC>
__global__ void __launch_bounds__(256)
k_up_df(const lzg_pass_args A_) {
    const lzg_pass_args& A = DESC ? c_pass : A_;
    __shared__ int s_ticket, s_head;
    const int tid = threadIdx.x;
    if (DESC && __ldcg(A.ctl + LZG_CTL_STOP)) return;
    const int epoch = __ldcg(A.ctl + LZG_CTL_EPOCH);
    const int total = __ldcg(A.ctl + LZG_CTL_NPATHS) * A.groups;
    const int4 r0 = __ldg(A.node4);
    const int nroot = (r0.z & 1) + ((r0.z >> 1) & 1);
    for (;;) {
        if (tid == 0) s_ticket = atomicAdd(A.ctl + LZG_CTL_TICKET, 1);
        __syncthreads();
        const int ticket = s_ticket;
        if (ticket >= total) break;
        const int g = ticket % A.groups, slot = ticket / A.groups;
        int head;
        if (slot < nroot) {
            head = slot == 0 ? r0.x : r0.y;
        } else {
            if (tid == 0) {
                const unsigned long long* q = A.queue + (size_t)g * A.qcap + (slot - nroot);
                unsigned long long e = ld_acquire_u64(q);
                long long spins = 0;
                while ((uint32_t)(e >> 32) != (uint32_t)epoch) {
                    __nanosleep(32);
                    if (++spins > (1ll << 24)) {  // > 0.5 s: the protocol is broken, do not hang the device
                        A.ctl[LZG_CTL_STOP] = 4;
                        e = ((unsigned long long)(uint32_t)epoch << 32) | 0xffffffffu;
                        break;
                    }
                    e = ld_acquire_u64(q);
                }
                s_head = (int)(uint32_t)(e & 0xffffffffu);
            }
            __syncthreads();
            head = s_head;
        }
        if (head >= 0) up_walk_df(A, head, g, tid, epoch);
        __syncthreads();  // s_ticket / s_head are rewritten by the next round
    }
    if (tid == 0) {
        __threadfence();
        if (atomicAdd(A.ctl + LZG_CTL_UPDONE, 1) == (int)gridDim.x - 1) {
            A.ctl[LZG_CTL_TICKET] = 0;
            A.ctl[LZG_CTL_UPDONE] = 0;
            for (int g = 0; g < A.groups; g++) A.tail[g] = 0;
            A.ctl[LZG_CTL_EPOCH] = epoch + 1;
        }
    }
}

// ---- accept a move -----------------------------------------------------------------------------------
#define LZG_PATH_CAP 1024  // steps of the path kept in shared memory for the parallel score update; longer paths finish sequentially

struct lzg_move {
    int u, w, v, z;
};

// CLIMB: the move is the winner k_nni_argmin left in ctl[BEST]; otherwise `m` (validated by the host).
template <bool CLIMB>
__global__ void __launch_bounds__(256)
k_apply2(const lzg_pass_args A_, const lzg_move m) {
    const lzg_pass_args& A = CLIMB ? c_pass : A_;
    __shared__ int2 s_path[LZG_PATH_CAP];  // (node written, sibling merged) per step
    __shared__ int32_t s_term[LZG_PATH_CAP];
    __shared__ int s_is_last;
    const int tid = threadIdx.x;
    int u, w, v, z, y;
    if (CLIMB) {
        if (__ldcg(A.ctl + LZG_CTL_STOP)) return;
        const int best = __ldcg(A.ctl + LZG_CTL_BEST);
        v = __ldg(A.pair_v + (best >> 1));
        const int a0 = A.child0[v], a1 = A.child1[v];
        z = (best & 1) ? a1 : a0;  // candidate cbase[v] swaps the child with the smaller id, cbase[v] + 1 the other
        y = (best & 1) ? a0 : a1;
        const int2 e = A.up[v];
        u = e.x;
        w = e.y;
    } else {
        u = m.u, w = m.w, v = m.v, z = m.z;
        y = A.child0[v] == z ? A.child1[v] : A.child0[v];
    }
    const int chunk = blockIdx.x * blockDim.x + tid;
    const bool active = chunk * 16 < A.l_pad;
    const size_t off = (size_t)chunk * 16, lp = (size_t)A.l_pad;
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    iv16 acc;
    uint32_t os[8], oe[8];
    // (parent, sibling) two steps ahead of the merge; the entries from u upwards are not touched by the move
    int2 e1 = A.up[u];
    int2 e2 = e1.x >= 0 ? A.up[e1.x] : make_int2(-1, -1);
    uint4 ys = z4, ye = z4, zs = z4, ze = z4, q1s = z4, q1e = z4, q2s = z4, q2e = z4;
    if (active) {
        widen(ldg16(A.start, w * lp + off), acc.s);
        widen(ldg16(A.end, w * lp + off), acc.e);
        ys = ldg16(A.start, y * lp + off), ye = ldg16(A.end, y * lp + off);
        zs = ldg16(A.start, z * lp + off), ze = ldg16(A.end, z * lp + off);
        if (e1.x >= 0) q1s = ldg16(A.start, e1.y * lp + off), q1e = ldg16(A.end, e1.y * lp + off);
    }
    int k = 0;
    auto step = [&](int node, int other, const uint4& rs, const uint4& re) {
        uint32_t c = 0;
        if (active) {
            widen(rs, os);
            widen(re, oe);
            c = merge16(acc, os, oe);
            *reinterpret_cast<uint4*>(A.wstart + node * lp + off) = narrow(acc.s);
            *reinterpret_cast<uint4*>(A.wend + node * lp + off) = narrow(acc.e);
        }
        c = warp_sum(c);
        if ((tid & 31) == 0 && c) atomicAdd(A.pcost + node, (int32_t)c);
        if (tid == 0 && k < LZG_PATH_CAP) s_path[k] = make_int2(node, other);
        k++;
    };
    step(v, y, ys, ye);  // the new v = merge(w, y)
    step(u, z, zs, ze);  // the new u = merge(v, z)
    while (e1.x >= 0) {
        int2 e3 = make_int2(-1, -1);
        if (e2.x >= 0) {
            e3 = A.up[e2.x];
            if (active) q2s = ldg16(A.start, e2.y * lp + off), q2e = ldg16(A.end, e2.y * lp + off);
        }
        step(e1.x, e1.y, q1s, q1e);
        e1 = e2;
        e2 = e3;
        q1s = q2s;
        q1e = q2e;
    }
    const int n_steps = k;
    __threadfence();
    __syncthreads();
    if (tid == 0) s_is_last = atomicAdd(A.ctl + LZG_CTL_DONE, 1) == (int)gridDim.x - 1;
    __syncthreads();
    if (!s_is_last) return;
    __threadfence();
    // cumulative scores along the path (src/copy_number.cxx:264): score[node_k] = cost_k + score[other_k] + score[node_(k-1)]
    const int n_listed = min(n_steps, LZG_PATH_CAP);
    for (int q = tid; q < n_listed; q += blockDim.x) {
        const int2 pq = s_path[q];
        s_term[q] = __ldcg(A.pcost + pq.x) + A.wscore[pq.y];
        A.pcost[pq.x] = 0;
    }
    __syncthreads();
    if (tid != 0) return;
    int32_t s = A.wscore[w];
    for (int q = 0; q < n_listed; q++) {
        s += s_term[q];
        A.wscore[s_path[q].x] = s;
    }
    for (int q = n_listed, b = s_path[LZG_PATH_CAP - 1].x; q < n_steps; q++) {  // beyond the shared list: follow the table
        const int2 ps = A.up[b];
        s += __ldcg(A.pcost + ps.x) + A.wscore[ps.y];
        A.pcost[ps.x] = 0;
        A.wscore[ps.x] = s;
        b = ps.x;
    }
    // nni(t, u, w, v, z), src/copy_number.cxx:279-284
    A.child0[v] = min(w, y), A.child1[v] = max(w, y);
    A.child0[u] = min(v, z), A.child1[u] = max(v, z);
    A.up[w] = make_int2(v, y);
    A.up[y] = make_int2(v, w);
    A.up[v] = make_int2(u, z);
    A.up[z] = make_int2(u, v);
    if (A.node4) {
        // heavy / light roles: only the subtree count of v changes (u keeps its node set)
        const int4 nu = A.node4[u], nv = A.node4[v];
        const int before = heads_below(u, nu) + heads_below(v, nv);
        A.node4[w].w = v;
        A.node4[z].w = u;
        const int4 nv2 = make_node4_dev(A.node4, min(w, y), max(w, y), u);
        A.node4[v] = nv2;
        const int4 nu2 = make_node4_dev(A.node4, min(v, z), max(v, z), nu.w);
        A.node4[u] = nu2;
        A.ctl[LZG_CTL_NPATHS] += heads_below(u, nu2) + heads_below(v, nv2) - before;
    }
    if (A.witems) {  // the items of v, w, z, y (parent, sibling or children changed) and of u (children changed)
        const int xs[5] = {v, w, z, y, u};
#pragma unroll
        for (int q = 0; q < 5; q++) {
            const int x = xs[q];
            if (x == 0 || A.child0[x] < 0) continue;  // the root and the leaves have no item
            const int slot = A.slot_of[x], cb = A.cbase[x];
            const int2 ps = A.up[x];
            A.witems[2 * slot] = make_int4(ps.x, ps.y, x, A.child0[x]);
            A.witems[2 * slot + 1] = make_int4(A.child1[x], cb, cb + 1, 0);
        }
    }
    A.ctl[LZG_CTL_DONE] = 0;
    A.ctl[LZG_CTL_PARTROOT] = s;
    if (CLIMB) {
        // under position sharding s is this rank's part; the new global score is the winner's reduced score
        const int32_t new_score = A.sharded ? A.ctl[LZG_CTL_SCORE] : s;
        const int nm = A.ctl[LZG_CTL_NMOVES];
        const int slot = A.ctl[LZG_CTL_NLOGGED];
        if (slot < A.log_cap) {
            int32_t* r = A.mlog + 5 * (size_t)slot;
            r[0] = u, r[1] = w, r[2] = v, r[3] = z, r[4] = new_score;
            A.ctl[LZG_CTL_NLOGGED] = slot + 1;
        }
        A.ctl[LZG_CTL_NMOVES] = nm + 1;
        int stop = 0;
        if (A.ctl[LZG_CTL_CURRENT] <= new_score) stop = 2;  // hill_climb :412
        else A.ctl[LZG_CTL_CURRENT] = new_score;
        if (!stop && (nm + 1 >= A.ctl[LZG_CTL_MAXMOVES] || slot + 1 >= A.log_cap)) stop = 3;
        if (stop) A.ctl[LZG_CTL_STOP] = stop;
    } else {
        A.ctl[LZG_CTL_SCORE] = s;
    }
}
