// lzg_tree_climb.cuh -- hill_climb (/root/reference/src/copy_number.cxx:375-417) with the host out of
// the loop and with INCREMENTAL passes.  Included by lzg_tree.cu inside its anonymous namespace, after
// lzg_tree_nni.cuh.
//
// One iteration of the reference's loop is greedy_nni (score the whole neighbourhood, pick) ->
// nni + unvisit + small_rectilinear (accept).  Here every kernel of an iteration reads its operands
// from device memory only, so the whole climb runs as ONE CUDA graph whose WHILE node repeats the
// iteration until the arg-min finds no improving candidate (cudaGraphSetConditional from the last
// kernel); where a collective sits inside the iteration (positions sharded over GPUs) the host
// queues the same kernels in batches.
//
// (1) Static work items.  Inside one hill_climb the shuffled
// edge order is fixed and an edge index stays with its child end (src/copy_number.cxx:381-388,
// 401-406), and an NNI never changes which nodes are internal.  So the candidates of the internal
// non-root node v are ALWAYS number cbase[v] (z = the child of v with the smaller id, the std::set
// order of :340-346) and cbase[v] + 1: the work items never move, an accepted move rewrites five of
// them (those of u, v, w, z and y), nothing is packed or validated on the host between passes.
//
// (2) Incremental passes.  An accepted move changes very little: the DOWN intervals of some nodes on
// the path v .. root, and through them the UP intervals of a few subtrees hanging off that path
// (measured on cfg4/cfg5-like inputs: ~1 % of the UP rows and 2-4 % of the work items change per
// move).  Every candidate score is a sum over positions of terms that depend on five rows only, so:
//  * k_apply2 compares every row it rewrites with the old one and stamps the chunk groups that
//    really changed (dn_stamp), and seeds the UP queue: the sibling of every changed path node, plus
//    v and z whose place in the tree changed (UP of the path nodes themselves cannot change: their
//    inputs are outside the path);
//  * k_up_delta recomputes exactly the queued UP rows, top-down: a row that comes out different is
//    stamped (up_stamp) and its internal children are queued -- the heavy child continues in
//    registers, the light one goes through the queue; a row that comes out equal ends the wave;
//  * k_dirty_scan lists the (item, chunk group) pairs that read a stamped row or a rewritten record,
//    k_items_wl re-scores only those and keeps their position sums in `part`;
//  * k_cand_assemble adds the cumulative-score terms (they change along the whole path) to the
//    position sums of EVERY item -- one thread per item -- and k_nni_argmin picks as before.
// The first pass of a climb runs the full UP pass (k_up_df) and marks everything dirty.  Scores are
// identical to a full pass by construction: an item is re-scored whenever one of its inputs changed.
//
//  * k_up_df   the full UP pass: heavy-path walkers, DATAFLOW scheduled -- a path's head is pushed to a
//              per-chunk-group ready queue by the walker that publishes the UP row it hangs off;
//              persistent CTAs draw tickets (slot, group) and wait for queue entry `slot`.  An entry
//              is always pushed while a smaller ticket is processed, and tickets are drawn by
//              running CTAs in order, so the waits cannot deadlock.  No host-built path tables.
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
__device__ __forceinline__ bool differs(const uint4& a, const uint4& b) {
    return ((a.x ^ b.x) | (a.y ^ b.y) | (a.z ^ b.z) | (a.w ^ b.w)) != 0;
}

// ---- full UP pass, dataflow --------------------------------------------------------------------------
// One heavy path x one chunk group.  head = first node of the path; its parent's UP row is published.
template <int W>
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
            widen_w<W>(ldcg16(A.ustart, p * lp + off), acc.s);
            widen_w<W>(ldcg16(A.uend, p * lp + off), acc.e);
            widen_w<W>(ss, os);
            widen_w<W>(se, oe);
            merge_w<W>(acc, os, oe);
        } else {
            widen_w<W>(ss, acc.s);
            widen_w<W>(se, acc.e);
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
            *reinterpret_cast<uint4*>(A.ustart + a * lp + off) = narrow_w<W>(acc.s);
            *reinterpret_cast<uint4*>(A.uend + a * lp + off) = narrow_w<W>(acc.e);
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
            widen_w<W>(l1s, os);
            widen_w<W>(l1e, oe);
            merge_w<W>(acc, os, oe);
        }
        a = na.x;
        na = nb;
        nb = nc;
        l1s = l2s;
        l1e = l2e;
    }
}

// Persistent CTAs.  Self-resetting: the last CTA to leave zeroes the ticket and the tails, so the next
// full pass needs no memset.  A wait that lasts implausibly long (a path count that does not match the
// node4 records would leave a ticket without a producer) raises STOP = 4 instead of hanging the GPU.
template <int W>
__global__ void __launch_bounds__(256)
k_up_df() {
    const lzg_pass_args& A = c_pass;
    __shared__ int s_ticket, s_head;
    const int tid = threadIdx.x;
    if (__ldcg(A.ctl + LZG_CTL_STOP)) return;
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
        if (head >= 0) up_walk_df<W>(A, head, g, tid, epoch);
        __syncthreads();  // s_ticket / s_head are rewritten by the next round
    }
    if (tid == 0) {
        __threadfence();
        if (atomicAdd(A.ctl + LZG_CTL_UPDONE, 1) == (int)gridDim.x - 1) {
            A.ctl[LZG_CTL_TICKET] = 0;
            A.ctl[LZG_CTL_UPDONE] = 0;
            for (int g = 0; g < A.groups; g++) A.tail[g] = 0;
        }
    }
}

// ---- the queue of UP rows to recompute (incremental pass) -------------------------------------------------
// Pushed by k_apply2 (before k_up_delta starts) and by the walkers of k_up_delta (while it runs: payload,
// fence, flag).
// LIVE = consumers may be running (k_up_delta's own pushes): the payload is fenced before the flag.  k_apply2's seeds are
// read by the next kernel only: plain stores.
template <bool LIVE>
__device__ __forceinline__ void uq_push(const lzg_pass_args& A, int stamp, int node_and_flags, int g, int parent, int sib) {
    const int pos = atomicAdd(A.xctl + LZG_X_UQ_PUSHED, 1);
    if (pos < A.ucap) {
        A.uq[pos] = make_int4(node_and_flags, g, parent, sib);
        if (LIVE) {
            __threadfence();
            st_release(A.uq_flag + pos, stamp);
        } else {
            A.uq_flag[pos] = stamp;
        }
    }
}

// ---- accept a move -----------------------------------------------------------------------------------
#define LZG_PATH_CAP 1024  // steps of the path kept in shared memory; longer paths continue in global scratch

// The move is the winner k_nni_argmin left in ctl[BEST].  One CTA per chunk group.
template <int W>
__global__ void __launch_bounds__(256)
k_apply2() {
    const lzg_pass_args& A = c_pass;
    __shared__ int2 s_path[LZG_PATH_CAP];  // (node written, sibling merged) per step
    __shared__ int32_t s_term[LZG_PATH_CAP];
    __shared__ int32_t s_chg[LZG_PATH_CAP];
    __shared__ int s_is_last;
    const int tid = threadIdx.x, g = blockIdx.x;
    if (__ldcg(A.ctl + LZG_CTL_STOP)) return;
    const int stamp = __ldcg(A.ctl + LZG_CTL_EPOCH);
    const int best = __ldcg(A.ctl + LZG_CTL_BEST);
    const int v = __ldg(A.pair_v + (best >> 1));
    const int a0 = A.child0[v], a1 = A.child1[v];
    const int z = (best & 1) ? a1 : a0;  // candidate cbase[v] swaps the child with the smaller id, cbase[v] + 1 the other
    const int y = (best & 1) ? a0 : a1;
    const int2 uw = A.up[v];
    const int u = uw.x, w = uw.y;
    const int chunk = g * blockDim.x + tid;
    const bool active = chunk * 16 < A.l_pad;
    const size_t off = (size_t)chunk * 16, lp = (size_t)A.l_pad;
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    int2* const deep_path = A.deep_path + (size_t)g * A.n_nodes;
    int32_t* const deep_chg = A.deep_chg + (size_t)g * A.n_nodes;
    iv16 acc;
    uint32_t os[8], oe[8];
    // (parent, sibling) two steps ahead of the merge; the entries from u upwards are not touched by the move
    int2 e1 = A.up[u];
    int2 e2 = e1.x >= 0 ? A.up[e1.x] : make_int2(-1, -1);
    uint4 ys = z4, ye = z4, zs = z4, ze = z4, q1s = z4, q1e = z4, q2s = z4, q2e = z4;
    uint4 ovs = z4, ove = z4, ous = z4, oue = z4, o1s = z4, o1e = z4, o2s = z4, o2e = z4;  // the rows about to be overwritten
    if (active) {
        widen_w<W>(ldg16(A.start, w * lp + off), acc.s);
        widen_w<W>(ldg16(A.end, w * lp + off), acc.e);
        ys = ldg16(A.start, y * lp + off), ye = ldg16(A.end, y * lp + off);
        zs = ldg16(A.start, z * lp + off), ze = ldg16(A.end, z * lp + off);
        ovs = ldg16(A.start, v * lp + off), ove = ldg16(A.end, v * lp + off);
        ous = ldg16(A.start, u * lp + off), oue = ldg16(A.end, u * lp + off);
        if (e1.x >= 0) {
            q1s = ldg16(A.start, e1.y * lp + off), q1e = ldg16(A.end, e1.y * lp + off);
            o1s = ldg16(A.start, e1.x * lp + off), o1e = ldg16(A.end, e1.x * lp + off);
        }
    }
    for (int q = tid; q < LZG_PATH_CAP; q += blockDim.x) s_term[q] = 0, s_chg[q] = 0;
    __syncthreads();
    int k = 0;
    // step k: node = merge(running interval, other); costs and "changed" flags are fire-and-forget
    // shared-memory reductions (no shuffle chain in the walk), read back once the walk is over
    auto step = [&](int node, int other, const uint4& rs, const uint4& re, const uint4& olds, const uint4& olde) {
        uint32_t c = 0;
        bool chg = false;
        if (active) {
            widen_w<W>(rs, os);
            widen_w<W>(re, oe);
            c = merge_w<W>(acc, os, oe);
            const uint4 ns = narrow_w<W>(acc.s), ne = narrow_w<W>(acc.e);
            chg = differs(ns, olds) || differs(ne, olde);
            *reinterpret_cast<uint4*>(A.wstart + node * lp + off) = ns;
            *reinterpret_cast<uint4*>(A.wend + node * lp + off) = ne;
        }
        if (k < LZG_PATH_CAP) {
            c = __reduce_add_sync(0xffffffffu, c);  // one REDUX per warp: 256 same-address shared atomics would serialise
            if ((tid & 31) == 0 && c) atomicAdd(&s_term[k], (int32_t)c);
            if (chg) s_chg[k] = 1;
            if (tid == 0) s_path[k] = make_int2(node, other);
        } else {
            c = warp_sum(c);
            if ((tid & 31) == 0 && c) atomicAdd(A.pcost + node, (int32_t)c);
            if (chg) deep_chg[k] = 1;
            if (tid == 0) deep_path[k] = make_int2(node, other);
        }
        k++;
    };
    step(v, y, ys, ye, ovs, ove);  // the new v = merge(w, y)
    step(u, z, zs, ze, ous, oue);  // the new u = merge(v, z)
    while (e1.x >= 0) {
        int2 e3 = make_int2(-1, -1);
        if (e2.x >= 0) {
            e3 = A.up[e2.x];
            if (active) {
                q2s = ldg16(A.start, e2.y * lp + off), q2e = ldg16(A.end, e2.y * lp + off);
                o2s = ldg16(A.start, e2.x * lp + off), o2e = ldg16(A.end, e2.x * lp + off);
            }
        }
        step(e1.x, e1.y, q1s, q1e, o1s, o1e);
        e1 = e2;
        e2 = e3;
        q1s = q2s, q1e = q2e;
        o1s = o2s, o1e = o2e;
    }
    const int n_steps = k;
    __threadfence();  // the deep-path records of this CTA, read back below by other threads of it
    __syncthreads();
    auto path_at = [&](int q) -> int2 { return q < LZG_PATH_CAP ? s_path[q] : deep_path[q]; };
    // this chunk group's share: costs, stamps of the rows that changed, UP rows to recompute
    for (int q = tid; q < n_steps; q += blockDim.x) {
        const bool chg = q < LZG_PATH_CAP ? s_chg[q] != 0 : __ldcg(deep_chg + q) != 0;
        if (q >= LZG_PATH_CAP) deep_chg[q] = 0;
        if (q < LZG_PATH_CAP && s_term[q]) atomicAdd(A.pcost + s_path[q].x, s_term[q]);
        if (!chg) continue;
        const int2 pq = path_at(q);
        A.dn_stamp[(size_t)pq.x * A.groups + g] = stamp;
        // DOWN(node_q) is an input of UP(sibling of node_q) = UP(other_(q+1)); node_0 = v: its sibling z is queued anyway
        if (q >= 1 && q + 1 < n_steps) {
            const int2 nx = path_at(q + 1);
            if (__ldg(A.node4 + nx.y).x >= 0) uq_push<false>(A, stamp, nx.y, g, nx.x, pq.x);
        }
    }
    if (tid == 0) {
        // v hangs under u beside z now, and its children w, y got a new parent / sibling: their UP rows are
        // recomputed whether or not UP(v) comes out different
        uq_push<false>(A, stamp, v | LZG_UQ_FORCE, g, u, z);
        if (__ldg(A.node4 + z).x >= 0) uq_push<false>(A, stamp, z, g, u, v);
    }
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
    for (int q = n_listed; q < n_steps; q++) {  // beyond the shared list
        const int2 pq = deep_path[q];
        s += __ldcg(A.pcost + pq.x) + A.wscore[pq.y];
        A.pcost[pq.x] = 0;
        A.wscore[pq.x] = s;
    }
    // nni(t, u, w, v, z), src/copy_number.cxx:279-284
    A.child0[v] = min(w, y), A.child1[v] = max(w, y);
    A.child0[u] = min(v, z), A.child1[u] = max(v, z);
    A.up[w] = make_int2(v, y);
    A.up[y] = make_int2(v, w);
    A.up[v] = make_int2(u, z);
    A.up[z] = make_int2(u, v);
    {
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
    {   // the items of v, w, z, y (parent, sibling or children changed) and of u (children changed)
        const int xs[5] = {v, w, z, y, u};
#pragma unroll
        for (int q = 0; q < 5; q++) {
            const int x = xs[q];
            if (x == 0 || A.child0[x] < 0) continue;  // the root and the leaves have no item
            const int slot = A.slot_of[x], cb = A.cbase[x];
            const int2 ps = A.up[x];
            A.witems[2 * slot] = make_int4(ps.x, ps.y, x, A.child0[x]);
            A.witems[2 * slot + 1] = make_int4(A.child1[x], cb, cb + 1, 0);
            A.rec_stamp[slot] = stamp;
        }
    }
    A.ctl[LZG_CTL_DONE] = 0;
    A.ctl[LZG_CTL_PARTROOT] = s;
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
}

// ---- incremental UP pass ---------------------------------------------------------------------------------
// Persistent CTAs drain the queue.  A unit = recompute UP(x) for one chunk group from UP(parent) and
// DOWN(sibling); if the row comes out different (or the unit is forced) its internal children follow: the
// heavy one in registers, the light one through the queue.  A CTA leaves when every pushed unit is done
// (done is read before pushed: pushed only grows and done <= pushed, so equality means nothing is in flight).
template <int W>
__global__ void __launch_bounds__(256)
k_up_delta() {
    const lzg_pass_args& A = c_pass;
    __shared__ int s_unit;
    const int tid = threadIdx.x;
    if (__ldcg(A.ctl + LZG_CTL_STOP)) return;
    const int stamp = __ldcg(A.ctl + LZG_CTL_EPOCH);
    const size_t lp = (size_t)A.l_pad;
    for (;;) {
        if (tid == 0) {
            const int t = atomicAdd(A.xctl + LZG_X_UQ_TICKET, 1);
            int got = -1;
            long long spins = 0;
            for (;;) {
                if (t < A.ucap && ld_acquire(A.uq_flag + t) == stamp) {
                    got = t;
                    break;
                }
                const int done = __ldcg(A.xctl + LZG_X_UQ_DONE);
                const int pushed = __ldcg(A.xctl + LZG_X_UQ_PUSHED);
                if (done == pushed && t >= pushed) break;
                if (++spins > (1ll << 24)) {
                    A.ctl[LZG_CTL_STOP] = 4;
                    break;
                }
                __nanosleep(20);
            }
            s_unit = got;
        }
        __syncthreads();
        const int unit = s_unit;
        if (unit < 0) break;
        const int4 e = __ldcg(A.uq + unit);
        int x = e.x & ~LZG_UQ_FORCE;
        bool force = (e.x & LZG_UQ_FORCE) != 0;
        const int g = e.y, p = e.z, sib = e.w;
        const int chunk = g * blockDim.x + tid;
        const bool active = chunk * 16 < A.l_pad;
        const size_t off = (size_t)chunk * 16;
        iv16 acc;
        uint32_t os[8], oe[8];
        if (active) {
            const uint4 ss = ldg16(A.start, sib * lp + off), se = ldg16(A.end, sib * lp + off);
            if (p != 0) {
                widen_w<W>(ldcg16(A.ustart, p * lp + off), acc.s);
                widen_w<W>(ldcg16(A.uend, p * lp + off), acc.e);
                widen_w<W>(ss, os);
                widen_w<W>(se, oe);
                merge_w<W>(acc, os, oe);
            } else {
                widen_w<W>(ss, acc.s);
                widen_w<W>(se, acc.e);
            }
        }
        for (;;) {
            const int4 nx = __ldg(A.node4 + x);
            bool chg = false;
            uint4 ns = make_uint4(0, 0, 0, 0), ne = ns;
            if (active) {
                ns = narrow_w<W>(acc.s), ne = narrow_w<W>(acc.e);
                chg = differs(ns, ldcg16(A.ustart, x * lp + off)) || differs(ne, ldcg16(A.uend, x * lp + off));
            }
            const int any = __syncthreads_or(chg);
            if (any) {
                if (active) {
                    *reinterpret_cast<uint4*>(A.ustart + x * lp + off) = ns;
                    *reinterpret_cast<uint4*>(A.uend + x * lp + off) = ne;
                }
                if (tid == 0) A.up_stamp[(size_t)x * A.groups + g] = stamp;
            }
            if (!any && !force) break;
            if (nx.z & 2) {  // light child internal: queue it (it reads UP(x) from memory)
                __syncthreads();
                if (tid == 0) uq_push<true>(A, stamp, nx.y, g, x, nx.x);
            }
            if (!(nx.z & 1)) break;
            if (active) {  // continue with the heavy child: UP(heavy) = merge(UP(x), DOWN(light))
                widen_w<W>(ldg16(A.start, nx.y * lp + off), os);
                widen_w<W>(ldg16(A.end, nx.y * lp + off), oe);
                merge_w<W>(acc, os, oe);
            }
            x = nx.x;
            force = false;  // a forced unit (v) recomputes BOTH children whatever came out for v; below them the wave decides
        }
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            atomicAdd(A.xctl + LZG_X_UQ_DONE, 1);
        }
    }
}

// ---- which items read something that changed ------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_dirty_scan() {
    const lzg_pass_args& A = c_pass;
    if (__ldcg(A.ctl + LZG_CTL_STOP)) return;
    const int stamp = __ldcg(A.ctl + LZG_CTL_EPOCH);
    const bool all = __ldcg(A.xctl + LZG_X_FULL) != 0;  // first pass of a climb: everything
    const int q = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per (item slot, chunk group)
    if (q == 0) {  // k_up_delta is done
        A.xctl[LZG_X_UQ_TOTAL] += A.xctl[LZG_X_UQ_PUSHED];
        A.xctl[LZG_X_UQ_PUSHED] = 0, A.xctl[LZG_X_UQ_DONE] = 0, A.xctl[LZG_X_UQ_TICKET] = 0;
    }
    const int G = A.groups;
    if (q >= A.n_items * G) return;
    const int slot = q / G, g = q % G;
    const int4 i0 = __ldg(A.items + 2 * slot), i1 = __ldg(A.items + 2 * slot + 1);
    const int u = i0.x, w = i0.y, a = i0.w, b = i1.x;
    // all six stamps at once (independent loads): the record, DOWN(w), DOWN(a), DOWN(b), DOWN(u), UP(u)
    const int s0 = __ldcg(A.rec_stamp + slot), s1 = __ldcg(A.dn_stamp + (size_t)w * G + g), s2 = __ldcg(A.dn_stamp + (size_t)a * G + g),
              s3 = __ldcg(A.dn_stamp + (size_t)b * G + g), s4 = __ldcg(A.dn_stamp + (size_t)u * G + g), s5 = __ldcg(A.up_stamp + (size_t)u * G + g);
    const bool d = all | (s0 == stamp) | (s1 == stamp) | (s2 == stamp) | (s3 == stamp) | ((u != 0) & ((s4 == stamp) | (s5 == stamp)));
    if (d) {
        reinterpret_cast<int2*>(A.part)[q] = make_int2(0, 0);
        A.wl[atomicAdd(A.xctl + LZG_X_WL_COUNT, 1)] = q;
    }
}

// ---- re-score the listed (item, chunk group) pairs ----------------------------------------------------------------
// Persistent CTAs of `threads` threads (the chunk-group width of the UP / apply kernels).  Position sums
// only: the cumulative-score terms are added per pass by k_cand_assemble.
template <int W>
__global__ void __launch_bounds__(256, 4)
k_items_wl() {
    const lzg_pass_args& A = c_pass;
    __shared__ int s_unit;
    const int tid = threadIdx.x;
    if (__ldcg(A.ctl + LZG_CTL_STOP)) return;
    const int count = __ldcg(A.xctl + LZG_X_WL_COUNT);
    for (;;) {
        if (tid == 0) s_unit = atomicAdd(A.xctl + LZG_X_WL_TICKET, 1);
        __syncthreads();
        const int t = s_unit;
        if (t >= count) break;
        const int e = __ldcg(A.wl + t);
        const int slot = e / A.groups, g = e % A.groups;
        const int4 i0 = __ldg(A.items + 2 * slot), i1 = __ldg(A.items + 2 * slot + 1);
        const int u = i0.x, w = i0.y, a = i0.w, b = i1.x;
        const size_t off = ((size_t)g * blockDim.x + tid) * 16;
        int32_t pA = 0, pB = 0;
        if (off < (size_t)A.l_pad) {
            if (W == 1) {
                if (u) score_pair<16, true>(A, u, w, a, b, off, pA, pB);
                else score_pair<16, false>(A, u, w, a, b, off, pA, pB);
            } else {
                if (u) score_pair_w2<true>(A, u, w, a, b, off, pA, pB);
                else score_pair_w2<false>(A, u, w, a, b, off, pA, pB);
            }
        }
        pA = warp_sum_i(pA);
        pB = warp_sum_i(pB);
        if ((tid & 31) == 0) {
            int32_t* cell = A.part + 2 * ((size_t)slot * A.groups + g);
            if (pA) atomicAdd(cell, pA);
            if (pB) atomicAdd(cell + 1, pB);
        }
        __syncthreads();  // s_unit is rewritten by the next round
    }
}

// ---- every candidate's score = its position sums + the cumulative-score terms -------------------------------------
__global__ void __launch_bounds__(256)
k_cand_assemble() {
    const lzg_pass_args& A = c_pass;
    if (__ldcg(A.ctl + LZG_CTL_STOP)) return;
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot == 0) {
        A.ctl[LZG_CTL_ROOT] = A.score[0];
        A.xctl[LZG_X_WL_TOTAL] += A.xctl[LZG_X_WL_COUNT];
        A.xctl[LZG_X_WL_COUNT] = 0, A.xctl[LZG_X_WL_TICKET] = 0, A.xctl[LZG_X_FULL] = 0;  // k_items_wl is done
    }
    if (slot >= A.n_items) return;
    const int4 i0 = __ldg(A.items + 2 * slot), i1 = __ldg(A.items + 2 * slot + 1);
    const int u = i0.x, w = i0.y, a = i0.w, b = i1.x, ia = i1.y, ib = i1.z;
    int32_t pA = 0, pB = 0;
    const int2* cell = reinterpret_cast<const int2*>(A.part) + (size_t)slot * A.groups;
    for (int g = 0; g < A.groups; g++) {
        const int2 c = __ldcg(cell + g);
        pA += c.x, pB += c.y;
    }
    const int32_t base = A.score[w] + A.score[a] + A.score[b] + (u ? A.score[0] - A.score[u] : 0);
    A.cand_score[ia] = pA + base;
    A.cand_score[ib] = pB + base;
}
