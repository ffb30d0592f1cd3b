// lzg_tree_pass.cuh -- one NNI neighbourhood pass: k_up_pass (UP intervals, heavy-path walkers),
// k_nni_items (both candidates of every edge (u, v) in one work item), k_nni_argmin (the
// reference's choice).  Included by lzg_tree_nni.cuh inside its anonymous namespace.
//
// Measured on B200 (profiles/README.md, round 1): the scoring arithmetic is ~350 integer
// instructions per 16 positions of an item, so the pass is bound by instruction issue and
// round-trip latency rather than by HBM; a persistent producer/consumer ring fed by bulk async
// copies (tried: one producer warp could not describe entries fast enough, and 148 walkers at a
// time starved the dependency chains) lost to plain short-lived CTAs at high occupancy.  So:
//  * k_up_pass: one CTA per (heavy path, chunk group), drawn from a ticket counter in
//    parents-first order; the running UP interval stays in registers, sibling rows are
//    prefetched two nodes ahead, every node's UP row is published with a per-(node, group) epoch
//    flag (stores, bar.sync, st.release by one thread / ld.acquire by one thread, bar.sync, ld.cg) --
//    only where another path's head hangs off the node (one node in three at cfg5): the release
//    fence costs a store round trip (200 -> 185 us at cfg5).  Cross-CTA hops are light edges only
//    (<= log2 n on any root-leaf path).  A ticket is drawn by a running CTA and only waits for
//    smaller tickets: no deadlock.  At cfg5 the kernel moves 0.88 GB of DRAM traffic (0.51 read,
//    0.37 written) in 185 us = 4.8 TB/s, 0.73 of the measured copy peak.  (Tried: persistent warp
//    walkers -- one warp per (path, 512-position slice), static item order, no ticket and no CTA
//    barrier: 212 us, eight times the fences.)
//  * k_nni_items: one CTA per (item, chunk group) behind the kernel boundary -- no flags, no
//    tickets: the item record, then all ten row loads at once, the arithmetic on u16x2 lanes,
//    per-warp partial scores added to cand_score with fire-and-forget REDs.
//  * k_nni_argmin: min (score, index) over the candidates below the current root score
//    (greedy: min index) -- the reference's strict '<', first-wins rule (src/copy_number.cxx:353-362).
#pragma once

// even bytes / odd bytes of a word as u16x2 lanes (the merge is per position, so any fixed lane
// assignment works; the even half is a full-rate LOP3, the odd half one PRMT)
__device__ __forceinline__ void split2(uint32_t x, uint32_t (&o)[2]) {
    o[0] = x & 0x00ff00ffu;
    o[1] = __byte_perm(x, 0, 0x4341);
}

struct lzg_pass_args {
    const int4* path_rec;  // [n_paths] {first index into order, nodes on the path, parent of the head, sibling of the head}
    const int2* order;     // {node, its sibling | LZG_ORDER_PUBLISH}: internal non-root nodes, heavy paths back to back, parents-first
    const int4* items;     // [n_items][2]: {u, w, v, a}, {b, index of the candidate z = a, index of the candidate z = b, 0}
    int n_paths, n_items, n_cand, groups, l_pad, epoch, greedy;
    const uint8_t *start, *end;
    uint8_t *ustart, *uend;
    const int32_t* score;
    int* ready;           // [n_nodes][groups] epoch flags of the UP rows
    int* ctl;             // [0] ticket, [2] best index, [3] best score, [4] last candidate's score, [15] root score
    int32_t* cand_score;  // [n_cand], zeroed
    // ---- the device-driven hill climb (lzg_tree_climb.cuh): everything a whole climb needs, so that its
    //      kernels take no per-tree launch arguments (they read this struct from constant memory) ----
    int4* node4;          // [n_nodes] {heavy child, light child, internal nodes below << 2 | flags, parent}
    int32_t *child0, *child1;
    int2* up;             // (parent, sibling)
    uint8_t *wstart, *wend;  // the DOWN intervals, writable (== start / end)
    int32_t* wscore;
    int32_t* pcost;       // [n_nodes] per-node merge cost of the move being applied; all zero between launches
    int4* witems;         // == items, writable: five records change per accepted move
    const int32_t* slot_of;  // [n_nodes] item slot of an internal non-root node
    const int32_t* cbase;    // [n_nodes] index of the node's first candidate in the host's enumeration (2 x rank among the internal edges)
    const int32_t* pair_v;   // [n_items] v of the candidates 2k, 2k + 1
    unsigned long long* queue;  // [groups][qcap] ready path heads of the dataflow UP pass: epoch << 32 | node
    int* tail;            // [groups] heads pushed so far
    int32_t* mlog;        // [log_cap][5] accepted moves (u, w, v, z, new score)
    unsigned long long* tlog;  // [log_cap + 2] %globaltimer at the end of every pass
    int* xctl;            // counters of the incremental pass (LZG_X_*)
    int32_t *dn_stamp, *up_stamp;  // [n_nodes][groups] iteration in which a chunk group of a DOWN / UP row last changed value
    int32_t* rec_stamp;   // [n_items] iteration in which an item record was rewritten
    int32_t* part;        // [n_items][groups][2] position sums (candidate z = a, z = b; net of c0) of one chunk group of an item
    int4* uq;             // [ucap] UP rows to recompute: {node | LZG_UQ_FORCE, chunk group, parent, sibling}
    int32_t* uq_flag;     // [ucap] stamp of the iteration the entry belongs to
    int32_t* wl;          // [n_items * groups] dirty (item slot, chunk group) pairs: slot * groups + group
    int2* deep_path;      // [groups][n_nodes] (node, sibling) per step of a path longer than LZG_PATH_CAP
    int32_t* deep_chg;    // [groups][n_nodes] "row changed" per such step; all zero between launches
    int n_nodes, ucap, threads;
    int qcap, log_cap, sharded, use_cond;
    unsigned long long cond_handle;  // cudaGraphConditionalHandle of the WHILE node the climb runs in
};

// The climb's copy of the arguments (cudaMemcpyToSymbolAsync before the climb is launched): a CUDA graph
// freezes kernel arguments, and one graph serves every tree of a shape.
__constant__ lzg_pass_args c_pass;

// ctl[LZG_CTL_ROOT] sits directly below cand_score: {root score, candidate scores} is ONE vector of
// n_cand + 1 ints, every entry a sum over positions -- what position-sharded ranks sum-reduce.
enum {
    LZG_CTL_TICKET = 0, LZG_CTL_DONE = 1, LZG_CTL_BEST = 2, LZG_CTL_SCORE = 3, LZG_CTL_LAST = 4,
    // the climb's own state (lzg_tree_climb.cuh)
    LZG_CTL_STOP = 5,      // 0 running, 1 no improving move, 2 the accepted move did not improve, 3 move budget used up, 4 protocol error
    LZG_CTL_NMOVES = 6, LZG_CTL_CURRENT = 7, LZG_CTL_NPATHS = 8, LZG_CTL_PASSES = 9, LZG_CTL_MAXMOVES = 10,
    LZG_CTL_EPOCH = 11, LZG_CTL_UPDONE = 12, LZG_CTL_PARTROOT = 13, LZG_CTL_NLOGGED = 14,
    LZG_CTL_ROOT = 15, LZG_CTL_INTS = 16
};
// xctl: the work queues of the incremental pass
enum { LZG_X_UQ_PUSHED = 0, LZG_X_UQ_DONE = 1, LZG_X_UQ_TICKET = 2, LZG_X_WL_COUNT = 3, LZG_X_WL_TICKET = 4, LZG_X_FULL = 5,
       LZG_X_WL_TOTAL = 6, LZG_X_UQ_TOTAL = 7, LZG_X_INTS = 8 };  // the totals: units processed since the climb started
enum { LZG_UQ_FORCE = 1 << 30 };

// order[].y: the sibling, and LZG_ORDER_PUBLISH if another path's head hangs off the node -- only
// those UP rows are waited for across walkers, only they pay the release fence.
enum { LZG_ORDER_PUBLISH = 1 << 30, LZG_ORDER_NODE = LZG_ORDER_PUBLISH - 1 };

// One heavy path x one chunk group: UP(a) for every node a of the path, top-down, running
// interval in registers; sibling intervals are prefetched two nodes ahead.
template <int W>
__device__ __forceinline__ void walk_path(const lzg_pass_args& A, int item, int tid) {
    const int g = item % A.groups, p = item / A.groups;
    const int chunk = g * blockDim.x + tid;
    const bool active = chunk * 16 < A.l_pad;
    const size_t off = (size_t)chunk * 16;
    const int l_pad = A.l_pad;
    const int4 pr = A.path_rec[p];
    const int k0 = pr.x, k1 = pr.x + pr.y;
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    uint4 ss = z4, se = z4, n1s = z4, n1e = z4, n2s = z4, n2e = z4;
    if (active) {
        ss = ldg16(A.start, (size_t)pr.w * l_pad + off);
        se = ldg16(A.end, (size_t)pr.w * l_pad + off);
    }
    int2 cur = A.order[k0];
    int2 nx1 = k0 + 1 < k1 ? A.order[k0 + 1] : make_int2(0, 0);
    int2 nx2 = k0 + 2 < k1 ? A.order[k0 + 2] : make_int2(0, 0);
    if (active && k0 + 1 < k1) {
        n1s = ldg16(A.start, (size_t)(nx1.y & LZG_ORDER_NODE) * l_pad + off);
        n1e = ldg16(A.end, (size_t)(nx1.y & LZG_ORDER_NODE) * l_pad + off);
    }
    iv16 acc;
    uint32_t os[8], oe[8];
    if (pr.z != 0) {
        if (tid == 0) {
            const int* flag = A.ready + (size_t)pr.z * A.groups + g;
            while (ld_acquire(flag) != A.epoch) __nanosleep(32);
        }
        __syncthreads();
        if (active) {
            widen_w<W>(ldcg16(A.ustart, (size_t)pr.z * l_pad + off), acc.s);
            widen_w<W>(ldcg16(A.uend, (size_t)pr.z * l_pad + off), acc.e);
            widen_w<W>(ss, os);
            widen_w<W>(se, oe);
            merge_w<W>(acc, os, oe);
        }
    } else if (active) {
        widen_w<W>(ss, acc.s);
        widen_w<W>(se, acc.e);
    }
    for (int k = k0;; k++) {
        int2 nx3 = make_int2(0, 0);
        if (k + 2 < k1) {
            if (active) {
                n2s = ldg16(A.start, (size_t)(nx2.y & LZG_ORDER_NODE) * l_pad + off);
                n2e = ldg16(A.end, (size_t)(nx2.y & LZG_ORDER_NODE) * l_pad + off);
            }
            if (k + 3 < k1) nx3 = A.order[k + 3];
        }
        if (active) {
            *reinterpret_cast<uint4*>(A.ustart + (size_t)cur.x * l_pad + off) = narrow_w<W>(acc.s);
            *reinterpret_cast<uint4*>(A.uend + (size_t)cur.x * l_pad + off) = narrow_w<W>(acc.e);
        }
        if (cur.y & LZG_ORDER_PUBLISH) {  // a path head hangs off this node (uniform over the CTA)
            __syncthreads();
            if (tid == 0) st_release(A.ready + (size_t)cur.x * A.groups + g, A.epoch);
        }
        if (k + 1 >= k1) break;
        if (active) {
            widen_w<W>(n1s, os);
            widen_w<W>(n1e, oe);
            merge_w<W>(acc, os, oe);
        }
        cur = nx1;
        nx1 = nx2;
        nx2 = nx3;
        n1s = n2s;
        n1e = n2e;
    }
}


template <int W>
__global__ void __launch_bounds__(256)
k_up_pass(const lzg_pass_args A) {
    __shared__ int s_item;
    if (threadIdx.x == 0) s_item = atomicAdd(A.ctl + LZG_CTL_TICKET, 1);
    __syncthreads();
    walk_path<W>(A, s_item, threadIdx.x);
}

// ---- items -------------------------------------------------------------------------------------------
template <int CPT>
struct lzg_vec {  // a thread's CPT positions of one row
    uint32_t w[CPT / 4];
};
template <int CPT>
__device__ __forceinline__ lzg_vec<CPT> ldg_vec(const uint8_t* p) {
    lzg_vec<CPT> v;
    if (CPT == 16) {
        const uint4 x = __ldg(reinterpret_cast<const uint4*>(p));
        v.w[0] = x.x, v.w[1] = x.y, v.w[CPT / 4 - 2] = x.z, v.w[CPT / 4 - 1] = x.w;
    } else {
        const uint2 x = __ldg(reinterpret_cast<const uint2*>(p));
        v.w[0] = x.x, v.w[1] = x.y;
    }
    return v;
}

__device__ __forceinline__ int32_t warp_sum_i(int32_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Both candidates of an edge (u, v) on CPT positions: cost parts of z = a and z = b, each already
// net of c0 (the current tree re-rooted above u).  rows: w, a, b[, DOWN(u), UP(u)] x {start, end}.
template <int CPT, bool HAS_UP>
__device__ __forceinline__ void score_pair(const lzg_pass_args& A, int u, int w, int a, int b, size_t off, int32_t& pA, int32_t& pB) {
    const size_t lp = (size_t)A.l_pad;
    lzg_vec<CPT> row[HAS_UP ? 10 : 6];
    row[0] = ldg_vec<CPT>(A.start + w * lp + off), row[1] = ldg_vec<CPT>(A.end + w * lp + off);
    row[2] = ldg_vec<CPT>(A.start + a * lp + off), row[3] = ldg_vec<CPT>(A.end + a * lp + off);
    row[4] = ldg_vec<CPT>(A.start + b * lp + off), row[5] = ldg_vec<CPT>(A.end + b * lp + off);
    if (HAS_UP) {
        row[6] = ldg_vec<CPT>(A.start + u * lp + off), row[7] = ldg_vec<CPT>(A.end + u * lp + off);
        row[8] = ldg_vec<CPT>(A.ustart + u * lp + off), row[9] = ldg_vec<CPT>(A.uend + u * lp + off);
    }
    uint32_t cA = 0, cB = 0, c0 = 0;
#pragma unroll
    for (int q = 0; q < CPT / 4; q++) {  // four positions (one word of every row) at a time: short live ranges
        uint32_t ws[2], we[2], as[2], ae[2], bs[2], be[2], ds[2], de[2], us[2], ue[2];
        split2(row[0].w[q], ws), split2(row[1].w[q], we);
        split2(row[2].w[q], as), split2(row[3].w[q], ae);
        split2(row[4].w[q], bs), split2(row[5].w[q], be);
        if (HAS_UP) {
            split2(row[6].w[q], ds), split2(row[7].w[q], de);
            split2(row[8].w[q], us), split2(row[9].w[q], ue);
        }
#pragma unroll
        for (int h = 0; h < 2; h++) {
            // Integer min/max issues at half rate (VIMNMX and the three-input VIMNMX3 alike, 63 lanes/clk/SM
            // measured), IADD3 at full rate, so: end = lo + hi - start instead of max(lo, hi), and the
            // cost of a merge whose interval is not needed, max(0, s - ue, us - e) (both intervals are
            // proper: s <= e), as ONE VIMNMX3 over differences biased by 256 per lane (no borrow between the
            // u16 lanes; each of cA, cB, c0 takes exactly one biased term per word, so the bias cancels in
            // cA - c0 and cB - c0).
            const uint32_t K = 0x01000100u;
            {   // z = a: v keeps b -- merge(w, b), then a, then UP(u)
                const uint32_t lo1 = __vmaxu2(ws[h], bs[h]), hi1 = __vminu2(we[h], be[h]);
                const uint32_t s1 = __vminu2(lo1, hi1), e1 = lo1 + hi1 - s1;
                const uint32_t lo2 = __vmaxu2(s1, as[h]), hi2 = __vminu2(e1, ae[h]);
                const uint32_t s2 = __vminu2(lo2, hi2);
                cA = cA + lo1 - s1;
                cA = cA + lo2 - s2;
                if (HAS_UP) {
                    const uint32_t e2 = lo2 + hi2 - s2;
                    cA += __vimax3_u16x2(s2 + K - ue[h], us[h] + K - e2, K);
                }
            }
            {   // z = b: v keeps a
                const uint32_t lo1 = __vmaxu2(ws[h], as[h]), hi1 = __vminu2(we[h], ae[h]);
                const uint32_t s1 = __vminu2(lo1, hi1), e1 = lo1 + hi1 - s1;
                const uint32_t lo2 = __vmaxu2(s1, bs[h]), hi2 = __vminu2(e1, be[h]);
                const uint32_t s2 = __vminu2(lo2, hi2);
                cB = cB + lo1 - s1;
                cB = cB + lo2 - s2;
                if (HAS_UP) {
                    const uint32_t e2 = lo2 + hi2 - s2;
                    cB += __vimax3_u16x2(s2 + K - ue[h], us[h] + K - e2, K);
                }
            }
            if (HAS_UP)  // the current tree re-rooted above u
                c0 += __vimax3_u16x2(ds[h] + K - ue[h], us[h] + K - de[h], K);
        }
    }
    // per u16 lane: <= 4 words x (2 x 255 + 511) -- no carry between the lanes
    const int32_t z = (int32_t)((c0 & 0xffffu) + (c0 >> 16));
    pA = (int32_t)((cA & 0xffffu) + (cA >> 16)) - z;
    pB = (int32_t)((cB & 0xffffu) + (cB >> 16)) - z;
}

// The same on 16-bit positions (W = 2): a 16-byte chunk holds 8 positions that already are four u16x2 words, so nothing
// is widened; lanes reach 65535, so an interval end is max(lo, hi) (lo + hi would overflow the lane), the distance to
// UP(u) is (max(s, ue) - ue) + (max(us, e) - e) (at most one term is positive) and every word's costs are folded into
// 32 bits at once.  The plain form: the path exists for value ranges no named configuration has.
__device__ __forceinline__ uint32_t fold16(uint32_t d) { return (d & 0xffffu) + (d >> 16); }

template <bool HAS_UP>
__device__ __forceinline__ void score_pair_w2(const lzg_pass_args& A, int u, int w, int a, int b, size_t off, int32_t& pA, int32_t& pB) {
    const size_t lp = (size_t)A.l_pad;
    lzg_vec<16> row[HAS_UP ? 10 : 6];
    row[0] = ldg_vec<16>(A.start + w * lp + off), row[1] = ldg_vec<16>(A.end + w * lp + off);
    row[2] = ldg_vec<16>(A.start + a * lp + off), row[3] = ldg_vec<16>(A.end + a * lp + off);
    row[4] = ldg_vec<16>(A.start + b * lp + off), row[5] = ldg_vec<16>(A.end + b * lp + off);
    if (HAS_UP) {
        row[6] = ldg_vec<16>(A.start + u * lp + off), row[7] = ldg_vec<16>(A.end + u * lp + off);
        row[8] = ldg_vec<16>(A.ustart + u * lp + off), row[9] = ldg_vec<16>(A.uend + u * lp + off);
    }
    uint32_t cA = 0, cB = 0, c0 = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const uint32_t ws = row[0].w[q], we = row[1].w[q], as = row[2].w[q], ae = row[3].w[q], bs = row[4].w[q], be = row[5].w[q];
        uint32_t ds = 0, de = 0, us = 0, ue = 0;
        if (HAS_UP) ds = row[6].w[q], de = row[7].w[q], us = row[8].w[q], ue = row[9].w[q];
        auto gap = [&](uint32_t s, uint32_t e) {  // distance between [s, e] and UP(u) = [us, ue]
            return fold16(__vmaxu2(s, ue) - ue) + fold16(__vmaxu2(us, e) - e);
        };
        {   // z = a: v keeps b -- merge(w, b), then a, then UP(u)
            const uint32_t lo1 = __vmaxu2(ws, bs), hi1 = __vminu2(we, be);
            const uint32_t s1 = __vminu2(lo1, hi1), e1 = __vmaxu2(lo1, hi1);
            const uint32_t lo2 = __vmaxu2(s1, as), hi2 = __vminu2(e1, ae);
            const uint32_t s2 = __vminu2(lo2, hi2), e2 = __vmaxu2(lo2, hi2);
            cA += fold16(lo1 - s1) + fold16(lo2 - s2);
            if (HAS_UP) cA += gap(s2, e2);
        }
        {   // z = b: v keeps a
            const uint32_t lo1 = __vmaxu2(ws, as), hi1 = __vminu2(we, ae);
            const uint32_t s1 = __vminu2(lo1, hi1), e1 = __vmaxu2(lo1, hi1);
            const uint32_t lo2 = __vmaxu2(s1, bs), hi2 = __vminu2(e1, be);
            const uint32_t s2 = __vminu2(lo2, hi2), e2 = __vmaxu2(lo2, hi2);
            cB += fold16(lo1 - s1) + fold16(lo2 - s2);
            if (HAS_UP) cB += gap(s2, e2);
        }
        if (HAS_UP) c0 += gap(ds, de);  // the current tree re-rooted above u
    }
    pA = (int32_t)cA - (int32_t)c0;
    pB = (int32_t)cB - (int32_t)c0;
}

template <int CPT, int W>
__global__ void __launch_bounds__(256, 4)  // 64 registers: four CTAs per SM (three at 72-80 registers were 10 % slower, five spill)
k_nni_items(const lzg_pass_args A, int groups_i) {
    const int tid = threadIdx.x;
    const int it = blockIdx.x / groups_i, g = blockIdx.x % groups_i;
    const int4 i0 = __ldg(A.items + 2 * it), i1 = __ldg(A.items + 2 * it + 1);
    const int u = i0.x, w = i0.y, a = i0.w, b = i1.x, ia = i1.y, ib = i1.z;
    const size_t off = ((size_t)g * 256 + tid) * CPT;
    int32_t base = 0;
    if (tid == 0 && g == 0) base = A.score[w] + A.score[a] + A.score[b] + (u ? A.score[0] - A.score[u] : 0);
    if (tid == 0 && blockIdx.x == 0) A.ctl[LZG_CTL_ROOT] = A.score[0];
    int32_t pA = 0, pB = 0;
    if (off < (size_t)A.l_pad) {
        if (W == 1) {
            if (u) score_pair<CPT, true>(A, u, w, a, b, off, pA, pB);
            else score_pair<CPT, false>(A, u, w, a, b, off, pA, pB);
        } else {
            if (u) score_pair_w2<true>(A, u, w, a, b, off, pA, pB);
            else score_pair_w2<false>(A, u, w, a, b, off, pA, pB);
        }
    }
    pA = warp_sum_i(pA) + base;
    pB = warp_sum_i(pB) + base;
    if ((tid & 31) == 0) {
        if (ia >= 0 && pA) atomicAdd(A.cand_score + ia, pA);
        if (ib >= 0 && pB) atomicAdd(A.cand_score + ib, pB);
    }
}

template <bool CLIMB>
__device__ __forceinline__ void argmin_body(const lzg_pass_args& A) {
    __shared__ unsigned long long s_key[32];
    __shared__ int32_t s_last_score;
    const int tid = threadIdx.x;
    const int32_t root = __ldcg(A.ctl + LZG_CTL_ROOT);  // written by k_nni_items; summed over the ranks when positions are sharded
    unsigned long long best = ~0ull;
    // non-greedy: min (score, index); greedy: min (index) with the score riding in the low half
    auto key_of = [&](int32_t sc, int i) -> unsigned long long {
        return A.greedy ? (((unsigned long long)(uint32_t)i) << 32) | (uint32_t)sc : (((unsigned long long)(uint32_t)sc) << 32) | (uint32_t)i;
    };
    // cand_score is 64-byte aligned (d_cand + LZG_CTL_INTS): four candidates per load, all loads of a
    // thread independent
    const int n4 = A.n_cand >> 2;
    const int4* sc4 = reinterpret_cast<const int4*>(A.cand_score);
    for (int q = tid; q < n4; q += blockDim.x) {
        const int4 v = __ldcg(sc4 + q);
        const int32_t sc[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const int i = 4 * q + e;
            const unsigned long long key = key_of(sc[e], i);
            best = (sc[e] < root && key < best) ? key : best;
            if (i == A.n_cand - 1) s_last_score = sc[e];
        }
    }
    for (int i = 4 * n4 + tid; i < A.n_cand; i += blockDim.x) {
        const int32_t sc = __ldcg(A.cand_score + i);
        const unsigned long long key = key_of(sc, i);
        best = (sc < root && key < best) ? key : best;
        if (i == A.n_cand - 1) s_last_score = sc;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
        best = other < best ? other : best;
    }
    if ((tid & 31) == 0) s_key[tid >> 5] = best;
    __syncthreads();
    if (tid == 0) {
        for (int k = 1; k < (int)(blockDim.x >> 5); k++) best = s_key[k] < best ? s_key[k] : best;
        A.ctl[LZG_CTL_LAST] = s_last_score;  // see lzg_tree_last_candidate_score
        if (best == ~0ull) {
            A.ctl[LZG_CTL_BEST] = -1;
            A.ctl[LZG_CTL_SCORE] = root;
        } else {
            A.ctl[LZG_CTL_BEST] = A.greedy ? (int)(best >> 32) : (int)(best & 0xffffffffu);
            A.ctl[LZG_CTL_SCORE] = A.greedy ? (int)(best & 0xffffffffu) : (int)(best >> 32);
        }
        if (CLIMB) {
            const int pass = A.ctl[LZG_CTL_PASSES];
            if (pass == 0) A.ctl[LZG_CTL_CURRENT] = root;  // hill_climb's current_score = t[0].data.score (src/copy_number.cxx:390)
            A.ctl[LZG_CTL_PASSES] = pass + 1;
            A.ctl[LZG_CTL_EPOCH] += 1;  // the stamp of the next iteration's changes (lzg_tree_climb.cuh)
            if (pass <= A.log_cap) {
                unsigned long long now;
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
                A.tlog[pass] = now;
            }
            if (best == ~0ull) A.ctl[LZG_CTL_STOP] = 1;  // greedy_nni found no move: hill_climb breaks (:394)
            if (A.use_cond) cudaGraphSetConditional(A.cond_handle, best == ~0ull ? 0u : 1u);
        }
    }
}

template <bool CLIMB>
__global__ void __launch_bounds__(1024)
k_nni_argmin(const lzg_pass_args A_) {
    const lzg_pass_args& A = CLIMB ? c_pass : A_;
    if (CLIMB && __ldcg(A.ctl + LZG_CTL_STOP)) {
        if (A.use_cond && threadIdx.x == 0) cudaGraphSetConditional(A.cond_handle, 0);
        return;
    }
    argmin_body<CLIMB>(A);
}
