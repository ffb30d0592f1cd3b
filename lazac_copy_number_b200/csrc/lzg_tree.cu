// lzg_tree.cu -- ZCNT small parsimony on a resident tree (sm_100a).
//
// Per position the reference's sankoff()/overlap() (/root/reference/src/copy_number.cxx:116-158)
// merges child intervals [s1,e1], [s2,e2] into
//     lo = max(s1,s2), hi = min(e1,e2), start = min(lo,hi), end = max(lo,hi), cost = max(0, lo-hi)
// (checked branch by branch against the reference in tests/).  Positions are
// independent, values are biased bytes in HBM, so every thread owns 16 consecutive positions
// (one 128-bit load per interval vector), widens them to eight u16x2 words with PRMT and runs the
// merge with the native VIMNMX.U16x2 min/max (there is no native u8x4 min/max on sm_100a).
#include "lzg_tree.cuh"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace {

struct iv16 {  // 16 positions of an interval vector, widened to u16x2 lanes
    uint32_t s[8], e[8];
};

__device__ __forceinline__ void widen(const uint4& v, uint32_t* o) {
    o[0] = __byte_perm(v.x, 0, 0x4140);
    o[1] = __byte_perm(v.x, 0, 0x4342);
    o[2] = __byte_perm(v.y, 0, 0x4140);
    o[3] = __byte_perm(v.y, 0, 0x4342);
    o[4] = __byte_perm(v.z, 0, 0x4140);
    o[5] = __byte_perm(v.z, 0, 0x4342);
    o[6] = __byte_perm(v.w, 0, 0x4140);
    o[7] = __byte_perm(v.w, 0, 0x4342);
}

__device__ __forceinline__ uint4 narrow(const uint32_t* o) {
    uint4 v;
    v.x = __byte_perm(o[0], o[1], 0x6420);
    v.y = __byte_perm(o[2], o[3], 0x6420);
    v.z = __byte_perm(o[4], o[5], 0x6420);
    v.w = __byte_perm(o[6], o[7], 0x6420);
    return v;
}

__device__ __forceinline__ uint4 ldg16(const uint8_t* base, size_t off) {
    return __ldg(reinterpret_cast<const uint4*>(base + off));
}

// acc := merge(acc, other); returns the cost of these 16 positions.
__device__ __forceinline__ uint32_t merge16(iv16& acc, const uint32_t* os, const uint32_t* oe) {
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        uint32_t lo = __vmaxu2(acc.s[k], os[k]);
        uint32_t hi = __vminu2(acc.e[k], oe[k]);
        uint32_t st = __vminu2(lo, hi);
        acc.e[k] = __vmaxu2(lo, hi);
        acc.s[k] = st;
        c += lo - st;  // per-lane lo >= st: no borrow between the two 16-bit lanes
    }
    return (c & 0xffffu) + (c >> 16);  // each lane <= 8 * 254
}

// ---- the two device formats behind one set of kernels (template parameter W = bytes per position) ----------------
// A thread always owns one 16-byte chunk of a row: 16 positions widened to eight u16x2 words (W = 1), or 8 positions that
// already ARE four u16x2 words (W = 2, values biased by 32768).  Row offsets and pitches are in bytes either way.
template <int W>
__device__ __forceinline__ void widen_w(const uint4& v, uint32_t* o) {
    if (W == 1) {
        widen(v, o);
    } else {
        o[0] = v.x, o[1] = v.y, o[2] = v.z, o[3] = v.w;
    }
}
template <int W>
__device__ __forceinline__ uint4 narrow_w(const uint32_t* o) {
    if (W == 1) return narrow(o);
    return make_uint4(o[0], o[1], o[2], o[3]);
}
// acc := merge(acc, other); returns the cost of the chunk.  16-bit lanes hold differences up to 65535, so their costs are
// folded into 32 bits word by word; 8-bit data can add eight words per lane first (<= 8 * 254).
template <int W>
__device__ __forceinline__ uint32_t merge_w(iv16& acc, const uint32_t* os, const uint32_t* oe) {
    if (W == 1) return merge16(acc, os, oe);
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint32_t lo = __vmaxu2(acc.s[k], os[k]);
        uint32_t hi = __vminu2(acc.e[k], oe[k]);
        uint32_t st = __vminu2(lo, hi);
        acc.e[k] = __vmaxu2(lo, hi);
        acc.s[k] = st;
        const uint32_t d = lo - st;  // per-lane lo >= st: no borrow between the lanes
        c += (d & 0xffffu) + (d >> 16);
    }
    return c;
}

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- leaves ------------------------------------------------------------------------------------
__global__ void k_gather_leaves(const uint8_t* __restrict__ prof, int prof_lpad, const int32_t* __restrict__ leaf_row,
                                uint8_t* __restrict__ start, uint8_t* __restrict__ end, int l_pad, uint32_t pad_word) {
    int u = blockIdx.y;
    int row = leaf_row[u];
    int chunk = blockIdx.x * blockDim.x + threadIdx.x;
    if (chunk * 16 >= l_pad) return;
    uint4 v = row >= 0 ? ldg16(prof, (size_t)row * prof_lpad + chunk * 16)
                       : make_uint4(pad_word, pad_word, pad_word, pad_word);
    *reinterpret_cast<uint4*>(start + (size_t)u * l_pad + chunk * 16) = v;
    *reinterpret_cast<uint4*>(end + (size_t)u * l_pad + chunk * 16) = v;
}

// ---- bottom-up, one launch per level (small_rectilinear, src/copy_number.cxx:227-277) -------------
template <int W>
__global__ void k_merge_level(const int32_t* __restrict__ order, const int32_t* __restrict__ child0,
                              const int32_t* __restrict__ child1, uint8_t* __restrict__ start,
                              uint8_t* __restrict__ end, int32_t* __restrict__ cost, int l_pad) {
    int u = order[blockIdx.y];
    int c0 = child0[u], c1 = child1[u];
    int chunk = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t c = 0;
    if (chunk * 16 < l_pad) {
        size_t o0 = (size_t)c0 * l_pad + chunk * 16, o1 = (size_t)c1 * l_pad + chunk * 16;
        uint4 s0 = ldg16(start, o0), e0 = ldg16(end, o0), s1 = ldg16(start, o1), e1 = ldg16(end, o1);
        iv16 acc;
        uint32_t os[8], oe[8];
        widen_w<W>(s0, acc.s);
        widen_w<W>(e0, acc.e);
        widen_w<W>(s1, os);
        widen_w<W>(e1, oe);
        c = merge_w<W>(acc, os, oe);
        size_t ou = (size_t)u * l_pad + chunk * 16;
        *reinterpret_cast<uint4*>(start + ou) = narrow_w<W>(acc.s);
        *reinterpret_cast<uint4*>(end + ou) = narrow_w<W>(acc.e);
    }
    c = warp_sum(c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(cost + u, (int32_t)c);
}

// score[u] = cost[u] + score[c0] + score[c1], level by level (src/copy_number.cxx:264).
__global__ void k_level_scores(const int32_t* __restrict__ order, const int32_t* __restrict__ level_off, int n_levels,
                               const int32_t* __restrict__ child0, const int32_t* __restrict__ child1,
                               const int32_t* __restrict__ cost, int32_t* __restrict__ score) {
    for (int lv = 0; lv < n_levels; lv++) {
        int b = level_off[lv], e = level_off[lv + 1];
        for (int i = b + threadIdx.x; i < e; i += blockDim.x) {
            int u = order[i];
            int c0 = child0[u];
            score[u] = c0 < 0 ? 0 : cost[u] + score[c0] + score[child1[u]];
        }
        __syncthreads();
    }
}

}  // namespace

#include "lzg_tree_nni.cuh"

// launches a kernel template for the tree's device format
#define LZG_BY_WIDTH(width_, ...)      \
    do {                               \
        if ((width_) == 2) {           \
            constexpr int W = 2;       \
            __VA_ARGS__;               \
        } else {                       \
            constexpr int W = 1;       \
            __VA_ARGS__;               \
        }                              \
    } while (0)

static inline int chunk_threads(int l_pad, int* groups) {
    int chunks = l_pad / 16;
    int g = (chunks + 255) / 256;
    int per = (chunks + g - 1) / g;
    *groups = g;
    return ((per + 31) / 32) * 32;
}

int lzg_tree_upload_topology(lzg_ctx* ctx, lzg_tree* t) {
    const int n = t->n_nodes;
    std::vector<int2> up(n, make_int2(-1, -1));
    t->parent.assign(n, -1);
    for (int u = 0; u < n; u++) {
        int a = t->child0[u], b = t->child1[u];
        if (a >= 0) {
            t->parent[a] = u;
            up[a] = make_int2(u, b);
        }
        if (b >= 0) {
            t->parent[b] = u;
            up[b] = make_int2(u, a);
        }
    }
    LZG_CUDA(ctx, cudaMemcpyAsync(t->d_child0, t->child0.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice, ctx->stream));
    LZG_CUDA(ctx, cudaMemcpyAsync(t->d_child1, t->child1.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice, ctx->stream));
    LZG_CUDA(ctx, cudaMemcpyAsync(t->d_up, up.data(), sizeof(int2) * n, cudaMemcpyHostToDevice, ctx->stream));
    LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // sources are pageable host vectors
    return LZG_OK;
}

int lzg_tree_gather_leaves(lzg_ctx* ctx, lzg_tree* t, const lzg_profiles* p, const int32_t* leaf_row) {
    LZG_CUDA(ctx, cudaMemcpyAsync(t->d_order, leaf_row, sizeof(int32_t) * t->n_nodes, cudaMemcpyHostToDevice, ctx->stream));
    int groups, threads = chunk_threads(t->l_pad, &groups);
    k_gather_leaves<<<dim3(groups, t->n_nodes), threads, 0, ctx->stream>>>(p->d, p->pitch, t->d_order, t->d_start, t->d_end, t->l_pad, t->width == 2 ? 0x80008000u : 0x80808080u);
    LZG_LAUNCH_CHECK(ctx);
    LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LZG_OK;
}

// Level schedule by height above the leaves; one merge launch per level, deepest first.
int lzg_tree_full_score(lzg_ctx* ctx, lzg_tree* t) {
    const int n = t->n_nodes;
    std::vector<int32_t> height(n, 0), order;
    std::vector<int32_t> stack;
    // post-order without recursion
    std::vector<int32_t> post;
    post.reserve(n);
    stack.push_back(0);
    while (!stack.empty()) {
        int u = stack.back();
        stack.pop_back();
        post.push_back(u);
        if (t->child0[u] >= 0) {
            stack.push_back(t->child0[u]);
            stack.push_back(t->child1[u]);
        }
    }
    if ((int)post.size() != n) return lzg_fail(ctx, LZG_ERR_ARG, "tree is not connected from node 0");
    int max_h = 0;
    for (int i = n - 1; i >= 0; i--) {
        int u = post[i];
        if (t->child0[u] >= 0) height[u] = 1 + std::max(height[t->child0[u]], height[t->child1[u]]);
        max_h = std::max(max_h, height[u]);
    }
    std::vector<int32_t> level_off(max_h + 2, 0);
    for (int u = 0; u < n; u++) level_off[height[u] + 1]++;
    for (int h = 0; h <= max_h; h++) level_off[h + 1] += level_off[h];
    order.resize(n);
    {
        std::vector<int32_t> fill(level_off.begin(), level_off.end() - 1);
        for (int u = 0; u < n; u++) order[fill[height[u]]++] = u;
    }
    LZG_CUDA(ctx, lzg_ensure_scratch(ctx, level_off.size()) == LZG_OK ? cudaSuccess : cudaErrorMemoryAllocation);
    LZG_CUDA(ctx, cudaMemcpyAsync(t->d_order, order.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice, ctx->stream));
    LZG_CUDA(ctx, cudaMemcpyAsync(ctx->d_scratch, level_off.data(), sizeof(int32_t) * level_off.size(), cudaMemcpyHostToDevice, ctx->stream));
    LZG_CUDA(ctx, cudaMemsetAsync(t->d_cost, 0, sizeof(int32_t) * n, ctx->stream));
    int groups, threads = chunk_threads(t->l_pad, &groups);
    for (int h = 1; h <= max_h; h++) {
        int cnt = level_off[h + 1] - level_off[h];
        for (int done = 0; done < cnt; done += 65535) {  // gridDim.y limit
            int part = std::min(65535, cnt - done);
            LZG_BY_WIDTH(t->width, k_merge_level<W><<<dim3(groups, part), threads, 0, ctx->stream>>>(t->d_order + level_off[h] + done, t->d_child0, t->d_child1,
                                                                                                    t->d_start, t->d_end, t->d_cost, t->l_pad));
            LZG_LAUNCH_CHECK(ctx);
        }
    }
    k_level_scores<<<1, 1024, 0, ctx->stream>>>(t->d_order, ctx->d_scratch, max_h + 1, t->d_child0, t->d_child1, t->d_cost, t->d_score);
    LZG_LAUNCH_CHECK(ctx);
    LZG_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned, t->d_score, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
    LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    t->root_score = ctx->h_pinned[0];
    t->scored = true;
    t->up_valid = false;
    return LZG_OK;
}

// Packed host mirror for the heavy-path tables: {child0, child1, internal nodes in the subtree
// (0 for leaves), parent} per node -- one cache line touch per visited node.  Rebuilt from
// scratch after a topology replacement; an accepted NNI (u, w, v, z) rewrites four entries
// (lzg_tree_apply), the subtree count changes for v only.
// node4 record of an internal node with children c0, c1 (whose records are current):
// {heavy child, light child, (internal nodes in the subtree) << 2 | light child internal << 1 | heavy
// child internal, parent}; heavy = more internal nodes below, child0 on a tie.  A leaf is {-1, -1, 0, parent}.
static inline int4 make_node4(const int4* nd, int c0, int c1, int parent) {
    if (c0 < 0) return make_int4(-1, -1, 0, parent);
    const int z0 = nd[c0].z >> 2, z1 = nd[c1].z >> 2;
    const bool h0 = z0 >= z1;
    return make_int4(h0 ? c0 : c1, h0 ? c1 : c0, ((1 + z0 + z1) << 2) | (((h0 ? z1 : z0) > 0) << 1) | ((h0 ? z0 : z1) > 0), parent);
}

static void rebuild_node4(lzg_tree* t) {
    const int n = t->n_nodes;
    std::vector<int32_t>& pre = t->tmp_pre;
    t->node4.resize(n);
    pre.clear();
    pre.reserve(n);
    pre.push_back(0);
    for (size_t i = 0; i < pre.size(); i++) {  // BFS order: parents before children
        const int a = pre[i];
        if (t->child0[a] >= 0) {
            pre.push_back(t->child0[a]);
            pre.push_back(t->child1[a]);
        }
    }
    for (int i = n - 1; i >= 0; i--) {
        const int a = pre[i];
        t->node4[a] = make_node4(t->node4.data(), t->child0[a], t->child1[a], t->parent[a]);
    }
    t->node4_valid = true;
}

// Heavy-path decomposition of the internal non-root nodes, paths listed parents-first (a path's
// head hangs off a node of an earlier path or off the root).  order[] = {node, sibling} path
// after path, rec[] = {first index into order, length, parent of the head, sibling of the head}.
// Also fills t->up_rank (position of a node in order, -1 for the root and the leaves).
// One 16-byte record per node visited and no comparison: node4 lists the heavy child first and
// carries "is internal" bits for both (the walk is one dependent load per node; 65 us for 10^4
// internal nodes, half of what comparing the children's counts on the fly took).
static int build_up_paths(lzg_tree* t, int2* order, int4* rec, int* n_order) {
    const int n = t->n_nodes;
    if (!t->node4_valid) rebuild_node4(t);
    const int4* nd = t->node4.data();
    t->up_rank.assign(n, -1);
    int32_t* rank = t->up_rank.data();
    int cnt = 0, paths = 0;
    t->tmp_heads.resize(3 * (size_t)n / 2 + 12);  // (head, its parent, its sibling) triples, a queue (+ one speculative triple)
    int32_t* heads = t->tmp_heads.data();
    size_t n_heads = 0;
    // item_start[u]: first slot of the work items (u, v) of node u, in an array
    // sorted by the walker rank of u (the root's items first); an item per internal child, heavy first
    t->item_start.resize((size_t)n);
    int32_t* item_start = t->item_start.data();  // indexed by node id (the root's items start at 0)
    item_start[0] = 0;
    int slots = 0;
    {
        const int4 r = nd[0];
        if (r.z & 1) heads[n_heads++] = r.x, heads[n_heads++] = 0, heads[n_heads++] = r.y;
        if (r.z & 2) heads[n_heads++] = r.y, heads[n_heads++] = 0, heads[n_heads++] = r.x;
        slots = (r.z & 1) + ((r.z >> 1) & 1);
    }
    for (size_t h = 0; h < n_heads; h += 3) {
        const int first = cnt;
        int a = heads[h], sib = heads[h + 2];
        while (true) {
            const int4 na = nd[a];
            rank[a] = cnt;
            // the light child's triple is always written, the queue only grows if it is internal;
            // its path will wait for UP(a): that node publishes
            const int light_internal = (na.z >> 1) & 1;
            item_start[a] = slots;
            slots += (na.z & 1) + light_internal;
            heads[n_heads] = na.y, heads[n_heads + 1] = a, heads[n_heads + 2] = na.x;
            n_heads += 3 * (size_t)light_internal;
            order[cnt++] = make_int2(a, sib | (light_internal ? LZG_ORDER_PUBLISH : 0));
            if (!(na.z & 1)) break;
            a = na.x;
            sib = na.y;
        }
        rec[paths++] = make_int4(first, cnt - first, heads[h + 1], heads[h + 2]);
    }
    *n_order = cnt;
    return paths;
}

static int ensure_cand_buffers(lzg_ctx* ctx, lzg_tree* t, int n_moves) {
    if (n_moves <= t->cand_cap && t->d_pass) return LZG_OK;
    const int cap = std::max(n_moves, 2 * t->n_nodes);
    lzg_dev_free(ctx, t->d_pass);
    lzg_dev_free(ctx, t->d_cand);
    t->d_pass = nullptr;
    t->d_cand = nullptr;
    t->cand_cap = 0;
    // d_pass: [path records | order | items] packed per call; d_cand: [ctl (16) | cand_score (cap)]
    LZG_CUDA(ctx, lzg_dev_alloc(ctx, (void**)&t->d_pass, sizeof(int32_t) * ((size_t)6 * t->n_nodes + 12 + 8 * (size_t)cap)));
    LZG_CUDA(ctx, lzg_dev_alloc(ctx, (void**)&t->d_cand, sizeof(int32_t) * (LZG_CTL_INTS + (size_t)cap)));
    t->cand_cap = cap;
    // the item records are written on the copy stream: order it behind this allocation point
    LZG_CUDA(ctx, cudaEventRecord(ctx->ev_items, ctx->stream));
    return LZG_OK;
}

// UP intervals + the epoch flags of the table-driven UP pass, allocated on first use
static int ensure_up_buffers(lzg_ctx* ctx, lzg_tree* t, int groups) {
    if (t->d_ustart) return LZG_OK;
    const int n = t->n_nodes;
    const size_t nb = (size_t)n * t->l_pad;
    LZG_CUDA(ctx, lzg_dev_alloc(ctx, (void**)&t->d_ustart, nb));
    LZG_CUDA(ctx, lzg_dev_alloc(ctx, (void**)&t->d_uend, nb));
    const size_t n_flags = (size_t)n * groups;  // per (node, chunk group)
    LZG_CUDA(ctx, lzg_dev_alloc(ctx, (void**)&t->d_ready, sizeof(int) * n_flags));
    LZG_CUDA(ctx, cudaMemsetAsync(t->d_ready, 0, sizeof(int) * n_flags, ctx->stream));
    t->epoch = 0;
    t->up_valid = false;
    return LZG_OK;
}

// The host mirror's share of an accepted move (topology, parents, the packed node records).
static void host_apply_move(lzg_tree* t, int u, int w, int v, int z) {
    const int y = t->child0[v] == z ? t->child1[v] : t->child0[v];
    t->child0[v] = std::min(w, y);
    t->child1[v] = std::max(w, y);
    t->child0[u] = std::min(v, z);
    t->child1[u] = std::max(v, z);
    t->parent[w] = v;
    t->parent[z] = u;
    if (t->node4_valid) {  // u's subtree keeps its node set: the count changes for v only, the heavy / light roles for v and u
        int4* nd = t->node4.data();
        nd[w].w = v;
        nd[z].w = u;
        nd[v] = make_node4(nd, t->child0[v], t->child1[v], u);
        nd[u] = make_node4(nd, t->child0[u], t->child1[u], nd[u].w);
    }
}

static bool is_move(const lzg_tree* t, int u, int w, int v, int z) {
    const int n = t->n_nodes;
    if (u < 0 || u >= n || v < 0 || v >= n || w < 0 || w >= n || z < 0 || z >= n) return false;
    if (t->child0[v] < 0) return false;                       // v must be internal
    if (t->parent[v] != u || t->parent[w] != u || w == v) return false;
    if (t->parent[z] != v) return false;
    return true;
}

// Queues the UP pass of the current topology on the stream: heavy-path tables (host), one H2D
// copy, k_up_pass.  Called by lzg_tree_nni_scores when the UP intervals are stale and by
// lzg_tree_apply right behind the move, so that the GPU rebuilds them while the host
// enumerates and packs the next neighbourhood.
static int launch_up_pass(lzg_ctx* ctx, lzg_tree* t, int groups, int threads) {
    const int n = t->n_nodes;
    int rc = lzg_ensure_pinned(ctx, 16 + (size_t)6 * n + 12);
    if (rc != LZG_OK) return rc;
    int32_t* stage = ctx->h_pinned + 16;
    int n_order = 0;
    const bool trace = getenv("LZG_PASS_DEBUG") && (atoi(getenv("LZG_PASS_DEBUG")) & 4);
    const auto t0 = std::chrono::steady_clock::now();
    // lzg_tree_apply returns without waiting for the table copy it queued: a second move right
    // behind the first must not rewrite the staging area under it
    LZG_CUDA(ctx, cudaEventSynchronize(ctx->ev_stage));
    // records are written at the front, the order behind the worst-case record area, then compacted
    const int n_paths = build_up_paths(t, reinterpret_cast<int2*>(stage + 4 * (size_t)n), reinterpret_cast<int4*>(stage), &n_order);
    memmove(stage + 4 * (size_t)n_paths, stage + 4 * (size_t)n, sizeof(int2) * n_order);
    const auto t1 = std::chrono::steady_clock::now();
    t->up_nodes = n_order;
    t->up_paths = n_paths;
    t->epoch++;
    t->up_ran = true;
    if (n_paths > 0) {
        LZG_CUDA(ctx, cudaMemsetAsync(t->d_cand, 0, sizeof(int32_t) * LZG_CTL_INTS, ctx->stream));  // the ticket counter
        LZG_CUDA(ctx, cudaMemcpyAsync(t->d_pass, stage, sizeof(int32_t) * (4 * (size_t)n_paths + 2 * (size_t)n_order), cudaMemcpyHostToDevice,
                                      ctx->stream));
        LZG_CUDA(ctx, cudaEventRecord(ctx->ev_stage, ctx->stream));
        lzg_pass_args A{};
        A.path_rec = reinterpret_cast<const int4*>(t->d_pass);
        A.order = reinterpret_cast<const int2*>(t->d_pass + 4 * (size_t)n_paths);
        A.n_paths = n_paths;
        A.groups = groups;
        A.l_pad = t->l_pad;
        A.epoch = t->epoch;
        A.start = t->d_start;
        A.end = t->d_end;
        A.ustart = t->d_ustart;
        A.uend = t->d_uend;
        A.ready = t->d_ready;
        A.ctl = t->d_cand;
        LZG_BY_WIDTH(t->width, k_up_pass<W><<<n_paths * groups, threads, 0, ctx->stream>>>(A));
        LZG_LAUNCH_CHECK(ctx);
    }
    if (trace)
        fprintf(stderr, "[lzg up] paths %d nodes %d: tables %.1f us, queueing %.1f us\n", n_paths, n_order,
                std::chrono::duration<double, std::micro>(t1 - t0).count(),
                std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t1).count());
    t->up_valid = true;
    return LZG_OK;
}

static cudaEvent_t g_apply_tl[3] = {nullptr, nullptr, nullptr};  // LZG_PASS_DEBUG=16: before / behind k_apply_move, behind the queued UP pass

// One neighbourhood pass.  The UP pass is launched first (it needs the path tables only), so the
// host's validation and packing of the move list run while the GPU builds the UP intervals; then
// items, argmin and one 64-byte read-back.
//
// Two halves: lzg_tree_nni_launch queues everything up to the candidate scores (no read-back),
// lzg_tree_nni_finish the arg-min and the read-back.  lzg_nni_best runs them back to back; with the
// positions sharded over ranks the caller sum-reduces {root score, scores} between the two.
int lzg_tree_nni_launch(lzg_ctx* ctx, lzg_tree* t, int32_t n_moves, const int32_t* moves) {
    const auto t_enter = std::chrono::steady_clock::now();
    t->last_up_ran = false;
    t->last_cand_score = t->root_score;
    t->last_items = t->last_root_items = 0;
    t->last_v.resize(n_moves);
    t->pass_pending = false;
    if (n_moves == 0) {
        t->pass_pending = true;
        return LZG_OK;
    }
    const int n = t->n_nodes;
    int groups, threads = chunk_threads(t->l_pad, &groups);
    int rc = ensure_cand_buffers(ctx, t, n_moves);
    if (rc != LZG_OK) return rc;
    rc = ensure_up_buffers(ctx, t, groups);
    if (rc != LZG_OK) return rc;
    // staging (pinned; every public entry point ends with a synchronising read-back, so the area is
    // free again when the next call fills it): [result (16) | path records + order (6 n + 8) | items]
    rc = lzg_ensure_pinned(ctx, 16 + (size_t)6 * n + 12 + 8 * (size_t)n_moves);
    if (rc != LZG_OK) return rc;
    int32_t* stage = ctx->h_pinned + 16;
    const size_t o_items = ((size_t)6 * n + 8 + 3) / 4 * 4;  // 16-byte aligned, fixed offset of the items in the staging area and in d_pass
    const int dbg = getenv("LZG_PASS_DEBUG") ? atoi(getenv("LZG_PASS_DEBUG")) : 0;
    const bool trace = (dbg & 4) != 0;  // LZG_PASS_DEBUG=4: host preparation / kernel time of every pass on stderr
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (trace) {
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0, ctx->stream);
    }
    lzg_pass_args A{};
    A.n_cand = n_moves;
    A.groups = groups;
    A.l_pad = t->l_pad;
    A.start = t->d_start;
    A.end = t->d_end;
    A.ustart = t->d_ustart;
    A.uend = t->d_uend;
    A.score = t->d_score;
    A.ready = t->d_ready;
    A.ctl = t->d_cand;
    A.cand_score = t->d_cand + LZG_CTL_INTS;
    A.items = reinterpret_cast<const int4*>(t->d_pass + o_items);
    A.n_items = 0;
    int n_paths = 0;
    if (!t->up_valid) {
        rc = launch_up_pass(ctx, t, groups, threads);
        if (rc != LZG_OK) return rc;
        n_paths = t->up_paths;
    }
    t->last_up_ran = t->up_ran;  // also when lzg_tree_apply already queued it behind the move
    t->up_ran = false;
    LZG_CUDA(ctx, cudaMemsetAsync(t->d_cand, 0, sizeof(int32_t) * (LZG_CTL_INTS + (size_t)n_moves), ctx->stream));
    const auto t_up = std::chrono::steady_clock::now();
    // items: the two candidates of an edge (u, v) as one record, sorted by the walker rank of u so
    // that neighbouring CTAs share rows (both items under one u read DOWN(u) and UP(u))
    int n_items = 0;
    {
        if (!t->node4_valid) rebuild_node4(t);
        const int4* nd = t->node4.data();
        const unsigned un = (unsigned)n;
        int32_t* dst = stage + o_items;
        // One candidate pair -> one record.  Validity (is_move) and the children of v come from the packed
        // node mirror (current here: the UP pass of this topology was listed from it).
        // direct = true: the list is expected to be a whole neighbourhood (one item per internal
        // non-root v, what greedy_nni enumerates): every record goes straight to its slot in the
        // rank-of-u order (item_start, laid out with the path tables), no sort.  Returns 2 if the
        // list turns out not to be one (an item missing or listed twice): the caller then runs the
        // general form, a counting sort by the rank of u.
        // The whole-neighbourhood form: candidates 2k and 2k + 1 are the two moves of one edge, fixed
        // stride -- no iteration waits for the records the one before it loaded.
        auto pack_direct = [&]() -> int {
            std::vector<uint64_t>& seen = t->tmp_seen;  // one bit per item slot
            seen.assign(((size_t)t->up_nodes + 64) / 64, 0);
            const int32_t* item_start = t->item_start.data();
            const int pairs = n_moves / 2;
            int bad = 0, again = 0, roots = 0;
            for (int k = 0; k < pairs; k++) {
                const int32_t* m = moves + 8 * (size_t)k;
                const int u = m[0], w = m[1], v = m[2], z0 = m[3], z1 = m[7];
                if (!((unsigned)u < un && (unsigned)w < un && (unsigned)v < un && (unsigned)z0 < un && (unsigned)z1 < un)) return 1;
                if (m[4] != u || m[5] != w || m[6] != v) return 2;  // not a pair: the general form decides
                const int4 nv = nd[v], nu = nd[u];
                // v internal, {v, w} = the children of u, {z0, z1} = the children of v
                const bool vw = (nu.x == v && nu.y == w) || (nu.x == w && nu.y == v);
                const bool zz = (nv.x == z0 && nv.y == z1) || (nv.x == z1 && nv.y == z0);
                bad |= !(nv.x >= 0 && vw && v != w);
                again |= !zz;
                if (nu.x < 0) return 2;  // u is a leaf: no item slots of its own
                const int slot = item_start[u] + (nu.x == v ? 0 : (nu.z & 1));
                if ((unsigned)slot >= (unsigned)t->up_nodes) return 2;
                const uint64_t bit = 1ull << (slot & 63);
                again |= (seen[slot >> 6] & bit) != 0;
                seen[slot >> 6] |= bit;
                roots += u == 0;
                int32_t* r = dst + 8 * (size_t)slot;
                r[0] = u, r[1] = w, r[2] = v, r[3] = nv.x, r[4] = nv.y;
                r[5] = nv.x == z0 ? 2 * k : 2 * k + 1, r[6] = nv.x == z0 ? 2 * k + 1 : 2 * k, r[7] = 0;
            }
            if (bad || again) return 2;  // the general form tells a malformed list from an unusual one
            n_items = pairs;
            t->last_root_items = roots;
            t->last_whole = true;  // the v of the candidates = every internal non-root node, twice (lzg_tree_last_nni_work)
            return 0;
        };
        // The general form: any list of moves; adjacent candidates of one edge share a record; counting sort
        // by the walker rank of u.
        auto pack_general = [&]() -> int {
            std::vector<int32_t>& raw = t->tmp_items;
            std::vector<int32_t>& cnt = t->tmp_count;
            raw.resize(8 * (size_t)n_moves);
            cnt.assign((size_t)t->up_nodes + 2, 0);
            n_items = 0;
            t->last_root_items = 0;
            t->last_whole = false;
            for (int i = 0; i < n_moves;) {
                const int32_t* m = moves + 4 * (size_t)i;
                const int u = m[0], w = m[1], v = m[2], z = m[3];
                bool ok = (unsigned)u < un && (unsigned)w < un && (unsigned)v < un && (unsigned)z < un;
                int4 nv = make_int4(-1, -1, 0, -1);
                if (ok) {
                    nv = nd[v];
                    const int4 nu = nd[u];
                    // v internal, {v, w} = the children of u, z a child of v
                    ok = nv.x >= 0 && ((nu.x == v && nu.y == w) || (nu.x == w && nu.y == v)) && v != w && nd[z].w == v;
                }
                if (!ok) return 1;
                t->last_v[i] = v;
                const int a = nv.x, b = nv.y;  // the children of v
                int ia = z == a ? i : -1, ib = z == b ? i : -1, used = 1;
                if (i + 1 < n_moves && m[4] == u && m[5] == w && m[6] == v && m[7] != z && (m[7] == a || m[7] == b)) {
                    (m[7] == a ? ia : ib) = i + 1;
                    t->last_v[i + 1] = v;
                    used = 2;
                }
                int32_t* r = raw.data() + 8 * (size_t)n_items++;
                r[0] = u, r[1] = w, r[2] = v, r[3] = a, r[4] = b, r[5] = ia, r[6] = ib, r[7] = 0;
                cnt[t->up_rank[u] + 2]++;
                t->last_root_items += u == 0;
                i += used;
            }
            for (size_t k = 1; k < cnt.size(); k++) cnt[k] += cnt[k - 1];
            for (int k = 0; k < n_items; k++) {
                const int32_t* r = raw.data() + 8 * (size_t)k;
                memcpy(dst + 8 * (size_t)cnt[t->up_rank[r[0]] + 1]++, r, sizeof(int32_t) * 8);
            }
            return 0;
        };
        // a whole neighbourhood has two candidates per internal non-root node
        int st = n_moves == 2 * t->up_nodes ? pack_direct() : 2;
        const bool direct_ok = st == 0;
        if (st == 2) st = pack_general();
        if (st != 0) {
            cudaStreamSynchronize(ctx->stream);
            return lzg_fail(ctx, LZG_ERR_NOT_A_MOVE, "candidate is not an NNI move of this topology");
        }
        t->last_items = n_items;
        if (trace)
            fprintf(stderr, "[lzg pack] %d records in %.1f us (%s)\n", n_items, std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_up).count(),
                    direct_ok ? "placed directly" : "counting sort");
    }
    static cudaEvent_t tl[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // LZG_PASS_DEBUG=16: stream timeline of the pass, no host waits
    if (dbg & 16) {
        for (auto& e : tl)
            if (!e) cudaEventCreate(&e);
        cudaEventRecord(tl[0], ctx->stream);  // behind the UP pass (and whatever was queued before this call)
    }
    // the item records go up on the copy stream, beside the UP kernel (not behind it); the copy only has
    // to wait for the previous reader of the record area
    LZG_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_items, 0));
    LZG_CUDA(ctx, cudaMemcpyAsync(t->d_pass + o_items, stage + o_items, sizeof(int32_t) * 8 * (size_t)n_items, cudaMemcpyHostToDevice, ctx->copy_stream));
    LZG_CUDA(ctx, cudaEventRecord(ctx->ev_items, ctx->copy_stream));
    LZG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_items, 0));
    if (dbg & 16) cudaEventRecord(tl[1], ctx->stream);
    A.n_items = n_items;
    const auto t_launch = std::chrono::steady_clock::now();
    {
        const int gi = (t->l_pad + 256 * 16 - 1) / (256 * 16);
        LZG_BY_WIDTH(t->width, k_nni_items<16, W><<<n_items * gi, 256, 0, ctx->stream>>>(A, gi));
    }
    LZG_LAUNCH_CHECK(ctx);
    LZG_CUDA(ctx, cudaEventRecord(ctx->ev_items, ctx->stream));
    if (dbg & 16) {
        cudaEventRecord(tl[2], ctx->stream);
        cudaEventSynchronize(tl[2]);
        float a = 0, b = 0;
        cudaEventElapsedTime(&a, tl[0], tl[1]);
        cudaEventElapsedTime(&b, tl[1], tl[2]);
        fprintf(stderr, "[lzg timeline] host: %.1f us from entry to the items launch; stream: wait for the item records %.1f us, k_nni_items %.1f us\n",
                std::chrono::duration<double, std::micro>(t_launch - t_enter).count(), a * 1e3, b * 1e3);
        if (g_apply_tl[2] && t->last_up_ran && n_paths == 0) {  // the move before this pass queued the UP pass
            float c = 0, d = 0, e = 0;
            cudaEventElapsedTime(&c, g_apply_tl[0], g_apply_tl[1]);
            cudaEventElapsedTime(&d, g_apply_tl[1], g_apply_tl[2]);
            cudaEventElapsedTime(&e, g_apply_tl[2], tl[0]);
            fprintf(stderr, "[lzg timeline]   before it: k_apply_move + read-back %.1f us, then idle + tables H2D + k_up_pass %.1f us, then %.1f us to this pass\n",
                    c * 1e3, d * 1e3, e * 1e3);
        }
    }
    if (trace) {
        cudaEventRecord(e1, ctx->stream);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        fprintf(stderr, "[lzg pass] moves %d items %d paths %d: host %.1f us to the UP launch + %.1f us to the items launch, stream %.1f us\n", n_moves,
                n_items, n_paths, std::chrono::duration<double, std::micro>(t_up - t_enter).count(),
                std::chrono::duration<double, std::micro>(t_launch - t_up).count(), ms * 1e3);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    }
    t->pass_pending = true;
    return LZG_OK;
}

int lzg_tree_nni_finish(lzg_ctx* ctx, lzg_tree* t, int greedy, int32_t* best_index, int32_t* best_score, int32_t* all_scores) {
    if (!t->pass_pending) return lzg_fail(ctx, LZG_ERR_STATE, "no neighbourhood pass is pending on this tree");
    t->pass_pending = false;
    const int n_moves = (int)t->last_v.size();
    if (n_moves == 0) {
        *best_index = -1;
        *best_score = t->root_score;
        return LZG_OK;
    }
    lzg_pass_args A{};
    A.n_cand = n_moves;
    A.greedy = greedy;
    A.ctl = t->d_cand;
    A.cand_score = t->d_cand + LZG_CTL_INTS;
    k_nni_argmin<false><<<1, 1024, 0, ctx->stream>>>(A);
    LZG_LAUNCH_CHECK(ctx);
    LZG_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned, t->d_cand, sizeof(int32_t) * LZG_CTL_INTS, cudaMemcpyDeviceToHost, ctx->stream));
    if (all_scores)
        LZG_CUDA(ctx, cudaMemcpyAsync(all_scores, t->d_cand + LZG_CTL_INTS, sizeof(int32_t) * (size_t)n_moves, cudaMemcpyDeviceToHost, ctx->stream));
    LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *best_index = ctx->h_pinned[LZG_CTL_BEST];
    *best_score = ctx->h_pinned[LZG_CTL_SCORE];
    t->last_cand_score = ctx->h_pinned[LZG_CTL_LAST];
    return LZG_OK;
}

int32_t* lzg_tree_exchange_vector(lzg_tree* t) { return t->d_cand + LZG_CTL_ROOT; }

int lzg_tree_nni_scores(lzg_ctx* ctx, lzg_tree* t, int32_t n_moves, const int32_t* moves, int greedy,
                        int32_t* best_index, int32_t* best_score, int32_t* all_scores) {
    const int rc = lzg_tree_nni_launch(ctx, t, n_moves, moves);
    if (rc != LZG_OK) return rc;
    return lzg_tree_nni_finish(ctx, t, greedy, best_index, best_score, all_scores);
}

// Accept a move: one memset, one launch (k_apply_move), one read-back.
int lzg_tree_apply(lzg_ctx* ctx, lzg_tree* t, int u, int w, int v, int z) {
    if (!is_move(t, u, w, v, z)) return lzg_fail(ctx, LZG_ERR_NOT_A_MOVE, "not an NNI move of this topology");
    const int y = t->child0[v] == z ? t->child1[v] : t->child0[v];
    // the path in the OLD topology: v, u, parent(u), ..., root; step k writes node[k] after merging other[k]
    lzg_apply_args P;
    int depth = 0;
    for (int a = v; a != 0; a = t->parent[a]) depth++;
    P.n_steps = depth + 1;
    P.u = u, P.w = w, P.v = v, P.z = z, P.y = y;
    P.node[0] = v, P.other[0] = y;
    P.node[1] = u, P.other[1] = z;
    for (int k = 2, b = u; k < std::min(P.n_steps, LZG_APPLY_LISTED); k++) {
        const int pb = t->parent[b];
        P.node[k] = pb;
        P.other[k] = t->child0[pb] == b ? t->child1[pb] : t->child0[pb];
        b = pb;
    }
    host_apply_move(t, u, w, v, z);
    t->up_valid = false;
    t->d_node4_valid = false;  // k_apply_move does not maintain the device twin
    const bool timeline = getenv("LZG_PASS_DEBUG") && (atoi(getenv("LZG_PASS_DEBUG")) & 16);
    if (timeline) {
        for (auto& e : g_apply_tl)
            if (!e) cudaEventCreate(&e);
        cudaEventRecord(g_apply_tl[0], ctx->stream);
    }
    LZG_CUDA(ctx, cudaMemsetAsync(t->d_ctl, 0, sizeof(int32_t) * (LZG_CTL_INTS + (size_t)depth + 1), ctx->stream));
    int groups, threads = chunk_threads(t->l_pad, &groups);
    LZG_BY_WIDTH(t->width, k_apply_move<W><<<groups, threads, 0, ctx->stream>>>(t->d_child0, t->d_child1, t->d_up, t->d_start, t->d_end, t->d_score, t->d_cost,
                                                                               t->d_ctl, P, t->l_pad));
    LZG_LAUNCH_CHECK(ctx);
    LZG_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned, t->d_ctl, sizeof(int32_t) * LZG_CTL_INTS, cudaMemcpyDeviceToHost, ctx->stream));
    LZG_CUDA(ctx, cudaEventRecord(ctx->ev_group, ctx->stream));
    if (timeline) cudaEventRecord(g_apply_tl[1], ctx->stream);
    // a neighbourhood pass follows an accepted move (hill_climb, src/copy_number.cxx:393-409): queue its
    // UP pass now, the caller only waits for the root score
    if (t->d_ustart && t->d_pass && !(getenv("LZG_EAGER_UP") && atoi(getenv("LZG_EAGER_UP")) == 0)) {
        int rc = launch_up_pass(ctx, t, groups, threads);
        if (rc != LZG_OK) return rc;
    }
    if (timeline) cudaEventRecord(g_apply_tl[2], ctx->stream);
    LZG_CUDA(ctx, cudaEventSynchronize(ctx->ev_group));
    t->root_score = ctx->h_pinned[LZG_CTL_SCORE];
    return LZG_OK;
}

// Rows of the device format (bytes v + 128, or 16-bit words v + 32768) to the caller's int32 matrix: the narrow values cross
// PCIe (a quarter or half of the int32 bytes) into two halves of the pinned staging area in turn -- the copy of block k + 1 runs
// while a few host threads widen block k -- instead of a device-side widening, a pageable int32 copy and a sync per block
// (98 ms for cfg4's 39 MB of labels in round 1).
static void widen_rows(const uint8_t* src, size_t src_pitch, int32_t* dst, int rows, int L, int width) {
    for (int r = 0; r < rows; r++) {
        const uint8_t* s = src + (size_t)r * src_pitch;
        int32_t* d = dst + (size_t)r * L;
        if (width == 1)
            for (int j = 0; j < L; j++) d[j] = (int32_t)s[j] - LZG_BIAS;
        else {
            const uint16_t* s16 = reinterpret_cast<const uint16_t*>(s);
            for (int j = 0; j < L; j++) d[j] = (int32_t)s16[j] - LZG_BIAS16;
        }
    }
}

static int download_bytes_as_i32(lzg_ctx* ctx, const uint8_t* d_src, int rows, int L, int l_pad, int32_t* h_dst, int width) {
    if (rows <= 0 || L <= 0) return LZG_OK;
    const size_t row_bytes = ((size_t)L * width + 15) / 16 * 16;  // staged pitch
    const size_t half = std::max<size_t>((size_t)8 << 20, row_bytes);
    int rc = lzg_ensure_pinned(ctx, 16 + 2 * half / sizeof(int32_t));
    if (rc != LZG_OK) return rc;
    uint8_t* stage[2] = {reinterpret_cast<uint8_t*>(ctx->h_pinned + 16), reinterpret_cast<uint8_t*>(ctx->h_pinned + 16) + half};
    const int block_rows = (int)std::max<size_t>(1, std::min<size_t>(rows, half / row_bytes));
    const int n_blocks = (rows + block_rows - 1) / block_rows;
    auto issue = [&](int k) {
        const int r0 = k * block_rows, nr = std::min(block_rows, rows - r0);
        return cudaMemcpy2DAsync(stage[k & 1], row_bytes, d_src + (size_t)r0 * l_pad, l_pad, (size_t)L * width, nr, cudaMemcpyDeviceToHost, ctx->stream);
    };
    cudaError_t e = issue(0);
    const int max_threads = (int)std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
    for (int k = 0; k < n_blocks && e == cudaSuccess; k++) {
        e = cudaStreamSynchronize(ctx->stream);  // block k is in its half
        if (e == cudaSuccess && k + 1 < n_blocks) e = issue(k + 1);
        if (e != cudaSuccess) break;
        const int r0 = k * block_rows, nr = std::min(block_rows, rows - r0);
        const int n_thr = (size_t)nr * L < ((size_t)1 << 18) ? 1 : std::min(max_threads, nr);
        if (n_thr == 1) {
            widen_rows(stage[k & 1], row_bytes, h_dst + (size_t)r0 * L, nr, L, width);
        } else {
            std::vector<std::thread> pool;
            for (int w = 0; w < n_thr; w++) {
                const int a = (int)((long long)nr * w / n_thr), b = (int)((long long)nr * (w + 1) / n_thr);
                pool.emplace_back(widen_rows, stage[k & 1] + (size_t)a * row_bytes, row_bytes, h_dst + (size_t)(r0 + a) * L, b - a, L, width);
            }
            for (auto& th : pool) th.join();
        }
    }
    if (e != cudaSuccess) {
        cudaStreamSynchronize(ctx->stream);
        return lzg_fail(ctx, LZG_ERR_CUDA, "interval download", e);
    }
    return LZG_OK;
}

int lzg_tree_download(lzg_ctx* ctx, const lzg_tree* t, int32_t* start, int32_t* end, int32_t* score) {
    int rc = LZG_OK;
    if (start) rc = download_bytes_as_i32(ctx, t->d_start, t->n_nodes, t->L, t->l_pad, start, t->width);
    if (rc == LZG_OK && end) rc = download_bytes_as_i32(ctx, t->d_end, t->n_nodes, t->L, t->l_pad, end, t->width);
    if (rc == LZG_OK && score) {
        LZG_CUDA(ctx, cudaMemcpyAsync(score, t->d_score, sizeof(int32_t) * t->n_nodes, cudaMemcpyDeviceToHost, ctx->stream));
        LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return rc;
}

// start/end rows of the listed nodes only ([n_ids][L] each): trees beyond a few GB are inspected node by node.
int lzg_tree_download_rows(lzg_ctx* ctx, const lzg_tree* t, int n_ids, const int32_t* ids, int32_t* start, int32_t* end) {
    for (int k = 0; k < n_ids; k++) {
        if (ids[k] < 0 || ids[k] >= t->n_nodes) return lzg_fail(ctx, LZG_ERR_ARG, "node id out of range");
        int rc = LZG_OK;
        if (start) rc = download_bytes_as_i32(ctx, t->d_start + (size_t)ids[k] * t->l_pad, 1, t->L, t->l_pad, start + (size_t)k * t->L, t->width);
        if (rc == LZG_OK && end) rc = download_bytes_as_i32(ctx, t->d_end + (size_t)ids[k] * t->l_pad, 1, t->L, t->l_pad, end + (size_t)k * t->L, t->width);
        if (rc != LZG_OK) return rc;
    }
    return LZG_OK;
}

// Depth-level schedule, root first.
int lzg_tree_label(lzg_ctx* ctx, const lzg_tree* t, int32_t* labels, int32_t* branch) {
    const int n = t->n_nodes;
    std::vector<int32_t> order, level_off;
    order.reserve(n);
    order.push_back(0);
    level_off.push_back(0);
    size_t b = 0;
    while (b < order.size()) {
        size_t e = order.size();
        level_off.push_back((int32_t)e);
        for (size_t i = b; i < e; i++) {
            int u = order[i];
            if (t->child0[u] >= 0) {
                order.push_back(t->child0[u]);
                order.push_back(t->child1[u]);
            }
        }
        b = e;
    }
    uint8_t* d_label = nullptr;
    int32_t* d_branch = nullptr;
    LZG_CUDA(ctx, lzg_dev_alloc(ctx, (void**)&d_label, (size_t)n * t->l_pad));
    cudaError_t e = lzg_dev_alloc(ctx, (void**)&d_branch, sizeof(int32_t) * n);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_branch, 0, sizeof(int32_t) * n, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(t->d_order, order.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice, ctx->stream);
    int rc = LZG_OK;
    if (e != cudaSuccess) rc = lzg_fail(ctx, LZG_ERR_CUDA, "labeling setup", e);
    int groups, threads = chunk_threads(t->l_pad, &groups);
    for (size_t lv = 0; rc == LZG_OK && lv + 1 < level_off.size(); lv++) {
        int cnt = level_off[lv + 1] - level_off[lv];
        for (int done = 0; done < cnt; done += 65535) {
            int part = std::min(65535, cnt - done);
            LZG_BY_WIDTH(t->width, k_label_level<W><<<dim3(groups, part), threads, 0, ctx->stream>>>(t->d_order + level_off[lv] + done, t->d_up, t->d_start, t->d_end,
                                                                                                    d_label, d_branch, t->l_pad));
            ctx->launches++;
        }
    }
    if (rc == LZG_OK) {
        e = cudaGetLastError();
        if (e != cudaSuccess) rc = lzg_fail(ctx, LZG_ERR_CUDA, "labeling launch", e);
    }
    if (rc == LZG_OK && labels) rc = download_bytes_as_i32(ctx, d_label, n, t->L, t->l_pad, labels, t->width);
    if (rc == LZG_OK && branch) {
        e = cudaMemcpyAsync(branch, d_branch, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = lzg_fail(ctx, LZG_ERR_CUDA, "branch download", e);
        else branch[0] = -1;
    }
    cudaStreamSynchronize(ctx->stream);
    lzg_dev_free(ctx, d_label);
    lzg_dev_free(ctx, d_branch);
    return rc;
}

// ---- hill_climb on the device (kernels: lzg_tree_climb.cuh) ----------------------------------------------
// Host side of lzg_hill_climb: lays the static work items out once per climb, uploads them with the
// candidate numbering of the caller's shuffled edge order, launches the climb (one WHILE graph, or
// batches of queued iterations where a collective sits inside the iteration), reads back one move
// log and replays it on the host mirror.
namespace {

constexpr int LZG_CLIMB_LOG = 4096;    // accepted moves logged on the device before the host drains the log
constexpr int LZG_UQ_SLACK = 4096;     // queue entries beyond one per (node, chunk group): tickets drawn by idle CTAs

struct climb_layout {  // offsets (bytes) into t->d_climb
    size_t items, slot_of, cbase, pair_v, upload_bytes, pcost, tail, xctl, dn_stamp, up_stamp, rec_stamp, part, uq_flag, wl, deep_chg, zero_end,
        mlog, uq, deep_path, queue, tlog, total;
    int ucap;
};

climb_layout climb_offsets(int n_nodes, int n_items, int groups) {
    climb_layout o;
    auto up16 = [](size_t x) { return (x + 15) / 16 * 16; };
    const size_t ni = (size_t)std::max(1, n_items), G = (size_t)groups, nn = (size_t)n_nodes;
    o.ucap = (int)std::min<size_t>(ni * G + LZG_UQ_SLACK, (size_t)1 << 30);
    size_t p = 0;
    o.items = p, p = up16(p + sizeof(int4) * 2 * ni);
    o.slot_of = p, p = up16(p + sizeof(int32_t) * nn);
    o.cbase = p, p = up16(p + sizeof(int32_t) * nn);
    o.pair_v = p, p = up16(p + sizeof(int32_t) * ni);
    o.upload_bytes = p;  // everything above is rewritten by the host at the start of a climb
    // zero at allocation and kept consistent by the kernels
    o.pcost = p, p = up16(p + sizeof(int32_t) * nn);
    o.tail = p, p = up16(p + sizeof(int32_t) * std::max<size_t>(4, G));
    o.xctl = p, p = up16(p + sizeof(int32_t) * LZG_X_INTS);
    o.dn_stamp = p, p = up16(p + sizeof(int32_t) * nn * G);
    o.up_stamp = p, p = up16(p + sizeof(int32_t) * nn * G);
    o.rec_stamp = p, p = up16(p + sizeof(int32_t) * ni);
    o.part = p, p = up16(p + sizeof(int32_t) * 2 * ni * G);
    o.uq_flag = p, p = up16(p + sizeof(int32_t) * (size_t)o.ucap);
    o.wl = p, p = up16(p + sizeof(int32_t) * ni * G);
    o.deep_chg = p, p = up16(p + sizeof(int32_t) * nn * G);
    o.queue = p, p = up16(p + sizeof(unsigned long long) * G * (ni + 1));
    o.zero_end = p;
    o.mlog = p, p = up16(p + sizeof(int32_t) * 5 * (size_t)LZG_CLIMB_LOG);
    o.uq = p, p = up16(p + sizeof(int4) * (size_t)o.ucap);
    o.deep_path = p, p = up16(p + sizeof(int2) * nn * G);
    o.tlog = p, p = up16(p + sizeof(unsigned long long) * ((size_t)LZG_CLIMB_LOG + 2));
    o.total = p;
    return o;
}

int ensure_climb_buffers(lzg_ctx* ctx, lzg_tree* t, int n_items, int groups) {
    const climb_layout o = climb_offsets(t->n_nodes, n_items, groups);
    if (t->d_climb && t->climb_items == n_items && t->climb_bytes == o.total) return LZG_OK;
    lzg_dev_free(ctx, t->d_climb);
    t->d_climb = nullptr;
    LZG_CUDA(ctx, lzg_dev_alloc(ctx, &t->d_climb, o.total));
    // cost cells, counters, stamps, queue flags start at zero (stamps and epochs start at 1); the upload area is written per climb
    LZG_CUDA(ctx, cudaMemsetAsync((uint8_t*)t->d_climb + o.pcost, 0, o.zero_end - o.pcost, ctx->stream));
    t->climb_bytes = o.total;
    t->climb_items = n_items;
    t->df_epoch = 1;
    if (!t->d_node4) {
        LZG_CUDA(ctx, lzg_dev_alloc(ctx, (void**)&t->d_node4, sizeof(int4) * (size_t)t->n_nodes));
        t->d_node4_valid = false;
    }
    return LZG_OK;
}

struct climb_geometry {
    int n_items, n_cand, groups, threads, up_grid, delta_grid, wl_grid, scan_grid, item_grid, width;
};

enum { K_UP_FULL = 0, K_DIRTY, K_ITEMS, K_ASSEMBLE, K_ARGMIN, K_APPLY, K_UP_DELTA };

struct climb_launch {
    void* fn;
    dim3 grid, block;
    bool has_arg;  // k_nni_argmin takes a (dummy) lzg_pass_args by value; the others take nothing
};

climb_launch climb_kernel(int which, const climb_geometry& G) {
    switch (which) {
        case K_UP_FULL: return {G.width == 2 ? (void*)k_up_df<2> : (void*)k_up_df<1>, dim3(G.up_grid), dim3(G.threads), false};
        case K_DIRTY: return {(void*)k_dirty_scan, dim3(G.scan_grid), dim3(256), false};
        case K_ITEMS: return {G.width == 2 ? (void*)k_items_wl<2> : (void*)k_items_wl<1>, dim3(G.wl_grid), dim3(G.threads), false};
        case K_ASSEMBLE: return {(void*)k_cand_assemble, dim3(G.item_grid), dim3(256), false};
        case K_ARGMIN: return {(void*)k_nni_argmin<true>, dim3(1), dim3(1024), true};
        case K_APPLY: return {G.width == 2 ? (void*)k_apply2<2> : (void*)k_apply2<1>, dim3(G.groups), dim3(G.threads), false};
        default: return {G.width == 2 ? (void*)k_up_delta<2> : (void*)k_up_delta<1>, dim3(G.delta_grid), dim3(G.threads), false};
    }
}

const int PASS0[] = {K_UP_FULL, K_DIRTY, K_ITEMS, K_ASSEMBLE, K_ARGMIN};
const int ITERATION[] = {K_APPLY, K_UP_DELTA, K_DIRTY, K_ITEMS, K_ASSEMBLE, K_ARGMIN};

// One pass as plain stream launches; the all-reduce of the position-sharded climb sits between the
// assembled scores and the arg-min.
int climb_enqueue(lzg_ctx* ctx, const climb_geometry& G, const int* seq, int n_seq, lzg_comm* comm, lzg_tree* t, cudaEvent_t* marks = nullptr) {
    static lzg_pass_args dummy{};
    for (int q = 0; q < n_seq; q++) {
        if (marks) cudaEventRecord(marks[q], ctx->stream);
        if (seq[q] == K_ARGMIN && comm && lzg_comm_world(comm) > 1) {
            const int rc = lzg_comm_allreduce_sum(ctx, comm, t->d_cand + LZG_CTL_ROOT, (size_t)G.n_cand + 1);
            if (rc != LZG_OK) return rc;
        }
        const climb_launch L = climb_kernel(seq[q], G);
        void* args[] = {(void*)&dummy};
        LZG_CUDA(ctx, cudaLaunchKernel(L.fn, L.grid, L.block, L.has_arg ? args : nullptr, 0, ctx->stream));
        ctx->launches++;
    }
    if (marks) cudaEventRecord(marks[n_seq], ctx->stream);
    return LZG_OK;
}

// The climb as one graph: pass 0, then a WHILE node whose body is one iteration; the arg-min kernel sets
// the loop condition.  The kernels read c_pass, so one graph serves every tree of this launch geometry.
lzg_ctx::climb_graph* climb_graph_for(lzg_ctx* ctx, const climb_geometry& G) {
    for (auto& g : ctx->climb_graphs)
        if (g.n_items == G.n_items && g.groups == G.groups && g.threads == G.threads && g.up_grid == G.up_grid && g.n_cand == G.n_cand && g.width == G.width) return &g;
    if (ctx->climb_graph_broken) return nullptr;
    lzg_ctx::climb_graph cg;
    cg.n_items = G.n_items, cg.groups = G.groups, cg.threads = G.threads, cg.up_grid = G.up_grid, cg.n_cand = G.n_cand, cg.width = G.width;
    static lzg_pass_args dummy{};
    auto add_chain = [&](cudaGraph_t graph, const int* seq, int n_seq, cudaGraphNode_t* last) -> cudaError_t {
        cudaGraphNode_t prev = nullptr;
        for (int q = 0; q < n_seq; q++) {
            const climb_launch L = climb_kernel(seq[q], G);
            cudaKernelNodeParams kp{};
            void* args[] = {(void*)&dummy};
            kp.func = L.fn;
            kp.gridDim = L.grid;
            kp.blockDim = L.block;
            kp.sharedMemBytes = 0;
            kp.kernelParams = L.has_arg ? args : nullptr;
            kp.extra = nullptr;
            cudaGraphNode_t node = nullptr;
            const cudaError_t e = cudaGraphAddKernelNode(&node, graph, prev ? &prev : nullptr, prev ? 1 : 0, &kp);
            if (e != cudaSuccess) return e;
            prev = node;
        }
        *last = prev;
        return cudaSuccess;
    };
    cudaError_t e = cudaGraphCreate(&cg.graph, 0);
    cudaGraphNode_t tail0 = nullptr, nw = nullptr;
    if (e == cudaSuccess) e = cudaGraphConditionalHandleCreate(&cg.handle, cg.graph, 1, cudaGraphCondAssignDefault);
    if (e == cudaSuccess) e = add_chain(cg.graph, PASS0, 5, &tail0);
    if (e == cudaSuccess) {
        cudaGraphNodeParams cp{};
        cp.type = cudaGraphNodeTypeConditional;
        cp.conditional.handle = cg.handle;
        cp.conditional.type = cudaGraphCondTypeWhile;
        cp.conditional.size = 1;
        e = cudaGraphAddNode(&nw, cg.graph, &tail0, 1, &cp);
        if (e == cudaSuccess) {
            cudaGraphNode_t tail1 = nullptr;
            e = add_chain(cp.conditional.phGraph_out[0], ITERATION, 6, &tail1);
        }
    }
    if (e == cudaSuccess) e = cudaGraphInstantiate(&cg.exec, cg.graph, 0);
    if (e != cudaSuccess) {
        cudaGetLastError();
        if (cg.graph) cudaGraphDestroy(cg.graph);
        ctx->climb_graph_broken = 1;
        if (getenv("LZG_CLIMB_DEBUG")) fprintf(stderr, "[lzg climb] WHILE graph unavailable (%s): queued batches instead\n", cudaGetErrorString(e));
        return nullptr;
    }
    ctx->climb_graphs.push_back(cg);
    return &ctx->climb_graphs.back();
}

}  // namespace

int lzg_tree_climb(lzg_ctx* ctx, lzg_comm* comm, lzg_tree* t, int32_t n_edges, const int32_t* edge_order, int32_t greedy, int32_t max_moves,
                   int32_t moves_cap, int32_t* moves, lzg_climb_result* out) {
    const int n = t->n_nodes;
    const bool sharded = comm && lzg_comm_world(comm) > 1;
    lzg_climb_result R{};
    R.true_score = R.returned_score = t->root_score;
    R.reason = 1;
    t->climb_stamps.clear();
    t->pass0_timed = false;
    if (n_edges != n - 1) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_hill_climb: n_edges must be n_nodes - 1");
    // child end of every edge index (t.edges(): by parent id, then child id) and the candidate numbering
    std::vector<int32_t>& child_of = t->tmp_pre;
    child_of.clear();
    for (int u = 0; u < n; u++)
        if (t->child0[u] >= 0) child_of.push_back(t->child0[u]), child_of.push_back(t->child1[u]);
    if ((int)child_of.size() != n_edges) return lzg_fail(ctx, LZG_ERR_STATE, "lzg_hill_climb: topology mirror is inconsistent");
    std::vector<int32_t> cbase(n, -1), pair_v;
    {
        std::vector<char> seen(n_edges, 0);
        for (int q = 0; q < n_edges; q++) {
            const int e = edge_order[q];
            if (e < 0 || e >= n_edges || seen[e]) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_hill_climb: edge_order is not a permutation");
            seen[e] = 1;
            const int v = child_of[e];
            if (t->child0[v] < 0) continue;  // greedy_nni skips edges into leaves (:334)
            cbase[v] = 2 * (int)pair_v.size();
            pair_v.push_back(v);
        }
    }
    const int n_items = (int)pair_v.size(), n_cand = 2 * n_items;
    R.n_passes = 0;
    if (n_items == 0 || max_moves <= 0) {  // nothing to enumerate: greedy_nni returns no move, the tree is untouched
        R.n_passes = n_items == 0 ? 1 : 0;
        R.reason = n_items == 0 ? 1 : 3;
        if (out) *out = R;
        return LZG_OK;
    }
    int groups, threads = chunk_threads(t->l_pad, &groups);
    int rc = ensure_cand_buffers(ctx, t, n_cand);
    if (rc == LZG_OK) rc = ensure_up_buffers(ctx, t, groups);
    if (rc == LZG_OK) rc = ensure_climb_buffers(ctx, t, n_items, groups);
    if (rc != LZG_OK) return rc;
    const climb_layout o = climb_offsets(n, n_items, groups);
    // ---- layout: item slots in the walker order of the starting topology (neighbours share rows in L2) ----
    rc = lzg_ensure_pinned(ctx, 16 + (size_t)6 * n + 12 + o.upload_bytes / 4 + 64 + 5 * (size_t)LZG_CLIMB_LOG);
    if (rc != LZG_OK) return rc;
    LZG_CUDA(ctx, cudaEventSynchronize(ctx->ev_stage));
    int32_t* scratch = ctx->h_pinned + 16;                                   // path tables of build_up_paths (not uploaded here)
    uint8_t* blob = reinterpret_cast<uint8_t*>(ctx->h_pinned + 16 + (size_t)6 * n + 12 + 4);
    blob = reinterpret_cast<uint8_t*>(((uintptr_t)blob + 15) / 16 * 16);
    int n_order = 0;
    const int n_paths = build_up_paths(t, reinterpret_cast<int2*>(scratch + 4 * (size_t)n), reinterpret_cast<int4*>(scratch), &n_order);
    if (n_order != n_items) return lzg_fail(ctx, LZG_ERR_STATE, "lzg_hill_climb: internal node count mismatch");
    {
        const int4* nd = t->node4.data();
        int32_t* slot_of = reinterpret_cast<int32_t*>(blob + o.slot_of);
        int32_t* h_cbase = reinterpret_cast<int32_t*>(blob + o.cbase);
        int32_t* h_pair = reinterpret_cast<int32_t*>(blob + o.pair_v);
        int4* items = reinterpret_cast<int4*>(blob + o.items);
        memcpy(h_cbase, cbase.data(), sizeof(int32_t) * n);
        memcpy(h_pair, pair_v.data(), sizeof(int32_t) * n_items);
        for (int x = 0; x < n; x++) slot_of[x] = -1;
        for (int k = 0; k < n_items; k++) {
            const int v = pair_v[k], u = nd[v].w;
            const int4 nu = nd[u];
            const int slot = t->item_start[u] + (nu.x == v ? 0 : (nu.z & 1));
            if (slot < 0 || slot >= n_items) return lzg_fail(ctx, LZG_ERR_STATE, "lzg_hill_climb: item slot out of range");
            slot_of[v] = slot;
            const int w = nu.x == v ? nu.y : nu.x;
            items[2 * slot] = make_int4(u, w, v, t->child0[v]);
            items[2 * slot + 1] = make_int4(t->child1[v], cbase[v], cbase[v] + 1, 0);
        }
    }
    uint8_t* dc = (uint8_t*)t->d_climb;
    LZG_CUDA(ctx, cudaMemcpyAsync(dc, blob, o.upload_bytes, cudaMemcpyHostToDevice, ctx->stream));
    if (!t->d_node4_valid) {
        LZG_CUDA(ctx, cudaMemcpyAsync(t->d_node4, t->node4.data(), sizeof(int4) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
        t->d_node4_valid = true;
    }
    // ---- the descriptor (constant memory) and the control block ----
    climb_geometry G;
    G.n_items = n_items, G.n_cand = n_cand, G.groups = groups, G.threads = threads, G.width = t->width;
    const int per_sm = std::max(1, 1024 / threads);
    // persistent grids; functions of the data set only (one graph per data set)
    G.up_grid = std::max(1, std::min(n_items * groups, ctx->sm_count * per_sm));
    G.delta_grid = std::max(1, std::min(n_items * groups, ctx->sm_count * 2));
    G.wl_grid = std::max(1, std::min(n_items * groups, ctx->sm_count * per_sm));
    G.scan_grid = (int)(((long long)n_items * groups + 255) / 256);
    G.item_grid = (n_items + 255) / 256;
    const bool want_graph = !sharded && !(getenv("LZG_CLIMB_GRAPH") && atoi(getenv("LZG_CLIMB_GRAPH")) == 0);
    lzg_ctx::climb_graph* cg = want_graph ? climb_graph_for(ctx, G) : nullptr;
    lzg_pass_args A{};
    A.n_items = n_items, A.n_cand = n_cand, A.groups = groups, A.l_pad = t->l_pad, A.greedy = greedy;
    A.n_nodes = n, A.threads = threads, A.ucap = o.ucap;
    A.start = t->d_start, A.end = t->d_end, A.wstart = t->d_start, A.wend = t->d_end;
    A.ustart = t->d_ustart, A.uend = t->d_uend;
    A.score = t->d_score, A.wscore = t->d_score;
    A.ctl = t->d_cand, A.cand_score = t->d_cand + LZG_CTL_INTS;
    A.node4 = t->d_node4, A.child0 = t->d_child0, A.child1 = t->d_child1, A.up = t->d_up;
    A.items = reinterpret_cast<const int4*>(dc + o.items), A.witems = reinterpret_cast<int4*>(dc + o.items);
    A.slot_of = reinterpret_cast<const int32_t*>(dc + o.slot_of);
    A.cbase = reinterpret_cast<const int32_t*>(dc + o.cbase);
    A.pair_v = reinterpret_cast<const int32_t*>(dc + o.pair_v);
    A.pcost = reinterpret_cast<int32_t*>(dc + o.pcost);
    A.tail = reinterpret_cast<int*>(dc + o.tail);
    A.xctl = reinterpret_cast<int*>(dc + o.xctl);
    A.dn_stamp = reinterpret_cast<int32_t*>(dc + o.dn_stamp);
    A.up_stamp = reinterpret_cast<int32_t*>(dc + o.up_stamp);
    A.rec_stamp = reinterpret_cast<int32_t*>(dc + o.rec_stamp);
    A.part = reinterpret_cast<int32_t*>(dc + o.part);
    A.uq = reinterpret_cast<int4*>(dc + o.uq);
    A.uq_flag = reinterpret_cast<int32_t*>(dc + o.uq_flag);
    A.wl = reinterpret_cast<int32_t*>(dc + o.wl);
    A.deep_path = reinterpret_cast<int2*>(dc + o.deep_path);
    A.deep_chg = reinterpret_cast<int32_t*>(dc + o.deep_chg);
    A.mlog = reinterpret_cast<int32_t*>(dc + o.mlog);
    A.queue = reinterpret_cast<unsigned long long*>(dc + o.queue);
    A.tlog = reinterpret_cast<unsigned long long*>(dc + o.tlog);
    A.qcap = n_items + 1, A.sharded = sharded ? 1 : 0;
    A.log_cap = getenv("LZG_CLIMB_LOGCAP") ? std::max(1, std::min(LZG_CLIMB_LOG, atoi(getenv("LZG_CLIMB_LOGCAP")))) : LZG_CLIMB_LOG;  // tests shrink it to exercise the drain-and-continue path
    A.use_cond = cg ? 1 : 0, A.cond_handle = cg ? (unsigned long long)cg->handle : 0ull;
    LZG_CUDA(ctx, cudaMemcpyToSymbolAsync(c_pass, &A, sizeof(A), 0, cudaMemcpyHostToDevice, ctx->stream));
    int32_t* h_ctl = ctx->h_pinned;  // the control block goes up and comes back through the head of the staging area
    for (int q = 0; q < LZG_CTL_INTS; q++) h_ctl[q] = 0;
    h_ctl[LZG_CTL_BEST] = -1;
    h_ctl[LZG_CTL_NPATHS] = n_paths;
    h_ctl[LZG_CTL_MAXMOVES] = max_moves;
    h_ctl[LZG_CTL_EPOCH] = t->df_epoch;
    h_ctl[LZG_CTL_CURRENT] = h_ctl[LZG_CTL_PARTROOT] = t->root_score;
    LZG_CUDA(ctx, cudaMemcpyAsync(t->d_cand, h_ctl, sizeof(int32_t) * LZG_CTL_INTS, cudaMemcpyHostToDevice, ctx->stream));
    LZG_CUDA(ctx, cudaEventRecord(ctx->ev_stage, ctx->stream));
    // ---- run ----
    int32_t* h_log = reinterpret_cast<int32_t*>(blob + ((o.upload_bytes + 63) / 64 * 64));  // behind the upload area (read by the H2D above only)
    const bool debug = getenv("LZG_CLIMB_DEBUG") != nullptr;
    int total_moves = 0, stop = 0, passes = 0;
    long long replayed = 0, item_units_before = 0, up_units_before = 0;
    t->up_valid = false;
    while (true) {
        // every (re)start scores the whole neighbourhood once: queues empty, everything dirty
        static const int32_t x_init[LZG_X_INTS] = {0, 0, 0, 0, 0, 1, 0, 0};
        if (passes > 0) {  // a restart after a drained log: keep the work counters
            int32_t h_x[LZG_X_INTS];
            LZG_CUDA(ctx, cudaMemcpyAsync(h_x, A.xctl, sizeof(h_x), cudaMemcpyDeviceToHost, ctx->stream));
            LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            item_units_before += h_x[LZG_X_WL_TOTAL], up_units_before += h_x[LZG_X_UQ_TOTAL];
        }
        LZG_CUDA(ctx, cudaMemcpyAsync(A.xctl, x_init, sizeof(x_init), cudaMemcpyHostToDevice, ctx->stream));
        const int passes_before = passes;
        if (cg) {
            LZG_CUDA(ctx, cudaGraphLaunch(cg->exec, ctx->stream));
        } else {
            // queued mode also times the kernels of the first (full) pass one by one: the bench's per-kernel rooflines
            cudaEvent_t marks[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
            const bool timed = passes == 0;
            if (timed)
                for (auto& m : marks) cudaEventCreate(&m);
            rc = climb_enqueue(ctx, G, PASS0, 5, comm, t, timed ? marks : nullptr);
            if (timed && rc == LZG_OK) {
                cudaEventSynchronize(marks[5]);
                for (int q = 0; q < 5; q++) cudaEventElapsedTime(&t->pass0_ms[q], marks[q], marks[q + 1]);
                t->pass0_timed = true;
            }
            if (timed)
                for (auto& m : marks) cudaEventDestroy(m);
            int batch = 2;
            while (rc == LZG_OK) {
                for (int q = 0; q < batch && rc == LZG_OK; q++) rc = climb_enqueue(ctx, G, ITERATION, 6, comm, t);
                if (rc != LZG_OK) break;
                LZG_CUDA(ctx, cudaMemcpyAsync(h_ctl, t->d_cand, sizeof(int32_t) * LZG_CTL_INTS, cudaMemcpyDeviceToHost, ctx->stream));
                LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
                if (h_ctl[LZG_CTL_STOP]) break;
                batch = std::min(2 * batch, 64);
            }
            if (rc != LZG_OK) {
                cudaStreamSynchronize(ctx->stream);
                return rc;
            }
        }
        LZG_CUDA(ctx, cudaMemcpyAsync(h_ctl, t->d_cand, sizeof(int32_t) * LZG_CTL_INTS, cudaMemcpyDeviceToHost, ctx->stream));
        LZG_CUDA(ctx, cudaMemcpyAsync(h_log, A.mlog, sizeof(int32_t) * 5 * (size_t)LZG_CLIMB_LOG, cudaMemcpyDeviceToHost, ctx->stream));
        LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        stop = h_ctl[LZG_CTL_STOP];
        const int logged = h_ctl[LZG_CTL_NLOGGED];
        passes = h_ctl[LZG_CTL_PASSES];
        if (cg) ctx->launches += 5 + 6 * (int64_t)std::max(0, passes - passes_before - 1);  // kernels the graph ran: pass 0 + the iterations
        total_moves = h_ctl[LZG_CTL_NMOVES];
        if (debug) fprintf(stderr, "[lzg climb] %s: stop %d, moves %d (logged %d), passes %d, current %d, paths %d\n", cg ? "graph" : "batches", stop, total_moves, logged, passes, h_ctl[LZG_CTL_CURRENT], h_ctl[LZG_CTL_NPATHS]);
        for (int k = 0; k < logged; k++) {
            const int32_t* r = h_log + 5 * (size_t)k;
            if (!is_move(t, r[0], r[1], r[2], r[3])) return lzg_fail(ctx, LZG_ERR_STATE, "lzg_hill_climb: the device logged a move the host mirror cannot apply");
            host_apply_move(t, r[0], r[1], r[2], r[3]);
            if (moves && replayed < moves_cap) memcpy(moves + 5 * replayed, r, sizeof(int32_t) * 5);
            replayed++;
        }
        if (stop == 3 && total_moves < max_moves) {  // only the log was full: drain it and go on
            LZG_CUDA(ctx, cudaMemsetAsync(t->d_cand + LZG_CTL_STOP, 0, sizeof(int32_t), ctx->stream));
            LZG_CUDA(ctx, cudaMemsetAsync(t->d_cand + LZG_CTL_NLOGGED, 0, sizeof(int32_t), ctx->stream));
            continue;
        }
        break;
    }
    t->df_epoch = h_ctl[LZG_CTL_EPOCH] + 1;
    if (stop == 4 || stop == 0) return lzg_fail(ctx, LZG_ERR_STATE, stop == 4 ? "lzg_hill_climb: a work queue waited for an entry that was never pushed" : "lzg_hill_climb: the climb did not terminate");
    if (replayed != total_moves) return lzg_fail(ctx, LZG_ERR_STATE, "lzg_hill_climb: move log incomplete");
    {   // the path count the device maintained incrementally must match the final topology
        int heads = 0;
        const int4* nd = t->node4.data();
        for (int p = 0; p < n; p++)
            if (nd[p].x >= 0) heads += p == 0 ? (nd[p].z & 1) + ((nd[p].z >> 1) & 1) : ((nd[p].z & 3) == 3);
        if (heads != h_ctl[LZG_CTL_NPATHS]) return lzg_fail(ctx, LZG_ERR_STATE, "lzg_hill_climb: path count diverged from the topology");
    }
    // time stamps of the passes
    {
        const int n_st = std::min(passes, A.log_cap + 1);
        t->climb_stamps.resize(n_st);
        if (n_st > 0) {
            LZG_CUDA(ctx, cudaMemcpyAsync(t->climb_stamps.data(), A.tlog, sizeof(uint64_t) * n_st, cudaMemcpyDeviceToHost, ctx->stream));
            LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        }
    }
    if (total_moves > 0) t->root_score = h_ctl[LZG_CTL_PARTROOT];
    t->last_cand_score = h_ctl[LZG_CTL_LAST];
    t->up_valid = false;  // the table-driven pass of the stepwise calls keeps its own books
    R.n_moves = total_moves;
    R.n_passes = passes;
    R.reason = stop;
    R.true_score = h_ctl[LZG_CTL_CURRENT];
    R.returned_score = stop == 1 ? h_ctl[LZG_CTL_LAST] : h_ctl[LZG_CTL_CURRENT];
    R.candidates = (int64_t)passes * n_cand;
    {   // how much of the neighbourhood was really re-scored / how many UP rows were recomputed (chunk-group units)
        int32_t h_x[LZG_X_INTS];
        LZG_CUDA(ctx, cudaMemcpyAsync(h_x, A.xctl, sizeof(h_x), cudaMemcpyDeviceToHost, ctx->stream));
        LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        R.item_units = h_x[LZG_X_WL_TOTAL] + item_units_before;
        R.up_units = h_x[LZG_X_UQ_TOTAL] + up_units_before;
        R.units_per_full_pass = (int64_t)n_items * groups;
    }
    if (out) *out = R;
    return LZG_OK;
}
