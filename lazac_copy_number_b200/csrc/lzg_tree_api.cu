// lzg_tree_api.cu -- C ABI for the resident-tree small-parsimony path (include/lazac_b200.h).
#include <algorithm>
#include <new>

#include "lzg_tree.cuh"

static void tree_release(lzg_tree* t) {
    if (!t) return;
    // stream-ordered frees into the context's pool (lzg_dev_alloc): the search loop clones and drops a
    // 40 MB .. GB tree per iteration; cudaMalloc / cudaFree cost it 5.7 ms per clone at cfg4's size
    lzg_dev_free(t->ctx, t->d_child0);
    lzg_dev_free(t->ctx, t->d_child1);
    lzg_dev_free(t->ctx, t->d_up);
    lzg_dev_free(t->ctx, t->d_start);
    lzg_dev_free(t->ctx, t->d_end);
    lzg_dev_free(t->ctx, t->d_score);
    lzg_dev_free(t->ctx, t->d_ctl);
    lzg_dev_free(t->ctx, t->d_order);
    lzg_dev_free(t->ctx, t->d_pass);
    lzg_dev_free(t->ctx, t->d_cand);
    lzg_dev_free(t->ctx, t->d_ustart);
    lzg_dev_free(t->ctx, t->d_uend);
    lzg_dev_free(t->ctx, t->d_ready);
    lzg_dev_free(t->ctx, t->d_node4);
    lzg_dev_free(t->ctx, t->d_climb);
    delete t;
}

static int tree_alloc(lzg_ctx* ctx, int n_nodes, int L, int l_pad, int width, lzg_tree** out) {
    lzg_tree* t = new (std::nothrow) lzg_tree();
    if (!t) return LZG_ERR_ARG;
    t->ctx = ctx;
    t->n_nodes = n_nodes;
    t->L = L;
    t->l_pad = l_pad;
    t->width = width;
    const size_t nb = (size_t)n_nodes * l_pad;
    cudaError_t e = lzg_dev_alloc(ctx, (void**)&t->d_child0, sizeof(int32_t) * n_nodes);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&t->d_child1, sizeof(int32_t) * n_nodes);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&t->d_up, sizeof(int2) * n_nodes);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&t->d_start, nb);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&t->d_end, nb);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&t->d_score, sizeof(int32_t) * n_nodes);
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&t->d_ctl, sizeof(int32_t) * (16 + n_nodes + 1));
    if (e == cudaSuccess) t->d_cost = t->d_ctl + 16;
    if (e == cudaSuccess) e = lzg_dev_alloc(ctx, (void**)&t->d_order, sizeof(int32_t) * n_nodes);
    if (e == cudaSuccess) e = cudaMemsetAsync(t->d_score, 0, sizeof(int32_t) * n_nodes, ctx->stream);
    if (e != cudaSuccess) {
        tree_release(t);
        return lzg_fail(ctx, LZG_ERR_CUDA, "tree allocation", e);
    }
    *out = t;
    return LZG_OK;
}

// Normalises children to ascending ids and checks the reference's structural requirement
// (every node has 0 or 2 children, src/copy_number.cxx:241-242) plus reachability from node 0.
static int set_topology_host(lzg_ctx* ctx, lzg_tree* t, const int32_t* child0, const int32_t* child1) {
    const int n = t->n_nodes;
    t->child0.resize(n);
    t->child1.resize(n);
    std::vector<int> indeg(n, 0);
    for (int u = 0; u < n; u++) {
        int a = child0[u], b = child1[u];
        if ((a < 0) != (b < 0)) return lzg_fail(ctx, LZG_ERR_NOT_BINARY, "every child must have exactly two children");
        if (a >= 0) {
            if (a >= n || b >= n || a == b || a == u || b == u) return lzg_fail(ctx, LZG_ERR_ARG, "child id out of range");
            indeg[a]++;
            indeg[b]++;
        } else {
            a = b = -1;
        }
        t->child0[u] = std::min(a, b);
        t->child1[u] = std::max(a, b);
    }
    if (n > 0 && indeg[0] != 0) return lzg_fail(ctx, LZG_ERR_ARG, "node 0 must be the root");
    for (int u = 1; u < n; u++)
        if (indeg[u] != 1) return lzg_fail(ctx, LZG_ERR_ARG, "every non-root node needs exactly one parent");
    // reachability (rules out cycles that are detached from the root)
    std::vector<int32_t> stack(1, 0);
    int seen = 0;
    while (!stack.empty()) {
        int u = stack.back();
        stack.pop_back();
        seen++;
        if (t->child0[u] >= 0) {
            stack.push_back(t->child0[u]);
            stack.push_back(t->child1[u]);
        }
    }
    if (seen != n) return lzg_fail(ctx, LZG_ERR_ARG, "tree is not connected from node 0");
    return LZG_OK;
}

extern "C" {

int lzg_tree_create(lzg_ctx* ctx, const lzg_profiles* p, int32_t n_nodes, const int32_t* child0, const int32_t* child1,
                    const int32_t* leaf_row, lzg_tree** out) {
    if (!ctx || !p || p->ctx != ctx || !out || n_nodes < 1 || !child0 || !child1 || !leaf_row)
        return lzg_fail(ctx, LZG_ERR_ARG, "lzg_tree_create: bad argument");
    *out = nullptr;
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    for (int u = 0; u < n_nodes; u++) {
        bool leaf = child0[u] < 0 && child1[u] < 0;
        if (leaf && (leaf_row[u] < 0 || leaf_row[u] >= p->n))
            return lzg_fail(ctx, LZG_ERR_ARG, "leaf without a profile row");  // bad_optional_access in the reference
    }
    lzg_tree* t = nullptr;
    int rc = tree_alloc(ctx, n_nodes, p->L, p->pitch, p->width, &t);
    if (rc != LZG_OK) return rc;
    rc = set_topology_host(ctx, t, child0, child1);
    if (rc == LZG_OK) rc = lzg_tree_upload_topology(ctx, t);
    if (rc == LZG_OK) {
        std::vector<int32_t> rows(n_nodes);
        for (int u = 0; u < n_nodes; u++) rows[u] = t->child0[u] < 0 ? leaf_row[u] : -1;
        rc = lzg_tree_gather_leaves(ctx, t, p, rows.data());
    }
    if (rc != LZG_OK) {
        tree_release(t);
        return rc;
    }
    *out = t;
    return LZG_OK;
}

int lzg_tree_clone(lzg_ctx* ctx, const lzg_tree* s, lzg_tree** out) {
    if (!ctx || !s || s->ctx != ctx || !out) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_tree_clone: bad argument");
    *out = nullptr;
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    lzg_tree* t = nullptr;
    int rc = tree_alloc(ctx, s->n_nodes, s->L, s->l_pad, s->width, &t);
    if (rc != LZG_OK) return rc;
    t->child0 = s->child0;
    t->child1 = s->child1;
    t->parent = s->parent;
    t->scored = s->scored;
    t->root_score = s->root_score;
    if (s->node4_valid) {  // the packed node records of the topology: a clone is usually climbed right away
        t->node4 = s->node4;
        t->node4_valid = true;
    }
    const int n = s->n_nodes;
    const size_t nb = (size_t)n * s->l_pad;
    cudaError_t e = cudaMemcpyAsync(t->d_child0, s->d_child0, sizeof(int32_t) * n, cudaMemcpyDeviceToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(t->d_child1, s->d_child1, sizeof(int32_t) * n, cudaMemcpyDeviceToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(t->d_up, s->d_up, sizeof(int2) * n, cudaMemcpyDeviceToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(t->d_start, s->d_start, nb, cudaMemcpyDeviceToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(t->d_end, s->d_end, nb, cudaMemcpyDeviceToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(t->d_score, s->d_score, sizeof(int32_t) * n, cudaMemcpyDeviceToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) {
        tree_release(t);
        return lzg_fail(ctx, LZG_ERR_CUDA, "tree clone", e);
    }
    *out = t;
    return LZG_OK;
}

int lzg_tree_free(lzg_tree* t) {
    if (!t) return LZG_ERR_ARG;
    cudaSetDevice(t->ctx->device);
    cudaStreamSynchronize(t->ctx->stream);
    tree_release(t);
    return LZG_OK;
}

int lzg_tree_set_topology(lzg_ctx* ctx, lzg_tree* t, const int32_t* child0, const int32_t* child1) {
    if (!ctx || !t || t->ctx != ctx || !child0 || !child1) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_tree_set_topology: bad argument");
    // the leaf set must not change: leaf intervals stay where they are
    for (int u = 0; u < t->n_nodes; u++)
        if ((child0[u] < 0 && child1[u] < 0) != (t->child0[u] < 0)) return lzg_fail(ctx, LZG_ERR_ARG, "leaf set changed");
    std::vector<int32_t> old0 = t->child0, old1 = t->child1;
    int rc = set_topology_host(ctx, t, child0, child1);
    if (rc != LZG_OK) {
        t->child0 = old0;
        t->child1 = old1;
        return rc;
    }
    t->scored = false;
    t->up_valid = false;
    t->node4_valid = false;
    t->d_node4_valid = false;
    return lzg_tree_upload_topology(ctx, t);
}

int lzg_tree_get_topology(const lzg_tree* t, int32_t* child0, int32_t* child1) {
    if (!t || !child0 || !child1) return LZG_ERR_ARG;
    std::copy(t->child0.begin(), t->child0.end(), child0);
    std::copy(t->child1.begin(), t->child1.end(), child1);
    return LZG_OK;
}

int lzg_tree_score(lzg_ctx* ctx, lzg_tree* t, int32_t* root_score) {
    if (!ctx || !t || t->ctx != ctx || !root_score) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_tree_score: bad argument");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!t->scored) {
        int rc = lzg_tree_full_score(ctx, t);
        if (rc != LZG_OK) return rc;
    }
    *root_score = t->root_score;
    return LZG_OK;
}

int lzg_nni_best(lzg_ctx* ctx, lzg_tree* t, int32_t n_moves, const int32_t* moves, int32_t greedy, int32_t* best_index,
                 int32_t* best_score, int32_t* all_scores) {
    if (!ctx || !t || t->ctx != ctx || n_moves < 0 || (n_moves > 0 && !moves) || !best_index || !best_score)
        return lzg_fail(ctx, LZG_ERR_ARG, "lzg_nni_best: bad argument");
    if (!t->scored) return lzg_fail(ctx, LZG_ERR_STATE, "tree must be scored (lzg_tree_score) before lzg_nni_best");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    return lzg_tree_nni_scores(ctx, t, n_moves, moves, greedy, best_index, best_score, all_scores);
}

int lzg_nni_scores_launch(lzg_ctx* ctx, lzg_tree* t, int32_t n_moves, const int32_t* moves, int32_t** d_exchange,
                          int32_t* n_exchange) {
    if (!ctx || !t || t->ctx != ctx || n_moves < 0 || (n_moves > 0 && !moves) || !d_exchange || !n_exchange)
        return lzg_fail(ctx, LZG_ERR_ARG, "lzg_nni_scores_launch: bad argument");
    if (!t->scored) return lzg_fail(ctx, LZG_ERR_STATE, "tree must be scored (lzg_tree_score) before lzg_nni_scores_launch");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    const int rc = lzg_tree_nni_launch(ctx, t, n_moves, moves);
    if (rc != LZG_OK) return rc;
    *d_exchange = n_moves ? lzg_tree_exchange_vector(t) : nullptr;
    *n_exchange = n_moves ? n_moves + 1 : 0;
    return LZG_OK;
}

int lzg_nni_scores_finish(lzg_ctx* ctx, lzg_tree* t, int32_t greedy, int32_t* best_index, int32_t* best_score,
                          int32_t* all_scores) {
    if (!ctx || !t || t->ctx != ctx || !best_index || !best_score)
        return lzg_fail(ctx, LZG_ERR_ARG, "lzg_nni_scores_finish: bad argument");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    return lzg_tree_nni_finish(ctx, t, greedy, best_index, best_score, all_scores);
}

int lzg_tree_apply_nni(lzg_ctx* ctx, lzg_tree* t, int32_t u, int32_t w, int32_t v, int32_t z, int32_t* root_score) {
    if (!ctx || !t || t->ctx != ctx) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_tree_apply_nni: bad argument");
    if (!t->scored) return lzg_fail(ctx, LZG_ERR_STATE, "tree must be scored before lzg_tree_apply_nni");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = lzg_tree_apply(ctx, t, u, w, v, z);
    if (rc == LZG_OK && root_score) *root_score = t->root_score;
    return rc;
}

int lzg_hill_climb(lzg_ctx* ctx, lzg_tree* t, int32_t n_edges, const int32_t* edge_order, int32_t greedy, int32_t max_moves,
                   int32_t moves_cap, int32_t* moves, lzg_climb_result* out) {
    if (!ctx || !t || t->ctx != ctx || n_edges < 0 || (n_edges > 0 && !edge_order) || moves_cap < 0 || (moves_cap > 0 && !moves))
        return lzg_fail(ctx, LZG_ERR_ARG, "lzg_hill_climb: bad argument");
    if (!t->scored) return lzg_fail(ctx, LZG_ERR_STATE, "tree must be scored (lzg_tree_score) before lzg_hill_climb");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    return lzg_tree_climb(ctx, nullptr, t, n_edges, edge_order, greedy, max_moves, moves_cap, moves, out);
}

int lzg_tree_last_climb_stamps(const lzg_tree* t, int32_t cap, uint64_t* stamps, int32_t* n) {
    if (!t || !n || cap < 0 || (cap > 0 && !stamps)) return LZG_ERR_ARG;
    *n = (int32_t)t->climb_stamps.size();
    for (int k = 0; k < cap && k < *n; k++) stamps[k] = t->climb_stamps[k];
    return LZG_OK;
}

int lzg_tree_last_climb_pass0_ms(const lzg_tree* t, float* ms5) {
    if (!t || !ms5) return LZG_ERR_ARG;
    if (!t->pass0_timed) return LZG_ERR_STATE;
    for (int q = 0; q < 5; q++) ms5[q] = t->pass0_ms[q];
    return LZG_OK;
}

int lzg_tree_get_intervals(lzg_ctx* ctx, const lzg_tree* t, int32_t* start, int32_t* end, int32_t* score) {
    if (!ctx || !t || t->ctx != ctx) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_tree_get_intervals: bad argument");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    return lzg_tree_download(ctx, t, start, end, score);
}

int lzg_tree_get_node_intervals(lzg_ctx* ctx, const lzg_tree* t, int32_t n_ids, const int32_t* ids, int32_t* start, int32_t* end) {
    if (!ctx || !t || t->ctx != ctx || n_ids < 0 || (n_ids > 0 && !ids)) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_tree_get_node_intervals: bad argument");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    return lzg_tree_download_rows(ctx, t, n_ids, ids, start, end);
}

int lzg_tree_ancestral(lzg_ctx* ctx, const lzg_tree* t, int32_t* labels, int32_t* in_branch_len) {
    if (!ctx || !t || t->ctx != ctx) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_tree_ancestral: bad argument");
    if (!t->scored) return lzg_fail(ctx, LZG_ERR_STATE, "tree must be scored before lzg_tree_ancestral");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    return lzg_tree_label(ctx, t, labels, in_branch_len);
}

int lzg_tree_last_nni_work(const lzg_tree* t, int64_t* merges, int64_t* bytes) {
    if (!t) return LZG_ERR_ARG;
    std::vector<int32_t> depth(t->n_nodes, -1);
    depth[0] = 0;
    int64_t m = 0;
    std::vector<int32_t> whole;
    if (t->last_whole)
        for (int v = 1; v < t->n_nodes; v++)
            if (t->child0[v] >= 0) whole.push_back(v), whole.push_back(v);
    for (int v : (t->last_whole ? whole : t->last_v)) {
        // depth by walking up with memoisation
        int a = v, steps = 0;
        while (depth[a] < 0) {
            a = t->parent[a];
            steps++;
        }
        int d = depth[a] + steps;
        int b = v, k = d;
        while (depth[b] < 0) {
            depth[b] = k--;
            b = t->parent[b];
        }
        m += d + 1;
    }
    if (merges) *merges = m;
    if (bytes) *bytes = (m + (int64_t)t->last_v.size()) * 2 * (int64_t)t->L;
    return LZG_OK;
}

int lzg_tree_last_candidate_score(const lzg_tree* t, int32_t* score) {
    if (!t || !score) return LZG_ERR_ARG;
    *score = t->last_cand_score;
    return LZG_OK;
}

int lzg_tree_last_nni_traffic(const lzg_tree* t, int64_t* cand_bytes, int64_t* up_bytes) {
    if (!t) return LZG_ERR_ARG;
    const int64_t c = (int64_t)t->last_items, L = t->L;  // one item = both candidates of an edge (u, v)
    if (cand_bytes) *cand_bytes = (c * 10 - (int64_t)t->last_root_items * 4) * L;
    if (up_bytes) *up_bytes = t->last_up_ran ? ((int64_t)t->up_nodes * 4 + (int64_t)t->up_paths * 2) * L : 0;
    return LZG_OK;
}

}  // extern "C"
