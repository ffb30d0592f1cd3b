// host/search.cpp -- see search.hpp.
#include "search.hpp"

#include <algorithm>
#include <numeric>
#include <stdexcept>

namespace lazac_host {

void throw_lzg(lzg_ctx* ctx, int rc) {
    const std::string text = lzg_last_error(ctx);
    if (rc == LZG_ERR_NOT_BINARY) throw std::logic_error("every child must have exactly two children");
    throw std::runtime_error(std::string(lzg_status_string(rc)) + (text.empty() ? "" : ": " + text));
}

std::vector<std::pair<int, int>> Topology::edges() const {
    std::vector<std::pair<int, int>> e;
    e.reserve(size());
    for (int u = 0; u < size(); u++)
        if (c0[u] >= 0) {
            e.emplace_back(u, c0[u]);
            e.emplace_back(u, c1[u]);
        }
    return e;
}

void Topology::set_children(int u, int a, int b) {
    c0[u] = std::min(a, b);
    c1[u] = std::max(a, b);
}

void Topology::nni(int u, int w, int v, int z) {
    const int y = c0[v] == z ? c1[v] : c0[v];
    set_children(v, w, y);
    set_children(u, v, z);
    parent[w] = v;
    parent[z] = u;
}

void Topology::rebuild_parents() {
    parent.assign(size(), -1);
    for (int u = 0; u < size(); u++)
        if (c0[u] >= 0) parent[c0[u]] = parent[c1[u]] = u;
}

namespace {
int rand_int(std::ranlux48_base& gen, int a, int b) {
    std::uniform_int_distribution<int> distrib(a, b);
    return distrib(gen);
}

// internal edges (u, v with v internal) in edges() order
void internal_edges(const Topology& t, std::vector<std::pair<int, int>>& out) {
    out.clear();
    for (int u = 0; u < t.size(); u++) {
        if (t.c0[u] < 0) continue;
        if (!t.leaf(t.c0[u])) out.emplace_back(u, t.c0[u]);
        if (!t.leaf(t.c1[u])) out.emplace_back(u, t.c1[u]);
    }
}
}  // namespace

Topology stochastic_nni(const Topology& t, std::ranlux48_base& gen, float aggression) {
    Topology p = t;
    std::vector<std::pair<int, int>> ie;
    internal_edges(p, ie);
    const int num_perturbations = ie.size() * aggression;  // size_t * float -> float -> int, as the reference
    for (int i = 0; i < num_perturbations; i++) {
        internal_edges(p, ie);
        const auto [u, v] = ie[rand_int(gen, 0, (int)ie.size() - 1)];
        const int w_only = p.c0[u] == v ? p.c1[u] : p.c0[u];
        const int w = (rand_int(gen, 0, 0), w_only);          // one child of u besides v: a draw from {0}
        const int z = rand_int(gen, 0, 1) == 0 ? p.c0[v] : p.c1[v];
        p.nni(u, w, v, z);
    }
    return p;
}

int hill_climb(lzg_ctx* ctx, GpuTree& t, std::ranlux48_base& gen, bool greedy, ClimbStats* stats) {
    // edge index == index of the edge's child end in the initial edges() order: an NNI re-hangs
    // w and z but keeps their indices (src/copy_number.cxx:401-406), so the enumeration only needs
    // "child of edge idx" plus the current topology
    const std::vector<std::pair<int, int>> e0 = t.topo.edges();
    std::vector<int> child_of(e0.size());
    std::vector<int> random_indices(e0.size());
    for (size_t i = 0; i < e0.size(); i++) {
        child_of[i] = e0[i].second;
        random_indices[i] = (int)i;
    }
    std::shuffle(random_indices.begin(), random_indices.end(), gen);

    int current = t.score;
    std::vector<int32_t> moves;
    while (true) {
        moves.clear();
        for (int idx : random_indices) {
            const int v = child_of[idx];
            if (t.topo.leaf(v)) continue;
            const int u = t.topo.parent[v];
            const int w = t.topo.c0[u] == v ? t.topo.c1[u] : t.topo.c0[u];
            for (int z : {t.topo.c0[v], t.topo.c1[v]}) moves.insert(moves.end(), {u, w, v, z});
        }
        int32_t best = -1, best_score = 0;
        int rc = lzg_nni_best(ctx, t.h, (int32_t)(moves.size() / 4), moves.data(), greedy ? 1 : 0, &best, &best_score, nullptr);
        if (rc != LZG_OK) throw_lzg(ctx, rc);
        if (stats) {
            stats->passes++;
            stats->candidates += (long)(moves.size() / 4);
        }
        if (best < 0) {
            // the reference returns the tree with t[0].score left at the LAST candidate's score
            // (greedy_nni does not restore it), and do_nni's acceptance test reads that number
            int32_t stale = t.score;
            rc = lzg_tree_last_candidate_score(t.h, &stale);
            if (rc != LZG_OK) throw_lzg(ctx, rc);
            t.score = stale;
            break;
        }
        const int u = moves[4 * best], w = moves[4 * best + 1], v = moves[4 * best + 2], z = moves[4 * best + 3];
        int32_t new_score = 0;
        rc = lzg_tree_apply_nni(ctx, t.h, u, w, v, z, &new_score);
        if (rc != LZG_OK) throw_lzg(ctx, rc);
        t.topo.nni(u, w, v, z);
        t.score = new_score;
        if (stats) stats->moves++;
        if (current <= new_score) break;
        current = new_score;
    }
    return t.score;
}

}  // namespace lazac_host
