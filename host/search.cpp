// host/search.cpp -- see search.hpp.
#include "search.hpp"

#include <algorithm>
#include <climits>
#include <cstdlib>
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

// Order statistics over "internal edges (u, v with v internal) in edges() order" = by u, child0
// before child1: a Fenwick tree over u of the number of internal children.  The reference
// rebuilds the whole edge list before every draw (src/copy_number.cxx:425-440, O(n log n) per
// step, O(n^2 log n) per call -- SURVEY.md 8(f) rank 4); the k-th entry of that list is all a
// step needs, and an NNI changes the count of u and v only.
struct InternalEdgeIndex {
    std::vector<int> fen;  // 1-based
    int n = 0, total = 0, top = 1;
    static int weight(const Topology& t, int u) { return t.c0[u] < 0 ? 0 : (int)!t.leaf(t.c0[u]) + (int)!t.leaf(t.c1[u]); }
    explicit InternalEdgeIndex(const Topology& t) : fen(t.size() + 1, 0), n(t.size()) {
        for (int u = 0; u < n; u++) {
            const int w = weight(t, u);
            fen[u + 1] += w;
            total += w;
            const int up = (u + 1) + ((u + 1) & -(u + 1));
            if (up <= n) fen[up] += fen[u + 1];
        }
        while (top * 2 <= n) top *= 2;
    }
    void add(int u, int delta) {
        for (int i = u + 1; i <= n; i += i & -i) fen[i] += delta;
    }
    // the k-th (0-based) internal edge
    std::pair<int, int> kth(const Topology& t, int k) const {
        int pos = 0, rem = k;
        for (int step = top; step > 0; step >>= 1)
            if (pos + step <= n && fen[pos + step] <= rem) {
                pos += step;
                rem -= fen[pos];
            }
        const int u = pos;  // 0-based node: prefix(u) <= k < prefix(u + 1)
        if (!t.leaf(t.c0[u]) && rem == 0) return {u, t.c0[u]};
        return {u, t.c1[u]};
    }
};
}  // namespace

Topology stochastic_nni(const Topology& t, std::ranlux48_base& gen, float aggression) {
    Topology p = t;
    InternalEdgeIndex index(p);
    const size_t n_edges = (size_t)index.total;
    const int num_perturbations = n_edges * aggression;  // size_t * float -> float -> int, as the reference
    for (int i = 0; i < num_perturbations; i++) {
        const auto [u, v] = index.kth(p, rand_int(gen, 0, (int)n_edges - 1));  // an NNI keeps the number of internal edges
        const int w_only = p.c0[u] == v ? p.c1[u] : p.c0[u];
        const int w = (rand_int(gen, 0, 0), w_only);          // one child of u besides v: a draw from {0}
        const int z = rand_int(gen, 0, 1) == 0 ? p.c0[v] : p.c1[v];
        const int wu = InternalEdgeIndex::weight(p, u), wv = InternalEdgeIndex::weight(p, v);
        p.nni(u, w, v, z);
        index.add(u, InternalEdgeIndex::weight(p, u) - wu);
        index.add(v, InternalEdgeIndex::weight(p, v) - wv);
    }
    return p;
}

// The reference's procedure word for word (edge list rebuilt before every draw): what the indexed
// version above must reproduce.  Only the self-test calls it.
static Topology stochastic_nni_rebuild(const Topology& t, std::ranlux48_base& gen, float aggression) {
    Topology p = t;
    std::vector<std::pair<int, int>> ie;
    auto list = [&]() {
        ie.clear();
        for (int u = 0; u < p.size(); u++) {
            if (p.c0[u] < 0) continue;
            if (!p.leaf(p.c0[u])) ie.emplace_back(u, p.c0[u]);
            if (!p.leaf(p.c1[u])) ie.emplace_back(u, p.c1[u]);
        }
    };
    list();
    const int num_perturbations = ie.size() * aggression;
    for (int i = 0; i < num_perturbations; i++) {
        list();
        const auto [u, v] = ie[rand_int(gen, 0, (int)ie.size() - 1)];
        const int w_only = p.c0[u] == v ? p.c1[u] : p.c0[u];
        const int w = (rand_int(gen, 0, 0), w_only);
        const int z = rand_int(gen, 0, 1) == 0 ? p.c0[v] : p.c1[v];
        p.nni(u, w, v, z);
    }
    return p;
}

bool stochastic_nni_selftest(int n_leaves, unsigned seed, float aggression) {
    // random topology by coalescence; internal ids 0 .. n_leaves-2 (root 0), leaves behind them
    const int n = 2 * n_leaves - 1;
    Topology t;
    t.c0.assign(n, -1);
    t.c1.assign(n, -1);
    std::ranlux48_base build(seed * 7919u + 1);
    std::vector<int> active;
    for (int i = 0; i < n_leaves; i++) active.push_back(n_leaves - 1 + i);
    int next = n_leaves - 2;
    while (active.size() > 1) {
        const int i = rand_int(build, 0, (int)active.size() - 1);
        const int a = active[i];
        active[i] = active.back();
        active.pop_back();
        const int j = rand_int(build, 0, (int)active.size() - 1);
        const int b = active[j];
        t.set_children(next, a, b);
        active[j] = next--;
    }
    t.rebuild_parents();
    std::ranlux48_base g1(seed), g2(seed), g3(seed);
    const Topology a = stochastic_nni(t, g1, aggression), b = stochastic_nni_rebuild(t, g2, aggression);
    stochastic_nni_draws(t, g3, aggression);
    const auto x1 = g1(), x2 = g2(), x3 = g3();  // the generators must be left in the same state
    return a.c0 == b.c0 && a.c1 == b.c1 && a.parent == b.parent && x1 == x2 && x1 == x3;
}

void stochastic_nni_draws(const Topology& t, std::ranlux48_base& gen, float aggression) {
    size_t n_edges = 0;
    for (int u = 0; u < t.size(); u++) n_edges += (size_t)InternalEdgeIndex::weight(t, u);
    const int num_perturbations = n_edges * aggression;
    for (int i = 0; i < num_perturbations; i++) {
        (void)rand_int(gen, 0, (int)n_edges - 1);
        (void)rand_int(gen, 0, 0);
        (void)rand_int(gen, 0, 1);
    }
}

int hill_climb(lzg_ctx* ctx, GpuTree& t, std::ranlux48_base& gen, bool greedy, ClimbStats* stats) {
    // The shuffle is the host's (the generator is): random_indices over the edges() order of the
    // current topology, src/copy_number.cxx:381-388.  Everything after it -- greedy_nni over the whole
    // neighbourhood, accepting the winner, re-scoring its path, the in-place edge-index patch of
    // :401-406 (an edge index stays with its child end) and the stopping rule -- runs on the device
    // inside ONE call (lzg_hill_climb).  LAZAC_STEPWISE=1 keeps the older call-per-pass loop.
    const int n_edges = t.topo.size() - 1;
    std::vector<int> random_indices((size_t)std::max(0, n_edges));
    for (int i = 0; i < n_edges; i++) random_indices[i] = i;
    std::shuffle(random_indices.begin(), random_indices.end(), gen);
    static const bool stepwise = getenv("LAZAC_STEPWISE") && atoi(getenv("LAZAC_STEPWISE")) != 0;
    if (stepwise) return hill_climb_stepwise(ctx, t, random_indices, greedy, stats);
    lzg_climb_result res{};
    int rc = lzg_hill_climb(ctx, t.h, n_edges, random_indices.data(), greedy ? 1 : 0, INT32_MAX, 0, nullptr, &res);
    if (rc != LZG_OK) throw_lzg(ctx, rc);
    if (res.n_moves > 0) {
        std::vector<int32_t> c0(t.topo.size()), c1(t.topo.size());
        rc = lzg_tree_get_topology(t.h, c0.data(), c1.data());
        if (rc != LZG_OK) throw_lzg(ctx, rc);
        t.topo.c0.assign(c0.begin(), c0.end());
        t.topo.c1.assign(c1.begin(), c1.end());
        t.topo.rebuild_parents();
    }
    // the reference returns the tree with t[0].score left at the LAST candidate's score of the final,
    // fruitless pass (greedy_nni does not restore it), and do_nni's acceptance test reads that number
    t.score = res.returned_score;
    if (stats) {
        stats->passes += res.n_passes;
        stats->candidates += (long)res.candidates;
        stats->moves += res.n_moves;
    }
    return t.score;
}

int hill_climb_stepwise(lzg_ctx* ctx, GpuTree& t, const std::vector<int>& random_indices, bool greedy, ClimbStats* stats) {
    // edge index == index of the edge's child end in the initial edges() order: an NNI re-hangs
    // w and z but keeps their indices (src/copy_number.cxx:401-406), so the enumeration only needs
    // "child of edge idx" plus the current topology
    const std::vector<std::pair<int, int>> e0 = t.topo.edges();
    std::vector<int> child_of(e0.size());
    for (size_t i = 0; i < e0.size(); i++) child_of[i] = e0[i].second;

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
