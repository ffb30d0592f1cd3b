// host/capi.cpp -- C entry points of the host-side search layer (liblazac_host.so), so that library users
// (lazac_copy_number_b200/search.py) run the same code as the `lazac` binary: the RNG-driven parts of
// /root/reference/src/copy_number.cxx (stochastic_nni :419-462, the shuffle and loop of hill_climb :375-417)
// and of the NJ seed tree (src/clearcut.cxx).  Plain pointers and sizes; exceptions become status codes.
#include <algorithm>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "search.hpp"
#include "seed_tree.hpp"

using namespace lazac_host;

namespace {
thread_local std::string g_error;
int fail(const std::exception& e) {
    g_error = e.what();
    return 1;
}
Topology topology_of(int n, const int32_t* c0, const int32_t* c1) {
    Topology t;
    t.c0.assign(n, -1);
    t.c1.assign(n, -1);
    for (int u = 0; u < n; u++)
        if (c0[u] >= 0) t.set_children(u, c0[u], c1[u]);
    t.rebuild_parents();
    return t;
}
}  // namespace

extern "C" {

const char* lzh_last_error(void) { return g_error.c_str(); }

// std::ranlux48_base, the generator of do_nni (src/lazac.cxx:237)
void* lzh_gen_new(int32_t seed) { return new std::ranlux48_base(seed); }
void lzh_gen_free(void* g) { delete static_cast<std::ranlux48_base*>(g); }
uint64_t lzh_gen_next(void* g) { return (*static_cast<std::ranlux48_base*>(g))(); }
int32_t lzh_gen_uniform_int(void* g, int32_t lo, int32_t hi) {
    std::uniform_int_distribution<int> d(lo, hi);
    return d(*static_cast<std::ranlux48_base*>(g));
}
double lzh_gen_uniform_real(void* g, double lo, double hi) {
    std::uniform_real_distribution<double> d(lo, hi);
    return d(*static_cast<std::ranlux48_base*>(g));
}

// stochastic_nni(t, gen, aggression), src/copy_number.cxx:419-462: topology in, topology out
int lzh_stochastic_nni(void* g, int32_t n_nodes, const int32_t* child0, const int32_t* child1, float aggression, int32_t* out0, int32_t* out1) {
    try {
        const Topology p = stochastic_nni(topology_of(n_nodes, child0, child1), *static_cast<std::ranlux48_base*>(g), aggression);
        for (int u = 0; u < n_nodes; u++) out0[u] = p.c0[u], out1[u] = p.c1[u];
        return 0;
    } catch (const std::exception& e) {
        return fail(e);
    }
}

// hill_climb(t, gen, greedy), src/copy_number.cxx:375-417, on a scored GPU tree: the shuffle with the
// caller's generator, then the device-side climb.  score_io: in = the tree's score, out = what the reference's
// returned tree reports as t[0].data.score.  stats: {passes, candidates, moves} (optional).
int lzh_hill_climb(lzg_ctx* ctx, lzg_tree* tree, void* g, int32_t n_nodes, int32_t greedy, int32_t* score_io, int32_t* child0_out,
                   int32_t* child1_out, int64_t* stats) {
    try {
        GpuTree t;
        t.h = tree;
        std::vector<int32_t> c0(n_nodes), c1(n_nodes);
        int rc = lzg_tree_get_topology(tree, c0.data(), c1.data());
        if (rc != LZG_OK) throw_lzg(ctx, rc);
        t.topo = topology_of(n_nodes, c0.data(), c1.data());
        t.score = *score_io;
        ClimbStats st;
        *score_io = hill_climb(ctx, t, *static_cast<std::ranlux48_base*>(g), greedy != 0, &st);
        for (int u = 0; u < n_nodes; u++) child0_out[u] = t.topo.c0[u], child1_out[u] = t.topo.c1[u];
        if (stats) stats[0] = st.passes, stats[1] = st.candidates, stats[2] = st.moves;
        return 0;
    } catch (const std::exception& e) {
        return fail(e);
    }
}

// NJ_shuffle_distance_matrix's taxon order for a seed (MT19937), and Clearcut-exact NJ on the CPU for a given
// order: the same (left, right) arrays as lzg_nj_seed_tree (leaves 0 .. n-1 = shuffled positions, root 2n-2).
int lzh_clearcut_shuffle(int32_t n, uint32_t seed, int32_t* perm) {
    const std::vector<long> p = clearcut_taxon_shuffle(n, seed);
    for (int i = 0; i < n; i++) perm[i] = (int32_t)p[i];
    return 0;
}

int lzh_nj_seed_tree_cpu(const int32_t* D, int32_t n, const int32_t* perm, int32_t* left, int32_t* right) {
    try {
        std::vector<int32_t> d(D, D + (size_t)n * n);
        std::vector<std::string> names(n, "x");
        std::vector<long> p(perm, perm + n);
        const SeedTree t = neighbor_joining_perm(d, names, p);
        for (size_t i = 0; i < t.left.size(); i++) left[i] = t.left[i], right[i] = t.right[i];
        return 0;
    } catch (const std::exception& e) {
        return fail(e);
    }
}

}  // extern "C"
