// host/search.hpp -- host side of `lazac nni`: topology mirror, RNG-driven perturbation and the
// hill climb that drives the GPU through the C ABI (include/lazac_b200.h).
// Follows /root/reference/src/copy_number.cxx:279-313 (nni, unvisit), :322-373 (greedy_nni),
// :375-417 (hill_climb), :419-462 (stochastic_nni) and the search loop of src/lazac.cxx:237-306.
#pragma once
#include <random>
#include <string>
#include <utility>
#include <vector>

#include "lazac_b200.h"

namespace lazac_host {

// Rooted binary tree with root 0; children kept in ascending id order (the iteration order of the
// reference's std::set successors), -1 for leaves.
struct Topology {
    std::vector<int> c0, c1, parent;
    int size() const { return (int)c0.size(); }
    bool leaf(int u) const { return c0[u] < 0; }
    // digraph::edges(): by parent id, then child id
    std::vector<std::pair<int, int>> edges() const;
    void set_children(int u, int a, int b);
    void nni(int u, int w, int v, int z);  // edges (u,w),(v,z) -> (v,w),(u,z)
    void rebuild_parents();
};

// stochastic_nni (src/copy_number.cxx:419-462): floor(#internal edges * aggression) random NNIs,
// three uniform_int_distribution draws each; the float product is the reference's.
Topology stochastic_nni(const Topology& t, std::ranlux48_base& gen, float aggression);
// The same RNG draws without the tree: do_nni perturbs a temporary and drops it
// (src/lazac.cxx:284); the draws do not depend on the topology (an NNI keeps the number of
// internal edges, the other two ranges are constants).
void stochastic_nni_draws(const Topology& t, std::ranlux48_base& gen, float aggression);
// `lazac selftest`: the indexed stochastic_nni against the reference's list-rebuilding procedure
// on a random tree (same result, same generator state afterwards; also for the draws-only variant).
bool stochastic_nni_selftest(int n_leaves, unsigned seed, float aggression);

struct GpuTree {
    lzg_tree* h = nullptr;
    Topology topo;
    int score = 0;
};

struct ClimbStats {
    long passes = 0, candidates = 0, moves = 0;
};

// hill_climb (src/copy_number.cxx:375-417) on a scored tree.  Returns what the reference's returned
// tree reports as t[0].data.score: the score of the last candidate probed by the final, fruitless
// pass (see lzg_tree_last_candidate_score) -- that is what do_nni's acceptance test compares.
int hill_climb(lzg_ctx* ctx, GpuTree& t, std::ranlux48_base& gen, bool greedy, ClimbStats* stats);
// The same climb with one library call per pass and per accepted move (lzg_nni_best + lzg_tree_apply_nni),
// the host enumerating the neighbourhood in between; kept as the cross-check of the device-side loop.
int hill_climb_stepwise(lzg_ctx* ctx, GpuTree& t, const std::vector<int>& random_indices, bool greedy, ClimbStats* stats);

[[noreturn]] void throw_lzg(lzg_ctx* ctx, int rc);

}  // namespace lazac_host
