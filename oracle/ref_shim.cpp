// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE ONLY (never on the product path).
//
// A C-ABI window onto the UNMODIFIED reference implementation.  This file contains no
// algorithm of its own: it only converts flat int32 arrays to/from the reference's
// containers and calls the reference's functions, which are compiled in place from
// /root/reference/src/copy_number.cxx (see oracle/Makefile; outputs go to oracle/_ref/).
//
// Reference entry points called (all in namespace copynumber):
//   convert_to_breakpoint_profile   src/copy_number.cxx:25
//   convert_to_copynumber_profile   src/copy_number.cxx:61
//   breakpoint_magnitude            src/copy_number.cxx:100, operator- include/copy_number.hpp:74
//   small_rectilinear               src/copy_number.cxx:227
//   nni / undo_nni / unvisit        src/copy_number.cxx:279-313
//   greedy_nni (external linkage)   src/copy_number.cxx:322
//   hill_climb                      src/copy_number.cxx:375
//   stochastic_nni                  src/copy_number.cxx:419
//   ancestral_labeling              src/copy_number.cxx:187
#include "copy_number.hpp"

#include <cstdint>
#include <cstring>
#include <map>
#include <optional>
#include <random>
#include <string>
#include <tuple>
#include <vector>

namespace copynumber {
// Defined (not declared in the header) at src/copy_number.cxx:322-325.
std::optional<std::tuple<int, int, int, int>> greedy_nni(digraph<rectilinear_vertex_data>& t,
                                                         const std::map<int, std::pair<int, int>>& indexed_edges,
                                                         const std::vector<int>& edge_indices,
                                                         bool greedy);
}  // namespace copynumber

using namespace copynumber;
typedef digraph<rectilinear_vertex_data> rtree;

struct ref_tree {
    rtree t;
    int n_nodes;
    int L;
};

static const char* kAlleles[2] = {"cn_a", "cn_b"};

static std::vector<genomic_bin> make_bins(int L, const int32_t* chrom, const int32_t* allele,
                                          const int32_t* start, const int32_t* end) {
    std::vector<genomic_bin> bins;
    bins.reserve(L);
    for (int i = 0; i < L; i++) {
        bins.emplace_back(std::to_string(chrom[i]), kAlleles[allele[i] & 1], start[i], end[i]);
    }
    return bins;
}

// Index of each output bin in the caller's input order (bins are unique per profile).
static void bins_to_perm(const std::vector<genomic_bin>& in, const std::vector<genomic_bin>& out,
                         int32_t* perm) {
    std::map<genomic_bin, int> where;
    for (size_t i = 0; i < in.size(); i++) where[in[i]] = (int)i;
    for (size_t i = 0; i < out.size(); i++) perm[i] = where.at(out[i]);
}

extern "C" {

// cn[L] with bins (chrom id, allele id, start, end) in arbitrary order -> bp[L] in the
// reference's (chromosome string, allele string, start, end) order; perm[k] = input index of
// output position k.
int ref_bp_from_cn(int L, const int32_t* chrom, const int32_t* allele, const int32_t* start,
                   const int32_t* end, const int32_t* cn, int32_t* bp_out, int32_t* perm_out) {
    copynumber_profile p;
    p.bins = make_bins(L, chrom, allele, start, end);
    p.profile.assign(cn, cn + L);
    breakpoint_profile bp = convert_to_breakpoint_profile(p, 2);
    if ((int)bp.profile.size() != L) return -1;
    std::memcpy(bp_out, bp.profile.data(), sizeof(int32_t) * L);
    if (perm_out) bins_to_perm(p.bins, bp.bins, perm_out);
    return 0;
}

int ref_cn_from_bp(int L, const int32_t* chrom, const int32_t* allele, const int32_t* start,
                   const int32_t* end, const int32_t* bp, int32_t* cn_out, int32_t* perm_out) {
    breakpoint_profile p;
    p.bins = make_bins(L, chrom, allele, start, end);
    p.profile.assign(bp, bp + L);
    copynumber_profile cn = convert_to_copynumber_profile(p, 2);
    if ((int)cn.profile.size() != L) return -1;
    std::memcpy(cn_out, cn.profile.data(), sizeof(int32_t) * L);
    if (perm_out) bins_to_perm(p.bins, cn.bins, perm_out);
    return 0;
}

// The pair loop of src/lazac.cxx:81-91 over rows already in sorted-name order.
int ref_distance_matrix(int N, int L, const int32_t* bp, int32_t* D) {
    std::vector<breakpoint_profile> profs(N);
    for (int i = 0; i < N; i++) {
        profs[i].profile.assign(bp + (size_t)i * L, bp + (size_t)(i + 1) * L);
        profs[i].bins.resize(L);  // breakpoint_magnitude returns 0 when bins is empty
    }
    for (int i = 0; i < N; i++) {
        for (int j = i; j < N; j++) {
            int d = breakpoint_magnitude(profs[i] - profs[j]);
            D[(size_t)i * N + j] = d;
            D[(size_t)j * N + i] = d;
        }
    }
    return 0;
}

// Rows rows[0..n_rows) of the same pair loop against every cell, from a compact int8 matrix
// (the large configurations: cfg3 is 400 MB as int8).  D_out is [n_rows][N].
int ref_distance_rows_i8(int N, int L, const int8_t* bp, int n_rows, const int32_t* rows, int32_t* D_out) {
    breakpoint_profile a, b;
    a.bins.resize(L);
    b.bins.resize(L);
    a.profile.resize(L);
    b.profile.resize(L);
    for (int r = 0; r < n_rows; r++) {
        const int8_t* ra = bp + (size_t)rows[r] * L;
        for (int k = 0; k < L; k++) a.profile[k] = ra[k];
        for (int j = 0; j < N; j++) {
            const int8_t* rb = bp + (size_t)j * L;
            for (int k = 0; k < L; k++) b.profile[k] = rb[k];
            D_out[(size_t)r * N + j] = breakpoint_magnitude(a - b);
        }
    }
    return 0;
}

// ---- trees -------------------------------------------------------------------------------
// Nodes are 0..n_nodes-1, root = 0.  child0/child1 = -1 for leaves.  leaf_bp is
// [n_nodes][L]; rows of internal nodes are ignored.
ref_tree* ref_tree_new(int n_nodes, int L, const int32_t* child0, const int32_t* child1,
                       const int32_t* leaf_bp) {
    ref_tree* h = new ref_tree();
    h->n_nodes = n_nodes;
    h->L = L;
    for (int u = 0; u < n_nodes; u++) {
        rectilinear_vertex_data r;
        r.name = std::to_string(u);
        if (child0[u] < 0 && child1[u] < 0) {
            std::vector<int> p(leaf_bp + (size_t)u * L, leaf_bp + (size_t)(u + 1) * L);
            r.start = p;
            r.end = p;
        }
        h->t.add_vertex(r);
    }
    for (int u = 0; u < n_nodes; u++) {
        if (child0[u] >= 0) h->t.add_edge(u, child0[u]);
        if (child1[u] >= 0) h->t.add_edge(u, child1[u]);
    }
    return h;
}

// Same, leaves taken from a compact int8 matrix: leaf_row[u] = row of bp for leaf u (-1 internal).
ref_tree* ref_tree_new_i8(int n_nodes, int L, const int32_t* child0, const int32_t* child1,
                          const int32_t* leaf_row, const int8_t* bp) {
    ref_tree* h = new ref_tree();
    h->n_nodes = n_nodes;
    h->L = L;
    for (int u = 0; u < n_nodes; u++) {
        rectilinear_vertex_data r;
        r.name = std::to_string(u);
        if (child0[u] < 0 && child1[u] < 0) {
            const int8_t* row = bp + (size_t)leaf_row[u] * L;
            std::vector<int> p(row, row + L);
            r.start = p;
            r.end = p;
        }
        h->t.add_vertex(r);
    }
    for (int u = 0; u < n_nodes; u++) {
        if (child0[u] >= 0) h->t.add_edge(u, child0[u]);
        if (child1[u] >= 0) h->t.add_edge(u, child1[u]);
    }
    return h;
}

ref_tree* ref_tree_clone(const ref_tree* h) { return new ref_tree(*h); }
void ref_tree_free(ref_tree* h) { delete h; }

// Returns 0, or 1 when the reference throws std::logic_error (non-binary node).
int ref_small_rectilinear(ref_tree* h, int32_t* root_score) {
    try {
        small_rectilinear(h->t, 0);
    } catch (const std::logic_error&) {
        return 1;
    }
    *root_score = h->t[0].data.score;
    return 0;
}

void ref_unvisit_all(ref_tree* h) { unvisit(h->t, 0); }

int ref_get_node(const ref_tree* h, int u, int32_t* start, int32_t* end, int32_t* score,
                 int32_t* visited) {
    const rectilinear_vertex_data& d = h->t[u].data;
    if (score) *score = d.score;
    if (visited) *visited = d.visited ? 1 : 0;
    if (!d.start.has_value()) return 1;
    if (start) std::memcpy(start, d.start->data(), sizeof(int32_t) * h->L);
    if (end) std::memcpy(end, d.end->data(), sizeof(int32_t) * h->L);
    return 0;
}

// score of every node
void ref_get_scores(const ref_tree* h, int32_t* score) {
    for (int u = 0; u < h->n_nodes; u++) score[u] = h->t[u].data.score;
}

// Children in ascending id order (std::set order), -1 for leaves.
void ref_get_topology(const ref_tree* h, int32_t* child0, int32_t* child1) {
    for (int u = 0; u < h->n_nodes; u++) {
        child0[u] = child1[u] = -1;
        int k = 0;
        for (int c : h->t.successors(u)) {
            if (k == 0) child0[u] = c;
            if (k == 1) child1[u] = c;
            k++;
        }
    }
}

// t.edges() in the reference's iteration order; returns the edge count.
int ref_edges(const ref_tree* h, int32_t* eu, int32_t* ev) {
    int k = 0;
    for (auto [u, v] : h->t.edges()) {
        eu[k] = u;
        ev[k] = v;
        k++;
    }
    return k;
}

void ref_apply_nni(ref_tree* h, int u, int w, int v, int z) {
    nni(h->t, u, w, v, z);
    unvisit(h->t, 0, v);
    small_rectilinear(h->t, 0);
}

// Scores every candidate in enumeration order exactly as the loop body of
// src/copy_number.cxx:347-366 does (mutate, un-visit path, re-score, undo).  Returns the
// number of candidates; moves[k] = (u,w,v,z), scores[k] = root score of candidate k.
int ref_candidate_scores(ref_tree* h, int n_edges, const int32_t* eu, const int32_t* ev,
                         int n_idx, const int32_t* order, int32_t* moves, int32_t* scores,
                         int max_out) {
    rtree& t = h->t;
    int k = 0;
    for (int q = 0; q < n_idx; q++) {
        int idx = order[q];
        if (idx < 0 || idx >= n_edges) return -1;
        int u = eu[idx], v = ev[idx];
        if (t.successors(v).empty()) continue;
        std::vector<int> uc, vc;
        for (int w : t.successors(u))
            if (w != v) uc.push_back(w);
        for (int w : t.successors(v)) vc.push_back(w);
        for (int w : uc) {
            for (int z : vc) {
                nni(t, u, w, v, z);
                unvisit(t, 0, v);
                small_rectilinear(t, 0);
                if (k < max_out) {
                    moves[4 * k + 0] = u;
                    moves[4 * k + 1] = w;
                    moves[4 * k + 2] = v;
                    moves[4 * k + 3] = z;
                    scores[k] = t[0].data.score;
                }
                k++;
                undo_nni(t, u, w, v, z);
                unvisit(t, 0, v);
            }
        }
    }
    small_rectilinear(t, 0);
    return k;
}

// The reference's own neighbourhood search.  Returns 1 and fills move[4] if a move was found.
int ref_greedy_nni(ref_tree* h, int n_edges, const int32_t* eu, const int32_t* ev, int n_idx,
                   const int32_t* order, int greedy, int32_t* move) {
    std::map<int, std::pair<int, int>> indexed;
    for (int i = 0; i < n_edges; i++) indexed[i] = std::make_pair(eu[i], ev[i]);
    std::vector<int> idx(order, order + n_idx);
    auto best = greedy_nni(h->t, indexed, idx, greedy != 0);
    if (!best) return 0;
    auto [u, w, v, z] = *best;
    move[0] = u;
    move[1] = w;
    move[2] = v;
    move[3] = z;
    return 1;
}

// ---- RNG-driven host logic ----------------------------------------------------------------
std::ranlux48_base* ref_gen_new(int seed) { return new std::ranlux48_base(seed); }
void ref_gen_free(std::ranlux48_base* g) { delete g; }
uint64_t ref_gen_next(std::ranlux48_base* g) { return (*g)(); }

void ref_hill_climb(ref_tree* h, std::ranlux48_base* g, int greedy) {
    h->t = hill_climb(h->t, *g, greedy != 0);
}

ref_tree* ref_stochastic_nni(const ref_tree* h, std::ranlux48_base* g, float aggression) {
    ref_tree* out = new ref_tree();
    out->n_nodes = h->n_nodes;
    out->L = h->L;
    out->t = stochastic_nni(h->t, *g, aggression);
    return out;
}

// ---- labeling -----------------------------------------------------------------------------
// Output is indexed by the NEW ids ancestral_labeling assigns (src/copy_number.cxx:215):
// old_id[k] = source node, parent_new[k] = new id of the parent (-1 for the root),
// labels[k][L], branch[k] (-1 for the root).
int ref_ancestral_labeling(ref_tree* h, int32_t* old_id, int32_t* parent_new, int32_t* labels,
                           int32_t* branch) {
    std::vector<genomic_bin> bins(h->L);
    auto bt = ancestral_labeling(h->t, 0, bins);
    int n = (int)bt.nodes().size();
    for (int k = 0; k < n; k++) {
        const auto& d = bt[k].data;
        old_id[k] = std::stoi(d.name);
        parent_new[k] = bt.predecessors(k).empty() ? -1 : *bt.predecessors(k).begin();
        std::memcpy(labels + (size_t)k * h->L, d.profile.profile.data(), sizeof(int32_t) * h->L);
        branch[k] = d.in_branch_length.has_value() ? d.in_branch_length.value() : -1;
    }
    return n;
}

}  // extern "C"
