// host/seed_tree.hpp -- NJ seed tree (see seed_tree.cpp).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "lazac_b200.h"

namespace lazac_host {

struct SeedTree {
    std::vector<int> left, right;        // -1 for leaves
    std::vector<std::string> leaf_name;  // empty for internal nodes
    int root = -1;
    std::string newick() const;          // "(...);\n" exactly as the reference's PREFIX_nj_tree.newick
};

// D: n x n symmetric integer matrix, rows/columns in the order of `names`; seed = -s.
SeedTree neighbor_joining_clearcut(const std::vector<int32_t>& D, const std::vector<std::string>& names, uint32_t seed);

// The same with the taxon shuffle given (position i of the working matrix is taxon perm[i]), and the shuffle itself
// (NJ_shuffle_distance_matrix, src/clearcut.cxx:1644, MT19937).
SeedTree neighbor_joining_perm(const std::vector<int32_t>& D, const std::vector<std::string>& names, const std::vector<long>& perm);
std::vector<long> clearcut_taxon_shuffle(long n, uint32_t seed);

// The same tree from the GPU (lzg_nj_seed_tree): the O(n^3) minimum scans run as device reductions.
SeedTree neighbor_joining_gpu(lzg_ctx* ctx, const std::vector<int32_t>& D, const std::vector<std::string>& names, uint32_t seed,
                              const lzg_profiles* prof = nullptr);

}  // namespace lazac_host
