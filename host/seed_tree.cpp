// host/seed_tree.cpp -- neighbour-joining seed tree, bit-compatible with the Clearcut code the
// reference links (/root/reference/src/clearcut.cxx NJ_neighbor_joining :925, NJ_min_transform
// :1938, NJ_compute_r :417, NJ_collapse :744, NJ_init_r :1413, NJ_shuffle_distance_matrix :1644,
// NJ_permute :338; /root/reference/src/prng.cxx MT19937 + NJ_genrand_int31_top :191).
//
// The reference writes the integer matrix to PREFIX_dist_matrix.txt, re-parses it with atof()
// into float32 (src/dmat.cxx:435) and runs NJ on the file.  Distances are integers below 2^24,
// so the float32 values are exact and the matrix can be handed over in memory;
// what has to be reproduced exactly is everything that decides ties and rounding:
//   * the MT19937 Fisher-Yates shuffle of the taxa before NJ,
//   * the strict '<' row-major scan for the minimum transformed distance d(i,j) - (r2[i] + r2[j]),
//   * float32 accumulation ORDER of the row sums r (initial sums, the two subtraction sweeps
//     after a join, the additions while the joined row is rebuilt -- including the harmless term
//     the reference adds for the vanishing row b itself),
//   * r2 = r / (size - 2) initially (the last entry is never initialised and stays 0) and
//     r / (size - 3) during a collapse,
//   * the index bookkeeping: the new node takes index a, old index 0 moves to index b, and
//     every index then shifts down by one.
// This file keeps a square float matrix and performs exactly that relabelling physically.
#include "seed_tree.hpp"

#include <cmath>
#include <cstdint>
#include <limits>
#include <random>
#include <stdexcept>

namespace lazac_host {

namespace {

// NJ_genrand_int31_top: unbiased draw in [0, top) from 31-bit MT19937 outputs.
long draw_below(std::mt19937& mt, long top) {
    if (top <= 0) return 0;
    const long limit = (0x7fffffffL / top) * top;
    long r;
    do {
        r = (long)(mt() >> 1);
    } while (r >= limit);
    return r % top;
}

struct Joiner {
    long n0;                 // number of taxa
    long size;               // active rows
    long base = 0;           // rows/columns dropped so far
    std::vector<float> m;    // n0 x n0, logical (i, j) lives at (i + base, j + base)
    std::vector<float> r, r2;
    std::vector<int> node;   // tree node id per logical index (offset by base as well)

    float& at(long i, long j) { return m[(size_t)(i + base) * n0 + (j + base)]; }
    float& R(long i) { return r[i + base]; }
    float& R2(long i) { return r2[i + base]; }
    int& N(long i) { return node[i + base]; }
};

}  // namespace

// NJ_shuffle_distance_matrix: position i of the working matrix is taxon perm[i].
static std::vector<long> clearcut_shuffle(long n, uint32_t seed) {
    std::mt19937 mt(seed);
    std::vector<long> perm(n);
    for (long i = 0; i < n; i++) perm[i] = i;
    for (long i = 0; i < n; i++) {
        long k = i + draw_below(mt, n - i);
        if (k != i) std::swap(perm[i], perm[k]);
    }
    return perm;
}

SeedTree neighbor_joining_gpu(lzg_ctx* ctx, const std::vector<int32_t>& D, const std::vector<std::string>& names, uint32_t seed, const lzg_profiles* prof) {
    const long n = (long)names.size();
    if ((size_t)n * n != D.size()) throw std::runtime_error("distance matrix size mismatch");
    if (n == 0) throw std::runtime_error("Failed to parse distance matrix.");
    const std::vector<long> perm = clearcut_shuffle(n, seed);
    std::vector<int32_t> p32(perm.begin(), perm.end()), left(2 * n - 1), right(2 * n - 1);
    // prof: the uploaded profiles D was computed from -- its device copy of the matrix is used, no second trip over PCIe
    const int rc = prof ? lzg_nj_seed_tree_profiles(ctx, prof, p32.data(), left.data(), right.data())
                        : lzg_nj_seed_tree(ctx, D.data(), (int32_t)n, p32.data(), left.data(), right.data());
    if (rc != LZG_OK) throw std::runtime_error(std::string("lzg_nj_seed_tree: ") + lzg_last_error(ctx));
    SeedTree out;
    out.left.assign(left.begin(), left.end());
    out.right.assign(right.begin(), right.end());
    out.leaf_name.resize(2 * n - 1);
    for (long i = 0; i < n; i++) out.leaf_name[i] = names[perm[i]];
    out.root = (int)(2 * n - 2);
    return out;
}

std::vector<long> clearcut_taxon_shuffle(long n, uint32_t seed) { return clearcut_shuffle(n, seed); }

SeedTree neighbor_joining_clearcut(const std::vector<int32_t>& D, const std::vector<std::string>& names, uint32_t seed) {
    return neighbor_joining_perm(D, names, clearcut_shuffle((long)names.size(), seed));
}

SeedTree neighbor_joining_perm(const std::vector<int32_t>& D, const std::vector<std::string>& names, const std::vector<long>& perm) {
    const long n = (long)names.size();
    if ((size_t)n * n != D.size() || (long)perm.size() != n) throw std::runtime_error("distance matrix size mismatch");
    SeedTree out;
    if (n == 0) throw std::runtime_error("Failed to parse distance matrix.");

    Joiner J;
    J.n0 = n;
    J.size = n;
    J.m.assign((size_t)n * n, 0.0f);
    J.r.assign(n, 0.0f);
    J.r2.assign(n, 0.0f);
    J.node.resize(n);
    for (long i = 0; i < n; i++)
        for (long j = 0; j < n; j++)
            if (i != j) J.m[(size_t)i * n + j] = (float)D[(size_t)perm[i] * n + perm[j]];

    // tree nodes: 0 .. n-1 leaves (shuffled positions), then internal nodes in creation order
    out.left.assign(n, -1);
    out.right.assign(n, -1);
    out.leaf_name.resize(n);
    for (long i = 0; i < n; i++) {
        out.leaf_name[i] = names[perm[i]];
        J.node[i] = (int)i;
    }
    auto join = [&](int x, int y) {
        out.left.push_back(x);
        out.right.push_back(y);
        out.leaf_name.push_back(std::string());
        return (int)out.left.size() - 1;
    };

    // ---- initial row sums (NJ_init_r): pairs in row-major order, r2 for all but the last index
    {
        const float denom = (float)(n - 2);
        for (long i = 0; i < n - 1; i++) {
            for (long j = i + 1; j < n; j++) {
                const float v = J.at(i, j);
                J.R(i) += v;
                J.R(j) += v;
            }
            J.R2(i) = J.R(i) / denom;
        }
    }

    while (J.size > 2) {
        const long size = J.size;
        // ---- global minimum of the transformed distances, first one in row-major order wins
        long a = 0, b = 0;
        float smallest = std::numeric_limits<float>::infinity();
        for (long i = 0; i < size; i++) {
            const float* row = &J.at(i, 0);
            const float r2i = J.R2(i);
            for (long j = i + 1; j < size; j++) {
                const float cur = row[j] - (r2i + J.R2(j));
                if (cur < smallest) {
                    smallest = cur;
                    a = i;
                    b = j;
                }
            }
        }
        const int joined = join(J.N(a), J.N(b));

        // ---- take d(., a) and d(., b) out of the row sums (NJ_compute_r), in the reference's sweep order
        for (long i = a + 1; i < size; i++) {
            J.R(i) -= J.at(a, i);
            if (i > b) J.R(i) -= J.at(b, i);
        }
        for (long i = 0; i < b; i++) {
            if (i < a) J.R(i) -= J.at(i, a);
            J.R(i) -= J.at(i, b);
        }

        // ---- rebuild row a as the joined node (NJ_collapse)
        const float dab = J.at(a, b);
        const float a2clade = (float)((dab + (J.R2(a) - J.R2(b))) / 2.0);
        const float b2clade = (float)((dab + (J.R2(b) - J.R2(a))) / 2.0);
        const float denom = (float)(size - 3);
        J.R(a) = 0.0f;
        for (long i = a + 1; i < size; i++) {
            // at i == b this reads the zero diagonal d(b, b): the reference adds that term too
            const float cval = (float)(((J.at(a, i) - a2clade) + (J.at(b, i) - b2clade)) / 2.0);
            J.at(a, i) = cval;
            J.at(i, a) = cval;
            J.R(a) += cval;
            J.R(i) += cval;
            J.R2(i) = J.R(i) / denom;
        }
        for (long i = 0; i < a; i++) {
            const float cval = (float)(((J.at(i, a) - a2clade) + (J.at(i, b) - b2clade)) / 2.0);
            J.at(i, a) = cval;
            J.at(a, i) = cval;
            J.R(a) += cval;
            J.R(i) += cval;
            J.R2(i) = J.R(i) / denom;
        }
        J.R2(a) = J.R(a) / denom;
        J.N(a) = joined;

        // ---- index 0 takes over index b, then index 0 is dropped
        for (long i = 0; i < size; i++) {
            if (i == b) continue;
            const float v = J.at(0, i);
            J.at(b, i) = v;
            J.at(i, b) = v;
        }
        J.at(b, b) = 0.0f;
        J.R(b) = J.R(0);
        J.R2(b) = J.R2(0);
        J.N(b) = J.N(0);
        J.base++;
        J.size--;
    }
    if (n == 1) {
        out.root = 0;
    } else {
        out.root = join(J.N(0), J.N(1));
    }
    return out;
}

// NJ_output_tree with expblen = 0 (src/lazac.cxx:179, src/clearcut.cxx:1300-1400): topology only.
std::string SeedTree::newick() const {
    std::string s;
    // explicit stack instead of recursion: NJ trees on clonal data can be thousands of levels deep
    struct Frame { int u; int stage; };
    std::vector<Frame> st;
    st.push_back({root, 0});
    while (!st.empty()) {
        Frame& f = st.back();
        if (left[f.u] < 0) {
            s += leaf_name[f.u];
            st.pop_back();
        } else if (f.stage == 0) {
            s += '(';
            f.stage = 1;
            st.push_back({left[f.u], 0});
        } else if (f.stage == 1) {
            s += ',';
            f.stage = 2;
            st.push_back({right[f.u], 0});
        } else {
            s += ')';
            st.pop_back();
        }
    }
    s += ";\n";
    return s;
}

}  // namespace lazac_host
