// lzg_tree.cuh -- resident tree state for ZCNT small parsimony (sm_100a).
#pragma once
#include "lzg_common.cuh"

// Replaces digraph<rectilinear_vertex_data> (/root/reference/include/copy_number.hpp:98-105,
// include/digraph.hpp:24-109): topology as flat arrays (host mirror + device copy), the
// per-node interval vectors start/end as biased bytes [n_nodes][l_pad] in HBM, cumulative
// subtree scores as int32.
struct lzg_tree {
    lzg_ctx* ctx = nullptr;
    int32_t n_nodes = 0, L = 0;
    int32_t l_pad = 0;   // BYTES per interval row (padded positions x bytes per position); a thread owns one 16-byte chunk of it
    int32_t width = 1;   // bytes per position: the profiles' device format (lzg_profiles::width)
    std::vector<int32_t> child0, child1, parent;  // host mirror (child0 < child1, -1 for leaves)
    int32_t *d_child0 = nullptr, *d_child1 = nullptr;
    int2* d_up = nullptr;       // (parent, sibling) per node, (-1,-1) for the root
    uint8_t *d_start = nullptr, *d_end = nullptr;
    int32_t* d_score = nullptr;  // cumulative subtree score (t[u].data.score)
    int32_t* d_ctl = nullptr;    // 16 control ints (finished-CTA counter, results) right in front of d_cost
    int32_t* d_cost = nullptr;   // = d_ctl + 16: per-node merge cost / per-step path cost scratch [n_nodes + 1]
    int32_t* d_order = nullptr;  // level-ordered node list scratch [n_nodes]
    // UP intervals (rest of the tree seen from the edge above a node), allocated on first use
    uint8_t *d_ustart = nullptr, *d_uend = nullptr;
    int* d_ready = nullptr;      // per (node, chunk group) epoch flags of the dependency-driven passes, + 1 ticket counter
    int epoch = 0;
    bool up_valid = false;       // cleared by every topology / interval change
    bool scored = false;         // == every node visited
    int32_t root_score = 0;
    // candidate buffers, grown on demand
    int32_t cand_cap = 0;
    int32_t* d_pass = nullptr;        // per-pass upload: [path_off | order | moves]
    int32_t* d_cand = nullptr;        // [ctl (16) | cand_score (cand_cap)]
    std::vector<int32_t> last_v;      // v of every candidate of the last lzg_nni_best
    int32_t last_cand_score = 0;      // score of the last candidate of that call (root score when it had none)
    int32_t last_items = 0;           // work items of that call (an item = the two candidates of one edge (u, v))
    int32_t last_root_items = 0;      // items whose u is the root (no UP / DOWN(u) rows: 6 L instead of 10 L bytes)
    bool last_up_ran = false;         // the UP intervals were rebuilt for that call (by it, or by the lzg_tree_apply before it)
    bool up_ran = false;              // an UP pass was queued since the last lzg_nni_best
    int32_t up_nodes = 0;             // internal non-root nodes (UP rows written by an UP pass)
    int32_t up_paths = 0;             // heavy paths they form
    std::vector<int32_t> up_rank;     // position of a node in the walker order (-1: root, leaves); valid while up_valid
    std::vector<int4> node4;          // {heavy child, light child, internal nodes in the subtree << 2 | flags, parent}: packed mirror for the path tables (make_node4)
    std::vector<int32_t> item_start;  // per node: first item slot of its internal children (build_up_paths)
    std::vector<uint64_t> tmp_seen;   // one bit per item slot (direct placement of a whole neighbourhood)
    bool last_whole = false;          // the last pass scored a whole neighbourhood: last_v is not filled, v = every internal non-root node twice
    bool pass_pending = false;        // lzg_tree_nni_launch queued a pass whose scores lzg_tree_nni_finish has not reduced yet
    bool node4_valid = false;         // cleared by lzg_tree_set_topology; kept current by lzg_tree_apply
    std::vector<int32_t> tmp_pre, tmp_heads, tmp_items, tmp_count;  // per-pass host scratch, kept to avoid reallocation
    // ---- device-driven climb (lzg_tree_climb.cuh) ----
    int4* d_node4 = nullptr;          // device twin of node4, kept current by k_apply2
    bool d_node4_valid = false;
    void* d_climb = nullptr;          // one block: [items | slot_of | cbase | pair_v | pcost | tail | mlog | queue | tlog]
    size_t climb_bytes = 0;
    int32_t climb_items = 0;          // internal non-root nodes the block was sized for
    int32_t df_epoch = 1;             // epoch of the dataflow UP pass's queue entries (monotonic per tree)
    std::vector<uint64_t> climb_stamps;  // %globaltimer at the end of every pass of the last climb
    float pass0_ms[5] = {0, 0, 0, 0, 0};  // device time of the kernels of the last climb's first pass (queued mode only): UP, dirty scan, items, assemble, arg-min
    bool pass0_timed = false;
};

struct lzg_comm;
struct lzg_climb_result;
int lzg_tree_climb(lzg_ctx* ctx, lzg_comm* comm, lzg_tree* t, int32_t n_edges, const int32_t* edge_order, int32_t greedy, int32_t max_moves,
                   int32_t moves_cap, int32_t* moves, lzg_climb_result* out);
// the all-reduce of the position-sharded pass (lzg_comm.cu): sums vec[0 .. n) over the ranks on ctx->stream; 0 = ok
int lzg_comm_allreduce_sum(lzg_ctx* ctx, lzg_comm* comm, int32_t* d_vec, size_t n);
int lzg_comm_world(const lzg_comm* comm);

int lzg_tree_upload_topology(lzg_ctx* ctx, lzg_tree* t);
int lzg_tree_gather_leaves(lzg_ctx* ctx, lzg_tree* t, const lzg_profiles* p, const int32_t* leaf_row);
int lzg_tree_full_score(lzg_ctx* ctx, lzg_tree* t);
int lzg_tree_nni_launch(lzg_ctx* ctx, lzg_tree* t, int32_t n_moves, const int32_t* moves);
int32_t* lzg_tree_exchange_vector(lzg_tree* t);  // {root score, candidate scores} of the pending pass (device)
int lzg_tree_nni_finish(lzg_ctx* ctx, lzg_tree* t, int greedy, int32_t* best_index, int32_t* best_score, int32_t* all_scores);
int lzg_tree_nni_scores(lzg_ctx* ctx, lzg_tree* t, int32_t n_moves, const int32_t* moves, int greedy,
                        int32_t* best_index, int32_t* best_score, int32_t* all_scores);
int lzg_tree_apply(lzg_ctx* ctx, lzg_tree* t, int u, int w, int v, int z);
int lzg_tree_label(lzg_ctx* ctx, const lzg_tree* t, int32_t* labels, int32_t* branch);
int lzg_tree_download(lzg_ctx* ctx, const lzg_tree* t, int32_t* start, int32_t* end, int32_t* score);
int lzg_tree_download_rows(lzg_ctx* ctx, const lzg_tree* t, int n_ids, const int32_t* ids, int32_t* start, int32_t* end);
