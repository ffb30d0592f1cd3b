/* lazac_b200.h -- C ABI of the B200-native ZCNT hot path (liblazac_b200.so).
 *
 * The reference (raphael-group/lazac-copy-number) has no FFI layer; its hot path is reached by
 * direct C++ calls with value semantics.  Each entry point below names the reference call site
 * it replaces (paths relative to the reference repository).  INTEGRATION.md shows the few
 * lines a maintainer of the reference would change to bind them.
 *
 * Conventions
 *   - every function returns an lzg_status (0 = ok); no exception crosses this boundary;
 *   - plain pointers and sizes only; the caller owns every host buffer, the library owns all
 *     device memory behind opaque handles; no host pointer is retained after a call returns;
 *   - one context per (process, GPU); a context is not thread-safe; calls are synchronous on
 *     return unless the name ends in _async;
 *   - breakpoint vectors are int32 [N][L] row-major on the host (rows in the caller's order:
 *     the reference sorts cell names, src/lazac.cxx:69) and int8 (+128 bias) in HBM;
 *   - trees are rooted, binary, node ids 0..n_nodes-1 with root 0 (as the reference assumes,
 *     include/copy_number.hpp:183), child0/child1 = -1 for leaves.
 */
#ifndef LAZAC_B200_H
#define LAZAC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum lzg_status {
    LZG_OK = 0,
    LZG_ERR_ARG = 1,        /* null pointer, bad size, bad handle, id out of range            */
    LZG_ERR_NOT_BINARY = 2, /* std::logic_error("every child must have exactly two children"),
                               src/copy_number.cxx:241-242                                     */
    LZG_ERR_RANGE = 3,      /* a breakpoint value does not fit the int8 device format          */
    LZG_ERR_CUDA = 4,       /* CUDA runtime error; text in lzg_last_error                      */
    LZG_ERR_NO_DEVICE = 5,  /* no usable GPU: there is no CPU fallback                         */
    LZG_ERR_STATE = 6,      /* tree not scored yet / profiles freed / wrong context            */
    LZG_ERR_NOT_A_MOVE = 7  /* (u,w,v,z) is not an NNI move of the current topology            */
} lzg_status;

typedef struct lzg_ctx lzg_ctx;
typedef struct lzg_profiles lzg_profiles;
typedef struct lzg_tree lzg_tree;

/* ---- context ------------------------------------------------------------------------------ */
int lzg_init(int device, lzg_ctx** out);
int lzg_shutdown(lzg_ctx* ctx);
const char* lzg_last_error(const lzg_ctx* ctx); /* valid until the next failing call */
const char* lzg_status_string(int status);
const char* lzg_version(void);
/* Kernel launches issued by this context since lzg_init (bench.py's gpu_launches). */
int lzg_launch_count(const lzg_ctx* ctx, int64_t* out);
/* Device timing on the context's own stream (CUDA events): start, ..., stop -> milliseconds. */
int lzg_timer_start(lzg_ctx* ctx);
int lzg_timer_stop(lzg_ctx* ctx, float* ms);
int lzg_sync(lzg_ctx* ctx);
/* The CUDA stream (cudaStream_t) every call on this context is ordered on. */
int lzg_stream(lzg_ctx* ctx, void** stream);
/* Writes a scratch buffer larger than L2 (timing hygiene between timed iterations). */
int lzg_flush_l2(lzg_ctx* ctx);
/* Pinned host memory for callers that want asynchronous copies to overlap. */
int lzg_host_alloc(lzg_ctx* ctx, uint64_t bytes, void** out);
int lzg_host_free(lzg_ctx* ctx, void* p);

/* ---- profiles: the output of convert_to_breakpoint_profile (src/copy_number.cxx:25-59) for
 *      all N cells, uploaded once.  Replaces the std::map<std::string, breakpoint_profile>
 *      handed to build_distance_matrix (src/lazac.cxx:63) and the leaf start/end vectors set at
 *      src/lazac.cxx:222-226. ------------------------------------------------------------- */
int lzg_profiles_upload(lzg_ctx* ctx, const int32_t* bp, int32_t n_cells, int32_t n_pos,
                        lzg_profiles** out);
/* Same, from a compact int8 host matrix (large synthetic configs). */
int lzg_profiles_upload_i8(lzg_ctx* ctx, const int8_t* bp, int32_t n_cells, int32_t n_pos,
                           lzg_profiles** out);
int lzg_profiles_free(lzg_profiles* p);
int lzg_profiles_shape(const lzg_profiles* p, int32_t* n_cells, int32_t* n_pos);
/* Which kernel the whole-matrix distance of these profiles runs on one GPU (decided at the first distance call from
 * the value range): info[0] = 1 the exact tensor-core product (thermometer code, tcgen05 kind::i8), 0 the VABSDIFF4
 * kernel; info[1] = K = sum over positions of (max - min); info[2] = padded code width; info[3] = tensor-core
 * launches so far; with LZG_TC_TIMING=1 in the environment info[4] / info[5] = microseconds of the code's preparation
 * and of the last product launch (CUDA events; call lzg_sync first), else 0; info[6] = tiles run.  int64 info[8]. */
int lzg_profiles_distance_path(const lzg_profiles* p, int64_t* info);
/* Host only -- no GPU, no context: the column layout of the thermometer code behind the tensor-core path, for tests.
 * lo / hi = lowest / highest value of every position (-128 .. 127).  Column c = 4 w + e of the code belongs to word w;
 * words[4w .. 4w+3] = {word index of the word's 4-byte-aligned base position, selector (nibble e = byte offset of
 * column e's position from the base, 0 .. 7), four threshold bytes (value + 128), 0x01 per valid column}:
 * T[i][c] = [bp[i][position] + 128 >= threshold].  Returns the number of words (entries beyond cap_words are not
 * written), -1 on a bad argument; *K = sum of (hi - lo). */
int64_t lzg_distance_tc_layout(const int32_t* lo, const int32_t* hi, int32_t L, uint32_t* words, int64_t cap_words, int64_t* K);

/* ---- all-pairs ZCNT distance: replaces the pair loop of build_distance_matrix,
 *      src/lazac.cxx:81-91 (operator- include/copy_number.hpp:74-83 + breakpoint_magnitude
 *      src/copy_number.cxx:100-113).  D is the full symmetric N x N int32 matrix. ---------- */
int lzg_distance_matrix(lzg_ctx* ctx, const lzg_profiles* p, int32_t* D_host);
/* Result left in HBM (d_D = device pointer to N*N int32); asynchronous on the context stream. */
int lzg_distance_matrix_device_async(lzg_ctx* ctx, const lzg_profiles* p, int32_t* d_D);
/* One shard of the tile rows (multi-GPU: shard = rank, n_shards = world size).  Tile rows are
 * dealt in pairs (r, R-1-r) so every shard gets the same triangle area.  Only tiles on or above
 * the diagonal of the shard's tile rows are computed; both D[i][j] and D[j][i] are written into
 * the caller's full N x N host matrix, other entries are left untouched. */
int lzg_distance_shard(lzg_ctx* ctx, const lzg_profiles* p, int32_t shard, int32_t n_shards,
                       int32_t* D_host, int64_t* pairs_done);
int lzg_distance_shard_device_async(lzg_ctx* ctx, const lzg_profiles* p, int32_t shard,
                                    int32_t n_shards, int32_t* d_D, int64_t* pairs_done);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * lzg_distance_matrix_device_async -- the whole matrix into the caller's d_D (device, N*N int32), asynchronous
 * on the context stream -- but the call allocates nothing: the kernel's scratch (its work plan: segments,
 * per-CTA offsets, split tiles) goes into a file-scope static device array of
 * lzg_distance_matrix_noalloc_max_bytes() = 1 MiB, which every GPU that loads the library holds until exit.
 * A shape whose plan is larger (beyond roughly N = 40,000 cells) is rejected with LZG_ERR_ARG and a message
 * naming both sizes.  All calls share the array: calls that overlap on the GPU race; serialise them. */
int lzg_distance_matrix_noalloc(lzg_ctx* ctx, const lzg_profiles* p, int32_t* d_D);
int lzg_distance_matrix_noalloc_max_bytes(void);

/* ---- several GPUs of one box, one process: build_distance_matrix ends with the WHOLE matrix in the
 *      caller's buffer (src/lazac.cxx:63-95 returns one matrix; do_nni feeds it to NJ, :153-200).
 * lzg_multi_init opens one context per listed device and enables peer access between them (NVLink /
 * NVSwitch).  lzg_distance_matrix_multi: every GPU uploads 1/G of the profile rows over its own
 * PCIe link and pushes them to its peers, computes its tile rows (dealt in pairs (r, R-1-r)), writes
 * the mirror of every tile straight into the owner's memory, and copies the output rows it owns into
 * D_host -- all GPUs at once.  Pinned host buffers (lzg_host_alloc) make the copies asynchronous. */
typedef struct lzg_multi lzg_multi;
int lzg_multi_init(const int32_t* devices, int32_t n_devices, lzg_multi** out);
int lzg_multi_shutdown(lzg_multi* m);
int lzg_multi_size(const lzg_multi* m);
/* The i-th device's context, for the single-GPU entry points (trees, NJ); owned by m. */
int lzg_multi_context(lzg_multi* m, int32_t i, lzg_ctx** ctx);
const char* lzg_multi_last_error(const lzg_multi* m);
int lzg_distance_matrix_multi(lzg_multi* m, const int32_t* bp, int32_t n_cells, int32_t n_pos, int32_t* D_host,
                              int64_t* pairs_done);
int lzg_distance_matrix_multi_i8(lzg_multi* m, const int8_t* bp, int32_t n_cells, int32_t n_pos, int32_t* D_host,
                                 int64_t* pairs_done);

/* Host only -- no GPU, no context: the work plan lzg_distance_* launches for an n x L input on a
 * device of sm_count SMs (one persistent CTA per SM).  segs receives four ints per
 * segment {tile row, tile column, k0, k1}: 16-byte position chunks [k0, |k1|) of that tile, k1 < 0
 * where the tile is cut across CTAs and accumulated with atomics; CTA b runs segments
 * seg_off[b] .. seg_off[b+1].  info = {CTAs, segments, tile edge, chunks per row} is always filled;
 * LZG_ERR_ARG when cap_segs / cap_ctas are too small (call again with info's sizes).  row_lo/row_hi
 * restrict the tile rows (0, INT32_MAX for all).  The CPU tests check with it that every chunk of
 * every tile of the reference's i <= j loop (src/lazac.cxx:81-91) is covered exactly once. */
int lzg_distance_plan(int32_t n, int32_t L, int32_t sm_count, int32_t shard, int32_t n_shards,
                      int32_t row_lo, int32_t row_hi, int32_t* segs, int64_t cap_segs,
                      int32_t* seg_off, int32_t cap_ctas, int32_t* info);

/* ---- small parsimony on a resident tree: replaces digraph<rectilinear_vertex_data>
 *      (include/copy_number.hpp:98-105) and the functions that walk it. -------------------- */
/* leaf_row[u] = profile row of leaf u, -1 for internal nodes (src/lazac.cxx:212-235). */
int lzg_tree_create(lzg_ctx* ctx, const lzg_profiles* p, int32_t n_nodes, const int32_t* child0,
                    const int32_t* child1, const int32_t* leaf_row, lzg_tree** out);
int lzg_tree_clone(lzg_ctx* ctx, const lzg_tree* t, lzg_tree** out); /* digraph copy-ctor */
int lzg_tree_free(lzg_tree* t);
/* New topology from the host (after stochastic_nni, src/copy_number.cxx:419-462, whose
 * unvisit(t,0) invalidates every interval). */
int lzg_tree_set_topology(lzg_ctx* ctx, lzg_tree* t, const int32_t* child0, const int32_t* child1);
int lzg_tree_get_topology(const lzg_tree* t, int32_t* child0, int32_t* child1);
/* small_rectilinear(t, 0), src/copy_number.cxx:227-277: level-scheduled bottom-up interval
 * DP over every un-scored node; root_score = t[0].data.score. */
int lzg_tree_score(lzg_ctx* ctx, lzg_tree* t, int32_t* root_score);
/* The candidate loop of greedy_nni, src/copy_number.cxx:331-370: scores all n_moves candidates
 * (u,w,v,z) (moves = int32 [n_moves][4], in the host's enumeration order) in one launch and
 * reduces them on the device.  best_index = first candidate with the strictly smallest score
 * below the current root score (greedy != 0: first candidate below it), -1 if none.
 * all_scores (optional, [n_moves]) receives every candidate's root score. */
int lzg_nni_best(lzg_ctx* ctx, lzg_tree* t, int32_t n_moves, const int32_t* moves, int32_t greedy,
                 int32_t* best_index, int32_t* best_score, int32_t* all_scores);
/* The same pass in two halves, for trees whose POSITIONS are sharded over GPUs (every quantity of
 * the pass is a sum over positions; each rank holds a column slice of the profiles and the same
 * topology): _launch queues the UP pass and the candidate scores of this rank's columns and
 * returns, without waiting, the DEVICE vector d_exchange = {root score, score of candidate 0, ...}
 * (n_exchange = n_moves + 1 ints; NULL / 0 for an empty move list); the caller sum-reduces it in
 * place over the ranks on lzg_stream(ctx) (NCCL all-reduce); _finish runs the arg-min of
 * lzg_nni_best on the reduced vector.  lzg_nni_best == _launch + _finish with nothing between. */
int lzg_nni_scores_launch(lzg_ctx* ctx, lzg_tree* t, int32_t n_moves, const int32_t* moves,
                          int32_t** d_exchange, int32_t* n_exchange);
int lzg_nni_scores_finish(lzg_ctx* ctx, lzg_tree* t, int32_t greedy, int32_t* best_index,
                          int32_t* best_score, int32_t* all_scores);
/* The three steps in one call, the all-reduce inside the library: ncclAllReduce(sum, int32) on
 * lzg_stream(ctx) through the process's libnccl.so.2 (loaded with dlopen on first use; the library
 * has no link-time dependency on NCCL).  lzg_comm_init is collective over the ranks: unique_id =
 * the 128 bytes of one ncclGetUniqueId() call, the same on every rank (world == 1: may be NULL,
 * no NCCL).  One communicator serves every tree of the context. */
typedef struct lzg_comm lzg_comm;
int lzg_comm_init(lzg_ctx* ctx, int32_t rank, int32_t world, const void* unique_id, lzg_comm** out);
int lzg_comm_free(lzg_comm* comm);
int lzg_nni_best_sharded(lzg_ctx* ctx, lzg_comm* comm, lzg_tree* t, int32_t n_moves, const int32_t* moves,
                         int32_t greedy, int32_t* best_index, int32_t* best_score, int32_t* all_scores);
/* Score of the LAST candidate of the last lzg_nni_best on this tree (the tree's root score if it
 * had no candidates).  The reference's greedy_nni leaves t[0].data.score at that value when a pass
 * finds nothing (src/copy_number.cxx:351 sets it per candidate, :365-366 do not restore it), and
 * hill_climb returns the tree like that, so do_nni's acceptance test (src/lazac.cxx:287) compares
 * this number, not the tree's true score.  A host that wants the same search trajectory needs it. */
int lzg_tree_last_candidate_score(const lzg_tree* t, int32_t* score);
/* nni(t,u,w,v,z); unvisit(t,0,v); small_rectilinear(t,0) -- src/copy_number.cxx:397,408-409. */
int lzg_tree_apply_nni(lzg_ctx* ctx, lzg_tree* t, int32_t u, int32_t w, int32_t v, int32_t z,
                       int32_t* root_score);
/* hill_climb(t, gen, greedy), src/copy_number.cxx:375-417, with the host out of the loop: every
 * iteration (greedy_nni over the whole neighbourhood, accept the winner, re-score its path) runs on the
 * device from device-resident state; on one GPU the whole climb is ONE CUDA graph launch (a WHILE
 * node repeats the iteration until a pass finds no improving candidate).
 *   edge_order = random_indices of :381-388: a permutation of 0 .. n_edges-1 (n_edges = n_nodes - 1),
 *   index i = the i-th edge of t.edges() (by parent id, then child id) of the tree's CURRENT topology;
 *   the shuffle itself (std::shuffle with the caller's generator) is host logic.  An edge index stays
 *   with its child end while moves are accepted (:401-406), which the library reproduces.
 *   greedy != 0: a pass accepts the first improving candidate in enumeration order (:358-362).
 *   max_moves: stop after that many accepted moves (INT32_MAX: until no candidate improves).
 *   moves (optional): the first moves_cap accepted moves, int32 [moves_cap][5] = (u, w, v, z, score after).
 * The tree ends with the final topology, intervals and scores; lzg_tree_get_topology returns it. */
typedef struct lzg_climb_result {
    int32_t n_moves;         /* accepted moves                                                       */
    int32_t n_passes;        /* neighbourhood passes (= greedy_nni calls)                            */
    int32_t reason;          /* 1 no improving candidate, 2 accepted move did not improve (:412),
                                3 max_moves reached                                                  */
    int32_t true_score;      /* small-parsimony score of the final tree                              */
    int32_t returned_score;  /* t[0].data.score as the reference's hill_climb returns it: the score of
                                the LAST candidate of the final, fruitless pass (see
                                lzg_tree_last_candidate_score); equals true_score unless reason == 1 */
    int64_t candidates;      /* candidates scored (every pass yields the score of every candidate)   */
    int64_t item_units;      /* (work item, chunk group) pairs whose position sums were re-computed: the
                                first pass computes all of them, later passes only those that read a
                                row an accepted move changed                                         */
    int64_t up_units;        /* (node, chunk group) UP rows recomputed by the incremental passes     */
    int64_t units_per_full_pass; /* work items x chunk groups                                        */
} lzg_climb_result;
int lzg_hill_climb(lzg_ctx* ctx, lzg_tree* t, int32_t n_edges, const int32_t* edge_order, int32_t greedy,
                   int32_t max_moves, int32_t moves_cap, int32_t* moves, lzg_climb_result* out);
/* The same climb on a tree whose positions are sharded over the ranks of comm: every rank runs the
 * iteration on its columns, {root score, candidate scores} is sum-reduced with ncclAllReduce on the
 * context's stream between the scoring and the arg-min, and every rank takes the same decisions. */
int lzg_hill_climb_sharded(lzg_ctx* ctx, lzg_comm* comm, lzg_tree* t, int32_t n_edges, const int32_t* edge_order,
                           int32_t greedy, int32_t max_moves, int32_t moves_cap, int32_t* moves, lzg_climb_result* out);
/* Device time stamps (ns, %globaltimer) at the end of every pass of the last climb on this tree:
 * stamps[k] - stamps[k-1] = one iteration (accept + UP pass + all candidates + arg-min). */
int lzg_tree_last_climb_stamps(const lzg_tree* t, int32_t cap, uint64_t* stamps, int32_t* n);
/* Device time (ms) of the five kernels of the last climb's FIRST pass -- the full one: UP pass, dirty scan,
 * all work items, score assembly, arg-min -- measured with CUDA events when the climb ran as queued launches
 * (LZG_CLIMB_GRAPH=0 or a sharded climb); LZG_ERR_STATE when it ran as a graph. */
int lzg_tree_last_climb_pass0_ms(const lzg_tree* t, float* ms5);
/* start/end: int32 [n_nodes][L]; score: int32 [n_nodes] (any may be NULL). */
int lzg_tree_get_intervals(lzg_ctx* ctx, const lzg_tree* t, int32_t* start, int32_t* end,
                           int32_t* score);
/* The same for the listed nodes only: start/end int32 [n_ids][L] (either may be NULL). */
int lzg_tree_get_node_intervals(lzg_ctx* ctx, const lzg_tree* t, int32_t n_ids, const int32_t* ids,
                                int32_t* start, int32_t* end);
/* ancestral_labeling's per-position math, src/copy_number.cxx:165-225: labels int32
 * [n_nodes][L] and in_branch_len int32 [n_nodes] (-1 for the root), indexed by THIS tree's node
 * ids; the pre-order renumbering of :215 is host logic. */
int lzg_tree_ancestral(lzg_ctx* ctx, const lzg_tree* t, int32_t* labels, int32_t* in_branch_len);
/* Work the REFERENCE's path algorithm does for the candidates of the last lzg_nni_best on this
 * tree: merges (= sum over candidates of depth(v)+1) and bytes read (= (merges + n_moves) * 2 * L). */
int lzg_tree_last_nni_work(const lzg_tree* t, int64_t* merges, int64_t* bytes);
/* Algorithmic bytes of the same call as executed here: work items (the two candidates of an edge
 * (u, v) listed back to back are ONE item of 10 L bytes: the intervals of w, both children of v, u
 * and the UP interval of u; 6 L when u is the root) and the UP pass that was run for it
 * (4 L per internal non-root node: read DOWN(sibling), write UP; + 2 L per heavy path for the UP
 * interval its head hangs off; 0 when the UP intervals were still valid). */
int lzg_tree_last_nni_traffic(const lzg_tree* t, int64_t* cand_bytes, int64_t* up_bytes);

/* ---- NJ seed tree ---------------------------------------------------
 * Replaces the file hop + Clearcut run of do_nni, /root/reference/src/lazac.cxx:155-200
 * (NJ_neighbor_joining, src/clearcut.cxx:925): neighbour joining on the integer matrix D (host,
 * N x N, values < 2^24), float32 operation order, tie rule and index bookkeeping as the reference.
 * The caller supplies the taxon shuffle (NJ_shuffle_distance_matrix, src/clearcut.cxx:1644, is
 * MT19937 host logic): position i of the working matrix is taxon perm[i].
 * left/right: int32 [2N-1], -1 for nodes 0 .. N-1 (the leaves, = shuffled positions), internal
 * nodes N .. 2N-2 in creation order, the root is node 2N-2. */
int lzg_nj_seed_tree(lzg_ctx* ctx, const int32_t* D_host, int32_t n, const int32_t* perm, int32_t* left,
                     int32_t* right);
/* The same with the matrix already in HBM (d_D: device pointer to N*N int32, e.g. what
 * lzg_distance_matrix_device_async left there; not modified): the matrix do_nni hands from
 * build_distance_matrix to NJ (src/lazac.cxx:153-200) never crosses PCIe. */
int lzg_nj_seed_tree_device(lzg_ctx* ctx, const int32_t* d_D, int32_t n, const int32_t* perm, int32_t* left,
                            int32_t* right);
/* ... and of the uploaded profiles' own matrix: the device copy lzg_distance_matrix left behind if there is
 * one, else computed now (do_nni: build_distance_matrix -> NJ, src/lazac.cxx:153-200, in one call). */
int lzg_nj_seed_tree_profiles(lzg_ctx* ctx, const lzg_profiles* p, const int32_t* perm, int32_t* left,
                              int32_t* right);

#ifdef __cplusplus
}
#endif
#endif /* LAZAC_B200_H */
