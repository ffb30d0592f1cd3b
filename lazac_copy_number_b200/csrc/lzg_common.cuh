// lzg_common.cuh -- shared definitions of liblazac_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "lazac_b200.h"

#define LZG_BIAS 128     // int8 value v is stored as the byte v + 128 (exact unsigned abs-diff / min / max)
#define LZG_BIAS16 32768  // wider values (|v| <= 32767) as the 16-bit word v + 32768: the second device format

// Persistent-grid work schedule of the distance kernel for one (tile size, shard) (lzg_distance.cu).
struct lzg_schedule {
    int n = 0, l_pad = 0;           // the input shape it was built for (schedules are cached per context)
    int bm = 0, shard = 0, n_shards = 0, row_lo = 0, row_hi = 0, own_mode = 0;
    int n_ctas = 0, n_segs = 0, n_split = 0;
    int64_t pairs = 0;
    void* d_block = nullptr;        // one pooled allocation holding the three arrays below
    size_t bytes = 0;
    void* d_segs = nullptr;         // lzg_seg[n_segs]
    int32_t* d_seg_off = nullptr;   // [n_ctas + 1]
    int2* d_split = nullptr;        // tiles accumulated with atomics (zeroed first)
};

struct lzg_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;
    int64_t launches = 0;
    void* flush_buf = nullptr;
    size_t flush_bytes = 0;
    int* d_flag = nullptr;          // range-check flag
    int32_t* h_pinned = nullptr;    // small pinned staging area (results, path arrays)
    size_t h_pinned_ints = 0;
    int32_t* d_scratch = nullptr;   // small device scratch (path arrays, results)
    size_t d_scratch_ints = 0;
    cudaStream_t copy_stream = nullptr;  // D2H of finished row groups while later groups compute
    cudaEvent_t ev_group = nullptr;
    cudaEvent_t ev_items = nullptr;  // the last k_nni_items launch (reader of the item records) / the item copy on copy_stream
    cudaEvent_t ev_stage = nullptr;  // the last asynchronous copy OUT of the pinned staging area that no synchronising call has waited for
    std::vector<lzg_schedule> schedules;  // distance schedules by (shape, tile, shard); freed at shutdown
    struct climb_graph {                  // one WHILE graph per launch geometry of the climb (lzg_tree.cu)
        int n_items = 0, groups = 0, threads = 0, up_grid = 0, n_cand = 0, width = 1;
        cudaGraph_t graph = nullptr;
        cudaGraphExec_t exec = nullptr;
        cudaGraphConditionalHandle handle = 0;
    };
    std::vector<climb_graph> climb_graphs;
    int climb_graph_broken = 0;           // conditional nodes refused by this driver: batches of queued iterations instead
};

// Stream-ordered device allocation from the device's default memory pool, whose release threshold
// lzg_init raises so that freed blocks stay cached: upload -> distance -> free cycles then cost no
// cudaMalloc/cudaFree.  Everything is ordered on ctx->stream.
inline cudaError_t lzg_dev_alloc(lzg_ctx* ctx, void** out, size_t bytes) {
    return cudaMallocAsync(out, bytes ? bytes : 1, ctx->stream);
}
inline void lzg_dev_free(lzg_ctx* ctx, void* p) {
    if (p) cudaFreeAsync(p, ctx->stream);
}

struct lzg_profiles {
    lzg_ctx* ctx = nullptr;
    int32_t n = 0, L = 0;
    int32_t n_pad = 0, l_pad = 0;  // rows padded to 128, positions padded to 128 (pad value = the bias)
    int32_t width = 1;             // bytes per position: 1 (|v| <= 127, every named configuration) or 2 (|v| <= 32767)
    int32_t pitch = 0;             // bytes per row = l_pad * width
    uint8_t* d = nullptr;          // [n_pad][pitch]
    size_t d_bytes = 0;
    int32_t* d_D = nullptr;        // lazily allocated N x N result used by the host-output entry points
    bool d_D_whole = false;        // d_D holds the whole matrix (set by lzg_distance_matrix; lzg_nj_seed_tree_profiles reads it)
    struct lzg_tc_state* tc = nullptr;  // the tensor-core form of these profiles (lzg_distance_tc.cu), built on first use
};

// Thermometer-coded profiles for the tensor-core distance (lzg_distance_tc.cu).
struct lzg_tc_state {
    int usable = 0;                 // 0: the value range rules the path out (k_zcnt_distance runs)
    int begun = 0;                  // the range kernel is queued, lzg_distance_tc_prepare_finish has not run yet
    int64_t K = 0;                  // sum over positions of (max - min)
    int k_pad = 0, n_pad = 0;       // columns padded to 128, rows to 256
    uint8_t* d_T = nullptr;         // [n_pad][k_pad] bytes 0 / 1
    int32_t* d_s = nullptr;         // [n_pad] row sums of T
    int2* d_tiles = nullptr;        // (row block of 128, column block of 256) of the current launch
    int tiles_cap = 0;
    std::vector<int2> h_tiles;
    std::vector<uint4> table;       // host copy of the word table of k_tc_expand
    alignas(64) unsigned char tmap[128];  // CUtensorMap of T
    int64_t launches = 0;           // k_tc_distance launches so far (tests, bench)
    int64_t tiles_run = 0;          // tiles of those launches
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};  // LZG_TC_TIMING=1: around the preparation and around the last k_tc_distance
};
int lzg_distance_tc_prepare(lzg_ctx* ctx, lzg_profiles* p);  // 1: usable, 0: not for these profiles, < 0: error
int lzg_distance_tc_prepare_begin(lzg_ctx* ctx, lzg_profiles* p);   // 2: queued, call _finish
int lzg_distance_tc_prepare_finish(lzg_ctx* ctx, lzg_profiles* p);
int lzg_distance_tc_run(lzg_ctx* ctx, lzg_profiles* p, int shard, int n_shards, int own_mode, int row_lo, int row_hi, int32_t* d_D, int32_t* const* d_row_base,
                        int64_t* pairs_done);
void lzg_distance_tc_free(lzg_ctx* ctx, lzg_profiles* p);
void lzg_distance_tc_timing(const lzg_profiles* p, int64_t* info);

int lzg_fail(lzg_ctx* ctx, int status, const char* what, cudaError_t e = cudaSuccess);

#define LZG_CUDA(ctx, call)                                               \
    do {                                                                  \
        cudaError_t e__ = (call);                                         \
        if (e__ != cudaSuccess) return lzg_fail((ctx), LZG_ERR_CUDA, #call, e__); \
    } while (0)

#define LZG_LAUNCH_CHECK(ctx)                                             \
    do {                                                                  \
        (ctx)->launches++;                                                \
        cudaError_t e__ = cudaGetLastError();                             \
        if (e__ != cudaSuccess) return lzg_fail((ctx), LZG_ERR_CUDA, "kernel launch", e__); \
    } while (0)

// distance module (lzg_distance.cu)
// d_row_base (optional, device array of one pointer per tile row): where the output rows of each tile row live
// instead of the one N x N matrix d_D (the row blocks of several GPUs, lzg_multi.cu)
int lzg_distance_run(lzg_ctx* ctx, lzg_profiles* p, int shard, int n_shards, int row_lo, int row_hi,
                     int32_t* d_D, int64_t* pairs_done, int32_t* const* d_row_base = nullptr, int own_mode = 0);
// owner of a tile row's output rows under own_mode 1 (snake order): what lzg_multi.cu's phases use
inline int lzg_snake_owner(int r, int n_shards) { const int k = r % n_shards; return (r / n_shards) % 2 == 0 ? k : n_shards - 1 - k; }
int lzg_pack_rows_i32(lzg_ctx* ctx, const int32_t* d_src, uint8_t* d_dst, int rows, int L, int l_pad, int width = 1);
int lzg_fill_bytes(lzg_ctx* ctx, uint8_t* d, size_t n, uint8_t v);
int lzg_fill_pad(lzg_ctx* ctx, uint8_t* d, size_t n_bytes, int width);
int lzg_pack_rows_i8_w2(lzg_ctx* ctx, const int8_t* d_src, uint8_t* d_dst, int rows, int L, int l_pad);  // the padding value of the format (bias) over whole rows
int lzg_bias_rows_i8(lzg_ctx* ctx, uint8_t* d, int n, int n_pad, int L, int l_pad);
// tile edge of the pair kernel for this input: one 64 x 64 tile when n <= 64, else 128 x 128
inline int lzg_distance_tile(const lzg_profiles* p) { return p->n <= 64 ? 64 : 128; }
// pinned / scratch helpers (lzg_api.cu)
int lzg_ensure_scratch(lzg_ctx* ctx, size_t ints);
int lzg_ensure_pinned(lzg_ctx* ctx, size_t ints);  // synchronises the stream when it has to grow
