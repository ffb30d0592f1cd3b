// lzg_comm.cu -- the one exchange step of the position-sharded neighbourhood pass, inside the
// library: lzg_nni_best_sharded = lzg_nni_scores_launch + ncclAllReduce(sum, int32) of
// {root score, candidate scores} on the context's stream + lzg_nni_scores_finish, with no host
// code (and no Python) between the three.  NCCL is reached through dlopen of the process's
// libnccl.so.2 (torch's bundled one when the caller is Python, the system's otherwise): the library
// itself has no link-time dependency on it and everything else works without it.
#include <dlfcn.h>

#include <cstring>
#include <new>

#include "lzg_tree.cuh"

namespace {

// the four entry points used, declared here as NCCL's public header declares them (nccl.h 2.x)
struct lzg_nccl_unique_id {
    char internal[128];
};
typedef struct ncclComm* lzg_nccl_comm;
typedef int (*fn_comm_init_rank)(lzg_nccl_comm*, int, lzg_nccl_unique_id, int);
typedef int (*fn_all_reduce)(const void*, void*, size_t, int, int, lzg_nccl_comm, cudaStream_t);
typedef int (*fn_comm_destroy)(lzg_nccl_comm);
typedef const char* (*fn_get_error_string)(int);
enum { LZG_NCCL_INT32 = 2, LZG_NCCL_SUM = 0 };  // ncclInt32, ncclSum

struct lzg_nccl_api {
    void* handle = nullptr;
    fn_comm_init_rank comm_init_rank = nullptr;
    fn_all_reduce all_reduce = nullptr;
    fn_comm_destroy comm_destroy = nullptr;
    fn_get_error_string get_error_string = nullptr;
};

lzg_nccl_api* nccl_api() {
    static lzg_nccl_api api;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (h) {
            api.comm_init_rank = (fn_comm_init_rank)dlsym(h, "ncclCommInitRank");
            api.all_reduce = (fn_all_reduce)dlsym(h, "ncclAllReduce");
            api.comm_destroy = (fn_comm_destroy)dlsym(h, "ncclCommDestroy");
            api.get_error_string = (fn_get_error_string)dlsym(h, "ncclGetErrorString");
            if (api.comm_init_rank && api.all_reduce && api.comm_destroy) api.handle = h;
        }
    }
    return api.handle ? &api : nullptr;
}

}  // namespace

struct lzg_comm {
    lzg_ctx* ctx = nullptr;
    lzg_nccl_comm comm = nullptr;
    int rank = 0, world = 1;
};

int lzg_comm_world(const lzg_comm* c) { return c ? c->world : 1; }

int lzg_comm_allreduce_sum(lzg_ctx* ctx, lzg_comm* c, int32_t* d_vec, size_t n) {
    if (!c || !c->comm) return LZG_OK;
    const int nrc = nccl_api()->all_reduce(d_vec, d_vec, n, LZG_NCCL_INT32, LZG_NCCL_SUM, c->comm, ctx->stream);
    if (nrc != 0) {
        cudaStreamSynchronize(ctx->stream);
        return lzg_fail(ctx, LZG_ERR_STATE, nccl_api()->get_error_string ? nccl_api()->get_error_string(nrc) : "ncclAllReduce failed");
    }
    return LZG_OK;
}

extern "C" {

int lzg_comm_init(lzg_ctx* ctx, int32_t rank, int32_t world, const void* unique_id, lzg_comm** out) {
    if (!ctx || !out || world < 1 || rank < 0 || rank >= world || (world > 1 && !unique_id))
        return lzg_fail(ctx, LZG_ERR_ARG, "lzg_comm_init: bad argument");
    *out = nullptr;
    lzg_comm* c = new (std::nothrow) lzg_comm();
    if (!c) return LZG_ERR_ARG;
    c->ctx = ctx;
    c->rank = rank;
    c->world = world;
    if (world > 1) {
        lzg_nccl_api* api = nccl_api();
        if (!api) {
            delete c;
            return lzg_fail(ctx, LZG_ERR_STATE, "libnccl.so.2 is not available in this process");
        }
        LZG_CUDA(ctx, cudaSetDevice(ctx->device));
        lzg_nccl_unique_id id;
        memcpy(id.internal, unique_id, sizeof(id.internal));
        const int rc = api->comm_init_rank(&c->comm, world, id, rank);
        if (rc != 0) {
            delete c;
            return lzg_fail(ctx, LZG_ERR_STATE, api->get_error_string ? api->get_error_string(rc) : "ncclCommInitRank failed");
        }
    }
    *out = c;
    return LZG_OK;
}

int lzg_comm_free(lzg_comm* c) {
    if (!c) return LZG_ERR_ARG;
    if (c->comm) {
        cudaStreamSynchronize(c->ctx->stream);
        nccl_api()->comm_destroy(c->comm);
    }
    delete c;
    return LZG_OK;
}

int lzg_nni_best_sharded(lzg_ctx* ctx, lzg_comm* c, lzg_tree* t, int32_t n_moves, const int32_t* moves, int32_t greedy,
                         int32_t* best_index, int32_t* best_score, int32_t* all_scores) {
    if (!ctx || !c || c->ctx != ctx || !t || t->ctx != ctx || n_moves < 0 || (n_moves > 0 && !moves) || !best_index || !best_score)
        return lzg_fail(ctx, LZG_ERR_ARG, "lzg_nni_best_sharded: bad argument");
    if (!t->scored) return lzg_fail(ctx, LZG_ERR_STATE, "tree must be scored (lzg_tree_score) before lzg_nni_best_sharded");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = lzg_tree_nni_launch(ctx, t, n_moves, moves);
    if (c->comm && lzg_comm_world(c) > 1) {
        // The ranks agree on the launch status BEFORE the data collective: a rank whose launch failed (not a move of its
        // topology, an allocation failure) must not leave the others waiting in ncclAllReduce.  One int, summed.
        int32_t* d_status = reinterpret_cast<int32_t*>(ctx->d_flag);
        ctx->h_pinned[1] = rc == LZG_OK ? 0 : 1;
        cudaError_t e = cudaMemcpyAsync(d_status, ctx->h_pinned + 1, sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream);
        int nrc = e == cudaSuccess ? nccl_api()->all_reduce(d_status, d_status, 1, LZG_NCCL_INT32, LZG_NCCL_SUM, c->comm, ctx->stream) : 0;
        if (e == cudaSuccess && nrc == 0) e = cudaMemcpyAsync(ctx->h_pinned + 1, d_status, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        const int failed = ctx->h_pinned[1];
        cudaMemsetAsync(ctx->d_flag, 0, sizeof(int), ctx->stream);  // the range-check flag's resting value
        if (e != cudaSuccess || nrc != 0) return lzg_fail(ctx, LZG_ERR_STATE, "lzg_nni_best_sharded: status exchange failed");
        if (rc != LZG_OK) return rc;
        if (failed) {
            if (n_moves > 0) {  // drop the pass this rank queued
                int32_t bi, bs;
                lzg_tree_nni_finish(ctx, t, greedy, &bi, &bs, nullptr);
            }
            return lzg_fail(ctx, LZG_ERR_STATE, "lzg_nni_best_sharded: the pass failed to launch on another rank");
        }
    }
    if (rc != LZG_OK) return rc;
    if (n_moves > 0 && c->comm) {
        int32_t* vec = lzg_tree_exchange_vector(t);
        const int nrc = nccl_api()->all_reduce(vec, vec, (size_t)n_moves + 1, LZG_NCCL_INT32, LZG_NCCL_SUM, c->comm, ctx->stream);
        if (nrc != 0) {
            cudaStreamSynchronize(ctx->stream);
            return lzg_fail(ctx, LZG_ERR_STATE, nccl_api()->get_error_string ? nccl_api()->get_error_string(nrc) : "ncclAllReduce failed");
        }
    }
    return lzg_tree_nni_finish(ctx, t, greedy, best_index, best_score, all_scores);
}

int lzg_hill_climb_sharded(lzg_ctx* ctx, lzg_comm* c, lzg_tree* t, int32_t n_edges, const int32_t* edge_order, int32_t greedy,
                           int32_t max_moves, int32_t moves_cap, int32_t* moves, lzg_climb_result* out) {
    if (!ctx || !c || c->ctx != ctx || !t || t->ctx != ctx || n_edges < 0 || (n_edges > 0 && !edge_order) || moves_cap < 0 || (moves_cap > 0 && !moves))
        return lzg_fail(ctx, LZG_ERR_ARG, "lzg_hill_climb_sharded: bad argument");
    if (!t->scored) return lzg_fail(ctx, LZG_ERR_STATE, "tree must be scored (lzg_tree_score) before lzg_hill_climb_sharded");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    return lzg_tree_climb(ctx, c, t, n_edges, edge_order, greedy, max_moves, moves_cap, moves, out);
}

}  // extern "C"
