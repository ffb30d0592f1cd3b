// lzg_api.cu -- the C ABI declared in include/lazac_b200.h.
#include <algorithm>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "lzg_common.cuh"
#include "lzg_tree.cuh"


static thread_local std::string g_init_error;

int lzg_fail(lzg_ctx* ctx, int status, const char* what, cudaError_t e) {
    std::string msg = what ? what : "";
    if (e != cudaSuccess) {
        msg += ": ";
        msg += cudaGetErrorString(e);
        cudaGetLastError();  // clear the sticky-less error state
    }
    if (ctx) ctx->err = msg;
    else g_init_error = msg;
    return status;
}

int lzg_ensure_scratch(lzg_ctx* ctx, size_t ints) {
    if (ints <= ctx->d_scratch_ints) return LZG_OK;
    if (ctx->d_scratch) cudaFree(ctx->d_scratch);
    ctx->d_scratch = nullptr;
    ctx->d_scratch_ints = 0;
    size_t want = std::max<size_t>(ints, 1 << 16);
    LZG_CUDA(ctx, cudaMalloc(&ctx->d_scratch, sizeof(int32_t) * want));
    ctx->d_scratch_ints = want;
    return LZG_OK;
}

int lzg_ensure_pinned(lzg_ctx* ctx, size_t ints) {
    if (ints <= ctx->h_pinned_ints) return LZG_OK;
    LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->h_pinned) cudaFreeHost(ctx->h_pinned);
    ctx->h_pinned = nullptr;
    ctx->h_pinned_ints = 0;
    size_t want = std::max<size_t>(ints * 2, 1 << 16);
    LZG_CUDA(ctx, cudaHostAlloc(&ctx->h_pinned, sizeof(int32_t) * want, cudaHostAllocDefault));
    ctx->h_pinned_ints = want;
    return LZG_OK;
}

extern "C" {

const char* lzg_version(void) { return "lazac_b200 0.1 (sm_100a)"; }

const char* lzg_status_string(int s) {
    switch (s) {
        case LZG_OK: return "ok";
        case LZG_ERR_ARG: return "bad argument";
        case LZG_ERR_NOT_BINARY: return "every child must have exactly two children";
        case LZG_ERR_RANGE: return "breakpoint value outside the device formats (|v| <= 32767)";
        case LZG_ERR_CUDA: return "CUDA error";
        case LZG_ERR_NO_DEVICE: return "no usable GPU (there is no CPU fallback)";
        case LZG_ERR_STATE: return "invalid state";
        case LZG_ERR_NOT_A_MOVE: return "not an NNI move of the current topology";
    }
    return "unknown status";
}

const char* lzg_last_error(const lzg_ctx* ctx) { return ctx ? ctx->err.c_str() : g_init_error.c_str(); }

int lzg_init(int device, lzg_ctx** out) {
    if (!out) return LZG_ERR_ARG;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return lzg_fail(nullptr, LZG_ERR_NO_DEVICE, "cudaGetDeviceCount", e);
    if (device < 0 || device >= count) return lzg_fail(nullptr, LZG_ERR_ARG, "device index out of range");
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return lzg_fail(nullptr, LZG_ERR_NO_DEVICE, "cudaSetDevice", e);
    lzg_ctx* ctx = new (std::nothrow) lzg_ctx();
    if (!ctx) return LZG_ERR_ARG;
    ctx->device = device;
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e == cudaSuccess) ctx->sm_count = prop.multiProcessorCount;
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_group, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_stage, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_items, cudaEventDisableTiming);
    if (e == cudaSuccess) {
        // keep freed blocks in the device's default pool (see lzg_dev_alloc)
        cudaMemPool_t pool;
        uint64_t keep = UINT64_MAX;
        e = cudaDeviceGetDefaultMemPool(&pool, device);
        if (e == cudaSuccess) e = cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev1);
    if (e == cudaSuccess) e = cudaMalloc(&ctx->d_flag, sizeof(int));
    if (e == cudaSuccess) e = cudaMemset(ctx->d_flag, 0, sizeof(int));
    ctx->h_pinned_ints = 1 << 16;
    if (e == cudaSuccess) e = cudaHostAlloc(&ctx->h_pinned, sizeof(int32_t) * ctx->h_pinned_ints, cudaHostAllocDefault);
    if (e == cudaSuccess && lzg_ensure_scratch(ctx, 1 << 16) != LZG_OK) e = cudaErrorMemoryAllocation;
    if (e != cudaSuccess) {
        int rc = lzg_fail(nullptr, LZG_ERR_CUDA, "lzg_init", e);
        delete ctx;
        return rc;
    }
    *out = ctx;
    return LZG_OK;
}

int lzg_shutdown(lzg_ctx* ctx) {
    if (!ctx) return LZG_ERR_ARG;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto& sc : ctx->schedules) lzg_dev_free(ctx, sc.d_block);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->flush_buf) cudaFree(ctx->flush_buf);
    if (ctx->d_flag) cudaFree(ctx->d_flag);
    if (ctx->d_scratch) cudaFree(ctx->d_scratch);
    if (ctx->h_pinned) cudaFreeHost(ctx->h_pinned);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->ev_group) cudaEventDestroy(ctx->ev_group);
    if (ctx->ev_stage) cudaEventDestroy(ctx->ev_stage);
    if (ctx->ev_items) cudaEventDestroy(ctx->ev_items);
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return LZG_OK;
}

int lzg_launch_count(const lzg_ctx* ctx, int64_t* out) {
    if (!ctx || !out) return LZG_ERR_ARG;
    *out = ctx->launches;
    return LZG_OK;
}

int lzg_timer_start(lzg_ctx* ctx) {
    if (!ctx) return LZG_ERR_ARG;
    LZG_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    return LZG_OK;
}

int lzg_timer_stop(lzg_ctx* ctx, float* ms) {
    if (!ctx || !ms) return LZG_ERR_ARG;
    LZG_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    LZG_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
    LZG_CUDA(ctx, cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
    return LZG_OK;
}

int lzg_sync(lzg_ctx* ctx) {
    if (!ctx) return LZG_ERR_ARG;
    LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LZG_OK;
}

int lzg_stream(lzg_ctx* ctx, void** stream) {
    if (!ctx || !stream) return LZG_ERR_ARG;
    *stream = (void*)ctx->stream;
    return LZG_OK;
}

int lzg_flush_l2(lzg_ctx* ctx) {
    if (!ctx) return LZG_ERR_ARG;
    if (!ctx->flush_buf) {
        ctx->flush_bytes = (size_t)256 << 20;  // 2x the 126 MB L2
        LZG_CUDA(ctx, cudaMalloc(&ctx->flush_buf, ctx->flush_bytes));
    }
    return lzg_fill_bytes(ctx, (uint8_t*)ctx->flush_buf, ctx->flush_bytes, 0);
}

int lzg_host_alloc(lzg_ctx* ctx, uint64_t bytes, void** out) {
    if (!ctx || !out) return LZG_ERR_ARG;
    LZG_CUDA(ctx, cudaHostAlloc(out, bytes, cudaHostAllocPortable));  // pinned for every GPU of the process (lzg_multi)
    return LZG_OK;
}

int lzg_host_free(lzg_ctx* ctx, void* p) {
    if (!ctx) return LZG_ERR_ARG;
    LZG_CUDA(ctx, cudaFreeHost(p));
    return LZG_OK;
}

// ---- profiles ----------------------------------------------------------------------------------
// Two device formats: one byte per position (v + 128) when every |v| <= 127 -- all named
// configurations -- else 16 bits (v + 32768) up to |v| <= 32767; beyond that LZG_ERR_RANGE.  The 8-bit form is
// tried first; the range flag of its packing kernel decides.
// int8 input: rows are copied straight into the padded device layout (one pitched H2D copy) and
// biased in place.  int32 input: staged through a bounded device buffer (<= 256 MiB) and packed.
static int profiles_upload_width(lzg_ctx* ctx, const void* host, bool is_i8, int32_t n, int32_t L, int width, lzg_profiles** out, bool* out_of_range) {
    *out = nullptr;
    *out_of_range = false;
    lzg_profiles* p = new (std::nothrow) lzg_profiles();
    if (!p) return LZG_ERR_ARG;
    p->ctx = ctx;
    p->n = n;
    p->L = L;
    p->n_pad = std::max(128, (n + 127) / 128 * 128);
    p->l_pad = std::max(128, (L + 127) / 128 * 128);
    p->width = width;
    p->pitch = p->l_pad * width;
    p->d_bytes = (size_t)p->n_pad * p->pitch;
    cudaError_t e = lzg_dev_alloc(ctx, (void**)&p->d, p->d_bytes);
    if (e != cudaSuccess) {
        delete p;
        return lzg_fail(ctx, LZG_ERR_CUDA, "profiles allocation", e);
    }
    int rc = LZG_OK;
    e = cudaMemsetAsync(ctx->d_flag, 0, sizeof(int), ctx->stream);
    if (e != cudaSuccess) rc = lzg_fail(ctx, LZG_ERR_CUDA, "flag reset", e);
    void* d_stage = nullptr;
    if (rc == LZG_OK && is_i8 && width == 1) {
        if (n > 0 && L > 0) {
            e = cudaMemcpy2DAsync(p->d, p->pitch, host, L, L, n, cudaMemcpyHostToDevice, ctx->stream);
            if (e != cudaSuccess) rc = lzg_fail(ctx, LZG_ERR_CUDA, "profiles H2D", e);
        }
        if (rc == LZG_OK) rc = lzg_bias_rows_i8(ctx, p->d, L > 0 ? n : 0, p->n_pad, L, p->l_pad);
    } else if (rc == LZG_OK) {
        rc = lzg_fill_pad(ctx, p->d + (size_t)n * p->pitch, p->d_bytes - (size_t)n * p->pitch, width);
        if (rc == LZG_OK && n > 0 && L > 0) {
            const size_t row_bytes = (size_t)L * (is_i8 ? 1 : 4);
            int block_rows = (int)std::max<size_t>(1, std::min<size_t>(n, ((size_t)256 << 20) / row_bytes));
            e = lzg_dev_alloc(ctx, &d_stage, (size_t)block_rows * row_bytes);
            if (e != cudaSuccess) rc = lzg_fail(ctx, LZG_ERR_CUDA, "staging allocation", e);
            for (int r0 = 0; rc == LZG_OK && r0 < n; r0 += block_rows) {
                int nr = std::min(block_rows, n - r0);
                e = cudaMemcpyAsync(d_stage, (const uint8_t*)host + (size_t)r0 * row_bytes, (size_t)nr * row_bytes, cudaMemcpyHostToDevice, ctx->stream);
                if (e != cudaSuccess) {
                    rc = lzg_fail(ctx, LZG_ERR_CUDA, "profiles H2D", e);
                    break;
                }
                if (is_i8) rc = lzg_pack_rows_i8_w2(ctx, (const int8_t*)d_stage, p->d + (size_t)r0 * p->pitch, nr, L, p->l_pad);
                else rc = lzg_pack_rows_i32(ctx, (const int32_t*)d_stage, p->d + (size_t)r0 * p->pitch, nr, L, p->l_pad, width);
            }
        } else if (rc == LZG_OK && n > 0) {
            rc = lzg_fill_pad(ctx, p->d, (size_t)n * p->pitch, width);
        }
    }
    if (rc == LZG_OK) {
        e = cudaMemcpyAsync(ctx->h_pinned, ctx->d_flag, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);  // the host buffer is not retained past this call
        if (e != cudaSuccess) rc = lzg_fail(ctx, LZG_ERR_CUDA, "profiles pack", e);
        else if (n > 0 && L > 0 && ctx->h_pinned[0]) *out_of_range = true;
    }
    lzg_dev_free(ctx, d_stage);
    if (rc != LZG_OK || *out_of_range) {
        lzg_dev_free(ctx, p->d);
        delete p;
        return rc;
    }
    *out = p;
    return LZG_OK;
}

static int profiles_upload_any(lzg_ctx* ctx, const void* host, bool is_i8, int32_t n, int32_t L, lzg_profiles** out) {
    if (!ctx || !out || n < 0 || L < 0 || ((n > 0 && L > 0) && !host)) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_profiles_upload: bad argument");
    *out = nullptr;
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    const int forced = getenv("LZG_FORCE_WIDTH") ? atoi(getenv("LZG_FORCE_WIDTH")) : 0;  // tests: the 16-bit kernels on small values
    for (int width = forced == 2 ? 2 : 1; width <= 2; width++) {
        bool out_of_range = false;
        const int rc = profiles_upload_width(ctx, host, is_i8, n, L, width, out, &out_of_range);
        if (rc != LZG_OK) return rc;
        if (!out_of_range) return LZG_OK;
    }
    return lzg_fail(ctx, LZG_ERR_RANGE, "breakpoint value outside [-32767, 32767]");
}

int lzg_profiles_upload(lzg_ctx* ctx, const int32_t* bp, int32_t n, int32_t L, lzg_profiles** out) {
    return profiles_upload_any(ctx, bp, false, n, L, out);
}

int lzg_profiles_upload_i8(lzg_ctx* ctx, const int8_t* bp, int32_t n, int32_t L, lzg_profiles** out) {
    return profiles_upload_any(ctx, bp, true, n, L, out);
}

int lzg_profiles_free(lzg_profiles* p) {
    if (!p) return LZG_ERR_ARG;
    lzg_ctx* ctx = p->ctx;
    cudaSetDevice(ctx->device);
    // stream-ordered frees: no synchronisation, the blocks go back to the pool
    lzg_dev_free(ctx, p->d_D);
    lzg_distance_tc_free(ctx, p);
    lzg_dev_free(ctx, p->d);
    delete p;
    return LZG_OK;
}

// Which kernel lzg_distance_* runs for these profiles' whole matrix on one GPU: info[0] = 1 the tensor-core product
// (k_tc_distance), 0 k_zcnt_distance; info[1] = K, the summed value range of the positions (0 before the first distance call
// and for the 16-bit format); info[2] = the thermometer code's padded width; info[3] = k_tc_distance launches so far;
// info[4], info[5] = microseconds of the preparation and of the last k_tc_distance launch (LZG_TC_TIMING=1, else 0); info[6] = tiles run.
int lzg_profiles_distance_path(const lzg_profiles* p, int64_t* info) {
    if (!p || !info) return LZG_ERR_ARG;
    info[0] = info[1] = info[2] = info[3] = 0;
    if (p->tc) info[0] = p->tc->usable, info[1] = p->tc->K, info[2] = p->tc->k_pad, info[3] = p->tc->launches;
    lzg_distance_tc_timing(p, info);
    return LZG_OK;
}

int lzg_profiles_shape(const lzg_profiles* p, int32_t* n, int32_t* L) {
    if (!p) return LZG_ERR_ARG;
    if (n) *n = p->n;
    if (L) *L = p->L;
    return LZG_OK;
}

// ---- distance ----------------------------------------------------------------------------------
int lzg_distance_shard_device_async(lzg_ctx* ctx, const lzg_profiles* p, int32_t shard, int32_t n_shards, int32_t* d_D, int64_t* pairs_done) {
    if (!ctx || !p || p->ctx != ctx || n_shards < 1 || shard < 0 || shard >= n_shards || (p->n > 0 && !d_D))
        return lzg_fail(ctx, LZG_ERR_ARG, "lzg_distance: bad argument");
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    // the thermometer code of the tensor-core path is part of every call's work, not a cache across calls
    lzg_distance_tc_free(ctx, const_cast<lzg_profiles*>(p));
    return lzg_distance_run(ctx, const_cast<lzg_profiles*>(p), shard, n_shards, 0, INT32_MAX, d_D, pairs_done);
}

int lzg_distance_matrix_device_async(lzg_ctx* ctx, const lzg_profiles* p, int32_t* d_D) {
    return lzg_distance_shard_device_async(ctx, p, 0, 1, d_D, nullptr);
}

static int ensure_dD(lzg_ctx* ctx, lzg_profiles* p) {
    if (p->d_D || p->n == 0) return LZG_OK;
    LZG_CUDA(ctx, lzg_dev_alloc(ctx, (void**)&p->d_D, sizeof(int32_t) * (size_t)p->n * p->n));
    return LZG_OK;
}

// Whole matrix to the host.  Large matrices are computed in groups of tile rows; rows of a group
// are final once the group is done (mirrors of earlier groups are already in place), so their
// device-to-host copy runs on the copy stream while the next group computes.
int lzg_distance_matrix(lzg_ctx* ctx, const lzg_profiles* cp, int32_t* D) {
    if (!ctx || !cp || cp->ctx != ctx || (cp->n > 0 && !D)) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_distance_matrix: bad argument");
    lzg_profiles* p = const_cast<lzg_profiles*>(cp);
    if (p->n == 0) return LZG_OK;
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = ensure_dD(ctx, p);
    if (rc != LZG_OK) return rc;
    lzg_distance_tc_free(ctx, p);  // rebuilt by the first row group below: part of this call's work
    const int n = p->n, bm = lzg_distance_tile(p), R = (n + bm - 1) / bm;
    const size_t bytes = sizeof(int32_t) * (size_t)n * n;
    const int groups = bytes < ((size_t)64 << 20) ? 1 : std::min(16, R);
    if (groups == 1) {
        rc = lzg_distance_run(ctx, p, 0, 1, 0, INT32_MAX, p->d_D, nullptr);
        if (rc != LZG_OK) return rc;
        LZG_CUDA(ctx, cudaMemcpyAsync(D, p->d_D, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        p->d_D_whole = true;
        return LZG_OK;
    }
    p->d_D_whole = false;
    for (int g = 0; g < groups; g++) {
        const int ra = (int)((int64_t)R * g / groups), rb = (int)((int64_t)R * (g + 1) / groups);
        if (ra == rb) continue;
        rc = lzg_distance_run(ctx, p, 0, 1, ra, rb, p->d_D, nullptr);
        if (rc != LZG_OK) break;
        const size_t r0 = (size_t)ra * bm, r1 = std::min<size_t>(n, (size_t)rb * bm);
        cudaError_t e = cudaEventRecord(ctx->ev_group, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_group, 0);
        if (e == cudaSuccess) e = cudaMemcpyAsync(D + r0 * n, p->d_D + r0 * n, sizeof(int32_t) * (r1 - r0) * n, cudaMemcpyDeviceToHost, ctx->copy_stream);
        if (e != cudaSuccess) {
            rc = lzg_fail(ctx, LZG_ERR_CUDA, "row-group D2H", e);
            break;
        }
    }
    cudaError_t e1 = cudaStreamSynchronize(ctx->copy_stream), e2 = cudaStreamSynchronize(ctx->stream);
    if (rc == LZG_OK && (e1 != cudaSuccess || e2 != cudaSuccess)) rc = lzg_fail(ctx, LZG_ERR_CUDA, "lzg_distance_matrix", e1 != cudaSuccess ? e1 : e2);
    p->d_D_whole = rc == LZG_OK;
    return rc;
}

// NJ seed tree of the profiles' own distance matrix without the matrix leaving HBM: the copy lzg_distance_matrix left on the
// device if there is one, else computed now.
int lzg_nj_seed_tree_profiles(lzg_ctx* ctx, const lzg_profiles* cp, const int32_t* perm, int32_t* left, int32_t* right) {
    if (!ctx || !cp || cp->ctx != ctx || cp->n < 1) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_nj_seed_tree_profiles: bad argument");
    lzg_profiles* p = const_cast<lzg_profiles*>(cp);
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = ensure_dD(ctx, p);
    if (rc != LZG_OK) return rc;
    if (!p->d_D_whole) {
        lzg_distance_tc_free(ctx, p);
        rc = lzg_distance_run(ctx, p, 0, 1, 0, INT32_MAX, p->d_D, nullptr);
        if (rc != LZG_OK) return rc;
        p->d_D_whole = true;
    }
    return lzg_nj_seed_tree_device(ctx, p->d_D, p->n, perm, left, right);
}

int lzg_distance_shard(lzg_ctx* ctx, const lzg_profiles* cp, int32_t shard, int32_t n_shards, int32_t* D, int64_t* pairs_done) {
    if (!ctx || !cp || cp->ctx != ctx || n_shards < 1 || shard < 0 || shard >= n_shards || (cp->n > 0 && !D))
        return lzg_fail(ctx, LZG_ERR_ARG, "lzg_distance_shard: bad argument");
    lzg_profiles* p = const_cast<lzg_profiles*>(cp);
    if (pairs_done) *pairs_done = 0;
    if (p->n == 0) return LZG_OK;
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = ensure_dD(ctx, p);
    if (rc != LZG_OK) return rc;
    p->d_D_whole = false;
    lzg_distance_tc_free(ctx, p);
    rc = lzg_distance_run(ctx, p, shard, n_shards, 0, INT32_MAX, p->d_D, pairs_done);
    if (rc != LZG_OK) return rc;
    // copy back the strips this shard owns: for tile row r the row strip [r*bm, ..) x [r*bm, n) and
    // its mirror, the column strip [r*bm, n) x [r*bm, ..)
    const int bm = lzg_distance_tile(p), n = p->n, R = (n + bm - 1) / bm;
    const size_t pitch = sizeof(int32_t) * (size_t)n;
    for (int r = 0; r < R; r++) {
        if (std::min(r, R - 1 - r) % n_shards != shard) continue;
        int r0 = r * bm, nr = std::min(n, r0 + bm) - r0;
        size_t o_row = (size_t)r0 * n + r0;
        LZG_CUDA(ctx, cudaMemcpy2DAsync(D + o_row, pitch, p->d_D + o_row, pitch, sizeof(int32_t) * (size_t)(n - r0), nr, cudaMemcpyDeviceToHost, ctx->stream));
        if (r0 + nr < n) {
            size_t o_col = (size_t)(r0 + nr) * n + r0;
            LZG_CUDA(ctx, cudaMemcpy2DAsync(D + o_col, pitch, p->d_D + o_col, pitch, sizeof(int32_t) * (size_t)nr, n - r0 - nr, cudaMemcpyDeviceToHost, ctx->stream));
        }
    }
    LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return LZG_OK;
}

}  // extern "C"
