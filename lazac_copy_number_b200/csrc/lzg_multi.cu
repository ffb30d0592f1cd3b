// lzg_multi.cu -- build_distance_matrix (/root/reference/src/lazac.cxx:63-95) on several GPUs of one box,
// from one process, ending with the whole N x N matrix in the caller's host buffer.
//
// The reference returns one matrix; a rank-per-process sharding leaves every rank with its own strips
// and needs a gather between processes.  Here one host thread drives one context per GPU:
//   1. profiles: GPU g receives rows [g n / G, (g+1) n / G) over ITS OWN PCIe link, packs them (bias /
//      padding) and pushes the packed block to every peer over NVLink (cudaMemcpyPeerAsync) -- an
//      all-gather that replaces G full uploads through host memory;
//   2. tiles: tile rows are dealt in snake order (0 .. G-1, G-1 .. 0, ...): equal triangle area per GPU over
//      every 2 G consecutive tile rows, no communication during the compute;
//   3. output: GPU g OWNS the output rows of its tile rows.  The kernel writes a tile into the owner's
//      row block and the MIRROR of the tile straight into the row block of the tile column's owner --
//      a peer store over NVLink from the epilogue (a table of one base pointer per tile row) -- so when
//      all kernels are done every GPU holds whole, final rows of D;
//   4. gather: every GPU copies its row blocks to their place in the host matrix over its own PCIe link,
//      all GPUs at once (10 MB contiguous per tile row at cfg3) -- in phases of tile rows, so that the
//      copies of one phase run while the next one computes.
// Nothing here computes: the kernels are lzg_distance.cu's.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "lzg_common.cuh"

struct lzg_multi {
    std::vector<lzg_ctx*> ctx;
    std::vector<uint8_t*> d_prof;       // per GPU: packed profiles [n_pad][l_pad]
    std::vector<size_t> prof_cap;
    std::vector<int32_t*> d_rows;       // per GPU: the output rows it owns, tile row after tile row
    std::vector<size_t> rows_cap;
    std::vector<int32_t**> d_table;     // per GPU: base pointer of every tile row's output rows
    std::vector<size_t> table_cap;
    std::vector<void*> d_stage;         // per GPU: int32 staging of an upload block
    std::vector<size_t> stage_cap;
    std::vector<cudaEvent_t> ev_up, ev_done;
    std::string err;
};

namespace {

// the entry points visit every GPU; the caller's current device is put back when they return
struct device_guard {
    int dev = -1;
    device_guard() { if (cudaGetDevice(&dev) != cudaSuccess) dev = -1; }
    ~device_guard() { if (dev >= 0) cudaSetDevice(dev); }
};

int multi_fail(lzg_multi* m, int status, const char* what, cudaError_t e = cudaSuccess) {
    std::string msg = what ? what : "";
    if (e != cudaSuccess) {
        msg += ": ";
        msg += cudaGetErrorString(e);
        cudaGetLastError();
    }
    if (m) m->err = msg;
    return status;
}

#define LZM_CUDA(m, call)                                                   \
    do {                                                                    \
        cudaError_t e__ = (call);                                           \
        if (e__ != cudaSuccess) return multi_fail((m), LZG_ERR_CUDA, #call, e__); \
    } while (0)

template <typename T>
int grow(lzg_multi* m, T** p, size_t* cap, size_t bytes) {
    if (bytes <= *cap && *p) return LZG_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    LZM_CUDA(m, cudaMalloc((void**)p, std::max<size_t>(bytes, 256)));  // plain cudaMalloc: peers may map it
    *cap = bytes;
    return LZG_OK;
}

}  // namespace

extern "C" {

int lzg_multi_init(const int32_t* devices, int32_t n_devices, lzg_multi** out) {
    if (!out || n_devices < 1 || !devices) return LZG_ERR_ARG;
    *out = nullptr;
    device_guard restore;
    lzg_multi* m = new (std::nothrow) lzg_multi();
    if (!m) return LZG_ERR_ARG;
    for (int g = 0; g < n_devices; g++) {
        for (int h = 0; h < g; h++)
            if (devices[h] == devices[g]) {
                for (auto* c : m->ctx) lzg_shutdown(c);
                delete m;
                return LZG_ERR_ARG;
            }
        lzg_ctx* c = nullptr;
        const int rc = lzg_init(devices[g], &c);
        if (rc != LZG_OK) {
            for (auto* q : m->ctx) lzg_shutdown(q);
            delete m;
            return rc;
        }
        m->ctx.push_back(c);
    }
    const int G = n_devices;
    m->d_prof.assign(G, nullptr), m->prof_cap.assign(G, 0);
    m->d_rows.assign(G, nullptr), m->rows_cap.assign(G, 0);
    m->d_table.assign(G, nullptr), m->table_cap.assign(G, 0);
    m->d_stage.assign(G, nullptr), m->stage_cap.assign(G, 0);
    m->ev_up.assign(G, nullptr), m->ev_done.assign(G, nullptr);
    int rc = LZG_OK;
    for (int g = 0; g < G && rc == LZG_OK; g++) {
        cudaError_t e = cudaSetDevice(devices[g]);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev_up[g], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev_done[g], cudaEventDisableTiming);
        for (int h = 0; h < G && e == cudaSuccess; h++) {
            if (h == g) continue;
            int can = 0;
            e = cudaDeviceCanAccessPeer(&can, devices[g], devices[h]);
            if (e == cudaSuccess && !can) {
                rc = LZG_ERR_STATE;  // no NVLink / PCIe peer path between two of the GPUs: the mirror stores need one
                break;
            }
            if (e == cudaSuccess) {
                e = cudaDeviceEnablePeerAccess(devices[h], 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) {
                    cudaGetLastError();
                    e = cudaSuccess;
                }
            }
        }
        if (e != cudaSuccess) rc = LZG_ERR_CUDA;
    }
    if (rc != LZG_OK) {
        lzg_multi_shutdown(m);
        return rc;
    }
    *out = m;
    return LZG_OK;
}

int lzg_multi_shutdown(lzg_multi* m) {
    if (!m) return LZG_ERR_ARG;
    device_guard restore;
    for (size_t g = 0; g < m->ctx.size(); g++) {
        cudaSetDevice(m->ctx[g]->device);
        cudaStreamSynchronize(m->ctx[g]->stream);
        cudaStreamSynchronize(m->ctx[g]->copy_stream);
        if (m->d_prof[g]) cudaFree(m->d_prof[g]);
        if (m->d_rows[g]) cudaFree(m->d_rows[g]);
        if (m->d_table[g]) cudaFree(m->d_table[g]);
        if (m->d_stage[g]) cudaFree(m->d_stage[g]);
        if (m->ev_up[g]) cudaEventDestroy(m->ev_up[g]);
        if (m->ev_done[g]) cudaEventDestroy(m->ev_done[g]);
    }
    for (auto* c : m->ctx) lzg_shutdown(c);
    delete m;
    return LZG_OK;
}

int lzg_multi_size(const lzg_multi* m) { return m ? (int)m->ctx.size() : 0; }

int lzg_multi_context(lzg_multi* m, int32_t i, lzg_ctx** ctx) {
    if (!m || !ctx || i < 0 || i >= (int)m->ctx.size()) return LZG_ERR_ARG;
    *ctx = m->ctx[i];
    return LZG_OK;
}

const char* lzg_multi_last_error(const lzg_multi* m) { return m ? m->err.c_str() : ""; }

// One attempt in one device format (width = bytes per position); *out_of_range tells the caller to try the wider one.
static int distance_multi_width(lzg_multi* m, const void* host, bool is_i8, int32_t n, int32_t L, int32_t* D_host, int64_t* pairs_out, int width,
                                bool* out_of_range_) {
    *out_of_range_ = false;
    const int G = (int)m->ctx.size();
    const int n_pad = std::max(128, (n + 127) / 128 * 128), l_pad = std::max(128, (L + 127) / 128 * 128);
    const int pitch = l_pad * width;
    lzg_profiles shape;  // for lzg_distance_tile / the schedule cache
    shape.n = n, shape.L = L, shape.n_pad = n_pad, shape.l_pad = l_pad, shape.width = width, shape.pitch = pitch;
    const int bm = lzg_distance_tile(&shape), R = (n + bm - 1) / bm;
    auto owner = [&](int r) { return lzg_snake_owner(r, G); };  // equal triangle area over every 2 G consecutive tile rows
    // output rows: slot of every tile row inside its owner's block
    std::vector<int> slot(R, 0), owned(G, 0);
    for (int r = 0; r < R; r++) slot[r] = owned[owner(r)]++;
    const size_t tile_row_ints = (size_t)bm * n;
    for (int g = 0; g < G; g++) {
        LZM_CUDA(m, cudaSetDevice(m->ctx[g]->device));
        int rc = grow(m, &m->d_prof[g], &m->prof_cap[g], (size_t)n_pad * pitch);
        if (rc == LZG_OK) rc = grow(m, &m->d_rows[g], &m->rows_cap[g], sizeof(int32_t) * tile_row_ints * std::max(1, owned[g]));
        if (rc == LZG_OK) rc = grow(m, &m->d_table[g], &m->table_cap[g], sizeof(int32_t*) * (size_t)R);
        if (rc != LZG_OK) return rc;
    }
    std::vector<int32_t*> table(R);
    for (int r = 0; r < R; r++) table[r] = m->d_rows[owner(r)] + tile_row_ints * slot[r];
    // ---- 1. upload: one row block per GPU, packed there, pushed to the peers ----
    const int blk = ((n_pad / 128 + G - 1) / G) * 128;  // rows per block, a multiple of 128
    int rc = LZG_OK;
    for (int g = 0; g < G && rc == LZG_OK; g++) {
        lzg_ctx* c = m->ctx[g];
        LZM_CUDA(m, cudaSetDevice(c->device));
        LZM_CUDA(m, cudaMemcpyAsync(m->d_table[g], table.data(), sizeof(int32_t*) * (size_t)R, cudaMemcpyHostToDevice, c->stream));
        LZM_CUDA(m, cudaMemsetAsync(c->d_flag, 0, sizeof(int), c->stream));
        const int r0 = std::min(n_pad, g * blk), r1 = std::min(n_pad, r0 + blk), rows = r1 - r0;
        const int valid = std::max(0, std::min(n, r1) - r0);
        uint8_t* dst = m->d_prof[g] + (size_t)r0 * pitch;
        if (rows > 0) {
            if (is_i8 && width == 1) {
                if (valid > 0 && L > 0)
                    LZM_CUDA(m, cudaMemcpy2DAsync(dst, l_pad, (const int8_t*)host + (size_t)r0 * L, L, L, valid, cudaMemcpyHostToDevice, c->stream));
                rc = lzg_bias_rows_i8(c, dst, L > 0 ? valid : 0, rows, L, l_pad);
            } else {
                rc = lzg_fill_pad(c, dst + (size_t)valid * pitch, (size_t)(rows - valid) * pitch, width);
                if (rc == LZG_OK && valid > 0 && L > 0) {
                    const size_t row_bytes = (size_t)L * (is_i8 ? 1 : 4);
                    const int step = (int)std::max<size_t>(1, std::min<size_t>(valid, ((size_t)128 << 20) / row_bytes));
                    rc = grow(m, &m->d_stage[g], &m->stage_cap[g], (size_t)step * row_bytes);
                    for (int q = 0; rc == LZG_OK && q < valid; q += step) {
                        const int nr = std::min(step, valid - q);
                        LZM_CUDA(m, cudaMemcpyAsync(m->d_stage[g], (const uint8_t*)host + (size_t)(r0 + q) * row_bytes, (size_t)nr * row_bytes, cudaMemcpyHostToDevice, c->stream));
                        if (is_i8) rc = lzg_pack_rows_i8_w2(c, (const int8_t*)m->d_stage[g], dst + (size_t)q * pitch, nr, L, l_pad);
                        else rc = lzg_pack_rows_i32(c, (const int32_t*)m->d_stage[g], dst + (size_t)q * pitch, nr, L, l_pad, width);
                    }
                } else if (rc == LZG_OK && valid > 0) {
                    rc = lzg_fill_pad(c, dst, (size_t)valid * pitch, width);
                }
            }
            if (rc != LZG_OK) {
                m->err = c->err;
                break;
            }
            for (int h = 0; h < G; h++) {
                if (h == g) continue;
                LZM_CUDA(m, cudaMemcpyPeerAsync(m->d_prof[h] + (size_t)r0 * pitch, m->ctx[h]->device, dst, c->device, (size_t)rows * pitch, c->stream));
            }
        }
        LZM_CUDA(m, cudaEventRecord(m->ev_up[g], c->stream));
    }
    // ---- 2. - 4. in PHASES of tile rows: every GPU computes its tile rows of the phase (the epilogue writes into the owners'
    // row blocks, mirrors over NVLink); once all GPUs are through a phase the output rows of its tile rows are final -- they
    // only receive mirrors from tile rows above them -- and every GPU copies the ones it owns to the host matrix while the
    // next phase computes.  Phase boundaries are multiples of 2 G tile rows (the snake deal is balanced there); as many phases
    // as leave a few tiles per CTA.
    int64_t pairs = 0;
    const int blocks2g = (R + 2 * G - 1) / (2 * G);
    const long long tiles_per_gpu = (long long)R * (R + 1) / 2 / G;
    const int sm = m->ctx[0]->sm_count;
    int phases = (int)std::max<long long>(1, std::min<long long>(G >= 4 ? 16 : 8, tiles_per_gpu / std::max(1, sm)));  // at least a tile per CTA and phase: the copies, not the kernels, set the pace (profiles/r02_multi_phases_n8.txt)
    if (getenv("LZG_MULTI_PHASES")) phases = std::max(1, atoi(getenv("LZG_MULTI_PHASES")));
    phases = std::min(phases, blocks2g);
    for (int g = 0; g < G && rc == LZG_OK; g++) {
        lzg_ctx* c = m->ctx[g];
        LZM_CUDA(m, cudaSetDevice(c->device));
        for (int h = 0; h < G; h++)
            if (h != g) LZM_CUDA(m, cudaStreamWaitEvent(c->stream, m->ev_up[h], 0));
    }
    // per-GPU profile handles for the whole call: the tensor-core path keeps its thermometer code in them.  The range
    // kernels of all GPUs are queued first, then each GPU's code is laid out and its expansion queued.
    std::vector<lzg_profiles> profs(G, shape);
    for (int g = 0; g < G && rc == LZG_OK; g++) {
        lzg_ctx* c = m->ctx[g];
        LZM_CUDA(m, cudaSetDevice(c->device));
        profs[g].ctx = c;
        profs[g].d = m->d_prof[g];
        profs[g].d_bytes = (size_t)n_pad * pitch;
        const int tc = lzg_distance_tc_prepare_begin(c, &profs[g]);
        if (tc < 0) rc = -tc, m->err = c->err;
    }
    for (int g = 0; g < G && rc == LZG_OK; g++) {
        lzg_ctx* c = m->ctx[g];
        LZM_CUDA(m, cudaSetDevice(c->device));
        const int tc = lzg_distance_tc_prepare_finish(c, &profs[g]);
        if (tc < 0) rc = -tc, m->err = c->err;
    }
    for (int ph = 0; ph < phases && rc == LZG_OK; ph++) {
        const int ra = std::min(R, (int)((long long)blocks2g * ph / phases) * 2 * G);
        const int rb = ph + 1 == phases ? R : std::min(R, (int)((long long)blocks2g * (ph + 1) / phases) * 2 * G);
        for (int g = 0; g < G && rc == LZG_OK; g++) {
            lzg_ctx* c = m->ctx[g];
            LZM_CUDA(m, cudaSetDevice(c->device));
            int64_t mine = 0;
            rc = lzg_distance_run(c, &profs[g], g, G, ra, rb, nullptr, &mine, m->d_table[g], 1);
            if (rc != LZG_OK) m->err = c->err;
            pairs += mine;
            LZM_CUDA(m, cudaEventRecord(m->ev_done[g], c->stream));
        }
        for (int g = 0; g < G && rc == LZG_OK; g++) {
            lzg_ctx* c = m->ctx[g];
            LZM_CUDA(m, cudaSetDevice(c->device));
            for (int h = 0; h < G; h++) LZM_CUDA(m, cudaStreamWaitEvent(c->copy_stream, m->ev_done[h], 0));
            for (int r = ra; r < rb; r++) {
                if (owner(r) != g) continue;
                const size_t row0 = (size_t)r * bm, nr = std::min<size_t>(n, row0 + bm) - row0;
                LZM_CUDA(m, cudaMemcpyAsync(D_host + row0 * n, table[r], sizeof(int32_t) * nr * n, cudaMemcpyDeviceToHost, c->copy_stream));
            }
        }
    }
    for (int g = 0; g < G && rc == LZG_OK; g++) {
        lzg_ctx* c = m->ctx[g];
        LZM_CUDA(m, cudaSetDevice(c->device));
        LZM_CUDA(m, cudaMemcpyAsync(c->h_pinned, c->d_flag, sizeof(int), cudaMemcpyDeviceToHost, c->copy_stream));
    }
    bool out_of_range = false;
    for (int g = 0; g < G; g++) {
        lzg_ctx* c = m->ctx[g];
        cudaSetDevice(c->device);
        lzg_distance_tc_free(c, &profs[g]);  // stream-ordered frees behind the last kernel
        cudaError_t e1 = cudaStreamSynchronize(c->stream), e2 = cudaStreamSynchronize(c->copy_stream);
        if (rc == LZG_OK && (e1 != cudaSuccess || e2 != cudaSuccess)) rc = multi_fail(m, LZG_ERR_CUDA, "lzg_distance_matrix_multi", e1 != cudaSuccess ? e1 : e2);
        if (rc == LZG_OK && L > 0 && c->h_pinned[0]) out_of_range = true;
    }
    if (rc == LZG_OK && out_of_range) *out_of_range_ = true;
    if (pairs_out) *pairs_out = pairs;
    return rc;
}

static int distance_multi(lzg_multi* m, const void* host, bool is_i8, int32_t n, int32_t L, int32_t* D_host, int64_t* pairs_out) {
    if (!m || n < 0 || L < 0 || (n > 0 && L > 0 && !host) || (n > 0 && !D_host)) return multi_fail(m, LZG_ERR_ARG, "lzg_distance_matrix_multi: bad argument");
    if (pairs_out) *pairs_out = 0;
    if (n == 0) return LZG_OK;
    device_guard restore;
    for (int width = 1; width <= 2; width++) {  // the 8-bit format first; its range flag sends wider data through the 16-bit one
        bool out_of_range = false;
        const int rc = distance_multi_width(m, host, is_i8, n, L, D_host, pairs_out, width, &out_of_range);
        if (rc != LZG_OK) return rc;
        if (!out_of_range) return LZG_OK;
    }
    if (pairs_out) *pairs_out = 0;
    return multi_fail(m, LZG_ERR_RANGE, "breakpoint value outside [-32767, 32767]");
}

int lzg_distance_matrix_multi(lzg_multi* m, const int32_t* bp, int32_t n_cells, int32_t n_pos, int32_t* D_host, int64_t* pairs) {
    return distance_multi(m, bp, false, n_cells, n_pos, D_host, pairs);
}

int lzg_distance_matrix_multi_i8(lzg_multi* m, const int8_t* bp, int32_t n_cells, int32_t n_pos, int32_t* D_host, int64_t* pairs) {
    return distance_multi(m, bp, true, n_cells, n_pos, D_host, pairs);
}

}  // extern "C"
