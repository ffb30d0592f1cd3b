// lzg_distance.cu -- all-pairs ZCNT distance (L1 between breakpoint vectors) for sm_100a.
//
// Replaces the pair loop of build_distance_matrix, /root/reference/src/lazac.cxx:81-91
// (breakpoint_profile::operator-, include/copy_number.hpp:74-83, then breakpoint_magnitude,
// src/copy_number.cxx:100-113): D[i][j] = D[j][i] = sum_k |bp_i[k] - bp_j[k]|.
//
// Data: profiles are bytes (value + 128) in HBM, row-major [n_pad][l_pad]; padding bytes are 128
// on both sides so they add |128 - 128| = 0.  The work is integer ALU, not a contraction: one
// VABSDIFF4.U8.ACC does four positions of one cell pair, so the kernel is tiled like a GEMM to
// feed that instruction from registers: a BM x BN tile of cell pairs per CTA, an 8 x 8 pair tile
// per thread (64 int32 accumulators), the position axis streamed through a multi-stage
// cp.async (LDGSTS) ring of XOR-swizzled shared-memory tiles read with conflict-free LDS.128.
// Only tiles on or above the diagonal are computed; the epilogue transposes through shared memory
// so that both D[i][j] and D[j][i] are written with coalesced stores.
#include "lzg_common.cuh"

namespace {

constexpr int KC = 128;  // bytes of the position axis per pipeline stage (8 chunks of 16 B)

__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N));
}

// int32 [rows][L] -> biased bytes [rows][l_pad]; flags values outside [-127, 127].
__global__ void k_pack_i32(const int32_t* __restrict__ src, uint8_t* __restrict__ dst, int rows,
                           int L, int l_pad, int* flag) {
    int r = blockIdx.y;
    const int32_t* s = src + (size_t)r * L;
    uint32_t* d = reinterpret_cast<uint32_t*>(dst + (size_t)r * l_pad);
    int bad = 0;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < l_pad / 4; w += gridDim.x * blockDim.x) {
        uint32_t out = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            int k = w * 4 + b;
            int v = k < L ? s[k] : 0;
            bad |= (v < -127) | (v > 127);
            out |= (uint32_t)((v + LZG_BIAS) & 0xff) << (8 * b);
        }
        d[w] = out;
    }
    if (bad) atomicOr(flag, 1);
}

// int8 [rows][L] -> biased bytes [rows][l_pad]; -128 is out of range (keeps |a-b| <= 254).
__global__ void k_pack_i8(const int8_t* __restrict__ src, uint8_t* __restrict__ dst, int rows, int L,
                          int l_pad, int* flag) {
    int r = blockIdx.y;
    const int8_t* s = src + (size_t)r * L;
    uint8_t* d = dst + (size_t)r * l_pad;
    int bad = 0;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < l_pad; k += gridDim.x * blockDim.x) {
        int v = k < L ? s[k] : 0;
        bad |= (v == -128);
        d[k] = (uint8_t)(v + LZG_BIAS);
    }
    if (bad) atomicOr(flag, 1);
}

__global__ void k_fill_bytes(uint8_t* dst, size_t n, uint8_t v) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[i] = v;
}

// One CTA = one BM x BN tile of cell pairs (rows of tile-row bi against rows of tile-row bj),
// over k-iterations [k_begin, k_end) of KC positions each (split-K when the tile count is small).
template <int BM, int BN, int STAGES, bool ATOMIC>
__global__ void __launch_bounds__((BM / 8) * (BN / 8))
k_zcnt_distance(const uint8_t* __restrict__ prof, int l_pad, int n, const int2* __restrict__ tiles,
                int k_iters_total, int k_splits, int32_t* __restrict__ D) {
    constexpr int TX = BN / 8, TY = BM / 8, T = TX * TY;
    constexpr int A_BYTES = BM * KC, B_BYTES = BN * KC, STAGE_BYTES = A_BYTES + B_BYTES;
    extern __shared__ __align__(128) uint8_t smem[];

    const int tid = threadIdx.x;
    const int tx = tid % TX, ty = tid / TX;
    const int2 tile = tiles[blockIdx.x];
    const int row0 = tile.x * BM, col0 = tile.y * BN;
    const int per = (k_iters_total + k_splits - 1) / k_splits;
    const int k_begin = blockIdx.y * per;
    const int k_end = min(k_iters_total, k_begin + per);
    const int n_it = k_end - k_begin;
    if (n_it <= 0) return;

    const uint8_t* gA = prof + (size_t)row0 * l_pad + (size_t)k_begin * KC;
    const uint8_t* gB = prof + (size_t)col0 * l_pad + (size_t)k_begin * KC;

    auto load_stage = [&](int it, int slot) {
        uint8_t* sA = smem + slot * STAGE_BYTES;
        uint8_t* sB = sA + A_BYTES;
        const uint8_t* a = gA + (size_t)it * KC;
        const uint8_t* b = gB + (size_t)it * KC;
#pragma unroll
        for (int q = tid; q < BM * 8; q += T) {
            int r = q >> 3, c = q & 7;
            cp_async16(sA + r * KC + ((c ^ (r & 7)) << 4), a + (size_t)r * l_pad + (c << 4));
        }
#pragma unroll
        for (int q = tid; q < BN * 8; q += T) {
            int r = q >> 3, c = q & 7;
            cp_async16(sB + r * KC + ((c ^ (r & 7)) << 4), b + (size_t)r * l_pad + (c << 4));
        }
    };

    uint32_t acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[i][j] = 0;

#pragma unroll
    for (int s = 0; s < STAGES - 1; s++) {
        if (s < n_it) load_stage(s, s);
        cp_async_commit();
    }

    const int swa = ty & 7, swb = tx & 7;  // (row & 7) of every row this thread reads
    for (int it = 0; it < n_it; it++) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        {
            int nx = it + STAGES - 1;
            if (nx < n_it) load_stage(nx, nx % STAGES);
            cp_async_commit();
        }
        const uint8_t* sA = smem + (it % STAGES) * STAGE_BYTES;
        const uint8_t* sB = sA + A_BYTES;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            uint4 a[8], b[8];
#pragma unroll
            for (int i = 0; i < 8; i++)
                a[i] = *reinterpret_cast<const uint4*>(sA + (ty + TY * i) * KC + ((c ^ swa) << 4));
#pragma unroll
            for (int j = 0; j < 8; j++)
                b[j] = *reinterpret_cast<const uint4*>(sB + (tx + TX * j) * KC + ((c ^ swb) << 4));
#pragma unroll
            for (int i = 0; i < 8; i++)
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    acc[i][j] = sad4(a[i].x, b[j].x, acc[i][j]);
                    acc[i][j] = sad4(a[i].y, b[j].y, acc[i][j]);
                    acc[i][j] = sad4(a[i].z, b[j].z, acc[i][j]);
                    acc[i][j] = sad4(a[i].w, b[j].w, acc[i][j]);
                }
        }
    }
    cp_async_wait<0>();
    __syncthreads();

    // Epilogue: stage the tile in shared memory, then coalesced stores of the tile and its mirror.
    int32_t* tbuf = reinterpret_cast<int32_t*>(smem);
    constexpr int LD = BN + 1;
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) tbuf[(ty + TY * i) * LD + tx + TX * j] = (int32_t)acc[i][j];
    __syncthreads();
    for (int q = tid; q < BM * BN; q += T) {
        int r = q / BN, c = q % BN;
        int gi = row0 + r, gj = col0 + c;
        if (gi < n && gj < n) {
            if (ATOMIC) atomicAdd(D + (size_t)gi * n + gj, tbuf[r * LD + c]);
            else D[(size_t)gi * n + gj] = tbuf[r * LD + c];
        }
    }
    if (tile.x != tile.y) {
        for (int q = tid; q < BM * BN; q += T) {
            int c = q / BM, r = q % BM;
            int gi = row0 + r, gj = col0 + c;
            if (gi < n && gj < n) {
                if (ATOMIC) atomicAdd(D + (size_t)gj * n + gi, tbuf[r * LD + c]);
                else D[(size_t)gj * n + gi] = tbuf[r * LD + c];
            }
        }
    }
}

template <int BM, int BN, int STAGES>
constexpr int dist_smem_bytes() {
    constexpr int pipe = STAGES * (BM + BN) * KC;
    constexpr int epi = BM * (BN + 1) * 4;
    return pipe > epi ? pipe : epi;
}

// Tile rows are dealt to shards in pairs (r, R-1-r): every pair has R+1 tiles.
const lzg_tile_list* get_tiles(lzg_ctx* ctx, lzg_profiles* p, int bm, int shard, int n_shards, int* status) {
    for (auto& tl : p->tile_lists)
        if (tl.bm == bm && tl.shard == shard && tl.n_shards == n_shards) return &tl;
    const int R = (p->n + bm - 1) / bm;
    std::vector<int2> tiles;
    int64_t pairs = 0;
    auto rows_in = [&](int r) { return (int64_t)(std::min(p->n, (r + 1) * bm) - r * bm); };
    for (int r = 0; r < R; r++) {
        int pr = std::min(r, R - 1 - r);
        if (pr % n_shards != shard) continue;
        for (int c = r; c < R; c++) {
            tiles.push_back(make_int2(r, c));
            pairs += (c == r) ? rows_in(r) * (rows_in(r) + 1) / 2 : rows_in(r) * rows_in(c);
        }
    }
    lzg_tile_list tl;
    tl.bm = bm;
    tl.shard = shard;
    tl.n_shards = n_shards;
    tl.n_tiles = (int)tiles.size();
    tl.pairs = pairs;
    if (tl.n_tiles > 0) {
        cudaError_t e = cudaMalloc(&tl.d_tiles, sizeof(int2) * tiles.size());
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(tl.d_tiles, tiles.data(), sizeof(int2) * tiles.size(), cudaMemcpyHostToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) {
            *status = lzg_fail(ctx, LZG_ERR_CUDA, "tile list upload", e);
            return nullptr;
        }
    }
    p->tile_lists.push_back(tl);
    return &p->tile_lists.back();
}

template <int BM, int BN, int STAGES>
int launch_distance(lzg_ctx* ctx, const lzg_profiles* p, const lzg_tile_list* tl, int k_splits, int32_t* d_D) {
    constexpr int smem = dist_smem_bytes<BM, BN, STAGES>();
    constexpr int T = (BM / 8) * (BN / 8);
    const int k_iters = p->l_pad / KC;
    dim3 grid(tl->n_tiles, k_splits);
    if (k_splits > 1) {
        auto kern = k_zcnt_distance<BM, BN, STAGES, true>;
        LZG_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        kern<<<grid, T, smem, ctx->stream>>>(p->d, p->l_pad, p->n, tl->d_tiles, k_iters, k_splits, d_D);
    } else {
        auto kern = k_zcnt_distance<BM, BN, STAGES, false>;
        LZG_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        kern<<<grid, T, smem, ctx->stream>>>(p->d, p->l_pad, p->n, tl->d_tiles, k_iters, k_splits, d_D);
    }
    LZG_LAUNCH_CHECK(ctx);
    return LZG_OK;
}

}  // namespace

int lzg_pack_rows_i32(lzg_ctx* ctx, const int32_t* d_src, uint8_t* d_dst, int rows, int L, int l_pad) {
    dim3 grid(std::max(1, std::min(64, (l_pad / 4 + 255) / 256)), rows);
    k_pack_i32<<<grid, 256, 0, ctx->stream>>>(d_src, d_dst, rows, L, l_pad, ctx->d_flag);
    LZG_LAUNCH_CHECK(ctx);
    return LZG_OK;
}

int lzg_pack_rows_i8(lzg_ctx* ctx, const int8_t* d_src, uint8_t* d_dst, int rows, int L, int l_pad) {
    dim3 grid(std::max(1, std::min(64, (l_pad + 255) / 256)), rows);
    k_pack_i8<<<grid, 256, 0, ctx->stream>>>(d_src, d_dst, rows, L, l_pad, ctx->d_flag);
    LZG_LAUNCH_CHECK(ctx);
    return LZG_OK;
}

int lzg_fill_bytes(lzg_ctx* ctx, uint8_t* d, size_t n, uint8_t v) {
    if (n == 0) return LZG_OK;
    int blocks = (int)std::min<size_t>(4096, (n + 255) / 256);
    k_fill_bytes<<<blocks, 256, 0, ctx->stream>>>(d, n, v);
    LZG_LAUNCH_CHECK(ctx);
    return LZG_OK;
}

// Computes the shard's tiles into the full N x N device matrix d_D (asynchronous).
int lzg_distance_run(lzg_ctx* ctx, lzg_profiles* p, int shard, int n_shards, int32_t* d_D, int64_t* pairs_done) {
    if (p->n == 0) {
        if (pairs_done) *pairs_done = 0;
        return LZG_OK;
    }
    const int R128 = (p->n + 127) / 128;
    const int64_t tiles128 = (int64_t)R128 * (R128 + 1) / 2 / n_shards;
    int status = LZG_OK;
    if (tiles128 >= 2 * (int64_t)ctx->sm_count) {
        const lzg_tile_list* tl = get_tiles(ctx, p, 128, shard, n_shards, &status);
        if (!tl) return status;
        if (pairs_done) *pairs_done = tl->pairs;
        if (tl->n_tiles == 0) return LZG_OK;
        return launch_distance<128, 128, 3>(ctx, p, tl, 1, d_D);
    }
    // Few cells: 64 x 64 tiles and split the position axis so that every SM gets several CTAs.
    const lzg_tile_list* tl = get_tiles(ctx, p, 64, shard, n_shards, &status);
    if (!tl) return status;
    if (pairs_done) *pairs_done = tl->pairs;
    if (tl->n_tiles == 0) return LZG_OK;
    const int k_iters = p->l_pad / KC;
    int want = (ctx->sm_count * 4 + tl->n_tiles - 1) / tl->n_tiles;
    int k_splits = std::max(1, std::min(want, k_iters / 4 > 0 ? k_iters / 4 : 1));
    if (k_splits > 1) {
        // the split-K partial sums are accumulated with integer atomics: clear the shard's cells first
        // (whole matrix for the single-shard case; per-tile clearing is not worth a kernel here)
        LZG_CUDA(ctx, cudaMemsetAsync(d_D, 0, sizeof(int32_t) * (size_t)p->n * p->n, ctx->stream));
    }
    return launch_distance<64, 64, 4>(ctx, p, tl, k_splits, d_D);
}
