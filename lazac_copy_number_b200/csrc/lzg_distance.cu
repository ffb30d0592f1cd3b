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
#include <algorithm>
#include <cstdlib>
#include <cstring>

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

// int32 [rows][L] -> biased 16-bit words [rows][l_pad]; flags values outside [-32767, 32767].
__global__ void k_pack_i32_w2(const int32_t* __restrict__ src, uint8_t* __restrict__ dst, int rows, int L, int l_pad, int* flag) {
    int r = blockIdx.y;
    const int32_t* s = src + (size_t)r * L;
    uint32_t* d = reinterpret_cast<uint32_t*>(dst + (size_t)r * l_pad * 2);
    int bad = 0;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < l_pad / 2; w += gridDim.x * blockDim.x) {
        uint32_t out = 0;
#pragma unroll
        for (int b = 0; b < 2; b++) {
            int k = w * 2 + b;
            int v = k < L ? s[k] : 0;
            bad |= (v < -32767) | (v > 32767);
            out |= (uint32_t)((v + LZG_BIAS16) & 0xffff) << (16 * b);
        }
        d[w] = out;
    }
    if (bad) atomicOr(flag, 1);
}

// int8 [rows][L] -> biased 16-bit words [rows][l_pad] (an int8 matrix that holds -128: the one value the 8-bit format lacks)
__global__ void k_pack_i8_w2(const int8_t* __restrict__ src, uint8_t* __restrict__ dst, int rows, int L, int l_pad) {
    int r = blockIdx.y;
    const int8_t* s = src + (size_t)r * L;
    uint16_t* d = reinterpret_cast<uint16_t*>(dst + (size_t)r * l_pad * 2);
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < l_pad; k += gridDim.x * blockDim.x) d[k] = (uint16_t)((k < L ? (int)s[k] : 0) + LZG_BIAS16);
}

__global__ void k_fill_bytes(uint8_t* dst, size_t n, uint8_t v) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[i] = v;
}

// In-place finish of an int8 upload: rows were copied raw (two's complement) into the padded
// [n_pad][l_pad] layout; v + 128 == v ^ 0x80 for a byte, padding columns and rows become 128.
// -128 (0x80) is out of range (keeps |a - b| <= 254).
__global__ void k_bias_rows_i8(uint8_t* __restrict__ d, int n, int L, int l_pad, int* flag) {
    const int r = blockIdx.y;
    uint4* row = reinterpret_cast<uint4*>(d + (size_t)r * l_pad);
    int bad = 0;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < l_pad / 16; w += gridDim.x * blockDim.x) {
        uint4 v = make_uint4(0x80808080u, 0x80808080u, 0x80808080u, 0x80808080u);
        const int k0 = w * 16;
        if (r < n && k0 < L) {
            uint4 x = row[w];
            uint32_t in[4] = {x.x, x.y, x.z, x.w}, out[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                uint32_t word = in[q], o = 0x80808080u;
                if (k0 + 4 * q + 4 <= L) {
                    bad |= __vcmpeq4(word, 0x80808080u) != 0;
                    o = word ^ 0x80808080u;
                } else {
#pragma unroll
                    for (int b = 0; b < 4; b++) {
                        if (k0 + 4 * q + b < L) {
                            uint32_t byte = (word >> (8 * b)) & 0xffu;
                            bad |= byte == 0x80u;
                            o = (o & ~(0xffu << (8 * b))) | ((byte ^ 0x80u) << (8 * b));
                        }
                    }
                }
                out[q] = o;
            }
            v = make_uint4(out[0], out[1], out[2], out[3]);
        }
        row[w] = v;
    }
    if (bad) atomicOr(flag, 1);
}

// ---- the pair kernel ---------------------------------------------------------------------------
// Work is described by SEGMENTS built on the host (get_schedule below): a segment is one
// BM x BN tile of cell pairs (rows of tile-row r against rows of tile-row c, c >= r) over the
// 16-byte position chunks [k0, k1).  The grid is persistent -- one CTA per SM -- and CTA b
// executes segments seg_off[b] .. seg_off[b+1].  The host cuts the (tile, chunk) space into
// ranges of equal cost (stream-K at chunk granularity, a fixed cost per segment for its
// pipeline fill and epilogue), so every SM finishes together whatever the number of tiles; a
// tile cut across CTAs is accumulated with integer atomics (exact, order-free) into cells zeroed
// by k_zero_tiles.  For large matrices the plan also decides which tiles run side by side (whole
// tiles in lock step, one super-block of tiles at a time) so that operand strips are shared
// through the L2 -- plan_schedule below.
//
// Diagonal tiles (r == c) skip their lower-left quadrant: thread (ty, tx) owns rows ty + TY*i and
// columns tx + TX*j, so "row >= BM/2 and column < BN/2" is i >= 4 && j < 4 for every thread
// alike -- a quarter of the SADs of those tiles, with no divergence; the quadrant is filled
// from the mirror of the upper-right one.
struct lzg_seg {
    int32_t r, c, k0, k1;  // chunks [k0, k1); k1 < 0: accumulate with atomics into pre-zeroed cells, range [k0, -k1)
};

// One k-iteration (KC positions) of a thread's 8 x 8 pairs.  DIAG skips the lower-left quadrant.
// (Measured on B200: reordering the SADs, streaming the row operands, or 8-byte operands with
// two CTAs per SM are all within +-2 % or slower -- the loop is VABSDIFF4-issue bound.)
// 16-bit positions: one VABSDIFF.U32 per position of a pair (there is no packed 2 x 16 SAD on sm_100a: vabsdiff2 is
// emulated), the halves of a word extracted once per row / column and reused over the 8 x 8 pairs.
__device__ __forceinline__ void sad_w2(uint32_t (&acc)[8][8], const uint32_t (&a)[8], const uint32_t (&b)[8], bool diag) {
    uint32_t alo[8], ahi[8], blo[8], bhi[8];
#pragma unroll
    for (int i = 0; i < 8; i++) alo[i] = a[i] & 0xffffu, ahi[i] = a[i] >> 16, blo[i] = b[i] & 0xffffu, bhi[i] = b[i] >> 16;
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (diag && i >= 4 && j < 4) continue;
            acc[i][j] = __usad(alo[i], blo[j], acc[i][j]);
            acc[i][j] = __usad(ahi[i], bhi[j], acc[i][j]);
        }
}

template <int BM, int BN, bool DIAG, int W>
__device__ __forceinline__ void tile_step(uint32_t (&acc)[8][8], const uint8_t* sA, const uint8_t* sB, int tx, int ty) {
    constexpr int TX = BN / 8, TY = BM / 8;
    const int swa = ty & 7, swb = tx & 7;  // (row & 7) of every row this thread reads
#pragma unroll
    for (int c = 0; c < 8; c++) {
        uint4 a[8], b[8];
#pragma unroll
        for (int i = 0; i < 8; i++)
            a[i] = *reinterpret_cast<const uint4*>(sA + (ty + TY * i) * KC + ((c ^ swa) << 4));
#pragma unroll
        for (int j = 0; j < 8; j++)
            b[j] = *reinterpret_cast<const uint4*>(sB + (tx + TX * j) * KC + ((c ^ swb) << 4));
        if (W == 2) {
            uint32_t aw[8], bw[8];
#pragma unroll
            for (int q = 0; q < 4; q++) {
#pragma unroll
                for (int i = 0; i < 8; i++) aw[i] = q == 0 ? a[i].x : q == 1 ? a[i].y : q == 2 ? a[i].z : a[i].w;
#pragma unroll
                for (int j = 0; j < 8; j++) bw[j] = q == 0 ? b[j].x : q == 1 ? b[j].y : q == 2 ? b[j].z : b[j].w;
                sad_w2(acc, aw, bw, DIAG);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++)
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if (DIAG && i >= 4 && j < 4) continue;  // lower-left quadrant: taken from the mirror
                    acc[i][j] = sad4(a[i].x, b[j].x, acc[i][j]);
                    acc[i][j] = sad4(a[i].y, b[j].y, acc[i][j]);
                    acc[i][j] = sad4(a[i].z, b[j].z, acc[i][j]);
                    acc[i][j] = sad4(a[i].w, b[j].w, acc[i][j]);
                }
        }
    }
}

// The last stage of a segment may hold fewer than eight chunks: same arithmetic, rolled loop.
template <int BM, int BN, bool DIAG, int W>
__device__ __forceinline__ void tile_step_partial(uint32_t (&acc)[8][8], const uint8_t* sA, const uint8_t* sB, int tx, int ty, int nc) {
    constexpr int TX = BN / 8, TY = BM / 8;
    const int swa = ty & 7, swb = tx & 7;
#pragma unroll 1
    for (int c = 0; c < nc; c++) {
        uint4 a[8], b[8];
#pragma unroll
        for (int i = 0; i < 8; i++)
            a[i] = *reinterpret_cast<const uint4*>(sA + (ty + TY * i) * KC + ((c ^ swa) << 4));
#pragma unroll
        for (int j = 0; j < 8; j++)
            b[j] = *reinterpret_cast<const uint4*>(sB + (tx + TX * j) * KC + ((c ^ swb) << 4));
        if (W == 2) {
            uint32_t aw[8], bw[8];
#pragma unroll
            for (int q = 0; q < 4; q++) {
#pragma unroll
                for (int i = 0; i < 8; i++) aw[i] = q == 0 ? a[i].x : q == 1 ? a[i].y : q == 2 ? a[i].z : a[i].w;
#pragma unroll
                for (int j = 0; j < 8; j++) bw[j] = q == 0 ? b[j].x : q == 1 ? b[j].y : q == 2 ? b[j].z : b[j].w;
                sad_w2(acc, aw, bw, DIAG);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++)
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if (DIAG && i >= 4 && j < 4) continue;
                    acc[i][j] = sad4(a[i].x, b[j].x, acc[i][j]);
                    acc[i][j] = sad4(a[i].y, b[j].y, acc[i][j]);
                    acc[i][j] = sad4(a[i].z, b[j].z, acc[i][j]);
                    acc[i][j] = sad4(a[i].w, b[j].w, acc[i][j]);
                }
        }
    }
}

// l_pad = bytes per profile row (positions x bytes per position); W = bytes per position.
template <int BM, int BN, int STAGES, int W>
__global__ void __launch_bounds__((BM / 8) * (BN / 8), 1)
k_zcnt_distance(const uint8_t* __restrict__ prof, int l_pad, int n, const lzg_seg* __restrict__ segs,
                const int32_t* __restrict__ seg_off, int32_t* __restrict__ D, int32_t* const* __restrict__ row_base) {
    constexpr int TX = BN / 8, TY = BM / 8, T = TX * TY;
    constexpr int A_BYTES = BM * KC, B_BYTES = BN * KC, STAGE_BYTES = A_BYTES + B_BYTES;
    extern __shared__ __align__(128) uint8_t smem[];

    const int tid = threadIdx.x;
    const int tx = tid % TX, ty = tid / TX;
    const int s_begin = seg_off[blockIdx.x], s_end = seg_off[blockIdx.x + 1];

    for (int s = s_begin; s < s_end; s++) {
        const lzg_seg sg = segs[s];
        const bool atomic = sg.k1 < 0;
        const int k_begin = sg.k0, n_chunks = (atomic ? -sg.k1 : sg.k1) - sg.k0, n_it = (n_chunks + 7) >> 3;
        const bool diag = sg.r == sg.c;
        const int row0 = sg.r * BM, col0 = sg.c * BN;
        // where the output rows of tile row r live: one N x N matrix, or (several GPUs, lzg_multi.cu) the
        // owner's row block -- the mirror of a tile then goes straight into a peer's memory over NVLink
        int32_t* const out_r = row_base ? row_base[sg.r] : D + (size_t)row0 * n;
        int32_t* const out_c = row_base ? row_base[sg.c] : D + (size_t)col0 * n;
        const uint8_t* gA = prof + (size_t)row0 * l_pad + (size_t)k_begin * 16;
        const uint8_t* gB = prof + (size_t)col0 * l_pad + (size_t)k_begin * 16;

        auto load_stage = [&](int it, int slot) {
            uint8_t* sA = smem + slot * STAGE_BYTES;
            uint8_t* sB = sA + A_BYTES;
            const uint8_t* a = gA + (size_t)it * KC;
            const uint8_t* b = gB + (size_t)it * KC;
            const int c = tid & 7;  // T is a multiple of 8: a thread always copies the same chunk column
            if (it * 8 + c < n_chunks) {  // the last stage of a segment may be partial
#pragma unroll
                for (int q = tid; q < BM * 8; q += T) {
                    const int r = q >> 3;
                    cp_async16(sA + r * KC + ((c ^ (r & 7)) << 4), a + (size_t)r * l_pad + (c << 4));
                }
#pragma unroll
                for (int q = tid; q < BN * 8; q += T) {
                    const int r = q >> 3;
                    cp_async16(sB + r * KC + ((c ^ (r & 7)) << 4), b + (size_t)r * l_pad + (c << 4));
                }
            }
        };

        uint32_t acc[8][8];
#pragma unroll
        for (int i = 0; i < 8; i++)
#pragma unroll
            for (int j = 0; j < 8; j++) acc[i][j] = 0;

#pragma unroll
        for (int st = 0; st < STAGES - 1; st++) {
            if (st < n_it) load_stage(st, st);
            cp_async_commit();
        }

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
            const int nc = n_chunks - it * 8;
            if (nc >= 8) {
                if (!diag) tile_step<BM, BN, false, W>(acc, sA, sB, tx, ty);
                else tile_step<BM, BN, true, W>(acc, sA, sB, tx, ty);
            } else {
                if (!diag) tile_step_partial<BM, BN, false, W>(acc, sA, sB, tx, ty, nc);
                else tile_step_partial<BM, BN, true, W>(acc, sA, sB, tx, ty, nc);
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
        {
            // tile: thread owns column tid % BN, rows (tid / BN) + k * (T / BN); pointers advance by rows
            constexpr int RS = T / BN;
            const int c = tid % BN, rmax = min(BM, n - row0);
            const bool col_ok = col0 + c < n;
            const int32_t* src = tbuf + (tid / BN) * LD + c;
            int32_t* dst = out_r + (size_t)(tid / BN) * n + col0 + c;
            const int r_skip = (diag && c < BN / 2) ? BM / 2 : BM;  // lower-left quadrant of a diagonal tile
            const int r_end = col_ok ? min(rmax, r_skip) : 0;
            if (atomic) {
#pragma unroll 4
                for (int r = tid / BN; r < r_end; r += RS, src += RS * LD, dst += (size_t)RS * n) atomicAdd(dst, *src);
            } else {
#pragma unroll 4
                for (int r = tid / BN; r < r_end; r += RS, src += RS * LD, dst += (size_t)RS * n) *dst = *src;
            }
        }
        {
            // mirror: thread owns tile row tid % BM (a column of D), tile columns (tid / BM) + k * (T / BM)
            constexpr int CS = T / BM;
            const int r = tid % BM, cmax = min(BN, n - col0);
            const bool row_ok = row0 + r < n && !(diag && r >= BM / 2);
            int c = tid / BM;
            if (diag) c += BN / 2;  // only the upper-right quadrant of a diagonal tile is mirrored
            const int32_t* src = tbuf + r * LD + c;
            int32_t* dst = out_c + (size_t)c * n + row0 + r;
            const int c_end = row_ok ? cmax : 0;
            if (atomic) {
#pragma unroll 4
                for (; c < c_end; c += CS, src += CS, dst += (size_t)CS * n) atomicAdd(dst, *src);
            } else {
#pragma unroll 4
                for (; c < c_end; c += CS, src += CS, dst += (size_t)CS * n) *dst = *src;
            }
        }
        __syncthreads();  // tbuf aliases the pipeline buffers of the next segment
    }
}

// Zeroes the cells (tile and mirror) of the tiles that are accumulated with atomics.
template <int BM, int BN>
__global__ void k_zero_tiles(const int2* __restrict__ tiles, int n, int32_t* __restrict__ D, int32_t* const* __restrict__ row_base) {
    const int2 t = tiles[blockIdx.x];
    const int row0 = t.x * BM, col0 = t.y * BN;
    int32_t* const out_r = row_base ? row_base[t.x] : D + (size_t)row0 * n;
    int32_t* const out_c = row_base ? row_base[t.y] : D + (size_t)col0 * n;
    for (int q = blockIdx.y * blockDim.x + threadIdx.x; q < BM * BN; q += gridDim.y * blockDim.x) {
        int r = q / BN, c = q % BN;
        int gi = row0 + r, gj = col0 + c;
        if (gi < n && gj < n) {
            out_r[(size_t)r * n + gj] = 0;
            out_c[(size_t)c * n + gi] = 0;
        }
    }
}

template <int BM, int BN, int STAGES>
constexpr int dist_smem_bytes() {
    constexpr int pipe = STAGES * (BM + BN) * KC;
    constexpr int epi = BM * (BN + 1) * 4;
    return pipe > epi ? pipe : epi;
}

// Cuts the shard's (tile, chunk) space into at most n_ctas contiguous ranges of equal cost.
// Tile rows are dealt to shards in pairs (r, R-1-r): every pair has R+1 tiles, so every shard
// gets the same triangle area.  Cost of one 16-byte chunk of a tile: 4 units, 3 for a diagonal
// tile; every segment adds LZG_SEG_COST units (pipeline fill + epilogue, about 0.8 of a 128-byte
// stage measured on cfg2).  The smallest feasible per-CTA budget is found by bisection.
struct lzg_plan {
    std::vector<lzg_seg> segs;
    std::vector<int32_t> seg_off;  // [n_ctas + 1]
    std::vector<int2> split;       // tiles accumulated with atomics
    int64_t pairs = 0;
    int n_ctas = 0;
};

// Host only (no CUDA call): lzg_distance_plan exports it so the CPU tests can check coverage and balance.
// own_mode 0: tile rows dealt in pairs (r, R-1-r) -- equal area over the WHOLE triangle; own_mode 1: in snake order
// (0 .. G-1, G-1 .. 0, ...) -- equal area over every block of 2 G consecutive tile rows, which is what lets several
// GPUs work through the triangle in row-group phases (lzg_multi.cu).
static inline int tile_row_owner(int r, int R, int n_shards, int own_mode) {
    if (own_mode == 0) return std::min(r, R - 1 - r) % n_shards;
    const int k = r % n_shards;
    return (r / n_shards) % 2 == 0 ? k : n_shards - 1 - k;
}

void plan_schedule(int n, int l_pad, int bm, int sm_count, int slots_per_sm, int shard, int n_shards, int row_lo, int row_hi,
                   lzg_plan& out, int own_mode = 0) {
    const int R = (n + bm - 1) / bm;
    const int K = l_pad / 16;  // chunks per row
    std::vector<int2> tiles;
    int64_t pairs = 0, units = 0;
    auto rows_in = [&](int r) { return (int64_t)(std::min(n, (r + 1) * bm) - r * bm); };
    for (int r = 0; r < R; r++) {
        if (tile_row_owner(r, R, n_shards, own_mode) != shard || r < row_lo || r >= row_hi) continue;
        for (int c = r; c < R; c++) {
            tiles.push_back(make_int2(r, c));
            pairs += (c == r) ? rows_in(r) * (rows_in(r) + 1) / 2 : rows_in(r) * rows_in(c);
            units += (int64_t)K * (c == r ? 3 : 4);
        }
    }
    const int64_t seg_cost = getenv("LZG_SEG_COST") ? atoi(getenv("LZG_SEG_COST")) : 26;
    const int min_take = 8;  // never cut a segment shorter than one stage (unless the tile ends)
    const int G = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)sm_count * slots_per_sm, (int64_t)tiles.size() * ((K + min_take - 1) / min_take)));
    // Large matrices: the CTAs run side by side at the same speed, so what CTA b does as its s-th
    // tile decides what is in L2 together.  In row-major order the 148 CTAs sit in 148 different
    // tile rows and every operand strip comes from DRAM (cfg3: 54.8 GB read for 0.4 GB of
    // profiles, profiles/r01_ncu_dist_cfg3.txt).  Here the off-diagonal tiles are listed by
    // SB x SB super-blocks (rows by their rank among the shard's rows) and, once the cuts are
    // known, dealt so that step s of all CTAs is one run of that list: one super-block's 2 SB
    // strips at a time.  Only off-diagonal tiles change places, and they all cost the same, so
    // the cuts stay where the row-major list put them; the diagonal tiles keep their positions
    // (spread over the CTAs: their cost is the least certain, and a CTA holding only diagonal
    // tiles finished 1.6 ms late at cfg3).
    const int SB = getenv("LZG_DIST_SB") ? atoi(getenv("LZG_DIST_SB")) : 12;
    size_t n_off = 0;
    for (const int2& t : tiles) n_off += t.x != t.y;
    const bool reorder = SB > 0 && n_off >= (size_t)2 * G;
    std::vector<int2> ord;
    if (reorder) {
        std::vector<int> rank_of(R, 0);
        for (int r = 0, k = 0; r < R; r++) {
            if (tile_row_owner(r, R, n_shards, own_mode) != shard || r < row_lo || r >= row_hi) continue;
            rank_of[r] = k++;
        }
        for (const int2& t : tiles)
            if (t.x != t.y) ord.push_back(t);
        std::stable_sort(ord.begin(), ord.end(), [&](const int2& a, const int2& b) {
            const int ba = rank_of[a.x] / SB, bb = rank_of[b.x] / SB;
            if (ba != bb) return ba < bb;
            return a.y / SB < b.y / SB;  // inside a super-block: row-major, as listed
        });
    }
    // Phase-aligned form (default; LZG_DIST_ALIGN=0 keeps the contiguous cuts; cfg3: 13.0 -> 7.7 GB read, same time): every CTA first runs m = n_off / G WHOLE
    // off-diagonal tiles -- CTA b's s-th is ord[s * G + b], so all CTAs start every tile together
    // and the CTAs sharing a strip want the same chunk at the same time -- and only the rest
    // (fewer than G off-diagonal tiles plus the diagonal ones) is cut into G equal-cost tails.
    const bool aligned = reorder && !(getenv("LZG_DIST_ALIGN") && atoi(getenv("LZG_DIST_ALIGN")) == 0);
    const size_t m_whole = aligned ? n_off / (size_t)G : 0;
    if (aligned) {
        std::vector<int2> rest(ord.begin() + m_whole * G, ord.end());
        for (const int2& t : tiles)
            if (t.x == t.y) rest.push_back(t);
        tiles.swap(rest);
        units = 0;
        for (const int2& t : tiles) units += (int64_t)K * (t.x == t.y ? 3 : 4);
    }
    std::vector<lzg_seg> segs;
    std::vector<int32_t> seg_off, seg_tile;
    // greedy fill with a per-CTA budget; returns the number of CTAs used (segments written when emit)
    auto fill = [&](int64_t budget, bool emit) {
        int ctas = 1;
        int64_t load = 0;
        if (emit) {
            segs.clear();
            seg_tile.clear();
            seg_off.assign(1, 0);
        }
        for (size_t ti = 0; ti < tiles.size(); ti++) {
            const int2 t = tiles[ti];
            const int u = t.x == t.y ? 3 : 4;
            int k = 0;
            while (k < K) {
                const int want = std::min(K - k, min_take);
                if (load > 0 && load + seg_cost + (int64_t)want * u > budget) {  // this CTA is full
                    ctas++;
                    load = 0;
                    if (emit) seg_off.push_back((int32_t)segs.size());
                }
                int take = (int)std::min<int64_t>(K - k, std::max<int64_t>(want, (budget - load - seg_cost) / u));
                if (K - k - take < min_take) take = K - k;  // no sliver at the end of a tile
                if (emit) {
                    segs.push_back(lzg_seg{t.x, t.y, k, k + take});
                    seg_tile.push_back((int32_t)ti);
                }
                load += seg_cost + (int64_t)take * u;
                k += take;
            }
        }
        return ctas;
    };
    int64_t lo = (units + (int64_t)tiles.size() * seg_cost) / G, hi = units + (int64_t)tiles.size() * seg_cost + seg_cost;
    while (lo < hi) {
        const int64_t mid = (lo + hi) / 2;
        if (fill(mid, false) <= G) hi = mid;
        else lo = mid + 1;
    }
    int used = tiles.empty() ? 0 : fill(lo, true);
    if (reorder && !aligned) {
        // x0[b] = list position of the tile CTA b starts in; its s-th off-diagonal tile becomes ord[s * used + b]
        std::vector<int64_t> x0;
        for (int b = 0; b < used; b++) x0.push_back(seg_tile[seg_off[b]]);
        x0.push_back((int64_t)tiles.size());
        std::vector<char> taken(n_off, 0), placed(tiles.size(), 0);
        for (int b = 0; b < used; b++) {
            int64_t step = 0;
            for (int64_t pos = x0[b]; pos < x0[b + 1]; pos++) {
                if (tiles[pos].x == tiles[pos].y) {
                    placed[pos] = 1;
                    continue;
                }
                const int64_t idx = step++ * used + b;
                if (idx < (int64_t)n_off && !taken[idx]) {
                    tiles[pos] = ord[idx];
                    taken[idx] = placed[pos] = 1;
                }
            }
        }
        size_t next = 0;
        for (size_t pos = 0; pos < tiles.size(); pos++) {
            if (placed[pos]) continue;
            while (taken[next]) next++;
            tiles[pos] = ord[next];
            taken[next++] = 1;
        }
        used = fill(lo, true);
    }
    if (!tiles.empty()) seg_off.push_back((int32_t)segs.size());
    // tiles cut into several segments are accumulated with atomics
    std::vector<int2> split;
    for (size_t q = 0; q < segs.size();) {
        size_t e = q + 1;
        while (e < segs.size() && segs[e].r == segs[q].r && segs[e].c == segs[q].c) e++;
        if (e - q > 1) {
            for (size_t i = q; i < e; i++) segs[i].k1 = -segs[i].k1;
            split.push_back(make_int2(segs[q].r, segs[q].c));
        }
        q = e;
    }
    if (aligned) {  // CTA b: its whole tiles, then its tail
        std::vector<lzg_seg> all;
        std::vector<int32_t> all_off;
        for (int b = 0; b < G; b++) {
            all_off.push_back((int32_t)all.size());
            for (size_t st = 0; st < m_whole; st++) {
                const int2 t = ord[st * G + b];
                all.push_back(lzg_seg{t.x, t.y, 0, K});
            }
            if (b < used) all.insert(all.end(), segs.begin() + seg_off[b], segs.begin() + seg_off[b + 1]);
        }
        all_off.push_back((int32_t)all.size());
        segs.swap(all);
        seg_off.swap(all_off);
        used = G;
    }
    if (seg_off.empty()) seg_off.assign(2, 0);
    out.segs.swap(segs);
    out.seg_off.swap(seg_off);
    out.split.swap(split);
    out.pairs = pairs;
    out.n_ctas = std::max(used, 1);
}

const lzg_schedule* get_schedule(lzg_ctx* ctx, lzg_profiles* p, int bm, int slots_per_sm, int shard, int n_shards,
                                 int row_lo, int row_hi, int* status, int own_mode) {
    for (auto& sc : ctx->schedules)
        if (sc.n == p->n && sc.l_pad == p->pitch && sc.bm == bm && sc.shard == shard && sc.n_shards == n_shards && sc.row_lo == row_lo && sc.row_hi == row_hi && sc.own_mode == own_mode) return &sc;
    lzg_plan plan;
    plan_schedule(p->n, p->pitch, bm, ctx->sm_count, slots_per_sm, shard, n_shards, row_lo, row_hi, plan, own_mode);
    std::vector<lzg_seg>& segs = plan.segs;
    std::vector<int32_t>& seg_off = plan.seg_off;
    std::vector<int2>& split = plan.split;
    lzg_schedule sc;
    sc.n = p->n;
    sc.l_pad = p->pitch;  // the plan is cut in 16-byte chunks of a row
    sc.bm = bm;
    sc.shard = shard;
    sc.own_mode = own_mode;
    sc.n_shards = n_shards;
    sc.row_lo = row_lo;
    sc.row_hi = row_hi;
    sc.pairs = plan.pairs;
    sc.n_ctas = plan.n_ctas;
    sc.n_segs = (int)segs.size();
    sc.n_split = (int)split.size();
    if (sc.n_segs > 0) {
        // one pooled device block: [segs | seg_off | split tiles], one copy
        const size_t b_segs = sizeof(lzg_seg) * segs.size(), b_off = sizeof(int32_t) * seg_off.size();
        const size_t o_off = b_segs, o_split = (o_off + b_off + 15) / 16 * 16;
        sc.bytes = o_split + sizeof(int2) * split.size();
        std::vector<uint8_t> blob(sc.bytes);
        memcpy(blob.data(), segs.data(), b_segs);
        memcpy(blob.data() + o_off, seg_off.data(), b_off);
        if (!split.empty()) memcpy(blob.data() + o_split, split.data(), sizeof(int2) * split.size());
        cudaError_t e = lzg_dev_alloc(ctx, &sc.d_block, sc.bytes);
        if (e == cudaSuccess) e = cudaMemcpyAsync(sc.d_block, blob.data(), sc.bytes, cudaMemcpyHostToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);  // the source is a pageable host vector
        if (e != cudaSuccess) {
            lzg_dev_free(ctx, sc.d_block);
            *status = lzg_fail(ctx, LZG_ERR_CUDA, "distance schedule upload", e);
            return nullptr;
        }
        sc.d_segs = sc.d_block;
        sc.d_seg_off = reinterpret_cast<int32_t*>((uint8_t*)sc.d_block + o_off);
        sc.d_split = reinterpret_cast<int2*>((uint8_t*)sc.d_block + o_split);
    }
    if (ctx->schedules.size() >= 64) {  // a small cache: same-shape matrices reuse their schedule
        lzg_dev_free(ctx, ctx->schedules.front().d_block);
        ctx->schedules.erase(ctx->schedules.begin());
    }
    ctx->schedules.push_back(sc);
    return &ctx->schedules.back();
}

template <int BM, int BN, int STAGES, int W>
int launch_distance(lzg_ctx* ctx, const lzg_profiles* p, const lzg_schedule* sc, int32_t* d_D, int32_t* const* d_row_base) {
    constexpr int smem = dist_smem_bytes<BM, BN, STAGES>();
    constexpr int T = (BM / 8) * (BN / 8);
    if (sc->n_split > 0) {
        k_zero_tiles<BM, BN><<<dim3(sc->n_split, 16), 256, 0, ctx->stream>>>(sc->d_split, p->n, d_D, d_row_base);
        LZG_LAUNCH_CHECK(ctx);
    }
    auto kern = k_zcnt_distance<BM, BN, STAGES, W>;
    LZG_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    LZG_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    kern<<<sc->n_ctas, T, smem, ctx->stream>>>(p->d, p->pitch, p->n, (const lzg_seg*)sc->d_segs, sc->d_seg_off, d_D, d_row_base);
    LZG_LAUNCH_CHECK(ctx);
    return LZG_OK;
}

}  // namespace

int lzg_pack_rows_i32(lzg_ctx* ctx, const int32_t* d_src, uint8_t* d_dst, int rows, int L, int l_pad, int width) {
    dim3 grid(std::max(1, std::min(64, (l_pad / 4 + 255) / 256)), rows);
    if (width == 2) k_pack_i32_w2<<<grid, 256, 0, ctx->stream>>>(d_src, d_dst, rows, L, l_pad, ctx->d_flag);
    else k_pack_i32<<<grid, 256, 0, ctx->stream>>>(d_src, d_dst, rows, L, l_pad, ctx->d_flag);
    LZG_LAUNCH_CHECK(ctx);
    return LZG_OK;
}

int lzg_pack_rows_i8_w2(lzg_ctx* ctx, const int8_t* d_src, uint8_t* d_dst, int rows, int L, int l_pad) {
    dim3 grid(std::max(1, std::min(64, (l_pad + 255) / 256)), rows);
    k_pack_i8_w2<<<grid, 256, 0, ctx->stream>>>(d_src, d_dst, rows, L, l_pad);
    LZG_LAUNCH_CHECK(ctx);
    return LZG_OK;
}

int lzg_bias_rows_i8(lzg_ctx* ctx, uint8_t* d, int n, int n_pad, int L, int l_pad) {
    dim3 grid(std::max(1, std::min(16, (l_pad / 16 + 255) / 256)), n_pad);
    k_bias_rows_i8<<<grid, 256, 0, ctx->stream>>>(d, n, L, l_pad, ctx->d_flag);
    LZG_LAUNCH_CHECK(ctx);
    return LZG_OK;
}

// whole padding rows of a format: bytes 0x80 (width 1), little-endian words 0x8000 (width 2: bytes 0x00, 0x80 alternating)
__global__ void k_fill_pad16(uint32_t* dst, size_t n_words) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_words; i += (size_t)gridDim.x * blockDim.x) dst[i] = 0x80008000u;
}

int lzg_fill_pad(lzg_ctx* ctx, uint8_t* d, size_t n_bytes, int width) {
    if (width == 1) return lzg_fill_bytes(ctx, d, n_bytes, LZG_BIAS);
    if (n_bytes == 0) return LZG_OK;
    const size_t words = n_bytes / 4;
    k_fill_pad16<<<(int)std::min<size_t>(4096, (words + 255) / 256), 256, 0, ctx->stream>>>(reinterpret_cast<uint32_t*>(d), words);
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

// Computes the shard's tiles of tile rows [row_lo, row_hi) into the full N x N device matrix d_D
// (asynchronous on ctx->stream).
int lzg_distance_run(lzg_ctx* ctx, lzg_profiles* p, int shard, int n_shards, int row_lo, int row_hi, int32_t* d_D,
                     int64_t* pairs_done, int32_t* const* d_row_base, int own_mode) {
    if (pairs_done) *pairs_done = 0;
    if (p->n == 0) return LZG_OK;
    // narrow value ranges: the exact tensor-core product (lzg_distance_tc.cu)
    {
        const int tc = lzg_distance_tc_prepare(ctx, p);
        if (tc < 0) return -tc;
        if (tc == 1) return lzg_distance_tc_run(ctx, p, shard, n_shards, own_mode, row_lo, row_hi, d_D, d_row_base, pairs_done);
    }
    int status = LZG_OK;
    const int bm = lzg_distance_tile(p);
    // one persistent CTA per SM: 8 warps x 254 registers fill the register file (3 CTAs of 2 warps for
    // the single-tile 64 x 64 case)
    const lzg_schedule* sc = get_schedule(ctx, p, bm, bm == 64 ? 3 : 1, shard, n_shards, row_lo, row_hi, &status, own_mode);
    if (!sc) return status;
    if (pairs_done) *pairs_done = sc->pairs;
    if (sc->n_segs == 0) return LZG_OK;
    if (p->width == 2) return bm == 64 ? launch_distance<64, 64, 4, 2>(ctx, p, sc, d_D, d_row_base) : launch_distance<128, 128, 3, 2>(ctx, p, sc, d_D, d_row_base);
    return bm == 64 ? launch_distance<64, 64, 4, 1>(ctx, p, sc, d_D, d_row_base) : launch_distance<128, 128, 3, 1>(ctx, p, sc, d_D, d_row_base);
}

// Host-only view of the work plan (no GPU, no context): see include/lazac_b200.h.
int lzg_distance_plan(int32_t n, int32_t L, int32_t sm_count, int32_t shard, int32_t n_shards, int32_t row_lo, int32_t row_hi,
                      int32_t* segs, int64_t cap_segs, int32_t* seg_off, int32_t cap_ctas, int32_t* info) {
    if (n < 1 || L < 1 || sm_count < 1 || n_shards < 1 || shard < 0 || shard >= n_shards || !info) return LZG_ERR_ARG;
    const int l_pad = std::max(128, (L + 127) / 128 * 128);  // as lzg_profiles_upload pads (the 8-bit format: one byte per position)
    const int bm = n <= 64 ? 64 : 128;                       // lzg_distance_tile
    lzg_plan plan;
    plan_schedule(n, l_pad, bm, sm_count, bm == 64 ? 3 : 1, shard, n_shards, row_lo, row_hi, plan);
    info[0] = plan.n_ctas;
    info[1] = (int32_t)plan.segs.size();
    info[2] = bm;
    info[3] = l_pad / 16;
    if ((int64_t)plan.segs.size() > cap_segs || plan.n_ctas > cap_ctas || (plan.segs.size() && (!segs || !seg_off))) return LZG_ERR_ARG;
    for (size_t i = 0; i < plan.segs.size(); i++) {
        segs[4 * i] = plan.segs[i].r;
        segs[4 * i + 1] = plan.segs[i].c;
        segs[4 * i + 2] = plan.segs[i].k0;
        segs[4 * i + 3] = plan.segs[i].k1;
    }
    for (int b = 0; b <= plan.n_ctas; b++) seg_off[b] = plan.seg_off[b];
    return LZG_OK;
}

// ---- lzg_distance_matrix_noalloc: the distance kernel with its work plan in a file-scope device array -------
// For callers that cannot call cudaMalloc (CUDA graph capture around them, tight serving loops, harnesses that
// forbid it).  The kernel's scratch -- the plan [segments | per-CTA offsets | split tiles] that lzg_distance_run
// keeps in a pooled device block per shape -- lives here in LZG_NOALLOC_SCRATCH_BYTES of static device memory
// (held by every GPU that loads the library until the process exits).  Every call rewrites the part of the array
// it uses and shares it with every other call: calls that overlap on the GPU (two streams, two contexts on one
// device) race.  Nothing in it has to start at zero; the split tiles of D are zeroed by k_zero_tiles as usual.
#define LZG_NOALLOC_SCRATCH_BYTES (1 << 20)
static __device__ __align__(256) unsigned char g_noalloc_scratch[LZG_NOALLOC_SCRATCH_BYTES];

namespace {
// In host code g_noalloc_scratch is a host-side stand-in, not a device address: the address the kernels get comes
// from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kMaxDevices = 64;
unsigned char* g_noalloc_ptr[kMaxDevices] = {};

int noalloc_scratch(lzg_ctx* ctx, unsigned char** out) {
    if (ctx->device < 0 || ctx->device >= kMaxDevices) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_distance_matrix_noalloc: device ordinal beyond 63");
    if (!g_noalloc_ptr[ctx->device]) {
        unsigned char* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr) return lzg_fail(ctx, LZG_ERR_CUDA, "lzg_distance_matrix_noalloc: cudaGetSymbolAddress", e);
        g_noalloc_ptr[ctx->device] = ptr;
    }
    *out = g_noalloc_ptr[ctx->device];
    return LZG_OK;
}
}  // namespace

extern "C" int lzg_distance_matrix_noalloc_max_bytes(void) { return LZG_NOALLOC_SCRATCH_BYTES; }

extern "C" int lzg_distance_matrix_noalloc(lzg_ctx* ctx, const lzg_profiles* cp, int32_t* d_D) {
    if (!ctx || !cp || cp->ctx != ctx || (cp->n > 0 && !d_D)) return lzg_fail(ctx, LZG_ERR_ARG, "lzg_distance_matrix_noalloc: bad argument");
    lzg_profiles* p = const_cast<lzg_profiles*>(cp);
    if (p->n == 0) return LZG_OK;
    LZG_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned char* scratch = nullptr;
    int rc = noalloc_scratch(ctx, &scratch);
    if (rc != LZG_OK) return rc;
    const int bm = lzg_distance_tile(p);
    lzg_plan plan;  // host vectors only
    plan_schedule(p->n, p->pitch, bm, ctx->sm_count, bm == 64 ? 3 : 1, 0, 1, 0, INT32_MAX, plan, 0);
    if (plan.segs.empty()) return LZG_OK;
    const size_t b_segs = sizeof(lzg_seg) * plan.segs.size(), b_off = sizeof(int32_t) * plan.seg_off.size();
    const size_t o_off = b_segs, o_split = (o_off + b_off + 15) / 16 * 16;
    const size_t bytes = o_split + sizeof(int2) * plan.split.size();
    if (bytes > LZG_NOALLOC_SCRATCH_BYTES) {
        char msg[224];
        snprintf(msg, sizeof(msg), "lzg_distance_matrix_noalloc: the work plan of %d x %d profiles needs %zu bytes of scratch, the static array holds %d -- use lzg_distance_matrix_device_async",
                 p->n, p->L, bytes, LZG_NOALLOC_SCRATCH_BYTES);
        return lzg_fail(ctx, LZG_ERR_ARG, msg);
    }
    // three copies from the plan's own (pageable) vectors: the runtime stages them itself, nothing is allocated here
    LZG_CUDA(ctx, cudaMemcpyAsync(scratch, plan.segs.data(), b_segs, cudaMemcpyHostToDevice, ctx->stream));
    LZG_CUDA(ctx, cudaMemcpyAsync(scratch + o_off, plan.seg_off.data(), b_off, cudaMemcpyHostToDevice, ctx->stream));
    if (!plan.split.empty()) LZG_CUDA(ctx, cudaMemcpyAsync(scratch + o_split, plan.split.data(), sizeof(int2) * plan.split.size(), cudaMemcpyHostToDevice, ctx->stream));
    LZG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the sources are pageable and go out of scope
    lzg_schedule sc;
    sc.n = p->n, sc.l_pad = p->pitch, sc.bm = bm, sc.n_shards = 1, sc.row_hi = INT32_MAX;
    sc.pairs = plan.pairs, sc.n_ctas = plan.n_ctas, sc.n_segs = (int)plan.segs.size(), sc.n_split = (int)plan.split.size();
    sc.bytes = bytes;
    sc.d_segs = scratch;
    sc.d_seg_off = reinterpret_cast<int32_t*>(scratch + o_off);
    sc.d_split = reinterpret_cast<int2*>(scratch + o_split);
    if (p->width == 2) return bm == 64 ? launch_distance<64, 64, 4, 2>(ctx, p, &sc, d_D, nullptr) : launch_distance<128, 128, 3, 2>(ctx, p, &sc, d_D, nullptr);
    return bm == 64 ? launch_distance<64, 64, 4, 1>(ctx, p, &sc, d_D, nullptr) : launch_distance<128, 128, 3, 1>(ctx, p, &sc, d_D, nullptr);
}
