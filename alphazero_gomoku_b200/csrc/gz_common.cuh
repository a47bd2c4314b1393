// gz_common.cuh — shared device helpers: 256-bit bitboards, Philox, warp reductions.
//
// Bitboard layout (all board sizes up to 15x15): cell (h, w) -> bit h*16 + w of a 256-bit
// word stored as 4 x u64 (row stride 16, so column 15 is an always-empty guard column and
// rows >= H are empty).  With the guard column, line scans by shift 1 (E), 16 (N),
// 17 (NE) and 15 (NW) never wrap between rows, so the reference's explicit range guards
// (game.py:144-158) hold by construction.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GZ_WARP 32

struct BB {
    uint64_t w[4];
};

__host__ __device__ __forceinline__ BB bb_zero()
{
    BB b;
    b.w[0] = b.w[1] = b.w[2] = b.w[3] = 0ull;
    return b;
}
__host__ __device__ __forceinline__ BB bb_or(const BB &a, const BB &b)
{
    BB r;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.w[i] = a.w[i] | b.w[i];
    return r;
}
__host__ __device__ __forceinline__ BB bb_and(const BB &a, const BB &b)
{
    BB r;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.w[i] = a.w[i] & b.w[i];
    return r;
}
__host__ __device__ __forceinline__ BB bb_andnot(const BB &a, const BB &b) // a & ~b
{
    BB r;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.w[i] = a.w[i] & ~b.w[i];
    return r;
}
__host__ __device__ __forceinline__ bool bb_any(const BB &a)
{
    return (a.w[0] | a.w[1] | a.w[2] | a.w[3]) != 0ull;
}
__host__ __device__ __forceinline__ int bb_bit(int move, int width)
{
    return (move / width) * 16 + (move % width);
}
// word select / update written as select chains so a BB always stays in registers (a dynamic
// index into w[] would push it to local memory)
__host__ __device__ __forceinline__ bool bb_test(const BB &a, int bit)
{
    const int i = bit >> 6;
    const uint64_t w = (i == 0) ? a.w[0] : (i == 1) ? a.w[1] : (i == 2) ? a.w[2] : a.w[3];
    return (w >> (bit & 63)) & 1ull;
}
__host__ __device__ __forceinline__ void bb_set_if(BB &a, int bit, bool cond)
{
    const int i = bit >> 6;
    const uint64_t m = cond ? (1ull << (bit & 63)) : 0ull;
    a.w[0] |= (i == 0) ? m : 0ull;
    a.w[1] |= (i == 1) ? m : 0ull;
    a.w[2] |= (i == 2) ? m : 0ull;
    a.w[3] |= (i == 3) ? m : 0ull;
}
__host__ __device__ __forceinline__ void bb_set(BB &a, int bit) { bb_set_if(a, bit, true); }
__host__ __device__ __forceinline__ BB bb_pick(const BB &a, const BB &b, bool first)
{
    BB r;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.w[i] = first ? a.w[i] : b.w[i];
    return r;
}
__device__ __forceinline__ int bb_popc(const BB &a)
{
    return __popcll(a.w[0]) + __popcll(a.w[1]) + __popcll(a.w[2]) + __popcll(a.w[3]);
}
// logical right shift of the 256-bit word by s, 0 < s < 64 (branch-free funnel shifts)
__host__ __device__ __forceinline__ BB bb_shr(const BB &a, int s)
{
    BB r;
    r.w[0] = (a.w[0] >> s) | (a.w[1] << (64 - s));
    r.w[1] = (a.w[1] >> s) | (a.w[2] << (64 - s));
    r.w[2] = (a.w[2] >> s) | (a.w[3] << (64 - s));
    r.w[3] = (a.w[3] >> s);
    return r;
}
// n-in-a-row detector for one colour: t_k = b & (t_{k-1} >> s); non-zero after n-1 steps
// iff some line of n stones exists in direction s (has_a_winner, game.py:129-160;
// overlines count, as in the reference).
__host__ __device__ __forceinline__ bool bb_has_line(const BB &b, int n)
{
    const int dirs[4] = {1, 16, 17, 15};
    bool win = false;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        BB t = b;
        for (int k = 1; k < n; ++k) t = bb_and(b, bb_shr(t, dirs[d]));
        win = win || bb_any(t);
    }
    return win;
}
// one direction of the detector above (warp-parallel callers give each of 8 lanes a colour x direction)
__host__ __device__ __forceinline__ bool bb_has_line_dir(const BB &b, int n, int s)
{
    BB t = b;
    for (int k = 1; k < n; ++k) t = bb_and(b, bb_shr(t, s));
    return bb_any(t);
}
// The same test for boards of at most 8 x 8 cells, on ONE 64-bit word: w0 / w1 are words 0 and 1 of the colour's
// 16-stride bitboard (rows 0-3 / 4-7, 16 bits apart); the rows are squeezed to one byte each (pitch 8), a direction is
// a shift by 1 / 8 / 9 / 7, and a run may only START where it stays on the board (columns <= W - n for E and NE,
// >= n - 1 for NW: the range guards of game.py:144-158) - with pitch 8 there is no guard column when W = 8.
// d4 = 0 E, 1 N, 2 NE, 3 NW, as the 16-stride shifts 1 / 16 / 17 / 15 of the callers above.
__host__ __device__ __forceinline__ uint64_t bb8_squeeze(uint64_t x) // bytes 0, 2, 4, 6 -> 0, 1, 2, 3
{
    x &= 0x00FF00FF00FF00FFull;
    x = (x | (x >> 8)) & 0x0000FFFF0000FFFFull;
    return (x | (x >> 16)) & 0xFFFFFFFFull;
}
__host__ __device__ __forceinline__ bool bb8_has_line_dir(uint64_t w0, uint64_t w1, int W, int n, int d4)
{
    const uint64_t b = bb8_squeeze(w0) | (bb8_squeeze(w1) << 32);
    const int s = d4 == 0 ? 1 : (d4 == 1 ? 8 : (d4 == 2 ? 9 : 7));
    uint64_t t = b;
    for (int k = 1; k < n; ++k) t = b & (t >> s);
    const uint32_t ce = (W - n >= 0) ? ((2u << (W - n)) - 1u) : 0u;                       // columns 0 .. W - n
    const uint32_t cw = (n - 1 < W) ? (((1u << W) - 1u) & ~((1u << (n - 1)) - 1u)) : 0u;  // columns n - 1 .. W - 1
    const uint64_t col = d4 == 1 ? 0xFFull : (d4 == 3 ? (uint64_t)cw : (uint64_t)ce);
    return (t & (col * 0x0101010101010101ull)) != 0ull;
}
// full-board mask for a W x H board
__host__ __device__ __forceinline__ BB bb_board_mask(int W, int H)
{
    BB m = bb_zero();
    uint64_t row = (1ull << W) - 1ull;
    for (int h = 0; h < H; ++h) m.w[h >> 2] |= row << ((h & 3) * 16);
    return m;
}
__host__ __device__ __forceinline__ uint32_t bb_row(const BB &a, int h)
{
    const int i = h >> 2; // select chain: a dynamic index into w[] would push the board to local memory
    const uint64_t w = (i == 0) ? a.w[0] : (i == 1) ? a.w[1] : (i == 2) ? a.w[2] : a.w[3];
    return (uint32_t)(w >> ((h & 3) * 16)) & 0xFFFFu;
}

// Bounding box of the stones of a 256-bit occupancy word, branch-free: rows from the lowest / highest
// set bit (bit = 16 * row + col), columns from the OR of all rows folded onto 16 bits.
struct BBox {
    int rmin, rmax, cmin, cmax; // rmax < 0: no stone
};
__device__ __forceinline__ BBox bb_bbox(const BB &occ)
{
    BBox b;
    const uint64_t any = occ.w[0] | occ.w[1] | occ.w[2] | occ.w[3];
    const int lo = occ.w[0] ? __ffsll((long long)occ.w[0]) - 1
                 : occ.w[1] ? 63 + __ffsll((long long)occ.w[1])
                 : occ.w[2] ? 127 + __ffsll((long long)occ.w[2])
                            : 191 + __ffsll((long long)occ.w[3]);
    const int hi = occ.w[3] ? 255 - __clzll((long long)occ.w[3])
                 : occ.w[2] ? 191 - __clzll((long long)occ.w[2])
                 : occ.w[1] ? 127 - __clzll((long long)occ.w[1])
                            : 63 - __clzll((long long)occ.w[0]);
    uint32_t c = (uint32_t)any | (uint32_t)(any >> 32);
    c = (c | (c >> 16)) & 0xFFFFu;
    b.rmin = lo >> 4;
    b.rmax = any ? (hi >> 4) : -1;
    b.cmin = __ffs((int)c) - 1;
    b.cmax = 31 - __clz((int)c);
    return b;
}
__device__ __forceinline__ void bbox_add(BBox &b, int bit)
{
    const int r = bit >> 4, c = bit & 15;
    if (b.rmax < 0) { b.rmin = b.rmax = r; b.cmin = b.cmax = c; return; }
    b.rmin = min(b.rmin, r); b.rmax = max(b.rmax, r);
    b.cmin = min(b.cmin, c); b.cmax = max(b.cmax, c);
}
// Board.get_move_target (game.py:35-58) from a bounding box: the box grown by 2 with the reference's
// clipping (:48-52; `top` / `right` may land one past the board, where `legal` is empty), no stone ->
// rows 5..9 x cols 5..9 (:41).  Returns box & legal.  Straight-line code: every row of the board is
// 16 bits, so the box is one 16-bit column pattern replicated over the rows bottom..top.
__device__ __forceinline__ BB bb_target_from_bbox(const BBox &bx, const BB &legal, int W, int H)
{
    int top = 9, bottom = 5, left = 5, right = 9;
    if (bx.rmax >= 0) {
        top = (bx.rmax <= H - 2) ? bx.rmax + 2 : H;
        bottom = (bx.rmin >= 2) ? bx.rmin - 2 : 0;
        left = (bx.cmin >= 2) ? bx.cmin - 2 : 0;
        right = (bx.cmax <= W - 2) ? bx.cmax + 2 : W;
    }
    right = min(right, 15);
    top = min(top, 15);
    const uint32_t rowmask = ((2u << right) - 1u) & ~((1u << left) - 1u) & 0xFFFFu;
    const uint64_t rep = (uint64_t)rowmask * 0x0001000100010001ull;
    BB t;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int lo = max(bottom - 4 * i, 0), hi = min(top - 4 * i, 3);
        const uint64_t m = (hi >= lo) ? ((~0ull >> (48 - 16 * max(hi, 0))) & (~0ull << (16 * min(lo, 3)))) : 0ull;
        t.w[i] = rep & m & legal.w[i];
    }
    return t;
}
__device__ __forceinline__ BB bb_move_target(const BB &occ, const BB &legal, int W, int H)
{
    return bb_target_from_bbox(bb_bbox(occ), legal, W, H);
}

// the same box computed by a warp: lane h holds board row h (every lane returns the full result)
// (seg = member mask of the 32- or 16-lane segment working on this board, sh = its first lane)
__device__ __forceinline__ BB bb_move_target_warp(const BB &occ, const BB &legal, int W, int H, int lane,
                                                  unsigned seg = 0xffffffffu, int sh = 0)
{
    const uint32_t r = lane < H ? bb_row(occ, lane & 15) : 0u;
    const unsigned rows = (__ballot_sync(seg, r != 0u) & seg) >> sh;
    const uint32_t cols = __reduce_or_sync(seg, r);
    int top = 9, bottom = 5, left = 5, right = 9; // empty board: the reference's hard-coded box (game.py:41)
    if (rows) {
        const int rmin = __ffs(rows) - 1, rmax = 31 - __clz(rows), cmin = __ffs(cols) - 1, cmax = 31 - __clz(cols);
        top = (rmax <= H - 2) ? rmax + 2 : H;
        bottom = (rmin >= 2) ? rmin - 2 : 0;
        left = (cmin >= 2) ? cmin - 2 : 0;
        right = (cmax <= W - 2) ? cmax + 2 : W;
    }
    BB box = bb_zero();
    if (right > 15) right = 15;
    if (left > 15) return box;
    const uint64_t rowmask = (((1ull << (right + 1)) - 1ull) & ~((1ull << left) - 1ull)) & 0xFFFFull;
#pragma unroll
    for (int h = 0; h < 16; ++h)
        if (h >= bottom && h <= top) box.w[h >> 2] |= rowmask << ((h & 3) * 16);
    return bb_and(box, legal);
}

// index (bit position) of the k-th set bit (k from 0) of a 256-bit word, -1 when there are not that
// many: word by prefix popcounts, then a popcount binary search inside the word (32 / 16 / 8 / 4 / 2 / 1),
// straight-line selects instead of a clear-lowest-bit loop (up to 63 dependent, divergent iterations).
__device__ __forceinline__ int bb_select(const BB &a, int k)
{
    const int c0 = __popcll(a.w[0]), c1 = __popcll(a.w[1]), c2 = __popcll(a.w[2]), c3 = __popcll(a.w[3]);
    const int p1 = c0, p2 = c0 + c1, p3 = p2 + c2;
    if (k < 0 || k >= p3 + c3) return -1;
    const int i = (k >= p1) + (k >= p2) + (k >= p3);
    k -= (i == 0) ? 0 : (i == 1) ? p1 : (i == 2) ? p2 : p3;
    const uint64_t x = (i == 0) ? a.w[0] : (i == 1) ? a.w[1] : (i == 2) ? a.w[2] : a.w[3];
    uint32_t y = (uint32_t)x;
    int base = 64 * i;
    int c = __popc(y);
    if (k >= c) { k -= c; y = (uint32_t)(x >> 32); base += 32; }
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        const uint32_t lowm = (1u << s) - 1u;
        c = __popc(y & lowm);
        const bool up = k >= c;
        k -= up ? c : 0;
        y = up ? (y >> s) : y;
        y &= lowm;
        base += up ? s : 0;
    }
    return base;
}
// logical right shift by a compile-time amount 0 < S < 256
template <int S> __device__ __forceinline__ BB bb_shr_c(const BB &a)
{
    constexpr int q = S >> 6, r = S & 63;
    BB o;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint64_t lo = (i + q < 4) ? a.w[(i + q) & 3] : 0ull;
        const uint64_t hi = (i + q + 1 < 4) ? a.w[(i + q + 1) & 3] : 0ull;
        o.w[i] = r ? ((lo >> r) | (hi << ((64 - r) & 63))) : lo;
    }
    return o;
}
// n-in-a-row of one colour in direction S with n a compile-time constant: run lengths double
// (pairs, quads, ...) instead of growing by one per step: 3 shift/AND steps for five in a row.
template <int N, int S> __device__ __forceinline__ BB bb_runs_c(const BB &b)
{
    static_assert(N >= 2 && N <= 8, "n_in_row");
    BB t = bb_and(b, bb_shr_c<S>(b)); // runs of 2
    if (N == 2) return t;
    if (N == 3) return bb_and(t, bb_shr_c<2 * S>(b));
    BB u = bb_and(t, bb_shr_c<2 * S>(t)); // runs of 4
    if (N == 4) return u;
    if (N == 5) return bb_and(u, bb_shr_c<4 * S>(b));
    if (N == 6) return bb_and(u, bb_shr_c<4 * S>(t));
    if (N == 7) return bb_and(bb_and(u, bb_shr_c<4 * S>(t)), bb_shr_c<6 * S>(b));
    return bb_and(u, bb_shr_c<4 * S>(u));
}
template <int N> __device__ __forceinline__ bool bb_has_line_c(const BB &b)
{
    const BB e = bb_runs_c<N, 1>(b), n = bb_runs_c<N, 16>(b), ne = bb_runs_c<N, 17>(b), nw = bb_runs_c<N, 15>(b);
    return bb_any(bb_or(bb_or(e, n), bb_or(ne, nw)));
}
// n-in-a-row for the side that just moved: compile-time n for the usual 3..6, else the generic detector
__device__ __forceinline__ bool bb_has_line_n(const BB &b, int n)
{
    switch (n) {
    case 3: return bb_has_line_c<3>(b);
    case 4: return bb_has_line_c<4>(b);
    case 5: return bb_has_line_c<5>(b);
    case 6: return bb_has_line_c<6>(b);
    default: return bb_has_line(b, n);
    }
}
__host__ __device__ __forceinline__ int bb_bit_to_move(int bit, int W) { return (bit >> 4) * W + (bit & 15); }

// Philox4x32-10 (identical to oracle/orc_board.c:orc_philox4x32)
struct Philox4 {
    uint32_t x, y, z, w;
};
__host__ __device__ __forceinline__ Philox4 philox4x32(uint32_t k0, uint32_t k1, uint32_t c0,
                                                       uint32_t c1, uint32_t c2, uint32_t c3)
{
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)M0 * c0;
        uint64_t p1 = (uint64_t)M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    Philox4 o;
    o.x = c0; o.y = c1; o.z = c2; o.w = c3;
    return o;
}

// per-game board record in HBM: 2 colours x 256 bit = 64 B (4 x 128-bit loads) + 16 B meta
struct __align__(16) GameBoard {
    uint64_t stones[2][4]; // [player-1][word]
};
struct __align__(16) GameMeta {
    int32_t n_stones;
    int32_t last_move;   // reference move id, -1 when empty
    int32_t start_player; // 0/1: players[start_player] moved first
    int32_t flags;        // bit0: illegal move was attempted
};

__device__ __forceinline__ BB load_bb(const uint64_t *p)
{
    // two 128-bit loads
    const ulonglong2 *q = reinterpret_cast<const ulonglong2 *>(p);
    ulonglong2 a = q[0], b = q[1];
    BB r;
    r.w[0] = a.x; r.w[1] = a.y; r.w[2] = b.x; r.w[3] = b.y;
    return r;
}
__device__ __forceinline__ void store_bb(uint64_t *p, const BB &v)
{
    ulonglong2 *q = reinterpret_cast<ulonglong2 *>(p);
    q[0] = make_ulonglong2(v.w[0], v.w[1]);
    q[1] = make_ulonglong2(v.w[2], v.w[3]);
}

// game_end from two colour boards (game.py:162-169): returns ended, sets winner (1/2/-1).
// has_a_winner's early-out (< n+2 stones -> no winner, game.py:136-137) is kept.
__host__ __device__ __forceinline__ bool bb_game_end(const BB &s0, const BB &s1, int n_stones, int W,
                                                     int H, int n, int *winner, bool *has_win)
{
    bool w0 = false, w1 = false;
    if (n_stones >= n + 2) {
        w0 = bb_has_line(s0, n);
        w1 = bb_has_line(s1, n);
    }
    *has_win = w0 || w1;
    if (w0 || w1) {
        // both colours winning is unreachable by legal play; the reference would report
        // the owner of the lowest-index winning stone it meets first — pick player 1.
        *winner = w0 ? 1 : 2;
        return true;
    }
    *winner = -1;
    return n_stones >= W * H;
}

// current_state planes (game.py:92-116) in the RETURNED orientation [4][W][H] (row reversed).
__device__ __forceinline__ void gz_write_planes(const BB &own, const BB &opp, int last_move,
                                                int n_stones, int W, int H, float *out, int lane,
                                                int stride = GZ_WARP)
{
    const int S = W * H;
    const float colour = (n_stones % 2 == 0) ? 1.f : 0.f;
    const int lr = last_move >= 0 ? last_move / W : -1;
    const int lc = last_move >= 0 ? last_move % H : -1;
    for (int idx = lane; idx < S; idx += stride) {
        int rs = idx / H, c = idx % H;
        int r = W - 1 - rs;
        int bit = r * 16 + c;
        out[idx] = bb_test(own, bit) ? 1.f : 0.f;
        out[S + idx] = bb_test(opp, bit) ? 1.f : 0.f;
        out[2 * S + idx] = (n_stones > 0 && r == lr && c == lc) ? 1.f : 0.f;
        out[3 * S + idx] = colour;
    }
}

// ---- programmatic dependent launch (the per-wave kernel chain trunk -> heads -> tree -> trunk) ----------
// A kernel of the chain calls gz_pdl_wait() before it touches anything its predecessor produces or still reads
// (everything before the wait is private set-up, or reads of state only kernels TWO steps back write) and
// gz_pdl_trigger() right AFTER its own wait: its successor may then be scheduled as SMs free up, so launch
// latency, block scheduling and the successor's set-up overlap this kernel's body.  Triggering only after the
// wait keeps the chain transitive (when my successor starts, my predecessor is complete); triggering first
// thing let a small tree kernel start while the tree kernel two steps back was still running.
// Without the launch attribute both are no-ops.
__device__ __forceinline__ void gz_pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void gz_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
static cudaError_t gz_launch_chain(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                   Args &&...args)
{
    static const bool no_pdl = getenv("GZ_NO_PDL") != nullptr; // A/B switch
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = no_pdl ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// the same, as thread-block clusters of `cluster` CTAs along x (grid.x must be a multiple of it)
template <typename... KArgs, typename... Args>
static cudaError_t gz_launch_chain_cluster(void (*kern)(KArgs...), unsigned cluster, dim3 grid, dim3 block, size_t smem,
                                           cudaStream_t stream, Args &&...args)
{
    static const bool no_pdl = getenv("GZ_NO_PDL") != nullptr;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = no_pdl ? 1 : 2;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

