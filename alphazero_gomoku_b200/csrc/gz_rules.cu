// gz_rules.cu — Board rules on bitboards (reference game.py:11-172).
//   K3: do_move, has_a_winner / game_end, legal mask, current_state planes, move_target.
// A warp (or half a warp) per position/game: lanes share the (tiny) bitboard work and write the per-cell
// outputs coalesced.
#include "gz_engine.h"

// A warp per position.  The move list is read 32 moves at a time (one coalesced load): a lane owns one
// move, earlier occurrences of the same cell inside the chunk are found with match.any, the first
// illegal move of the list (out of range or occupied, list.remove's ValueError in game.py:122) ends the
// replay exactly where the reference's sequential loop would, and the stones are OR-reduced into the two
// bitboards with redux.sync.  Per-cell outputs leave as 4-byte / 16-byte stores when S is a multiple of 4
// (the bulk of the traffic: 18 * S bytes per position against 4 bytes per ply read).
// The per-position scalar work is spread over the lanes too (8 lanes test one colour x direction each,
// lane h holds board row h for the candidate box); WT > 0 makes the board side a compile-time constant
// for the usual square sizes (cell <-> bit conversions become shifts / multiplies).
// LP = lanes per position: 16 puts two positions in a warp (every warp-wide primitive runs on the half-warp's
// mask), which halves the issued instructions per position - the kernel is issue-bound, not HBM-bound.
template <int WT, int LP>
__global__ void __launch_bounds__(128)
rules_eval_kernel(const int32_t *__restrict__ moves, const int64_t *__restrict__ offsets,
                  int64_t n_pos, const int8_t *__restrict__ start, int W_, int H_, int n, const BB board_mask,
                  int32_t *status, int8_t *ended, int8_t *winner, int8_t *haswin,
                  uint8_t *legal, float *planes, uint8_t *target, int8_t *cur)
{
    static_assert(LP == 32 || LP == 16, "a position takes a warp or half a warp");
    const int lane = threadIdx.x & (LP - 1);                     // lane inside the position's segment
    const int sh = (threadIdx.x & 31) & ~(LP - 1);               // first warp lane of the segment
    const unsigned seg = LP == 32 ? 0xffffffffu : (0xffffu << sh); // the segment's member mask
    const int64_t pos = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / LP;
    if (pos >= n_pos) return;
    const int W = WT > 0 ? WT : W_, H = WT > 0 ? WT : H_;
    const int S = W * H;
    constexpr int NPIECE = WT > 0 ? ((WT - 1) * 16 + WT + 31) / 32 : 8; // 32-bit pieces of a bitboard that hold cells
    const int64_t b0 = offsets[pos], b1 = offsets[pos + 1];
    const int sp = start ? start[pos] : 0; // index 0 -> player 1
    uint32_t acc[2][8]; // the two 256-bit boards as 32-bit pieces (this lane's contribution, reduced per chunk)
    BB st[2] = {bb_zero(), bb_zero()};
    int last = -1, ns = 0, bad = 0;
    for (int64_t base = b0; base < b1 && !bad; base += LP) {
        const bool have = base + lane < b1;
        const int mv = have ? moves[base + lane] : -1;
        const bool inr = have && mv >= 0 && mv < S;
        const int bit = inr ? bb_bit(mv, W) : 0;
        const bool occupied = inr && (bb_test(st[0], bit) || bb_test(st[1], bit));
        const unsigned peers = __match_any_sync(seg, inr ? mv : -1 - lane) >> sh;
        const bool dup = inr && (peers & ((1u << lane) - 1u)) != 0u; // an earlier move of this chunk took the cell
        const unsigned badmask = (__ballot_sync(seg, have && (!inr || occupied || dup)) & seg) >> sh;
        const unsigned havemask = (__ballot_sync(seg, have) & seg) >> sh;
        const int nvalid = badmask ? __ffs(badmask) - 1 : __popc(havemask);
        const bool apply = lane < nvalid;
        const int pl = (sp + ns + lane) & 1;
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int q = 0; q < 8; ++q) // pieces above the board's last row stay empty
                acc[c][q] = q < NPIECE ? __reduce_or_sync(seg, (apply && pl == c && (bit >> 5) == q) ? (1u << (bit & 31)) : 0u) : 0u;
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int q = 0; q < 4; ++q) st[c].w[q] |= (uint64_t)acc[c][2 * q] | ((uint64_t)acc[c][2 * q + 1] << 32);
        if (nvalid > 0) last = __shfl_sync(seg, mv, nvalid - 1, LP);
        ns += nvalid;
        bad = badmask != 0u;
    }
    const int player = (sp + ns) & 1;
    // game_end / has_a_winner (game.py:129-169) as bb_game_end does it, one colour x direction per lane
    bool w0 = false, w1 = false;
    if (ns >= n + 2) {
        const int d4 = lane & 3;
        const int dir = (d4 == 0) ? 1 : (d4 == 1) ? 16 : (d4 == 2) ? 17 : 15;
        bool line = false;
        if (lane < 8) line = bb_has_line_dir(bb_pick(st[0], st[1], lane < 4), n, dir);
        const unsigned b = (__ballot_sync(seg, line) & seg) >> sh;
        w0 = (b & 0x0Fu) != 0u;
        w1 = (b & 0xF0u) != 0u;
    }
    const bool hw = w0 || w1;
    const int win = w0 ? 1 : (w1 ? 2 : -1);
    const bool end = hw || ns >= S;
    BB occ = bb_or(st[0], st[1]);
    BB leg = bb_andnot(board_mask, occ);
    const BB own = bb_pick(st[0], st[1], player == 0), opp = bb_pick(st[1], st[0], player == 0); // side to move / other
    if (lane == 0) {
        if (status) status[pos] = bad ? GZ_E_ILLEGAL : 0;
        if (ended) ended[pos] = end ? 1 : 0;
        if (winner) winner[pos] = (int8_t)win;
        if (haswin) haswin[pos] = hw ? 1 : 0;
        if (cur) cur[pos] = (int8_t)(player + 1);
    }
    const bool s4 = (S & 3) == 0;
    auto write_mask = [&](uint8_t *out, const BB &m) { // uint8[S] membership of m
        if (s4 && (reinterpret_cast<uintptr_t>(out) & 3) == 0) {
            uint32_t *o = reinterpret_cast<uint32_t *>(out + pos * S);
            for (int q = lane; q < (S >> 2); q += LP) {
                uint32_t v = 0;
                if ((W & 3) == 0) { // the four cells sit in one board row: a nibble of that row
                    const int c4 = 4 * q, row = c4 / W, col = c4 - row * W;
                    const uint32_t nb = (bb_row(m, row) >> col) & 0xFu;
                    v = (nb & 1u) | ((nb & 2u) << 7) | ((nb & 4u) << 14) | ((nb & 8u) << 21);
                } else {
#pragma unroll
                    for (int k = 0; k < 4; ++k) v |= (bb_test(m, bb_bit(4 * q + k, W)) ? 1u : 0u) << (8 * k);
                }
                o[q] = v;
            }
        } else {
            for (int c = lane; c < S; c += LP) out[pos * S + c] = bb_test(m, bb_bit(c, W)) ? 1 : 0;
        }
    };
    if (legal) write_mask(legal, leg);
    if (target) {
        const BB tg = bb_move_target_warp(occ, leg, W, H, lane, seg, sh);
        write_mask(target, tg);
    }
    if (planes) {
        float *out = planes + pos * 4 * S;
        if (s4 && (reinterpret_cast<uintptr_t>(planes) & 15) == 0) {
            // current_state (game.py:92-116), returned orientation, four cells per 16-byte store
            const float colour = (ns % 2 == 0) ? 1.f : 0.f;
            const int lr = last >= 0 ? last / W : -1, lc = last >= 0 ? last % H : -1;
            const int S4 = S >> 2;
            for (int q = lane; q < S; q += LP) { // q = plane * S4 + group of 4 cells
                const int plane = q / S4, idx0 = (q - plane * S4) * 4;
                float v[4];
                if ((H & 3) == 0) { // one board row per group of four: a nibble per plane
                    const int rs = idx0 / H, c = idx0 - rs * H, r = W - 1 - rs;
                    uint32_t nb;
                    if (plane < 2) nb = (bb_row(bb_pick(own, opp, plane == 0), r) >> c) & 0xFu;
                    else if (plane == 2) nb = (ns > 0 && r == lr && lc >= c && lc < c + 4) ? (1u << (lc - c)) : 0u;
                    else nb = (ns % 2 == 0) ? 0xFu : 0u;
#pragma unroll
                    for (int k = 0; k < 4; ++k) v[k] = ((nb >> k) & 1u) ? 1.f : 0.f;
                } else {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int idx = idx0 + k, rs = idx / H, c = idx % H, r = W - 1 - rs, bit = r * 16 + c;
                        v[k] = plane == 0 ? (bb_test(own, bit) ? 1.f : 0.f)
                             : plane == 1 ? (bb_test(opp, bit) ? 1.f : 0.f)
                             : plane == 2 ? ((ns > 0 && r == lr && c == lc) ? 1.f : 0.f) : colour;
                    }
                }
                reinterpret_cast<float4 *>(out)[q] = make_float4(v[0], v[1], v[2], v[3]);
            }
        } else {
            gz_write_planes(own, opp, last, ns, W, H, out, lane, LP);
        }
    }
}

int gzk_rules_eval(gz_engine *e, const int32_t *d_moves, const int64_t *d_offsets, int64_t n_pos,
                   const int8_t *d_start, int32_t *d_status, int8_t *d_ended, int8_t *d_winner,
                   int8_t *d_haswin, uint8_t *d_legal, float *d_planes, uint8_t *d_target,
                   int8_t *d_cur)
{
    if (n_pos <= 0) return GZ_OK;
    // half a warp per position where the per-cell outputs leave as packed stores (S a multiple of 4: measured
    // 2.19 vs 1.79 TB/s at 8x8), a whole warp otherwise (1.92 vs 1.73 TB/s at 15x15)
    const bool half = (e->S & 3) == 0;
    const int wpb = half ? 8 : 4; // positions per 128-thread block
    int64_t blocks = (n_pos + wpb - 1) / wpb;
    ProfScope ps(e, PC_OTHER);
    const BB mask = bb_board_mask(e->W, e->H);
#define GZ_RULES(WT) (half ? rules_eval_kernel<WT, 16> : rules_eval_kernel<WT, 32>)<<<(unsigned)blocks, 128, 0, e->stream>>>( \
        d_moves, d_offsets, n_pos, d_start, e->W, e->H, e->n_in_row, mask, d_status, d_ended, d_winner, \
        d_haswin, d_legal, d_planes, d_target, d_cur)
    const int side = e->W == e->H ? e->W : 0;
    if (side == 6) GZ_RULES(6);
    else if (side == 8) GZ_RULES(8);
    else if (side == 9) GZ_RULES(9);
    else if (side == 11) GZ_RULES(11);
    else if (side == 15) GZ_RULES(15);
    else GZ_RULES(0);
#undef GZ_RULES
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

// ---- engine-resident boards -------------------------------------------------------------
__global__ void boards_reset_kernel(GameBoard *boards, GameMeta *meta, const int8_t *start, int G)
{
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    store_bb(boards[g].stones[0], bb_zero());
    store_bb(boards[g].stones[1], bb_zero());
    GameMeta m;
    m.n_stones = 0;
    m.last_move = -1;
    m.start_player = start ? start[g] : 0;
    m.flags = 0;
    meta[g] = m;
}

int gzk_boards_reset(gz_engine *e, const int8_t *d_start)
{
    boards_reset_kernel<<<(e->G + 127) / 128, 128, 0, e->stream>>>(e->boards, e->meta, d_start, e->G);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

// Board.do_move (game.py:119-127): set the mover's bit, toggle player, remember last move.
__global__ void boards_do_move_kernel(GameBoard *boards, GameMeta *meta, const int32_t *moves,
                                      int32_t *status, int G, int W, int S)
{
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    int mv = moves[g];
    if (mv < 0) {
        if (status) status[g] = 0;
        return;
    }
    GameMeta m = meta[g];
    BB s0 = load_bb(boards[g].stones[0]), s1 = load_bb(boards[g].stones[1]);
    int bit = mv < S ? bb_bit(mv, W) : 0;
    if (mv >= S || bb_test(s0, bit) || bb_test(s1, bit)) {
        if (status) status[g] = GZ_E_ILLEGAL;
        m.flags |= 1;
        meta[g] = m;
        return;
    }
    int player = (m.start_player + m.n_stones) & 1;
    if (player == 0) { bb_set(s0, bit); store_bb(boards[g].stones[0], s0); }
    else { bb_set(s1, bit); store_bb(boards[g].stones[1], s1); }
    m.n_stones += 1;
    m.last_move = mv;
    meta[g] = m;
    if (status) status[g] = 0;
}

int gzk_boards_do_move(gz_engine *e, const int32_t *d_moves, int32_t *d_status)
{
    boards_do_move_kernel<<<(e->G + 127) / 128, 128, 0, e->stream>>>(e->boards, e->meta, d_moves,
                                                                      d_status, e->G, e->W, e->S);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

__global__ void boards_query_kernel(const GameBoard *boards, const GameMeta *meta, int G, int W, int H,
                                    int n, int8_t *ended, int8_t *winner, int8_t *cur,
                                    int32_t *nstones, int32_t *last, uint8_t *legal, float *planes,
                                    int8_t *cells)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (g >= G) return;
    const int S = W * H;
    GameMeta m = meta[g];
    BB s0 = load_bb(boards[g].stones[0]), s1 = load_bb(boards[g].stones[1]);
    int player = (m.start_player + m.n_stones) & 1;
    if (lane == 0) {
        int win;
        bool hw;
        bool end = bb_game_end(s0, s1, m.n_stones, W, H, n, &win, &hw);
        if (ended) ended[g] = end ? 1 : 0;
        if (winner) winner[g] = (int8_t)win;
        if (cur) cur[g] = (int8_t)(player + 1);
        if (nstones) nstones[g] = m.n_stones;
        if (last) last[g] = m.last_move;
    }
    if (legal || cells) {
        for (int mv = lane; mv < S; mv += GZ_WARP) {
            int bit = bb_bit(mv, W);
            bool a = bb_test(s0, bit), b = bb_test(s1, bit);
            if (legal) legal[(size_t)g * S + mv] = (!a && !b) ? 1 : 0;
            if (cells) cells[(size_t)g * S + mv] = a ? 1 : (b ? 2 : 0);
        }
    }
    if (planes)
        gz_write_planes(player == 0 ? s0 : s1, player == 0 ? s1 : s0, m.last_move, m.n_stones, W, H,
                        planes + (size_t)g * 4 * S, lane);
}

int gzk_boards_query(gz_engine *e, int8_t *d_ended, int8_t *d_winner, int8_t *d_cur,
                     int32_t *d_nstones, int32_t *d_last, uint8_t *d_legal, float *d_planes,
                     int8_t *d_cells)
{
    const int wpb = 4;
    boards_query_kernel<<<(e->G + wpb - 1) / wpb, wpb * 32, 0, e->stream>>>(
        e->boards, e->meta, e->G, e->W, e->H, e->n_in_row, d_ended, d_winner, d_cur, d_nstones,
        d_last, d_legal, d_planes, d_cells);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}
