// gz_rules.cu — Board rules on bitboards (reference game.py:11-172).
//   K3: do_move, has_a_winner / game_end, legal mask, current_state planes, move_target.
// One warp per position/game: lanes share the (tiny) bitboard work and write the per-cell
// outputs coalesced.
#include "gz_engine.h"

__global__ void rules_eval_kernel(const int32_t *__restrict__ moves, const int64_t *__restrict__ offsets,
                                  int64_t n_pos, const int8_t *__restrict__ start, int W, int H, int n,
                                  int32_t *status, int8_t *ended, int8_t *winner, int8_t *haswin,
                                  uint8_t *legal, float *planes, uint8_t *target, int8_t *cur)
{
    const int lane = threadIdx.x & 31;
    const int64_t pos = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (pos >= n_pos) return;
    const int S = W * H;
    const int64_t b0 = offsets[pos], b1 = offsets[pos + 1];
    BB st[2] = {bb_zero(), bb_zero()};
    int sp = start ? start[pos] : 0;
    int player = sp; // index 0 -> player 1
    int last = -1, ns = 0, bad = 0;
    for (int64_t i = b0; i < b1; ++i) {
        int mv = moves[i];
        if (mv < 0 || mv >= S) { bad = 1; break; }
        int bit = bb_bit(mv, W);
        if (bb_test(st[0], bit) || bb_test(st[1], bit)) { bad = 1; break; }
        bb_set(st[player], bit);
        player ^= 1;
        last = mv;
        ++ns;
    }
    int win;
    bool hw;
    bool end = bb_game_end(st[0], st[1], ns, W, H, n, &win, &hw);
    BB occ = bb_or(st[0], st[1]);
    BB leg = bb_andnot(bb_board_mask(W, H), occ);
    if (lane == 0) {
        if (status) status[pos] = bad ? GZ_E_ILLEGAL : 0;
        if (ended) ended[pos] = end ? 1 : 0;
        if (winner) winner[pos] = (int8_t)win;
        if (haswin) haswin[pos] = hw ? 1 : 0;
        if (cur) cur[pos] = (int8_t)(player + 1);
    }
    if (legal)
        for (int m = lane; m < S; m += GZ_WARP) legal[pos * S + m] = bb_test(leg, bb_bit(m, W)) ? 1 : 0;
    if (target) {
        BB tg = bb_move_target(occ, leg, W, H);
        for (int m = lane; m < S; m += GZ_WARP) target[pos * S + m] = bb_test(tg, bb_bit(m, W)) ? 1 : 0;
    }
    if (planes) gz_write_planes(st[player], st[player ^ 1], last, ns, W, H, planes + pos * 4 * S, lane);
}

int gzk_rules_eval(gz_engine *e, const int32_t *d_moves, const int64_t *d_offsets, int64_t n_pos,
                   const int8_t *d_start, int32_t *d_status, int8_t *d_ended, int8_t *d_winner,
                   int8_t *d_haswin, uint8_t *d_legal, float *d_planes, uint8_t *d_target,
                   int8_t *d_cur)
{
    if (n_pos <= 0) return GZ_OK;
    const int wpb = 4;
    int64_t blocks = (n_pos + wpb - 1) / wpb;
    rules_eval_kernel<<<(unsigned)blocks, wpb * 32, 0, e->stream>>>(
        d_moves, d_offsets, n_pos, d_start, e->W, e->H, e->n_in_row, d_status, d_ended, d_winner,
        d_haswin, d_legal, d_planes, d_target, d_cur);
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
