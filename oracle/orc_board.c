/*
 * ORACLE — TEST INFRASTRUCTURE ONLY (see orc_board.h).
 * Board rules restated from reference game.py:11-172.
 */
#include "orc_board.h"
#include <string.h>

int orc_board_sizeof(void) { return (int)sizeof(orc_board); }

/* game.py:60-68 init_board: raises if width or height < n_in_row (-> -1 here). */
int orc_board_init(orc_board *b, int width, int height, int n_in_row, int start_player)
{
    if (width * height > ORC_MAXS || width < 1 || height < 1) return -2;
    if (start_player != 0 && start_player != 1) return -3;
    b->width = width;
    b->height = height;
    b->n_in_row = n_in_row;
    if (width < n_in_row || height < n_in_row) return -1;
    memset(b->states, 0, sizeof(b->states));
    b->n_avail = width * height;
    for (int i = 0; i < b->n_avail; ++i) b->availables[i] = i;
    b->n_moved = 0;
    b->last_move = -1;
    b->current_player = start_player == 0 ? 1 : 2;
    b->n_target = 0;
    return 0;
}

/* game.py:119-127 do_move: states[move] = player; availables.remove(move) (ValueError
 * when absent -> -1 here, board untouched apart from nothing); toggle player; last_move. */
int orc_board_do_move(orc_board *b, int move)
{
    int pos = -1;
    for (int i = 0; i < b->n_avail; ++i)
        if (b->availables[i] == move) { pos = i; break; }
    if (pos < 0) return -1;
    b->states[move] = b->current_player;
    for (int i = pos; i + 1 < b->n_avail; ++i) b->availables[i] = b->availables[i + 1];
    b->n_avail--;
    b->n_moved++;
    b->current_player = (b->current_player == 2) ? 1 : 2;
    b->last_move = move;
    return 0;
}

static int line_same(const orc_board *b, int m, int step, int n)
{
    int p = b->states[m];
    for (int k = 0; k < n; ++k) {
        int idx = m + k * step;
        /* states.get(i, -1): indexes beyond the board cannot occur under the guards */
        if (idx < 0 || idx >= b->width * b->height) return 0;
        if (b->states[idx] != p) return 0;
    }
    return 1;
}

/* game.py:129-160 has_a_winner.  Early-out when fewer than n+2 stones (:136-137);
 * then for every placed stone test E, N, NE, NW runs of length n with the reference's
 * guards (:144-158).  Iteration is ascending move order (the reference iterates a set;
 * the order only matters in positions with two winners, unreachable by legal play). */
int orc_board_has_a_winner(const orc_board *b, int *winner)
{
    int W = b->width, H = b->height, n = b->n_in_row;
    *winner = -1;
    if (b->n_moved < n + 2) return 0;
    for (int m = 0; m < W * H; ++m) {
        if (!b->states[m]) continue;
        int h = m / W, w = m % W;
        int ok_e = (w >= 0 && w < W - n + 1);
        int ok_n = (h >= 0 && h < H - n + 1);
        int ok_w = (w >= n - 1 && w < W);
        if (ok_e && line_same(b, m, 1, n)) { *winner = b->states[m]; return 1; }
        if (ok_n && line_same(b, m, W, n)) { *winner = b->states[m]; return 1; }
        if (ok_e && ok_n && line_same(b, m, W + 1, n)) { *winner = b->states[m]; return 1; }
        if (ok_w && ok_n && line_same(b, m, W - 1, n)) { *winner = b->states[m]; return 1; }
    }
    return 0;
}

/* game.py:162-169 game_end. */
int orc_board_game_end(const orc_board *b, int *winner)
{
    int w;
    if (orc_board_has_a_winner(b, &w)) { *winner = w; return 1; }
    *winner = -1;
    return b->n_avail == 0 ? 1 : 0;
}

/* game.py:92-116 current_state.  Allocated (4, width, height), indexed
 * [move // width, move % height]; returned as the row-reversed view [:, ::-1, :].
 * `out` receives the RETURNED orientation as contiguous double[4][width][height].
 * Square boards only (the reference's indexing is only self-consistent there). */
int orc_board_current_state(const orc_board *b, double *out)
{
    int W = b->width, H = b->height;
    if (W != H) return -1;
    memset(out, 0, sizeof(double) * 4 * W * H);
    if (b->n_moved > 0) {
        for (int m = 0; m < W * H; ++m) {
            if (!b->states[m]) continue;
            int r = m / W, c = m % H;
            int plane = (b->states[m] == b->current_player) ? 0 : 1;
            out[(plane * W + (W - 1 - r)) * H + c] = 1.0;
        }
        int r = b->last_move / W, c = b->last_move % H;
        out[(2 * W + (W - 1 - r)) * H + c] = 1.0;
    }
    if (b->n_moved % 2 == 0)
        for (int i = 0; i < W * H; ++i) out[3 * W * H + i] = 1.0;
    return 0;
}

/* game.py:25-58 get_move_target / move_in_rect: bounding box of all stones grown by 2
 * with the reference's exact clipping (:48-52); empty board -> rows 5..9 x cols 5..9
 * hard-coded (:41); inclusive membership test (:29). */
void orc_board_get_move_target(orc_board *b)
{
    int W = b->width, H = b->height;
    int top, bottom, left, right;
    if (b->n_moved == 0) {
        top = 9; bottom = 5; left = 5; right = 9;
    } else {
        top = -1; bottom = 1 << 30; left = 1 << 30; right = -1;
        for (int m = 0; m < W * H; ++m) {
            if (!b->states[m]) continue;
            int r = m / W, c = m % W;
            if (r > top) top = r;
            if (r < bottom) bottom = r;
            if (c < left) left = c;
            if (c > right) right = c;
        }
        const int step = 2;
        top = (top <= H - step) ? top + step : H;
        bottom = (bottom >= step) ? bottom - step : 0;
        left = (left >= step) ? left - step : 0;
        right = (right <= W - step) ? right + step : W;
    }
    b->n_target = 0;
    for (int i = 0; i < b->n_avail; ++i) {
        int x = b->availables[i];
        int r = x / W, c = x % W;
        if (r >= bottom && r <= top && c >= left && c <= right)
            b->move_target[b->n_target++] = x;
    }
}

/* Philox4x32-10 (Salmon et al. 2011), standard constants. */
void orc_philox4x32(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t c2,
                    uint32_t c3, uint32_t out[4])
{
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    const uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)M0 * c0;
        uint64_t p1 = (uint64_t)M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
