// Host-side check (no GPU needed) of the one-word win test used by the fused tree kernels on boards up to 8 x 8
// (gz_common.cuh: bb8_has_line_dir) against the 256-bit detector it replaces (bb_has_line_dir), direction by
// direction, and against a plain scan of the board written from game.py:131-160 - random positions of every density,
// all board sides 3..8 and run lengths 3..6.  Built and run by tests/test_cpu_host.py.
#include <cstdio>
#include <cstdlib>
#include "../../alphazero_gomoku_b200/csrc/gz_common.cuh"

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint32_t rnd()
{
    rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17;
    return (uint32_t)(rng_state >> 32);
}

// has_a_winner's four scans for one colour (game.py:144-158), one direction: 0 E, 1 N, 2 NE, 3 NW
static bool plain_dir(const bool *cell, int W, int H, int n, int d4)
{
    const int dh = d4 == 0 ? 0 : 1, dw = d4 == 0 ? 1 : (d4 == 1 ? 0 : (d4 == 2 ? 1 : -1));
    for (int h = 0; h < H; ++h)
        for (int w = 0; w < W; ++w) {
            bool all = true;
            for (int k = 0; k < n && all; ++k) {
                const int hh = h + k * dh, ww = w + k * dw;
                all = hh >= 0 && hh < H && ww >= 0 && ww < W && cell[hh * W + ww];
            }
            if (all) return true;
        }
    return false;
}

int main()
{
    long checked = 0;
    for (int W = 3; W <= 8; ++W)
        for (int n = 3; n <= 6 && n <= W; ++n)
            for (int rep = 0; rep < 4000; ++rep) {
                const int H = W, dens = 1 + (int)(rnd() % 9); // stones with probability dens / 10
                bool cell[64];
                BB b = bb_zero();
                for (int i = 0; i < W * H; ++i) {
                    cell[i] = (int)(rnd() % 10) < dens;
                    if (cell[i]) {
                        const int bit = bb_bit(i, W);
                        b.w[bit >> 6] |= 1ull << (bit & 63);
                    }
                }
                if (b.w[2] | b.w[3]) { printf("rows 0-7 must live in words 0 and 1\n"); return 1; }
                for (int d4 = 0; d4 < 4; ++d4) {
                    const int s16 = d4 == 0 ? 1 : (d4 == 1 ? 16 : (d4 == 2 ? 17 : 15));
                    const bool want = plain_dir(cell, W, H, n, d4);
                    const bool wide = bb_has_line_dir(b, n, s16);
                    const bool one = bb8_has_line_dir(b.w[0], b.w[1], W, n, d4);
                    if (wide != want || one != want) {
                        printf("mismatch W=%d n=%d dir=%d: plain %d, 256-bit %d, one-word %d\n", W, n, d4, want, wide, one);
                        return 1;
                    }
                    ++checked;
                }
            }
    printf("ok %ld\n", checked);
    return 0;
}
