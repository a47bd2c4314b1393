// gz_rollout.cu — pure MCTS: uniform priors over Board.move_target and random rollouts
// (reference mcts_pure.py:15-46, :156-212).  K4: one thread per (game, rollout) plays
// uniformly random moves inside the current bounding box until the game ends; the random
// index comes from the counter-based Philox4x32-10 stream, ONE call per four plies:
//     r = philox(key = (seed.lo, seed.hi ^ game_id.hi), counter = (playout_idx, ply / 4, game_id.lo, search_id))
//     index = mulhi32(r[ply % 4], n_candidates) -> k-th set bit of the candidate mask.
// The oracle (oracle/orc_mcts.c:evaluate_rollout) uses the same stream, so rollouts are bit-exact.
// A ply is straight-line code (the rollout is latency-bound: ~1 warp per scheduler at 16384 games):
// the bounding box only grows, so it is kept as four integers (bbox_add); the candidate mask is one
// 16-bit column pattern replicated over the box rows; the k-th candidate is found by a popcount binary
// search; only the side that just moved is tested for a line, with run lengths doubling
// (bb_has_line_n); the Philox block of the next four plies is computed while this one is consumed.
// Included by gz_api.cu after gz_tree.cu (single translation unit).

// ---- boards of at most 64 cells (6x6, 8x8: BASELINE configs C1-C3): the whole board is ONE 64-bit word -----------
// Dense packing, bit = move id = h * W + w (no guard column): the candidate mask, the k-th-candidate search and
// the line test all work on single registers, about a third of the instructions of the 256-bit path.  Without a
// guard column a shifted run can wrap into the next row, so a run of n starting at column c only counts when the
// reference's own range guards hold (game.py:144-158): E and NE need c <= W - n, NW needs c >= n - 1.
__device__ __forceinline__ int u64_select(uint64_t x, int k) // bit position of the k-th set bit (k from 0, k < popc)
{
    uint32_t y = (uint32_t)x;
    int base = 0;
    int c = __popc(y);
    if (k >= c) { k -= c; y = (uint32_t)(x >> 32); base = 32; }
#pragma unroll
    for (int sft = 16; sft >= 1; sft >>= 1) {
        const uint32_t lowm = (1u << sft) - 1u;
        c = __popc(y & lowm);
        const bool up = k >= c;
        k -= up ? c : 0;
        y = (up ? (y >> sft) : y) & lowm;
        base += up ? sft : 0;
    }
    return base;
}
template <int N> __device__ __forceinline__ uint64_t u64_runs(uint64_t b, int sft) // starts of runs of N in direction sft
{
    if (N < 2 || N > 8) { // generic: one stone per step
        uint64_t t = b;
        for (int k = 1; k < N; ++k) t = b & (t >> sft);
        return t;
    }
    const uint64_t t2 = b & (b >> sft);
    if (N == 2) return t2;
    if (N == 3) return t2 & (b >> (2 * sft));
    const uint64_t t4 = t2 & (t2 >> (2 * sft));
    if (N == 4) return t4;
    if (N == 5) return t4 & (b >> (4 * sft));
    if (N == 6) return t4 & (t2 >> (4 * sft));
    if (N == 7) return t4 & (t2 >> (4 * sft)) & (b >> (6 * sft));
    return t4 & (t4 >> (4 * sft));
}
struct U64Geom {
    uint64_t rep;     // bit 0 of every row: sum of 1 << (W * h)
    uint64_t m_e, m_w; // run starts allowed for E / NE (column <= W - n) and for NW (column >= n - 1)
    int W, H, S;
    uint32_t invW;    // (bit * invW) >> 16 == bit / W for bit < 64
};
__device__ __forceinline__ U64Geom u64_geom(int W, int H, int n)
{
    U64Geom g;
    g.W = W; g.H = H; g.S = W * H;
    g.rep = 0ull;
    for (int h = 0; h < H; ++h) g.rep |= 1ull << (W * h);
    const uint32_t roww = (1u << W) - 1u;
    const uint32_t ce = (W - n >= 0) ? ((2u << (W - n)) - 1u) : 0u;               // columns 0 .. W - n
    const uint32_t cw = (n - 1 < W) ? (roww & ~((1u << (n - 1)) - 1u)) : 0u;      // columns n - 1 .. W - 1
    g.m_e = (uint64_t)ce * g.rep;
    g.m_w = (uint64_t)cw * g.rep;
    g.invW = (65536u + (uint32_t)W - 1u) / (uint32_t)W;
    return g;
}
template <int N> __device__ __forceinline__ bool u64_has_line(uint64_t b, const U64Geom &g)
{
    const uint64_t e = u64_runs<N>(b, 1) & g.m_e, nn = u64_runs<N>(b, g.W), ne = u64_runs<N>(b, g.W + 1) & g.m_e,
                   nw = (g.W > 1) ? (u64_runs<N>(b, g.W - 1) & g.m_w) : 0ull;
    return (e | nn | ne | nw) != 0ull;
}
__device__ __forceinline__ bool u64_has_line_n(uint64_t b, int n, const U64Geom &g)
{
    switch (n) {
    case 3: return u64_has_line<3>(b, g);
    case 4: return u64_has_line<4>(b, g);
    case 5: return u64_has_line<5>(b, g);
    case 6: return u64_has_line<6>(b, g);
    default: { // any n: one stone per step (the template's generic branch needs a constant; do it here)
        bool any = false;
        const int dirs[4] = {1, g.W, g.W + 1, g.W - 1};
        for (int d = 0; d < 4; ++d) {
            if (d == 3 && g.W <= 1) continue;
            uint64_t t = b;
            for (int k = 1; k < n; ++k) t = b & (t >> dirs[d]);
            t &= (d == 1) ? ~0ull : (d == 3 ? g.m_w : g.m_e);
            any = any || t != 0ull;
        }
        return any;
    }
    }
}
__device__ __forceinline__ uint64_t u64_from_bb(const BB &a, int W, int H)
{
    uint64_t d = 0ull;
    for (int h = 0; h < H; ++h) d |= (uint64_t)(bb_row(a, h) & ((1u << W) - 1u)) << (W * h);
    return d;
}
// Board.get_move_target (game.py:35-58) on the dense word: box rows bottom..top x columns left..right, & open
__device__ __forceinline__ uint64_t u64_target(int rmin, int rmax, int cmin, int cmax, uint64_t open, const U64Geom &g)
{
    int top = 9, bottom = 5, left = 5, right = 9; // no stone: the reference's hard-coded box (:41)
    if (rmax >= 0) {
        top = (rmax <= g.H - 2) ? rmax + 2 : g.H;
        bottom = (rmin >= 2) ? rmin - 2 : 0;
        left = (cmin >= 2) ? cmin - 2 : 0;
        right = (cmax <= g.W - 2) ? cmax + 2 : g.W;
    }
    if (bottom >= g.H || left >= g.W) return 0ull;
    top = min(top, g.H - 1);
    right = min(right, g.W - 1);
    const uint32_t rowmask = ((2u << right) - 1u) & ~((1u << left) - 1u);
    const int lo = bottom * g.W, hi = (top + 1) * g.W; // bit range of the box rows, hi <= 64
    const uint64_t rows = ((hi >= 64) ? ~0ull : ((1ull << hi) - 1ull)) & ~((1ull << lo) - 1ull);
    return ((uint64_t)rowmask * g.rep) & rows & open;
}

__device__ int gz_rollout_u64(const BB &s0b, const BB &s1b, int ns, int player, int W, int H, int n, uint32_t k0,
                              uint32_t k1, uint32_t playout_idx, uint32_t gid, uint32_t search_id,
                              unsigned int *plies_out, unsigned int *empty_out)
{
    const int me = player;
    const U64Geom g = u64_geom(W, H, n);
    const uint64_t s0 = u64_from_bb(s0b, W, H), s1 = u64_from_bb(s1b, W, H);
    int win = -1;
    bool end = false;
    if (ns >= n + 2) { // has_a_winner's early-out (game.py:136-137); the start position is tested for both colours
        const bool w0 = u64_has_line_n(s0, n, g), w1 = u64_has_line_n(s1, n, g);
        if (w0 || w1) { end = true; win = w0 ? 1 : 2; }
    }
    if (!end && ns >= g.S) end = true; // tie: win stays -1
    uint64_t own = player == 0 ? s0 : s1, opp = player == 0 ? s1 : s0; // own = side to move
    const uint64_t full = (g.S >= 64) ? ~0ull : ((1ull << g.S) - 1ull);
    uint64_t open = full & ~(s0 | s1);
    int rmin = 0, rmax = -1, cmin = 0, cmax = 0;
    {
        const uint64_t occ = s0 | s1;
        if (occ) {
            rmin = (int)(((uint32_t)(__ffsll((long long)occ) - 1) * g.invW) >> 16);
            rmax = (int)(((uint32_t)(63 - __clzll((long long)occ)) * g.invW) >> 16);
            uint32_t cols = 0u;
            for (int h = 0; h < H; ++h) cols |= (uint32_t)(occ >> (W * h));
            cols &= (1u << W) - 1u;
            cmin = __ffs((int)cols) - 1;
            cmax = 31 - __clz((int)cols);
        }
    }
    unsigned int plies = 0, empty = 0;
    if (!end) {
        Philox4 cur = philox4x32(k0, k1, playout_idx, 0u, gid, search_id);
        for (uint32_t blk = 0; blk < 250u && !end; ++blk) { // limit=1000 plies (mcts_pure.py:193)
            const Philox4 nxt = philox4x32(k0, k1, playout_idx, blk + 1u, gid, search_id);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (end) break;
                const uint64_t tg = u64_target(rmin, rmax, cmin, cmax, open, g);
                const int nt = __popcll(tg);
                if (nt == 0) { empty = 1u; end = true; win = -1; break; } // see gz_rollout
                const uint32_t r = (j == 0) ? cur.x : (j == 1) ? cur.y : (j == 2) ? cur.z : cur.w;
                const int bit = u64_select(tg, (int)__umulhi(r, (uint32_t)nt));
                const uint64_t m = 1ull << bit;
                own |= m;
                open &= ~m;
                const int rr = (int)(((uint32_t)bit * g.invW) >> 16), cc = bit - rr * W;
                if (rmax < 0) { rmin = rmax = rr; cmin = cmax = cc; }
                else { rmin = min(rmin, rr); rmax = max(rmax, rr); cmin = min(cmin, cc); cmax = max(cmax, cc); }
                ++ns;
                ++plies;
                if (ns >= n + 2 && u64_has_line_n(own, n, g)) { end = true; win = player + 1; }
                else if (ns >= g.S) { end = true; win = -1; }
                const uint64_t t = own; own = opp; opp = t;
                player ^= 1;
            }
            cur = nxt;
        }
    }
    if (plies_out) *plies_out = plies;
    if (empty_out) *empty_out = empty;
    if (!end || win == -1) return 0;
    return (win == me + 1) ? 1 : -1;
}

// _evaluate_rollout (mcts_pure.py:193-212): returns +1 / -1 / 0 from `player`'s view.
__device__ int gz_rollout(BB s0, BB s1, int ns, int player, int W, int H, int n, uint32_t k0,
                          uint32_t k1, uint32_t playout_idx, uint64_t game_id, uint32_t search_id,
                          unsigned int *plies_out, unsigned int *empty_out)
{
    const int me = player;
    const uint32_t gid = (uint32_t)game_id;
    k1 ^= (uint32_t)(game_id >> 32);
    if (W * H <= 64 && n >= 2) // one-word boards (same stream, same candidate order: ascending move id)
        return gz_rollout_u64(s0, s1, ns, player, W, H, n, k0, k1, playout_idx, gid, search_id, plies_out, empty_out);
    int win;
    bool hw;
    bool end = bb_game_end(s0, s1, ns, W, H, n, &win, &hw); // the start position: both colours
    BB own = bb_pick(s0, s1, player == 0), opp = bb_pick(s1, s0, player == 0); // own = side to move
    BB open = bb_andnot(bb_board_mask(W, H), bb_or(s0, s1));
    BBox bx = bb_bbox(bb_or(s0, s1));
    const int S = W * H;
    unsigned int plies = 0, empty = 0;
    if (!end) {
        Philox4 cur = philox4x32(k0, k1, playout_idx, 0u, gid, search_id);
        for (uint32_t blk = 0; blk < 250u && !end; ++blk) { // limit=1000 plies (mcts_pure.py:193)
            const Philox4 nxt = philox4x32(k0, k1, playout_idx, blk + 1u, gid, search_id);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (end) break;
                const BB tg = bb_target_from_bbox(bx, open, W, H);
                const int nt = bb_popc(tg);
                if (nt == 0) { // empty board narrower than 6: the reference's 5..9 box holds no cell (game.py:41)
                    empty = 1u; end = true; win = -1;
                    break;
                }
                const uint32_t r = (j == 0) ? cur.x : (j == 1) ? cur.y : (j == 2) ? cur.z : cur.w;
                const int bit = bb_select(tg, (int)__umulhi(r, (uint32_t)nt));
                const int wi = bit >> 6;
                const uint64_t m = 1ull << (bit & 63);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint64_t mi = (wi == i) ? m : 0ull;
                    own.w[i] |= mi;
                    open.w[i] &= ~mi;
                }
                bbox_add(bx, bit);
                ++ns;
                ++plies;
                // only the side that just moved can have completed a line
                if (ns >= n + 2 && bb_has_line_n(own, n)) { end = true; win = player + 1; }
                else if (ns >= S) { end = true; win = -1; }
                const BB t = own; own = opp; opp = t;
                player ^= 1;
            }
            cur = nxt;
        }
    }
    if (plies_out) *plies_out = plies;
    if (empty_out) *empty_out = empty;
    if (!end || win == -1) return 0;
    return (win == me + 1) ? 1 : -1;
}

// expand a pure-MCTS leaf: children = move_target cells, P slot = k (prior 1/k, float64)
__device__ __forceinline__ void gz_pure_expand_game(const TreeParams &p, int g, int lane)
{
    if (p.leaf_state[g] != 0) return;
    const int d = p.path_len[g];
    BB occ = bb_or(load_bb(p.leaf_board[g].stones[0]), load_bb(p.leaf_board[g].stones[1]));
    BB legal = bb_andnot(bb_board_mask(p.W, p.H), occ);
    BB tg = bb_move_target(occ, legal, p.W, p.H);
    const int k = bb_popc(tg);
    if (k == 0) { // empty board narrower than 6 (game.py:41): nothing to expand, the rollout reports it
        if (lane == 0) atomicAdd(&p.ctr->empty_box, 1ull);
        return;
    }
    int n = gz_alloc_node(p, g, lane);
    if (n < 0) return;
    NodeView nv = gz_node_view(p.pool, (size_t)g * p.cap + n, p.Sp);
    for (int i = lane; i < p.Sp; i += GZ_WARP) {
        bool in = (i < p.S) && bb_test(tg, bb_bit(i, p.W));
        nv.P[i] = in ? (float)k : -1.f;
        nv.N[i] = 0u;
        nv.Q[i] = 0.0;
        nv.child[i] = -1;
    }
    gz_link_node(p, g, d, n, lane);
    if (lane == 0) atomicAdd(&p.ctr->expanded_children, (unsigned long long)k);
}

// One pure-MCTS wave boundary per game and warp: backup of the previous playout's rollout result (when
// there is one), descent to the next leaf, expansion of that leaf (mcts_pure.py:156-191 expands before it
// rolls out; the rollout does not look at the tree).  With pure_rollout_kernel this makes a playout two
// launches instead of four and a memset.  Plain stream-ordered launches: in a dependent-launch chain the
// early-resident wave blocks (24 warps per SM parked in griddepcontrol.wait) slow the 4-warps-per-SM
// rollout kernel down by 20-25 % (94 vs 114 M playouts/s at 6x6).
__global__ void __launch_bounds__(128, 6) pure_wave_kernel(TreeParams p, int do_backup)
{
    const int lane = threadIdx.x & 31;
    const int g = p.g0 + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    __shared__ SelStats ss;
    if (threadIdx.x == 0) { ss.levels = 0; ss.children = 0; ss.playouts = 0; ss.done = 0; }
    __syncthreads();
    unsigned lv = 0, ch = 0, pl = 0;
    if (g < p.G) {
        if (do_backup && p.leaf_state[g] != 2) gz_backup(p, g, p.path_len[g], p.leaf_value[g], lane);
        __syncwarp(); // the descent reads the statistics other lanes have just written
        gz_select_game<true>(p, g, lane, lv, ch, pl);
        __syncwarp(); // leaf state / board / path length written by lane 0
        gz_pure_expand_game(p, g, lane);
    }
    gz_stats_finish(&ss, p.ctr, lane, lv, ch, pl);
}

// The same wave boundary written around dependent memory round trips like the AlphaZero kernel
// (gz_expand_select_game): trip 1 = everything addressed by the game index (flags, path length, the
// rollout's value, root, free-list top, position, this lane's path entry), trip 2 = everything addressed by those
// (this lane's path statistics, the free node index - fetched whether or not the leaf will be expanded - and the
// whole root block); backup from registers with the root block patched, one trip per level below the root, and
// the new leaf's block written from registers (uniform priors over Board.move_target).  Same arithmetic in the
// same order per node as gz_backup + gz_select_game<true> + gz_pure_expand_game: bit-identical trees.
template <int NSL, int WT>
__device__ __forceinline__ void gz_pure_wave_game(const TreeParams &p, int g, int lane, int do_backup, unsigned &levels_out,
                                                  unsigned &children_out, unsigned &playouts_out)
{
    constexpr int SP = 32 * NSL;
    const int BW = WT > 0 ? WT : p.W, BH = WT > 0 ? WT : p.H, BS = WT > 0 ? WT * WT : p.S;
    levels_out = children_out = playouts_out = 0;
    const size_t nbase = (size_t)g * p.cap;
    // ---- trip 1
    const int active = p.active[g];
    const int st = p.leaf_state[g];
    const int d = p.path_len[g];
    const int top = p.free_top[g];
    const int root = p.root_node[g];
    uint32_t rootN = p.root_N[g];
    const double rootQ = p.root_Q[g];
    const double lval = p.leaf_value[g];
    const GameMeta m = p.meta[g];
    // one 64-bit word of the position per lane on the way down (lane l < 8: word l & 3 of colour l >> 2), as in
    // gz_expand_select_game; the boards are put together again at the leaf
    uint64_t bw = lane < 8 ? p.boards[g].stones[lane >> 2][lane & 3] : 0ull;
    const uint32_t e_mine = p.path[(size_t)g * BS + (lane < BS ? lane : 0)];
    // ---- trip 2
    const bool do_bk = do_backup && st != 2;
    const bool has_e = do_bk && lane < d;
    NodeView ev = gz_node_view(p.pool, nbase + (has_e ? e_mine / SP : 0u), SP);
    const int cell_e = (int)(e_mine % SP);
    uint32_t n_e = 0u;
    double q_e = 0.0;
    if (has_e) { n_e = ev.N[cell_e]; q_e = ev.Q[cell_e]; }
    int newn = -1;
    if (active && top > 0) newn = p.free_stack[nbase + top - 1];
    float rP[NSL];
    uint32_t rN[NSL];
    double rQ[NSL];
    int rC[NSL];
    if (root >= 0 && active) {
        NodeView rv = gz_node_view(p.pool, nbase + root, SP);
#pragma unroll
        for (int j = 0; j < NSL; ++j) {
            const int i = lane + 32 * j;
            if (i < BS) { rP[j] = rv.P[i]; rN[j] = rv.N[i]; rQ[j] = rv.Q[i]; rC[j] = rv.child[i]; }
            else rP[j] = -1.f; // padding slots are never children: their sectors are not fetched
        }
    }
    // ---- backup of the previous playout's rollout value (mcts_pure.py:87-103, :191)
    if (do_bk) {
        const double v = lval;
        uint32_t n_new = 0u;
        double q_new = 0.0;
        if (has_e) {
            const double val = ((d - (lane + 1)) & 1) ? v : -v;
            n_new = n_e + 1u;
            q_new = q_e + 1.0 * (val - q_e) / (double)n_new;
            ev.N[cell_e] = n_new;
            ev.Q[cell_e] = q_new;
        }
        for (int k = lane + 32; k < d; k += 32) {
            const uint32_t e = p.path[(size_t)g * BS + k];
            NodeView nv = gz_node_view(p.pool, nbase + e / SP, SP);
            const int cell = (int)(e % SP);
            const double val = ((d - (k + 1)) & 1) ? v : -v;
            const uint32_t n = nv.N[cell] + 1u;
            nv.N[cell] = n;
            const double q = nv.Q[cell];
            nv.Q[cell] = q + 1.0 * (val - q) / (double)n;
        }
        {
            const double val = (d & 1) ? v : -v;
            rootN += 1u;
            if (lane == 0) {
                p.root_N[g] = rootN;
                p.root_Q[g] = rootQ + 1.0 * (val - rootQ) / (double)rootN;
            }
        }
        if (d >= 1) { // path entry 0 is an edge of the prefetched root block
            const int c0 = __shfl_sync(0xffffffffu, cell_e, 0);
            const uint32_t nn = __shfl_sync(0xffffffffu, n_new, 0);
            const double qq = __shfl_sync(0xffffffffu, q_new, 0);
#pragma unroll
            for (int j = 0; j < NSL; ++j)
                if (lane + 32 * j == c0) { rN[j] = nn; rQ[j] = qq; }
        }
    }
    __syncwarp();
    if (!active) {
        if (lane == 0) { p.leaf_state[g] = 2; p.path_len[g] = 0; }
        return;
    }
    // ---- select (mcts_pure.py:79-85, :105-113): prior 1 / k in float64
    int player = (m.start_player + m.n_stones) & 1;
    int ns = m.n_stones, last = m.last_move;
    int node = root, prev_node = -1, prev_mv = 0;
    uint32_t Np = rootN;
    int depth = 0;
    unsigned int nchild = 0;
    while (node >= 0) {
        const double sq = gz_sqrt_visits(Np);
        double best = -1.0e300;
        int besti = 0x7fffffff, bestc = -1;
        uint32_t bestn = 0u;
        // every child of a node carries the same prior 1 / k (mcts_pure.py:45): one division per lane and level
        float kf = -1.f;
#pragma unroll
        for (int j = 0; j < NSL; ++j) kf = rP[j] >= 0.f ? rP[j] : kf;
        const double cp = (double)p.c_puct * (1.0 / (double)kf), cpsq = cp * sq;
#pragma unroll
        for (int j = 0; j < NSL; ++j) {
            if (rP[j] >= 0.f) {
                const double sc = rQ[j] + cpsq / (double)(1u + rN[j]);
                if (sc > best) { best = sc; besti = lane + 32 * j; bestn = rN[j]; bestc = rC[j]; }
                ++nchild;
            }
        }
        const int mv = warp_argmax_redux(best, besti);
        const int owner = mv & (GZ_WARP - 1);
        Np = __shfl_sync(0xffffffffu, bestn, owner);
        const int next = __shfl_sync(0xffffffffu, bestc, owner);
        if (lane == 0) p.path[(size_t)g * BS + depth] = (uint32_t)node * SP + mv;
        ++depth;
        if (next >= 0) {
            NodeView v = gz_node_view(p.pool, nbase + next, SP);
#pragma unroll
            for (int j = 0; j < NSL; ++j) {
                const int i = lane + 32 * j;
                if (i < BS) { rP[j] = v.P[i]; rN[j] = v.N[i]; rQ[j] = v.Q[i]; rC[j] = v.child[i]; }
                else rP[j] = -1.f;
            }
        }
        {
            const int bit = bb_bit(mv, BW);
            if (lane == 4 * player + (bit >> 6)) bw |= 1ull << (bit & 63);
        }
        player ^= 1;
        last = mv;
        ++ns;
        prev_node = node;
        prev_mv = mv;
        node = next;
    }
    BB s0, s1;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        s0.w[k] = __shfl_sync(0xffffffffu, bw, k);
        s1.w[k] = __shfl_sync(0xffffffffu, bw, 4 + k);
    }
    bool w0 = false, w1 = false;
    if (ns >= p.n_in_row + 2) {
        const int d4 = lane & 3;
        const int dir = (d4 == 0) ? 1 : (d4 == 1) ? 16 : (d4 == 2) ? 17 : 15;
        bool line = false;
        if (WT > 0 && WT <= 8) { // one squeezed 64-bit word per colour (gz_common.cuh: bb8_has_line_dir)
            if (lane < 8) line = bb8_has_line_dir(lane < 4 ? s0.w[0] : s1.w[0], lane < 4 ? s0.w[1] : s1.w[1], BW, p.n_in_row, d4);
        } else if (lane < 8) line = bb_has_line_dir(bb_pick(s0, s1, lane < 4), p.n_in_row, dir);
        const unsigned b = __ballot_sync(0xffffffffu, line);
        w0 = (b & 0x0Fu) != 0u;
        w1 = (b & 0xF0u) != 0u;
    }
    const int win = w0 ? 1 : (w1 ? 2 : -1);
    const bool end = (w0 || w1) || ns >= BW * BH;
    if (lane == 0) {
        p.path_len[g] = depth;
        GameMeta lm;
        lm.n_stones = ns; lm.last_move = last; lm.start_player = m.start_player; lm.flags = 0;
        p.leaf_meta[g] = lm;
        store_bb(p.leaf_board[g].stones[0], s0);
        store_bb(p.leaf_board[g].stones[1], s1);
        p.leaf_state[g] = end ? 1 : 0;
        if (end) p.leaf_value[g] = (win == -1) ? 0.0 : ((win == player + 1) ? 1.0 : -1.0);
    }
    // ---- expand the new leaf (mcts_pure.py:181-183 with policy_value_fn :33-46)
    if (!end) {
        const BB occ = bb_or(s0, s1);
        const BB tg = bb_move_target(occ, bb_andnot(bb_board_mask(BW, BH), occ), BW, BH);
        const int k = bb_popc(tg);
        if (k == 0) {
            if (lane == 0) atomicAdd(&p.ctr->empty_box, 1ull);
        } else if (newn < 0) {
            if (lane == 0) atomicAdd(&p.ctr->pool_overflows, 1ull);
        } else {
            NodeView nv = gz_node_view(p.pool, nbase + newn, SP);
#pragma unroll
            for (int j = 0; j < NSL; ++j) {
                const int i = lane + 32 * j;
                const bool in = (i < BS) && bb_test(tg, bb_bit(i, BW));
                nv.P[i] = in ? (float)k : -1.f;
                nv.N[i] = 0u; nv.Q[i] = 0.0; nv.child[i] = -1;
            }
            if (lane == 0) {
                if (depth == 0) p.root_node[g] = newn;
                else gz_node_view(p.pool, nbase + prev_node, SP).child[prev_mv] = newn;
                p.free_top[g] = top - 1;
                atomicAdd(&p.ctr->expanded_children, (unsigned long long)k);
            }
        }
    }
    levels_out = (unsigned)depth;
    children_out = __reduce_add_sync(0xffffffffu, nchild);
    playouts_out = 1u;
}

#ifndef GZ_PW_MINB
#define GZ_PW_MINB 7
#endif
template <int NSL, int WT>
__global__ void __launch_bounds__(128, NSL <= 2 ? GZ_PW_MINB : (NSL <= 4 ? 4 : 3)) pure_wave_small_kernel(TreeParams p, int do_backup)
{
    const int lane = threadIdx.x & 31;
    const int g = p.g0 + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    __shared__ SelStats ss;
    if (threadIdx.x == 0) { ss.levels = 0; ss.children = 0; ss.playouts = 0; ss.done = 0; }
    __syncthreads();
    unsigned lv = 0, ch = 0, pl = 0;
    if (g < p.G) gz_pure_wave_game<NSL, WT>(p, g, lane, do_backup, lv, ch, pl);
    gz_stats_finish(&ss, p.ctr, lane, lv, ch, pl);
}

// rollout from every leaf of the wave (thread per game); writes leaf_value, marks state 1
__global__ void pure_rollout_kernel(TreeParams p, uint32_t k0, uint32_t k1, uint32_t playout_idx,
                                    uint32_t search_id, uint64_t game_id_base)
{
    const int g = p.g0 + blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int plies = 0, empty = 0;
    if (g < p.G && p.leaf_state[g] != 2) {
        GameMeta lm = p.leaf_meta[g];
        BB s0 = load_bb(p.leaf_board[g].stones[0]), s1 = load_bb(p.leaf_board[g].stones[1]);
        int player = (lm.start_player + lm.n_stones) & 1;
        int r = gz_rollout(s0, s1, lm.n_stones, player, p.W, p.H, p.n_in_row, k0, k1, playout_idx,
                           game_id_base + (uint64_t)g, search_id, &plies, &empty);
        p.leaf_value[g] = (double)r;
        p.leaf_state[g] = 1;
    }
    plies = __reduce_add_sync(0xffffffffu, plies);
    empty = __reduce_add_sync(0xffffffffu, empty);
    if ((threadIdx.x & 31) == 0 && plies) atomicAdd(&p.ctr->rollout_plies, (unsigned long long)plies);
    if ((threadIdx.x & 31) == 0 && empty) atomicAdd(&p.ctr->empty_box, (unsigned long long)empty);
}

__global__ void __launch_bounds__(128) backup_only_kernel(TreeParams p)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (g >= p.G) return;
    if (p.leaf_state[g] == 2) return;
    gz_backup(p, g, p.path_len[g], p.leaf_value[g], lane);
}

// Threads per block of the thread-per-rollout kernels: single warps while they all fit on the SMs at once
// (a rollout kernel of 16384 games is 512 warps: 32-thread blocks put 3-4 on every SM, one per scheduler;
// 128-thread blocks would leave 20 SMs idle and stack four warps on one scheduler pair elsewhere).
static int gz_rollout_block(const gz_engine *e, long long threads)
{
    return threads <= (long long)(e->n_sm > 0 ? e->n_sm : 148) * 32 * 24 ? 32 : 128;
}

// n playouts in every game: wave kernel + rollout kernel per playout, one trailing backup.
// The games are cut into 4 parts that run the same sequence on their own streams, one wave apart: the rollout
// kernel is a few warps per SM waiting on their own instruction latency (5 % of the warp slots, ncu), the wave
// kernel is bound by how many warps the SMs hold - so the rollouts of one part run underneath the waves of the
// others.  The rollout kernels go to an urgent-priority stream (its few blocks take the next free slots instead
// of queueing behind a whole wave grid).  Measured at 16384 games (profiles/r2_pure_*): 6x6 one stream 269 M
// playouts/s, 2 parts 307 M, 3 parts 339 M, 4 parts 346 M; without the priority 314 M.  Games are independent
// and each part sees exactly the launches it saw before: results are bit-identical.
// GZ_PURE_PARTS=1..4 / GZ_PURE_PRIO=0 are A/B switches.
static int gz_pure_streams(gz_engine *e)
{
    if (e->pure_st[0]) return GZ_OK;
    int lo = 0, hi = 0;
    GZ_CUDA(e, cudaDeviceGetStreamPriorityRange(&lo, &hi)); // hi = numerically lowest = most urgent
    for (int i = 0; i < GZ_PURE_MAXPARTS; ++i) {
        GZ_CUDA(e, cudaStreamCreateWithPriority(&e->pure_st[i], cudaStreamNonBlocking, lo));
        GZ_CUDA(e, cudaStreamCreateWithPriority(&e->pure_hi[i], cudaStreamNonBlocking, hi));
        for (cudaEvent_t *ev : {&e->pure_ev[i][0], &e->pure_ev[i][1], &e->pure_ev[i][2], &e->pure_ev[i][3]})
            GZ_CUDA(e, cudaEventCreateWithFlags(ev, cudaEventDisableTiming));
    }
    GZ_CUDA(e, cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
    return GZ_OK;
}

static void gz_pure_launch_wave(gz_engine *e, const TreeParams &p, unsigned wgrid, int bk, cudaStream_t st)
{
    const bool sq = e->W == e->H;
    static const bool generic = getenv("GZ_PURE_GENERIC") != nullptr; // A/B switch: the un-fused wave kernel
#define GZ_PW(NSL, WT) pure_wave_small_kernel<NSL, WT><<<wgrid, 128, 0, st>>>(p, bk)
    if (generic) pure_wave_kernel<<<wgrid, 128, 0, st>>>(p, bk);
    else if (e->Sp == 64 && sq && e->W == 6) GZ_PW(2, 6);
    else if (e->Sp == 64 && sq && e->W == 8) GZ_PW(2, 8);
    else if (e->Sp == 32) GZ_PW(1, 0);
    else if (e->Sp == 64) GZ_PW(2, 0);
    else if (e->Sp == 96) GZ_PW(3, 0);
    else if (e->Sp == 128) GZ_PW(4, 0);
    else if (e->Sp == 256) GZ_PW(8, 0);
    else pure_wave_kernel<<<wgrid, 128, 0, st>>>(p, bk);
#undef GZ_PW
}

int gzk_pure_playouts(gz_engine *e, int n, uint32_t search_id)
{
    if (n <= 0) return GZ_OK;
    static const int parts_env = getenv("GZ_PURE_PARTS") ? atoi(getenv("GZ_PURE_PARTS")) : 0;
    static const int prio_env = getenv("GZ_PURE_PRIO") ? atoi(getenv("GZ_PURE_PRIO")) : 1;
    int parts = parts_env > 0 ? parts_env : 4;
    if (parts > GZ_PURE_MAXPARTS) parts = GZ_PURE_MAXPARTS;
    if (e->G < 2048 || n < 2) parts = 1;
    if (parts == 1) { // everything on the engine stream
        TreeParams p = gz_tree_params(e);
        const int rtpb = gz_rollout_block(e, e->G);
        for (int i = 0; i < n; ++i) {
            { ProfScope ps(e, PC_SELECT); gz_pure_launch_wave(e, p, gz_warp_blocks(e->G, 4), i > 0 ? 1 : 0, e->stream); }
            {
                ProfScope ps(e, PC_ROLLOUT);
                pure_rollout_kernel<<<(e->G + rtpb - 1) / rtpb, rtpb, 0, e->stream>>>(
                    p, (uint32_t)e->cfg.seed, (uint32_t)(e->cfg.seed >> 32), (uint32_t)i, search_id, e->cfg.game_id_base);
            }
            e->launches += 2;
        }
    } else {
        GZ_TRY(gz_pure_streams(e));
        TreeParams pp[GZ_PURE_MAXPARTS];
        const int per = (((e->G + parts - 1) / parts) + 3) & ~3; // whole blocks of 4 games
        for (int h = 0; h < parts; ++h) {
            pp[h] = gz_tree_params(e);
            pp[h].g0 = min(h * per, e->G);
            pp[h].G = min((h + 1) * per, e->G);
        }
        GZ_CUDA(e, cudaEventRecord(e->ev_fork, e->stream));
        for (int h = 0; h < parts; ++h) GZ_CUDA(e, cudaStreamWaitEvent(e->pure_st[h], e->ev_fork, 0));
        for (int i = 0; i < n; ++i) {
            for (int h = 0; h < parts; ++h) {
                const TreeParams &p = pp[h];
                const int cnt = p.G - p.g0;
                if (cnt <= 0) continue;
                cudaStream_t sw = e->pure_st[h], sr = prio_env ? e->pure_hi[h] : sw;
                if (i == 0 && h > 0) GZ_CUDA(e, cudaStreamWaitEvent(sw, e->pure_ev[h - 1][2], 0)); // start one wave later than the part before
                gz_pure_launch_wave(e, p, gz_warp_blocks(cnt, 4), i > 0 ? 1 : 0, sw);
                if (i == 0) GZ_CUDA(e, cudaEventRecord(e->pure_ev[h][2], sw));
                if (sr != sw) { // the rollout of this part on the urgent stream: its few blocks get the next free slots
                    GZ_CUDA(e, cudaEventRecord(e->pure_ev[h][0], sw));
                    GZ_CUDA(e, cudaStreamWaitEvent(sr, e->pure_ev[h][0], 0));
                }
                const int rtpb = gz_rollout_block(e, cnt);
                pure_rollout_kernel<<<(cnt + rtpb - 1) / rtpb, rtpb, 0, sr>>>(
                    p, (uint32_t)e->cfg.seed, (uint32_t)(e->cfg.seed >> 32), (uint32_t)i, search_id, e->cfg.game_id_base);
                if (sr != sw) {
                    GZ_CUDA(e, cudaEventRecord(e->pure_ev[h][1], sr));
                    GZ_CUDA(e, cudaStreamWaitEvent(sw, e->pure_ev[h][1], 0));
                }
                e->launches += 2;
            }
        }
        for (int h = 0; h < parts; ++h) {
            GZ_CUDA(e, cudaEventRecord(e->pure_ev[h][3], e->pure_st[h]));
            GZ_CUDA(e, cudaStreamWaitEvent(e->stream, e->pure_ev[h][3], 0));
        }
    }
    { ProfScope ps(e, PC_BACKUP); backup_only_kernel<<<gz_warp_blocks(e->G, 4), 128, 0, e->stream>>>(gz_tree_params(e)); }
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

// rollouts straight from the resident boards: results[g][r]
__global__ void rollouts_kernel(const GameBoard *boards, const GameMeta *meta, int G, int n_rollouts,
                                int W, int H, int n, uint32_t k0, uint32_t k1, uint32_t search_id,
                                uint64_t game_id_base, int8_t *results, DevCounters *ctr)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int plies = 0, empty = 0;
    if (t < (long long)G * n_rollouts) {
        int g = (int)(t / n_rollouts), r = (int)(t % n_rollouts);
        GameMeta m = meta[g];
        BB s0 = load_bb(boards[g].stones[0]), s1 = load_bb(boards[g].stones[1]);
        int player = (m.start_player + m.n_stones) & 1;
        results[t] = (int8_t)gz_rollout(s0, s1, m.n_stones, player, W, H, n, k0, k1, (uint32_t)r,
                                        game_id_base + (uint64_t)g, search_id, &plies, &empty);
    }
    plies = __reduce_add_sync(0xffffffffu, plies);
    empty = __reduce_add_sync(0xffffffffu, empty);
    if ((threadIdx.x & 31) == 0 && plies) atomicAdd(&ctr->rollout_plies, (unsigned long long)plies);
    if ((threadIdx.x & 31) == 0 && empty) atomicAdd(&ctr->empty_box, (unsigned long long)empty);
}

int gzk_rollouts(gz_engine *e, int n_rollouts, uint32_t search_id, int8_t *d_results)
{
    long long total = (long long)e->G * n_rollouts;
    if (total <= 0) return GZ_OK;
    ProfScope ps(e, PC_ROLLOUT);
    const int tpb = gz_rollout_block(e, total);
    rollouts_kernel<<<(unsigned)((total + tpb - 1) / tpb), tpb, 0, e->stream>>>(
        e->boards, e->meta, e->G, n_rollouts, e->W, e->H, e->n_in_row, (uint32_t)e->cfg.seed,
        (uint32_t)(e->cfg.seed >> 32), search_id, e->cfg.game_id_base, d_results, e->counters);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}
