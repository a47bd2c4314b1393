// gz_rollout.cu — pure MCTS: uniform priors over Board.move_target and random rollouts
// (reference mcts_pure.py:15-46, :156-212).  K4: one thread per (game, rollout) plays
// uniformly random moves inside the current bounding box until the game ends; the random
// index comes from the counter-based Philox4x32-10 stream
//     philox(key = seed, counter = (playout_idx, rollout_ply, game_id, search_id)).x
// via mulhi32(r, n_candidates) -> k-th set bit of the candidate mask.  The oracle
// (oracle/orc_mcts.c:evaluate_rollout) uses the same stream, so rollouts are bit-exact.
// Included by gz_api.cu after gz_tree.cu (single translation unit).

// _evaluate_rollout (mcts_pure.py:193-212): returns +1 / -1 / 0 from `player`'s view.
__device__ int gz_rollout(BB s0, BB s1, int ns, int player, int W, int H, int n, uint32_t k0,
                          uint32_t k1, uint32_t playout_idx, uint32_t game_id, uint32_t search_id,
                          unsigned int *plies_out)
{
    const int me = player;
    const BB full = bb_board_mask(W, H);
    int win;
    bool hw;
    bool end = bb_game_end(s0, s1, ns, W, H, n, &win, &hw);
    unsigned int plies = 0;
    for (int i = 0; i < 1000 && !end; ++i) {
        BB occ = bb_or(s0, s1);
        BB legal = bb_andnot(full, occ);
        BB tg = bb_move_target(occ, legal, W, H);
        int nt = bb_popc(tg);
        Philox4 r = philox4x32(k0, k1, playout_idx, (uint32_t)i, game_id, search_id);
        int idx = (int)__umulhi(r.x, (uint32_t)nt);
        int bit = bb_select(tg, idx);
        if (player == 0) bb_set(s0, bit); else bb_set(s1, bit);
        ++ns;
        ++plies;
        // only the side that just moved can have completed a line (the start position was
        // checked for both colours above)
        bool w = (ns >= n + 2) && bb_has_line(player == 0 ? s0 : s1, n);
        if (w) { end = true; win = player + 1; }
        else if (ns >= W * H) { end = true; win = -1; }
        player ^= 1;
    }
    if (plies_out) *plies_out = plies;
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
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
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

// rollout from every leaf of the wave (thread per game); writes leaf_value, marks state 1
__global__ void pure_rollout_kernel(TreeParams p, uint32_t k0, uint32_t k1, uint32_t playout_idx,
                                    uint32_t search_id, uint64_t game_id_base)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int plies = 0;
    if (g < p.G && p.leaf_state[g] != 2) {
        GameMeta lm = p.leaf_meta[g];
        BB s0 = load_bb(p.leaf_board[g].stones[0]), s1 = load_bb(p.leaf_board[g].stones[1]);
        int player = (lm.start_player + lm.n_stones) & 1;
        int r = gz_rollout(s0, s1, lm.n_stones, player, p.W, p.H, p.n_in_row, k0, k1, playout_idx,
                           (uint32_t)(game_id_base + (uint64_t)g), search_id, &plies);
        p.leaf_value[g] = (double)r;
        p.leaf_state[g] = 1;
    }
    plies = __reduce_add_sync(0xffffffffu, plies);
    if ((threadIdx.x & 31) == 0 && plies) atomicAdd(&p.ctr->rollout_plies, (unsigned long long)plies);
}

__global__ void __launch_bounds__(128) backup_only_kernel(TreeParams p)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (g >= p.G) return;
    if (p.leaf_state[g] == 2) return;
    gz_backup(p, g, p.path_len[g], p.leaf_value[g], lane);
}

// n playouts in every game: wave kernel + rollout kernel per playout, one trailing backup
int gzk_pure_playouts(gz_engine *e, int n, uint32_t search_id)
{
    if (n <= 0) return GZ_OK;
    TreeParams p = gz_tree_params(e);
    const unsigned wgrid = gz_warp_blocks(e->G, 4);
    for (int i = 0; i < n; ++i) {
        {
            ProfScope ps(e, PC_SELECT);
            pure_wave_kernel<<<wgrid, 128, 0, e->stream>>>(p, i > 0 ? 1 : 0);
        }
        {
            ProfScope ps(e, PC_ROLLOUT);
            pure_rollout_kernel<<<(e->G + 127) / 128, 128, 0, e->stream>>>(
                p, (uint32_t)e->cfg.seed, (uint32_t)(e->cfg.seed >> 32), (uint32_t)i, search_id, e->cfg.game_id_base);
        }
        e->launches += 2;
    }
    { ProfScope ps(e, PC_BACKUP); backup_only_kernel<<<wgrid, 128, 0, e->stream>>>(p); }
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
    unsigned int plies = 0;
    if (t < (long long)G * n_rollouts) {
        int g = (int)(t / n_rollouts), r = (int)(t % n_rollouts);
        GameMeta m = meta[g];
        BB s0 = load_bb(boards[g].stones[0]), s1 = load_bb(boards[g].stones[1]);
        int player = (m.start_player + m.n_stones) & 1;
        results[t] = (int8_t)gz_rollout(s0, s1, m.n_stones, player, W, H, n, k0, k1, (uint32_t)r,
                                        (uint32_t)(game_id_base + (uint64_t)g), search_id, &plies);
    }
    plies = __reduce_add_sync(0xffffffffu, plies);
    if ((threadIdx.x & 31) == 0 && plies) atomicAdd(&ctr->rollout_plies, (unsigned long long)plies);
}

int gzk_rollouts(gz_engine *e, int n_rollouts, uint32_t search_id, int8_t *d_results)
{
    long long total = (long long)e->G * n_rollouts;
    if (total <= 0) return GZ_OK;
    ProfScope ps(e, PC_ROLLOUT);
    rollouts_kernel<<<(unsigned)((total + 127) / 128), 128, 0, e->stream>>>(
        e->boards, e->meta, e->G, n_rollouts, e->W, e->H, e->n_in_row, (uint32_t)e->cfg.seed,
        (uint32_t)(e->cfg.seed >> 32), search_id, e->cfg.game_id_base, d_results, e->counters);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}
