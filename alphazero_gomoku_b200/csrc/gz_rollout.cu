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

// _evaluate_rollout (mcts_pure.py:193-212): returns +1 / -1 / 0 from `player`'s view.
__device__ int gz_rollout(BB s0, BB s1, int ns, int player, int W, int H, int n, uint32_t k0,
                          uint32_t k1, uint32_t playout_idx, uint64_t game_id, uint32_t search_id,
                          unsigned int *plies_out, unsigned int *empty_out)
{
    const int me = player;
    const uint32_t gid = (uint32_t)game_id;
    k1 ^= (uint32_t)(game_id >> 32);
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

// n playouts in every game: wave kernel + rollout kernel per playout, one trailing backup
int gzk_pure_playouts(gz_engine *e, int n, uint32_t search_id)
{
    if (n <= 0) return GZ_OK;
    TreeParams p = gz_tree_params(e);
    const unsigned wgrid = gz_warp_blocks(e->G, 4);
    const int rtpb = gz_rollout_block(e, e->G);
    for (int i = 0; i < n; ++i) {
        {
            ProfScope ps(e, PC_SELECT);
            pure_wave_kernel<<<wgrid, 128, 0, e->stream>>>(p, i > 0 ? 1 : 0);
        }
        {
            ProfScope ps(e, PC_ROLLOUT);
            pure_rollout_kernel<<<(e->G + rtpb - 1) / rtpb, rtpb, 0, e->stream>>>(
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
