// gz_move.cu — K6 root move selection and K7 self-play bookkeeping.
//   gz_root_select: MCTSPlayer.get_action after the search (mcts_alphaZero.py:212-262):
//     pi = softmax(log(visits + 1e-10) / temp) in float64 (:179), self-play move ~
//     0.75 pi + 0.25 Dirichlet(0.1) (:236-240), match play move ~ pi (:247); pure MCTS
//     takes the most visited child, first maximum (mcts_pure.py:230).
//   selfplay_advance: Game.start_self_play's loop body (game.py:371-416): record
//     (state, pi, player), do_move, update_with_move (tree reuse), game_end, and on the end
//     of a game winners_z (:396-399), sample emission, reset_player / init_board.
// Included by gz_api.cu after gz_tree.cu (single translation unit).

__device__ __forceinline__ double gz_u01(uint32_t r) { return ((double)r + 0.5) * (1.0 / 4294967296.0); }

// Gamma(alpha, 1) by Marsaglia-Tsang (with the U^(1/alpha) boost for alpha < 1), Philox-driven.
__device__ double gz_gamma(double alpha, uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1,
                           uint32_t c2, uint32_t tag)
{
    const bool boost = alpha < 1.0;
    const double a = boost ? alpha + 1.0 : alpha;
    const double d = a - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * d);
    double g = d;
    for (uint32_t attempt = 0; attempt < 64u; ++attempt) {
        Philox4 r = philox4x32(k0, k1, c0, c1, c2, tag + attempt);
        double u1 = gz_u01(r.x), u2 = gz_u01(r.y), u3 = gz_u01(r.z);
        double x = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
        double v = 1.0 + c * x;
        if (v <= 0.0) continue;
        v = v * v * v;
        if (log(u3) < 0.5 * x * x + d - d * v + d * log(v)) {
            g = d * v;
            if (boost) g *= pow(gz_u01(r.w), 1.0 / alpha);
            break;
        }
    }
    return g;
}

struct MoveParams {
    float temp;
    int selfplay;     // 1: Dirichlet-mixed sampling
    int best_visits;  // 1: argmax visits (pure MCTS)
    float alpha, eps;
    uint32_t k0, k1;
    const int64_t *game_id; // [G] or NULL -> game_id_base + g
    uint64_t game_id_base;
};

#define GZ_MAXCH 8 // Sp <= 256 -> at most 8 cells per lane

__global__ void __launch_bounds__(128)
root_select_kernel(TreeParams p, MoveParams mp, int32_t *moves, float *probs)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (g >= p.G) return;
    if (!p.active[g]) {
        if (lane == 0) moves[g] = -2;
        return;
    }
    const int r = p.root_node[g];
    if (r < 0) { // no children: "the board is full" branch (mcts_alphaZero.py:261-262)
        if (lane == 0) moves[g] = -1;
        if (probs) for (int i = lane; i < p.S; i += GZ_WARP) probs[(size_t)g * p.S + i] = 0.f;
        return;
    }
    NodeView v = gz_node_view(p.pool, (size_t)g * p.cap + r, p.Sp);
    const int nch = p.Sp / GZ_WARP;
    double x[GZ_MAXCH];
    int vis[GZ_MAXCH];
    bool isch[GZ_MAXCH];
    double mx = -1.0e300;
    int bestv = -1, besti = 0x7fffffff;
    const double invt = 1.0 / (double)mp.temp;
#pragma unroll
    for (int j = 0; j < GZ_MAXCH; ++j) {
        isch[j] = false; x[j] = 0.0; vis[j] = 0;
        if (j < nch) {
            int i = j * GZ_WARP + lane;
            if (v.P[i] >= 0.f) {
                isch[j] = true;
                vis[j] = (int)v.N[i];
                x[j] = invt * log((double)vis[j] + 1e-10);
                mx = fmax(mx, x[j]);
                if (vis[j] > bestv) { bestv = vis[j]; besti = i; }
            }
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, off));
        int ov = __shfl_xor_sync(0xffffffffu, bestv, off);
        int oi = __shfl_xor_sync(0xffffffffu, besti, off);
        if (ov > bestv || (ov == bestv && oi < besti)) { bestv = ov; besti = oi; }
    }
    double sum = 0.0;
#pragma unroll
    for (int j = 0; j < GZ_MAXCH; ++j)
        if (isch[j]) { x[j] = exp(x[j] - mx); sum += x[j]; }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
#pragma unroll
    for (int j = 0; j < GZ_MAXCH; ++j) {
        if (j < nch) {
            int i = j * GZ_WARP + lane;
            x[j] = isch[j] ? x[j] / sum : 0.0;
            if (probs && i < p.S) probs[(size_t)g * p.S + i] = (float)x[j];
        }
    }
    if (mp.best_visits) {
        if (lane == 0) moves[g] = besti;
        return;
    }
    const uint32_t ply = (uint32_t)p.meta[g].n_stones;
    // both halves of the 64-bit game id enter the stream: low word in the counter, high word in the key
    // (ranks use game_id_base = rank << 40, which a 32-bit truncation would drop)
    const uint64_t gid64 = mp.game_id ? (uint64_t)mp.game_id[g] : mp.game_id_base + (uint64_t)g;
    const uint32_t gid = (uint32_t)gid64;
    const uint32_t k1g = mp.k1 ^ (uint32_t)(gid64 >> 32);
    if (mp.selfplay) {
        double gm[GZ_MAXCH];
        double gsum = 0.0;
#pragma unroll
        for (int j = 0; j < GZ_MAXCH; ++j) {
            gm[j] = 0.0;
            if (isch[j]) {
                gm[j] = gz_gamma((double)mp.alpha, mp.k0, k1g, (uint32_t)(j * GZ_WARP + lane), ply, gid,
                                 0xD1A00000u);
                gsum += gm[j];
            }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) gsum += __shfl_xor_sync(0xffffffffu, gsum, off);
        int cnt = 0;
#pragma unroll
        for (int j = 0; j < GZ_MAXCH; ++j) cnt += isch[j];
        cnt = __reduce_add_sync(0xffffffffu, cnt);
#pragma unroll
        for (int j = 0; j < GZ_MAXCH; ++j)
            if (isch[j]) {
                double dir = gsum > 0.0 ? gm[j] / gsum : 1.0 / (double)cnt;
                x[j] = (1.0 - (double)mp.eps) * x[j] + (double)mp.eps * dir;
            }
    }
    // np.random.choice(acts, p): cdf = cumsum(p) / cdf[-1]; first index with cdf > u
    double tot = 0.0;
#pragma unroll
    for (int j = 0; j < GZ_MAXCH; ++j) tot += x[j];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, off);
    Philox4 rr = philox4x32(mp.k0, k1g, 0u, ply, gid, 0x5A3B0000u);
    const double target = gz_u01(rr.x) * tot;
    double run = 0.0;
    int chosen = -1, lastch = -1;
#pragma unroll
    for (int j = 0; j < GZ_MAXCH; ++j) {
        if (j < nch && chosen < 0) {
            double c = x[j];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                double t = __shfl_up_sync(0xffffffffu, c, off);
                if (lane >= off) c += t;
            }
            c += run;
            unsigned hit = __ballot_sync(0xffffffffu, isch[j] && c > target);
            unsigned any = __ballot_sync(0xffffffffu, isch[j]);
            if (any) lastch = j * GZ_WARP + (31 - __clz(any));
            if (hit) chosen = j * GZ_WARP + (__ffs(hit) - 1);
            run = __shfl_sync(0xffffffffu, c, 31);
        }
    }
    if (chosen < 0) chosen = lastch; // rounding at the top of the cdf
    if (lane == 0) moves[g] = chosen;
}

int gzk_root_select_move(gz_engine *e, float temp, int selfplay, int32_t *d_moves, float *d_probs,
                         bool best_visits)
{
    MoveParams mp;
    mp.temp = temp; mp.selfplay = selfplay; mp.best_visits = best_visits ? 1 : 0;
    mp.alpha = e->cfg.dirichlet_alpha; mp.eps = e->cfg.dirichlet_eps;
    mp.k0 = (uint32_t)e->cfg.seed; mp.k1 = (uint32_t)(e->cfg.seed >> 32);
    mp.game_id = e->sp_begun ? e->sp_game_id : nullptr;
    mp.game_id_base = e->cfg.game_id_base;
    ProfScope ps(e, PC_MOVE);
    root_select_kernel<<<gz_warp_blocks(e->G, 4), 128, 0, e->stream>>>(gz_tree_params(e), mp, d_moves,
                                                                        d_probs);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

// ---- self-play advance ------------------------------------------------------------------
struct SelfplayParams {
    GameBoard *boards;
    GameMeta *meta;
    int8_t *active;
    const int32_t *sel_moves;
    const float *sel_probs;
    int32_t *sp_moves;
    float *sp_pi;
    int64_t *sp_game_id;
    int recycle;          // 0 idle after a game, 1 next id = id + G, 2 next id = next start serial (selfplay_assign_ids_kernel)
    int32_t *fin_flag;    // [G] recycle == 2: slot finished a game in this ply and waits for its next id
    int64_t out_cap;
    float *out_states, *out_pi, *out_z;
    int64_t *out_gid;
    int32_t *out_ply;
    unsigned long long *out_count;
    int64_t gout_cap;
    int32_t *gout_moves, *gout_len;
    int8_t *gout_winner;
    int64_t *gout_gid;
    unsigned long long *gout_count;
    unsigned long long *dropped;
};

__global__ void __launch_bounds__(128) selfplay_advance_kernel(TreeParams p, SelfplayParams sp)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (g >= p.G) return;
    if (!sp.active[g]) return;
    const int mv = sp.sel_moves[g];
    if (mv < 0) return;
    const int S = p.S;
    // high-water mark of tree nodes in use (the tree is largest right after a search)
    if (lane == 0) atomicMax(&p.ctr->max_nodes, (unsigned long long)(p.cap - p.free_top[g]));
    GameMeta m = sp.meta[g];
    const int ply = m.n_stones;
    // record (state before the move is implied by the move list), pi, then play the move
    if (lane == 0) sp.sp_moves[(size_t)g * S + ply] = mv;
    for (int i = lane; i < S; i += GZ_WARP)
        sp.sp_pi[((size_t)g * S + ply) * S + i] = sp.sel_probs[(size_t)g * S + i];
    BB st[2];
    st[0] = load_bb(sp.boards[g].stones[0]);
    st[1] = load_bb(sp.boards[g].stones[1]);
    const int player = (m.start_player + ply) & 1;
    bb_set(st[player], bb_bit(mv, p.W));
    const int ns = ply + 1;
    int win;
    bool hw;
    const bool end = bb_game_end(st[0], st[1], ns, p.W, p.H, p.n_in_row, &win, &hw);
    __syncwarp();
    if (!end) {
        if (lane == 0) {
            store_bb(sp.boards[g].stones[player], st[player]);
            m.n_stones = ns;
            m.last_move = mv;
            sp.meta[g] = m;
        }
        gz_reroot(p, g, mv, lane); // update_with_move(move): keep the chosen subtree
        if (lane == 0) atomicAdd(&p.ctr->plies, 1ull);
        return;
    }
    // ---- game over: winners_z, emit samples, reset the slot -------------------------------
    const int T = ns;
    const int64_t gid = sp.sp_game_id[g];
    unsigned long long base = ~0ull;
    if (lane == 0) { // reserve T records only if they fit (count never runs past out_cap)
        unsigned long long old = *sp.out_count, assumed;
        do {
            if ((int64_t)(old + T) > sp.out_cap) { old = ~0ull; break; }
            assumed = old;
            old = atomicCAS(sp.out_count, assumed, assumed + (unsigned long long)T);
        } while (old != assumed);
        base = old;
    }
    base = __shfl_sync(0xffffffffu, base, 0);
    const bool room = base != ~0ull;
    if (room) {
        BB rb[2] = {bb_zero(), bb_zero()};
        int last = -1;
        for (int t = 0; t < T; ++t) {
            const int pl = (m.start_player + t) & 1;
            const size_t o = (size_t)base + t;
            gz_write_planes(rb[pl], rb[pl ^ 1], last, t, p.W, p.H, sp.out_states + o * 4 * S, lane);
            for (int i = lane; i < S; i += GZ_WARP)
                sp.out_pi[o * S + i] = sp.sp_pi[((size_t)g * S + t) * S + i];
            if (lane == 0) {
                sp.out_z[o] = (win == -1) ? 0.f : ((pl + 1 == win) ? 1.f : -1.f);
                sp.out_gid[o] = gid;
                sp.out_ply[o] = t;
            }
            last = sp.sp_moves[(size_t)g * S + t];
            bb_set(rb[pl], bb_bit(last, p.W));
        }
    } else if (lane == 0) {
        atomicAdd(sp.dropped, (unsigned long long)T);
    }
    unsigned long long gi = ~0ull;
    if (lane == 0) {
        unsigned long long old = *sp.gout_count, assumed;
        do {
            if ((int64_t)(old + 1) > sp.gout_cap) { old = ~0ull; break; }
            assumed = old;
            old = atomicCAS(sp.gout_count, assumed, assumed + 1ull);
        } while (old != assumed);
        gi = old;
    }
    gi = __shfl_sync(0xffffffffu, gi, 0);
    if (gi != ~0ull) {
        for (int t = lane; t < T; t += GZ_WARP)
            sp.gout_moves[(size_t)gi * S + t] = sp.sp_moves[(size_t)g * S + t];
        if (lane == 0) {
            sp.gout_len[gi] = T;
            sp.gout_winner[gi] = (int8_t)win;
            sp.gout_gid[gi] = gid;
        }
    }
    // reset_player() + init_board(): fresh tree, empty board
    for (int i = lane; i < p.cap; i += GZ_WARP) p.free_stack[(size_t)g * p.cap + i] = p.cap - 1 - i;
    if (lane == 0) {
        p.free_top[g] = p.cap;
        p.root_node[g] = -1;
        p.root_N[g] = 0u;
        p.root_Q[g] = 0.0;
        store_bb(sp.boards[g].stones[0], bb_zero());
        store_bb(sp.boards[g].stones[1], bb_zero());
        GameMeta z;
        z.n_stones = 0; z.last_move = -1; z.start_player = 0; z.flags = 0;
        sp.meta[g] = z;
        if (sp.recycle == 2) sp.fin_flag[g] = 1;
        else if (sp.recycle) sp.sp_game_id[g] = gid + (int64_t)p.G;
        else sp.active[g] = 0;
        atomicAdd(&p.ctr->plies, 1ull);
        atomicAdd(&p.ctr->games_finished, 1ull);
        atomicAdd(&p.ctr->samples_out, (unsigned long long)T);
        atomicAdd(&p.ctr->wins[win == -1 ? 0 : win], 1ull);
    }
}

// recycle mode 2 ("serial ids", the continuous collect mode of train.py): the slots that finished a game in this
// ply get the next game serials in SLOT ORDER (block-wide scan over the flags: deterministic, unlike an atomic
// counter), so that game ids order the games by when they STARTED.  A consumer that releases finished games in id
// order then hands out every game exactly once, in start order, however long it took - short games are not
// over-represented the way they are when the first n games to FINISH are taken.
__global__ void __launch_bounds__(1024) selfplay_assign_ids_kernel(int32_t *fin_flag, int64_t *game_id,
                                                                   unsigned long long *serial, int G, uint64_t base)
{
    __shared__ unsigned warp_sum[32];
    __shared__ unsigned long long next;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) next = *serial;
    __syncthreads();
    for (int g0 = 0; g0 < G; g0 += 1024) {
        const int g = g0 + (int)threadIdx.x;
        const bool f = g < G && fin_flag[g] != 0;
        const unsigned b = __ballot_sync(0xffffffffu, f);
        if (lane == 0) warp_sum[wid] = __popc(b);
        __syncthreads();
        unsigned before = 0, total = 0;
        for (int w = 0; w < 32; ++w) { before += w < wid ? warp_sum[w] : 0u; total += warp_sum[w]; }
        if (f) {
            game_id[g] = (int64_t)(base + next + before + __popc(b & ((1u << lane) - 1u)));
            fin_flag[g] = 0;
        }
        __syncthreads();
        if (threadIdx.x == 0) next += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *serial = next;
}

int gzk_selfplay_advance(gz_engine *e)
{
    SelfplayParams sp;
    sp.boards = e->boards; sp.meta = e->meta; sp.active = e->active; sp.sel_moves = e->sel_moves;
    sp.sel_probs = e->sel_probs; sp.sp_moves = e->sp_moves; sp.sp_pi = e->sp_pi;
    sp.sp_game_id = e->sp_game_id; sp.recycle = e->sp_recycle; sp.fin_flag = e->sp_fin_flag; sp.out_cap = e->out_cap;
    sp.out_states = e->out_states; sp.out_pi = e->out_pi; sp.out_z = e->out_z; sp.out_gid = e->out_gid;
    sp.out_ply = e->out_ply; sp.out_count = e->out_count; sp.gout_cap = e->gout_cap;
    sp.gout_moves = e->gout_moves; sp.gout_len = e->gout_len; sp.gout_winner = e->gout_winner;
    sp.gout_gid = e->gout_gid; sp.gout_count = e->gout_count;
    sp.dropped = reinterpret_cast<unsigned long long *>(&e->counters->dropped);
    ProfScope ps(e, PC_ADVANCE);
    selfplay_advance_kernel<<<gz_warp_blocks(e->G, 4), 128, 0, e->stream>>>(gz_tree_params(e), sp);
    e->launches++;
    if (e->sp_recycle == 2) {
        selfplay_assign_ids_kernel<<<1, 1024, 0, e->stream>>>(e->sp_fin_flag, e->sp_game_id, e->sp_serial, e->G,
                                                              e->cfg.game_id_base);
        e->launches++;
    }
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}
