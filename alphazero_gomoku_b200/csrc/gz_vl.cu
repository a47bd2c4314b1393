// gz_vl.cu — several leaves in flight per game (virtual loss), BASELINE north_star subsystem (1); opt-in, K = 1 is
// the default path of gz_tree.cu and is untouched.  The reference's playouts are strictly sequential
// (mcts_alphaZero.py:161-163): with ONE game on the device that is one 4-board MMA group per wave and a launch-latency
// bound 33 us per playout.  Here a wave carries up to K playouts of every game:
//   select   one warp per game makes up to K descents one after another.  Descent j scores a child as if every earlier
//            descent of this wave through it had come back as a loss for the player choosing there: n_eff = n + c,
//            q_eff = c running-mean updates of Q with -1 (the reference's update rule, :64-65), N_parent likewise.
//            The virtual losses live in shared memory for the duration of the kernel (the earlier paths' edges, level
//            by level: a path passes through a node iff its edge at that depth starts there) - the tree in HBM never
//            holds one, so there is nothing to revert and no atomic to order: the result is deterministic.
//            A descent that reaches a non-terminal leaf an earlier descent of the wave holds ends the wave early.
//   net      all leaves of all games in one batch (up to G * K boards).
//   backup   the same warp expands and backs up its leaves IN ORDER with the reference's arithmetic (:38-73).
// The exact specification is restated on a pointer tree in oracle/orc_mcts.c:orc_mcts_search_vl; the kernels match it
// bit for bit under a shared evaluator (tests/test_gpu_vl.py).  What K > 1 costs is fidelity to the reference's
// SEQUENTIAL search: paths 2..K of a wave are chosen without the true results of the earlier ones.  Measured against
// the reference's visit counts (400 playouts, shipped 8x8 weights): total variation 0.008 mean at K = 2, 0.031 at
// K = 4, 0.057 at K = 8 (same best move in 15 of 16 positions) - outside the 0.02 tier, which is why it is opt-in.
// Included by gz_api.cu after gz_tree.cu and gz_move.cu (single translation unit).

#define GZ_VL_MAXK 16
#define GZ_VL_DEPTH 32 // levels of a path that carry virtual losses (oracle: ORC_VL_DEPTH)

struct VlParams {
    uint32_t *path;        // [G*K][S]
    int32_t *path_len;     // [G*K]
    int8_t *leaf_state;    // [G*K] 0 eval, 1 terminal
    double *leaf_value;    // [G*K]
    int32_t *leaf_slot;    // [G*K]
    GameBoard *leaf_board; // [G*K]
    int32_t *keff;         // [G] leaves of the wave in flight
    int32_t *remaining;    // [G] playouts still to start
    float *planes;         // [G*K][4][S] compacted rows, or null: the evaluator stages a leaf from leaf_board / leaf_meta
    GameMeta *leaf_meta;   // [G*K]
    int32_t *slot_leaf;    // [G*K] leaf record (g * K + k) of a batch slot
    int32_t *count, *count_zero; // leaf counter of this wave / the one the net consumed last (reset here)
    int K;
};

template <int NSL>
__device__ __forceinline__ void gz_vl_game(const TreeParams &p, const VlParams &vp, int g, int lane,
                                           const float *__restrict__ priors, const float *__restrict__ values,
                                           int do_expand, int do_select, uint32_t (*sh_edge)[GZ_VL_DEPTH],
                                           unsigned &levels_out, unsigned &children_out, unsigned &playouts_out)
{
    constexpr int SP = 32 * NSL;
    const int K = vp.K, S = p.S;
    const size_t nbase = (size_t)g * p.cap;
    levels_out = children_out = playouts_out = 0;
    int root = p.root_node[g];
    uint32_t rootN = p.root_N[g];
    double rootQ = p.root_Q[g];
    int top = p.free_top[g];
    int remaining = vp.remaining[g];
    int cellbit[NSL]; // bitboard bit of this lane's child slots (the board side is a run-time value here: divide once)
#pragma unroll
    for (int q = 0; q < NSL; ++q) cellbit[q] = (lane + 32 * q < S) ? bb_bit(lane + 32 * q, p.W) : 0;
    // ---- expand + backup of the wave in flight, leaf by leaf in descent order (mcts_alphaZero.py:131-148)
    if (do_expand) {
        const int keff = vp.keff[g];
        // trip 1: the per-leaf scalars of all leaves (lane j holds leaf j)
        int my_st = 2, my_d = 0, my_slot = 0;
        double my_val = 0.0;
        if (lane < keff) {
            const size_t v = (size_t)g * K + lane;
            my_st = vp.leaf_state[v]; my_d = vp.path_len[v]; my_slot = vp.leaf_slot[v];
        }
        const bool deep = __any_sync(0xffffffffu, my_d > 32);
        if (!deep) {
            // trip 2, everything at once: values, every leaf's path (lane = level), the free nodes in pop order, the
            // leaf positions; trip 3: the statistics of every path edge.  After that the leaves are processed in order
            // FROM REGISTERS: an edge shared by two paths of the wave sits at the same level, i.e. in the same lane,
            // so the later path finds the earlier one's update by comparing edge ids - no reload, no wait.
            if (lane < keff) my_val = my_st == 0 ? (double)values[my_slot] : vp.leaf_value[(size_t)g * K + lane];
            uint32_t e_all[GZ_VL_MAXK], n_all[GZ_VL_MAXK];
            double q_all[GZ_VL_MAXK];
#pragma unroll
            for (int j = 0; j < GZ_VL_MAXK; ++j) {
                const int d = __shfl_sync(0xffffffffu, my_d, j);
                e_all[j] = (j < keff && lane < d) ? vp.path[((size_t)g * K + j) * S + lane] : 0xffffffffu;
            }
            int fs_mine = -1;
            if (lane < keff && lane < top) fs_mine = p.free_stack[nbase + top - 1 - lane];
            BB occ_mine = bb_zero();
            if (lane < keff) {
                const size_t v = (size_t)g * K + lane;
                occ_mine = bb_or(load_bb(vp.leaf_board[v].stones[0]), load_bb(vp.leaf_board[v].stones[1]));
            }
#pragma unroll
            for (int j = 0; j < GZ_VL_MAXK; ++j) {
                n_all[j] = 0u; q_all[j] = 0.0;
                if (e_all[j] != 0xffffffffu) {
                    NodeView nv = gz_node_view(p.pool, nbase + e_all[j] / SP, SP);
                    n_all[j] = nv.N[e_all[j] % SP];
                    q_all[j] = nv.Q[e_all[j] % SP];
                }
            }
            int n_exp = 0;
#pragma unroll
            for (int j = 0; j < GZ_VL_MAXK; ++j) {
                if (j < keff) {
                    const int st = __shfl_sync(0xffffffffu, my_st, j), d = __shfl_sync(0xffffffffu, my_d, j);
                    const int slot = __shfl_sync(0xffffffffu, my_slot, j);
                    const double val = __shfl_sync(0xffffffffu, my_val, j);
                    if (st == 0) {
                        const int newn = __shfl_sync(0xffffffffu, fs_mine, n_exp & 31);
                        if (n_exp < top && newn >= 0) {
                            BB occ;
#pragma unroll
                            for (int wq = 0; wq < 4; ++wq) occ.w[wq] = __shfl_sync(0xffffffffu, occ_mine.w[wq], j);
                            NodeView nv = gz_node_view(p.pool, nbase + newn, SP);
                            unsigned cnt = 0;
#pragma unroll
                            for (int q = 0; q < NSL; ++q) {
                                const int i = lane + 32 * q;
                                const bool leg = (i < S) && !bb_test(occ, cellbit[q]);
                                nv.P[i] = leg ? priors[(size_t)slot * S + i] : -1.f;
                                nv.N[i] = 0u; nv.Q[i] = 0.0; nv.child[i] = -1;
                                cnt += leg;
                            }
                            if (d == 0) root = newn;
                            else {
                                const uint32_t e = __shfl_sync(0xffffffffu, e_all[j], (d - 1) & 31);
                                if (lane == 0) gz_node_view(p.pool, nbase + e / SP, SP).child[e % SP] = newn;
                            }
                            ++n_exp;
                            cnt = __reduce_add_sync(0xffffffffu, cnt);
                            if (lane == 0) {
                                atomicAdd(&p.ctr->expanded_children, (unsigned long long)cnt);
                                atomicAdd(&p.ctr->leaf_evals, 1ull);
                            }
                        } else if (lane == 0) {
                            atomicAdd(&p.ctr->pool_overflows, 1ull);
                        }
                    }
                    if (lane < d) { // backup of this lane's level (:57-73): latest statistics of the edge, updated, stored
                        uint32_t n = n_all[j];
                        double q = q_all[j];
#pragma unroll
                        for (int i = 0; i < j; ++i)
                            if (e_all[i] == e_all[j]) { n = n_all[i]; q = q_all[i]; }
                        const double x = ((d - (lane + 1)) & 1) ? val : -val;
                        n += 1u;
                        q = q + 1.0 * (x - q) / (double)n;
                        n_all[j] = n; q_all[j] = q;
                        NodeView nv = gz_node_view(p.pool, nbase + e_all[j] / SP, SP);
                        nv.N[e_all[j] % SP] = n;
                        nv.Q[e_all[j] % SP] = q;
                    }
                    {
                        const double x = (d & 1) ? val : -val;
                        rootN += 1u;
                        rootQ = rootQ + 1.0 * (x - rootQ) / (double)rootN;
                    }
                }
            }
            top -= n_exp;
            __syncwarp(); // the descents below read what other lanes have just written
        } else {
        // paths deeper than a warp (rare): leaf by leaf with reloads
        if (lane < keff) my_val = my_st == 0 ? (double)values[my_slot] : vp.leaf_value[(size_t)g * K + lane];
        for (int j = 0; j < keff; ++j) {
            const size_t v = (size_t)g * K + j;
            const int st = __shfl_sync(0xffffffffu, my_st, j), d = __shfl_sync(0xffffffffu, my_d, j);
            const int slot = __shfl_sync(0xffffffffu, my_slot, j);
            const double val = __shfl_sync(0xffffffffu, my_val, j);
            const uint32_t e_mine = (lane < d) ? vp.path[v * S + lane] : 0u;
            if (st == 0) {
                int newn = -1;
                if (top > 0) newn = p.free_stack[nbase + top - 1]; // (same address for all lanes: one transaction)
                if (newn >= 0) {
                    const BB occ = bb_or(load_bb(vp.leaf_board[v].stones[0]), load_bb(vp.leaf_board[v].stones[1]));
                    NodeView nv = gz_node_view(p.pool, nbase + newn, SP);
                    unsigned cnt = 0;
#pragma unroll
                    for (int q = 0; q < NSL; ++q) {
                        const int i = lane + 32 * q;
                        const bool leg = (i < S) && !bb_test(occ, cellbit[q]);
                        nv.P[i] = leg ? priors[(size_t)slot * S + i] : -1.f;
                        nv.N[i] = 0u; nv.Q[i] = 0.0; nv.child[i] = -1;
                        cnt += leg;
                    }
                    if (d == 0) root = newn;
                    else {
                        const uint32_t e_last = __shfl_sync(0xffffffffu, e_mine, (d - 1) & 31);
                        const uint32_t e = d <= 32 ? e_last : vp.path[v * S + d - 1];
                        if (lane == 0) gz_node_view(p.pool, nbase + e / SP, SP).child[e % SP] = newn;
                    }
                    --top;
                    cnt = __reduce_add_sync(0xffffffffu, cnt);
                    if (lane == 0) {
                        atomicAdd(&p.ctr->expanded_children, (unsigned long long)cnt);
                        atomicAdd(&p.ctr->leaf_evals, 1ull);
                    }
                } else if (lane == 0) {
                    atomicAdd(&p.ctr->pool_overflows, 1ull);
                }
            }
            for (int k = lane; k < d; k += 32) {
                const uint32_t e = k < 32 ? e_mine : vp.path[v * S + k];
                NodeView nv = gz_node_view(p.pool, nbase + e / SP, SP);
                const int cell = (int)(e % SP);
                const double x = ((d - (k + 1)) & 1) ? val : -val;
                const uint32_t n = nv.N[cell] + 1u;
                nv.N[cell] = n;
                const double q = nv.Q[cell];
                nv.Q[cell] = q + 1.0 * (x - q) / (double)n;
            }
            {
                const double x = (d & 1) ? val : -val;
                rootN += 1u;
                rootQ = rootQ + 1.0 * (x - rootQ) / (double)rootN;
            }
            __syncwarp(); // the next leaf's backup (and the descents below) read what other lanes have just written
        }
        }
        if (lane == 0) {
            p.root_node[g] = root; p.root_N[g] = rootN; p.root_Q[g] = rootQ; p.free_top[g] = top;
            if (keff) atomicAdd(&p.ctr->playouts, (unsigned long long)keff);
        }
    }
    if (g == 0 && lane == 0 && vp.count_zero) *vp.count_zero = 0;
    if (!do_select) {
        if (lane == 0) vp.keff[g] = 0;
        return;
    }
    // ---- up to K descents
    const int kmax = (p.active[g] && remaining > 0) ? min(K, remaining) : 0;
    if (kmax == 0) {
        if (lane == 0) vp.keff[g] = 0;
        return;
    }
    const GameMeta m = p.meta[g];
    // one 64-bit word of the position per lane on the way down (lane l < 8: word l & 3 of colour l >> 2), as in
    // gz_expand_select_game
    const uint64_t bw0 = lane < 8 ? p.boards[g].stones[lane >> 2][lane & 3] : 0ull;
    float rP0[NSL];
    uint32_t rN0[NSL];
    double rQ0[NSL];
    int rC0[NSL];
    if (root >= 0) {
        NodeView rv = gz_node_view(p.pool, nbase + root, SP);
#pragma unroll
        for (int q = 0; q < NSL; ++q) {
            const int i = lane + 32 * q;
            if (i < S) { rP0[q] = rv.P[i]; rN0[q] = rv.N[i]; rQ0[q] = rv.Q[i]; rC0[q] = rv.child[i]; }
            else { rP0[q] = -1.f; rN0[q] = 0u; rQ0[q] = 0.0; rC0[q] = -1; }
        }
    }
    int my_plen = 0;                  // lane i: length and last edge of kept descent i of this wave
    uint32_t my_plast = 0xffffffffu;
    unsigned term_mask = 0u;          // lane-uniform: descents that ended in a terminal leaf
    int keff = 0;
    unsigned nchild = 0, nlevels = 0;
    for (int j = 0; j < kmax; ++j) {
        const size_t v = (size_t)g * K + j;
        uint64_t bw = bw0;
        int player = (m.start_player + m.n_stones) & 1;
        int ns = m.n_stones, last = m.last_move;
        int node = root, depth = 0;
        uint32_t Np = rootN + (uint32_t)j; // every earlier descent of the wave passed through the root
        uint32_t edge = 0xffffffffu;
        float rP[NSL];
        uint32_t rN[NSL];
        double rQ[NSL];
        int rC[NSL];
#pragma unroll
        for (int q = 0; q < NSL; ++q) { rP[q] = rP0[q]; rN[q] = rN0[q]; rQ[q] = rQ0[q]; rC[q] = rC0[q]; }
        while (node >= 0) {
            int c[NSL];
#pragma unroll
            for (int q = 0; q < NSL; ++q) c[q] = 0;
            if (j > 0 && depth < GZ_VL_DEPTH) {
                // lane i < j looks at kept descent i: does it pass through this node at this depth, and by which child?
                uint32_t e_i = 0xffffffffu;
                if (lane < j && depth < my_plen) e_i = sh_edge[lane][depth];
                unsigned mm = __ballot_sync(0xffffffffu, e_i != 0xffffffffu && (int)(e_i / SP) == node);
                while (mm) { // one round per descent that does (all of them at the root, few further down)
                    const int i = __ffs(mm) - 1;
                    mm &= mm - 1u;
                    const int cell = (int)(__shfl_sync(0xffffffffu, e_i, i) % SP);
#pragma unroll
                    for (int q = 0; q < NSL; ++q) c[q] += (cell == lane + 32 * q);
                }
            }
            const double sq = gz_sqrt_visits(Np);
            double best = -1.0e300;
            int besti = 0x7fffffff, bestc = -1;
            uint32_t bestn = 0u;
#pragma unroll
            for (int q = 0; q < NSL; ++q) {
                if (rP[q] >= 0.f) {
                    double qe = rQ[q];
                    for (int t = 0; t < c[q]; ++t) qe += 1.0 * (-1.0 - qe) / (double)(rN[q] + (uint32_t)t + 1u);
                    const uint32_t ne = rN[q] + (uint32_t)c[q];
                    const double cp = (double)(p.c_puct * rP[q]);
                    const double sc = qe + cp * sq / (double)(1u + ne);
                    if (sc > best) { best = sc; besti = lane + 32 * q; bestn = ne; bestc = rC[q]; }
                    ++nchild;
                }
            }
            const int mv = warp_argmax_redux(best, besti);
            const int owner = mv & 31;
            Np = __shfl_sync(0xffffffffu, bestn, owner);
            const int next = __shfl_sync(0xffffffffu, bestc, owner);
            edge = (uint32_t)node * SP + (uint32_t)mv;
            if (lane == 0) {
                vp.path[v * S + depth] = edge;
                if (depth < GZ_VL_DEPTH) sh_edge[j][depth] = edge;
            }
            ++depth;
            if (next >= 0) {
                NodeView nv = gz_node_view(p.pool, nbase + next, SP);
#pragma unroll
                for (int q = 0; q < NSL; ++q) {
                    const int i = lane + 32 * q;
                    if (i < S) { rP[q] = nv.P[i]; rN[q] = nv.N[i]; rQ[q] = nv.Q[i]; rC[q] = nv.child[i]; }
                    else rP[q] = -1.f;
                }
            }
            {
                int cb = cellbit[0]; // the bit of cell mv sits in its owner lane
#pragma unroll
                for (int q = 1; q < NSL; ++q) cb = (mv >> 5) == q ? cellbit[q] : cb;
                const int bit = __shfl_sync(0xffffffffu, cb, owner);
                if (lane == 4 * player + (bit >> 6)) bw |= 1ull << (bit & 63);
            }
            player ^= 1;
            last = mv;
            ++ns;
            node = next;
        }
        __syncwarp(); // sh_edge[j] written by lane 0 is read by every lane in the next descent
        bool w0 = false, w1 = false;
        if (ns >= p.n_in_row + 2) {
            // lanes 0-3 test colour 0, lanes 4-7 colour 1, one direction each: gather that colour's four words
            const int d4 = lane & 3;
            const int dir = (d4 == 0) ? 1 : (d4 == 1) ? 16 : (d4 == 2) ? 17 : 15;
            BB cc;
#pragma unroll
            for (int k = 0; k < 4; ++k) cc.w[k] = __shfl_sync(0xffffffffu, bw, (lane & 4) + k);
            bool line = false;
            if (lane < 8) line = bb_has_line_dir(cc, p.n_in_row, dir);
            const unsigned b = __ballot_sync(0xffffffffu, line);
            w0 = (b & 0x0Fu) != 0u;
            w1 = (b & 0xF0u) != 0u;
        }
        const int win = w0 ? 1 : (w1 ? 2 : -1);
        const bool end = (w0 || w1) || ns >= p.W * p.H;
        if (!end) { // a non-terminal leaf an earlier descent holds is about to be expanded: the wave ends here
            const bool mine = lane < j && !((term_mask >> lane) & 1u) && my_plen == depth && my_plast == edge;
            if (__any_sync(0xffffffffu, mine)) break;
        }
        if (lane == j) { my_plen = depth; my_plast = edge; }
        term_mask |= end ? (1u << j) : 0u;
        keff = j + 1;
        nlevels += (unsigned)depth;
        int slot = 0;
        if (lane < 8) vp.leaf_board[v].stones[lane >> 2][lane & 3] = bw;
        if (lane == 0) {
            if (!end) slot = atomicAdd(vp.count, 1);
            vp.path_len[v] = depth;
            vp.leaf_state[v] = end ? 1 : 0;
            vp.leaf_slot[v] = slot;
            if (end) vp.leaf_value[v] = (win == -1) ? 0.0 : ((win == player + 1) ? 1.0 : -1.0);
            else if (!vp.planes) {
                GameMeta lm;
                lm.n_stones = ns; lm.last_move = last; lm.start_player = m.start_player; lm.flags = 0;
                vp.leaf_meta[v] = lm;
                vp.slot_leaf[slot] = (int32_t)v;
            }
        }
        if (!end && vp.planes) {
            BB s0, s1;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                s0.w[k] = __shfl_sync(0xffffffffu, bw, k);
                s1.w[k] = __shfl_sync(0xffffffffu, bw, 4 + k);
            }
            slot = __shfl_sync(0xffffffffu, slot, 0);
            gz_write_planes(bb_pick(s0, s1, player == 0), bb_pick(s1, s0, player == 0), last, ns, p.W, p.H,
                            vp.planes + (size_t)slot * 4 * S, lane);
        }
    }
    if (lane == 0) {
        vp.keff[g] = keff;
        vp.remaining[g] = remaining - keff;
    }
    levels_out = nlevels;
    children_out = __reduce_add_sync(0xffffffffu, nchild);
    playouts_out = 0u; // counted at backup
}

template <int NSL>
__global__ void __launch_bounds__(128, NSL <= 2 ? 4 : 2)
vl_wave_kernel(TreeParams p, VlParams vp, const float *__restrict__ priors, const float *__restrict__ values,
               int do_expand, int do_select)
{
    __shared__ uint32_t sh_edge[4][GZ_VL_MAXK][GZ_VL_DEPTH];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int g = blockIdx.x * 4 + wid;
    // link of the dependent-launch chain heads -> this kernel -> trunk (gz_common.cuh): scheduled while the heads
    // kernel drains, and the trunk after it may set itself up while this one runs
    gz_pdl_wait();
    gz_pdl_trigger();
    unsigned lv = 0, ch = 0, pl = 0;
    if (g < p.G) gz_vl_game<NSL>(p, vp, g, lane, priors, values, do_expand, do_select, sh_edge[wid], lv, ch, pl);
    if (lane == 0 && (lv | ch)) {
        atomicAdd(&p.ctr->select_levels, (unsigned long long)lv);
        atomicAdd(&p.ctr->select_children, (unsigned long long)ch);
    }
}

__global__ void vl_init_kernel(const int8_t *active, int32_t *remaining, int32_t *keff, int G, int n)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g < G) { remaining[g] = active[g] ? n : 0; keff[g] = 0; }
}

__global__ void vl_max_kernel(const int32_t *remaining, int G, int32_t *out)
{
    int mx = 0;
    for (int g = threadIdx.x; g < G; g += blockDim.x) mx = max(mx, remaining[g]);
    mx = __reduce_max_sync(0xffffffffu, mx);
    if ((threadIdx.x & 31) == 0) atomicMax(out, mx);
}

static int gz_vl_ensure(gz_engine *e, int K)
{
    if (e->vl_K >= K) return GZ_OK;
    void *old[] = {e->vl_path, e->vl_path_len, e->vl_leaf_state, e->vl_leaf_value, e->vl_leaf_slot, e->vl_leaf_board,
                   e->vl_keff, e->vl_remaining, e->vl_planes, e->vl_probs, e->vl_values, e->vl_count, e->vl_leaf_meta,
                   e->vl_slot_leaf};
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    for (void *q : old) if (q) cudaFree(q);
    const size_t GK = (size_t)e->G * K, S = e->S;
    e->vl_K = 0;
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_path, GK * S * sizeof(uint32_t)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_path_len, GK * sizeof(int32_t)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_leaf_state, GK));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_leaf_value, GK * sizeof(double)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_leaf_slot, GK * sizeof(int32_t)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_leaf_board, GK * sizeof(GameBoard)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_leaf_meta, GK * sizeof(GameMeta)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_slot_leaf, GK * sizeof(int32_t)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_keff, (size_t)e->G * sizeof(int32_t)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_remaining, (size_t)e->G * sizeof(int32_t)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_planes, GK * 4 * S * sizeof(float)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_probs, GK * S * sizeof(float)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_values, GK * sizeof(float)));
    GZ_CUDA(e, cudaMalloc((void **)&e->vl_count, 4 * sizeof(int32_t)));
    e->vl_K = K;
    return GZ_OK;
}

static int gz_vl_launch(gz_engine *e, const TreeParams &p, const VlParams &vp, int do_expand, int do_select)
{
    const unsigned grid = gz_warp_blocks(e->G, 4);
    ProfScope ps(e, PC_SELECT);
#define GZ_VLK(NSL) GZ_CUDA(e, gz_launch_chain(vl_wave_kernel<NSL>, dim3(grid), dim3(128), 0, e->stream, p, vp, (const float *)e->vl_probs, (const float *)e->vl_values, do_expand, do_select))
    switch (e->Sp / 32) {
    case 1: GZ_VLK(1); break;
    case 2: GZ_VLK(2); break;
    case 3: GZ_VLK(3); break;
    case 4: GZ_VLK(4); break;
    case 8: GZ_VLK(8); break;
    default: e->err = "gz_search_vl: unsupported board size"; return GZ_E_INVALID;
    }
#undef GZ_VLK
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

// n_playout playouts in every active game with up to K leaves of a game in flight per wave
extern "C" int gz_search_vl(gz_engine *e, int32_t n_playout, int32_t K, int32_t *waves_out)
{
    if (!e || n_playout < 0 || K < 1 || K > GZ_VL_MAXK) return GZ_E_INVALID;
    if (!e->wts.loaded && !e->dbg_hash_on) { e->err = "gz_set_weights has not been called"; return GZ_E_NOWEIGHTS; }
    if (e->W != e->H) { e->err = "the net needs a square board"; return GZ_E_INVALID; }
    if (e->Sp / 32 > 8 || (e->Sp / 32 > 4 && e->Sp / 32 != 8)) { e->err = "gz_search_vl: unsupported board size"; return GZ_E_INVALID; }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    if (waves_out) *waves_out = 0;
    if (n_playout == 0) return GZ_OK;
    GZ_TRY(gz_vl_ensure(e, K));
    const TreeParams p = gz_tree_params(e);
    VlParams vp;
    vp.path = e->vl_path; vp.path_len = e->vl_path_len; vp.leaf_state = e->vl_leaf_state; vp.leaf_value = e->vl_leaf_value;
    vp.leaf_slot = e->vl_leaf_slot; vp.leaf_board = e->vl_leaf_board; vp.keff = e->vl_keff; vp.remaining = e->vl_remaining;
    vp.leaf_meta = e->vl_leaf_meta; vp.slot_leaf = e->vl_slot_leaf; vp.K = K;
    // leaves for the fused tcgen05 trunk are staged from their bitboards, as in gz_waves (GZ_LEAF_FLOAT=1: fp32 planes)
    static const bool leaf_float = getenv("GZ_LEAF_FLOAT") != nullptr;
    struct PackedScope {
        gz_engine *e;
        ~PackedScope() { e->leaf_packed = false; e->pk_slot = nullptr; }
    } scope{e};
    e->leaf_packed = !leaf_float && !e->dbg_hash_on && e->wts.loaded && e->cfg.net_impl != 1 && gz_tc_supported(e);
    e->pk_slot = e->vl_slot_leaf; e->pk_board = e->vl_leaf_board; e->pk_meta = e->vl_leaf_meta;
    vp.planes = e->leaf_packed ? nullptr : e->vl_planes;
    GZ_CUDA(e, cudaMemsetAsync(e->vl_count, 0, 4 * sizeof(int32_t), e->stream));
    vl_init_kernel<<<(e->G + 255) / 256, 256, 0, e->stream>>>(e->active, e->vl_remaining, e->vl_keff, e->G, n_playout);
    e->launches++;
    const int64_t GK = (int64_t)e->G * K;
    int waves = 0, planned = (n_playout + K - 1) / K + 1; // wave 0 of a fresh root carries one leaf
    bool first = true;
    for (;;) {
        for (int w = 0; w < planned; ++w, ++waves) {
            vp.count = e->vl_count + (waves & 1);
            vp.count_zero = e->vl_count + ((waves + 1) & 1);
            GZ_TRY(gz_vl_launch(e, p, vp, first ? 0 : 1, 1));
            first = false;
            GZ_TRY(gzk_net_forward(e, e->vl_planes, vp.count, GK, e->vl_probs, e->vl_values));
        }
        // does any game still have playouts to start?  (the leaves in flight are backed up by the next kernel either way)
        int32_t left = 0;
        GZ_CUDA(e, cudaMemsetAsync(e->vl_count + 2, 0, sizeof(int32_t), e->stream));
        vl_max_kernel<<<1, 1024, 0, e->stream>>>(e->vl_remaining, e->G, e->vl_count + 2);
        e->launches++;
        GZ_CUDA(e, cudaMemcpyAsync(&left, e->vl_count + 2, sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
        GZ_CUDA(e, cudaStreamSynchronize(e->stream));
        if (left == 0) break;
        planned = (left + K - 1) / K + 1;
    }
    // the last wave's leaves (if any are in flight) are backed up here
    vp.count = e->vl_count + (waves & 1);
    vp.count_zero = e->vl_count + ((waves + 1) & 1);
    GZ_TRY(gz_vl_launch(e, p, vp, 1, 0));
    if (waves_out) *waves_out = waves;
    return GZ_OK;
}
