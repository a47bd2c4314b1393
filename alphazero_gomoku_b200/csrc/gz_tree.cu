// gz_tree.cu — GPU-resident MCTS trees (reference mcts_alphaZero.py:24-192, mcts_pure.py:49-243).
//   K1 select: one warp per game walks root -> leaf; at each level the lanes score the
//      children (Q + u, u = c_puct * P * sqrt(N_parent) / (1 + N), all in the reference's
//      mixed float32/float64 arithmetic) and a shuffle argmax with
//      lowest-index tie-break reproduces Python's first-max (mcts_alphaZero.py:55).
//   K2 expand + backup: writes the priors of the legal moves into a fresh node block and
//      walks the recorded path updating N and the running-mean Q with alternating sign
//      (update_recursive, mcts_alphaZero.py:57-73).  One playout in flight per game, so the
//      tree sees exactly the reference's sequential statistics.
//   K2+K1 fused across the wave boundary (expand_select*_kernel): the expand + backup of playout k
//      and the descent of playout k+1 by the same warp, organised around dependent memory round
//      trips; bit-identical to running K2 then K1.
//   re-root with subtree reuse (update_with_move, :183-192) frees the discarded nodes onto a
//      per-game free stack that doubles as the BFS queue.
#include "gz_engine.h"

struct TreeParams {
    const GameBoard *boards;
    const GameMeta *meta;
    const int8_t *active;
    uint8_t *pool;
    int32_t *free_stack, *free_top, *root_node;
    uint32_t *root_N;
    double *root_Q;
    uint32_t *path;
    int32_t *path_len;
    int8_t *leaf_state;
    double *leaf_value;
    int32_t *leaf_slot, *leaf_count;
    int32_t *leaf_zero; // fused wave kernels: the counter of the wave whose leaves were just evaluated, reset for re-use (or null)
    float *leaf_planes;   // null: the evaluator reads leaf_board / leaf_meta through slot_game instead
    int32_t *slot_game;   // [G] game of a leaf-batch slot
    GameBoard *leaf_board;
    GameMeta *leaf_meta;
    DevCounters *ctr;
    int G, W, H, S, Sp, n_in_row, cap;
    float c_puct;
    int g0; // first game of the launch (kernels that work on a slice of the games: [g0, G))
};

// sqrt(N_parent) of the PUCT score (mcts_alphaZero.py:81) for the visit counts a search actually meets: the correctly
// rounded double, filled by the host at gz_create (IEEE sqrt on both sides, so the values ARE the device's); a
// warp-uniform constant-bank read instead of the 13-instruction double-precision sequence, once per level.
#define GZ_SQRT_TAB 4096
__constant__ double gz_sqrt_tab[GZ_SQRT_TAB];
__device__ __forceinline__ double gz_sqrt_visits(uint32_t n) { return n < GZ_SQRT_TAB ? gz_sqrt_tab[n] : sqrt((double)n); }
static int gz_sqrt_tab_init(int device)
{
    static bool done[64] = {};
    if (device >= 0 && device < 64 && done[device]) return 0;
    static double h[GZ_SQRT_TAB];
    for (int i = 0; i < GZ_SQRT_TAB; ++i) h[i] = std::sqrt((double)i);
    if (cudaMemcpyToSymbol(gz_sqrt_tab, h, sizeof(h)) != cudaSuccess) return -1;
    if (device >= 0 && device < 64) done[device] = true;
    return 0;
}

__device__ __forceinline__ void warp_argmax(double &s, int &i)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        double os = __shfl_xor_sync(0xffffffffu, s, off);
        int oi = __shfl_xor_sync(0xffffffffu, i, off);
        if (os > s || (os == s && oi < i)) { s = os; i = oi; }
    }
}

// Same result (largest score, lowest index on ties) with three warp-wide integer reductions: the
// double is mapped to an order-preserving 64-bit key and reduced high word first.
__device__ __forceinline__ int warp_argmax_redux(double s, int i)
{
    const unsigned long long bits = (unsigned long long)__double_as_longlong(s + 0.0); // -0.0 -> +0.0
    const unsigned long long key = (bits >> 63) ? ~bits : (bits | 0x8000000000000000ull);
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
    const unsigned ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
    return (int)__reduce_min_sync(0xffffffffu, (hi == mh && lo == ml) ? (unsigned)i : 0x7fffffffu);
}

// n-in-a-row test of one colour in one direction (shift 1 / 16 / 17 / 15), see bb_has_line

// per-block statistics without a trailing barrier: warps add into shared counters and the last
// warp to finish flushes them to the device counters
struct SelStats {
    unsigned int levels, children, playouts, done;
};
__device__ __forceinline__ void gz_stats_finish(SelStats *ss, DevCounters *ctr, int lane, unsigned levels,
                                                unsigned children, unsigned playouts)
{
    if (lane == 0) {
        if (playouts) {
            atomicAdd(&ss->levels, levels);
            atomicAdd(&ss->children, children);
            atomicAdd(&ss->playouts, playouts);
        }
        __threadfence_block();
        const unsigned nw = blockDim.x >> 5;
        if (atomicAdd(&ss->done, 1u) == nw - 1u) {
            __threadfence_block();
            const unsigned pl = atomicAdd(&ss->playouts, 0u);
            if (pl) {
                atomicAdd(&ctr->select_levels, (unsigned long long)atomicAdd(&ss->levels, 0u));
                atomicAdd(&ctr->select_children, (unsigned long long)atomicAdd(&ss->children, 0u));
                atomicAdd(&ctr->playouts, (unsigned long long)pl);
            }
        }
    }
}

// K1 for one game (one warp): root -> leaf, leaf position, terminal test, leaf batch slot + planes
template <bool PURE>
__device__ __forceinline__ void gz_select_game(const TreeParams &p, int g, int lane, unsigned &levels_out,
                                               unsigned &children_out, unsigned &playouts_out)
{
    levels_out = children_out = playouts_out = 0;
    if (!p.active[g]) {
        if (lane == 0) { p.leaf_state[g] = 2; p.path_len[g] = 0; }
        return;
    }
    GameMeta m = p.meta[g];
    BB s0 = load_bb(p.boards[g].stones[0]), s1 = load_bb(p.boards[g].stones[1]);
    int player = (m.start_player + m.n_stones) & 1;
    int ns = m.n_stones, last = m.last_move;
    int node = p.root_node[g];
    uint32_t Np = p.root_N[g];
    int depth = 0;
    unsigned int nchild = 0;
    const size_t nbase = (size_t)g * p.cap;
    while (node >= 0) {
        NodeView v = gz_node_view(p.pool, nbase + node, p.Sp);
        const double sq = gz_sqrt_visits(Np);
        double best = -1.0e300;
        int besti = 0x7fffffff, bestc = -1;
        uint32_t bestn = 0u;
        for (int i0 = lane; i0 < p.Sp; i0 += 2 * GZ_WARP) {
            // two child slots per lane per trip, all eight loads in flight before any use
            const int i1 = i0 + GZ_WARP;
            const bool two = i1 < p.Sp;
            const float pr0 = v.P[i0], pr1 = two ? v.P[i1] : -1.f;
            const uint32_t n0 = v.N[i0], n1 = two ? v.N[i1] : 0u;
            const double q0 = v.Q[i0], q1 = two ? v.Q[i1] : 0.0;
            const int c0 = v.child[i0], c1 = two ? v.child[i1] : -1;
            if (pr0 >= 0.f) {
                const double cp = PURE ? ((double)p.c_puct * (1.0 / (double)pr0)) : (double)(p.c_puct * pr0);
                const double s = q0 + cp * sq / (double)(1u + n0);
                if (s > best) { best = s; besti = i0; bestn = n0; bestc = c0; }
                ++nchild;
            }
            if (pr1 >= 0.f) {
                const double cp = PURE ? ((double)p.c_puct * (1.0 / (double)pr1)) : (double)(p.c_puct * pr1);
                const double s = q1 + cp * sq / (double)(1u + n1);
                if (s > best) { best = s; besti = i1; bestn = n1; bestc = c1; }
                ++nchild;
            }
        }
        warp_argmax(best, besti);
        const int mv = besti;
        const int owner = mv & (GZ_WARP - 1); // lane whose local best is the global best
        Np = __shfl_sync(0xffffffffu, bestn, owner);
        const int next = __shfl_sync(0xffffffffu, bestc, owner);
        if (lane == 0) p.path[(size_t)g * p.S + depth] = (uint32_t)node * p.Sp + mv;
        ++depth;
        {
            const int bit = bb_bit(mv, p.W);
            bb_set_if(s0, bit, player == 0);
            bb_set_if(s1, bit, player == 1);
        }
        player ^= 1;
        last = mv;
        ++ns;
        node = next; // a terminal child is never expanded, so child stays -1 and we stop here
    }
    // game_end at the leaf (game.py:162-169): lanes 0-7 each test one (colour, direction)
    bool w0 = false, w1 = false;
    if (ns >= p.n_in_row + 2) {
        const int d4 = lane & 3;
        const int dir = (d4 == 0) ? 1 : (d4 == 1) ? 16 : (d4 == 2) ? 17 : 15;
        bool line = false;
        if (lane < 8) line = bb_has_line_dir(bb_pick(s0, s1, lane < 4), p.n_in_row, dir);
        const unsigned b = __ballot_sync(0xffffffffu, line);
        w0 = (b & 0x0Fu) != 0u;
        w1 = (b & 0xF0u) != 0u;
    }
    const int win = w0 ? 1 : (w1 ? 2 : -1);
    const bool end = (w0 || w1) || ns >= p.W * p.H;
    if (lane == 0) {
        p.path_len[g] = depth;
        GameMeta lm;
        lm.n_stones = ns; lm.last_move = last; lm.start_player = m.start_player; lm.flags = 0;
        p.leaf_meta[g] = lm;
        store_bb(p.leaf_board[g].stones[0], s0);
        store_bb(p.leaf_board[g].stones[1], s1);
    }
    if (end) {
        // mcts_alphaZero.py:140-145: 0 on a tie, else +1/-1 for the side to move
        if (lane == 0) {
            p.leaf_state[g] = 1;
            p.leaf_value[g] = (win == -1) ? 0.0 : ((win == player + 1) ? 1.0 : -1.0);
        }
    } else {
        int slot = 0;
        if (lane == 0) {
            if (!PURE) slot = atomicAdd(p.leaf_count, 1); // pure MCTS rolls out in place: no leaf batch
            p.leaf_state[g] = 0;
            p.leaf_slot[g] = slot;
        }
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (!PURE) {
            if (p.leaf_planes)
                gz_write_planes(bb_pick(s0, s1, player == 0), bb_pick(s1, s0, player == 0), last, ns, p.W, p.H,
                                p.leaf_planes + (size_t)slot * 4 * p.S, lane);
            else if (lane == 0)
                p.slot_game[slot] = g;
        }
    }
    levels_out = (unsigned)depth;
    children_out = __reduce_add_sync(0xffffffffu, nchild);
    playouts_out = 1u;
}

template <bool PURE>
__global__ void __launch_bounds__(128, 8) select_kernel(TreeParams p)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    __shared__ SelStats ss;
    if (threadIdx.x == 0) { ss.levels = 0; ss.children = 0; ss.playouts = 0; ss.done = 0; }
    __syncthreads();
    unsigned lv = 0, ch = 0, pl = 0;
    if (g < p.G) gz_select_game<PURE>(p, g, lane, lv, ch, pl);
    gz_stats_finish(&ss, p.ctr, lane, lv, ch, pl);
}

// backup along the recorded path (lanes = levels), then the root
__device__ __forceinline__ void gz_backup(const TreeParams &p, int g, int d, double v, int lane)
{
    const size_t nbase = (size_t)g * p.cap;
    for (int k = lane; k < d; k += GZ_WARP) {
        uint32_t e = p.path[(size_t)g * p.S + k];
        NodeView nv = gz_node_view(p.pool, nbase + e / p.Sp, p.Sp);
        int cell = e % p.Sp;
        double val = ((d - (k + 1)) & 1) ? v : -v; // leaf node gets -v, alternating upwards
        uint32_t n = nv.N[cell] + 1u;
        nv.N[cell] = n;
        double q = nv.Q[cell];
        nv.Q[cell] = q + 1.0 * (val - q) / (double)n;
    }
    if (lane == 0) {
        double val = (d & 1) ? v : -v;
        uint32_t n = p.root_N[g] + 1u;
        p.root_N[g] = n;
        double q = p.root_Q[g];
        p.root_Q[g] = q + 1.0 * (val - q) / (double)n;
    }
}

// allocate a node block for game g; returns -1 (and counts an overflow) if the pool is empty
__device__ __forceinline__ int gz_alloc_node(const TreeParams &p, int g, int lane)
{
    int n = -1;
    if (lane == 0) {
        int top = p.free_top[g];
        if (top > 0) {
            n = p.free_stack[(size_t)g * p.cap + top - 1];
            p.free_top[g] = top - 1;
        } else {
            atomicAdd(&p.ctr->pool_overflows, 1ull);
        }
    }
    return __shfl_sync(0xffffffffu, n, 0);
}

__device__ __forceinline__ void gz_link_node(const TreeParams &p, int g, int d, int n, int lane)
{
    if (lane != 0) return;
    if (d == 0) {
        p.root_node[g] = n;
    } else {
        uint32_t e = p.path[(size_t)g * p.S + d - 1];
        NodeView pv = gz_node_view(p.pool, (size_t)g * p.cap + e / p.Sp, p.Sp);
        pv.child[e % p.Sp] = n;
    }
}

// K2 (AlphaZero) for one game: expand with the evaluator's priors (mcts_alphaZero.py:38-45, :137)
// + backup.
__device__ __forceinline__ void gz_expand_backup_game(const TreeParams &p, int g, int lane,
                                                      const float *__restrict__ priors,
                                                      const float *__restrict__ values, int compacted)
{
    const int st = p.leaf_state[g];
    if (st == 2) return;
    const int d = p.path_len[g];
    double v;
    if (st == 0) {
        const int slot = compacted ? p.leaf_slot[g] : g;
        v = (double)values[slot];
        int n = gz_alloc_node(p, g, lane);
        if (n >= 0) {
            BB occ = bb_or(load_bb(p.leaf_board[g].stones[0]), load_bb(p.leaf_board[g].stones[1]));
            NodeView nv = gz_node_view(p.pool, (size_t)g * p.cap + n, p.Sp);
            const float *pr = priors + (size_t)slot * p.S;
            unsigned int cnt = 0;
            for (int i = lane; i < p.Sp; i += GZ_WARP) {
                bool leg = (i < p.S) && !bb_test(occ, bb_bit(i, p.W));
                nv.P[i] = leg ? pr[i] : -1.f;
                nv.N[i] = 0u;
                nv.Q[i] = 0.0;
                nv.child[i] = -1;
                cnt += leg;
            }
            gz_link_node(p, g, d, n, lane);
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            if (lane == 0) {
                atomicAdd(&p.ctr->expanded_children, (unsigned long long)cnt);
                atomicAdd(&p.ctr->leaf_evals, 1ull);
            }
        }
    } else {
        v = p.leaf_value[g];
    }
    gz_backup(p, g, d, v, lane);
}

__global__ void __launch_bounds__(128)
expand_backup_kernel(TreeParams p, const float *__restrict__ priors, const float *__restrict__ values,
                     int compacted)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (g >= p.G) return;
    gz_expand_backup_game(p, g, lane, priors, values, compacted);
}

// Fused wave boundary (NSL = Sp / 32 child slots per lane, instantiated for 1-4 and 8), written
// around the number of dependent memory round trips, which is what bounds these kernels:
//   trip 1  everything addressed by the game index alone (flags, path length, leaf slot, root,
//           free-list top, both positions, this lane's path entry)
//   trip 2  everything addressed by trip 1's values: the evaluator's value and priors, the free
//           node index, this lane's path statistics (N, Q) and the whole root block
//   then    expand / link / backup from registers, patch the prefetched root block with the one
//           edge the backup changed, and descend: one trip per level below the root.
// Same arithmetic, in the same order per node, as gz_expand_backup_game + gz_select_game.
template <int NSL, int WT>
__device__ __forceinline__ void gz_expand_select_game(const TreeParams &p, int g, int lane,
                                                      const float *__restrict__ priors,
                                                      const float *__restrict__ values, unsigned &levels_out,
                                                      unsigned &children_out, unsigned &playouts_out)
{
    constexpr int SP = 32 * NSL;
    // board side as a compile-time constant where the launcher has one (WT > 0; the net path is
    // square): cell <-> bit conversions and the plane writer become shifts instead of divisions
    const int BW = WT > 0 ? WT : p.W, BH = WT > 0 ? WT : p.H, BS = WT > 0 ? WT * WT : p.S;
    levels_out = children_out = playouts_out = 0;
    const size_t nbase = (size_t)g * p.cap;
    // ---- trip 1
    const int active = p.active[g];
    const int st = p.leaf_state[g];
    const int d = p.path_len[g];
    const int slot = p.leaf_slot[g];
    const int top = p.free_top[g];
    int root = p.root_node[g];
    uint32_t rootN = p.root_N[g];
    const double rootQ = p.root_Q[g];
    const double lval = p.leaf_value[g];
    const GameMeta m = p.meta[g];
    // the position travels down the tree as ONE 64-bit word per lane: lane l < 8 holds word l & 3 of colour l >> 2
    // (a stone is one OR in one lane; the eight words are put together again only for the win test at the leaf)
    uint64_t bw = lane < 8 ? p.boards[g].stones[lane >> 2][lane & 3] : 0ull;
    const BB occ = bb_or(load_bb(p.leaf_board[g].stones[0]), load_bb(p.leaf_board[g].stones[1]));
    const uint32_t e_mine = p.path[(size_t)g * BS + (lane < BS ? lane : 0)];
    // ---- trip 2
    const bool do_exp = st == 0, do_bk = st != 2;
    float vnet = 0.f, prr[NSL];
    int newn = -1;
    // Everything above reads state only tree kernels write; the evaluator's outputs below come from the net
    // kernels before this one in the chain, and from here on this kernel writes what they read (leaf planes,
    // leaf counter): programmatic-dependent-launch boundary.
    gz_pdl_wait();
    gz_pdl_trigger();
    if (g == 0 && lane == 0 && p.leaf_zero) *p.leaf_zero = 0; // the counter the net just consumed (gzk_expand_select)
    if (do_exp) {
        vnet = values[slot];
#pragma unroll
        for (int j = 0; j < NSL; ++j) {
            const int i = lane + 32 * j;
            prr[j] = i < BS ? priors[(size_t)slot * BS + i] : 0.f;
        }
        if (top > 0) newn = p.free_stack[nbase + top - 1];
    }
    const bool has_e = do_bk && lane < d;
    NodeView ev = gz_node_view(p.pool, nbase + (has_e ? e_mine / SP : 0u), SP);
    const int cell_e = (int)(e_mine % SP);
    uint32_t n_e = 0u;
    double q_e = 0.0;
    if (has_e) { n_e = ev.N[cell_e]; q_e = ev.Q[cell_e]; }
    float rP[NSL];
    uint32_t rN[NSL];
    double rQ[NSL];
    int rC[NSL];
    if (root >= 0) {
        NodeView rv = gz_node_view(p.pool, nbase + root, SP);
#pragma unroll
        for (int j = 0; j < NSL; ++j) {
            const int i = lane + 32 * j;
            if (i < BS) { rP[j] = rv.P[i]; rN[j] = rv.N[i]; rQ[j] = rv.Q[i]; rC[j] = rv.child[i]; }
            else rP[j] = -1.f; // padding slots are never children: their sectors are not fetched
        }
    }
    // ---- expand (mcts_alphaZero.py:38-45) and link
    if (do_exp) {
        if (newn >= 0) {
            NodeView nv = gz_node_view(p.pool, nbase + newn, SP);
            unsigned cnt = 0;
#pragma unroll
            for (int j = 0; j < NSL; ++j) {
                const int i = lane + 32 * j;
                const bool leg = (i < BS) && !bb_test(occ, bb_bit(i, BW));
                const float pv = leg ? prr[j] : -1.f;
                nv.P[i] = pv; nv.N[i] = 0u; nv.Q[i] = 0.0; nv.child[i] = -1;
                cnt += leg;
                if (d == 0) { rP[j] = pv; rN[j] = 0u; rQ[j] = 0.0; rC[j] = -1; } // the new node is the root
            }
            if (d == 0) {
                root = newn;
                if (lane == 0) p.root_node[g] = newn;
            } else if (d <= 32) {
                if (lane == d - 1) ev.child[cell_e] = newn;
            } else if (lane == 0) {
                const uint32_t e = p.path[(size_t)g * BS + d - 1];
                gz_node_view(p.pool, nbase + e / SP, SP).child[e % SP] = newn;
            }
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            if (lane == 0) {
                p.free_top[g] = top - 1;
                atomicAdd(&p.ctr->expanded_children, (unsigned long long)cnt);
                atomicAdd(&p.ctr->leaf_evals, 1ull);
            }
        } else if (lane == 0) {
            atomicAdd(&p.ctr->pool_overflows, 1ull);
        }
    }
    // ---- backup (mcts_alphaZero.py:62-79): lanes = levels, then the root
    if (do_bk) {
        const double v = do_exp ? (double)vnet : lval;
        uint32_t n_new = 0u;
        double q_new = 0.0;
        if (has_e) {
            const double val = ((d - (lane + 1)) & 1) ? v : -v; // leaf node gets -v, alternating upwards
            n_new = n_e + 1u;
            q_new = q_e + 1.0 * (val - q_e) / (double)n_new;
            ev.N[cell_e] = n_new;
            ev.Q[cell_e] = q_new;
        }
        for (int k = lane + 32; k < d; k += 32) { // paths deeper than a warp: the plain way
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
        if (d >= 1) { // path entry 0 is an edge of the (prefetched) root block: patch it
            const int c0 = __shfl_sync(0xffffffffu, cell_e, 0);
            const uint32_t nn = __shfl_sync(0xffffffffu, n_new, 0);
            const double qq = __shfl_sync(0xffffffffu, q_new, 0);
#pragma unroll
            for (int j = 0; j < NSL; ++j)
                if (lane + 32 * j == c0) {
                    rN[j] = nn; rQ[j] = qq;
                    if (d == 1 && newn >= 0) rC[j] = newn;
                }
        }
    }
    __syncwarp(); // this warp's tree writes are ordered before the reads of the descent below
    // ---- select (mcts_alphaZero.py:47-60, 118-129)
    if (!active) {
        if (lane == 0) { p.leaf_state[g] = 2; p.path_len[g] = 0; }
        return;
    }
    int player = (m.start_player + m.n_stones) & 1;
    int ns = m.n_stones, last = m.last_move;
    int node = root;
    uint32_t Np = rootN;
    int depth = 0;
    unsigned int nchild = 0;
    while (node >= 0) {
        const double sq = gz_sqrt_visits(Np);
        double best = -1.0e300;
        int besti = 0x7fffffff, bestc = -1;
        uint32_t bestn = 0u;
#pragma unroll
        for (int j = 0; j < NSL; ++j) {
            if (rP[j] >= 0.f) {
                const double cp = (double)(p.c_puct * rP[j]);
                const double sc = rQ[j] + cp * sq / (double)(1u + rN[j]);
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
        if (next >= 0) { // next level's block: issue the loads before the board bookkeeping
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
        node = next;
    }
    bool w0 = false, w1 = false;
    if (ns >= p.n_in_row + 2) {
        // lanes 0-3 test colour 0, lanes 4-7 colour 1, one direction each: gather that colour's words
        const int d4 = lane & 3;
        bool line = false;
        if (WT > 0 && WT <= 8) { // rows 0-7 are words 0 and 1: the test runs on one squeezed 64-bit word
            const uint64_t c0 = __shfl_sync(0xffffffffu, bw, lane & 4), c1 = __shfl_sync(0xffffffffu, bw, (lane & 4) + 1);
            if (lane < 8) line = bb8_has_line_dir(c0, c1, BW, p.n_in_row, d4);
        } else {
            const int dir = (d4 == 0) ? 1 : (d4 == 1) ? 16 : (d4 == 2) ? 17 : 15;
            BB c;
#pragma unroll
            for (int k = 0; k < 4; ++k) c.w[k] = __shfl_sync(0xffffffffu, bw, (lane & 4) + k);
            if (lane < 8) line = bb_has_line_dir(c, p.n_in_row, dir);
        }
        const unsigned b = __ballot_sync(0xffffffffu, line);
        w0 = (b & 0x0Fu) != 0u;
        w1 = (b & 0xF0u) != 0u;
    }
    const int win = w0 ? 1 : (w1 ? 2 : -1);
    const bool end = (w0 || w1) || ns >= BW * BH;
    int nslot = 0;
    if (lane < 8) p.leaf_board[g].stones[lane >> 2][lane & 3] = bw;
    if (lane == 0) {
        if (!end) nslot = atomicAdd(p.leaf_count, 1);
        p.path_len[g] = depth;
        GameMeta lm;
        lm.n_stones = ns; lm.last_move = last; lm.start_player = m.start_player; lm.flags = 0;
        p.leaf_meta[g] = lm;
        if (end) {
            p.leaf_state[g] = 1;
            p.leaf_value[g] = (win == -1) ? 0.0 : ((win == player + 1) ? 1.0 : -1.0);
        } else {
            p.leaf_state[g] = 0;
            p.leaf_slot[g] = nslot;
            if (!p.leaf_planes) p.slot_game[nslot] = g;
        }
    }
    if (p.leaf_planes) { // evaluators that take fp32 planes (not the fused tcgen05 search path): rebuild both boards
        BB s0, s1;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            s0.w[k] = __shfl_sync(0xffffffffu, bw, k);
            s1.w[k] = __shfl_sync(0xffffffffu, bw, 4 + k);
        }
        nslot = __shfl_sync(0xffffffffu, nslot, 0);
        if (!end)
            gz_write_planes(bb_pick(s0, s1, player == 0), bb_pick(s1, s0, player == 0), last, ns, BW, BH,
                            p.leaf_planes + (size_t)nslot * 4 * BS, lane);
    }
    levels_out = (unsigned)depth;
    children_out = __reduce_add_sync(0xffffffffu, nchild);
    playouts_out = 1u;
}

#ifndef GZ_ES_MINB
#define GZ_ES_MINB 7 // 72 registers, 28 warps per SM: 8 blocks (64 registers) spill on the dependent chain, C2 53.4 -> 55.0 M playouts/s
#endif
template <int NSL, int WT>
__global__ void __launch_bounds__(128, NSL <= 2 ? GZ_ES_MINB : (NSL <= 4 ? 5 : 3))
expand_select_small_kernel(TreeParams p, const float *__restrict__ priors, const float *__restrict__ values)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    __shared__ SelStats ss;
    if (threadIdx.x == 0) { ss.levels = 0; ss.children = 0; ss.playouts = 0; ss.done = 0; }
    __syncthreads();
    unsigned lv = 0, ch = 0, pl = 0;
    if (g < p.G) gz_expand_select_game<NSL, WT>(p, g, lane, priors, values, lv, ch, pl);
    else { gz_pdl_wait(); gz_pdl_trigger(); }
    gz_stats_finish(&ss, p.ctr, lane, lv, ch, pl);
}

// Wave boundary fused: expand + backup of playout k, then select of playout k+1, by the same warp
// of the same game (the path nodes it just updated are still in cache).
__global__ void __launch_bounds__(128, 6)
expand_select_kernel(TreeParams p, const float *__restrict__ priors, const float *__restrict__ values)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    __shared__ SelStats ss;
    if (threadIdx.x == 0) { ss.levels = 0; ss.children = 0; ss.playouts = 0; ss.done = 0; }
    if (blockIdx.x == 0 && threadIdx.x == 0 && p.leaf_zero) *p.leaf_zero = 0; // see gzk_expand_select
    __syncthreads();
    unsigned lv = 0, ch = 0, pl = 0;
    if (g < p.G) {
        gz_expand_backup_game(p, g, lane, priors, values, 1);
        __syncwarp(); // the tree writes of this warp are ordered before the reads of the next descent
        gz_select_game<false>(p, g, lane, lv, ch, pl);
    }
    gz_stats_finish(&ss, p.ctr, lane, lv, ch, pl);
}

__global__ void tree_reset_kernel(TreeParams p)
{
    const int g = blockIdx.x;
    for (int i = threadIdx.x; i < p.cap; i += blockDim.x)
        p.free_stack[(size_t)g * p.cap + i] = p.cap - 1 - i;
    if (threadIdx.x == 0) {
        p.free_top[g] = p.cap;
        p.root_node[g] = -1;
        p.root_N[g] = 0u;
        p.root_Q[g] = 0.0;
    }
}

__global__ void root_visits_kernel(TreeParams p, int32_t *visits, double *q, int32_t *rootn)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (g >= p.G) return;
    int r = p.root_node[g];
    if (lane == 0 && rootn) rootn[g] = (int32_t)p.root_N[g];
    for (int i = lane; i < p.S; i += GZ_WARP) {
        int32_t vis = -1;
        double qq = 0.0;
        if (r >= 0) {
            NodeView v = gz_node_view(p.pool, (size_t)g * p.cap + r, p.Sp);
            if (v.P[i] >= 0.f) { vis = (int32_t)v.N[i]; qq = v.Q[i]; }
        }
        if (visits) visits[(size_t)g * p.S + i] = vis;
        if (q) q[(size_t)g * p.S + i] = qq;
    }
}

// MCTS.update_with_move (mcts_alphaZero.py:183-192).  moves[g] == -2: leave this game alone.
__device__ void gz_reroot(const TreeParams &p, int g, int mv, int lane)
{
    const size_t nbase = (size_t)g * p.cap;
    int r = p.root_node[g];
    if (r < 0) {
        // a leaf root has no children: "last_move in children" is false -> fresh TreeNode
        if (lane == 0) { p.root_N[g] = 0u; p.root_Q[g] = 0.0; }
        return;
    }
    NodeView rv = gz_node_view(p.pool, nbase + r, p.Sp);
    bool is_child = (mv >= 0 && mv < p.S) && (rv.P[mv] >= 0.f);
    int keep = -1;
    uint32_t newN = 0u;
    double newQ = 0.0;
    if (is_child) { keep = rv.child[mv]; newN = rv.N[mv]; newQ = rv.Q[mv]; }
    int32_t *fs = p.free_stack + nbase;
    int top = p.free_top[g];
    int scan = top;
    __syncwarp();
    if (lane == 0) fs[top] = r;
    ++top;
    __syncwarp();
    // Breadth-first over the dropped subtree, RU queued nodes per round: their child arrays are loaded
    // together (one dependent DRAM round trip per round instead of one per node - a steady-state tree
    // drops ~1500 nodes per move), then pushed from registers.
    constexpr int RU = 8;
    while (scan < top) {
        const int nb = min(RU, top - scan);
        const int mine = lane < nb ? fs[scan + lane] : -1;
        for (int i0 = 0; i0 < p.Sp; i0 += GZ_WARP) {
            const int i = i0 + lane;
            int c[RU];
#pragma unroll
            for (int u = 0; u < RU; ++u) {
                const int node = __shfl_sync(0xffffffffu, mine, u);
                c[u] = node >= 0 ? gz_node_view(p.pool, nbase + node, p.Sp).child[i] : -1;
                if (node == r && is_child && i == mv) c[u] = -1; // the kept subtree stays
            }
#pragma unroll
            for (int u = 0; u < RU; ++u) {
                const bool push = c[u] >= 0;
                const unsigned b = __ballot_sync(0xffffffffu, push);
                if (push) fs[top + __popc(b & ((1u << lane) - 1u))] = c[u];
                top += __popc(b);
            }
        }
        scan += nb;
        __syncwarp();
    }
    if (lane == 0) {
        p.free_top[g] = top;
        p.root_node[g] = keep;
        p.root_N[g] = newN;
        p.root_Q[g] = newQ;
    }
}

__global__ void __launch_bounds__(128) update_with_move_kernel(TreeParams p, const int32_t *moves)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (g >= p.G) return;
    int mv = moves[g];
    if (mv == -2) return;
    gz_reroot(p, g, mv, lane);
}

// ---- host launchers -----------------------------------------------------------------------
static TreeParams gz_tree_params(gz_engine *e)
{
    TreeParams p;
    p.boards = e->boards; p.meta = e->meta; p.active = e->active; p.pool = e->pool;
    p.free_stack = e->free_stack; p.free_top = e->free_top; p.root_node = e->root_node;
    p.root_N = e->root_N; p.root_Q = e->root_Q; p.path = e->path; p.path_len = e->path_len;
    p.leaf_state = e->leaf_state; p.leaf_value = e->leaf_value; p.leaf_slot = e->leaf_slot;
    p.leaf_count = e->leaf_count; p.leaf_zero = nullptr; p.leaf_planes = e->leaf_packed ? nullptr : e->leaf_planes; p.slot_game = e->slot_game;
    p.leaf_board = e->leaf_board;
    p.leaf_meta = e->leaf_meta; p.ctr = e->counters;
    p.G = e->G; p.W = e->W; p.H = e->H; p.S = e->S; p.Sp = e->Sp; p.n_in_row = e->n_in_row;
    p.cap = e->cap; p.c_puct = e->cfg.c_puct; p.g0 = 0;
    return p;
}

static inline unsigned gz_warp_blocks(int G, int wpb) { return (unsigned)((G + wpb - 1) / wpb); }

int gzk_tree_reset(gz_engine *e)
{
    tree_reset_kernel<<<e->G, 128, 0, e->stream>>>(gz_tree_params(e));
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

int gzk_select(gz_engine *e, bool pure)
{
    // the leaf counter is one of a pair (see gzk_expand_select): a search starts on the first with both clear
    e->leaf_count = e->leaf_count_pair;
    GZ_CUDA(e, cudaMemsetAsync(e->leaf_count_pair, 0, 2 * sizeof(int32_t), e->stream));
    TreeParams p = gz_tree_params(e);
    ProfScope ps(e, PC_SELECT);
    if (pure) select_kernel<true><<<gz_warp_blocks(e->G, 4), 128, 0, e->stream>>>(p);
    else select_kernel<false><<<gz_warp_blocks(e->G, 4), 128, 0, e->stream>>>(p);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

int gzk_expand_backup(gz_engine *e, const float *d_priors, const float *d_values, bool compacted)
{
    ProfScope ps(e, PC_EXPAND);
    expand_backup_kernel<<<gz_warp_blocks(e->G, 4), 128, 0, e->stream>>>(gz_tree_params(e), d_priors,
                                                                          d_values, compacted ? 1 : 0);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

int gzk_expand_select(gz_engine *e, const float *d_priors, const float *d_values)
{
    // No memset between the net kernels and this one: the leaves of the next wave are counted in the OTHER
    // counter of the pair (left at zero by the fused kernel before this one, or by gzk_select), and this
    // kernel resets the counter the net has just consumed.  e->leaf_count always names the counter the next
    // net call reads.
    int32_t *consumed = e->leaf_count;
    e->leaf_count = consumed == e->leaf_count_pair ? e->leaf_count_pair + 1 : e->leaf_count_pair;
    ProfScope ps(e, PC_SELECT);
    const unsigned grid = gz_warp_blocks(e->G, 4);
    TreeParams tp = gz_tree_params(e);
    tp.leaf_zero = consumed;
    const bool sq = e->W == e->H;
    // round-trip-minimal kernel: NSL = Sp / 32 child slots per lane, board side compile-time for the
    // reference's usual sizes; other sizes take the generic fused kernel
#define GZ_ES(NSL, WT) GZ_CUDA(e, gz_launch_chain(expand_select_small_kernel<NSL, WT>, dim3(grid), dim3(128), 0, e->stream, tp, d_priors, d_values))
    if (e->Sp == 64 && sq && e->W == 8) GZ_ES(2, 8);
    else if (e->Sp == 64 && sq && e->W == 6) GZ_ES(2, 6);
    else if (e->Sp == 64) GZ_ES(2, 0);
    else if (e->Sp == 32) GZ_ES(1, 0);
    else if (e->Sp == 96 && sq && e->W == 9) GZ_ES(3, 9);
    else if (e->Sp == 96) GZ_ES(3, 0);
    else if (e->Sp == 128) GZ_ES(4, 0);
    else if (e->Sp == 256 && sq && e->W == 15) GZ_ES(8, 15);
    else if (e->Sp == 256) GZ_ES(8, 0);
#undef GZ_ES
    else expand_select_kernel<<<grid, 128, 0, e->stream>>>(tp, d_priors, d_values);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

int gzk_root_visits(gz_engine *e, int32_t *d_visits, double *d_q, int32_t *d_rootn)
{
    root_visits_kernel<<<gz_warp_blocks(e->G, 4), 128, 0, e->stream>>>(gz_tree_params(e), d_visits, d_q,
                                                                        d_rootn);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

int gzk_update_with_move(gz_engine *e, const int32_t *d_moves)
{
    update_with_move_kernel<<<gz_warp_blocks(e->G, 4), 128, 0, e->stream>>>(gz_tree_params(e), d_moves);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}
