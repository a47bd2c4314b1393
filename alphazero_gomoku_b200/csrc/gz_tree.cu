// gz_tree.cu — GPU-resident MCTS trees (reference mcts_alphaZero.py:24-192, mcts_pure.py:49-243).
//   K1 select: one warp per game walks root -> leaf; at each level the lanes score the
//      children (Q + u, u = c_puct * P * sqrt(N_parent) / (1 + N), all in the reference's
//      mixed float32/float64 arithmetic) and a shuffle argmax with
//      lowest-index tie-break reproduces Python's first-max (mcts_alphaZero.py:55).
//   K2 expand + backup: writes the priors of the legal moves into a fresh node block and
//      walks the recorded path updating N and the running-mean Q with alternating sign
//      (update_recursive, mcts_alphaZero.py:57-73).  One playout in flight per game, so the
//      tree sees exactly the reference's sequential statistics.
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
    float *leaf_planes;
    GameBoard *leaf_board;
    GameMeta *leaf_meta;
    DevCounters *ctr;
    int G, W, H, S, Sp, n_in_row, cap;
    float c_puct;
};

__device__ __forceinline__ void warp_argmax(double &s, int &i)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        double os = __shfl_xor_sync(0xffffffffu, s, off);
        int oi = __shfl_xor_sync(0xffffffffu, i, off);
        if (os > s || (os == s && oi < i)) { s = os; i = oi; }
    }
}

template <bool PURE>
__global__ void __launch_bounds__(128) select_kernel(TreeParams p)
{
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const int g = blockIdx.x * (blockDim.x >> 5) + wib;
    __shared__ unsigned int s_levels, s_children, s_playouts;
    if (threadIdx.x == 0) { s_levels = 0; s_children = 0; s_playouts = 0; }
    __syncthreads();
    if (g < p.G) {
        if (!p.active[g]) {
            if (lane == 0) { p.leaf_state[g] = 2; p.path_len[g] = 0; }
        } else {
            GameMeta m = p.meta[g];
            BB st[2];
            st[0] = load_bb(p.boards[g].stones[0]);
            st[1] = load_bb(p.boards[g].stones[1]);
            int player = (m.start_player + m.n_stones) & 1;
            int ns = m.n_stones, last = m.last_move;
            int node = p.root_node[g];
            uint32_t Np = p.root_N[g];
            int depth = 0;
            unsigned int nchild = 0;
            const size_t nbase = (size_t)g * p.cap;
            while (node >= 0) {
                NodeView v = gz_node_view(p.pool, nbase + node, p.Sp);
                const double sq = sqrt((double)Np);
                double best = -1.0e300;
                int besti = 0x7fffffff;
                for (int i = lane; i < p.Sp; i += GZ_WARP) {
                    float pr = v.P[i];
                    if (pr >= 0.f) {
                        uint32_t n = v.N[i];
                        double q = v.Q[i];
                        double cp = PURE ? ((double)p.c_puct * (1.0 / (double)pr))
                                         : (double)(p.c_puct * pr);
                        double s = q + cp * sq / (double)(1u + n);
                        if (s > best) { best = s; besti = i; }
                        ++nchild;
                    }
                }
                warp_argmax(best, besti);
                const int mv = besti;
                if (lane == 0) p.path[(size_t)g * p.S + depth] = (uint32_t)node * p.Sp + mv;
                ++depth;
                bb_set(st[player], bb_bit(mv, p.W));
                player ^= 1;
                last = mv;
                ++ns;
                Np = v.N[mv];
                node = v.child[mv];
                // a terminal child is never expanded, so child stays -1 and we stop here
            }
            int win;
            bool hw;
            bool end = bb_game_end(st[0], st[1], ns, p.W, p.H, p.n_in_row, &win, &hw);
            if (lane == 0) {
                p.path_len[g] = depth;
                GameMeta lm;
                lm.n_stones = ns; lm.last_move = last; lm.start_player = m.start_player; lm.flags = 0;
                p.leaf_meta[g] = lm;
                store_bb(p.leaf_board[g].stones[0], st[0]);
                store_bb(p.leaf_board[g].stones[1], st[1]);
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
                    slot = atomicAdd(p.leaf_count, 1);
                    p.leaf_state[g] = 0;
                    p.leaf_slot[g] = slot;
                }
                slot = __shfl_sync(0xffffffffu, slot, 0);
                if (!PURE)
                    gz_write_planes(st[player], st[player ^ 1], last, ns, p.W, p.H,
                                    p.leaf_planes + (size_t)slot * 4 * p.S, lane);
            }
            nchild = __reduce_add_sync(0xffffffffu, nchild);
            if (lane == 0) {
                atomicAdd(&s_levels, (unsigned)depth);
                atomicAdd(&s_children, nchild);
                atomicAdd(&s_playouts, 1u);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0 && s_playouts) {
        atomicAdd(&p.ctr->select_levels, (unsigned long long)s_levels);
        atomicAdd(&p.ctr->select_children, (unsigned long long)s_children);
        atomicAdd(&p.ctr->playouts, (unsigned long long)s_playouts);
    }
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

// K2 (AlphaZero): expand with the evaluator's priors (mcts_alphaZero.py:38-45, :137) + backup.
__global__ void __launch_bounds__(128)
expand_backup_kernel(TreeParams p, const float *__restrict__ priors, const float *__restrict__ values,
                     int compacted)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (g >= p.G) return;
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
    while (scan < top) {
        int node = fs[scan++];
        NodeView v = gz_node_view(p.pool, nbase + node, p.Sp);
        for (int i0 = 0; i0 < p.Sp; i0 += GZ_WARP) {
            int i = i0 + lane;
            int c = v.child[i];
            bool push = c >= 0 && !(node == r && is_child && i == mv);
            unsigned b = __ballot_sync(0xffffffffu, push);
            if (push) fs[top + __popc(b & ((1u << lane) - 1u))] = c;
            top += __popc(b);
        }
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
    p.leaf_count = e->leaf_count; p.leaf_planes = e->leaf_planes; p.leaf_board = e->leaf_board;
    p.leaf_meta = e->leaf_meta; p.ctr = e->counters;
    p.G = e->G; p.W = e->W; p.H = e->H; p.S = e->S; p.Sp = e->Sp; p.n_in_row = e->n_in_row;
    p.cap = e->cap; p.c_puct = e->cfg.c_puct;
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
    GZ_CUDA(e, cudaMemsetAsync(e->leaf_count, 0, sizeof(int32_t), e->stream));
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
