/*
 * ORACLE — TEST INFRASTRUCTURE ONLY (see orc_board.h).
 * Pointer-tree MCTS restated from reference mcts_alphaZero.py:24-192 (TreeNode, MCTS)
 * and mcts_pure.py:15-243 (rollout_policy_fn, policy_value_fn, TreeNode, MCTS).
 * Numeric types follow what the reference actually computes with a Python-float leaf
 * value: prior P float32 (AlphaZero) or float64 (pure), c_puct*P in
 * P's type, sqrt/div/Q in float64.
 */
#include "orc_board.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int W, H, c1, c2, c3;
    const float *p[16];
} orc_net;
int orc_net_forward(const orc_net *net, const float *planes, int B, float *logp, float *value);

enum { ORC_EVAL_NET = 0, ORC_EVAL_HASH = 1, ORC_EVAL_PURE = 2, ORC_EVAL_CALLBACK = 3 };

/* host callback evaluator: fills probs[n_avail] for board->availables order and *value */
typedef void (*orc_eval_cb)(const orc_board *b, float *probs, float *value);

typedef struct orc_node {
    struct orc_node *parent;
    int n_children;
    int *acts;               /* insertion order = evaluator order (ascending moves) */
    struct orc_node **kids;
    int n_visits;
    double Q, u, P;
    int p_f32;               /* prior is np.float32 (AlphaZero) vs float64 (pure) */
    int vl, taken;           /* virtual-loss mode only (orc_mcts_search_vl): in-flight paths through / ending at this node */
} orc_node;

typedef struct {
    orc_node *root;
    double c_puct;
    int n_playout;
    int eval_kind;
    const orc_net *net;
    orc_eval_cb cb;
    uint64_t salt;           /* hash evaluator salt / philox seed for rollouts */
    uint32_t game_id, search_id, playout_idx;
    long long n_evals, n_terminal, n_rollout_plies, n_nodes;
} orc_mcts;

static orc_node *node_new(orc_mcts *m, orc_node *parent, double P, int p_f32)
{
    orc_node *n = (orc_node *)calloc(1, sizeof(orc_node));
    n->parent = parent;
    n->P = P;
    n->p_f32 = p_f32;
    m->n_nodes++;
    return n;
}

static void node_free(orc_node *n)
{
    if (!n) return;
    for (int i = 0; i < n->n_children; ++i) node_free(n->kids[i]);
    free(n->acts);
    free(n->kids);
    free(n);
}

/* mcts_alphaZero.py:38-45 expand (children keyed by action, inserted in evaluator order) */
static void node_expand(orc_mcts *m, orc_node *n, const int *acts, const double *probs, int k,
                        int p_f32)
{
    if (k == 0) return;
    n->acts = (int *)malloc(sizeof(int) * k);
    n->kids = (orc_node **)malloc(sizeof(orc_node *) * k);
    n->n_children = k;
    for (int i = 0; i < k; ++i) {
        n->acts[i] = acts[i];
        n->kids[i] = node_new(m, n, probs[i], p_f32);
    }
}

/* mcts_alphaZero.py:75-83 get_value: u = c_puct * P * sqrt(N_parent) / (1 + N); Q + u */
static double node_get_value(orc_node *n, double c_puct)
{
    double cp;
    if (n->p_f32) cp = (double)((float)c_puct * (float)n->P);
    else cp = c_puct * n->P;
    n->u = cp * sqrt((double)n->parent->n_visits) / (double)(1 + n->n_visits);
    return n->Q + n->u;
}

/* mcts_alphaZero.py:48-55 select: Python max -> first maximum in insertion order */
static int node_select(orc_node *n, double c_puct)
{
    int best = 0;
    double bv = node_get_value(n->kids[0], c_puct);
    for (int i = 1; i < n->n_children; ++i) {
        double v = node_get_value(n->kids[i], c_puct);
        if (v > bv) { bv = v; best = i; }
    }
    return best;
}

/* mcts_alphaZero.py:57-73 update / update_recursive (parent first, with -value) */
static void node_update_recursive(orc_node *n, double leaf_value)
{
    if (n->parent) node_update_recursive(n->parent, -leaf_value);
    n->n_visits += 1;
    n->Q += 1.0 * (leaf_value - n->Q) / (double)n->n_visits;
}

static uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

/* Synthetic, integer-defined evaluator used to drive oracle and CUDA tree kernels with
 * bit-identical priors/values (not in the reference; numpy twin in tests/helpers.py). */
void orc_hash_eval(const orc_board *b, uint64_t salt, float *probs /*[n_avail]*/, float *value)
{
    uint64_t key = splitmix64(salt + 1000ull + (uint64_t)b->current_player);
    int S = b->width * b->height;
    for (int m = 0; m < S; ++m)
        if (b->states[m]) key ^= splitmix64(salt + (uint64_t)(m * 2 + (b->states[m] - 1)) + 1ull);
    uint32_t sum = 0;
    uint32_t w[ORC_MAXS];
    for (int i = 0; i < b->n_avail; ++i) {
        uint64_t h = splitmix64(key + (uint64_t)b->availables[i] + 1ull);
        uint32_t t = (uint32_t)(h >> 56) + 1u;
        w[i] = t * t * t;
        sum += w[i];
    }
    for (int i = 0; i < b->n_avail; ++i) probs[i] = (float)w[i] / (float)sum;
    int vi = (int)(splitmix64(key ^ 0x5555555555555555ull) % 2001ull) - 1000;
    *value = (float)vi / 1000.0f;
}

/* policy_value_net_pytorch.py:96-115 policy_value_fn (value wrapped as float) */
static void eval_net(orc_mcts *m, const orc_board *b, float *probs, float *value)
{
    int S = b->width * b->height;
    double st[4 * ORC_MAXS];
    float pl[4 * ORC_MAXS], logp[ORC_MAXS];
    orc_board_current_state(b, st);
    for (int i = 0; i < 4 * S; ++i) pl[i] = (float)st[i];
    orc_net_forward(m->net, pl, 1, logp, value);
    for (int i = 0; i < b->n_avail; ++i) probs[i] = expf(logp[b->availables[i]]);
}

/* mcts_pure.py:193-212 _evaluate_rollout with rollout_policy_fn (:15-30): each ply a
 * uniformly random move of the current move_target box.  The reference draws
 * np.random.rand per candidate and takes the argmax (= uniform choice); here the choice
 * is index = mulhi32(r[ply % 4], n_target) with r = philox(key, (playout, ply / 4, game, search)),
 * one Philox4x32 call per four plies, bit-identical to the CUDA rollout kernel
 * (csrc/gz_rollout.cu).  An empty candidate box (empty board narrower than 6, game.py:41: the
 * reference raises in max([])) ends the rollout as a tie. */
static int evaluate_rollout(orc_mcts *m, orc_board *b, int limit)
{
    int player = b->current_player;
    int winner = -1, end = 0;
    uint32_t r[4] = {0, 0, 0, 0};
    for (int i = 0; i < limit; ++i) {
        end = orc_board_game_end(b, &winner);
        if (end) break;
        orc_board_get_move_target(b);
        if (b->n_target == 0) { end = 1; winner = -1; break; }
        if ((i & 3) == 0)
            orc_philox4x32((uint32_t)m->salt, (uint32_t)(m->salt >> 32), m->playout_idx, (uint32_t)(i >> 2),
                           m->game_id, m->search_id, r);
        uint32_t idx = (uint32_t)(((uint64_t)r[i & 3] * (uint64_t)b->n_target) >> 32);
        orc_board_do_move(b, b->move_target[idx]);
        m->n_rollout_plies++;
    }
    if (!end) winner = -1; /* limit reached: reference keeps the last winner (-1) */
    if (winner == -1) return 0;
    return winner == player ? 1 : -1;
}

static void playout_leaf(orc_mcts *m, orc_board *state, orc_node *node);

/* mcts_alphaZero.py:113-148 _playout / mcts_pure.py:156-191 _playout. */
static void playout(orc_mcts *m, orc_board *state)
{
    orc_node *node = m->root;
    while (node->n_children != 0) {
        int i = node_select(node, m->c_puct);
        orc_board_do_move(state, node->acts[i]);
        node = node->kids[i];
    }
    playout_leaf(m, state, node);
}

/* the part of _playout after the descent: evaluate, expand, back up (:131-148) */
static void playout_leaf(orc_mcts *m, orc_board *state, orc_node *node)
{
    int acts[ORC_MAXS];
    double probs[ORC_MAXS];
    float pf[ORC_MAXS];
    float vf = 0.f;
    int k = 0, p_f32 = 1;
    double leaf_value = 0.0;
    /* the evaluator is called BEFORE the end check (:132-134 / :176-179) */
    if (m->eval_kind == ORC_EVAL_PURE) {
        orc_board_get_move_target(state);               /* mcts_pure.py:33-46 */
        k = state->n_target;
        for (int i = 0; i < k; ++i) { acts[i] = state->move_target[i]; probs[i] = 1.0 / (double)k; }
        p_f32 = 0;
    } else {
        if (m->eval_kind == ORC_EVAL_NET) eval_net(m, state, pf, &vf);
        else if (m->eval_kind == ORC_EVAL_HASH) orc_hash_eval(state, m->salt, pf, &vf);
        else m->cb(state, pf, &vf);
        k = state->n_avail;
        for (int i = 0; i < k; ++i) { acts[i] = state->availables[i]; probs[i] = (double)pf[i]; }
        leaf_value = (double)vf;
    }
    m->n_evals++;
    int winner;
    int end = orc_board_game_end(state, &winner);
    if (!end) node_expand(m, node, acts, probs, k, p_f32);
    if (m->eval_kind == ORC_EVAL_PURE) {
        leaf_value = (double)evaluate_rollout(m, state, 1000);
    } else if (end) {
        m->n_terminal++;
        if (winner == -1) leaf_value = 0.0;
        else leaf_value = (winner == state->current_player) ? 1.0 : -1.0;
    }
    node_update_recursive(node, -leaf_value);
}

/* Decision-level replay (test infrastructure for the search tier): run one playout whose descent follows
 * a RECORDED path (the moves another implementation's select calls returned: the reference's, captured by
 * tests/golden/make_golden2.py, or the CUDA engine's, gz_search_paths) instead of this tree's own argmax,
 * and measure at every level how far the recorded choice is from the oracle's:
 *   regret  = score(oracle argmax) - score(recorded move)  >= 0, in the reference's arithmetic (:75-83)
 *   gap     = score(best) - score(second best)               (SURVEY App. B near-tie measure)
 * both relative to max(1, |best score|), and divided by the sensitivity of a score difference to an error
 * eps of the evaluator, sens = c_puct * sqrt(N_parent) * (1/(1+n_a) + 1/(1+n_b)) + 2 ("margin": the largest
 * uniform net error that provably cannot flip this decision).  n_tie_order counts recorded moves that tie
 * the best score exactly but are not the FIRST maximum (mcts_alphaZero.py:55: Python max keeps the first).
 * The leaf is then evaluated, expanded and backed up exactly like playout(), so after n playouts the oracle
 * tree is the tree the recorded search built.  Returns 0, or <0 when the path is not a root-to-leaf path
 * of this tree (a structural divergence: that is a parity failure, not a near-tie). */
typedef struct {
    double max_regret_rel, max_regret_margin, min_gap_rel, min_gap_margin;
    long long n_decisions, n_deviations, n_tie_order;
} orc_replay_stats;

void orc_replay_stats_init(orc_replay_stats *st)
{
    memset(st, 0, sizeof(*st));
    st->min_gap_rel = 1e300;
    st->min_gap_margin = 1e300;
}

int orc_mcts_playout_forced(orc_mcts *m, const orc_board *b, const uint8_t *path, int len,
                            orc_replay_stats *st)
{
    orc_board state = *b;
    orc_node *node = m->root;
    for (int j = 0; j < len; ++j) {
        if (node->n_children == 0) return -1;            /* recorded path runs past this tree's leaf */
        int best = 0, second = -1, forced = -1;
        double bv = -1e300, sv = -1e300;
        for (int i = 0; i < node->n_children; ++i) {
            double v = node_get_value(node->kids[i], m->c_puct);
            if (node->acts[i] == (int)path[j]) forced = i;
            if (i == 0) { bv = v; best = 0; }
            else if (v > bv) { sv = bv; second = best; bv = v; best = i; }   /* strict: first maximum wins */
            else if (second < 0 || v > sv) { sv = v; second = i; }
        }
        if (forced < 0) return -2;                       /* recorded move is not a child here */
        double fv = node->kids[forced]->Q + node->kids[forced]->u;
        double sq = sqrt((double)node->n_visits);
        double scale = fabs(bv) > 1.0 ? fabs(bv) : 1.0;
        st->n_decisions++;
        if (forced != best) {
            double regret = bv - fv;
            double sens = m->c_puct * sq * (1.0 / (1 + node->kids[best]->n_visits) +
                                            1.0 / (1 + node->kids[forced]->n_visits)) + 2.0;
            st->n_deviations++;
            if (regret == 0.0) st->n_tie_order++;        /* equal score, but not the first maximum */
            if (regret / scale > st->max_regret_rel) st->max_regret_rel = regret / scale;
            if (regret / sens > st->max_regret_margin) st->max_regret_margin = regret / sens;
        }
        if (second >= 0) {
            double gap = bv - sv;
            double sens = m->c_puct * sq * (1.0 / (1 + node->kids[best]->n_visits) +
                                            1.0 / (1 + node->kids[second]->n_visits)) + 2.0;
            if (gap / scale < st->min_gap_rel) st->min_gap_rel = gap / scale;
            if (gap / sens < st->min_gap_margin) st->min_gap_margin = gap / sens;
        }
        orc_board_do_move(&state, node->acts[forced]);
        node = node->kids[forced];
    }
    if (node->n_children != 0) return -3;                /* recorded path stops above this tree's leaf */
    playout_leaf(m, &state, node);
    return 0;
}

/* n playouts along recorded paths: tlen[k] moves of playout k, stored back to back in trace.
 * Returns 0, or -(8 * (k + 1) + c) for the first bad playout k with c = 1..3 (see orc_mcts_playout_forced). */
int orc_mcts_replay(orc_mcts *m, const orc_board *b, const uint8_t *tlen, const uint8_t *trace, int n,
                    orc_replay_stats *st)
{
    int off = 0;
    for (int k = 0; k < n; ++k) {
        m->playout_idx = (uint32_t)k;
        int rc = orc_mcts_playout_forced(m, b, trace + off, (int)tlen[k], st);
        if (rc != 0) return -((k + 1) * 8) + rc;        /* rc in -1..-3 */
        off += (int)tlen[k];
    }
    return 0;
}

orc_mcts *orc_mcts_new(double c_puct, int n_playout, int eval_kind, const orc_net *net,
                       orc_eval_cb cb, uint64_t salt)
{
    orc_mcts *m = (orc_mcts *)calloc(1, sizeof(orc_mcts));
    m->c_puct = c_puct;
    m->n_playout = n_playout;
    m->eval_kind = eval_kind;
    m->net = net;
    m->cb = cb;
    m->salt = salt;
    m->root = node_new(m, NULL, 1.0, 0);
    return m;
}

void orc_mcts_free(orc_mcts *m)
{
    if (!m) return;
    node_free(m->root);
    free(m);
}

void orc_mcts_set_stream(orc_mcts *m, uint32_t game_id, uint32_t search_id)
{
    m->game_id = game_id;
    m->search_id = search_id;
}

/* mcts_alphaZero.py:154-181 get_move_probs up to the visit counts (the softmax is applied
 * by the caller in float64 numpy exactly like :179); mcts_pure.py:214-231 get_move's loop.
 * Returns the number of root children; acts/visits/q are filled in insertion order. */
int orc_mcts_search(orc_mcts *m, const orc_board *b, int n_playout, int *acts, int *visits,
                    double *q, double *prior)
{
    if (n_playout < 0) n_playout = m->n_playout;
    for (int n = 0; n < n_playout; ++n) {
        orc_board copy = *b;                             /* copy.deepcopy(state) :162 */
        m->playout_idx = (uint32_t)n;
        playout(m, &copy);
    }
    orc_node *r = m->root;
    for (int i = 0; i < r->n_children; ++i) {
        if (acts) acts[i] = r->acts[i];
        if (visits) visits[i] = r->kids[i]->n_visits;
        if (q) q[i] = r->kids[i]->Q;
        if (prior) prior[i] = r->kids[i]->P;
    }
    return r->n_children;
}

/* ---- several leaves in flight per search (virtual loss) ----------------------------------------------------------
 * NOT in the reference (its loop mcts_alphaZero.py:161-163 is strictly sequential); this is the build's own
 * specification of the K-leaves-per-wave mode of the CUDA engine (gz_search_vl), restated on the pointer tree so that
 * the kernels can be checked bit for bit.  A wave: up to K descents from the root, one after another; descent j scores a
 * child as if every earlier descent of this wave that passed through it had already come back as a LOSS for the player
 * choosing there: n_eff = n + c, q_eff = c running-mean updates of Q with -1 (the reference's own update rule :64-65),
 * N_parent = the parent's visits + the paths through the parent; only the first ORC_VL_DEPTH levels of a path carry
 * virtual losses.  A descent that ends at a non-terminal leaf an earlier descent of the wave already holds ends the
 * wave early (that leaf is about to be expanded).  Then the leaves are evaluated, expanded and backed up IN ORDER with
 * the reference's arithmetic - the statistics never contain a virtual loss.  K = 1 is the reference's loop. */
#define ORC_VL_DEPTH 32
#define ORC_VL_MAXK 16

static double vl_score(const orc_node *c, double c_puct, double np_eff)
{
    double cp = c->p_f32 ? (double)((float)c_puct * (float)c->P) : c_puct * c->P;
    double q = c->Q;
    for (int t = 0; t < c->vl; ++t) q += 1.0 * (-1.0 - q) / (double)(c->n_visits + t + 1);
    return q + cp * sqrt(np_eff) / (double)(1 + c->n_visits + c->vl);
}

int orc_mcts_search_vl(orc_mcts *m, const orc_board *b, int n_playout, int K, int *waves_out)
{
    if (K < 1 || K > ORC_VL_MAXK) return -1;
    int done = 0, waves = 0;
    orc_node *path[ORC_VL_MAXK][ORC_MAXS + 1];
    int plen[ORC_VL_MAXK];
    orc_board leafb[ORC_VL_MAXK];
    while (done < n_playout) {
        int kmax = n_playout - done < K ? n_playout - done : K, keff = 0;
        for (int j = 0; j < kmax; ++j) {
            orc_board st = *b;
            orc_node *node = m->root;
            int d = 0;
            path[j][0] = node;
            while (node->n_children != 0) {
                double np_eff = (double)(node->n_visits + node->vl);
                int best = 0;
                double bv = vl_score(node->kids[0], m->c_puct, np_eff);
                for (int i = 1; i < node->n_children; ++i) {
                    double v = vl_score(node->kids[i], m->c_puct, np_eff);
                    if (v > bv) { bv = v; best = i; }
                }
                orc_board_do_move(&st, node->acts[best]);
                node = node->kids[best];
                path[j][++d] = node;
            }
            int winner, end = orc_board_game_end(&st, &winner);
            if (!end && node->taken) break;              /* an earlier descent of this wave holds this leaf */
            for (int i = 0; i <= d && i <= ORC_VL_DEPTH; ++i) path[j][i]->vl++;   /* node at depth i = edge i - 1 */
            if (!end) node->taken = 1;
            plen[j] = d;
            leafb[j] = st;
            keff = j + 1;
        }
        for (int j = 0; j < keff; ++j) {
            for (int i = 0; i <= plen[j] && i <= ORC_VL_DEPTH; ++i) path[j][i]->vl--;
            path[j][plen[j]]->taken = 0;
        }
        for (int j = 0; j < keff; ++j) {
            m->playout_idx = (uint32_t)(done + j);
            playout_leaf(m, &leafb[j], path[j][plen[j]]);
        }
        done += keff;
        ++waves;
    }
    if (waves_out) *waves_out = waves;
    return 0;
}

int orc_mcts_root_visits(const orc_mcts *m) { return m->root->n_visits; }
double orc_mcts_root_q(const orc_mcts *m) { return m->root->Q; }
long long orc_mcts_stat(const orc_mcts *m, int which)
{
    switch (which) {
    case 0: return m->n_evals;
    case 1: return m->n_terminal;
    case 2: return m->n_rollout_plies;
    case 3: return m->n_nodes;
    }
    return -1;
}

/* mcts_alphaZero.py:183-192 update_with_move */
void orc_mcts_update_with_move(orc_mcts *m, int last_move)
{
    orc_node *r = m->root;
    for (int i = 0; i < r->n_children; ++i)
        if (r->acts[i] == last_move) {
            orc_node *c = r->kids[i];
            r->kids[i] = NULL;
            c->parent = NULL;
            node_free(r);
            m->root = c;
            return;
        }
    node_free(r);
    m->root = node_new(m, NULL, 1.0, 0);
}

/* mcts_pure.py:214-231 get_move: n_playout playouts, most-visited child (first max). */
int orc_pure_get_move(orc_mcts *m, const orc_board *b)
{
    int acts[ORC_MAXS], visits[ORC_MAXS];
    int k = orc_mcts_search(m, b, -1, acts, visits, NULL, NULL);
    if (k == 0) return -1;
    int best = 0;
    for (int i = 1; i < k; ++i) if (visits[i] > visits[best]) best = i;
    return acts[best];
}

/* One rollout from a position (for distribution tests against the reference). */
int orc_rollout_once(const orc_board *b, uint64_t seed, uint32_t game_id, uint32_t search_id,
                     uint32_t playout_idx)
{
    orc_mcts m;
    memset(&m, 0, sizeof(m));
    m.salt = seed; m.game_id = game_id; m.search_id = search_id; m.playout_idx = playout_idx;
    orc_board copy = *b;
    return evaluate_rollout(&m, &copy, 1000);
}
