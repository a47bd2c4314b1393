// gz_engine.h — internal engine state (host struct owning the device arrays).
#pragma once
#include "../../include/gomoku_b200.h"
#include "gz_common.cuh"
#include <string>
#include <unordered_map>
#include <vector>

// Tree node block (one per expanded node, SoA inside the block, Sp = S rounded up to 32):
//   double Q[Sp] | float P[Sp] | uint32 N[Sp] | int32 child[Sp]      = 20 * Sp bytes
// P[i] < 0 marks "cell i is not a child".  AlphaZero mode: P = prior (float32, as the
// reference's np.float32 prob).  Pure mode: P = number of children k (prior = 1/k in
// float64, mcts_pure.py:45).  child[i] = node index of the expanded child or -1.
struct NodeView {
    double *Q;
    float *P;
    uint32_t *N;
    int32_t *child;
};
__host__ __device__ __forceinline__ size_t gz_node_bytes(int Sp) { return (size_t)20 * Sp; }
__host__ __device__ __forceinline__ NodeView gz_node_view(uint8_t *pool, size_t node_index, int Sp)
{
    uint8_t *b = pool + node_index * gz_node_bytes(Sp);
    NodeView v;
    v.Q = reinterpret_cast<double *>(b);
    v.P = reinterpret_cast<float *>(b + (size_t)8 * Sp);
    v.N = reinterpret_cast<uint32_t *>(b + (size_t)12 * Sp);
    v.child = reinterpret_cast<int32_t *>(b + (size_t)16 * Sp);
    return v;
}

// device-side counters (one struct in device memory, updated with warp-aggregated atomics)
struct DevCounters {
    unsigned long long select_levels, select_children, playouts, leaf_evals, expanded_children,
        rollout_plies, pool_overflows, plies, games_finished, samples_out, dropped, wins[3], max_nodes, empty_box;
};

struct NetWeights {
    // device copies, torch layout fp32 (for gz_get_weights) and kernel layouts
    float *raw[16] = {};
    size_t raw_elems[16] = {};
    float *w1 = nullptr;   // conv1 [tap][ci(4)][c1]
    float *w2 = nullptr;   // conv2 [tap][ci][c2]
    float *w3 = nullptr;   // conv3 [tap][ci][c3]
    float *wh = nullptr;   // heads 1x1 [c3][8] (cols 0..3 policy, 4..5 value, 6..7 zero)
    float *wpfc = nullptr; // policy fc transposed [4S][S]
    uint8_t *tc_fcimg = nullptr; // fp16 hi/lo image of both FC-head weight matrices (tcgen05 path)
    float *wvfc = nullptr; // value fc1 transposed [2S][64]
    float *bias = nullptr; // packed: b1 | b2 | b3 | bh[8] | bp[S] | bv[64] | wv2[64] | bv2[1]
    int off_b1 = 0, off_b2 = 0, off_b3 = 0, off_bh = 0, off_bp = 0, off_bv = 0, off_wv2 = 0, off_bv2 = 0;
    std::vector<float> tc_consts; // host copy of the tcgen05 trunk's epilogue constants (kernel parameter)
    uint8_t *tc_img = nullptr; // tcgen05 path: fp16 hi/lo conv weight image (gz_net_tc.cu layout)
    uint8_t *tc_imgp = nullptr; // the same weights as two per-rank half images for the CTA-pair (cta_group::2) trunk
    float tc_inv[5] = {1.f, 1.f, 1.f, 1.f, 1.f}; // 1 / power-of-two weight scale of conv1..3, policy FC, value FC1
    uint8_t *tcw_img1 = nullptr, *tcw_img2 = nullptr, *tcw_img3 = nullptr; // wide-trunk tcgen05 path: conv weight images (gz_net_tcw.cu)
    float tcw_inv[3] = {1.f, 1.f, 1.f}; // 1 / weight scale of conv2, conv3, conv1
    bool loaded = false;
};

struct TrainState; // gz_train.cu
#define GZ_PURE_MAXPARTS 4

struct gz_engine {
    gz_config cfg;
    int W, H, S, Sp, n_in_row, G, cap;
    int c1, c2, c3;
    cudaStream_t stream;
    bool own_stream;
    // pure-MCTS search split over parts of the games (gz_rollout.cu), created on first use: per part a stream for the
    // wave kernels, an urgent one for the rollout kernels, and events (wave done, rollout done, first wave done, part done)
    cudaStream_t pure_st[GZ_PURE_MAXPARTS], pure_hi[GZ_PURE_MAXPARTS];
    cudaEvent_t pure_ev[GZ_PURE_MAXPARTS][4], ev_fork;
    std::string err;
    long long launches;

    // boards
    GameBoard *boards;
    GameMeta *meta;
    int8_t *active; // [G] 1 = slot participates in waves
    // tree
    uint8_t *pool;
    int32_t *free_stack; // [G][cap]
    int32_t *free_top;   // [G]
    int32_t *root_node;  // [G] -1 = root is an unexpanded leaf
    uint32_t *root_N;    // [G]
    double *root_Q;      // [G]
    uint32_t *path;      // [G][S] edge ids node*Sp + cell
    int32_t *path_len;   // [G]
    // leaves of the current wave
    int8_t *leaf_state;  // [G] 0 eval, 1 terminal, 2 inactive
    double *leaf_value;  // [G] terminal value (side to move at the leaf)
    int32_t *leaf_slot;  // [G] row in the compacted leaf batch
    int32_t *leaf_count; // the counter of the current wave: one of leaf_count_pair[2] (gzk_expand_select)
    int32_t *leaf_count_pair;
    float *leaf_planes;  // [G][4][S] compacted rows
    GameBoard *leaf_board; // [G] leaf position (by game)
    int32_t *slot_game = nullptr; // [G] game of a leaf-batch slot: the fused tcgen05 trunk stages a leaf straight from
                                  // leaf_board / leaf_meta of that game instead of from fp32 planes (leaf_packed)
    bool leaf_packed = false;     // set for the duration of gz_waves / gz_search_vl when that path is taken
    const int32_t *pk_slot = nullptr;    // ... and then: slot -> leaf record, the records' boards and meta
    const GameBoard *pk_board = nullptr; //     (gz_waves: slot_game / leaf_board / leaf_meta, by game;
    const GameMeta *pk_meta = nullptr;   //      gz_search_vl: vl_slot_leaf / vl_leaf_board / vl_leaf_meta, by game * K + k)
    GameMeta *leaf_meta;   // [G]
    float *net_probs;    // [G][S] compacted rows
    float *net_values;   // [G]
    // net
    NetWeights wts;
    float *act1, *act2, *act3; // NHWC activations for the fp32 path
    int64_t act_cap;           // boards the activation buffers hold
    float *tc_head;            // tcgen05 path: ReLU(1x1 head convs), fp16 hi/lo image (see trunk_tc_kernel)
    float *tc_dbg;             // optional per-layer dump for parity tests
    int64_t tc_cap, tc_dbg_cap;
    uint8_t *tcw_a0, *tcw_act1, *tcw_act2; // wide-trunk path: fp16 hi/lo images of the planes and of conv1 / conv2's outputs (per group)
    float *tcw_dbg;
    int64_t tcw_cap, tcw_dbg_cap; // groups / boards
    int n_sm;
    unsigned attr_done;        // bit i: cudaFuncSetAttribute done for kernel i
    // staging for host<->device argument copies
    void *stage;
    size_t stage_bytes;
    std::unordered_map<void *, size_t> upload_bytes; // sizes of the arrays gz_upload allocated (reused on equal size)
    // self-play
    float sp_temp;
    int sp_recycle;
    bool sp_begun;
    int32_t *sp_moves;   // [G][S] moves of the game in progress
    float *sp_pi;        // [G][S][S] pi per ply of the game in progress
    int64_t *sp_game_id; // [G]
    int32_t *sp_fin_flag;           // [G] recycle mode 2: finished this ply, next id pending
    unsigned long long *sp_serial;  // [1] recycle mode 2: next game serial on the device
    int64_t sp_next_serial; // next unused game serial (ids = game_id_base + serial): never reset
    bool sp_ids_live;       // sp_game_id holds ids of an earlier begin
    int32_t *sel_moves;  // [G] scratch: chosen moves
    float *sel_probs;    // [G][S] scratch: pi
    // finished-game output ring (device), drained by the host
    int64_t out_cap;     // samples
    float *out_states;   // [out_cap][4][S]
    float *out_pi;       // [out_cap][S]
    float *out_z;        // [out_cap]
    int64_t *out_gid;    // [out_cap]
    int32_t *out_ply;    // [out_cap]
    unsigned long long *out_count; // [1]
    int64_t out_drained;
    int64_t gout_cap;    // finished games log
    int32_t *gout_moves; // [gout_cap][S]
    int32_t *gout_len;   // [gout_cap]
    int8_t *gout_winner; // [gout_cap]
    int64_t *gout_gid;   // [gout_cap]
    unsigned long long *gout_count; // [1]
    int64_t gout_drained;
    bool dbg_hash_on;       // diagnostic: the oracle's integer-hash evaluator instead of the net (gz_debug_hash_evaluator)
    uint64_t dbg_hash_salt;
    // several leaves in flight per game (gz_vl.cu), allocated on first use for K = vl_K
    int vl_K;
    uint32_t *vl_path;
    int32_t *vl_path_len, *vl_leaf_slot, *vl_keff, *vl_remaining, *vl_count;
    int8_t *vl_leaf_state;
    double *vl_leaf_value;
    GameBoard *vl_leaf_board;
    GameMeta *vl_leaf_meta = nullptr;  // [G*K]
    int32_t *vl_slot_leaf = nullptr;   // [G*K] leaf record of a batch slot
    float *vl_planes, *vl_probs, *vl_values;
    DevCounters *counters;
    TrainState *tr;      // device train step state (master weights, Adam moments), created on first use
    // rollouts
    int8_t *roll_results;
    int64_t roll_cap;
    // per-kernel-class timing
    bool prof_on;
    int prof_stride;                     // time every prof_stride-th launch of a class (events cost GPU idle time)
    long long prof_seen[GZ_PROF_CLASSES]; // launches of the class since the last read, timed or not
    std::vector<cudaEvent_t> prof_ev[GZ_PROF_CLASSES]; // start/stop pairs
    size_t prof_used[GZ_PROF_CLASSES];
    double prof_ms[GZ_PROF_CLASSES];
    long long prof_n[GZ_PROF_CLASSES];
};

enum { PC_SELECT = 0, PC_CONV1, PC_CONV2, PC_CONV3, PC_HEADS, PC_EXPAND, PC_MOVE, PC_ADVANCE,
       PC_PURE_EXPAND, PC_ROLLOUT, PC_BACKUP, PC_OTHER, PC_TRUNK_TC, PC_HEADS_FC };

// RAII: records a start event now and a stop event at scope exit (only when profiling is on)
struct ProfScope {
    gz_engine *e;
    int c;
    bool live;
    ProfScope(gz_engine *eng, int cls) : e(eng), c(cls), live(false)
    {
        if (!e->prof_on) return;
        if ((e->prof_seen[c]++ % e->prof_stride) != 0) return;
        size_t u = e->prof_used[c];
        if (u + 2 > e->prof_ev[c].size()) {
            if (e->prof_ev[c].size() >= 400000) return;
            for (int i = 0; i < 1024; ++i) {
                cudaEvent_t ev;
                if (cudaEventCreate(&ev) != cudaSuccess) return;
                e->prof_ev[c].push_back(ev);
            }
        }
        cudaEventRecord(e->prof_ev[c][u], e->stream);
        live = true;
    }
    ~ProfScope()
    {
        if (!live) return;
        cudaEventRecord(e->prof_ev[c][e->prof_used[c] + 1], e->stream);
        e->prof_used[c] += 2;
    }
};

#define GZ_CUDA(e, call)                                                                     \
    do {                                                                                     \
        cudaError_t _err = (call);                                                           \
        if (_err != cudaSuccess) {                                                           \
            (e)->err = std::string(#call) + ": " + cudaGetErrorString(_err);                 \
            return GZ_E_CUDA;                                                                \
        }                                                                                    \
    } while (0)

#define GZ_TRY(x) do { int _rc = (x); if (_rc != GZ_OK) return _rc; } while (0)

// kernels' launchers (each returns a GZ_* code)
int gzk_rules_eval(gz_engine *e, const int32_t *d_moves, const int64_t *d_offsets, int64_t n_pos,
                   const int8_t *d_start, int32_t *d_status, int8_t *d_ended, int8_t *d_winner,
                   int8_t *d_haswin, uint8_t *d_legal, float *d_planes, uint8_t *d_target,
                   int8_t *d_cur);
int gzk_boards_reset(gz_engine *e, const int8_t *d_start);
int gzk_boards_do_move(gz_engine *e, const int32_t *d_moves, int32_t *d_status);
int gzk_boards_query(gz_engine *e, int8_t *d_ended, int8_t *d_winner, int8_t *d_cur,
                     int32_t *d_nstones, int32_t *d_last, uint8_t *d_legal, float *d_planes,
                     int8_t *d_cells);
int gzk_tree_reset(gz_engine *e);
int gzk_select(gz_engine *e, bool pure);
int gzk_expand_backup(gz_engine *e, const float *d_priors, const float *d_values, bool compacted);
int gzk_expand_select(gz_engine *e, const float *d_priors, const float *d_values);
int gzk_pure_playouts(gz_engine *e, int n, uint32_t search_id);
int gzk_root_visits(gz_engine *e, int32_t *d_visits, double *d_q, int32_t *d_rootn);
int gzk_update_with_move(gz_engine *e, const int32_t *d_moves);
int gzk_root_select_move(gz_engine *e, float temp, int selfplay, int32_t *d_moves, float *d_probs,
                         bool best_visits);
int gzk_selfplay_advance(gz_engine *e);
int gzk_net_forward(gz_engine *e, const float *d_planes, const int32_t *d_count, int64_t maxB,
                    float *d_probs, float *d_values);
int gzk_rollouts(gz_engine *e, int n_rollouts, uint32_t search_id, int8_t *d_results);
int gz_net_ensure_buffers(gz_engine *e, int64_t B);
