/*
 * gomoku_b200 — C ABI of the B200-native Gomoku self-play engine (libgomoku_b200.so).
 *
 * The reference (AbcEric/AlphaZero_Gomoku) is pure Python and has no FFI layer; its
 * boundary is a duck-typed Python protocol.  These entry points are what a
 * ctypes binding of that protocol needs; each cites the reference interface it replaces
 * (file:line relative to the reference repo).  The Python classes in
 * alphazero_gomoku_b200/ (Board, Game, MCTS, MCTSPlayer, PolicyValueNet, TrainPipeline)
 * are thin ctypes wrappers over exactly these symbols; INTEGRATION.md shows the stub a
 * maintainer of the reference would add.
 *
 * Conventions
 *   - every function returns 0 on success, <0 on error (GZ_E_*); gz_last_error() gives text.
 *   - array arguments are caller-owned, valid only for the call, and may be HOST or DEVICE
 *     pointers (resolved through UVA); engine-owned device memory is never handed out.
 *   - one host thread per engine, one engine per GPU; not re-entrant (like the reference).
 *   - all device work runs on the engine's stream (gz_config.stream, or an owned one).
 *   - moves are reference move ids: move = h * width + w (game.py:70-90); players are 1, 2.
 */
#ifndef GOMOKU_B200_H
#define GOMOKU_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GZ_OK 0
#define GZ_E_INVALID -1   /* bad argument / bad config (reference raises Exception) */
#define GZ_E_ILLEGAL -2   /* illegal move (reference: ValueError from list.remove, game.py:122) */
#define GZ_E_CUDA -3      /* CUDA runtime failure */
#define GZ_E_NOWEIGHTS -4 /* net used before gz_set_weights */
#define GZ_E_POOL -5      /* tree pool exhausted (raise nodes_per_game) */
#define GZ_E_NOGPU -6     /* no CUDA device: there is no CPU fallback */

#define GZ_MAX_SIDE 15
#define GZ_MAX_CELLS 225

#define GZ_LAYOUT_TORCH 0  /* conv OIHW cross-correlation, fc [out,in] (policy_value_net_pytorch.py:31-41) */
#define GZ_LAYOUT_THEANO 1 /* conv OIHW flipped, fc [in,out] (policy_value_net.py:118-121 pickles) */

typedef struct gz_engine gz_engine;

/* Mirrors the constructor arguments of Board (game.py:14-23), MCTS / MCTSPlayer
 * (mcts_alphaZero.py:98-111, :201-203), Net (policy_value_net_pytorch.py:25-41) and the
 * hard-coded self-play constants (Dirichlet 0.1 / 0.25, mcts_alphaZero.py:236-240). */
typedef struct {
    int32_t width, height, n_in_row;
    int32_t n_games;         /* concurrent game slots G resident on this GPU */
    int32_t n_playout;       /* playouts per move (used to size the tree pool) */
    int32_t nodes_per_game;  /* tree pool capacity per game; 0 = 8 * n_playout + 8 (measured high-water mark with subtree
                              * reuse: 5.4 * n_playout over 4096 games; an exhausted pool skips the expansion and
                              * counts a pool_overflow) */
    float c_puct;            /* 5 */
    float dirichlet_alpha;   /* 0.1 */
    float dirichlet_eps;     /* 0.25 */
    int32_t trunk[3];        /* conv widths, reference 32/64/128; 0 = reference.  Wider trunks (every width a
                                multiple of 32, c1, c2 <= 128, c3 <= 256, e.g. 64/128/256) also run on tcgen05 */
    uint64_t seed;           /* Philox key */
    uint64_t game_id_base;   /* global id of slot 0's first game (rank * G in multi-GPU) */
    int32_t device;          /* CUDA ordinal */
    int32_t net_impl;        /* 0 = default (tcgen05 where the trunk is supported, else fp32 CUDA cores),
                                1 = fp32 CUDA-core path, 2 = tcgen05 path or GZ_E_INVALID */
    void *stream;            /* cudaStream_t to run on; NULL = engine creates its own */
} gz_config;

/* ---- lifecycle ---------------------------------------------------------------------- */
int gz_create(const gz_config *cfg, gz_engine **out);
void gz_destroy(gz_engine *e);
const char *gz_last_error(const gz_engine *e); /* e may be NULL (error of gz_create) */
int gz_sync(gz_engine *e);                      /* cudaStreamSynchronize on the engine stream */
int gz_device_count(void);
const char *gz_version(void);

/* ---- weights: PolicyValueNet(model_file=...) / load_state_dict
 *      (policy_value_net_pytorch.py:76-78); 16 arrays in state_dict order -------------- */
int gz_set_weights(gz_engine *e, const float *const ptrs[16], int32_t layout);
int gz_get_weights(gz_engine *e, float *const ptrs[16]); /* torch layout (get_policy_param, :150) */

/* ---- stateless rules over positions given as move lists (Board, game.py:11-172) ------
 * Position i is moves[offsets[i] .. offsets[i+1]) played from an empty board with
 * start_player[i] (0/1, NULL = 0).  Outputs (any may be NULL), all per position:
 *   status  int32  : 0 ok, GZ_E_ILLEGAL if some move was occupied / out of range
 *   ended   int8, winner int8 : Board.game_end() (game.py:162-169); winner 1/2/-1
 *   has_win int8   : Board.has_a_winner()[0] (game.py:129-160)
 *   legal   uint8[S]: 1 where the cell is in Board.availables
 *   planes  float[4*S]: Board.current_state() in the returned (row-reversed) orientation
 *                    (game.py:92-116); square boards only
 *   target  uint8[S]: Board.get_move_target() membership (game.py:35-58)
 *   cur_player int8 : Board.get_current_player()                                         */
int gz_rules_eval(gz_engine *e, const int32_t *moves, const int64_t *offsets, int64_t n_pos,
                  const int8_t *start_player, int32_t *status, int8_t *ended, int8_t *winner,
                  int8_t *has_win, uint8_t *legal, float *planes, uint8_t *target,
                  int8_t *cur_player);

/* ---- engine-resident boards, one per game slot (Board.init_board / do_move / ...) ---- */
int gz_boards_reset(gz_engine *e, const int8_t *start_player /*[G] or NULL*/);
int gz_boards_set(gz_engine *e, int32_t game, const int32_t *moves, int32_t n_moves,
                  int32_t start_player);
int gz_boards_do_move(gz_engine *e, const int32_t *moves /*[G], <0 = leave*/,
                      int32_t *status /*[G] or NULL*/);
int gz_boards_status(gz_engine *e, int8_t *ended, int8_t *winner, int8_t *cur_player,
                     int32_t *n_stones, int32_t *last_move); /* each [G] or NULL */
int gz_boards_legal_mask(gz_engine *e, uint8_t *out /*[G,S]*/);
int gz_boards_planes(gz_engine *e, float *out /*[G,4,W,H]*/);
int gz_boards_cells(gz_engine *e, int8_t *out /*[G,S]: 0 empty, 1, 2*/);

/* ---- leaf evaluator: PolicyValueNet.policy_value (policy_value_net_pytorch.py:80-94) -
 * planes float[B,4,W,H]; probs float[B,S] = exp(log_softmax); values float[B] = tanh.    */
int gz_net_forward(gz_engine *e, const float *planes, int64_t B, float *probs, float *values);

/* gz_net_forward for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.
 * Same arithmetic and outputs as gz_net_forward on the fused tcgen05 path (square board, trunk 32/64/128), but the
 * scratch between its two kernels (the head-activation image, ceil(B / 128) * J * 8192 bytes; 1.5 KB per board at
 * 8x8) is a file-scope device array of 8 MiB that every GPU loading the library holds until exit, so the call
 * makes no cudaMalloc / cudaFree / allocating library call.  Limits: planes, probs, values must be DEVICE pointers;
 * B is rejected (GZ_E_INVALID, gz_last_error names the limit) when its image exceeds 8 MiB - at most 8192 boards
 * at 6x6, 5376 at 8x8, 3840 at 9x9, 1408 at 15x15; not after device-side training until gz_net_forward /
 * gz_get_weights has refreshed the inference weights once.  Stream-ordered on the engine stream WITHOUT a trailing
 * synchronisation: gz_sync() before the outputs are read from elsewhere.  All calls on a device share the array
 * (the part a call uses is cleared first; it keeps the last call's contents): calls that overlap on the GPU
 * (several engines or streams) race and must be serialised by the caller.                                  */
int gz_net_forward_noalloc(gz_engine *e, const float *planes, int64_t B, float *probs, float *values);
/* Diagnostic for parity tests: on != 0 replaces the net inside gz_search / gz_search_vl / gz_selfplay_step by the
 * integer-hash evaluator of oracle/orc_mcts.c:orc_hash_eval computed on the device from the leaf planes (positions
 * with start_player 0), so the fused search paths can be compared with the oracle bit for bit.  Off by default. */
int gz_debug_hash_evaluator(gz_engine *e, int32_t on, uint64_t salt);
/* Diagnostic for parity tests: runs the tensor-core path and also returns the activations after
 * each ReLU: act1 [B,S,32], act2 [B,S,64], act3 [B,S,128] (cell-major, channel-minor) and
 * head [B,6,S] (the two 1x1 head convs, policy_value_net_pytorch.py:49,53).  Any output may be NULL. */
int gz_debug_net_layers(gz_engine *e, const float *planes, int64_t B, float *act1, float *act2,
                        float *act3, float *head, float *probs, float *values);

/* ---- search: MCTS.get_move_probs / _playout (mcts_alphaZero.py:113-181) ---------------
 * gz_search runs n_playout waves; each wave = one playout in every active game
 * (select -> net on the gathered leaves -> expand + backup).                             */
int gz_search(gz_engine *e, int32_t n_playout);
/* gz_search with up to K (1..16) leaves of a game in flight per wave (virtual loss): a single game or a small arena
 * fills a leaf batch K times larger per kernel chain.  The reference's loop (mcts_alphaZero.py:161-163) is strictly
 * sequential; K = 1 reproduces it (and gz_search) exactly, K > 1 picks descents 2..K of a wave without the results of
 * the earlier ones (each counted as a virtual loss while in flight) - deterministic, specified in
 * oracle/orc_mcts.c:orc_mcts_search_vl, but outside the TV <= 0.02 tier (0.057 mean at K = 8, 400 playouts).
 * Every game still gets exactly n_playout playouts; *waves_out (or NULL) = kernel chains launched.          */
int gz_search_vl(gz_engine *e, int32_t n_playout, int32_t K, int32_t *waves_out);
/* Same wave split in two so that a caller-supplied evaluator (the reference's
 * policy_value_fn protocol, mcts_alphaZero.py:98-107) can sit in the middle:
 * select fills leaf_* (each [G]; leaf_state 0 = needs evaluation, 1 = terminal, 2 = game
 * inactive) and the leaf boards; expand_backup consumes priors[G,S] (only legal cells
 * read) and values[G].                                                                   */
int gz_search_select(gz_engine *e, int8_t *leaf_state, float *leaf_planes /*[G,4,W,H]*/,
                     int8_t *leaf_cells /*[G,S] or NULL*/, int8_t *leaf_player /*[G] or NULL*/);
int gz_search_expand_backup(gz_engine *e, const float *priors, const float *values);
/* The descent of the last gz_search_select, i.e. the (action, node) = node.select(c_puct) sequence of
 * one _playout (mcts_alphaZero.py:118-129): path_moves[G,S] holds the chosen move of every level (first
 * path_len[g] entries of row g), path_len[G] the depth.  Lets a test replay every single select decision
 * of the device against the reference's arithmetic, not only compare final visit counts.            */
int gz_search_paths(gz_engine *e, int32_t *path_moves /*[G,S]*/, int32_t *path_len /*[G]*/);
int gz_root_visits(gz_engine *e, int32_t *visits /*[G,S], -1 = not a child*/,
                   double *q /*[G,S] or NULL*/, int32_t *root_n /*[G] or NULL*/);
/* MCTS.update_with_move (mcts_alphaZero.py:183-192): moves[G]; a move that is a root
 * child re-roots (subtree kept), anything else (e.g. -1) resets to a fresh root.
 * Does NOT touch the boards.                                                             */
int gz_update_with_move(gz_engine *e, const int32_t *moves);
/* Which slots take part in the next searches / move choices: mask[G] (0 = sit out), NULL = all.
 * Used by the batched arena (train.py:216-276 policy_evaluate): the two players of a match search
 * on alternate plies, each on the games where it is to move.                                */
int gz_set_active(gz_engine *e, const int8_t *mask);
/* MCTSPlayer.get_action after the search (mcts_alphaZero.py:212-262): visit counts ->
 * softmax(log(v + 1e-10) / temp) -> (self-play) 0.75 p + 0.25 Dir(alpha) -> sampled move.
 * Writes moves[G] and probs[G,S] (noise-free pi); does not apply the move.               */
int gz_root_select_move(gz_engine *e, float temp, int32_t selfplay, int32_t *moves,
                        float *probs);

/* ---- self-play: Game.start_self_play / TrainPipeline.collect_selfplay_data
 *      (game.py:360-416, train.py:137-159) --------------------------------------------- */
typedef struct {
    int64_t n_games;     /* finished games reported */
    int64_t n_samples;   /* plies recorded over those games */
    int64_t n_playouts;  /* playouts executed by this call */
    int64_t n_leaf_evals;/* leaves that went through the net (terminal leaves skip it) */
    int64_t n_plies;     /* moves played by this call (all slots) */
    int64_t pool_overflows; /* expansions dropped because the tree pool was full (must be 0) */
    int64_t n_wins[3];   /* finished games by winner: [tie, player1, player2] */
    int64_t max_tree_nodes; /* high-water mark of node blocks in use by one game (<= nodes_per_game) */
} gz_selfplay_stats;

/* Start (or restart) self-play in every slot.  recycle = 1: a slot that finishes a game
 * immediately starts another with game id = id + G (throughput mode); 2: the same, but the next id is the next
 * game START serial (slots that finish in one ply are numbered in slot order), so ids order the games by when
 * they began - the continuous collect mode of train.py releases finished games in id order; 0: the slot goes idle
 * (collect n = G games).  Game ids (which key the exploration-noise streams) never repeat on an engine: a later
 * begin continues after the largest id handed out.                                                     */
int gz_selfplay_begin(gz_engine *e, float temp, int32_t recycle);
/* Play n_plies moves in every active slot (each = n_playout waves + move choice + record). */
int gz_selfplay_step(gz_engine *e, int32_t n_plies);
/* Number of finished-game samples waiting to be drained / finished games so far. */
int gz_selfplay_pending(gz_engine *e, int64_t *n_samples, int64_t *n_games_finished,
                        int64_t *n_active_slots);
/* Drain finished-game samples (at most max_samples, whole records): states float[n,4,W,H]
 * (current_state orientation), pis float[n,S], zs float[n] (winners_z, game.py:396-399),
 * game_ids int64[n], plies int32[n].  Returns the count in *n_out.                        */
int gz_selfplay_drain(gz_engine *e, int64_t max_samples, float *states, float *pis, float *zs,
                      int64_t *game_ids, int32_t *plies, int64_t *n_out);
/* Finished-game move lists for the reference's {'steps':[...],'winner':w} log line
 * (game.py:411): moves int32[max_moves], offsets int64[max_games+1], winners int8, ids. */
int gz_selfplay_drain_games(gz_engine *e, int64_t max_games, int64_t max_moves, int32_t *moves,
                            int64_t *offsets, int8_t *winners, int64_t *game_ids,
                            int64_t *n_games_out);
int gz_selfplay_get_stats(gz_engine *e, gz_selfplay_stats *out, int32_t reset);
/* Bookkeeping of the running self-play: game_ids[G] (or NULL) = id of the game in progress in every slot (the
 * reference has no ids: they are this build's handle on "which start_self_play call of train.py:146 a sample
 * came from"); *dropped_samples (or NULL) = samples lost because the device ring was full (must stay 0: drain). */
int gz_selfplay_state(gz_engine *e, int64_t *game_ids, int64_t *dropped_samples);
/* What MCTSPlayer.get_action(board, temp, return_prob=1) returned in the last ply of the last
 * gz_selfplay_step: moves int32[G] (-2 = idle slot) and move_probs float[G,S]
 * (mcts_alphaZero.py:257-260).                                                             */
int gz_selfplay_last(gz_engine *e, int32_t *moves, float *probs);

/* TrainPipeline.get_equi_data (train.py:116-135) on the device: the 8 dihedral images of n samples
 * in the reference's order (quarter turns 1..4, each followed by its left-right mirror).
 * states float[n,4,W,H], pis float[n,S], zs float[n] -> out_states float[8n,4,W,H],
 * out_pis float[8n,S], out_zs float[8n].  Square boards; host or device pointers.           */
int gz_equi_data(gz_engine *e, const float *states, const float *pis, const float *zs, int64_t n,
                 float *out_states, float *out_pis, float *out_zs);

/* ---- train step on the device (policy_value_net_pytorch.py:117-148, train.py:161-214) --- */
/* PolicyValueNet.train_step: forward, loss = mean((v - z)^2) - mean(sum(pi * log p, 1)), backward and
 * one torch.optim.Adam(weight_decay=1e-4) step at learning rate lr, all fp32.  states [n,4,W,H]
 * (current_state layout), pis [n,S], zs [n]; host or device pointers.  *loss / *entropy receive
 * loss.item() and -mean(sum(p * log p, 1)) (:139-148), both taken BEFORE the update as in the
 * reference.  The optimiser state (Adam moments, step count) lives in the engine from the first call
 * until gz_set_weights or gz_train_reset; inference entry points see the new weights on their next
 * call.  Needs the reference trunk 32/64/128 (GZ_E_INVALID otherwise).                             */
int gz_train_step(gz_engine *e, const float *states, const float *pis, const float *zs, int64_t n, float lr,
                  float *loss, float *entropy);
/* The two halves of gz_train_step for data-parallel training: backward leaves the flat gradient
 * (state_dict order, torch layout, n_params floats) in the engine; gz_train_grads copies it out (host or
 * device `out`; out = NULL only reports n_params); gz_train_apply does the Adam step with `grads`
 * (e.g. the all-reduced mean) or, if NULL, with the engine's own gradient.                          */
int gz_train_backward(gz_engine *e, const float *states, const float *pis, const float *zs, int64_t n,
                      float *loss, float *entropy);
int gz_train_grads(gz_engine *e, float *out, int64_t capacity, int64_t *n_params);
int gz_train_apply(gz_engine *e, const float *grads, float lr);
/* PolicyValueNet.policy_value (:80-94) with the weights as trained so far, on the fp32 path: the call
 * policy_update makes between epochs for its KL early stop (train.py:178-193).                      */
int gz_train_forward(gz_engine *e, const float *states, int64_t n, float *probs /*[n,S]*/, float *values /*[n]*/);
/* Drops the optimiser state; the next train call restarts Adam from the current weights. */
int gz_train_reset(gz_engine *e);

/* ---- pure MCTS with random rollouts (mcts_pure.py:156-231) ---------------------------- */
/* n_playout waves of: select -> uniform priors over Board.move_target -> Philox rollout
 * to the end of the game -> backup; search_id salts the Philox counter.                   */
int gz_pure_search(gz_engine *e, int32_t n_playout, uint32_t search_id);
/* MCTS.get_move (mcts_pure.py:230): most visited root child, first maximum. */
int gz_pure_best_move(gz_engine *e, int32_t *moves /*[G]*/);
/* Rollout kernel alone (_evaluate_rollout, mcts_pure.py:193-212) from the resident
 * boards: n_rollouts per game, results int8[G, n_rollouts] in {-1,0,+1} from the side to
 * move's view; rollout r of game g uses Philox counter (r, ply, game_id(g), search_id). */
int gz_rollouts(gz_engine *e, int32_t n_rollouts, uint32_t search_id, int8_t *results,
                int64_t *total_plies);

/* ---- counters for bench / roofline accounting ---------------------------------------- */
typedef struct {
    int64_t kernel_launches;  /* kernels of this library launched since the last reset */
    int64_t select_levels;    /* tree levels descended (sum over playouts) */
    int64_t select_children;  /* child slots scored (sum of A_l over levels) */
    int64_t playouts;
    int64_t leaf_evals;
    int64_t expanded_children;/* sum of A_leaf over expansions */
    int64_t rollout_plies;
} gz_counters;
int gz_get_counters(gz_engine *e, gz_counters *out, int32_t reset);


/* ---- per-kernel-class device timing (CUDA events on the engine stream) for bench.py ----- */
#define GZ_PROF_CLASSES 14
/* on = 0 off, 1 time every launch, N > 1 time every N-th launch of each class (totals are scaled) */
int gz_profile_enable(gz_engine *e, int32_t on);
/* ms[c] = summed device time, launches[c] = timed launches of class c since the last reset */
int gz_profile_read(gz_engine *e, double *ms /*[GZ_PROF_CLASSES]*/, int64_t *launches, int32_t reset);
const char *gz_profile_class_name(int32_t c);

#ifdef __cplusplus
}
#endif
#endif /* GOMOKU_B200_H */
