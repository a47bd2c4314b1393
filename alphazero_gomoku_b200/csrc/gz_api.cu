// gz_api.cu — the C ABI (include/gomoku_b200.h) over the kernels.  Single translation unit:
// the kernel files are included here so device helpers are shared without relocatable code.
#include "gz_engine.h"
#include <cmath>
#include <cstdio>
#include <algorithm>
#include <cstring>
#include <new>

#include "gz_rules.cu"
#include "gz_tree.cu"
#include "gz_rollout.cu"
#include "gz_move.cu"
#include "gz_net_ffma.cu"

static int gz_train_flush(gz_engine *e); // gz_train.cu
static void gz_train_free(gz_engine *e);
static void gz_train_invalidate(gz_engine *e);

static std::string g_create_error;

// ---- host/device argument marshalling -----------------------------------------------------
static bool gz_is_device_ptr(const void *p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

struct ArgStage {
    gz_engine *e;
    size_t off = 0;
    struct Back { void *host; void *dev; size_t bytes; };
    std::vector<Back> backs;
    int rc = GZ_OK;
    explicit ArgStage(gz_engine *eng) : e(eng) {}
    // must be called once, before in()/out(), with an upper bound of all host-side bytes
    int reserve(size_t bytes)
    {
        bytes += 4096;
        if (bytes <= e->stage_bytes) return GZ_OK;
        if (e->stage) {
            cudaStreamSynchronize(e->stream);
            cudaFree(e->stage);
            e->stage = nullptr;
            e->stage_bytes = 0;
        }
        size_t want = bytes + bytes / 4;
        GZ_CUDA(e, cudaMalloc(&e->stage, want));
        e->stage_bytes = want;
        return GZ_OK;
    }
    void *carve(size_t bytes)
    {
        size_t o = (off + 255) & ~(size_t)255;
        if (o + bytes > e->stage_bytes) { rc = GZ_E_INVALID; e->err = "internal: stage overflow"; return nullptr; }
        off = o + bytes;
        return (char *)e->stage + o;
    }
    template <typename T> const T *in(const T *p, size_t n)
    {
        if (!p || n == 0) return p;
        if (gz_is_device_ptr(p)) return p;
        void *d = carve(n * sizeof(T));
        if (!d) return nullptr;
        if (cudaMemcpyAsync(d, p, n * sizeof(T), cudaMemcpyHostToDevice, e->stream) != cudaSuccess) {
            rc = GZ_E_CUDA; e->err = "H2D copy failed";
        }
        return (const T *)d;
    }
    template <typename T> T *out(T *p, size_t n)
    {
        if (!p || n == 0) return p;
        if (gz_is_device_ptr(p)) return p;
        void *d = carve(n * sizeof(T));
        if (!d) return nullptr;
        backs.push_back({(void *)p, d, n * sizeof(T)});
        return (T *)d;
    }
    int finish(bool sync_always = true)
    {
        for (auto &b : backs)
            if (cudaMemcpyAsync(b.host, b.dev, b.bytes, cudaMemcpyDeviceToHost, e->stream) != cudaSuccess) {
                rc = GZ_E_CUDA; e->err = "D2H copy failed";
            }
        if (!backs.empty() || sync_always) {
            cudaError_t er = cudaStreamSynchronize(e->stream);
            if (er != cudaSuccess) { rc = GZ_E_CUDA; e->err = std::string("stream sync: ") + cudaGetErrorString(er); }
        }
        return rc;
    }
};

template <typename T> static int gz_dalloc(gz_engine *e, T **p, size_t n)
{
    GZ_CUDA(e, cudaMalloc((void **)p, (n ? n : 1) * sizeof(T)));
    GZ_CUDA(e, cudaMemsetAsync(*p, 0, (n ? n : 1) * sizeof(T), e->stream));
    return GZ_OK;
}

// ---- lifecycle ------------------------------------------------------------------------------
extern "C" const char *gz_version(void) { return "gomoku_b200 0.1 (sm_100a)"; }

extern "C" int gz_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" const char *gz_last_error(const gz_engine *e) { return e ? e->err.c_str() : g_create_error.c_str(); }

static int gz_alloc_all(gz_engine *e)
{
    const size_t G = e->G, S = e->S;
    GZ_TRY(gz_dalloc(e, &e->boards, G));
    GZ_TRY(gz_dalloc(e, &e->meta, G));
    GZ_TRY(gz_dalloc(e, &e->active, G));
    GZ_CUDA(e, cudaMemsetAsync(e->active, 1, G, e->stream));
    GZ_CUDA(e, cudaMalloc((void **)&e->pool, G * (size_t)e->cap * gz_node_bytes(e->Sp)));
    GZ_TRY(gz_dalloc(e, &e->free_stack, G * (size_t)e->cap));
    GZ_TRY(gz_dalloc(e, &e->free_top, G));
    GZ_TRY(gz_dalloc(e, &e->root_node, G));
    GZ_TRY(gz_dalloc(e, &e->root_N, G));
    GZ_TRY(gz_dalloc(e, &e->root_Q, G));
    GZ_TRY(gz_dalloc(e, &e->path, G * S));
    GZ_TRY(gz_dalloc(e, &e->path_len, G));
    GZ_TRY(gz_dalloc(e, &e->leaf_state, G));
    GZ_TRY(gz_dalloc(e, &e->leaf_value, G));
    GZ_TRY(gz_dalloc(e, &e->leaf_slot, G));
    GZ_TRY(gz_dalloc(e, &e->leaf_count_pair, 2));
    e->leaf_count = e->leaf_count_pair;
    GZ_TRY(gz_dalloc(e, &e->leaf_planes, G * 4 * S));
    GZ_TRY(gz_dalloc(e, &e->leaf_board, G));
    GZ_TRY(gz_dalloc(e, &e->slot_game, G));
    GZ_TRY(gz_dalloc(e, &e->leaf_meta, G));
    GZ_TRY(gz_dalloc(e, &e->net_probs, G * S));
    GZ_TRY(gz_dalloc(e, &e->net_values, G));
    GZ_TRY(gz_dalloc(e, &e->sel_moves, G));
    GZ_TRY(gz_dalloc(e, &e->sel_probs, G * S));
    GZ_TRY(gz_dalloc(e, &e->counters, 1));
    GZ_TRY(gz_dalloc(e, &e->out_count, 1));
    GZ_TRY(gz_dalloc(e, &e->gout_count, 1));
    GZ_TRY(gzk_boards_reset(e, nullptr));
    GZ_TRY(gzk_tree_reset(e));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    return GZ_OK;
}

extern "C" int gz_create(const gz_config *cfg, gz_engine **out)
{
    if (!cfg || !out) { g_create_error = "null argument"; return GZ_E_INVALID; }
    *out = nullptr;
    if (gz_device_count() <= 0) {
        g_create_error = "no CUDA device visible: gomoku_b200 has no CPU fallback";
        return GZ_E_NOGPU;
    }
    const int W = cfg->width, H = cfg->height, n = cfg->n_in_row;
    if (W < 1 || H < 1 || W > GZ_MAX_SIDE || H > GZ_MAX_SIDE) { g_create_error = "board side must be in 1..15"; return GZ_E_INVALID; }
    if (W < n || H < n) { // game.py:61-62
        g_create_error = "board width and height can not be less than n_in_row";
        return GZ_E_INVALID;
    }
    if (n < 1 || cfg->n_games < 1) { g_create_error = "n_in_row and n_games must be positive"; return GZ_E_INVALID; }
    gz_engine *e = new (std::nothrow) gz_engine();
    if (!e) { g_create_error = "out of host memory"; return GZ_E_INVALID; }
    e->cfg = *cfg;
    if (e->cfg.c_puct <= 0.f) e->cfg.c_puct = 5.f;
    if (e->cfg.dirichlet_alpha <= 0.f) e->cfg.dirichlet_alpha = 0.1f;
    if (e->cfg.dirichlet_eps < 0.f) e->cfg.dirichlet_eps = 0.25f;
    if (e->cfg.n_playout <= 0) e->cfg.n_playout = 400;
    e->W = W; e->H = H; e->S = W * H; e->Sp = (e->S + 31) / 32 * 32; e->n_in_row = n; e->G = cfg->n_games;
    e->cap = cfg->nodes_per_game > 0 ? cfg->nodes_per_game : 8 * e->cfg.n_playout + 8;
    e->c1 = cfg->trunk[0] > 0 ? cfg->trunk[0] : 32;
    e->c2 = cfg->trunk[1] > 0 ? cfg->trunk[1] : 64;
    e->c3 = cfg->trunk[2] > 0 ? cfg->trunk[2] : 128;
    if (e->c1 % 8 || e->c2 % 64 || e->c3 % 64) {
        g_create_error = "trunk widths: c1 multiple of 8, c2 and c3 multiples of 64";
        delete e;
        return GZ_E_INVALID;
    }
    if (cudaSetDevice(cfg->device) != cudaSuccess) {
        g_create_error = std::string("cudaSetDevice: ") + cudaGetErrorString(cudaGetLastError());
        delete e;
        return GZ_E_CUDA;
    }
    if (cudaDeviceGetAttribute(&e->n_sm, cudaDevAttrMultiProcessorCount, cfg->device) != cudaSuccess) e->n_sm = 0;
    if (gz_sqrt_tab_init(cfg->device) != 0) {
        g_create_error = std::string("cudaMemcpyToSymbol(gz_sqrt_tab): ") + cudaGetErrorString(cudaGetLastError());
        delete e;
        return GZ_E_CUDA;
    }
    if (cfg->stream) { e->stream = (cudaStream_t)cfg->stream; e->own_stream = false; }
    else {
        if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess) {
            g_create_error = "cudaStreamCreate failed";
            delete e;
            return GZ_E_CUDA;
        }
        e->own_stream = true;
    }
    int rc = gz_alloc_all(e);
    if (rc != GZ_OK) {
        g_create_error = e->err;
        gz_destroy(e);
        return rc;
    }
    *out = e;
    return GZ_OK;
}

extern "C" void gz_destroy(gz_engine *e)
{
    if (!e) return;
    cudaSetDevice(e->cfg.device);
    cudaStreamSynchronize(e->stream);
    void *ptrs[] = {e->boards, e->meta, e->active, e->pool, e->free_stack, e->free_top, e->root_node,
                    e->root_N, e->root_Q, e->path, e->path_len, e->leaf_state, e->leaf_value, e->leaf_slot,
                    e->leaf_count_pair, e->leaf_planes, e->leaf_board, e->slot_game, e->leaf_meta, e->net_probs, e->net_values,
                    e->act1, e->act2, e->act3, e->stage, e->sp_moves, e->sp_pi, e->sp_game_id, e->sp_fin_flag, e->sp_serial, e->sel_moves,
                    e->sel_probs, e->out_states, e->out_pi, e->out_z, e->out_gid, e->out_ply, e->out_count,
                    e->gout_moves, e->gout_len, e->gout_winner, e->gout_gid, e->gout_count, e->counters,
                    e->roll_results, e->wts.w1, e->wts.w2, e->wts.w3, e->wts.wh, e->wts.wpfc, e->wts.wvfc,
                    e->wts.bias, e->wts.tc_img, e->wts.tc_imgp, e->wts.tc_fcimg, e->tc_head, e->tc_dbg,
                    e->wts.tcw_img1, e->wts.tcw_img2, e->wts.tcw_img3, e->tcw_a0, e->tcw_act1, e->tcw_act2, e->tcw_dbg,
                    e->vl_path, e->vl_path_len, e->vl_leaf_state, e->vl_leaf_value, e->vl_leaf_slot, e->vl_leaf_board,
                    e->vl_keff, e->vl_remaining, e->vl_planes, e->vl_probs, e->vl_values, e->vl_count,
                    e->vl_leaf_meta, e->vl_slot_leaf};
    for (void *p : ptrs) if (p) cudaFree(p);
    gz_train_free(e);
    for (int i = 0; i < 16; ++i) if (e->wts.raw[i]) cudaFree(e->wts.raw[i]);
    for (int c = 0; c < GZ_PROF_CLASSES; ++c) for (cudaEvent_t ev : e->prof_ev[c]) cudaEventDestroy(ev);
    if (e->pure_st[0]) {
        for (int i = 0; i < GZ_PURE_MAXPARTS; ++i) {
            if (e->pure_st[i]) cudaStreamDestroy(e->pure_st[i]);
            if (e->pure_hi[i]) cudaStreamDestroy(e->pure_hi[i]);
            for (int k = 0; k < 4; ++k) if (e->pure_ev[i][k]) cudaEventDestroy(e->pure_ev[i][k]);
        }
        if (e->ev_fork) cudaEventDestroy(e->ev_fork);
    }
    if (e->own_stream) cudaStreamDestroy(e->stream);
    delete e;
}

extern "C" int gz_sync(gz_engine *e)
{
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    return GZ_OK;
}

// ---- weights --------------------------------------------------------------------------------
static void gz_weight_sizes(const gz_engine *e, size_t n[16])
{
    const size_t S = e->S, c1 = e->c1, c2 = e->c2, c3 = e->c3;
    n[0] = c1 * 4 * 9; n[1] = c1; n[2] = c2 * c1 * 9; n[3] = c2; n[4] = c3 * c2 * 9; n[5] = c3;
    n[6] = 4 * c3; n[7] = 4; n[8] = S * 4 * S; n[9] = S; n[10] = 2 * c3; n[11] = 2;
    n[12] = 64 * 2 * S; n[13] = 64; n[14] = 64; n[15] = 1;
}

// Device copy of a host array.  A weight update (gz_set_weights every step of the bench's e2e region,
// once per policy update in train.py) re-uploads arrays of unchanged size: the allocation is reused then,
// which keeps cudaFree (a device-wide synchronisation) and cudaMalloc off that path.
template <typename T> static int gz_upload(gz_engine *e, T **dst, const std::vector<T> &src)
{
    const size_t bytes = src.size() * sizeof(T);
    if (*dst) {
        auto it = e->upload_bytes.find((void *)*dst);
        if (it == e->upload_bytes.end() || it->second != bytes) {
            if (it != e->upload_bytes.end()) e->upload_bytes.erase(it);
            cudaFree(*dst);
            *dst = nullptr;
        }
    }
    if (!*dst) {
        GZ_CUDA(e, cudaMalloc((void **)dst, bytes ? bytes : 1));
        e->upload_bytes[(void *)*dst] = bytes;
    }
    GZ_CUDA(e, cudaMemcpy(*dst, src.data(), bytes, cudaMemcpyHostToDevice));
    return GZ_OK;
}

#include "gz_net_tc.cu"
#include "gz_net_tcw.cu"


static void gz_conv_to_tap_major(const std::vector<float> &w, int cout, int cin, std::vector<float> &o)
{
    o.assign((size_t)9 * cin * cout, 0.f);
    for (int co = 0; co < cout; ++co)
        for (int ci = 0; ci < cin; ++ci)
            for (int t = 0; t < 9; ++t)
                o[((size_t)t * cin + ci) * cout + co] = w[((size_t)co * cin + ci) * 9 + t];
}

int gz_net_prepare_weights(gz_engine *e, std::vector<float> h[16])
{
    const int S = e->S, c1 = e->c1, c2 = e->c2, c3 = e->c3;
    std::vector<float> t;
    gz_conv_to_tap_major(h[0], c1, 4, t);  GZ_TRY(gz_upload(e, &e->wts.w1, t));
    gz_conv_to_tap_major(h[2], c2, c1, t); GZ_TRY(gz_upload(e, &e->wts.w2, t));
    gz_conv_to_tap_major(h[4], c3, c2, t); GZ_TRY(gz_upload(e, &e->wts.w3, t));
    t.assign((size_t)c3 * 8, 0.f);
    for (int c = 0; c < c3; ++c) {
        for (int j = 0; j < 4; ++j) t[(size_t)c * 8 + j] = h[6][(size_t)j * c3 + c];
        for (int j = 0; j < 2; ++j) t[(size_t)c * 8 + 4 + j] = h[10][(size_t)j * c3 + c];
    }
    GZ_TRY(gz_upload(e, &e->wts.wh, t));
    t.assign((size_t)4 * S * S, 0.f);
    for (int o = 0; o < S; ++o)
        for (int k = 0; k < 4 * S; ++k) t[(size_t)k * S + o] = h[8][(size_t)o * 4 * S + k];
    GZ_TRY(gz_upload(e, &e->wts.wpfc, t));
    t.assign((size_t)2 * S * 64, 0.f);
    for (int o = 0; o < 64; ++o)
        for (int k = 0; k < 2 * S; ++k) t[(size_t)k * 64 + o] = h[12][(size_t)o * 2 * S + k];
    GZ_TRY(gz_upload(e, &e->wts.wvfc, t));
    // packed biases, each segment 4-float aligned
    NetWeights &w = e->wts;
    auto al = [](int x) { return (x + 3) & ~3; };
    int off = 0;
    w.off_b1 = off; off = al(off + c1);
    w.off_b2 = off; off = al(off + c2);
    w.off_b3 = off; off = al(off + c3);
    w.off_bh = off; off = al(off + 8);
    w.off_bp = off; off = al(off + S);
    w.off_bv = off; off = al(off + 64);
    w.off_wv2 = off; off = al(off + 64);
    w.off_bv2 = off; off = al(off + 1);
    t.assign(off, 0.f);
    std::copy(h[1].begin(), h[1].end(), t.begin() + w.off_b1);
    std::copy(h[3].begin(), h[3].end(), t.begin() + w.off_b2);
    std::copy(h[5].begin(), h[5].end(), t.begin() + w.off_b3);
    for (int j = 0; j < 4; ++j) t[w.off_bh + j] = h[7][j];
    for (int j = 0; j < 2; ++j) t[w.off_bh + 4 + j] = h[11][j];
    std::copy(h[9].begin(), h[9].end(), t.begin() + w.off_bp);
    std::copy(h[13].begin(), h[13].end(), t.begin() + w.off_bv);
    std::copy(h[14].begin(), h[14].end(), t.begin() + w.off_wv2);
    t[w.off_bv2] = h[15][0];
    GZ_TRY(gz_upload(e, &e->wts.bias, t));
    if (gz_tc_supported(e)) GZ_TRY(gz_tc_prepare_weights(e, h));
    else if (gz_tcw_supported(e)) GZ_TRY(gz_tcw_prepare_weights(e, h));
    return GZ_OK;
}

extern "C" int gz_set_weights(gz_engine *e, const float *const ptrs[16], int32_t layout)
{
    if (!e || !ptrs) return GZ_E_INVALID;
    if (layout != GZ_LAYOUT_TORCH && layout != GZ_LAYOUT_THEANO) { e->err = "unknown weight layout"; return GZ_E_INVALID; }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    size_t n[16];
    gz_weight_sizes(e, n);
    std::vector<float> h[16];
    for (int i = 0; i < 16; ++i) {
        if (!ptrs[i]) { e->err = "null weight array"; return GZ_E_INVALID; }
        h[i].resize(n[i]);
        GZ_CUDA(e, cudaMemcpy(h[i].data(), ptrs[i], n[i] * sizeof(float), cudaMemcpyDefault));
    }
    if (layout == GZ_LAYOUT_THEANO) {
        // conv filters are stored for a true convolution (flip both
        // spatial axes), FC weights are [in, out] (transpose).
        const int conv_idx[3] = {0, 2, 4};
        for (int q = 0; q < 3; ++q) {
            std::vector<float> &w = h[conv_idx[q]];
            for (size_t f = 0; f < w.size() / 9; ++f)
                for (int t = 0; t < 4; ++t) std::swap(w[f * 9 + t], w[f * 9 + 8 - t]);
        }
        const int fc_idx[3] = {8, 12, 14};
        const size_t fc_out[3] = {(size_t)e->S, 64, 1};
        for (int q = 0; q < 3; ++q) {
            std::vector<float> &w = h[fc_idx[q]];
            const size_t no = fc_out[q], ni = w.size() / no;
            std::vector<float> t(w.size());
            for (size_t i = 0; i < ni; ++i)
                for (size_t o = 0; o < no; ++o) t[o * ni + i] = w[i * no + o];
            w.swap(t);
        }
    }
    for (int i = 0; i < 16; ++i) {
        GZ_TRY(gz_upload(e, &e->wts.raw[i], h[i]));
        e->wts.raw_elems[i] = n[i];
    }
    GZ_TRY(gz_net_prepare_weights(e, h));
    e->wts.loaded = true;
    gz_train_invalidate(e); // new weights from outside: the optimiser state no longer describes them
    return GZ_OK;
}

extern "C" int gz_get_weights(gz_engine *e, float *const ptrs[16])
{
    if (!e || !ptrs) return GZ_E_INVALID;
    if (!e->wts.loaded) { e->err = "weights not set"; return GZ_E_NOWEIGHTS; }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_TRY(gz_train_flush(e));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    for (int i = 0; i < 16; ++i)
        GZ_CUDA(e, cudaMemcpy(ptrs[i], e->wts.raw[i], e->wts.raw_elems[i] * sizeof(float), cudaMemcpyDefault));
    return GZ_OK;
}

// ---- diagnostic evaluator (parity tests of the tree kernels) -------------------------------------------------------
// The integer-hash evaluator of oracle/orc_mcts.c:orc_hash_eval computed on the device from the leaf planes (own /
// opponent stones, colour plane; start_player 0 assumed: an even stone count means player 1 is to move), so that the
// FUSED search paths (gz_search, gz_search_vl) can be compared with the oracle bit for bit - the real net differs
// from the oracle's by ~1e-6, which flips near-ties.  Off unless gz_debug_hash_evaluator switches it on.
__device__ __forceinline__ uint64_t gz_splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void __launch_bounds__(128) hash_eval_kernel(const float *planes, const int32_t *count, long long maxB, int W,
                                                        int S, uint64_t salt, float *probs, float *values)
{
    const int lane = threadIdx.x & 31;
    const long long b = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
    const long long n = count ? min((long long)*count, maxB) : maxB;
    if (b >= n) return;
    const float *pl = planes + b * 4 * S;
    const int cur = pl[3 * S] > 0.5f ? 1 : 2; // plane 3 is all ones iff the stone count is even
    uint64_t key = 0ull;
    for (int mv = lane; mv < S; mv += 32) {
        const int idx = (W - 1 - mv / W) * W + mv % W; // current_state is stored row-reversed (game.py:116)
        const int st = pl[idx] > 0.5f ? cur : (pl[S + idx] > 0.5f ? 3 - cur : 0);
        if (st) key ^= gz_splitmix64(salt + (uint64_t)(mv * 2 + (st - 1)) + 1ull);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) key ^= __shfl_xor_sync(0xffffffffu, key, off);
    key ^= gz_splitmix64(salt + 1000ull + (uint64_t)cur);
    uint32_t wsum = 0u;
    for (int mv = lane; mv < S; mv += 32) {
        const int idx = (W - 1 - mv / W) * W + mv % W;
        if (pl[idx] > 0.5f || pl[S + idx] > 0.5f) continue;
        const uint32_t t = (uint32_t)(gz_splitmix64(key + (uint64_t)mv + 1ull) >> 56) + 1u;
        wsum += t * t * t;
    }
    wsum = __reduce_add_sync(0xffffffffu, wsum);
    for (int mv = lane; mv < S; mv += 32) {
        const int idx = (W - 1 - mv / W) * W + mv % W;
        float pr = 0.f;
        if (!(pl[idx] > 0.5f || pl[S + idx] > 0.5f)) {
            const uint32_t t = (uint32_t)(gz_splitmix64(key + (uint64_t)mv + 1ull) >> 56) + 1u;
            pr = __fdiv_rn((float)(t * t * t), (float)wsum);
        }
        probs[b * S + mv] = pr;
    }
    if (lane == 0) {
        const int vi = (int)(gz_splitmix64(key ^ 0x5555555555555555ull) % 2001ull) - 1000;
        values[b] = __fdiv_rn((float)vi, 1000.0f);
    }
}

extern "C" int gz_debug_hash_evaluator(gz_engine *e, int32_t on, uint64_t salt)
{
    if (!e) return GZ_E_INVALID;
    e->dbg_hash_on = on != 0;
    e->dbg_hash_salt = salt;
    return GZ_OK;
}

int gzk_net_forward(gz_engine *e, const float *d_planes, const int32_t *d_count, int64_t maxB,
                    float *d_probs, float *d_values)
{
    if (e->dbg_hash_on) {
        if (maxB <= 0) return GZ_OK;
        hash_eval_kernel<<<(unsigned)((maxB + 3) / 4), 128, 0, e->stream>>>(d_planes, d_count, (long long)maxB, e->W, e->S,
                                                                            e->dbg_hash_salt, d_probs, d_values);
        e->launches++;
        GZ_CUDA(e, cudaGetLastError());
        return GZ_OK;
    }
    if (!e->wts.loaded) { e->err = "gz_set_weights has not been called"; return GZ_E_NOWEIGHTS; }
    GZ_TRY(gz_train_flush(e)); // weights trained on the device since the last forward
    // net_impl: 0 = tcgen05 path (fused kernel for the reference trunk 32/64/128, layer-wise kernels of
    // gz_net_tcw.cu for wider trunks) when it supports the configuration, 1 = fp32 CUDA-core path,
    // 2 = tcgen05 path or an error
    const int impl = e->cfg.net_impl;
    if (impl != 1 && !gz_tc_supported(e) && gz_tcw_supported(e))
        return gz_net_forward_tcw(e, d_planes, d_count, maxB, d_probs, d_values, false);
    if (impl == 2 || (impl == 0 && gz_tc_supported(e)))
    {
        const bool pk = e->leaf_packed && e->pk_slot && (d_planes == e->leaf_planes || d_planes == e->vl_planes);
        return gz_net_forward_tc(e, d_planes, d_count, maxB, d_probs, d_values, false, pk ? e->pk_slot : nullptr,
                                 e->pk_board, e->pk_meta);
    }
    return gz_net_forward_ffma(e, d_planes, d_count, maxB, d_probs, d_values);
}

// ---- stateless rules ------------------------------------------------------------------------
extern "C" int gz_rules_eval(gz_engine *e, const int32_t *moves, const int64_t *offsets, int64_t n_pos,
                             const int8_t *start_player, int32_t *status, int8_t *ended, int8_t *winner,
                             int8_t *has_win, uint8_t *legal, float *planes, uint8_t *target,
                             int8_t *cur_player)
{
    if (!e || !offsets || n_pos < 0) return GZ_E_INVALID;
    if (n_pos == 0) return GZ_OK;
    if (planes && e->W != e->H) { e->err = "current_state needs a square board (game.py:97-110)"; return GZ_E_INVALID; }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    int64_t total = 0;
    if (gz_is_device_ptr(offsets))
        GZ_CUDA(e, cudaMemcpy(&total, offsets + n_pos, sizeof(int64_t), cudaMemcpyDeviceToHost));
    else
        total = offsets[n_pos];
    if (total > 0 && !moves) return GZ_E_INVALID;
    const size_t S = e->S, N = (size_t)n_pos;
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)total * 4 + (N + 1) * 8 + N * (1 + 4 + 1 + 1 + 1 + 1) + N * S * (1 + 16 + 1) + 16 * 256));
    const int32_t *dm = a.in(moves, (size_t)total);
    const int64_t *dof = a.in(offsets, N + 1);
    const int8_t *dsp = a.in(start_player, N);
    int32_t *dst = a.out(status, N);
    int8_t *den = a.out(ended, N), *dwi = a.out(winner, N), *dhw = a.out(has_win, N);
    uint8_t *dle = a.out(legal, N * S);
    float *dpl = a.out(planes, N * 4 * S);
    uint8_t *dtg = a.out(target, N * S);
    int8_t *dcu = a.out(cur_player, N);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_rules_eval(e, dm, dof, n_pos, dsp, dst, den, dwi, dhw, dle, dpl, dtg, dcu));
    return a.finish();
}

// ---- resident boards --------------------------------------------------------------------------
extern "C" int gz_boards_reset(gz_engine *e, const int8_t *start_player)
{
    if (!e) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    ArgStage a(e);
    GZ_TRY(a.reserve(e->G));
    const int8_t *d = a.in(start_player, e->G);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_boards_reset(e, d));
    GZ_CUDA(e, cudaMemsetAsync(e->active, 1, e->G, e->stream));
    GZ_TRY(gzk_tree_reset(e));
    e->sp_begun = false;
    return a.finish();
}

__global__ void boards_set_kernel(GameBoard *boards, GameMeta *meta, int game, const int32_t *moves,
                                  int n_moves, int start_player, int W, int S, int32_t *status)
{
    BB st[2] = {bb_zero(), bb_zero()};
    int player = start_player, last = -1, bad = 0;
    for (int i = 0; i < n_moves; ++i) {
        int mv = moves[i];
        if (mv < 0 || mv >= S) { bad = 1; break; }
        int bit = bb_bit(mv, W);
        if (bb_test(st[0], bit) || bb_test(st[1], bit)) { bad = 1; break; }
        bb_set(st[player], bit);
        player ^= 1;
        last = mv;
    }
    *status = bad ? GZ_E_ILLEGAL : 0;
    if (bad) return;
    store_bb(boards[game].stones[0], st[0]);
    store_bb(boards[game].stones[1], st[1]);
    GameMeta m;
    m.n_stones = n_moves; m.last_move = last; m.start_player = start_player; m.flags = 0;
    meta[game] = m;
}

extern "C" int gz_boards_set(gz_engine *e, int32_t game, const int32_t *moves, int32_t n_moves,
                             int32_t start_player)
{
    if (!e || game < 0 || game >= e->G || n_moves < 0 || n_moves > e->S || (n_moves && !moves)) return GZ_E_INVALID;
    if (start_player != 0 && start_player != 1) { e->err = "start_player should be either 0 or 1"; return GZ_E_INVALID; }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)n_moves * 4 + 256));
    const int32_t *dm = a.in(moves, (size_t)n_moves);
    int32_t hstatus = 0;
    int32_t *dst = a.out(&hstatus, 1);
    if (a.rc) return a.rc;
    boards_set_kernel<<<1, 1, 0, e->stream>>>(e->boards, e->meta, game, dm, n_moves, start_player, e->W, e->S, dst);
    e->launches++;
    GZ_TRY(a.finish());
    if (hstatus) { e->err = "illegal move in position"; return GZ_E_ILLEGAL; }
    return GZ_OK;
}

extern "C" int gz_boards_do_move(gz_engine *e, const int32_t *moves, int32_t *status)
{
    if (!e || !moves) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)e->G * 8 + 512));
    const int32_t *dm = a.in(moves, e->G);
    int32_t *ds = a.out(status, e->G);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_boards_do_move(e, dm, ds));
    return a.finish();
}

extern "C" int gz_boards_status(gz_engine *e, int8_t *ended, int8_t *winner, int8_t *cur_player,
                                int32_t *n_stones, int32_t *last_move)
{
    if (!e) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)e->G * 11 + 2048));
    int8_t *de = a.out(ended, e->G), *dw = a.out(winner, e->G), *dc = a.out(cur_player, e->G);
    int32_t *dn = a.out(n_stones, e->G), *dl = a.out(last_move, e->G);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_boards_query(e, de, dw, dc, dn, dl, nullptr, nullptr, nullptr));
    return a.finish();
}

extern "C" int gz_boards_legal_mask(gz_engine *e, uint8_t *out)
{
    if (!e || !out) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)e->G * e->S + 256));
    uint8_t *d = a.out(out, (size_t)e->G * e->S);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_boards_query(e, nullptr, nullptr, nullptr, nullptr, nullptr, d, nullptr, nullptr));
    return a.finish();
}

extern "C" int gz_boards_planes(gz_engine *e, float *out)
{
    if (!e || !out) return GZ_E_INVALID;
    if (e->W != e->H) { e->err = "current_state needs a square board (game.py:97-110)"; return GZ_E_INVALID; }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)e->G * e->S * 16 + 256));
    float *d = a.out(out, (size_t)e->G * e->S * 4);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_boards_query(e, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, d, nullptr));
    return a.finish();
}

extern "C" int gz_boards_cells(gz_engine *e, int8_t *out)
{
    if (!e || !out) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)e->G * e->S + 256));
    int8_t *d = a.out(out, (size_t)e->G * e->S);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_boards_query(e, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, d));
    return a.finish();
}

// ---- net --------------------------------------------------------------------------------------
extern "C" int gz_net_forward(gz_engine *e, const float *planes, int64_t B, float *probs, float *values)
{
    if (!e || !planes || !probs || !values || B < 0) return GZ_E_INVALID;
    if (B == 0) return GZ_OK;
    if (e->W != e->H) { e->err = "the net needs a square board"; return GZ_E_INVALID; }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    const size_t S = e->S;
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)B * S * 20 + (size_t)B * 4 + 1024));
    const float *dp = a.in(planes, (size_t)B * 4 * S);
    float *dpr = a.out(probs, (size_t)B * S);
    float *dv = a.out(values, (size_t)B);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_net_forward(e, dp, nullptr, B, dpr, dv));
    return a.finish();
}

// Diagnostic: the tcgen05 path with its per-layer activations exported (parity tests).
extern "C" int gz_debug_net_layers(gz_engine *e, const float *planes, int64_t B, float *act1, float *act2,
                                   float *act3, float *head, float *probs, float *values)
{
    if (!e || !planes || B <= 0) return GZ_E_INVALID;
    if (!e->wts.loaded) { e->err = "gz_set_weights has not been called"; return GZ_E_NOWEIGHTS; }
    const bool wide = !gz_tc_supported(e) && gz_tcw_supported(e);
    if (!gz_tc_supported(e) && !wide) { e->err = "tcgen05 net path: unsupported board or trunk widths"; return GZ_E_INVALID; }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    const size_t S = e->S, N = (size_t)B;
    if (wide) { // layer-wise kernels: act1 / act3 / head from the kernels' dumps, act2 decoded from its fp16 image
        const size_t c1 = e->c1, c2 = e->c2, c3 = e->c3;
        ArgStage a(e);
        GZ_TRY(a.reserve(N * S * 20 + N * 8 + 2048));
        const float *dp = a.in(planes, N * 4 * S);
        float *dpr = a.out(probs, N * S);
        float *dv = a.out(values, N);
        if (a.rc) return a.rc;
        if (!dpr || !dv) { e->err = "gz_debug_net_layers: probs and values are required"; return GZ_E_INVALID; }
        GZ_TRY(gz_train_flush(e));
        GZ_TRY(gz_net_forward_tcw(e, dp, nullptr, B, dpr, dv, true));
        GZ_TRY(a.finish());
        GZ_CUDA(e, cudaStreamSynchronize(e->stream));
        if (act1) GZ_CUDA(e, cudaMemcpy(act1, e->tcw_dbg, N * S * c1 * 4, cudaMemcpyDefault));
        if (act2) GZ_TRY(gz_tcw_decode_act2(e, B, act2));
        if (act3) GZ_CUDA(e, cudaMemcpy(act3, e->tcw_dbg + N * S * (c1 + c2), N * S * c3 * 4, cudaMemcpyDefault));
        if (head) GZ_CUDA(e, cudaMemcpy(head, e->tcw_dbg + N * S * (c1 + c2 + c3), N * 6 * S * 4, cudaMemcpyDefault));
        return GZ_OK;
    }
    ArgStage a(e);
    GZ_TRY(a.reserve(N * S * 20 + N * 8 + 2048));
    const float *dp = a.in(planes, N * 4 * S);
    float *dpr = a.out(probs, N * S);
    float *dv = a.out(values, N);
    if (a.rc) return a.rc;
    float *tp = nullptr, *tv = nullptr;
    if (!dpr) { GZ_CUDA(e, cudaMalloc(&tp, N * S * sizeof(float))); dpr = tp; }
    if (!dv) { GZ_CUDA(e, cudaMalloc(&tv, N * sizeof(float))); dv = tv; }
    int rc = gz_tc_ensure_buffers(e, B, true);
    if (rc == GZ_OK) GZ_CUDA(e, cudaMemsetAsync(e->tc_dbg, 0, N * S * 230 * sizeof(float), e->stream));
    if (rc == GZ_OK) rc = gz_net_forward_tc(e, dp, nullptr, B, dpr, dv, true);
    if (rc == GZ_OK) rc = a.finish();
    if (rc == GZ_OK) {
        if (act1) GZ_CUDA(e, cudaMemcpy(act1, e->tc_dbg, N * S * 32 * 4, cudaMemcpyDefault));
        if (act2) GZ_CUDA(e, cudaMemcpy(act2, e->tc_dbg + N * S * 32, N * S * 64 * 4, cudaMemcpyDefault));
        if (act3) GZ_CUDA(e, cudaMemcpy(act3, e->tc_dbg + N * S * 96, N * S * 128 * 4, cudaMemcpyDefault));
        if (head) GZ_CUDA(e, cudaMemcpy(head, e->tc_dbg + N * S * 224, N * 6 * S * 4, cudaMemcpyDefault));
    }
    if (tp) cudaFree(tp);
    if (tv) cudaFree(tv);
    return rc;
}

// ---- search -----------------------------------------------------------------------------------
// n playouts in every active game.  One wave = select -> net on the gathered leaves -> expand +
// backup; the expand/backup of wave k and the select of wave k+1 run as one fused kernel.
static int gz_waves(gz_engine *e, int n)
{
    if (n <= 0) return GZ_OK;
    // When the leaves go to the fused tcgen05 trunk, the tree kernels do not write fp32 planes (1 KB per 8x8 leaf): they
    // record which game owns a batch slot and the trunk's staging warps build the four planes from that game's leaf
    // bitboards (64 B + 16 B, which the tree kernels store anyway).  GZ_LEAF_FLOAT=1 keeps the planes (A/B switch).
    static const bool leaf_float = getenv("GZ_LEAF_FLOAT") != nullptr;
    struct PackedScope {
        gz_engine *e;
        ~PackedScope() { e->leaf_packed = false; e->pk_slot = nullptr; }
    } scope{e};
    e->leaf_packed = !leaf_float && !e->dbg_hash_on && e->wts.loaded && e->cfg.net_impl != 1 && gz_tc_supported(e);
    e->pk_slot = e->slot_game; e->pk_board = e->leaf_board; e->pk_meta = e->leaf_meta;
    GZ_TRY(gzk_select(e, false));
    for (int i = 0; i < n; ++i) {
        GZ_TRY(gzk_net_forward(e, e->leaf_planes, e->leaf_count, e->G, e->net_probs, e->net_values));
        if (i + 1 < n) GZ_TRY(gzk_expand_select(e, e->net_probs, e->net_values));
        else GZ_TRY(gzk_expand_backup(e, e->net_probs, e->net_values, true));
    }
    return GZ_OK;
}

extern "C" int gz_search(gz_engine *e, int32_t n_playout)
{
    if (!e || n_playout < 0) return GZ_E_INVALID;
    if (!e->wts.loaded && !e->dbg_hash_on) { e->err = "gz_set_weights has not been called"; return GZ_E_NOWEIGHTS; }
    if (e->W != e->H) { e->err = "the net needs a square board"; return GZ_E_INVALID; }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_TRY(gz_waves(e, n_playout));
    return GZ_OK;
}

__global__ void leaf_export_kernel(const int8_t *leaf_state, const int32_t *leaf_slot, const float *leaf_planes,
                                   const GameBoard *leaf_board, const GameMeta *leaf_meta, int G, int W, int S,
                                   float *planes_out, int8_t *cells_out, int8_t *player_out)
{
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (g >= G) return;
    const int st = leaf_state[g];
    if (planes_out) {
        for (int i = lane; i < 4 * S; i += GZ_WARP)
            planes_out[(size_t)g * 4 * S + i] = (st == 0) ? leaf_planes[(size_t)leaf_slot[g] * 4 * S + i] : 0.f;
    }
    if (st == 2) return;
    GameMeta m = leaf_meta[g];
    if (cells_out) {
        BB s0 = load_bb(leaf_board[g].stones[0]), s1 = load_bb(leaf_board[g].stones[1]);
        for (int mv = lane; mv < S; mv += GZ_WARP) {
            int bit = bb_bit(mv, W);
            cells_out[(size_t)g * S + mv] = bb_test(s0, bit) ? 1 : (bb_test(s1, bit) ? 2 : 0);
        }
    }
    if (player_out && lane == 0) player_out[g] = (int8_t)(((m.start_player + m.n_stones) & 1) + 1);
}

extern "C" int gz_search_select(gz_engine *e, int8_t *leaf_state, float *leaf_planes, int8_t *leaf_cells,
                                int8_t *leaf_player)
{
    if (!e) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    const size_t G = e->G, S = e->S;
    ArgStage a(e);
    GZ_TRY(a.reserve(G * (2 + S * 17) + 2048));
    int8_t *dls = a.out(leaf_state, G);
    float *dpl = a.out(leaf_planes, G * 4 * S);
    int8_t *dce = a.out(leaf_cells, G * S);
    int8_t *dpy = a.out(leaf_player, G);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_select(e, false));
    if (dce) GZ_CUDA(e, cudaMemsetAsync(dce, 0, G * S, e->stream));
    if (dpy) GZ_CUDA(e, cudaMemsetAsync(dpy, 0, G, e->stream));
    leaf_export_kernel<<<gz_warp_blocks(e->G, 4), 128, 0, e->stream>>>(
        e->leaf_state, e->leaf_slot, e->leaf_planes, e->leaf_board, e->leaf_meta, e->G, e->W, e->S, dpl, dce, dpy);
    e->launches++;
    if (dls) GZ_CUDA(e, cudaMemcpyAsync(dls, e->leaf_state, G, cudaMemcpyDeviceToDevice, e->stream));
    return a.finish();
}

extern "C" int gz_search_expand_backup(gz_engine *e, const float *priors, const float *values)
{
    if (!e || !priors || !values) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    const size_t G = e->G, S = e->S;
    ArgStage a(e);
    GZ_TRY(a.reserve(G * S * 4 + G * 4 + 1024));
    const float *dp = a.in(priors, G * S);
    const float *dv = a.in(values, G);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_expand_backup(e, dp, dv, false));
    return a.finish();
}

__global__ void paths_export_kernel(const uint32_t *path, const int32_t *path_len, int G, int S, int Sp,
                                    int32_t *moves_out, int32_t *len_out)
{
    const int g = blockIdx.x;
    const int d = path_len[g];
    for (int k = threadIdx.x; k < S; k += blockDim.x)
        moves_out[(size_t)g * S + k] = k < d ? (int32_t)(path[(size_t)g * S + k] % (uint32_t)Sp) : -1;
    if (threadIdx.x == 0 && len_out) len_out[g] = d;
}

extern "C" int gz_search_paths(gz_engine *e, int32_t *path_moves, int32_t *path_len)
{
    if (!e || !path_moves) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    const size_t G = e->G, S = e->S;
    ArgStage a(e);
    GZ_TRY(a.reserve(G * S * 4 + G * 4 + 1024));
    int32_t *dm = a.out(path_moves, G * S);
    int32_t *dl = a.out(path_len, G);
    if (a.rc) return a.rc;
    paths_export_kernel<<<e->G, 64, 0, e->stream>>>(e->path, e->path_len, e->G, e->S, e->Sp, dm, dl);
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return a.finish();
}

extern "C" int gz_root_visits(gz_engine *e, int32_t *visits, double *q, int32_t *root_n)
{
    if (!e) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    const size_t G = e->G, S = e->S;
    ArgStage a(e);
    GZ_TRY(a.reserve(G * S * 12 + G * 4 + 1024));
    int32_t *dv = a.out(visits, G * S);
    double *dq = a.out(q, G * S);
    int32_t *dn = a.out(root_n, G);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_root_visits(e, dv, dq, dn));
    return a.finish();
}

extern "C" int gz_update_with_move(gz_engine *e, const int32_t *moves)
{
    if (!e || !moves) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)e->G * 4 + 256));
    const int32_t *dm = a.in(moves, e->G);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_update_with_move(e, dm));
    return a.finish();
}

extern "C" int gz_set_active(gz_engine *e, const int8_t *mask)
{
    if (!e) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    if (!mask) {
        GZ_CUDA(e, cudaMemsetAsync(e->active, 1, e->G, e->stream));
        return GZ_OK;
    }
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)e->G + 256));
    const int8_t *dm = a.in(mask, e->G);
    if (a.rc) return a.rc;
    GZ_CUDA(e, cudaMemcpyAsync(e->active, dm, e->G, cudaMemcpyDeviceToDevice, e->stream));
    return a.finish();
}

extern "C" int gz_root_select_move(gz_engine *e, float temp, int32_t selfplay, int32_t *moves, float *probs)
{
    if (!e || !moves || !(temp > 0.f)) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    const size_t G = e->G, S = e->S;
    ArgStage a(e);
    GZ_TRY(a.reserve(G * S * 4 + G * 4 + 1024));
    int32_t *dm = a.out(moves, G);
    float *dp = a.out(probs, G * S);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_root_select_move(e, temp, selfplay, dm, dp, false));
    return a.finish();
}

// ---- self-play --------------------------------------------------------------------------------
extern "C" int gz_selfplay_begin(gz_engine *e, float temp, int32_t recycle)
{
    if (!e || !(temp > 0.f)) return GZ_E_INVALID;
    if (!e->wts.loaded) { e->err = "gz_set_weights has not been called"; return GZ_E_NOWEIGHTS; }
    if (e->W != e->H) { e->err = "the net needs a square board"; return GZ_E_INVALID; }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    const size_t G = e->G, S = e->S;
    if (!e->sp_moves) {
        GZ_TRY(gz_dalloc(e, &e->sp_moves, G * S));
        GZ_TRY(gz_dalloc(e, &e->sp_pi, G * S * S));
        GZ_TRY(gz_dalloc(e, &e->sp_game_id, G));
        GZ_TRY(gz_dalloc(e, &e->sp_fin_flag, G));
        GZ_TRY(gz_dalloc(e, &e->sp_serial, 1));
        e->out_cap = (int64_t)(G * S);
        GZ_TRY(gz_dalloc(e, &e->out_states, (size_t)e->out_cap * 4 * S));
        GZ_TRY(gz_dalloc(e, &e->out_pi, (size_t)e->out_cap * S));
        GZ_TRY(gz_dalloc(e, &e->out_z, (size_t)e->out_cap));
        GZ_TRY(gz_dalloc(e, &e->out_gid, (size_t)e->out_cap));
        GZ_TRY(gz_dalloc(e, &e->out_ply, (size_t)e->out_cap));
        e->gout_cap = (int64_t)(4 * G);
        GZ_TRY(gz_dalloc(e, &e->gout_moves, (size_t)e->gout_cap * S));
        GZ_TRY(gz_dalloc(e, &e->gout_len, (size_t)e->gout_cap));
        GZ_TRY(gz_dalloc(e, &e->gout_winner, (size_t)e->gout_cap));
        GZ_TRY(gz_dalloc(e, &e->gout_gid, (size_t)e->gout_cap));
    }
    // Game ids key the Philox streams of the Dirichlet noise and the move sampling (gz_move.cu), so they must
    // never repeat on one engine: a second begin continues after the largest id any slot has reached
    // (recycling hands out id + G), instead of replaying ids 0..G-1 and with them the same noise.
    std::vector<int64_t> ids(G);
    if (e->sp_ids_live) {
        GZ_CUDA(e, cudaStreamSynchronize(e->stream));
        GZ_CUDA(e, cudaMemcpy(ids.data(), e->sp_game_id, G * sizeof(int64_t), cudaMemcpyDeviceToHost));
        int64_t mx = -1;
        for (size_t g = 0; g < G; ++g) mx = std::max(mx, ids[g] - (int64_t)e->cfg.game_id_base);
        e->sp_next_serial = std::max(e->sp_next_serial, mx + 1);
    }
    for (size_t g = 0; g < G; ++g) ids[g] = (int64_t)(e->cfg.game_id_base + (uint64_t)e->sp_next_serial + g);
    e->sp_next_serial += (int64_t)G;
    e->sp_ids_live = true;
    const unsigned long long dev_serial = (unsigned long long)e->sp_next_serial; // recycle mode 2 continues from here
    GZ_CUDA(e, cudaMemcpyAsync(e->sp_serial, &dev_serial, sizeof(dev_serial), cudaMemcpyHostToDevice, e->stream));
    GZ_CUDA(e, cudaMemsetAsync(e->sp_fin_flag, 0, G * sizeof(int32_t), e->stream));
    GZ_CUDA(e, cudaMemcpyAsync(e->sp_game_id, ids.data(), G * sizeof(int64_t), cudaMemcpyHostToDevice, e->stream));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    GZ_TRY(gzk_boards_reset(e, nullptr));
    GZ_TRY(gzk_tree_reset(e));
    GZ_CUDA(e, cudaMemsetAsync(e->active, 1, G, e->stream));
    GZ_CUDA(e, cudaMemsetAsync(e->out_count, 0, sizeof(unsigned long long), e->stream));
    GZ_CUDA(e, cudaMemsetAsync(e->gout_count, 0, sizeof(unsigned long long), e->stream));
    e->out_drained = 0;
    e->gout_drained = 0;
    e->sp_temp = temp;
    e->sp_recycle = recycle == 2 ? 2 : (recycle ? 1 : 0);
    e->sp_begun = true;
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    return GZ_OK;
}

extern "C" int gz_selfplay_step(gz_engine *e, int32_t n_plies)
{
    if (!e || n_plies < 0) return GZ_E_INVALID;
    if (!e->sp_begun) { e->err = "gz_selfplay_begin has not been called"; return GZ_E_INVALID; }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    for (int p = 0; p < n_plies; ++p) {
        GZ_TRY(gz_waves(e, e->cfg.n_playout));
        GZ_TRY(gzk_root_select_move(e, e->sp_temp, 1, e->sel_moves, e->sel_probs, false));
        GZ_TRY(gzk_selfplay_advance(e));
    }
    return GZ_OK;
}

extern "C" int gz_selfplay_pending(gz_engine *e, int64_t *n_samples, int64_t *n_games_finished,
                                   int64_t *n_active_slots)
{
    if (!e || !e->sp_begun) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    unsigned long long cnt = 0;
    DevCounters c;
    GZ_CUDA(e, cudaMemcpy(&cnt, e->out_count, sizeof(cnt), cudaMemcpyDeviceToHost));
    GZ_CUDA(e, cudaMemcpy(&c, e->counters, sizeof(c), cudaMemcpyDeviceToHost));
    if (n_samples) *n_samples = (int64_t)cnt - e->out_drained;
    if (n_games_finished) *n_games_finished = (int64_t)c.games_finished;
    if (n_active_slots) {
        std::vector<int8_t> act(e->G);
        GZ_CUDA(e, cudaMemcpy(act.data(), e->active, e->G, cudaMemcpyDeviceToHost));
        int64_t n = 0;
        for (int8_t a : act) n += a ? 1 : 0;
        *n_active_slots = n;
    }
    return GZ_OK;
}

extern "C" int gz_selfplay_drain(gz_engine *e, int64_t max_samples, float *states, float *pis, float *zs,
                                 int64_t *game_ids, int32_t *plies, int64_t *n_out)
{
    if (!e || !e->sp_begun || !n_out || max_samples < 0) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    unsigned long long cnt = 0;
    GZ_CUDA(e, cudaMemcpy(&cnt, e->out_count, sizeof(cnt), cudaMemcpyDeviceToHost));
    int64_t n = (int64_t)cnt - e->out_drained;
    if (n > max_samples) n = max_samples;
    const size_t S = e->S, o = (size_t)e->out_drained;
    if (n > 0) {
        if (states) GZ_CUDA(e, cudaMemcpyAsync(states, e->out_states + o * 4 * S, (size_t)n * 4 * S * 4, cudaMemcpyDefault, e->stream));
        if (pis) GZ_CUDA(e, cudaMemcpyAsync(pis, e->out_pi + o * S, (size_t)n * S * 4, cudaMemcpyDefault, e->stream));
        if (zs) GZ_CUDA(e, cudaMemcpyAsync(zs, e->out_z + o, (size_t)n * 4, cudaMemcpyDefault, e->stream));
        if (game_ids) GZ_CUDA(e, cudaMemcpyAsync(game_ids, e->out_gid + o, (size_t)n * 8, cudaMemcpyDefault, e->stream));
        if (plies) GZ_CUDA(e, cudaMemcpyAsync(plies, e->out_ply + o, (size_t)n * 4, cudaMemcpyDefault, e->stream));
        GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    }
    e->out_drained += n;
    if (e->out_drained == (int64_t)cnt) {
        GZ_CUDA(e, cudaMemset(e->out_count, 0, sizeof(unsigned long long)));
        e->out_drained = 0;
    }
    *n_out = n;
    return GZ_OK;
}

extern "C" int gz_selfplay_drain_games(gz_engine *e, int64_t max_games, int64_t max_moves, int32_t *moves,
                                       int64_t *offsets, int8_t *winners, int64_t *game_ids,
                                       int64_t *n_games_out)
{
    if (!e || !e->sp_begun || !n_games_out || !moves || !offsets || max_games < 0) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    unsigned long long cnt = 0;
    GZ_CUDA(e, cudaMemcpy(&cnt, e->gout_count, sizeof(cnt), cudaMemcpyDeviceToHost));
    int64_t avail = (int64_t)cnt - e->gout_drained;
    const size_t S = e->S, o = (size_t)e->gout_drained;
    std::vector<int32_t> len((size_t)avail);
    if (avail > 0) GZ_CUDA(e, cudaMemcpy(len.data(), e->gout_len + o, (size_t)avail * 4, cudaMemcpyDeviceToHost));
    int64_t n = 0, total = 0;
    while (n < avail && n < max_games && total + len[(size_t)n] <= max_moves) total += len[(size_t)n++];
    std::vector<int32_t> mv((size_t)n * S);
    if (n > 0) GZ_CUDA(e, cudaMemcpy(mv.data(), e->gout_moves + o * S, (size_t)n * S * 4, cudaMemcpyDeviceToHost));
    int64_t off = 0;
    offsets[0] = 0;
    for (int64_t i = 0; i < n; ++i) {
        memcpy(moves + off, mv.data() + (size_t)i * S, (size_t)len[(size_t)i] * 4);
        off += len[(size_t)i];
        offsets[i + 1] = off;
    }
    if (n > 0) {
        if (winners) GZ_CUDA(e, cudaMemcpy(winners, e->gout_winner + o, (size_t)n, cudaMemcpyDeviceToHost));
        if (game_ids) GZ_CUDA(e, cudaMemcpy(game_ids, e->gout_gid + o, (size_t)n * 8, cudaMemcpyDeviceToHost));
    }
    e->gout_drained += n;
    if (e->gout_drained == (int64_t)cnt) {
        GZ_CUDA(e, cudaMemset(e->gout_count, 0, sizeof(unsigned long long)));
        e->gout_drained = 0;
    }
    *n_games_out = n;
    return GZ_OK;
}

extern "C" int gz_selfplay_get_stats(gz_engine *e, gz_selfplay_stats *out, int32_t reset)
{
    if (!e || !out) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    DevCounters c;
    GZ_CUDA(e, cudaMemcpy(&c, e->counters, sizeof(c), cudaMemcpyDeviceToHost));
    out->n_games = (int64_t)c.games_finished;
    out->n_samples = (int64_t)c.samples_out;
    out->n_playouts = (int64_t)c.playouts;
    out->n_leaf_evals = (int64_t)c.leaf_evals;
    out->n_plies = (int64_t)c.plies;
    out->pool_overflows = (int64_t)c.pool_overflows;
    for (int i = 0; i < 3; ++i) out->n_wins[i] = (int64_t)c.wins[i];
    out->max_tree_nodes = (int64_t)c.max_nodes;
    if (reset) GZ_CUDA(e, cudaMemset(e->counters, 0, sizeof(DevCounters)));
    return GZ_OK;
}

extern "C" int gz_selfplay_state(gz_engine *e, int64_t *game_ids, int64_t *dropped_samples)
{
    if (!e || !e->sp_begun) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    if (game_ids) GZ_CUDA(e, cudaMemcpy(game_ids, e->sp_game_id, (size_t)e->G * sizeof(int64_t), cudaMemcpyDefault));
    if (dropped_samples) {
        unsigned long long d = 0;
        GZ_CUDA(e, cudaMemcpy(&d, &e->counters->dropped, sizeof(d), cudaMemcpyDeviceToHost));
        *dropped_samples = (int64_t)d;
    }
    return GZ_OK;
}

extern "C" int gz_selfplay_last(gz_engine *e, int32_t *moves, float *probs)
{
    if (!e || !e->sp_begun) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    const size_t G = e->G, S = e->S;
    if (moves) GZ_CUDA(e, cudaMemcpyAsync(moves, e->sel_moves, G * 4, cudaMemcpyDefault, e->stream));
    if (probs) GZ_CUDA(e, cudaMemcpyAsync(probs, e->sel_probs, G * S * 4, cudaMemcpyDefault, e->stream));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    return GZ_OK;
}

// ---- pure MCTS ----------------------------------------------------------------------------------
// Board.get_move_target on an EMPTY board is the hard-coded box rows/cols 5..9 (game.py:41); on a board narrower
// than 6 it holds no cell and the reference raises (max() of an empty sequence, mcts_pure.py:205).  The
// kernels end such a rollout / skip such an expansion and count it; the next synchronising pure-MCTS call
// reports it as an error instead of returning a move chosen from nothing.
static int gz_check_empty_box(gz_engine *e)
{
    unsigned long long n = 0;
    GZ_CUDA(e, cudaMemcpy(&n, &e->counters->empty_box, sizeof(n), cudaMemcpyDeviceToHost));
    if (n == 0) return GZ_OK;
    GZ_CUDA(e, cudaMemset(&e->counters->empty_box, 0, sizeof(n)));
    e->err = "pure MCTS: empty candidate box (empty board narrower than 6: game.py:41 hard-codes rows/cols 5..9)";
    return GZ_E_INVALID;
}

extern "C" int gz_pure_search(gz_engine *e, int32_t n_playout, uint32_t search_id)
{
    if (!e || n_playout < 0) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    return gzk_pure_playouts(e, n_playout, search_id);
}

extern "C" int gz_pure_best_move(gz_engine *e, int32_t *moves)
{
    if (!e || !moves) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)e->G * 4 + 256));
    int32_t *dm = a.out(moves, e->G);
    if (a.rc) return a.rc;
    GZ_TRY(gzk_root_select_move(e, 1.0f, 0, dm, nullptr, true));
    GZ_TRY(a.finish());
    return gz_check_empty_box(e);
}

extern "C" int gz_rollouts(gz_engine *e, int32_t n_rollouts, uint32_t search_id, int8_t *results,
                           int64_t *total_plies)
{
    if (!e || !results || n_rollouts < 1) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    const size_t n = (size_t)e->G * n_rollouts;
    ArgStage a(e);
    GZ_TRY(a.reserve(n + 256));
    int8_t *dr = a.out(results, n);
    if (a.rc) return a.rc;
    DevCounters c0, c1;
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    GZ_CUDA(e, cudaMemcpy(&c0, e->counters, sizeof(c0), cudaMemcpyDeviceToHost));
    GZ_TRY(gzk_rollouts(e, n_rollouts, search_id, dr));
    GZ_TRY(a.finish());
    GZ_CUDA(e, cudaMemcpy(&c1, e->counters, sizeof(c1), cudaMemcpyDeviceToHost));
    if (total_plies) *total_plies = (int64_t)(c1.rollout_plies - c0.rollout_plies);
    return gz_check_empty_box(e);
}

extern "C" int gz_get_counters(gz_engine *e, gz_counters *out, int32_t reset)
{
    if (!e || !out) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    DevCounters c;
    GZ_CUDA(e, cudaMemcpy(&c, e->counters, sizeof(c), cudaMemcpyDeviceToHost));
    out->kernel_launches = e->launches;
    out->select_levels = (int64_t)c.select_levels;
    out->select_children = (int64_t)c.select_children;
    out->playouts = (int64_t)c.playouts;
    out->leaf_evals = (int64_t)c.leaf_evals;
    out->expanded_children = (int64_t)c.expanded_children;
    out->rollout_plies = (int64_t)c.rollout_plies;
    if (reset) {
        GZ_CUDA(e, cudaMemset(e->counters, 0, sizeof(DevCounters)));
        e->launches = 0;
    }
    return GZ_OK;
}

// ---- per-kernel-class timing ----------------------------------------------------------------------
static const char *g_prof_names[GZ_PROF_CLASSES] = {"select", "conv1", "conv2", "conv3", "heads", "expand_backup",
                                                    "root_select_move", "selfplay_advance", "pure_expand",
                                                    "rollout", "backup", "other", "trunk_tc", "heads_fc"};

extern "C" const char *gz_profile_class_name(int32_t c)
{
    return (c >= 0 && c < GZ_PROF_CLASSES) ? g_prof_names[c] : "";
}

extern "C" int gz_profile_enable(gz_engine *e, int32_t on)
{
    if (!e) return GZ_E_INVALID;
    e->prof_on = on != 0;
    e->prof_stride = on > 1 ? on : 1; // on = N > 1: time every N-th launch of each class and scale
    return GZ_OK;
}

extern "C" int gz_profile_read(gz_engine *e, double *ms, int64_t *launches, int32_t reset)
{
    if (!e) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    for (int c = 0; c < GZ_PROF_CLASSES; ++c) {
        double sum = 0.0;
        long long n = 0;
        for (size_t i = 0; i + 1 < e->prof_used[c]; i += 2) {
            float t = 0.f;
            if (cudaEventElapsedTime(&t, e->prof_ev[c][i], e->prof_ev[c][i + 1]) == cudaSuccess) {
                sum += t;
                n += 1;
            }
        }
        if (n > 0) { // scale the timed sample up to every launch seen
            e->prof_ms[c] += sum * (double)e->prof_seen[c] / (double)n;
            e->prof_n[c] += e->prof_seen[c];
        }
        e->prof_seen[c] = 0;
        e->prof_used[c] = 0;
        if (ms) ms[c] = e->prof_ms[c];
        if (launches) launches[c] = e->prof_n[c];
        if (reset) { e->prof_ms[c] = 0.0; e->prof_n[c] = 0; }
    }
    return GZ_OK;
}

#include "gz_augment.cu"
#include "gz_train.cu"
#include "gz_noalloc.cu"
#include "gz_vl.cu"
