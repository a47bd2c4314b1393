// gz_train.cu — PolicyValueNet.train_step on the device (reference policy_value_net_pytorch.py:117-148,
// SURVEY 8f-1): forward with saved activations, loss = MSE(v, z) - mean(sum(pi * log p)), backward,
// Adam with L2 weight decay (torch.optim.Adam(weight_decay=l2_const), :74-75), all fp32 on CUDA cores.
//
// The master parameters live in ONE flat device vector in state_dict order and torch layout (what
// gz_get_weights returns); gradients and both Adam moments are flat vectors of the same shape, so a
// data-parallel caller can all-reduce the gradient vector between gz_train_backward and
// gz_train_apply.  Every pass starts by packing the master vector into the kernel layouts
// (tap-major conv filters, their flipped/transposed twins for the data gradients, transposed FC
// matrices).  Activations are NHWC fp32 [B*S][C] like the fp32 inference path, whose conv kernels
// (gz_net_ffma.cu) are reused for the forward convs and, with flipped filters and a ReLU-mask
// epilogue, for the data gradients.  Weight gradients are split-K partial sums reduced in a fixed
// order (bit-reproducible run to run).
// Included by gz_api.cu (single translation unit).

struct TrainPack {
    int S, c1, c2, c3;
    // offsets (floats) into the packed buffer
    int o_w1t, o_w2t, o_w3t, o_w3d, o_w2d, o_wh, o_wpT, o_wvT, o_bias, total;
    // offsets inside the packed bias block (same order as NetWeights.bias)
    int b_b1, b_b2, b_b3, b_bh, b_bp, b_bv, b_wv2, b_bv2, b_total;
    // offsets of the 16 state_dict tensors in the flat master vector
    int p[17];
};

struct TrainState {
    TrainPack pk;
    int64_t cap = 0;             // boards the activation buffers hold
    float *p = nullptr, *g = nullptr, *m = nullptr, *v = nullptr; // flat [np]
    float *packed = nullptr;
    long long t = 0;             // Adam steps taken
    bool init = false;           // p/m/v hold a live optimiser state
    bool dirty = false;          // p is newer than the engine's inference images
    bool have_grads = false;
    float *a1 = nullptr, *a2 = nullptr, *a3 = nullptr, *g1 = nullptr, *g2 = nullptr, *g3 = nullptr;
    float *h6 = nullptr, *hv = nullptr, *dlog = nullptr, *dhv = nullptr, *dvpre = nullptr, *dh6 = nullptr,
          *terms = nullptr, *probs = nullptr, *values = nullptr;
    float *part = nullptr;       // split-K partial sums
    size_t part_n = 0;
    float *out2 = nullptr;       // loss, entropy
};

// ---- master vector -> kernel layouts -----------------------------------------------------------
__global__ void train_pack_kernel(const float *__restrict__ p, float *__restrict__ out, TrainPack k)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k.total) return;
    const int S = k.S, c1 = k.c1, c2 = k.c2, c3 = k.c3;
    float v = 0.f;
    if (i < k.o_w2t) { // conv1 [tap][4][c1]
        const int j = i - k.o_w1t, tap = j / (4 * c1), ci = (j / c1) % 4, co = j % c1;
        v = p[k.p[0] + (co * 4 + ci) * 9 + tap];
    } else if (i < k.o_w3t) { // conv2 [tap][c1][c2]
        const int j = i - k.o_w2t, tap = j / (c1 * c2), ci = (j / c2) % c1, co = j % c2;
        v = p[k.p[2] + (co * c1 + ci) * 9 + tap];
    } else if (i < k.o_w3d) { // conv3 [tap][c2][c3]
        const int j = i - k.o_w3t, tap = j / (c2 * c3), ci = (j / c3) % c2, co = j % c3;
        v = p[k.p[4] + (co * c2 + ci) * 9 + tap];
    } else if (i < k.o_w2d) { // conv3 data-gradient filter [tap'][c3][c2] = W3[8 - tap'][ci][co]
        const int j = i - k.o_w3d, tap = j / (c3 * c2), co = (j / c2) % c3, ci = j % c2;
        v = p[k.p[4] + (co * c2 + ci) * 9 + (8 - tap)];
    } else if (i < k.o_wh) { // conv2 data-gradient filter [tap'][c2][c1]
        const int j = i - k.o_w2d, tap = j / (c2 * c1), co = (j / c1) % c2, ci = j % c1;
        v = p[k.p[2] + (co * c1 + ci) * 9 + (8 - tap)];
    } else if (i < k.o_wpT) { // 1x1 heads [c3][8]
        const int j = i - k.o_wh, c = j / 8, q = j % 8;
        v = q < 4 ? p[k.p[6] + q * c3 + c] : (q < 6 ? p[k.p[10] + (q - 4) * c3 + c] : 0.f);
    } else if (i < k.o_wvT) { // policy FC transposed [4S][S]
        const int j = i - k.o_wpT, kk = j / S, o = j % S;
        v = p[k.p[8] + o * 4 * S + kk];
    } else if (i < k.o_bias) { // value FC1 transposed [2S][64]
        const int j = i - k.o_wvT, kk = j / 64, o = j % 64;
        v = p[k.p[12] + o * 2 * S + kk];
    } else {
        const int j = i - k.o_bias;
        if (j < k.b_b2) { if (j < c1) v = p[k.p[1] + j]; }
        else if (j < k.b_b3) { if (j - k.b_b2 < c2) v = p[k.p[3] + j - k.b_b2]; }
        else if (j < k.b_bh) { if (j - k.b_b3 < c3) v = p[k.p[5] + j - k.b_b3]; }
        else if (j < k.b_bp) { const int q = j - k.b_bh; v = q < 4 ? p[k.p[7] + q] : (q < 6 ? p[k.p[11] + q - 4] : 0.f); }
        else if (j < k.b_bv) { if (j - k.b_bp < S) v = p[k.p[9] + j - k.b_bp]; }
        else if (j < k.b_wv2) { if (j - k.b_bv < 64) v = p[k.p[13] + j - k.b_bv]; }
        else if (j < k.b_bv2) { if (j - k.b_wv2 < 64) v = p[k.p[14] + j - k.b_wv2]; }
        else if (j == k.b_bv2) v = p[k.p[15]];
    }
    out[i] = v;
}

// ---- heads: forward, loss terms and the gradient down to the 1x1-conv outputs, one CTA per board --
__device__ __forceinline__ float tr_block_sum(float v, float *red, int tid)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    return (red[0] + red[1]) + (red[2] + red[3]);
}
__device__ __forceinline__ float tr_block_max(float v, float *red, int tid)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    return fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
}

struct TrainHeadArgs {
    const float *a3, *wh, *bh, *wpT, *bp, *wvT, *bv, *wv2, *bv2; // packed layouts
    const float *wp, *wv;                                        // torch layouts [S][4S], [64][2S]
    const float *pi, *z;
    float *h6, *hv, *dlog, *dhv, *dvpre, *dh6, *terms, *probs, *values;
    int B, S, c3, backward;
};

__global__ void __launch_bounds__(128) train_heads_kernel(TrainHeadArgs a)
{
    extern __shared__ float sm[];
    const int S = a.S, c3 = a.c3, tid = threadIdx.x;
    float *s_wh = sm;              // c3*8
    float *s_h = s_wh + c3 * 8;    // 6S  post-ReLU head planes ([4][S] policy, [2][S] value)
    float *s_l = s_h + 6 * S;      // S   logits -> log-probs
    float *s_d = s_l + S;          // S   dL/dlogits
    float *s_v = s_d + S;          // 64  post-ReLU value hidden
    float *s_dv = s_v + 64;        // 64
    float *s_red = s_dv + 64;      // 8
    const long long b = blockIdx.x;
    for (int i = tid; i < c3 * 8; i += blockDim.x) s_wh[i] = a.wh[i];
    __syncthreads();
    for (int cell = tid; cell < S; cell += blockDim.x) {
        float acc[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) acc[j] = a.bh[j];
        const float4 *x = reinterpret_cast<const float4 *>(a.a3 + (b * S + cell) * c3);
        for (int c4 = 0; c4 < c3 / 4; ++c4) {
            const float4 v = x[c4];
            const float av[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float *wr = s_wh + (c4 * 4 + q) * 8;
#pragma unroll
                for (int j = 0; j < 6; ++j) acc[j] = fmaf(av[q], wr[j], acc[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < 6; ++j) s_h[j * S + cell] = fmaxf(acc[j], 0.f);
    }
    __syncthreads();
    for (int o = tid; o < S; o += blockDim.x) {
        float acc = a.bp[o];
        for (int k = 0; k < 4 * S; ++k) acc = fmaf(s_h[k], a.wpT[(size_t)k * S + o], acc);
        s_l[o] = acc;
    }
    if (tid < 64) {
        float acc = a.bv[tid];
        const float *hv = s_h + 4 * S;
        for (int k = 0; k < 2 * S; ++k) acc = fmaf(hv[k], a.wvT[(size_t)k * 64 + tid], acc);
        s_v[tid] = fmaxf(acc, 0.f);
    }
    __syncthreads();
    // log_softmax (:51): x - max - log(sum(exp(x - max)))
    float mx = -3.0e38f;
    for (int o = tid; o < S; o += blockDim.x) mx = fmaxf(mx, s_l[o]);
    mx = tr_block_max(mx, s_red, tid);
    float sum = 0.f;
    for (int o = tid; o < S; o += blockDim.x) sum += expf(s_l[o] - mx);
    sum = tr_block_sum(sum, s_red, tid);
    const float lse = mx + logf(sum);
    float vacc = tid < 64 ? s_v[tid] * a.wv2[tid] : 0.f;
    vacc = tr_block_sum(vacc, s_red, tid);
    const float val = tanhf(vacc + a.bv2[0]);
    if (!a.backward) {
        for (int o = tid; o < S; o += blockDim.x) a.probs[b * S + o] = expf(s_l[o] - lse);
        if (tid == 0) a.values[b] = val;
        return;
    }
    // loss terms of this board and dL/dlogits
    float pl = 0.f, ent = 0.f, spi = 0.f;
    for (int o = tid; o < S; o += blockDim.x) {
        const float lp = s_l[o] - lse, pr = expf(lp), t = a.pi[b * S + o];
        s_l[o] = lp;
        pl -= t * lp;
        ent -= pr * lp;
        spi += t;
    }
    pl = tr_block_sum(pl, s_red, tid);
    ent = tr_block_sum(ent, s_red, tid);
    spi = tr_block_sum(spi, s_red, tid);
    const float invB = 1.f / (float)a.B, zb = a.z[b];
    const float dvp = 2.f * (val - zb) * invB * (1.f - val * val);
    for (int o = tid; o < S; o += blockDim.x) {
        const float d = (expf(s_l[o]) * spi - a.pi[b * S + o]) * invB;
        s_d[o] = d;
        a.dlog[b * S + o] = d;
    }
    if (tid < 64) {
        const float d = s_v[tid] > 0.f ? dvp * a.wv2[tid] : 0.f;
        s_dv[tid] = d;
        a.dhv[b * 64 + tid] = d;
        a.hv[b * 64 + tid] = s_v[tid];
    }
    if (tid == 0) {
        a.dvpre[b] = dvp;
        a.terms[b * 3 + 0] = (val - zb) * (val - zb);
        a.terms[b * 3 + 1] = pl;
        a.terms[b * 3 + 2] = ent;
    }
    __syncthreads();
    // gradient w.r.t. the (pre-ReLU) 1x1 conv outputs
    for (int k = tid; k < 4 * S; k += blockDim.x) {
        float acc = 0.f;
        for (int o = 0; o < S; ++o) acc = fmaf(s_d[o], a.wp[(size_t)o * 4 * S + k], acc);
        a.dh6[b * 6 * S + k] = s_h[k] > 0.f ? acc : 0.f;
        a.h6[b * 6 * S + k] = s_h[k];
    }
    for (int k = tid; k < 2 * S; k += blockDim.x) {
        float acc = 0.f;
        for (int j = 0; j < 64; ++j) acc = fmaf(s_dv[j], a.wv[(size_t)j * 2 * S + k], acc);
        a.dh6[b * 6 * S + 4 * S + k] = s_h[4 * S + k] > 0.f ? acc : 0.f;
        a.h6[b * 6 * S + 4 * S + k] = s_h[4 * S + k];
    }
}

// loss = mean(vl) + mean(pl), entropy = mean(ent) (:139-146)
__global__ void __launch_bounds__(128) train_loss_kernel(const float *__restrict__ terms, int B, float *__restrict__ out2)
{
    __shared__ float red[8];
    float vl = 0.f, pl = 0.f, en = 0.f;
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        vl += terms[b * 3];
        pl += terms[b * 3 + 1];
        en += terms[b * 3 + 2];
    }
    vl = tr_block_sum(vl, red, threadIdx.x);
    pl = tr_block_sum(pl, red, threadIdx.x);
    en = tr_block_sum(en, red, threadIdx.x);
    if (threadIdx.x == 0) {
        out2[0] = vl / (float)B + pl / (float)B;
        out2[1] = en / (float)B;
    }
}

// ---- FC weight gradients: one thread per (parameter, batch chunk), fixed-order sums ----------------
// parameter index i enumerates act_fc1.weight [S][4S], act_fc1.bias, val_fc1.weight [64][2S],
// val_fc1.bias, val_fc2.weight [64], val_fc2.bias; fc_map gives the two factor streams and the
// position in the flat gradient vector.
struct FcMap {
    const float *x, *d;
    int xs, ds, dst;
};
__device__ __forceinline__ FcMap train_fc_map(int i, const TrainPack &k, const float *h6, const float *hv,
                                              const float *dlog, const float *dhv, const float *dvpre)
{
    const int S = k.S, n_wp = 4 * S * S, n_wv = 128 * S;
    FcMap f;
    f.x = nullptr; f.xs = 0; f.dst = -1; f.d = nullptr; f.ds = 0;
    if (i < n_wp) { const int o = i / (4 * S), kk = i % (4 * S); f.x = h6 + kk; f.xs = 6 * S; f.d = dlog + o; f.ds = S; f.dst = k.p[8] + i; }
    else if (i < n_wp + S) { const int o = i - n_wp; f.d = dlog + o; f.ds = S; f.dst = k.p[9] + o; }
    else if (i < n_wp + S + n_wv) { const int j = i - n_wp - S, o = j / (2 * S), kk = j % (2 * S); f.x = h6 + 4 * S + kk; f.xs = 6 * S; f.d = dhv + o; f.ds = 64; f.dst = k.p[12] + j; }
    else if (i < n_wp + S + n_wv + 64) { const int o = i - n_wp - S - n_wv; f.d = dhv + o; f.ds = 64; f.dst = k.p[13] + o; }
    else if (i < n_wp + S + n_wv + 128) { const int o = i - n_wp - S - n_wv - 64; f.x = hv + o; f.xs = 64; f.d = dvpre; f.ds = 1; f.dst = k.p[14] + o; }
    else if (i == n_wp + S + n_wv + 128) { f.d = dvpre; f.ds = 1; f.dst = k.p[15]; }
    return f;
}

// part [nchunk][n_fc]
__global__ void __launch_bounds__(256)
train_fc_wgrad_kernel(const float *__restrict__ h6, const float *__restrict__ hv, const float *__restrict__ dlog,
                      const float *__restrict__ dhv, const float *__restrict__ dvpre, float *__restrict__ part,
                      TrainPack k, int B, int chunk, int n_fc)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_fc) return;
    const FcMap f = train_fc_map(i, k, h6, hv, dlog, dhv, dvpre);
    const int b0 = blockIdx.y * chunk, b1 = min(B, b0 + chunk);
    float acc = 0.f;
    if (f.x) {
#pragma unroll 8
        for (int b = b0; b < b1; ++b) acc = fmaf(f.x[(size_t)b * f.xs], f.d[(size_t)b * f.ds], acc);
    } else {
#pragma unroll 8
        for (int b = b0; b < b1; ++b) acc += f.d[(size_t)b * f.ds];
    }
    part[(size_t)blockIdx.y * n_fc + i] = acc;
}

// ---- 1x1 head convs backward: g3 = mask(a3) * (dh6 . wh^T), partial sums of dWh and dbh ----------
// part [nblk][7][c3]: rows 0..5 = sum_m a3[m][c] * dh6[m][j], row 6 cols 0..5 = sum_m dh6[m][j]
__global__ void __launch_bounds__(128)
train_head_dgrad_kernel(const float *__restrict__ a3, const float *__restrict__ dh6, const float *__restrict__ wh,
                        float *__restrict__ g3, float *__restrict__ part, long long M, int S, int c3, int R)
{
    const long long m0 = (long long)blockIdx.x * R;
    for (int c = threadIdx.x; c < c3; c += blockDim.x) {
        float w[6], acc[6], bs = 0.f;
#pragma unroll
        for (int j = 0; j < 6; ++j) { w[j] = wh[c * 8 + j]; acc[j] = 0.f; }
        for (int r = 0; r < R; ++r) {
            const long long m = m0 + r;
            if (m >= M) break;
            const long long b = m / S;
            const int cell = (int)(m % S);
            const float *dp = dh6 + b * 6 * S + cell;
            float d[6], s = 0.f;
#pragma unroll
            for (int j = 0; j < 6; ++j) { d[j] = dp[j * S]; s = fmaf(d[j], w[j], s); }
            const float x = a3[m * c3 + c];
            g3[m * c3 + c] = x > 0.f ? s : 0.f;
#pragma unroll
            for (int j = 0; j < 6; ++j) acc[j] = fmaf(x, d[j], acc[j]);
            if (c < 6) bs += dp[c * S];
        }
#pragma unroll
        for (int j = 0; j < 6; ++j) part[((size_t)blockIdx.x * 7 + j) * c3 + c] = acc[j];
        part[((size_t)blockIdx.x * 7 + 6) * c3 + c] = bs;
    }
}

// ---- 3x3 conv weight gradient: dW[tap][ci][co] = sum_m in[m + tap][ci] * gout[m][co] --------------
// grid (9 taps, nsplit); each CTA owns a row range and writes one partial [CIN][COUT] tile for its
// tap (+ the bias partial from the centre tap).  part [nsplit][9*CIN*COUT + COUT].
template <int CIN, int COUT, int TM>
__global__ void __launch_bounds__((CIN / TM) * (COUT / 8))
train_conv_wgrad_kernel(const float *__restrict__ in, const float *__restrict__ gout, float *__restrict__ part,
                        long long M, int rows_per_split, int S, int W, int H)
{
    constexpr int KC = 32, TXN = COUT / 8, NT = (CIN / TM) * TXN;
    __shared__ __align__(16) float As[KC][CIN];
    __shared__ __align__(16) float Gs[KC][COUT];
    const int tap = blockIdx.x, dy = tap / 3 - 1, dx = tap % 3 - 1;
    const long long m_begin = (long long)blockIdx.y * rows_per_split;
    const long long m_end = (m_begin + rows_per_split < M) ? m_begin + rows_per_split : M;
    const int tid = threadIdx.x, tx = tid % TXN, ty = tid / TXN;
    const bool do_bias = (tap == 4 && ty == 0);
    float acc[TM][8], bacc[8];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) bacc[j] = 0.f;
    for (long long mc = m_begin; mc < m_end; mc += KC) {
        for (int i = tid; i < KC * CIN / 4; i += NT) {
            const int r = i / (CIN / 4), c4 = i % (CIN / 4);
            const long long m = mc + r;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (m < m_end) {
                const int cell = (int)(m % S), yy = cell / W + dy, xx = cell % W + dx;
                if (yy >= 0 && yy < H && xx >= 0 && xx < W)
                    v = *reinterpret_cast<const float4 *>(in + (m + dy * W + dx) * CIN + c4 * 4);
            }
            *reinterpret_cast<float4 *>(&As[r][c4 * 4]) = v;
        }
        for (int i = tid; i < KC * COUT / 4; i += NT) {
            const int r = i / (COUT / 4), c4 = i % (COUT / 4);
            const long long m = mc + r;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (m < m_end) v = *reinterpret_cast<const float4 *>(gout + m * COUT + c4 * 4);
            *reinterpret_cast<float4 *>(&Gs[r][c4 * 4]) = v;
        }
        __syncthreads();
#pragma unroll 4
        for (int k = 0; k < KC; ++k) {
            const float4 b0 = *reinterpret_cast<const float4 *>(&Gs[k][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4 *>(&Gs[k][COUT / 2 + tx * 4]);
            const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < TM; ++i) {
                const float av = As[k][ty * TM + i];
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av, bv[j], acc[i][j]);
            }
            if (do_bias) {
#pragma unroll
                for (int j = 0; j < 8; ++j) bacc[j] += bv[j];
            }
        }
        __syncthreads();
    }
    float *o = part + (size_t)blockIdx.y * (9 * CIN * COUT + COUT);
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        float *row = o + ((size_t)tap * CIN + ty * TM + i) * COUT;
        *reinterpret_cast<float4 *>(row + tx * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        *reinterpret_cast<float4 *>(row + COUT / 2 + tx * 4) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
    }
    if (do_bias) {
        float *row = o + (size_t)9 * CIN * COUT;
        *reinterpret_cast<float4 *>(row + tx * 4) = make_float4(bacc[0], bacc[1], bacc[2], bacc[3]);
        *reinterpret_cast<float4 *>(row + COUT / 2 + tx * 4) = make_float4(bacc[4], bacc[5], bacc[6], bacc[7]);
    }
}

// conv1: the input is the NCHW plane stack (4 channels).  Each CTA owns a range of boards, stages one
// board's planes zero-padded in shared memory, and every thread (ci = warp, co = lane) accumulates its
// 9 taps; part [nsplit][36*c1 + c1], blockDim = 4*c1.
__global__ void __launch_bounds__(128)
train_conv1_wgrad_kernel(const float *__restrict__ planes, const float *__restrict__ gout, float *__restrict__ part,
                         int B, int boards_per_split, int S, int W, int H, int c1)
{
    extern __shared__ float sp[]; // [4][(H+2)*(W+2)]
    const int PW = W + 2, PS = (H + 2) * PW;
    const int b0 = blockIdx.x * boards_per_split, b1 = min(B, b0 + boards_per_split);
    const int co = threadIdx.x % c1, ci = threadIdx.x / c1;
    float acc[9], bacc = 0.f;
#pragma unroll
    for (int t = 0; t < 9; ++t) acc[t] = 0.f;
    for (int i = threadIdx.x; i < 4 * PS; i += blockDim.x) sp[i] = 0.f;
    for (int b = b0; b < b1; ++b) {
        __syncthreads();
        for (int i = threadIdx.x; i < 4 * S; i += blockDim.x) {
            const int c = i / S, cell = i % S;
            sp[c * PS + (cell / W + 1) * PW + cell % W + 1] = planes[(size_t)b * 4 * S + i];
        }
        __syncthreads();
        const float *g = gout + (size_t)b * S * c1 + co;
        const float *win = sp + ci * PS;
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                const float gv = g[(y * W + x) * c1];
                bacc += gv;
                const float *q = win + y * PW + x;
#pragma unroll
                for (int t = 0; t < 9; ++t) acc[t] = fmaf(q[(t / 3) * PW + t % 3], gv, acc[t]);
            }
    }
    float *o = part + (size_t)blockIdx.x * (36 * c1 + c1);
#pragma unroll
    for (int t = 0; t < 9; ++t) o[(t * 4 + ci) * c1 + co] = acc[t];
    if (ci == 0) o[36 * c1 + co] = bacc;
}

// ---- fixed-order reduction of the split partials into the flat gradient vector --------------------
// mode 0: conv partial [9][Cin][Cout] + [Cout]  -> OIHW weight at offW, bias at offB
// mode 1: head partial [7][c3] (see train_head_dgrad_kernel) -> act_conv1 / val_conv1 weights + biases
// mode 2: FC partial [n_fc] -> train_fc_map destinations
struct TrainReduce {
    int mode, n, nsplit, Cin, Cout, offW, offB, offW2, offB2;
};
__global__ void __launch_bounds__(256)
train_reduce_kernel(const float *__restrict__ part, float *__restrict__ g, TrainReduce r, TrainPack k)
{
    __shared__ float s[8][33];
    const int li = threadIdx.x & 31, ls = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + li;
    float acc = 0.f;
    if (i < r.n)
        for (int sp = ls; sp < r.nsplit; sp += 8) acc += part[(size_t)sp * r.n + i];
    s[ls][li] = acc;
    __syncthreads();
    if (ls != 0 || i >= r.n) return;
    float tot = ((s[0][li] + s[1][li]) + (s[2][li] + s[3][li])) + ((s[4][li] + s[5][li]) + (s[6][li] + s[7][li]));
    int dst = -1;
    if (r.mode == 0) {
        const int nw = 9 * r.Cin * r.Cout;
        if (i < nw) {
            const int tap = i / (r.Cin * r.Cout), ci = (i / r.Cout) % r.Cin, co = i % r.Cout;
            dst = r.offW + (co * r.Cin + ci) * 9 + tap;
        } else
            dst = r.offB + (i - nw);
    } else if (r.mode == 2) {
        dst = train_fc_map(i, k, nullptr, nullptr, nullptr, nullptr, nullptr).dst;
    } else {
        const int c3 = r.Cin, j = i / c3, c = i % c3;
        if (j < 4) dst = r.offW + j * c3 + c;
        else if (j < 6) dst = r.offW2 + (j - 4) * c3 + c;
        else if (c < 4) dst = r.offB + c;
        else if (c < 6) dst = r.offB2 + (c - 4);
    }
    if (dst >= 0) g[dst] = tot;
}

// ---- Adam with L2 weight decay, torch.optim.Adam semantics (single tensor path) -------------------
__global__ void __launch_bounds__(256)
train_adam_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v,
                  int n, float wd, float beta1, float beta2, float step_size, float sqrt_bc2, float eps)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float w = p[i];
    const float gr = fmaf(wd, w, g[i]);            // grad.add(param, alpha=weight_decay)
    const float mi = m[i] + (gr - m[i]) * (1.f - beta1); // exp_avg.lerp_(grad, 1 - beta1)
    const float vi = fmaf(1.f - beta2, gr * gr, v[i] * beta2);
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / sqrt_bc2 + eps;
    p[i] = w - step_size * (mi / denom);
}

// ---- host side ---------------------------------------------------------------------------------
static void gz_train_free(gz_engine *e)
{
    TrainState *t = e->tr;
    if (!t) return;
    float *ptrs[] = {t->p, t->g, t->m, t->v, t->packed, t->a1, t->a2, t->a3, t->g1, t->g2, t->g3, t->h6, t->hv,
                     t->dlog, t->dhv, t->dvpre, t->dh6, t->terms, t->probs, t->values, t->part, t->out2};
    for (float *q : ptrs)
        if (q) cudaFree(q);
    delete t;
    e->tr = nullptr;
}

static void gz_train_invalidate(gz_engine *e)
{
    if (!e->tr) return;
    e->tr->init = false;
    e->tr->dirty = false;
    e->tr->have_grads = false;
}

static bool gz_train_supported(const gz_engine *e)
{
    return e->c1 == 32 && e->c2 == 64 && e->c3 == 128;
}

// creates the state (first use) and (re)initialises the master vector from the engine's weights
static int gz_train_ensure(gz_engine *e, int64_t B)
{
    if (!e->wts.loaded) { e->err = "gz_set_weights has not been called"; return GZ_E_NOWEIGHTS; }
    if (!gz_train_supported(e)) { e->err = "the device train step needs the reference trunk 32/64/128"; return GZ_E_INVALID; }
    if (!e->tr) e->tr = new TrainState();
    if (e->n_sm <= 0) {
        int dev = 0;
        GZ_CUDA(e, cudaGetDevice(&dev));
        GZ_CUDA(e, cudaDeviceGetAttribute(&e->n_sm, cudaDevAttrMultiProcessorCount, dev));
    }
    TrainState *t = e->tr;
    const int S = e->S, c1 = e->c1, c2 = e->c2, c3 = e->c3;
    if (!t->p) {
        TrainPack &k = t->pk;
        k.S = S; k.c1 = c1; k.c2 = c2; k.c3 = c3;
        size_t n[16];
        gz_weight_sizes(e, n);
        k.p[0] = 0;
        for (int i = 0; i < 16; ++i) k.p[i + 1] = k.p[i] + (int)n[i];
        auto al = [](int x) { return (x + 3) & ~3; };
        int off = 0;
        k.b_b1 = off; off = al(off + c1);
        k.b_b2 = off; off = al(off + c2);
        k.b_b3 = off; off = al(off + c3);
        k.b_bh = off; off = al(off + 8);
        k.b_bp = off; off = al(off + S);
        k.b_bv = off; off = al(off + 64);
        k.b_wv2 = off; off = al(off + 64);
        k.b_bv2 = off; off = al(off + 1);
        k.b_total = off;
        off = 0;
        k.o_w1t = off; off += 36 * c1;
        k.o_w2t = off; off += 9 * c1 * c2;
        k.o_w3t = off; off += 9 * c2 * c3;
        k.o_w3d = off; off += 9 * c2 * c3;
        k.o_w2d = off; off += 9 * c1 * c2;
        k.o_wh = off; off += c3 * 8;
        k.o_wpT = off; off += 4 * S * S;
        k.o_wvT = off; off += 2 * S * 64;
        k.o_bias = off; off += k.b_total;
        k.total = off;
        const size_t np = (size_t)k.p[16];
        GZ_CUDA(e, cudaMalloc(&t->p, np * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->g, np * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->m, np * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->v, np * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->packed, (size_t)k.total * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->out2, 2 * sizeof(float)));
    }
    if (!t->init) {
        const size_t np = (size_t)t->pk.p[16];
        for (int i = 0; i < 16; ++i)
            GZ_CUDA(e, cudaMemcpyAsync(t->p + t->pk.p[i], e->wts.raw[i], e->wts.raw_elems[i] * sizeof(float),
                                       cudaMemcpyDeviceToDevice, e->stream));
        GZ_CUDA(e, cudaMemsetAsync(t->g, 0, np * sizeof(float), e->stream));
        GZ_CUDA(e, cudaMemsetAsync(t->m, 0, np * sizeof(float), e->stream));
        GZ_CUDA(e, cudaMemsetAsync(t->v, 0, np * sizeof(float), e->stream));
        t->t = 0;
        t->init = true;
        t->dirty = false;
        t->have_grads = false;
    }
    if (B > t->cap) {
        float **bufs[] = {&t->a1, &t->a2, &t->a3, &t->g1, &t->g2, &t->g3, &t->h6, &t->hv, &t->dlog, &t->dhv,
                          &t->dvpre, &t->dh6, &t->terms, &t->probs, &t->values, &t->part};
        GZ_CUDA(e, cudaStreamSynchronize(e->stream));
        for (float **q : bufs)
            if (*q) { cudaFree(*q); *q = nullptr; }
        t->cap = 0;
        const size_t rows = (size_t)B * S, b = (size_t)B;
        GZ_CUDA(e, cudaMalloc(&t->a1, rows * c1 * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->a2, rows * c2 * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->a3, rows * c3 * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->g1, rows * c1 * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->g2, rows * c2 * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->g3, rows * c3 * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->h6, rows * 6 * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->dh6, rows * 6 * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->hv, b * 64 * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->dhv, b * 64 * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->dlog, rows * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->probs, rows * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->dvpre, b * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->values, b * sizeof(float)));
        GZ_CUDA(e, cudaMalloc(&t->terms, b * 3 * sizeof(float)));
        // split-K partials: the largest user is conv3 (GZ_TR_SPLITS x (9*c2*c3 + c3)) or the head
        // reduction (ceil(rows / 32) x 7*c3)
        const size_t n_head = ((rows + 31) / 32) * 7 * c3;
        const size_t n_conv = (size_t)64 * (9 * c2 * c3 + c3);
        const size_t n_fcp = (size_t)16 * (4 * S * S + S + 128 * S + 129);
        const size_t n_c1 = (size_t)(2 * e->n_sm + 1) * (37 * c1);
        t->part_n = n_head > n_conv ? n_head : n_conv;
        if (n_fcp > t->part_n) t->part_n = n_fcp;
        if (n_c1 > t->part_n) t->part_n = n_c1;
        GZ_CUDA(e, cudaMalloc(&t->part, t->part_n * sizeof(float)));
        t->cap = B;
    }
    return GZ_OK;
}

static int gz_train_forward_pass(gz_engine *e, const float *d_states, int64_t B)
{
    TrainState *t = e->tr;
    const TrainPack &k = t->pk;
    train_pack_kernel<<<(k.total + 255) / 256, 256, 0, e->stream>>>(t->p, t->packed, k);
    NetDims d = {e->W, e->H, e->S, e->c1, e->c2, e->c3};
    const float *bias = t->packed + k.o_bias;
    const long long threads = (long long)B * e->S * (e->c1 / 8);
    conv1_kernel<<<(unsigned)((threads + 255) / 256), 256, (size_t)(37 * e->c1) * sizeof(float), e->stream>>>(
        d_states, t->packed + k.o_w1t, bias + k.b_b1, t->a1, nullptr, B, d);
    const unsigned gx = (unsigned)(((long long)B * e->S + IG_BM - 1) / IG_BM);
    conv3x3_igemm_kernel<64, 0><<<dim3(gx, e->c2 / 64), 128, 0, e->stream>>>(
        t->a1, t->packed + k.o_w2t, bias + k.b_b2, t->a2, nullptr, B, e->S, e->W, e->H, e->c1, e->c2);
    conv3x3_igemm_kernel<128, 0><<<dim3(gx, e->c3 / 128), 256, 0, e->stream>>>(
        t->a2, t->packed + k.o_w3t, bias + k.b_b3, t->a3, nullptr, B, e->S, e->W, e->H, e->c2, e->c3);
    e->launches += 4;
    return GZ_OK;
}

static void gz_train_head_args(gz_engine *e, TrainHeadArgs &a, const float *d_pi, const float *d_z, int64_t B,
                               int backward)
{
    TrainState *t = e->tr;
    const TrainPack &k = t->pk;
    const float *bias = t->packed + k.o_bias;
    a.a3 = t->a3; a.wh = t->packed + k.o_wh; a.bh = bias + k.b_bh; a.wpT = t->packed + k.o_wpT; a.bp = bias + k.b_bp;
    a.wvT = t->packed + k.o_wvT; a.bv = bias + k.b_bv; a.wv2 = bias + k.b_wv2; a.bv2 = bias + k.b_bv2;
    a.wp = t->p + k.p[8]; a.wv = t->p + k.p[12];
    a.pi = d_pi; a.z = d_z;
    a.h6 = t->h6; a.hv = t->hv; a.dlog = t->dlog; a.dhv = t->dhv; a.dvpre = t->dvpre; a.dh6 = t->dh6;
    a.terms = t->terms; a.probs = t->probs; a.values = t->values;
    a.B = (int)B; a.S = e->S; a.c3 = e->c3; a.backward = backward;
}

// forward + backward of one batch; leaves the gradient in tr->g and (loss, entropy) in tr->out2
static int gz_train_backward_pass(gz_engine *e, const float *d_states, const float *d_pi, const float *d_z, int64_t B)
{
    TrainState *t = e->tr;
    const TrainPack &k = t->pk;
    const int S = e->S, W = e->W, H = e->H, c1 = e->c1, c2 = e->c2, c3 = e->c3;
    const long long M = (long long)B * S;
    GZ_TRY(gz_train_forward_pass(e, d_states, B));
    TrainHeadArgs ha;
    gz_train_head_args(e, ha, d_pi, d_z, B, 1);
    const size_t hsm = (size_t)(c3 * 8 + 8 * S + 128 + 8) * sizeof(float);
    train_heads_kernel<<<(unsigned)B, 128, hsm, e->stream>>>(ha);
    train_loss_kernel<<<1, 128, 0, e->stream>>>(t->terms, (int)B, t->out2);
    const int n_fc = 4 * S * S + S + 128 * S + 64 + 64 + 1;
    TrainReduce r;
    {
        int chunk = (int)((B + 15) / 16);
        if (chunk < 32) chunk = 32;
        const int nchunk = (int)((B + chunk - 1) / chunk);
        train_fc_wgrad_kernel<<<dim3((n_fc + 255) / 256, nchunk), 256, 0, e->stream>>>(
            t->h6, t->hv, t->dlog, t->dhv, t->dvpre, t->part, k, (int)B, chunk, n_fc);
        r.mode = 2; r.n = n_fc; r.nsplit = nchunk; r.Cin = r.Cout = r.offW = r.offB = r.offW2 = r.offB2 = 0;
        train_reduce_kernel<<<(r.n + 31) / 32, 256, 0, e->stream>>>(t->part, t->g, r, k);
    }
    // 1x1 head convs
    const int R = 32, nblk = (int)((M + R - 1) / R);
    train_head_dgrad_kernel<<<nblk, 128, 0, e->stream>>>(t->a3, t->dh6, t->packed + k.o_wh, t->g3, t->part, M, S, c3, R);
    r.mode = 1; r.n = 7 * c3; r.nsplit = nblk; r.Cin = c3; r.Cout = 0;
    r.offW = k.p[6]; r.offB = k.p[7]; r.offW2 = k.p[10]; r.offB2 = k.p[11];
    train_reduce_kernel<<<(r.n + 31) / 32, 256, 0, e->stream>>>(t->part, t->g, r, k);
    // conv3: weight gradient, then the data gradient masked by ReLU(conv2)
    int nsplit = (int)((M + 511) / 512);
    if (nsplit > 64) nsplit = 64;
    if (nsplit < 1) nsplit = 1;
    int rows = (int)((M + nsplit - 1) / nsplit);
    rows = (rows + 31) & ~31;
    nsplit = (int)((M + rows - 1) / rows);
    train_conv_wgrad_kernel<64, 128, 4><<<dim3(9, nsplit), 256, 0, e->stream>>>(t->a2, t->g3, t->part, M, rows, S, W, H);
    r.mode = 0; r.Cin = c2; r.Cout = c3; r.n = 9 * c2 * c3 + c3; r.nsplit = nsplit; r.offW = k.p[4]; r.offB = k.p[5];
    train_reduce_kernel<<<(r.n + 31) / 32, 256, 0, e->stream>>>(t->part, t->g, r, k);
    const unsigned gx = (unsigned)((M + IG_BM - 1) / IG_BM);
    conv3x3_igemm_kernel<64, 1><<<dim3(gx, 1), 128, 0, e->stream>>>(
        t->g3, t->packed + k.o_w3d, t->a2, t->g2, nullptr, B, S, W, H, c3, c2);
    // conv2
    train_conv_wgrad_kernel<32, 64, 2><<<dim3(9, nsplit), 128, 0, e->stream>>>(t->a1, t->g2, t->part, M, rows, S, W, H);
    r.Cin = c1; r.Cout = c2; r.n = 9 * c1 * c2 + c2; r.offW = k.p[2]; r.offB = k.p[3];
    train_reduce_kernel<<<(r.n + 31) / 32, 256, 0, e->stream>>>(t->part, t->g, r, k);
    conv3x3_igemm_kernel<32, 1><<<dim3(gx, 1), 64, 0, e->stream>>>(
        t->g2, t->packed + k.o_w2d, t->a1, t->g1, nullptr, B, S, W, H, c2, c1);
    // conv1 (no data gradient: the input is the position)
    const int bps = (int)((B + 2 * e->n_sm - 1) / (2 * e->n_sm));
    const int ns1 = (int)((B + bps - 1) / bps);
    train_conv1_wgrad_kernel<<<ns1, 4 * c1, (size_t)4 * (H + 2) * (W + 2) * sizeof(float), e->stream>>>(
        d_states, t->g1, t->part, (int)B, bps, S, W, H, c1);
    r.Cin = 4; r.Cout = c1; r.n = 36 * c1 + c1; r.nsplit = ns1; r.offW = k.p[0]; r.offB = k.p[1];
    train_reduce_kernel<<<(r.n + 31) / 32, 256, 0, e->stream>>>(t->part, t->g, r, k);
    e->launches += 14;
    GZ_CUDA(e, cudaGetLastError());
    t->have_grads = true;
    return GZ_OK;
}

static int gz_train_apply_pass(gz_engine *e, const float *d_grads, float lr)
{
    TrainState *t = e->tr;
    const int np = t->pk.p[16];
    t->t += 1;
    const double b1 = 0.9, b2 = 0.999;
    const double bc1 = 1.0 - pow(b1, (double)t->t), bc2 = 1.0 - pow(b2, (double)t->t);
    train_adam_kernel<<<(np + 255) / 256, 256, 0, e->stream>>>(
        t->p, d_grads ? d_grads : t->g, t->m, t->v, np, 1e-4f, (float)b1, (float)b2, (float)((double)lr / bc1),
        (float)sqrt(bc2), 1e-8f);
    e->launches += 1;
    GZ_CUDA(e, cudaGetLastError());
    t->dirty = true;
    return GZ_OK;
}

// pushes the trained master vector back into the engine's inference images (tcgen05 and fp32 paths)
static int gz_train_flush(gz_engine *e)
{
    TrainState *t = e->tr;
    if (!t || !t->init || !t->dirty) return GZ_OK;
    size_t n[16];
    gz_weight_sizes(e, n);
    std::vector<float> flat((size_t)t->pk.p[16]);
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    GZ_CUDA(e, cudaMemcpy(flat.data(), t->p, flat.size() * sizeof(float), cudaMemcpyDeviceToHost));
    std::vector<float> h[16];
    for (int i = 0; i < 16; ++i) {
        h[i].assign(flat.begin() + t->pk.p[i], flat.begin() + t->pk.p[i + 1]);
        GZ_CUDA(e, cudaMemcpy(e->wts.raw[i], t->p + t->pk.p[i], n[i] * sizeof(float), cudaMemcpyDeviceToDevice));
    }
    GZ_TRY(gz_net_prepare_weights(e, h));
    t->dirty = false;
    return GZ_OK;
}

extern "C" int gz_train_backward(gz_engine *e, const float *states, const float *pis, const float *zs, int64_t n,
                                 float *loss, float *entropy)
{
    if (!e || !states || !pis || !zs || n <= 0) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_TRY(gz_train_ensure(e, n));
    const size_t S = e->S, N = (size_t)n;
    ArgStage a(e);
    GZ_TRY(a.reserve(N * S * 20 + N * 4 + 4096));
    const float *ds = a.in(states, N * 4 * S);
    const float *dp = a.in(pis, N * S);
    const float *dz = a.in(zs, N);
    if (a.rc) return a.rc;
    ProfScope ps(e, PC_OTHER);
    GZ_TRY(gz_train_backward_pass(e, ds, dp, dz, n));
    float h2[2] = {0.f, 0.f};
    if (loss || entropy) {
        GZ_CUDA(e, cudaMemcpyAsync(h2, e->tr->out2, sizeof(h2), cudaMemcpyDeviceToHost, e->stream));
        GZ_CUDA(e, cudaStreamSynchronize(e->stream));
        if (loss) *loss = h2[0];
        if (entropy) *entropy = h2[1];
    }
    return GZ_OK;
}

extern "C" int gz_train_apply(gz_engine *e, const float *grads, float lr)
{
    if (!e) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    if (grads) GZ_TRY(gz_train_ensure(e, 0)); // a caller-supplied gradient needs no backward pass of ours
    if (!e->tr || !e->tr->init || (!grads && !e->tr->have_grads)) { e->err = "gz_train_apply: no gradient (call gz_train_backward first)"; return GZ_E_INVALID; }
    ArgStage a(e);
    GZ_TRY(a.reserve((size_t)e->tr->pk.p[16] * 4 + 4096));
    const float *dg = a.in(grads, (size_t)e->tr->pk.p[16]);
    if (a.rc) return a.rc;
    return gz_train_apply_pass(e, dg, lr);
}

extern "C" int gz_train_step(gz_engine *e, const float *states, const float *pis, const float *zs, int64_t n,
                             float lr, float *loss, float *entropy)
{
    if (!e || !states || !pis || !zs || n <= 0) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_TRY(gz_train_ensure(e, n));
    const size_t S = e->S, N = (size_t)n;
    ArgStage a(e);
    GZ_TRY(a.reserve(N * S * 20 + N * 4 + 4096));
    const float *ds = a.in(states, N * 4 * S);
    const float *dp = a.in(pis, N * S);
    const float *dz = a.in(zs, N);
    if (a.rc) return a.rc;
    {
        ProfScope ps(e, PC_OTHER);
        GZ_TRY(gz_train_backward_pass(e, ds, dp, dz, n));
        GZ_TRY(gz_train_apply_pass(e, nullptr, lr));
    }
    float h2[2] = {0.f, 0.f};
    if (loss || entropy) {
        GZ_CUDA(e, cudaMemcpyAsync(h2, e->tr->out2, sizeof(h2), cudaMemcpyDeviceToHost, e->stream));
        GZ_CUDA(e, cudaStreamSynchronize(e->stream));
        if (loss) *loss = h2[0];
        if (entropy) *entropy = h2[1];
    }
    return GZ_OK;
}

extern "C" int gz_train_grads(gz_engine *e, float *out, int64_t capacity, int64_t *n_params)
{
    if (!e) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    size_t n[16], np = 0;
    gz_weight_sizes(e, n);
    for (int i = 0; i < 16; ++i) np += n[i];
    if (n_params) *n_params = (int64_t)np;
    if (!out) return GZ_OK;
    if (!e->tr || !e->tr->init || !e->tr->have_grads) { e->err = "gz_train_grads: no gradient (call gz_train_backward first)"; return GZ_E_INVALID; }
    if (capacity < (int64_t)np) { e->err = "gz_train_grads: buffer too small"; return GZ_E_INVALID; }
    GZ_CUDA(e, cudaMemcpyAsync(out, e->tr->g, np * sizeof(float), cudaMemcpyDefault, e->stream));
    GZ_CUDA(e, cudaStreamSynchronize(e->stream));
    return GZ_OK;
}

// policy_value (:80-94) with the master (possibly just-updated) weights on the fp32 path: what
// policy_update's KL check calls between epochs (train.py:178-193) without refreshing the inference images
extern "C" int gz_train_forward(gz_engine *e, const float *states, int64_t n, float *probs, float *values)
{
    if (!e || !states || n <= 0) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_TRY(gz_train_ensure(e, n));
    const size_t S = e->S, N = (size_t)n;
    ArgStage a(e);
    GZ_TRY(a.reserve(N * S * 24 + N * 8 + 4096));
    const float *ds = a.in(states, N * 4 * S);
    float *dpr = a.out(probs, N * S);
    float *dv = a.out(values, N);
    if (a.rc) return a.rc;
    GZ_TRY(gz_train_forward_pass(e, ds, n));
    TrainHeadArgs ha;
    gz_train_head_args(e, ha, nullptr, nullptr, n, 0);
    if (dpr) ha.probs = dpr;
    if (dv) ha.values = dv;
    const size_t hsm = (size_t)(e->c3 * 8 + 8 * S + 128 + 8) * sizeof(float);
    train_heads_kernel<<<(unsigned)n, 128, hsm, e->stream>>>(ha);
    e->launches += 1;
    GZ_CUDA(e, cudaGetLastError());
    return a.finish();
}

// forgets the optimiser state (the next train call restarts Adam from the engine's current weights)
extern "C" int gz_train_reset(gz_engine *e)
{
    if (!e) return GZ_E_INVALID;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    GZ_TRY(gz_train_flush(e));
    if (e->tr) e->tr->init = false;
    return GZ_OK;
}
