// gz_net_ffma.cu — PolicyValueNet forward (reference policy_value_net_pytorch.py:25-57),
// fp32 CUDA-core path (net_impl = 1): the numerics anchor for the tcgen05 path.
//   conv1 (4 -> c1) direct; conv2 / conv3 as implicit GEMM (M = boards*S cells, N = Cout,
//   K = 9*Cin) with 128 x BN x 8 smem tiles, 8x8 register tiles and register-prefetch double
//   buffering; heads kernel = both 1x1 convs + FCs + softmax + tanh per board.
// Activations are NHWC fp32 [boards*S][C].  The number of boards is read from device memory
// (`count`), so the leaf batch of a wave never round-trips to the host.
// Included by gz_api.cu (single translation unit).

struct NetDims {
    int W, H, S, c1, c2, c3;
};

// ---- conv1: planes NCHW [B][4][H][W] -> act NHWC [B*S][c1], bias + ReLU --------------------
__global__ void __launch_bounds__(256)
conv1_kernel(const float *__restrict__ planes, const float *__restrict__ w /*[9][4][c1]*/,
             const float *__restrict__ bias, float *__restrict__ out, const int32_t *__restrict__ count,
             long long maxB, NetDims d)
{
    extern __shared__ float s_w1[]; // 36*c1 + c1
    const int nw = 36 * d.c1;
    for (int i = threadIdx.x; i < nw; i += blockDim.x) s_w1[i] = w[i];
    for (int i = threadIdx.x; i < d.c1; i += blockDim.x) s_w1[nw + i] = bias[i];
    __syncthreads();
    const long long B = count ? (long long)*count : maxB;
    const int cg = d.c1 / 8;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long m = t / cg;
    const int c0 = (int)(t % cg) * 8;
    if (m >= B * d.S) return;
    const long long b = m / d.S;
    const int cell = (int)(m % d.S), y = cell / d.W, x = cell % d.W;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = s_w1[nw + c0 + j];
    const float *pin = planes + b * 4 * d.S;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
        const int yy = y + ky - 1;
        if (yy < 0 || yy >= d.H) continue;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
            const int xx = x + kx - 1;
            if (xx < 0 || xx >= d.W) continue;
#pragma unroll
            for (int ci = 0; ci < 4; ++ci) {
                const float a = pin[ci * d.S + yy * d.W + xx];
                const float *wr = s_w1 + ((ky * 3 + kx) * 4 + ci) * d.c1 + c0;
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fmaf(a, wr[j], acc[j]);
            }
        }
    }
    float4 *o = reinterpret_cast<float4 *>(out + m * d.c1 + c0);
    o[0] = make_float4(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f), fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
    o[1] = make_float4(fmaxf(acc[4], 0.f), fmaxf(acc[5], 0.f), fmaxf(acc[6], 0.f), fmaxf(acc[7], 0.f));
}

// ---- conv 3x3 implicit GEMM, bias + ReLU ---------------------------------------------------
#define IG_BM 128
#define IG_BK 8
#define IG_AS (IG_BM + 4) // padded row stride of the transposed A tile

// EPI 0: out = ReLU(acc + bias) (forward).  EPI 1: out = acc where mask > 0, else 0 (data gradient of
// a ReLU layer, gz_train.cu: `bias` then points at the layer's forward activations [M][Cout]).
template <int BN, int EPI = 0>
__global__ void __launch_bounds__(2 * BN)
conv3x3_igemm_kernel(const float *__restrict__ in /*[M][Cin]*/, const float *__restrict__ w /*[9][Cin][Cout]*/,
                     const float *__restrict__ bias, float *__restrict__ out /*[M][Cout]*/,
                     const int32_t *__restrict__ count, long long maxB, int S, int W, int H, int Cin, int Cout)
{
    constexpr int NT = 2 * BN;        // threads
    constexpr int TXN = BN / 8;       // threads along N
    constexpr int AROWS = 256 / NT;   // A rows (float4 halves) loaded per thread: 1 (BN=128) or 2
    __shared__ __align__(16) float As[2][IG_BK][IG_AS];
    __shared__ __align__(16) float Bs[2][IG_BK][BN];
    const long long B = count ? (long long)*count : maxB;
    const long long M = B * S;
    const long long m0 = (long long)blockIdx.x * IG_BM;
    if (m0 >= M) return;
    const int n0 = blockIdx.y * BN;
    const int tid = threadIdx.x, tx = tid % TXN, ty = tid / TXN;

    // A-load assignment: float4 #idx covers row idx/2, channels (idx%2)*4..+3 of the chunk
    int a_row[AROWS], a_half[AROWS], a_y[AROWS], a_x[AROWS];
    long long a_m[AROWS];
    bool a_ok[AROWS];
#pragma unroll
    for (int r = 0; r < AROWS; ++r) {
        int idx = tid + r * NT;
        a_row[r] = idx >> 1;
        a_half[r] = idx & 1;
        a_m[r] = m0 + a_row[r];
        a_ok[r] = a_m[r] < M;
        int cell = (int)(a_m[r] % S);
        a_y[r] = cell / W;
        a_x[r] = cell % W;
    }
    // B-load assignment: one float4 per thread: row k = tid / (BN/4), col4 = tid % (BN/4)
    const int b_k = tid / (BN / 4), b_c = (tid % (BN / 4)) * 4;

    const int chunks_per_tap = Cin / IG_BK;
    const int n_chunks = 9 * chunks_per_tap;
    float4 ra[AROWS], rb;
    auto gload = [&](int ch) {
        const int tap = ch / chunks_per_tap, c0 = (ch % chunks_per_tap) * IG_BK;
        const int dy = tap / 3 - 1, dx = tap % 3 - 1;
#pragma unroll
        for (int r = 0; r < AROWS; ++r) {
            const int yy = a_y[r] + dy, xx = a_x[r] + dx;
            if (a_ok[r] && yy >= 0 && yy < H && xx >= 0 && xx < W)
                ra[r] = *reinterpret_cast<const float4 *>(in + (a_m[r] + dy * W + dx) * Cin + c0 + a_half[r] * 4);
            else
                ra[r] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        rb = *reinterpret_cast<const float4 *>(w + ((long long)tap * Cin + c0 + b_k) * Cout + n0 + b_c);
    };
    auto sstore = [&](int buf) {
#pragma unroll
        for (int r = 0; r < AROWS; ++r) {
            const int k = a_half[r] * 4;
            As[buf][k + 0][a_row[r]] = ra[r].x;
            As[buf][k + 1][a_row[r]] = ra[r].y;
            As[buf][k + 2][a_row[r]] = ra[r].z;
            As[buf][k + 3][a_row[r]] = ra[r].w;
        }
        *reinterpret_cast<float4 *>(&Bs[buf][b_k][b_c]) = rb;
    };

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    gload(0);
    sstore(0);
    __syncthreads();
    for (int ch = 0; ch < n_chunks; ++ch) {
        const int buf = ch & 1;
        if (ch + 1 < n_chunks) gload(ch + 1);
#pragma unroll
        for (int k = 0; k < IG_BK; ++k) {
            const float4 a0 = *reinterpret_cast<const float4 *>(&As[buf][k][ty * 8]);
            const float4 a1 = *reinterpret_cast<const float4 *>(&As[buf][k][ty * 8 + 4]);
            const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][k][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][k][BN / 2 + tx * 4]);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (ch + 1 < n_chunks) sstore(buf ^ 1);
        __syncthreads();
    }
    if (EPI == 1) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const long long m = m0 + ty * 8 + i;
            if (m >= M) continue;
            float *o = out + m * Cout + n0;
            const float4 k0 = *reinterpret_cast<const float4 *>(bias + m * Cout + n0 + tx * 4);
            const float4 k1 = *reinterpret_cast<const float4 *>(bias + m * Cout + n0 + BN / 2 + tx * 4);
            float4 v0, v1;
            v0.x = k0.x > 0.f ? acc[i][0] : 0.f; v0.y = k0.y > 0.f ? acc[i][1] : 0.f;
            v0.z = k0.z > 0.f ? acc[i][2] : 0.f; v0.w = k0.w > 0.f ? acc[i][3] : 0.f;
            v1.x = k1.x > 0.f ? acc[i][4] : 0.f; v1.y = k1.y > 0.f ? acc[i][5] : 0.f;
            v1.z = k1.z > 0.f ? acc[i][6] : 0.f; v1.w = k1.w > 0.f ? acc[i][7] : 0.f;
            *reinterpret_cast<float4 *>(o + tx * 4) = v0;
            *reinterpret_cast<float4 *>(o + BN / 2 + tx * 4) = v1;
        }
        return;
    }
    const float4 bi0 = *reinterpret_cast<const float4 *>(bias + n0 + tx * 4);
    const float4 bi1 = *reinterpret_cast<const float4 *>(bias + n0 + BN / 2 + tx * 4);
    const float bb[8] = {bi0.x, bi0.y, bi0.z, bi0.w, bi1.x, bi1.y, bi1.z, bi1.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const long long m = m0 + ty * 8 + i;
        if (m >= M) continue;
        float *o = out + m * Cout + n0;
        float4 v0, v1;
        v0.x = fmaxf(acc[i][0] + bb[0], 0.f); v0.y = fmaxf(acc[i][1] + bb[1], 0.f);
        v0.z = fmaxf(acc[i][2] + bb[2], 0.f); v0.w = fmaxf(acc[i][3] + bb[3], 0.f);
        v1.x = fmaxf(acc[i][4] + bb[4], 0.f); v1.y = fmaxf(acc[i][5] + bb[5], 0.f);
        v1.z = fmaxf(acc[i][6] + bb[6], 0.f); v1.w = fmaxf(acc[i][7] + bb[7], 0.f);
        *reinterpret_cast<float4 *>(o + tx * 4) = v0;
        *reinterpret_cast<float4 *>(o + BN / 2 + tx * 4) = v1;
    }
}

// ---- heads: 1x1 convs + FCs + softmax / tanh, one CTA per board ---------------------------
// wh [c3][8] (cols 0-3 policy conv, 4-5 value conv), bh[8]; wp [4S][S] (transposed act_fc1),
// bp[S]; wv [2S][64] (transposed val_fc1), bv[64]; wv2[64], bv2[1].
struct HeadWeights {
    const float *wh, *bh, *wp, *bp, *wv, *bv, *wv2, *bv2;
};

__global__ void __launch_bounds__(128)
heads_kernel(const float *__restrict__ act3 /*[B*S][c3]*/, HeadWeights hw, float *__restrict__ probs,
             float *__restrict__ values, const int32_t *__restrict__ count, long long maxB, int S, int c3)
{
    extern __shared__ float sm[];
    float *s_wh = sm;              // c3*8
    float *s_h = s_wh + c3 * 8;    // 6*S   ([4][S] policy planes then [2][S] value planes)
    float *s_l = s_h + 6 * S;      // S logits
    float *s_v = s_l + S;          // 64
    float *s_red = s_v + 64;       // 8
    const long long B = count ? (long long)*count : maxB;
    const long long b = blockIdx.x;
    if (b >= B) return;
    const int tid = threadIdx.x;
    for (int i = tid; i < c3 * 8; i += blockDim.x) s_wh[i] = hw.wh[i];
    __syncthreads();
    for (int cell = tid; cell < S; cell += blockDim.x) {
        float acc[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) acc[j] = hw.bh[j];
        const float4 *a = reinterpret_cast<const float4 *>(act3 + (b * S + cell) * c3);
        for (int c4 = 0; c4 < c3 / 4; ++c4) {
            const float4 v = a[c4];
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
    // policy FC (flatten order (C,H,W): k = c*S + cell, policy_value_net_pytorch.py:50)
    for (int o = tid; o < S; o += blockDim.x) {
        float acc = hw.bp[o];
        for (int k = 0; k < 4 * S; ++k) acc = fmaf(s_h[k], hw.wp[(size_t)k * S + o], acc);
        s_l[o] = acc;
    }
    // value FC1 + ReLU
    if (tid < 64) {
        float acc = hw.bv[tid];
        const float *hv = s_h + 4 * S;
        for (int k = 0; k < 2 * S; ++k) acc = fmaf(hv[k], hw.wv[(size_t)k * 64 + tid], acc);
        s_v[tid] = fmaxf(acc, 0.f);
    }
    __syncthreads();
    // softmax over S (exp(log_softmax), :51 and :91)
    float mx = -3.0e38f;
    for (int o = tid; o < S; o += blockDim.x) mx = fmaxf(mx, s_l[o]);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    if ((tid & 31) == 0) s_red[tid >> 5] = mx;
    __syncthreads();
    mx = fmaxf(fmaxf(s_red[0], s_red[1]), fmaxf(s_red[2], s_red[3]));
    float sum = 0.f;
    for (int o = tid; o < S; o += blockDim.x) {
        float ex = expf(s_l[o] - mx);
        s_l[o] = ex;
        sum += ex;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
    if ((tid & 31) == 0) s_red[4 + (tid >> 5)] = sum;
    __syncthreads();
    sum = s_red[4] + s_red[5] + s_red[6] + s_red[7];
    const float inv = 1.f / sum;
    for (int o = tid; o < S; o += blockDim.x) probs[b * S + o] = s_l[o] * inv;
    if (tid < 32) {
        float acc = s_v[tid] * hw.wv2[tid] + s_v[tid + 32] * hw.wv2[tid + 32];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        if (tid == 0) values[b] = tanhf(acc + hw.bv2[0]);
    }
}

// ---- host side ------------------------------------------------------------------------------
int gz_net_ensure_buffers(gz_engine *e, int64_t B)
{
    if (B <= e->act_cap) return GZ_OK;
    if (e->act1) cudaFree(e->act1);
    if (e->act2) cudaFree(e->act2);
    if (e->act3) cudaFree(e->act3);
    e->act1 = e->act2 = e->act3 = nullptr;
    e->act_cap = 0;
    size_t rows = (size_t)B * e->S;
    GZ_CUDA(e, cudaMalloc(&e->act1, rows * e->c1 * sizeof(float)));
    GZ_CUDA(e, cudaMalloc(&e->act2, rows * e->c2 * sizeof(float)));
    GZ_CUDA(e, cudaMalloc(&e->act3, rows * e->c3 * sizeof(float)));
    e->act_cap = B;
    return GZ_OK;
}

// d_count: device pointer to the number of boards (NULL -> maxB boards)
static int gz_net_forward_ffma(gz_engine *e, const float *d_planes, const int32_t *d_count, int64_t maxB,
                               float *d_probs, float *d_values)
{
    if (maxB <= 0) return GZ_OK;
    int rc = gz_net_ensure_buffers(e, maxB);
    if (rc) return rc;
    NetDims d = {e->W, e->H, e->S, e->c1, e->c2, e->c3};
    const NetWeights &w = e->wts;
    {
        ProfScope ps(e, PC_CONV1);
        long long threads = (long long)maxB * e->S * (e->c1 / 8);
        size_t smem = (size_t)(37 * e->c1) * sizeof(float);
        conv1_kernel<<<(unsigned)((threads + 255) / 256), 256, smem, e->stream>>>(
            d_planes, w.w1, w.bias + w.off_b1, e->act1, d_count, maxB, d);
    }
    const unsigned gx = (unsigned)(((long long)maxB * e->S + IG_BM - 1) / IG_BM);
    {
    ProfScope ps(e, PC_CONV2);
    if (e->c2 % 128 == 0)
        conv3x3_igemm_kernel<128><<<dim3(gx, e->c2 / 128), 256, 0, e->stream>>>(
            e->act1, w.w2, w.bias + w.off_b2, e->act2, d_count, maxB, e->S, e->W, e->H, e->c1, e->c2);
    else
        conv3x3_igemm_kernel<64><<<dim3(gx, e->c2 / 64), 128, 0, e->stream>>>(
            e->act1, w.w2, w.bias + w.off_b2, e->act2, d_count, maxB, e->S, e->W, e->H, e->c1, e->c2);
    }
    {
    ProfScope ps(e, PC_CONV3);
    if (e->c3 % 128 == 0)
        conv3x3_igemm_kernel<128><<<dim3(gx, e->c3 / 128), 256, 0, e->stream>>>(
            e->act2, w.w3, w.bias + w.off_b3, e->act3, d_count, maxB, e->S, e->W, e->H, e->c2, e->c3);
    else
        conv3x3_igemm_kernel<64><<<dim3(gx, e->c3 / 64), 128, 0, e->stream>>>(
            e->act2, w.w3, w.bias + w.off_b3, e->act3, d_count, maxB, e->S, e->W, e->H, e->c2, e->c3);
    }
    HeadWeights hw;
    hw.wh = w.wh; hw.bh = w.bias + w.off_bh; hw.wp = w.wpfc; hw.bp = w.bias + w.off_bp;
    hw.wv = w.wvfc; hw.bv = w.bias + w.off_bv; hw.wv2 = w.bias + w.off_wv2; hw.bv2 = w.bias + w.off_bv2;
    size_t hsm = (size_t)(e->c3 * 8 + 7 * e->S + 64 + 8) * sizeof(float);
    {
    ProfScope ps(e, PC_HEADS);
    heads_kernel<<<(unsigned)maxB, 128, hsm, e->stream>>>(e->act3, hw, d_probs, d_values, d_count, maxB,
                                                           e->S, e->c3);
    }
    e->launches += 4;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}
