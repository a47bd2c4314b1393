// gz_net_tcw.cu — PolicyValueNet forward (reference policy_value_net_pytorch.py:25-57) on the
// sm_100a tensor cores for trunks WIDER than the reference's 32/64/128 (BASELINE config C4:
// "wider conv trunk", e.g. 64/128/256).  The fused kernel of gz_net_tc.cu keeps every activation of a
// group in shared memory, which stops fitting at these widths, so this path runs layer by layer:
//
//   conv1_tcw_kernel     conv1 (K = 36: too thin for the tensor core) in fp32 on the CUDA cores;
//                        writes ReLU(conv1) split into fp16 hi/lo directly as the A-operand image
//                        of the next layer.
//   conv_tcw_kernel<0>   conv2 as an implicit GEMM on tcgen05 (M = 128 line-space rows, N = Cout,
//                        K = 16 input channels of one filter tap); epilogue writes the next image.
//   conv_tcw_kernel<1>   conv3 the same way; the epilogue also applies both 1x1 head convs and
//                        writes heads_tc_kernel's A operand, so the FC heads of gz_net_tc.cu run
//                        unchanged on its output.
//
// Activation image of a group of boards (global memory, L2-resident between layers): the padded
// line space of gz_net_tc.cu (Lay<false>: cell (i, j) of board b -> row b * BS + i * P + j, LEAD
// zero rows in front, RALLOC rows in all) stored as [Cin/8 hi chunk planes | Cin/8 lo chunk
// planes][row][8 halves].  That IS the no-swizzle K-major UMMA layout, so a CTA brings the A operand
// of a whole group into shared memory with two bulk copies per K step and a filter tap is a shift
// of the descriptor's start address.  Padding rows are written as zeros by the producing epilogue;
// the LEAD / trailing rows are zeroed once at allocation and never written.
//
// Same fp32-grade numerics as the fused kernel: x = hi + lo (two fp16), three MMAs per product
// (hi*hi + lo*hi + hi*lo), fp32 accumulation in TMEM, weights pre-scaled by a power of two.
// Included by gz_api.cu after gz_net_tc.cu (shares its PTX wrappers, TcGeom and heads_tc_kernel).

namespace tcw {
constexpr int LEAD = tc::Lay<false>::LEAD, RALLOC = tc::Lay<false>::RALLOC, CH = tc::Lay<false>::CH;
constexpr int THREADS = 352;      // warps 0-7 epilogue, 8 MMA issue, 9 weight stages, 10 A-operand loads
constexpr int MAX_NS = 8;         // weight ring depth (slots of 64 * Cout bytes)
constexpr int OFF_BAR = 0;        // a_full[8] @0, full[8] @64, empty[8] @128, mma_done @192, tmem_free @200, tmem slot @208
constexpr int OFF_BIAS = 256;     // float bias[256]
constexpr int OFF_WH = OFF_BIAS + 1024;  // float wh[256][8] (1x1 head convs: 0-3 policy, 4-5 value) | bh[8]
constexpr int OFF_A = OFF_WH + 8192 + 128; // A operand of the group: (Cin / 4) chunk planes of CH bytes
constexpr int SMEM_MAX = 232448;
static_assert(OFF_A % 128 == 0, "A operand alignment");
} // namespace tcw

struct TcwLayer {
    const uint8_t *wimg; // stage (k step, tap): [pass hi|lo][8-channel chunk 2][Cout][8 halves]
    const float *bias;   // [Cout]
    const float *wh;     // MODE 1: [Cout][8] 1x1 head conv weights
    const float *bh;     // MODE 1: [8] head biases
    float inv_s;         // 1 / weight pre-scale
    int cin, cout, ns;   // ns = ring depth
};

// ---- conv1 on the CUDA cores --------------------------------------------------------------------
// One block per group, one thread per line-space row; w1 = [tap][plane 4][c1] fp32.
__global__ void __launch_bounds__(256)
conv1_tcw_kernel(const float *__restrict__ planes, const float *__restrict__ w1, const float *__restrict__ b1,
                 uint8_t *__restrict__ act_out, const int32_t *__restrict__ count, long long maxB, TcGeom g, int c1,
                 float *__restrict__ dbg)
{
    using namespace tcw;
    extern __shared__ __align__(16) float s_w[]; // [36][c1] | bias[c1]
    const long long B = count ? (long long)*count : maxB;
    const long long grp = blockIdx.x;
    if (grp * g.NB >= B) return;
    for (int i = threadIdx.x; i < 36 * c1; i += 256) s_w[i] = w1[i];
    for (int i = threadIdx.x; i < c1; i += 256) s_w[36 * c1 + i] = b1[i];
    __syncthreads();
    const int r = threadIdx.x;
    const int bq = r / g.BS, rr = r % g.BS;
    const int ci = rr / g.P, cj = rr % g.P;
    const long long board = grp * g.NB + bq;
    const bool valid = bq < g.NB && ci < g.W && cj < g.W && board < B;
    float in[36];
#pragma unroll
    for (int t = 0; t < 9; ++t) {
        const int y = ci + t / 3 - 1, x = cj + t % 3 - 1;
        const bool ok = valid && y >= 0 && y < g.W && x >= 0 && x < g.W;
#pragma unroll
        for (int p = 0; p < 4; ++p) in[t * 4 + p] = ok ? planes[(board * 4 + p) * g.S + y * g.W + x] : 0.f;
    }
    uint8_t *out = act_out + (size_t)grp * (c1 / 4) * CH + (size_t)(LEAD + r) * 16;
    for (int c8 = 0; c8 < c1 / 8; ++c8) {
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = s_w[36 * c1 + c8 * 8 + j];
#pragma unroll
        for (int k = 0; k < 36; ++k) {
            const float4 wa = *reinterpret_cast<const float4 *>(s_w + k * c1 + c8 * 8);
            const float4 wb = *reinterpret_cast<const float4 *>(s_w + k * c1 + c8 * 8 + 4);
            acc[0] = fmaf(in[k], wa.x, acc[0]); acc[1] = fmaf(in[k], wa.y, acc[1]);
            acc[2] = fmaf(in[k], wa.z, acc[2]); acc[3] = fmaf(in[k], wa.w, acc[3]);
            acc[4] = fmaf(in[k], wb.x, acc[4]); acc[5] = fmaf(in[k], wb.y, acc[5]);
            acc[6] = fmaf(in[k], wb.z, acc[6]); acc[7] = fmaf(in[k], wb.w, acc[7]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = valid ? fmaxf(acc[j], 0.f) : 0.f;
        uint4 hi, lo;
        tc_split8(acc, hi, lo);
        *reinterpret_cast<uint4 *>(out + (size_t)c8 * CH) = hi;
        *reinterpret_cast<uint4 *>(out + (size_t)(c1 / 8 + c8) * CH) = lo;
        if (dbg && valid) {
            float *d = dbg + ((size_t)board * g.S + ci * g.W + cj) * c1 + c8 * 8;
#pragma unroll
            for (int j = 0; j < 8; ++j) d[j] = acc[j];
        }
    }
}

// ---- 3x3 conv as an implicit GEMM on tcgen05 ---------------------------------------------------
// Persistent over groups (grid = min(groups, SMs), 1 CTA per SM).  Per group: warp 10 brings the A
// operand (the whole activation image of the group, (Cin / 4) * CH bytes) into shared memory with two
// bulk copies per K step, each K step on its own mbarrier so the MMAs start when the first 16 input
// channels have landed; warp 9 streams the weight stages (k step major, 9 taps each) through a ring;
// warp 8 issues 3 passes x 2 M-tiles per stage with N = Cout; warps 0-7 run the epilogue from TMEM
// (2 x Cout fp32 columns) while warp 10 already loads the next group's A operand.
//   MODE 0: bias + ReLU, hi/lo split, written as the next layer's image (16-byte coalesced stores).
//   MODE 1: bias + ReLU, both 1x1 head convs (+ ReLU), hi/lo halves written as heads_tc_kernel's operand.
template <int MODE>
__global__ void __launch_bounds__(tcw::THREADS, 1)
conv_tcw_kernel(const uint8_t *__restrict__ act_in, const TcwLayer L, uint8_t *__restrict__ act_out,
                __half *__restrict__ head_act, const int32_t *__restrict__ count, long long maxB, TcGeom g,
                float *__restrict__ dbg_act, float *__restrict__ dbg_head)
{
    using namespace tcw;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const long long B = __shfl_sync(0xffffffffu, count ? (long long)*count : maxB, 0);
    const long long n_groups = (B + g.NB - 1) / g.NB;
    const uint32_t sbase = tc_smem_u32(smem);
    const uint32_t bar_a = sbase + OFF_BAR, bar_full = sbase + OFF_BAR + 64, bar_empty = sbase + OFF_BAR + 128;
    const uint32_t bar_done = sbase + OFF_BAR + 192, bar_free = sbase + OFF_BAR + 200;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(smem + OFF_BAR + 208);
    float *s_bias = reinterpret_cast<float *>(smem + OFF_BIAS);
    float *s_wh = reinterpret_cast<float *>(smem + OFF_WH);
    const int cin = L.cin, cout = L.cout, KS = cin >> 4;
    const uint32_t stage_bytes = 64u * (uint32_t)cout;
    const uint32_t a_bytes = (uint32_t)(cin >> 2) * CH;
    const uint32_t ring = sbase + OFF_A + a_bytes;

    for (int i = tid; i < cout; i += THREADS) s_bias[i] = L.bias[i];
    if (MODE == 1) {
        for (int i = tid; i < cout * 8; i += THREADS) s_wh[i] = L.wh[i];
        if (tid < 8) s_wh[cout * 8 + tid] = L.bh[tid];
    }
    if (tid == 0) {
        for (int s = 0; s < 8; ++s) {
            tc_mbar_init(bar_a + 8 * s, 1);
            tc_mbar_init(bar_full + 8 * s, 1);
            tc_mbar_init(bar_empty + 8 * s, 1);
        }
        tc_mbar_init(bar_done, 1);
        tc_mbar_init(bar_free, 8); // one arrival per epilogue warp
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + OFF_BAR + 208),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);

    if (warp == 10) {
        // ---- A operand: the group's activation image, one mbarrier per K step ---------------------
        if (lane == 0) {
            uint32_t gi = 0;
            for (long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++gi) {
                if (gi > 0) tc_mbar_wait(bar_done, (gi - 1) & 1u); // every MMA that read the previous image is complete
                const uint8_t *src = act_in + (size_t)grp * a_bytes;
                for (int ks = 0; ks < KS; ++ks) {
                    tc_mbar_expect_tx(bar_a + 8 * ks, 4u * CH);
                    const uint32_t o_hi = (uint32_t)(2 * ks) * CH, o_lo = (uint32_t)((cin >> 3) + 2 * ks) * CH;
                    tc_bulk_g2s(sbase + OFF_A + o_hi, src + o_hi, 2u * CH, bar_a + 8 * ks);
                    tc_bulk_g2s(sbase + OFF_A + o_lo, src + o_lo, 2u * CH, bar_a + 8 * ks);
                }
            }
        }
    } else if (warp == 9) {
        // ---- weight stages: the same KS * 9 stages for every group, streamed from L2 -------------
        if (lane == 0) {
            uint32_t it = 0;
            const uint32_t ns = (uint32_t)L.ns;
            for (long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x)
                for (int s = 0; s < KS * 9; ++s, ++it) {
                    const uint32_t slot = it % ns, ph = (it / ns) & 1u;
                    tc_mbar_wait(bar_empty + 8 * slot, ph ^ 1u);
                    tc_mbar_expect_tx(bar_full + 8 * slot, stage_bytes);
                    tc_bulk_g2s(ring + slot * stage_bytes, L.wimg + (size_t)s * stage_bytes, stage_bytes, bar_full + 8 * slot);
                }
        }
    } else if (warp == 8) {
        // ---- MMA issue ---------------------------------------------------------------------------
        const uint32_t leader = tc_elect();
        const uint32_t ns = (uint32_t)L.ns;
        const uint32_t idesc = tc_idesc(cout);
        const uint64_t dA = tc_desc(sbase + OFF_A + LEAD * 16, CH);
        const uint64_t dB = tc_desc(ring, 16u * (uint32_t)cout);
        uint32_t it = 0, gi = 0;
        for (long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++gi) {
            if (gi > 0) { // the epilogue warps have drained the previous group's accumulators
                tc_mbar_wait(bar_free, (gi - 1) & 1u);
                tc_fence_after();
            }
            for (int ks = 0; ks < KS; ++ks) {
                tc_mbar_wait(bar_a + 8 * ks, gi & 1u);
                tc_fence_after();
                for (int t = 0; t < 9; ++t, ++it) {
                    const uint32_t slot = it % ns, ph = (it / ns) & 1u;
                    tc_mbar_wait(bar_full + 8 * slot, ph);
                    tc_fence_after();
                    const int shift = (t / 3 - 1) * g.P + (t % 3 - 1);
                    const uint64_t a_hi = dA + (int64_t)(shift + (2 * ks) * (CH / 16));
                    const uint64_t a_lo = dA + (int64_t)(shift + ((cin >> 3) + 2 * ks) * (CH / 16));
                    const uint64_t b_hi = dB + (uint64_t)(slot * (stage_bytes / 16)), b_lo = b_hi + (uint64_t)(2 * cout);
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const uint32_t d = tmem + (uint32_t)(j * cout);
                        tc_umma_f16(leader, d, a_hi + (uint32_t)(j * 128), b_hi, idesc, (ks | t) != 0);
                        tc_umma_f16(leader, d, a_lo + (uint32_t)(j * 128), b_hi, idesc, 1u);
                        tc_umma_f16(leader, d, a_hi + (uint32_t)(j * 128), b_lo, idesc, 1u);
                    }
                    tc_umma_commit(leader, bar_empty + 8 * slot);
                }
            }
            tc_umma_commit(leader, bar_done);
        }
    } else {
        // ---- epilogue warps 0-7: warp w owns TMEM lanes 32 * (w % 4) of M-tile w / 4 ----------------
        const int tile = warp >> 2;
        const int r = tid;
        const int bq = r / g.BS, rr = r % g.BS;
        const int ci = rr / g.P, cj = rr % g.P;
        const bool geo_ok = bq < g.NB && ci < g.W && cj < g.W;
        const int cell = ci * g.W + cj;
        const uint32_t t_lane = ((uint32_t)((warp & 3) * 32)) << 16;
        uint32_t gi = 0;
        for (long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++gi) {
            const long long board = grp * g.NB + bq;
            const bool valid = geo_ok && board < B;
            tc_mbar_wait(bar_done, gi & 1u);
            tc_fence_after();
            float hacc[6];
            if (MODE == 1) {
#pragma unroll
                for (int q = 0; q < 6; ++q) hacc[q] = s_wh[cout * 8 + q];
            }
            uint8_t *out = MODE == 0 ? act_out + (size_t)grp * (cout >> 2) * CH + (size_t)(LEAD + r) * 16 : nullptr;
            for (int part = 0; part < (cout >> 5); ++part) {
                uint32_t v[32];
                tc_tmem_ld32(tmem + t_lane + (uint32_t)(tile * cout + part * 32), v);
                float f[32];
#pragma unroll
                for (int c = 0; c < 32; ++c)
                    f[c] = valid ? fmaxf(fmaf(__uint_as_float(v[c]), L.inv_s, s_bias[part * 32 + c]), 0.f) : 0.f;
                if (MODE == 0) {
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        uint4 hi, lo;
                        tc_split8(f + 8 * c, hi, lo);
                        *reinterpret_cast<uint4 *>(out + (size_t)(part * 4 + c) * CH) = hi;
                        *reinterpret_cast<uint4 *>(out + (size_t)((cout >> 3) + part * 4 + c) * CH) = lo;
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        const float4 wa = *reinterpret_cast<const float4 *>(s_wh + (part * 32 + c) * 8);
                        const float2 wb = *reinterpret_cast<const float2 *>(s_wh + (part * 32 + c) * 8 + 4);
                        hacc[0] = fmaf(f[c], wa.x, hacc[0]); hacc[1] = fmaf(f[c], wa.y, hacc[1]);
                        hacc[2] = fmaf(f[c], wa.z, hacc[2]); hacc[3] = fmaf(f[c], wa.w, hacc[3]);
                        hacc[4] = fmaf(f[c], wb.x, hacc[4]); hacc[5] = fmaf(f[c], wb.y, hacc[5]);
                    }
                }
                if (dbg_act && valid) {
                    float *d = dbg_act + ((size_t)board * g.S + cell) * cout + part * 32;
#pragma unroll
                    for (int c = 0; c < 32; ++c) d[c] = f[c];
                }
            }
            // accumulators are in registers: release TMEM to the next group's MMAs
            tc_fence_before();
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_free) : "memory");
            if (MODE == 1 && valid) {
                // heads_tc_kernel's A operand: [tile of 128 boards][K step][pass hi|lo][8-k chunk 2][board 128][8 halves]
                __half *ha = head_act + (size_t)(board >> 7) * g.J * 4096 + (size_t)(board & 127) * 8;
#pragma unroll
                for (int q = 0; q < 6; ++q) {
                    const float x = fmaxf(hacc[q], 0.f);
                    const int k = (q < 4 ? q * g.S : 16 * g.Jp + (q - 4) * g.S) + cell;
                    const __half hi = __float2half_rn(x);
                    const size_t o = (size_t)((k >> 4) * 4 + ((k >> 3) & 1)) * 1024 + (k & 7);
                    ha[o] = hi;
                    ha[o + 2048] = __float2half_rn(x - __half2float(hi));
                    if (dbg_head) dbg_head[((size_t)board * 6 + q) * g.S + cell] = x;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    }
}

// ---- host side --------------------------------------------------------------------------------
// Trunks this path takes: square boards of side 3-15; conv2 / conv3 input widths (c1, c2) multiples of 16
// and at most 128 (the A operand of a group must fit shared memory beside the weight ring); output widths
// (c2, c3) multiples of 32 and at most 256 (one N per MMA, 2 x Cout TMEM columns).
static bool gz_tcw_supported(const gz_engine *e)
{
    return e->W == e->H && e->W >= 3 && e->W <= 15 && e->c1 % 16 == 0 && e->c1 >= 16 && e->c1 <= 128 &&
           e->c2 % 32 == 0 && e->c2 >= 32 && e->c2 <= 128 && e->c3 % 32 == 0 && e->c3 >= 32 && e->c3 <= 256;
}

static int gz_tcw_ring_depth(int cin, int cout)
{
    const int left = tcw::SMEM_MAX - tcw::OFF_A - (cin / 4) * tcw::CH;
    int ns = left / (64 * cout);
    if (ns > tcw::MAX_NS) ns = tcw::MAX_NS;
    return ns;
}

// weight image of one conv layer: stage (k step ks, tap t) = [pass hi|lo][chunk 2][Cout][8 halves]
static void gz_tcw_conv_image(const std::vector<float> &w, int cin, int cout, float s, std::vector<uint8_t> &img)
{
    const size_t stage = (size_t)64 * cout;
    img.assign((size_t)(cin / 16) * 9 * stage, 0);
    for (int n = 0; n < cout; ++n)
        for (int ci = 0; ci < cin; ++ci)
            for (int t = 0; t < 9; ++t) {
                float back;
                __half hi, lo;
                gz_tc_split(w[((size_t)n * cin + ci) * 9 + t] * s, back, hi, lo);
                const int ks = ci / 16, c = (ci / 8) & 1, el = ci % 8;
                const size_t o = (size_t)(ks * 9 + t) * stage + ((size_t)c * cout + n) * 16 + el * 2;
                memcpy(img.data() + o, &hi, 2);
                memcpy(img.data() + o + (size_t)32 * cout, &lo, 2);
            }
}

int gz_tcw_prepare_weights(gz_engine *e, std::vector<float> h[16])
{
    const float s2 = gz_tc_scale(h[2]), s3 = gz_tc_scale(h[4]);
    std::vector<uint8_t> img;
    gz_tcw_conv_image(h[2], e->c1, e->c2, s2, img);
    GZ_TRY(gz_upload(e, &e->wts.tcw_img2, img));
    gz_tcw_conv_image(h[4], e->c2, e->c3, s3, img);
    GZ_TRY(gz_upload(e, &e->wts.tcw_img3, img));
    e->wts.tcw_inv[0] = 1.f / s2; e->wts.tcw_inv[1] = 1.f / s3;
    return gz_tc_prepare_fc(e, h);
}

static int gz_tcw_ensure_buffers(gz_engine *e, int64_t groups, int64_t B, bool want_dbg)
{
    if (groups > e->tcw_cap) {
        if (e->tcw_act1) cudaFree(e->tcw_act1);
        if (e->tcw_act2) cudaFree(e->tcw_act2);
        e->tcw_act1 = e->tcw_act2 = nullptr;
        e->tcw_cap = 0;
        // zeroed once: the LEAD / trailing rows of every group image are never written afterwards
        const size_t b1 = (size_t)groups * (e->c1 / 4) * tcw::CH, b2 = (size_t)groups * (e->c2 / 4) * tcw::CH;
        GZ_CUDA(e, cudaMalloc(&e->tcw_act1, b1));
        GZ_CUDA(e, cudaMalloc(&e->tcw_act2, b2));
        GZ_CUDA(e, cudaMemsetAsync(e->tcw_act1, 0, b1, e->stream));
        GZ_CUDA(e, cudaMemsetAsync(e->tcw_act2, 0, b2, e->stream));
        e->tcw_cap = groups;
    }
    if (want_dbg && B > e->tcw_dbg_cap) {
        if (e->tcw_dbg) cudaFree(e->tcw_dbg);
        e->tcw_dbg = nullptr;
        GZ_CUDA(e, cudaMalloc(&e->tcw_dbg, (size_t)B * e->S * (e->c1 + e->c2 + e->c3 + 6) * sizeof(float)));
        e->tcw_dbg_cap = B;
    }
    return GZ_OK;
}

// dbg layout (want_dbg): act1 [B][S][c1] | act2 [B][S][c2] | act3 [B][S][c3] | head [B][6][S]; act2 is
// decoded from its fp16 hi/lo image by gz_debug_net_layers.
static int gz_net_forward_tcw(gz_engine *e, const float *d_planes, const int32_t *d_count, int64_t maxB,
                              float *d_probs, float *d_values, bool want_dbg)
{
    using namespace tcw;
    if (maxB <= 0) return GZ_OK;
    if (!gz_tcw_supported(e)) { e->err = "tcgen05 wide-trunk path: unsupported trunk widths or board"; return GZ_E_INVALID; }
    const TcGeom g = gz_tc_geom(e, false);
    const long long groups = (maxB + g.NB - 1) / g.NB;
    GZ_TRY(gz_tc_ensure_buffers(e, maxB, false));
    GZ_TRY(gz_tcw_ensure_buffers(e, groups, maxB, want_dbg));
    if (e->n_sm <= 0) {
        int dev = 0;
        GZ_CUDA(e, cudaGetDevice(&dev));
        GZ_CUDA(e, cudaDeviceGetAttribute(&e->n_sm, cudaDevAttrMultiProcessorCount, dev));
    }
    if (!(e->attr_done & 4u)) {
        GZ_CUDA(e, cudaFuncSetAttribute(conv_tcw_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX));
        GZ_CUDA(e, cudaFuncSetAttribute(conv_tcw_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX));
        e->attr_done |= 4u;
    }
    const NetWeights &w = e->wts;
    const size_t BS_ = (size_t)maxB * e->S;
    float *dbg1 = want_dbg ? e->tcw_dbg : nullptr;
    float *dbg3 = want_dbg ? e->tcw_dbg + BS_ * (e->c1 + e->c2) : nullptr;
    float *dbgh = want_dbg ? e->tcw_dbg + BS_ * (e->c1 + e->c2 + e->c3) : nullptr;
    const unsigned grid = (unsigned)(groups < e->n_sm ? groups : e->n_sm);
    TcwLayer L2 = {w.tcw_img2, w.bias + w.off_b2, nullptr, nullptr, w.tcw_inv[0], e->c1, e->c2, gz_tcw_ring_depth(e->c1, e->c2)};
    TcwLayer L3 = {w.tcw_img3, w.bias + w.off_b3, w.wh, w.bias + w.off_bh, w.tcw_inv[1], e->c2, e->c3, gz_tcw_ring_depth(e->c2, e->c3)};
    if (L2.ns < 2 || L3.ns < 2) { e->err = "tcgen05 wide-trunk path: weight ring does not fit"; return GZ_E_INVALID; }
    {
        ProfScope ps(e, PC_TRUNK_TC);
        conv1_tcw_kernel<<<(unsigned)groups, 256, (size_t)37 * e->c1 * sizeof(float), e->stream>>>(
            d_planes, w.w1, w.bias + w.off_b1, e->tcw_act1, d_count, maxB, g, e->c1, dbg1);
        conv_tcw_kernel<0><<<grid, THREADS, OFF_A + (size_t)(L2.cin / 4) * CH + (size_t)L2.ns * 64 * L2.cout, e->stream>>>(
            e->tcw_act1, L2, e->tcw_act2, nullptr, d_count, maxB, g, nullptr, nullptr);
        conv_tcw_kernel<1><<<grid, THREADS, OFF_A + (size_t)(L3.cin / 4) * CH + (size_t)L3.ns * 64 * L3.cout, e->stream>>>(
            e->tcw_act2, L3, nullptr, reinterpret_cast<__half *>(e->tc_head), d_count, maxB, g, dbg3, dbgh);
    }
    GZ_TRY(gz_tc_launch_heads(e, g, d_probs, d_values, d_count, maxB));
    e->launches += 4;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK;
}

// parity tests: decode the fp16 hi/lo image of conv2's output into [B][S][c2] fp32 (host side)
static int gz_tcw_decode_act2(gz_engine *e, int64_t B, float *out)
{
    const TcGeom g = gz_tc_geom(e, false);
    const int64_t groups = (B + g.NB - 1) / g.NB;
    const size_t gbytes = (size_t)(e->c2 / 4) * tcw::CH;
    std::vector<uint8_t> img(gbytes * groups);
    GZ_CUDA(e, cudaMemcpy(img.data(), e->tcw_act2, img.size(), cudaMemcpyDeviceToHost));
    for (int64_t b = 0; b < B; ++b)
        for (int i = 0; i < g.W; ++i)
            for (int j = 0; j < g.W; ++j) {
                const size_t row = tcw::LEAD + (size_t)(b % g.NB) * g.BS + i * g.P + j;
                const uint8_t *gp = img.data() + (size_t)(b / g.NB) * gbytes;
                for (int c = 0; c < e->c2; ++c) {
                    __half hi, lo;
                    memcpy(&hi, gp + (size_t)(c / 8) * tcw::CH + row * 16 + (c % 8) * 2, 2);
                    memcpy(&lo, gp + (size_t)(e->c2 / 8 + c / 8) * tcw::CH + row * 16 + (c % 8) * 2, 2);
                    out[((size_t)b * e->S + i * g.W + j) * e->c2 + c] = __half2float(hi) + __half2float(lo);
                }
            }
    return GZ_OK;
}
