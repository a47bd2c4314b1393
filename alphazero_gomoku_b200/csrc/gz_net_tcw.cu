// gz_net_tcw.cu — PolicyValueNet forward (reference policy_value_net_pytorch.py:25-57) on the
// sm_100a tensor cores for trunks WIDER than the reference's 32/64/128 (BASELINE config C4:
// "wider conv trunk", e.g. 64/128/256).  The fused kernel of gz_net_tc.cu keeps every activation of a
// group in shared memory, which stops fitting at these widths, so this path runs layer by layer:
//
//   stage_tcw_kernel     input planes -> fp16 hi/lo operand image (4 real + 12 zero input channels).
//   conv_tcw_kernel<0>   conv1 and conv2 as implicit GEMMs on tcgen05 (M = 128 line-space rows, N = up to
//                        128 output channels, K = 16 input channels of one filter tap); the epilogue
//                        writes ReLU(conv) split into fp16 hi/lo directly as the next layer's image.
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
constexpr int MAX_NS = 8;         // weight ring depth (slots of 64 * NC bytes)
constexpr int OFF_BAR = 0;        // a_full[2][8] @0, full[8] @128, empty[8] @192, done[2] @256, free[2] @272, gdone[2] @288, tmem slot @304
constexpr int OFF_BIAS = 384;     // float bias[256]
constexpr int OFF_WH = OFF_BIAS + 1024;  // float wh[256][8] (1x1 head convs: 0-3 policy, 4-5 value) | bh[8]
constexpr int OFF_A = OFF_WH + 8192 + 128; // A operand of the group: (Cin / 4) chunk planes of CH bytes
constexpr int SMEM_MAX = 232448;
static_assert(OFF_A % 128 == 0, "A operand alignment");
} // namespace tcw

struct TcwLayer {
    const uint8_t *wimg; // stage (output chunk, k step, tap): [pass hi|lo][8-channel chunk 2][NC][8 halves]
    const float *bias;   // [Cout]
    const float *wh;     // MODE 1: [Cout][8] 1x1 head conv weights
    const float *bh;     // MODE 1: [8] head biases
    float inv_s;         // 1 / weight pre-scale
    int cin, cout, nc, ns, nabuf; // nc = output channels per chunk (128 or 64), ns = ring depth, nabuf = A-operand buffers (1 or 2)
};

// ---- input planes -> operand image of conv1 ------------------------------------------------------
// conv1 runs through conv_tcw_kernel<0> as a 16-input-channel layer (channels 4-15 are zero), so the
// planes only have to be laid out as an activation image: [hi chunk 0 | hi chunk 1 = 0 | lo chunk 0 |
// lo chunk 1 = 0][row][8 halves].  One block per group, one thread per line-space row.  Board planes
// are exactly {0, 1} (lo = 0); arbitrary float states of policy_value(state_batch) are split hi/lo.
__global__ void __launch_bounds__(256)
stage_tcw_kernel(const float *__restrict__ planes, uint8_t *__restrict__ a0, const int32_t *__restrict__ count,
                 long long maxB, TcGeom g)
{
    using namespace tcw;
    gz_pdl_wait(); // the leaf count and the planes come from the tree kernel before this one
    gz_pdl_trigger();
    const long long B = count ? (long long)*count : maxB;
    const long long grp = blockIdx.x;
    if (grp * g.NB >= B) return;
    const int r = threadIdx.x;
    const int bq = r / g.BS, rr = r % g.BS;
    const int ci = rr / g.P, cj = rr % g.P;
    const long long board = grp * g.NB + bq;
    const bool valid = bq < g.NB && ci < g.W && cj < g.W && board < B;
    uint4 hi = make_uint4(0, 0, 0, 0), lo = make_uint4(0, 0, 0, 0);
    if (valid) {
        const float *pp = planes + board * 4 * g.S + ci * g.W + cj;
        const float x0 = pp[0], x1 = pp[g.S], x2 = pp[2 * g.S], x3 = pp[3 * g.S];
        const __half2 a = __floats2half2_rn(x0, x1), b = __floats2half2_rn(x2, x3);
        const float2 fa = __half22float2(a), fb = __half22float2(b);
        const __half2 la = __floats2half2_rn(x0 - fa.x, x1 - fa.y), lb = __floats2half2_rn(x2 - fb.x, x3 - fb.y);
        hi.x = *reinterpret_cast<const uint32_t *>(&a);
        hi.y = *reinterpret_cast<const uint32_t *>(&b);
        lo.x = *reinterpret_cast<const uint32_t *>(&la);
        lo.y = *reinterpret_cast<const uint32_t *>(&lb);
    }
    uint8_t *out = a0 + (size_t)grp * 4 * CH + (size_t)(LEAD + r) * 16;
    *reinterpret_cast<uint4 *>(out) = hi;
    *reinterpret_cast<uint4 *>(out + 2 * (size_t)CH) = lo;
}

// ---- 3x3 conv as an implicit GEMM on tcgen05 ---------------------------------------------------
// Persistent over groups (grid = min(groups, SMs), 1 CTA per SM).  The output channels of a group are
// produced in chunks of NC = 128 (64 when Cout is not a multiple of 128) so that TMEM holds TWO
// accumulator buffers (2 M-tiles x NC fp32 columns each): the epilogue of one chunk runs under the MMAs
// of the next chunk / the next group.  Per group: warp 10 brings the A operand (the whole activation
// image of the group, (Cin / 4) * CH bytes) into shared memory with two bulk copies per K step, each K
// step on its own mbarrier so the MMAs start when the first 16 input channels have landed; warp 9
// streams the weight stages (chunk, k step, tap) through a ring; warp 8 issues 3 passes x 2 M-tiles per
// stage; warps 0-7 run the epilogues from TMEM.
//   MODE 0: bias + ReLU, hi/lo split, written as the next layer's image (16-byte coalesced stores).
//   MODE 1: bias + ReLU, both 1x1 head convs (+ ReLU), hi/lo halves written as heads_tc_kernel's operand.
template <int MODE>
__global__ void __launch_bounds__(tcw::THREADS, 1)
conv_tcw_kernel(const uint8_t *__restrict__ act_in, const TcwLayer L, uint8_t *__restrict__ act_out,
                __half *__restrict__ head_act, const int32_t *__restrict__ count, long long maxB, TcGeom g,
                float *__restrict__ dbg_act, float *__restrict__ dbg_head, long long *__restrict__ tl)
{
    using namespace tcw;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const uint32_t sbase = tc_smem_u32(smem);
    const uint32_t bar_a = sbase + OFF_BAR, bar_full = sbase + OFF_BAR + 128, bar_empty = sbase + OFF_BAR + 192;
    const uint32_t bar_done = sbase + OFF_BAR + 256, bar_free = sbase + OFF_BAR + 272; // one per accumulator buffer
    const uint32_t bar_gdone = sbase + OFF_BAR + 288;                                  // one per A-operand buffer
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(smem + OFF_BAR + 304);
    float *s_bias = reinterpret_cast<float *>(smem + OFF_BIAS);
    float *s_wh = reinterpret_cast<float *>(smem + OFF_WH);
    const int cin = L.cin, cout = L.cout, KS = cin >> 4;
    const int NC = L.nc, NCH = cout / NC;
    const uint32_t stage_bytes = 64u * (uint32_t)NC;
    const uint32_t a_bytes = (uint32_t)(cin >> 2) * CH;
    const uint32_t nabuf = (uint32_t)L.nabuf;
    const uint32_t ring = sbase + OFF_A + nabuf * a_bytes;

    for (int i = tid; i < cout; i += THREADS) s_bias[i] = L.bias[i];
    if (MODE == 1) {
        for (int i = tid; i < cout * 8; i += THREADS) s_wh[i] = L.wh[i];
        if (tid < 8) s_wh[cout * 8 + tid] = L.bh[tid];
    }
    if (tid == 0) {
        for (int s = 0; s < 8; ++s) {
            tc_mbar_init(bar_a + 8 * s, 1);
            tc_mbar_init(bar_a + 64 + 8 * s, 1);
            tc_mbar_init(bar_full + 8 * s, 1);
            tc_mbar_init(bar_empty + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            tc_mbar_init(bar_done + 8 * b, 1);
            tc_mbar_init(bar_free + 8 * b, 8); // one arrival per epilogue warp
            tc_mbar_init(bar_gdone + 8 * b, 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sbase + OFF_BAR + 304),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    // the set-up above read only this layer's constants; the input image and the leaf count come from the
    // kernels before this one in the chain (gz_common.cuh, gz_launch_chain)
    gz_pdl_wait();
    gz_pdl_trigger();
    const long long B = __shfl_sync(0xffffffffu, count ? (long long)*count : maxB, 0);
    const long long n_groups = (B + g.NB - 1) / g.NB;

    // Work units u = 0, 1, ... of this CTA = (group, chunk) pairs in order; unit u accumulates in TMEM
    // buffer u & 1 and is phase u >> 1 of that buffer's done / free barriers.
    if (warp == 10) {
        // ---- A operand: the group's activation image, one mbarrier per K step ---------------------
        if (lane == 0) {
            uint32_t gi = 0;
            for (long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++gi) {
                const uint32_t ab = gi % nabuf, ak = gi / nabuf; // buffer and its use count
                if (ak > 0) tc_mbar_wait(bar_gdone + 8 * ab, (ak - 1) & 1u); // MMAs of the group that used this buffer are complete
                const uint8_t *src = act_in + (size_t)grp * a_bytes;
                const uint32_t dst = sbase + OFF_A + ab * a_bytes, bar = bar_a + 64 * ab;
                for (int ks = 0; ks < KS; ++ks) {
                    tc_mbar_expect_tx(bar + 8 * ks, 4u * CH);
                    const uint32_t o_hi = (uint32_t)(2 * ks) * CH, o_lo = (uint32_t)((cin >> 3) + 2 * ks) * CH;
                    tc_bulk_g2s(dst + o_hi, src + o_hi, 2u * CH, bar + 8 * ks);
                    tc_bulk_g2s(dst + o_lo, src + o_lo, 2u * CH, bar + 8 * ks);
                }
            }
        }
    } else if (warp == 9) {
        // ---- weight stages: the same NCH * KS * 9 stages for every group, streamed from L2 ---------
        if (lane == 0) {
            const uint32_t ns = (uint32_t)L.ns;
            const int n_st = NCH * KS * 9;
            uint32_t slot = 0, ph = 0;
            for (long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x)
                for (int s = 0; s < n_st; ++s) {
                    tc_mbar_wait(bar_empty + 8 * slot, ph ^ 1u);
                    tc_mbar_expect_tx(bar_full + 8 * slot, stage_bytes);
                    tc_bulk_g2s(ring + slot * stage_bytes, L.wimg + (size_t)s * stage_bytes, stage_bytes, bar_full + 8 * slot);
                    if (++slot == ns) { slot = 0; ph ^= 1u; }
                }
        }
    } else if (warp == 8) {
        // ---- MMA issue ---------------------------------------------------------------------------
        const uint32_t leader = tc_elect();
        const uint32_t ns = (uint32_t)L.ns;
        const uint32_t idesc = tc_idesc(NC);
        const uint64_t dA0 = tc_desc(sbase + OFF_A + LEAD * 16, CH);
        const uint64_t dB = tc_desc(ring, 16u * (uint32_t)NC);
        uint32_t gi = 0, u = 0, slot = 0, ph = 0;
        int shift[9];
#pragma unroll
        for (int t = 0; t < 9; ++t) shift[t] = (t / 3 - 1) * g.P + (t % 3 - 1);
        const uint32_t lo_off = (uint32_t)(cin >> 3) * (CH / 16), sb16 = stage_bytes / 16, nc2 = 2u * (uint32_t)NC;
        // GZ_TCW_TIMELINE: cycles the MMA warp of CTA 0 spends waiting for a free accumulator buffer, the A
        // operand and the weight stages, and its total loop time
        long long w_free = 0, w_a = 0, w_full = 0, t0 = 0, t1;
        const bool mark = tl != nullptr && blockIdx.x == 0;
#define TCW_T0() do { if (mark) t0 = clock64(); } while (0)
#define TCW_ACC(v) do { if (mark) { t1 = clock64(); v += t1 - t0; } } while (0)
        const long long t_begin = mark ? clock64() : 0;
        for (long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++gi) {
            const uint32_t ab = gi % nabuf, ak = gi / nabuf;
            const uint64_t dA = dA0 + (uint64_t)(ab * (a_bytes / 16));
            const uint32_t bar_ag = bar_a + 64 * ab;
            for (int ch = 0; ch < NCH; ++ch, ++u) {
                const uint32_t buf = u & 1u;
                if (u >= 2) { // the epilogue warps have drained this buffer's previous unit
                    TCW_T0();
                    tc_mbar_wait(bar_free + 8 * buf, ((u >> 1) & 1u) ^ 1u);
                    TCW_ACC(w_free);
                    tc_fence_after();
                }
                for (int ks = 0; ks < KS; ++ks) {
                    TCW_T0();
                    tc_mbar_wait(bar_ag + 8 * ks, ak & 1u);
                    TCW_ACC(w_a);
                    tc_fence_after();
                    const uint64_t a_ks = dA + (uint64_t)((2 * ks) * (CH / 16));
                    const uint32_t d0 = tmem + (uint32_t)(buf * 2 * NC), d1 = d0 + (uint32_t)NC;
#pragma unroll
                    for (int t = 0; t < 9; ++t) {
                        TCW_T0();
                        tc_mbar_wait(bar_full + 8 * slot, ph);
                        TCW_ACC(w_full);
                        tc_fence_after();
                        const uint64_t a_hi = a_ks + (int64_t)shift[t], a_lo = a_hi + lo_off;
                        const uint64_t b_hi = dB + (uint64_t)(slot * sb16), b_lo = b_hi + nc2;
                        const uint32_t acc = (ks | t) != 0;
                        tc_umma_f16(leader, d0, a_hi, b_hi, idesc, acc);
                        tc_umma_f16(leader, d0, a_lo, b_hi, idesc, 1u);
                        tc_umma_f16(leader, d0, a_hi, b_lo, idesc, 1u);
                        tc_umma_f16(leader, d1, a_hi + 128u, b_hi, idesc, acc);
                        tc_umma_f16(leader, d1, a_lo + 128u, b_hi, idesc, 1u);
                        tc_umma_f16(leader, d1, a_hi + 128u, b_lo, idesc, 1u);
                        tc_umma_commit(leader, bar_empty + 8 * slot);
                        if (++slot == ns) { slot = 0; ph ^= 1u; }
                    }
                }
                tc_umma_commit(leader, bar_done + 8 * buf);
            }
            tc_umma_commit(leader, bar_gdone + 8 * ab);
        }
        if (mark && lane == 0) {
            tl[0] = w_free; tl[1] = w_a; tl[2] = w_full; tl[3] = clock64() - t_begin; tl[4] = gi;
        }
#undef TCW_T0
#undef TCW_ACC
    } else {
        // ---- epilogue warps 0-7: warp w owns TMEM lanes 32 * (w % 4) of M-tile w / 4 ----------------
        const int tile = warp >> 2;
        const int r = tid;
        const int bq = r / g.BS, rr = r % g.BS;
        const int ci = rr / g.P, cj = rr % g.P;
        const bool geo_ok = bq < g.NB && ci < g.W && cj < g.W;
        const int cell = ci * g.W + cj;
        const uint32_t t_lane = ((uint32_t)((warp & 3) * 32)) << 16;
        uint32_t u = 0;
        for (long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x) {
            const long long board = grp * g.NB + bq;
            const bool valid = geo_ok && board < B;
            float hacc[6];
            if (MODE == 1) {
#pragma unroll
                for (int q = 0; q < 6; ++q) hacc[q] = s_wh[cout * 8 + q];
            }
            uint8_t *out = MODE == 0 ? act_out + (size_t)grp * (cout >> 2) * CH + (size_t)(LEAD + r) * 16 : nullptr;
            for (int ch = 0; ch < NCH; ++ch, ++u) {
                const uint32_t buf = u & 1u;
                tc_mbar_wait(bar_done + 8 * buf, (u >> 1) & 1u);
                tc_fence_after();
                for (int part = 0; part < (NC >> 5); ++part) {
                    const int c0 = ch * NC + part * 32; // first output channel of this 32-column slice
                    uint32_t v[32];
                    tc_tmem_ld32(tmem + t_lane + (uint32_t)((buf * 2 + tile) * NC + part * 32), v);
                    float f[32];
#pragma unroll
                    for (int c = 0; c < 32; ++c)
                        f[c] = valid ? fmaxf(fmaf(__uint_as_float(v[c]), L.inv_s, s_bias[c0 + c]), 0.f) : 0.f;
                    if (MODE == 0) {
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            uint4 hi, lo;
                            tc_split8(f + 8 * c, hi, lo);
                            *reinterpret_cast<uint4 *>(out + (size_t)((c0 >> 3) + c) * CH) = hi;
                            *reinterpret_cast<uint4 *>(out + (size_t)((cout >> 3) + (c0 >> 3) + c) * CH) = lo;
                        }
                    } else {
#pragma unroll
                        for (int c = 0; c < 32; ++c) {
                            const float4 wa = *reinterpret_cast<const float4 *>(s_wh + (c0 + c) * 8);
                            const float2 wb = *reinterpret_cast<const float2 *>(s_wh + (c0 + c) * 8 + 4);
                            hacc[0] = fmaf(f[c], wa.x, hacc[0]); hacc[1] = fmaf(f[c], wa.y, hacc[1]);
                            hacc[2] = fmaf(f[c], wa.z, hacc[2]); hacc[3] = fmaf(f[c], wa.w, hacc[3]);
                            hacc[4] = fmaf(f[c], wb.x, hacc[4]); hacc[5] = fmaf(f[c], wb.y, hacc[5]);
                        }
                    }
                    if (dbg_act && valid) {
                        float *d = dbg_act + ((size_t)board * g.S + cell) * cout + c0;
#pragma unroll
                        for (int c = 0; c < 32; ++c) d[c] = f[c];
                    }
                }
                // this unit's accumulators are consumed: release the TMEM buffer
                tc_fence_before();
                __syncwarp();
                if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_free + 8 * buf) : "memory");
            }
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
// Trunks this path takes: square boards of side 3-15; every width a multiple of 32; conv2 / conv3 input
// widths (c1, c2) at most 128 (the A operand of a group must fit shared memory beside the weight ring),
// c3 at most 256.
static bool gz_tcw_supported(const gz_engine *e)
{
    return e->W == e->H && e->W >= 3 && e->W <= 15 && e->c1 % 32 == 0 && e->c1 >= 32 && e->c1 <= 128 &&
           e->c2 % 32 == 0 && e->c2 >= 32 && e->c2 <= 128 && e->c3 % 32 == 0 && e->c3 >= 32 && e->c3 <= 256;
}

static int gz_tcw_chunk(int cout) { return cout % 128 == 0 ? 128 : (cout % 64 == 0 ? 64 : 32); }

// two A-operand buffers (the next group's image loads under this group's MMAs) when they fit beside a
// ring of at least 4 weight stages
static int gz_tcw_abufs(int cin, int cout)
{
    return tcw::SMEM_MAX - tcw::OFF_A - 2 * (cin / 4) * tcw::CH >= 4 * 64 * gz_tcw_chunk(cout) ? 2 : 1;
}

static int gz_tcw_ring_depth(int cin, int cout)
{
    const int left = tcw::SMEM_MAX - tcw::OFF_A - gz_tcw_abufs(cin, cout) * (cin / 4) * tcw::CH;
    int ns = left / (64 * gz_tcw_chunk(cout));
    if (ns > tcw::MAX_NS) ns = tcw::MAX_NS;
    return ns;
}

// weight image of one conv layer: stage (output chunk, k step ks, tap t) = [pass hi|lo][chunk 2][NC][8 halves]
static void gz_tcw_conv_image(const std::vector<float> &w, int cin, int cout, float s, std::vector<uint8_t> &img)
{
    const int NC = gz_tcw_chunk(cout), KS = cin / 16;
    const size_t stage = (size_t)64 * NC;
    img.assign((size_t)(cout / NC) * KS * 9 * stage, 0);
    for (int n = 0; n < cout; ++n)
        for (int ci = 0; ci < cin; ++ci)
            for (int t = 0; t < 9; ++t) {
                float back;
                __half hi, lo;
                gz_tc_split(w[((size_t)n * cin + ci) * 9 + t] * s, back, hi, lo);
                const int ks = ci / 16, c = (ci / 8) & 1, el = ci % 8;
                const size_t o = (size_t)(((n / NC) * KS + ks) * 9 + t) * stage + ((size_t)c * NC + n % NC) * 16 + el * 2;
                memcpy(img.data() + o, &hi, 2);
                memcpy(img.data() + o + (size_t)32 * NC, &lo, 2);
            }
}

int gz_tcw_prepare_weights(gz_engine *e, std::vector<float> h[16])
{
    const float s1 = gz_tc_scale(h[0]), s2 = gz_tc_scale(h[2]), s3 = gz_tc_scale(h[4]);
    std::vector<uint8_t> img;
    {   // conv1 as a 16-input-channel layer: channels 4-15 carry zero weights
        std::vector<float> w16((size_t)e->c1 * 16 * 9, 0.f);
        for (int n = 0; n < e->c1; ++n)
            for (int ci = 0; ci < 4; ++ci)
                for (int t = 0; t < 9; ++t) w16[((size_t)n * 16 + ci) * 9 + t] = h[0][((size_t)n * 4 + ci) * 9 + t];
        gz_tcw_conv_image(w16, 16, e->c1, s1, img);
        GZ_TRY(gz_upload(e, &e->wts.tcw_img1, img));
    }
    gz_tcw_conv_image(h[2], e->c1, e->c2, s2, img);
    GZ_TRY(gz_upload(e, &e->wts.tcw_img2, img));
    gz_tcw_conv_image(h[4], e->c2, e->c3, s3, img);
    GZ_TRY(gz_upload(e, &e->wts.tcw_img3, img));
    e->wts.tcw_inv[0] = 1.f / s2; e->wts.tcw_inv[1] = 1.f / s3; e->wts.tcw_inv[2] = 1.f / s1;
    return gz_tc_prepare_fc(e, h);
}

static int gz_tcw_ensure_buffers(gz_engine *e, int64_t groups, int64_t B, bool want_dbg)
{
    if (groups > e->tcw_cap) {
        if (e->tcw_a0) cudaFree(e->tcw_a0);
        if (e->tcw_act1) cudaFree(e->tcw_act1);
        if (e->tcw_act2) cudaFree(e->tcw_act2);
        e->tcw_a0 = e->tcw_act1 = e->tcw_act2 = nullptr;
        e->tcw_cap = 0;
        // zeroed once: the LEAD / trailing rows of every group image are never written afterwards
        const size_t b1 = (size_t)groups * (e->c1 / 4) * tcw::CH, b2 = (size_t)groups * (e->c2 / 4) * tcw::CH;
        const size_t b0 = (size_t)groups * 4 * tcw::CH;
        GZ_CUDA(e, cudaMalloc(&e->tcw_a0, b0));
        GZ_CUDA(e, cudaMemsetAsync(e->tcw_a0, 0, b0, e->stream));
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
    static const bool want_tl = getenv("GZ_TCW_TIMELINE") != nullptr; // profiling aid, off by default
    long long *d_tl = nullptr;
    if (want_tl) {
        GZ_CUDA(e, cudaMalloc(&d_tl, 24 * sizeof(long long)));
        GZ_CUDA(e, cudaMemsetAsync(d_tl, 0, 24 * sizeof(long long), e->stream));
    }
    TcwLayer L1 = {w.tcw_img1, w.bias + w.off_b1, nullptr, nullptr, w.tcw_inv[2], 16, e->c1, gz_tcw_chunk(e->c1), gz_tcw_ring_depth(16, e->c1), gz_tcw_abufs(16, e->c1)};
    TcwLayer L2 = {w.tcw_img2, w.bias + w.off_b2, nullptr, nullptr, w.tcw_inv[0], e->c1, e->c2, gz_tcw_chunk(e->c2), gz_tcw_ring_depth(e->c1, e->c2), gz_tcw_abufs(e->c1, e->c2)};
    TcwLayer L3 = {w.tcw_img3, w.bias + w.off_b3, w.wh, w.bias + w.off_bh, w.tcw_inv[1], e->c2, e->c3, gz_tcw_chunk(e->c3), gz_tcw_ring_depth(e->c2, e->c3), gz_tcw_abufs(e->c2, e->c3)};
    if (L2.ns < 2 || L3.ns < 2) { e->err = "tcgen05 wide-trunk path: weight ring does not fit"; return GZ_E_INVALID; }
    {
        ProfScope ps(e, PC_TRUNK_TC);
        auto lsmem = [](const TcwLayer &L) { return OFF_A + (size_t)L.nabuf * (L.cin / 4) * CH + (size_t)L.ns * 64 * L.nc; };
        GZ_CUDA(e, gz_launch_chain(stage_tcw_kernel, dim3((unsigned)groups), dim3(256), 0, e->stream, d_planes, e->tcw_a0,
                                   d_count, maxB, g));
        GZ_CUDA(e, gz_launch_chain(conv_tcw_kernel<0>, dim3(grid), dim3(THREADS), lsmem(L1), e->stream, e->tcw_a0, L1,
                                   e->tcw_act1, nullptr, d_count, maxB, g, dbg1, nullptr, d_tl));
        GZ_CUDA(e, gz_launch_chain(conv_tcw_kernel<0>, dim3(grid), dim3(THREADS), lsmem(L2), e->stream, e->tcw_act1, L2,
                                   e->tcw_act2, nullptr, d_count, maxB, g, nullptr, nullptr, d_tl ? d_tl + 8 : nullptr));
        GZ_CUDA(e, gz_launch_chain(conv_tcw_kernel<1>, dim3(grid), dim3(THREADS), lsmem(L3), e->stream, e->tcw_act2, L3,
                                   nullptr, reinterpret_cast<__half *>(e->tc_head), d_count, maxB, g, dbg3, dbgh,
                                   d_tl ? d_tl + 16 : nullptr));
    }
    if (d_tl) {
        long long h[24];
        GZ_CUDA(e, cudaStreamSynchronize(e->stream));
        GZ_CUDA(e, cudaMemcpy(h, d_tl, sizeof(h), cudaMemcpyDeviceToHost));
        cudaFree(d_tl);
        for (int l = 0; l < 3; ++l)
            fprintf(stderr, "tcw conv%d CTA0 MMA warp: %lld groups, loop %lld cycles; waiting: accumulators %lld, A operand %lld, weight stages %lld\n",
                    l + 1, h[8 * l + 4], h[8 * l + 3], h[8 * l], h[8 * l + 1], h[8 * l + 2]);
    }
    GZ_TRY(gz_tc_launch_heads(e, g, d_probs, d_values, d_count, maxB));
    e->launches += 5;
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
