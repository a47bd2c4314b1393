/*
 * ORACLE — TEST INFRASTRUCTURE ONLY (see orc_board.h).
 * PolicyValueNet forward restated from reference policy_value_net_pytorch.py:25-57
 * (Net.__init__/Net.forward).  The reference's arithmetic lives in torch/ATen (third
 * party, unpinned: "PyTorch >= 0.2.0", README.md:23); this is the published algorithm:
 * cross-correlation 3x3 with zero padding 1, ReLU, 1x1 convs, Linear, log_softmax, tanh.
 * Accumulation is in double, rounded to float per layer output, so the oracle sits
 * within ~1e-6 of any fp32 summation order.  Pinned against torch fp32 CPU outputs in
 * tests/golden/ and (live) against reference Net when /root/reference is present.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* Parameter order = torch state_dict order (policy_value_net_pytorch.py:31-41):
 * 0 conv1.w [c1,4,3,3] 1 conv1.b | 2 conv2.w [c2,c1,3,3] 3 b | 4 conv3.w [c3,c2,3,3] 5 b
 * 6 act_conv1.w [4,c3,1,1] 7 b | 8 act_fc1.w [S,4S] 9 b
 * 10 val_conv1.w [2,c3,1,1] 11 b | 12 val_fc1.w [64,2S] 13 b | 14 val_fc2.w [1,64] 15 b */
#define ORC_NET_MAXS 225

typedef struct {
    int W, H, c1, c2, c3;
    const float *p[16];
} orc_net;

/* Textbook cross-correlation with zero padding; per output (co, y, x) the terms are added in the order
 * ci -> ky -> kx in double.  The loops are arranged pixel-outermost with the output channels innermost over
 * a [ci][ky][kx][co] copy of the filter (contiguous doubles: vectorisable), and a zero input (about half
 * of the post-ReLU activations) skips its row of multiply-adds, which leaves every sum unchanged. */
#if defined(__GNUC__) && defined(__x86_64__)
__attribute__((target_clones("avx2", "default")))
#endif
static void conv3x3_relu(const float *in, int cin, const double *wt, const float *bias,
                         int cout, int H, int W, float *out)
{
    double acc[256];
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            for (int co = 0; co < cout; ++co) acc[co] = bias[co];
            for (int ci = 0; ci < cin; ++ci)
                for (int ky = 0; ky < 3; ++ky) {
                    const int yy = y + ky - 1;
                    if (yy < 0 || yy >= H) continue;
                    for (int kx = 0; kx < 3; ++kx) {
                        const int xx = x + kx - 1;
                        if (xx < 0 || xx >= W) continue;
                        const double v = (double)in[(ci * H + yy) * W + xx];
                        if (v == 0.0) continue;
                        const double *wr = wt + ((size_t)ci * 9 + ky * 3 + kx) * cout;
                        for (int co = 0; co < cout; ++co) acc[co] += v * wr[co];
                    }
                }
            for (int co = 0; co < cout; ++co) {
                float v = (float)acc[co];
                out[co * H * W + y * W + x] = v > 0.f ? v : 0.f;
            }
        }
}

/* OIHW float filter -> [ci][tap][co] double.  Searches call the net one board at a time, so the
 * converted filters are kept in a small per-thread cache, validated against a copy of the float
 * weights on every use (weights may be updated in place by the caller). */
#define ORC_WCACHE 6
typedef struct { const float *src; int cin, cout; float *copy; double *wt; } orc_wcache;
static __thread orc_wcache g_wc[ORC_WCACHE];
static __thread int g_wc_next;

static const double *conv_filter_t(const float *w, int cin, int cout)
{
    const size_t n = (size_t)cin * 9 * cout;
    for (int i = 0; i < ORC_WCACHE; ++i)
        if (g_wc[i].src == w && g_wc[i].cin == cin && g_wc[i].cout == cout &&
            memcmp(g_wc[i].copy, w, n * sizeof(float)) == 0)
            return g_wc[i].wt;
    orc_wcache *c = &g_wc[g_wc_next];
    g_wc_next = (g_wc_next + 1) % ORC_WCACHE;
    free(c->copy); free(c->wt);
    c->copy = (float *)malloc(n * sizeof(float));
    c->wt = (double *)malloc(n * sizeof(double));
    if (!c->copy || !c->wt) { free(c->copy); free(c->wt); memset(c, 0, sizeof(*c)); return NULL; }
    memcpy(c->copy, w, n * sizeof(float));
    c->src = w; c->cin = cin; c->cout = cout;
    for (int co = 0; co < cout; ++co)
        for (int ci = 0; ci < cin; ++ci)
            for (int t = 0; t < 9; ++t)
                c->wt[((size_t)ci * 9 + t) * cout + co] = (double)w[((size_t)co * cin + ci) * 9 + t];
    return c->wt;
}

static void conv1x1_relu(const float *in, int cin, const float *w, const float *bias,
                         int cout, int S, float *out)
{
    for (int co = 0; co < cout; ++co)
        for (int s = 0; s < S; ++s) {
            double acc = bias[co];
            for (int ci = 0; ci < cin; ++ci) acc += (double)in[ci * S + s] * (double)w[co * cin + ci];
            float v = (float)acc;
            out[co * S + s] = v > 0.f ? v : 0.f;
        }
}

static void linear(const float *in, int nin, const float *w, const float *bias, int nout,
                   float *out)
{
    for (int o = 0; o < nout; ++o) {
        double acc = bias[o];
        for (int i = 0; i < nin; ++i) acc += (double)in[i] * (double)w[o * nin + i];
        out[o] = (float)acc;
    }
}

/* planes: float[B][4][W][H] (the array policy_value_fn feeds, :103-111);
 * logp: float[B][S] log-softmax (:51); value: float[B] tanh (:56). */
int orc_net_forward(const orc_net *net, const float *planes, int B, float *logp, float *value)
{
    int H = net->W, W = net->H; /* tensor dims 2,3; square boards */
    int S = H * W;
    int cmax = net->c3 > net->c2 ? net->c3 : net->c2;
    if (net->c1 > cmax) cmax = net->c1;
    float *a = (float *)malloc(sizeof(float) * cmax * S);
    float *b = (float *)malloc(sizeof(float) * cmax * S);
    float *hp = (float *)malloc(sizeof(float) * 4 * S);
    float *hv = (float *)malloc(sizeof(float) * 2 * S);
    float *lg = (float *)malloc(sizeof(float) * S);
    float v1[64], v2[1];
    if (net->c1 > 256 || net->c2 > 256 || net->c3 > 256) return -2;
    const double *w1 = conv_filter_t(net->p[0], 4, net->c1), *w2 = conv_filter_t(net->p[2], net->c1, net->c2),
                 *w3 = conv_filter_t(net->p[4], net->c2, net->c3);
    if (!a || !b || !hp || !hv || !lg || !w1 || !w2 || !w3) return -1;
    for (int n = 0; n < B; ++n) {
        const float *x = planes + (size_t)n * 4 * S;
        conv3x3_relu(x, 4, w1, net->p[1], net->c1, H, W, a);
        conv3x3_relu(a, net->c1, w2, net->p[3], net->c2, H, W, b);
        conv3x3_relu(b, net->c2, w3, net->p[5], net->c3, H, W, a);
        /* policy head (:49-51) */
        conv1x1_relu(a, net->c3, net->p[6], net->p[7], 4, S, hp);
        linear(hp, 4 * S, net->p[8], net->p[9], S, lg);
        double mx = lg[0];
        for (int i = 1; i < S; ++i) if (lg[i] > mx) mx = lg[i];
        double sum = 0.0;
        for (int i = 0; i < S; ++i) sum += exp((double)lg[i] - mx);
        double lse = mx + log(sum);
        for (int i = 0; i < S; ++i) logp[(size_t)n * S + i] = (float)((double)lg[i] - lse);
        /* value head (:53-56) */
        conv1x1_relu(a, net->c3, net->p[10], net->p[11], 2, S, hv);
        linear(hv, 2 * S, net->p[12], net->p[13], 64, v1);
        for (int i = 0; i < 64; ++i) v1[i] = v1[i] > 0.f ? v1[i] : 0.f;
        linear(v1, 64, net->p[14], net->p[15], 1, v2);
        value[n] = (float)tanh((double)v2[0]);
    }
    free(a); free(b); free(hp); free(hv); free(lg);
    return 0;
}
