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

static void conv3x3_relu(const float *in, int cin, const float *w, const float *bias,
                         int cout, int H, int W, float *out)
{
    /* same sums as the textbook 7-deep loop, reordered so the innermost loop runs over a
     * contiguous row (vectorisable); accumulation stays in double */
    double acc[ORC_NET_MAXS];
    for (int co = 0; co < cout; ++co) {
        for (int i = 0; i < H * W; ++i) acc[i] = bias[co];
        for (int ci = 0; ci < cin; ++ci)
            for (int ky = 0; ky < 3; ++ky)
                for (int kx = 0; kx < 3; ++kx) {
                    const double wv = (double)w[((co * cin + ci) * 3 + ky) * 3 + kx];
                    const int x0 = kx == 0 ? 1 : 0, x1 = kx == 2 ? W - 1 : W;
                    for (int y = 0; y < H; ++y) {
                        const int yy = y + ky - 1;
                        if (yy < 0 || yy >= H) continue;
                        const float *src = in + (ci * H + yy) * W + (kx - 1);
                        double *dst = acc + y * W;
                        for (int x = x0; x < x1; ++x) dst[x] += (double)src[x] * wv;
                    }
                }
        for (int i = 0; i < H * W; ++i) {
            float v = (float)acc[i];
            out[co * H * W + i] = v > 0.f ? v : 0.f;
        }
    }
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
    if (!a || !b || !hp || !hv || !lg) return -1;
    for (int n = 0; n < B; ++n) {
        const float *x = planes + (size_t)n * 4 * S;
        conv3x3_relu(x, 4, net->p[0], net->p[1], net->c1, H, W, a);
        conv3x3_relu(a, net->c1, net->p[2], net->p[3], net->c2, H, W, b);
        conv3x3_relu(b, net->c2, net->p[4], net->p[5], net->c3, H, W, a);
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
