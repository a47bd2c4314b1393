// gz_augment.cu — TrainPipeline.get_equi_data (reference train.py:116-135) on the device: the 8
// dihedral images of every (state, pi, z) sample, in the reference's order
//   for i in 1..4:  (rot90^i(state), flipud(rot90^i(flipud(pi))))  then its left-right mirror.
// Pure index permutations: the host builds two gather tables per image (state cell, pi cell) by
// composing the numpy definitions (np.rot90(A)[r][c] = A[c][w-1-r], flipud, fliplr) and the kernel
// is a streaming gather, bound by HBM: 5 S + 1 floats read (once: the 8 images of a sample are made
// by one CTA) and 8 (5 S + 1) floats written per sample.
// Included by gz_api.cu (single translation unit).

// one 256-thread block per sample (grid-stride), one warp per image: the lanes sweep the cells, so the
// stores are coalesced rows of the output and the two table entries of a cell serve all five planes
__global__ void __launch_bounds__(256)
equi_gather_kernel(const float *__restrict__ states, const float *__restrict__ pis, const float *__restrict__ zs,
                   long long n, int S, const int32_t *__restrict__ tab /*[2][8][S]: state table, pi table*/,
                   float *__restrict__ out_states, float *__restrict__ out_pis, float *__restrict__ out_zs)
{
    const int t = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int32_t *ts = tab + (size_t)t * S, *tp = tab + (size_t)(8 + t) * S;
    for (long long smp = blockIdx.x; smp < n; smp += gridDim.x) {
        const float *si = states + smp * 4 * S, *pi = pis + smp * S;
        float *so = out_states + (smp * 8 + t) * 4 * S, *po = out_pis + (smp * 8 + t) * S;
        for (int j = lane; j < S; j += 32) {
            const int a = ts[j], b = tp[j];
#pragma unroll
            for (int pl = 0; pl < 4; ++pl) so[pl * S + j] = si[pl * S + a];
            po[j] = pi[b];
        }
        if (lane == 0) out_zs[smp * 8 + t] = zs[smp];
    }
}

// gather tables: image t = 2 * (i - 1) + f  (i = number of quarter turns 1..4, f = mirrored)
static void gz_equi_tables(int w, std::vector<int32_t> &tab)
{
    const int S = w * w;
    tab.assign((size_t)16 * S, 0);
    auto rot1 = [&](int &r, int &c) { const int rr = c, cc = w - 1 - r; r = rr; c = cc; }; // np.rot90: out[r][c] = in[c][w-1-r]
    for (int i = 1; i <= 4; ++i)
        for (int f = 0; f < 2; ++f) {
            const int t = 2 * (i - 1) + f;
            for (int r0 = 0; r0 < w; ++r0)
                for (int c0 = 0; c0 < w; ++c0) {
                    // state: fliplr^f(rot90^i(s))
                    int r = r0, c = c0;
                    if (f) c = w - 1 - c;
                    for (int k = 0; k < i; ++k) rot1(r, c);
                    tab[(size_t)t * S + r0 * w + c0] = r * w + c;
                    // pi: flipud(fliplr^f(rot90^i(flipud(grid))))
                    r = w - 1 - r0; c = c0;
                    if (f) c = w - 1 - c;
                    for (int k = 0; k < i; ++k) rot1(r, c);
                    r = w - 1 - r;
                    tab[(size_t)(8 + t) * S + r0 * w + c0] = r * w + c;
                }
        }
}

extern "C" int gz_equi_data(gz_engine *e, const float *states, const float *pis, const float *zs, int64_t n,
                            float *out_states, float *out_pis, float *out_zs)
{
    if (!e || n < 0 || (n > 0 && (!states || !pis || !zs || !out_states || !out_pis || !out_zs))) return GZ_E_INVALID;
    if (e->W != e->H) { e->err = "get_equi_data needs a square board (train.py:120-133 rotates the planes)"; return GZ_E_INVALID; }
    if (n == 0) return GZ_OK;
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    const size_t S = e->S, N = (size_t)n;
    std::vector<int32_t> tab;
    gz_equi_tables(e->W, tab);
    ArgStage a(e);
    GZ_TRY(a.reserve(N * (5 * S + 1) * 4 * 9 + tab.size() * 4 + 4096));
    const float *ds = a.in(states, N * 4 * S), *dp = a.in(pis, N * S), *dz = a.in(zs, N);
    const int32_t *dt = a.in(tab.data(), tab.size());
    float *os = a.out(out_states, N * 8 * 4 * S), *op = a.out(out_pis, N * 8 * S), *oz = a.out(out_zs, N * 8);
    if (a.rc) return a.rc;
    if (e->n_sm <= 0) {
        int dev = 0;
        GZ_CUDA(e, cudaGetDevice(&dev));
        GZ_CUDA(e, cudaDeviceGetAttribute(&e->n_sm, cudaDevAttrMultiProcessorCount, dev));
    }
    const long long cap_blocks = (long long)e->n_sm * 8; // 8 resident CTAs per SM, grid-stride over the samples
    const unsigned grid = (unsigned)((long long)N < cap_blocks ? (long long)N : cap_blocks);
    {
        ProfScope ps(e, PC_OTHER);
        equi_gather_kernel<<<grid, 256, 0, e->stream>>>(ds, dp, dz, (long long)N, (int)S, dt, os, op, oz);
    }
    e->launches++;
    GZ_CUDA(e, cudaGetLastError());
    return a.finish();
}
