// gz_noalloc.cu — gz_net_forward_noalloc: the tcgen05 net forward (trunk_tc_kernel -> heads_tc_kernel) for callers
// that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// The only scratch the forward needs between its two kernels is the head-activation image the trunk's back end
// writes and heads_tc_kernel reads (fp16 hi/lo, heads_tc's A-operand layout: ceil(B / 128) tiles x J x 8192 bytes,
// 1.5 KB per board at 8x8).  gz_net_forward keeps it in a cudaMalloc'ed buffer that grows with B; here it lives in
// a file-scope device array of GZ_NOALLOC_SCRATCH_BYTES (8 MiB), resident on every GPU that loads the library
// until the process exits.  The entry point calls no cudaMalloc / cudaFree / allocating library routine, takes
// DEVICE pointers only (host arrays would need a staging buffer) and rejects batches whose image does not fit.
//
// Sharing: every call on a device - from any engine, any stream - uses the same array, and the array keeps the
// previous call's contents.  The part a call uses is cleared first (the K padding of the image and the rows of a
// partial last tile must be finite zeros, and another engine may have left a different geometry's image behind);
// calls that OVERLAP on the GPU (two streams, two engines) race on it: serialise them.
//
// In HOST code the array's name is a host-side stand-in, not a device address; the address a kernel can use comes
// from cudaGetSymbolAddress, looked up once per device with that device current.
// Included by gz_api.cu (single translation unit) after gz_net_tc.cu and gz_train.cu.

#define GZ_NOALLOC_SCRATCH_BYTES (8u << 20)
#define GZ_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char gz_noalloc_scratch[GZ_NOALLOC_SCRATCH_BYTES];

static void *g_noalloc_ptr[GZ_NOALLOC_MAX_DEVICES]; // device address of gz_noalloc_scratch, per device ordinal

// device address of the scratch array on the engine's device (that device is current here)
static int gz_noalloc_scratch_ptr(gz_engine *e, void **out)
{
    const int dev = e->cfg.device;
    if (dev < 0 || dev >= GZ_NOALLOC_MAX_DEVICES) { e->err = "gz_net_forward_noalloc: device ordinal out of range"; return GZ_E_INVALID; }
    if (!g_noalloc_ptr[dev]) {
        void *p = nullptr;
        cudaError_t rc = cudaGetSymbolAddress(&p, gz_noalloc_scratch);
        if (rc != cudaSuccess || !p) {
            e->err = std::string("gz_net_forward_noalloc: cudaGetSymbolAddress(gz_noalloc_scratch): ") + cudaGetErrorString(rc);
            return GZ_E_CUDA;
        }
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return GZ_OK;
}

extern "C" int gz_net_forward_noalloc(gz_engine *e, const float *planes, int64_t B, float *probs, float *values)
{
    using namespace tc;
    if (!e || !planes || !probs || !values || B < 0) return GZ_E_INVALID;
    if (B == 0) return GZ_OK;
    if (!e->wts.loaded) { e->err = "gz_set_weights has not been called"; return GZ_E_NOWEIGHTS; }
    if (!gz_tc_supported(e)) {
        e->err = "gz_net_forward_noalloc: only the fused tcgen05 path (square board, trunk 32/64/128) has a no-allocation variant";
        return GZ_E_INVALID;
    }
    if (e->tr && e->tr->init && e->tr->dirty) {
        e->err = "gz_net_forward_noalloc: weights were trained on the device since the last forward; refreshing the "
                 "inference images allocates - call gz_net_forward or gz_get_weights once first";
        return GZ_E_INVALID;
    }
    GZ_CUDA(e, cudaSetDevice(e->cfg.device));
    if (!gz_is_device_ptr(planes) || !gz_is_device_ptr(probs) || !gz_is_device_ptr(values)) {
        e->err = "gz_net_forward_noalloc: planes, probs and values must be device pointers (a host array needs a staging buffer)";
        return GZ_E_INVALID;
    }
    const TcGeom g = gz_tc_geom(e);
    const size_t tile_bytes = (size_t)g.J * 8192; // head-activation image of 128 boards
    const size_t tiles = (size_t)((B + 127) / 128);
    if (tiles * tile_bytes > (size_t)GZ_NOALLOC_SCRATCH_BYTES) {
        char msg[256];
        snprintf(msg, sizeof(msg), "gz_net_forward_noalloc: B = %lld needs %zu bytes of scratch, the static array holds %u "
                 "(at most %zu boards on a %dx%d board); use gz_net_forward", (long long)B, tiles * tile_bytes,
                 (unsigned)GZ_NOALLOC_SCRATCH_BYTES, (size_t)(GZ_NOALLOC_SCRATCH_BYTES / tile_bytes) * 128, e->W, e->H);
        e->err = msg;
        return GZ_E_INVALID;
    }
    void *scratch = nullptr;
    GZ_TRY(gz_noalloc_scratch_ptr(e, &scratch));
    GZ_CUDA(e, cudaMemsetAsync(scratch, 0, tiles * tile_bytes, e->stream));

    const bool dense = gz_tc_dense(e);
    const int smem_bytes = dense ? Lay<true>::SMEM_BYTES : Lay<false>::SMEM_BYTES;
    if (!(e->attr_done & 1u)) {
        GZ_CUDA(e, cudaFuncSetAttribute(trunk_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Lay<false>::SMEM_BYTES));
        GZ_CUDA(e, cudaFuncSetAttribute(trunk_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, Lay<true>::SMEM_BYTES));
        e->attr_done |= 1u;
    }
    if (!(e->attr_done & 2u)) {
        GZ_CUDA(e, cudaFuncSetAttribute(heads_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        e->attr_done |= 2u;
    }
    const int n_sm = e->n_sm > 0 ? e->n_sm : 148;
    const NetWeights &w = e->wts;
    const long long groups = (B + g.NB - 1) / g.NB;
    const unsigned grid = (unsigned)(groups < n_sm ? groups : n_sm);
    TcConsts cst;
    memcpy(&cst, w.tc_consts.data(), sizeof(TcConsts));
    auto kern = dense ? trunk_tc_kernel<true> : trunk_tc_kernel<false>;
    GZ_CUDA(e, gz_launch_chain(kern, dim3(grid), dim3(THREADS), (size_t)smem_bytes, e->stream, planes, w.tc_img, cst,
                               w.tc_inv[0], w.tc_inv[1], w.tc_inv[2], reinterpret_cast<__half *>(scratch),
                               (const int32_t *)nullptr, (long long)B, g, (float *)nullptr, (long long *)nullptr,
                               (const int32_t *)nullptr, (const GameBoard *)nullptr, (const GameMeta *)nullptr));
    FcTc fw;
    fw.wimg = w.tc_fcimg; fw.bp = w.bias + w.off_bp; fw.bv = w.bias + w.off_bv;
    fw.wv2 = w.bias + w.off_wv2; fw.bv2 = w.bias + w.off_bv2;
    fw.inv_p = w.tc_inv[3]; fw.inv_v = w.tc_inv[4];
    fw.S = e->S; fw.Np = (e->S + 15) & ~15; fw.stages_p = g.Jp / 2; fw.stages_v = (g.J - g.Jp) / 2;
    const int slot_bytes = fct::A_STAGE + 128 * (fw.Np > 64 ? fw.Np : 64);
    fw.ns = (200 * 1024 - fct::OFF_RING) / slot_bytes;
    if (fw.ns > fct::MAX_NS) fw.ns = fct::MAX_NS;
    if (fw.ns > fw.stages_p + fw.stages_v) fw.ns = fw.stages_p + fw.stages_v;
    const dim3 hgrid((unsigned)tiles, 2 * tiles <= (size_t)n_sm ? 2 : 1);
    GZ_CUDA(e, gz_launch_chain(heads_tc_kernel, hgrid, dim3(fct::THREADS), fct::OFF_RING + (size_t)fw.ns * slot_bytes,
                               e->stream, reinterpret_cast<const uint8_t *>(scratch), fw, probs, values,
                               (const int32_t *)nullptr, (long long)B));
    e->launches += 2;
    GZ_CUDA(e, cudaGetLastError());
    return GZ_OK; // stream-ordered: no synchronisation here (see the header); gz_sync() before reading the outputs elsewhere
}
