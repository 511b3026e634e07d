// Layer-0 module convolution on the 5th-gen tensor cores (tcgen05, kind::i8, accumulators in TMEM).
//
// Reference semantics: relu(conv2d(s, W_conv[0,j], stride) + b_conv[0,j]) * gate_j summed over the
// path's active modules in ascending j (game_ac_network.py:186-193), s = k/255 u8 frames
// (game_state.py:49-50).
//
// Formulation (k = 2*stride, stride*C = 16 bytes; 84x84x4 frames, 8x8 stride-4 filters):
//  * space-to-depth: the 8x8/4 convolution is a 2x2/1 convolution over a 21x21 grid of 4x4x4-byte
//    blocks.  A frame is staged in shared memory as 4 regions (g = raw row within a block); region g
//    holds raw rows g, g+4, ... back to back, i.e. 16 bytes per block, blocks in raster order.  The
//    copy is one cp.async.bulk per raw row: no im2col, no conversion, every frame byte is read from HBM
//    once.
//  * the A operand of tap (dy,dx) is that SAME buffer read through a shifted descriptor: row m of the
//    MMA (block index m = oy*21+ox) starts at byte (m + dy*21 + dx)*16 of each region (K-major, no
//    swizzle: 8 blocks x 16 B core matrices, SBO = 128 B, LBO = region size).  Rows with ox == 20 or
//    oy == 20 are computed and discarded (41 of 441).
//  * frames are exact u8 integers; the fp32 weights are split per (module, channel) into three signed
//    base-256 digits of round(w / max|w_c| * 2^22) (conv0_wprep_kernel, run every call since RMSProp
//    changed the weights).  Integer MMAs are exact, so the only deviation from fp32 arithmetic is the
//    2^-23 (relative to the channel's largest weight) rounding of that split.  B = [digit][channel]
//    rows per active module (24 columns per module), gathered per path from the genotype bitmask.
//  * epilogue (thread = output pixel = TMEM lane): Q = D0 + 256*D1 + 65536*D2 in int64,
//    pre = Q * scale_c + b_c, ReLU, ascending-j sum, ReLU bit mask for the backward pass.
//
// Warp roles (192 threads, one CTA per SM, persistent over a contiguous range of transitions):
//   warp 0  bulk-copy producer (frames -> smem ring), warp 1  TMEM owner + B gather + MMA issue,
//   warps 2-5  epilogue (one TMEM lane quadrant each).
#include "pn_internal.cuh"
#include "pn_tc.cuh"

using namespace pntc;

namespace {

constexpr int kTcThreads = 192;
constexpr int kAccStages = 2;
constexpr int kAccCols = 128;            // TMEM columns per accumulator stage (max N of one pass = 96)
constexpr int kSlotBytes = 3 * 2048;     // one module's B rows: 3 digit groups x (8 channels x 256 K bytes)
constexpr int kPassSlots = 4;            // modules per MMA pass
constexpr int kQMax = 1 << 22;

struct TcGeom {
    int H, W, C, s;                  // frame, stride (k = 2s, s*C = 16)
    int Hs, Ws, OH, OW;              // block grid and output grid
    int tiles;                       // 128-row MMA tiles per transition
    int region_bytes, stage_bytes;   // smem bytes of one region / one frame stage (s regions)
    int row_bytes, frame_bytes;      // W*C, H*W*C
    int nst;                         // frame stages in the ring
    int ksteps;                      // MMAs (K = 32) per tap = s*16/32
};

// ---- weight split: fp32 -> 3 signed base-256 digits per (module, channel) ---------------------------
// w8 [M][kSlotBytes] in the canonical K-major layout: element (n = digit*8 + c, kidx) at
// digit*2048 + (kidx/16)*128 + c*16 + kidx%16 with kidx = tap*64 + g*16 + b,
// (kh, kw, ci) = (dy*s + g, dx*s + b/C, b%C), tap = dy*2 + dx.   wscale [M][8] = max|w_c| / (2^22 * 255).
__global__ void __launch_bounds__(256)
conv0_wprep_kernel(const float* __restrict__ params, PnNetDesc nd, int s, int8_t* __restrict__ w8,
                   float* __restrict__ wscale) {
    __shared__ float red[8][8];
    __shared__ float mx[8];
    const int j = blockIdx.x, KK = nd.KK[0], C = nd.Cin[0], k = nd.k[0];
    const float* W = params + nd.convW_base[0] + (long long)j * nd.conv_stride[0];
    const int kk = threadIdx.x;                 // KK == 4*s*s*C == 256 for the supported geometry
    float w[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) w[c] = kk < KK ? W[kk * 8 + c] : 0.f;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        float m = fabsf(w[c]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) red[warp][c] = m;
    }
    __syncthreads();
    if (threadIdx.x < 8) {
        float m = 0.f;
        for (int i = 0; i < 8; ++i) m = fmaxf(m, red[i][threadIdx.x]);
        mx[threadIdx.x] = m;
        wscale[j * 8 + threadIdx.x] = m / ((float)kQMax * 255.0f);
    }
    __syncthreads();
    if (kk >= KK) return;
    const int kh = kk / (k * C), rem = kk % (k * C), kw = rem / C, ci = rem % C;
    const int dy = kh / s, g = kh % s, dx = kw / s, b = (kw % s) * C + ci;
    const int kidx = (dy * 2 + dx) * (s * 16) + g * 16 + b;
    int8_t* out = w8 + (size_t)j * kSlotBytes + (kidx >> 4) * 128 + (kidx & 15);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const float m = mx[c];
        int q = m > 0.f ? __float2int_rn(w[c] * ((float)kQMax / m)) : 0;
        const int d0 = ((q + 128) & 255) - 128;
        q = (q - d0) >> 8;
        const int d1 = ((q + 128) & 255) - 128;
        const int d2 = (q - d1) >> 8;
        out[0 * 2048 + c * 16] = (int8_t)d0;
        out[1 * 2048 + c * 16] = (int8_t)d1;
        out[2 * 2048 + c * 16] = (int8_t)d2;
    }
}

// ---- the convolution -----------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTcThreads, 1)
conv0_fwd_tc_kernel(const uint8_t* __restrict__ frames, const int8_t* __restrict__ w8,
                    const float* __restrict__ wscale, const float* __restrict__ params, PnNetDesc nd,
                    const uint32_t* __restrict__ masks, int B, int T, TcGeom gm, float* __restrict__ Y,
                    uint8_t* __restrict__ R) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* frame_s = smem;
    uint8_t* Bs = smem + (size_t)gm.nst * gm.stage_bytes;
    float* sbuf = reinterpret_cast<float*>(Bs + (size_t)(nd.M + 1) * kSlotBytes);   // [M][16] scale[8], bias[8]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sbuf + nd.M * 16);
    uint64_t* frame_full = bars;
    uint64_t* frame_empty = bars + gm.nst;
    uint64_t* acc_full = bars + 2 * gm.nst;
    uint64_t* acc_empty = acc_full + kAccStages;
    uint64_t* path_ready = acc_empty + kAccStages;
    uint64_t* drain = path_ready + 1;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t_lo = (int)((long long)T * blockIdx.x / gridDim.x);
    const int t_hi = (int)((long long)T * (blockIdx.x + 1) / gridDim.x);
    const int npix = gm.OH * gm.OW;

    if (threadIdx.x == 0) {
        for (int i = 0; i < gm.nst; ++i) { mbar_init(&frame_full[i], 1); mbar_init(&frame_empty[i], 1); }
        for (int i = 0; i < kAccStages; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
        mbar_init(path_ready, 1);
        mbar_init(drain, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(&tmem_base_s, kAccStages * kAccCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tmem_base_s;

    if (warp == 0) {
        // ================= producer: one bulk copy per raw frame row =================
        int stage = 0;
        uint32_t phase = 0;
        for (int t = t_lo; t < t_hi; ++t) {
            mbar_wait(&frame_empty[stage], phase ^ 1);
            if (lane == 0) mbar_expect_tx(&frame_full[stage], (uint32_t)gm.frame_bytes);
            __syncwarp();
            const uint8_t* src = frames + (size_t)t * gm.frame_bytes;
            uint8_t* dst = frame_s + (size_t)stage * gm.stage_bytes;
            for (int ry = lane; ry < gm.H; ry += 32)
                bulk_g2s(dst + (ry % gm.s) * gm.region_bytes + (ry / gm.s) * gm.row_bytes, src + (size_t)ry * gm.row_bytes,
                         (uint32_t)gm.row_bytes, &frame_full[stage]);
            if (++stage == gm.nst) { stage = 0; phase ^= 1; }
        }
    } else if (warp == 1) {
        // ================= B gather + MMA issue =================
        int stage = 0, acc = 0, cur_path = -1, ns = 0;
        uint32_t fphase = 0, drain_phase = 0;
        uint32_t fills[kAccStages] = {0, 0};
        for (int t = t_lo; t < t_hi; ++t) {
            const int p = t / B;
            if (p != cur_path) {
                if (cur_path >= 0) {
                    // every MMA that reads the old B has completed, every accumulator has been drained
                    if (lane == 0) umma_commit(drain);
                    __syncwarp();
                    mbar_wait(drain, drain_phase);
                    drain_phase ^= 1;
                    for (int a = 0; a < kAccStages; ++a) mbar_wait(&acc_empty[a], (fills[a] & 1) ^ 1);
                }
                const uint32_t mask = masks[p * nd.L];
                ns = __popc(mask);
                for (int slot = 0; slot < ns; ++slot) {
                    const int j = __fns(mask, 0, slot + 1);
                    const int4* src = reinterpret_cast<const int4*>(w8 + (size_t)j * kSlotBytes);
                    int4* dst = reinterpret_cast<int4*>(Bs + (size_t)slot * kSlotBytes);
                    for (int i = lane; i < kSlotBytes / 16; i += 32) dst[i] = src[i];
                    if (lane < 8) sbuf[slot * 16 + lane] = wscale[j * 8 + lane];
                    else if (lane < 16)
                        sbuf[slot * 16 + lane] = params[nd.convW_base[0] + (long long)j * nd.conv_stride[0] + nd.convb_off[0] + lane - 8];
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(path_ready);
                cur_path = p;
            }
            mbar_wait(&frame_full[stage], fphase);
            tc_fence_after();
            const int npass = (ns + kPassSlots - 1) / kPassSlots;
            const uint32_t fbase = smem_u32(frame_s + (size_t)stage * gm.stage_bytes);
            for (int tile = 0; tile < gm.tiles; ++tile) {
                for (int pass = 0; pass < npass; ++pass) {
                    mbar_wait(&acc_empty[acc], (fills[acc] & 1) ^ 1);
                    tc_fence_after();
                    if (lane == 0) {
                        const int nsl = min(kPassSlots, ns - pass * kPassSlots);
                        const int N = (24 * nsl + 15) & ~15;
                        const uint32_t idesc = make_idesc_i8(128, N, 0, 1, 0, 0);
                        const uint32_t bbase = smem_u32(Bs + (size_t)pass * kPassSlots * kSlotBytes);
                        uint32_t accum = 0;
                        for (int tap = 0; tap < 4; ++tap) {
                            const int off = (tap >> 1) * gm.Ws + (tap & 1);
                            for (int ks = 0; ks < gm.ksteps; ++ks) {
                                const uint64_t ad = make_smem_desc(
                                    fbase + (uint32_t)(2 * ks * gm.region_bytes + (tile * 128 + off) * 16),
                                    (uint32_t)gm.region_bytes, 128);
                                const uint64_t bd =
                                    make_smem_desc(bbase + (uint32_t)((tap * gm.s + 2 * ks) * 128), 128, 2048);
                                umma_i8(td + acc * kAccCols, ad, bd, idesc, accum);
                                accum = 1;
                            }
                        }
                        umma_commit(&acc_full[acc]);
                    }
                    __syncwarp();
                    ++fills[acc];
                    acc ^= 1;
                }
            }
            // the frame stage is free once every MMA issued so far has completed
            if (lane == 0) {
                if (npass > 0) umma_commit(&frame_empty[stage]);
                else mbar_arrive(&frame_empty[stage]);
            }
            __syncwarp();
            if (++stage == gm.nst) { stage = 0; fphase ^= 1; }
        }
    } else {
        // ================= epilogue: thread = TMEM lane = output pixel of the tile =================
        const int quad = warp & 3;
        const int L = quad * 32 + lane;
        int acc = 0, cur_path = -1, ns = 0;
        uint32_t path_phase = 0, mask = 0;
        uint32_t fills[kAccStages] = {0, 0};
        for (int t = t_lo; t < t_hi; ++t) {
            const int p = t / B;
            if (p != cur_path) {
                mbar_wait(path_ready, path_phase);
                path_phase ^= 1;
                cur_path = p;
                mask = masks[p * nd.L];
                ns = __popc(mask);
            }
            const int npass = (ns + kPassSlots - 1) / kPassSlots;
            if (npass == 0) {     // no active module in layer 0: the layer output is exactly zero
                float4* yo = reinterpret_cast<float4*>(Y + (size_t)t * npix * PN_F);
                for (int i = threadIdx.x - 64; i < npix * 2; i += 128) yo[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                continue;
            }
            for (int tile = 0; tile < gm.tiles; ++tile) {
                const int m = tile * 128 + L;
                const int oy = m / gm.Ws, ox = m - oy * gm.Ws;
                const bool valid = ox < gm.OW && oy < gm.OH;
                const int pix = oy * gm.OW + ox;
                float sum[8];
#pragma unroll
                for (int c = 0; c < 8; ++c) sum[c] = 0.f;
                for (int pass = 0; pass < npass; ++pass) {
                    mbar_wait(&acc_full[acc], fills[acc] & 1);
                    tc_fence_after();
                    const int nsl = min(kPassSlots, ns - pass * kPassSlots);
                    const uint32_t tbase = td + ((uint32_t)(quad * 32) << 16) + acc * kAccCols;
                    for (int sl = 0; sl < nsl; ++sl) {
                        uint32_t d0[8], d1[8], d2[8];
                        tmem_ld8(tbase + sl * 24, d0);
                        tmem_ld8(tbase + sl * 24 + 8, d1);
                        tmem_ld8(tbase + sl * 24 + 16, d2);
                        tmem_ld_wait();
                        const int slot = pass * kPassSlots + sl;
                        const float* sb = sbuf + slot * 16;
                        unsigned bits = 0;
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            const long long Q = (long long)(int)d0[c] + ((long long)(int)d1[c] << 8) +
                                                ((long long)(int)d2[c] << 16);
                            const float pre = fmaf((float)Q, sb[c], sb[8 + c]);
                            if (pre > 0.f) { sum[c] += pre; bits |= 1u << c; }
                        }
                        if (valid) {
                            const int j = __fns(mask, 0, slot + 1);
                            R[((size_t)t * nd.M + j) * npix + pix] = (uint8_t)bits;
                        }
                    }
                    tc_fence_before();
                    mbar_arrive(&acc_empty[acc]);
                    ++fills[acc];
                    acc ^= 1;
                }
                if (valid) {
                    float4* yo = reinterpret_cast<float4*>(Y + ((size_t)t * npix + pix) * PN_F);
                    yo[0] = make_float4(sum[0], sum[1], sum[2], sum[3]);
                    yo[1] = make_float4(sum[4], sum[5], sum[6], sum[7]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(td, kAccStages * kAccCols);
}

bool make_geom(const pn_handle* h, TcGeom& g, size_t& smem_bytes) {
    const PnNetDesc& nd = h->nd;
    if (nd.nconv < 1) return false;
    const int s = nd.s[0], k = nd.k[0], C = nd.Cin[0], H = nd.IH[0], W = nd.IW[0];
    if (k != 2 * s || s * C != 16 || (s & 1) || H % s || W % s || nd.KK[0] > 256) return false;
    g.H = H; g.W = W; g.C = C; g.s = s;
    g.Hs = H / s; g.Ws = W / s; g.OH = nd.OH[0]; g.OW = nd.OW[0];
    if (g.OH != g.Hs - 1 || g.OW != g.Ws - 1) return false;
    const int rows = (g.OH - 1) * g.Ws + g.OW;
    g.tiles = (rows + 127) / 128;
    const int region_blocks = (g.tiles * 128 + g.Ws + 1 + 7) & ~7;
    g.region_bytes = region_blocks * 16;
    g.stage_bytes = s * g.region_bytes;
    g.row_bytes = W * C;
    g.frame_bytes = H * W * C;
    g.ksteps = s / 2;
    if (g.row_bytes % 16 || g.Hs * g.Ws > region_blocks) return false;
    const size_t fixed = (size_t)(nd.M + 1) * kSlotBytes + (size_t)nd.M * 16 * sizeof(float) + 64 * sizeof(uint64_t);
    for (g.nst = 3; g.nst >= 2; --g.nst) {
        smem_bytes = (size_t)g.nst * g.stage_bytes + fixed;
        if (smem_bytes <= 220 * 1024) return true;
    }
    return false;
}

}  // namespace

bool pn_conv0_tc_supported(const pn_handle* h) {
    TcGeom g;
    size_t sm;
    return make_geom(h, g, sm);
}

size_t pn_conv0_tc_scratch_bytes(const pn_handle* h) { return (size_t)h->nd.M * (kSlotBytes + 8 * sizeof(float)); }

int pn_launch_conv0_fwd_tc(pn_handle* h, const uint8_t* frames, int B, cudaStream_t st) {
    TcGeom g;
    size_t smem_bytes = 0;
    if (!make_geom(h, g, smem_bytes)) return pn_fail(h, PN_ERR_STATE, "tensor-core conv0 not supported for this geometry");
    const PnNetDesc& nd = h->nd;
    const int T = h->P * B;
    int8_t* w8 = reinterpret_cast<int8_t*>(h->tc_w8);
    float* wscale = reinterpret_cast<float*>(h->tc_w8 + (size_t)nd.M * kSlotBytes);
    {
        PnProfScope prof(h, PN_PROF_CONV0_WPREP, st);
        conv0_wprep_kernel<<<nd.M, 256, 0, st>>>(h->params, nd, g.s, w8, wscale);
    }
    PN_LAUNCH_CHECK(h, "conv0_wprep_kernel");
    PN_CUDA(h, cudaFuncSetAttribute(conv0_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    const int grid = T < h->num_sms ? T : h->num_sms;
    {
        PnProfScope prof(h, PN_PROF_CONV_FWD + 0, st);
        conv0_fwd_tc_kernel<<<grid, kTcThreads, smem_bytes, st>>>(frames, w8, wscale, h->params, nd, h->d_masks, B, T, g,
                                                                 h->act[0], h->rmask[0]);
    }
    PN_LAUNCH_CHECK(h, "conv0_fwd_tc_kernel");
    return PN_OK;
}
