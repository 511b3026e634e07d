// Module convolutions of the middle layers (k = 4 / stride 2 and k = 3 / stride 1, 8 -> 8 channels per module) on the
// tensor cores (the text below describes the stride-2 case; the stride-1 differences are noted at make_mid_tc_geom): tcgen05 kind::f16 with bf16 hi/lo operands (three MMAs per product, as in pn_lin_tc.cu) and fp32
// accumulation in TMEM.
//
// Reference semantics: relu(conv2d(net, W_conv[l,j], stride, VALID) + b_conv[l,j]) * gate_j summed over the
// active modules in ascending j (game_ac_network.py:186-193); gradients = tf.gradients
// (a3c_training_thread.py:55-60).
//
// Formulation.  Space-to-depth by the stride turns the 4x4/s2 convolution into a 2x2/s1 convolution over
// "blocks" of 2x2 pixels x 8 channels = 32 values; a path's whole batch is one raster of blocks
//     row R = b * RB + Y * Xp + X        (RB = Yp * Xp blocks per transition, e.g. 10 x 10)
// and the output pixel (b, oy, ox) sits on raster row m = b * RB + oy * Xp + ox, so that
//     out[m][slot, co] = sum over taps (dy, dx) and the 32 block values of  Xs2d[m + dy * Xp + dx][.] * W[tap][.][slot, co]
// i.e. four GEMMs whose A operands are the SAME staged rows shifted by dy * Xp + dx rows.  Rows with
// oy = Yp - 1 or ox = Xp - 1 are junk and never stored (19 % for 10 x 10: the tensor pipe has time to spare).
//
// Shared-memory operand ("chunk-major"): a chunk = the 8 channels of one (py, px) pixel of a block = 16 bytes of
// bf16; chunk q of row r lives at q * chunk_bytes + r * 16, so 8 consecutive rows of one chunk are one
// 128-byte tcgen05 core matrix (K-major: SBO = 128, LBO = chunk_bytes) and a tap shift is a plain
// start-address offset of 16 * (dy * Xp + dx) bytes.  Converter warps read the fp32 NHWC activations straight
// from global memory (64 contiguous bytes per (block, py)), split them into bf16 hi/lo and store the chunks;
// there is no im2col and no separate conversion pass.
#include "pn_internal.cuh"
#include "pn_tc.cuh"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <algorithm>

using namespace pntc;

namespace {

constexpr int kTileM = 128;
constexpr int kXStages = 3;
constexpr int kFwdStages = 4;          // forward kernel: two CTAs per SM, four operand stages each
constexpr int kFwdThreads = 320;      // warps 0-3 epilogue, 4-8 operand converters, 9 MMA issue + weight loads
constexpr int kConvWarps = 5;
constexpr int kConvThreads = kConvWarps * 32;
constexpr int kModBytes = 2048;       // one module's bf16 weights (hi or lo): 128 k x 8 co x 2 bytes
constexpr bool kEnableS1 = true;      // 3x3 / stride-1 layers on the tensor cores too

struct MidTcGeom {
    int IH, IW, Yp, Xp, RB, OH, OW, npix, npix_in;
    int xrows;            // staged raster rows per tile: 128 + Xp + 1 rounded up to a multiple of 8
    int chunk_bytes;      // xrows * 16
    int hl_bytes;         // 4 chunks
    int stage_bytes;      // hi + lo
    uint32_t rcp_RB, rcp_Xp;   // ceil(2^32 / d): m / d == __umulhi(m, rcp) for every m with m * d < 2^32 (raster rows here < 2^24)
    // forward K steps (16 bf16 = two 16-byte chunks each): A start offset inside a stage, distance between the two
    // chunks (the descriptor's LBO), offset of the weight block pair inside a module's operand block
    int s1;               // 1: the 3x3 / stride-1 variant (block = pixel, one chunk per raster row, Xp = IW, Yp = IH)
    int nk, max_shift;
    int a_off[8], a_lbo[8], b_off[8];
};

bool make_mid_tc_geom(const PnNetDesc& nd, int l, MidTcGeom& g) {
    if (l < 1 || l >= nd.nconv) return false;
    if (nd.Cin[l] != 8) return false;
    g.IH = nd.IH[l]; g.IW = nd.IW[l];
    g.OH = nd.OH[l]; g.OW = nd.OW[l]; g.npix = g.OH * g.OW; g.npix_in = g.IH * g.IW;
    if (nd.k[l] == 3 && nd.s[l] == 1 && kEnableS1) {
        // 3x3 / stride 1: the raster is the image itself; taps t = kh*3 + kw shift by kh*IW + kw rows.  A K step pairs
        // taps (2i, 2i+1): chunk 0 = rows at shift(2i), chunk 1 = the SAME chunk at shift(2i+1), i.e. the descriptor's
        // LBO is the difference of the two shifts; the ninth tap is paired with a zero weight block.
        g.s1 = 1; g.Yp = g.IH; g.Xp = g.IW; g.RB = g.IH * g.IW;
        if (g.OH != g.IH - 2 || g.OW != g.IW - 2 || g.IW < 3) return false;
        g.max_shift = 2 * g.IW + 2;
        g.xrows = (kTileM + g.max_shift + 7) & ~7;
        if (g.xrows > kConvThreads) return false;             // one converter thread per staged row
        g.chunk_bytes = g.xrows * 16;
        g.hl_bytes = g.chunk_bytes;
        g.stage_bytes = 2 * g.hl_bytes;
        g.nk = 5;
        for (int i = 0; i < 5; ++i) {
            const int t0 = 2 * i, t1 = 2 * i + 1;
            const int s0 = (t0 / 3) * g.IW + t0 % 3, sh1 = t1 < 9 ? (t1 / 3) * g.IW + t1 % 3 : s0 + 1;
            g.a_off[i] = s0 * 16; g.a_lbo[i] = (sh1 - s0) * 16; g.b_off[i] = t0 * 128;
        }
        g.rcp_RB = (uint32_t)(((1ull << 32) + g.RB - 1) / g.RB);
        g.rcp_Xp = (uint32_t)(((1ull << 32) + g.Xp - 1) / g.Xp);
        return true;
    }
    if (nd.k[l] != 4 || nd.s[l] != 2 || (nd.IH[l] & 1) || (nd.IW[l] & 1)) return false;
    g.s1 = 0; g.Yp = g.IH / 2; g.Xp = g.IW / 2; g.RB = g.Yp * g.Xp;
    if (g.OH != g.Yp - 1 || g.OW != g.Xp - 1 || g.Xp > 62) return false;
    g.max_shift = g.Xp + 1;
    g.xrows = (kTileM + g.Xp + 1 + 7) & ~7;
    g.chunk_bytes = g.xrows * 16;
    g.hl_bytes = 4 * g.chunk_bytes;
    g.stage_bytes = 2 * g.hl_bytes;
    g.nk = 8;
    for (int i = 0; i < 8; ++i) {
        const int tap = i >> 1, kk = i & 1;
        g.a_off[i] = 2 * kk * g.chunk_bytes + ((tap >> 1) * g.Xp + (tap & 1)) * 16;
        g.a_lbo[i] = g.chunk_bytes;
        g.b_off[i] = (tap * 4 + 2 * kk) * 128;
    }
    g.rcp_RB = (uint32_t)(((1ull << 32) + g.RB - 1) / g.RB);
    g.rcp_Xp = (uint32_t)(((1ull << 32) + g.Xp - 1) / g.Xp);
    return true;
}

// raster row -> (transition b, block row Y, block column X) without integer division (two multiply-high instead of ~40
// instructions: these decompositions were a fifth of the kernels' issued instructions)
__device__ __forceinline__ void raster_split(int R, const MidTcGeom& g, int& b, int& Y, int& X) {
    b = (int)__umulhi((uint32_t)R, g.rcp_RB);
    const int rem = R - b * g.RB;
    Y = (int)__umulhi((uint32_t)rem, g.rcp_Xp);
    X = rem - Y * g.Xp;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&v);
}
// 8 consecutive floats -> one 16-byte chunk of bf16 hi and one of bf16 lo (x ~ hi + lo to ~2^-17 relative)
__device__ __forceinline__ void split8(const float4& a, const float4& b, uint4& hi, uint4& lo) {
    const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    float h[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) h[i] = __bfloat162float(__float2bfloat16_rn(v[i]));
    hi = make_uint4(pack_bf16x2(h[0], h[1]), pack_bf16x2(h[2], h[3]), pack_bf16x2(h[4], h[5]), pack_bf16x2(h[6], h[7]));
    lo = make_uint4(pack_bf16x2(v[0] - h[0], v[1] - h[1]), pack_bf16x2(v[2] - h[2], v[3] - h[3]),
                    pack_bf16x2(v[4] - h[4], v[5] - h[5]), pack_bf16x2(v[6] - h[6], v[7] - h[7]));
}

// Forward operands are fp16 hi/lo instead (x * scale ~ hi + lo to ~2^-22 relative: 11 + 11 mantissa bits), because the
// forward pass decides the ReLU masks: with bf16 hi/lo (2^-16) a few pre-activations per million land on the other side of
// zero than in fp32 arithmetic, and one flipped unit moves a small tensor's gradient by more than the 1e-3 budget.  The
// power-of-two scales keep the lo parts out of fp16's subnormal range (activations x 2^4, weights x 2^8; the epilogue
// multiplies by 2^-12, all exact); activations must stay below 65504 / 16.
constexpr float kActScale = 16.0f, kWgtScale = 256.0f, kFwdUnscale = 1.0f / (16.0f * 256.0f);
__device__ __forceinline__ uint32_t pack_f16x2(float a, float b) {
    const __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ void split8_f16(const float4& a, const float4& b, float scale, uint4& hi, uint4& lo) {
    const float v[8] = {a.x * scale, a.y * scale, a.z * scale, a.w * scale, b.x * scale, b.y * scale, b.z * scale, b.w * scale};
    float h[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) h[i] = __half2float(__float2half_rn(v[i]));
    hi = make_uint4(pack_f16x2(h[0], h[1]), pack_f16x2(h[2], h[3]), pack_f16x2(h[4], h[5]), pack_f16x2(h[6], h[7]));
    lo = make_uint4(pack_f16x2(v[0] - h[0], v[1] - h[1]), pack_f16x2(v[2] - h[2], v[3] - h[3]),
                    pack_f16x2(v[4] - h[4], v[5] - h[5]), pack_f16x2(v[6] - h[6], v[7] - h[7]));
}

// ---- weights -> hi/lo operand blocks (wf fp16 x 2^8 for the forward pass, wd bf16 for the data gradient), every step (RMSProp changed them) ------------------------------------
// wf [M][hl][kc 16][co 8][ci 8]   forward B operand (K-major over k = kc * 8 + ci;  kc = tap * 4 + py * 2 + px)
// wd [M][hl][kc 16][ci 8][co 8]   data-gradient B operand (K-major over co)
// with (kh, kw) = (2 dy + py, 2 dx + px), tap = dy * 2 + dx; source W[(kh * 4 + kw) * 8 + ci][co].
__global__ void __launch_bounds__(256)
conv_mid_wprep_kernel(const float* __restrict__ params, PnNetDesc nd, int l, int s1, uint8_t* __restrict__ wf,
                      uint8_t* __restrict__ wd) {
    const int j = blockIdx.x;
    const float* W = params + nd.convW_base[l] + (long long)j * nd.conv_stride[l];
    const int t = threadIdx.x & 127, kc = t >> 3, e = t & 7;
    if (s1) {
        // 3x3 / stride 1: forward block kc = tap kh*3 + kw ([co][ci]), block 9 = zeros (partner of the ninth tap)
        // data-gradient block kc = tap ([ci][co], the tensor's own order), 9 blocks
        const bool fwd = threadIdx.x < 128;
        if (kc > (fwd ? 9 : 8)) return;
        float v1[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) v1[q] = kc < 9 ? (fwd ? W[(size_t)kc * 64 + q * 8 + e] : W[(size_t)kc * 64 + e * 8 + q]) : 0.f;
        uint4 h1, l1;
        if (fwd) split8_f16(make_float4(v1[0], v1[1], v1[2], v1[3]), make_float4(v1[4], v1[5], v1[6], v1[7]), kWgtScale, h1, l1);
        else split8(make_float4(v1[0], v1[1], v1[2], v1[3]), make_float4(v1[4], v1[5], v1[6], v1[7]), h1, l1);
        uint8_t* d1 = (fwd ? wf : wd) + (size_t)j * 2 * kModBytes + kc * 128 + e * 16;
        *reinterpret_cast<uint4*>(d1) = h1;
        *reinterpret_cast<uint4*>(d1 + kModBytes) = l1;
        return;
    }
    const int tap = kc >> 2, pq = kc & 3;
    const int kh = (tap >> 1) * 2 + (pq >> 1), kw = (tap & 1) * 2 + (pq & 1);
    const float* src = W + (size_t)(kh * 4 + kw) * 64;
    float v[8];
    if (threadIdx.x < 128) {
#pragma unroll
        for (int ci = 0; ci < 8; ++ci) v[ci] = src[ci * 8 + e];          // e = co
    } else {
#pragma unroll
        for (int co = 0; co < 8; ++co) v[co] = src[e * 8 + co];          // e = ci
    }
    uint4 hi, lo;
    if (threadIdx.x < 128) split8_f16(make_float4(v[0], v[1], v[2], v[3]), make_float4(v[4], v[5], v[6], v[7]), kWgtScale, hi, lo);
    else split8(make_float4(v[0], v[1], v[2], v[3]), make_float4(v[4], v[5], v[6], v[7]), hi, lo);
    uint8_t* dst = (threadIdx.x < 128 ? wf : wd) + (size_t)j * 2 * kModBytes + kc * 128 + e * 16;
    *reinterpret_cast<uint4*>(dst) = hi;
    *reinterpret_cast<uint4*>(dst + kModBytes) = lo;
}

// =====================================================================================================
// forward.  Work list = (path, tile of 128 raster rows), path-major; CTA c owns a contiguous range.
// =====================================================================================================
__global__ void __launch_bounds__(kFwdThreads, 2)
conv_mid_fwd_tc_kernel(const float* __restrict__ X, const uint8_t* __restrict__ wf, const float* __restrict__ params,
                       PnNetDesc nd, int l, MidTcGeom g, const uint32_t* __restrict__ masks, int P, int B, int ntile,
                       int w_slots, uint32_t tmem_cols, float* __restrict__ Y, uint8_t* __restrict__ R) {
    pn_pdl_sync();
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* x_full = reinterpret_cast<uint64_t*>(smem);      // [kFwdStages]
    uint64_t* x_empty = x_full + kFwdStages;                     // [kFwdStages]
    uint64_t* acc_full = x_empty + kFwdStages;                   // [2]
    uint64_t* acc_empty = acc_full + 2;                        // [2]
    uint64_t* w_full = acc_empty + 2;
    uint64_t* w_free = w_full + 1;
    uint8_t* w_s = smem + 128;                                 // [hl][w_slots][kModBytes]
    uint8_t* x_s = w_s + (size_t)2 * w_slots * kModBytes;      // [kFwdStages][stage_bytes]
    __shared__ uint32_t tmem_base_s;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const long long NT = (long long)P * ntile;
    const int g0 = (int)(NT * blockIdx.x / gridDim.x), g1 = (int)(NT * (blockIdx.x + 1) / gridDim.x);
    const int NR = B * g.RB;
    const uint32_t acc_stride = 8u * (uint32_t)w_slots;        // TMEM columns per accumulator stage

    if (threadIdx.x == 0) {
        for (int i = 0; i < kFwdStages; ++i) { mbar_init(&x_full[i], kConvWarps); mbar_init(&x_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
        mbar_init(w_full, 1); mbar_init(w_free, 1);
        fence_mbar_init();
    }
    if (warp == 9) tmem_alloc(&tmem_base_s, tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tmem_base_s;

    if (warp >= 4 && warp < 4 + kConvWarps) {
        // ================= operand converters: global fp32 NHWC -> fp16 hi/lo chunks =================
        const int cidx = threadIdx.x - 128;
        if (g.s1) {
            // stride 1: one raster row = one pixel = one chunk, and the raster IS the NHWC tensor of the path.
            // Software-pipelined like the stride-2 case below: tile i + 1 is being loaded while tile i is converted.
            float4 a[2], c[2];
            auto fetch = [&](int gt, int set) {
                const int p = gt / ntile, tile = gt - p * ntile;
                a[set] = c[set] = make_float4(0.f, 0.f, 0.f, 0.f);
                const int Rr = tile * kTileM + cidx;
                if (cidx < g.xrows && Rr < NR) {
                    const float4* src = reinterpret_cast<const float4*>(X + ((size_t)p * NR + Rr) * 8);
                    a[set] = __ldg(src); c[set] = __ldg(src + 1);
                }
            };
            auto convert = [&](int set, int it) {
                const int stage = it % kFwdStages;
                mbar_wait(&x_empty[stage], ((it / kFwdStages) & 1) ^ 1);
                if (cidx < g.xrows) {
                    uint4 hi, lo;
                    split8_f16(a[set], c[set], kActScale, hi, lo);
                    uint8_t* dst = x_s + (size_t)stage * g.stage_bytes + cidx * 16;
                    *reinterpret_cast<uint4*>(dst) = hi;
                    *reinterpret_cast<uint4*>(dst + g.hl_bytes) = lo;
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(&x_full[stage]);
            };
            if (g0 < g1) fetch(g0, 0);
            int it2 = 0;
            for (int gt = g0; gt < g1; gt += 2, it2 += 2) {
                if (gt + 1 < g1) fetch(gt + 1, 1);
                convert(0, it2);
                if (gt + 1 < g1) {
                    if (gt + 2 < g1) fetch(gt + 2, 0);
                    convert(1, it2 + 1);
                }
            }
        }
        if (!g.s1) {
            // stride 2, software-pipelined: the loads of tile i + 1 are in flight while tile i is converted.  One tile at a
            // time left every converter thread waiting out a full memory round trip per tile (long-scoreboard stalls were
            // half of this kernel's samples); two register sets cost the third resident CTA, which was worth less.
            float4 v[2][2][4];
            int rr[2][2], pyy[2][2];
            auto fetch = [&](int gt, int set) {
                const int p = gt / ntile, tile = gt - p * ntile;
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const int item = cidx + u * kConvThreads;
                    const int py = item >= g.xrows ? 1 : 0, r = item - py * g.xrows;
                    rr[set][u] = item < 2 * g.xrows ? r : -1;
                    pyy[set][u] = py;
                    const int Rr = tile * kTileM + r;
                    if (rr[set][u] >= 0 && Rr < NR) {
                        int b, Yb, Xb;
                        raster_split(Rr, g, b, Yb, Xb);
                        const float4* src = reinterpret_cast<const float4*>(
                            X + ((((size_t)p * B + b) * g.IH + 2 * Yb + py) * g.IW + 2 * Xb) * 8);
#pragma unroll
                        for (int q = 0; q < 4; ++q) v[set][u][q] = __ldg(src + q);
                    } else {
#pragma unroll
                        for (int q = 0; q < 4; ++q) v[set][u][q] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
            };
            auto convert = [&](int set, int it) {
                const int stage = it % kFwdStages;
                mbar_wait(&x_empty[stage], ((it / kFwdStages) & 1) ^ 1);
                uint8_t* xs = x_s + (size_t)stage * g.stage_bytes;
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    if (rr[set][u] >= 0) {
#pragma unroll
                        for (int px = 0; px < 2; ++px) {
                            uint4 hi, lo;
                            split8_f16(v[set][u][2 * px], v[set][u][2 * px + 1], kActScale, hi, lo);
                            uint8_t* dst = xs + (size_t)(pyy[set][u] * 2 + px) * g.chunk_bytes + rr[set][u] * 16;
                            *reinterpret_cast<uint4*>(dst) = hi;
                            *reinterpret_cast<uint4*>(dst + g.hl_bytes) = lo;
                        }
                    }
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(&x_full[stage]);
            };
            if (g0 < g1) fetch(g0, 0);
            int it2 = 0;
            for (int gt = g0; gt < g1; gt += 2, it2 += 2) {      // unrolled by two so that the register sets are static
                if (gt + 1 < g1) fetch(gt + 1, 1);
                convert(0, it2);
                if (gt + 1 < g1) {
                    if (gt + 2 < g1) fetch(gt + 2, 0);
                    convert(1, it2 + 1);
                }
            }
        }
    } else if (warp == 9) {
        // ================= weight loads + MMA issue (whole warp, uniform values, elected instructions) =================
        int cur_p = -1, ns = 0, it = 0;
        uint32_t wphase = 0, fphase = 0, idesc = 0;
        const uint32_t ws = smem_u32(w_s);
        const uint32_t w_hl = (uint32_t)w_slots * kModBytes;
        // Descriptor templates, built ONCE: this warp's dependent uniform-datapath instructions (~7 cycles each) are the
        // pace of the kernel's MMA stream, and building four descriptors from the geometry tables inside the tile loop
        // cost ~25 such instructions per K step.  Per tile only the stage offset is added (descriptor unit: 16 bytes).
        uint64_t a_t[8], b_t[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            a_t[i] = make_smem_desc(smem_u32(x_s) + (uint32_t)g.a_off[i < g.nk ? i : 0], (uint32_t)g.a_lbo[i < g.nk ? i : 0], 128);
            b_t[i] = make_smem_desc(ws + (uint32_t)g.b_off[i < g.nk ? i : 0], 128, kModBytes);
        }
        const uint64_t a_lo_adv = (uint64_t)(g.hl_bytes >> 4), b_lo_adv = (uint64_t)(w_hl >> 4);
        const uint64_t stage_adv = (uint64_t)(g.stage_bytes >> 4);
        for (int gt = g0; gt < g1; ++gt, ++it) {
            const int p = gt / ntile;
            if (p != cur_p) {
                const uint32_t mask = __reduce_or_sync(0xffffffffu, masks[p * nd.L + l]);
                ns = __popc(mask);
                if (it > 0) {                       // the previous path's MMAs still read the weight buffer
                    if (elect_one()) umma_commit(w_free);
                    __syncwarp();
                    mbar_wait(w_free, fphase);
                    fphase ^= 1;
                }
                if (ns > 0) {
                    if (elect_one()) {
                        mbar_expect_tx(w_full, (uint32_t)(ns * 2 * kModBytes));
                        for (int slot = 0; slot < ns; ++slot) {
                            const int j = __fns(mask, 0, slot + 1);
                            bulk_g2s(w_s + (size_t)slot * kModBytes, wf + (size_t)j * 2 * kModBytes, kModBytes, w_full);
                            bulk_g2s(w_s + w_hl + (size_t)slot * kModBytes, wf + (size_t)j * 2 * kModBytes + kModBytes,
                                     kModBytes, w_full);
                        }
                    }
                    __syncwarp();
                    mbar_wait(w_full, wphase);
                    wphase ^= 1;
                }
                idesc = make_idesc_f16(kTileM, 8 * ((ns + 1) & ~1), 0, 0, 0, 0);   // fp16 x fp16, both K-major
                cur_p = p;
            }
            const int stage = it % kFwdStages, as = it & 1;
            mbar_wait(&acc_empty[as], ((it >> 1) & 1) ^ 1);
            mbar_wait(&x_full[stage], (it / kFwdStages) & 1);
            tc_fence_after();
            if (ns > 0) {
                const uint64_t s_adv = stage_adv * (uint64_t)stage;
                const uint32_t dcol = td + as * acc_stride;
                if (elect_one()) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {              // K steps of the layer's geometry (g.nk of them)
                        if (i < g.nk) {
                            const uint64_t a_hi = a_t[i] + s_adv, a_lo = a_hi + a_lo_adv;
                            const uint64_t b_hi = b_t[i], b_lo = b_hi + b_lo_adv;
                            umma_f16(dcol, a_hi, b_hi, idesc, i ? 1u : 0u);
                            umma_f16(dcol, a_hi, b_lo, idesc, 1u);
                            umma_f16(dcol, a_lo, b_hi, idesc, 1u);
                        }
                    }
                    umma_commit(&x_empty[stage]);
                    umma_commit(&acc_full[as]);
                }
            } else {
                if (elect_one()) { mbar_arrive(&x_empty[stage]); mbar_arrive(&acc_full[as]); }
            }
            __syncwarp();
        }
    } else if (warp < 4) {
        // ================= epilogue: thread = raster row of the tile =================
        // The path's first four biases and module ids live in registers and are reloaded only when the path changes:
        // fetched per tile, the two dependent global loads per module sat in front of every tile's arithmetic.
        int it = 0, cur_p = -1, ns = 0;
        uint32_t mask = 0;
        float bias4[4][8];
        int j4[4] = {0, 0, 0, 0};
        for (int gt = g0; gt < g1; ++gt, ++it) {
            const int p = gt / ntile, tile = gt - p * ntile;
            if (p != cur_p) {
                cur_p = p;
                mask = masks[p * nd.L + l];
                ns = __popc(mask);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    j4[q] = q < ns ? __fns(mask, 0, q + 1) : 0;
                    const float4* bp = reinterpret_cast<const float4*>(
                        params + nd.convW_base[l] + (long long)j4[q] * nd.conv_stride[l] + nd.convb_off[l]);
                    const float4 b0 = __ldg(bp), b1 = __ldg(bp + 1);
                    bias4[q][0] = b0.x; bias4[q][1] = b0.y; bias4[q][2] = b0.z; bias4[q][3] = b0.w;
                    bias4[q][4] = b1.x; bias4[q][5] = b1.y; bias4[q][6] = b1.z; bias4[q][7] = b1.w;
                }
            }
            const int Rr = tile * kTileM + warp * 32 + lane;
            int b, Yb, Xb;
            raster_split(Rr, g, b, Yb, Xb);
            const bool valid = Rr < NR && Yb < g.OH && Xb < g.OW;
            const size_t t = (size_t)p * B + b;
            const int pix = Yb * g.OW + Xb;
            const int as = it & 1;
            float sum[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) sum[c] = 0.f;
            mbar_wait(&acc_full[as], (it >> 1) & 1);
            tc_fence_after();
#pragma unroll
            for (int s0 = 0; s0 < 4; s0 += 2) {                 // slots 0..3 from the registers
                if (s0 < ns) {
                    uint32_t d[16];
                    tmem_ld16(td + ((uint32_t)(warp * 32) << 16) + as * acc_stride + s0 * 8, d);
                    tmem_ld_wait();
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        if (s0 + u < ns) {
                            unsigned bits = 0;
#pragma unroll
                            for (int c = 0; c < 8; ++c) {
                                const float pre = fmaf(__uint_as_float(d[u * 8 + c]), kFwdUnscale, bias4[s0 + u][c]);
                                if (pre > 0.f) { sum[c] += pre; bits |= 1u << c; }
                            }
                            if (valid) R[(t * nd.M + j4[s0 + u]) * g.npix + pix] = (uint8_t)bits;
                        }
                    }
                }
            }
            for (int s0 = 4; s0 < ns; s0 += 2) {                // further slots (a mutated or OR'ed path): biases from global memory
                uint32_t d[16];
                tmem_ld16(td + ((uint32_t)(warp * 32) << 16) + as * acc_stride + s0 * 8, d);
                tmem_ld_wait();
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    if (s0 + u < ns) {
                        const int j = __fns(mask, 0, s0 + u + 1);
                        const float4* bp = reinterpret_cast<const float4*>(
                            params + nd.convW_base[l] + (long long)j * nd.conv_stride[l] + nd.convb_off[l]);
                        const float4 b0 = __ldg(bp), b1 = __ldg(bp + 1);
                        const float bias[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                        unsigned bits = 0;
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            const float pre = fmaf(__uint_as_float(d[u * 8 + c]), kFwdUnscale, bias[c]);
                            if (pre > 0.f) { sum[c] += pre; bits |= 1u << c; }
                        }
                        if (valid) R[(t * nd.M + j) * g.npix + pix] = (uint8_t)bits;
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(&acc_empty[as]);
            if (valid) {
                float4* yo = reinterpret_cast<float4*>(Y + (t * g.npix + pix) * 8);
                yo[0] = make_float4(sum[0], sum[1], sum[2], sum[3]);
                yo[1] = make_float4(sum[4], sum[5], sum[6], sum[7]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) tmem_dealloc(td, tmem_cols);
}

// =====================================================================================================
// data gradient.  In raster form
//     dX[R][(py, px, ci)] = sum over taps (dy, dx), slots, co of  dpre_slot[R - (dy * Xp + dx)][co] * W_slot[2dy+py, 2dx+px][ci][co]
// with dpre = dY * relu-mask on the output raster (junk rows, rows before the path's first transition and rows
// past its last are zero), i.e. the forward structure again: A = dpre chunks (chunk = slot, 8 co), shifted
// BACK by the tap offset; B = wd; N = 32 = one block of the input raster; K = 8 * slots per tap.
// The staged A rows start kLead = Xp + 1 rows before the tile so that every shift stays inside the stage.
// =====================================================================================================
__global__ void __launch_bounds__(kFwdThreads, 3)
conv_mid_dgrad_tc_kernel(const float* __restrict__ dY, const uint8_t* __restrict__ R, const uint8_t* __restrict__ wd,
                         PnNetDesc nd, int l, MidTcGeom g, const uint32_t* __restrict__ masks, int P, int B, int ntile,
                         int w_slots, float* __restrict__ dX, unsigned* __restrict__ absmax_out) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* x_full = reinterpret_cast<uint64_t*>(smem);      // [kXStages]
    uint64_t* x_empty = x_full + kXStages;                     // [kXStages]
    uint64_t* acc_full = x_empty + kXStages;                   // [2]
    uint64_t* acc_empty = acc_full + 2;                        // [2]
    uint64_t* w_full = acc_empty + 2;
    uint64_t* w_free = w_full + 1;
    uint8_t* w_s = smem + 128;                                 // [hl][w_slots][kModBytes]
    const int d_hl = w_slots * g.chunk_bytes;                  // bytes of the hi (or lo) chunks of one stage
    uint8_t* d_s = w_s + (size_t)2 * w_slots * kModBytes;      // [kXStages][hl][w_slots][xrows] 16-byte rows
    __shared__ uint32_t tmem_base_s;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const long long NT = (long long)P * ntile;
    const int g0 = (int)(NT * blockIdx.x / gridDim.x), g1 = (int)(NT * (blockIdx.x + 1) / gridDim.x);
    const int NR = B * g.RB;
    const int lead = g.max_shift;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kXStages; ++i) { mbar_init(&x_full[i], kConvWarps); mbar_init(&x_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
        mbar_init(w_full, 1); mbar_init(w_free, 1);
        fence_mbar_init();
    }
    if (warp == 9) tmem_alloc(&tmem_base_s, 64);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tmem_base_s;

    if (warp >= 4 && warp < 4 + kConvWarps) {
        // ================= converters: masked output gradient -> bf16 hi/lo chunks, one raster row per thread =================
        const int r = threadIdx.x - 128;
        int it = 0, cur_p = -1, ns = 0, ns_pad = 0;
        uint32_t mask = 0;
        int jq[4] = {0, 0, 0, 0};
        for (int gt = g0; gt < g1; ++gt, ++it) {
            const int p = gt / ntile, tile = gt - p * ntile;
            if (p != cur_p) {
                mask = masks[p * nd.L + l];
                ns = __popc(mask); ns_pad = (ns + 1) & ~1;
#pragma unroll
                for (int q = 0; q < 4; ++q) jq[q] = q < ns ? __fns(mask, 0, q + 1) : 0;
                cur_p = p;
            }
            const int m = tile * kTileM - lead + r;
            bool valid = r < g.xrows && m >= 0 && m < NR;
            int b = 0, pix = 0;
            if (valid) {
                int oy, ox;
                raster_split(m, g, b, oy, ox);
                valid = oy < g.OH && ox < g.OW;
                pix = oy * g.OW + ox;
            }
            const size_t t = (size_t)p * B + b;
            float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
            unsigned mk[4] = {0, 0, 0, 0};
            if (valid) {
                const float4* src = reinterpret_cast<const float4*>(dY + (t * g.npix + pix) * 8);
                v0 = __ldg(src); v1 = __ldg(src + 1);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (q < ns) mk[q] = R[(t * nd.M + jq[q]) * g.npix + pix];
            }
            const int stage = it % kXStages;
            mbar_wait(&x_empty[stage], ((it / kXStages) & 1) ^ 1);
            if (r < g.xrows) {
                uint8_t* dst = d_s + (size_t)stage * 2 * d_hl + r * 16;
                for (int slot = 0; slot < ns_pad; ++slot) {
                    unsigned bits = 0;
                    if (slot < 4) bits = slot == 0 ? mk[0] : (slot == 1 ? mk[1] : (slot == 2 ? mk[2] : mk[3]));
                    else if (valid && slot < ns) bits = R[(t * nd.M + __fns(mask, 0, slot + 1)) * g.npix + pix];
                    const float4 a = make_float4((bits & 1u) ? v0.x : 0.f, (bits & 2u) ? v0.y : 0.f, (bits & 4u) ? v0.z : 0.f,
                                                 (bits & 8u) ? v0.w : 0.f);
                    const float4 c = make_float4((bits & 16u) ? v1.x : 0.f, (bits & 32u) ? v1.y : 0.f, (bits & 64u) ? v1.z : 0.f,
                                                 (bits & 128u) ? v1.w : 0.f);
                    uint4 hi, lo;
                    split8(a, c, hi, lo);
                    *reinterpret_cast<uint4*>(dst + (size_t)slot * g.chunk_bytes) = hi;
                    *reinterpret_cast<uint4*>(dst + d_hl + (size_t)slot * g.chunk_bytes) = lo;
                }
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&x_full[stage]);
        }
    } else if (warp == 9) {
        // ================= weight loads + MMA issue =================
        int cur_p = -1, ns = 0, it = 0;
        uint32_t wphase = 0, fphase = 0;
        const uint32_t ws = smem_u32(w_s);
        const uint32_t w_hl = (uint32_t)w_slots * kModBytes;
        // stride 2: N = 32 = one 2x2x8 block of the input raster, 4 taps of 4 weight blocks each;
        // stride 1: N = 8 input channels (issued as N = 16 with SBO = 0: columns 8..15 repeat 0..7 and are ignored), 9 taps
        const uint32_t idesc = make_idesc_f16(kTileM, g.s1 ? 16 : 32, 1, 1, 0, 0);
        const int ntap = g.s1 ? 9 : 4;
        const uint32_t b_tap = g.s1 ? 128u : 512u, b_sbo = g.s1 ? 0u : 128u;
        for (int gt = g0; gt < g1; ++gt, ++it) {
            const int p = gt / ntile;
            if (p != cur_p) {
                const uint32_t mask = __reduce_or_sync(0xffffffffu, masks[p * nd.L + l]);
                ns = __popc(mask);
                if (it > 0) {
                    if (elect_one()) umma_commit(w_free);
                    __syncwarp();
                    mbar_wait(w_free, fphase);
                    fphase ^= 1;
                }
                if (ns > 0) {
                    if (ns & 1) {                   // zero the pad slot (its A chunk is zero too; stale bits could be NaN)
                        uint4* z0 = reinterpret_cast<uint4*>(w_s + (size_t)ns * kModBytes);
                        uint4* z1 = reinterpret_cast<uint4*>(w_s + w_hl + (size_t)ns * kModBytes);
                        for (int i = lane; i < kModBytes / 16; i += 32) { z0[i] = make_uint4(0, 0, 0, 0); z1[i] = make_uint4(0, 0, 0, 0); }
                        fence_proxy_async();
                        __syncwarp();
                    }
                    if (elect_one()) {
                        mbar_expect_tx(w_full, (uint32_t)(ns * 2 * kModBytes));
                        for (int slot = 0; slot < ns; ++slot) {
                            const int j = __fns(mask, 0, slot + 1);
                            bulk_g2s(w_s + (size_t)slot * kModBytes, wd + (size_t)j * 2 * kModBytes, kModBytes, w_full);
                            bulk_g2s(w_s + w_hl + (size_t)slot * kModBytes, wd + (size_t)j * 2 * kModBytes + kModBytes,
                                     kModBytes, w_full);
                        }
                    }
                    __syncwarp();
                    mbar_wait(w_full, wphase);
                    wphase ^= 1;
                }
                cur_p = p;
            }
            const int stage = it % kXStages, as = it & 1;
            mbar_wait(&acc_empty[as], ((it >> 1) & 1) ^ 1);
            mbar_wait(&x_full[stage], (it / kXStages) & 1);
            tc_fence_after();
            if (ns > 0) {
                const uint32_t ds = smem_u32(d_s + (size_t)stage * 2 * d_hl);
                const uint32_t dcol = td + as * 32;
                const int nk = (ns + 1) >> 1;
                for (int tap = 0; tap < ntap; ++tap) {
                    const int tshift = g.s1 ? (tap / 3) * g.Xp + tap % 3 : (tap >> 1) * g.Xp + (tap & 1);
                    const uint32_t shift = (uint32_t)((lead - tshift) * 16);
                    for (int kk = 0; kk < nk; ++kk) {
                        const uint32_t a_addr = ds + (uint32_t)(2 * kk) * g.chunk_bytes + shift;
                        const uint32_t b_addr = ws + (uint32_t)tap * b_tap + (uint32_t)(2 * kk) * kModBytes;
                        const uint64_t a_hi = make_smem_desc(a_addr, g.chunk_bytes, 128);
                        const uint64_t a_lo = make_smem_desc(a_addr + d_hl, g.chunk_bytes, 128);
                        const uint64_t b_hi = make_smem_desc(b_addr, kModBytes, b_sbo);
                        const uint64_t b_lo = make_smem_desc(b_addr + w_hl, kModBytes, b_sbo);
                        if (elect_one()) {
                            umma_f16(dcol, a_hi, b_hi, idesc, (tap | kk) ? 1u : 0u);
                            umma_f16(dcol, a_hi, b_lo, idesc, 1u);
                            umma_f16(dcol, a_lo, b_hi, idesc, 1u);
                        }
                    }
                }
                if (elect_one()) { umma_commit(&x_empty[stage]); umma_commit(&acc_full[as]); }
            } else {
                if (elect_one()) { mbar_arrive(&x_empty[stage]); mbar_arrive(&acc_full[as]); }
            }
            __syncwarp();
        }
    } else if (warp < 4) {
        // ================= epilogue: thread = input raster row; 32 columns = (py, px, ci) =================
        int it = 0;
        for (int gt = g0; gt < g1; ++gt, ++it) {
            const int p = gt / ntile, tile = gt - p * ntile;
            const int ns = __popc(masks[p * nd.L + l]);
            const int Rr = tile * kTileM + warp * 32 + lane;
            float amax = 0.f;                                 // max |dX| of this tile: the first conv layer's quantiser scale is per path
            int b, Yb, Xb;
            raster_split(Rr, g, b, Yb, Xb);
            const bool valid = Rr < NR;
            const int as = it & 1;
            mbar_wait(&acc_full[as], (it >> 1) & 1);
            tc_fence_after();
            uint32_t d[32];
            if (ns > 0) {
                const uint32_t ta = td + ((uint32_t)(warp * 32) << 16) + as * 32;
                tmem_ld16(ta, *reinterpret_cast<uint32_t(*)[16]>(&d[0]));
                tmem_ld16(ta + 16, *reinterpret_cast<uint32_t(*)[16]>(&d[16]));
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int c = 0; c < 32; ++c) d[c] = 0u;
            }
            tc_fence_before();
            mbar_arrive(&acc_empty[as]);
            if (valid && g.s1) {
                // stride 1: columns 0..7 = the pixel's 8 input channels; the raster is the NHWC tensor itself
                float4* o = reinterpret_cast<float4*>(dX + ((size_t)p * NR + Rr) * 8);
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const float4 v = make_float4(__uint_as_float(d[q * 4]), __uint_as_float(d[q * 4 + 1]),
                                                 __uint_as_float(d[q * 4 + 2]), __uint_as_float(d[q * 4 + 3]));
                    o[q] = v;
                    amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
                }
            } else if (valid) {
#pragma unroll
                for (int py = 0; py < 2; ++py) {
                    float4* o = reinterpret_cast<float4*>(
                        dX + ((((size_t)p * B + b) * g.IH + 2 * Yb + py) * g.IW + 2 * Xb) * 8);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float4 v = make_float4(__uint_as_float(d[py * 16 + q * 4]), __uint_as_float(d[py * 16 + q * 4 + 1]),
                                                     __uint_as_float(d[py * 16 + q * 4 + 2]), __uint_as_float(d[py * 16 + q * 4 + 3]));
                        o[q] = v;
                        amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
                    }
                }
            }
            if (absmax_out) {                                 // (a tile never straddles two paths)
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
                if (lane == 0 && amax > 0.f) atomicMax(absmax_out + p, __float_as_uint(amax));   // non-negative floats order like uints
            }
        }

    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) tmem_dealloc(td, 64);
}

// =====================================================================================================
// data gradient of the 3x3 / stride-1 layers ("col2im" form).  In raster form
//     dX[R][(py, px, ci)] = sum over taps (dy, dx), slots, co of  dpre_slot[R - shift_tap][co] * W_slot[tap][(py,px,ci)][co]
// with dpre = dY * relu-mask on the output raster (junk rows, rows before the path's first transition and rows
// past its last are zero) and shift_tap = dy * Xp + dx (stride 2) or kh * IW + kw (stride 1).
// Round 1 issued one GEMM per tap with the A operand shifted BACK by the tap (N = 32 or 8 columns: 24 .. 54 MMAs per
// 128-row tile, each costing its 4 KB A fetch).  Here the taps are COLUMNS of one GEMM on the unshifted rows,
//     D[R'][tap * nb + (py,px,ci)] = sum over (slot, co) of dpre_slot[R'][co] * W_slot[tap][(py,px,ci)][co]       (nb = 32 / 8)
// K = 8 * slots, N = 4 * 32 = 128 (stride 2) or 9 * 8 = 72 -> 80 (stride 1): 3 * slots / 2 MMAs per tile, and the epilogue
// adds the shifted rows:  dX[R] = sum over taps of D[R - shift_tap][tap block]  (through shared memory, one tap group at
// a time).  A tile computes D for 128 consecutive rows starting `lead` = max shift rows before its first output row and
// produces 128 - lead output rows.  The weight operand blocks are the ones round 1 used (block = tap, [ci][co]).
// Measured (256 paths x 32): 61 -> 40 us for the stride-1 layer.  For the 4x4 / stride-2 layer the same form reads four
// times as many accumulator columns out of TMEM (128 per row) and measured SLOWER (61 -> 88 us): that layer keeps the
// per-tap kernel above, whose tap sum happens inside the tensor core.
// =====================================================================================================
__global__ void __launch_bounds__(kFwdThreads, 2)
conv_mid_dgrad_col2im_tc_kernel(const float* __restrict__ dY, const uint8_t* __restrict__ R, const uint8_t* __restrict__ wd,
                         PnNetDesc nd, int l, MidTcGeom g, const uint32_t* __restrict__ masks, int P, int B, int ntile,
                         int w_slots, float* __restrict__ dX, unsigned* __restrict__ absmax_out) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* x_full = reinterpret_cast<uint64_t*>(smem);      // [kXStages]
    uint64_t* x_empty = x_full + kXStages;                     // [kXStages]
    uint64_t* acc_full = x_empty + kXStages;                   // [2]
    uint64_t* acc_empty = acc_full + 2;                        // [2]
    uint64_t* w_full = acc_empty + 2;
    uint64_t* w_free = w_full + 1;
    uint8_t* w_s = smem + 128;                                 // [hl][w_slots][kModBytes]
    const int d_chunk = kTileM * 16;                           // bytes of one slot's chunk of a stage (128 rows)
    const int d_hl = w_slots * d_chunk;                        // bytes of the hi (or lo) chunks of one stage
    uint8_t* d_s = w_s + (size_t)2 * w_slots * kModBytes;      // [kXStages][hl][w_slots][128] 16-byte rows
    float* c_s = reinterpret_cast<float*>(d_s + (size_t)kXStages * 2 * d_hl);    // [128][pitch] tap group of D for the shifted sum
    __shared__ uint32_t tmem_base_s;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const long long NT = (long long)P * ntile;
    const int g0 = (int)(NT * blockIdx.x / gridDim.x), g1 = (int)(NT * (blockIdx.x + 1) / gridDim.x);
    const int NR = B * g.RB;
    const int lead = g.max_shift, tstride = kTileM - lead;    // output rows per tile
    const int ncols = g.s1 ? 80 : 128;                         // accumulator columns (stride 1: 72 used)

    if (threadIdx.x == 0) {
        for (int i = 0; i < kXStages; ++i) { mbar_init(&x_full[i], kConvWarps); mbar_init(&x_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
        mbar_init(w_full, 1); mbar_init(w_free, 1);
        fence_mbar_init();
    }
    if (warp == 9) tmem_alloc(&tmem_base_s, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tmem_base_s;

    if (warp >= 4 && warp < 4 + kConvWarps) {
        // ================= converters: masked output gradient -> bf16 hi/lo chunks, one raster row per thread =================
        const int r = threadIdx.x - 128;
        int it = 0, cur_p = -1, ns = 0, ns_pad = 0;
        uint32_t mask = 0;
        int jq[4] = {0, 0, 0, 0};
        for (int gt = g0; gt < g1; ++gt, ++it) {
            const int p = gt / ntile, tile = gt - p * ntile;
            if (p != cur_p) {
                mask = masks[p * nd.L + l];
                ns = __popc(mask); ns_pad = (ns + 1) & ~1;
#pragma unroll
                for (int q = 0; q < 4; ++q) jq[q] = q < ns ? __fns(mask, 0, q + 1) : 0;
                cur_p = p;
            }
            const int m = tile * tstride - lead + r;
            bool valid = r < kTileM && m >= 0 && m < NR;
            int b = 0, pix = 0;
            if (valid) {
                int oy, ox;
                raster_split(m, g, b, oy, ox);
                valid = oy < g.OH && ox < g.OW;
                pix = oy * g.OW + ox;
            }
            const size_t t = (size_t)p * B + b;
            float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
            unsigned mk[4] = {0, 0, 0, 0};
            if (valid) {
                const float4* src = reinterpret_cast<const float4*>(dY + (t * g.npix + pix) * 8);
                v0 = __ldg(src); v1 = __ldg(src + 1);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (q < ns) mk[q] = R[(t * nd.M + jq[q]) * g.npix + pix];
            }
            const int stage = it % kXStages;
            mbar_wait(&x_empty[stage], ((it / kXStages) & 1) ^ 1);
            if (r < kTileM) {
                uint8_t* dst = d_s + (size_t)stage * 2 * d_hl + r * 16;
                for (int slot = 0; slot < ns_pad; ++slot) {
                    unsigned bits = 0;
                    if (slot < 4) bits = slot == 0 ? mk[0] : (slot == 1 ? mk[1] : (slot == 2 ? mk[2] : mk[3]));
                    else if (valid && slot < ns) bits = R[(t * nd.M + __fns(mask, 0, slot + 1)) * g.npix + pix];
                    const float4 a = make_float4((bits & 1u) ? v0.x : 0.f, (bits & 2u) ? v0.y : 0.f, (bits & 4u) ? v0.z : 0.f,
                                                 (bits & 8u) ? v0.w : 0.f);
                    const float4 c = make_float4((bits & 16u) ? v1.x : 0.f, (bits & 32u) ? v1.y : 0.f, (bits & 64u) ? v1.z : 0.f,
                                                 (bits & 128u) ? v1.w : 0.f);
                    uint4 hi, lo;
                    split8(a, c, hi, lo);
                    *reinterpret_cast<uint4*>(dst + (size_t)slot * d_chunk) = hi;
                    *reinterpret_cast<uint4*>(dst + d_hl + (size_t)slot * d_chunk) = lo;
                }
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&x_full[stage]);
        }
    } else if (warp == 9) {
        // ================= weight loads + MMA issue =================
        int cur_p = -1, ns = 0, it = 0;
        uint32_t wphase = 0, fphase = 0;
        const uint32_t ws = smem_u32(w_s);
        const uint32_t w_hl = (uint32_t)w_slots * kModBytes;
        // B: rows n = tap block * 8 + ci (stride 2: tap * 32 + (py,px,ci): 16 consecutive blocks of the module's operand
        // buffer), K = (slot, co): 8-row groups 128 B apart, the two slots of a K step kModBytes apart
        const uint32_t idesc = make_idesc_f16(kTileM, ncols, 1, 1, 0, 0);
        for (int gt = g0; gt < g1; ++gt, ++it) {
            const int p = gt / ntile;
            if (p != cur_p) {
                const uint32_t mask = __reduce_or_sync(0xffffffffu, masks[p * nd.L + l]);
                ns = __popc(mask);
                if (it > 0) {
                    if (elect_one()) umma_commit(w_free);
                    __syncwarp();
                    mbar_wait(w_free, fphase);
                    fphase ^= 1;
                }
                if (ns > 0) {
                    if (ns & 1) {                   // zero the pad slot (its A chunk is zero too; stale bits could be NaN)
                        uint4* z0 = reinterpret_cast<uint4*>(w_s + (size_t)ns * kModBytes);
                        uint4* z1 = reinterpret_cast<uint4*>(w_s + w_hl + (size_t)ns * kModBytes);
                        for (int i = lane; i < kModBytes / 16; i += 32) { z0[i] = make_uint4(0, 0, 0, 0); z1[i] = make_uint4(0, 0, 0, 0); }
                        fence_proxy_async();
                        __syncwarp();
                    }
                    if (elect_one()) {
                        mbar_expect_tx(w_full, (uint32_t)(ns * 2 * kModBytes));
                        for (int slot = 0; slot < ns; ++slot) {
                            const int j = __fns(mask, 0, slot + 1);
                            bulk_g2s(w_s + (size_t)slot * kModBytes, wd + (size_t)j * 2 * kModBytes, kModBytes, w_full);
                            bulk_g2s(w_s + w_hl + (size_t)slot * kModBytes, wd + (size_t)j * 2 * kModBytes + kModBytes,
                                     kModBytes, w_full);
                        }
                    }
                    __syncwarp();
                    mbar_wait(w_full, wphase);
                    wphase ^= 1;
                }
                cur_p = p;
            }
            const int stage = it % kXStages, as = it & 1;
            mbar_wait(&acc_empty[as], ((it >> 1) & 1) ^ 1);
            mbar_wait(&x_full[stage], (it / kXStages) & 1);
            tc_fence_after();
            if (ns > 0) {
                const uint32_t ds = smem_u32(d_s + (size_t)stage * 2 * d_hl);
                const uint32_t dcol = td + as * 128;
                const int nk = (ns + 1) >> 1;
                if (elect_one()) {
                    for (int kk = 0; kk < nk; ++kk) {
                        const uint32_t a_addr = ds + (uint32_t)(2 * kk) * d_chunk;
                        const uint32_t b_addr = ws + (uint32_t)(2 * kk) * kModBytes;
                        const uint64_t a_hi = make_smem_desc(a_addr, d_chunk, 128);
                        const uint64_t a_lo = make_smem_desc(a_addr + d_hl, d_chunk, 128);
                        const uint64_t b_hi = make_smem_desc(b_addr, kModBytes, 128);
                        const uint64_t b_lo = make_smem_desc(b_addr + w_hl, kModBytes, 128);
                        umma_f16(dcol, a_hi, b_hi, idesc, kk ? 1u : 0u);
                        umma_f16(dcol, a_hi, b_lo, idesc, 1u);
                        umma_f16(dcol, a_lo, b_hi, idesc, 1u);
                    }
                    umma_commit(&x_empty[stage]);
                    umma_commit(&acc_full[as]);
                }
            } else {
                if (elect_one()) { mbar_arrive(&x_empty[stage]); mbar_arrive(&acc_full[as]); }
            }
            __syncwarp();
        }
    } else if (warp < 4) {
        // ================= epilogue: thread = row of D; output row o = thread (o < tstride) =================
        const int rD = warp * 32 + lane;
        const int pitch = g.s1 ? 28 : 36;                      // floats per row of c_s (16-byte aligned, conflict-free for LDS.128 / STS.128)
        int it = 0;
        for (int gt = g0; gt < g1; ++gt, ++it) {
            const int p = gt / ntile, tile = gt - p * ntile;
            const int ns = __popc(masks[p * nd.L + l]);
            const int Rr = tile * tstride + rD;                 // output raster row of this thread
            const bool valid = rD < tstride && Rr < NR;
            const int as = it & 1;
            float amax = 0.f;                                 // max |dX| of this tile: the first conv layer's quantiser scale is per path
            mbar_wait(&acc_full[as], (it >> 1) & 1);
            tc_fence_after();
            const uint32_t ta = td + ((uint32_t)(warp * 32) << 16) + as * 128;
            if (g.s1) {
                float acc[8];
#pragma unroll
                for (int c = 0; c < 8; ++c) acc[c] = 0.f;
                for (int kh = 0; kh < 3; ++kh) {
                    if (ns > 0) {
                        uint32_t d[24];
                        tmem_ld8(ta + kh * 24, *reinterpret_cast<uint32_t(*)[8]>(&d[0]));
                        tmem_ld8(ta + kh * 24 + 8, *reinterpret_cast<uint32_t(*)[8]>(&d[8]));
                        tmem_ld8(ta + kh * 24 + 16, *reinterpret_cast<uint32_t(*)[8]>(&d[16]));
                        tmem_ld_wait();
                        float4* row = reinterpret_cast<float4*>(c_s + rD * pitch);
#pragma unroll
                        for (int q = 0; q < 6; ++q)
                            row[q] = make_float4(__uint_as_float(d[4 * q]), __uint_as_float(d[4 * q + 1]), __uint_as_float(d[4 * q + 2]),
                                                 __uint_as_float(d[4 * q + 3]));
                    }
                    if (kh == 2) { tc_fence_before(); mbar_arrive(&acc_empty[as]); }      // the accumulator stage has been read
                    asm volatile("bar.sync 2, 128;" ::: "memory");
                    if (valid && ns > 0) {
#pragma unroll
                        for (int kw = 0; kw < 3; ++kw) {
                            const float4* src = reinterpret_cast<const float4*>(c_s + (rD + lead - (kh * g.Xp + kw)) * pitch + kw * 8);
                            const float4 a = src[0], b4 = src[1];
                            acc[0] += a.x; acc[1] += a.y; acc[2] += a.z; acc[3] += a.w;
                            acc[4] += b4.x; acc[5] += b4.y; acc[6] += b4.z; acc[7] += b4.w;
                        }
                    }
                    asm volatile("bar.sync 2, 128;" ::: "memory");
                }
                if (valid) {
                    // stride 1: the raster is the NHWC tensor of the path itself
                    float4* o = reinterpret_cast<float4*>(dX + ((size_t)p * NR + Rr) * 8);
                    o[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
                    o[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
#pragma unroll
                    for (int c = 0; c < 8; ++c) amax = fmaxf(amax, fabsf(acc[c]));
                }
            } else {
                float acc[32];
#pragma unroll
                for (int c = 0; c < 32; ++c) acc[c] = 0.f;
                for (int tap = 0; tap < 4; ++tap) {
                    if (ns > 0) {
                        uint32_t d[32];
                        tmem_ld16(ta + tap * 32, *reinterpret_cast<uint32_t(*)[16]>(&d[0]));
                        tmem_ld16(ta + tap * 32 + 16, *reinterpret_cast<uint32_t(*)[16]>(&d[16]));
                        tmem_ld_wait();
                        float4* row = reinterpret_cast<float4*>(c_s + rD * pitch);
#pragma unroll
                        for (int q = 0; q < 8; ++q)
                            row[q] = make_float4(__uint_as_float(d[4 * q]), __uint_as_float(d[4 * q + 1]), __uint_as_float(d[4 * q + 2]),
                                                 __uint_as_float(d[4 * q + 3]));
                    }
                    if (tap == 3) { tc_fence_before(); mbar_arrive(&acc_empty[as]); }
                    asm volatile("bar.sync 2, 128;" ::: "memory");
                    if (valid && ns > 0) {
                        const float4* src = reinterpret_cast<const float4*>(c_s + (rD + lead - ((tap >> 1) * g.Xp + (tap & 1))) * pitch);
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const float4 a = src[q];
                            acc[4 * q] += a.x; acc[4 * q + 1] += a.y; acc[4 * q + 2] += a.z; acc[4 * q + 3] += a.w;
                        }
                    }
                    asm volatile("bar.sync 2, 128;" ::: "memory");
                }
                if (valid) {
                    int b, Yb, Xb;
                    raster_split(Rr, g, b, Yb, Xb);
#pragma unroll
                    for (int py = 0; py < 2; ++py) {
                        float4* o = reinterpret_cast<float4*>(
                            dX + ((((size_t)p * B + b) * g.IH + 2 * Yb + py) * g.IW + 2 * Xb) * 8);
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            o[q] = make_float4(acc[py * 16 + q * 4], acc[py * 16 + q * 4 + 1], acc[py * 16 + q * 4 + 2], acc[py * 16 + q * 4 + 3]);
                    }
#pragma unroll
                    for (int c = 0; c < 32; ++c) amax = fmaxf(amax, fabsf(acc[c]));
                }
            }
            if (absmax_out) {                                 // (a tile never straddles two paths)
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
                if (lane == 0 && amax > 0.f) atomicMax(absmax_out + p, __float_as_uint(amax));   // non-negative floats order like uints
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) tmem_dealloc(td, 256);
}

// =====================================================================================================
// weight gradient.  d W_slot[(2dy+py, 2dx+px)][ci][co] = sum over raster rows m of  dpre_slot[m][co] * Xs2d[m + dy*Xp + dx][(py,px,ci)]
// as D[(hl, slot, co) = 64 rows][tap * 32 + (py, px, ci) = 128 columns], accumulated in TMEM over all tiles of
// one (path, chunk) item: A = dpre chunks (MN-major, chunk = hi slot 0..3 | lo slot 0..3), B = the forward's X
// chunks (MN-major, N = 32 per tap, the tap being a start-address shift), K = 16 raster rows per MMA.
//   D rows 0..31  = dpre_hi * (X_hi + X_lo),  rows 32..63 = dpre_lo * (X_hi + X_lo); the flush adds the halves.
// The bias gradient (plain row sums of dpre) is accumulated in fp32 registers by the dpre converter threads
// (a fixed row per thread) and reduced in a fixed order.  Four active modules per pass; a path with more owns
// further blocks (one per pass), so the dispatch is per path, not per population.
// Warps 0-3: dpre converters, then flush; warps 4-8: X converters; warp 9: MMA issue.
// =====================================================================================================
constexpr int kWgStages = 2;
constexpr int kDChunk = kTileM * 16;          // bytes of one dpre chunk of a tile

__global__ void __launch_bounds__(kFwdThreads, 3)
conv_mid_wgrad_tc_kernel(const float* __restrict__ X, const float* __restrict__ dY, const uint8_t* __restrict__ R,
                         PnNetDesc nd, int l, MidTcGeom g, const uint32_t* __restrict__ masks, int B, int ntile, int nch,
                         int nch_alloc, int npass, float* __restrict__ part) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* x_full = reinterpret_cast<uint64_t*>(smem);      // [kWgStages]
    uint64_t* x_empty = x_full + kWgStages;                    // [kWgStages]
    uint64_t* acc_full = x_empty + kWgStages;
    uint8_t* st_s = smem + 128;                                // [kWgStages][ X hi/lo chunks | dpre 8 chunks ]
    const int stage_bytes = g.stage_bytes + 8 * kDChunk;
    __shared__ uint32_t tmem_base_s;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    // block -> (path, chunk of tiles, pass): a pass covers four of the path's active modules, so a path that a mutation
    // or the frozen path of an earlier task has given more than four modules in this layer simply owns more blocks
    const int pass = blockIdx.x % npass, pc = blockIdx.x / npass;
    const int p = pc / nch, ch = pc - p * nch;
    const int t_lo = (int)((long long)ntile * ch / nch), t_hi = (int)((long long)ntile * (ch + 1) / nch);
    const uint32_t mask = masks[p * nd.L + l];
    const int s0 = 4 * pass;
    const int ns = min(4, __popc(mask) - s0);                  // modules of this pass
    if (ns <= 0 || t_lo >= t_hi) return;                       // nothing to do: these partials are never read
    const int NR = B * g.RB;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kWgStages; ++i) { mbar_init(&x_full[i], 9); mbar_init(&x_empty[i], 1); }
        mbar_init(acc_full, 1);
        fence_mbar_init();
    }
    if (warp == 9) tmem_alloc(&tmem_base_s, 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tmem_base_s;

    if (warp < 4) {
        // ================= dpre converters (one raster row per thread), then flush =================
        const int r = threadIdx.x;
        int jq[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) jq[q] = q < ns ? __fns(mask, 0, s0 + q + 1) : 0;
        float bacc[4][8];
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int c = 0; c < 8; ++c) bacc[q][c] = 0.f;
        int it = 0;
        for (int tile = t_lo; tile < t_hi; ++tile, ++it) {
            const int m = tile * kTileM + r;
            bool valid = m < NR;
            int b = 0, pix = 0;
            if (valid) {
                int oy, ox;
                raster_split(m, g, b, oy, ox);
                valid = oy < g.OH && ox < g.OW;
                pix = oy * g.OW + ox;
            }
            const size_t t = (size_t)p * B + b;
            float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
            unsigned mk[4] = {0, 0, 0, 0};
            if (valid) {
                const float4* src = reinterpret_cast<const float4*>(dY + (t * g.npix + pix) * 8);
                v0 = __ldg(src); v1 = __ldg(src + 1);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (q < ns) mk[q] = R[(t * nd.M + jq[q]) * g.npix + pix];
            }
            const int stage = it % kWgStages;
            mbar_wait(&x_empty[stage], ((it / kWgStages) & 1) ^ 1);
            uint8_t* dst = st_s + (size_t)stage * stage_bytes + g.stage_bytes + r * 16;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const unsigned bits = mk[q];
                const float4 a = make_float4((bits & 1u) ? v0.x : 0.f, (bits & 2u) ? v0.y : 0.f, (bits & 4u) ? v0.z : 0.f,
                                             (bits & 8u) ? v0.w : 0.f);
                const float4 c = make_float4((bits & 16u) ? v1.x : 0.f, (bits & 32u) ? v1.y : 0.f, (bits & 64u) ? v1.z : 0.f,
                                             (bits & 128u) ? v1.w : 0.f);
                bacc[q][0] += a.x; bacc[q][1] += a.y; bacc[q][2] += a.z; bacc[q][3] += a.w;
                bacc[q][4] += c.x; bacc[q][5] += c.y; bacc[q][6] += c.z; bacc[q][7] += c.w;
                uint4 hi, lo;
                split8(a, c, hi, lo);
                *reinterpret_cast<uint4*>(dst + (size_t)q * kDChunk) = hi;
                *reinterpret_cast<uint4*>(dst + (size_t)(4 + q) * kDChunk) = lo;
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&x_full[stage]);
        }
        // ---- flush: TMEM -> shared (fp32 [64 rows][129]) -> hi + lo -> partial buffer ----
        mbar_wait(acc_full, 0);
        tc_fence_after();
        float* fs = reinterpret_cast<float*>(st_s);            // the stages are free: every MMA has completed
        constexpr int kPitch = 129;
        {
            // M = 64: accumulator row q sits in TMEM lane (q / 16) * 32 + q % 16 -> lanes 0..15 of each warp's quadrant
            const int row = warp * 16 + lane;
#pragma unroll
            for (int c0 = 0; c0 < 128; c0 += 16) {
                uint32_t d[16];
                tmem_ld16(td + ((uint32_t)(warp * 32) << 16) + c0, d);
                tmem_ld_wait();
                if (lane < 16) {
#pragma unroll
                    for (int c = 0; c < 16; ++c) fs[row * kPitch + c0 + c] = __uint_as_float(d[c]);
                }
            }
        }
        float* bs = fs + 64 * kPitch;                          // [128 threads][33] bias partials
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int c = 0; c < 8; ++c) bs[r * 33 + q * 8 + c] = bacc[q][c];
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const int KK = nd.KK[l];                               // 128 (4x4x8) or 72 (3x3x8)
        const int stride_part = KK * 8 + 8;
        for (int slot = 0; slot < ns; ++slot) {
            float* out = part + ((size_t)(p * nch_alloc + ch) * nd.M + __fns(mask, 0, s0 + slot + 1)) * stride_part;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int idx = i * 128 + r;                   // = ((kh * k + kw) * 8 + ci) * 8 + co
                if (idx < KK * 8) {
                    const int co = idx & 7;
                    int n;
                    if (g.s1) {
                        n = idx >> 3;                          // column kh * 24 + kw * 8 + ci == im2col row (kh * 3 + kw) * 8 + ci
                    } else {
                        const int ci = (idx >> 3) & 7, khkw = idx >> 6, kh = khkw >> 2, kw = khkw & 3;
                        n = ((kh >> 1) * 2 + (kw >> 1)) * 32 + ((kh & 1) * 2 + (kw & 1)) * 8 + ci;
                    }
                    const int row = slot * 8 + co;
                    out[idx] = fs[row * kPitch + n] + fs[(32 + row) * kPitch + n];
                }
            }
            if (r < 8) {
                float sacc = 0.f;
                for (int k = 0; k < 128; ++k) sacc += bs[k * 33 + slot * 8 + r];
                out[KK * 8 + r] = sacc;
            }
        }
    } else if (warp < 4 + kConvWarps) {
        // ================= X converters (as in the forward kernel) =================
        const int cidx = threadIdx.x - 128;
        int it = 0;
        for (int tile = t_lo; tile < t_hi; ++tile, ++it) {
            if (g.s1) {                 // stride 1: one raster row = one pixel = one chunk (as in the forward kernel)
                float4 a = make_float4(0.f, 0.f, 0.f, 0.f), c = a;
                const int Rr = tile * kTileM + cidx;
                if (cidx < g.xrows && Rr < NR) {
                    const float4* src = reinterpret_cast<const float4*>(X + ((size_t)p * NR + Rr) * 8);
                    a = __ldg(src); c = __ldg(src + 1);
                }
                const int stage = it % kWgStages;
                mbar_wait(&x_empty[stage], ((it / kWgStages) & 1) ^ 1);
                if (cidx < g.xrows) {
                    uint4 hi, lo;
                    split8(a, c, hi, lo);
                    uint8_t* dst = st_s + (size_t)stage * stage_bytes + cidx * 16;
                    *reinterpret_cast<uint4*>(dst) = hi;
                    *reinterpret_cast<uint4*>(dst + g.hl_bytes) = lo;
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(&x_full[stage]);
                continue;
            }
            float4 v[2][4];
            int rr[2], pyy[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int item = cidx + u * kConvThreads;
                const int py = item >= g.xrows ? 1 : 0, r = item - py * g.xrows;
                rr[u] = item < 2 * g.xrows ? r : -1;
                pyy[u] = py;
                const int Rr = tile * kTileM + r;
                if (rr[u] >= 0 && Rr < NR) {
                    int b, Yb, Xb;
            raster_split(Rr, g, b, Yb, Xb);
                    const float4* src = reinterpret_cast<const float4*>(
                        X + ((((size_t)p * B + b) * g.IH + 2 * Yb + py) * g.IW + 2 * Xb) * 8);
#pragma unroll
                    for (int q = 0; q < 4; ++q) v[u][q] = __ldg(src + q);
                } else {
#pragma unroll
                    for (int q = 0; q < 4; ++q) v[u][q] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            const int stage = it % kWgStages;
            mbar_wait(&x_empty[stage], ((it / kWgStages) & 1) ^ 1);
            uint8_t* xs = st_s + (size_t)stage * stage_bytes;
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                if (rr[u] >= 0) {
#pragma unroll
                    for (int px = 0; px < 2; ++px) {
                        uint4 hi, lo;
                        split8(v[u][2 * px], v[u][2 * px + 1], hi, lo);
                        uint8_t* dst = xs + (size_t)(pyy[u] * 2 + px) * g.chunk_bytes + rr[u] * 16;
                        *reinterpret_cast<uint4*>(dst) = hi;
                        *reinterpret_cast<uint4*>(dst + g.hl_bytes) = lo;
                    }
                }
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&x_full[stage]);
        }
    } else {
        // ================= MMA issue =================
        // stride 2: four taps of N = 32 (one block); stride 1: three kernel rows of N = 24 = (kw, ci): the three kw taps are one
        // row apart, i.e. 16 bytes, which is the B operand's SBO (n / 8 = kw)
        const uint32_t idesc = make_idesc_f16(64, g.s1 ? 24 : 32, 1, 1, 1, 1);         // bf16, A and B MN-major
        const int ngrp = g.s1 ? 3 : 4;
        const uint32_t b_sbo = g.s1 ? 16u : (uint32_t)g.chunk_bytes, dstep = g.s1 ? 24u : 32u;
        int it = 0;
        for (int tile = t_lo; tile < t_hi; ++tile, ++it) {
            const int stage = it % kWgStages;
            mbar_wait(&x_full[stage], (it / kWgStages) & 1);
            tc_fence_after();
            const uint32_t xs = smem_u32(st_s + (size_t)stage * stage_bytes);
            const uint32_t ds = xs + (uint32_t)g.stage_bytes;
#pragma unroll 2
            for (int ks = 0; ks < kTileM / 16; ++ks) {
                const uint64_t a_d = make_smem_desc(ds + (uint32_t)(ks * 256), 128, kDChunk);
#pragma unroll
                for (int tap = 0; tap < 4; ++tap) {
                    if (tap < ngrp) {
                        const int tshift = g.s1 ? tap * g.Xp : (tap >> 1) * g.Xp + (tap & 1);
                        const uint32_t b_addr = xs + (uint32_t)(tshift * 16 + ks * 256);
                        const uint64_t b_hi = make_smem_desc(b_addr, 128, b_sbo);
                        const uint64_t b_lo = make_smem_desc(b_addr + g.hl_bytes, 128, b_sbo);
                        if (elect_one()) {
                            umma_f16(td + tap * dstep, a_d, b_hi, idesc, (it | ks) ? 1u : 0u);
                            umma_f16(td + tap * dstep, a_d, b_lo, idesc, 1u);
                        }
                    }
                }
            }
            if (elect_one()) umma_commit(&x_empty[stage]);
            __syncwarp();
        }
        if (elect_one()) umma_commit(acc_full);
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) tmem_dealloc(td, 128);
}

uint32_t pow2_cols(uint32_t n) {
    uint32_t c = 32;
    while (c < n) c <<= 1;
    return c;
}

}  // namespace

bool pn_conv_mid_tc_supported(const pn_handle* h, int l) {
    MidTcGeom g;
    if (!make_mid_tc_geom(h->nd, l, g)) return false;
    const int w_slots = (h->nd.M + 1) & ~1;
    const size_t smem = 128 + (size_t)2 * w_slots * kModBytes + (size_t)kFwdStages * g.stage_bytes;
    return smem <= 113 * 1024 && 16 * w_slots <= 512 && h->nd.convb_off[l] % 4 == 0;
}

size_t pn_conv_mid_tc_scratch_bytes(const pn_handle* h, int l) {
    (void)l;
    return (size_t)h->nd.M * 2 * kModBytes * 2;       // wf | wd
}

int pn_launch_conv_mid_wprep(pn_handle* h, int l, cudaStream_t st) {
    const PnNetDesc& nd = h->nd;
    uint8_t* wf = h->tc_mid_buf[l];
    uint8_t* wd = wf + (size_t)nd.M * 2 * kModBytes;
    {
        PnProfScope prof(h, PN_PROF_CONV_MID_PREP, st);
        MidTcGeom g;
        if (!make_mid_tc_geom(nd, l, g)) return pn_fail(h, PN_ERR_STATE, "tensor-core middle conv layer not supported");
        conv_mid_wprep_kernel<<<nd.M, 256, 0, st>>>(h->params, nd, l, g.s1, wf, wd);
    }
    PN_LAUNCH_CHECK(h, "conv_mid_wprep_kernel");
    return PN_OK;
}

int pn_launch_conv_mid_fwd_tc(pn_handle* h, int l, int B, cudaStream_t st) {
    const PnNetDesc& nd = h->nd;
    MidTcGeom g;
    if (!make_mid_tc_geom(nd, l, g)) return pn_fail(h, PN_ERR_STATE, "tensor-core middle conv layer not supported");
    const uint8_t* wf = h->tc_mid_buf[l];
    const int w_slots = std::max(2, (h->max_active[l] + 1) & ~1);
    const uint32_t tmem_cols = pow2_cols(16u * w_slots);
    const size_t smem = 128 + (size_t)2 * w_slots * kModBytes + (size_t)kFwdStages * g.stage_bytes;
    PN_SET_SMEM(h, conv_mid_fwd_tc_kernel, smem);
    const int ntile = (B * g.RB + kTileM - 1) / kTileM;
    const long long NT = (long long)h->P * ntile;
    const int grid = (int)std::min<long long>(NT, 2ll * h->num_sms);
    {
        PnProfScope prof(h, PN_PROF_CONV_FWD + l, st);
        pn_launch(pn_use_pdl(h), conv_mid_fwd_tc_kernel, dim3(grid), dim3(kFwdThreads), smem, st, (const float*)h->act[l - 1], wf,
                  (const float*)h->params, nd, l, g, (const uint32_t*)h->d_masks, h->P, B, ntile, w_slots, tmem_cols, h->act[l], h->rmask[l]);
    }
    PN_LAUNCH_CHECK(h, "conv_mid_fwd_tc_kernel");
    return PN_OK;
}

static int launch_dgrad_col2im(pn_handle* h, int l, int B, cudaStream_t st) {
    const PnNetDesc& nd = h->nd;
    MidTcGeom g;
    if (!make_mid_tc_geom(nd, l, g)) return pn_fail(h, PN_ERR_STATE, "tensor-core middle conv layer not supported");
    const uint8_t* wd = h->tc_mid_buf[l] + (size_t)nd.M * 2 * kModBytes;     // written by this step's forward launch
    const int w_slots = std::max(2, (h->max_active[l] + 1) & ~1);
    const int pitch = g.s1 ? 28 : 36;
    const size_t smem = 128 + (size_t)2 * w_slots * kModBytes + (size_t)kXStages * 2 * w_slots * (kTileM * 16) +
                        (size_t)kTileM * pitch * sizeof(float);
    PN_SET_SMEM(h, conv_mid_dgrad_col2im_tc_kernel, smem);
    const int tstride = kTileM - g.max_shift;                                // output rows per tile (see the kernel)
    const int ntile = (B * g.RB + tstride - 1) / tstride;
    const long long NT = (long long)h->P * ntile;
    const int grid = (int)std::min<long long>(NT, 2ll * h->num_sms);
    // layer 1 writes the first conv layer's output gradient: fold the abs-max its quantiser needs into this pass
    unsigned* amax = l == 1 ? pn_conv0_absmax_slot(h) : nullptr;
    if (amax) PN_CUDA(h, cudaMemsetAsync(amax, 0, (size_t)h->P * sizeof(unsigned), st));
    {
        PnProfScope prof(h, PN_PROF_CONV_DGRAD + l, st);
        conv_mid_dgrad_col2im_tc_kernel<<<grid, kFwdThreads, smem, st>>>(h->dact[l], h->rmask[l], wd, nd, l, g, h->d_masks, h->P, B,
                                                                  ntile, w_slots, h->dact[l - 1], amax);
    }
    h->absmax_fused = amax != nullptr;
    PN_LAUNCH_CHECK(h, "conv_mid_dgrad_col2im_tc_kernel");
    return PN_OK;
}

int pn_launch_conv_mid_dgrad_tc(pn_handle* h, int l, int B, cudaStream_t st) {
    const PnNetDesc& nd = h->nd;
    MidTcGeom g;
    if (!make_mid_tc_geom(nd, l, g)) return pn_fail(h, PN_ERR_STATE, "tensor-core middle conv layer not supported");
    if (g.s1) return launch_dgrad_col2im(h, l, B, st);
    const uint8_t* wd = h->tc_mid_buf[l] + (size_t)nd.M * 2 * kModBytes;     // written by this step's forward launch
    const int w_slots = std::max(2, (h->max_active[l] + 1) & ~1);
    const size_t smem = 128 + (size_t)2 * w_slots * kModBytes + (size_t)kXStages * 2 * w_slots * g.chunk_bytes;
    PN_SET_SMEM(h, conv_mid_dgrad_tc_kernel, smem);
    const int ntile = (B * g.RB + kTileM - 1) / kTileM;
    const long long NT = (long long)h->P * ntile;
    const int grid = (int)std::min<long long>(NT, 3ll * h->num_sms);
    // layer 1 writes the first conv layer's output gradient: fold the abs-max its quantiser needs into this pass
    unsigned* amax = l == 1 ? pn_conv0_absmax_slot(h) : nullptr;
    if (amax) PN_CUDA(h, cudaMemsetAsync(amax, 0, (size_t)h->P * sizeof(unsigned), st));
    {
        PnProfScope prof(h, PN_PROF_CONV_DGRAD + l, st);
        conv_mid_dgrad_tc_kernel<<<grid, kFwdThreads, smem, st>>>(h->dact[l], h->rmask[l], wd, nd, l, g, h->d_masks, h->P, B,
                                                                  ntile, w_slots, h->dact[l - 1], amax);
    }
    h->absmax_fused = amax != nullptr;
    PN_LAUNCH_CHECK(h, "conv_mid_dgrad_tc_kernel");
    return PN_OK;
}

// data gradient on the tensor cores: both geometries
bool pn_conv_mid_tc_has_bwd(const pn_handle* h, int l) {
    MidTcGeom g;
    return h->tc_mid[l] && make_mid_tc_geom(h->nd, l, g);
}
bool pn_conv_mid_wgrad_tc_usable(const pn_handle* h, int l) {
    MidTcGeom g;
    return h->tc_mid[l] && make_mid_tc_geom(h->nd, l, g);
}

int pn_launch_conv_mid_wgrad_tc(pn_handle* h, int l, int B, cudaStream_t st, int* nch_used) {
    const PnNetDesc& nd = h->nd;
    MidTcGeom g;
    if (!make_mid_tc_geom(nd, l, g)) return pn_fail(h, PN_ERR_STATE, "tensor-core middle conv layer not supported");
    const int ntile = (B * g.RB + kTileM - 1) / kTileM;
    const int nch = std::max(1, std::min(std::min(h->nch, ntile), 3 * h->num_sms / std::max(1, h->P)));
    // the flush reuses the stages as scratch: fp32 [64][129] accumulator copy + [128][33] bias partials
    const size_t smem = 128 + std::max<size_t>((size_t)kWgStages * (g.stage_bytes + 8 * kDChunk), (64 * 129 + 128 * 33) * sizeof(float));
    PN_SET_SMEM(h, conv_mid_wgrad_tc_kernel, smem);
    {
        PnProfScope prof(h, PN_PROF_CONV_WGRAD + l, st);
        const int npass = std::max(1, (h->max_active[l] + 3) / 4);
        conv_mid_wgrad_tc_kernel<<<h->P * nch * npass, kFwdThreads, smem, st>>>(h->act[l - 1], h->dact[l], h->rmask[l], nd, l, g,
                                                                               h->d_masks, B, ntile, nch, h->nch, npass,
                                                                               h->wpart[l]);
    }
    PN_LAUNCH_CHECK(h, "conv_mid_wgrad_tc_kernel");
    *nch_used = nch;
    return PN_OK;
}
