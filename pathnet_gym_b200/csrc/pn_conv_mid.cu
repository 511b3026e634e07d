// Module convolutions of the middle layers (layers 1 .. L-2: 8 input channels, 8 output channels per module)
// as register-tiled fp32 CUDA-core kernels.
//
// The k = 4 / stride 2 layers normally run on the tensor cores (pn_conv_mid_tc.cu).  These kernels cover what that
// path does not take -- other kernel/stride geometries (3x3/s1 in the L = 4 configuration), weight gradients of paths
// with more than 4 active modules -- and are the comparison arm of the tests (cfg.flags bit 3).
//
// Reference semantics: relu(conv2d(net, W_conv[l,j], stride, VALID) + b_conv[l,j]) * gate_j summed over the
// active modules in ascending j (game_ac_network.py:186-193); gradients = tf.gradients
// (a3c_training_thread.py:55-60).
//
// Shared-memory input layout (per transition): the 8-channel pixels are split by stride phase so that
// neighbouring output pixels read neighbouring float4s:  plane = ((iy % s) * s + ix % s) * 2 + half,
// position (iy / s) * Xp + ix / s, one float4 = channels half*4 .. half*4+3.  All weight reads are warp
// broadcasts.  Every thread works on two transitions at once so a weight float4 feeds 8 FMAs x 2.
#include "pn_internal.cuh"

#include <algorithm>

namespace {

constexpr int kG = 4;                 // transitions per CTA (forward / dgrad) and per round (wgrad)
constexpr int kPass = 4;              // modules per pass (accumulators in registers)

struct MidGeom {
    int IH, IW, k, s, OH, OW, KK, npix_in, npix;
    int Yp, Xp, PS, in_f4;            // phase-plane grid, float4 per plane, float4 per staged transition
};

MidGeom make_mid_geom(const PnNetDesc& nd, int l) {
    MidGeom g;
    g.IH = nd.IH[l]; g.IW = nd.IW[l]; g.k = nd.k[l]; g.s = nd.s[l]; g.OH = nd.OH[l]; g.OW = nd.OW[l]; g.KK = nd.KK[l];
    g.npix_in = g.IH * g.IW; g.npix = g.OH * g.OW;
    g.Yp = (g.IH + g.s - 1) / g.s; g.Xp = (g.IW + g.s - 1) / g.s;
    g.PS = g.Yp * g.Xp;
    g.in_f4 = g.s * g.s * 2 * g.PS;
    return g;
}

__device__ __forceinline__ void stage_input(float4* __restrict__ dst, const float* __restrict__ src, const MidGeom& g) {
    for (int i = threadIdx.x; i < g.npix_in * 2; i += blockDim.x) {
        const int pix = i >> 1, half = i & 1;
        const int iy = pix / g.IW, ix = pix - iy * g.IW;
        const int plane = ((iy % g.s) * g.s + ix % g.s) * 2 + half;
        dst[plane * g.PS + (iy / g.s) * g.Xp + ix / g.s] = reinterpret_cast<const float4*>(src)[i];
    }
}

__device__ __forceinline__ float f4c(const float4& v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w)); }

// acc[0..7] += a * {w0, w1} with the packed fp32x2 FMA of sm_100 (FFMA2): 4 instructions instead of 8.
// These kernels are issue-bound (about half of the issued instructions were FFMA), so halving the FMA
// instruction count is worth more than anything else here.
#define FMA8(acc, a, w0, w1)                                                                         \
    do {                                                                                             \
        const float2 aa__ = make_float2(a, a);                                                       \
        float2* a2__ = reinterpret_cast<float2*>(&acc[0]);                                           \
        a2__[0] = __ffma2_rn(aa__, make_float2(w0.x, w0.y), a2__[0]);                                \
        a2__[1] = __ffma2_rn(aa__, make_float2(w0.z, w0.w), a2__[1]);                                \
        a2__[2] = __ffma2_rn(aa__, make_float2(w1.x, w1.y), a2__[2]);                                \
        a2__[3] = __ffma2_rn(aa__, make_float2(w1.z, w1.w), a2__[3]);                                \
    } while (0)

// ---------------------------------------------------------------------------------------------------
// forward: CTA = (path, 4 transitions); thread = (output pixel, pair of transitions)
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512)
conv_mid_fwd_kernel(const float* __restrict__ X, const float* __restrict__ params, PnNetDesc nd, int l, MidGeom g,
                    const uint32_t* __restrict__ masks, int B, float* __restrict__ Y, uint8_t* __restrict__ R) {
    extern __shared__ __align__(16) float4 sm4[];
    float4* Xs = sm4;                                           // [kG][in_f4]
    float4* Ws = Xs + (size_t)kG * g.in_f4;                     // [kPass][KK*2]   (8 floats per im2col row)
    float* bs = reinterpret_cast<float*>(Ws + (size_t)kPass * g.KK * 2);   // [kPass][8]
    const int p = blockIdx.x, b0 = blockIdx.y * kG;
    const uint32_t mask = masks[p * nd.L + l];
    const int ns = __popc(mask);
    for (int tg = 0; tg < kG; ++tg)
        if (b0 + tg < B) stage_input(Xs + (size_t)tg * g.in_f4, X + ((size_t)p * B + b0 + tg) * g.npix_in * 8, g);
    const int units = g.npix * (kG / 2);
    for (int u0 = 0; u0 < units; u0 += blockDim.x) {
        const int u = u0 + threadIdx.x;
        const bool live = u < units;
        const int pix = live ? u % g.npix : 0, th = live ? u / g.npix : 0;
        const int oy = pix / g.OW, ox = pix - oy * g.OW;
        const float4* xa = Xs + (size_t)th * g.in_f4;
        const float4* xb = Xs + (size_t)(th + kG / 2) * g.in_f4;
        float sum0[8], sum1[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) sum0[c] = sum1[c] = 0.f;
        for (int sb = 0; sb < ns; sb += kPass) {
            const int nsl = min(kPass, ns - sb);
            __syncthreads();
            for (int sl = 0; sl < nsl; ++sl) {
                const int j = __fns(mask, 0, sb + sl + 1);
                const float* Wg = params + nd.convW_base[l] + (long long)j * nd.conv_stride[l];
                for (int i = threadIdx.x; i < g.KK * 2; i += blockDim.x) Ws[sl * g.KK * 2 + i] = reinterpret_cast<const float4*>(Wg)[i];
                if (threadIdx.x < 8) bs[sl * 8 + threadIdx.x] = Wg[nd.convb_off[l] + threadIdx.x];
            }
            __syncthreads();
            __align__(8) float acc0[kPass][8], acc1[kPass][8];
#pragma unroll
            for (int sl = 0; sl < kPass; ++sl)
#pragma unroll
                for (int c = 0; c < 8; ++c) acc0[sl][c] = acc1[sl][c] = 0.f;
            if (live) {
                for (int kh = 0; kh < g.k; ++kh) {
                    const int py = kh % g.s, dy = kh / g.s;
                    for (int kw = 0; kw < g.k; ++kw) {
                        const int px = kw % g.s, dx = kw / g.s;
                        const int base = ((py * g.s + px) * 2) * g.PS + (oy + dy) * g.Xp + ox + dx;
                        const float4* wrow = Ws + (size_t)(kh * g.k + kw) * 16;      // 8 ci x 2 float4
#pragma unroll
                        for (int half = 0; half < 2; ++half) {
                            const float4 va = xa[base + half * g.PS], vb = xb[base + half * g.PS];
#pragma unroll
                            for (int ci = 0; ci < 4; ++ci) {
                                const float a = f4c(va, ci), b = f4c(vb, ci);
#pragma unroll
                                for (int sl = 0; sl < kPass; ++sl) {
                                    if (sl < nsl) {
                                        const float4 w0 = wrow[sl * g.KK * 2 + (half * 4 + ci) * 2];
                                        const float4 w1 = wrow[sl * g.KK * 2 + (half * 4 + ci) * 2 + 1];
                                        FMA8(acc0[sl], a, w0, w1);
                                        FMA8(acc1[sl], b, w0, w1);
                                    }
                                }
                            }
                        }
                    }
                }
#pragma unroll
                for (int sl = 0; sl < kPass; ++sl) {
                    if (sl < nsl) {
                        const int j = __fns(mask, 0, sb + sl + 1);
                        unsigned bits0 = 0, bits1 = 0;
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            const float bb = bs[sl * 8 + c];
                            const float p0 = acc0[sl][c] + bb, p1 = acc1[sl][c] + bb;
                            if (p0 > 0.f) { sum0[c] += p0; bits0 |= 1u << c; }
                            if (p1 > 0.f) { sum1[c] += p1; bits1 |= 1u << c; }
                        }
                        const int ba = b0 + th, bbk = b0 + th + kG / 2;
                        if (ba < B) R[(((size_t)p * B + ba) * nd.M + j) * g.npix + pix] = (uint8_t)bits0;
                        if (bbk < B) R[(((size_t)p * B + bbk) * nd.M + j) * g.npix + pix] = (uint8_t)bits1;
                    }
                }
            }
        }
        if (live) {
            const int ba = b0 + th, bbk = b0 + th + kG / 2;
            if (ba < B) {
                float4* yo = reinterpret_cast<float4*>(Y + (((size_t)p * B + ba) * g.npix + pix) * 8);
                yo[0] = make_float4(sum0[0], sum0[1], sum0[2], sum0[3]);
                yo[1] = make_float4(sum0[4], sum0[5], sum0[6], sum0[7]);
            }
            if (bbk < B) {
                float4* yo = reinterpret_cast<float4*>(Y + (((size_t)p * B + bbk) * g.npix + pix) * 8);
                yo[0] = make_float4(sum1[0], sum1[1], sum1[2], sum1[3]);
                yo[1] = make_float4(sum1[4], sum1[5], sum1[6], sum1[7]);
            }
        }
    }
}

// masked output gradient of one pass of modules -> shared memory [tg][slot][half][pix] float4
__device__ __forceinline__ void stage_dpre(float4* __restrict__ Ds, const float* __restrict__ dY, const uint8_t* __restrict__ R,
                                           const PnNetDesc& nd, const MidGeom& g, uint32_t mask, int sb, int nsl, size_t t0,
                                           int ntg) {
    const int per = nsl * g.npix * 2;
    for (int i = threadIdx.x; i < ntg * per; i += blockDim.x) {
        const int tg = i / per, r = i - tg * per;
        const int sl = r / (g.npix * 2), q = r - sl * g.npix * 2;
        const int pix = q >> 1, half = q & 1;
        const size_t t = t0 + tg;
        const int j = __fns(mask, 0, sb + sl + 1);
        const unsigned bits = R[(t * nd.M + j) * g.npix + pix] >> (half * 4);
        float4 v = reinterpret_cast<const float4*>(dY + (t * g.npix + pix) * 8)[half];
        v.x = (bits & 1u) ? v.x : 0.f; v.y = (bits & 2u) ? v.y : 0.f; v.z = (bits & 4u) ? v.z : 0.f; v.w = (bits & 8u) ? v.w : 0.f;
        Ds[((size_t)(tg * kPass + sl) * 2 + half) * g.npix + pix] = v;
    }
}

// ---------------------------------------------------------------------------------------------------
// data gradient: CTA = (path, 4 transitions); thread = (input pixel in phase-plane order, pair of transitions)
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512, 2)
conv_mid_dgrad_kernel(const float* __restrict__ dY, const uint8_t* __restrict__ R, const float* __restrict__ params,
                      PnNetDesc nd, int l, MidGeom g, const uint32_t* __restrict__ masks, int B, float* __restrict__ dX,
                      unsigned* __restrict__ absmax_out) {
    extern __shared__ __align__(16) float4 sm4[];
    float4* Ds = sm4;                                            // [kG][kPass][2][npix]
    float4* Ws = Ds + (size_t)kG * kPass * 2 * g.npix;           // [kPass][KK*2]
    float amax = 0.f;                                            // max |dX| written by this thread (first conv layer's quantiser scale)
    const int p = blockIdx.x, b0 = blockIdx.y * kG;
    const int ntg = min(kG, B - b0);
    const uint32_t mask = masks[p * nd.L + l];
    const int ns = __popc(mask);
    const int nq = g.s * g.s * g.PS;                             // input positions in phase-plane order
    const int units = nq * (kG / 2);
    for (int u0 = 0; u0 < units; u0 += blockDim.x) {
        const int u = u0 + threadIdx.x;
        const int q = u % nq, th = u / nq;
        const int plane = q / g.PS, r = q - plane * g.PS;
        const int py = plane / g.s, px = plane - py * g.s, Yb = r / g.Xp, Xb = r - Yb * g.Xp;
        const int iy = Yb * g.s + py, ix = Xb * g.s + px;
        const bool live = u < units && iy < g.IH && ix < g.IW;
        float2 acc0[8], acc1[8];                                 // (even-co, odd-co) partial sums per input channel
#pragma unroll
        for (int c = 0; c < 8; ++c) acc0[c] = acc1[c] = make_float2(0.f, 0.f);
        for (int sb = 0; sb < ns; sb += kPass) {
            const int nsl = min(kPass, ns - sb);
            __syncthreads();
            for (int sl = 0; sl < nsl; ++sl) {
                const int j = __fns(mask, 0, sb + sl + 1);
                const float* Wg = params + nd.convW_base[l] + (long long)j * nd.conv_stride[l];
                for (int i = threadIdx.x; i < g.KK * 2; i += blockDim.x) Ws[sl * g.KK * 2 + i] = reinterpret_cast<const float4*>(Wg)[i];
            }
            stage_dpre(Ds, dY, R, nd, g, mask, sb, nsl, (size_t)p * B + b0, ntg);
            __syncthreads();
            if (live) {
                for (int dy = 0; dy * g.s + py < g.k; ++dy) {
                    const int oy = Yb - dy;
                    if (oy < 0 || oy >= g.OH) continue;
                    const int kh = dy * g.s + py;
                    for (int dx = 0; dx * g.s + px < g.k; ++dx) {
                        const int ox = Xb - dx;
                        if (ox < 0 || ox >= g.OW) continue;
                        const int kw = dx * g.s + px;
                        const int opix = oy * g.OW + ox;
                        for (int sl = 0; sl < nsl; ++sl) {
                            const float4* da = Ds + ((size_t)(th * kPass + sl) * 2) * g.npix + opix;
                            const float4* db = Ds + ((size_t)((th + kG / 2) * kPass + sl) * 2) * g.npix + opix;
                            const float4 a0 = da[0], a1 = da[g.npix], b0v = db[0], b1v = db[g.npix];
                            const float4* w = Ws + (size_t)sl * g.KK * 2 + (size_t)(kh * g.k + kw) * 16;
#pragma unroll
                            for (int ci = 0; ci < 8; ++ci) {
                                // partial sums over even / odd output channels in the two halves of a float2
                                const float4 w0 = w[ci * 2], w1 = w[ci * 2 + 1];
                                const float2 wa = make_float2(w0.x, w0.y), wb = make_float2(w0.z, w0.w);
                                const float2 wc = make_float2(w1.x, w1.y), wd = make_float2(w1.z, w1.w);
                                float2 s0 = acc0[ci], s1 = acc1[ci];
                                s0 = __ffma2_rn(make_float2(a0.x, a0.y), wa, s0); s0 = __ffma2_rn(make_float2(a0.z, a0.w), wb, s0);
                                s0 = __ffma2_rn(make_float2(a1.x, a1.y), wc, s0); s0 = __ffma2_rn(make_float2(a1.z, a1.w), wd, s0);
                                s1 = __ffma2_rn(make_float2(b0v.x, b0v.y), wa, s1); s1 = __ffma2_rn(make_float2(b0v.z, b0v.w), wb, s1);
                                s1 = __ffma2_rn(make_float2(b1v.x, b1v.y), wc, s1); s1 = __ffma2_rn(make_float2(b1v.z, b1v.w), wd, s1);
                                acc0[ci] = s0; acc1[ci] = s1;
                            }
                        }
                    }
                }
            }
        }
        if (live) {
            const int ba = b0 + th, bb = b0 + th + kG / 2;
            const int ipix = iy * g.IW + ix;
            if (ba < B) {
                float4* o = reinterpret_cast<float4*>(dX + (((size_t)p * B + ba) * g.npix_in + ipix) * 8);
                o[0] = make_float4(acc0[0].x + acc0[0].y, acc0[1].x + acc0[1].y, acc0[2].x + acc0[2].y, acc0[3].x + acc0[3].y);
                o[1] = make_float4(acc0[4].x + acc0[4].y, acc0[5].x + acc0[5].y, acc0[6].x + acc0[6].y, acc0[7].x + acc0[7].y);
#pragma unroll
                for (int c = 0; c < 8; ++c) amax = fmaxf(amax, fabsf(acc0[c].x + acc0[c].y));
            }
            if (bb < B) {
                float4* o = reinterpret_cast<float4*>(dX + (((size_t)p * B + bb) * g.npix_in + ipix) * 8);
                o[0] = make_float4(acc1[0].x + acc1[0].y, acc1[1].x + acc1[1].y, acc1[2].x + acc1[2].y, acc1[3].x + acc1[3].y);
                o[1] = make_float4(acc1[4].x + acc1[4].y, acc1[5].x + acc1[5].y, acc1[6].x + acc1[6].y, acc1[7].x + acc1[7].y);
#pragma unroll
                for (int c = 0; c < 8; ++c) amax = fmaxf(amax, fabsf(acc1[c].x + acc1[c].y));
            }
        }
    }
    if (absmax_out) {                    // warp -> block -> one global atomic per CTA (non-negative floats order like uints)
        __shared__ unsigned blk_amax;
        if (threadIdx.x == 0) blk_amax = 0u;
        __syncthreads();
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        if ((threadIdx.x & 31) == 0) atomicMax(&blk_amax, __float_as_uint(amax));
        __syncthreads();
        if (threadIdx.x == 0 && blk_amax) atomicMax(absmax_out + p, blk_amax);          // one scale per path
    }
}

// ---------------------------------------------------------------------------------------------------
// weight-gradient partials: CTA = (path, chunk of transitions); thread = (pair of im2col rows kk, transition
// group r of 4); rounds of 4 transitions; fixed-order reduction over the 4 groups at the end of a pass.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512)
conv_mid_wgrad_kernel(const float* __restrict__ X, const float* __restrict__ dY, const uint8_t* __restrict__ R,
                      PnNetDesc nd, int l, MidGeom g, const uint32_t* __restrict__ masks, int B, int cs, int nch_alloc,
                      float* __restrict__ part) {
    extern __shared__ __align__(16) float4 sm4[];
    float4* Xs = sm4;                                            // [kG][in_f4]
    float4* Ds = Xs + (size_t)kG * g.in_f4;                      // [kG][kPass][2][npix]
    float* red = reinterpret_cast<float*>(sm4);                  // reduction scratch reuses both (after a barrier)
    const int p = blockIdx.x, ch = blockIdx.y;
    const int b_lo = ch * cs, b_hi = min(B, b_lo + cs);
    const int nq = g.KK / 2;
    const int q = threadIdx.x % nq, r = threadIdx.x / nq;        // r in [0, kG)
    const int kk0 = 2 * q;
    const int tap = kk0 >> 3, ci = kk0 & 7, half = ci >> 2, cc = ci & 3;    // cc in {0, 2}
    const int kh = tap / g.k, kw = tap - kh * g.k;
    const int py = kh % g.s, dy = kh / g.s, px = kw % g.s, dx = kw / g.s;
    const int xbase = ((py * g.s + px) * 2 + half) * g.PS + dy * g.Xp + dx;
    const uint32_t mask = masks[p * nd.L + l];
    const int ns = __popc(mask);
    const int stride_part = g.KK * 8 + 8;
    for (int sb = 0; sb < ns; sb += kPass) {
        const int nsl = min(kPass, ns - sb);
        __align__(8) float acc[2][kPass][8];
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int sl = 0; sl < kPass; ++sl)
#pragma unroll
                for (int c = 0; c < 8; ++c) acc[a][sl][c] = 0.f;
        float bacc = 0.f;
        const bool bias_thread = q < nsl * 8;
        for (int bb = b_lo; bb < b_hi; bb += kG) {
            const int ntg = min(kG, b_hi - bb);
            __syncthreads();
            for (int tg = 0; tg < ntg; ++tg) stage_input(Xs + (size_t)tg * g.in_f4, X + ((size_t)p * B + bb + tg) * g.npix_in * 8, g);
            stage_dpre(Ds, dY, R, nd, g, mask, sb, nsl, (size_t)p * B + bb, ntg);
            __syncthreads();
            if (r < ntg) {
                const float* xs = reinterpret_cast<const float*>(Xs + (size_t)r * g.in_f4 + xbase) + cc;
                const float4* ds = Ds + (size_t)r * kPass * 2 * g.npix;
                for (int oy = 0; oy < g.OH; ++oy) {
                    for (int ox = 0; ox < g.OW; ++ox) {
                        const float2 xv = *reinterpret_cast<const float2*>(xs + (size_t)(oy * g.Xp + ox) * 4);
                        const int pix = oy * g.OW + ox;
#pragma unroll
                        for (int sl = 0; sl < kPass; ++sl) {
                            if (sl < nsl) {
                                const float4 d0 = ds[(size_t)(sl * 2) * g.npix + pix], d1 = ds[(size_t)(sl * 2 + 1) * g.npix + pix];
                                FMA8(acc[0][sl], xv.x, d0, d1);
                                FMA8(acc[1][sl], xv.y, d0, d1);
                            }
                        }
                    }
                }
                if (bias_thread) {
                    const int sl = q >> 3, c = q & 7;
                    const float* dcol = reinterpret_cast<const float*>(ds + (size_t)(sl * 2 + (c >> 2)) * g.npix) + (c & 3);
                    for (int pix = 0; pix < g.npix; ++pix) bacc += dcol[(size_t)pix * 4];
                }
            }
        }
        // fixed-order sum over the 4 transition groups
        __syncthreads();
        float* mine = red + ((size_t)r * nq + q) * (2 * kPass * 8 + 1);
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int sl = 0; sl < kPass; ++sl)
#pragma unroll
                for (int c = 0; c < 8; ++c) mine[(a * kPass + sl) * 8 + c] = acc[a][sl][c];
        mine[2 * kPass * 8] = bacc;
        __syncthreads();
        if (r == 0) {
            for (int sl = 0; sl < nsl; ++sl) {
                const int j = __fns(mask, 0, sb + sl + 1);
                float* out = part + ((size_t)(p * nch_alloc + ch) * nd.M + j) * stride_part;
#pragma unroll
                for (int a = 0; a < 2; ++a) {
                    float o[8];
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        float s = 0.f;
                        for (int rr = 0; rr < kG; ++rr) s += red[((size_t)rr * nq + q) * (2 * kPass * 8 + 1) + (a * kPass + sl) * 8 + c];
                        o[c] = s;
                    }
                    float4* dst = reinterpret_cast<float4*>(out + (size_t)(kk0 + a) * 8);
                    dst[0] = make_float4(o[0], o[1], o[2], o[3]);
                    dst[1] = make_float4(o[4], o[5], o[6], o[7]);
                }
            }
            if (bias_thread) {
                float s = 0.f;
                for (int rr = 0; rr < kG; ++rr) s += red[((size_t)rr * nq + q) * (2 * kPass * 8 + 1) + 2 * kPass * 8];
                const int j = __fns(mask, 0, sb + (q >> 3) + 1);
                part[((size_t)(p * nch_alloc + ch) * nd.M + j) * stride_part + (size_t)g.KK * 8 + (q & 7)] = s;
            }
        }
    }
}

template <typename K>
int set_smem_mid(pn_handle* h, K kern, size_t bytes) {
    PN_SET_SMEM(h, kern, bytes);
    return PN_OK;
}

}  // namespace

bool pn_conv_mid_supported(const pn_handle* h, int l) {
    const PnNetDesc& nd = h->nd;
    if (l < 1 || nd.Cin[l] != 8 || (nd.KK[l] & 1)) return false;
    const MidGeom g = make_mid_geom(nd, l);
    const size_t fwd = ((size_t)kG * g.in_f4 + (size_t)kPass * g.KK * 2) * 16 + kPass * 8 * 4;
    const size_t dgr = ((size_t)kG * kPass * 2 * g.npix + (size_t)kPass * g.KK * 2) * 16;
    const size_t wgr = std::max(((size_t)kG * g.in_f4 + (size_t)kG * kPass * 2 * g.npix) * 16,
                                (size_t)kG * (g.KK / 2) * (2 * kPass * 8 + 1) * 4);
    return std::max(fwd, std::max(dgr, wgr)) <= 200 * 1024 && kG * (g.KK / 2) <= 512 && g.KK / 2 >= kPass * 8;
}

int pn_launch_conv_mid_fwd(pn_handle* h, int l, int B, cudaStream_t st) {
    const PnNetDesc& nd = h->nd;
    const MidGeom g = make_mid_geom(nd, l);
    const size_t smem = ((size_t)kG * g.in_f4 + (size_t)kPass * g.KK * 2) * 16 + kPass * 8 * 4;
    if (int rc = set_smem_mid(h, conv_mid_fwd_kernel, smem)) return rc;
    const int threads = std::min(512, (g.npix * (kG / 2) + 31) / 32 * 32);
    dim3 grid(h->P, (B + kG - 1) / kG);
    {
        PnProfScope prof(h, PN_PROF_CONV_FWD + l, st);
        conv_mid_fwd_kernel<<<grid, threads, smem, st>>>(h->act[l - 1], h->params, nd, l, g, h->d_masks, B, h->act[l], h->rmask[l]);
    }
    PN_LAUNCH_CHECK(h, "conv_mid_fwd_kernel");
    return PN_OK;
}

int pn_launch_conv_mid_bwd(pn_handle* h, int l, int B, cudaStream_t st, int* nch_used, bool do_wgrad, bool do_dgrad) {
    const PnNetDesc& nd = h->nd;
    const MidGeom g = make_mid_geom(nd, l);
    const int P = h->P;
    // weight gradient: chunks of a multiple of 4 transitions, about two CTAs per SM
    int nch = std::max(1, std::min(h->nch, (2 * h->num_sms + P - 1) / P));
    int cs = ((B + nch - 1) / nch + kG - 1) / kG * kG;
    nch = (B + cs - 1) / cs;
    if (do_wgrad) {
        const size_t smem = std::max(((size_t)kG * g.in_f4 + (size_t)kG * kPass * 2 * g.npix) * 16,
                                     (size_t)kG * (g.KK / 2) * (2 * kPass * 8 + 1) * 4);
        if (int rc = set_smem_mid(h, conv_mid_wgrad_kernel, smem)) return rc;
        dim3 grid(P, nch);
        PnProfScope prof(h, PN_PROF_CONV_WGRAD + l, st);
        conv_mid_wgrad_kernel<<<grid, kG * (g.KK / 2), smem, st>>>(h->act[l - 1], h->dact[l], h->rmask[l], nd, l, g, h->d_masks,
                                                                  B, cs, h->nch, h->wpart[l]);
        PN_LAUNCH_CHECK(h, "conv_mid_wgrad_kernel");
    }
    *nch_used = nch;
    if (do_dgrad) {
        const size_t smem = ((size_t)kG * kPass * 2 * g.npix + (size_t)kPass * g.KK * 2) * 16;
        if (int rc = set_smem_mid(h, conv_mid_dgrad_kernel, smem)) return rc;
        const int units = g.s * g.s * g.PS * (kG / 2);
        const int threads = std::min(512, (units + 31) / 32 * 32);
        dim3 grid(P, (B + kG - 1) / kG);
        // layer 1 writes the first conv layer's output gradient: fold the abs-max its quantiser needs into this pass
        unsigned* amax = l == 1 ? pn_conv0_absmax_slot(h) : nullptr;
        if (amax) PN_CUDA(h, cudaMemsetAsync(amax, 0, (size_t)h->P * sizeof(unsigned), st));
        PnProfScope prof(h, PN_PROF_CONV_DGRAD + l, st);
        conv_mid_dgrad_kernel<<<grid, threads, smem, st>>>(h->dact[l], h->rmask[l], h->params, nd, l, g, h->d_masks, B,
                                                          h->dact[l - 1], amax);
        h->absmax_fused = amax != nullptr;
        PN_LAUNCH_CHECK(h, "conv_mid_dgrad_kernel");
    }
    return PN_OK;
}
