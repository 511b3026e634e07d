// Forward / loss / backward kernels of the module super-net (fp32 CUDA-core path).
//
// Semantics follow the reference graph (game_ac_network.py:182-201 forward,
// :38-59 loss, a3c_training_thread.py:55-60 gradients); only the modules a
// path's genotype enables are evaluated -- the reference multiplies the others
// by an exact 0.0 gate, so the sums are identical.  Module outputs are summed in
// ascending module index, like np.sum over the reference's object array.
//
// Layouts (t = p*B + b):
//   act[l]   [T][OH*OW][8]   fp32  sum over active modules of relu(conv+b)
//   rmask[l] [T][M][OH*OW]   u8    bit c = (pre-activation of channel c > 0)
//   hbuf     [T][hidden]     fp32  (sum of relu(linear))/M
//   lmask    [T][M][hidden]  u8    0/1
#include "pn_internal.cuh"
#include "pn_tc.cuh"

#include <algorithm>

namespace {

__device__ __forceinline__ float to_f(uint8_t v) { return (float)v; }
__device__ __forceinline__ float to_f(float v) { return v; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---------------------------------------------------------------------------
// conv layer forward: one CTA = one transition x one band of output rows
// ---------------------------------------------------------------------------
template <typename TIn>
__global__ void __launch_bounds__(128)
conv_fwd_kernel(const TIn* __restrict__ X, float in_scale, const float* __restrict__ params, PnNetDesc nd, int l,
                const uint32_t* __restrict__ masks, int B, int rows_per_band, float* __restrict__ Y,
                uint8_t* __restrict__ R) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int t = blockIdx.x, band = blockIdx.y, p = t / B;
    const int IH = nd.IH[l], IW = nd.IW[l], Cin = nd.Cin[l], k = nd.k[l], s = nd.s[l];
    const int OH = nd.OH[l], OW = nd.OW[l], KK = nd.KK[l];
    const int r0 = band * rows_per_band;
    const int nrows = min(rows_per_band, OH - r0);
    const int npix_band = nrows * OW, npix = OH * OW;
    const int in_rows = (nrows - 1) * s + k, row_elems = IW * Cin;
    float* Ws = reinterpret_cast<float*>(smem_raw);   // [KK*8]
    float* bs = Ws + KK * PN_F;                        // [8]
    TIn* Xs = reinterpret_cast<TIn*>(bs + PN_F);       // [in_rows*row_elems]
    const TIn* Xg = X + ((size_t)t * IH + (size_t)r0 * s) * row_elems;
    for (int i = threadIdx.x; i < in_rows * row_elems; i += blockDim.x) Xs[i] = Xg[i];
    const uint32_t mask = masks[p * nd.L + l];
    const int iters = (npix_band + blockDim.x - 1) / blockDim.x;
    const int kc = k * Cin;
    for (int it = 0; it < iters; ++it) {
        const int pix = it * blockDim.x + threadIdx.x;
        const bool live = pix < npix_band;
        const int oy = live ? pix / OW : 0, ox = live ? pix % OW : 0;
        float sum[PN_F];
#pragma unroll
        for (int c = 0; c < PN_F; ++c) sum[c] = 0.f;
        for (int j = 0; j < nd.M; ++j) {
            if (!((mask >> j) & 1u)) continue;   // block-uniform
            __syncthreads();
            const float* Wg = params + nd.convW_base[l] + (long long)j * nd.conv_stride[l];
            for (int i = threadIdx.x; i < KK * PN_F; i += blockDim.x) Ws[i] = Wg[i];
            if (threadIdx.x < PN_F) bs[threadIdx.x] = Wg[nd.convb_off[l] + threadIdx.x];
            __syncthreads();
            if (live) {
                float acc[PN_F];
#pragma unroll
                for (int c = 0; c < PN_F; ++c) acc[c] = 0.f;
                for (int kh = 0; kh < k; ++kh) {
                    const TIn* xr = Xs + (oy * s + kh) * row_elems + ox * s * Cin;
                    const float4* wr = reinterpret_cast<const float4*>(Ws + kh * kc * PN_F);
                    for (int q = 0; q < kc; ++q) {
                        const float a = to_f(xr[q]);
                        const float4 w0 = wr[2 * q], w1 = wr[2 * q + 1];
                        acc[0] = fmaf(a, w0.x, acc[0]); acc[1] = fmaf(a, w0.y, acc[1]);
                        acc[2] = fmaf(a, w0.z, acc[2]); acc[3] = fmaf(a, w0.w, acc[3]);
                        acc[4] = fmaf(a, w1.x, acc[4]); acc[5] = fmaf(a, w1.y, acc[5]);
                        acc[6] = fmaf(a, w1.z, acc[6]); acc[7] = fmaf(a, w1.w, acc[7]);
                    }
                }
                unsigned bits = 0;
#pragma unroll
                for (int c = 0; c < PN_F; ++c) {
                    const float pre = fmaf(acc[c], in_scale, bs[c]);
                    if (pre > 0.f) { sum[c] += pre; bits |= 1u << c; }
                }
                R[((size_t)t * nd.M + j) * npix + r0 * OW + pix] = (uint8_t)bits;
            }
        }
        if (live) {
            float4* yo = reinterpret_cast<float4*>(Y + ((size_t)t * npix + r0 * OW + pix) * PN_F);
            yo[0] = make_float4(sum[0], sum[1], sum[2], sum[3]);
            yo[1] = make_float4(sum[4], sum[5], sum[6], sum[7]);
        }
    }
}

// ---------------------------------------------------------------------------
// linear modules forward: CTA = (path, 64 hidden units, 32 transitions)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
lin_fwd_kernel(const float* __restrict__ X, const float* __restrict__ params, PnNetDesc nd,
               const uint32_t* __restrict__ masks, int B, float* __restrict__ Hout, uint8_t* __restrict__ Rl) {
    __shared__ float Xs[32][33];
    __shared__ __align__(16) float Ws[32][64];
    const int p = blockIdx.x, c0 = blockIdx.y * 64, b0 = blockIdx.z * 32;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int K = nd.last_lin, Hd = nd.hidden;
    const uint32_t mask = masks[p * nd.L + nd.L - 1];
    float sum[2][4];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) sum[a][b] = 0.f;
    for (int j = 0; j < nd.M; ++j) {
        if (!((mask >> j) & 1u)) continue;
        const float* W = params + nd.linW_base + (long long)j * nd.lin_stride;
        float acc[2][4];
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
        for (int k0 = 0; k0 < K; k0 += 32) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int idx = tid + i * 256, row = idx >> 5, kk = idx & 31, b = b0 + row;
                Xs[row][kk] = (b < B && k0 + kk < K) ? X[(size_t)(p * B + b) * K + k0 + kk] : 0.f;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int idx = tid + i * 256, kk = idx >> 6, c = idx & 63;
                Ws[kk][c] = (k0 + kk < K && c0 + c < Hd) ? W[(size_t)(k0 + kk) * Hd + c0 + c] : 0.f;
            }
            __syncthreads();
#pragma unroll 8
            for (int kk = 0; kk < 32; ++kk) {
                const float x0 = Xs[ty * 2][kk], x1 = Xs[ty * 2 + 1][kk];
                const float4 w = *reinterpret_cast<const float4*>(&Ws[kk][tx * 4]);
                acc[0][0] = fmaf(x0, w.x, acc[0][0]); acc[0][1] = fmaf(x0, w.y, acc[0][1]);
                acc[0][2] = fmaf(x0, w.z, acc[0][2]); acc[0][3] = fmaf(x0, w.w, acc[0][3]);
                acc[1][0] = fmaf(x1, w.x, acc[1][0]); acc[1][1] = fmaf(x1, w.y, acc[1][1]);
                acc[1][2] = fmaf(x1, w.z, acc[1][2]); acc[1][3] = fmaf(x1, w.w, acc[1][3]);
            }
            __syncthreads();
        }
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            const int b = b0 + ty * 2 + a;
            if (b >= B) continue;
            const size_t t = (size_t)p * B + b;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int col = c0 + tx * 4 + q;
                if (col >= Hd) continue;
                const float pre = acc[a][q] + W[nd.linb_off + col];
                const bool on = pre > 0.f;
                if (on) sum[a][q] += pre;
                Rl[(t * nd.M + j) * Hd + col] = on ? 1 : 0;
            }
        }
    }
    const float Mf = nd.out_div;
#pragma unroll
    for (int a = 0; a < 2; ++a) {
        const int b = b0 + ty * 2 + a;
        if (b >= B) continue;
        const size_t t = (size_t)p * B + b;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int col = c0 + tx * 4 + q;
            if (col < Hd) Hout[t * Hd + col] = sum[a][q] / Mf;   // net/FLAGS.M, game_ac_network.py:194
        }
    }
}

// ---------------------------------------------------------------------------
// heads + A3C loss + closed-form dL/dlogits, dL/dv, dL/dh : one warp per transition
// (game_ac_network.py:198-201 heads, :38-59 loss)
// ---------------------------------------------------------------------------
// One block = one chunk of PN_HEAD_CHUNK (16) transitions, 256 threads.  The chunk's hidden activations, logits
// and dL/dz stay in shared memory, so the block runs its small dense phases back to back:
//   1. z[tt][k]  = sum_i h[tt][i] Wc[i][k]     thread = (transition tt, 1/16 slice of i), 16-lane shuffle reduce
//   2. softmax / loss / dL/dz                  one warp per transition
//   3. dh[tt][i] = sum_k dz[tt][k] Wc[i][k]  and the chunk's partial head-weight gradient
//      d Wc[i][k] = sum_tt h[tt][i] dz[tt][k] (fixed order)       thread = hidden unit i
// Wc[i][0..A-1] = W_fc2[i][:], Wc[i][A] = W_fc3[i], zero up to NP (a multiple of 4): every inner loop is
// float4 shared loads feeding packed fp32x2 FMAs with no per-column predicate.
template <int NP>
__global__ void __launch_bounds__(256, 4)       // 64 registers: all 512 chunks of an 8192-transition step are resident at once
head_kernel(const float* __restrict__ Hin, const float* __restrict__ params, PnNetDesc nd, int T, int training,
            const int32_t* __restrict__ actions, const float* __restrict__ td, const float* __restrict__ r,
            float beta, float* __restrict__ pi_out, float* __restrict__ v_out, float* __restrict__ dzv,
            float* __restrict__ dh, float* __restrict__ loss_t, float* __restrict__ part) {
    pn_pdl_sync();
    static_assert(PN_HEAD_CHUNK == 16 && NP % 4 == 0 && NP <= 32, "phase 1 maps 16 transitions x 16 slices onto 256 threads");
    extern __shared__ __align__(16) float hsm[];
    const int Hd = nd.hidden, A = nd.A, A1 = A + 1;
    const int hpitch = ((Hd + 3) & ~3) + 16;   // +16 floats: the two transitions of a warp hit different banks
    float* Wc = hsm;                           // [Hd][NP]
    float* b2s = Wc + (size_t)Hd * NP;         // [32] (b3 at index A)
    float* hs = b2s + 32;                      // [16][hpitch]
    float* zs = hs + PN_HEAD_CHUNK * hpitch;   // [16][32] logits (k < A), value (k == A)
    float* dzs = zs + PN_HEAD_CHUNK * 32;      // [16][32]
    // A CTA walks over several 16-transition chunks (chunk = blockIdx.x, + gridDim.x, ...): the head weights are staged once
    // per CTA, and the CTA's partial of the head weight gradient is accumulated over its chunks (first chunk writes, later
    // chunks add, same CTA, fixed order), so the reduction that follows reads one partial per CTA instead of one per chunk.
    const int nchunks_all = (T + PN_HEAD_CHUNK - 1) / PN_HEAD_CHUNK;
    {
        for (int i = threadIdx.x; i < Hd; i += blockDim.x) {       // one weight row per thread, loads issued together
            float wr[NP];
            const float* src = params + nd.fc2W + (size_t)i * A;
            if ((A & 1) == 0) {
#pragma unroll
                for (int k = 0; k < NP; k += 2) {
                    const float2 v = k < A ? *reinterpret_cast<const float2*>(src + k) : make_float2(0.f, 0.f);
                    wr[k] = v.x; wr[k + 1] = v.y;
                }
            } else {
#pragma unroll
                for (int k = 0; k < NP; ++k) wr[k] = k < A ? src[k] : 0.f;
            }
            const float w3 = params[nd.fc3W + i];
#pragma unroll
            for (int k = 0; k < NP; ++k) wr[k] = k == A ? w3 : wr[k];
#pragma unroll
            for (int q = 0; q < NP / 4; ++q)
                reinterpret_cast<float4*>(Wc + (size_t)i * NP)[q] = make_float4(wr[4 * q], wr[4 * q + 1], wr[4 * q + 2], wr[4 * q + 3]);
        }
        if (threadIdx.x < 32) b2s[threadIdx.x] = threadIdx.x < A ? params[nd.fc2b + threadIdx.x]
                                                                  : (threadIdx.x == A ? params[nd.fc3b] : 0.f);
    }
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    bool first = true;
    for (int chunk = blockIdx.x; chunk < nchunks_all; chunk += gridDim.x, first = false) {
    const int t0 = chunk * PN_HEAD_CHUNK, nt = min(PN_HEAD_CHUNK, T - t0);
    __syncthreads();                                               // the previous chunk's readers are done with hs / zs / dzs
    {
        const int m4 = Hd >> 2;
        if ((Hd & 3) == 0) {
            for (int i = threadIdx.x; i < nt * m4; i += blockDim.x) {
                const int tt = i / m4, c = i - tt * m4;
                pntc::cp_async16(hs + tt * hpitch + 4 * c, Hin + (size_t)(t0 + tt) * Hd + 4 * c);
            }
        } else {
            for (int i = threadIdx.x; i < nt * Hd; i += blockDim.x) {
                const int tt = i / Hd, c = i - tt * Hd;
                hs[tt * hpitch + c] = Hin[(size_t)(t0 + tt) * Hd + c];
            }
        }
        for (int i = threadIdx.x; i < PN_HEAD_CHUNK * 32; i += blockDim.x) dzs[i] = 0.f;
        pntc::cp_async_wait_all();
    }
    __syncthreads();
    {   // phase 1
        const int tt = threadIdx.x >> 4, ks = threadIdx.x & 15;
        float2 acc[NP / 2];
#pragma unroll
        for (int k = 0; k < NP / 2; ++k) acc[k] = make_float2(0.f, 0.f);
        if (tt < nt) {
            const float* hrow = hs + tt * hpitch;
            for (int i = ks; i < Hd; i += 16) {
                const float hv = hrow[i];
                const float2 h2 = make_float2(hv, hv);
                const float4* wrow = reinterpret_cast<const float4*>(Wc + (size_t)i * NP);
#pragma unroll
                for (int q = 0; q < NP / 4; ++q) {
                    const float4 wv = wrow[q];
                    acc[2 * q] = __ffma2_rn(h2, make_float2(wv.x, wv.y), acc[2 * q]);
                    acc[2 * q + 1] = __ffma2_rn(h2, make_float2(wv.z, wv.w), acc[2 * q + 1]);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < NP / 2; ++k) {
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                acc[k].x += __shfl_xor_sync(0xffffffffu, acc[k].x, o);
                acc[k].y += __shfl_xor_sync(0xffffffffu, acc[k].y, o);
            }
        }
        if (ks == 0) {
#pragma unroll
            for (int q = 0; q < NP / 4; ++q)
                reinterpret_cast<float4*>(zs + tt * 32)[q] = make_float4(acc[2 * q].x, acc[2 * q].y, acc[2 * q + 1].x, acc[2 * q + 1].y);
        }
    }
    __syncthreads();
    for (int tt = w; tt < nt; tt += 8) {       // phase 2
        const int t = t0 + tt;
        const float acc = lane < A1 ? zs[tt * 32 + lane] + b2s[lane] : 0.f;
        const float v = __shfl_sync(0xffffffffu, acc, A);
        const float zk = lane < A ? acc : -INFINITY;
        const float m = warp_max(zk);
        const float e = lane < A ? expf(zk - m) : 0.f;
        const float ssum = warp_sum(e);
        const float pik = e / ssum;
        if (pi_out && lane < A) pi_out[(size_t)t * A + lane] = pik;
        if (v_out && lane == 0) v_out[t] = v;
        if (training) {
            const int a = actions[t];
            const float tdv = td[t], rv = r[t];
            const float clipped = fminf(fmaxf(pik, 1e-20f), 1.0f);
            const float logp = lane < A ? logf(clipped) : 0.f;
            const bool inside = (pik >= 1e-20f) && (pik <= 1.0f);
            const float Hent = -warp_sum(lane < A ? pik * logp : 0.f);
            const bool a_ok = a >= 0 && a < A;
            const float logp_a = __shfl_sync(0xffffffffu, logp, a_ok ? a : 0);
            const float diff = rv - v;
            const float loss = -((a_ok ? logp_a * tdv : 0.f) + Hent * beta) + 0.25f * diff * diff;
            float dpi = 0.f;
            if (lane < A) {
                dpi = beta * (logp + (inside ? 1.f : 0.f));
                if (a_ok && lane == a && inside) dpi -= tdv / clipped;
            }
            const float dot = warp_sum(dpi * pik);
            const float dz = pik * (dpi - dot);
            const float dv = -0.5f * diff;
            // action < 0 marks a padded transition (a path whose rollout ended before B steps): no loss, no gradient
            const float val = a < 0 ? 0.f : (lane < A ? dz : (lane == A ? dv : 0.f));
            dzv[(size_t)t * 32 + lane] = val;
            dzs[tt * 32 + lane] = val;
            if (lane == 0) loss_t[t] = a < 0 ? 0.f : loss;
        }
    }
    if (!training) continue;
    __syncthreads();
    const float rM = 1.0f / nd.out_div;
    float* out = part ? part + (size_t)blockIdx.x * ((size_t)Hd * A1 + A1) : nullptr;
    for (int i = threadIdx.x; i < Hd; i += blockDim.x) {       // phase 3
        float2 wr[NP / 2], acc[NP / 2];
#pragma unroll
        for (int q = 0; q < NP / 4; ++q) {
            const float4 wv = reinterpret_cast<const float4*>(Wc + (size_t)i * NP)[q];
            wr[2 * q] = make_float2(wv.x, wv.y); wr[2 * q + 1] = make_float2(wv.z, wv.w);
            acc[2 * q] = acc[2 * q + 1] = make_float2(0.f, 0.f);
        }
#pragma unroll 4
        for (int tt = 0; tt < nt; ++tt) {
            const float hv = hs[tt * hpitch + i];
            const float2 h2 = make_float2(hv, hv);
            const float4* dz4 = reinterpret_cast<const float4*>(dzs + tt * 32);
            float2 g = make_float2(0.f, 0.f);
#pragma unroll
            for (int q = 0; q < NP / 4; ++q) {
                const float4 d = dz4[q];
                const float2 d0 = make_float2(d.x, d.y), d1 = make_float2(d.z, d.w);
                g = __ffma2_rn(d0, wr[2 * q], g);
                g = __ffma2_rn(d1, wr[2 * q + 1], g);
                acc[2 * q] = __ffma2_rn(h2, d0, acc[2 * q]);
                acc[2 * q + 1] = __ffma2_rn(h2, d1, acc[2 * q + 1]);
            }
            dh[(size_t)(t0 + tt) * Hd + i] = (g.x + g.y) * rM;
        }
        if (out) {
#pragma unroll
            // column-major partial (element k * Hd + i): consecutive threads write consecutive floats
            for (int k = 0; k < NP / 2; ++k) {
                if (2 * k < A1) out[(size_t)(2 * k) * Hd + i] = (first ? 0.f : out[(size_t)(2 * k) * Hd + i]) + acc[k].x;
                if (2 * k + 1 < A1) out[(size_t)(2 * k + 1) * Hd + i] = (first ? 0.f : out[(size_t)(2 * k + 1) * Hd + i]) + acc[k].y;
            }
        }
    }
    if (out && threadIdx.x < A1) {
        float s = 0.f;
        for (int tt = 0; tt < nt; ++tt) s += dzs[tt * 32 + threadIdx.x];
        out[(size_t)Hd * A1 + threadIdx.x] = (first ? 0.f : out[(size_t)Hd * A1 + threadIdx.x]) + s;
    }
    }   // chunks of this CTA
}

// stage 2: fixed-order sum over chunks, scatter into W_fc2/b_fc2/W_fc3/b_fc3.
// block = 32 elements x 8 chunk groups; group g adds chunks g, g+8, ... in order, then the 8 group sums are
// added in group order (deterministic, 8x more loads in flight than one thread per element)
__device__ __forceinline__ void head_reduce_body(int bid, const float* __restrict__ part, int nchunks, const PnNetDesc& nd,
                                                 float* __restrict__ grads, int nred_blocks, const float* __restrict__ loss_t,
                                                 int P, int B, float* __restrict__ loss, float (*red)[33]) {
    const int Hd = nd.hidden, A = nd.A, A1 = A + 1;
    const int n = Hd * A1 + A1;
    if (bid >= nred_blocks) {                        // trailing blocks: per-path loss = sum over its B transitions
        const int p = (bid - nred_blocks) * blockDim.x + threadIdx.x;
        if (p < P) {
            float s = 0.f;
            for (int b = 0; b < B; ++b) s += loss_t[(size_t)p * B + b];
            loss[p] = s;
        }
        return;
    }
    const int el = threadIdx.x & 31, grp = threadIdx.x >> 5;
    const int e = bid * 32 + el;
    float s = 0.f;
    if (e < n) {
        // sixteen independent partial sums (chunks c, c+8, ..., c+120 of this group's sequence): sixteen loads in flight
        // instead of a chain of dependent round trips (512 partials per element at 8192 transitions: the chain WAS this
        // launch's duration); combined in a fixed order
        float s16[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) s16[u] = 0.f;
        int c = grp;
        for (; c + 120 < nchunks; c += 128) {
#pragma unroll
            for (int u = 0; u < 16; ++u) s16[u] += part[(size_t)(c + 8 * u) * n + e];
        }
#pragma unroll
        for (int u = 0; u < 16; ++u)
            if (c + 8 * u < nchunks) s16[u] += part[(size_t)(c + 8 * u) * n + e];
#pragma unroll
        for (int w = 8; w > 0; w >>= 1)
#pragma unroll
            for (int u = 0; u < w; ++u) s16[u] += s16[u + w];
        s = s16[0];
    }
    red[grp][el] = s;
    __syncthreads();
    if (grp != 0 || e >= n) return;
    s = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) s += red[q][el];
    if (e < Hd * A1) {
        const int kx = e / Hd, i = e - kx * Hd;          // the partials are column-major (see head_kernel)
        if (kx < A) grads[nd.fc2W + (long long)i * A + kx] = s;
        else grads[nd.fc3W + i] = s;
    } else {
        const int kx = e - Hd * A1;
        if (kx < A) grads[nd.fc2b + kx] = s;
        else grads[nd.fc3b] = s;
    }
}

// ---------------------------------------------------------------------------
// linear modules backward
// ---------------------------------------------------------------------------
// d(flat)[t][k] = sum_{j active} sum_c dh[t][c]*mask_j[t][c] * W_j[k][c]
__global__ void __launch_bounds__(256)
lin_dgrad_kernel(const float* __restrict__ dh, const uint8_t* __restrict__ Rl, const float* __restrict__ params,
                 PnNetDesc nd, const uint32_t* __restrict__ masks, int B, float* __restrict__ dflat) {
    __shared__ float Ds[32][33];
    __shared__ float Wt[64][33];
    const int p = blockIdx.x, k0 = blockIdx.y * 64, b0 = blockIdx.z * 32;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int K = nd.last_lin, Hd = nd.hidden;
    const uint32_t mask = masks[p * nd.L + nd.L - 1];
    float acc[2][4];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
    for (int j = 0; j < nd.M; ++j) {
        if (!((mask >> j) & 1u)) continue;
        const float* W = params + nd.linW_base + (long long)j * nd.lin_stride;
        for (int c0 = 0; c0 < Hd; c0 += 32) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int idx = tid + i * 256, row = idx >> 5, c = idx & 31, b = b0 + row;
                float val = 0.f;
                if (b < B && c0 + c < Hd) {
                    const size_t t = (size_t)p * B + b;
                    val = Rl[(t * nd.M + j) * Hd + c0 + c] ? dh[t * Hd + c0 + c] : 0.f;
                }
                Ds[row][c] = val;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int idx = tid + i * 256, kk = idx >> 5, c = idx & 31;
                Wt[kk][c] = (k0 + kk < K && c0 + c < Hd) ? W[(size_t)(k0 + kk) * Hd + c0 + c] : 0.f;
            }
            __syncthreads();
#pragma unroll 8
            for (int c = 0; c < 32; ++c) {
                const float d0 = Ds[ty * 2][c], d1 = Ds[ty * 2 + 1][c];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float wv = Wt[tx + 16 * q][c];
                    acc[0][q] = fmaf(d0, wv, acc[0][q]);
                    acc[1][q] = fmaf(d1, wv, acc[1][q]);
                }
            }
            __syncthreads();
        }
    }
#pragma unroll
    for (int a = 0; a < 2; ++a) {
        const int b = b0 + ty * 2 + a;
        if (b >= B) continue;
        const size_t t = (size_t)p * B + b;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int kg = k0 + tx + 16 * q;
            if (kg < K) dflat[t * K + kg] = acc[a][q];
        }
    }
}

// dW_lin[j][k][c] = sum over paths using j (ascending) and their transitions of flat[t][k]*dpre_j[t][c]
__global__ void __launch_bounds__(256)
lin_wgrad_kernel(const float* __restrict__ X, const float* __restrict__ dh, const uint8_t* __restrict__ Rl,
                 PnNetDesc nd, const int* __restrict__ list, const int* __restrict__ list_count, int P, int B,
                 float* __restrict__ grads) {
    __shared__ __align__(16) float Xs[16][64];
    __shared__ __align__(16) float Ds[16][64];
    const int j = blockIdx.x, k0 = blockIdx.y * 64, c0 = blockIdx.z * 64;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int K = nd.last_lin, Hd = nd.hidden, lm = (nd.L - 1) * nd.M + j;
    const int total = list_count[lm] * B;
    const int* lst = list + (size_t)lm * P;
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
    float bacc = 0.f;
    const bool do_bias = (blockIdx.y == 0) && tid < 64;
    for (int r0 = 0; r0 < total; r0 += 16) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = tid + i * 256, rr = idx >> 6, cc = idx & 63, row = r0 + rr;
            float xv = 0.f, dv = 0.f;
            if (row < total) {
                const size_t t = (size_t)lst[row / B] * B + (row % B);
                if (k0 + cc < K) xv = X[t * K + k0 + cc];
                if (c0 + cc < Hd) dv = Rl[(t * nd.M + j) * Hd + c0 + cc] ? dh[t * Hd + c0 + cc] : 0.f;
            }
            Xs[rr][cc] = xv;
            Ds[rr][cc] = dv;
        }
        __syncthreads();
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) {
            const float4 x4 = *reinterpret_cast<const float4*>(&Xs[rr][ty * 4]);
            const float4 d4 = *reinterpret_cast<const float4*>(&Ds[rr][tx * 4]);
            const float xa[4] = {x4.x, x4.y, x4.z, x4.w}, da[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(xa[a], da[b], acc[a][b]);
            if (do_bias) bacc += Ds[rr][tid];
        }
        __syncthreads();
    }
    float* G = grads + nd.linW_base + (long long)j * nd.lin_stride;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int kg = k0 + ty * 4 + a;
        if (kg >= K) continue;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int col = c0 + tx * 4 + b;
            if (col < Hd) G[(size_t)kg * Hd + col] = acc[a][b];
        }
    }
    if (do_bias && c0 + tid < Hd) G[nd.linb_off + c0 + tid] = bacc;
}

// ---------------------------------------------------------------------------
// conv layers backward
// ---------------------------------------------------------------------------
// data gradient of conv layer l >= 1 (Cin == 8): one CTA per transition
__global__ void __launch_bounds__(256)
conv_dgrad_kernel(const float* __restrict__ dY, const uint8_t* __restrict__ R, const float* __restrict__ params,
                  PnNetDesc nd, int l, const uint32_t* __restrict__ masks, int B, float* __restrict__ dX) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int t = blockIdx.x, p = t / B;
    const int IH = nd.IH[l], IW = nd.IW[l], k = nd.k[l], s = nd.s[l], OH = nd.OH[l], OW = nd.OW[l], KK = nd.KK[l];
    const int npix = OH * OW;
    const uint32_t mask = masks[p * nd.L + l];
    const int nact = __popc(mask);
    float* dYs = reinterpret_cast<float*>(smem_raw);          // [npix*8]
    float* Ws = dYs + ((npix * PN_F + 3) & ~3);                // [nact][KK*8]
    uint8_t* Rs = reinterpret_cast<uint8_t*>(Ws + (size_t)nact * KK * PN_F);   // [nact][npix]
    for (int i = threadIdx.x; i < npix * PN_F; i += blockDim.x) dYs[i] = dY[(size_t)t * npix * PN_F + i];
    int slot = 0;
    for (int j = 0; j < nd.M; ++j) {
        if (!((mask >> j) & 1u)) continue;
        const float* Wg = params + nd.convW_base[l] + (long long)j * nd.conv_stride[l];
        for (int i = threadIdx.x; i < KK * PN_F; i += blockDim.x) Ws[(size_t)slot * KK * PN_F + i] = Wg[i];
        for (int i = threadIdx.x; i < npix; i += blockDim.x) Rs[slot * npix + i] = R[((size_t)t * nd.M + j) * npix + i];
        ++slot;
    }
    __syncthreads();
    for (int ipix = threadIdx.x; ipix < IH * IW; ipix += blockDim.x) {
        const int iy = ipix / IW, ix = ipix % IW;
        float acc[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[c] = 0.f;
        for (int sl = 0; sl < nact; ++sl) {
            const float* Wsl = Ws + (size_t)sl * KK * PN_F;
            for (int kh = 0; kh < k; ++kh) {
                const int ry = iy - kh;
                if (ry < 0 || (ry % s) != 0) continue;
                const int oy = ry / s;
                if (oy >= OH) continue;
                for (int kw = 0; kw < k; ++kw) {
                    const int rx = ix - kw;
                    if (rx < 0 || (rx % s) != 0) continue;
                    const int ox = rx / s;
                    if (ox >= OW) continue;
                    const int opix = oy * OW + ox;
                    const unsigned bits = Rs[sl * npix + opix];
                    if (!bits) continue;
                    float d[8];
#pragma unroll
                    for (int co = 0; co < 8; ++co) d[co] = ((bits >> co) & 1u) ? dYs[opix * PN_F + co] : 0.f;
                    const float4* w4 = reinterpret_cast<const float4*>(Wsl + (size_t)(kh * k + kw) * 8 * PN_F);
#pragma unroll
                    for (int ci = 0; ci < 8; ++ci) {
                        const float4 a = w4[2 * ci], b = w4[2 * ci + 1];
                        acc[ci] += d[0] * a.x + d[1] * a.y + d[2] * a.z + d[3] * a.w + d[4] * b.x + d[5] * b.y +
                                   d[6] * b.z + d[7] * b.w;
                    }
                }
            }
        }
        float4* o = reinterpret_cast<float4*>(dX + ((size_t)t * IH * IW + ipix) * 8);
        o[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
        o[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
}

// weight-gradient partials of conv layer l: CTA = (path, chunk of its transitions),
// thread = one im2col row kk = (kh, kw, ci); the layer input is read through L1.
template <typename TIn>
__global__ void __launch_bounds__(1024)
conv_wgrad_kernel(const TIn* __restrict__ X, float in_scale, const float* __restrict__ dY,
                  const uint8_t* __restrict__ R, PnNetDesc nd, int l, const uint32_t* __restrict__ masks, int B,
                  int cs, int nch_alloc, float* __restrict__ part) {
    extern __shared__ __align__(16) float Dsm[];   // [npix*8]
    const int p = blockIdx.x, ch = blockIdx.y;
    const int IH = nd.IH[l], IW = nd.IW[l], Cin = nd.Cin[l], k = nd.k[l], s = nd.s[l];
    const int OH = nd.OH[l], OW = nd.OW[l], KK = nd.KK[l], npix = OH * OW;
    const int b_lo = ch * cs, b_hi = min(B, b_lo + cs);
    const int kk = threadIdx.x;
    const bool valid = kk < KK;
    const int kh = valid ? kk / (k * Cin) : 0, rem = valid ? kk % (k * Cin) : 0;
    const int base = kh * IW * Cin + rem;
    const uint32_t mask = masks[p * nd.L + l];
    const int stride_part = KK * PN_F + PN_F;
    for (int j = 0; j < nd.M; ++j) {
        if (!((mask >> j) & 1u)) continue;
        float acc[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[c] = 0.f;
        float bacc = 0.f;
        for (int b = b_lo; b < b_hi; ++b) {
            const size_t t = (size_t)p * B + b;
            __syncthreads();
            for (int i = threadIdx.x; i < npix * PN_F; i += blockDim.x) {
                const unsigned bits = R[(t * nd.M + j) * npix + (i >> 3)];
                Dsm[i] = ((bits >> (i & 7)) & 1u) ? dY[t * npix * PN_F + i] : 0.f;
            }
            __syncthreads();
            if (valid) {
                const TIn* xt = X + t * (size_t)IH * IW * Cin + base;
                for (int oy = 0; oy < OH; ++oy) {
                    const TIn* xr = xt + (size_t)oy * s * IW * Cin;
                    const float4* d4 = reinterpret_cast<const float4*>(Dsm + (size_t)oy * OW * PN_F);
                    for (int ox = 0; ox < OW; ++ox) {
                        const float a = to_f(__ldg(xr + ox * s * Cin));
                        const float4 d0 = d4[2 * ox], d1 = d4[2 * ox + 1];
                        acc[0] = fmaf(a, d0.x, acc[0]); acc[1] = fmaf(a, d0.y, acc[1]);
                        acc[2] = fmaf(a, d0.z, acc[2]); acc[3] = fmaf(a, d0.w, acc[3]);
                        acc[4] = fmaf(a, d1.x, acc[4]); acc[5] = fmaf(a, d1.y, acc[5]);
                        acc[6] = fmaf(a, d1.z, acc[6]); acc[7] = fmaf(a, d1.w, acc[7]);
                    }
                }
            }
            if (threadIdx.x < PN_F) {
                float sb = 0.f;
                for (int px = 0; px < npix; ++px) sb += Dsm[px * PN_F + threadIdx.x];
                bacc += sb;
            }
        }
        float* out = part + ((size_t)(p * nch_alloc + ch) * nd.M + j) * stride_part;
        if (valid) {
            float4* o = reinterpret_cast<float4*>(out + (size_t)kk * PN_F);
            o[0] = make_float4(acc[0] * in_scale, acc[1] * in_scale, acc[2] * in_scale, acc[3] * in_scale);
            o[1] = make_float4(acc[4] * in_scale, acc[5] * in_scale, acc[6] * in_scale, acc[7] * in_scale);
        }
        if (threadIdx.x < PN_F) out[KK * PN_F + threadIdx.x] = bacc;
    }
}

// fixed-order reduction of the partials over the paths that use module (l, j).
// block = 32 elements x 8 term groups (term = (path in the module's list, chunk)); group g adds terms
// g, g+8, ... in order, then the 8 group sums are added in group order.
__device__ __forceinline__ void conv_reduce_body(int j, int by, const float* __restrict__ part, const PnNetDesc& nd, int l,
                                                 const int* __restrict__ list, const int* __restrict__ list_count, int P,
                                                 int nch_alloc, int nch_active, float* __restrict__ grads, float (*red)[33]) {
    const int KK = nd.KK[l];
    const int n = KK * PN_F + PN_F;
    const int el = threadIdx.x & 31, grp = threadIdx.x >> 5;
    const int e = by * 32 + el;
    const int lm = l * nd.M + j, cnt = list_count[lm];
    const int* lst = list + (size_t)lm * P;
    const int nterms = cnt * nch_active;
    // the module's path list goes through shared memory first: a list load feeding the address of every partial load
    // would put two dependent global round trips into each term
    __shared__ int lst_s[1024];
    const bool staged = cnt <= 1024;
    if (staged) {
        for (int i = threadIdx.x; i < cnt; i += blockDim.x) lst_s[i] = lst[i];
        __syncthreads();
    }
    float s = 0.f;
    if (e < n) {
        // group g adds terms g, g + 8, ... (term i = (path lst[i / nch], chunk i % nch)); the (path, chunk) pair is advanced
        // incrementally -- an integer division per term made this loop the launch's whole instruction count -- and eight
        // loads are kept in flight; the eight partial sums are combined in a fixed order
        float s8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        const float* part_je = part + (size_t)j * n + e;
        const size_t stride_pc = (size_t)nd.M * n;
        const int dpi = 8 / nch_active, dc = 8 % nch_active;
        int pi = grp / nch_active, c = grp % nch_active;
        auto next_term = [&]() -> const float* {
            const int p = staged ? lst_s[pi] : lst[pi];
            const float* a = part_je + (size_t)(p * nch_alloc + c) * stride_pc;
            pi += dpi; c += dc;
            if (c >= nch_active) { c -= nch_active; ++pi; }
            return a;
        };
        int i = grp;
        for (; i + 56 < nterms; i += 64) {      // eight independent loads per round trip
            float v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = __ldg(next_term());
#pragma unroll
            for (int u = 0; u < 8; ++u) s8[u] += v[u];
        }
        for (int u = 0; i < nterms; i += 8, ++u) s8[u] += __ldg(next_term());
        s = ((s8[0] + s8[1]) + (s8[2] + s8[3])) + ((s8[4] + s8[5]) + (s8[6] + s8[7]));
    }
    red[grp][el] = s;
    __syncthreads();
    if (grp != 0 || e >= n) return;
    s = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) s += red[q][el];
    const long long wb = nd.convW_base[l] + (long long)j * nd.conv_stride[l];
    if (e < KK * PN_F) grads[wb + e] = s;
    else grads[wb + nd.convb_off[l] + (e - KK * PN_F)] = s;
}

// One launch finishes every gradient that is accumulated through partial buffers: the conv layers' weight
// gradients (sum over the paths of each module's list and their chunks), the head weights (sum over chunks) and
// the per-path loss.  All sums run in a fixed order.
struct PnReduceJobs {
    int nconv;                                   // conv jobs in use
    int l[PN_MAX_LAYERS], nch_active[PN_MAX_LAYERS], nby[PN_MAX_LAYERS], blk0[PN_MAX_LAYERS];
    const float* part[PN_MAX_LAYERS];
    int head_blk0, head_nred, head_nchunks;
};

__global__ void __launch_bounds__(256)
grad_reduce_kernel(PnReduceJobs jobs, PnNetDesc nd, const int* __restrict__ list, const int* __restrict__ list_count, int P,
                   int B, int nch_alloc, const float* __restrict__ head_part, const float* __restrict__ loss_t,
                   float* __restrict__ loss, float* __restrict__ grads) {
    pn_pdl_sync();
    __shared__ float red[8][33];
    const int bid = blockIdx.x;
    if (bid >= jobs.head_blk0) {
        head_reduce_body(bid - jobs.head_blk0, head_part, jobs.head_nchunks, nd, grads, jobs.head_nred, loss_t, P, B, loss, red);
        return;
    }
    int q = 0;
    while (q + 1 < jobs.nconv && bid >= jobs.blk0[q + 1]) ++q;
    const int local = bid - jobs.blk0[q];
    conv_reduce_body(local / jobs.nby[q], local % jobs.nby[q], jobs.part[q], nd, jobs.l[q], list, list_count, P, nch_alloc,
                     jobs.nch_active[q], grads, red);
}

template <typename K>
int set_smem(pn_handle* h, K kern, size_t bytes) {
    if (bytes > 48 * 1024) {
        PN_SET_SMEM(h, kern, bytes);
    }
    return PN_OK;
}

}  // namespace

// ===========================================================================
// host launchers
// ===========================================================================
int pn_launch_forward(pn_handle* h, const void* frames, int is_u8, int B, bool training, cudaStream_t st) {
    const PnNetDesc& nd = h->nd;
    const int T = h->P * B;
    (void)training;
    // operand preparation that depends on the parameters only (bf16 hi/lo weight blocks of the middle conv layers and
    // the linear modules) runs on the side stream, next to the first conv layer
    bool forked = false;
    {
        bool any = h->tc_lin;
        for (int l = 1; l < nd.nconv; ++l) any = any || h->tc_mid[l];
        if (any) {
            PN_CUDA(h, cudaEventRecord(h->ev_fork, st));
            PN_CUDA(h, cudaStreamWaitEvent(h->side, h->ev_fork, 0));
            for (int l = 1; l < nd.nconv; ++l)
                if (h->tc_mid[l])
                    if (int rc = pn_launch_conv_mid_wprep(h, l, h->side)) return rc;
            if (h->tc_lin)
                if (int rc = pn_launch_lin_wsplit(h, h->side)) return rc;
            PN_CUDA(h, cudaEventRecord(h->ev_join, h->side));
            forked = true;
        }
    }
    for (int l = 0; l < nd.nconv; ++l) {
        if (l == 1 && forked) { PN_CUDA(h, cudaStreamWaitEvent(st, h->ev_join, 0)); forked = false; }
        if (l == 0 && is_u8 && h->tc_conv0) {
            if (int rc = pn_launch_conv0_fwd_tc(h, (const uint8_t*)frames, B, st)) return rc;
            continue;
        }
        if (l >= 1 && h->tc_mid[l]) {
            if (int rc = pn_launch_conv_mid_fwd_tc(h, l, B, st)) return rc;
            continue;
        }
        if (l >= 1 && !(h->cfg.flags & 4) && pn_conv_mid_supported(h, l)) {
            if (int rc = pn_launch_conv_mid_fwd(h, l, B, st)) return rc;
            continue;
        }
        const int rpb = max(1, 128 / nd.OW[l]);
        const int nbands = (nd.OH[l] + rpb - 1) / rpb;
        const int in_rows = (min(rpb, nd.OH[l]) - 1) * nd.s[l] + nd.k[l];
        const size_t wbytes = ((size_t)nd.KK[l] * PN_F + PN_F) * sizeof(float);
        dim3 grid(T, nbands);
        PnProfScope prof(h, PN_PROF_CONV_FWD + l, st);
        if (l == 0 && is_u8) {
            const size_t sm = wbytes + (size_t)in_rows * nd.IW[l] * nd.Cin[l];
            if (int rc = set_smem(h, conv_fwd_kernel<uint8_t>, sm)) return rc;
            conv_fwd_kernel<uint8_t><<<grid, 128, sm, st>>>((const uint8_t*)frames, 1.0f / 255.0f, h->params, nd, l,
                                                            h->d_masks, B, rpb, h->act[l], h->rmask[l]);
        } else {
            const float* in = l == 0 ? (const float*)frames : h->act[l - 1];
            const size_t sm = wbytes + (size_t)in_rows * nd.IW[l] * nd.Cin[l] * sizeof(float);
            if (int rc = set_smem(h, conv_fwd_kernel<float>, sm)) return rc;
            conv_fwd_kernel<float><<<grid, 128, sm, st>>>(in, 1.0f, h->params, nd, l, h->d_masks, B, rpb, h->act[l],
                                                          h->rmask[l]);
        }
        PN_LAUNCH_CHECK(h, "conv_fwd_kernel");
    }
    if (forked) { PN_CUDA(h, cudaStreamWaitEvent(st, h->ev_join, 0)); forked = false; }
    if (h->tc_lin) return pn_launch_lin_fwd_tc(h, B, st);
    dim3 grid(h->P, (nd.hidden + 63) / 64, (B + 31) / 32);
    {
        PnProfScope prof(h, PN_PROF_LIN_FWD, st);
        lin_fwd_kernel<<<grid, 256, 0, st>>>(h->act[nd.nconv - 1], h->params, nd, h->d_masks, B, h->hbuf, h->lmask);
    }
    PN_LAUNCH_CHECK(h, "lin_fwd_kernel");
    (void)T;
    return PN_OK;
}

int pn_launch_heads(pn_handle* h, int B, bool training, const int32_t* actions, const float* td, const float* r,
                    float beta, float* pi, float* v, float* loss, cudaStream_t st) {
    const PnNetDesc& nd = h->nd;
    const int T = h->P * B;
    const size_t hpitch = (((size_t)nd.hidden + 3) & ~(size_t)3) + 16;
    const int np = nd.A + 1 <= 8 ? 8 : (nd.A + 1 <= 20 ? 20 : 32);
    const size_t sm = ((size_t)nd.hidden * np + 32 + PN_HEAD_CHUNK * hpitch + 2 * PN_HEAD_CHUNK * 32) * sizeof(float);
    // one CTA per chunk while they all fit the partial buffer (measured: letting a CTA walk over several chunks to amortise
    // the weight staging made the kernel slower, 35 -> 41 us at 8192 transitions: it is latency-, not traffic-bound)
    const int nchunks = std::min((T + PN_HEAD_CHUNK - 1) / PN_HEAD_CHUNK, h->head_chunks);
    h->head_ctas = nchunks;
    auto kern = np == 8 ? head_kernel<8> : (np == 20 ? head_kernel<20> : head_kernel<32>);
    if (int rc = set_smem(h, kern, sm)) return rc;
    {
        PnProfScope prof(h, PN_PROF_HEAD, st);
        pn_launch(pn_use_pdl(h), kern, dim3(nchunks), dim3(256), sm, st, (const float*)pn_head_input(h), (const float*)h->params, nd, T,
                  training ? 1 : 0, actions, td, r, beta, pi, v, h->dzv, h->dh, h->loss_t, h->head_part);
    }
    PN_LAUNCH_CHECK(h, "head_kernel");
    h->pending_loss = training ? loss : nullptr;     // summed per path by the head-gradient reduce launch
    return PN_OK;
}

int pn_launch_backward(pn_handle* h, const void* frames, int is_u8, int B, cudaStream_t st) {
    const PnNetDesc& nd = h->nd;
    const int T = h->P * B, P = h->P;
    PnReduceJobs jobs;
    jobs.nconv = 0;
    auto add_job = [&](int l, int nch_active) {
        const int q = jobs.nconv++;
        jobs.l[q] = l; jobs.nch_active[q] = nch_active; jobs.part[q] = h->wpart[l];
        jobs.nby[q] = (nd.KK[l] * PN_F + PN_F + 31) / 32;
    };
    // linear modules
    if (h->tc_lin) {
        if (int rc = pn_launch_lin_bwd_tc(h, B, st)) return rc;
    } else {
        const int K = nd.last_lin;
        dim3 g1(P, (K + 63) / 64, (B + 31) / 32);
        {
            PnProfScope prof(h, PN_PROF_LIN_DGRAD, st);
            lin_dgrad_kernel<<<g1, 256, 0, st>>>(h->dh, h->lmask, h->params, nd, h->d_masks, B, h->dact[nd.nconv - 1]);
        }
        PN_LAUNCH_CHECK(h, "lin_dgrad_kernel");
        dim3 g2(nd.M, (K + 63) / 64, (nd.hidden + 63) / 64);
        {
            PnProfScope prof(h, PN_PROF_LIN_WGRAD, st);
            lin_wgrad_kernel<<<g2, 256, 0, st>>>(h->act[nd.nconv - 1], h->dh, h->lmask, nd, h->d_list, h->d_list_count,
                                                 P, B, h->grads);
        }
        PN_LAUNCH_CHECK(h, "lin_wgrad_kernel");
    }
    if (!h->tc_lin) PN_CUDA(h, cudaEventRecord(h->ev_lin_grads, st));     // (the tensor-core launcher records it itself)
    // conv layers, last to first
    const int nch_simt = std::min(h->nch, std::max(1, (296 + P - 1) / P));
    const int cs = (B + nch_simt - 1) / nch_simt;
    int nch_active = (B + cs - 1) / cs;
    // The partials of the middle conv layers and of the heads are complete before the first conv layer's weight-gradient
    // kernel starts: their reduction runs on the side stream underneath it (that kernel is the longest of the step), and
    // only layer 0's own partials are reduced after it.
    bool early_done = false, side_used = false;
    auto launch_reduce = [&](PnReduceJobs& jb, bool with_head, cudaStream_t rs) -> int {
        int blk = 0;
        for (int q = 0; q < jb.nconv; ++q) { jb.blk0[q] = blk; blk += nd.M * jb.nby[q]; }
        const int nh = nd.hidden * (nd.A + 1) + nd.A + 1;
        jb.head_blk0 = blk;
        jb.head_nred = with_head ? (nh + 31) / 32 : 0;
        jb.head_nchunks = h->head_ctas;               // one partial of the head weight gradient per CTA of the head kernel
        const int nloss = (with_head && h->pending_loss) ? (P + 255) / 256 : 0;
        if (blk + jb.head_nred + nloss == 0) return PN_OK;
        PnProfScope prof(h, PN_PROF_HEAD_WGRAD, rs);
        pn_launch(pn_use_pdl(h), grad_reduce_kernel, dim3(blk + jb.head_nred + nloss), dim3(256), 0, rs, jb, nd, (const int*)h->d_list,
                  (const int*)h->d_list_count, P, B, h->nch, (const float*)h->head_part, (const float*)h->loss_t, h->pending_loss, h->grads);
        if (with_head) h->pending_loss = nullptr;
        PN_LAUNCH_CHECK(h, "grad_reduce_kernel");
        return PN_OK;
    };
    for (int l = nd.nconv - 1; l >= 0; --l) {
        const int npix = nd.OH[l] * nd.OW[l];
        if (l == 0 && nd.nconv > 1 && !h->prof.on) {
            // fork: everything launched so far on `st` has produced the partials below
            PN_CUDA(h, cudaEventRecord(h->ev_fork, st));
            PN_CUDA(h, cudaStreamWaitEvent(h->side, h->ev_fork, 0));
            if (int rc = launch_reduce(jobs, true, h->side)) return rc;
            PN_CUDA(h, cudaEventRecord(h->ev_join, h->side));
            PN_CUDA(h, cudaEventRecord(h->ev_early_grads, h->side));
            jobs.nconv = 0;
            early_done = true;
        }
        if (l >= 1 && !(h->cfg.flags & 4) && pn_conv_mid_supported(h, l)) {
            int nch_used = 1;
            const bool wg_tc = pn_conv_mid_wgrad_tc_usable(h, l);
            const bool dg_tc = pn_conv_mid_tc_has_bwd(h, l);
            // The layer's weight gradient and data gradient both need only dact[l]: the weight gradient goes to the side
            // stream and runs beside the data-gradient chain (they stress different units: MMA pipe vs converters / epilogue).
            // With per-kernel timing on, everything stays on one stream so that the event pairs mean what they say.
            cudaStream_t wst = st;
            if (wg_tc && dg_tc && !h->prof.on && !(h->cfg.flags & PN_FLAG_SERIAL_BWD)) {
                PN_CUDA(h, cudaEventRecord(h->ev_fork, st));
                PN_CUDA(h, cudaStreamWaitEvent(h->side, h->ev_fork, 0));
                wst = h->side;
                side_used = true;
            }
            if (wg_tc)
                if (int rc = pn_launch_conv_mid_wgrad_tc(h, l, B, wst, &nch_used)) return rc;
            if (!wg_tc || !dg_tc)
                if (int rc = pn_launch_conv_mid_bwd(h, l, B, st, &nch_used, !wg_tc, !dg_tc)) return rc;
            if (dg_tc)
                if (int rc = pn_launch_conv_mid_dgrad_tc(h, l, B, st)) return rc;
            add_job(l, nch_used);
            continue;
        }
        if (l == 0 && is_u8 && h->tc_conv0 && pn_conv0_wgrad_tc_supported(h)) {
            if (int rc = pn_launch_conv0_wgrad_tc(h, (const uint8_t*)frames, B, st)) return rc;
            add_job(l, h->last_wgrad_nch);
            continue;
        }
        const int threads = (nd.KK[l] + 31) / 32 * 32;
        const size_t sm = (size_t)npix * PN_F * sizeof(float);
        dim3 grid(P, nch_active);
        if (int rc = set_smem(h, conv_wgrad_kernel<uint8_t>, sm)) return rc;
        if (int rc = set_smem(h, conv_wgrad_kernel<float>, sm)) return rc;
        {
        PnProfScope prof(h, PN_PROF_CONV_WGRAD + l, st);
        if (l == 0 && is_u8) {
            conv_wgrad_kernel<uint8_t><<<grid, threads, sm, st>>>((const uint8_t*)frames, 1.0f / 255.0f, h->dact[l],
                                                                  h->rmask[l], nd, l, h->d_masks, B, cs, h->nch,
                                                                  h->wpart[l]);
        } else {
            const float* in = l == 0 ? (const float*)frames : h->act[l - 1];
            conv_wgrad_kernel<float><<<grid, threads, sm, st>>>(in, 1.0f, h->dact[l], h->rmask[l], nd, l, h->d_masks,
                                                                B, cs, h->nch, h->wpart[l]);
        }
        }
        PN_LAUNCH_CHECK(h, "conv_wgrad_kernel");
        add_job(l, nch_active);
        if (l > 0) {
            const size_t smd = (((size_t)npix * PN_F + 3) & ~(size_t)3) * sizeof(float) +
                               (size_t)nd.M * nd.KK[l] * PN_F * sizeof(float) + (size_t)nd.M * npix;
            if (int rc = set_smem(h, conv_dgrad_kernel, smd)) return rc;
            PnProfScope prof(h, PN_PROF_CONV_DGRAD + l, st);
            conv_dgrad_kernel<<<T, 256, smd, st>>>(h->dact[l], h->rmask[l], h->params, nd, l, h->d_masks, B,
                                                   h->dact[l - 1]);
            PN_LAUNCH_CHECK(h, "conv_dgrad_kernel");
        }
    }
    if (side_used && !early_done) {                      // (cannot happen today: a middle layer implies nconv > 1)
        PN_CUDA(h, cudaEventRecord(h->ev_join, h->side));
        PN_CUDA(h, cudaStreamWaitEvent(st, h->ev_join, 0));
    }
    if (int rc = launch_reduce(jobs, !early_done, st)) return rc;
    if (early_done) PN_CUDA(h, cudaStreamWaitEvent(st, h->ev_join, 0));
    else PN_CUDA(h, cudaEventRecord(h->ev_early_grads, st));
    return PN_OK;
}
