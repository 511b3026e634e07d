// Linear modules (last PathNet layer) on the tensor cores: tcgen05 kind::f16 with bf16 operands and
// fp32 accumulation in TMEM.
//
// Reference semantics (game_ac_network.py:184-194): net = (sum over active j, ascending, of
// relu(flat @ W_lin[j] + b_lin[j])) / M;  backward through tf.gradients (a3c_training_thread.py:55-60).
//
// Precision: every fp32 operand x is split into bf16 hi = rn(x) and lo = rn(x - hi); a product is
// evaluated as hi*hi + hi*lo + lo*hi (three MMAs into the same fp32 accumulator).  The dropped lo*lo term
// and the bf16 rounding of lo leave a relative error of ~2^-16 per product, far inside the 1e-3 budget.
//
// Operand storage: 8x8 bf16 blocks of 128 bytes (8 rows of 16 bytes).  One block layout serves every use
// because a tcgen05 "core matrix" is 8 rows x 16 bytes in both the K-major and the MN-major descriptor
// form -- only the roles of the two strides change:
//   Wp  [j][hl][h][kb 0..KB-1][cb16 0..15][8 k][8 c]   W_lin[j] rows k, columns c = h*128 + cb16*8 + ..
//   Xp  [p][bc][hl][kb 0..KB-1][bb 0..3][8 b][8 k]     flat activations of path p, rows b = bc*32 + bb*8 + ..
//   Dp  [p][bc][slot][hl][cb 0..31][bb 0..3][8 b][8 c] dh * relu-mask of the slot's module
// (hl = 0 hi / 1 lo; KB = K/8 + 1 rounded up to even: at least one extra k-block; in Xp its first k holds 1.0 so that the weight
// gradient GEMM also produces the bias gradient, in Wp it is zero.)
//   forward : D[c, b]  = sum_k  W[k, c] X[b, k]       A = Wp (MN-major), B = Xp (K-major)
//   dgrad   : D[k, b]  = sum_c  W[k, c] dpre[b, c]    A = Wp (K-major),  B = Dp (K-major)
//   wgrad   : D[k, c]  = sum_b  X[b, k] dpre[b, c]    A = Xp (MN-major), B = Dp (MN-major)
// All loads are cp.async.bulk of contiguous pieces of those buffers; no conversion work in the GEMM kernels.
#include "pn_internal.cuh"
#include "pn_tc.cuh"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <algorithm>

using namespace pntc;

namespace {

constexpr int kBlk = 128;                  // bytes of one 8x8 bf16 block
constexpr int kNB = 32;                    // batch rows per GEMM (N of forward / dgrad)
constexpr int kWChunkKb = 6;               // k-blocks per forward W stage: 3 K steps, one per hi*hi accumulator (see lin_fwd_tc_kernel)
constexpr int kWStages = 5;

struct LinGeom {
    int K, KB, Hd, M, L;                   // last_lin, K/8 + 1, hidden, modules, layers
    int nbc;                               // 32-row batch chunks per path = ceil(B/32)
    size_t wp_module;                      // bytes of Wp per module
    size_t xp_chunk, xp_path;              // bytes of Xp per (path, chunk) and per path
    size_t dp_slot, dp_chunk, dp_path;     // bytes of Dp per (path, chunk, slot), per chunk (M slots), per path
};

__device__ __forceinline__ void split_bf16(float x, __nv_bfloat16& hi, __nv_bfloat16& lo) {
    hi = __float2bfloat16_rn(x);
    lo = __float2bfloat16_rn(x - __bfloat162float(hi));
}

// The weights (Wp) and the activations (Xp) are stored as fp16 hi/lo instead: x * scale ~ hi + lo to ~2^-22 relative
// (11 + 11 mantissa bits).  The forward GEMM decides the ReLU masks, and with bf16 hi/lo (2^-16) a few pre-activations
// per million fall on the other side of zero than in fp32 arithmetic, which moves small gradient tensors by more than the
// 1e-3 budget.  Power-of-two scales (exact) keep the lo parts out of fp16's subnormal range: weights x 2^8, activations
// x 2^4 (so activations must stay below 65504 / 16); the forward epilogue undoes them.  The backward GEMMs keep bf16 copies
// (Wp, Xp) next to the gradient operand Dp, whose dynamic range needs bf16: one tcgen05.mma cannot mix an fp16 operand with a
// bf16 one (the GPU reports an illegal instruction), so the prep kernels write both forms from one read.
constexpr float kWScale = 256.0f, kXScale = 16.0f;
__device__ __forceinline__ void split_row8_f16(const float* v, float scale, uint4& hi, uint4& lo) {
    __half h[8], l[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float x = v[i] * scale;
        h[i] = __float2half_rn(x);
        l[i] = __float2half_rn(x - __half2float(h[i]));
    }
    hi = *reinterpret_cast<uint4*>(h);
    lo = *reinterpret_cast<uint4*>(l);
}

// 8 consecutive floats -> one 16-byte row of the hi block and of the lo block
__device__ __forceinline__ void split_row8(const float* v, uint4& hi, uint4& lo) {
    __nv_bfloat16 h[8], l[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) split_bf16(v[i], h[i], l[i]);
    hi = *reinterpret_cast<uint4*>(h);
    lo = *reinterpret_cast<uint4*>(l);
}

// ---- prep: W_lin -> Wp (every step: RMSProp changed the weights) and flat activations -> Xp, one launch ----
// One thread per 16-byte output row.  Lane mapping: 8 consecutive lanes fill the 8 rows of one 128-byte block
// and groups of 4 lanes read 4 x 32 = 128 contiguous source bytes, so both sides move whole 128-byte lines.
__device__ __forceinline__ void wsplit_item(long long i, const float* __restrict__ params, const PnNetDesc& nd,
                                            const LinGeom& g, uint8_t* __restrict__ wp, uint8_t* __restrict__ wph) {
    const int ncb4 = g.Hd / 32;
    const int r = (int)(i & 7), cbq = (int)((i >> 3) & 3);
    long long rest = i >> 5;
    const int cb4 = (int)(rest % ncb4); rest /= ncb4;
    const int kb = (int)(rest % g.KB), j = (int)(rest / g.KB);
    if (j >= g.M) return;
    const int cb = cb4 * 4 + cbq, h = cb / 16, cb16 = cb % 16, k = kb * 8 + r;
    float v[8];
    if (k < g.K) {
        const float* W = params + nd.linW_base + (long long)j * nd.lin_stride + (size_t)k * g.Hd + cb * 8;
        const float4 a = *reinterpret_cast<const float4*>(W), b = *reinterpret_cast<const float4*>(W + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = 0.f;
    }
    uint4 hi, lo;
    const size_t off = (size_t)j * g.wp_module + ((size_t)h * g.KB + kb) * 16 * kBlk + (size_t)cb16 * kBlk + r * 16;
    split_row8(v, hi, lo);
    *reinterpret_cast<uint4*>(wp + off) = hi;
    *reinterpret_cast<uint4*>(wp + off + (size_t)2 * g.KB * 16 * kBlk) = lo;
    split_row8_f16(v, kWScale, hi, lo);
    *reinterpret_cast<uint4*>(wph + off) = hi;
    *reinterpret_cast<uint4*>(wph + off + (size_t)2 * g.KB * 16 * kBlk) = lo;
}

__device__ __forceinline__ void xsplit_item(long long i, const float* __restrict__ X, const LinGeom& g, int P, int B,
                                            uint8_t* __restrict__ xp, uint8_t* __restrict__ xph) {
    const int nkb4 = (g.KB + 3) / 4;
    const int kbq = (int)(i & 3), r8 = (int)((i >> 2) & 7), bb = (int)((i >> 5) & 3);
    long long rest = i >> 7;
    const int kb4 = (int)(rest % nkb4); rest /= nkb4;
    const int bc = (int)(rest % g.nbc), p = (int)(rest / g.nbc);
    const int kb = kb4 * 4 + kbq;
    if (p >= P || kb >= g.KB) return;
    const int b = bc * kNB + bb * 8 + r8;
    float v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = 0.f;
    if (b < B) {
        if (kb * 8 < g.K) {
            const float* src = X + ((size_t)p * B + b) * g.K + kb * 8;
            const float4 a = *reinterpret_cast<const float4*>(src), c = *reinterpret_cast<const float4*>(src + 4);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
        } else {
            v[0] = 1.0f;          // the "ones" k-row: the weight-gradient GEMM turns it into the bias gradient
        }
    }
    uint4 hi, lo;
    const size_t off = (size_t)p * g.xp_path + (size_t)bc * g.xp_chunk + ((size_t)kb * 4 + bb) * kBlk + r8 * 16;
    split_row8(v, hi, lo);
    *reinterpret_cast<uint4*>(xp + off) = hi;
    *reinterpret_cast<uint4*>(xp + off + g.xp_chunk / 2) = lo;
    split_row8_f16(v, kXScale, hi, lo);
    *reinterpret_cast<uint4*>(xph + off) = hi;
    *reinterpret_cast<uint4*>(xph + off + g.xp_chunk / 2) = lo;
}

__global__ void __launch_bounds__(256)
lin_prep_kernel(const float* __restrict__ params, PnNetDesc nd, LinGeom g, uint8_t* __restrict__ wp,
                uint8_t* __restrict__ wph, int wblocks, const float* __restrict__ X, int P, int B, uint8_t* __restrict__ xp,
                uint8_t* __restrict__ xph) {
    pn_pdl_sync();
    if ((int)blockIdx.x < wblocks) wsplit_item((long long)blockIdx.x * 256 + threadIdx.x, params, nd, g, wp, wph);
    else xsplit_item((long long)(blockIdx.x - wblocks) * 256 + threadIdx.x, X, g, P, B, xp, xph);
}

// ---- forward GEMM ----------------------------------------------------------------------------------------------------
// item = (path, batch chunk, half h of the hidden units); D[c = h*128 + lane, b] per active slot in TMEM columns
// slot*32.  Warps 0-3 epilogue, warp 4 bulk-copy producer, warp 5 MMA issue (warp-uniform code).
__device__ __forceinline__ long long lin_clock() {      // clock read that stays where it is written (debug brackets)
    long long t;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory");
    return t;
}

// Accumulators: the tensor core adds into its fp32 accumulator with truncation, so a long chain of accumulations leaves a
// small one-sided error (measured: ~3e-6 relative over the 75..123 MMAs of one output).  It cancels nowhere downstream --
// a sum over the batch such as the value head's bias gradient adds it up coherently -- so every slot owns FOUR accumulators of
// kNB columns: one for the small cross terms (hi*lo + lo*hi, magnitude 2^-11 of the result) and three that take the hi*hi
// MMAs of the K steps round-robin; the epilogue adds the four in fp32 (round-to-nearest).  Chain length at full magnitude:
// a third of the K steps instead of three times the K steps.  The slots rotate over a ring of kAccRing TMEM stages.
constexpr int kAccRing = 4;                // TMEM stages (slots in flight)
constexpr int kAccPerSlot = 4;             // cross | hh (ks % 3 == 0) | hh (1) | hh (2)
constexpr int kSlotCols = kAccPerSlot * kNB;

__global__ void __launch_bounds__(192, 1)
lin_fwd_tc_kernel(const uint8_t* __restrict__ wp, const uint8_t* __restrict__ xp, const float* __restrict__ params,
                  PnNetDesc nd, LinGeom g, const uint32_t* __restrict__ masks, int P, int B, float* __restrict__ Hout,
                  uint8_t* __restrict__ Rl, long long* __restrict__ dbg_arg) {
    pn_pdl_sync();
    long long* const dbg = kTcDebugBuild ? dbg_arg : nullptr;   // the instrumentation below compiles away in the product build
    const bool dbg9 = dbg && dbg[148 * 8] == 9;              // PN_TC_DEBUG=9: where the forward GEMM's roles wait
    const long long t_entry = dbg9 ? lin_clock() : 0;
    extern __shared__ __align__(1024) uint8_t smem[];
    const int w_stage_bytes = 2 * kWChunkKb * 16 * kBlk;             // hi + lo, 8 k-blocks x 16 c-blocks
    uint8_t* w_s = smem;
    uint8_t* x_s = smem + (size_t)kWStages * w_stage_bytes;          // [hl][KB][4][128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(x_s + g.xp_chunk);
    uint64_t* w_full = bars;                  // [kWStages]
    uint64_t* w_empty = bars + kWStages;      // [kWStages]
    uint64_t* x_full = bars + 2 * kWStages;
    uint64_t* x_empty = x_full + 1;
    uint64_t* slot_full = x_full + 2;         // [kAccRing] the slot in this TMEM stage is complete -> its epilogue runs under the next slot's MMAs
    uint64_t* slot_empty = slot_full + kAccRing;   // [kAccRing] the epilogue has read the stage
    __shared__ uint32_t tmem_base_s;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int nitems = P * g.nbc * 2;
    const int nkstep = (g.KB + 1) / 2;                               // K steps of 16
    const int nchunk = (g.KB + kWChunkKb - 1) / kWChunkKb;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kWStages; ++i) { mbar_init(&w_full[i], 1); mbar_init(&w_empty[i], 1); }
        mbar_init(x_full, 1); mbar_init(x_empty, 1);
        for (int i = 0; i < kAccRing; ++i) { mbar_init(&slot_full[i], 1); mbar_init(&slot_empty[i], 128); }
        fence_mbar_init();
    }
    if (warp == 5) tmem_alloc(&tmem_base_s, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tmem_base_s;

    if (warp == 4) {
        // ================= producer =================
        int stage = 0;
        uint32_t wphase = 0, it = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x, ++it) {
            const int h = item & 1, bc = (item >> 1) % g.nbc, p = (item >> 1) / g.nbc;
            const uint32_t mask = __reduce_or_sync(0xffffffffu, masks[p * g.L + g.L - 1]);
            const int ns = __popc(mask);
            mbar_wait(x_empty, (it & 1) ^ 1);
            if (elect_one()) {
                mbar_expect_tx(x_full, (uint32_t)g.xp_chunk);
                bulk_g2s(x_s, xp + (size_t)p * g.xp_path + (size_t)bc * g.xp_chunk, (uint32_t)g.xp_chunk, x_full);
            }
            __syncwarp();
            for (int slot = 0; slot < ns; ++slot) {
                const int j = __fns(mask, 0, slot + 1);
                const uint8_t* src = wp + (size_t)j * g.wp_module + (size_t)h * g.KB * 16 * kBlk;
                const size_t hl_stride = (size_t)2 * g.KB * 16 * kBlk;
                for (int c = 0; c < nchunk; ++c) {
                    const int nkb = min(kWChunkKb, g.KB - c * kWChunkKb);
                    const uint32_t bytes = (uint32_t)(nkb * 16 * kBlk);
                    mbar_wait(&w_empty[stage], wphase ^ 1);
                    if (elect_one()) {
                        uint8_t* dst = w_s + (size_t)stage * w_stage_bytes;
                        mbar_expect_tx(&w_full[stage], 2 * bytes);
                        bulk_g2s(dst, src + (size_t)c * kWChunkKb * 16 * kBlk, bytes, &w_full[stage]);
                        bulk_g2s(dst + w_stage_bytes / 2, src + hl_stride + (size_t)c * kWChunkKb * 16 * kBlk, bytes, &w_full[stage]);
                    }
                    __syncwarp();
                    if (++stage == kWStages) { stage = 0; wphase ^= 1; }
                }
            }
        }
    } else if (warp == 5) {
        // ================= MMA issue (whole warp, uniform values, elected instructions) =================
        int stage = 0;
        uint32_t wphase = 0, it = 0, sc = 0;                              // sc: slots issued so far (ring position)
        long long d_wx = 0, d_ww = 0;
        const uint32_t idesc = make_idesc_f16(128, kNB, 0, 0, 1, 0);      // fp16 x fp16, A MN-major, B K-major
        // The issuing warp's own instruction stream limited this kernel (ncu: tensor pipe 21 % active, ~100 cycles per MMA):
        // it rebuilt four descriptors per K step on the uniform datapath.  Now a stage is exactly three K steps -- the hi*hi
        // accumulator of a step is a compile-time column -- and every descriptor is a per-stage base plus an immediate.
        // B: Xp blocks [kb][bb]: b-groups 128 B apart, k-groups 512 B apart
        const uint64_t b_hi0 = make_smem_desc(smem_u32(x_s), 4 * kBlk, kBlk);
        const uint64_t b_lo_off = (uint64_t)((g.xp_chunk / 2) >> 4);
        constexpr uint64_t kAStep = (2 * 16 * kBlk) >> 4, kBStep = (2 * 4 * kBlk) >> 4, kALoOff = (uint64_t)(kWChunkKb * 16 * kBlk) >> 4;
        static_assert(kWChunkKb == 6, "a W stage must hold exactly three K steps");
        for (int item = blockIdx.x; item < nitems; item += gridDim.x, ++it) {
            const int p = (item >> 1) / g.nbc;
            const int ns = __popc(__reduce_or_sync(0xffffffffu, masks[p * g.L + g.L - 1]));
            const long long cx = dbg9 ? lin_clock() : 0;
            mbar_wait(x_full, it & 1);
            if (dbg9) d_wx += lin_clock() - cx;
            tc_fence_after();
            for (int slot = 0; slot < ns; ++slot, ++sc) {
                const uint32_t ring = sc % kAccRing;
                mbar_wait(&slot_empty[ring], ((sc / kAccRing) & 1) ^ 1);
                tc_fence_after();
                const uint32_t dcol = td + ring * kSlotCols;
                uint64_t b_hi = b_hi0;
                for (int c = 0; c < nchunk; ++c) {
                    const long long cw = dbg9 ? lin_clock() : 0;
                    mbar_wait(&w_full[stage], wphase);
                    if (dbg9) d_ww += lin_clock() - cw;
                    tc_fence_after();
                    // A: Wp blocks [kb][cb16]: mn-groups (c) 128 B apart, k-groups 2048 B apart
                    const uint64_t a_hi = make_smem_desc(smem_u32(w_s + (size_t)stage * w_stage_bytes), 16 * kBlk, kBlk);
                    const uint64_t a_lo = a_hi + kALoOff, b_lo = b_hi + b_lo_off;
                    const int nk = min(3, nkstep - c * 3);
                    if (elect_one()) {
#pragma unroll
                        for (int kk = 0; kk < 3; ++kk) {
                            if (kk < nk) {
                                umma_f16(dcol + (uint32_t)(1 + kk) * kNB, a_hi + kk * kAStep, b_hi + kk * kBStep, idesc, c ? 1u : 0u);
                                umma_f16(dcol, a_hi + kk * kAStep, b_lo + kk * kBStep, idesc, (c | kk) ? 1u : 0u);
                                umma_f16(dcol, a_lo + kk * kAStep, b_hi + kk * kBStep, idesc, 1u);
                            }
                        }
                        umma_commit(&w_empty[stage]);
                    }
                    b_hi += 3 * kBStep;
                    __syncwarp();
                    if (++stage == kWStages) { stage = 0; wphase ^= 1; }
                }
                if (elect_one()) umma_commit(&slot_full[ring]);
                __syncwarp();
            }
            if (elect_one()) {
                if (ns > 0) umma_commit(x_empty);
                else mbar_arrive(x_empty);
            }
            __syncwarp();
        }
        if (dbg9 && lane == 0) {
            dbg[blockIdx.x * 8 + 0] = d_wx; dbg[blockIdx.x * 8 + 1] = d_ww; dbg[blockIdx.x * 8 + 2] = lin_clock() - t_entry;
        }
    } else {
        // ================= epilogue: thread = hidden unit c, registers = the 32 batch rows =================
        const int quad = warp & 3;
        uint32_t sc = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int h = item & 1, bc = (item >> 1) % g.nbc, p = (item >> 1) / g.nbc;
            const uint32_t mask = masks[p * g.L + g.L - 1];
            const int ns = __popc(mask);
            const int c = h * 128 + quad * 32 + lane;
            float sum[kNB];
#pragma unroll
            for (int b = 0; b < kNB; ++b) sum[b] = 0.f;
            for (int slot = 0; slot < ns; ++slot, ++sc) {
                const uint32_t ring = sc % kAccRing;
                const long long ce = dbg9 ? lin_clock() : 0;
                mbar_wait(&slot_full[ring], (sc / kAccRing) & 1);
                if (dbg9 && threadIdx.x == 0 && slot == ns - 1) { dbg[blockIdx.x * 8 + 3] = lin_clock() - ce; dbg[blockIdx.x * 8 + 4] = lin_clock() - t_entry; }
                tc_fence_after();
                const int j = __fns(mask, 0, slot + 1);
                const float bias = params[nd.linW_base + (long long)j * nd.lin_stride + nd.linb_off + c];
                const uint32_t ta = td + ((uint32_t)(quad * 32) << 16) + ring * kSlotCols;
                float acc[kNB];
                {   // ((hh0 + hh1) + hh2) + cross, in fp32 registers
                    uint32_t d[kNB];
                    tmem_ld16(ta + kNB, *reinterpret_cast<uint32_t(*)[16]>(&d[0]));
                    tmem_ld16(ta + kNB + 16, *reinterpret_cast<uint32_t(*)[16]>(&d[16]));
                    tmem_ld_wait();
#pragma unroll
                    for (int b = 0; b < kNB; ++b) acc[b] = __uint_as_float(d[b]);
#pragma unroll
                    for (int a = 2; a <= kAccPerSlot; ++a) {
                        const uint32_t col = ta + (uint32_t)(a % kAccPerSlot) * kNB;      // hh1, hh2, then the cross terms (a = 4 -> 0)
                        tmem_ld16(col, *reinterpret_cast<uint32_t(*)[16]>(&d[0]));
                        tmem_ld16(col + 16, *reinterpret_cast<uint32_t(*)[16]>(&d[16]));
                        tmem_ld_wait();
#pragma unroll
                        for (int b = 0; b < kNB; ++b) acc[b] += __uint_as_float(d[b]);
                    }
                }
                tc_fence_before();
                mbar_arrive(&slot_empty[ring]);
#pragma unroll
                for (int b = 0; b < kNB; ++b) {
                    const float pre = fmaf(acc[b], 1.0f / (kWScale * kXScale), bias);
                    const bool on = pre > 0.f;
                    if (on) sum[b] += pre;
                    const int bg = bc * kNB + b;
                    if (bg < B) Rl[(((size_t)p * B + bg) * g.M + j) * g.Hd + c] = on ? 1 : 0;   // 32 consecutive bytes per warp
                }
            }
            const float Mf = nd.out_div;
#pragma unroll
            for (int b = 0; b < kNB; ++b) {
                const int bg = bc * kNB + b;
                if (bg < B) Hout[((size_t)p * B + bg) * g.Hd + c] = __fdividef(sum[b], Mf);   // net / FLAGS.M (game_ac_network.py:194)
            }
            if (dbg9 && threadIdx.x == 0) dbg[blockIdx.x * 8 + 5] = lin_clock() - t_entry;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 5) tmem_dealloc(td, 512);
}


// ---- prep: dh * relu-mask -> Dp --------------------------------------------------------------------------------
// grid = (row groups, slot, path); blocks of inactive slots leave at once.  Same lane mapping as above.
__global__ void __launch_bounds__(256)
lin_dsplit_kernel(const float* __restrict__ dh, const uint8_t* __restrict__ Rl, LinGeom g, const uint32_t* __restrict__ masks,
                  int P, int B, uint8_t* __restrict__ dp) {
    pn_pdl_sync();
    const int p = blockIdx.z, slot = blockIdx.y;
    const uint32_t mask = masks[p * g.L + g.L - 1];
    if (slot >= __popc(mask)) return;
    const int j = __fns(mask, 0, slot + 1);
    const int ncb4 = g.Hd / 32;
    const int i = blockIdx.x * 256 + threadIdx.x;
    const int cbq = i & 3, r8 = (i >> 2) & 7, bb = (i >> 5) & 3;
    const int cb4 = (i >> 7) % ncb4, bc = (i >> 7) / ncb4;
    if (bc >= g.nbc) return;
    const int cb = cb4 * 4 + cbq, b = bc * kNB + bb * 8 + r8;
    float v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = 0.f;
    if (b < B) {
        const size_t t = (size_t)p * B + b;
        const float* src = dh + t * g.Hd + cb * 8;
        const float4 a = *reinterpret_cast<const float4*>(src), c = *reinterpret_cast<const float4*>(src + 4);
        const uint2 mk = *reinterpret_cast<const uint2*>(Rl + (t * g.M + j) * g.Hd + cb * 8);
        v[0] = (mk.x & 0xFFu) ? a.x : 0.f;       v[1] = (mk.x & 0xFF00u) ? a.y : 0.f;
        v[2] = (mk.x & 0xFF0000u) ? a.z : 0.f;   v[3] = (mk.x & 0xFF000000u) ? a.w : 0.f;
        v[4] = (mk.y & 0xFFu) ? c.x : 0.f;       v[5] = (mk.y & 0xFF00u) ? c.y : 0.f;
        v[6] = (mk.y & 0xFF0000u) ? c.z : 0.f;   v[7] = (mk.y & 0xFF000000u) ? c.w : 0.f;
    }
    uint4 hi, lo;
    split_row8(v, hi, lo);
    uint8_t* dst = dp + (size_t)p * g.dp_path + (size_t)bc * g.dp_chunk + (size_t)slot * g.dp_slot +
                   ((size_t)cb * 4 + bb) * kBlk + r8 * 16;
    *reinterpret_cast<uint4*>(dst) = hi;
    *reinterpret_cast<uint4*>(dst + g.dp_slot / 2) = lo;
}

// ---- data gradient: d(flat)[b, k] = sum_j sum_c dpre_j[b, c] W_j[k, c] ------------------------------------------
// item = (path, batch chunk, tile kt of 128 rows k).  A = Wp rows k (K-major over c), B = Dp (K-major over c).
// One accumulator (32 columns) sums over both halves of c and over all active slots; two accumulator stages.
__global__ void __launch_bounds__(192, 1)
lin_dgrad_tc_kernel(const uint8_t* __restrict__ wp, const uint8_t* __restrict__ dp, LinGeom g,
                    const uint32_t* __restrict__ masks, int P, int B, int KT, float* __restrict__ dflat) {
    pn_pdl_sync();
    extern __shared__ __align__(1024) uint8_t smem[];
    const int w_stage_bytes = 2 * 16 * 16 * kBlk;                    // (hi + lo) x 16 kb x 16 cb16 = 64 KB
    uint8_t* w_s = smem;                                             // 2 stages, one per (slot, half)
    uint8_t* d_s = smem + 2 * (size_t)w_stage_bytes;                 // 2 stages of one slot's Dp (32 KB)
    uint64_t* bars = reinterpret_cast<uint64_t*>(d_s + 2 * g.dp_slot);
    uint64_t* w_full = bars;          // [2]
    uint64_t* w_empty = bars + 2;     // [2]
    uint64_t* d_full = bars + 4;      // [2]
    uint64_t* d_empty = bars + 6;     // [2]
    uint64_t* acc_full = bars + 8;    // [2]
    uint64_t* acc_empty = bars + 10;  // [2]
    __shared__ uint32_t tmem_base_s;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int nitems = P * g.nbc * KT;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&w_full[i], 1); mbar_init(&w_empty[i], 1); mbar_init(&d_full[i], 1); mbar_init(&d_empty[i], 1);
            mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128);
        }
        fence_mbar_init();
    }
    if (warp == 5) tmem_alloc(&tmem_base_s, 64);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tmem_base_s;

    if (warp == 4) {
        // ================= producer =================
        uint32_t wc = 0, dc = 0;                                     // W stages / Dp stages filled so far
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int kt = item % KT, bc = (item / KT) % g.nbc, p = item / (KT * g.nbc);
            const uint32_t mask = __reduce_or_sync(0xffffffffu, masks[p * g.L + g.L - 1]);
            const int ns = __popc(mask);
            const int nkb = min(16, g.KB - kt * 16);
            const uint32_t wbytes = (uint32_t)(nkb * 16 * kBlk);
            for (int slot = 0; slot < ns; ++slot) {
                const int j = __fns(mask, 0, slot + 1);
                {
                    const int st = dc & 1;
                    mbar_wait(&d_empty[st], ((dc >> 1) & 1) ^ 1);
                    if (elect_one()) {
                        mbar_expect_tx(&d_full[st], (uint32_t)g.dp_slot);
                        bulk_g2s(d_s + (size_t)st * g.dp_slot,
                                 dp + (size_t)p * g.dp_path + (size_t)bc * g.dp_chunk + (size_t)slot * g.dp_slot,
                                 (uint32_t)g.dp_slot, &d_full[st]);
                    }
                    __syncwarp();
                    ++dc;
                }
                for (int h = 0; h < 2; ++h) {
                    const int st = wc & 1;
                    mbar_wait(&w_empty[st], ((wc >> 1) & 1) ^ 1);
                    if (elect_one()) {
                        const uint8_t* src = wp + (size_t)j * g.wp_module + ((size_t)h * g.KB + kt * 16) * 16 * kBlk;
                        const size_t hl_stride = (size_t)2 * g.KB * 16 * kBlk;
                        uint8_t* dst = w_s + (size_t)st * w_stage_bytes;
                        mbar_expect_tx(&w_full[st], 2 * wbytes);
                        bulk_g2s(dst, src, wbytes, &w_full[st]);
                        bulk_g2s(dst + w_stage_bytes / 2, src + hl_stride, wbytes, &w_full[st]);
                    }
                    __syncwarp();
                    ++wc;
                }
            }
        }
    } else if (warp == 5) {
        // ================= MMA issue =================
        uint32_t wc = 0, dc = 0, it = 0;
        const uint32_t idesc = make_idesc_f16(128, kNB, 1, 1, 0, 0);      // bf16 x bf16, A K-major, B K-major
        for (int item = blockIdx.x; item < nitems; item += gridDim.x, ++it) {
            const int p = item / (KT * g.nbc);
            const int ns = __popc(__reduce_or_sync(0xffffffffu, masks[p * g.L + g.L - 1]));
            const uint32_t acc = it & 1;
            mbar_wait(&acc_empty[acc], ((it >> 1) & 1) ^ 1);
            tc_fence_after();
            const uint32_t dcol = td + acc * kNB;
            uint32_t accum = 0;
            for (int slot = 0; slot < ns; ++slot) {
                const int dst_ = dc & 1;
                mbar_wait(&d_full[dst_], (dc >> 1) & 1);
                const uint32_t ds = smem_u32(d_s + (size_t)dst_ * g.dp_slot);
                for (int h = 0; h < 2; ++h) {
                    const int st = wc & 1;
                    mbar_wait(&w_full[st], (wc >> 1) & 1);
                    tc_fence_after();
                    const uint32_t ws = smem_u32(w_s + (size_t)st * w_stage_bytes);
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk) {
                        // A: rows k in groups of 8 (kb) 2048 B apart, K-groups (cb16) 128 B apart
                        const uint64_t a_hi = make_smem_desc(ws + (uint32_t)(kk * 2 * kBlk), kBlk, 16 * kBlk);
                        const uint64_t a_lo = make_smem_desc(ws + (uint32_t)(w_stage_bytes / 2 + kk * 2 * kBlk), kBlk, 16 * kBlk);
                        // B: Dp blocks [cb][bb]: b-groups 128 B apart, K-groups (cb) 512 B apart
                        const uint64_t b_hi = make_smem_desc(ds + (uint32_t)((h * 16 + kk * 2) * 4 * kBlk), 4 * kBlk, kBlk);
                        const uint64_t b_lo = make_smem_desc(ds + (uint32_t)(g.dp_slot / 2 + (h * 16 + kk * 2) * 4 * kBlk), 4 * kBlk, kBlk);
                        if (elect_one()) {
                            umma_f16(dcol, a_hi, b_hi, idesc, accum | (uint32_t)(kk > 0));
                            umma_f16(dcol, a_hi, b_lo, idesc, 1u);
                            umma_f16(dcol, a_lo, b_hi, idesc, 1u);
                        }
                    }
                    accum = 1;
                    if (elect_one()) umma_commit(&w_empty[st]);
                    __syncwarp();
                    ++wc;
                }
                if (elect_one()) umma_commit(&d_empty[dst_]);
                __syncwarp();
                ++dc;
            }
            if (elect_one()) {
                if (ns > 0) umma_commit(&acc_full[acc]);
                else mbar_arrive(&acc_full[acc]);
            }
            __syncwarp();
        }
    } else {
        // ================= epilogue: thread = row k, registers = the 32 batch rows =================
        const int quad = warp & 3;
        uint32_t it = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x, ++it) {
            const int kt = item % KT, bc = (item / KT) % g.nbc, p = item / (KT * g.nbc);
            const int ns = __popc(masks[p * g.L + g.L - 1]);
            const uint32_t acc = it & 1;
            const int k = kt * 128 + quad * 32 + lane;
            mbar_wait(&acc_full[acc], (it >> 1) & 1);
            tc_fence_after();
            uint32_t d[kNB];
            if (ns > 0) {
                const uint32_t ta = td + ((uint32_t)(quad * 32) << 16) + acc * kNB;
                tmem_ld16(ta, *reinterpret_cast<uint32_t(*)[16]>(&d[0]));
                tmem_ld16(ta + 16, *reinterpret_cast<uint32_t(*)[16]>(&d[16]));
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int b = 0; b < kNB; ++b) d[b] = 0u;
            }
            tc_fence_before();
            mbar_arrive(&acc_empty[acc]);
            if (k < g.K) {
#pragma unroll
                for (int b = 0; b < kNB; ++b) {
                    const int bg = bc * kNB + b;
                    if (bg < B) dflat[((size_t)p * B + bg) * g.K + k] = __uint_as_float(d[b]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 5) tmem_dealloc(td, 64);
}

// ---- weight gradient: dW_j[k, c] = sum over the module's paths and their rows b of X[b, k] dpre_j[b, c] ----------
// item = (module j, tile kt of 128 rows k, half h of the columns c).  One stage = one (path, batch chunk):
// A = Xp tile (MN-major, mn = k), B = Dp half (MN-major, mn = c); 2 K-steps of 16 rows, 3 MMAs each.
// Row K (the "ones" row of Xp) of the last tile is the bias gradient.
__global__ void __launch_bounds__(192, 1)
lin_wgrad_tc_kernel(const uint8_t* __restrict__ xp, const uint8_t* __restrict__ dp, PnNetDesc nd, LinGeom g,
                    const uint32_t* __restrict__ masks, const int* __restrict__ list, const int* __restrict__ list_count,
                    int P, int KT, float* __restrict__ grads) {
    pn_pdl_sync();
    extern __shared__ __align__(1024) uint8_t smem[];
    constexpr int kStages = 4;
    const int a_bytes = 16 * 4 * kBlk, b_bytes = 16 * 4 * kBlk;      // per hl: 16 kb (or cb) x 4 bb blocks = 8 KB
    const int stage_bytes = 2 * a_bytes + 2 * b_bytes;               // A hi, A lo, B hi, B lo = 32 KB
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)kStages * stage_bytes);
    uint64_t* full = bars;                 // [kStages]
    uint64_t* empty = bars + kStages;      // [kStages]
    uint64_t* acc_full = bars + 2 * kStages;       // [2]
    uint64_t* acc_empty = acc_full + 2;            // [2]
    __shared__ uint32_t tmem_base_s;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int nitems = g.M * KT * 2;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
        fence_mbar_init();
    }
    if (warp == 5) tmem_alloc(&tmem_base_s, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tmem_base_s;

    if (warp == 4) {
        // ================= producer =================
        uint32_t sc = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int h = item & 1, kt = (item >> 1) % KT, j = (item >> 1) / KT;
            const int lm = (g.L - 1) * g.M + j;
            const int cnt = __reduce_max_sync(0xffffffffu, list_count[lm]);
            const int nkb = min(16, g.KB - kt * 16);
            const uint32_t abytes = (uint32_t)(nkb * 4 * kBlk);
            // The module's path list and each path's slot of this module are fetched 32 paths at a time (lane i holds path
            // i0 + i): one list load + one mask load per path, done serially inside the loop, used to cost ~1 us per path and
            // left the four-stage ring empty -- the producer, not the tensor pipe, set this kernel's pace.
            int pv = 0, sv = 0;
            for (int i = 0; i < cnt; ++i) {
                if ((i & 31) == 0) {
                    const int ii = i + lane;
                    pv = ii < cnt ? list[(size_t)lm * P + ii] : 0;
                    const uint32_t mk = ii < cnt ? masks[pv * g.L + g.L - 1] : 0u;
                    sv = __popc(mk & ((1u << j) - 1u));
                }
                const int p = __shfl_sync(0xffffffffu, pv, i & 31);
                const int slot = __shfl_sync(0xffffffffu, sv, i & 31);
                for (int bc = 0; bc < g.nbc; ++bc) {
                    const int st = sc % kStages;
                    mbar_wait(&empty[st], ((sc / kStages) & 1) ^ 1);
                    if (elect_one()) {
                        uint8_t* dst = smem + (size_t)st * stage_bytes;
                        const uint8_t* xa = xp + (size_t)p * g.xp_path + (size_t)bc * g.xp_chunk + (size_t)kt * 16 * 4 * kBlk;
                        const uint8_t* db = dp + (size_t)p * g.dp_path + (size_t)bc * g.dp_chunk + (size_t)slot * g.dp_slot +
                                            (size_t)h * 16 * 4 * kBlk;
                        mbar_expect_tx(&full[st], 2 * abytes + 2 * (uint32_t)b_bytes);
                        bulk_g2s(dst, xa, abytes, &full[st]);
                        bulk_g2s(dst + a_bytes, xa + g.xp_chunk / 2, abytes, &full[st]);
                        bulk_g2s(dst + 2 * a_bytes, db, (uint32_t)b_bytes, &full[st]);
                        bulk_g2s(dst + 2 * a_bytes + b_bytes, db + g.dp_slot / 2, (uint32_t)b_bytes, &full[st]);
                    }
                    __syncwarp();
                    ++sc;
                }
            }
        }
    } else if (warp == 5) {
        // ================= MMA issue =================
        uint32_t sc = 0, it = 0;
        const uint32_t idesc = make_idesc_f16(128, 128, 1, 1, 1, 1);      // bf16 x bf16, A MN-major, B MN-major
        for (int item = blockIdx.x; item < nitems; item += gridDim.x, ++it) {
            const int j = (item >> 1) / KT;
            const int cnt = __reduce_max_sync(0xffffffffu, list_count[(g.L - 1) * g.M + j]);
            const uint32_t acc = it & 1;
            mbar_wait(&acc_empty[acc], ((it >> 1) & 1) ^ 1);
            tc_fence_after();
            const uint32_t dcol = td + acc * 128;
            uint32_t accum = 0;
            for (int i = 0; i < cnt * g.nbc; ++i) {
                const int st = sc % kStages;
                mbar_wait(&full[st], (sc / kStages) & 1);
                tc_fence_after();
                const uint32_t base = smem_u32(smem + (size_t)st * stage_bytes);
#pragma unroll
                for (int ks = 0; ks < 2; ++ks) {
                    // A: mn-groups (kb) 512 B apart, K-groups (bb) 128 B apart;  B: mn-groups (cb) 512 B apart
                    const uint64_t a_hi = make_smem_desc(base + (uint32_t)(ks * 2 * kBlk), kBlk, 4 * kBlk);
                    const uint64_t a_lo = make_smem_desc(base + (uint32_t)(a_bytes + ks * 2 * kBlk), kBlk, 4 * kBlk);
                    const uint64_t b_hi = make_smem_desc(base + (uint32_t)(2 * a_bytes + ks * 2 * kBlk), kBlk, 4 * kBlk);
                    const uint64_t b_lo = make_smem_desc(base + (uint32_t)(2 * a_bytes + b_bytes + ks * 2 * kBlk), kBlk, 4 * kBlk);
                    if (elect_one()) {
                        umma_f16(dcol, a_hi, b_hi, idesc, accum | (uint32_t)ks);
                        umma_f16(dcol, a_hi, b_lo, idesc, 1u);
                        umma_f16(dcol, a_lo, b_hi, idesc, 1u);
                    }
                }
                accum = 1;
                if (elect_one()) umma_commit(&empty[st]);
                __syncwarp();
                ++sc;
            }
            if (elect_one()) {
                if (cnt > 0) umma_commit(&acc_full[acc]);
                else mbar_arrive(&acc_full[acc]);
            }
            __syncwarp();
        }
    } else {
        // ================= epilogue: thread = row k of the tile, 128 columns c =================
        const int quad = warp & 3;
        uint32_t it = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x, ++it) {
            const int h = item & 1, kt = (item >> 1) % KT, j = (item >> 1) / KT;
            const int cnt = list_count[(g.L - 1) * g.M + j];
            const uint32_t acc = it & 1;
            const int k = kt * 128 + quad * 32 + lane;
            float* G = grads + nd.linW_base + (long long)j * nd.lin_stride;
            float* dst = k < g.K ? G + (size_t)k * g.Hd + h * 128 : (k == g.K ? G + nd.linb_off + h * 128 : nullptr);
            mbar_wait(&acc_full[acc], (it >> 1) & 1);
            tc_fence_after();
#pragma unroll
            for (int c0 = 0; c0 < 128; c0 += 16) {
                uint32_t d[16];
                if (cnt > 0) {
                    tmem_ld16(td + ((uint32_t)(quad * 32) << 16) + acc * 128 + c0, d);
                    tmem_ld_wait();
                } else {
#pragma unroll
                    for (int q = 0; q < 16; ++q) d[q] = 0u;
                }
                if (dst) {
#pragma unroll
                    for (int q = 0; q < 16; q += 4)
                        *reinterpret_cast<float4*>(dst + c0 + q) = make_float4(__uint_as_float(d[q]), __uint_as_float(d[q + 1]),
                                                                               __uint_as_float(d[q + 2]), __uint_as_float(d[q + 3]));
                }
            }
            tc_fence_before();
            mbar_arrive(&acc_empty[acc]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 5) tmem_dealloc(td, 256);
}

bool lin_geom(const pn_handle* h, int B, LinGeom& g) {
    const PnNetDesc& nd = h->nd;
    if (nd.hidden != 256 || nd.last_lin % 8 || nd.last_lin < 48 || nd.M > 16) return false;   // (>= 3 K steps: every hi*hi accumulator is written)
    g.K = nd.last_lin; g.KB = (g.K / 8 + 2) & ~1; g.Hd = nd.hidden; g.M = nd.M; g.L = nd.L;   // >= K/8 + 1, even
    g.nbc = (B + kNB - 1) / kNB;
    g.wp_module = (size_t)2 * 2 * g.KB * 16 * kBlk;
    g.xp_chunk = (size_t)2 * g.KB * 4 * kBlk;
    g.xp_path = g.xp_chunk * g.nbc;
    g.dp_slot = (size_t)2 * 32 * 4 * kBlk;
    g.dp_chunk = g.dp_slot * g.M;
    g.dp_path = g.dp_chunk * g.nbc;
    return true;
}

}  // namespace

bool pn_lin_tc_supported(const pn_handle* h) {
    LinGeom g;
    if (!lin_geom(h, h->Bmax, g)) return false;
    const size_t smem = (size_t)kWStages * 2 * kWChunkKb * 16 * kBlk + g.xp_chunk + 256;
    return smem <= 220 * 1024;
}

size_t pn_lin_tc_scratch_bytes(const pn_handle* h) {
    LinGeom g;
    if (!lin_geom(h, h->Bmax, g)) return 0;
    // Wp | Xp | Dp (bf16, backward)  |  Wph | Xph (fp16, forward)
    return 2 * (g.wp_module * g.M + g.xp_path * h->P) + g.dp_path * h->P;
}

// W_lin -> Wp: depends on the parameters only, so the forward launcher runs it on the side stream
int pn_launch_lin_wsplit(pn_handle* h, cudaStream_t st) {
    LinGeom g;
    if (!lin_geom(h, h->Bmax, g)) return pn_fail(h, PN_ERR_STATE, "tensor-core linear path not supported");
    {
        PnProfScope prof(h, PN_PROF_LIN_WPREP, st);
        const long long wthreads = (long long)g.M * g.KB * (g.Hd / 32) * 32;
        const int wblocks = (int)((wthreads + 255) / 256);
        uint8_t* wph = h->tc_lin_buf + g.wp_module * g.M + g.xp_path * h->P + g.dp_path * h->P;
        lin_prep_kernel<<<wblocks, 256, 0, st>>>(h->params, h->nd, g, h->tc_lin_buf, wph, wblocks, nullptr, h->P, 1, nullptr, nullptr);
    }
    PN_LAUNCH_CHECK(h, "lin_prep_kernel");
    return PN_OK;
}

int pn_launch_lin_fwd_tc(pn_handle* h, int B, cudaStream_t st) {
    LinGeom g;
    if (!lin_geom(h, B, g)) return pn_fail(h, PN_ERR_STATE, "tensor-core linear path not supported");
    LinGeom gmax;
    lin_geom(h, h->Bmax, gmax);
    const PnNetDesc& nd = h->nd;
    uint8_t* wp = h->tc_lin_buf;
    uint8_t* xp = wp + gmax.wp_module * g.M;
    uint8_t* wph = xp + gmax.xp_path * h->P + gmax.dp_path * h->P;
    uint8_t* xph = wph + gmax.wp_module * g.M;
    {
        PnProfScope prof(h, PN_PROF_LIN_PREP, st);
        const long long xthreads = (long long)h->P * g.nbc * ((g.KB + 3) / 4) * 128;
        const int xblocks = (int)((xthreads + 255) / 256);
        pn_launch(pn_use_pdl(h), lin_prep_kernel, dim3(xblocks), dim3(256), 0, st, (const float*)h->params, nd, g, wp, wph, 0,
                  (const float*)h->act[nd.nconv - 1], h->P, B, xp, xph);
    }
    PN_LAUNCH_CHECK(h, "lin_prep_kernel");
    const size_t smem = (size_t)kWStages * 2 * kWChunkKb * 16 * kBlk + g.xp_chunk + 256;
    PN_SET_SMEM(h, lin_fwd_tc_kernel, smem);
    const int items = h->P * g.nbc * 2;
    const int grid = items < h->num_sms ? items : h->num_sms;
    {
        PnProfScope prof(h, PN_PROF_LIN_FWD, st);
        pn_launch(pn_use_pdl(h), lin_fwd_tc_kernel, dim3(grid), dim3(192), smem, st, (const uint8_t*)wph, (const uint8_t*)xph,
                  (const float*)h->params, nd, g, (const uint32_t*)h->d_masks, h->P, B, h->hbuf, h->lmask, h->tc_dbg);
    }
    PN_LAUNCH_CHECK(h, "lin_fwd_tc_kernel");
    return PN_OK;
}

int pn_launch_lin_bwd_tc(pn_handle* h, int B, cudaStream_t st) {
    LinGeom g, gmax;
    if (!lin_geom(h, B, g) || !lin_geom(h, h->Bmax, gmax)) return pn_fail(h, PN_ERR_STATE, "tensor-core linear path not supported");
    const PnNetDesc& nd = h->nd;
    uint8_t* wp = h->tc_lin_buf;
    uint8_t* xp = wp + gmax.wp_module * g.M;
    uint8_t* dp = xp + gmax.xp_path * h->P;
    {
        PnProfScope prof(h, PN_PROF_LIN_PREP, st);
        dim3 dgrid((g.nbc * (g.Hd / 32) * 128 + 255) / 256, g.M, h->P);
        pn_launch(pn_use_pdl(h), lin_dsplit_kernel, dgrid, dim3(256), 0, st, (const float*)h->dh, (const uint8_t*)h->lmask, g,
                  (const uint32_t*)h->d_masks, h->P, B, dp);
    }
    PN_LAUNCH_CHECK(h, "lin_dsplit_kernel");
    const int KT = (g.K / 8 + 1 + 15) / 16;          // tiles of 16 k-blocks covering the K rows plus the ones row
    // weight gradient first: its result (97 % of the gradient bytes) is what a multi-GPU host wants to start summing,
    // and the data-gradient GEMM that follows gives that all-reduce SMs to share (one 192-thread CTA per SM)
    {
        const size_t smem = 4 * (size_t)(4 * 16 * 4 * kBlk) + 256;
        PN_SET_SMEM(h, lin_wgrad_tc_kernel, smem);
        const int items = g.M * KT * 2;
        const int grid = items < h->num_sms ? items : h->num_sms;
        PnProfScope prof(h, PN_PROF_LIN_WGRAD, st);
        pn_launch(pn_use_pdl(h), lin_wgrad_tc_kernel, dim3(grid), dim3(192), smem, st, (const uint8_t*)xp, (const uint8_t*)dp, nd, g,
                  (const uint32_t*)h->d_masks, (const int*)h->d_list, (const int*)h->d_list_count, h->P, KT, h->grads);
    }
    PN_LAUNCH_CHECK(h, "lin_wgrad_tc_kernel");
    PN_CUDA(h, cudaEventRecord(h->ev_lin_grads, st));     // W_lin / b_lin gradients are final from here on
    {
        const size_t smem = 2 * (size_t)(2 * 16 * 16 * kBlk) + 2 * g.dp_slot + 256;
        PN_SET_SMEM(h, lin_dgrad_tc_kernel, smem);
        const int items = h->P * g.nbc * KT;
        const int grid = items < h->num_sms ? items : h->num_sms;
        PnProfScope prof(h, PN_PROF_LIN_DGRAD, st);
        pn_launch(pn_use_pdl(h), lin_dgrad_tc_kernel, dim3(grid), dim3(192), smem, st, (const uint8_t*)wp, (const uint8_t*)dp, g,
                  (const uint32_t*)h->d_masks, h->P, B, KT, h->dact[nd.nconv - 1]);
    }
    PN_LAUNCH_CHECK(h, "lin_dgrad_tc_kernel");
    return PN_OK;
}
