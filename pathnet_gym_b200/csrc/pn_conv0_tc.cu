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
//    copy is 16-byte cp.async straight from the u8 frame: no im2col, no conversion, every frame byte
//    is read from HBM once.
//  * the A operand of tap (dy,dx) is that SAME buffer read through a shifted descriptor: row m of the
//    MMA (block index m = oy*21+ox) starts at byte (m + dy*21 + dx)*16 of each region (K-major, no
//    swizzle: 8 blocks x 16 B core matrices, SBO = 128 B, LBO = region size).  Rows with ox == 20 or
//    oy == 20 are computed and discarded (41 of 441).
//  * frames are exact u8 integers; the fp32 weights are split per (module, channel) into three signed
//    base-256 digits of round(w / max|w_c| * 2^22) (conv0_wprep_kernel, run every call since RMSProp
//    changed the weights).  Integer MMAs are exact, so the only deviation from fp32 arithmetic is the
//    2^-23 (relative to the channel's largest weight) rounding of that split.  B = [digit][channel]
//    rows per active module (24 columns per module), gathered per path from the genotype bitmask.
//  * epilogue (thread = output pixel = TMEM lane): pre = D2*(65536 s_c) + D1*(256 s_c) + D0*s_c + b_c
//    in fp32 (the integer sums convert exactly), ReLU, ascending-j sum, ReLU bit mask for backward.
//
// Warp roles (320 threads, one CTA per SM, persistent over a contiguous range of transitions):
//   warps 0-15  four epilogue groups of four warps (one TMEM lane quadrant each); consecutive 128-row
//   tiles rotate over the groups, each group owning one TMEM accumulator stage,
//   warp 16  TMA producer (frames -> smem ring), warp 17  TMEM owner + B gather + MMA issue.
#include "pn_internal.cuh"
#include "pn_tc.cuh"

#include <cuda.h>

#include <algorithm>
#include <vector>

using namespace pntc;

namespace {

constexpr int kEpiGroups = 4;            // forward kernel: 4 groups of 4 epilogue warps, then producer and MMA warps
constexpr int kFwdThreads = (4 * kEpiGroups + 3) * 32;    // epilogue groups, producer, MMA warp A, MMA warp B
// One TMEM accumulator stage per epilogue group: tile c -> group c%4 -> stage c%4.  (A stage must not be shared
// between groups: mbarrier parity waits only distinguish the current from the preceding phase, so a group that
// ran ahead would take another group's older phase for its own.)
constexpr int kAccStages = kEpiGroups;
constexpr int kAccCols = 128;            // TMEM columns per accumulator stage (max N of one pass = 96)
constexpr int kSlotBytes = 3 * 2048;     // one module's B rows: 3 digit groups x (8 channels x 256 K bytes)
constexpr int kPassSlots = 4;            // modules per MMA pass
constexpr int kQMax = 1 << 22;

// Work split over the CTAs.  A transition's cost depends on how many modules its path has active in layer 0 (and jumps when
// a second MMA pass is needed), so equal COUNTS per CTA leave the kernel waiting for the CTA that drew the expensive
// paths: with 10 % of the paths at five modules the forward kernel took 241 us against 130 us for the same work spread
// evenly.  The host cuts the path-major sequence of work units at equal cumulative cost instead (cost table measured on
// B200 with uniform populations, tools/layer0_cost.py); CTA c owns units [b[c], b[c+1]).
struct TcBounds { int b[160]; };
constexpr int kMaxCtas = 159;

struct TcGeom {
    int H, W, C, s;                  // frame, stride (k = 2s, s*C = 16)
    int Hs, Ws, OH, OW;              // block grid and output grid
    int tiles;                       // 128-row MMA tiles per transition
    int region_bytes, stage_bytes;   // smem bytes of one region / one frame stage (s regions)
    int row_bytes, frame_bytes;      // W*C, H*W*C
    int nst;                         // frame stages in the ring
    int ksteps;                      // MMAs (K = 32) per tap = s*16/32
    int mq;                          // rows of the weight-gradient reduction (block raster, multiple of 32)
    int rcp_ws;                      // ceil(65536 / Ws): m / Ws == (m * rcp_ws) >> 16 for every row index used
};

// ---- weight split: fp32 -> 3 signed base-256 digits per (module, channel) ---------------------------
// w8 [M][kSlotBytes] in the canonical K-major layout: element (n = digit*8 + c, kidx) at
// digit*2048 + (kidx/16)*128 + c*16 + kidx%16 with kidx = tap*64 + g*16 + b,
// (kh, kw, ci) = (dy*s + g, dx*s + b/C, b%C), tap = dy*2 + dx.   wscale [M][8] = max|w_c| / (2^22 * 255).
__global__ void __launch_bounds__(256)
conv0_wprep_kernel(const float* __restrict__ params, PnNetDesc nd, int s, int8_t* __restrict__ w8,
                   float* __restrict__ wscale) {
    pn_pdl_sync();                // the weights may have been written by the kernel just before (the RMSProp update)
    __shared__ float red[8][8];
    __shared__ float mx[8];
    const int j = blockIdx.x, KK = nd.KK[0], C = nd.Cin[0], k = nd.k[0];
    const float* W = params + nd.convW_base[0] + (long long)j * nd.conv_stride[0];
    const int kk = threadIdx.x;                 // KK == 4*s*s*C == 256 for the supported geometry
    float w[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) w[c] = kk < KK ? W[kk * 8 + c] : 0.f;
    // broadcast through a shuffle so that the compiler knows the warp index is warp-uniform (role branches
    // then count as uniform control flow and the MMA operands can live in uniform registers)
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
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

// Stage one frame with ONE 4-D TMA load.  The tensor map views the u8 frames as
// [t][g = raw row % s][Y = raw row / s][W*C/4 u32] (strides 336*s, 336, H*W*C bytes for 84x84x4), so a box
// {W*C/4, Hs, s, 1} lands in shared memory as s regions of Hs*Ws 16-byte blocks in raster order.
// (One cp.async.bulk per 336-byte raw row was measured at ~100 cycles of issue per copy -- 84 copies per
// frame held the tensor pipe at 16 % -- and a single warp of 16-byte cp.async was slower still.)
__device__ __forceinline__ void stage_frame_tma(uint8_t* dst, const CUtensorMap* tmap, int t, const TcGeom& gm,
                                                uint64_t* bar) {
    mbar_expect_tx(bar, (uint32_t)gm.frame_bytes);
    tma_load_4d(dst, tmap, 0, 0, 0, t, bar);
}

// ---- the convolution -----------------------------------------------------------------------------------
// clock read that the compiler cannot move across memory operations / other volatile asm (debug brackets)
__device__ __forceinline__ long long pinned_clock() {
    long long t;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory");
    return t;
}

// kFast: a frame is exactly one 128-row tile per accumulator stage and a tap is two K steps (84x84x4 frames, 8x8 / 4 filters);
// the generic variant handles every other supported geometry.  Two instantiations instead of one kernel with both code
// paths: the issuing warp keeps its descriptors in uniform registers, and with all variants in one body some of them
// spilled to vector registers (an R2UR chain in front of every tile's MMAs: +8 % kernel time).
template <bool kFast>
__global__ void __launch_bounds__(kFwdThreads, 1)
conv0_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmap, const int8_t* __restrict__ w8,
                    const float* __restrict__ wscale, const float* __restrict__ params, PnNetDesc nd,
                    const uint32_t* __restrict__ masks, int B, int T, TcGeom gm, const __grid_constant__ TcBounds bounds,
                    float* __restrict__ Y, uint8_t* __restrict__ R, long long* __restrict__ dbg_arg) {
    long long* const dbg = kTcDebugBuild ? dbg_arg : nullptr;   // the instrumentation below compiles away in the product build
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* frame_s = smem;
    uint8_t* Bs = smem + (size_t)gm.nst * gm.stage_bytes;
    float* sbuf = reinterpret_cast<float*>(Bs + (size_t)(nd.M + 1) * kSlotBytes);   // [M][32] s0[8] s1[8] s2[8] bias[8]
    int* jtab = reinterpret_cast<int*>(sbuf + nd.M * 32);                            // [M] module id of each slot
    uint64_t* bars = reinterpret_cast<uint64_t*>(jtab + ((nd.M + 1) & ~1));
    uint64_t* frame_full = bars;
    uint64_t* frame_empty = bars + gm.nst;
    uint64_t* acc_full = bars + 2 * gm.nst;
    uint64_t* acc_empty = acc_full + kAccStages;
    uint64_t* path_ready = acc_empty + kAccStages;
    uint64_t* drain = path_ready + 1;
    uint64_t* b_full = drain + 1;
    uint64_t* epi_idle = b_full + 1;      // all epilogue threads have finished the previous path (they still read
                                          // its scales / module ids after releasing the accumulator stage)
    __shared__ uint32_t tmem_base_s;

    // broadcast through a shuffle so that the compiler knows the warp index is warp-uniform (role branches
    // then count as uniform control flow and the MMA operands can live in uniform registers)
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int t_lo = bounds.b[blockIdx.x], t_hi = bounds.b[blockIdx.x + 1];     // equal cumulative cost, not equal counts
    const int npix = gm.OH * gm.OW;
    const bool dbg_detail = dbg && dbg[gridDim.x * 8] == 1;
    const long long t_entry = dbg ? clock64() : 0;

    if (threadIdx.x == 0) {
        // kFast: two warps issue MMAs (tiles 0, 2 and tiles 1, 3 of every frame), so a frame stage is free after two commits
        for (int i = 0; i < gm.nst; ++i) { mbar_init(&frame_full[i], 1); mbar_init(&frame_empty[i], kFast ? 2 : 1); }
        // (one arrival per thread: electing one lane per warp after a __syncwarp measured 7 % SLOWER here)
        for (int i = 0; i < kAccStages; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
        mbar_init(path_ready, 1);
        mbar_init(drain, 1);
        mbar_init(b_full, 1);
        mbar_init(epi_idle, 128 * kEpiGroups);
        fence_mbar_init();
    }
    // Warp ids matter: within an SM sub-partition the scheduler prefers the highest warp id, so the two
    // single-thread roles that everything else waits for (TMA producer = warp 16, MMA issue = warp 17) sit
    // above the sixteen epilogue warps.  (With the MMA warp at id 1 it was measured at ~205 cycles per MMA
    // against 52 in isolation: it lost arbitration to the epilogue warps of its sub-partition.)
    if (warp == 4 * kEpiGroups + 1) tmem_alloc(&tmem_base_s, kAccStages * kAccCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tmem_base_s;
    pn_pdl_sync();                // everything above touches this CTA's shared memory and TMEM only

    if (warp == 4 * kEpiGroups) {
        // ================= producer: one TMA load per frame =================
        if (lane == 0) tma_prefetch_desc(&tmap);
        int stage = 0;
        uint32_t phase = 0;
        long long w_empty = 0, t_begin = clock64();
        for (int t = t_lo; t < t_hi; ++t) {
            const long long c0 = clock64();
            mbar_wait(&frame_empty[stage], phase ^ 1);
            w_empty += clock64() - c0;
            if (lane == 0) stage_frame_tma(frame_s + (size_t)stage * gm.stage_bytes, &tmap, t, gm, &frame_full[stage]);
            __syncwarp();
            if (++stage == gm.nst) { stage = 0; phase ^= 1; }
        }
        if (dbg && lane == 0 && dbg[gridDim.x * 8] == 4) { dbg[blockIdx.x * 8 + 0] = w_empty; dbg[blockIdx.x * 8 + 1] = clock64() - t_begin; }
    } else if (warp == 4 * kEpiGroups + 1) {
        // ================= B gather + MMA issue =================
        // Everything here is executed by the whole (converged) warp on warp-uniform values; only the
        // tcgen05 / mbarrier-arrive instructions themselves are elected to one lane (see elect_one()).
        int stage = 0, cur_path = -1, ns = 0;
        uint32_t fphase = 0, drain_phase = 0, b_phase = 0, idle_phase = 0;
        uint32_t tilec = 0, fill_par = 0;            // tiles issued; bit a = parity of the fills of accumulator stage a
        const long long t_begin = clock64();
        long long w_frame = 0, w_acc = 0, w_path = 0, w_issue = 0, w_fence = 0, w_commit = 0, w_desc = 0, w_fend = 0;
        const bool dbg_nomma = dbg && __reduce_or_sync(0xffffffffu, (unsigned)(dbg[gridDim.x * 8] == 6));   // experiment: no MMAs at all
        const bool dbg_twice = dbg && __reduce_or_sync(0xffffffffu, (unsigned)(dbg[gridDim.x * 8] == 3));   // experiment: every MMA twice
        // (path index by counting: an integer division per frame is ~25 dependent instructions on these roles' critical paths)
        for (int t = t_lo, p = t_lo / B, p_next = (t_lo / B + 1) * B; t < t_hi; ++t) {
            if (t == p_next) { ++p; p_next += B; }
            const long long cp0 = dbg_detail ? clock64() : 0;
            if (p != cur_path) {
                if (cur_path >= 0) {
                    // every MMA that reads the old B has completed, every accumulator has been drained
                    if (elect_one()) umma_commit(drain);
                    __syncwarp();
                    mbar_wait(drain, drain_phase);
                    drain_phase ^= 1;
                    for (uint32_t a = 0; a < kAccStages; ++a) mbar_wait(&acc_empty[a], ((fill_par >> a) & 1u) ^ 1u);
                    mbar_wait(epi_idle, idle_phase);
                    idle_phase ^= 1;
                }
                // the OR-reduction hands the compiler a value it knows to be warp-uniform
                const uint32_t mask = __reduce_or_sync(0xffffffffu, masks[p * nd.L]);
                ns = __popc(mask);
                if (ns > 0 && elect_one()) mbar_expect_tx(b_full, (uint32_t)(ns * kSlotBytes));
                __syncwarp();
                for (int slot = 0; slot < ns; ++slot) {
                    const int j = __fns(mask, 0, slot + 1);
                    // the module's digit rows arrive by one bulk copy (async proxy: no staging registers)
                    if (lane == 0) bulk_g2s(Bs + (size_t)slot * kSlotBytes, w8 + (size_t)j * kSlotBytes, kSlotBytes, b_full);
                    // per slot: the 8 channel scales, then the 8 biases (the digit weights 256 and 65536 are immediates in
                    // the epilogue: every shared-memory wavefront the epilogue spends competes with the tensor core's
                    // operand fetch for the same port, and the scale table used to be half of the epilogue's loads)
                    if (lane < 8) sbuf[slot * 32 + lane] = wscale[j * 8 + lane];
                    else if (lane < 16)
                        sbuf[slot * 32 + lane] = params[nd.convW_base[0] + (long long)j * nd.conv_stride[0] + nd.convb_off[0] + lane - 8];
                    if (lane == 0) jtab[slot] = j;
                }
                if (ns > 0) { mbar_wait(b_full, b_phase); b_phase ^= 1; }
                __syncwarp();
                if (elect_one()) mbar_arrive(path_ready);
                cur_path = p;
            }
            if (dbg_detail) w_path += clock64() - cp0;
            const long long cf0 = dbg_detail ? clock64() : 0;
            mbar_wait(&frame_full[stage], fphase);
            if (dbg_detail) w_frame += clock64() - cf0;
            const long long cd0 = dbg_detail ? clock64() : 0;
            tc_fence_after();
            const int npass = (ns + kPassSlots - 1) / kPassSlots;
            const uint32_t fbase = smem_u32(frame_s + (size_t)stage * gm.stage_bytes);
            const uint32_t bsm = smem_u32(Bs);
            // low descriptor words of tile 0 / pass 0 for the 8 (tap, K-step) MMAs; a tile advances A by 128
            // blocks, a pass advances B by 4 modules (units of 16 bytes) -- one uniform add per MMA
            uint32_t alo[8], blo[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int tap = i >> 1, ks = i & 1;                          // 2 K-steps of 32 bytes per tap (s = 4)
                alo[i] = desc_lo(fbase + (uint32_t)(2 * ks * gm.region_bytes + ((tap >> 1) * gm.Ws + (tap & 1)) * 16),
                                 (uint32_t)gm.region_bytes);
                blo[i] = desc_lo(bsm + (uint32_t)((tap * gm.s + 2 * ks) * 128), 128);
            }
            const uint32_t ahi = desc_hi(128), bhi = desc_hi(2048);
            const int nmma = 4 * gm.ksteps;
            if (dbg_detail) w_desc += clock64() - cd0;
            if constexpr (kFast) {
                // Common geometry (84x84x4 frames): a frame is exactly one tile per accumulator stage.  This warp's scalar
                // instruction stream is the critical path of the whole kernel (one warp, dependent uniform-datapath ops at
                // ~7 cycles each: 120 instructions per tile were measured at ~870 cycles against 424 cycles of tensor time),
                // so everything is hoisted: the 16 descriptors are built once per frame as 64-bit values and a tile costs one
                // add per MMA.  Paths with more than four active modules (mutation adds modules, task 2 ORs the frozen
                // path in) take further passes over the same frame, pass-major: by the time stage t is filled again its
                // epilogue group has drained the previous pass.
                uint64_t a64[8], b64[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) { a64[i] = desc_join(alo[i], ahi); b64[i] = desc_join(blo[i], bhi); }
                if (npass == 1) {
                    // (kept apart from the multi-pass loop below: with the pass loop around it this code measured 6-10 %
                    // slower for the common case -- the single issuing warp's instruction stream is that critical)
                    const int N = (24 * ns + 15) & ~15;
                    const uint32_t idesc = make_idesc_i8(128, N, 0, 1, 0, 0);
#pragma unroll
                    for (int tile = 0; tile < kAccStages; tile += 2) {       // (tiles 1 and 3: warp B)
                        mbar_wait(&acc_empty[tile], ((fill_par >> tile) & 1u) ^ 1u);
                        tc_fence_after();
                        const uint32_t dcol = td + tile * kAccCols;
                        if (!dbg_nomma && elect_one()) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) umma_i8(dcol, a64[i] + (uint64_t)(tile * 128), b64[i], idesc, i ? 1u : 0u);
                            umma_commit(&acc_full[tile]);
                        } else if (dbg_nomma && elect_one()) {
                            umma_commit(&acc_full[tile]);
                        }
                        __syncwarp();
                    }
                    fill_par ^= (1u << kAccStages) - 1u;
                } else
                for (int pass = 0; pass < npass; ++pass) {
                    const int nsl = min(kPassSlots, ns - pass * kPassSlots);
                    const int N = (24 * nsl + 15) & ~15;
                    const uint32_t idesc = make_idesc_i8(128, N, 0, 1, 0, 0);
                    const uint64_t b_adv = (uint64_t)((pass * kPassSlots * kSlotBytes) >> 4);
#pragma unroll
                    for (int tile = 0; tile < kAccStages; tile += 2) {       // (tiles 1 and 3: warp B)
                        mbar_wait(&acc_empty[tile], ((fill_par >> tile) & 1u) ^ 1u);
                        tc_fence_after();
                        const uint32_t dcol = td + tile * kAccCols;
                        if (elect_one()) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) umma_i8(dcol, a64[i] + (uint64_t)(tile * 128), b64[i] + b_adv, idesc, i ? 1u : 0u);
                            umma_commit(&acc_full[tile]);
                        }
                        __syncwarp();
                    }
                    fill_par ^= (1u << kAccStages) - 1u;
                }
                tilec += kAccStages;
            } else
            for (int tile = 0; tile < gm.tiles && npass > 0; ++tile, ++tilec) {
                const uint32_t acc = tilec & (kAccStages - 1);
                for (int pass = 0; pass < npass; ++pass) {
                    const long long ca0 = dbg_detail ? clock64() : 0;
                    mbar_wait(&acc_empty[acc], ((fill_par >> acc) & 1u) ^ 1u);
                    if (dbg_detail) w_acc += clock64() - ca0;
                    const long long cq0 = dbg_detail ? clock64() : 0;
                    tc_fence_after();
                    if (dbg_detail) w_fence += clock64() - cq0;
                    const int nsl = min(kPassSlots, ns - pass * kPassSlots);
                    const int N = (24 * nsl + 15) & ~15;
                    const uint32_t idesc = make_idesc_i8(128, N, 0, 1, 0, 0);
                    const uint32_t dcol = td + acc * kAccCols;
                    const uint32_t a_adv = (uint32_t)(tile * 128), b_adv = (uint32_t)((pass * kPassSlots * kSlotBytes) >> 4);
                    const long long ci0 = dbg_detail ? clock64() : 0;
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        if (i < nmma && !dbg_nomma && elect_one())
                            umma_i8(dcol, desc_join(alo[i] + a_adv, ahi), desc_join(blo[i] + b_adv, bhi), idesc, i ? 1u : 0u);
                    if (dbg_detail) w_issue += clock64() - ci0;
                    if (dbg_twice) {
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            if (i < nmma && elect_one())
                                umma_i8(dcol, desc_join(alo[i] + a_adv, ahi), desc_join(blo[i] + b_adv, bhi), idesc, 1u);
                    }
                    const long long cc0 = dbg_detail ? clock64() : 0;
                    if (elect_one()) umma_commit(&acc_full[acc]);
                    __syncwarp();
                    if (dbg_detail) w_commit += clock64() - cc0;
                    fill_par ^= 1u << acc;
                }
            }
            // the frame stage is free once every MMA issued so far has completed
            const long long ce0 = dbg_detail ? clock64() : 0;
            if (elect_one()) {
                if (npass > 0) umma_commit(&frame_empty[stage]);
                else mbar_arrive(&frame_empty[stage]);
            }
            __syncwarp();
            if (dbg_detail) w_fend += clock64() - ce0;
            if (++stage == gm.nst) { stage = 0; fphase ^= 1; }
        }
        if (dbg && lane == 0) {
            dbg[blockIdx.x * 8 + 2] = w_desc; dbg[blockIdx.x * 8 + 3] = w_fend; dbg[blockIdx.x * 8 + 5] = clock64() - t_begin;
            dbg[blockIdx.x * 8 + 4] = w_path; dbg[blockIdx.x * 8 + 0] = dbg_detail ? w_issue : t_begin - t_entry;
            dbg[blockIdx.x * 8 + 1] = w_fence; dbg[blockIdx.x * 8 + 6] = dbg_detail ? w_commit : clock64();
        }
    } else if (warp == 4 * kEpiGroups + 2) {
        // ================= second MMA-issue warp (kFast only): tiles 1 and 3 of every frame =================
        // One issuing warp could not keep the tensor pipe fed: between two tiles it waits for the accumulator stage, fences,
        // elects and commits, ~400 cycles during which the pipe drains (the kernel ran at 3300 cycles per frame with one
        // module active against 1664 cycles of MMA time).  With two warps one warp's gap is covered by the other's MMAs.
        // Warp A still owns the B operand: this warp learns about a new path from path_ready, like the epilogue warps.
        if constexpr (kFast) {
            int stage = 0, cur_path = -1, ns = 0;
            uint32_t fphase = 0, path_phase = 0, fill_par = 0;
            // (path index by counting: an integer division per frame is ~25 dependent instructions on these roles' critical paths)
            for (int t = t_lo, p = t_lo / B, p_next = (t_lo / B + 1) * B; t < t_hi; ++t) {
                if (t == p_next) { ++p; p_next += B; }
                if (p != cur_path) {
                    mbar_wait(path_ready, path_phase);
                    path_phase ^= 1;
                    cur_path = p;
                    ns = __popc(__reduce_or_sync(0xffffffffu, masks[p * nd.L]));
                }
                mbar_wait(&frame_full[stage], fphase);
                tc_fence_after();
                const int npass = (ns + kPassSlots - 1) / kPassSlots;
                const uint32_t fbase = smem_u32(frame_s + (size_t)stage * gm.stage_bytes);
                const uint32_t bsm = smem_u32(Bs);
                const uint32_t ahi = desc_hi(128), bhi = desc_hi(2048);
                uint64_t a64[8], b64[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int tap = i >> 1, ks = i & 1;
                    a64[i] = desc_join(desc_lo(fbase + (uint32_t)(2 * ks * gm.region_bytes + ((tap >> 1) * gm.Ws + (tap & 1)) * 16),
                                               (uint32_t)gm.region_bytes), ahi);
                    b64[i] = desc_join(desc_lo(bsm + (uint32_t)((tap * gm.s + 2 * ks) * 128), 128), bhi);
                }
                if (npass == 1) {
                    const int N = (24 * ns + 15) & ~15;
                    const uint32_t idesc = make_idesc_i8(128, N, 0, 1, 0, 0);
#pragma unroll
                    for (int tile = 1; tile < kAccStages; tile += 2) {
                        mbar_wait(&acc_empty[tile], ((fill_par >> tile) & 1u) ^ 1u);
                        tc_fence_after();
                        const uint32_t dcol = td + tile * kAccCols;
                        if (elect_one()) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) umma_i8(dcol, a64[i] + (uint64_t)(tile * 128), b64[i], idesc, i ? 1u : 0u);
                            umma_commit(&acc_full[tile]);
                        }
                        __syncwarp();
                    }
                    fill_par ^= (1u << kAccStages) - 1u;
                } else
                for (int pass = 0; pass < npass; ++pass) {
                    const int nsl = min(kPassSlots, ns - pass * kPassSlots);
                    const int N = (24 * nsl + 15) & ~15;
                    const uint32_t idesc = make_idesc_i8(128, N, 0, 1, 0, 0);
                    const uint64_t b_adv = (uint64_t)((pass * kPassSlots * kSlotBytes) >> 4);
#pragma unroll
                    for (int tile = 1; tile < kAccStages; tile += 2) {
                        mbar_wait(&acc_empty[tile], ((fill_par >> tile) & 1u) ^ 1u);
                        tc_fence_after();
                        const uint32_t dcol = td + tile * kAccCols;
                        if (elect_one()) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) umma_i8(dcol, a64[i] + (uint64_t)(tile * 128), b64[i] + b_adv, idesc, i ? 1u : 0u);
                            umma_commit(&acc_full[tile]);
                        }
                        __syncwarp();
                    }
                    fill_par ^= (1u << kAccStages) - 1u;
                }
                if (elect_one()) {
                    if (npass > 0) umma_commit(&frame_empty[stage]);
                    else mbar_arrive(&frame_empty[stage]);
                }
                __syncwarp();
                if (++stage == gm.nst) { stage = 0; fphase ^= 1; }
            }
        }
    } else {
        // ================= epilogue: thread = TMEM lane = output pixel of the tile =================
        const int quad = warp & 3;
        const int group = warp >> 2;                          // warps 4g .. 4g+3
        const int L = quad * 32 + lane;
        int cur_path = -1, ns = 0;
        uint32_t path_phase = 0, mask = 0, tilec = 0, gi = 0;   // gi: items this group has consumed from its stage
        long long w_full = 0, t_begin = clock64(), e_ld = 0, e_math = 0, e_t0 = 0, e_path = 0, e_ystore = 0;
        const bool dbg_epi = dbg && dbg[gridDim.x * 8] == 7;
        // (path index by counting: an integer division per frame is ~25 dependent instructions on these roles' critical paths)
        for (int t = t_lo, p = t_lo / B, p_next = (t_lo / B + 1) * B; t < t_hi; ++t) {
            if (t == p_next) { ++p; p_next += B; }
            if (p != cur_path) {
                if (cur_path >= 0) mbar_arrive(epi_idle);
                const long long cpw = dbg_epi ? pinned_clock() : 0;
                mbar_wait(path_ready, path_phase);
                if (dbg_epi) e_path += pinned_clock() - cpw;
                path_phase ^= 1;
                cur_path = p;
                mask = masks[p * nd.L];
                ns = __popc(mask);
            }
            const int npass = (ns + kPassSlots - 1) / kPassSlots;
            if (npass == 0) {     // no active module in layer 0: the layer output is exactly zero
                float4* yo = reinterpret_cast<float4*>(Y + (size_t)t * npix * PN_F);
                for (int i = threadIdx.x; i < npix * 2; i += 128 * kEpiGroups) yo[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                continue;
            }
            // kFast: a frame is exactly one tile per epilogue group, so the group's tile (and with it the thread's pixel)
            // never changes -- no loop, no per-frame index arithmetic
            for (int tile = kFast ? group : 0; tile < (kFast ? group + 1 : gm.tiles); ++tile, ++tilec) {
                if (!kFast && (int)(tilec & (kEpiGroups - 1)) != group) continue;
                const int m = tile * 128 + L;
                const int oy = (int)(((uint32_t)m * (uint32_t)gm.rcp_ws) >> 16), ox = m - oy * gm.Ws;   // m / Ws
                const bool valid = ox < gm.OW && oy < gm.OH;
                const int pix = oy * gm.OW + ox;
                float sum[8];
#pragma unroll
                for (int c = 0; c < 8; ++c) sum[c] = 0.f;
                for (int pass = 0; pass < npass; ++pass, ++gi) {
                    const uint32_t acc = (uint32_t)group;
                    const long long ce0 = dbg_epi ? pinned_clock() : 0;
                    mbar_wait(&acc_full[acc], gi & 1u);
                    if (dbg_epi) { const long long c = pinned_clock(); w_full += c - ce0; e_t0 = c; }
                    tc_fence_after();
                    const int nsl = min(kPassSlots, ns - pass * kPassSlots);
                    const uint32_t tbase = td + ((uint32_t)(quad * 32) << 16) + acc * kAccCols;
                    // A warp whose 32 raster rows all lie past the end of the frame's raster (the last tile covers 36 useful
                    // rows of 128) skips the read-out and the arithmetic.
                    const bool warp_all_junk = tile * 128 + quad * 32 >= gm.OH * gm.Ws;
                    if (warp_all_junk || (dbg && dbg[gridDim.x * 8] == 2)) {          // (debug mode 2: handshake only)
                        tc_fence_before();
                        mbar_arrive(&acc_empty[acc]);
                        continue;
                    }
                    // This loop is the kernel: the epilogue warps are issue-bound (ncu: 2.6 of 4 warp instructions per cycle
                    // and SM).  It is unrolled over the pass's slots so that the TMEM column, the scale / bias rows and the
                    // module index are immediates off per-tile bases, and it spends 5 instructions per digit triple:
                    // lo = d1 * 256 + d0 stays inside int32 (|d| <= 256 * 255 * 128, so |lo| < 2^31) and costs one LEA
                    // instead of a conversion and an fma; two conversions, fma(d2, 65536, lo), fma(.., scale, bias).
                    uint8_t* const rrow = R + (size_t)t * nd.M * npix + pix;
                    const float* const sbp = sbuf + pass * kPassSlots * 32;
                    const int* const jtp = jtab + pass * kPassSlots;
#pragma unroll
                    for (int sl = 0; sl < kPassSlots; ++sl) {
                        if (sl >= nsl) break;
                        uint32_t d0[8], d1[8], d2[8];
                        tmem_ld8(tbase + sl * 24, d0);
                        tmem_ld8(tbase + sl * 24 + 8, d1);
                        tmem_ld8(tbase + sl * 24 + 16, d2);
                        const float4* sb = reinterpret_cast<const float4*>(sbp + sl * 32);
                        float s0[8], bi[8];
                        *reinterpret_cast<float4*>(&s0[0]) = sb[0]; *reinterpret_cast<float4*>(&s0[4]) = sb[1];
                        *reinterpret_cast<float4*>(&bi[0]) = sb[2]; *reinterpret_cast<float4*>(&bi[4]) = sb[3];
                        const int j = jtp[sl];
                        tmem_ld_wait();
                        if (dbg_epi) { const long long c = pinned_clock(); e_ld += c - e_t0; e_t0 = c; }
                        if (sl == nsl - 1) {        // last read of this accumulator stage: hand it back to the MMA warp
                            tc_fence_before();
                            mbar_arrive(&acc_empty[acc]);
                        }
                        unsigned bits = 0;
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            const int lo = (int)d1[c] * 256 + (int)d0[c];
                            const float acc = fmaf((float)(int)d2[c], 65536.0f, (float)lo);
                            const float pre = fmaf(acc, s0[c], bi[c]);
                            if (pre > 0.f) { sum[c] += pre; bits |= 1u << c; }
                        }
                        if (valid) rrow[(size_t)j * npix] = (uint8_t)bits;
                        if (dbg_epi) { const long long c = pinned_clock(); e_math += c - e_t0; e_t0 = c; }
                    }
                }
                const long long cys = dbg_epi ? pinned_clock() : 0;
                if (valid) {
                    float4* yo = reinterpret_cast<float4*>(Y + ((size_t)t * npix + pix) * PN_F);
                    yo[0] = make_float4(sum[0], sum[1], sum[2], sum[3]);
                    yo[1] = make_float4(sum[4], sum[5], sum[6], sum[7]);
                }
                if (dbg_epi) e_ystore += pinned_clock() - cys;
            }
        }
        if (dbg && threadIdx.x == 0) { if (dbg[gridDim.x * 8] == 4) dbg[blockIdx.x * 8 + 6] = w_full; dbg[blockIdx.x * 8 + 7] = clock64() - t_begin; }
        if (dbg_epi && threadIdx.x == 0) {      // epilogue group 0, warp 0: where its time goes
            dbg[blockIdx.x * 8 + 1] = w_full; dbg[blockIdx.x * 8 + 2] = e_ld; dbg[blockIdx.x * 8 + 3] = e_math;
            dbg[blockIdx.x * 8 + 4] = clock64() - t_begin;
            dbg[blockIdx.x * 8 + 0] = e_path; dbg[blockIdx.x * 8 + 5] = e_ystore;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4 * kEpiGroups + 1) tmem_dealloc(td, kAccStages * kAccCols);
    if (dbg && !dbg_detail && threadIdx.x == 0) {        // timeline: MMA-loop end -> kernel exit, entry -> exit
        dbg[blockIdx.x * 8 + 6] = clock64() - dbg[blockIdx.x * 8 + 6];
        dbg[blockIdx.x * 8 + 7] = clock64() - t_entry;
    }
}


// =====================================================================================================
// Layer-0 weight gradient on the tensor cores.
//   dW_j[kh,kw,ci,co] = sum over the transitions t of the paths that use module j and over output
//   pixels of  x[t, oy*s+kh, ox*s+kw, ci] * dpre_j[t, oy, ox, co],   dpre_j = dY * relu_mask_j
// (tf.gradients through Conv2D, a3c_training_thread.py:55-60).  Per tap (dy,dx) this is
//   D_tap[64 x N] += A_tap^T[64 x m] * dq[m x N]      (m = block raster index, K of the MMA)
// with A_tap^T the staged frame read MN-major through a shifted descriptor and dq the masked output gradient quantised to
// kWgDigits signed base-256 digits of round(dpre * 2^(22-e_p)), 2^e_p >= max|dY| over the PATH's batch (one power-of-two
// scale per path, so the integer accumulation is exact and order-independent; resolution 2^-22 of the path's largest
// gradient entry).
//
// M = 128: the two taps of a kernel row share one MMA.  A tcgen05 instruction with M = 64 costs what one with M = 128
// costs (the pipe time is the operand fetch, 128 B per clock), and it fills only half of the TMEM lanes.  Rows 64..127 of
// the A operand must sit SBO * 4 bytes behind rows 0..63 in shared memory, so every frame is staged TWICE: copy 0 as it
// is and copy 1 shifted by one block (the TMA load starts 4 u32 further right; the columns it then reads past the row end are
// zero-filled, they only feed discarded positions).  One MMA per kernel row dy then covers (dy, dx = 0) in accumulator rows
// 0..63 and (dy, dx = 1) in rows 64..127: 28 MMAs per transition instead of 56.  The TMA box carries one extra (zero-filled)
// block row per region so that four regions are a multiple of 128 bytes and both copies start 128-byte aligned.
//
// One work item = (path, chunk of its transitions); the two accumulators (one per dy, N <= 96 columns each) live in
// TMEM for the whole chunk.  There are TWO accumulator sets: the flush of item i (dedicated warps) runs underneath the
// MMAs of item i+1, which used to wait for it (20 % of the kernel).
// Warp roles (704 threads): warps 0-15 quantise dq, one raster row per thread (and keep the bias-gradient partial sums),
// warps 16-19 flush (one TMEM lane quadrant each), warp 20 frame producer, warp 21 MMA issue (highest id = scheduler
// priority).  With 8 quantiser warps the MMA warp waited for quantised operands 58 % of the time.
// =====================================================================================================
constexpr int kWgDigits = 3;
constexpr int kWgQuantWarps = 16;                // at most: one raster row per quantiser thread (mq <= 512)
constexpr int kWgSetCols = 256;                  // TMEM columns per accumulator set: two accumulators of 128
constexpr int kWgAccCols = 128;

// max |x| per path (x = the layer-0 output gradient [P][per_path floats]) when no earlier kernel has left it behind
__global__ void absmax_kernel(const float* __restrict__ x, size_t n4_per_path, unsigned* __restrict__ out) {
    const int p = blockIdx.y;
    const float4* xp = reinterpret_cast<const float4*>(x) + (size_t)p * n4_per_path;
    float m = 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4_per_path; i += (size_t)gridDim.x * blockDim.x) {
        const float4 v = xp[i];
        m = fmaxf(m, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(out + p, __float_as_uint(m));   // non-negative floats order like uints
}

struct WgGeom {
    int region_bytes;     // (Hs + 1) * row_bytes: one region of a staged frame copy (multiple of 32)
    int stage_bytes;      // two copies of four regions + slack for the shifted reads
    int region_q;         // mq * 16: one 16-column group of dq
};

// QW = quantiser warps: 14 cover the 84x84 frame's 448 raster rows; then the block has 20 warps and 96 registers per thread
// (22 warps are allocated as 24: 80 registers, at which the quantisers had no room to keep their loads in flight).
template <int QW>
__global__ void __launch_bounds__((QW + 6) * 32, 1)
conv0_wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmap, const float* __restrict__ dY, const uint8_t* __restrict__ R,
                      const unsigned* __restrict__ absmax, PnNetDesc nd, const uint32_t* __restrict__ masks, int B,
                      int P, int cs, int nch, int nch_alloc, TcGeom gm, WgGeom wg, float* __restrict__ part, long long* __restrict__ dbg_arg) {
    long long* const dbg = kTcDebugBuild ? dbg_arg : nullptr;   // the instrumentation below compiles away in the product build
    const bool dbg8 = dbg && dbg[148 * 8] == 8;                 // PN_TC_DEBUG=8: where the weight-gradient kernel's roles wait
    const long long t_entry = dbg8 ? pinned_clock() : 0;
    extern __shared__ __align__(1024) uint8_t smem[];
    const int dq_stage_bytes = 2 * kWgDigits * wg.region_q;     // up to 4 modules: 3 digits x 32 columns = 6 groups of 16
    uint8_t* frame_s = smem;
    uint8_t* dq_s = smem + 2 * (size_t)wg.stage_bytes;
    float* bias_s = reinterpret_cast<float*>(dq_s + 2 * (size_t)dq_stage_bytes);     // [quantiser warps][32]
    uint64_t* bars = reinterpret_cast<uint64_t*>(bias_s + kWgQuantWarps * 32);
    uint64_t* frame_full = bars;          // [2]
    uint64_t* frame_empty = bars + 2;     // [2]
    uint64_t* dq_full = bars + 4;         // [2]
    uint64_t* dq_empty = bars + 6;        // [2]
    uint64_t* acc_full = bars + 8;        // [2] accumulator set s holds a finished item
    uint64_t* acc_empty = bars + 10;      // [2] ... and has been flushed
    __shared__ uint32_t tmem_base_s;

    // broadcast through a shuffle so that the compiler knows the warp index is warp-uniform (role branches
    // then count as uniform control flow and the MMA operands can live in uniform registers)
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int npix = gm.OH * gm.OW;
    const int nitems = P * nch;      // round-robin over the CTAs: a path's chunks land on different CTAs, which spreads expensive
                                     // paths by itself (a contiguous cost-balanced split measured worse here)
    const int stride_part = nd.KK[0] * PN_F + PN_F;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&frame_full[i], 1); mbar_init(&frame_empty[i], 1);
            // one arrival per WARP (after __syncwarp): 512 threads arriving on one barrier serialise on its shared-memory word
            mbar_init(&dq_full[i], QW); mbar_init(&dq_empty[i], 1);
            mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4);
        }
        fence_mbar_init();
    }
    if (warp == QW + 5) tmem_alloc(&tmem_base_s, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tmem_base_s;

    if (warp == QW + 4) {
        // ================= frame producer: two copies per frame (see above) =================
        if (lane == 0) tma_prefetch_desc(&tmap);
        int stage = 0;
        uint32_t phase = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int p = item / nch, ch = item % nch;
            const int b_lo = ch * cs, b_hi = min(B, b_lo + cs);
            const int ns = __popc(masks[p * nd.L]);
            const int npass = (ns + kPassSlots - 1) / kPassSlots;
            for (int pass = 0; pass < npass; ++pass)
                for (int b = b_lo; b < b_hi; ++b) {
                    const int t = p * B + b;
                    mbar_wait(&frame_empty[stage], phase ^ 1);
                    if (lane == 0) {
                        uint8_t* dst = frame_s + (size_t)stage * wg.stage_bytes;
                        mbar_expect_tx(&frame_full[stage], (uint32_t)(8 * wg.region_bytes));
                        tma_load_4d(dst, &tmap, 0, 0, 0, t, &frame_full[stage]);
                        tma_load_4d(dst + 4 * wg.region_bytes, &tmap, 4, 0, 0, t, &frame_full[stage]);     // one block = 4 u32 further right
                    }
                    __syncwarp();
                    if (++stage == 2) { stage = 0; phase ^= 1; }
                }
        }
    } else if (warp == QW + 5) {
        // ================= MMA issue =================
        // whole-warp, warp-uniform code; only the tcgen05 instructions are elected to one lane
        int stage = 0;
        uint32_t phase = 0, nacc = 0;                           // nacc: (item, pass) accumulations started so far
        long long m_wframe = 0, m_wdq = 0, m_wacc = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int p = item / nch, ch = item % nch;
            const int b_lo = ch * cs, b_hi = min(B, b_lo + cs);
            const int ns = __popc(__reduce_or_sync(0xffffffffu, masks[p * nd.L]));
            const int npass = (ns + kPassSlots - 1) / kPassSlots;
            for (int pass = 0; pass < npass; ++pass) {
                if (b_lo >= b_hi) continue;
                const int nsl = min(kPassSlots, ns - pass * kPassSlots);
                const int N = kWgDigits * 8 * ((nsl + 1) & ~1);                  // columns (digit, slot, channel), slots padded to even
                const uint32_t idesc = make_idesc_i8(128, N, 0, 1, 1, 1);
                const uint32_t set = nacc & 1u;
                const long long ca = dbg8 ? pinned_clock() : 0;
                mbar_wait(&acc_empty[set], ((nacc >> 1) & 1u) ^ 1u);           // the flush of the item that used this set is done
                if (dbg8) m_wacc += pinned_clock() - ca;
                tc_fence_after();
                const uint32_t d0 = td + set * kWgSetCols;
                uint32_t accum = 0;
                for (int b = b_lo; b < b_hi; ++b) {
                    const long long c0 = dbg8 ? pinned_clock() : 0;
                    mbar_wait(&frame_full[stage], phase);
                    const long long c1 = dbg8 ? pinned_clock() : 0;
                    mbar_wait(&dq_full[stage], phase);
                    if (dbg8) { const long long c2 = pinned_clock(); m_wframe += c1 - c0; m_wdq += c2 - c1; }
                    tc_fence_after();
                    const uint32_t fbase = smem_u32(frame_s + (size_t)stage * wg.stage_bytes);
                    const uint32_t qbase = smem_u32(dq_s + (size_t)stage * dq_stage_bytes);
                    // a K step advances both start addresses by 32 blocks (descriptor unit: 16 bytes): one add each
                    uint32_t alo0 = desc_lo(fbase, 128), alo1 = desc_lo(fbase + (uint32_t)(gm.Ws * 16), 128);
                    uint32_t blo = desc_lo(qbase, 128);
                    const uint32_t ahi = desc_hi((uint32_t)wg.region_bytes), bhi = desc_hi((uint32_t)wg.region_q);
                    for (int ks = 0; ks < gm.mq / 32; ++ks) {
                        const uint64_t bd = desc_join(blo, bhi);
                        if (elect_one()) {
                            umma_i8(d0, desc_join(alo0, ahi), bd, idesc, accum);
                            umma_i8(d0 + kWgAccCols, desc_join(alo1, ahi), bd, idesc, accum);
                        }
                        alo0 += 32; alo1 += 32; blo += 32;
                        accum = 1;
                    }
                    if (elect_one()) {
                        umma_commit(&frame_empty[stage]);
                        umma_commit(&dq_empty[stage]);
                    }
                    __syncwarp();
                    if (++stage == 2) { stage = 0; phase ^= 1; }
                }
                if (elect_one()) umma_commit(&acc_full[set]);
                __syncwarp();
                ++nacc;
            }
        }
        if (dbg8 && lane == 0) {
            dbg[blockIdx.x * 8 + 0] = m_wframe; dbg[blockIdx.x * 8 + 1] = m_wdq; dbg[blockIdx.x * 8 + 2] = m_wacc;
            dbg[blockIdx.x * 8 + 3] = pinned_clock() - t_entry;
        }
    } else if (warp >= QW) {
        // ================= flush: TMEM lane = accumulator row = (dx, g, byte) =================
        const int quad = warp & 3;
        const int r = quad * 32 + lane;                         // 0..127
        const int dx = r >> 6, g = (r >> 4) & 3, bb = r & 15;
        uint32_t nacc = 0;
        long long f_wait = 0, f_work = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int p = item / nch, ch = item % nch;
            const int b_lo = ch * cs, b_hi = min(B, b_lo + cs);
            const uint32_t mask = masks[p * nd.L];
            const int ns = __popc(mask);
            const int npass = (ns + kPassSlots - 1) / kPassSlots;
            const unsigned am = absmax[p];
            int e = 0;
            if (am) frexpf(__uint_as_float(am), &e);            // max = f * 2^e, f in [0.5, 1)  =>  2^e > max
            const double dscale = am ? ldexp(1.0, e - (8 * kWgDigits - 2)) / 255.0 : 0.0;
            for (int pass = 0; pass < npass; ++pass) {
                if (b_lo >= b_hi) continue;
                const int nsl = min(kPassSlots, ns - pass * kPassSlots);
                const int dstride = 8 * ((nsl + 1) & ~1);       // columns per digit
                const uint32_t set = nacc & 1u;
                const long long cf = dbg8 ? pinned_clock() : 0;
                mbar_wait(&acc_full[set], (nacc >> 1) & 1u);
                const long long cf1 = dbg8 ? pinned_clock() : 0;
                tc_fence_after();
                for (int dy = 0; dy < 2; ++dy) {
                    const int kh = dy * gm.s + g, kw = dx * gm.s + bb / gm.C, ci = bb % gm.C;
                    const int kk = (kh * nd.k[0] + kw) * gm.C + ci;
                    for (int sl = 0; sl < nsl; ++sl) {
                        uint32_t q0[8], q1[8], q2[8];
                        const uint32_t ta = td + ((uint32_t)(quad * 32) << 16) + set * kWgSetCols + dy * kWgAccCols + sl * 8;
                        tmem_ld8(ta, q0); tmem_ld8(ta + dstride, q1); tmem_ld8(ta + 2 * dstride, q2);
                        tmem_ld_wait();
                        float o[8];
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            // digits combined in 64-bit integers (exact), then ONE conversion
                            const long long v = (long long)(int)q0[c] + ((long long)(int)q1[c] << 8) + ((long long)(int)q2[c] << 16);
                            o[c] = (float)((double)v * dscale);
                        }
                        const int j = __fns(mask, 0, pass * kPassSlots + sl + 1);
                        float4* dst = reinterpret_cast<float4*>(
                            part + ((size_t)(p * nch_alloc + ch) * nd.M + j) * stride_part + (size_t)kk * PN_F);
                        dst[0] = make_float4(o[0], o[1], o[2], o[3]);
                        dst[1] = make_float4(o[4], o[5], o[6], o[7]);
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[set]);
                if (dbg8) { f_wait += cf1 - cf; f_work += pinned_clock() - cf1; }
                ++nacc;
            }
        }
        if (dbg8 && threadIdx.x == QW * 32) { dbg[blockIdx.x * 8 + 5] = f_wait; dbg[blockIdx.x * 8 + 6] = f_work; }
    } else {
        // ================= quantise dq per transition; bias-gradient partials per (item, pass) =================
        const int m = threadIdx.x;                             // this thread's raster row, 0..511 (rows >= mq: nothing to do)
        const int oy = m / gm.Ws, ox = m - oy * gm.Ws;
        const bool inside = m < gm.mq, valid = inside && ox < gm.OW && oy < gm.OH;
        const int pix = oy * gm.OW + ox;
        int stage = 0;
        uint32_t phase = 0;
        long long q_wempty = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int p = item / nch, ch = item % nch;
            const int b_lo = ch * cs, b_hi = min(B, b_lo + cs);
            const uint32_t mask = masks[p * nd.L];
            const int ns = __popc(mask);
            const int npass = (ns + kPassSlots - 1) / kPassSlots;
            const unsigned am = absmax[p];
            int e = 0;
            if (am) frexpf(__uint_as_float(am), &e);
            const float qscale = am ? ldexpf(1.0f, (8 * kWgDigits - 2) - e) : 0.f;     // |q| <= 2^22
            for (int pass = 0; pass < npass; ++pass) {
                if (b_lo >= b_hi) continue;
                const int nsl = min(kPassSlots, ns - pass * kPassSlots);
                const int npair = (nsl + 1) >> 1;               // 16-column groups per digit
                int jl[kPassSlots];
#pragma unroll
                for (int sl = 0; sl < kPassSlots; ++sl) jl[sl] = sl < nsl ? __fns(mask, 0, pass * kPassSlots + sl + 1) : 0;
                float bsum[kPassSlots][8];
#pragma unroll
                for (int sl = 0; sl < kPassSlots; ++sl)
#pragma unroll
                    for (int c = 0; c < 8; ++c) bsum[sl][c] = 0.f;
                // Software pipeline over the transitions: the loads of transition b + 1 are issued before transition b is
                // processed.  All quantiser warps run in lock step (they meet at the ring's barriers), so without this
                // every warp sat out the same L2 round trip once per transition.
                // (The mask bytes stay in their own registers until the next iteration: any arithmetic on a loaded value right
                // here -- the shift that packs them was the top stall of the kernel -- makes the warp wait for the load.)
                float4 ga_n = make_float4(0.f, 0.f, 0.f, 0.f), gb_n = ga_n;
                uint32_t rraw_n[kPassSlots];
                auto fetch = [&](int b) {
                    ga_n = gb_n = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int sl = 0; sl < kPassSlots; ++sl) rraw_n[sl] = 0;
                    if (valid) {
                        const size_t t = (size_t)p * B + b;
                        const float4* g4 = reinterpret_cast<const float4*>(dY + (t * npix + pix) * PN_F);
                        ga_n = __ldg(g4); gb_n = __ldg(g4 + 1);
#pragma unroll
                        for (int sl = 0; sl < kPassSlots; ++sl)
                            if (sl < nsl) rraw_n[sl] = __ldg(R + (t * nd.M + jl[sl]) * npix + pix);
                    }
                };
                fetch(b_lo);
                for (int b = b_lo; b < b_hi; ++b) {
                    const float4 ga = ga_n, gb = gb_n;
                    uint32_t rb = 0;
#pragma unroll
                    for (int sl = 0; sl < kPassSlots; ++sl) rb |= rraw_n[sl] << (8 * sl);
                    if (b + 1 < b_hi) fetch(b + 1);
                    const long long cq = dbg8 ? pinned_clock() : 0;
                    mbar_wait(&dq_empty[stage], phase ^ 1);
                    if (dbg8) q_wempty += pinned_clock() - cq;
                    if (inside) {
                        uint8_t* dq = dq_s + (size_t)stage * dq_stage_bytes + m * 16;
                        const float dv[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w};
                        // balanced base-256 digits: the low byte of q is digit 0, of q1 = (q + 128) >> 8 digit 1, of
                        // q2 = (q1 + 128) >> 8 digit 2 (two's complement bytes are what the s8 operand wants)
                        int q0[8], q1[8], q2[8];
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            q0[c] = __float2int_rn(dv[c] * qscale);
                            q1[c] = (q0[c] + 128) >> 8;
                            q2[c] = (q1[c] + 128) >> 8;
                        }
                        uint32_t w[kWgDigits][2];               // [digit][channels 0-3 / 4-7], byte k = channel
#pragma unroll
                        for (int hf = 0; hf < 2; ++hf) {
                            const int* a0 = &q0[4 * hf]; const int* a1 = &q1[4 * hf]; const int* a2 = &q2[4 * hf];
                            w[0][hf] = __byte_perm(__byte_perm(a0[0], a0[1], 0x0040), __byte_perm(a0[2], a0[3], 0x0040), 0x5410);
                            w[1][hf] = __byte_perm(__byte_perm(a1[0], a1[1], 0x0040), __byte_perm(a1[2], a1[3], 0x0040), 0x5410);
                            w[2][hf] = __byte_perm(__byte_perm(a2[0], a2[1], 0x0040), __byte_perm(a2[2], a2[3], 0x0040), 0x5410);
                        }
                        // columns (digit, slot, channel): a 16-column group = two slots of one digit
#pragma unroll
                        for (int pr = 0; pr < kPassSlots / 2; ++pr) {
                            if (pr >= npair) break;
                            const unsigned bits0 = (rb >> (16 * pr)) & 255u, bits1 = (rb >> (16 * pr + 8)) & 255u;   // 0 outside the grid / for the pad slot
                            // bit k -> byte k = 0xFF: spread the four bits to the bytes' low bits, then multiply by 255
                            const uint32_t m0lo = (((bits0 & 15u) * 0x00204081u) & 0x01010101u) * 255u;
                            const uint32_t m0hi = (((bits0 >> 4) * 0x00204081u) & 0x01010101u) * 255u;
                            const uint32_t m1lo = (((bits1 & 15u) * 0x00204081u) & 0x01010101u) * 255u;
                            const uint32_t m1hi = (((bits1 >> 4) * 0x00204081u) & 0x01010101u) * 255u;
#pragma unroll
                            for (int d = 0; d < kWgDigits; ++d) {
                                const uint4 gq = make_uint4(w[d][0] & m0lo, w[d][1] & m0hi, w[d][0] & m1lo, w[d][1] & m1hi);
                                *reinterpret_cast<uint4*>(dq + (size_t)(d * npair + pr) * wg.region_q) = gq;
                            }
#pragma unroll
                            for (int c = 0; c < 8; ++c) {
                                if ((bits0 >> c) & 1u) bsum[2 * pr][c] += dv[c];
                                if ((bits1 >> c) & 1u) bsum[2 * pr + 1][c] += dv[c];
                            }
                        }
                    }
                    fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&dq_full[stage]);
                    if (++stage == 2) { stage = 0; phase ^= 1; }
                }
                // ---- bias gradient of this (item, pass): warp shuffle trees, then the warp sums in warp order ----
                float mine = 0.f;                               // lane l ends up with the warp's sum of value l = (slot, channel)
#pragma unroll
                for (int sl = 0; sl < kPassSlots; ++sl)
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        float v = bsum[sl][c];
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                        if (lane == sl * 8 + c) mine = v;
                    }
                asm volatile("bar.sync 1, %0;" ::"n"(QW * 32) : "memory");   // the previous (item, pass) has been read out of bias_s
                bias_s[warp * 32 + lane] = mine;
                asm volatile("bar.sync 1, %0;" ::"n"(QW * 32) : "memory");
                if (threadIdx.x < nsl * 8) {
                    float sacc = 0.f;
#pragma unroll
                    for (int q = 0; q < QW; ++q) sacc += bias_s[q * 32 + threadIdx.x];
                    part[((size_t)(p * nch_alloc + ch) * nd.M + jl[threadIdx.x >> 3]) * stride_part + (size_t)nd.KK[0] * PN_F + (threadIdx.x & 7)] = sacc;
                }
            }
        }
        if (dbg8 && threadIdx.x == 0) {
            dbg[blockIdx.x * 8 + 4] = q_wempty;
            dbg[blockIdx.x * 8 + 7] = pinned_clock() - t_entry;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == QW + 5) tmem_dealloc(td, 512);
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
    // regions are packed (TMA box layout); shifted descriptors read up to `extent` blocks from a region's
    // start, running into the next region / the padding behind the last one: any bytes are harmless there
    // because those MMA rows are discarded (forward) or multiplied by zero gradients (weight gradient)
    const int region_blocks = g.Hs * g.Ws;
    g.region_bytes = region_blocks * 16;
    g.mq = (rows + 31) & ~31;
    const int extent = std::max(g.tiles * 128, g.mq) + g.Ws + 2;
    g.stage_bytes = (((s - 1) * g.region_bytes + extent * 16) + 1023) & ~1023;
    g.row_bytes = W * C;
    g.frame_bytes = H * W * C;
    g.ksteps = s / 2;
    g.rcp_ws = (65536 + g.Ws - 1) / g.Ws;
    for (int m = 0; m < g.tiles * 128; ++m)
        if ((int)(((unsigned)m * (unsigned)g.rcp_ws) >> 16) != m / g.Ws) return false;
    if (g.row_bytes % 16 || g.row_bytes / 4 > 256 || g.Hs > 256) return false;
    const size_t fixed = (size_t)(nd.M + 1) * kSlotBytes + (size_t)nd.M * 32 * sizeof(float) +
                         (size_t)(nd.M + 2) * sizeof(int) + 64 * sizeof(uint64_t);
    for (g.nst = 3; g.nst >= 2; --g.nst) {      // (5 stages measured no faster: the frame loads are not the limit)
        smem_bytes = (size_t)g.nst * g.stage_bytes + fixed;
        if (smem_bytes <= 220 * 1024) return true;
    }
    return false;
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_frame_tmap(pn_handle* h, const TcGeom& g, const uint8_t* frames, int T, CUtensorMap* out, int extra_rows = 0) {
    static PFN_encodeTiled encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
            return pn_fail(h, PN_ERR_CUDA, "cuTensorMapEncodeTiled is not available: %s", cudaGetErrorString(e));
        encode = reinterpret_cast<PFN_encodeTiled>(fn);
    }
    if (reinterpret_cast<uintptr_t>(frames) & 15) return pn_fail(h, PN_ERR_INVALID, "frames must be 16-byte aligned");
    const cuuint64_t dims[4] = {(cuuint64_t)(g.row_bytes / 4), (cuuint64_t)g.Hs, (cuuint64_t)g.s, (cuuint64_t)T};
    const cuuint64_t strides[3] = {(cuuint64_t)g.row_bytes * g.s, (cuuint64_t)g.row_bytes, (cuuint64_t)g.frame_bytes};
    // extra_rows > 0: the box is taller than the tensor; the rows past the end are zero-filled in shared memory
    const cuuint32_t box[4] = {(cuuint32_t)(g.row_bytes / 4), (cuuint32_t)(g.Hs + extra_rows), (cuuint32_t)g.s, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, const_cast<uint8_t*>(frames), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return pn_fail(h, PN_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return PN_OK;
}

// relative cost of one transition by the number of active layer-0 modules of its path (us per 8192 transitions, B200)
const float kFwdCost[11] = {50.f, 62.f, 75.f, 87.f, 101.f, 148.f, 155.f, 164.f, 179.f, 226.f, 240.f};   // tools/layer0_cost.py, current kernel

// cut `units` work units (unit u belongs to path u / per_path) into `ctas` contiguous ranges of equal cumulative cost
void balance_units(const pn_handle* h, const float* table, int per_path, int ctas, TcBounds& out) {
    const int P = h->P, L = h->cfg.L;
    std::vector<double> cum((size_t)P + 1, 0.0);
    for (int p = 0; p < P; ++p) {
        const int ns = std::min(10, __builtin_popcount(h->h_masks[(size_t)p * L]));
        cum[p + 1] = cum[p] + (double)per_path * (double)table[ns];
    }
    const double total = cum[P];
    const long long units = (long long)P * per_path;
    out.b[0] = 0;
    int p = 0;
    for (int c = 1; c < ctas; ++c) {
        const double want = total * c / ctas;
        while (p < P && cum[p + 1] <= want) ++p;
        long long u;
        if (p >= P || total <= 0.0) u = units * c / ctas;
        else {
            const double per = (cum[p + 1] - cum[p]) / per_path;           // cost of one unit of path p
            u = (long long)p * per_path + (per > 0.0 ? (long long)((want - cum[p]) / per + 0.5) : 0);
        }
        u = std::max<long long>(u, out.b[c - 1]);
        out.b[c] = (int)std::min(u, units);
    }
    out.b[ctas] = (int)units;
}

}  // namespace

bool pn_conv0_tc_supported(const pn_handle* h) {
    TcGeom g;
    size_t sm;
    return make_geom(h, g, sm);
}

size_t pn_conv0_tc_scratch_bytes(const pn_handle* h) {
    // int8 weight digits + fp32 scales per module, then one word per path: max |dY| of layer 0 over the path's batch
    return (((size_t)h->nd.M * (kSlotBytes + 8 * sizeof(float)) + 15) & ~(size_t)15) + (((size_t)h->P * 4 + 15) & ~(size_t)15);
}

unsigned* pn_conv0_absmax_slot(const pn_handle* h) {        // [P] words
    if (!h->tc_conv0 || !h->tc_w8) return nullptr;
    return reinterpret_cast<unsigned*>(h->tc_w8 + (((size_t)h->nd.M * (kSlotBytes + 8 * sizeof(float)) + 15) & ~(size_t)15));
}

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
        pn_launch(pn_use_pdl(h), conv0_wprep_kernel, dim3(nd.M), dim3(256), 0, st, (const float*)h->params, nd, g.s, w8, wscale);
    }
    PN_LAUNCH_CHECK(h, "conv0_wprep_kernel");
    const bool fast = g.tiles == kAccStages && 4 * g.ksteps == 8;
    if (fast) PN_SET_SMEM(h, conv0_fwd_tc_kernel<true>, smem_bytes);
    else PN_SET_SMEM(h, conv0_fwd_tc_kernel<false>, smem_bytes);
    const int grid = std::min(std::min(T, h->num_sms), kMaxCtas);
    TcBounds bounds;
    balance_units(h, kFwdCost, B, grid, bounds);
    CUtensorMap tmap;
    if (int rc = make_frame_tmap(h, g, frames, T, &tmap)) return rc;
    {
        PnProfScope prof(h, PN_PROF_CONV_FWD + 0, st);
        // (programmatic dependent launch behind the weight-digit kernel: see pn_pdl_sync)
        if (fast)
            pn_launch(pn_use_pdl(h), conv0_fwd_tc_kernel<true>, dim3(grid), dim3(kFwdThreads), smem_bytes, st, tmap, w8, wscale, h->params, nd,
                      h->d_masks, B, T, g, bounds, h->act[0], h->rmask[0], h->tc_dbg);
        else
            pn_launch(pn_use_pdl(h), conv0_fwd_tc_kernel<false>, dim3(grid), dim3(kFwdThreads), smem_bytes, st, tmap, w8, wscale, h->params, nd,
                      h->d_masks, B, T, g, bounds, h->act[0], h->rmask[0], h->tc_dbg);
    }
    PN_LAUNCH_CHECK(h, "conv0_fwd_tc_kernel");
    return PN_OK;
}

// geometry of the weight-gradient kernel's staging; false if the frame does not fit its scheme
static bool make_wg_geom(const TcGeom& g, WgGeom& w, size_t& smem_bytes) {
    w.region_bytes = (g.Hs + 1) * g.row_bytes;                       // one zero-filled block row appended per region
    if (w.region_bytes % 32) return false;                           // four regions must be a multiple of 128 bytes
    // shifted reads reach block mq - 1 + Ws + 1 of the last region of copy 1
    const int over = std::max(0, (g.mq + g.Ws + 2) * 16 - w.region_bytes);
    w.stage_bytes = (8 * w.region_bytes + over + 1023) & ~1023;
    w.region_q = g.mq * 16;
    if (g.mq > 512) return false;                                    // the quantiser covers 512 raster rows per round
    smem_bytes = 2 * (size_t)w.stage_bytes + 2 * (size_t)(2 * kWgDigits * w.region_q) + kWgQuantWarps * 32 * sizeof(float) + 16 * sizeof(uint64_t);
    return smem_bytes <= 227 * 1024;
}

bool pn_conv0_wgrad_tc_supported(const pn_handle* h) {
    TcGeom g;
    WgGeom w;
    size_t sm = 0;
    return make_geom(h, g, sm) && make_wg_geom(g, w, sm);
}

int pn_launch_conv0_wgrad_tc(pn_handle* h, const uint8_t* frames, int B, cudaStream_t st) {
    TcGeom g;
    WgGeom wg;
    size_t fwd_smem = 0, smem_bytes = 0;
    if (!make_geom(h, g, fwd_smem) || !make_wg_geom(g, wg, smem_bytes))
        return pn_fail(h, PN_ERR_STATE, "tensor-core conv0 weight gradient not supported for this geometry");
    const PnNetDesc& nd = h->nd;
    const int P = h->P, T = P * B;
    const int npix = g.OH * g.OW;
    unsigned* absmax = pn_conv0_absmax_slot(h);
    if (!h->absmax_fused) {          // the layer-1 data-gradient kernel normally leaves the per-path abs-max behind
        PN_CUDA(h, cudaMemsetAsync(absmax, 0, (size_t)P * sizeof(unsigned), st));
        PnProfScope prof(h, PN_PROF_CONV0_DQUANT, st);
        dim3 grid(std::max(1, 2 * h->num_sms / P), P);
        absmax_kernel<<<grid, 256, 0, st>>>(h->dact[0], (size_t)B * npix * PN_F / 4, absmax);
        PN_LAUNCH_CHECK(h, "absmax_kernel");
    }
    h->absmax_fused = false;
    // chunks per path: balance P*nch items over the SMs, at most 32 transitions per chunk (int32 accumulators)
    int best_nch = 1;
    double best_eff = -1.0;
    for (int nch = 1; nch <= h->nch && nch <= B; ++nch) {
        const int cs = (B + nch - 1) / nch;
        if (cs > 32 || (B + cs - 1) / cs != nch) continue;
        const long long items = (long long)P * nch, rounds = (items + h->num_sms - 1) / h->num_sms;
        // cost of an item ~ cs transitions of MMAs; its flush runs underneath the next item's MMAs, so only the
        // pipeline fill (~half a transition) is charged per item
        const double eff = (double)T / ((double)rounds * h->num_sms * (cs + 0.5));
        if (eff > best_eff) { best_eff = eff; best_nch = nch; }
    }
    if (best_eff < 0) return pn_fail(h, PN_ERR_STATE, "no valid chunking for the conv0 weight-gradient kernel");
    const int nch = best_nch, cs = (B + nch - 1) / nch;
    const bool narrow = g.mq <= 14 * 32;
    if (narrow) PN_SET_SMEM(h, conv0_wgrad_tc_kernel<14>, smem_bytes);
    else PN_SET_SMEM(h, conv0_wgrad_tc_kernel<kWgQuantWarps>, smem_bytes);
    const int items = P * nch;
    const int grid = std::min(std::min(items, h->num_sms), kMaxCtas);
    CUtensorMap tmap;
    if (int rc = make_frame_tmap(h, g, frames, T, &tmap, 1)) return rc;
    {
        PnProfScope prof(h, PN_PROF_CONV_WGRAD + 0, st);
        if (narrow)
            conv0_wgrad_tc_kernel<14><<<grid, (14 + 6) * 32, smem_bytes, st>>>(tmap, h->dact[0], h->rmask[0], absmax, nd, h->d_masks,
                                                                             B, P, cs, nch, h->nch, g, wg, h->wpart[0], h->tc_dbg);
        else
            conv0_wgrad_tc_kernel<kWgQuantWarps><<<grid, (kWgQuantWarps + 6) * 32, smem_bytes, st>>>(
                tmap, h->dact[0], h->rmask[0], absmax, nd, h->d_masks, B, P, cs, nch, h->nch, g, wg, h->wpart[0], h->tc_dbg);
    }
    PN_LAUNCH_CHECK(h, "conv0_wgrad_tc_kernel");
    h->last_wgrad_nch = nch;
    return PN_OK;
}
