// The LSTM variant of the PathNet actor-critic (GameACPathNetLSTMNetwork, game_ac_network.py:303-521), population-batched.
//
// Reference semantics.  The module super-net is the same as the feed-forward net's, but the last layer's module sum is
// NOT divided by M (game_ac_network.py:394); it is the input of tf.contrib.rnn.BasicLSTMCell(256, state_is_tuple=True)
// unrolled by tf.nn.dynamic_rnn over the rollout's steps, batch 1 per path (:396-416); the cell's outputs feed the two
// heads (:420-426).  BasicLSTMCell (TF 1.x rnn_cell_impl.py), forget_bias = 1.0:
//     [i | j | f | o] = [x_t, h_{t-1}] @ kernel[512, 1024] + bias[1024]        (gate order i, j, f, o)
//     c_t = c_{t-1} * sigmoid(f + 1.0) + sigmoid(i) * tanh(j)
//     h_t = tanh(c_t) * sigmoid(o)
// kernel rows 0..255 multiply the input, rows 256..511 the previous output.  Variables W_lstm = ".../kernel",
// b_lstm = ".../bias" are appended to get_vars() and are never frozen (:437-438,502-520).  The state (c, h) is carried by
// the host between calls (run_policy_and_value updates it, run_value rolls it back, reset_state zeroes it: :441-481); a
// training call starts from the state at the beginning of the rollout and feeds the batch in chronological order
// (a3c_training_thread.py:113-114,184-203).
//
// Here every path is one sequence: transitions t = p*B + b, b = 0..B-1 in time order.  The input projection of ALL
// steps is one GEMM, the recurrence is one small kernel per time step over all paths, and the backward pass (BPTT)
// mirrors it: one kernel per step for the gate gradients, then two GEMMs for the input gradient and the weight gradient.
// Plain fp32 FFMA kernels (the recurrent GEMM is [P, 256] x [256, 1024] per step); fixed summation order everywhere.
#include "pn_internal.cuh"

namespace {

constexpr int kH = 256;          // cell width (game_ac_network.py:397)
constexpr int kG = 4 * kH;       // gate columns

// ---- fp32 tile GEMM core: acc[TM][TN] += A[m0.., k] * B[k, n0..] over k in [k_lo, k_hi) --------------------------------
// 256 threads as 16 x 16, thread (ty, tx) owns rows ty*TM.. and columns tx*TN.. of a (16 TM) x (16 TN) tile; 16-deep K
// slices.  The loads of slice s + 1 are in flight (in registers) while slice s is multiplied out of shared memory: the
// first version fetched, synchronised and multiplied one slice at a time, and every kernel built on it ran at one
// exposed memory round trip per slice (the K = 8192 weight-gradient GEMM: 512 slices = 0.56 ms).
// Operands are addressed with element strides, so one core serves A x B, A^T x B and A x B^T; the loader walks whichever
// dimension is contiguous.  Fixed summation order (k ascending), so results are run-to-run reproducible.
struct Operand {
    const float* p; long long s_mn, s_k;      // element (mn, k) at p[mn * s_mn + k * s_k]
    int n_mn;                                  // rows (A) / columns (B) that exist: the rest reads as zero
    int gate_stride;                           // != 0: column mn stands for gate (mn & 3) of unit (mn >> 2), at offset
                                               // (mn & 3) * gate_stride + (mn >> 2) (the four gates of a unit side by side)
};
__device__ __forceinline__ long long operand_mn_offset(const Operand& o, int mn) {
    return o.gate_stride ? (long long)(mn & 3) * o.gate_stride + (mn >> 2) : (long long)mn * o.s_mn;
}

template <int TM, int TN>
struct TileGemm {
    static constexpr int BM = 16 * TM, BN = 16 * TN, KS = 16;
    static constexpr int LA = BM * KS / 256, LB = BN * KS / 256;      // elements per thread and slice
    float (*As)[BM + 4];
    float (*Bs)[BN + 4];
    float ra[LA], rb[LB];

    template <int BT, int L>
    __device__ __forceinline__ void fetch(const Operand& o, int mn0, int k0, int k_hi, float (&r)[L]) {
        const bool kfast = o.s_k == 1;
#pragma unroll
        for (int q = 0; q < L; ++q) {
            const int e = threadIdx.x + q * 256;
            int kk, mm;
            if (kfast) { kk = e % KS; mm = e / KS; } else { mm = e % BT; kk = e / BT; }
            const int mn = mn0 + mm, k = k0 + kk;
            r[q] = (mn < o.n_mn && k < k_hi) ? __ldg(o.p + operand_mn_offset(o, mn) + (long long)k * o.s_k) : 0.f;
        }
    }
    template <int BT, int L>
    __device__ __forceinline__ void stash(const Operand& o, const float (&r)[L], float (*S)[BT + 4]) {
        const bool kfast = o.s_k == 1;
#pragma unroll
        for (int q = 0; q < L; ++q) {
            const int e = threadIdx.x + q * 256;
            int kk, mm;
            if (kfast) { kk = e % KS; mm = e / KS; } else { mm = e % BT; kk = e / BT; }
            S[kk][mm] = r[q];
        }
    }
    __device__ __forceinline__ void run(const Operand& A, const Operand& B, int m0, int n0, int k_lo, int k_hi,
                                        float (&acc)[TM][TN]) {
        const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
        if (k_lo >= k_hi) return;
        fetch<BM, LA>(A, m0, k_lo, k_hi, ra);
        fetch<BN, LB>(B, n0, k_lo, k_hi, rb);
        for (int k0 = k_lo; k0 < k_hi; k0 += KS) {
            __syncthreads();                                   // the previous slice has been read out of shared memory
            stash<BM, LA>(A, ra, As);
            stash<BN, LB>(B, rb, Bs);
            __syncthreads();
            if (k0 + KS < k_hi) {
                fetch<BM, LA>(A, m0, k0 + KS, k_hi, ra);
                fetch<BN, LB>(B, n0, k0 + KS, k_hi, rb);
            }
#pragma unroll
            for (int kk = 0; kk < KS; ++kk) {
                float av[TM], bv[TN];
                if constexpr (TM % 4 == 0) {
#pragma unroll
                    for (int i = 0; i < TM; i += 4) *reinterpret_cast<float4*>(&av[i]) = *reinterpret_cast<const float4*>(&As[kk][ty * TM + i]);
                } else {
#pragma unroll
                    for (int i = 0; i < TM; i += 2) *reinterpret_cast<float2*>(&av[i]) = *reinterpret_cast<const float2*>(&As[kk][ty * TM + i]);
                }
                if constexpr (TN % 4 == 0) {
#pragma unroll
                    for (int j = 0; j < TN; j += 4) *reinterpret_cast<float4*>(&bv[j]) = *reinterpret_cast<const float4*>(&Bs[kk][tx * TN + j]);
                } else {
#pragma unroll
                    for (int j = 0; j < TN; j += 2) *reinterpret_cast<float2*>(&bv[j]) = *reinterpret_cast<const float2*>(&Bs[kk][tx * TN + j]);
                }
#pragma unroll
                for (int i = 0; i < TM; ++i)
#pragma unroll
                    for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
            }
        }
    }
};

// ---- generic GEMM launch: C[M x N] = A * B (+ bias[N]); blockIdx.z = K split (partials [z][M][N], summed by splitk_sum_kernel) ---
struct GemmArgs {
    Operand A, B;                              // A.n_mn = M, B.n_mn = N
    float* C; long long c_m;                   // C(m, n) = C[m * c_m + n]   (split z: C + z * M * c_m)
    const float* bias;                        // optional, per column (only without K split)
    int M, N, K, k_per_split;
};

template <int TM, int TN>
__global__ void __launch_bounds__(256)
gemm_kernel(GemmArgs g) {
    using TG = TileGemm<TM, TN>;
    __shared__ __align__(16) float As[TG::KS][TG::BM + 4];
    __shared__ __align__(16) float Bs[TG::KS][TG::BN + 4];
    TG tg;
    tg.As = As; tg.Bs = Bs;
    const int m0 = blockIdx.y * TG::BM, n0 = blockIdx.x * TG::BN;
    const int k_lo = blockIdx.z * g.k_per_split, k_hi = min(g.K, k_lo + g.k_per_split);
    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
    tg.run(g.A, g.B, m0, n0, k_lo, k_hi, acc);
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    float* C = g.C + (size_t)blockIdx.z * g.M * g.c_m;
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int m = m0 + ty * TM + i;
        if (m >= g.M) continue;
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int n = n0 + tx * TN + j;
            if (n < g.N) C[(long long)m * g.c_m + n] = acc[i][j] + (g.bias ? g.bias[n] : 0.f);
        }
    }
}

// out[e] = sum over the K splits, in split order (fixed summation order)
__global__ void __launch_bounds__(256)
splitk_sum_kernel(const float* __restrict__ part, int nsplit, long long n, float* __restrict__ out) {
    const long long i = ((long long)blockIdx.x * 256 + threadIdx.x) * 4;
    if (i >= n) return;
    float4 s = *reinterpret_cast<const float4*>(part + i);
    for (int z = 1; z < nsplit; ++z) {
        const float4 v = *reinterpret_cast<const float4*>(part + (long long)z * n + i);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    *reinterpret_cast<float4*>(out + i) = s;
}

// Programmatic dependent launch for the chains of per-step kernels: step t + 1 is launched while step t still runs and waits
// (griddepcontrol.wait, first thing in the kernel) until step t has completed, so the launch latency between two dependent
// steps (~3 us, 64 times per training step) overlaps with the step before.
__device__ __forceinline__ void pdl_trigger_then_wait() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
template <class... KArgs, class... Args>
cudaError_t launch_step_pdl(void (*kernel)(KArgs...), dim3 grid, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(256);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ---- tile GEMM core for the per-step kernels: acc[TM][TN] += sum_k A[m][k] * B[n][k], both operands k-contiguous ---------
// A time step's GEMM is tiny (134 MFLOP over the whole GPU) and strictly ordered after the previous step, so what it costs is
// memory round trips: with one K slice in flight a block paid one per slice (K = 1024: 64 of them, 55 us per step).  Here the
// slices stream through NS shared-memory stages with 16-byte cp.async copies, NS - 1 slices in flight per block.
// Thread (ty, tx) owns rows ty*TM.. of A and rows tx, tx + 16, .. of B (acc[i][j] = row ty*TM + i x row tx + 16 j).
// Rows past n_rows read as zero.  K is a multiple of 16, row strides are multiples of 4 floats.  Fixed summation order.
__device__ __forceinline__ void cp_async_16(void* dst_smem, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int TM, int TN, int NS>
struct TileGemmNT {
    static constexpr int BM = 16 * TM, BN = 16 * TN, KS = 16, LD = KS + 4;
    static constexpr size_t kSmemFloats = (size_t)NS * (BM + BN) * LD;
    float* smem;                                                          // [NS][BM + BN][LD]
    __device__ __forceinline__ float* a_row(int stage, int m) { return smem + ((size_t)stage * (BM + BN) + m) * LD; }
    __device__ __forceinline__ float* b_row(int stage, int n) { return smem + ((size_t)stage * (BM + BN) + BM + n) * LD; }
    __device__ __forceinline__ void issue(const float* A, long long lda, int a_rows, int m0, const float* Bm, long long ldb,
                                          int b_rows, int n0, int k0, int stage) {
        for (int c = threadIdx.x; c < (BM + BN) * 4; c += 256) {
            const int r = c >> 2, k4 = (c & 3) * 4;
            float* dst = smem + ((size_t)stage * (BM + BN) + r) * LD + k4;
            const bool isA = r < BM;
            const int row = isA ? m0 + r : n0 + r - BM;
            const bool ok = row < (isA ? a_rows : b_rows);
            const float* src = isA ? A + (long long)row * lda + k0 + k4 : Bm + (long long)row * ldb + k0 + k4;
            if (ok) cp_async_16(dst, src);
            else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    __device__ __forceinline__ void run(const float* A, long long lda, int a_rows, int m0, const float* Bm, long long ldb,
                                        int b_rows, int n0, int K, float (&acc)[TM][TN]) {
        const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
        const int nslice = K / KS;
#pragma unroll
        for (int s = 0; s < NS - 1; ++s) {
            if (s < nslice) issue(A, lda, a_rows, m0, Bm, ldb, b_rows, n0, s * KS, s);
            cp_async_commit();
        }
        for (int s = 0; s < nslice; ++s) {
            cp_async_wait<NS - 2>();                 // this thread's copies of slice s have landed
            __syncthreads();                         // ... everybody's, and everybody has finished multiplying slice s - 1
            if (s + NS - 1 < nslice) issue(A, lda, a_rows, m0, Bm, ldb, b_rows, n0, (s + NS - 1) * KS, (s + NS - 1) % NS);
            cp_async_commit();
            const int st = s % NS;
#pragma unroll
            for (int kq = 0; kq < KS; kq += 4) {
                float4 a4[TM], b4[TN];
#pragma unroll
                for (int i = 0; i < TM; ++i) a4[i] = *reinterpret_cast<const float4*>(a_row(st, ty * TM + i) + kq);
#pragma unroll
                for (int j = 0; j < TN; ++j) b4[j] = *reinterpret_cast<const float4*>(b_row(st, tx + 16 * j) + kq);   // (rows 20 floats apart: 8 lanes, 8 bank groups)
#pragma unroll
                for (int i = 0; i < TM; ++i)
#pragma unroll
                    for (int j = 0; j < TN; ++j) {
                        float v = acc[i][j];
                        v = fmaf(a4[i].x, b4[j].x, v); v = fmaf(a4[i].y, b4[j].y, v);
                        v = fmaf(a4[i].z, b4[j].z, v); v = fmaf(a4[i].w, b4[j].w, v);
                        acc[i][j] = v;
                    }
            }
        }
        cp_async_wait<0>();
    }
};

// ---- the same product with K split over the block's 8 warps ----------------------------------------------------------------
// out[BM x BN] = sum_k A[m][k] * B[n][k] for one BM-row x BN-column tile (BM = 16 or 32, BN = 32 or 64); warp w multiplies k in
// [w K/8, (w+1) K/8) for the WHOLE tile, streaming its operands through its own ring of shared-memory stages, then the 8
// partial tiles are added in warp order (fixed summation order).  A block's serial chain is K/8 deep instead of K: the
// step kernels are latency-bound (one block per SM, 64-128 blocks per step), so this is what their time follows.
// Lane (ly = lane / 4, lx = lane % 4) owns rows ly + 8 i (i < BM / 8) and columns lx + 4 j (j < BN / 4): with rows 20 floats apart
// the 8 (4) distinct rows of a 128-bit load fall into distinct bank groups, equal addresses broadcast.
template <int BM, int BN, int NSW>
struct WarpSplitGemmNT {
    static constexpr int KS = 16, LD = KS + 4, TI = BM / 8, TJ = BN / 4;
    static constexpr int kStageFloats = (BM + BN) * LD;
    static constexpr size_t kSmemBytes = sizeof(float) * (size_t)8 * NSW * kStageFloats;      // >= 8 * BM * BN floats for the partial tiles
    // returns with part[w][m][n] (w < 8) in `smem` (layout [8][BM][BN]); the caller sums over w
    __device__ __forceinline__ void run(float* smem, const float* A, long long lda, int a_rows, int m0, const float* Bm, long long ldb,
                                        int b_rows, int n0, int K) {
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, ly = lane >> 2, lx = lane & 3;
        float* ws = smem + (size_t)warp * NSW * kStageFloats;
        const int kw = K / 8, k_lo = warp * kw, nslice = kw / KS;
        auto issue = [&](int slice, int stage) {
            float* base = ws + (size_t)stage * kStageFloats;
            const int k0 = k_lo + slice * KS;
#pragma unroll
            for (int q = 0; q < (BM + BN) * 4 / 32; ++q) {
                const int c = lane + q * 32, r = c >> 2, k4 = (c & 3) * 4;
                const bool isA = r < BM;
                const int row = isA ? m0 + r : n0 + r - BM;
                float* dst = base + r * LD + k4;
                if (row < (isA ? a_rows : b_rows)) cp_async_16(dst, isA ? A + (long long)row * lda + k0 + k4 : Bm + (long long)row * ldb + k0 + k4);
                else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        };
        float acc[TI][TJ];
#pragma unroll
        for (int i = 0; i < TI; ++i)
#pragma unroll
            for (int j = 0; j < TJ; ++j) acc[i][j] = 0.f;
#pragma unroll
        for (int s = 0; s < NSW - 1; ++s) {
            if (s < nslice) issue(s, s);
            cp_async_commit();
        }
        for (int s = 0; s < nslice; ++s) {
            cp_async_wait<NSW - 2>();
            __syncwarp();                            // the whole warp's copies of slice s have landed; slice s - 1 has been multiplied
            if (s + NSW - 1 < nslice) issue(s + NSW - 1, (s + NSW - 1) % NSW);
            cp_async_commit();
            const float* base = ws + (size_t)(s % NSW) * kStageFloats;
#pragma unroll
            for (int kq = 0; kq < KS; kq += 4) {
                float4 a4[TI], b4[TJ];
#pragma unroll
                for (int i = 0; i < TI; ++i) a4[i] = *reinterpret_cast<const float4*>(base + (ly + 8 * i) * LD + kq);
#pragma unroll
                for (int j = 0; j < TJ; ++j) b4[j] = *reinterpret_cast<const float4*>(base + (BM + lx + 4 * j) * LD + kq);
#pragma unroll
                for (int i = 0; i < TI; ++i)
#pragma unroll
                    for (int j = 0; j < TJ; ++j) {
                        float v = acc[i][j];
                        v = fmaf(a4[i].x, b4[j].x, v); v = fmaf(a4[i].y, b4[j].y, v);
                        v = fmaf(a4[i].z, b4[j].z, v); v = fmaf(a4[i].w, b4[j].w, v);
                        acc[i][j] = v;
                    }
            }
        }
        cp_async_wait<0>();
        __syncthreads();                             // every warp is done with its stages: they become the partial tiles
        float* part = smem + (size_t)warp * BM * BN;
#pragma unroll
        for (int i = 0; i < TI; ++i)
#pragma unroll
            for (int j = 0; j < TJ; ++j) part[(ly + 8 * i) * BN + lx + 4 * j] = acc[i][j];
        __syncthreads();
    }
    // sum over the 8 warps' partials of element (m, n), in warp order
    __device__ __forceinline__ static float total(const float* smem, int m, int n) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) s += smem[(size_t)w * BM * BN + m * BN + n];
        return s;
    }
};

// Wh [256 k][4 gates x 256 units] -> WhR [(unit / 16) * 64 + gate * 16 + unit % 16][256 k]: the forward step's B operand,
// k-contiguous; a block's 64 rows are the four gates of its 16 units, gate-major, so that thread tx owns rows tx + 16 * gate.
// 32 x 32 tile transpose per gate.
__global__ void __launch_bounds__(256)
lstm_wh_reorder_kernel(const float* __restrict__ Wh, float* __restrict__ WhR) {
    __shared__ float tile[32][33];
    const int g = blockIdx.z, u0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) tile[r][tx] = Wh[(size_t)(k0 + r) * kG + g * kH + u0 + tx];
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int u = u0 + r;
        WhR[((size_t)(u >> 4) * 64 + g * 16 + (u & 15)) * kH + k0 + tx] = tile[tx][r];
    }
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// ---- one time step of the recurrence, all paths ------------------------------------------------------------------------
// block = 32 paths x 16 cell units (x 4 gates, side by side as GEMM columns); thread (ty, tx): paths 2*ty, 2*ty+1, unit u0 + tx.
// gates = gx[t] + h_{t-1} @ Wh;  writes the activated gates (i, g = tanh(j), f, o), c_t, tanh(c_t), h_t and h_{t-1}.
__global__ void __launch_bounds__(256)
lstm_step_fwd_kernel(const float* __restrict__ gx, const float* __restrict__ WhR, const float* __restrict__ state_in,
                     int P, int B, int t, float* __restrict__ act, float* __restrict__ cbuf, float* __restrict__ tcbuf,
                     float* __restrict__ hout, float* __restrict__ hprev_buf, float* __restrict__ state_out) {
    using WG = WarpSplitGemmNT<32, 64, 3>;
    extern __shared__ __align__(16) float tile_s[];
    pdl_trigger_then_wait();
    const int p0 = blockIdx.y * 32, u0 = blockIdx.x * 16;
    const bool have_h = t > 0 || state_in;
    if (have_h) {
        // A = h_{t-1}: the previous step's outputs, or the bound initial state (c, h per path);
        // B = WhR: row u0 * 4 + 16 * gate + ul = gate `gate` of unit u0 + ul
        const float* A = t > 0 ? hout + (size_t)(t - 1) * kH : state_in + kH;
        WG wg;
        wg.run(tile_s, A, t > 0 ? (long long)B * kH : 2LL * kH, P, p0, WhR, kH, kG, u0 * 4, kH);
    }
    const int ul = threadIdx.x & 15, u = u0 + ul;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int m = (threadIdx.x >> 4) + 16 * i, p = p0 + m;
        if (p >= P) continue;
        const size_t row = (size_t)p * B + t;
        const float* gr = gx + row * kG;
        float ai = gr[u], aj = gr[kH + u], af = gr[2 * kH + u], ao = gr[3 * kH + u];
        if (have_h) {
            ai += WG::total(tile_s, m, ul); aj += WG::total(tile_s, m, 16 + ul);
            af += WG::total(tile_s, m, 32 + ul); ao += WG::total(tile_s, m, 48 + ul);
        }
        const float c_prev = t == 0 ? (state_in ? state_in[(size_t)p * 2 * kH + u] : 0.f) : cbuf[(row - 1) * kH + u];
        const float h_prev = t == 0 ? (state_in ? state_in[((size_t)p * 2 + 1) * kH + u] : 0.f) : hout[(row - 1) * kH + u];
        const float ig = sigmoidf_(ai), gg = tanhf(aj), fg = sigmoidf_(af + 1.0f), og = sigmoidf_(ao);   // forget_bias = 1.0
        const float c = c_prev * fg + ig * gg;
        const float tc = tanhf(c);
        const float hnew = tc * og;
        float* a = act + row * kG;
        a[u] = ig; a[kH + u] = gg; a[2 * kH + u] = fg; a[3 * kH + u] = og;
        cbuf[row * kH + u] = c;
        tcbuf[row * kH + u] = tc;
        hout[row * kH + u] = hnew;
        hprev_buf[row * kH + u] = h_prev;
        if (state_out && t == B - 1) {
            state_out[(size_t)p * 2 * kH + u] = c;
            state_out[((size_t)p * 2 + 1) * kH + u] = hnew;
        }
    }
}

// ---- one step of BPTT, all paths: dA_t (gradient at the gate pre-activations) ----------------------------------------
// dh_t = dh_out[t] + dA_{t+1} @ Wh^T;  block = 16 paths x 32 units (128 blocks at 256 paths: every SM gets one), K = 1024 gate columns.
__global__ void __launch_bounds__(256)
lstm_step_bwd_kernel(const float* __restrict__ dh_out, const float* __restrict__ Wh, const float* __restrict__ act,
                     const float* __restrict__ cbuf, const float* __restrict__ tcbuf, const float* __restrict__ state_in,
                     int P, int B, int t, float* __restrict__ dA, float* __restrict__ dc_carry) {
    using WG = WarpSplitGemmNT<16, 32, 4>;
    extern __shared__ __align__(16) float tile_s[];
    pdl_trigger_then_wait();
    const int p0 = blockIdx.y * 16, u0 = blockIdx.x * 32;
    // dA_{t+1}[p][k] x Wh^T: (k = gate column, n = unit) = Wh[n][k], both k-contiguous
    if (t < B - 1) {
        WG wg;
        wg.run(tile_s, dA + (size_t)(t + 1) * kG, (long long)B * kG, P, p0, Wh, kG, kH, u0, kG);
    }
    {
        const int n = threadIdx.x & 31, u = u0 + n;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int m = (threadIdx.x >> 5) + 8 * i, p = p0 + m;
            if (p >= P) continue;
            const size_t row = (size_t)p * B + t;
            const float dh = dh_out[row * kH + u] + (t < B - 1 ? WG::total(tile_s, m, n) : 0.f);
            const float* a = act + row * kG;
            const float ig = a[u], gg = a[kH + u], fg = a[2 * kH + u], og = a[3 * kH + u];
            const float tc = tcbuf[row * kH + u];
            const float c_prev = t == 0 ? (state_in ? state_in[(size_t)p * 2 * kH + u] : 0.f) : cbuf[(row - 1) * kH + u];
            const float dcn = t < B - 1 ? dc_carry[(size_t)p * kH + u] : 0.f;
            const float dc = dh * og * (1.0f - tc * tc) + dcn;
            float* d = dA + row * kG;
            d[u] = dc * gg * ig * (1.0f - ig);                     // i
            d[kH + u] = dc * ig * (1.0f - gg * gg);                // j (tanh)
            d[2 * kH + u] = dc * c_prev * fg * (1.0f - fg);        // f
            d[3 * kH + u] = dh * tc * og * (1.0f - og);            // o
            dc_carry[(size_t)p * kH + u] = dc * fg;
        }
    }
}

// column sums of dA [T][1024] in a fixed order, two levels: block (x, y) adds rows y, y + gridDim.y, ... of 32 columns in 8 row
// groups (then the 8 partials in order) into part[y][c]; colsum_finish_kernel adds the gridDim.y partials in order.
constexpr int kColsumChunks = 32;
__global__ void __launch_bounds__(256)
colsum_kernel(const float* __restrict__ dA, int T, float* __restrict__ part) {
    __shared__ float red[8][33];
    const int c = blockIdx.x * 32 + (threadIdx.x & 31), g = threadIdx.x >> 5;
    float s = 0.f;
    for (int r = blockIdx.y * 8 + g; r < T; r += 8 * gridDim.y) s += dA[(size_t)r * kG + c];
    red[g][threadIdx.x & 31] = s;
    __syncthreads();
    if (g == 0) {
        float tot = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) tot += red[q][threadIdx.x & 31];
        part[(size_t)blockIdx.y * kG + c] = tot;
    }
}
__global__ void __launch_bounds__(256)
colsum_finish_kernel(const float* __restrict__ part, int nchunks, float* __restrict__ out) {
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c >= kG) return;
    float s = 0.f;
    for (int y = 0; y < nchunks; ++y) s += part[(size_t)y * kG + c];
    out[c] = s;
}

// C = A * B (+ bias).  Big tiles (128 x 128, 8 x 8 per thread) when the output has enough of them to fill the GPU, else
// 64 x 64; nsplit > 1 cuts K into equal pieces whose partial products go to `partials` and are added in split order.
int gemm(pn_handle* h, GemmArgs g, int nsplit, float* partials, cudaStream_t st, const char* what) {
    float* out = g.C;
    g.k_per_split = ((g.K + nsplit - 1) / nsplit + 15) & ~15;
    nsplit = (g.K + g.k_per_split - 1) / g.k_per_split;
    if (nsplit > 1) { g.C = partials; g.c_m = g.N; g.bias = nullptr; }
    const long long big_tiles = (long long)((g.M + 127) / 128) * ((g.N + 127) / 128) * nsplit;
    if (big_tiles >= h->num_sms) {
        dim3 grid((g.N + 127) / 128, (g.M + 127) / 128, nsplit);
        gemm_kernel<8, 8><<<grid, 256, 0, st>>>(g);
    } else {
        dim3 grid((g.N + 63) / 64, (g.M + 63) / 64, nsplit);
        gemm_kernel<4, 4><<<grid, 256, 0, st>>>(g);
    }
    PN_LAUNCH_CHECK(h, what);
    if (nsplit > 1) {
        const long long n = (long long)g.M * g.N;
        splitk_sum_kernel<<<(int)((n / 4 + 255) / 256), 256, 0, st>>>(partials, nsplit, n, out);
        PN_LAUNCH_CHECK(h, "splitk_sum_kernel");
    }
    return PN_OK;
}

}  // namespace

constexpr int kLstmSplitK = 16;

namespace {
struct LstmPlanes { uint8_t *x, *hp, *da, *wx; };
LstmPlanes lstm_planes(const pn_handle* h) {
    const size_t T = (size_t)h->T;
    float* base = h->lstm_ws + T * kG * 3 + T * kH * 4 + (size_t)h->P * kH + (size_t)kLstmSplitK * kH * kG + (size_t)kColsumChunks * kG +
                  (size_t)kH * kG;
    LstmPlanes p;
    p.x = reinterpret_cast<uint8_t*>(base);
    p.hp = p.x + pn_gemm_planes_bytes((int)T, kH);
    p.da = p.hp + pn_gemm_planes_bytes((int)T, kH);
    p.wx = p.da + pn_gemm_planes_bytes((int)T, kG);
    return p;
}
}  // namespace

size_t pn_lstm_workspace_floats(const pn_handle* h) {
    // gx [T,1024], act [T,1024], dA [T,1024], c, tanh c, h_out, h_prev [T,256] each, dc carry [P,256],
    // split-K partials of a weight-gradient GEMM [16][256][1024], column-sum partials [32][1024]
    const size_t T = (size_t)h->T;
    return T * kG * 3 + T * kH * 4 + (size_t)h->P * kH + (size_t)kLstmSplitK * kH * kG + (size_t)kColsumChunks * kG +
           (size_t)kH * kG +                         // + the recurrent weights reordered for the forward step (WhR)
           // + bf16 hi/lo block planes of X, Hprev [T,256], dA [T,1024] and the input half of the kernel (pn_gemm_tc.cu)
           (pn_gemm_planes_bytes((int)T, kH) * 2 + pn_gemm_planes_bytes((int)T, kG) + pn_gemm_planes_bytes(kH, kG)) / sizeof(float);
}

// hbuf (module sum, no /M) -> lstm_out; reads the bound initial state, writes the final state if one is bound
int pn_launch_lstm_fwd(pn_handle* h, int B, cudaStream_t st) {
    const PnNetDesc& nd = h->nd;
    if (nd.hidden != kH) return pn_fail(h, PN_ERR_STATE, "the LSTM variant needs hidden = 256 (BasicLSTMCell(256))");
    const int P = h->P, T = P * B;
    const float* W = h->params + nd.lstmW;
    float* gx = h->lstm_ws;
    float* act = gx + (size_t)h->T * kG;
    float* cbuf = act + 2 * (size_t)h->T * kG;       // (dA sits between act and cbuf)
    float* tcbuf = cbuf + (size_t)h->T * kH;
    float* hout = tcbuf + (size_t)h->T * kH;
    float* hprev = hout + (size_t)h->T * kH;
    PnProfScope prof(h, PN_PROF_LSTM_FWD, st);
    // rows 0..255 of the kernel multiply the input: gx = X @ W[0:256] + b
    const LstmPlanes pl = lstm_planes(h);
    // the planes of X are made whenever the weight-gradient GEMM of this step could use them, too
    if (int rc = pn_gemm_split(h, h->hbuf, T, kH, kH, pl.x, st)) return rc;
    if (int rc = pn_gemm_split(h, W, kH, kG, kG, pl.wx, st)) return rc;
    if (int rc = pn_gemm_tc(h, PnGemmOperand{pl.x, T, kH, false}, PnGemmOperand{pl.wx, kH, kG, true}, gx, kG, h->params + nd.lstmb, T,
                            kG, kH, 1, st))
        return rc;
    const float* Wh = W + (size_t)kH * kG;
    float* WhR = hprev + (size_t)h->T * kH + (size_t)h->P * kH + (size_t)kLstmSplitK * kH * kG + (size_t)kColsumChunks * kG;
    lstm_wh_reorder_kernel<<<dim3(kH / 32, kH / 32, 4), 256, 0, st>>>(Wh, WhR);
    PN_LAUNCH_CHECK(h, "lstm_wh_reorder_kernel");
    dim3 grid(kH / 16, (P + 31) / 32);
    const size_t smem = WarpSplitGemmNT<32, 64, 3>::kSmemBytes;
    PN_SET_SMEM(h, lstm_step_fwd_kernel, smem);
    for (int t = 0; t < B; ++t) {
        cudaError_t e = launch_step_pdl(lstm_step_fwd_kernel, grid, smem, st, gx, WhR, h->lstm_state_in, P, B, t, act, cbuf, tcbuf, hout,
                                        hprev, h->lstm_state_out);
        if (e != cudaSuccess) return pn_fail(h, PN_ERR_CUDA, "launch of lstm_step_fwd_kernel failed: %s", cudaGetErrorString(e));
        PN_LAUNCH_CHECK(h, "lstm_step_fwd_kernel");
    }
    return PN_OK;
}

// dh (gradient at the cell outputs, written by the head kernel) -> dh (gradient at the module sum), W_lstm / b_lstm gradients
int pn_launch_lstm_bwd(pn_handle* h, int B, cudaStream_t st) {
    const PnNetDesc& nd = h->nd;
    const int P = h->P, T = P * B;
    const float* W = h->params + nd.lstmW;
    const float* Wh = W + (size_t)kH * kG;
    float* gx = h->lstm_ws;
    float* act = gx + (size_t)h->T * kG;
    float* dA = act + (size_t)h->T * kG;
    float* cbuf = dA + (size_t)h->T * kG;
    float* tcbuf = cbuf + (size_t)h->T * kH;
    float* hout = tcbuf + (size_t)h->T * kH;
    float* hprev = hout + (size_t)h->T * kH;
    float* dcc = hprev + (size_t)h->T * kH;
    float* partials = dcc + (size_t)h->P * kH;
    float* colpart = partials + (size_t)kLstmSplitK * kH * kG;
    (void)hout;
    PnProfScope prof(h, PN_PROF_LSTM_BWD, st);
    dim3 grid(kH / 32, (P + 15) / 16);
    const size_t smem = WarpSplitGemmNT<16, 32, 4>::kSmemBytes;
    PN_SET_SMEM(h, lstm_step_bwd_kernel, smem);
    for (int t = B - 1; t >= 0; --t) {
        cudaError_t e = launch_step_pdl(lstm_step_bwd_kernel, grid, smem, st, h->dh, Wh, act, cbuf, tcbuf, h->lstm_state_in, P, B, t, dA,
                                        dcc);
        if (e != cudaSuccess) return pn_fail(h, PN_ERR_CUDA, "launch of lstm_step_bwd_kernel failed: %s", cudaGetErrorString(e));
        PN_LAUNCH_CHECK(h, "lstm_step_bwd_kernel");
    }
    float* gW = h->grads + nd.lstmW;
    // dW rows 0..255 = X^T dA, rows 256..511 = Hprev^T dA (sum over every transition of every path): K = T is cut into
    // pieces so that the 8 output tiles become 128 blocks
    const LstmPlanes pl = lstm_planes(h);
    if (int rc = pn_gemm_split(h, dA, T, kG, kG, pl.da, st)) return rc;
    int nsplit = 0;                                 // tensor-core weight gradient: K = T in pieces of whole 64-deep chunks
    for (int s2 = kLstmSplitK; s2 >= 1; --s2)
        if (T % (64 * s2) == 0 && 16 * s2 <= 2 * h->num_sms) { nsplit = s2; break; }
    if (nsplit) {
        if (int rc = pn_gemm_split(h, hprev, T, kH, kH, pl.hp, st)) return rc;
        const uint8_t* a_planes[2] = {pl.x, pl.hp};
        for (int half = 0; half < 2; ++half) {
            float* out = gW + (size_t)half * kH * kG;
            if (int rc = pn_gemm_tc(h, PnGemmOperand{a_planes[half], T, kH, true}, PnGemmOperand{pl.da, T, kG, true},
                                    nsplit > 1 ? partials : out, kG, nullptr, kH, kG, T, nsplit, st))
                return rc;
            if (nsplit > 1) {
                const long long n = (long long)kH * kG;
                splitk_sum_kernel<<<(int)((n / 4 + 255) / 256), 256, 0, st>>>(partials, nsplit, n, out);
                PN_LAUNCH_CHECK(h, "splitk_sum_kernel");
            }
        }
    } else {
        GemmArgs gw1{{h->hbuf, 1, kH, kH, 0}, {dA, 1, kG, kG, 0}, gW, kG, nullptr, kH, kG, T, 0};
        if (int rc = gemm(h, gw1, kLstmSplitK, partials, st, "lstm weight gradient (input rows)")) return rc;
        GemmArgs gw2{{hprev, 1, kH, kH, 0}, {dA, 1, kG, kG, 0}, gW + (size_t)kH * kG, kG, nullptr, kH, kG, T, 0};
        if (int rc = gemm(h, gw2, kLstmSplitK, partials, st, "lstm weight gradient (recurrent rows)")) return rc;
    }
    colsum_kernel<<<dim3(kG / 32, kColsumChunks), 256, 0, st>>>(dA, T, colpart);
    PN_LAUNCH_CHECK(h, "colsum_kernel");
    colsum_finish_kernel<<<kG / 256, 256, 0, st>>>(colpart, kColsumChunks, h->grads + nd.lstmb);
    PN_LAUNCH_CHECK(h, "colsum_finish_kernel");
    // gradient at the cell input = dA @ Wx^T, overwriting dh (every step has consumed its dh_out by now)
    return pn_gemm_tc(h, PnGemmOperand{pl.da, T, kG, false}, PnGemmOperand{pl.wx, kH, kG, false}, h->dh, kH, nullptr, T, kH, kG, 1, st);
}
