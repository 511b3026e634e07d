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

// ---- generic fp32 GEMM: C[M x N] = A[M x K] * B[K x N] (+ bias[N]) with arbitrary element strides -------------------
// 64 x 64 tile, 16-deep K slices, 256 threads, 4 x 4 outputs per thread.
struct GemmArgs {
    const float* A; long long a_m, a_k;       // A(m, k) = A[m * a_m + k * a_k]
    const float* B; long long b_k, b_n;       // B(k, n) = B[k * b_k + n * b_n]
    float* C; long long c_m;                  // C(m, n) = C[m * c_m + n]
    const float* bias;                        // optional, per column
    int M, N, K;
};

__global__ void __launch_bounds__(256)
gemm_kernel(GemmArgs g) {
    __shared__ float As[16][64 + 4];
    __shared__ float Bs[16][64 + 4];
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    // loaders: choose the thread mapping so that the contiguous dimension of each operand is the fast one
    const bool a_kfast = g.a_k == 1, b_nfast = g.b_n == 1;
    for (int k0 = 0; k0 < g.K; k0 += 16) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int e = threadIdx.x + q * 256;                       // 1024 elements per tile
            int mm, kk;
            if (a_kfast) { kk = e & 15; mm = e >> 4; } else { mm = e & 63; kk = e >> 6; }
            const int m = m0 + mm, k = k0 + kk;
            As[kk][mm] = (m < g.M && k < g.K) ? g.A[m * g.a_m + k * g.a_k] : 0.f;
            int nn, kb;
            if (b_nfast) { nn = e & 63; kb = e >> 6; } else { kb = e & 15; nn = e >> 4; }
            const int n = n0 + nn, k2 = k0 + kb;
            Bs[kb][nn] = (n < g.N && k2 < g.K) ? g.B[k2 * g.b_k + n * g.b_n] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= g.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n < g.N) g.C[m * g.c_m + n] = acc[i][j] + (g.bias ? g.bias[n] : 0.f);
        }
    }
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// ---- one time step of the recurrence, all paths ------------------------------------------------------------------------
// block = 64 paths x 16 cell units (x 4 gates); thread (ty, tx): paths 4*ty..4*ty+3, unit u0 + tx.
// gates = gx[t] + h_{t-1} @ Wh;  writes the activated gates (i, g = tanh(j), f, o), c_t, tanh(c_t), h_t and h_{t-1}.
__global__ void __launch_bounds__(256)
lstm_step_fwd_kernel(const float* __restrict__ gx, const float* __restrict__ Wh, const float* __restrict__ state_in,
                     int P, int B, int t, float* __restrict__ act, float* __restrict__ cbuf, float* __restrict__ tcbuf,
                     float* __restrict__ hout, float* __restrict__ hprev_buf, float* __restrict__ state_out) {
    __shared__ float Hs[16][64 + 4];
    __shared__ float Ws[16][64 + 4];           // [k][gate * 16 + unit]
    const int p0 = blockIdx.y * 64, u0 = blockIdx.x * 16;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    float acc[4][4];                            // [path][gate]
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int k0 = 0; k0 < kH; k0 += 16) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int e = threadIdx.x + q * 256;
            const int kk = e & 15, pp = e >> 4, p = p0 + pp;
            float hv = 0.f;
            if (p < P) {
                if (t == 0) hv = state_in ? state_in[((size_t)p * 2 + 1) * kH + k0 + kk] : 0.f;
                else hv = hout[((size_t)p * B + t - 1) * kH + k0 + kk];
            }
            Hs[kk][pp] = hv;
            const int nn = e & 63, kb = e >> 6;
            Ws[kb][nn] = Wh[(size_t)(k0 + kb) * kG + (nn >> 4) * kH + u0 + (nn & 15)];
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            const float4 a = *reinterpret_cast<const float4*>(&Hs[kk][ty * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w};
            const float w[4] = {Ws[kk][tx], Ws[kk][16 + tx], Ws[kk][32 + tx], Ws[kk][48 + tx]};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], w[j], acc[i][j]);
        }
        __syncthreads();
    }
    const int u = u0 + tx;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int p = p0 + ty * 4 + i;
        if (p >= P) continue;
        const size_t row = (size_t)p * B + t;
        const float* gr = gx + row * kG;
        const float ai = acc[i][0] + gr[u], aj = acc[i][1] + gr[kH + u], af = acc[i][2] + gr[2 * kH + u],
                    ao = acc[i][3] + gr[3 * kH + u];
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
// dh_t = dh_out[t] + dA_{t+1} @ Wh^T;  block = 64 paths x 64 units, K = 1024 gate columns.
__global__ void __launch_bounds__(256)
lstm_step_bwd_kernel(const float* __restrict__ dh_out, const float* __restrict__ Wh, const float* __restrict__ act,
                     const float* __restrict__ cbuf, const float* __restrict__ tcbuf, const float* __restrict__ state_in,
                     int P, int B, int t, float* __restrict__ dA, float* __restrict__ dc_carry) {
    __shared__ float As[16][64 + 4];
    __shared__ float Bs[16][64 + 4];
    const int p0 = blockIdx.y * 64, u0 = blockIdx.x * 64;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    if (t < B - 1) {
        for (int k0 = 0; k0 < kG; k0 += 16) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int e = threadIdx.x + q * 256;
                const int kk = e & 15, mm = e >> 4, p = p0 + mm;
                As[kk][mm] = p < P ? dA[((size_t)p * B + t + 1) * kG + k0 + kk] : 0.f;
                Bs[kk][mm] = Wh[(size_t)(u0 + mm) * kG + k0 + kk];        // Wh^T: (k = gate column, n = unit)
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < 16; ++kk) {
                const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
                const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
                const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
            }
            __syncthreads();
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int p = p0 + ty * 4 + i;
        if (p >= P) continue;
        const size_t row = (size_t)p * B + t;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int u = u0 + tx * 4 + j;
            const float dh = dh_out[row * kH + u] + acc[i][j];
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

// column sums of dA [T][1024] in a fixed order: block = 32 columns, 8 row groups, then the 8 partials in order
__global__ void __launch_bounds__(256)
colsum_kernel(const float* __restrict__ dA, int T, float* __restrict__ out) {
    __shared__ float red[8][33];
    const int c = blockIdx.x * 32 + (threadIdx.x & 31), g = threadIdx.x >> 5;
    float s = 0.f;
    for (int r = g; r < T; r += 8) s += dA[(size_t)r * kG + c];
    red[g][threadIdx.x & 31] = s;
    __syncthreads();
    if (g == 0) {
        float tot = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) tot += red[q][threadIdx.x & 31];
        out[c] = tot;
    }
}

int gemm(pn_handle* h, const GemmArgs& g, cudaStream_t st, const char* what) {
    dim3 grid((g.N + 63) / 64, (g.M + 63) / 64);
    gemm_kernel<<<grid, 256, 0, st>>>(g);
    PN_LAUNCH_CHECK(h, what);
    return PN_OK;
}

}  // namespace

size_t pn_lstm_workspace_floats(const pn_handle* h) {
    // gx / dA share nothing: gx [T,1024], act [T,1024], dA [T,1024], c, tanh c, h_out, h_prev [T,256] each, dc carry [P,256]
    const size_t T = (size_t)h->T;
    return T * kG * 3 + T * kH * 4 + (size_t)h->P * kH;
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
    GemmArgs g{h->hbuf, kH, 1, W, kG, 1, gx, kG, h->params + nd.lstmb, T, kG, kH};      // rows 0..255 of the kernel: the input
    if (int rc = gemm(h, g, st, "lstm input projection")) return rc;
    const float* Wh = W + (size_t)kH * kG;
    dim3 grid(kH / 16, (P + 63) / 64);
    for (int t = 0; t < B; ++t) {
        lstm_step_fwd_kernel<<<grid, 256, 0, st>>>(gx, Wh, h->lstm_state_in, P, B, t, act, cbuf, tcbuf, hout, hprev,
                                                  h->lstm_state_out);
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
    (void)hout;
    PnProfScope prof(h, PN_PROF_LSTM_BWD, st);
    dim3 grid(kH / 64, (P + 63) / 64);
    for (int t = B - 1; t >= 0; --t) {
        lstm_step_bwd_kernel<<<grid, 256, 0, st>>>(h->dh, Wh, act, cbuf, tcbuf, h->lstm_state_in, P, B, t, dA, dcc);
        PN_LAUNCH_CHECK(h, "lstm_step_bwd_kernel");
    }
    float* gW = h->grads + nd.lstmW;
    // dW rows 0..255 = X^T dA, rows 256..511 = Hprev^T dA (sum over every transition of every path)
    GemmArgs gw1{h->hbuf, 1, kH, dA, kG, 1, gW, kG, nullptr, kH, kG, T};
    if (int rc = gemm(h, gw1, st, "lstm weight gradient (input rows)")) return rc;
    GemmArgs gw2{hprev, 1, kH, dA, kG, 1, gW + (size_t)kH * kG, kG, nullptr, kH, kG, T};
    if (int rc = gemm(h, gw2, st, "lstm weight gradient (recurrent rows)")) return rc;
    colsum_kernel<<<kG / 32, 256, 0, st>>>(dA, T, h->grads + nd.lstmb);
    PN_LAUNCH_CHECK(h, "colsum_kernel");
    // gradient at the cell input = dA @ Wx^T, overwriting dh (every step has consumed its dh_out by now)
    GemmArgs gd{dA, kG, 1, W, 1, kG, h->dh, kH, nullptr, T, kH, kG};
    return gemm(h, gd, st, "lstm input gradient");
}
