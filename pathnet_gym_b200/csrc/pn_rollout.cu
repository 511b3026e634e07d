// Rollout bookkeeping on the device: what a3c_training_thread.py does on the host between two
// network calls of a rollout, for a whole population at once.
//   * choose_action (a3c_training_thread.py:89-90): np.random.choice(range(A), p=pi) -- numpy's legacy algorithm is
//     cdf = cumsum(float64(p)); cdf /= cdf[-1]; index = searchsorted(cdf, u, side='right') with ONE uniform u.  Here u
//     comes from the counter-based Philox stream (key = seed, counter = (epoch, id, 0, DOMAIN_ACTION<<24)), so a draw
//     is a pure function of (seed, epoch, id): bit-exact against the oracle restatement, which in turn is pinned
//     against numpy's own RandomState.choice (tests/test_oracle.py).
//   * batch assembly (a3c_training_thread.py:136,155-180): rewards clipped to [-1, 1]; R <- r_i + gamma * R backwards from
//     the bootstrap value (0 after a terminal step); td_i = R_i - V_i.  Python floats are fp64: the scan runs in fp64
//     with separate multiply and add (no FMA contraction) and rounds once to fp32 at the end.
// Paths whose rollout ended early (terminal) are padded to the common length B: padded entries get R = td = 0 here and
// the caller marks them with action = -1, which the loss kernel treats as "contributes nothing".
#include "pn_internal.cuh"

namespace {

enum { DOMAIN_ACTION = 3 };

__global__ void sample_actions_kernel(const float* __restrict__ pi, int n, int A, uint64_t seed, uint32_t epoch, uint32_t id0,
                                      int32_t* __restrict__ actions) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const float* p = pi + (size_t)t * A;
    double cdf[32];
    double s = 0.0;
    for (int k = 0; k < A; ++k) { s = __dadd_rn(s, (double)p[k]); cdf[k] = s; }      // np.cumsum: sequential fp64 adds
    const double total = cdf[A - 1];
    const double u = pn_philox_uniform(seed, epoch, id0 + (uint32_t)t, 0u, DOMAIN_ACTION, 0u);
    int idx = 0;
    for (int k = 0; k < A; ++k) idx += (__ddiv_rn(cdf[k], total) <= u) ? 1 : 0;        // searchsorted(..., side='right')
    actions[t] = idx < A ? idx : A - 1;
}

__global__ void nstep_returns_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                                     const float* __restrict__ bootstrap, const int32_t* __restrict__ lengths, int P, int B,
                                     double gamma, float* __restrict__ R_out, float* __restrict__ td_out) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    int len = lengths ? lengths[p] : B;
    len = len < 0 ? 0 : (len > B ? B : len);
    double R = (double)bootstrap[p];
    for (int i = B - 1; i >= len; --i) { R_out[(size_t)p * B + i] = 0.f; td_out[(size_t)p * B + i] = 0.f; }
    for (int i = len - 1; i >= 0; --i) {
        double r = (double)rewards[(size_t)p * B + i];
        r = r < -1.0 ? -1.0 : (r > 1.0 ? 1.0 : r);                                     // np.clip(reward, -1, 1)
        R = __dadd_rn(r, __dmul_rn(gamma, R));
        R_out[(size_t)p * B + i] = (float)R;
        td_out[(size_t)p * B + i] = (float)__dsub_rn(R, (double)values[(size_t)p * B + i]);
    }
}

// ---- frame stacks from the rollout's unique planes ------------------------------------------------------------------
// GameState keeps s_t as the stack of the C most recent preprocessed frames (game_state.py:62-66 reset: C copies of the
// first frame; :75-81 step: s_t1 = append(s_t[:, :, 1:], x_t1)), so consecutive transitions of one rollout share C-1 of
// their C channels.  A batch of B transitions therefore consists of only B + C - 1 distinct H x W planes per path:
//     planes[p][0 .. C-1] = the channels of s_0 (after a reset: C copies of the first frame),  planes[p][C-1+i] = x_i
//     frames[p][b][y][x][c] = planes[p][b + c][y][x]
// The host ships the planes (3.7x fewer bytes over PCIe at B = 32, C = 4); this kernel rebuilds the stacks the layer-0
// kernels read.  One thread = 4 pixels x C channels (C == 4: a 4x4 byte transpose, 16-byte coalesced stores).
__global__ void __launch_bounds__(256)
expand_planes_kernel(const uint8_t* __restrict__ planes, int P, int B, int HW, int C, uint8_t* __restrict__ frames) {
    const int t = blockIdx.y;                                   // transition p * B + b
    const int p = t / B, b = t - p * B;
    const int q = blockIdx.x * blockDim.x + threadIdx.x;        // group of 4 pixels
    if (q * 4 >= HW) return;
    const uint8_t* src = planes + ((size_t)p * (B + C - 1) + b) * HW + (size_t)q * 4;
    uint8_t* dst = frames + ((size_t)t * HW + (size_t)q * 4) * C;
    if (C == 4) {
        const uint32_t a0 = *reinterpret_cast<const uint32_t*>(src), a1 = *reinterpret_cast<const uint32_t*>(src + HW);
        const uint32_t a2 = *reinterpret_cast<const uint32_t*>(src + 2 * (size_t)HW), a3 = *reinterpret_cast<const uint32_t*>(src + 3 * (size_t)HW);
        // byte k of plane word a_c is pixel k, channel c -> output word k holds bytes (a0.k, a1.k, a2.k, a3.k)
        const uint32_t lo01 = __byte_perm(a0, a1, 0x5140), hi01 = __byte_perm(a0, a1, 0x7362);   // (a0.0 a1.0 a0.1 a1.1), (a0.2 a1.2 a0.3 a1.3)
        const uint32_t lo23 = __byte_perm(a2, a3, 0x5140), hi23 = __byte_perm(a2, a3, 0x7362);
        uint4 o;
        o.x = __byte_perm(lo01, lo23, 0x5410); o.y = __byte_perm(lo01, lo23, 0x7632);
        o.z = __byte_perm(hi01, hi23, 0x5410); o.w = __byte_perm(hi01, hi23, 0x7632);
        *reinterpret_cast<uint4*>(dst) = o;
    } else {
        for (int k = 0; k < 4; ++k)
            for (int c = 0; c < C; ++c) dst[k * C + c] = src[(size_t)c * HW + k];
    }
}

}  // namespace

// paths [p0, p0 + np) of the slot's plane and frame buffers
int pn_launch_expand_planes(pn_handle* h, const uint8_t* planes_dev, int B, uint8_t* frames_dev, int p0, int np, cudaStream_t st) {
    const int HW = h->cfg.H * h->cfg.W, C = h->cfg.C;
    if (HW % 4) return pn_fail(h, PN_ERR_INVALID, "H*W = %d must be a multiple of 4", HW);
    if (np <= 0) return PN_OK;
    dim3 grid((HW / 4 + 255) / 256, np * B);
    expand_planes_kernel<<<grid, 256, 0, st>>>(planes_dev + (size_t)p0 * (B + C - 1) * HW, np, B, HW, C,
                                               frames_dev + (size_t)p0 * B * HW * C);
    PN_LAUNCH_CHECK(h, "expand_planes_kernel");
    return PN_OK;
}

extern "C" PN_API int pn_sample_actions(const float* pi_dev, int n, int A, uint64_t seed, uint32_t epoch, uint32_t id0,
                                        int32_t* actions_dev, void* stream) {
    if (!pi_dev || !actions_dev) return pn_fail_global(PN_ERR_INVALID, "pn_sample_actions: null argument");
    if (n < 0 || A < 1 || A > 32) return pn_fail_global(PN_ERR_INVALID, "pn_sample_actions: n = %d, A = %d (A must be 1..32)", n, A);
    if (n == 0) return PN_OK;
    sample_actions_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(pi_dev, n, A, seed, epoch, id0, actions_dev);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return pn_fail_global(PN_ERR_CUDA, "sample_actions_kernel: %s", cudaGetErrorString(e));
    return PN_OK;
}

extern "C" PN_API int pn_nstep_returns(const float* rewards_dev, const float* values_dev, const float* bootstrap_dev,
                                       const int32_t* lengths_dev, int P, int B, double gamma, float* R_dev, float* td_dev,
                                       void* stream) {
    if (!rewards_dev || !values_dev || !bootstrap_dev || !R_dev || !td_dev)
        return pn_fail_global(PN_ERR_INVALID, "pn_nstep_returns: null argument");
    if (P < 0 || B < 1) return pn_fail_global(PN_ERR_INVALID, "pn_nstep_returns: P = %d, B = %d", P, B);
    if (P == 0) return PN_OK;
    nstep_returns_kernel<<<(P + 127) / 128, 128, 0, (cudaStream_t)stream>>>(rewards_dev, values_dev, bootstrap_dev, lengths_dev,
                                                                           P, B, gamma, R_dev, td_dev);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return pn_fail_global(PN_ERR_CUDA, "nstep_returns_kernel: %s", cudaGetErrorString(e));
    return PN_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Frame preprocessing (game_state.py:37-51,62-66,75-81): the step right before the hot path.
//   observation [n, Hin, Win, 3] u8  --cv2.cvtColor(COLOR_BGR2GRAY)-->  gray u8  --cv2.resize(INTER_LINEAR)-->  [H, W] u8
//   --> pushed into the 4-frame stack [n, H, W, C] u8 (s_t1 = append(s_t[:, :, 1:], x_t1); on reset all C channels = x_t).
// The reference then multiplies by 1/255 into fp32; the product's u8 frame k IS k/255 (pathnet_b200.h), so nothing is lost.
// Bit-exact restatement of OpenCV's 8-bit paths (checked against cv2 itself in tests/test_oracle.py):
//   gray = (3735 B + 19235 G + 9798 R + 2^14) >> 15
//   resize: fx = float((dx + 0.5) * (Win / W) - 0.5); sx = floor(fx); fx -= sx; clamp to the image (weights (1, 0));
//           ia = round_half_even((1 - fx) * 2048), ib = round_half_even(fx * 2048); same in y;
//           row value r = S[sx] * ia + S[sx + 1] * ib;  out = (((by0 * (r0 >> 4)) >> 16) + ((by1 * (r1 >> 4)) >> 16) + 2) >> 2
// ---------------------------------------------------------------------------------------------------------------
namespace {

__device__ __forceinline__ void resize_tap(int d, int dn, int sn, int& s0, int& s1, int& a0, int& a1) {
    const double scale = (double)sn / (double)dn;
    float f = (float)__dsub_rn(__dmul_rn((double)d + 0.5, scale), 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= sn - 1) { s = sn - 1; f = 0.f; }
    s0 = s; s1 = min(s + 1, sn - 1);
    a0 = __float2int_rn((1.f - f) * 2048.f);
    a1 = __float2int_rn(f * 2048.f);
}

__device__ __forceinline__ int gray_at(const uint8_t* __restrict__ img, int Win, int y, int x) {
    const uint8_t* px = img + ((size_t)y * Win + x) * 3;
    return (3735 * (int)px[0] + 19235 * (int)px[1] + 9798 * (int)px[2] + (1 << 14)) >> 15;
}

__global__ void preprocess_frames_kernel(const uint8_t* __restrict__ obs, int n, int Hin, int Win, uint8_t* __restrict__ frames,
                                         int H, int W, int C, const uint8_t* __restrict__ reset) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n * H * W) return;
    const int x = (int)(i % W), y = (int)((i / W) % H), e = (int)(i / ((long long)W * H));
    int x0, x1, ax0, ax1, y0, y1, ay0, ay1;
    resize_tap(x, W, Win, x0, x1, ax0, ax1);
    resize_tap(y, H, Hin, y0, y1, ay0, ay1);
    const uint8_t* img = obs + (size_t)e * Hin * Win * 3;
    const int r0 = gray_at(img, Win, y0, x0) * ax0 + gray_at(img, Win, y0, x1) * ax1;
    const int r1 = gray_at(img, Win, y1, x0) * ax0 + gray_at(img, Win, y1, x1) * ax1;
    int v = (((ay0 * (r0 >> 4)) >> 16) + ((ay1 * (r1 >> 4)) >> 16) + 2) >> 2;
    v = v < 0 ? 0 : (v > 255 ? 255 : v);
    uint8_t* px = frames + (size_t)i * C;
    if (reset && reset[e]) {
        for (int c = 0; c < C; ++c) px[c] = (uint8_t)v;
    } else {
        for (int c = 0; c + 1 < C; ++c) px[c] = px[c + 1];
        px[C - 1] = (uint8_t)v;
    }
}

}  // namespace

extern "C" PN_API int pn_preprocess_frames(const uint8_t* obs_dev, int n, int Hin, int Win, uint8_t* frames_dev, int H, int W,
                                           int C, const uint8_t* reset_dev, void* stream) {
    if (!obs_dev || !frames_dev) return pn_fail_global(PN_ERR_INVALID, "pn_preprocess_frames: null argument");
    if (n < 0 || Hin < 1 || Win < 1 || H < 1 || W < 1 || C < 1 || C > 16)
        return pn_fail_global(PN_ERR_INVALID, "pn_preprocess_frames: n = %d, %dx%d -> %dx%dx%d", n, Hin, Win, H, W, C);
    const long long total = (long long)n * H * W;
    if (total == 0) return PN_OK;
    preprocess_frames_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(obs_dev, n, Hin, Win, frames_dev,
                                                                                             H, W, C, reset_dev);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return pn_fail_global(PN_ERR_CUDA, "preprocess_frames_kernel: %s", cudaGetErrorString(e));
    return PN_OK;
}
