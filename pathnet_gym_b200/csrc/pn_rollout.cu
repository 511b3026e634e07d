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

}  // namespace

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
