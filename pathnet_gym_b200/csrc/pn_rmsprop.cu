// Per-tensor clip_by_norm + RMSProp over the fp32 parameter / accumulator slabs.
//
// Reference: RMSPropApplier.apply_gradients (rmsprop_applier.py:92-106) =
//   g' = tf.clip_by_norm(g, clip)                 per tensor (W and b separately)
//   training_ops.apply_rms_prop(var, rms, mom, lr, decay, momentum, eps, g')
//       ms  += (g'^2 - ms) * (1 - decay)
//       mom  = mom * momentum + (g' * lr) / sqrt(eps + ms)
//       var -= mom
// Tensors whose update op the caller does not run (frozen path,
// a3c_training_thread.py:205-209) are skipped: var, ms, mom untouched.
//
// HBM-bound: 20 B per non-frozen parameter when momentum == 0 (read g, var, ms;
// write var, ms; the momentum slot is elided because it is overwritten by
// lr*g'/sqrt(ms+eps) and never read).  The norm pre-pass re-reads g from L2.
// The slab is cut into chunks of PN_RMS_CHUNK floats that never straddle a
// tensor; phase 1 writes one partial sum of squares per chunk, phase 2 adds a
// tensor's partials in index order (deterministic) and updates.
#include "pn_internal.cuh"

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ float block_sum(float v, float* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) red[w] = v;
    __syncthreads();
    float s = 0.f;
    if (threadIdx.x < 32) {
        s = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    return s;   // valid in warp 0
}

__global__ void __launch_bounds__(kThreads)
rms_sumsq_kernel(const float* __restrict__ grads, const PnRmsChunk* __restrict__ chunks,
                 const uint8_t* __restrict__ frozen, float* __restrict__ chunk_ss, int c0, int n0, int c1) {
    __shared__ float red[32];
    // two chunk segments in one launch: [c0, c0 + n0) then [c1, ...)
    const int cid = (int)blockIdx.x < n0 ? c0 + (int)blockIdx.x : c1 + ((int)blockIdx.x - n0);
    const PnRmsChunk ck = chunks[cid];
    if (frozen[ck.tensor]) return;
    const float* g = grads + ck.off;
    const int n4 = ck.len >> 2;
    float s = 0.f;
    for (int i = threadIdx.x; i < n4; i += kThreads) {
        const float4 v = reinterpret_cast<const float4*>(g)[i];
        s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
    }
    for (int i = (n4 << 2) + threadIdx.x; i < ck.len; i += kThreads) s = fmaf(g[i], g[i], s);
    s = block_sum(s, red);
    if (threadIdx.x == 0) chunk_ss[cid] = s;
}

template <bool HAS_MOM>
__device__ __forceinline__ void rms_elem(float g, float& var, float& ms, float& mom, float clip, float scale,
                                         float lr, float one_minus_decay, float momentum, float eps) {
    const float gc = (g * clip) * scale;                       // tf.clip_by_norm
    ms = ms + (gc * gc - ms) * one_minus_decay;
    const float upd = (gc * lr) / sqrtf(eps + ms);
    if (HAS_MOM) { mom = mom * momentum + upd; var -= mom; }
    else { var -= upd; }
}

template <bool HAS_MOM>
__global__ void __launch_bounds__(kThreads)
rms_update_kernel(float* __restrict__ params, float* __restrict__ msl, float* __restrict__ moml,
                  const float* __restrict__ grads, const PnRmsChunk* __restrict__ chunks,
                  const uint8_t* __restrict__ frozen, const int* __restrict__ tchunk0,
                  const int* __restrict__ tnchunks, const float* __restrict__ chunk_ss, float lr, float decay,
                  float momentum, float eps, float clip, int c0, int n0, int c1) {
    __shared__ float s_scale;
    const PnRmsChunk ck = chunks[(int)blockIdx.x < n0 ? c0 + (int)blockIdx.x : c1 + ((int)blockIdx.x - n0)];
    if (frozen[ck.tensor]) return;
    if (threadIdx.x == 0) {
        const int c0 = tchunk0[ck.tensor], nc = tnchunks[ck.tensor];
        float ss = 0.f;
        for (int c = 0; c < nc; ++c) ss += chunk_ss[c0 + c];
        // t*clip * min(rsqrt(sum t^2), 1/clip); rsqrt(0) = inf -> 1/clip
        const float l2inv = 1.0f / sqrtf(ss);
        s_scale = fminf(l2inv, 1.0f / clip);
    }
    __syncthreads();
    const float scale = s_scale;
    const float omd = 1.0f - decay;
    float* pv = params + ck.off;
    float* pm = msl + ck.off;
    float* pq = HAS_MOM ? moml + ck.off : nullptr;
    const float* g = grads + ck.off;
    const int n4 = ck.len >> 2;
    for (int i = threadIdx.x; i < n4; i += kThreads) {
        const float4 gv = reinterpret_cast<const float4*>(g)[i];
        float4 v = reinterpret_cast<float4*>(pv)[i];
        float4 m = reinterpret_cast<float4*>(pm)[i];
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        if (HAS_MOM) q = reinterpret_cast<float4*>(pq)[i];
        rms_elem<HAS_MOM>(gv.x, v.x, m.x, q.x, clip, scale, lr, omd, momentum, eps);
        rms_elem<HAS_MOM>(gv.y, v.y, m.y, q.y, clip, scale, lr, omd, momentum, eps);
        rms_elem<HAS_MOM>(gv.z, v.z, m.z, q.z, clip, scale, lr, omd, momentum, eps);
        rms_elem<HAS_MOM>(gv.w, v.w, m.w, q.w, clip, scale, lr, omd, momentum, eps);
        reinterpret_cast<float4*>(pv)[i] = v;
        reinterpret_cast<float4*>(pm)[i] = m;
        if (HAS_MOM) reinterpret_cast<float4*>(pq)[i] = q;
    }
    for (int i = (n4 << 2) + threadIdx.x; i < ck.len; i += kThreads) {
        float v = pv[i], m = pm[i], q = HAS_MOM ? pq[i] : 0.f;
        rms_elem<HAS_MOM>(g[i], v, m, q, clip, scale, lr, omd, momentum, eps);
        pv[i] = v; pm[i] = m;
        if (HAS_MOM) pq[i] = q;
    }
}

// Both phases in ONE launch for selections of small tensors (conv and head tensors: at most a few chunks each): one block
// per tensor computes the per-chunk sums of squares exactly like rms_sumsq_kernel (same thread mapping, same block
// reduction), adds them in chunk order like rms_update_kernel, and updates.  Bit-identical to the two-kernel path; two
// dependent ~6 us launches become one.
constexpr int kFusedMaxChunks = 32;

template <bool HAS_MOM>
__global__ void __launch_bounds__(kThreads)
rms_fused_tensor_kernel(float* __restrict__ params, float* __restrict__ msl, float* __restrict__ moml,
                        const float* __restrict__ grads, const PnRmsChunk* __restrict__ chunks,
                        const uint8_t* __restrict__ frozen, const int* __restrict__ tchunk0, const int* __restrict__ tnchunks,
                        int first, int n, int complement, float lr, float decay, float momentum, float eps, float clip) {
    __shared__ float red[32];
    __shared__ float ssum[kFusedMaxChunks];
    __shared__ float s_scale;
    const int b = blockIdx.x;
    const int t = complement ? (b < first ? b : b + n) : first + b;
    if (frozen[t]) return;
    const int c0 = tchunk0[t], nc = tnchunks[t];
    for (int c = 0; c < nc; ++c) {
        const PnRmsChunk ck = chunks[c0 + c];
        const float* g = grads + ck.off;
        const int n4 = ck.len >> 2;
        float s = 0.f;
        for (int i = threadIdx.x; i < n4; i += kThreads) {
            const float4 v = reinterpret_cast<const float4*>(g)[i];
            s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
        }
        for (int i = (n4 << 2) + threadIdx.x; i < ck.len; i += kThreads) s = fmaf(g[i], g[i], s);
        s = block_sum(s, red);
        if (threadIdx.x == 0) ssum[c] = s;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        float ss = 0.f;
        for (int c = 0; c < nc; ++c) ss += ssum[c];
        const float l2inv = 1.0f / sqrtf(ss);
        s_scale = fminf(l2inv, 1.0f / clip);
    }
    __syncthreads();
    const float scale = s_scale;
    const float omd = 1.0f - decay;
    for (int c = 0; c < nc; ++c) {
        const PnRmsChunk ck = chunks[c0 + c];
        float* pv = params + ck.off;
        float* pm = msl + ck.off;
        float* pq = HAS_MOM ? moml + ck.off : nullptr;
        const float* g = grads + ck.off;
        const int n4 = ck.len >> 2;
        for (int i = threadIdx.x; i < n4; i += kThreads) {
            const float4 gv = reinterpret_cast<const float4*>(g)[i];
            float4 v = reinterpret_cast<float4*>(pv)[i];
            float4 m = reinterpret_cast<float4*>(pm)[i];
            float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
            if (HAS_MOM) q = reinterpret_cast<float4*>(pq)[i];
            rms_elem<HAS_MOM>(gv.x, v.x, m.x, q.x, clip, scale, lr, omd, momentum, eps);
            rms_elem<HAS_MOM>(gv.y, v.y, m.y, q.y, clip, scale, lr, omd, momentum, eps);
            rms_elem<HAS_MOM>(gv.z, v.z, m.z, q.z, clip, scale, lr, omd, momentum, eps);
            rms_elem<HAS_MOM>(gv.w, v.w, m.w, q.w, clip, scale, lr, omd, momentum, eps);
            reinterpret_cast<float4*>(pv)[i] = v;
            reinterpret_cast<float4*>(pm)[i] = m;
            if (HAS_MOM) reinterpret_cast<float4*>(pq)[i] = q;
        }
        for (int i = (n4 << 2) + threadIdx.x; i < ck.len; i += kThreads) {
            float v = pv[i], m = pm[i], q = HAS_MOM ? pq[i] : 0.f;
            rms_elem<HAS_MOM>(g[i], v, m, q, clip, scale, lr, omd, momentum, eps);
            pv[i] = v; pm[i] = m;
            if (HAS_MOM) pq[i] = q;
        }
    }
}

__global__ void __launch_bounds__(kThreads)
reinit_kernel(float* __restrict__ params, const float* __restrict__ backup, const PnRmsChunk* __restrict__ chunks,
              const uint8_t* __restrict__ frozen) {
    const PnRmsChunk ck = chunks[blockIdx.x];
    if (frozen[ck.tensor]) return;
    for (int i = threadIdx.x; i < ck.len; i += kThreads) params[ck.off + i] = backup[ck.off + i];
}

}  // namespace

int pn_launch_rmsprop(pn_handle* h, int first_tensor, int n_tensors, bool complement, float lr, float decay,
                      float momentum, float eps, float clip, cudaStream_t st) {
    // chunks are stored in tensor order: a contiguous range of tensors is a contiguous range of chunks
    const int nt = (int)h->tensors.size(), nall = (int)h->h_chunks.size();
    const int ca = first_tensor < nt ? h->h_tensor_chunk0[first_tensor] : nall;
    const int cb = first_tensor + n_tensors < nt ? h->h_tensor_chunk0[first_tensor + n_tensors] : nall;
    // segment 0 = [c0, c0 + n0), segment 1 = [c1, c1 + n1)
    const int c0 = complement ? 0 : ca, n0 = complement ? ca : cb - ca;
    const int c1 = cb, n1 = complement ? nall - cb : 0;
    const int nchunks = n0 + n1;
    if (nchunks <= 0) return PN_OK;
    // a selection of small tensors only (conv + head tensors): one fused launch, one block per tensor
    {
        const int nsel = complement ? nt - n_tensors : n_tensors;
        bool small = nsel > 0 && nsel <= 4096;
        for (int t = 0; t < nt && small; ++t) {
            const bool in_range = t >= first_tensor && t < first_tensor + n_tensors;
            if (in_range == complement) continue;
            const int tc = (t + 1 < nt ? h->h_tensor_chunk0[t + 1] : nall) - h->h_tensor_chunk0[t];
            if (tc > kFusedMaxChunks) small = false;
        }
        if (small) {
            PnProfScope prof(h, PN_PROF_RMS_UPDATE, st);
            if (momentum != 0.0f)
                rms_fused_tensor_kernel<true><<<nsel, kThreads, 0, st>>>(h->params, h->ms, h->mom, h->grads, h->d_chunks, h->d_frozen,
                                                                         h->d_tensor_chunk0, h->d_tensor_nchunks, first_tensor,
                                                                         n_tensors, complement ? 1 : 0, lr, decay, momentum, eps, clip);
            else
                rms_fused_tensor_kernel<false><<<nsel, kThreads, 0, st>>>(h->params, h->ms, nullptr, h->grads, h->d_chunks, h->d_frozen,
                                                                          h->d_tensor_chunk0, h->d_tensor_nchunks, first_tensor,
                                                                          n_tensors, complement ? 1 : 0, lr, decay, momentum, eps, clip);
            PN_LAUNCH_CHECK(h, "rms_fused_tensor_kernel");
            return PN_OK;
        }
    }
    {
        PnProfScope prof(h, PN_PROF_RMS_SUMSQ, st);
        rms_sumsq_kernel<<<nchunks, kThreads, 0, st>>>(h->grads, h->d_chunks, h->d_frozen, h->d_chunk_ss, c0, n0, c1);
    }
    PN_LAUNCH_CHECK(h, "rms_sumsq_kernel");
    PnProfScope prof(h, PN_PROF_RMS_UPDATE, st);
    if (momentum != 0.0f) {
        rms_update_kernel<true><<<nchunks, kThreads, 0, st>>>(h->params, h->ms, h->mom, h->grads, h->d_chunks,
                                                              h->d_frozen, h->d_tensor_chunk0, h->d_tensor_nchunks,
                                                              h->d_chunk_ss, lr, decay, momentum, eps, clip, c0, n0, c1);
    } else {
        rms_update_kernel<false><<<nchunks, kThreads, 0, st>>>(h->params, h->ms, nullptr, h->grads, h->d_chunks,
                                                               h->d_frozen, h->d_tensor_chunk0, h->d_tensor_nchunks,
                                                               h->d_chunk_ss, lr, decay, momentum, eps, clip, c0, n0, c1);
    }
    PN_LAUNCH_CHECK(h, "rms_update_kernel");
    return PN_OK;
}

int pn_launch_reinit(pn_handle* h, const float* backup, cudaStream_t st) {
    const int nchunks = (int)h->h_chunks.size();
    reinit_kernel<<<nchunks, kThreads, 0, st>>>(h->params, backup, h->d_chunks, h->d_frozen);
    PN_LAUNCH_CHECK(h, "reinit_kernel");
    return PN_OK;
}
