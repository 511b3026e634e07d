// Per-tensor clip_by_norm + RMSProp: the two kernels shared by pn_rmsprop.cu (handle: tables in device memory) and
// pn_noalloc.cu (no handle: the tensor table travels as a kernel argument).
//
// Reference: RMSPropApplier.apply_gradients (rmsprop_applier.py:92-106) =
//   g' = tf.clip_by_norm(g, clip)                 per tensor (W and b separately)
//   training_ops.apply_rms_prop(var, rms, mom, lr, decay, momentum, eps, g')
//       ms  += (g'^2 - ms) * (1 - decay)
//       mom  = mom * momentum + (g' * lr) / sqrt(eps + ms)
//       var -= mom
// Tensors whose update op the caller does not run (frozen path, a3c_training_thread.py:205-209) are skipped.
//
// Work unit = a chunk of PN_RMS_CHUNK floats that never straddles a tensor, one 256-thread block per chunk, two
// float4 per thread.  Kernel 1 writes one sum of squares per chunk.  Kernel 2 is launched with programmatic stream
// serialisation: its blocks start while kernel 1 drains, issue the loads of their chunk's gradient, parameters and rms
// slots (none of which kernel 1 writes), and only then wait for kernel 1 (griddepcontrol.wait); warp 0 adds the
// tensor's chunk sums -- lane i takes chunks i, i+32, ... in order, then a fixed shuffle tree, so the result is
// deterministic -- and the block updates from registers.  HBM traffic: 20 B per parameter (g, var, ms in; var, ms
// out; the momentum slot is elided when the coefficient is 0); the second read of g is an L2 hit.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#ifndef PN_RMS_CHUNK
#define PN_RMS_CHUNK 2048
#endif

namespace pnrms {

constexpr int kThreads = 256;
constexpr int kPerThread = PN_RMS_CHUNK / (4 * kThreads);      // float4 per thread = 2

struct ChunkRef { int tensor; int len; long long off; bool frozen; int chunk0; int nchunks; };

__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// sum over the block of one value per thread; the result is valid in warp 0 (all its lanes)
__device__ __forceinline__ float block_sum(float v, float* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) red[w] = v;
    __syncthreads();
    float s = 0.f;
    if (threadIdx.x < 32) {
        s = threadIdx.x < (kThreads >> 5) ? red[threadIdx.x] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    return s;
}

// ---- kernel 1: one sum of squares per chunk ---------------------------------------------------------------------------
template <class Tab>
__device__ __forceinline__ void sumsq_body(const Tab& tb, int cid, const float* __restrict__ grads, float* __restrict__ chunk_ss) {
    __shared__ float red[32];
    pdl_launch_dependents();                                    // the update kernel's blocks may start their loads now
    const ChunkRef ck = tb.chunk(cid);
    if (ck.frozen) return;
    const float* g = grads + ck.off;
    const bool vec = (ck.off & 3) == 0;
    const int n4 = vec ? (ck.len >> 2) : 0;
    float s = 0.f;
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) {
        const int i = threadIdx.x + u * kThreads;
        if (i < n4) {
            const float4 v = reinterpret_cast<const float4*>(g)[i];
            s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
        }
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

// ---- kernel 2: the update -----------------------------------------------------------------------------------------------
template <bool HAS_MOM, class Tab>
__device__ __forceinline__ void update_body(const Tab& tb, int cid, float* __restrict__ params, float* __restrict__ msl,
                                            float* __restrict__ moml, const float* __restrict__ grads,
                                            const float* __restrict__ chunk_ss, float lr, float decay, float momentum,
                                            float eps, float clip) {
    __shared__ float s_scale;
    const ChunkRef ck = tb.chunk(cid);
    if (ck.frozen) return;
    float* pv = params + ck.off;
    float* pm = msl + ck.off;
    float* pq = HAS_MOM ? moml + ck.off : nullptr;
    const float* g = grads + ck.off;
    const bool vec = (ck.off & 3) == 0;
    const int n4 = vec ? (ck.len >> 2) : 0;
    // everything this block reads except the chunk sums is independent of kernel 1: issue the loads first
    float4 gv[kPerThread], vv[kPerThread], mv[kPerThread], qv[kPerThread];
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) {
        const int i = threadIdx.x + u * kThreads;
        if (i < n4) {
            gv[u] = reinterpret_cast<const float4*>(g)[i];
            vv[u] = reinterpret_cast<const float4*>(pv)[i];
            mv[u] = reinterpret_cast<const float4*>(pm)[i];
            if (HAS_MOM) qv[u] = reinterpret_cast<const float4*>(pq)[i];
        }
    }
    pdl_wait();                                                 // kernel 1 has completed: the chunk sums are visible
    if (threadIdx.x < 32) {
        float ss = 0.f;
        for (int c = threadIdx.x; c < ck.nchunks; c += 32) ss += __ldcg(chunk_ss + ck.chunk0 + c);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        // t*clip * min(rsqrt(sum t^2), 1/clip); rsqrt(0) = inf -> 1/clip
        if (threadIdx.x == 0) s_scale = fminf(1.0f / sqrtf(ss), 1.0f / clip);
    }
    __syncthreads();
    const float scale = s_scale;
    const float omd = 1.0f - decay;
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) {
        const int i = threadIdx.x + u * kThreads;
        if (i < n4) {
            float4 v = vv[u], m = mv[u], q = make_float4(0.f, 0.f, 0.f, 0.f);
            if (HAS_MOM) q = qv[u];
            rms_elem<HAS_MOM>(gv[u].x, v.x, m.x, q.x, clip, scale, lr, omd, momentum, eps);
            rms_elem<HAS_MOM>(gv[u].y, v.y, m.y, q.y, clip, scale, lr, omd, momentum, eps);
            rms_elem<HAS_MOM>(gv[u].z, v.z, m.z, q.z, clip, scale, lr, omd, momentum, eps);
            rms_elem<HAS_MOM>(gv[u].w, v.w, m.w, q.w, clip, scale, lr, omd, momentum, eps);
            reinterpret_cast<float4*>(pv)[i] = v;
            reinterpret_cast<float4*>(pm)[i] = m;
            if (HAS_MOM) reinterpret_cast<float4*>(pq)[i] = q;
        }
    }
    for (int i = (n4 << 2) + threadIdx.x; i < ck.len; i += kThreads) {
        float v = pv[i], m = pm[i], q = HAS_MOM ? pq[i] : 0.f;
        rms_elem<HAS_MOM>(g[i], v, m, q, clip, scale, lr, omd, momentum, eps);
        pv[i] = v; pm[i] = m;
        if (HAS_MOM) pq[i] = q;
    }
}

// Launch `kernel` so that it may begin while the preceding kernel in `st` is still draining (programmatic dependent launch).
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), int grid, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

}  // namespace pnrms
