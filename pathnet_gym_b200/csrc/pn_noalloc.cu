// pn_rmsprop_apply_noalloc -- the clipped RMSProp update for callers that must not allocate.
//
// Who this is for.  Some hosts of this library cannot call cudaMalloc at the point where they
// want to update the weights:
//   * a trainer embedded in a server that reserved the whole GPU for its own pool at start-up, so
//     a later cudaMalloc fails (or is forbidden by the host's policy) even for a few kilobytes;
//   * a control loop that cannot tolerate cudaMalloc / cudaFree stalling the device -- both
//     synchronise it, which would break the overlap the loop was built around;
//   * a plugin that is handed device pointers by its host and is not allowed to own device memory.
// The regular entry point (pn_rmsprop_apply) needs a handle, and pn_create allocates the handle's
// workspace.  This variant takes no handle: the caller passes the slabs and the tensor table, and
// the one piece of scratch the kernels need -- a partial sum of squares per 2048-float chunk -- is
// a file-scope device array that every GPU which loads this library carries until the process
// exits (64 KiB per device).  No cudaMalloc, cudaFree or other allocating routine is called.
//
// Limits.  At most 128 tensors (the table travels as a kernel argument) and 16384 chunks of 2048
// floats (the size of the scratch array): about 33.5 M parameters when the tensors are large.
// Anything bigger is refused with PN_ERR_INVALID and a message; nothing is launched.
//
// Sharing.  There is ONE scratch array per device and it keeps whatever the previous call left in
// it.  Nothing in it has to start at zero: every slot a call reads is written earlier in that same
// call.  Calls issued one after another on one stream are therefore safe; calls whose kernels can
// run at the same time on one GPU (different streams, different host threads) race on the array and
// must be ordered by the caller.
//
// First call on a device.  It looks up the array's device address (cudaGetSymbolAddress, with the
// device current) and caches it.  Issue that first call before any stream capture begins.
#include "pn_internal.cuh"

#include <mutex>

#define PN_NOALLOC_MAX_TENSORS 128
#define PN_NOALLOC_MAX_CHUNKS 16384
#define PN_NOALLOC_MAX_DEVICES 64

namespace {

// Scratch shared by every call on this device.  In host code this name is only a handle for the
// runtime (its host-side shadow), never a device address: the address is looked up below.
static __device__ __align__(256) float g_noalloc_chunk_ss[PN_NOALLOC_MAX_CHUNKS];

struct NoallocTable {                       // passed by value (kernel parameter space, < 4 KB)
    int n;
    int chunk0[PN_NOALLOC_MAX_TENSORS + 1]; // first chunk of tensor t; chunk0[n] = total chunks
    int len[PN_NOALLOC_MAX_TENSORS];
    long long off[PN_NOALLOC_MAX_TENSORS];
    unsigned char frozen[PN_NOALLOC_MAX_TENSORS];
};

constexpr int kThreads = 256;

__device__ __forceinline__ int find_tensor(const NoallocTable& tb, int chunk) {
    int lo = 0, hi = tb.n - 1;               // last t with chunk0[t] <= chunk
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (tb.chunk0[mid] <= chunk) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__device__ __forceinline__ float block_sum_na(float v, float* red) {
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
    return s;
}

__global__ void __launch_bounds__(kThreads)
noalloc_sumsq_kernel(const float* __restrict__ grads, const __grid_constant__ NoallocTable tb,
                     float* __restrict__ chunk_ss) {
    __shared__ float red[32];
    const int t = find_tensor(tb, blockIdx.x);
    if (tb.frozen[t]) return;
    const int c = blockIdx.x - tb.chunk0[t];
    const long long o = (long long)c * PN_RMS_CHUNK;
    const int len = (int)min((long long)PN_RMS_CHUNK, (long long)tb.len[t] - o);
    const float* g = grads + tb.off[t] + o;
    const bool vec = ((tb.off[t] & 3) == 0);
    const int n4 = vec ? (len >> 2) : 0;
    float s = 0.f;
    for (int i = threadIdx.x; i < n4; i += kThreads) {
        const float4 v = reinterpret_cast<const float4*>(g)[i];
        s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
    }
    for (int i = (n4 << 2) + threadIdx.x; i < len; i += kThreads) s = fmaf(g[i], g[i], s);
    s = block_sum_na(s, red);
    if (threadIdx.x == 0) chunk_ss[blockIdx.x] = s;
}

template <bool HAS_MOM>
__device__ __forceinline__ void rms_elem_na(float g, float& var, float& ms, float& mom, float clip, float scale,
                                            float lr, float omd, float momentum, float eps) {
    const float gc = (g * clip) * scale;
    ms = ms + (gc * gc - ms) * omd;
    const float upd = (gc * lr) / sqrtf(eps + ms);
    if (HAS_MOM) { mom = mom * momentum + upd; var -= mom; }
    else { var -= upd; }
}

template <bool HAS_MOM>
__global__ void __launch_bounds__(kThreads)
noalloc_update_kernel(float* __restrict__ params, float* __restrict__ msl, float* __restrict__ moml,
                      const float* __restrict__ grads, const __grid_constant__ NoallocTable tb,
                      const float* __restrict__ chunk_ss, float lr, float decay, float momentum, float eps,
                      float clip) {
    __shared__ float s_scale;
    const int t = find_tensor(tb, blockIdx.x);
    if (tb.frozen[t]) return;
    if (threadIdx.x == 0) {
        float ss = 0.f;
        for (int c = tb.chunk0[t]; c < tb.chunk0[t + 1]; ++c) ss += chunk_ss[c];
        s_scale = fminf(1.0f / sqrtf(ss), 1.0f / clip);
    }
    __syncthreads();
    const float scale = s_scale, omd = 1.0f - decay;
    const int c = blockIdx.x - tb.chunk0[t];
    const long long o = tb.off[t] + (long long)c * PN_RMS_CHUNK;
    const int len = (int)min((long long)PN_RMS_CHUNK, (long long)tb.len[t] - (long long)c * PN_RMS_CHUNK);
    float* pv = params + o;
    float* pm = msl + o;
    float* pq = HAS_MOM ? moml + o : nullptr;
    const float* g = grads + o;
    const bool vec = ((tb.off[t] & 3) == 0);
    const int n4 = vec ? (len >> 2) : 0;
    for (int i = threadIdx.x; i < n4; i += kThreads) {
        const float4 gv = reinterpret_cast<const float4*>(g)[i];
        float4 v = reinterpret_cast<float4*>(pv)[i];
        float4 m = reinterpret_cast<float4*>(pm)[i];
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        if (HAS_MOM) q = reinterpret_cast<float4*>(pq)[i];
        rms_elem_na<HAS_MOM>(gv.x, v.x, m.x, q.x, clip, scale, lr, omd, momentum, eps);
        rms_elem_na<HAS_MOM>(gv.y, v.y, m.y, q.y, clip, scale, lr, omd, momentum, eps);
        rms_elem_na<HAS_MOM>(gv.z, v.z, m.z, q.z, clip, scale, lr, omd, momentum, eps);
        rms_elem_na<HAS_MOM>(gv.w, v.w, m.w, q.w, clip, scale, lr, omd, momentum, eps);
        reinterpret_cast<float4*>(pv)[i] = v;
        reinterpret_cast<float4*>(pm)[i] = m;
        if (HAS_MOM) reinterpret_cast<float4*>(pq)[i] = q;
    }
    for (int i = (n4 << 2) + threadIdx.x; i < len; i += kThreads) {
        float v = pv[i], m = pm[i], q = HAS_MOM ? pq[i] : 0.f;
        rms_elem_na<HAS_MOM>(g[i], v, m, q, clip, scale, lr, omd, momentum, eps);
        pv[i] = v; pm[i] = m;
        if (HAS_MOM) pq[i] = q;
    }
}

std::mutex g_setup_mutex;
float* g_scratch_ptr[PN_NOALLOC_MAX_DEVICES] = {};

// Once per device, with that device current: turn the symbol into the address kernels can use.
int scratch_for_current_device(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return pn_fail_global(PN_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= PN_NOALLOC_MAX_DEVICES)
        return pn_fail_global(PN_ERR_INVALID, "device ordinal %d outside 0..%d", dev, PN_NOALLOC_MAX_DEVICES - 1);
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_scratch_ptr[dev]) {
        float* p = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_chunk_ss);
        if (e != cudaSuccess || !p)
            return pn_fail_global(PN_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", dev,
                                  cudaGetErrorString(e));
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return PN_OK;
}

}  // namespace

extern "C" PN_API int pn_rmsprop_apply_noalloc(float* params_dev, float* ms_dev, float* mom_dev, const float* grads_dev,
                                               const pn_tensor_desc* tensors_host, const uint8_t* frozen_host,
                                               int num_tensors, float lr, float decay, float momentum, float epsilon,
                                               float clip_norm, void* stream) {
    if (!params_dev || !ms_dev || !grads_dev || !tensors_host)
        return pn_fail_global(PN_ERR_INVALID, "pn_rmsprop_apply_noalloc: null slab or tensor table");
    if (momentum != 0.0f && !mom_dev)
        return pn_fail_global(PN_ERR_INVALID, "pn_rmsprop_apply_noalloc: momentum != 0 needs a momentum slab");
    if (!(clip_norm > 0.0f)) return pn_fail_global(PN_ERR_INVALID, "clip_norm must be > 0");
    for (const void* p : {(const void*)params_dev, (const void*)ms_dev, (const void*)mom_dev, (const void*)grads_dev})
        if (p && (reinterpret_cast<uintptr_t>(p) & 15))
            return pn_fail_global(PN_ERR_INVALID, "pn_rmsprop_apply_noalloc: slabs must be 16-byte aligned");
    if (num_tensors < 1 || num_tensors > PN_NOALLOC_MAX_TENSORS)
        return pn_fail_global(PN_ERR_INVALID,
                              "pn_rmsprop_apply_noalloc: %d tensors, this entry point handles 1..%d (its tensor table "
                              "travels as a kernel argument); use pn_rmsprop_apply", num_tensors,
                              PN_NOALLOC_MAX_TENSORS);
    NoallocTable tb;
    tb.n = num_tensors;
    long long chunks = 0;
    for (int t = 0; t < num_tensors; ++t) {
        const pn_tensor_desc& d = tensors_host[t];
        if (d.offset < 0 || d.length < 1 || d.length > 0x7fffffffLL)
            return pn_fail_global(PN_ERR_INVALID, "tensor %d: bad offset/length %lld/%lld", t, (long long)d.offset,
                                  (long long)d.length);
        tb.chunk0[t] = (int)chunks;
        tb.len[t] = (int)d.length;
        tb.off[t] = d.offset;
        tb.frozen[t] = (frozen_host && frozen_host[t]) ? 1 : 0;
        chunks += (d.length + PN_RMS_CHUNK - 1) / PN_RMS_CHUNK;
        if (chunks > PN_NOALLOC_MAX_CHUNKS)
            return pn_fail_global(PN_ERR_INVALID,
                                  "pn_rmsprop_apply_noalloc: the tensors need more than %d chunks of %d floats, the size "
                                  "of the built-in scratch array; use pn_rmsprop_apply", PN_NOALLOC_MAX_CHUNKS,
                                  PN_RMS_CHUNK);
    }
    tb.chunk0[num_tensors] = (int)chunks;
    float* scratch = nullptr;
    if (int rc = scratch_for_current_device(&scratch)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int nblk = (int)chunks;
    noalloc_sumsq_kernel<<<nblk, kThreads, 0, st>>>(grads_dev, tb, scratch);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return pn_fail_global(PN_ERR_CUDA, "noalloc_sumsq_kernel: %s", cudaGetErrorString(e));
    if (momentum != 0.0f)
        noalloc_update_kernel<true><<<nblk, kThreads, 0, st>>>(params_dev, ms_dev, mom_dev, grads_dev, tb, scratch, lr,
                                                               decay, momentum, epsilon, clip_norm);
    else
        noalloc_update_kernel<false><<<nblk, kThreads, 0, st>>>(params_dev, ms_dev, nullptr, grads_dev, tb, scratch,
                                                                lr, decay, momentum, epsilon, clip_norm);
    e = cudaGetLastError();
    if (e != cudaSuccess) return pn_fail_global(PN_ERR_CUDA, "noalloc_update_kernel: %s", cudaGetErrorString(e));
    return PN_OK;
}
