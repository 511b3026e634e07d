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
#include "pn_rms_core.cuh"

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

    // chunk -> the same description the handle form reads from its device tables, so both forms run the same
    // kernel bodies (pn_rms_core.cuh) and give bit-identical results
    __device__ __forceinline__ pnrms::ChunkRef chunk(int cid) const {
        int lo = 0, hi = n - 1;             // last t with chunk0[t] <= cid
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (chunk0[mid] <= cid) lo = mid; else hi = mid - 1;
        }
        const int t = lo, c = cid - chunk0[t];
        pnrms::ChunkRef r;
        r.tensor = t;
        r.off = off[t] + (long long)c * PN_RMS_CHUNK;
        r.len = (int)min((long long)PN_RMS_CHUNK, (long long)len[t] - (long long)c * PN_RMS_CHUNK);
        r.frozen = frozen[t] != 0;
        r.chunk0 = chunk0[t];
        r.nchunks = chunk0[t + 1] - chunk0[t];
        return r;
    }
};

using pnrms::kThreads;

__global__ void __launch_bounds__(kThreads)
noalloc_sumsq_kernel(const float* __restrict__ grads, const __grid_constant__ NoallocTable tb,
                     float* __restrict__ chunk_ss) {
    pnrms::sumsq_body(tb, (int)blockIdx.x, grads, chunk_ss);
}

template <bool HAS_MOM>
__global__ void __launch_bounds__(kThreads)
noalloc_update_kernel(float* __restrict__ params, float* __restrict__ msl, float* __restrict__ moml,
                      const float* __restrict__ grads, const __grid_constant__ NoallocTable tb,
                      const float* __restrict__ chunk_ss, float lr, float decay, float momentum, float eps,
                      float clip) {
    pnrms::update_body<HAS_MOM>(tb, (int)blockIdx.x, params, msl, moml, grads, chunk_ss, lr, decay, momentum, eps, clip);
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
    // (launched so that its blocks may start their loads while the first kernel drains; see pn_rms_core.cuh)
    if (momentum != 0.0f)
        e = pnrms::launch_pdl(noalloc_update_kernel<true>, nblk, st, params_dev, ms_dev, mom_dev, grads_dev, tb,
                              (const float*)scratch, lr, decay, momentum, epsilon, clip_norm);
    else
        e = pnrms::launch_pdl(noalloc_update_kernel<false>, nblk, st, params_dev, ms_dev, (float*)nullptr, grads_dev, tb,
                              (const float*)scratch, lr, decay, momentum, epsilon, clip_norm);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return pn_fail_global(PN_ERR_CUDA, "noalloc_update_kernel: %s", cudaGetErrorString(e));
    return PN_OK;
}
