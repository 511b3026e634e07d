// Per-tensor clip_by_norm + RMSProp over the fp32 parameter / accumulator slabs (handle form).
// The kernels' bodies, the reference semantics they follow (rmsprop_applier.py:79-106) and the traffic model are in
// pn_rms_core.cuh; this file supplies the chunk table from the handle's device arrays and the launcher.
#include "pn_internal.cuh"
#include "pn_rms_core.cuh"

namespace {

using pnrms::kThreads;

struct HandleTab {
    const PnRmsChunk* chunks; const uint8_t* frozen; const int* tchunk0; const int* tnchunks;
    __device__ __forceinline__ pnrms::ChunkRef chunk(int cid) const {
        const PnRmsChunk ck = chunks[cid];
        pnrms::ChunkRef r;
        r.tensor = ck.tensor; r.len = ck.len; r.off = ck.off; r.frozen = frozen[ck.tensor] != 0;
        r.chunk0 = tchunk0[ck.tensor]; r.nchunks = tnchunks[ck.tensor];
        return r;
    }
};

// two chunk segments in one launch: [c0, c0 + n0) then [c1, ...)
__global__ void __launch_bounds__(kThreads)
rms_sumsq_kernel(HandleTab tb, const float* __restrict__ grads, float* __restrict__ chunk_ss, int c0, int n0, int c1) {
    pnrms::pdl_wait();          // launched programmatically behind the kernel that finishes the gradients (pn_pdl_sync)
    const int cid = (int)blockIdx.x < n0 ? c0 + (int)blockIdx.x : c1 + ((int)blockIdx.x - n0);
    pnrms::sumsq_body(tb, cid, grads, chunk_ss);
}

template <bool HAS_MOM>
__global__ void __launch_bounds__(kThreads)
rms_update_kernel(HandleTab tb, float* __restrict__ params, float* __restrict__ msl, float* __restrict__ moml,
                  const float* __restrict__ grads, const float* __restrict__ chunk_ss, float lr, float decay,
                  float momentum, float eps, float clip, int c0, int n0, int c1) {
    const int cid = (int)blockIdx.x < n0 ? c0 + (int)blockIdx.x : c1 + ((int)blockIdx.x - n0);
    pnrms::update_body<HAS_MOM>(tb, cid, params, msl, moml, grads, chunk_ss, lr, decay, momentum, eps, clip);
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
    // which selection this is, for the per-kernel timing table: the linear modules' tensors, everything else, or other
    int lin_first = -1, lin_count = 0;
    for (int t = 0; t < nt; ++t)
        if (h->tensors[t].kind == 2 || h->tensors[t].kind == 3) { if (lin_first < 0) lin_first = t; ++lin_count; }
    const bool is_lin = first_tensor == lin_first && n_tensors == lin_count;
    const int slot = is_lin ? (complement ? PN_PROF_RMS_UPDATE_SMALL : PN_PROF_RMS_UPDATE_LIN) : PN_PROF_RMS_UPDATE;
    const HandleTab tb{h->d_chunks, h->d_frozen, h->d_tensor_chunk0, h->d_tensor_nchunks};
    PnProfScope prof(h, slot, st);                  // ONE bracket around the pair: an event between them would undo the overlap
    pn_launch(pn_use_pdl(h), rms_sumsq_kernel, dim3(nchunks), dim3(kThreads), 0, st, tb, (const float*)h->grads, h->d_chunk_ss, c0, n0, c1);
    PN_LAUNCH_CHECK(h, "rms_sumsq_kernel");
    cudaError_t e;
    if (momentum != 0.0f)
        e = pnrms::launch_pdl(rms_update_kernel<true>, nchunks, st, tb, h->params, h->ms, h->mom, (const float*)h->grads,
                              (const float*)h->d_chunk_ss, lr, decay, momentum, eps, clip, c0, n0, c1);
    else
        e = pnrms::launch_pdl(rms_update_kernel<false>, nchunks, st, tb, h->params, h->ms, (float*)nullptr, (const float*)h->grads,
                              (const float*)h->d_chunk_ss, lr, decay, momentum, eps, clip, c0, n0, c1);
    if (e != cudaSuccess) return pn_fail(h, PN_ERR_CUDA, "launch of rms_update_kernel failed: %s", cudaGetErrorString(e));
    PN_LAUNCH_CHECK(h, "rms_update_kernel");
    return PN_OK;
}

int pn_launch_reinit(pn_handle* h, const float* backup, cudaStream_t st) {
    const int nchunks = (int)h->h_chunks.size();
    reinit_kernel<<<nchunks, kThreads, 0, st>>>(h->params, backup, h->d_chunks, h->d_frozen);
    PN_LAUNCH_CHECK(h, "reinit_kernel");
    return PN_OK;
}
