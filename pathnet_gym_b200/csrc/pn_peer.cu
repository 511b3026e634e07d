// Cross-GPU sum of the small gradient regions (conv layers + heads, ~140 KB) over NVLink peer memory, in ONE kernel.
//
// Why: in a sharded population the linear modules' gradients (97 % of the bytes) are all-reduced by NCCL
// underneath the conv layers' backward pass, but the conv and head gradients only exist when the step's last kernel has
// run, so their all-reduce sits exposed at the end of every step and is pure latency (two dependent NCCL kernels'
// worth of launch + handshake for 140 KB).  Here every GPU publishes its values in a buffer its peers can read
// (symmetric memory: the same allocation mapped into every process of the node), raises a flag on every peer, waits
// for the peers' flags, and adds up all copies itself in rank order -- one launch, and bit-identical on every rank.
//
//   exchange buffer of a rank: [256 x u32 control words][slot 0: n floats][slot 1: n floats]
//   control words: [0..31] flags of the small-region sum, [32] its CTA counter ([136..151], [152] and [160..175], [176] for its
//                  early and late parts when the sum is split: pn_peer_sum_part);
//                  [64..95] / [96..127] "gradients ready" / "sums written" flags of the linear-region all-reduce, [128] its CTA counter
//   flags[r] = number of steps rank r has published (monotonic).  Step s uses slot s & 1: a rank publishes step s + 1 only
//   after it has finished reading every peer's step-s slot, and nobody can overwrite slot s & 1 (step s + 2) before all
//   ranks have published s + 1, so two slots are enough.
// The grid must be co-resident (the CTAs spin): it is capped at one CTA per SM, far below that in practice.
#include "pn_internal.cuh"

#include <algorithm>

namespace {

constexpr int kPeerFlagWords = 256;    // control words in front of the data slots (see the layout above)
constexpr int kPeerMaxWorld = 16;

struct PeerPtrs {
    uint64_t buf[kPeerMaxWorld];
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float ld_peer(const float* p) {       // never from a stale L1 line
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

// Bounded wait for a peer's flag.  Every rank must make the same call with the same step; if one does not (it raised an
// error before launching, died, or passed another step), an unbounded spin would hold this GPU inside a kernel forever
// with no host-side error.  After kSpinBudgetNs the waiter records the failure in control word kPeerErrWord of ITS OWN
// buffer and goes on (the sums of that step are then garbage); pn_peer_check reports it to the host.  The loop backs off
// with nanosleep so that the spinning CTAs leave issue slots to the kernels they share the SMs with.
constexpr unsigned long long kSpinBudgetNs = 2000000000ull;     // 2 s
constexpr int kPeerErrWord = 200;
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void wait_flag(const unsigned* f, unsigned target, unsigned* err_word, unsigned code) {
    if ((int)(ld_acquire_sys(f) - target) >= 0) return;
    const unsigned long long t0 = global_timer_ns();
    unsigned it = 0;
    while ((int)(ld_acquire_sys(f) - target) < 0) {
        __nanosleep(64);
        if ((++it & 255u) == 0 && global_timer_ns() - t0 > kSpinBudgetNs) { atomicExch(err_word, code); return; }
    }
}

// slab index of element e of the small set: the conv region [0, lin_off) followed by the head region after the linear one
__device__ __forceinline__ long long small_index(int e, long long lin_off, long long lin_cnt) {
    return e < lin_off ? (long long)e : (long long)e + lin_cnt;
}

__global__ void __launch_bounds__(256)
peer_sum_small_kernel(float* __restrict__ grads, long long lin_off, long long lin_cnt, int n, int e_lo, int e_hi, int flag_word,
                      int counter_word, PeerPtrs peers, int rank, int world, unsigned step) {
    // elements [e_lo, e_hi) of the small set; flags and the CTA counter of this part at control words flag_word / counter_word
    const int slot = step & 1u;
    unsigned* cta_counter = reinterpret_cast<unsigned*>(peers.buf[rank]) + counter_word;   // lives in the buffer: one count per buffer
    float* mine = reinterpret_cast<float*>(peers.buf[rank]) + kPeerFlagWords + (size_t)slot * n;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
    // A: publish my values (16 bytes per thread when the regions allow it: n, lin_off, lin_cnt multiples of 4)
    const bool vec = ((n | lin_off | lin_cnt | e_lo | e_hi) & 3) == 0;
    if (vec) {
        for (int e = e_lo + tid * 4; e < e_hi; e += nthreads * 4)
            *reinterpret_cast<float4*>(mine + e) = *reinterpret_cast<const float4*>(grads + small_index(e, lin_off, lin_cnt));
    } else {
        for (int e = e_lo + tid; e < e_hi; e += nthreads) mine[e] = grads[small_index(e, lin_off, lin_cnt)];
    }
    __threadfence_system();
    __syncthreads();
    __shared__ unsigned last;
    if (threadIdx.x == 0) last = atomicAdd(cta_counter, 1u) == step * gridDim.x + gridDim.x - 1 ? 1u : 0u;   // counter is monotonic
    __syncthreads();
    if (last) __threadfence_system();     // every CTA's values (fenced before its atomicAdd) are ordered before the flag
    if (last && threadIdx.x < world) {
        // the last CTA to finish writing raises my flag on every rank (including myself)
        unsigned* f = reinterpret_cast<unsigned*>(peers.buf[threadIdx.x]) + flag_word + rank;
        st_release_sys(f, step + 1u);
    }
    // B: wait until every rank has published this step
    if (threadIdx.x < world) {
        unsigned* mine_ctrl = reinterpret_cast<unsigned*>(peers.buf[rank]);
        wait_flag(mine_ctrl + flag_word + threadIdx.x, step + 1u, mine_ctrl + kPeerErrWord, 1u);
    }
    __syncthreads();
    // C: fixed-order sum over the ranks.  All ranks' loads are issued before the first add (an NVLink round trip is
    // ~2 us: summing rank by rank would pay it `world` times per element).
    if (vec) {
        for (int e = e_lo + tid * 4; e < e_hi; e += nthreads * 4) {
            float4 v[kPeerMaxWorld];
#pragma unroll
            for (int r = 0; r < kPeerMaxWorld; ++r) {
                v[r] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < world)
                    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                                 : "=f"(v[r].x), "=f"(v[r].y), "=f"(v[r].z), "=f"(v[r].w)
                                 : "l"(reinterpret_cast<const float*>(peers.buf[r]) + kPeerFlagWords + (size_t)slot * n + e)
                                 : "memory");
            }
            float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int r = 0; r < kPeerMaxWorld; ++r)
                if (r < world) { s4.x += v[r].x; s4.y += v[r].y; s4.z += v[r].z; s4.w += v[r].w; }
            *reinterpret_cast<float4*>(grads + small_index(e, lin_off, lin_cnt)) = s4;
        }
    } else {
        for (int e = e_lo + tid; e < e_hi; e += nthreads) {
            float v[kPeerMaxWorld];
#pragma unroll
            for (int r = 0; r < kPeerMaxWorld; ++r)
                v[r] = r < world ? ld_peer(reinterpret_cast<const float*>(peers.buf[r]) + kPeerFlagWords + (size_t)slot * n + e) : 0.f;
            float s1 = 0.f;
#pragma unroll
            for (int r = 0; r < kPeerMaxWorld; ++r)
                if (r < world) s1 += v[r];
            grads[small_index(e, lin_off, lin_cnt)] = s1;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// All-reduce of the linear modules' gradient region (97 % of the slab) in place, over peer memory.  The gradient slabs
// themselves are symmetric memory here, so nothing is staged:
//   1. rank r tells every peer "my linear gradients are final" (stream order guarantees it at kernel start)
//   2. waits for all peers, then owns shard r of the region: reads shard r of every rank's slab over NVLink, adds the
//      copies in rank order and writes the sum into shard r of EVERY rank's slab (reduce-scatter and all-gather in one pass)
//   3. tells every peer "my shard is written everywhere" and waits for all peers before the kernel ends, so the update
//      kernel that follows in the stream reads complete sums and nobody overwrites a slab a peer still reads.
// Only rank r reads or writes shard r between the two barriers, so there is no write/read overlap; the sums are
// bit-identical on all ranks.  The CTAs spin: the grid is kept far below one CTA per SM.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
peer_allreduce_lin_kernel(PeerPtrs slabs, PeerPtrs ctrl, long long lin_off, long long lin_cnt, int rank, int world, unsigned step) {
    unsigned* my_ctrl = reinterpret_cast<unsigned*>(ctrl.buf[rank]);
    // 1. gradients ready
    if (blockIdx.x == 0 && threadIdx.x < world) {
        __threadfence_system();
        st_release_sys(reinterpret_cast<unsigned*>(ctrl.buf[threadIdx.x]) + 64 + rank, step + 1u);
    }
    if (threadIdx.x < world) wait_flag(my_ctrl + 64 + threadIdx.x, step + 1u, my_ctrl + kPeerErrWord, 2u);
    __syncthreads();
    // 2. my shard: float4 granularity
    const long long n4 = lin_cnt >> 2;
    const long long lo4 = n4 * rank / world, hi4 = n4 * (rank + 1) / world;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (long long)gridDim.x * blockDim.x;
    // four float4 per thread and rank in flight (NVLink latency ~2 us: the loop is all about bytes in flight)
    constexpr int kU = 4;
    for (long long e0 = lo4 + tid; e0 < hi4; e0 += nthreads * kU) {
        float4 s[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) s[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int r = 0; r < world; ++r) {
            const float4* base = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(slabs.buf[r]) + lin_off);
            float4 v[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const long long e = e0 + u * nthreads;
                v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (e < hi4)
                    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                                 : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(base + e) : "memory");
            }
#pragma unroll
            for (int u = 0; u < kU; ++u) { s[u].x += v[u].x; s[u].y += v[u].y; s[u].z += v[u].z; s[u].w += v[u].w; }
        }
        for (int r = 0; r < world; ++r) {
            float4* base = reinterpret_cast<float4*>(reinterpret_cast<float*>(slabs.buf[r]) + lin_off);
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const long long e = e0 + u * nthreads;
                if (e < hi4) base[e] = s[u];
            }
        }
    }
    if (rank == world - 1 && blockIdx.x == 0)             // tail that is not a multiple of 4 floats (never with 32-float aligned tensors)
        for (long long e = (n4 << 2) + threadIdx.x; e < lin_cnt; e += blockDim.x) {
            float s = 0.f;
            for (int r = 0; r < world; ++r) s += ld_peer(reinterpret_cast<const float*>(slabs.buf[r]) + lin_off + e);
            for (int r = 0; r < world; ++r) (reinterpret_cast<float*>(slabs.buf[r]) + lin_off)[e] = s;
        }
    // 3. sums written everywhere
    __threadfence_system();
    __syncthreads();
    __shared__ unsigned last;
    if (threadIdx.x == 0) last = atomicAdd(my_ctrl + 128, 1u) == step * gridDim.x + gridDim.x - 1 ? 1u : 0u;
    __syncthreads();
    if (last) __threadfence_system();
    if (last && threadIdx.x < world) st_release_sys(reinterpret_cast<unsigned*>(ctrl.buf[threadIdx.x]) + 96 + rank, step + 1u);
    if (threadIdx.x < world) wait_flag(my_ctrl + 96 + threadIdx.x, step + 1u, my_ctrl + kPeerErrWord, 3u);
    __syncthreads();
}

}  // namespace

extern "C" PN_API int pn_peer_allreduce_lin(pn_handle* h, const uint64_t* peer_slabs_host, const uint64_t* peer_bufs_host, int rank,
                                            int world, uint32_t step, void* stream) {
    if (!h || !peer_slabs_host || !peer_bufs_host) return pn_fail_global(PN_ERR_INVALID, "pn_peer_allreduce_lin: null argument");
    if (world < 1 || world > kPeerMaxWorld || rank < 0 || rank >= world)
        return pn_fail(h, PN_ERR_INVALID, "pn_peer_allreduce_lin: rank %d of %d (at most %d ranks)", rank, world, kPeerMaxWorld);
    if (!h->grads) return pn_fail(h, PN_ERR_STATE, "gradient slab not bound");
    if ((uint64_t)(uintptr_t)h->grads != peer_slabs_host[rank])
        return pn_fail(h, PN_ERR_STATE, "pn_peer_allreduce_lin: peer_slabs[rank] is not the bound gradient slab");
    const long long lin_off = h->nd.linW_base, lin_cnt = h->nd.fc2W - h->nd.linW_base;
    if (lin_off & 3) return pn_fail(h, PN_ERR_STATE, "linear region not 16-byte aligned");
    PeerPtrs ps, pc;
    for (int r = 0; r < kPeerMaxWorld; ++r) { ps.buf[r] = r < world ? peer_slabs_host[r] : 0; pc.buf[r] = r < world ? peer_bufs_host[r] : 0; }
    const int grid = std::max(1, std::min(h->num_sms / 4, (int)((lin_cnt / world + 4095) / 4096)));
    peer_allreduce_lin_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(ps, pc, lin_off, lin_cnt, rank, world, step);
    PN_LAUNCH_CHECK(h, "peer_allreduce_lin_kernel");
    return PN_OK;
}

extern "C" PN_API int pn_peer_check(pn_handle* h, const void* my_peer_buf_dev) {
    if (!h || !my_peer_buf_dev) return pn_fail_global(PN_ERR_INVALID, "pn_peer_check: null argument");
    unsigned code = 0;
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    PN_CUDA(h, cudaMemcpy(&code, reinterpret_cast<const unsigned*>(my_peer_buf_dev) + kPeerErrWord, sizeof(code), cudaMemcpyDeviceToHost));
    if (code)
        return pn_fail(h, PN_ERR_STATE, "a peer-memory exchange kernel gave up waiting for another rank after 2 s (wait %u: "
                                        "1 = small regions published, 2 = linear gradients ready, 3 = linear sums written): "
                                        "the ranks did not all make the same call with the same step", code);
    return PN_OK;
}

extern "C" PN_API int64_t pn_peer_buffer_bytes(const pn_handle* h) {
    if (!h) return -1;
    const long long n = h->slab_floats - (h->nd.fc2W - h->nd.linW_base);
    return (int64_t)(kPeerFlagWords * 4 + 2 * n * 4);
}

namespace {
// part 0: all small regions; part 1: those final before the first conv layer's weight-gradient kernel (middle conv layers,
// heads, LSTM); part 2: the first conv layer's tensors.  Each part has its own flags and CTA counter.
int launch_peer_sum(pn_handle* h, const uint64_t* peer_bufs_host, int rank, int world, uint32_t step, int part, cudaStream_t st) {
    if (!h || !peer_bufs_host) return pn_fail_global(PN_ERR_INVALID, "pn_peer_sum: null argument");
    if (world < 1 || world > kPeerMaxWorld || rank < 0 || rank >= world)
        return pn_fail(h, PN_ERR_INVALID, "pn_peer_sum: rank %d of %d (at most %d ranks)", rank, world, kPeerMaxWorld);
    if (part < 0 || part > 2) return pn_fail(h, PN_ERR_INVALID, "pn_peer_sum: part %d (0, 1 or 2)", part);
    if (!h->grads) return pn_fail(h, PN_ERR_STATE, "gradient slab not bound");
    const long long lin_off = h->nd.linW_base, lin_cnt = h->nd.fc2W - h->nd.linW_base;
    const int n = (int)(h->slab_floats - lin_cnt);
    const int late = (int)pn_late_grads_floats(h);
    const int e_lo = part == 1 ? late : 0, e_hi = part == 2 ? late : n;
    if (e_hi <= e_lo) return PN_OK;
    static const int kFlagWord[3] = {0, 136, 160}, kCounterWord[3] = {32, 152, 176};
    PeerPtrs pp;
    for (int r = 0; r < kPeerMaxWorld; ++r) pp.buf[r] = r < world ? peer_bufs_host[r] : 0;
    // a function of the handle and the part only: the same every step (the CTA counter relies on it)
    const int grid = std::max(1, std::min(h->num_sms / 2, (e_hi - e_lo + 1023) / 1024));
    peer_sum_small_kernel<<<grid, 256, 0, st>>>(h->grads, lin_off, lin_cnt, n, e_lo, e_hi, kFlagWord[part], kCounterWord[part], pp,
                                                rank, world, step);
    PN_LAUNCH_CHECK(h, "peer_sum_small_kernel");
    return PN_OK;
}
}  // namespace

extern "C" PN_API int pn_peer_sum_small(pn_handle* h, const uint64_t* peer_bufs_host, int rank, int world, uint32_t step,
                                        void* stream) {
    return launch_peer_sum(h, peer_bufs_host, rank, world, step, 0, (cudaStream_t)stream);
}

extern "C" PN_API int pn_peer_sum_part(pn_handle* h, const uint64_t* peer_bufs_host, int rank, int world, uint32_t step, int part,
                                       void* stream) {
    return launch_peer_sum(h, peer_bufs_host, rank, world, step, part, (cudaStream_t)stream);
}
