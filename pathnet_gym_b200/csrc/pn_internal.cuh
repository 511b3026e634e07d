// Internal declarations shared by the translation units of libpathnet_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <unordered_map>
#include <vector>

#include "pathnet_b200.h"

#define PN_F 8            // conv output channels per module (feature_num, game_ac_network.py:132)
#define PN_ALIGN 32       // slab alignment of every tensor, in floats
#define PN_HEAD_CHUNK 16  // transitions per partial of the head weight gradient
#define PN_RMS_CHUNK 2048 // floats per RMSProp work chunk (never straddles a tensor)

struct PnNetDesc {        // passed by value to kernels
    int L, M, hidden, A, nconv, last_lin;
    int IH[PN_MAX_LAYERS], IW[PN_MAX_LAYERS], Cin[PN_MAX_LAYERS], k[PN_MAX_LAYERS], s[PN_MAX_LAYERS];
    int OH[PN_MAX_LAYERS], OW[PN_MAX_LAYERS], KK[PN_MAX_LAYERS];
    long long convW_base[PN_MAX_LAYERS];   // float offset of W_conv[l,0]
    int conv_stride[PN_MAX_LAYERS];        // floats between module j and j+1 of layer l
    int convb_off[PN_MAX_LAYERS];          // b_conv[l,j] = W_conv[l,j] + convb_off[l]
    long long linW_base; int lin_stride; int linb_off;
    long long fc2W, fc2b, fc3W, fc3b;
    long long lstmW, lstmb;                // LSTM variant only: kernel [512,1024], bias [1024] (game_ac_network.py:437-438)
    float out_div;                         // the last layer's module sum is divided by this: M (feed-forward net,
                                           // game_ac_network.py:194) or 1 (LSTM net, :394)
};

struct PnRmsChunk { int tensor; int len; long long off; };

// per-kernel CUDA-event timing (pn_profile_enable / pn_profile_read)
enum PnProfSlot {
    PN_PROF_CONV_FWD = 0,                              // + layer
    PN_PROF_CONV_WGRAD = PN_MAX_LAYERS,                // + layer
    PN_PROF_CONV_WGRAD_REDUCE = 2 * PN_MAX_LAYERS,     // + layer
    PN_PROF_CONV_DGRAD = 3 * PN_MAX_LAYERS,            // + layer
    PN_PROF_LIN_FWD = 4 * PN_MAX_LAYERS,
    PN_PROF_HEAD,
    PN_PROF_HEAD_WGRAD,
    PN_PROF_LIN_DGRAD,
    PN_PROF_LIN_WGRAD,
    PN_PROF_RMS_SUMSQ,
    PN_PROF_RMS_UPDATE,
    PN_PROF_CONV0_WPREP,
    PN_PROF_CONV0_DQUANT,
    PN_PROF_LIN_PREP,
    PN_PROF_CONV_MID_PREP,
    PN_PROF_LIN_WPREP,
    PN_PROF_RMS_UPDATE_LIN,       // sumsq + update pair over the linear modules' tensors
    PN_PROF_RMS_UPDATE_SMALL,     // the same pair over everything else (conv + head tensors)
    PN_PROF_LSTM_FWD,             // input projection + the per-step recurrence kernels
    PN_PROF_LSTM_BWD,             // BPTT step kernels + weight / input gradient GEMMs
    PN_PROF_NSLOTS
};
#define PN_PROF_RING 64
struct PnProf {
    bool on = false;
    std::vector<cudaEvent_t> ev;      // [slot][ring][2], created lazily
    int count[PN_PROF_NSLOTS] = {};
};

#define PN_STAGE_PARTS 4

struct pn_handle {
    pn_config cfg{};
    PnNetDesc nd{};
    int P = 0, Bmax = 0, T = 0;
    std::vector<pn_tensor_desc> tensors;
    long long slab_floats = 0;
    // bound slabs (borrowed)
    float *params = nullptr, *ms = nullptr, *mom = nullptr, *grads = nullptr;
    // genotype state
    bool have_masks = false;
    std::vector<uint32_t> h_masks;
    uint32_t* d_masks = nullptr;        // [P*L]
    int* d_list = nullptr;              // [L*M*P] paths using module (l,j), ascending
    int* d_list_count = nullptr;        // [L*M]
    // freeze table
    std::vector<uint8_t> h_frozen;
    uint8_t* d_frozen = nullptr;        // [num_tensors]
    // workspace
    float* act[PN_MAX_LAYERS] = {};     // conv layer outputs [T, npix, F]
    float* dact[PN_MAX_LAYERS] = {};    // gradients w.r.t. act[l]
    uint8_t* rmask[PN_MAX_LAYERS] = {}; // [T][M][npix] one byte = 8 channel bits
    float* hbuf = nullptr;              // [T, hidden]
    uint8_t* lmask = nullptr;           // [T][M][hidden] 0/1
    float* dzv = nullptr;               // [T][32]: dL/dlogits (k<A), dL/dv at k==A
    float* dh = nullptr;                // [T, hidden] gradient w.r.t. the module sum (already /M)
    float* loss_t = nullptr;            // [T]
    float* wpart[PN_MAX_LAYERS] = {};   // conv wgrad partials [P*nch][M][KK*F+F]
    int nch = 1;                        // transition chunks per path in conv wgrad
    float* pending_loss = nullptr;      // per-path loss output of the step in flight (written by the head reduce launch)
    float* head_part = nullptr;         // [head_chunks][hidden*(A+1) + (A+1)]
    int head_chunks = 0;
    int head_ctas = 0;                  // CTAs of the last head-kernel launch = partial slices it wrote
    // rmsprop
    std::vector<PnRmsChunk> h_chunks;
    std::vector<int> h_tensor_chunk0;   // host copy of d_tensor_chunk0
    PnRmsChunk* d_chunks = nullptr;
    int* d_tensor_chunk0 = nullptr;     // first chunk of each tensor
    int* d_tensor_nchunks = nullptr;
    float* d_chunk_ss = nullptr;        // [n_chunks] partial sums of squares
    // tensor-core conv0 (pn_conv0_tc.cu)
    bool tc_mid[PN_MAX_LAYERS] = {};    // middle conv layer l runs on the tensor cores (cfg.flags bit 3 clear)
    uint8_t* tc_mid_buf[PN_MAX_LAYERS] = {};   // bf16 hi/lo weight operand blocks of that layer
    int max_active[PN_MAX_LAYERS] = {}; // largest number of active modules of any path, per layer (from pn_set_genotypes)
    cudaStream_t side = nullptr;        // weight-operand preparation runs here, beside the first conv layer
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaEvent_t ev_early_grads = nullptr;   // every gradient except the first conv layer's is final (see pn_stream_wait_early_grads)
    cudaEvent_t ev_lin_grads = nullptr; // recorded after the linear weight-gradient launch of every training step
    bool absmax_fused = false;          // the layer-1 data-gradient launch of this step already wrote max |dY0|
    bool tc_conv0 = false;              // geometry supported and not disabled by cfg.flags bit 0
    int num_sms = 148;
    uint8_t* tc_w8 = nullptr;           // [M][6144] int8 weight digits + [M][8] fp32 scales + batch max|dY|
    bool tc_lin = false;                // tensor-core linear modules (pn_lin_tc.cu)
    uint8_t* tc_lin_buf = nullptr;      // Wp | Xp | Dp  bf16 hi/lo block buffers
    long long* tc_dbg = nullptr;        // optional [num_sms][8] cycle counters (PN_TC_DEBUG=1), else null
    int last_wgrad_nch = 1;             // chunks per path used by the last tensor-core weight-gradient launch
    // host-call staging
    void* st_frames = nullptr; size_t st_frames_bytes = 0;
    int32_t* st_actions = nullptr; float* st_td = nullptr; float* st_r = nullptr; float* st_loss = nullptr;
    // double-buffered staging for the pipelined host path (pn_stage_inputs_host / pn_train_step_staged)
    cudaStream_t stage_aux = nullptr;   // rebuilds the stacks of one group of paths while the next group crosses PCIe
    struct StageSlot {
        void* frames = nullptr; size_t frames_bytes = 0;
        uint8_t* planes = nullptr; size_t planes_bytes = 0;     // pn_stage_rollout_host: the rollout's unique planes
        int32_t* actions = nullptr; float* td = nullptr; float* r = nullptr; float* loss = nullptr;
        cudaEvent_t copied = nullptr, consumed = nullptr;
        cudaEvent_t part[PN_STAGE_PARTS] = {};                  // pn_stage_rollout_host: one per group of paths copied
        bool used = false; int B = 0; int is_u8 = 1;
    } slot[2];
    long long workspace_bytes = 0;
    // LSTM variant (cfg.flags bit 4; pn_lstm.cu)
    bool lstm = false;
    float* lstm_ws = nullptr;           // gx | gate activations | dA | c | tanh c | h_out | h_prev | dc carry
    const float* lstm_state_in = nullptr;   // [P][2][256] (c, h) at the first step of every path, or null = zeros (borrowed)
    float* lstm_state_out = nullptr;        // [P][2][256] state after the last step, or null (borrowed)
    bool no_pdl = false;                // PN_NO_PDL: ordinary launches everywhere (comparison)
    long long launches = 0;             // kernels launched through this handle
    PnProf prof;
    std::string err;
};

struct PnProfScope {                    // brackets one kernel launch with events when profiling is on
    pn_handle* h; int slot; cudaStream_t st;
    PnProfScope(pn_handle* h_, int slot_, cudaStream_t st_) : h(h_), slot(slot_), st(st_) {
        if (h->prof.on) cudaEventRecord(h->prof.ev[((size_t)slot * PN_PROF_RING + h->prof.count[slot] % PN_PROF_RING) * 2], st);
    }
    ~PnProfScope() {
        if (h->prof.on) {
            cudaEventRecord(h->prof.ev[((size_t)slot * PN_PROF_RING + h->prof.count[slot] % PN_PROF_RING) * 2 + 1], st);
            ++h->prof.count[slot];
        }
    }
};

int pn_fail(pn_handle* h, int code, const char* fmt, ...);
int pn_fail_global(int code, const char* fmt, ...);

#define PN_CUDA(h, call)                                                                      \
    do {                                                                                      \
        cudaError_t e__ = (call);                                                             \
        if (e__ != cudaSuccess)                                                               \
            return pn_fail((h), PN_ERR_CUDA, "%s failed: %s (%s:%d)", #call,                  \
                           cudaGetErrorString(e__), __FILE__, __LINE__);                      \
    } while (0)

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) only when this process has not yet set at least `bytes` for the kernel
// on the handle's device (the call costs several microseconds of host time; a training step used to make twelve)
int pn_set_smem_cached(pn_handle* h, const void* kern, int bytes);
#define PN_SET_SMEM(h, kern, bytes)                                                        \
    do {                                                                                   \
        if (int rc__ = pn_set_smem_cached((h), (const void*)(kern), (int)(bytes))) return rc__; \
    } while (0)

// Programmatic dependent launch along a chain of kernels on one stream: the next kernel is launched while this one runs
// and waits -- pn_pdl_sync(), the first thing it does -- until this one has completed and its writes are visible, so
// the launch latency between two dependent kernels (~2.7 us) overlaps with the kernel before.  A kernel launched the
// ordinary way executes pn_pdl_sync() as a no-op.  Off while per-kernel timing is on and with PN_NO_PDL set.
#ifdef __CUDACC__
__device__ __forceinline__ void pn_pdl_sync() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
template <class... KArgs, class... Args>
inline cudaError_t pn_launch(bool pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif
inline bool pn_use_pdl(const pn_handle* h) { return !h->prof.on && !h->no_pdl; }

#define PN_LAUNCH_CHECK(h, name)                                                              \
    do {                                                                                      \
        ++(h)->launches;                                                                      \
        cudaError_t e__ = cudaGetLastError();                                                 \
        if (e__ != cudaSuccess)                                                               \
            return pn_fail((h), PN_ERR_CUDA, "launch of %s failed: %s", name,                 \
                           cudaGetErrorString(e__));                                          \
    } while (0)

// kernels' host launchers (pn_net_kernels.cu)
int pn_launch_forward(pn_handle* h, const void* frames, int is_u8, int B, bool training, cudaStream_t st);
int pn_launch_heads(pn_handle* h, int B, bool training, const int32_t* actions, const float* td, const float* r,
                    float beta, float* pi, float* v, float* loss, cudaStream_t st);
int pn_launch_backward(pn_handle* h, const void* frames, int is_u8, int B, cudaStream_t st);
// pn_conv0_tc.cu
bool pn_conv0_tc_supported(const pn_handle* h);
size_t pn_conv0_tc_scratch_bytes(const pn_handle* h);
unsigned* pn_conv0_absmax_slot(const pn_handle* h);   // nullptr when the tensor-core conv0 path is off
int pn_launch_conv0_fwd_tc(pn_handle* h, const uint8_t* frames, int B, cudaStream_t st);
bool pn_conv0_wgrad_tc_supported(const pn_handle* h);
int pn_launch_conv0_wgrad_tc(pn_handle* h, const uint8_t* frames, int B, cudaStream_t st);
// pn_conv_mid_tc.cu
bool pn_conv_mid_tc_supported(const pn_handle* h, int l);
size_t pn_conv_mid_tc_scratch_bytes(const pn_handle* h, int l);
int pn_launch_conv_mid_wprep(pn_handle* h, int l, cudaStream_t st);
int pn_launch_conv_mid_fwd_tc(pn_handle* h, int l, int B, cudaStream_t st);
int pn_launch_conv_mid_dgrad_tc(pn_handle* h, int l, int B, cudaStream_t st);
bool pn_conv_mid_wgrad_tc_usable(const pn_handle* h, int l);
bool pn_conv_mid_tc_has_bwd(const pn_handle* h, int l);
int pn_launch_conv_mid_wgrad_tc(pn_handle* h, int l, int B, cudaStream_t st, int* nch_used);
// pn_conv_mid.cu
bool pn_conv_mid_supported(const pn_handle* h, int l);
int pn_launch_conv_mid_fwd(pn_handle* h, int l, int B, cudaStream_t st);
int pn_launch_conv_mid_bwd(pn_handle* h, int l, int B, cudaStream_t st, int* nch_used, bool do_wgrad, bool do_dgrad);
// pn_lin_tc.cu
bool pn_lin_tc_supported(const pn_handle* h);
size_t pn_lin_tc_scratch_bytes(const pn_handle* h);
int pn_launch_lin_wsplit(pn_handle* h, cudaStream_t st);
int pn_launch_lin_fwd_tc(pn_handle* h, int B, cudaStream_t st);
int pn_launch_lin_bwd_tc(pn_handle* h, int B, cudaStream_t st);
// pn_lstm.cu
size_t pn_lstm_workspace_floats(const pn_handle* h);
int pn_launch_lstm_fwd(pn_handle* h, int B, cudaStream_t st);
int pn_launch_lstm_bwd(pn_handle* h, int B, cudaStream_t st);
inline const float* pn_head_input(const pn_handle* h) {      // what the two heads read: the module sum, or the cell outputs
    return h->lstm ? h->lstm_ws + (size_t)h->T * 1024 * 3 + (size_t)h->T * 256 * 2 : h->hbuf;
}
// pn_rollout.cu
// pn_gemm_tc.cu: plain GEMMs on the tensor cores (bf16 hi/lo block planes), used by the LSTM variant
struct PnGemmOperand { const uint8_t* planes; int R, C; bool mn; };
size_t pn_gemm_planes_bytes(int R, int C);
int pn_gemm_split(pn_handle* h, const float* S, int R, int C, long long ld, uint8_t* planes, cudaStream_t st);
bool pn_gemm_tc_usable(int M, int N, int K, int nsplit);
int pn_gemm_tc(pn_handle* h, const PnGemmOperand& A, const PnGemmOperand& B, float* C, long long ldc, const float* bias, int M,
               int N, int K, int nsplit, cudaStream_t st);
long long pn_late_grads_floats(const pn_handle* h);   // floats at the start of the slab that ev_early_grads does not cover
int pn_launch_expand_planes(pn_handle* h, const uint8_t* planes_dev, int B, uint8_t* frames_dev, int p0, int np, cudaStream_t st);
// pn_rmsprop.cu
int pn_launch_rmsprop(pn_handle* h, int first_tensor, int n_tensors, bool complement, float lr, float decay,
                      float momentum, float eps, float clip, cudaStream_t st);
int pn_launch_reinit(pn_handle* h, const float* backup, cudaStream_t st);

// ---------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al.), shared by host-less genotype kernels
// ---------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ void pn_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                          uint32_t k0, uint32_t k1, uint32_t& o0, uint32_t& o1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    o0 = c0; o1 = c1;
}
__device__ __forceinline__ uint64_t pn_philox_u64(uint64_t seed, uint32_t epoch, uint32_t path, uint32_t pos,
                                                  uint32_t domain, uint32_t slot) {
    uint32_t o0, o1;
    pn_philox(epoch, path, pos, (domain << 24) | (slot & 0xFFFFFFu), (uint32_t)seed, (uint32_t)(seed >> 32), o0, o1);
    return ((uint64_t)o0 << 32) | o1;
}
__device__ __forceinline__ double pn_philox_uniform(uint64_t seed, uint32_t epoch, uint32_t path, uint32_t pos,
                                                    uint32_t domain, uint32_t slot) {
    return (double)(pn_philox_u64(seed, epoch, path, pos, domain, slot) >> 11) * 0x1.0p-53;
}
#endif
