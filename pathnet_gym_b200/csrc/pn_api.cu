// C ABI of libpathnet_b200.so (see include/pathnet_b200.h): handle lifecycle,
// slab layout, genotype / freeze tables and the entry points of the hot path.
#include "pn_internal.cuh"

#include <map>
#include <mutex>
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

int pn_fail(pn_handle* h, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (h) h->err = buf;
    pn_fail_global(code, "%s", buf);
    return code;
}

namespace {

inline long long align_up(long long v, long long a) { return (v + a - 1) / a * a; }

template <typename T>
int dev_alloc(pn_handle* h, T** p, size_t count) {
    const size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(p), bytes);
    if (e != cudaSuccess) {
        *p = nullptr;
        return pn_fail(h, PN_ERR_NOMEM, "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
    }
    h->workspace_bytes += (long long)bytes;
    return PN_OK;
}

int build_layout(pn_handle* h) {
    const pn_config& c = h->cfg;
    PnNetDesc& nd = h->nd;
    nd.L = c.L; nd.M = c.M; nd.hidden = c.hidden; nd.A = c.A; nd.nconv = c.L - 1;
    int ih = c.H, iw = c.W, cin = c.C;
    long long off = 0;
    for (int l = 0; l < nd.nconv; ++l) {
        const int k = c.kernel[l], s = c.stride[l];
        if (k < 1 || s < 1 || ih < k || iw < k)
            return pn_fail(h, PN_ERR_INVALID, "conv layer %d: kernel %d / stride %d does not fit input %dx%d", l, k, s,
                           ih, iw);
        nd.IH[l] = ih; nd.IW[l] = iw; nd.Cin[l] = cin; nd.k[l] = k; nd.s[l] = s;
        nd.OH[l] = (ih - k) / s + 1; nd.OW[l] = (iw - k) / s + 1;   // VALID padding, game_ac_network.py:110
        nd.KK[l] = k * k * cin;
        if (nd.KK[l] > 1024) return pn_fail(h, PN_ERR_INVALID, "conv layer %d: k*k*Cin = %d > 1024", l, nd.KK[l]);
        const long long wlen = (long long)nd.KK[l] * PN_F;
        nd.convW_base[l] = off;
        nd.convb_off[l] = (int)align_up(wlen, PN_ALIGN);
        nd.conv_stride[l] = nd.convb_off[l] + (int)align_up(PN_F, PN_ALIGN);
        for (int j = 0; j < c.M; ++j) {
            const long long wb = off + (long long)j * nd.conv_stride[l];
            h->tensors.push_back(pn_tensor_desc{wb, wlen, l, j, 0, 0});
            h->tensors.push_back(pn_tensor_desc{wb + nd.convb_off[l], PN_F, l, j, 1, 0});
        }
        off += (long long)c.M * nd.conv_stride[l];
        ih = nd.OH[l]; iw = nd.OW[l]; cin = PN_F;
    }
    nd.last_lin = ih * iw * cin;
    {
        const long long wlen = (long long)nd.last_lin * c.hidden;
        nd.linW_base = off;
        nd.linb_off = (int)align_up(wlen, PN_ALIGN);
        nd.lin_stride = nd.linb_off + (int)align_up(c.hidden, PN_ALIGN);
        for (int j = 0; j < c.M; ++j) {
            const long long wb = off + (long long)j * nd.lin_stride;
            h->tensors.push_back(pn_tensor_desc{wb, wlen, c.L - 1, j, 2, 0});
            h->tensors.push_back(pn_tensor_desc{wb + nd.linb_off, c.hidden, c.L - 1, j, 3, 0});
        }
        off += (long long)c.M * nd.lin_stride;
    }
    nd.fc2W = off; h->tensors.push_back(pn_tensor_desc{off, (long long)c.hidden * c.A, c.L, -1, 4, 0});
    off += align_up((long long)c.hidden * c.A, PN_ALIGN);
    nd.fc2b = off; h->tensors.push_back(pn_tensor_desc{off, c.A, c.L, -1, 5, 0});
    off += align_up(c.A, PN_ALIGN);
    nd.fc3W = off; h->tensors.push_back(pn_tensor_desc{off, c.hidden, c.L, -1, 6, 0});
    off += align_up(c.hidden, PN_ALIGN);
    nd.fc3b = off; h->tensors.push_back(pn_tensor_desc{off, 1, c.L, -1, 7, 0});
    off += PN_ALIGN;
    nd.out_div = (float)c.M;
    nd.lstmW = nd.lstmb = 0;
    if (c.flags & PN_FLAG_LSTM) {
        // GameACPathNetLSTMNetwork.get_vars(): ..., W_fc3, b_fc3, W_lstm, b_lstm (game_ac_network.py:502-520); no /M (:394)
        nd.out_div = 1.0f;
        nd.lstmW = off; h->tensors.push_back(pn_tensor_desc{off, 2ll * c.hidden * 4 * c.hidden, c.L + 1, -1, 8, 0});
        off += align_up(2ll * c.hidden * 4 * c.hidden, PN_ALIGN);
        nd.lstmb = off; h->tensors.push_back(pn_tensor_desc{off, 4ll * c.hidden, c.L + 1, -1, 9, 0});
        off += align_up(4ll * c.hidden, PN_ALIGN);
    }
    h->slab_floats = off;
    return PN_OK;
}

int check_ready(pn_handle* h, bool need_grads) {
    if (!h) return pn_fail_global(PN_ERR_INVALID, "null handle");
    if (!h->params) return pn_fail(h, PN_ERR_STATE, "pn_bind_slabs has not been called");
    if (need_grads && !h->grads) return pn_fail(h, PN_ERR_STATE, "no gradient slab bound");
    if (!h->have_masks) return pn_fail(h, PN_ERR_STATE, "pn_set_genotypes has not been called");
    return PN_OK;
}

int check_B(pn_handle* h, int B) {
    if (B < 1 || B > h->Bmax) return pn_fail(h, PN_ERR_INVALID, "B = %d outside 1..B_max = %d", B, h->Bmax);
    return PN_OK;
}

}  // namespace

int pn_set_smem_cached(pn_handle* h, const void* kern, int bytes) {
    static std::mutex mu;
    static std::map<std::pair<int, const void*>, int> have;       // (device, kernel) -> largest size set so far
    std::lock_guard<std::mutex> lock(mu);
    int dev = h->cfg.device;
    cudaGetDevice(&dev);                                            // the attribute applies to the CURRENT device
    int& cur = have[std::make_pair(dev, kern)];
    if (bytes > cur) {
        PN_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        cur = bytes;
    }
    return PN_OK;
}

extern "C" PN_API int pn_abi_version(void) { return PN_ABI_VERSION; }

extern "C" PN_API int pn_create(const pn_config* cfg, pn_handle** out) {
    if (!cfg || !out) return pn_fail_global(PN_ERR_INVALID, "pn_create: null argument");
    *out = nullptr;
    const pn_config& c = *cfg;
    if (c.L < 2 || c.L > PN_MAX_LAYERS) return pn_fail_global(PN_ERR_INVALID, "L = %d outside 2..%d", c.L, PN_MAX_LAYERS);
    if (c.M < 1 || c.M > PN_MAX_MODULES) return pn_fail_global(PN_ERR_INVALID, "M = %d outside 1..%d", c.M, PN_MAX_MODULES);
    if (c.F != PN_F) return pn_fail_global(PN_ERR_INVALID, "F = %d: only feature_num = 8 is supported", c.F);
    if (c.A < 1 || c.A > 31) return pn_fail_global(PN_ERR_INVALID, "A = %d outside 1..31", c.A);
    if (c.hidden < 4 || c.hidden > 1024 || (c.hidden % 4))
        return pn_fail_global(PN_ERR_INVALID, "hidden = %d must be a multiple of 4 in 4..1024", c.hidden);
    if (c.P < 1 || c.B_max < 1 || c.H < 1 || c.W < 1 || c.C < 1)
        return pn_fail_global(PN_ERR_INVALID, "P, B_max, H, W, C must be positive");
    if ((long long)c.P * c.B_max > (1ll << 24)) return pn_fail_global(PN_ERR_INVALID, "P*B_max too large");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return pn_fail_global(PN_ERR_CUDA, "no CUDA device: %s (this library has no CPU fallback)",
                              cudaGetErrorString(e));
    if (c.device < 0 || c.device >= ndev) return pn_fail_global(PN_ERR_INVALID, "device %d of %d", c.device, ndev);
    pn_handle* h = new (std::nothrow) pn_handle();
    if (!h) return pn_fail_global(PN_ERR_NOMEM, "out of host memory");
    h->cfg = c;
    h->P = c.P; h->Bmax = c.B_max; h->T = c.P * c.B_max;
    h->lstm = (c.flags & PN_FLAG_LSTM) != 0;
    if (h->lstm && c.hidden != 256) { delete h; return pn_fail_global(PN_ERR_INVALID, "the LSTM variant needs hidden = 256"); }
    int rc = build_layout(h);
    if (rc) { pn_fail_global(rc, "%s", h->err.c_str()); delete h; return rc; }
    const PnNetDesc& nd = h->nd;
#define PN_TRY(x) do { rc = (x); if (rc) { pn_fail_global(rc, "%s", h->err.c_str()); pn_destroy(h); return rc; } } while (0)
#define PN_TRYC(x) do { cudaError_t e2 = (x); if (e2 != cudaSuccess) { rc = pn_fail(h, PN_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e2)); pn_destroy(h); return rc; } } while (0)
    PN_TRYC(cudaSetDevice(c.device));
    const size_t T = (size_t)h->T;
    for (int l = 0; l < nd.nconv; ++l) {
        const size_t npix = (size_t)nd.OH[l] * nd.OW[l];
        PN_TRY(dev_alloc(h, &h->act[l], T * npix * PN_F));
        PN_TRY(dev_alloc(h, &h->dact[l], T * npix * PN_F));
        PN_TRY(dev_alloc(h, &h->rmask[l], T * c.M * npix));
    }
    h->nch = std::min(c.B_max, std::max(8, (296 + c.P - 1) / c.P));
    for (int l = 0; l < nd.nconv; ++l)
        PN_TRY(dev_alloc(h, &h->wpart[l], (size_t)c.P * h->nch * c.M * ((size_t)nd.KK[l] * PN_F + PN_F)));
    PN_TRY(dev_alloc(h, &h->hbuf, T * c.hidden));
    PN_TRY(dev_alloc(h, &h->lmask, T * c.M * c.hidden));
    PN_TRY(dev_alloc(h, &h->dzv, T * 32));
    PN_TRY(dev_alloc(h, &h->dh, T * c.hidden));
    PN_TRY(dev_alloc(h, &h->loss_t, T));
    if (h->lstm) PN_TRY(dev_alloc(h, &h->lstm_ws, pn_lstm_workspace_floats(h)));
    h->head_chunks = (int)((T + PN_HEAD_CHUNK - 1) / PN_HEAD_CHUNK);
    PN_TRY(dev_alloc(h, &h->head_part, (size_t)h->head_chunks * ((size_t)c.hidden * (c.A + 1) + c.A + 1)));
    PN_TRY(dev_alloc(h, &h->d_masks, (size_t)c.P * c.L));
    PN_TRY(dev_alloc(h, &h->d_list, (size_t)c.L * c.M * c.P));
    PN_TRY(dev_alloc(h, &h->d_list_count, (size_t)c.L * c.M));
    const int nt = (int)h->tensors.size();
    PN_TRY(dev_alloc(h, &h->d_frozen, (size_t)nt));
    h->h_frozen.assign(nt, 0);
    PN_TRYC(cudaMemset(h->d_frozen, 0, nt));
    // RMSProp chunk table
    std::vector<int> chunk0(nt), nchunks(nt);
    for (int t = 0; t < nt; ++t) {
        chunk0[t] = (int)h->h_chunks.size();
        for (long long o = 0; o < h->tensors[t].length; o += PN_RMS_CHUNK)
            h->h_chunks.push_back(PnRmsChunk{t, (int)std::min<long long>(PN_RMS_CHUNK, h->tensors[t].length - o),
                                             h->tensors[t].offset + o});
        nchunks[t] = (int)h->h_chunks.size() - chunk0[t];
    }
    PN_TRY(dev_alloc(h, &h->d_chunks, h->h_chunks.size()));
    PN_TRY(dev_alloc(h, &h->d_tensor_chunk0, (size_t)nt));
    PN_TRY(dev_alloc(h, &h->d_tensor_nchunks, (size_t)nt));
    PN_TRY(dev_alloc(h, &h->d_chunk_ss, h->h_chunks.size()));
    PN_TRYC(cudaMemcpy(h->d_chunks, h->h_chunks.data(), h->h_chunks.size() * sizeof(PnRmsChunk), cudaMemcpyHostToDevice));
    PN_TRYC(cudaMemcpy(h->d_tensor_chunk0, chunk0.data(), nt * sizeof(int), cudaMemcpyHostToDevice));
    h->h_tensor_chunk0 = chunk0;
    PN_TRYC(cudaMemcpy(h->d_tensor_nchunks, nchunks.data(), nt * sizeof(int), cudaMemcpyHostToDevice));
    // tensor-core path for layer 0 (u8 frames); cfg.flags bit 0 forces the CUDA-core kernels
    {
        cudaDeviceProp prop;
        PN_TRYC(cudaGetDeviceProperties(&prop, c.device));
        h->num_sms = prop.multiProcessorCount;
        h->tc_conv0 = !(c.flags & 1) && prop.major == 10 && pn_conv0_tc_supported(h);
        if (h->tc_conv0) PN_TRY(dev_alloc(h, &h->tc_w8, pn_conv0_tc_scratch_bytes(h)));
        h->tc_lin = !(c.flags & 2) && prop.major == 10 && pn_lin_tc_supported(h);
        if (h->tc_lin) PN_TRY(dev_alloc(h, &h->tc_lin_buf, pn_lin_tc_scratch_bytes(h)));
        for (int l = 1; l < nd.nconv; ++l) {
            h->tc_mid[l] = !(c.flags & 8) && !(c.flags & 4) && prop.major == 10 && pn_conv_mid_tc_supported(h, l);
            if (h->tc_mid[l]) PN_TRY(dev_alloc(h, &h->tc_mid_buf[l], pn_conv_mid_tc_scratch_bytes(h, l)));
        }
        if (h->tc_conv0 && getenv("PN_TC_DEBUG")) {
            PN_TRY(dev_alloc(h, &h->tc_dbg, (size_t)h->num_sms * 8 + 8));
            PN_TRYC(cudaMemset(h->tc_dbg, 0, ((size_t)h->num_sms * 8 + 8) * sizeof(long long)));
            const long long mode = atoll(getenv("PN_TC_DEBUG"));
            PN_TRYC(cudaMemcpy(h->tc_dbg + (size_t)h->num_sms * 8, &mode, sizeof(mode), cudaMemcpyHostToDevice));
        }
    }
    h->no_pdl = getenv("PN_NO_PDL") != nullptr;
    PN_TRYC(cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking));
    PN_TRYC(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
    PN_TRYC(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
    PN_TRYC(cudaEventCreateWithFlags(&h->ev_lin_grads, cudaEventDisableTiming));
    PN_TRYC(cudaEventCreateWithFlags(&h->ev_early_grads, cudaEventDisableTiming));
    // staging for the host-pointer entry point
    PN_TRY(dev_alloc(h, &h->st_actions, T));
    PN_TRY(dev_alloc(h, &h->st_td, T));
    PN_TRY(dev_alloc(h, &h->st_r, T));
    PN_TRY(dev_alloc(h, &h->st_loss, (size_t)c.P));
    PN_TRYC(cudaDeviceSynchronize());
#undef PN_TRY
#undef PN_TRYC
    *out = h;
    return PN_OK;
}

extern "C" PN_API void pn_destroy(pn_handle* h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    for (int l = 0; l < PN_MAX_LAYERS; ++l) {
        cudaFree(h->act[l]); cudaFree(h->dact[l]); cudaFree(h->rmask[l]); cudaFree(h->wpart[l]);
    }
    cudaFree(h->lstm_ws); cudaFree(h->hbuf); cudaFree(h->lmask); cudaFree(h->dzv); cudaFree(h->dh); cudaFree(h->loss_t);
    cudaFree(h->head_part); cudaFree(h->d_masks); cudaFree(h->d_list); cudaFree(h->d_list_count);
    cudaFree(h->d_frozen); cudaFree(h->d_chunks); cudaFree(h->d_tensor_chunk0); cudaFree(h->d_tensor_nchunks);
    cudaFree(h->tc_w8); cudaFree(h->tc_dbg); cudaFree(h->tc_lin_buf);
    for (int l = 0; l < PN_MAX_LAYERS; ++l) cudaFree(h->tc_mid_buf[l]);
    cudaFree(h->d_chunk_ss); cudaFree(h->st_frames); cudaFree(h->st_actions); cudaFree(h->st_td);
    cudaFree(h->st_r); cudaFree(h->st_loss);
    for (auto& sl : h->slot) {
        cudaFree(sl.frames); cudaFree(sl.planes); cudaFree(sl.actions); cudaFree(sl.td); cudaFree(sl.r); cudaFree(sl.loss);
        if (sl.copied) cudaEventDestroy(sl.copied);
        if (sl.consumed) cudaEventDestroy(sl.consumed);
        for (auto& e : sl.part) if (e) cudaEventDestroy(e);
    }
    if (h->stage_aux) cudaStreamDestroy(h->stage_aux);
    if (h->side) cudaStreamDestroy(h->side);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    if (h->ev_lin_grads) cudaEventDestroy(h->ev_lin_grads);
    if (h->ev_early_grads) cudaEventDestroy(h->ev_early_grads);
    for (auto& e : h->prof.ev) cudaEventDestroy(e);
    delete h;
}

extern "C" PN_API const char* pn_last_error(const pn_handle* h) { return h ? h->err.c_str() : pn_last_global_error(); }
extern "C" PN_API int64_t pn_slab_floats(const pn_handle* h) { return h ? h->slab_floats : 0; }
extern "C" PN_API int pn_num_tensors(const pn_handle* h) { return h ? (int)h->tensors.size() : 0; }
extern "C" PN_API int pn_last_lin(const pn_handle* h) { return h ? h->nd.last_lin : 0; }
extern "C" PN_API int64_t pn_workspace_bytes(const pn_handle* h) { return h ? h->workspace_bytes : 0; }

extern "C" PN_API int pn_debug_read(pn_handle* h, long long* out_host, int capacity) {
    if (!h || !h->tc_dbg) return 0;
    const int n = h->num_sms * 8;
    if (capacity < n) return -1;
    cudaDeviceSynchronize();
    cudaMemcpy(out_host, h->tc_dbg, (size_t)n * sizeof(long long), cudaMemcpyDeviceToHost);
    return n;
}

extern "C" PN_API int64_t pn_debug_read_relu_masks(pn_handle* h, int layer, int B, uint8_t* out_host, int64_t capacity_bytes) {
    if (!h || !out_host) return pn_fail_global(PN_ERR_INVALID, "pn_debug_read_relu_masks: null argument");
    if (layer < 0 || layer >= h->cfg.L) return pn_fail(h, PN_ERR_INVALID, "layer %d outside 0..%d", layer, h->cfg.L - 1);
    if (int rc = check_B(h, B)) return rc;
    const PnNetDesc& nd = h->nd;
    const bool lin = layer == nd.nconv;
    const size_t per = lin ? (size_t)nd.hidden : (size_t)nd.OH[layer] * nd.OW[layer];
    const size_t bytes = (size_t)h->P * B * nd.M * per;
    if ((int64_t)bytes > capacity_bytes) return pn_fail(h, PN_ERR_INVALID, "capacity %lld < %zu bytes", (long long)capacity_bytes, bytes);
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    PN_CUDA(h, cudaDeviceSynchronize());
    PN_CUDA(h, cudaMemcpy(out_host, lin ? h->lmask : h->rmask[layer], bytes, cudaMemcpyDeviceToHost));
    return (int64_t)bytes;
}

extern "C" PN_API int pn_engine_paths(const pn_handle* h, char* out, int capacity) {
    if (!h || !out || capacity < 1) return pn_fail_global(PN_ERR_INVALID, "pn_engine_paths: null argument");
    std::string s;
    for (int l = 0; l < h->nd.nconv; ++l) {
        const bool tc = l == 0 ? h->tc_conv0 : h->tc_mid[l];
        s += "conv" + std::to_string(l) + (tc ? "=tcgen05" : "=cuda-core") + (l == 0 && tc ? " (u8 frames; fp32 frames take the cuda-core kernel)," : ",");
    }
    s += std::string("lin=") + (h->tc_lin ? "tcgen05" : "cuda-core");
    snprintf(out, (size_t)capacity, "%s", s.c_str());
    return (int)s.size();
}

extern "C" PN_API int64_t pn_launch_count(const pn_handle* h) { return h ? h->launches : 0; }

extern "C" PN_API int pn_profile_enable(pn_handle* h, int on) {
    if (!h) return pn_fail_global(PN_ERR_INVALID, "null handle");
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    if (on && h->prof.ev.empty()) {
        h->prof.ev.resize((size_t)PN_PROF_NSLOTS * PN_PROF_RING * 2);
        for (auto& e : h->prof.ev) PN_CUDA(h, cudaEventCreate(&e));
    }
    for (int& c : h->prof.count) c = 0;
    h->prof.on = on != 0;
    return PN_OK;
}

extern "C" PN_API int pn_profile_read(pn_handle* h, pn_profile_entry* out, int capacity) {
    if (!h || !out) return pn_fail_global(PN_ERR_INVALID, "pn_profile_read: null argument");
    if (h->prof.ev.empty()) return 0;
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    PN_CUDA(h, cudaDeviceSynchronize());
    int n = 0;
    for (int slot = 0; slot < PN_PROF_NSLOTS; ++slot) {
        const int cnt = std::min(h->prof.count[slot], (int)PN_PROF_RING);
        if (cnt == 0) continue;
        if (n >= capacity) return pn_fail(h, PN_ERR_INVALID, "pn_profile_read: capacity %d too small", capacity);
        float total = 0.f;
        for (int i = 0; i < cnt; ++i) {
            float ms = 0.f;
            PN_CUDA(h, cudaEventElapsedTime(&ms, h->prof.ev[((size_t)slot * PN_PROF_RING + i) * 2],
                                            h->prof.ev[((size_t)slot * PN_PROF_RING + i) * 2 + 1]));
            total += ms;
        }
        const char* base; int layer = -1;
        if (slot < PN_PROF_CONV_WGRAD) { base = "conv_fwd"; layer = slot - PN_PROF_CONV_FWD; }
        else if (slot < PN_PROF_CONV_WGRAD_REDUCE) { base = "conv_wgrad"; layer = slot - PN_PROF_CONV_WGRAD; }
        else if (slot < PN_PROF_CONV_DGRAD) { base = "conv_wgrad_reduce"; layer = slot - PN_PROF_CONV_WGRAD_REDUCE; }
        else if (slot < PN_PROF_LIN_FWD) { base = "conv_dgrad"; layer = slot - PN_PROF_CONV_DGRAD; }
        else {
            static const char* names[] = {"lin_fwd", "head_loss", "grad_reduce", "lin_dgrad", "lin_wgrad",
                                          "rms_sumsq", "rms_update", "conv0_wprep", "conv0_dquant", "lin_prep",
                                          "conv_mid_prep", "lin_wprep", "rms_update_lin", "rms_update_small", "lstm_fwd", "lstm_bwd"};
            base = names[slot - PN_PROF_LIN_FWD];
        }
        if (layer >= 0) snprintf(out[n].name, sizeof(out[n].name), "%s_l%d", base, layer);
        else snprintf(out[n].name, sizeof(out[n].name), "%s", base);
        out[n].total_ms = total;
        out[n].launches = cnt;
        ++n;
    }
    return n;
}

extern "C" PN_API int pn_tensor_table(const pn_handle* h, pn_tensor_desc* out, int capacity) {
    if (!h || !out) return pn_fail_global(PN_ERR_INVALID, "pn_tensor_table: null argument");
    const int n = (int)h->tensors.size();
    if (capacity < n) return pn_fail_global(PN_ERR_INVALID, "pn_tensor_table: capacity %d < %d", capacity, n);
    std::memcpy(out, h->tensors.data(), n * sizeof(pn_tensor_desc));
    return n;
}

extern "C" PN_API int pn_bind_slabs(pn_handle* h, float* params_dev, float* ms_dev, float* mom_dev, float* grads_dev) {
    if (!h) return pn_fail_global(PN_ERR_INVALID, "null handle");
    if (!params_dev) return pn_fail(h, PN_ERR_INVALID, "params slab is null");
    for (const void* p : {(const void*)params_dev, (const void*)ms_dev, (const void*)mom_dev, (const void*)grads_dev})
        if (p && (reinterpret_cast<uintptr_t>(p) & 15)) return pn_fail(h, PN_ERR_INVALID, "slabs must be 16-byte aligned");
    h->params = params_dev; h->ms = ms_dev; h->mom = mom_dev; h->grads = grads_dev;
    return PN_OK;
}

extern "C" PN_API int pn_set_genotypes(pn_handle* h, const uint32_t* masks_host, void* stream) {
    if (!h || !masks_host) return pn_fail_global(PN_ERR_INVALID, "pn_set_genotypes: null argument");
    const int P = h->P, L = h->cfg.L, M = h->cfg.M;
    const uint32_t valid = M == 32 ? 0xFFFFFFFFu : ((1u << M) - 1u);
    for (int i = 0; i < P * L; ++i)
        if (masks_host[i] & ~valid) return pn_fail(h, PN_ERR_INVALID, "mask[%d] = 0x%x has bits >= M = %d", i, masks_host[i], M);
    h->h_masks.assign(masks_host, masks_host + (size_t)P * L);
    for (int l = 0; l < L; ++l) {
        int mx = 0;
        for (int p = 0; p < P; ++p) mx = std::max(mx, __builtin_popcount(masks_host[p * L + l]));
        h->max_active[l] = mx;
    }
    std::vector<int> list((size_t)L * M * P, 0), cnt((size_t)L * M, 0);
    for (int p = 0; p < P; ++p)
        for (int l = 0; l < L; ++l)
            for (int j = 0; j < M; ++j)
                if ((masks_host[p * L + l] >> j) & 1u) list[(size_t)(l * M + j) * P + cnt[l * M + j]++] = p;
    cudaStream_t st = (cudaStream_t)stream;
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    PN_CUDA(h, cudaMemcpyAsync(h->d_masks, h->h_masks.data(), (size_t)P * L * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    PN_CUDA(h, cudaMemcpyAsync(h->d_list, list.data(), list.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    PN_CUDA(h, cudaMemcpyAsync(h->d_list_count, cnt.data(), cnt.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    PN_CUDA(h, cudaStreamSynchronize(st));   // the host vectors above go out of scope
    h->have_masks = true;
    return PN_OK;
}

extern "C" PN_API int pn_lin_grads_region(const pn_handle* h, int64_t* offset_floats, int64_t* count_floats) {
    if (!h || !offset_floats || !count_floats) return pn_fail_global(PN_ERR_INVALID, "pn_lin_grads_region: null argument");
    *offset_floats = h->nd.linW_base;
    *count_floats = h->nd.fc2W - h->nd.linW_base;      // the head tensors follow the linear modules in the slab
    return PN_OK;
}

extern "C" PN_API int pn_stream_wait_lin_grads(pn_handle* h, void* stream) {
    if (!h) return pn_fail_global(PN_ERR_INVALID, "pn_stream_wait_lin_grads: null handle");
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    PN_CUDA(h, cudaStreamWaitEvent((cudaStream_t)stream, h->ev_lin_grads, 0));
    return PN_OK;
}

long long pn_late_grads_floats(const pn_handle* h) { return h->nd.nconv > 1 ? h->nd.convW_base[1] : h->nd.linW_base; }

extern "C" PN_API int pn_stream_wait_early_grads(pn_handle* h, void* stream, int64_t* late_floats) {
    if (!h) return pn_fail_global(PN_ERR_INVALID, "pn_stream_wait_early_grads: null handle");
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    PN_CUDA(h, cudaStreamWaitEvent((cudaStream_t)stream, h->ev_early_grads, 0));
    if (late_floats) *late_floats = pn_late_grads_floats(h);
    return PN_OK;
}

extern "C" PN_API int pn_lstm_bind_state(pn_handle* h, const float* state_in_dev, float* state_out_dev) {
    if (!h) return pn_fail_global(PN_ERR_INVALID, "null handle");
    if (!h->lstm) return pn_fail(h, PN_ERR_STATE, "this handle was not created with PN_FLAG_LSTM");
    h->lstm_state_in = state_in_dev;
    h->lstm_state_out = state_out_dev;
    return PN_OK;
}

extern "C" PN_API int pn_set_frozen(pn_handle* h, const uint8_t* frozen_host) {
    if (!h || !frozen_host) return pn_fail_global(PN_ERR_INVALID, "pn_set_frozen: null argument");
    const int nt = (int)h->tensors.size();
    for (int t = 0; t < nt; ++t) h->h_frozen[t] = frozen_host[t] ? 1 : 0;
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    // rare (task switch): wait for every update / re-init kernel in flight on any stream before the table changes
    PN_CUDA(h, cudaDeviceSynchronize());
    PN_CUDA(h, cudaMemcpy(h->d_frozen, h->h_frozen.data(), nt, cudaMemcpyHostToDevice));
    return PN_OK;
}

extern "C" PN_API int pn_forward(pn_handle* h, const void* frames_dev, int frames_are_u8, int B, float* pi_dev, float* v_dev,
                          void* stream) {
    if (int rc = check_ready(h, false)) return rc;
    if (int rc = check_B(h, B)) return rc;
    if (!frames_dev) return pn_fail(h, PN_ERR_INVALID, "frames is null");
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t st = (cudaStream_t)stream;
    if (int rc = pn_launch_forward(h, frames_dev, frames_are_u8, B, false, st)) return rc;
    if (h->lstm)
        if (int rc = pn_launch_lstm_fwd(h, B, st)) return rc;
    return pn_launch_heads(h, B, false, nullptr, nullptr, nullptr, 0.f, pi_dev, v_dev, nullptr, st);
}

extern "C" PN_API int pn_train_step(pn_handle* h, const void* frames_dev, int frames_are_u8, const int32_t* actions_dev,
                             const float* td_dev, const float* r_dev, int B, float entropy_beta, float* loss_dev,
                             float* pi_dev, float* v_dev, void* stream) {
    if (int rc = check_ready(h, true)) return rc;
    if (int rc = check_B(h, B)) return rc;
    if (!frames_dev || !actions_dev || !td_dev || !r_dev) return pn_fail(h, PN_ERR_INVALID, "null input array");
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t st = (cudaStream_t)stream;
    if (int rc = pn_launch_forward(h, frames_dev, frames_are_u8, B, true, st)) return rc;
    if (h->lstm)
        if (int rc = pn_launch_lstm_fwd(h, B, st)) return rc;
    if (int rc = pn_launch_heads(h, B, true, actions_dev, td_dev, r_dev, entropy_beta, pi_dev, v_dev, loss_dev, st))
        return rc;
    if (h->lstm)
        if (int rc = pn_launch_lstm_bwd(h, B, st)) return rc;
    return pn_launch_backward(h, frames_dev, frames_are_u8, B, st);
}

extern "C" PN_API int pn_train_step_host(pn_handle* h, const void* frames_host, int frames_are_u8, const int32_t* actions_host,
                                  const float* td_host, const float* r_host, int B, float entropy_beta,
                                  float* loss_host, void* stream) {
    if (int rc = check_ready(h, true)) return rc;
    if (int rc = check_B(h, B)) return rc;
    if (!frames_host || !actions_host || !td_host || !r_host) return pn_fail(h, PN_ERR_INVALID, "null input array");
    cudaStream_t st = (cudaStream_t)stream;
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    const size_t T = (size_t)h->P * B;
    const size_t fbytes = T * h->cfg.H * h->cfg.W * h->cfg.C * (frames_are_u8 ? 1 : sizeof(float));
    if (h->st_frames_bytes < fbytes) {
        const size_t want = (size_t)h->T * h->cfg.H * h->cfg.W * h->cfg.C * (frames_are_u8 ? 1 : sizeof(float));
        PN_CUDA(h, cudaStreamSynchronize(st));
        cudaFree(h->st_frames);
        h->st_frames = nullptr; h->st_frames_bytes = 0;
        cudaError_t e = cudaMalloc(&h->st_frames, want);
        if (e != cudaSuccess) return pn_fail(h, PN_ERR_NOMEM, "frame staging cudaMalloc(%zu): %s", want, cudaGetErrorString(e));
        h->st_frames_bytes = want;
        h->workspace_bytes += (long long)want;
    }
    PN_CUDA(h, cudaMemcpyAsync(h->st_frames, frames_host, fbytes, cudaMemcpyHostToDevice, st));
    PN_CUDA(h, cudaMemcpyAsync(h->st_actions, actions_host, T * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    PN_CUDA(h, cudaMemcpyAsync(h->st_td, td_host, T * sizeof(float), cudaMemcpyHostToDevice, st));
    PN_CUDA(h, cudaMemcpyAsync(h->st_r, r_host, T * sizeof(float), cudaMemcpyHostToDevice, st));
    if (int rc = pn_train_step(h, h->st_frames, frames_are_u8, h->st_actions, h->st_td, h->st_r, B, entropy_beta,
                               h->st_loss, nullptr, nullptr, stream))
        return rc;
    if (loss_host)
        PN_CUDA(h, cudaMemcpyAsync(loss_host, h->st_loss, (size_t)h->P * sizeof(float), cudaMemcpyDeviceToHost, st));
    PN_CUDA(h, cudaStreamSynchronize(st));
    return PN_OK;
}

namespace {
// staging slot set-up shared by the two host-input forms: events, small arrays, frame buffer of the full [P, B_max] size
int prepare_slot(pn_handle* h, pn_handle::StageSlot& s, size_t want_frames) {
    const size_t Tmax = (size_t)h->T;
    if (!s.copied) {
        PN_CUDA(h, cudaEventCreateWithFlags(&s.copied, cudaEventDisableTiming));
        PN_CUDA(h, cudaEventCreateWithFlags(&s.consumed, cudaEventDisableTiming));
        PN_CUDA(h, cudaMalloc(&s.actions, Tmax * sizeof(int32_t)));
        PN_CUDA(h, cudaMalloc(&s.td, Tmax * sizeof(float)));
        PN_CUDA(h, cudaMalloc(&s.r, Tmax * sizeof(float)));
        PN_CUDA(h, cudaMalloc(&s.loss, (size_t)h->P * sizeof(float)));
        h->workspace_bytes += (long long)(Tmax * 12 + (size_t)h->P * 4);
    }
    if (s.frames_bytes < want_frames) {
        PN_CUDA(h, cudaDeviceSynchronize());
        cudaFree(s.frames);
        s.frames = nullptr; s.frames_bytes = 0;
        cudaError_t e = cudaMalloc(&s.frames, want_frames);
        if (e != cudaSuccess) return pn_fail(h, PN_ERR_NOMEM, "frame staging cudaMalloc(%zu): %s", want_frames, cudaGetErrorString(e));
        s.frames_bytes = want_frames;
        h->workspace_bytes += (long long)want_frames;
    }
    return PN_OK;
}
}  // namespace

extern "C" PN_API int pn_stage_inputs_host(pn_handle* h, int slot, const void* frames_host, int frames_are_u8,
                                           const int32_t* actions_host, const float* td_host, const float* r_host, int B,
                                           void* copy_stream) {
    if (!h) return pn_fail_global(PN_ERR_INVALID, "null handle");
    if (slot < 0 || slot > 1) return pn_fail(h, PN_ERR_INVALID, "slot must be 0 or 1");
    if (int rc = check_B(h, B)) return rc;
    if (!frames_host || !actions_host || !td_host || !r_host) return pn_fail(h, PN_ERR_INVALID, "null input array");
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    pn_handle::StageSlot& s = h->slot[slot];
    cudaStream_t cs = (cudaStream_t)copy_stream;
    const size_t Tmax = (size_t)h->T, T = (size_t)h->P * B;
    const size_t esz = frames_are_u8 ? 1 : sizeof(float);
    const size_t want = Tmax * h->cfg.H * h->cfg.W * h->cfg.C * esz, fbytes = T * h->cfg.H * h->cfg.W * h->cfg.C * esz;
    if (int rc = prepare_slot(h, s, want)) return rc;
    // do not overwrite the slot while the step that last used it is still reading it
    if (s.used) PN_CUDA(h, cudaStreamWaitEvent(cs, s.consumed, 0));
    PN_CUDA(h, cudaMemcpyAsync(s.frames, frames_host, fbytes, cudaMemcpyHostToDevice, cs));
    PN_CUDA(h, cudaMemcpyAsync(s.actions, actions_host, T * sizeof(int32_t), cudaMemcpyHostToDevice, cs));
    PN_CUDA(h, cudaMemcpyAsync(s.td, td_host, T * sizeof(float), cudaMemcpyHostToDevice, cs));
    PN_CUDA(h, cudaMemcpyAsync(s.r, r_host, T * sizeof(float), cudaMemcpyHostToDevice, cs));
    PN_CUDA(h, cudaEventRecord(s.copied, cs));
    s.B = B; s.is_u8 = frames_are_u8; s.used = true;
    return PN_OK;
}

extern "C" PN_API int pn_stage_rollout_host(pn_handle* h, int slot, const uint8_t* planes_host, const int32_t* actions_host,
                                            const float* td_host, const float* r_host, int B, void* copy_stream) {
    if (!h) return pn_fail_global(PN_ERR_INVALID, "null handle");
    if (slot < 0 || slot > 1) return pn_fail(h, PN_ERR_INVALID, "slot must be 0 or 1");
    if (int rc = check_B(h, B)) return rc;
    if (!planes_host || !actions_host || !td_host || !r_host) return pn_fail(h, PN_ERR_INVALID, "null input array");
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    pn_handle::StageSlot& s = h->slot[slot];
    cudaStream_t cs = (cudaStream_t)copy_stream;
    const size_t Tmax = (size_t)h->T, T = (size_t)h->P * B;
    const size_t HW = (size_t)h->cfg.H * h->cfg.W, C = (size_t)h->cfg.C;
    if (int rc = prepare_slot(h, s, Tmax * HW * C)) return rc;
    const size_t pwant = (size_t)h->P * (h->Bmax + C - 1) * HW;
    if (s.planes_bytes < pwant) {
        PN_CUDA(h, cudaDeviceSynchronize());
        cudaFree(s.planes);
        s.planes = nullptr; s.planes_bytes = 0;
        cudaError_t e = cudaMalloc(&s.planes, pwant);
        if (e != cudaSuccess) return pn_fail(h, PN_ERR_NOMEM, "plane staging cudaMalloc(%zu): %s", pwant, cudaGetErrorString(e));
        s.planes_bytes = pwant;
        h->workspace_bytes += (long long)pwant;
    }
    if (!h->stage_aux) PN_CUDA(h, cudaStreamCreateWithFlags(&h->stage_aux, cudaStreamNonBlocking));
    for (auto& e : s.part)
        if (!e) PN_CUDA(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    if (s.used) PN_CUDA(h, cudaStreamWaitEvent(cs, s.consumed, 0));
    PN_CUDA(h, cudaMemcpyAsync(s.actions, actions_host, T * sizeof(int32_t), cudaMemcpyHostToDevice, cs));
    PN_CUDA(h, cudaMemcpyAsync(s.td, td_host, T * sizeof(float), cudaMemcpyHostToDevice, cs));
    PN_CUDA(h, cudaMemcpyAsync(s.r, r_host, T * sizeof(float), cudaMemcpyHostToDevice, cs));
    // All host reads are on `cs`; the stacks of one group of paths are rebuilt on the auxiliary stream while the next
    // group's planes cross PCIe.  `copied` (what pn_train_step_staged waits for) follows the last rebuild.
    const size_t per_path = (size_t)(B + C - 1) * HW;
    const int parts = h->P < PN_STAGE_PARTS ? h->P : PN_STAGE_PARTS;
    for (int j = 0; j < parts; ++j) {
        const int p0 = (int)((long long)h->P * j / parts), p1 = (int)((long long)h->P * (j + 1) / parts);
        PN_CUDA(h, cudaMemcpyAsync(s.planes + p0 * per_path, planes_host + p0 * per_path, (p1 - p0) * per_path,
                                   cudaMemcpyHostToDevice, cs));
        PN_CUDA(h, cudaEventRecord(s.part[j], cs));
        PN_CUDA(h, cudaStreamWaitEvent(h->stage_aux, s.part[j], 0));
        if (int rc = pn_launch_expand_planes(h, s.planes, B, (uint8_t*)s.frames, p0, p1 - p0, h->stage_aux)) return rc;
    }
    PN_CUDA(h, cudaEventRecord(s.copied, h->stage_aux));
    s.B = B; s.is_u8 = 1; s.used = true;
    return PN_OK;
}

extern "C" PN_API int pn_train_step_staged(pn_handle* h, int slot, float entropy_beta, float* loss_host, void* stream) {
    if (int rc = check_ready(h, true)) return rc;
    if (slot < 0 || slot > 1 || !h->slot[slot].used) return pn_fail(h, PN_ERR_STATE, "slot %d has not been staged", slot);
    pn_handle::StageSlot& s = h->slot[slot];
    cudaStream_t st = (cudaStream_t)stream;
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    PN_CUDA(h, cudaStreamWaitEvent(st, s.copied, 0));
    if (int rc = pn_train_step(h, s.frames, s.is_u8, s.actions, s.td, s.r, s.B, entropy_beta, s.loss, nullptr, nullptr, stream))
        return rc;
    PN_CUDA(h, cudaEventRecord(s.consumed, st));
    if (loss_host) PN_CUDA(h, cudaMemcpyAsync(loss_host, s.loss, (size_t)h->P * sizeof(float), cudaMemcpyDeviceToHost, st));
    return PN_OK;
}

extern "C" PN_API int pn_rmsprop_apply(pn_handle* h, float lr, float decay, float momentum, float epsilon, float clip_norm,
                                void* stream) {
    if (!h) return pn_fail_global(PN_ERR_INVALID, "null handle");
    if (!h->params || !h->ms || !h->grads) return pn_fail(h, PN_ERR_STATE, "params / rms / grads slabs must be bound");
    if (momentum != 0.0f && !h->mom) return pn_fail(h, PN_ERR_STATE, "momentum != 0 needs a momentum slab");
    if (!(clip_norm > 0.0f)) return pn_fail(h, PN_ERR_INVALID, "clip_norm must be > 0");
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    return pn_launch_rmsprop(h, 0, (int)h->tensors.size(), false, lr, decay, momentum, epsilon, clip_norm, (cudaStream_t)stream);
}

extern "C" PN_API int pn_rmsprop_apply_range(pn_handle* h, int first_tensor, int n_tensors, int complement, float lr, float decay,
                                             float momentum, float epsilon, float clip_norm, void* stream) {
    if (!h) return pn_fail_global(PN_ERR_INVALID, "null handle");
    if (!h->params || !h->ms || !h->grads) return pn_fail(h, PN_ERR_STATE, "params / rms / grads slabs must be bound");
    if (momentum != 0.0f && !h->mom) return pn_fail(h, PN_ERR_STATE, "momentum != 0 needs a momentum slab");
    if (!(clip_norm > 0.0f)) return pn_fail(h, PN_ERR_INVALID, "clip_norm must be > 0");
    const int nt = (int)h->tensors.size();
    if (first_tensor < 0 || n_tensors < 0 || first_tensor + n_tensors > nt)
        return pn_fail(h, PN_ERR_INVALID, "tensor range [%d, %d) outside 0..%d", first_tensor, first_tensor + n_tensors, nt);
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    return pn_launch_rmsprop(h, first_tensor, n_tensors, complement != 0, lr, decay, momentum, epsilon, clip_norm,
                             (cudaStream_t)stream);
}

extern "C" PN_API int pn_reinit_unfrozen(pn_handle* h, const float* backup_params_dev, void* stream) {
    if (!h) return pn_fail_global(PN_ERR_INVALID, "null handle");
    if (!h->params || !backup_params_dev) return pn_fail(h, PN_ERR_STATE, "params slab and backup must be set");
    PN_CUDA(h, cudaSetDevice(h->cfg.device));
    return pn_launch_reinit(h, backup_params_dev, (cudaStream_t)stream);
}
