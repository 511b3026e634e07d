// Development probe (not part of the library): dense tcgen05 kind::i8 throughput on this GPU, as the denominator for the
// layer-0 kernels' tensor-side roofline.  One CTA per SM issues M = 128, N = 256, K = 32 MMAs (u8 x s8 -> s32, operands in
// shared memory in the no-swizzle canonical layout this library uses) back to back into two alternating accumulators.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -I.. i8_peak.cu -o i8_peak && ./i8_peak
// Prints one JSON line: burst (50 ms) and sustained (3 s) rates in dense int8 TOP/s (2 ops per multiply-accumulate).
#include <cstdio>
#include <cstdlib>
#include "../pn_tc.cuh"
using namespace pntc;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__global__ void __launch_bounds__(128, 1) i8_loop(int reps, int ksteps) {
    extern __shared__ __align__(1024) uint8_t sm[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase;
    // A: 128 rows x (ksteps * 32) bytes, B: 256 rows x (ksteps * 32) bytes; K-major, no swizzle (core matrix = 8 rows x 16 B)
    const int a_bytes = 128 * ksteps * 32, b_bytes = 256 * ksteps * 32;
    for (int i = threadIdx.x; i < (a_bytes + b_bytes) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = (i * 2654435761u) & 0x03030303u;
    fence_proxy_async();
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    if (threadIdx.x < 32) tmem_alloc(&tbase, 512);
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t td = tbase;
    if (threadIdx.x < 32) {
        const uint32_t idesc = make_idesc_i8(128, 256, 0, 1, 0, 0);
        const uint32_t a0 = smem_u32(sm), b0 = a0 + a_bytes;
        // K-major: 16-byte K chunks are LBO apart, 8-row groups SBO = 128 bytes apart; a K step of 32 bytes = 2 chunks
        const uint32_t a_lbo = 128 * 16, b_lbo = 256 * 16;
        uint32_t phase = 0;
        for (int r = 0; r < reps; ++r) {
            if (elect_one()) {
                for (int ks = 0; ks < ksteps; ++ks) {
                    const uint64_t ad = make_smem_desc(a0 + ks * 2 * a_lbo, a_lbo, 128);
                    const uint64_t bd = make_smem_desc(b0 + ks * 2 * b_lbo, b_lbo, 128);
                    umma_i8(td + (r & 1) * 256, ad, bd, idesc, ks ? 1u : 0u);
                }
            }
            __syncwarp();
            if ((r & 63) == 63) {                                  // bound the queue: wait for everything issued so far
                if (elect_one()) umma_commit(&bar);
                __syncwarp();
                mbar_wait(&bar, phase); phase ^= 1;
                tc_fence_after();
            }
        }
        if (elect_one()) umma_commit(&bar);
        __syncwarp();
        mbar_wait(&bar, phase);
    }
    tc_fence_before(); __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(td, 512);
}

int main() {
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int ksteps = 8;
    const size_t smem = (128 + 256) * ksteps * 32 + 1024;
    CK(cudaFuncSetAttribute(i8_loop, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    auto run = [&](int reps, int launches) {
        CK(cudaEventRecord(e0));
        for (int i = 0; i < launches; ++i) i8_loop<<<sms, 128, smem>>>(reps, ksteps);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double ops = 2.0 * 128 * 256 * 32 * (double)ksteps * reps * launches * sms;
        return ops / (ms * 1e-3) / 1e12;
    };
    run(2000, 2);                                   // warm-up
    const double burst = run(40000, 1);             // ~50 ms
    const double sustained = run(40000, 60);        // ~3 s under the power cap
    printf("{\"kind\": \"tcgen05 kind::i8, M=128 N=256 K=32, no-swizzle smem operands, one CTA per SM\", \"sms\": %d, "
           "\"i8_tops_dense_burst\": %.1f, \"i8_tops_dense_sustained\": %.1f}\n", sms, burst, sustained);
    return 0;
}
