// Development probe: cycles per tcgen05.mma kind::i8 for the operand layouts / shapes of the conv0 kernels.
#include <cstdio>
#include <cstdlib>
#include "../pn_tc.cuh"
using namespace pntc;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

// cfg: M, N, a_off (bytes added to the A start), a_lbo, a_sbo, b_lbo, b_sbo, a_mn, b_mn, reps
// cfg[11]: number of extra warps that hammer TMEM loads (tcgen05.ld) while the MMAs run; cfg[12]: extra warps doing LDS
__global__ void __launch_bounds__(416) timing(const int* cfg, long long* out, const uint8_t* gsrc) {
    extern __shared__ __align__(1024) uint8_t sm[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase;
    __shared__ volatile int done;
    if (threadIdx.x == 0) done = 0;
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = i * 2654435761u;
    fence_proxy_async();
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    if (threadIdx.x < 32) tmem_alloc(&tbase, 512);
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t td = tbase;
    if (threadIdx.x == 0) {
        const int M = cfg[0], N = cfg[1], reps = cfg[9];
        const uint32_t idesc = make_idesc_i8(M, N, 0, 1, cfg[7], cfg[8]);
        const uint64_t ad = make_smem_desc(smem_u32(sm) + cfg[2], cfg[3], cfg[4]);
        const uint64_t bd = make_smem_desc(smem_u32(sm) + 96 * 1024, cfg[5], cfg[6]);
        const long long t0 = clock64();
        const int nacc = cfg[10];     // MMAs rotate over this many independent accumulators
        __shared__ uint64_t bar2;
        if (cfg[9] == 2001) {          // commit + fence every 8 MMAs, descriptors advanced per MMA (like the conv kernel)
            mbar_init(&bar2, 1);
            for (int i = 0; i < reps; i += 8) {
                tc_fence_after();
#pragma unroll
                for (int j = 0; j < 8; ++j) umma_i8(td + ((i >> 3) & 3) * 128, ad + (uint64_t)(j * 8), bd + (uint64_t)(j * 8), idesc, j ? 1u : 0u);
                umma_commit(&bar2);
            }
        } else if (cfg[9] == 2002) {   // as 2001 but additionally wait for every commit before the next group
            mbar_init(&bar2, 1);
            uint32_t ph = 0;
            for (int i = 0; i < reps; i += 8) {
#pragma unroll
                for (int j = 0; j < 8; ++j) umma_i8(td + ((i >> 3) & 3) * 128, ad + (uint64_t)(j * 8), bd + (uint64_t)(j * 8), idesc, j ? 1u : 0u);
                umma_commit(&bar2);
                mbar_wait(&bar2, ph); ph ^= 1;
                tc_fence_after();
            }
        } else if (cfg[9] >= 3000 && cfg[9] < 3004) {   // distinct operand data per MMA: A and/or B start address cycles through its region
            const uint64_t am = (cfg[9] & 1) ? 1 : 0, bm = (cfg[9] & 2) ? 1 : 0;
            for (int i = 0; i < 4000; i += 20) {
#pragma unroll
                for (int j = 0; j < 20; ++j)
                    umma_i8(td + (j & 3) * 128, ad + am * (uint64_t)((j * 4096) >> 4), bd + bm * (uint64_t)(((j % 5) * 256) >> 4), idesc, (i | j) >= 4);
            }
        } else
        if (nacc == 4) { for (int i = 0; i < reps; ++i) umma_i8(td + (i & 3) * 128, ad, bd, idesc, i >= 4); }
        else if (nacc == 2) { for (int i = 0; i < reps; ++i) umma_i8(td + (i & 1) * 128, ad, bd, idesc, i >= 2); }
        else { for (int i = 0; i < reps; ++i) umma_i8(td, ad, bd, idesc, i >= 1); }
        const long long t1 = clock64();
        umma_commit(&bar);
        mbar_wait(&bar, 0);
        const long long t2 = clock64();
        out[0] = t1 - t0; out[1] = t2 - t0; if (cfg[9] >= 3000) { out[0] = out[0] * cfg[9] / 4000; out[1] = out[1] * cfg[9] / 4000; }
        done = 1;
    } else if (threadIdx.x >= 128) {
        const int w = (threadIdx.x >> 5) - 4;            // 0..7
        unsigned acc = 0;
        if (w < cfg[11]) {
            while (!done) {
                uint32_t r0[8], r1[8], r2[8];
                const uint32_t ta = td + ((uint32_t)((w & 3) * 32) << 16) + 256;
                tmem_ld8(ta, r0); tmem_ld8(ta + 8, r1); tmem_ld8(ta + 16, r2);
                tmem_ld_wait();
                for (int i = 0; i < 8; ++i) acc += r0[i] ^ r1[i] ^ r2[i];
            }
        } else if (w < cfg[11] + cfg[12]) {
            while (!done) {
                const uint4 v = reinterpret_cast<const uint4*>(sm)[(threadIdx.x * 37 + acc) & 4095];
                acc += v.x ^ v.y ^ v.z ^ v.w;
            }
        }
        if (acc == 0x12345) out[2] = acc;
    }
    if (threadIdx.x == 384 && cfg[13]) {
        // stream 28 KB "frames" into the upper part of shared memory, like the conv kernels' TMA producer
        __shared__ uint64_t pbar;
        mbar_init(&pbar, 1); fence_mbar_init();
        uint32_t ph = 0; long long n = 0;
        while (!done) {
            mbar_expect_tx(&pbar, 28224);
            bulk_g2s(sm + 120 * 1024, gsrc + (n % 64) * 28224, 28224, &pbar);
            mbar_wait(&pbar, ph); ph ^= 1; ++n;
        }
        out[3] = n;
    }
    __syncthreads();
    tc_fence_before(); __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(td, 512);
}

int main() {
    int* cfg_d; long long* out_d;
    CK(cudaMalloc(&cfg_d, 64)); CK(cudaMalloc(&out_d, 32)); uint8_t* gsrc; CK(cudaMalloc(&gsrc, 64 * 28224)); CK(cudaMemset(gsrc, 1, 64 * 28224));
    CK(cudaFuncSetAttribute(timing, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    struct C { const char* name; int v[14]; };
    const int R = 7056;
    C cases[] = {
        {"M128 N80 same A, same B        ", {128, 80, 0, 128, 128, 128, 2048, 0, 0, 3000, 4, 0, 0, 0}},
        {"M128 N80 distinct A (80 KB)    ", {128, 80, 0, 128, 128, 128, 2048, 0, 0, 3001, 4, 0, 0, 0}},
        {"M128 N80 distinct B            ", {128, 80, 0, 128, 128, 128, 2048, 0, 0, 3002, 4, 0, 0, 0}},
        {"M128 N80 distinct A and B      ", {128, 80, 0, 128, 128, 128, 2048, 0, 0, 3003, 4, 0, 0, 0}},
        {"M128 N80 distinct A, LBO 7168  ", {128, 80, 0, 7168, 128, 128, 2048, 0, 0, 3001, 4, 0, 0, 0}},
    };
    for (auto& c : cases) {
        CK(cudaMemcpy(cfg_d, c.v, sizeof(c.v), cudaMemcpyHostToDevice));
        for (int grid : {1, 148}) {
            timing<<<grid, 416, 160 * 1024>>>(cfg_d, out_d, gsrc);
            CK(cudaDeviceSynchronize());
            long long o[2];
            CK(cudaMemcpy(o, out_d, 16, cudaMemcpyDeviceToHost));
            printf("%s grid %3d: issue %.1f cyc/mma, complete %.1f cyc/mma\n", c.name, grid, (double)o[0] / c.v[9], (double)o[1] / c.v[9]);
        }
    }
    return 0;
}
