// Development probe (not part of the library): checks on real sm_100a hardware that the shared-memory
// descriptor / instruction-descriptor encodings in pn_tc.cuh mean what the kernels assume.
//   nvcc -gencode arch=compute_100a,code=sm_100a -I.. -o umma_probe umma_probe.cu && ./umma_probe
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../pn_tc.cuh"
using namespace pntc;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int REGION = 536 * 16;

// mode 0: K-major forward form (tile 1, 4 taps x 2 k-steps, N = 48)
// mode 1/2: MN-major A and B, K = 64 pixels, M = 128 (mode 1) or M = 64 (mode 2)
__global__ void __launch_bounds__(128) probe(const uint8_t* a_g, int a_bytes, const int8_t* b_g, int b_bytes, int mode,
                                             int* out /*[128][64]*/) {
    extern __shared__ __align__(1024) uint8_t sm[];
    uint8_t* As = sm;
    uint8_t* Bs = sm + ((a_bytes + 1023) / 1024) * 1024;
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase;
    for (int i = threadIdx.x; i < a_bytes; i += 128) As[i] = a_g[i];
    for (int i = threadIdx.x; i < b_bytes; i += 128) Bs[i] = (uint8_t)b_g[i];
    fence_proxy_async();
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    if (threadIdx.x < 32) tmem_alloc(&tbase, 64);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tbase;
    if (threadIdx.x == 0) {
        if (mode == 0) {
            const uint32_t idesc = make_idesc_i8(128, 48, 0, 1, 0, 0);
            int first = 1;
            for (int tap = 0; tap < 4; ++tap)
                for (int kk = 0; kk < 2; ++kk) {
                    const int off = (tap >> 1) * 21 + (tap & 1);
                    const uint64_t ad = make_smem_desc(smem_u32(As) + (2 * kk) * REGION + (128 + off) * 16, REGION, 128);
                    const uint64_t bd = make_smem_desc(smem_u32(Bs) + (tap * 4 + 2 * kk) * 128, 128, 2048);
                    umma_i8(td, ad, bd, idesc, first ? 0 : 1);
                    first = 0;
                }
        } else {
            const int M = mode == 1 ? 128 : 64;
            const uint32_t idesc = make_idesc_i8(M, 48, 0, 1, 1, 1);
            for (int ks = 0; ks < 2; ++ks) {
                const uint64_t ad = make_smem_desc(smem_u32(As) + ks * 512, 128, 1024);
                const uint64_t bd = make_smem_desc(smem_u32(Bs) + ks * 512, 128, 1024);
                umma_i8(td, ad, bd, idesc, ks);
            }
        }
        umma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int c = 0; c < 48; c += 8) {
        uint32_t r[8];
        tmem_ld8(td + ((uint32_t)(32 * w) << 16) + c, r);
        tmem_ld_wait();
        for (int i = 0; i < 8; ++i) out[(32 * w + lane) * 64 + c + i] = (int)r[i];
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(td, 64);
}

int main() {
    srand(1);
    int* out_d;
    CK(cudaMalloc(&out_d, 128 * 64 * 4));
    std::vector<int> out(128 * 64);
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    // ---------------- mode 0 ----------------
    {
        std::vector<uint8_t> S(4 * REGION);
        for (auto& v : S) v = rand() & 255;
        std::vector<int8_t> Bw(48 * 256), Bc(6 * 2048);
        for (auto& v : Bw) v = (int8_t)((rand() & 255) - 128);
        for (int n = 0; n < 48; ++n)
            for (int k = 0; k < 256; ++k) Bc[(n / 8) * 2048 + (k / 16) * 128 + (n % 8) * 16 + (k % 16)] = Bw[n * 256 + k];
        uint8_t* a_d; int8_t* b_d;
        CK(cudaMalloc(&a_d, S.size())); CK(cudaMalloc(&b_d, Bc.size()));
        CK(cudaMemcpy(a_d, S.data(), S.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(b_d, Bc.data(), Bc.size(), cudaMemcpyHostToDevice));
        CK(cudaMemset(out_d, 0xff, 128 * 64 * 4));
        probe<<<1, 128, 64 * 1024>>>(a_d, (int)S.size(), b_d, (int)Bc.size(), 0, out_d);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(out.data(), out_d, out.size() * 4, cudaMemcpyDeviceToHost));
        long long bad = 0, maxd = 0;
        for (int m = 0; m < 128; ++m)
            for (int n = 0; n < 48; ++n) {
                long long ref = 0;
                for (int tap = 0; tap < 4; ++tap) {
                    const int off = (tap >> 1) * 21 + (tap & 1);
                    for (int g = 0; g < 4; ++g)
                        for (int b = 0; b < 16; ++b)
                            ref += (long long)S[g * REGION + (128 + m + off) * 16 + b] * Bw[n * 256 + (tap * 4 + g) * 16 + b];
                }
                const long long d = llabs(ref - out[m * 64 + n]);
                if (d) ++bad;
                if (d > maxd) maxd = d;
            }
        printf("mode0 K-major i8 (fwd form): mismatches=%lld maxdiff=%lld  sample out[0][0]=%d\n", bad, maxd, out[0]);
    }
    // ---------------- mode 1 / 2 ----------------
    for (int mode = 1; mode <= 2; ++mode) {
        std::vector<uint8_t> A(128 * 64), Ac(8 * 1024);     // A[mn][k]
        std::vector<int8_t> B(48 * 64), Bc(3 * 1024);       // B[n][k]
        for (auto& v : A) v = rand() & 255;
        for (auto& v : B) v = (int8_t)((rand() & 255) - 128);
        for (int mn = 0; mn < 128; ++mn)
            for (int k = 0; k < 64; ++k) Ac[(mn / 16) * 1024 + (k / 8) * 128 + (k % 8) * 16 + (mn % 16)] = A[mn * 64 + k];
        for (int n = 0; n < 48; ++n)
            for (int k = 0; k < 64; ++k) Bc[(n / 16) * 1024 + (k / 8) * 128 + (k % 8) * 16 + (n % 16)] = B[n * 64 + k];
        uint8_t* a_d; int8_t* b_d;
        CK(cudaMalloc(&a_d, Ac.size())); CK(cudaMalloc(&b_d, Bc.size()));
        CK(cudaMemcpy(a_d, Ac.data(), Ac.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(b_d, Bc.data(), Bc.size(), cudaMemcpyHostToDevice));
        CK(cudaMemset(out_d, 0xff, 128 * 64 * 4));
        probe<<<1, 128, 64 * 1024>>>(a_d, (int)Ac.size(), b_d, (int)Bc.size(), mode, out_d);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(out.data(), out_d, out.size() * 4, cudaMemcpyDeviceToHost));
        const int M = mode == 1 ? 128 : 64;
        std::vector<long long> ref(M * 48);
        for (int mn = 0; mn < M; ++mn)
            for (int n = 0; n < 48; ++n) {
                long long r = 0;
                for (int k = 0; k < 64; ++k) r += (long long)A[mn * 64 + k] * B[n * 64 + k];
                ref[mn * 48 + n] = r;
            }
        if (mode == 1) {
            long long bad = 0;
            for (int mn = 0; mn < 128; ++mn)
                for (int n = 0; n < 48; ++n) bad += ref[mn * 48 + n] != out[mn * 64 + n];
            printf("mode1 MN-major i8 M=128: mismatches=%lld\n", bad);
        } else {
            // discover which TMEM lane holds each of the 64 rows
            printf("mode2 MN-major i8 M=64: row->lane map:");
            int found = 0;
            for (int mn = 0; mn < 64; ++mn) {
                int hit = -1;
                for (int lane = 0; lane < 128; ++lane) {
                    bool ok = true;
                    for (int n = 0; n < 48 && ok; ++n) ok = ref[mn * 48 + n] == out[lane * 64 + n];
                    if (ok) { hit = lane; break; }
                }
                if (hit >= 0) ++found;
                if (mn % 16 == 0) printf(" [%d->%d]", mn, hit);
            }
            printf("  found %d/64\n", found);
        }
    }
    printf("probe done\n");
    return 0;
}
