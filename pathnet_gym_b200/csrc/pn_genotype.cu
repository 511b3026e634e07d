// Genotype initialisation, mutation and tournament as integer kernels driven by
// a counter-based RNG (Philox4x32-10), bit-exact against the oracle restatement
// and against the reference functions themselves (tests/golden/genotype_golden.npz
// was produced by /root/reference/pathnet.py:50-63,78-87 fed with this stream).
//
// Counter layout: key = seed; counter = (epoch, path_id, pos,
// domain<<24 | slot).  u = (x >> 11) * 2^-53 in fp64.  Products such as
// (u*L)*N are evaluated left to right in fp64 exactly like the Python
// expressions `int(np.random.rand()*L*N)`; there are no additions, so no FMA
// contraction can change them.
#include "pn_internal.cuh"

#include <cstdarg>
#include <cstdio>

namespace {

enum { DOMAIN_MUTATE = 0, DOMAIN_INIT = 1, DOMAIN_PERM = 2 };

// pathnet.mutation (pathnet.py:50-63): row-major scan of the LIVE row.
__device__ void mutate_rows(uint32_t* g, int L, int M, int N, uint64_t seed, uint32_t epoch, uint32_t pid) {
    for (int l = 0; l < L; ++l) {
        uint32_t row = g[l];
        for (int j = 0; j < M; ++j) {
            const uint32_t pos = (uint32_t)(l * M + j);
            const double u0 = pn_philox_uniform(seed, epoch, pid, pos, DOMAIN_MUTATE, 0);
            bool hit;
            if ((row >> j) & 1u) {
                hit = (long long)((u0 * (double)L) * (double)N) <= 1;          // int(rand()*L*N) <= 1
                if (hit) row &= ~(1u << j);
            } else {
                hit = (long long)(((u0 * (double)L) * (double)(M - N)) * (double)M) <= 1;   // int(rand()*L*(M-N)*M) <= 1
            }
            if (hit) {
                const double u1 = pn_philox_uniform(seed, epoch, pid, pos, DOMAIN_MUTATE, 1);
                const int c = (int)(u1 * (double)M);                            // np.random.randint(0, M)
                row |= 1u << c;
            }
        }
        g[l] = row;
    }
}

__global__ void mutate_kernel(uint32_t* masks, int n, int pid0, int L, int M, int N, uint64_t seed, uint32_t epoch) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t g[PN_MAX_LAYERS];
    for (int l = 0; l < L; ++l) g[l] = masks[(size_t)i * L + l];
    mutate_rows(g, L, M, N, seed, epoch, (uint32_t)(pid0 + i));
    for (int l = 0; l < L; ++l) masks[(size_t)i * L + l] = g[l];
}

// pathnet.get_geopath (pathnet.py:78-87): per layer, rejection-sample int(rand()*M) until N distinct.
__global__ void geopath_init_kernel(uint32_t* masks, int n, int pid0, int L, int M, int N, uint64_t seed,
                                    uint32_t epoch) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t pid = (uint32_t)(pid0 + i);
    for (int l = 0; l < L; ++l) {
        uint32_t m = 0;
        int cnt = 0;
        uint32_t slot = 0;
        while (cnt < N) {
            const double u = pn_philox_uniform(seed, epoch, pid, (uint32_t)l, DOMAIN_INIT, slot++);
            const int rv = (int)(u * (double)M);
            if (!((m >> rv) & 1u)) { m |= 1u << rv; ++cnt; }
        }
        masks[(size_t)i * L + l] = m;
    }
}

// One controller pass (doom_pathnet.py:225-270) over n_groups disjoint candidate groups.
// Single CTA; dynamic smem: keys[P] u64, perm[P] i32.
__global__ void __launch_bounds__(256)
tournament_kernel(uint32_t* genotypes, float* scores, const uint32_t* fixed, uint32_t* effective, int32_t* winners,
                  int P, int L, int M, int N, int Bc, int n_groups, uint64_t seed, uint32_t epoch) {
    extern __shared__ __align__(8) unsigned char tsm[];
    uint64_t* keys = reinterpret_cast<uint64_t*>(tsm);
    int* perm = reinterpret_cast<int*>(keys + P);
    for (int i = threadIdx.x; i < P; i += blockDim.x) keys[i] = pn_philox_u64(seed, epoch, (uint32_t)i, 0, DOMAIN_PERM, 0);
    __syncthreads();
    // rank of path i under (key, id) ordering = its position in the permutation
    for (int i = threadIdx.x; i < P; i += blockDim.x) {
        const uint64_t ki = keys[i];
        int rank = 0;
        for (int q = 0; q < P; ++q) {
            const uint64_t kq = keys[q];
            rank += (kq < ki) || (kq == ki && q < i);
        }
        perm[rank] = i;
    }
    __syncthreads();
    for (int grp = threadIdx.x; grp < n_groups; grp += blockDim.x) {
        int win = -1;
        if ((grp + 1) * Bc <= P) {
            bool ready = true;
            float best = 0.f;
            for (int q = 0; q < Bc; ++q) {
                const int m = perm[grp * Bc + q];
                const float sc = scores[m];
                if (sc == -1000.0f) ready = false;
                if (q == 0 || sc > best) { best = sc; win = m; }   // first arg-max (np.argmax)
            }
            if (!ready) win = -1;
        }
        winners[grp] = win;
        if (win < 0) continue;
        uint32_t wg[PN_MAX_LAYERS];
        for (int l = 0; l < L; ++l) wg[l] = genotypes[(size_t)win * L + l];
        for (int q = 0; q < Bc; ++q) {
            const int m = perm[grp * Bc + q];
            if (m != win) {
                uint32_t g[PN_MAX_LAYERS];
                for (int l = 0; l < L; ++l) g[l] = wg[l];                 // np.copy(geopath_set[winner_idx])
                mutate_rows(g, L, M, N, seed, epoch, (uint32_t)m);
                for (int l = 0; l < L; ++l) genotypes[(size_t)m * L + l] = g[l];
            }
            scores[m] = -1000.0f;                                         // winner too (doom_pathnet.py:267)
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < P * L; i += blockDim.x) effective[i] = genotypes[i] | fixed[i % L];
}

thread_local char g_err[512] = "";

int check_geno_args(int n, int L, int M, int N) {
    if (n < 0 || L < 1 || L > PN_MAX_LAYERS || M < 1 || M > PN_MAX_MODULES || N < 0 || N > M)
        return pn_fail_global(PN_ERR_INVALID, "bad genotype shape: n=%d L=%d M=%d N=%d", n, L, M, N);
    return PN_OK;
}

}  // namespace

int pn_fail_global(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

extern "C" PN_API const char* pn_last_global_error(void) { return g_err; }

extern "C" PN_API int pn_geopath_init(uint32_t* masks_dev, int n_paths, int path_id0, int L, int M, int N, uint64_t seed,
                               uint32_t epoch, void* stream) {
    if (int rc = check_geno_args(n_paths, L, M, N)) return rc;
    if (n_paths == 0) return PN_OK;
    geopath_init_kernel<<<(n_paths + 127) / 128, 128, 0, (cudaStream_t)stream>>>(masks_dev, n_paths, path_id0, L, M, N,
                                                                                 seed, epoch);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return pn_fail_global(PN_ERR_CUDA, "geopath_init_kernel: %s", cudaGetErrorString(e));
    return PN_OK;
}

extern "C" PN_API int pn_mutate(uint32_t* masks_dev, int n_paths, int path_id0, int L, int M, int N, uint64_t seed,
                         uint32_t epoch, void* stream) {
    if (int rc = check_geno_args(n_paths, L, M, N)) return rc;
    if (n_paths == 0) return PN_OK;
    mutate_kernel<<<(n_paths + 127) / 128, 128, 0, (cudaStream_t)stream>>>(masks_dev, n_paths, path_id0, L, M, N, seed,
                                                                           epoch);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return pn_fail_global(PN_ERR_CUDA, "mutate_kernel: %s", cudaGetErrorString(e));
    return PN_OK;
}

extern "C" PN_API int pn_tournament(uint32_t* genotypes_dev, float* scores_dev, const uint32_t* fixed_dev,
                             uint32_t* effective_dev, int32_t* winners_dev, int P_total, int L, int M, int N, int Bcand,
                             int n_groups, uint64_t seed, uint32_t epoch, void* stream) {
    if (int rc = check_geno_args(P_total, L, M, N)) return rc;
    if (P_total < 1 || P_total > 8192)
        return pn_fail_global(PN_ERR_INVALID, "pn_tournament supports 1..8192 paths, got %d", P_total);
    if (Bcand < 1 || n_groups < 0 || (long long)n_groups * Bcand > P_total)
        return pn_fail_global(PN_ERR_INVALID, "n_groups*Bcand = %d*%d exceeds the population %d", n_groups, Bcand,
                              P_total);
    const size_t sm = (size_t)P_total * (sizeof(uint64_t) + sizeof(int));
    if (sm > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(tournament_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if (e != cudaSuccess) return pn_fail_global(PN_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    tournament_kernel<<<1, 256, sm, (cudaStream_t)stream>>>(genotypes_dev, scores_dev, fixed_dev, effective_dev,
                                                            winners_dev, P_total, L, M, N, Bcand, n_groups, seed, epoch);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return pn_fail_global(PN_ERR_CUDA, "tournament_kernel: %s", cudaGetErrorString(e));
    return PN_OK;
}
