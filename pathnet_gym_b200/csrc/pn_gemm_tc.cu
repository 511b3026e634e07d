// Plain fp32-in / fp32-out GEMMs on the tensor cores (tcgen05, bf16 hi/lo operands, fp32 accumulation in TMEM) for the LSTM
// variant's four large products (pn_lstm.cu): the input projection of all steps, the gradient at the cell input and the two
// halves of the weight gradient -- 4.3 GFLOP each at 256 paths x 32 steps, which fp32 FFMA kernels did in ~190 us each.
//
// Precision: every fp32 operand x is split into bf16 hi = rn(x), lo = rn(x - hi); a product is hi*hi + hi*lo + lo*hi (three
// MMAs into one accumulator), ~2^-16 relative per product like the backward GEMMs of the other layers.
//
// Operand storage ("planes"): a row-major fp32 matrix S[R][C] becomes two arrays (hi, lo) of 8 x 8 bf16 blocks of 128 bytes,
// block (rb, cb) = rows 8 rb .., columns 8 cb .., stored [plane][rb][cb][8 r][8 c].  A block is a tcgen05 core matrix in both
// descriptor forms -- 8 rows of 16 bytes -- so one split serves every use of the matrix:
//   the matrix's contiguous dimension is the GEMM's K   -> K-major operand  (rows = M or N index)
//   the matrix's contiguous dimension is M (or N)       -> MN-major operand (rows = K index)
// A CTA computes one 128 x 128 tile of C; the operands of a 64-deep K chunk arrive by ONE 3-D TMA load per operand (box =
// [plane 2][row blocks][column blocks x 128 B]), three stages in flight; warp 4 loads, warp 5 issues the MMAs, warps 0-3 read
// the accumulator, stage the tile in shared memory and write it out in whole 512-byte rows.  blockIdx.z splits K; the
// partial products are added in split order by the caller (fixed summation order).
// TMA loads here read only this GPU's memory.
#include "pn_internal.cuh"
#include "pn_tc.cuh"

#include <cuda.h>
#include <cuda_bf16.h>

using namespace pntc;

namespace {

constexpr int kBlkB = 128;               // bytes of one 8 x 8 bf16 block
constexpr int kTile = 128;               // C tile (rows and columns)
constexpr int kKC = 64;                  // K per stage
constexpr int kStagesG = 3;
constexpr int kOperandStage = 2 * (kTile / 8) * (kKC / 8) * kBlkB;      // hi + lo of one operand: 32 KB
constexpr int kStageBytesG = 2 * kOperandStage;

// ---- fp32 row-major [R][C] (leading dimension ld) -> bf16 hi / lo block planes; rows >= R of the last row block are zero ----
// one thread per 16-byte row of a block; 8 consecutive threads fill one 128-byte block
__global__ void __launch_bounds__(256)
plane_split_kernel(const float* __restrict__ S, int R, int C, long long ld, uint8_t* __restrict__ planes) {
    const int ncb = C / 8, nrb = (R + 7) / 8;
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= (long long)nrb * ncb * 8) return;
    const int r8 = (int)(i & 7);
    const long long blk = i >> 3;
    const int cb = (int)(blk % ncb), rb = (int)(blk / ncb);
    const int r = rb * 8 + r8;
    float v[8];
    if (r < R) {
        const float4 a = *reinterpret_cast<const float4*>(S + (long long)r * ld + cb * 8);
        const float4 b = *reinterpret_cast<const float4*>(S + (long long)r * ld + cb * 8 + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = 0.f;
    }
    __nv_bfloat16 hi[8], lo[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        hi[q] = __float2bfloat16_rn(v[q]);
        lo[q] = __float2bfloat16_rn(v[q] - __bfloat162float(hi[q]));
    }
    const size_t off = (size_t)blk * kBlkB + r8 * 16, plane = (size_t)nrb * ncb * kBlkB;
    *reinterpret_cast<uint4*>(planes + off) = *reinterpret_cast<uint4*>(hi);
    *reinterpret_cast<uint4*>(planes + plane + off) = *reinterpret_cast<uint4*>(lo);
}

typedef CUresult (*PFN_encodeTiledG)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                     CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 3-D view of the planes: x = 8-byte words along a row of blocks (16 per block), y = row block, z = plane
int make_plane_tmap(pn_handle* h, const uint8_t* planes, int nrb, int ncb, int box_cb, int box_rb, CUtensorMap* out) {
    static PFN_encodeTiledG encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
            return pn_fail(h, PN_ERR_CUDA, "cuTensorMapEncodeTiled is not available: %s", cudaGetErrorString(e));
        encode = reinterpret_cast<PFN_encodeTiledG>(fn);
    }
    const cuuint64_t dims[3] = {(cuuint64_t)ncb * 16, (cuuint64_t)nrb, 2};
    const cuuint64_t strides[2] = {(cuuint64_t)ncb * kBlkB, (cuuint64_t)nrb * ncb * kBlkB};
    const cuuint32_t box[3] = {(cuuint32_t)box_cb * 16, (cuuint32_t)box_rb, 2};
    const cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<uint8_t*>(planes), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return pn_fail(h, PN_ERR_CUDA, "cuTensorMapEncodeTiled (planes) failed with CUresult %d", (int)r);
    return PN_OK;
}

__device__ __forceinline__ void tma_load_3d(void* dst_smem, const void* tmap, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::
            "r"(smem_u32(dst_smem)),
        "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
        : "memory");
}

struct GemmTcArgs {
    float* C; long long ldc;                  // C(m, n) = C[m * ldc + n]; K split z writes C + z * M * ldc
    const float* bias;                        // optional, per column (only without a K split)
    int M, N, k_chunks;                       // k_chunks = 64-deep K chunks per CTA (per split)
};

template <bool A_MN, bool B_MN>
__global__ void __launch_bounds__(192, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB, GemmTcArgs g) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)kStagesG * kStageBytesG);
    uint64_t* full = bars;                    // [kStagesG]
    uint64_t* empty = bars + kStagesG;        // [kStagesG]
    uint64_t* acc_full = bars + 2 * kStagesG;
    __shared__ uint32_t tmem_base_s;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int nt = blockIdx.x, mt = blockIdx.y, z = blockIdx.z;
    const int chunk0 = z * g.k_chunks;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kStagesG; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        mbar_init(acc_full, 1);
        fence_mbar_init();
    }
    if (warp == 5) tmem_alloc(&tmem_base_s, 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t td = tmem_base_s;

    if (warp == 4) {
        // ================= producer: one 3-D TMA load per operand and stage =================
        if (lane == 0) { tma_prefetch_desc(&tmapA); tma_prefetch_desc(&tmapB); }
        for (int c = 0; c < g.k_chunks; ++c) {
            const int st = c % kStagesG;
            mbar_wait(&empty[st], ((c / kStagesG) & 1) ^ 1);
            if (elect_one()) {
                uint8_t* dst = smem + (size_t)st * kStageBytesG;
                const int kc = chunk0 + c;
                mbar_expect_tx(&full[st], (uint32_t)kStageBytesG);
                // K-major: x = column blocks of the chunk (k), y = the tile's row blocks;  MN-major: x = the tile's column
                // blocks (m or n), y = the chunk's row blocks (k).  16 words of 8 bytes per block.
                if (A_MN) tma_load_3d(dst, &tmapA, mt * (kTile / 8) * 16, kc * (kKC / 8), 0, &full[st]);
                else tma_load_3d(dst, &tmapA, kc * (kKC / 8) * 16, mt * (kTile / 8), 0, &full[st]);
                if (B_MN) tma_load_3d(dst + kOperandStage, &tmapB, nt * (kTile / 8) * 16, kc * (kKC / 8), 0, &full[st]);
                else tma_load_3d(dst + kOperandStage, &tmapB, kc * (kKC / 8) * 16, nt * (kTile / 8), 0, &full[st]);
            }
            __syncwarp();
        }
    } else if (warp == 5) {
        // ================= MMA issue =================
        const uint32_t idesc = make_idesc_f16(kTile, kTile, 1, 1, A_MN ? 1 : 0, B_MN ? 1 : 0);     // bf16 x bf16
        // descriptor of K step 0 relative to the operand's stage base; a K step (16) advances by two K groups
        //   K-major  [row block][k block]: K groups 128 B apart, row groups (kKC / 8) * 128 B apart
        //   MN-major [k block][col block]: K groups (kTile / 8) * 128 B apart, column groups 128 B apart
        constexpr uint32_t kLboA = A_MN ? (kTile / 8) * kBlkB : kBlkB, kSboA = A_MN ? kBlkB : (kKC / 8) * kBlkB;
        constexpr uint32_t kLboB = B_MN ? (kTile / 8) * kBlkB : kBlkB, kSboB = B_MN ? kBlkB : (kKC / 8) * kBlkB;
        constexpr uint64_t kStepA = (2 * kLboA) >> 4, kStepB = (2 * kLboB) >> 4, kLoOff = (kOperandStage / 2) >> 4;
        for (int c = 0; c < g.k_chunks; ++c) {
            const int st = c % kStagesG;
            mbar_wait(&full[st], (c / kStagesG) & 1);
            tc_fence_after();
            const uint32_t base = smem_u32(smem + (size_t)st * kStageBytesG);
            const uint64_t a_hi = make_smem_desc(base, kLboA, kSboA), b_hi = make_smem_desc(base + kOperandStage, kLboB, kSboB);
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < kKC / 16; ++ks) {
                    umma_f16(td, a_hi + ks * kStepA, b_hi + ks * kStepB, idesc, (c | ks) ? 1u : 0u);
                    umma_f16(td, a_hi + ks * kStepA, b_hi + kLoOff + ks * kStepB, idesc, 1u);
                    umma_f16(td, a_hi + kLoOff + ks * kStepA, b_hi + ks * kStepB, idesc, 1u);
                }
                umma_commit(&empty[st]);
            }
            __syncwarp();
        }
        if (elect_one()) umma_commit(acc_full);
        __syncwarp();
    } else {
        // ================= epilogue: thread = row of the tile; tile staged in shared memory, written out in whole rows =====
        const int quad = warp & 3, row = quad * 32 + lane;
        constexpr int kLdS = kTile + 4;                                  // staging row stride (floats): conflict-free 16-byte accesses
        float* Cs = reinterpret_cast<float*>(smem);                      // the operand stages are free once acc_full has completed
        mbar_wait(acc_full, 0);
        tc_fence_after();
        const uint32_t ta = td + ((uint32_t)(quad * 32) << 16);
#pragma unroll
        for (int c0 = 0; c0 < kTile; c0 += 32) {
            uint32_t d[32];
            tmem_ld16(ta + c0, *reinterpret_cast<uint32_t(*)[16]>(&d[0]));
            tmem_ld16(ta + c0 + 16, *reinterpret_cast<uint32_t(*)[16]>(&d[16]));
            tmem_ld_wait();
#pragma unroll
            for (int q = 0; q < 32; q += 4)
                *reinterpret_cast<uint4*>(Cs + (size_t)row * kLdS + c0 + q) = make_uint4(d[q], d[q + 1], d[q + 2], d[q + 3]);
        }
        tc_fence_before();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        float* C = g.C + (size_t)z * g.M * g.ldc;
        const int n = nt * kTile + lane * 4;
        float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (g.bias && n < g.N) bv = *reinterpret_cast<const float4*>(g.bias + n);
        for (int r = quad; r < kTile; r += 4) {
            const int m = mt * kTile + r;
            if (m >= g.M || n >= g.N) continue;
            float4 v = *reinterpret_cast<const float4*>(Cs + (size_t)r * kLdS + lane * 4);
            v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
            *reinterpret_cast<float4*>(C + (size_t)m * g.ldc + n) = v;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 5) tmem_dealloc(td, 128);
}
}  // namespace

size_t pn_gemm_planes_bytes(int R, int C) { return (size_t)2 * ((R + 7) / 8) * (C / 8) * kBlkB; }

// S[R][C] fp32 (leading dimension ld floats, 16-byte aligned rows) -> planes (pn_gemm_planes_bytes(R, C) bytes)
int pn_gemm_split(pn_handle* h, const float* S, int R, int C, long long ld, uint8_t* planes, cudaStream_t st) {
    if (C % 8 || ld % 4) return pn_fail(h, PN_ERR_INVALID, "pn_gemm_split: C = %d, ld = %lld", C, ld);
    const long long n = (long long)((R + 7) / 8) * (C / 8) * 8;
    plane_split_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>(S, R, C, ld, planes);
    PN_LAUNCH_CHECK(h, "plane_split_kernel");
    return PN_OK;
}

bool pn_gemm_tc_usable(int M, int N, int K, int nsplit) {
    return M > 0 && N > 0 && N % 4 == 0 && K > 0 && nsplit >= 1 && K % (kKC * nsplit) == 0;
}

// C[M x N] (+ bias) = A * B on the tensor cores.  An operand is the planes of a row-major matrix S[R][C]:
//   mn == false: GEMM row (m or n) = S row, k = S column (K-major);   mn == true: k = S row, GEMM row = S column (MN-major).
// nsplit > 1: split z writes its partial product to C + z * M * ldc (the caller adds them up); bias must be null then.
int pn_gemm_tc(pn_handle* h, const PnGemmOperand& A, const PnGemmOperand& B, float* C, long long ldc, const float* bias, int M,
               int N, int K, int nsplit, cudaStream_t st) {
    if (!pn_gemm_tc_usable(M, N, K, nsplit) || (nsplit > 1 && bias))
        return pn_fail(h, PN_ERR_INVALID, "pn_gemm_tc: M = %d, N = %d, K = %d, %d splits", M, N, K, nsplit);
    if ((A.mn ? A.R : A.C) != K || (B.mn ? B.R : B.C) != K || (A.mn ? A.C : A.R) < M || (B.mn ? B.C : B.R) < N)
        return pn_fail(h, PN_ERR_INVALID, "pn_gemm_tc: operand shapes do not match M = %d, N = %d, K = %d", M, N, K);
    CUtensorMap ta, tb;
    if (int rc = make_plane_tmap(h, A.planes, (A.R + 7) / 8, A.C / 8, A.mn ? kTile / 8 : kKC / 8, A.mn ? kKC / 8 : kTile / 8, &ta)) return rc;
    if (int rc = make_plane_tmap(h, B.planes, (B.R + 7) / 8, B.C / 8, B.mn ? kTile / 8 : kKC / 8, B.mn ? kKC / 8 : kTile / 8, &tb)) return rc;
    GemmTcArgs g{C, ldc, bias, M, N, K / kKC / nsplit};
    const size_t smem = (size_t)kStagesG * kStageBytesG + 64;
    dim3 grid((N + kTile - 1) / kTile, (M + kTile - 1) / kTile, nsplit);
#define PN_GEMM_TC_LAUNCH(AM, BM_)                                                   \
    do {                                                                             \
        PN_SET_SMEM(h, (gemm_tc_kernel<AM, BM_>), smem);                             \
        gemm_tc_kernel<AM, BM_><<<grid, 192, smem, st>>>(ta, tb, g);                 \
    } while (0)
    if (A.mn && B.mn) PN_GEMM_TC_LAUNCH(true, true);
    else if (A.mn) PN_GEMM_TC_LAUNCH(true, false);
    else if (B.mn) PN_GEMM_TC_LAUNCH(false, true);
    else PN_GEMM_TC_LAUNCH(false, false);
#undef PN_GEMM_TC_LAUNCH
    PN_LAUNCH_CHECK(h, "gemm_tc_kernel");
    return PN_OK;
}
