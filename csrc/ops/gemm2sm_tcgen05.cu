// Two-SM (cta_group::2) persistent GEMM: C[M,N] = A[M,K] * B[N,K]^T, bf16/fp16 in, fp32 accumulate.
//
// EXPERIMENTAL — compiled into the library and reachable through dpt_gemm2sm_bf16(), but not
// used by any model path yet and not yet validated on hardware (tests/test_gpu_gemm2sm.py is
// opt-in, DPT_TEST_2SM=1).  It is the proving ground for moving the large-shape conv / GEMM
// path from 1-CTA 128x128 tiles (L2-bandwidth bound at ~50 % of cuBLAS, profiles/README.md §4)
// to CTA pairs.
//
// Structure (one thread-block cluster of 2 CTAs = one SM pair works on a 256 x 256 output tile):
//   * tcgen05.mma.cta_group::2, M = 256, N = 256, K = 16 per instruction, issued by one thread
//     of the leader CTA (cluster rank 0) only.  Each CTA stages ITS half of A (128 rows) and ITS
//     half of B (128 of the 256 N rows) — 32 KB per k-step per CTA instead of 48 KB for the same
//     math with two independent 128 x 256 tiles — and holds its 128 accumulator rows in its own
//     TMEM (2 x 256 columns, double buffered).
//   * Both CTAs' producer threads issue their TMA loads with .cta_group::2 and the LEADER's
//     mbarrier as completion target; the leader arms that barrier with the bytes of both CTAs.
//   * tcgen05.commit ... .multicast::cluster releases the smem stage in both CTAs and publishes
//     the finished accumulator to both epilogues; the epilogue warps of both CTAs hand the
//     accumulator back by arriving (remotely for the peer) on the leader's tmem_empty barrier.
//   * Persistent: 74 clusters walk the tile list; the epilogue of tile i overlaps the mainloop
//     of tile i+1.
// PTX forms follow the vendored CUTLASS headers (cute/arch/copy_sm100_tma.hpp,
// cutlass/arch/barrier.h, cute/arch/tmem_allocator_sm100.hpp, cute/arch/mma_sm100_umma.hpp).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>

#include "../common/ptx.cuh"
#include "../../include/dptb200.h"

namespace dpt2sm {
using namespace dpt;

constexpr int kBM = 128;          // rows of A (and of the accumulator) per CTA
constexpr int kBNHalf = 128;      // rows of B staged per CTA
constexpr int kBN = 256;          // tile N of the pair
constexpr int kBK = 64;           // k elements per stage (one 128-byte swizzle row of 16-bit data)
constexpr int kStages = 6;
constexpr int kAStage = kBM * kBK * 2;        // 16 KB
constexpr int kBStage = kBNHalf * kBK * 2;    // 16 KB
constexpr int kStage = kAStage + kBStage;     // 32 KB per CTA per k-step
constexpr int kThreads = 192;                 // warp 0 TMA, warp 1 MMA + TMEM, warps 2..5 epilogue
constexpr uint32_t kTmemCols = 512;           // 2 accumulators x 256 fp32 columns
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // clears the CTA-pair bit of a shared::cluster address

struct Params {
    void* C;
    long long ldc;
    int M, N, K;
    int tiles_m, tiles_n;
    int out_f32;
    int f16;   // operands are fp16 instead of bf16
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_result, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(smem_result)),
                 "r"(ncols)
                 : "memory");
}
__device__ __forceinline__ void tmem_relinquish_2sm() {
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
                 : "memory");
}
// TMA tile load issued by either CTA of the pair; the transaction bytes are counted on the
// barrier at the same shared-memory offset in the LEADER CTA (peer bit cleared).
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const void* tmap, uint64_t* bar,
                                                int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void umma_f16_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                             uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Arrive (once the MMAs issued so far have retired) on the barrier at this offset in every CTA of
// `cta_mask` (bit r = cluster rank r).
__device__ __forceinline__ void umma_commit_2sm_mc(uint64_t* bar, uint16_t cta_mask) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 "
        "[%0], %1;" ::"r"(smem_u32(bar)),
        "h"(cta_mask)
        : "memory");
}
// Arrive on the barrier at the same offset in cluster rank `rank`.
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t rank) {
    asm volatile(
        "{\n\t"
        ".reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(rank)
        : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm2sm_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
               const __grid_constant__ Params prm) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~uintptr_t(1023));
    __shared__ __align__(8) uint64_t full_bar[kStages];      // used in the leader only
    __shared__ __align__(8) uint64_t empty_bar[kStages];     // per CTA (multicast commit)
    __shared__ __align__(8) uint64_t tmem_full_bar[2];       // per CTA (multicast commit)
    __shared__ __align__(8) uint64_t tmem_empty_bar[2];      // used in the leader only
    __shared__ uint32_t tmem_base_smem;

    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int pair = static_cast<int>(blockIdx.x >> 1);
    const int npairs = static_cast<int>(gridDim.x >> 1);
    const int ntiles = prm.tiles_m * prm.tiles_n;
    const int nk = (prm.K + kBK - 1) / kBK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_a);
        tma_prefetch_desc(&tm_b);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);    // the leader producer's arrive.expect_tx
            mbar_init(&empty_bar[s], 1);   // one multicast commit per use
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&tmem_full_bar[a], 1);
            mbar_init(&tmem_empty_bar[a], 8);   // 4 epilogue warps x 2 CTAs
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        // joint allocation: the same warp of both CTAs issues it, result at the same smem offset
        tmem_alloc_2sm(&tmem_base_smem, kTmemCols);
        tmem_relinquish_2sm();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();   // barriers initialised and TMEM allocated in BOTH CTAs before any
                          // remote arrive / multicast commit / pair MMA
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    if (warp == 0) {
        // ================================================================ TMA producer (both CTAs)
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = pair; tile < ntiles; tile += npairs) {
                const int tm = tile % prm.tiles_m, tn = tile / prm.tiles_m;
                const int row_a = tm * 2 * kBM + static_cast<int>(rank) * kBM;
                const int row_b = tn * kBN + static_cast<int>(rank) * kBNHalf;
                for (int k = 0; k < nk; ++k) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * kStage;
                    uint8_t* sb = sa + kAStage;
                    if (leader) mbar_expect_tx(&full_bar[stage], 2 * kStage);
                    tma_load_2d_2sm(sa, &tm_a, &full_bar[stage], k * kBK, row_a);
                    tma_load_2d_2sm(sb, &tm_b, &full_bar[stage], k * kBK, row_b);
                    if (++stage == kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================================================================ MMA issuer (leader only)
        if (leader && lane == 0) {
            constexpr uint32_t kIdescBf16 = make_idesc(2 * kBM, kBN, 0, 0, 1);
            constexpr uint32_t kIdescF16 = make_idesc(2 * kBM, kBN, 0, 0, 0);
            const uint32_t idesc = prm.f16 ? kIdescF16 : kIdescBf16;
            int stage = 0;
            uint32_t phase = 0;
            uint32_t nacc = 0;
            for (int tile = pair; tile < ntiles; tile += npairs) {
                const uint32_t acc = nacc & 1u;
                mbar_wait(&tmem_empty_bar[acc], ((nacc >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t tmem_acc = tmem_base + acc * kBN;
                for (int k = 0; k < nk; ++k) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * kStage);
                    const uint32_t sb = sa + kAStage;
#pragma unroll
                    for (int kk = 0; kk < kBK / 16; ++kk) {
                        const uint64_t da = make_smem_desc_sw128(sa + kk * 32, 16, 1024);
                        const uint64_t db = make_smem_desc_sw128(sb + kk * 32, 16, 1024);
                        umma_f16_2sm(tmem_acc, da, db, idesc, (k | kk) != 0 ? 1u : 0u);
                    }
                    umma_commit_2sm_mc(&empty_bar[stage], 3);   // frees the stage in both CTAs
                    if (++stage == kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit_2sm_mc(&tmem_full_bar[acc], 3);     // accumulator ready in both CTAs
                ++nacc;
            }
        }
    } else {
        // ================================================================ epilogue (both CTAs)
        const uint32_t sub = warp & 3;
        uint32_t nacc = 0;
        const bool vec32 = prm.out_f32 ? (prm.ldc % 8 == 0) : (prm.ldc % 16 == 0);
        for (int tile = pair; tile < ntiles; tile += npairs) {
            const int tm = tile % prm.tiles_m, tn = tile / prm.tiles_m;
            const uint32_t acc = nacc & 1u;
            mbar_wait(&tmem_full_bar[acc], (nacc >> 1) & 1u);
            tc_fence_after();
            const int row = tm * 2 * kBM + static_cast<int>(rank) * kBM + static_cast<int>(sub * 32 + lane);
            const bool row_ok = row < prm.M;
            const long long row_off = static_cast<long long>(row) * prm.ldc;
#pragma unroll 1
            for (int c = 0; c < kBN / 16; ++c) {
                uint32_t v[16];
                tmem_ld16(tmem_base + ((sub * 32u) << 16) + acc * kBN + c * 16, v);
                tmem_ld_wait();
                const int col = tn * kBN + c * 16;
                if (col >= prm.N) continue;   // warp-uniform; the loads above stay collective
                if (!row_ok) continue;
                const bool full = col + 16 <= prm.N;
                if (prm.out_f32) {
                    float* o = reinterpret_cast<float*>(prm.C) + row_off + col;
                    if (full && vec32 && (reinterpret_cast<uintptr_t>(o) % 32) == 0) {
                        uint32_t lo[8], hi[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            lo[i] = v[i];
                            hi[i] = v[8 + i];
                        }
                        st_global_v8(o, lo);
                        st_global_v8(o + 8, hi);
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (col + i < prm.N) o[i] = __uint_as_float(v[i]);
                    }
                } else {
                    uint16_t* o = reinterpret_cast<uint16_t*>(prm.C) + row_off + col;
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float a = __uint_as_float(v[2 * i]), b = __uint_as_float(v[2 * i + 1]);
                        if (prm.f16) {
                            __half2 h = __floats2half2_rn(a, b);
                            pk[i] = *reinterpret_cast<uint32_t*>(&h);
                        } else {
                            __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
                            pk[i] = *reinterpret_cast<uint32_t*>(&h);
                        }
                    }
                    if (full && vec32 && (reinterpret_cast<uintptr_t>(o) % 32) == 0) {
                        st_global_v8(o, pk);
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (col + i < prm.N)
                                o[i] = static_cast<uint16_t>((i & 1) ? (pk[i >> 1] >> 16)
                                                                     : (pk[i >> 1] & 0xFFFFu));
                    }
                }
            }
            // all tcgen05.ld of this warp have completed: give the accumulator back to the
            // leader's MMA warp (a remote arrive for the peer CTA)
            tc_fence_before();
            if (lane == 0) mbar_arrive_cluster(&tmem_empty_bar[acc], 0);
            ++nacc;
        }
    }

    // Nobody may leave while the peer can still multicast-commit into this CTA's barriers, read
    // its operand stages or write its TMEM.
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) tmem_dealloc_2sm(tmem_base, kTmemCols);
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
    static PFN_encodeTiled fn = nullptr;
    if (fn == nullptr) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
                cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(p);
    }
    return fn;
}

// [rows][ld] 16-bit row-major matrix, box = 64 (k) x 128 (rows), SWIZZLE_128B, OOB reads as zero.
static int make_map(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld,
                    bool f16) {
    PFN_encodeTiled enc = get_encode();
    if (enc == nullptr) return DPT_ERR_DRIVER;
    const cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    const cuuint64_t gstr[1] = {static_cast<cuuint64_t>(ld) * 2};
    const cuuint32_t box[2] = {kBK, kBM};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                     const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : DPT_ERR_TMAP;
}

}  // namespace dpt2sm

// C[M,N] (bf16/fp16 or fp32, row-major, ldc) = A[M,K] * B[N,K]^T with A, B 16-bit row-major
// (lda, ldb in elements, multiples of 8).  `f16` selects fp16 operands / output instead of bf16.
extern "C" int dpt_gemm2sm_bf16(const void* A, const void* B, void* C, int M, int N, int K,
                                long long lda, long long ldb, long long ldc, int out_f32, int f16,
                                void* stream) {
    using namespace dpt2sm;
    if (M <= 0 || N <= 0 || K <= 0 || lda % 8 != 0 || ldb % 8 != 0) return DPT_ERR_SHAPE;
    CUtensorMap ta, tb;
    int e = make_map(&ta, A, M, K, lda, f16 != 0);
    if (e) return e;
    e = make_map(&tb, B, N, K, ldb, f16 != 0);
    if (e) return e;
    Params prm;
    memset(&prm, 0, sizeof(prm));
    prm.C = C;
    prm.ldc = ldc;
    prm.M = M;
    prm.N = N;
    prm.K = K;
    prm.tiles_m = (M + 2 * kBM - 1) / (2 * kBM);
    prm.tiles_n = (N + kBN - 1) / kBN;
    prm.out_f32 = out_f32;
    prm.f16 = f16;
    const int smem = kStages * kStage + 1024;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t ce = cudaFuncSetAttribute(gemm2sm_kernel,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (ce != cudaSuccess) return static_cast<int>(ce);
        attr_done = true;
    }
    int pairs = prm.tiles_m * prm.tiles_n;
    if (pairs > 74) pairs = 74;   // one cluster per SM pair
    gemm2sm_kernel<<<dim3(2 * pairs), dim3(kThreads), smem, static_cast<cudaStream_t>(stream)>>>(
        ta, tb, prm);
    return static_cast<int>(cudaGetLastError());
}
