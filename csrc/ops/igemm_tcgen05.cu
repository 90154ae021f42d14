// Implicit-GEMM convolution + plain GEMM on the 5th-gen tensor cores (sm_100a).
//
// One warp-specialised kernel template covers the three GEMM-shaped passes of a
// convolution (and the equivalent plain GEMMs of a linear layer):
//
//   MODE_FPROP  Y[m, co]  = sum_{tap, ci} X[m @ tap, ci] * W[co, tap, ci]
//               A = activations, K-major, fetched by 4-D TMA boxes straight out of the NHWC
//               tensor (one box per filter tap; TMA zero-fills the padding halo);
//               B = weights [Cout][R*S*Cin], K-major, 2-D TMA.
//   MODE_DGRAD  dX[m, ci] = sum_{tap, co} dY[m @ tap, co] * W[co, tap, ci]
//               A = dY, K-major 4-D TMA; B = the *same* weight tensor read as an N-major
//               operand (no transposed weight copy exists anywhere).
//               Stride-2 convs are decomposed into 4 output-parity classes (blockIdx.z) so
//               no multiply-by-zero work is done; a second conv (the 1x1 downsample that
//               shares the block input) can be accumulated in the same TMEM tile.
//   MODE_WGRAD  dW[co, tap, ci] = sum_m dY[m, co] * X[m @ tap, ci]
//               both operands MN-major (the reduction runs over pixels), split over pixel
//               tiles across CTAs, reduced with vector fp32 RED straight into the gradient
//               arena slot (no partial buffers, no copy-in for the reducer).
//
// Pipeline: warp 0 = TMA producer, warp 1 = tcgen05.mma issuer (one elected lane) and TMEM
// owner, warps 2..5 = epilogue (tcgen05.ld -> registers -> global), ring of NSTAGES smem
// stages guarded by full/empty mbarriers; accumulator tile 128 x BLOCK_N fp32 in TMEM.
//
// Replaces, on the reference's path, cuDNN convolution fwd/bwd and the cuBLAS fc GEMMs that
// torchvision's resnet18 dispatches to (reference train_ddp.py:153-156, 205-213).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>

#include "../common/ptx.cuh"
#include "../../include/dptb200.h"

namespace dpt {

constexpr int MODE_FPROP = 0;
constexpr int MODE_DGRAD = 1;
constexpr int MODE_WGRAD = 2;

constexpr int OUT_BF16 = 0;
constexpr int OUT_F32 = 1;
constexpr int OUT_F32_ATOMIC = 2;

constexpr int kTileM = 128;
constexpr int kKBlk = 64;
constexpr int kMaxTaps = 52;
constexpr int kNumThreads = 192;         // TMA warp + MMA warp + 4 epilogue warps
constexpr int kNumThreadsPersist = 320;  // TMA warp + MMA warp + 8 epilogue warps
constexpr int kAStageBytes = kTileM * kKBlk * 2;  // 16 KiB in either major-ness

struct IgemmTap {
    int8_t a_map, b_map;
    int16_t dw, dh;
    int16_t pad_;
    int32_t koff;  // column offset of this tap inside the weight matrix
};

struct IgemmMaps {
    CUtensorMap a[4];
    CUtensorMap b[4];
};

struct IgemmParams {
    int bq, bp, bn;                 // rows of a tile enumerate (n, p, q) inside this box
    int tiles_q, tiles_p, tiles_n;  // tile grid over the row space
    int Q, P, NB;                   // row-space extents (for masking)
    int chunks;                     // 64-wide k chunks per tap (fprop/dgrad)
    int splits;                     // k splits (fprop/dgrad) or pixel-tile splits (wgrad)
    int nclasses;                   // 1, or 4 for the stride-2 dgrad parity decomposition
    int cls_begin[5];               // tap range per class
    int n_total;                    // valid output columns
    int m_total;                    // wgrad: valid output rows (Cout)
    int ntile_per_tap;              // wgrad: N tiles per tap
    // output addressing
    void* out;
    int out_mode;
    long long ldo;
    int OH, OW, osh, osw;
    float* stat_sum;
    float* stat_sq;
    const float* bias;
    int pgx, pgy, ptiles;           // persistent kernels: logical tile grid (x fastest) and its size
    IgemmTap taps[kMaxTaps];
};

// What one output tile works on: its position in the row / column space and its k range.
struct TileCoord {
    int cls;                 // dgrad parity class
    int n0, p0, q0;          // row-space origin (fprop / dgrad)
    int col0;                // first output column
    int tap_w, row0_w;       // wgrad: filter tap and first output row (cout)
    int k_begin, k_end;      // k-loop range of this tile (after splitting)
};

template <int MODE, int BLOCK_N>
__device__ __forceinline__ TileCoord tile_coord(const IgemmParams& prm, int bx, int by, int bz) {
    TileCoord tc;
    tc.cls = 0;
    tc.n0 = tc.p0 = tc.q0 = 0;
    tc.tap_w = tc.row0_w = 0;
    int split = bz;
    int total;
    if (MODE != MODE_WGRAD) {
        tc.cls = bz / prm.splits;
        split = bz % prm.splits;
        int t = bx;
        tc.q0 = (t % prm.tiles_q) * prm.bq;
        t /= prm.tiles_q;
        tc.p0 = (t % prm.tiles_p) * prm.bp;
        t /= prm.tiles_p;
        tc.n0 = t * prm.bn;
        tc.col0 = by * BLOCK_N;
        const int ntap = prm.cls_begin[tc.cls + 1] - prm.cls_begin[tc.cls];
        total = ntap * prm.chunks;
    } else {
        tc.tap_w = bx / prm.ntile_per_tap;
        tc.col0 = (bx % prm.ntile_per_tap) * BLOCK_N;
        tc.row0_w = by * kTileM;
        total = prm.tiles_q * prm.tiles_p * prm.tiles_n;
    }
    const int per = (total + prm.splits - 1) / prm.splits;
    tc.k_begin = split * per;
    tc.k_end = min(total, tc.k_begin + per);
    return tc;
}

template <int W>
__device__ __forceinline__ void xpose_reduce_step(float (&v)[16], uint32_t lane, uint32_t bit) {
    const bool upper = (lane & bit) != 0;
#pragma unroll
    for (int i = 0; i < W / 2; ++i) {
        float send = upper ? v[i] : v[i + W / 2];
        float keep = upper ? v[i + W / 2] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
    }
}
// Sum 16 per-lane values over the 32 lanes; afterwards v[0] on lane l holds the total of
// column ((l>>4)&1)*8 + ((l>>3)&1)*4 + ((l>>2)&1)*2 + ((l>>1)&1) (duplicated on l and l^1).
__device__ __forceinline__ float warp_colsum16(float (&v)[16], uint32_t lane) {
    xpose_reduce_step<16>(v, lane, 16);
    xpose_reduce_step<8>(v, lane, 8);
    xpose_reduce_step<4>(v, lane, 4);
    xpose_reduce_step<2>(v, lane, 2);
    return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
}

// DT: element type of both operands: 0 = bf16, 1 = fp16 (kind::f16), 2 = fp32 read as TF32
// (kind::tf32; the reference's FP32 mode also runs its convolutions in TF32 through cuDNN).
// PERSIST: the CTA walks a list of output tiles (tile = blockIdx.x, += gridDim.x) instead of
// owning one.  The accumulator is double buffered in TMEM (2 x BLOCK_N columns) and handed back
// and forth through tmem_full / tmem_empty mbarriers, so the epilogue of tile i (TMEM -> registers
// -> global, BN statistics) overlaps the loads and MMAs of tile i+1, the TMA producer runs up to
// NSTAGES k-steps ahead across tile boundaries, and barrier initialisation / TMEM allocation /
// descriptor prefetch are paid once per CTA instead of once per tile.
// The persistent variant also doubles the epilogue: 8 warps (two per TMEM lane quarter, each
// taking half of the tile's columns).  With short k loops the kernel is bound by the issue rate
// of the epilogue warps (ncu on the 1x1 64->256 conv of ResNet-50 layer1: tensor pipe 4 % busy,
// 12 resident warps, issue slots 42 % busy), so more of them is what buys throughput.
template <int MODE, int BLOCK_N, int NSTAGES, int DT, bool PERSIST = false>
__global__ void __launch_bounds__(PERSIST ? kNumThreadsPersist : kNumThreads, 1)
igemm_kernel(const __grid_constant__ IgemmMaps maps, const __grid_constant__ IgemmParams prm) {
    constexpr bool A_MN = (MODE == MODE_WGRAD);
    constexpr bool B_MN = (MODE != MODE_FPROP);
    constexpr bool F16 = (DT == 1);
    constexpr int ESZ = (DT == 2) ? 4 : 2;
    constexpr int kKB = 128 / ESZ;             // k elements (K-major) or k rows (MN-major) per stage
    constexpr int kMnBlk = 128 / ESZ;          // MN elements in one 128-byte swizzle block
    constexpr int kMnBlkBytes = kKB * 128;     // one MN block = kKB rows x 128 B
    constexpr int kUmmaK = 32 / ESZ;           // 16 (16-bit) or 8 (tf32)
    constexpr int kBStageBytes = BLOCK_N * 128;
    constexpr int kStageBytes = kAStageBytes + kBStageBytes;
    constexpr uint32_t kAccCols = BLOCK_N < 32 ? 32 : BLOCK_N;
    constexpr uint32_t kTmemCols = PERSIST ? 2 * kAccCols : kAccCols;
    // operand format field: 1 = BF16, 0 = F16 (bits 7-9 for A, 10-12 for B)
    constexpr uint32_t kIdesc =
        make_idesc(kTileM, BLOCK_N, A_MN ? 1 : 0, B_MN ? 1 : 0, DT == 0 ? 1 : (DT == 1 ? 0 : 2));

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // SWIZZLE_128B needs 1024-byte aligned stage bases.
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~uintptr_t(1023));
    __shared__ __align__(8) uint64_t full_bar[NSTAGES];
    __shared__ __align__(8) uint64_t empty_bar[NSTAGES];
    __shared__ __align__(8) uint64_t tmem_full_bar[2];
    __shared__ __align__(8) uint64_t tmem_empty_bar[2];
    __shared__ uint32_t tmem_base_smem;
    __shared__ __align__(16) float s_stat[2][4][BLOCK_N];  // per-epilogue-warp column partials

    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31;

    // ---------------------------------------------------------------- tile list of this CTA
    const int tile_first = PERSIST ? static_cast<int>(blockIdx.x) : 0;
    const int tile_step = PERSIST ? static_cast<int>(gridDim.x) : 1;
    const int tile_end = PERSIST ? prm.ptiles : 1;
    auto coord_of = [&](int tile) {
        if (PERSIST) {
            const int bx = tile % prm.pgx;
            const int r = tile / prm.pgx;
            return tile_coord<MODE, BLOCK_N>(prm, bx, r % prm.pgy, r / prm.pgy);
        }
        return tile_coord<MODE, BLOCK_N>(prm, blockIdx.x, blockIdx.y, blockIdx.z);
    };

    // ---------------------------------------------------------------- one-time setup
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&maps.a[0]);
        tma_prefetch_desc(&maps.b[0]);
        for (int s = 0; s < NSTAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&tmem_full_bar[a], 1);
            mbar_init(&tmem_empty_bar[a], PERSIST ? 8 : 4);   // one arrival per epilogue warp
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(&tmem_base_smem, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;
    // everything above (barrier init, TMEM allocation, descriptor prefetch) overlapped the tail
    // of the previous kernel; from here on we read what it produced
    pdl_launch_dependents();
    pdl_wait();

    if (warp == 0) {
        // ============================================================ TMA producer
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = tile_first; tile < tile_end; tile += tile_step) {
            const TileCoord tc = coord_of(tile);
            const int cls = tc.cls, n0 = tc.n0, p0 = tc.p0, q0 = tc.q0, col0 = tc.col0;
            const int tap_w = tc.tap_w, row0_w = tc.row0_w;
            for (int k = tc.k_begin; k < tc.k_end; ++k) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                uint8_t* sa = smem + stage * kStageBytes;
                uint8_t* sb = sa + kAStageBytes;
                mbar_expect_tx(&full_bar[stage], kStageBytes);
                if (MODE == MODE_FPROP) {
                    const int ti = prm.cls_begin[cls] + k / prm.chunks;
                    const int j = k % prm.chunks;
                    const IgemmTap tp = prm.taps[ti];
                    tma_load_4d(sa, &maps.a[tp.a_map], &full_bar[stage], j * kKB, q0 + tp.dw,
                                p0 + tp.dh, n0);
                    tma_load_2d(sb, &maps.b[tp.b_map], &full_bar[stage], tp.koff + j * kKB, col0);
                } else if (MODE == MODE_DGRAD) {
                    const int ti = prm.cls_begin[cls] + k / prm.chunks;
                    const int j = k % prm.chunks;
                    const IgemmTap tp = prm.taps[ti];
                    tma_load_4d(sa, &maps.a[tp.a_map], &full_bar[stage], j * kKB, q0 + tp.dw,
                                p0 + tp.dh, n0);
#pragma unroll
                    for (int nb = 0; nb < BLOCK_N / kMnBlk; ++nb)
                        tma_load_2d(sb + nb * kMnBlkBytes, &maps.b[tp.b_map], &full_bar[stage],
                                    tp.koff + col0 + nb * kMnBlk, j * kKB);
                } else {
                    int t = k;
                    const int wq = (t % prm.tiles_q) * prm.bq;
                    t /= prm.tiles_q;
                    const int wp = (t % prm.tiles_p) * prm.bp;
                    t /= prm.tiles_p;
                    const int wn = t * prm.bn;
                    const IgemmTap tp = prm.taps[tap_w];
#pragma unroll
                    for (int mb = 0; mb < kTileM / kMnBlk; ++mb)
                        tma_load_4d(sa + mb * kMnBlkBytes, &maps.a[0], &full_bar[stage],
                                    row0_w + mb * kMnBlk, wq, wp, wn);
#pragma unroll
                    for (int nb = 0; nb < BLOCK_N / kMnBlk; ++nb)
                        tma_load_4d(sb + nb * kMnBlkBytes, &maps.b[tp.b_map], &full_bar[stage],
                                    col0 + nb * kMnBlk, wq + tp.dw, wp + tp.dh, wn);
                }
                if (++stage == NSTAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            }
        }
    } else if (warp == 1) {
        // ============================================================ MMA issuer
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            uint32_t nacc = 0;   // accumulator tiles produced so far (tiles with an empty k range
                                 // use no accumulator; the epilogue counts the same way)
            for (int tile = tile_first; tile < tile_end; tile += tile_step) {
            const TileCoord tc = coord_of(tile);
            const int nk = tc.k_end - tc.k_begin;
            if (nk <= 0) continue;
            const uint32_t acc = PERSIST ? (nacc & 1u) : 0u;
            if (PERSIST) {
                // the epilogue must have drained this accumulator (its previous use)
                mbar_wait(&tmem_empty_bar[acc], ((nacc >> 1) & 1u) ^ 1u);
                tc_fence_after();
            }
            const uint32_t tmem_acc = tmem_base + acc * kAccCols;
            for (int k = 0; k < nk; ++k) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + stage * kStageBytes);
                const uint32_t sb = sa + kAStageBytes;
#pragma unroll
                for (int kk = 0; kk < kKB / kUmmaK; ++kk) {
                    // K-major: one instruction's K (32 bytes) further inside the 128 B swizzle row.
                    // MN-major: kUmmaK k-rows of 128 B further; LBO = next MN block.
                    // 32-bit MN-major operands use the 32-byte-base swizzle (4-row K atoms)
                    constexpr uint32_t kMnSbo = (DT == 2) ? 512 : 1024;
                    constexpr uint32_t kMnLayout = (DT == 2) ? 1 : 2;
                    const uint64_t da =
                        A_MN ? make_smem_desc(sa + kk * (kUmmaK * 128), kMnBlkBytes, kMnSbo, kMnLayout)
                             : make_smem_desc_sw128(sa + kk * 32, 16, 1024);
                    const uint64_t db =
                        B_MN ? make_smem_desc(sb + kk * (kUmmaK * 128), kMnBlkBytes, kMnSbo, kMnLayout)
                             : make_smem_desc_sw128(sb + kk * 32, 16, 1024);
                    if (DT == 2)
                        umma_tf32(tmem_acc, da, db, kIdesc, (k | kk) != 0 ? 1u : 0u);
                    else
                        umma_f16(tmem_acc, da, db, kIdesc, (k | kk) != 0 ? 1u : 0u);
                }
                umma_commit(&empty_bar[stage]);  // frees the smem slot when the MMAs retire
                if (++stage == NSTAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            umma_commit(&tmem_full_bar[acc]);
            ++nacc;
            }
        }
    } else {
        // ============================================================ epilogue (4 or 8 warps)
        const uint32_t sub = warp & 3;  // TMEM sub-partition this warp may read
        const int r = sub * 32 + lane;  // accumulator row
        // with 8 epilogue warps, warps 2..5 take the first half of the columns, 6..9 the second
        constexpr int kColHalves = PERSIST ? 2 : 1;
        constexpr int kChunksPerWarp = BLOCK_N / 16 / kColHalves;
        constexpr int kEpiThreads = PERSIST ? 256 : 128;
        const int half = PERSIST ? static_cast<int>((warp - 2) >> 2) : 0;
        uint32_t nacc = 0;
        for (int tile = tile_first; tile < tile_end; tile += tile_step) {
        const TileCoord tc = coord_of(tile);
        const int cls = tc.cls, n0 = tc.n0, p0 = tc.p0, q0 = tc.q0, col0 = tc.col0;
        const int tap_w = tc.tap_w, row0_w = tc.row0_w;
        const int nk = max(0, tc.k_end - tc.k_begin);
        const uint32_t acc = PERSIST ? (nacc & 1u) : 0u;
        const uint32_t tmem_acc = tmem_base + acc * kAccCols;
        if (nk > 0) {
            mbar_wait(&tmem_full_bar[acc], PERSIST ? ((nacc >> 1) & 1u) : 0u);
            tc_fence_after();
        }
        bool row_ok;
        long long row_off;
        int koff_out = 0;
        if (MODE != MODE_WGRAD) {
            const int q = q0 + r % prm.bq;
            const int p = p0 + (r / prm.bq) % prm.bp;
            const int n = n0 + r / (prm.bq * prm.bp);
            row_ok = (q < prm.Q) && (p < prm.P) && (n < prm.NB);
            const int oh0 = (prm.nclasses > 1) ? (cls >> 1) : 0;
            const int ow0 = (prm.nclasses > 1) ? (cls & 1) : 0;
            row_off = ((static_cast<long long>(n) * prm.OH + (p * prm.osh + oh0)) * prm.OW +
                       (q * prm.osw + ow0)) *
                      prm.ldo;
        } else {
            const int row = row0_w + r;
            row_ok = row < prm.m_total;
            row_off = static_cast<long long>(row) * prm.ldo;
            koff_out = prm.taps[tap_w].koff;
        }
        const bool vec_ok = (prm.ldo % 8) == 0 && (koff_out % 8) == 0;
        // 32-byte stores need every 16-column chunk of every row on a 32-byte boundary
        const bool vec32_ok = (prm.ldo % 16) == 0 && (koff_out % 16) == 0 &&
                              (reinterpret_cast<uintptr_t>(prm.out) % 32) == 0;
#pragma unroll 1
        for (int c = half * kChunksPerWarp; c < (half + 1) * kChunksPerWarp; ++c) {
            float f[16];
            if (nk > 0) {
                uint32_t v[16];
                tmem_ld16(tmem_acc + ((sub * 32u) << 16) + c * 16, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) f[i] = 0.f;
            }
            const int cbase = col0 + c * 16;
            if (cbase >= prm.n_total) break;  // warp-uniform
            if (prm.bias != nullptr) {
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    if (cbase + i < prm.n_total) f[i] += __ldg(prm.bias + cbase + i);
            }
            const bool full = (cbase + 16 <= prm.n_total);
            if (row_ok) {
                const long long off = row_off + koff_out + cbase;
                if (prm.out_mode == OUT_BF16) {
                    uint16_t* o = reinterpret_cast<uint16_t*>(prm.out) + off;
                    if (full && vec_ok) {
                        uint32_t pk[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            if (F16) {
                                __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
                                pk[i] = *reinterpret_cast<uint32_t*>(&h);
                            } else {
                                __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
                                pk[i] = *reinterpret_cast<uint32_t*>(&h);
                            }
                        }
                        if (vec32_ok) {
                            st_global_v8(o, pk);   // one full 32-byte sector
                        } else {
                            uint4* o4 = reinterpret_cast<uint4*>(o);
                            o4[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                            o4[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (cbase + i < prm.n_total) {
                                if (F16) {
                                    const __half h = __float2half_rn(f[i]);
                                    o[i] = *reinterpret_cast<const uint16_t*>(&h);
                                } else {
                                    const __nv_bfloat16 h = __float2bfloat16_rn(f[i]);
                                    o[i] = *reinterpret_cast<const uint16_t*>(&h);
                                }
                            }
                    }
                } else if (prm.out_mode == OUT_F32) {
                    float* o = reinterpret_cast<float*>(prm.out) + off;
                    if (full && vec_ok) {
                        if (vec32_ok) {
                            uint32_t lo[8], hi[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                lo[i] = __float_as_uint(f[i]);
                                hi[i] = __float_as_uint(f[8 + i]);
                            }
                            st_global_v8(o, lo);
                            st_global_v8(o + 8, hi);
                        } else {
                            float4* o4 = reinterpret_cast<float4*>(o);
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                o4[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (cbase + i < prm.n_total) o[i] = f[i];
                    }
                } else {
                    float* o = reinterpret_cast<float*>(prm.out) + off;
                    if (full && vec_ok) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            red_add_f32x4(o + 4 * i, f[4 * i], f[4 * i + 1], f[4 * i + 2],
                                          f[4 * i + 3]);
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (cbase + i < prm.n_total) atomicAdd(o + i, f[i]);
                    }
                }
            }
            if (MODE == MODE_FPROP && prm.stat_sum != nullptr) {
                // Per-channel sum / sum-of-squares of this tile for the following BatchNorm
                // (statistics come from the fp32 accumulators, before bf16 rounding).
                float s1[16], s2[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float x = row_ok ? f[i] : 0.f;
                    s1[i] = x;
                    s2[i] = x * x;
                }
                const float t1 = warp_colsum16(s1, lane);
                const float t2 = warp_colsum16(s2, lane);
                const int colsel = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 +
                                   ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
                if ((lane & 1) == 0) {
                    s_stat[0][sub][c * 16 + colsel] = t1;
                    s_stat[1][sub][c * 16 + colsel] = t2;
                }
            }
        }
        if (PERSIST && nk > 0) {
            // every tcgen05.ld of this warp has completed (tmem_ld_wait): hand the accumulator
            // back to the MMA warp so the next-but-one tile can start while we finish up
            tc_fence_before();
            if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
        }
        if (nk > 0) ++nacc;
        if (MODE == MODE_FPROP && prm.stat_sum != nullptr) {
            // combine the 4 epilogue warps in shared memory, then 16-byte vector RED: the
            // 2*Cout statistics live on a few cache lines and L2 serialises atomics per line
            asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
            if (sub == 0) {
                constexpr int kColsPerWarp = BLOCK_N / kColHalves;
                for (int j = half * kColsPerWarp + lane * 4; j < (half + 1) * kColsPerWarp; j += 128) {
                    const int cg = col0 + j;
                    if (cg >= prm.n_total) break;
                    float a[4], b[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        a[e] = s_stat[0][0][j + e] + s_stat[0][1][j + e] + s_stat[0][2][j + e] +
                               s_stat[0][3][j + e];
                        b[e] = s_stat[1][0][j + e] + s_stat[1][1][j + e] + s_stat[1][2][j + e] +
                               s_stat[1][3][j + e];
                    }
                    if (cg + 4 <= prm.n_total) {
                        red_add_f32x4(prm.stat_sum + cg, a[0], a[1], a[2], a[3]);
                        red_add_f32x4(prm.stat_sq + cg, b[0], b[1], b[2], b[3]);
                    } else {
                        for (int e = 0; e < 4 && cg + e < prm.n_total; ++e) {
                            atomicAdd(prm.stat_sum + cg + e, a[e]);
                            atomicAdd(prm.stat_sq + cg + e, b[e]);
                        }
                    }
                }
            }
            // the partials are overwritten by the next tile: wait until warp `sub == 0` has
            // read them
            if (PERSIST) asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
        }
        }
    }

    // Reductions (split-K accumulators, BN statistics) are fire-and-forget: make sure they have
    // been performed before this CTA counts as complete for a programmatically dependent kernel.
    if (warp >= 2 && (prm.out_mode == OUT_F32_ATOMIC || prm.stat_sum != nullptr)) __threadfence();
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

// ================================================================================ host side
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

// Element type of the operands (0 = bf16, 1 = fp16, 2 = fp32/TF32), set through
// dpt_set_lowp_dtype().
static int g_lowp_f16 = 0;
static inline int esz() { return g_lowp_f16 == 2 ? 4 : 2; }
static inline int kblk() { return 128 / esz(); }      // elements per 128-byte swizzle row
static inline int lowp_out_mode() { return g_lowp_f16 == 2 ? OUT_F32 : OUT_BF16; }

// 16-bit tensor map, up to 4-D, inner box = 64 elements (128 B) with SWIZZLE_128B.
// dims/strides are in elements; strides[i] is the stride of dimension i+1.
static int make_tmap(CUtensorMap* m, const void* base, int rank, const long long* dims,
                     const long long* strides, const int* box, bool mn_major = false) {
    PFN_encodeTiled enc = get_encode();
    if (enc == nullptr) return DPT_ERR_DRIVER;
    cuuint64_t gdim[4];
    cuuint64_t gstr[3];
    cuuint32_t bdim[4];
    cuuint32_t estr[4] = {1, 1, 1, 1};
    for (int i = 0; i < rank; ++i) {
        gdim[i] = static_cast<cuuint64_t>(dims[i] > 0 ? dims[i] : 1);
        bdim[i] = static_cast<cuuint32_t>(box[i]);
    }
    for (int i = 0; i + 1 < rank; ++i) gstr[i] = static_cast<cuuint64_t>(strides[i]) * esz();
    const CUtensorMapDataType tdt = g_lowp_f16 == 2   ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                    : g_lowp_f16 == 1 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                                      : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    CUresult r = enc(m, tdt, rank, const_cast<void*>(base), gdim, gstr,
                     bdim, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     (mn_major && g_lowp_f16 == 2) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                                   : CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        fprintf(stderr,
                "[dptb200] cuTensorMapEncodeTiled failed (%d): rank %d base %p dims %lld %lld %lld "
                "%lld strides %lld %lld %lld box %d %d %d %d\n",
                int(r), rank, base, dims[0], rank > 1 ? dims[1] : 0, rank > 2 ? dims[2] : 0,
                rank > 3 ? dims[3] : 0, strides[0], rank > 2 ? strides[1] : 0,
                rank > 3 ? strides[2] : 0, box[0], rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0,
                rank > 3 ? box[3] : 0);
        return DPT_ERR_TMAP;
    }
    return 0;
}

static int pow2_ceil(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

// Choose (bq, bp, bn) with bq*bp*bn == rows covering a (Q, P, N) row space.
static void pick_box(int rows, int Q, int P, int N, int* bq, int* bp, int* bn) {
    int q = pow2_ceil(Q);
    if (q > rows) q = rows;
    int p = pow2_ceil(P);
    if (p > rows / q) p = rows / q;
    int n = rows / (q * p);
    (void)N;
    *bq = q;
    *bp = p;
    *bn = n;
}

// persistent-variant policy knobs (dpt_igemm_set_persist): minimum number of tiles from which the
// persistent kernels are used, and the CTA count they are launched with
static long long g_persist_min_tiles = 4 * 148;
static int g_persist_grid = 2 * 148;

// Opt-in 128 x 256 tiles (DPT_IGEMM_BN256=1 / dpt_igemm_set_bn256): per k-step a CTA stages
// 16 KB of A and 32 KB of B for twice the math of a 128 x 128 tile, i.e. 25 % less operand traffic
// per FLOP — the 1-CTA kernel is L2-bandwidth bound on long k loops (profiles/README.md).  Not
// yet measured on hardware, hence off by default.
static int g_bn256 = -1;
static bool bn256_enabled() {
    if (g_bn256 < 0) {
        const char* e = std::getenv("DPT_IGEMM_BN256");
        g_bn256 = (e != nullptr && e[0] == '1') ? 1 : 0;
    }
    return g_bn256 != 0;
}
// Output-column tile width: 64 below 128 columns, else 128, or 256 when enabled, the column count
// is a multiple of 256 and `other_tiles` x (cols / 256) still fills two waves of the 148 SMs.
static int pick_block_n(int cols, long long other_tiles) {
    if (bn256_enabled() && esz() == 2 && cols % 256 == 0 &&
        other_tiles * (cols / 256) >= 2 * 148)
        return 256;
    return cols >= 128 ? 128 : 64;
}

template <int MODE, int BLOCK_N, int NST, bool PERSIST>
static int launch_cfg(const IgemmMaps& maps, const IgemmParams& prm_in, dim3 grid, cudaStream_t st) {
    constexpr int kBStageBytes = BLOCK_N * 128;
    const int smem = NST * (kAStageBytes + kBStageBytes) + 1024;
    auto kern = g_lowp_f16 == 2   ? igemm_kernel<MODE, BLOCK_N, NST, 2, PERSIST>
                : g_lowp_f16 == 1 ? igemm_kernel<MODE, BLOCK_N, NST, 1, PERSIST>
                                  : igemm_kernel<MODE, BLOCK_N, NST, 0, PERSIST>;
    static bool attr_done[3] = {false, false, false};
    if (!attr_done[g_lowp_f16]) {
        cudaError_t e =
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return int(e);
        attr_done[g_lowp_f16] = true;
    }
    {
        static int igemm_pdl = -1;
        if (igemm_pdl < 0) {
            const char* e = std::getenv("DPT_PDL_IGEMM");
            igemm_pdl = (e == nullptr || e[0] != '0') ? 1 : 0;
        }
        if (!igemm_pdl) pdl_skip_flag() = 1;
    }
    IgemmParams prm = prm_in;
    if (PERSIST) {
        // logical tile grid in the order the hardware would have scheduled it (x fastest);
        // two CTAs per SM (2 x 2 x BLOCK_N TMEM columns, 2 x ~97 KB smem)
        prm.pgx = static_cast<int>(grid.x);
        prm.pgy = static_cast<int>(grid.y);
        prm.ptiles = static_cast<int>(grid.x * grid.y * grid.z);
        // 256-wide tiles take all 512 TMEM columns when double buffered: one CTA per SM
        const int resident = (BLOCK_N == 256 && g_persist_grid > 148) ? 148 : g_persist_grid;
        grid = dim3(prm.ptiles < resident ? prm.ptiles : resident, 1, 1);
    }
    cudaError_t le = launch_pdl(kern, grid, dim3(PERSIST ? kNumThreadsPersist : kNumThreads),
                                static_cast<size_t>(smem), st, maps, prm);
    if (le != cudaSuccess) return int(le);
    return int(cudaGetLastError());
}

// Launch policy.
//   * Grids of up to a few waves (every ResNet-18 / CIFAR shape): one tile per CTA, one CTA per
//     SM with a deep ring (6 x 32 KB or 8 x 24 KB) — the CTA has to cover the TMA latency on its
//     own and there is no second tile to overlap with.
//   * Grids of >= 4 waves (ResNet-50 @224: 6272 row tiles in layer1; large plain GEMMs): a
//     128 x BLOCK_N tile with a short k loop is latency bound end to end (load -> MMA -> TMEM
//     read -> store, nothing overlapping inside a one-tile CTA), so the persistent variant runs:
//     2 CTAs per SM, each looping over tiles with a double-buffered TMEM accumulator.
// DPT_IGEMM_PERSIST=0 forces the one-tile-per-CTA configuration.
template <int MODE>
static int launch(const IgemmMaps& maps, const IgemmParams& prm, dim3 grid, int block_n,
                  cudaStream_t st) {
    static int persist = -1;
    if (persist < 0) {
        const char* e = std::getenv("DPT_IGEMM_PERSIST");
        persist = (e == nullptr || e[0] != '0') ? 1 : 0;
    }
    const long long ctas = static_cast<long long>(grid.x) * grid.y * grid.z;
    const bool many = persist && ctas >= g_persist_min_tiles && ctas < (1LL << 31);
    if (block_n == 256) {   // 4 x 48 KB stages, one CTA per SM either way
        if (many) return launch_cfg<MODE, 256, 4, true>(maps, prm, grid, st);
        return launch_cfg<MODE, 256, 4, false>(maps, prm, grid, st);
    }
    if (block_n == 128) {
        if (many) return launch_cfg<MODE, 128, 3, true>(maps, prm, grid, st);
        return launch_cfg<MODE, 128, 6, false>(maps, prm, grid, st);
    }
    if (many) return launch_cfg<MODE, 64, 4, true>(maps, prm, grid, st);
    return launch_cfg<MODE, 64, 8, false>(maps, prm, grid, st);
}

// A-side activation maps for a conv whose taps read X[n, p*stride + r - pad, q*stride + s - pad].
// For stride 2 one strided "parity view" map per (h parity, w parity) is built so that every
// tap is a unit-stride box in its view.  Returns taps (map index + box offsets) for all (r, s)
// whose receptive row/column range intersects the tensor at all.
struct TapBuilder {
    IgemmTap taps[kMaxTaps];
    int ntaps = 0;
};

static int build_act_maps(CUtensorMap* maps4, const void* x, int N, int H, int W, int C, int stride,
                          int rows, int bq, int bp, int bn, bool mn_major = false) {
    (void)rows;
    const int box[4] = {kblk(), bq, bp, bn};
    if (stride == 1) {
        const long long dims[4] = {C, W, H, N};
        const long long str[3] = {C, (long long)W * C, (long long)H * W * C};
        int e = make_tmap(&maps4[0], x, 4, dims, str, box, mn_major);
        if (e) return e;
        for (int i = 1; i < 4; ++i) maps4[i] = maps4[0];
        return 0;
    }
    for (int ph = 0; ph < 2; ++ph)
        for (int pw = 0; pw < 2; ++pw) {
            const char* base =
                reinterpret_cast<const char*>(x) + ((long long)ph * W + pw) * C * esz();
            const long long dims[4] = {C, (W - pw + 1) / 2, (H - ph + 1) / 2, N};
            const long long str[3] = {2LL * C, 2LL * W * C, (long long)H * W * C};
            if (dims[1] <= 0 || dims[2] <= 0) {
                maps4[ph * 2 + pw] = maps4[0];
                continue;
            }
            int e = make_tmap(&maps4[ph * 2 + pw], base, 4, dims, str, box, mn_major);
            if (e) return e;
        }
    return 0;
}

// floor division / modulo for possibly negative numerators
static inline int fdiv(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }
static inline int fmod_pos(int a, int b) { return a - fdiv(a, b) * b; }

static void fprop_taps(TapBuilder* tb, int H, int W, int P, int Q, int Cin, int R, int S, int stride,
                       int pad, int b_map) {
    for (int r = 0; r < R; ++r)
        for (int s = 0; s < S; ++s) {
            // skip taps that only ever see padding
            const int hmin = r - pad, hmax = (P - 1) * stride + r - pad;
            const int wmin = s - pad, wmax = (Q - 1) * stride + s - pad;
            if (hmax < 0 || hmin >= H || wmax < 0 || wmin >= W) continue;
            IgemmTap t{};
            const int dh = r - pad, dw = s - pad;
            if (stride == 1) {
                t.a_map = 0;
                t.dh = dh;
                t.dw = dw;
            } else {
                const int ph = fmod_pos(dh, stride), pw = fmod_pos(dw, stride);
                t.a_map = ph * 2 + pw;
                t.dh = fdiv(dh, stride);
                t.dw = fdiv(dw, stride);
            }
            t.b_map = b_map;
            t.koff = (r * S + s) * Cin;
            tb->taps[tb->ntaps++] = t;
        }
}

}  // namespace dpt

using namespace dpt;

// called by dpt_set_lowp_dtype (norm_pool_loss.cu) so that every translation unit agrees
extern "C" int dpt_igemm_set_lowp_dtype(int is_f16) {
    dpt::g_lowp_f16 = (is_f16 >= 0 && is_f16 <= 2) ? is_f16 : 0;
    return 0;
}

// Policy of the persistent kernel variant: used from `min_tiles` output tiles upwards (<= 0
// restores the default of 4 waves) with `grid_ctas` CTAs (<= 0 restores 2 per SM).  Tests use
// it to push small problems through the multi-tile path.
extern "C" int dpt_igemm_set_persist(long long min_tiles, int grid_ctas) {
    dpt::g_persist_min_tiles = min_tiles > 0 ? min_tiles : 4 * 148;
    dpt::g_persist_grid = grid_ctas > 0 ? grid_ctas : 2 * 148;
    return 0;
}

// 128 x 256 output tiles where the column count allows it (see pick_block_n); < 0 re-reads the
// DPT_IGEMM_BN256 environment variable.
extern "C" int dpt_igemm_set_bn256(int on) {
    dpt::g_bn256 = on < 0 ? -1 : (on ? 1 : 0);
    return 0;
}

extern "C" int dpt_conv_fprop(const void* x, const void* w, void* y, int N, int H, int W, int Cin,
                              int Cout, int R, int S, int stride, int pad, float* stat_sum,
                              float* stat_sq, int out_f32_atomic, int splits, void* stream) {
    if (Cin % (16 / esz()) != 0 || (stride != 1 && stride != 2) || R * S > kMaxTaps) return DPT_ERR_SHAPE;
    const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
    IgemmMaps maps;
    IgemmParams prm;
    memset(&prm, 0, sizeof(prm));
    pick_box(kTileM, Q, P, N, &prm.bq, &prm.bp, &prm.bn);
    int e = build_act_maps(maps.a, x, N, H, W, Cin, stride, kTileM, prm.bq, prm.bp, prm.bn);
    if (e) return e;
    const int block_n =
        pick_block_n(Cout, ((long long)N * P * Q + kTileM - 1) / kTileM);
    {
        const long long dims[2] = {(long long)R * S * Cin, Cout};
        const long long str[1] = {(long long)R * S * Cin};
        const int box[2] = {kblk(), block_n};
        e = make_tmap(&maps.b[0], w, 2, dims, str, box);
        if (e) return e;
        for (int i = 1; i < 4; ++i) maps.b[i] = maps.b[0];
    }
    TapBuilder tb;
    fprop_taps(&tb, H, W, P, Q, Cin, R, S, stride, pad, 0);
    if (tb.ntaps == 0) return DPT_ERR_SHAPE;
    memcpy(prm.taps, tb.taps, sizeof(IgemmTap) * tb.ntaps);
    prm.tiles_q = (Q + prm.bq - 1) / prm.bq;
    prm.tiles_p = (P + prm.bp - 1) / prm.bp;
    prm.tiles_n = (N + prm.bn - 1) / prm.bn;
    prm.Q = Q;
    prm.P = P;
    prm.NB = N;
    prm.chunks = (Cin + kblk() - 1) / kblk();
    const int ksteps = tb.ntaps * prm.chunks;
    if (splits < 1) splits = 1;
    if (splits > ksteps) splits = ksteps;
    if (!out_f32_atomic) splits = 1;
    // make sure no split is empty
    {
        int per = (ksteps + splits - 1) / splits;
        splits = (ksteps + per - 1) / per;
    }
    prm.splits = splits;
    prm.nclasses = 1;
    prm.cls_begin[0] = 0;
    prm.cls_begin[1] = tb.ntaps;
    prm.n_total = Cout;
    prm.out = y;
    prm.out_mode = out_f32_atomic ? OUT_F32_ATOMIC : lowp_out_mode();
    prm.ldo = Cout;
    prm.OH = P;
    prm.OW = Q;
    prm.osh = 1;
    prm.osw = 1;
    prm.stat_sum = out_f32_atomic ? nullptr : stat_sum;
    prm.stat_sq = out_f32_atomic ? nullptr : stat_sq;
    dim3 grid(prm.tiles_q * prm.tiles_p * prm.tiles_n, (Cout + block_n - 1) / block_n, splits);
    return launch<MODE_FPROP>(maps, prm, grid, block_n, static_cast<cudaStream_t>(stream));
}

// dX = conv_dgrad(dY, W) [+ conv_dgrad(dY2, W2) for a second conv reading the same input].
extern "C" int dpt_conv_dgrad_ex(const void* dy, const void* w, void* dx, int N, int H, int W,
                                 int Cin, int Cout, int R, int S, int stride, int pad,
                                 const void* dy2, const void* w2, int R2, int S2, int pad2,
                                 int out_f32_atomic, int splits, void* stream);

extern "C" int dpt_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin,
                              int Cout, int R, int S, int stride, int pad, const void* dy2,
                              const void* w2, int R2, int S2, int pad2, void* stream) {
    return dpt_conv_dgrad_ex(dy, w, dx, N, H, W, Cin, Cout, R, S, stride, pad, dy2, w2, R2, S2, pad2,
                             0, 1, stream);
}

// out_f32_atomic = 1: dx is fp32 (pre-zeroed) and the k loop is split over `splits` CTAs per
// tile, reduced with vector RED (fills the machine when the tile grid is tiny).
extern "C" int dpt_conv_dgrad_ex(const void* dy, const void* w, void* dx, int N, int H, int W,
                                 int Cin, int Cout, int R, int S, int stride, int pad,
                                 const void* dy2, const void* w2, int R2, int S2, int pad2,
                                 int out_f32_atomic, int splits, void* stream) {
    if (Cin % (16 / esz()) != 0 || Cout % (16 / esz()) != 0 || (stride != 1 && stride != 2)) return DPT_ERR_SHAPE;
    const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
    IgemmMaps maps;
    IgemmParams prm;
    memset(&prm, 0, sizeof(prm));
    // Row space: dX pixels of one parity class (all pixels for stride 1).
    const int HC = (H + stride - 1) / stride, WC = (W + stride - 1) / stride;
    pick_box(kTileM, WC, HC, N, &prm.bq, &prm.bp, &prm.bn);
    const int box[4] = {kblk(), prm.bq, prm.bp, prm.bn};
    {
        const long long dims[4] = {Cout, Q, P, N};
        const long long str[3] = {Cout, (long long)Q * Cout, (long long)P * Q * Cout};
        int e = make_tmap(&maps.a[0], dy, 4, dims, str, box);
        if (e) return e;
        maps.a[1] = maps.a[0];
        if (dy2 != nullptr) {
            const int P2 = (H + 2 * pad2 - R2) / stride + 1, Q2 = (W + 2 * pad2 - S2) / stride + 1;
            if (P2 != P || Q2 != Q) return DPT_ERR_SHAPE;
            e = make_tmap(&maps.a[1], dy2, 4, dims, str, box);
            if (e) return e;
        }
        maps.a[2] = maps.a[3] = maps.a[0];
    }
    const int block_n = pick_block_n(Cin, ((long long)N * H * W + kTileM - 1) / kTileM);
    {
        const long long dims[2] = {(long long)R * S * Cin, Cout};
        const long long str[1] = {(long long)R * S * Cin};
        const int bbox[2] = {kblk(), kblk()};
        int e = make_tmap(&maps.b[0], w, 2, dims, str, bbox, /*mn_major=*/true);
        if (e) return e;
        maps.b[1] = maps.b[0];
        if (dy2 != nullptr) {
            const long long dims2[2] = {(long long)R2 * S2 * Cin, Cout};
            const long long str2[1] = {(long long)R2 * S2 * Cin};
            e = make_tmap(&maps.b[1], w2, 2, dims2, str2, bbox, /*mn_major=*/true);
            if (e) return e;
        }
        maps.b[2] = maps.b[3] = maps.b[0];
    }
    // taps per parity class: dX[h] gets dY[(h + pad - r)/stride] * W[r] when divisible
    const int ncls = (stride == 1) ? 1 : 4;
    int nt = 0;
    for (int c = 0; c < ncls; ++c) {
        prm.cls_begin[c] = nt;
        const int hp = (stride == 1) ? 0 : (c >> 1), wp = (stride == 1) ? 0 : (c & 1);
        for (int which = 0; which < (dy2 ? 2 : 1); ++which) {
            const int RR = which ? R2 : R, SS = which ? S2 : S, PD = which ? pad2 : pad;
            for (int r = 0; r < RR; ++r)
                for (int s = 0; s < SS; ++s) {
                    if (fmod_pos(hp + PD - r, stride) != 0 || fmod_pos(wp + PD - s, stride) != 0)
                        continue;
                    const int dh = fdiv(hp + PD - r, stride), dw = fdiv(wp + PD - s, stride);
                    // range of dY rows touched: i + dh for i in [0, HC)
                    if (dh + HC - 1 < 0 || dh >= P || dw + WC - 1 < 0 || dw >= Q) continue;
                    if (nt >= kMaxTaps) return DPT_ERR_SHAPE;
                    IgemmTap t{};
                    t.a_map = which;
                    t.b_map = which;
                    t.dh = dh;
                    t.dw = dw;
                    t.koff = (r * SS + s) * Cin;
                    prm.taps[nt++] = t;
                }
        }
    }
    for (int c = ncls; c < 5; ++c) prm.cls_begin[c] = nt;
    prm.tiles_q = (WC + prm.bq - 1) / prm.bq;
    prm.tiles_p = (HC + prm.bp - 1) / prm.bp;
    prm.tiles_n = (N + prm.bn - 1) / prm.bn;
    prm.Q = WC;
    prm.P = HC;
    prm.NB = N;
    // (odd H/W with stride 2: classes hp=1 have one row fewer; mask through OH/OW below)
    prm.chunks = (Cout + kblk() - 1) / kblk();
    if (!out_f32_atomic || splits < 1) splits = 1;
    prm.splits = splits;
    prm.nclasses = ncls;
    prm.n_total = Cin;
    prm.out = dx;
    prm.out_mode = out_f32_atomic ? OUT_F32_ATOMIC : lowp_out_mode();
    prm.ldo = Cin;
    prm.OH = H;
    prm.OW = W;
    prm.osh = stride;
    prm.osw = stride;
    if (stride == 2 && ((H & 1) || (W & 1))) return DPT_ERR_SHAPE;  // TODO: odd sizes
    dim3 grid(prm.tiles_q * prm.tiles_p * prm.tiles_n, (Cin + block_n - 1) / block_n, ncls * splits);
    return launch<MODE_DGRAD>(maps, prm, grid, block_n, static_cast<cudaStream_t>(stream));
}

// dW[Cout][R][S][Cin] (fp32) += dY^T * im2col(X), reduced across pixel-tile splits with RED.
extern "C" int dpt_conv_wgrad(const void* dy, const void* x, float* dw, int N, int H, int W,
                              int Cin, int Cout, int R, int S, int stride, int pad, int splits,
                              void* stream) {
    if (Cin % (16 / esz()) != 0 || Cout % (16 / esz()) != 0 || (stride != 1 && stride != 2)) return DPT_ERR_SHAPE;
    const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
    IgemmMaps maps;
    IgemmParams prm;
    memset(&prm, 0, sizeof(prm));
    pick_box(kblk(), Q, P, N, &prm.bq, &prm.bp, &prm.bn);
    const int box[4] = {kblk(), prm.bq, prm.bp, prm.bn};
    {
        const long long dims[4] = {Cout, Q, P, N};
        const long long str[3] = {Cout, (long long)Q * Cout, (long long)P * Q * Cout};
        int e = make_tmap(&maps.a[0], dy, 4, dims, str, box, /*mn_major=*/true);
        if (e) return e;
        for (int i = 1; i < 4; ++i) maps.a[i] = maps.a[0];
    }
    int e = build_act_maps(maps.b, x, N, H, W, Cin, stride, kblk(), prm.bq, prm.bp, prm.bn,
                           /*mn_major=*/true);
    if (e) return e;
    TapBuilder tb;
    fprop_taps(&tb, H, W, P, Q, Cin, R, S, stride, pad, 0);
    if (tb.ntaps == 0) return DPT_ERR_SHAPE;
    for (int i = 0; i < tb.ntaps; ++i) {
        tb.taps[i].b_map = tb.taps[i].a_map;  // activation parity view lives on the B side here
    }
    memcpy(prm.taps, tb.taps, sizeof(IgemmTap) * tb.ntaps);
    const int block_n = pick_block_n(Cin, 1 << 20);   // wgrad fills the SMs through its splits
    prm.tiles_q = (Q + prm.bq - 1) / prm.bq;
    prm.tiles_p = (P + prm.bp - 1) / prm.bp;
    prm.tiles_n = (N + prm.bn - 1) / prm.bn;
    const int mtiles = prm.tiles_q * prm.tiles_p * prm.tiles_n;
    prm.ntile_per_tap = (Cin + block_n - 1) / block_n;
    const int out_tiles = tb.ntaps * prm.ntile_per_tap * ((Cout + kTileM - 1) / kTileM);
    if (splits <= 0) {
        // fill ~2 waves of the 148 SMs, but keep at least 2 pixel tiles per CTA
        splits = (296 + out_tiles - 1) / out_tiles;
        if (splits > (mtiles + 1) / 2) splits = (mtiles + 1) / 2;
    }
    if (splits < 1) splits = 1;
    if (splits > mtiles) splits = mtiles;
    {
        int per = (mtiles + splits - 1) / splits;
        splits = (mtiles + per - 1) / per;
    }
    prm.splits = splits;
    prm.nclasses = 1;
    prm.n_total = Cin;
    prm.m_total = Cout;
    prm.out = dw;
    prm.out_mode = OUT_F32_ATOMIC;
    prm.ldo = (long long)R * S * Cin;
    dim3 grid(tb.ntaps * prm.ntile_per_tap, (Cout + kTileM - 1) / kTileM, splits);
    return launch<MODE_WGRAD>(maps, prm, grid, block_n, static_cast<cudaStream_t>(stream));
}

// Plain bf16 GEMMs, fp32 accumulate.  Row-major everywhere.
//   op 0 (NT): C[M,N] = A[M,K] * B[N,K]^T        (linear forward, im2col conv forward)
//   op 1 (NN): C[M,N] = A[M,K] * B[K,N]          (linear dgrad)
//   op 2 (TN): C[M,N] = A[K,M]^T * B[K,N]        (linear / im2col wgrad; split over K with RED)
// out_mode: 0 bf16 store, 1 fp32 store, 2 fp32 atomic accumulate.
extern "C" int dpt_gemm_bf16(int op, const void* A, const void* B, void* C, int M, int N, int K,
                             long long lda, long long ldb, long long ldc, int out_mode,
                             const float* bias, float* stat_sum, float* stat_sq, int splits,
                             void* stream) {
    if (lda % (16 / esz()) != 0 || ldb % (16 / esz()) != 0 || op < 0 || op > 2) return DPT_ERR_SHAPE;
    if (out_mode == OUT_BF16) out_mode = lowp_out_mode();
    IgemmMaps maps;
    IgemmParams prm;
    memset(&prm, 0, sizeof(prm));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int block_n = pick_block_n(N, op == 2 ? (1 << 20) : (M + kTileM - 1) / kTileM);
    prm.nclasses = 1;
    prm.cls_begin[0] = 0;
    for (int c = 1; c < 5; ++c) prm.cls_begin[c] = 1;
    prm.n_total = N;
    prm.out = C;
    prm.ldo = ldc;
    prm.bias = bias;
    prm.OH = 1;
    prm.osh = prm.osw = 1;
    prm.bp = prm.bn = 1;
    prm.tiles_p = prm.tiles_n = 1;
    prm.P = prm.NB = 1;
    if (splits < 1) splits = 1;
    if (op == 0 || op == 1) {
        const long long dims[4] = {K, M, 1, 1};
        const long long str[3] = {lda, lda * M, lda * M};
        const int box[4] = {kblk(), kTileM, 1, 1};
        int e = make_tmap(&maps.a[0], A, 4, dims, str, box);
        if (e) return e;
        if (op == 0) {
            const long long bd[2] = {K, N};
            const long long bs[1] = {ldb};
            const int bb[2] = {kblk(), block_n};
            e = make_tmap(&maps.b[0], B, 2, bd, bs, bb);
        } else {
            const long long bd[2] = {N, K};
            const long long bs[1] = {ldb};
            const int bb[2] = {kblk(), kblk()};
            e = make_tmap(&maps.b[0], B, 2, bd, bs, bb, /*mn_major=*/true);
        }
        if (e) return e;
        for (int i = 1; i < 4; ++i) {
            maps.a[i] = maps.a[0];
            maps.b[i] = maps.b[0];
        }
        prm.bq = kTileM;
        prm.tiles_q = (M + kTileM - 1) / kTileM;
        prm.Q = M;
        prm.OW = M;
        prm.chunks = (K + kblk() - 1) / kblk();
        if (out_mode != OUT_F32_ATOMIC) splits = 1;
        if (splits > prm.chunks) splits = prm.chunks;
        {
            int per = (prm.chunks + splits - 1) / splits;
            splits = (prm.chunks + per - 1) / per;
        }
        prm.splits = splits;
        prm.out_mode = out_mode;
        prm.stat_sum = (op == 0 && out_mode != OUT_F32_ATOMIC) ? stat_sum : nullptr;
        prm.stat_sq = (op == 0 && out_mode != OUT_F32_ATOMIC) ? stat_sq : nullptr;
        dim3 grid(prm.tiles_q, (N + block_n - 1) / block_n, splits);
        if (op == 0) return launch<MODE_FPROP>(maps, prm, grid, block_n, st);
        return launch<MODE_DGRAD>(maps, prm, grid, block_n, st);
    }
    // op 2: reduction over the rows of both operands
    {
        const long long ad[4] = {M, K, 1, 1};
        const long long as[3] = {lda, lda * K, lda * K};
        const long long bd[4] = {N, K, 1, 1};
        const long long bs[3] = {ldb, ldb * K, ldb * K};
        const int box[4] = {kblk(), kblk(), 1, 1};
        int e = make_tmap(&maps.a[0], A, 4, ad, as, box, /*mn_major=*/true);
        if (e) return e;
        e = make_tmap(&maps.b[0], B, 4, bd, bs, box, /*mn_major=*/true);
        if (e) return e;
        for (int i = 1; i < 4; ++i) {
            maps.a[i] = maps.a[0];
            maps.b[i] = maps.b[0];
        }
    }
    prm.bq = kblk();
    prm.tiles_q = (K + kblk() - 1) / kblk();
    prm.m_total = M;
    prm.ntile_per_tap = (N + block_n - 1) / block_n;
    if (out_mode != OUT_F32_ATOMIC) splits = 1;
    if (splits > prm.tiles_q) splits = prm.tiles_q;
    {
        int per = (prm.tiles_q + splits - 1) / splits;
        splits = (prm.tiles_q + per - 1) / per;
    }
    prm.splits = splits;
    prm.out_mode = out_mode;
    dim3 grid(prm.ntile_per_tap, (M + kTileM - 1) / kTileM, splits);
    return launch<MODE_WGRAD>(maps, prm, grid, block_n, st);
}
