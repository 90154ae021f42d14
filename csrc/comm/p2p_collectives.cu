// NVLink-5 / NVSwitch peer-memory collectives for the gradient path (sm_100a).
//
// Every rank owns a "symmetric" arena (same layout on every GPU) that all peers have mapped
// through CUDA IPC.  The kernels below read and write peers' arenas directly with 16-byte
// loads/stores; cross-GPU ordering uses monotonically increasing 32-bit epochs in a per-rank
// signal pad (st.release.sys / ld.acquire.sys), one slot per (CTA, peer), so no flag is ever
// reset and a captured CUDA graph can be replayed (the epoch counters live on the device).
//
//   two-shot all-reduce : rank r reduces slice r from all peers, applies the fused epilogue,
//                         and stores the result into every peer (reduce-scatter + all-gather
//                         in one kernel, 2(N-1)/N * S bytes per GPU per direction)
//   one-shot all-reduce : every rank reads all peers and reduces locally (small messages)
//   broadcast           : the source rank stores into every peer (initial parameter sync and
//                         the per-step BatchNorm buffer sync)
//
// Fused epilogues on the reduced value while it is still in registers:
//   F1: x (1/world) x inv_scale, non-finite check -> found_inf word on every rank
//   F2: F1 + SGD-momentum update of the fp32 master slice, new fp32 weights and bf16 shadow
//       pushed to every peer (momentum stays sharded on the owner, ZeRO-1 style)
//
// Replaces torch DDP's Reducer -> ProcessGroupNCCL allreduce (reference train_ddp.py:305-310,
// 207/213), GradScaler's unscale_ + found_inf.item() (:208) and, in F2, optimizer.step() (:214).
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>

#include "../../include/dptb200.h"
#include "../common/ptx.cuh"

namespace dpt {

constexpr int kMaxWorld = 8;
constexpr int kMaxCommCtas = 64;
constexpr int kCommThreads = 512;

struct CommCtx {
    int rank, world;
    uint32_t* flags[kMaxWorld];  // signal pad of every rank: [kMaxCommCtas][kMaxWorld] u32
    uint32_t* epoch;             // local: [kMaxCommCtas] barrier counters
    uint32_t* err;               // local (host-mapped) error word, 0 = ok
    unsigned long long timeout_ns;
    // optional timeline buffer (dpt_comm_set_trace): every collective kernel appends
    // {kind/bytes, start, after first barrier, end} globaltimer stamps (CTA 0, thread 0)
    unsigned long long* trace;
    uint32_t trace_cap;          // records of 4 u64; trace[0] is the atomic cursor
};

__device__ __forceinline__ unsigned long long globaltimer_ns();

// Timeline record for the step-trace tool (benchmarks/step_trace.py): 4 u64 per record.
struct TraceRec {
    unsigned long long* slot;
    __device__ __forceinline__ void begin(const CommCtx& c, uint32_t kind, long long bytes) {
        slot = nullptr;
        if (c.trace == nullptr || blockIdx.x != 0 || threadIdx.x != 0) return;
        const unsigned long long i = atomicAdd(c.trace, 1ULL);
        if (i >= c.trace_cap) return;
        slot = c.trace + 4 + 4 * i;
        slot[0] = (static_cast<unsigned long long>(kind) << 56) | static_cast<unsigned long long>(bytes);
        slot[1] = globaltimer_ns();
        slot[2] = 0;
        slot[3] = 0;
    }
    __device__ __forceinline__ void mark(int k) {
        if (slot != nullptr) slot[k] = globaltimer_ns();
    }
};

struct Peers {
    void* p[kMaxWorld];
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// All-ranks barrier for CTA index blockIdx.x.  Establishes: everything this CTA wrote (to any
// GPU) before the call is visible to the matching CTA of every peer after its call returns.
__device__ __forceinline__ void xgpu_barrier(const CommCtx& c, uint32_t e) {
    __syncthreads();
    if (threadIdx.x < c.world) {
        const int peer = threadIdx.x;
        uint32_t* remote = c.flags[peer] + blockIdx.x * kMaxWorld + c.rank;
        __threadfence_system();
        st_release_sys_u32(remote, e);
        const uint32_t* mine = c.flags[c.rank] + blockIdx.x * kMaxWorld + peer;
        const unsigned long long t0 = globaltimer_ns();
        while (static_cast<int32_t>(ld_acquire_sys_u32(mine) - e) < 0) {
            if (c.timeout_ns != 0 && globaltimer_ns() - t0 > c.timeout_ns) {
                // peer never arrived: report (rank, peer, epoch) and stop waiting
                atomicCAS(c.err, 0u, 0x80000000u | (uint32_t(peer) << 24) | (e & 0xffffffu));
                break;
            }
        }
    }
    __syncthreads();
}

__device__ __forceinline__ float4 ld_stream(const float4* p) {
    float4 r;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ void st_stream(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x),
                 "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

struct Epilogue {
    float scale;              // 1/world (times a static factor)
    const float* inv_scale;   // optional device scalar (AMP 1/loss_scale)
    float* found_inf[kMaxWorld];  // found_inf word of every rank (may be null)
    // F2 (fused SGD) -- all null/0 for F1
    float* param[kMaxWorld];          // fp32 master weights of every rank
    __nv_bfloat16* shadow[kMaxWorld]; // bf16 shadow of every rank
    float* momentum;                  // local momentum (only the owner's slice is touched)
    float lr, mu, wd;
    int nesterov;
    int fused_sgd;
    // NVLS + fused SGD: multicast mappings of the param / shadow arenas
    float* param_mc;
    __nv_bfloat16* shadow_mc;
    int shadow_f16;  // 16-bit shadow format: 0 = bf16, 1 = fp16
    // 1 = leave without the closing cross-GPU barrier.  Legal for every bucket kernel of a step
    // but the last: buckets are disjoint, and the opening barrier of the NEXT collective on the
    // same stream makes this kernel's peer stores visible (a rank only arrives there after this
    // kernel completed); only the step's final collective has to prove that everything landed
    // before the step's consumers (next forward, next gradient memset) run.
    int skip_closing_barrier;
};

static int g_comm_shadow_f16 = 0;
static thread_local int g_skip_closing_next = 0;   // consumed by the next fused all-reduce launch
__device__ __forceinline__ uint32_t comm_pack_lp2(float a, float b, int f16) {
    if (f16) {
        __half2 h = __floats2half2_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&h);
    }
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}


// NVLS: multimem.ld_reduce returns the in-switch sum of every rank's copy, multimem.st writes
// every copy.  Rank r still owns slice r (so each element is reduced and broadcast exactly
// once); link traffic per GPU is ~S instead of 2(N-1)/N*S and the reduction costs no SM adds.
__device__ __forceinline__ float4 multimem_ld_reduce_f32x4(const float4* mc) {
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(mc)
                 : "memory");
    return r;
}
__device__ __forceinline__ void multimem_st_b64(void* mc, unsigned long long v) {
    asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(mc), "l"(v) : "memory");
}
__device__ __forceinline__ void multimem_st_f32x4(float4* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x),
                 "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// grads: symmetric buffers (float), n4 = number of float4.  In place.
template <int WORLD>
__global__ void __launch_bounds__(kCommThreads, 1)
allreduce_twoshot_kernel(CommCtx c, Peers grads, long long n4, Epilogue ep) {
    __shared__ uint32_t s_epoch;
    if (threadIdx.x == 0) s_epoch = c.epoch[blockIdx.x];
    __syncthreads();
    const uint32_t e0 = s_epoch;
    TraceRec tr;
    tr.begin(c, 2u, n4 * 16);
    xgpu_barrier(c, e0 + 1);  // every peer's bucket is complete and its consumers are done
    tr.mark(2);
    constexpr int kUnroll = WORLD > 4 ? 2 : 4;  // keep WORLD * kUnroll float4 in registers

    const long long per = (n4 + WORLD - 1) / WORLD;
    const long long lo = per * c.rank;
    const long long hi = min(n4, lo + per);
    const float gs = ep.scale * (ep.inv_scale ? *ep.inv_scale : 1.f);
    bool bad = false;
    const long long step = static_cast<long long>(gridDim.x) * kCommThreads;
    for (long long i0 = lo + static_cast<long long>(blockIdx.x) * kCommThreads + threadIdx.x;
         i0 < hi; i0 += step * kUnroll) {
        float4 v[kUnroll][WORLD];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const long long i = i0 + u * step;
            if (i < hi) {
#pragma unroll
                for (int p = 0; p < WORLD; ++p)
                    v[u][p] = ld_stream(reinterpret_cast<const float4*>(grads.p[p]) + i);
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const long long i = i0 + u * step;
            if (i >= hi) continue;
            // fixed summation order => bitwise identical on every rank
            float4 s = v[u][0];
#pragma unroll
            for (int p = 1; p < WORLD; ++p) {
                s.x += v[u][p].x;
                s.y += v[u][p].y;
                s.z += v[u][p].z;
                s.w += v[u][p].w;
            }
            bad |= !(isfinite(s.x) && isfinite(s.y) && isfinite(s.z) && isfinite(s.w));
            s.x *= gs;
            s.y *= gs;
            s.z *= gs;
            s.w *= gs;
            if (!ep.fused_sgd) {
#pragma unroll
                for (int p = 0; p < WORLD; ++p)
                    st_stream(reinterpret_cast<float4*>(grads.p[p]) + i, s);
            } else {
                float4 pv = reinterpret_cast<float4*>(ep.param[c.rank])[i];
                float4 mv = reinterpret_cast<float4*>(ep.momentum)[i];
                float pa[4] = {pv.x, pv.y, pv.z, pv.w};
                float ma[4] = {mv.x, mv.y, mv.z, mv.w};
                const float ga[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float d = fmaf(ep.wd, pa[k], ga[k]);
                    ma[k] = fmaf(ep.mu, ma[k], d);
                    d = ep.nesterov ? fmaf(ep.mu, ma[k], d) : ma[k];
                    pa[k] = fmaf(-ep.lr, d, pa[k]);
                }
                reinterpret_cast<float4*>(ep.momentum)[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
                const float4 np = make_float4(pa[0], pa[1], pa[2], pa[3]);
                uint2 pk;
                pk.x = comm_pack_lp2(pa[0], pa[1], ep.shadow_f16);
                pk.y = comm_pack_lp2(pa[2], pa[3], ep.shadow_f16);
                if (ep.param_mc != nullptr) {
                    // NVSwitch multicast push (dpt_allreduce_fused_mcpush): ONE store each for the
                    // new weights and their 16-bit shadow, replicated to every rank by the switch,
                    // instead of WORLD peer stores each — 8x fewer store instructions and 8x less
                    // outbound link traffic for this phase at 8 GPUs
                    multimem_st_f32x4(reinterpret_cast<float4*>(ep.param_mc) + i, np);
                    multimem_st_b64(reinterpret_cast<uint2*>(ep.shadow_mc) + i,
                                    static_cast<unsigned long long>(pk.x) |
                                        (static_cast<unsigned long long>(pk.y) << 32));
                } else {
#pragma unroll
                    for (int p = 0; p < WORLD; ++p) {
                        st_stream(reinterpret_cast<float4*>(ep.param[p]) + i, np);
                        reinterpret_cast<uint2*>(ep.shadow[p])[i] = pk;
                    }
                }
            }
        }
    }
    if (ep.found_inf[0] != nullptr && __any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) {
#pragma unroll
        for (int p = 0; p < WORLD; ++p) *ep.found_inf[p] = 1.f;
    }
    if (ep.skip_closing_barrier) {
        tr.mark(3);
        if (threadIdx.x == 0) c.epoch[blockIdx.x] = e0 + 1;
        return;
    }
    xgpu_barrier(c, e0 + 2);  // all slices have landed everywhere
    tr.mark(3);
    if (threadIdx.x == 0) c.epoch[blockIdx.x] = e0 + 2;
}

// One-shot: out (local, may alias nothing symmetric that peers read) = epilogue(sum_p in_p).
template <int WORLD>
__global__ void __launch_bounds__(kCommThreads, 1)
allreduce_oneshot_kernel(CommCtx c, Peers in, float* __restrict__ out, long long n4, Epilogue ep) {
    __shared__ uint32_t s_epoch;
    if (threadIdx.x == 0) s_epoch = c.epoch[blockIdx.x];
    __syncthreads();
    const uint32_t e0 = s_epoch;
    xgpu_barrier(c, e0 + 1);
    const float gs = ep.scale * (ep.inv_scale ? *ep.inv_scale : 1.f);
    bool bad = false;
    const long long step = static_cast<long long>(gridDim.x) * kCommThreads;
    for (long long i = static_cast<long long>(blockIdx.x) * kCommThreads + threadIdx.x; i < n4;
         i += step) {
        float4 v[WORLD];
#pragma unroll
        for (int p = 0; p < WORLD; ++p) v[p] = ld_stream(reinterpret_cast<const float4*>(in.p[p]) + i);
        float4 s = v[0];
#pragma unroll
        for (int p = 1; p < WORLD; ++p) {
            s.x += v[p].x;
            s.y += v[p].y;
            s.z += v[p].z;
            s.w += v[p].w;
        }
        bad |= !(isfinite(s.x) && isfinite(s.y) && isfinite(s.z) && isfinite(s.w));
        s.x *= gs;
        s.y *= gs;
        s.z *= gs;
        s.w *= gs;
        reinterpret_cast<float4*>(out)[i] = s;
    }
    if (ep.found_inf[0] != nullptr && __any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0)
        *ep.found_inf[c.rank] = 1.f;
    xgpu_barrier(c, e0 + 2);  // peers are done reading my input: it may be overwritten
    if (threadIdx.x == 0) c.epoch[blockIdx.x] = e0 + 2;
}

// Broadcast n16 16-byte words from rank `src` into every peer's buffer (in place).
__global__ void __launch_bounds__(kCommThreads, 1)
broadcast_kernel(CommCtx c, Peers bufs, long long n16, int src) {
    __shared__ uint32_t s_epoch;
    if (threadIdx.x == 0) s_epoch = c.epoch[blockIdx.x];
    __syncthreads();
    const uint32_t e0 = s_epoch;
    xgpu_barrier(c, e0 + 1);  // receivers are ready to be overwritten
    if (c.rank == src) {
        const long long step = static_cast<long long>(gridDim.x) * kCommThreads;
        for (long long i = static_cast<long long>(blockIdx.x) * kCommThreads + threadIdx.x; i < n16;
             i += step) {
            const float4 v = reinterpret_cast<const float4*>(bufs.p[src])[i];
            for (int p = 0; p < c.world; ++p)
                if (p != src) st_stream(reinterpret_cast<float4*>(bufs.p[p]) + i, v);
        }
    }
    xgpu_barrier(c, e0 + 2);
    if (threadIdx.x == 0) c.epoch[blockIdx.x] = e0 + 2;
}

__global__ void __launch_bounds__(kCommThreads, 1)
allreduce_nvls_kernel(CommCtx c, float* __restrict__ mc, long long n4, Epilogue ep) {
    __shared__ uint32_t s_epoch;
    if (threadIdx.x == 0) s_epoch = c.epoch[blockIdx.x];
    __syncthreads();
    const uint32_t e0 = s_epoch;
    TraceRec tr;
    tr.begin(c, 3u, n4 * 16);
    xgpu_barrier(c, e0 + 1);
    tr.mark(2);
    const long long per = (n4 + c.world - 1) / c.world;
    const long long lo = per * c.rank;
    const long long hi = min(n4, lo + per);
    const float gs = ep.scale * (ep.inv_scale ? *ep.inv_scale : 1.f);
    bool bad = false;
    float4* mc4 = reinterpret_cast<float4*>(mc);
    const long long step = static_cast<long long>(gridDim.x) * kCommThreads;
    constexpr int kU = 4;
    for (long long i0 = lo + static_cast<long long>(blockIdx.x) * kCommThreads + threadIdx.x;
         i0 < hi; i0 += step * kU) {
        float4 v[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u)
            if (i0 + u * step < hi) v[u] = multimem_ld_reduce_f32x4(mc4 + i0 + u * step);
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const long long i = i0 + u * step;
            if (i >= hi) continue;
            float4 s = v[u];
            bad |= !(isfinite(s.x) && isfinite(s.y) && isfinite(s.z) && isfinite(s.w));
            s.x *= gs;
            s.y *= gs;
            s.z *= gs;
            s.w *= gs;
            if (!ep.fused_sgd) {
                multimem_st_f32x4(mc4 + i, s);
            } else {
                // owner updates its slice of the fp32 master weights + momentum, then one
                // multicast store each publishes the new weights and their bf16 shadow everywhere
                float4 pv = reinterpret_cast<float4*>(ep.param[c.rank])[i];
                float4 mv = reinterpret_cast<float4*>(ep.momentum)[i];
                float pa[4] = {pv.x, pv.y, pv.z, pv.w};
                float ma[4] = {mv.x, mv.y, mv.z, mv.w};
                const float ga[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float d = fmaf(ep.wd, pa[k], ga[k]);
                    ma[k] = fmaf(ep.mu, ma[k], d);
                    d = ep.nesterov ? fmaf(ep.mu, ma[k], d) : ma[k];
                    pa[k] = fmaf(-ep.lr, d, pa[k]);
                }
                reinterpret_cast<float4*>(ep.momentum)[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
                multimem_st_f32x4(reinterpret_cast<float4*>(ep.param_mc) + i,
                                  make_float4(pa[0], pa[1], pa[2], pa[3]));
                const unsigned long long pk =
                    static_cast<unsigned long long>(comm_pack_lp2(pa[0], pa[1], ep.shadow_f16)) |
                    (static_cast<unsigned long long>(comm_pack_lp2(pa[2], pa[3], ep.shadow_f16)) << 32);
                multimem_st_b64(reinterpret_cast<uint2*>(ep.shadow_mc) + i, pk);
            }
        }
    }
    if (ep.found_inf[0] != nullptr && __any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) {
        for (int p = 0; p < c.world; ++p) *ep.found_inf[p] = 1.f;
    }
    if (ep.skip_closing_barrier) {
        tr.mark(3);
        if (threadIdx.x == 0) c.epoch[blockIdx.x] = e0 + 1;
        return;
    }
    xgpu_barrier(c, e0 + 2);
    tr.mark(3);
    if (threadIdx.x == 0) c.epoch[blockIdx.x] = e0 + 2;
}

// One-shot "push" all-reduce for small buckets (the exposed tail of backward): every rank stores
// its contribution into slot [rank] of a staging area on every peer, ONE cross-GPU barrier, then
// every rank reduces the WORLD slots it received from its own memory and applies the epilogue to
// its own copy — no second barrier, no remote loads, (WORLD-1) x S bytes pushed per rank.
//   * every rank sums in the same fixed order, so results are bitwise identical everywhere;
//   * with the SGD update fused, each rank updates its own fp32 masters, momentum and 16-bit
//     shadow for the WHOLE bucket (replicated update: nothing to publish, momentum not sharded);
//   * the staging area is double buffered on a per-CTA use counter (epoch[kMaxCommCtas + cta]),
//     so a fast rank pushing for step t+1 can never overwrite slots a slow rank still reduces
//     for step t (it cannot be two uses ahead: each use ends behind a barrier with that rank).
// Replaces, for small messages, the NCCL all-reduce torch DDP issues per bucket
// (train_ddp.py:305-310; LL protocol in NCCL) + unscale + optimizer step.
template <int WORLD>
__global__ void __launch_bounds__(kCommThreads, 1)
allreduce_oneshot_push_kernel(CommCtx c, float* __restrict__ grads, Peers staging, long long cap4,
                              long long n4, Epilogue ep) {
    __shared__ uint32_t s_epoch, s_use;
    if (threadIdx.x == 0) {
        s_epoch = c.epoch[blockIdx.x];
        s_use = c.epoch[kMaxCommCtas + blockIdx.x];
    }
    __syncthreads();
    const uint32_t e0 = s_epoch;
    TraceRec tr;
    tr.begin(c, 1u, n4 * 16);
    const long long half = static_cast<long long>(s_use & 1u) * WORLD * cap4;
    const long long step = static_cast<long long>(gridDim.x) * kCommThreads;
    const long long first = static_cast<long long>(blockIdx.x) * kCommThreads + threadIdx.x;
    const float4* g4 = reinterpret_cast<const float4*>(grads);
    for (long long i = first; i < n4; i += step) {
        const float4 v = g4[i];
#pragma unroll
        for (int p = 0; p < WORLD; ++p)
            st_stream(reinterpret_cast<float4*>(staging.p[p]) + half + c.rank * cap4 + i, v);
    }
    xgpu_barrier(c, e0 + 1);  // CTA b of every peer has pushed exactly the elements CTA b reduces
    tr.mark(2);
    const float gs = ep.scale * (ep.inv_scale ? *ep.inv_scale : 1.f);
    bool bad = false;
    const float4* mine = reinterpret_cast<const float4*>(staging.p[c.rank]) + half;
    for (long long i = first; i < n4; i += step) {
        float4 s = ld_stream(mine + i);
#pragma unroll
        for (int p = 1; p < WORLD; ++p) {
            const float4 v = ld_stream(mine + p * cap4 + i);
            s.x += v.x;
            s.y += v.y;
            s.z += v.z;
            s.w += v.w;
        }
        bad |= !(isfinite(s.x) && isfinite(s.y) && isfinite(s.z) && isfinite(s.w));
        s.x *= gs;
        s.y *= gs;
        s.z *= gs;
        s.w *= gs;
        if (!ep.fused_sgd) {
            reinterpret_cast<float4*>(grads)[i] = s;
        } else {
            float4 pv = reinterpret_cast<float4*>(ep.param[c.rank])[i];
            float4 mv = reinterpret_cast<float4*>(ep.momentum)[i];
            float pa[4] = {pv.x, pv.y, pv.z, pv.w};
            float ma[4] = {mv.x, mv.y, mv.z, mv.w};
            const float ga[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float d = fmaf(ep.wd, pa[k], ga[k]);
                ma[k] = fmaf(ep.mu, ma[k], d);
                d = ep.nesterov ? fmaf(ep.mu, ma[k], d) : ma[k];
                pa[k] = fmaf(-ep.lr, d, pa[k]);
            }
            reinterpret_cast<float4*>(ep.momentum)[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
            reinterpret_cast<float4*>(ep.param[c.rank])[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
            uint2 pk;
            pk.x = comm_pack_lp2(pa[0], pa[1], ep.shadow_f16);
            pk.y = comm_pack_lp2(pa[2], pa[3], ep.shadow_f16);
            reinterpret_cast<uint2*>(ep.shadow[c.rank])[i] = pk;
        }
    }
    if (ep.found_inf[c.rank] != nullptr && __any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0)
        *ep.found_inf[c.rank] = 1.f;   // every rank computes the same sums: a local flag suffices
    tr.mark(3);
    if (threadIdx.x == 0) {
        c.epoch[blockIdx.x] = e0 + 1;
        c.epoch[kMaxCommCtas + blockIdx.x] = s_use + 1;
    }
}

__global__ void barrier_kernel(CommCtx c) {
    __shared__ uint32_t s_epoch;
    if (threadIdx.x == 0) s_epoch = c.epoch[blockIdx.x];
    __syncthreads();
    const uint32_t e0 = s_epoch;
    xgpu_barrier(c, e0 + 1);
    if (threadIdx.x == 0) c.epoch[blockIdx.x] = e0 + 1;
}

// Local (world == 1) version of the F1/F2 epilogue so single-GPU runs share the code path.
__global__ void __launch_bounds__(256)
local_epilogue_kernel(float* __restrict__ g, long long n4, Epilogue ep) {
    const float gs = ep.scale * (ep.inv_scale ? *ep.inv_scale : 1.f);
    bool bad = false;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float4 s = reinterpret_cast<float4*>(g)[i];
        bad |= !(isfinite(s.x) && isfinite(s.y) && isfinite(s.z) && isfinite(s.w));
        s.x *= gs;
        s.y *= gs;
        s.z *= gs;
        s.w *= gs;
        reinterpret_cast<float4*>(g)[i] = s;
    }
    if (ep.found_inf[0] != nullptr && __any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0)
        *ep.found_inf[0] = 1.f;
}

}  // namespace dpt

using namespace dpt;

// ---------------------------------------------------------------------------- host objects
struct dpt_comm {
    CommCtx ctx;
    int device;
    uint32_t* flags_local;
    uint32_t* err_host;  // pinned, mapped
    void* opened[kMaxWorld];
};

extern "C" int dpt_comm_set_lowp_dtype(int is_f16) {
    dpt::g_comm_shadow_f16 = (is_f16 == 1) ? 1 : 0;
    return 0;
}

extern "C" int dpt_symm_alloc(size_t bytes, void** ptr, void* ipc_handle_64) {
    cudaError_t e = cudaMalloc(ptr, bytes);
    if (e != cudaSuccess) return static_cast<int>(e);
    e = cudaMemset(*ptr, 0, bytes);
    if (e != cudaSuccess) return static_cast<int>(e);
    cudaIpcMemHandle_t h;
    e = cudaIpcGetMemHandle(&h, *ptr);
    if (e != cudaSuccess) return static_cast<int>(e);
    static_assert(sizeof(h) == 64, "ipc handle size");
    memcpy(ipc_handle_64, &h, 64);
    return 0;
}

extern "C" int dpt_symm_open(const void* ipc_handle_64, void** ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle_64, 64);
    return static_cast<int>(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
}

extern "C" int dpt_symm_close(void* ptr) { return static_cast<int>(cudaIpcCloseMemHandle(ptr)); }
extern "C" int dpt_symm_free(void* ptr) { return static_cast<int>(cudaFree(ptr)); }

extern "C" size_t dpt_comm_flag_bytes() { return sizeof(uint32_t) * kMaxCommCtas * kMaxWorld; }

// flags[r] = this process's mapping of rank r's signal pad (flags[rank] is the local one).
extern "C" int dpt_comm_create(int rank, int world, void* const* flags, double timeout_s,
                               dpt_comm** out) {
    if (world < 1 || world > kMaxWorld || rank < 0 || rank >= world) return DPT_ERR_ARG;
    dpt_comm* c = new dpt_comm();
    memset(c, 0, sizeof(*c));
    c->ctx.rank = rank;
    c->ctx.world = world;
    for (int r = 0; r < world; ++r) c->ctx.flags[r] = static_cast<uint32_t*>(flags[r]);
    // [0, kMaxCommCtas): barrier epochs; [kMaxCommCtas, 2*kMaxCommCtas): one-shot use counters
    cudaError_t e = cudaMalloc(&c->ctx.epoch, sizeof(uint32_t) * 2 * kMaxCommCtas);
    if (e != cudaSuccess) return static_cast<int>(e);
    cudaMemset(c->ctx.epoch, 0, sizeof(uint32_t) * 2 * kMaxCommCtas);
    e = cudaHostAlloc(reinterpret_cast<void**>(&c->err_host), sizeof(uint32_t), cudaHostAllocMapped);
    if (e != cudaSuccess) return static_cast<int>(e);
    *c->err_host = 0;
    void* dptr = nullptr;
    e = cudaHostGetDevicePointer(&dptr, c->err_host, 0);
    if (e != cudaSuccess) return static_cast<int>(e);
    c->ctx.err = static_cast<uint32_t*>(dptr);
    c->ctx.timeout_ns = static_cast<unsigned long long>(timeout_s * 1e9);
    cudaGetDevice(&c->device);
    *out = c;
    return 0;
}

extern "C" int dpt_comm_destroy(dpt_comm* c) {
    if (c == nullptr) return 0;
    cudaFree(c->ctx.epoch);
    cudaFreeHost(c->err_host);
    delete c;
    return 0;
}

// 0 = healthy; otherwise bit31 set, bits 24..30 = peer that never arrived, low 24 bits = epoch.
extern "C" unsigned dpt_comm_error(const dpt_comm* c) { return c ? *c->err_host : 0u; }

// Debug aid: copies the per-CTA barrier epoch counters (up to kMaxCommCtas words) to host memory
// after synchronising the device.  Every rank launches the same collectives with the same CTA
// counts, so after a completed step the counters must be identical on all ranks; a difference
// means a rank skipped or repeated a collective.  Returns the number of words written.
extern "C" int dpt_comm_epochs(const dpt_comm* c, unsigned* out, int n) {
    if (c == nullptr || out == nullptr || n <= 0) return DPT_ERR_ARG;
    if (n > kMaxCommCtas) n = kMaxCommCtas;
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return -static_cast<int>(e) - 100;
    e = cudaMemcpy(out, c->ctx.epoch, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return -static_cast<int>(e) - 100;
    return n;
}

static int clamp_ctas(int ctas, long long n4, int world) {
    if (ctas <= 0) ctas = 32;
    if (ctas > kMaxCommCtas) ctas = kMaxCommCtas;
    long long need = (n4 / (world > 0 ? world : 1) + kCommThreads - 1) / kCommThreads;
    if (need < 1) need = 1;
    if (ctas > need) ctas = static_cast<int>(need);
    return ctas;
}

static void fill_epilogue(Epilogue* ep, int world, float scale, const float* inv_scale,
                          void* const* found_inf) {
    memset(ep, 0, sizeof(*ep));
    ep->shadow_f16 = g_comm_shadow_f16;
    ep->skip_closing_barrier = g_skip_closing_next;
    g_skip_closing_next = 0;
    ep->scale = scale;
    ep->inv_scale = inv_scale;
    if (found_inf != nullptr)
        for (int r = 0; r < world; ++r) ep->found_inf[r] = static_cast<float*>(found_inf[r]);
}

template <int W>
static void launch_twoshot(const dpt_comm* c, const Peers& g, long long n4, const Epilogue& ep,
                           int ctas, cudaStream_t st) {
    allreduce_twoshot_kernel<W><<<ctas, kCommThreads, 0, st>>>(c->ctx, g, n4, ep);
}
template <int W>
static void launch_oneshot(const dpt_comm* c, const Peers& g, float* out, long long n4,
                           const Epilogue& ep, int ctas, cudaStream_t st) {
    allreduce_oneshot_kernel<W><<<ctas, kCommThreads, 0, st>>>(c->ctx, g, out, n4, ep);
}

// In-place fused all-reduce of n floats (n % 4 == 0, 16-byte aligned) living at bufs[r] on rank r.
//   algo: 0 auto, 1 one-shot (needs `scratch_out`... see dpt_allreduce_oneshot), 2 two-shot
//   scale: static multiplier (1/world); inv_scale: optional device scalar; found_inf[r]: optional
//   sgd: optional fused update (param[r], shadow[r] for every rank; local momentum).
static int allreduce_fused_impl(dpt_comm* c, void* const* bufs, long long n, float scale,
                                const float* inv_scale, void* const* found_inf,
                                void* const* param, void* const* shadow, float* momentum,
                                void* param_mc, void* shadow_mc, float lr, float mu, float wd,
                                int nesterov, int ctas, void* stream) {
    if (n % 4 != 0) return DPT_ERR_SHAPE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int world = c->ctx.world;
    const long long n4 = n / 4;
    Epilogue ep;
    fill_epilogue(&ep, world, scale, inv_scale, found_inf);
    if (param != nullptr) {
        ep.fused_sgd = 1;
        for (int r = 0; r < world; ++r) {
            ep.param[r] = static_cast<float*>(param[r]);
            ep.shadow[r] = static_cast<__nv_bfloat16*>(shadow[r]);
        }
        ep.momentum = momentum;
        ep.lr = lr;
        ep.mu = mu;
        ep.wd = wd;
        ep.nesterov = nesterov;
        if (param_mc != nullptr && shadow_mc != nullptr) {
            ep.param_mc = static_cast<float*>(param_mc);
            ep.shadow_mc = static_cast<__nv_bfloat16*>(shadow_mc);
        }
    }
    if (world == 1) {
        if (ep.fused_sgd) return DPT_ERR_ARG;  // single GPU: use dpt_sgd_flat (same math)
        long long b = (n4 + 255) / 256;
        if (b > 148 * 4) b = 148 * 4;
        if (b < 1) b = 1;
        local_epilogue_kernel<<<static_cast<int>(b), 256, 0, st>>>(static_cast<float*>(bufs[0]), n4,
                                                                   ep);
        return static_cast<int>(cudaGetLastError());
    }
    Peers g;
    memset(&g, 0, sizeof(g));
    for (int r = 0; r < world; ++r) g.p[r] = bufs[r];
    ctas = clamp_ctas(ctas, n4, world);
    switch (world) {
        case 2: launch_twoshot<2>(c, g, n4, ep, ctas, st); break;
        case 3: launch_twoshot<3>(c, g, n4, ep, ctas, st); break;
        case 4: launch_twoshot<4>(c, g, n4, ep, ctas, st); break;
        case 5: launch_twoshot<5>(c, g, n4, ep, ctas, st); break;
        case 6: launch_twoshot<6>(c, g, n4, ep, ctas, st); break;
        case 7: launch_twoshot<7>(c, g, n4, ep, ctas, st); break;
        case 8: launch_twoshot<8>(c, g, n4, ep, ctas, st); break;
        default: return DPT_ERR_ARG;
    }
    return static_cast<int>(cudaGetLastError());
}

extern "C" int dpt_allreduce_fused(dpt_comm* c, void* const* bufs, long long n, float scale,
                                   const float* inv_scale, void* const* found_inf,
                                   void* const* param, void* const* shadow, float* momentum,
                                   float lr, float mu, float wd, int nesterov, int ctas,
                                   void* stream) {
    return allreduce_fused_impl(c, bufs, n, scale, inv_scale, found_inf, param, shadow, momentum,
                                nullptr, nullptr, lr, mu, wd, nesterov, ctas, stream);
}

// EXPERIMENTAL (not yet run on hardware): dpt_allreduce_fused with the SGD update fused, where
// the owner publishes the new weights / 16-bit shadow with one NVSwitch multicast store each
// (`param_mc` / `shadow_mc` = this rank's multicast mappings of the same arenas at the bucket's
// offset) instead of one store per peer.  The reduce phase still pulls the peers' gradients P2P.
extern "C" int dpt_allreduce_fused_mcpush(dpt_comm* c, void* const* bufs, long long n, float scale,
                                          const float* inv_scale, void* const* found_inf,
                                          void* const* param, void* const* shadow,
                                          float* momentum, void* param_mc, void* shadow_mc,
                                          float lr, float mu, float wd, int nesterov, int ctas,
                                          void* stream) {
    if (param == nullptr || param_mc == nullptr || shadow_mc == nullptr) return DPT_ERR_ARG;
    return allreduce_fused_impl(c, bufs, n, scale, inv_scale, found_inf, param, shadow, momentum,
                                param_mc, shadow_mc, lr, mu, wd, nesterov, ctas, stream);
}

// One-shot variant: out (local, n floats) = scale * sum_r in[r].  `in` buffers may be reused by
// their owners as soon as the call completes on every rank.
extern "C" int dpt_allreduce_oneshot(dpt_comm* c, void* const* in, float* out, long long n,
                                     float scale, const float* inv_scale, void* const* found_inf,
                                     int ctas, void* stream) {
    if (n % 4 != 0) return DPT_ERR_SHAPE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int world = c->ctx.world;
    const long long n4 = n / 4;
    Epilogue ep;
    fill_epilogue(&ep, world, scale, inv_scale, found_inf);
    Peers g;
    memset(&g, 0, sizeof(g));
    for (int r = 0; r < world; ++r) g.p[r] = in[r];
    ctas = clamp_ctas(ctas, n4 * world, world);
    switch (world) {
        case 1: launch_oneshot<1>(c, g, out, n4, ep, ctas, st); break;
        case 2: launch_oneshot<2>(c, g, out, n4, ep, ctas, st); break;
        case 3: launch_oneshot<3>(c, g, out, n4, ep, ctas, st); break;
        case 4: launch_oneshot<4>(c, g, out, n4, ep, ctas, st); break;
        case 5: launch_oneshot<5>(c, g, out, n4, ep, ctas, st); break;
        case 6: launch_oneshot<6>(c, g, out, n4, ep, ctas, st); break;
        case 7: launch_oneshot<7>(c, g, out, n4, ep, ctas, st); break;
        case 8: launch_oneshot<8>(c, g, out, n4, ep, ctas, st); break;
        default: return DPT_ERR_ARG;
    }
    return static_cast<int>(cudaGetLastError());
}

// One-shot push all-reduce (small buckets).  `grads`: this rank's n floats (in place when the SGD
// update is not fused); `staging[r]`: rank r's staging area of 2 * world * cap floats (symmetric
// memory, cap % 4 == 0, n <= cap); `found_inf`, `param`, `shadow` are LOCAL pointers (the update
// is replicated, nothing is published); param == NULL selects the F1 epilogue.
extern "C" int dpt_allreduce_oneshot_push(dpt_comm* c, float* grads, void* const* staging,
                                          long long cap, long long n, float scale,
                                          const float* inv_scale, float* found_inf, float* param,
                                          void* shadow, float* momentum, float lr, float mu,
                                          float wd, int nesterov, int ctas, void* stream) {
    if (n % 4 != 0 || cap % 4 != 0) return DPT_ERR_SHAPE;
    if (n > cap) return DPT_ERR_TOO_LARGE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int world = c->ctx.world;
    if (world < 2) return DPT_ERR_ARG;
    const long long n4 = n / 4;
    Epilogue ep;
    memset(&ep, 0, sizeof(ep));
    ep.shadow_f16 = g_comm_shadow_f16;
    ep.scale = scale;
    ep.inv_scale = inv_scale;
    ep.found_inf[c->ctx.rank] = found_inf;
    if (param != nullptr) {
        ep.fused_sgd = 1;
        ep.param[c->ctx.rank] = param;
        ep.shadow[c->ctx.rank] = static_cast<__nv_bfloat16*>(shadow);
        ep.momentum = momentum;
        ep.lr = lr;
        ep.mu = mu;
        ep.wd = wd;
        ep.nesterov = nesterov;
    }
    Peers sg;
    memset(&sg, 0, sizeof(sg));
    for (int r = 0; r < world; ++r) sg.p[r] = staging[r];
    if (ctas <= 0) ctas = 8;
    ctas = clamp_ctas(ctas, n4, 1);
    float* g = grads;
    const long long cap4 = cap / 4;
#define DPT_OS_PUSH(W)                                                                          \
    case W:                                                                                     \
        allreduce_oneshot_push_kernel<W><<<ctas, kCommThreads, 0, st>>>(c->ctx, g, sg, cap4, n4, ep); \
        break;
    switch (world) {
        DPT_OS_PUSH(2) DPT_OS_PUSH(3) DPT_OS_PUSH(4) DPT_OS_PUSH(5) DPT_OS_PUSH(6) DPT_OS_PUSH(7)
        DPT_OS_PUSH(8)
        default: return DPT_ERR_ARG;
    }
#undef DPT_OS_PUSH
    return static_cast<int>(cudaGetLastError());
}

// The next dpt_allreduce_fused* / dpt_allreduce_nvls* launch from this thread leaves without its
// closing barrier (see Epilogue::skip_closing_barrier).  Only for a collective that is followed,
// on the same stream and on every rank, by another collective of this communicator before
// anything consumes its result; the gradient reducer uses it for all buckets but the last.
extern "C" int dpt_comm_skip_closing_barrier_next(void) {
    g_skip_closing_next = 1;
    return 0;
}

// Timeline buffer for the step-trace tool: `buf` = device memory of (1 + records) * 32 bytes that
// the caller zeroed; every collective kernel launched afterwards appends one record (see
// TraceRec).  NULL switches tracing off.  Kernels already captured in a CUDA graph keep the
// setting they were captured with.
extern "C" int dpt_comm_set_trace(dpt_comm* c, void* buf, int records) {
    if (c == nullptr) return DPT_ERR_ARG;
    c->ctx.trace = static_cast<unsigned long long*>(buf);
    c->ctx.trace_cap = buf != nullptr && records > 0 ? static_cast<uint32_t>(records) : 0u;
    return 0;
}

// A single-thread kernel that appends a {kind = 255 | tag, time} record to a trace buffer from
// any stream: marks phase boundaries of the compute streams on the same clock as the collectives.
__global__ void trace_stamp_kernel(unsigned long long* trace, uint32_t cap, uint32_t tag) {
    const unsigned long long i = atomicAdd(trace, 1ULL);
    if (i >= cap) return;
    unsigned long long* slot = trace + 4 + 4 * i;
    slot[0] = (255ULL << 56) | tag;
    slot[1] = dpt::globaltimer_ns();
    slot[2] = 0;
    slot[3] = 0;
}
extern "C" int dpt_trace_stamp(void* buf, int records, int tag, void* stream) {
    if (buf == nullptr || records <= 0) return DPT_ERR_ARG;
    trace_stamp_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<unsigned long long*>(buf), static_cast<uint32_t>(records),
        static_cast<uint32_t>(tag));
    return static_cast<int>(cudaGetLastError());
}

// NVLS in-place fused all-reduce: `mc` is this rank's mapping of the multicast object covering
// every rank's buffer (plus the same element offset on every rank).
extern "C" int dpt_allreduce_nvls_sgd(dpt_comm* c, void* mc, long long n, float scale,
                                      const float* inv_scale, void* const* found_inf,
                                      float* param_local, float* momentum, void* param_mc,
                                      void* shadow_mc, float lr, float mu, float wd, int nesterov,
                                      int ctas, void* stream);

extern "C" int dpt_allreduce_nvls(dpt_comm* c, void* mc, long long n, float scale,
                                  const float* inv_scale, void* const* found_inf, int ctas,
                                  void* stream) {
    return dpt_allreduce_nvls_sgd(c, mc, n, scale, inv_scale, found_inf, nullptr, nullptr, nullptr,
                                  nullptr, 0.f, 0.f, 0.f, 0, ctas, stream);
}

// NVLS all-reduce with the SGD-momentum update fused in (param_local != NULL): rank r reduces
// slice r in the switch, updates its slice of the local fp32 masters + momentum and publishes
// the new weights / bf16 shadow to every rank through the multicast mappings.
extern "C" int dpt_allreduce_nvls_sgd(dpt_comm* c, void* mc, long long n, float scale,
                                      const float* inv_scale, void* const* found_inf,
                                      float* param_local, float* momentum, void* param_mc,
                                      void* shadow_mc, float lr, float mu, float wd, int nesterov,
                                      int ctas, void* stream) {
    if (n % 4 != 0) return DPT_ERR_SHAPE;
    const int world = c->ctx.world;
    if (world < 2) return DPT_ERR_ARG;
    const long long n4 = n / 4;
    Epilogue ep;
    fill_epilogue(&ep, world, scale, inv_scale, found_inf);
    if (param_local != nullptr) {
        ep.fused_sgd = 1;
        ep.param[c->ctx.rank] = param_local;
        ep.momentum = momentum;
        ep.param_mc = static_cast<float*>(param_mc);
        ep.shadow_mc = static_cast<__nv_bfloat16*>(shadow_mc);
        ep.lr = lr;
        ep.mu = mu;
        ep.wd = wd;
        ep.nesterov = nesterov;
    }
    if (ctas <= 0) ctas = 16;
    ctas = clamp_ctas(ctas, n4, world);
    allreduce_nvls_kernel<<<ctas, kCommThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        c->ctx, static_cast<float*>(mc), n4, ep);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int dpt_broadcast(dpt_comm* c, void* const* bufs, long long bytes, int src, int ctas,
                             void* stream) {
    if (bytes % 16 != 0) return DPT_ERR_SHAPE;
    const int world = c->ctx.world;
    if (world == 1) return 0;
    Peers g;
    memset(&g, 0, sizeof(g));
    for (int r = 0; r < world; ++r) g.p[r] = bufs[r];
    ctas = clamp_ctas(ctas, bytes / 16, 1);
    broadcast_kernel<<<ctas, kCommThreads, 0, static_cast<cudaStream_t>(stream)>>>(c->ctx, g,
                                                                                   bytes / 16, src);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int dpt_barrier(dpt_comm* c, void* stream) {
    if (c->ctx.world == 1) return 0;
    barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(c->ctx);
    return static_cast<int>(cudaGetLastError());
}
