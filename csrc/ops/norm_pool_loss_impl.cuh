// Element-type generic implementation of the bandwidth-bound model kernels; included twice by
// norm_pool_loss.cu (bf16 and fp16 instantiations in separate namespaces).  See that file.
// Requires: DPT_NS (namespace name) and the lowp / lowp2 types + conversion helpers in scope.
namespace DPT_NS {
using namespace dpt;


// 8 activations as whole 16-byte registers: every global / shared access of a vector is ONE
// 128-bit instruction (two for fp32 activations).  (A member-wise `lowp2 v[4]` definition made
// nvcc emit four 32-bit LDG / STG per vector — 4x the memory instructions and 4-byte
// partial-sector stores; measured on B200: ResNet-18 step 0.734 -> 0.687 ms with this layout.)
struct alignas(16) bf16x8 {
    uint4 q[sizeof(lowp2) / 4];   // 1 x 16 B for 16-bit element types, 2 x 16 B for fp32
};

__device__ __forceinline__ void unpack8(const bf16x8& b, float (&f)[8]) {
    if constexpr (sizeof(lowp2) == 4) {
        const uint4 q = b.q[0];   // one 128-bit load
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            lowp2 h;
            memcpy(&h, &w[i], sizeof(h));
            float2 t = lp2_to_float2(h);
            f[2 * i] = t.x;
            f[2 * i + 1] = t.y;
        }
    } else {
        const uint4 q0 = b.q[0], q1 = b.q[sizeof(lowp2) / 4 - 1];
        f[0] = __uint_as_float(q0.x); f[1] = __uint_as_float(q0.y);
        f[2] = __uint_as_float(q0.z); f[3] = __uint_as_float(q0.w);
        f[4] = __uint_as_float(q1.x); f[5] = __uint_as_float(q1.y);
        f[6] = __uint_as_float(q1.z); f[7] = __uint_as_float(q1.w);
    }
}
__device__ __forceinline__ bf16x8 pack8(const float (&f)[8]) {
    bf16x8 b;
    if constexpr (sizeof(lowp2) == 4) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const lowp2 h = floats_to_lp2(f[2 * i], f[2 * i + 1]);
            memcpy(&w[i], &h, sizeof(h));
        }
        b.q[0] = make_uint4(w[0], w[1], w[2], w[3]);
    } else {
        b.q[0] = make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                            __float_as_uint(f[3]));
        b.q[sizeof(lowp2) / 4 - 1] = make_uint4(__float_as_uint(f[4]), __float_as_uint(f[5]),
                                                __float_as_uint(f[6]), __float_as_uint(f[7]));
    }
    return b;
}

constexpr int kEwThreads = 256;

static inline int ew_grid(long long nvec, int cvec) {
    // Threads are bound to a fixed channel group: total thread count must be a multiple of
    // cvec (= C/8).  256 % cvec == 0 is checked by the callers.
    (void)cvec;
    long long blocks = (nvec + kEwThreads - 1) / kEwThreads;
    const long long cap = 148LL * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}


// Block-level reduction for kernels whose threads are bound to a channel group
// (thread t handles channels 8*(t % cvec) .. +7; cvec is a power of two dividing 256).
// Lanes of a warp that share a group are combined with shuffles, the <= 8 remaining partials
// per group through shared memory.  On return threads t < cvec hold the block totals.
__device__ __forceinline__ void block_reduce_cg(float (&s1)[8], float (&s2)[8], int cvec,
                                                float (*sh)[kEwThreads / 32][32][8 + 1]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int off = 16; off >= cvec; off >>= 1) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            s1[k] += __shfl_xor_sync(0xffffffffu, s1[k], off);
            s2[k] += __shfl_xor_sync(0xffffffffu, s2[k], off);
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        sh[0][warp][lane][k] = s1[k];
        sh[1][warp][lane][k] = s2[k];
    }
    __syncthreads();
    if (threadIdx.x < cvec) {
#pragma unroll
        for (int k = 0; k < 8; ++k) s1[k] = s2[k] = 0.f;
        if (cvec <= 32) {
            // group g lives in lane g of every warp
            for (int w = 0; w < kEwThreads / 32; ++w) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    s1[k] += sh[0][w][threadIdx.x][k];
                    s2[k] += sh[1][w][threadIdx.x][k];
                }
            }
        } else {
            // group g lives in threads g, g + cvec, ...
            for (int t = threadIdx.x; t < kEwThreads; t += cvec) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    s1[k] += sh[0][t >> 5][t & 31][k];
                    s2[k] += sh[1][t >> 5][t & 31][k];
                }
            }
        }
    }
}

// Per-channel reductions finish with atomics on a handful of cache lines (2*C floats), which the
// L2 serialises per line: keep the CTA count at one per SM and use 16-byte vector RED.
// That rule is for the small tensors of the CIFAR models, where the kernel is a few microseconds
// long and the atomic tail would dominate.  Long streaming reductions get up to 4 CTAs per SM
// (all co-resident: <= 64 registers, 18 KB smem); the 4x REDs are amortised over >= 64 loop trips.
// (The ResNet-50-sized tensors go through the bulk-copy kernels further down instead.)
static inline int reduce_grid_resident(long long nvec) {
    // for kernels with a grid-wide spin barrier: never more than one CTA per SM
    long long blocks = (nvec + kEwThreads - 1) / kEwThreads;
    if (blocks > 148) blocks = 148;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}
static inline int reduce_grid(long long nvec) {
    long long blocks = (nvec + kEwThreads - 1) / kEwThreads;
    const long long sms = 148;
    if (blocks >= sms * 64) return static_cast<int>(sms * 4);
    if (blocks >= sms * 16) return static_cast<int>(sms * 2);
    if (blocks > sms) blocks = sms;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}
__device__ __forceinline__ void red_add8(float* dst, const float (&v)[8]) {
    red_add_f32x4(dst, v[0], v[1], v[2], v[3]);
    red_add_f32x4(dst + 4, v[4], v[5], v[6], v[7]);
}

// ------------------------------------------------------------------------------------ BN stats
// sum[c] += sum_m x[m,c]; sq[c] += sum_m x[m,c]^2   (buffers zeroed by the caller)
__global__ void __launch_bounds__(kEwThreads) bn_stats_kernel(const bf16x8* __restrict__ x,
                                                              float* __restrict__ sum,
                                                              float* __restrict__ sq, long long nvec,
                                                              int cvec) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sh[2][kEwThreads / 32][32][8 + 1];
    const long long stride = static_cast<long long>(gridDim.x) * kEwThreads;
    long long i = static_cast<long long>(blockIdx.x) * kEwThreads + threadIdx.x;
    float s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (; i < nvec; i += stride) {
        float f[8];
        unpack8(x[i], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            s1[k] += f[k];
            s2[k] += f[k] * f[k];
        }
    }
    block_reduce_cg(s1, s2, cvec, sh);
    if (threadIdx.x < cvec) {
        const int cg = threadIdx.x;
        red_add8(sum + cg * 8, s1);
        red_add8(sq + cg * 8, s2);
        // The consumer of these sums is launched with the programmatic attribute and reads them
        // right after griddepcontrol.wait; REDs are fire-and-forget, and with hundreds of CTAs
        // queueing on a few hot cache lines the last ones can still be in flight when the grid
        // counts as complete (same effect as measured for split-K accumulators).  Wait for them.
        __threadfence();
    }
}

// ------------------------------------------------------------------------------------ BN fwd
// y = relu?( (x - mean) * invstd * gamma + beta (+ residual) ), batch statistics from sum/sq.
// Block 0 also writes save_mean / save_invstd and updates the running statistics.
__global__ void __launch_bounds__(kEwThreads)
bn_fwd_kernel(const bf16x8* __restrict__ x, const float* __restrict__ sum,
              const float* __restrict__ sq, const float* __restrict__ gamma,
              const float* __restrict__ beta, const bf16x8* __restrict__ residual,
              bf16x8* __restrict__ y, float* __restrict__ save_mean,
              float* __restrict__ save_invstd, float* __restrict__ running_mean,
              float* __restrict__ running_var, long long* __restrict__ num_batches_tracked,
              long long nvec, int cvec, float inv_m, float unbias, float momentum, float eps,
              int relu) {
    pdl_launch_dependents();
    pdl_wait();
    const long long stride = static_cast<long long>(gridDim.x) * kEwThreads;
    long long i = static_cast<long long>(blockIdx.x) * kEwThreads + threadIdx.x;
    const int cg = static_cast<int>(i % cvec);
    float sc[8], sf[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = cg * 8 + k;
        const float mean = sum[c] * inv_m;
        const float var = fmaxf(sq[c] * inv_m - mean * mean, 0.f);
        const float invstd = rsqrtf(var + eps);
        sc[k] = gamma[c] * invstd;
        sf[k] = beta[c] - mean * sc[k];
        if (blockIdx.x == 0 && threadIdx.x < cvec) {
            save_mean[c] = mean;
            save_invstd[c] = invstd;
            if (running_mean != nullptr) {
                running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * var * unbias;
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && num_batches_tracked != nullptr)
        *num_batches_tracked += 1;
    for (; i < nvec; i += stride) {
        float f[8];
        unpack8(x[i], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], sf[k]);
        if (residual != nullptr) {
            float r[8];
            unpack8(residual[i], r);
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] += r[k];
        }
        if (relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
        }
        y[i] = pack8(f);
    }
}

// Eval mode: y = relu?(x * scale + shift (+ residual)) with scale/shift folded from running stats.
__global__ void __launch_bounds__(kEwThreads)
bn_eval_kernel(const bf16x8* __restrict__ x, const float* __restrict__ gamma,
               const float* __restrict__ beta, const float* __restrict__ running_mean,
               const float* __restrict__ running_var, const bf16x8* __restrict__ residual,
               bf16x8* __restrict__ y, long long nvec, int cvec, float eps, int relu) {
    pdl_launch_dependents();
    pdl_wait();
    const long long stride = static_cast<long long>(gridDim.x) * kEwThreads;
    long long i = static_cast<long long>(blockIdx.x) * kEwThreads + threadIdx.x;
    const int cg = static_cast<int>(i % cvec);
    float sc[8], sf[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = cg * 8 + k;
        sc[k] = gamma[c] * rsqrtf(running_var[c] + eps);
        sf[k] = beta[c] - running_mean[c] * sc[k];
    }
    for (; i < nvec; i += stride) {
        float f[8];
        unpack8(x[i], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], sf[k]);
        if (residual != nullptr) {
            float r[8];
            unpack8(residual[i], r);
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] += r[k];
        }
        if (relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
        }
        y[i] = pack8(f);
    }
}

// Column-strip BatchNorm training forward for small tensors whose producer is a split-K conv with
// fp32 atomic output: one CTA owns 8 channels for ALL rows, so statistics and normalisation
// happen in a single launch.  Writes y (bf16) and a bf16 copy of x for the backward pass.
__global__ void __launch_bounds__(kEwThreads)
bn_fwd_small_f32_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                        const float* __restrict__ beta, const bf16x8* __restrict__ residual,
                        bf16x8* __restrict__ y, bf16x8* __restrict__ x_bf16,
                        float* __restrict__ save_mean, float* __restrict__ save_invstd,
                        float* __restrict__ running_mean, float* __restrict__ running_var,
                        long long* __restrict__ num_batches_tracked, int M, int cvec, float inv_m,
                        float unbias, float momentum, float eps, int relu) {
    __shared__ float sh[2][kEwThreads / 32][8];
    __shared__ float s_scale[8], s_shift[8];
    const int cg = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    // rows are independent: issue 4 rows' loads before consuming them (latency, not bandwidth,
    // bounds this kernel)
    constexpr int kRows = 4;
    for (int m0 = threadIdx.x; m0 < M; m0 += kEwThreads * kRows) {
        float4 a[kRows], b[kRows];
#pragma unroll
        for (int j = 0; j < kRows; ++j) {
            const int m = m0 + j * kEwThreads;
            if (m < M) {
                const float4* p4 = reinterpret_cast<const float4*>(
                    x + (static_cast<long long>(m) * cvec + cg) * 8);
                a[j] = p4[0];
                b[j] = p4[1];
            } else {
                a[j] = b[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
#pragma unroll
        for (int j = 0; j < kRows; ++j) {
            const float f[8] = {a[j].x, a[j].y, a[j].z, a[j].w, b[j].x, b[j].y, b[j].z, b[j].w};
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                s1[k] += f[k];
                s2[k] += f[k] * f[k];
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s1[k] += __shfl_xor_sync(0xffffffffu, s1[k], o);
            s2[k] += __shfl_xor_sync(0xffffffffu, s2[k], o);
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            sh[0][warp][k] = s1[k];
            sh[1][warp][k] = s2[k];
        }
    }
    __syncthreads();
    if (threadIdx.x < 8) {
        const int k = threadIdx.x, c = cg * 8 + k;
        float t1 = 0.f, t2 = 0.f;
        for (int w = 0; w < kEwThreads / 32; ++w) {
            t1 += sh[0][w][k];
            t2 += sh[1][w][k];
        }
        const float mean = t1 * inv_m;
        const float var = fmaxf(t2 * inv_m - mean * mean, 0.f);
        const float invstd = rsqrtf(var + eps);
        const float sc = gamma[c] * invstd;
        s_scale[k] = sc;
        s_shift[k] = beta[c] - mean * sc;
        save_mean[c] = mean;
        save_invstd[c] = invstd;
        if (running_mean != nullptr) {
            running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * var * unbias;
        }
        if (cg == 0 && k == 0 && num_batches_tracked != nullptr) *num_batches_tracked += 1;
    }
    __syncthreads();
    float sc[8], sf[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        sc[k] = s_scale[k];
        sf[k] = s_shift[k];
    }
    for (int m0 = threadIdx.x; m0 < M; m0 += kEwThreads * kRows) {
        float4 a[kRows], b[kRows];
        bf16x8 rres[kRows];
#pragma unroll
        for (int j = 0; j < kRows; ++j) {
            const int m = m0 + j * kEwThreads;
            if (m < M) {
                const long long v = static_cast<long long>(m) * cvec + cg;
                const float4* p4 = reinterpret_cast<const float4*>(x + v * 8);
                a[j] = p4[0];
                b[j] = p4[1];
                if (residual != nullptr) rres[j] = residual[v];
            }
        }
#pragma unroll
        for (int j = 0; j < kRows; ++j) {
            const int m = m0 + j * kEwThreads;
            if (m >= M) continue;
            const long long v = static_cast<long long>(m) * cvec + cg;
            float f[8] = {a[j].x, a[j].y, a[j].z, a[j].w, b[j].x, b[j].y, b[j].z, b[j].w};
            x_bf16[v] = pack8(f);
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], sf[k]);
            if (residual != nullptr) {
                float r[8];
                unpack8(rres[j], r);
#pragma unroll
                for (int k = 0; k < 8; ++k) f[k] += r[k];
            }
            if (relu) {
#pragma unroll
                for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
            }
            y[v] = pack8(f);
        }
    }
}

// ------------------------------------------------------------------------------------ BN bwd
// Pass 1: g = (dy1 + dy2) * [y > 0];  dbeta[c] += sum g;  dgamma[c] += sum g * xhat.
// Writes g (bf16) for pass 2 and for the residual branch.
template <bool DY_F32>
__global__ void __launch_bounds__(kEwThreads)
bn_bwd_reduce_kernel(const void* __restrict__ dy1_raw, const bf16x8* __restrict__ dy2,
                     const bf16x8* __restrict__ y, const bf16x8* __restrict__ x,
                     const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
                     bf16x8* __restrict__ g_out, float* __restrict__ dgamma,
                     float* __restrict__ dbeta, long long nvec, int cvec) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sh[2][kEwThreads / 32][32][8 + 1];
    const long long stride = static_cast<long long>(gridDim.x) * kEwThreads;
    long long i = static_cast<long long>(blockIdx.x) * kEwThreads + threadIdx.x;
    const int cg = static_cast<int>(i % cvec);
    float mean[8], istd[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        mean[k] = save_mean[cg * 8 + k];
        istd[k] = save_invstd[cg * 8 + k];
    }
    float s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (; i < nvec; i += stride) {
        float g[8], xv[8];
        if (DY_F32) {
            const float4* p4 = reinterpret_cast<const float4*>(dy1_raw) + 2 * i;
            const float4 a = p4[0], b = p4[1];
            g[0] = a.x; g[1] = a.y; g[2] = a.z; g[3] = a.w;
            g[4] = b.x; g[5] = b.y; g[6] = b.z; g[7] = b.w;
        } else {
            unpack8(reinterpret_cast<const bf16x8*>(dy1_raw)[i], g);
        }
        if (dy2 != nullptr) {
            float t[8];
            unpack8(dy2[i], t);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] += t[k];
        }
        if (y != nullptr) {
            float yv[8];
            unpack8(y[i], yv);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] = yv[k] > 0.f ? g[k] : 0.f;
        }
        if (g_out != nullptr) {
            const bf16x8 gb = pack8(g);
            g_out[i] = gb;
            unpack8(gb, g);  // reduce exactly what pass 2 will read
        }
        unpack8(x[i], xv);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            s1[k] += g[k];
            s2[k] += g[k] * (xv[k] - mean[k]) * istd[k];
        }
    }
    block_reduce_cg(s1, s2, cvec, sh);
    if (threadIdx.x < cvec) {
        red_add8(dbeta + cg * 8, s1);
        red_add8(dgamma + cg * 8, s2);
        __threadfence();   // see bn_stats_kernel: the PDL-chained apply kernel reads these first
    }
}

// Pass 2: dx = gamma * invstd * (g - dbeta/M - xhat * dgamma/M)
__global__ void __launch_bounds__(kEwThreads)
bn_bwd_apply_kernel(const bf16x8* __restrict__ g, const bf16x8* __restrict__ x,
                    const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
                    const float* __restrict__ gamma, const float* __restrict__ dgamma,
                    const float* __restrict__ dbeta, bf16x8* __restrict__ dx, long long nvec,
                    int cvec, float inv_m, float grad_unscale) {
    pdl_launch_dependents();
    pdl_wait();
    const long long stride = static_cast<long long>(gridDim.x) * kEwThreads;
    long long i = static_cast<long long>(blockIdx.x) * kEwThreads + threadIdx.x;
    const int cg = static_cast<int>(i % cvec);
    float mean[8], istd[8], a[8], b[8], c2[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = cg * 8 + k;
        mean[k] = save_mean[c];
        istd[k] = save_invstd[c];
        a[k] = gamma[c] * istd[k];
        b[k] = dbeta[c] * inv_m * grad_unscale;
        c2[k] = dgamma[c] * inv_m * grad_unscale;
    }
    for (; i < nvec; i += stride) {
        float gv[8], xv[8], o[8];
        unpack8(g[i], gv);
        unpack8(x[i], xv);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float xh = (xv[k] - mean[k]) * istd[k];
            o[k] = a[k] * (gv[k] - b[k] - xh * c2[k]);
        }
        dx[i] = pack8(o);
    }
}

// ------------------------------------------------------------------ bulk-copy streaming variants
// For tensors of hundreds of megabytes (ResNet-50 @224) the register-file-bound kernels above top
// out at ~3 TB/s: every byte in flight needs a destination register, and 64 registers x 1024
// threads leave room for ~48 KB of outstanding loads per SM.  These variants move the operands
// with the TMA unit instead: one thread issues 4 KB (8 KB in FP32 mode) 1-D bulk copies (cp.async.bulk, completion on
// an mbarrier) into a ring of kBulkStages shared-memory slots per operand, so 48-64 KB per CTA
// (2-3 CTAs per SM) are in flight regardless of registers; the 256 threads then read their 16
// bytes from shared memory, `__syncthreads()` hands the slot back to the producer thread, and the
// math overlaps the copies already in flight.  A chunk is 256 consecutive vectors, so the
// thread <-> channel-group binding (256 % cvec == 0) is the same as in the kernels above.
constexpr int kBulkStages = 4;
constexpr int kBulkStagesReduce = 4;   // 4 operands: 64 KB ring + 18 KB reduction scratch -> 2 CTAs / SM
constexpr int kBulkVec = static_cast<int>(sizeof(bf16x8));   // 16 B (16-bit types) or 32 B (fp32)
constexpr int kBulkChunk = kEwThreads * kBulkVec;             // bytes per operand per stage

template <int NOPS, int STAGES, typename Body>
__device__ __forceinline__ void bulk_stream(const bf16x8* const (&src)[NOPS], long long nvec,
                                            uint8_t* ring, uint64_t* full, Body&& body) {
    const long long nchunks = (nvec + kEwThreads - 1) / kEwThreads;
    const long long first = blockIdx.x, step = gridDim.x;
    auto issue = [&](long long chunk, int stage) {
        const long long v0 = chunk * kEwThreads;
        const long long left = nvec - v0;
        const uint32_t bytes = static_cast<uint32_t>(left < kEwThreads ? left : kEwThreads) * kBulkVec;
        uint32_t total = 0;
#pragma unroll
        for (int o = 0; o < NOPS; ++o)
            if (src[o] != nullptr) total += bytes;
        mbar_expect_tx(&full[stage], total);
#pragma unroll
        for (int o = 0; o < NOPS; ++o)
            if (src[o] != nullptr)
                bulk_load_1d(ring + (stage * NOPS + o) * kBulkChunk, src[o] + v0, bytes, &full[stage]);
    };
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            const long long chunk = first + s * step;
            if (chunk < nchunks) issue(chunk, s);
        }
    }
    int stage = 0;
    uint32_t phase = 0;
    for (long long chunk = first; chunk < nchunks; chunk += step) {
        mbar_wait(&full[stage], phase);
        const long long idx = chunk * kEwThreads + threadIdx.x;
        const bool live = idx < nvec;
        bf16x8 v[NOPS];
        if (live) {
#pragma unroll
            for (int o = 0; o < NOPS; ++o)
                if (src[o] != nullptr)
                    v[o] = *reinterpret_cast<const bf16x8*>(ring + (stage * NOPS + o) * kBulkChunk +
                                                            threadIdx.x * kBulkVec);
        }
        __syncthreads();   // every thread has copied its vectors out: the slot can be refilled
        if (threadIdx.x == 0) {
            const long long next = chunk + static_cast<long long>(STAGES) * step;
            // Order the generic-proxy reads of this slot (all threads, before the barrier) ahead
            // of the async-proxy write that refills it.  Without the fence the refill was
            // observed to overtake the LDS.128 reads on B200 (wrong dx for the first chunks of a
            // CTA, reproducible on every run): __syncthreads() orders threads, not proxies.
            fence_proxy_async();
            if (next < nchunks) issue(next, stage);
        }
        if (live) body(idx, v);
        if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
        }
    }
}

__device__ __forceinline__ void bulk_ring_init(uint64_t* full, int stages = kBulkStages) {
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) mbar_init(&full[s], 1);
        fence_mbar_init();
    }
    __syncthreads();
}

static inline int bulk_grid(long long nvec, int ctas_per_sm) {
    long long chunks = (nvec + kEwThreads - 1) / kEwThreads;
    long long g = 148LL * ctas_per_sm;
    if (g > chunks) g = chunks;
    return static_cast<int>(g < 1 ? 1 : g);
}
// The bulk variants pay off once every CTA streams many chunks (ResNet-50 bs256: 30.5 vs 32.1
// ms/step).  DPT_BULK_EW is a debugging bit mask (1 forward, 2 backward reduce, 4 backward apply;
// 0 = grid-stride kernels everywhere); default all on.
// History: without the fence.proxy.async in bulk_stream the backward-apply kernel wrote, for its
// first chunks, the result of the chunk that REFILLS the slot (C = 256 shapes, every run) — the
// 128-bit shared loads were overtaken by the TMA write of the refill.
static inline bool bulk_worthwhile(long long nvec, int which = 7) {
    static int on = -1;
    if (on < 0) {
        const char* e = std::getenv("DPT_BULK_EW");   // bit mask: 1 forward, 2 backward reduce, 4 backward apply
        on = (e != nullptr) ? std::atoi(e) : 7;
    }
    return (on & which) && nvec >= 148LL * 2 * kEwThreads * 16;
}

// bn_fwd_kernel for large tensors (same arguments and results).
__global__ void __launch_bounds__(kEwThreads)
bn_fwd_bulk_kernel(const bf16x8* __restrict__ x, const float* __restrict__ sum,
                   const float* __restrict__ sq, const float* __restrict__ gamma,
                   const float* __restrict__ beta, const bf16x8* __restrict__ residual,
                   bf16x8* __restrict__ y, float* __restrict__ save_mean,
                   float* __restrict__ save_invstd, float* __restrict__ running_mean,
                   float* __restrict__ running_var, long long* __restrict__ num_batches_tracked,
                   long long nvec, int cvec, float inv_m, float unbias, float momentum, float eps,
                   int relu) {
    extern __shared__ __align__(128) uint8_t bulk_ring[];
    __shared__ __align__(8) uint64_t full[kBulkStages];
    bulk_ring_init(full);
    pdl_launch_dependents();
    pdl_wait();
    const int cg = static_cast<int>(threadIdx.x % cvec);
    float sc[8], sf[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = cg * 8 + k;
        const float mean = sum[c] * inv_m;
        const float var = fmaxf(sq[c] * inv_m - mean * mean, 0.f);
        const float invstd = rsqrtf(var + eps);
        sc[k] = gamma[c] * invstd;
        sf[k] = beta[c] - mean * sc[k];
        if (blockIdx.x == 0 && threadIdx.x < cvec) {
            save_mean[c] = mean;
            save_invstd[c] = invstd;
            if (running_mean != nullptr) {
                running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * var * unbias;
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && num_batches_tracked != nullptr)
        *num_batches_tracked += 1;
    const bool has_res = residual != nullptr;
    const bf16x8* const src[2] = {x, residual};
    bulk_stream<2, kBulkStages>(src, nvec, bulk_ring, full, [&](long long idx, const bf16x8 (&v)[2]) {
        float f[8];
        unpack8(v[0], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], sf[k]);
        if (has_res) {
            float r[8];
            unpack8(v[1], r);
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] += r[k];
        }
        if (relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
        }
        y[idx] = pack8(f);
    });
}

// bn_bwd_reduce_kernel<false> for large tensors (16-bit dy1).
__global__ void __launch_bounds__(kEwThreads)
bn_bwd_reduce_bulk_kernel(const bf16x8* __restrict__ dy1, const bf16x8* __restrict__ dy2,
                          const bf16x8* __restrict__ y, const bf16x8* __restrict__ x,
                          const float* __restrict__ save_mean,
                          const float* __restrict__ save_invstd, bf16x8* __restrict__ g_out,
                          float* __restrict__ dgamma, float* __restrict__ dbeta, long long nvec,
                          int cvec) {
    extern __shared__ __align__(128) uint8_t bulk_ring[];
    __shared__ __align__(8) uint64_t full[kBulkStages];
    __shared__ float sh[2][kEwThreads / 32][32][8 + 1];
    bulk_ring_init(full, kBulkStagesReduce);
    pdl_launch_dependents();
    pdl_wait();
    const int cg = static_cast<int>(threadIdx.x % cvec);
    float mean[8], istd[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        mean[k] = save_mean[cg * 8 + k];
        istd[k] = save_invstd[cg * 8 + k];
    }
    float s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const bool has_dy2 = dy2 != nullptr, has_y = y != nullptr;
    const bf16x8* const src[4] = {dy1, dy2, y, x};
    bulk_stream<4, kBulkStagesReduce>(src, nvec, bulk_ring, full, [&](long long idx, const bf16x8 (&v)[4]) {
        float g[8], xv[8];
        unpack8(v[0], g);
        if (has_dy2) {
            float t[8];
            unpack8(v[1], t);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] += t[k];
        }
        if (has_y) {
            float yv[8];
            unpack8(v[2], yv);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] = yv[k] > 0.f ? g[k] : 0.f;
        }
        if (g_out != nullptr) {
            const bf16x8 gb = pack8(g);
            g_out[idx] = gb;
            unpack8(gb, g);  // reduce exactly what pass 2 will read
        }
        unpack8(v[3], xv);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            s1[k] += g[k];
            s2[k] += g[k] * (xv[k] - mean[k]) * istd[k];
        }
    });
    block_reduce_cg(s1, s2, cvec, sh);
    if (threadIdx.x < cvec) {
        red_add8(dbeta + cg * 8, s1);
        red_add8(dgamma + cg * 8, s2);
        __threadfence();   // see bn_stats_kernel: the PDL-chained apply kernel reads these first
    }
}

// bn_bwd_apply_kernel for large tensors.
__global__ void __launch_bounds__(kEwThreads)
bn_bwd_apply_bulk_kernel(const bf16x8* __restrict__ g, const bf16x8* __restrict__ x,
                         const float* __restrict__ save_mean,
                         const float* __restrict__ save_invstd, const float* __restrict__ gamma,
                         const float* __restrict__ dgamma, const float* __restrict__ dbeta,
                         bf16x8* __restrict__ dx, long long nvec, int cvec, float inv_m,
                         float grad_unscale) {
    extern __shared__ __align__(128) uint8_t bulk_ring[];
    __shared__ __align__(8) uint64_t full[kBulkStages];
    bulk_ring_init(full);
    pdl_launch_dependents();
    pdl_wait();
    const int cg = static_cast<int>(threadIdx.x % cvec);
    float mean[8], istd[8], a[8], b[8], c2[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = cg * 8 + k;
        mean[k] = save_mean[c];
        istd[k] = save_invstd[c];
        a[k] = gamma[c] * istd[k];
        b[k] = dbeta[c] * inv_m * grad_unscale;
        c2[k] = dgamma[c] * inv_m * grad_unscale;
    }
    const bf16x8* const src[2] = {g, x};
    bulk_stream<2, kBulkStages>(src, nvec, bulk_ring, full, [&](long long idx, const bf16x8 (&v)[2]) {
        float gv[8], xv[8], o[8];
        unpack8(v[0], gv);
        unpack8(v[1], xv);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float xh = (xv[k] - mean[k]) * istd[k];
            o[k] = a[k] * (gv[k] - b[k] - xh * c2[k]);
        }
        dx[idx] = pack8(o);
    });
}

template <typename K>
static inline int bulk_smem_attr(K kern, int nops, int stages = kBulkStages) {
    // always opt in: static (reduction scratch) + dynamic shared memory may exceed the 48 KB
    // default even when the dynamic part alone does not
    const int smem = stages * nops * kBulkChunk;
    return static_cast<int>(
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
}

// Single-launch BatchNorm backward: pass 1 (masked-gradient sums) and pass 2 (dx) in one kernel
// separated by a grid-wide barrier.  The grid is at most one CTA per SM, so every CTA is
// resident and the spin barrier cannot deadlock.  `scratch` = 2*C floats + `counter` must be zero
// on entry (the engine clears them with the per-step statistics memset); the per-channel sums
// are added into dgamma / dbeta (gradient accumulation semantics) by CTA 0.
template <bool DY_F32>
__global__ void __launch_bounds__(kEwThreads)
bn_bwd_fused_kernel(const void* __restrict__ dy1_raw, const bf16x8* __restrict__ dy2,
                    const bf16x8* __restrict__ y, const bf16x8* __restrict__ x,
                    const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
                    const float* __restrict__ gamma, bf16x8* __restrict__ g_out,
                    float* __restrict__ dgamma, float* __restrict__ dbeta,
                    float* __restrict__ scratch, unsigned int* __restrict__ counter,
                    bf16x8* __restrict__ dx, long long nvec, int cvec, float inv_m) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sh[2][kEwThreads / 32][32][8 + 1];
    const long long stride = static_cast<long long>(gridDim.x) * kEwThreads;
    const long long i0 = static_cast<long long>(blockIdx.x) * kEwThreads + threadIdx.x;
    const int cg = static_cast<int>(i0 % cvec);
    const int C = cvec * 8;
    float mean[8], istd[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        mean[k] = save_mean[cg * 8 + k];
        istd[k] = save_invstd[cg * 8 + k];
    }
    float s1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, s2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (long long i = i0; i < nvec; i += stride) {
        float g[8], xv[8];
        if (DY_F32) {
            const float4* p4 = reinterpret_cast<const float4*>(dy1_raw) + 2 * i;
            const float4 a = p4[0], b = p4[1];
            g[0] = a.x; g[1] = a.y; g[2] = a.z; g[3] = a.w;
            g[4] = b.x; g[5] = b.y; g[6] = b.z; g[7] = b.w;
        } else {
            unpack8(reinterpret_cast<const bf16x8*>(dy1_raw)[i], g);
        }
        if (dy2 != nullptr) {
            float t[8];
            unpack8(dy2[i], t);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] += t[k];
        }
        if (y != nullptr) {
            float yv[8];
            unpack8(y[i], yv);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] = yv[k] > 0.f ? g[k] : 0.f;
        }
        const bf16x8 gb = pack8(g);
        g_out[i] = gb;
        unpack8(gb, g);
        unpack8(x[i], xv);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            s1[k] += g[k];
            s2[k] += g[k] * (xv[k] - mean[k]) * istd[k];
        }
    }
    block_reduce_cg(s1, s2, cvec, sh);
    if (threadIdx.x < cvec) {
        red_add8(scratch + cg * 8, s1);          // sum g
        red_add8(scratch + C + cg * 8, s2);      // sum g * xhat
    }
    // ---- grid barrier ----------------------------------------------------------------------
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        atomicAdd(counter, 1u);
        unsigned int seen;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(counter) : "memory");
        } while (seen < gridDim.x);
    }
    __syncthreads();
    // ---- pass 2 ----------------------------------------------------------------------------
    float a[8], b[8], c2[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = cg * 8 + k;
        const float sg = __ldcg(scratch + c), sgx = __ldcg(scratch + C + c);
        a[k] = gamma[c] * istd[k];
        b[k] = sg * inv_m;
        c2[k] = sgx * inv_m;
        if (blockIdx.x == 0 && threadIdx.x < cvec) {
            dbeta[c] += sg;
            dgamma[c] += sgx;
        }
    }
    for (long long i = i0; i < nvec; i += stride) {
        float gv[8], xv[8], o[8];
        unpack8(g_out[i], gv);
        unpack8(x[i], xv);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float xh = (xv[k] - mean[k]) * istd[k];
            o[k] = a[k] * (gv[k] - b[k] - xh * c2[k]);
        }
        dx[i] = pack8(o);
    }
}

// ------------------------------------------------------------------------------------ max pool
// 3x3, stride 2, pad 1 (NHWC).  idx holds the winning tap (0..8) for the backward pass.
__global__ void __launch_bounds__(kEwThreads)
maxpool_fwd_kernel(const bf16x8* __restrict__ x, bf16x8* __restrict__ y, uint8_t* __restrict__ idx,
                   int N, int H, int W, int cvec, int P, int Q) {
    pdl_launch_dependents();
    pdl_wait();
    const long long total = static_cast<long long>(N) * P * Q * cvec;
    for (long long i = static_cast<long long>(blockIdx.x) * kEwThreads + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * kEwThreads) {
        const int cg = static_cast<int>(i % cvec);
        long long t = i / cvec;
        const int q = static_cast<int>(t % Q);
        t /= Q;
        const int p = static_cast<int>(t % P);
        const int n = static_cast<int>(t / P);
        float best[8];
        int bi[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            best[k] = -INFINITY;
            bi[k] = 0;
        }
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int h = p * 2 - 1 + r;
            if (h < 0 || h >= H) continue;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const int w = q * 2 - 1 + s;
                if (w < 0 || w >= W) continue;
                float f[8];
                unpack8(x[((static_cast<long long>(n) * H + h) * W + w) * cvec + cg], f);
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (f[k] > best[k]) {
                        best[k] = f[k];
                        bi[k] = r * 3 + s;
                    }
            }
        }
        y[i] = pack8(best);
        uint2 packed;
        packed.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
        packed.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
        reinterpret_cast<uint2*>(idx)[i] = packed;
    }
}

__global__ void __launch_bounds__(kEwThreads)
maxpool_bwd_kernel(const bf16x8* __restrict__ dy, const bf16x8* __restrict__ dy2,
                   const uint8_t* __restrict__ idx,
                   bf16x8* __restrict__ dx, int N, int H, int W, int cvec, int P, int Q) {
    pdl_launch_dependents();
    pdl_wait();
    const long long total = static_cast<long long>(N) * H * W * cvec;
    for (long long i = static_cast<long long>(blockIdx.x) * kEwThreads + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * kEwThreads) {
        const int cg = static_cast<int>(i % cvec);
        long long t = i / cvec;
        const int w = static_cast<int>(t % W);
        t /= W;
        const int h = static_cast<int>(t % H);
        const int n = static_cast<int>(t / H);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        // output rows p with 2p-1 <= h <= 2p+1
        for (int p = h / 2; p <= (h + 1) / 2; ++p) {
            if (p < 0 || p >= P) continue;
            const int r = h - (p * 2 - 1);
            if (r < 0 || r > 2) continue;
            for (int q = w / 2; q <= (w + 1) / 2; ++q) {
                if (q < 0 || q >= Q) continue;
                const int s = w - (q * 2 - 1);
                if (s < 0 || s > 2) continue;
                const long long o = ((static_cast<long long>(n) * P + p) * Q + q) * cvec + cg;
                const uint2 pk = reinterpret_cast<const uint2*>(idx)[o];
                float g[8];
                unpack8(dy[o], g);
                if (dy2 != nullptr) {
                    float g2[8];
                    unpack8(dy2[o], g2);
#pragma unroll
                    for (int k = 0; k < 8; ++k) g[k] += g2[k];
                }
                const int tap = r * 3 + s;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint32_t word = k < 4 ? pk.x : pk.y;
                    const int sel = (word >> (8 * (k & 3))) & 0xff;
                    if (sel == tap) acc[k] += g[k];
                }
            }
        }
        dx[i] = pack8(acc);
    }
}

// ------------------------------------------------------------------------------------ head
// pooled[n,c] = mean_{hw} x[n,hw,c];  logits[n,k] = sum_c pooled[n,c] * w[k,c] + b[k]
__global__ void __launch_bounds__(256)
head_fwd_kernel(const lowp* __restrict__ x, const float* __restrict__ w,
                const float* __restrict__ b, float* __restrict__ pooled,
                float* __restrict__ logits, int HW, int C, int K) {
    pdl_launch_dependents();
    pdl_wait();
    extern __shared__ float sp[];
    const int n = blockIdx.x;
    const float inv = 1.f / HW;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float s = 0.f;
        for (int i = 0; i < HW; ++i) s += lp_to_float(x[(static_cast<long long>(n) * HW + i) * C + c]);
        s *= inv;
        sp[c] = s;
        pooled[static_cast<long long>(n) * C + c] = s;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int k = warp; k < K; k += nw) {
        float s = 0.f;
        for (int c = lane; c < C; c += 32) s += sp[c] * w[static_cast<long long>(k) * C + c];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) logits[static_cast<long long>(n) * K + k] = s + (b ? b[k] : 0.f);
    }
}

// dx[n,hw,c] = (sum_k dlogits[n,k] w[k,c]) / HW
__global__ void __launch_bounds__(256)
head_bwd_dx_kernel(const float* __restrict__ dlogits, const float* __restrict__ w,
                   lowp* __restrict__ dx, int HW, int C, int K) {
    pdl_launch_dependents();
    pdl_wait();
    extern __shared__ float sd[];
    const int n = blockIdx.x;
    for (int k = threadIdx.x; k < K; k += blockDim.x) sd[k] = dlogits[static_cast<long long>(n) * K + k];
    __syncthreads();
    const float inv = 1.f / HW;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float s = 0.f;
        for (int k = 0; k < K; ++k) s += sd[k] * w[static_cast<long long>(k) * C + c];
        const lowp v = float_to_lp(s * inv);
        for (int i = 0; i < HW; ++i) dx[(static_cast<long long>(n) * HW + i) * C + c] = v;
    }
}

// dw[k,c] += sum_n dlogits[n,k] pooled[n,c];  db[k] += sum_n dlogits[n,k]
__global__ void __launch_bounds__(64)
head_bwd_dw_kernel(const float* __restrict__ dlogits, const float* __restrict__ pooled,
                   float* __restrict__ dw, float* __restrict__ db, int B, int C, int K,
                   float unscale) {
    pdl_launch_dependents();
    pdl_wait();
    const int k = blockIdx.y;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < C) {
        float s = 0.f;
#pragma unroll 16
        for (int n = 0; n < B; ++n)
            s += __ldg(dlogits + static_cast<long long>(n) * K + k) *
                 __ldg(pooled + static_cast<long long>(n) * C + c);
        dw[static_cast<long long>(k) * C + c] += s * unscale;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && db != nullptr) {
        float s = 0.f;
        for (int n = 0; n < B; ++n) s += dlogits[static_cast<long long>(n) * K + k];
        db[k] += s * unscale;
    }
}

// ------------------------------------------------------------------------------------ CE
// loss = mean_n (logsumexp(z_n) - z_n[t_n]);  dz = (softmax - onehot) * scale / B.
// One block, one warp per row (looped).  Also accumulates epoch metrics on the device:
// metrics[0] += loss * B (f64), metrics[1] += #correct, metrics[2] += B.
__global__ void __launch_bounds__(1024)
ce_fwd_bwd_kernel(const float* __restrict__ logits, const long long* __restrict__ target,
                  float* __restrict__ loss_out, float* __restrict__ dlogits,
                  const float* __restrict__ loss_scale, double* __restrict__ metrics, int B, int K) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float s_loss[32];
    __shared__ int s_corr[32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const float scale = (loss_scale != nullptr ? *loss_scale : 1.f) / B;
    float wl = 0.f;
    int wc = 0;
#pragma unroll 4
    for (int n = warp; n < B; n += nw) {
        const float* z = logits + static_cast<long long>(n) * K;
        float mx = -INFINITY;
        int am = 0;
        for (int k = lane; k < K; k += 32) {
            const float v = z[k];
            if (v > mx) {
                mx = v;
                am = k;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float om = __shfl_xor_sync(0xffffffffu, mx, o);
            const int oa = __shfl_xor_sync(0xffffffffu, am, o);
            if (om > mx || (om == mx && oa < am)) {
                mx = om;
                am = oa;
            }
        }
        float se = 0.f;
        for (int k = lane; k < K; k += 32) se += __expf(z[k] - mx);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
        const int t = static_cast<int>(target[n]);
        const float lse = mx + __logf(se);
        if (dlogits != nullptr) {
            const float inv = 1.f / se;
            for (int k = lane; k < K; k += 32) {
                const float p = __expf(z[k] - mx) * inv;
                dlogits[static_cast<long long>(n) * K + k] = (p - (k == t ? 1.f : 0.f)) * scale;
            }
        }
        if (lane == 0) {
            wl += lse - z[t];
            wc += (am == t);
        }
    }
    if (lane == 0) {
        s_loss[warp] = wl;
        s_corr[warp] = wc;
    }
    __syncthreads();
    if (warp == 0) {
        float l = lane < nw ? s_loss[lane] : 0.f;
        int c = lane < nw ? s_corr[lane] : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            l += __shfl_xor_sync(0xffffffffu, l, o);
            c += __shfl_xor_sync(0xffffffffu, c, o);
        }
        if (lane == 0) {
            loss_out[0] = l / B;
            if (metrics != nullptr) {
                metrics[0] += static_cast<double>(l);
                metrics[1] += static_cast<double>(c);
                metrics[2] += static_cast<double>(B);
            }
        }
    }
}

// ------------------------------------------------------------------------------------ stem
// NCHW fp32 image -> im2col rows [N*P*Q][ldk] bf16, column (r*S+s)*C + c (KRSC order);
// columns >= R*S*C are zero.
__global__ void __launch_bounds__(kEwThreads)
im2col_nchw_kernel(const float* __restrict__ x, lowp* __restrict__ out, int N, int C,
                   int H, int W, int R, int S, int stride, int pad, int P, int Q, int ldk) {
    pdl_launch_dependents();
    pdl_wait();
    // one thread = 8 consecutive columns of one row (ldk % 8 == 0): a single 16-byte store
    const int cv = ldk / 8;
    const long long total = static_cast<long long>(N) * P * Q * cv;
    const int K = R * S * C;
    // column -> (input offset, tap) table: keeps the four runtime-divisor divisions per element
    // out of the main loop (the 224x224 stem builds 3.2M rows x 160 columns per step)
    constexpr int kTab = 512;
    __shared__ int s_off[kTab];
    __shared__ short s_rs[kTab];
    const bool use_tab = ldk <= kTab;
    if (use_tab) {
        for (int col = threadIdx.x; col < ldk; col += kEwThreads) {
            if (col < K) {
                const int c = col % C;
                const int rs = col / C;
                const int s = rs % S, r = rs / S;
                s_off[col] = (c * H + r) * W + s;
                s_rs[col] = static_cast<short>((r << 8) | s);
            } else {
                s_off[col] = 0;
                s_rs[col] = -1;
            }
        }
        __syncthreads();
    }
    for (long long i = static_cast<long long>(blockIdx.x) * kEwThreads + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * kEwThreads) {
        const int col0 = static_cast<int>(i % cv) * 8;
        long long t = i / cv;
        const int q = static_cast<int>(t % Q);
        t /= Q;
        const int p = static_cast<int>(t % P);
        const int n = static_cast<int>(t / P);
        float v[8];
        if (use_tab) {
            const int h0 = p * stride - pad, w0 = q * stride - pad;
            const float* base = x + (static_cast<long long>(n) * C * H + h0) * W + w0;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int rs = s_rs[col0 + e];
                float val = 0.f;
                if (rs >= 0) {
                    const int h = h0 + (rs >> 8), w = w0 + (rs & 255);
                    if (h >= 0 && h < H && w >= 0 && w < W) val = __ldg(base + s_off[col0 + e]);
                }
                v[e] = val;
            }
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int col = col0 + e;
                float val = 0.f;
                if (col < K) {
                    const int c = col % C;
                    const int rs = col / C;
                    const int s = rs % S, r = rs / S;
                    const int h = p * stride - pad + r, w = q * stride - pad + s;
                    if (h >= 0 && h < H && w >= 0 && w < W)
                        val = __ldg(x + ((static_cast<long long>(n) * C + c) * H + h) * W + w);
                }
                v[e] = val;
            }
        }
        reinterpret_cast<bf16x8*>(out)[i] = pack8(v);
    }
}

// fp32 [rows][cols] -> bf16 [rows][ld] (zero padded); used for the stem weight whose 147-wide
// rows are not 16-byte aligned.
__global__ void __launch_bounds__(kEwThreads)
pad_rows_kernel(const float* __restrict__ src, lowp* __restrict__ dst, int rows, int cols,
                int ld) {
    pdl_launch_dependents();
    pdl_wait();
    const int total = rows * ld;
    for (int i = blockIdx.x * kEwThreads + threadIdx.x; i < total; i += gridDim.x * kEwThreads) {
        const int c = i % ld, r = i / ld;
        dst[i] = float_to_lp(c < cols ? src[r * cols + c] : 0.f);
    }
}

// NCHW fp32 -> NHWC bf16 (general layout change for inputs with C % 8 == 0 handled by callers)
__global__ void __launch_bounds__(kEwThreads)
nchw_to_nhwc_kernel(const float* __restrict__ x, lowp* __restrict__ y, int N, int C, int H,
                    int W) {
    pdl_launch_dependents();
    pdl_wait();
    const long long total = static_cast<long long>(N) * C * H * W;
    for (long long i = static_cast<long long>(blockIdx.x) * kEwThreads + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * kEwThreads) {
        const int c = static_cast<int>(i % C);
        long long t = i / C;
        const int w = static_cast<int>(t % W);
        t /= W;
        const int h = static_cast<int>(t % H);
        const int n = static_cast<int>(t / H);
        y[i] = float_to_lp(x[((static_cast<long long>(n) * C + c) * H + h) * W + w]);
    }
}

// ------------------------------------------------------------------------------------ augment
// CIFAR training transform on the device: RandomCrop(S, padding) + RandomHorizontalFlip +
// ToTensor + Normalize in one pass (reference train_ddp.py:91-96 does this per sample on the
// host with PIL).  in: uint8 [N, C, S, S]; params: int32 [N, 3] = (dy, dx, flip) with dy, dx in
// [0, 2*pad]; out: fp32 [N, C, S, S].
__global__ void __launch_bounds__(kEwThreads)
augment_crop_flip_norm_kernel(const uint8_t* __restrict__ in, const int* __restrict__ params,
                              float* __restrict__ out, int N, int C, int S, int pad, float m0,
                              float m1, float m2, float is0, float is1, float is2) {
    pdl_launch_dependents();
    pdl_wait();
    const long long total = static_cast<long long>(N) * C * S * S;
    for (long long i = static_cast<long long>(blockIdx.x) * kEwThreads + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * kEwThreads) {
        const int w = static_cast<int>(i % S);
        long long t = i / S;
        const int h = static_cast<int>(t % S);
        t /= S;
        const int c = static_cast<int>(t % C);
        const int n = static_cast<int>(t / C);
        const int dy = params[n * 3 + 0], dx = params[n * 3 + 1], flip = params[n * 3 + 2];
        const int ws = flip ? (S - 1 - w) : w;
        const int hs = h + dy - pad, wsrc = ws + dx - pad;
        float v = 0.f;
        if (hs >= 0 && hs < S && wsrc >= 0 && wsrc < S)
            v = static_cast<float>(in[((static_cast<long long>(n) * C + c) * S + hs) * S + wsrc]) *
                (1.f / 255.f);
        const float mean = c == 0 ? m0 : (c == 1 ? m1 : m2);
        const float istd = c == 0 ? is0 : (c == 1 ? is1 : is2);
        out[i] = (v - mean) * istd;
    }
}




#define DPT_CHECK_LAUNCH() return static_cast<int>(cudaGetLastError())

// debugging knob: DPT_PDL_EW=0 launches the kernels of this file without PDL
static inline void ew_pdl_gate() {
    static int on = -1;
    if (on < 0) {
        const char* e = std::getenv("DPT_PDL_EW");
        on = (e == nullptr || e[0] != '0') ? 1 : 0;
    }
    if (!on) pdl_skip_flag() = 1;
}
static inline bool cvec_ok(int C) { return C % 8 == 0 && (kEwThreads % (C / 8)) == 0; }

inline int dpt_bn_stats(const void* x, float* sum, float* sq, long long M, int C, void* stream) {
    if (!cvec_ok(C)) return DPT_ERR_SHAPE;
    const long long nvec = M * (C / 8);
    ew_pdl_gate();
    launch_pdl(bn_stats_kernel, dim3(reduce_grid(nvec)), dim3(kEwThreads), 0, static_cast<cudaStream_t>(stream), 
        static_cast<const bf16x8*>(x), sum, sq, nvec, C / 8);
    DPT_CHECK_LAUNCH();
}

inline int dpt_bn_fwd_train(const void* x, const float* sum, const float* sq, const float* gamma,
                                const float* beta, const void* residual, void* y, float* save_mean,
                                float* save_invstd, float* running_mean, float* running_var,
                                long long* num_batches_tracked, long long M, int C, float momentum,
                                float eps, int relu, void* stream) {
    if (!cvec_ok(C)) return DPT_ERR_SHAPE;
    const long long nvec = M * (C / 8);
    const float unbias = M > 1 ? static_cast<float>(M) / static_cast<float>(M - 1) : 1.f;
    ew_pdl_gate();
    if (bulk_worthwhile(nvec, 1)) {
        static int attr = bulk_smem_attr(bn_fwd_bulk_kernel, 2);
        if (attr != 0) return attr;
        launch_pdl(bn_fwd_bulk_kernel, dim3(bulk_grid(nvec, 3)), dim3(kEwThreads),
                   static_cast<size_t>(kBulkStages * 2 * kBulkChunk), static_cast<cudaStream_t>(stream),
                   static_cast<const bf16x8*>(x), sum, sq, gamma, beta,
                   static_cast<const bf16x8*>(residual), static_cast<bf16x8*>(y), save_mean,
                   save_invstd, running_mean, running_var, num_batches_tracked, nvec, C / 8,
                   1.f / static_cast<float>(M), unbias, momentum, eps, relu);
        DPT_CHECK_LAUNCH();
    }
    launch_pdl(bn_fwd_kernel, dim3(ew_grid(nvec, C / 8)), dim3(kEwThreads), 0, static_cast<cudaStream_t>(stream), 
        static_cast<const bf16x8*>(x), sum, sq, gamma, beta, static_cast<const bf16x8*>(residual),
        static_cast<bf16x8*>(y), save_mean, save_invstd, running_mean, running_var,
        num_batches_tracked, nvec, C / 8, 1.f / static_cast<float>(M), unbias, momentum, eps, relu);
    DPT_CHECK_LAUNCH();
}

inline int dpt_bn_fwd_train_small_f32(const float* x, const float* gamma, const float* beta,
                                          const void* residual, void* y, void* x_bf16,
                                          float* save_mean, float* save_invstd,
                                          float* running_mean, float* running_var,
                                          long long* num_batches_tracked, long long M, int C,
                                          float momentum, float eps, int relu, void* stream) {
    if (C % 8 != 0 || M <= 0 || M > (1 << 24)) return DPT_ERR_SHAPE;
    const float unbias = M > 1 ? static_cast<float>(M) / static_cast<float>(M - 1) : 1.f;
    ew_pdl_gate();
    // Measured on B200: launched with the programmatic attribute right behind a split-K conv,
    // this kernel observed the accumulator before the conv's reductions had landed (CUDA-graph
    // replay only).  Consumers of RED-accumulated fp32 data are therefore ordinary launches.
    pdl_skip_flag() = 1;
    launch_pdl(bn_fwd_small_f32_kernel, dim3(C / 8), dim3(kEwThreads), 0,
               static_cast<cudaStream_t>(stream), x, gamma, beta,
               static_cast<const bf16x8*>(residual), static_cast<bf16x8*>(y),
               static_cast<bf16x8*>(x_bf16), save_mean, save_invstd, running_mean, running_var,
               num_batches_tracked, static_cast<int>(M), C / 8, 1.f / static_cast<float>(M), unbias,
               momentum, eps, relu);
    DPT_CHECK_LAUNCH();
}

inline int dpt_bn_fwd_eval(const void* x, const float* gamma, const float* beta,
                               const float* running_mean, const float* running_var,
                               const void* residual, void* y, long long M, int C, float eps,
                               int relu, void* stream) {
    if (!cvec_ok(C)) return DPT_ERR_SHAPE;
    const long long nvec = M * (C / 8);
    ew_pdl_gate();
    launch_pdl(bn_eval_kernel, dim3(ew_grid(nvec, C / 8)), dim3(kEwThreads), 0, static_cast<cudaStream_t>(stream), 
        static_cast<const bf16x8*>(x), gamma, beta, running_mean, running_var,
        static_cast<const bf16x8*>(residual), static_cast<bf16x8*>(y), nvec, C / 8, eps, relu);
    DPT_CHECK_LAUNCH();
}

inline int dpt_bn_bwd_reduce_ex(const void* dy1, int dy1_is_f32, const void* dy2, const void* y,
                                    const void* x, const float* save_mean,
                                    const float* save_invstd, void* g_out, float* dgamma,
                                    float* dbeta, long long M, int C, void* stream);

inline int dpt_bn_bwd_reduce(const void* dy1, const void* dy2, const void* y, const void* x,
                                 const float* save_mean, const float* save_invstd, void* g_out,
                                 float* dgamma, float* dbeta, long long M, int C, void* stream) {
    return dpt_bn_bwd_reduce_ex(dy1, 0, dy2, y, x, save_mean, save_invstd, g_out, dgamma, dbeta, M,
                                C, stream);
}

// dy1 may be fp32 (output of a split-K dgrad); g_out is then mandatory (pass 2 reads bf16 g).
inline int dpt_bn_bwd_reduce_ex(const void* dy1, int dy1_is_f32, const void* dy2, const void* y,
                                    const void* x, const float* save_mean,
                                    const float* save_invstd, void* g_out, float* dgamma,
                                    float* dbeta, long long M, int C, void* stream) {
    if (!cvec_ok(C)) return DPT_ERR_SHAPE;
    if (dy1_is_f32 && g_out == nullptr) return DPT_ERR_ARG;
    const long long nvec = M * (C / 8);
    auto kern = dy1_is_f32 ? bn_bwd_reduce_kernel<true> : bn_bwd_reduce_kernel<false>;
    ew_pdl_gate();
    if (!dy1_is_f32 && bulk_worthwhile(nvec, 2)) {
        static int attr = bulk_smem_attr(bn_bwd_reduce_bulk_kernel, 4, kBulkStagesReduce);
        if (attr != 0) return attr;
        launch_pdl(bn_bwd_reduce_bulk_kernel, dim3(bulk_grid(nvec, 2)), dim3(kEwThreads),
                   static_cast<size_t>(kBulkStagesReduce * 4 * kBulkChunk), static_cast<cudaStream_t>(stream),
                   static_cast<const bf16x8*>(dy1), static_cast<const bf16x8*>(dy2),
                   static_cast<const bf16x8*>(y), static_cast<const bf16x8*>(x), save_mean,
                   save_invstd, static_cast<bf16x8*>(g_out), dgamma, dbeta, nvec, C / 8);
        DPT_CHECK_LAUNCH();
    }
    if (dy1_is_f32) pdl_skip_flag() = 1;  // dy1 was accumulated with REDs by a split-K dgrad
    launch_pdl(kern, dim3(reduce_grid(nvec)), dim3(kEwThreads), 0, static_cast<cudaStream_t>(stream), 
        dy1, static_cast<const bf16x8*>(dy2),
        static_cast<const bf16x8*>(y), static_cast<const bf16x8*>(x), save_mean, save_invstd,
        static_cast<bf16x8*>(g_out), dgamma, dbeta, nvec, C / 8);
    DPT_CHECK_LAUNCH();
}

// One-launch BatchNorm backward (see bn_bwd_fused_kernel).  scratch: 2*C zeroed floats; counter:
// one zeroed uint32; g_out is mandatory (pass 2 re-reads the masked gradient from it).
inline int dpt_bn_bwd_fused(const void* dy1, int dy1_is_f32, const void* dy2, const void* y,
                                const void* x, const float* save_mean, const float* save_invstd,
                                const float* gamma, void* g_out, float* dgamma, float* dbeta,
                                float* scratch, unsigned int* counter, void* dx, long long M, int C,
                                void* stream) {
    if (!cvec_ok(C) || g_out == nullptr) return DPT_ERR_SHAPE;
    const long long nvec = M * (C / 8);
    auto kern = dy1_is_f32 ? bn_bwd_fused_kernel<true> : bn_bwd_fused_kernel<false>;
    ew_pdl_gate();
    if (dy1_is_f32) pdl_skip_flag() = 1;  // dy1 was accumulated with REDs by a split-K dgrad
    launch_pdl(kern, dim3(reduce_grid_resident(nvec)), dim3(kEwThreads), 0, static_cast<cudaStream_t>(stream),
               dy1, static_cast<const bf16x8*>(dy2), static_cast<const bf16x8*>(y),
               static_cast<const bf16x8*>(x), save_mean, save_invstd, gamma,
               static_cast<bf16x8*>(g_out), dgamma, dbeta, scratch, counter,
               static_cast<bf16x8*>(dx), nvec, C / 8, 1.f / static_cast<float>(M));
    DPT_CHECK_LAUNCH();
}

inline int dpt_bn_bwd_apply(const void* g, const void* x, const float* save_mean,
                                const float* save_invstd, const float* gamma, const float* dgamma,
                                const float* dbeta, void* dx, long long M, int C, void* stream) {
    if (!cvec_ok(C)) return DPT_ERR_SHAPE;
    const long long nvec = M * (C / 8);
    ew_pdl_gate();
    if (bulk_worthwhile(nvec, 4)) {
        static int attr = bulk_smem_attr(bn_bwd_apply_bulk_kernel, 2);
        if (attr != 0) return attr;
        launch_pdl(bn_bwd_apply_bulk_kernel, dim3(bulk_grid(nvec, 3)), dim3(kEwThreads),
                   static_cast<size_t>(kBulkStages * 2 * kBulkChunk), static_cast<cudaStream_t>(stream),
                   static_cast<const bf16x8*>(g), static_cast<const bf16x8*>(x), save_mean,
                   save_invstd, gamma, dgamma, dbeta, static_cast<bf16x8*>(dx), nvec, C / 8,
                   1.f / static_cast<float>(M), 1.f);
        DPT_CHECK_LAUNCH();
    }
    launch_pdl(bn_bwd_apply_kernel, dim3(ew_grid(nvec, C / 8)), dim3(kEwThreads), 0, static_cast<cudaStream_t>(stream), 
        static_cast<const bf16x8*>(g), static_cast<const bf16x8*>(x), save_mean, save_invstd, gamma,
        dgamma, dbeta, static_cast<bf16x8*>(dx), nvec, C / 8, 1.f / static_cast<float>(M), 1.f);
    DPT_CHECK_LAUNCH();
}

inline int dpt_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C,
                                    void* stream) {
    if (C % 8 != 0) return DPT_ERR_SHAPE;
    const int P = (H + 2 - 3) / 2 + 1, Q = (W + 2 - 3) / 2 + 1;
    const long long total = static_cast<long long>(N) * P * Q * (C / 8);
    ew_pdl_gate();
    launch_pdl(maxpool_fwd_kernel, dim3(ew_grid(total, 1)), dim3(kEwThreads), 0, static_cast<cudaStream_t>(stream), 
        static_cast<const bf16x8*>(x), static_cast<bf16x8*>(y), static_cast<uint8_t*>(idx), N, H, W,
        C / 8, P, Q);
    DPT_CHECK_LAUNCH();
}

inline int dpt_maxpool3x3s2_bwd(const void* dy, const void* dy2, const void* idx, void* dx,
                                    int N, int H, int W, int C, void* stream) {
    if (C % 8 != 0) return DPT_ERR_SHAPE;
    const int P = (H + 2 - 3) / 2 + 1, Q = (W + 2 - 3) / 2 + 1;
    const long long total = static_cast<long long>(N) * H * W * (C / 8);
    ew_pdl_gate();
    launch_pdl(maxpool_bwd_kernel, dim3(ew_grid(total, 1)), dim3(kEwThreads), 0, static_cast<cudaStream_t>(stream), 
        static_cast<const bf16x8*>(dy), static_cast<const bf16x8*>(dy2),
        static_cast<const uint8_t*>(idx), static_cast<bf16x8*>(dx), N, H, W, C / 8, P, Q);
    DPT_CHECK_LAUNCH();
}

inline int dpt_head_fwd(const void* x, const float* w, const float* b, float* pooled,
                            float* logits, int B, int HW, int C, int K, void* stream) {
    ew_pdl_gate();
    launch_pdl(head_fwd_kernel, dim3(B), dim3(256), C * sizeof(float), static_cast<cudaStream_t>(stream), 
        static_cast<const lowp*>(x), w, b, pooled, logits, HW, C, K);
    DPT_CHECK_LAUNCH();
}

inline int dpt_head_bwd(const float* dlogits, const float* w, const float* pooled, void* dx,
                            float* dw, float* db, int B, int HW, int C, int K, void* stream) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // dx == NULL or dw == NULL skips that half (the two kernels are independent, so callers may
    // issue them on different streams)
    if (dx != nullptr) {
        ew_pdl_gate();
        launch_pdl(head_bwd_dx_kernel, dim3(B), dim3(256), K * sizeof(float), st, dlogits, w,
                   static_cast<lowp*>(dx), HW, C, K);
    }
    if (dw != nullptr) {
        dim3 grid((C + 63) / 64, K);
        ew_pdl_gate();
        launch_pdl(head_bwd_dw_kernel, dim3(grid), dim3(64), 0, st, dlogits, pooled, dw, db, B, C, K,
                   1.f);
    }
    DPT_CHECK_LAUNCH();
}

inline int dpt_cross_entropy(const float* logits, const long long* target, float* loss,
                                 float* dlogits, const float* loss_scale, double* metrics, int B,
                                 int K, void* stream) {
    int threads = B * 32;
    if (threads > 1024) threads = 1024;
    if (threads < 32) threads = 32;
    ew_pdl_gate();
    launch_pdl(ce_fwd_bwd_kernel, dim3(1), dim3(threads), 0, static_cast<cudaStream_t>(stream), 
        logits, target, loss, dlogits, loss_scale, metrics, B, K);
    DPT_CHECK_LAUNCH();
}

inline int dpt_im2col_nchw_f32(const float* x, void* out, int N, int C, int H, int W, int R,
                                   int S, int stride, int pad, int ldk, void* stream) {
    const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
    if (ldk % 8 != 0) return DPT_ERR_SHAPE;
    const long long total = static_cast<long long>(N) * P * Q * (ldk / 8);
    ew_pdl_gate();
    launch_pdl(im2col_nchw_kernel, dim3(ew_grid(total, 1)), dim3(kEwThreads), 0, static_cast<cudaStream_t>(stream), 
        x, static_cast<lowp*>(out), N, C, H, W, R, S, stride, pad, P, Q, ldk);
    DPT_CHECK_LAUNCH();
}

inline int dpt_pad_rows_f32_to_bf16(const float* src, void* dst, int rows, int cols, int ld,
                                        void* stream) {
    ew_pdl_gate();
    launch_pdl(pad_rows_kernel, dim3(ew_grid(static_cast<long long>(rows) * ld, 1)), dim3(kEwThreads), 0, static_cast<cudaStream_t>(stream), src, static_cast<lowp*>(dst),
                                                           rows, cols, ld);
    DPT_CHECK_LAUNCH();
}

inline int dpt_nchw_f32_to_nhwc_bf16(const float* x, void* y, int N, int C, int H, int W,
                                         void* stream) {
    const long long total = static_cast<long long>(N) * C * H * W;
    ew_pdl_gate();
    launch_pdl(nchw_to_nhwc_kernel, dim3(ew_grid(total, 1)), dim3(kEwThreads), 0, static_cast<cudaStream_t>(stream), 
        x, static_cast<lowp*>(y), N, C, H, W);
    DPT_CHECK_LAUNCH();
}


// Statistics + normalisation with caller-provided (already zeroed) scratch of 2*C floats: the
// per-element-type half of dpt_bn_fwd_train_noalloc (norm_pool_loss.cu).
inline int bn_fwd_train_with_scratch(float* scratch, const void* x, const float* gamma,
                                     const float* beta, const void* residual, void* y,
                                     float* save_mean, float* save_invstd, float* running_mean,
                                     float* running_var, long long* num_batches_tracked,
                                     long long M, int C, float momentum, float eps, int relu,
                                     void* stream) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float* sum = scratch;
    float* sq = scratch + C;
    const long long nvec = M * (C / 8);
    ew_pdl_gate();
    launch_pdl(bn_stats_kernel, dim3(reduce_grid(nvec)), dim3(kEwThreads), 0, st,
               static_cast<const bf16x8*>(x), sum, sq, nvec, C / 8);
    const float unbias = M > 1 ? static_cast<float>(M) / static_cast<float>(M - 1) : 1.f;
    ew_pdl_gate();
    launch_pdl(bn_fwd_kernel, dim3(ew_grid(nvec, C / 8)), dim3(kEwThreads), 0, st,
        static_cast<const bf16x8*>(x), sum, sq, gamma, beta, static_cast<const bf16x8*>(residual),
        static_cast<bf16x8*>(y), save_mean, save_invstd, running_mean, running_var,
        num_batches_tracked, nvec, C / 8, 1.f / static_cast<float>(M), unbias, momentum, eps, relu);
    DPT_CHECK_LAUNCH();
}

inline int dpt_augment_crop_flip_norm(const void* in_u8, const int* params, float* out, int N,
                                          int C, int S, int pad, const float* mean3,
                                          const float* std3, void* stream) {
    if (C > 3 || C < 1) return DPT_ERR_SHAPE;
    const long long total = static_cast<long long>(N) * C * S * S;
    ew_pdl_gate();
    launch_pdl(augment_crop_flip_norm_kernel, dim3(ew_grid(total, 1)), dim3(kEwThreads), 0,
               static_cast<cudaStream_t>(stream), static_cast<const uint8_t*>(in_u8), params, out, N,
               C, S, pad, mean3[0], mean3[1], mean3[2], 1.f / std3[0], 1.f / std3[1], 1.f / std3[2]);
    DPT_CHECK_LAUNCH();
}

}  // namespace DPT_NS
