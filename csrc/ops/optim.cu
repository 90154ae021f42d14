// Flat-arena optimizer and AMP kernels (sm_100a).
//
// The whole model lives in three parallel fp32 arenas (param / grad / momentum) plus a bf16
// "shadow" arena holding the low-precision weight copy the tensor-core kernels consume.  One
// launch updates everything: it replaces torch.optim.SGD's foreach path (4 multi-tensor launches
// per step, reference train_ddp.py:339-344 -> torch/optim/sgd.py:425-473), the per-forward
// autocast weight casts (train_ddp.py:204), GradScaler's unscale/inf-check
// (torch/amp/grad_scaler.py:292-298) and its host-synchronising found_inf.item() (:371).
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/dptb200.h"
#include "../common/ptx.cuh"

namespace dpt {

static int g_shadow_f16 = 0;  // element type of the 16-bit shadow weights (dpt_set_lowp_dtype)

// pack 2 floats into the 16-bit shadow format
__device__ __forceinline__ uint32_t pack_lp2(float a, float b, int f16) {
    if (f16) {
        __half2 h = __floats2half2_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&h);
    }
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint16_t pack_lp1(float a, int f16) {
    if (f16) {
        __half h = __float2half_rn(a);
        return *reinterpret_cast<uint16_t*>(&h);
    }
    __nv_bfloat16 h = __float2bfloat16_rn(a);
    return *reinterpret_cast<uint16_t*>(&h);
}

__device__ __forceinline__ bool finite4(float4 v) {
    return isfinite(v.x) && isfinite(v.y) && isfinite(v.z) && isfinite(v.w);
}

// g *= inv_scale; found_inf |= any(!finite(g))
__global__ void __launch_bounds__(256)
amp_unscale_kernel(float* __restrict__ g, long long n, const float* __restrict__ inv_scale,
                   float* __restrict__ found_inf) {
    pdl_launch_dependents();
    pdl_wait();
    const float s = inv_scale ? *inv_scale : 1.f;
    bool bad = false;
    const long long n4 = n / 4;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float4 v = reinterpret_cast<float4*>(g)[i];
        bad |= !finite4(v);
        v.x *= s; v.y *= s; v.z *= s; v.w *= s;
        reinterpret_cast<float4*>(g)[i] = v;
    }
    for (long long i = n4 * 4 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float v = g[i];
        bad |= !isfinite(v);
        g[i] = v * s;
    }
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) *found_inf = 1.f;
}

// GradScaler.update(): scale/growth_tracker state machine, entirely on the device.
// inv_scale_out (optional) receives 1/new_scale for the next step's fused unscale.
__global__ void amp_update_scale_kernel(float* scale, int* growth_tracker, float* found_inf,
                                        float* inv_scale_out, float growth_factor,
                                        float backoff_factor, int growth_interval) {
    pdl_launch_dependents();
    pdl_wait();
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        if (*found_inf > 0.f) {
            *scale = *scale * backoff_factor;
            *growth_tracker = 0;
        } else {
            const int t = *growth_tracker + 1;
            if (t == growth_interval) {
                const float ns = *scale * growth_factor;
                if (isfinite(ns)) *scale = ns;
                *growth_tracker = 0;
            } else {
                *growth_tracker = t;
            }
        }
        *found_inf = 0.f;
        if (inv_scale_out != nullptr) *inv_scale_out = 1.f / *scale;
    }
}

// SGD with momentum / weight decay / nesterov (torch.optim.SGD semantics with the momentum
// buffer starting at zero, which is identical to torch's "clone the first gradient" for
// dampening = 0).  Optionally: g is scaled by *grad_scale first (AMP unscale / 1/world), the
// whole update is skipped when *found_inf != 0, and the bf16 shadow copy is refreshed.
__global__ void __launch_bounds__(256)
sgd_flat_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                uint16_t* __restrict__ shadow, long long n, float lr, float momentum,
                float dampening, float wd, int nesterov, const float* __restrict__ grad_scale,
                const float* __restrict__ found_inf, int shadow_f16) {
    pdl_launch_dependents();
    pdl_wait();
    if (found_inf != nullptr && *found_inf > 0.f) return;
    const float gs = grad_scale ? *grad_scale : 1.f;
    const long long n4 = n / 4;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float4 pv = reinterpret_cast<float4*>(p)[i];
        const float4 gv = reinterpret_cast<const float4*>(g)[i];
        float4 mv = reinterpret_cast<float4*>(m)[i];
        float pa[4] = {pv.x, pv.y, pv.z, pv.w};
        const float ga[4] = {gv.x, gv.y, gv.z, gv.w};
        float ma[4] = {mv.x, mv.y, mv.z, mv.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float d = fmaf(wd, pa[k], ga[k] * gs);
            if (momentum != 0.f) {
                ma[k] = fmaf(momentum, ma[k], (1.f - dampening) * d);
                d = nesterov ? fmaf(momentum, ma[k], d) : ma[k];
            }
            pa[k] = fmaf(-lr, d, pa[k]);
        }
        reinterpret_cast<float4*>(p)[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
        reinterpret_cast<float4*>(m)[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
        if (shadow != nullptr) {
            uint2 pk;
            pk.x = pack_lp2(pa[0], pa[1], shadow_f16);
            pk.y = pack_lp2(pa[2], pa[3], shadow_f16);
            reinterpret_cast<uint2*>(shadow)[i] = pk;
        }
    }
    for (long long i = n4 * 4 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float d = fmaf(wd, p[i], g[i] * gs);
        if (momentum != 0.f) {
            m[i] = fmaf(momentum, m[i], (1.f - dampening) * d);
            d = nesterov ? fmaf(momentum, m[i], d) : m[i];
        }
        p[i] = fmaf(-lr, d, p[i]);
        if (shadow != nullptr) shadow[i] = pack_lp1(p[i], shadow_f16);
    }
}

// Adam / AdamW.  step_count is a device int incremented by block 0 *after* use by everyone
// (each thread reads the pre-increment value + 1; the increment happens in a tail kernel).
__global__ void __launch_bounds__(256)
adam_flat_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                 float* __restrict__ v, uint16_t* __restrict__ shadow, long long n, float lr,
                 float beta1, float beta2, float eps, float wd, int decoupled,
                 const int* __restrict__ step_count, const float* __restrict__ grad_scale,
                 const float* __restrict__ found_inf, int shadow_f16) {
    pdl_launch_dependents();
    pdl_wait();
    if (found_inf != nullptr && *found_inf > 0.f) return;
    const float gs = grad_scale ? *grad_scale : 1.f;
    const float t = static_cast<float>(*step_count + 1);
    const float bc1 = 1.f - powf(beta1, t), bc2 = 1.f - powf(beta2, t);
    const float step_size = lr / bc1;
    const float inv_sqrt_bc2 = rsqrtf(bc2);
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float pv = p[i];
        float gv = g[i] * gs;
        if (decoupled)
            pv *= (1.f - lr * wd);
        else
            gv = fmaf(wd, pv, gv);
        const float mv = fmaf(beta1, m[i], (1.f - beta1) * gv);
        const float vv = fmaf(beta2, v[i], (1.f - beta2) * gv * gv);
        m[i] = mv;
        v[i] = vv;
        const float denom = sqrtf(vv) * inv_sqrt_bc2 + eps;
        pv -= step_size * mv / denom;
        p[i] = pv;
        if (shadow != nullptr) shadow[i] = pack_lp1(pv, shadow_f16);
    }
}

__global__ void step_inc_kernel(int* step_count, const float* found_inf) {
    pdl_launch_dependents();
    pdl_wait();
    if (found_inf != nullptr && *found_inf > 0.f) return;
    *step_count += 1;
}

__global__ void __launch_bounds__(256)
cast_f32_bf16_kernel(const float* __restrict__ src, uint16_t* __restrict__ dst, long long n,
                     int f16) {
    pdl_launch_dependents();
    pdl_wait();
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x)
        dst[i] = pack_lp1(src[i], f16);
}

static inline int flat_grid(long long n) {
    long long b = (n / 4 + 255) / 256;
    if (b > 148 * 8) b = 148 * 8;
    if (b < 1) b = 1;
    return static_cast<int>(b);
}

}  // namespace dpt

using namespace dpt;

extern "C" int dpt_amp_unscale(float* g, long long n, const float* inv_scale, float* found_inf,
                               void* stream) {
    launch_pdl(amp_unscale_kernel, dim3(flat_grid(n)), dim3(256), 0, static_cast<cudaStream_t>(stream), g, n, inv_scale,
                                                                                   found_inf);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int dpt_amp_update_scale(float* scale, int* growth_tracker, float* found_inf,
                                    float* inv_scale_out, float growth_factor, float backoff_factor,
                                    int growth_interval, void* stream) {
    launch_pdl(amp_update_scale_kernel, dim3(1), dim3(32), 0, static_cast<cudaStream_t>(stream), 
        scale, growth_tracker, found_inf, inv_scale_out, growth_factor, backoff_factor,
        growth_interval);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int dpt_sgd_flat(float* p, const float* g, float* m, void* shadow, long long n, float lr,
                            float momentum, float dampening, float wd, int nesterov,
                            const float* grad_scale, const float* found_inf, void* stream) {
    launch_pdl(sgd_flat_kernel, dim3(flat_grid(n)), dim3(256), 0, static_cast<cudaStream_t>(stream), 
        p, g, m, static_cast<uint16_t*>(shadow), n, lr, momentum, dampening, wd, nesterov,
        grad_scale, found_inf, g_shadow_f16);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int dpt_adam_flat(float* p, const float* g, float* m, float* v, void* shadow, long long n,
                             float lr, float beta1, float beta2, float eps, float wd, int decoupled,
                             int* step_count, const float* grad_scale, const float* found_inf,
                             void* stream) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    launch_pdl(adam_flat_kernel, dim3(flat_grid(n)), dim3(256), 0, st, p, g, m, v, static_cast<uint16_t*>(shadow),
                                                   n, lr, beta1, beta2, eps, wd, decoupled,
                                                   step_count, grad_scale, found_inf, g_shadow_f16);
    launch_pdl(step_inc_kernel, dim3(1), dim3(1), 0, st, step_count, found_inf);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int dpt_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream) {
    launch_pdl(cast_f32_bf16_kernel, dim3(flat_grid(n * 4)), dim3(256), 0, static_cast<cudaStream_t>(stream), 
        src, static_cast<uint16_t*>(dst), n, g_shadow_f16);
    return static_cast<int>(cudaGetLastError());
}

// The next kernel launched through this library follows a cross-stream event wait: launch it
// without the programmatic-dependent-launch attribute (see csrc/common/ptx.cuh).
extern "C" int dpt_optim_set_lowp_dtype(int is_f16) {
    dpt::g_shadow_f16 = (is_f16 == 1) ? 1 : 0;
    return 0;
}

extern "C" int dpt_pdl_skip_next() {
    dpt::pdl_skip_flag() = 1;
    return 0;
}
