// C++ gradient reducer over CUDA streams/events (no torch Reducer, no ProcessGroup on the hot
// path).  Stands in for torch's c10d::Reducer as driven by the reference through
// DistributedDataParallel (train_ddp.py:303-311; torch reducer.hpp:45-584):
//   * parameters are grouped into size-capped buckets (first bucket small, then bucket_cap);
//   * a bucket is launched the moment its last gradient is ready, strictly in bucket order;
//   * the launch is ONE peer-memory all-reduce kernel (csrc/comm) on a dedicated high-priority
//     stream, ordered after the producing kernels with an event, so it overlaps the rest of
//     backward; 1/world averaging, AMP unscale, the non-finite check and optionally the SGD
//     update are fused into that kernel;
//   * gradients are born in the bucket (the arena slice), so there is no copy-in / copy-out;
//   * per-bucket device timing (CUDA events) feeds the gradient-sync profiler.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

struct dpt_comm;

namespace dptr {

// Greedy size-capped bucket assignment, same contract as
// torch.distributed._compute_bucket_assignment_by_size (reducer.hpp:586-598) for dense
// same-device tensors: tensors are appended in order; when the accumulated bytes reach the
// current limit the bucket is closed and the next limit (if any) becomes current.
// `dtype_key[i]` separates tensors that may not share a bucket.
std::vector<std::vector<int64_t>> compute_bucket_assignment(const std::vector<int64_t>& nbytes,
                                                            const std::vector<int64_t>& dtype_key,
                                                            const std::vector<int64_t>& limits);

struct BucketSpec {
    int64_t offset;  // first element (fp32) of the bucket inside the gradient arena
    int64_t numel;   // padded to a multiple of 4
    std::vector<int> params;
};

struct SgdFusion {
    bool enabled = false;
    std::vector<uintptr_t> param_ptrs;   // per rank: fp32 master arena base
    std::vector<uintptr_t> shadow_ptrs;  // per rank: bf16 shadow arena base
    uintptr_t momentum = 0;              // local fp32 momentum arena base
    float lr = 0, mu = 0, wd = 0;
    int nesterov = 0;
    // NVLS: multicast mappings of the param / shadow arenas (0 = not available)
    uintptr_t param_mc = 0, shadow_mc = 0;
};

struct BucketTiming {
    float ready_ms;  // bwd start -> bucket ready (compute stream)
    float start_ms;  // bwd start -> all-reduce kernel start (comm stream)
    float end_ms;    // bwd start -> all-reduce kernel end
    int64_t bytes;
};

class Reducer {
  public:
    Reducer(std::vector<BucketSpec> buckets, int num_params, uintptr_t comm, int world,
            std::vector<uintptr_t> grad_ptrs, std::vector<uintptr_t> found_inf_ptrs,
            uintptr_t inv_scale_ptr, int ctas, bool profile);
    ~Reducer();

    void set_fusion(const SgdFusion& f) { fusion_ = f; }
    void set_lr(float lr) { fusion_.lr = lr; }
    void set_scale_options(uintptr_t inv_scale_ptr, bool check_finite) {
        inv_scale_ = inv_scale_ptr;
        check_finite_ = check_finite;
    }
    void set_enabled(bool on) { enabled_ = on; }  // no_sync()
    void set_rank(int rank) { rank_ = rank; }
    // NVLS: base of this rank's multicast mapping of the gradient arena; buckets of at least
    // `min_bytes` use the multimem kernel (unless the fused-SGD epilogue is active).
    void set_nvls(uintptr_t mc_grad_base, int64_t min_bytes, int rank, bool with_fused_sgd) {
        nvls_mc_ = mc_grad_base;
        nvls_min_bytes_ = min_bytes;
        rank_ = rank;
        nvls_with_fusion_ = with_fused_sgd;
    }

    // Fused-SGD two-shot buckets publish through the NVSwitch multicast mappings in the SgdFusion
    // (param_mc / shadow_mc) instead of per-peer stores.  Experimental.
    void set_mc_push(bool on) { mc_push_ = on; }

    // One-shot push all-reduce (dpt_allreduce_oneshot_push) for buckets of at most `max_bytes`:
    // `staging_ptrs[r]` = rank r's staging area of 2 * world * cap_elems floats.
    void set_oneshot(std::vector<uintptr_t> staging_ptrs, int64_t cap_elems, int64_t max_bytes,
                     int rank, int ctas) {
        os_staging_ = std::move(staging_ptrs);
        os_cap_ = cap_elems;
        os_max_bytes_ = max_bytes;
        rank_ = rank;
        os_ctas_ = ctas;
    }
    // Measurement aid: launch every bucket as if this were the only rank (same gradient scaling
    // and, with the fused optimizer, a local SGD update of the whole bucket) — the step without
    // any cross-GPU traffic or waiting, on the same GPU, under the same graph.  The difference to
    // the real step is the exposed gradient-sync time.
    void set_local_only(bool on) { local_only_ = on; }
    // closing barrier only in the last bucket kernel of a step (default on)
    void set_relaxed_barriers(bool on) { relaxed_barriers_ = on; }
    // 0 two-shot, 1 one-shot push, 2 NVLS, 3 two-shot + multicast publish (per bucket, as planned)
    std::vector<int> bucket_algos() const;

    void prepare_for_backward(uintptr_t compute_stream);
    void mark_ready(int param_index, uintptr_t compute_stream);
    // Makes `compute_stream` wait for every launched bucket.  Returns the number launched.
    int finalize(uintptr_t compute_stream);
    bool all_launched() const { return next_bucket_ == static_cast<int>(buckets_.size()); }

    std::vector<BucketTiming> timings();  // synchronises; valid when profile == true
    float backward_ms();                  // bwd start -> finalize on the compute stream
    int64_t launches() const { return launches_; }
    uintptr_t comm_stream() const { return reinterpret_cast<uintptr_t>(comm_stream_); }
    const std::vector<BucketSpec>& buckets() const { return buckets_; }

  private:
    void launch_bucket(int b, cudaStream_t compute);

    std::vector<BucketSpec> buckets_;
    std::vector<int> param_bucket_;
    std::vector<int> pending_;
    std::vector<char> ready_;
    std::vector<char> param_ready_;
    int next_bucket_ = 0;
    dpt_comm* comm_;
    int world_;
    std::vector<uintptr_t> grad_ptrs_, found_inf_ptrs_;
    uintptr_t inv_scale_;
    bool check_finite_ = false;
    int ctas_;
    bool profile_;
    bool enabled_ = true;
    bool armed_ = false;
    SgdFusion fusion_;
    cudaStream_t comm_stream_ = nullptr;
    std::vector<cudaEvent_t> ev_ready_, ev_start_, ev_done_;
    cudaEvent_t ev_bwd_start_ = nullptr, ev_bwd_end_ = nullptr;
    int64_t launches_ = 0;
    uintptr_t nvls_mc_ = 0;
    int64_t nvls_min_bytes_ = 0;
    int rank_ = 0;
    bool nvls_with_fusion_ = false;
    bool mc_push_ = false;
    std::vector<uintptr_t> os_staging_;
    int64_t os_cap_ = 0, os_max_bytes_ = 0;
    int os_ctas_ = 0;
    bool local_only_ = false;
    bool relaxed_barriers_ = true;
    int algo_for(const BucketSpec& bk) const;
};

}  // namespace dptr
