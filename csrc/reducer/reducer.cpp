#include "reducer.hpp"

#include <algorithm>
#include <map>
#include <stdexcept>

#include "../../include/dptb200.h"

namespace dptr {

#define DPTR_CUDA(x)                                                                      \
    do {                                                                                  \
        cudaError_t e_ = (x);                                                             \
        if (e_ != cudaSuccess)                                                            \
            throw std::runtime_error(std::string("CUDA error in reducer: ") +             \
                                     cudaGetErrorString(e_));                             \
    } while (0)

std::vector<std::vector<int64_t>> compute_bucket_assignment(const std::vector<int64_t>& nbytes,
                                                            const std::vector<int64_t>& dtype_key,
                                                            const std::vector<int64_t>& limits) {
    if (limits.empty()) throw std::invalid_argument("bucket size limits must not be empty");
    struct Acc {
        std::vector<int64_t> idx;
        int64_t size = 0;
        size_t limit_pos = 0;
    };
    std::map<int64_t, Acc> open;  // per dtype/device key
    std::vector<std::vector<int64_t>> out;
    for (size_t i = 0; i < nbytes.size(); ++i) {
        Acc& a = open[dtype_key.empty() ? 0 : dtype_key[i]];
        a.idx.push_back(static_cast<int64_t>(i));
        a.size += nbytes[i];
        const int64_t limit = limits[std::min(a.limit_pos, limits.size() - 1)];
        if (a.size >= limit) {
            out.push_back(std::move(a.idx));
            a.idx.clear();
            a.size = 0;
            if (a.limit_pos + 1 < limits.size()) ++a.limit_pos;
        }
    }
    for (auto& kv : open)
        if (!kv.second.idx.empty()) out.push_back(std::move(kv.second.idx));
    // torch orders the result by the smallest tensor index of each bucket
    std::sort(out.begin(), out.end(), [](const std::vector<int64_t>& a, const std::vector<int64_t>& b) {
        return *std::min_element(a.begin(), a.end()) < *std::min_element(b.begin(), b.end());
    });
    return out;
}

Reducer::Reducer(std::vector<BucketSpec> buckets, int num_params, uintptr_t comm, int world,
                 std::vector<uintptr_t> grad_ptrs, std::vector<uintptr_t> found_inf_ptrs,
                 uintptr_t inv_scale_ptr, int ctas, bool profile)
    : buckets_(std::move(buckets)),
      param_bucket_(num_params, -1),
      comm_(reinterpret_cast<dpt_comm*>(comm)),
      world_(world),
      grad_ptrs_(std::move(grad_ptrs)),
      found_inf_ptrs_(std::move(found_inf_ptrs)),
      inv_scale_(inv_scale_ptr),
      ctas_(ctas),
      profile_(profile) {
    for (size_t b = 0; b < buckets_.size(); ++b)
        for (int p : buckets_[b].params) {
            if (p < 0 || p >= num_params) throw std::invalid_argument("bad parameter index");
            param_bucket_[p] = static_cast<int>(b);
        }
    pending_.resize(buckets_.size());
    ready_.resize(buckets_.size());
    param_ready_.resize(num_params);
    int lo = 0, hi = 0;
    DPTR_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    DPTR_CUDA(cudaStreamCreateWithPriority(&comm_stream_, cudaStreamNonBlocking, hi));
    const unsigned flags = profile_ ? cudaEventDefault : cudaEventDisableTiming;
    for (size_t b = 0; b < buckets_.size(); ++b) {
        cudaEvent_t a, s, d;
        DPTR_CUDA(cudaEventCreateWithFlags(&a, flags));
        DPTR_CUDA(cudaEventCreateWithFlags(&s, flags));
        DPTR_CUDA(cudaEventCreateWithFlags(&d, flags));
        ev_ready_.push_back(a);
        ev_start_.push_back(s);
        ev_done_.push_back(d);
    }
    DPTR_CUDA(cudaEventCreateWithFlags(&ev_bwd_start_, flags));
    DPTR_CUDA(cudaEventCreateWithFlags(&ev_bwd_end_, flags));
}

Reducer::~Reducer() {
    for (auto e : ev_ready_) cudaEventDestroy(e);
    for (auto e : ev_start_) cudaEventDestroy(e);
    for (auto e : ev_done_) cudaEventDestroy(e);
    if (ev_bwd_start_) cudaEventDestroy(ev_bwd_start_);
    if (ev_bwd_end_) cudaEventDestroy(ev_bwd_end_);
    if (comm_stream_) cudaStreamDestroy(comm_stream_);
}

void Reducer::prepare_for_backward(uintptr_t compute_stream) {
    for (size_t b = 0; b < buckets_.size(); ++b) {
        pending_[b] = static_cast<int>(buckets_[b].params.size());
        ready_[b] = 0;
    }
    std::fill(param_ready_.begin(), param_ready_.end(), 0);
    next_bucket_ = 0;
    armed_ = true;
    if (profile_)
        DPTR_CUDA(cudaEventRecord(ev_bwd_start_, reinterpret_cast<cudaStream_t>(compute_stream)));
}

void Reducer::mark_ready(int p, uintptr_t compute_stream) {
    if (!enabled_ || !armed_) return;
    if (p < 0 || p >= static_cast<int>(param_bucket_.size()) || param_bucket_[p] < 0) return;
    if (param_ready_[p])
        throw std::runtime_error("reducer: parameter " + std::to_string(p) +
                                 " marked ready twice in one backward pass");
    param_ready_[p] = 1;
    const int b = param_bucket_[p];
    if (--pending_[b] == 0) {
        ready_[b] = 1;
        cudaStream_t cs = reinterpret_cast<cudaStream_t>(compute_stream);
        // launch strictly in bucket order (every rank must issue the same sequence)
        while (next_bucket_ < static_cast<int>(buckets_.size()) && ready_[next_bucket_]) {
            launch_bucket(next_bucket_, cs);
            ++next_bucket_;
        }
    }
}

int Reducer::algo_for(const BucketSpec& bk) const {
    const bool fuse = fusion_.enabled && world_ > 1;
    if (world_ > 1 && os_cap_ > 0 && bk.numel <= os_cap_ && bk.numel * 4 <= os_max_bytes_) return 1;
    // measured at 8 GPUs: with the SGD update fused, the two-shot kernel (0.855 ms/step) beats the
    // NVLS one (0.871); without fusion NVLS wins (0.882 vs 0.926) -> NVLS+fusion only on request
    const bool nvls_fuse_ok =
        !fuse || (nvls_with_fusion_ && fusion_.param_mc != 0 && fusion_.shadow_mc != 0);
    if (nvls_mc_ != 0 && nvls_fuse_ok && world_ > 1 && bk.numel * 4 >= nvls_min_bytes_) return 2;
    if (fuse && mc_push_ && fusion_.param_mc != 0 && fusion_.shadow_mc != 0) return 3;
    return 0;
}

std::vector<int> Reducer::bucket_algos() const {
    std::vector<int> out;
    for (const auto& bk : buckets_) out.push_back(algo_for(bk));
    return out;
}

void Reducer::launch_bucket(int b, cudaStream_t compute) {
    const BucketSpec& bk = buckets_[b];
    DPTR_CUDA(cudaEventRecord(ev_ready_[b], compute));
    DPTR_CUDA(cudaStreamWaitEvent(comm_stream_, ev_ready_[b], 0));
    if (profile_) DPTR_CUDA(cudaEventRecord(ev_start_[b], comm_stream_));
    const int algo = algo_for(bk);
    if (local_only_) {
        // single-rank stand-in (see set_local_only): same math, no peer traffic
        const int me0 = rank_;
        void* one[8] = {reinterpret_cast<void*>(grad_ptrs_[me0] + bk.offset * 4)};
        int rc0 = 0;
        if (fusion_.enabled) {
            rc0 = dpt_sgd_flat(reinterpret_cast<float*>(fusion_.param_ptrs[me0] + bk.offset * 4),
                               reinterpret_cast<const float*>(grad_ptrs_[me0] + bk.offset * 4),
                               reinterpret_cast<float*>(fusion_.momentum + bk.offset * 4),
                               reinterpret_cast<void*>(fusion_.shadow_ptrs[me0] + bk.offset * 2),
                               bk.numel, fusion_.lr, fusion_.mu, 0.f, fusion_.wd, fusion_.nesterov,
                               nullptr, nullptr, comm_stream_);
        } else {
            (void)one;
        }
        if (rc0 != 0) throw std::runtime_error("local-only bucket update failed: " + std::to_string(rc0));
        DPTR_CUDA(cudaEventRecord(ev_done_[b], comm_stream_));
        ++launches_;
        return;
    }
    if (algo == 1) {
        const int me = rank_;
        void* stg[8] = {nullptr};
        for (int r = 0; r < world_; ++r) stg[r] = reinterpret_cast<void*>(os_staging_[r]);
        const bool fuse1 = fusion_.enabled;
        int rc1 = dpt_allreduce_oneshot_push(
            comm_, reinterpret_cast<float*>(grad_ptrs_[me] + bk.offset * 4), stg, os_cap_, bk.numel,
            1.0f / static_cast<float>(world_), reinterpret_cast<const float*>(inv_scale_),
            (check_finite_ && !found_inf_ptrs_.empty())
                ? reinterpret_cast<float*>(found_inf_ptrs_[me])
                : nullptr,
            fuse1 ? reinterpret_cast<float*>(fusion_.param_ptrs[me] + bk.offset * 4) : nullptr,
            fuse1 ? reinterpret_cast<void*>(fusion_.shadow_ptrs[me] + bk.offset * 2) : nullptr,
            fuse1 ? reinterpret_cast<float*>(fusion_.momentum + bk.offset * 4) : nullptr, fusion_.lr,
            fusion_.mu, fusion_.wd, fusion_.nesterov, os_ctas_, comm_stream_);
        if (rc1 != 0)
            throw std::runtime_error("dpt_allreduce_oneshot_push failed with code " +
                                     std::to_string(rc1));
        DPTR_CUDA(cudaEventRecord(ev_done_[b], comm_stream_));
        ++launches_;
        return;
    }
    void* bufs[8] = {nullptr};
    void* finf[8] = {nullptr};
    void* prm[8] = {nullptr};
    void* shd[8] = {nullptr};
    for (int r = 0; r < world_; ++r) {
        bufs[r] = reinterpret_cast<void*>(grad_ptrs_[r] + bk.offset * 4);
        if (check_finite_ && !found_inf_ptrs_.empty())
            finf[r] = reinterpret_cast<void*>(found_inf_ptrs_[r]);
        if (fusion_.enabled) {
            prm[r] = reinterpret_cast<void*>(fusion_.param_ptrs[r] + bk.offset * 4);
            shd[r] = reinterpret_cast<void*>(fusion_.shadow_ptrs[r] + bk.offset * 2);
        }
    }
    const bool fuse = fusion_.enabled && world_ > 1;
    // every bucket kernel but the step's last leaves without its closing barrier: the next
    // bucket's opening barrier covers its peer stores (dptb200.h)
    if (relaxed_barriers_ && world_ > 1 && b + 1 < static_cast<int>(buckets_.size()))
        dpt_comm_skip_closing_barrier_next();
    if (algo == 2) {
        const int me = rank_;
        int rc2 = dpt_allreduce_nvls_sgd(
            comm_, reinterpret_cast<void*>(nvls_mc_ + bk.offset * 4), bk.numel,
            1.0f / static_cast<float>(world_), reinterpret_cast<const float*>(inv_scale_),
            (check_finite_ && !found_inf_ptrs_.empty()) ? finf : nullptr,
            fuse ? reinterpret_cast<float*>(fusion_.param_ptrs[me] + bk.offset * 4) : nullptr,
            fuse ? reinterpret_cast<float*>(fusion_.momentum + bk.offset * 4) : nullptr,
            fuse ? reinterpret_cast<void*>(fusion_.param_mc + bk.offset * 4) : nullptr,
            fuse ? reinterpret_cast<void*>(fusion_.shadow_mc + bk.offset * 2) : nullptr, fusion_.lr,
            fusion_.mu, fusion_.wd, fusion_.nesterov, ctas_, comm_stream_);
        if (rc2 != 0)
            throw std::runtime_error("dpt_allreduce_nvls failed with code " + std::to_string(rc2));
        DPTR_CUDA(cudaEventRecord(ev_done_[b], comm_stream_));
        ++launches_;
        return;
    }
    if (algo == 3) {
        int rc3 = dpt_allreduce_fused_mcpush(
            comm_, bufs, bk.numel, 1.0f / static_cast<float>(world_),
            reinterpret_cast<const float*>(inv_scale_),
            (check_finite_ && !found_inf_ptrs_.empty()) ? finf : nullptr, prm, shd,
            reinterpret_cast<float*>(fusion_.momentum + bk.offset * 4),
            reinterpret_cast<void*>(fusion_.param_mc + bk.offset * 4),
            reinterpret_cast<void*>(fusion_.shadow_mc + bk.offset * 2), fusion_.lr, fusion_.mu,
            fusion_.wd, fusion_.nesterov, ctas_, comm_stream_);
        if (rc3 != 0)
            throw std::runtime_error("dpt_allreduce_fused_mcpush failed with code " +
                                     std::to_string(rc3));
        DPTR_CUDA(cudaEventRecord(ev_done_[b], comm_stream_));
        ++launches_;
        return;
    }
    int rc = dpt_allreduce_fused(
        comm_, bufs, bk.numel, 1.0f / static_cast<float>(world_),
        reinterpret_cast<const float*>(inv_scale_),
        (check_finite_ && !found_inf_ptrs_.empty()) ? finf : nullptr, fuse ? prm : nullptr,
        fuse ? shd : nullptr,
        fuse ? reinterpret_cast<float*>(fusion_.momentum + bk.offset * 4) : nullptr, fusion_.lr,
        fusion_.mu, fusion_.wd, fusion_.nesterov, ctas_, comm_stream_);
    if (rc != 0) throw std::runtime_error("dpt_allreduce_fused failed with code " + std::to_string(rc));
    DPTR_CUDA(cudaEventRecord(ev_done_[b], comm_stream_));
    ++launches_;
}

int Reducer::finalize(uintptr_t compute_stream) {
    if (!armed_) return 0;
    cudaStream_t cs = reinterpret_cast<cudaStream_t>(compute_stream);
    if (enabled_ && next_bucket_ != static_cast<int>(buckets_.size())) {
        // Some parameter never reported a gradient (unused in this iteration).  The collective
        // sequence must stay identical on every rank, so flush the remaining buckets now.
        while (next_bucket_ < static_cast<int>(buckets_.size())) {
            launch_bucket(next_bucket_, cs);
            ++next_bucket_;
        }
    }
    const int n = enabled_ ? next_bucket_ : 0;
    for (int b = 0; b < n; ++b) DPTR_CUDA(cudaStreamWaitEvent(cs, ev_done_[b], 0));
    if (profile_) DPTR_CUDA(cudaEventRecord(ev_bwd_end_, cs));
    armed_ = false;
    return n;
}

std::vector<BucketTiming> Reducer::timings() {
    std::vector<BucketTiming> out;
    if (!profile_) return out;
    DPTR_CUDA(cudaEventSynchronize(ev_bwd_end_));
    for (size_t b = 0; b < buckets_.size(); ++b) {
        BucketTiming t{0, 0, 0, buckets_[b].numel * 4};
        if (cudaEventElapsedTime(&t.ready_ms, ev_bwd_start_, ev_ready_[b]) != cudaSuccess) continue;
        cudaEventElapsedTime(&t.start_ms, ev_bwd_start_, ev_start_[b]);
        cudaEventElapsedTime(&t.end_ms, ev_bwd_start_, ev_done_[b]);
        out.push_back(t);
    }
    cudaGetLastError();
    return out;
}

float Reducer::backward_ms() {
    if (!profile_) return 0.f;
    float ms = 0.f;
    DPTR_CUDA(cudaEventSynchronize(ev_bwd_end_));
    DPTR_CUDA(cudaEventElapsedTime(&ms, ev_bwd_start_, ev_bwd_end_));
    return ms;
}

}  // namespace dptr
