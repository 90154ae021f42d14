"""Measured crossovers behind the per-bucket algorithm choice of the native gradient path.

NCCL picks ring / tree / NVLS and its protocol per message from built-in tuning tables
(torch DDP hands it one bucket at a time, reference train_ddp.py:305-310).  The native backend
chooses between its own kernels the same way; the numbers below are MEASUREMENTS on 8xB200
(NVLink 5 / NVSwitch), each with the file that holds the raw data, so that the defaults in
``DistributedDataParallel`` are traceable and can be re-derived (``allreduce_algo="autotune"``
re-measures the two-shot / NVLS crossover at construction time on the machine at hand).

  kernel                when                                  evidence
  --------------------  ------------------------------------  ------------------------------------------
  one-shot push         bucket <= 256 KB (any N)              38 KB tail: 8-9 us at N=2, 20 us at N=8 of
                                                              which 19 us is rank skew (profiles/r2/
                                                              step_trace_*.json); 9.4 MB: 261 us vs 64 us
                                                              two-shot at N=2 -> only for small buckets
  two-shot              fused SGD update, N = 2..3            N=2 step 0.709 ms (step_trace_2_c16.json)
  two-shot + multicast  fused SGD update, N >= 4, multicast   N=8 step 0.712 vs 0.728 ms with per-peer
    publish             bound arenas                          stores (step_trace_8_mc16 / _c16.json)
  NVLS (multimem)       no fused update, N >= 4, >= 1 MB      all-reduce sweep: NVLS beats two-shot from
                                                              ~1 MB (profiles/sweep8_nvls.txt); at N=2 it
                                                              is slower (profiles/sweep2_nvls.json)
  comm CTAs per bucket  16                                    N=8: 0.730 ms vs 0.753 with 32
                                                              (profiles/r2/g8_bench8*.json)
  bucket cap            4 MiB (+ 1 MiB first, 64 KB tail)     launch cost ~15 us per bucket kernel, exposed
                                                              tail shrinks with the last bucket
"""
from __future__ import annotations

ONESHOT_MAX_BYTES = 256 * 1024        # buckets up to this size: one-shot push kernel
TAIL_BUCKET_BYTES = 64 * 1024         # last-ready parameters peeled into their own bucket
NVLS_MIN_WORLD = 4                    # in-switch reduction pays from 4 GPUs
NVLS_MIN_BYTES = 1 << 20              # ... and from ~1 MB per bucket
MC_PUBLISH_MIN_WORLD = 4              # one multimem.st instead of N peer stores
NATIVE_BUCKET_CAP_MB = 4
TORCH_BUCKET_CAP_MB = 25              # torch reducer.hpp:31
COMM_CTAS = 16


def choose_algorithm(nbytes: int, world: int, fused_sgd: bool, multicast: bool,
                     oneshot_max_bytes: int = ONESHOT_MAX_BYTES,
                     nvls_min_bytes: int = NVLS_MIN_BYTES) -> str:
    """The kernel ``allreduce_algo="auto"`` uses for one bucket — the Python mirror of
    ``Reducer::algo_for`` (csrc/reducer/reducer.cpp), kept in step with it by a test."""
    if world < 2:
        return "none"
    if 0 < nbytes <= oneshot_max_bytes:
        return "oneshot"
    if multicast and world >= NVLS_MIN_WORLD:
        if not fused_sgd and nbytes >= nvls_min_bytes:
            return "nvls"
        if fused_sgd and world >= MC_PUBLISH_MIN_WORLD:
            return "twoshot_mc"
    return "twoshot"
