"""Bucket planning (thin Python face of the C++ planner in csrc/reducer/reducer.cpp)."""
from __future__ import annotations

import importlib.util
from pathlib import Path

_EXT = None
_EXT_PATH = Path(__file__).resolve().parent.parent / "_native" / "_dptb200_C.so"

DEFAULT_FIRST_BUCKET_BYTES = 1024 * 1024          # torch reducer.hpp:30
DEFAULT_BUCKET_BYTES_CAP = 25 * 1024 * 1024       # torch reducer.hpp:31


def ext():
    """The compiled C++ runtime (reducer + planner)."""
    global _EXT
    if _EXT is None:
        if not _EXT_PATH.exists():
            raise RuntimeError(f"{_EXT_PATH} is missing: run `python build_native.py`")
        import torch  # noqa: F401  (libtorch must be loaded first)
        spec = importlib.util.spec_from_file_location("_dptb200_C", str(_EXT_PATH))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _EXT = mod
    return _EXT


def ext_available() -> bool:
    return _EXT_PATH.exists()


def compute_bucket_assignment_by_size(tensors, bucket_size_limits):
    """Same result as ``torch.distributed._compute_bucket_assignment_by_size`` for dense tensors:
    returns ``(bucket_indices, per_bucket_size_limits)``."""
    nbytes = [t.numel() * t.element_size() for t in tensors]
    keys = []
    seen: dict = {}
    for t in tensors:
        k = (str(t.dtype), str(t.device))
        keys.append(seen.setdefault(k, len(seen)))
    limits = [int(x) for x in bucket_size_limits]
    if ext_available():
        buckets = ext().compute_bucket_assignment(nbytes, keys, limits)
    else:  # pure-python mirror (identical algorithm) for environments without the build
        buckets = _py_assignment(nbytes, keys, limits)
    per_limits = []
    pos: dict = {}
    order = sorted(range(len(buckets)), key=lambda i: min(buckets[i]))
    # recover which limit each bucket was closed under (first bucket of a key -> limits[0], ...)
    for i in order:
        k = keys[buckets[i][0]]
        j = pos.get(k, 0)
        per_limits.append(limits[min(j, len(limits) - 1)])
        pos[k] = j + 1
    return [list(b) for b in buckets], per_limits


def _py_assignment(nbytes, keys, limits):
    open_: dict = {}
    out = []
    for i, nb in enumerate(nbytes):
        a = open_.setdefault(keys[i], {"idx": [], "size": 0, "pos": 0})
        a["idx"].append(i)
        a["size"] += nb
        if a["size"] >= limits[min(a["pos"], len(limits) - 1)]:
            out.append(a["idx"])
            a["idx"], a["size"] = [], 0
            if a["pos"] + 1 < len(limits):
                a["pos"] += 1
    for a in open_.values():
        if a["idx"]:
            out.append(a["idx"])
    out.sort(key=min)
    return out


def split_tail(idx_buckets, nbytes, tail_bucket_bytes):
    """Peel the last-ready parameters off the final bucket into a bucket of their own (at most
    ``tail_bucket_bytes``).  The final bucket closes only after the very last weight gradient of
    backward (the stem), so everything in it is exposed; a tail of a few tens of KB goes through
    the low-latency one-shot kernel while the rest of the former last bucket is reduced under the
    stem's backward kernels."""
    if tail_bucket_bytes <= 0 or not idx_buckets or len(idx_buckets[-1]) < 2:
        return idx_buckets
    last = list(idx_buckets[-1])
    if sum(nbytes[i] for i in last) <= tail_bucket_bytes:
        return idx_buckets          # the whole last bucket already is a small tail
    size, cut = 0, len(last)
    while cut > 1 and size + nbytes[last[cut - 1]] <= tail_bucket_bytes:
        size += nbytes[last[cut - 1]]
        cut -= 1
    if cut == len(last):
        return idx_buckets
    return [list(b) for b in idx_buckets[:-1]] + [last[:cut], last[cut:]]


def plan_arena_buckets(arena, bucket_cap_bytes=DEFAULT_BUCKET_BYTES_CAP,
                       first_bucket_bytes=DEFAULT_FIRST_BUCKET_BYTES, tail_bucket_bytes=0):
    """Buckets over the arena order (= gradient-ready order).  Each bucket is a contiguous,
    16-byte aligned slice ``[offset, offset + numel)`` of the flat gradient arena.
    Returns a list of dicts {offset, numel, params (indices into arena.order), bytes}.
    ``tail_bucket_bytes`` > 0 additionally splits a small last bucket off (see split_tail)."""
    params = arena.order
    idx_buckets, _ = compute_bucket_assignment_by_size(
        [p for p in params], [first_bucket_bytes, bucket_cap_bytes])
    idx_buckets = split_tail(idx_buckets, [p.numel() * 4 for p in params], tail_bucket_bytes)
    plans = []
    for b in idx_buckets:
        first, last = params[b[0]], params[b[-1]]
        start = arena.slot(first).offset
        end = arena.slot(last).offset + (arena.slot(last).numel + 7) // 8 * 8
        plans.append({"offset": start, "numel": end - start, "params": list(b),
                      "bytes": sum(params[i].numel() * 4 for i in b)})
    return plans
