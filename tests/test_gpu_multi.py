"""Multi-GPU tests (one process per GPU, >= 2 B200s): NVLink peer-memory collectives vs
reference sums, fused epilogues, the native DDP path vs torch DDP + NCCL, TrainStep under DDP."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

needs2 = pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2,
                            reason="needs >= 2 GPUs")


def _world():
    """Ranks for the kernel-level tests: every GPU of the box (DPT_TEST_WORLD caps it)."""
    return min(torch.cuda.device_count(), int(os.environ.get("DPT_TEST_WORLD", "8")))


def _model_world():
    """Ranks for the tests that train whole models through every algorithm variant: 2 by
    default (what they were validated on: profiles/r2/multi_gpu_tests_2gpu_*.log; the 8-GPU
    paths are checked for exactness by bench.py's multi_gpu_checks at every N of the scaling
    run); DPT_TEST_WORLD_MODEL=8 runs them on the whole box."""
    return min(torch.cuda.device_count(), int(os.environ.get("DPT_TEST_WORLD_MODEL", "2")))


def _init(rank, world, port):
    import torch.distributed as dist
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
                       "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank)})
    sys.path.insert(0, ROOT)
    torch.cuda.set_device(rank)
    import datetime
    # a rank that fails an assertion leaves the others inside a collective: keep that wait short
    dist.init_process_group("nccl", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank),
                            timeout=datetime.timedelta(seconds=120))


def _run(worker, world=None):
    import torch.multiprocessing as mp
    from dpt_b200.launcher import free_port
    world = world or _world()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_guard, args=(worker, r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    import queue as _queue
    import time
    results = []
    deadline = time.time() + 600          # whole test
    failed_at = None
    while len(results) < world and time.time() < deadline:
        try:
            results.append(q.get(timeout=2.0))
        except _queue.Empty:
            if all(not p.is_alive() for p in procs) and q.empty():
                break
            # once one rank has failed the others are usually stuck in a collective with it:
            # give them 30 s, then stop the test instead of sitting out a communicator timeout
            if failed_at is not None and time.time() - failed_at > 30:
                break
            continue
        if results[-1][1] != "ok" and failed_at is None:
            failed_at = time.time()
    for p in procs:
        p.join(5 if failed_at is not None else 60)
    for p in procs:
        if p.is_alive():
            p.kill()
    bad = [(r, m) for r, m in results if m != "ok"]
    assert not bad, "\n".join(f"rank {r}:\n{m}" for r, m in bad[:2])
    assert len(results) == world, f"only {len(results)} of {world} ranks reported (hung or died)"


def _rank_mean(values):
    """Element-wise mean over ranks of a list of floats (collective)."""
    import torch.distributed as dist
    t = torch.tensor(values, device="cuda", dtype=torch.float64)
    dist.all_reduce(t)
    return (t / dist.get_world_size()).tolist()


def _agree(ok: bool, msg: str = "") -> None:
    """Collective assertion: every rank calls it with its local verdict; if ANY rank failed, all
    ranks raise.  Conditions on rank-local data (each rank trains on its own batch) must go
    through this — a plain assert that fails on one rank only leaves the others waiting for it
    in the next collective."""
    import torch.distributed as dist
    flag = torch.tensor([0 if ok else 1], device="cuda", dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MAX)
    if int(flag.item()) != 0:
        raise AssertionError(msg if not ok else "another rank failed: " + msg.split(":")[0])


def _guard(worker, rank, world, port, q):
    import torch.distributed as dist
    try:
        _init(rank, world, port)
        worker(rank, world)
        torch.cuda.synchronize()
        q.put((rank, "ok"))
    except Exception:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        try:
            if dist.is_initialized():
                dist.destroy_process_group()
        except Exception:  # noqa: BLE001
            pass


# --------------------------------------------------------------------------------- collectives
def _collectives_worker(rank, world):
    import torch.distributed as dist
    from dpt_b200.parallel.comm import Communicator
    comm = Communicator(timeout_s=20.0)
    dev = torch.device("cuda", rank)
    sizes = [4, 64, 1024, 4100, 65536, 1 << 20, (1 << 22) + 12, 11181648]
    buf = comm.alloc(max(sizes) * 4)
    for n in sizes:
        for rep in range(2):
            g = torch.Generator(device=dev).manual_seed(1000 * rank + n % 977 + rep)
            x = torch.randn(n, device=dev, generator=g)
            view = buf.view(torch.float32, 0, n)
            view.copy_(x)
            ref = x.clone()
            dist.all_reduce(ref)                     # NCCL reference sum
            torch.cuda.synchronize()
            comm.allreduce_(buf, 0, n, scale=1.0 / world)
            torch.cuda.synchronize()
            err = (view - ref / world).abs().max().item()
            assert err <= 1e-5 * max(1.0, ref.abs().max().item()), (n, rep, err)
            # bitwise identical on every rank (fixed summation order)
            lo, hi = view.clone(), view.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            assert torch.equal(lo, hi), n
    # one-shot (out of place)
    for n in (4, 1024, 262144):
        x = torch.full((n,), float(rank + 1), device=dev)
        buf.view(torch.float32, 0, n).copy_(x)
        out = torch.empty(n, device=dev)
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce_oneshot(buf, 0, out, scale=1.0)
        torch.cuda.synchronize()
        assert torch.allclose(out, torch.full_like(out, world * (world + 1) / 2)), n
    # broadcast from every source, odd sizes of 16-byte words
    for src in range(world):
        n = 9600 + 16 * src
        v = buf.view(torch.float32, 0, n)
        v.fill_(float(rank))
        torch.cuda.synchronize()
        dist.barrier()
        comm.broadcast_(buf, 0, n * 4, src=src)
        torch.cuda.synchronize()
        assert torch.all(v == float(src)), src
    # repeated barriers (epoch counters keep increasing, no resets)
    for _ in range(200):
        comm.barrier()
    torch.cuda.synchronize()
    comm.check_health()
    # fused epilogue: scale * inv_scale, inf detection visible on every rank
    n = 8192
    view = buf.view(torch.float32, 0, n)
    view.fill_(2.0)
    inv = torch.tensor(0.25, device=dev)
    comm.found_inf.zero_()
    torch.cuda.synchronize()
    dist.barrier()
    comm.allreduce_(buf, 0, n, inv_scale=inv, check_finite=True)
    torch.cuda.synchronize()
    assert torch.allclose(view, torch.full_like(view, 2.0 * 0.25)) and comm.found_inf.item() == 0.0
    view.fill_(1.0)
    if rank == world - 1:
        view[n - 3] = float("inf")
    torch.cuda.synchronize()
    dist.barrier()
    comm.allreduce_(buf, 0, n, inv_scale=inv, check_finite=True)
    torch.cuda.synchronize()
    dist.barrier()
    assert comm.found_inf.item() == 1.0, "non-finite gradient must be flagged on every rank"
    comm.close()


@needs2
def test_peer_memory_collectives():
    _run(_collectives_worker)


# --------------------------------------------------------------------------------- one-shot push
def _oneshot_push_worker(rank, world):
    """dpt_allreduce_oneshot_push vs the NCCL sum: in place (F1 epilogue incl. inv_scale and the
    non-finite flag), back-to-back calls with nothing in between (the double-buffered staging
    area), and the replicated fused SGD update vs all-reduce -> torch.optim.SGD."""
    import torch.distributed as dist
    from dpt_b200.parallel.comm import Communicator
    comm = Communicator(timeout_s=20.0)
    dev = torch.device("cuda", rank)
    cap = 65536
    staging = comm.alloc(2 * world * cap * 4)
    for n in (4, 1024, 9536, 65536):
        for rep in range(5):
            x = torch.randn(n, device=dev,
                            generator=torch.Generator(device=dev).manual_seed(977 * rank + n + rep))
            ref = x.clone()
            dist.all_reduce(ref)
            g = x.clone()
            torch.cuda.synchronize()
            comm.allreduce_oneshot_push_(g, staging, cap)
            torch.cuda.synchronize()
            err = (g - ref / world).abs().max().item()
            assert err <= 1e-5 * max(1.0, ref.abs().max().item()), (n, rep, err)
            lo, hi = g.clone(), g.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            assert torch.equal(lo, hi), (n, rep)
    # 50 back-to-back calls, different data every time, no host sync in between
    n = 9536
    outs, refs = [], []
    for it in range(50):
        x = torch.full((n,), float(rank + 1 + it), device=dev)
        outs.append(x)
        refs.append(sum(float(r + 1 + it) for r in range(world)) / world)
    torch.cuda.synchronize()
    dist.barrier()
    for x in outs:
        comm.allreduce_oneshot_push_(x, staging, cap)
    torch.cuda.synchronize()
    for it, (x, r) in enumerate(zip(outs, refs)):
        assert torch.allclose(x, torch.full_like(x, r)), (it, x[:2].tolist(), r)
    # n beyond the staging capacity is refused, not truncated
    big = torch.zeros(cap + 4, device=dev)
    with pytest.raises(RuntimeError):
        comm.allreduce_oneshot_push_(big, staging, cap)
    # inv_scale + non-finite detection (the flag is local: every rank sees the same sums)
    inv = torch.tensor(0.5, device=dev)
    finf = torch.zeros(1, device=dev)
    g = torch.full((1024,), 2.0, device=dev)
    comm.allreduce_oneshot_push_(g, staging, cap, inv_scale=inv, found_inf=finf)
    torch.cuda.synchronize()
    assert torch.allclose(g, torch.full_like(g, 1.0)) and finf.item() == 0.0
    g = torch.ones(1024, device=dev)
    if rank == world - 1:
        g[77] = float("nan")
    comm.allreduce_oneshot_push_(g, staging, cap, inv_scale=inv, found_inf=finf)
    torch.cuda.synchronize()
    assert finf.item() == 1.0
    # replicated fused SGD
    n = 9536
    p = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(5))
    shadow = torch.zeros(n, device=dev, dtype=torch.bfloat16)
    m = torch.zeros(n, device=dev)
    ref_p = torch.nn.Parameter(p.clone())
    opt = torch.optim.SGD([ref_p], lr=0.1, momentum=0.9, weight_decay=5e-4)
    for it in range(4):
        gi = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(100 * rank + it))
        gsum = gi.clone()
        dist.all_reduce(gsum)
        ref_p.grad = gsum / world
        opt.step()
        comm.allreduce_oneshot_push_(gi, staging, cap, sgd={"param": p, "shadow": shadow, "momentum": m,
                                                            "lr": 0.1, "momentum_coef": 0.9,
                                                            "weight_decay": 5e-4})
    torch.cuda.synchronize()
    assert torch.allclose(p, ref_p.data, atol=1e-5, rtol=1e-5)
    assert torch.allclose(m, opt.state[ref_p]["momentum_buffer"], atol=1e-5, rtol=1e-5)
    assert torch.equal(shadow, p.bfloat16())
    lo, hi = p.clone(), p.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert torch.equal(lo, hi)
    comm.check_health()
    comm.validate_epochs()
    comm.close()


@needs2
def test_oneshot_push_allreduce():
    _run(_oneshot_push_worker)


# --------------------------------------------------------------------------------- fused SGD
def _fused_sgd_worker(rank, world):
    import torch.distributed as dist
    from dpt_b200.parallel.comm import Communicator
    comm = Communicator(timeout_s=20.0)
    dev = torch.device("cuda", rank)
    n = 100000
    gbuf, pbuf, sbuf = comm.alloc(n * 4), comm.alloc(n * 4), comm.alloc(n * 2)
    g, p, s = gbuf.view(torch.float32, 0, n), pbuf.view(torch.float32, 0, n), sbuf.view(torch.bfloat16, 0, n)
    m = torch.zeros(n, device=dev)
    p.copy_(torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(5)))
    ref_p = torch.nn.Parameter(p.clone())
    opt = torch.optim.SGD([ref_p], lr=0.1, momentum=0.9, weight_decay=5e-4)
    for it in range(3):
        gi = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(100 * rank + it))
        g.copy_(gi)
        gsum = gi.clone()
        dist.all_reduce(gsum)
        ref_p.grad = gsum / world
        opt.step()
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce_(gbuf, 0, n, sgd={"param": pbuf, "shadow": sbuf, "byte_offset": 0,
                                         "momentum": m.data_ptr(), "lr": 0.1, "momentum_coef": 0.9,
                                         "weight_decay": 5e-4})
        torch.cuda.synchronize()
        dist.barrier()
    assert torch.allclose(p, ref_p.data, atol=1e-5, rtol=1e-5)
    assert torch.equal(s, p.bfloat16())
    lo, hi = p.clone(), p.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert torch.equal(lo, hi)
    comm.close()


@needs2
def test_fused_sgd_allreduce_matches_allreduce_then_sgd():
    _run(_fused_sgd_worker)


# --------------------------------------------------------------------------------- DDP
def _ddp_worker(rank, world):
    import copy
    import torch.distributed as dist
    import torch.nn.functional as F
    from torch.nn.parallel import DistributedDataParallel as TorchDDP
    from dpt_b200.models.resnet import resnet18
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
    dev = torch.device("cuda", rank)
    torch.manual_seed(100 + rank)
    # IEEE fp32 + deterministic cuDNN algorithms on both sides: the two runs then differ only in
    # the gradient reduction order (none at world 2: a + b is commutative and 1/2 is exact) and in
    # the optimizer's rounding (fused multiply-add vs torch's separate foreach ops)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False
    base = resnet18(10).to(dev)
    base.native = "never"        # same ATen model math on both sides: isolates the gradient path
    ours = OurDDP(copy.deepcopy(base), device_ids=[rank], backend="native", profile=True)
    ref = TorchDDP(copy.deepcopy(base), device_ids=[rank])
    assert ours.backend == "native"
    for a, b in zip(ours.module.state_dict().values(), ref.module.state_dict().values()):
        assert torch.equal(a, b), "construction broadcast differs from torch DDP"
    # small lr keeps the comparison well conditioned (TF32 cuDNN kernels are not run-to-run
    # deterministic and lr 0.05 on the random-init stem amplifies that chaotically)
    opt_o = SGD(ours.parameters(), lr=0.002, momentum=0.9, weight_decay=5e-4)
    opt_r = torch.optim.SGD(ref.parameters(), lr=0.002, momentum=0.9, weight_decay=5e-4)
    g = torch.Generator(device=dev).manual_seed(7 + rank)
    init = {n: p.detach().clone() for n, p in ref.module.named_parameters()}   # rank 0's weights
    for it in range(4):
        x = torch.randn(16, 3, 32, 32, device=dev, generator=g)
        t = torch.randint(0, 10, (16,), device=dev, generator=g)
        for model, opt in ((ours, opt_o), (ref, opt_r)):
            opt.zero_grad(set_to_none=True)
            F.cross_entropy(model(x), t).backward()
            if it == 0 and model is ref:
                # averaged gradients of the first step (identical weights): only the reduction
                # order differs — exact at world 2, a few ulp beyond
                torch.cuda.synchronize()
                tol0 = 1e-6 if world == 2 else 2e-5
                for (n, a), b in zip(ours.module.named_parameters(), ref.module.parameters()):
                    err = (a.grad - b.grad).abs().max().item()
                    den = b.grad.abs().max().item() + 1e-12
                    assert err / den <= tol0, f"{n}: averaged grad differs from torch DDP + NCCL ({err/den:.3g})"
            opt.step()
    torch.cuda.synchronize()
    ours.comm.check_health()
    worst = 0.0
    for (n, a), b in zip(ours.module.named_parameters(), ref.module.parameters()):
        # absolute agreement after 4 optimizer steps, measured against the size of the UPDATE
        # (so zero-initialised biases are judged as strictly as the weights)
        upd = (b - init[n]).norm().item()
        rel = (a - b).norm().item() / (upd + 1e-12)
        worst = max(worst, rel)
        assert rel < 2e-2, f"{n}: update differs from torch DDP + NCCL after 4 steps (rel {rel:.3g})"
    if rank == 0:
        print(f"[ddp parity] worst per-tensor error relative to the 4-step update: {worst:.3g}")
    for (n, a), b in zip(ours.module.named_buffers(), ref.module.buffers()):
        assert torch.allclose(a.float(), b.float(), atol=1e-4, rtol=1e-3), f"buffer {n}"
    t = ours.sync_timings()
    assert len(t["buckets"]) == len(ours.plans) and all(b["end_ms"] >= b["start_ms"] for b in t["buckets"])
    assert ours.reducer.launches() == 4 * len(ours.plans)
    ours.shutdown()


@needs2
def test_native_ddp_matches_torch_ddp_nccl():
    _run(_ddp_worker)


def _trainstep_worker(rank, world):
    """TrainStep under the native DDP wrapper, every (graph, fused update, scaler, algorithm)
    combination against ONE oracle: the same native model stepped with gradients averaged by
    torch.distributed (NCCL) and the plain flat SGD kernel.  The parameter UPDATE (p_final -
    p_init) of every rank-slice of every bucket must match the oracle's — a reduction that is
    wrong but symmetric (a dropped slice, a wrong 1/world) fails here even though the replicas
    agree with each other."""
    import torch.distributed as dist
    from dpt_b200.amp import GradScaler
    from dpt_b200.graph_step import TrainStep
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
    dev = torch.device("cuda", rank)
    shape = (32, 3, 32, 32)
    steps = 6
    g = torch.Generator().manual_seed(3 + rank)
    x = torch.randn(shape, generator=g).pin_memory()
    t = torch.randint(0, 10, (32,), generator=g).pin_memory()

    def run(backend, graph, fused, fp16_scaler, algo):
        torch.manual_seed(11 + rank)
        model = OurDDP(build_model(dev, "resnet18", 10), device_ids=[rank], backend=backend,
                       allreduce_algo=algo)
        if algo == "twoshot_mc" and not getattr(model, "nvls", False):
            model.shutdown()           # NVSwitch multicast not exposed on this host
            return None
        init = model.arena.param.clone()
        opt = SGD(model.parameters(), lr=0.02, momentum=0.9, weight_decay=5e-4)
        if fused:
            assert model.enable_fused_sgd(opt)
        scaler = GradScaler(device=dev, init_scale=256.0, enabled=fp16_scaler)
        step = TrainStep(model, opt, scaler, batch_shape=shape, cuda_graph=graph)
        losses = []
        first = None
        for i in range(steps):
            step(x, t)
            losses.append(step.read_loss(lag=0))
            if i == 0:
                torch.cuda.synchronize()
                first = (model.arena.param - init).clone()   # the update of step 1
        torch.cuda.synchronize()
        # every rank trains on its own batch: judge the loss curve averaged over ranks, so all
        # ranks reach the same verdict
        return model, opt, scaler, init, _rank_mean(losses), first

    oracle, _, _, init0, losses0, first0 = run("nccl", False, False, False, "auto")
    delta0 = (oracle.arena.param - init0).clone()
    assert losses0[-1] < losses0[0], losses0
    slices = []
    for pl in oracle.plans:
        n4 = pl["numel"] // 4
        per = (n4 + world - 1) // world
        for r in range(world):
            lo, hi = min(n4, per * r) * 4, min(n4, per * (r + 1)) * 4
            if hi - lo >= 256:
                slices.append((pl["offset"] + lo, pl["offset"] + hi))

    def slice_err(a, b):
        worst = 0.0
        for lo, hi in slices:
            den = b[lo:hi].norm().item()
            if den > 1e-7:
                worst = max(worst, (a[lo:hi] - b[lo:hi]).norm().item() / den)
        return worst

    # noise floor: the oracle against itself (atomics order in the native model kernels)
    oracle2, _, _, _, _, first0b = run("nccl", False, False, False, "auto")
    floor1 = slice_err(first0b, first0)
    floor6 = slice_err(oracle2.arena.param - init0, delta0)
    oracle2.shutdown()
    # (random-init bf16 ResNet on noise: the model's own gradient varies by tens of percent from
    # run to run in the worst slice, so this check only catches O(1) errors such as a dropped
    # slice; test_reducer_paths_are_exact below pins the gradient path to 1e-6)
    tol1 = max(0.05, 4.0 * floor1)
    if rank == 0:
        print(f"[trainstep] run-to-run noise of the NCCL oracle: step-1 update {floor1:.4f}, "
              f"{steps}-step update {floor6:.4f} (worst slice, relative)")

    configs = [(False, False, False, "auto"), (True, False, False, "twoshot"),
               (True, True, False, "twoshot"), (True, True, False, "auto"),
               (True, True, False, "oneshot"), (True, False, True, "auto"),
               (True, False, False, "nvls"), (True, True, False, "nvls"),
               (True, True, False, "twoshot_mc")]
    for graph, fused, fp16_scaler, algo in configs:
        res = run("native", graph, fused, fp16_scaler, algo)
        if res is None:
            continue
        model, opt, scaler, init, losses, first = res
        tag = f"graph={graph} fused={fused} scaler={fp16_scaler} algo={algo} -> {model.bucket_algorithms()}"
        model.comm.check_health()
        assert torch.equal(init, init0), "construction broadcast differs between backends"
        assert losses[-1] < losses[0], (tag, losses)
        assert abs(losses[-1] - losses0[-1]) < 0.05 * abs(losses0[-1]) + 0.02, (tag, losses, losses0)
        assert model.verify_replicas()                 # bit-exact parameter digest, all ranks
        ep = model.comm.validate_epochs()              # same collectives issued everywhere
        assert max(ep) > 0
        # step 1 (same weights, same batch on both sides): every rank-slice of every bucket must
        # carry the oracle's update — a dropped slice or a wrong scale is an O(1) error here
        for lo, hi in slices:
            den = first0[lo:hi].norm().item()
            if den < 1e-7:
                continue
            rel = (first[lo:hi] - first0[lo:hi]).norm().item() / den
            assert rel < tol1, f"{tag}: step-1 update of arena[{lo}:{hi}] differs from the NCCL oracle by {rel:.4f}"
        worst = slice_err(first, first0)
        worst6 = slice_err(model.arena.param - init, delta0)
        cos = torch.nn.functional.cosine_similarity(model.arena.param - init, delta0, dim=0).item()
        shc = model.arena.shadow.double().sum().reshape(1)
        lo2, hi2 = shc.clone(), shc.clone()
        dist.all_reduce(lo2, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi2, op=dist.ReduceOp.MAX)
        assert torch.equal(lo2, hi2), f"bf16 shadows diverged ({tag})"
        assert torch.equal(model.arena.shadow, model.arena.param.bfloat16()), "shadow != bf16(param)"
        if fp16_scaler:
            assert scaler.get_scale() == 256.0
        if algo == "auto":
            assert model.bucket_algorithms()[-1] == "oneshot" and model.plans[-1]["bytes"] <= 65536
        if rank == 0:
            print(f"[trainstep] {tag}: final loss {losses[-1]:.4f} (oracle {losses0[-1]:.4f}), worst slice "
                  f"error step 1 {worst:.4f} / step {steps} {worst6:.4f}, cos {cos:.5f}")
        model.shutdown()
    oracle.shutdown()


@needs2
def test_trainstep_under_native_ddp_graph_and_fused_optimizer():
    _run(_trainstep_worker, world=_model_world())


# --------------------------------------------------------------------------------- exact reducer
def _reducer_exact_worker(rank, world):
    """The whole native gradient path with KNOWN gradients: the arena is filled with per-rank
    random values, every parameter is reported ready in arena order (what the engine does), and
    the result is compared element by element with NCCL all-reduce / world (+ torch.optim.SGD for
    the fused update).  Covers every bucket algorithm, the tail bucket, the relaxed closing
    barriers and the momentum shards; tolerance 1e-6 relative."""
    import torch.distributed as dist
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
    dev = torch.device("cuda", rank)
    configs = [("auto", False), ("auto", True), ("twoshot", False), ("twoshot", True),
               ("oneshot", True), ("nvls", False), ("nvls", True), ("twoshot_mc", True)]
    for algo, fused in configs:
        torch.manual_seed(50 + rank)
        model = OurDDP(build_model(dev, "resnet18", 10), device_ids=[rank], backend="native",
                       allreduce_algo=algo, bucket_cap_mb=2)
        if algo in ("nvls", "twoshot_mc") and not getattr(model, "nvls", False):
            model.shutdown()
            continue
        arena = model.arena
        opt = SGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-3)
        if fused:
            assert model.enable_fused_sgd(opt)
        ref_p = torch.nn.Parameter(arena.param.clone())
        ref_opt = torch.optim.SGD([ref_p], lr=0.05, momentum=0.9, weight_decay=1e-3)
        order = list(arena.order)
        tag = f"algo={algo} fused={fused} -> {model.bucket_algorithms()}"
        if algo == "auto":
            # the documented decision table (parallel/tuning.py) is what the C++ reducer does
            from dpt_b200.parallel import tuning
            mc = bool(getattr(model, "nvls", False))
            want = [tuning.choose_algorithm(pl["numel"] * 4, world, fused, mc) for pl in model.plans]
            assert want == model.bucket_algorithms(), (want, model.bucket_algorithms())
        for it in range(3):
            g = torch.randn(arena.numel, device=dev,
                            generator=torch.Generator(device=dev).manual_seed(1000 * it + rank))
            ref = g.clone()
            dist.all_reduce(ref)
            ref /= world
            arena.grad.copy_(g)
            torch.cuda.synchronize()
            dist.barrier()
            model.prepare_for_backward()
            for p in order:
                model._sink([p])
            assert model.finalize_backward() == len(model.plans)
            if not fused:
                torch.cuda.synchronize()
                # padding between parameters is not reduced: compare parameter slots only
                for pl in model.plans:
                    a, b = pl["offset"], pl["offset"] + pl["numel"]
                    err = (arena.grad[a:b] - ref[a:b]).abs().max().item()
                    assert err <= 1e-6 * max(1.0, ref[a:b].abs().max().item()), (tag, it, a, err)
                opt.step()
            ref_p.grad = ref
            ref_opt.step()
            torch.cuda.synchronize()
            for pl in model.plans:
                a, b = pl["offset"], pl["offset"] + pl["numel"]
                err = (arena.param[a:b] - ref_p.data[a:b]).abs().max().item()
                assert err <= 2e-6 * max(1.0, ref_p.data[a:b].abs().max().item()), \
                    f"{tag}: step {it}: parameters of bucket at {a} differ from all-reduce + SGD by {err:.3g}"
            assert torch.equal(arena.shadow, arena.param.bfloat16()), (tag, "shadow != bf16(param)")
            assert model.verify_replicas()
        # momentum: complete after the shard gather, equal to torch's buffer
        model.gather_optimizer_state(opt)
        m = opt._flat_momentum()
        mref = ref_opt.state[ref_p]["momentum_buffer"]
        for pl in model.plans:
            a, b = pl["offset"], pl["offset"] + pl["numel"]
            err = (m[a:b] - mref[a:b]).abs().max().item()
            assert err <= 2e-6 * max(1.0, mref[a:b].abs().max().item()), (tag, "momentum", a, err)
        model.comm.check_health()
        model.comm.validate_epochs()
        if rank == 0:
            print(f"[reducer exact] {tag}: ok")
        model.shutdown()


@needs2
def test_reducer_paths_are_exact():
    _run(_reducer_exact_worker, world=_model_world())


# --------------------------------------------------------------------------------- fault path
def _stalled_rank_worker(rank, world):
    """A rank that never enters the collective must produce a diagnostic on the others within
    the timeout instead of a hang; DPT_FAULT_STALL_RANK injects the fault."""
    import time
    import torch.distributed as dist
    from dpt_b200.parallel.comm import Communicator
    comm = Communicator(timeout_s=1.5)
    comm.start_watchdog(interval_s=0.2)
    dev = torch.device("cuda", rank)
    buf = comm.alloc(1 << 20)
    stalled = int(os.environ.get("DPT_FAULT_STALL_RANK", world - 1))
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.time()
    if rank != stalled:
        comm.allreduce_(buf, 0, 1 << 18)
        torch.cuda.synchronize()               # returns once the device-side wait timed out
        elapsed = time.time() - t0
        assert elapsed < 20.0, f"collective hung for {elapsed:.1f}s"
        raised = False
        try:
            comm.check_health()
        except RuntimeError as e:
            raised = True
            assert f"rank {stalled}" in str(e) or "timed out" in str(e), str(e)
        assert raised, "a missing peer must surface as an error"
        time.sleep(0.5)
        assert comm.fault is not None and comm.fault[0] == rank, "watchdog thread did not report"
    else:
        time.sleep(4.0)                         # the "dead" rank
    hb = comm.peer_heartbeats()
    assert all(v < 30 for v in hb.values()), hb
    dist.barrier()
    comm.stop_watchdog()


@needs2
def test_stalled_rank_times_out_with_diagnostic():
    _run(_stalled_rank_worker, world=2)


# --------------------------------------------------------------------------------- NVLS
def _nvls_worker(rank, world):
    import torch.distributed as dist
    from dpt_b200.parallel.comm import Communicator
    comm = Communicator(timeout_s=20.0)
    dev = torch.device("cuda", rank)
    buf = comm.alloc_multicast(64 << 20)
    if buf is None:
        if rank == 0:
            print("NVLS multicast not exposed on this host: skipped")
        comm.close()
        return
    for n in (4, 1024, 4100, 1 << 20, 11181648):
        for rep in range(2):
            x = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(31 * rank + rep + n % 97))
            view = buf.view(torch.float32, 0, n)
            view.copy_(x)
            ref = x.clone()
            dist.all_reduce(ref)
            torch.cuda.synchronize()
            dist.barrier()
            comm.allreduce_nvls_(buf, 0, n)
            torch.cuda.synchronize()
            err = (view - ref / world).abs().max().item()
            assert err <= 2e-5 * max(1.0, ref.abs().max().item()), (n, rep, err)
            lo, hi = view.clone(), view.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            assert torch.equal(lo, hi), n
    # init-time autotune: measured table + crossover, identical on every rank
    res = comm.autotune(buf, sizes=(1 << 16, 1 << 20, 1 << 22), iters=5)
    assert set(res["table"]) == {1 << 16, 1 << 20, 1 << 22}
    assert all(r["twoshot_us"] > 0 and r["nvls_us"] > 0 for r in res["table"].values())
    gathered = [None] * world
    dist.all_gather_object(gathered, res["nvls_min_bytes"])
    assert all(g == gathered[0] for g in gathered)
    # the same buffer also works with the P2P kernels (unicast peer mappings)
    n = 65536
    view = buf.view(torch.float32, 0, n)
    view.fill_(float(rank + 1))
    torch.cuda.synchronize()
    dist.barrier()
    comm.allreduce_(buf, 0, n, scale=1.0)
    torch.cuda.synchronize()
    assert torch.allclose(view, torch.full_like(view, world * (world + 1) / 2))
    comm.check_health()
    comm.close()


@needs2
def test_nvls_multicast_allreduce():
    _run(_nvls_worker)


# --------------------------------------------------------------------------------- compat loop
def _compat_loop_worker(rank, world):
    """Reference-style loop (autocast -> model(x) -> criterion -> backward -> step) through the
    native engine under the native DDP wrapper: replicas stay identical, loss goes down."""
    import torch.distributed as dist
    from dpt_b200.models.resnet import build_model
    from dpt_b200.ops.loss import CrossEntropyLoss
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
    dev = torch.device("cuda", rank)
    torch.manual_seed(5 + rank)
    model = OurDDP(build_model(dev, "resnet18", 10), device_ids=[rank], backend="native", profile=True)
    opt = SGD(model.parameters(), lr=0.02, momentum=0.9, weight_decay=5e-4)
    crit = CrossEntropyLoss()
    g = torch.Generator(device=dev).manual_seed(rank)
    x = torch.randn(32, 3, 32, 32, device=dev, generator=g)
    t = torch.randint(0, 10, (32,), device=dev, generator=g)
    losses = []
    for _ in range(6):
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = crit(model(x), t)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    torch.cuda.synchronize()
    model.comm.check_health()
    losses = _rank_mean(losses)           # same verdict on every rank
    assert losses[-1] < losses[0], losses
    cs = model.arena.param.double().sum().reshape(1)
    lo, hi = cs.clone(), cs.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert torch.equal(lo, hi), "replicas diverged in the autograd-bridge path"
    tm = model.sync_timings()
    assert len(tm["buckets"]) == len(model.plans)
    model.shutdown()


@needs2
def test_reference_style_loop_under_native_ddp():
    _run(_compat_loop_worker, world=2)


# --------------------------------------------------------------------------------- checkpoint
def _fused_checkpoint_worker(rank, world):
    """With the SGD update fused into the two-shot all-reduce each rank holds momentum only for
    the slices it owns (one-shot buckets are replicated).  save_checkpoint must gather the shards:
    driven with KNOWN gradients, the file rank 0 writes must hold exactly the momentum buffers and
    weights of torch.optim.SGD fed with the NCCL-averaged gradients, and a run resumed from it
    (unfused this time) must take exactly the reference's next step."""
    import tempfile
    import torch.distributed as dist
    from dpt_b200.checkpoint import load_checkpoint, save_checkpoint
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
    dev = torch.device("cuda", rank)

    def build(fused):
        torch.manual_seed(31 + rank)
        model = OurDDP(build_model(dev, "resnet18", 10), device_ids=[rank], backend="native",
                       allreduce_algo="auto")
        opt = SGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-3)
        if fused:
            assert model.enable_fused_sgd(opt)
        return model, opt

    def known_step(model, opt, it, fused):
        arena = model.arena
        g = torch.randn(arena.numel, device=dev,
                        generator=torch.Generator(device=dev).manual_seed(1000 * it + rank))
        ref = g.clone()
        dist.all_reduce(ref)
        ref /= world
        arena.grad.copy_(g)
        torch.cuda.synchronize()
        dist.barrier()
        model.prepare_for_backward()
        for p in list(arena.order):
            model._sink([p])
        assert model.finalize_backward() == len(model.plans)
        if not fused:
            opt.step()
        torch.cuda.synchronize()
        return ref

    def close(got, want, what):
        err = (got.float() - want.float()).abs().max().item()
        assert err <= 2e-6 * max(1.0, want.float().abs().max().item()), f"{what}: max abs err {err:.3g}"

    path = [None]
    if rank == 0:
        path[0] = os.path.join(tempfile.mkdtemp(prefix="dpt_ckpt_"), "checkpoint.pt")
    dist.broadcast_object_list(path, src=0)
    model, opt = build(True)
    assert len(model.momentum_shards()) >= world and "oneshot" in model.bucket_algorithms()
    arena = model.arena
    ref_p = torch.nn.Parameter(arena.param.clone())
    ref_opt = torch.optim.SGD([ref_p], lr=0.05, momentum=0.9, weight_decay=1e-3)
    for it in range(3):
        ref_p.grad = known_step(model, opt, it, True)
        ref_opt.step()
    own = opt._flat_momentum().clone()
    save_checkpoint(path[0], model, opt, None, 1, None, rank)     # collective: gathers the shards
    dist.barrier()
    mref = ref_opt.state[ref_p]["momentum_buffer"]
    for a, b, owner in model.momentum_shards():
        if owner == rank:
            assert torch.equal(opt._flat_momentum()[a:b], own[a:b]), "the owner's shard must be untouched"
    # the FILE (written by rank 0, read here by every rank) against the torch reference
    payload = torch.load(path[0], map_location="cpu", weights_only=False)
    assert payload["epoch"] == 1 and len(payload["optimizer"]["state"]) == len(opt._params)
    for i, st in payload["optimizer"]["state"].items():
        sl = arena.slot(opt._params[int(i)])
        close(st["momentum_buffer"], mref.as_strided(sl.shape, sl.stride, sl.offset).cpu(),
              f"saved momentum_buffer of {sl.name}")
    for name, p in model.module.named_parameters():
        sl = arena.slot(p)
        close(payload["model"]["module." + name],
              ref_p.data.as_strided(sl.shape, sl.stride, sl.offset).cpu(), f"saved weight {name}")
    model.shutdown()
    dist.barrier()
    # resume into a fresh wrapper WITHOUT the fused update and take the reference's next step
    model2, opt2 = build(False)
    load_checkpoint(path[0], model2, opt2, None)
    # (per parameter: the padding between arena slots is not part of a checkpoint)
    slots = [model2.arena.slot(p) for p in model2.arena.order]
    for sl in slots:
        a, b = sl.offset, sl.offset + sl.numel
        close(opt2._flat_momentum()[a:b], mref[a:b], f"restored momentum of {sl.name}")
        close(model2.arena.param[a:b], ref_p.data[a:b], f"restored weights of {sl.name}")
    ref_p.grad = known_step(model2, opt2, 3, False)
    ref_opt.step()
    for sl in slots:
        a, b = sl.offset, sl.offset + sl.numel
        close(model2.arena.param[a:b], ref_p.data[a:b], f"weights of {sl.name} one step after resume")
        close(opt2._flat_momentum()[a:b], mref[a:b], f"momentum of {sl.name} one step after resume")
    assert model2.verify_replicas()
    model2.comm.check_health()
    model2.shutdown()


@needs2
def test_checkpoint_after_fused_sgd_gathers_momentum_shards():
    _run(_fused_checkpoint_worker, world=2)


# --------------------------------------------------------------------------------- killed rank
def _killed_rank_worker(rank, world, port, q):
    """Rank 1's process disappears in the middle of training (os._exit inside the step loop).
    Rank 0 must END with a diagnostic naming the missing rank — through the watchdog DDP starts
    (exit code 70) or the health check — instead of spinning in a collective forever."""
    import time
    _init(rank, world, port)
    from dpt_b200.graph_step import TrainStep
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
    dev = torch.device("cuda", rank)
    shape = (16, 3, 32, 32)
    model = OurDDP(build_model(dev, "resnet18", 10), device_ids=[rank], backend="native",
                   comm_timeout=2.0)
    opt = SGD(model.parameters(), lr=0.01, momentum=0.9)
    model.enable_fused_sgd(opt)
    step = TrainStep(model, opt, None, batch_shape=shape, cuda_graph=True)
    x = torch.randn(shape).pin_memory()
    t = torch.randint(0, 10, (16,)).pin_memory()
    q.put((rank, "started"))
    t0 = time.time()
    for i in range(100000):
        if rank == 1 and i == 5:
            torch.cuda.synchronize()
            os._exit(0)                       # the rank is gone: no teardown, no goodbye
        step(x, t)
        if i % 4 == 3:
            step.read_loss(lag=0)
            try:
                model.comm.check_health()
            except RuntimeError as e:
                q.put((rank, "detected: " + str(e)))
                os._exit(3)
        if time.time() - t0 > 60:
            q.put((rank, "no error after 60 s"))
            os._exit(4)


@needs2
def test_rank_killed_mid_step_ends_the_job_with_a_diagnostic():
    import time
    import torch.multiprocessing as mp
    from dpt_b200.launcher import free_port
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_killed_rank_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    t0 = time.time()
    procs[0].join(150)
    took = time.time() - t0
    alive = procs[0].is_alive()
    msgs = []
    while not q.empty():
        msgs.append(q.get())
    for p in procs:
        if p.is_alive():
            p.kill()
    assert not alive, f"rank 0 still running {took:.0f} s after its peer vanished: {msgs}"
    assert procs[1].exitcode == 0
    detected = [m for r, m in msgs if r == 0 and m.startswith("detected")]
    assert procs[0].exitcode == 70 or detected, (procs[0].exitcode, msgs)
    if detected:
        assert "rank 1" in detected[0], detected
