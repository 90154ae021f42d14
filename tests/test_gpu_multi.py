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
    return min(torch.cuda.device_count(), int(os.environ.get("DPT_TEST_WORLD", "8")))


def _init(rank, world, port):
    import torch.distributed as dist
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
                       "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank)})
    sys.path.insert(0, ROOT)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))


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
    results = []
    for _ in procs:
        results.append(q.get(timeout=900))
    for p in procs:
        p.join(120)
    for rank, msg in results:
        assert msg == "ok", f"rank {rank}:\n{msg}"


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
    # IEEE fp32 convolutions: with TF32 a 1e-7 weight difference (cuDNN picks differently
    # rounded kernels for differently aligned weights) can flip a TF32 rounding and the
    # random-init net amplifies that to O(0.1) gradient differences within two steps
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
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
    for it in range(4):
        x = torch.randn(16, 3, 32, 32, device=dev, generator=g)
        t = torch.randint(0, 10, (16,), device=dev, generator=g)
        for model, opt in ((ours, opt_o), (ref, opt_r)):
            opt.zero_grad(set_to_none=True)
            F.cross_entropy(model(x), t).backward()
            if it == 0 and model is ref:
                # averaged gradients of the first step: only the reduction order differs
                torch.cuda.synchronize()
                for (n, a), b in zip(ours.module.named_parameters(), ref.module.parameters()):
                    err = (a.grad - b.grad).abs().max().item()
                    den = b.grad.abs().max().item() + 1e-12
                    assert err / den < 2e-3, f"{n}: averaged grad differs from torch DDP + NCCL ({err/den:.3g})"
            opt.step()
    torch.cuda.synchronize()
    ours.comm.check_health()
    for (n, a), b in zip(ours.module.named_parameters(), ref.module.parameters()):
        # TF32 cuDNN kernels are not run-to-run deterministic; trajectories stay parallel
        c = torch.nn.functional.cosine_similarity(a.flatten(), b.flatten(), dim=0).item()
        assert c > 0.999, f"{n} diverged from torch DDP + NCCL (cos {c:.6f})"
    for (n, a), b in zip(ours.module.named_buffers(), ref.module.buffers()):
        assert torch.allclose(a.float(), b.float(), atol=5e-3, rtol=5e-2), f"buffer {n}"
    t = ours.sync_timings()
    assert len(t["buckets"]) == len(ours.plans) and all(b["end_ms"] >= b["start_ms"] for b in t["buckets"])
    assert ours.reducer.launches() == 4 * len(ours.plans)
    ours.shutdown()


@needs2
def test_native_ddp_matches_torch_ddp_nccl():
    _run(_ddp_worker)


def _trainstep_worker(rank, world):
    import torch.distributed as dist
    from dpt_b200.amp import GradScaler
    from dpt_b200.graph_step import TrainStep
    from dpt_b200.models.resnet import build_model
    from dpt_b200.optim import SGD
    from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
    dev = torch.device("cuda", rank)
    shape = (32, 3, 32, 32)
    configs = [(False, False, False, "twoshot"), (True, False, False, "twoshot"),
               (True, True, False, "twoshot"), (True, False, True, "twoshot"),
               (True, False, False, "nvls"), (True, True, False, "nvls")]
    if os.environ.get("DPT_TEST_MCPUSH", "0") == "1":
        # experimental: fused-SGD two-shot whose publish phase is one multicast store
        configs.append((True, True, False, "twoshot_mc"))
    for graph, fused, fp16_scaler, algo in configs:
        torch.manual_seed(11 + rank)
        model = OurDDP(build_model(dev, "resnet18", 10), device_ids=[rank], backend="native",
                       allreduce_algo=algo)
        opt = SGD(model.parameters(), lr=0.02, momentum=0.9, weight_decay=5e-4)
        if fused:
            assert model.enable_fused_sgd(opt)
        scaler = GradScaler(device=dev, init_scale=256.0, enabled=fp16_scaler)
        step = TrainStep(model, opt, scaler, batch_shape=shape, cuda_graph=graph)
        g = torch.Generator().manual_seed(3 + rank)
        x = torch.randn(shape, generator=g).pin_memory()
        t = torch.randint(0, 10, (32,), generator=g).pin_memory()
        losses = []
        for _ in range(8):
            step(x, t)
            losses.append(step.read_loss(lag=0))
        torch.cuda.synchronize()
        model.comm.check_health()
        assert losses[-1] < losses[0], (graph, fused, algo, losses)
        assert model.verify_replicas()                 # bit-exact parameter digest, all ranks
        ep = model.comm.validate_epochs()              # same collectives issued everywhere
        assert max(ep) > 0
        cs = model.arena.param.double().sum().reshape(1)
        lo, hi = cs.clone(), cs.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert torch.equal(lo, hi), f"replicas diverged (graph={graph}, fused={fused}, algo={algo})"
        shc = model.arena.shadow.double().sum().reshape(1)
        lo2, hi2 = shc.clone(), shc.clone()
        dist.all_reduce(lo2, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi2, op=dist.ReduceOp.MAX)
        assert torch.equal(lo2, hi2), f"bf16 shadows diverged (fused={fused}, algo={algo})"
        assert torch.equal(model.arena.shadow, model.arena.param.bfloat16()), "shadow != bf16(param)"
        sh = model.arena.shadow.float().sum().reshape(1)
        assert torch.isfinite(sh).all()
        if fp16_scaler:
            assert scaler.get_scale() == 256.0
        model.shutdown()


@needs2
def test_trainstep_under_native_ddp_graph_and_fused_optimizer():
    _run(_trainstep_worker)


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
