"""Multi-process CPU tests (gloo, world_size 2): the wrapper's bucketing / hooks / broadcast
logic against torch DDP, and the entry point end to end (BASELINE.json config 1)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _init(rank, world, port):
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
                       "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank)})
    sys.path.insert(0, ROOT)
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)


def _ddp_parity_worker(rank, world, port, q):
    try:
        _init(rank, world, port)
        import copy
        from dpt_b200.models.resnet import resnet18
        from dpt_b200.optim import SGD
        from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
        from torch.nn.parallel import DistributedDataParallel as TorchDDP
        torch.manual_seed(100 + rank)          # replicas start different (seed + rank semantics)
        base = resnet18(10)
        ours = OurDDP(copy.deepcopy(base), backend="gloo", bucket_cap_mb=1)
        ref = TorchDDP(copy.deepcopy(base), bucket_cap_mb=1)
        # both broadcast rank 0's weights at construction
        for a, b in zip(ours.module.state_dict().values(), ref.module.state_dict().values()):
            assert torch.allclose(a.float(), b.float()), "init broadcast mismatch"
        assert len(ours.plans) > 3
        opt_o = SGD(ours.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4)
        opt_r = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4)
        g = torch.Generator().manual_seed(7 + rank)  # each rank its own shard
        losses = []
        for it in range(3):
            x = torch.randn(4, 3, 32, 32, generator=g)
            t = torch.randint(0, 10, (4,), generator=g)
            for model, opt in ((ours, opt_o), (ref, opt_r)):
                opt.zero_grad(set_to_none=True)
                loss = torch.nn.functional.cross_entropy(model(x), t)
                loss.backward()
                opt.step()
                losses.append(loss.item())
        for (n, a), b in zip(ours.module.named_parameters(), ref.module.parameters()):
            assert torch.allclose(a, b, atol=2e-4, rtol=1e-3), f"param {n} diverged from torch DDP"
        # replicas identical across ranks (checksum all-reduce)
        cs = torch.stack([p.detach().double().sum() for p in ours.parameters()])
        lo, hi = cs.clone(), cs.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert torch.allclose(lo, hi, atol=1e-9), "replicas diverged across ranks"
        # running stats follow rank 0 (broadcast_buffers)
        rm = ours.module.bn1.running_mean.clone()
        rm0 = rm.clone()
        dist.broadcast(rm0, 0)
        # no_sync skips the reduction
        before = ours.reducer.launch_count
        with ours.no_sync():
            torch.nn.functional.cross_entropy(ours(x), t).backward()
        assert ours.reducer.launch_count == before
        # state_dict carries the DDP prefix
        assert all(k.startswith("module.") for k in ours.state_dict())
        log = ours._get_ddp_logging_data()
        assert log["world_size"] == world and log["num_parameter_tensors"] == 62
        assert log["total_parameter_size_bytes"] == 44726568
        for k in ("avg_forward_compute_time", "avg_backward_compute_time", "avg_backward_comm_time",
                  "avg_backward_compute_comm_overlap_time"):
            assert k in log                               # torch Logger's runtime keys (ns)
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def _comm_hook_worker(rank, world, port, q):
    try:
        _init(rank, world, port)
        from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
        torch.manual_seed(0)
        net = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 4))
        ddp = OurDDP(net, backend="gloo")
        seen = []

        def hook(state, bucket):
            seen.append((bucket.index(), bucket.is_last(), len(bucket.parameters())))
            buf = bucket.buffer()
            buf.div_(dist.get_world_size())
            fut = dist.all_reduce(buf, async_op=True).get_future()
            return fut.then(lambda f: f.value()[0])

        ddp.register_comm_hook(None, hook)
        x = torch.full((2, 8), float(rank + 1))
        ddp(x).sum().backward()
        assert seen and seen[-1][1] is True
        g = [p.grad.clone() for p in ddp.parameters()]
        for t in g:
            lo, hi = t.clone(), t.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            assert torch.allclose(lo, hi)
        q.put((rank, "ok"))
    except Exception:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def _join_worker(rank, world, port, q):
    """Uneven inputs: rank 0 has 3 batches, rank 1 has 5.  Under ``with model.join():`` nobody
    hangs and the parameters follow torch DDP's Join algorithm step by step."""
    try:
        _init(rank, world, port)
        import copy
        from dpt_b200.optim import SGD
        from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
        from torch.nn.parallel import DistributedDataParallel as TorchDDP
        torch.manual_seed(3)
        base = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3, padding=1), torch.nn.BatchNorm2d(8),
                                   torch.nn.ReLU(), torch.nn.Flatten(), torch.nn.Linear(8 * 8 * 8, 10))
        ours = OurDDP(copy.deepcopy(base), backend="gloo", bucket_cap_mb=0.01, first_bucket_mb=0.001)
        ref = TorchDDP(copy.deepcopy(base))
        assert len(ours.plans) >= 2
        opt_o = SGD(ours.parameters(), lr=0.05, momentum=0.9)
        opt_r = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
        n_batches = 3 if rank == 0 else 5
        g = torch.Generator().manual_seed(20 + rank)
        data = [(torch.randn(4, 3, 8, 8, generator=g), torch.randint(0, 10, (4,), generator=g))
                for _ in range(n_batches)]
        for model, opt in ((ours, opt_o), (ref, opt_r)):
            with model.join():
                for x, t in data:
                    opt.zero_grad(set_to_none=True)
                    torch.nn.functional.cross_entropy(model(x), t).backward()
                    opt.step()
        for (n, a), b in zip(ours.module.named_parameters(), ref.module.parameters()):
            assert torch.allclose(a, b, atol=1e-5, rtol=1e-4), f"{n} differs from torch DDP under join()"
        # both implementations finish by handing every rank the longest-trained rank's model
        cs = torch.stack([p.detach().double().sum() for p in ours.parameters()])
        lo, hi = cs.clone(), cs.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert torch.equal(lo, hi), "replicas differ after join()"
        q.put((rank, "ok"))
    except Exception:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def _run(worker, world=2):
    from dpt_b200.launcher import free_port
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(60)
    for rank, msg in results:
        assert msg == "ok", f"rank {rank}:\n{msg}"


def _mismatch_worker(rank, world, port, q):
    try:
        _init(rank, world, port)
        from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
        net = torch.nn.Linear(8, 16 + rank)          # different shapes on different ranks
        try:
            OurDDP(net, backend="gloo")
            q.put((rank, "constructor accepted mismatching replicas"))
        except RuntimeError as e:
            q.put((rank, "ok" if "same model across all ranks" in str(e) else str(e)))
    except Exception:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def _builtin_hooks_worker(rank, world, port, q):
    try:
        _init(rank, world, port)
        from dpt_b200.parallel import comm_hooks
        from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
        x = torch.full((2, 8), float(rank + 1))
        grads = {}
        for name in ("allreduce_hook", "fp16_compress_hook", "bf16_compress_hook", "noop_hook"):
            torch.manual_seed(0)
            net = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 4))
            ddp = OurDDP(net, backend="gloo")
            ddp.register_comm_hook(None, getattr(comm_hooks, name))
            ddp(x).sum().backward()
            grads[name] = torch.cat([p.grad.flatten() for p in ddp.parameters()]).clone()
        ref = grads["allreduce_hook"]
        both = [torch.empty_like(ref) for _ in range(world)]
        dist.all_gather(both, ref)
        assert torch.equal(both[0], both[1]), "all-reduce hook left replicas different"
        assert torch.allclose(grads["fp16_compress_hook"], ref, rtol=2e-3, atol=1e-3)
        assert torch.allclose(grads["bf16_compress_hook"], ref, rtol=2e-2, atol=1e-2)
        local = [torch.empty_like(ref) for _ in range(world)]
        dist.all_gather(local, grads["noop_hook"])
        assert not torch.equal(local[0], local[1]), "noop hook must not communicate"
        assert torch.allclose((local[0] + local[1]) / world, ref, rtol=1e-5, atol=1e-6)
        q.put((rank, "ok"))
    except Exception:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def _verify_replicas_worker(rank, world, port, q):
    try:
        _init(rank, world, port)
        from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
        torch.manual_seed(5 + rank)
        net = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 4))
        ddp = OurDDP(net, backend="gloo")
        assert ddp.verify_replicas()                     # rank 0's weights were broadcast
        ddp(torch.randn(2, 8)).sum().backward()
        with torch.no_grad():
            for p in ddp.parameters():
                p -= 0.1 * p.grad
        assert ddp.verify_replicas()                     # same reduced gradient everywhere
        if rank == 1:
            with torch.no_grad():
                next(ddp.parameters()).view(-1)[3] += 1e-7   # one-ulp-scale silent desync
        assert ddp.verify_replicas(raise_on_mismatch=False) is False
        try:
            ddp.verify_replicas()
            q.put((rank, "desync not reported"))
        except RuntimeError as e:
            q.put((rank, "ok" if "diverged" in str(e) else str(e)))
    except Exception:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def test_builtin_comm_hooks_gloo():
    _run(_builtin_hooks_worker)


def test_verify_replicas_detects_desync():
    _run(_verify_replicas_worker)


def test_parameter_shape_verification_across_ranks():
    _run(_mismatch_worker)


def test_ddp_wrapper_matches_torch_ddp_gloo():
    _run(_ddp_parity_worker)


def test_comm_hook_escape_hatch_gloo():
    _run(_comm_hook_worker)


def _entry(outdir):
    sys.path.insert(0, ROOT)
    torch.set_num_threads(2)
    from dpt_b200.train_ddp import main
    main(["--synthetic", "--synthetic-size", "32", "--epochs", "1", "--batch-size", "4",
          "--workers", "0", "--print-freq", "2", "--output-dir", outdir, "--backend", "gloo",
          "--steps", "3", "--save-every", "1", "--check-sync-every", "1", "--nvtx"])


def test_entry_point_two_process_gloo(tmp_path):
    """Config 1: default model FP32 on CPU/gloo world_size=2 with synthetic data."""
    from dpt_b200.launcher import spawn
    out = str(tmp_path / "exp")
    spawn(_entry, 2, args=(out,))
    lines = open(os.path.join(out, "metrics_rank0.csv")).read().splitlines()
    assert lines[0].startswith("epoch,train_loss") and len(lines) == 2
    ck = torch.load(os.path.join(out, "checkpoint.pt"), weights_only=False)
    assert all(k.startswith("module.") for k in ck["model"]) and ck["epoch"] == 1


def test_join_with_uneven_inputs_matches_torch_ddp_gloo():
    _run(_join_worker)


def test_torchrun_restart_resumes_from_the_checkpoint(tmp_path):
    """Elastic recovery as far as the reference's stack offers it: a rank dies at
    the start of the second epoch, torchrun (--max-restarts) tears the group down and starts it
    again with the same command line, ``--resume auto`` picks up the checkpoint written after the
    first epoch, and the job finishes all epochs — the first epoch is not trained twice.

    (With plain ``env://`` the restarted ranks read the dead incarnation's rendezvous keys from
    the agent's store about every other run — "connection refused" or a stalled connect;
    ``runtime._restart_safe_store`` gives every attempt its own key prefix.  One retry is kept
    for a loaded machine; a second rendezvous failure fails the test.)"""
    import subprocess
    from dpt_b200.launcher import free_port
    last = ""
    for attempt in range(2):
        out = tmp_path / f"exp{attempt}"
        env = dict(os.environ, DPT_FAULT_EXIT="1:1", OMP_NUM_THREADS="2")
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
            env.pop(k, None)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
               "--max-restarts=1", "--monitor-interval=1", "--master-addr", "127.0.0.1",
               "--master-port", str(free_port()), os.path.join(ROOT, "train_ddp.py"),
               "--synthetic", "--synthetic-size", "32", "--epochs", "3", "--batch-size", "4", "--workers", "0",
               "--backend", "gloo", "--print-freq", "100", "--output-dir", str(out), "--save-every", "1",
               "--resume", "auto", "--steps", "2"]
        try:
            r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=150, cwd=ROOT)
        except subprocess.TimeoutExpired as e:
            last = "timed out: " + str(e.stdout)[-500:] + str(e.stderr)[-1500:]
            continue
        log = r.stdout + r.stderr
        if r.returncode != 0:
            last = log[-3000:]
            if "connectFullMesh" in log or "Connection refused" in log or "timed out" in log:
                continue                                  # rendezvous trouble, not a resume bug
            raise AssertionError(last)
        assert "[fault injection] rank 1 exits at the start of epoch 1" in log
        assert "Resumed from" in log and "at epoch 1" in log, log[-3000:]
        lines = open(out / "metrics_rank0.csv").read().splitlines()
        assert [ln.split(",")[0] for ln in lines] == ["epoch", "1", "2", "3"], lines
        ck = torch.load(out / "checkpoint.pt", weights_only=False)
        assert ck["epoch"] == 3
        return
    raise AssertionError("torchrun restart did not complete (twice): " + last[-2000:])
