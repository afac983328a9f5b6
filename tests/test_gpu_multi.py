"""Multi-GPU tests (need >= 2 GPUs; skipped otherwise): one process per GPU, NCCL for the [P+1] bucket.
The W-GPU run is compared with the W-process gloo DDP run of the oracle (the reference's topology)."""
import os

import numpy as np
import pytest
import torch

from oracle import preprocess_ref, synth, train_ref

pytestmark = pytest.mark.gpu


def _need(n):
    if not torch.cuda.is_available() or torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


def _nccl_worker(rank, world, port, feats, labels, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from wxddp import trainer

    h = trainer.fit(feats, labels, device=torch.device("cuda", rank), rank=rank, world=world, dist=dist, batch_size=4,
                    max_epochs=2, parity=True, verbose=False)
    q.put((rank, h["train_loss"], h["val_loss"], h["val_acc"], h["params"].numpy()))
    dist.destroy_process_group()


def _gloo_worker(rank, world, port, feats, labels, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ref = train_ref.fit(feats, labels, rank=rank, world=world, use_ddp=True, max_epochs=2)
    q.put((rank, ref["train_loss"], ref["val_loss"], ref["val_acc"], ref["params"].numpy()))
    dist.destroy_process_group()


def _run(worker, world, port, feats, labels):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=worker, args=(r, world, port, feats, labels, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=600) for _ in range(world)], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    return res


@pytest.mark.parametrize("world", [2, 4])
def test_ddp_on_gpus_matches_gloo_ddp_oracle(world):
    _need(world)
    t = synth.synth_table(31, 203, 5)
    pp = preprocess_ref.preprocess_table(t["raw"], t["labels"])
    feats, labels = torch.from_numpy(pp["features"]), torch.from_numpy(pp["labels"])
    got = _run(_nccl_worker, world, _free_port(), feats, labels)
    ref = _run(_gloo_worker, world, _free_port(), feats, labels)
    for g, r in zip(got, ref):
        np.testing.assert_allclose(g[1], r[1], rtol=2e-4, atol=1e-6)      # per-step rank-mean train_loss
        np.testing.assert_allclose(g[2], r[2], rtol=2e-4)
        np.testing.assert_allclose(g[3], r[3], rtol=1e-6)
        np.testing.assert_allclose(g[4], r[4], rtol=2e-3, atol=2e-5)
    for r in range(1, world):
        np.testing.assert_array_equal(got[0][4], got[r][4])               # replicas stay bit-identical


def _free_port() -> int:
    import socket

    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as sk:
        sk.bind(("127.0.0.1", 0))
        return int(sk.getsockname()[1])


def _peer_worker(rank, world, port, q):
    """Fused in-kernel exchange (NVLink peer memory) vs the NCCL bucket path (same data, same dropout bits), and vs a
    one-GPU run on the concatenated batch (dropout off: SURVEY 8(e) "equal up to summation order")."""
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    import wxddp
    from wxddp.engine import StepEngine

    f, b, steps = 20, 8192, 6
    g = torch.Generator().manual_seed(100)
    x_all = torch.randn(steps, world, b, f, generator=g)                    # [step][rank][row][feature]
    y_all = (torch.rand(steps, world, b, generator=g) < 0.3).long()
    x = x_all[:, rank].reshape(steps * b, f).contiguous().to(dev)
    y = y_all[:, rank].reshape(steps * b).contiguous().to(dev)
    res = {}
    for name, use_peers, mode in (("nccl", False, wxddp.DROPOUT_PHILOX), ("peer", True, wxddp.DROPOUT_PHILOX),
                                  ("peer_nodrop", True, wxddp.DROPOUT_OFF)):
        torch.manual_seed(5)
        eng = StepEngine(f, device=dev, world=world, rank=rank, process_group=dist)
        if use_peers:
            eng.connect_peers()
        losses = []
        for s in range(steps):
            l = eng.step(x, y, row0=s * b, batch=b, dropout=mode)
            losses.append(float(l.item()))
        torch.cuda.synchronize()
        res[name] = (losses, eng.params.cpu().numpy(), eng.ctx.tc_status(), int(eng.state[0].item()) if use_peers else steps)
        dist.barrier()
    # raw-column form (what bench.py launches): standardisation fused, statistics merged across ranks
    raw = (x.t().contiguous() * 7.0 + 50.0)
    cols = [raw[j] for j in range(f)]
    torch.manual_seed(5)
    eng = StepEngine(f, device=dev, world=world, rank=rank, process_group=dist)
    eng.fit_stats(cols)
    eng.connect_peers()
    raw_losses = [float(eng.step_raw(cols, y, row0=s * b, batch=b, dropout=wxddp.DROPOUT_OFF).item()) for s in range(steps)]
    res["peer_raw"] = (raw_losses, eng.params.cpu().numpy(), eng.ctx.tc_status(), int(eng.state[0].item()))
    dist.barrier()
    if rank == 0:
        # the same optimisation on ONE GPU over the concatenated batches
        torch.manual_seed(5)
        one = StepEngine(f, device=dev, world=1, rank=0)
        xc = x_all.reshape(steps * world * b, f).contiguous().to(dev)
        yc = y_all.reshape(steps * world * b).contiguous().to(dev)
        l1 = [float(one.step(xc, yc, row0=s * world * b, batch=world * b, dropout=wxddp.DROPOUT_OFF).item())
              for s in range(steps)]
        res["single"] = (l1, one.params.cpu().numpy(), one.ctx.tc_status(), steps)
    q.put((rank, res))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_in_kernel_peer_allreduce_matches_nccl_bucket_path(world):
    _need(world)
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_peer_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=900) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, res in out:
        for name in ("peer", "peer_nodrop", "peer_raw"):
            assert res[name][2] == 0 and res[name][3] == 6, (rank, name, res[name][2:])
        np.testing.assert_allclose(res["peer"][0], res["nccl"][0], rtol=1e-5)              # rank-mean train_loss per step
        np.testing.assert_allclose(res["peer"][1], res["nccl"][1], rtol=1e-4, atol=1e-6)   # parameters after 6 Adam steps
    for name in ("peer", "peer_nodrop", "peer_raw"):                                       # replicas stay bit-identical
        for rank in range(1, world):
            np.testing.assert_array_equal(out[0][1][name][1], out[rank][1][name][1])
            assert out[0][1][name][0] == out[rank][1][name][0]
    single = out[0][1]["single"]
    np.testing.assert_allclose(out[0][1]["peer_nodrop"][0], single[0], rtol=2e-5)          # W ranks == one GPU, all rows
    np.testing.assert_allclose(out[0][1]["peer_nodrop"][1], single[1], rtol=2e-4, atol=2e-6)
    # raw columns standardised inside the step == the standardised rows (x*7+50 is undone by the statistics up to scale)
    assert abs(out[0][1]["peer_raw"][0][0] - out[0][1]["peer_nodrop"][0][0]) < 0.2


def _small_worker(rank, world, port, q):
    """The reference's own regime on GPUs (batch 4 per rank, train_lightning_ddp.py:122, WORLD_SIZE ranks): many optimizer
    steps per launch with the DDP exchange inside the kernel, against the same steps taken one by one through the NCCL
    bucket path (same rows, same Philox streams)."""
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    import wxddp
    from wxddp.engine import StepEngine

    f, batch, steps, n = 5, 4, 300, 4000
    g = torch.Generator().manual_seed(7)
    x = torch.randn(n, f, generator=g).to(dev)
    y = (torch.rand(n, generator=g) < 0.3).long().to(dev)
    order = torch.randperm(n, generator=torch.Generator().manual_seed(100 + rank))[:steps * batch].to(dev)
    out = {}
    for name in ("nccl", "fused"):
        torch.manual_seed(5)
        eng = StepEngine(f, device=dev, world=world, rank=rank, process_group=dist)
        losses = torch.zeros(steps, device=dev)
        if name == "fused":
            eng.connect_peers()
            eng.steps_small(x, y, order, steps, batch, losses)
            torch.cuda.synchronize()
            t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
            dist.barrier()
            t0.record()
            eng.steps_small(x, y, order, steps, batch, torch.zeros(steps, device=dev))      # timed second launch
            t1.record()
            torch.cuda.synchronize()
            us = 1e3 * t0.elapsed_time(t1) / steps
            eng.check()
            out["us_per_step"] = us
            out["state"] = int(eng.state[0].item())
        else:
            for s_ in range(steps):
                l = eng.step(x, y, row_idx=order[s_ * batch:(s_ + 1) * batch].contiguous(), dropout=wxddp.DROPOUT_PHILOX)
                losses[s_] = l[0]
        if name == "nccl":
            out["nccl"] = (losses.cpu().numpy(), eng.params.cpu().numpy())
        else:
            out["fused_losses"] = losses.cpu().numpy()
        dist.barrier()
    # parameters after the FIRST fused launch are compared: rerun it on a fresh engine (the timed launch moved them on)
    torch.manual_seed(5)
    eng = StepEngine(f, device=dev, world=world, rank=rank, process_group=dist)
    eng.connect_peers()
    eng.steps_small(x, y, order, steps, batch, torch.zeros(steps, device=dev))
    torch.cuda.synchronize()
    out["fused_params"] = eng.params.cpu().numpy()
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_reference_regime_many_steps_per_launch_with_in_kernel_ddp(world):
    _need(world)
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_small_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=900) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, o in out:
        assert o["state"] == 600                                                       # two launches of 300 steps
        np.testing.assert_allclose(o["fused_losses"], o["nccl"][0], rtol=2e-4, atol=1e-6)   # rank-mean train_loss, 300 steps
        d = np.abs(o["fused_params"] - o["nccl"][1])
        assert np.quantile(d, 0.99) <= 2e-3 and d.max() <= 0.05                        # free-running Adam, see SURVEY 7(4)
    for rank in range(1, world):
        np.testing.assert_array_equal(out[0][1]["fused_params"], out[rank][1]["fused_params"])   # replicas bit-identical
    print(f"world {world}: {out[0][1]['us_per_step']:.2f} us per optimizer step (batch 4 per rank, in-kernel exchange)")
    assert out[0][1]["us_per_step"] < 25.0


def _wide_worker(rank, world, port, q):
    """Wide MLP under DDP: per-layer all-reduce overlapped with backward vs one all-reduce at the end; replicas equal;
    the reduced gradient equals the oracle's gradient of the concatenated batch (dropout off: masks are per-rank)."""
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from wxddp.wide import WideEngine

    f, h, nl, b = 20, 256, 2, 256
    torch.manual_seed(42)
    eng = WideEngine(f, h, nl, max_batch=b, device=rank, world=world, rank=rank, process_group=dist)
    g = torch.Generator().manual_seed(100)
    x_all = torch.randn(world * b, f, generator=g)
    y_all = (torch.rand(world * b, generator=g) < 0.3).long()
    x, y = x_all[rank * b:(rank + 1) * b].cuda(), y_all[rank * b:(rank + 1) * b].cuda()
    p0 = eng.params.cpu().clone()
    g_overlap = eng.loss_and_grads(x, y, dropout=0, overlap_reduce=True).clone()
    g_local = eng.loss_and_grads(x, y, dropout=0, overlap_reduce=False).clone()
    dist.all_reduce(g_local)
    losses = [float(eng.step(x, y).item()) for _ in range(3)]
    q.put((rank, p0.numpy(), g_overlap.cpu().numpy(), g_local.cpu().numpy(), losses, eng.params.cpu().numpy(),
           x_all.numpy(), y_all.numpy(), eng.status()))
    dist.destroy_process_group()


_PORT = {}


def _free_port_once() -> int:
    if "wide" not in _PORT:
        _PORT["wide"] = _free_port()
    return _PORT["wide"]


def test_wide_ddp_overlapped_reduce():
    _need(2)
    import torch.multiprocessing as mp
    from oracle import wide_ref

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_wide_worker, args=(r, 2, _free_port_once(), q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=600) for _ in range(2)], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    (r0, r1) = res
    assert r0[8] == 0 and r1[8] == 0
    np.testing.assert_array_equal(r0[1], r1[1])                      # broadcast initial parameters
    np.testing.assert_array_equal(r0[2], r0[3])                      # overlapped == reduce-at-the-end, bit for bit
    np.testing.assert_array_equal(r0[2], r1[2])                      # every rank holds the same reduced bucket
    np.testing.assert_array_equal(r0[5], r1[5])                      # replicas stay identical after 3 steps
    assert r0[4] == r1[4]
    want = wide_ref.loss_and_grads(torch.from_numpy(r0[1]), torch.from_numpy(r0[6]), torch.from_numpy(r0[7]), 20, 256, 2,
                                   dropout=False)
    got = torch.from_numpy(r0[2])
    np.testing.assert_allclose(float(got[-1]), float(want[-1]), rtol=1e-4)
    assert float((got[:-1] - want[:-1]).norm() / want[:-1].norm()) < 1e-3 + 1.5 / 512


def test_two_contexts_on_two_devices_in_one_process():
    """The opt-in to large dynamic shared memory is a per-device function attribute: a second context on another GPU of
    the same process must be able to launch the tensor-core kernels too."""
    _need(2)
    import wxddp

    outs = []
    g = torch.Generator().manual_seed(3)
    x = torch.randn(4096, 20, generator=g); y = (torch.rand(4096, generator=g) < 0.3).long()
    p = torch.randn(64 * 20 + 194, generator=g) * 0.2
    for d in (0, 1):
        c = wxddp.Context(20, device=d)
        dev = torch.device("cuda", d)
        with torch.cuda.device(dev):
            outs.append(c.tc_mlp_grad(p.to(dev), x.to(dev), y.to(dev), dropout=wxddp.DROPOUT_PHILOX, seed=5, step=1).cpu())
            assert c.tc_status() == 0
    np.testing.assert_array_equal(outs[0].numpy(), outs[1].numpy())
