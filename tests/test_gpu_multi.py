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


@pytest.mark.parametrize("world", [2])
def test_ddp_on_gpus_matches_gloo_ddp_oracle(world):
    _need(world)
    t = synth.synth_table(31, 203, 5)
    pp = preprocess_ref.preprocess_table(t["raw"], t["labels"])
    feats, labels = torch.from_numpy(pp["features"]), torch.from_numpy(pp["labels"])
    got = _run(_nccl_worker, world, 29671, feats, labels)
    ref = _run(_gloo_worker, world, 29673, feats, labels)
    for g, r in zip(got, ref):
        np.testing.assert_allclose(g[1], r[1], rtol=2e-4, atol=1e-6)      # per-step rank-mean train_loss
        np.testing.assert_allclose(g[2], r[2], rtol=2e-4)
        np.testing.assert_allclose(g[3], r[3], rtol=1e-6)
        np.testing.assert_allclose(g[4], r[4], rtol=2e-3, atol=2e-5)
    np.testing.assert_array_equal(got[0][4], got[1][4])                   # replicas stay bit-identical


def _peer_worker(rank, world, port, q):
    """Fused in-kernel exchange (NVLink peer memory) vs the NCCL bucket path, same data, same dropout bits."""
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    import wxddp
    from wxddp.engine import StepEngine

    f, b, steps = 20, 4096 + 128 * rank, 6        # per-rank batches of different size are fine for the exchange itself
    b = 4096
    g = torch.Generator().manual_seed(100 + rank)
    x = torch.randn(steps * b, f, generator=g).to(dev); y = (torch.rand(steps * b, generator=g) < 0.3).long().to(dev)
    torch.manual_seed(5)
    res = {}
    for name, use_peers in (("nccl", False), ("peer", True)):
        torch.manual_seed(5)
        eng = StepEngine(f, device=dev, world=world, rank=rank, process_group=dist)
        if use_peers:
            eng.connect_peers()
        losses = []
        for s in range(steps):
            l = eng.step(x, y, row0=s * b, batch=b, dropout=wxddp.DROPOUT_PHILOX)
            losses.append(float(l.item()))
        torch.cuda.synchronize()
        res[name] = (losses, eng.params.cpu().numpy(), eng.ctx.tc_status(), int(eng.state[0].item()) if use_peers else steps)
        dist.barrier()
    q.put((rank, res))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_in_kernel_peer_allreduce_matches_nccl_bucket_path(world):
    _need(world)
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_peer_worker, args=(r, world, 29691, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=600) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, res in out:
        assert res["peer"][2] == 0 and res["peer"][3] == 6
        np.testing.assert_allclose(res["peer"][0], res["nccl"][0], rtol=1e-5)              # rank-mean train_loss per step
        np.testing.assert_allclose(res["peer"][1], res["nccl"][1], rtol=1e-4, atol=1e-6)   # parameters after 6 Adam steps
    np.testing.assert_array_equal(out[0][1]["peer"][1], out[1][1]["peer"][1])              # replicas bit-identical
    np.testing.assert_allclose(out[0][1]["peer"][0], out[1][1]["peer"][0], rtol=0, atol=0)


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


def test_wide_ddp_overlapped_reduce():
    _need(2)
    import torch.multiprocessing as mp
    from oracle import wide_ref

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_wide_worker, args=(r, 2, 29681, q)) for r in range(2)]
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
