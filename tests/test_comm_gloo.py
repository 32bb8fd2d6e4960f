"""Host-side data-parallel logic on CPU: world size 2, gloo backend.  Checks that
the bucketed, hook-driven gradient all-reduce yields the mean of the per-rank gradients, that
parameters are broadcast from rank 0, and metric_average's sum/average semantics
(comm/distributed.py:12-34,195-199)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from bias_gan_b200.comm import comm as Comm
    cm = Comm(backend="gloo")
    try:
        assert cm.size() == world and cm.rank() == rank
        torch.manual_seed(100 + rank)  # different init per rank: broadcast must equalise
        model = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.Tanh(), torch.nn.Linear(32, 8),
                                    torch.nn.Linear(8, 1))
        cm.DistributedModel(model, bucket_mb=0)  # tiny buckets: one per parameter, exercised from hooks
        ref = [p.detach().clone() for p in model.parameters()]
        gathered = [torch.zeros_like(ref[0]) for _ in range(world)]
        dist.all_gather(gathered, ref[0])
        assert torch.equal(gathered[0], gathered[1]), "parameters were not broadcast from rank 0"
        torch.manual_seed(7 + rank)
        x = torch.randn(4, 16)
        loss = model(x).pow(2).mean()
        local = torch.autograd.grad(loss, list(model.parameters()), retain_graph=True)
        loss.backward()
        model.grad_sync.finish()
        for p, g in zip(model.parameters(), local):
            parts = [torch.zeros_like(g) for _ in range(world)]
            dist.all_gather(parts, g.contiguous())
            torch.testing.assert_close(p.grad, sum(parts) / world, rtol=1e-6, atol=1e-7)
        # second backward re-arms the buckets
        model.zero_grad()
        model(x).sum().backward()
        model.grad_sync.finish()
        assert all(p.grad is not None for p in model.parameters())
        s = cm.metric_average(torch.tensor(float(rank + 1)))
        a = cm.metric_average(torch.tensor(float(rank + 1)), op_name="average")
        assert s == 3.0 and a == 1.5
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_bucketed_allreduce_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=150) for _ in procs]
    for p in procs:
        p.join(timeout=30)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res


class _TinyG(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a, self.b = torch.nn.Linear(6, 12), torch.nn.Linear(12, 6)

    def forward_field(self, x):
        return self.b(torch.tanh(self.a(x)))


class _TinyD(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a, self.b = torch.nn.Linear(6, 10), torch.nn.Linear(10, 1)

    def forward_field(self, x):
        lo = self.b(torch.tanh(self.a(x)))
        return lo, torch.sigmoid(lo)


def _step_worker(rank, world, port, q, lean):
    """gan_step under DistributedModel: after each phase the stepped network's gradients must be the
    rank-mean of the local gradients of *that phase's* loss -- also on the reference-faithful schedule
    (lean=False), where d_loss.backward() reaches G and g_loss.backward() reaches D."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import copy
    from bias_gan_b200.comm import comm as Comm
    from bias_gan_b200.losses import GANLoss
    from bias_gan_b200.step import gan_step
    cm = Comm(backend="gloo")
    try:
        torch.manual_seed(3)
        G, D = _TinyG(), _TinyD()
        cm.DistributedModel(G, bucket_mb=0)
        cm.DistributedModel(D, bucket_mb=0)
        torch.manual_seed(50 + rank)
        x, y = torch.randn(4, 6), torch.randn(4, 6)
        crit = GANLoss("ModifiedMinMax", 4, "cpu")
        crit.draw_labels()   # static labels: the same in the step and in the local recomputation below
        l1 = lambda a, b: (a - b).abs().mean()

        class Spy(torch.optim.SGD):   # records the gradients the optimizer sees
            def step(self, closure=None):
                self.seen = [p.grad.detach().clone() for g in self.param_groups for p in g["params"]]

        go, do = Spy(G.parameters(), lr=0.1), Spy(D.parameters(), lr=0.1)
        G0, D0 = copy.deepcopy(G), copy.deepcopy(D)
        gan_step(G, D, go, do, crit, l1, x, y, lean=lean)
        # local gradients of each phase's loss (parameters did not move: Spy.step is a no-op)
        fake = G0.forward_field(x)
        d_loss = crit.d_loss(D0.forward_field(y)[0], D0.forward_field(fake.detach())[0])
        d_local = torch.autograd.grad(d_loss, list(D0.parameters()))
        fake = G0.forward_field(x)
        g_loss = crit.g_loss(D0.forward_field(fake)[0]) + l1(fake, y)
        g_local = torch.autograd.grad(g_loss, list(G0.parameters()))
        for seen, local in ((do.seen, d_local), (go.seen, g_local)):
            for got, g in zip(seen, local):
                parts = [torch.zeros_like(g) for _ in range(world)]
                dist.all_gather(parts, g.contiguous())
                torch.testing.assert_close(got, sum(parts) / world, rtol=1e-5, atol=1e-7)
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
@pytest.mark.parametrize("lean", [True, False])
def test_gan_step_syncs_the_phase_gradients_world2_gloo(lean):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_step_worker, args=(r, 2, port, q, lean)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=150) for _ in procs]
    for p in procs:
        p.join(timeout=30)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res


def test_dummy_mode_single_process():
    from bias_gan_b200.comm import comm as Comm
    cm = Comm(mode="dummy")
    assert cm.size() == 1 and cm.rank() == 0
    assert cm.metric_average(torch.tensor(2.5)) == 2.5
