"""Data-parallel correctness on real GPUs (SURVEY §8c last tolerance, §8e): two ranks over NCCL.

  * after a graphed G+D step under comm.DistributedModel, the gradients the optimizers consumed equal the
    mean of the per-rank local gradients (recomputed without synchronisation from the same state);
  * after three replays of the captured step (all-reduce hooks inside the CUDA graph) the parameters of
    the two ranks are bit-identical.
Skipped on boxes with fewer than two GPUs (the driver's 1-GPU test tier); run with `gpurun --gpus 2`.
"""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch.distributed as dist
    try:
        torch.cuda.set_device(rank)
        from bias_gan_b200 import kernels as K
        from bias_gan_b200.comm import comm as Comm
        from bias_gan_b200.deeplab_gan import Discriminator, Generator
        from bias_gan_b200.losses import GANLoss, L1LossField
        from bias_gan_b200.step import GraphedGanStep, gan_step
        from oracle.seeded import seed_state, seeded_fields
        cm = Comm(backend="nccl")
        dev = torch.device("cuda", rank)
        Bn, C, H, W = 2, 16, 64, 96
        x, y = seeded_fields(40 + rank, Bn, C, H, W)      # different data per rank
        xf, yf = K.nchw_to_field(x.to(dev)), K.nchw_to_field(y.to(dev))
        opt = lambda m: torch.optim.AdamW(m.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5, fused=True, capturable=True)

        def build(sync):
            G = seed_state(Generator(C, C, "Interpolate", "Normal", 0), 21 + (0 if sync else 0)).to(dev).train()
            D = seed_state(Discriminator(C, input_size=(H, W)), 22).to(dev).train()
            if sync:
                cm.DistributedModel(G)
                cm.DistributedModel(D)
            return G, D

        class Spy(torch.optim.SGD):   # sees the gradients, leaves the parameters alone
            def step(self, closure=None):
                self.seen = [p.grad.detach().clone() for g in self.param_groups for p in g["params"]]

        # (1) averaged gradients = mean of local gradients (kernels are deterministic, so the local
        # recomputation from the same state reproduces exactly what each rank fed the all-reduce)
        outs = {}
        for sync in (True, False):
            G, D = build(sync)
            go, do = Spy(G.parameters(), lr=0.0), Spy(D.parameters(), lr=0.0)
            torch.manual_seed(5)   # same label noise on both ranks and in both runs
            gan_step(G, D, go, do, GANLoss("ModifiedMinMax", Bn, dev), L1LossField(), xf, yf)
            outs[sync] = (go.seen, do.seen)
        for synced, local in zip(outs[True], outs[False]):
            for a, g in zip(synced, local):
                parts = [torch.zeros_like(g) for _ in range(world)]
                dist.all_gather(parts, g.contiguous())
                want = sum(parts) / world
                torch.testing.assert_close(a, want, rtol=1e-5, atol=1e-6 * float(want.abs().max() + 1e-30) + 1e-12)
        # (2) parameters stay identical across ranks over graphed steps with the all-reduce captured
        G, D = build(True)
        go, do = opt(G), opt(D)
        torch.manual_seed(5)
        gs = GraphedGanStep(G, D, go, do, GANLoss("ModifiedMinMax", Bn, dev), L1LossField(), tuple(xf.shape), dev)
        for _ in range(3):
            dl, gl = gs(xf, yf)
        torch.cuda.synchronize()
        assert bool(torch.isfinite(dl)) and bool(torch.isfinite(gl))
        for m in (G, D):
            for n, p in m.named_parameters():
                parts = [torch.zeros_like(p) for _ in range(world)]
                dist.all_gather(parts, p.detach().contiguous())
                assert torch.equal(parts[0], parts[1]), "parameter %s differs between ranks" % n
        moved = max(float((p - q).abs().max()) for p, q in zip(G.parameters(), build(False)[0].parameters()))
        assert moved > 1e-5, "three optimizer steps did not move the generator"
        q.put((rank, "ok"))
        q.close(); q.join_thread()
        # no destroy_process_group(): tearing NCCL down while CUDA graphs that captured its kernels are alive can
        # block for minutes, and a worker that never exits keeps the pytest process from exiting
        os._exit(0)
    except Exception:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))
        q.close(); q.join_thread()
        os._exit(1)     # the peer may be blocked in a collective: do not wait for it in destroy_process_group


@pytest.mark.timeout(900)
def test_two_rank_nccl_gradients_and_parameters():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = []
    try:
        for _ in procs:
            res.append(q.get(timeout=600))
            if res[-1][1] != "ok":      # a failed rank leaves its peer blocked in a collective: stop both now
                break
    finally:
        for p in procs:
            p.join(timeout=20 if len(res) == 2 and all(r[1] == "ok" for r in res) else 1)
            if p.is_alive():
                p.kill()
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
