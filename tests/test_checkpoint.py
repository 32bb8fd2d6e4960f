"""Checkpoint interchange (comm/distributed.py:130-157, train_gan.py:401-431): a `.cpt`
in the reference's dictionary format with DDP-prefixed keys, written from the *oracle's* modules,
restores the product's modules and optimizers; the product's own checkpoints load back into the oracle."""
import torch

from oracle import ref_model as R
from oracle.seeded import seed_state

C, H, W = 16, 64, 96


def _opt(m):
    return torch.optim.AdamW(m.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5)


def test_reference_checkpoint_round_trip(tmp_path):
    from bias_gan_b200.comm import comm
    from bias_gan_b200.deeplab_gan import Discriminator, Generator
    Gr, Dr = seed_state(R.GeneratorRef(C, C), 31), seed_state(R.DiscriminatorRef(C, (H, W)), 32)
    gro, dro = _opt(Gr), _opt(Dr)
    for m, o in ((Gr, gro), (Dr, dro)):          # give the optimizers state to carry
        for p in m.parameters():
            p.grad = torch.full_like(p, 1e-3)
        o.step()
    ckpt = {"step": 17, "epoch": 3, "generator": {"module." + k: v for k, v in Gr.state_dict().items()},
            "discriminator": {"module." + k: v for k, v in Dr.state_dict().items()},
            "g_opt": gro.state_dict(), "d_opt": dro.state_dict(), "amp": {}}
    path = str(tmp_path / "ref_step_17.cpt")
    torch.save(ckpt, path)

    c = comm(mode="dummy")
    Gp, Dp = Generator(C, C, "Interpolate", "Normal", 0), Discriminator(C, input_size=(H, W))
    gpo, dpo = _opt(Gp), _opt(Dp)
    assert c.init_gan_training_state(Gp, Dp, gpo, dpo, path, "cpu") == (17, 3)
    for prod, ref in ((Gp, Gr), (Dp, Dr)):
        sp, sr = prod.state_dict(), ref.state_dict()
        assert list(sp.keys()) == list(sr.keys())                      # names and order are the contract
        assert all(torch.equal(sp[k], sr[k]) for k in sp)
    for po, ro in ((gpo, gro), (dpo, dro)):
        a, b = po.state_dict()["state"], ro.state_dict()["state"]
        assert a.keys() == b.keys() and all(torch.equal(a[k]["exp_avg"], b[k]["exp_avg"]) for k in a)
    assert c.init_gan_training_state(Gp, Dp, gpo, dpo, str(tmp_path / "missing.cpt"), "cpu") == (0, 0)

    out = str(tmp_path / "prod_step_18.cpt")
    c.save_gan_checkpoint(out, 18, 3, Gp, Dp, gpo, dpo)
    back = torch.load(out, map_location="cpu")
    assert set(back) == {"step", "epoch", "generator", "discriminator", "g_opt", "d_opt", "amp"}
    G2, D2 = R.GeneratorRef(C, C), R.DiscriminatorRef(C, (H, W))
    G2.load_state_dict(back["generator"]); D2.load_state_dict(back["discriminator"])   # strict: every key matches
    assert all(torch.equal(a, b) for a, b in zip(G2.state_dict().values(), Gr.state_dict().values()))
