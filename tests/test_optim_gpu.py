"""GPU: FusedAdamW (moss_adamw_step through the C ABI) against torch.optim.AdamW configured like the reference
(scene/gaussian_model.py:203-225: six groups with their own lr, lr=0.0 default, eps=1e-15), on the same seeded
parameters and gradients over several steps, including a learning-rate change between steps (the xyz scheduler,
scene/gaussian_model.py:238-246) and a state edit of the kind densification performs."""
import pytest
import torch

from moss_b200.optim import FusedAdamW

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SHAPES = {"xyz": (5003, 3), "f_dc": (5003, 1, 3), "f_rest": (5003, 15, 3), "opacity": (5003, 1), "scaling": (5003, 3),
          "rotation": (5003, 4)}
LRS = {"xyz": 1.6e-4 * 5.0, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20.0, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}


def _groups(seed):
    g = torch.Generator().manual_seed(seed)
    return [{"params": [torch.nn.Parameter(torch.randn(*s, generator=g).to(DEV))], "lr": LRS[n], "name": n} for n, s in SHAPES.items()]


def test_fused_adamw_matches_torch_adamw():
    ga, gb = _groups(3), _groups(3)
    ref = torch.optim.AdamW(ga, lr=0.0, eps=1e-15)
    ours = FusedAdamW(gb, lr=0.0, eps=1e-15)
    gen = torch.Generator().manual_seed(9)
    for it in range(6):
        for a, b in zip(ga, gb):
            gr = (torch.randn(a["params"][0].shape, generator=gen) * (10.0 ** (it - 3))).to(DEV)
            if it == 2 and a["name"] == "opacity":
                gr.zero_()                                         # zero gradient: eps = 1e-15 decides the step
            a["params"][0].grad, b["params"][0].grad = gr.clone(), gr.clone()
        if it == 3:                                                # the xyz learning-rate schedule
            for opt in (ref, ours):
                for grp in opt.param_groups:
                    if grp["name"] == "xyz":
                        grp["lr"] = 3.3e-4
        ref.step(); ours.step()
        ref.zero_grad(set_to_none=True); ours.zero_grad(set_to_none=True)
        for a, b in zip(ga, gb):
            pa, pb = a["params"][0], b["params"][0]
            sa, sb = ref.state[pa], ours.state[pb]
            assert int(sa["step"]) == int(sb["step"]) == it + 1
            torch.testing.assert_close(sb["exp_avg"], sa["exp_avg"], rtol=2e-6, atol=1e-30)
            torch.testing.assert_close(sb["exp_avg_sq"], sa["exp_avg_sq"], rtol=2e-6, atol=1e-38)
            # the update is lr-sized: compare the parameters to a fraction of one step
            assert float((pa - pb).abs().max()) <= 2e-6 * max(a["lr"], 1e-6) + 2e-7 * float(pa.abs().max()), a["name"]


def test_fused_adamw_survives_a_densification_style_state_edit():
    """cat_tensors_to_optimizer (scene/gaussian_model.py): the parameter is replaced by a longer one and its moments
    are concatenated with zeros; both optimizers must carry on identically from the edited state."""
    gen = torch.Generator().manual_seed(5)
    p0, g1, ext, g2 = (torch.randn(n, 3, generator=gen).to(DEV) for n in (101, 101, 7, 108))
    finals = []
    for cls in (torch.optim.AdamW, FusedAdamW):
        grp = [{"params": [torch.nn.Parameter(p0.clone())], "lr": 1e-2, "name": "xyz"}]
        opt = cls(grp, lr=0.0, eps=1e-15)
        p = grp[0]["params"][0]
        p.grad = g1.clone()
        opt.step()
        stored = opt.state.pop(p)
        stored["exp_avg"] = torch.cat((stored["exp_avg"], torch.zeros_like(ext)), dim=0)
        stored["exp_avg_sq"] = torch.cat((stored["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
        new_p = torch.nn.Parameter(torch.cat((p.detach(), ext), dim=0))
        opt.param_groups[0]["params"][0] = new_p
        opt.state[new_p] = stored
        new_p.grad = g2.clone()
        opt.step()
        assert int(opt.state[new_p]["step"]) == 2
        finals.append(new_p.detach().clone())
    torch.testing.assert_close(finals[1], finals[0], rtol=1e-6, atol=1e-7)


def test_fused_adamw_handles_unaligned_and_ragged_sizes():
    ps = [torch.nn.Parameter(torch.randn(n, device=DEV)) for n in (1, 3, 1023, 1025, 4099)]
    qs = [torch.nn.Parameter(p.detach().clone()) for p in ps]
    # views at odd offsets: not 16-byte aligned -> the scalar tail path
    base = torch.randn(1000, device=DEV)
    ps.append(torch.nn.Parameter(base[1:998])); qs.append(torch.nn.Parameter(ps[-1].detach().clone()))
    assert ps[-1].data_ptr() % 16 != 0
    ref = torch.optim.AdamW([{"params": qs, "lr": 1e-2}], lr=0.0, eps=1e-15)
    ours = FusedAdamW([{"params": ps, "lr": 1e-2}], lr=0.0, eps=1e-15)
    for _ in range(3):
        for p, q in zip(ps, qs):
            g = torch.randn_like(p)
            p.grad, q.grad = g.clone(), g.clone()
        ref.step(); ours.step()
    for p, q in zip(ps, qs):
        torch.testing.assert_close(p, q, rtol=1e-6, atol=1e-7)
