"""Pins the CPU oracle (oracle/slider_oracle.py) to vectors produced by the reference's own classes
(tests/golden/make_golden.py).  CPU only."""
from pathlib import Path

import pytest
import torch
import torch.nn as nn

from helpers import ROOT  # noqa: F401  (path setup)
from oracle import slider_oracle as O

GOLD = torch.load(Path(__file__).parent / "golden" / "reference_vectors.pt", weights_only=False)


def _modules(c):
    W, A, B = c["W"], c["A"], c["B"]
    if c["kind"] == "linear":
        org = nn.Linear(W.shape[1], W.shape[0], bias=c["bias"] is not None)
        down, up = nn.Linear(A.shape[1], A.shape[0], bias=False), nn.Linear(B.shape[1], B.shape[0], bias=False)
    else:
        k = W.shape[-1]
        stride = 2 if c["kind"] == "conv3s2" else 1
        org = nn.Conv2d(W.shape[1], W.shape[0], k, stride=stride, padding=(k - 1) // 2, bias=c["bias"] is not None)
        down = nn.Conv2d(A.shape[1], A.shape[0], k, stride=stride, padding=(k - 1) // 2, bias=False)
        up = nn.Conv2d(B.shape[1], B.shape[0], 1, bias=False)
    with torch.no_grad():
        org.weight.copy_(W); down.weight.copy_(A); up.weight.copy_(B)
        if c["bias"] is not None:
            org.bias.copy_(c["bias"])
    return org, down, up


@pytest.mark.parametrize("i", range(len(GOLD["lora_cases"])))
def test_lora_forward_backward(i):
    c = GOLD["lora_cases"][i]
    org, down, up = _modules(c)
    x = c["x"].clone().requires_grad_(True)
    y = O.lora_forward(org.forward, down, up, c["mult"], c["scale"], x)
    assert torch.equal(y, c["y"])
    (y * c["gy"]).sum().backward()
    assert torch.allclose(x.grad, c["dx"], rtol=0, atol=0)
    assert torch.equal(down.weight.grad, c["dA"]) and torch.equal(up.weight.grad, c["dB"])


@pytest.mark.parametrize("i", range(len(GOLD["loss_cases"])))
def test_guidance_loss(i):
    c = GOLD["loss_cases"][i]
    t = c["t"].clone().requires_grad_(True)
    loss = O.guidance_loss(t, c["p"], c["n"], c["q"], c["guidance"], c["action"])
    loss.backward()
    assert torch.equal(loss.detach(), c["loss"]) and torch.equal(t.grad, c["grad"])


@pytest.mark.parametrize("i", range(len(GOLD["ddim_cases"])))
def test_ddim_prev_step(i):
    c = GOLD["ddim_cases"][i]
    s = O.OracleScheduler("ddim")
    s.set_timesteps(c["n_inf"])
    assert torch.equal(s.step(c["eps"], c["t"], c["x"]).prev_sample, c["prev"])


@pytest.mark.parametrize("i", range(len(GOLD["cfg_cases"])))
def test_cfg(i):
    c = GOLD["cfg_cases"][i]
    assert torch.equal(O.cfg_combine(c["eps2"], c["guidance"]), c["guided"])


def test_scheduler_tables():
    s = O.OracleScheduler("ddim")
    s.set_timesteps(50)
    assert s.timesteps[:3].tolist() == [980, 960, 940] and s.timesteps[-1].item() == 0
    s.set_timesteps(1000)
    # train_lora.py:209-213: current_timestep = timesteps[int(timesteps_to*1000/50)] = 999 - 20*timesteps_to
    for tt in (1, 25, 49):
        assert s.timesteps[int(tt * 1000 / 50)].item() == 999 - 20 * tt
    ac = O.alphas_cumprod()
    assert abs(ac[0].item() - 0.99915) < 1e-5 and abs(ac[-1].item() - 0.0046601) < 1e-5


def test_train_loop_matches_reference_minitrain():
    """The oracle's train_iteration, iterated with torch AdamW, against the losses and final LoRA weights produced by
    the reference's own unmodified `train()` loop (tests/golden/make_golden_minitrain.py)."""
    from sliders_conceptmod_b200 import model_util, train_util, workload
    from sliders_conceptmod_b200.lora import LoRANetwork
    gold = torch.load(Path(__file__).parent / "golden" / "reference_minitrain.pt", weights_only=False)
    cfg = gold["config"]
    _, enc, unet, _ = model_util.load_models("tiny-sd", scheduler_name="ddim")
    unet.eval().requires_grad_(False)
    checksum = float(sum(p.detach().double().abs().sum() for p in unet.parameters()))
    assert abs(checksum - gold["unet_checksum"]) < 1e-6 * gold["unet_checksum"], "UNet stand-in init changed"
    net = LoRANetwork(unet, rank=4, multiplier=1.0, alpha=1.0, target_replace=workload.target_modules(True),
                      train_method="noxattn")
    assert set(net.state_dict()) == set(gold["init"])
    net.load_state_dict(gold["init"])
    O.attach(net)
    opt = torch.optim.AdamW(net.prepare_optimizer_params(), lr=cfg["train"]["lr"])
    sched = O.OracleScheduler("ddim")
    n_den = cfg["train"]["max_denoising_steps"]
    embs = [{k: enc.encode([p[k]]) for k in ("target", "positive", "neutral", "unconditional")} for p in gold["prompts"]]
    torch.manual_seed(gold["loop_seed"])
    for i, ref_loss in enumerate(gold["losses"]):
        sched.set_timesteps(n_den)
        opt.zero_grad()
        idx = torch.randint(0, len(embs), (1,)).item()
        timesteps_to = torch.randint(1, n_den, (1,)).item()
        p = gold["prompts"][idx]
        latents = train_util.get_initial_latents(sched, p["batch_size"], p["resolution"], p["resolution"], 1)
        loss, _ = O.train_iteration(unet, net, sched, embs[idx], latents, timesteps_to, p["guidance_scale"], p["action"],
                                    max_denoising_steps=n_den)
        opt.step()
        assert abs(loss.item() - ref_loss) <= 1e-5 * abs(ref_loss), f"iteration {i}: {loss.item()} vs {ref_loss}"
    sd = net.state_dict()
    for k, v in gold["final"].items():
        assert torch.allclose(sd[k], v, rtol=1e-4, atol=1e-6), k


def test_dora_oracle_and_fold_match_reference_vectors():
    """DoRA ("next" row): the oracle restatement of dora.py:122-163 reproduces the reference module's
    outputs (tests/golden/make_golden_dora.py), and the per-input-channel fold of sliders_conceptmod_b200/dora.py —
    one fused base + low-rank pass on the pre-scaled input — equals it."""
    import torch.nn.functional as F
    from sliders_conceptmod_b200 import dora
    cases = torch.load(Path(__file__).parent / "golden" / "reference_dora.pt", weights_only=False)["dora_cases"]
    assert len(cases) == 4
    for c in cases:
        conv = c["W"].dim() == 4
        stride, padding = (c["stride"], c["padding"]) if conv else (1, 0)
        y = O.dora_forward(c["x"], c["W"], c["bias"], c["A"], c["B"], c["dora_scale"], c["mult"], c["scale"],
                           stride=stride, padding=padding)
        assert torch.allclose(y, c["y"], rtol=0, atol=1e-6), c["kind"]
        s = c["scale"] * c["mult"]
        v, a2 = dora.dora_fold(c["W"].double(), c["A"].double(), c["B"].double(), c["dora_scale"].double(), s)
        x = c["x"].double()
        bias = None if c["bias"] is None else c["bias"].double()
        if conv:
            xv = x * v.reshape(1, -1, 1, 1)
            h = F.conv2d(xv, a2, stride=stride, padding=padding)
            yf = F.conv2d(xv, c["W"].double(), bias, stride=stride, padding=padding) + F.conv2d(h, c["B"].double())
        else:
            xv = x * v
            yf = F.linear(xv, c["W"].double(), bias) + F.linear(F.linear(xv, a2), c["B"].double())
        assert torch.allclose(yf.float(), c["y"], rtol=1e-5, atol=2e-5), (c["kind"], (yf.float() - c["y"]).abs().max())
        # operand-side form (what ops.dora_prepare computes on the GPU): W_eff = W*v, A_s = A*(s c), adapter scale 1
        w_eff, a_s, _, _ = dora.dora_operands(c["W"].double(), c["A"].double(), c["B"].double(), c["dora_scale"].double(), s)
        if conv:
            yo = F.conv2d(x, w_eff, bias, stride=stride, padding=padding) + \
                F.conv2d(F.conv2d(x, a_s, stride=stride, padding=padding), c["B"].double())
        else:
            yo = F.linear(x, w_eff, bias) + F.linear(F.linear(x, a_s), c["B"].double())
        assert torch.allclose(yo.float(), c["y"], rtol=1e-5, atol=2e-5), (c["kind"], (yo.float() - c["y"]).abs().max())
