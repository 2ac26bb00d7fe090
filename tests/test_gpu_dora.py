"""DoRA (conceptmod/textsliders/dora.py:52-330, `--peft_type dora` of train_lora_xl.py:87-90) on the CUDA path:
forward against the vectors produced by the reference's own DoRAModule (tests/golden/reference_dora.pt), gradients
(dx, lora_down, lora_up, dora_scale) against fp32 autograd of the oracle restatement of the reference forward."""
from pathlib import Path

import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu

GOLD = torch.load(Path(__file__).parent / "golden" / "reference_dora.pt", weights_only=False)["dora_cases"]


def _module(c, dtype):
    from sliders_conceptmod_b200.dora import DoRAModule
    W, A = c["W"], c["A"]
    if c["kind"] == "linear":
        org = nn.Linear(W.shape[1], W.shape[0], bias=c["bias"] is not None)
    else:
        k = W.shape[-1]
        org = nn.Conv2d(W.shape[1], W.shape[0], k, padding=(k - 1) // 2, bias=c["bias"] is not None)
    with torch.no_grad():
        org.weight.copy_(W)
        if c["bias"] is not None:
            org.bias.copy_(c["bias"])
    m = DoRAModule("g", org, multiplier=c["mult"], lora_dim=A.shape[0], alpha=c["scale"] * A.shape[0])
    with torch.no_grad():
        m.lora_down.weight.copy_(A); m.lora_up.weight.copy_(c["B"]); m.dora_scale.copy_(c["dora_scale"])
    org.requires_grad_(False)
    m.apply_to()
    org.to("cuda", dtype); m.to("cuda", dtype)
    return org, m


@pytest.mark.parametrize("i", range(len(GOLD)))
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-3), (torch.bfloat16, 4e-2)])
def test_dora_module_matches_reference_vectors(i, dtype, tol):
    c = GOLD[i]
    org, m = _module(c, dtype)
    assert abs(m.scale - c["scale"]) < 1e-12
    with torch.no_grad():
        y = org(c["x"].to("cuda", dtype))
    err = (y.float().cpu() - c["y"]).abs().max().item()
    assert err <= tol * c["y"].abs().max().item(), f"{c['kind']}: max-abs {err:.3e}"
    # adapter off (multiplier 0, outside `with network:`): the frozen base layer
    m.multiplier = 0
    with torch.no_grad():
        y0 = org(c["x"].to("cuda", dtype))
    ref0 = torch.nn.functional.linear(c["x"], c["W"], c["bias"]) if c["kind"] == "linear" else \
        torch.nn.functional.conv2d(c["x"], c["W"], c["bias"], padding=(c["W"].shape[-1] - 1) // 2)
    assert (y0.float().cpu() - ref0).abs().max().item() <= tol * ref0.abs().max().item()


@pytest.mark.parametrize("i", range(len(GOLD)))
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 3e-3), (torch.bfloat16, 5e-2)])
def test_dora_gradients_match_oracle_autograd(i, dtype, tol):
    """dx, d lora_down, d lora_up, d dora_scale against autograd through oracle.dora_forward (fp32, on the values the
    CUDA module actually holds after the dtype cast)."""
    from oracle import slider_oracle as O
    c = GOLD[i]
    org, m = _module(c, dtype)
    conv = c["W"].dim() == 4
    g = torch.Generator().manual_seed(100 + i)
    x = c["x"].to("cuda", dtype).requires_grad_(True)
    y = org(x)
    gy = torch.randn(y.shape, generator=g).to(dtype)
    y.backward(gy.cuda())
    # oracle on the same (rounded) operands
    f = lambda t: t.detach().float().cpu()
    xr = f(x).requires_grad_(True)
    Ar, Br, Mr = f(m.lora_down.weight).requires_grad_(True), f(m.lora_up.weight).requires_grad_(True), f(m.dora_scale).requires_grad_(True)
    bias = None if c["bias"] is None else f(org.bias)
    yr = O.dora_forward(xr, f(org.weight), bias, Ar, Br, Mr, c["mult"], m.scale, stride=1 if not conv else c["stride"],
                        padding=0 if not conv else c["padding"])
    yr.backward(gy.float())

    def close(got, ref, name):
        got = got.float().cpu()
        err = (got - ref).abs().max().item()
        cos = torch.nn.functional.cosine_similarity(got.flatten(), ref.flatten(), dim=0).item()
        assert err <= tol * ref.abs().max().item() + 1e-7 and cos > 0.999, f"{c['kind']} {name}: err {err:.3e} scale {ref.abs().max().item():.3e} cos {cos:.5f}"

    close(x.grad, xr.grad, "dx")
    close(m.lora_down.weight.grad, Ar.grad, "dA")
    close(m.lora_up.weight.grad, Br.grad, "dB")
    close(m.dora_scale.grad, Mr.grad, "d dora_scale")


def test_dora_network_trains_tiny_unet_like_the_oracle():
    """DoRANetwork on the tiny UNet stand-in: module set / key names equal LoRANetwork's plus dora_scale, the F10
    constructor quirk (alpha lands in `prefix`, scale = 1/rank), and three AdamW steps follow an fp32 replay of the
    reference arithmetic (oracle.dora_forward patched into a CPU copy)."""
    from oracle import slider_oracle as O
    from sliders_conceptmod_b200 import model_util
    from sliders_conceptmod_b200.dora import DoRANetwork
    from sliders_conceptmod_b200.lora import LoRANetwork
    from sliders_conceptmod_b200.unet import UNet2DConditionModel, UNetConfig
    cfg = UNetConfig.tiny()
    torch.manual_seed(3)
    unet_g = UNet2DConditionModel(cfg)
    unet_c = UNet2DConditionModel(cfg)
    unet_c.load_state_dict(unet_g.state_dict())
    unet_g.requires_grad_(False); unet_c.requires_grad_(False)
    torch.manual_seed(5)
    net_g = DoRANetwork(unet_g, rank=4, multiplier=1.0, delimiter="_", alpha=8.0, target_replace=["Attention"],
                        prefix="lora_unet", train_method="noxattn")
    names = [m.lora_name for m in net_g.unet_loras]
    torch.manual_seed(5)
    ref_names = [m.lora_name for m in LoRANetwork(unet_c, rank=4, multiplier=1.0, delimiter="_", alpha=8.0,
                                                  target_replace=["Attention"], train_method="noxattn").unet_loras]
    assert names == ref_names and len(names) > 0
    assert all(abs(m.scale - 0.25) < 1e-12 for m in net_g.unet_loras)          # alpha = 1 whatever was passed (F10)
    keys = set(net_g.state_dict().keys())
    assert any(k.endswith("dora_scale") for k in keys) and any(k.endswith("lora_down.weight") for k in keys)
    # CPU oracle twin: same parameters, reference arithmetic
    unet_o = UNet2DConditionModel(cfg)
    unet_o.load_state_dict(unet_g.state_dict())
    unet_o.requires_grad_(False)
    torch.manual_seed(5)
    net_o = DoRANetwork(unet_o, rank=4, multiplier=1.0, delimiter="_", alpha=8.0, target_replace=["Attention"],
                        prefix="lora_unet", train_method="noxattn")
    with torch.no_grad():
        for mo, mg in zip(net_o.unet_loras, net_g.unet_loras):
            mg.lora_up.weight.normal_(0, 0.05, generator=torch.Generator().manual_seed(len(mg.lora_name)))
            mo.lora_up.weight.copy_(mg.lora_up.weight); mo.lora_down.weight.copy_(mg.lora_down.weight)
            mo.dora_scale.copy_(mg.dora_scale)
    for mo in net_o.unet_loras:
        org = mo._org
        base_w, base_b = org.weight.detach(), (None if org.bias is None else org.bias.detach())

        def fwd(x, mo=mo, base_w=base_w, base_b=base_b):
            return O.dora_forward(x, base_w, base_b, mo.lora_down.weight, mo.lora_up.weight, mo.dora_scale, mo.multiplier,
                                  mo.scale)
        org.forward = fwd
    unet_g.to("cuda"); net_g.to("cuda")
    opt_g = torch.optim.AdamW(net_g.parameters(), lr=1e-3, weight_decay=1e-6)
    opt_o = torch.optim.AdamW(net_o.parameters(), lr=1e-3, weight_decay=1e-6)
    gen = torch.Generator().manual_seed(9)
    for step in range(3):
        lat = torch.randn(2, 4, 16, 16, generator=gen)
        emb = torch.randn(2, 77, cfg.cross_attention_dim, generator=gen)
        tgt = torch.randn(2, 4, 16, 16, generator=gen)
        t = torch.tensor([500.0])
        losses = []
        for unet, net, opt, dev in ((unet_o, net_o, opt_o, "cpu"), (unet_g, net_g, opt_g, "cuda")):
            opt.zero_grad()
            with net:
                out = unet(lat.to(dev), t.to(dev), encoder_hidden_states=emb.to(dev)).sample
            loss = torch.nn.functional.mse_loss(out.float(), tgt.to(dev))
            loss.backward()
            opt.step()
            losses.append(loss.item())
        assert abs(losses[0] - losses[1]) <= 2e-3 * abs(losses[0]), f"step {step}: oracle {losses[0]} cuda {losses[1]}"
    for mo, mg in zip(net_o.unet_loras, net_g.unet_loras):
        for name in ("lora_down.weight", "lora_up.weight", "dora_scale"):
            po, pg = mo.get_parameter(name), mg.get_parameter(name).cpu()
            cos = torch.nn.functional.cosine_similarity((po - 0).flatten(), pg.flatten(), dim=0).item()
            assert cos > 0.999, f"{mg.lora_name}.{name}: cosine {cos}"
