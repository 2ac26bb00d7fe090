"""GPU parity of the bandwidth-bound kernels against the same arithmetic in PyTorch fp32."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n", [16384, 65536, 4 * 4 * 32 * 32 + 3, 1 << 22])
@pytest.mark.parametrize("sign", [1, -1])
def test_guidance_loss(dtype, n, sign):
    from sliders_conceptmod_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(n + sign)
    t, p, nn, q = (torch.randn(n, device="cuda", generator=gen).to(dtype) for _ in range(4))
    g = 4.0
    loss, grad = ops.guidance_loss(t, p, nn, q, g, sign)
    tgt = nn.float() + sign * g * (p.float() - q.float())
    d = t.float() - tgt
    ref = (d * d).mean()
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
    gref = 2 * d / n
    tol = 1e-6 if dtype == torch.float32 else 1e-2
    assert (grad.float() - gref).abs().max().item() <= tol * gref.abs().max().item() + 1e-12
    loss2, _ = ops.guidance_loss(t, p, nn, q, g, sign, want_grad=False)
    assert loss2.item() == loss.item()  # deterministic reduction
    lm, gm = ops.guidance_loss(t, None, nn, None, 0.0, 1)
    dm = t.float() - nn.float()
    assert abs(lm.item() - (dm * dm).mean().item()) <= 1e-5 * (dm * dm).mean().item()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n", [16384, 65536 + 5])
def test_cfg_step(dtype, n):
    from sliders_conceptmod_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(n)
    x, eu, ec, nz = (torch.randn(n, device="cuda", generator=gen).to(dtype) for _ in range(4))
    g, cx, ce, cn = 7.5, 1.01, -0.03, 0.2
    tol = 1e-6 if dtype == torch.float32 else 1e-2
    eps = eu.float() + g * (ec.float() - eu.float())
    out = ops.cfg_step(x, eu, ec, nz, g, cx, ce, cn)
    ref = cx * x.float() + ce * eps + cn * nz.float()
    assert (out.float() - ref).abs().max().item() <= tol * ref.abs().max().item()
    out = ops.cfg_step(None, eu, ec, guidance=g)
    assert (out.float() - eps).abs().max().item() <= tol * eps.abs().max().item()
    out = ops.cfg_step(x, None, None, nz, c_x=0.7, c_noise=0.3)
    ref = 0.7 * x.float() + 0.3 * nz.float()
    assert (out.float() - ref).abs().max().item() <= tol * ref.abs().max().item()
    coef = torch.tensor([g, cx, ce, cn], device="cuda")
    out = ops.cfg_step(x, eu, ec, nz, 0, 0, 0, 0, dev_coef=coef)
    ref = cx * x.float() + ce * eps + cn * nz.float()
    assert (out.float() - ref).abs().max().item() <= tol * ref.abs().max().item()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_adamw_flat_and_clip(dtype):
    from sliders_conceptmod_b200 import ops
    n = 100003
    gen = torch.Generator(device="cuda").manual_seed(7)
    p0 = torch.randn(n, device="cuda", generator=gen)
    p = p0.to(dtype).clone()
    pref = torch.nn.Parameter(p.float().clone())
    opt = torch.optim.AdamW([pref], lr=2e-3, weight_decay=1e-2)
    m = torch.zeros(n, device="cuda"); v = torch.zeros(n, device="cuda")
    for step in range(1, 4):
        g = torch.randn(n, device="cuda", generator=gen)
        pref.grad = g.clone()
        opt.step()
        ops.adamw_flat(p, g, m, v, 2e-3, 0.9, 0.999, 1e-8, 1e-2, step)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert (p.float() - pref.detach()).abs().max().item() <= tol
    g = torch.randn(n, device="cuda", generator=gen)
    c = ops.clip_coef(g, 0.2, 0.5)
    ref = min(1.0, 0.2 / (g.norm().item() * 0.5 + 1e-6)) * 0.5
    assert abs(c.item() - ref) <= 1e-5 * ref


@pytest.mark.parametrize("pred", ["epsilon", "v_prediction"])
def test_lms_trajectory_matches_oracle(pred):
    """LMSDiscreteScheduler (model_util.py:415-423; SD-1 inference default, generate_images_sd1.py:51): 10 fused
    CFG + multistep updates on the device against the oracle's restatement of the diffusers rule on the CPU (fp32;
    the model output is a fixed function of the latents so that errors would compound)."""
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
    from oracle import slider_oracle as O
    from sliders_conceptmod_b200 import model_util
    s = model_util.create_noise_scheduler("lms", prediction_type=pred)
    o = O.OracleScheduler("lms", prediction_type=pred)
    s.set_timesteps(10); o.set_timesteps(10)
    g = torch.Generator().manual_seed(3)
    x_c = torch.randn(2, 4, 16, 16, generator=g) * float(o.init_noise_sigma)
    w_u, w_c = torch.randn(4, 4, generator=g) * 0.3, torch.randn(4, 4, generator=g) * 0.3
    x_d = x_c.cuda()
    guidance = 3.0

    def model(x, w):      # channel mix, bounded: stands in for the UNet
        return torch.tanh(torch.einsum("oc,bchw->bohw", w.to(x.device), x))

    for t in s.timesteps:
        xin_c = o.scale_model_input(x_c, t)
        eu, ec = model(xin_c, w_u), model(xin_c, w_c)
        x_c = o.step(eu + guidance * (ec - eu), t, x_c).prev_sample
        xin_d = s.scale_model_input(x_d, t)
        x_d = s.fused_cfg_step(x_d, model(xin_d, w_u).contiguous(), model(xin_d, w_c).contiguous(), guidance, t)
    torch.cuda.synchronize()
    err = (x_d.cpu() - x_c).abs().max().item()
    assert err <= 2e-4 * x_c.abs().max().item(), err
    # the diffusers-style surface: step(model_output) on an already combined prediction
    s.set_timesteps(10)
    y = s.step(torch.ones(1, 4, 8, 8, device="cuda"), s.timesteps[0], torch.zeros(1, 4, 8, 8, device="cuda")).prev_sample
    assert y.shape == (1, 4, 8, 8)
