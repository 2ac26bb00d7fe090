"""fp16 storage (SLIDER_F16) on the forward / inference entry points: the reference's inference scripts run the UNet and
the adapter in fp16 (eval-scripts/generate_images_xl.py:443, generate_images_sd1.py).  Kernels against fp32 torch on the
fp16-rounded operands; the inference sweep of the tiny UNet against the fp32 CPU oracle; training entry points reject
the dtype."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
H = torch.float16


@pytest.mark.parametrize("case", [(1, 1, 8192, 320, 320, 1, 4), (2, 32, 32, 640, 640, 3, 4), (2, 8, 8, 1280, 1280, 3, 8),
                                  (1, 1, 300, 64, 72, 1, 16), (1, 1, 2, 1280, 320, 1, 4), (1, 1, 512, 1280, 1280, 1, 4)])
def test_lora_forward_fp16(case):
    from sliders_conceptmod_b200 import ops
    B, Hh, W, C, N, k, r = case
    g = torch.Generator(device="cuda").manual_seed(sum(case))
    x = torch.randn(B, Hh, W, C, device="cuda", generator=g).to(H)
    w = ((torch.rand(N, k * k, C, device="cuda", generator=g) * 2 - 1) / (C * k * k) ** 0.5).to(H)
    bias = (torch.randn(N, device="cuda", generator=g) * 0.1).to(H)
    a = ((torch.rand(r, k * k, C, device="cuda", generator=g) * 2 - 1) / (C * k * k) ** 0.5).to(H)
    b = (torch.randn(N, r, device="cuda", generator=g) * 0.2).to(H)
    lin = k == 1 and B == 1 and Hh == 1
    xin = x.reshape(W, C) if lin else x
    for s, res in ((0.75, False), (0.0, False), (0.75, True)):
        if res and W * Hh * B <= 16:
            continue
        r0 = torch.randn(*xin.shape[:-1], N, device="cuda", generator=g).to(H) if res else None
        y, _ = ops.lora_forward(xin, w, bias, a, b, s, ksize=k, residual=None if r0 is None else r0.clone())
        xn = x.float().permute(0, 3, 1, 2)
        ref = F.conv2d(xn, w.float().reshape(N, k, k, C).permute(0, 3, 1, 2), bias.float(), padding=(k - 1) // 2)
        if s:
            hh = F.conv2d(xn, a.float().reshape(r, k, k, C).permute(0, 3, 1, 2), padding=(k - 1) // 2)
            ref = ref + s * F.conv2d(hh, b.float()[:, :, None, None])
        ref = ref.permute(0, 2, 3, 1).reshape(y.shape)
        if res:
            ref = ref + r0.float()
        err = (y.float() - ref).abs().max().item()
        assert y.dtype == H and err <= 4e-3 * ref.abs().max().item(), f"s={s} res={res}: {err:.3e} vs {ref.abs().max().item():.3e}"


def test_elementwise_forward_kernels_fp16():
    from sliders_conceptmod_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(3)
    x = (torch.randn(2, 32, 32, 320, device="cuda", generator=g) * 1.5 + 0.3).to(H)
    gamma = (1 + 0.2 * torch.randn(320, device="cuda", generator=g)).to(H)
    beta = (0.1 * torch.randn(320, device="cuda", generator=g)).to(H)
    add = torch.randn(2, 320, device="cuda", generator=g).to(H)
    y = ops.groupnorm_nhwc(x, 32, gamma, beta, 1e-5, True, add)
    xr = (x.float() + add.float()[:, None, None, :]).permute(0, 3, 1, 2)
    ref = F.silu(F.group_norm(xr, 32, gamma.float(), beta.float(), 1e-5)).permute(0, 2, 3, 1)
    assert (y.float() - ref).abs().max().item() <= 4e-3 * ref.abs().max().item()
    rows = x.reshape(-1, 320)
    ln = ops.layernorm(rows, gamma, beta, 1e-5)
    assert (ln.float() - F.layer_norm(rows.float(), (320,), gamma.float(), beta.float(), 1e-5)).abs().max().item() <= 5e-3
    hx = torch.randn(512, 2560, device="cuda", generator=g).to(H)
    a_, g_ = hx.float().chunk(2, dim=-1)
    assert (ops.geglu(hx).float() - a_ * F.gelu(g_)).abs().max().item() <= 4e-3 * (a_ * F.gelu(g_)).abs().max().item()
    wg = (torch.randn(2560, 320, device="cuda", generator=g) / 320 ** 0.5).to(H)
    bg = (0.1 * torch.randn(2560, device="cuda", generator=g)).to(H)
    idx = ops.geglu_permutation(1280, "cuda")
    fz = ops.linear_geglu(rows, wg.index_select(0, idx).contiguous(), bg.index_select(0, idx).contiguous())
    a2, g2 = F.linear(rows.float(), wg.float(), bg.float()).chunk(2, dim=-1)
    assert (fz.float() - a2 * F.gelu(g2)).abs().max().item() <= 4e-3 * (a2 * F.gelu(g2)).abs().max().item()
    up = ops.upsample_nearest2x_nhwc(x)
    assert torch.equal(up, F.interpolate(x.permute(0, 3, 1, 2), scale_factor=2.0, mode="nearest").permute(0, 2, 3, 1).contiguous())
    assert torch.equal(ops.concat_channels_nhwc(x, x), torch.cat([x, x], dim=-1))
    e0, e1, z = (torch.randn(1, 4, 64, 64, device="cuda", generator=g).to(H) for _ in range(3))
    out = ops.cfg_step(z, e0, e1, None, guidance=3.0, c_x=0.9, c_eps=-0.2)
    ref = 0.9 * z.float() - 0.2 * (e0.float() + 3.0 * (e1.float() - e0.float()))
    assert (out.float() - ref).abs().max().item() <= 2e-3 * ref.abs().max().item()


def test_training_entry_points_reject_fp16():
    from sliders_conceptmod_b200 import ops
    t = torch.randn(1024, device="cuda").to(H)
    with pytest.raises(RuntimeError):
        ops.guidance_loss(t, t, t, t, 1.0, 1)
    gy = torch.randn(256, 64, device="cuda").to(H)
    with pytest.raises(RuntimeError):
        ops.lora_grad(gy, gy, torch.zeros(256, 4, device="cuda"), torch.zeros(64, 4, device="cuda").to(H), 1.0, 1,
                      torch.zeros(4, 1, 64, device="cuda"), torch.zeros(64, 4, device="cuda"))


@pytest.mark.parametrize("scheduler,steps", [("ddim", 8), ("lms", 6)])
def test_inference_sweep_fp16_matches_oracle(scheduler, steps):
    """generate_images.denoise_with_slider with the UNet stand-in and the adapter in fp16 against the fp32 CPU oracle."""
    from helpers import build_pair
    from oracle import slider_oracle
    from sliders_conceptmod_b200 import generate_images, model_util
    from sliders_conceptmod_b200.unet import UNetConfig
    cfg = UNetConfig.tiny()
    unet_c, net_c, unet_g, net_g = build_pair(cfg, "cuda", H)
    g = torch.Generator().manual_seed(4)
    emb = torch.randn(2, 77, cfg.cross_attention_dim, generator=g).to(H)
    lat = torch.randn(1, 4, 16, 16, generator=g).to(H)
    sched = model_util.create_noise_scheduler(scheduler)
    osch = slider_oracle.OracleScheduler(scheduler)
    sched.set_timesteps(steps); osch.set_timesteps(steps)
    for scale in (2.0, -1.0):
        out_g = generate_images.denoise_with_slider(unet_g, net_g, sched, emb.cuda(), (lat * float(sched.init_noise_sigma)).cuda(),
                                                    scale, 500, 7.5, steps)
        out_c = slider_oracle.inference_sweep(unet_c, net_c, osch, emb.float(), lat.float() * float(osch.init_noise_sigma),
                                              scale, 500, 7.5, steps)
        err = (out_g.float().cpu() - out_c).abs().max().item()
        assert out_g.dtype == H and err <= 3e-2 * out_c.abs().max().item(), f"{scheduler} {scale}: {err:.3e} of {out_c.abs().max().item():.3e}"
