"""Generates tests/golden/*.pt by EXECUTING THE REFERENCE's own classes (read-only import from
/root/reference) on seeded inputs.  Run in the authoring container only; the vectors travel, the
reference does not.

    python tests/golden/make_golden.py

What is pinned:
  lora_linear / lora_conv : reference LoRAModule.forward + autograd   (trainscripts/imagesliders/lora.py:49-111,
                                                                      conceptmod/textsliders/lora.py:76-138)
  loss                    : reference PromptEmbedsPair.loss, both dialects, enhance/erase, + autograd
                            (trainscripts/imagesliders/prompt_util.py:71-148, conceptmod/textsliders/prompt_util.py:91-174)
  keys                    : reference LoRANetwork.create_modules key lists on the UNet stand-in
                            (SD-1.4 noxattn+c3lier, SD-XL attention-only, SD-XL c3lier, tiny; all train methods on tiny)
  ddim_prev_step          : reference prev_step (trainscripts/imagesliders/train_lora-scale.py:31-39)
  cfg                     : reference train_util.predict_noise CFG tail via a stub UNet (train_util.py:276-302)
"""
import importlib.util
import sys
import types
from pathlib import Path

import torch
import torch.nn as nn

REF = Path("/root/reference")
HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT))


def _stub_diffusers():
    class _Any(types.ModuleType):
        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return type(name, (), {})
    for name in ["diffusers", "diffusers.models", "diffusers.models.transformers", "diffusers.utils",
                 "diffusers.utils.torch_utils", "diffusers.pipelines", "diffusers.pipelines.flux",
                 "diffusers.pipelines.flux.pipeline_flux", "diffusers.image_processor", "diffusers.schedulers",
                 "diffusers.pipelines.stable_diffusion.convert_from_ckpt", "diffusers.pipelines.stable_diffusion"]:
        m = _Any(name)
        m.__path__ = []
        sys.modules.setdefault(name, m)


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def main():
    _stub_diffusers()
    up_lora = _load("ref_up_lora", REF / "trainscripts/imagesliders/lora.py")
    fk_lora = _load("ref_fk_lora", REF / "conceptmod/textsliders/lora.py")
    up_prompt = _load("ref_up_prompt", REF / "trainscripts/imagesliders/prompt_util.py")
    fk_prompt = _load("ref_fk_prompt", REF / "conceptmod/textsliders/prompt_util.py")
    out = {}

    # ---- LoRAModule forward/backward -----------------------------------------------------------------
    def lora_case(kind, org, x, mult, rank, alpha, seed):
        torch.manual_seed(seed)
        m = fk_lora.LoRAModule("g", org, multiplier=mult, lora_dim=rank, alpha=alpha)
        with torch.no_grad():
            m.lora_up.weight.normal_(0, 0.05)
        m.apply_to()
        x = x.clone().requires_grad_(True)
        y = org(x)
        w = torch.randn(y.shape, generator=torch.Generator().manual_seed(seed + 7))
        (y * w).sum().backward()
        return dict(kind=kind, x=x.detach(), W=org.weight.detach(), bias=None if org.bias is None else org.bias.detach(),
                    A=m.lora_down.weight.detach(), B=m.lora_up.weight.detach(), mult=mult, scale=m.scale,
                    y=y.detach(), gy=w, dx=x.grad, dA=m.lora_down.weight.grad, dB=m.lora_up.weight.grad)

    g = torch.Generator().manual_seed(1234)
    cases = []
    torch.manual_seed(1)
    cases.append(lora_case("linear", nn.Linear(64, 96), torch.randn(2, 40, 64, generator=g), 1.0, 4, 1.0, 10))
    cases.append(lora_case("linear", nn.Linear(128, 64, bias=False), torch.randn(3, 128, generator=g), -2.0, 8, 4.0, 11))
    cases.append(lora_case("conv3", nn.Conv2d(32, 64, 3, padding=1), torch.randn(2, 32, 12, 10, generator=g), 1.5, 4, 1.0, 12))
    cases.append(lora_case("conv1", nn.Conv2d(64, 32, 1), torch.randn(2, 64, 8, 8, generator=g), 1.0, 4, None, 13))
    cases.append(lora_case("conv3s2", nn.Conv2d(32, 32, 3, stride=2, padding=1), torch.randn(2, 32, 16, 16, generator=g), 1.0, 4, 1.0, 14))
    out["lora_cases"] = cases

    # ---- loss ------------------------------------------------------------------------------------------
    losses = []
    for dialect, mod in (("upstream", up_prompt), ("fork", fk_prompt)):
        for action in ("enhance", "erase"):
            for gs in (1.0, 4.0):
                for dt in (torch.float32, torch.bfloat16):
                    t, p, n, q = (torch.randn(1, 4, 16, 16, generator=g).to(dt) for _ in range(4))
                    st = mod.PromptSettings(target="a", positive="b", unconditional="c", neutral="d", action=action,
                                            guidance_scale=gs)
                    if dialect == "upstream":
                        pair = mod.PromptEmbedsPair(torch.nn.MSELoss(), None, None, None, None, st)
                        kw = dict(unconditional_latents=q)
                    else:
                        pair = mod.PromptEmbedsPair(torch.nn.MSELoss(), None, None, None, None, None, st)
                        kw = dict(negative_latents=q)
                    tt = t.clone().requires_grad_(True)
                    loss = pair.loss(target_latents=tt, positive_latents=p, neutral_latents=n, **kw)
                    loss.backward()
                    losses.append(dict(dialect=dialect, action=action, guidance=gs, t=t, p=p, n=n, q=q,
                                       loss=loss.detach(), grad=tt.grad))
    out["loss_cases"] = losses

    # ---- key lists ---------------------------------------------------------------------------------------
    from sliders_conceptmod_b200.unet import UNet2DConditionModel, UNetConfig
    keys = {}
    conv = list(fk_lora.UNET_TARGET_REPLACE_MODULE_CONV)
    attn = list(fk_lora.UNET_TARGET_REPLACE_MODULE_TRANSFORMER)

    def ref_keys(cfg, target, method, rank=4):
        with torch.device("meta"):
            u = UNet2DConditionModel(cfg)
        net = fk_lora.LoRANetwork(u, rank=rank, alpha=1.0, target_replace=target, train_method=method)
        sd = net.state_dict()
        return {k: list(v.shape) for k, v in sd.items()}

    keys["sd14_noxattn_c3lier"] = ref_keys(UNetConfig.sd14(), attn + conv, "noxattn")
    keys["sd14_noxattn_lierla"] = ref_keys(UNetConfig.sd14(), attn, "noxattn")
    keys["sdxl_noxattn_attn"] = ref_keys(UNetConfig.sdxl(), attn, "noxattn")
    keys["sdxl_noxattn_c3lier"] = ref_keys(UNetConfig.sdxl(), attn + conv, "noxattn")
    for method in ("noxattn", "innoxattn", "selfattn", "xattn", "full", "xattn-strict", "noxattn-hspace",
                   "noxattn-hspace-last"):
        keys["tiny_" + method] = ref_keys(UNetConfig.tiny(), attn + conv, method)
    out["keys"] = keys

    # ---- DDIM prev_step (reference's own restatement) ----------------------------------------------------------
    src = (REF / "trainscripts/imagesliders/train_lora-scale.py").read_text().splitlines()
    start = next(i for i, l in enumerate(src) if l.startswith("def prev_step"))
    end = next(i for i in range(start + 1, len(src)) if src[i].startswith("def "))
    ns = {}
    exec("\n".join(src[start:end]), ns)       # executes the reference function text; nothing is copied into the repo
    from oracle.slider_oracle import alphas_cumprod
    ddim = []
    for n_inf in (50, 1000):
        for t in (980, 500, 20, 0):
            sched = types.SimpleNamespace(config=types.SimpleNamespace(num_train_timesteps=1000), num_inference_steps=n_inf,
                                          alphas_cumprod=alphas_cumprod(), final_alpha_cumprod=torch.tensor(1.0))
            eps, x = torch.randn(1, 4, 8, 8, generator=g), torch.randn(1, 4, 8, 8, generator=g)
            ddim.append(dict(n_inf=n_inf, t=t, eps=eps, x=x, prev=ns["prev_step"](eps, t, sched, x)))
    out["ddim_cases"] = ddim

    # ---- CFG tail of predict_noise (reference train_util) ----------------------------------------------------------
    try:
        sys.path.insert(0, str(REF / "trainscripts/imagesliders"))
        tu = _load("ref_up_train_util", REF / "trainscripts/imagesliders/train_util.py")
        cfgs = []
        for gs in (1.0, 3.0, 7.5):
            eps2 = torch.randn(2, 4, 8, 8, generator=g)
            unet = lambda x, t, encoder_hidden_states=None: types.SimpleNamespace(sample=eps2)
            sched = types.SimpleNamespace(scale_model_input=lambda x, t: x)
            guided = tu.predict_noise(unet, sched, 10, torch.zeros(1, 4, 8, 8), None, guidance_scale=gs)
            cfgs.append(dict(guidance=gs, eps2=eps2, guided=guided))
        out["cfg_cases"] = cfgs
    except Exception as e:  # pragma: no cover
        print("train_util import failed, cfg cases skipped:", repr(e))

    torch.save(out, HERE / "reference_vectors.pt")
    print("wrote", HERE / "reference_vectors.pt", {k: len(v) for k, v in out.items()})


if __name__ == "__main__":
    main()
