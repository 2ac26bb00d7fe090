"""CPU tests of the host-side mirror of the reference API (no kernels are launched)."""
import math
import re
from pathlib import Path

import pytest
import torch
import torch.nn as nn

from helpers import ROOT
from oracle import slider_oracle as O
from sliders_conceptmod_b200 import config_util, lora, model_util, prompt_util, schedulers, train_util, workload
from sliders_conceptmod_b200.unet import UNet2DConditionModel, UNetConfig

GOLD = torch.load(Path(__file__).parent / "golden" / "reference_vectors.pt", weights_only=False)
ATTN = list(lora.UNET_TARGET_REPLACE_MODULE_TRANSFORMER)
CONV = list(lora.UNET_TARGET_REPLACE_MODULE_CONV)


def _keys(cfg, target, method):
    with torch.device("meta"):
        u = UNet2DConditionModel(cfg)
    net = lora.LoRANetwork(u, rank=4, alpha=1.0, target_replace=target, train_method=method)
    return {k: list(v.shape) for k, v in net.state_dict().items()}, net


@pytest.mark.parametrize("name,cfg,target,method,n_mod", [
    ("sd14_noxattn_c3lier", UNetConfig.sd14(), ATTN + CONV, "noxattn", 150),
    ("sd14_noxattn_lierla", UNetConfig.sd14(), ATTN, "noxattn", 64),
    ("sdxl_noxattn_attn", UNetConfig.sdxl(), ATTN, "noxattn", 280),
    ("sdxl_noxattn_c3lier", UNetConfig.sdxl(), ATTN + CONV, "noxattn", 346),
])
def test_key_layout_matches_reference(name, cfg, target, method, n_mod):
    keys, net = _keys(cfg, target, method)
    assert keys == GOLD["keys"][name]          # same names, same shapes, same order as the reference class
    assert len(net.unet_loras) == n_mod


@pytest.mark.parametrize("method", ["noxattn", "innoxattn", "selfattn", "xattn", "full", "xattn-strict",
                                    "noxattn-hspace", "noxattn-hspace-last"])
def test_train_method_filters(method):
    keys, _ = _keys(UNetConfig.tiny(), ATTN + CONV, method)
    assert keys == GOLD["keys"]["tiny_" + method]


def test_unknown_train_method_raises():
    with pytest.raises(NotImplementedError):
        _keys(UNetConfig.tiny(), ATTN, "outxattn")


def test_param_counts_match_survey():
    _, net = _keys(UNetConfig.sd14(), ATTN + CONV, "noxattn")
    assert sum(p.numel() for p in net.parameters()) == 2_906_880
    _, net = _keys(UNetConfig.sdxl(), ATTN, "noxattn")
    assert sum(p.numel() for p in net.parameters()) == 2_662_400


class Attention(nn.Module):          # the class NAME is what create_modules matches (lora.py:225)
    def __init__(self, child_name, child):
        super().__init__()
        self.add_module(child_name, child)


class Root(nn.Module):
    def __init__(self, holder):
        super().__init__()
        self.attn1 = holder


def test_multiplier_context_and_scale():
    lin = nn.Linear(16, 24)
    root = Root(Attention("to_q", lin))
    net = lora.LoRANetwork(root, rank=4, multiplier=1.0, alpha=1.0, train_method="full")
    m = net.unet_loras[0]
    assert m.lora_name == "lora_unet_attn1_to_q"
    assert m.multiplier == 1.0                      # before the first `with` (lora.py:126)
    assert m.scale == 0.25 and float(m.alpha) == 1.0
    net.set_lora_slider(3)
    with net:
        assert m.multiplier == 3.0
    assert m.multiplier == 0                         # lora.py:299-301
    assert torch.count_nonzero(m.lora_up.weight) == 0           # zero-init up (lora.py:124)
    assert lin.forward.__self__ is m                 # forward patched, module identity kept (lora.py:129-132)
    assert "org_module" not in dict(net.named_modules()) and not any("org" in k for k in net.state_dict())
    # upstream positional signature (unet, rank, multiplier, alpha, train_method)
    root2 = Root(Attention("to_k", nn.Linear(8, 8)))
    net2 = lora.LoRANetwork(root2, 4, 1.0, 2.0, "full")
    assert net2.alpha == 2.0 and net2.unet_loras[0].scale == 0.5
    with pytest.raises(RuntimeError):
        lin(torch.randn(2, 16))                      # no CPU arithmetic path


def test_alpha_none_and_conv_rank_clamp():
    conv = nn.Conv2d(2, 8, 3, padding=1)
    m = lora.LoRAModule("c", conv, 1.0, 4, None)
    assert m.lora_dim == 2 and m.lora_down.weight.shape == (2, 2, 3, 3) and m.lora_up.weight.shape == (8, 2, 1, 1)
    assert m.scale == 4 / 2                          # alpha None -> lora_dim argument (lora.py:118-119)
    with pytest.raises(NotImplementedError):
        lora.LoRAModule("c5", nn.Conv2d(8, 8, 5, padding=2), 1.0, 4, 1.0)


def test_save_weights_formats(tmp_path):
    _, net = _keys(UNetConfig.tiny(), ATTN + CONV, "noxattn")
    net = net.to_empty(device="cpu")
    for p in net.parameters():
        nn.init.normal_(p)
    net.save_weights(str(tmp_path / "a.pt"), dtype=torch.bfloat16)
    net.save_weights(str(tmp_path / "a.safetensors"), dtype=torch.float32, metadata={"k": "v"})
    sd = torch.load(tmp_path / "a.pt")
    from safetensors.torch import load_file
    sf = load_file(str(tmp_path / "a.safetensors"))
    assert list(sd) == list(net.state_dict()) and set(sf) == set(sd)
    assert all(v.dtype == torch.bfloat16 for v in sd.values()) and all(v.dtype == torch.float32 for v in sf.values())
    k = next(k for k in sd if k.endswith("lora_down.weight"))
    assert torch.equal(sf[k], net.state_dict()[k].float())


def test_prompt_settings_and_yaml(tmp_path):
    ps = prompt_util.PromptSettings(target="x")
    assert (ps.positive, ps.unconditional, ps.neutral, ps.action, ps.guidance_scale) == ("x", "", "", "erase", 1.0)
    with pytest.raises(Exception):
        prompt_util.PromptSettings(positive="y")
    f = tmp_path / "p.yaml"
    f.write_text("- target: 'person'\n  positive: 'old person'\n  unconditional: 'young person'\n  neutral: 'person'\n"
                 "  action: 'enhance'\n  guidance_scale: 4\n")
    out = prompt_util.load_prompts_from_yaml(str(f), ["male", "female"])
    assert [p.target for p in out] == ["male person", "female person"] and out[1].unconditional == "female young person"
    assert len(prompt_util.load_prompts_from_yaml(str(f))) == 1
    (tmp_path / "e.yaml").write_text("[]")
    with pytest.raises(ValueError):
        prompt_util.load_prompts_from_yaml(str(tmp_path / "e.yaml"))
    with pytest.raises(NotImplementedError):
        prompt_util.PromptEmbedsPair(torch.nn.L1Loss(), None, None, None, None, ps)
    pair = prompt_util.PromptEmbedsPair(torch.nn.MSELoss(), 1, 2, 3, 4, 5, ps)     # fork form (7 args)
    assert pair.negative == 5 and pair.action == "erase"
    with pytest.raises(TypeError):
        pair.loss(target_latents=None, positive_latents=None, neutral_latents=None)


def test_config_schema(tmp_path):
    f = tmp_path / "c.yaml"
    f.write_text("prompts_file: 'p.yaml'\npretrained_model:\n  name_or_path: 'CompVis/stable-diffusion-v1-4'\n"
                 "network:\n  type: 'c3lier'\n  rank: 4\n  alpha: 1.0\n  training_method: 'noxattn'\n")
    c = config_util.load_config_from_yaml(str(f))
    assert c.train.precision == "bfloat16" and c.train.max_denoising_steps == 50 and c.save.per_steps == 200
    assert config_util.parse_precision("bf16") == torch.bfloat16
    with pytest.raises(ValueError):
        config_util.parse_precision("int8")
    assert "c3lier" in c.json()


@pytest.mark.parametrize("kind", ["ddim", "ddpm", "euler_a"])
@pytest.mark.parametrize("pred", ["epsilon", "v_prediction"])
def test_scheduler_coefficients_match_oracle_steps(kind, pred):
    s = model_util.create_noise_scheduler(kind, pred)
    o = O.OracleScheduler(kind, pred)
    n = 12 if kind == "euler_a" else 50
    s.set_timesteps(n); o.set_timesteps(n)
    assert torch.equal(s.timesteps.float(), o.timesteps.float())
    assert abs(float(s.init_noise_sigma) - float(o.init_noise_sigma)) < 1e-6
    g = torch.Generator().manual_seed(0)
    x, e, z = (torch.randn(1, 4, 8, 8, generator=g, dtype=torch.float64) for _ in range(3))
    for t in (s.timesteps[0], s.timesteps[n // 2], s.timesteps[-1]):
        cx, ce, cn = s.step_coefficients(t)
        ours = cx * x + ce * e + cn * z
        ref = o.step(e, t, x, noise=z).prev_sample
        assert torch.allclose(ours, ref.double(), rtol=1e-5, atol=1e-5), (kind, pred, float(t))
        xin = o.scale_model_input(x, t)
        assert torch.allclose(xin, x * s.input_scale(t), rtol=1e-6, atol=1e-6)


def test_lms_scheduler_tables_and_coefficients():
    """LMS (model_util.py:415-423): sigma table / timesteps / input scale equal the oracle restatement, order-1
    coefficient is the Euler step sigma[t+1] - sigma[t], the order-k coefficients sum to it (Lagrange bases sum to 1),
    and they match the oracle's quadrature."""
    s = model_util.create_noise_scheduler("lms")
    o = O.OracleScheduler("lms")
    for n in (30, 50):
        s.set_timesteps(n); o.set_timesteps(n)
        assert torch.allclose(s.timesteps, o.timesteps) and torch.allclose(s.sigmas, o.sigmas)
        assert abs(float(s.init_noise_sigma) - float(o.init_noise_sigma)) < 1e-6
        sig = s.sigmas.double()
        for t in (0, 1, 5, n - 1):
            assert abs(s.lms_coefficient(1, t, 0) - float(sig[t + 1] - sig[t])) < 1e-4 * abs(float(sig[t]))
            order = min(t + 1, 4)
            cs = [s.lms_coefficient(order, t, c) for c in range(order)]
            assert abs(sum(cs) - float(sig[t + 1] - sig[t])) < 1e-3 * abs(float(sig[t]))
            for c in range(order):
                assert abs(cs[c] - O.lms_coefficient(o, order, t, c)) < 1e-6 * max(1.0, abs(cs[c]))
        x = torch.randn(1, 4, 8, 8, dtype=torch.float64)
        t = s.timesteps[3]
        assert torch.allclose(o.scale_model_input(x, t), x * s.input_scale(t), rtol=1e-6, atol=1e-6)
    with pytest.raises(NotImplementedError):
        s.step_coefficients(s.timesteps[0])         # multistep: no single affine rule (CUDA-graph engine: ddim/ddpm/euler_a)
    with pytest.raises(ValueError):
        model_util.create_noise_scheduler("nope")


def test_train_util_host_helpers():
    s = model_util.create_noise_scheduler("ddim")
    lat = train_util.get_initial_latents(s, 2, 512, 256, 1, generator=torch.Generator().manual_seed(3))
    assert lat.shape == (2, 4, 64, 32)
    emb = train_util.concat_embeddings(torch.zeros(1, 77, 8), torch.ones(1, 77, 8), 2)
    assert emb.shape == (4, 77, 8) and emb[:2].sum() == 0 and emb[2:].min() == 1
    ids = train_util.get_add_time_ids(1024, 1024)
    assert ids.tolist() == [[1024, 1024, 0, 0, 1024, 1024]]
    assert train_util.get_optimizer("AdamW") is torch.optim.AdamW
    with pytest.raises(ValueError):
        train_util.get_optimizer("sgd")
    p = nn.Parameter(torch.zeros(1)); opt = torch.optim.AdamW([p], lr=1.0)
    assert isinstance(train_util.get_lr_scheduler("cosine", opt, 100, 0.01), torch.optim.lr_scheduler.CosineAnnealingLR)
    with pytest.raises(ValueError):
        train_util.get_lr_scheduler("bogus", opt, 100, 0.01)


def test_workload_definitions():
    wl = workload.WORKLOADS["sd14_text_slider_rank4_512px"]
    pairs = workload.prompt_pairs(wl, 768)
    assert len(pairs) == 4 and pairs[0]["target"].shape == (1, 77, 768)
    assert torch.equal(pairs[0]["target"], workload.prompt_pairs(wl, 768)[0]["target"])     # deterministic


def test_groupnorm_single_launch_plan_table():
    """Host-side plan of the single-launch GroupNorm (csrc/groupnorm.cu gn_group_plan; no device needed): groups per
    slab (rows must be whole 16-byte vectors), cluster size (<= 24 KB of the slab per CTA, at most 8 CTAs), block size
    (a multiple of the 32-bit words per row, so a thread keeps one column), and the shapes that keep the two-kernel form."""
    import ctypes
    from sliders_conceptmod_b200 import _lib
    fn = _lib.lib().slider_debug_gn_group_plan
    fn.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    fn.restype = ctypes.c_int

    def plan(hw, c, g, dtype=_lib.SLIDER_BF16):
        out = (ctypes.c_int * 6)()
        rc = fn(hw, c, g, dtype, out)
        return rc, dict(zip(("k", "gp", "wr", "rows", "threads", "smem"), out))

    # SD-1.x / SD-XL shapes (32 groups, bf16): hw, channels -> (groups per slab, cluster size)
    for hw, c, gp, k in [(4096, 320, 4, 8), (4096, 640, 2, 8), (1024, 640, 2, 4), (1024, 1920, 2, 8),
                         (256, 1280, 1, 1), (256, 2560, 1, 2), (64, 1280, 1, 1), (64, 2560, 1, 1)]:
        rc, p = plan(hw, c, 32)
        assert rc == 1 and (p["gp"], p["k"]) == (gp, k), (hw, c, p)
        seg = p["gp"] * (c // 32) * 2
        assert seg % 16 == 0 and p["wr"] == seg // 4 and p["rows"] * p["k"] == hw
        assert p["threads"] % p["wr"] == 0 and p["threads"] % 32 == 0 and 256 < p["threads"] <= 512
        assert p["rows"] * seg <= p["smem"] <= p["rows"] * seg + 2048
    assert plan(4096, 320, 32)[1]["threads"] == 480
    # more than 100 KB of slab per CTA at k = 8 (one CTA per SM: clusters start in waves) keeps the two-kernel form;
    # a 10-byte group is not a whole number of 32-bit words; odd pixel counts fall back to smaller clusters
    assert plan(16384, 320, 32)[0] == 0 and plan(4096, 960, 32)[0] == 0 and plan(16384, 640, 32)[0] == 0
    rc, p = plan(4096, 640, 32, _lib.SLIDER_F32)
    assert rc == 1 and p["k"] == 8 and p["rows"] * 160 <= 100 * 1024
    assert plan(16, 40, 8)[0] == 0
    rc, p = plan(35, 64, 8)
    assert rc == 1 and p["k"] == 1 and p["rows"] == 35
    rc, p = plan(4096, 320, 32, _lib.SLIDER_F32)
    assert rc == 1 and p["gp"] == 2 and p["k"] == 8
    assert fn(0, 320, 32, _lib.SLIDER_BF16, (ctypes.c_int * 6)()) == -1


def test_combine_loras_matches_reference_merge(tmp_path):
    """f4 edge: sliders_conceptmod_b200/combine_loras.py against the reference's own merge()
    (tests/golden/make_golden_combine.py ran conceptmod/notrigger/combine_loras.py:26-62), plus the file round trip."""
    from pathlib import Path
    from safetensors.torch import load_file, save_file
    from sliders_conceptmod_b200 import combine_loras
    gold = torch.load(Path(__file__).parent / "golden" / "reference_combine.pt", weights_only=False)
    out = {}
    for fam, sd in gold["sources"].items():
        combine_loras.merge(out, (combine_loras.BinDataHandler(sd), set(sd.keys())), gold["strengths"][fam], fam)
    assert set(out) == set(gold["merged"])
    for k, v in gold["merged"].items():
        assert torch.equal(out[k], v), k
    assert combine_loras.rename_key("lora_unet-a-b.lora_down.weight", "transformer") == "transformer.a.b.lora_A.weight"
    # CLI path: three files in, one file out
    paths = []
    for fam, sd in gold["sources"].items():
        p = tmp_path / f"{fam}.safetensors"
        save_file({k: v.contiguous() for k, v in sd.items()}, str(p))
        paths.append(str(p))
    s = gold["strengths"]
    combine_loras.main(paths + [str(tmp_path / "merged.safetensors"), str(s["transformer"]), str(s["CLIP"]), str(s["T5"])])
    merged = load_file(str(tmp_path / "merged.safetensors"))
    assert set(merged) == set(gold["merged"])
    for k, v in gold["merged"].items():
        assert torch.allclose(merged[k], v), k
    import pytest
    with pytest.raises(AssertionError):
        combine_loras.merge({}, (combine_loras.BinDataHandler({"weird": torch.zeros(1)}), {"weird"}), 1.0, "CLIP")


def test_attention_view_rule_and_cpu_refusal():
    """ops._rows_ld / attention_usable decide which q / k / v views the short-key attention kernel takes (dense tensors
    and last-dim slices of ONE fused projection); nothing on the CPU is ever 'usable' and the op itself refuses CPU
    tensors instead of falling back."""
    from sliders_conceptmod_b200 import ops
    qkv = torch.zeros(2, 64, 3 * 320, dtype=torch.bfloat16)
    q, k, v = qkv.split(320, dim=-1)
    assert ops._rows_ld(q) == 960 and ops._rows_ld(k) == 960            # slices of the fused projection: row stride kept
    assert ops._rows_ld(torch.zeros(2, 64, 320)) == 320
    assert ops._rows_ld(torch.zeros(64, 320)) is None                   # not [B, len, C]
    assert ops._rows_ld(torch.zeros(2, 64, 640)[:, :, ::2]) is None     # rows not dense
    assert ops._rows_ld(torch.zeros(2, 128, 320)[:, ::2]) == 640        # every other row: still ONE row stride over the batch
    assert ops._rows_ld(torch.zeros(2, 100, 320)[:, :64]) is None       # batch items do not follow each other
    assert ops._rows_ld(torch.zeros(2, 64, 324)[:, :, :320]) is None    # row stride not a multiple of 16 bytes (bf16)
    assert not ops.attention_usable(q, k, v, heads=8)                   # CPU tensors
    with pytest.raises(Exception):
        ops.attention(q.contiguous(), k.contiguous(), v.contiguous(), 8)


def test_workspace_lanes_nest_and_restore():
    from sliders_conceptmod_b200 import ops
    assert ops.current_lane() == 0
    with ops.workspace_lane(1):
        assert ops.current_lane() == 1
        with ops.workspace_lane(3):                                     # the shortcut branch of a block on lane 1
            assert ops.current_lane() == 3
            assert ops._ws_key(torch.device("cuda", 0)) == (0, 3)
        assert ops.current_lane() == 1
    assert ops.current_lane() == 0
    try:
        with ops.workspace_lane(2):
            raise RuntimeError("x")
    except RuntimeError:
        pass
    assert ops.current_lane() == 0                                      # restored on the error path too


@pytest.mark.parametrize("name,params,tensors", [("sd14", 859_520_964, 686), ("sdxl", 2_567_463_684, 1680)])
def test_unet_stand_in_has_the_published_size(name, params, tensors):
    """The random-init stand-in is the architecture the reference loads (`UNet2DConditionModel.from_pretrained`,
    model_util.py:75-85, 142-186): parameter and tensor counts of the published SD-1.x / SD-XL UNet checkpoints, so the
    benched FLOPs, the wrapped-layer shapes and the state-dict layout are the real workload's."""
    with torch.device("meta"):
        u = UNet2DConditionModel(getattr(UNetConfig, name)())
    assert sum(p.numel() for p in u.parameters()) == params
    assert len(u.state_dict()) == tensors


def test_oracle_is_test_infrastructure_only():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/: no module of the product package
    (nor the C sources) imports, loads or names a file under it, and nothing at run time reads /root/reference."""
    import ast
    pkg = ROOT / "sliders_conceptmod_b200"
    for path in sorted(pkg.glob("*.py")):
        tree = ast.parse(path.read_text())
        for node in ast.walk(tree):
            mods = []
            if isinstance(node, ast.Import):
                mods = [a.name for a in node.names]
            elif isinstance(node, ast.ImportFrom):
                mods = [node.module or ""]
            assert not any(m == "oracle" or m.startswith("oracle.") for m in mods), f"{path.name} imports the oracle"
        docstrings = {id(n.body[0].value) for n in ast.walk(tree)
                      if isinstance(n, (ast.Module, ast.FunctionDef, ast.ClassDef)) and n.body
                      and isinstance(n.body[0], ast.Expr) and isinstance(n.body[0].value, ast.Constant)}
        for node in ast.walk(tree):       # citations in docstrings are fine; a path used as data is not
            if isinstance(node, ast.Constant) and isinstance(node.value, str) and id(node) not in docstrings:
                assert "/root/reference" not in node.value, f"{path.name} names the reference checkout"
    for path in sorted((pkg / "csrc").glob("*")):
        assert "oracle" not in path.read_text().lower(), f"{path.name} mentions the oracle"
    entry = ast.parse((ROOT / "__graft_entry__.py").read_text())
    users = {f.name for f in ast.walk(entry) if isinstance(f, ast.FunctionDef)
             and any(isinstance(n, (ast.Import, ast.ImportFrom)) and "oracle" in ast.dump(n) for n in ast.walk(f))}
    assert users <= {"smoke"}, users
    for name in ("bench.py", "__graft_entry__.py"):
        assert "/root/reference" not in (ROOT / name).read_text(), f"{name} reads the reference at run time"
