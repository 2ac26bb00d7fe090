"""Size-independent properties of the oracle's arithmetic (CPU, fp64 so that a property holds to rounding, not to a
tolerance tuned per case).  The GPU suite checks the CUDA path against the oracle; these check that the oracle itself
has the algebra the reference's formulas imply, over random shapes (hypothesis)."""
import torch
import torch.nn as nn
from hypothesis import given, settings, strategies as st

from oracle import slider_oracle as O

settings.register_profile("oracle", max_examples=25, deadline=None, derandomize=True, database=None)
settings.load_profile("oracle")
F64 = torch.float64


def _lin(cin, cout, r, seed):
    g = torch.Generator().manual_seed(seed)
    org = nn.Linear(cin, cout).to(F64)
    down, up = nn.Linear(cin, r, bias=False).to(F64), nn.Linear(r, cout, bias=False).to(F64)
    with torch.no_grad():
        for p in (*org.parameters(), *down.parameters(), *up.parameters()):
            p.copy_(torch.randn(p.shape, generator=g, dtype=F64))
    return org, down, up


@given(st.integers(1, 40), st.integers(1, 24), st.integers(1, 24), st.integers(1, 8), st.floats(-4, 4), st.integers(0, 999))
def test_lora_update_is_linear_in_the_slider_scale(m, cin, cout, r, s, seed):
    """lora.py:134-138: y(s) - y(0) = s * (y(1) - y(0)); multiplier 0 is the frozen layer (what `with network:` exits to)."""
    org, down, up = _lin(cin, cout, r, seed)
    x = torch.randn(m, cin, generator=torch.Generator().manual_seed(seed + 1), dtype=F64)
    y = lambda mult: O.lora_forward(org.forward, down, up, mult, 0.25, x)
    assert torch.equal(y(0.0), org(x))
    torch.testing.assert_close(y(s) - y(0.0), s * (y(1.0) - y(0.0)), rtol=1e-10, atol=1e-10)


@given(st.integers(1, 4), st.integers(1, 6), st.floats(0.5, 8), st.integers(0, 999))
def test_guidance_loss_erase_is_enhance_with_the_opposite_sign(b, hw, g, seed):
    """prompt_util.py:134-174: erase(g) == enhance(-g); the loss is zero exactly at the guided target and its gradient
    w.r.t. the target latents is 2 (t - tgt) / n."""
    gen = torch.Generator().manual_seed(seed)
    t, p, n, q = (torch.randn(b, 4, hw, hw, generator=gen, dtype=F64) for _ in range(4))
    assert torch.equal(O.guidance_loss(t, p, n, q, g, "erase"), O.guidance_loss(t, p, n, q, -g, "enhance"))
    tgt = n + g * (p - q)
    assert O.guidance_loss(tgt, p, n, q, g, "enhance").item() == 0.0
    tr = t.clone().requires_grad_(True)
    O.guidance_loss(tr, p, n, q, g, "enhance").backward()
    torch.testing.assert_close(tr.grad, 2 * (t - tgt) / t.numel(), rtol=1e-12, atol=1e-14)


@given(st.integers(1, 3), st.integers(1, 6), st.floats(0, 12), st.integers(0, 999))
def test_cfg_combine_fixed_points(b, hw, g, seed):
    """train_util.py:297-300: guidance 1 returns the conditional half, guidance 0 the unconditional half, and equal
    halves are a fixed point for every guidance scale."""
    gen = torch.Generator().manual_seed(seed)
    u, c = torch.randn(b, 4, hw, hw, generator=gen, dtype=F64), torch.randn(b, 4, hw, hw, generator=gen, dtype=F64)
    torch.testing.assert_close(O.cfg_combine(torch.cat([u, c]), 1.0), c, rtol=1e-14, atol=1e-14)
    assert torch.equal(O.cfg_combine(torch.cat([u, c]), 0.0), u)
    assert torch.equal(O.cfg_combine(torch.cat([u, u]), g), u)


@given(st.integers(1, 12), st.integers(1, 10), st.integers(1, 4), st.integers(0, 999))
def test_dora_with_zero_update_and_unit_column_scale_is_the_base_layer(cin, cout, r, seed):
    """dora.py:122-163: with lora_up = 0 and dora_scale = ||W|| per input channel the decomposed weight is W itself,
    so the module adds nothing whatever the multiplier — the state the reference initialises to."""
    gen = torch.Generator().manual_seed(seed)
    w = torch.randn(cout, cin, generator=gen, dtype=F64)
    x = torch.randn(5, cin, generator=gen, dtype=F64)
    a = torch.randn(r, cin, generator=gen, dtype=F64)
    b = torch.zeros(cout, r, dtype=F64)
    m = w.norm(dim=0, keepdim=True)
    y = O.dora_forward(x, w, None, a, b, m, 3.0, 0.5)
    torch.testing.assert_close(y, x @ w.t(), rtol=1e-12, atol=1e-12)


@given(st.integers(2, 50), st.integers(0, 999))
def test_ddim_step_reconstructs_the_clean_sample(steps, seed):
    """train_lora-scale.py:31-39: when the model output is the TRUE noise of x_t = sqrt(a) x0 + sqrt(1-a) eps, one
    eta = 0 step lands on sqrt(a_prev) x0 + sqrt(1-a_prev) eps: every step of the loop stays on the same (x0, eps) ray."""
    sch = O.OracleScheduler("ddim")
    sch.set_timesteps(steps)
    gen = torch.Generator().manual_seed(seed)
    x0, eps = torch.randn(1, 4, 4, 4, generator=gen, dtype=F64), torch.randn(1, 4, 4, 4, generator=gen, dtype=F64)
    ac = sch.alphas_cumprod.to(F64)
    for t in sch.timesteps[: min(len(sch.timesteps), 6)]:
        t = int(t)
        xt = ac[t].sqrt() * x0 + (1 - ac[t]).sqrt() * eps
        prev = t - sch.config.num_train_timesteps // sch.num_inference_steps
        a_prev = ac[prev] if prev >= 0 else torch.as_tensor(sch.final_alpha_cumprod, dtype=F64)
        want = a_prev.sqrt() * x0 + (1 - a_prev).sqrt() * eps
        got = O.ddim_prev_step(eps, t, sch, xt)
        torch.testing.assert_close(got.to(F64), want, rtol=1e-5, atol=1e-5)
