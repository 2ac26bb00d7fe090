"""Generates tests/golden/reference_dora.pt by EXECUTING THE REFERENCE's DoRAModule (read-only import from
/root/reference/conceptmod/textsliders/dora.py:52-163) on seeded inputs.  Run in the authoring container only.

    python tests/golden/make_golden_dora.py

DoRA is a "next" row; the vectors pin the oracle restatement (oracle/slider_oracle.py dora_forward) and
the per-input-channel fold that maps DoRA onto the fused base + low-rank kernel (sliders_conceptmod_b200/dora.py).
"""
import sys
from pathlib import Path

import torch
import torch.nn as nn

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
from make_golden import REF, _load, _stub_diffusers  # noqa: E402


def main():
    _stub_diffusers()
    ref = _load("ref_fk_dora", REF / "conceptmod/textsliders/dora.py")
    g = torch.Generator().manual_seed(4321)
    cases = []

    def case(kind, org, x, mult, rank, alpha, seed):
        torch.manual_seed(seed)
        m = ref.DoRAModule("g", org, multiplier=mult, lora_dim=rank, alpha=alpha)
        with torch.no_grad():
            m.lora_up.weight.normal_(0, 0.05)
            m.dora_scale.mul_(1 + 0.1 * torch.randn(m.dora_scale.shape, generator=torch.Generator().manual_seed(seed + 3)))
        m.apply_to()
        with torch.no_grad():
            y = org(x)          # the patched forward
        cases.append(dict(kind=kind, x=x, W=m.org_module_weight.clone(), bias=None if org.bias is None else org.bias.detach().clone(),
                          A=m.lora_down.weight.detach().clone(), B=m.lora_up.weight.detach().clone(),
                          dora_scale=m.dora_scale.detach().clone(), mult=mult, scale=float(m.scale),
                          stride=getattr(m, "stride", None), padding=getattr(m, "padding", None), y=y))

    torch.manual_seed(2)
    case("linear", nn.Linear(64, 96), torch.randn(2, 40, 64, generator=g), 1.0, 4, 1.0, 20)
    case("linear", nn.Linear(128, 64, bias=False), torch.randn(3, 128, generator=g), -2.0, 8, 4.0, 21)
    case("conv3", nn.Conv2d(32, 64, 3, padding=1), torch.randn(2, 32, 12, 10, generator=g), 1.5, 4, 1.0, 22)
    case("conv1", nn.Conv2d(64, 32, 1), torch.randn(2, 64, 8, 8, generator=g), 0.5, 4, 2.0, 23)
    torch.save({"dora_cases": cases}, HERE / "reference_dora.pt")
    print("wrote", HERE / "reference_dora.pt", len(cases), "cases")


if __name__ == "__main__":
    main()
