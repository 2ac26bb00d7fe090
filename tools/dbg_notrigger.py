import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import config_util, train_notrigger, lora
from sliders_conceptmod_b200.text_encoder import load_text_encoders
from oracle import slider_oracle

gold = torch.load(Path(__file__).resolve().parent.parent / "tests/golden/reference_notrigger.pt")
config = config_util.RootConfig(**gold["config"])


def enc():
    return [load_text_encoders("SDXL", seed=7, tiny=True), load_text_encoders("SDXL", seed=7, tiny=True)]


for case in ("positive_only_te2", "both_te2"):
    c = gold["cases"][case]
    torch.manual_seed(c["torch_seed"])
    tc = train_notrigger.NoTriggerTrainer(config, "cpu", clip_index=1, peft_type="lora", rank=4, encoders=enc(), **c["kw"])
    slider_oracle.attach(tc.network)
    torch.manual_seed(c["torch_seed"])
    tg = train_notrigger.NoTriggerTrainer(config, "cuda", clip_index=1, peft_type="lora", rank=4, encoders=enc(), **c["kw"])
    oc, og = tc.iteration(0), tg.iteration(0)
    print(case, "obj", oc, og)
    for (n, pc), (_, pg) in zip(tc.network.named_parameters(), tg.network.named_parameters()):
        print("   %-60s cpu |g| %.3e max %.3e   gpu |g| %.3e max %.3e   diff %.3e" % (n[-58:], float(pc.grad.norm()), float(pc.grad.abs().max()), float(pg.grad.norm()), float(pg.grad.abs().max()), float((pc.grad - pg.grad.cpu()).abs().max())))
