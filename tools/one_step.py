"""Run N steps of the path on one config (for ncu metric sweeps): python tools/one_step.py [config] [steps]."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "speech-integration_b200"))
import torch
from ssi_b200 import synth
from ssi_b200.step import PathStep
cfg = synth.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else "c3"]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda", 0)
D, V = synth.HIDDEN, cfg.vocab
W = (torch.randn(V, D, generator=torch.Generator().manual_seed(1)) * 0.02).to(torch.bfloat16).to(dev)
tokens_h, labels_h = synth.make_tokens_labels(cfg, torch.Generator().manual_seed(2), batch=cfg.batch)
gd = torch.Generator(device=dev).manual_seed(3)
X = torch.randn(cfg.batch, cfg.seq, D, generator=gd, device=dev).to(torch.bfloat16)
dH0 = (torch.randn(cfg.batch, cfg.seq, D, generator=gd, device=dev) * 1e-3).to(torch.bfloat16)
path = PathStep(W, mode=os.environ.get("SSI_B200_MODE"))
for _ in range(steps):
    path.zero_grad(set_to_none=True)
    out = path(tokens_h.to(dev), labels_h.to(dev), X, dH0)
torch.cuda.synchronize()
print("loss", out[0].item())
