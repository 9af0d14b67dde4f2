"""GPU-box experiment: logits error vs the reference golden (tests/golden/model_seed7.npz) and GRU launch time for the
state modes of the one-pass bf16 path (DP_GRU_STATE = f32 | mixed | bf16; read once per process)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from deepore_b200 import model, modelio
gold = np.load(os.path.join(ROOT, "tests", "golden", "model_seed7.npz"))
m = model.from_state(modelio.seeded_state(7), "cuda", "bf16")
img = torch.from_numpy(gold["img"]).cuda()
logits, pred = m.forward_windows(img)
torch.cuda.synchronize()
lg = logits.cpu().numpy()[:, :148]
d = np.abs(lg - gold["logits"])
print("DP_GRU_STATE=%s: logits max abs err %.4g, mean abs err %.4g, rms %.4g" % (os.environ.get("DP_GRU_STATE", "f32"), d.max(), d.mean(), np.sqrt((d ** 2).mean())))
