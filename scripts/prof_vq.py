import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200"))
import torch
from vaekit import functional as F
lat = torch.nn.functional.leaky_relu(torch.randn(1024, 64, 16, 16, device="cuda")); cb = torch.randn(512, 64, device="cuda")
for _ in range(3):
    F.vector_quantize(lat, cb, 0.25)
torch.cuda.synchronize()
