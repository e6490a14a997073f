"""Host-side profile of the VQ step (python bench workload): where do the ~300 us per step go?"""
import cProfile, os, pstats, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pytorch-vae_b200")); sys.path.insert(0, ROOT)
import bench
wl = bench.OP_WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "vq"]()
dev = torch.device("cuda", 0)
m = wl.model().to(dev)
t = wl.make(dev, 0, 1)
for _ in range(20):
    wl.step(m, t)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200):
    wl.step(m, t)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("host issue time %.1f us/step, with final sync %.1f us/step" % ((t1 - t0) / 200 * 1e6, (t2 - t0) / 200 * 1e6))
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    wl.step(m, t)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
