"""BASELINE config 1 batch (32 x 1024): eager launches vs the same forward replayed from a CUDA graph (torch.cuda.CUDAGraph
captures the library's launches on torch's capture stream).  Also the pure-GPU time of one forward (CUDA events around a
single graph replay) to show how much of the latency is launch overhead."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from oracle import pointnet_numpy as onp          # parameter init only
from tests import synth
from pointcloudprocessing_b200.pointnet import PointNet

b, n = 32, 1024
m = PointNet(23, 12, 0.3, 42, precision="bf16")
m.set_weights_dict(onp.make_params(23, 12, seed=7))
x = torch.from_numpy(synth.gaussian_clouds(4004, b, n)).cuda()
for heads in (("cls",), ("cls", "seg")):
    for _ in range(3):
        ref = m(x, heads=heads)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(50):
        out = m(x, heads=heads)
    torch.cuda.synchronize()
    eager = (time.perf_counter() - t0) / 50 * 1e6
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(2):
            m(x, heads=heads)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        gout = m(x, heads=heads)
    g.replay(); torch.cuda.synchronize()
    same = all(torch.equal(a, c) for a, c in zip(ref, gout) if a is not None)
    t0 = time.perf_counter()
    for _ in range(50):
        g.replay()
    torch.cuda.synchronize()
    graph = (time.perf_counter() - t0) / 50 * 1e6
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    print(heads, "eager %.1f us  graph %.1f us  single replay %.1f us  identical outputs: %s" % (eager, graph, e0.elapsed_time(e1) * 1e3, same), flush=True)
