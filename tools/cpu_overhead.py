"""Host-side cost of one learner step (python + ctypes + launches) against its device time: the step is GPU bound as
long as the issue time stays below the device time."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import bench  # noqa: E402
from gnn_hex_b200 import ModernTwoHeaded  # noqa: E402
from helpers import loss_inputs  # noqa: E402

dev = torch.device("cuda", 0)
m = ModernTwoHeaded(110, 15, 2).to(dev)
b = bench.make_batch("play", 0)
hb = b.pin_memory()
a, t, w = [u.to(dev) for u in loss_inputs(b)]


def step():
    m.zero_grad(set_to_none=True)
    d = hb.to(dev, non_blocking=True)
    return m.train_step(d.x, d.edge_index, d.batch, d.ptr, action=a, target=t, weight=w)[0]


for _ in range(20):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(40):
    step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("host issue %.1f us/step, host+device %.1f us/step" % ((t1 - t0) / 40 * 1e6, (t2 - t0) / 40 * 1e6))
if len(sys.argv) > 1:
    import cProfile, pstats
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(200):
        step()
    pr.disable()
    torch.cuda.synchronize()
    pstats.Stats(pr).sort_stats("tottime").print_stats(14)
