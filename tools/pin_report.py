"""Diagnostic (not a test): every error of the CUDA path against the reference's golden vectors (tests/golden/model_*.npz)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers, positions
from gnn_hex_b200 import ModernTwoHeaded, data
CASES = {"S": (35, 10, 2, False, "pos_7x7_seed1234.npz"), "L": (110, 15, 2, False, "pos_11x11_seed1234.npz"),
         "S_noisy": (35, 10, 2, True, "pos_7x7_seed1234.npz")}
rel = lambda a, b: float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / max(np.abs(b).max(), 1e-30))
for tag, (H, L, Lh, noisy, pos) in CASES.items():
    fx = helpers.load_model_fixture(tag)
    for maker in (1, 0):
        b = data.collate(positions.load_positions(pos, maker=maker)[:int(fx["count"])])
        m = ModernTwoHeaded(H, L, Lh, noisy_dqn=noisy, noisy_sigma0=0.5)
        helpers.set_hash_weights(m, seed=0, gain=float(fx["gain"]))
        names = [str(n) for n in fx["names"]]
        m = m.cuda(); d = b.to("cuda")
        action, target, weight = (torch.from_numpy(fx[k]).cuda() for k in ("action", "target", "weight"))
        q = m(d.x, d.edge_index, d.batch, d.ptr)
        loss = torch.mean(weight * (target - q[action + d.ptr[:-1]]) ** 2); loss.backward()
        with torch.no_grad():
            adv = m(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True)
            v, c = m(d.x, d.edge_index, d.batch, d.ptr, seperate=True)
        g = helpers.flat_named(m, "grad", names)[:: int(fx["gstride"])]
        gmax = float(fx[f"gradmax_{maker}"]); qs = float(np.abs(fx[f"q64_{maker}"]).max())
        print(tag, maker, "q %.2e q64 %.2e adv %.2e value/qs %.2e centred/qs %.2e loss %.2e grad64 %.2e grad32 %.2e | ref32-vs-64: q %.2e grad %.2e" % (
            rel(q.detach().cpu().numpy(), fx[f"q_{maker}"]), rel(q.detach().cpu().numpy(), fx[f"q64_{maker}"]),
            rel(adv.cpu().numpy(), fx[f"adv_{maker}"]), np.abs(v.cpu().numpy() - fx[f"value_{maker}"]).max() / qs,
            np.abs(c.cpu().numpy() - fx[f"centred_{maker}"]).max() / qs, rel(loss.detach().cpu().numpy(), fx[f"loss_{maker}"]),
            np.abs(g - fx[f"grad64_{maker}"]).max() / gmax, np.abs(g - fx[f"grad_{maker}"]).max() / gmax,
            rel(fx[f"q_{maker}"], fx[f"q64_{maker}"]), np.abs(fx[f"grad_{maker}"] - fx[f"grad64_{maker}"]).max() / gmax))
