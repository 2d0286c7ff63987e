"""Golden vectors of the MODEL half of the hot path, produced by the reference's OWN code.

Run in the authoring container only (needs /root/reference):   python tests/golden/make_model_fixtures.py

`GN0/models.py` is imported unmodified, where it lies, on the test-only stand-in for torch_geometric / torch_scatter
(oracle/pyg_stub/, see its README: only the third-party primitives SAGEConv / LayerNorm / scatter are restated there).
For each case the reference's `get_pre_defined("modern_two_headed", args)` module gets name-hashed weights
(tests/helpers.set_hash_weights: reproducible anywhere, so they are not stored), runs on committed positions
(tests/golden/pos_*.npz, played by the reference's compiled C++ game) in the three output modes x both heads, and
`loss.backward()` of the learner loss (rainbow.py:301-312).  Saved per case and head flag m in {1, 0}:
    q_m, adv_m, value_m, centred_m   fp32 outputs of forward / advantages_only / seperate          (models.py:537-584)
    loss_m, grad_m                   fp32 loss and gradient (all named parameters, sorted-name order; strided by `gstride`)
    q64_m, grad64_m                  the same from the reference module in fp64 (the arbiter of fp32 round-off)
    embeds_m                         final_conv_acts (models.py:553), fp32, every k-th row
plus action / target / weight, count, and the parameter-name list.
"""
import os
import sys
from argparse import Namespace

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers  # noqa: E402
import positions  # noqa: E402
from gnn_hex_b200 import data  # noqa: E402

CASES = {
    # tag: (hidden, layers, head layers, norm, noisy, positions, count, gradient stride, weight gain)
    "S": (35, 10, 2, False, False, "pos_7x7_seed1234.npz", 256, 1, 1.5),        # BASELINE configs[0] at full size
    "L": (110, 15, 2, False, False, "pos_11x11_seed1234.npz", 24, 7, 1.7),      # GNN-L (configs[1] model), 24 positions
    "S_norm": (35, 10, 2, True, False, "pos_7x7_seed1234.npz", 48, 1, 1.5),     # --norm True (PyG LayerNorm, whole batch)
    "S_noisy": (35, 10, 2, False, True, "pos_7x7_seed1234.npz", 48, 1, 1.5),    # --noisy_dqn True
}


def args_of(case):
    H, L, Lh, norm, noisy = case[:5]
    return Namespace(hidden_channels=H, num_layers=L, num_head_layers=Lh, norm=norm, noisy_dqn=noisy, noisy_sigma0=0.5)


def batch_of(case, maker):
    return data.collate(positions.load_positions(case[5], maker=maker)[:case[6]])


def run_case(ref, tag):
    case = CASES[tag]
    out = {"gstride": np.int64(case[7]), "count": np.int64(case[6]), "gain": np.float64(case[8])}
    for maker in (1, 0):
        b = batch_of(case, maker)
        action, target, weight = helpers.loss_inputs(b)
        res = {}
        for dt in (torch.float32, torch.float64):
            model = ref.get_pre_defined("modern_two_headed", args_of(case))
            helpers.set_hash_weights(model, seed=0, gain=case[8])
            model = model.to(dt)
            x = b.x.to(dt)
            q = model(x, b.edge_index, b.batch, b.ptr)
            emb = model.final_conv_acts.detach().clone()
            loss = torch.mean(weight.to(dt) * (target.to(dt) - q[action + b.ptr[:-1]]) ** 2)   # rainbow.py:301-312 (MSE)
            loss.backward()
            names = sorted(dict(model.named_parameters()))
            g = helpers.flat_named(model, "grad", names)
            # (grad mode stays on: the reference registers a hook on the embeddings in every call, models.py:554)
            adv = model(x, b.edge_index, b.batch, b.ptr, advantages_only=True)
            v, c = model(x, b.edge_index, b.batch, b.ptr, seperate=True)
            res[dt] = (q.detach(), adv.detach(), v.detach(), c.detach(), loss.detach(), g, emb, names)
        q, adv, v, c, loss, g, emb, names = res[torch.float32]
        s = int(case[7])
        out[f"q_{maker}"] = q.numpy()
        out[f"adv_{maker}"] = adv.numpy()
        out[f"value_{maker}"] = v.numpy()
        out[f"centred_{maker}"] = c.numpy()
        out[f"loss_{maker}"] = loss.numpy()
        out[f"grad_{maker}"] = g[::s].astype(np.float32)
        out[f"embeds_{maker}"] = emb.numpy()[:: max(1, emb.shape[0] // 64)]
        out[f"q64_{maker}"] = res[torch.float64][0].numpy()
        out[f"grad64_{maker}"] = res[torch.float64][5][::s]
        out[f"gradmax_{maker}"] = np.float64(np.abs(res[torch.float64][5]).max())
        if maker == 1:
            out["action"], out["target"], out["weight"] = action.numpy(), target.numpy(), weight.numpy()
            out["names"] = np.array(names)
    return out


def main():
    ref = helpers.import_reference_models()
    if ref is None:
        raise SystemExit("make_model_fixtures.py needs /root/reference (authoring container only)")
    for tag in CASES:
        out = run_case(ref, tag)
        path = os.path.join(HERE, f"model_{tag}.npz")
        np.savez_compressed(path, **out)
        print(tag, "->", path, f"{os.path.getsize(path) / 1024:.0f} KB")


if __name__ == "__main__":
    main()
