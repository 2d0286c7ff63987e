"""Golden vectors of the HexAra policy / value network, produced by the reference's OWN code.

Run in the authoring container only (needs /root/reference):   python tests/golden/make_hexara_fixtures.py

`GN0/torch_script_models.py` is imported unmodified, where it lies, on the test-only stand-in for torch_geometric /
torch_scatter (oracle/pyg_stub/).  `get_current_model("SAGE", ...)` (:495-500) builds `SAGE_torch_script` (:286-379), which gets
name-hashed weights (tests/helpers.set_hash_weights) and runs on committed positions in the C++ feature convention
(tests/positions.load_positions_cpp).  Saved per case and swap_allowed s in {0, 1}, fp32 and fp64:
    pi_s, value_s, gi_s, ptr_s (+ pi64_s, value64_s) = the module's 4-tuple (:379);  embeds_s = final_conv_acts (:314), strided
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

import helpers  # noqa: E402
import positions  # noqa: E402
from gnn_hex_b200 import data  # noqa: E402

CASES = {
    # tag: (hidden, body layers, policy layers, value layers, positions, count, gain)
    "ara_S": (28, 5, 2, 2, "pos_7x7_seed1234.npz", 64, 1.5),
    "ara_L": (60, 15, 2, 2, "pos_11x11_seed1234.npz", 32, 1.7),      # get_current_model's defaults
    "ara_P1": (44, 3, 1, 3, "pos_7x7_seed1234.npz", 24, 1.5),        # one-conv policy stack (hidden -> 1 directly)
}


def main():
    ref = helpers.import_reference_torch_script_models()
    assert ref is not None, "needs /root/reference"
    for tag, (H, L, Lp, Lv, pos, count, gain) in CASES.items():
        b = data.collate(positions.load_positions_cpp(pos)[:count])
        out = {"count": np.int64(count), "gain": np.float64(gain)}
        for swap in (0, 1):
            for dt in (torch.float32, torch.float64):
                m = ref.get_current_model("SAGE", hidden_channels=H, hidden_layers=L, policy_layers=Lp, value_layers=Lv,
                                          swap_allowed=bool(swap))
                helpers.set_hash_weights(m, seed=0, gain=gain)
                m = m.to(dt)
                pi, value, gi, ptr = m(b.x.to(dt), b.edge_index, b.batch, b.ptr)   # (grad mode on: the module registers a hook)
                sfx = "" if dt == torch.float32 else "64"
                out[f"pi{sfx}_{swap}"] = pi.detach().numpy()
                out[f"value{sfx}_{swap}"] = value.detach().numpy()
                if dt == torch.float32:
                    out[f"gi_{swap}"] = gi.numpy()
                    out[f"ptr_{swap}"] = ptr.numpy()
                    emb = m.final_conv_acts.detach().numpy()
                    out[f"embeds_{swap}"] = emb[:: max(1, emb.shape[0] // 64)]
            out["names"] = np.array(list(m.state_dict().keys()))
        np.savez_compressed(os.path.join(HERE, f"model_{tag}.npz"), **out)
        print(tag, {k: v.shape for k, v in out.items() if hasattr(v, "shape") and v.ndim}, "value spread", float(np.ptp(out["value64_1"])))


if __name__ == "__main__":
    sys.exit(main())
