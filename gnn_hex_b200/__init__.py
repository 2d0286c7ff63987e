"""gnn_hex_b200 -- B200-native fused forward/backward of GNN_Hex's `modern_two_headed` Q-network.

The package is a thin host layer over libhexgnn.so (include/hexgnn.h).  Importing it never builds
or falls back: a missing library raises on first use (python -m gnn_hex_b200.build).
"""
from ._lib import HexGnnError  # noqa: F401
from .models import FusedAdam, ModernTwoHeaded, get_pre_defined  # noqa: F401
from .ops import (CsrBlob, adam_step, build_csr, fused_backward, fused_forward, fused_train_step, mask_argmax,  # noqa: F401
                  td_loss, td_target)
from . import data  # noqa: F401
from . import custom_ops  # noqa: F401  (registers torch.ops.hexgnn.*)
from .learner import RainbowLearner  # noqa: F401
from .hexara import SAGETorchScript, get_current_model  # noqa: F401
