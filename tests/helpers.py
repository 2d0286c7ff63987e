"""Shared helpers for the parity tests (test infrastructure; may import oracle/)."""
import os
import sys
import zlib

import numpy as np
import torch

from gnn_hex_b200 import ModernTwoHeaded, data
from oracle import hexgnn_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
REFERENCE_ROOT = "/root/reference"
PYG_STUB = os.path.join(ROOT, "oracle", "pyg_stub")

REL_TOL = 1e-5  # BASELINE.json north_star: 1e-5 relative (fp32) on Q-values and gradients
# Per-tensor bound on the kernel's distance from the fp64 truth, in units of the fp32 ORACLE's own
# distance from it.  The 3xTF32 tensor-core contraction has exact products but the tensor core's fp32
# accumulator rounds 3x as often (and by truncation), measured at <= ~12x the torch fp32 round-off on
# the tiniest (1e-3 of the largest) gradient tensors; the whole-gradient 1e-5 bar is unaffected.
ROUNDOFF_FACTOR = 16.0
# Floor of the per-tensor bound (relative to that tensor's own largest gradient entry).  The fp32 CUDA-core
# path meets 1e-5 here as well; the tensor core's truncating fp32 accumulator leaves a ~1e-6/layer
# systematic bias in the activations that shows up as <= ~2e-5 on individual head tensors.
PER_TENSOR_TOL = 4e-5


def make_pair(hidden, layers, head_layers=2, seed=0, dtype=torch.float32, noisy_dqn=False, noisy_sigma0=0.5):
    """Oracle + product module with identical weights (default torch init, seeded)."""
    torch.manual_seed(seed)
    o = orc.ModernTwoHeadedOracle(hidden, layers, head_layers, noisy_dqn=noisy_dqn, noisy_sigma0=noisy_sigma0)
    m = ModernTwoHeaded(hidden, layers, head_layers, noisy_dqn=noisy_dqn, noisy_sigma0=noisy_sigma0)
    m.load_state_dict(o.state_dict())
    if dtype != torch.float32:
        o = o.to(dtype)
    return o, m


def loss_inputs(batch, seed=1):
    """a_g ~ randint(2, n_g), t_g ~ U(-1,1), w_g ~ U(0.5,1)  (SURVEY.md §8d config 1)."""
    g = torch.Generator().manual_seed(seed)
    n = (batch.ptr[1:] - batch.ptr[:-1])
    u = torch.rand(batch.num_graphs, generator=g)
    action = (2 + (u * (n - 2).clamp(min=1)).long()).clamp(max=n - 1).clamp(min=0)
    target = torch.rand(batch.num_graphs, generator=g) * 2 - 1
    weight = torch.rand(batch.num_graphs, generator=g) * 0.5 + 0.5
    return action, target, weight


def rel_err(a, b):
    """max|a-b| / max|b| (b = reference)."""
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    denom = b.abs().max().clamp(min=1e-30)
    return float((a - b).abs().max() / denom)


def oracle64(o):
    """fp64 copy of an oracle module: the arbiter of fp32 round-off."""
    h = o.gnn.hidden_channels
    noisy = isinstance(o.maker_head.linear, orc.NoisyLinearOracle)
    o64 = orc.ModernTwoHeadedOracle(h, o.gnn.num_layers, o.maker_head.gnn.num_layers, noisy_dqn=noisy).double()
    o64.load_state_dict({k: v.double() for k, v in o.state_dict().items()})
    return o64


def grad_errs(model, oracle, oracle_fp64=None):
    """Gradient parity of the product module against the oracle.

    Returns (global_err, worst_excess, name):
      global_err   max|g - ref| / max|ref| over the WHOLE flat gradient -- the 1e-5 bar of BASELINE.json.
      worst_excess per tensor, with scale_t = max(|ref_t|_max, 1e-3 * |ref|_max):
                   err_t / max(REL_TOL, ROUNDOFF_FACTOR * err_t(fp32 oracle vs fp64 oracle)); <= 1 means the kernel is
                   within fp32 round-off of the truth wherever the fp32 oracle itself is.  Deep-layer
                   gradients are ~1e-3 of the largest one, so their per-tensor relative round-off (of
                   ANY fp32 implementation, the oracle included) exceeds 1e-5; fp64 arbitrates.
    ref = the fp64 oracle when given, else the fp32 oracle."""
    r32 = dict(oracle.named_parameters())
    ref = dict(oracle_fp64.named_parameters()) if oracle_fp64 is not None else r32
    gmax = max(float(p.grad.abs().max()) for p in ref.values() if p.grad is not None)
    glob, worst, name_worst = 0.0, 0.0, None
    for name, p in model.named_parameters():
        rg = ref[name].grad
        rg = torch.zeros_like(ref[name]) if rg is None else rg
        rg = rg.double()
        g = torch.zeros_like(rg) if p.grad is None else p.grad.detach().cpu().double()
        scale = max(float(rg.abs().max()), 1e-3 * gmax)
        d = float((g - rg).abs().max())
        glob = max(glob, d / gmax)
        allow = PER_TENSOR_TOL
        if oracle_fp64 is not None and r32[name].grad is not None:
            allow = max(PER_TENSOR_TOL, ROUNDOFF_FACTOR * float((r32[name].grad.double() - rg).abs().max()) / scale)
        e = d / scale / allow
        if e > worst:
            worst, name_worst = e, name
    return glob, worst, name_worst


# ---- the reference's own module code (GN0/models.py, unmodified) on the test-only PyG stand-in -----------------------
def import_reference_models():
    """`GN0.models` of the reference, imported where it lies with oracle/pyg_stub/ standing in for torch_geometric /
    torch_scatter (absent from the image).  Returns None when /root/reference is not mounted (the GPU box): callers skip
    and rely on the committed golden vectors (tests/golden/model_*.npz) instead."""
    if not os.path.isdir(os.path.join(REFERENCE_ROOT, "GN0")):
        return None
    for p in (REFERENCE_ROOT, PYG_STUB):
        if p not in sys.path:
            sys.path.insert(0, p)
    import GN0.models as ref_models
    return ref_models


def import_reference_torch_script_models():
    """`GN0.torch_script_models` of the reference (HexAra networks), imported where it lies on the PyG stand-in; None when
    /root/reference is not mounted."""
    if not os.path.isdir(os.path.join(REFERENCE_ROOT, "GN0")):
        return None
    for p in (REFERENCE_ROOT, PYG_STUB):
        if p not in sys.path:
            sys.path.insert(0, p)
    import GN0.torch_script_models as ref
    return ref


def hash_uniform(name: str, shape, seed: int = 0) -> torch.Tensor:
    """U(-1, 1) values that depend only on (name, seed, element index): splitmix64 in integer arithmetic, so the golden
    weights are reproducible on any box / library version without being stored."""
    n = int(np.prod(shape)) if len(shape) else 1
    start = (zlib.crc32(name.encode()) * 0x9E3779B97F4A7C15 + seed * 0xD1B54A32D192ED03) & 0xFFFFFFFFFFFFFFFF
    with np.errstate(over="ignore"):
        z = np.arange(n, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(start)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    u = (z >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)
    return torch.from_numpy((2.0 * u - 1.0).astype(np.float32)).reshape(tuple(shape))


@torch.no_grad()
def set_hash_weights(module: torch.nn.Module, seed: int = 0, gain: float = 1.0) -> None:
    """Fill every parameter / buffer of `module` by NAME (same names in the reference model, the oracle and the product
    module): matrices U(+-gain/sqrt(fan_in)) (gain 1 = the default init's range), vectors U(+-0.1), LayerNorm weights
    1 + U(+-0.2).  With gain 1 a 12..17-layer ReLU stack contracts to node-independent embeddings (advantage spread 1e-5 of
    its mean: Q = v + a - mean(a) is then pure cancellation); the golden cases use gains 1.5 / 1.7, which keep the
    activations O(1) and the advantages spread over the graph."""
    for name, t in list(module.named_parameters()) + list(module.named_buffers()):
        u = hash_uniform(name, t.shape, seed)
        if t.dim() >= 2:
            v = gain * u / float(t.shape[-1]) ** 0.5
        elif "norm" in name and name.endswith("weight"):
            v = 1.0 + 0.2 * u
        else:
            v = 0.1 * u
        t.copy_(v.to(t.dtype))


def load_model_fixture(tag: str):
    """tests/golden/model_<tag>.npz as a dict (made by tests/golden/make_model_fixtures.py from the reference's code)."""
    z = np.load(os.path.join(GOLDEN_DIR, f"model_{tag}.npz"), allow_pickle=False)
    return {k: z[k] for k in z.files}


def flat_named(module, what="grad", names=None):
    """Concatenate `p.grad` (zeros for None) or `p` over named_parameters, in sorted-name order -> float64 numpy."""
    d = dict(module.named_parameters())
    out = []
    for k in (names if names is not None else sorted(d)):
        p = d[k]
        t = p if what == "param" else (p.grad if p.grad is not None else torch.zeros_like(p))
        out.append(t.detach().double().cpu().reshape(-1))
    return torch.cat(out).numpy()
