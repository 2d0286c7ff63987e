"""The Rainbow learner step around the fused kernels (SURVEY 8 f1 + f2): what `Rainbow.train` does between sampling a
batch and returning its statistics (GN0/RainbowDQN/Rainbow/common/rainbow.py:259-277 td_target, :279-338 train), with
every piece on the device and no host synchronisation inside the step:

    next_state: ONE CSR build shared by q_policy(advantages_only) and q_target       rainbow.py:263,268
                terminal mask + per-graph arg-max (hexgnn_mask_argmax)               rainbow.py:264-266
                td_tgt = r + gamma^n * q_target[best] * (1 - done) (hexgnn_td_target) rainbow.py:268-269
    state:      hexgnn_train_step = forward + PER-weighted loss + backward           rainbow.py:301-316
                gradient all-reduce when data parallel (no reference counterpart)
                clip_grad_norm_ + Adam (hexgnn_adam_step)                            rainbow.py:320-326
    returns     loss, |td_err| + 1e-6 priorities, grad_norm, td_est / td_tgt means   rainbow.py:307-308,331-338

The replay buffer, its priority tree and the environment stay the caller's (out of scope); they hand in the sampled
`state`, `next_state` batches (x, edge_index, batch, ptr) and `action`, `reward`, `done`, `weights` tensors."""
import torch

from . import ops
from .models import FusedAdam


class RainbowLearner:
    def __init__(self, q_policy, q_target, lr, adam_eps=1.5e-4, max_grad_norm=10.0, gamma=0.97, n_step=1,
                 double_dqn=True, loss_fn="mse", skip_untaken=False, world_size=1):
        self.q_policy, self.q_target = q_policy, q_target
        self.double_dqn = bool(double_dqn)
        self.n_step_gamma = float(gamma) ** int(n_step)            # rainbow.py:75
        self.loss_fn = "mse" if loss_fn == "mse" else "smooth_l1"   # rainbow.py:84-85
        self.world_size = int(world_size)
        self.max_grad_norm = float(max_grad_norm)
        if getattr(q_policy, "noisy_dqn", False) or getattr(q_policy, "norm", False):
            # noisy nets / norm=True: the advantage layer's mu / sigma and the LayerNorm weights live outside the flat
            # parameter block FusedAdam walks, so the optimizer is the reference's own pair (rainbow.py:79,320-326) over
            # the named parameters
            self.opt = torch.optim.Adam(q_policy.parameters(), lr=lr, eps=adam_eps)
            self._fused_opt = False
        else:
            self.opt = FusedAdam(q_policy, lr=lr, eps=adam_eps, max_grad_norm=max_grad_norm, skip_untaken=skip_untaken)
            self._fused_opt = True

    def sync_Q_target(self):
        """rainbow.py:163-164."""
        self.q_target.copy_parameters_from(self.q_policy)

    @torch.no_grad()
    def td_target(self, reward, next_state, done):
        """rainbow.py:259-277.  `next_state`: x, edge_index, batch, ptr on the device."""
        ns = next_state
        if getattr(self.q_target, "noisy_dqn", False):
            self.q_target.reset_noise()                              # rainbow.py:261
        csr = self.q_policy.build_csr(ns.edge_index, ns.batch, ns.ptr, num_nodes=int(ns.x.shape[0]),
                                      num_graphs=int(ns.ptr.shape[0]) - 1)
        q_t = self.q_target(ns.x, ns.edge_index, ns.batch, ns.ptr, csr=csr)
        if self.double_dqn:
            adv = self.q_policy(ns.x, ns.edge_index, ns.batch, ns.ptr, advantages_only=True, csr=csr).reshape(-1)
            _, best = ops.mask_argmax(adv, ns.ptr, mask_terminals=True, mask_value=-5.0)
            return ops.td_target(q_t, best, reward, done, self.n_step_gamma)
        mx, _ = ops.mask_argmax(q_t, ns.ptr, mask_terminals=True, mask_value=-5.0)
        return reward.float() + self.n_step_gamma * mx * (1.0 - done.float())

    @torch.no_grad()
    def act(self, batch, eps: float = 0.0):
        """rainbow.py:214-239 on an already collated device batch: greedy action per graph from the advantages with the two
        terminal rows masked (-5), as LOCAL node indices; `eps` > 0 replaces actions by uniform random legal nodes on the
        host exactly like the reference's loop (one sync).  Returns (actions [B] int64 on the device, exploratories np.bool)."""
        import random
        import numpy as np
        adv = self.q_policy(batch.x, batch.edge_index, batch.batch, batch.ptr, advantages_only=True).reshape(-1)
        _, best = ops.mask_argmax(adv, batch.ptr, mask_terminals=True, mask_value=-5.0)
        actions = best - batch.ptr[:-1].to(best.device)
        exploratories = np.zeros(actions.shape[0], dtype=bool)
        if eps > 0:
            sizes = (batch.ptr[1:] - batch.ptr[:-1]).tolist()
            for i, n in enumerate(sizes):
                if random.random() < eps and n > 2:
                    actions[i] = random.randint(2, n - 1)
                    exploratories[i] = True
        return actions, exploratories

    def train(self, state, next_state, action, reward, done, weights=None):
        """rainbow.py:279-338 without the buffer: returns a dict of DEVICE tensors (no sync):
        loss, grad_norm, priorities (|td_err| + 1e-6, :307-308), est_mean, gt_mean (:331-336)."""
        td_tgt = self.td_target(reward, next_state, done)
        self.q_policy.zero_grad(set_to_none=True)
        loss, td_err, q = self.q_policy.train_step(state.x, state.edge_index, state.batch, state.ptr, action=action,
                                                   target=td_tgt, weight=weights, kind=self.loss_fn,
                                                   scale=1.0 / self.world_size)
        if self.world_size > 1:
            self.q_policy.allreduce_gradients()
        if self._fused_opt:
            grad_norm = self.opt.step(state.x)
        else:
            grad_norm = torch.nn.utils.clip_grad_norm_(self.q_policy.parameters(), self.max_grad_norm)   # rainbow.py:320
            # same guard as the fused optimizer, without a host sync: a non-finite norm zeroes the gradients, which leaves
            # Adam's parameters untouched up to its momentum (torch.optim.Adam has no device-side skip)
            ok = torch.isfinite(grad_norm)
            for p in self.q_policy.parameters():
                if p.grad is not None:
                    p.grad.mul_(ok).nan_to_num_(nan=0.0, posinf=0.0, neginf=0.0)
            self.opt.step()
        td_est = td_err + td_tgt
        # `status` (device int32 [1], read it lazily): non-zero when a kernel flagged the batch -- bit 1 = a graph exceeds
        # the tile capacity (128 nodes / 1024 directed edges / 64 graphs per tile: its rows come back NaN).  The fused
        # optimizer skips a step whose gradient norm is not finite (parameters, moments and step counters untouched), so
        # one oversize position cannot poison the network; `grad_norm` then reports the non-finite value.
        return {"loss": loss, "grad_norm": grad_norm, "priorities": td_err.abs() + 1e-6,
                "est_mean": td_est.mean(), "gt_mean": td_tgt.mean(), "td_target": td_tgt, "q": q,
                "status": self.q_policy.last_csr.status}
