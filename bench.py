"""bench.py -- headline benchmark of the fused hot path (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload play|empty]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One "step" = one pass of the hot path over one replay batch of 256 11x11 Hex positions per GPU:
device CSR build + fused forward (with stash) + fused TD loss + fused backward + deterministic gradient reduction
(+ NCCL all-reduce of the flat gradient when N > 1).  The optimizer is NOT part of the path
 and is not timed.  Prints ONE JSON line on rank 0.

  value  graphs/s with the batch already resident in HBM (CUDA events, L2 flushed between steps)
  e2e    graphs/s through the public module API with HOST (pinned) inputs: H2D of x / edge_index /
         batch / ptr + forward + loss + backward + D2H of the loss, every step
  roofline / cpu_baseline: see DESIGN.md "Measurement".
`--impl reference` times the CPU restatement of the reference PyG path (oracle/, kind "port":
torch_geometric is not installable here) on the host cores for the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
if os.path.join(ROOT, "tests") not in sys.path:
    sys.path.insert(1, os.path.join(ROOT, "tests"))     # tests/positions.py: the committed position fixtures (bench inputs)

METRIC = "Hex graphs/sec (fwd+bwd, GNN-L 11x11 b256)"
UNIT = "graphs/s"
HIDDEN, LAYERS, HEAD_LAYERS, BATCH = 110, 15, 2, 256
FIXTURE = "pos_11x11_seed1234.npz"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="fused", choices=["fused", "reference"])
    ap.add_argument("--workload", default="play", choices=["play", "empty"],
                    help="play: 256 seeded random-playout positions (mean 77 nodes); empty: 256 empty boards (123 nodes)")
    ap.add_argument("--path", default="auto", choices=["auto", "simt", "tc"])
    ap.add_argument("--allreduce", default="auto", choices=["auto", "p2p", "pull", "mc", "nccl"],
                    help="N > 1: auto = the own multimem kernel from 4 ranks on where a multicast address exists (measured: "
                         "0.643 vs 0.652 ms/step at N = 8, a tie at N = 4, NCCL ahead at N = 2), else NCCL.  NCCL, or the library's own all-reduce over NVLink peer memory -- mc: two-shot in place "
                         "through the NVSwitch (multimem.ld_reduce / multimem.st), pull: one-shot rank-ordered pull, "
                         "p2p: mc where a multicast address exists, else pull")
    ap.add_argument("--api", default="train_step", choices=["train_step", "autograd"],
                    help="train_step: model.train_step (hexgnn_train_step: forward + TD loss + backward in one call, the "
                         "three tensor-core kernels overlapped); autograd: model(...) -> ops.td_loss -> loss.backward()")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        d["source"] = "measured"
        return d
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def make_batch(workload, rank=0, world=1):
    """This rank's shard of the global replay batch of 256 * world positions (weak scaling: 256 per GPU).  The global
    batch is the fixture repeated once per rank, rotated by 32 positions each time so the shards differ; it is split into
    contiguous graph ranges balanced by node count (data.shard_by_nodes)."""
    from gnn_hex_b200 import data
    import positions
    if workload == "empty":
        base = [positions.empty_board(11, 1)] * BATCH
    else:
        base = positions.load_positions(FIXTURE, maker=1)[:BATCH]
    gs = []
    for r in range(world):
        k = (32 * r) % len(base)
        gs += base[k:] + base[:k]
    return data.collate(data.shard_graphs(gs, world, rank))


def loss_inputs(batch, seed=1):
    """Rainbow learner loss inputs (rainbow.py:301-312): action per graph, TD target, PER weight."""
    import torch
    g = torch.Generator().manual_seed(seed)
    n = batch.ptr[1:] - batch.ptr[:-1]
    u = torch.rand(batch.num_graphs, generator=g)
    action = (2 + (u * (n - 2).clamp(min=1)).long()).clamp(max=n - 1).clamp(min=0)
    target = torch.rand(batch.num_graphs, generator=g) * 2 - 1
    weight = torch.rand(batch.num_graphs, generator=g) * 0.5 + 0.5
    return action, target, weight


def flops_train(batch):
    """Algorithmic FLOPs of one fwd+bwd over `batch`: 3 x forward."""
    H, L, Lh = HIDDEN, LAYERS, HEAD_LAYERS
    per_node = 4 * 2 * H + (L - 1 + Lh) * 4 * H * H + 2 * H
    fwd = int(batch.x.shape[0]) * per_node + batch.num_graphs * 2 * (4 * H * (H // 2) + (H // 2))
    return fwd, 3 * fwd


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 20 ms; stop(t0, t1) keeps the samples whose
    timestamp falls inside the timed region [t0, t1] (B200_PROFILING.md clocks line)."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "20"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self, t0, t1):
        import datetime
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.1)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        rows = []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [t.strip() for t in line.split(",")]
            if len(c) < 9:
                continue
            try:
                ts = datetime.datetime.strptime(c[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                rows.append((ts, float(c[1]), float(c[2]), float(c[3]),
                             [nm for nm, v in zip(names, c[5:9]) if v.lower().startswith("active")]))
            except ValueError:
                continue
        self.f.close()
        os.unlink(self.f.name)
        inside = [r for r in rows if t0 - 0.02 <= r[0] <= t1 + 0.02]
        where = "timed region"
        if not inside:
            inside, where = rows, "whole run (no sample fell inside the timed region)"
        sm = sorted(r[1] for r in inside)
        reasons = sorted({x for r in inside for x in r[4]})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max((r[2] for r in inside), default=None),
                "power_w_max": max((r[3] for r in inside), default=None), "samples": len(inside),
                "window": where, "reasons": reasons}


def read_ncu_summary():
    """Newest profiles/ncu_full_r*.csv (written by tools/ncu_summary.py from an `ncu --set full` capture of this command):
    per kernel the DRAM bytes per launch and sm__pipe_tensor_cycles_active (% of peak while the SM is active)."""
    import csv
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "ncu_full_r*.csv")), key=os.path.getmtime)
    for path in reversed(sorted(files, key=lambda f: os.path.basename(f))):
        try:
            with open(path) as f:
                rows = list(csv.reader(f))
            head = rows[0]
            table = {r[0]: r for r in rows[1:] if r}
            out = {}
            for col in range(2, len(head)):
                name = head[col]
                for key in ("sage_fwd_tc_kernel", "sage_bwd_tc_kernel", "sage_dw_tc_kernel", "csr_build_fused_kernel", "reduce_tc_kernel"):
                    if key in name:
                        rd, wr = float(table["dram__bytes_read.sum"][col]), float(table["dram__bytes_write.sum"][col])
                        unit = table["dram__bytes_read.sum"][1].lower()
                        mul = 1e6 if unit.startswith("mbyte") else 1e3 if unit.startswith("kbyte") else 1e9 if unit.startswith("gbyte") else 1.0
                        out[key] = {"dram_bytes": (rd + wr) * mul,
                                    "tensor_pct": float(table["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"][col])}
            if out:
                return {"file": os.path.relpath(path, ROOT), "kernels": out}
        except (OSError, KeyError, ValueError, IndexError):
            continue
    return None


def time_cpu_port(batch, steps, warmup, max_seconds=None, device=None):
    """fwd+bwd of the CPU restatement of the reference PyG path (oracle/, test infrastructure used
    here only as the reported CPU baseline) on all host threads.  Returns (graphs/s, steps run, cores).
    `device` = a CUDA device: the same torch-ops restatement on the GPU's STOCK kernels (index_select / index_add /
    addmm ...), the "reference-equivalent on the same box" bar of SURVEY.md 8(d) -- reported beside the CPU number."""
    import torch
    from oracle import hexgnn_oracle as orc
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    o = orc.ModernTwoHeadedOracle(HIDDEN, LAYERS, HEAD_LAYERS)
    action, target, weight = loss_inputs(batch)
    if device is not None:
        o = o.to(device)
        batch = batch.to(device)
        action, target, weight = action.to(device), target.to(device), weight.to(device)

    def step():
        o.zero_grad(set_to_none=True)
        q = o(batch.x, batch.edge_index, batch.batch, batch.ptr)
        loss = orc.rainbow_loss(q, batch.ptr, action, target, weight)
        loss.backward()
        return float(loss.detach())      # (on the GPU: the host sync that ends the step)

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        step()
        done += 1
        if max_seconds is not None and time.perf_counter() - t0 > max_seconds:
            break
    dt = time.perf_counter() - t0
    return batch.num_graphs * done / dt, done, cores, dt


def run_reference(args, rank, world):
    """Reference arm: the reference's own CPU implementation of the path.  torch_geometric /
    torch_scatter are absent and not installable (no network), and the model is Python, so this is
    the oracle port (kind "port") on all host cores; rank 0 only."""
    if rank != 0:
        return
    batch = make_batch(args.workload)
    value, done, cores, dt = time_cpu_port(batch, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": done, "warmup": args.warmup, "ms_per_step": 1e3 * dt / done, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, batch, 1),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{done} steps of the full {batch.num_graphs}-graph batch, torch fp32 oracle"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(args, batch, world):
    return {"workload": f"GNN-L modern_two_headed fwd+bwd, hidden {HIDDEN}, {LAYERS}+{HEAD_LAYERS} SAGE layers, "
                        f"{BATCH} 11x11 Hex positions per GPU ({args.workload})",
            "graphs_per_gpu": batch.num_graphs, "global_batch": batch.num_graphs * world,
            "nodes_per_gpu": int(batch.x.shape[0]), "edges_per_gpu": int(batch.edge_index.shape[1]),
            "positions": ("tests/golden/" + FIXTURE + " (reference C++ game, seeded random playouts)")
            if args.workload == "play" else "256 empty 11x11 boards (123 nodes / 644 edges each)",
            "parallelism": f"dp{world}", "l2": "flushed between steps (256 MiB write)",
            "optimizer": "excluded (not on the path)", "path": args.path}


def run_fused(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from gnn_hex_b200 import ModernTwoHeaded, _lib, ops

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the fused path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    torch.manual_seed(0)
    model = ModernTwoHeaded(HIDDEN, LAYERS, HEAD_LAYERS, path=args.path).to(dev)
    collective = "none"
    if world > 1:
        collective = "nccl all_reduce"
        if args.allreduce == "auto":
            args.allreduce = "mc_or_nccl" if world >= 4 else "nccl"
        if args.allreduce == "mc_or_nccl":
            try:
                ok = model.enable_p2p_allreduce(mode="mc")
            except Exception:  # noqa: BLE001 -- no multicast address on this box: NCCL (every rank takes the same branch)
                ok = False
            if ok:
                collective = "hexgnn_allreduce_mc (own kernel: two-shot in place through the NVSwitch, multimem.ld_reduce / .st)"
        elif args.allreduce != "nccl" and model.enable_p2p_allreduce(mode={"p2p": "auto"}.get(args.allreduce, args.allreduce)):
            collective = ("hexgnn_allreduce_mc (own kernel: two-shot in place through the NVSwitch, multimem.ld_reduce / .st)"
                          if model._p2p["mc"] else "hexgnn_allreduce_p2p (own kernel: one-shot pull over NVLink peer memory)")
    batch = make_batch(args.workload, rank, world)
    hb = batch.pin_memory()                      # HOST copy of the step's inputs (e2e arm)
    db = batch.to(dev)                           # resident copy (value arm)
    action, target, weight = [t.to(dev) for t in loss_inputs(batch)]
    global_graphs = batch.num_graphs
    if world > 1:
        t = torch.tensor([batch.num_graphs], dtype=torch.int64, device=dev)
        dist.all_reduce(t)
        global_graphs = int(t.item())
    scale = batch.num_graphs / global_graphs     # loss = mean over the GLOBAL batch (rainbow.py:312)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2
    # what the learner reads back every step: the loss and the B TD errors that become the PER priorities (rainbow.py:306-309)
    out_host = torch.empty(1 + batch.num_graphs, dtype=torch.float32).pin_memory()

    fused_api = args.api == "train_step" and args.path != "simt"

    def fwd_loss_bwd(b):
        if fused_api:
            # rainbow.py:301-316 in one call; sets p.grad exactly like the autograd route (bit-identical, tested)
            loss, td_err, _ = model.train_step(b.x, b.edge_index, b.batch, b.ptr, action=action, target=target,
                                               weight=weight, kind="mse", scale=scale)
            return loss, td_err
        q = model(b.x, b.edge_index, b.batch, b.ptr)
        loss, td_err = ops.td_loss(q, b.ptr, action, target, weight, "mse", scale)
        loss.backward()
        return loss, td_err

    def step_resident():
        model.zero_grad(set_to_none=True)
        loss, _ = fwd_loss_bwd(db)
        if world > 1:
            model.allreduce_gradients()
        return loss

    def step_e2e():
        model.zero_grad(set_to_none=True)
        b = hb.to(dev, non_blocking=True)
        loss, td_err = fwd_loss_bwd(b)
        if world > 1:
            model.allreduce_gradients()
        if fused_api:
            out_host.copy_(ops.loss_and_td_err(loss), non_blocking=True)     # loss + TD errors: one buffer, one D2H copy
        else:
            out_host[:1].copy_(loss.detach().reshape(1), non_blocking=True)
            out_host[1:].copy_(td_err.detach(), non_blocking=True)
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(step_fn, steps, sync_each=False):
        """Sum of per-step CUDA-event durations on the current stream; L2 flushed between steps."""
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        for s, e in evs:
            flush.fill_(1)
            s.record()
            step_fn()
            e.record()
            if sync_each:
                e.synchronize()       # e2e: the caller reads the loss before the next step
        barrier()
        ms = sum(s.elapsed_time(e) for s, e in evs)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    sampler = ClockSampler(local_rank) if rank == 0 else None
    for _ in range(args.warmup):
        step_resident()
    for _ in range(max(3, args.warmup // 2)):
        step_e2e()
    barrier()

    t_wall0 = time.time()
    _lib.profile_enable(True)
    _lib.profile_read(0); _lib.profile_read(1); _lib.profile_read(2)
    l0 = ops.launch_counter
    ms_total = timed(step_resident, args.steps)
    launches = ops.launch_counter - l0
    fwd_ms, fwd_n = _lib.profile_read(0)
    bwd_ms, bwd_n = _lib.profile_read(1)
    trio_ms, trio_n = _lib.profile_read(2)
    _lib.profile_enable(False)
    ms_e2e = timed(step_e2e, args.steps, sync_each=True)
    # the HBM-bound integer phase on its own (north_star: "achieved HBM GB/s for the edge / feature streams"): the device
    # CSR builder over the resident batch, CUDA events per call, L2 flushed between calls; reported beside the main roofline
    csr_n = min(50, args.steps)
    csr_evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(csr_n)]
    for s_ev, e_ev in csr_evs:
        flush.fill_(1)
        s_ev.record()
        model.build_csr(db.edge_index, db.batch, db.ptr, num_nodes=int(db.x.shape[0]), num_graphs=batch.num_graphs)
        e_ev.record()
    torch.cuda.synchronize(dev)
    csr_ms = sum(s_ev.elapsed_time(e_ev) for s_ev, e_ev in csr_evs) / csr_n
    clocks = sampler.stop(t_wall0, time.time()) if sampler else None

    status = int(model.last_csr.status.item()) if getattr(model, "last_csr", None) is not None else 0
    if status:
        raise SystemExit(f"bench.py: device status {status} (a graph did not fit a tile?)")

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    graphs = global_graphs
    f_fwd, f_train = flops_train(batch)
    pk = peaks()
    # The node-linear contractions are fp32-parity emulations on the tensor cores: forward and dH chain as three fp16
    # products per algorithmic product (hi = fp16(x), lo = fp16(2^11 (x - hi)); kind::f16, K = 16), the dW GEMM as 3xTF32.
    # The denominator SURVEY 8(d) fixes for an fp32-parity result is the dense TF32 peak = 1/2 of the measured bf16 peak
    # (B200_PROFILING.md); the fraction of the 16-bit peak (what the fp16 MMAs could reach) is reported beside it.
    tf32_peak = 0.5 * pk["bf16_tflops"]
    # Timed unit of the roofline (one CUDA-event pair inside the library, live in the timed region):
    #  * api train_step: the forward kernel, the backward dH chain and the dW GEMM are programmatic dependent launches
    #    of each other and overlap (tiles handed over through flags), so the TRIO is the unit: 3 x F_fwd algorithmic
    #    FLOPs (forward, dX = dZ W, dW = h^T dZ -- one forward's worth each) per trio;
    #  * api autograd: forward alone (F_fwd) and the overlapped backward pair chain + dW (2 x F_fwd).
    # An event between overlapped kernels would serialise them; per-kernel times are in the ncu launch list (profiles/).
    tf = lambda flops, ms: (flops / (ms * 1e-3) / 1e12) if ms > 0 else None
    fwd_avg_ms = fwd_ms / max(fwd_n, 1)
    bwd_avg_ms = bwd_ms / max(bwd_n, 1)
    trio_avg_ms = trio_ms / max(trio_n, 1)
    if trio_n:
        unit_name = ("sage_fwd_tc_kernel + sage_bwd_tc_kernel + sage_dw_tc_kernel (forward, backward dH chain, dW GEMM; "
                     "overlapped by programmatic dependent launch: one event pair)")
        unit_flops, unit_ms, kernels_ms = 3 * f_fwd, trio_avg_ms, trio_ms
        parts = None
    else:
        unit_name = "sage_bwd_tc_kernel + sage_dw_tc_kernel (backward dH chain and dW GEMM, overlapped: one event pair)"
        unit_flops, unit_ms, kernels_ms = 2 * f_fwd, bwd_avg_ms, bwd_ms + fwd_ms
        parts = {"avg_launch_ms": fwd_avg_ms, "flops_per_launch": f_fwd, "achieved": tf(f_fwd, fwd_avg_ms),
                 "frac": (tf(f_fwd, fwd_avg_ms) or 0) / tf32_peak}
    achieved = tf(unit_flops, unit_ms)
    # ncu --set full (same workload): dram read + write per launch, summed over the kernels of the unit, and the tensor-pipe
    # activity per kernel -- READ from the newest committed capture under profiles/ (never measured under the profiler here)
    ncu = read_ncu_summary() if (args.workload == "play" and args.path != "simt") else None
    traffic = None
    if ncu:
        names = ("sage_fwd_tc_kernel", "sage_bwd_tc_kernel", "sage_dw_tc_kernel") if trio_n else ("sage_bwd_tc_kernel", "sage_dw_tc_kernel")
        if all(k in ncu["kernels"] for k in names):
            traffic = sum(ncu["kernels"][k]["dram_bytes"] for k in names)
    line = {
        "metric": METRIC, "value": graphs * args.steps / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(workload_config(args, batch, world), collective=collective),
        "clocks": clocks,
        "e2e": {"value": graphs * args.steps / (ms_e2e * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": int(hb._packed.numel()) if getattr(hb, "_packed", None) is not None else
                int(sum(t.numel() * t.element_size() for t in (hb.x, hb.edge_index, hb.batch, hb.ptr))),
                "d2h_bytes_per_step": int(out_host.numel()) * 4, "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "kernel": unit_name,
                     "achieved": achieved, "peak": tf32_peak, "unit": "TFLOP/s",
                     "frac": (achieved / tf32_peak) if achieved else None, "traffic": traffic,
                     "traffic_source": ("ncu dram__bytes_read+write per launch, summed over the unit's kernels, " + ncu["file"]) if ncu else None,
                     "tensor_pipe_active_pct": ({k: v["tensor_pct"] for k, v in ncu["kernels"].items() if k.startswith("sage_")}
                                                if ncu else None),
                     "peak_source": f"0.5 x bf16_tflops ({pk['source']}) = dense TF32",
                     "arithmetic": ("fp32-parity 3-product emulation: forward + dH chain = 3 fp16 MMAs (kind::f16, K = 16) per "
                                    "algorithmic product = 1.5 TF32-pass equivalents (ceiling 2/3 of this peak), "
                                    "dW GEMM = 3xTF32 (ceiling 1/3)"),
                     "peak_f16": pk["bf16_tflops"],
                     "frac_of_f16_peak": (achieved / pk["bf16_tflops"]) if achieved else None,
                     "flops_per_launch": unit_flops, "avg_launch_ms": unit_ms,
                     "step_frac": tf(f_train, ms_total / args.steps) / tf32_peak,
                     "kernel_share_of_step": kernels_ms / ms_total},
    }
    if parts:
        line["roofline"]["fwd_kernel"] = parts
    # SURVEY 8(d) algorithmic bytes of the CSR / feature stream: 16 B per edge (int64 edge_index) + 8 B per node (batch) +
    # 8 (B+1) (ptr) read, 4 B per edge (col) + 4 B per node (rowptr) + 4 (B+1) (gptr) written
    n_nodes, n_edges = int(batch.x.shape[0]), int(batch.edge_index.shape[1])
    csr_bytes = 16 * n_edges + 8 * n_nodes + 8 * (batch.num_graphs + 1) + 4 * n_edges + 4 * (n_nodes + 1) + 4 * (batch.num_graphs + 1)
    line["roofline"]["csr_phase"] = {
        "bound": "hbm", "kernel": "csr_build_fused_kernel (hexgnn_build_csr: init, histogram, scan, fill, rank + tile packing in one cooperative launch)",
        "bytes_per_launch": csr_bytes, "avg_launch_ms": csr_ms, "achieved": csr_bytes / (csr_ms * 1e-3) / 1e9,
        "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": csr_bytes / (csr_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
        "note": "2.3 MB of integer work in five dependent phases (grid barriers, atomics, a serial tile-packing chain): latency bound, not bandwidth bound"}
    line["config"]["api"] = ("model.train_step (hexgnn_train_step)" if fused_api
                             else "model.forward -> ops.td_loss -> loss.backward()")
    if not args.no_cpu_baseline and world == 1:
        v, done, cores, dt = time_cpu_port(batch, 1000, 2, max_seconds=args.cpu_seconds)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{done} fwd+bwd steps of the same {batch.num_graphs}-graph batch "
                                          f"({dt:.1f} s), torch fp32 restatement of the PyG path"}
        try:   # the same restatement on this GPU's stock torch kernels (not the product, not the target)
            gv, gdone, _, gdt = time_cpu_port(batch, 50, 5, max_seconds=10.0, device=dev)
            line["cpu_baseline"]["same_port_on_this_gpu_stock_kernels"] = {
                "value": gv, "unit": UNIT, "ms_per_step": 1e3 * gdt / gdone, "steps": gdone}
        except Exception as e:  # noqa: BLE001 -- a baseline leg must not take the bench line down
            line["cpu_baseline"]["same_port_on_this_gpu_stock_kernels"] = {"error": repr(e)[:200]}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


_JSON_FD = None


def protect_stdout():
    """The contract is ONE JSON line on stdout.  Native libraries write there too (NCCL prints its version line when the
    box exports NCCL_DEBUG), so fd 1 is pointed at stderr for the whole run and the line goes to the saved descriptor."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under torch.distributed.run
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29531"),
               os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    protect_stdout()
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_fused(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
