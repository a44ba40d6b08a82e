#!/usr/bin/env python
"""DynEdge events/s on B200 (BASELINE.json metric), next to the host-CPU reference path.

    python bench.py --gpus 1 --steps K --warmup W            # this repo, N=1
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...                      # CPU reference arm

One "step" = one DynEdge inference pass (StandardModel: DynEdge backbone +
EnergyReconstruction head) over one collated batch of synthetic IceCube-86
events (BASELINE.json configs[1]: 1024 events/GPU, ~60 pulses/event, k=8).
Events are independent, so with N GPUs every rank processes its own batch
(weak scaling, no data-path collective).

Printed JSON (rank 0, one line): see the bench contract in the task statement;
`value` = device-timed events/s with inputs resident in HBM, `e2e` = the same
through the public API with host (pinned) inputs and the prediction read back.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "DynEdge inference events/sec"
TRAIN_METRIC = "DynEdge training events/sec"
UNIT = "events/s"
POOLS = ["min", "max", "mean", "sum"]


_REAL_STDOUT = None


def claim_stdout():
    """Keep stdout for the ONE JSON line: whatever else writes to file descriptor 1 (NCCL's
    version banner, library chatter) goes to stderr from here on."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT if _REAL_STDOUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=None, choices=[None, "fp32", "bf16"])
    ap.add_argument("--events", type=int, default=1024, help="events per GPU per step")
    ap.add_argument("--pulses", type=float, default=60.0)
    ap.add_argument("--cpu-events", type=int, default=128,
                    help="events per CPU-baseline step (reference distribute step is O(N*B))")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-riders", action="store_true",
                    help="skip the fp32 / C4 / C5 / sharded-stream riders of the default line")
    ap.add_argument("--workload", default="c2", choices=["c2", "c4", "c5"],
                    help="c2: BASELINE configs[1] (default); c4: configs[3] shapes (IceCube-Upgrade "
                         "geometry, F=14, k=16, pools min/max/mean; use --events 256 --pulses 500); "
                         "c5: one point of configs[4] (direction + energy heads, exactly --pulses "
                         "pulses in every event; tools/sweep_c5.sh walks 16..2048)")
    ap.add_argument("--mode", default="infer", choices=["infer", "train"],
                    help="infer: BASELINE configs[1] (default); train: configs[2] (bf16 training "
                         "step: forward, backward, gradient all-reduce, Adam)")
    return ap.parse_args()


def workload_cfg(args):
    """(detector, k, pools) of the selected workload."""
    if getattr(args, "workload", "c2") == "c4":
        return "upgrade", 16, ["min", "max", "mean"]
    return "icecube86", 8, POOLS


def workload_name(args):
    if getattr(args, "workload", "c2") == "c4":
        return (f"C4: DynEdge inference, synthetic IceCube-Upgrade pulse maps, ~{args.pulses:g} "
                f"pulses/event, F=14, k=16, batch {args.events}/GPU, pools min/max/mean")
    if getattr(args, "workload", "c2") == "c5":
        return (f"C5: DynEdge direction + energy multi-task inference, synthetic IceCube-86 pulse "
                f"maps, exactly {int(args.pulses)} pulses/event, k=8, batch {args.events}/GPU, pools "
                f"min/max/mean/sum")
    if getattr(args, "mode", "infer") == "train":
        return (f"C3: DynEdge training step (forward, backward, gradient all-reduce, Adam), bf16, "
                f"synthetic IceCube-86 events, ~{args.pulses:g} pulses/event, k=8, batch "
                f"{args.events}/GPU, LogCosh on log10-energy targets ~U(0,4)")
    return (f"C2: DynEdge inference, synthetic IceCube-86 pulse maps, ~{args.pulses:g} pulses/event, "
            f"k=8, batch {args.events}/GPU, pools min/max/mean/sum, EnergyReconstruction head")


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"],
                "bf16_tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0,
            "source": "fallback"}


# --------------------------------------------------------------------------- #
# CPU reference arm (the oracle port of the reference's CPU PyG path)
# --------------------------------------------------------------------------- #
def cpu_reference_events_per_s(num_events, pulses, steps, warmup, seed=1234):
    import oracle
    from oracle.ops import knn_graph_padded
    from oracle import tasks as otasks
    from graphnet_b200.synthetic import make_pulse_batch

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    pb = make_pulse_batch(num_events, pulses, seed=seed)
    ei = knn_graph_padded(pb.x[:, :3], 8, pb.batch)  # dataset-side graph, not timed
    spec = oracle.DynEdgeSpec(nb_inputs=7, global_pooling_schemes=POOLS)
    sd = oracle.init_dynedge_state_dict(spec, seed=0)
    torch.manual_seed(1)
    head = torch.nn.Linear(128, 1)

    # The reference's dynamic kNN runs in torch_cluster (C++); the vectorised padded
    # restatement is the closest CPU equivalent, so it replaces the per-event loop here.
    import oracle.ops as oops
    loop_knn = oops.knn_graph
    oops.knn_graph = lambda x, k, batch=None, loop=False: knn_graph_padded(x, k, batch)
    try:
        def step():
            with torch.no_grad():
                y = oracle.dynedge_forward(spec, sd, pb.x, ei, pb.batch, pb.n_pulses,
                                           dense_distribute=True)
                return otasks.task_forward("EnergyReconstruction", head.weight, head.bias, y)
        for _ in range(warmup):
            step()
        times = []
        for _ in range(steps):
            t0 = time.perf_counter()
            step()
            times.append(time.perf_counter() - t0)
    finally:
        oops.knn_graph = loop_knn
    total = sum(times)
    return num_events * steps / total, cores, total / steps


def cpu_reference_train_events_per_s(num_events, pulses, steps, warmup, seed=1234):
    """Oracle port of the reference's CPU training step (torch autograd + Adam)."""
    import oracle
    from oracle.ops import knn_graph_padded
    from oracle import tasks as otasks
    from graphnet_b200.synthetic import make_pulse_batch
    import oracle.ops as oops

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    pb = make_pulse_batch(num_events, pulses, seed=seed)
    ei = knn_graph_padded(pb.x[:, :3], 8, pb.batch)
    spec = oracle.DynEdgeSpec(nb_inputs=7, global_pooling_schemes=POOLS)
    sd = {k: v.clone().requires_grad_(True) for k, v in
          oracle.init_dynedge_state_dict(spec, seed=0).items()}
    torch.manual_seed(1)
    head = torch.nn.Linear(128, 1)
    target = torch.rand(num_events, 1, generator=torch.Generator().manual_seed(7)) * 4
    opt = torch.optim.Adam(list(sd.values()) + list(head.parameters()), lr=1e-3, eps=1e-3)
    loop_knn = oops.knn_graph
    oops.knn_graph = lambda x, k, batch=None, loop=False: knn_graph_padded(x, k, batch)
    try:
        def step():
            opt.zero_grad(set_to_none=True)
            y = oracle.dynedge_forward(spec, sd, pb.x, ei, pb.batch, pb.n_pulses,
                                       dense_distribute=True)
            pred = otasks.task_forward("EnergyReconstruction", head.weight, head.bias, y)
            d = pred - target
            loss = (d + torch.nn.functional.softplus(-2.0 * d) - 0.6931471805599453).mean()
            loss.backward()
            opt.step()
        for _ in range(warmup):
            step()
        times = []
        for _ in range(steps):
            t0 = time.perf_counter()
            step()
            times.append(time.perf_counter() - t0)
    finally:
        oops.knn_graph = loop_knn
    total = sum(times)
    return num_events * steps / total, cores, total / steps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, args.steps), args.warmup
    # bound the run: ~0.3 s per 128-event step on 8 cores
    steps = min(steps, 40)
    warmup = min(warmup, 3)
    train = args.mode == "train"
    fn = cpu_reference_train_events_per_s if train else cpu_reference_events_per_s
    if train:
        steps = min(steps, 15)
    v, cores, sec = fn(args.cpu_events, args.pulses, steps, warmup)
    sample = (f"{args.cpu_events}-event batch of the same workload per step (reference distribute "
              f"step allocates N*B*G floats), {steps} steps, oracle port of the reference's CPU path "
              f"incl. its dense one-hot distribute")
    line = {
        "impl": "reference", "metric": TRAIN_METRIC if train else METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "cpu_events_per_step": args.cpu_events},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


# --------------------------------------------------------------------------- #
# clocks sampler
# --------------------------------------------------------------------------- #
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100", "-i", str(self.index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                smax = float(r[1])
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": smax,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------- #
# our arm
# --------------------------------------------------------------------------- #
def algorithmic_flops_per_event_batch(N, E, F=7):
    """Unfactored FLOPs the reference executes."""
    d0 = 2 * F + 5
    conv = 2 * E * (2 * d0 * 128 + 128 * 256) + 3 * 2 * E * (2 * 256 * 336 + 336 * 256)
    post = 2 * N * ((d0 + 1024) * 336 + 336 * 256)
    return conv + post



# --------------------------------------------------------------------------- #
# riders: the other BASELINE configurations, a few steps each, in the same JSON line
# --------------------------------------------------------------------------- #
def _timed_steps(fn, steps, warmup, dev, world, flush_buf):
    """`steps` calls of fn, CUDA events per step, L2 flushed in between; ms = max over ranks."""
    import torch.distributed as dist
    for _ in range(warmup):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    for i in range(steps):
        flush_buf.fill_(i & 0xFF)
        starts[i].record()
        fn()
        ends[i].record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ms = sum(a.elapsed_time(b) for a, b in zip(starts, ends))
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def rider_infer(dev, world, rank, flush_buf, *, workload, events, pulses, precision, steps=5,
                warmup=3):
    """Inference events/s of one BASELINE configuration (inputs resident, every rank its own
    batch, value = all ranks' events / max-over-ranks time)."""
    import graphnet_b200 as g
    from graphnet_b200.synthetic import make_pulse_batch
    ns = argparse.Namespace(workload=workload, events=events, pulses=pulses, mode="infer")
    det, knn_k, pools = workload_cfg(ns)
    pb = make_pulse_batch(events, pulses, detector=det, seed=1234 + rank,
                          fixed_pulses=int(pulses) if workload == "c5" else None)
    x_dev, batch_dev, npulse_dev = pb.x.to(dev), pb.batch.to(dev), pb.n_pulses.to(dev)
    ei_dev = g.knn_graph(x_dev[:, :3].contiguous(), knn_k, batch_dev, num_events=events)
    N, E = x_dev.shape[0], ei_dev.shape[1]
    torch.manual_seed(0)
    backbone = g.DynEdge(x_dev.shape[1], nb_neighbours=knn_k, global_pooling_schemes=pools,
                         precision=precision)
    tasks = [g.EnergyReconstruction(hidden_size=backbone.nb_outputs, target_labels="energy",
                                    loss_function=g.LogCoshLoss())]
    if workload == "c5":
        tasks.insert(0, g.DirectionReconstructionWithKappa(
            hidden_size=backbone.nb_outputs, target_labels="direction",
            loss_function=g.VonMisesFisher3DLoss()))
    model = g.StandardModel(tasks=tasks, graph_definition=object(), gnn=backbone).to(dev)
    model.inference()
    model.train(False)
    backbone.validate_inputs = False
    data_dev = g.Data(x=x_dev, edge_index=ei_dev, batch=batch_dev, n_pulses=npulse_dev)
    backbone.check_inputs(data_dev)
    ms = _timed_steps(lambda: model(data_dev)[0], steps, warmup, dev, world, flush_buf)
    out = {"workload": workload_name(ns), "value": events * world * steps / (ms * 1e-3),
           "unit": UNIT, "n_gpus": world, "ms_per_step": ms / steps, "steps": steps,
           "warmup": warmup, "dtype": "bf16" if precision == "bf16" else
           "f32 (tensor cores: 16-bit operand pieces summing to the fp32 operands, fp32 accumulate)",
           "nodes_per_batch": N, "edges_per_batch": E,
           "nodes_per_s": N * world * steps / (ms * 1e-3)}
    if precision == "fp32" and workload == "c2":
        roof = dominant_kernel_roofline(backbone, data_dev, "fp32", N, E, dev, flush_buf)
        r = roof(peaks())
        out["edge_kernel"] = {k_: r.get(k_) for k_ in ("kernel", "ms_per_launch", "achieved",
                                                       "unit", "error") if k_ in r}
    return out


def rider_sharded(dev, world, rank, flush_buf, *, events_per_gpu=1024, steps=5, warmup=3):
    """One RAGGED event stream (every rank holds the same host batch), split over the ranks by
    `sharding.predict_sharded(balance="pulses")`: each rank copies and runs its contiguous share,
    the per-event outputs come back with one NCCL all_gather.  Timed per step on the device:
    H2D of the share + forward + all_gather; value = all events / max-over-ranks time."""
    import numpy as np
    import graphnet_b200 as g
    from graphnet_b200.sharding import predict_sharded, shard_ranges
    from graphnet_b200.synthetic import make_pulse_batch
    B = events_per_gpu * world
    rng = np.random.default_rng(99)
    counts = np.exp(rng.uniform(np.log(8.0), np.log(400.0), size=B)).astype(np.int64)  # ~8..400
    pb = make_pulse_batch(B, seed=4321, counts=counts)       # the same stream on every rank
    host = g.Data(x=pb.x.pin_memory(), batch=pb.batch.pin_memory(),
                  n_pulses=pb.n_pulses.pin_memory())
    torch.manual_seed(0)
    backbone = g.DynEdge(7, global_pooling_schemes=POOLS, precision="bf16")
    task = g.EnergyReconstruction(hidden_size=backbone.nb_outputs, target_labels="energy",
                                  loss_function=g.LogCoshLoss())
    model = g.StandardModel(tasks=[task], graph_definition=object(), gnn=backbone).to(dev)
    model.inference()
    model.train(False)
    backbone.validate_inputs = False
    outs = []

    def step():
        outs[:] = [predict_sharded(model, host, device=dev, balance="pulses")]
    ms = _timed_steps(step, steps, warmup, dev, world, flush_buf)
    ranges = shard_ranges(pb.n_pulses, world, "pulses")
    share = [int(pb.n_pulses[lo:hi].sum()) for lo, hi in ranges]
    assert outs[0].shape[0] == B, outs[0].shape
    return {"workload": f"one ragged stream of {B} events (8..400 pulses, log-uniform), "
                        f"sharding.predict_sharded(balance='pulses'), input graph built on device",
            "value": B * steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "ms_per_step": ms / steps, "steps": steps, "events": B, "nodes": int(pb.x.shape[0]),
            "events_per_rank": [hi - lo for lo, hi in ranges], "pulses_per_rank": share,
            "collective": "all_gather of [B_local, 1] outputs (NCCL)" if world > 1 else "none",
            "h2d_bytes_per_step_rank0": int(share[0] * (7 * 4 + 8) + (ranges[0][1] - ranges[0][0]) * 4)}


RIDERS = (
    ("fp32", dict(workload="c2", events=1024, pulses=60.0, precision="fp32")),
    ("c4", dict(workload="c4", events=256, pulses=500.0, precision="bf16")),
    ("c5_n16", dict(workload="c5", events=65536, pulses=16, precision="bf16")),
    ("c5_n256", dict(workload="c5", events=4096, pulses=256, precision="bf16")),
    ("c5_n2048", dict(workload="c5", events=512, pulses=2048, precision="bf16")),
)


def run_riders(dev, world, rank, flush_buf):
    out = {}
    for key, kw in RIDERS:
        try:
            out[key] = rider_infer(dev, world, rank, flush_buf, **kw)
        except Exception as exc:  # noqa: BLE001 - a rider never costs the main line
            out[key] = {"error": repr(exc)}
        torch.cuda.synchronize(dev)
        torch.cuda.empty_cache()
    try:
        out["sharded"] = rider_sharded(dev, world, rank, flush_buf)
    except Exception as exc:  # noqa: BLE001
        out["sharded"] = {"error": repr(exc)}
    torch.cuda.empty_cache()
    return out


def run_ours(args):
    import torch.distributed as dist
    import graphnet_b200 as g
    from graphnet_b200 import _lib
    from graphnet_b200.synthetic import make_pulse_batch
    import oracle  # only for the dataset-side initial graph (not timed) and cpu_baseline

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    precision = args.precision
    if precision is None:
        try:
            import graphnet_b200.dynedge_bf16  # noqa: F401
            precision = "bf16"
        except ImportError:
            precision = "fp32"
    lib = _lib.load()

    # ---- inputs: one synthetic batch per rank, host (pinned) and device copies
    det, knn_k, pools = workload_cfg(args)
    pb = make_pulse_batch(args.events, args.pulses, detector=det, seed=1234 + rank,
                          fixed_pulses=int(args.pulses) if args.workload == "c5" else None)
    x_dev = pb.x.to(dev)
    batch_dev = pb.batch.to(dev)
    npulse_dev = pb.n_pulses.to(dev)
    # the initial graph is built by the dataset in the reference (KNNGraph in the
    # dataloader workers), not by the model: build it once here, on the device
    ei_dev = g.knn_graph(x_dev[:, :3].contiguous(), knn_k, batch_dev, num_events=args.events)
    N, E = x_dev.shape[0], ei_dev.shape[1]
    host = {"x": pb.x.pin_memory(), "edge_index": ei_dev.cpu().pin_memory(),
            "batch": pb.batch.pin_memory(), "n_pulses": pb.n_pulses.pin_memory()}

    torch.manual_seed(0)
    backbone = g.DynEdge(x_dev.shape[1], nb_neighbours=knn_k, global_pooling_schemes=pools,
                         precision=precision)
    task = g.EnergyReconstruction(hidden_size=backbone.nb_outputs, target_labels="energy",
                                  loss_function=g.LogCoshLoss())
    tasks = [task]
    if args.workload == "c5":   # BASELINE configs[4]: two heads on one backbone pass
        tasks = [g.DirectionReconstructionWithKappa(hidden_size=backbone.nb_outputs,
                                                    target_labels="direction",
                                                    loss_function=g.VonMisesFisher3DLoss()), task]
    model = g.StandardModel(tasks=tasks, graph_definition=object(), gnn=backbone).to(dev)
    model.inference()
    model.train(False)
    backbone.validate_inputs = False  # inputs are validated once below; keeps steps async
    data_dev = g.Data(x=x_dev, edge_index=ei_dev, batch=batch_dev, n_pulses=npulse_dev)
    backbone.check_inputs(data_dev)

    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step_resident():
        return model(data_dev)[0]

    # end to end: every step copies its own inputs from pinned host memory and reads its
    # prediction back; the copy of step i+1 is issued before step i's result is awaited
    # (graphnet_b200.data.DevicePrefetcher - what a pinned-memory dataloader does)
    from graphnet_b200.data import DevicePrefetcher
    import itertools
    feeder = DevicePrefetcher(itertools.repeat(g.Data(**host)), dev)

    # ... and the read of step i's prediction is awaited after step i+1 has been enqueued
    # (graphnet_b200.data.ResultReader; StandardModel.predict_to_host is this loop): every step's
    # result is copied to pinned host memory inside the timed region, the device does not idle
    # for the host's launch time after each read
    from graphnet_b200.data import ResultReader
    reader = ResultReader()

    def step_e2e():
        return reader.push(model(next(feeder)))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        launches0 = lib.gn_launch_count()
        for i in range(steps):
            flush_buf.fill_(i & 0xFF)  # L2 flush between timed iterations (not timed)
            starts[i].record()
            fn()
            ends[i].record()
        barrier()
        launches = lib.gn_launch_count() - launches0
        ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_res, launches = timed(step_resident, args.steps, max(3, args.warmup))
    ms_e2e, _ = timed(step_e2e, args.steps, max(3, args.warmup))
    # the same end-to-end step when the host sends pulses only (x, batch, n_pulses: what
    # graphnet_b200.io.SQLitePulseSource produces) and the input graph is built on the device
    host_dg = {k_: v for k_, v in host.items() if k_ != "edge_index"}
    feeder_dg = DevicePrefetcher(itertools.repeat(g.Data(**host_dg)), dev)
    backbone.input_graph_neighbours = knn_k

    reader.flush()

    def step_e2e_dg():
        return reader.push(model(next(feeder_dg)))
    ms_e2e_dg, _ = timed(step_e2e_dg, args.steps, max(3, args.warmup))
    reader.flush()
    del feeder_dg
    clocks = sampler.stop() if rank == 0 else None

    # the same step replayed from a CUDA graph (fixed batch shape: graphnet_b200.GraphedInference)
    ms_graph = None
    try:
        runner = g.GraphedInference(model, data_dev)
        ms_graph, _ = timed(lambda: runner.replay()[0], args.steps, max(3, args.warmup))
        del runner
    except Exception as exc:  # noqa: BLE001 - an extra, never worth the bench line
        print(f"graph replay skipped: {exc!r}", file=sys.stderr)

    # ---- roofline of the dominant kernel (conv layers 2-4 fused edge kernel), timed alone
    roof = dominant_kernel_roofline(backbone, data_dev, precision, N, E, dev, flush_buf)
    roof_hbm = pooling_roofline(batch_dev, args.events, N, pools, dev, flush_buf)

    # BASELINE's metric is "inference+train": a short measurement of the training step (configs[2];
    # `--mode train` prints it as the main line) rides along; it can never cost the line above
    train = None
    try:
        if args.workload != "c2":
            raise RuntimeError("training summary only rides along the C2 workload")
        del model, backbone, feeder
        torch.cuda.synchronize(dev)
        torch.cuda.empty_cache()   # same allocator state as a stand-alone `--mode train` run
        targs = argparse.Namespace(**vars(args))
        targs.mode = "train"
        train = train_measure(targs, min(args.steps, 10), 12)
    except Exception as exc:  # noqa: BLE001
        train = {"error": repr(exc)}

    riders = None
    if args.workload == "c2" and not args.no_riders:
        riders = run_riders(dev, world, rank, flush_buf)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    pk = peaks()
    events_total = args.events * world
    value = events_total * args.steps / (ms_res * 1e-3)
    e2e_value = events_total * args.steps / (ms_e2e * 1e-3)
    h2d = sum(t.numel() * t.element_size() for t in host.values())
    h2d_dg = sum(t.numel() * t.element_size() for k_, t in host.items() if k_ != "edge_index")
    d2h = args.events * 4 * (5 if args.workload == "c5" else 1)   # [B,1] energy (+ [B,4] direction)
    flops = algorithmic_flops_per_event_batch(N, E, F=int(x_dev.shape[1]))
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms_res / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if precision == "bf16" else "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "nodes_per_batch": N, "edges_per_batch": E,
                   "l2": "256 MiB buffer written between timed steps (L2 flush)",
                   "parallelism": f"events sharded, {world} independent rank(s)",
                   "algorithmic_tflop_per_step": flops / 1e12,
                   "model_tflops_unfactored": flops * world * args.steps / (ms_res * 1e-3) / 1e12},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / args.steps,
                "note": "StandardModel call on host batches: inputs copied from pinned memory every "
                        "step (copy of step i+1 under step i's kernels), every step's prediction "
                        "copied to pinned host memory, awaited one step behind the launches"},
        "e2e_device_graph": {
            "value": events_total * args.steps / (ms_e2e_dg * 1e-3), "unit": UNIT,
            "h2d_bytes_per_step": h2d_dg, "d2h_bytes_per_step": d2h,
            "ms_per_step": ms_e2e_dg / args.steps,
            "note": "host sends x, batch, n_pulses only; the input kNN graph is built on the device "
                    "(no edge_index in the batch)"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roof(pk) if roof else None,
        "roofline_hbm": roof_hbm(pk) if roof_hbm else None,
    }
    if ms_graph is not None:
        line["graph_replay"] = {"value": events_total * args.steps / (ms_graph * 1e-3), "unit": UNIT,
                                "ms_per_step": ms_graph / args.steps,
                                "note": "same step, inputs resident, replayed from one CUDA graph"}
    if rank == 0 and train is not None and args.workload == "c2":
        line["train"] = ({k: train[k] for k in ("metric", "value", "unit", "ms_per_step", "dtype",
                                                "steps", "gpu_launches")}
                         | {"e2e": train["e2e"], "workload": train["config"]["workload"]}
                         if "error" not in train else train)
    if riders:
        # BASELINE configs beside the headline one: C2 in fp32 (the reference's own arithmetic),
        # C4, three points of the C5 sweep, and one event stream sharded over the ranks
        c5 = {k_[3:]: v_ for k_, v_ in riders.items() if k_.startswith("c5_")}
        line["fp32"] = riders.get("fp32")
        line["c4"] = riders.get("c4")
        line["c5"] = c5
        line["sharded"] = riders.get("sharded")
    if world == 1 and not args.no_cpu_baseline and args.workload == "c2":
        v, cores, sec = cpu_reference_events_per_s(args.cpu_events, args.pulses, steps=10, warmup=2)
        line["cpu_baseline"] = {
            "value": v, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"10 steps of a {args.cpu_events}-event batch of the same workload "
                      f"({sec * 1e3:.0f} ms/step), oracle port incl. the reference's dense distribute"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_ours_train(args):
    """BASELINE configs[2]: one step = forward + backward + gradient all-reduce + Adam on one
    collated batch per GPU (data parallel, weak scaling; the all-reduce is the one exchange)."""
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    line = train_measure(args, args.steps, max(3, args.warmup), with_clocks=True)
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sec = cpu_reference_train_events_per_s(args.cpu_events, args.pulses, steps=5,
                                                             warmup=1)
            line["cpu_baseline"] = {
                "value": v, "unit": UNIT, "cores": cores, "kind": "port",
                "sample": f"5 training steps of a {args.cpu_events}-event batch of the same "
                          f"workload ({sec * 1e3:.0f} ms/step), oracle port (torch autograd + "
                          f"Adam on CPU)"}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def train_measure(args, steps, warmup, with_clocks=False):
    """Times `steps` training steps on every rank (process group, if any, already up); returns
    the JSON object on rank 0 (None elsewhere)."""
    import torch.distributed as dist
    import graphnet_b200 as g
    from graphnet_b200 import _lib
    from graphnet_b200.sharding import allreduce_gradients
    from graphnet_b200.synthetic import make_pulse_batch

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dev = torch.device("cuda", local_rank)
    lib = _lib.load()

    pb = make_pulse_batch(args.events, args.pulses, seed=1234 + rank)
    x_dev, batch_dev, npulse_dev = pb.x.to(dev), pb.batch.to(dev), pb.n_pulses.to(dev)
    ei_dev = g.knn_graph(x_dev[:, :3].contiguous(), 8, batch_dev, num_events=args.events)
    target = torch.rand(args.events, generator=torch.Generator().manual_seed(7 + rank)) * 4
    host = {"x": pb.x.pin_memory(), "edge_index": ei_dev.cpu().pin_memory(),
            "batch": pb.batch.pin_memory(), "n_pulses": pb.n_pulses.pin_memory(),
            "energy": target.pin_memory()}
    N, E = x_dev.shape[0], ei_dev.shape[1]

    torch.manual_seed(0)   # same initial weights on every rank (DDP broadcasts them; same seed here)
    backbone = g.DynEdge(7, global_pooling_schemes=POOLS, precision="bf16")
    task = g.EnergyReconstruction(hidden_size=backbone.nb_outputs, target_labels="energy",
                                  loss_function=g.LogCoshLoss())
    model = g.StandardModel(tasks=[task], graph_definition=object(), gnn=backbone,
                            optimizer_kwargs={"lr": 1e-3, "eps": 1e-3}).to(dev)
    model.train()
    backbone.validate_inputs = False
    params = [p for p in model.parameters() if p.requires_grad]
    opt = model.configure_optimizers()["optimizer"]
    data_dev = g.Data(x=x_dev, edge_index=ei_dev, batch=batch_dev, n_pulses=npulse_dev,
                      energy=target.to(dev))
    backbone.check_inputs(data_dev)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    grad_bytes = [0]

    def train_step(d):
        opt.zero_grad(set_to_none=True)
        loss = model.training_step(d, 0)
        loss.backward()
        grad_bytes[0] = allreduce_gradients(params)
        opt.step()
        return loss

    def step_resident():
        return train_step(data_dev)

    from graphnet_b200.data import DevicePrefetcher
    import itertools
    feeder = DevicePrefetcher(itertools.repeat(g.Data(**host)), dev)

    def step_e2e():
        return float(train_step(next(feeder)).item())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        launches0 = lib.gn_launch_count()
        for i in range(steps):
            flush_buf.fill_(i & 0xFF)
            starts[i].record()
            fn()
            ends[i].record()
        barrier()
        launches = lib.gn_launch_count() - launches0
        ms = sum(a.elapsed_time(b) for a, b in zip(starts, ends))
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    sampler = ClockSampler(local_rank) if with_clocks else None
    if rank == 0 and sampler:
        sampler.start()
    ms_res, launches = timed(step_resident, steps, warmup)
    ms_e2e, _ = timed(step_e2e, steps, warmup)
    clocks = sampler.stop() if (rank == 0 and sampler) else None
    peak_mem = torch.cuda.max_memory_allocated(dev)
    if rank != 0:
        return None
    events_total = args.events * world
    h2d = sum(t.numel() * t.element_size() for t in host.values())
    # executed tensor work per step (factored forward; backward = 3 slot GEMMs per EdgeConv +
    # two GEMMs per Linear), for orientation only
    line = {
        "metric": TRAIN_METRIC, "value": events_total * steps / (ms_res * 1e-3), "unit": UNIT,
        "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": ms_res / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(args), "nodes_per_batch": N, "edges_per_batch": E,
                   "l2": "256 MiB buffer written between timed steps (L2 flush)",
                   "parallelism": f"data parallel, {world} rank(s), one flat NCCL all-reduce of "
                                  f"{grad_bytes[0]} gradient bytes per step",
                   "optimizer": "torch.optim.Adam(lr=1e-3, eps=1e-3) on fp32 master weights",
                   "peak_device_memory_mb": peak_mem / 1e6},
        "e2e": {"value": events_total * steps / (ms_e2e * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                "ms_per_step": ms_e2e / steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": None,
    }
    return line


def ncu_traffic_bytes(kernel_name):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the
    committed `ncu --set full` capture of this workload (profiles/r01_edge_kernel_ncu.json);
    None when the capture is of another kernel."""
    for fname in ("r02_edge_kernel_ncu.json", "r01_edge_kernel_ncu.json"):
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", fname)
        try:
            with open(path) as f:
                cap = json.load(f)
        except (OSError, ValueError):
            continue
        if cap.get("kernel_prefix") and cap["kernel_prefix"] in kernel_name:
            return cap.get("dram_bytes_per_launch")
    return None


def dominant_kernel_roofline(backbone, data, precision, N, E, dev, flush_buf):
    """Time the fused EdgeConv of layer 2 alone, CUDA events on the launch stream."""
    import graphnet_b200.ops as ops
    try:
        if precision == "bf16":
            from graphnet_b200.dynedge_bf16 import time_edge_kernel
            ms, executed_flops, name = time_edge_kernel(backbone, data, flush_buf)
        else:
            from graphnet_b200.dynedge_x3 import time_edge_kernel_f32
            ms, executed_flops, name = time_edge_kernel_f32(backbone, data, flush_buf)
    except Exception as exc:  # roofline is diagnostic; never lose the bench line
        return lambda pk: {"error": repr(exc)}
    algorithmic = 2 * E * (2 * 256 * 336 + 336 * 256)

    def fill(pk):
        peak = pk["bf16_tflops"]
        ach = executed_flops / (ms * 1e-3) / 1e12
        return {"bound": "tensor", "kernel": name, "achieved": ach, "peak": peak,
                "unit": "TFLOP/s", "frac": ach / peak, "traffic": ncu_traffic_bytes(name),
                "peak_source": f"{pk['source']} cuBLAS bf16 burst (kernel timed alone)",
                "ms_per_launch": ms, "executed_gflop_per_launch": executed_flops / 1e9,
                "algorithmic_gflop_per_launch": algorithmic / 1e9,
                "achieved_unfactored": algorithmic / (ms * 1e-3) / 1e12,
                "note": "achieved / frac count the FLOPs the kernel issues (factored form: the "
                        "first Linear is the node GEMM, this kernel runs 2*slots*Kp*b16 per node); "
                        "achieved_unfactored counts the reference's 2E(2d*a+a*b) per SURVEY.md 8(d) "
                        "and is not a utilisation"}
    return fill


def pooling_roofline(batch_dev, num_events, N, pools, dev, flush_buf):
    """The largest pure byte mover of the forward, `gn_segment_pool` (all schemes from one read of
    the [N, 256] fp32 node matrix), timed alone with a flushed L2 against the measured HBM peak.
    Algorithmic bytes: 4*N*256 in + 4*B*256*P out; DRAM traffic per launch from
    the committed counters (profiles/r01_hbm_kernels.md: 63.0 MB at C2)."""
    import graphnet_b200.ops as ops
    try:
        ptr, _ = ops.batch_to_ptr(batch_dev, num_events)
        h = torch.randn(N, 256, device=dev)
        out = torch.empty(num_events, 256 * len(pools), device=dev)
        for _ in range(3):
            ops.segment_pool(h, ptr, pools, out=out)
        reps = 10
        s = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]
        e = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]
        for i in range(reps):
            flush_buf.fill_(i)
            s[i].record()
            ops.segment_pool(h, ptr, pools, out=out)
            e[i].record()
        torch.cuda.synchronize(dev)
        ms = sum(a.elapsed_time(b) for a, b in zip(s, e)) / reps
    except Exception as exc:  # diagnostic; never lose the bench line
        return lambda pk: {"error": repr(exc)}
    nbytes = 4 * N * 256 + 4 * num_events * 256 * len(pools)

    def fill(pk):
        ach = nbytes / (ms * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": "gn::segment_pool_kernel<float,4> (min/max/mean/sum from one read)",
                "achieved": ach, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": ach / pk["hbm_gbs"],
                "traffic": None, "peak_source": f"{pk['source']} copy bandwidth",
                "ms_per_launch": ms, "algorithmic_mb_per_launch": nbytes / 1e6}
    return fill


if __name__ == "__main__":
    a = parse_args()
    claim_stdout()
    if a.impl == "reference":
        run_reference(a)
    elif a.mode == "train":
        run_ours_train(a)
    else:
        run_ours(a)
