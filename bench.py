#!/usr/bin/env python
"""bench.py - EnergyEquivGNN forward + stiffness throughput (lattices/s) on B200, BASELINE.json config 2:
1024 synthetic periodic lattices (30 nodes / 60 struts -> 120 directed edges each) per GPU, default hyper-parameters
(scripts/train_main.py:25-52), fp32, random-init weights (torch.manual_seed(0)).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--lattices 1024]

One JSON line on stdout (rank 0).  `--impl reference` times the CPU restatement of the reference (oracle/) on the
host cores on a bounded sample of the same workload (the reference's own stack - e3nn/torch_scatter - is not
installable on this box).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "lattices/sec fwd+stiffness"
UNIT = "lattices/s"
N_NODES = 30
CPU_SAMPLE = 16  # lattices per CPU-baseline step (BASELINE.json config 1)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_reference_run(steps: int, warmup: int, seed0: int):
    """Oracle forward on CPU_SAMPLE lattices, all host threads. Returns (lattices/s, ms/step, threads)."""
    from energy_equiv_lattice_gnn_b200.data import synthetic_batch
    from tests.util import make_pair
    batch, rmax = synthetic_batch(range(seed0, seed0 + CPU_SAMPLE), N_NODES)
    oracle, _ = make_pair(rmax, dtype=torch.float32)
    oracle.eval()
    with torch.no_grad():
        for _ in range(warmup):
            oracle(batch)
        ts = []
        for _ in range(steps):
            t0 = time.perf_counter()
            oracle(batch)
            ts.append(time.perf_counter() - t0)
    ms = 1e3 * sum(ts) / len(ts)
    return CPU_SAMPLE / (ms / 1e3), ms, torch.get_num_threads()


class ClockSampler:
    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "20", "-i", str(gpu_index)],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name).read().strip().splitlines() if r.strip()]
        os.unlink(self.f.name)
        sm, reasons = [], set()
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                out["sm_max_mhz"] = float(r[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.strip().lower().startswith("active"):
                    reasons.add(name)
        if sm:
            sm.sort()
            out["sm_mhz"] = sm[len(sm) // 2]
        out["reasons"] = sorted(reasons)
        out["samples"] = len(sm)
        return out


def fp32_peak():
    """Measured fp32 FMA ceiling (TFLOP/s) of the B200: profiles/ffma2_rate.json, written by tools/micro/measure_fp32_peak.py."""
    p = os.path.join(ROOT, "profiles", "ffma2_rate.json")
    try:
        with open(p) as f:
            d = json.load(f)
        return float(d["fp32_tflops_peak"]), f"profiles/ffma2_rate.json ({d.get('source', '')}, {d.get('when', '')})"
    except (OSError, ValueError, KeyError):
        return 68.3, "fallback: round-1 measurement of tools/micro/ffma2_rate.cu (profiles/ffma2_rate.json missing)"


def algorithmic_model(model, n_nodes: int, n_edges: int):
    """Per-kernel ALGORITHMIC bytes / flops for one launch, keyed by timing tag.

    L1.tp_scatter follows section 8(d)'s message+scatter figure: read h' [N,D,C] + edge inputs (Y, sender, CSR) + write the
    message with interaction.linear folded ([N,D,C]) = 211 KB / 30-node lattice, + the streamed per-edge weights [E,P,C] =
    856 KB / lattice (`bytes`: the built variant streams the weights).  `bytes_fused` is the 211 KB figure, `bytes_as_built`
    adds what this implementation really moves (the un-folded M rows [N, 230, C])."""
    C, D = model.C, (model.lmax + 1) ** 2
    P1 = len(model._tp[-1]["paths"])
    rows1 = model._tp[-1]["rows_total"]
    cg1 = sum(len(v) for v in model._tp[-1]["cg"].values())
    ncoef = model._wmap["n_coef"]
    N, E = n_nodes, n_edges
    f4 = 4
    fused = f4 * (N * D * C + E * (D + 12) + (N + 1) + N * D * C)
    out = {
        "L1.tp_scatter": dict(bytes=fused + f4 * E * P1 * C, bytes_fused=fused,
                              bytes_as_built=f4 * (N * D * C + E * (P1 * C + D + 12) + N * rows1 * C + N + 1),
                              flops=E * C * (2 * cg1 + 2 * rows1), bound="hbm"),
        "L1.sym_contract": dict(bytes=f4 * (2 * N * D * C + ncoef * C), flops=N * C * 2 * ncoef, bound="fp32"),
        "L0.sym_contract": dict(bytes=f4 * (2 * N * D * C + ncoef * C), flops=N * C * 2 * ncoef, bound="fp32"),
        "L1.radial_mlp": dict(bytes=f4 * (E * 12 + E * P1 * C + 2 * 64 * (64 + P1 * C)), flops=2 * E * (12 * 64 + 64 * 64 + 64 * P1 * C), bound="hbm"),
        "L1.interaction_linear": dict(bytes=f4 * (N * rows1 * C + N * D * C + rows1 * C * C // D), flops=2 * N * rows1 * C * C, bound="hbm"),
    }
    return out


def ncu_traffic(tag, n_nodes, n_edges):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the kernel behind `tag`, from the committed
    `ncu --set full` capture (profiles/ncu_traffic.json); None when no capture exists for this workload size."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_traffic.json")
    try:
        with open(path) as f:
            ent = json.load(f).get(tag)
    except (OSError, ValueError):
        return None, None
    if not ent or ent.get("n_nodes") != n_nodes or ent.get("n_edges") != n_edges:
        return None, None
    return ent["dram_bytes_read"] + ent["dram_bytes_write"], ent.get("source")


_REAL_STDOUT = None


def _guard_stdout():
    """stdout must hold exactly ONE JSON line: point fd 1 at stderr for the whole run (NCCL prints its version banner to
    stdout from C at NCCL_DEBUG >= VERSION, libraries may print warnings) and keep the real stdout for `_emit`."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def _emit(line: dict):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        os.write(1, data)
    else:
        os.write(_REAL_STDOUT, data)


def main():
    _guard_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--lattices", type=int, default=None, help="lattices per GPU per step (default: the config's)")
    ap.add_argument("--config", type=int, default=2, choices=[2, 4, 5],
                    help="BASELINE.json workload: 2 = headline (1024 x 30-node lattices, default model); 4 = lmax 3, C = 128, 256 lattices; "
                         "5 = 1024 x 216-node lattices (inference only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the training-step leg (BASELINE.json configs[2])")
    ap.add_argument("--train-lattices", type=int, default=512, help="lattices per GPU per training step")
    ap.add_argument("--sharded", action="store_true",
                    help="build ONE global batch of gpus x lattices lattices and give every rank its edge-balanced shard through "
                         "sharding.shard_batch (default for --config 5); otherwise every rank generates its own batch (weak scaling)")
    ap.add_argument("--no-small-batch", action="store_true", help="skip the batch-64 CUDA-graph leg")
    ap.add_argument("--small-batch", type=int, default=64, help="lattices per step of the small-batch leg (scripts/train_main.py:42)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # workload table: lattices/GPU, nodes/lattice, first seed, model overrides, BASELINE.json index
    cfg = {2: (1024, 30, 2000, {}, "default hyper-parameters (lmax=4, 32 channels, correlation 3, 2 message passes)", 1),
           4: (256, 30, 20000, dict(lmax=3, hidden_irreps="128x0e+128x1o+128x2e+128x3o", readout_irreps="64x0e+64x1o+64x2e+64x3o"),
               "lmax=3, 128 channels, correlation 3, 2 message passes", 3),
           5: (1024, 216, 30000, {}, "default hyper-parameters, 216-node lattices (32 distinct lattices tiled 32x per batch)", 4)}[args.config]
    if args.lattices is None:
        args.lattices = cfg[0]
    n_nodes_lat = cfg[1]
    config = {"workload": f"EnergyEquivGNN inference fwd+stiffness, batch {args.lattices} synthetic lattices/GPU "
                          f"({n_nodes_lat} nodes, {2 * n_nodes_lat} struts, {4 * n_nodes_lat} directed edges each), {cfg[4]}, "
                          f"fp32 [BASELINE.json configs[{cfg[5]}]]",
              "lattices_per_gpu": args.lattices, "nodes_per_lattice": n_nodes_lat, "edges_per_lattice": 4 * n_nodes_lat,
              "l2": "per-step working set (GBs of intermediates) >> 126 MB L2; inputs rotate over 2 distinct batches"}
    if args.config != 2:
        args.no_train = True
        args.no_cpu_baseline = True
        args.no_small_batch = True
    if args.config == 5:
        args.sharded = True
    config["sharding"] = ("one global batch of gpus x lattices_per_gpu lattices, contiguous edge-balanced shards (sharding.shard_batch), no "
                          "data-path collective") if args.sharded else "every rank generates its own lattices_per_gpu batch (independent units)"

    if args.impl == "reference":
        if rank != 0:
            return
        torch.set_num_threads(os.cpu_count() or 1)
        val, ms, thr = cpu_reference_run(max(1, args.steps), max(1, args.warmup), 2000)
        line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": val, "unit": UNIT, "cores": thr, "kind": "port",
                                 "sample": f"{CPU_SAMPLE} lattices/step of the same synthetic workload (seeds 2000..), oracle/ CPU restatement of the "
                                           "reference formulation (dense w3j tensor product, U.W.x contraction), torch fp32"},
                "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        _emit(line)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the hot path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"  # the version banner goes to stdout, which must hold exactly one JSON line
        dist.init_process_group("nccl", device_id=dev)
    from energy_equiv_lattice_gnn_b200 import EnergyEquivGNN, native
    from energy_equiv_lattice_gnn_b200.data import synthetic_batch
    from energy_equiv_lattice_gnn_b200.params import default_params
    native.lib()

    G = args.lattices
    host_batches = []
    rmax = 0.0
    shard_info = None
    for k in range(2):
        from energy_equiv_lattice_gnn_b200.data import collate, synthetic_lattice
        if args.sharded:
            # the SAME global batch on every rank (seeded), then this rank's shard
            from energy_equiv_lattice_gnn_b200.sharding import shard_batch
            Gg = G * world
            n_distinct = min(32 if n_nodes_lat > 64 else 256, Gg)
            distinct = [synthetic_lattice(cfg[2] + k * n_distinct + i, n_nodes_lat) for i in range(n_distinct)]
            gb = collate([distinct[i % n_distinct] for i in range(Gg)])
            r = max(float(d.edge_attr.max()) for d in distinct)
            b, (g0, g1) = shard_batch(gb, rank, world)
            shard_info = {"global_lattices": Gg, "range": [g0, g1], "nodes": int(b.positions.shape[0]), "edges": int(b.edge_index.shape[1])}
            del gb
        elif args.config == 5:   # 216-node lattices are slow to generate on the host: 32 distinct ones, tiled
            distinct = [synthetic_lattice(cfg[2] + (2 * rank + k) * 32 + i, n_nodes_lat) for i in range(min(32, G))]
            data = [distinct[i % len(distinct)] for i in range(G)]
            b, r = collate(data), max(float(d.edge_attr.max()) for d in distinct)
        else:
            seed0 = cfg[2] + (2 * rank + k) * G
            b, r = synthetic_batch(range(seed0, seed0 + G), n_nodes_lat)
        host_batches.append(b.pin_memory())
        rmax = max(rmax, r)
    G = int(host_batches[0].num_graphs)   # sharded: this rank's share
    params = default_params(max_edge_radius=rmax, **cfg[3])
    torch.manual_seed(0)
    model = EnergyEquivGNN(params).to(dev).eval()
    dev_batches = [b.to(dev) for b in host_batches]
    n_nodes, n_edges = dev_batches[0].positions.shape[0], dev_batches[0].edge_index.shape[1]

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        with torch.no_grad():
            for i in range(warmup):
                fn(i)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(steps):
                fn(i)
            e1.record()
            barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    results = []

    def step_resident(i):
        results.append(model(dev_batches[i % 2])["stiffness"])
        if len(results) > 4:
            results.pop(0)

    # end to end through the public streaming API: every step uploads its own pinned host batch and brings its [G,6,6]
    # result back to pinned host memory; upload / forward / download of neighbouring steps overlap on separate streams
    from energy_equiv_lattice_gnn_b200.serve import StreamingPredictor
    predictor = StreamingPredictor(model)

    def timed_e2e(steps, warmup):
        def run(k):
            n = 0
            for res in predictor.run(host_batches[i % 2] for i in range(k)):
                n += res.shape[0]
            torch.cuda.current_stream().wait_event(predictor.last_done)
            return n
        run(warmup)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        assert run(steps) == steps * G
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    sampler = ClockSampler(local_rank) if rank == 0 else None
    l0 = native.launch_count
    ms_total = timed(step_resident, args.steps, max(3, args.warmup))
    launches = native.launch_count - l0
    launches_timed = launches * args.steps // (args.steps + max(3, args.warmup))
    clocks = sampler.stop() if sampler else {}
    ms_e2e = timed_e2e(args.steps, max(3, args.warmup))
    hb = host_batches[0]
    h2d = sum(getattr(hb, k).numel() * getattr(hb, k).element_size() for k in ("positions", "node_attrs", "edge_index", "shifts", "edge_attr", "batch"))
    d2h = G * 36 * 4

    # per-kernel durations, CUDA events on the launching stream, same steps
    native.start_profile()
    with torch.no_grad():
        for i in range(args.steps):
            model(dev_batches[i % 2])
    prof = native.stop_profile()
    kern = {k: sum(v) / len(v) for k, v in prof.items()}
    step_sum = sum(kern.values())
    alg = algorithmic_model(model, n_nodes, n_edges)
    peak, peak_src = measured_peaks()
    # north_star: the message / scatter kernel against the HBM roofline (bytes of SURVEY.md section 8(d)); the symmetric
    # contraction is fp32-FMA bound and gets its own entry against the measured fp32 ceiling (roofline_kernels)
    top = "L1.tp_scatter" if "L1.tp_scatter" in kern else max((k for k in kern if k in alg), key=lambda k: kern[k])
    top_ms = kern[top]
    achieved = alg[top]["bytes"] / (top_ms * 1e-3) / 1e9
    traffic, traffic_src = ncu_traffic(top, n_nodes, n_edges)
    fpeak, fpeak_src = fp32_peak()
    roofline = {"kernel": top, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "ms_per_launch": top_ms, "share_of_step": top_ms / step_sum,
                "algorithmic_bytes_per_launch": alg[top]["bytes"],
                "algorithmic_bytes_note": "SURVEY 8(d) message+scatter, streamed per-edge weights: 856 KB per 30-node lattice",
                "frac_fused_211KB": alg[top].get("bytes_fused", 0) / (top_ms * 1e-3) / 1e9 / peak,
                "as_built_bytes_per_launch": alg[top].get("bytes_as_built"),
                "algorithmic_flops_per_launch": alg[top]["flops"],
                "fp32_tflops_achieved": alg[top]["flops"] / (top_ms * 1e-3) / 1e12, "fp32_tflops_peak_measured": fpeak,
                "fp32_frac": alg[top]["flops"] / (top_ms * 1e-3) / 1e12 / fpeak, "fp32_peak_source": fpeak_src,
                "limiter": "issue / shared-memory broadcast bound (lane = channel: one LDS.128 of the per-edge G matrix per 4 FMAs), "
                           "not HBM: profiles/r01_ncu_tp_fwd_s13_summary.txt"}
    roofline_kernels = []
    for k in sorted((k for k in kern if k in alg), key=lambda k: -kern[k]):
        t = kern[k] * 1e-3
        ent = {"kernel": k, "ms_per_launch": kern[k], "share_of_step": kern[k] / step_sum, "bound": alg[k]["bound"],
               "GBps": alg[k]["bytes"] / t / 1e9, "hbm_frac": alg[k]["bytes"] / t / 1e9 / peak,
               "fp32_tflops": alg[k]["flops"] / t / 1e12, "fp32_frac": alg[k]["flops"] / t / 1e12 / fpeak}
        roofline_kernels.append(ent)
    kernels = {k: {"ms": round(v, 4), **({"GBps": round(alg[k]["bytes"] / v / 1e6, 1), "TFLOPs": round(alg[k]["flops"] / v / 1e9, 2)} if k in alg else {})}
               for k, v in sorted(kern.items(), key=lambda kv: -kv[1])}

    # ---- training step (BASELINE.json configs[2]): fwd + loss + explicit backward + flat all-reduce + clip + AdamW(amsgrad)
    train = None
    if not args.no_train:
        from energy_equiv_lattice_gnn_b200.train import Trainer, random_psd_targets
        Gt = args.train_lattices
        tb_, _ = synthetic_batch(range(10000 + Gt * rank, 10000 + Gt * rank + Gt), N_NODES)
        tb_ = tb_.to(dev)
        tb_.stiffness = random_psd_targets(Gt, seed=rank).to(dev)
        torch.manual_seed(0)
        tmodel = EnergyEquivGNN(params).to(dev)
        trainer = Trainer(tmodel)
        k_train = min(args.steps, 10)
        l1 = native.launch_count
        for _ in range(5):
            trainer.training_step(tb_)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k_train):
            loss = trainer.training_step(tb_)
        e1.record()
        barrier()
        ms_t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if dist is not None:
            dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
        native.start_profile()
        trainer.training_step(tb_)
        tprof = {k: sum(v) for k, v in native.stop_profile().items()}
        ms_train = float(ms_t.item()) / k_train
        train = {"metric": "lattices/sec training step (fwd + loss + backward + gradient all-reduce + clip + AdamW amsgrad)",
                 "value": world * Gt / (ms_train * 1e-3), "unit": UNIT, "ms_per_step": ms_train, "lattices_per_gpu": Gt, "steps": k_train,
                 "warmup": 5, "loss_last": float(loss), "gpu_launches_per_step": (native.launch_count - l1) // (k_train + 6),
                 "allreduce_bytes_per_step": trainer.opt.grad.numel() * 4 if world > 1 else 0,
                 "kernels_ms": {k: round(v, 4) for k, v in sorted(tprof.items(), key=lambda kv: -kv[1])[:24]},
                 "deterministic": "fixed-order reductions everywhere (no atomics): two runs give bitwise identical gradients",
                 "workload": "BASELINE.json configs[2]: default hyper-parameters, random PSD targets, relative-MSE loss, no accumulation"}
        del trainer, tmodel
    # ---- small-batch leg: the reference's batch size (scripts/train_main.py:42), eager vs CUDA-graph replay, pinned host in / out
    small = None
    if not args.no_small_batch:
        from energy_equiv_lattice_gnn_b200.serve import GraphedPredictor
        Gs = args.small_batch
        sb = [synthetic_batch(range(40000 + (2 * rank + k) * Gs, 40000 + (2 * rank + k) * Gs + Gs), N_NODES)[0].pin_memory() for k in range(2)]
        host_out = torch.empty(Gs, 6, 6, pin_memory=True)
        gp = GraphedPredictor(model)

        def eager(i):
            host_out.copy_(model(sb[i % 2].to(dev, non_blocking=True))["stiffness"], non_blocking=True)

        def graphed(i):
            gp(sb[i % 2], sync=False)

        k_small = max(args.steps, 50)
        l2 = native.launch_count
        ms_eager = timed(eager, k_small, 5) / k_small
        launches_small = (native.launch_count - l2) // (k_small + 5)
        ms_graph = timed(graphed, k_small, 5) / k_small
        small = {"lattices_per_step": Gs, "eager": {"value": world * Gs / (ms_eager * 1e-3), "unit": UNIT, "ms_per_step": ms_eager,
                                                    "kernel_launches_per_step": launches_small},
                 "cuda_graph": {"value": world * Gs / (ms_graph * 1e-3), "unit": UNIT, "ms_per_step": ms_graph,
                                "graph_launches_per_step": 1, "kernels_in_graph": launches_small},
                 "speedup": ms_eager / ms_graph,
                 "note": "pinned host batch -> H2D -> forward -> D2H of [G,6,6] every step; same kernels, replayed as one CUDA graph"}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    ms_step = ms_total / args.steps
    total_lattices = shard_info["global_lattices"] if shard_info else world * G
    value = total_lattices / (ms_step * 1e-3)
    if shard_info:
        config["shard_rank0"] = shard_info
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": config, "clocks": clocks,
            "e2e": {"value": total_lattices / (ms_e2e / args.steps * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches_timed, "roofline": roofline, "roofline_kernels": roofline_kernels, "kernels": kernels}
    if small is not None:
        line["small_batch"] = small
    if train is not None:
        line["train_step"] = train
    if world == 1 and not args.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        val, ms, thr = cpu_reference_run(3, 1, 2000)
        line["cpu_baseline"] = {"value": val, "unit": UNIT, "cores": thr, "kind": "port", "ms_per_step": ms,
                                "sample": f"{CPU_SAMPLE} lattices/step (seeds 2000..{2000 + CPU_SAMPLE - 1}) of the same workload, 1 warm-up + 3 timed, "
                                          "oracle/ CPU restatement of the reference formulation, torch fp32"}
    _emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
