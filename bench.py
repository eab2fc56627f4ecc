#!/usr/bin/env python
"""bench.py -- train samples/s of the MS-GAT hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config pemsd7]

A "step" is what the train branch of the reference engine does per batch
(src/engine.py:54-63): forward, Huber loss, zero_grad, backward, (gradient all-reduce when
N > 1), Adam step -- on synthetic PEMS-shaped data (the datasets are Drive-hosted,
SURVEY.md 8d).  Default workload: BASELINE.json configs[3], PEMSD7-shaped (N=883, C=1, M=5),
batch 32 PER GPU (weak scaling), msgat72, fp32 path.

`value`   : samples/s with the batch already resident in HBM (CUDA events, max over ranks).
`e2e`     : same metric through the public module call with the batch in pinned HOST memory:
            H2D copies of X/H/D/Y and the D2H read of the loss are inside the timed region.
`roofline`: the kernel with the largest share of device time (per-kernel CUDA events recorded
            by the library on its launch stream, in a separate pass right after the timed one),
            algorithmic HBM bytes per launch / mean launch time vs MEASURED_PEAKS.json.
`cpu_baseline` / --impl reference : the CPU restatement of the reference (oracle/, "port":
            the reference itself is a Python repo that cannot travel to the GPU box) timed on
            the host cores on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "train samples/sec (fwd+bwd+Adam) at PEMSD7 shape"
DELTA = 50.0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="pemsd7")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the config's)")
    ap.add_argument("--model", default="msgat72")
    ap.add_argument("--precision", default="fp32", choices=["fp32", "tf32"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="samples per step of the CPU baseline (0: the full per-GPU batch)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile-steps", type=int, default=3)
    ap.add_argument("--graph", default="auto", choices=["auto", "on", "off"],
                    help="replay the step as one CUDA graph (msgat_b200.train.GraphedStep); auto: on, eager if capture fails")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], False
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def cpu_reference_arm(args, shape, steps, warmup):
    """The reference's CPU path (oracle port of src/models + the train branch of
    src/engine.py:54-63 incl. Adam), all host threads, bounded sample of the workload."""
    from msgat_b200 import synthetic
    from oracle import msgat_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    Bs = args.cpu_sample if args.cpu_sample > 0 else (args.batch or shape["B"])
    adj = synthetic.make_adjacency(args.config)
    p = O.init_params(args.model, shape["M"], shape["C"], adj)
    train = [v.requires_grad_(True) for k, v in p.items() if k != "adj"]
    opt = torch.optim.Adam(train, lr=1e-3, weight_decay=5e-4)
    batch = synthetic.make_batch(Bs, shape["M"], shape["C"], shape["N"], seed=1234)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        loss = O.train_step_loss(p, args.model, batch.X, batch.H, batch.D, batch.Y, DELTA)
        opt.zero_grad()
        loss.backward()
        opt.step()
        float(loss.detach())
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    dt = sum(times) / len(times)
    return {"value": Bs / dt, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": f"{Bs} samples/step of the {args.config} workload ({args.model}, N={shape['N']}, fp32), "
                      f"{len(times)} timed steps of {dt * 1e3:.0f} ms after {warmup} warm-up"}, dt


def algorithmic_bytes(kernel: str, B: int, N: int):
    """HBM bytes one launch of `kernel` must move: activations read
    and written once, fp32, nothing else spilled.  kernel = "name<..>@C<in>O<out>"."""
    name, _, tag = kernel.partition("@")
    name = name.strip("()").split("<")[0]
    try:
        C = int(tag[1:tag.index("O")]); O = int(tag[tag.index("O") + 1:])
    except ValueError:
        return None
    A = B * N * 12 * 4  # bytes of one channel plane of the batch
    b = O // 3
    Cz = b if C > b else C
    table = {
        "k_prep": (C + Cz + 2) * A,
        "k_apply": (C + Cz + O) * A,
        "k_tc_apply": (C + Cz + O + 3 * b) * A,        # + Y1 and the first-conv taps kept for the backward
        "k_tc_apply3": (C + Cz + O + 3 * b) * A,
        "k_tc_dgrad": (2 * C + O + Cz + 5 * b) * A,    # x, dout, Z gather, saved taps; dx, dA_now/dA_past/GZ
        "k_bwd_main": (2 * C + O + Cz + 3 * b) * A,    # x, dout, Z gather, dx + dA_now/dA_past/GZ handed to k_tc_wgrad
        "k_tc_wgrad": (C + O + b + 3 * b) * A,         # x, dout, Y1, dA_now/dA_past/GZ
        "k_bwd_fix": 3 * C * A,
        "k_relu_mask": 3 * O * A,
        "k_head_fwd": C * A, "k_head_fwd12": C * A,
        "k_head_bwd": 2 * C * A, "k_head_bwd12": 2 * C * A,
        "k_sstats": 15 * B * N * 4 + B * N * 12 * 4,
        "k_scol": (12 + 3 + 12) * B * N * 4,
    }
    return table.get(name)


def main():
    args = parse()
    from msgat_b200 import synthetic
    shape = dict(synthetic.PEMS_SHAPES[args.config])
    Bper = args.batch or shape["B"]
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    workload = (f"{args.config}-shaped synthetic: N={shape['N']} sensors, C={shape['C']}, M={shape['M']} components, "
                f"T=12, {args.model}, batch {Bper} per GPU")
    config = {"workload": workload, "global_batch": Bper * max(world, args.gpus if world == 1 else world),
              "parallelism": f"dp{world} (batch shard + flat NCCL all-reduce)" if world > 1 else "single GPU",
              "precision_path": args.precision,
              "l2": "activations per MEAM (98 MB at layer 2) are streamed, the step touches > 2 GB >> 126 MB L2"}

    if args.impl == "reference":
        if rank != 0:
            return
        # bounded: at ~12 s per sample-step on 8 cores the arm caps its timed steps at 6
        args.steps = max(1, min(args.steps, 6))
        base, dt = cpu_reference_arm(args, shape, args.steps, max(1, min(args.warmup, 1)))
        line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": "samples/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": base, "gpu_launches": 0,
                "e2e": {"value": base["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    import torch.distributed as dist
    from msgat_b200 import _lib, models
    from msgat_b200.distributed import FlatGradients
    from oracle.msgat_oracle import huber  # the loss of engine.py:56 (test/bench infrastructure only)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev, timeout=__import__("datetime").timedelta(seconds=180))
    torch.manual_seed(0)
    adj = synthetic.make_adjacency(args.config)
    model = getattr(models, args.model)(n_components=shape["M"], in_channels=shape["C"], in_timesteps=12,
                                        out_timesteps=12, use_te=True, adj=adj.clone()).to(dev)
    if args.precision == "tf32":
        model.set_precision(_lib.PREC_TF32)
    fg = FlatGradients(model)
    opt = torch.optim.Adam(fg.params, lr=1e-3, weight_decay=5e-4, fused=True, capturable=args.graph != "off")  # engine.py:106
    host = synthetic.make_batch(Bper, shape["M"], shape["C"], shape["N"], seed=1234 + rank, pin=True)
    devb = [t.to(dev) for t in (host.X, host.H, host.D, host.Y)]
    h2d = sum(t.numel() * t.element_size() for t in (host.X, host.H, host.D, host.Y))
    loss_host = torch.empty((), dtype=torch.float32).pin_memory()

    def step(X, H, D, Y):
        out = model(X, H, D)
        loss = huber(out, Y, DELTA)
        fg.zero()
        loss.backward()
        fg.all_reduce_mean()
        opt.step()
        return loss

    feeder = None

    def step_e2e():
        # the NEXT step's batch is already on its way (side stream) while this step computes; every
        # step's H2D copy and loss read-back happen inside the timed region
        i, (X, H, D, Y) = feeder.take()
        feeder.stage((host.X, host.H, host.D, host.Y))
        loss = graphed(X, H, D, Y) if graphed is not None else step(X, H, D, Y)
        feeder.release(i)
        loss_host.copy_(loss.detach(), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(loss_host)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    def timed(fn, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms

    for _ in range(max(args.warmup, 3)):
        step(*devb)
    # launches of this library per step, counted on an eager step (a graph replay launches the same kernels)
    torch.cuda.synchronize()
    l0 = _lib.launch_count()
    step(*devb)
    launches = _lib.launch_count() - l0
    graphed = None
    if args.graph != "off":
        from msgat_b200.train import GraphedStep
        try:
            graphed = GraphedStep(model, lambda o, y: huber(o, y, DELTA), opt, fg, devb, warmup=1)
            for _ in range(2):
                graphed(*graphed.inputs)
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001 - a failed capture must not take the bench down
            if args.graph == "on":
                raise
            print(f"[bench] CUDA graph capture failed, running eager: {e!r}", file=sys.stderr)
            graphed = None
            torch.cuda.synchronize()
    config["cuda_graph"] = graphed is not None
    run = (lambda: graphed(*graphed.inputs)) if graphed is not None else (lambda: step(*devb))
    with ClockSampler(local) as clk:
        ms = timed(run, args.steps)
    from msgat_b200.train import DeviceFeeder
    feeder = DeviceFeeder(devb, dev)
    feeder.stage((host.X, host.H, host.D, host.Y))
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    gB = Bper * world
    value = gB * args.steps / (ms * 1e-3)
    e2e = gB * args.steps / (ms_e2e * 1e-3)

    # per-kernel device time (CUDA events recorded by the library on the launch stream)
    roof = None
    kernels = {}
    # every rank runs the profiled steps (they contain the gradient all-reduce); rank 0 records
    torch.cuda.synchronize()
    model.concurrent_components = False   # one stream: a kernel's events then time that kernel alone
    if rank == 0:
        _lib.profile(True)
    for _ in range(args.profile_steps):
        step(*devb)
    torch.cuda.synchronize()
    if rank == 0:
        _lib.profile(False)
    if rank == 0:
        rep = _lib.profile_report()
        total = sum(v[1] for v in rep.values()) or 1.0
        top = max(rep.items(), key=lambda kv: kv[1][1])
        peak, src = peaks()
        for k, (cnt, tms) in sorted(rep.items(), key=lambda kv: -kv[1][1]):
            ab = algorithmic_bytes(k, Bper, shape["N"])
            kernels[k] = {"launches_per_step": cnt / args.profile_steps, "ms_per_launch": tms / cnt,
                          "share": tms / total,
                          "GBps": None if ab is None else ab / (tms / cnt * 1e-3) / 1e9}
        ab = algorithmic_bytes(top[0], Bper, shape["N"])
        avg_ms = top[1][1] / top[1][0]
        ach = None if ab is None else ab / (avg_ms * 1e-3) / 1e9
        traffic = None  # dram read+write bytes per launch from the committed ncu --set full capture
        tpath = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_traffic.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                traffic = json.load(f).get(f"{args.config}:{args.precision}", {}).get(top[0])
        roof = {"kernel": top[0], "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": None if ach is None else ach / peak, "traffic": traffic, "peak_source": src,
                "algorithmic_bytes_per_launch": ab, "avg_launch_ms": avg_ms, "share_of_device_time": top[1][1] / total,
                "my_kernels_ms_per_step": total / args.profile_steps}
    def finish():
        sys.stdout.flush()
        sys.stderr.flush()
        if world > 1:
            torch.cuda.synchronize()
            if graphed is not None:
                # tearing NCCL down while a captured graph still holds its collectives can hang: every
                # collective of the run has completed on every rank here, so leave without the teardown
                os._exit(0)
            dist.destroy_process_group()

    if rank != 0:
        finish()
        return
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        cpu, _ = cpu_reference_arm(args, shape, 2, 1)
    line = {"metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32" if args.precision == "fp32" else "tf32", "data": "synthetic",
            "config": config, "clocks": clk.summary(),
            "e2e": {"value": e2e, "unit": "samples/s", "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4},
            "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu, "kernels": kernels}
    print(json.dumps(line))
    finish()


if __name__ == "__main__":
    main()
