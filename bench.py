#!/usr/bin/env python
"""bench.py -- train samples/s of the MS-GAT hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config pemsd7]
                    [--scaling weak|strong] [--precision fp32|tf32]

A "step" is what the train branch of the reference engine does per batch (src/engine.py:54-70): forward,
Huber loss, zero_grad, backward, (gradient all-reduce when N > 1), Adam step, metrics update -- on synthetic
PEMS-shaped data (the datasets are Drive-hosted).  Default workload: BASELINE.json configs[3],
PEMSD7-shaped (N=883, C=1, M=5), msgat72, fp32-grade path, batch 32 PER GPU (weak scaling); ``--scaling strong``
keeps the GLOBAL batch at 32 (README.md:33 ``-b 32``) and shards it, and at N > 1 the default line carries a
``strong`` sub-record measured in the same run.

Every kernel inside the timed region is this library's (model, loss + metrics, Adam) or NCCL's.
`value`   : samples/s with the batch already resident in HBM: K steps timed with CUDA events between barriers,
            max over ranks; the K-step window is repeated (``--windows``) and the MEDIAN window is reported
            (`windows_ms_per_step` lists all of them, `ms_per_step_min` the fastest).
`e2e`     : same metric through the public module call with the batch in pinned HOST memory: H2D copies of
            X/H/D/Y and the D2H read of the loss are inside the timed region.
`roofline`: the kernel with the largest share of device time (per-kernel CUDA events recorded by the library on its
            launch stream, in a separate pass right after the timed one), algorithmic HBM bytes per launch / mean
            launch time vs MEASURED_PEAKS.json.
`cpu_baseline` / --impl reference : the reference's own CPU implementation of the step on the host cores -- the
            unmodified reference modules from oracle/_ref ("reference") when that git-ignored copy travelled with
            the snapshot, else the restatement oracle/msgat_oracle.py ("port") -- on a bounded sample.
"""
import argparse
import datetime
import gc
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
    ap.add_argument("--windows", type=int, default=5, help="repetitions of the K-step timed window (median reported)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="pemsd7")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the config's; strong: config's / N)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--model", default="msgat72")
    ap.add_argument("--precision", default="fp32", choices=["fp32", "tf32"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="samples per step of the CPU baseline (0: the full per-GPU batch)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling sub-record at N > 1")
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
            time.sleep(0.05)

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


def cpu_reference_arm(args, shape, Bs, steps, warmup):
    """The reference's CPU path for one training step (engine.py:54-70: forward, Huber, zero_grad, backward, Adam,
    metrics) on all host threads, on a bounded sample of the workload.  Runs the UNMODIFIED reference modules from
    oracle/_ref when that copy is present, else the restatement in oracle/msgat_oracle.py."""
    from msgat_b200 import synthetic
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import make_ref
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    adj = synthetic.make_adjacency(args.config)
    batch = synthetic.make_batch(Bs, shape["M"], shape["C"], shape["N"], seed=1234)
    torch.manual_seed(0)
    if make_ref.available():
        kind = "reference"
        ref_models, ref_loss, ref_metrics = make_ref.load()
        factory = {"msgat48": ref_models.msgat48, "msgat72": ref_models.msgat72, "msgat96": ref_models.msgat96}[args.model]
        model = factory(n_components=shape["M"], in_channels=shape["C"], in_timesteps=12, out_timesteps=12, use_te=True,
                        adj=adj.clone())
        loss_fn = ref_loss.HuberLoss(DELTA)
        opt = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=5e-4)   # engine.py:106
        metrics = ref_metrics.Metrics()

        def one_step():
            pred = model(batch.X, batch.H, batch.D)
            loss = loss_fn(pred, batch.Y)
            opt.zero_grad()
            loss.backward()
            opt.step()
            loss.item()
            metrics.update(pred.detach(), batch.Y)
    else:
        kind = "port"
        from oracle import msgat_oracle as O
        p = O.init_params(args.model, shape["M"], shape["C"], adj)
        train = [v.requires_grad_(True) for k, v in p.items() if k != "adj"]
        opt = torch.optim.Adam(train, lr=1e-3, weight_decay=5e-4)
        state = {}

        def one_step():
            pred = O.msgat(batch.X, batch.H, batch.D, p, O.MODEL_DILATIONS[args.model])
            loss = O.huber(pred, batch.Y, DELTA)
            opt.zero_grad()
            loss.backward()
            opt.step()
            float(loss.detach())
            O.metrics_update(state, pred.detach(), batch.Y)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        one_step()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    dt = sum(times) / len(times)
    what = "unmodified reference modules (oracle/_ref)" if kind == "reference" else "oracle port (oracle/msgat_oracle.py)"
    return {"value": Bs / dt, "unit": "samples/s", "cores": cores, "kind": kind,
            "sample": f"{Bs} samples/step of the {args.config} workload ({args.model}, N={shape['N']}, fp32), {what}, "
                      f"{len(times)} timed steps of {dt * 1e3:.0f} ms after {warmup} warm-up"}, dt


def algorithmic_bytes(kernel: str, S: int, N: int):
    """HBM bytes one launch of `kernel` must move: activations read and written once, fp32,
    nothing else spilled.  kernel = "name<..>@C<in>O<out>"; S = samples the launch covers (components x batch)."""
    name, _, tag = kernel.partition("@")
    name = name.strip("()").split("<")[0]
    try:
        C = int(tag[1:tag.index("O")]); O = int(tag[tag.index("O") + 1:])
    except ValueError:
        return None
    A = S * N * 12 * 4  # bytes of one channel plane of every sample of the launch
    b = O // 3
    Cz = b if C > b else C
    table = {
        "k_prep": (C + Cz + 2) * A,
        "k_apply": (C + Cz + O) * A,
        "k_tc_apply": (C + Cz + O + 3 * b) * A,        # + Y1 and the first-conv taps kept for the backward
        "k_tc_apply3": (C + Cz + O + 3 * b) * A,
        "k_tc_dgrad": (2 * C + O + 5 * b) * A,         # x, G, taps, GZ (from k_l2_gz) in; dx and dA_now | dA_past out
        "k_bwd_main": (2 * C + O + Cz + 3 * b) * A,    # x, dout, Z gather, dx + dA_now/dA_past/GZ handed to k_tc_wgrad
        "k_tc_wgrad": (C + O + b + 3 * b) * A,         # x, dout, Y1, dA_now/dA_past/GZ
        "k_bwd_fix": 3 * C * A,
        "k_relu_mask": 3 * O * A,
        "k_head_fwd": C * A, "k_head_fwd12": C * A,
        "k_head_bwd": 2 * C * A, "k_head_bwd12": 2 * C * A,
        "k_sstats": 15 * S * N * 4 + S * N * 12 * 4,
        "k_scol": (12 + 3 + 12) * S * N * 4,
        # layer-1 streaming kernels (csrc/meam_l1.cu): x and the node statistics in, the 72-channel output out
        "k_l1_fwd": (C + O) * A,
        "k_l1_bwd": (C + 2 * O) * A,                   # x, out (ReLU mask), dout
        # layer-2 pipelined tcgen05 kernels (csrc/l2_*.cu)
        "k_l2_mix": (C + 3 * b) * A,                   # x in; first-conv taps a_now | a_past and Z out
        "k_l2_out": (C + 2 * b + Cz + O + b) * A,      # x, taps, Z gather; out and Y1
        "k_l2_gz": 3 * b * A,                          # G_s and Z in (the per-edge gathers hit L2), GZ out
        "k_l2_dy": (2 * b + 2 * b + b + 3 * b) * A,    # G_t, G_s, taps, Z gather; dA_now | dA_past | GZ
        "k_l2_dx": (2 * C + O + 3 * b) * A,            # G, dA/GZ, x; dx
    }
    return table.get(name)


class Run:
    """One model + optimizer + (graphed) step at a given per-GPU batch."""

    def __init__(self, args, shape, Bper, rank, world, local):
        from msgat_b200 import _lib, models, synthetic
        from msgat_b200.distributed import FlatGradients
        from msgat_b200.loss import Metrics, huber_loss
        from msgat_b200.optim import FlatAdam
        self.args, self.world, self.rank, self.Bper = args, world, rank, Bper
        self.dev = torch.device("cuda", local)
        torch.manual_seed(0)
        adj = synthetic.make_adjacency(args.config)
        self.model = getattr(models, args.model)(n_components=shape["M"], in_channels=shape["C"], in_timesteps=12,
                                                 out_timesteps=12, use_te=True, adj=adj.clone()).to(self.dev)
        if args.precision == "tf32":
            self.model.set_precision(_lib.PREC_TF32)
        self.fg = FlatGradients(self.model, direct=True)
        self.opt = FlatAdam(self.fg, lr=1e-3, weight_decay=5e-4, grad_scale=1.0 / world)   # engine.py:106
        self.metrics = Metrics()
        self.loss_fn = lambda o, y: huber_loss(o, y, DELTA, self.metrics)                  # engine.py:56,70
        self.host = synthetic.make_batch(Bper, shape["M"], shape["C"], shape["N"], seed=1234 + rank, pin=True)
        self.devb = [t.to(self.dev) for t in (self.host.X, self.host.H, self.host.D, self.host.Y)]
        self.h2d = sum(t.numel() * t.element_size() for t in (self.host.X, self.host.H, self.host.D, self.host.Y))
        self.loss_host = torch.empty((), dtype=torch.float32).pin_memory()
        self.graphed = None
        self.feeder = None

    def step(self, X, H, D, Y):
        out = self.model(X, H, D)
        loss = self.loss_fn(out, Y)
        self.fg.zero()
        loss.backward()
        self.fg.all_reduce_sum()
        self.opt.step()
        return loss

    def prepare(self):
        from msgat_b200 import _lib
        args = self.args
        for _ in range(max(args.warmup, 3)):
            self.step(*self.devb)
        torch.cuda.synchronize()
        l0 = _lib.launch_count()   # launches of this library per step, counted on an eager step
        self.step(*self.devb)
        self.launches = _lib.launch_count() - l0
        if args.graph != "off":
            from msgat_b200.train import GraphedStep
            try:
                self.graphed = GraphedStep(self.model, self.loss_fn, self.opt, self.fg, self.devb, warmup=1)
                for _ in range(2):
                    self.graphed(*self.graphed.inputs)
                torch.cuda.synchronize()
            except Exception as e:  # noqa: BLE001 - a failed capture must not take the bench down
                if args.graph == "on":
                    raise
                print(f"[bench] CUDA graph capture failed, running eager: {e!r}", file=sys.stderr)
                self.graphed = None
                torch.cuda.synchronize()

    def run_resident(self):
        if self.graphed is not None:
            self.graphed(*self.graphed.inputs)
        else:
            self.step(*self.devb)

    def run_e2e(self):
        # the NEXT step's batch is already on its way (side stream) while this step computes; every
        # step's H2D copy and loss read-back happen inside the timed region
        i, (X, H, D, Y) = self.feeder.take()
        self.feeder.stage((self.host.X, self.host.H, self.host.D, self.host.Y))
        loss = self.graphed(X, H, D, Y) if self.graphed is not None else self.step(X, H, D, Y)
        self.feeder.release(i)
        self.loss_host.copy_(loss.detach(), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(self.loss_host)

    def start_e2e(self):
        from msgat_b200.train import DeviceFeeder
        self.feeder = DeviceFeeder(self.devb, self.dev)
        self.feeder.stage((self.host.X, self.host.H, self.host.D, self.host.Y))
        for _ in range(2):
            self.run_e2e()

    def close(self):
        if self.graphed is not None:
            try:
                self.graphed.graph.reset()
            except Exception:
                pass
            self.graphed = None
        self.feeder = None
        gc.collect()
        torch.cuda.synchronize()


def main():
    args = parse()
    from msgat_b200 import synthetic
    shape = dict(synthetic.PEMS_SHAPES[args.config])
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.scaling == "strong":
        if shape["B"] % world:
            raise SystemExit(f"strong scaling: global batch {shape['B']} does not divide over {world} ranks")
        Bper = args.batch or shape["B"] // world
    else:
        Bper = args.batch or shape["B"]
    workload = (f"{args.config}-shaped synthetic: N={shape['N']} sensors, C={shape['C']}, M={shape['M']} components, "
                f"T=12, {args.model}, batch {shape['B']} per GPU")
    if args.scaling == "strong":
        workload = workload.replace(f"batch {shape['B']} per GPU", f"GLOBAL batch {Bper * world} sharded over the GPUs")
    elif Bper != shape["B"]:
        workload = workload.replace(f"batch {shape['B']} per GPU", f"batch {Bper} per GPU")
    config = {"workload": workload, "global_batch": Bper * max(world, args.gpus if world == 1 else world),
              "parallelism": f"dp{world} (batch shard + flat NCCL all-reduce)" if world > 1 else "single GPU",
              "precision_path": args.precision,
              "l2": "activations per MEAM (98 MB at layer 2) are streamed, the step touches > 2 GB >> 126 MB L2"}

    if args.impl == "reference":
        if rank != 0:
            return
        # bounded: at ~3 s per step on 16 cores the arm caps its timed steps at 6
        steps = max(1, min(args.steps, 6))
        Bs = args.cpu_sample if args.cpu_sample > 0 else Bper
        base, dt = cpu_reference_arm(args, shape, Bs, steps, max(1, min(args.warmup, 1)))
        line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": "samples/s", "n_gpus": args.gpus,
                "steps": steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": base, "gpu_launches": 0,
                "e2e": {"value": base["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    import torch.distributed as dist
    from msgat_b200 import _lib
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))

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

    def windows(fn, n, reps):
        w = sorted(timed(fn, n) for _ in range(max(1, reps)))
        return w[len(w) // 2], w

    run = Run(args, shape, Bper, rank, world, local)
    run.prepare()
    config["cuda_graph"] = run.graphed is not None
    config["timing"] = f"{args.windows} windows of {args.steps} steps, median window reported"
    with ClockSampler(local) as clk:
        ms, wins = windows(run.run_resident, args.steps, args.windows)
    run.start_e2e()
    ms_e2e, wins_e2e = windows(run.run_e2e, args.steps, min(args.windows, 3))
    gB = Bper * world
    value = gB * args.steps / (ms * 1e-3)
    e2e = gB * args.steps / (ms_e2e * 1e-3)

    # per-kernel device time (CUDA events recorded by the library on the launch stream), eager launches, one stream
    roof = None
    kernels = {}
    torch.cuda.synchronize()
    if hasattr(run.model, "concurrent_components"):
        run.model.concurrent_components = False
    if rank == 0:
        _lib.profile(True)
    for _ in range(args.profile_steps):   # every rank runs them (they contain the gradient all-reduce); rank 0 records
        run.step(*run.devb)
    torch.cuda.synchronize()
    if rank == 0:
        _lib.profile(False)
        rep = _lib.profile_report()
        total = sum(v[1] for v in rep.values()) or 1.0
        peak, src = peaks()
        S = Bper * shape["M"]
        for k, (cnt, tms) in sorted(rep.items(), key=lambda kv: -kv[1][1]):
            per = cnt / args.profile_steps
            Sk = S if per <= 2.5 else Bper   # a grouped launch covers all M components, a per-component one a batch
            ab = algorithmic_bytes(k, Sk, shape["N"])
            kernels[k] = {"launches_per_step": per, "ms_per_launch": tms / cnt, "share": tms / total,
                          "GBps": None if ab is None else ab / (tms / cnt * 1e-3) / 1e9}
        top = max(rep.items(), key=lambda kv: kv[1][1])
        per = top[1][0] / args.profile_steps
        ab = algorithmic_bytes(top[0], S if per <= 2.5 else Bper, shape["N"])
        avg_ms = top[1][1] / top[1][0]
        ach = None if ab is None else ab / (avg_ms * 1e-3) / 1e9
        traffic = None  # dram read+write bytes per launch from the committed ncu --set full capture
        tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                traffic = json.load(f).get(f"{args.config}:{args.precision}", {}).get(top[0])
        roof = {"kernel": top[0], "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": None if ach is None else ach / peak, "traffic": traffic, "peak_source": src,
                "algorithmic_bytes_per_launch": ab, "avg_launch_ms": avg_ms, "share_of_device_time": top[1][1] / total,
                "my_kernels_ms_per_step": total / args.profile_steps}
    launches = run.launches
    run.close()

    # strong-scaling sub-record (BASELINE.json configs[3]: GLOBAL batch 32 sharded over the GPUs), same run
    strong = None
    if world > 1 and args.scaling == "weak" and not args.no_strong and shape["B"] % world == 0:
        srun = Run(args, shape, shape["B"] // world, rank, world, local)
        srun.prepare()
        sms, swins = windows(srun.run_resident, args.steps, args.windows)
        strong = {"global_batch": shape["B"], "per_gpu_batch": shape["B"] // world,
                  "value": shape["B"] * args.steps / (sms * 1e-3), "unit": "samples/s", "ms_per_step": sms / args.steps,
                  "windows_ms_per_step": [w / args.steps for w in swins], "cuda_graph": srun.graphed is not None}
        srun.close()

    def finish():
        sys.stdout.flush()
        sys.stderr.flush()
        if world > 1:
            # orderly: the captured graphs were reset in Run.close(), nothing of NCCL's is in flight any more
            wd = threading.Timer(60.0, lambda: os._exit(3))   # watchdog only (exit code 3 = the teardown hung); daemon: it
            wd.daemon = True                                   # does not keep a finished process alive
            wd.start()
            barrier()
            dist.destroy_process_group()

    if rank != 0:
        finish()
        return
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        cpu, _ = cpu_reference_arm(args, shape, args.cpu_sample if args.cpu_sample > 0 else Bper, 2, 1)
    line = {"metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "ms_per_step_min": wins[0] / args.steps,
            "windows_ms_per_step": [w / args.steps for w in wins], "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f32" if args.precision == "fp32" else "tf32", "data": "synthetic",
            "config": config, "clocks": clk.summary(),
            "e2e": {"value": e2e, "unit": "samples/s", "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": run.h2d, "d2h_bytes_per_step": 4},
            "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu, "kernels": kernels}
    if strong is not None:
        line["strong"] = strong
    print(json.dumps(line))
    finish()


if __name__ == "__main__":
    main()
