#!/usr/bin/env python
"""bench.py -- x4 upsampled points/sec of the PU-Flow discrete upsampling path on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU)
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (oracle port), rank 0 only

Workload (BASELINE.json configs[1]): a batch of 128 synthetic 2,048-point clouds per GPU -> 128 x 8,192 points.
One "step" = one pass of the whole x4 path over that batch through the drop-in module API
(modules.utils.patch.PatchHelper.upsample + remove_outliers around modules.discrete.interpflow.PointInterpFlow,
i.e. the body of the reference's per-file loop, /root/reference/modules/discrete/upsample.py:42-56):
normalise -> seed FPS -> patch kNN-256 -> [kNN-16 -> 6x EdgeConv+merge -> flow fwd -> interpolation -> flow inverse]
-> merge FPS -> Chamfer outlier removal.

  value : points/s with the input clouds already resident in HBM (CUDA events, max over ranks).
  e2e   : same metric through modules.discrete.upsample.upsample_array with HOST buffers: pinned host -> device copy
          of the clouds and device -> pinned host copy of the 128 x 8,192 x 3 result inside the timed region.
  network_only : the five north_star subsystems alone (PointInterpFlow.sample on pre-extracted patches).
  roofline : the kernel with the largest share of the step, algorithmic work / its average launch duration
             measured live with CUDA events on the launching stream (see DESIGN.md for the per-kernel figures).
  cpu_baseline : the CPU restatement of the reference (oracle/, kind "port"; the reference's Python cannot travel
             to the GPU box) timed on this box's host cores on a bounded sample of the same workload.

Multi-GPU: clouds are independent -> every rank processes its own 128 clouds (weak scaling), no collective on
the data path; barrier + synchronize on both sides of the timed region, max over ranks.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CKPT = os.path.join(ROOT, "tests", "golden", "puflow-x4-pu1k.pt")
METRIC = "x4_upsampled_points_per_sec"
UNIT = "points/s"
N_PTS, UP, N_OUTLIER, PATCH = 2048, 4, 24, 256


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--clouds", type=int, default=128, help="clouds per GPU per step (config 2: 128)")
    ap.add_argument("--npts", type=int, default=N_PTS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-breakdown", action="store_true")
    ap.add_argument("--ref-clouds", type=int, default=1, help="clouds per step of the CPU reference arm")
    return ap.parse_args()


def workload_name(clouds, npts):
    return f"PU1K-shape batch: {clouds} synthetic {npts}-pt clouds -> {npts * UP} pts, x4, per B200"


def make_clouds(clouds, npts, rank):
    import numpy as np
    from tools import synth    # input generation only (numpy)
    c = synth.make_clouds(clouds, npts, seed=2021 + rank * clouds)
    rng = np.random.default_rng(7 + rank)
    return np.stack([x[rng.permutation(npts)] for x in c], 0)


# ---- clocks during the timed region (B200_PROFILING.md recipe) -------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index, enabled=True):
        # One sampler per JOB (rank 0, all GPUs of the node in one nvidia-smi process): eight per-rank nvidia-smi loops at
        # 10 Hz measurably stalled kernel launches in every process (gaps of 3.4 ms per step at N = 8 against 1.7 ms at
        # N = 1, same kernels), i.e. the measurement disturbed the run.
        self.gpu, self.proc, self.lines, self.enabled = gpu_index, None, [], enabled

    def start(self):
        if not self.enabled:
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.enabled:
            return None
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.th.join(timeout=2)
        sm, mx, pw, reasons = [], [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])), mx.append(float(f[2])), pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        busy = [s for s, p in zip(sm, pw) if p > 0.5 * max(pw)] or sm
        return {"sm_mhz": statistics.median(busy), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---- the reference's CPU path (oracle port), timed on the host cores ----------------------------------------------
def cpu_reference_run(steps, warmup, clouds_per_step, npts):
    """Times oracle.puflow_oracle (the CPU restatement of the reference's Python: torch CPU ops + plain C for
    kNN / FPS / Chamfer) on `clouds_per_step` clouds of the bench workload per step.  Returns (points/s, ms/step,
    cores)."""
    import numpy as np
    import torch
    from oracle import puflow_oracle as po      # the one place bench.py executes oracle/: as the timed CPU baseline
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = torch.load(CKPT, map_location="cpu")
    orc = po.PUFlowOracle(sd)
    clouds = make_clouds(max(clouds_per_step, 1), npts, 0)
    n_out = npts * UP + N_OUTLIER

    def step():
        outs = []
        for c in clouds[:clouds_per_step]:
            pc = torch.from_numpy(c)[None]
            sr = po.upsample_cloud(orc, pc, n_out, UP, PATCH, 4)
            outs.append(po.remove_outliers(sr, pc, N_OUTLIER))
        return outs

    with torch.no_grad():
        for _ in range(warmup):
            step()
        t0 = time.perf_counter()
        for _ in range(steps):
            out = step()
        dt = time.perf_counter() - t0
    assert tuple(out[0].shape[-2:]) == (npts * UP, 3), out[0].shape
    pts = steps * clouds_per_step * npts * UP
    return pts / dt, dt / steps * 1e3, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    v, ms, cores = cpu_reference_run(args.steps, args.warmup, args.ref_clouds, args.npts)
    sample = (f"{args.ref_clouds} cloud(s) of the {args.clouds}-cloud batch per step, full pipeline "
              f"(patching, network, merge FPS, outlier removal), torch CPU ops + C kNN/FPS/Chamfer")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args.clouds, args.npts), "checkpoint": "pretrain/puflow-x4-pu1k.pt layout",
                   "sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


# ---- algorithmic work per launch of each kernel (DESIGN.md "Kernels" table) ----------------------------------------
def algorithmic_work(tag, Bc, npts):
    """(bound, amount per launch, unit) for a launch tagged `tag` at Bc clouds of npts points per step.
    Bytes are the compulsory HBM traffic of the launch (inputs read once + outputs written once); flops are the
    reference's own MACs x2 that the launch is responsible for (no factoring credit)."""
    n_patch = int(npts / PATCH * 4)
    Bp, M = Bc * n_patch, Bc * n_patch * PATCH          # patches, input points in patches
    if tag.startswith("knn_topk_k"):
        K = int(tag.split("_k")[-1])
        return "hbm", M * (12 + 4 * K), "B"
    if tag.startswith("edgeconv_g") or tag.startswith("edgeconv_tc_g"):
        g, c, k = (int(x[1:]) for x in tag.replace("edgeconv_tc_", "edgeconv_").split("_")[1:4])
        R = 4 * g + c
        return "tensor", 2.0 * (6 * g * g + 4 * g * c) * M * k, "FLOP"          # hidden-channel MACs of the dense block
    if tag.startswith("linear_k") or tag.startswith("linear_tc_k") or tag.startswith("linear_split_k"):
        f = tag.replace("linear_tc_", "linear_").replace("linear_split_", "linear_").split("_")
        K, N = int(f[1][1:]), int(f[2][1:])
        return "hbm", M * 4.0 * (K + N), "B"        # the pre-split input (fp16 hi + lo) is 4 B per element as well
    if tag.startswith("cond_chain_tc_c"):
        return "hbm", M * 4.0 * (int(tag.rsplit("_c", 1)[1]) + 64 + 8), "B"
    if tag in ("flow_cond", "flow_cond_tc"):
        return "hbm", 6 * M * 4.0 * (192 + 8), "B"
    if tag in ("flow_fwd", "flow_fwd_tc"):
        return "hbm", M * (24 + 4.0 * 6 * (64 + 8)), "B"
    if tag in ("flow_inv", "flow_inv_tc"):
        return "hbm", M * (UP * 24 + 4.0 * 6 * (64 + 8)), "B"
    if tag in ("interp", "interp_tc"):
        return "tensor", 2.0 * 81792 * 8 * M, "FLOP"
    if tag.startswith("fps_n"):
        n = int(tag[5:])
        return "hbm", Bc * n * 12.0, "B"
    if tag == "knn_patch":
        return "hbm", Bc * (npts * 12.0 + n_patch * (12 + 4 * PATCH)), "B"
    if tag == "chamfer_nn":
        return "hbm", Bc * ((npts * UP + N_OUTLIER) * 20.0 + npts * 12.0), "B"       # a -> b direction only
    return "hbm", 0.0, "B"


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm_gbs": float(p["hbm_gbs"]), "tensor_tflops": float(p.get("bf16_tflops_sustained", p["bf16_tflops"])),
                "source": "MEASURED_PEAKS.json (hbm copy; bf16 sustained, kernel timed inside a long step)"}
    except (OSError, KeyError, ValueError):
        return {"hbm_gbs": 6650.0, "tensor_tflops": 1590.0, "source": "fallback (B200_PROFILING.md)"}


# ---- this repo's CUDA path ---------------------------------------------------------------------------------------
def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the puflow_b200 path has no CPU fallback)")
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from puflow_b200 import ops                       # raises ImportError if the CUDA library is missing
    from modules.discrete.interpflow import PointInterpFlow
    from modules.discrete.upsample import upsample_array
    from modules.utils.patch import PatchHelper

    net = PointInterpFlow(3)
    net.load_state_dict(torch.load(CKPT, map_location="cpu"))
    net.set_to_initialized_state()
    net = net.to(dev).eval()
    helper = PatchHelper(PATCH, patch_expand_ratio=4)

    Bc, npts = args.clouds, args.npts
    n_out = npts * UP + N_OUTLIER
    host_in = torch.from_numpy(make_clouds(Bc, npts, rank)).pin_memory()
    host_out = torch.empty((Bc, npts * UP, 3), dtype=torch.float32).pin_memory()
    clouds = host_in.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    @torch.no_grad()
    def step():
        pred = helper.upsample(net, clouds, npoint=n_out, upratio=UP, jitter=False)
        return PatchHelper.remove_outliers(pred, clouds, N_OUTLIER)

    for _ in range(max(args.warmup, 3)):
        out = step()
    torch.cuda.synchronize(dev)
    assert tuple(out.shape) == (Bc, npts * UP, 3) and bool(torch.isfinite(out).all()), "bad output"

    # ---- timed region: exactly K steps, device-timed, per-launch events on ---------------------------------
    log = ops.LaunchLog(events=not args.no_breakdown)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    sampler = ClockSampler(",".join(str(i) for i in range(world)), enabled=(rank == 0))   # this job's GPUs (local ranks)
    sampler.start()
    barrier()
    # per-launch CUDA events: the network's come from inside the one C call (pf_profile_*), the caller-side kernels'
    # (FPS, patch kNN, Chamfer) from the Python wrappers -- the kernel-by-kernel Python path cost 0.6 ms per step of host
    # time at N = 1 and 3 ms at N = 8 (eight ranks on 16 host cores), which the timed region would have charged to the GPU
    prev = ops.set_launch_log(log)
    ops.c_profile(not args.no_breakdown)
    for e0, e1 in ev:
        flush.zero_()
        e0.record()
        step()
        e1.record()
    ops.set_launch_log(prev)
    barrier()
    c_times = ops.c_profile_read() if not args.no_breakdown else {}
    ops.c_profile(False)
    clocks = sampler.stop()
    t_ms = sum(e0.elapsed_time(e1) for e0, e1 in ev)
    tt = torch.tensor([t_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_ms = float(tt.item())
    pts_per_step = world * Bc * npts * UP
    value = pts_per_step * args.steps / (t_ms * 1e-3)

    # ---- per-kernel breakdown + roofline of the dominant kernel ----------------------------------------------
    breakdown, roofline = None, None
    if not args.no_breakdown:
        peaks = load_peaks()
        times = log.times_ms()
        for k, v in c_times.items():
            if k != "logp":                                   # two tiny reduction kernels, counted with the plumbing
                times.setdefault(k, []).extend(v)
        per = {k: (sum(v) / args.steps, len(v) // args.steps) for k, v in times.items()}
        tot = sum(v[0] for v in per.values())
        step_ms = t_ms / args.steps
        breakdown = {k: {"ms_per_step": round(v[0], 4), "launches_per_step": v[1], "share": round(v[0] / step_ms, 4)}
                     for k, v in sorted(per.items(), key=lambda kv: -kv[1][0])}
        breakdown["(torch plumbing + gaps)"] = {"ms_per_step": round(step_ms - tot, 4), "launches_per_step": None,
                                                "share": round((step_ms - tot) / step_ms, 4)}
        top = max(per, key=lambda k: per[k][0])
        bound, amount, unit = algorithmic_work(top, Bc, npts)
        dur_s = per[top][0] / per[top][1] * 1e-3
        if bound == "hbm":
            ach, peak, u = amount / dur_s / 1e9, peaks["hbm_gbs"], "GB/s"
        else:
            ach, peak, u = amount / dur_s / 1e12, peaks["tensor_tflops"], "TFLOP/s"
        roofline = {"kernel": top, "bound": bound, "achieved": ach, "peak": peak, "unit": u, "frac": ach / peak,
                    "traffic": None, "avg_launch_ms": dur_s * 1e3, "algorithmic_per_launch": amount,
                    "algorithmic_unit": unit, "peak_source": peaks["source"]}
        if bound == "tensor":
            roofline["fp32_ffma_peak_tflops"] = 74.4
            roofline["fp32_ffma_frac"] = ach / 74.4
        rp = os.path.join(ROOT, "profiles", "traffic.json")          # dram bytes per launch from the ncu --set full capture
        if os.path.exists(rp):
            try:
                roofline["traffic"] = json.load(open(rp)).get(top)
            except ValueError:
                pass
    gpu_launches = log.count

    # ---- network only (the five north_star subsystems on pre-extracted patches) ------------------------------
    with torch.no_grad():
        pcn, _, _ = PatchHelper.normalize_pc(clouds)
        patches = PatchHelper.extract_knn_patch(pcn, helper.knn, PATCH, 4).reshape(-1, PATCH, 3)
        patches, _, _ = PatchHelper.normalize_pc(patches)
        patches = patches.contiguous()
        net.sample(patches, UP)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        nrep = max(2, min(args.steps, 5))
        barrier()
        e0.record()
        for _ in range(nrep):
            net.sample(patches, UP)
        e1.record()
        barrier()
    tn = torch.tensor([e0.elapsed_time(e1) / nrep], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tn, op=dist.ReduceOp.MAX)
    net_ms = float(tn.item())
    network_only = {"value": world * patches.shape[0] * PATCH * UP / (net_ms * 1e-3), "unit": "generated points/s",
                    "ms_per_step": net_ms, "patches_per_gpu": int(patches.shape[0])}
    # the same network through the single C-ABI entry point (pf_upsample_patches), rank-local figure
    with torch.no_grad():
        eng = net.engine(dev)
        eng.forward_c(patches, UP)
        e0.record()
        for _ in range(nrep):
            eng.forward_c(patches, UP)
        e1.record()
        torch.cuda.synchronize()
    network_only["c_abi_one_call_ms"] = e0.elapsed_time(e1) / nrep

    # ---- e2e: host buffers in and out through the public API ----------------------------------------------------
    for _ in range(2):
        upsample_array(net, helper, host_in, UP, N_OUTLIER, None, dev, out=host_out)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        upsample_array(net, helper, host_in, UP, N_OUTLIER, None, dev, out=host_out)
    torch.cuda.synchronize(dev)
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    barrier()
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e = {"value": pts_per_step * args.steps / float(te.item()), "unit": UNIT,
           "h2d_bytes_per_step": host_in.numel() * 4, "d2h_bytes_per_step": host_out.numel() * 4,
           "ms_per_step": float(te.item()) / args.steps * 1e3}
    assert bool(torch.isfinite(host_out).all())

    # ---- CPU baseline on rank 0 at N=1 (bounded sample) ----------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, ms, cores = cpu_reference_run(steps=3, warmup=1, clouds_per_step=1, npts=npts)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "ms_per_cloud": ms,
               "sample": "1 cloud of the batch per run (full pipeline), 1 warm-up + 3 timed runs"}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": t_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(Bc, npts), "clouds_per_gpu": Bc, "points_per_cloud": npts,
                       "patches_per_gpu": int(patches.shape[0]), "up_ratio": UP, "num_outlier": N_OUTLIER,
                       "weights": "pretrain/puflow-x4-pu1k.pt (408-key layout)",
                       "l2": "256 MiB flush between steps; per-step working set is several GB (>> 126 MB L2)",
                       "sharding": f"{world} rank(s), clouds sharded, no collective on the data path"},
            "e2e": e2e, "gpu_launches": gpu_launches, "clocks": clocks, "roofline": roofline,
            "cpu_baseline": cpu, "network_only": network_only, "breakdown": breakdown,
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
