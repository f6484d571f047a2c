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
  e2e   : same metric through modules.discrete.upsample.upsample_arrays (the loop over batches a user runs) with HOST
          buffers: every step's pinned host -> device copy of the clouds and device -> pinned host copy of the
          128 x 8,192 x 3 result are inside the timed region; the copies of neighbouring steps run on copy streams
          under the kernels.  (One synchronous call, upsample_array: latency_one_cloud_ms.)
  network_only : the five north_star subsystems alone (PointInterpFlow.sample on pre-extracted patches).
  roofline : the kernel with the largest share of the step, algorithmic work / its average launch duration
             measured live with CUDA events on the launching stream (see DESIGN.md for the per-kernel figures);
  roofline_all : the same figure for every kernel of the step (kNN, the EdgeConv shapes, projections, conditioning
             chain, flow fwd / inv, interpolation, FPS, Chamfer, the caller-side glue kernels).
  parity   : one cloud of the timed batch against the CPU oracle, outside the timed region: kNN index mismatches,
             normwise error of the network output on that cloud's patches, Chamfer distance of the final clouds.
  latency_one_cloud_ms : BASELINE.json configs[0] on the GPU (one 2,048-pt cloud through upsample_array, host buffers).
  --config 3 : BASELINE.json configs[2], 5,000-pt clouds -> 20,000 points: (a) `--clouds` clouds per rank (cloud-sharded,
             weak scaling), (b) ONE cloud with its 78 patches split over the ranks (all_gather of the generated points
             timed separately; the merge FPS is sequential per cloud and replicated: Amdahl).
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
    ap.add_argument("--config", type=int, default=2, choices=[2, 3],
                    help="BASELINE.json config: 2 = 128 x 2,048-pt clouds per GPU (default, the metric's config); "
                         "3 = 5,000-pt clouds (8 per GPU unless --clouds is given) + one cloud patch-sharded")
    a = ap.parse_args()
    if a.config == 3:
        a.npts = 5000 if a.npts == N_PTS else a.npts
        a.clouds = 8 if a.clouds == 128 else a.clouds
    return a


DTYPE = "f32 (3xfp16-split tcgen05, fp32 accumulate)"


def workload_name(clouds, npts):
    shape = "PU1K-shape" if npts == N_PTS else ("PU-GAN-shape" if npts == 5000 else "synthetic")
    return f"{shape} batch: {clouds} synthetic {npts}-pt clouds -> {npts * UP} pts, x4, per B200"


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
def cpu_reference_run(steps, warmup, clouds_per_step, npts, want_outputs=False):
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
    if want_outputs:
        return pts / dt, dt / steps * 1e3, cores, orc, out
    return pts / dt, dt / steps * 1e3, cores


def parity_block(orc, oracle_out, gpu_out0, gpu_patches0, gpu_knn0, gpu_x0):
    """Cloud 0 of the timed batch, GPU path against the CPU oracle (checker), outside the timed region.
    kNN: index mismatches on the cloud's normalised patches (bit-exact bar); x: normwise relative error of the network
    output per patch (bar 1e-4); cd: Chamfer distance of the final clouds, mean(d1) + mean(d2) of squared distances on
    coordinates normalised like evaluate.py:230-235 (bar 1e-5).  The final clouds need not line up row by row: the merge
    FPS is chaotic under 1e-7 perturbations of near-duplicate candidates."""
    import numpy as np
    import torch
    from oracle import puflow_oracle as po
    with torch.no_grad():
        ref = orc.forward(torch.from_numpy(gpu_patches0), UP, return_all=True)
    knn_bad = int((ref["knn_idx"].numpy() != gpu_knn0).sum())
    xr = ref["x"].numpy()
    err = np.abs(gpu_x0 - xr).reshape(xr.shape[0], -1).max(1) / np.abs(xr).reshape(xr.shape[0], -1).max(1)
    a, b = gpu_out0.astype(np.float32), oracle_out.astype(np.float32)
    c = b.mean(0, keepdims=True)
    r = np.sqrt(((b - c) ** 2).sum(-1)).max()
    d1, d2, _, _ = po.chamfer_nn(((a - c) / r)[None], ((b - c) / r)[None])
    return {"cloud": 0, "patches": int(xr.shape[0]), "knn_index_mismatches": knn_bad, "knn_indices": int(gpu_knn0.size),
            "x_normwise_rel_err_max": float(err.max()), "x_tolerance": 1e-4,
            "chamfer_to_oracle": float(d1.mean() + d2.mean()), "chamfer_tolerance": 1e-5,
            "ok": bool(knn_bad == 0 and err.max() <= 1e-4 and d1.mean() + d2.mean() <= 1e-5)}


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
    n_out, ncand = npts * UP + N_OUTLIER, n_patch * PATCH * (UP + 1)
    if tag == "normalize_pc":
        return "hbm", Bc * npts * 24.0, "B"
    if tag == "gather_normalize":
        return "hbm", Bc * (npts * 12.0 + n_patch * PATCH * 16.0), "B"
    if tag == "denorm_concat":
        return "hbm", Bc * (n_patch * PATCH * 12.0 + 2 * ncand * 12.0 - n_patch * PATCH * 12.0), "B"
    if tag == "gather_points":      # two launches per step: the 32 seeds (negligible) and the merged cloud
        return "hbm", Bc * (n_out * 16.0 + n_out * 12.0) / 2, "B"
    if tag == "remove_outliers":
        return "hbm", Bc * (n_out * 16.0 + (n_out - N_OUTLIER) * 12.0), "B"
    return "hbm", 0.0, "B"


def pairs_per_launch(tag, Bc, npts):
    """Pair-distance count of the brute-force neighbour kernels (their binding resource is the ALU)."""
    n_patch = int(npts / PATCH * 4)
    if tag.startswith("knn_topk_k"):
        return float(Bc * n_patch * PATCH * PATCH)
    if tag == "chamfer_nn":          # what the brute-force kernel would compute: the grid-pruned kernel (csrc/nn.cu) visits ~2 % of them
        return float(Bc * (npts * UP + N_OUTLIER) * npts)
    if tag == "knn_patch":
        return float(Bc * n_patch * npts)
    return None


def roofline_entry(tag, avg_ms, Bc, npts, peaks, traffic):
    bound, amount, unit = algorithmic_work(tag, Bc, npts)
    dur_s = avg_ms * 1e-3
    if bound == "hbm":
        ach, peak, u = amount / dur_s / 1e9, peaks["hbm_gbs"], "GB/s"
    else:
        ach, peak, u = amount / dur_s / 1e12, peaks["tensor_tflops"], "TFLOP/s"
    e = {"kernel": tag, "bound": bound, "achieved": ach, "peak": peak, "unit": u, "frac": ach / peak,
         "traffic": traffic.get(tag), "avg_launch_ms": avg_ms, "algorithmic_per_launch": amount, "algorithmic_unit": unit}
    pr = pairs_per_launch(tag, Bc, npts)
    if pr:
        e["pair_distances_per_launch"] = pr
        if tag == "chamfer_nn":
            e["pair_distances_note"] = "brute-force count; the grid-pruned kernel evaluates a few per cent of them"
        else:
            e["tera_pairs_per_s"] = pr / dur_s / 1e12
    return e


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
    from modules.discrete.upsample import upsample_array, upsample_arrays
    from modules.utils.patch import PatchHelper

    net = PointInterpFlow(3)
    net.load_state_dict(torch.load(CKPT, map_location="cpu"))
    net.set_to_initialized_state()
    net = net.to(dev).eval()
    helper = PatchHelper(PATCH, patch_expand_ratio=4)
    eng = net.engine(dev)

    Bc, npts = args.clouds, args.npts
    n_out = npts * UP + N_OUTLIER
    n_patch = int(npts / PATCH * 4)
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
        # PatchHelper.upsample + remove_outliers (upsample.py:46-54) as the product path runs them: ONE C call
        return eng.upsample_clouds(clouds, UP, n_out, N_OUTLIER, PATCH, n_patch)

    for _ in range(max(args.warmup, 3)):
        out = step()
    torch.cuda.synchronize(dev)
    assert tuple(out.shape) == (Bc, npts * UP, 3) and bool(torch.isfinite(out).all()), "bad output"

    # ---- timed region: exactly K steps, device-timed, per-launch events recorded inside the C call ----------------
    log = ops.LaunchLog(events=False)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    sampler = ClockSampler(",".join(str(i) for i in range(world)), enabled=(rank == 0))   # this job's GPUs (local ranks)
    sampler.start()
    barrier()
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

    # ---- per-kernel breakdown, roofline of every kernel and of the dominant one ------------------------------------
    breakdown, roofline, roofline_all = None, None, None
    if not args.no_breakdown:
        peaks = load_peaks()
        traffic = {}
        rp = os.path.join(ROOT, "profiles", "traffic.json")          # dram bytes per launch from ncu (tools/ncu_traffic.py)
        if os.path.exists(rp):
            try:
                traffic = json.load(open(rp))
            except ValueError:
                pass
        per = {k: (sum(v) / args.steps, len(v) // args.steps) for k, v in c_times.items() if k != "logp"}
        tot = sum(v[0] for v in per.values())
        step_ms = t_ms / args.steps
        breakdown = {k: {"ms_per_step": round(v[0], 4), "launches_per_step": v[1], "share": round(v[0] / step_ms, 4)}
                     for k, v in sorted(per.items(), key=lambda kv: -kv[1][0])}
        breakdown["(launch gaps + logp)"] = {"ms_per_step": round(step_ms - tot, 4), "launches_per_step": None,
                                             "share": round((step_ms - tot) / step_ms, 4)}
        roofline_all = [roofline_entry(k, v[0] / v[1], Bc, npts, peaks, traffic)
                        for k, v in sorted(per.items(), key=lambda kv: -kv[1][0])]
        roofline = dict(roofline_all[0])
        roofline["peak_source"] = peaks["source"]
        if roofline["bound"] == "tensor":
            roofline["fp32_ffma_peak_tflops"] = 74.4
            roofline["fp32_ffma_frac"] = roofline["achieved"] / 74.4
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

    # ---- e2e: host buffers in and out through the public API (modules.discrete.upsample.upsample_arrays: the loop over
    # batches a user runs -- every step's pinned-host -> device copy, one C call replayed from a CUDA graph, and the
    # device -> host copy of its result are inside the timed region; the copies of neighbouring steps overlap the kernels) ----
    host_outs = [host_out, torch.empty_like(host_out).pin_memory()]
    for _ in upsample_arrays(net, helper, (host_in for _ in range(3)), UP, N_OUTLIER, None, dev, outs=host_outs):
        pass
    barrier()
    t0 = time.perf_counter()
    n_done = 0
    for res in upsample_arrays(net, helper, (host_in for _ in range(args.steps)), UP, N_OUTLIER, None, dev, outs=host_outs):
        n_done += 1                                  # (res has landed in pinned host memory when it is yielded)
    torch.cuda.synchronize(dev)
    assert n_done == args.steps
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    barrier()
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e = {"value": pts_per_step * args.steps / float(te.item()), "unit": UNIT,
           "h2d_bytes_per_step": host_in.numel() * 4, "d2h_bytes_per_step": host_out.numel() * 4,
           "ms_per_step": float(te.item()) / args.steps * 1e3}
    assert bool(torch.isfinite(host_outs[0]).all()) and bool(torch.isfinite(host_outs[1]).all())
    e2e["api"] = "upsample_arrays (pipelined loop over batches); a single synchronous upsample_array call: see latency_one_cloud_ms"

    # ---- one cloud alone (configs[0] on the GPU): latency through the public API, host buffers --------------------------
    latency = None
    if rank == 0:
        one_in, one_out = host_in[:1].clone().pin_memory(), torch.empty((1, npts * UP, 3)).pin_memory()
        latency = {}
        for name, g in (("eager", False), ("graph", True)):
            for _ in range(3):
                upsample_array(net, helper, one_in, UP, N_OUTLIER, None, dev, out=one_out, graph=g)
            ts = []
            for _ in range(15):
                t0 = time.perf_counter()
                upsample_array(net, helper, one_in, UP, N_OUTLIER, None, dev, out=one_out, graph=g)
                ts.append((time.perf_counter() - t0) * 1e3)
            latency[name] = statistics.median(ts)
        latency["note"] = f"one {npts}-pt cloud -> {npts * UP} pts, pinned host in / out, median of 15"

    # ---- config 3 (b): ONE cloud, its patches split over the ranks ----------------------------------------------------
    patch_sharded = None
    if args.config == 3:
        one = clouds[:1].contiguous() if rank == 0 else torch.empty((1, npts, 3), device=dev)
        if world > 1:
            dist.broadcast(one, 0)                     # every rank holds the same cloud (as when every rank reads the file)

        @torch.no_grad()
        def sharded():
            pred = helper.upsample(net, one, npoint=n_out, upratio=UP, jitter=False, shard_patches=True)
            return PatchHelper.remove_outliers(pred, one, N_OUTLIER)

        for _ in range(3):
            ref = sharded()
        full = eng.upsample_clouds(one, UP, n_out, N_OUTLIER, PATCH, n_patch)
        same = bool(torch.equal(ref, full))
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            sharded()
        e1.record()
        barrier()
        tp = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tp, op=dist.ReduceOp.MAX)
        # the exchange alone: all_gather of ceil(n_patch / world) x 1024 x 3 floats per rank
        ag_ms = 0.0
        if world > 1:
            per = -(-n_patch // world)
            buf = torch.zeros((per, PATCH * UP, 3), device=dev)
            parts = [torch.empty_like(buf) for _ in range(world)]
            for _ in range(3):
                dist.all_gather(parts, buf)
            barrier()
            e0.record()
            for _ in range(20):
                dist.all_gather(parts, buf)
            e1.record()
            barrier()
            ag_ms = e0.elapsed_time(e1) / 20
        # the part that does not shard: merge FPS of one cloud (+ Chamfer / outlier pass), from the one-call path's timers
        ops.c_profile(True)
        eng.upsample_clouds(one, UP, n_out, N_OUTLIER, PATCH, n_patch)
        ct = ops.c_profile_read()
        ops.c_profile(False)
        serial = sum(v[0] for k, v in ct.items() if k.startswith("fps_n") or k in ("chamfer_nn", "remove_outliers", "normalize_pc",
                                                                                 "knn_patch", "gather_points", "gather_normalize",
                                                                                 "denorm_concat"))
        total1 = sum(v[0] for v in ct.values())
        patch_sharded = {"ms_per_cloud": float(tp.item()), "points_per_s": npts * UP / (float(tp.item()) * 1e-3),
                         "all_gather_ms": ag_ms, "patches": n_patch, "patches_per_rank": -(-n_patch // world),
                         "bit_identical_to_unsharded": same,
                         "one_call_unsharded_ms": total1, "replicated_ms": serial,
                         "amdahl_serial_fraction": serial / total1,
                         "note": "seed FPS, patch kNN and the merge FPS (sequential per cloud) run on every rank; only the "
                                 "network shards"}

    # ---- CPU baseline + parity on rank 0 at N=1 (bounded sample) ---------------------------------------------------
    cpu, par = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, ms, cores, orc, oracle_out = cpu_reference_run(steps=3, warmup=1, clouds_per_step=1, npts=npts, want_outputs=True)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "ms_per_cloud": ms,
               "sample": "1 cloud of the batch per run (full pipeline), 1 warm-up + 3 timed runs"}
        with torch.no_grad():
            p0 = patches[:n_patch].contiguous()
            r0 = eng.forward(p0, UP, return_all=True)
        par = parity_block(orc, np.asarray(oracle_out[0]).reshape(-1, 3), out[0].cpu().numpy(), p0.cpu().numpy(),
                           r0["knn_idx"].cpu().numpy().astype(np.int64), r0["x"].cpu().numpy())

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": t_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
            "config": {"workload": workload_name(Bc, npts), "clouds_per_gpu": Bc, "points_per_cloud": npts,
                       "patches_per_gpu": int(patches.shape[0]), "up_ratio": UP, "num_outlier": N_OUTLIER,
                       "weights": "pretrain/puflow-x4-pu1k.pt (408-key layout)",
                       "l2": "256 MiB flush between steps; per-step working set is several GB (>> 126 MB L2)",
                       "sharding": f"{world} rank(s), clouds sharded, no collective on the data path"},
            "e2e": e2e, "gpu_launches": gpu_launches, "clocks": clocks, "roofline": roofline,
            "cpu_baseline": cpu, "parity": par, "latency_one_cloud_ms": latency, "network_only": network_only,
            "breakdown": breakdown, "roofline_all": roofline_all,
        }
        if patch_sharded is not None:
            line["patch_sharded_one_cloud"] = patch_sharded
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
