"""Discrete PU-Flow x4 upsampling entry point -- same `upsampling()` signature, CLI flags and file
formats as the reference's modules/discrete/upsample.py:20-86 (whitespace .xyz in, '%.6f' .xyz out,
same file name), running on the puflow_b200 kernels.

Additions that keep the reference behaviour as the default:
  --device cuda:N      (the reference hard-codes cuda:0, upsample.py:27)
  torchrun / WORLD_SIZE > 1: the file list is sharded round-robin over ranks, one process per GPU,
  no collective (every rank writes its own output files, exactly like the reference's per-file loop).
"""
import os
import sys
import torch
import numpy as np

from argparse import ArgumentParser
from typing import List
from pathlib import Path

sys.path.append(os.getcwd())

from modules.utils.patch import PatchHelper
from puflow_b200 import ops
from puflow_b200.xyzio import load_xyz, save_xyz
from modules.discrete.interpflow import PointInterpFlow


# -----------------------------------------------------------------------------------------
@torch.no_grad()
def upsampling(data_paths: List[str], target_path: str, checkpoint_path: str, up_ratio: int, num_outlier: int,
               num_patch: int, num_upsampling: int = None, seed=None, device=None, batch_clouds: int = 128):

    if seed is not None:
        np.random.seed(seed)
        torch.random.manual_seed(seed)
        torch.cuda.manual_seed(seed)

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if device is None:
        device = torch.device(f"cuda:{int(os.environ.get('LOCAL_RANK', 0))}")
    device = torch.device(device)
    torch.cuda.set_device(device)

    pc_channel = 3
    network = PointInterpFlow(pc_channel)

    network.load_state_dict(torch.load(checkpoint_path, map_location="cpu"))
    network.set_to_initialized_state()
    network = network.to(device)
    network.eval()
    patch_helper = PatchHelper(num_patch, patch_expand_ratio=4)

    paths = sorted(data_paths)[rank::world] if world > 1 else list(data_paths)
    # Files are read (and permuted: one torch.randperm per file, in list order, like the reference's loop) first, then
    # clouds of the same size go through the GPU together: a 2,048-point cloud alone is 32 patches and an 8,216-round FPS
    # on one SM, a batch of them fills the device.  Outputs are the reference's: same file names, '%.6f' rows.
    clouds = []
    for path in paths:
        pt_input = load_xyz(path)                       # np.loadtxt(path, dtype=np.float32), upsample.py:42
        clouds.append(pt_input[torch.randperm(pt_input.shape[0]).numpy()])
    by_size = {}
    for i, c in enumerate(clouds):
        by_size.setdefault(c.shape[0], []).append(i)
    chunks = [members[j:j + batch_clouds] for members in by_size.values() for j in range(0, len(members), batch_clouds)]
    batches = (torch.from_numpy(np.stack([clouds[i] for i in chunk], 0)).pin_memory() for chunk in chunks)
    # pipelined: batch i's result is copied out and its files are written while batch i + 1 is on the GPU
    for chunk, pred in zip(chunks, upsample_arrays(network, patch_helper, batches, up_ratio, num_outlier, num_upsampling, device)):
        for i, p in zip(chunk, pred):
            save_xyz(Path(target_path) / os.path.split(paths[i])[1], p)  # np.savetxt(..., fmt='%.6f'), upsample.py:57


# -----------------------------------------------------------------------------------------
def _check_clouds(clouds, who):
    src = torch.from_numpy(clouds) if isinstance(clouds, np.ndarray) else clouds
    if src.dim() != 3 or src.shape[2] != 3 or src.dtype != torch.float32:
        raise ValueError(f"{who}: expected fp32 clouds[B,N,3], got {tuple(src.shape)} {src.dtype}")
    return src


def _upsample_device(network, patch_helper, pt_input, up_ratio, num_outlier, num_upsampling, device, graph):
    """The device part of the per-file body of the reference's loop (upsample.py:46-54) for clouds[B,N,3] on the device:
    PatchHelper.upsample + remove_outliers, on the current stream; returns a device tensor the caller owns."""
    n_pts = pt_input.shape[1] * up_ratio if num_upsampling is None else num_upsampling
    n_drop = num_outlier or 0
    eng = network.engine(device) if hasattr(network, "engine") else None
    k = patch_helper.npoint_patch
    n_patch = int(pt_input.shape[1] / k * patch_helper.patch_expand_ratio)
    from modules.utils import patch as _patch
    if eng is not None and eng.flow_tc and pt_input.shape[1] >= k >= 16 and 1 <= n_patch <= pt_input.shape[1] \
            and not ops.per_launch_events_wanted() and not _patch.FPS_SKIP_NEAR_ORIGIN:
        # the whole per-file body of the reference's loop in one C call, replayed from a CUDA graph per shape
        return eng.upsample_clouds(pt_input, up_ratio, n_pts + n_drop, n_drop, k, n_patch, graph=graph)
    pred = patch_helper.upsample(network, pt_input, npoint=n_pts + n_drop, upratio=up_ratio, jitter=False)
    if n_drop > 0:
        pred = PatchHelper.remove_outliers(pred, pt_input, n_drop)
    return pred


@torch.no_grad()
def upsample_array(network, patch_helper, clouds, up_ratio: int = 4, num_outlier: int = 24,
                   num_upsampling: int = None, device=None, out: torch.Tensor = None, graph: bool = True):
    """Host-buffer form of the per-file body of the reference's loop (upsample.py:42-56): clouds[B,N,3] fp32 in HOST
    memory (numpy array or CPU tensor, pinned or not) -> upsampled clouds [B, N*up_ratio (or num_upsampling), 3] in
    host memory (numpy; or written into the pinned CPU tensor `out`).  The host->device copy, the whole
    PatchHelper.upsample + remove_outliers pipeline and the device->host copy run on the current stream; the call
    returns after the result has landed in host memory.  With the drop-in PointInterpFlow the device part is ONE C call
    (pf_upsample_clouds) replayed from a CUDA graph captured per cloud shape (graph=False launches it directly)."""
    device = torch.device(device) if device is not None else next(network.parameters()).device
    src = _check_clouds(clouds, "upsample_array")
    pt_input = src.to(device, non_blocking=True).contiguous()
    pred = _upsample_device(network, patch_helper, pt_input, up_ratio, num_outlier, num_upsampling, device, graph)
    if out is not None:
        out.copy_(pred, non_blocking=True)
        torch.cuda.current_stream(device).synchronize()
        return out
    return pred.cpu().numpy()


@torch.no_grad()
def upsample_arrays(network, patch_helper, batches, up_ratio: int = 4, num_outlier: int = 24, num_upsampling: int = None,
                    device=None, outs=None, graph: bool = True):
    """upsample_array over a sequence of host batches, pipelined: a generator that yields one result per batch, in order.
    The kernels of batch i run on the current stream while batch i + 1's host->device copy and batch i - 1's
    device->host copy run on two copy streams, so a loop over files / batches (the reference's, upsample.py:40-57) costs
    the kernels alone once it is rolling.  `batches`: iterable of clouds[B,N,3] fp32 host arrays (pinned CPU tensors
    overlap; pageable memory and numpy arrays work, their copies just synchronise).  `outs`: optional sequence of pinned
    CPU tensors used round-robin for the results (at least two; a yielded tensor is overwritten `len(outs)` batches
    later); without it every result is a fresh numpy array.  Results are identical to upsample_array's."""
    device = torch.device(device) if device is not None else next(network.parameters()).device
    cur = torch.cuda.current_stream(device)
    s_in, s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
    if outs is not None and len(outs) < 2:
        raise ValueError("upsample_arrays: outs needs at least two buffers (one is being filled while the previous is read)")

    def stage_in(clouds):
        src = _check_clouds(clouds, "upsample_arrays")
        with torch.cuda.stream(s_in):
            t = src.to(device, non_blocking=True).contiguous()
            ev = torch.cuda.Event()
            ev.record(s_in)
        return t, ev

    it = iter(batches)
    nxt = next(it, None)
    staged = stage_in(nxt) if nxt is not None else None
    pending, i = None, 0
    while staged is not None:
        pt_input, ev_in = staged
        cur.wait_event(ev_in)
        pt_input.record_stream(cur)
        pred = _upsample_device(network, patch_helper, pt_input, up_ratio, num_outlier, num_upsampling, device, graph)
        ready = torch.cuda.Event()
        ready.record(cur)
        nxt = next(it, None)                        # batch i + 1 goes up while batch i computes
        staged = stage_in(nxt) if nxt is not None else None
        host = outs[i % len(outs)] if outs is not None else torch.empty(pred.shape, dtype=pred.dtype, pin_memory=True)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ready)
            host.copy_(pred, non_blocking=True)
            pred.record_stream(s_out)
            done = torch.cuda.Event()
            done.record(s_out)
        if pending is not None:
            pending[0].synchronize()
            yield pending[1] if outs is not None else pending[1].numpy()
        pending, i = (done, host), i + 1
    if pending is not None:
        pending[0].synchronize()
        yield pending[1] if outs is not None else pending[1].numpy()


# -----------------------------------------------------------------------------------------
if __name__ == "__main__":

    # usage: python modules/discrete/upsample.py --source=path/to/input --target=path/to/output --checkpoint=path/to/checkpoint --up_ratio=4 --num_out=20000

    parser = ArgumentParser()
    parser.add_argument('--source', type=str, help='Path of input directory')
    parser.add_argument('--target', type=str, help='Path of output directory')
    parser.add_argument('--seed', type=int, default=2021)
    parser.add_argument('--checkpoint', type=str, help='Path of checkpoint')
    parser.add_argument('--up_ratio', type=int, help='upsampling ratio', default=4)
    parser.add_argument('--num_patch', type=int, help='number of point in each patch', default=256)  # 192
    parser.add_argument('--num_out', type=int, default=None, help='number of point of output point cloud')
    parser.add_argument('--device', type=str, default=None, help='CUDA device (default cuda:LOCAL_RANK)')
    parser.add_argument('--batch_clouds', type=int, default=128, help='same-sized clouds per GPU batch')

    args = parser.parse_args()

    os.makedirs(args.target, exist_ok=True)

    data_paths = []
    for root, dirs, files in os.walk(args.source):
        data_paths.extend([os.path.join(root, f) for f in files if '.xyz' in f])

    upsampling(data_paths, args.target, args.checkpoint, up_ratio=args.up_ratio,
        num_outlier=24, num_patch=args.num_patch, num_upsampling=args.num_out, seed=args.seed, device=args.device,
        batch_clouds=args.batch_clouds)
