"""torchrun check of the config-3 patch-sharded mode: one 5,000-pt cloud, its 78 patches split over the ranks, one
all_gather of the generated patches, merge FPS replicated.  Every rank compares with its own unsharded result.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dev_shard.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from tools import synth

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
dev = torch.device(f"cuda:{local}")
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
from modules.discrete.interpflow import PointInterpFlow
from modules.utils.patch import PatchHelper

net = PointInterpFlow(3)
net.load_state_dict(torch.load("tests/golden/puflow-x4-pu1k.pt", map_location="cpu"))
net.set_to_initialized_state()
net = net.to(dev).eval()
ph = PatchHelper(256, patch_expand_ratio=4)
pc = torch.from_numpy(synth.make_cloud(5000, 77, "torus"))[None].to(dev)


def run(shard):
    with torch.no_grad():
        pred = ph.upsample(net, pc, npoint=20024, upratio=4, jitter=False, shard_patches=shard)
        return PatchHelper.remove_outliers(pred, pc, 24)


for shard in (False, True):
    for _ in range(2):
        out = run(shard)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(5):
        out = run(shard)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    if shard:
        same = bool(torch.equal(out, ref))
        print(f"rank {rank}/{world}: sharded {dt * 1e3:.2f} ms/cloud, equal to unsharded: {same}", flush=True)
        assert same
    else:
        ref = out
        print(f"rank {rank}/{world}: unsharded {dt * 1e3:.2f} ms/cloud", flush=True)
if world > 1:
    dist.destroy_process_group()
