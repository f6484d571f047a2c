"""One step of the bench workload (config 2: 128 x 2,048-pt clouds) through pf_upsample_clouds, after 2 warm-up steps:
the target of the ncu launch list / traffic capture (tools/gpu_profile_step.sh).  32 launches per step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from modules.discrete.interpflow import PointInterpFlow

clouds_n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
npts = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
dev = torch.device("cuda:0")
net = PointInterpFlow(3)
net.load_state_dict(torch.load(bench.CKPT, map_location="cpu"))
net.set_to_initialized_state()
net = net.to(dev).eval()
eng = net.engine(dev)
clouds = torch.from_numpy(bench.make_clouds(clouds_n, npts, 0)).to(dev)
n_patch = int(npts / bench.PATCH * 4)
for _ in range(3):
    out = eng.upsample_clouds(clouds, bench.UP, npts * bench.UP + bench.N_OUTLIER, bench.N_OUTLIER, bench.PATCH, n_patch)
torch.cuda.synchronize()
print("ok", tuple(out.shape))
