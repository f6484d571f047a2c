"""Farthest point sampling helpers with the call signatures of the reference's modules/utils/fps.py
(farthest_point_sampling :40-73, index_points :20-37), running on the cluster FPS kernel (pf_fps)."""
import torch
from torch import Tensor

from puflow_b200 import ops


def farthest_point_sampling(xyz: Tensor, npoint: int) -> Tensor:
    """xyz[B,N,3] CUDA -> int64 [B,npoint]; starts at index 0 like the reference (fps.py:54)."""
    return ops.fps(xyz.contiguous().float(), npoint).long()


def index_points(points: Tensor, idx: Tensor) -> Tensor:
    """points[B,N,C], idx[B,S] -> [B,S,C] (fps.py:20-37)."""
    B = points.shape[0]
    return points[torch.arange(B, device=points.device).view(-1, 1), idx.long()]
