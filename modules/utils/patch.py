"""PatchHelper -- same class / method names / argument meaning as the reference's
modules/utils/patch.py:17-215, with knn_cuda.KNN, pointnet2_ops.furthest_point_sample/gather_operation and
the chamfer_3D extension replaced by puflow_b200 kernels (pf_knn_patch, pf_fps, pf_chamfer_nn) and the tensor
glue in between (normalise, patch gather, de-normalise + concat, merge gather, outlier selection) by the kernels of
csrc/caller.cu.  `upsample` with a PointInterpFlow and jitter=False is ONE C call (pf_upsample_clouds)."""
import torch
import torch.nn as nn
from torch import Tensor

from puflow_b200 import ops
from puflow_b200.sharding import patch_sharded


class KNN(object):
    """knn_cuda.KNN stand-in with the same call convention (transpose_mode=False): ref[B,C,N], query[B,C,M]
    -> (dist[B,k,M], idx[B,k,M] int64), L2 (sqrt) distances ascending, ties to the lower index."""

    def __init__(self, k: int, transpose_mode: bool = False):
        assert not transpose_mode
        self.k = k

    def __call__(self, ref: Tensor, query: Tensor):
        p = ref.transpose(1, 2).contiguous().float()
        q = query.transpose(1, 2).contiguous().float()
        idx = ops.knn_patch(q, p, self.k).long()                           # [B, M, k]
        B = p.shape[0]
        nb = p[torch.arange(B, device=p.device).view(-1, 1, 1), idx]       # [B, M, k, 3]
        d = q[:, :, None, :] - nb
        dist = torch.sqrt((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2])
        return dist.transpose(1, 2).contiguous(), idx.transpose(1, 2).contiguous()


# Upstream pointnet2_ops.furthest_point_sample never selects a candidate with |p|^2 <= 1e-3; the reference's own
# modules/utils/fps.py (the semantics followed by default, by the oracle and by the golden fixtures) has no such rule.  Set
# this to 1e-3 for the upstream rule: the FPS calls below then go through pf_fps_park_near_origin, and
# PatchHelper.upsample / upsample_array take the staged path (the one-call C entry point has the default rule built in).
FPS_SKIP_NEAR_ORIGIN = 0.0


def furthest_point_sample(xyz: Tensor, npoint: int) -> Tensor:
    """pointnet2_ops signature: xyz[B,N,3] -> int32 [B,npoint]."""
    return ops.fps(xyz.contiguous().float(), npoint, skip_near_origin=FPS_SKIP_NEAR_ORIGIN)


def gather_operation(features: Tensor, idx: Tensor) -> Tensor:
    """pointnet2_ops signature: features[B,C,N], idx[B,M] int32 -> [B,C,M]."""
    return torch.gather(features, 2, idx.long()[:, None, :].expand(-1, features.shape[1], -1))


# -----------------------------------------------------------------------------------------
class PatchHelper(object):

    def __init__(self, npoint_patch: int, patch_expand_ratio: float, extract='knn'):
        super(PatchHelper, self).__init__()
        self.__npoint_patch = npoint_patch
        self.__patch_expand_ratio = patch_expand_ratio
        if extract == 'knn':
            self.__extract_func_name__ = 'extract_knn_patch'
            self.knn = KNN(k=self.__npoint_patch, transpose_mode=False)

    @property
    def npoint_patch(self):
        return self.__npoint_patch

    @property
    def patch_expand_ratio(self):
        return self.__patch_expand_ratio

    def upsample(self, upsampler: nn.Module, pc: Tensor, npoint: int, upratio=None, jitter=False, **kwargs):
        """pc[B,N,3] -> [B,npoint,3]; reference patch.py:35-79."""
        B, N, C = pc.shape
        extract_func = getattr(self, self.__extract_func_name__)
        # the whole method as one C call (pf_upsample_clouds: same kernels in the same order, no Python in between) when
        # the upsampler is the drop-in PointInterpFlow and nothing below needs a Python hook
        eng = upsampler.engine(pc.device) if hasattr(upsampler, "engine") and pc.is_cuda else None
        if eng is not None and eng.flow_tc and not jitter and not kwargs and C == 3 and not ops.per_launch_events_wanted() \
                and not FPS_SKIP_NEAR_ORIGIN \
                and N >= self.__npoint_patch >= 16:
            n_patch = int(N / self.__npoint_patch * self.__patch_expand_ratio)
            if 1 <= n_patch <= N:
                return eng.upsample_clouds(pc, upratio or 4, npoint, 0, self.__npoint_patch, n_patch)

        pc, g_centroid, g_furthest_distance = PatchHelper.normalize_pc(pc)
        if jitter:
            pc = PatchHelper.jitter_perturbation_point_cloud(pc)

        patches = extract_func(pc, self.knn, self.__npoint_patch, self.__patch_expand_ratio)  # [B, n_patch, k1, 3]
        patches = patches.reshape(B, -1, self.__npoint_patch, C)

        predict_patches = PatchHelper._upsampling_patches(upsampler, patches, upratio, **kwargs)
        predict_pc = PatchHelper.merge_patches(predict_patches, npoint)       # [B, 3, npoint]

        predict_pc = predict_pc * g_furthest_distance + g_centroid.transpose(1, 2)
        predict_pc = predict_pc.transpose(1, 2).contiguous()
        return predict_pc  # [B, npoint, 3]

    @staticmethod
    def _upsampling_patches(upsampler: nn.Module, patches: Tensor, upratio=None, **kwargs):
        """reference patch.py:82-93 (name-mangled __upsampling_patches there).
        Addition: shard_patches=True under an initialised torch.distributed group splits the B * n_patch patches over
        the ranks (every rank must hold the same `patches`) and all_gathers the generated points -- the config-3
        single-large-cloud mode, puflow_b200/sharding.py; default False = the reference behaviour."""
        B, n_patch, k1, C = patches.shape
        patches = patches.reshape(B * n_patch, k1, C)
        patches, centroids, furthest_distance = PatchHelper.normalize_pc(patches)

        if kwargs.pop("shard_patches", False):
            predict_patches = patch_sharded(lambda p: upsampler.sample(p, upratio=(upratio or 4), **kwargs), patches)
        else:
            predict_patches = upsampler.sample(patches, upratio=(upratio or 4), **kwargs)  # [B * n_patch, k2, C]
        if predict_patches.is_cuda and C == 3:          # cat + scale + shift in one kernel (pf_denorm_concat)
            predict_patches = ops.denorm_concat(predict_patches.contiguous().float(), patches, centroids, furthest_distance)
        else:
            predict_patches = torch.cat([predict_patches, patches], dim=1)
            predict_patches = predict_patches * furthest_distance + centroids
        return predict_patches.reshape(B, n_patch, -1, C)

    @staticmethod
    def _extract_idx_patches(pc: Tensor, knn_searcher, npoint_patch: int, expand_ratio: float, seed_centroids_idx=None):
        """reference patch.py:96-109 -> (idx[B, n_patch, k] int64, n_patch)."""
        _, N, _ = pc.shape
        if seed_centroids_idx is None:
            n_patch = int(N / npoint_patch * expand_ratio)
            patch_centroids_idx = furthest_point_sample(pc, n_patch)  # [B, n_patch]
        else:
            _, n_patch = seed_centroids_idx.shape
            patch_centroids_idx = seed_centroids_idx
        pc32 = pc.contiguous().float()
        seeds = ops.gather_points(pc32, patch_centroids_idx.int().contiguous())
        idx = ops.knn_patch(seeds, pc32, npoint_patch)                                     # [B, n_patch, k]
        return idx.long(), n_patch

    @staticmethod
    def extract_knn_patch(pc: Tensor, knn_searcher, npoint_patch: int, expand_ratio: float, seed_centroids_idx=None):
        """pc[B,N,C] -> patches [B, n_patch, k, C]; reference patch.py:112-125."""
        B, _, C = pc.shape
        idx_b = torch.arange(B, device=pc.device).view(-1, 1)
        idx_patches, n_patch = PatchHelper._extract_idx_patches(pc, knn_searcher, npoint_patch, expand_ratio,
                                                                seed_centroids_idx)
        if pc.is_cuda and C == 3 and pc.dtype == torch.float32:
            patches = ops.gather_points(pc.contiguous(), idx_patches.flatten(start_dim=1).int().contiguous())
        else:
            patches = pc[idx_b, idx_patches.flatten(start_dim=1)]  # [B, n_patch * k, C]
        return patches.reshape(B, n_patch, npoint_patch, C)

    @staticmethod
    def fps(pc: Tensor, n_point: int, transpose=True):
        centroids_idx = furthest_point_sample(pc, n_point)
        centroids = gather_operation(pc.transpose(1, 2).contiguous(), centroids_idx)
        return centroids.transpose(1, 2).contiguous() if transpose is True else centroids

    @staticmethod
    def merge_patches(patches: Tensor, npoint: int, origins: Tensor = None):
        """patches[B, n_patch, k, 3] -> [B, 3, npoint] by FPS over all candidates; reference patch.py:142-158."""
        B, _, _, C = patches.shape
        if origins is None:
            patches = patches.reshape(B, -1, C).contiguous()
        else:
            patches = torch.cat([patches.reshape(B, -1, C), origins], dim=1).contiguous()
        idx_predict_pc = furthest_point_sample(patches, npoint)
        if patches.is_cuda and C == 3 and patches.dtype == torch.float32:
            return ops.gather_points(patches, idx_predict_pc).transpose(1, 2)  # [B, 3, npoint]
        return gather_operation(patches.transpose(1, 2).contiguous(), idx_predict_pc)  # [B, 3, npoint]

    @staticmethod
    def merge_pc(pc1: Tensor, pc2: Tensor, npoint: int):
        from modules.utils.fps import farthest_point_sampling, index_points
        tmp = torch.cat([pc1, pc2], dim=1)
        return index_points(tmp, farthest_point_sampling(tmp, npoint))

    @staticmethod
    def normalize_pc(pc: Tensor):
        """centroid / max-radius normalisation; reference patch.py:168-178.  [B,n,3] fp32 CUDA tensors go through
        pf_normalize_pc (one kernel); anything else (other widths, CPU tensors of host-side tools) through the torch
        expressions of the reference."""
        if pc.is_cuda and pc.dim() == 3 and pc.shape[2] == 3 and pc.dtype == torch.float32:
            return ops.normalize_pc(pc.contiguous())
        centroid = torch.mean(pc, dim=1, keepdim=True)
        pc = pc - centroid
        dist_square = torch.sum(pc ** 2, dim=-1, keepdim=True).sqrt()
        furthest_distance, _ = torch.max(dist_square, dim=1, keepdim=True)
        pc = pc / furthest_distance
        return pc, centroid, furthest_distance

    @staticmethod
    def jitter_perturbation_point_cloud(pc, sigma=0.010, clip=0.020):
        if sigma > 0:
            B, N, C = pc.shape
            assert (clip > 0)
            jittered_data = torch.clamp(sigma * torch.randn(B, N, C, device=pc.device), -1 * clip, clip)
            jittered_data[:, :, 3:] = 0
            jittered_data += pc
            return jittered_data
        return pc

    @staticmethod
    def remove_outliers(sr: Tensor, lr: Tensor, num_outliers: int) -> Tensor:
        """Drop the num_outliers generated points farthest from the input, keeping the survivors' order;
        reference patch.py:198-214 with chamfer_3DDist -> pf_chamfer_nn (only the sr -> lr direction is read) and
        argsort + nonzero -> pf_remove_outliers (a top-k select on the distances and an ordered compaction: the points
        dropped are the first num_outliers of a stable descending argsort; no sort, no host round trip)."""
        sr = sr.contiguous().float()
        dist1, _, _, _ = ops.chamfer_nn(sr, lr.contiguous().float(), both=False)
        return ops.remove_outliers(sr, dist1, num_outliers)
# -----------------------------------------------------------------------------------------
