"""Mirror of the operator part of hf/core/feature_extractors/pointnet_util.py: what the RPN / RCNN PointNet++
set-abstraction and feature-propagation layers call before their MLPs (which are out of scope).

Same function names, argument order and return values as the reference (:23-62, :286-308); eager on torch
CUDA tensors.  Each function is a short sequence of libhfops calls with the glue fused into the kernels.
"""
from . import spatial
from .grouping.tf_grouping import group_point, group_point_concat, knn_point, query_ball_point
from .interpolate.tf_interpolate import three_interpolate, three_nn_weights
from .sampling.tf_sampling import farthest_point_sample, gather_point


def sample_and_group(npoint, radius, nsample, xyz, points, knn=False, use_xyz=True):
    """
    Input:
        npoint: int32
        radius: float32
        nsample: int32
        xyz: (batch_size, ndataset, 3) tensor
        points: (batch_size, ndataset, channel) tensor, if None will just use xyz as points
        knn: bool, if True use kNN instead of radius search
        use_xyz: bool, if True concat XYZ with local point features, otherwise just use point features
    Output:
        new_xyz: (batch_size, npoint, 3) tensor
        new_points: (batch_size, npoint, nsample, 3+channel) tensor
        idx: (batch_size, npoint, nsample) tensor, indices of local points as in ndataset points
        grouped_xyz: (batch_size, npoint, nsample, 3) tensor, normalized point XYZs
            (subtracted by seed point XYZ) in local regions

    Reference: hf/core/feature_extractors/pointnet_util.py:23-62.
    """
    new_xyz = gather_point(xyz, farthest_point_sample(npoint, xyz))
    if knn:
        _, idx = knn_point(nsample, xyz, new_xyz)
    else:
        idx, _ = query_ball_point(radius, nsample, xyz, new_xyz)
    if points is not None and use_xyz:
        new_points = group_point_concat(xyz, points, idx, new_xyz)  # [grouped_xyz - new_xyz, grouped_points]
        grouped_xyz = new_points[..., :3]
    else:
        grouped_xyz = group_point_concat(xyz, None, idx, new_xyz)
        new_points = group_point(points, idx) if points is not None else grouped_xyz
    return new_xyz, new_points, idx, grouped_xyz


def sample_and_group_msg(npoint, radius_list, nsample_list, xyz, points, use_xyz=True):
    """The operator part of pointnet_sa_module_msg (hf/core/feature_extractors/pointnet_util.py:223-286, the module the
    paper config uses: rpn_cars_pointnet_paper.config:62-111): ONE farthest-point sampling, then per scale i

        idx, _         = query_ball_point(radius_list[i], nsample_list[i], xyz, new_xyz)
        grouped_xyz    = group_point(xyz, idx) - new_xyz[:, :, None, :]
        grouped_points = concat([group_point(points, idx), grouped_xyz], -1)    # features FIRST, unlike sample_and_group
                         (points None: grouped_xyz; use_xyz False: group_point(points, idx))

    Returns new_xyz (b,npoint,3) and the list of grouped_points, (b,npoint,nsample_i,channel+3) each - the tensors the
    module's per-scale MLPs consume.  Both clouds are indexed once and the index is shared by every scale's search.
    """
    new_xyz = gather_point(xyz, farthest_point_sample(npoint, xyz))
    ix = (None, None)
    if spatial.applies(npoint, xyz.shape[1]):
        ix = (spatial.build_cloud_index(xyz), spatial.build_cloud_index(new_xyz))
    grouped = []
    for radius, nsample in zip(radius_list, nsample_list):
        idx, _ = query_ball_point(radius, nsample, xyz, new_xyz, index1=ix[0], index2=ix[1])
        if points is not None and not use_xyz:
            grouped.append(group_point(points, idx))
        else:
            grouped.append(group_point_concat(xyz, points, idx, new_xyz, xyz_last=True))
    return new_xyz, grouped


def interpolate_points(xyz1, xyz2, points2):
    """The operator part of pointnet_fp_module (hf/core/feature_extractors/pointnet_util.py:301-308):
    three_nn, inverse-distance weights (clamped at 1e-10), three_interpolate.
    xyz1 (b,n1,3) dense, xyz2 (b,n2,3) sparse, points2 (b,n2,c) -> interpolated_points (b,n1,c)."""
    ix = (None, None)
    if spatial.applies(xyz1.shape[1], xyz2.shape[1]):
        ix = (spatial.build_cloud_index(xyz1), spatial.build_cloud_index(xyz2))
    _, idx, weight = three_nn_weights(xyz1, xyz2, index1=ix[0], index2=ix[1])
    return three_interpolate(points2, idx, weight)
