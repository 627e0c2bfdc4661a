"""pose_estimation_b200 -- B200-native (sm_100a) pose recovery + ADD(-S) evaluation hot path.

Drop-in for the reference's call sites (see api.py); every numeric entry point launches
hand-written CUDA kernels from libposeb200.so through the C ABI in include/poseb200.h.
There is no CPU fallback: importing the API is cheap, calling it without the built library or
without a GPU raises.
"""
from .api import (BPnP, BPnPModle, Metric, TransparentMetric, PossLoss, PoseLoss, RansacConfig, batch_intrinsic_transform, cal_dis,  # noqa: F401
                  choose_ransac_batch, estimateSimilarityTransform, estimateSimilarityUmeyama, get_pose, get_xyz, knn,
                  vote_translation, trans_estimateSimilarityTransform, best_batch_fit_transform, icp_refine, uvd2cloud)
from .sweep import EvalSweep, format_report, reduce_table, shard_range, summarize  # noqa: F401

__all__ = ["Metric", "PoseLoss", "RansacConfig", "batch_intrinsic_transform", "cal_dis",
           "choose_ransac_batch", "get_pose", "knn", "vote_translation", "EvalSweep", "reduce_table",
           "shard_range", "summarize", "estimateSimilarityTransform", "estimateSimilarityUmeyama", "BPnP", "BPnPModle", "PossLoss", "TransparentMetric", "get_xyz",
           "trans_estimateSimilarityTransform", "best_batch_fit_transform", "icp_refine", "uvd2cloud", "format_report"]
