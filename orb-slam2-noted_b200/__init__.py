"""orb-slam2-noted_b200: B200-native ORB front end (extraction + Hamming matching).

Host-side mirror of the reference interface (include/ORBextractor.h, include/ORBmatcher.h of
QiuYue-bit/ORB-SLAM2-noted) over the C ABI in include/orbx.h (liborbx.so, hand-written sm_100a
CUDA).  There is no CPU fallback: importing works anywhere, but creating an extractor without the
built library or without a CUDA device raises.

The package directory has a hyphen (the repository's naming contract); import it with
    orbx = importlib.import_module("orb-slam2-noted_b200")
"""
from .binding import (KP_DTYPE, OrbxError, lib, lib_path, ORBextractor, descriptor_distance, hamming_best2,
                      stage_names, cvt_gray, undistort_keypoints, FMT_RGB, FMT_BGR, FMT_RGBA, FMT_BGRA)

from .matcher import (ORBmatcher, Vocabulary, FrameView, feature_vector, assign_features_to_grid, compute_stereo_matches,
                      compute_stereo_matches_batch, window_best, distinctive_descriptors, bow_score_l1, PROJ_QUERY_DTYPE)

from .kfdb import KeyFrameDatabase
__all__ = ["KeyFrameDatabase", "ORBmatcher", "Vocabulary", "FrameView", "feature_vector", "assign_features_to_grid", "compute_stereo_matches", "compute_stereo_matches_batch",
           "window_best", "distinctive_descriptors", "bow_score_l1", "PROJ_QUERY_DTYPE", "KP_DTYPE", "OrbxError", "lib", "lib_path", "ORBextractor", "descriptor_distance", "hamming_best2",
           "stage_names", "cvt_gray", "undistort_keypoints", "FMT_RGB", "FMT_BGR", "FMT_RGBA", "FMT_BGRA"]
