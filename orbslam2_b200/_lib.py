"""ctypes loader for liborbx.so — the C-ABI CUDA library declared in include/orbx.h. No fallback of any kind:
if the library is missing it is built with nvcc; if that fails, or no CUDA device exists, calls raise."""
import ctypes as C
import os

from . import build as _build

_LIB = None

ORBX_OK, ORBX_EINVAL, ORBX_ESHAPE, ORBX_ECAP, ORBX_ECUDA, ORBX_ESTATE = range(6)
MEM_HOST, MEM_DEVICE = 0, 1
FMT_GRAY8, FMT_RGB8, FMT_BGR8, FMT_RGBA8, FMT_BGRA8 = range(5)
_NAMES = {1: "ORBX_EINVAL", 2: "ORBX_ESHAPE", 3: "ORBX_ECAP", 4: "ORBX_ECUDA", 5: "ORBX_ESTATE"}


class OrbxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{_NAMES.get(code, code)}: {msg}")
        self.code = code


class OrbxParams(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scaleFactor", C.c_float), ("nlevels", C.c_int), ("iniThFAST", C.c_int),
                ("minThFAST", C.c_int), ("patchSize", C.c_int), ("halfPatchSize", C.c_int), ("edgeThreshold", C.c_int)]


# every symbol include/orbx.h declares (tests check the export table against this list and the header)
SYMBOLS = ["orbx_create", "orbx_destroy", "orbx_set_memory_space", "orbx_extract", "orbx_extract_batch",
           "orbx_extract_batch_submit", "orbx_extract_batch_wait", "orbx_extract_batch_multi", "orbx_set_input_format", "orbx_set_rectify_maps",
           "orbx_undistort_keypoints",
           "orbx_max_keypoints", "orbx_set_chunk", "orbx_level_count", "orbx_tables", "orbx_level_size", "orbx_pyramid_level",
           "orbx_blurred_level", "orbx_pyramid_bytes", "orbx_pyramid_all", "orbx_candidates", "orbx_stereo_match", "orbx_stereo_match_batch", "orbm_create",
           "orbm_destroy", "orbm_set_memory_space", "orbm_set_train", "orbm_knn2", "orbm_set_knn_mode", "orbm_last_knn_path", "orbm_knn2_static", "orbm_window_search", "orbm_distance",
           "orbm_search_by_projection", "orbm_search_by_bow", "orbm_search_by_bow_batch", "orbm_fuse_search", "orbm_keyframe_hint", "orbm_keyframe_cache_hits", "orbm_search_kf_projection", "orbm_search_for_initialization", "orbm_search_for_triangulation", "orbm_search_for_triangulation_batch", "orbm_search_local_points", "orbm_search_stats", "orbm_distinctive_descriptors",
           "orbx_stream", "orbm_stream", "orbx_launch_count", "orbm_launch_count", "orbx_stage_ms", "orbx_set_profiling",
           "orbv_create", "orbv_destroy", "orbv_info", "orbv_set_memory_space", "orbv_transform_batch", "orbv_score_batch",
           "orbv_stream", "orbv_launch_count", "orbx_last_error_string", "orbx_version"]


class ProjParams(C.Structure):
    """orbm_proj_params (include/orbx.h)."""
    _fields_ = [("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("bf", C.c_float),
                ("minX", C.c_float), ("maxX", C.c_float), ("minY", C.c_float), ("maxY", C.c_float),
                ("grid_w_inv", C.c_float), ("grid_h_inv", C.c_float), ("nlevels", C.c_int),
                ("scale", C.c_float * 16), ("Rcw", C.c_float * 9), ("tcw", C.c_float * 3),
                ("forward", C.c_int), ("backward", C.c_int), ("th", C.c_float), ("th_high", C.c_int),
                ("check_orientation", C.c_int), ("skip_depth_test", C.c_int)]


class LastFrameArgs(C.Structure):
    _fields_ = [("n", C.c_int), ("world_pos", C.c_void_p), ("flags", C.c_void_p), ("octave", C.c_void_p),
                ("angle", C.c_void_p), ("desc", C.c_void_p)]


class MapPointsArgs(C.Structure):
    """orbm_map_points (include/orbx.h)."""
    _fields_ = [("n", C.c_int), ("proj_x", C.c_void_p), ("proj_y", C.c_void_p), ("proj_xr", C.c_void_p),
                ("level", C.c_void_p), ("view_cos", C.c_void_p), ("flags", C.c_void_p), ("desc", C.c_void_p)]


class BowSide(C.Structure):
    """orbm_bow_side (include/orbx.h)."""
    _fields_ = [("n", C.c_int), ("eligible", C.c_void_p), ("angle", C.c_void_p), ("desc", C.c_void_p),
                ("n_nodes", C.c_int), ("node_id", C.c_void_p), ("node_off", C.c_void_p), ("node_feat", C.c_void_p)]


class Triangulation(C.Structure):
    """orbm_triangulation (include/orbx.h)."""
    _fields_ = [("pt_a", C.c_void_p), ("pt_b", C.c_void_p), ("octave_b", C.c_void_p), ("stereo_a", C.c_void_p),
                ("stereo_b", C.c_void_p), ("F12", C.c_float * 9), ("ex", C.c_float), ("ey", C.c_float),
                ("nlevels", C.c_int), ("scale_factors", C.c_float * 16), ("level_sigma2", C.c_float * 16)]


class FusePoints(C.Structure):
    """orbm_fuse_points (include/orbx.h)."""
    _fields_ = [("n", C.c_int), ("proj_x", C.c_void_p), ("proj_y", C.c_void_p), ("proj_xr", C.c_void_p),
                ("level", C.c_void_p), ("valid", C.c_void_p), ("desc", C.c_void_p)]


class CurFrameArgs(C.Structure):
    _fields_ = [("n", C.c_int), ("keys_un", C.c_void_p), ("u_right", C.c_void_p), ("desc", C.c_void_p),
                ("taken", C.c_void_p)]


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    path = _build.LIB
    if not os.path.exists(path):
        _build.build()
    L = C.CDLL(path)
    vp, ci, cf, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    L.orbx_create.argtypes = [C.POINTER(OrbxParams), ci, C.POINTER(vp)]
    L.orbx_destroy.argtypes = [vp]
    L.orbx_set_memory_space.argtypes = [vp, ci, ci]
    L.orbx_extract.argtypes = [vp, vp, ci, ci, sz, vp, vp, ci, vp]
    L.orbx_extract_batch.argtypes = [vp, vp, ci, ci, ci, sz, sz, vp, vp, ci, vp]
    L.orbx_extract_batch_submit.argtypes = [vp, vp, ci, ci, ci, sz, sz, vp, vp, ci]
    L.orbx_extract_batch_wait.argtypes = [vp, vp]
    L.orbx_pyramid_bytes.argtypes = [vp, ci, vp]
    L.orbx_pyramid_all.argtypes = [vp, ci, vp, sz, ci, vp]
    L.orbx_extract_batch_multi.argtypes = [vp, ci, vp, ci, ci, ci, sz, sz, vp, vp, ci, vp]
    L.orbx_set_input_format.argtypes = [vp, ci]
    L.orbx_set_rectify_maps.argtypes = [vp, vp, vp, ci, ci]
    L.orbx_undistort_keypoints.argtypes = [vp, vp, vp, ci, vp, ci]
    L.orbx_max_keypoints.argtypes = [vp, ci, ci, C.POINTER(ci)]
    L.orbx_set_chunk.argtypes = [vp, ci]
    L.orbx_level_count.argtypes = [vp]
    L.orbx_tables.argtypes = [vp] + [vp] * 6
    L.orbx_level_size.argtypes = [vp, ci, C.POINTER(ci), C.POINTER(ci)]
    L.orbx_pyramid_level.argtypes = [vp, ci, ci, vp, sz, ci]
    L.orbx_blurred_level.argtypes = [vp, ci, ci, vp, sz]
    L.orbx_candidates.argtypes = [vp, ci, ci, vp, ci, C.POINTER(ci)]
    L.orbx_stereo_match.argtypes = [vp, vp, ci, cf, cf, cf, vp, vp]
    L.orbx_stereo_match_batch.argtypes = [vp, vp, cf, cf, cf, vp, vp, ci]
    L.orbm_create.argtypes = [ci, C.POINTER(vp)]
    L.orbm_destroy.argtypes = [vp]
    L.orbm_set_memory_space.argtypes = [vp, ci, ci]
    L.orbm_set_train.argtypes = [vp, vp, ci]
    L.orbm_knn2.argtypes = [vp, vp, ci, vp, ci, vp, vp, vp]
    L.orbm_knn2_static.argtypes = [ci, vp, vp, ci, vp, ci, vp, vp, vp]
    L.orbm_set_knn_mode.argtypes = [vp, ci]
    L.orbm_last_knn_path.argtypes = [vp]
    L.orbm_window_search.argtypes = [vp, vp, ci, vp, vp, vp, vp, ci, vp, vp, vp, vp]
    L.orbm_distance.argtypes = [vp, vp, vp, ci, vp]
    L.orbm_search_by_projection.argtypes = [vp, C.POINTER(ProjParams), C.POINTER(LastFrameArgs), C.POINTER(CurFrameArgs),
                                            vp, vp, vp, C.POINTER(ci)]
    L.orbm_search_by_bow.argtypes = [vp, C.POINTER(BowSide), C.POINTER(BowSide), ci, ci, cf, ci, ci, ci, vp, C.POINTER(ci)]
    L.orbm_search_for_triangulation.argtypes = [vp, C.POINTER(BowSide), C.POINTER(BowSide), ci, ci, C.POINTER(Triangulation),
                                                ci, ci, vp, C.POINTER(ci)]
    L.orbm_fuse_search.argtypes = [vp, C.POINTER(ProjParams), vp, ci, C.POINTER(FusePoints), C.POINTER(CurFrameArgs), vp, vp]
    L.orbm_search_kf_projection.argtypes = [vp, C.POINTER(ProjParams), C.POINTER(FusePoints), C.POINTER(CurFrameArgs), vp, vp, C.POINTER(ci)]
    L.orbm_search_for_initialization.argtypes = [vp, C.POINTER(ProjParams), cf, ci, ci, ci, vp, vp, vp,
                                                 C.POINTER(CurFrameArgs), vp, C.POINTER(ci)]
    L.orbm_search_by_bow_batch.argtypes = [vp, ci, vp, vp, vp, vp, cf, ci, ci, ci, vp, vp]
    L.orbm_search_for_triangulation_batch.argtypes = [vp, ci, vp, vp, vp, vp, vp, ci, ci, vp, vp]
    L.orbm_search_stats.argtypes = [vp, C.POINTER(ci), C.POINTER(C.c_int64)]
    L.orbm_distinctive_descriptors.argtypes = [vp, vp, vp, ci, vp, vp]
    L.orbm_search_local_points.argtypes = [vp, C.POINTER(ProjParams), cf, C.POINTER(MapPointsArgs), C.POINTER(CurFrameArgs),
                                           vp, vp, C.POINTER(ci)]
    L.orbx_stream.argtypes = [vp, C.POINTER(vp)]
    L.orbm_stream.argtypes = [vp, C.POINTER(vp)]
    L.orbx_launch_count.argtypes = [vp]; L.orbx_launch_count.restype = C.c_int64
    L.orbm_launch_count.argtypes = [vp]; L.orbm_launch_count.restype = C.c_int64
    L.orbv_create.argtypes = [C.c_char_p, sz, ci, C.POINTER(vp)]
    L.orbv_destroy.argtypes = [vp]
    L.orbv_info.argtypes = [vp, vp, vp, vp, vp]
    L.orbv_set_memory_space.argtypes = [vp, ci, ci]
    L.orbv_transform_batch.argtypes = [vp, vp, vp, ci, ci, ci, ci] + [vp] * 7
    L.orbv_score_batch.argtypes = [vp, vp, vp, ci, vp, vp, vp, ci, vp]
    L.orbv_stream.argtypes = [vp, C.POINTER(vp)]
    L.orbv_launch_count.argtypes = [vp]; L.orbv_launch_count.restype = C.c_int64
    L.orbx_stage_ms.argtypes = [vp, vp]
    L.orbx_set_profiling.argtypes = [vp, ci]
    L.orbx_last_error_string.restype = C.c_char_p
    L.orbx_version.restype = C.c_char_p
    _LIB = L
    return L


def check(rc):
    if rc != ORBX_OK:
        raise OrbxError(rc, lib().orbx_last_error_string().decode())
