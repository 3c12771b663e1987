"""ctypes bindings of include/hg_b200.h (libhg_b200.so). No compute happens here."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HG_LIB_PATH") or os.path.join(_HERE, "libhg_b200.so")  # HG_LIB_PATH: A/B of two builds
_LIB = None

vp = C.c_void_p
c_i = C.c_int
c_i64 = C.c_int64
c_f = C.c_float
c_d = C.c_double

KINDS = {
    "occupied": 0, "interpolated_occupied": 1, "tsdf": 2, "interpolated_tsdf": 3,
    "interpolated_tsdf_per_point": 4, "multires_tsdf": 5, "interpolated_multires_tsdf": 6,
    "interpolated_multires_tsdf_per_point": 7,
}


class HgError(RuntimeError):
    pass


class FcsmOptions(C.Structure):
    _fields_ = [("branch_and_bound_depth", C.c_int32), ("full_resolution_depth", C.c_int32),
                ("use_rotational_scan_matcher", C.c_int32), ("min_rotational_score", c_d),
                ("min_low_resolution_score", c_d), ("linear_xy_search_window", c_d),
                ("linear_z_search_window", c_d), ("angular_search_window", c_d),
                ("float_flavor", C.c_int32), ("strict_visit_order", C.c_int32)]


class NodeData(C.Structure):
    _fields_ = [("high_resolution_point_cloud", vp), ("num_high_resolution_points", C.c_int32),
                ("low_resolution_point_cloud", vp), ("num_low_resolution_points", C.c_int32),
                ("rotational_scan_matcher_histogram", vp), ("histogram_size", C.c_int32),
                ("gravity_alignment", c_d * 4)]


class MatchResult(C.Structure):
    _fields_ = [("found", C.c_int32), ("score", c_f), ("pose_translation", c_d * 3),
                ("pose_rotation", c_d * 4), ("rotational_score", c_f), ("low_resolution_score", c_f),
                ("best_scan_index", C.c_int32), ("best_offset", C.c_int32 * 3),
                ("num_discrete_scans", C.c_int32), ("num_candidates_scored", c_i64),
                ("num_candidates_visited", c_i64), ("num_score_launches", C.c_int32),
                ("num_lowest_resolution_candidates", c_i64), ("lowest_resolution_kernel_ms", c_f),
                ("best_ancestor_score", c_f), ("best_ancestor_index", c_i64), ("dive_bound", c_f),
                ("order_unresolved", C.c_int32)]


class OccupancyInserterOptions(C.Structure):
    _fields_ = [("hit_probability", c_d), ("miss_probability", c_d), ("num_free_space_voxels", C.c_int32),
                ("reserved", C.c_int32)]


class TsdfInserterOptions(C.Structure):
    _fields_ = [("relative_truncation_distance", c_d), ("maximum_weight", c_d),
                ("weight_function_epsilon", c_d), ("weight_function_sigma", c_d), ("min_range", c_d),
                ("max_range", c_d), ("insertion_ratio", c_d), ("num_free_space_voxels", C.c_int32),
                ("reserved", C.c_int32)]


class GridProtoInfo(C.Structure):
    _fields_ = [("resolution", c_f), ("relative_truncation_distance", c_f), ("max_weight", c_f),
                ("is_tsdf", C.c_int32), ("num_voxels", c_i64)]


class SubmapInfo(C.Structure):
    _fields_ = [("trajectory_id", C.c_int32), ("submap_index", C.c_int32), ("local_pose", c_d * 7),
                ("num_range_data", C.c_int32), ("finished", C.c_int32), ("high_is_tsdf", C.c_int32),
                ("low_is_tsdf", C.c_int32), ("histogram_size", C.c_int32), ("reserved", C.c_int32)]


class ConstraintJob(C.Structure):
    _fields_ = [("submap_index", C.c_int32), ("match_full_submap", C.c_int32),
                ("global_node_pose", c_d * 7), ("global_submap_pose", c_d * 7), ("data", NodeData)]


class ConstraintOptions(C.Structure):
    _fields_ = [("min_score", c_f), ("global_localization_min_score", c_f),
                ("occupied_space_weights", c_d * 2), ("translation_weight", c_d),
                ("rotation_weight", c_d), ("only_optimize_yaw", C.c_int32),
                ("max_num_iterations", C.c_int32), ("use_nonmonotonic_steps", C.c_int32)]


class ConstraintResult(C.Structure):
    _fields_ = [("found", C.c_int32), ("score", c_f), ("pose", c_d * 7), ("rotational_score", c_f),
                ("low_resolution_score", c_f), ("num_candidates_scored", c_i64)]


# Every symbol include/hg_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "hg_last_error": (C.c_char_p, []),
    "hg_version": (C.c_char_p, []),
    "hg_kernel_launch_count": (C.c_uint64, []),
    "hg_set_device": (c_i, [c_i]),
    "hg_grid_upload_prob": (c_i, [c_f, c_i, c_i64, vp, vp, vp]),
    "hg_grid_upload_tsdf": (c_i, [c_f, c_f, c_f, c_i, c_i64, vp, vp, vp, vp]),
    "hg_grid_destroy": (None, [vp]),
    "hg_grid_size": (c_i, [vp]),
    "hg_grid_num_bricks": (c_i64, [vp]),
    "hg_grid_lookup": (c_i, [vp, c_i, c_i64, vp, vp]),
    "hg_grid_stencil_bytes": (c_i64, [vp]),
    "hg_grid_interpolated_weight": (c_i, [vp, c_i64, vp, vp]),
    "hg_get_cell_index": (c_i, [c_f, c_i64, vp, vp]),
    "hg_cost_create": (c_i, [c_i, c_d, vp, c_i, c_i, vp, c_i, c_d, vp]),
    "hg_cost_destroy": (None, [vp]),
    "hg_cost_num_residuals": (c_i, [vp]),
    "hg_cost_num_parameter_blocks": (c_i, [vp]),
    "hg_cost_evaluate": (c_i, [vp, vp, vp, vp]),
    "hg_cost_evaluate_device": (c_i, [vp, vp, vp, vp, vp]),
    "hg_cost_normal_equations": (c_i, [vp, vp, vp, vp, vp]),
    "hg_cost_normal_equations_static": (c_i, [vp, vp, vp, vp, vp]),
    "hg_batch_create": (c_i, [c_i, c_i, vp, c_i, vp, vp, vp, vp, c_i, vp]),
    "hg_batch_destroy": (None, [vp]),
    "hg_batch_num_residuals": (c_i64, [vp]),
    "hg_batch_update_points": (c_i, [vp, vp, c_i]),
    "hg_batch_evaluate": (c_i, [vp, vp, c_i, vp, vp, c_i, vp, vp, vp, vp, vp, vp]),
    "hg_batch_normal_equations": (c_i, [vp, vp, c_i, vp, vp, vp, vp, vp]),
    "hg_batch_level_usage": (c_i, [vp, vp, c_i, vp, vp, vp]),
    "hg_scan_match_evaluate": (c_i, [c_i, vp, vp, vp, vp, c_d, c_d, c_i, vp, vp, vp, vp, vp]),
    "hg_scan_match": (c_i, [c_i, vp, vp, vp, vp, c_d, c_d, c_i, c_i, c_i, vp, vp, vp, vp]),
    "hg_stack_build": (c_i, [vp, c_i, c_i, vp]),
    "hg_stack_destroy": (None, [vp]),
    "hg_stack_depth": (c_i, [vp]),
    "hg_stack_lookup": (c_i, [vp, c_i, c_i64, vp, vp]),
    "hg_stack_num_nonzero": (c_i64, [vp, c_i]),
    "hg_stack_export": (c_i, [vp, c_i, c_i64, vp, vp, vp]),
    "hg_fcsm_create": (c_i, [vp, vp, vp, c_i, vp, vp]),
    "hg_fcsm_destroy": (None, [vp]),
    "hg_fcsm_stack": (vp, [vp]),
    "hg_fcsm_match": (c_i, [vp, vp, vp, vp, c_f, vp]),
    "hg_fcsm_match_full_submap": (c_i, [vp, vp, vp, vp, c_f, c_i, c_i, vp]),
    "hg_fcsm_match_batch": (c_i, [vp, c_i, vp, vp, vp, c_f, vp]),
    "hg_fcsm_full_submap_begin": (c_i, [vp, vp, vp, vp, c_f, c_i, c_i, vp, vp]),
    "hg_fcsm_full_submap_begin_async": (c_i, [vp, vp, vp, vp, c_f, c_i, c_i, vp, vp, vp]),
    "hg_fcsm_full_submap_finish": (c_i, [vp, c_f, vp]),
    "hg_fcsm_search_abandon": (None, [vp]),
    "hg_pack_best_record": (c_i, [vp, vp]),
    "hg_unpack_best_record": (c_i, [vp, vp]),
    "hg_select_best_record": (c_i, [vp, c_i, vp, vp]),
    "hg_fcsm_discrete_scans": (c_i, [vp, c_i, c_i, c_d, vp, vp, vp, c_i, vp, vp, vp, vp]),
    "hg_fcsm_score_candidates": (c_i, [vp, c_i, c_i, c_d, vp, vp, vp, c_i, c_i64, vp, vp, vp, vp]),
    "hg_low_resolution_scores": (c_i, [vp, vp, c_i, c_i64, vp, vp]),
    "hg_constraints_run_batch": (c_i, [vp, vp, vp, c_i, vp, c_i64, vp, c_i, c_i, vp]),
    "hg_grid_export": (c_i, [vp, c_i64, vp, vp, vp, vp]),
    "hg_grid_proto_decode": (c_i, [vp, c_i64, c_i, vp, c_i64, vp, vp, vp]),
    "hg_grid_from_proto": (c_i, [vp, c_i64, c_i, vp]),
    "hg_grid_to_proto": (c_i, [vp, vp, c_i64, vp]),
    "hg_pbstream_open": (c_i, [C.c_char_p, vp]),
    "hg_pbstream_next_submap": (c_i, [vp, vp, vp]),
    "hg_pbstream_submap_grid": (c_i, [vp, c_i, vp]),
    "hg_pbstream_submap_histogram": (c_i, [vp, c_i, vp]),
    "hg_pbstream_close": (None, [vp]),
    "hg_grid_insert_occupancy": (c_i, [vp, vp, vp, vp, c_i, c_i64]),
    "hg_grid_insert_tsdf": (c_i, [vp, vp, vp, vp, c_i, c_i64, vp]),
    "hg_voxel_filter": (c_i, [c_f, vp, c_i, c_i64, vp, vp]),
    "hg_adaptive_voxel_filter": (c_i, [c_f, c_i, c_f, vp, c_i, c_i64, vp, vp]),
}


def build(verbose=False):
    """Compiles libhg_b200.so in-tree (nvcc, sm_100a)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-j8"]
    if not verbose:
        cmd.append("-s")
    subprocess.check_call(cmd)


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise HgError("libhg_b200.so is missing: run hectorgrapher_b200.build() / "
                          "__graft_entry__.build(); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            if "HG_LIB_PATH" in os.environ and not hasattr(L, name):
                continue  # an older build under A/B: calling what it lacks fails loudly at the call
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def _check(rc):
    if rc != 0:
        raise HgError("hg error %d: %s" % (rc, lib().hg_last_error().decode()))


def _p(a):
    return None if a is None else a.ctypes.data_as(vp)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def set_device(device):
    _check(lib().hg_set_device(device))


def kernel_launch_count():
    return int(lib().hg_kernel_launch_count())


class Grid:
    def __init__(self, handle, tsdf):
        self.h = handle
        self.is_tsdf = tsdf

    @staticmethod
    def prob(resolution, xyz, values, grid_size=0):
        xyz = _i32(xyz).reshape(-1, 3)
        values = np.ascontiguousarray(values, dtype=np.uint16)
        out = vp()
        _check(lib().hg_grid_upload_prob(resolution, grid_size, len(xyz), _p(xyz), _p(values), C.byref(out)))
        return Grid(out, False)

    @staticmethod
    def tsdf(resolution, relative_truncation_distance, max_weight, xyz, tsd, weight, grid_size=0):
        xyz = _i32(xyz).reshape(-1, 3)
        tsd = np.ascontiguousarray(tsd, dtype=np.uint16)
        weight = np.ascontiguousarray(weight, dtype=np.uint16)
        out = vp()
        _check(lib().hg_grid_upload_tsdf(resolution, relative_truncation_distance, max_weight, grid_size,
                                         len(xyz), _p(xyz), _p(tsd), _p(weight), C.byref(out)))
        return Grid(out, True)

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().hg_grid_destroy(self.h)
                self.h = None
        except Exception:  # interpreter shutdown: module globals may already be gone
            pass

    @property
    def grid_size(self):
        return lib().hg_grid_size(self.h)

    @property
    def num_bricks(self):
        return lib().hg_grid_num_bricks(self.h)

    @property
    def stencil_bytes(self):
        return lib().hg_grid_stencil_bytes(self.h)

    def lookup(self, what, xyz):
        xyz = _i32(xyz).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        code = {"probability": 0, "tsd": 1, "weight": 2}[what]
        _check(lib().hg_grid_lookup(self.h, code, len(xyz), _p(xyz), _p(out)))
        return out

    def interpolated_weight(self, xyz):
        """InterpolatedTSDF::GetWeight (interpolated_tsdf.h:129-156) at real-valued points."""
        xyz = _f64(xyz).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float64)
        _check(lib().hg_grid_interpolated_weight(self.h, len(xyz), _p(xyz), _p(out)))
        return out

    @staticmethod
    def from_proto(data, tsdf):
        """HybridGrid(proto) / HybridGridTSDF(proto) from a serialized message (bytes)."""
        buf = (C.c_char * len(data)).from_buffer_copy(data)
        out = vp()
        _check(lib().hg_grid_from_proto(buf, len(data), int(tsdf), C.byref(out)))
        return Grid(out, bool(tsdf))

    def to_proto(self):
        """Serialized proto::HybridGrid / proto::HybridGridTSDF (bytes)."""
        n = c_i64(0)
        _check(lib().hg_grid_to_proto(self.h, None, 0, C.byref(n)))
        buf = (C.c_char * max(n.value, 1))()
        _check(lib().hg_grid_to_proto(self.h, buf, n.value, C.byref(n)))
        return bytes(buf[:n.value])

    def export(self):
        """(xyz [n,3] int32, values_a [n] uint16, values_b [n] uint16 or None): the non-default voxels."""
        n = c_i64(0)
        _check(lib().hg_grid_export(self.h, 0, C.byref(n), None, None, None))
        xyz = np.zeros((n.value, 3), np.int32)
        a = np.zeros(n.value, np.uint16)
        b = np.zeros(n.value, np.uint16) if self.is_tsdf else None
        if n.value:
            _check(lib().hg_grid_export(self.h, n.value, C.byref(n), _p(xyz), _p(a),
                                        _p(b) if b is not None else None))
        return xyz, a, b

    def insert_occupancy(self, origin, pts, hit_probability, miss_probability, num_free_space_voxels, stride=3):
        """OccupancyGridRangeDataInserter3D::Insert on the device."""
        origin = _f32(origin)
        pts = _f32(pts).reshape(-1, stride)
        o = OccupancyInserterOptions(hit_probability, miss_probability, num_free_space_voxels, 0)
        _check(lib().hg_grid_insert_occupancy(self.h, C.byref(o), _p(origin), _p(pts), stride, len(pts)))

    def insert_tsdf(self, origin, pts, relative_truncation_distance=2.5, maximum_weight=1000.0,
                    num_free_space_voxels=0, epsilon=0.1, sigma=4.0, min_range=0.0, max_range=1e30,
                    insertion_ratio=1.0, stride=3):
        """TSDFRangeDataInserter3D::Insert (no normal projection) on the device; returns the number of
        (return, sample) cell updates applied."""
        origin = _f32(origin)
        pts = _f32(pts).reshape(-1, stride)
        o = TsdfInserterOptions(relative_truncation_distance, maximum_weight, epsilon, sigma, min_range,
                                max_range, insertion_ratio, num_free_space_voxels, 0)
        n = c_i64(0)
        _check(lib().hg_grid_insert_tsdf(self.h, C.byref(o), _p(origin), _p(pts), stride, len(pts), C.byref(n)))
        return n.value


def grid_proto_decode(data, tsdf):
    """Host-only decode of a serialized grid message: (info dict, xyz, values_a, values_b or None)."""
    buf = (C.c_char * max(len(data), 1)).from_buffer_copy(data if data else b"\0")
    info = GridProtoInfo()
    _check(lib().hg_grid_proto_decode(buf, len(data), int(tsdf), C.byref(info), 0, None, None, None))
    n = info.num_voxels
    xyz = np.zeros((n, 3), np.int32)
    a = np.zeros(n, np.uint16)
    b = np.zeros(n, np.uint16) if tsdf else None
    if n:
        _check(lib().hg_grid_proto_decode(buf, len(data), int(tsdf), C.byref(info), n, _p(xyz), _p(a),
                                          _p(b) if tsdf else None))
    d = dict(resolution=info.resolution, relative_truncation_distance=info.relative_truncation_distance,
             max_weight=info.max_weight, num_voxels=n)
    return d, xyz, a, b


class PbStream:
    """Iterates the 3D submaps of a .pbstream file: yields (info dict, high Grid or None, low Grid or
    None, histogram)."""

    def __init__(self, path, upload=True):
        self.h = vp()
        self.upload = upload
        _check(lib().hg_pbstream_open(str(path).encode(), C.byref(self.h)))

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().hg_pbstream_close(self.h)
                self.h = None
        except Exception:  # interpreter shutdown: module globals may already be gone
            pass

    def __iter__(self):
        while True:
            has = C.c_int(0)
            info = SubmapInfo()
            _check(lib().hg_pbstream_next_submap(self.h, C.byref(has), C.byref(info)))
            if not has.value:
                return
            d = dict(trajectory_id=info.trajectory_id, submap_index=info.submap_index,
                     local_pose=np.array(list(info.local_pose)), num_range_data=info.num_range_data,
                     finished=bool(info.finished), high_is_tsdf=info.high_is_tsdf, low_is_tsdf=info.low_is_tsdf)
            hist = np.zeros(info.histogram_size, np.float32)
            _check(lib().hg_pbstream_submap_histogram(self.h, info.histogram_size, _p(hist) if len(hist) else None))
            grids = [None, None]
            if self.upload:
                for which, flag in enumerate((info.high_is_tsdf, info.low_is_tsdf)):
                    if flag >= 0:
                        out = vp()
                        _check(lib().hg_pbstream_submap_grid(self.h, which, C.byref(out)))
                        grids[which] = Grid(out, bool(flag))
            yield d, grids[0], grids[1], hist


def get_cell_index(resolution, pts):
    pts = _f32(pts).reshape(-1, 3)
    out = np.zeros((len(pts), 3), np.int32)
    _check(lib().hg_get_cell_index(resolution, len(pts), _p(pts), _p(out)))
    return out


def _grid_array(grids):
    return (vp * len(grids))(*[g.h for g in grids])


class Cost:
    """One residual block; evaluate() has ceres::CostFunction::Evaluate semantics."""

    def __init__(self, kind, scaling, pts, grids, ratio=0.0, stride=3):
        self.pts = _f32(pts).reshape(-1, stride)
        self.grids = list(grids)
        out = vp()
        k = KINDS[kind] if isinstance(kind, str) else kind
        _check(lib().hg_cost_create(k, scaling, _p(self.pts), stride, len(self.pts), _grid_array(grids),
                                    len(grids), ratio, C.byref(out)))
        self.h = out
        self.n = lib().hg_cost_num_residuals(self.h)
        self.nblocks = lib().hg_cost_num_parameter_blocks(self.h)
        self.sizes = [3, 4] * (self.nblocks // 2)

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().hg_cost_destroy(self.h)
                self.h = None
        except Exception:  # interpreter shutdown: module globals may already be gone
            pass

    def _blocks(self, params):
        params = _f64(params).reshape(-1)
        blocks, off = [], 0
        for s in self.sizes:
            blocks.append(np.ascontiguousarray(params[off:off + s]))
            off += s
        return blocks

    def evaluate(self, params, jacobians=True, skip_blocks=()):
        blocks = self._blocks(params)
        parr = (vp * len(blocks))(*[b.ctypes.data for b in blocks])
        r = np.zeros(self.n, np.float64)
        if not jacobians:
            _check(lib().hg_cost_evaluate(self.h, parr, _p(r), None))
            return r, None
        J = [np.zeros((self.n, s), np.float64) for s in self.sizes]
        jarr = (vp * len(J))(*[None if i in skip_blocks else J[i].ctypes.data for i in range(len(J))])
        _check(lib().hg_cost_evaluate(self.h, parr, _p(r), jarr))
        return r, J

    def normal_equations(self, params, static_scratch=False):
        """static_scratch=True goes through hg_cost_normal_equations_static (no allocation at call time)."""
        blocks = self._blocks(params)
        parr = (vp * len(blocks))(*[b.ctypes.data for b in blocks])
        n = sum(self.sizes)
        JtJ = np.zeros((n, n))
        Jtr = np.zeros(n)
        cost = np.zeros(1)
        fn = lib().hg_cost_normal_equations_static if static_scratch else lib().hg_cost_normal_equations
        _check(fn(self.h, parr, _p(JtJ), _p(Jtr), _p(cost)))
        return JtJ, Jtr, float(cost[0])


class Batch:
    """Many residual blocks of one kind over the same grids, one launch."""

    def __init__(self, kind, clouds_or_points, offsets, scales, grids, ratios=None, stride=3):
        self.kind = KINDS[kind] if isinstance(kind, str) else kind
        self.two_pose = self.kind in (1, 3, 4, 6, 7)
        pts = _f32(clouds_or_points).reshape(-1, stride)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        self.num_blocks = len(self.offsets) - 1
        self.N = int(self.offsets[-1])
        self.stride = stride
        scales = _f64(scales)
        ratios = None if ratios is None else _f64(ratios)
        self.grids = list(grids)
        out = vp()
        _check(lib().hg_batch_create(self.kind, self.num_blocks, _p(pts), stride, _p(self.offsets), _p(scales),
                                     _p(ratios), _grid_array(grids), len(grids), C.byref(out)))
        self.h = out

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().hg_batch_destroy(self.h)
                self.h = None
        except Exception:  # interpreter shutdown: module globals may already be gone
            pass

    def update_points(self, pts):
        """pts: host array (pinned torch tensors are passed via .numpy())."""
        pts = _f32(pts)
        _check(lib().hg_batch_update_points(self.h, _p(pts), self.stride))

    def evaluate_raw(self, poses_ptr, num_poses, i0_ptr, i1_ptr, device_outputs, r, jt0, jq0, jt1, jq1,
                     stream=None):
        """All arguments are raw addresses (ints) or None."""
        _check(lib().hg_batch_evaluate(self.h, poses_ptr, num_poses, i0_ptr, i1_ptr, int(device_outputs), r,
                                       jt0, jq0, jt1, jq1, stream))

    def evaluate(self, poses, i0, i1=None):
        """Host outputs as numpy arrays (parity tests)."""
        poses = _f64(poses).reshape(-1, 7)
        i0 = _i32(i0)
        i1 = None if i1 is None else _i32(i1)
        N = self.N
        r = np.zeros(N)
        jt0, jq0 = np.zeros((N, 3)), np.zeros((N, 4))
        jt1 = np.zeros((N, 3)) if self.two_pose else None
        jq1 = np.zeros((N, 4)) if self.two_pose else None
        _check(lib().hg_batch_evaluate(self.h, _p(poses), len(poses), _p(i0), _p(i1), 0, _p(r), _p(jt0),
                                       _p(jq0), _p(jt1), _p(jq1), None))
        return r, [jt0, jq0] + ([jt1, jq1] if self.two_pose else [])

    def normal_equations(self, poses, i0, i1=None):
        poses = _f64(poses).reshape(-1, 7)
        i0 = _i32(i0)
        i1 = None if i1 is None else _i32(i1)
        n = 14 if self.two_pose else 7
        JtJ = np.zeros((self.num_blocks, n, n))
        Jtr = np.zeros((self.num_blocks, n))
        cost = np.zeros(self.num_blocks)
        _check(lib().hg_batch_normal_equations(self.h, _p(poses), len(poses), _p(i0), _p(i1), _p(JtJ), _p(Jtr),
                                               _p(cost)))
        return JtJ, Jtr, cost


    def level_usage(self, poses, i0, i1=None):
        """Points answered by each pyramid level, last entry = by none (multi-resolution kinds)."""
        poses = _f64(poses).reshape(-1, 7)
        i0 = _i32(i0)
        i1 = None if i1 is None else _i32(i1)
        counts = np.zeros(len(self.grids) + 1, np.int64)
        _check(lib().hg_batch_level_usage(self.h, _p(poses), len(poses), _p(i0), _p(i1), _p(counts)))
        return counts


def scan_match(clouds, grids, weights, translation_weight, rotation_weight, initial_pose,
               target_translation=None, only_optimize_yaw=False, max_num_iterations=10,
               use_nonmonotonic_steps=True):
    clouds = [_f32(c).reshape(-1, 3) for c in clouds]
    carr = (vp * len(clouds))(*[c.ctypes.data for c in clouds])
    sizes = _i32([len(c) for c in clouds])
    weights = _f64(weights)
    initial_pose = _f64(initial_pose)
    target = _f64(initial_pose[:3] if target_translation is None else target_translation)
    pose = np.zeros(7)
    summary = np.zeros(6)
    _check(lib().hg_scan_match(len(clouds), carr, _p(sizes), _grid_array(grids), _p(weights),
                               translation_weight, rotation_weight, int(only_optimize_yaw),
                               max_num_iterations, int(use_nonmonotonic_steps), _p(target), _p(initial_pose),
                               _p(pose), _p(summary)))
    return pose, summary


def scan_match_evaluate(clouds, grids, weights, translation_weight, rotation_weight, pose, target_translation=None,
                        only_optimize_yaw=False):
    """hg_scan_match_evaluate (CeresScanMatcher3D::Evaluate): (cost, residuals, gradient in the tangent space)."""
    clouds = [_f32(c).reshape(-1, 3) for c in clouds]
    carr = (vp * len(clouds))(*[c.ctypes.data for c in clouds])
    sizes = _i32([len(c) for c in clouds])
    weights = _f64(weights)
    pose = _f64(pose)
    target = _f64(pose[:3] if target_translation is None else target_translation)
    cost = np.zeros(1)
    r = np.zeros(int(sizes.sum()) + 6)
    g = np.zeros(4 if only_optimize_yaw else 6)
    _check(lib().hg_scan_match_evaluate(len(clouds), carr, _p(sizes), _grid_array(grids), _p(weights),
                                        translation_weight, rotation_weight, int(only_optimize_yaw), _p(target),
                                        _p(pose), _p(cost), _p(r), _p(g)))
    return float(cost[0]), r, g


class Stack:
    def __init__(self, grid, depth, full_resolution_depth, handle=None, owner=None):
        self.grid = grid
        self.owner = owner
        if handle is None:
            out = vp()
            _check(lib().hg_stack_build(grid.h, depth, full_resolution_depth, C.byref(out)))
            self.h = out
            self.owned = True
        else:
            self.h = handle
            self.owned = False
        self.depth = lib().hg_stack_depth(self.h)

    def __del__(self):
        try:
            if getattr(self, "h", None) and self.owned:
                lib().hg_stack_destroy(self.h)
            self.h = None
        except Exception:  # interpreter shutdown: module globals may already be gone
            pass

    def lookup(self, d, xyz):
        xyz = _i32(xyz).reshape(-1, 3)
        out = np.zeros(len(xyz), np.uint8)
        _check(lib().hg_stack_lookup(self.h, d, len(xyz), _p(xyz), _p(out)))
        return out

    def export(self, d):
        n = lib().hg_stack_num_nonzero(self.h, d)
        if n < 0:
            raise HgError(lib().hg_last_error().decode())
        xyz = np.zeros((n, 3), np.int32)
        v = np.zeros(n, np.uint8)
        written = c_i64(0)
        _check(lib().hg_stack_export(self.h, d, n, _p(xyz), _p(v), C.byref(written)))
        assert written.value == n
        return xyz, v


FCSM_DEFAULTS = dict(branch_and_bound_depth=8, full_resolution_depth=3, use_rotational_scan_matcher=True,
                     min_rotational_score=0.77, min_low_resolution_score=0.55,
                     linear_xy_search_window=5.0, linear_z_search_window=1.0,
                     angular_search_window=float(np.deg2rad(15.0)), float_flavor=1, strict_visit_order=0)


def make_node_data(high, low, hist, gravity=(1, 0, 0, 0)):
    """Returns (NodeData, keepalive tuple)."""
    high = _f32(high).reshape(-1, 3)
    low = _f32(low).reshape(-1, 3)
    hist = _f32(hist)
    d = NodeData()
    d.high_resolution_point_cloud = high.ctypes.data
    d.num_high_resolution_points = len(high)
    d.low_resolution_point_cloud = low.ctypes.data if len(low) else None
    d.num_low_resolution_points = len(low)
    d.rotational_scan_matcher_histogram = hist.ctypes.data if len(hist) else None
    d.histogram_size = len(hist)
    for k in range(4):
        d.gravity_alignment[k] = float(gravity[k])
    return d, (high, low, hist)


def _result_dict(r):
    return dict(found=float(r.found), score=float(r.score), tx=r.pose_translation[0], ty=r.pose_translation[1],
                tz=r.pose_translation[2], qw=r.pose_rotation[0], qx=r.pose_rotation[1],
                qy=r.pose_rotation[2], qz=r.pose_rotation[3], rotational_score=float(r.rotational_score),
                low_resolution_score=float(r.low_resolution_score), scan_index=float(r.best_scan_index),
                ox=float(r.best_offset[0]), oy=float(r.best_offset[1]), oz=float(r.best_offset[2]),
                num_scans=float(r.num_discrete_scans), num_scored=int(r.num_candidates_scored),
                num_visited=int(r.num_candidates_visited), num_launches=int(r.num_score_launches),
                num_lattice=int(r.num_lowest_resolution_candidates),
                lattice_kernel_ms=float(r.lowest_resolution_kernel_ms),
                ancestor_score=float(r.best_ancestor_score), ancestor_index=int(r.best_ancestor_index),
                dive_bound=float(r.dive_bound), order_unresolved=int(r.order_unresolved))


BEST_RECORD_DOUBLES = 20


def _result_struct(d):
    """MatchResult from a result dict (the fields the best-record exchange carries)."""
    r = MatchResult()
    r.found = int(d["found"])
    r.score = d["score"]
    for k, n in enumerate(("tx", "ty", "tz")):
        r.pose_translation[k] = d[n]
    for k, n in enumerate(("qw", "qx", "qy", "qz")):
        r.pose_rotation[k] = d[n]
    r.rotational_score = d["rotational_score"]
    r.low_resolution_score = d["low_resolution_score"]
    r.best_scan_index = int(d["scan_index"])
    for k, n in enumerate(("ox", "oy", "oz")):
        r.best_offset[k] = int(d[n])
    r.num_candidates_scored = int(d.get("num_scored", 0))
    r.best_ancestor_score = d.get("ancestor_score", 0.0)
    r.best_ancestor_index = int(d.get("ancestor_index", 0))
    r.order_unresolved = int(d.get("order_unresolved", 0))
    return r


def pack_best_record(result):
    """hg_pack_best_record of a result dict -> numpy array of BEST_RECORD_DOUBLES doubles."""
    rec = np.zeros(BEST_RECORD_DOUBLES)
    r = _result_struct(result)
    _check(lib().hg_pack_best_record(C.byref(r), _p(rec)))
    return rec


def select_best_record(records):
    """hg_select_best_record over [n, BEST_RECORD_DOUBLES] records -> (result dict or None, order_unresolved)."""
    records = _f64(records).reshape(-1, BEST_RECORD_DOUBLES)
    winner, unresolved = C.c_int(-1), C.c_int(0)
    _check(lib().hg_select_best_record(_p(records), len(records), C.byref(winner), C.byref(unresolved)))
    if winner.value < 0:
        return None, False
    r = MatchResult()
    _check(lib().hg_unpack_best_record(_p(np.ascontiguousarray(records[winner.value])), C.byref(r)))
    return _result_dict(r), bool(unresolved.value)


class Fcsm:
    def __init__(self, grid, low_grid, submap_hist, **options):
        o = dict(FCSM_DEFAULTS)
        o.update(options)
        self.o = o
        self.grid, self.low_grid = grid, low_grid
        self.hist = _f32(submap_hist)
        opt = FcsmOptions()
        for k, v in o.items():
            setattr(opt, k, int(v) if k in ("branch_and_bound_depth", "full_resolution_depth",
                                            "use_rotational_scan_matcher", "float_flavor",
                                            "strict_visit_order") else float(v))
        self.opt = opt
        out = vp()
        _check(lib().hg_fcsm_create(grid.h, low_grid.h, _p(self.hist) if len(self.hist) else None,
                                    len(self.hist), C.byref(opt), C.byref(out)))
        self.h = out

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().hg_fcsm_destroy(self.h)
                self.h = None
        except Exception:  # interpreter shutdown: module globals may already be gone
            pass

    def stack(self):
        return Stack(self.grid, 0, 0, handle=lib().hg_fcsm_stack(self.h), owner=self)

    def match(self, node_pose, submap_pose, high, low, hist, gravity=(1, 0, 0, 0), min_score=0.55,
              full=False, shard_index=0, num_shards=1):
        node_pose, submap_pose = _f64(node_pose), _f64(submap_pose)
        data, keep = make_node_data(high, low, hist, gravity)
        res = MatchResult()
        if full:
            nq = np.ascontiguousarray(node_pose[3:])
            sq = np.ascontiguousarray(submap_pose[3:])
            _check(lib().hg_fcsm_match_full_submap(self.h, _p(nq), _p(sq), C.byref(data), min_score,
                                                   shard_index, num_shards, C.byref(res)))
        else:
            _check(lib().hg_fcsm_match(self.h, _p(node_pose), _p(submap_pose), C.byref(data), min_score,
                                       C.byref(res)))
        del keep
        return _result_dict(res)

    def match_batch(self, node_poses, submap_poses, nodes, min_score=0.55):
        """nodes: list of (high, low, hist[, gravity]) tuples; one device-resident search for all."""
        n = len(nodes)
        node_poses = _f64(node_poses).reshape(n, 7)
        submap_poses = _f64(submap_poses).reshape(n, 7)
        datas = (NodeData * n)()
        keep = []
        for k, node in enumerate(nodes):
            d, ka = make_node_data(*node)
            datas[k] = d
            keep.append(ka)
        res = (MatchResult * n)()
        _check(lib().hg_fcsm_match_batch(self.h, n, _p(node_poses), _p(submap_poses), datas, min_score, res))
        del keep
        return [_result_dict(r) for r in res]

    def full_submap_begin(self, node_rotation, submap_rotation, high, low, hist, gravity=(1, 0, 0, 0),
                          min_score=0.55, shard_index=0, num_shards=1):
        """First half of a sharded MatchFullSubmap; returns (handle, dive bound). The handle keeps the
        node data alive; pass it to full_submap_finish on the same thread."""
        nq, sq = _f64(node_rotation), _f64(submap_rotation)
        data, keep = make_node_data(high, low, hist, gravity)
        out = vp()
        bound = C.c_float(0)
        _check(lib().hg_fcsm_full_submap_begin(self.h, _p(nq), _p(sq), C.byref(data), min_score, shard_index,
                                               num_shards, C.byref(out), C.byref(bound)))
        return (out, data, keep), float(bound.value)

    def full_submap_begin_async(self, node_rotation, submap_rotation, high, low, hist, gravity=(1, 0, 0, 0),
                                min_score=0.55, shard_index=0, num_shards=1):
        """First half without a host round trip (hg_fcsm_full_submap_begin_async): returns (handle,
        device address of the float holding the dive bound or 0, cudaStream_t the work is queued on or 0).
        All-reduce-max the float in place on that stream, then full_submap_finish(handle, 0.0)."""
        nq, sq = _f64(node_rotation), _f64(submap_rotation)
        data, keep = make_node_data(high, low, hist, gravity)
        out, d_bound, stream = vp(), vp(), vp()
        _check(lib().hg_fcsm_full_submap_begin_async(self.h, _p(nq), _p(sq), C.byref(data), min_score, shard_index,
                                                     num_shards, C.byref(out), C.byref(d_bound), C.byref(stream)))
        return (out, data, keep), int(d_bound.value or 0), int(stream.value or 0)

    def full_submap_finish(self, handle, bound):
        res = MatchResult()
        _check(lib().hg_fcsm_full_submap_finish(handle[0], float(bound), C.byref(res)))
        return _result_dict(res)

    def discrete_scans(self, wxy, wz, angular_window, node_pose, submap_pose, high, hist, gravity=(1, 0, 0, 0)):
        node_pose, submap_pose = _f64(node_pose), _f64(submap_pose)
        data, keep = make_node_data(high, np.zeros((0, 3)), hist, gravity)
        S = C.c_int32(0)
        _check(lib().hg_fcsm_discrete_scans(self.h, wxy, wz, angular_window, _p(node_pose), _p(submap_pose),
                                            C.byref(data), 0, C.byref(S), None, None, None))
        D = self.o["branch_and_bound_depth"]
        n = data.num_high_resolution_points
        poses = np.zeros((S.value, 7), np.float32)
        rot = np.zeros(S.value, np.float32)
        cells = np.zeros((S.value, D, n, 3), np.int32)
        _check(lib().hg_fcsm_discrete_scans(self.h, wxy, wz, angular_window, _p(node_pose), _p(submap_pose),
                                            C.byref(data), S.value, C.byref(S), _p(poses), _p(rot), _p(cells)))
        del keep
        return poses, rot, cells

    def score_candidates(self, wxy, wz, angular_window, node_pose, submap_pose, high, hist, depth, cand,
                         gravity=(1, 0, 0, 0)):
        node_pose, submap_pose = _f64(node_pose), _f64(submap_pose)
        data, keep = make_node_data(high, np.zeros((0, 3)), hist, gravity)
        cand = _i32(cand).reshape(-1, 4)
        sums = np.zeros(len(cand), np.int32)
        scores = np.zeros(len(cand), np.float32)
        ms = c_f(0)
        _check(lib().hg_fcsm_score_candidates(self.h, wxy, wz, angular_window, _p(node_pose), _p(submap_pose),
                                              C.byref(data), depth, len(cand), _p(cand), _p(sums), _p(scores),
                                              C.byref(ms)))
        del keep
        return sums, scores, ms.value


def low_resolution_scores(low_grid, pts, poses7):
    pts = _f32(pts).reshape(-1, 3)
    poses7 = _f32(poses7).reshape(-1, 7)
    out = np.zeros(len(poses7), np.float32)
    _check(lib().hg_low_resolution_scores(low_grid.h, _p(pts), len(pts), len(poses7), _p(poses7), _p(out)))
    return out


class ConstraintBatch:
    """Marshals a list of jobs once (dicts: submap_index, full, node_pose, submap_pose, high, low, hist,
    gravity) so that run() is just the hg_constraints_run_batch call."""

    def __init__(self, matchers, grids, low_grids, jobs, options):
        n = len(jobs)
        self.n = n
        self.arr = (ConstraintJob * n)()
        self.keep = [matchers, grids, low_grids]
        cache = {}
        for i, j in enumerate(jobs):
            self.arr[i].submap_index = j["submap_index"]
            self.arr[i].match_full_submap = int(j.get("full", False))
            for k in range(7):
                self.arr[i].global_node_pose[k] = float(j["node_pose"][k])
                self.arr[i].global_submap_pose[k] = float(j["submap_pose"][k])
            key = (id(j["high"]), id(j["low"]), id(j["hist"]), tuple(j.get("gravity", (1, 0, 0, 0))))
            if key not in cache:  # jobs that share their clouds share one marshalled copy
                cache[key] = make_node_data(j["high"], j["low"], j["hist"], j.get("gravity", (1, 0, 0, 0)))
                self.keep.append((j["high"], j["low"], j["hist"]))
            d, ka = cache[key]
            self.arr[i].data = d
            self.keep.append(ka)
        opt = ConstraintOptions()
        opt.min_score = options.get("min_score", 0.55)
        opt.global_localization_min_score = options.get("global_localization_min_score", 0.6)
        w = options.get("occupied_space_weights", (5.0, 30.0))
        opt.occupied_space_weights[0], opt.occupied_space_weights[1] = w
        opt.translation_weight = options.get("translation_weight", 10.0)
        opt.rotation_weight = options.get("rotation_weight", 1.0)
        opt.only_optimize_yaw = int(options.get("only_optimize_yaw", False))
        opt.max_num_iterations = options.get("max_num_iterations", 10)
        opt.use_nonmonotonic_steps = int(options.get("use_nonmonotonic_steps", False))
        self.opt = opt
        self.res = (ConstraintResult * n)()
        self.marr = (vp * len(matchers))(*[m.h for m in matchers])
        self.garr = _grid_array(grids)
        self.larr = _grid_array(low_grids)
        self.num_submaps = len(matchers)

    def run(self, shard_index=0, num_shards=1):
        _check(lib().hg_constraints_run_batch(self.marr, self.garr, self.larr, self.num_submaps, self.arr, self.n,
                                              C.byref(self.opt), shard_index, num_shards, self.res))
        return self

    def results(self):
        out = []
        for i in range(self.n):
            r = self.res[i]
            out.append(dict(found=int(r.found), score=float(r.score), pose=np.array(list(r.pose)),
                            rotational_score=float(r.rotational_score),
                            low_resolution_score=float(r.low_resolution_score),
                            num_scored=int(r.num_candidates_scored)))
        return out

    def totals(self):
        """(constraints found, candidates scored) over all jobs."""
        return (sum(int(self.res[i].found) for i in range(self.n)),
                sum(int(self.res[i].num_candidates_scored) for i in range(self.n)))


def constraints_run_batch(matchers, grids, low_grids, jobs, options, shard_index=0, num_shards=1):
    """jobs: list of dicts(submap_index, full, node_pose, submap_pose, high, low, hist, gravity)."""
    return ConstraintBatch(matchers, grids, low_grids, jobs, options).run(shard_index, num_shards).results()


def voxel_filter(size, pts, stride=3):
    """Indices of the points sensor::VoxelFilter(size).Filter keeps."""
    pts = _f32(pts).reshape(-1, stride)
    kept = np.zeros(len(pts), np.int32)
    n = c_i64(0)
    _check(lib().hg_voxel_filter(size, _p(pts), stride, len(pts), C.byref(n), _p(kept)))
    return kept[:n.value].copy()


def adaptive_voxel_filter(max_length, min_num_points, max_range, pts, stride=3):
    pts = _f32(pts).reshape(-1, stride)
    kept = np.zeros(len(pts), np.int32)
    n = c_i64(0)
    _check(lib().hg_adaptive_voxel_filter(max_length, min_num_points, max_range, _p(pts), stride, len(pts),
                                          C.byref(n), _p(kept)))
    return kept[:n.value].copy()
