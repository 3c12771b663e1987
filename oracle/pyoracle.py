"""ORACLE — TEST INFRASTRUCTURE ONLY. ctypes view of oracle/liboracle.so.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module. The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_f = C.c_float
c_d = C.c_double
c_i = C.c_int
c_ll = C.c_longlong
vp = C.c_void_p


def build():
    subprocess.check_call(["make", "-C", _HERE, "-s"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.orc_prob_grid_new.restype = vp
        L.orc_prob_grid_new.argtypes = [c_f]
        L.orc_tsdf_grid_new.restype = vp
        L.orc_tsdf_grid_new.argtypes = [c_f, c_f, c_f]
        L.orc_grid_free.argtypes = [vp]
        L.orc_grid_size.argtypes = [vp]
        L.orc_grid_resolution.restype = c_f
        L.orc_grid_resolution.argtypes = [vp]
        L.orc_tsdf_grid_max_tsd.restype = c_f
        L.orc_tsdf_grid_max_tsd.argtypes = [vp]
        L.orc_grid_num_voxels.restype = c_ll
        L.orc_grid_num_voxels.argtypes = [vp]
        L.orc_grid_export.restype = c_ll
        L.orc_grid_export.argtypes = [vp, vp, vp, vp]
        L.orc_prob_grid_set_raw.argtypes = [vp, c_i, vp, vp]
        L.orc_prob_grid_set_probability.argtypes = [vp, c_i, vp, vp]
        L.orc_tsdf_grid_set_raw.argtypes = [vp, c_i, vp, vp, vp]
        L.orc_tsdf_grid_set_cell.argtypes = [vp, c_i, vp, vp, vp]
        L.orc_get_cell_index.argtypes = [c_f, c_i, vp, vp]
        L.orc_get_center_of_cell.argtypes = [c_f, c_i, vp, vp]
        L.orc_grid_lookup.argtypes = [vp, c_i, c_i, vp, vp]
        L.orc_value_tables.argtypes = [c_f, c_f, c_f, vp]
        L.orc_probability_to_value.restype = C.c_uint16
        L.orc_probability_to_value.argtypes = [c_f]
        L.orc_value_to_probability.restype = c_f
        L.orc_value_to_probability.argtypes = [C.c_uint16]
        L.orc_apply_odds_table.argtypes = [c_f, vp]
        L.orc_odds.restype = c_f
        L.orc_odds.argtypes = [c_f]
        L.orc_prob_grid_apply_odds.argtypes = [vp, vp, c_f]
        L.orc_prob_grid_finish_update.argtypes = [vp]
        L.orc_insert_occupancy.argtypes = [vp, vp, c_i, vp, c_f, c_f, c_i]
        L.orc_insert_tsdf.argtypes = [vp, vp, c_i, vp, c_d, c_d, c_i, c_d, c_d, c_d, c_d, c_d]
        L.orc_interpolate.argtypes = [vp, c_i, vp, vp]
        L.orc_interpolate_weight.argtypes = [vp, c_i, vp, vp]
        L.orc_cost_new.restype = vp
        L.orc_cost_new.argtypes = [c_i, c_d, vp, c_i, c_i, vp, c_i, c_d]
        L.orc_cost_free.argtypes = [vp]
        L.orc_cost_num_residuals.argtypes = [vp]
        L.orc_cost_num_blocks.argtypes = [vp]
        L.orc_cost_evaluate.argtypes = [vp, vp, vp, vp]
        L.orc_cost_time.restype = c_d
        L.orc_cost_time.argtypes = [vp, vp, c_i, c_i]
        L.orc_scan_match.argtypes = [c_i, vp, vp, vp, vp, c_d, c_d, c_i, c_i, c_i, vp, vp, vp, vp]
        L.orc_scan_match_trace.argtypes = [c_i, vp, vp, vp, vp, c_d, c_d, c_i, c_i, c_i, vp, vp, vp, vp, vp, c_i]
        L.orc_stack_new.restype = vp
        L.orc_stack_new.argtypes = [vp, c_i, c_i]
        L.orc_stack_free.argtypes = [vp]
        L.orc_stack_num_voxels.restype = c_ll
        L.orc_stack_num_voxels.argtypes = [vp, c_i]
        L.orc_stack_export.restype = c_ll
        L.orc_stack_export.argtypes = [vp, c_i, vp, vp]
        L.orc_stack_lookup.argtypes = [vp, c_i, c_i, vp, vp]
        L.orc_precompute_fullres_lookup.argtypes = [vp, c_i, c_i, vp, vp]
        L.orc_precomp_to_probability.restype = c_f
        L.orc_precomp_to_probability.argtypes = [c_f]
        L.orc_rotational_match.argtypes = [vp, vp, c_i, c_f, vp, c_i, vp]
        L.orc_rotate_histogram.argtypes = [vp, c_i, c_f, vp]
        L.orc_fcsm_new.restype = vp
        L.orc_fcsm_new.argtypes = [vp, vp, vp, c_i, c_i, c_i, c_i, c_d, c_d, c_d, c_d, c_d]
        L.orc_fcsm_free.argtypes = [vp]
        L.orc_fcsm_match.argtypes = [vp, c_i, vp, vp, vp, c_i, vp, c_i, vp, c_i, vp, c_f, vp]
        L.orc_fcsm_discrete_scans.argtypes = [vp, c_i, c_i, c_d, vp, vp, vp, c_i, vp, c_i, vp,
                                              c_i, vp, vp, vp]
        L.orc_fcsm_score_candidates.restype = c_d
        L.orc_fcsm_score_candidates.argtypes = [vp, c_i, c_i, c_d, vp, vp, vp, c_i, vp, c_i, vp,
                                                c_i, c_ll, vp, vp, vp, c_i]
        L.orc_low_resolution_score.restype = c_f
        L.orc_low_resolution_score.argtypes = [vp, vp, c_i, vp]
        L.orc_set_float_flavor.argtypes = [c_i]
        L.orc_voxel_filter.argtypes = [c_f, vp, c_i, c_i, vp]
        L.orc_adaptive_voxel_filter.argtypes = [c_f, c_i, c_f, vp, c_i, c_i, vp]
        _LIB = L
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(vp)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class Grid:
    """HybridGrid (probability) or HybridGridTSDF."""

    def __init__(self, handle, tsdf):
        self.h = handle
        self.is_tsdf = tsdf

    @staticmethod
    def prob(resolution):
        return Grid(lib().orc_prob_grid_new(resolution), False)

    @staticmethod
    def tsdf(resolution, relative_truncation_distance, max_weight):
        return Grid(lib().orc_tsdf_grid_new(resolution, relative_truncation_distance, max_weight), True)

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().orc_grid_free(self.h)
                self.h = None
        except Exception:  # interpreter shutdown: module globals may already be gone
            pass

    @property
    def grid_size(self):
        return lib().orc_grid_size(self.h)

    @property
    def resolution(self):
        return lib().orc_grid_resolution(self.h)

    @property
    def max_tsd(self):
        return lib().orc_tsdf_grid_max_tsd(self.h)

    def set_raw(self, xyz, a, b=None):
        xyz = _i32(xyz).reshape(-1, 3)
        a = np.ascontiguousarray(a, dtype=np.uint16)
        if self.is_tsdf:
            b = np.ascontiguousarray(b, dtype=np.uint16)
            lib().orc_tsdf_grid_set_raw(self.h, len(xyz), _p(xyz), _p(a), _p(b))
        else:
            lib().orc_prob_grid_set_raw(self.h, len(xyz), _p(xyz), _p(a))

    def set_probability(self, xyz, p):
        xyz = _i32(xyz).reshape(-1, 3)
        p = _f32(np.broadcast_to(p, (len(xyz),)))
        lib().orc_prob_grid_set_probability(self.h, len(xyz), _p(xyz), _p(p))

    def set_cell(self, xyz, tsd, weight):
        xyz = _i32(xyz).reshape(-1, 3)
        tsd = _f32(np.broadcast_to(tsd, (len(xyz),)))
        weight = _f32(np.broadcast_to(weight, (len(xyz),)))
        lib().orc_tsdf_grid_set_cell(self.h, len(xyz), _p(xyz), _p(tsd), _p(weight))

    def export(self):
        n = lib().orc_grid_num_voxels(self.h)
        xyz = np.zeros((n, 3), np.int32)
        a = np.zeros(n, np.uint16)
        b = np.zeros(n, np.uint16)
        lib().orc_grid_export(self.h, _p(xyz), _p(a), _p(b))
        return (xyz, a, b) if self.is_tsdf else (xyz, a)

    def lookup(self, what, xyz):
        xyz = _i32(xyz).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        code = {"probability": 0, "tsd": 1, "weight": 2, "raw_a": 3, "raw_b": 4}[what]
        lib().orc_grid_lookup(self.h, code, len(xyz), _p(xyz), _p(out))
        return out

    def cell_index(self, pts):
        return get_cell_index(self.resolution, pts)

    def apply_odds(self, xyz, probability):
        xyz = _i32(xyz).reshape(3)
        return bool(lib().orc_prob_grid_apply_odds(self.h, _p(xyz), probability))

    def finish_update(self):
        lib().orc_prob_grid_finish_update(self.h)

    def insert_occupancy(self, origin, pts, hit_p, miss_p, num_free_space_voxels):
        origin = _f32(origin)
        pts = _f32(pts).reshape(-1, 3)
        lib().orc_insert_occupancy(self.h, _p(origin), len(pts), _p(pts), hit_p, miss_p,
                                   num_free_space_voxels)

    def insert_tsdf(self, origin, pts, relative_truncation_distance=2.5, maximum_weight=1000.0,
                    num_free_space_voxels=0, epsilon=0.1, sigma=4.0, min_range=0.0, max_range=1e30,
                    insertion_ratio=1.0):
        origin = _f32(origin)
        pts = _f32(pts).reshape(-1, 3)
        lib().orc_insert_tsdf(self.h, _p(origin), len(pts), _p(pts), relative_truncation_distance,
                              maximum_weight, num_free_space_voxels, epsilon, sigma, min_range, max_range,
                              insertion_ratio)

    def interpolate(self, xyz):
        xyz = _f64(xyz).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float64)
        lib().orc_interpolate(self.h, len(xyz), _p(xyz), _p(out))
        return out

    def interpolate_weight(self, xyz):
        xyz = _f64(xyz).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float64)
        lib().orc_interpolate_weight(self.h, len(xyz), _p(xyz), _p(out))
        return out


def get_cell_index(resolution, pts):
    pts = _f32(pts).reshape(-1, 3)
    out = np.zeros((len(pts), 3), np.int32)
    lib().orc_get_cell_index(resolution, len(pts), _p(pts), _p(out))
    return out


def get_center_of_cell(resolution, xyz):
    xyz = _i32(xyz).reshape(-1, 3)
    out = np.zeros((len(xyz), 3), np.float32)
    lib().orc_get_center_of_cell(resolution, len(xyz), _p(xyz), _p(out))
    return out


def value_table(unknown, lo, hi):
    out = np.zeros(65536, np.float32)
    lib().orc_value_tables(unknown, lo, hi, _p(out))
    return out


KINDS = {
    "occupied": 0, "interpolated_occupied": 1, "tsdf": 2, "interpolated_tsdf": 3,
    "interpolated_tsdf_per_point": 4, "multires_tsdf": 5, "interpolated_multires_tsdf": 6,
    "interpolated_multires_tsdf_per_point": 7,
}


class Cost:
    def __init__(self, kind, scaling, pts, grids, ratio=0.0, stride=3):
        self.pts = _f32(pts).reshape(-1, stride)
        self.grids = list(grids)
        arr = (vp * len(grids))(*[g.h for g in grids])
        self.h = lib().orc_cost_new(KINDS[kind] if isinstance(kind, str) else kind, scaling,
                                    _p(self.pts), len(self.pts), stride, arr, len(grids), ratio)
        self.n = lib().orc_cost_num_residuals(self.h)
        self.nblocks = lib().orc_cost_num_blocks(self.h)
        self.sizes = [3, 4] * (self.nblocks // 2)

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().orc_cost_free(self.h)
                self.h = None
        except Exception:  # interpreter shutdown: module globals may already be gone
            pass

    def _params(self, params):
        params = _f64(params).reshape(-1)
        blocks, off = [], 0
        for s in self.sizes:
            blocks.append(np.ascontiguousarray(params[off:off + s]))
            off += s
        return blocks

    def evaluate(self, params, jacobians=True, skip_blocks=()):
        blocks = self._params(params)
        parr = (vp * len(blocks))(*[b.ctypes.data for b in blocks])
        r = np.zeros(self.n, np.float64)
        if not jacobians:
            lib().orc_cost_evaluate(self.h, parr, _p(r), None)
            return r, None
        J = [np.zeros((self.n, s), np.float64) for s in self.sizes]
        jarr = (vp * len(J))(*[None if i in skip_blocks else J[i].ctypes.data for i in range(len(J))])
        lib().orc_cost_evaluate(self.h, parr, _p(r), jarr)
        return r, J

    def time(self, params, repeats, threads=1):
        blocks = self._params(params)
        parr = (vp * len(blocks))(*[b.ctypes.data for b in blocks])
        return lib().orc_cost_time(self.h, parr, repeats, threads)


def interpolate_transform(start, end, factor):
    """transform::InterpolateTransform on (t3, q4 = w, x, y, z) poses."""
    out = np.zeros(7)
    f = lib().orc_interpolate_transform
    f.argtypes = [vp, vp, c_d, vp]
    f.restype = None
    f(_p(_f64(start)), _p(_f64(end)), float(factor), _p(out))
    return out


def rotation_delta(scaling_factor, target_wxyz, rotation_wxyz, jacobian=False):
    r = np.zeros(3)
    J = np.zeros((3, 4)) if jacobian else None
    f = lib().orc_rotation_delta
    f.argtypes = [c_d, vp, vp, vp, vp]
    f.restype = None
    f(float(scaling_factor), _p(_f64(target_wxyz)), _p(_f64(rotation_wxyz)), _p(r), _p(J))
    return (r, J) if jacobian else r


def translation_delta(scaling_factor, target, translation, jacobian=False):
    r = np.zeros(3)
    J = np.zeros((3, 3)) if jacobian else None
    f = lib().orc_translation_delta
    f.argtypes = [c_d, vp, vp, vp, vp]
    f.restype = None
    f(float(scaling_factor), _p(_f64(target)), _p(_f64(translation)), _p(r), _p(J))
    return (r, J) if jacobian else r


def scan_match(clouds, grids, weights, translation_weight, rotation_weight, initial_pose,
               target_translation=None, only_optimize_yaw=False, max_num_iterations=10,
               use_nonmonotonic_steps=True, trace=None):
    """trace: a list that receives one row per solver iteration (iteration, cost, candidate cost,
    model cost change, step quality, radius, successful)."""
    clouds = [_f32(c).reshape(-1, 3) for c in clouds]
    carr = (vp * len(clouds))(*[c.ctypes.data for c in clouds])
    sizes = _i32([len(c) for c in clouds])
    garr = (vp * len(grids))(*[g.h for g in grids])
    weights = _f64(weights)
    initial_pose = _f64(initial_pose)
    target = _f64(initial_pose[:3] if target_translation is None else target_translation)
    pose = np.zeros(7)
    summary = np.zeros(6)
    rows = np.zeros((max_num_iterations + 2, 7))
    term = lib().orc_scan_match_trace(len(clouds), carr, _p(sizes), garr, _p(weights), translation_weight,
                                      rotation_weight, int(only_optimize_yaw), max_num_iterations,
                                      int(use_nonmonotonic_steps), _p(target), _p(initial_pose), _p(pose),
                                      _p(summary), _p(rows), len(rows))
    if trace is not None:
        trace.extend(rows[:int(summary[2]) + 1].tolist())
    return pose, summary, term


def scan_match_evaluate(clouds, grids, weights, translation_weight, rotation_weight, pose, target_translation=None,
                        only_optimize_yaw=False):
    """CeresScanMatcher3D::Evaluate: (cost, residuals, gradient in the tangent space)."""
    clouds = [_f32(c).reshape(-1, 3) for c in clouds]
    carr = (vp * len(clouds))(*[c.ctypes.data for c in clouds])
    sizes = _i32([len(c) for c in clouds])
    garr = (vp * len(grids))(*[g.h for g in grids])
    weights = _f64(weights)
    pose = _f64(pose)
    target = _f64(pose[:3] if target_translation is None else target_translation)
    cost = np.zeros(1)
    r = np.zeros(int(sizes.sum()) + 6)
    g = np.zeros(4 if only_optimize_yaw else 6)
    f = lib().orc_scan_match_evaluate
    f.argtypes = [c_i, vp, vp, vp, vp, c_d, c_d, c_i, vp, vp, vp, vp, vp]
    rc = f(len(clouds), carr, _p(sizes), garr, _p(weights), translation_weight, rotation_weight,
           int(only_optimize_yaw), _p(target), _p(pose), _p(cost), _p(r), _p(g))
    assert rc == 0
    return float(cost[0]), r, g


class Stack:
    def __init__(self, grid, depth, full_resolution_depth):
        self.grid = grid
        self.depth = depth
        self.h = lib().orc_stack_new(grid.h, depth, full_resolution_depth)

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().orc_stack_free(self.h)
                self.h = None
        except Exception:  # interpreter shutdown: module globals may already be gone
            pass

    def export(self, d):
        n = lib().orc_stack_num_voxels(self.h, d)
        xyz = np.zeros((n, 3), np.int32)
        v = np.zeros(n, np.uint8)
        lib().orc_stack_export(self.h, d, _p(xyz), _p(v))
        return xyz, v

    def lookup(self, d, xyz):
        xyz = _i32(xyz).reshape(-1, 3)
        out = np.zeros(len(xyz), np.uint8)
        lib().orc_stack_lookup(self.h, d, len(xyz), _p(xyz), _p(out))
        return out


def precompute_fullres_lookup(grid, levels, xyz):
    xyz = _i32(xyz).reshape(-1, 3)
    out = np.zeros(len(xyz), np.uint8)
    lib().orc_precompute_fullres_lookup(grid.h, levels, len(xyz), _p(xyz), _p(out))
    return out


def rotational_match(submap_hist, scan_hist, initial_angle, angles):
    a, b, ang = _f32(submap_hist), _f32(scan_hist), _f32(angles)
    out = np.zeros(len(ang), np.float32)
    lib().orc_rotational_match(_p(a), _p(b), len(a), initial_angle, _p(ang), len(ang), _p(out))
    return out


def rotate_histogram(hist, angle):
    h = _f32(hist)
    out = np.zeros(len(h), np.float32)
    lib().orc_rotate_histogram(_p(h), len(h), angle, _p(out))
    return out


FCSM_DEFAULTS = dict(branch_and_bound_depth=8, full_resolution_depth=3,
                     use_rotational_scan_matcher=True, min_rotational_score=0.77,
                     min_low_resolution_score=0.55, linear_xy_search_window=5.0,
                     linear_z_search_window=1.0, angular_search_window=np.deg2rad(15.0))


class Fcsm:
    RESULT_FIELDS = ["found", "score", "tx", "ty", "tz", "qw", "qx", "qy", "qz", "rotational_score",
                     "low_resolution_score", "scan_index", "ox", "oy", "oz", "num_scored", "num_scans"]

    def __init__(self, grid, low_grid, submap_hist, **options):
        o = dict(FCSM_DEFAULTS)
        o.update(options)
        self.o = o
        self.grid, self.low_grid = grid, low_grid
        self.hist = _f32(submap_hist)
        self.h = lib().orc_fcsm_new(grid.h, low_grid.h, _p(self.hist), len(self.hist),
                                    o["branch_and_bound_depth"], o["full_resolution_depth"],
                                    int(o["use_rotational_scan_matcher"]), o["min_rotational_score"],
                                    o["min_low_resolution_score"], o["linear_xy_search_window"],
                                    o["linear_z_search_window"], o["angular_search_window"])

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().orc_fcsm_free(self.h)
                self.h = None
        except Exception:  # interpreter shutdown: module globals may already be gone
            pass

    def match(self, node_pose, submap_pose, high, low, hist, gravity=(1, 0, 0, 0), min_score=0.55,
              full=False):
        node_pose, submap_pose = _f64(node_pose), _f64(submap_pose)
        high, low, hist = _f32(high).reshape(-1, 3), _f32(low).reshape(-1, 3), _f32(hist)
        g = _f64(gravity)
        out = np.zeros(17)
        lib().orc_fcsm_match(self.h, int(full), _p(node_pose), _p(submap_pose), _p(high), len(high),
                             _p(low), len(low), _p(hist), len(hist), _p(g), min_score, _p(out))
        return dict(zip(self.RESULT_FIELDS, out))

    def discrete_scans(self, wxy, wz, angular_window, node_pose, submap_pose, high, hist,
                       gravity=(1, 0, 0, 0)):
        node_pose, submap_pose = _f64(node_pose), _f64(submap_pose)
        high, hist, g = _f32(high).reshape(-1, 3), _f32(hist), _f64(gravity)
        S = lib().orc_fcsm_discrete_scans(self.h, wxy, wz, angular_window, _p(node_pose),
                                          _p(submap_pose), _p(high), len(high), _p(hist), len(hist),
                                          _p(g), 0, None, None, None)
        D = self.o["branch_and_bound_depth"]
        poses = np.zeros((S, 7), np.float32)
        rot = np.zeros(S, np.float32)
        cells = np.zeros((S, D, len(high), 3), np.int32)
        lib().orc_fcsm_discrete_scans(self.h, wxy, wz, angular_window, _p(node_pose), _p(submap_pose),
                                      _p(high), len(high), _p(hist), len(hist), _p(g), S, _p(poses),
                                      _p(rot), _p(cells))
        return poses, rot, cells

    def score_candidates(self, wxy, wz, angular_window, node_pose, submap_pose, high, hist, depth,
                         cand, gravity=(1, 0, 0, 0), threads=1, want_outputs=True):
        node_pose, submap_pose = _f64(node_pose), _f64(submap_pose)
        high, hist, g = _f32(high).reshape(-1, 3), _f32(hist), _f64(gravity)
        cand = _i32(cand).reshape(-1, 4)
        sums = np.zeros(len(cand), np.int32) if want_outputs else None
        scores = np.zeros(len(cand), np.float32) if want_outputs else None
        sec = lib().orc_fcsm_score_candidates(self.h, wxy, wz, angular_window, _p(node_pose),
                                              _p(submap_pose), _p(high), len(high), _p(hist),
                                              len(hist), _p(g), depth, len(cand), _p(cand),
                                              _p(sums), _p(scores), threads)
        return sums, scores, sec


def low_resolution_score(low_grid, pts, pose7):
    pts, pose7 = _f32(pts).reshape(-1, 3), _f32(pose7)
    return lib().orc_low_resolution_score(low_grid.h, _p(pts), len(pts), _p(pose7))


def voxel_filter(size, pts, stride=3):
    """Indices of the points sensor::VoxelFilter(size).Filter keeps (first point per voxel, input order)."""
    pts = _f32(pts).reshape(-1, stride)
    kept = np.zeros(len(pts), np.int32)
    n = lib().orc_voxel_filter(size, _p(pts), stride, len(pts), _p(kept))
    return kept[:n].copy()


def adaptive_voxel_filter(max_length, min_num_points, max_range, pts, stride=3):
    pts = _f32(pts).reshape(-1, stride)
    kept = np.zeros(len(pts), np.int32)
    n = lib().orc_adaptive_voxel_filter(max_length, min_num_points, max_range, _p(pts), stride, len(pts), _p(kept))
    return kept[:n].copy()
