"""ctypes mirror of include/mvba.h.  The same structs are accepted by libmvba.so (the product) and
by the CPU checker of the tests (oracle/liboracle.so, prefix ``oracle_``): its binding lives under oracle/ (oracle/binding.py,
test infrastructure) and plugs in through register_checker(); nothing in this package names or loads the checker."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "multiview_b200", "libmvba.so")

c_double_p = C.POINTER(C.c_double)
c_u8_p = C.POINTER(C.c_uint8)
c_i32_p = C.POINTER(C.c_int32)
c_i64_p = C.POINTER(C.c_int64)
c_f32_p = C.POINTER(C.c_float)


class Config(C.Structure):
    _fields_ = [("num_sensors", C.c_int32), ("num_ref_frames", C.c_int32), ("num_cameras", C.c_int32),
                ("num_landmarks", C.c_int32), ("no_rig", C.c_int32), ("affine_depth_to_image", C.c_int32),
                ("robust_threshold", C.c_double), ("depth_tri_weight", C.c_double), ("depth_mesh_weight", C.c_double)]


class Params(C.Structure):
    _fields_ = [(n, c_double_p) for n in ("world_to_ref", "ref_stamps", "ref_to_cam", "offsets", "offset_lo", "offset_hi",
                                          "focal", "center", "distortion", "depth_to_image", "depth_scale",
                                          "world_to_cam", "xyz")]


class FloatMask(C.Structure):
    _fields_ = [(n, c_u8_p) for n in ("focal", "center", "distortion", "rig", "depth_to_image", "poses")] + \
               [("float_timestamp_offsets", C.c_int32), ("float_scale", C.c_int32)]


class SolveOptions(C.Structure):
    _fields_ = [("max_num_iterations", C.c_int32), ("parameter_tolerance", C.c_double), ("function_tolerance", C.c_double),
                ("gradient_tolerance", C.c_double), ("initial_trust_region_radius", C.c_double),
                ("max_trust_region_radius", C.c_double), ("min_trust_region_radius", C.c_double),
                ("min_relative_decrease", C.c_double), ("min_lm_diagonal", C.c_double), ("max_lm_diagonal", C.c_double),
                ("eta", C.c_double), ("max_linear_solver_iterations", C.c_int32), ("min_linear_solver_iterations", C.c_int32),
                ("jacobi_scaling", C.c_int32), ("linear_solver", C.c_int32), ("preconditioner", C.c_int32), ("verbose", C.c_int32)]


class Summary(C.Structure):
    _fields_ = [("initial_cost", C.c_double), ("final_cost", C.c_double), ("num_iterations", C.c_int32),
                ("num_successful_steps", C.c_int32), ("num_linear_iterations", C.c_int32), ("termination", C.c_int32),
                ("num_jacobian_evals", C.c_int32), ("num_residual_evals", C.c_int32), ("num_kernel_launches", C.c_int64),
                ("seconds_total", C.c_double), ("seconds_device", C.c_double), ("seconds_build", C.c_double),
                ("seconds_linear", C.c_double), ("seconds_update", C.c_double), ("final_gradient_max_norm", C.c_double),
                ("final_radius", C.c_double), ("seconds_build_kernel", C.c_double), ("num_build_launches", C.c_int32),
                ("reserved", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class Iteration(C.Structure):
    _fields_ = [("cost", C.c_double), ("cost_change", C.c_double), ("gradient_max_norm", C.c_double), ("step_norm", C.c_double),
                ("relative_decrease", C.c_double), ("radius", C.c_double), ("linear_iterations", C.c_int32),
                ("step_accepted", C.c_int32)]


def default_options(**kw) -> SolveOptions:
    """ceres::Solver::Options as the reference sets them (rig_calibrator.cc:1910-1918) + Ceres 1.14 defaults."""
    o = SolveOptions(20, 1e-12, 1e-16, 1e-16, 1e4, 1e16, 1e-32, 1e-3, 1e-6, 1e32, 0.1, 500, 0, 1, 0, 0, 0)
    for k, v in kw.items():
        setattr(o, k, v)
    return o


def _p(a, ty):
    if a is None or (hasattr(a, "size") and a.size == 0):
        return C.cast(None, ty)
    return a.ctypes.data_as(ty)


class MvbaError(RuntimeError):
    pass


class Binding:
    """Thin object wrapper over one handle of either library."""

    def __init__(self, lib, prefix, problem, device=0):
        self.lib, self.prefix, self.pr = lib, prefix, problem
        self.h = None
        self._keep = []
        pr = problem
        T = len(pr.sensors)
        cfg = Config(T, len(pr.ref_stamps), len(pr.cam_sensor), pr.num_landmarks, int(pr.no_rig), int(pr.affine_depth),
                     pr.robust_threshold, pr.depth_tri_weight, pr.depth_mesh_weight)
        self._create(cfg, device)
        f = self._f
        for t, s in enumerate(pr.sensors):
            self._ck(f("set_sensor")(self.h, C.c_int32(t), C.c_double(s.width), C.c_double(s.height), C.c_int32(s.nd)))
        prm = Params(_p(pr.world_to_ref, c_double_p), _p(pr.ref_stamps, c_double_p), _p(pr.ref_to_cam, c_double_p),
                     _p(pr.offsets, c_double_p), _p(pr.offset_lo, c_double_p), _p(pr.offset_hi, c_double_p),
                     _p(pr.focal, c_double_p), _p(pr.center, c_double_p), _p(pr.distortion, c_double_p),
                     _p(pr.depth_to_image, c_double_p), _p(pr.depth_scale, c_double_p), _p(pr.world_to_cam, c_double_p),
                     _p(pr.xyz, c_double_p))
        self._keep.append(prm)
        self._ck(f("set_params")(self.h, C.byref(prm)))
        self._ck(f("set_cameras")(self.h, C.c_int32(len(pr.cam_sensor)), _p(pr.cam_sensor, c_i32_p), _p(pr.cam_beg, c_i32_p),
                                  _p(pr.cam_end, c_i32_p), _p(pr.cam_stamp, c_double_p)))
        self._ck(f("set_observations")(self.h, C.c_int64(pr.num_obs), _p(pr.pid, c_i32_p), _p(pr.cid, c_i32_p), _p(pr.pix, c_f32_p),
                                       _p(pr.has_depth, c_u8_p), _p(pr.depth_xyz, c_f32_p), _p(pr.has_mesh, c_u8_p),
                                       _p(pr.mesh_xyz, c_double_p)))
        for kind, (ppid, ref, w, th) in sorted(pr.priors.items()):
            self._keep += [ppid, ref]
            self._ck(f("set_point_priors")(self.h, C.c_int32(kind), C.c_int64(len(ppid)), _p(ppid, c_i32_p), _p(ref, c_double_p),
                                           C.c_double(w), C.c_double(th)))
        m = FloatMask(_p(pr.float_focal, c_u8_p), _p(pr.float_center, c_u8_p), _p(pr.float_dist, c_u8_p), _p(pr.float_rig, c_u8_p),
                      _p(pr.float_d2i, c_u8_p), _p(pr.float_poses, c_u8_p), int(pr.float_offsets), int(pr.float_scale))
        self._keep.append(m)
        self._ck(f("set_float_mask")(self.h, C.byref(m)))

    def __new__(cls, lib, prefix="mvba_", *args, **kw):
        # a prefix other than the product's selects a registered checker binding (a subclass; the tests register the CPU oracle's)
        if cls is Binding and prefix != "mvba_":
            if prefix not in _CHECKERS:
                raise MvbaError(f"no binding registered for prefix {prefix!r} (the tests register the CPU checker: oracle/binding.py)")
            return object.__new__(_CHECKERS[prefix])
        return object.__new__(cls)

    def _create(self, cfg, device):
        h = C.c_void_p()
        self._ck(self.lib.mvba_create(C.byref(cfg), int(device), C.byref(h)))
        self.h = h

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    def _ck(self, rc):
        if rc != 0:
            msg = ""
            if self.prefix == "mvba_":
                self.lib.mvba_last_error.restype = C.c_char_p
                msg = (self.lib.mvba_last_error() or b"").decode()
            raise MvbaError(f"{self.prefix} call failed rc={rc}: {msg}")

    def close(self):
        if self.h is not None:
            self._f("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- queries -------------------------------------------------------------------------
    def num_residuals(self):
        a, b = C.c_int64(), C.c_int64()
        self._ck(self._f("num_residuals")(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def residual_layout(self):
        _, nb = self.num_residuals()
        kind = np.zeros(nb, np.int32); row0 = np.zeros(nb, np.int64); src = np.zeros(nb, np.int64)
        self._ck(self._f("residual_layout")(self.h, _p(kind, c_i32_p), _p(row0, c_i64_p), _p(src, c_i64_p)))
        return kind, row0, src

    def evaluate(self, apply_loss=False, want_residuals=True):
        nres, _ = self.num_residuals()
        cost = C.c_double()
        r = np.zeros(nres) if want_residuals else None
        self._ck(self._f("evaluate")(self.h, C.c_int32(int(apply_loss)), C.byref(cost), _p(r, c_double_p)))
        return cost.value, r

    def jacobian_width(self, kind):
        w = C.c_int32()
        self._ck(self._f("jacobian_width")(self.h, C.c_int32(kind), C.byref(w)))
        return w.value

    def evaluate_jacobian(self, kind, first, count, mode=1):
        W = self.jacobian_width(kind)
        rows = 2 if kind == 0 else 3
        r = np.zeros((count, rows)); J = np.zeros((count, rows, W))
        self._evaluate_jacobian(kind, first, count, mode, r, J)
        return r, J

    def _evaluate_jacobian(self, kind, first, count, mode, r, J):
        self._ck(self.lib.mvba_evaluate_jacobian(self.h, C.c_int32(kind), C.c_int64(first), C.c_int64(count), _p(r, c_double_p), _p(J, c_double_p)))

    # -- per-track steps around a pass (N1, N2) ------------------------------------------
    def triangulate(self, inlier=None, xyz0=None):
        """multiViewTriangulation at the current parameters: returns (xyz[P,3], inlier[N])."""
        pr = self.pr
        inl = np.ones(pr.num_obs, np.uint8) if inlier is None else np.ascontiguousarray(inlier, np.uint8).copy()
        xyz = np.zeros(3 * pr.num_landmarks) if xyz0 is None else np.ascontiguousarray(xyz0, np.float64).ravel().copy()
        self._ck(self._f("triangulate")(self.h, _p(inl, c_u8_p), _p(xyz, c_double_p)))
        return xyz.reshape(-1, 3), inl

    def flag_outliers(self, min_triangulation_angle, max_reprojection_error, inlier=None):
        """flagOutliersByTriAngleAndReprojErr at the current parameters: returns (inlier[N], counts[4])."""
        inl = np.ones(self.pr.num_obs, np.uint8) if inlier is None else np.ascontiguousarray(inlier, np.uint8).copy()
        counts = np.zeros(4, np.int64)
        self._ck(self._f("flag_outliers")(self.h, C.c_double(min_triangulation_angle), C.c_double(max_reprojection_error),
                                          _p(inl, c_u8_p), _p(counts, c_i64_p)))
        return inl, counts

    def flag_exclusion(self, image_size, crop_size, inlier=None):
        """flagOutlierByExclusionDist: image_size / crop_size are [T,2]."""
        inl = np.ones(self.pr.num_obs, np.uint8) if inlier is None else np.ascontiguousarray(inlier, np.uint8).copy()
        a = np.ascontiguousarray(image_size, np.float64).ravel(); b = np.ascontiguousarray(crop_size, np.float64).ravel()
        self._ck(self._f("flag_exclusion")(self.h, _p(a, c_double_p), _p(b, c_double_p), _p(inl, c_u8_p)))
        return inl

    def mesh_intersect(self, vertices, faces, min_ray_dist, max_ray_dist, inlier=None):
        """meshTriangulations at the current parameters: (has_mesh[N], mesh_xyz[N,3], pid_mesh_xyz[P,3])."""
        pr = self.pr
        inl = np.ones(pr.num_obs, np.uint8) if inlier is None else np.ascontiguousarray(inlier, np.uint8)
        v = np.ascontiguousarray(vertices, np.float32); f = np.ascontiguousarray(faces, np.uint32)
        has = np.zeros(pr.num_obs, np.uint8); xyz = np.zeros(3 * pr.num_obs); mean = np.zeros(3 * pr.num_landmarks)
        self._ck(self._f("mesh_intersect")(self.h, C.c_int64(len(v)), _p(v, c_f32_p), C.c_int64(len(f)), f.ctypes.data_as(C.POINTER(C.c_uint32)),
                                           C.c_double(min_ray_dist), C.c_double(max_ray_dist), _p(inl, c_u8_p), _p(has, c_u8_p),
                                           _p(xyz, c_double_p), _p(mean, c_double_p)))
        return has, xyz.reshape(-1, 3), mean.reshape(-1, 3)

    def build_plan(self):
        """mvba_build_plan of the last solve as a dict (product library only)."""
        plan = (C.c_int32 * 8)()
        self._ck(self.lib.mvba_build_plan(self.h, plan))
        keys = ("pass_a_accumulators", "pass_b", "blocks_per_warp", "tracks_per_round", "tiles", "rounds", "fallback_tracks", "l8max")
        return dict(zip(keys, list(plan)))

    def solve(self, options: SolveOptions | None = None, max_log=256, jac_mode=1, num_threads=0):
        o = options or default_options()
        s = Summary()
        its = (Iteration * max_log)()
        self._solve(o, s, its, max_log, jac_mode, num_threads)
        n = min(max_log, s.num_iterations + 1)
        log = [{k: getattr(its[i], k) for k, _ in Iteration._fields_} for i in range(n)]
        return s, log


    def _solve(self, o, s, its, max_log, jac_mode, num_threads):
        self._ck(self.lib.mvba_solve(self.h, C.byref(o), C.byref(s), its, C.c_int32(max_log)))


_CHECKERS = {}      # prefix -> Binding subclass
_RPC_FITTERS = {}   # prefix -> function(lib, a, b, deg, num_iterations, parameter_tolerance, num_threads, c, st) -> rc


def register_checker(prefix, binding_cls, rpc_fit_fn=None):
    """Test infrastructure hook: a library with another symbol prefix (the CPU oracle) binds through its own subclass."""
    _CHECKERS[prefix] = binding_cls
    if rpc_fit_fn is not None:
        _RPC_FITTERS[prefix] = rpc_fit_fn


_LIB = None


def load_product():
    """Loads multiview_b200/libmvba.so.  Fails loudly when it is missing: there is no fallback."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise MvbaError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                            "(there is no CPU or PyTorch fallback for this path)")
        _LIB = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        _LIB.mvba_last_error.restype = C.c_char_p
    return _LIB


def bind_product(problem, device=0) -> Binding:
    return Binding(load_product(), "mvba_", problem, device)


def rpc_num_params(deg: int) -> int:
    return 2 * (deg + 1) * (deg + 2) - 2


def rpc_fit(lib, prefix, in_xy, out_xy, deg, num_iterations=100, parameter_tolerance=1e-12, jacobian_mode=0, num_threads=0, device=0):
    """fitRpcDist-style fit of out = rpc(in) through `mvba_rpc_fit` (product) or a registered checker:
    (coeffs, {initial_cost, final_cost, iterations, successful})."""
    a = np.ascontiguousarray(in_xy, np.float64); b = np.ascontiguousarray(out_xy, np.float64)
    assert a.shape == b.shape and a.ndim == 2 and a.shape[1] == 2
    c = np.zeros(rpc_num_params(deg)); st = np.zeros(4)
    if prefix == "mvba_":
        rc = lib.mvba_rpc_fit(C.c_int32(device), C.c_int64(len(a)), _p(a, c_double_p), _p(b, c_double_p), C.c_int32(deg), C.c_int32(num_iterations),
                              C.c_double(parameter_tolerance), C.c_int32(jacobian_mode), _p(c, c_double_p), _p(st, c_double_p))
        if rc != 0:
            raise MvbaError(f"mvba_rpc_fit failed ({rc}): {lib.mvba_last_error().decode()}")
    else:
        if prefix not in _RPC_FITTERS:
            raise MvbaError(f"no RPC fitter registered for prefix {prefix!r}")
        rc = _RPC_FITTERS[prefix](lib, a, b, deg, num_iterations, parameter_tolerance, num_threads, c, st)
        if rc != 0:
            raise MvbaError(f"{prefix}rpc_fit failed ({rc})")
    return c, dict(initial_cost=st[0], final_cost=st[1], iterations=int(st[2]), successful=int(st[3]))
