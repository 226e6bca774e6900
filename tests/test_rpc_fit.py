"""the RPC distortion fit (fitRpcDist / fitRpcUndist, rpc_distortion.cc:556-728).

CPU: the oracle (Ceres-rule central differences + dense LM, oracle/rpc_fit.h) on models it can represent exactly and on a
radtan lens.  GPU: mvba_rpc_fit against the oracle -- the fitted FUNCTION and the final cost are compared (the coefficients of
a ratio of polynomials are only determined up to the conditioning of the fit)."""
import numpy as np
import pytest

from multiview_b200 import _abi
from oracle.binding import compute_rpc


def rpc_eval(pts, c):
    """numpy restatement of compute_rpc's polynomial layout (rpc_distortion.cc:116-160), independent of the C code."""
    n = len(c); deg = int(round(np.sqrt(2.0 * n + 5.0) / 2.0 - 1.5)); T = (deg + 1) * (deg + 2) // 2
    mono = np.stack([pts[:, 0] ** (d - i) * pts[:, 1] ** i for d in range(deg + 1) for i in range(d + 1)], 1)
    num_x, den_x, num_y, den_y = c[:T], c[T:2 * T - 1], c[2 * T - 1:3 * T - 1], c[3 * T - 1:]
    fx = (mono @ num_x) / (1.0 + mono[:, 1:] @ den_x)
    fy = (mono @ num_y) / (1.0 + mono[:, 1:] @ den_y)
    return np.stack([fx, fy], 1)


def grid(ns, half=(640.0, 480.0)):
    xs = np.linspace(-half[0], half[0], ns); ys = np.linspace(-half[1], half[1], ns)
    X, Y = np.meshgrid(xs, ys, indexing="ij")
    return np.ascontiguousarray(np.stack([X.ravel(), Y.ravel()], 1))


def radtan(u, f=600.0, k=(-0.25, 0.08, 0.0012, -0.0008)):
    n = u / f; r2 = (n ** 2).sum(1)
    d = n * (1 + k[0] * r2 + k[1] * r2 * r2)[:, None]
    d[:, 0] += 2 * k[2] * n[:, 0] * n[:, 1] + k[3] * (r2 + 2 * n[:, 0] ** 2)
    d[:, 1] += k[2] * (r2 + 2 * n[:, 1] ** 2) + 2 * k[3] * n[:, 0] * n[:, 1]
    return np.ascontiguousarray(d * f)


def test_layout_matches_compute_rpc(oracle_lib):
    rng = np.random.default_rng(0)
    for deg in (1, 2, 4):
        c = rng.normal(size=_abi.rpc_num_params(deg)) * 1e-3 ** np.arange(_abi.rpc_num_params(deg)).clip(0, 3)
        pts = rng.uniform(-3, 3, (5, 2))
        assert np.allclose(compute_rpc(oracle_lib, pts, c), rpc_eval(pts, c), rtol=1e-13, atol=1e-13)


def test_oracle_recovers_an_affine_map_in_one_degree(oracle_lib):
    u = grid(20)
    A = np.array([[0.98, 0.01], [-0.02, 1.03]]); b = np.array([3.5, -2.25])
    c, st = _abi.rpc_fit(oracle_lib, "oracle_", u, u @ A.T + b, 1, num_threads=4)
    want = np.array([b[0], A[0, 0], A[0, 1], 0, 0, b[1], A[1, 0], A[1, 1], 0, 0])
    assert st["final_cost"] <= 1e-18 * st["initial_cost"]
    assert np.abs(c - want).max() <= 1e-9


def test_oracle_degree_continuation_on_a_radtan_lens(oracle_lib):
    u = grid(40); d = radtan(u)
    costs = []
    for deg in (1, 2, 3):
        c, st = _abi.rpc_fit(oracle_lib, "oracle_", u, d, deg, num_threads=8)
        assert len(c) == _abi.rpc_num_params(deg) and st["successful"] >= 1
        assert np.isclose(0.5 * ((rpc_eval(u, c) - d) ** 2).sum(), st["final_cost"], rtol=1e-9)
        costs.append(st["final_cost"])
    assert costs[0] > costs[1] > costs[2]                         # every degree starts from the previous fit
    assert np.sqrt(2 * costs[2] / len(u)) < 2.0                   # cubic ratio: under 2 px rms on a strongly distorted lens


def test_central_difference_of_a_ratio_has_the_closed_form_the_kernel_uses(oracle_lib):
    """rpc_fit.cuh evaluates d/d den_k as -N m_k / (D^2 - (delta_k m_k)^2) and d/d num_k as m_k / D instead of differencing:
    here the Ceres-rule quotient (f(c + delta) - f(c - delta)) / (2 delta), delta = max(|c| 1e-6, sqrt(eps)), is formed with the
    oracle's compute_rpc and compared -- including monomials for which delta * m_k is NOT small against D."""
    rng = np.random.default_rng(11)
    deg = 3; T = (deg + 1) * (deg + 2) // 2; n = _abi.rpc_num_params(deg)
    c = np.zeros(n); c[1] = 1.0; c[2 * T - 1 + 2] = 1.0
    c += rng.normal(size=n) * np.concatenate([10.0 ** -np.array([0, 1, 1] + [4] * 3 + [7] * 4), 10.0 ** -np.array([3, 3] + [6] * 3 + [9] * 4)] * 2)
    pts = rng.uniform(-600, 600, (40, 2))
    mono = np.stack([pts[:, 0] ** (d - i) * pts[:, 1] ** i for d in range(deg + 1) for i in range(d + 1)], 1)
    N = mono @ c[:T]; D = 1.0 + mono[:, 1:] @ c[T:2 * T - 1]
    delta = np.maximum(np.abs(c) * 1e-6, np.sqrt(np.finfo(float).eps))
    worst_linearisation_error = 0.0
    for k in list(range(0, T, 3)) + list(range(T, 2 * T - 1, 2)):           # x-numerator and x-denominator coefficients
        cp, cm = c.copy(), c.copy(); cp[k] += delta[k]; cm[k] -= delta[k]
        fd = (compute_rpc(oracle_lib, pts, cp)[:, 0] - compute_rpc(oracle_lib, pts, cm)[:, 0]) / (2 * delta[k])
        if k < T:
            closed = mono[:, k] / D
        else:
            m = mono[:, k - T + 1]
            closed = -N * m / (D * D - (delta[k] * m) ** 2)
            worst_linearisation_error = max(worst_linearisation_error, np.abs(closed / (-N * m / (D * D)) - 1).max())
        scale = np.abs(closed).max()
        assert np.abs(fd - closed).max() <= 1e-5 * scale + 1e-9 * np.abs(N / D).max() / delta[k]   # differencing noise: eps |f| / delta
    assert worst_linearisation_error > 1e-3      # the quotient is NOT the derivative for the high monomials: that is the point


@pytest.mark.parametrize("deg", [1, 2])
def test_oracle_fit_is_the_minimum_an_independent_lm_finds(deg, oracle_lib):
    """MINPACK's Levenberg-Marquardt (scipy.optimize.least_squares, method="lm") on the same residuals: from the identity model it
    reaches the oracle's cost, and started AT the oracle's coefficients it cannot improve them."""
    from scipy.optimize import least_squares
    u = grid(40); d = radtan(u)
    c, st = _abi.rpc_fit(oracle_lib, "oracle_", u, d, deg, num_iterations=200, num_threads=8)
    f = lambda x: (rpc_eval(u, x) - d).ravel()
    T = (deg + 1) * (deg + 2) // 2
    x0 = np.zeros(_abi.rpc_num_params(deg)); x0[1] = 1.0; x0[2 * T - 1 + 2] = 1.0
    kw = dict(method="lm", xtol=1e-15, ftol=1e-15, gtol=1e-15, max_nfev=20000, x_scale="jac")
    from_identity = least_squares(f, x0, **kw)
    polished = least_squares(f, c, **kw)
    assert abs(from_identity.cost - st["final_cost"]) <= 1e-5 * st["final_cost"]
    assert st["final_cost"] - polished.cost <= 1e-9 * st["final_cost"]


@pytest.mark.gpu
@pytest.mark.parametrize("deg", [1, 2, 3, 4])
def test_gpu_fit_follows_the_oracle(deg, oracle_lib, product_lib):
    u = grid(60); d = radtan(u)
    co, so = _abi.rpc_fit(oracle_lib, "oracle_", u, d, deg, num_threads=0)
    cg, sg = _abi.rpc_fit(product_lib, "mvba_", u, d, deg, jacobian_mode=0)
    assert np.isclose(sg["initial_cost"], so["initial_cost"], rtol=1e-12)
    # the oracle differentiates numerically (noise ~1e-6 relative in J); the kernel evaluates the same difference quotient in
    # closed form, so the two LM trajectories agree to that noise (a numpy replay of the kernel's arithmetic gives 1e-9 on the
    # cost, 1e-6 px on the fitted function, 5e-7 on the coefficients)
    assert abs(sg["final_cost"] - so["final_cost"]) <= 1e-6 * so["final_cost"]
    assert np.abs(rpc_eval(u, cg) - rpc_eval(u, co)).max() <= 1e-4
    assert np.abs(cg - co).max() <= 1e-4 * np.abs(co).max()


@pytest.mark.gpu
def test_gpu_exact_derivatives_fit_at_least_as_well(oracle_lib, product_lib):
    u = grid(60); d = radtan(u)
    _, so = _abi.rpc_fit(oracle_lib, "oracle_", u, d, 4, num_threads=0)
    cg, sg = _abi.rpc_fit(product_lib, "mvba_", u, d, 4, jacobian_mode=1)
    assert sg["final_cost"] <= 0.5 * so["final_cost"]        # the difference quotient stalls at degree 3; exact derivatives do not
    assert np.isclose(0.5 * ((rpc_eval(u, cg) - d) ** 2).sum(), sg["final_cost"], rtol=1e-9)


@pytest.mark.gpu
def test_gpu_recovers_an_exact_rational_model(product_lib):
    """Data generated by a degree-2 RPC are fitted to rounding level by the degree-2 fit with exact derivatives."""
    u = grid(50, half=(2.0, 1.5))
    c0 = np.zeros(22); c0[[1, 13]] = 1.0
    c0[[0, 3, 4, 6, 8]] = [0.01, 0.02, -0.03, 0.05, -0.04]; c0[[11, 14, 17, 19]] = [-0.02, 0.03, 0.01, 0.02]
    d = rpc_eval(u, c0)
    cg, sg = _abi.rpc_fit(product_lib, "mvba_", u, d, 2, jacobian_mode=1)
    assert sg["final_cost"] <= 1e-20 * len(u)
    assert np.abs(rpc_eval(u, cg) - d).max() <= 1e-9
