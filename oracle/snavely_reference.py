"""ORACLE — TEST INFRASTRUCTURE ONLY. Never imported by the product.

Independent numpy restatement of the bundle-adjustment residual and its automatic derivatives, written
from the reference's sources and sharing NO header with the product's csrc/snavely_model.h:

  SnavelyReprojectionError::operator()            examples/snavely_reprojection_error.h:57-92
  AngleAxisRotatePoint                            include/ceres/rotation.h:792-850
  ceres::Jet<double, 12> arithmetic (+ - * /, sin, cos, hypot(x, y, z), fpclassify on the scalar part)
                                                  include/ceres/jet.h
  AutoDiffCostFunction<SnavelyReprojectionError, 2, 9, 3>: partials 0..8 = camera, 9..11 = point
  ProgramEvaluator::Evaluate + BlockJacobianWriter layout (all E cells 2x3, then all F cells 2x9, rows
  grouped by point)                               program_evaluator.h:189-260, block_jacobian_writer.cc:102-165
  BlockSparseMatrix::SquaredColumnNorm / ScaleColumns / RightMultiplyAndAccumulate
                                                  block_sparse_matrix.cc:351-450, :239-274

Pinned by: two hand-computed known-answer projections (tests/test_oracle_kat.py), complex-step
differentiation of the plain residual (which needs no Jet arithmetic at all), and the near-zero
rotation branch switching at EXACTLY zero (rotation.h:799-801: fpclassify(theta) != FP_ZERO).
"""
import numpy as np


class Jet:
    """A batch of ceres::Jet<double, K>: a [n] scalar parts, v [n, K] infinitesimal parts."""

    __slots__ = ("a", "v")

    def __init__(self, a, v):
        self.a, self.v = a, v

    @staticmethod
    def variable(a, k, K):
        v = np.zeros((len(a), K))
        v[:, k] = 1.0
        return Jet(np.asarray(a, dtype=np.float64).copy(), v)

    def _lift(self, o):
        if isinstance(o, Jet):
            return o
        return Jet(np.broadcast_to(np.asarray(o, dtype=np.float64), self.a.shape).copy(), np.zeros_like(self.v))

    def __add__(self, o):
        o = self._lift(o)
        return Jet(self.a + o.a, self.v + o.v)

    __radd__ = __add__

    def __sub__(self, o):
        o = self._lift(o)
        return Jet(self.a - o.a, self.v - o.v)

    def __rsub__(self, o):
        return self._lift(o) - self

    def __neg__(self):
        return Jet(-self.a, -self.v)

    def __mul__(self, o):
        o = self._lift(o)  # jet.h: (f g).v = f.a g.v + f.v g.a
        return Jet(self.a * o.a, self.a[:, None] * o.v + self.v * o.a[:, None])

    __rmul__ = __mul__

    def __truediv__(self, o):
        o = self._lift(o)  # jet.h: g_inv = 1/g.a; q = f.a g_inv; (f/g).v = (f.v - q g.v) g_inv
        g_inv = 1.0 / o.a
        q = self.a * g_inv
        return Jet(q, (self.v - q[:, None] * o.v) * g_inv[:, None])

    def __rtruediv__(self, o):
        return self._lift(o) / self


def jet_sin(f):
    return Jet(np.sin(f.a), np.cos(f.a)[:, None] * f.v)


def jet_cos(f):
    return Jet(np.cos(f.a), -np.sin(f.a)[:, None] * f.v)


def jet_hypot3(x, y, z):
    """jet.h hypot(x, y, z): value std::hypot, partials (x.a/h) x.v + (y.a/h) y.v + (z.a/h) z.v."""
    h = np.sqrt(x.a * x.a + y.a * y.a + z.a * z.a)
    with np.errstate(divide="ignore", invalid="ignore"):
        v = (x.a / h)[:, None] * x.v + (y.a / h)[:, None] * y.v + (z.a / h)[:, None] * z.v
    return Jet(h, v)


def _select(mask, f, g):
    return Jet(np.where(mask, f.a, g.a), np.where(mask[:, None], f.v, g.v))


def angle_axis_rotate_point(aa, pt):
    """rotation.h:792-850 on batches of Jets; the Taylor branch is taken only where theta == 0 exactly."""
    theta = jet_hypot3(aa[0], aa[1], aa[2])
    nonzero = theta.a != 0.0  # fpclassify(theta) != FP_ZERO (subnormals count as non-zero)
    safe = Jet(np.where(nonzero, theta.a, 1.0), theta.v)
    safe.v = np.where(nonzero[:, None], safe.v, 0.0)
    costheta, sintheta = jet_cos(safe), jet_sin(safe)
    theta_inverse = 1.0 / safe
    w = [aa[k] * theta_inverse for k in range(3)]
    w_cross_pt = [w[1] * pt[2] - w[2] * pt[1], w[2] * pt[0] - w[0] * pt[2], w[0] * pt[1] - w[1] * pt[0]]
    tmp = (w[0] * pt[0] + w[1] * pt[1] + w[2] * pt[2]) * (1.0 - costheta)
    far = [pt[k] * costheta + w_cross_pt[k] * sintheta + w[k] * tmp for k in range(3)]
    a_cross_pt = [aa[1] * pt[2] - aa[2] * pt[1], aa[2] * pt[0] - aa[0] * pt[2], aa[0] * pt[1] - aa[1] * pt[0]]
    near = [pt[k] + a_cross_pt[k] for k in range(3)]
    return [_select(nonzero, far[k], near[k]) for k in range(3)]


def snavely_residual_jets(camera, point, observed):
    """snavely_reprojection_error.h:57-92 with T = Jet<double, 12>. camera [n, 9], point [n, 3],
    observed [n, 2] -> two Jets (residual x, residual y)."""
    cam = [Jet.variable(camera[:, k], k, 12) for k in range(9)]
    pt = [Jet.variable(point[:, k], 9 + k, 12) for k in range(3)]
    p = angle_axis_rotate_point(cam[0:3], pt)
    p = [p[0] + cam[3], p[1] + cam[4], p[2] + cam[5]]
    xp = -p[0] / p[2]
    yp = -p[1] / p[2]
    l1, l2 = cam[7], cam[8]
    r2 = xp * xp + yp * yp
    distortion = 1.0 + r2 * (l1 + l2 * r2)
    focal = cam[6]
    predicted_x = focal * distortion * xp
    predicted_y = focal * distortion * yp
    return predicted_x - observed[:, 0], predicted_y - observed[:, 1]


def snavely_residual_plain(camera, point, observed):
    """The same operator with T = double (or complex, for complex-step differentiation away from the
    theta == 0 branch). camera [n, 9], point [n, 3] -> residuals [n, 2]."""
    aa = camera[:, 0:3]
    theta = np.sqrt((aa * aa).sum(axis=1))
    costheta, sintheta = np.cos(theta), np.sin(theta)
    w = aa / theta[:, None]
    w_cross_pt = np.cross(w, point)
    tmp = (w * point).sum(axis=1) * (1.0 - costheta)
    p = point * costheta[:, None] + w_cross_pt * sintheta[:, None] + w * tmp[:, None] + camera[:, 3:6]
    xp = -p[:, 0] / p[:, 2]
    yp = -p[:, 1] / p[:, 2]
    r2 = xp * xp + yp * yp
    distortion = 1.0 + r2 * (camera[:, 7] + camera[:, 8] * r2)
    return np.stack([camera[:, 6] * distortion * xp, camera[:, 6] * distortion * yp], axis=1) - observed


def evaluate(point_row_start, cam_idx, observations, cameras, points, jacobian=True, chunk=1 << 18):
    """Evaluator::Evaluate for a BAL problem whose rows are grouped by point.
    Returns (cost = 1/2 |r|^2, residuals [2N], values [24N] = [E (N x 6) | F (N x 18)] or None)."""
    prs = np.asarray(point_row_start, dtype=np.int64)
    cam_idx = np.asarray(cam_idx, dtype=np.int64)
    N, P = len(cam_idx), len(prs) - 1
    row_point = np.repeat(np.arange(P, dtype=np.int64), np.diff(prs))
    cams = np.asarray(cameras, dtype=np.float64).reshape(-1, 9)
    pts = np.asarray(points, dtype=np.float64).reshape(-1, 3)
    obs = np.asarray(observations, dtype=np.float64).reshape(-1, 2)
    residuals = np.empty(2 * N)
    values = np.empty(24 * N) if jacobian else None
    for s in range(0, N, chunk):
        e = min(N, s + chunk)
        c, x, o = cams[cam_idx[s:e]], pts[row_point[s:e]], obs[s:e]
        if jacobian:
            rx, ry = snavely_residual_jets(c, x, o)
            residuals[2 * s:2 * e:2], residuals[2 * s + 1:2 * e:2] = rx.a, ry.a
            E = values[6 * s:6 * e].reshape(-1, 2, 3)
            F = values[6 * N + 18 * s:6 * N + 18 * e].reshape(-1, 2, 9)
            E[:, 0, :], E[:, 1, :] = rx.v[:, 9:12], ry.v[:, 9:12]
            F[:, 0, :], F[:, 1, :] = rx.v[:, 0:9], ry.v[:, 0:9]
        else:
            zero_aa = (c[:, 0:3] == 0.0).all(axis=1)
            if zero_aa.any():  # the plain form divides by theta: route exact zeros through the Jets
                rx, ry = snavely_residual_jets(c, x, o)
                r = np.stack([rx.a, ry.a], axis=1)
            else:
                r = snavely_residual_plain(c, x, o)
            residuals[2 * s:2 * e] = r.reshape(-1)
    return 0.5 * float(residuals @ residuals), residuals, values


class ReferenceProblem:
    """Everything oracle/lm_reference.py needs from a bundle-adjustment problem, in numpy, with no
    product code behind it: evaluate, squared_column_norm, scale_columns, jacobian_times."""

    def __init__(self, point_row_start, cam_idx, num_cameras, observations, cameras=None, points=None):
        self.point_row_start = np.asarray(point_row_start, dtype=np.int64)
        self.cam_idx = np.asarray(cam_idx, dtype=np.int64)
        self.C, self.P, self.N = int(num_cameras), len(self.point_row_start) - 1, len(self.cam_idx)
        self.observations = np.asarray(observations, dtype=np.float64)
        self.cameras, self.points = cameras, points
        self.row_point = np.repeat(np.arange(self.P, dtype=np.int64), np.diff(self.point_row_start))

    @classmethod
    def of(cls, p):
        """From any object with point_row_start / cam_idx / C / observations / cameras / points."""
        return cls(p.point_row_start, p.cam_idx, p.C, p.observations, p.cameras, p.points)

    def evaluate(self, cameras=None, points=None, jacobian=True):
        return evaluate(self.point_row_start, self.cam_idx, self.observations,
                        self.cameras if cameras is None else cameras,
                        self.points if points is None else points, jacobian=jacobian)

    def _blocks(self, values):
        N = self.N
        return values[:6 * N].reshape(N, 2, 3), values[6 * N:].reshape(N, 2, 9)

    def squared_column_norm(self, values):  # block_sparse_matrix.cc:351-391
        E, F = self._blocks(values)
        out = np.zeros(3 * self.P + 9 * self.C)
        e2, f2 = (E * E).sum(axis=1), (F * F).sum(axis=1)
        for k in range(3):
            out[k:3 * self.P:3] = np.bincount(self.row_point, weights=e2[:, k], minlength=self.P)
        for k in range(9):
            out[3 * self.P + k::9] = np.bincount(self.cam_idx, weights=f2[:, k], minlength=self.C)
        return out

    def scale_columns(self, values, scale):  # block_sparse_matrix.cc:393-450
        E, F = self._blocks(values)
        E *= scale[:3 * self.P].reshape(self.P, 3)[self.row_point][:, None, :]
        F *= scale[3 * self.P:].reshape(self.C, 9)[self.cam_idx][:, None, :]
        return values

    def jacobian_times(self, values, x):  # block_sparse_matrix.cc:239-274
        E, F = self._blocks(values)
        xe = x[:3 * self.P].reshape(self.P, 3)[self.row_point]
        xf = x[3 * self.P:].reshape(self.C, 9)[self.cam_idx]
        return (np.einsum("nij,nj->ni", E, xe) + np.einsum("nij,nj->ni", F, xf)).reshape(-1)
