"""TEST INFRASTRUCTURE ONLY (see oracle/README): numpy restatement of the BAL example
loader's geometry passes, to check csrc/bal_io.h.

Reference: examples/bal_problem.cc:226-300 (CameraToAngleAxisAndCenter,
AngleAxisAndCenterToCamera, Normalize), include/ceres/rotation.h:792-850
(AngleAxisRotatePoint). Plain loops over cameras: small inputs only.
"""
import numpy as np


def angle_axis_rotate_point(aa, pt):
    """rotation.h:792-850, including the first-order branch near zero."""
    aa, pt = np.asarray(aa, float), np.asarray(pt, float)
    theta2 = float(aa @ aa)
    if theta2 > np.finfo(float).eps:
        theta = np.sqrt(theta2)
        w = aa / theta
        c, s = np.cos(theta), np.sin(theta)
        return pt * c + np.cross(w, pt) * s + w * (w @ pt) * (1.0 - c)
    return pt + np.cross(aa, pt)


def camera_center(camera):
    """c = -R' t (bal_problem.cc:226-240)."""
    return -angle_axis_rotate_point(-camera[:3], camera[3:6])


def upper_median(v):
    """bal_problem.cc:65-69: nth_element at size / 2."""
    v = np.asarray(v)
    return np.partition(v, len(v) // 2)[len(v) // 2]


def normalize(cameras, points):
    """bal_problem.cc:257-300. Returns new (cameras [C x 9], points [P x 3])."""
    cameras = np.array(cameras, float).reshape(-1, 9)
    points = np.array(points, float).reshape(-1, 3)
    median = np.array([upper_median(points[:, i]) for i in range(3)])
    scale = 100.0 / upper_median(np.abs(points - median).sum(axis=1))
    points = scale * (points - median)
    for cam in cameras:
        center = scale * (camera_center(cam) - median)
        cam[3:6] = -angle_axis_rotate_point(cam[:3], center)  # t = -R c (:242-255)
    return cameras, points
