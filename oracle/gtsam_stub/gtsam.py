"""Minimal pure-numpy stand-in for the `gtsam` wheel (absent from this image), enough to import the
reference's verifier (`gtsfm/frontend/verifier/*`, `gtsfm/utils/{features,verification}.py`) and its
tests.  ORACLE-side test infrastructure: put this directory on PYTHONPATH *only* when the real gtsam
is not importable.  Semantics follow gtsam 4.2 (environment_linux.yml:55)."""
import numpy as np


class Cal3Bundler:
    def __init__(self, fx=1.0, k1=0.0, k2=0.0, u0=0.0, v0=0.0):
        self._f, self._k1, self._k2, self._u0, self._v0 = float(fx), float(k1), float(k2), float(u0), float(v0)

    def fx(self): return self._f
    def fy(self): return self._f
    def k1(self): return self._k1
    def k2(self): return self._k2
    def px(self): return self._u0
    def py(self): return self._v0

    def K(self):
        return np.array([[self._f, 0.0, self._u0], [0.0, self._f, self._v0], [0.0, 0.0, 1.0]])

    def uncalibrate(self, p):
        p = np.asarray(p, float).ravel()
        r2 = p[0] * p[0] + p[1] * p[1]
        g = 1.0 + (self._k1 + self._k2 * r2) * r2
        return np.array([self._f * g * p[0] + self._u0, self._f * g * p[1] + self._v0])

    def calibrate(self, pi):
        pi = np.asarray(pi, float).ravel()
        px, py = (pi[0] - self._u0) / self._f, (pi[1] - self._v0) / self._f
        rr = px * px + py * py
        g = 1.0 + (self._k1 + self._k2 * rr) * rr
        pn = np.array([px / g, py / g])
        if self._k1 != 0.0 or self._k2 != 0.0:
            for _ in range(10):
                d = self.uncalibrate(pn) - pi
                if d @ d <= 1e-5:
                    break
                r2 = pn @ pn
                g = 1.0 + (self._k1 + self._k2 * r2) * r2
                pn = np.array([px / g, py / g])
        return pn


class Rot3:
    def __init__(self, R=None):
        self._R = np.eye(3) if R is None else np.array(R, float).reshape(3, 3)

    @staticmethod
    def RzRyRx(x, y, z):
        cx, sx, cy, sy, cz, sz = np.cos(x), np.sin(x), np.cos(y), np.sin(y), np.cos(z), np.sin(z)
        Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
        Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
        Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
        return Rot3(Rz @ Ry @ Rx)

    def matrix(self): return self._R.copy()
    def inverse(self): return Rot3(self._R.T)
    def compose(self, o): return Rot3(self._R @ o._R)
    def between(self, o): return Rot3(self._R.T @ o._R)
    def rotate(self, p): return self._R @ np.asarray(p, float).ravel()

    def axisAngle(self):
        ang = np.arccos(np.clip((np.trace(self._R) - 1) / 2, -1, 1))
        ax = np.array([self._R[2, 1] - self._R[1, 2], self._R[0, 2] - self._R[2, 0], self._R[1, 0] - self._R[0, 1]])
        n = np.linalg.norm(ax)
        return Unit3(ax / n if n > 0 else np.array([1.0, 0, 0])), float(ang)


class Unit3:
    def __init__(self, v=(1.0, 0.0, 0.0)):
        v = np.asarray(v, float).ravel()
        self._v = v / np.linalg.norm(v)

    def point3(self): return self._v.copy()
    def dot(self, o): return float(self._v @ o._v)


class Pose3:
    def __init__(self, R=None, t=(0.0, 0.0, 0.0)):
        self._R = R if isinstance(R, Rot3) else Rot3(R)
        self._t = np.asarray(t, float).ravel()

    def rotation(self): return self._R
    def translation(self): return self._t.copy()
    def inverse(self): return Pose3(self._R.inverse(), -(self._R.matrix().T @ self._t))
    def compose(self, o): return Pose3(self._R.compose(o._R), self._t + self._R.matrix() @ o._t)
    def between(self, o): return self.inverse().compose(o)
    def transformTo(self, p): return self._R.matrix().T @ (np.asarray(p, float).ravel() - self._t)


class EssentialMatrix:
    def __init__(self, R, U):
        self._R, self._U = R, U

    def rotation(self): return self._R
    def direction(self): return self._U

    def matrix(self):
        t = self._U.point3()
        tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
        return tx @ self._R.matrix()


class PinholeCameraCal3Bundler:
    def __init__(self, pose=None, cal=None):
        self._pose = pose or Pose3()
        self._cal = cal or Cal3Bundler()

    def pose(self): return self._pose
    def calibration(self): return self._cal

    def project(self, p):
        q = self._pose.transformTo(p)
        return self._cal.uncalibrate(q[:2] / q[2])


Point3 = np.array
Point2 = np.array


class _Placeholder:
    """Names imported (but not exercised) by reference modules on the verifier tests' import path."""

    def __init__(self, *a, **k):
        raise NotImplementedError("gtsam stub: not needed on the two-view hot path")


class Pose3Pairs(_Placeholder): pass
class Rot3Vector(_Placeholder): pass
class Similarity3(_Placeholder): pass
class Pose3Vector(_Placeholder): pass
class SfmTrack(_Placeholder): pass
class Values(_Placeholder): pass
class NonlinearFactorGraph(_Placeholder): pass
