"""Minimal stand-in for the subset of spatialmath-python the reference touches
(se3_helpers.py:1,24-67; renderer.py:30,39,96): SO3 / SE3 with Rx/Rz/AngleAxis/Rt, `*`,
`.inv()` and `.data[0]`.  spatialmath is not installed on the target machines; the pose
sampler only needs these few closed forms."""
import numpy as np


def _rot_axis(axis, theta):
    c, s = np.cos(theta), np.sin(theta)
    if axis == 0:
        return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], dtype=np.float64)
    if axis == 1:
        return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=np.float64)


def _angle(theta, unit):
    return np.deg2rad(theta) if unit == "deg" else float(theta)


class SO3:
    def __init__(self, R=None, check=True):
        R = np.eye(3) if R is None else np.asarray(R, dtype=np.float64)
        if check:
            if R.shape != (3, 3) or not np.allclose(R @ R.T, np.eye(3), atol=1e-8) or np.linalg.det(R) < 0:
                raise ValueError("bad argument to constructor")
        self.data = [R]

    @property
    def R(self):
        return self.data[0]

    @classmethod
    def Rx(cls, theta, unit="rad"):
        return cls(_rot_axis(0, _angle(theta, unit)), check=False)

    @classmethod
    def Ry(cls, theta, unit="rad"):
        return cls(_rot_axis(1, _angle(theta, unit)), check=False)

    @classmethod
    def Rz(cls, theta, unit="rad"):
        return cls(_rot_axis(2, _angle(theta, unit)), check=False)

    @classmethod
    def AngleAxis(cls, theta, v, *, unit="rad"):
        """Rodrigues; the axis is normalised (spatialmath's angvec2r uses the unit vector)."""
        th = _angle(theta, unit)
        v = np.asarray(v, dtype=np.float64)
        n = np.linalg.norm(v)
        if n < 10 * np.finfo(np.float64).eps or abs(th) < 10 * np.finfo(np.float64).eps:
            return cls(np.eye(3), check=False)
        k = v / n
        Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
        return cls(np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * (Kx @ Kx), check=False)

    def __mul__(self, other):
        return SO3(self.R @ other.R, check=False)

    def inv(self):
        return SO3(self.R.T, check=False)


class SE3:
    def __init__(self, T=None, check=True):
        T = np.eye(4) if T is None else np.asarray(T, dtype=np.float64)
        if check and T.shape != (4, 4):
            raise ValueError("bad argument to constructor")
        self.data = [T]

    @property
    def A(self):
        return self.data[0]

    @property
    def R(self):
        return self.data[0][:3, :3]

    @property
    def t(self):
        return self.data[0][:3, 3]

    @classmethod
    def Rt(cls, R, t=None):
        T = np.eye(4)
        T[:3, :3] = R.R if isinstance(R, SO3) else np.asarray(R, dtype=np.float64)
        if t is not None:
            T[:3, 3] = np.asarray(t, dtype=np.float64).reshape(3)
        return cls(T, check=False)

    @classmethod
    def Rx(cls, theta, unit="rad"):
        return cls.Rt(SO3.Rx(theta, unit))

    @classmethod
    def Ry(cls, theta, unit="rad"):
        return cls.Rt(SO3.Ry(theta, unit))

    @classmethod
    def Rz(cls, theta, unit="rad"):
        return cls.Rt(SO3.Rz(theta, unit))

    def __mul__(self, other):
        return SE3(self.A @ other.A, check=False)

    def inv(self):
        T = np.eye(4)
        T[:3, :3] = self.R.T
        T[:3, 3] = -self.R.T @ self.t
        return SE3(T, check=False)

    def __repr__(self):
        return "SE3(\n%s)" % np.array2string(self.A, precision=6, suppress_small=True)
