"""2-D Taylor-Green vortex (reference: src/functions/taylor_green.py:4-36)."""
from math import cos, exp, pi, sin

import numpy as np

from ._taylor_green_common import decay

Lx = Ly = 1


def alpha(nu, t):
    return decay(nu, t, 2)


def velocity(coords, alpha=1):
    x_ = 2 * pi * coords[:, 0] / Lx
    y_ = 2 * pi * coords[:, 1] / Ly
    vel = np.zeros(coords.shape)
    vel[:, 0] = np.cos(x_) * np.sin(y_) * alpha
    vel[:, 1] = -np.sin(x_) * np.cos(y_) * alpha
    return vel.flatten()


def vorticity(coord, alpha):
    x_ = 2 * pi * coord[:, 0] / Lx
    y_ = 2 * pi * coord[:, 1] / Ly
    return -2 * pi * (1.0 / Lx + 1.0 / Ly) * np.cos(x_) * np.cos(y_) * alpha


def velocity_test(coord, t, nu):
    a = decay(nu, t, 2)
    x_, y_ = 2 * pi * coord[0] / Lx, 2 * pi * coord[1] / Ly
    return [cos(x_) * sin(y_) * a, -sin(x_) * cos(y_) * a]


def vorticity_test(coord, nu, t):
    a = decay(nu, t, 2)
    x_, y_ = 2 * pi * coord[0] / Lx, 2 * pi * coord[1] / Ly
    return [-2 * pi * (1.0 / Lx + 1.0 / Ly) * cos(x_) * cos(y_) * a]


def convective(coord, alpha):
    pass


def diffusive(coord, alpha):
    pass
