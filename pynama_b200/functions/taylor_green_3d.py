"""2-D Taylor-Green field embedded in 3-D (reference: src/functions/taylor_green_3d.py:4-29)."""
from math import pi

import numpy as np

from ._taylor_green_common import decay

Lx = Ly = 1


def alpha(nu, t):
    return decay(nu, t, 2)


def velocity(coord, alpha):
    x_ = 2 * pi * coord[:, 0] / Lx
    y_ = 2 * pi * coord[:, 1] / Ly
    vel = np.zeros(coord.shape)
    vel[:, 0] = np.cos(x_) * np.sin(y_) * Lx * alpha
    vel[:, 1] = -np.sin(x_) * np.cos(y_) * Ly * alpha
    return vel.flatten()


def vorticity(coord, alpha):
    x_ = 2 * pi * coord[:, 0] / Lx
    y_ = 2 * pi * coord[:, 1] / Ly
    vort = np.zeros(coord.shape)
    vort[:, 2] = -2 * pi * (Ly / Lx + Lx / Ly) * np.cos(x_) * np.cos(y_) * alpha
    return vort.flatten()
