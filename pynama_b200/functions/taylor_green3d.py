"""3-D Taylor-Green field (reference: src/functions/taylor_green3d.py:4-55)."""
from math import pi

import numpy as np

from ._taylor_green_common import decay

Lx = Ly = Lz = 1


def alpha(nu, t):
    return decay(nu, t, 3)


def _xyz(coord):
    return 2 * pi * coord[:, 0] / Lx, 2 * pi * coord[:, 1] / Ly, 2 * pi * coord[:, 2] / Lz


def velocity(coord, alpha):
    x_, y_, z_ = _xyz(coord)
    vel = np.zeros(coord.shape)
    vel[:, 0] = np.cos(x_) * np.sin(y_) * np.sin(z_) * Lx * alpha
    vel[:, 1] = np.sin(x_) * np.cos(y_) * np.sin(z_) * Ly * alpha
    vel[:, 2] = -2 * np.sin(x_) * np.sin(y_) * np.cos(z_) * Lz * alpha
    return vel.flatten()


def vorticity(coord, alpha):
    x_, y_, z_ = _xyz(coord)
    vort = np.zeros(coord.shape)
    vort[:, 0] = -2 * pi * (Ly / Lz + 2 * Lz / Ly) * np.sin(x_) * np.cos(y_) * np.cos(z_) * alpha
    vort[:, 1] = 2 * pi * (Lx / Lz + 2 * Lz / Lx) * np.cos(x_) * np.sin(y_) * np.cos(z_) * alpha
    vort[:, 2] = 2 * pi * (Ly / Lx - Lx / Ly) * np.cos(x_) * np.cos(y_) * np.sin(z_) * alpha
    return vort.flatten()


def convective(coord, alpha):
    x_, y_, z_ = _xyz(coord)
    k = (2 * pi * alpha) ** 2
    conv = np.zeros(coord.shape)
    conv[:, 0] = -2 * (2 * Lz / Ly + Ly / Lz) * k * np.sin(y_) * np.cos(y_) * np.sin(z_) * np.cos(z_)
    conv[:, 1] = 2 * (2 * Lz / Lx + Lx / Lz) * k * np.sin(x_) * np.cos(x_) * np.sin(z_) * np.cos(z_)
    conv[:, 2] = 2 * (2 * Lx / Ly - 2 * Ly / Lx) * k * np.sin(y_) * np.cos(y_) * np.sin(x_) * np.cos(x_)
    return conv.flatten()


def diffusive(coord, alpha):
    x_, y_, z_ = _xyz(coord)
    k = (2 * pi) ** 3 * alpha
    difv = np.zeros(coord.shape)
    difv[:, 0] = k * np.sin(x_) * np.cos(y_) * np.cos(z_) * (
        2 * (Lz / (Lx * Lx * Ly) + Lz / (Ly * Ly * Ly) + Lz / (Lz * Lz * Ly))
        + Ly / (Lx * Lx * Lz) + Ly / (Ly * Ly * Lz) + Ly / (Lz * Lz * Lz))
    difv[:, 1] = -k * np.cos(x_) * np.sin(y_) * np.cos(z_) * (
        2 * (Lz / (Lx * Lx * Lx) + Lz / (Ly * Ly * Lx) + Lz / (Lz * Lz * Lx))
        + Lx / (Lx * Lx * Lz) + Lx / (Ly * Ly * Lz) + Lx / (Lz * Lz * Lz))
    difv[:, 2] = k * np.cos(x_) * np.cos(y_) * np.sin(z_) * (
        Lx / (Lx * Lx * Ly) + Lx / (Ly * Ly * Ly) + Lx / (Lz * Lz * Ly)
        - Ly / (Lx * Lx * Lx) - Ly / (Ly * Ly * Lx) - Ly / (Lz * Lz * Lx))
    return difv.flatten()
