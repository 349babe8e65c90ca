"""Analytic Taylor-Green fields used as initial / boundary / exact data.

Same formulas as the reference's src/functions/taylor_green*.py (they are the
analytic solution, not an algorithm), written once for the three variants.
All take ``coord`` [n, dim] and the decay factor ``alpha`` and return the
flattened node-major array the reference returns.
"""
from math import exp, pi

import numpy as np


def decay(nu, t, ndirs, uref=1.0, lengths=(1.0, 1.0, 1.0)):
    s = sum(1.0 / lengths[d] ** 2 for d in range(ndirs))
    return uref * exp(-4 * (pi ** 2) * nu * t * s)


def angles(coord, lengths):
    return [2 * pi * coord[:, d] / lengths[d] for d in range(coord.shape[1])]
