"""Golovin analytic solution for the additive kernel, restating soln_additive_exp
(partmc/src/coag_kernel_additive.F90:101-167).  TEST INFRASTRUCTURE ONLY.

The Fortran evaluates I_1 with bessi1/calci1 (:172-381) and switches to the
asymptotic form for x >= 500; here scipy's exponentially scaled ive(1, x) gives
exp(-x) I_1(x) directly, which covers both branches without overflow."""
import numpy as np
from scipy.special import ive

BETA_1 = 1000.0  # coag_kernel_additive.F90:23


def soln_additive_exp(centers, time, num_conc, radius_at_mean_vol):
    """n(ln r) per bin centre (radius grid), like aero_binned%num_conc."""
    mean_vol = 4.0 * np.pi * radius_at_mean_vol ** 3 / 3.0
    v = 4.0 * np.pi * centers ** 3 / 3.0
    if time == 0.0:
        return np.pi / 2.0 * (2.0 * centers) ** 3 * num_conc / mean_vol * np.exp(-(v / mean_vol))
    tau = num_conc * mean_vol * BETA_1 * time
    T = 1.0 - np.exp(-tau)
    rat_v = v / mean_vol
    x = 2.0 * rat_v * np.sqrt(T)
    nn = num_conc / v * (1.0 - T) / np.sqrt(T) * np.exp(-((1.0 + T) * rat_v) + x) * ive(1, x)
    return np.pi / 2.0 * (2.0 * centers) ** 3 * nn


def compare(particle_vol, magnitude, time, num_conc, radius_at_mean_vol):
    """Relative 2-norm errors (number, volume) on the test/additive grid:
    160 log bins over diameters 1e-8..1e-3 m (run_part.spec / run_exact.spec)."""
    edges = np.logspace(-8, -3, 161) / 2.0
    centers = np.sqrt(edges[1:] * edges[:-1])
    r = (3.0 * particle_vol / 4.0 / np.pi) ** (1.0 / 3.0)
    h, _ = np.histogram(r, edges)
    dln = np.log(edges[1] / edges[0])
    num = h * magnitude / dln
    ex = soln_additive_exp(centers, time, num_conc, radius_at_mean_vol)
    v = 4.0 * np.pi * centers ** 3 / 3.0
    err_n = np.linalg.norm(num - ex) / np.linalg.norm(ex)
    err_v = np.linalg.norm(num * v - ex * v) / np.linalg.norm(ex * v)
    return err_n, err_v
