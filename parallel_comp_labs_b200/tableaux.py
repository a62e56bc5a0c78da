"""Butcher tableaux for nbody_step_rk / nbody_step_embedded ("adaptive/other
tableaux (RK3/RK45)"), and a step-size controller over the embedded pairs.

Not in the reference: its only scheme is the explicit midpoint rule (lab_4
.../schemes/ode_RK2_MPI.hpp:72-92, lab_5 N_body.cu:299-316), which is `midpoint` below and the
hard-wired NBODY_RK2 of nbody_step.  Every scheme runs over the same force kernels and exchange;
only the fused stage kernel differs (all stage derivatives resident).

A tableau is (a, b, b_err, order): a (s, s) strictly lower triangular, b (s,) the weights of the
propagated solution, b_err (s,) the weights of the embedded companion or None, order = order of
the propagated solution (the controller's exponent uses min(order, companion order) + 1)."""
from __future__ import annotations

from fractions import Fraction as F

import numpy as np


def _t(rows, b, b_err=None, order=1, err_order=None):
    s = len(b)
    a = np.zeros((s, s))
    for i, r in enumerate(rows):
        for j, v in enumerate(r):
            a[i + 1, j] = float(v)
    return {"a": a, "b": np.array([float(x) for x in b]),
            "b_err": None if b_err is None else np.array([float(x) for x in b_err]),
            "order": order, "err_order": err_order}


TABLEAUX = {
    "euler": _t([], [1], order=1),
    "midpoint": _t([[F(1, 2)]], [0, 1], order=2),                       # the reference's scheme
    "heun": _t([[1]], [F(1, 2), F(1, 2)], order=2),
    "ralston": _t([[F(2, 3)]], [F(1, 4), F(3, 4)], order=2),
    "kutta3": _t([[F(1, 2)], [-1, 2]], [F(1, 6), F(2, 3), F(1, 6)], order=3),
    "rk4": _t([[F(1, 2)], [0, F(1, 2)], [0, 0, 1]], [F(1, 6), F(1, 3), F(1, 3), F(1, 6)], order=4),
    "rk38": _t([[F(1, 3)], [F(-1, 3), 1], [1, -1, 1]], [F(1, 8), F(3, 8), F(3, 8), F(1, 8)], order=4),
    # Bogacki-Shampine 3(2)
    "bs23": _t([[F(1, 2)], [0, F(3, 4)], [F(2, 9), F(1, 3), F(4, 9)]],
               [F(2, 9), F(1, 3), F(4, 9), 0],
               [F(7, 24), F(1, 4), F(1, 3), F(1, 8)], order=3, err_order=2),
    # Runge-Kutta-Fehlberg 4(5): the 4th-order solution is propagated
    "rkf45": _t([[F(1, 4)], [F(3, 32), F(9, 32)], [F(1932, 2197), F(-7200, 2197), F(7296, 2197)],
                 [F(439, 216), -8, F(3680, 513), F(-845, 4104)],
                 [F(-8, 27), 2, F(-3544, 2565), F(1859, 4104), F(-11, 40)]],
                [F(25, 216), 0, F(1408, 2565), F(2197, 4104), F(-1, 5), 0],
                [F(16, 135), 0, F(6656, 12825), F(28561, 56430), F(-9, 50), F(2, 55)],
                order=4, err_order=5),
    # Cash-Karp 5(4)
    "cashkarp": _t([[F(1, 5)], [F(3, 40), F(9, 40)], [F(3, 10), F(-9, 10), F(6, 5)],
                    [F(-11, 54), F(5, 2), F(-70, 27), F(35, 27)],
                    [F(1631, 55296), F(175, 512), F(575, 13824), F(44275, 110592), F(253, 4096)]],
                   [F(37, 378), 0, F(250, 621), F(125, 594), 0, F(512, 1771)],
                   [F(2825, 27648), 0, F(18575, 48384), F(13525, 55296), F(277, 14336), F(1, 4)],
                   order=5, err_order=4),
    # Dormand-Prince 5(4), 7 stages (the 7th is f(y_new): used for the error estimate only)
    "dopri5": _t([[F(1, 5)], [F(3, 40), F(9, 40)], [F(44, 45), F(-56, 15), F(32, 9)],
                  [F(19372, 6561), F(-25360, 2187), F(64448, 6561), F(-212, 729)],
                  [F(9017, 3168), F(-355, 33), F(46732, 5247), F(49, 176), F(-5103, 18656)],
                  [F(35, 384), 0, F(500, 1113), F(125, 192), F(-2187, 6784), F(11, 84)]],
                 [F(35, 384), 0, F(500, 1113), F(125, 192), F(-2187, 6784), F(11, 84), 0],
                 [F(5179, 57600), 0, F(7571, 16695), F(393, 640), F(-92097, 339200), F(187, 2100),
                  F(1, 40)], order=5, err_order=4),
}


def is_fsal(name: str) -> bool:
    """First same as last: the last stage is f(y_new), i.e. the next step's first stage.  The
    library detects this from the coefficients and skips that force evaluation."""
    t = TABLEAUX[name]
    a, b = t["a"], t["b"]
    return len(b) > 1 and b[-1] == 0.0 and bool(np.all(a[-1, :-1] == b[:-1]))


def check_order_conditions(name: str) -> float:
    """Largest violation of the order conditions up to order 3 plus row-sum consistency (used by
    the CPU tests to guard the table against typos)."""
    t = TABLEAUX[name]
    a, b = t["a"], t["b"]
    c = a.sum(axis=1)
    worst = abs(b.sum() - 1)
    if t["order"] >= 2:
        worst = max(worst, abs(b @ c - 0.5))
    if t["order"] >= 3:
        worst = max(worst, abs(b @ c ** 2 - 1 / 3), abs(b @ (a @ c) - 1 / 6))
    if t["order"] >= 4:
        worst = max(worst, abs(b @ c ** 3 - 0.25), abs((b * c) @ (a @ c) - 0.125),
                    abs(b @ (a @ c ** 2) - 1 / 12), abs(b @ (a @ (a @ c)) - 1 / 24))
    if t["b_err"] is not None:
        be = t["b_err"]
        worst = max(worst, abs(be.sum() - 1), abs(be @ c - 0.5))
    return float(worst)


def integrate_adaptive(ctx, name: str, t_start: float, t_final: float, tau0: float,
                       rtol: float = 1e-9, atol: float = 0.0, max_steps: int = 100000,
                       safety: float = 0.9, on_accept=None) -> dict:
    """Advance `ctx` from t_start to t_final with an embedded pair and the standard controller
    tau_new = tau * min(5, max(0.2, safety * err^(-1/(q+1)))), q = min(order, err_order);
    err = max(|err_r| / (atol + rtol |r|), |err_v| / (atol + rtol |v|)) in max-norms.
    Rejected steps are undone on the device (nbody_step_undo)."""
    t = TABLEAUX[name]
    if t["b_err"] is None:
        raise ValueError(f"{name} has no embedded companion")
    q = min(t["order"], t["err_order"])
    time, tau = float(t_start), float(tau0)
    accepted = rejected = 0
    taus = []
    while time < t_final and accepted + rejected < max_steps:
        tau = min(tau, t_final - time)
        n = ctx.step_embedded(t["a"], t["b"], t["b_err"], tau)
        err = max(n[0] / (atol + rtol * n[2]), n[1] / (atol + rtol * n[3]))
        if not np.isfinite(err):
            raise FloatingPointError("non-finite error estimate")
        if err <= 1.0:
            time += tau
            accepted += 1
            taus.append(tau)
            if on_accept:
                on_accept(time, tau, err)
        else:
            ctx.step_undo()
            rejected += 1
        tau *= min(5.0, max(0.2, safety * (err if err > 0 else 1e-16) ** (-1.0 / (q + 1))))
    return {"t": time, "accepted": accepted, "rejected": rejected, "taus": taus}
