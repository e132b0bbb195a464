"""Synthetic collocation meshes for measurement.

One phase of N equal segments with a smooth reference trajectory perturbed by 1e-2*U(-1,1) (seed 12345) so the defects
are non-zero, and multipliers 100*U(-1,1) (seed 67890, the reference's NLPTest choice, NonLinearProgram.cpp:708-719).
The ODEs are the ones the reference's examples use: two-body low thrust (examples/SimpleLowThrust.py:32-39),
VanDerPol (examples/VanDerPol.py:20-27), HyperSens (examples/HyperSens.py:23-26), Brachistochrone
(examples/Brachistochrone.py:23-31).
"""
import numpy as np

from . import OptimalControl as oc
from . import VectorFunctions as vf

CARDINALS = {"LGL3": 2, "LGL5": 3, "LGL7": 4, "Trapezoidal": 2}


def lowthrust_ode(acc=0.02):
    args = oc.ODEArguments(6, 3)
    r, v, u = args.head3(), args.segment3(3), args.tail3()
    return oc.ODEBase(vf.stack([v, r.normalized_power3() * (-1.0) + u * acc]), 6, 3)


def vanderpol_ode(mu=1.0):
    args = oc.ODEArguments(2, 1)
    x0, x1, u = args.XVar(0), args.XVar(1), args.UVar(0)
    return oc.ODEBase(vf.stack([x1, (1.0 - x0 * x0) * x1 * mu - x0 + u]), 2, 1)


def hypersens_ode():
    args = oc.ODEArguments(1, 1)
    x, u = args.XVar(0), args.UVar(0)
    return oc.ODEBase(vf.stack([u - x]), 1, 1)


def brachistochrone_ode(g=9.81):
    args = oc.ODEArguments(3, 1)
    v, theta = args.XVar(2), args.UVar(0)
    return oc.ODEBase(vf.stack([vf.sin(theta) * v, vf.cos(theta) * v * (-1.0), vf.cos(theta) * g]), 3, 1)


def _perturb(T, cols, scale, rng):
    T[:, cols] += scale * rng.uniform(-1.0, 1.0, (T.shape[0], len(cols)))


def reference_trajectory(name, S, rng):
    """[S, XV+1+UV] node states on t in [0, T] (time strictly increasing, unperturbed)."""
    if name == "lowthrust":
        th = np.linspace(0.0, 6.4 * np.pi, S)
        T = np.zeros((S, 10))
        T[:, 0], T[:, 1] = np.cos(th), np.sin(th)
        T[:, 3], T[:, 4] = -np.sin(th), np.cos(th)
        T[:, 6] = th
        T[:, 7], T[:, 8] = -0.8 * np.sin(th), 0.8 * np.cos(th)  # u = 0.8 * v_hat
        _perturb(T, list(range(6)) + [7, 8, 9], 1e-2, rng)
    elif name == "vanderpol":
        t = np.linspace(0.0, 10.0, S)
        T = np.stack([np.cos(t), -np.sin(t), t, 0.1 * np.sin(2 * t)], axis=1)
        _perturb(T, [0, 1, 3], 1e-2, rng)
    elif name == "hypersens":
        t = np.linspace(0.0, 1.0e4, S)
        T = np.stack([1.0 + 0.5 * np.sin(t * 1e-3), t, 0.5 * np.cos(t * 1e-3)], axis=1)
        _perturb(T, [0, 2], 1e-2, rng)
    elif name == "brachistochrone":
        t = np.linspace(0.0, 1.0, S)
        T = np.stack([t, 1.0 - t, 1.0 + 3.0 * t, t, 0.3 + t], axis=1)
        _perturb(T, [0, 1, 2, 4], 1e-2, rng)
    else:
        raise ValueError(f"unknown workload ODE '{name}'")
    return T


ODES = {"lowthrust": lowthrust_ode, "vanderpol": vanderpol_ode, "hypersens": hypersens_ode,
        "brachistochrone": brachistochrone_ode}


def make_phase(name, mode, nseg, control_mode=None, extras=True, seed=12345):
    """A single phase with the functions the reference's example scripts add (bounds, norm bound, integral objective,
    boundary values) when ``extras`` is true; defect-only otherwise."""
    rng = np.random.default_rng(seed)
    ode = ODES[name]()
    S = (CARDINALS[mode] - 1) * nseg + 1
    T = reference_trajectory(name, S, rng)
    ph = ode.phase(mode)
    if control_mode is not None:
        ph.setControlMode(control_mode)
    ph.setTraj(T, nseg)
    if extras:
        xv, uv = ode.XV, ode.UV
        ph.addBoundaryValue("Front", list(range(xv + 1)), T[0, :xv + 1])
        ph.addBoundaryValue("Back", list(range(xv)), T[-1, :xv])
        ucols = list(range(xv + 1, xv + 1 + uv))
        if name == "lowthrust":
            ph.addLUNormBound("Path", ucols, 0.001, 1.0)
            ph.addIntegralObjective(vf.Arguments(3).squared_norm() * 0.5, ucols)
        else:
            ph.addLUVarBound("Path", ucols[0], -10.0, 10.0)
            ph.addIntegralObjective(vf.Arguments(1).squared_norm() * 0.5, ucols[:1])
    return ph


def multipliers(nlp, seed=67890):
    d = nlp.dims()
    rng = np.random.default_rng(seed)
    return 100.0 * rng.uniform(-1, 1, d["neq"]), 100.0 * rng.uniform(-1, 1, d["niq"])


def algorithmic_bytes(nlp):
    """every input read once, every output written once, indices excluded."""
    d = nlp.dims()
    nnz_user = d["nnz"] - _solver_only_slots(nlp)
    return 8 * (d["nvars"] + d["neq"] + d["niq"] + nnz_user + d["neq"] + d["niq"] + 2 * d["nvars"])


def _solver_only_slots(nlp):
    d = nlp.dims()
    loc = nlp.kkt_locations()
    user = np.zeros(d["nnz"], dtype=bool)
    user[loc[:d["n_user"]]] = True
    return int(d["nnz"] - user.sum())
