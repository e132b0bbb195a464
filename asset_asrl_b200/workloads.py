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


def climb_tables():
    """Interpolation tables with the shapes of the minimum-time-to-climb example's (examples/MinimumTimeToClimb.py:10,
    examples/MinimumTimeToClimbTables.py:103-108: five cubic 1-D tables over altitude / Mach number, one cubic 2-D
    thrust table over Mach number x unevenly spaced altitude).  The data are SYNTHETIC: closed-form atmosphere, drag
    polar and thrust-lapse models sampled on grids of the example's sizes, not the example's tabulated measurements."""
    alts = np.concatenate([np.arange(0.0, 12000.0, 1000.0), np.arange(12000.0, 24001.0, 2000.0)])  # uneven
    temp = np.where(alts < 11000.0, 288.15 - 6.5e-3 * alts, 216.65) + 3.0 * np.exp(-((alts - 11000.0) / 1500.0) ** 2)
    rho = 1.225 * np.exp(-alts / 7400.0) * (1.0 + 0.02 * np.sin(alts / 3000.0))
    sos = np.sqrt(1.4 * 287.05 * temp)
    mach = np.array([0.0, 0.4, 0.8, 0.9, 1.0, 1.2, 1.4, 1.6, 1.8])  # uneven
    bump = np.exp(-((mach - 1.05) / 0.18) ** 2)
    clalpha = 3.44 + 0.9 * bump - 0.35 * mach
    cd0 = 0.013 + 0.03 * bump + 0.004 * mach
    eta = 0.54 + 0.25 * np.tanh(3.0 * (mach - 0.95)) + 0.1 * mach
    tmach = np.linspace(0.0, 1.8, 10)  # even
    talt = 304.8 * np.array([-0.5, 0.0, 5.0, 10.0, 15.0, 20.0, 25.0, 30.0, 40.0, 50.0, 70.0])  # uneven
    thrust = 4448.2 * 24.0 * (1.0 + 0.45 * tmach[None, :] + 0.3 * tmach[None, :] ** 2) * np.exp(-talt[:, None] / 8500.0) \
        * (1.0 - 0.25 * np.exp(-((tmach[None, :] - 1.0) / 0.3) ** 2))
    return {"rho": vf.InterpTable1D(alts, rho, kind="cubic"), "sos": vf.InterpTable1D(alts, sos, kind="cubic"),
            "Clalpha": vf.InterpTable1D(mach, clalpha, kind="cubic"), "eta": vf.InterpTable1D(mach, eta, kind="cubic"),
            "CD0": vf.InterpTable1D(mach, cd0, kind="cubic"), "Thrust": vf.InterpTable2D(tmach, talt, thrust, kind="cubic")}


# non-dimensional units of the example (examples/MinimumTimeToClimb.py:27-45)
_CL, _CT, _CM = 10000.0, 250.0, 19050.864
_CV, _CRHO, _CF = _CL / _CT, _CM / _CL ** 3, (_CL / _CT ** 2) * _CM
_MU, _RE, _SREF, _VEX = 3.986012e14 / (_CL ** 3 / _CT ** 2), 6378145.0 / _CL, 49.2386 / _CL ** 2, 1600.0 * 9.80665 / _CV


def minclimb_ode():
    """Supersonic-aircraft climb dynamics of examples/MinimumTimeToClimb.py:48-88 (states altitude, speed, flight-path
    angle, mass; control angle of attack) over the tables of ``climb_tables``."""
    T = climb_tables()
    args = oc.ODEArguments(4, 1)
    h, v, fpa, mass = args.XVar(0), args.XVar(1), args.XVar(2), args.XVar(3)
    alpha = args.UVar(0)
    rho = T["rho"](h * _CL) / _CRHO
    sos = T["sos"](h * _CL) / _CV
    mach = v / sos
    cd0, clalpha, eta = T["CD0"](mach), T["Clalpha"](mach), T["eta"](mach)
    thrust = T["Thrust"](mach, h * _CL) / _CF
    cd = cd0 + eta * clalpha * (alpha ** 2)
    cl = clalpha * alpha
    q = rho * (v ** 2) * 0.5
    drag, lift = q * _SREF * cd, q * _SREF * cl
    r = h + _RE
    hdot = v * vf.sin(fpa)
    vdot = (thrust * vf.cos(alpha) - drag) / mass - vf.sin(fpa) * _MU / (r ** 2)
    fpadot = (thrust * vf.sin(alpha) + lift) / (mass * v) + vf.cos(fpa) * (v / r - (v * (r ** 2)).inverse() * _MU)
    mdot = thrust * (-1.0 / _VEX)
    return oc.ODEBase(vf.stack([hdot, vdot, fpadot, mdot]), 4, 1)


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
    elif name == "minclimb":
        # the example's straight-line initial guess (examples/MinimumTimeToClimb.py:143-166), inside every table
        s_ = np.linspace(0.0, 1.0, S)
        h = (0.01 + (19994.88 - 0.01) * s_) / _CL
        v = (129.314 + (295.092 - 129.314) * s_) / _CV
        T = np.stack([h, v, 0.05 * np.sin(np.pi * s_), np.full(S, 1.0) - 0.05 * s_, (200.0 / _CT) * s_, 0.05 * np.cos(3.0 * s_)], axis=1)
        _perturb(T, [0, 1, 2, 3, 5], 2e-3, rng)
        T[:, 0] = np.maximum(T[:, 0], 1e-6)
    else:
        raise ValueError(f"unknown workload ODE '{name}'")
    return T


ODES = {"lowthrust": lowthrust_ode, "vanderpol": vanderpol_ode, "hypersens": hypersens_ode,
        "brachistochrone": brachistochrone_ode, "minclimb": minclimb_ode}


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
        elif name == "minclimb":  # the example's bounds and objective (examples/MinimumTimeToClimb.py:171-179)
            ph.addLUVarBound("Path", 0, 0.0, 21000.0 / _CL)
            ph.addLUVarBound("Path", 1, 5.0 / _CV, 600.0 / _CV)
            ph.addLUVarBound("Path", 2, -20.0 * np.pi / 180.0, 40.0 * np.pi / 180.0)
            ph.addLowerVarBound("Back", 3, 16500.0 / _CM)
            ph.addLUVarBound("Path", 5, -np.pi / 4.0, np.pi / 4.0)
            ph.addDeltaTimeObjective(1.0)
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
