"""Optimal-control front end with the reference's Python surface for the transcription path.

Mirrors ``oc.ODEArguments`` (/root/reference/src/OptimalControl/ODEArguments.h:22-39), ``oc.ODEBase``
(/root/reference/asset_asrl/OptimalControl/ODEBaseClass.py:7-254), ``ode.phase(...)`` (ODE.h:138-166) and the phase
methods bound in ODEPhaseBaseBind.cpp.  The transcription itself (index maps, function list) runs in C++
(csrc/phase.cpp); this file only forwards.  PSIOPT (the interior-point solver and its sparse LDLT) is out of scope:
``phase.optimize()`` raises, while ``phase.transcribe()`` returns the NonLinearProgram whose eval* methods run on
the GPU.
"""
import ctypes as C

import numpy as np

from . import _lib
from . import VectorFunctions as vf
from ._lib import lib
from .Solvers import NonLinearProgram
from .VectorFunctions import GRAPH, Arguments, ScalarFunction, VectorFunction


class _Enum:
    @classmethod
    def parse(cls, v):
        if isinstance(v, str):
            for k, val in cls.__dict__.items():
                if not k.startswith("_") and k.lower() == v.lower():
                    return val
            alias = getattr(cls, "_alias", {})
            if v in alias:
                return alias[v]
            raise ValueError(f"Unrecognized {cls.__name__} string: {v}")
        return int(v)


class PhaseRegionFlags(_Enum):
    NotSet, Front, Back, FrontandBack, BackandFront, Path, InnerPath, NodalPath, DefectPath, PairWisePath, \
        DefectPairWisePath, FrontNodalBackPath, Params, ODEParams, StaticParams, Accumulate, BlockDefectPath = range(17)
    _alias = {"First": 1, "Last": 2, "FirstandLast": 3, "LastandFirst": 4}


class TranscriptionModes(_Enum):
    LGL3, LGL5, LGL7, Trapezoidal, CentralShooting = range(5)


class ControlModes(_Enum):
    HighestOrderSpline, FirstOrderSpline, NoSpline, BlockConstant = range(4)


class IntegralModes(_Enum):
    BaseIntegral, SimpsonIntegral, TrapIntegral = range(3)


class ODEArguments(Arguments):
    """Arguments laid out as [X, t, U, P] (ODEArguments.h:22-39)."""

    def __init__(self, Xv, Uv=0, Pv=0):
        self._xv, self._uv, self._pv = int(Xv), int(Uv or 0), int(Pv or 0)
        super().__init__(self._xv + 1 + self._uv + self._pv)

    def XVars(self):
        return self._xv

    def UVars(self):
        return self._uv

    def PVars(self):
        return self._pv

    def XVec(self):
        return self.head(self._xv)

    def XtVec(self):
        return self.head(self._xv + 1)

    def XVar(self, i):
        return self.coeff(i)

    def TVar(self):
        return self.coeff(self._xv)

    def UVec(self):
        return self.segment(self._xv + 1, self._uv)

    def UVar(self, i):
        return self.coeff(self._xv + 1 + i)

    def PVec(self):
        return self.tail(self._pv)

    def PVar(self, i):
        return self.coeff(self._xv + 1 + self._uv + i)


def _idx(v):
    if v is None:
        return []
    if isinstance(v, (int, np.integer)):
        return [int(v)]
    return [int(k) for k in v]


def _rawnode(kind, ir, orows, ch=(), d=(), ip=()):
    """Node with the type a C++-built helper has in the reference: not wrapped in a GenericFunction."""
    n = GRAPH.node(kind, ir, orows, [c.node for c in ch], d, ip, -1, 1)
    return (ScalarFunction if orows == 1 else VectorFunction)(n, ir, orows)


def _lu_bound(lb, ub, lbs, ubs):
    """StackedOutputs{(lb - x) * lbscale, (x - ub) * ubscale} (ODEPhaseBase.cpp:110-130)."""
    x = _rawnode(vf.VK_ARGS, 1, 1)
    low = _rawnode(vf.VK_SCALED, 1, 1, [_rawnode(vf.VK_VECMINUS, 1, 1, [x], d=[lb])], d=[lbs])
    up = _rawnode(vf.VK_SCALED, 1, 1, [_rawnode(vf.VK_ADDVEC, 1, 1, [x], d=[-ub])], d=[ubs])
    return _rawnode(vf.VK_STACK, 1, 2, [low, up])


class ODEBase:
    """oc.ODEBase(odefunc, Xvars, Uvars, Pvars) (ODEBaseClass.py:7-41)."""

    def __init__(self, odefunc, Xvars, Uvars=None, Pvars=None, Vgroups=None):
        self.XV, self.UV, self.PV = int(Xvars), int(Uvars or 0), int(Pvars or 0)
        if not isinstance(odefunc, VectorFunction):
            raise TypeError("odefunc must be a VectorFunction")
        if odefunc.ir != self.XV + 1 + self.UV + self.PV or odefunc.orows != self.XV:
            raise ValueError("ODE function sizes do not match Xvars/Uvars/Pvars")
        self.odefunc = odefunc
        self.ode = self

    def XVars(self):
        return self.XV

    def UVars(self):
        return self.UV

    def PVars(self):
        return self.PV

    def TVar(self):
        return self.XV

    def XtUVars(self):
        return self.XV + 1 + self.UV

    def IRows(self):
        return self.odefunc.ir

    def ORows(self):
        return self.XV

    def vf(self):
        return self.odefunc

    def phase(self, mode, Traj=None, numDefects=None, LerpIG=True):
        ph = ODEPhase(self, mode)
        if Traj is not None:
            ph.setTraj(Traj, numDefects)
        return ph

    def integrator(self, *a, **k):
        raise NotImplementedError("integrators are outside the KKT evaluation path (SURVEY.md section 2, row 15)")


class ODEPhase:
    """The phase object returned by ``ode.phase(mode, Traj, n)`` (ODEPhaseBase / ODEPhase<DODE>)."""

    def __init__(self, ode, mode):
        self.ode = ode
        self.tmode = TranscriptionModes.parse(mode)
        if self.tmode == TranscriptionModes.CentralShooting:
            raise NotImplementedError("CentralShooting needs the sequential integrators, which are outside the KKT path")
        self.h = None
        self._nseg = 0
        self._cmode = ControlModes.FirstOrderSpline
        self._imode = IntegralModes.BaseIntegral
        self._funcs = []  # (role, fun, region, xtuv, opv, spv, ext) in insertion order
        self._spv = 0
        self._static = np.zeros(0)
        self._bins = (np.array([0.0, 1.0]), None)
        self._traj = None
        self.Threads = 1
        self.nlp = None

    # ------------------------------------------------------------------ mesh / trajectory
    def _rebuild(self):
        if self.h:
            lib.ab_phase_destroy(self.h)
        self.h = lib.ab_phase_create(GRAPH.h, self.ode.odefunc.node, self.ode.XV, self.ode.UV, self.ode.PV, self.tmode, self._nseg)
        if not self.h:
            raise ValueError(_lib.last_error())
        _lib.check(lib.ab_phase_set_control_mode(self.h, self._cmode))
        _lib.check(lib.ab_phase_set_integral_mode(self.h, self._imode))
        _lib.check(lib.ab_phase_set_static_params(self.h, self._spv))
        sp, defs = self._bins
        if defs is not None:
            a, ap = _lib.darr(sp)
            b, bp = _lib.iarr(defs)
            _lib.check(lib.ab_phase_set_mesh(self.h, len(b), ap, bp))
        for (role, fun, reg, x, o, s, e) in self._funcs:
            xa, xp = _lib.iarr(x)
            oa, op = _lib.iarr(o)
            sa, sp_ = _lib.iarr(s)
            ea, ep = _lib.iarr(e)
            _lib.check(lib.ab_phase_add_function(self.h, role, fun.node, reg, xp, len(xa), op, len(oa), sp_, len(sa), ep, len(ea)))

    def __del__(self):
        try:
            if self.h:
                lib.ab_phase_destroy(self.h)
        except Exception:
            pass

    def dims(self):
        out = (C.c_int * 8)()
        _lib.check(lib.ab_phase_dims(self.h, out))
        return dict(zip(["numPhaseVars", "numStates", "numDefects", "cardinalStates", "XtUVars", "PVars", "StaticPVars",
                         "blocked"], list(out)))

    def state_spacing(self):
        n = self.dims()["numStates"]
        out = np.zeros(n)
        _lib.check(lib.ab_phase_state_spacing(self.h, _lib.dptr(out)))
        return out

    def setTraj(self, Traj, numDefects=None, DefBinSpacing=None, DefsPerBin=None, LerpIG=True):
        """Resample ``Traj`` (list of [X,t,U,P] states) onto the collocation mesh by linear interpolation in time.

        The reference's default is its LGLInterpTable; only the ``LerpIG`` behaviour is provided because the
        interpolation tables are outside the KKT path.  A trajectory that already has exactly numStates nodes is
        taken as is."""
        T = np.asarray(Traj, dtype=np.float64)
        if T.ndim != 2 or T.shape[1] != self.ode.XtUVars() + self.ode.PV:
            raise ValueError("Trajectory states must have XVars+1+UVars+PVars entries")
        if not np.all(np.isfinite(T)):
            raise ValueError("NaN or Inf detected in trajectory")
        if DefBinSpacing is not None:
            self._bins = (np.asarray(DefBinSpacing, dtype=np.float64), np.asarray(DefsPerBin, dtype=np.int32))
            self._nseg = int(np.sum(DefsPerBin))
        else:
            if numDefects is None:
                raise ValueError("number of segments required")
            self._nseg = int(numDefects)
            self._bins = (np.array([0.0, 1.0]), None)
        self._rebuild()
        d = self.dims()
        S = d["numStates"]
        if T.shape[0] == S:
            self._traj = T.copy()
        else:
            tv = self.ode.XV
            t0, tf = T[0, tv], T[-1, tv]
            ts = t0 + (tf - t0) * self.state_spacing()
            self._traj = np.stack([np.interp(ts, T[:, tv], T[:, k]) for k in range(T.shape[1])], axis=1)
            self._traj[:, tv] = ts
        self.nlp = None

    def returnTraj(self):
        return [r.copy() for r in self._traj]

    def setStaticParams(self, params):
        self._static = np.asarray(params, dtype=np.float64).ravel()
        self._spv = self._static.size
        if self.h:
            self._rebuild()
        self.nlp = None

    def setControlMode(self, m):
        self._cmode = ControlModes.parse(m)
        if self.h:
            self._rebuild()
        self.nlp = None

    def setIntegralMode(self, m):
        self._imode = IntegralModes.parse(m)
        if self.h:
            self._rebuild()
        self.nlp = None

    def setThreads(self, *a):
        self.Threads = int(a[0]) if a else 1

    # ------------------------------------------------------------------ function registration
    def _add(self, role, fun, reg, x, o=(), s=(), e=()):
        reg = PhaseRegionFlags.parse(reg)
        x, o, s, e = _idx(x), _idx(o), _idx(s), _idx(e)
        if reg == PhaseRegionFlags.ODEParams and not o:
            x, o = [], x
        elif reg == PhaseRegionFlags.StaticParams and not s:
            x, s = [], x
        if not isinstance(fun, VectorFunction):
            raise TypeError("expected a VectorFunction")
        self._funcs.append((role, fun, reg, x, o, s, e))
        if self.h:
            xa, xp = _lib.iarr(x)
            oa, op = _lib.iarr(o)
            sa, sp_ = _lib.iarr(s)
            ea, ep = _lib.iarr(e)
            _lib.check(lib.ab_phase_add_function(self.h, role, fun.node, reg, xp, len(xa), op, len(oa), sp_, len(sa), ep, len(ea)))
        self.nlp = None
        return len(self._funcs) - 1

    def addEqualCon(self, reg, fun, XtUvars, OPvars=(), SPvars=()):
        return self._add(1, fun, reg, XtUvars, OPvars, SPvars)

    def addInequalCon(self, reg, fun, XtUvars, OPvars=(), SPvars=()):
        return self._add(2, fun, reg, XtUvars, OPvars, SPvars)

    def addStateObjective(self, reg, fun, XtUvars, OPvars=(), SPvars=()):
        return self._add(0, fun, reg, XtUvars, OPvars, SPvars)

    def addIntegralObjective(self, fun, XtUvars, OPvars=(), SPvars=()):
        return self._add(3, fun, PhaseRegionFlags.Path, XtUvars, OPvars, SPvars)

    def addIntegralParamFunction(self, fun, XtUvars, accum_param, OPvars=(), SPvars=()):
        return self._add(4, fun, PhaseRegionFlags.Path, XtUvars, OPvars, SPvars, _idx(accum_param))

    def addBoundaryValue(self, reg, args, value):
        value = np.atleast_1d(np.asarray(value, dtype=np.float64))
        args = _idx(args)
        x = _rawnode(vf.VK_ARGS, value.size, value.size)
        return self.addEqualCon(reg, _rawnode(vf.VK_ADDVEC, value.size, value.size, [x], d=-value), args)

    def addDeltaVarEqualCon(self, var, value, scale=1.0):
        a = _rawnode(vf.VK_ARGS, 2, 2)
        d = _rawnode(vf.VK_DIFF, 2, 1, [_rawnode(vf.VK_SEGMENT, 2, 1, [a], ip=[1, 1]), _rawnode(vf.VK_SEGMENT, 2, 1, [a], ip=[0, 1])])
        f = _rawnode(vf.VK_SCALED, 2, 1, [_rawnode(vf.VK_ADDVEC, 2, 1, [d], d=[-float(value)])], d=[float(scale)])
        return self.addEqualCon(PhaseRegionFlags.FrontandBack, f, var)

    def addDeltaTimeEqualCon(self, value, scale=1.0):
        return self.addDeltaVarEqualCon(self.ode.TVar(), value, scale)

    def addLUVarBound(self, reg, var, lowerbound, upperbound, lbscale=1.0, ubscale=None):
        if lowerbound > upperbound:
            raise ValueError("Lower-bound greater than Upper-bound")
        ubscale = lbscale if ubscale is None else ubscale
        return self.addInequalCon(reg, _lu_bound(lowerbound, upperbound, lbscale, ubscale), var)

    def addLUVarBounds(self, reg, vars_, lowerbound, upperbound, scale=1.0):
        return [self.addLUVarBound(reg, v, lowerbound, upperbound, scale) for v in _idx(vars_)]

    def addLowerVarBound(self, reg, var, lowerbound, lbscale=1.0):
        x = _rawnode(vf.VK_ARGS, 1, 1)
        return self.addInequalCon(reg, _rawnode(vf.VK_SCALED, 1, 1, [_rawnode(vf.VK_VECMINUS, 1, 1, [x], d=[lowerbound])], d=[lbscale]), var)

    def addUpperVarBound(self, reg, var, upperbound, ubscale=1.0):
        x = _rawnode(vf.VK_ARGS, 1, 1)
        return self.addInequalCon(reg, _rawnode(vf.VK_SCALED, 1, 1, [_rawnode(vf.VK_ADDVEC, 1, 1, [x], d=[-upperbound])], d=[ubscale]), var)

    def addLUFuncBound(self, reg, func, XtUvars, lowerbound, upperbound, lbscale=1.0, ubscale=None, OPvars=(), SPvars=()):
        if lowerbound > upperbound:
            raise ValueError("Lower-bound greater than Upper-bound")
        ubscale = lbscale if ubscale is None else ubscale
        lub = _lu_bound(lowerbound, upperbound, lbscale, ubscale)
        n = GRAPH.node(vf.VK_NESTED, func.ir, 2, [lub.node, func.node], (), (), -1, 1)
        return self.addInequalCon(reg, VectorFunction(n, func.ir, 2), XtUvars, OPvars, SPvars)

    def _norm_bound(self, reg, XtUvars, lb, ub, lbs, ubs, p):
        XtUvars = _idx(XtUvars)
        size = len(XtUvars) * (2 if PhaseRegionFlags.parse(reg) in (3, 4) else 1)
        args = _rawnode(vf.VK_ARGS, size, size)
        nrm = _rawnode(vf.VK_NORM, size, 1, [args], ip=[p])
        return self.addLUFuncBound(reg, nrm, XtUvars, lb, ub, lbs, ubs)

    def addLUNormBound(self, reg, XtUvars, lowerbound, upperbound, lbscale=1.0, ubscale=None):
        return self._norm_bound(reg, XtUvars, lowerbound, upperbound, lbscale, ubscale, 1)

    def addLUSquaredNormBound(self, reg, XtUvars, lowerbound, upperbound, lbscale=1.0, ubscale=None):
        return self._norm_bound(reg, XtUvars, lowerbound, upperbound, lbscale, ubscale, 2)

    def addValueObjective(self, reg, var, scale):
        x = _rawnode(vf.VK_ARGS, 1, 1)
        return self.addStateObjective(reg, _rawnode(vf.VK_SCALED, 1, 1, [x], d=[float(scale)]), var)

    def addDeltaVarObjective(self, var, scale):
        a = _rawnode(vf.VK_ARGS, 2, 2)
        d = _rawnode(vf.VK_DIFF, 2, 1, [_rawnode(vf.VK_SEGMENT, 2, 1, [a], ip=[1, 1]), _rawnode(vf.VK_SEGMENT, 2, 1, [a], ip=[0, 1])])
        return self.addStateObjective(PhaseRegionFlags.FrontandBack, _rawnode(vf.VK_SCALED, 2, 1, [d], d=[float(scale)]), var)

    def addDeltaTimeObjective(self, scale):
        return self.addDeltaVarObjective(self.ode.TVar(), scale)

    # ------------------------------------------------------------------ transcription / evaluation
    def makeSolverInput(self):
        d = self.dims()
        X = np.zeros(d["numPhaseVars"])
        traj = np.ascontiguousarray(self._traj)
        sp = np.ascontiguousarray(self._static) if self._spv else np.zeros(1)
        _lib.check(lib.ab_phase_make_input(self.h, _lib.dptr(traj), _lib.dptr(sp), _lib.dptr(X)))
        return X

    def transcribe(self, showstats=False, showfuns=False):
        """ODEPhaseBase::transcribe (ODEPhaseBase.cpp:1383-1418): builds the NonLinearProgram for this phase."""
        if not self.h:
            raise RuntimeError("setTraj must be called before transcribe")
        nlp = NonLinearProgram()
        out = (C.c_int * 3)()
        _lib.check(lib.ab_phase_transcribe(self.h, nlp.h, 0, 0, 0, out))
        nlp.set_dims(out[0], out[1], out[2])
        nlp.analyze()
        self.nlp = nlp
        return nlp

    def transcribe_window(self, rank, world, halo=8):
        """This rank's share of the phase, analysed locally: the transcription of the whole phase restricted to the
        applications within ``halo`` segments of mesh segments [N*rank/world, N*(rank+1)/world) before the sparsity
        analysis (replaces the whole-problem pass of NonLinearProgram.cpp:267-344; cost and memory follow the share,
        not the phase).  Returns a NonLinearProgram for ``to_device(device, rank=rank, world=world)``; indices and
        dimensions are those of the whole problem, the CSR pattern is the window's own."""
        if not self.h:
            raise RuntimeError("setTraj must be called before transcribe")
        d = self.dims()
        if d["blocked"]:
            raise NotImplementedError("windowed transcription of block-constant controls")
        nlp = NonLinearProgram()
        out = (C.c_int * 3)()
        _lib.check(lib.ab_phase_transcribe(self.h, nlp.h, 0, 0, 0, out))
        nlp.set_dims(out[0], out[1], out[2])
        n = d["numDefects"]
        seg_vars = d["XtUVars"] * (d["cardinalStates"] - 1)
        mesh_vars = d["numStates"] * d["XtUVars"]
        nlp.window(mesh_vars, seg_vars, n, n * rank // world, n * (rank + 1) // world, halo)
        nlp.analyze()
        return nlp

    def optimize(self, *a, **k):
        raise NotImplementedError("PSIOPT's interior-point loop and sparse LDLT are out of scope for this build "
                                  "(SURVEY.md section 2 row 13); use phase.transcribe() and nlp.evalKKT(...)")

    solve = optimize
    solve_optimize = optimize

    def test_threads(self, i, j, n):
        """The reference's own KKT micro-benchmark protocol (ODEPhaseBase.cpp:1420-1441, NonLinearProgram.cpp:686-820):
        n repetitions of evalKKT with random multipliers; here the GPU evaluation is timed and compared run-to-run."""
        nlp = self.nlp or self.transcribe()
        return nlp.NLPTest(self.makeSolverInput(), n)


class LinkFlags(_Enum):
    """OptimalControlFlags.h LinkFlags: which phase regions a two-phase link function reads."""
    BackToFront, BackToBack, FrontToBack, FrontToFront, ParamsToParams, PathToPath, LinkParams, ReadRegions = range(8)
    _regions = {0: (2, 1), 1: (2, 2), 2: (1, 2), 3: (1, 1), 4: (12, 12), 5: (5, 5), 6: ()}


class _LinkFunction:
    """LinkFunction<FuncType> (/root/reference/src/OptimalControl/LinkFunction.h:11-300): a function applied once per
    entry of ``ptl`` to the concatenation of one region of each listed phase and some link parameters."""

    def __init__(self, fun, flag, regs, ptl, xtv, opv, spv, lv):
        self.fun, self.flag, self.regs = fun, flag, [PhaseRegionFlags.parse(r) for r in regs]
        self.ptl = [[int(p) for p in col] for col in ptl]
        self.xtv, self.opv, self.spv = [_idx(v) for v in xtv], [_idx(v) for v in opv], [_idx(v) for v in spv]
        self.lv = [_idx(v) for v in lv]
        ncols = max(len(self.ptl), len(self.lv))
        if not self.lv:
            self.lv = [[] for _ in range(ncols)]
        if not self.ptl:
            self.ptl = [[] for _ in range(ncols)]
        for col in self.ptl:
            if len(col) != len(self.regs):
                raise ValueError("every entry of PhasesToLink must name one phase per region")
        if not (len(self.xtv) == len(self.opv) == len(self.spv) == len(self.regs)):
            raise ValueError("one XtUVars / OPVars / SPVars list per linked region is required")


class OptimalControlProblem:
    """Multi-phase problem: phase offsets, link parameters, link functions and one shared NonLinearProgram
    (OptimalControlProblem.cpp:115-155 transcribe_phases, :271-387 transcribe_links, :659-764 make_link_Vindex_Cindex).
    Variable order: every phase's variables, then the link parameters; link constraint rows follow all phase rows."""

    def __init__(self):
        self.Phases = []
        self.nlp = None
        self._linkparams = np.zeros(0)
        self._links = {1: [], 2: [], 0: []}  # role -> [_LinkFunction]; insertion order (the reference's std::map by index)

    def addPhase(self, ph):
        self.Phases.append(ph)
        return len(self.Phases) - 1

    def getPhaseNum(self, ph):
        if isinstance(ph, (int, np.integer)):
            n = int(ph)
            return n + len(self.Phases) if n < 0 else n
        for k, q in enumerate(self.Phases):
            if q is ph:
                return k
        raise ValueError("phase is not part of this OptimalControlProblem")

    # ------------------------------------------------------------------ link parameters
    def setLinkParams(self, parm, units=None):
        self._linkparams = np.asarray(parm, dtype=np.float64).ravel().copy()
        self.nlp = None

    def returnLinkParams(self):
        return self._linkparams.copy()

    # ------------------------------------------------------------------ link functions
    def _make_link(self, fun, *a):
        """Argument forms of the reference (OptimalControlProblem.h:371-493, :685-817):
        (fun, packs, lv) with packs = [(phase, region, XtUVars, OPVars, SPVars), ...];
        (fun, p0, reg0, v0, p1, reg1, v1[, lv]);  (fun, p0, reg0, XtUV0, OPV0, SPV0, p1, reg1, XtUV1, OPV1, SPV1[, lv]);
        (fun, LinkFlag, PTL, XtUVars[, OPVars, SPVars][, LinkParams]) -- the flag/PTL interface with many columns."""
        if not isinstance(fun, VectorFunction):
            raise TypeError("expected a VectorFunction")
        if a and isinstance(a[0], (list, tuple)) and a[0] and isinstance(a[0][0], (list, tuple)) and len(a[0][0]) == 5:
            packs, lv = a[0], (a[1] if len(a) > 1 else [])
            phs = [self.getPhaseNum(pk[0]) for pk in packs]
            return _LinkFunction(fun, LinkFlags.ReadRegions, [pk[1] for pk in packs], [phs], [pk[2] for pk in packs],
                                 [pk[3] for pk in packs], [pk[4] for pk in packs], [lv])
        def is_seq(q):
            return isinstance(q, (list, tuple, range, np.ndarray))
        flag_form = bool(a) and len(a) >= 2 and is_seq(a[1]) and len(a[1]) > 0 and is_seq(a[1][0]) and (
            (isinstance(a[0], str) and a[0] in LinkFlags.__dict__) or isinstance(a[0], (int, np.integer)))
        if flag_form:
            flag = LinkFlags.parse(a[0])
            regs = LinkFlags._regions[flag]
            ptl = [[self.getPhaseNum(p) for p in col] for col in a[1]]
            rest = list(a[2:])
            nreg = len(regs)

            def per_region(v):
                v = list(v)
                if nreg and len(v) == nreg and all(isinstance(q, (list, tuple, range, np.ndarray)) for q in v):
                    return [list(q) for q in v]
                return [list(v) for _ in range(nreg)]
            xtv = per_region(rest[0]) if rest else [[] for _ in range(nreg)]
            opv = per_region(rest[1]) if len(rest) >= 4 else [[] for _ in range(nreg)]
            spv = per_region(rest[2]) if len(rest) >= 4 else [[] for _ in range(nreg)]
            lv = rest[3] if len(rest) >= 4 else (rest[1] if len(rest) == 2 else [])
            return _LinkFunction(fun, flag, regs, ptl, xtv, opv, spv, lv)
        if len(a) in (6, 7):
            p0, r0, v0, p1, r1, v1 = a[:6]
            lv = a[6] if len(a) == 7 else []

            def split(reg, v):
                reg = PhaseRegionFlags.parse(reg)
                if reg == PhaseRegionFlags.ODEParams:
                    return [], v, []
                if reg == PhaseRegionFlags.StaticParams:
                    return [], [], v
                return v, [], []
            x0, o0, s0 = split(r0, v0)
            x1, o1, s1 = split(r1, v1)
            return self._make_link(fun, [(p0, r0, x0, o0, s0), (p1, r1, x1, o1, s1)], lv)
        if len(a) in (10, 11):
            lv = a[10] if len(a) == 11 else []
            return self._make_link(fun, [tuple(a[0:5]), tuple(a[5:10])], lv)
        raise TypeError("unrecognised link function arguments")

    def _add_link(self, role, lf):
        self._links[role].append(lf)
        self.nlp = None
        return len(self._links[role]) - 1

    def addLinkEqualCon(self, fun, *a):
        return self._add_link(1, self._make_link(fun, *a))

    def addLinkInequalCon(self, fun, *a):
        return self._add_link(2, self._make_link(fun, *a))

    def addLinkObjective(self, fun, *a):
        if fun.ORows() != 1:
            raise ValueError("a link objective must be scalar")
        return self._add_link(0, self._make_link(fun, *a))

    def _lp_link(self, fun, lpvs):
        lpvs = list(lpvs)
        if lpvs and not isinstance(lpvs[0], (list, tuple, range, np.ndarray)):
            lpvs = [lpvs]
        return _LinkFunction(fun, LinkFlags.LinkParams, [], [], [], [], [], lpvs)

    def addLinkParamEqualCon(self, fun, lpvs):
        """fun applied to link parameters only, once per entry of lpvs (OptimalControlProblem.h:671-680, :825-834)."""
        return self._add_link(1, self._lp_link(fun, lpvs))

    def addLinkParamInequalCon(self, fun, lpvs):
        return self._add_link(2, self._lp_link(fun, lpvs))

    def addLinkParamObjective(self, fun, lpvs):
        return self._add_link(0, self._lp_link(fun, lpvs))

    def addForwardLinkEqualCon(self, iphase, fphase, vars_):
        """x_back(phase i) - x_front(phase i+1) = 0 for i in [iphase, fphase) (OptimalControlProblem.h:580-610)."""
        i0, i1 = self.getPhaseNum(iphase), self.getPhaseNum(fphase)
        v = _idx(vars_)
        args = Arguments(2 * len(v))
        fun = args.head(len(v)) - args.tail(len(v))
        return [self.addLinkEqualCon(fun, i, "Last", v, i + 1, "First", v) for i in range(i0, i1)]

    def addDirectLinkEqualCon(self, *a):
        """(p0, reg0, v0, p1, reg1, v1): v0 of region reg0 of p0 equals v1 of reg1 of p1 (OptimalControlProblem.h:614-635);
        (fun, p0, reg0, v0, p1, reg1, v1) is the same as addLinkEqualCon."""
        if isinstance(a[0], VectorFunction):
            return self.addLinkEqualCon(*a)
        p0, r0, v0, p1, r1, v1 = a
        n = len(_idx(v0))
        args = Arguments(2 * n)
        return self.addLinkEqualCon(args.head(n) - args.tail(n), p0, r0, v0, p1, r1, v1)

    # ------------------------------------------------------------------ transcription
    def _region_vindex(self, pnum, reg, x, o, s):
        ph = self.Phases[pnum]
        xa, xp = _lib.iarr(x)
        oa, op = _lib.iarr(o)
        sa, sp = _lib.iarr(s)
        dims = (C.c_int * 2)()
        _lib.check(lib.ab_phase_region_vindex(ph.h, reg, xp, len(xa), op, len(oa), sp, len(sa), None, 0, dims))
        V = np.zeros((dims[0], dims[1]), dtype=np.int32)
        if V.size:
            _lib.check(lib.ab_phase_region_vindex(ph.h, reg, xp, len(xa), op, len(oa), sp, len(sa),
                                                  V.ctypes.data_as(C.POINTER(C.c_int)), V.size, dims))
        return V

    def _link_vindex(self, lf, lp0):
        """make_link_Vindex_Cindex (OptimalControlProblem.cpp:659-764): [napplications][IRows] variable indices."""
        if lf.flag == LinkFlags.PathToPath:
            cols = []
            for col in lf.ptl:
                parts = [self._region_vindex(col[j], lf.regs[j], lf.xtv[j], lf.opv[j], lf.spv[j]) for j in range(len(lf.regs))]
                if any(p.shape[0] != parts[0].shape[0] for p in parts):
                    raise ValueError("Phases cannot be linked path to path")
                cols.append(np.concatenate(parts, axis=1))
            return np.concatenate(cols, axis=0)
        rows = []
        for i in range(max(len(lf.ptl), len(lf.lv))):
            parts = []
            for j in range(len(lf.regs)):
                V = self._region_vindex(lf.ptl[i][j], lf.regs[j], lf.xtv[j], lf.opv[j], lf.spv[j])
                if V.shape[0] != 1:
                    raise ValueError("link functions read single-application regions (Front, Back, Params, ...)")
                parts.append(V[0])
            lv = np.asarray(lf.lv[i], dtype=np.int32)
            if lv.size and (lv.min() < 0 or lv.max() >= self._linkparams.size):
                raise ValueError("link function references a non-existent link parameter")
            parts.append(lp0 + lv)
            rows.append(np.concatenate(parts).astype(np.int32))
        return np.stack(rows, axis=0)

    def transcribe(self, showstats=False, showfuns=False):
        nlp = NonLinearProgram()
        v = e = i = 0
        self.offsets = []
        for ph in self.Phases:
            if not ph.h:
                raise RuntimeError("setTraj must be called on every phase before transcribe")
            out = (C.c_int * 3)()
            _lib.check(lib.ab_phase_transcribe(ph.h, nlp.h, v, e, i, out))
            self.offsets.append((v, e, i))
            v, e, i = v + out[0], e + out[1], i + out[2]
        self.LinkParamLocs = v + np.arange(self._linkparams.size)
        lp0 = v
        v += self._linkparams.size
        nxt = {1: e, 2: i}
        # equalities, inequalities, objectives -- the order of transcribe_links; rows are numbered application-major
        for role in (1, 2, 0):
            for lf in self._links[role]:
                V = self._link_vindex(lf, lp0)
                if V.shape[1] != lf.fun.IRows():
                    raise ValueError(f"link function expects {lf.fun.IRows()} inputs but its regions supply {V.shape[1]}")
                tape = nlp.add_tape(lf.fun)
                cidx = None
                if role:
                    n = V.shape[0] * lf.fun.ORows()
                    cidx = (nxt[role] + np.arange(n, dtype=np.int32)).reshape(V.shape[0], lf.fun.ORows())
                    nxt[role] += n
                nlp.add_function(role, tape, lf.fun.is_linear(), V, cidx)
        nlp.set_dims(v, nxt[1], nxt[2])
        nlp.analyze()
        self.nlp = nlp
        return nlp

    def makeSolverInput(self):
        return np.concatenate([ph.makeSolverInput() for ph in self.Phases] + [self._linkparams])

    def optimize(self, *a, **k):
        raise NotImplementedError("PSIOPT is out of scope for this build; use ocp.transcribe() and nlp.evalKKT(...)")

    solve = optimize
