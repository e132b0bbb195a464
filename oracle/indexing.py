"""Independent restatement of the reference's index maps (test infrastructure; never imported by the product).

Written from the reference sources only -- it calls nothing in csrc/phase.cpp -- so that "bit-exact index maps" is a
comparison between two implementations and not the product with itself:

  * PhaseIndexer::set_dimensions              /root/reference/src/OptimalControl/PhaseIndexer.h:63-99
  * PhaseIndexer::begin_indexing              PhaseIndexer.h:100-126
  * PhaseIndexer::getXTUVarLoc (both forms)   PhaseIndexer.h:168-197
  * PhaseIndexer::make_Vindex_Cindex          PhaseIndexer.cpp:132-410 (all region flags, blocked and non-blocked)
  * addEquality / addPartitionedEquality / addAccumulation / addInequality / addObjective   PhaseIndexer.cpp:3-130
  * ODEPhase::transcribe_dynamics             ODEPhase.h:218-341
  * transcribe_axis_funcs / control / integrals / basic   ODEPhaseBase.cpp:743-1061
  * transcribe_phase emission order           ODEPhaseBase.cpp:1364-1381
  * OptimalControlProblem::transcribe_phases  OptimalControlProblem.cpp:115-155
  * transcribe_links / make_link_Vindex_Cindex   OptimalControlProblem.cpp:271-387, :659-764

Input is the *user-level* description a Python phase object carries (ODE sizes, transcription / control / integral mode,
number of segments, the functions as the user added them); output is the reference's three function lists
(Objectives, EqualityConstraints, InequalityConstraints), each entry a dict with Vindex [applications, inputs] and Cindex
[applications, rows] -- the transposes of the reference's column-per-application matrices.
"""
import numpy as np

# PhaseRegionFlags (OptimalControlFlags.h), numbered as the product's Python enum numbers them
(NotSet, Front, Back, FrontandBack, BackandFront, Path, InnerPath, NodalPath, DefectPath, PairWisePath,
 DefectPairWisePath, FrontNodalBackPath, Params, ODEParams, StaticParams, Accumulate, BlockDefectPath) = range(17)
LGL3, LGL5, LGL7, Trapezoidal, CentralShooting = range(5)
HighestOrderSpline, FirstOrderSpline, NoSpline, BlockConstant = range(4)
BaseIntegral, SimpsonIntegral, TrapIntegral = range(3)
CARDINAL = {LGL3: 2, LGL5: 3, LGL7: 4, Trapezoidal: 2}


class Indexer:
    """PhaseIndexer for one phase."""

    def __init__(self, XV, UV, PV, SPV):
        self.XV, self.UV, self.PV, self.SPV = XV, UV, PV, SPV
        self.XtV = XV + 1
        self.XtUV = XV + 1 + UV

    def set_dimensions(self, DCS, Dnum, blocked):
        self.numDefects, self.DCS, self.blocked = Dnum, DCS, blocked
        self.numStates = (DCS - 1) * Dnum + 1
        self.numNodalStates = Dnum + 1
        first = list(range(self.XtUV))
        if blocked:
            self.numPhaseVars = self.numStates * self.XtV + Dnum * self.UV + self.PV + self.SPV
            for k in range(self.XtV, self.XtUV):
                first[k] += self.XtV * self.numStates - self.XtV
            last = [v + self.XtV * (self.numStates - 1) for v in first]
            for k in range(self.XtV, self.XtUV):
                last[k] = first[k] + self.UV * (Dnum - 1)
        else:
            self.numPhaseVars = self.numStates * self.XtUV + self.PV + self.SPV
            last = [v + self.XtUV * (self.numStates - 1) for v in first]
        self.first, self.last = first, last
        self.odep = [k + self.numPhaseVars - self.PV - self.SPV for k in range(self.PV)]
        self.statp = [k + self.numPhaseVars - self.SPV for k in range(self.SPV)]

    def begin_indexing(self, n, ep, ip):
        self.first = [v + n for v in self.first]
        self.last = [v + n for v in self.last]
        self.odep = [v + n for v in self.odep]
        self.statp = [v + n for v in self.statp]
        self.nextEq, self.nextIq = ep, ip
        self.numEq = self.numIq = 0

    def loc(self, v, state, defect=None):
        if self.blocked:
            if v < self.XtV:
                return self.first[v] + state * self.XtV
            if defect is None:
                defect = min(state // (self.DCS - 1), self.numDefects - 1)
            return self.first[v] + defect * self.UV
        return self.first[v] + state * self.XtUV

    def vindex_cindex(self, reg, x, op, sp, orows, next_c):
        """Returns (Vindex columns, Cindex columns, next constraint row)."""
        x, op, sp = list(x), list(op), list(sp)
        only_control = all(v >= self.XtV for v in x)
        D, N, S = self.DCS, self.numDefects, self.numStates
        tail = [self.odep[k] for k in op] + [self.statp[k] for k in sp]

        def at(state):
            return [self.loc(v, state) for v in x]

        def at_defect(state, defect):
            return [self.loc(v, state, defect) for v in x]

        blk_u = only_control and self.blocked
        if reg in (ODEParams, StaticParams, Params):
            cols = [tail]
        elif reg == Front:
            cols = [[self.first[v] for v in x] + tail]
        elif reg == Back:
            cols = [[self.last[v] for v in x] + tail]
        elif reg == FrontandBack:
            cols = [[self.first[v] for v in x] + [self.last[v] for v in x] + tail]
        elif reg == BackandFront:
            cols = [[self.last[v] for v in x] + [self.first[v] for v in x] + tail]
        elif reg == Path:
            cols = [at(i * (D - 1)) + tail for i in range(N)] if blk_u else [at(i) + tail for i in range(S)]
        elif reg == InnerPath:
            cols = ([at(i * (D - 1)) + tail for i in range(1, N - 1)] if blk_u else [at(i) + tail for i in range(1, S - 1)])
        elif reg == NodalPath:
            cols = [at(i * (D - 1)) + tail for i in range(N if blk_u else N + 1)]
        elif reg == PairWisePath:
            if blk_u:
                cols = [at(i * (D - 1)) + at((i + 1) * (D - 1)) + tail for i in range(N - 1)]
            else:
                cols = [at(i) + at(i + 1) + tail for i in range(S - 1)]
        elif reg == FrontNodalBackPath:
            if blk_u:
                # the reference writes the parameters of this branch into column i instead of i - 1
                # (PhaseIndexer.cpp:334-335); only reachable with parameters, which no caller passes
                if tail:
                    raise NotImplementedError("FrontNodalBackPath on block-constant controls with parameters")
                rng = range(1, N - 1)
            else:
                rng = range(1, N)
            cols = [[self.first[v] for v in x] + at(i * (D - 1)) + [self.last[v] for v in x] + tail for i in rng]
        elif reg == DefectPath:
            cols = []
            for i in range(N):
                c = []
                for j in range(D):
                    c += at_defect(i * (D - 1) + j, i)
                cols.append(c + tail)
        elif reg == BlockDefectPath:
            cols = []
            for i in range(N):
                c = []
                for j in range(D):
                    c += [self.loc(k, i * (D - 1) + j, i) for k in range(self.XtV)]
                c += [self.loc(k, i * (D - 1), i) for k in range(self.XtV, self.XtUV)]
                cols.append(c + [self.odep[k] for k in op])  # static parameters are not appended (PhaseIndexer.cpp:388)
        elif reg == DefectPairWisePath:
            cols = []
            for i in range(N - 1):
                c = []
                for j in range(2 * D - 1):
                    c += at_defect(i * (D - 1) + j, i)
                cols.append(c + tail)
        else:
            raise ValueError("Indexing for requested phase region has not been implemented")
        cin = []
        for _ in cols:
            cin.append(list(range(next_c, next_c + orows)))
            next_c += orows
        return cols, cin, next_c


def _fn(name, V, Cc, unique=True):
    V = np.asarray(V, dtype=np.int64).reshape(len(V), -1)
    Cc = np.asarray(Cc, dtype=np.int64).reshape(len(V), -1)
    return {"name": name, "Vindex": V, "Cindex": Cc, "unique_constraints": unique}


class Lists:
    def __init__(self):
        self.obj, self.eq, self.iq = [], [], []


def make_indexer(ph):
    """initIndexing for a product-side Python phase object (reads only what the user set)."""
    ix = Indexer(ph.ode.XV, ph.ode.UV, ph.ode.PV, ph._spv)
    ix.set_dimensions(CARDINAL[ph.tmode], ph._nseg, ph._cmode == BlockConstant)
    return ix


def transcribe_phase(ph, ix, vo, eqo, iqo, out):
    """ODEPhaseBase::transcribe_phase: dynamics, axis, control, integrals, basic functions -- in that order."""
    XV, UV, PV = ph.ode.XV, ph.ode.UV, ph.ode.PV
    ix.begin_indexing(vo, eqo, iqo)
    CS = ix.DCS
    tmode, cmode, imode = ph.tmode, ph._cmode, ph._imode

    def add_eq(name, reg, x, op, sp, orows, unique=True):
        V, Cc, ix.nextEq = ix.vindex_cindex(reg, x, op, sp, orows, ix.nextEq)
        out.eq.append(_fn(name, V, Cc, unique))
        return V, Cc

    def add_iq(name, reg, x, op, sp, orows):
        V, Cc, ix.nextIq = ix.vindex_cindex(reg, x, op, sp, orows, ix.nextIq)
        out.iq.append(_fn(name, V, Cc))

    def add_obj(name, reg, x, op, sp):
        V, _, _ = ix.vindex_cindex(reg, x, op, sp, 1, 0)
        out.obj.append(_fn(name, V, [[] for _ in V]))

    # transcribe_dynamics: LGL / trapezoidal defects, XV rows per segment (CS - 1 blocks of XV rows for LGL)
    stateT, oparT = list(range(XV + 1 + UV)), list(range(PV))
    blocked_ode = cmode == BlockConstant and not (UV == 0 and PV == 0)
    defect_rows = XV * (CS - 1)
    add_eq("defect", BlockDefectPath if blocked_ode else DefectPath, stateT, oparT, [], defect_rows)
    # transcribe_axis_funcs
    tloc = [XV]
    if tmode == LGL7:
        add_eq("lgl_mesh_spacing", DefectPath, tloc, [], [], 2)       # LGLMeshSpacing<4>: CSC - 2 rows (MeshSpacingConstraints.h:103)
    elif tmode == LGL5:
        add_eq("lgl_mesh_spacing", DefectPath, tloc, [], [], 1)       # LGLMeshSpacing<3>
    # addPartitionedEquality: one indexing call, then one single-column function per interior node
    V, Cc, ix.nextEq = ix.vindex_cindex(FrontNodalBackPath, tloc, [], [], 1, ix.nextEq)
    for v, c in zip(V, Cc):
        out.eq.append(_fn("single_mesh_spacing", [v], [c]))
    # transcribe_control_funcs
    if UV > 0 and tmode in (LGL5, LGL7) and cmode in (HighestOrderSpline, FirstOrderSpline):
        order = 2 if (tmode == LGL7 and cmode == HighestOrderSpline) else 1
        tu = list(range(XV, XV + 1 + UV))
        add_eq("control_spline", DefectPairWisePath, tu, [], [], _spline_rows(CS, UV, order))
    # transcribe_integrals
    def integral_region():
        if imode == BaseIntegral and tmode in (LGL5, LGL7):
            return DefectPath
        return None
    for (role, fun, reg, x, o, s, e) in ph._funcs:
        if role != 3:
            continue
        r = integral_region()
        if r is None:
            r = DefectPath if tmode in (LGL3, Trapezoidal) else PairWisePath
        add_obj("integral_objective", r, list(x) + [XV], o, s)
    for (role, fun, reg, x, o, s, e) in ph._funcs:
        if role != 4:
            continue
        r = integral_region() or PairWisePath
        # addAccumulation: the accumulator row first (Params region over the named static parameter), then the integral
        # with every application writing that same row
        Vacc, Cacc = add_eq("accumulation_param", Params, [], [], e, 1, unique=False)
        V, _, _ = ix.vindex_cindex(r, list(x) + [XV], o, s, 1, 0)
        out.eq.append(_fn("accumulation_integral", V, [Cacc[0] for _ in V], False))
    # transcribe_basic_funcs: equalities, inequalities, state objectives, each in the user's insertion order
    for (role, fun, reg, x, o, s, e) in ph._funcs:
        if role == 1:
            add_eq("user_equality", reg, x, o, s, fun.orows)
    for (role, fun, reg, x, o, s, e) in ph._funcs:
        if role == 2:
            add_iq("user_inequality", reg, x, o, s, fun.orows)
    for (role, fun, reg, x, o, s, e) in ph._funcs:
        if role == 0:
            add_obj("user_objective", reg, x, o, s)
    ix.numEq, ix.numIq = ix.nextEq - eqo, ix.nextIq - iqo


def _spline_rows(CS, UV, order):
    """Output rows of LGLControlSpline<CSC, USZ, Order> (LGLControlSplines.h:9): USZ * Order."""
    return UV * order


def transcribe_single_phase(ph):
    out = Lists()
    ix = make_indexer(ph)
    transcribe_phase(ph, ix, 0, 0, 0, out)
    return out, {"nvars": ix.numPhaseVars, "neq": ix.numEq, "niq": ix.numIq}


def transcribe_ocp(ocp):
    """OptimalControlProblem::transcribe_phases then transcribe_links."""
    out = Lists()
    ixs = [make_indexer(ph) for ph in ocp.Phases]
    nv = [ix.numPhaseVars for ix in ixs]
    ne, ni = [], []
    for k, (ph, ix) in enumerate(zip(ocp.Phases, ixs)):
        transcribe_phase(ph, ix, sum(nv[:k]), sum(ne), sum(ni), out)
        ne.append(ix.numEq)
        ni.append(ix.numIq)
    next_eq, next_iq = sum(ne), sum(ni)
    link_start = sum(nv)
    nlink = len(ocp._linkparams)

    def link_vc(lf, orows, next_c):
        PATH_TO_PATH = 5
        cols = []
        if lf.flag == PATH_TO_PATH:
            for col in lf.ptl:
                parts = [ixs[p].vindex_cindex(r, lf.xtv[j], lf.opv[j], lf.spv[j], 1, 0)[0] for j, (p, r) in enumerate(zip(col, lf.regs))]
                n = len(parts[0])
                if any(len(q) != n for q in parts):
                    raise ValueError("Phases cannot be linked path to path")
                for a in range(n):
                    c = []
                    for q in parts:
                        c += q[a]
                    cols.append(c)
        else:
            ncols = max(len(lf.ptl), len(lf.lv))
            for i in range(ncols):
                c = []
                for j, r in enumerate(lf.regs):
                    c += ixs[lf.ptl[i][j]].vindex_cindex(r, lf.xtv[j], lf.opv[j], lf.spv[j], 1, 0)[0][0]
                c += [link_start + k for k in lf.lv[i]]
                cols.append(c)
        cin = []
        for _ in cols:
            cin.append(list(range(next_c, next_c + orows)))
            next_c += orows
        return cols, cin, next_c

    for lf in ocp._links[1]:
        V, Cc, next_eq = link_vc(lf, lf.fun.orows, next_eq)
        out.eq.append(_fn("link_equality", V, Cc))
    for lf in ocp._links[2]:
        V, Cc, next_iq = link_vc(lf, lf.fun.orows, next_iq)
        out.iq.append(_fn("link_inequality", V, Cc))
    for lf in ocp._links[0]:
        V, _, _ = link_vc(lf, 1, 0)
        out.obj.append(_fn("link_objective", V, [[] for _ in V]))
    return out, {"nvars": link_start + nlink, "neq": next_eq, "niq": next_iq}
