#include "phase.hpp"

#include <cmath>
#include <stdexcept>

namespace ab {

void PhaseIndexer::set_dimensions(int dcs, int ndef, bool block) {
  numDefects = ndef;
  DCS = dcs;
  numStates = (dcs - 1) * ndef + 1;
  numNodalStates = ndef + 1;
  blocked = block;
  const int n = XtUV();
  firstLocs.resize(n);
  lastLocs.resize(n);
  for (int i = 0; i < n; i++) firstLocs[i] = i;
  if (blocked) {
    numPhaseVars = numStates * XtV() + numDefects * UV + PV + SPV;
    for (int i = XtV(); i < n; i++) firstLocs[i] += XtV() * numStates - XtV();
    for (int i = 0; i < n; i++) lastLocs[i] = firstLocs[i] + XtV() * (numStates - 1);
    for (int i = XtV(); i < n; i++) lastLocs[i] = firstLocs[i] + UV * (numDefects - 1);
  } else {
    numPhaseVars = numStates * n + PV + SPV;
    for (int i = 0; i < n; i++) lastLocs[i] = firstLocs[i] + n * (numStates - 1);
  }
  odeParamLocs.resize(PV);
  for (int i = 0; i < PV; i++) odeParamLocs[i] = i + numPhaseVars - PV - SPV;
  staticParamLocs.resize(SPV);
  for (int i = 0; i < SPV; i++) staticParamLocs[i] = i + numPhaseVars - SPV;
}

void PhaseIndexer::begin_indexing(int voff, int eoff, int ioff) {
  for (int& v: firstLocs) v += voff;
  for (int& v: lastLocs) v += voff;
  for (int& v: odeParamLocs) v += voff;
  for (int& v: staticParamLocs) v += voff;
  nextEq = eoff;
  nextIq = ioff;
  numEq = numIq = 0;
}

int PhaseIndexer::loc(int v, int state) const {
  if (blocked) {
    if (v < XtV()) return firstLocs[v] + state * XtV();
    int unum = state / (DCS - 1);
    if (unum > numDefects - 1) unum = numDefects - 1;
    return firstLocs[v] + unum * UV;
  }
  return firstLocs[v] + state * XtUV();
}

int PhaseIndexer::loc(int v, int state, int defect) const {
  if (blocked) {
    if (v < XtV()) return firstLocs[v] + state * XtV();
    return firstLocs[v] + defect * UV;
  }
  return firstLocs[v] + state * XtUV();
}

void PhaseIndexer::make_vindex_cindex(int region, const std::vector<int>& xtuv, const std::vector<int>& opv,
                                      const std::vector<int>& spv, int orows, int& next_c, std::vector<int>& V,
                                      std::vector<int>& C, int& nappl) const {
  const int xs = (int) xtuv.size();
  bool only_control = true;
  for (int v: xtuv)
    if (v < XtV()) only_control = false;
  for (int v: xtuv)
    if (v < 0 || v >= XtUV()) throw std::out_of_range("make_vindex_cindex: XtU variable index out of range");
  for (int v: opv)
    if (v < 0 || v >= PV) throw std::out_of_range("make_vindex_cindex: ODE parameter index out of range");
  for (int v: spv)
    if (v < 0 || v >= SPV) throw std::out_of_range("make_vindex_cindex: static parameter index out of range");
  V.clear();
  C.clear();
  auto from = [&](const std::vector<int>& locs, const std::vector<int>& vars) {
    for (int v: vars) V.push_back(locs[v]);
  };
  auto path = [&](int state) {
    for (int v: xtuv) V.push_back(loc(v, state));
  };
  auto dpath = [&](int state, int defect) {
    for (int v: xtuv) V.push_back(loc(v, state, defect));
  };
  auto params = [&]() {
    from(odeParamLocs, opv);
    from(staticParamLocs, spv);
  };
  const bool ctl_blk = only_control && blocked;
  const int step = DCS - 1;
  nappl = 0;
  switch (region) {
    case R_ODEPARAMS:
    case R_STATICPARAMS:
    case R_PARAMS:
      params();
      nappl = 1;
      break;
    case R_FRONT:
      from(firstLocs, xtuv);
      params();
      nappl = 1;
      break;
    case R_BACK:
      from(lastLocs, xtuv);
      params();
      nappl = 1;
      break;
    case R_FRONTANDBACK:
      from(firstLocs, xtuv);
      from(lastLocs, xtuv);
      params();
      nappl = 1;
      break;
    case R_BACKANDFRONT:
      from(lastLocs, xtuv);
      from(firstLocs, xtuv);
      params();
      nappl = 1;
      break;
    case R_PATH:
      if (ctl_blk) {
        for (int i = 0; i < numDefects; i++) {
          path(i * step);
          params();
        }
        nappl = numDefects;
      } else {
        for (int i = 0; i < numStates; i++) {
          path(i);
          params();
        }
        nappl = numStates;
      }
      break;
    case R_INNERPATH:
      if (ctl_blk) {
        for (int i = 1; i < numDefects - 1; i++) {
          path(i * step);
          params();
        }
        nappl = std::max(numDefects - 2, 0);
      } else {
        for (int i = 1; i < numStates - 1; i++) {
          path(i);
          params();
        }
        nappl = std::max(numStates - 2, 0);
      }
      break;
    case R_NODALPATH:
      if (ctl_blk) {
        for (int i = 0; i < numDefects; i++) {
          path(i * step);
          params();
        }
        nappl = numDefects;
      } else {
        for (int i = 0; i < numDefects + 1; i++) {
          path(i * step);
          params();
        }
        nappl = numDefects + 1;
      }
      break;
    case R_PAIRWISEPATH:
      if (ctl_blk) {
        for (int i = 0; i < numDefects - 1; i++) {
          path(i * step);
          path((i + 1) * step);
          params();
        }
        nappl = numDefects - 1;
      } else {
        for (int i = 0; i < numStates - 1; i++) {
          path(i);
          path(i + 1);
          params();
        }
        nappl = numStates - 1;
      }
      break;
    case R_FRONTNODALBACKPATH:
      if (ctl_blk) {
        if (!opv.empty() || !spv.empty())
          throw std::invalid_argument("FrontNodalBackPath with parameters is not supported for block-constant controls");
        for (int i = 1; i < numDefects - 1; i++) {
          from(firstLocs, xtuv);
          path(i * step);
          from(lastLocs, xtuv);
        }
        nappl = std::max(numDefects - 2, 0);
      } else {
        for (int i = 1; i < numDefects; i++) {
          from(firstLocs, xtuv);
          path(i * step);
          from(lastLocs, xtuv);
          params();
        }
        nappl = numDefects - 1;
      }
      break;
    case R_DEFECTPATH:
      for (int i = 0; i < numDefects; i++) {
        for (int j = 0; j < DCS; j++) dpath(i * step + j, i);
        params();
      }
      nappl = numDefects;
      break;
    case R_BLOCKDEFECTPATH:
      for (int i = 0; i < numDefects; i++) {
        for (int j = 0; j < DCS; j++)
          for (int k = 0; k < XtV(); k++) V.push_back(loc(k, i * step + j, i));
        for (int k = XtV(); k < XtUV(); k++) V.push_back(loc(k, i * step, i));
        from(odeParamLocs, opv);
      }
      nappl = numDefects;
      break;
    case R_DEFECTPAIRWISEPATH:
      for (int i = 0; i < numDefects - 1; i++) {
        for (int j = 0; j < 2 * DCS - 1; j++) dpath(i * step + j, i);
        params();
      }
      nappl = numDefects - 1;
      break;
    default: throw std::invalid_argument("Indexing for requested phase region has not been implemented");
  }
  (void) xs;
  C.resize((size_t) nappl * orows);
  for (size_t q = 0; q < C.size(); q++) C[q] = next_c++;
}

void Phase::init_indexing() {
  idx.XV = XV;
  idx.UV = UV;
  idx.PV = PV;
  idx.SPV = SPV;
  idx.set_dimensions(cardinal_states(), num_defects(), blocked());
}

std::vector<double> Phase::state_spacing() const {
  const int CS = cardinal_states();
  const LGLTable& T = lgl_table(CS);
  std::vector<double> s;
  for (size_t b = 0; b < defs_per_bin.size(); b++) {
    const double lo = bin_spacing[b], hi = bin_spacing[b + 1];
    const int nd = defs_per_bin[b];
    for (int d = 0; d < nd; d++) {
      const double t0 = lo + (hi - lo) * d / nd, t1 = lo + (hi - lo) * (d + 1) / nd;
      for (int j = 0; j < CS - 1; j++) s.push_back(t0 + (t1 - t0) * T.tau_card[j]);
    }
  }
  s.push_back(bin_spacing.back());
  return s;
}

namespace {
int add_tape(Problem& prob, ScalarFn& f, bool linear, const std::string& desc) {
  if (!linear) f.make_gradient();
  prob.tapes.push_back(emit_tape(f, !linear));
  prob.tapes.back().desc = desc;
  prob.tapes.back().fn = std::make_shared<ScalarFn>(f);
  return (int) prob.tapes.size() - 1;
}
}  // namespace

void Phase::transcribe(Problem& prob, int voff, int eoff, int ioff) {
  if (!g || ode_root < 0) throw std::invalid_argument("Phase::transcribe: no ODE set");
  init_indexing();
  idx.begin_indexing(voff, eoff, ioff);
  const int CS = cardinal_states();
  const int N = num_defects();
  const std::vector<int> empty;
  std::vector<int> stateT(idx.XtUV()), oparT(PV);
  for (int i = 0; i < idx.XtUV(); i++) stateT[i] = i;
  for (int i = 0; i < PV; i++) oparT[i] = i;

  auto add_fn = [&](int role, int tape, bool linear, int region, const std::vector<int>& xtuv, const std::vector<int>& opv,
                    const std::vector<int>& spv, const std::string& name) -> FuncRec& {
    FuncRec fr;
    fr.role = role;
    fr.tape = tape;
    fr.ir = prob.tapes[tape].nin;
    fr.orows = prob.tapes[tape].nout;
    fr.nappconst = prob.tapes[tape].nappconst;
    fr.linear = linear;
    fr.name = name;
    int dummy = 0;
    int& next = (role == 1) ? idx.nextEq : (role == 2 ? idx.nextIq : dummy);
    const int before = next;
    idx.make_vindex_cindex(region, xtuv, opv, spv, role == 0 ? 0 : fr.orows, next, fr.vindex, fr.cindex, fr.nappl);
    if (fr.nappl > 0 && (int) (fr.vindex.size() / fr.nappl) != fr.ir)
      throw std::invalid_argument("function '" + name + "' expects " + std::to_string(fr.ir) + " inputs but its region supplies "
                                  + std::to_string(fr.vindex.size() / fr.nappl));
    if (role == 1) idx.numEq += next - before;
    if (role == 2) idx.numIq += next - before;
    fr.appconst.assign((size_t) fr.nappl * fr.nappconst, 0.0);
    prob.funcs.push_back(std::move(fr));
    return prob.funcs.back();
  };

  // (1) dynamics
  {
    const bool blk = blocked() && !(UV == 0 && PV == 0);
    ScalarFn f;
    if (tmode == TM_TRAPEZOIDAL)
      f = blk ? make_trap_defect(*g, ode_root, XV, 0, UV + PV) : make_trap_defect(*g, ode_root, XV, UV, PV);
    else
      f = blk ? make_lgl_defect(*g, ode_root, CS, XV, 0, UV + PV) : make_lgl_defect(*g, ode_root, CS, XV, UV, PV);
    const std::string sizes = blk ? (std::to_string(CS) + " " + std::to_string(XV) + " 0 " + std::to_string(UV + PV))
                                  : (std::to_string(CS) + " " + std::to_string(XV) + " " + std::to_string(UV) + " " + std::to_string(PV));
    int tp = add_tape(prob, f, false, std::string(tmode == TM_TRAPEZOIDAL ? "trap_defect " : "lgl_defect ") + sizes + " " + g->describe(ode_root));
    add_fn(1, tp, false, blk ? R_BLOCKDEFECTPATH : R_DEFECTPATH, stateT, oparT, empty, "defect");
  }
  // (2) axis functions
  {
    std::vector<int> tloc {XV};
    if (tmode == TM_LGL7 || tmode == TM_LGL5) {
      ScalarFn f = make_lgl_mesh_spacing(CS);
      int tp = add_tape(prob, f, false, "lgl_mesh_spacing " + std::to_string(CS));
      add_fn(1, tp, false, R_DEFECTPATH, tloc, empty, empty, "lgl_mesh_spacing");
    }
    if (idx.numNodalStates - 2 > 0) {
      // N-1 single-application SingleMeshSpacing(c_{i+1}) functions in the reference
      // (addPartitionedEquality, PhaseIndexer.cpp:20-39); batched here as one function with a per-application constant.
      std::vector<double> cspace((size_t) N + 1);
      int start = 0;
      for (size_t b = 0; b < defs_per_bin.size(); b++) {
        // Eigen's setLinSpaced(size, low, high) arithmetic (ODEPhaseBase.cpp:963-969)
        const int sz = defs_per_bin[b] + 1;
        const double lo = bin_spacing[b], hi = bin_spacing[b + 1];
        const double step = (sz == 1) ? 0.0 : (hi - lo) / (double) (sz - 1);
        const bool flip = std::fabs(hi) < std::fabs(lo);
        for (int k = 0; k < sz; k++) {
          double v;
          if (sz == 1)
            v = hi;
          else if (flip)
            v = (k == 0) ? lo : hi - (double) (sz - 1 - k) * step;
          else
            v = (k == sz - 1) ? hi : lo + (double) k * step;
          cspace[(size_t) start + k] = v;
        }
        start += defs_per_bin[b];
      }
      ScalarFn f = make_single_mesh_spacing();
      int tp = add_tape(prob, f, true, "single_mesh_spacing");
      FuncRec& fr = add_fn(1, tp, true, R_FRONTNODALBACKPATH, tloc, empty, empty, "single_mesh_spacing");
      for (int a = 0; a < fr.nappl; a++) fr.appconst[a] = cspace[(size_t) a + 1];
    }
  }
  // (3) control splines
  if (UV > 0 && (tmode == TM_LGL5 || tmode == TM_LGL7) && (cmode == CM_HIGHEST_ORDER_SPLINE || cmode == CM_FIRST_ORDER_SPLINE)) {
    int order = 1;
    if (tmode == TM_LGL7 && cmode == CM_HIGHEST_ORDER_SPLINE) order = 2;
    std::vector<int> tu;
    for (int i = XV; i < idx.XtUV(); i++) tu.push_back(i);
    ScalarFn f = make_control_spline(CS, UV, order);
    int tp = add_tape(prob, f, false, "control_spline " + std::to_string(CS) + " " + std::to_string(UV) + " " + std::to_string(order));
    add_fn(1, tp, false, R_DEFECTPAIRWISEPATH, tu, empty, empty, "control_spline");
  }
  // (4) integrals
  std::string integral_desc;
  auto integral_fn = [&](const StateFn& sf, int& region) {
    int xp = (int) sf.xtuv.size();
    int sop = (int) (sf.opv.size() + sf.spv.size());
    int ics = 2;
    region = R_PAIRWISEPATH;
    if (imode == IM_BASE && (tmode == TM_LGL5 || tmode == TM_LGL7)) {
      ics = CS;
      region = R_DEFECTPATH;
    }
    integral_desc = "lgl_integral " + std::to_string(ics) + " " + std::to_string(xp) + " " + std::to_string(sop) + " " + g->describe(sf.root);
    return make_lgl_integral(*g, sf.root, ics, xp, sop);
  };
  for (const StateFn& sf: integrals) {
    int region;
    ScalarFn f = integral_fn(sf, region);
    if (!(imode == IM_BASE && (tmode == TM_LGL5 || tmode == TM_LGL7)))
      region = (tmode == TM_LGL3 || tmode == TM_TRAPEZOIDAL) ? R_DEFECTPATH : R_PAIRWISEPATH;
    std::vector<int> xtrap = sf.xtuv;
    xtrap.push_back(XV);
    int tp = add_tape(prob, f, false, integral_desc);
    add_fn(0, tp, false, region, xtrap, sf.opv, sf.spv, "integral_objective");
  }
  for (const StateFn& sf: param_integrals) {
    int region;
    ScalarFn f = integral_fn(sf, region);
    std::vector<int> xtrap = sf.xtuv;
    xtrap.push_back(XV);
    // accumulation (PhaseIndexer.cpp:41-76): -AccScale*sp_k and every segment's integral share one row
    ScalarFn acc;
    acc.nin = 1;
    acc.nout = 1;
    acc.outs.push_back(acc.dag.mulc(acc.dag.input(0), -1.0));
    int tacc = add_tape(prob, acc, true, "expr (scaled 1 1 1 [] [-1] (args 1 1 1 [] []))");
    FuncRec& fa = add_fn(1, tacc, true, R_PARAMS, empty, empty, sf.ext, "accumulation_param");
    fa.unique_constraints = false;
    const int row = fa.cindex[0];
    int tp = add_tape(prob, f, false, integral_desc);
    int save = idx.nextEq, saven = idx.numEq;
    FuncRec& fi = add_fn(1, tp, false, region, xtrap, sf.opv, sf.spv, "accumulation_integral");
    idx.nextEq = save;  // the integral claims no rows of its own
    idx.numEq = saven;
    for (int& c: fi.cindex) c = row;
    fi.unique_constraints = false;
  }
  // (5) user functions
  auto user = [&](const std::vector<StateFn>& fns, int role, const char* nm) {
    for (const StateFn& sf: fns) {
      ScalarFn f = lower_function(*g, sf.root);
      const bool lin = g->at(sf.root).tlin;
      int tp = add_tape(prob, f, lin, "expr " + g->describe(sf.root));
      add_fn(role, tp, lin, sf.region, sf.xtuv, sf.opv, sf.spv, nm);
    }
  };
  user(eqs, 1, "user_equality");
  user(iqs, 2, "user_inequality");
  user(objs, 0, "user_objective");
}

}  // namespace ab
