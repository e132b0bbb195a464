// Host-side transcription: phase variable layout, region index generation and the ordered function list.
//
// Restates /root/reference/src/OptimalControl/PhaseIndexer.{h,cpp} (layout h:63-99, index generation
// cpp:132-410), ODEPhase::transcribe_dynamics (ODEPhase.h:218-341) and
// ODEPhaseBase::transcribe_axis_funcs / control_funcs / integrals / basic_funcs
// (ODEPhaseBase.cpp:962-1005, :1007-1061, :743-890, :891-961) in the order fixed by
// transcribe_phase (:1364-1381).  Runs once per transcription on the host.
#pragma once
#include <map>
#include <string>
#include <vector>

#include "builtins.hpp"
#include "nlp.hpp"

namespace ab {

enum Region : int {
  R_NOTSET = 0,
  R_FRONT,
  R_BACK,
  R_FRONTANDBACK,
  R_BACKANDFRONT,
  R_PATH,
  R_INNERPATH,
  R_NODALPATH,
  R_DEFECTPATH,
  R_PAIRWISEPATH,
  R_DEFECTPAIRWISEPATH,
  R_FRONTNODALBACKPATH,
  R_PARAMS,
  R_ODEPARAMS,
  R_STATICPARAMS,
  R_ACCUMULATE,
  R_BLOCKDEFECTPATH,
};
enum TranscriptionMode : int { TM_LGL3 = 0, TM_LGL5 = 1, TM_LGL7 = 2, TM_TRAPEZOIDAL = 3 };
enum ControlMode : int { CM_HIGHEST_ORDER_SPLINE = 0, CM_FIRST_ORDER_SPLINE = 1, CM_NO_SPLINE = 2, CM_BLOCK_CONSTANT = 3 };
enum IntegralMode : int { IM_BASE = 0, IM_SIMPSON = 1, IM_TRAP = 2 };

struct PhaseIndexer {
  int XV = 0, UV = 0, PV = 0, SPV = 0;
  int numDefects = 0, numStates = 0, numNodalStates = 0, DCS = 2;
  bool blocked = false;
  int numPhaseVars = 0;
  std::vector<int> firstLocs, lastLocs, odeParamLocs, staticParamLocs;
  int nextEq = 0, nextIq = 0, numEq = 0, numIq = 0;

  int XtV() const { return XV + 1; }
  int XtUV() const { return XV + 1 + UV; }
  void set_dimensions(int dcs, int ndef, bool block);
  void begin_indexing(int voff, int eoff, int ioff);
  int loc(int v, int state) const;
  int loc(int v, int state, int defect) const;
  // vindex is [nappl][ir], cindex [nappl][orows]; next_c advances by nappl*orows
  void make_vindex_cindex(int region, const std::vector<int>& xtuv, const std::vector<int>& opv,
                          const std::vector<int>& spv, int orows, int& next_c, std::vector<int>& vindex,
                          std::vector<int>& cindex, int& nappl) const;
};

struct StateFn {
  int root = -1;  // expression node in the phase's graph
  int region = R_NOTSET;
  std::vector<int> xtuv, opv, spv, ext;
};

struct Phase {
  Graph* g = nullptr;
  int ode_root = -1;
  int XV = 0, UV = 0, PV = 0, SPV = 0;
  int tmode = TM_LGL3;
  int cmode = CM_FIRST_ORDER_SPLINE;
  int imode = IM_BASE;
  std::vector<double> bin_spacing {0.0, 1.0};
  std::vector<int> defs_per_bin {1};
  std::vector<StateFn> eqs, iqs, objs, integrals, param_integrals;
  PhaseIndexer idx;

  int cardinal_states() const { return tmode == TM_LGL7 ? 4 : (tmode == TM_LGL5 ? 3 : 2); }
  int num_defects() const {
    int n = 0;
    for (int d: defs_per_bin) n += d;
    return n;
  }
  bool blocked() const { return cmode == CM_BLOCK_CONSTANT; }
  void init_indexing();
  // appends this phase's functions to prob (tapes + function records) with the given offsets
  void transcribe(Problem& prob, int voff, int eoff, int ioff);
  // normalised cardinal spacing of every trajectory state, size numStates
  std::vector<double> state_spacing() const;
};

}  // namespace ab
