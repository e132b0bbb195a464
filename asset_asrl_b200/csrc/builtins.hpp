// Transcription functions the phase adds by itself (defects, mesh spacing, control splines, integrals),
// built directly on the scalar DAG so they lower to the same device tape as user expressions.
#pragma once
#include "expr.hpp"

namespace ab {

// Collocation coefficients; numeric literals are those of
// /root/reference/src/OptimalControl/LGLCoeffs.h:14-393 (decimal literals copied, not recomputed).
struct LGLTable {
  int CS;                 // cardinal states per segment: 2 (LGL3), 3 (LGL5), 4 (LGL7)
  double tau_card[4];     // CardinalSpacings
  double tau_int[3];      // InteriorSpacings
  double xi[3][4];        // Cardinal_XInterp_Weights
  double dxi[3][4];       // Cardinal_DXInterp_Weights
  double up[3][4];        // Cardinal_UPoly_Weights
  double xd[3][4];        // Cardinal_XDef_Weights
  double dxd[3][4];       // Cardinal_DXDef_Weights
  double idxd[3];         // Interior_DXDef_Weights
  double redw[4];         // Reduced_Integral_Weights
  double uzero[2][4];     // UZeroSpline_Weights
  double uone[2][4];      // UOneSpline_Weights
};
const LGLTable& lgl_table(int CS);

// Defect constraints.  Input layout [ (x,t,u)_0 .. (x,t,u)_{CS-1} , p ], output (CS-1)*XV
// (LGLDefects.h:57-122; TrapezoidalDefects.h:146-183).
ScalarFn make_lgl_defect(const Graph& g, int ode_root, int CS, int XV, int UV, int PV);
ScalarFn make_trap_defect(const Graph& g, int ode_root, int XV, int UV, int PV);
// Interior cardinal-node time ratios (MeshSpacingConstraints.h:107-193), IR = CS, OR = CS-2.
ScalarFn make_lgl_mesh_spacing(int CS);
// c*(t2-t0) - (t1-t0), c = per-application constant 0 (MeshSpacingConstraints.h:8-103), linear.
ScalarFn make_single_mesh_spacing();
// Control derivative continuity across a segment boundary (LGLControlSplines.h:64-312).
ScalarFn make_control_spline(int CS, int UV, int order);
// h * sum_i W_i g([x_i ; p]) over the cardinal nodes (LGLIntegrals.h:9-73, TrapezoidalIntegrals.h:8-58).
ScalarFn make_lgl_integral(const Graph& g, int integrand_root, int CS, int xv, int pv);

}  // namespace ab
