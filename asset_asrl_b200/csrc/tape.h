// Device instruction tape: the compact scalar program every user expression is lowered to once.
// Shared between host lowering (expr.cpp) and the sm_100a interpreter (kernels.cu).
//
// A tape has two sections.  The *primal* section computes the function outputs f(x); the *adjoint*
// section (emitted by symbolic reverse accumulation on the host, expr.cpp) computes
// g = d/dx (lambda . f(x)).  The device interprets both sections on forward-mode dual numbers with one
// tangent direction per lane, so d f_k / d x_dir is the Jacobian column and d g_i / d x_dir the
// column of the multiplier-weighted Hessian.  This is the B200 replacement for the reference's
// expression templates with analytic 1st/2nd derivatives
// (/root/reference/src/VectorFunctions/CommonFunctions/*.h, NestedFunction.h:140-270).
#pragma once
#include <stdint.h>

namespace ab {

enum Op : uint16_t {
  OP_CONST = 0,  // v = imm
  OP_INPUT,      // v = x[a]           (tangent = [a == dir])
  OP_LAMBDA,     // v = lambda[a]      (multiplier; constant w.r.t. x)
  OP_APPCONST,   // v = appconst[a]    (per-application constant, e.g. SingleMeshSpacing's c)
  OP_ADD,
  OP_SUB,
  OP_MUL,
  OP_DIV,
  OP_NEG,
  OP_ADDC,  // v = a + imm
  OP_MULC,  // v = a * imm
  OP_SQUARE,
  OP_INV,
  OP_SQRT,
  OP_EXP,
  OP_LOG,
  OP_SIN,
  OP_COS,
  OP_TAN,
  OP_ASIN,
  OP_ACOS,
  OP_ATAN,
  OP_SINH,
  OP_COSH,
  OP_TANH,
  OP_ASINH,
  OP_ACOSH,
  OP_ATANH,
  OP_ABS,
  OP_SIGN,
  OP_POWC,   // v = a ^ imm
  OP_ATAN2,  // v = atan2(a, b)
  OP_LT,     // v = (a <  b) ? 1 : 0   (zero derivative)
  OP_LE,     // v = (a <= b) ? 1 : 0
  OP_SELECT, // v = (a != 0) ? b : c   (c index is stored in the imm bits)
  OP_FMA,    // v = a * b + c          (c index is stored in the imm bits)
  OP_OUT_F,  // f[dst] = value(a), J[dst][dir] = tangent(a)
  OP_OUT_G,  // g[dst] = value(a), H[dst][dir] = tangent(a)
  // Interpolation tables (InterpTable1D.h / InterpTable2D.h of the reference).  The table data of a function sit in one
  // pool of doubles (SDag::tabpool); `imm` is an offset into it.  Both have zero derivative: an interpolant is
  // differentiated through its local coordinate, not through the choice of the interval.
  OP_TABIDX,   // v = index of the interval of axis `imm` that holds a: pool[imm] = n, pool[imm+1] = evenly spaced ? 1 : 0,
               //     pool[imm+2 ..] = n ascending coordinates; InterpTable1D::get_telem (InterpTable1D.h:180-197)
  OP_TABLOAD,  // v = pool[imm + (int) a]
  OP_COUNT
};

// 16 bytes, so one LDS.128 / LDG.128 fetches a whole instruction.
struct Instr {
  uint16_t op;
  uint16_t dst;  // destination slot (or output row for OP_OUT_*)
  uint16_t a;    // slot, or input / lambda / appconst index for the leaf ops
  uint16_t b;
  union {
    double imm;
    uint64_t c;  // third operand slot for OP_SELECT / OP_FMA
  };
};
static_assert(sizeof(Instr) == 16, "Instr must be 16 bytes");

}  // namespace ab
