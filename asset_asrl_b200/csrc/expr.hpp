// Host-side expression graph -> scalar DAG -> device tape.
//
// The reference builds C++ expression-template types behind type erasure and has no walkable graph
// (/root/reference/src/VectorFunctions/VectorFunctionTypeErasure/GenericFunction.h:87-405); this
// build records its own DAG at construction time and lowers it once to the tape of tape.h.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "tape.h"

namespace ab {

// ---------------------------------------------------------------------------------------------
// Vector-level expression nodes (what the Python / C++ user composes).
// ---------------------------------------------------------------------------------------------
enum VKind : int {
  VK_ARGS = 0,      // Arguments(ir): identity
  VK_SEGMENT,       // child[start : start+size]          ip = {start, size}
  VK_ELEMENTS,      // child[ip...] gather
  VK_CONSTANT,      // constant vector d, input size ir
  VK_STACK,         // concatenate children outputs
  VK_SUM,           // sum of children (same orows)
  VK_DIFF,          // ch0 - ch1
  VK_SCALED,        // ch0 * d[0]
  VK_ROWSCALED,     // ch0 .* d   (per-output scale)
  VK_ADDVEC,        // ch0 + d
  VK_VECMINUS,      // d - ch0
  VK_UNARY,         // cwise scalar op ip[0] (an ab::Op) on every output of ch0, d[0] = power for POWC
  VK_NORM,          // ||ch0||^p      ip = {p}
  VK_NORMALIZED,    // ch0 / ||ch0||^p ip = {p}
  VK_DOT,           // ch0 . ch1
  VK_CROSS,         // ch0 x ch1 (3-vectors)
  VK_CWISEPROD,     // ch0 .* ch1
  VK_CWISEQUOT,     // ch0 ./ ch1
  VK_VSPROD,        // vector ch0 * scalar ch1
  VK_VSDIV,         // vector ch0 / scalar ch1
  VK_CWISESUM,      // sum of the elements of ch0
  VK_ATAN2,         // atan2(ch0, ch1), scalars
  VK_NESTED,        // ch0(ch1(x)): outer evaluated at inner
  VK_IFELSE,        // ch0 (condition, scalar 0/1) ? ch1 : ch2
  VK_COMPARE,       // ip[0] in {0:<,1:<=,2:>,3:>=}, ch0 vs ch1 (scalars) -> 0/1
  VK_BOOL,          // ip[0] in {0:and,1:or}
  VK_IOSCALED,      // out_scale .* ch0(in_scale .* x)    d = [in_scales(ir) , out_scales(orows)]
  VK_APPCONST,      // per-application constant vector of size orows, ip = {first index}
  VK_MATVEC,        // matrix(d, orows x ch.orows, row-major) * ch0
  VK_QUATPROD,      // Hamilton product ch0 (x) ch1 of two quaternions stored vector-first (x, y, z, w)
  VK_INTERP1D,      // InterpTable1D(ch0 = t): ip = {n, vlen, kind (0 cubic, 1 linear), even}, d = [ts(n), vs(vlen x n), dvs(vlen x n)]
  VK_INTERP2D,      // InterpTable2D(ch0 = (x, y)): ip = {nx, ny, kind, xeven, yeven}, d = [xs, ys, z, dz/dx, dz/dy, d2z/dxdy (ny x nx each)]
  VK_MATPROD,       // matrix product A(ch0) * B(ch1): ip = {r, k, c, A row-major?, B row-major?}; output r x c column-major
  VK_MATINV,        // inverse of the n x n matrix ch0: ip = {n, row-major?}; output in the same storage order
  VK_INTERPND,      // InterpTable3D / 4D (ch0 = the D coordinates): ip = {D, kind, n_0.., even_0..}, d = [axis_0 .. axis_{D-1},
                    //   2^D arrays (cubic; array b holds the derivative along every axis whose bit is set in b) or one (linear),
                    //   each of shape (n_0, ..., n_{D-1}) in C order]
};

struct VNode {
  int kind = 0;
  int ir = 0;     // input rows
  int orows = 0;  // output rows
  std::vector<int> ch;
  std::vector<double> d;
  std::vector<int> ip;
  // Type-level linearity exactly as the reference's C++ types would report it (SURVEY.md section 8a,
  // note L).  `tlin` is the concrete type's static IsLinearFunction; `raw` says the value is NOT
  // wrapped in a GenericFunction (only Arguments/Segment/Element or C++-built helpers), i.e. a parent
  // type sees `tlin` rather than GenericFunction's `false` (ComputableBase.h:90).
  bool tlin = false;
  bool raw = false;
};

struct Graph {
  std::vector<VNode> nodes;
  int add(const VNode& n) {
    nodes.push_back(n);
    return (int) nodes.size() - 1;
  }
  const VNode& at(int i) const { return nodes.at(i); }
  // linearity a parent type observes for child i
  bool seen_lin(int i) const { return nodes[i].raw && nodes[i].tlin; }
  std::string describe(int root) const;  // s-expression consumed by the CPU oracle (tests only)
};

// ---------------------------------------------------------------------------------------------
// Scalar DAG with hash-consing, constant folding and light algebraic simplification.
// ---------------------------------------------------------------------------------------------
struct SNode {
  uint16_t op;
  int a, b, c;
  double imm;
};

struct SDag {
  std::vector<SNode> n;
  struct Key {
    uint16_t op;
    int a, b, c;
    uint64_t imm;
    bool operator==(const Key& o) const { return op == o.op && a == o.a && b == o.b && c == o.c && imm == o.imm; }
  };
  struct KeyHash {
    size_t operator()(const Key& k) const {
      uint64_t h = k.op * 0x9E3779B97F4A7C15ull;
      h ^= (uint64_t) (uint32_t) k.a + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
      h ^= (uint64_t) (uint32_t) k.b + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
      h ^= (uint64_t) (uint32_t) k.c + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
      h ^= k.imm + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
      return (size_t) h;
    }
  };
  std::unordered_map<Key, int, KeyHash> cons;

  // interpolation-table data referenced by OP_TABIDX / OP_TABLOAD nodes (offsets in `imm`); copies of a DAG carry it along
  struct TabEntry {
    size_t off, len;
  };
  std::vector<double> tabpool;
  std::vector<TabEntry> tabs;
  size_t tab_add(const double* p, size_t n);  // returns the offset; identical data are stored once
  const TabEntry& tab_find(size_t off) const;  // the entry an offset falls in
  // pool offset of a table node of `src` after its data have been copied into this DAG
  double tab_import(const SDag& src, double imm);
  int tabidx(int a, size_t axis_off);
  int tabload(int idx, size_t base_off);

  int raw(uint16_t op, int a, int b, int c, double imm);
  bool is_const(int i) const { return n[i].op == OP_CONST; }
  double cval(int i) const { return n[i].imm; }

  int constant(double v) { return raw(OP_CONST, -1, -1, -1, v); }
  int input(int k) { return raw(OP_INPUT, k, -1, -1, 0.0); }
  int lambda(int k) { return raw(OP_LAMBDA, k, -1, -1, 0.0); }
  int appconst(int k) { return raw(OP_APPCONST, k, -1, -1, 0.0); }
  int add(int a, int b);
  int sub(int a, int b);
  int mul(int a, int b);
  int div(int a, int b);
  int neg(int a);
  int addc(int a, double c);
  int mulc(int a, double c);
  int unary(uint16_t op, int a, double imm = 0.0);
  int atan2(int y, int x);
  int cmp(uint16_t op, int a, int b);
  int select(int c, int a, int b);
};

// A lowered function: outputs and (optionally) the adjoint gradient of lambda . f.
struct ScalarFn {
  int nin = 0, nout = 0;
  SDag dag;
  std::vector<int> outs;   // nout scalar node ids
  std::vector<int> grads;  // nin scalar node ids (d(lambda.f)/dx_i), empty until make_gradient()
  std::vector<int> jac;    // nout*nin node ids dF_k/dx_j (row-major), empty until make_second_order()
  std::vector<int> hess;   // nin*nin node ids dg_i/dx_j (row-major; symmetric, the emitter uses j <= i)
  void make_gradient();
  void make_second_order();  // deriv.cpp: symbolic sparse forward sweep over outs and grads
};

// Device tape + bookkeeping.
struct Tape {
  std::shared_ptr<ScalarFn> fn;  // the lowered function the device programs are generated from (plan build)
  std::vector<Instr> code;
  int n_primal = 0;       // instructions in the primal section (values + Jacobian)
  int slots_primal = 0;   // slots live in the primal section
  int slots_total = 0;    // slots for the whole tape
  int nin = 0, nout = 0, nappconst = 0;
  bool has_adjoint = false;
  // what the tape was lowered from, as text: "expr <s-expression>" or a built-in transcription function with its
  // sizes.  Introspection only (tests rebuild the same function in the CPU oracle from it).
  std::string desc;
};

// ---------------------------------------------------------------------------------------------
// Scalar program: what the device actually runs (program.cpp / kernels.cu).  One thread evaluates one function
// application; the 32 lanes of a warp are 32 applications (mesh segments) executing the same instruction, so decode is
// amortised over the warp and every value lives in a lane-private shared-memory slot.  Instructions are grouped in
// bundles of four mutually independent operations of one kind (instruction-level parallelism for the interpreter).
// ---------------------------------------------------------------------------------------------
enum PKind : uint32_t { PK_FMA = 0, PK_LOAD = 1, PK_OUT = 2, PK_SPECIAL = 3, PK_END = 4 };
enum POut : uint32_t { PO_F = 0, PO_J = 1, PO_G = 2, PO_H = 3, PO_TABLE = 4 /* index into a destination table */ };
enum PLoad : uint32_t { PL_X = 0, PL_LAMBDA = 1, PL_APPCONST = 2 };
// Operand word: bit0 = 1 lane-private slot / 0 shared constant, bit1 = negate, bits[31:3] = index (unlinked) or
// byte offset (linked for a block size).  The destination word of the first instruction of a bundle carries the
// bundle kind in bits[31:28] and the number of live instructions in bits[27:24].
struct PInstr {
  uint32_t d, a, b, c;
};
static_assert(sizeof(PInstr) == 16, "PInstr must be 16 bytes");
constexpr uint32_t P_SLOT = 1u, P_NEG = 2u;
// the block streams the program through two shared-memory buffers of this many bundles (cp.async double buffering)
constexpr uint32_t PROG_CHUNK_BUNDLES = 64u;

struct Program {
  std::vector<PInstr> code;  // 4 instructions per bundle, PK_END bundle last
  std::vector<double> consts;
  std::vector<double> tables;  // the function's interpolation-table pool (SDag::tabpool); empty for most programs
  int nslots = 0;
  int nin = 0, nout = 0, nappconst = 0;
  int n_ops = 0;       // live instructions (for reporting)
  int n_fma = 0, n_special = 0, n_load = 0, n_out = 0;
  int bundles() const { return (int) code.size() / 4; }
};

// what a program computes; mirrors FnMode in device_types.h
struct ProgramSpec {
  bool want_f = true, want_j = false, want_g = false, want_h = false;
  bool is_objective = false;  // objectives scatter no Jacobian triplets
  bool linear = false;        // type-level linear: no Hessian triplets exist
};
// Triplet offset of (row, col i) inside one application's block, DenseFunctionBase::getKKTSpace order.
int triplet_offset_h(int i, int j, int ir, int orows, bool lay_h, bool lay_j);
int triplet_offset_j(int i, int k, int ir, int orows, bool lay_h, bool lay_j);
Program emit_program(ScalarFn& f, const ProgramSpec& spec);
// The same scheduler over an explicit output list (in the order given); used for fused programs whose inputs and
// outputs are entries of address tables rather than one function's argument / triplet numbers.
struct ProgOut {
  int node;
  uint32_t kind, index;
};
Program emit_program_outs(ScalarFn& f, const std::vector<ProgOut>& outs);
struct LinkedProgram {
  std::vector<PInstr> code;
  int nthreads = 0, nslots = 0;
  unsigned smem_bytes = 0;
};
// Affine index maps of one function over its bulk application range [a0, a1): for those applications every gather /
// scatter index is base + (app - a0) * stride, so the kernel computes addresses instead of loading index arrays.
// Applications outside the range (mesh ends) use the explicit maps.  |stride| < 2^23.
struct AffineMaps {
  int a0 = 0, a1 = 0;
  std::vector<int> vbase, vstride;  // per input: global variable index
  std::vector<int> cbase, cstride;  // per output: multiplier (constraint row) index; empty for objectives
  std::vector<int> fbase, fstride;  // per output: f_dest (>= 0 direct, < 0 ~stage); empty for objectives
  std::vector<int> kbase, kstride;  // per triplet: kkt_dest
};
LinkedProgram link_program(const Program& P, unsigned smem_budget, const AffineMaps* am = nullptr);

// interval index of t on an axis stored as [n, even, coordinates...] (host copy of what the device does for OP_TABIDX)
int table_interval(const double* axis, double t);

std::vector<int> lower(const Graph& g, int root, const std::vector<int>& in, SDag& dag);
ScalarFn lower_function(const Graph& g, int root);
Tape emit_tape(const ScalarFn& f, bool with_adjoint);

}  // namespace ab
