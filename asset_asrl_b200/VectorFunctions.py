"""Expression front end with the reference's Python surface (``vf.Arguments``, operators, ``vf.sin`` ...).

Mirrors the names bound in /root/reference/src/VectorFunctions/DenseFunctionBase.h:1643-2263 and
re-exported by /root/reference/asset_asrl/VectorFunctions/__init__.py:6-56.  Every object is a node id in a
C++ expression graph (asset_asrl_b200/csrc/expr.hpp); nothing is evaluated in Python.
"""
import numbers

import numpy as np

from . import _lib
from ._lib import lib

# ab::VKind (csrc/expr.hpp)
(VK_ARGS, VK_SEGMENT, VK_ELEMENTS, VK_CONSTANT, VK_STACK, VK_SUM, VK_DIFF, VK_SCALED, VK_ROWSCALED, VK_ADDVEC,
 VK_VECMINUS, VK_UNARY, VK_NORM, VK_NORMALIZED, VK_DOT, VK_CROSS, VK_CWISEPROD, VK_CWISEQUOT, VK_VSPROD, VK_VSDIV,
 VK_CWISESUM, VK_ATAN2, VK_NESTED, VK_IFELSE, VK_COMPARE, VK_BOOL, VK_IOSCALED, VK_APPCONST, VK_MATVEC, VK_QUATPROD,
 VK_INTERP1D, VK_INTERP2D, VK_MATPROD, VK_MATINV, VK_INTERPND) = range(35)

# ab::Op (csrc/tape.h)
OPS = ["CONST", "INPUT", "LAMBDA", "APPCONST", "ADD", "SUB", "MUL", "DIV", "NEG", "ADDC", "MULC", "SQUARE", "INV", "SQRT",
       "EXP", "LOG", "SIN", "COS", "TAN", "ASIN", "ACOS", "ATAN", "SINH", "COSH", "TANH", "ASINH", "ACOSH", "ATANH", "ABS",
       "SIGN", "POWC", "ATAN2", "LT", "LE", "SELECT", "FMA", "OUT_F", "OUT_G", "TABIDX", "TABLOAD"]
OP = {n: i for i, n in enumerate(OPS)}


class _Graph:
    """Process-wide expression graph (the C++ arena every function lives in)."""

    def __init__(self):
        self.h = lib.ab_graph_create()
        if not self.h:
            raise MemoryError("ab_graph_create failed")

    def node(self, kind, ir, orows, ch=(), d=(), ip=(), tlin=-1, raw=0):
        cha, chp = _lib.iarr(list(ch))
        da, dp = _lib.darr(list(d))
        ipa, ipp = _lib.iarr(list(ip))
        return _lib.check(lib.ab_graph_node(self.h, kind, ir, orows, chp, len(cha), dp, len(da), ipp, len(ipa), tlin, raw))


GRAPH = _Graph()


def _is_num(x):
    return isinstance(x, numbers.Real) and not isinstance(x, bool)


def _vec(x):
    return np.atleast_1d(np.asarray(x, dtype=np.float64)).ravel()


class VectorFunction:
    """Generic vector function f: R^IRows -> R^ORows (the reference's GenericFunction<-1,-1>)."""

    __array_priority__ = 1000

    def __init__(self, node, ir=None, orows=None):
        if isinstance(node, VectorFunction):  # allow class MyFun(vf.VectorFunction): super().__init__(expr)
            node, ir, orows = node.node, node.ir, node.orows
        self.node = int(node)
        self.ir = ir
        self.orows = orows

    # ---- sizing / meta
    def IRows(self):
        return self.ir

    def ORows(self):
        return self.orows

    def is_linear(self):
        import ctypes as C
        t = C.c_int()
        _lib.check(lib.ab_graph_node_info(GRAPH.h, self.node, None, None, C.byref(t), None))
        return bool(t.value)

    def describe(self):
        n = lib.ab_graph_describe(GRAPH.h, self.node, None, 0)
        import ctypes as C
        buf = C.create_string_buffer(int(n) + 1)
        lib.ab_graph_describe(GRAPH.h, self.node, buf, n + 1)
        return buf.value.decode()

    def vf(self):
        return VectorFunction(self.node, self.ir, self.orows)

    def sf(self):
        if self.orows != 1:
            raise ValueError("sf(): function is not scalar valued")
        return ScalarFunction(self.node, self.ir, 1)

    # ---- construction helpers
    def _new(self, kind, orows, ch, d=(), ip=(), raw=0):
        cls = ScalarFunction if orows == 1 else VectorFunction
        for c in ch:
            if c.ir != self.ir:
                raise ValueError(f"Input Size of Func1 (IRows = {self.ir}) does not match Input Size of Func2 (IRows = {c.ir}).")
        n = GRAPH.node(kind, self.ir, orows, [c.node for c in ch], d, ip, -1, raw)
        return cls(n, self.ir, orows)

    def _raw(self):
        import ctypes as C
        r = C.c_int()
        _lib.check(lib.ab_graph_node_info(GRAPH.h, self.node, None, None, None, C.byref(r)))
        return r.value

    def _coerce(self, other):
        """number / vector -> None (handled by caller); function -> function"""
        if isinstance(other, VectorFunction):
            return other
        return None

    # ---- element access
    def segment(self, start, size):
        if start < 0 or size < 0 or start + size > self.orows:
            raise ValueError("Segment index out of bounds.")
        return self._new(VK_SEGMENT, size, [self], ip=[start, size], raw=self._raw())

    def head(self, size):
        return self.segment(0, size)

    def tail(self, size):
        return self.segment(self.orows - size, size)

    def coeff(self, i):
        return self.segment(i, 1)

    def head2(self):
        return self.head(2)

    def head3(self):
        return self.head(3)

    def tail2(self):
        return self.tail(2)

    def tail3(self):
        return self.tail(3)

    def segment2(self, s):
        return self.segment(s, 2)

    def segment3(self, s):
        return self.segment(s, 3)

    head_2, head_3, tail_2, tail_3, segment_2, segment_3 = head2, head3, tail2, tail3, segment2, segment3

    def tolist(self, spec=None):
        if spec is None:
            return [self.coeff(i) for i in range(self.orows)]
        return [self.segment(s, n) for (s, n) in spec]

    def __getitem__(self, i):
        if isinstance(i, slice):
            idx = list(range(*i.indices(self.orows)))
            if not idx:
                raise ValueError("empty slice")
            if idx == list(range(idx[0], idx[0] + len(idx))):
                return self.segment(idx[0], len(idx))
            return self._new(VK_ELEMENTS, len(idx), [self], ip=idx, raw=self._raw())
        if isinstance(i, (list, tuple, np.ndarray)):
            idx = [int(k) for k in i]
            return self._new(VK_ELEMENTS, len(idx), [self], ip=idx, raw=self._raw())
        i = int(i)
        if i < 0:
            i += self.orows
        return self.coeff(i)

    def __len__(self):
        return self.orows

    def __iter__(self):
        return iter(self.tolist())

    # ---- arithmetic
    def __neg__(self):
        return self._new(VK_SCALED, self.orows, [self], d=[-1.0])

    def __add__(self, o):
        f = self._coerce(o)
        if f is not None:
            if f.orows != self.orows:
                raise ValueError("Output sizes do not match in sum")
            return self._new(VK_SUM, self.orows, [self, f])
        v = _vec(o)
        if v.size == 1 and self.orows > 1:
            v = np.full(self.orows, v[0])
        if v.size != self.orows:
            raise ValueError("vector size does not match function output size")
        return self._new(VK_ADDVEC, self.orows, [self], d=v)

    __radd__ = __add__

    def __sub__(self, o):
        f = self._coerce(o)
        if f is not None:
            if f.orows != self.orows:
                raise ValueError("Output sizes do not match in difference")
            return self._new(VK_DIFF, self.orows, [self, f])
        return self.__add__(-_vec(o))

    def __rsub__(self, o):
        v = _vec(o)
        if v.size == 1 and self.orows > 1:
            v = np.full(self.orows, v[0])
        if v.size != self.orows:
            raise ValueError("vector size does not match function output size")
        return self._new(VK_VECMINUS, self.orows, [self], d=v)

    def __mul__(self, o):
        f = self._coerce(o)
        if f is not None:
            if f.orows == 1 and self.orows == 1:
                return self._new(VK_CWISEPROD, 1, [self, f])
            if f.orows == 1:
                return self._new(VK_VSPROD, self.orows, [self, f])
            if self.orows == 1:
                return f._new(VK_VSPROD, f.orows, [f, self])
            raise ValueError("vector * vector is ambiguous: use dot / cross / cwiseProduct")
        if _is_num(o):
            return self._new(VK_SCALED, self.orows, [self], d=[float(o)])
        v = _vec(o)
        if self.orows == 1:  # scalar function times constant vector
            c = self._new(VK_CONSTANT, v.size, [], d=v)
            return c._new(VK_VSPROD, v.size, [c, self])
        if v.size != self.orows:
            raise ValueError("vector size does not match function output size")
        return self._new(VK_ROWSCALED, self.orows, [self], d=v)

    __rmul__ = __mul__

    def __truediv__(self, o):
        f = self._coerce(o)
        if f is not None:
            if f.orows != 1:
                raise ValueError("can only divide by a scalar function")
            if self.orows == 1:
                return self._new(VK_CWISEQUOT, 1, [self, f])
            return self._new(VK_VSDIV, self.orows, [self, f])
        if _is_num(o):
            return self._new(VK_SCALED, self.orows, [self], d=[1.0 / float(o)])
        v = _vec(o)
        return self._new(VK_ROWSCALED, self.orows, [self], d=1.0 / v)

    def __rtruediv__(self, o):
        if self.orows != 1:
            raise ValueError("can only divide by a scalar function")
        inv = self._unary("INV")
        return inv.__mul__(o)

    def __pow__(self, p):
        return self.pow(p)

    def __abs__(self):
        return self._unary("ABS")

    # ---- cwise math (scalar functions in the reference; applied elementwise here)
    def _unary(self, name, imm=None):
        d = [] if imm is None else [float(imm)]
        return self._new(VK_UNARY, self.orows, [self], d=d, ip=[OP[name]])

    def sin(self):
        return self._unary("SIN")

    def cos(self):
        return self._unary("COS")

    def tan(self):
        return self._unary("TAN")

    def sqrt(self):
        return self._unary("SQRT")

    def exp(self):
        return self._unary("EXP")

    def log(self):
        return self._unary("LOG")

    def squared(self):
        return self._unary("SQUARE")

    def inverse(self):
        return self._unary("INV")

    def arcsin(self):
        return self._unary("ASIN")

    def arccos(self):
        return self._unary("ACOS")

    def arctan(self):
        return self._unary("ATAN")

    def sinh(self):
        return self._unary("SINH")

    def cosh(self):
        return self._unary("COSH")

    def tanh(self):
        return self._unary("TANH")

    def arcsinh(self):
        return self._unary("ASINH")

    def arccosh(self):
        return self._unary("ACOSH")

    def arctanh(self):
        return self._unary("ATANH")

    def sign(self):
        return self._unary("SIGN")

    def pow(self, p):
        if not _is_num(p):
            raise TypeError("pow: exponent must be a number")
        return self._unary("POWC", float(p))

    # ---- norms
    def _norm(self, p):
        return self._new(VK_NORM, 1, [self], ip=[p])

    def norm(self):
        return self._norm(1)

    def squared_norm(self):
        return self._norm(2)

    def cubed_norm(self):
        return self._norm(3)

    def inverse_norm(self):
        return self._norm(-1)

    def inverse_squared_norm(self):
        return self._norm(-2)

    def inverse_cubed_norm(self):
        return self._norm(-3)

    def inverse_four_norm(self):
        return self._norm(-4)

    def _normalized(self, p):
        return self._new(VK_NORMALIZED, self.orows, [self], ip=[p])

    def normalized(self):
        return self._normalized(1)

    def normalized_power2(self):
        return self._normalized(2)

    def normalized_power3(self, shift=None, scale=None):
        f = self
        if shift is not None:
            f = self + _vec(shift)
        r = f._normalized(3)
        return r if scale is None else r * float(scale)

    def normalized_power4(self):
        return self._normalized(4)

    def normalized_power5(self):
        return self._normalized(5)

    def sum(self):
        return self._new(VK_CWISESUM, 1, [self])

    # ---- products
    def _other(self, o):
        if isinstance(o, VectorFunction):
            return o
        v = _vec(o)
        return self._new(VK_CONSTANT, v.size, [], d=v, raw=1)

    def dot(self, o):
        o = self._other(o)
        if o.orows != self.orows:
            raise ValueError("dot: size mismatch")
        return self._new(VK_DOT, 1, [self, o])

    def cross(self, o):
        o = self._other(o)
        if o.orows != 3 or self.orows != 3:
            raise ValueError("cross: both operands must have 3 outputs")
        return self._new(VK_CROSS, 3, [self, o])

    def quatProduct(self, o):
        """Hamilton product of two quaternions stored vector-first (FunctionQuatProduct, VectorProducts.h:28-30,229-236)."""
        o = self._other(o)
        if o.orows != 4 or self.orows != 4:
            raise ValueError("quatProduct: both operands must have 4 outputs")
        return self._new(VK_QUATPROD, 4, [self, o])

    def padded_lower(self, n):
        """The outputs followed by n zeros (PaddedOutput, Padded.h:24-60; type-level linear when the function is)."""
        return self.padded(0, n)

    def padded_upper(self, n):
        return self.padded(n, 0)

    def padded(self, upper, lower):
        upper, lower = int(upper), int(lower)
        if upper < 0 or lower < 0:
            raise ValueError("padded: negative padding")
        parts = []
        if upper:
            parts.append(self._new(VK_CONSTANT, upper, [], d=np.zeros(upper), raw=1))
        parts.append(self)
        if lower:
            parts.append(self._new(VK_CONSTANT, lower, [], d=np.zeros(lower), raw=1))
        if len(parts) == 1:
            return self
        return self._new(VK_STACK, self.orows + upper + lower, parts)

    def cwiseProduct(self, o):
        o = self._other(o)
        if o.orows != self.orows:
            raise ValueError("cwiseProduct: size mismatch")
        return self._new(VK_CWISEPROD, self.orows, [self, o])

    def cwiseQuotient(self, o):
        o = self._other(o)
        if o.orows != self.orows:
            raise ValueError("cwiseQuotient: size mismatch")
        return self._new(VK_CWISEQUOT, self.orows, [self, o])

    # ---- composition
    def eval(self, inner, *rest):
        """self(inner(x)); ``eval(ir, indices)`` selects elements of a larger argument vector."""
        if rest:
            ir, idx = int(inner), [int(k) for k in rest[0]]
            inner = Arguments(ir)[idx]
        if not isinstance(inner, VectorFunction):
            raise TypeError("eval expects a VectorFunction")
        if inner.orows != self.ir:
            raise ValueError(f"eval: inner function has {inner.orows} outputs but outer expects {self.ir} inputs")
        cls = ScalarFunction if self.orows == 1 else VectorFunction
        n = GRAPH.node(VK_NESTED, inner.ir, self.orows, [self.node, inner.node], (), (), -1, 0)
        return cls(n, inner.ir, self.orows)

    def __call__(self, inner):
        return self.eval(inner)

    # ---- comparisons (scalar) -> conditional objects for ifelse
    def _cmp(self, o, code):
        if self.orows != 1:
            raise ValueError("comparison needs scalar functions")
        if not isinstance(o, VectorFunction):
            o = self._new(VK_CONSTANT, 1, [], d=[float(o)], raw=1)
        return self._new(VK_COMPARE, 1, [self, o], ip=[code])

    def __lt__(self, o):
        return self._cmp(o, 0)

    def __le__(self, o):
        return self._cmp(o, 1)

    def __gt__(self, o):
        return self._cmp(o, 2)

    def __ge__(self, o):
        return self._cmp(o, 3)

    def __and__(self, o):
        return self._new(VK_BOOL, 1, [self, o], ip=[0])

    def __or__(self, o):
        return self._new(VK_BOOL, 1, [self, o], ip=[1])


class ScalarFunction(VectorFunction):
    """f: R^IRows -> R (the reference's GenericFunction<-1,1>)."""

    def __init__(self, node, ir=None, orows=1):
        super().__init__(node, ir, orows)
        if self.orows != 1:
            raise ValueError("ScalarFunction needs a single output")


class Arguments(VectorFunction):
    """Arguments(n): the identity on R^n, root of every expression (Segment.h / Arguments)."""

    def __init__(self, n):
        n = int(n)
        if n <= 0:
            raise ValueError("Arguments size must be positive")
        node = GRAPH.node(VK_ARGS, n, n, raw=1)
        super().__init__(node, n, n)


def ConstantVector(ir, values):
    v = _vec(values)
    return VectorFunction(GRAPH.node(VK_CONSTANT, int(ir), v.size, d=v, raw=1), int(ir), v.size)


def ConstantScalar(ir, value):
    return ScalarFunction(GRAPH.node(VK_CONSTANT, int(ir), 1, d=[float(value)], raw=1), int(ir), 1)


def _as_fun(f):
    if not isinstance(f, VectorFunction):
        raise TypeError("expected a VectorFunction")
    return f


def stack(*fs):
    """vf.stack([f1, f2, ...]) (BulkOperationsBuild.cpp:6-15)."""
    if len(fs) == 1 and isinstance(fs[0], (list, tuple)):
        fs = fs[0]
    fs = [_as_fun(f) for f in fs]
    if not fs:
        raise ValueError("stack of nothing")
    return fs[0]._new(VK_STACK, builtins_sum(f.orows for f in fs), fs)


def builtins_sum(it):
    t = 0
    for v in it:
        t += v
    return t


stack_scalar = stack
Stack = stack
StackScalar = stack


def sum(*fs):  # noqa: A001 - the reference exports vf.sum
    if len(fs) == 1 and isinstance(fs[0], (list, tuple)):
        fs = fs[0]
    fs = [_as_fun(f) for f in fs]
    for f in fs:
        if f.orows != fs[0].orows:
            raise ValueError("sum: output sizes differ")
    return fs[0]._new(VK_SUM, fs[0].orows, fs)


Sum = sum
SumScalar = sum


def SumElems(f):
    return f.sum()


def _u(name):
    def fn(f):
        return _as_fun(f)._unary(name)
    fn.__name__ = name.lower()
    return fn


sin, cos, tan = _u("SIN"), _u("COS"), _u("TAN")
arcsin, arccos, arctan = _u("ASIN"), _u("ACOS"), _u("ATAN")
sinh, cosh, tanh = _u("SINH"), _u("COSH"), _u("TANH")
arcsinh, arccosh, arctanh = _u("ASINH"), _u("ACOSH"), _u("ATANH")
sqrt, exp, log, squared, sign = _u("SQRT"), _u("EXP"), _u("LOG"), _u("SQUARE"), _u("SIGN")
abs = _u("ABS")  # noqa: A001


def pow(f, p):  # noqa: A001
    return _as_fun(f).pow(p)


def arctan2(y, x):
    y, x = _as_fun(y), _as_fun(x)
    return y._new(VK_ATAN2, 1, [y, x])


def norm(f):
    return f.norm()


def inverse_norm(f):
    return f.inverse_norm()


def normalize(f):
    return f.normalized()


def dot(a, b):
    return (a if isinstance(a, VectorFunction) else b).dot(b if isinstance(a, VectorFunction) else a)


def cross(a, b):
    if isinstance(a, VectorFunction):
        return a.cross(b)
    return -(b.cross(a))


def quatProduct(a, b):
    """vf.quatProduct(q1, q2) (FreeFunctionsBuild.cpp:371-378)."""
    if isinstance(a, VectorFunction):
        return a.quatProduct(b)
    return b._other(a).quatProduct(b)


def quatRotate(q, v):
    """vf.quatRotate(q, v): the vector part of q (x) (v, 0) (x) q^-1 for a unit quaternion q stored vector-first
    (QuatRotation::Definition, FreeFunctionsBuild.cpp:310-321)."""
    q = _as_fun(q)
    v = v if isinstance(v, VectorFunction) else q._other(v)
    if q.orows != 4 or v.orows != 3:
        raise ValueError("quatRotate: needs a quaternion and a 3-vector")
    qinv = stack(q.head(3) * (-1.0), q.coeff(3))
    return q.quatProduct(v.padded_lower(1)).quatProduct(qinv).head(3)


def cwiseProduct(a, b):
    return (a if isinstance(a, VectorFunction) else b).cwiseProduct(b if isinstance(a, VectorFunction) else a)


def cwiseQuotient(a, b):
    if isinstance(a, VectorFunction):
        return a.cwiseQuotient(b)
    c = b._other(a)
    return c.cwiseQuotient(b)


def doublecross(a, b, c):
    return cross(a, cross(b, c))


def ifelse(cond, a, b):
    """vf.ifelse(condition, f_true, f_false) (Conditional.h:19-235)."""
    a = a if isinstance(a, VectorFunction) else cond._other(a)
    b = b if isinstance(b, VectorFunction) else cond._other(b)
    if a.orows != b.orows:
        raise ValueError("ifelse: branch sizes differ")
    return cond._new(VK_IFELSE, a.orows, [cond, a, b])


class _MatrixView:
    """A vector function read as a matrix (MatrixFunctionView, CommonFunctions/MatrixFunction.h:9-19): ``ColMatrix`` stores
    its columns one after another, ``RowMatrix`` its rows.  Python surface of MatrixFunctionBuild.cpp:17-215."""
    rowmajor = 0

    def __init__(self, f, rows=None, cols=None):
        if isinstance(f, (list, tuple)):  # a list of column (row) functions
            if len(f) == 0:
                raise ValueError("List must contain at least one function.")
            parts = [_as_fun(q) for q in f]
            for q in parts:
                if q.orows != parts[0].orows:
                    raise ValueError(("Row" if self.rowmajor else "Column") + " Functions must have same output size")
            n_other, n_list = parts[0].orows, len(parts)
            rows, cols = (n_list, n_other) if self.rowmajor else (n_other, n_list)
            f = stack(parts)
        f = _as_fun(f)
        if rows is None or cols is None or int(rows) * int(cols) != f.orows:
            raise ValueError("matrix view: rows * cols must equal the function's output size")
        self.f, self.MatrixRows, self.MatrixCols = f, int(rows), int(cols)

    def vf(self):
        return self.f

    def IRows(self):
        return self.f.ir

    def transpose(self):
        other = ColMatrix if self.rowmajor else RowMatrix
        return other(self.f, self.MatrixCols, self.MatrixRows)

    def inverse(self):
        if self.MatrixRows != self.MatrixCols:
            raise ValueError("Matrix must be square to be invertible")
        n = self.MatrixRows
        inv = self.f._new(VK_MATINV, n * n, [self.f], ip=[n, self.rowmajor])
        return type(self)(inv, n, n)

    def _prod(self, o):
        if self.MatrixCols != o.MatrixRows:
            raise ValueError("Invalid matrix product. Number of columns in matrix 1 does not match number of rows in matrix 2.")
        r, k, c = self.MatrixRows, self.MatrixCols, o.MatrixCols
        return self.f._new(VK_MATPROD, r * c, [self.f, o.f], ip=[r, k, c, self.rowmajor, o.rowmajor])

    def __mul__(self, o):
        if isinstance(o, _MatrixView):
            return ColMatrix(self._prod(o), self.MatrixRows, o.MatrixCols)  # products come back column-major
        if isinstance(o, VectorFunction):
            return self._prod(ColMatrix(o, o.orows, 1))  # a vector
        if _is_num(o):
            return type(self)(self.f * float(o), self.MatrixRows, self.MatrixCols)
        v = _vec(o)  # constant vector
        return self * self.f._new(VK_CONSTANT, v.size, [], d=v, raw=1)

    def __rmul__(self, o):
        if _is_num(o):
            return self * o
        return NotImplemented

    def __add__(self, o):
        if isinstance(o, _MatrixView):
            if (o.MatrixRows, o.MatrixCols) != (self.MatrixRows, self.MatrixCols):
                raise ValueError("Matrices must have the same dimensions to be added.")
            if o.rowmajor != self.rowmajor:
                raise TypeError("matrices of different storage order cannot be added")
            return type(self)(self.f + o.f, self.MatrixRows, self.MatrixCols)
        m = np.asarray(o, dtype=np.float64)
        if m.shape != (self.MatrixRows, self.MatrixCols):
            raise ValueError("Matrices must have the same dimensions to be added.")
        return type(self)(self.f + (m.ravel() if self.rowmajor else m.T.ravel()), self.MatrixRows, self.MatrixCols)

    __radd__ = __add__
    __array_ufunc__ = None


class ColMatrix(_MatrixView):
    rowmajor = 0


class RowMatrix(_MatrixView):
    rowmajor = 1


def matmul(M, f):
    """vf.matmul (MatrixFunctionBuild.cpp:218-271): matrix function x matrix function / vector function / constant vector,
    or a constant matrix x vector function (MatrixScaled)."""
    if isinstance(M, _MatrixView):
        return M * f
    M = np.asarray(M, dtype=np.float64)
    if M.ndim != 2 or M.shape[1] != f.orows:
        raise ValueError("matmul: shape mismatch")
    return f._new(VK_MATVEC, M.shape[0], [f], d=M.ravel())


def ParsedInput(f, varlocs, irr):
    """ParsedInput(f, varlocs, irr) (CommonFunctions/ParsedInput.h:6-27): f applied to the listed elements of a larger input."""
    return _as_fun(f).eval(int(irr), [int(v) for v in varlocs])


def ioscaled(f, in_scales, out_scales):
    """IOScaled wrapper (IOScaled.h:8-158): out_scales .* f(in_scales .* x)."""
    i, o = _vec(in_scales), _vec(out_scales)
    if i.size != f.ir or o.size != f.orows:
        raise ValueError("ioscaled: scale sizes do not match")
    return f._new(VK_IOSCALED, f.orows, [f], d=np.concatenate([i, o]))


def _stencil_derivs(cs, V, even):
    """d V / d c along the last axis by five-point stencils fitted to the actual spacing: central where two neighbours
    exist on both sides, one-sided at the ends (InterpTable1D::calc_derivs, InterpTable1D.h:133-178; an evenly spaced
    axis reuses its first central stencil, :151-153)."""
    n = cs.size
    out = np.zeros_like(V, dtype=np.float64)
    rhs = np.array([0.0, 1.0, 0.0, 0.0, 0.0])
    coeffs, hitcent = None, False
    for i in range(n):
        recalc = True
        if i + 2 <= n - 1 and i - 2 >= 0:
            if even and hitcent:
                recalc = False
            hitcent = True
            start = i - 2
        elif i < n - 1 - i:
            start = 0
        else:
            start = n - 5
        step = np.abs(cs[i + (1 if i < n - 1 else -1)] - cs[i])
        if recalc:
            tau = (cs[start:start + 5] - cs[i]) / step
            coeffs = np.linalg.solve(np.vander(tau, 5, increasing=True).T, rhs)
        out[..., i] = V[..., start:start + 5] @ (coeffs / step)
    return out


def _is_even(cs):
    return not (np.abs(cs - np.linspace(cs[0], cs[-1], cs.size)).max() > np.abs(cs[-1] - cs[0]) * 1.0e-12)


def _kind(kind):
    if kind in ("cubic", "Cubic"):
        return 0
    if kind in ("linear", "Linear"):
        return 1
    raise ValueError("Unrecognized interpolation type")


def _ascending(cs, what):
    if cs.size < 5:
        raise ValueError(f"{what} coordinates must be larger than 4")
    if np.any(np.diff(cs) < 0):
        raise ValueError(f"{what} Coordinates must be in ascending order")


def _interval(cs, even, t):
    """InterpTable*::get_*elems: truncating division on an evenly spaced axis, upper_bound otherwise; clamped."""
    if even:
        e = int((t - cs[0]) / (cs[1] - cs[0]))
    else:
        e = int(np.searchsorted(cs, t, side="right")) - 1
    return max(min(e, cs.size - 2), 0)


def _hermite_nd(axes, even, arrays, cubic, point):
    """Host evaluation of the tensor-product interpolant the device programs implement (value only): ``arrays[b]`` holds
    the derivative along every axis whose bit is set in b (cubic) or just the values (linear)."""
    D = len(axes)
    cell, weights = [], []
    for d in range(D):
        e = _interval(axes[d], even[d], float(point[d]))
        h = axes[d][e + 1] - axes[d][e]
        u = (float(point[d]) - axes[d][e]) / h
        cell.append(e)
        if cubic:
            u2, u3 = u * u, u * u * u
            weights.append((2 * u3 - 3 * u2 + 1, -2 * u3 + 3 * u2, (u3 - 2 * u2 + u) * h, (u3 - u2) * h))
        else:
            weights.append((1 - u, u))
    total = 0.0
    for fb in range(1 << D if cubic else 1):
        for cb in range(1 << D):
            idx = tuple(cell[d] + (cb >> d & 1) for d in range(D))
            w = 1.0
            for d in range(D):
                w *= weights[d][((fb >> d & 1) << 1) | (cb >> d & 1)]
            total += arrays[fb][idx] * w
    return total


class InterpTable1D:
    """vf.InterpTable1D(ts, Vs, axis=0, kind='cubic') (InterpTable1D.h:9-131): piecewise cubic-Hermite (slopes from
    five-point stencils) or linear interpolation of vector data over a scalar coordinate.  ``tab(f)`` with a scalar
    function ``f`` gives the interpolant as a function evaluated on the device; ``tab(number)`` evaluates on the host."""

    def __init__(self, ts, Vs=None, axis=0, kind="cubic", tvar=None):
        if Vs is None or isinstance(Vs, (int, np.integer)):  # InterpTable1D(Vts, tvar=-1, kind): rows of [values..., t]
            if isinstance(axis, str):  # InterpTable1D(Vts, tvar, kind) given positionally
                kind = axis
            rows = np.asarray(ts, dtype=np.float64)
            if rows.ndim != 2 or rows.shape[0] == 0:
                raise ValueError("Input is empty")
            if rows.shape[1] < 2:
                raise ValueError("Invalid sized value-time data.")
            tv = -1 if Vs is None and tvar is None else int(tvar if tvar is not None else Vs)
            tv = tv + rows.shape[1] if tv < 0 else tv
            if not 0 <= tv < rows.shape[1]:
                raise ValueError("Invalid time variable index")
            ts, Vs, axis = rows[:, tv], np.delete(rows, tv, axis=1).T, 1
        ts = np.asarray(ts, dtype=np.float64).ravel()
        Vs = np.asarray(Vs, dtype=np.float64)
        if Vs.ndim == 1:
            Vs, axis = Vs[None, :], 1
        if axis == 0:
            Vs = Vs.T
        elif axis != 1:
            raise ValueError("Interpolation axis must be 0 or 1")
        self.kind = _kind(kind)
        _ascending(ts, "t")
        if ts.size != Vs.shape[1]:
            raise ValueError("Length of t coordinates must match length of interpolation axis")
        self.ts, self.vs = ts, np.ascontiguousarray(Vs)
        self.vlen, self.tsize = Vs.shape[0], ts.size
        self.teven = _is_even(ts)
        self.dvs_dts = _stencil_derivs(ts, self.vs, self.teven) if self.kind == 0 else None
        self.WarnOutOfBounds, self.ThrowOutOfBounds = True, False

    def _data(self):
        parts = [self.ts, self.vs.ravel()] + ([self.dvs_dts.ravel()] if self.kind == 0 else [])
        return np.concatenate(parts), [self.tsize, self.vlen, self.kind, int(self.teven)]

    def vf(self):
        t = Arguments(1)
        d, ip = self._data()
        return VectorFunction(GRAPH.node(VK_INTERP1D, 1, self.vlen, [t.node], d, ip, -1, 0), 1, self.vlen)

    def sf(self):
        if self.vlen != 1:
            raise ValueError("InterpTable1D storing Vector data cannot be converted to Scalar Function.")
        return ScalarFunction(self.vf().node, 1, 1)

    def _telem(self, t):
        if self.teven:
            e = int((t - self.ts[0]) / (self.ts[1] - self.ts[0]))
        else:
            e = int(np.searchsorted(self.ts, t, side="right")) - 1
        return max(min(e, self.tsize - 2), 0)

    def interp_deriv2(self, t):
        """(v, dv/dt, d2v/dt2) on the host (InterpTable1D::interp_impl, InterpTable1D.h:199-262)."""
        t = float(t)
        e = self._telem(t)
        h = self.ts[e + 1] - self.ts[e]
        u = (t - self.ts[e]) / h
        v0, v1 = self.vs[:, e], self.vs[:, e + 1]
        if self.kind == 1:
            return v0 * (1 - u) + v1 * u, (v1 - v0) / h, np.zeros(self.vlen)
        d0, d1 = self.dvs_dts[:, e], self.dvs_dts[:, e + 1]
        u2, u3 = u * u, u * u * u
        v = v0 * (2 * u3 - 3 * u2 + 1) + v1 * (-2 * u3 + 3 * u2) + d0 * ((u3 - 2 * u2 + u) * h) + d1 * ((u3 - u2) * h)
        dv = v0 * ((6 * u2 - 6 * u) / h) + v1 * ((-6 * u2 + 6 * u) / h) + d0 * (3 * u2 - 4 * u + 1) + d1 * (3 * u2 - 2 * u)
        d2v = v0 * ((12 * u - 6) / (h * h)) + v1 * ((-12 * u + 6) / (h * h)) + d0 * ((6 * u - 4) / h) + d1 * ((6 * u - 2) / h)
        return v, dv, d2v

    def interp_deriv1(self, t):
        return self.interp_deriv2(t)[:2]

    def interp(self, t):
        if np.ndim(t) == 0:
            return self.interp_deriv2(t)[0]
        return np.stack([self.interp_deriv2(x)[0] for x in np.asarray(t).ravel()], axis=1)

    def __call__(self, t):
        if isinstance(t, VectorFunction):
            if t.orows != 1:
                raise ValueError("InterpTable1D: the coordinate must be a scalar function")
            f = self.vf().eval(t)
            return ScalarFunction(f.node, f.ir, 1) if self.vlen == 1 else f
        return self.interp(t)


class InterpTable2D:
    """vf.InterpTable2D(xs, ys, Z, kind='cubic') (InterpTable2D.h:7-120): bicubic (Hermite patches with stencil
    derivatives) or bilinear interpolation of Z[iy][ix] over (x, y)."""

    def __init__(self, xs, ys, Z, kind="cubic"):
        self.kind = _kind(kind)
        xs, ys = np.asarray(xs, dtype=np.float64).ravel(), np.asarray(ys, dtype=np.float64).ravel()
        Z = np.ascontiguousarray(np.asarray(Z, dtype=np.float64))
        if xs.size < 5:
            raise ValueError("X coordinates must be larger than 4")
        if ys.size < 5:
            raise ValueError("Y  coordinates must be larger than 4")
        if Z.ndim != 2 or xs.size != Z.shape[1]:
            raise ValueError("X coordinates must match cols in Z matrix")
        if ys.size != Z.shape[0]:
            raise ValueError("Y coordinates must match rows in Z matrix")
        _ascending(xs, "X")
        _ascending(ys, "Y")
        self.xs, self.ys, self.zs = xs, ys, Z
        self.xeven, self.yeven = _is_even(xs), _is_even(ys)
        if self.kind == 0:  # InterpTable2D::calc_derivs (InterpTable2D.h:122-211)
            self.dzys = _stencil_derivs(ys, Z.T, self.yeven).T
            self.dzxs = _stencil_derivs(xs, Z, self.xeven)
            self.dzys_dxs = _stencil_derivs(xs, self.dzys, self.xeven)
        self.WarnOutOfBounds, self.ThrowOutOfBounds = True, False

    def _data(self):
        parts = [self.xs, self.ys, self.zs.ravel()]
        if self.kind == 0:
            parts += [self.dzxs.ravel(), self.dzys.ravel(), self.dzys_dxs.ravel()]
        return np.concatenate(parts), [self.xs.size, self.ys.size, self.kind, int(self.xeven), int(self.yeven)]

    def vf(self):
        d, ip = self._data()
        return ScalarFunction(GRAPH.node(VK_INTERP2D, 2, 1, [Arguments(2).node], d, ip, -1, 0), 2, 1)

    sf = vf

    def __call__(self, x, y=None):
        if isinstance(x, VectorFunction):
            xy = x if y is None else stack([x, y])
            if xy.orows != 2:
                raise ValueError("InterpTable2D: needs two scalar coordinates")
            f = self.vf().eval(xy)
            return ScalarFunction(f.node, f.ir, 1)
        xa, ya = np.asarray(x, dtype=np.float64), np.asarray(y, dtype=np.float64)
        if xa.shape != ya.shape:
            raise ValueError("InterpTable2D: x and y must have the same shape")
        # host evaluation (InterpTable2D::interp, InterpTable2D.h:392-408); arrays are indexed [iy][ix]: axis order (y, x)
        arrays = [self.zs] if self.kind else [self.zs, self.dzys, self.dzxs, self.dzys_dxs]
        f = lambda a, b: _hermite_nd([self.ys, self.xs], [self.yeven, self.xeven], arrays, self.kind == 0, (b, a))  # noqa: E731
        if xa.ndim == 0:
            return f(float(xa), float(ya))
        return np.vectorize(f)(xa, ya)

    interp = __call__


class _InterpTableND:
    """InterpTable3D / InterpTable4D (CommonFunctions/InterpTable3D.h:9-390, InterpTable4D.h): tri- / quadri-cubic Hermite
    patches through the values and every mixed first derivative at the cell corners (derivatives by the same five-point
    stencils as the lower-dimensional tables), or multilinear interpolation.  ``fs[i, j, k(, l)]`` is the value at
    ``(xs[i], ys[j], zs[k](, ws[l]))``."""
    D = 0

    def __init__(self, *args, kind="cubic", cache=False):
        args = list(args)
        if len(args) > self.D + 1 and isinstance(args[self.D + 1], str):
            kind = args[self.D + 1]
        axes, fs = [np.asarray(a, dtype=np.float64).ravel() for a in args[:self.D]], np.asarray(args[self.D], dtype=np.float64)
        self.kind = _kind(kind)
        names = "XYZW"
        for d, c in enumerate(axes):
            _ascending(c, names[d])
        if fs.shape != tuple(c.size for c in axes):
            raise ValueError("coordinates must match the dimensions of the value tensor")
        self.axes, self.fs = axes, np.ascontiguousarray(fs)
        self.even = [_is_even(c) for c in axes]
        self.arrays = [self.fs]
        if self.kind == 0:
            self.arrays = [None] * (1 << self.D)
            self.arrays[0] = self.fs
            for b in range(1, 1 << self.D):
                d = b.bit_length() - 1  # differentiate the array without the highest bit along that axis
                src = np.moveaxis(self.arrays[b & ~(1 << d)], d, -1)
                self.arrays[b] = np.ascontiguousarray(np.moveaxis(_stencil_derivs(axes[d], src, self.even[d]), -1, d))
        self.WarnOutOfBounds, self.ThrowOutOfBounds = True, False

    def vf(self):
        d = np.concatenate(self.axes + [a.ravel() for a in self.arrays])
        ip = [self.D, self.kind] + [c.size for c in self.axes] + [int(e) for e in self.even]
        return ScalarFunction(GRAPH.node(VK_INTERPND, self.D, 1, [Arguments(self.D).node], d, ip, -1, 0), self.D, 1)

    sf = vf

    def __call__(self, *xs):
        if len(xs) == 1 and isinstance(xs[0], VectorFunction):
            x = xs[0]
        elif all(isinstance(x, VectorFunction) for x in xs):
            x = stack(list(xs))
        else:
            pt = [float(v) for v in xs]
            if len(pt) != self.D:
                raise ValueError(f"InterpTable{self.D}D: needs {self.D} coordinates")
            return _hermite_nd(self.axes, self.even, self.arrays, self.kind == 0, pt)  # host evaluation, value only
        if x.orows != self.D:
            raise ValueError(f"InterpTable{self.D}D: needs {self.D} scalar coordinates")
        f = self.vf().eval(x)
        return ScalarFunction(f.node, f.ir, 1)

    interp = __call__


class InterpTable3D(_InterpTableND):
    D = 3


class InterpTable4D(_InterpTableND):
    D = 4


Element = ScalarFunction
Segment = VectorFunction
Segment2 = VectorFunction
Segment3 = VectorFunction


class PyVectorFunction:
    def __init__(self, *a, **k):
        raise NotImplementedError("PyVectorFunction is a host callback and cannot be lowered to a device tape; "
                                  "there is no CPU fallback on this path")


PyScalarFunction = PyVectorFunction
