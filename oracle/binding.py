"""ctypes binding of oracle/_build/liboracle.so (test infrastructure only)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
_ip = C.POINTER(C.c_int)
_dp = C.POINTER(C.c_double)
_lib = None

FL_OBJ, FL_CON, FL_JAC, FL_HESS, FL_GRAD = 1, 2, 4, 8, 16
# flags per NonLinearProgram::eval* mode (NonLinearProgram.cpp:347-683)
MODE_FLAGS = {
    "KKT": FL_OBJ | FL_CON | FL_JAC | FL_HESS | FL_GRAD,
    "KKTNO": FL_CON | FL_JAC | FL_HESS | FL_GRAD,
    "SOE": FL_CON | FL_JAC,
    "AUG": FL_OBJ | FL_CON | FL_JAC | FL_GRAD,
    "RHS": FL_OBJ | FL_CON | FL_GRAD,
    "OGC": FL_OBJ | FL_CON | FL_GRAD,
    "OCC": FL_OBJ | FL_CON,
    "OBJ": FL_OBJ,
}


def build(force=False):
    src = os.path.join(_HERE, "oracle.cpp")
    if force or not os.path.exists(LIB_PATH) or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(LIB_PATH)):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return LIB_PATH


def use_variant(name):
    """Switch every Oracle created from now on to another build of oracle.cpp: "ieee" (default, the parity copy) or
    "fast" (-O2 -march=native -ffast-math, the reference's flags; timing only, built on first use)."""
    global _lib, LIB_PATH
    if name == "fast":
        path = os.path.join(_HERE, "_build", "liboracle_fast.so")
        # -march=native: rebuild when the copy was made on another machine (built files travel with the repository)
        stamp, cpu = os.path.join(_HERE, "_build", "fast.host"), ""
        try:
            with open("/proc/cpuinfo") as f:
                cpu = next((ln for ln in f if ln.startswith("flags")), "")
        except OSError:
            pass
        import hashlib
        tag = hashlib.sha1(cpu.encode()).hexdigest()
        if not os.path.exists(stamp) or open(stamp).read() != tag:
            if os.path.exists(path):
                os.remove(path)
        subprocess.check_call(["make", "-C", _HERE, "-s", "fast"])
        with open(stamp, "w") as f:
            f.write(tag)
    elif name == "ieee":
        path = os.path.join(_HERE, "_build", "liboracle.so")
    else:
        raise ValueError(name)
    LIB_PATH = path
    _lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.orc_last_error.restype = C.c_char_p
        L.orc_problem_create.restype = C.c_void_p
        L.orc_problem_create.argtypes = [C.c_int] * 3
        L.orc_problem_destroy.argtypes = [C.c_void_p]
        L.orc_problem_add_function.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.c_int, _ip, _ip, _dp, C.c_int]
        L.orc_problem_analyze.argtypes = [C.c_void_p]
        L.orc_problem_dims.argtypes = [C.c_void_p, C.POINTER(C.c_longlong)]
        L.orc_problem_pattern.argtypes = [C.c_void_p, _ip, _ip, _ip]
        L.orc_problem_eval.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double] + [_dp] * 10
        L.orc_function_computeall.argtypes = [C.c_char_p] + [_dp] * 7 + [_ip, _ip]
        _lib = L
    return _lib


def _chk(rc):
    if rc < 0:
        raise RuntimeError("oracle: " + lib().orc_last_error().decode())
    return rc


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def hardware_threads():
    return int(lib().orc_hardware_threads())


def computeall(desc, x, l, appconst=None):
    """fx, jx, gx, hx of one application -- the reference's ``Fun.computeall(X, L)``."""
    L = lib()
    ir, orr = C.c_int(), C.c_int()
    _chk(L.orc_function_computeall(desc.encode(), None, None, None, None, None, None, None, C.byref(ir), C.byref(orr)))
    ir, orr = ir.value, orr.value
    x = np.ascontiguousarray(x, dtype=np.float64)
    l = np.ascontiguousarray(l, dtype=np.float64)
    assert x.size == ir and l.size == orr, (x.size, ir, l.size, orr)
    ac = None if appconst is None else np.ascontiguousarray(appconst, dtype=np.float64)
    fx, jx, gx, hx = np.zeros(orr), np.zeros((orr, ir)), np.zeros(ir), np.zeros((ir, ir))
    _chk(L.orc_function_computeall(desc.encode(), _d(x), _d(l), _d(ac), _d(fx), _d(jx), _d(gx), _d(hx), None, None))
    return fx, jx, gx, hx


class Oracle:
    """CPU NonLinearProgram built from the same function list as a product problem."""

    def __init__(self, nvars, neq, niq):
        self.L = lib()
        self.h = C.c_void_p(self.L.orc_problem_create(nvars, neq, niq))
        self.nvars, self.neq, self.niq = nvars, neq, niq
        self._dims = None

    def __del__(self):
        try:
            self.L.orc_problem_destroy(self.h)
        except Exception:
            pass

    @classmethod
    def from_problem(cls, nlp):
        """Mirror a product ``NonLinearProgram`` (asset_asrl_b200.Solvers): same functions, same index maps."""
        d = nlp.dims()
        o = cls(d["nvars"], d["neq"], d["niq"])
        for f in range(d["nfuncs"]):
            fi = nlp.function_info(f)
            V, Cc = nlp.function_indices(f)
            ac = nlp.function_appconst(f)
            o.add_function(fi["role"], nlp.tape_describe(fi["tape"]), fi["linear"], V, Cc, ac)
        o.analyze()
        return o

    def add_function(self, role, desc, linear, V, Cc=None, appconst=None):
        V = np.ascontiguousarray(V, dtype=np.int32)
        nappl = V.shape[0]
        Cc = np.ascontiguousarray(Cc if Cc is not None else np.zeros((nappl, 0)), dtype=np.int32)
        nac = 0 if appconst is None or appconst.size == 0 else appconst.shape[1]
        ac = None if nac == 0 else np.ascontiguousarray(appconst, dtype=np.float64)
        return _chk(self.L.orc_problem_add_function(self.h, role, desc.encode(), int(linear), nappl, V.ctypes.data_as(_ip),
                                                    Cc.ctypes.data_as(_ip), _d(ac), nac))

    def analyze(self):
        _chk(self.L.orc_problem_analyze(self.h))
        out = (C.c_longlong * 4)()
        self.L.orc_problem_dims(self.h, out)
        self._dims = dict(kktdim=int(out[0]), nnz=int(out[1]), n_user=int(out[2]), n_solver=int(out[3]))

    def dims(self):
        return self._dims

    def pattern(self):
        d = self._dims
        indptr = np.zeros(d["kktdim"] + 1, dtype=np.int32)
        indices = np.zeros(d["nnz"], dtype=np.int32)
        loc = np.zeros(d["n_user"] + d["n_solver"], dtype=np.int32)
        self.L.orc_problem_pattern(self.h, indptr.ctypes.data_as(_ip), indices.ctypes.data_as(_ip), loc.ctypes.data_as(_ip))
        return indptr, indices, loc

    def eval(self, mode, obj_scale, X, LE=None, LI=None, solver_coeffs=None, threads=1, outputs=True):
        """Returns dict(val, PGX, AGX, FXE, FXI, KKT) for a NonLinearProgram::eval* mode name."""
        d = self._dims
        X = np.ascontiguousarray(X, dtype=np.float64)
        LE = np.zeros(self.neq) if LE is None else np.ascontiguousarray(LE, dtype=np.float64)
        LI = np.zeros(self.niq) if LI is None else np.ascontiguousarray(LI, dtype=np.float64)
        sc = None if solver_coeffs is None else np.ascontiguousarray(solver_coeffs, dtype=np.float64)
        val = np.zeros(1)
        out = dict(PGX=np.zeros(self.nvars), AGX=np.zeros(self.nvars), FXE=np.zeros(self.neq), FXI=np.zeros(self.niq),
                   KKT=np.zeros(d["nnz"]))
        _chk(self.L.orc_problem_eval(self.h, threads, MODE_FLAGS[mode], float(obj_scale), _d(X), _d(LE), _d(LI), _d(sc), _d(val),
                                     _d(out["PGX"]), _d(out["AGX"]), _d(out["FXE"]), _d(out["FXI"]), _d(out["KKT"])))
        out["val"] = float(val[0])
        return out
