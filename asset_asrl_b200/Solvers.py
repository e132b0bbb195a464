"""NonLinearProgram: the KKT evaluation object PSIOPT calls, evaluated on the GPU.

Mirrors the eight eval entry points of /root/reference/src/Solvers/NonLinearProgram.h:411-478 (argument meaning
and output set per mode) and the set-up products of make_NLP / analyzeSparsity (NonLinearProgram.cpp:25-39,
:267-344).  Outputs are *assigned* sums (the reference accumulates into caller-zeroed buffers; the results are
the same numbers).  There is no CPU fallback: creating the device plan without a GPU raises.
"""
import ctypes as C
import time

import numpy as np

from . import _lib
from ._lib import lib

EVAL_KKT, EVAL_KKTNO, EVAL_SOE, EVAL_AUG, EVAL_RHS, EVAL_OGC, EVAL_OCC, EVAL_OBJ = range(8)
OUT_ARRAYS = ("KKT", "PGX", "AGX", "FXE", "FXI", "val")  # order of the gather lists / cut sets in the C ABI


class NonLinearProgram:
    def __init__(self, nvars=0, neq=0, niq=0):
        self.h = lib.ab_problem_create(nvars, neq, niq)
        if not self.h:
            raise MemoryError("ab_problem_create failed")
        self.plan = None
        self.device = 0
        self._dims = None

    def __del__(self):
        try:
            if self.plan:
                lib.ab_plan_destroy(self.plan)
            if self.h:
                lib.ab_problem_destroy(self.h)
        except Exception:
            pass

    # ------------------------------------------------------------------ set-up
    def set_dims(self, nvars, neq, niq):
        _lib.check(lib.ab_problem_set_dims(self.h, nvars, neq, niq))
        self._dims = None

    def add_tape(self, fun, with_adjoint=True):
        from .VectorFunctions import GRAPH
        return _lib.check(lib.ab_problem_add_tape(self.h, GRAPH.h, fun.node, 1 if with_adjoint else 0))

    def add_function(self, role, tape, linear, vindex, cindex=None, appconst=None, unique_constraints=True):
        """vindex: [nappl][ir] global variable indices; cindex: [nappl][orows] rows within the eq / iq block."""
        V = np.ascontiguousarray(vindex, dtype=np.int32)
        nappl = V.shape[0]
        va, vp = _lib.iarr(V)
        ca, cp = _lib.iarr(cindex if cindex is not None else [])
        aa, ap = _lib.darr(appconst if appconst is not None else [])
        return _lib.check(lib.ab_problem_add_function(self.h, role, tape, int(linear), int(unique_constraints), nappl, vp,
                                                      cp if cindex is not None else None, ap if appconst is not None else None))

    def window(self, mesh_vars, seg_vars, nseg, seg_lo, seg_hi, halo=8):
        """Rank-local analysis (ab_problem_window): keep the applications homed within ``halo`` segments of
        [seg_lo, seg_hi), own those homed inside it.  Call before analyze()."""
        _lib.check(lib.ab_problem_window(self.h, int(mesh_vars), int(seg_vars), int(nseg), int(seg_lo), int(seg_hi), int(halo)))
        self._dims = None
        self.windowed = True

    def window_info(self, f):
        out = (C.c_int * 4)()
        _lib.check(lib.ab_problem_window_info(self.h, f, out))
        return dict(zip(["app_base", "nappl_whole", "own_lo", "own_hi"], list(out)))

    def window_far_vars(self):
        n = int(lib.ab_problem_window_far_vars(self.h, None, 0))
        out = np.zeros(max(n, 1), dtype=np.int32)
        lib.ab_problem_window_far_vars(self.h, out.ctypes.data_as(_lib._ip), n)
        return out[:n]

    def analyze(self):
        _lib.check(lib.ab_problem_analyze(self.h))
        self._dims = None

    def dims(self):
        if self._dims is None:
            out = (C.c_longlong * 10)()
            _lib.check(lib.ab_problem_dims(self.h, out))
            self._dims = dict(zip(["nvars", "neq", "niq", "kktdim", "nnz", "n_user", "n_solver", "stage_size", "nfuncs",
                                   "ntapes"], [int(v) for v in out]))
        return self._dims

    PrimalVars = property(lambda s: s.dims()["nvars"])
    EqualCons = property(lambda s: s.dims()["neq"])
    InequalCons = property(lambda s: s.dims()["niq"])
    KKTdim = property(lambda s: s.dims()["kktdim"])

    def pattern(self):
        """(indptr, indices) of the upper-triangular row-major CSR KKT pattern (analyzeSparsity)."""
        d = self.dims()
        indptr = np.zeros(d["kktdim"] + 1, dtype=np.int32)
        indices = np.zeros(d["nnz"], dtype=np.int32)
        _lib.check(lib.ab_problem_pattern(self.h, indptr.ctypes.data_as(_lib._ip), indices.ctypes.data_as(_lib._ip)))
        return indptr, indices

    def kkt_locations(self):
        d = self.dims()
        out = np.zeros(d["n_user"] + d["n_solver"], dtype=np.int32)
        _lib.check(lib.ab_problem_kkt_locations(self.h, out.ctypes.data_as(_lib._ip)))
        return out

    def function_info(self, f):
        out = (C.c_int * 8)()
        _lib.check(lib.ab_problem_function_info(self.h, f, out))
        return dict(zip(["role", "tape", "ir", "orows", "linear", "unique_constraints", "nappl", "nappconst"], list(out)))

    def function_indices(self, f):
        fi = self.function_info(f)
        V = np.zeros((fi["nappl"], fi["ir"]), dtype=np.int32)
        Cc = np.zeros((fi["nappl"], fi["orows"] if fi["role"] != 0 else 0), dtype=np.int32)
        _lib.check(lib.ab_problem_function_indices(self.h, f, V.ctypes.data_as(_lib._ip),
                                                   Cc.ctypes.data_as(_lib._ip) if Cc.size else None))
        return V, Cc

    def function_appconst(self, f):
        fi = self.function_info(f)
        out = np.zeros((fi["nappl"], fi["nappconst"]))
        if out.size:
            _lib.check(lib.ab_problem_function_appconst(self.h, f, _lib.dptr(out)))
        return out

    def tape(self, t):
        out = (C.c_int * 8)()
        _lib.check(lib.ab_problem_tape_info(self.h, t, out))
        info = dict(zip(["n_instr", "n_primal", "slots_primal", "slots_total", "nin", "nout", "nappconst", "has_adjoint"],
                        list(out)))
        dt = np.dtype([("op", "<u2"), ("dst", "<u2"), ("a", "<u2"), ("b", "<u2"), ("imm", "<f8")])
        code = np.zeros(info["n_instr"], dtype=dt)
        if info["n_instr"]:
            _lib.check(lib.ab_problem_tape_code(self.h, t, code.ctypes.data_as(C.c_void_p)))
        return info, code

    def program_stats(self, f, mode=4):
        """Size of the device program of function ``f`` in an evaluation mode (4 = values + Jacobian + Hessian)."""
        out = (C.c_longlong * 8)()
        _lib.check(lib.ab_problem_program_stats(self.h, f, mode, out))
        return dict(zip(["ops", "fma", "special", "loads", "outs", "bundles", "slots", "consts"], [int(v) for v in out]))

    def program(self, f, mode=4):
        """(stats, code, consts) of the unlinked device program: code is a structured array of 16-byte instructions."""
        st = self.program_stats(f, mode)
        dt = np.dtype([("d", "<u4"), ("a", "<u4"), ("b", "<u4"), ("c", "<u4")])
        code = np.zeros(4 * st["bundles"], dtype=dt)
        consts = np.zeros(st["consts"])
        _lib.check(lib.ab_problem_program_code(self.h, f, mode, code.ctypes.data_as(C.c_void_p), _lib.dptr(consts)))
        return st, code, consts

    def program_tables(self, f, mode=4):
        """The interpolation-table pool the program of function ``f`` reads (empty for a function without tables)."""
        n = _lib.check(lib.ab_problem_program_tables(self.h, f, mode, None, 0))
        out = np.zeros(n)
        if n:
            _lib.check(lib.ab_problem_program_tables(self.h, f, mode, _lib.dptr(out), n))
        return out

    def fused_info(self, min_threads=32):
        """Host-side description of the owner-writes fused interior (no GPU needed)."""
        out = (C.c_longlong * 12)()
        why = C.create_string_buffer(512)
        _lib.check(lib.ab_problem_fused_info(self.h, min_threads, out, why, 512))
        keys = ["ok", "anchor", "t_lo", "t_hi", "vars_per_segment", "slots_per_segment", "families", "home_families", "loads",
                "outputs", "fma", "pieces"]
        d = dict(zip(keys, [int(v) for v in out]))
        d["why"] = why.value.decode()
        return d

    def fused_ptx(self, piece=0, min_threads=32):
        n = _lib.check(lib.ab_problem_fused_ptx(self.h, min_threads, piece, None, 0))
        if n == 0:
            return None
        buf = C.create_string_buffer(n + 1)
        _lib.check(lib.ab_problem_fused_ptx(self.h, min_threads, piece, buf, n + 1))
        return buf.value.decode()

    def fused_residual_ptx(self, mode, min_threads=32):
        """PTX of the residual-only fused kernel of eval mode 4 (RHS), 5 (OGC), 6 (OCC) or 7 (OBJ); None if there is none."""
        n = _lib.check(lib.ab_problem_fused_residual_ptx(self.h, min_threads, mode, None, 0))
        if n == 0:
            return None
        buf = C.create_string_buffer(n + 1)
        _lib.check(lib.ab_problem_fused_residual_ptx(self.h, min_threads, mode, buf, n + 1))
        return buf.value.decode()

    def fused_layout(self, min_threads=32):
        """{"vb", "vs", "kb", "ks", "ex_lo", "ex_hi"} of the fused interior: owned variables / CSR slots per thread and, per
        function, the applications the fused threads evaluate."""
        nf = self.dims()["nfuncs"]
        out = (C.c_longlong * 4)()
        lo, hi = np.zeros(nf, dtype=np.int32), np.zeros(nf, dtype=np.int32)
        _lib.check(lib.ab_problem_fused_layout(self.h, min_threads, out, lo.ctypes.data_as(_lib._ip), hi.ctypes.data_as(_lib._ip)))
        return {"vb": int(out[0]), "vs": int(out[1]), "kb": int(out[2]), "ks": int(out[3]), "ex_lo": lo, "ex_hi": hi}

    def dest_maps(self, f):
        """(kkt_dest [nappl, E], f_dest [nappl, orows], first gradient staging slot, first objective-value staging slot)."""
        fi = self.function_info(f)
        E = (0 if fi["linear"] else fi["ir"] * (fi["ir"] + 1) // 2) + (fi["ir"] * fi["orows"] if fi["role"] else 0)
        kd = np.zeros((fi["nappl"], E), dtype=np.int32)
        fd = np.zeros((fi["nappl"], fi["orows"] if fi["role"] else 0), dtype=np.int32)
        out = (C.c_longlong * 2)()
        _lib.check(lib.ab_problem_dest_maps(self.h, f, kd.ctypes.data_as(_lib._ip), fd.ctypes.data_as(_lib._ip), out))
        return kd, fd, int(out[0]), int(out[1])

    def fused_mode_ptx(self, mode, piece=0, min_threads=32):
        """PTX of piece ``piece`` of the fused kernels of eval mode 1 (KKTNO), 2 (SOE), 3 (AUG) -- the partial-matrix
        evaluations, same row blocks as the evalKKT pieces -- or 4..7 (piece 0); None past the last piece."""
        n = _lib.check(lib.ab_problem_fused_mode_ptx(self.h, min_threads, mode, piece, None, 0))
        if n == 0:
            return None
        buf = C.create_string_buffer(n + 1)
        _lib.check(lib.ab_problem_fused_mode_ptx(self.h, min_threads, mode, piece, buf, n + 1))
        return buf.value.decode()

    def kernel_ptx(self, f, mode=4, indexed=False):
        """PTX of the specialised kernel for function ``f``: the affine variant (None if the function has no affine
        application range) or the indexed variant that reads explicit index maps."""
        n = _lib.check(lib.ab_problem_kernel_ptx(self.h, f, mode, int(indexed), None, 0))
        if n == 0:
            return None
        buf = C.create_string_buffer(n + 1)
        _lib.check(lib.ab_problem_kernel_ptx(self.h, f, mode, int(indexed), buf, n + 1))
        return buf.value.decode()

    @staticmethod
    def assemble_ptx(ptx):
        """Assemble PTX for sm_100a in-process (no GPU needed): {cubin_bytes, registers, spill_bytes}."""
        out = (C.c_longlong * 3)()
        _lib.check(lib.ab_ptx_assemble(ptx.encode(), out))
        return {"cubin_bytes": int(out[0]), "registers": int(out[1]), "spill_bytes": int(out[2])}

    def tape_describe(self, t):
        """Text form of what tape ``t`` was lowered from (introspection; the CPU oracle in tests/ rebuilds it)."""
        n = _lib.check(lib.ab_problem_tape_describe(self.h, t, None, 0))
        buf = C.create_string_buffer(n + 1)
        _lib.check(lib.ab_problem_tape_describe(self.h, t, buf, n + 1))
        return buf.value.decode()

    # ------------------------------------------------------------------ device
    def to_device(self, device=0, rank=0, world=1, group=None):
        """Build the device plan.  ``world > 1``: this process evaluates rank's share of every function (contiguous
        application ranges); the cut set is summed over ``group`` (a torch.distributed process
        group, NCCL) after every evaluation and every output ends up non-zero on exactly one rank."""
        if self.plan:
            lib.ab_plan_destroy(self.plan)
            self.plan = None
        if self._dims is None or not self.dims()["kktdim"]:
            self.analyze()
        p = C.c_void_p()
        if world > 1:
            _lib.check0(lib.ab_plan_create_sharded(self.h, device, rank, world, C.byref(p)))
        else:
            _lib.check0(lib.ab_plan_create(self.h, device, C.byref(p)))
        self.plan = p
        self.device = device
        self.rank, self.world, self.group = rank, world, group
        self._cut = None
        return self

    def partition(self, rank, world):
        """Host-side description of one rank's share (no GPU needed): application ranges per function, the gather
        lists restricted to the rank's own sources and the cut set {array: (target, owner, mine)}."""
        q = lib.ab_partition_create(self.h, rank, world)
        if not q:
            raise ValueError(_lib.last_error())
        try:
            nf = self.dims()["nfuncs"]
            rng = np.zeros((nf, 2), dtype=np.int32)
            for f in range(nf):
                _lib.check(lib.ab_partition_app_range(q, f, rng[f].ctypes.data_as(_lib._ip)))
            gathers, cuts = {}, {}
            for a, name in enumerate(OUT_ARRAYS):
                lists = []
                for lvl in range(2):
                    sz = (C.c_longlong * 2)()
                    _lib.check(lib.ab_partition_gather_sizes(q, a, lvl, sz))
                    ptr = np.zeros(sz[0] + 1, dtype=np.int64)
                    src = np.zeros(max(sz[1], 1), dtype=np.int32)
                    tgt = np.zeros(max(sz[0], 1), dtype=np.int32)
                    _lib.check(lib.ab_partition_gather(q, a, lvl, ptr.ctypes.data_as(_lib._llp), src.ctypes.data_as(_lib._ip),
                                                       tgt.ctypes.data_as(_lib._ip)))
                    lists.append((ptr, src[:sz[1]], tgt[:sz[0]]))
                gathers[name] = lists
                n = int(lib.ab_partition_cut_size(q, a))
                t, o, m = (np.zeros(max(n, 1), dtype=np.int32) for _ in range(3))
                _lib.check(lib.ab_partition_cut(q, a, t.ctypes.data_as(_lib._ip), o.ctypes.data_as(_lib._ip),
                                                m.ctypes.data_as(_lib._ip)))
                cuts[name] = (t[:n], o[:n], m[:n])
            return {"app_ranges": rng, "gathers": gathers, "cuts": cuts}
        finally:
            lib.ab_partition_destroy(q)

    def gather_lists(self):
        """The unpartitioned fixed-order gather lists {array: [(ptr, src, target)] * 2} and the staging buffer size."""
        out, stage = {}, 0
        for a, name in enumerate(OUT_ARRAYS):
            lists = []
            for lvl in range(2):
                sz = (C.c_longlong * 3)()
                _lib.check(lib.ab_problem_gather_sizes(self.h, a, lvl, sz))
                stage = int(sz[2])
                ptr = np.zeros(sz[0] + 1, dtype=np.int64)
                src = np.zeros(max(sz[1], 1), dtype=np.int32)
                tgt = np.zeros(max(sz[0], 1), dtype=np.int32)
                _lib.check(lib.ab_problem_gather(self.h, a, lvl, ptr.ctypes.data_as(_lib._llp), src.ctypes.data_as(_lib._ip),
                                                 tgt.ctypes.data_as(_lib._ip)))
                lists.append((ptr, src[:sz[1]], tgt[:sz[0]]))
            out[name] = lists
        return out, stage

    def reduce_cut(self):
        """Sum the cut set over the ranks (asynchronous on the plan's stream): pack, NCCL all-reduce, unpack."""
        if getattr(self, "world", 1) <= 1:
            return
        n = int(lib.ab_plan_cut_count(self.plan))
        if n == 0:
            return
        import torch
        import torch.distributed as dist
        cut = self.cut_tensor()
        _lib.check0(lib.ab_plan_cut_pack(self.plan))
        with torch.cuda.stream(self._tstream):
            dist.all_reduce(cut, op=dist.ReduceOp.SUM, group=self.group)
        _lib.check0(lib.ab_plan_cut_unpack(self.plan))

    def check_guards(self):
        """0 if no kernel wrote outside a device array of this plan (needs ASSET_B200_GUARD=1 when the plan was made)."""
        rc = int(lib.ab_plan_check_guards(self.plan))
        if rc != 0:
            raise _lib.AssetB200Error(f"guard check {rc}: {_lib.last_error()}")
        return 0

    def device_tensor(self, name):
        """Output / input array `name` (X, LE, LI, KKT, FXE, FXI, PGX, AGX, val) as a torch tensor aliasing the plan's
        device memory (ab_plan_device_ptr; no copy) -- what a device-resident consumer of the values would read."""
        import torch
        self._need_plan()
        d = self.dims()
        which = {"X": (0, d["nvars"]), "LE": (1, d["neq"]), "LI": (2, d["niq"]), "KKT": (3, d["nnz"]), "FXE": (4, d["neq"]),
                 "FXI": (5, d["niq"]), "PGX": (6, d["nvars"]), "AGX": (7, d["nvars"]), "val": (8, 1)}[name]
        if which[1] == 0:
            return torch.zeros(0, dtype=torch.float64, device=torch.device("cuda", self.device))

        class _Raw:
            pass
        raw = _Raw()
        raw.__cuda_array_interface__ = {"shape": (int(which[1]),), "typestr": "<f8", "version": 3, "strides": None,
                                        "data": (int(lib.ab_plan_device_ptr(self.plan, which[0])), False)}
        return torch.as_tensor(raw, device=torch.device("cuda", self.device))

    def cut_tensor(self):
        """The packed cut buffer as a torch tensor aliasing the plan's device memory (no copy)."""
        import torch
        n = int(lib.ab_plan_cut_count(self.plan))  # depends on which evaluation ran last (fused / per-function)
        if self._cut is None or self._cut.numel() != n:

            class _Raw:
                pass
            raw = _Raw()
            raw.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (int(lib.ab_plan_cut_buffer(self.plan)), False),
                                            "version": 3, "strides": None}
            self._cut = torch.as_tensor(raw, device=torch.device("cuda", self.device))
            self._tstream = torch.cuda.ExternalStream(self.stream(), device=self.device)
        return self._cut

    def output_ranges(self):
        """{array: [(lo, hi), ...]} of what this (sharded) plan writes and reads back."""
        res = {}
        for a, nm in enumerate(OUT_ARRAYS):
            n = _lib.check(lib.ab_plan_output_ranges(self.plan, a, None, 0))
            out = (C.c_longlong * max(2 * n, 2))()
            _lib.check(lib.ab_plan_output_ranges(self.plan, a, out, n))
            res[nm] = [(int(out[2 * k]), int(out[2 * k + 1])) for k in range(n)]
        return res

    def input_ranges(self):
        """{"X" | "LE" | "LI": [(lo, hi), ...]}: what the functions of this problem read of the iterate -- everything for an
        ordinary problem, the rank's segments + halo + far variables for a windowed one (ab_problem_input_ranges; the plan
        keeps its own copy, ab_plan_input_ranges, which ``eval_host_sharded(sliced=True)`` uploads)."""
        res = {}
        for w, nm in enumerate(("X", "LE", "LI")):
            if self.plan:
                call = lambda out, cap: lib.ab_plan_input_ranges(self.plan, w, out, cap)  # noqa: E731
            else:
                call = lambda out, cap: lib.ab_problem_input_ranges(self.h, w, out, cap)  # noqa: E731
            n = _lib.check(call(None, 0))
            out = (C.c_longlong * max(2 * n, 2))()
            _lib.check(call(out, n))
            res[nm] = [(int(out[2 * k]), int(out[2 * k + 1])) for k in range(n)]
        return res

    def cut_pack(self):
        _lib.check0(lib.ab_plan_cut_pack(self.plan))

    def cut_unpack(self):
        _lib.check0(lib.ab_plan_cut_unpack(self.plan))

    def _need_plan(self):
        if not self.plan:
            self.to_device(self.device)

    def set_solver_coeffs(self, coeffs):
        self._need_plan()
        a, ap = _lib.darr(coeffs)
        assert a.size == self.dims()["n_solver"]
        _lib.check0(lib.ab_plan_set_solver_coeffs(self.plan, ap))

    def _eval(self, mode, ObjScale, X, LE, LI, want):
        self._need_plan()
        d = self.dims()
        X = np.ascontiguousarray(X, dtype=np.float64)
        LE = np.ascontiguousarray(LE if LE is not None else np.zeros(d["neq"]), dtype=np.float64)
        LI = np.ascontiguousarray(LI if LI is not None else np.zeros(d["niq"]), dtype=np.float64)
        if X.size != d["nvars"] or LE.size != d["neq"] or LI.size != d["niq"]:
            raise ValueError("X / LE / LI sizes do not match the problem")
        out = {}
        if "val" in want:
            out["val"] = np.zeros(1)
        if "PGX" in want:
            out["PGX"] = np.zeros(d["nvars"])
        if "AGX" in want:
            out["AGX"] = np.zeros(d["nvars"])
        if "FX" in want:
            out["FXE"] = np.zeros(d["neq"])
            out["FXI"] = np.zeros(d["niq"])
        if "KKT" in want:
            out["KKT"] = np.zeros(d["nnz"])
        g = lambda k: _lib.dptr(out[k]) if k in out else None  # noqa: E731
        if getattr(self, "world", 1) > 1:  # sharded: the cut set is reduced between the evaluation and the read-back
            self.set_inputs(X, LE, LI)
            self.eval_device(mode, ObjScale)
            _lib.check0(lib.ab_plan_get_outputs(self.plan, g("val"), g("PGX"), g("AGX"), g("FXE"), g("FXI"), g("KKT")))
        else:
            _lib.check0(lib.ab_plan_eval_host(self.plan, mode, float(ObjScale), _lib.dptr(X), _lib.dptr(LE), _lib.dptr(LI),
                                              g("val"), g("PGX"), g("AGX"), g("FXE"), g("FXI"), g("KKT")))
        if "val" in out:
            out["val"] = float(out["val"][0])
        return out

    def evalKKT(self, ObjScale, X, LE, LI):
        return self._eval(EVAL_KKT, ObjScale, X, LE, LI, {"val", "PGX", "AGX", "FX", "KKT"})

    def evalKKTNO(self, ObjScale, X, LE, LI):
        return self._eval(EVAL_KKTNO, ObjScale, X, LE, LI, {"AGX", "FX", "KKT"})

    def evalSOE(self, ObjScale, X, LE, LI):
        return self._eval(EVAL_SOE, ObjScale, X, LE, LI, {"FX", "KKT"})

    def evalAUG(self, ObjScale, X, LE, LI):
        return self._eval(EVAL_AUG, ObjScale, X, LE, LI, {"val", "PGX", "AGX", "FX", "KKT"})

    def evalRHS(self, ObjScale, X, LE, LI):
        return self._eval(EVAL_RHS, ObjScale, X, LE, LI, {"val", "PGX", "AGX", "FX"})

    def evalOGC(self, ObjScale, X):
        return self._eval(EVAL_OGC, ObjScale, X, None, None, {"val", "PGX", "FX"})

    def evalOCC(self, ObjScale, X):
        return self._eval(EVAL_OCC, ObjScale, X, None, None, {"val", "FX"})

    def evalOBJ(self, ObjScale, X):
        return self._eval(EVAL_OBJ, ObjScale, X, None, None, {"val"})

    # ------------------------------------------------------------------ device-resident interface
    def set_inputs(self, X, LE, LI):
        """Host -> device copy of the iterate (asynchronous on the plan's stream)."""
        self._need_plan()
        _lib.check0(lib.ab_plan_set_inputs(self.plan, _lib.dptr(X), _lib.dptr(LE), _lib.dptr(LI)))

    def eval_device(self, mode, ObjScale):
        """Launch one evaluation on the device-resident inputs (asynchronous); outputs stay on the device."""
        _lib.check0(lib.ab_plan_eval(self.plan, mode, float(ObjScale)))
        if getattr(self, "world", 1) > 1:
            self.reduce_cut()

    def get_outputs(self, want=("val", "PGX", "AGX", "FXE", "FXI", "KKT")):
        """Device -> host copy of the last evaluation's outputs (synchronises the plan's stream)."""
        d = self.dims()
        size = {"val": 1, "PGX": d["nvars"], "AGX": d["nvars"], "FXE": d["neq"], "FXI": d["niq"], "KKT": d["nnz"]}
        out = {k: np.zeros(size[k]) for k in want}
        g = lambda k: _lib.dptr(out[k]) if k in out else None  # noqa: E731
        _lib.check0(lib.ab_plan_get_outputs(self.plan, g("val"), g("PGX"), g("AGX"), g("FXE"), g("FXI"), g("KKT")))
        if "val" in out:
            out["val"] = float(out["val"][0])
        return out

    def sync(self):
        _lib.check0(lib.ab_plan_sync(self.plan))

    def stream(self):
        return int(lib.ab_plan_stream(self.plan) or 0)

    def eval_host(self, mode, ObjScale, X, LE, LI, val, PGX, AGX, FXE, FXI, KKT):
        """H2D + eval + D2H into caller-owned (ideally pinned) float64 arrays; None skips an output."""
        g = _lib.dptr
        _lib.check0(lib.ab_plan_eval_host(self.plan, mode, float(ObjScale), g(X), g(LE), g(LI), g(val), g(PGX), g(AGX),
                                          g(FXE), g(FXI), g(KKT)))

    def eval_host_sharded(self, mode, ObjScale, X, LE, LI, val, PGX, AGX, FXE, FXI, KKT, compact=False, sliced=False):
        """eval_host for a sharded plan: H2D of the iterate (the whole of it, or with ``sliced`` only the ranges this
        rank's functions read, see input_ranges), evaluation, cut-set all-reduce, D2H of this rank's output ranges only."""
        g = _lib.dptr
        _lib.check0((lib.ab_plan_set_inputs_sliced if sliced else lib.ab_plan_set_inputs)(self.plan, g(X), g(LE), g(LI)))
        self.eval_device(mode, ObjScale)
        if compact:  # the arrays hold this rank's output ranges back to back (see output_ranges)
            _lib.check0(lib.ab_plan_get_outputs_packed(self.plan, g(KKT), g(PGX), g(AGX), g(FXE), g(FXI), g(val)))
        else:
            _lib.check0(lib.ab_plan_get_outputs(self.plan, g(val), g(PGX), g(AGX), g(FXE), g(FXI), g(KKT)))

    def profile(self, enable=True):
        self._need_plan()
        _lib.check0(lib.ab_plan_profile(self.plan, int(enable)))

    def profile_times(self):
        """[(label, ms)] per kernel launch group of the last eval (device time from CUDA events)."""
        n = _lib.check(lib.ab_plan_profile_read(self.plan, None, 0))
        buf = (C.c_float * max(n, 1))()
        n = _lib.check(lib.ab_plan_profile_read(self.plan, buf, n))
        out = []
        for k in range(n):
            ln = lib.ab_plan_profile_label(self.plan, k, None, 0)
            sb = C.create_string_buffer(ln + 1)
            lib.ab_plan_profile_label(self.plan, k, sb, ln + 1)
            out.append((sb.value.decode(), float(buf[k])))
        return out

    def launches(self):
        return int(lib.ab_plan_launches(self.plan)) if self.plan else 0

    def fused_segments(self):
        """Segments evaluated by the fused owner-writes kernel in the last eval (0: per-function path only)."""
        return int(lib.ab_plan_fused_segments(self.plan)) if self.plan else 0

    def specialised_launches(self):
        """How many launches of the last eval were specialised (run-time assembled) kernels."""
        return int(lib.ab_plan_specialised_launches(self.plan)) if self.plan else 0

    def kernel_info(self, k, mode=4):
        out = (C.c_int * 4)()
        _lib.check(lib.ab_plan_kernel_info(self.plan, k, mode, out))
        return dict(zip(["state", "registers", "spill_bytes", "affine_apps"], [int(v) for v in out]))

    def NLPTest(self, X, n, seed=67890):
        """NLPTest protocol (NonLinearProgram.cpp:686-820): multipliers 100*U(-1,1), n repetitions of evalKKT;
        returns mean ms per evalKKT (host buffers in and out) and the max run-to-run difference (0 = deterministic)."""
        d = self.dims()
        rng = np.random.default_rng(seed)
        LE = 100.0 * rng.uniform(-1, 1, d["neq"])
        LI = 100.0 * rng.uniform(-1, 1, d["niq"])
        ref = self.evalKKT(1.0, X, LE, LI)
        t0 = time.perf_counter()
        maxdiff = 0.0
        for _ in range(n):
            r = self.evalKKT(1.0, X, LE, LI)
            maxdiff = max(maxdiff, float(np.max(np.abs(r["KKT"] - ref["KKT"]))) if d["nnz"] else 0.0)
        ms = (time.perf_counter() - t0) * 1e3 / max(n, 1)
        return {"ms_per_evalKKT": ms, "max_run_to_run_diff": maxdiff}
