#include "plan.hpp"

#include <cstring>

namespace ab {

#define AB_CUDA(call)                                                        \
  do {                                                                       \
    cudaError_t e__ = (call);                                                \
    if (e__ != cudaSuccess) {                                                \
      last_error = std::string(#call) + ": " + cudaGetErrorString(e__);      \
      return (int) e__;                                                      \
    }                                                                        \
  } while (0)

template<class T>
T* Plan::dalloc(size_t n) {
  void* p = nullptr;
  if (n == 0) n = 1;
  if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) return nullptr;
  owned.push_back(p);
  return (T*) p;
}
template<class T>
T* Plan::dupload(const std::vector<T>& v) {
  T* p = dalloc<T>(v.size());
  if (p && !v.empty()) {
    if (cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) return nullptr;
  }
  return p;
}

Plan::~Plan() {
  for (cudaEvent_t e: prof_events) cudaEventDestroy(e);
  for (void* p: owned) cudaFree(p);
  if (stream) cudaStreamDestroy(stream);
}

// Largest application range around the middle of the function over which every index map is affine in the application
// number (true for all interior mesh segments of a phase: the transcription and the CSR pattern are periodic there).
static AffineMaps find_affine(const int* V, int ir, const int* C, int orr, const int* FD, const int* KD, int E, int napp) {
  AffineMaps am;
  if (napp < 8) return am;
  const long long mid = napp / 2;
  struct Col {
    const int* p;
    int w;
    std::vector<int>*base, *stride;
  };
  std::vector<Col> cols;
  cols.push_back({V, ir, &am.vbase, &am.vstride});
  if (C) cols.push_back({C, orr, &am.cbase, &am.cstride});
  if (FD) cols.push_back({FD, orr, &am.fbase, &am.fstride});
  cols.push_back({KD, E, &am.kbase, &am.kstride});
  for (Col& c: cols) {
    c.base->resize(c.w);
    c.stride->resize(c.w);
    for (int q = 0; q < c.w; q++) {
      const long long st = (long long) c.p[(mid + 1) * c.w + q] - c.p[mid * c.w + q];
      if (st >= (1 << 23) || st <= -(1 << 23)) return AffineMaps();
      (*c.stride)[q] = (int) st;
    }
  }
  auto row_ok = [&](long long a) {
    for (const Col& c: cols)
      for (int q = 0; q < c.w; q++)
        if ((long long) c.p[a * c.w + q] != (long long) c.p[mid * c.w + q] + (a - mid) * (*c.stride)[q]) return false;
    return true;
  };
  long long a0 = mid, a1 = mid + 1;
  while (a0 > 0 && row_ok(a0 - 1)) a0--;
  while (a1 < napp && row_ok(a1)) a1++;
  if (a1 - a0 < 4) return AffineMaps();
  for (Col& c: cols)
    for (int q = 0; q < c.w; q++) (*c.base)[q] = c.p[a0 * c.w + q];
  am.a0 = (int) a0;
  am.a1 = (int) a1;
  return am;
}

static bool up_gather(Plan& P, const GatherList gl[2], DevGather& dg) {
  for (int w = 0; w < 2; w++) {
    dg.dev[w].ntargets = gl[w].ntargets();
    long long* ptr = P.dupload(gl[w].ptr);
    int* src = P.dupload(gl[w].src);
    int* tgt = P.dupload(gl[w].target);
    if (!ptr || !src || !tgt) return false;
    dg.dev[w].ptr = ptr;
    dg.dev[w].src = src;
    dg.dev[w].target = tgt;
  }
  return true;
}

int Plan::build(int dev) {
  if (!prob.analyzed) prob.analyze();
  device = dev;
  AB_CUDA(cudaSetDevice(dev));
  AB_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  int* d_kkt_dest = dupload(prob.kkt_dest);
  int* d_f_dest = dupload(prob.f_dest);
  if (!d_kkt_dest || !d_f_dest) {
    last_error = "out of device memory for the destination maps";
    return (int) cudaErrorMemoryAllocation;
  }
  fns.clear();
  progs.clear();
  for (int fi: prob.order) {
    const FuncRec& fr = prob.funcs[fi];
    const Tape& t = prob.tapes.at(fr.tape);
    if (t.nin != fr.ir || t.nout != fr.orows || !t.fn) {
      last_error = "tape/function size mismatch in " + fr.name;
      return (int) cudaErrorInvalidValue;
    }
    FnDev f;
    std::memset(&f, 0, sizeof(f));
    // device programs: one per evaluation mode this role can be asked for
    std::array<ProgDev, 5> pd;
    std::memset(pd.data(), 0, sizeof(ProgDev) * 5);
    AffineMaps am;
    if (fr.nappl >= 8)
      am = find_affine(fr.vindex.data(), fr.ir, fr.role ? fr.cindex.data() : nullptr, fr.orows,
                       fr.role ? prob.f_dest.data() + prob.fdest_base[fi] : nullptr, prob.kkt_dest.data() + prob.trip_base[fi],
                       fr.triplets_per_app(), fr.nappl);
    try {
      ScalarFn work = *t.fn;
      for (int mode = 0; mode < 5; mode++) {
        if (fr.role == ROLE_OBJ && (mode == FM_JAC || mode == FM_JAC_GRAD)) continue;
        ProgramSpec sp;
        sp.is_objective = fr.role == ROLE_OBJ;
        sp.linear = fr.linear;
        sp.want_g = (mode == FM_GRAD || mode == FM_JAC_GRAD || mode == FM_ALL);
        sp.want_j = (mode >= FM_JAC);
        sp.want_h = (mode == FM_ALL);
        Program P = emit_program(work, sp);
        LinkedProgram L = link_program(P, 220u * 1024u, &am);
        std::vector<uint4> code(L.code.size());
        std::memcpy(code.data(), L.code.data(), L.code.size() * sizeof(PInstr));
        pd[mode].code = dupload(code);
        pd[mode].consts = dupload(P.consts);
        if (!pd[mode].code || !pd[mode].consts) {
          last_error = "out of device memory for the program of " + fr.name;
          return (int) cudaErrorMemoryAllocation;
        }
        pd[mode].nconst = (int) P.consts.size();
        pd[mode].nslots = L.nslots;
        pd[mode].nthreads = L.nthreads;
        pd[mode].nbundles = P.bundles();
        pd[mode].smem_bytes = L.smem_bytes;
      }
    } catch (const std::exception& e) {
      last_error = std::string(e.what()) + " (" + fr.name + ")";
      return (int) cudaErrorInvalidValue;
    }
    progs.push_back(pd);
    f.aff_begin = am.a0;
    f.aff_end = am.a1;
    f.ir = fr.ir;
    f.orows = fr.orows;
    f.nappconst = fr.nappconst;
    f.role = fr.role;
    f.linear = fr.linear ? 1 : 0;
    f.app_begin = 0;
    f.app_end = fr.nappl;
    f.E = fr.triplets_per_app();
    f.vindex = dupload(fr.vindex);
    f.cindex = dupload(fr.cindex);
    f.appconst = dupload(fr.appconst);
    if (!f.vindex || !f.cindex || !f.appconst) {
      last_error = "out of device memory for index maps of " + fr.name;
      return (int) cudaErrorMemoryAllocation;
    }
    f.kkt_dest = d_kkt_dest + prob.trip_base[fi];
    f.f_dest = d_f_dest + prob.fdest_base[fi];
    f.g_stage = prob.g_stage[fi];
    f.v_stage = prob.v_stage[fi];
    fns.push_back(f);
  }
  d_X = dalloc<double>(prob.nvars);
  d_LE = dalloc<double>(prob.neq);
  d_LI = dalloc<double>(prob.niq);
  d_kkt = dalloc<double>((size_t) prob.nnz());
  d_fxe = dalloc<double>(prob.neq);
  d_fxi = dalloc<double>(prob.niq);
  d_pgx = dalloc<double>(prob.nvars);
  d_agx = dalloc<double>(prob.nvars);
  d_val = dalloc<double>(1);
  d_stage = dalloc<double>((size_t) prob.stage_size);
  if (!d_X || !d_LE || !d_LI || !d_kkt || !d_fxe || !d_fxi || !d_pgx || !d_agx || !d_val || !d_stage) {
    last_error = "out of device memory for the value arrays";
    return (int) cudaErrorMemoryAllocation;
  }
  AB_CUDA(cudaMemset(d_stage, 0, sizeof(double) * (size_t) std::max<long long>(prob.stage_size, 1)));
  AB_CUDA(cudaMemset(d_kkt, 0, sizeof(double) * (size_t) std::max<long long>(prob.nnz(), 1)));
  if (!up_gather(*this, prob.g_kkt, gk) || !up_gather(*this, prob.g_pgx, gp) || !up_gather(*this, prob.g_agx, ga)
      || !up_gather(*this, prob.g_fxe, ge) || !up_gather(*this, prob.g_fxi, gi) || !up_gather(*this, prob.g_val, gv)) {
    last_error = "out of device memory for the gather lists";
    return (int) cudaErrorMemoryAllocation;
  }
  AB_CUDA(cudaDeviceSynchronize());
  return 0;
}

int Plan::set_solver_coeffs(const double* host) {
  AB_CUDA(cudaSetDevice(device));
  AB_CUDA(cudaMemcpyAsync(d_stage + prob.stage_solver, host, sizeof(double) * (size_t) prob.n_solver,
                          cudaMemcpyHostToDevice, stream));
  AB_CUDA(cudaStreamSynchronize(stream));
  return 0;
}

int Plan::prof_begin(const std::string& label) {
  if (!profiling) return 0;
  const size_t k = prof_labels.size();
  while (prof_events.size() < 2 * (k + 1)) {
    cudaEvent_t e;
    AB_CUDA(cudaEventCreate(&e));
    prof_events.push_back(e);
  }
  prof_labels.push_back(label);
  AB_CUDA(cudaEventRecord(prof_events[2 * k], stream));
  return 0;
}
int Plan::prof_end() {
  if (!profiling) return 0;
  AB_CUDA(cudaEventRecord(prof_events[2 * (prof_labels.size() - 1) + 1], stream));
  return 0;
}
int Plan::profile_read(float* ms, int cap) {
  AB_CUDA(cudaSetDevice(device));
  AB_CUDA(cudaStreamSynchronize(stream));
  const int n = (int) prof_labels.size();
  for (int k = 0; k < n && k < cap; k++) AB_CUDA(cudaEventElapsedTime(&ms[k], prof_events[2 * k], prof_events[2 * k + 1]));
  return n;
}

int Plan::eval(int mode, double obj_scale) {
  AB_CUDA(cudaSetDevice(device));
  int obj_mode = -1, con_mode = -1;
  bool out_kkt = false, out_pgx = false, out_agx = false, out_fx = false, out_val = false;
  switch (mode) {
    case EM_KKT: obj_mode = FM_ALL; con_mode = FM_ALL; out_kkt = out_pgx = out_agx = out_fx = out_val = true; break;
    case EM_KKTNO: con_mode = FM_ALL; out_kkt = out_agx = out_fx = true; break;
    case EM_SOE: con_mode = FM_JAC; out_kkt = out_fx = true; break;
    case EM_AUG: obj_mode = FM_GRAD; con_mode = FM_JAC_GRAD; out_kkt = out_pgx = out_agx = out_fx = out_val = true; break;
    case EM_RHS: obj_mode = FM_GRAD; con_mode = FM_GRAD; out_pgx = out_agx = out_fx = out_val = true; break;
    case EM_OGC: obj_mode = FM_GRAD; con_mode = FM_VALUES; out_pgx = out_fx = out_val = true; break;
    case EM_OCC: obj_mode = FM_VALUES; con_mode = FM_VALUES; out_fx = out_val = true; break;
    case EM_OBJ: obj_mode = FM_VALUES; out_val = true; break;
    default: last_error = "unknown eval mode"; return (int) cudaErrorInvalidValue;
  }
  launches = 0;
  if (profiling != 2) prof_labels.clear();
  if (out_kkt && mode != EM_KKT) {
    // partial matrix modes: entries this mode does not compute must read as zero
    AB_CUDA(cudaMemsetAsync(d_kkt, 0, sizeof(double) * (size_t) prob.nnz(), stream));
    AB_CUDA(cudaMemsetAsync(d_stage, 0, sizeof(double) * (size_t) prob.stage_solver, stream));
  }
  for (const FnDev& f: fns) {
    int fm = (f.role == ROLE_OBJ) ? obj_mode : con_mode;
    if (fm < 0) continue;
    EvalIO io;
    io.X = d_X;
    io.L = (f.role == ROLE_EQ) ? d_LE : d_LI;
    io.obj_scale = obj_scale;
    io.kkt = d_kkt;
    io.fx = (f.role == ROLE_EQ) ? d_fxe : d_fxi;
    io.stage = d_stage;
    if (int rc = prof_begin(prob.funcs[prob.order[&f - fns.data()]].name)) return rc;
    FnDev fd = f;
    fd.prog = progs[&f - fns.data()][fm];
    if (fd.prog.nthreads == 0) {
      last_error = "no device program for this mode";
      return (int) cudaErrorInvalidValue;
    }
    AB_CUDA(launch_eval(fd, stream, io));
    if (int rc = prof_end()) return rc;
    launches++;
  }
  auto G = [&](DevGather& g, double* out, const char* label) -> int {
    const int n = (g.dev[0].ntargets > 0) + (g.dev[1].ntargets > 0);
    if (n == 0) return 0;
    if (int rc = prof_begin(label)) return rc;
    AB_CUDA(launch_gather(g.dev[0], g.dev[1], d_stage, out, stream));
    launches += n;
    return prof_end();
  };
  int rc = 0;
  if (out_kkt && (rc = G(gk, d_kkt, "gather:KKT"))) return rc;
  if (out_pgx && (rc = G(gp, d_pgx, "gather:PGX"))) return rc;
  if (out_agx && (rc = G(ga, d_agx, "gather:AGX"))) return rc;
  if (out_fx && ((rc = G(ge, d_fxe, "gather:FXE")) || (rc = G(gi, d_fxi, "gather:FXI")))) return rc;
  if (out_val && (rc = G(gv, d_val, "gather:val"))) return rc;
  return 0;
}

}  // namespace ab
