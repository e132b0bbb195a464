// A C++ client of include/asset_b200.h that knows nothing else of this repository: what the reference-side adapter
//  does, reduced to one VanDerPol LGL3 phase.
//   * without a GPU: builds the expression graph, transcribes the phase, analyses the sparsity, checks the pattern's
//     invariants and that plan creation fails with AB_ERR_NO_DEVICE (there is no CPU fallback);
//   * with a GPU: evaluates evalKKT through ab_plan_eval_host and checks A^T lambda == AGX through the CSR values
//     (the reference's own derivative self-check, asset_asrl/test/test_VectorFunctions/__init__.py:34-67).
// Exit code 0 on success; prints one line "cabi client ok ...".
#include <cmath>
#include <cstdio>
#include <vector>

#include "asset_b200.h"

#define CHECK(call)                                                                   \
  do {                                                                                \
    long long rc_ = (long long) (call);                                               \
    if (rc_ < 0) {                                                                    \
      std::fprintf(stderr, "%s failed (%lld): %s\n", #call, rc_, ab_last_error());    \
      return 1;                                                                       \
    }                                                                                 \
  } while (0)

static int node(ab_graph* g, int kind, int ir, int orows, std::vector<int> ch, std::vector<double> d = {}, std::vector<int> ip = {},
                int raw = 0) {
  return ab_graph_node(g, kind, ir, orows, ch.data(), (int) ch.size(), d.data(), (int) d.size(), ip.data(), (int) ip.size(), -1, raw);
}

int main() {
  if (ab_version() < 100) return 1;
  ab_graph* g = ab_graph_create();
  // VanDerPol: x0' = x1, x1' = (1 - x0^2) x1 - x0 + u; arguments [x0, x1, t, u]
  const int ir = 4;
  const int args = node(g, AB_VK_ARGS, ir, ir, {}, {}, {}, 1);
  auto elem = [&](int i) { return node(g, AB_VK_SEGMENT, ir, 1, {args}, {}, {i, 1}, 1); };
  const int x0 = elem(0), x1 = elem(1), u = elem(3);
  const int x0sq = node(g, AB_VK_CWISEPROD, ir, 1, {x0, x0});
  const int one_minus = node(g, AB_VK_VECMINUS, ir, 1, {x0sq}, {1.0});
  const int damp = node(g, AB_VK_CWISEPROD, ir, 1, {one_minus, x1});
  const int rhs1 = node(g, AB_VK_SUM, ir, 1, {node(g, AB_VK_DIFF, ir, 1, {damp, x0}), u});
  const int ode = node(g, AB_VK_STACK, ir, 2, {x1, rhs1});
  CHECK(ode);
  const int nseg = 40;
  ab_phase* ph = ab_phase_create(g, ode, 2, 1, 0, AB_TM_LGL3, nseg);
  if (!ph) {
    std::fprintf(stderr, "ab_phase_create: %s\n", ab_last_error());
    return 1;
  }
  // u in [-2, 2] along the path: stack{-2 - u, u - 2}
  const int a1 = node(g, AB_VK_ARGS, 1, 1, {}, {}, {}, 1);
  const int lo = node(g, AB_VK_VECMINUS, 1, 1, {a1}, {-2.0}, {}, 1), hi = node(g, AB_VK_ADDVEC, 1, 1, {a1}, {-2.0}, {}, 1);
  const int bound = node(g, AB_VK_STACK, 1, 2, {lo, hi}, {}, {}, 1);
  const int uvar[1] = {3};
  CHECK(ab_phase_add_function(ph, AB_ROLE_INEQUALITY, bound, AB_REGION_PATH, uvar, 1, nullptr, 0, nullptr, 0, nullptr, 0));
  // integral of u^2
  const int usq = node(g, AB_VK_CWISEPROD, 1, 1, {a1, a1});
  CHECK(ab_phase_add_function(ph, AB_ROLE_INTEGRAL_OBJECTIVE, usq, AB_REGION_PATH, uvar, 1, nullptr, 0, nullptr, 0, nullptr, 0));
  ab_problem* p = ab_problem_create(0, 0, 0);
  int sizes[3];
  CHECK(ab_phase_transcribe(ph, p, 0, 0, 0, sizes));
  CHECK(ab_problem_set_dims(p, sizes[0], sizes[1], sizes[2]));
  CHECK(ab_problem_analyze(p));
  long long d[10];
  CHECK(ab_problem_dims(p, d));
  const long long nvars = d[0], neq = d[1], niq = d[2], kktdim = d[3], nnz = d[4];
  std::vector<int> indptr((size_t) kktdim + 1), indices((size_t) nnz);
  CHECK(ab_problem_pattern(p, indptr.data(), indices.data()));
  for (long long r = 0; r < kktdim; r++) {
    if (indptr[r + 1] <= indptr[r] || indices[indptr[r]] != r) return 2;  // every row starts with its diagonal (upper triangle)
    for (int q = indptr[r] + 1; q < indptr[r + 1]; q++)
      if (indices[q] <= indices[q - 1]) return 2;
  }
  ab_plan* plan = nullptr;
  if (ab_device_count() <= 0) {
    if (ab_plan_create(p, 0, &plan) != AB_ERR_NO_DEVICE || plan) return 3;
    std::printf("cabi client ok (no device): nvars %lld neq %lld niq %lld nnz %lld\n", nvars, neq, niq, nnz);
    return 0;
  }
  CHECK(ab_plan_create(p, 0, &plan));
  std::vector<double> X((size_t) nvars), LE((size_t) neq), LI((size_t) niq), sc((size_t) d[6], 0.0);
  for (long long i = 0; i < nvars; i++) X[(size_t) i] = 0.3 + 0.01 * std::sin(0.7 * i);
  for (long long i = 0; i < nvars; i += 4) X[(size_t) i + 2 < X.size() ? (size_t) i + 2 : 0] = 0.05 * (double) (i / 4);  // time increases
  for (long long i = 0; i < neq; i++) LE[(size_t) i] = std::cos(0.3 * i);
  for (long long i = 0; i < niq; i++) LI[(size_t) i] = 0.5 + 0.25 * std::sin(0.2 * i);
  CHECK(ab_plan_set_solver_coeffs(plan, sc.data()));
  double val = 0;
  std::vector<double> pgx((size_t) nvars), agx((size_t) nvars), fxe((size_t) neq), fxi((size_t) niq), kkt((size_t) nnz);
  if (int rc = ab_plan_eval_host(plan, AB_EVAL_KKT, 1.0, X.data(), LE.data(), LI.data(), &val, pgx.data(), agx.data(), fxe.data(),
                                 fxi.data(), kkt.data())) {
    std::fprintf(stderr, "ab_plan_eval_host (%d): %s\n", rc, ab_last_error());
    return 1;
  }
  // AGX = J^T lambda: the Jacobian sits in the columns >= nvars + niq of the variable rows (upper triangle)
  double worst = 0, scale = 0;
  for (long long r = 0; r < nvars; r++) {
    double s = 0;
    for (int q = indptr[r]; q < indptr[r + 1]; q++) {
      const long long c = indices[q];
      if (c >= nvars + niq + neq) s += kkt[(size_t) q] * LI[(size_t) (c - nvars - niq - neq)];
      else if (c >= nvars + niq) s += kkt[(size_t) q] * LE[(size_t) (c - nvars - niq)];
    }
    worst = std::fmax(worst, std::fabs(s - agx[(size_t) r]));
    scale = std::fmax(scale, std::fabs(agx[(size_t) r]));
  }
  if (!(worst <= 1e-12 * std::fmax(scale, 1.0)) || ab_plan_launches(plan) <= 0) {
    std::fprintf(stderr, "J^T lambda differs from AGX by %g (scale %g)\n", worst, scale);
    return 4;
  }
  std::printf("cabi client ok (device): nvars %lld nnz %lld, |J^T L - AGX| = %.2e, %lld launches, val %.6f\n", nvars, nnz, worst,
              ab_plan_launches(plan), val);
  ab_plan_destroy(plan);
  ab_problem_destroy(p);
  ab_phase_destroy(ph);
  ab_graph_destroy(g);
  return 0;
}
