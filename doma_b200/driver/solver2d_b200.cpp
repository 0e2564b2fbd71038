// solver2d_b200.cpp — binds the C ABI under the reference's OWN 2D driver loops.
//
// The reference has one extension point: the template parameter T of TransportSolver2D<T>
// (include/TransportSolver2D.h:12-16,112-113).  This translation unit
//   1. defines a new operator type B200Op2D (its per-cell solve() is never called),
//   2. declares explicit specialisations of the hot-path members for TransportSolver2D<B200Op2D>,
//   3. #includes the reference's src/TransportSolver2D.cpp VERBATIM (compiled from where it lies,
//      -I $DOMA_REF/src), so solveFixedSource / solveEigenvalue / solveTransient / sourceIteration
//      are the reference's code, not a copy,
//   4. defines the specialisations: each is ONE call into include/doma_b200.h,
//   5. instantiates the class.
// No reference file is edited.  The unmodified TWDDTransportSolver2D is instantiated by the included
// file as well, so the CPU oracle lives in the same binary (main.cpp --impl reference).
#include "TransportSolver2D.h"

#include "flatten.h"

struct B200Op2D
{
  // the T concept's per-cell call (include/TWDiamondDifference2D.h:15-21); unused: sweep() is replaced
  void solve(CartesianCell2D &, const double &, const double &, const double &, const double &, unsigned int, unsigned int,
             CertesianFaceSide, CertesianFaceSide, CertesianFaceSide, CertesianFaceSide, unsigned int,
             RunMode = RunMode::FixedSrc, const double & = 0.0) const
  {
  }
  ~B200Op2D()
  {
    if (ctx)
      doma_b200_destroy(ctx);
  }
  mutable doma_b200_ctx * ctx = nullptr; // _eq_system is a const member of the solver
  mutable std::vector<double> staging;
};

typedef TransportSolver2D<B200Op2D> B200TransportSolver2D;

template <> void B200TransportSolver2D::initializeSolve();
template <> void B200TransportSolver2D::initZeroIC();
template <> bool B200TransportSolver2D::initSteadyIC();
template <> void B200TransportSolver2D::updateStepFluxes();
template <> void B200TransportSolver2D::stepDNPs();
template <> void B200TransportSolver2D::updateMultigroupSource(unsigned int, double);
template <> void B200TransportSolver2D::updateMultigroupSourceEigen(unsigned int);
template <> void B200TransportSolver2D::updateEigenvalue();
template <> void B200TransportSolver2D::updateScatteringSource(unsigned int);
template <> double B200TransportSolver2D::computeScatteringResidual(unsigned int);
template <> double B200TransportSolver2D::computeMGFluxResidual();
template <> void B200TransportSolver2D::sweep(unsigned int);

#include "TransportSolver2D.cpp" // the reference's file, verbatim ($DOMA_REF/src)

using namespace doma_b200_driver;

// initializeSolve (2D:499-529) doubles as the lazy context creation: it is the first member every
// steady / eigen solve calls, and it has _mesh, _angular_quad, _mode, _dt in scope.
static doma_b200_ctx *
ensureContext2D(const B200Op2D & op, BrickMesh2D & mesh_unused, FlatProblem & flat)
{
  (void)mesh_unused;
  if (op.ctx)
    return op.ctx;
  const auto p = flat.view();
  doma_b200_ctx * ctx = nullptr;
  const int rc = doma_b200_create(&p, &ctx);
  if (rc != 0)
  {
    std::fprintf(stderr, "Error: doma_b200_create failed (%d): %s\n", rc, doma_b200_last_error(nullptr));
    std::exit(1);
  }
  if (config().tile_y > 0 && config().tile_z > 0)
    check(doma_b200_set_tile(ctx, 8, config().tile_y, config().tile_z), ctx, "doma_b200_set_tile");
  bootstrapComm(ctx, config().rank);
  op.ctx = ctx;
  config().live_ctx = ctx;
  return ctx;
}

#define B200_CTX()                                                                                               \
  [this]() -> doma_b200_ctx * {                                                                                  \
    if (_eq_system.ctx)                                                                                          \
      return _eq_system.ctx;                                                                                     \
    FlatProblem flat;                                                                                            \
    flat.dims = 2, flat.scheme = config().scheme, flat.mode = static_cast<int>(_mode), flat.device = config().rank.local; \
    flat.dt = _mode == RunMode::Transient ? _dt : 0.0; /* _dt is garbage outside transients */        \
    const unsigned int nx = _mesh._tot_num_x, ny = _mesh._tot_num_y;                                              \
    for (unsigned int i = 0; i < nx; ++i) flat.dx.push_back(_mesh._cells[i]._l_x);                                \
    for (unsigned int j = 0; j < ny; ++j) flat.dy.push_back(_mesh._cells[j * nx]._l_y);                           \
    flat.setMaterials(_mesh._block_mat_info, _mesh._block_step_src, _num_groups);                                \
    flat.mat_of_cell.resize(_mesh._cells.size());                                                                \
    for (std::size_t c = 0; c < _mesh._cells.size(); ++c)                                                        \
      flat.mat_of_cell[c] = flat.materialIndex(_mesh._cells[c]._block_id);                                       \
    flat.setQuadrature(_angular_quad, config().rank);                                                            \
    if (config().fast_dump)                                                                                      \
    {                                                                                                            \
      auto & cf = config();                                                                                      \
      cf.geom_dims = flat.dims, cf.geom_groups = static_cast<int>(_num_groups);                                  \
      cf.geom_dx = flat.dx, cf.geom_dy = flat.dy, cf.geom_dz = flat.dz;                                          \
      cf.geom_block_of_cell.clear();                                                                             \
      for (const auto & cell : _mesh._cells) cf.geom_block_of_cell.push_back(cell._block_id);                    \
    }                                                                                                            \
    return ensureContext2D(_eq_system, _mesh, flat);                                                             \
  }()

template <>
void
B200TransportSolver2D::initializeSolve()
{
  std::cout << "Initializing the solver..." << std::endl;
  // host vectors sized exactly as the generic member does: the mesh dumps read them
  for (auto & cell : _mesh._cells)
  {
    cell._total_scalar_flux.resize(_num_groups, _mode == RunMode::Eigen ? 1.0 : 0.0);
    cell._prev_mg_scalar_flux.resize(_num_groups, 0.0);
    cell._current_iteration_source = 0.0;
    cell._current_scalar_flux = 0.0;
  }
  auto * ctx = B200_CTX();
  check(doma_b200_reset_state(ctx, _mode == RunMode::Eigen ? 1.0 : 0.0), ctx, "doma_b200_reset_state");
}

template <>
void
B200TransportSolver2D::initZeroIC()
{
  for (auto & cell : _mesh._cells)
  {
    const auto & p = cell.getMatProps();
    cell._total_scalar_flux.resize(_num_groups, 0.0);
    cell._prev_mg_scalar_flux.resize(_num_groups, 0.0);
    cell._last_t_scalar_flux.resize(_num_groups, 0.0);
    cell._current_t_dnps.resize(p._num_d_groups, 0.0);
    cell._last_t_dnps.resize(p._num_d_groups, 0.0);
  }
  auto * ctx = B200_CTX();
  check(doma_b200_init_zero_ic(ctx), ctx, "doma_b200_init_zero_ic");
}

template <>
bool
B200TransportSolver2D::initSteadyIC()
{
  // in the reference this path indexes vectors that have not been sized (undefined behaviour)
  std::cerr << "Error: ics=\"steady\" is undefined behaviour in the reference and is not supported." << std::endl;
  std::exit(1);
  return false;
}

template <>
void
B200TransportSolver2D::updateStepFluxes()
{
  auto * ctx = B200_CTX();
  check(doma_b200_update_step_fluxes(ctx), ctx, "doma_b200_update_step_fluxes");
}

template <>
void
B200TransportSolver2D::stepDNPs()
{
  auto * ctx = B200_CTX();
  check(doma_b200_step_dnps(ctx), ctx, "doma_b200_step_dnps");
  // dumpDNPsToTextFile reads cell._current_t_dnps right after this call (2D:125-128)
  auto & buf = _eq_system.staging;
  buf.resize(_mesh._cells.size());
  unsigned int max_d = 0;
  for (const auto & kv : _mesh._block_mat_info)
    max_d = std::max(max_d, kv.second._num_d_groups);
  for (unsigned int d = 0; d < max_d; ++d)
  {
    check(doma_b200_get_array(ctx, DOMA_ARR_DNP_CUR, d, buf.data()), ctx, "doma_b200_get_array");
    for (std::size_t c = 0; c < _mesh._cells.size(); ++c)
      if (_mesh._cells[c]._current_t_dnps.size() > d)
        _mesh._cells[c]._current_t_dnps[d] = _mesh._cells[c]._last_t_dnps[d] = buf[c];
  }
}

template <>
void
B200TransportSolver2D::updateMultigroupSource(unsigned int g, double t)
{
  auto * ctx = B200_CTX();
  check(doma_b200_update_mg_source(ctx, g, t), ctx, "doma_b200_update_mg_source");
}

template <>
void
B200TransportSolver2D::updateMultigroupSourceEigen(unsigned int g)
{
  auto * ctx = B200_CTX();
  check(doma_b200_update_mg_source_eigen(ctx, g, _k), ctx, "doma_b200_update_mg_source_eigen");
}

template <>
void
B200TransportSolver2D::updateEigenvalue()
{
  auto * ctx = B200_CTX();
  double k = _k, k_prev = 0.0;
  check(doma_b200_update_k(ctx, &k, &k_prev), ctx, "doma_b200_update_k");
  _k_prev = k_prev;
  _k = k;
  config().last_k = k;
}

template <>
void
B200TransportSolver2D::updateScatteringSource(unsigned int g)
{
  auto * ctx = B200_CTX();
  check(doma_b200_update_scatter_source(ctx, g), ctx, "doma_b200_update_scatter_source");
}

template <>
double
B200TransportSolver2D::computeScatteringResidual(unsigned int g)
{
  auto * ctx = B200_CTX();
  double r = 0.0;
  check(doma_b200_scatter_residual(ctx, g, &r), ctx, "doma_b200_scatter_residual");
  return r;
}

template <>
double
B200TransportSolver2D::computeMGFluxResidual()
{
  auto * ctx = B200_CTX();
  double r = 0.0;
  check(doma_b200_mg_residual(ctx, &r), ctx, "doma_b200_mg_residual");
  // the loops dump the mesh right after this member (2D:61-75, 292-306): refresh the host copy
  // (not with --fast-dump in steady / eigenvalue runs: the final dump then comes from doma_b200_dump_text)
  if (config().fast_dump && _mode != RunMode::Transient)
    return r;
  auto & buf = _eq_system.staging;
  buf.resize(_mesh._cells.size());
  for (unsigned int g = 0; g < _num_groups; ++g)
  {
    check(doma_b200_get_array(ctx, DOMA_ARR_PHI_TOT, g, buf.data()), ctx, "doma_b200_get_array");
    for (std::size_t c = 0; c < _mesh._cells.size(); ++c)
      _mesh._cells[c]._total_scalar_flux[g] = buf[c];
  }
  return r;
}

template <>
void
B200TransportSolver2D::sweep(unsigned int g)
{
  auto * ctx = B200_CTX();
  check(doma_b200_sweep(ctx, g), ctx, "doma_b200_sweep");
  check(doma_b200_allreduce_phi_cur(ctx), ctx, "doma_b200_allreduce_phi_cur"); // no-op on one rank
}

template class TransportSolver2D<B200Op2D>;

// ---- entry points main.cpp uses (keeps the operator type private to this file) ----
namespace doma_b200_driver
{
bool
runB200Solver2D(BrickMesh2D & mesh, const InputParameters & params, bool verbose, const std::string & out_base, double * k_out,
                long long * launches)
{
  B200TransportSolver2D solver(mesh, params, verbose, 1u);
  bool ok = false;
  const bool fast = config().fast_dump && params._mode != RunMode::Transient && !out_base.empty();
  switch (params._mode)
  {
    case RunMode::FixedSrc: ok = solver.solveFixedSource(fast ? std::string("") : out_base); break;
    case RunMode::Eigen: ok = solver.solveEigenvalue(fast ? std::string("") : out_base); break;
    case RunMode::Transient: ok = solver.solveTransient(out_base); break;
  }
  if (fast && ok && config().live_ctx)
  {
    // the dump the reference's loop would have made on convergence (3D:74-75, 305-306), same bytes
    writeGeometryFiles(config(), out_base);
    check(doma_b200_dump_text(config().live_ctx, out_base.c_str(), 0, 0), config().live_ctx, "doma_b200_dump_text");
  }
  if (k_out)
    *k_out = config().last_k;
  if (launches && config().live_ctx)
  {
    int64_t v = 0;
    doma_b200_get_info(config().live_ctx, DOMA_INFO_KERNEL_LAUNCHES, &v);
    *launches = v;
  }
  config().live_ctx = nullptr; // the solver's operator owns and destroys the context
  return ok;
}

template <typename Solver>
static bool
runRef2D(Solver & solver, const InputParameters & params, const std::string & out_base, double * k_out)
{
  bool ok = false;
  switch (params._mode)
  {
    case RunMode::FixedSrc: ok = solver.solveFixedSource(out_base); break;
    case RunMode::Eigen: ok = solver.solveEigenvalue(out_base); break;
    case RunMode::Transient: ok = solver.solveTransient(out_base); break;
  }
  (void)k_out;
  return ok;
}

bool
runReferenceSolver2D(BrickMesh2D & mesh, const InputParameters & params, bool verbose, unsigned int threads,
                     const std::string & out_base)
{
  if (params._eq_type == EquationType::DD)
  {
    DDTransportSolver2D solver(mesh, params, verbose, threads);
    return runRef2D(solver, params, out_base, nullptr);
  }
  TWDDTransportSolver2D solver(mesh, params, verbose, threads);
  return runRef2D(solver, params, out_base, nullptr);
}
} // namespace doma_b200_driver
