// solver2d_b200.cpp — binds the C ABI under the reference's OWN 2D driver loops (see seam.inc).
//
// The operator type B200Op2D is the T of TransportSolver2D<T> (include/TransportSolver2D.h:12-16,112-113);
// its per-cell solve() is never called because sweep() itself is specialised.
#include <chrono>

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
  mutable std::vector<double> staging, history;
};

#define SEAM_DIMS 2
#define SEAM_SOLVER TransportSolver2D
#define SEAM_MESH BrickMesh2D
#define SEAM_OP B200Op2D
#define SEAM_REF_CPP "TransportSolver2D.cpp"
#define SEAM_TWDD TWDDTransportSolver2D
#define SEAM_DD DDTransportSolver2D
#define SEAM_RUN runB200Solver2D
#define SEAM_RUN_ORACLE runReferenceSolver2D
#include "seam.inc"
