// solver3d_b200.cpp — binds the C ABI under the reference's OWN 3D driver loops (see seam.inc).
//
// The operator type B200Op3D is the T of TransportSolver3D<T> (include/TransportSolver3D.h:12-16,112-113);
// its per-cell solve() is never called because sweep() itself is specialised.  The unmodified
// TWDDTransportSolver3D is instantiated by the included reference file as well, so the CPU original
// lives in the same binary (main.cpp --oracle, an A/B aid).
#include <chrono>

#include "TransportSolver3D.h"

#include "flatten.h"

struct B200Op3D
{
  // the T concept's per-cell call (include/TWDiamondDifference3D.h:15-22); unused: sweep() is replaced
  void solve(CartesianCell3D &, const double &, const double &, const double &, const double &, unsigned int, unsigned int,
             CertesianFaceSide, CertesianFaceSide, CertesianFaceSide, CertesianFaceSide, CertesianFaceSide, CertesianFaceSide,
             unsigned int, RunMode = RunMode::FixedSrc, const double & = 0.0) const
  {
  }
  ~B200Op3D()
  {
    if (ctx)
      doma_b200_destroy(ctx);
  }
  mutable doma_b200_ctx * ctx = nullptr; // _eq_system is a const member of the solver
  mutable std::vector<double> staging, history;
};

#define SEAM_DIMS 3
#define SEAM_SOLVER TransportSolver3D
#define SEAM_MESH BrickMesh3D
#define SEAM_OP B200Op3D
#define SEAM_REF_CPP "TransportSolver3D.cpp"
#define SEAM_TWDD TWDDTransportSolver3D
#define SEAM_DD DDTransportSolver3D
#define SEAM_RUN runB200Solver3D
#define SEAM_RUN_ORACLE runReferenceSolver3D
#include "seam.inc"
