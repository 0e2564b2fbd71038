// doma_b200_driver — DOMA's command line for the 2D/3D cases with the sweep hot path on a B200.
//
// Mirrors the 2D/3D branches of the reference's src/main.cpp:106-292: same flags (-i, -n, --verbose),
// same parser (parseInputParameters), same meshes (BrickMesh2D/3D with all-vacuum boundaries), same
// solver loops; the only difference is the operator type the solver is instantiated with
// (solver{2,3}d_b200.cpp).  Extra flags:
//   --impl reference     run the unmodified OpenMP solver instead (same binary, for A/B checks)
//   --no-output          do not write the text dumps
//   --fast-dump          steady / eigenvalue runs: final dump through doma_b200_dump_text (same bytes, multi-threaded),
//                        and no per-outer-iteration copy of the flux into the per-cell host vectors
//   --tile TYxTZ         KBA tile override
// One process per GPU: with RANK / WORLD_SIZE / LOCAL_RANK in the environment (torchrun --no-python,
// or any launcher) the ordinates are sharded over the ranks and phi is all-reduced with NCCL;
// rank 0 prints and writes.
#include <chrono>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <thread>

#include <omp.h>

#include "BrickMesh2D.h"
#include "BrickMesh3D.h"
#include "InputParameters.h"
#include "flatten.h"

namespace doma_b200_driver
{
bool runB200Solver3D(BrickMesh3D &, const InputParameters &, bool, const std::string &, double *, long long *);
bool runB200Solver2D(BrickMesh2D &, const InputParameters &, bool, const std::string &, double *, long long *);
bool runReferenceSolver3D(BrickMesh3D &, const InputParameters &, bool, unsigned int, const std::string &);
bool runReferenceSolver2D(BrickMesh2D &, const InputParameters &, bool, unsigned int, const std::string &);

void
bootstrapComm(doma_b200_ctx * ctx, const RankInfo & r)
{
  if (r.world <= 1)
    return;
  const char * port = std::getenv("MASTER_PORT");
  const std::string path = std::string("/tmp/doma_b200_nccl_id.") + (port ? port : "0");
  char id[128];
  if (r.rank == 0)
  {
    check(doma_b200_comm_unique_id(id, 128), nullptr, "doma_b200_comm_unique_id");
    std::ofstream f(path + ".tmp", std::ios::binary);
    f.write(id, 128);
    f.close();
    std::filesystem::rename(path + ".tmp", path);
  }
  else
  {
    for (int tries = 0; !std::filesystem::exists(path) || std::filesystem::file_size(path) < 128; ++tries)
    {
      if (tries > 6000)
      {
        std::cerr << "Error: timed out waiting for " << path << std::endl;
        std::exit(1);
      }
      std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    std::ifstream f(path, std::ios::binary);
    f.read(id, 128);
  }
  check(doma_b200_comm_init(ctx, r.rank, r.world, id, 128), ctx, "doma_b200_comm_init");
  if (r.rank == 0)
    std::filesystem::remove(path);
}
} // namespace doma_b200_driver

using namespace doma_b200_driver;

int
main(int argc, char ** argv)
{
  std::string input, impl = "b200";
  int threads = 1;
  bool verbose = false, output = true;
  for (int i = 1; i < argc; ++i)
  {
    const std::string a = argv[i];
    auto next = [&]() -> std::string {
      if (i + 1 >= argc)
      {
        std::cerr << "missing value after " << a << std::endl;
        std::exit(1);
      }
      return argv[++i];
    };
    if (a == "-i" || a == "--input_file") input = next();
    else if (a == "-n" || a == "--n-threads") threads = std::max(1, std::atoi(next().c_str()));
    else if (a == "--verbose") verbose = true;
    else if (a == "--impl") impl = next();
    else if (a == "--no-output") output = false;
    else if (a == "--fast-dump") config().fast_dump = true;
    else if (a == "--tile")
    {
      const std::string t = next();
      if (std::sscanf(t.c_str(), "%dx%d", &config().tile_y, &config().tile_z) != 2)
      {
        std::cerr << "--tile expects TYxTZ" << std::endl;
        std::exit(1);
      }
    }
    else
    {
      std::cerr << "usage: doma_b200_driver -i case.xml [-n threads] [--verbose] [--impl b200|reference] [--no-output] [--fast-dump] [--tile TYxTZ]"
                << std::endl;
      std::exit(1);
    }
  }
  if (input.empty())
  {
    std::cerr << "-i/--input_file is required" << std::endl;
    std::exit(1);
  }
  config().rank = rankFromEnv();
  const bool root = config().rank.rank == 0;
  if (!root)
    std::cout.setstate(std::ios::failbit); // rank 0 prints

  auto inp_path = std::filesystem::path(input);
  const auto params = parseInputParameters(inp_path.string());
  omp_set_num_threads(threads);
  config().scheme = params._eq_type == EquationType::DD ? DOMA_SCHEME_DD : DOMA_SCHEME_TWDD;
  const std::string base = (output && root) ? (inp_path.parent_path() / inp_path.stem()).string() : std::string("");
  if (params._mode == RunMode::Transient && base.empty() && root)
    std::cerr << "note: transient runs always write per-step dumps (reference behaviour); they go to ./_t*" << std::endl;

  auto start = std::chrono::high_resolution_clock::now();
  bool ok = false;
  double k = 1.0;
  long long launches = 0;
  if (params._num_dims == 3)
  {
    std::array<BoundaryCondition, 6u> bcs;
    bcs.fill(BoundaryCondition::Vacuum);
    // the GPU path never touches the per-thread host scratch: build the mesh with one thread's worth
    BrickMesh3D mesh(params, bcs, impl == "reference" ? threads : 1);
    mesh.validateProps();
    ok = impl == "reference" ? runReferenceSolver3D(mesh, params, verbose, threads, base)
                             : runB200Solver3D(mesh, params, verbose, base, &k, &launches);
  }
  else if (params._num_dims == 2)
  {
    std::array<BoundaryCondition, 4u> bcs;
    bcs.fill(BoundaryCondition::Vacuum);
    BrickMesh2D mesh(params, bcs, impl == "reference" ? threads : 1);
    mesh.validateProps();
    ok = impl == "reference" ? runReferenceSolver2D(mesh, params, verbose, threads, base)
                             : runB200Solver2D(mesh, params, verbose, base, &k, &launches);
  }
  else
  {
    std::cerr << "Only the 2D and 3D solvers run on the B200 path; use the reference's doma for 1D cases." << std::endl;
    std::exit(1);
  }
  auto stop = std::chrono::high_resolution_clock::now();
  const double sec = std::chrono::duration<double>(stop - start).count();
  if (ok)
  {
    switch (params._mode)
    {
      case RunMode::FixedSrc: std::cout << "Fixed source solver finished executing." << std::endl; break;
      case RunMode::Eigen: std::cout << "Eigenvalue solver finished executing." << std::endl; break;
      case RunMode::Transient: std::cout << "Transient solver finished executing." << std::endl; break;
    }
    std::cout << "Time elapsed: " << std::chrono::duration_cast<std::chrono::seconds>(stop - start).count() << " (s)." << std::endl;
  }
  std::cout << std::setprecision(17) << "DRIVER impl=" << impl << " ok=" << (ok ? 1 : 0) << " seconds=" << sec << " k=" << k
            << " kernel_launches=" << launches << std::endl;
  return ok ? 0 : 3;
}
