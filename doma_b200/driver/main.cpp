// doma_b200_driver — DOMA's command line for the 2D/3D cases with the sweep hot path on a B200.
//
// Mirrors the 2D/3D branches of the reference's src/main.cpp:106-292: same flags (-i, -n, --verbose),
// same parser (parseInputParameters), same meshes (BrickMesh2D/3D with all-vacuum boundaries), same
// solver loops; the only difference is the operator type the solver is instantiated with
// (solver{2,3}d_b200.cpp, seam.inc).  Extra flags:
//   --no-output          do not write the text dumps
//   --fast-dump          steady / eigenvalue runs: final dump through doma_b200_dump_text (same bytes, multi-threaded),
//                        and no per-outer-iteration copy of the flux into the per-cell host vectors
//   --flat / --no-flat   flat mesh builder: widths and block ids straight from the parsed input, the solver
//                        gets a stub mesh of one cell per block; implies --fast-dump.  Default: on above 2 M cells in
//                        steady / eigenvalue runs (the reference's per-cell construction takes tens of seconds there)
//   --check-flat         compare the flat builder with the mesh the reference built, bit for bit, and exit (no GPU needed)
//   --dump-raw DIR       final flux (and precursors) as raw doubles + meta.json, full precision
//   --separate-si        the reference's own sweep / residual / update sequence instead of doma_b200_source_iteration
//   --check              poison the sweep scratch before every sweep and verify phi afterwards (doma_b200_set_check)
//   --tile TYxTZ         KBA tile override
// One process per GPU: with RANK / WORLD_SIZE / LOCAL_RANK in the environment (torchrun --no-python,
// or any launcher) the ordinates are sharded over the ranks and phi is summed with NCCL;
// rank 0 prints and writes.
#include <arpa/inet.h>
#include <netinet/in.h>
#include <sys/socket.h>
#include <unistd.h>

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

// The 128-byte NCCL unique id goes from rank 0 to the other ranks over a TCP connection to
// MASTER_ADDR:(MASTER_PORT + 1) (DOMA_B200_COMM_PORT overrides): nothing is left behind that a later
// launch could mistake for its own id, and a rank that cannot reach rank 0 within a minute says so.
static void
die(const std::string & what)
{
  std::cerr << "Error: " << what << ": " << std::strerror(errno) << std::endl;
  std::exit(1);
}

static bool
transferAll(int fd, char * buf, std::size_t n, bool send_it)
{
  std::size_t done = 0;
  while (done < n)
  {
    const ssize_t r = send_it ? ::send(fd, buf + done, n - done, MSG_NOSIGNAL) : ::recv(fd, buf + done, n - done, 0);
    if (r <= 0)
      return false;
    done += static_cast<std::size_t>(r);
  }
  return true;
}

void
bootstrapComm(doma_b200_ctx * ctx, const RankInfo & r)
{
  if (r.world <= 1)
    return;
  const char * addr_env = std::getenv("MASTER_ADDR");
  const char * port_env = std::getenv("DOMA_B200_COMM_PORT");
  const char * mport = std::getenv("MASTER_PORT");
  const int port = port_env ? std::atoi(port_env) : (mport ? std::atoi(mport) + 1 : 29601);
  sockaddr_in sa{};
  sa.sin_family = AF_INET;
  sa.sin_port = htons(static_cast<uint16_t>(port));
  char id[128];
  if (r.rank == 0)
  {
    check(doma_b200_comm_unique_id(id, 128), nullptr, "doma_b200_comm_unique_id");
    const int ls = ::socket(AF_INET, SOCK_STREAM, 0);
    if (ls < 0)
      die("socket");
    const int one = 1;
    ::setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sa.sin_addr.s_addr = htonl(INADDR_ANY);
    if (::bind(ls, reinterpret_cast<sockaddr *>(&sa), sizeof(sa)) != 0 || ::listen(ls, r.world) != 0)
      die("cannot listen on port " + std::to_string(port) + " for the NCCL id exchange");
    for (int peer = 1; peer < r.world; ++peer)
    {
      const int fd = ::accept(ls, nullptr, nullptr);
      if (fd < 0 || !transferAll(fd, id, 128, true))
        die("sending the NCCL id");
      ::close(fd);
    }
    ::close(ls);
  }
  else
  {
    if (::inet_pton(AF_INET, (addr_env && *addr_env) ? addr_env : "127.0.0.1", &sa.sin_addr) != 1)
      sa.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
    bool got = false;
    for (int tries = 0; tries < 600 && !got; ++tries) // rank 0 may still be creating its context: retry for a minute
    {
      const int fd = ::socket(AF_INET, SOCK_STREAM, 0);
      if (fd < 0)
        die("socket");
      if (::connect(fd, reinterpret_cast<sockaddr *>(&sa), sizeof(sa)) == 0)
        got = transferAll(fd, id, 128, false);
      ::close(fd);
      if (!got)
        std::this_thread::sleep_for(std::chrono::milliseconds(100));
    }
    if (!got)
    {
      std::cerr << "Error: rank " << r.rank << " could not fetch the NCCL id from rank 0 on port " << port << std::endl;
      std::exit(1);
    }
  }
  check(doma_b200_comm_init(ctx, r.rank, r.world, id, 128), ctx, "doma_b200_comm_init");
}
} // namespace doma_b200_driver

using namespace doma_b200_driver;

int
main(int argc, char ** argv)
{
  std::string input;
  int threads = 1, flat_opt = -1;
  bool verbose = false, output = true, oracle = false;
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
    else if (a == "--oracle") oracle = true; // A/B aid: the unmodified OpenMP solver from the same binary (not a fallback)
    else if (a == "--no-output") output = false;
    else if (a == "--fast-dump") config().fast_dump = true;
    else if (a == "--flat") flat_opt = 1;
    else if (a == "--no-flat") flat_opt = 0;
    else if (a == "--check-flat") config().check_flat = true;
    else if (a == "--separate-si") config().separate_si = true;
    else if (a == "--check") config().poison = true;
    else if (a == "--dump-raw") config().raw_dir = next();
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
      std::cerr << "usage: doma_b200_driver -i case.xml [-n threads] [--verbose] [--no-output] [--fast-dump] [--flat|--no-flat] "
                   "[--check-flat] [--dump-raw DIR] [--separate-si] [--check] [--tile TYxTZ]"
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
  if (!config().raw_dir.empty() && root)
    std::filesystem::create_directories(config().raw_dir);

  auto inp_path = std::filesystem::path(input);
  const auto t_parse = std::chrono::steady_clock::now();
  const auto params = parseInputParameters(inp_path.string());
  omp_set_num_threads(threads);
  config().scheme = params._eq_type == EquationType::DD ? DOMA_SCHEME_DD : DOMA_SCHEME_TWDD;
  config().output_enabled = output;
  const std::string base = (output && root) ? (inp_path.parent_path() / inp_path.stem()).string() : std::string("");
  if (params._mode == RunMode::Transient && base.empty() && root)
    std::cerr << "note: transient runs always write per-step dumps (reference behaviour); they go to ./_t*" << std::endl;
  if (params._num_dims != 2 && params._num_dims != 3)
  {
    std::cerr << "Only the 2D and 3D solvers run on the B200 path; use the reference's doma for 1D cases." << std::endl;
    std::exit(1);
  }

  // the flat builder always runs (it is cheap); whether the solver also gets the reference's full mesh depends on the mode
  config().flat.setGeometryFromParams(params);
  const std::size_t n_cells = config().flat.block_of_cell.size();
  const bool steady = params._mode != RunMode::Transient;
  config().flat_mesh = !oracle && !config().check_flat && steady && (flat_opt == 1 || (flat_opt < 0 && n_cells > 2000000u));
  if (flat_opt == 1 && !steady && root)
    std::cerr << "note: --flat applies to steady and eigenvalue runs only (the transient loop dumps the mesh every step)" << std::endl;
  InputParameters mesh_params = params;
  if (config().flat_mesh)
  {
    // stub mesh: one cell per block keeps every block id and material lookup the solver's loops could ask for
    for (auto & v : mesh_params._x_intervals) v = 1u;
    for (auto & v : mesh_params._y_intervals) v = 1u;
    for (auto & v : mesh_params._z_intervals) v = 1u;
  }

  auto start = std::chrono::high_resolution_clock::now();
  bool ok = false;
  double k = 1.0;
  long long launches = 0;
  double mesh_seconds = 0.0;
  if (params._num_dims == 3)
  {
    std::array<BoundaryCondition, 6u> bcs;
    bcs.fill(BoundaryCondition::Vacuum);
    // the GPU path never touches the per-thread host scratch: build the mesh with one thread's worth
    BrickMesh3D mesh(mesh_params, bcs, oracle ? threads : 1);
    mesh.validateProps();
    mesh_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_parse).count();
    ok = oracle ? runReferenceSolver3D(mesh, params, verbose, threads, base) : runB200Solver3D(mesh, params, verbose, base, &k, &launches);
  }
  else
  {
    std::array<BoundaryCondition, 4u> bcs;
    bcs.fill(BoundaryCondition::Vacuum);
    BrickMesh2D mesh(mesh_params, bcs, oracle ? threads : 1);
    mesh.validateProps();
    mesh_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_parse).count();
    ok = oracle ? runReferenceSolver2D(mesh, params, verbose, threads, base) : runB200Solver2D(mesh, params, verbose, base, &k, &launches);
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
  std::cout << std::setprecision(17) << "DRIVER impl=" << (oracle ? "oracle" : "b200") << " ok=" << (ok ? 1 : 0) << " seconds=" << sec
            << " k=" << k << " kernel_launches=" << launches << " cells=" << n_cells << " flat_mesh=" << (config().flat_mesh ? 1 : 0)
            << " setup_seconds=" << mesh_seconds << " solve_seconds=" << config().solve_seconds << std::endl;
  return ok ? 0 : 3;
}
