// flatten.h — the reference's problem objects -> the flat, caller-owned doma_b200_problem.
//
// Everything here reads the reference's own containers (InputParameters, MaterialProps,
// SourceStep: include/InputParameters.h:10-117; GCAngularQuadrature tables through
// order/direction/weight: include/GCAngularQuadrature.h:17-19), so the GPU path is fed exactly what
// the reference's driver built.  The per-cell geometry (CartesianCell3D::_l_x/_l_y/_l_z, _block_id)
// is private to the mesh; the seam translation units (solver3d_b200.cpp / solver2d_b200.cpp) read it
// from inside TransportSolver{2,3}D<T> members, which the mesh befriends, and hand it over here.
#pragma once
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "GCAngularQuadrature.h"
#include "InputParameters.h"
#include "doma_b200.h"

namespace doma_b200_driver
{
struct RankInfo
{
  int rank = 0, world = 1, local = 0;
};

inline RankInfo
rankFromEnv()
{
  RankInfo r;
  if (const char * s = std::getenv("RANK")) r.rank = std::atoi(s);
  if (const char * s = std::getenv("WORLD_SIZE")) r.world = std::max(1, std::atoi(s));
  if (const char * s = std::getenv("LOCAL_RANK")) r.local = std::atoi(s);
  else r.local = r.rank;
  return r;
}

struct FlatProblem
{
  int dims = 3, scheme = DOMA_SCHEME_TWDD, mode = DOMA_MODE_FIXED, device = 0;
  double dt = 0.0;
  std::vector<double> dx, dy, dz;
  std::vector<unsigned int> block_of_cell; // XML block id per cell, x fastest (what CartesianCell::_block_id holds)
  std::vector<int32_t> mat_of_cell;        // dense material index per cell (mapMaterials)
  int G = 1, M = 0, D = 0;
  std::vector<unsigned int> block_of_mat;
  std::vector<int32_t> mat_flags, n_delayed, step_type;
  std::vector<double> sigma_t, sigma_s, nu_sigma_f, chi_p, src, inv_v, chi_d, beta, lambda, step_t_insert, step_t_remove, step_src;
  std::vector<int32_t> ord_octant;
  std::vector<double> mu, eta, xi, w;

  // Flat builder: per-cell widths and block ids straight from the parsed input, without constructing
  // the reference's per-cell objects.  Same arithmetic as the mesh constructors (src/BrickMesh3D.cpp:44-81,
  // src/BrickMesh2D.cpp:38-63): every interval i contributes n_i cells of width d_i / n_i (n_i already carries the
  // refinement factor, src/InputParameters.cpp:212-214), and the block id of interval (i,j,k) is repeated over its cells.
  // --check-flat compares the result, bit for bit, with what the mesh built.
  void setGeometryFromParams(const InputParameters & p)
  {
    dims = static_cast<int>(p._num_dims);
    auto widths = [](const std::vector<double> & d, const std::vector<unsigned int> & n, std::vector<double> & out) {
      out.clear();
      for (std::size_t i = 0; i < n.size(); ++i)
        out.insert(out.end(), n[i], d[i] / static_cast<double>(n[i]));
    };
    widths(p._dx, p._x_intervals, dx);
    widths(p._dy, p._y_intervals, dy);
    const std::vector<unsigned int> one(1, 1u);
    const std::vector<unsigned int> & nzv = dims == 3 ? p._z_intervals : one;
    if (dims == 3)
      widths(p._dz, p._z_intervals, dz);
    else
      dz.clear();
    const std::size_t bx = p._x_intervals.size(), by = p._y_intervals.size();
    if (bx * by * nzv.size() != p._blocks.size())
    {
      std::fprintf(stderr, "Error: The blocks vector is not equal to the number of subdivisions!\n");
      std::exit(1);
    }
    block_of_cell.clear();
    block_of_cell.reserve(dx.size() * dy.size() * (dims == 3 ? dz.size() : 1));
    std::vector<unsigned int> row;
    for (std::size_t k = 0; k < nzv.size(); ++k)
      for (unsigned int kk = 0; kk < nzv[k]; ++kk)
        for (std::size_t j = 0; j < by; ++j)
        {
          row.clear();
          for (std::size_t i = 0; i < bx; ++i)
            row.insert(row.end(), p._x_intervals[i], p._blocks[(k * by + j) * bx + i]);
          for (unsigned int jj = 0; jj < p._y_intervals[j]; ++jj)
            block_of_cell.insert(block_of_cell.end(), row.begin(), row.end());
        }
  }

  // block ids -> dense material indices (after setMaterials)
  void mapMaterials()
  {
    mat_of_cell.resize(block_of_cell.size());
    unsigned int last_block = 0;
    int last_index = -1;
    for (std::size_t c = 0; c < block_of_cell.size(); ++c)
    {
      if (last_index < 0 || block_of_cell[c] != last_block)
        last_block = block_of_cell[c], last_index = materialIndex(last_block);
      mat_of_cell[c] = last_index;
    }
  }

  // materials: dense index = rank of the block id (the reference keys an unordered_map by block id)
  void setMaterials(const std::unordered_map<unsigned int, MaterialProps> & mats,
                    const std::unordered_map<unsigned int, SourceStep> & steps, unsigned int groups)
  {
    G = static_cast<int>(groups);
    block_of_mat.clear();
    std::map<unsigned int, const MaterialProps *> sorted;
    for (const auto & kv : mats)
      sorted[kv.first] = &kv.second;
    M = static_cast<int>(sorted.size());
    D = 0;
    for (const auto & kv : sorted)
      D = std::max(D, static_cast<int>(kv.second->_num_d_groups));
    const size_t g = G, m = M, d = D;
    mat_flags.assign(m, 0), n_delayed.assign(m, 0), step_type.assign(m, DOMA_STEP_NONE);
    sigma_t.assign(m * g, 0.0), sigma_s.assign(m * g * g, 0.0), nu_sigma_f.assign(m * g, 0.0), chi_p.assign(m * g, 0.0);
    src.assign(m * g, 0.0), inv_v.assign(m * g, 0.0), chi_d.assign(m * g * d + 1, 0.0), beta.assign(m * g * d + 1, 0.0);
    lambda.assign(m * d + 1, 0.0), step_t_insert.assign(m, 0.0), step_t_remove.assign(m, 0.0), step_src.assign(m * g, 0.0);
    size_t i = 0;
    for (const auto & kv : sorted)
    {
      const MaterialProps & p = *kv.second;
      block_of_mat.push_back(kv.first);
      std::copy(p._g_total.begin(), p._g_total.end(), sigma_t.begin() + i * g);
      std::copy(p._g_g_scatter_mat.begin(), p._g_g_scatter_mat.end(), sigma_s.begin() + i * g * g);
      if (!p._g_src.empty()) { mat_flags[i] |= DOMA_MAT_HAS_SRC; std::copy(p._g_src.begin(), p._g_src.end(), src.begin() + i * g); }
      if (!p._g_chi_p.empty()) { mat_flags[i] |= DOMA_MAT_HAS_CHI_P; std::copy(p._g_chi_p.begin(), p._g_chi_p.end(), chi_p.begin() + i * g); }
      if (!p._g_prod.empty()) { mat_flags[i] |= DOMA_MAT_HAS_NU_SIGMA_F; std::copy(p._g_prod.begin(), p._g_prod.end(), nu_sigma_f.begin() + i * g); }
      if (!p._g_chi_p.empty() && p._g_prod.empty())
      {
        // the reference indexes an empty _g_prod here (undefined behaviour); reject instead
        std::fprintf(stderr, "Block %u provides chi_p but no nu_sigma_f!\n", kv.first);
        std::exit(1);
      }
      if (p._g_inv_v.size() == g) std::copy(p._g_inv_v.begin(), p._g_inv_v.end(), inv_v.begin() + i * g);
      n_delayed[i] = static_cast<int32_t>(p._num_d_groups);
      const size_t nd = p._num_d_groups;
      for (size_t gg = 0; gg < g && nd; ++gg)
        for (size_t dd = 0; dd < nd; ++dd)
        {
          chi_d[(i * g + gg) * d + dd] = p._n_g_chi_d[gg * nd + dd];
          beta[(i * g + gg) * d + dd] = p._g_n_beta[gg * nd + dd];
        }
      for (size_t dd = 0; dd < nd; ++dd)
        lambda[i * d + dd] = p._n_lambda[dd];
      auto st = steps.find(kv.first);
      if (st != steps.end())
      {
        const SourceStep & s = st->second;
        step_type[i] = s._type == StepType::Both ? DOMA_STEP_BOTH : (s._type == StepType::Insert ? DOMA_STEP_INSERT : DOMA_STEP_REMOVE);
        if (s._type != StepType::Remove) step_t_insert[i] = s._insert_time;
        if (s._type != StepType::Insert) step_t_remove[i] = s._remove_time;
        std::copy(s._g_src.begin(), s._g_src.end(), step_src.begin() + i * g);
      }
      ++i;
    }
  }

  int materialIndex(unsigned int block) const
  {
    auto it = std::lower_bound(block_of_mat.begin(), block_of_mat.end(), block);
    if (it == block_of_mat.end() || *it != block)
    {
      std::fprintf(stderr, "Block %u (defined on the mesh) has no material properties!\n", block);
      std::exit(1);
    }
    return static_cast<int>(it - block_of_mat.begin());
  }

  // quadrature in the reference's octant-major order; this rank keeps the sweep units i with
  // i % world == rank, a unit being a batch of <= 32 ordinates of one octant (as api.cu build_units)
  void setQuadrature(const GCAngularQuadrature & quad, const RankInfo & r)
  {
    ord_octant.clear(), mu.clear(), eta.clear(), xi.clear(), w.clear();
    int unit = 0;
    for (unsigned int o = 0u; o < 8u; ++o)
    {
      const int cnt = static_cast<int>(quad.order(static_cast<Octant>(o)));
      if (cnt == 0)
        continue;
      const int nb = (cnt + 31) / 32;
      int pos = 0;
      for (int b = 0; b < nb; ++b, ++unit)
      {
        const int sz = cnt / nb + (b < cnt % nb ? 1 : 0);
        if (unit % r.world == r.rank)
          for (int n = pos; n < pos + sz; ++n)
          {
            double m = 0.0, e = 0.0, x = 0.0;
            quad.direction(static_cast<Octant>(o), n, m, e, x);
            ord_octant.push_back(static_cast<int32_t>(o));
            mu.push_back(m), eta.push_back(e), xi.push_back(x);
            w.push_back(quad.weight(static_cast<Octant>(o), n));
          }
        pos += sz;
      }
    }
  }

  doma_b200_problem view() const
  {
    doma_b200_problem p{};
    p.dims = dims;
    p.nx = static_cast<int32_t>(dx.size()), p.ny = static_cast<int32_t>(dy.size()), p.nz = dims == 3 ? static_cast<int32_t>(dz.size()) : 1;
    p.dx = dx.data(), p.dy = dy.data(), p.dz = dims == 3 ? dz.data() : nullptr;
    p.mat_of_cell = mat_of_cell.data();
    p.n_groups = G, p.n_mat = M, p.n_delayed_max = D;
    p.mat_flags = mat_flags.data(), p.n_delayed = n_delayed.data();
    p.sigma_t = sigma_t.data(), p.sigma_s = sigma_s.data(), p.nu_sigma_f = nu_sigma_f.data(), p.chi_p = chi_p.data();
    p.src = src.data(), p.inv_v = inv_v.data(), p.chi_d = chi_d.data(), p.beta = beta.data(), p.lambda = lambda.data();
    p.step_type = step_type.data(), p.step_t_insert = step_t_insert.data(), p.step_t_remove = step_t_remove.data();
    p.step_src = step_src.data();
    p.n_ord = static_cast<int32_t>(mu.size());
    p.ord_octant = ord_octant.data(), p.mu = mu.data(), p.eta = eta.data(), p.xi = xi.data(), p.w = w.data();
    p.scheme = scheme, p.mode = mode, p.dt = dt, p.device = device;
    return p;
  }
};

// Settings the driver's main() fixes before it constructs a solver (the operator type is
// default-constructed inside the reference's constructor, so it cannot take arguments).
struct DriverConfig
{
  int scheme = DOMA_SCHEME_TWDD;
  RankInfo rank;
  int tile_y = 0, tile_z = 0; // 0 = library default
  double last_k = 1.0;                // TransportSolver::_k after the last updateEigenvalue
  double solve_seconds = 0.0;         // wall time of the solve*() call alone
  doma_b200_ctx * live_ctx = nullptr; // context of the solver currently alive (owned by its operator)
  FlatProblem flat;                   // what the context was created from (kept for the dumps)
  // --fast-dump: steady / eigenvalue runs write their final text dump through doma_b200_dump_text instead of
  // BrickMesh::dumpToTextFile (same bytes, tests/test_text_output.py), and computeMGFluxResidual no longer copies
  // the flux into the per-cell host vectors after every outer iteration.
  bool fast_dump = false;
  bool output_enabled = true; // false with --no-output (the same on every rank; only rank 0 has a file base)
  // flat mesh (--flat, automatic above 2 M cells in steady / eigenvalue runs): geometry and block ids come from
  // FlatProblem::setGeometryFromParams, the solver is handed a stub mesh of one cell per block; implies --fast-dump
  bool flat_mesh = false;
  bool check_flat = false;   // --check-flat: compare the flat builder with the mesh the reference built, then exit
  bool separate_si = false;  // --separate-si: the reference's own three-call source iteration instead of doma_b200_source_iteration
  bool poison = false;       // --check: DOMA_B200_CHECK-style poisoning of the sweep scratch (doma_b200_set_check)
  std::string raw_dir;       // --dump-raw <dir>: final state as raw doubles, full precision
};
inline DriverConfig & config()
{
  static DriverConfig c;
  return c;
}

inline void
check(int rc, doma_b200_ctx * ctx, const char * what)
{
  if (rc != 0)
  {
    // the reference's fatal-error style: print and exit(1) (src/BrickMesh3D.cpp:26-30)
    std::fprintf(stderr, "Error: %s failed (%d): %s\n", what, rc, doma_b200_last_error(ctx));
    std::exit(1);
  }
}

inline std::FILE *
openOrDie(const std::string & path, const char * mode)
{
  std::FILE * f = std::fopen(path.c_str(), mode);
  if (!f)
  {
    std::fprintf(stderr, "Error: cannot open %s\n", path.c_str());
    std::exit(1);
  }
  return f;
}

// <base>_dims.txt, _blocks.txt, _meshx/y/z.txt exactly as BrickMesh3D::dumpToTextFile writes them
// (src/BrickMesh3D.cpp:216-258; 2D: src/BrickMesh2D.cpp:169-205), from the flattened geometry.
inline void
writeGeometryFiles(const FlatProblem & fp, int groups, const std::string & base)
{
  const std::size_t nx = fp.dx.size(), ny = fp.dy.size(), nz = fp.dims == 3 ? fp.dz.size() : 1;
  auto centres = [](const std::vector<double> & w) { // the reference's running centre (src/BrickMesh3D.cpp:84-101)
    std::vector<double> v(w.size());
    double acc = 0.5 * w[0];
    for (std::size_t i = 0; i < w.size(); ++i)
    {
      v[i] = acc;
      acc += w[i];
    }
    return v;
  };
  const auto xc = centres(fp.dx), yc = centres(fp.dy), zc = fp.dims == 3 ? centres(fp.dz) : std::vector<double>();
  std::FILE * f = openOrDie(base + "_dims.txt", "w");
  std::fprintf(f, "num_x: %zu\nnum_y: %zu\n", nx, ny);
  if (fp.dims == 3)
    std::fprintf(f, "num_z: %zu\n", nz);
  std::fprintf(f, "num_g: %d\n", groups);
  std::fclose(f);
  f = openOrDie(base + "_blocks.txt", "w");
  for (unsigned int b : fp.block_of_cell)
    std::fprintf(f, "%u\n", b);
  std::fclose(f);
  std::vector<double> v(nx * ny * nz);
  for (int axis = 0; axis < fp.dims; ++axis)
  {
    for (std::size_t k = 0, n = 0; k < nz; ++k)
      for (std::size_t j = 0; j < ny; ++j)
        for (std::size_t i = 0; i < nx; ++i, ++n)
          v[n] = axis == 0 ? xc[i] : (axis == 1 ? yc[j] : zc[k]);
    const char * names[] = {"_meshx.txt", "_meshy.txt", "_meshz.txt"};
    if (doma_b200_write_values_text(v.data(), (int64_t)v.size(), (base + names[axis]).c_str(), 0) != 0)
    {
      std::fprintf(stderr, "Error: cannot write %s%s\n", base.c_str(), names[axis]);
      std::exit(1);
    }
  }
}

// --check-flat: 0 when the flat builder reproduced the mesh's widths and block ids bit for bit
inline int
reportFlatCheck(const FlatProblem & flat, const FlatProblem & mesh)
{
  auto same = [](const std::vector<double> & a, const std::vector<double> & b) {
    return a.size() == b.size() && (a.empty() || std::memcmp(a.data(), b.data(), a.size() * sizeof(double)) == 0);
  };
  const bool ok = same(flat.dx, mesh.dx) && same(flat.dy, mesh.dy) && same(flat.dz, mesh.dz) && flat.block_of_cell == mesh.block_of_cell;
  std::printf("FLATCHECK %s nx=%zu ny=%zu nz=%zu cells=%zu\n", ok ? "identical" : "DIFFERENT", flat.dx.size(), flat.dy.size(),
              flat.dz.size(), flat.block_of_cell.size());
  return ok ? 0 : 4;
}

// NCCL unique-id exchange between the ranks of one launch (main.cpp).
void bootstrapComm(doma_b200_ctx * ctx, const RankInfo & r);

inline doma_b200_ctx *
createContext(const FlatProblem & flat)
{
  const auto p = flat.view();
  doma_b200_ctx * ctx = nullptr;
  const int rc = doma_b200_create(&p, &ctx);
  if (rc != 0)
  {
    std::fprintf(stderr, "Error: doma_b200_create failed (%d): %s\n", rc, doma_b200_last_error(nullptr));
    std::exit(1);
  }
  auto & cf = config();
  if (cf.tile_y > 0 && cf.tile_z > 0)
    check(doma_b200_set_tile(ctx, 8, cf.tile_y, cf.tile_z), ctx, "doma_b200_set_tile");
  if (cf.poison)
    check(doma_b200_set_check(ctx, 1), ctx, "doma_b200_set_check");
  bootstrapComm(ctx, cf.rank);
  cf.live_ctx = ctx;
  return ctx;
}

// --dump-raw <dir>: the final state at full precision, the files oracle/ref_harness.cpp writes for the reference
// (phi_g<g>.f64, dnp_d<d>.f64, meta.json), so the same digests check both.
inline void
dumpRaw(doma_b200_ctx * ctx, const FlatProblem & fp, bool ok, const std::string & dir)
{
  std::vector<double> buf(fp.mat_of_cell.size());
  auto put = [&](int which, int index, const std::string & name) {
    check(doma_b200_get_array(ctx, which, index, buf.data()), ctx, "doma_b200_get_array");
    if (config().rank.rank != 0)
      return;
    std::FILE * f = openOrDie(dir + "/" + name, "wb");
    if (std::fwrite(buf.data(), sizeof(double), buf.size(), f) != buf.size())
    {
      std::fprintf(stderr, "Error: short write to %s/%s\n", dir.c_str(), name.c_str());
      std::exit(1);
    }
    std::fclose(f);
  };
  for (int g = 0; g < fp.G; ++g)
    put(DOMA_ARR_PHI_TOT, g, "phi_g" + std::to_string(g) + ".f64");
  if (fp.mode == DOMA_MODE_TRANSIENT)
    for (int d = 0; d < fp.D; ++d)
      put(DOMA_ARR_DNP_CUR, d, "dnp_d" + std::to_string(d) + ".f64");
  if (config().rank.rank != 0)
    return;
  std::FILE * f = openOrDie(dir + "/meta.json", "w");
  std::fprintf(f, "{\"dims\": %d, \"nx\": %zu, \"ny\": %zu, \"nz\": %zu, \"groups\": %d, \"n_delayed\": %d, \"ok\": %s, \"k\": %.17g, "
                  "\"solve_seconds\": %.6f}\n",
               fp.dims, fp.dx.size(), fp.dy.size(), fp.dims == 3 ? fp.dz.size() : (std::size_t)1, fp.G, fp.D, ok ? "true" : "false",
               config().last_k, config().solve_seconds);
  std::fclose(f);
}
} // namespace doma_b200_driver
