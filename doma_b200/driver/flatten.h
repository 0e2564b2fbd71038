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
  std::vector<int32_t> mat_of_cell;
  int G = 1, M = 0, D = 0;
  std::vector<unsigned int> block_of_mat;
  std::vector<int32_t> mat_flags, n_delayed, step_type;
  std::vector<double> sigma_t, sigma_s, nu_sigma_f, chi_p, src, inv_v, chi_d, beta, lambda, step_t_insert, step_t_remove, step_src;
  std::vector<int32_t> ord_octant;
  std::vector<double> mu, eta, xi, w;

  // materials: dense index = rank of the block id (the reference keys an unordered_map by block id)
  void setMaterials(const std::unordered_map<unsigned int, MaterialProps> & mats,
                    const std::unordered_map<unsigned int, SourceStep> & steps, unsigned int groups)
  {
    G = static_cast<int>(groups);
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
  doma_b200_ctx * live_ctx = nullptr; // context of the solver currently alive (owned by its operator)
  // --fast-dump: steady / eigenvalue runs write their final text dump through doma_b200_dump_text instead of
  // BrickMesh::dumpToTextFile (same bytes, tests/test_text_output.py), and computeMGFluxResidual no longer copies
  // the flux into the per-cell host vectors after every outer iteration.  The geometry kept for the static files:
  bool fast_dump = false;
  int geom_dims = 3, geom_groups = 1;
  std::vector<double> geom_dx, geom_dy, geom_dz;
  std::vector<unsigned int> geom_block_of_cell;
};
inline DriverConfig & config()
{
  static DriverConfig c;
  return c;
}

// <base>_dims.txt, _blocks.txt, _meshx/y/z.txt exactly as BrickMesh3D::dumpToTextFile writes them
// (src/BrickMesh3D.cpp:216-258; 2D: src/BrickMesh2D.cpp:169-205), from the flattened geometry.
inline void
writeGeometryFiles(const DriverConfig & c, const std::string & base)
{
  const std::size_t nx = c.geom_dx.size(), ny = c.geom_dy.size(), nz = c.geom_dims == 3 ? c.geom_dz.size() : 1;
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
  const auto xc = centres(c.geom_dx), yc = centres(c.geom_dy), zc = c.geom_dims == 3 ? centres(c.geom_dz) : std::vector<double>();
  std::FILE * f = std::fopen((base + "_dims.txt").c_str(), "w");
  std::fprintf(f, "num_x: %zu\nnum_y: %zu\n", nx, ny);
  if (c.geom_dims == 3)
    std::fprintf(f, "num_z: %zu\n", nz);
  std::fprintf(f, "num_g: %d\n", c.geom_groups);
  std::fclose(f);
  f = std::fopen((base + "_blocks.txt").c_str(), "w");
  for (unsigned int b : c.geom_block_of_cell)
    std::fprintf(f, "%u\n", b);
  std::fclose(f);
  std::vector<double> v(nx * ny * nz);
  for (int axis = 0; axis < c.geom_dims; ++axis)
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

// NCCL unique-id exchange between the ranks of one box through a file (rank 0 writes, others poll).
void bootstrapComm(doma_b200_ctx * ctx, const RankInfo & r);

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
} // namespace doma_b200_driver
