/*
 * doma_b200.h — C ABI of the B200-native DOMA sweep hot path.
 *
 * One context = one problem (mesh + materials + this rank's ordinates) resident on ONE GPU.
 * Every entry point replaces one private member of the reference's
 * TransportSolver3D<T> / TransportSolver2D<T> (citations are relative to the DOMA source tree;
 * "3D" = src/TransportSolver3D.cpp, "2D" = src/TransportSolver2D.cpp).  The reference has no FFI;
 * its only extension point is the template parameter T (include/TransportSolver3D.h:12-16,112-113),
 * so the boundary sits on the members that call T::solve.  INTEGRATION.md shows the
 * explicit-specialisation seam that binds these calls under the reference's unmodified loops.
 *
 * Conventions
 *  - plain pointers and sizes only; every function returns 0 on success, non-zero on error
 *    (message via doma_b200_last_error); nothing throws across the boundary.
 *  - all host pointers are read (or written) during the call only; the context owns device memory.
 *  - one calling thread per context; calls are ordered on the context's CUDA stream; calls that
 *    return a scalar or copy to the host synchronise that stream.
 *  - all floating point is IEEE binary64, as in the reference.
 *  - there is NO CPU fallback: create fails if no CUDA device is usable.
 */
#ifndef DOMA_B200_H
#define DOMA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* include/TransportBase.h:20-30 (Octant), :38-43 (RunMode), :45-49 (EquationType), :57-62 (StepType) */
enum { DOMA_OCT_PPP = 0, DOMA_OCT_PPM, DOMA_OCT_PMP, DOMA_OCT_PMM, DOMA_OCT_MPP, DOMA_OCT_MPM, DOMA_OCT_MMP, DOMA_OCT_MMM };
enum { DOMA_MODE_FIXED = 0, DOMA_MODE_EIGEN = 1, DOMA_MODE_TRANSIENT = 2 };
enum { DOMA_SCHEME_DD = 0, DOMA_SCHEME_TWDD = 1 };
enum { DOMA_STEP_NONE = -1, DOMA_STEP_INSERT = 0, DOMA_STEP_REMOVE = 1, DOMA_STEP_BOTH = 2 };
/* mat_flags bits: which optional MaterialProps vectors are non-empty (include/InputParameters.h:26-52) */
enum { DOMA_MAT_HAS_SRC = 1, DOMA_MAT_HAS_CHI_P = 2, DOMA_MAT_HAS_NU_SIGMA_F = 4 };

/*
 * Flat, caller-owned problem description; copied at create.
 * Cell order is the reference's: index = k*ny*nx + j*nx + i, x fastest (src/BrickMesh3D.cpp:86-102).
 * G = n_groups, D = n_delayed_max, M = n_mat.
 */
typedef struct doma_b200_problem
{
  int32_t dims;               /* 2 or 3 */
  int32_t nx, ny, nz;         /* total cells per axis (BrickMesh3D::_tot_num_x/y/z); nz = 1 in 2D */
  const double * dx;          /* [nx] per-cell x widths (CartesianCell3D::_l_x, src/BrickMesh3D.cpp:44-49) */
  const double * dy;          /* [ny] */
  const double * dz;          /* [nz] (ignored in 2D) */
  const int32_t * mat_of_cell;/* [nx*ny*nz] dense material index of each cell (from _block_id) */

  int32_t n_groups;
  int32_t n_mat;
  int32_t n_delayed_max;      /* D: max MaterialProps::_num_d_groups over materials (0 outside transients) */
  const int32_t * mat_flags;  /* [M] DOMA_MAT_HAS_* */
  const int32_t * n_delayed;  /* [M] MaterialProps::_num_d_groups */
  const double * sigma_t;     /* [M*G]   _g_total */
  const double * sigma_s;     /* [M*G*G] _g_g_scatter_mat, [to*G+from] after any transpose (src/InputParameters.cpp:301-318) */
  const double * nu_sigma_f;  /* [M*G]   _g_prod   (zeros when absent) */
  const double * chi_p;       /* [M*G]   _g_chi_p  (zeros when absent) */
  const double * src;         /* [M*G]   _g_src    (zeros when absent) */
  const double * inv_v;       /* [M*G]   _g_inv_v  (zeros when absent) */
  const double * chi_d;       /* [M*G*D] _n_g_chi_d re-strided to [g*D+d] */
  const double * beta;        /* [M*G*D] _g_n_beta  re-strided to [g*D+d] */
  const double * lambda;      /* [M*D]   _n_lambda */
  const int32_t * step_type;  /* [M] DOMA_STEP_* of the block's SourceStep (include/InputParameters.h:10-22) */
  const double * step_t_insert; /* [M] */
  const double * step_t_remove; /* [M] */
  const double * step_src;    /* [M*G] */

  /* This context's ordinates (the caller shards them across ranks), signed exactly as
   * GCAngularQuadrature::direction returns them (src/GCAngularQuadrature.cpp:58-64). */
  int32_t n_ord;
  const int32_t * ord_octant; /* [n_ord] DOMA_OCT_* */
  const double * mu;          /* [n_ord] */
  const double * eta;         /* [n_ord] */
  const double * xi;          /* [n_ord] */
  const double * w;           /* [n_ord] */

  int32_t scheme;             /* DOMA_SCHEME_* (XML scheme=, src/main.cpp:208,250) */
  int32_t mode;               /* DOMA_MODE_*   (TransportSolver3D::_mode) */
  double dt;                  /* TransportSolver3D::_dt; read only in transient mode */
  int32_t device;             /* CUDA device ordinal */
} doma_b200_problem;

typedef struct doma_b200_ctx doma_b200_ctx;

/* lifetime -------------------------------------------------------------------------------------- */
int doma_b200_create(const doma_b200_problem * problem, doma_b200_ctx ** out);
int doma_b200_destroy(doma_b200_ctx * ctx);
const char * doma_b200_last_error(const doma_b200_ctx * ctx); /* ctx may be NULL: last create error */
/* run the context on a caller-provided CUDA stream (cudaStream_t as void*); NULL = the context's own */
int doma_b200_set_stream(doma_b200_ctx * ctx, void * cuda_stream);
int doma_b200_synchronize(doma_b200_ctx * ctx);

/* state management ------------------------------------------------------------------------------ */
/* initializeSolve (3D:499-529): phi_tot[g] = phi_init (1.0 eigen / 0.0 otherwise), phi_prev = 0, q = 0, phi_cur = 0 */
int doma_b200_reset_state(doma_b200_ctx * ctx, double phi_init);
/* initZeroIC (3D:140-156): phi_tot, phi_prev, phi_last, dnp_cur, dnp_last = 0 (q and phi_cur untouched) */
int doma_b200_init_zero_ic(doma_b200_ctx * ctx);
/* updateStepFluxes (3D:212-234): phi_last = phi_tot; phi_tot = 0; q = 0; phi_cur = 0 */
int doma_b200_update_step_fluxes(doma_b200_ctx * ctx);
/* stepDNPs (3D:236-260): analytic precursor step with exp(-lambda*dt) */
int doma_b200_step_dnps(doma_b200_ctx * ctx);

/* source build ---------------------------------------------------------------------------------- */
/* updateMultigroupSource(g,t) (3D:339-402, 2D same lines with 0.5/pi) */
int doma_b200_update_mg_source(doma_b200_ctx * ctx, int32_t g, double t);
/* updateMultigroupSourceEigen(g) (3D:404-430); q is accumulated ("+="); k = TransportSolver3D::_k */
int doma_b200_update_mg_source_eigen(doma_b200_ctx * ctx, int32_t g, double k);
/* updateScatteringSource(g) (3D:531-545) */
int doma_b200_update_scatter_source(doma_b200_ctx * ctx, int32_t g);

/* sweep ----------------------------------------------------------------------------------------- */
/* sweep(g) (3D:565-622, 2D:565-602): phi_cur[c] += sum over this context's ordinates of w*psi_c.
 * With ordinates sharded over ranks the caller sums phi_cur across ranks afterwards
 * (doma_b200_allreduce_phi_cur, or doma_b200_device_ptr(ctx, DOMA_ARR_PHI_CUR, 0, &p) + its own collective on
 * the same stream); doma_b200_source_iteration does both. */
int doma_b200_sweep(doma_b200_ctx * ctx, int32_t g);

/* Host-buffer form of sweep(g), the call a host-resident driver makes: copies q (N doubles, reference
 * cell order) to the device, zeroes phi_cur, sweeps, copies phi_cur back into phi_cur_out. Synchronous. */
int doma_b200_sweep_from_host(doma_b200_ctx * ctx, int32_t g, const double * q_host, double * phi_cur_out);

/* Asynchronous form of the host-buffer sweep, for callers that pipeline groups: the same work as
 * doma_b200_sweep_from_host, enqueued and returned from at once.  The host-to-device copy of q runs on a copy
 * stream (doma_b200_stage_q_from_host may start it early, while the previous sweep is still running; then pass
 * q_host = NULL here), the device-to-host copy of phi_cur on another one, under whatever the caller enqueues next.
 * q_host must stay valid and unchanged, and phi_cur_out is only complete, once doma_b200_host_wait (or
 * doma_b200_synchronize) has returned.
 * One staged source and one pending result at a time per context (a second stage before its sweep is an error;
 * a second sweep waits for the previous result to leave the staging buffer). */
int doma_b200_stage_q_from_host(doma_b200_ctx * ctx, const double * q_host);
int doma_b200_sweep_from_host_async(doma_b200_ctx * ctx, int32_t g, const double * q_host, double * phi_cur_out);
int doma_b200_host_wait(doma_b200_ctx * ctx);

/* text output -------------------------------------------------------------------------------------- */
/* The flux (and precursor) files of BrickMesh3D::dumpToTextFile / dumpDNPsToTextFile (src/BrickMesh3D.cpp:216-295;
 * 2D: src/BrickMesh2D.cpp): "<base>_g<g>_flux.txt" for every group and, with with_dnps != 0 in a transient context,
 * "<base>_d<d>_dnps.txt", one "%.6g" value per line in reference cell order - byte for byte what the reference's
 * `stream << std::setprecision(6) << v << std::endl` loop writes, formatted by n_threads host threads (0 = all cores).
 * The geometry files (_dims, _blocks, _meshx/y/z) do not depend on the solve and stay with the host front end. */
int doma_b200_dump_text(doma_b200_ctx * ctx, const char * file_base, int32_t with_dnps, int32_t n_threads);
/* The same, overlapped with whatever the caller does next: the state is copied to the host before the call returns
 * (so the solve may continue), the files are formatted and written on a worker thread.  One dump in flight per context:
 * the next dump, doma_b200_dump_wait and doma_b200_destroy wait for it; its error, if any, is returned by that call. */
int doma_b200_dump_text_async(doma_b200_ctx * ctx, const char * file_base, int32_t with_dnps, int32_t n_threads);
int doma_b200_dump_wait(doma_b200_ctx * ctx);
/* the formatter on its own (pure host code, needs no context or device) */
int doma_b200_write_values_text(const double * values, int64_t n, const char * path, int32_t n_threads);

/* sourceIteration(g) (3D:458-497) as ONE call (the fused inner iteration of SURVEY.md 8f-1):
 *   do { sweep(g); all-reduce phi_cur over ranks; residual = computeScatteringResidual(g); updateScatteringSource(g); it++ }
 *   while (it < max_it && tol < residual);
 * with the residual and the scattering-source update fused into one pass over the cells (identical values).
 * iters_out = iterations executed, residual_out = the last residual; the caller applies the reference's
 * "converged = iters < max_it" (3D:481).  residual_history (may be NULL, else >= max_it doubles) receives the
 * residual of every iteration, for the reference's --verbose trace. */
int doma_b200_source_iteration(doma_b200_ctx * ctx, int32_t g, double tol, int32_t max_it, int32_t * iters_out,
                               double * residual_out, double * residual_history);

/* angle sharding across ranks (one process per GPU): the scalar-flux increment is summed once per
 * sweep with an in-place ncclAllReduce(phi_cur, N, ncclDouble, ncclSum) on the context's stream
 * (replaces the serial thread-sum of 3D:609-621).  Rank 0 creates the 128-byte NCCL unique id, the
 * host layer distributes it, every rank calls comm_init once.  Without comm_init, allreduce is a no-op. */
int doma_b200_comm_unique_id(char * out128, int32_t len);
int doma_b200_comm_init(doma_b200_ctx * ctx, int32_t rank, int32_t nranks, const char * unique_id128, int32_t len);
int doma_b200_allreduce_phi_cur(doma_b200_ctx * ctx);
/* Cell sharding (default ON after comm_init with nranks > 1): only the sweep needs every cell on every rank, so rank r
 * keeps phi_tot / phi_prev / phi_last / the precursors of cells [lo, hi) = DOMA_INFO_CELL_LO / _HI only and runs the
 * source build, residual, update and k kernels on those cells:
 *     ncclAllGather(q) -> sweep -> ncclReduceScatter(phi_cur) -> fused residual + update on the own cells
 * (the same NVLink bytes as one all-reduce, the elementwise work divided by nranks; the two partial sums of a reduction
 * are summed with a 16-byte ncclAllReduce, so every rank sees identical bits and takes identical branches).
 * Consequences for the caller, all ranks alike: doma_b200_get_array, doma_b200_dump_text[_async] and
 * doma_b200_set_cell_sharding are COLLECTIVE (every rank calls them; dump_text with an empty file base takes part in the
 * gather and writes nothing); the host-buffer sweeps move the own cells only (q_host[lo..hi) up, phi_cur_out[lo..hi) down).
 * on = 0 restores the r01 scheme: in-place ncclAllReduce(phi_cur), every rank repeats the elementwise work. */
int doma_b200_set_cell_sharding(doma_b200_ctx * ctx, int32_t on);

/* reductions (each synchronises and returns scalars) ------------------------------------------- */
/* computeScatteringResidual(g) (3D:547-563) */
int doma_b200_scatter_residual(doma_b200_ctx * ctx, int32_t g, double * out);
/* The same residual for callers that cannot let the library allocate (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * It needs no context and calls no cudaMalloc / cudaFree / allocating routine: the reduction's scratch (one pair of
 * partial sums per block of DOMA_NOALLOC_CELLS_PER_BLOCK cells + the final pair) is a file-scope __device__ array of
 * DOMA_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.  Limit: nx*ny*nz <=
 * DOMA_NOALLOC_MAX_CELLS, more is rejected (error 40 + message).  All calls on a device share the array: a call reads
 * only slots it has written itself, so nothing needs clearing between calls, but calls that OVERLAP on the GPU (other
 * streams, concurrent graph branches) race on it and must be serialised by the caller.
 *   inputs  (device): phi_tot_g[N], phi_cur[N] (reference cell order), dx[nx], dy[ny], dz[nz] (dz ignored for dims = 2)
 *   outputs: sums_out_dev[2] (device) = { sum V cur^2, sum V (tot+cur)^2 }, written on cuda_stream (a cudaStream_t, may
 *            be NULL = the default stream); residual_out_host (may be NULL): when given, the call synchronises the
 *            stream and stores total > 1e-8 ? sqrt(diff)/sqrt(total) : 0 (3D:562) there.
 *   err_out (may be NULL): receives the message of a non-zero return.  1 = bad argument, 40 = too many cells, 100 + e = CUDA. */
#define DOMA_NOALLOC_CELLS_PER_BLOCK 4096
#define DOMA_NOALLOC_MAX_CELLS 33554432LL  /* 8 192 blocks x 4 096 cells: twice the 256^3 bench mesh */
#define DOMA_NOALLOC_SCRATCH_BYTES 131328  /* 2 x 8 192 partial sums + 32 doubles of tail, per GPU */
int doma_b200_scatter_residual_noalloc(const double * phi_tot_g_dev, const double * phi_cur_dev, const double * dx_dev,
                                       const double * dy_dev, const double * dz_dev, int32_t dims, int32_t nx, int32_t ny, int32_t nz,
                                       void * cuda_stream, double * sums_out_dev, double * residual_out_host, char * err_out,
                                       int32_t err_len);
/* computeMGFluxResidual() (3D:318-337) */
int doma_b200_mg_residual(doma_b200_ctx * ctx, double * out);
/* updateEigenvalue() (3D:432-456): k_prev_out = *k_inout; *k_inout = num/den */
int doma_b200_update_k(doma_b200_ctx * ctx, double * k_inout, double * k_prev_out);

/* host <-> device state ------------------------------------------------------------------------- */
enum
{
  DOMA_ARR_PHI_TOT = 0,  /* CartesianCell3D::_total_scalar_flux[g]      index = group */
  DOMA_ARR_PHI_PREV = 1, /* _prev_mg_scalar_flux[g]                     index = group */
  DOMA_ARR_PHI_LAST = 2, /* _last_t_scalar_flux[g]                      index = group */
  DOMA_ARR_DNP_CUR = 3,  /* _current_t_dnps[d]                          index = delayed group */
  DOMA_ARR_DNP_LAST = 4, /* _last_t_dnps[d]                             index = delayed group */
  DOMA_ARR_Q = 5,        /* _current_iteration_source                   index ignored */
  DOMA_ARR_PHI_CUR = 6   /* _current_scalar_flux                        index ignored */
};
/* copy one per-cell array (N doubles, reference cell order) device -> host / host -> device */
int doma_b200_get_array(doma_b200_ctx * ctx, int32_t which, int32_t index, double * host_out);
int doma_b200_set_array(doma_b200_ctx * ctx, int32_t which, int32_t index, const double * host_in);
/* raw device pointer of a per-cell array (for stream-ordered collectives by the host layer) */
int doma_b200_device_ptr(doma_b200_ctx * ctx, int32_t which, int32_t index, void ** out);

/* introspection --------------------------------------------------------------------------------- */
enum
{
  DOMA_INFO_KERNEL_LAUNCHES = 0, /* kernels launched by this context so far */
  DOMA_INFO_SWEEP_VARIANT = 1,   /* face-exchange protocol of the next sweep: DOMA_SWEEP_FLAGS or DOMA_SWEEP_LL (see set_sweep_variant) */
  DOMA_INFO_N_CELLS = 2,
  DOMA_INFO_N_ORD = 3,
  DOMA_INFO_FACE_BYTES = 4,
  DOMA_INFO_DEVICE_BYTES = 5,
  DOMA_INFO_N_UNITS = 6,         /* sweep units = (octant, batch of <= 32 ordinates) */
  DOMA_INFO_N_TILES = 7,         /* tiles of the KBA decomposition */
  DOMA_INFO_TILE_Y = 8,          /* tile extents in cells (x extent: DOMA_INFO_TILE_X); chosen by the library unless set_tile was called */
  DOMA_INFO_TILE_Z = 9,
  DOMA_INFO_CELL_LO = 10,        /* cells [lo, hi) whose state this rank keeps (0, N unless cell-sharded) */
  DOMA_INFO_CELL_HI = 11,
  DOMA_INFO_ROWS_PER_STEP = 12,  /* rows of a slab the next sweep advances together: 1 or 2 (see set_rows_per_step) */
  DOMA_INFO_SWEEP_BLOCKS = 13,   /* persistent CTAs (4 warps each) of the last sweep launch */
  DOMA_INFO_TILE_X = 14          /* x extent of a tile: 8 or 4 cells */
};
int doma_b200_get_info(doma_b200_ctx * ctx, int32_t what, int64_t * out);
/* How the persistent sweep kernel hands tile-boundary face fluxes from the warp that produced them to the warp that
 * consumes them (sweep.cu):
 *   DOMA_SWEEP_FLAGS  8-byte face values + one done-flag per (unit, tile) behind a device-scope fence: least traffic,
 *                     best when the wavefront saturates the GPU (many units per GPU);
 *   DOMA_SWEEP_LL     16-byte self-validating face values (two 8-byte {half, tag} words, no flag, no fence): a consumer
 *                     follows its producer row by row instead of tile by tile; best when the sweep is a chain of task
 *                     latencies (few units per GPU: 8-way angle sharding, the shipped assembly and 2D cases);
 *   DOMA_SWEEP_AUTO   (default) the library picks per problem from units x tiles per hyperplane vs resident warps.
 * Results are identical (same arithmetic, same order); only the timing differs. */
enum { DOMA_SWEEP_AUTO = 0, DOMA_SWEEP_FLAGS = 1, DOMA_SWEEP_LL = 2 };
int doma_b200_set_sweep_variant(doma_b200_ctx * ctx, int32_t variant);
/* Rows of a tile slab a warp advances together (sweep.cu, template parameter R2): 1 = row after row, 16 warps per SM,
 * best throughput when the wavefront fills the GPU; 2 = two rows skewed by one cell, two independent dependency chains
 * per warp, 12 warps per SM, best when the sweep is a chain of task latencies; 0 (default) = the library picks from
 * the tasks in flight per resident warp.  Results are identical bit for bit. */
int doma_b200_set_rows_per_step(doma_b200_ctx * ctx, int32_t rows);
/* Octant pairing: with <= 16 ordinates per octant two octants that differ only in the x direction share a warp,
 * 16 lanes each (default on; off = one octant per warp, for A/B tests).  Rebuilds the sweep units. */
int doma_b200_set_pairing(doma_b200_ctx * ctx, int32_t on);
/* Debug mode (also enabled by the environment variable DOMA_B200_CHECK=1 at create): before every sweep the face planes
 * are poisoned (NaN / an impossible tag) and after it the library verifies that every tile of every unit completed and
 * that phi_cur is finite; a missed wavefront dependency then fails the call (error 30) instead of reading last
 * sweep's value.  Costs a synchronisation per sweep. */
int doma_b200_set_check(doma_b200_ctx * ctx, int32_t on);
/* Measurement mode, environment variable DOMA_B200_COMM_TIMING=1 at create (multi-rank contexts): every collective
 * (all-gather of q, reduce-scatter / all-reduce of phi_cur, the 16-byte pair all-reduce, the accumulator clear) is
 * timed on the device in isolation - ranks aligned first, the stream synchronised - and the means are printed to stderr
 * by doma_b200_destroy.  Not for timed runs: it serialises the step. */
/* KBA tile extents in cells; tx is 8 or 4 (the row a warp keeps in registers), tx * ty <= 32; (tx, 0, 0) returns to the
 * library's own choice. Rebuilds the tile table.
 * Without this call the library picks the tile from the mesh size and the number of sweep units (api.cu). */
int doma_b200_set_tile(doma_b200_ctx * ctx, int32_t tx, int32_t ty, int32_t tz);
/* device time in ms of the most recent doma_b200_sweep (CUDA events on the context's stream; synchronises) */
int doma_b200_last_sweep_ms(doma_b200_ctx * ctx, float * out);

#ifdef __cplusplus
}
#endif
#endif /* DOMA_B200_H */
