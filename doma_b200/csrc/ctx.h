// ctx.h — device-resident problem state shared by the translation units of libdoma_b200.so.
//
// Data layout in HBM (structure of arrays, x-fastest like the reference's cell index
// k*ny*nx + j*nx + i, src/BrickMesh3D.cpp:86-102) replacing the reference's 256-byte
// CartesianCell3D objects with five heap vectors each (include/CartesianCell3D.h:84-121):
//   q[N], phi_cur[N]                       _current_iteration_source, _current_scalar_flux
//   phi_tot[G][N], phi_prev[G][N]          _total_scalar_flux[g], _prev_mg_scalar_flux[g]
//   phi_last[G][N], dnp_cur/last[D][N]     transient mode only
//   mat[N] (int32 dense material index), dx[nx] dy[ny] dz[nz] and their reciprocals
//   material tables [M][G] / [M][G][G] / [M][G][D]
//   sweep scratch: per (octant, angle-batch) "unit" three face planes, angle fastest (32 lanes),
//   tile-done flags, the hyperplane-ordered tile table and the work-queue counter.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "doma_b200.h"

#define DOMA_MAX_ERR 512
#define DOMA_LANES 32  // ordinates of one octant batched across the lanes of a warp

struct SweepUnit
{
  int octant;  // DOMA_OCT_*
  int n_lanes; // ordinates in this unit, <= 32
};

#define DOMA_SI_HIST 2048
struct SiState
{
  int it;    // iterations completed
  int stop;  // set by the closing kernel once !(it < max_it && tol < residual): the remaining enqueued kernels return at once
  double res;
  double hist[DOMA_SI_HIST];
};

struct doma_b200_ctx
{
  int device;
  int dims, nx, ny, nz, G, M, D, n_ord, scheme, mode;
  int64_t N;
  double dt;
  cudaStream_t stream;
  bool own_stream;
  int n_sms;

  // geometry / materials
  double *dx, *dy, *dz, *rdx, *rdy, *rdz;
  int32_t * mat;
  int32_t *mat_flags, *n_delayed, *step_type;
  double *sigma_t, *sigma_s, *nu_sigma_f, *chi_p, *src, *inv_v, *chi_d, *beta, *lambda;
  double *step_t_insert, *step_t_remove, *step_src;
  double * sig_cell;  // [G][N] total cross section the cell solve uses, expanded per cell (built once, elementwise.cu)
  int sig_cell_group; // G once sig_cell has been built, -1 before

  // state
  double *phi_tot, *phi_prev, *phi_last, *dnp_cur, *dnp_last, *q, *phi_cur;

  // quadrature: host copy of this context's ordinates as given at create, and the same regrouped
  // into sweep units (octant x V batches of <=32 ordinates), lane-major
  int32_t * h_ord_octant;
  double *h_mu, *h_eta, *h_xi, *h_w; // [n_ord] signed, as passed in
  int n_units;
  SweepUnit * h_units; // host copy
  int * unit_octant;   // [U]
  bool units_paired;   // every unit carries two octants that differ in x, 16 lanes each (sweep.cu, PAIR)
  bool pairing_enabled; // doma_b200_set_pairing (default on)
  bool check;           // doma_b200_set_check / DOMA_B200_CHECK=1
  double *u_mu, *u_eta, *u_xi, *u_w; // [U][32] absolute direction cosines; padding lanes: 1,1,1,w=0

  // sweep scratch
  int tile_x, tile_y, tile_z; // tile_x: 8 or 4, the row a warp holds in registers (template parameter of the kernel)
  bool tile_auto;             // tile (and protocol) chosen by doma_choose_tile (sweep.cu), not by set_tile
  int variant;                // DOMA_SWEEP_AUTO / _FLAGS / _LL as requested (doma_b200_set_sweep_variant)
  bool use_ll;                // doma_choose_tile's protocol choice when variant is AUTO
  bool last_sweep_ll;         // protocol of the most recent sweep
  int rows2;                  // two rows per step in the sweep (R2): -1 chooser, 0 off, 1 on (DOMA_B200_ROWS2 at create)
  bool use_rows2;             // the chooser's answer
  int max_blocks;             // > 0: cap on the CTAs of a sweep launch (experiments: DOMA_B200_MAX_BLOCKS at create)
  int last_blocks;            // CTAs of the last sweep launch (DOMA_INFO_SWEEP_BLOCKS)
  int ll_epoch;               // LL tag epoch (20 bits; the planes are cleared when it wraps)
  bool planes_raw;            // the face planes were last written by a FLAGS sweep (raw doubles, no tags)
  int ntx, nty, ntz, n_tiles;
  uint32_t * tile_order; // [n_tiles] packed (I',J',K') sorted by I'+J'+K'
  double *face_x, *face_y, *face_z; // [U][nz][ny][32], [U][nz][nx][32], [U][ny][nx][32]
  int * tile_done;                   // [U][n_tiles] epoch of completion
  int * work_counter;                // [1]
  int epoch;
  int64_t face_bytes, device_bytes;

  // reductions
  double * red_partial; // [2 * RED_BLOCKS]
  double * red_out;     // [2] device
  double * h_red;       // [2] pinned, mapped host memory
  double * h_red_dev;   // the device-side address of h_red (final_reduce_kernel stores into it)

  // device-side source iteration (doma_b200_source_iteration on a single rank): iteration count, stop flag and the
  // residual history live in pinned, mapped host memory that the closing kernel of every iteration writes
  struct SiState * si;      // pinned host mirror (written by the closing kernel with zero-copy stores)
  struct SiState * si_host_dev; // device address of the mirror
  struct SiState * si_dev;  // the state the kernels read, in device memory

  // timing / accounting
  cudaEvent_t ev0, ev1;
  bool comm_timing;           // DOMA_B200_COMM_TIMING=1 at create: every collective is bracketed by events and synchronised (a measurement mode)
  double comm_ms[4];          // accumulated ms: all-gather(q), reduce-scatter / all-reduce(phi_cur), pair all-reduce, accumulator clear
  int comm_calls[4];
  double * comm_scratch;      // 16 bytes for the rank-aligning all-reduce of the timing mode
  bool have_sweep_time;
  int64_t launches;

  // asynchronous host-buffer sweeps (api.cu, doma_b200_sweep_from_host_async): staging buffers, copy streams
  // and the events that order them against the context's stream; created on first use
  double *q_stage, *phi_snap;
  cudaStream_t copy_in, copy_out;
  cudaEvent_t ev_q_staged, ev_q_consumed, ev_snap, ev_copied;
  bool async_ready, q_is_staged, copy_pending;

  void * dump_job; // text dump in flight on a worker thread (api.cu, DumpJob), or null

  // NCCL (resolved with dlopen at comm_init; see api.cu)
  void * nccl_comm;
  int rank, nranks;
  // Cell sharding of the elementwise phases across the ranks of an angle-sharded run (api.cu, comm_init): the sweep
  // needs q on every cell and adds into phi_cur on every cell, but everything else is per cell, so rank r keeps
  // the state of cells [cell_lo, cell_hi) only:  all-gather(q) -> sweep -> reduce-scatter(phi_cur) -> residual /
  // update / source build on the own cells.  Same NVLink bytes as one all-reduce, elementwise work / P.
  bool cell_sharded;
  int64_t cell_lo, cell_hi, slice; // slice = cells per rank, nranks * slice >= N (arrays are padded for it)
  double * phi_red;                // [slice] the summed increment of the own cells (what the residual / update kernels read)
  double * gather_tmp;             // [nranks * slice] scratch: get_array / dumps gather the slices here
  bool q_slice_only;               // q is current on the own cells only: all-gather before the next sweep
  bool phi_cur_reduced;            // phi_red holds the summed increment; the local accumulator phi_cur is zero again

  char err[DOMA_MAX_ERR];
};

#define DOMA_CUDA(ctx, call)                                                                                     \
  do                                                                                                             \
  {                                                                                                              \
    cudaError_t e_ = (call);                                                                                     \
    if (e_ != cudaSuccess)                                                                                       \
    {                                                                                                            \
      snprintf((ctx)->err, DOMA_MAX_ERR, "%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return 100 + (int)e_;                                                                                      \
    }                                                                                                            \
  } while (0)

#define DOMA_PAD 1024 // doubles of padding behind every per-cell array (nranks * slice - N <= 2 * nranks)
// the cells the elementwise kernels of this rank work on, and the increment they read
static inline int64_t ew_lo(const doma_b200_ctx * c) { return c->cell_sharded ? c->cell_lo : 0; }
static inline int64_t ew_hi(const doma_b200_ctx * c) { return c->cell_sharded ? c->cell_hi : c->N; }
static inline double * ew_phi_cur(const doma_b200_ctx * c) { return c->cell_sharded ? c->phi_red - c->cell_lo : c->phi_cur; }

// api.cu: collectives of the cell-sharded mode (no-ops otherwise)
int doma_comm_gather_q(doma_b200_ctx * c);              // all-gather(q) if only the own cells are current
int doma_comm_sum_pair(doma_b200_ctx * c);              // red_out[0..1] summed over ranks (identical bits on every rank)

// elementwise.cu
int doma_launch_reset_state(doma_b200_ctx * c, double phi_init);
int doma_launch_init_zero_ic(doma_b200_ctx * c);
int doma_launch_update_step_fluxes(doma_b200_ctx * c);
int doma_launch_step_dnps(doma_b200_ctx * c);
int doma_launch_update_mg_source(doma_b200_ctx * c, int g, double t);
int doma_launch_update_mg_source_eigen(doma_b200_ctx * c, int g, double k);
int doma_launch_update_scatter_source(doma_b200_ctx * c, int g);
int doma_launch_stage_in(doma_b200_ctx * c);          // q = q_stage, phi_cur = 0
int doma_launch_snapshot_phi_cur(doma_b200_ctx * c);  // phi_snap = phi_cur
int doma_launch_scatter_residual(doma_b200_ctx * c, int g); // leaves (diff, total) in h_red after sync
int doma_launch_scatter_residual_update(doma_b200_ctx * c, int g); // fused pair, leaves (diff, total) in h_red
// one iteration of the device-side loop: fused residual + update (skipped when si->stop), then the closing kernel that
// turns the sums into the residual, appends it to the history and decides whether the loop goes on; no synchronisation
int doma_launch_si_iteration_tail(doma_b200_ctx * c, int g, double tol, int max_it);
int doma_launch_mg_residual(doma_b200_ctx * c);
int doma_launch_k_sums(doma_b200_ctx * c, double k);
int doma_launch_sig_eff(doma_b200_ctx * c, int g);

// api.cu: (re)build sweep units, face planes and the tile table
int doma_build_units(doma_b200_ctx * c);

// sweep.cu
int doma_sweep_setup_tiles(doma_b200_ctx * c, int tx, int ty, int tz);
int doma_choose_tile(doma_b200_ctx * c);
bool doma_sweep_uses_ll(const doma_b200_ctx * c);
bool doma_sweep_uses_rows2(const doma_b200_ctx * c);
void doma_choose_rows(doma_b200_ctx * c);
// stop: device flag, non-zero = the kernel returns at once; timed = false keeps the events of an earlier launch
int doma_launch_sweep(doma_b200_ctx * c, int g, const int * stop = nullptr, bool timed = true);
int doma_sweep_check(doma_b200_ctx * c); // check mode: every tile done, phi_cur finite (synchronises)
