// cell_solve.cuh - the per-cell arithmetic of the sweep (include/TWDiamondDifference3D.h:38-91, TWDiamondDifference2D.h:36-71,
// DiamondDifference3D.h:35-40, DiamondDifference2D.h:34-40), shared by sweep.cu and the issue-ceiling microbenchmark
// scripts/cell_ceiling.cu.  Included inside an anonymous namespace.
#pragma once

// 1/d to ~2 ulp without the IEEE slow path: MUFU.RCP64H seed (~2^-20 relative) + one cubic step
// (error^3 ~ 2^-60, below the rounding of the three FMAs).  All denominators on this path are
// normal, finite and positive when the result is used (see inv_weight_v / cell_solve).
#define VLOOP _Pragma("unroll") for (int v = 0; v < V; ++v)

// The cell solve works on V ordinates in lockstep: every step is written for all V before the next
// step, so the V dependency chains are interleaved in program order and the FP64 pipe always has an
// independent instruction to issue.

// 1/d for V values
template <int V>
__device__ __forceinline__ void
fast_rcp_v(const double (&d)[V], double (&r)[V])
{
#ifdef DOMA_IEEE_DIV
  // debug build (scripts/build_exp.sh ieee -DDOMA_IEEE_DIV): correctly rounded reciprocals, to attribute the difference
  // between this path and the reference to its sources
  VLOOP r[v] = __drcp_rn(d[v]);
#else
  double e[V];
  VLOOP asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r[v]) : "d"(d[v]));
  VLOOP e[v] = fma(-d[v], r[v], 1.0);
  VLOOP e[v] = fma(e[v], e[v], e[v]);
  VLOOP r[v] = fma(r[v], e[v], r[v]);
#endif
}

// 1/a of the theta-weighted scheme: a = 1 - num/den clamped to [1/2, 1]  ->  1/a = den/(den-num)
// clamped to [1, 2]; a = 1/2 when the upwind flux is <= 1e-8 (include/TWDiamondDifference3D.h:43-70).
// All the case analysis runs on the integer pipe, off the FP64 pipe that bounds the kernel:
//   * up > 1e-8 is an exact signed 64-bit compare of the bit patterns (order-preserving for
//     doubles >= 0; negative values have the sign bit and compare below);
//   * where the upwind flux fails that test the divisor's high word is zeroed, so the flushed
//     reciprocal is +inf and the quotient lands in the "2" case below whatever den is;
//   * the clamp of r = den/(den-num) to [1,2] looks at r's high word only: 1.0 and 2.0 are powers of
//     two, so r < 1 <=> hi < 0x3FF00000 and (r >= 2, r < 0, inf or NaN) <=> unsigned hi >= 0x40000000.
//     (den-num <= den/2, which includes <= 0, is exactly the r >= 2 or r < 0 case.)
template <int V>
__device__ __forceinline__ void
inv_weight_v(const double (&up)[V], const double (&den)[V], const double (&dmn)[V], double (&r)[V])
{
  double d[V];
  VLOOP
  {
    const bool use = __double_as_longlong(up[v]) > 0x3E45798EE2308C3ALL; // bits of 1e-8
    d[v] = __hiloint2double(use ? __double2hiint(dmn[v]) : 0, __double2loint(dmn[v]));
  }
  fast_rcp_v<V>(d, r);
  VLOOP r[v] = den[v] * r[v];
  VLOOP
  {
    const unsigned hi = (unsigned)__double2hiint(r[v]);
    const bool inside = hi - 0x3FF00000u < 0x00100000u;
    const unsigned chi = min(max(hi, 0x3FF00000u), 0x40000000u);
    r[v] = __hiloint2double((int)chi, inside ? __double2loint(r[v]) : 0);
  }
}

// num/den to ~3 ulp; den > 0 always (it contains |mu|/dx + |eta|/dy + ...)
template <int V>
__device__ __forceinline__ void
fast_div_v(const double (&num)[V], const double (&den)[V], double (&out)[V])
{
  double r[V];
  fast_rcp_v<V>(den, r);
  VLOOP out[v] = num[v] * r[v];
}

// One cell, V ordinates.  m = |mu|/dx, e = |eta|/dy, z_ = |xi|/dz (3D) or |xi|/(dx dy) (2D TWDD
// quirk).  x, y, z: upwind face fluxes in, downwind face fluxes out.
// mid(): called once, after the three weights and before the final division - the point where the fewest values are live;
// the sweep's row loop uses it to start the shared-memory loads of the NEXT cell a full reciprocal ahead of their use
// (the compiler cannot hoist them itself across this cell's shared-memory stores).
struct NoMid
{
  __device__ __forceinline__ void operator()() const {}
};

template <int DIMS, int SCHEME, int V, class Mid = NoMid>
__device__ __forceinline__ void
cell_solve(const double (&q)[V], const double (&sig)[V], const double (&m)[V], const double (&e)[V], const double (&z_)[V],
           double (&x)[V], double (&y)[V], double (&z)[V], double (&psi)[V], Mid mid = Mid())
{
  constexpr double th = 0.9, ths = 0.9, omt = 1.0 - 0.9;
  if (SCHEME == DOMA_SCHEME_TWDD)
  {
    if (DIMS == 3)
    {
      double tx[V], ty[V], tz[V], base[V], da[V], db[V], dc[V], ra[V], rb[V], rc[V], num[V], den[V], qs[V];
      VLOOP qs[v] = q[v] * ths;
      VLOOP tx[v] = m[v] * x[v];
      VLOOP ty[v] = e[v] * y[v];
      VLOOP tz[v] = z_[v] * z[v];
      VLOOP base[v] = fma(th, tx[v] + ty[v] + tz[v], qs[v]);
      VLOOP
      {
        const double e2 = e[v] + e[v], z2 = z_[v] + z_[v]; // row / slab invariants, hoisted by the compiler
        const double sm = fma(2.0, m[v], sig[v]);
        da[v] = (sig[v] + (e2 + z2)) * x[v];
        db[v] = (sm + z2) * y[v];
        dc[v] = (sm + e2) * z[v];
      }
      // den - num for the three weights
      VLOOP tx[v] = da[v] - fma(omt, tx[v], base[v]);
      VLOOP ty[v] = db[v] - fma(omt, ty[v], base[v]);
      VLOOP tz[v] = dc[v] - fma(omt, tz[v], base[v]);
      inv_weight_v<V>(x, da, tx, ra);
      inv_weight_v<V>(y, db, ty, rb);
      inv_weight_v<V>(z, dc, tz, rc);
      VLOOP
      {
        const double ca = m[v] * ra[v], cb = e[v] * rb[v], cc = z_[v] * rc[v];
        num[v] = fma(cc, z[v], fma(cb, y[v], fma(ca, x[v], q[v])));
        den[v] = sig[v] + ca + cb + cc;
      }
      mid();
      fast_div_v<V>(num, den, psi);
      VLOOP x[v] = fma(ra[v], psi[v] - x[v], x[v]);
      VLOOP y[v] = fma(rb[v], psi[v] - y[v], y[v]);
      VLOOP z[v] = fma(rc[v], psi[v] - z[v], z[v]);
    }
    else
    {
      // include/TWDiamondDifference2D.h:36-71; z_ carries the bare "+ abs_xi" term divided by the area
      double tx[V], ty[V], da[V], db[V], ra[V], rb[V], num[V], den[V], qs[V];
      VLOOP qs[v] = q[v] * ths;
      VLOOP tx[v] = m[v] * x[v];
      VLOOP ty[v] = e[v] * y[v];
      VLOOP da[v] = (sig[v] + (e[v] + e[v])) * x[v];
      VLOOP db[v] = (sig[v] + (m[v] + m[v])) * y[v];
      double dna[V], dnb[V];
      VLOOP dna[v] = da[v] - (fma(th, ty[v], qs[v]) + tx[v]);
      VLOOP dnb[v] = db[v] - (fma(th, tx[v] + z_[v], qs[v]) + ty[v]);
      inv_weight_v<V>(x, da, dna, ra);
      inv_weight_v<V>(y, db, dnb, rb);
      VLOOP
      {
        const double ca = m[v] * ra[v], cb = e[v] * rb[v];
        num[v] = fma(cb, y[v], fma(ca, x[v], q[v]));
        den[v] = sig[v] + ca + cb;
      }
      mid();
      fast_div_v<V>(num, den, psi);
      VLOOP x[v] = fma(ra[v], psi[v] - x[v], x[v]);
      VLOOP y[v] = fma(rb[v], psi[v] - y[v], y[v]);
    }
  }
  else
  {
    // plain diamond difference (include/DiamondDifference3D.h:35-40, DiamondDifference2D.h:34-40)
    double num[V], den[V];
    VLOOP
    {
      const double m2 = m[v] + m[v], e2 = e[v] + e[v], z2 = DIMS == 3 ? z_[v] + z_[v] : 0.0;
      num[v] = fma(e2, y[v], fma(m2, x[v], q[v]));
      den[v] = m2 + e2 + sig[v];
      if (DIMS == 3)
      {
        num[v] = fma(z2, z[v], num[v]);
        den[v] = m2 + e2 + z2 + sig[v];
      }
    }
    mid();
    fast_div_v<V>(num, den, psi);
    VLOOP x[v] = 2.0 * psi[v] - x[v];
    VLOOP y[v] = 2.0 * psi[v] - y[v];
    if (DIMS == 3)
      VLOOP z[v] = 2.0 * psi[v] - z[v];
  }
}
