// sedi.cuh -- the gravitational (sedimentation) coagulation kernel of partmc/src/coag_kernel_sedi.F90 for the device:
//   kernel_sedi_helper :79-103   k = E(r1, r2) pi (r1 + r2)^2 |w(r1) - w(r2)|
//   fall_g             :110-173  terminal velocity in three size regimes (Stokes-Cunningham, two polynomial fits in
//                                log space), cgs inside, m/s out
//   effic              :180-291  collision efficiency: bilinear interpolation in the 15 x 21 table of sedi_table.cuh
// The kernel depends on the two volumes only (kernel_sedi :29-46 passes the particle volumes; spherical particles),
// and kernel_sedi_minmax (:52-73) returns the same value as minimum and maximum.  Integer powers x**(i-1) are
// formed by square-and-multiply in the order gfortran's __powidf2 uses, as in the oracle.
// Kept out of line: mc_coag is instruction-fetch bound and this body is only reached for PMCGPU_KERNEL_SEDI.
#pragma once
#include "sedi_table.cuh"

__device__ __forceinline__ double sedi_powi(double x, int m) {
  unsigned n = (unsigned)m;
  double y = (n & 1u) ? x : 1.0;
  while (n >>= 1) {
    x = x * x;
    if (n & 1u) y = y * x;
  }
  return y;
}

static __device__ __noinline__ double sedi_fall_g(double r) {
  const double b[7] = {-0.318657e1, 0.992696e0, -0.153193e-2, -0.987059e-3, -0.578878e-3, 0.855176e-4, -0.327815e-5};
  const double c[6] = {-0.500015e1, 0.523778e1, -0.204914e1, 0.475294e0, -0.542819e-1, 0.238449e-2};
  const double eta = 1.818e-4, xlamb = 6.62e-6, rhow = 1.0, rhoa = 1.225e-3, grav = 980.665;
  const double cunh = 1.257 * xlamb, t0 = 273.15;
  const double sigma = 76.1 - 0.155 * (293.15 - t0);
  const double stok = 2.0 * grav * (rhow - rhoa) / (9.0 * eta);
  const double stb = 32.0 * rhoa * (rhow - rhoa) * grav / (3.0 * eta * eta);
  const double phy = sigma * sigma * sigma * rhoa * rhoa / (sedi_powi(eta, 4) * grav * (rhow - rhoa));
  const double py = pow(phy, 1.0 / 6.0);
  const double rr = r * 1e2;
  double w_inf = 0.0;
  if (rr <= 1e-3) {
    w_inf = stok * (rr * rr + cunh * rr);
  } else if (rr <= 5.35e-2) {
    const double x = log(stb * rr * rr * rr);
    double y = 0.0;
#pragma unroll 1
    for (int i = 1; i <= 7; i++) y = y + b[i - 1] * sedi_powi(x, i - 1);
    const double xrey = (1.0 + cunh / rr) * exp(y);
    w_inf = xrey * eta / (2.0 * rhoa * rr);
  } else {
    double bond = grav * (rhow - rhoa) * (rr * rr) / sigma;
    if (rr > 0.35) bond = grav * (rhow - rhoa) * (0.35 * 0.35) / sigma;
    const double x = log(16.0 * bond * py / 3.0);
    double y = 0.0;
#pragma unroll 1
    for (int i = 1; i <= 6; i++) y = y + c[i - 1] * sedi_powi(x, i - 1);
    const double xrey = py * exp(y);
    w_inf = xrey * eta / (2.0 * rhoa * rr);
    if (rr > 0.35) w_inf = xrey * eta / (2.0 * rhoa * 0.35);
  }
  return w_inf / 100.0;
}

static __device__ __noinline__ double sedi_effic(double r1, double r2) {
  const double r_small = fmin(r1 * 1e6, r2 * 1e6), r_big = fmax(r1 * 1e6, r2 * 1e6);
  const double rq = r_small / r_big;
  int ir = 1, iq = 1;
  for (int k = 1; k <= 15; k++)
    if (r_big > c_sedi_r0[k - 1]) ir = k + 1;
  for (int kk = 1; kk <= 21; kk++)
    if (rq > c_sedi_rat[kk - 1]) iq = kk + 1;
#define SEDI_E(i, j) c_sedi_ecoll[((j) - 1) * 15 + ((i) - 1)]
  const double q = (rq - c_sedi_rat[iq - 2]) / (c_sedi_rat[iq - 1] - c_sedi_rat[iq - 2]);
  double ec;
  if (ir < 16) {
    if (ir >= 2) {
      const double p = (r_big - c_sedi_r0[ir - 2]) / (c_sedi_r0[ir - 1] - c_sedi_r0[ir - 2]);
      ec = (1.0 - p) * (1.0 - q) * SEDI_E(ir - 1, iq - 1) + p * (1.0 - q) * SEDI_E(ir, iq - 1) +
           q * (1.0 - p) * SEDI_E(ir - 1, iq) + p * q * SEDI_E(ir, iq);
    } else {
      ec = (1.0 - q) * SEDI_E(1, iq - 1) + q * SEDI_E(1, iq);
    }
  } else {
    const double ek = (1.0 - q) * SEDI_E(15, iq - 1) + q * SEDI_E(15, iq);
    ec = fmin(ek, 1.0);
  }
#undef SEDI_E
  return ec;
}

static __device__ __noinline__ double kernel_sedi(double v1, double v2) {
  const double r1 = vol2rad(v1), r2 = vol2rad(v2);
  const double winf1 = sedi_fall_g(r1), winf2 = sedi_fall_g(r2);
  const double ec = sedi_effic(r1, r2);
  return ec * PmcConst::pi * ((r1 + r2) * (r1 + r2)) * fabs(winf1 - winf2);
}
