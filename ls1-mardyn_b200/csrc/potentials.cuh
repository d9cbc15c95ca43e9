// potentials.cuh — the seven site-site interactions of VectorizedCellProcessor as device functions.
//
// Each function follows the corresponding _loopBody* of the reference
// (src/particleContainer/adapter/VectorizedCellProcessor.cpp) and returns the force on site 1 ("f"), the
// torques on the oriented sites and the pair potential.  Newton-3 partners are obtained by the caller as
// -f (owner-computes: every molecule accumulates only its own sites).
//   _loopBodyLJ                :173-226      pot_lj
//   _loopBodyCharge            :229-279      pot_qq
//   _loopBodyChargeDipole      :281-349      pot_q_mu     (site 1 = charge, site 2 = dipole)
//   _loopBodyDipole            :351-444      pot_mu_mu
//   _loopBodyChargeQuadrupole  :446-526      pot_q_Q      (site 1 = charge, site 2 = quadrupole)
//   _loopBodyDipoleQuadrupole  :528-650      pot_mu_Q     (site 1 = dipole, site 2 = quadrupole)
//   _loopBodyQuadrupole        :652-794      pot_Q_Q
#pragma once

// reciprocal of a positive, normal double: hardware seed (MUFU.RCP64H) + two Newton-Raphson steps (within ~1 ulp).  The reference itself uses NR-refined estimates on AVX2/AVX-512
// (vectorization/RealVecDouble.h:411-541); parity tolerance is 1e-10 relative.
__device__ __forceinline__ double fast_rcp(double a) {
	double x;
	asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(a));
	// seed error <= 2^-20 (PTX ISA, rcp.approx.ftz.f64) -> 2^-40 -> 2^-80: two steps reach the rounding level
	double e = fma(-a, x, 1.0);
	x = fma(x, e, x);
	e = fma(-a, x, 1.0);
	x = fma(x, e, x);
	return x;
}
// same seed, ONE third-order step: x1 = x0 (1 + e + e^2), e = 1 - a x0; |e| <= 2^-20 -> error e^3 <= 2^-60: full double
// precision with 3 dependent FMAs instead of 4 (used by the list force kernel, where the FP64 pipe is the bound)
__device__ __forceinline__ double fast_rcp3(double a) {
	double x;
	asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(a));
	const double e = fma(-a, x, 1.0);
	const double t = fma(e, e, e);
	return fma(x, t, x);
}
// 1/sqrt(a): hardware seed (MUFU.RSQ64H, relative error <= 2^-20) + ONE third-order step:
// e = 1 - a y0^2,  y = y0 (1 + e/2 + 3 e^2/8),  error (5/16) e^3 <= 2^-61: full double precision with 5 FP64 instructions
__device__ __forceinline__ double fast_rsqrt(double a) {
	double x;
	asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(a));
	const double t = a * x;
	const double e = fma(-t, x, 1.0);
	const double p = fma(0.375, e, 0.5) * e;
	return fma(x, p, x);
}

struct V3 { double x, y, z; };
__device__ __forceinline__ V3 v3(double x, double y, double z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ double dot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(const V3& a, const V3& b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }

// d = r1 - r2 (site positions).  All functions: f = force on site 1.
__device__ __forceinline__ void pot_lj(const V3& d, double eps24, double sig2, double shift6, V3& f, double& u6) {
	double r2 = dot(d, d);
	double inv = fast_rcp(r2);
	double lj2 = sig2 * inv, lj4 = lj2 * lj2, lj6 = lj4 * lj2, lj12 = lj6 * lj6, lj12m6 = lj12 - lj6;
	double scale = (eps24 * inv) * (lj12 + lj12m6);
	f = v3(d.x * scale, d.y * scale, d.z * scale);
	u6 = eps24 * lj12m6 + shift6;
}

__device__ __forceinline__ void pot_qq(const V3& d, double q1q2, V3& f, double& u) {
	double r2 = dot(d, d);
	double inv1 = fast_rsqrt(r2), inv2 = inv1 * inv1;
	u = q1q2 * inv1;
	double fac = u * inv2;
	f = v3(d.x * fac, d.y * fac, d.z * fac);
}

// site 1 = charge q, site 2 = dipole (e, mu); M2 = torque on the dipole
__device__ __forceinline__ void pot_q_mu(const V3& d, double q, const V3& e, double mu, V3& f, V3& M2, double& u) {
	double r2 = dot(d, d);
	double inv1 = fast_rsqrt(r2), inv2 = inv1 * inv1, inv3 = inv2 * inv1;
	double re = dot(d, e);
	double qp3 = q * mu * inv3;
	double fac = inv2 * 3.0 * re;
	f = v3(qp3 * (e.x - d.x * fac), qp3 * (e.y - d.y * fac), qp3 * (e.z - d.z * fac));
	u = -qp3 * re;
	M2 = v3(qp3 * (e.y * d.z - e.z * d.y), qp3 * (e.z * d.x - e.x * d.z), qp3 * (e.x * d.y - e.y * d.x));
}

// dipole-dipole incl. reaction field; rf = rffac * (e1.e2) (caller accumulates it NEGATED, :2728-2729)
__device__ __forceinline__ void pot_mu_mu(const V3& d, const V3& e1, double mu1, const V3& e2, double mu2, double epsRFInvrc3,
                                           V3& f, V3& M1, V3& M2, double& u, double& rf) {
	double r2 = dot(d, d);
	double inv1 = fast_rsqrt(r2), inv2 = inv1 * inv1;
	double three_inv2 = 3.0 * inv2;
	double p1p2 = mu1 * mu2;
	double rffac = p1p2 * epsRFInvrc3;
	double p3 = p1p2 * inv1 * inv2;
	double p5x3 = p3 * three_inv2;
	double e1e2 = dot(e1, e2), re1 = dot(d, e1), re2 = dot(d, e2);
	double re1_3r2 = re1 * three_inv2, re2_3r2 = re2 * three_inv2;
	double re1re2r2 = inv2 * re1 * re2;
	double t = e1e2 - 5.0 * re1re2r2;
	f = v3(p5x3 * (d.x * t + e1.x * re2 + e2.x * re1),
	       p5x3 * (d.y * t + e1.y * re2 + e2.y * re1),
	       p5x3 * (d.z * t + e1.z * re2 + e2.z * re1));
	u = p3 * (e1e2 - 3.0 * re1re2r2);
	rf = rffac * e1e2;
	V3 c = cross(e1, e2);
	V3 e1d = cross(e1, d), e2d = cross(e2, d);
	M1 = v3(p3 * (re2_3r2 * e1d.x - c.x) + rffac * c.x, p3 * (re2_3r2 * e1d.y - c.y) + rffac * c.y, p3 * (re2_3r2 * e1d.z - c.z) + rffac * c.z);
	M2 = v3(p3 * (re1_3r2 * e2d.x + c.x) - rffac * c.x, p3 * (re1_3r2 * e2d.y + c.y) - rffac * c.y, p3 * (re1_3r2 * e2d.z + c.z) - rffac * c.z);
}

// site 1 = charge q, site 2 = quadrupole (e, Q); M2 = torque on the quadrupole
__device__ __forceinline__ void pot_q_Q(const V3& d, double q, const V3& e, double Q, V3& f, V3& M2, double& u) {
	double r2 = dot(d, d);
	double inv1 = fast_rsqrt(r2), inv2 = inv1 * inv1;
	double qQ05 = 0.5 * q * Q;
	double costj = dot(e, d) * inv1;
	double qQinv4dr3 = qQ05 * inv1 * inv2;
	u = qQinv4dr3 * (3.0 * costj * costj - 1.0);
	double minus_partialRijInvdr = 3.0 * u * inv2;
	double partialTjInvdr = 6.0 * costj * qQinv4dr3 * inv1;
	double fac = costj * partialTjInvdr * inv1 + minus_partialRijInvdr;
	f = v3(fac * d.x - partialTjInvdr * e.x, fac * d.y - partialTjInvdr * e.y, fac * d.z - partialTjInvdr * e.z);
	M2 = v3(partialTjInvdr * (e.z * d.y - e.y * d.z), partialTjInvdr * (e.x * d.z - e.z * d.x), partialTjInvdr * (e.y * d.x - e.x * d.y));
}

// shared tail of the two-axis interactions (:622-649, :767-793)
__device__ __forceinline__ void two_axis_tail(const V3& d, const V3& e1, const V3& e2, double fac, double mTi, double mTj, double pGij,
                                               V3& f, V3& M1, V3& M2) {
	f = v3(fac * d.x + mTi * e1.x + mTj * e2.x, fac * d.y + mTi * e1.y + mTj * e2.y, fac * d.z + mTi * e1.z + mTj * e2.z);
	V3 c = cross(e1, e2);
	V3 g = v3(pGij * c.x, pGij * c.y, pGij * c.z);
	V3 e1d = cross(e1, d), e2d = cross(e2, d);
	M1 = v3(mTi * e1d.x - g.x, mTi * e1d.y - g.y, mTi * e1d.z - g.z);
	M2 = v3(mTj * e2d.x + g.x, mTj * e2d.y + g.y, mTj * e2d.z + g.z);
}

// site 1 = dipole (e1, mu), site 2 = quadrupole (e2, Q)
__device__ __forceinline__ void pot_mu_Q(const V3& d, const V3& e1, double mu, const V3& e2, double Q, V3& f, V3& M1, V3& M2, double& u) {
	double r2 = dot(d, d);
	double inv1 = fast_rsqrt(r2), inv2 = inv1 * inv1;
	double myqfac = 1.5 * mu * Q * inv2 * inv2;
	double costi = dot(e1, d) * inv1, costj = dot(e2, d) * inv1;
	double cos2tj = costj * costj;
	double cosgij = dot(e1, e2);
	double f5 = 5.0 * cos2tj - 1.0;
	double two_costj = 2.0 * costj;
	u = myqfac * (two_costj * cosgij - costi * f5);
	double myqfacXinvdr = myqfac * inv1;
	double mRij = 4.0 * u * inv2;
	double mTi = myqfacXinvdr * f5;
	double mTj = myqfacXinvdr * 2.0 * (5.0 * costi * costj - cosgij);
	double pGij = myqfac * two_costj;
	double fac = mRij - (costi * mTi + costj * mTj) * inv1;
	two_axis_tail(d, e1, e2, fac, mTi, mTj, pGij, f, M1, M2);
}

__device__ __forceinline__ void pot_Q_Q(const V3& d, const V3& e1, double Q1, const V3& e2, double Q2, V3& f, V3& M1, V3& M2, double& u) {
	double r2 = dot(d, d);
	double inv1 = fast_rsqrt(r2), inv2 = inv1 * inv1;
	double qfac = 0.75 * inv1 * (Q1 * Q2) * (inv2 * inv2);
	double costi = dot(e1, d) * inv1, costj = dot(e2, d) * inv1;
	double cos2ti = costi * costi, cos2tj = costj * costj;
	double cosgij = dot(e1, e2);
	double term = cosgij - 5.0 * (costi * costj);
	u = qfac * ((1.0 + 2.0 * term * term) - (5.0 * (cos2ti + cos2tj) + 15.0 * cos2ti * cos2tj));
	double mRij = 5.0 * u * inv2;
	double part1 = qfac * 10.0 * inv1;
	double two_term = 2.0 * term;
	double mTi = part1 * (costi + (two_term * costj + 3.0 * costi * cos2tj));
	double mTj = part1 * (costj + (two_term * costi + 3.0 * costj * cos2ti));
	double pGij = qfac * 4.0 * term;
	double fac = mRij - (mTi * costi + mTj * costj) * inv1;
	two_axis_tail(d, e1, e2, fac, mTi, mTj, pGij, f, M1, M2);
}
