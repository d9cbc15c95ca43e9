/*
 * TEST INFRASTRUCTURE — NOT PRODUCT CODE (see ls1_oracle.h).
 *
 * Scalar C restatement of the reference algorithm, pinned against outputs of the unmodified
 * reference (tests/golden/, tests/test_oracle_golden.py).  Compile with -ffp-contract=off: every
 * fused multiply-add the reference build performs on this path that matters for BIT-EXACT results
 * (cell borders) is written as an explicit fma() below.
 * Paths cited are relative to the reference tree (/root/reference).
 */
#include "ls1_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

/* ======================================================================================
 * Grid geometry: src/particleContainer/LinkedCells.cpp:42-107 (constructor, double rc) and
 * :136-204 (rebuild, float rc = _cutoffRadius/_cellsInCutoff).
 * ==================================================================================== */
int o_grid_init(o_grid* g, const double box_min[3], const double box_max[3], double rc, int mode) {
	g->cutoff = rc;
	for (int d = 0; d < 3; ++d) {
		g->box_min[d] = box_min[d];
		g->box_max[d] = box_max[d];
		double diff = box_max[d] - box_min[d];
		int n;
		if (mode == 1) {
			n = (int)floor(diff / rc);                 /* LinkedCells.cpp:53-54 */
			if (n == 0) n = 1;                         /* :56-58 */
		} else {
			float rcf = (float)(rc / 1);               /* :152  float rc = _cutoffRadius/_cellsInCutoff */
			n = (int)floor(diff / rcf);                /* :155  (double / float -> double division) */
			if (n == 0) return 1;                      /* :160-163 "region too small" */
		}
		g->box_cells[d] = n;
		g->cpd[d] = n + 2;                             /* one halo layer each side */
		g->cell_len[d] = diff / n;                     /* :61 / :168 */
		g->cell_len_recip[d] = n / diff;               /* :62 / :169 */
		g->halo_min[d] = box_min[d] - g->cell_len[d];  /* _haloLength = 1*_cellLength */
		g->halo_max[d] = box_max[d] + g->cell_len[d];
	}
	return 0;
}

/* CellBorderAndFlagManager::getBorder (CellBorderAndFlagManager.h:134-157).  The general branch
 * "index * _cellLength + _haloBoundingBoxMin" is contracted to an FMA by the reference's g++ -O3
 * -mfma build (C++ keeps -ffp-contract=fast under -std=c++17); pinned by tests/golden/cellindex_*.npz. */
static double border(const o_grid* g, int index, int d) {
	if (index == 0) return g->halo_min[d];
	if (index == 1) return g->box_min[d];
	if (index == g->cpd[d] - 1) return g->box_max[d];
	if (index == g->cpd[d]) return g->halo_max[d];
	return fma((double)index, g->cell_len[d], g->halo_min[d]);
}

/* LinkedCells::getCellIndexOfPoint (LinkedCells.cpp:830-882) + testPointInCell (ParticleCellBase.h:68-73) */
uint32_t o_cell_index(const o_grid* g, const double point[3]) {
	int ci[3];
	double p[3] = {point[0], point[1], point[2]};
	for (int d = 0; d < 3; ++d) {
		if (p[d] <= g->halo_min[d]) p[d] += g->cell_len[d] * 0.5;          /* :838-840 */
		else if (p[d] >= g->halo_max[d]) p[d] -= g->cell_len[d] * 0.5;     /* :841-843 */
		int c = (int)floor((p[d] - g->box_min[d]) * g->cell_len_recip[d]) + 1;  /* :858 */
		if (c < 0) c = 0;
		if (c > g->cpd[d] - 1) c = g->cpd[d] - 1;
		ci[d] = c;
	}
	int in = 1;
	for (int d = 0; d < 3; ++d) in &= (border(g, ci[d], d) <= p[d]) && (p[d] < border(g, ci[d] + 1, d));
	if (!in) {                                                              /* :866-877 */
		int c0[3] = {ci[0], ci[1], ci[2]};
		for (int d = 0; d < 3; ++d) {
			if (p[d] < border(g, c0[d], d)) ci[d]--;
			else if (p[d] >= border(g, c0[d] + 1, d)) ci[d]++;
		}
	}
	return (uint32_t)((ci[2] * g->cpd[1] + ci[1]) * g->cpd[0] + ci[0]);     /* :884-886 */
}

void o_cell_indices(const o_grid* g, size_t n, const double* pts, uint32_t* out) {
	for (size_t i = 0; i < n; ++i) out[i] = o_cell_index(g, pts + 3 * i);
}

/* halo flag: LinkedCells::initializeCells (LinkedCells.cpp:664-674) */
int o_is_halo_cell(const o_grid* g, uint32_t cell) {
	int x = cell % g->cpd[0];
	int y = (cell / g->cpd[0]) % g->cpd[1];
	int z = cell / (g->cpd[0] * g->cpd[1]);
	return x == 0 || y == 0 || z == 0 || x == g->cpd[0] - 1 || y == g->cpd[1] - 1 || z == g->cpd[2] - 1;
}

/* C08CellPairTraversal::traverseCellPairs (C08CellPairTraversal.h:50-80, :202-227) driving
 * C08BasedTraversals::processBaseCell (C08BasedTraversals.h:48-102) with the 14 offset pairs of
 * computeOffsets (:104-150). */
size_t o_c08_pairs(const o_grid* g, uint32_t* out, size_t capacity) {
	const long nx = g->cpd[0], ny = g->cpd[1], nz = g->cpd[2];
	const long o = 0, x = 1, y = nx, z = nx * ny, xy = x + y, yz = y + z, xz = x + z, xyz = x + y + z;
	const long pairs[14][2] = {{o, o}, {o, y}, {y, z}, {o, z}, {o, yz}, {x, yz}, {x, y}, {x, z},
	                           {o, x}, {o, xy}, {xy, z}, {y, xz}, {o, xz}, {o, xyz}};
	size_t cnt = 0;
	for (int col = 0; col < 8; ++col) {
		/* oneToThreeD(col, {2,2,2}): x fastest (utils/threeDimensionalMapping.h) */
		long bx = col % 2, by = (col / 2) % 2, bz = col / 4;
		for (long cz = bz; cz < nz - 1; cz += 2)
			for (long cy = by; cy < ny - 1; cy += 2)
				for (long cx = bx; cx < nx - 1; cx += 2) {
					long base = (cz * ny + cy) * nx + cx;
					for (int j = 0; j < 14; ++j) {
						uint32_t c1 = (uint32_t)(base + pairs[j][0]), c2 = (uint32_t)(base + pairs[j][1]);
						int h1 = o_is_halo_cell(g, c1), h2 = o_is_halo_cell(g, c2);
						if (h1 && h2) continue;                         /* :82-84 */
						uint32_t a = c1, b = c2, kind = 1;
						if (c1 == c2) kind = 0;                         /* :86-88 processCell */
						else if (h1) { a = c2; b = c1; }                /* :90-95 non-halo cell first */
						if (out && cnt < capacity) { out[3 * cnt] = a; out[3 * cnt + 1] = b; out[3 * cnt + 2] = kind; }
						++cnt;
					}
				}
	}
	return cnt;
}

/* ======================================================================================
 * Parameter tables: Comp2Param::initialize (src/molecules/Comp2Param.cpp:10-100) streamed into
 * VectorizedCellProcessor::_eps_sig / _shift6 (VectorizedCellProcessor.cpp:56-79), indexed by the
 * global LJ-site id = component look-up id + site (EnsembleBase.cpp:91-103).
 * ==================================================================================== */
static int lj_offset(const o_params* p, int c) { int o = 0; for (int k = 0; k < c; ++k) o += (int)p->counts[6 * k + 0]; return o; }
static int ch_offset(const o_params* p, int c) { int o = 0; for (int k = 0; k < c; ++k) o += (int)p->counts[6 * k + 1]; return o; }
static int di_offset(const o_params* p, int c) { int o = 0; for (int k = 0; k < c; ++k) o += (int)p->counts[6 * k + 2]; return o; }
static int qu_offset(const o_params* p, int c) { int o = 0; for (int k = 0; k < c; ++k) o += (int)p->counts[6 * k + 3]; return o; }

int o_lj_tables(const o_params* p, double* eps24, double* sig2, double* shift6, int cap) {
	int S = lj_offset(p, p->nc);
	if (S * S > cap) return -S;
	int mixpos = 0;
	for (int ci = 0; ci < p->nc; ++ci) {
		int nci = (int)p->counts[6 * ci], oi = lj_offset(p, ci);
		for (int a = 0; a < nci; ++a) {                                   /* Comp2Param.cpp:24-41 same component */
			const double* si = p->lj + 7 * (oi + a);
			for (int b = 0; b < nci; ++b) {
				const double* sj = p->lj + 7 * (oi + b);
				double e24 = 24. * sqrt(si[4] * sj[4]);
				double s2 = .5 * (si[5] + sj[5]); s2 *= s2;
				int k = (oi + a) * S + (oi + b);
				eps24[k] = e24; sig2[k] = s2; shift6[k] = si[6];          /* shift6 of site i (:28,40) */
			}
		}
		for (int cj = ci + 1; cj < p->nc; ++cj) {                         /* :43-98 unlike components */
			int ncj = (int)p->counts[6 * cj], oj = lj_offset(p, cj);
			double xi = p->mix[mixpos++], eta = p->mix[mixpos++];
			for (int a = 0; a < nci; ++a)
				for (int b = 0; b < ncj; ++b) {
					const double* si = p->lj + 7 * (oi + a);
					const double* sj = p->lj + 7 * (oj + b);
					double e24 = 24. * xi * sqrt(si[4] * sj[4]);
					double s2 = eta * .5 * (si[5] + sj[5]); s2 *= s2;
					double sr2 = s2 / (p->rc_lj * p->rc_lj);
					double sr6 = sr2 * sr2 * sr2;
					double sh = e24 * (sr6 - sr6 * sr6);
					int k = (oi + a) * S + (oj + b), kt = (oj + b) * S + (oi + a);
					eps24[k] = e24; sig2[k] = s2; shift6[k] = sh;
					/* the j->i stream recomputes with sqrt(epsi*epsj), same operand order (:78-96) */
					eps24[kt] = e24; sig2[kt] = s2; shift6[kt] = sh;
				}
		}
	}
	return S;
}

/* ======================================================================================
 * Periodic images on one process: DomainDecompBase::exchangeMolecules
 * (src/parallel/DomainDecompBase.cpp:51-86): handleDomainLeavingParticles (:174-225) for x,y,z then
 * populateHaloLayerWithCopies (:293-348) for x,y,z.  Halo copies of halo copies give edges/corners.
 * ==================================================================================== */
size_t o_exchange(const o_grid* g, const double global_len[3], size_t n, double* r, size_t capacity,
                  uint32_t* halo_src, double* halo_r) {
	(void)global_len;
	/* leaving particles: shift magnitude is the CONTAINER's box extent (:175) */
	for (int d = 0; d < 3; ++d) {
		double L = g->box_max[d] - g->box_min[d];
		for (size_t i = 0; i < n; ++i) {
			double* x = r + 3 * i + d;
			if (*x < g->box_min[d]) {                                   /* crossed the lower boundary: +L (:196-198,216-221) */
				*x = *x + L;
				if (*x >= g->box_max[d]) *x = nextafter(g->box_max[d], g->box_max[d] - 1.0);
			} else if (*x >= g->box_max[d]) {                           /* crossed the upper boundary: -L (:211-215) */
				*x = *x - L;
				if (*x <= g->box_min[d]) *x = g->box_min[d];
			}
		}
	}
	/* halo copies, dimension by dimension over owned molecules AND copies made so far (ALL_CELLS, :325) */
	size_t h = 0;
	for (int d = 0; d < 3; ++d) {
		double L = g->box_max[d] - g->box_min[d];
		size_t h_before = h;
		for (int dir = -1; dir <= 1; dir += 2) {
			double shift = dir < 0 ? L : -L;                            /* direction<0: lower slab, +L (:304) */
			double lo = dir < 0 ? g->box_min[d] : g->box_max[d] - g->cutoff;
			double hi = dir < 0 ? g->box_min[d] + g->cutoff : g->box_max[d];
			for (size_t k = 0; k < n + h_before; ++k) {
				const double* src = k < n ? r + 3 * k : halo_r + 3 * (k - n);
				/* region test in the other two dims is [min-rc, max+rc) which every molecule here satisfies */
				if (!(src[d] >= lo && src[d] < hi)) continue;           /* Molecule::inBox (MoleculeInterface.h:300-310) */
				double x = src[d] + shift;
				if (shift < 0) { if (x >= g->box_min[d]) x = nextafter(g->box_min[d], g->box_min[d] - 1.0); }  /* :331-335 */
				else { if (x < g->box_max[d]) x = nextafter(g->box_max[d], g->box_max[d] + 1.0); }           /* :336-341 */
				if (h < capacity) {
					halo_src[h] = k < n ? (uint32_t)k : halo_src[k - n];
					halo_r[3 * h + 0] = src[0]; halo_r[3 * h + 1] = src[1]; halo_r[3 * h + 2] = src[2];
					halo_r[3 * h + d] = x;
				}
				++h;
			}
		}
	}
	return h;
}

/* ======================================================================================
 * Force pass.
 * Site caches: FullParticleCell::buildSoACaches (src/particleContainer/FullParticleCell.cpp:143-193)
 * + FullMolecule::setupSoACache (src/molecules/FullMolecule.cpp:714-770): q is normalised, site world
 * position = R + rotate(q, d_body), orientation e = rotate(q, e_body); F, V, M cleared (clearFM :479-524).
 * ==================================================================================== */
/* Quaternion::rotate (src/molecules/Quaternion.cpp:43-62) */
static void q_rotate(const double q[4], const double d[3], double out[3]) {
	double ww = q[0]*q[0], xx = q[1]*q[1], yy = q[2]*q[2], zz = q[3]*q[3];
	double wx = q[0]*q[1], wy = q[0]*q[2], wz = q[0]*q[3], xy = q[1]*q[2], xz = q[1]*q[3], yz = q[2]*q[3];
	out[0] = (ww+xx-yy-zz)*d[0] + 2.*(xy-wz)*d[1] + 2.*(wy+xz)*d[2];
	out[1] = 2.*(wz+xy)*d[0] + (ww-xx+yy-zz)*d[1] + 2.*(yz-wx)*d[2];
	out[2] = 2.*(xz-wy)*d[0] + 2.*(wx+yz)*d[1] + (ww-xx-yy+zz)*d[2];
}
/* Quaternion::rotateinv (Quaternion.cpp:64-83) */
static void q_rotateinv(const double q[4], const double d[3], double out[3]) {
	double ww = q[0]*q[0], xx = q[1]*q[1], yy = q[2]*q[2], zz = q[3]*q[3];
	double wx = q[0]*q[1], wy = q[0]*q[2], wz = q[0]*q[3], xy = q[1]*q[2], xz = q[1]*q[3], yz = q[2]*q[3];
	out[0] = (ww+xx-yy-zz)*d[0] + 2.*(xy+wz)*d[1] + 2.*(xz-wy)*d[2];
	out[1] = 2.*(xy-wz)*d[0] + (ww-xx+yy-zz)*d[1] + 2.*(yz+wx)*d[2];
	out[2] = 2.*(xz+wy)*d[0] + 2.*(yz-wx)*d[1] + (ww-xx-yy+zz)*d[2];
}

enum { T_LJ = 0, T_CH = 1, T_DI = 2, T_QU = 3 };

typedef struct {
	int    type;
	int    ljid;        /* global LJ site id */
	double val;         /* q | absMy | absQ */
	double r[3];        /* world position */
	double d[3];        /* rotated body offset (site_d) */
	double e[3];        /* world orientation */
	double f[3], V[3], M[3];
} o_site;

typedef struct {
	const o_params* p;
	int     S;
	double *eps24, *sig2, *shift6;
	double  epsRFInvrc3;
	double  rc2, rclj2;
	const double* R;       /* molecule centres [n][3] */
	const size_t* soff;    /* site offset per molecule [n+1] */
	o_site* sites;
	double  upot6lj, upotXpoles, virial, myRF;
	uint64_t n_dist, n_pairs;
} o_fctx;

/* _loopBodyLJ (VectorizedCellProcessor.cpp:173-226) */
static void body_lj(o_fctx* c, const double* Ri, o_site* a, const double* Rj, o_site* b, int macro) {
	double dx = a->r[0]-b->r[0], dy = a->r[1]-b->r[1], dz = a->r[2]-b->r[2];
	double r2 = dx*dx + dy*dy + dz*dz;
	double inv = 1.0 / r2;
	int k = a->ljid * c->S + b->ljid;
	double lj2 = c->sig2[k] * inv, lj4 = lj2*lj2, lj6 = lj4*lj2, lj12 = lj6*lj6, lj12m6 = lj12 - lj6;
	double scale = (c->eps24[k] * inv) * (lj12 + lj12m6);
	double f[3] = {dx*scale, dy*scale, dz*scale};
	double V[3] = {(Ri[0]-Rj[0])*f[0], (Ri[1]-Rj[1])*f[1], (Ri[2]-Rj[2])*f[2]};
	for (int d = 0; d < 3; ++d) { a->f[d] += f[d]; b->f[d] -= f[d]; a->V[d] += V[d]; b->V[d] += V[d]; }
	if (macro) { c->upot6lj += c->eps24[k]*lj12m6 + c->shift6[k]; c->virial += V[0]+V[1]+V[2]; }
}

/* _loopBodyCharge (:229-279): a = charge of i, b = charge of j */
static void body_qq(o_fctx* c, const double* Ri, o_site* a, const double* Rj, o_site* b, int macro) {
	double dx = a->r[0]-b->r[0], dy = a->r[1]-b->r[1], dz = a->r[2]-b->r[2];
	double r2 = dx*dx + dy*dy + dz*dz;
	double inv2 = 1.0 / r2, inv1 = sqrt(inv2);
	double upot = a->val * b->val * inv1;
	double fac = upot * inv2;
	double f[3] = {dx*fac, dy*fac, dz*fac};
	double V[3] = {(Ri[0]-Rj[0])*f[0], (Ri[1]-Rj[1])*f[1], (Ri[2]-Rj[2])*f[2]};
	for (int d = 0; d < 3; ++d) { a->f[d] += f[d]; b->f[d] -= f[d]; a->V[d] += V[d]; b->V[d] += V[d]; }
	if (macro) { c->upotXpoles += upot; c->virial += V[0]+V[1]+V[2]; }
}

/* _loopBodyChargeDipole (:281-349): ch = charge site (molecule centre Rc), dp = dipole site (centre Rd).
 * f is the force on the charge; the dipole gets -f and the torque M. */
static void body_qd(o_fctx* c, const double* Rc, o_site* ch, const double* Rd, o_site* dp, int macro) {
	double dx = ch->r[0]-dp->r[0], dy = ch->r[1]-dp->r[1], dz = ch->r[2]-dp->r[2];
	double r2 = dx*dx + dy*dy + dz*dz;
	double inv2 = 1.0 / r2, inv1 = sqrt(inv2), inv3 = inv2 * inv1;
	const double* e = dp->e;
	double re = dx*e[0] + dy*e[1] + dz*e[2];
	double qp3 = ch->val * dp->val * inv3;
	double fac = inv2 * 3.0 * re;
	double f[3] = {qp3*(e[0] - dx*fac), qp3*(e[1] - dy*fac), qp3*(e[2] - dz*fac)};
	double V[3] = {(Rc[0]-Rd[0])*f[0], (Rc[1]-Rd[1])*f[1], (Rc[2]-Rd[2])*f[2]};
	for (int d = 0; d < 3; ++d) { ch->f[d] += f[d]; dp->f[d] -= f[d]; ch->V[d] += V[d]; dp->V[d] += V[d]; }
	if (macro) { c->upotXpoles -= qp3 * re; c->virial += V[0]+V[1]+V[2]; }
	dp->M[0] += qp3 * (e[1]*dz - e[2]*dy);
	dp->M[1] += qp3 * (e[2]*dx - e[0]*dz);
	dp->M[2] += qp3 * (e[0]*dy - e[1]*dx);
}

/* _loopBodyDipole (:351-444) */
static void body_dd(o_fctx* c, const double* Ri, o_site* a, const double* Rj, o_site* b, int macro) {
	double dx = a->r[0]-b->r[0], dy = a->r[1]-b->r[1], dz = a->r[2]-b->r[2];
	double r2 = dx*dx + dy*dy + dz*dz;
	double inv2 = 1.0 / r2, inv1 = sqrt(inv2);
	double three_inv2 = 3.0 * inv2;
	double p1p2 = a->val * b->val;
	double rffac = p1p2 * c->epsRFInvrc3;
	double p3 = p1p2 * inv1 * inv2;
	double p5x3 = p3 * three_inv2;
	const double *e1 = a->e, *e2 = b->e;
	double e1e2 = e1[0]*e2[0] + e1[1]*e2[1] + e1[2]*e2[2];
	double re1 = dx*e1[0] + dy*e1[1] + dz*e1[2];
	double re2 = dx*e2[0] + dy*e2[1] + dz*e2[2];
	double re1_3r2 = re1 * three_inv2, re2_3r2 = re2 * three_inv2;
	double re1re2r2 = inv2 * re1 * re2;
	double t = e1e2 - 5.0 * re1re2r2;
	double f[3] = {p5x3 * (dx*t + e1[0]*re2 + e2[0]*re1),
	               p5x3 * (dy*t + e1[1]*re2 + e2[1]*re1),
	               p5x3 * (dz*t + e1[2]*re2 + e2[2]*re1)};
	double V[3] = {(Ri[0]-Rj[0])*f[0], (Ri[1]-Rj[1])*f[1], (Ri[2]-Rj[2])*f[2]};
	for (int d = 0; d < 3; ++d) { a->f[d] += f[d]; b->f[d] -= f[d]; a->V[d] += V[d]; b->V[d] += V[d]; }
	if (macro) {
		c->upotXpoles += p3 * (e1e2 - 3.0 * re1re2r2);
		c->virial += V[0]+V[1]+V[2];
		c->myRF -= rffac * e1e2;                      /* accumulated negated (:2728-2729) */
	}
	double cx = e1[1]*e2[2] - e1[2]*e2[1], cy = e1[2]*e2[0] - e1[0]*e2[2], cz = e1[0]*e2[1] - e1[1]*e2[0]; /* e1 x e2 */
	a->M[0] += p3 * (re2_3r2 * (e1[1]*dz - e1[2]*dy) - cx) + rffac * cx;
	a->M[1] += p3 * (re2_3r2 * (e1[2]*dx - e1[0]*dz) - cy) + rffac * cy;
	a->M[2] += p3 * (re2_3r2 * (e1[0]*dy - e1[1]*dx) - cz) + rffac * cz;
	b->M[0] += p3 * (re1_3r2 * (e2[1]*dz - e2[2]*dy) + cx) - rffac * cx;
	b->M[1] += p3 * (re1_3r2 * (e2[2]*dx - e2[0]*dz) + cy) - rffac * cy;
	b->M[2] += p3 * (re1_3r2 * (e2[0]*dy - e2[1]*dx) + cz) - rffac * cz;
}

/* _loopBodyChargeQuadrupole (:446-526): ch = charge (centre Rc), qd = quadrupole (centre Rq);
 * f = force on the charge, M = torque on the quadrupole */
static void body_qQ(o_fctx* c, const double* Rc, o_site* ch, const double* Rq, o_site* qd, int macro) {
	double dx = ch->r[0]-qd->r[0], dy = ch->r[1]-qd->r[1], dz = ch->r[2]-qd->r[2];
	double r2 = dx*dx + dy*dy + dz*dz;
	double inv2 = 1.0 / r2, inv1 = sqrt(inv2);
	const double* e = qd->e;
	double qQ05 = 0.5 * ch->val * qd->val;
	double costj = (e[0]*dx + e[1]*dy + e[2]*dz) * inv1;
	double qQinv4dr3 = qQ05 * inv1 * inv2;
	double upot = qQinv4dr3 * (3.0 * costj * costj - 1.0);
	double minus_partialRijInvdr = 3.0 * upot * inv2;
	double partialTjInvdr = 6.0 * costj * qQinv4dr3 * inv1;
	double fac = costj * partialTjInvdr * inv1 + minus_partialRijInvdr;
	double f[3] = {fac*dx - partialTjInvdr*e[0], fac*dy - partialTjInvdr*e[1], fac*dz - partialTjInvdr*e[2]};
	double V[3] = {(Rc[0]-Rq[0])*f[0], (Rc[1]-Rq[1])*f[1], (Rc[2]-Rq[2])*f[2]};
	for (int d = 0; d < 3; ++d) { ch->f[d] += f[d]; qd->f[d] -= f[d]; ch->V[d] += V[d]; qd->V[d] += V[d]; }
	if (macro) { c->upotXpoles += upot; c->virial += V[0]+V[1]+V[2]; }
	qd->M[0] += partialTjInvdr * (e[2]*dy - e[1]*dz);
	qd->M[1] += partialTjInvdr * (e[0]*dz - e[2]*dx);
	qd->M[2] += partialTjInvdr * (e[1]*dx - e[0]*dy);
}

/* shared tail of the two-orientation bodies: f = fac*d + mTi*e1 + mTj*e2, torques (:622-649, :767-793) */
static void two_axis_tail(o_fctx* c, const double* R1, o_site* s1, const double* R2, o_site* s2, int macro,
                          double dx, double dy, double dz, double fac, double mTi, double mTj, double pGij, double upot) {
	const double *e1 = s1->e, *e2 = s2->e;
	double f[3] = {fac*dx + mTi*e1[0] + mTj*e2[0], fac*dy + mTi*e1[1] + mTj*e2[1], fac*dz + mTi*e1[2] + mTj*e2[2]};
	double V[3] = {(R1[0]-R2[0])*f[0], (R1[1]-R2[1])*f[1], (R1[2]-R2[2])*f[2]};
	for (int d = 0; d < 3; ++d) { s1->f[d] += f[d]; s2->f[d] -= f[d]; s1->V[d] += V[d]; s2->V[d] += V[d]; }
	if (macro) { c->upotXpoles += upot; c->virial += V[0]+V[1]+V[2]; }
	double gx = pGij * (e1[1]*e2[2] - e1[2]*e2[1]);
	double gy = pGij * (e1[2]*e2[0] - e1[0]*e2[2]);
	double gz = pGij * (e1[0]*e2[1] - e1[1]*e2[0]);
	s1->M[0] += mTi * (e1[1]*dz - e1[2]*dy) - gx;
	s1->M[1] += mTi * (e1[2]*dx - e1[0]*dz) - gy;
	s1->M[2] += mTi * (e1[0]*dy - e1[1]*dx) - gz;
	s2->M[0] += mTj * (e2[1]*dz - e2[2]*dy) + gx;
	s2->M[1] += mTj * (e2[2]*dx - e2[0]*dz) + gy;
	s2->M[2] += mTj * (e2[0]*dy - e2[1]*dx) + gz;
}

/* _loopBodyDipoleQuadrupole (:528-650): s1 = dipole, s2 = quadrupole; f = force on the dipole */
static void body_dQ(o_fctx* c, const double* R1, o_site* s1, const double* R2, o_site* s2, int macro) {
	double dx = s1->r[0]-s2->r[0], dy = s1->r[1]-s2->r[1], dz = s1->r[2]-s2->r[2];
	double r2 = dx*dx + dy*dy + dz*dz;
	double inv2 = 1.0 / r2, inv1 = sqrt(inv2);
	const double *e1 = s1->e, *e2 = s2->e;
	double myqfac = 1.5 * s1->val * s2->val * inv2 * inv2;
	double costi = (e1[0]*dx + e1[1]*dy + e1[2]*dz) * inv1;
	double costj = (e2[0]*dx + e2[1]*dy + e2[2]*dz) * inv1;
	double cos2tj = costj * costj;
	double cosgij = e1[0]*e2[0] + e1[1]*e2[1] + e1[2]*e2[2];
	double f5 = 5.0 * cos2tj - 1.0;
	double two_costj = 2.0 * costj;
	double upot = myqfac * (two_costj * cosgij - costi * f5);
	double myqfacXinvdr = myqfac * inv1;
	double mRij = 4.0 * upot * inv2;
	double mTi = myqfacXinvdr * f5;
	double mTj = myqfacXinvdr * 2.0 * (5.0 * costi * costj - cosgij);
	double pGij = myqfac * two_costj;
	double fac = mRij - (costi * mTi + costj * mTj) * inv1;
	two_axis_tail(c, R1, s1, R2, s2, macro, dx, dy, dz, fac, mTi, mTj, pGij, upot);
}

/* _loopBodyQuadrupole (:652-794) */
static void body_QQ(o_fctx* c, const double* R1, o_site* s1, const double* R2, o_site* s2, int macro) {
	double dx = s1->r[0]-s2->r[0], dy = s1->r[1]-s2->r[1], dz = s1->r[2]-s2->r[2];
	double r2 = dx*dx + dy*dy + dz*dz;
	double inv2 = 1.0 / r2, inv1 = sqrt(inv2);
	const double *e1 = s1->e, *e2 = s2->e;
	double qfac = 0.75 * inv1 * (s1->val * s2->val) * (inv2 * inv2);
	double costi = (e1[0]*dx + e1[1]*dy + e1[2]*dz) * inv1;
	double costj = (e2[0]*dx + e2[1]*dy + e2[2]*dz) * inv1;
	double cos2ti = costi*costi, cos2tj = costj*costj;
	double cosgij = e1[0]*e2[0] + e1[1]*e2[1] + e1[2]*e2[2];
	double term = cosgij - 5.0 * (costi * costj);
	double upot = qfac * ((1.0 + 2.0*term*term) - (5.0*(cos2ti + cos2tj) + 15.0*cos2ti*cos2tj));
	double mRij = 5.0 * upot * inv2;
	double part1 = qfac * 10.0 * inv1;
	double two_term = 2.0 * term;
	double mTi = part1 * (costi + (two_term*costj + 3.0*costi*cos2tj));
	double mTj = part1 * (costj + (two_term*costi + 3.0*costj*cos2ti));
	double pGij = qfac * 4.0 * term;
	double fac = mRij - (mTi*costi + mTj*costj) * inv1;
	two_axis_tail(c, R1, s1, R2, s2, macro, dx, dy, dz, fac, mTi, mTj, pGij, upot);
}

/* one molecule pair (i in soa1, j in soa2): VectorizedCellProcessor::_calculatePairs (:796-2732).
 * The cutoff mask is on the molecule CENTRES (calcDistLookup, SIMD_VectorizedCellProcessorHelpers.h:
 * 427-472): rc_LJ^2 for LJ sites, rc^2 for charges/dipoles/quadrupoles. */
static void pair_molecules(o_fctx* c, size_t i, size_t j, int macro, int same_cell) {
	const double *Ri = c->R + 3*i, *Rj = c->R + 3*j;
	double mx = Ri[0]-Rj[0], my = Ri[1]-Rj[1], mz = Ri[2]-Rj[2];
	double m2 = mx*mx + my*my + mz*mz;
	c->n_dist++;
	if (same_cell && m2 == 0.0) return;                  /* SingleCellPolicy_::GetForceMask (Helpers.h:367-376) */
	int in_lj = m2 < c->rclj2, in_rc = m2 < c->rc2;
	if (in_lj || in_rc) c->n_pairs++;
	o_site *si = c->sites + c->soff[i], *sj = c->sites + c->soff[j];
	size_t ni = c->soff[i+1] - c->soff[i], nj = c->soff[j+1] - c->soff[j];
	for (size_t a = 0; a < ni; ++a)
		for (size_t b = 0; b < nj; ++b) {
			o_site *A = si + a, *B = sj + b;
			int ta = A->type, tb = B->type;
			if (ta == T_LJ || tb == T_LJ) { if (ta == T_LJ && tb == T_LJ && in_lj) body_lj(c, Ri, A, Rj, B, macro); continue; }
			if (!in_rc) continue;
			if (ta == T_CH && tb == T_CH) body_qq(c, Ri, A, Rj, B, macro);
			else if (ta == T_CH && tb == T_DI) body_qd(c, Ri, A, Rj, B, macro);   /* q_i - mu_j (:1847-2006) */
			else if (ta == T_DI && tb == T_CH) body_qd(c, Rj, B, Ri, A, macro);   /* mu_i - q_j (:1323-1481) */
			else if (ta == T_DI && tb == T_DI) body_dd(c, Ri, A, Rj, B, macro);
			else if (ta == T_CH && tb == T_QU) body_qQ(c, Ri, A, Rj, B, macro);   /* q_i - Q_j (:2380-2538) */
			else if (ta == T_QU && tb == T_CH) body_qQ(c, Rj, B, Ri, A, macro);   /* Q_i - q_j (:1485-1650) */
			else if (ta == T_DI && tb == T_QU) body_dQ(c, Ri, A, Rj, B, macro);   /* mu_i - Q_j (:2540-2720) */
			else if (ta == T_QU && tb == T_DI) body_dQ(c, Rj, B, Ri, A, macro);   /* Q_i - mu_j (:2008-2192) */
			else if (ta == T_QU && tb == T_QU) body_QQ(c, Ri, A, Rj, B, macro);
		}
}

static void normalize_q(double* q) {   /* Quaternion::normalize (Quaternion.h:39-45): scale(1/sqrt(|q|^2)) */
	double s = 1. / sqrt(q[0]*q[0] + q[1]*q[1] + q[2]*q[2] + q[3]*q[3]);
	q[0] *= s; q[1] *= s; q[2] *= s; q[3] *= s;
}

int o_forces(const o_params* p, const o_grid* g, size_t n, const uint32_t* cid, const double* r, double* q,
             double* F, double* M, double* Vi, double* upot, double* virial, uint64_t* stats) {
	o_fctx c; memset(&c, 0, sizeof c);
	c.p = p; c.R = r;
	c.S = lj_offset(p, p->nc);
	size_t SS = (size_t)(c.S > 0 ? c.S * c.S : 1);
	c.eps24 = (double*)calloc(SS, 8); c.sig2 = (double*)calloc(SS, 8); c.shift6 = (double*)calloc(SS, 8);
	o_lj_tables(p, c.eps24, c.sig2, c.shift6, (int)SS);
	/* VectorizedCellProcessor.cpp:24 */
	c.epsRFInvrc3 = 2. * (p->eps_rf - 1.) / ((p->rc * p->rc * p->rc) * (2. * p->eps_rf + 1.));
	c.rc2 = p->rc * p->rc; c.rclj2 = p->rc_lj * p->rc_lj;

	/* site caches */
	size_t* soff = (size_t*)malloc((n + 1) * sizeof(size_t));
	soff[0] = 0;
	for (size_t i = 0; i < n; ++i) { const int64_t* k = p->counts + 6 * cid[i]; soff[i+1] = soff[i] + k[0] + k[1] + k[2] + k[3]; }
	o_site* sites = (o_site*)calloc(soff[n] ? soff[n] : 1, sizeof(o_site));
	c.soff = soff; c.sites = sites;
	for (size_t i = 0; i < n; ++i) {
		int ci = cid[i];
		const int64_t* k = p->counts + 6 * ci;
		normalize_q(q + 4*i);                                   /* setupSoACache -> normalizeQuaternion (:723) */
		o_site* s = sites + soff[i];
		const double* R = r + 3*i;
		for (int a = 0; a < k[0]; ++a, ++s) { const double* t = p->lj + 7*(lj_offset(p, ci) + a);
			s->type = T_LJ; s->ljid = (int)k[5] + a; q_rotate(q + 4*i, t, s->d); }
		for (int a = 0; a < k[1]; ++a, ++s) { const double* t = p->ch + 5*(ch_offset(p, ci) + a);
			s->type = T_CH; s->val = t[4]; q_rotate(q + 4*i, t, s->d); }
		for (int a = 0; a < k[2]; ++a, ++s) { const double* t = p->di + 7*(di_offset(p, ci) + a);
			s->type = T_DI; s->val = t[6]; q_rotate(q + 4*i, t, s->d); q_rotate(q + 4*i, t + 3, s->e); }
		for (int a = 0; a < k[3]; ++a, ++s) { const double* t = p->qu + 7*(qu_offset(p, ci) + a);
			s->type = T_QU; s->val = t[6]; q_rotate(q + 4*i, t, s->d); q_rotate(q + 4*i, t + 3, s->e); }
		for (s = sites + soff[i]; s < sites + soff[i+1]; ++s) for (int d = 0; d < 3; ++d) s->r[d] = s->d[d] + R[d];
	}

	/* bin molecules into cells (LinkedCells::update), keeping input order inside a cell */
	size_t ncells = (size_t)g->cpd[0] * g->cpd[1] * g->cpd[2];
	uint32_t* cell = (uint32_t*)malloc((n ? n : 1) * sizeof(uint32_t));
	size_t* cstart = (size_t*)calloc(ncells + 1, sizeof(size_t));
	for (size_t i = 0; i < n; ++i) { cell[i] = o_cell_index(g, r + 3*i); cstart[cell[i] + 1]++; }
	for (size_t k = 0; k < ncells; ++k) cstart[k+1] += cstart[k];
	size_t* order = (size_t*)malloc((n ? n : 1) * sizeof(size_t));
	size_t* fill = (size_t*)calloc(ncells, sizeof(size_t));
	for (size_t i = 0; i < n; ++i) order[cstart[cell[i]] + fill[cell[i]]++] = i;

	/* traversal */
	size_t np = o_c08_pairs(g, NULL, 0);
	uint32_t* trip = (uint32_t*)malloc((np ? np : 1) * 3 * sizeof(uint32_t));
	o_c08_pairs(g, trip, np);
	for (size_t t = 0; t < np; ++t) {
		uint32_t c1 = trip[3*t], c2 = trip[3*t+1];
		size_t n1 = cstart[c1+1] - cstart[c1], n2 = cstart[c2+1] - cstart[c2];
		if (trip[3*t+2] == 0) {                                    /* processCell (:2734-2744) */
			if (o_is_halo_cell(g, c1) || n1 < 2) continue;
			for (size_t a = 0; a < n1; ++a) for (size_t b = a + 1; b < n1; ++b)
				pair_molecules(&c, order[cstart[c1] + a], order[cstart[c1] + b], 1, 1);
		} else {                                                   /* processCellPair (:2746-2821) */
			if (n1 == 0 || n2 == 0) continue;
			int h1 = o_is_halo_cell(g, c1), h2 = o_is_halo_cell(g, c2);
			if (h1 && h2) continue;
			int macro = (!h1 && !h2) || (c1 < c2);
			int first_is_1 = n1 <= n2;                             /* iterate the smaller cell as soa1 (:2765) */
			uint32_t ca = first_is_1 ? c1 : c2, cb = first_is_1 ? c2 : c1;
			size_t na = cstart[ca+1] - cstart[ca], nb = cstart[cb+1] - cstart[cb];
			for (size_t a = 0; a < na; ++a) for (size_t b = 0; b < nb; ++b)
				pair_molecules(&c, order[cstart[ca] + a], order[cstart[cb] + b], macro, 0);
		}
	}

	/* FullMolecule::calcFM (src/molecules/FullMolecule.cpp:551-631) */
	for (size_t i = 0; i < n; ++i) {
		double f[3] = {0, 0, 0}, m[3] = {0, 0, 0}, tm[3] = {0, 0, 0}, tv[3] = {0, 0, 0};
		for (o_site* s = sites + soff[i]; s < sites + soff[i+1]; ++s) {
			for (int d = 0; d < 3; ++d) f[d] += s->f[d];
			m[0] += s->d[1]*s->f[2] - s->d[2]*s->f[1];
			m[1] += s->d[2]*s->f[0] - s->d[0]*s->f[2];
			m[2] += s->d[0]*s->f[1] - s->d[1]*s->f[0];
			for (int d = 0; d < 3; ++d) { tv[d] += s->V[d]; tm[d] += s->M[d]; }
		}
		for (int d = 0; d < 3; ++d) { F[3*i+d] = f[d]; M[3*i+d] = m[d] + tm[d]; Vi[3*i+d] = 0.5 * tv[d]; }
	}
	/* VectorizedCellProcessor::endTraversal (:124-157) */
	*virial = c.virial + 3.0 * c.myRF;
	*upot = c.upot6lj / 6.0 + c.upotXpoles + c.myRF;
	if (stats) { stats[0] = c.n_dist; stats[1] = c.n_pairs; }

	free(trip); free(fill); free(order); free(cstart); free(cell); free(sites); free(soff);
	free(c.eps24); free(c.sig2); free(c.shift6);
	return 0;
}

/* ======================================================================================
 * Integrator: Leapfrog (src/integrators/Leapfrog.cpp:48-155) calling
 * FullMolecule::upd_preF (src/molecules/FullMolecule.cpp:334-364) and upd_postF (:366-390).
 * ==================================================================================== */
static void comp_inertia(const o_params* p, int c, double I[3], double invI[3]) {
	for (int d = 0; d < 3; ++d) {                     /* FullMolecule::updateMassInertia (FullMolecule.h:88-101) */
		I[d] = p->scal[4*c + 1 + d];
		invI[d] = I[d] != 0. ? 1. / I[d] : 0.;
	}
}
/* Quaternion::differentiate (Quaternion.cpp:94-99) */
static void q_diff(const double q[4], const double w[3], double dq[4]) {
	dq[0] = .5 * (-q[1]*w[0] - q[2]*w[1] - q[3]*w[2]);
	dq[1] = .5 * ( q[0]*w[0] - q[3]*w[1] + q[2]*w[2]);
	dq[2] = .5 * ( q[3]*w[0] + q[0]*w[1] - q[1]*w[2]);
	dq[3] = .5 * (-q[2]*w[0] + q[1]*w[1] + q[0]*w[2]);
}

void o_upd_preF(const o_params* p, size_t n, const uint32_t* cid, double dt, double* r, double* v,
                double* q, double* L, const double* F, const double* M) {
	for (size_t i = 0; i < n; ++i) {
		double m = p->scal[4*cid[i]], I[3], invI[3];
		comp_inertia(p, cid[i], I, invI);
		double dt_halve = .5 * dt, dtInv2m = dt_halve / m;
		for (int d = 0; d < 3; ++d) { v[3*i+d] += dtInv2m * F[3*i+d]; r[3*i+d] += dt * v[3*i+d]; }
		double w[3], qh[4], dq[4];
		double* qi = q + 4*i;
		q_rotateinv(qi, L + 3*i, w);
		for (int d = 0; d < 3; ++d) w[d] *= invI[d];
		q_diff(qi, w, qh);
		for (int d = 0; d < 4; ++d) qh[d] = qh[d] * dt_halve + qi[d];
		double qcorr = 1. / sqrt(qh[0]*qh[0] + qh[1]*qh[1] + qh[2]*qh[2] + qh[3]*qh[3]);
		for (int d = 0; d < 4; ++d) qh[d] *= qcorr;
		for (int d = 0; d < 3; ++d) L[3*i+d] += dt_halve * M[3*i+d];
		q_rotateinv(qh, L + 3*i, w);
		for (int d = 0; d < 3; ++d) w[d] *= invI[d];
		q_diff(qh, w, dq);
		for (int d = 0; d < 4; ++d) qi[d] += dq[d] * dt;
		qcorr = 1. / sqrt(qi[0]*qi[0] + qi[1]*qi[1] + qi[2]*qi[2] + qi[3]*qi[3]);
		for (int d = 0; d < 4; ++d) qi[d] *= qcorr;
	}
}

void o_upd_postF(const o_params* p, size_t n, const uint32_t* cid, double dt_half, double* v, const double* q,
                 double* L, const double* F, const double* M, double* summv2, double* sumIw2, uint64_t* rot_dof) {
	double s1 = 0, s2 = 0; uint64_t rd = 0;
	for (size_t i = 0; i < n; ++i) {
		double m = p->scal[4*cid[i]], I[3], invI[3];
		comp_inertia(p, cid[i], I, invI);
		double dtInv2m = dt_half / m, v2 = 0;
		for (int d = 0; d < 3; ++d) { v[3*i+d] += dtInv2m * F[3*i+d]; v2 += v[3*i+d]*v[3*i+d]; L[3*i+d] += dt_half * M[3*i+d]; }
		s1 += m * v2;
		double w[3], Iw2 = 0;
		q_rotateinv(q + 4*i, L + 3*i, w);
		for (int d = 0; d < 3; ++d) { w[d] *= invI[d]; Iw2 += I[d] * w[d] * w[d]; }
		s2 += Iw2;
		rd += (uint64_t)p->counts[6*cid[i] + 4];
	}
	*summv2 = s1; *sumIw2 = s2; if (rot_dof) *rot_dof = rd;
}

/* Simulation::simulate order (src/Simulation.cpp:979-1177) on one periodic process; thermostat =
 * Domain::calculateGlobalValues betas (Domain.cpp:226-247) + VelocityScalingThermostat::apply. */
int o_run(const o_params* p, const o_grid* g, const double global_len[3], size_t n, const uint32_t* cid,
          double* r, double* v, double* q, double* L, double* F, double* M, int steps, double dt,
          double target_T, double* series) {
	size_t cap = 27 * n + 64;
	uint32_t* hsrc = (uint32_t*)malloc(cap * sizeof(uint32_t));
	double* hr = (double*)malloc(cap * 3 * 8);
	for (int s = 0; s < steps; ++s) {
		o_upd_preF(p, n, cid, dt, r, v, q, L, F, M);
		size_t h = o_exchange(g, global_len, n, r, cap, hsrc, hr);
		if (h > cap) { free(hsrc); free(hr); return 1; }
		size_t nt = n + h;
		uint32_t* cid2 = (uint32_t*)malloc(nt * 4);
		double* r2 = (double*)malloc(nt * 24); double* q2 = (double*)malloc(nt * 32);
		double* F2 = (double*)malloc(nt * 24); double* M2 = (double*)malloc(nt * 24); double* V2 = (double*)malloc(nt * 24);
		memcpy(cid2, cid, n * 4); memcpy(r2, r, n * 24); memcpy(q2, q, n * 32);
		for (size_t k = 0; k < h; ++k) { cid2[n+k] = cid[hsrc[k]]; memcpy(r2 + 3*(n+k), hr + 3*k, 24); memcpy(q2 + 4*(n+k), q + 4*hsrc[k], 32); }
		double upot, virial;
		o_forces(p, g, nt, cid2, r2, q2, F2, M2, V2, &upot, &virial, NULL);
		memcpy(F, F2, n * 24); memcpy(M, M2, n * 24); memcpy(q, q2, n * 32);
		free(cid2); free(r2); free(q2); free(F2); free(M2); free(V2);
		double summv2, sumIw2; uint64_t rd;
		o_upd_postF(p, n, cid, 0.5 * dt, v, q, L, F, M, &summv2, &sumIw2, &rd);
		if (series) { series[4*s] = upot; series[4*s+1] = virial; series[4*s+2] = summv2; series[4*s+3] = sumIw2; }
		if (target_T > 0) {
			double bt = pow(3.0 * n * target_T / summv2, 0.4);
			double br = sumIw2 == 0.0 ? 1.0 : pow(rd * target_T / sumIw2, 0.4);
			for (size_t i = 0; i < 3*n; ++i) { v[i] *= bt; L[i] *= br; }
		}
	}
	free(hsrc); free(hr);
	return 0;
}
