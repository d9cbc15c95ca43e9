// pair_body.cuh — all site-site interactions of one molecule pair, seen from the owner molecule: the ten ordered loops of
// VectorizedCellProcessor::_calculatePairs (VectorizedCellProcessor.cpp:1031-2720) with calcFM folded in (FullMolecule.cpp:551-631).
// Shared by the cell-scan kernel (force.cu) and the list-driven kernel (lists.cu).
#pragma once
#include "potentials.cuh"

struct ForceArgs {
	Grid g;
	size_t n;
	const double4* pos; const uint32_t* cell; const uint32_t* cell_start;
	const CompDev* comps;
	const double4* cache; int stride;
	const double *eps24, *sig2, *shift6; int S;
	double rc2, rclj2, epsRFInvrc3;
	double *F, *M, *Vi;
	double* partials;               // [blocks][5]: upot6lj, upotXpoles, virial, myRF, pairs
	int bpr, nyi;                   // blocks per owned cell row, owned rows per z layer
	double eps24_c[16], sig2_c[16], shift6_c[16];   // single-component kernels (S <= 4): the LJ tables travel in the kernel parameters (constant bank)
};

__device__ __forceinline__ V3 ld3(const double4& a) { return v3(a.x, a.y, a.z); }

// Accumulators of one owned molecule (registers of its thread).
struct MolAcc {
	V3 F, M, V;
	double u6, uX, rf;
	unsigned long long np;
};

// All site-site interactions of the molecule pair (i, j): the ten ordered loops of _calculatePairs
// (VectorizedCellProcessor.cpp:1031-2720) seen from molecule i; m = R_i - R_j (molecule centres), in_lj / in_rc the two
// molecule-centre cutoff decisions (:1013-1024).
// lj = the LJ parameter tables [3][LJ_TAB_STRIDE or S * S]: eps24 | sig2 | shift6, in shared memory when they fit (lj_tables)
#define LJ_TAB_MAX 64       // S * S entries per table that the shared-memory copy holds (S <= 8 LJ site types; static shared memory is 48 KB)
struct LJTab { const double *eps24, *sig2, *shift6; };
// mixtures: every block copies the tables into shared memory once (tab = 3 * LJ_TAB_MAX doubles); larger tables stay in global memory
__device__ __forceinline__ LJTab lj_tables(const ForceArgs& a, double* tab) {
	LJTab t; t.eps24 = a.eps24; t.sig2 = a.sig2; t.shift6 = a.shift6;
	const int n = a.S * a.S;
	if (n <= LJ_TAB_MAX) {
		for (int k = threadIdx.x; k < n; k += blockDim.x) { tab[k] = a.eps24[k]; tab[LJ_TAB_MAX + k] = a.sig2[k]; tab[2 * LJ_TAB_MAX + k] = a.shift6[k]; }
		__syncthreads();
		t.eps24 = tab; t.sig2 = tab + LJ_TAB_MAX; t.shift6 = tab + 2 * LJ_TAB_MAX;
	}
	return t;
}

__device__ __forceinline__ void pair_body(const ForceArgs& a, const LJTab& lj, const V3& Ri, const CompDev& ci, const double4* __restrict__ si,
                                          const int i_ch0, const int i_di0, const int i_qu0,
                                          const uint32_t j, const double4& Pj, const V3& m, const bool in_lj, const bool in_rc, MolAcc& A_) {
	V3& F = A_.F; V3& M = A_.M; V3& V = A_.V; double& u6 = A_.u6; double& uX = A_.uX; double& rf = A_.rf;
	{
		const CompDev cj = a.comps[pos_cid(Pj)];
		const double4* sj = a.cache + (size_t)j * a.stride;
		const int j_ch0 = cj.n_lj, j_di0 = cj.n_lj + cj.n_ch, j_qu0 = j_di0 + 2 * cj.n_di;
		// ---- LJ - LJ (VectorizedCellProcessor.cpp:1031-1180)
		if (in_lj)
			for (int s1 = 0; s1 < ci.n_lj; ++s1) {
				const double4 A = si[s1];
				const int ida = (int)__double_as_longlong(A.w);
				V3 fa = v3(0, 0, 0);
				for (int s2 = 0; s2 < cj.n_lj; ++s2) {
					const double4 B = sj[s2];
					const int k = ida * a.S + (int)__double_as_longlong(B.w);
					V3 f; double u;
					pot_lj(v3(A.x - B.x, A.y - B.y, A.z - B.z), lj.eps24[k], lj.sig2[k], lj.shift6[k], f, u);
					fa.x += f.x; fa.y += f.y; fa.z += f.z; u6 += u;
				}
				F.x += fa.x; F.y += fa.y; F.z += fa.z;
				V.x += m.x * fa.x; V.y += m.y * fa.y; V.z += m.z * fa.z;
				V3 t = cross(v3(A.x - Ri.x, A.y - Ri.y, A.z - Ri.z), fa);
				M.x += t.x; M.y += t.y; M.z += t.z;
			}
		if (!in_rc) return;
		// ---- charges of i
		for (int s1 = 0; s1 < ci.n_ch; ++s1) {
			const double4 A = si[i_ch0 + s1];
			V3 fa = v3(0, 0, 0);
			for (int s2 = 0; s2 < cj.n_ch; ++s2) {                     // q - q (:1192-1319)
				const double4 B = sj[j_ch0 + s2];
				V3 f; double u;
				pot_qq(v3(A.x - B.x, A.y - B.y, A.z - B.z), A.w * B.w, f, u);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u;
			}
			for (int s2 = 0; s2 < cj.n_di; ++s2) {                     // q_i - mu_j (:1847-2006)
				const double4 B = sj[j_di0 + 2 * s2], E = sj[j_di0 + 2 * s2 + 1];
				V3 f, M2; double u;
				pot_q_mu(v3(A.x - B.x, A.y - B.y, A.z - B.z), A.w, ld3(E), B.w, f, M2, u);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u;
			}
			for (int s2 = 0; s2 < cj.n_qu; ++s2) {                     // q_i - Q_j (:2380-2538)
				const double4 B = sj[j_qu0 + 2 * s2], E = sj[j_qu0 + 2 * s2 + 1];
				V3 f, M2; double u;
				pot_q_Q(v3(A.x - B.x, A.y - B.y, A.z - B.z), A.w, ld3(E), B.w, f, M2, u);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u;
			}
			F.x += fa.x; F.y += fa.y; F.z += fa.z;
			V.x += m.x * fa.x; V.y += m.y * fa.y; V.z += m.z * fa.z;
			V3 t = cross(v3(A.x - Ri.x, A.y - Ri.y, A.z - Ri.z), fa);
			M.x += t.x; M.y += t.y; M.z += t.z;
		}
		// ---- dipoles of i
		for (int s1 = 0; s1 < ci.n_di; ++s1) {
			const double4 A = si[i_di0 + 2 * s1]; const V3 e1 = ld3(si[i_di0 + 2 * s1 + 1]);
			V3 fa = v3(0, 0, 0), ta = v3(0, 0, 0);
			for (int s2 = 0; s2 < cj.n_ch; ++s2) {                     // mu_i - q_j (:1323-1481): charge is site 1
				const double4 B = sj[j_ch0 + s2];
				V3 f, M2; double u;
				pot_q_mu(v3(B.x - A.x, B.y - A.y, B.z - A.z), B.w, e1, A.w, f, M2, u);
				fa.x -= f.x; fa.y -= f.y; fa.z -= f.z; uX += u;
				ta.x += M2.x; ta.y += M2.y; ta.z += M2.z;
			}
			for (int s2 = 0; s2 < cj.n_di; ++s2) {                     // mu - mu (:1667-1845)
				const double4 B = sj[j_di0 + 2 * s2]; const V3 e2 = ld3(sj[j_di0 + 2 * s2 + 1]);
				V3 f, M1, M2; double u, r;
				pot_mu_mu(v3(A.x - B.x, A.y - B.y, A.z - B.z), e1, A.w, e2, B.w, a.epsRFInvrc3, f, M1, M2, u, r);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u; rf -= r;
				ta.x += M1.x; ta.y += M1.y; ta.z += M1.z;
			}
			for (int s2 = 0; s2 < cj.n_qu; ++s2) {                     // mu_i - Q_j (:2540-2720)
				const double4 B = sj[j_qu0 + 2 * s2]; const V3 e2 = ld3(sj[j_qu0 + 2 * s2 + 1]);
				V3 f, M1, M2; double u;
				pot_mu_Q(v3(A.x - B.x, A.y - B.y, A.z - B.z), e1, A.w, e2, B.w, f, M1, M2, u);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u;
				ta.x += M1.x; ta.y += M1.y; ta.z += M1.z;
			}
			F.x += fa.x; F.y += fa.y; F.z += fa.z;
			V.x += m.x * fa.x; V.y += m.y * fa.y; V.z += m.z * fa.z;
			V3 t = cross(v3(A.x - Ri.x, A.y - Ri.y, A.z - Ri.z), fa);
			M.x += t.x + ta.x; M.y += t.y + ta.y; M.z += t.z + ta.z;
		}
		// ---- quadrupoles of i
		for (int s1 = 0; s1 < ci.n_qu; ++s1) {
			const double4 A = si[i_qu0 + 2 * s1]; const V3 e1 = ld3(si[i_qu0 + 2 * s1 + 1]);
			V3 fa = v3(0, 0, 0), ta = v3(0, 0, 0);
			for (int s2 = 0; s2 < cj.n_ch; ++s2) {                     // Q_i - q_j (:1485-1650): charge is site 1
				const double4 B = sj[j_ch0 + s2];
				V3 f, M2; double u;
				pot_q_Q(v3(B.x - A.x, B.y - A.y, B.z - A.z), B.w, e1, A.w, f, M2, u);
				fa.x -= f.x; fa.y -= f.y; fa.z -= f.z; uX += u;
				ta.x += M2.x; ta.y += M2.y; ta.z += M2.z;
			}
			for (int s2 = 0; s2 < cj.n_di; ++s2) {                     // Q_i - mu_j (:2008-2192): dipole is site 1
				const double4 B = sj[j_di0 + 2 * s2]; const V3 e2 = ld3(sj[j_di0 + 2 * s2 + 1]);
				V3 f, M1, M2; double u;
				pot_mu_Q(v3(B.x - A.x, B.y - A.y, B.z - A.z), e2, B.w, e1, A.w, f, M1, M2, u);
				fa.x -= f.x; fa.y -= f.y; fa.z -= f.z; uX += u;
				ta.x += M2.x; ta.y += M2.y; ta.z += M2.z;
			}
			for (int s2 = 0; s2 < cj.n_qu; ++s2) {                     // Q - Q (:2202-2378)
				const double4 B = sj[j_qu0 + 2 * s2]; const V3 e2 = ld3(sj[j_qu0 + 2 * s2 + 1]);
				V3 f, M1, M2; double u;
				pot_Q_Q(v3(A.x - B.x, A.y - B.y, A.z - B.z), e1, A.w, e2, B.w, f, M1, M2, u);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u;
				ta.x += M1.x; ta.y += M1.y; ta.z += M1.z;
			}
			F.x += fa.x; F.y += fa.y; F.z += fa.z;
			V.x += m.x * fa.x; V.y += m.y * fa.y; V.z += m.z * fa.z;
			V3 t = cross(v3(A.x - Ri.x, A.y - Ri.y, A.z - Ri.z), fa);
			M.x += t.x + ta.x; M.y += t.y + ta.y; M.z += t.z + ta.z;
		}
	}
}

// The same ten loops for a SINGLE-COMPONENT system whose site counts are compile-time constants: every loop unrolls, the
// site rows of both molecules live in registers, the LJ tables come from the constant bank.
template <int NLJ, int NCH, int NDI, int NQU>
__device__ __forceinline__ void pair_body_fixed(const ForceArgs& a, const V3& Ri, const double4* si, const double4* sj,
                                                const V3& m, const bool in_lj, const bool in_rc, MolAcc& A_) {
	constexpr int i_ch0 = NLJ, i_di0 = NLJ + NCH, i_qu0 = NLJ + NCH + 2 * NDI;
	constexpr int j_ch0 = i_ch0, j_di0 = i_di0, j_qu0 = i_qu0;
	V3& F = A_.F; V3& M = A_.M; V3& V = A_.V; double& u6 = A_.u6; double& uX = A_.uX; double& rf = A_.rf;
	{
		// ---- LJ - LJ (VectorizedCellProcessor.cpp:1031-1180)
		if (in_lj)
			_Pragma("unroll") for (int s1 = 0; s1 < NLJ; ++s1) {
				const double4 A = si[s1];
				V3 fa = v3(0, 0, 0);
				_Pragma("unroll") for (int s2 = 0; s2 < NLJ; ++s2) {
					const double4 B = sj[s2];
					const int k = s1 * NLJ + s2;      // single component: global LJ id == site index
					V3 f; double u;
					pot_lj(v3(A.x - B.x, A.y - B.y, A.z - B.z), a.eps24_c[k], a.sig2_c[k], a.shift6_c[k], f, u);
					fa.x += f.x; fa.y += f.y; fa.z += f.z; u6 += u;
				}
				F.x += fa.x; F.y += fa.y; F.z += fa.z;
				V.x += m.x * fa.x; V.y += m.y * fa.y; V.z += m.z * fa.z;
				V3 t = cross(v3(A.x - Ri.x, A.y - Ri.y, A.z - Ri.z), fa);
				M.x += t.x; M.y += t.y; M.z += t.z;
			}
		if (!in_rc) return;
		// ---- charges of i
		_Pragma("unroll") for (int s1 = 0; s1 < NCH; ++s1) {
			const double4 A = si[i_ch0 + s1];
			V3 fa = v3(0, 0, 0);
			_Pragma("unroll") for (int s2 = 0; s2 < NCH; ++s2) {                     // q - q (:1192-1319)
				const double4 B = sj[j_ch0 + s2];
				V3 f; double u;
				pot_qq(v3(A.x - B.x, A.y - B.y, A.z - B.z), A.w * B.w, f, u);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u;
			}
			_Pragma("unroll") for (int s2 = 0; s2 < NDI; ++s2) {                     // q_i - mu_j (:1847-2006)
				const double4 B = sj[j_di0 + 2 * s2], E = sj[j_di0 + 2 * s2 + 1];
				V3 f, M2; double u;
				pot_q_mu(v3(A.x - B.x, A.y - B.y, A.z - B.z), A.w, ld3(E), B.w, f, M2, u);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u;
			}
			_Pragma("unroll") for (int s2 = 0; s2 < NQU; ++s2) {                     // q_i - Q_j (:2380-2538)
				const double4 B = sj[j_qu0 + 2 * s2], E = sj[j_qu0 + 2 * s2 + 1];
				V3 f, M2; double u;
				pot_q_Q(v3(A.x - B.x, A.y - B.y, A.z - B.z), A.w, ld3(E), B.w, f, M2, u);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u;
			}
			F.x += fa.x; F.y += fa.y; F.z += fa.z;
			V.x += m.x * fa.x; V.y += m.y * fa.y; V.z += m.z * fa.z;
			V3 t = cross(v3(A.x - Ri.x, A.y - Ri.y, A.z - Ri.z), fa);
			M.x += t.x; M.y += t.y; M.z += t.z;
		}
		// ---- dipoles of i
		_Pragma("unroll") for (int s1 = 0; s1 < NDI; ++s1) {
			const double4 A = si[i_di0 + 2 * s1]; const V3 e1 = ld3(si[i_di0 + 2 * s1 + 1]);
			V3 fa = v3(0, 0, 0), ta = v3(0, 0, 0);
			_Pragma("unroll") for (int s2 = 0; s2 < NCH; ++s2) {                     // mu_i - q_j (:1323-1481): charge is site 1
				const double4 B = sj[j_ch0 + s2];
				V3 f, M2; double u;
				pot_q_mu(v3(B.x - A.x, B.y - A.y, B.z - A.z), B.w, e1, A.w, f, M2, u);
				fa.x -= f.x; fa.y -= f.y; fa.z -= f.z; uX += u;
				ta.x += M2.x; ta.y += M2.y; ta.z += M2.z;
			}
			_Pragma("unroll") for (int s2 = 0; s2 < NDI; ++s2) {                     // mu - mu (:1667-1845)
				const double4 B = sj[j_di0 + 2 * s2]; const V3 e2 = ld3(sj[j_di0 + 2 * s2 + 1]);
				V3 f, M1, M2; double u, r;
				pot_mu_mu(v3(A.x - B.x, A.y - B.y, A.z - B.z), e1, A.w, e2, B.w, a.epsRFInvrc3, f, M1, M2, u, r);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u; rf -= r;
				ta.x += M1.x; ta.y += M1.y; ta.z += M1.z;
			}
			_Pragma("unroll") for (int s2 = 0; s2 < NQU; ++s2) {                     // mu_i - Q_j (:2540-2720)
				const double4 B = sj[j_qu0 + 2 * s2]; const V3 e2 = ld3(sj[j_qu0 + 2 * s2 + 1]);
				V3 f, M1, M2; double u;
				pot_mu_Q(v3(A.x - B.x, A.y - B.y, A.z - B.z), e1, A.w, e2, B.w, f, M1, M2, u);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u;
				ta.x += M1.x; ta.y += M1.y; ta.z += M1.z;
			}
			F.x += fa.x; F.y += fa.y; F.z += fa.z;
			V.x += m.x * fa.x; V.y += m.y * fa.y; V.z += m.z * fa.z;
			V3 t = cross(v3(A.x - Ri.x, A.y - Ri.y, A.z - Ri.z), fa);
			M.x += t.x + ta.x; M.y += t.y + ta.y; M.z += t.z + ta.z;
		}
		// ---- quadrupoles of i
		_Pragma("unroll") for (int s1 = 0; s1 < NQU; ++s1) {
			const double4 A = si[i_qu0 + 2 * s1]; const V3 e1 = ld3(si[i_qu0 + 2 * s1 + 1]);
			V3 fa = v3(0, 0, 0), ta = v3(0, 0, 0);
			_Pragma("unroll") for (int s2 = 0; s2 < NCH; ++s2) {                     // Q_i - q_j (:1485-1650): charge is site 1
				const double4 B = sj[j_ch0 + s2];
				V3 f, M2; double u;
				pot_q_Q(v3(B.x - A.x, B.y - A.y, B.z - A.z), B.w, e1, A.w, f, M2, u);
				fa.x -= f.x; fa.y -= f.y; fa.z -= f.z; uX += u;
				ta.x += M2.x; ta.y += M2.y; ta.z += M2.z;
			}
			_Pragma("unroll") for (int s2 = 0; s2 < NDI; ++s2) {                     // Q_i - mu_j (:2008-2192): dipole is site 1
				const double4 B = sj[j_di0 + 2 * s2]; const V3 e2 = ld3(sj[j_di0 + 2 * s2 + 1]);
				V3 f, M1, M2; double u;
				pot_mu_Q(v3(B.x - A.x, B.y - A.y, B.z - A.z), e2, B.w, e1, A.w, f, M1, M2, u);
				fa.x -= f.x; fa.y -= f.y; fa.z -= f.z; uX += u;
				ta.x += M2.x; ta.y += M2.y; ta.z += M2.z;
			}
			_Pragma("unroll") for (int s2 = 0; s2 < NQU; ++s2) {                     // Q - Q (:2202-2378)
				const double4 B = sj[j_qu0 + 2 * s2]; const V3 e2 = ld3(sj[j_qu0 + 2 * s2 + 1]);
				V3 f, M1, M2; double u;
				pot_Q_Q(v3(A.x - B.x, A.y - B.y, A.z - B.z), e1, A.w, e2, B.w, f, M1, M2, u);
				fa.x += f.x; fa.y += f.y; fa.z += f.z; uX += u;
				ta.x += M1.x; ta.y += M1.y; ta.z += M1.z;
			}
			F.x += fa.x; F.y += fa.y; F.z += fa.z;
			V.x += m.x * fa.x; V.y += m.y * fa.y; V.z += m.z * fa.z;
			V3 t = cross(v3(A.x - Ri.x, A.y - Ri.y, A.z - Ri.z), fa);
			M.x += t.x + ta.x; M.y += t.y + ta.y; M.z += t.z + ta.z;
		}
	}
}

