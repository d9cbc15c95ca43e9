// force.cu — the short-range force pass (multi-site path) and the shared reduction epilogue.
//
// Replaces (reference, src/):
//   LinkedCells::updateMoleculeCaches -> FullParticleCell::buildSoACaches -> FullMolecule::setupSoACache
//        particleContainer/LinkedCells.cpp:1076-1086, FullParticleCell.cpp:143-193, molecules/FullMolecule.cpp:714-770
//   LinkedCells::traverseCells -> C08 traversal -> VectorizedCellProcessor::processCell/processCellPair/_calculatePairs
//        particleContainer/LinkedCells.cpp:564-575, adapter/VectorizedCellProcessor.cpp:796-2821
//   VectorizedCellProcessor::endTraversal                  adapter/VectorizedCellProcessor.cpp:124-157
//   Simulation::updateForces -> FullMolecule::calcFM        Simulation.cpp:752-762, molecules/FullMolecule.cpp:551-631
//
// GPU formulation: owner-computes full shell.  The reference visits every unordered cell pair once (C08, 8
// colours, Newton 3) because its threads share the per-cell SoA; on the GPU every owned molecule accumulates
// the forces on ITS OWN sites from all molecules in the 27 surrounding cells, so there is no colouring, no
// atomics and the result does not depend on scheduling.  Macroscopic sums: every real pair is visited from
// both sides (weight 1/2 each); a pair across the periodic boundary is visited once from the owned side of
// each of its two images (again 1/2 + 1/2), which equals the reference's "count the image pair only where
// index(nonHalo) < index(halo)" rule (VectorizedCellProcessor.cpp:2797-2818).
// calcFM is fused: site forces are folded into F, M = sum d_site x f_site + site torques, Vi = 1/2 sum V.
#include "ls1g_internal.cuh"
#include "potentials.cuh"

int ls1g_ensure_partials(ls1g_ctx* c, size_t blocks, int nv);
int k_forces_1clj(ls1g_ctx* c);   // force_1clj.cu

#define FORCE_THREADS 128

// ------------------------------------------------------------------------------------------------
// site cache: world-frame site data per molecule slot (the device analogue of CellDataSoA).
// row layout per molecule (stride = cache_stride double4):
//   n_lj x (x y z | ljid bits)   n_ch x (x y z | q)   n_di x [(x y z | mu), (ex ey ez | 0)]   n_qu x [(x y z | Q), (ex ey ez | 0)]
// q is normalised in place (FullMolecule::setupSoACache -> normalizeQuaternion, FullMolecule.cpp:723).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) site_cache_kernel(size_t n, const double4* __restrict__ pos, double4* __restrict__ q,
                                                          const CompDev* __restrict__ comps, const SiteDev* __restrict__ sites,
                                                          double4* __restrict__ cache, int stride) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	double4 p = pos[i];
	double4 qq = q[i];
	double s = 1. / sqrt(qq.x * qq.x + qq.y * qq.y + qq.z * qq.z + qq.w * qq.w);   // Quaternion::normalize (Quaternion.h:39-45)
	qq.x *= s; qq.y *= s; qq.z *= s; qq.w *= s;
	q[i] = qq;
	const CompDev cp = comps[pos_cid(p)];
	double4* row = cache + i * (size_t)stride;
	int k = 0;
	for (int a = 0; a < cp.n_sites; ++a) {
		const SiteDev st = sites[cp.site0 + a];
		double d[3];
		quat_rotate(qq, st.d, d);
		double w = st.type == SITE_LJ ? __longlong_as_double((long long)st.ljid) : st.val;
		row[k++] = make_double4(d[0] + p.x, d[1] + p.y, d[2] + p.z, w);
		if (st.type >= SITE_DI) {
			double e[3];
			quat_rotate(qq, st.e, e);
			row[k++] = make_double4(e[0], e[1], e[2], 0.);
		}
	}
}

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
};

__device__ __forceinline__ V3 ld3(const double4& a) { return v3(a.x, a.y, a.z); }

// generic multi-site, multi-component kernel: one thread per owned molecule.
__global__ void __launch_bounds__(FORCE_THREADS) force_generic_kernel(ForceArgs a) {
	__shared__ double red[5 * 32];
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	double acc[5] = {0., 0., 0., 0., 0.};
	uint32_t cflag = i < a.n ? a.cell[i] : HALO_FLAG;
	if (!(cflag & HALO_FLAG)) {
		const Grid& g = a.g;
		const double4 Pi = a.pos[i];
		const V3 Ri = ld3(Pi);
		const CompDev ci = a.comps[pos_cid(Pi)];
		const double4* si = a.cache + i * (size_t)a.stride;
		const int i_ch0 = ci.n_lj, i_di0 = ci.n_lj + ci.n_ch, i_qu0 = i_di0 + 2 * ci.n_di;
		V3 F = v3(0, 0, 0), M = v3(0, 0, 0), V = v3(0, 0, 0);
		double u6 = 0., uX = 0., rf = 0.; unsigned long long np = 0;
		const int cx = cflag % g.cpd[0], cy = (cflag / g.cpd[0]) % g.cpd[1], cz = cflag / (g.cpd[0] * g.cpd[1]);
		for (int dz = -1; dz <= 1; ++dz)
		for (int dy = -1; dy <= 1; ++dy) {
			// the three x-neighbours of a row are contiguous in the cell-sorted arrays
			uint32_t crow = (uint32_t)(((cz + dz) * g.cpd[1] + (cy + dy)) * g.cpd[0] + cx);
			uint32_t j0 = a.cell_start[crow - 1], j1 = a.cell_start[crow + 2];
			uint32_t own0 = a.cell_start[cflag], own1 = a.cell_start[cflag + 1];
			bool own_row = (dz == 0 && dy == 0);
			for (uint32_t j = j0; j < j1; ++j) {
				if (j == (uint32_t)i) continue;
				const double4 Pj = a.pos[j];
				const V3 m = v3(Ri.x - Pj.x, Ri.y - Pj.y, Ri.z - Pj.z);
				const double m2 = dot(m, m);
				// SingleCellPolicy_::GetForceMask: pairs at distance 0 inside one cell are skipped (Helpers.h:367-376)
				if (own_row && j >= own0 && j < own1 && m2 == 0.0) continue;
				const bool in_lj = m2 < a.rclj2, in_rc = m2 < a.rc2;
				if (!(in_lj || in_rc)) continue;
				++np;
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
							pot_lj(v3(A.x - B.x, A.y - B.y, A.z - B.z), a.eps24[k], a.sig2[k], a.shift6[k], f, u);
							fa.x += f.x; fa.y += f.y; fa.z += f.z; u6 += u;
						}
						F.x += fa.x; F.y += fa.y; F.z += fa.z;
						V.x += m.x * fa.x; V.y += m.y * fa.y; V.z += m.z * fa.z;
						V3 t = cross(v3(A.x - Ri.x, A.y - Ri.y, A.z - Ri.z), fa);
						M.x += t.x; M.y += t.y; M.z += t.z;
					}
				if (!in_rc) continue;
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
		// calcFM (FullMolecule.cpp:551-631): Vi = 1/2 sum V_site
		a.F[3 * i] = F.x; a.F[3 * i + 1] = F.y; a.F[3 * i + 2] = F.z;
		a.M[3 * i] = M.x; a.M[3 * i + 1] = M.y; a.M[3 * i + 2] = M.z;
		a.Vi[3 * i] = 0.5 * V.x; a.Vi[3 * i + 1] = 0.5 * V.y; a.Vi[3 * i + 2] = 0.5 * V.z;
		acc[0] = u6; acc[1] = uX; acc[2] = V.x + V.y + V.z; acc[3] = rf; acc[4] = (double)np;
	}
	block_sum<5>(acc, red);
	if (threadIdx.x == 0) {
#pragma unroll
		for (int k = 0; k < 5; ++k) a.partials[(size_t)blockIdx.x * 5 + k] = acc[k];
	}
}

// endTraversal (VectorizedCellProcessor.cpp:124-157) on the per-block partials; every pair was visited twice.
// out[0] = upot, out[1] = virial, out[2] = molecule pairs inside the cutoff (counted once)
__global__ void __launch_bounds__(256) force_finalize_kernel(const double* __restrict__ partials, size_t nb, double* __restrict__ out) {
	__shared__ double red[5 * 32];
	double s[5] = {0., 0., 0., 0., 0.};
	for (size_t b = threadIdx.x; b < nb; b += blockDim.x)
#pragma unroll
		for (int k = 0; k < 5; ++k) s[k] += partials[b * 5 + k];
	block_sum<5>(s, red);
	if (threadIdx.x == 0) {
		double upot6lj = 0.5 * s[0], upotXpoles = 0.5 * s[1], virial = 0.5 * s[2], myRF = 0.5 * s[3];
		out[0] = upot6lj / 6.0 + upotXpoles + myRF;
		out[1] = virial + 3.0 * myRF;
		out[2] = 0.5 * s[4];
		out[3] = 0.;
	}
}

static inline unsigned nblk(size_t n, unsigned t) { return (unsigned)((n + t - 1) / t); }

// first level of the deterministic two-level reduction: gridDim.x slices of the per-block partials -> out[gridDim.x][NV]
template <int NV>
__global__ void __launch_bounds__(256) reduce_partials_kernel(const double* __restrict__ partials, size_t nb, double* __restrict__ out) {
	__shared__ double red[NV * 32];
	const size_t per = (nb + gridDim.x - 1) / gridDim.x;
	const size_t b0 = (size_t)blockIdx.x * per, b1 = b0 + per < nb ? b0 + per : nb;
	double s[NV];
#pragma unroll
	for (int k = 0; k < NV; ++k) s[k] = 0.;
	for (size_t b = b0 + threadIdx.x; b < b1; b += blockDim.x)
#pragma unroll
		for (int k = 0; k < NV; ++k) s[k] += partials[b * NV + k];
	block_sum<NV>(s, red);
	if (threadIdx.x == 0)
#pragma unroll
		for (int k = 0; k < NV; ++k) out[(size_t)blockIdx.x * NV + k] = s[k];
}

// reduces partials[nb][nv] (nv = 4 or 5) to at most 256 rows when nb is large; returns the rows to finalize and their base
int ls1g_prereduce(ls1g_ctx* c, size_t& nb, int nv, const double*& base) {
	base = c->partials;
	if (nb <= 2048) return 0;
	double* out = c->partials + nb * nv;          // ensure_partials reserves twice the requested size
	if (nv == 5) reduce_partials_kernel<5><<<256, 256, 0, c->stream>>>(c->partials, nb, out);
	else reduce_partials_kernel<4><<<256, 256, 0, c->stream>>>(c->partials, nb, out);
	CUDA_TRY(cudaGetLastError());
	nb = 256; base = out;
	return 0;
}

int k_force_finalize(ls1g_ctx* c, size_t nb) {
	timer_start(c, LS1G_T_AUX);
	const double* base; int extra = nb > 2048;
	if (ls1g_prereduce(c, nb, 5, base)) return 1;
	force_finalize_kernel<<<1, 256, 0, c->stream>>>(base, nb, c->d_scalars);
	timer_stop(c, LS1G_T_AUX, 1 + extra);
	CUDA_TRY(cudaGetLastError());
	return 0;
}

int k_forces(ls1g_ctx* c) {
	if (!c->sorted) LS1G_FAIL("ls1g_traverse: container not sorted (call ls1g_update / ls1g_exchange first)");
	if (c->simple_1clj) return k_forces_1clj(c);
	size_t n = c->n_slots;
	unsigned nb = nblk(n, FORCE_THREADS);
	if (ls1g_ensure_partials(c, nb ? nb : 1, 5)) return 1;
	if (n) {
		timer_start(c, LS1G_T_AUX);
		site_cache_kernel<<<nblk(n, 256), 256, 0, c->stream>>>(n, c->cur.pos, c->cur.q, c->comps, c->sites, c->site_cache, c->cache_stride);
		timer_stop(c, LS1G_T_AUX, 1);
		ForceArgs a;
		a.g = c->bgrid; a.n = n; a.pos = c->cur.pos; a.cell = c->cur.cell; a.cell_start = c->cell_start;
		a.comps = c->comps; a.cache = c->site_cache; a.stride = c->cache_stride;
		a.eps24 = c->eps24; a.sig2 = c->sig2; a.shift6 = c->shift6; a.S = c->S_lj;
		a.rc2 = c->rc2; a.rclj2 = c->rclj2; a.epsRFInvrc3 = c->epsRFInvrc3;
		a.F = c->F; a.M = c->M; a.Vi = c->Vi; a.partials = c->partials;
		timer_start(c, LS1G_T_FORCE);
		force_generic_kernel<<<nb, FORCE_THREADS, 0, c->stream>>>(a);
		timer_stop(c, LS1G_T_FORCE, 1);
	}
	if (k_force_finalize(c, nb)) return 1;
	CUDA_TRY(cudaGetLastError());
	return 0;
}
