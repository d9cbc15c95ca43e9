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
#include "pair_body.cuh"

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

#define GQ_CAP 88        // queue entries per lane (molecule pairs inside the cutoff waiting for their site loops)
#define GQ_CHUNK 16      // candidates scanned between two "queue nearly full?" votes

// generic multi-site, multi-component kernel: one thread per owned molecule.
// The molecule-centre cutoff accepts only ~10 % of the candidates of the 27 cells, and the site loops behind it are long:
// evaluating them under that branch leaves ~3 of 32 lanes busy.  So every lane first SCANS candidates and queues the
// survivors (shared memory, one column per lane); when some lane's queue is nearly full - or the candidates are
// exhausted - all lanes DRAIN their queues through the site loops together.
// NLJ >= 0: single-component system with compile-time site counts (pair_body_fixed); NLJ < 0: anything (pair_body).
template <int NLJ, int NCH, int NDI, int NQU>
__global__ void __launch_bounds__(FORCE_THREADS) force_generic_kernel(ForceArgs a) {
	constexpr bool FIXED = NLJ >= 0;
	constexpr int NS = FIXED ? NLJ + NCH + 2 * NDI + 2 * NQU : 1;      // site-cache rows per molecule
	__shared__ double red[5 * 32];
	__shared__ uint32_t queue[GQ_CAP * FORCE_THREADS];
	__shared__ double ljtab[FIXED ? 1 : 3 * LJ_TAB_MAX];      // mixtures: the LJ tables in shared memory
	LJTab lj; lj.eps24 = lj.sig2 = lj.shift6 = nullptr;
	if constexpr (!FIXED) lj = lj_tables(a, ljtab);
	// block (row, b) = 128 consecutive OWNED slots of one cell row (y, z): the owned cells of a row are contiguous in the
	// cell-sorted arrays, so no lane is spent on a halo slot
	const Grid& g = a.g;
	double acc[5] = {0., 0., 0., 0., 0.};
	size_t i; bool owned;
	{
		const int row = blockIdx.x / a.bpr, bi = blockIdx.x % a.bpr;
		const uint32_t crow0 = (uint32_t)(((row / a.nyi + 1) * g.cpd[1] + (row % a.nyi + 1)) * g.cpd[0]);
		const uint32_t rs = a.cell_start[crow0 + 1], re = a.cell_start[crow0 + g.cpd[0] - 1];
		i = (size_t)rs + (size_t)bi * FORCE_THREADS + threadIdx.x;
		owned = i < re;
	}
	const uint32_t cflag = owned ? a.cell[i] : HALO_FLAG;
	double4 Pi = make_double4(0., 0., 0., 0.);
	if (owned) Pi = a.pos[i];
	const V3 Ri = ld3(Pi);
	CompDev ci; ci.n_lj = ci.n_ch = ci.n_di = ci.n_qu = 0;
	if (owned) ci = a.comps[pos_cid(Pi)];
	const double4* si = a.cache + i * (size_t)a.stride;
	const int i_ch0 = ci.n_lj, i_di0 = ci.n_lj + ci.n_ch, i_qu0 = i_di0 + 2 * ci.n_di;
	double4 sir[NS];                       // FIXED: the molecule's own site rows in registers
	if (FIXED && owned) {
#pragma unroll
		for (int k = 0; k < NS; ++k) sir[k] = si[k];
	}
	MolAcc A_; A_.F = v3(0, 0, 0); A_.M = v3(0, 0, 0); A_.V = v3(0, 0, 0); A_.u6 = 0.; A_.uX = 0.; A_.rf = 0.; A_.np = 0;
	const int cx = cflag % g.cpd[0], cy = (cflag / g.cpd[0]) % g.cpd[1], cz = cflag / (g.cpd[0] * g.cpd[1]);
	uint32_t own0 = 0, own1 = 0;
	if (owned) { own0 = a.cell_start[cflag]; own1 = a.cell_start[cflag + 1]; }
	uint32_t* q = queue + threadIdx.x;
	int cnt = 0;
	int row = owned ? 0 : 9;              // rows of the 27-cell neighbourhood: the three x-neighbours of a row are contiguous slots
	uint32_t j = 0, jend = 0;
	auto open_row = [&]() {
		const int dz = row / 3 - 1, dy = row % 3 - 1;
		const uint32_t crow = (uint32_t)(((cz + dz) * g.cpd[1] + (cy + dy)) * g.cpd[0] + cx);
		j = a.cell_start[crow - 1]; jend = a.cell_start[crow + 2];
	};
	if (owned) open_row();
	for (;;) {
		// ---- scan: molecule-centre distances only ----------------------------------------------------------------
		for (;;) {
			if (row < 9) {
				const uint32_t je = min(jend, j + GQ_CHUNK);
				const bool own_row = row == 4;
				for (; j < je; ++j) {
					const double4 Pj = a.pos[j];
					const double mx = Ri.x - Pj.x, my = Ri.y - Pj.y, mz = Ri.z - Pj.z;
					const double m2 = mx * mx + my * my + mz * mz;
					bool take = (m2 < a.rclj2 || m2 < a.rc2) && j != (uint32_t)i;
					// SingleCellPolicy_::GetForceMask: pairs at distance 0 inside one cell are skipped (Helpers.h:367-376)
					if (own_row && j >= own0 && j < own1 && m2 == 0.0) take = false;
					if (take) { q[cnt * FORCE_THREADS] = j; ++cnt; }
				}
				if (j == jend) { ++row; if (row < 9) open_row(); }
			}
			const bool any_full = __any_sync(0xffffffffu, cnt > GQ_CAP - GQ_CHUNK);
			const bool any_more = __any_sync(0xffffffffu, row < 9);
			if (any_full || !any_more) break;
		}
		// ---- drain: the site loops of the queued pairs -----------------------------------------------------------
		// Not everything: every lane takes the D newest entries of its queue, D = the smallest backlog in the warp (all
		// lanes busy), but at least what the fullest lane must shed to scan on.  Only the last round empties the queues.
		const bool last_round = !__any_sync(0xffffffffu, row < 9);
		int D = GQ_CAP;
		if (!last_round) {
			const int mn = __reduce_min_sync(0xffffffffu, owned && (row < 9 || cnt > 0) ? cnt : GQ_CAP);
			const int mx = __reduce_max_sync(0xffffffffu, cnt);
			D = max(mn, mx - (GQ_CAP - GQ_CHUNK));
		}
		const int take = min(cnt, D);
		for (int t = 0; t < take; ++t) {
			const uint32_t jj = q[(cnt - 1 - t) * FORCE_THREADS];
			const double4 Pj = a.pos[jj];
			const V3 m = v3(Ri.x - Pj.x, Ri.y - Pj.y, Ri.z - Pj.z);
			const double m2 = dot(m, m);
			++A_.np;
			if constexpr (FIXED) {
				double4 sjr[NS];
				const double4* sj = a.cache + (size_t)jj * a.stride;
#pragma unroll
				for (int k = 0; k < NS; ++k) sjr[k] = sj[k];
				pair_body_fixed<NLJ, NCH, NDI, NQU>(a, Ri, sir, sjr, m, m2 < a.rclj2, m2 < a.rc2, A_);
			} else {
				pair_body(a, lj, Ri, ci, si, i_ch0, i_di0, i_qu0, jj, Pj, m, m2 < a.rclj2, m2 < a.rc2, A_);
			}
		}
		cnt -= take;
		if (last_round) break;
	}
	if (owned) {
		// calcFM (FullMolecule.cpp:551-631): Vi = 1/2 sum V_site
		a.F[3 * i] = A_.F.x; a.F[3 * i + 1] = A_.F.y; a.F[3 * i + 2] = A_.F.z;
		a.M[3 * i] = A_.M.x; a.M[3 * i + 1] = A_.M.y; a.M[3 * i + 2] = A_.M.z;
		a.Vi[3 * i] = 0.5 * A_.V.x; a.Vi[3 * i + 1] = 0.5 * A_.V.y; a.Vi[3 * i + 2] = 0.5 * A_.V.z;
		acc[0] = A_.u6; acc[1] = A_.uX; acc[2] = A_.V.x + A_.V.y + A_.V.z; acc[3] = A_.rf; acc[4] = (double)A_.np;
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

// reduces partials[nb][nv] (nv = 4, 5 or 8) to at most 256 rows when nb is large; returns the rows to finalize and their base
int ls1g_prereduce(ls1g_ctx* c, size_t& nb, int nv, const double*& base) {
	base = c->partials;
	if (nb <= 2048) return 0;
	double* out = c->partials + nb * nv;          // ensure_partials reserves twice the requested size
	if (nv == 5) reduce_partials_kernel<5><<<256, 256, 0, c->stream>>>(c->partials, nb, out);
	else if (nv == 8) reduce_partials_kernel<8><<<256, 256, 0, c->stream>>>(c->partials, nb, out);
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

// world-frame site rows of every slot (owned and halo); normalises q in place like setupSoACache
int k_site_cache(ls1g_ctx* c) {
	if (!c->n_slots) return 0;
	timer_start(c, LS1G_T_AUX);
	site_cache_kernel<<<nblk(c->n_slots, 256), 256, 0, c->stream>>>(c->n_slots, c->cur.pos, c->cur.q, c->comps, c->sites, c->site_cache, c->cache_stride);
	timer_stop(c, LS1G_T_AUX, 1);
	CUDA_TRY(cudaGetLastError());
	return 0;
}

int k_forces(ls1g_ctx* c) {
	if (!c->sorted) LS1G_FAIL("ls1g_traverse: container not sorted (call ls1g_update / ls1g_exchange first)");
	if (c->simple_1clj) return k_forces_1clj(c);
	size_t n = c->n_slots;
	const int bpr = (int)((c->max_row + FORCE_THREADS - 1) / FORCE_THREADS);
	const int nyi = c->bgrid.cpd[1] - 2, nzi = c->bgrid.cpd[2] - 2;
	unsigned nb = n ? (unsigned)((size_t)nyi * nzi * (bpr > 0 ? bpr : 1)) : 0;
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
		a.bpr = bpr > 0 ? bpr : 1; a.nyi = nyi;
		// single-component systems of the common rigid models get a kernel with compile-time site counts
		const CompDev& c0 = c->comps_h[0];
		const int key = c->n_comp == 1 && c->S_lj <= 4 ? c0.n_lj * 1000 + c0.n_ch * 100 + c0.n_di * 10 + c0.n_qu : -1;
		if (key >= 0)
			for (int k = 0; k < c->S_lj * c->S_lj; ++k) { a.eps24_c[k] = c->eps24_h[k]; a.sig2_c[k] = c->sig2_h[k]; a.shift6_c[k] = c->shift6_h[k]; }
		timer_start(c, LS1G_T_FORCE);
		switch (key) {
#define LS1G_FIXED_CASE(L, Q, D, U) case L * 1000 + Q * 100 + D * 10 + U: force_generic_kernel<L, Q, D, U><<<nb, FORCE_THREADS, 0, c->stream>>>(a); break;
		LS1G_FIXED_CASE(3, 3, 0, 0)      // CO2 (3 LJ + 3 charges), examples/CO2
		LS1G_FIXED_CASE(3, 0, 1, 0)      // ethylene oxide (3 LJ + dipole), examples/EOX
		LS1G_FIXED_CASE(3, 0, 0, 1)      // CO2 Merker (3 LJ + quadrupole)
		LS1G_FIXED_CASE(2, 0, 0, 1)      // 2CLJQ
		LS1G_FIXED_CASE(2, 0, 1, 0)      // 2CLJD
		LS1G_FIXED_CASE(1, 0, 1, 0)      // Stockmayer
		LS1G_FIXED_CASE(2, 0, 0, 0)      // 2CLJ
		LS1G_FIXED_CASE(4, 0, 0, 0)      // 4 LJ sites
#undef LS1G_FIXED_CASE
		default: force_generic_kernel<-1, 0, 0, 0><<<nb, FORCE_THREADS, 0, c->stream>>>(a);
		}
		timer_stop(c, LS1G_T_FORCE, 1);
	}
	if (k_force_finalize(c, nb)) return 1;
	CUDA_TRY(cudaGetLastError());
	return 0;
}
