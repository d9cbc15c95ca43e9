// rdf.cu — the reference's RDFCellProcessor + the counters of its RDF plugin as a device cell processor.
//
// Replaces (reference, src/): particleContainer/adapter/RDFCellProcessor.cpp:18-133 (processCell / processCellPair),
// io/RDF.h:95-160 (observeRDF, observeRDFMolecule, observeRDFSite, observeARDFMolecule) and
// molecules/MoleculeInterface.h:157-190 (dist2, orientationAngle), i.e. `container->traverseCells(rdfCellProcessor)` as
// RDF::afterForces (io/RDF.cpp:573-581) does it.
//
// The reference walks unordered cell pairs (C08, non-halo cell first) and observes every molecule pair whose centre distance^2 is
// below rc^2 - except that a (boundary, halo) cell pair is only visited when index(boundary) < index(halo), so that a pair across
// the periodic boundary is seen from one of its two images only.  Here every OWNED molecule i scans its 27 cells and observes the
// partner j when   cell(j) is a halo cell ? cell(i) < cell(j)  :  (cell(i), i) < (cell(j), j),   which is the same set of pairs.
// All observers are symmetric in (i, j), so the histograms equal the reference's bin for bin.
//
// Arithmetic is written with explicit round-to-nearest intrinsics in the order of the reference's expressions (no FMA
// contraction): a distance that sits on a bin border must fall on the same side.
#include "ls1g_internal.cuh"

#define RDF_THREADS 128

struct RdfArgs {
	Grid g;
	size_t n;
	const double4* pos; const double4* q; const uint32_t* cell; const uint32_t* cell_start;
	const CompDev* comps;
	const double* sitepos;          // [n][max_sites][3] absolute site positions (rdf_sites_kernel), or null
	int max_sites;
	double rc2, max2, bw, aw;
	uint32_t bins, abins;
	int nc, S, do_site, do_ardf;
	size_t off_site, off_ardf, total;   // histogram layout: mol | site | ardf
	unsigned long long* hist;
};

__device__ __forceinline__ double rdf_dot3(double a0, double a1, double a2, double b0, double b1, double b2) {
	return __dadd_rn(__dadd_rn(__dmul_rn(a0, b0), __dmul_rn(a1, b1)), __dmul_rn(a2, b2));
}

// Quaternion::rotate (Quaternion.cpp:43-62) term by term, q = (w, x, y, z) in (q.x, q.y, q.z, q.w)
__device__ __forceinline__ void rdf_rotate(const double4& q, const double d[3], double out[3]) {
	const double ww = __dmul_rn(q.x, q.x), xx = __dmul_rn(q.y, q.y), yy = __dmul_rn(q.z, q.z), zz = __dmul_rn(q.w, q.w);
	const double wx = __dmul_rn(q.x, q.y), wy = __dmul_rn(q.x, q.z), wz = __dmul_rn(q.x, q.w);
	const double xy = __dmul_rn(q.y, q.z), xz = __dmul_rn(q.y, q.w), yz = __dmul_rn(q.z, q.w);
	const double a00 = __dsub_rn(__dsub_rn(__dadd_rn(ww, xx), yy), zz), a01 = __dmul_rn(2., __dsub_rn(xy, wz)), a02 = __dmul_rn(2., __dadd_rn(wy, xz));
	const double a10 = __dmul_rn(2., __dadd_rn(wz, xy)), a11 = __dsub_rn(__dadd_rn(__dsub_rn(ww, xx), yy), zz), a12 = __dmul_rn(2., __dsub_rn(yz, wx));
	const double a20 = __dmul_rn(2., __dsub_rn(xz, wy)), a21 = __dmul_rn(2., __dadd_rn(wx, yz)), a22 = __dadd_rn(__dsub_rn(__dsub_rn(ww, xx), yy), zz);
	out[0] = rdf_dot3(a00, a01, a02, d[0], d[1], d[2]);
	out[1] = rdf_dot3(a10, a11, a12, d[0], d[1], d[2]);
	out[2] = rdf_dot3(a20, a21, a22, d[0], d[1], d[2]);
}

// absolute site positions r + rotate(normalised q, d_body) of every slot (FullMolecule::setupSoACache, FullMolecule.cpp:714-770)
__global__ void __launch_bounds__(256) rdf_sites_kernel(size_t n, const double4* __restrict__ pos, const double4* __restrict__ q,
                                                         const CompDev* __restrict__ comps, const SiteDev* __restrict__ sites,
                                                         double* __restrict__ sitepos, int max_sites) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const double4 p = pos[i];
	double4 qq = q ? q[i] : make_double4(1., 0., 0., 0.);
	const double s = __dsqrt_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(qq.x, qq.x), __dmul_rn(qq.y, qq.y)), __dmul_rn(qq.z, qq.z)), __dmul_rn(qq.w, qq.w)));
	qq.x = __ddiv_rn(qq.x, s); qq.y = __ddiv_rn(qq.y, s); qq.z = __ddiv_rn(qq.z, s); qq.w = __ddiv_rn(qq.w, s);
	const CompDev cp = comps[pos_cid(p)];
	double* row = sitepos + i * (size_t)max_sites * 3;
	for (int a = 0; a < cp.n_sites; ++a) {
		const SiteDev st = sites[cp.site0 + a];
		double d[3];
		rdf_rotate(qq, st.d, d);
		row[3 * a + 0] = __dadd_rn(p.x, d[0]);
		row[3 * a + 1] = __dadd_rn(p.y, d[1]);
		row[3 * a + 2] = __dadd_rn(p.z, d[2]);
	}
}

template <bool SHARED>
__device__ __forceinline__ void rdf_hit(const RdfArgs& a, uint32_t* sh, size_t idx) {
	if (SHARED) atomicAdd(&sh[idx], 1u);
	else atomicAdd(&a.hist[idx], 1ull);
}

// MoleculeInterface::orientationAngle (MoleculeInterface.h:167-190) of the molecule with quaternion q towards dr, |dr|^2 = dd
__device__ __forceinline__ double rdf_cos_phi(const double4& q, double dr0, double dr1, double dr2, double dd) {
	const double o0 = __dmul_rn(2., __dadd_rn(__dmul_rn(q.y, q.w), __dmul_rn(q.x, q.z)));
	const double o1 = __dmul_rn(2., __dsub_rn(__dmul_rn(q.z, q.w), __dmul_rn(q.x, q.y)));
	const double o2 = __dsub_rn(1., __dmul_rn(2., __dadd_rn(__dmul_rn(q.y, q.y), __dmul_rn(q.z, q.z))));
	const double so = __dsqrt_rn(rdf_dot3(o0, o1, o2, o0, o1, o2)), sd = __dsqrt_rn(dd);
	double c = 0.;
	c = __dadd_rn(c, __ddiv_rn(__ddiv_rn(__dmul_rn(o0, dr0), so), sd));
	c = __dadd_rn(c, __ddiv_rn(__ddiv_rn(__dmul_rn(o1, dr1), so), sd));
	c = __dadd_rn(c, __ddiv_rn(__ddiv_rn(__dmul_rn(o2, dr2), so), sd));
	return c;
}

template <bool SHARED>
__global__ void __launch_bounds__(RDF_THREADS) rdf_kernel(RdfArgs a) {
	extern __shared__ uint32_t sh[];
	if (SHARED) {
		for (size_t k = threadIdx.x; k < a.total; k += blockDim.x) sh[k] = 0u;
		__syncthreads();
	}
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < a.n && !(a.cell[i] & HALO_FLAG)) {
		const double4 pi = a.pos[i];
		const int ci = (int)pos_cid(pi);
		const CompDev cpi = a.comps[ci];
		const double4 qi = (a.do_ardf && a.q) ? a.q[i] : make_double4(1., 0., 0., 0.);
		const uint32_t c = a.cell[i];
		const int nx = a.g.cpd[0], ny = a.g.cpd[1];
		const int cx = c % nx, cy = (c / nx) % ny, cz = c / (nx * ny);
		for (int dz = -1; dz <= 1; ++dz)
			for (int dy = -1; dy <= 1; ++dy)
				for (int dx = -1; dx <= 1; ++dx) {
					const uint32_t cj = (uint32_t)(((cz + dz) * ny + (cy + dy)) * nx + (cx + dx));
					const bool halo_j = grid_is_halo(a.g, cj);
					if (halo_j ? !(c < cj) : (cj < c)) continue;
					for (uint32_t j = a.cell_start[cj]; j < a.cell_start[cj + 1]; ++j) {
						if (cj == c && j <= i) continue;
						const double4 pj = a.pos[j];
						// molecule2.dist2(molecule1): dr = r1 - r2 with molecule1 = i
						const double e0 = __dsub_rn(pi.x, pj.x), e1 = __dsub_rn(pi.y, pj.y), e2 = __dsub_rn(pi.z, pj.z);
						const double dd = rdf_dot3(e0, e1, e2, e0, e1, e2);
						if (!(dd < a.rc2)) continue;
						const int cjd = (int)pos_cid(pj);
						if (!(dd > a.max2)) {
							const unsigned long long bin = (unsigned long long)floor(__ddiv_rn(__dsqrt_rn(dd), a.bw));
							if (bin < a.bins) {
								const int lo = min(ci, cjd), hi = max(ci, cjd);
								const size_t row = (size_t)lo * a.nc - (size_t)lo * (lo - 1) / 2 + (hi - lo);
								rdf_hit<SHARED>(a, sh, row * a.bins + bin);
								if (a.do_ardf) {
									const double4 qj = a.q ? a.q[j] : make_double4(1., 0., 0., 0.);
									const double c1 = rdf_cos_phi(qi, -e0, -e1, -e2, dd);       // i towards j: r_j - r_i
									const double c2 = rdf_cos_phi(qj, e0, e1, e2, dd);
									const unsigned long long a1 = (unsigned long long)floor(__ddiv_rn(__dadd_rn(-c1, 1.), a.aw));
									const unsigned long long a2 = (unsigned long long)floor(__ddiv_rn(__dadd_rn(-c2, 1.), a.aw));
									const size_t per = (size_t)a.bins * a.abins;
									const size_t f1 = bin * a.abins + a1, f2 = bin * a.abins + a2;
									if (f1 < per) rdf_hit<SHARED>(a, sh, a.off_ardf + ((size_t)ci * a.nc + cjd) * per + f1);
									if (f2 < per) rdf_hit<SHARED>(a, sh, a.off_ardf + ((size_t)cjd * a.nc + ci) * per + f2);
								}
							}
						}
						if (a.do_site) {
							const CompDev cpj = a.comps[cjd];
							if (cpi.n_sites + cpj.n_sites > 2) {
								const double* si = a.sitepos + i * (size_t)a.max_sites * 3;
								const double* sj = a.sitepos + (size_t)j * a.max_sites * 3;
								for (int m = 0; m < cpi.n_sites; ++m)
									for (int n = 0; n < cpj.n_sites; ++n) {
										const double s0 = __dsub_rn(si[3 * m], sj[3 * n]), s1 = __dsub_rn(si[3 * m + 1], sj[3 * n + 1]), s2 = __dsub_rn(si[3 * m + 2], sj[3 * n + 2]);
										const double d2 = rdf_dot3(s0, s1, s2, s0, s1, s2);
										if (d2 > a.max2) continue;
										const unsigned long long bin = (unsigned long long)floor(__ddiv_rn(__dsqrt_rn(d2), a.bw));
										if (bin >= a.bins) continue;
										const size_t ga = (size_t)cpi.site0 + m, gb = (size_t)cpj.site0 + n;
										const size_t first = ci <= cjd ? ga * a.S + gb : gb * a.S + ga;
										rdf_hit<SHARED>(a, sh, a.off_site + first * a.bins + bin);
										if (ci == cjd && m != n) rdf_hit<SHARED>(a, sh, a.off_site + (gb * a.S + ga) * a.bins + bin);
									}
							}
						}
					}
				}
	}
	if (SHARED) {
		__syncthreads();
		for (size_t k = threadIdx.x; k < a.total; k += blockDim.x)
			if (sh[k]) atomicAdd(&a.hist[k], (unsigned long long)sh[k]);
	}
}

int k_rdf(ls1g_ctx* c, const ls1g_rdf_params* p, uint64_t* mol, uint64_t* site, uint64_t* ardf) {
	const int nc = c->n_comp, S = c->n_site_rows;
	int max_sites = 1, do_site = 0;
	for (int i = 0; i < nc; ++i) {
		max_sites = std::max(max_sites, c->comps_h[i].n_sites);
		for (int j = i; j < nc; ++j) if (c->comps_h[i].n_sites + c->comps_h[j].n_sites > 2) do_site = 1;
	}
	if (do_site && !site) LS1G_FAIL("ls1g_rdf: site histogram missing (a component pair has more than two sites)");
	const int do_ardf = p->angular_bins > 1;
	if (do_ardf && !ardf) LS1G_FAIL("ls1g_rdf: angular histogram missing (angular_bins > 1)");
	RdfArgs a;
	a.g = c->bgrid; a.n = c->n_slots;
	a.pos = c->cur.pos; a.q = c->has_rot ? c->cur.q : nullptr; a.cell = c->cur.cell; a.cell_start = c->cell_start; a.comps = c->comps;
	a.sitepos = nullptr; a.max_sites = max_sites;
	a.rc2 = c->rc2;
	a.max2 = p->interval_length * p->interval_length * p->bins * p->bins;      // RDF::init (io/RDF.cpp:45)
	a.bw = p->interval_length; a.aw = do_ardf ? 2. / p->angular_bins : 2.;
	a.bins = p->bins; a.abins = do_ardf ? p->angular_bins : 1;
	a.nc = nc; a.S = S; a.do_site = do_site; a.do_ardf = do_ardf;
	const size_t n_mol = (size_t)nc * (nc + 1) / 2 * p->bins;
	const size_t n_site = do_site ? (size_t)S * S * p->bins : 0;
	const size_t n_ardf = do_ardf ? (size_t)nc * nc * p->bins * p->angular_bins : 0;
	a.off_site = n_mol; a.off_ardf = n_mol + n_site; a.total = n_mol + n_site + n_ardf;
	if (a.total == 0) return 0;
	unsigned long long* d_hist = nullptr;
	double* d_sites = nullptr;
	CUDA_TRY(cudaMalloc((void**)&d_hist, a.total * sizeof(unsigned long long)));
	cudaError_t e = cudaMemsetAsync(d_hist, 0, a.total * sizeof(unsigned long long), c->stream);
	if (e == cudaSuccess && do_site && c->n_slots) e = cudaMalloc((void**)&d_sites, c->n_slots * (size_t)max_sites * 3 * sizeof(double));
	a.hist = d_hist; a.sitepos = d_sites;
	std::vector<unsigned long long> h(a.total);
	if (e == cudaSuccess && c->n_slots) {
		timer_start(c, LS1G_T_AUX);
		int launches = 1;
		if (do_site) {
			rdf_sites_kernel<<<(unsigned)((c->n_slots + 255) / 256), 256, 0, c->stream>>>(c->n_slots, c->cur.pos, a.q, c->comps, c->sites, d_sites, max_sites);
			++launches;
		}
		const unsigned blocks = (unsigned)((c->n_slots + RDF_THREADS - 1) / RDF_THREADS);
		const size_t smem = a.total * sizeof(uint32_t);
		if (smem <= 96 * 1024) {       // block-private histogram, flushed once per block
			e = cudaFuncSetAttribute(rdf_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
			if (e == cudaSuccess) rdf_kernel<true><<<blocks, RDF_THREADS, smem, c->stream>>>(a);
		} else {
			rdf_kernel<false><<<blocks, RDF_THREADS, 0, c->stream>>>(a);
		}
		timer_stop(c, LS1G_T_AUX, launches);
	}
	if (e == cudaSuccess) e = cudaMemcpyAsync(h.data(), d_hist, a.total * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream);
	if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
	if (e == cudaSuccess) e = cudaGetLastError();
	cudaFree(d_hist);
	if (d_sites) cudaFree(d_sites);
	if (e != cudaSuccess) LS1G_FAIL(std::string("ls1g_rdf: ") + cudaGetErrorString(e));
	for (size_t k = 0; k < n_mol; ++k) mol[k] += h[k];
	for (size_t k = 0; k < n_site; ++k) site[k] += h[a.off_site + k];
	for (size_t k = 0; k < n_ardf; ++k) ardf[k] += h[a.off_ardf + k];
	return 0;
}
