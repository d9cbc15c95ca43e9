// api.cu — the C ABI declared in include/ls1g.h: context life cycle, parameter tables, storage, the per-step
// operators and the device-resident time loop.  Kernels are in cells.cu / force*.cu / integrate.cu / comm.cu.
#include "ls1g_internal.cuh"

#include <math.h>
#include <string.h>
#include <stdlib.h>
#include <algorithm>
#include <mutex>

static thread_local std::string g_err;
void ls1g_set_error(const std::string& msg) { g_err = msg; }
extern "C" const char* ls1g_last_error(void) { return g_err.c_str(); }
extern "C" int ls1g_version(void) { return 100; }
extern "C" int ls1g_device_count(int* n) {
	if (!n) LS1G_FAIL("null argument");
	*n = 0;
	cudaError_t e = cudaGetDeviceCount(n);
	if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) { *n = 0; (void)cudaGetLastError(); return 0; }      // a host without a usable GPU: 0 devices
	if (e != cudaSuccess) { *n = 0; LS1G_FAIL(std::string("ls1g_device_count: ") + cudaGetErrorString(e)); }
	return 0;
}


// ------------------------------------------------------------------------------------------------
// timers: event pairs are queued and resolved lazily (no host sync inside the step)
// ------------------------------------------------------------------------------------------------
static std::vector<PendingEvt>& pending(ls1g_ctx* c) { return c->pending; }
static std::vector<cudaEvent_t>& evpool(ls1g_ctx* c) { return c->evpool; }

static cudaEvent_t get_event(ls1g_ctx* c) {
	auto& pool = evpool(c);
	if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
	cudaEvent_t e; cudaEventCreate(&e); return e;
}
static void resolve_timers(ls1g_ctx* c) {
	auto& p = pending(c);
	for (auto& e : p) {
		if (!e.b) { evpool(c).push_back(e.a); continue; }
		cudaEventSynchronize(e.b);
		float ms = 0.f;
		if (cudaEventElapsedTime(&ms, e.a, e.b) == cudaSuccess) { c->timers[e.which].ms += ms; c->timers[e.which].launches += e.launches; }
		evpool(c).push_back(e.a); evpool(c).push_back(e.b);
	}
	p.clear();
}
void ls1g_resolve_timers(ls1g_ctx* c) { resolve_timers(c); }
void timer_start(ls1g_ctx* c, int which) {
	if (pending(c).size() > 8192) resolve_timers(c);
	PendingEvt e; e.a = get_event(c); e.b = nullptr; e.which = which; e.launches = 0;
	cudaEventRecord(e.a, c->stream);
	pending(c).push_back(e);
}
void timer_stop(ls1g_ctx* c, int which, int launches) {
	auto& p = pending(c);
	for (size_t k = p.size(); k-- > 0;)
		if (p[k].which == which && !p[k].b) { p[k].b = get_event(c); p[k].launches = launches; cudaEventRecord(p[k].b, c->stream); break; }
}

// ------------------------------------------------------------------------------------------------
// grid geometry: LinkedCells::rebuild (LinkedCells.cpp:136-204, float rc) / constructor (:42-107, double rc)
// ------------------------------------------------------------------------------------------------
// cut: cell size bound and halo slab width; mode LS1G_GRID_* or -1 (plain double arithmetic, internal bin grids)
int grid_init_cut(Grid& g, const ls1g_config& cfg, double cut, int mode) {
	memset(&g, 0, sizeof(g));
	g.cutoff = cut;
	long ncell = 1;
	for (int d = 0; d < 3; ++d) {
		g.box_min[d] = cfg.box_min[d]; g.box_max[d] = cfg.box_max[d];
		double diff = cfg.box_max[d] - cfg.box_min[d];
		int n;
		if (mode == LS1G_GRID_CTOR || mode < 0) {
			n = (int)floor(diff / cut);                   // LinkedCells.cpp:53-54
			if (n == 0) n = 1;                            // :56-58
		} else {
			float rcf = (float)(cut / 1);                 // :152 float rc = _cutoffRadius / _cellsInCutoff
			n = (int)floor(diff / rcf);                   // :155
			if (n == 0) LS1G_FAIL("ls1g_create: bounding box too small for the cutoff (LinkedCells.cpp:160-163)");
		}
		g.cpd[d] = n + 2;
		g.cell_len[d] = diff / n;                         // :168
		g.cell_len_recip[d] = n / diff;                   // :169
		g.halo_min[d] = cfg.box_min[d] - g.cell_len[d];
		g.halo_max[d] = cfg.box_max[d] + g.cell_len[d];
		g.box_len[d] = diff;
		g.self_periodic[d] = cfg.rank_grid[d] <= 1 ? 1 : 0;
		ncell *= g.cpd[d];
	}
	if (ncell > 0x7fffffffL - 8) LS1G_FAIL("ls1g_create: too many cells");
	g.n_cells = (int)ncell;
	return 0;
}
static int grid_init(Grid& g, const ls1g_config& cfg) { return grid_init_cut(g, cfg, cfg.cutoff, cfg.grid_mode); }

// ------------------------------------------------------------------------------------------------
// life cycle
// ------------------------------------------------------------------------------------------------
static void free_molecules(ls1g_ctx* c) {
	MolBuf* bufs[2] = {&c->cur, &c->alt};
	for (MolBuf* b : bufs) {
		cudaFree(b->pos); cudaFree(b->vel); cudaFree(b->q); cudaFree(b->L); cudaFree(b->id); cudaFree(b->cell);
		memset(b, 0, sizeof(MolBuf));
	}
	cudaFree(c->F); cudaFree(c->M); cudaFree(c->Vi); cudaFree(c->F_alt); cudaFree(c->M_alt); cudaFree(c->Vi_alt); cudaFree(c->site_cache);
	c->Vi_alt = nullptr;
	cudaFree(c->slot_src); cudaFree(c->slot_code); cudaFree(c->slot_key); cudaFree(c->slot_cell);
	cudaFree(c->stage); c->stage = nullptr; c->stage_bytes = 0;
	cudaFree(c->halo_src); cudaFree(c->halo_code); cudaFree(c->slot_dst); cudaFree(c->pos_build); cudaFree(c->lists); cudaFree(c->lcount);
	cudaFree(c->pos_prune); cudaFree(c->lists_in); cudaFree(c->lcount_in);
	c->pos_prune = nullptr; c->lists_in = nullptr; c->lcount_in = nullptr; c->lists_in_cap = c->lcount_in_cap = 0; c->inner_valid = 0;
	c->halo_src = nullptr; c->halo_code = nullptr; c->slot_dst = nullptr; c->pos_build = nullptr; c->lists = nullptr; c->lcount = nullptr;
	c->lists_cap = c->lcount_cap = 0;
	c->F = c->M = c->Vi = c->F_alt = c->M_alt = nullptr; c->site_cache = nullptr;
	c->slot_src = nullptr; c->slot_code = nullptr; c->slot_key = nullptr; c->slot_cell = nullptr;
	c->cap = 0;
}

static int alloc_molecules(ls1g_ctx* c, size_t cap) {
	free_molecules(c);
	if (cap >= 0x7fffff00ull) LS1G_FAIL("ls1g: more than 2^31 slots per GPU are not supported");
	MolBuf* bufs[2] = {&c->cur, &c->alt};
	for (MolBuf* b : bufs) {
		CUDA_TRY(cudaMalloc((void**)&b->pos, cap * sizeof(double4)));
		CUDA_TRY(cudaMalloc((void**)&b->vel, cap * 3 * sizeof(double)));
		CUDA_TRY(cudaMalloc((void**)&b->id, cap * sizeof(uint64_t)));
		CUDA_TRY(cudaMalloc((void**)&b->cell, cap * sizeof(uint32_t)));
		if (c->has_rot) {
			CUDA_TRY(cudaMalloc((void**)&b->q, cap * sizeof(double4)));
			CUDA_TRY(cudaMalloc((void**)&b->L, cap * 3 * sizeof(double)));
		}
	}
	CUDA_TRY(cudaMalloc((void**)&c->F, cap * 3 * sizeof(double)));
	CUDA_TRY(cudaMalloc((void**)&c->Vi, cap * 3 * sizeof(double)));
	CUDA_TRY(cudaMalloc((void**)&c->F_alt, cap * 3 * sizeof(double)));
	CUDA_TRY(cudaMalloc((void**)&c->Vi_alt, cap * 3 * sizeof(double)));
	CUDA_TRY(cudaMemset(c->F, 0, cap * 3 * sizeof(double)));
	CUDA_TRY(cudaMemset(c->Vi, 0, cap * 3 * sizeof(double)));
	if (c->has_rot) {
		CUDA_TRY(cudaMalloc((void**)&c->M, cap * 3 * sizeof(double)));
		CUDA_TRY(cudaMalloc((void**)&c->M_alt, cap * 3 * sizeof(double)));
		CUDA_TRY(cudaMemset(c->M, 0, cap * 3 * sizeof(double)));
		CUDA_TRY(cudaMalloc((void**)&c->site_cache, cap * (size_t)c->cache_stride * sizeof(double4)));
	}
	CUDA_TRY(cudaMalloc((void**)&c->slot_src, cap * sizeof(uint32_t)));
	CUDA_TRY(cudaMalloc((void**)&c->slot_code, cap * sizeof(uint8_t)));
	CUDA_TRY(cudaMalloc((void**)&c->slot_key, cap * sizeof(uint64_t)));
	CUDA_TRY(cudaMalloc((void**)&c->slot_cell, cap * sizeof(uint32_t)));
	CUDA_TRY(cudaMalloc((void**)&c->slot_dst, cap * sizeof(uint32_t)));
	CUDA_TRY(cudaMalloc((void**)&c->halo_src, cap * sizeof(uint32_t)));
	CUDA_TRY(cudaMalloc((void**)&c->halo_code, cap * sizeof(uint8_t)));
	c->cap = cap;
	return 0;
}

extern "C" int ls1g_create(const ls1g_config* cfg, ls1g_ctx** out) {
	if (!cfg || !out) LS1G_FAIL("ls1g_create: null argument");
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev == 0)
		LS1G_FAIL(std::string("ls1g_create: no CUDA device available (there is no CPU fallback): ") + cudaGetErrorString(e));
	if (cfg->device < 0 || cfg->device >= ndev) LS1G_FAIL("ls1g_create: invalid device ordinal");
	CUDA_TRY(cudaSetDevice(cfg->device));
	ls1g_ctx* c = new ls1g_ctx();
	// zero the POD part
	c->cfg = *cfg;
	for (int d = 0; d < 3; ++d) if (c->cfg.rank_grid[d] < 1) c->cfg.rank_grid[d] = 1;
	if (c->cfg.lj_cutoff <= 0) c->cfg.lj_cutoff = c->cfg.cutoff;
	c->stream = nullptr;
	c->n_comp = 0; c->n_site_rows = 0; c->S_lj = 0; c->max_sites = 0; c->has_rot = 0; c->simple_1clj = 0;
	c->comps = nullptr; c->sites = nullptr; c->eps24 = c->sig2 = c->shift6 = nullptr;
	c->cap = c->n_slots = c->n_owned = c->n_halo = 0; c->sorted = 0; c->forces_valid = 0;
	memset(&c->cur, 0, sizeof(MolBuf)); memset(&c->alt, 0, sizeof(MolBuf));
	c->F = c->M = c->Vi = c->F_alt = c->M_alt = nullptr; c->site_cache = nullptr; c->cache_stride = 1;
	c->stage = nullptr; c->stage_bytes = 0;
	c->cell_count = c->cell_fill = c->cell_start = c->scan_block = nullptr;
	c->slot_src = nullptr; c->slot_code = nullptr; c->slot_key = nullptr; c->slot_cell = nullptr;
	c->d_counters = c->h_counters = nullptr;
	c->partials = nullptr; c->partials_cap = 0; c->d_scalars = c->h_scalars = nullptr; c->d_pairs = nullptr; c->last_pairs = 0;
	c->nccl_comm = nullptr; c->n_ranks = 1; c->rank = 0; c->comm_mode = 0;
	c->send_buf[0] = c->send_buf[1] = c->recv_buf[0] = c->recv_buf[1] = nullptr; c->xbuf_cap = 0;
	c->gbuf = nullptr; c->gbuf_owner = 0; c->gseq = 0; c->d_rec = nullptr; c->d_rec_cap = 0;
	c->ev_pack = c->ev_read = c->ev_ar = nullptr;
	memset(c->timers, 0, sizeof(c->timers));
	{ const char* fk = getenv("LS1G_FORCE_KERNEL"); c->force_variant = (fk && !strcmp(fk, "simple")) ? 1 : 0; }
	if (grid_init(c->grid, c->cfg)) { delete c; return 1; }
	c->bgrid = c->grid;
	{ const char* e = getenv("LS1G_VI_EVERY_STEP"); c->vi_every_step = e && atoi(e) != 0; }
	{ const char* lr = getenv("LS1G_LIST_ROWS"); c->list_R = lr && atoi(lr) == 1 ? 1 : 4; }
	// bank-aware list order: 1 (default) the build kernel orders every 16-entry chunk of a list as it writes it (lists.cuh: emit);
	// 0 tile order; 2 tile order + the separate whole-list pass nl_permute_kernel (measurement only: 5.6 ms per build at 10^7 molecules,
	// profiles/r02c_permute_tradeoff.json)
	{ const char* lp = getenv("LS1G_LIST_PERMUTE"); c->list_permute = lp ? atoi(lp) : 1; if (c->list_permute < 0 || c->list_permute > 2) c->list_permute = 1; }
	{ const char* lm = getenv("LS1G_LIST_MULTISITE"); c->list_multisite = lm && atoi(lm) == 0 ? 0 : 1; }
	// skin: LS1G_SKIN, else 0.12 cutoff.  LS1G_SKIN_INNER > 0 (default 0 = off) adds pruned inner lists for single-centre systems
	// (lists.cuh: nl_prune_kernel); the outer skin then defaults to 0.2 cutoff (ls1g_set_components).  Measured at 10^7 molecules:
	// no gain - the pruned lists lose the bank-aware order of the build, DESIGN.md section 3
	{ const char* sk = getenv("LS1G_SKIN"); c->skin_user = sk != nullptr; c->skin = sk ? atof(sk) : 0.12 * c->cfg.cutoff; if (c->skin < 0.) c->skin = 0.; }
	{ const char* si = getenv("LS1G_SKIN_INNER"); c->skin_in = si ? atof(si) : 0.; if (c->skin_in < 0.) c->skin_in = 0.; }
	c->rc2 = c->cfg.cutoff * c->cfg.cutoff;
	c->rclj2 = c->cfg.lj_cutoff * c->cfg.lj_cutoff;
	// VectorizedCellProcessor.cpp:24
	c->epsRFInvrc3 = 2. * (c->cfg.eps_rf - 1.) / ((c->cfg.cutoff * c->cfg.cutoff * c->cfg.cutoff) * (2. * c->cfg.eps_rf + 1.));
	CUDA_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
	{
		int lo = 0, hi = 0;
		CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
		CUDA_TRY(cudaStreamCreateWithPriority(&c->stream_comm, cudaStreamNonBlocking, hi));
		CUDA_TRY(cudaEventCreateWithFlags(&c->ev_pos, cudaEventDisableTiming));
		CUDA_TRY(cudaEventCreateWithFlags(&c->ev_halo, cudaEventDisableTiming));
		CUDA_TRY(cudaEventCreateWithFlags(&c->ev_pack, cudaEventDisableTiming));
		CUDA_TRY(cudaEventCreateWithFlags(&c->ev_read, cudaEventDisableTiming));
		CUDA_TRY(cudaEventCreateWithFlags(&c->ev_ar, cudaEventDisableTiming));
	}
	size_t nc = (size_t)c->grid.n_cells + 1;
	CUDA_TRY(cudaMalloc((void**)&c->cell_count, nc * sizeof(uint32_t)));
	CUDA_TRY(cudaMalloc((void**)&c->cell_fill, nc * sizeof(uint32_t)));
	CUDA_TRY(cudaMalloc((void**)&c->cell_start, (nc + 8) * sizeof(uint32_t)));
	CUDA_TRY(cudaMemset(c->cell_start, 0, (nc + 8) * sizeof(uint32_t)));
	CUDA_TRY(cudaMalloc((void**)&c->scan_block, (nc / 2048 + 8) * sizeof(uint32_t)));
	CUDA_TRY(cudaMalloc((void**)&c->d_counters, 16 * sizeof(uint32_t)));
	CUDA_TRY(cudaMemset(c->d_counters, 0, 16 * sizeof(uint32_t)));
	CUDA_TRY(cudaMallocHost((void**)&c->h_counters, 16 * sizeof(uint32_t)));
	CUDA_TRY(cudaMalloc((void**)&c->d_scalars, 16 * sizeof(double)));
	CUDA_TRY(cudaMemset(c->d_scalars, 0, 16 * sizeof(double)));
	CUDA_TRY(cudaMallocHost((void**)&c->h_scalars, 16 * sizeof(double)));
	CUDA_TRY(cudaMalloc((void**)&c->d_flags, 16 * sizeof(unsigned long long)));
	CUDA_TRY(cudaMemset(c->d_flags, 0, 16 * sizeof(unsigned long long)));
	CUDA_TRY(cudaMallocHost((void**)&c->h_flags, 48 * sizeof(unsigned long long)));
	memset(c->h_flags, 0, 48 * sizeof(unsigned long long));
	for (int k = 0; k < 16; ++k) CUDA_TRY(cudaEventCreateWithFlags(&c->disp_ev[k], cudaEventDisableTiming));
	*out = c;
	return 0;
}

extern "C" int ls1g_destroy(ls1g_ctx* c) {
	if (!c) return 0;
	cudaSetDevice(c->cfg.device);
	if (c->stream) cudaStreamSynchronize(c->stream);
	resolve_timers(c);
	comm_destroy(c);
	for (cudaEvent_t e : evpool(c)) cudaEventDestroy(e);
	free_molecules(c);
	cudaFree(c->comps); cudaFree(c->sites); cudaFree(c->eps24); cudaFree(c->sig2); cudaFree(c->shift6);
	cudaFree(c->cell_count); cudaFree(c->cell_fill); cudaFree(c->cell_start); cudaFree(c->scan_block);
	cudaFree(c->d_flags); cudaFreeHost(c->h_flags);
	for (int k = 0; k < 16; ++k) if (c->disp_ev[k]) cudaEventDestroy(c->disp_ev[k]);
	cudaFree(c->d_counters); cudaFreeHost(c->h_counters); cudaFree(c->partials); cudaFree(c->d_scalars); cudaFreeHost(c->h_scalars);
	for (int k = 0; k < 2; ++k) { cudaFree(c->send_buf[k]); cudaFree(c->recv_buf[k]); }
	if (c->ev_pos) cudaEventDestroy(c->ev_pos);
	if (c->ev_halo) cudaEventDestroy(c->ev_halo);
	if (c->ev_pack) cudaEventDestroy(c->ev_pack);
	if (c->ev_read) cudaEventDestroy(c->ev_read);
	if (c->ev_ar) cudaEventDestroy(c->ev_ar);
	cudaFree(c->d_rec);
	if (c->stream_comm) cudaStreamDestroy(c->stream_comm);
	cudaFree(c->bmap); cudaFree(c->bflag); cudaFree(c->plans);
	if (c->stream) cudaStreamDestroy(c->stream);
	delete c;
	return 0;
}

// ------------------------------------------------------------------------------------------------
// components + mixing tables (Comp2Param::initialize, Comp2Param.cpp:10-100; VCP ctor :21-98)
// ------------------------------------------------------------------------------------------------
extern "C" int ls1g_set_components(ls1g_ctx* c, int nc, const ls1g_component* comps, const double* mix, int n_mix) {
	if (!c || nc <= 0 || !comps) LS1G_FAIL("ls1g_set_components: bad arguments");
	if (n_mix < nc * (nc - 1)) LS1G_FAIL("ls1g_set_components: mixcoeff needs (xi,eta) for every component pair i<j");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	c->comps_h.clear(); c->sites_h.clear();
	int S = 0, max_rows = 1;
	for (int k = 0; k < nc; ++k) {
		const ls1g_component& in = comps[k];
		CompDev cd; memset(&cd, 0, sizeof(cd));
		cd.mass = in.mass;
		for (int d = 0; d < 3; ++d) { cd.I[d] = in.inertia[d]; cd.invI[d] = in.inertia[d] != 0. ? 1. / in.inertia[d] : 0.; }  // FullMolecule.h:88-101
		cd.rot_dof = in.rot_dof;
		cd.n_lj = in.n_lj; cd.n_ch = in.n_charge; cd.n_di = in.n_dipole; cd.n_qu = in.n_quadrupole;
		cd.lj_id0 = S; cd.site0 = (int)c->sites_h.size();
		cd.n_sites = cd.n_lj + cd.n_ch + cd.n_di + cd.n_qu;
		for (int a = 0; a < in.n_lj; ++a) {
			SiteDev s; memset(&s, 0, sizeof(s)); const double* r = in.lj + 7 * a;
			s.d[0] = r[0]; s.d[1] = r[1]; s.d[2] = r[2]; s.type = SITE_LJ; s.ljid = S + a; c->sites_h.push_back(s);
		}
		for (int a = 0; a < in.n_charge; ++a) {
			SiteDev s; memset(&s, 0, sizeof(s)); const double* r = in.charge + 5 * a;
			s.d[0] = r[0]; s.d[1] = r[1]; s.d[2] = r[2]; s.val = r[4]; s.type = SITE_CH; c->sites_h.push_back(s);
		}
		for (int a = 0; a < in.n_dipole; ++a) {
			SiteDev s; memset(&s, 0, sizeof(s)); const double* r = in.dipole + 7 * a;
			for (int d = 0; d < 3; ++d) { s.d[d] = r[d]; s.e[d] = r[3 + d]; } s.val = r[6]; s.type = SITE_DI; c->sites_h.push_back(s);
		}
		for (int a = 0; a < in.n_quadrupole; ++a) {
			SiteDev s; memset(&s, 0, sizeof(s)); const double* r = in.quadrupole + 7 * a;
			for (int d = 0; d < 3; ++d) { s.d[d] = r[d]; s.e[d] = r[3 + d]; } s.val = r[6]; s.type = SITE_QU; c->sites_h.push_back(s);
		}
		S += in.n_lj;
		max_rows = std::max(max_rows, cd.n_lj + cd.n_ch + 2 * cd.n_di + 2 * cd.n_qu);
		c->comps_h.push_back(cd);
	}
	c->n_comp = nc; c->S_lj = S; c->cache_stride = max_rows; c->n_site_rows = (int)c->sites_h.size();
	// LJ tables indexed by global LJ-site id
	size_t SS = (size_t)std::max(S, 1) * std::max(S, 1);
	c->eps24_h.assign(SS, 0.); c->sig2_h.assign(SS, 0.); c->shift6_h.assign(SS, 0.);
	const double rclj = c->cfg.lj_cutoff;
	int mixpos = 0;
	for (int ci = 0; ci < nc; ++ci) {
		const ls1g_component& A = comps[ci]; int oi = c->comps_h[ci].lj_id0;
		for (int a = 0; a < A.n_lj; ++a)                                            // Comp2Param.cpp:24-41
			for (int b = 0; b < A.n_lj; ++b) {
				const double *si = A.lj + 7 * a, *sj = A.lj + 7 * b;
				double s2 = .5 * (si[5] + sj[5]); s2 *= s2;
				size_t k = (size_t)(oi + a) * S + (oi + b);
				c->eps24_h[k] = 24. * sqrt(si[4] * sj[4]); c->sig2_h[k] = s2; c->shift6_h[k] = si[6];
			}
		for (int cj = ci + 1; cj < nc; ++cj) {                                      // :43-98
			const ls1g_component& B = comps[cj]; int oj = c->comps_h[cj].lj_id0;
			double xi = mix[mixpos++], eta = mix[mixpos++];
			for (int a = 0; a < A.n_lj; ++a)
				for (int b = 0; b < B.n_lj; ++b) {
					const double *si = A.lj + 7 * a, *sj = B.lj + 7 * b;
					double e24 = 24. * xi * sqrt(si[4] * sj[4]);
					double s2 = eta * .5 * (si[5] + sj[5]); s2 *= s2;
					double sr2 = s2 / (rclj * rclj), sr6 = sr2 * sr2 * sr2;
					double sh = e24 * (sr6 - sr6 * sr6);
					size_t k = (size_t)(oi + a) * S + (oj + b), kt = (size_t)(oj + b) * S + (oi + a);
					c->eps24_h[k] = e24; c->sig2_h[k] = s2; c->shift6_h[k] = sh;
					c->eps24_h[kt] = e24; c->sig2_h[kt] = s2; c->shift6_h[kt] = sh;
				}
		}
	}
	const CompDev& c0 = c->comps_h[0];
	const SiteDev* s0 = c->sites_h.empty() ? nullptr : &c->sites_h[0];
	c->simple_1clj = (nc == 1 && c0.n_lj == 1 && c0.n_ch == 0 && c0.n_di == 0 && c0.n_qu == 0 && c0.rot_dof == 0 &&
	                  s0->d[0] == 0. && s0->d[1] == 0. && s0->d[2] == 0.);
	c->has_rot = !c->simple_1clj;
	// single-centre systems run on pruned inner lists (lists.cuh: nl_prune_kernel): the outer lists can afford the larger skin
	if (!c->skin_user) c->skin = (c->simple_1clj && c->skin_in > 0.) ? 0.2 * c->cfg.cutoff : 0.12 * c->cfg.cutoff;
	if (c->skin_in >= 0.5 * c->skin) c->skin_in = 0.;      // no room between the two radii: inner lists off
	cudaFree(c->comps); cudaFree(c->sites); cudaFree(c->eps24); cudaFree(c->sig2); cudaFree(c->shift6);
	CUDA_TRY(cudaMalloc((void**)&c->comps, nc * sizeof(CompDev)));
	CUDA_TRY(cudaMalloc((void**)&c->sites, std::max<size_t>(1, c->sites_h.size()) * sizeof(SiteDev)));
	CUDA_TRY(cudaMalloc((void**)&c->eps24, SS * sizeof(double)));
	CUDA_TRY(cudaMalloc((void**)&c->sig2, SS * sizeof(double)));
	CUDA_TRY(cudaMalloc((void**)&c->shift6, SS * sizeof(double)));
	CUDA_TRY(cudaMemcpy(c->comps, c->comps_h.data(), nc * sizeof(CompDev), cudaMemcpyHostToDevice));
	if (!c->sites_h.empty()) CUDA_TRY(cudaMemcpy(c->sites, c->sites_h.data(), c->sites_h.size() * sizeof(SiteDev), cudaMemcpyHostToDevice));
	CUDA_TRY(cudaMemcpy(c->eps24, c->eps24_h.data(), SS * sizeof(double), cudaMemcpyHostToDevice));
	CUDA_TRY(cudaMemcpy(c->sig2, c->sig2_h.data(), SS * sizeof(double), cudaMemcpyHostToDevice));
	CUDA_TRY(cudaMemcpy(c->shift6, c->shift6_h.data(), SS * sizeof(double), cudaMemcpyHostToDevice));
	free_molecules(c);   // storage layout depends on has_rot / cache_stride
	c->n_slots = c->n_owned = c->n_halo = 0; c->sorted = 0; c->forces_valid = 0;
	return 0;
}

// ------------------------------------------------------------------------------------------------
// grid queries
// ------------------------------------------------------------------------------------------------
extern "C" int ls1g_get_grid(const ls1g_ctx* c, int cells_per_dim[3], double cell_length[3]) {
	if (!c) LS1G_FAIL("null context");
	for (int d = 0; d < 3; ++d) { cells_per_dim[d] = c->grid.cpd[d]; cell_length[d] = c->grid.cell_len[d]; }
	return 0;
}
extern "C" int ls1g_cell_index_of_points_host(const ls1g_ctx* c, size_t n, const double* pts, uint32_t* out) {
	if (!c) LS1G_FAIL("null context");
	for (size_t i = 0; i < n; ++i) out[i] = grid_cell_index(c->grid, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
	return 0;
}
extern "C" int ls1g_cell_index_of_points(ls1g_ctx* c, size_t n, const double* pts, uint32_t* out) {
	if (!c) LS1G_FAIL("null context");
	if (n == 0) return 0;
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	double* d_p = nullptr; uint32_t* d_o = nullptr;
	CUDA_TRY(cudaMalloc((void**)&d_p, n * 3 * sizeof(double)));
	CUDA_TRY(cudaMalloc((void**)&d_o, n * sizeof(uint32_t)));
	int rc = 0;
	if (cudaMemcpyAsync(d_p, pts, n * 3 * sizeof(double), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) rc = 1;
	if (!rc) rc = k_cell_index(c, n, d_p, d_o);
	if (!rc && cudaMemcpyAsync(out, d_o, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess) rc = 1;
	cudaError_t e = cudaStreamSynchronize(c->stream);
	cudaFree(d_p); cudaFree(d_o);
	if (e != cudaSuccess) { ls1g_set_error(std::string("cell index kernel: ") + cudaGetErrorString(e)); return 1; }
	return rc;
}
extern "C" int ls1g_debug_cell_pairs(const ls1g_ctx* c, uint32_t* triplets, size_t capacity, size_t* n_out) {
	if (!c || !n_out) LS1G_FAIL("null argument");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	return k_debug_cell_pairs(c, triplets, capacity, n_out);
}

// ------------------------------------------------------------------------------------------------
// storage
// ------------------------------------------------------------------------------------------------
static size_t default_capacity(const ls1g_ctx* c, size_t n) {
	// halo copies live in the shell of halo cells: N * (cells incl. halo / real cells - 1), with head room.  The neighbour-list loop
	// bins into cells of cutoff + skin, whose shell is thicker: the larger of the two ratios counts.
	auto shell = [](const Grid& g) {
		double tot = 1., real = 1.;
		for (int d = 0; d < 3; ++d) { tot *= g.cpd[d]; real *= (g.cpd[d] - 2); }
		return tot / real - 1.;
	};
	double ratio = shell(c->grid);
	if (c->skin > 0.) {
		Grid g;
		if (grid_init_cut(g, c->cfg, std::max(c->cfg.cutoff, c->cfg.lj_cutoff) + c->skin, -1) == 0) ratio = std::max(ratio, shell(g));
	}
	double est = (double)n * ratio * 1.3 + 4096;
	double hard = 27.0 * (double)n + 4096;
	return n + (size_t)std::min(est, hard) + (size_t)(0.05 * n);
}

int k_reserve_stage(ls1g_ctx* c);
int k_list_reserve(ls1g_ctx* c, size_t n);      // lists.cu
extern "C" int ls1g_reserve(ls1g_ctx* c, size_t n) {
	if (!c) LS1G_FAIL("null context");
	if (c->n_comp == 0) LS1G_FAIL("ls1g_reserve: call ls1g_set_components first");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	const size_t want = c->cfg.capacity ? (size_t)c->cfg.capacity : default_capacity(c, n);
	if (want < n) LS1G_FAIL("ls1g_reserve: more molecules than the configured capacity");
	if (c->cap < want || c->cap == 0) {
		if (c->n_slots) LS1G_FAIL("ls1g_reserve: the context already holds molecules");
		if (alloc_molecules(c, std::max<size_t>(want, 1))) return 1;
	}
	if (k_reserve_stage(c)) return 1;
	if (k_list_reserve(c, n)) return 1;
	CUDA_TRY(cudaDeviceSynchronize());
	return 0;
}

extern "C" int ls1g_upload(ls1g_ctx* c, size_t n, const uint64_t* id, const uint32_t* cid,
                           const double* r, const double* v, const double* q, const double* L) {
	if (!c) LS1G_FAIL("null context");
	if (c->n_comp == 0) LS1G_FAIL("ls1g_upload: call ls1g_set_components first");
	if (n && (!id || !cid || !r)) LS1G_FAIL("ls1g_upload: id, cid and r are required");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	size_t want = c->cfg.capacity ? (size_t)c->cfg.capacity : default_capacity(c, n);
	if (want < n) LS1G_FAIL("ls1g_upload: more molecules than the configured capacity");
	if (c->cap < want || c->cap == 0) { if (alloc_molecules(c, std::max<size_t>(want, 1))) return 1; }
	if (c->in_list_mode) { c->in_list_mode = 0; c->track_disp = 0; c->bgrid = c->grid; }     // the content is replaced
	if (k_upload(c, n, id, cid, r, v, q, L)) return 1;
	c->n_uploaded = n;
	c->n_slots = c->n_owned = n; c->n_halo = 0; c->sorted = 0; c->forces_valid = 0;
	return 0;
}

// --- binary phase-space records (records.cu) ---------------------------------------------------------
int k_upload_records(ls1g_ctx* c, size_t n, const void* records, int format, uint64_t* n_kept, uint64_t* max_id);
int k_download_records(ls1g_ctx* c, double scale_v, double scale_D, size_t n_ret, void* records);

extern "C" int ls1g_record_size(int format, size_t* bytes) {
	if (!bytes) LS1G_FAIL("null argument");
	if (format < LS1G_REC_ICRVQD || format > LS1G_REC_IRV) LS1G_FAIL("Not a valid molecule format");      // BinaryReader.cpp:109-112
	*bytes = format == LS1G_REC_ICRVQD ? 116 : format == LS1G_REC_ICRV ? 60 : 56;
	return 0;
}

extern "C" int ls1g_upload_records(ls1g_ctx* c, size_t n, const void* records, int format, uint64_t* n_kept, uint64_t* max_id) {
	if (!c) LS1G_FAIL("null context");
	if (c->n_comp == 0) LS1G_FAIL("ls1g_upload_records: call ls1g_set_components first");
	if (format < LS1G_REC_ICRVQD || format > LS1G_REC_IRV) LS1G_FAIL("Not a valid molecule format");
	if (n && !records) LS1G_FAIL("ls1g_upload_records: records is required");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	// room: the configured capacity; else, on a rank that owns the whole box, the estimate for all n records as in ls1g_upload; on one
	// rank of several (most records fall outside its box) what an earlier ls1g_reserve left, and the estimate for n only without one
	const bool whole_box = c->cfg.rank_grid[0] * c->cfg.rank_grid[1] * c->cfg.rank_grid[2] <= 1;
	if (c->cap == 0 || (!c->cfg.capacity && whole_box && c->cap < default_capacity(c, n))) {
		const size_t want = c->cfg.capacity ? (size_t)c->cfg.capacity : default_capacity(c, n);
		if (alloc_molecules(c, std::max<size_t>(want, 1))) return 1;
	}
	if (c->in_list_mode) { c->in_list_mode = 0; c->track_disp = 0; c->bgrid = c->grid; }     // the content is replaced
	c->n_uploaded = 0; c->n_slots = c->n_owned = 0; c->n_halo = 0; c->sorted = 0; c->forces_valid = 0;
	uint64_t kept = 0, mx = 0;
	if (k_upload_records(c, n, records, format, &kept, &mx)) return 1;
	c->n_uploaded = (size_t)kept;
	c->n_slots = c->n_owned = (size_t)kept;
	if (n_kept) *n_kept = kept;
	if (max_id) *max_id = mx;
	return 0;
}

extern "C" int ls1g_download_records(ls1g_ctx* c, double scale_v, double scale_D, size_t capacity, void* records, size_t* n_out) {
	if (!c || !n_out) LS1G_FAIL("null argument");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	if (c->in_list_mode && ls1g_nl_leave(c)) return 1;
	const size_t n_ret = c->n_owned;
	*n_out = n_ret;
	if (!records) return 0;       // count query
	if (capacity < n_ret) LS1G_FAIL("ls1g_download_records: capacity too small");
	CUDA_TRY(cudaStreamSynchronize(c->stream));
	return k_download_records(c, scale_v, scale_D, n_ret, records);
}

extern "C" int ls1g_count(ls1g_ctx* c, uint64_t* n_owned, uint64_t* n_halo) {
	if (!c) LS1G_FAIL("null context");
	if (n_owned) *n_owned = c->n_owned;
	if (n_halo) *n_halo = c->n_halo;
	return 0;
}

extern "C" int ls1g_download(ls1g_ctx* c, int mask, int include_halo, size_t capacity, ls1g_fields* out, size_t* n_out) {
	if (!c || !out || !n_out) LS1G_FAIL("null argument");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	if (c->in_list_mode && ls1g_nl_leave(c)) return 1;
	size_t ns = c->n_slots;
	bool halo_ok = include_halo && c->n_halo > 0;
	size_t n_ret = halo_ok ? c->n_owned + c->n_halo : c->n_owned;
	*n_out = n_ret;
	if (mask == 0) return 0;      // count query: the number a following download returns (the list layout has been left above)
	if (capacity < n_ret) LS1G_FAIL("ls1g_download: capacity too small");
	(void)ns;
	CUDA_TRY(cudaStreamSynchronize(c->stream));
	return k_download(c, mask, halo_ok ? 1 : 0, n_ret, out);
}

int k_download_range(ls1g_ctx* c, int mask, size_t first, size_t count, ls1g_fields* out);
extern "C" int ls1g_download_range(ls1g_ctx* c, int mask, size_t first, size_t count, ls1g_fields* out) {
	if (!c || !out) LS1G_FAIL("null argument");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	if (c->in_list_mode && ls1g_nl_leave(c)) return 1;
	if (!c->sorted) LS1G_FAIL("ls1g_download_range: container not sorted (call ls1g_update / ls1g_exchange first)");
	if (first > c->n_slots || count > c->n_slots - first) LS1G_FAIL("ls1g_download_range: range outside the slots");
	return k_download_range(c, mask, first, count, out);
}

extern "C" int ls1g_slot_count(ls1g_ctx* c, uint64_t* n_slots) {
	if (!c || !n_slots) LS1G_FAIL("null argument");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	if (c->in_list_mode && ls1g_nl_leave(c)) return 1;
	*n_slots = c->n_slots;
	return 0;
}

// Molecule::setF / setM for the molecules of the last ls1g_upload, same order (restart with forces:
// the leapfrog half-kick of the next step needs F(t), Leapfrog.cpp:48-68)
extern "C" int ls1g_upload_forces(ls1g_ctx* c, size_t n, const double* F, const double* M) {
	if (!c) LS1G_FAIL("null context");
	if (n != c->n_uploaded || c->sorted || c->n_slots != n) LS1G_FAIL("ls1g_upload_forces: call directly after ls1g_upload with the same n");
	if (n && !F) LS1G_FAIL("ls1g_upload_forces: F is required");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	if (k_upload_forces(c, n, F, M)) return 1;
	c->forces_valid = 1;
	return 0;
}

int k_host_alloc(size_t bytes, void** out);
int k_host_free(void* p);
extern "C" int ls1g_host_alloc(ls1g_ctx* c, size_t bytes, void** out) {
	if (!c || !out) LS1G_FAIL("null argument");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	return k_host_alloc(bytes, out);
}
extern "C" int ls1g_host_free(ls1g_ctx* c, void* p) {
	if (!c) LS1G_FAIL("null context");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	return k_host_free(p);
}

// the CUDA stream all work of this context is issued on (cudaStream_t), for event timing by the caller
extern "C" int ls1g_stream(ls1g_ctx* c, void** stream) {
	if (!c || !stream) LS1G_FAIL("null argument");
	*stream = (void*)c->stream;
	return 0;
}
