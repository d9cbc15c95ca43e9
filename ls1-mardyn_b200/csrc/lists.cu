// lists.cu — host side of the neighbour-list mode of ls1g_run (kernels in lists.cuh).
#include "ls1g_internal.cuh"
#include "potentials.cuh"
#include "lists.cuh"

#include <math.h>
#include <stdlib.h>
#include <algorithm>
#include <string.h>

int ls1g_ensure_partials(ls1g_ctx* c, size_t blocks, int nv);
int k_force_finalize(ls1g_ctx* c, size_t nb);
int ls1g_prereduce(ls1g_ctx* c, size_t& nb, int nv, const double*& base);   // force.cu

static inline unsigned nblk(size_t n, unsigned t) { return (unsigned)((n + t - 1) / t); }

// list mode applies to: single-centre LJ, default kernels, skin > 0, not disabled by an earlier overflow; one rank or several with
// either transport (NCCL, in-process group)
int nl_applicable(const ls1g_ctx* c) {
	return c->simple_1clj && (c->n_ranks == 1 || c->comm_mode == 1 || c->comm_mode == 2) && c->force_variant == 0 && c->skin > 0. && !c->list_disabled;
}

// switch the bin grid: cells >= cutoff + skin (list mode) or the reference grid
int nl_set_bin_grid(ls1g_ctx* c, int list_mode) {
	if (!list_mode) { c->bgrid = c->grid; return 0; }
	Grid g;
	if (grid_init_cut(g, c->cfg, c->cfg.cutoff + c->skin, -1)) return 1;
	if (g.n_cells > c->grid.n_cells) LS1G_FAIL("neighbour-list grid larger than the reference grid");
	c->bgrid = g;
	return 0;
}

static NLArgs nl_args(ls1g_ctx* c) {
	NLArgs a;
	a.g = c->bgrid; a.pos = c->cur.pos; a.cell = c->cur.cell; a.cell_start = c->cell_start;
	a.lists = c->lists; a.lcount = c->lcount; a.plans = (NLPlan*)c->plans; a.flags = c->d_flags; a.bpr = c->list_bpr; a.nyi = c->bgrid.cpd[1] - 2; a.lcap4 = c->list_lcap / 4; a.btcap = c->build_tcap;
	a.nzi = c->bgrid.cpd[2] - 2; a.nyg = (a.nyi + 1) / 2;
	a.tcap = NL_TCAP;
	const double rl = c->cfg.lj_cutoff + c->skin;
	// FP32 filter: coordinates relative to the block centre are < 8 cell lengths: relative error on r^2 near rl^2 < 2e-5;
	// the 3e-4 margin can only over-accept (the force kernel applies the exact FP64 test)
	a.rl2f = (float)(rl * rl * (1.0 + 3e-4));
	a.eps24 = c->eps24_h[0]; a.sig2 = c->sig2_h[0]; a.shift6 = c->shift6_h[0]; a.rclj2 = c->rclj2;
	a.F = c->F; a.Vi = c->Vi; a.partials = c->partials;
	a.guard = c->guard_on ? c->d_flags + 3 : nullptr; a.guard_thr = (unsigned long long)host_double_bits(c->disp_thr2);
	a.bmap = c->bmap; a.bflag = c->bflag; a.phase = 0;
	return a;
}

// blocks = cell rows (R = 1) or 2 x 2 row bundles (R = 4) times the blocks per row
static size_t nl_blocks(const ls1g_ctx* c) {
	const size_t nyi = c->bgrid.cpd[1] - 2, nzi = c->bgrid.cpd[2] - 2;
	const size_t groups = c->list_R == 4 ? ((nyi + 1) / 2) * ((nzi + 1) / 2) : nyi * nzi;
	return groups * (size_t)c->list_bpr;
}
static size_t nl_bt(const ls1g_ctx* c) { return c->list_R == 4 ? NLGeom<4>::BT : NLGeom<1>::BT; }

// dynamic shared memory limits are per function AND per device: set once for every device a context lives on
#define NL_FORCE_SMEM_MAX (3 * (NL_TCAP + 64) * 8)
template <bool COUNT, bool VI, int FUSE>
static cudaError_t nl_attr() {
	cudaError_t e = cudaFuncSetAttribute(nl_force_kernel<COUNT, VI, FUSE, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, NL_FORCE_SMEM_MAX);
	if (e == cudaSuccess) e = cudaFuncSetAttribute(nl_force_kernel<COUNT, VI, FUSE, true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, NL_FORCE_SMEM_MAX);
	if (e == cudaSuccess) e = cudaFuncSetAttribute(nl_force_kernel<COUNT, VI, FUSE, false, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, NL_FORCE_SMEM_MAX);
	if (e == cudaSuccess) e = cudaFuncSetAttribute(nl_force_kernel<COUNT, VI, FUSE, true, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, NL_FORCE_SMEM_MAX);
	return e;
}
static int nl_force_attrs(const ls1g_ctx* c) {
	static unsigned long long done_mask = 0;       // bit d: attributes set on device d (ordinals < 64)
	const int d = c->cfg.device;
	if (d < 64 && (done_mask >> d & 1ull)) return 0;
	CUDA_TRY(cudaFuncSetAttribute(nl_build_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, NL_BUILD_SMEM(1, NL_TCAP)));
	CUDA_TRY(cudaFuncSetAttribute(nl_build_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, NL_BUILD_SMEM(4, NL_TCAP)));
	CUDA_TRY((cudaFuncSetAttribute(nl_permute_kernel<128, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NL_PERM_SMEM(128, 128))));
	CUDA_TRY((cudaFuncSetAttribute(nl_permute_kernel<256, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NL_PERM_SMEM(256, 128))));
	CUDA_TRY((cudaFuncSetAttribute(nl_permute_kernel<512, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NL_PERM_SMEM(512, 64))));
	CUDA_TRY((cudaFuncSetAttribute(nl_permute_kernel<1024, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NL_PERM_SMEM(1024, 32))));
	CUDA_TRY((nl_attr<true, true, 0>()));
	CUDA_TRY((nl_attr<true, false, 0>()));
	CUDA_TRY((nl_attr<false, false, 0>()));
	CUDA_TRY((nl_attr<true, false, 1>()));
	CUDA_TRY((nl_attr<false, false, 1>()));
	CUDA_TRY((nl_attr<true, true, 2>()));
	CUDA_TRY((nl_attr<true, false, 2>()));
	CUDA_TRY((nl_attr<false, false, 2>()));
	if (d < 64) done_mask |= 1ull << d;
	return 0;
}
// ZCHK variant only while the last build saw a (nearly) coincident pair; R as the lists were built (lists.cuh)
template <bool COUNT, bool VI, int FUSE>
static void nl_launch(const ls1g_ctx* c, const NLArgs& a, size_t grid) {
	const size_t sm = (size_t)3 * a.tcap * 8;
	if (c->list_R == 4) {
		if (c->list_zchk) nl_force_kernel<COUNT, VI, FUSE, true, 4><<<(unsigned)grid, NLGeom<4>::BT, sm, c->stream>>>(a);
		else nl_force_kernel<COUNT, VI, FUSE, false, 4><<<(unsigned)grid, NLGeom<4>::BT, sm, c->stream>>>(a);
	} else {
		if (c->list_zchk) nl_force_kernel<COUNT, VI, FUSE, true, 1><<<(unsigned)grid, NLGeom<1>::BT, sm, c->stream>>>(a);
		else nl_force_kernel<COUNT, VI, FUSE, false, 1><<<(unsigned)grid, NLGeom<1>::BT, sm, c->stream>>>(a);
	}
}

// buffers of the list loop for nb blocks (grown, never shrunk)
static int nl_ensure_buffers(ls1g_ctx* c, size_t nb) {
	const size_t want_l = nb * (size_t)c->list_lcap * nl_bt(c), want_c = nb * nl_bt(c);
	if (want_l > c->lists_cap) {
		cudaFree(c->lists); c->lists = nullptr; c->lists_cap = 0;
		CUDA_TRY(cudaMalloc((void**)&c->lists, want_l * sizeof(unsigned short)));
		c->lists_cap = want_l;
	}
	if (want_c > c->lcount_cap) {
		cudaFree(c->lcount); c->lcount = nullptr; c->lcount_cap = 0;
		CUDA_TRY(cudaMalloc((void**)&c->lcount, want_c * sizeof(uint32_t)));
		c->lcount_cap = want_c;
	}
	if (!c->pos_build) CUDA_TRY(cudaMalloc((void**)&c->pos_build, c->cap * sizeof(double4)));
	if (nb > c->bmap_cap) {
		cudaFree(c->plans); c->plans = nullptr;
		CUDA_TRY(cudaMalloc((void**)&c->plans, nb * sizeof(NLPlan)));
		cudaFree(c->bmap); cudaFree(c->bflag); c->bmap = nullptr; c->bflag = nullptr; c->bmap_cap = 0;
		CUDA_TRY(cudaMalloc((void**)&c->bmap, nb * sizeof(uint32_t)));
		CUDA_TRY(cudaMalloc((void**)&c->bflag, nb));
		c->bmap_cap = nb;
	}
	return ls1g_ensure_partials(c, nb ? nb : 1, 8);
}

// Set-up of the list loop ahead of the first step (called by ls1g_reserve): the kernels' modules are loaded (setting their
// attributes does that under CUDA's lazy loading) and the buffers are allocated from an ESTIMATE of the block count - n molecules
// spread evenly over the bin grid, 30 % head room; a build that needs more grows them as before.  Without this the first step of a
// host-driven loop pays for a 2.6 GB cudaMalloc (10^7 molecules) and the module loads inside the caller's loop timer.
int ls1g_ensure_rec(ls1g_ctx* c, size_t n_steps);
__global__ void fused_finalize_kernel(const double* __restrict__ partials, size_t nb, double* __restrict__ out);
int k_list_reserve(ls1g_ctx* c, size_t n) {
	if (!nl_applicable(c) || c->cap == 0) return 0;
	if (nl_force_attrs(c)) return 1;
	{
		cudaFuncAttributes fa;      // kernels of the loop that carry no attribute of their own
		CUDA_TRY(cudaFuncGetAttributes(&fa, halo_refresh_kernel));
		CUDA_TRY(cudaFuncGetAttributes(&fa, fused_finalize_kernel));
	}
	Grid g;
	if (grid_init_cut(g, c->cfg, c->cfg.cutoff + c->skin, -1)) return 1;
	if (c->list_lcap < NL_LCAP0) c->list_lcap = NL_LCAP0;
	const size_t nx = g.cpd[0] - 2, nyi = g.cpd[1] - 2, nzi = g.cpd[2] - 2;
	const int R = (c->list_R == 1 || nyi < 2 || nzi < 2) ? 1 : 4;
	const double per_row = 1.3 * (double)n / (double)(nyi * nzi);
	size_t nb;
	if (R == 4) {
		const size_t bpr = std::max((size_t)(4. * per_row / NL_BUNDLE_FILL) + 1, (nx + NL_XSPAN - 1) / NL_XSPAN);
		nb = ((nyi + 1) / 2) * ((nzi + 1) / 2) * bpr;
	} else nb = nyi * nzi * ((size_t)(per_row / NL_BT) + 1);
	const int keepR = c->list_R;
	c->list_R = R;                                    // nl_bt() sizes by it
	const int rc = nl_ensure_buffers(c, nb);
	c->list_R = keepR;
	if (rc) return 1;
	return ls1g_ensure_rec(c, 1024);
}

// after k_sort on the bin grid: size the block grid, build the lists, remember the build positions.
// The caller synchronises and reads h_flags[1] (overflow bits).
int k_list_build(ls1g_ctx* c) {
	if (c->list_lcap < NL_LCAP0) c->list_lcap = NL_LCAP0;
	if (c->list_R != 1 && c->list_R != 4) c->list_R = 4;
	if (c->bgrid.cpd[1] - 2 < 2 || c->bgrid.cpd[2] - 2 < 2) c->list_R = 1;      // nothing to bundle
	// R = 1: 128 slots of a row per block; R = 4: every bundle is cut into ceil(molecules / NL_BUNDLE_FILL) blocks (lists.cuh: nl_plan)
	c->list_bpr = c->list_R == 4 ? std::max((int)((c->max_bundle + NL_BUNDLE_FILL - 1) / NL_BUNDLE_FILL), (c->bgrid.cpd[0] - 2 + NL_XSPAN - 1) / NL_XSPAN)
	                             : (int)((c->max_row + NL_BT - 1) / NL_BT);
	if (c->list_bpr < 1) c->list_bpr = 1;
	if (c->list_R == 4) c->build_tcap = NL_TCAP;         // 16-row tiles: the large build tile from the start
	else if (c->build_tcap < NL_TCAP0) c->build_tcap = NL_TCAP0;
	const size_t nb = nl_blocks(c);
	if (nl_ensure_buffers(c, nb)) return 1;
	timer_start(c, LS1G_T_SORT);
	CUDA_TRY(cudaMemsetAsync(c->d_flags, 0, 3 * sizeof(unsigned long long), c->stream));
	CUDA_TRY(cudaMemsetAsync(c->d_flags + 5, 0, 2 * sizeof(unsigned long long), c->stream));      // boundary-block count, near-pair flag
	CUDA_TRY(cudaMemcpyAsync(c->pos_build, c->cur.pos, c->n_slots * sizeof(double4), cudaMemcpyDeviceToDevice, c->stream));
	CUDA_TRY(cudaMemcpyAsync(c->alt.pos, c->cur.pos, c->n_slots * sizeof(double4), cudaMemcpyDeviceToDevice, c->stream));   // fused steps ping-pong cur.pos / alt.pos
	NLArgs a = nl_args(c);
	if (nl_force_attrs(c)) return 1;
	CUDA_TRY(cudaMemsetAsync(c->lcount, 0, nb * nl_bt(c) * sizeof(uint32_t), c->stream));      // lanes without a molecule: empty lists
	if (c->list_R == 4) nl_build_kernel<4><<<(unsigned)nb, NLGeom<4>::BT, NL_BUILD_SMEM(4, a.btcap), c->stream>>>(a);
	else nl_build_kernel<1><<<(unsigned)nb, NLGeom<1>::BT, NL_BUILD_SMEM(1, a.btcap), c->stream>>>(a);
	CUDA_TRY(cudaGetLastError());
	const bool permute = c->list_permute == 1 || (c->list_permute == 2 && c->list_life >= 32);
	if (permute) {      // bank-aware order of every list (lists.cuh: nl_permute_kernel)
		const size_t lanes = nb * nl_bt(c);
		const int bt = (int)nl_bt(c);
		if (c->list_lcap <= 128) nl_permute_kernel<128, 128><<<nblk(lanes, 128), 128, NL_PERM_SMEM(128, 128), c->stream>>>(c->lists, c->lcount, a.lcap4, bt, lanes);
		else if (c->list_lcap <= 256) nl_permute_kernel<256, 128><<<nblk(lanes, 128), 128, NL_PERM_SMEM(256, 128), c->stream>>>(c->lists, c->lcount, a.lcap4, bt, lanes);
		else if (c->list_lcap <= 512) nl_permute_kernel<512, 64><<<nblk(lanes, 64), 64, NL_PERM_SMEM(512, 64), c->stream>>>(c->lists, c->lcount, a.lcap4, bt, lanes);
		else nl_permute_kernel<1024, 32><<<nblk(lanes, 32), 32, NL_PERM_SMEM(1024, 32), c->stream>>>(c->lists, c->lcount, a.lcap4, bt, lanes);
		CUDA_TRY(cudaGetLastError());
	}
	CUDA_TRY(cudaMemcpyAsync(c->h_flags, c->d_flags, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
	CUDA_TRY(cudaMemcpyAsync(c->h_flags + 3, c->d_flags + 5, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));   // read by the caller after its sync
	CUDA_TRY(cudaMemcpyAsync(c->h_flags + 4, c->d_flags + 6, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));   // near-pair flag
	timer_stop(c, LS1G_T_SORT, 2 + (permute ? 1 : 0));
	return 0;
}

int k_forces_list(ls1g_ctx* c, int last_step) {
	const size_t nb = nl_blocks(c);
	if (ls1g_ensure_partials(c, nb ? nb : 1, 5)) return 1;
	if (nl_force_attrs(c)) return 1;
	NLArgs a = nl_args(c);
	a.tcap = c->list_tcap;      // the largest tile of the last build: shared memory sized to what this system needs
	timer_start(c, LS1G_T_FORCE);
	// last step of a run: pair statistics + per-molecule virial; other steps: forces + the scalar sums of the step record
	if (last_step || c->vi_every_step) nl_launch<true, true, 0>(c, a, nb);
	else if (c->shift6_h[0] != 0.) nl_launch<true, false, 0>(c, a, nb);
	else nl_launch<false, false, 0>(c, a, nb);
	timer_stop(c, LS1G_T_FORCE, 1);
	if (k_force_finalize(c, nb)) return 1;
	CUDA_TRY(cudaGetLastError());
	return 0;
}

// endTraversal + the kinetic sums of the fused step: partials[nb][8] = u6, -, virial, -, pairs, sum m v^2, N, -
// out[0..3] = upot, virial, pairs, 0 (as force_finalize_kernel);  out[4..7] = summv2, sumIw2 = 0, N, rotDOF = 0
__global__ void __launch_bounds__(256) fused_finalize_kernel(const double* __restrict__ partials, size_t nb, double* __restrict__ out) {
	__shared__ double red[8 * 32];
	double s[8];
#pragma unroll
	for (int k = 0; k < 8; ++k) s[k] = 0.;
	for (size_t b = threadIdx.x; b < nb; b += blockDim.x)
#pragma unroll
		for (int k = 0; k < 8; ++k) s[k] += partials[b * 8 + k];
	block_sum<8>(s, red);
	if (threadIdx.x == 0) {
		out[0] = 0.5 * s[0] / 6.0; out[1] = 0.5 * s[2]; out[2] = 0.5 * s[4]; out[3] = 0.;
		out[4] = s[5]; out[5] = 0.; out[6] = s[6]; out[7] = 0.;
	}
}

// One NVE step of the list loop in ONE pass: forces, post(n), pre(n+1) (nl_force_kernel<.., FUSE>).  Reads c->cur.pos, writes
// the new owned positions to c->alt.pos (free between sorts) and swaps the two pointers.
// phase: 0 = one launch; 1 = inner blocks only (no finalize, no swap); 2 = boundary blocks, then finalize + swap (see NLArgs::phase)
int k_forces_list_fused(ls1g_ctx* c, double dt, int phase) {
	const size_t nb = nl_blocks(c);
	if (ls1g_ensure_partials(c, nb ? nb : 1, 8)) return 1;
	if (nl_force_attrs(c)) return 1;
	NLArgs a = nl_args(c);
	a.tcap = c->list_tcap;
	a.dt = dt; a.dt_half = 0.5 * dt; a.mass = c->comps_h[0].mass;
	a.vel = c->cur.vel; a.pos_out = c->alt.pos; a.pos_build = c->pos_build;
	a.phase = phase;
	const size_t grid = phase == 2 ? (size_t)c->n_boundary_blocks : nb;
	timer_start(c, LS1G_T_FORCE);
	if (grid) {
		if (c->shift6_h[0] != 0.) nl_launch<true, false, 1>(c, a, grid);
		else nl_launch<false, false, 1>(c, a, grid);
	}
	timer_stop(c, LS1G_T_FORCE, grid ? 1 : 0);
	CUDA_TRY(cudaGetLastError());
	if (phase == 1) return 0;
	std::swap(c->cur.pos, c->alt.pos);
	timer_start(c, LS1G_T_AUX);
	size_t nbr = nb; const double* base;
	const int extra = nb > 2048;
	if (ls1g_prereduce(c, nbr, 8, base)) return 1;
	fused_finalize_kernel<<<1, 256, 0, c->stream>>>(base, nbr, c->d_scalars);
	timer_stop(c, LS1G_T_AUX, 1 + extra);
	CUDA_TRY(cudaGetLastError());
	return 0;
}

// forces + post(n) in one pass (nl_force_kernel<.., FUSE = 2>): the last step of a run / the single step of ls1g_step.
// vi: per-molecule virial + pair statistics.  Leaves the step scalars in d_scalars[0..7] like the fused NVE step.
int k_forces_list_post(ls1g_ctx* c, double dt, int vi) {
	const size_t nb = nl_blocks(c);
	if (ls1g_ensure_partials(c, nb ? nb : 1, 8)) return 1;
	if (nl_force_attrs(c)) return 1;
	NLArgs a = nl_args(c);
	a.tcap = c->list_tcap;
	a.dt = dt; a.dt_half = 0.5 * dt; a.mass = c->comps_h[0].mass;
	a.vel = c->cur.vel; a.pos_out = nullptr; a.pos_build = c->pos_build;
	timer_start(c, LS1G_T_FORCE);
	if (nb) {
		if (vi || c->vi_every_step) nl_launch<true, true, 2>(c, a, nb);
		else if (c->shift6_h[0] != 0.) nl_launch<true, false, 2>(c, a, nb);
		else nl_launch<false, false, 2>(c, a, nb);
	}
	timer_stop(c, LS1G_T_FORCE, nb ? 1 : 0);
	CUDA_TRY(cudaGetLastError());
	timer_start(c, LS1G_T_AUX);
	size_t nbr = nb; const double* base;
	const int extra = nb > 2048;
	if (ls1g_prereduce(c, nbr, 8, base)) return 1;
	fused_finalize_kernel<<<1, 256, 0, c->stream>>>(base, nbr, c->d_scalars);
	timer_stop(c, LS1G_T_AUX, 1 + extra);
	CUDA_TRY(cudaGetLastError());
	return 0;
}

int k_halo_refresh(ls1g_ctx* c) {
	if (c->n_slots == 0) return 0;
	timer_start(c, LS1G_T_SORT);
	halo_refresh_kernel<<<nblk(c->n_slots, 256), 256, 0, c->stream>>>(c->bgrid, c->n_slots, c->cur.pos, c->cur.cell, c->halo_src, c->halo_code);
	timer_stop(c, LS1G_T_SORT, 1);
	CUDA_TRY(cudaGetLastError());
	return 0;
}
