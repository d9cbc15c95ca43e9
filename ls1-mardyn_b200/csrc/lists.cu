// lists.cu — host side of the neighbour-list mode of ls1g_run (kernels in lists.cuh).
#include "ls1g_internal.cuh"
#include "potentials.cuh"
#include "pair_body.cuh"
#include "lists.cuh"

#include <math.h>
#include <stdlib.h>
#include <algorithm>
#include <string.h>

int ls1g_ensure_partials(ls1g_ctx* c, size_t blocks, int nv);
int k_force_finalize(ls1g_ctx* c, size_t nb);
int ls1g_prereduce(ls1g_ctx* c, size_t& nb, int nv, const double*& base);   // force.cu

static inline unsigned nblk(size_t n, unsigned t) { return (unsigned)((n + t - 1) / t); }

// list mode applies to: single-centre LJ and (LS1G_LIST_MULTISITE=0 switches that off) rigid multi-site systems, default kernels, skin > 0, not disabled by an earlier overflow; one rank or several with
// either transport (NCCL, in-process group)
int nl_applicable(const ls1g_ctx* c) {
	// a block shorter than the list radius in some dimension would get ONE bin of less than that radius (grid_init_cut, as
	// LinkedCells.cpp:56-58 does for its cells): partners two periodic images away would be missed, so such boxes stay on the per-step path
	const double rl = std::max(c->cfg.cutoff, c->cfg.lj_cutoff) + c->skin;
	for (int d = 0; d < 3; ++d) if (c->cfg.box_max[d] - c->cfg.box_min[d] < rl) return 0;
	return (c->simple_1clj || c->list_multisite) && (c->n_ranks == 1 || c->comm_mode == 1 || c->comm_mode == 2) && c->force_variant == 0 && c->skin > 0. && !c->list_disabled;
}

// switch the bin grid: cells >= cutoff + skin (list mode) or the reference grid
int nl_set_bin_grid(ls1g_ctx* c, int list_mode) {
	if (!list_mode) { c->bgrid = c->grid; return 0; }
	Grid g;
	if (grid_init_cut(g, c->cfg, std::max(c->cfg.cutoff, c->cfg.lj_cutoff) + c->skin, -1)) return 1;
	if (g.n_cells > c->grid.n_cells) LS1G_FAIL("neighbour-list grid larger than the reference grid");
	c->bgrid = g;
	return 0;
}

static size_t nl_blocks(const ls1g_ctx* c);
static NLArgs nl_args(ls1g_ctx* c) {
	NLArgs a;
	a.g = c->bgrid; a.pos = c->cur.pos; a.cell = c->cur.cell; a.cell_start = c->cell_start;
	a.lists = c->lists; a.lcount = c->lcount; a.plans = (NLPlan*)c->plans; a.flags = c->d_flags; a.bpr = c->list_bpr; a.nyi = c->bgrid.cpd[1] - 2; a.lcap4 = c->list_lcap / 4; a.btcap = c->build_tcap;
	a.nzi = c->bgrid.cpd[2] - 2; a.nyg = (a.nyi + 1) / 2;
	a.order = c->list_permute == 1;
	a.tcap = NL_TCAP;
	const double rl = std::max(c->cfg.lj_cutoff, c->cfg.cutoff) + c->skin;
	// FP32 filter: coordinates relative to the block centre are < 8 cell lengths: relative error on r^2 near rl^2 < 2e-5;
	// the 3e-4 margin can only over-accept (the force kernel applies the exact FP64 test)
	a.rl2f = (float)(rl * rl * (1.0 + 3e-4));
	a.eps24 = c->eps24_h[0]; a.sig2 = c->sig2_h[0]; a.shift6 = c->shift6_h[0]; a.rclj2 = c->rclj2;
	a.F = c->F; a.Vi = c->Vi; a.partials = c->partials;
	a.guard = nl_guard_ptr(c); a.guard_thr = nl_guard_thr(c);
	if (c->inner_valid) { a.lists = c->lists_in; a.lcount = c->lcount_in; }      // the force kernels read the pruned lists
	a.pos_prune = c->inner_valid ? c->pos_prune : nullptr;
	a.lists_out = nullptr; a.lcount_out = nullptr; a.pos_prune_out = nullptr; a.rin2 = 0.;
	a.bmap = c->bmap; a.bflag = c->bflag; a.phase = 0; a.nb_total = (unsigned)nl_blocks(c);
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
	CUDA_TRY(cudaFuncSetAttribute(nl_prune_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, NL_FORCE_SMEM_MAX));
	CUDA_TRY(cudaFuncSetAttribute(nl_prune_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, NL_FORCE_SMEM_MAX));
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

// inner lists: single-centre systems only (the multi-site kernel reads the outer lists), LS1G_SKIN_INNER=0 switches them off
bool nl_inner_wanted(const ls1g_ctx* c) { return c->simple_1clj && c->skin_in > 0. && c->skin_in < 0.5 * c->skin; }

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
	if (nl_inner_wanted(c)) {      // the pruned copy of the lists (nl_prune_kernel)
		if (want_l > c->lists_in_cap) {
			cudaFree(c->lists_in); c->lists_in = nullptr; c->lists_in_cap = 0;
			CUDA_TRY(cudaMalloc((void**)&c->lists_in, want_l * sizeof(unsigned short)));
			c->lists_in_cap = want_l;
		}
		if (want_c > c->lcount_in_cap) {
			cudaFree(c->lcount_in); c->lcount_in = nullptr; c->lcount_in_cap = 0;
			CUDA_TRY(cudaMalloc((void**)&c->lcount_in, want_c * sizeof(uint32_t)));
			c->lcount_in_cap = want_c;
		}
		if (!c->pos_prune) CUDA_TRY(cudaMalloc((void**)&c->pos_prune, c->cap * sizeof(double4)));
	}
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
	if (grid_init_cut(g, c->cfg, std::max(c->cfg.cutoff, c->cfg.lj_cutoff) + c->skin, -1)) return 1;
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
	c->inner_valid = 0;                                                                            // the build replaces the outer lists
	CUDA_TRY(cudaMemsetAsync(c->d_flags, 0, 3 * sizeof(unsigned long long), c->stream));           // displacement (build, prune), max tile
	CUDA_TRY(cudaMemsetAsync(c->d_flags + 4, 0, 4 * sizeof(unsigned long long), c->stream));       // inner- and boundary-block count, near-pair flag, overflow bits
	CUDA_TRY(cudaMemcpyAsync(c->pos_build, c->cur.pos, c->n_slots * sizeof(double4), cudaMemcpyDeviceToDevice, c->stream));
	CUDA_TRY(cudaMemcpyAsync(c->alt.pos, c->cur.pos, c->n_slots * sizeof(double4), cudaMemcpyDeviceToDevice, c->stream));   // fused steps ping-pong cur.pos / alt.pos
	NLArgs a = nl_args(c);
	if (nl_force_attrs(c)) return 1;
	CUDA_TRY(cudaMemsetAsync(c->lcount, 0, nb * nl_bt(c) * sizeof(uint32_t), c->stream));      // lanes without a molecule: empty lists
	if (c->list_R == 4) nl_build_kernel<4><<<(unsigned)nb, NLGeom<4>::BT, NL_BUILD_SMEM(4, a.btcap), c->stream>>>(a);
	else nl_build_kernel<1><<<(unsigned)nb, NLGeom<1>::BT, NL_BUILD_SMEM(1, a.btcap), c->stream>>>(a);
	CUDA_TRY(cudaGetLastError());
	const bool permute = c->list_permute == 2;      // the separate whole-list pass (measurement only; the build kernel orders its chunks itself)
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
	CUDA_TRY(cudaMemcpyAsync(c->h_flags + 1, c->d_flags + NL_F_OVF, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));   // overflow bits
	CUDA_TRY(cudaMemcpyAsync(c->h_flags + 3, c->d_flags + 5, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));   // read by the caller after its sync
	CUDA_TRY(cudaMemcpyAsync(c->h_flags + 4, c->d_flags + 6, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));   // near-pair flag
	timer_stop(c, LS1G_T_SORT, 2 + (permute ? 1 : 0));
	return 0;
}

// outer lists -> inner lists at the current positions (lists.cuh: nl_prune_kernel); from the next force pass on the kernels read the
// inner lists and are guarded by the displacement since THIS prune.  Halo positions must be current (after the refresh of the step).
// guarded: a prune the host enqueues ahead of its knowledge (it does nothing when the lists turn out expired, like the step it
// precedes); unguarded: right after a build, or after the host has seen the inner lists expire.
int k_list_prune(ls1g_ctx* c, int guarded) {
	const size_t nb = nl_blocks(c);
	if (nl_force_attrs(c)) return 1;
	c->inner_valid = 0;                     // nl_args: read the OUTER lists
	const int keep_guard = c->guard_on;
	c->guard_on = guarded;
	NLArgs a = nl_args(c);
	c->guard_on = keep_guard;
	a.tcap = c->list_tcap;
	a.lists_out = c->lists_in; a.lcount_out = c->lcount_in; a.pos_prune_out = c->pos_prune;
	const double rin = std::max(c->cfg.lj_cutoff, c->cfg.cutoff) + c->skin_in;
	a.rin2 = rin * rin;
	timer_start(c, LS1G_T_SORT);
	const size_t sm = (size_t)3 * a.tcap * 8;
	if (nb) {
		if (c->list_R == 4) nl_prune_kernel<4><<<(unsigned)nb, NLGeom<4>::BT, sm, c->stream>>>(a);
		else nl_prune_kernel<1><<<(unsigned)nb, NLGeom<1>::BT, sm, c->stream>>>(a);
	}
	nl_prune_commit_kernel<<<1, 1, 0, c->stream>>>(c->d_flags, a.guard, a.guard_thr);
	timer_stop(c, LS1G_T_SORT, 2);
	CUDA_TRY(cudaGetLastError());
	c->inner_valid = 1;
	++c->list_prunes;
	return 0;
}

// the guard word of the next step from the displacement flags (after the all-reduce of a multi-rank step)
int k_list_guard(ls1g_ctx* c) {
	nl_guard_kernel<<<1, 1, 0, c->stream>>>(c->d_flags, c->disp_thr2, c->thr_in2, c->inner_valid);
	CUDA_TRY(cudaGetLastError());
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
// phase: 0 = one launch; 1 = inner blocks only (no finalize, no swap); 2 = boundary blocks, then finalize + swap (see NLArgs::phase);
//        3 = boundary blocks only (launched on the communication stream, next to the inner blocks); 4 = finalize + swap only
int k_forces_list_fused(ls1g_ctx* c, double dt, int phase) {
	const size_t nb = nl_blocks(c);
	if (ls1g_ensure_partials(c, nb ? nb : 1, 8)) return 1;
	if (nl_force_attrs(c)) return 1;
	if (phase != 4) {
		NLArgs a = nl_args(c);
		a.tcap = c->list_tcap;
		a.dt = dt; a.dt_half = 0.5 * dt; a.mass = c->comps_h[0].mass;
		a.vel = c->cur.vel; a.pos_out = c->alt.pos; a.pos_build = c->pos_build;
		a.phase = phase == 3 ? 2 : phase;
		const size_t grid = a.phase == 2 ? (size_t)c->n_boundary_blocks : (a.phase == 1 ? nb - (size_t)c->n_boundary_blocks : nb);
		// force_span (phases 1 + 3 run side by side on two streams): ONE timer interval on the main stream, from the start of the
		// inner blocks to the completion of both kernels (phase 4) - the sum of two overlapping durations would count time twice
		if (!(c->force_span && phase == 3)) timer_start(c, LS1G_T_FORCE);
		if (grid) {
			if (c->shift6_h[0] != 0.) nl_launch<true, false, 1>(c, a, grid);
			else nl_launch<false, false, 1>(c, a, grid);
		}
		if (!c->force_span) timer_stop(c, LS1G_T_FORCE, grid ? 1 : 0);
		CUDA_TRY(cudaGetLastError());
	}
	if (phase == 1 || phase == 3) return 0;
	if (phase == 4 && c->force_span) timer_stop(c, LS1G_T_FORCE, 2);
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

// ---------------------------------------------------------------------------------------------------------------
// multi-site molecules on the same lists
// ---------------------------------------------------------------------------------------------------------------
// The list build works on molecule centres with the radius max(rc, rc_LJ) + skin and does not care what a molecule is made of, so
// rigid multi-site systems (VectorizedCellProcessor.cpp:1013-1024: both cutoffs are decided on the molecule centres) use the same
// frozen layout, plans and 16-bit lists.  The force kernel differs: the site loops of ONE molecule pair are hundreds of FP64
// instructions (CO2: 9 LJ + 9 charge pairs), so lanes work independently - no shared tile; a partner's world-frame site rows come
// from the site cache (L2), its slot from the stored plan (tile index -> row -> slot).  128 threads per block over the build's lane
// numbering (a build block of 128 or 512 lanes is a whole number of these blocks).  Same pair functions and calcFM epilogue as
// force_generic_kernel (pair_body.cuh); per-block partials [5] as there.
#define NLM_THREADS 128
#define NLM_QCAP 80        // queue entries per lane (partners inside the cutoff waiting for their site loops)
#define NLM_CHUNK4 4       // list quads scanned between two "queue nearly full?" votes
#ifndef NLM_MINB
#define NLM_MINB 3         // resident blocks per SM the register allocation aims at
#endif
struct NLMArgs {
	const unsigned short* lists; const uint32_t* lcount; const NLPlan* plans;
	int lcap4, BT, R;
	size_t n_lanes;
	const unsigned long long* guard; unsigned long long guard_thr;
};

template <int NLJ, int NCH, int NDI, int NQU>
__global__ void __launch_bounds__(NLM_THREADS, NLM_MINB) nl_force_ms_kernel(ForceArgs a, NLMArgs m) {
	constexpr bool FIXED = NLJ >= 0;
	constexpr int NS = FIXED ? NLJ + NCH + 2 * NDI + 2 * NQU : 1;
	__shared__ double red[5 * 32];
	__shared__ NLPlan P;
	__shared__ uint32_t queue[NLM_QCAP * NLM_THREADS];
	__shared__ double ljtab[FIXED ? 1 : 3 * LJ_TAB_MAX];      // mixtures: the LJ tables in shared memory
	if (m.guard && *m.guard > m.guard_thr) return;      // speculatively enqueued step, lists expired (see run_list_steps)
	LJTab lj; lj.eps24 = lj.sig2 = lj.shift6 = nullptr;
	if constexpr (!FIXED) lj = lj_tables(a, ljtab);
	const int tid = threadIdx.x;
	const size_t gl = (size_t)blockIdx.x * NLM_THREADS + tid;
	const size_t blk = ((size_t)blockIdx.x * NLM_THREADS) / (size_t)m.BT;      // the same for the whole block
	if (tid < (int)(sizeof(NLPlan) / 4)) ((uint32_t*)&P)[tid] = ((const uint32_t*)(m.plans + blk))[tid];
	__syncthreads();
	const int t = (int)(gl - blk * (size_t)m.BT);
	const int NR = m.R == 4 ? 16 : 9;
	double acc[5] = {0., 0., 0., 0., 0.};
	int q = 0;
	if (m.R == 4) q = (t >= P.o[1]) + (t >= P.o[2]) + (t >= P.o[3]);
	const bool owned = P.live && t < P.o[m.R];
	const size_t i = (size_t)P.s0[q] + (size_t)(t - P.o[q]);
	double4 Pi = make_double4(0., 0., 0., 0.);
	if (owned) Pi = a.pos[i];
	const V3 Ri = ld3(Pi);
	CompDev ci; ci.n_lj = ci.n_ch = ci.n_di = ci.n_qu = 0;
	if (owned) ci = a.comps[pos_cid(Pi)];
	const double4* si = a.cache + i * (size_t)a.stride;
	const int i_ch0 = ci.n_lj, i_di0 = ci.n_lj + ci.n_ch, i_qu0 = i_di0 + 2 * ci.n_di;
	double4 sir[NS];
	if (FIXED && owned) {
#pragma unroll
		for (int k = 0; k < NS; ++k) sir[k] = si[k];
	}
	MolAcc A_; A_.F = v3(0, 0, 0); A_.M = v3(0, 0, 0); A_.V = v3(0, 0, 0); A_.u6 = 0.; A_.uX = 0.; A_.rf = 0.; A_.np = 0;
	// Two alternating phases, as in force_generic_kernel: every lane SCANS its list (slot of the entry, centre distance, both cutoffs:
	// ~30 % of the listed partners are in the skin) and queues the survivors in its shared-memory column; when some lane's queue is
	// nearly full all lanes DRAIN the same number of entries - the smallest backlog in the warp, so every lane is busy in the long site
	// loops; only the last round empties the queues.
	uint32_t* qq = queue + tid;
	int cnt = 0, o = 0;
	const int n4 = owned ? (int)(m.lcount[gl] & 0xffffu) : 0;
	const uint2* my = (const uint2*)(m.lists + ((size_t)blk * m.lcap4 * m.BT + t) * 4);
	for (;;) {
		for (;;) {
			const int oe = min(n4, o + NLM_CHUNK4);
			for (; o < oe; ++o) {
				const uint2 w = my[(size_t)o * m.BT];
				const uint32_t ee[4] = {w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16};
#pragma unroll
				for (int u = 0; u < 4; ++u) {
					const int k = (int)(ee[u] >> 3) - NL_DUMMY;         // tile index; pad entries give k < 0
					int r = 0;
					while (r + 1 < NR && k >= P.off[r + 1]) ++r;
					const uint32_t jj = P.beg[r] + (uint32_t)(k - P.off[r]);
					bool take = false;
					if (ee[u]) {
						const double4 Pj = a.pos[jj];
						const double mx = Ri.x - Pj.x, my_ = Ri.y - Pj.y, mz = Ri.z - Pj.z;
						const double m2 = mx * mx + my_ * my_ + mz * mz;
						// coincident centres lie in the same cell: the reference's SingleCellPolicy_ mask skips them (Helpers.h:367-376)
						take = (m2 < a.rclj2 || m2 < a.rc2) && m2 != 0.0;
					}
					if (take) { qq[cnt * NLM_THREADS] = jj; ++cnt; }
				}
			}
			const bool any_full = __any_sync(0xffffffffu, cnt > NLM_QCAP - 4 * NLM_CHUNK4);
			const bool any_more = __any_sync(0xffffffffu, o < n4);
			if (any_full || !any_more) break;
		}
		const bool last_round = !__any_sync(0xffffffffu, o < n4);
		int D = NLM_QCAP;
		if (!last_round) {
			const int mn = __reduce_min_sync(0xffffffffu, owned && (o < n4 || cnt > 0) ? cnt : NLM_QCAP);
			const int mx = __reduce_max_sync(0xffffffffu, cnt);
			D = max(mn, mx - (NLM_QCAP - 4 * NLM_CHUNK4));
		}
		const int take_n = min(cnt, D);
		for (int tt = 0; tt < take_n; ++tt) {
			const uint32_t jj = qq[(cnt - 1 - tt) * NLM_THREADS];
			const double4 Pj = a.pos[jj];
			const V3 mm = v3(Ri.x - Pj.x, Ri.y - Pj.y, Ri.z - Pj.z);
			const double m2 = dot(mm, mm);
			++A_.np;
			if constexpr (FIXED) {
				double4 sjr[NS];
				const double4* sj = a.cache + (size_t)jj * a.stride;
#pragma unroll
				for (int kk = 0; kk < NS; ++kk) sjr[kk] = sj[kk];
				pair_body_fixed<NLJ, NCH, NDI, NQU>(a, Ri, sir, sjr, mm, m2 < a.rclj2, m2 < a.rc2, A_);
			} else {
				pair_body(a, lj, Ri, ci, si, i_ch0, i_di0, i_qu0, jj, Pj, mm, m2 < a.rclj2, m2 < a.rc2, A_);
			}
		}
		cnt -= take_n;
		if (last_round) break;
	}
	if (owned) {
		a.F[3 * i] = A_.F.x; a.F[3 * i + 1] = A_.F.y; a.F[3 * i + 2] = A_.F.z;
		a.M[3 * i] = A_.M.x; a.M[3 * i + 1] = A_.M.y; a.M[3 * i + 2] = A_.M.z;
		a.Vi[3 * i] = 0.5 * A_.V.x; a.Vi[3 * i + 1] = 0.5 * A_.V.y; a.Vi[3 * i + 2] = 0.5 * A_.V.z;
		acc[0] = A_.u6; acc[1] = A_.uX; acc[2] = A_.V.x + A_.V.y + A_.V.z; acc[3] = A_.rf; acc[4] = (double)A_.np;
	}
	block_sum<5>(acc, red);
	if (tid == 0) {
#pragma unroll
		for (int k = 0; k < 5; ++k) a.partials[(size_t)blockIdx.x * 5 + k] = acc[k];
	}
}

int k_site_cache(ls1g_ctx* c);      // force.cu
// list force pass of a multi-site system: site cache of all slots, then the list kernel; leaves upot, virial, pairs in d_scalars[0..3]
int k_forces_list_ms(ls1g_ctx* c) {
	const size_t nb = nl_blocks(c);
	const size_t lanes = nb * nl_bt(c);
	const unsigned grid = nblk(lanes, NLM_THREADS);
	if (ls1g_ensure_partials(c, grid ? grid : 1, 5)) return 1;
	if (k_site_cache(c)) return 1;
	ForceArgs a;
	a.g = c->bgrid; a.n = c->n_slots; a.pos = c->cur.pos; a.cell = c->cur.cell; a.cell_start = c->cell_start;
	a.comps = c->comps; a.cache = c->site_cache; a.stride = c->cache_stride;
	a.eps24 = c->eps24; a.sig2 = c->sig2; a.shift6 = c->shift6; a.S = c->S_lj;
	a.rc2 = c->rc2; a.rclj2 = c->rclj2; a.epsRFInvrc3 = c->epsRFInvrc3;
	a.F = c->F; a.M = c->M; a.Vi = c->Vi; a.partials = c->partials;
	a.bpr = c->list_bpr; a.nyi = c->bgrid.cpd[1] - 2;
	NLMArgs m;
	m.lists = c->lists; m.lcount = c->lcount; m.plans = (const NLPlan*)c->plans; m.lcap4 = c->list_lcap / 4; m.BT = (int)nl_bt(c); m.R = c->list_R;
	m.n_lanes = lanes;
	m.guard = nl_guard_ptr(c); m.guard_thr = nl_guard_thr(c);
	const CompDev& c0 = c->comps_h[0];
	const int key = c->n_comp == 1 && c->S_lj <= 4 ? c0.n_lj * 1000 + c0.n_ch * 100 + c0.n_di * 10 + c0.n_qu : -1;
	if (key >= 0)
		for (int k = 0; k < c->S_lj * c->S_lj; ++k) { a.eps24_c[k] = c->eps24_h[k]; a.sig2_c[k] = c->sig2_h[k]; a.shift6_c[k] = c->shift6_h[k]; }
	timer_start(c, LS1G_T_FORCE);
	if (grid) {
		switch (key) {
#define LS1G_FIXED_CASE(L, Q, D, U) case L * 1000 + Q * 100 + D * 10 + U: nl_force_ms_kernel<L, Q, D, U><<<grid, NLM_THREADS, 0, c->stream>>>(a, m); break;
		LS1G_FIXED_CASE(3, 3, 0, 0)      // CO2 (3 LJ + 3 charges), examples/CO2
		LS1G_FIXED_CASE(3, 0, 1, 0)      // ethylene oxide (3 LJ + dipole), examples/EOX
		LS1G_FIXED_CASE(3, 0, 0, 1)      // CO2 Merker (3 LJ + quadrupole)
		LS1G_FIXED_CASE(2, 0, 0, 1)      // 2CLJQ
		LS1G_FIXED_CASE(2, 0, 1, 0)      // 2CLJD
		LS1G_FIXED_CASE(1, 0, 1, 0)      // Stockmayer
		LS1G_FIXED_CASE(2, 0, 0, 0)      // 2CLJ
		LS1G_FIXED_CASE(4, 0, 0, 0)      // 4 LJ sites
#undef LS1G_FIXED_CASE
		default: nl_force_ms_kernel<-1, 0, 0, 0><<<grid, NLM_THREADS, 0, c->stream>>>(a, m);
		}
	}
	timer_stop(c, LS1G_T_FORCE, grid ? 1 : 0);
	CUDA_TRY(cudaGetLastError());
	return k_force_finalize(c, grid);
}

int k_halo_refresh(ls1g_ctx* c) {
	if (c->n_slots == 0) return 0;
	timer_start(c, LS1G_T_SORT);
	halo_refresh_kernel<<<nblk(c->n_slots, 256), 256, 0, c->stream>>>(c->bgrid, c->n_slots, c->cur.pos, c->has_rot ? c->cur.q : nullptr, c->cur.cell, c->halo_src, c->halo_code);
	timer_stop(c, LS1G_T_SORT, 1);
	CUDA_TRY(cudaGetLastError());
	return 0;
}
