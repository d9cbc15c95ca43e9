// step.cu — per-step operators of the C ABI and the device-resident time loop (Simulation::simulate order,
// src/Simulation.cpp:979-1177).
#include "ls1g_internal.cuh"

#include <math.h>
#include <string.h>
#include <stdio.h>
#include <time.h>
#include <stdlib.h>

int k_post_pre(ls1g_ctx* c, double dt, int do_post, int do_pre, int thermostat, double target_T);
int k_post_sums(ls1g_ctx* c, double dt);                                       // integrate.cu: post kick + local kinetic sums
int k_betas_scale_pre(ls1g_ctx* c, double dt, int do_pre, double target_T);   // integrate.cu: betas from the (global) sums, scale (+ pre)
void ls1g_resolve_timers(ls1g_ctx* c);

#define NEED(c) do { if (!(c)) LS1G_FAIL("null context"); CUDA_TRY(cudaSetDevice((c)->cfg.device)); } while (0)
// per-step operators and transfers work on the reference-grid state: leave the neighbour-list layout first
#define NEED_REF(c) do { NEED(c); if ((c)->in_list_mode && ls1g_nl_leave(c)) return 1; CUDA_TRY(cudaSetDevice((c)->cfg.device)); } while (0)

int nl_set_bin_grid(ls1g_ctx* c, int list_mode);
static int sync_check(ls1g_ctx* c, const char* what);
static inline void set_dev(const ls1g_ctx* c) { cudaSetDevice(c->cfg.device); }
// CUDA call inside the step loops: 0 on success, else 1 with the error recorded (ls1g_last_error)
static int cu_fail(cudaError_t e, const char* what, int line) {
	if (e == cudaSuccess) return 0;
	ls1g_set_error(std::string("ls1g step loop: ") + what + " (step.cu:" + std::to_string(line) + "): " + cudaGetErrorString(e));
	return 1;
}
#define CU(expr) cu_fail((expr), #expr, __LINE__)
#define FOR_G(G, c) for (int _r = 0; _r < (G).n; ++_r) if (ls1g_ctx* c = (set_dev((G).c[_r]), (G).c[_r]))

// the group a context is driven with: itself, or - for a member of an in-process group - all ranks of that group
static Group group_of(ls1g_ctx*& self) {
	if (self->comm_mode == 2 && !self->peers.empty()) return Group{self->peers.data(), (int)self->peers.size()};
	return Group{&self, 1};
}

// ls1g_run leaves the container in the frozen neighbour-list layout so that the next ls1g_run continues without a
// rebuild.  Everything else first restores what the per-step loop leaves behind: wrapped positions, molecules sorted on
// the reference grid, halo copies dropped (one exchange; collective on multi-rank contexts, like the operators themselves).
int ls1g_nl_leave_group(const Group& G) {
	bool any = false;
	for (int r = 0; r < G.n; ++r) any = any || G.c[r]->in_list_mode;
	if (!any) return 0;
	FOR_G(G, c) {
		c->in_list_mode = 0; c->track_disp = 0;
		if (nl_set_bin_grid(c, 0)) return 1;
		c->sorted = 0;
	}
	if (G.multi() ? comm_exchange(G) : k_sort(G.c[0], 1)) return 1;
	FOR_G(G, c) { if (k_drop_halo(c)) return 1; if (sync_check(c, "leaving the neighbour-list layout")) return 1; }
	return 0;
}
int ls1g_nl_leave(ls1g_ctx* c) {
	if (!c->in_list_mode) return 0;
	return ls1g_nl_leave_group(group_of(c));
}

static int sync_check(ls1g_ctx* c, const char* what) {
	cudaError_t e = cudaStreamSynchronize(c->stream);
	if (e == cudaSuccess) e = cudaGetLastError();
	if (e != cudaSuccess) LS1G_FAIL(std::string(what) + ": " + cudaGetErrorString(e));
	return 0;
}

// LinkedCells::update (LinkedCells.cpp:243-356): re-bin every owned molecule; no wrap, no halo copies
extern "C" int ls1g_update(ls1g_ctx* c) {
	NEED_REF(c);
	if (k_sort(c, 0)) return 1;
	return sync_check(c, "ls1g_update");
}

// DomainDecompBase::exchangeMolecules (DomainDecompBase.cpp:51-86) / DomainDecomposition::balanceAndExchange
extern "C" int ls1g_exchange(ls1g_ctx* c) {
	NEED_REF(c);
	if (c->n_ranks > 1) {
		if (c->comm_mode == 2) LS1G_FAIL("ls1g_exchange: the ranks of a local group exchange together (ls1g_exchange_group)");
		Group G{&c, 1};
		if (comm_exchange(G)) return 1;
	} else if (k_sort(c, 1)) return 1;
	return sync_check(c, "ls1g_exchange");
}

extern "C" int ls1g_traverse(ls1g_ctx* c, double* upot, double* virial) {
	NEED_REF(c);
	if (k_forces(c)) return 1;
	CUDA_TRY(cudaMemcpyAsync(c->h_scalars, c->d_scalars, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
	if (sync_check(c, "ls1g_traverse")) return 1;
	c->forces_valid = 1;
	c->last_pairs = (uint64_t)llround(c->h_scalars[2]);
	if (upot) *upot = c->h_scalars[0];
	if (virial) *virial = c->h_scalars[1];
	return 0;
}

extern "C" int ls1g_delete_outer(ls1g_ctx* c) {
	NEED_REF(c);
	return k_drop_halo(c);
}

extern "C" int ls1g_integrate_pre(ls1g_ctx* c, double dt) {
	NEED_REF(c);
	if (k_post_pre(c, dt, 0, 1, 0, 0.)) return 1;
	c->sorted = 0;   // positions moved: cell_start is stale until the next update/exchange
	return sync_check(c, "ls1g_integrate_pre");
}

extern "C" int ls1g_integrate_post(ls1g_ctx* c, double dt_half, double* summv2, double* sumIw2, uint64_t* n, uint64_t* rot_dof) {
	NEED_REF(c);
	if (k_post_pre(c, 2. * dt_half, 1, 0, 0, 0.)) return 1;
	CUDA_TRY(cudaMemcpyAsync(c->h_scalars + 4, c->d_scalars + 4, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
	if (sync_check(c, "ls1g_integrate_post")) return 1;
	if (summv2) *summv2 = c->h_scalars[4];
	if (sumIw2) *sumIw2 = c->h_scalars[5];
	if (n) *n = (uint64_t)llround(c->h_scalars[6]);
	if (rot_dof) *rot_dof = (uint64_t)llround(c->h_scalars[7]);
	return 0;
}

extern "C" int ls1g_scale(ls1g_ctx* c, double bt, double br) {
	NEED(c);      // velocities only: valid in the frozen neighbour-list layout too (no ls1g_nl_leave)
	if (k_scale(c, bt, br)) return 1;
	return sync_check(c, "ls1g_scale");
}

// Domain::calculateGlobalValues explosion heuristics (Domain.cpp:254-295)
int k_selective(ls1g_ctx* c, double limit_energy, double Ti);
extern "C" int ls1g_selective_thermostat(ls1g_ctx* c, double limit_trans_energy, double temperature) {
	NEED(c);      // v, F, D, M only: valid in the frozen neighbour-list layout too
	if (!c->forces_valid) LS1G_FAIL("ls1g_selective_thermostat: forces missing (F and M are scaled together with v and D)");
	if (k_selective(c, limit_trans_energy, temperature)) return 1;
	return sync_check(c, "ls1g_selective_thermostat");
}

int nl_applicable(const ls1g_ctx* c);
int k_list_build(ls1g_ctx* c);
int k_forces_list(ls1g_ctx* c, int count_pairs);
int k_forces_list_ms(ls1g_ctx* c);               // lists.cu: rigid multi-site molecules on the same lists
int k_forces_list_fused(ls1g_ctx* c, double dt, int phase);
int k_forces_list_post(ls1g_ctx* c, double dt, int vi);
int k_halo_refresh(ls1g_ctx* c);

// [post(step n)] [+ thermostat] [+ pre(step n+1)] on every rank.  With a thermostat on several ranks the betas come from the GLOBAL
// kinetic sums (Domain::calculateGlobalValues, Domain.cpp:170-176,226-247): local sums, group all-reduce, then scale (+ pre).
static int g_post_pre(const Group& G, double dt, int do_post, int do_pre, int thermo, double target_T) {
	if (!(thermo && do_post && G.multi())) {
		FOR_G(G, c) if (k_post_pre(c, dt, do_post, do_pre, thermo, target_T)) return 1;
		return 0;
	}
	// d_scalars[4..7] keep this rank's sums (the step record); the global ones are formed in d_scalars[12..15]
	FOR_G(G, c) {
		if (k_post_sums(c, dt)) return 1;
		CUDA_TRY(cudaMemcpyAsync(c->d_scalars + 12, c->d_scalars + 4, 4 * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
	}
	if (comm_allreduce(G, COMM_SCALARS, 12, 4, COMM_SUM)) return 1;
	FOR_G(G, c) if (k_betas_scale_pre(c, dt, do_pre, target_T)) return 1;
	return 0;
}

// Neighbour-list steps of ls1g_run (lists.cuh).  On entry the pre-kick of step 0 is done.  Runs steps [0, n_steps) unless
// a buffer limit is hit while building lists for step s: then the bin grid is restored and *done = s tells the caller to
// continue with the rebuild-every-step loop (the state is exactly what that loop expects).
// Steps are enqueued ahead of the host's knowledge: the displacement that decides whether step s may still use the
// lists is produced by the integrate kernel of step s-1, and waiting for it before every launch would leave the device
// idle for a host round trip per step.  So step s is enqueued GUARDED: its force and integrate kernels first read a
// snapshot of the displacement flag (d_flags[3]) and return at once when the lists have expired.  The host reads the
// same value up to NL_AHEAD steps late (pinned copy + event), rebuilds, and enqueues the skipped steps again.  Device and host evaluate the
// identical comparison d2 > disp_thr2 on the identical bits; the flag is monotone between builds (atomicMax), so an
// expired step stays skipped however many kernels follow it.  Several ranks (NCCL: this rank and its peers in other processes;
// local group: all contexts of G) take every decision on the all-reduced maximum, so they rebuild at the same step.
#define NL_OVF_LIST_BIT 2ull      // lists.cuh: NL_OVF_LIST
#define NL_OVF_TILE_BIT 1ull      // lists.cuh: NL_OVF_TILE
#define NL_TCAP_MAX_H 5120        // lists.cuh: NL_TCAP
#define NL_LCAP_MAX_H 1020        // lists.cuh: NL_LCAP_MAX
#define NL_AHEAD 8      // steps the host may enqueue beyond the last displacement flag it has read (absorbs ~15 ms of host jitter)
#define NL_RING 16      // pinned flag copies / events in flight (> NL_AHEAD)
int k_list_guard(ls1g_ctx* c);      // lists.cu
static int flag_snapshot(const Group& G, int ring) {
	// displacement since the build and since the last prune: non-negative doubles, bit order == value order
	if (G.multi() && comm_allreduce(G, COMM_FLAGS, 0, 2, COMM_MAX)) return 1;
	FOR_G(G, c) {
		if (k_list_guard(c)) return 1;       // d_flags[3]: may the next step run?
		CUDA_TRY(cudaMemcpyAsync(c->h_flags + 8 + ring, c->d_flags, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
		CUDA_TRY(cudaMemcpyAsync(c->h_flags + 24 + ring, c->d_flags + 1, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
		CUDA_TRY(cudaEventRecord(c->disp_ev[ring], c->stream));
	}
	return 0;
}

static int run_list_steps(const Group& G, int n_steps, double dt, int thermo, double target_T, int* done, int want_vi = 1) {
	*done = 0;
	bool need_build = true;
	int rc = 0, s = 0;
	ls1g_ctx* c0 = G.c[0];
	const bool inner = nl_inner_wanted(c0);       // the force kernels read pruned inner lists (lists.cuh: nl_prune_kernel)
	for (int r = 0; r < G.n; ++r) {
		// the lists stay exact while no molecule moved more than skin/2; with inner lists the outer ones must also still be prunable,
		// i.e. (skin - skin_in)/2, and the inner ones are exact within skin_in/2 of the prune
		const double so = inner ? c0->skin - c0->skin_in : c0->skin;
		G.c[r]->disp_thr2 = (0.4995 * so) * (0.4995 * so);
		G.c[r]->thr_in2 = (0.4995 * c0->skin_in) * (0.4995 * c0->skin_in);
		G.c[r]->guard_on = 0;
	}
	const bool dbg = getenv("LS1G_DEBUG") != nullptr;
	// the integrator is fused into the force kernel on NVE steps (LS1G_FUSE=0 keeps the separate integrate kernel)
	const char* fe = getenv("LS1G_FUSE");
	const bool ms = !c0->simple_1clj;        // rigid multi-site molecules: list force kernel + the separate (rotational) integrate kernel
	const bool fuse_nve = !ms && !thermo && !c0->vi_every_step && !(fe && atoi(fe) == 0);
	const bool post_fuse = !ms && !thermo && c0->n_ranks == 1 && !(fe && atoi(fe) == 0);      // single rank: the kinetic sums need no all-reduce
	// several ranks: overlap the halo refresh with the inner blocks of the force pass (LS1G_OVERLAP=0: exchange first, then one launch)
	// (LS1G_OVERLAP=1: the boundary blocks wait on the main stream; default: they run on the communication stream right behind the
	// refresh, next to the inner blocks - one kernel tail and one stream hand-over less per step)
	const char* oe = getenv("LS1G_OVERLAP");
	const bool overlap = G.multi() && !(oe && atoi(oe) == 0);
	const bool overlap_concurrent = overlap && !(oe && atoi(oe) == 1);
	// what the flags of a finished step say about the next one (all ranks hold the same values): 0 = go, 1 = the inner lists expired
	// (prune again), 2 = the outer lists expired (build).  The same comparisons as nl_guard_kernel, on the same bits.
	auto expired = [&](int ring) {
		double dO, dI;
		memcpy(&dO, c0->h_flags + 8 + ring, sizeof(double)); memcpy(&dI, c0->h_flags + 24 + ring, sizeof(double));
		if (dO > c0->disp_thr2) return 2;
		return (c0->inner_valid && dI > c0->thr_in2) ? 1 : 0;
	};
	bool need_prune = false;
	auto fail = [&]() { FOR_G(G, c) { c->track_disp = 0; c->in_list_mode = 0; c->guard_on = 0; nl_set_bin_grid(c, 0); cudaStreamSynchronize(c->stream); cudaStreamSynchronize(c->stream_comm); } return 1; };
	bool cont = true;
	for (int r = 0; r < G.n; ++r) cont = cont && G.c[r]->in_list_mode;
	if (cont) {       // continuing the previous ls1g_run: the pre-kick above was tracked; rebuild only if the lists expired
		if (flag_snapshot(G, 0)) return fail();
		FOR_G(G, c) if (sync_check(c, "ls1g_run (displacement)")) return fail();
		const int st = expired(0);
		need_build = st == 2; need_prune = st == 1;
	} else {
		FOR_G(G, c) if (nl_set_bin_grid(c, 1)) return fail();
	}
	for (int r = 0; r < G.n; ++r) G.c[r]->in_list_mode = 1;
	int rebuilds = 0, repeats = 0, regrow = 0, prunes = 0;
	double t_build = 0., t_wait = 0., t_enq = 0., t_max_wait = 0., t_max_enq = 0.;     // host wall clock, LS1G_DEBUG only
	auto now = [&]() { if (!dbg) return 0.; struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; };
	auto rec_copy = [&](ls1g_ctx* c, int step, int off, int n) {
		if (c->d_rec && CU(cudaMemcpyAsync(c->d_rec + 8 * (size_t)step + off, c->d_scalars + off, n * sizeof(double), cudaMemcpyDeviceToDevice, c->stream))) return 1;
		return 0;
	};
	// s   = next step to enqueue
	// chk = next step whose displacement flag (written by its integrate kernel, deciding step chk + 1) the host has not read
	int chk = 0;
	while (!rc) {
		if (s < n_steps && s - chk < NL_AHEAD) {
			const bool built = need_build;
			bool overlapped = false;
			const bool prune_now = !need_build && inner && c0->inner_valid && (need_prune || (c0->inner_life > 1 && c0->since_prune >= c0->inner_life - 1));
			const double t0 = now();
			if (need_build) {
				// lifetime of the lists being replaced: steps [0, s) of this call have run with them (plus those of earlier calls)
				FOR_G(G, c) { const uint64_t now_steps = c->list_steps + (uint64_t)s; c->list_life = c->list_steps_at_build ? now_steps - c->list_steps_at_build : 0; c->list_steps_at_build = now_steps ? now_steps : 1; }
				FOR_G(G, c) { c->sorted = 0; c->track_disp = 0; c->guard_on = 0; c->forces_valid = 0; }
				rc = G.multi() ? comm_exchange_build(G) : k_sort(c0, 1);
				FOR_G(G, c) {
					c->forces_valid = 1;
					if (!rc) rc = k_halo_map(c);
					if (!rc) rc = k_list_build(c);             // zeroes d_flags[0..2]
					if (!rc && CU(cudaMemsetAsync(c->d_flags + 3, 0, sizeof(unsigned long long), c->stream))) rc = 1;
				}
				FOR_G(G, c) if (!rc && sync_check(c, "ls1g_run (list build)")) rc = 1;
				if (rc) break;
				// every rank must take the same decision: [hard overflow, list overflow, tile overflow] over all ranks
				unsigned long long hb = 0;
				for (int r = 0; r < G.n; ++r) hb |= G.c[r]->h_flags[1];
				if (c0->comm_mode == 1) {
					double ovf[3] = {(hb & ~(NL_OVF_LIST_BIT | NL_OVF_TILE_BIT)) ? 1. : 0., (hb & NL_OVF_LIST_BIT) ? 1. : 0., (hb & NL_OVF_TILE_BIT) ? 1. : 0.};
					if (CU(cudaMemcpyAsync(c0->d_scalars + 12, ovf, 3 * sizeof(double), cudaMemcpyHostToDevice, c0->stream))) rc = 1;
					if (!rc) rc = comm_allreduce(G, COMM_SCALARS, 12, 3, COMM_MAX);
					if (!rc && CU(cudaMemcpyAsync(ovf, c0->d_scalars + 12, 3 * sizeof(double), cudaMemcpyDeviceToHost, c0->stream))) rc = 1;
					if (!rc && sync_check(c0, "ls1g_run (list build)")) rc = 1;
					if (rc) break;
					if (ovf[0] != 0.) hb |= 8;
					if (ovf[1] != 0.) hb |= NL_OVF_LIST_BIT;
					if (ovf[2] != 0.) hb |= NL_OVF_TILE_BIT;
				}
				{
					// more neighbours per molecule than the lists hold / a denser neighbourhood than the build tile holds: raise
					// the capacity and build again (the sort is repeated too - it is idempotent on sorted data)
					bool grown = false;
					int lcap = c0->list_lcap, tcap = c0->build_tcap, R = c0->list_R;
					if ((hb & NL_OVF_LIST_BIT) && lcap < NL_LCAP_MAX_H) { lcap = lcap * 2 > NL_LCAP_MAX_H ? NL_LCAP_MAX_H : lcap * 2; grown = true; }
					if ((hb & NL_OVF_TILE_BIT) && R == 4) { R = 1; tcap = 0; grown = true; }      // rows that do not line up in x: single-row blocks
					else if ((hb & NL_OVF_TILE_BIT) && tcap < NL_TCAP_MAX_H) { tcap = tcap + 1024 > NL_TCAP_MAX_H ? NL_TCAP_MAX_H : tcap + 1024; grown = true; }
					for (int r = 0; r < G.n; ++r) { G.c[r]->list_lcap = lcap; G.c[r]->build_tcap = tcap; G.c[r]->list_R = R; }
					const bool fixable = !(hb & ~(NL_OVF_LIST_BIT | NL_OVF_TILE_BIT)) && grown;
					if (hb && fixable && ++regrow <= 6) {
						if (dbg) fprintf(stderr, "ls1g: list buffers raised to %d entries per molecule, build tile %d, rows per block %d (overflow bits %llu)\n", lcap, tcap, R, hb);
						continue;
					}
				}
				if (hb) {                 // tile / list / row capacity exceeded: this system is too dense for the list buffers
					FOR_G(G, c) { c->list_disabled = 1; c->in_list_mode = 0; c->track_disp = 0; nl_set_bin_grid(c, 0); }
					if (dbg) fprintf(stderr, "ls1g: neighbour lists disabled, overflow bits %llu (1 tile, 2 list, 4 row), bpr %d\n", hb, c0->list_bpr);
					*done = s;
					return 0;
				}
				FOR_G(G, c) {
					c->track_disp = 1;
					c->n_boundary_blocks = (uint32_t)c->h_flags[3];
					c->list_zchk = c->h_flags[4] != 0;
					c->list_tcap = (int)((c->h_flags[2] + 63) & ~63ull);
					if (c->list_tcap < 64) c->list_tcap = 64;
				}
				if (dbg) fprintf(stderr, "ls1g: list build at step %d: max tile %llu, rows per block %d, bpr %d, slots %zu, bin grid %d %d %d\n", s, c0->h_flags[2], c0->list_R, c0->list_bpr, c0->n_slots, c0->bgrid.cpd[0], c0->bgrid.cpd[1], c0->bgrid.cpd[2]);
				need_build = false; ++rebuilds;
				if (inner) {       // the first inner lists of these outer lists, at the build positions
					FOR_G(G, c) { if (!rc) rc = k_list_prune(c, 0); c->since_prune = 0; }
					FOR_G(G, c) if (!rc) rc = k_list_guard(c);
					need_prune = false;
				}
				t_build += now() - t0;
			} else if (overlap && fuse_nve && s + 1 < n_steps && !prune_now) {
				// several ranks, NVE step: the halo positions travel on the communication stream while the inner blocks run
				overlapped = true;
				FOR_G(G, c) {
					if (CU(cudaEventRecord(c->ev_pos, c->stream)) || CU(cudaStreamWaitEvent(c->stream_comm, c->ev_pos, 0))) rc = 1;
					std::swap(c->stream, c->stream_comm);
				}
				if (!rc) rc = comm_refresh(G);
				FOR_G(G, c) {
					if (!rc) rc = k_halo_refresh(c);
					if (overlap_concurrent) {      // boundary blocks: same stream, right behind the data they wait for
						c->guard_on = 1; c->force_span = 1;
						if (!rc) rc = k_forces_list_fused(c, dt, 3);
						c->guard_on = 0;
					}
					std::swap(c->stream, c->stream_comm);
					if (!rc && CU(cudaEventRecord(c->ev_halo, c->stream_comm))) rc = 1;
				}
			} else {
				if (G.multi()) rc = comm_refresh(G);
				FOR_G(G, c) if (!rc) rc = k_halo_refresh(c);
				if (prune_now) {
					// recovery (the host has seen the inner lists expire: the guard word says "expired", so the prune runs unguarded and
					// the word is recomputed) or proactive (one step before the inner lists are expected to expire: guarded like the step)
					const int guarded = need_prune ? 0 : 1;
					FOR_G(G, c) {
						c->guard_on = guarded;
						if (!rc) rc = k_list_prune(c, guarded);
						c->guard_on = 0;
						c->since_prune = 0;
						if (!rc && !guarded) rc = k_list_guard(c);
						if (guarded && ++c->inner_streak >= 16) { ++c->inner_life; c->inner_streak = 0; }      // no expiry for a while: try longer
					}
					need_prune = false; ++prunes;
				}
			}
			const double t1 = now();
			FOR_G(G, c) {
				c->guard_on = built ? 0 : 1;
				if (overlapped) {
					if (!rc) rc = k_forces_list_fused(c, dt, 1);
					if (!rc && CU(cudaStreamWaitEvent(c->stream, c->ev_halo, 0))) rc = 1;
					if (!rc) rc = k_forces_list_fused(c, dt, overlap_concurrent ? 4 : 2);
					c->force_span = 0;
					if (!rc) rc = rec_copy(c, s, 0, 8);
				} else if (fuse_nve && s + 1 < n_steps) {
					// NVE step that is not the last of the run: forces + post(s) + pre(s + 1) in one kernel
					if (!rc) rc = k_forces_list_fused(c, dt, 0);
					if (!rc) rc = rec_copy(c, s, 0, 8);
				} else if (post_fuse && s + 1 == n_steps) {
					// last step of the run (or the single step of ls1g_step): forces + post(s) in one kernel, no pre
					if (!rc) rc = k_forces_list_post(c, dt, want_vi);
					if (!rc) rc = rec_copy(c, s, 0, 8);
				} else {
					if (!rc) rc = ms ? k_forces_list_ms(c) : k_forces_list(c, s + 1 == n_steps && want_vi);      // 1CLJ: pair statistics + per-molecule virial only for the last step
					if (!rc) rc = rec_copy(c, s, 0, 4);
				}
			}
			if (!rc && !(overlapped || (fuse_nve && s + 1 < n_steps) || (post_fuse && s + 1 == n_steps))) {
				rc = g_post_pre(G, dt, 1, s + 1 < n_steps, thermo, target_T);
				FOR_G(G, c) if (!rc) rc = rec_copy(c, s, 4, 4);
			}
			FOR_G(G, c) { c->guard_on = 0; ++c->since_prune; }
			if (!rc && s + 1 < n_steps) rc = flag_snapshot(G, s % NL_RING);      // decides step s + 1
			const double t2 = now();
			t_enq += t2 - t1; if (t2 - t1 > t_max_enq) t_max_enq = t2 - t1;
			++s;
		} else if (chk < s && chk + 1 < n_steps) {
			// did the steps after chk really run?  step chk + 1 was guarded by the flag step chk wrote
			const double t2 = now();
			FOR_G(G, c) if (CU(cudaEventSynchronize(c->disp_ev[chk % NL_RING]))) rc = 1;
			if (rc) break;
			const double t3 = now();
			t_wait += t3 - t2; if (t3 - t2 > t_max_wait) t_max_wait = t3 - t2;
			const int st = expired(chk % NL_RING);
			if (st) {      // steps chk + 1 .. s - 1 were skipped on the device: rebuild (or prune again) and enqueue them again
				repeats += s - (chk + 1);
				FOR_G(G, c) c->since_prune -= s - (chk + 1);
				if (st == 1) {     // lifetime of the inner lists that just expired: the next ones are replaced one step earlier
					FOR_G(G, c) { c->inner_life = c->since_prune > 1 ? c->since_prune : 1; c->inner_streak = 0; }
					need_prune = true;
				} else need_build = true;
				s = chk + 1;
			}
			++chk;
		} else break;
	}
	FOR_G(G, c) c->guard_on = 0;
	FOR_G(G, c) if (!rc && sync_check(c, "ls1g_run (list steps)")) rc = 1;
	if (rc) return fail();
	if (dbg) fprintf(stderr, "ls1g: inner lists %s, %d prunes, expected lifetime %d steps\n", inner ? "on" : "off", prunes, c0->inner_life);
	if (dbg) fprintf(stderr, "ls1g: %d list steps, %d rebuilds, %d repeated steps; host ms: build %.2f, enqueue %.2f (max %.2f), wait %.2f (max %.2f)\n",
	                 n_steps, rebuilds, repeats, t_build, t_enq, t_max_enq, t_wait, t_max_wait);
	FOR_G(G, c) { c->list_rebuilds += (uint64_t)rebuilds; c->list_steps += (uint64_t)n_steps; }
	// stay in the frozen layout (in_list_mode): the next ls1g_run continues, any other entry point calls ls1g_nl_leave first
	FOR_G(G, c) if (k_drop_halo(c)) return 1;
	*done = n_steps;
	return 0;
}

// Device-resident loop.  Per step: pre -> exchange (wrap, images, sort) -> forces (+calcFM) -> post -> thermostat.
// post(n), the thermostat and pre(n+1) run as one kernel between two force passes.
// bt, br: betas applied to v, D before the first pre-kick (fused into it); want_vi: per-molecule virial + pair count on the last step.
// records: [n_steps], sums over the ranks of G (a NCCL rank reports its own part, as the reference's Domain::setLocal* values).
int k_scale_pre(ls1g_ctx* c, double dt, double bt, double br);
int ls1g_ensure_rec(ls1g_ctx* c, size_t n_steps);
static int run_impl(const Group& G, int n_steps, double dt, double target_T, ls1g_step_record* records, double bt, double br, int want_vi) {
	bool lists_ok = true;
	for (int r = 0; r < G.n; ++r) lists_ok = lists_ok && nl_applicable(G.c[r]);
	bool in_list = false;
	for (int r = 0; r < G.n; ++r) in_list = in_list || G.c[r]->in_list_mode;
	if (in_list && !lists_ok && ls1g_nl_leave_group(G)) return 1;
	for (int r = 0; r < G.n; ++r)
		if (!G.c[r]->forces_valid) LS1G_FAIL("ls1g_run: forces missing - call ls1g_exchange + ls1g_traverse once (prepare_start, Simulation.cpp:813-844)");
	const int thermo = target_T > 0.;
	int rc = 0;
	FOR_G(G, c) {
		if (!rc && records) rc = ls1g_ensure_rec(c, (size_t)n_steps);
		if (!records) { /* no records wanted: the kernels still reduce the step scalars, nothing is kept */ }
		if (!rc) rc = (bt != 1.0 || br != 1.0) ? k_scale_pre(c, dt, bt, br) : k_post_pre(c, dt, 0, 1, 0, 0.);
	}
	std::vector<double*> keep(G.n);
	for (int r = 0; r < G.n; ++r) { keep[r] = G.c[r]->d_rec; if (!records) G.c[r]->d_rec = nullptr; }      // rec_copy tests d_rec
	int s0 = 0;
	if (!rc && lists_ok) rc = run_list_steps(G, n_steps, dt, thermo, target_T, &s0, want_vi);
	for (int s = s0; s < n_steps && !rc; ++s) {
		FOR_G(G, c) { c->sorted = 0; c->forces_valid = 0; }     // F is recomputed before its next use: do not carry it through the sort
		rc = G.multi() ? comm_exchange(G) : k_sort(G.c[0], 1);
		FOR_G(G, c) {
			c->forces_valid = 1;
			if (!rc) rc = k_forces(c);
			if (!rc) rc = k_drop_halo(c);
			if (!rc && c->d_rec && CU(cudaMemcpyAsync(c->d_rec + 8 * (size_t)s, c->d_scalars, 4 * sizeof(double), cudaMemcpyDeviceToDevice, c->stream))) rc = 1;
		}
		if (!rc) rc = g_post_pre(G, dt, 1, s + 1 < n_steps, thermo, target_T);
		FOR_G(G, c)
			if (!rc && c->d_rec && CU(cudaMemcpyAsync(c->d_rec + 8 * (size_t)s + 4, c->d_scalars + 4, 4 * sizeof(double), cudaMemcpyDeviceToDevice, c->stream))) rc = 1;
	}
	std::vector<std::vector<double>> h(G.n);
	FOR_G(G, c) {
		if (!rc && c->d_rec) {
			h[_r].resize((size_t)n_steps * 8);
			if (CU(cudaMemcpyAsync(h[_r].data(), c->d_rec, h[_r].size() * sizeof(double), cudaMemcpyDeviceToHost, c->stream))) rc = 1;
		}
	}
	for (int r = 0; r < G.n; ++r) G.c[r]->d_rec = keep[r];
	if (rc) { FOR_G(G, c) cudaStreamSynchronize(c->stream); return 1; }
	FOR_G(G, c) if (sync_check(c, "ls1g_run")) return 1;
	if (records)
		for (int s = 0; s < n_steps; ++s) {
			double t[8] = {0., 0., 0., 0., 0., 0., 0., 0.};
			for (int r = 0; r < G.n; ++r) for (int k = 0; k < 8; ++k) t[k] += h[r][(size_t)s * 8 + k];
			records[s].upot = t[0]; records[s].virial = t[1]; records[s].summv2 = t[4]; records[s].sumIw2 = t[5];
			records[s].n = (uint64_t)llround(t[6]); records[s].rot_dof = (uint64_t)llround(t[7]);
			for (int r = 0; r < G.n; ++r) G.c[r]->last_pairs = (uint64_t)llround(h[r][(size_t)s * 8 + 2]);
		}
	return 0;
}

static int check_run_group(ls1g_ctx** ctxs, int n, const char* who) {
	if (!ctxs || n < 1) LS1G_FAIL(std::string(who) + ": bad arguments");
	for (int r = 0; r < n; ++r) if (!ctxs[r]) LS1G_FAIL(std::string(who) + ": null context");
	if (n == 1) return 0;
	for (int r = 0; r < n; ++r)
		if (ctxs[r]->comm_mode != 2 || (int)ctxs[r]->peers.size() != n || ctxs[r]->peers[r] != ctxs[r]) LS1G_FAIL(std::string(who) + ": call ls1g_comm_init_local with exactly these contexts first");
	return 0;
}

extern "C" int ls1g_run(ls1g_ctx* c, int n_steps, double dt, double target_T, ls1g_step_record* records) {
	NEED(c);
	if (n_steps <= 0) return 0;
	if (c->comm_mode == 2) LS1G_FAIL("ls1g_run: the ranks of a local group step together (ls1g_run_group)");
	Group G{&c, 1};
	return run_impl(G, n_steps, dt, target_T, records, 1.0, 1.0, 1);
}

// One time step for a host-driven loop: VelocityScalingThermostat::apply(beta_trans, beta_rot) of the PREVIOUS step fused into
// this step's pre-kick, then exchange / forces / deleteOuter / post as ls1g_run(1).  NVE inside: the betas are the caller's
// (Domain::calculateGlobalValues), so the reference's own thermostat logic stays in charge.
extern "C" int ls1g_step(ls1g_ctx* c, double dt, double beta_trans, double beta_rot, int flags, ls1g_step_record* record) {
	NEED(c);
	if (c->comm_mode == 2) LS1G_FAIL("ls1g_step: the ranks of a local group step together (ls1g_step_group)");
	Group G{&c, 1};
	return run_impl(G, 1, dt, -1., record, beta_trans, beta_rot, (flags & LS1G_STEP_VI) ? 1 : 0);
}

extern "C" int ls1g_run_group(ls1g_ctx** ctxs, int n, int n_steps, double dt, double target_T, ls1g_step_record* records) {
	if (check_run_group(ctxs, n, "ls1g_run_group")) return 1;
	if (n_steps <= 0) return 0;
	if (n == 1) return ls1g_run(ctxs[0], n_steps, dt, target_T, records);
	Group G{ctxs, n};
	return run_impl(G, n_steps, dt, target_T, records, 1.0, 1.0, 1);
}

extern "C" int ls1g_step_group(ls1g_ctx** ctxs, int n, double dt, double beta_trans, double beta_rot, int flags, ls1g_step_record* record) {
	if (check_run_group(ctxs, n, "ls1g_step_group")) return 1;
	if (n == 1) return ls1g_step(ctxs[0], dt, beta_trans, beta_rot, flags, record);
	Group G{ctxs, n};
	return run_impl(G, 1, dt, -1., record, beta_trans, beta_rot, (flags & LS1G_STEP_VI) ? 1 : 0);
}

int ls1g_ensure_rec(ls1g_ctx* c, size_t n_steps) {
	if (c->d_rec && c->d_rec_cap >= n_steps) return 0;
	cudaFree(c->d_rec); c->d_rec = nullptr; c->d_rec_cap = 0;
	const size_t cap = n_steps < 1024 ? 1024 : n_steps;
	CUDA_TRY(cudaMalloc((void**)&c->d_rec, cap * 8 * sizeof(double)));
	c->d_rec_cap = cap;
	return 0;
}

// ------------------------------------------------------------------------------------------------
// decomposition helpers (host only)
// ------------------------------------------------------------------------------------------------
// DomainDecomposition::getBoundingBoxMin/Max (DomainDecomposition.cpp:116-125)
extern "C" int ls1g_decomp_box(const int grid[3], const int coords[3], const double L[3], double box_min[3], double box_max[3]) {
	for (int d = 0; d < 3; ++d) {
		if (grid[d] < 1 || coords[d] < 0 || coords[d] >= grid[d]) LS1G_FAIL("ls1g_decomp_box: bad grid/coords");
		box_min[d] = coords[d] * L[d] / grid[d];
		box_max[d] = coords[d] + 1 == grid[d] ? L[d] : (coords[d] + 1) * L[d] / grid[d];
	}
	return 0;
}

// MPI_Dims_create(n, 3, dims) with dims = {0,0,0}: balanced factorisation, non-increasing order; MPI_Cart_coords:
// row-major, last dimension varies fastest (DomainDecomposition.cpp:35-39).
extern "C" int ls1g_decomp_grid(int n_ranks, int rank, int grid[3], int coords[3]) {
	if (n_ranks < 1 || rank < 0 || rank >= n_ranks) LS1G_FAIL("ls1g_decomp_grid: bad rank");
	int best[3] = {n_ranks, 1, 1};
	for (int a = 1; a <= n_ranks; ++a) {
		if (n_ranks % a) continue;
		int rest = n_ranks / a;
		for (int b = 1; b <= rest; ++b) {
			if (rest % b) continue;
			int cdim = rest / b;
			if (!(a >= b && b >= cdim)) continue;
			if (a - cdim < best[0] - best[2] || (a - cdim == best[0] - best[2] && a < best[0])) { best[0] = a; best[1] = b; best[2] = cdim; }
		}
	}
	for (int d = 0; d < 3; ++d) grid[d] = best[d];
	coords[2] = rank % grid[2];
	coords[1] = (rank / grid[2]) % grid[1];
	coords[0] = rank / (grid[2] * grid[1]);
	return 0;
}

// ------------------------------------------------------------------------------------------------
// measurement
// ------------------------------------------------------------------------------------------------
extern "C" int ls1g_timers(ls1g_ctx* c, double ms[LS1G_T_COUNT], uint64_t launches[LS1G_T_COUNT], int reset) {
	NEED(c);
	CUDA_TRY(cudaStreamSynchronize(c->stream));
	ls1g_resolve_timers(c);
	for (int k = 0; k < LS1G_T_COUNT; ++k) {
		if (ms) ms[k] = c->timers[k].ms;
		if (launches) launches[k] = c->timers[k].launches;
		if (reset) { c->timers[k].ms = 0.; c->timers[k].launches = 0; }
	}
	return 0;
}

int k_flop_count(ls1g_ctx* c, ls1g_flop_counts* out);
extern "C" int ls1g_flop_count(ls1g_ctx* c, ls1g_flop_counts* out) {
	if (!out) LS1G_FAIL("null argument");
	NEED_REF(c);
	if (!c->sorted) LS1G_FAIL("ls1g_flop_count: container not sorted (call ls1g_update / ls1g_exchange first)");
	return k_flop_count(c, out);
}

int k_rdf(ls1g_ctx* c, const ls1g_rdf_params* p, uint64_t* mol, uint64_t* site, uint64_t* ardf);
extern "C" int ls1g_rdf(ls1g_ctx* c, const ls1g_rdf_params* p, uint64_t* mol, uint64_t* site, uint64_t* ardf) {
	if (!p || !mol) LS1G_FAIL("null argument");
	if (p->bins == 0 || !(p->interval_length > 0.)) LS1G_FAIL("ls1g_rdf: bins and interval_length must be positive");
	NEED_REF(c);
	if (!c->sorted) LS1G_FAIL("ls1g_rdf: container not sorted (call ls1g_update / ls1g_exchange first)");
	return k_rdf(c, p, mol, site, ardf);
}

extern "C" int ls1g_measure_fp64_peak(int device, double* tflops) {
	if (!tflops) LS1G_FAIL("null argument");
	return k_fp64_peak(device, tflops);
}

extern "C" int ls1g_list_stats(ls1g_ctx* c, uint64_t* rebuilds, int* rows_per_block, int* entries_per_molecule, double* skin) {
	if (!c) LS1G_FAIL("null context");
	if (rebuilds) *rebuilds = c->list_rebuilds;
	if (rows_per_block) *rows_per_block = c->list_R;
	if (entries_per_molecule) *entries_per_molecule = c->list_lcap;
	if (skin) *skin = c->skin;
	return 0;
}

extern "C" int ls1g_last_pair_count(ls1g_ctx* c, uint64_t* pairs) {
	if (!c || !pairs) LS1G_FAIL("null argument");
	*pairs = c->last_pairs;
	return 0;
}
