// integrate.cu — Leapfrog translation + rigid-body quaternion rotation, kinetic sums, thermostat scaling.
//
// Replaces (reference, src/):
//   Leapfrog::transition1to2 -> FullMolecule::upd_preF    integrators/Leapfrog.cpp:48-68, molecules/FullMolecule.cpp:334-364
//   Leapfrog::transition2to3 -> FullMolecule::upd_postF   integrators/Leapfrog.cpp:70-155, molecules/FullMolecule.cpp:366-390
//   VelocityScalingThermostat::apply (global betas)       thermostats/VelocityScalingThermostat.cpp:45-80
//   Domain::calculateGlobalValues betas                   Domain.cpp:226-247
//
// One kernel template covers {post, scale, pre} in any combination the step order allows; in the
// device-resident loop (ls1g_run) NVE steps use the fused post+pre variant: one read of r,v,F (+q,L,M)
// and one write of r,v (+q,L) per molecule-update.  HBM-bound; bytes per molecule in DESIGN.md.
#include "ls1g_internal.cuh"

#define INT_THREADS 256

// Quaternion::differentiate (Quaternion.cpp:94-99)
__device__ __forceinline__ double4 quat_diff(const double4& q, const double w[3]) {
	double4 r;
	r.x = .5 * (-q.y * w[0] - q.z * w[1] - q.w * w[2]);
	r.y = .5 * ( q.x * w[0] - q.w * w[1] + q.z * w[2]);
	r.z = .5 * ( q.w * w[0] + q.x * w[1] - q.y * w[2]);
	r.w = .5 * (-q.z * w[0] + q.y * w[1] + q.x * w[2]);
	return r;
}

struct IntArgs {
	size_t n;
	double4* pos; double* vel; double4* q; double* L;
	const double* F; const double* M;
	const uint32_t* cell;
	const CompDev* comps;
	double dt;             // full step (pre)
	double dt_half;        // post
	const double* betas;   // device [2] beta_trans, beta_rot; nullptr = no scaling
	const double4* pos_build;          // neighbour-list mode: positions at the last list build (nullptr = not tracked)
	const double4* pos_prune;          // ... at the last prune of the inner lists (nullptr = no inner lists)
	unsigned long long* maxdisp;       // [0] max |r - r_build|^2, [1] max |r - r_prune|^2 as double bits (atomicMax)
	double* partials;      // [gridDim.x][4]
	const unsigned long long* guard; unsigned long long guard_thr;   // skip when *guard > guard_thr (see run_list_steps)
};

template <bool POST, bool PRE, bool ROT>
__global__ void __launch_bounds__(INT_THREADS) integrate_kernel(IntArgs a) {
	__shared__ double red[4 * 32];
	if (a.guard && *a.guard > a.guard_thr) return;
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	double acc[4] = {0., 0., 0., 0.};   // summv2, sumIw2, N, rotDOF
	double disp2 = 0., disp2i = 0.;
	bool active = i < a.n && !(a.cell[i] & HALO_FLAG);
	if (active) {
		double4 p = a.pos[i];
		const CompDev& cp = a.comps[pos_cid(p)];
		const double m = cp.mass;
		double v[3] = {a.vel[3 * i], a.vel[3 * i + 1], a.vel[3 * i + 2]};
		double F[3] = {a.F[3 * i], a.F[3 * i + 1], a.F[3 * i + 2]};
		double4 q; double L[3], M[3];
		if (ROT) {
			q = a.q[i];
			L[0] = a.L[3 * i]; L[1] = a.L[3 * i + 1]; L[2] = a.L[3 * i + 2];
			M[0] = a.M[3 * i]; M[1] = a.M[3 * i + 1]; M[2] = a.M[3 * i + 2];
		}
		if (POST) {                                             // FullMolecule::upd_postF
			double dtInv2m = a.dt_half / m, v2 = 0.;
#pragma unroll
			for (int d = 0; d < 3; ++d) { v[d] += dtInv2m * F[d]; v2 += v[d] * v[d]; }
			acc[0] = m * v2;
			if (ROT) {
#pragma unroll
				for (int d = 0; d < 3; ++d) L[d] += a.dt_half * M[d];
				double w[3], Iw2 = 0.;
				quat_rotateinv(q, L, w);
#pragma unroll
				for (int d = 0; d < 3; ++d) { w[d] *= cp.invI[d]; Iw2 += cp.I[d] * w[d] * w[d]; }
				acc[1] = Iw2;
			}
			acc[2] = 1.0;
			acc[3] = (double)cp.rot_dof;
		}
		if (a.betas) {                                          // VelocityScalingThermostat::apply
			double bt = a.betas[0], br = a.betas[1];
#pragma unroll
			for (int d = 0; d < 3; ++d) { v[d] *= bt; if (ROT) L[d] *= br; }
		}
		if (PRE) {                                              // FullMolecule::upd_preF
			double dt_halve = .5 * a.dt, dtInv2m = dt_halve / m;
			double r[3] = {p.x, p.y, p.z};
#pragma unroll
			for (int d = 0; d < 3; ++d) { v[d] += dtInv2m * F[d]; r[d] += a.dt * v[d]; }
			p.x = r[0]; p.y = r[1]; p.z = r[2];
			a.pos[i] = p;
			if (a.pos_build) { const double4 b = a.pos_build[i]; const double ex = p.x - b.x, ey = p.y - b.y, ez = p.z - b.z; disp2 = ex * ex + ey * ey + ez * ez; }
			if (a.pos_build && a.pos_prune) { const double4 b = a.pos_prune[i]; const double ex = p.x - b.x, ey = p.y - b.y, ez = p.z - b.z; disp2i = ex * ex + ey * ey + ez * ez; }
			if (ROT) {
				double w[3];
				quat_rotateinv(q, L, w);
#pragma unroll
				for (int d = 0; d < 3; ++d) w[d] *= cp.invI[d];
				double4 qh = quat_diff(q, w);
				qh.x = qh.x * dt_halve + q.x; qh.y = qh.y * dt_halve + q.y; qh.z = qh.z * dt_halve + q.z; qh.w = qh.w * dt_halve + q.w;
				double qcorr = 1. / sqrt(qh.x * qh.x + qh.y * qh.y + qh.z * qh.z + qh.w * qh.w);
				qh.x *= qcorr; qh.y *= qcorr; qh.z *= qcorr; qh.w *= qcorr;
#pragma unroll
				for (int d = 0; d < 3; ++d) L[d] += dt_halve * M[d];
				quat_rotateinv(qh, L, w);
#pragma unroll
				for (int d = 0; d < 3; ++d) w[d] *= cp.invI[d];
				double4 dq = quat_diff(qh, w);
				q.x += dq.x * a.dt; q.y += dq.y * a.dt; q.z += dq.z * a.dt; q.w += dq.w * a.dt;
				qcorr = 1. / sqrt(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
				q.x *= qcorr; q.y *= qcorr; q.z *= qcorr; q.w *= qcorr;
				a.q[i] = q;
			}
		}
		a.vel[3 * i] = v[0]; a.vel[3 * i + 1] = v[1]; a.vel[3 * i + 2] = v[2];
		if (ROT) { a.L[3 * i] = L[0]; a.L[3 * i + 1] = L[1]; a.L[3 * i + 2] = L[2]; }
	}
	if (PRE && a.pos_build) {          // displacement criterion of the neighbour lists (lists.cuh)
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) disp2 = fmax(disp2, __shfl_xor_sync(0xffffffffu, disp2, o));
		const unsigned long long bits = (unsigned long long)__double_as_longlong(disp2);
		if ((threadIdx.x & 31) == 0 && bits > *(volatile unsigned long long*)a.maxdisp) atomicMax(a.maxdisp, bits);
		if (a.pos_prune) {
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) disp2i = fmax(disp2i, __shfl_xor_sync(0xffffffffu, disp2i, o));
			const unsigned long long bi = (unsigned long long)__double_as_longlong(disp2i);
			if ((threadIdx.x & 31) == 0 && bi > *(volatile unsigned long long*)(a.maxdisp + 1)) atomicMax(a.maxdisp + 1, bi);
		}
	}
	if (POST) {
		block_sum<4>(acc, red);
		if (threadIdx.x == 0) {
#pragma unroll
			for (int k = 0; k < 4; ++k) a.partials[(size_t)blockIdx.x * 4 + k] = acc[k];
		}
	}
}

// deterministic final reduction of per-block partials: out[k] = sum_b partials[b][k]; optionally the
// thermostat betas of Domain::calculateGlobalValues (Domain.cpp:226-247) from the kinetic sums.
__global__ void __launch_bounds__(256) finalize_kernel(const double* __restrict__ partials, size_t nb, int nv, double* __restrict__ out,
                                                        double target_T, double* __restrict__ betas) {
	__shared__ double red[32];
	__shared__ double res[8];
	for (int k = 0; k < nv; ++k) {
		double s[1] = {0.};
		for (size_t b = threadIdx.x; b < nb; b += blockDim.x) s[0] += partials[b * nv + k];
		block_sum<1>(s, red);
		if (threadIdx.x == 0) { out[k] = s[0]; res[k] = s[0]; }
		__syncthreads();
	}
	if (betas && threadIdx.x == 0) {
		double summv2 = res[0], sumIw2 = res[1], N = res[2], rotDOF = res[3];
		double bt = 1.0, br = 1.0;
		if (target_T > 0.) {
			bt = summv2 > 0. ? pow(3.0 * N * target_T / summv2, 0.4) : 1.0;
			br = sumIw2 > 0. ? pow(rotDOF * target_T / sumIw2, 0.4) : 1.0;
		}
		betas[0] = bt; betas[1] = br;
	}
}

__global__ void __launch_bounds__(INT_THREADS) scale_kernel(size_t n, const uint32_t* __restrict__ cell, double* vel, double* L, double bt, double br) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n || (cell[i] & HALO_FLAG)) return;
#pragma unroll
	for (int d = 0; d < 3; ++d) { vel[3 * i + d] *= bt; if (L) L[3 * i + d] *= br; }
}

// Selective thermostat of the reference's "explosion heuristics" (Domain::calculateGlobalValues, Domain.cpp:254-295): a molecule
// whose translational energy exceeds limit_energy gets v and F scaled by sqrt(limit/U_trans); one whose rotational energy
// (FullMolecule::U_rot, FullMolecule.cpp:392-400) exceeds 3 * rotDOF * Ti gets D and M scaled by sqrt(limit/U_rot).
__global__ void __launch_bounds__(INT_THREADS) selective_kernel(size_t n, const uint32_t* __restrict__ cell, const double4* __restrict__ pos,
                                                                 double* vel, double* F, const double4* __restrict__ q, double* L, double* M,
                                                                 const CompDev* __restrict__ comps, double limit_energy, double Ti) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n || (cell[i] & HALO_FLAG)) return;
	const CompDev& cp = comps[pos_cid(pos[i])];
	const double v0 = vel[3 * i], v1 = vel[3 * i + 1], v2 = vel[3 * i + 2];
	const double Utrans = 0.5 * cp.mass * (v0 * v0 + v1 * v1 + v2 * v2);
	if (Utrans > limit_energy) {
		const double vcorr = sqrt(limit_energy / Utrans);
#pragma unroll
		for (int d = 0; d < 3; ++d) { vel[3 * i + d] *= vcorr; F[3 * i + d] *= vcorr; }
	}
	if (q && cp.rot_dof > 0) {
		const double limit_rot = 3.0 * cp.rot_dof * Ti;
		double Lm[3] = {L[3 * i], L[3 * i + 1], L[3 * i + 2]}, w[3], Iw2 = 0.;
		quat_rotateinv(q[i], Lm, w);
#pragma unroll
		for (int d = 0; d < 3; ++d) { w[d] *= cp.invI[d]; Iw2 += cp.I[d] * w[d] * w[d]; }
		const double Urot = 0.5 * Iw2;
		if (Urot > limit_rot) {
			const double Dcorr = sqrt(limit_rot / Urot);
#pragma unroll
			for (int d = 0; d < 3; ++d) { L[3 * i + d] *= Dcorr; M[3 * i + d] *= Dcorr; }
		}
	}
}

int ls1g_prereduce(ls1g_ctx* c, size_t& nb, int nv, const double*& base);   // force.cu

// ------------------------------------------------------------------------------------------------
static inline unsigned nblk(size_t n, unsigned t) { return (unsigned)((n + t - 1) / t); }

static IntArgs make_args(ls1g_ctx* c) {
	IntArgs a;
	a.n = c->n_slots; a.pos = c->cur.pos; a.vel = c->cur.vel; a.q = c->cur.q; a.L = c->cur.L;
	a.F = c->F; a.M = c->M; a.cell = c->cur.cell; a.comps = c->comps;
	a.dt = 0.; a.dt_half = 0.; a.betas = nullptr; a.partials = c->partials;
	a.pos_build = c->track_disp ? c->pos_build : nullptr; a.maxdisp = c->d_flags;
	a.pos_prune = c->track_disp && c->inner_valid ? c->pos_prune : nullptr;
	a.guard = nl_guard_ptr(c); a.guard_thr = nl_guard_thr(c);
	return a;
}

static int ensure_partials(ls1g_ctx* c, size_t blocks, int nv) {
	size_t need = blocks * nv + 64 + 256 * 8;   // + room for the first-level output of ls1g_prereduce
	if (need > c->partials_cap) {
		if (c->partials) cudaFree(c->partials);
		c->partials = nullptr; c->partials_cap = 0;
		CUDA_TRY(cudaMalloc((void**)&c->partials, need * 2 * sizeof(double)));
		c->partials_cap = need * 2;
	}
	return 0;
}
int ls1g_ensure_partials(ls1g_ctx* c, size_t blocks, int nv) { return ensure_partials(c, blocks, nv); }

template <bool POST, bool PRE>
static void launch_int(ls1g_ctx* c, const IntArgs& a, unsigned nb) {
	if (c->has_rot) integrate_kernel<POST, PRE, true><<<nb, INT_THREADS, 0, c->stream>>>(a);
	else integrate_kernel<POST, PRE, false><<<nb, INT_THREADS, 0, c->stream>>>(a);
}

int k_scale(ls1g_ctx* c, double bt, double br) {
	if (c->n_slots == 0) return 0;
	timer_start(c, LS1G_T_INTEGRATE);
	scale_kernel<<<nblk(c->n_slots, INT_THREADS), INT_THREADS, 0, c->stream>>>(c->n_slots, c->cur.cell, c->cur.vel,
	                                                                         c->has_rot ? c->cur.L : nullptr, bt, br);
	CUDA_TRY(cudaGetLastError());
	timer_stop(c, LS1G_T_INTEGRATE, 1);
	return 0;
}

int k_selective(ls1g_ctx* c, double limit_energy, double Ti) {
	if (c->n_slots == 0) return 0;
	timer_start(c, LS1G_T_INTEGRATE);
	selective_kernel<<<nblk(c->n_slots, INT_THREADS), INT_THREADS, 0, c->stream>>>(c->n_slots, c->cur.cell, c->cur.pos, c->cur.vel, c->F,
	                                                                             c->has_rot ? c->cur.q : nullptr, c->cur.L, c->M, c->comps, limit_energy, Ti);
	CUDA_TRY(cudaGetLastError());
	timer_stop(c, LS1G_T_INTEGRATE, 1);
	return 0;
}

// VelocityScalingThermostat::apply with host-given betas fused into the pre-kick (upd_preF) of the next step: one pass over v
int k_scale_pre(ls1g_ctx* c, double dt, double bt, double br) {
	unsigned nb = nblk(c->n_slots, INT_THREADS);
	if (!nb) return 0;
	timer_start(c, LS1G_T_INTEGRATE);
	c->h_scalars[8] = bt; c->h_scalars[9] = br;       // pinned; consumed by the copy below before any later call can rewrite it (callers sync)
	CUDA_TRY(cudaMemcpyAsync(c->d_scalars + 8, c->h_scalars + 8, 2 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
	IntArgs a = make_args(c); a.dt = dt; a.dt_half = 0.5 * dt; a.partials = c->partials;
	a.betas = c->d_scalars + 8;
	launch_int<false, true>(c, a, nb);
	CUDA_TRY(cudaGetLastError());
	timer_stop(c, LS1G_T_INTEGRATE, 1);
	return 0;
}

// [post(step n)] [+ thermostat] [+ pre(step n+1)].
//   no thermostat : ONE kernel (post+pre fused when both are requested) + the partial-sum finalize
//   thermostat    : post, finalize (betas computed on the device, Domain.cpp:226-247), then scale(+pre) in one kernel
// Kinetic sums of the post part end in d_scalars[4..7] (summv2, sumIw2, N, rotDOF), betas in d_scalars[8..9].
int k_post_pre(ls1g_ctx* c, double dt, int do_post, int do_pre, int thermostat, double target_T) {
	unsigned nb = nblk(c->n_slots, INT_THREADS);
	if (ensure_partials(c, nb ? nb : 1, 4)) return 1;
	timer_start(c, LS1G_T_INTEGRATE);
	IntArgs a = make_args(c); a.dt = dt; a.dt_half = 0.5 * dt; a.partials = c->partials;
	int launches = 0;
	size_t nbr = nb; const double* pbase = c->partials;
	if (do_post && !thermostat) {
		if (nb) { if (do_pre) launch_int<true, true>(c, a, nb); else launch_int<true, false>(c, a, nb); ++launches; }
		if (ls1g_prereduce(c, nbr, 4, pbase)) return 1;
		finalize_kernel<<<1, 256, 0, c->stream>>>(pbase, nbr, 4, c->d_scalars + 4, 0., nullptr); launches += 1 + (nbr != nb);
	} else if (do_post) {
		// one rank (several ranks: k_post_sums + group all-reduce + k_betas_scale_pre, see g_post_pre in step.cu)
		if (nb) { launch_int<true, false>(c, a, nb); ++launches; }
		if (ls1g_prereduce(c, nbr, 4, pbase)) return 1;
		finalize_kernel<<<1, 256, 0, c->stream>>>(pbase, nbr, 4, c->d_scalars + 4, target_T, c->d_scalars + 8); launches += 1 + (nbr != nb);
		a.betas = c->d_scalars + 8;
		if (nb) { if (do_pre) launch_int<false, true>(c, a, nb); else launch_int<false, false>(c, a, nb); ++launches; }
	} else if (do_pre) {
		if (nb) { launch_int<false, true>(c, a, nb); ++launches; }
	}
	CUDA_TRY(cudaGetLastError());
	timer_stop(c, LS1G_T_INTEGRATE, launches);
	return 0;
}

// Thermostat step on several ranks, first half: post kick + this rank's kinetic sums in d_scalars[4..7]
int k_post_sums(ls1g_ctx* c, double dt) {
	unsigned nb = nblk(c->n_slots, INT_THREADS);
	if (ensure_partials(c, nb ? nb : 1, 4)) return 1;
	timer_start(c, LS1G_T_INTEGRATE);
	IntArgs a = make_args(c); a.dt = dt; a.dt_half = 0.5 * dt; a.partials = c->partials;
	int launches = 0;
	size_t nbr = nb; const double* pbase = c->partials;
	if (nb) { launch_int<true, false>(c, a, nb); ++launches; }
	if (ls1g_prereduce(c, nbr, 4, pbase)) return 1;
	finalize_kernel<<<1, 256, 0, c->stream>>>(pbase, nbr, 4, c->d_scalars + 4, 0., nullptr); launches += 1 + (nbr != nb);
	CUDA_TRY(cudaGetLastError());
	timer_stop(c, LS1G_T_INTEGRATE, launches);
	return 0;
}

// ... second half, after the all-reduce of the sums into d_scalars[12..15]: betas of Domain::calculateGlobalValues from the global sums, then
// VelocityScalingThermostat::apply (+ the pre kick of the next step) in one kernel
int k_betas_scale_pre(ls1g_ctx* c, double dt, int do_pre, double target_T) {
	unsigned nb = nblk(c->n_slots, INT_THREADS);
	timer_start(c, LS1G_T_INTEGRATE);
	IntArgs a = make_args(c); a.dt = dt; a.dt_half = 0.5 * dt; a.partials = c->partials;
	int launches = 1;
	finalize_kernel<<<1, 256, 0, c->stream>>>(c->d_scalars + 12, 1, 4, c->d_scalars + 12, target_T, c->d_scalars + 8);
	a.betas = c->d_scalars + 8;
	if (nb) { if (do_pre) launch_int<false, true>(c, a, nb); else launch_int<false, false>(c, a, nb); ++launches; }
	CUDA_TRY(cudaGetLastError());
	timer_stop(c, LS1G_T_INTEGRATE, launches);
	return 0;
}
