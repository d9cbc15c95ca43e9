// force_1clj.cu — force pass specialised for a single-centre Lennard-Jones fluid (configs A, B, E).
//
// Same contract as force.cu (owner-computes full shell, calcFM fused); replaces for this component
// signature VectorizedCellProcessor::_calculatePairs' LJ loop (adapter/VectorizedCellProcessor.cpp:1031-1180,
// _loopBodyLJ :173-226) and calcDistLookup (vectorization/SIMD_VectorizedCellProcessorHelpers.h:427-472).
#include "ls1g_internal.cuh"
#include "potentials.cuh"

int ls1g_ensure_partials(ls1g_ctx* c, size_t blocks, int nv);
int k_force_finalize(ls1g_ctx* c, size_t nb);

struct LJ1Args {
	Grid g;
	size_t n;
	const double4* pos; const uint32_t* cell; const uint32_t* cell_start;
	double eps24, sig2, shift6, rclj2;
	double *F, *Vi;
	double* partials;   // [blocks][5]
};

#define LJ1_THREADS 128

// v1: one thread per owned molecule, neighbours streamed through L1/L2 (reference implementation of the
// GPU path; the tiled kernel below is validated against it).
__global__ void __launch_bounds__(LJ1_THREADS) force_1clj_simple_kernel(LJ1Args a) {
	__shared__ double red[5 * 32];
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	double acc[5] = {0., 0., 0., 0., 0.};
	uint32_t cflag = i < a.n ? a.cell[i] : HALO_FLAG;
	if (!(cflag & HALO_FLAG)) {
		const Grid& g = a.g;
		const double4 Pi = a.pos[i];
		double Fx = 0., Fy = 0., Fz = 0., Vx = 0., Vy = 0., Vz = 0., u6 = 0.;
		unsigned np = 0;
		const int cx = cflag % g.cpd[0], cy = (cflag / g.cpd[0]) % g.cpd[1], cz = cflag / (g.cpd[0] * g.cpd[1]);
		const uint32_t own0 = a.cell_start[cflag], own1 = a.cell_start[cflag + 1];
		for (int dz = -1; dz <= 1; ++dz)
		for (int dy = -1; dy <= 1; ++dy) {
			uint32_t crow = (uint32_t)(((cz + dz) * g.cpd[1] + (cy + dy)) * g.cpd[0] + cx);
			uint32_t j0 = a.cell_start[crow - 1], j1 = a.cell_start[crow + 2];
			for (uint32_t j = j0; j < j1; ++j) {
				const double4 Pj = a.pos[j];
				const double dx = Pi.x - Pj.x, dy_ = Pi.y - Pj.y, dz_ = Pi.z - Pj.z;
				const double r2 = dx * dx + dy_ * dy_ + dz_ * dz_;
				if (r2 < a.rclj2 && j != (uint32_t)i) {
					if (r2 == 0.0 && j >= own0 && j < own1) continue;   // SingleCellPolicy_ mask (Helpers.h:367-376)
					const double inv = fast_rcp(r2);
					const double lj2 = a.sig2 * inv, lj4 = lj2 * lj2, lj6 = lj4 * lj2, lj12 = lj6 * lj6, lj12m6 = lj12 - lj6;
					const double scale = (a.eps24 * inv) * (lj12 + lj12m6);
					const double fx = dx * scale, fy = dy_ * scale, fz = dz_ * scale;
					Fx += fx; Fy += fy; Fz += fz;
					Vx += dx * fx; Vy += dy_ * fy; Vz += dz_ * fz;
					u6 += a.eps24 * lj12m6 + a.shift6;
					++np;
				}
			}
		}
		a.F[3 * i] = Fx; a.F[3 * i + 1] = Fy; a.F[3 * i + 2] = Fz;
		a.Vi[3 * i] = 0.5 * Vx; a.Vi[3 * i + 1] = 0.5 * Vy; a.Vi[3 * i + 2] = 0.5 * Vz;
		acc[0] = u6; acc[2] = Vx + Vy + Vz; acc[4] = (double)np;
	}
	block_sum<5>(acc, red);
	if (threadIdx.x == 0) {
#pragma unroll
		for (int k = 0; k < 5; ++k) a.partials[(size_t)blockIdx.x * 5 + k] = acc[k];
	}
}


#include "force_1clj_tile.cuh"

static inline unsigned nblk(size_t n, unsigned t) { return (unsigned)((n + t - 1) / t); }

int k_forces_1clj(ls1g_ctx* c) {
	size_t n = c->n_slots;
	const Grid& g = c->bgrid;
	if (c->force_variant == 1) {   // v1 (LS1G_FORCE_KERNEL=simple): kept as the in-library cross-check of v2
		unsigned nb = nblk(n, LJ1_THREADS);
		if (ls1g_ensure_partials(c, nb ? nb : 1, 5)) return 1;
		if (n) {
			LJ1Args a;
			a.g = c->bgrid; a.n = n; a.pos = c->cur.pos; a.cell = c->cur.cell; a.cell_start = c->cell_start;
			a.eps24 = c->eps24_h[0]; a.sig2 = c->sig2_h[0]; a.shift6 = c->shift6_h[0]; a.rclj2 = c->rclj2;
			a.F = c->F; a.Vi = c->Vi; a.partials = c->partials;
			timer_start(c, LS1G_T_FORCE);
			force_1clj_simple_kernel<<<nb, LJ1_THREADS, 0, c->stream>>>(a);
			timer_stop(c, LS1G_T_FORCE, 1);
		}
		if (k_force_finalize(c, nb)) return 1;
		CUDA_TRY(cudaGetLastError());
		return 0;
	}
	LJ3Args a;
	a.g = g; a.pos = c->cur.pos; a.cell_start = c->cell_start;
	a.eps24 = c->eps24_h[0]; a.sig2 = c->sig2_h[0]; a.shift6 = c->shift6_h[0]; a.rclj2 = c->rclj2;
	// FP32 filter: coordinates relative to the segment centre are < 6 cell lengths, i.e. an absolute error of
	// ~2^-24 * 6 cells per coordinate and < 1e-5 relative on r^2 near rc^2; the 3e-4 margin is > 30x that bound, so the
	// filter never rejects a pair the FP64 test accepts
	a.rc2f = (float)(c->rclj2 * (1.0 + 3e-4));
	a.F = c->F; a.Vi = c->Vi;
	a.nxi = g.cpd[0] - 2; a.nyi = g.cpd[1] - 2;
	a.nbx = (a.nxi + LJ3_GX - 1) / LJ3_GX;
	unsigned nb = (unsigned)(a.nbx * a.nyi * (g.cpd[2] - 2));
	if (ls1g_ensure_partials(c, nb ? nb : 1, 5)) return 1;
	a.partials = c->partials;
	if (n && nb) {
		static unsigned long long attr_set = 0;      // bit d: done on device d (function attributes are per device)
		const int d = c->cfg.device;
		if (d >= 64 || !(attr_set >> d & 1ull)) {
			CUDA_TRY(cudaFuncSetAttribute(force_1clj_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LJ3_SMEM_BYTES));
			if (d < 64) attr_set |= 1ull << d;
		}
		timer_start(c, LS1G_T_FORCE);
		force_1clj_tile_kernel<<<nb, LJ3_W * 32, LJ3_SMEM_BYTES, c->stream>>>(a);
		timer_stop(c, LS1G_T_FORCE, 1);
	} else nb = 0;
	if (k_force_finalize(c, nb)) return 1;
	CUDA_TRY(cudaGetLastError());
	return 0;
}
