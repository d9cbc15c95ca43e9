// force_1clj_tile.cuh — the production 1CLJ force kernel (included by force_1clj.cu).
//
// Work decomposition
//   block  = one x-segment of LJ3_GX = LJ3_W * LJ3_G non-halo cells of one cell row (y, z)
//   warp   = LJ3_G x-adjacent cells of that segment; lane = one molecule i of those cells (cell-sorted slots are
//            contiguous along x, so the lanes' slots are consecutive)
//   tile   = the 9 neighbouring cell rows (dy, dz in -1..1), each cut to cells [x0-1, x0+GX]: 9 contiguous slot ranges
//            of the cell-sorted arrays, staged ONCE per block into shared memory in two forms:
//              xs/ys/zs  FP32 position relative to the segment centre  (candidate filter)
//              xd/yd/zd  the original FP64 position                    (exact test + force)
// Per warp
//   filter  all 32 lanes walk the warp's sub-range of every staged row in lockstep (every load is a shared-memory
//           BROADCAST of 4 candidates); each lane tests its own molecule in FP32 with a safety margin that never
//           rejects a pair the FP64 test accepts, and appends survivors to its PRIVATE queue q[t][lane]
//   force   each lane drains its queue: exact FP64 test on the original doubles (same decision as calcDistLookup,
//           SIMD_VectorizedCellProcessorHelpers.h:427-472), then _loopBodyLJ (VectorizedCellProcessor.cpp:173-226);
//           F, V, upot6 stay in the lane's registers: no cross-lane reduction, no atomics, fixed summation order
// The FP64 pipe only sees the ~15 % of the candidates that are inside the cutoff; the scan runs on the FP32 pipe.
// Dense systems (tile or queue larger than the buffers) are handled by staging the rows in several passes and by
// draining the queues early; molecules beyond 32 per warp are handled in chunks of 32.
#pragma once

#define LJ3_W 4          // warps per block
#define LJ3_G 2          // cells per warp
#define LJ3_GX (LJ3_W * LJ3_G)
#define LJ3_TCAP 1344    // candidates per staged tile (multiple of 4)
#define LJ3_QCAP 96      // queue entries per lane
#define LJ3_MAXP 12      // pieces per stage

struct LJ3Args {
	Grid g;
	const double4* pos; const uint32_t* cell_start;
	double eps24, sig2, shift6, rclj2;
	float rc2f;                 // rclj2 * (1 + margin) in FP32
	double *F, *Vi;
	double* partials;           // [blocks][5]
	int nbx, nxi, nyi;
};

#define LJ3_SMEM_BYTES (LJ3_TCAP * (3 * 8 + 3 * 4) + LJ3_W * LJ3_QCAP * 32 * 2)

__global__ void __launch_bounds__(LJ3_W * 32) force_1clj_tile_kernel(LJ3Args a) {
	extern __shared__ __align__(16) unsigned char lj3_smem[];
	double* xd = (double*)lj3_smem; double* yd = xd + LJ3_TCAP; double* zd = yd + LJ3_TCAP;
	float* xs = (float*)(zd + LJ3_TCAP); float* ys = xs + LJ3_TCAP; float* zs = ys + LJ3_TCAP;
	unsigned short* qall = (unsigned short*)(zs + LJ3_TCAP);
	__shared__ uint32_t s_rowbeg[9], s_rowend[9];
	__shared__ uint32_t s_wlo[LJ3_W][9], s_whi[LJ3_W][9];
	__shared__ int s_prow[LJ3_MAXP], s_poff[LJ3_MAXP], s_plen[LJ3_MAXP];
	__shared__ uint32_t s_pslot[LJ3_MAXP];
	__shared__ int s_np, s_next_r; __shared__ uint32_t s_next_s;
	__shared__ double red[5 * 32];

	const Grid& g = a.g;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const int nx = g.cpd[0], ny = g.cpd[1];
	const int bx = blockIdx.x % a.nbx, cy = (blockIdx.x / a.nbx) % a.nyi + 1, cz = blockIdx.x / (a.nbx * a.nyi) + 1;
	const int x0 = 1 + bx * LJ3_GX;                       // first own cell of the block
	const int xe = min(x0 + LJ3_GX, nx - 1);              // last staged cell (own + 1, clipped to the halo layer)
	const int xw = x0 + warp * LJ3_G;                     // first own cell of this warp
	const bool warp_valid = xw <= nx - 2;
	const uint32_t* cs = a.cell_start;
	if (tid < 9) {
		const int crow = ((cz + tid / 3 - 1) * ny + (cy + tid % 3 - 1)) * nx;
		s_rowbeg[tid] = cs[crow + x0 - 1];
		s_rowend[tid] = cs[crow + xe + 1];
	}
	if (lane < 9 && warp_valid) {
		const int crow = ((cz + lane / 3 - 1) * ny + (cy + lane % 3 - 1)) * nx;
		s_wlo[warp][lane] = cs[crow + xw - 1];
		s_whi[warp][lane] = cs[crow + min(xw + LJ3_G, nx - 1) + 1];
	}
	uint32_t ibeg = 0, iend = 0;
	if (warp_valid) {
		const int c4 = (cz * ny + cy) * nx;
		ibeg = cs[c4 + xw];
		iend = cs[c4 + min(xw + LJ3_G - 1, nx - 2) + 1];
	}
	if (tid == 0) { s_next_r = 0; }
	__syncthreads();
	if (tid == 0) s_next_s = s_rowbeg[0];
	// FP32 frame: centre of the staged segment
	const double ox = g.halo_min[0] + (x0 + 0.5 * LJ3_GX) * g.cell_len[0];
	const double oy = g.halo_min[1] + (cy + 0.5) * g.cell_len[1];
	const double oz = g.halo_min[2] + (cz + 0.5) * g.cell_len[2];
	unsigned short* q = qall + warp * (LJ3_QCAP * 32) + lane;     // q[t * 32]
	double acc_u6 = 0., acc_vir = 0., acc_np = 0.;
	bool first_stage = true;
	__syncthreads();
	for (;;) {
		// ---- plan one stage: as many (row, slot range) pieces as fit into the tile ---------------------
		if (tid == 0) {
			int r = s_next_r, np = 0, fill = 0;
			uint32_t s = s_next_s;
			while (r < 9 && np < LJ3_MAXP && fill + 4 <= LJ3_TCAP) {
				const uint32_t left = s_rowend[r] - s;
				const int take = (int)min(left, (uint32_t)(LJ3_TCAP - fill));
				if (take > 0) {
					s_prow[np] = r; s_pslot[np] = s; s_poff[np] = fill; s_plen[np] = take; ++np;
					fill += (take + 3) & ~3; s += take;
				}
				if (s == s_rowend[r]) { ++r; if (r < 9) s = s_rowbeg[r]; }
			}
			s_np = np; s_next_r = r; s_next_s = s;
		}
		__syncthreads();
		const int np = s_np;
		if (np == 0) break;
		// ---- stage ------------------------------------------------------------------------------------
		for (int p = 0; p < np; ++p) {
			const int off = s_poff[p], len = s_plen[p], padded = (len + 3) & ~3;
			const uint32_t slot = s_pslot[p];
			for (int t = tid; t < padded; t += LJ3_W * 32) {
				if (t < len) {
					const double4 P = a.pos[slot + t];
					xd[off + t] = P.x; yd[off + t] = P.y; zd[off + t] = P.z;
					xs[off + t] = (float)(P.x - ox); ys[off + t] = (float)(P.y - oy); zs[off + t] = (float)(P.z - oz);
				} else {
					xs[off + t] = 1e30f; ys[off + t] = 1e30f; zs[off + t] = 1e30f;
					xd[off + t] = 1e300; yd[off + t] = 1e300; zd[off + t] = 1e300;
				}
			}
		}
		__syncthreads();
		// ---- filter + force, per chunk of 32 molecules of this warp ---------------------------------------
		for (uint32_t ic = ibeg; ic < iend; ic += 32) {
			const uint32_t i = ic + lane;
			const bool active = i < iend;
			double4 Pi = make_double4(0., 0., 0., 0.);
			float fx = -1e30f, fy = -1e30f, fz = -1e30f;
			if (active) {
				Pi = a.pos[i];
				fx = (float)(Pi.x - ox); fy = (float)(Pi.y - oy); fz = (float)(Pi.z - oz);
			}
			double Fx = 0., Fy = 0., Fz = 0., Vx = 0., Vy = 0., Vz = 0., u6 = 0.;
			int np_i = 0;
			int cnt = 0;
			auto drain = [&]() {
				const int nmax = __reduce_max_sync(0xffffffffu, cnt);
#pragma unroll 2
				for (int t = 0; t < nmax; ++t) {
					if (t < cnt) {
						const int k = q[t * 32];
						const double dx = Pi.x - xd[k], dy = Pi.y - yd[k], dz = Pi.z - zd[k];
						const double r2 = dx * dx + dy * dy + dz * dz;
						// r2 == 0: the molecule itself, or a coincident molecule of the same cell (SingleCellPolicy_ mask, Helpers.h:367-376)
						if (r2 < a.rclj2 && r2 != 0.0) {
							const double inv = fast_rcp(r2);
							const double lj2 = a.sig2 * inv, lj4 = lj2 * lj2, lj6 = lj4 * lj2, lj12 = lj6 * lj6, lj12m6 = lj12 - lj6;
							const double scale = (a.eps24 * inv) * (lj12 + lj12m6);
							const double px = dx * scale, py = dy * scale, pz = dz * scale;
							Fx += px; Fy += py; Fz += pz;
							Vx += dx * px; Vy += dy * py; Vz += dz * pz;
							u6 += a.eps24 * lj12m6 + a.shift6;
							++np_i;
						}
					}
				}
				cnt = 0;
			};
			for (int p = 0; p < np; ++p) {
				const int row = s_prow[p];
				const uint32_t slot = s_pslot[p];
				const uint32_t lo = max(slot, s_wlo[warp][row]), hi = min(slot + (uint32_t)s_plen[p], s_whi[warp][row]);
				if (lo >= hi) continue;
				const int off = s_poff[p];
				const int k0 = off + (int)((lo - slot) & ~3u), k1 = off + (int)((hi - slot + 3u) & ~3u);
				for (int k = k0; k < k1; k += 4) {
					const float4 cx4 = *(const float4*)(xs + k), cy4 = *(const float4*)(ys + k), cz4 = *(const float4*)(zs + k);
					const float cxa[4] = {cx4.x, cx4.y, cx4.z, cx4.w}, cya[4] = {cy4.x, cy4.y, cy4.z, cy4.w}, cza[4] = {cz4.x, cz4.y, cz4.z, cz4.w};
#pragma unroll
					for (int u = 0; u < 4; ++u) {
						const float dx = fx - cxa[u], dy = fy - cya[u], dz = fz - cza[u];
						const float r2 = dx * dx + dy * dy + dz * dz;
						if (r2 < a.rc2f) { q[cnt * 32] = (unsigned short)(k + u); ++cnt; }
					}
					if (__any_sync(0xffffffffu, cnt > LJ3_QCAP - 4)) drain();
				}
			}
			drain();
			if (active) {
				double* Fp = a.F + 3 * (size_t)i; double* Vp = a.Vi + 3 * (size_t)i;
				if (first_stage) { Fp[0] = Fx; Fp[1] = Fy; Fp[2] = Fz; Vp[0] = 0.5 * Vx; Vp[1] = 0.5 * Vy; Vp[2] = 0.5 * Vz; }
				else { Fp[0] += Fx; Fp[1] += Fy; Fp[2] += Fz; Vp[0] += 0.5 * Vx; Vp[1] += 0.5 * Vy; Vp[2] += 0.5 * Vz; }
				acc_u6 += u6; acc_vir += Vx + Vy + Vz; acc_np += (double)np_i;
			}
		}
		first_stage = false;
		__syncthreads();
	}
	double acc[5] = {acc_u6, 0., acc_vir, 0., acc_np};
	block_sum<5>(acc, red);
	if (tid == 0) {
#pragma unroll
		for (int k = 0; k < 5; ++k) a.partials[(size_t)blockIdx.x * 5 + k] = acc[k];
	}
}
