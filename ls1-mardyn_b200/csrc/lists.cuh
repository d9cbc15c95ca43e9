// lists.cuh — neighbour-list kernels of the device-resident time loop (1CLJ): list build + list-driven force pass.
//
// Why: the reference re-enumerates all 27-cell candidate pairs every step (calcDistLookup,
// SIMD_VectorizedCellProcessorHelpers.h:427-472); only ~15 % of them are inside the cutoff.  On the GPU that candidate
// scan costs more than the force arithmetic.  Inside ls1g_run the scan is therefore done once every few steps with the
// radius rc + skin and its survivors are kept (a Verlet list); the steps in between only evaluate listed pairs.  A pair
// closer than rc at ANY later step was closer than rc + skin when the list was built as long as no molecule moved more
// than skin/2 since then - the integrator tracks the maximum displacement and the host rebuilds before that happens - so
// the set of pairs inside the cutoff, and with it forces, upot and virial, are exactly those of the per-step enumeration.
//
// Frozen layout: between rebuilds molecules keep their slot (no re-sort, no wrap); periodic images are refreshed from
// their source slot (halo_refresh_kernel).  The sort bins on c->bgrid whose cells are >= rc + skin.
//
// Work decomposition (both kernels, identical so that tile indices stay valid):
//   block (row, b)  = 128 consecutive owned slots [s0, s1) of one cell row (y, z); lane = one molecule
//   tile            = the 9 neighbouring rows cut to cells [x_first-1, x_last+1] of the block's molecules: 9 contiguous
//                     slot ranges, staged in shared memory (FP64 originals; the build adds FP32 copies for the filter)
//   list entry      = 16-bit tile index; lists[((blk*(LCAP/8) + t/8)*128 + thread)*8 + t%8]: one 16-byte load = 8 entries
#pragma once

#define NL_BT 128          // threads per block = molecules per block
#define NL_TCAP 2560       // tile capacity (candidates), multiple of 4
#define NL_LCAP 128        // list entries per molecule, multiple of 8

#define NL_FORCE_SMEM (3 * NL_TCAP * 8)

enum { NL_OVF_TILE = 1, NL_OVF_LIST = 2, NL_OVF_ROW = 4 };

struct NLArgs {
	Grid g;                     // bin grid
	const double4* pos; const uint32_t* cell; const uint32_t* cell_start;
	unsigned short* lists; uint32_t* lcount;
	unsigned long long* flags;  // [1] |= NL_OVF_*
	int bpr;                    // blocks per row
	int nyi;
	float rl2f;                 // (rc + skin)^2 * (1 + margin), FP32 filter
	double eps24, sig2, shift6, rclj2;
	double *F, *Vi;
	double* partials;           // [blocks][5]
};

struct NLBlock {
	uint32_t s0, s1;            // owned slots of this block
	int total;                  // tile entries (padded)
	bool live;
};

// Common prologue: slot range, tile plan (s_beg/s_off per row) - identical in build and force kernels.
__device__ __forceinline__ NLBlock nl_plan(const NLArgs& a, uint32_t* s_beg, int* s_off, int* s_len, int& x_first, int& x_last) {
	const Grid& g = a.g;
	const int nx = g.cpd[0], ny = g.cpd[1];
	const int row = blockIdx.x / a.bpr, bi = blockIdx.x % a.bpr;
	const int cy = row % a.nyi + 1, cz = row / a.nyi + 1;
	const int crow = (cz * ny + cy) * nx;
	const uint32_t rs = a.cell_start[crow + 1], re = a.cell_start[crow + nx - 1];
	NLBlock b;
	b.s0 = rs + (uint32_t)bi * NL_BT;
	b.live = b.s0 < re;
	b.s1 = min(b.s0 + NL_BT, re);
	b.total = 0;
	if (bi == a.bpr - 1 && b.s0 + NL_BT < re && threadIdx.x == 0) atomicOr(a.flags + 1, (unsigned long long)NL_OVF_ROW);
	if (!b.live) return b;
	x_first = (int)((a.cell[b.s0] & ~HALO_FLAG) % (uint32_t)nx);
	x_last = (int)((a.cell[b.s1 - 1] & ~HALO_FLAG) % (uint32_t)nx);
	if (threadIdx.x < 9) {
		const int r = threadIdx.x;
		const int cr = ((cz + r / 3 - 1) * ny + (cy + r % 3 - 1)) * nx;
		const uint32_t beg = a.cell_start[cr + x_first - 1], end = a.cell_start[cr + x_last + 2];
		s_beg[r] = beg; s_len[r] = (int)(end - beg);
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		int off = 0;
		for (int r = 0; r < 9; ++r) { s_off[r] = off; off += (s_len[r] + 3) & ~3; }
		s_off[9] = off;
	}
	__syncthreads();
	b.total = s_off[9];
	if (b.total > NL_TCAP) {
		if (threadIdx.x == 0) atomicOr(a.flags + 1, (unsigned long long)NL_OVF_TILE);
		b.live = false;
	}
	return b;
}

// ---------------------------------------------------------------------------------------------------------------
// build: FP32 candidate filter with radius rc + skin -> per-molecule lists of tile indices
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(NL_BT) nl_build_kernel(NLArgs a) {
	__shared__ __align__(16) float xs[NL_TCAP], ys[NL_TCAP], zs[NL_TCAP];
	__shared__ uint32_t s_beg[9]; __shared__ int s_off[10], s_len[9];
	const Grid& g = a.g;
	const int tid = threadIdx.x, lane = tid & 31;
	int x_first = 0, x_last = 0;
	const NLBlock b = nl_plan(a, s_beg, s_off, s_len, x_first, x_last);
	if (!b.live) return;
	const int nx = g.cpd[0], ny = g.cpd[1];
	const int row = blockIdx.x / a.bpr;
	const int cy = row % a.nyi + 1, cz = row / a.nyi + 1;
	const double ox = g.halo_min[0] + 0.5 * (x_first + x_last + 1) * g.cell_len[0];
	const double oy = g.halo_min[1] + (cy + 0.5) * g.cell_len[1];
	const double oz = g.halo_min[2] + (cz + 0.5) * g.cell_len[2];
	for (int r = 0; r < 9; ++r) {
		const int off = s_off[r], len = s_len[r], padded = (len + 3) & ~3;
		const uint32_t beg = s_beg[r];
		for (int t = tid; t < padded; t += NL_BT) {
			float x = 1e30f, y = 1e30f, z = 1e30f;
			if (t < len) { const double4 P = a.pos[beg + t]; x = (float)(P.x - ox); y = (float)(P.y - oy); z = (float)(P.z - oz); }
			xs[off + t] = x; ys[off + t] = y; zs[off + t] = z;
		}
	}
	__syncthreads();
	// this warp's molecules and the cells they span
	const uint32_t w0 = b.s0 + (uint32_t)(tid & ~31);
	if (w0 >= b.s1) return;
	const uint32_t w1 = min(w0 + 32u, b.s1);
	const int xw_first = (int)((a.cell[w0] & ~HALO_FLAG) % (uint32_t)nx), xw_last = (int)((a.cell[w1 - 1] & ~HALO_FLAG) % (uint32_t)nx);
	const uint32_t i = w0 + lane;
	const bool active = i < w1;
	float fx = -1e30f, fy = -1e30f, fz = -1e30f;
	if (active) { const double4 Pi = a.pos[i]; fx = (float)(Pi.x - ox); fy = (float)(Pi.y - oy); fz = (float)(Pi.z - oz); }
	unsigned short* my = a.lists + ((size_t)blockIdx.x * (NL_LCAP / 8) * NL_BT + tid) * 8;     // entry t at my[(t/8)*NL_BT*8 + t%8]
	int cnt = 0; bool ovf = false;
	const float rl2 = a.rl2f;
	for (int r = 0; r < 9; ++r) {
		const int cr = ((cz + r / 3 - 1) * ny + (cy + r % 3 - 1)) * nx;
		const uint32_t lo = a.cell_start[cr + xw_first - 1], hi = a.cell_start[cr + xw_last + 2];
		const uint32_t beg = s_beg[r];
		const int k0 = s_off[r] + (int)((lo - beg) & ~3u), k1 = s_off[r] + (int)((hi - beg + 3u) & ~3u);
		for (int k = k0; k < k1; k += 4) {
			const float4 cx4 = *(const float4*)(xs + k), cy4 = *(const float4*)(ys + k), cz4 = *(const float4*)(zs + k);
			const float cxa[4] = {cx4.x, cx4.y, cx4.z, cx4.w}, cya[4] = {cy4.x, cy4.y, cy4.z, cy4.w}, cza[4] = {cz4.x, cz4.y, cz4.z, cz4.w};
#pragma unroll
			for (int u = 0; u < 4; ++u) {
				const float dx = fx - cxa[u], dy = fy - cya[u], dz = fz - cza[u];
				const float r2 = dx * dx + dy * dy + dz * dz;
				if (r2 < rl2) { my[(cnt >> 3) * (NL_BT * 8) + (cnt & 7)] = (unsigned short)(k + u); ++cnt; }
			}
			if (cnt > NL_LCAP - 4) { ovf = true; cnt = 0; }
		}
	}
	if (active) a.lcount[(size_t)blockIdx.x * NL_BT + tid] = (uint32_t)cnt;
	if (ovf && active) atomicOr(a.flags + 1, (unsigned long long)NL_OVF_LIST);
}

// ---------------------------------------------------------------------------------------------------------------
// force: exact FP64 test + _loopBodyLJ over the listed pairs
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(NL_BT) nl_force_kernel(NLArgs a) {
	extern __shared__ __align__(16) double nl_tile[];
	double* xd = nl_tile; double* yd = xd + NL_TCAP; double* zd = yd + NL_TCAP;
	__shared__ uint32_t s_beg[9]; __shared__ int s_off[10], s_len[9];
	__shared__ double red[5 * 32];
	const int tid = threadIdx.x;
	int x_first = 0, x_last = 0;
	const NLBlock b = nl_plan(a, s_beg, s_off, s_len, x_first, x_last);
	double acc[5] = {0., 0., 0., 0., 0.};
	if (b.live) {
		for (int r = 0; r < 9; ++r) {
			const int off = s_off[r], len = s_len[r];
			const uint32_t beg = s_beg[r];
			for (int t = tid; t < len; t += NL_BT) { const double4 P = a.pos[beg + t]; xd[off + t] = P.x; yd[off + t] = P.y; zd[off + t] = P.z; }
		}
		__syncthreads();
		const uint32_t i = b.s0 + tid;
		if (i < b.s1) {
			const double4 Pi = a.pos[i];
			const int cnt = (int)a.lcount[(size_t)blockIdx.x * NL_BT + tid];
			const uint4* my = (const uint4*)(a.lists + ((size_t)blockIdx.x * (NL_LCAP / 8) * NL_BT + tid) * 8);   // 8 entries per load
			double Fx = 0., Fy = 0., Fz = 0., Vx = 0., Vy = 0., Vz = 0., u6 = 0.;
			int np_i = 0;
			// 4 pairs at a time, branch-free (invalid pairs contribute exact zeros) so that the four dependency chains overlap
			auto quad = [&](uint32_t w0, uint32_t w1, int t) {
				int k[4] = {(int)(w0 & 0xffffu), (int)(w0 >> 16), (int)(w1 & 0xffffu), (int)(w1 >> 16)};
				double dx[4], dy[4], dz[4], r2[4], sc[4], e6[4];
				bool ok[4];
#pragma unroll
				for (int u = 0; u < 4; ++u) {
					if (t + u >= cnt) k[u] = 0;
					dx[u] = Pi.x - xd[k[u]]; dy[u] = Pi.y - yd[k[u]]; dz[u] = Pi.z - zd[k[u]];
					r2[u] = dx[u] * dx[u] + dy[u] * dy[u] + dz[u] * dz[u];
					// r2 == 0: the molecule itself, or a coincident molecule of the same cell (SingleCellPolicy_ mask, Helpers.h:367-376)
					ok[u] = (t + u < cnt) && r2[u] < a.rclj2 && r2[u] != 0.0;
					r2[u] = ok[u] ? r2[u] : 1.0;
				}
#pragma unroll
				for (int u = 0; u < 4; ++u) {
					const double inv = fast_rcp(r2[u]);
					const double lj2 = a.sig2 * inv, lj4 = lj2 * lj2, lj6 = lj4 * lj2, lj12 = lj6 * lj6, lj12m6 = lj12 - lj6;
					const double scale = (a.eps24 * inv) * (lj12 + lj12m6);
					sc[u] = ok[u] ? scale : 0.0;
					e6[u] = ok[u] ? a.eps24 * lj12m6 + a.shift6 : 0.0;
				}
#pragma unroll
				for (int u = 0; u < 4; ++u) {
					const double px = dx[u] * sc[u], py = dy[u] * sc[u], pz = dz[u] * sc[u];
					Fx += px; Fy += py; Fz += pz;
					Vx += dx[u] * px; Vy += dy[u] * py; Vz += dz[u] * pz;
					u6 += e6[u];
					np_i += ok[u] ? 1 : 0;
				}
			};
			if (cnt > 0) {
				uint4 cur = my[0];
				for (int t = 0; t < cnt; t += 8) {
					uint4 nxt = cur;
					if (t + 8 < cnt) nxt = my[((t >> 3) + 1) * NL_BT];      // prefetch the next 8 entries
					quad(cur.x, cur.y, t);
					if (t + 4 < cnt) quad(cur.z, cur.w, t + 4);
					cur = nxt;
				}
			}
			double* Fp = a.F + 3 * (size_t)i; double* Vp = a.Vi + 3 * (size_t)i;
			Fp[0] = Fx; Fp[1] = Fy; Fp[2] = Fz; Vp[0] = 0.5 * Vx; Vp[1] = 0.5 * Vy; Vp[2] = 0.5 * Vz;
			acc[0] = u6; acc[2] = Vx + Vy + Vz; acc[4] = (double)np_i;
		}
	}
	block_sum<5>(acc, red);
	if (tid == 0) {
#pragma unroll
		for (int k = 0; k < 5; ++k) a.partials[(size_t)blockIdx.x * 5 + k] = acc[k];
	}
}

// periodic images follow their source molecule (frozen layout; plain +-L shift, DomainDecompBase.cpp:326-341 without the
// boundary guards, which only matter for cell assignment)
__global__ void __launch_bounds__(256) halo_refresh_kernel(Grid g, size_t n, double4* __restrict__ pos, const uint32_t* __restrict__ cell,
                                                            const uint32_t* __restrict__ halo_src, const uint8_t* __restrict__ halo_code) {
	size_t h = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (h >= n || !(cell[h] & HALO_FLAG)) return;
	const int code = halo_code[h];
	if (!code) return;
	double4 p = pos[halo_src[h]];
	const int o0 = code % 3, o1 = (code / 3) % 3, o2 = code / 9;
	if (o0) p.x = o0 == 1 ? p.x + g.box_len[0] : p.x - g.box_len[0];
	if (o1) p.y = o1 == 1 ? p.y + g.box_len[1] : p.y - g.box_len[1];
	if (o2) p.z = o2 == 1 ? p.z + g.box_len[2] : p.z - g.box_len[2];
	pos[h] = p;
}
