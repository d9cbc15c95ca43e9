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
//   row groups      = the force kernel stages the tile in three phases (less than half the shared memory, twice the
//                     resident warps).  Every group is point-symmetric about the own row: {own row + 2 opposite corner
//                     rows}, {the 4 edge rows}, {the other 2 corner rows}; wherever a molecule sits in its cell, what it
//                     loses in one row of a group it gains in the opposite one, so the lanes of a warp stay in step.  A
//                     molecule's list is grouped accordingly, each group padded to a multiple of 4 entries with index 0
//   list entry      = 16-bit index into the group's tile + NL_DUMMY; tile entries 0..NL_DUMMY-1 hold a far-away dummy
//                     position (r^2 = 3e300 fails the cutoff test), so padding needs no special case; lists[((blk*(LCAP/4) + t/4)*128 + thread)*4 + t%4]: one 8-byte
//                     load = 4 entries; lcount packs the three cumulative group ends (in units of 4 entries, 8 bits each)
#pragma once

#define NL_BT 128          // threads per block = molecules per block
#define NL_TCAP0 3072      // capacity of the build kernel's 9-row FP32 tile (candidates) to start with, multiple of 4
#define NL_TCAP 5120       // ... after a tile overflow (denser systems; fewer resident blocks)
#define NL_LCAP0 128       // list entries per molecule to start with (multiple of 4); doubled on overflow up to NL_LCAP_MAX
#define NL_LCAP_MAX 1020   // lcount packs the group ends in units of 4 entries, 8 bits each
#define NL_DUMMY 4         // dummy entries at the start of every group tile (keeps real entries 4-aligned)

// staging / scan order of the 9 neighbour rows (r = 3*(dz+1) + (dy+1)): three point-symmetric groups
__device__ __constant__ int NL_ORD[9] = {4, 0, 8, 1, 7, 3, 5, 2, 6};
__device__ __constant__ int NL_GRP[4] = {0, 3, 7, 9};      // group g = positions [NL_GRP[g], NL_GRP[g+1]) of NL_ORD


enum { NL_OVF_TILE = 1, NL_OVF_LIST = 2, NL_OVF_ROW = 4 };

struct NLArgs {
	Grid g;                     // bin grid
	const double4* pos; const uint32_t* cell; const uint32_t* cell_start;
	unsigned short* lists; uint32_t* lcount;
	unsigned long long* flags;  // [1] |= NL_OVF_*
	int bpr;                    // blocks per row
	int lcap4;                  // list capacity per molecule in units of 4 entries
	int btcap;                  // capacity of the build kernel's 9-row tile (NL_TCAP0 or NL_TCAP)
	int nyi;
	int tcap;                   // slab tile capacity of the force launch (max slab tile of the last build), multiple of 4
	float rl2f;                 // (rc + skin)^2 * (1 + margin), FP32 filter
	double eps24, sig2, shift6, rclj2;
	double *F, *Vi;
	double* partials;           // [blocks][5]
	const unsigned long long* guard; unsigned long long guard_thr;   // skip the launch when *guard > guard_thr (lists expired)
	// FUSE: leapfrog post(n) + pre(n+1) in the epilogue (NVE steps of ls1g_run)
	double dt, dt_half, mass;
	double* vel;                // [slots][3], updated in place
	double4* pos_out;           // the OTHER position buffer: new positions of the owned molecules (tiles still read pos)
	const double4* pos_build;   // positions at the last list build
	// multi-rank overlap of the halo exchange with the force pass: blocks whose tile touches no halo cell ("inner") need no
	// received data.  The build kernel classifies every block once per build; the force kernel runs as two launches:
	// phase 1 = all blocks, boundary blocks return at once; phase 2 = a compact grid over bmap (the boundary blocks).  phase 0 = one launch.
	uint32_t* bmap;             // [n_boundary] block ids of the boundary blocks (order irrelevant: every block owns its outputs)
	unsigned char* bflag;       // [blocks] 1 = boundary
	int phase, nzi;
};

struct NLBlock {
	uint32_t s0, s1;            // owned slots of this block
	int total;                  // tile entries (padded)
	bool live;
};

// Common prologue: slot range, tile plan (s_beg/s_off per row) - identical in build and force kernels.
__device__ __forceinline__ NLBlock nl_plan(const NLArgs& a, uint32_t* s_beg, int* s_off, int* s_len, int& x_first, int& x_last, const unsigned blk) {
	const Grid& g = a.g;
	const int nx = g.cpd[0], ny = g.cpd[1];
	const int row = blk / a.bpr, bi = blk % a.bpr;
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
		const int r = NL_ORD[threadIdx.x];               // s_beg / s_len / s_off are indexed by POSITION in NL_ORD
		const int cr = ((cz + r / 3 - 1) * ny + (cy + r % 3 - 1)) * nx;
		const uint32_t beg = a.cell_start[cr + x_first - 1], end = a.cell_start[cr + x_last + 2];
		s_beg[threadIdx.x] = beg; s_len[threadIdx.x] = (int)(end - beg);
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		int off = 0;
		for (int r = 0; r < 9; ++r) { s_off[r] = off; off += (s_len[r] + 3) & ~3; }
		s_off[9] = off;
	}
	__syncthreads();
	b.total = s_off[9];
	if (b.total > a.btcap) {
		if (threadIdx.x == 0) atomicOr(a.flags + 1, (unsigned long long)NL_OVF_TILE);
		b.live = false;
	}
	return b;
}

// ---------------------------------------------------------------------------------------------------------------
// build: FP32 candidate filter with radius rc + skin -> per-molecule lists of tile indices
// ---------------------------------------------------------------------------------------------------------------
#define NL_QCAP 64         // per-lane queue entries of one row group (shared memory), multiple of 4
#define NL_BUILD_SMEM(tcap) (3 * (tcap) * 4 + NL_QCAP * NL_BT * 2)

__global__ void __launch_bounds__(NL_BT) nl_build_kernel(NLArgs a) {
	extern __shared__ __align__(16) unsigned char nl_bsm[];
	float* xs = (float*)nl_bsm; float* ys = xs + a.btcap; float* zs = ys + a.btcap;
	unsigned short* q = (unsigned short*)(zs + a.btcap);          // q[t][tid]: survivors of the current row group, one column per lane
	__shared__ uint32_t s_beg[9]; __shared__ int s_off[10], s_len[9];
	const Grid& g = a.g;
	const int tid = threadIdx.x, lane = tid & 31;
	int x_first = 0, x_last = 0;
	const NLBlock b = nl_plan(a, s_beg, s_off, s_len, x_first, x_last, blockIdx.x);
	if (tid == 0 && a.bflag) {      // inner / boundary classification for the overlapped multi-rank force pass (see NLArgs)
		const int row_ = blockIdx.x / a.bpr, cy_ = row_ % a.nyi + 1, cz_ = row_ / a.nyi + 1;
		const bool boundary = b.live && (cy_ == 1 || cy_ == a.nyi || cz_ == 1 || cz_ == a.nzi || x_first <= 1 || x_last >= a.g.cpd[0] - 2);
		a.bflag[blockIdx.x] = boundary ? 1 : 0;
		if (boundary) a.bmap[atomicAdd(a.flags + 5, 1ull)] = blockIdx.x;
	}
	if (!b.live) return;
	if (tid == 0) {      // largest slab tile of this build: sizes the force kernel's shared memory
		const int m = max(max(s_off[3], s_off[7] - s_off[3]), s_off[9] - s_off[7]) + NL_DUMMY;
		if ((unsigned long long)m > *(volatile unsigned long long*)(a.flags + 2)) atomicMax(a.flags + 2, (unsigned long long)m);
	}
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
	uint2* my = (uint2*)(a.lists + ((size_t)blockIdx.x * a.lcap4 * NL_BT + tid) * 4);     // entries 4t..4t+3 at my[t * NL_BT]
	unsigned short* const q0 = q + tid;
	// the queue is addressed through one 32-bit shared-window register (append = st.shared.u16 + add, both predicated)
	const uint32_t qa0 = (uint32_t)__cvta_generic_to_shared(q0);
	const uint32_t qalim = qa0 + (NL_QCAP - 4) * NL_BT * 2;       // room for one more scan iteration
	uint32_t qa = qa0;
	int done4 = 0; bool ovf = false;       // groups of 4 entries already written to the list
	uint32_t ends = 0;
	const float rl2 = a.rl2f;
	int grp = 0;
	// queue -> list: complete quads go out as packed 8-byte stores; a mid-group flush (queue full) keeps the 0..3 left-over
	// entries at the front of the queue, the flush at the end of a row group pads them with the dummy entry
	auto flush = [&](bool final) {
		int cnt = (int)(qa - qa0) / (NL_BT * 2);
		if (final) while (cnt & 3) { q0[cnt * NL_BT] = 0; ++cnt; }
		const int n4 = cnt >> 2, rem = cnt & 3;
		if (done4 + n4 > a.lcap4) ovf = true;
		else {
			for (int t = 0; t < n4; ++t) {
				const unsigned short* e = q0 + 4 * t * NL_BT;
				uint2 w;
				w.x = (uint32_t)e[0] | ((uint32_t)e[NL_BT] << 16);
				w.y = (uint32_t)e[2 * NL_BT] | ((uint32_t)e[3 * NL_BT] << 16);
				my[(size_t)(done4 + t) * NL_BT] = w;
			}
			done4 += n4;
		}
		for (int t = 0; t < rem; ++t) q0[t * NL_BT] = q0[(4 * n4 + t) * NL_BT];
		qa = qa0 + rem * (NL_BT * 2);
	};
	for (int p = 0; p < 9; ++p) {
		const int r = NL_ORD[p];
		const int gbase = s_off[NL_GRP[grp]];          // indices are relative to the row group's tile
		const int cr = ((cz + r / 3 - 1) * ny + (cy + r % 3 - 1)) * nx;
		const uint32_t lo = a.cell_start[cr + xw_first - 1], hi = a.cell_start[cr + xw_last + 2];
		const uint32_t beg = s_beg[p];
		const int k0 = s_off[p] + (int)((lo - beg) & ~3u), k1 = s_off[p] + (int)((hi - beg + 3u) & ~3u);
		int val = k0 - gbase + NL_DUMMY;               // list entry of candidate k0
		for (int k = k0; k < k1; k += 4, val += 4) {
			const float4 cx4 = *(const float4*)(xs + k), cy4 = *(const float4*)(ys + k), cz4 = *(const float4*)(zs + k);
			const float cxa[4] = {cx4.x, cx4.y, cx4.z, cx4.w}, cya[4] = {cy4.x, cy4.y, cy4.z, cy4.w}, cza[4] = {cz4.x, cz4.y, cz4.z, cz4.w};
#pragma unroll
			for (int u = 0; u < 4; ++u) {
				const float dx = fx - cxa[u], dy = fy - cya[u], dz = fz - cza[u];
				const float r2 = dx * dx + dy * dy + dz * dz;
				if (r2 < rl2) { asm volatile("st.shared.u16 [%0], %1;" :: "r"(qa), "h"((unsigned short)(val + u)) : "memory"); qa += NL_BT * 2; }
			}
			if (qa > qalim) flush(false);
		}
		if (p + 1 == NL_GRP[grp + 1]) {                 // end of a row group: pad to a multiple of 4 with the dummy entry, flush
			flush(true);
			ends |= (uint32_t)done4 << (8 * grp);
			++grp;
		}
	}
	if (active) a.lcount[(size_t)blockIdx.x * NL_BT + tid] = ends;
	if (ovf && active) atomicOr(a.flags + 1, (unsigned long long)NL_OVF_LIST);
}

// ---------------------------------------------------------------------------------------------------------------
// force: exact FP64 test + _loopBodyLJ over the listed pairs
// ---------------------------------------------------------------------------------------------------------------
// COUNT: count the pairs inside the cutoff (pair statistics; the shift6 term of a shifted potential needs the count).
// VI:    accumulate and store the per-molecule virial Vi (FullMolecule::calcFM, FullMolecule.cpp:551-631).  Inside ls1g_run
//        nothing can read Vi between two steps, so only the last step of a run produces it; the other steps accumulate the
//        scalar virial sum(f . d) = sum(sc r^2) directly (the record of every step still gets upot and virial).
// FUSE:  the Leapfrog half-kick of this step and the half-kick + drift of the next one (integrate_kernel<POST, PRE>,
//        Leapfrog.cpp:42-68, FullMolecule::upd_postF / upd_preF) run in the epilogue, on the force still in registers:
//        v += dt/2m F; sum m v^2; v += dt/2m F; r += dt v.  F is not stored, the new position goes to the other position
//        buffer (other blocks still stage the old one), the displacement since the list build is tracked as in
//        integrate_kernel.  Saves one pass over r, v, F per step.  partials then hold 8 values per block.
//        FUSE == 2: only the half-kick of THIS step (upd_postF) - the last step of a run and the host-driven ls1g_step, where the
//        next pre-kick is a separate call; F (and Vi) are stored as well, positions stay where they are.
template <bool COUNT, bool VI, int FUSE>
__global__ void __launch_bounds__(NL_BT) nl_force_kernel(NLArgs a) {
	extern __shared__ __align__(16) double nl_tile[];
	double* tile = nl_tile;                         // [tcap][3]: x y z of candidate k at tile[3k..3k+2]
	__shared__ uint32_t s_beg[9]; __shared__ int s_off[10], s_len[9];
	__shared__ double red[(FUSE ? 8 : 5) * 32];
	const int tid = threadIdx.x;
	const unsigned blk = a.phase == 2 ? a.bmap[blockIdx.x] : blockIdx.x;
	if (a.phase == 1 && a.bflag[blk]) return;         // boundary block: runs in phase 2, after the halo positions arrived
	// speculatively enqueued step whose lists turned out expired: nothing is computed, the host rebuilds and repeats it
	if (FUSE != 1 && a.guard && *a.guard > a.guard_thr) return;
	int x_first = 0, x_last = 0;
	const NLBlock b = nl_plan(a, s_beg, s_off, s_len, x_first, x_last, blk);
	if (FUSE == 1 && a.guard && *a.guard > a.guard_thr) {      // the host swaps the two position buffers after every fused launch: carry the positions over
		const uint32_t i = b.s0 + tid;
		if (i < b.s1) a.pos_out[i] = a.pos[i];
		return;
	}
	constexpr int NV = FUSE ? 8 : 5;
	double acc[NV];
#pragma unroll
	for (int k = 0; k < NV; ++k) acc[k] = 0.;
	double disp2 = 0.;
	if (b.live) {
		const uint32_t i = b.s0 + tid;
		const bool active = i < b.s1;
		double4 Pi = make_double4(0., 0., 0., 0.);
		uint32_t ends = 0;
		if (active) { Pi = a.pos[i]; ends = a.lcount[(size_t)blk * NL_BT + tid]; }
		const uint2* my = (const uint2*)(a.lists + ((size_t)blk * a.lcap4 * NL_BT + tid) * 4);   // 4 entries per load
		double Fx = 0., Fy = 0., Fz = 0., Vx = 0., Vy = 0., Vz = 0., u6 = 0.;
		int np_i = 0;
		// 4 pairs at a time, branch-free (invalid pairs contribute exact zeros) so that the dependency chains overlap.
		// eps24 and shift6 are constants of the single component: F, V accumulate d * inv*(2 lj12 - lj6) and are scaled by
		// eps24 once; upot6 = eps24 * sum(lj12 - lj6) + shift6 * pairs.
		double s6 = 0.;
		// 0 < r2 < rc^2  <=>  (r2b - 1) <u (rc2b - 1) on the bit patterns (r2b == 0 wraps to the largest unsigned): one predicate
		const unsigned long long rc2m1 = (unsigned long long)__double_as_longlong(a.rclj2) - 1ull;
		auto quad = [&](const uint2& w) {
			const uint32_t ww[2] = {w.x, w.y};
			double dx[4], dy[4], dz[4], r2[4];
			bool ok[4];
#pragma unroll
			for (int u = 0; u < 4; ++u) {
				const int k = (int)((u & 1) ? (ww[u >> 1] >> 16) : (ww[u >> 1] & 0xffffu));
				const double* pk = tile + 3 * k;
				dx[u] = Pi.x - pk[0]; dy[u] = Pi.y - pk[1]; dz[u] = Pi.z - pk[2];
				r2[u] = dx[u] * dx[u] + dy[u] * dy[u] + dz[u] * dz[u];
				// r2 == 0: the molecule itself, or a coincident molecule of the same cell (SingleCellPolicy_ mask, Helpers.h:367-376).
				// r2 and rc^2 are non-negative doubles: their order is the order of their bit patterns, so both tests run on
				// the integer pipe and leave the FP64 pipe to the force arithmetic
				ok[u] = (unsigned long long)__double_as_longlong(r2[u]) - 1ull < rc2m1;
			}
#pragma unroll
			for (int u = 0; u < 4; ++u) {
				double inv = fast_rcp3(r2[u]);
				inv = ok[u] ? inv : 0.0;                  // rejected pair (possibly inf/nan): selected away, everything below is an exact 0
				const double lj2 = a.sig2 * inv, lj4 = lj2 * lj2, lj6 = lj4 * lj2, lj12 = lj6 * lj6, lj12m6 = lj12 - lj6;
				const double sc = inv * (lj12 + lj12m6);
				s6 += lj12m6;
				if (VI) {
					const double px = dx[u] * sc, py = dy[u] * sc, pz = dz[u] * sc;
					Fx += px; Fy += py; Fz += pz;
					Vx += dx[u] * px; Vy += dy[u] * py; Vz += dz[u] * pz;
				} else {
					Fx = fma(dx[u], sc, Fx); Fy = fma(dy[u], sc, Fy); Fz = fma(dz[u], sc, Fz);
					Vx = fma(r2[u], sc, Vx);                  // Vx carries the whole trace
				}
				if (COUNT) np_i += ok[u] ? 1 : 0;
			}
		};
		int o = 0;                                       // next group of 4 entries of this molecule
		// list entries stream from HBM: keep three loads in flight per lane, across the row-group boundaries
		const int oall = (int)((ends >> 16) & 0xffu);
		uint2 p0 = make_uint2(0, 0), p1 = p0, p2 = p0;
		if (0 < oall) p0 = my[0];
		if (1 < oall) p1 = my[NL_BT];
		if (2 < oall) p2 = my[2 * NL_BT];
		for (int gz = 0; gz < 3; ++gz) {
			if (gz) __syncthreads();                     // everyone is done with the previous row group
			const int gbase = s_off[NL_GRP[gz]];
			// dummy entries: far outside any cutoff, yet r^2 = 3e300 stays FINITE (the trace accumulation multiplies r^2 by an exact 0)
			if (tid < NL_DUMMY) { double* q = tile + 3 * tid; q[0] = 1e150; q[1] = 1e150; q[2] = 1e150; }
			for (int r = NL_GRP[gz]; r < NL_GRP[gz + 1]; ++r) {
				const int off = s_off[r] - gbase + NL_DUMMY, len = s_len[r];
				const uint32_t beg = s_beg[r];
				for (int t = tid; t < len; t += NL_BT) { const double4 P = a.pos[beg + t]; double* q = tile + 3 * (off + t); q[0] = P.x; q[1] = P.y; q[2] = P.z; }
			}
			__syncthreads();
			const int oend = (int)((ends >> (8 * gz)) & 0xffu);
			for (; o < oend; ++o) {
				const uint2 cur = p0;
				p0 = p1; p1 = p2;
				if (o + 3 < oall) p2 = my[(o + 3) * NL_BT];
				quad(cur);
			}
		}
		if (active) {
			Fx *= a.eps24; Fy *= a.eps24; Fz *= a.eps24; Vx *= a.eps24; Vy *= a.eps24; Vz *= a.eps24;
			u6 = a.eps24 * s6;       // + shift6 * pairs: added by the caller's finalize when the potential is shifted (COUNT)
			if (COUNT) u6 += a.shift6 * (double)np_i;
			if (FUSE == 2) {
				const double m = a.mass, dtInv2m = a.dt_half / m;         // upd_postF
				double* vp = a.vel + 3 * (size_t)i;
				double vx = vp[0], vy = vp[1], vz = vp[2];
				vx += dtInv2m * Fx; vy += dtInv2m * Fy; vz += dtInv2m * Fz;
				double v2 = 0.;
				v2 += vx * vx; v2 += vy * vy; v2 += vz * vz;
				acc[5] = m * v2; acc[6] = 1.0;
				vp[0] = vx; vp[1] = vy; vp[2] = vz;
				double* Fp = a.F + 3 * (size_t)i;
				Fp[0] = Fx; Fp[1] = Fy; Fp[2] = Fz;
				if (VI) { double* Vp = a.Vi + 3 * (size_t)i; Vp[0] = 0.5 * Vx; Vp[1] = 0.5 * Vy; Vp[2] = 0.5 * Vz; }
			} else if (FUSE == 1) {
				const double m = a.mass;
				double* vp = a.vel + 3 * (size_t)i;
				double vx = vp[0], vy = vp[1], vz = vp[2];       // (requesting these before the pair loops costs 16 registers = 2 resident blocks: slower)
				{	// upd_postF
					const double dtInv2m = a.dt_half / m;
					vx += dtInv2m * Fx; vy += dtInv2m * Fy; vz += dtInv2m * Fz;
					double v2 = 0.;
					v2 += vx * vx; v2 += vy * vy; v2 += vz * vz;
					acc[5] = m * v2; acc[6] = 1.0;
				}
				{	// upd_preF
					const double dt_halve = .5 * a.dt, dtInv2m = dt_halve / m;
					vx += dtInv2m * Fx; vy += dtInv2m * Fy; vz += dtInv2m * Fz;
					Pi.x += a.dt * vx; Pi.y += a.dt * vy; Pi.z += a.dt * vz;
				}
				vp[0] = vx; vp[1] = vy; vp[2] = vz;
				a.pos_out[i] = Pi;
				const double4 bp = a.pos_build[i];
				const double ex = Pi.x - bp.x, ey = Pi.y - bp.y, ez = Pi.z - bp.z;
				disp2 = ex * ex + ey * ey + ez * ez;
			} else {
				double* Fp = a.F + 3 * (size_t)i;
				Fp[0] = Fx; Fp[1] = Fy; Fp[2] = Fz;
				if (VI) { double* Vp = a.Vi + 3 * (size_t)i; Vp[0] = 0.5 * Vx; Vp[1] = 0.5 * Vy; Vp[2] = 0.5 * Vz; }
			}
			acc[0] = u6; acc[2] = Vx + Vy + Vz; acc[4] = (double)np_i;
		}
	}
	if (FUSE == 1) {          // displacement criterion of the neighbour lists, as in integrate_kernel
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) disp2 = fmax(disp2, __shfl_xor_sync(0xffffffffu, disp2, o));
		const unsigned long long bits = (unsigned long long)__double_as_longlong(disp2);
		if ((tid & 31) == 0 && bits > *(volatile unsigned long long*)a.flags) atomicMax(a.flags, bits);
	}
	block_sum<NV>(acc, red);
	if (tid == 0) {
#pragma unroll
		for (int k = 0; k < NV; ++k) a.partials[(size_t)blk * NV + k] = acc[k];
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
