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
// Work decomposition (both kernels, identical so that tile indices stay valid), template parameter R:
//   block           = R = 1: 128 consecutive owned slots of one cell row (y, z), 128 threads.
//                     R = 4 (default): a 2 x 2 bundle of cell rows (2gy+1.., 2gz+1..) holding M molecules is cut into
//                     nbk = max(ceil(M / 508), ceil(cells in x / 8)) blocks of 512 threads; block j owns the j-th nbk-th OF EVERY ROW (slots
//                     [j n_r / nbk, (j+1) n_r / nbk) of row r): rows of different density - the lanes of an fcc lattice - still
//                     line up in x as long as each row is about homogeneous along x.  lane = one molecule, rows back to back.
//                     The 4 x 4 rows around a bundle are 16 tile rows for ~500 molecules where four single-row blocks stage 36:
//                     less than half the shared memory and staging traffic per molecule, which is what lets 32 warps stay
//                     resident with the WHOLE tile on chip.  Where the rows do not line up the tile outgrows its capacity, the
//                     build reports it and the host falls back to R = 1
//   tile            = the 9 (16) neighbouring rows cut to cells [x_first-1, x_last+1] of the block's molecules: contiguous slot
//                     ranges.  The force kernel keeps the whole tile in shared memory as two arrays, Z[k] (8 B) and XY[k] (16 B):
//                     one LDS.64 + one LDS.128 per listed partner, entry k of both addressed from one 16-bit list entry.  The tile
//                     is staged with cp.async (LDGSTS: global -> shared without registers) in two batches, each signalling an
//                     mbarrier; a warp waits for the data it is about to read, never for other warps: no __syncthreads between
//                     the prologue and the final block sum, and the second batch lands while the first is being evaluated
//   list entry      = 16 bits, 8 * (tile index + NL_DUMMY) = byte offset of the partner in Z (twice that in XY); tile entries
//                     0..NL_DUMMY-1 hold a far-away dummy position (r^2 = 3e300 fails the cutoff test), so padding needs no
//                     special case.  lists[((blk*(LCAP/4) + t/4)*BT + thread)*4 + t%4]: one 8-byte load = 4 entries, padded
//                     to a multiple of 4 with entry 0.  The molecule itself is not in its list.  Order: tile rows in staging
//                     order, x ascending, then every chunk of 16 list positions re-ordered for the shared-memory banks (see the
//                     build kernel's emit).  lcount = quads of the molecule | quads that lie entirely in batch A << 16
#pragma once

#define NL_BT 128          // owned slots per cell row and block
#define NL_TCAP0 3072      // capacity of the build kernel's FP32 tile (candidates) to start with, multiple of 4
#define NL_TCAP 5120       // ... after a tile overflow (denser systems; fewer resident blocks)
#define NL_LCAP0 128       // list entries per molecule to start with (multiple of 4); doubled on overflow up to NL_LCAP_MAX
#define NL_LCAP_MAX 1020   // lcount packs quad counts in 16 bits
#define NL_DUMMY 4         // dummy entries at the start of the tile (keeps real entries 4-aligned)
#define NL_MAXROWS 16
#define NL_F_OVF 7         // d_flags word that collects the overflow bits of a build (ls1g_internal.cuh: d_flags)
#ifndef NL_R4_BT
#define NL_R4_BT 512       // R = 4: threads per block
#endif
#define NL_BUNDLE_FILL (NL_R4_BT - 4) // R = 4: molecules per block a bundle is cut for (4 rows x integer rounding stay within the block)
#define NL_XSPAN (NL_R4_BT / 64)      // R = 4: a block spans at most about this many cells in x (bundles at the upper faces hold fewer than 4 rows)

template <int R> struct NLGeom;
template <> struct NLGeom<1> { static constexpr int BT = NL_BT, NR = 9, ROWS_A = 3, QCAP = 64; };
template <> struct NLGeom<4> { static constexpr int BT = NL_R4_BT, NR = 16, ROWS_A = 8, QCAP = NL_R4_BT >= 512 ? 32 : 64; };

// R = 1: staging / scan order of the 9 neighbour rows (r = 3*(dz+1) + (dy+1)): own row first, then point-symmetric pairs
__device__ __constant__ int NL_ORD[9] = {4, 0, 8, 1, 7, 3, 5, 2, 6};

enum { NL_OVF_TILE = 1, NL_OVF_LIST = 2, NL_OVF_ROW = 4 };

struct NLArgs {
	Grid g;                     // bin grid
	const double4* pos; const uint32_t* cell; const uint32_t* cell_start;
	unsigned short* lists; uint32_t* lcount;
	struct NLPlan* plans;       // [blocks] the plan of every block as the build kernel computed it (the layout is frozen until the next build)
	unsigned long long* flags;  // [1] |= NL_OVF_*
	int bpr;                    // blocks per row (R = 4: per row bundle)
	int lcap4;                  // list capacity per molecule in units of 4 entries
	int order;                  // 1: bank-aware order of every list chunk (build kernel), 0: tile order
	int btcap;                  // capacity of the build kernel's tile (NL_TCAP0 or NL_TCAP)
	int nyi, nzi;               // owned rows per dimension
	int nyg;                    // R = 4: row bundles in y = ceil(nyi / 2)
	int tcap;                   // tile capacity of the force launch (largest tile of the last build), multiple of 4
	float rl2f;                 // (rc + skin)^2 * (1 + margin), FP32 filter
	double eps24, sig2, shift6, rclj2;
	double *F, *Vi;
	double* partials;           // [blocks][5 or 8]
	const unsigned long long* guard; unsigned long long guard_thr;   // skip the launch when *guard > guard_thr (lists expired)
	// FUSE: leapfrog post(n) + pre(n+1) in the epilogue (NVE steps of ls1g_run)
	double dt, dt_half, mass;
	double* vel;                // [slots][3], updated in place
	double4* pos_out;           // the OTHER position buffer: new positions of the owned molecules (tiles still read pos)
	const double4* pos_build;   // positions at the last list build
	const double4* pos_prune;   // positions at the last prune of the inner lists (nullptr: no inner lists); tracked into flags[1]
	// nl_prune_kernel: outer lists (lists, lcount) -> inner lists
	unsigned short* lists_out; uint32_t* lcount_out; double4* pos_prune_out;
	double rin2;                // (cutoff + inner skin)^2
	// multi-rank overlap of the halo exchange with the force pass: blocks whose tile touches no halo cell ("inner") need no
	// received data.  The build kernel classifies every block once per build; the force kernel runs as two launches:
	// phase 1 = a compact grid over the inner blocks; phase 2 = a compact grid over the boundary blocks.  phase 0 = one launch over all.
	uint32_t* bmap;             // [blocks] block ids: the boundary blocks from the front, the inner ones from the back (order irrelevant:
	                            // every block owns its outputs)
	unsigned nb_total;          // blocks of the build
	unsigned char* bflag;       // [blocks] 1 = boundary
	int phase;
};

// Shared-memory plan of one block, identical in build and force kernels
struct NLPlan {
	uint32_t beg[NL_MAXROWS];   // first slot of tile row p
	int len[NL_MAXROWS];        // slots of tile row p (0: row outside the grid or next to no live owned row)
	int off[NL_MAXROWS + 1];    // tile index of its first entry (rows padded to multiples of 4)
	uint32_t s0[4], s1[4];      // owned slots of owned row q (s0 == s1: none)
	int o[5];                   // first thread of owned row q (rows back to back); o[R] = molecules of the block
	int xf[4], xl[4];           // first / last cell (x) of those slots
	int ty[4], tz[4];           // cell row (y, z) of owned row q
	int x_first, x_last, live, boundary;
};

// tile row p of block (gy, gz) -> cell row (y, z), or y < 0 when outside the grid
template <int R>
__device__ __forceinline__ void nl_tile_row(const NLArgs& a, int gy, int gz, int p, int& y, int& z) {
	if (R == 1) { const int r = NL_ORD[p]; y = gy + r % 3 - 1; z = gz + r / 3 - 1; }
	else { y = 2 * gy + (p & 3); z = 2 * gz + (p >> 2); if (y > a.g.cpd[1] - 1 || z > a.g.cpd[2] - 1) y = -1; }
}

// Common prologue: owned slot ranges, tile plan.  All threads call it; three __syncthreads.
template <int R>
__device__ __forceinline__ void nl_plan(const NLArgs& a, NLPlan& P, const unsigned blk) {
	const Grid& g = a.g;
	const int nx = g.cpd[0], ny = g.cpd[1];
	const int grp = blk / a.bpr, bi = blk % a.bpr;
	// R = 1: gy, gz = the cell row itself; R = 4: bundle coordinates, owned rows (2gy+1+(q&1), 2gz+1+(q>>1))
	const int gy = R == 1 ? grp % a.nyi + 1 : grp % a.nyg, gz = R == 1 ? grp / a.nyi + 1 : grp / a.nyg;
	const int tid = threadIdx.x;
	if (tid < R) {
		const int q = tid;
		const int cy = R == 1 ? gy : 2 * gy + 1 + (q & 1), cz = R == 1 ? gz : 2 * gz + 1 + (q >> 1);
		uint32_t s0 = 0, s1 = 0; int xf = 0x7fffffff, xl = -1;
		if (R == 1) {
			const int crow = (cz * ny + cy) * nx;
			const uint32_t rs = a.cell_start[crow + 1], re = a.cell_start[crow + nx - 1];
			s0 = rs + (uint32_t)bi * NL_BT; s1 = min(s0 + NL_BT, re);
			if (bi == a.bpr - 1 && s0 + NL_BT < re) atomicOr(a.flags + NL_F_OVF, (unsigned long long)NL_OVF_ROW);
			if (s0 >= re) s1 = s0;
		} else {
			// every thread q sums the whole bundle (4 x 2 loads) so that no barrier separates M from the cut
			uint32_t M = 0, rs = 0, n = 0;
#pragma unroll
			for (int r = 0; r < 4; ++r) {
				const int ry = 2 * gy + 1 + (r & 1), rz = 2 * gz + 1 + (r >> 1);
				if (ry > a.nyi || rz > a.nzi) continue;
				const int crow = (rz * ny + ry) * nx;
				const uint32_t b0 = a.cell_start[crow + 1], b1 = a.cell_start[crow + nx - 1];
				M += b1 - b0;
				if (r == q) { rs = b0; n = b1 - b0; }
			}
			const uint32_t nbk = max(max(1u, (M + NL_BUNDLE_FILL - 1) / NL_BUNDLE_FILL), (uint32_t)(nx - 2 + NL_XSPAN - 1) / NL_XSPAN);
			if ((uint32_t)bi < nbk) {
				s0 = rs + (uint32_t)((unsigned long long)bi * n / nbk);
				s1 = rs + (uint32_t)((unsigned long long)(bi + 1) * n / nbk);
			}
			if (q == 0 && nbk > (uint32_t)a.bpr) atomicOr(a.flags + NL_F_OVF, (unsigned long long)NL_OVF_ROW);
		}
		if (s0 < s1) {
			xf = (int)((a.cell[s0] & ~HALO_FLAG) % (uint32_t)nx);
			xl = (int)((a.cell[s1 - 1] & ~HALO_FLAG) % (uint32_t)nx);
		}
		P.s0[q] = s0; P.s1[q] = s1; P.xf[q] = xf; P.xl[q] = xl; P.ty[q] = cy; P.tz[q] = cz;
	}
	__syncthreads();
	int x_first = 0x7fffffff, x_last = -1;
#pragma unroll
	for (int q = 0; q < R; ++q) { x_first = min(x_first, P.xf[q]); x_last = max(x_last, P.xl[q]); }
	const bool live = x_last >= 0;
	if (tid < NLGeom<R>::NR) {
		int y, z;
		nl_tile_row<R>(a, gy, gz, tid, y, z);
		uint32_t beg = 0; int len = 0;
		bool need = live && y >= 0;
		if (need && R > 1) {          // only rows next to an owned row that has molecules in this block
			need = false;
#pragma unroll
			for (int q = 0; q < R; ++q) if (P.xl[q] >= 0 && abs(P.ty[q] - y) <= 1 && abs(P.tz[q] - z) <= 1) need = true;
		}
		if (need) {
			const int cr = (z * ny + y) * nx;
			beg = a.cell_start[cr + x_first - 1];
			len = (int)(a.cell_start[cr + x_last + 2] - beg);
		}
		P.beg[tid] = beg; P.len[tid] = len;
	}
	__syncthreads();
	if (tid == 0) {
		int off = 0;
		for (int r = 0; r < NLGeom<R>::NR; ++r) { P.off[r] = off; off += (P.len[r] + 3) & ~3; }
		P.off[NLGeom<R>::NR] = off;
		P.x_first = x_first; P.x_last = x_last;
		P.o[0] = 0;
		for (int q = 0; q < R; ++q) P.o[q + 1] = P.o[q] + (int)(P.s1[q] - P.s0[q]);
		bool lv = live;
		if (lv && off > a.btcap) { atomicOr(a.flags + NL_F_OVF, (unsigned long long)NL_OVF_TILE); lv = false; }
		P.live = lv ? 1 : 0;
		// a block is "boundary" when its tile reaches into a halo cell (the positions there arrive with the halo refresh)
		const int ylo = R == 1 ? gy - 1 : 2 * gy, yhi = R == 1 ? gy + 1 : min(2 * gy + 3, ny - 1);
		const int zlo = R == 1 ? gz - 1 : 2 * gz, zhi = R == 1 ? gz + 1 : min(2 * gz + 3, g.cpd[2] - 1);
		P.boundary = lv && (ylo == 0 || zlo == 0 || yhi == ny - 1 || zhi == g.cpd[2] - 1 || x_first <= 1 || x_last >= nx - 2);
	}
	__syncthreads();
}

// thread -> (owned row q, slot i); has = the thread has a molecule
template <int R>
__device__ __forceinline__ void nl_my_slot(const NLPlan& P, int tid, int& q, uint32_t& i, bool& has) {
	if (R == 1) { q = 0; i = P.s0[0] + (uint32_t)tid; has = i < P.s1[0]; }
	else {
		q = (tid >= P.o[1]) + (tid >= P.o[2]) + (tid >= P.o[3]);
		i = P.s0[q] + (uint32_t)(tid - P.o[q]);
		has = tid < P.o[4];
	}
}

// ---------------------------------------------------------------------------------------------------------------
// build: FP32 candidate filter with radius rc + skin -> per-molecule lists of tile indices
// ---------------------------------------------------------------------------------------------------------------
#define NL_BUILD_SMEM(R, tcap) (3 * (tcap) * 4 + NLGeom<R>::QCAP * NLGeom<R>::BT * 2)

template <int R>
__global__ void __launch_bounds__(NLGeom<R>::BT) nl_build_kernel(NLArgs a) {
	constexpr int BT = NLGeom<R>::BT, NR = NLGeom<R>::NR, QCAP = NLGeom<R>::QCAP;
	extern __shared__ __align__(16) unsigned char nl_bsm[];
	// candidates in groups of four: [x0..x3 | y0..y3 | z0..z3], 48 bytes per group (three 16-byte loads off one address register)
	float* cs = (float*)nl_bsm;
	unsigned short* q = (unsigned short*)(cs + 3 * a.btcap);      // q[t][tid]: queued survivors, one column per lane
	__shared__ NLPlan P;
	const Grid& g = a.g;
	const int tid = threadIdx.x;
	nl_plan<R>(a, P, blockIdx.x);
	if (tid == 0 && a.bflag) {      // inner / boundary classification for the overlapped multi-rank force pass (see NLArgs)
		a.bflag[blockIdx.x] = P.boundary ? 1 : 0;
		// boundary blocks fill bmap from the front, the others from the back: two compact grids out of one array
		if (P.boundary) a.bmap[atomicAdd(a.flags + 5, 1ull)] = blockIdx.x;
		else a.bmap[gridDim.x - 1 - (unsigned)atomicAdd(a.flags + 4, 1ull)] = blockIdx.x;
	}
	// the force kernel loads the plan instead of re-deriving it (three dependent global loads behind three barriers)
	if (tid < (int)(sizeof(NLPlan) / 4)) ((uint32_t*)(a.plans + blockIdx.x))[tid] = ((const uint32_t*)&P)[tid];
	if (!P.live) return;
	if (tid == 0) {      // largest tile of this build: sizes the force kernel's shared memory
		const int m = P.off[NR] + NL_DUMMY;
		if ((unsigned long long)m > *(volatile unsigned long long*)(a.flags + 2)) atomicMax(a.flags + 2, (unsigned long long)m);
	}
	const int nx = g.cpd[0], ny = g.cpd[1];
	const int x_first = P.x_first, x_last = P.x_last;
	// origin of the FP32 coordinates: the block centre
	const double ox = g.halo_min[0] + 0.5 * (x_first + x_last + 1) * g.cell_len[0];
	const double oy = g.halo_min[1] + (R == 1 ? P.ty[0] + 0.5 : P.ty[0] + 1.0) * g.cell_len[1];
	const double oz = g.halo_min[2] + (R == 1 ? P.tz[0] + 0.5 : P.tz[0] + 1.0) * g.cell_len[2];
	for (int r = 0; r < NR; ++r) {
		const int off = P.off[r], len = P.len[r], padded = (len + 3) & ~3;
		const uint32_t beg = P.beg[r];
		for (int t = tid; t < padded; t += BT) {
			float x = 1e30f, y = 1e30f, z = 1e30f;
			if (t < len) { const double4 Q = a.pos[beg + t]; x = (float)(Q.x - ox); y = (float)(Q.y - oy); z = (float)(Q.z - oz); }
			float* c4 = cs + ((off + t) >> 2) * 12 + ((off + t) & 3);
			c4[0] = x; c4[4] = y; c4[8] = z;
		}
	}
	__syncthreads();
	// this thread's molecule; a warp may hold molecules of more than one owned row (R = 4: rows back to back)
	int qr; uint32_t i; bool active;
	nl_my_slot<R>(P, tid, qr, i, active);
	unsigned todo = __ballot_sync(0xffffffffu, active);
	if (!todo) return;
	float fx = -1e30f, fy = -1e30f, fz = -1e30f;
	if (active) { const double4 Pi = a.pos[i]; fx = (float)(Pi.x - ox); fy = (float)(Pi.y - oy); fz = (float)(Pi.z - oz); }
	uint2* my = (uint2*)(a.lists + ((size_t)blockIdx.x * a.lcap4 * BT + tid) * 4);     // entries 4t..4t+3 at my[t * BT]
	unsigned short* const q0 = q + tid;
	// the queue is addressed through one 32-bit shared-window register (append = st.shared.u16 + add, both predicated)
	const uint32_t qa0 = (uint32_t)__cvta_generic_to_shared(q0), cs_a = (uint32_t)__cvta_generic_to_shared(cs);
	const uint32_t qalim = qa0 + (QCAP - 4) * BT * 2;       // room for one more scan iteration
	uint32_t qa = qa0;
	int done4 = 0; bool ovf = false;       // groups of 4 entries already written to the list
	int nA4 = 0;                           // quads that only hold entries of staging batch A
	float rmin = 1e30f;                    // a partner closer than 1e-4 (possibly coincident): the force kernel keeps its r2 == 0 test
	const float rl2 = a.rl2f;
	// queue -> list, in CHUNKS of up to 16 entries = 16 consecutive list positions, and in a bank-aware order.  The force kernel reads
	// one tile entry per lane and list position: a random 8-byte (Z) and 16-byte (XY) shared-memory read per lane, ~13 wavefronts
	// per warp and position in tile order against 6 without conflicts - more than the FP64 pipe needs for the same position.  The
	// ORDER of a molecule's list is free, so lane l wants at list position t an entry whose tile index is = (l + t) mod 16: then the
	// 16 lanes of a half-warp hit 16 different 8-byte banks of Z and the 8 lanes of a quarter-warp 8 different 16-byte banks of XY.
	// The 16 positions of a chunk want every residue once, so a chunk is a 16-slot hash table (rows 0..15 of the lane's queue
	// column, the chunk's entries wait in registers, a 16-bit mask tracks the free slots): an entry goes to the slot that wants its
	// residue, else to the slot 8 further (keeps XY conflict-free), else to the first free slot.  A fixed function of the build: runs stay bitwise reproducible.
	// The whole warp flushes together (mode 0: some lane's queue is nearly full -> every lane with >= 16 entries emits one chunk;
	// mode 1: end of staging batch A -> all complete quads; mode 2: end of the list -> everything, padded with the dummy entry 0).
	// (the appends are asm statements without a memory clobber, so that the scan loop keeps its addresses in registers; the clobbers
	// here order them against the queue reads and writes of the flush)
	const int order = a.order;
	auto emit = [&](const int n) {       // n <= 16 entries from the queue front -> (n + 3) / 4 quads of the list
		const int n4 = (n + 3) >> 2, np = 4 * n4;
		if (done4 + n4 > a.lcap4) { ovf = true; return; }
		uint32_t e[16];
#pragma unroll
		for (int j = 0; j < 16; ++j) e[j] = j < n ? (uint32_t)q0[j * BT] : 0u;
		const uint32_t base = (uint32_t)(tid + 4 * done4);      // lane + list position of slot 0
		uint32_t free_slots = np >= 16 ? 0xffffu : (1u << np) - 1u;            // bit s: slot s of the chunk is still empty
#pragma unroll
		for (int j = 0; j < 16; ++j) {
			if (j < n) {
				uint32_t s = ((e[j] >> 3) - base) & 15u;
				if (!(free_slots >> s & 1u)) s ^= 8u;
				if (!order || !(free_slots >> s & 1u)) s = (uint32_t)__ffs((int)free_slots) - 1u;
				free_slots &= ~(1u << s);
				q0[s * BT] = (unsigned short)e[j];
			}
		}
		while (free_slots) { q0[(__ffs((int)free_slots) - 1) * BT] = 0; free_slots &= free_slots - 1u; }      // final chunk: the pad entries
		for (int t = 0; t < n4; ++t) {
			const unsigned short* w4 = q0 + 4 * t * BT;
			uint2 w;
			w.x = (uint32_t)w4[0] | ((uint32_t)w4[BT] << 16);
			w.y = (uint32_t)w4[2 * BT] | ((uint32_t)w4[3 * BT] << 16);
			my[(size_t)(done4 + t) * BT] = w;
		}
		done4 += n4;
	};
	auto flush = [&](const int mode) {
		asm volatile("" ::: "memory");
		int cnt = (int)(qa - qa0) / (BT * 2);
		for (;;) {
			const int n = cnt >= 16 ? 16 : (mode == 0 ? 0 : (mode == 1 ? (cnt & ~3) : cnt));
			if (!__any_sync(0xffffffffu, n > 0)) break;
			if (n > 0) {
				emit(n);
				const int rem = cnt - n;
				for (int j = 0; j < rem; ++j) q0[j * BT] = q0[(n + j) * BT];
				cnt = rem;
			}
			if (mode == 0) break;
		}
		qa = qa0 + cnt * (BT * 2);
		asm volatile("" ::: "memory");
	};
	const int gy = R == 1 ? P.ty[0] : (P.ty[0] - 1) / 2, gz = R == 1 ? P.tz[0] : (P.tz[0] - 1) / 2;
	while (todo) {      // one pass per owned row present in this warp (warp-uniform control flow)
		const int qq = __shfl_sync(0xffffffffu, qr, __ffs(todo) - 1);
		const unsigned grp = __ballot_sync(0xffffffffu, active && qr == qq);
		todo &= ~grp;
		const bool mine = active && qr == qq;
		// the cells the lanes of this row span (slots ascend with the lane)
		const uint32_t iw0 = __shfl_sync(0xffffffffu, i, __ffs(grp) - 1), iw1 = __shfl_sync(0xffffffffu, i, 31 - __clz(grp));
		const int xw_first = (int)((a.cell[iw0] & ~HALO_FLAG) % (uint32_t)nx), xw_last = (int)((a.cell[iw1] & ~HALO_FLAG) % (uint32_t)nx);
		const int cy = P.ty[qq], cz = P.tz[qq];
		const float gx = mine ? fx : -1e30f, gyf = mine ? fy : -1e30f, gzf = mine ? fz : -1e30f;
		float nx1 = -gx, ny1 = -gyf, nz1 = -gzf;
		asm volatile("" : "+f"(nx1), "+f"(ny1), "+f"(nz1));      // opaque: kept in registers instead of being re-derived in every scan iteration
		const float2 ngx = make_float2(nx1, nx1), ngy = make_float2(ny1, ny1), ngz = make_float2(nz1, nz1);
		for (int p = 0; p < NR; ++p) {
			int y, z;
			nl_tile_row<R>(a, gy, gz, p, y, z);
			if (y >= 0 && abs(y - cy) <= 1 && abs(z - cz) <= 1 && P.len[p] > 0) {
				const int cr = (z * ny + y) * nx;
				const uint32_t lo = a.cell_start[cr + xw_first - 1], hi = a.cell_start[cr + xw_last + 2];
				const uint32_t beg = P.beg[p];
				const int k0 = P.off[p] + (int)((lo - beg) & ~3u), k1 = P.off[p] + (int)((hi - beg + 3u) & ~3u);
				int val = (k0 + NL_DUMMY) * 8;                 // list entry of candidate k0
				const bool own = y == cy && z == cz;         // the row that holds the lanes' own molecules (warp-uniform)
				const int selfv = (own && mine) ? (P.off[p] + (int)(i - beg) + NL_DUMMY) * 8 : -1;      // the molecule itself
				// OWN = false (8 of 9 rows): no candidate is the molecule itself, the index compare drops out of the loop
				auto scan = [&](auto own_row) {
					constexpr bool OWN = decltype(own_row)::value;
					// candidate groups through one 32-bit shared-window address register (3 loads off it, one add per iteration)
					for (uint32_t ca = cs_a + 12u * (uint32_t)k0, ce = cs_a + 12u * (uint32_t)k1; ca < ce; ca += 48u, val += 32) {
						float4 cx4, cy4, cz4;
						asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(cx4.x), "=f"(cx4.y), "=f"(cx4.z), "=f"(cx4.w) : "r"(ca));
						asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+16];" : "=f"(cy4.x), "=f"(cy4.y), "=f"(cy4.z), "=f"(cy4.w) : "r"(ca));
						asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+32];" : "=f"(cz4.x), "=f"(cz4.y), "=f"(cz4.z), "=f"(cz4.w) : "r"(ca));
						// two candidates per instruction: the packed FP32 pipe of sm_100 (FADD2 / FMUL2 / FFMA2); the filter is issue-bound
						const float2 dxa = __fadd2_rn(make_float2(cx4.x, cx4.y), ngx), dxb = __fadd2_rn(make_float2(cx4.z, cx4.w), ngx);
						const float2 dya = __fadd2_rn(make_float2(cy4.x, cy4.y), ngy), dyb = __fadd2_rn(make_float2(cy4.z, cy4.w), ngy);
						const float2 dza = __fadd2_rn(make_float2(cz4.x, cz4.y), ngz), dzb = __fadd2_rn(make_float2(cz4.z, cz4.w), ngz);
						const float2 ra = __ffma2_rn(dza, dza, __ffma2_rn(dya, dya, __fmul2_rn(dxa, dxa)));
						const float2 rb = __ffma2_rn(dzb, dzb, __ffma2_rn(dyb, dyb, __fmul2_rn(dxb, dxb)));
						float r2a[4] = {ra.x, ra.y, rb.x, rb.y};
#pragma unroll
						for (int u = 0; u < 4; ++u) {
							if (OWN && val + 8 * u == selfv) r2a[u] = 1e30f;
							if (r2a[u] < rl2) { asm volatile("st.shared.u16 [%0], %1;" :: "r"(qa), "h"((unsigned short)(val + 8 * u))); qa += BT * 2; }
						}
						rmin = fminf(fminf(rmin, r2a[0]), fminf(fminf(r2a[1], r2a[2]), r2a[3]));
						if (__any_sync(0xffffffffu, qa > qalim)) flush(0);
					}
				};
				if (own) scan(std::true_type{}); else scan(std::false_type{});
			}
			if (p + 1 == NLGeom<R>::ROWS_A) {      // everything listed so far lies in batch A: its complete quads go out now
				flush(1);
				if (mine) nA4 = done4;
			}
		}
	}
	flush(2);
	if (active) a.lcount[(size_t)blockIdx.x * BT + tid] = (uint32_t)done4 | ((uint32_t)nA4 << 16);
	if (ovf && active) atomicOr(a.flags + NL_F_OVF, (unsigned long long)NL_OVF_LIST);
	if (rmin < 1e-8f && active) a.flags[6] = 1ull;
}

// ---------------------------------------------------------------------------------------------------------------
// permute: bank-aware order of every molecule's list
// ---------------------------------------------------------------------------------------------------------------
// The force kernel reads one tile entry per lane and list position: a random 8-byte (Z) and 16-byte (XY) shared-memory read per
// lane.  In tile order these cost ~13 wavefronts per warp and position (conflict-free: 6), and the shared-memory pipe, not the FP64
// pipe, bounds the kernel (ncu: 80 % busy, profiles/r02a_ncu_nl_force_bundled_10M_summary.csv).  The ORDER of a molecule's list is
// free, so after the build every lane re-orders its entries: at list position t lane l wants an entry whose tile index is
// = (l + t) mod 16 - then the 16 lanes of a half-warp hit 16 different 8-byte banks of Z and the 8 lanes of a quarter-warp 8
// different 16-byte banks of XY.  Greedy per lane: the wanted residue if the lane still has such an entry, else the residue 8
// further (keeps XY conflict-free), else from its fullest bucket.  ~7.9 wavefronts per position on a liquid (simulation), i.e. below
// the 10 SM-cycles the FP64 pipe needs.  Entries of staging batch A stay in front of the others (lcount's second field), the pad
// entries at the end.  The result is a fixed function of the build, so runs stay bitwise reproducible.
// Working arrays in shared memory, one column per thread ([index][thread]: a dynamically indexed access of a warp is at most a
// 2-way bank conflict; the same arrays in local memory cost 32 wavefronts per access): the bucket-sorted entries and, per segment
// (0 = batch A, 1 = the rest) and residue, the read position and the end of the bucket.  The list is read from global memory twice
// (count, scatter) and written once.
#define NL_PERM_SMEM(CAP, TPB) ((size_t)(CAP) * (TPB) * 2 + 2 * 32 * (TPB) * 2)
template <int CAP, int TPB>
__global__ void __launch_bounds__(TPB) nl_permute_kernel(unsigned short* __restrict__ lists, const uint32_t* __restrict__ lcount, int lcap4, int BT, size_t n_lanes) {
	extern __shared__ __align__(16) unsigned short nl_psm[];
	unsigned short* srt = nl_psm + threadIdx.x;                       // srt[t * TPB]
	unsigned short* pos = nl_psm + (size_t)CAP * TPB + threadIdx.x;   // pos[(16 seg + b) * TPB]
	unsigned short* end = pos + 32 * TPB;
	const size_t gl = (size_t)blockIdx.x * TPB + threadIdx.x;         // lane = blk * BT + tid
	if (gl >= n_lanes) return;
	const uint32_t cnt = lcount[gl];
	const int n4 = (int)(cnt & 0xffffu), nA4 = min((int)(cnt >> 16), n4);
	if (n4 == 0 || 4 * n4 > CAP) return;
	const size_t blk = gl / (size_t)BT; const int tid = (int)(gl % (size_t)BT);
	uint2* my = (uint2*)(lists + ((size_t)blk * lcap4 * BT + tid) * 4);
	// pass 1: bucket sizes per segment (pad entries are 0 and are not counted)
#pragma unroll
	for (int k = 0; k < 32; ++k) pos[k * TPB] = 0;
	for (int o = 0; o < n4; ++o) {
		const uint2 w = my[(size_t)o * BT];
		const int sb = o < nA4 ? 0 : 16;
		const uint32_t ee[4] = {w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16};
#pragma unroll
		for (int u = 0; u < 4; ++u) if (ee[u]) ++pos[(sb + ((ee[u] >> 3) & 15)) * TPB];
	}
	int nrealA = 0, nrealB = 0;
	{
		int run = 0;
#pragma unroll
		for (int k = 0; k < 32; ++k) {
			if (k == 16) { nrealA = run; run = 4 * nA4; }
			const int c = pos[k * TPB];
			pos[k * TPB] = (unsigned short)run; run += c; end[k * TPB] = (unsigned short)run;
		}
		nrealB = run - 4 * nA4;
	}
	// pass 2: scatter into the buckets (end[] serves as the fill pointer and is restored afterwards)
#pragma unroll
	for (int k = 0; k < 32; ++k) end[k * TPB] = pos[k * TPB];
	for (int o = 0; o < n4; ++o) {
		const uint2 w = my[(size_t)o * BT];
		const int sb = o < nA4 ? 0 : 16;
		const uint32_t ee[4] = {w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16};
#pragma unroll
		for (int u = 0; u < 4; ++u)
			if (ee[u]) { unsigned short* f = end + (sb + ((ee[u] >> 3) & 15)) * TPB; srt[(size_t)(*f) * TPB] = (unsigned short)ee[u]; ++*f; }
	}
	// pass 3: emit in the wanted order, four entries per store
	const int l16 = tid & 15;
	for (int o = 0; o < n4; ++o) {
		uint32_t q4[4];
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			const int t = 4 * o + u;
			const int seg = o < nA4 ? 0 : 1, sb = 16 * seg;
			const int rel = t - (seg ? 4 * nA4 : 0);
			uint32_t ev = 0;
			if (rel < (seg ? nrealB : nrealA)) {
				int b = (l16 + t) & 15;
				if (pos[(sb + b) * TPB] == end[(sb + b) * TPB]) {
					b = (b + 8) & 15;
					if (pos[(sb + b) * TPB] == end[(sb + b) * TPB]) {
						int best = -1;
						for (int k = 0; k < 16; ++k) { const int left = (int)end[(sb + k) * TPB] - (int)pos[(sb + k) * TPB]; if (left > best) { best = left; b = k; } }
					}
				}
				unsigned short* pp = pos + (sb + b) * TPB;
				ev = srt[(size_t)(*pp) * TPB]; ++*pp;
			}
			q4[u] = ev;
		}
		my[(size_t)o * BT] = make_uint2(q4[0] | (q4[1] << 16), q4[2] | (q4[3] << 16));
	}
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
// ZCHK:  keep the reference's "skip a coincident molecule of the same cell" test (see the comment at the cutoff predicate).

// ---- asynchronous staging primitives (sm_80+: cp.async = LDGSTS; mbarrier completion instead of wait_group + __syncthreads) ----
__device__ __forceinline__ uint32_t nl_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void nl_mbar_init(uint64_t* bar, unsigned count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(nl_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void nl_cp_async_16(uint32_t dst, const void* src) {
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void nl_cp_async_8(uint32_t dst, const void* src) {
	asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(dst), "l"(src) : "memory");
}
// one arrival on `bar` once all cp.async operations this thread has issued so far have landed in shared memory
__device__ __forceinline__ void nl_cp_async_arrive(uint64_t* bar) {
	asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" :: "r"(nl_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void nl_mbar_wait(uint64_t* bar, unsigned parity) {
	uint32_t done;
	do {
		asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
		             : "=r"(done) : "r"(nl_smem_u32(bar)), "r"(parity) : "memory");
	} while (!done);
}

// Stages the tile rows of a block: Z[k] (8 B) and XY[k] (16 B) per entry, straight from the double4 positions (x y | z w).
// Batch A = rows [0, ROWS_A) -> mbar[0], batch B = the rest -> mbar[1].  Every thread arrives on both barriers.
template <int R>
__device__ __forceinline__ void nl_stage_tile(const NLArgs& a, const NLPlan& P, double* Z, double2* XY, uint64_t* mbar) {
	const uint32_t z0 = nl_smem_u32(Z), xy0 = nl_smem_u32(XY);
	const int tid = threadIdx.x;
#pragma unroll 1
	for (int r = 0; r < NLGeom<R>::NR; ++r) {
		const int off = P.off[r] + NL_DUMMY, len = P.len[r];
		const double4* src = a.pos + P.beg[r];
		for (int t = tid; t < len; t += NLGeom<R>::BT) {
			nl_cp_async_16(xy0 + 16u * (uint32_t)(off + t), src + t);
			nl_cp_async_8(z0 + 8u * (uint32_t)(off + t), (const char*)(src + t) + 16);
		}
		if (r + 1 == NLGeom<R>::ROWS_A) nl_cp_async_arrive(mbar);
	}
	nl_cp_async_arrive(mbar + 1);
}

// 1/a for a positive, normal double: hardware seed (MUFU.RCP64H writes the HIGH word of the estimate, relative error <= 2^-20) + ONE
// third-order step x1 = x0 (1 + e + e^2), e = 1 - a x0.  The low word of the seed is taken from a itself instead of being
// zeroed (one instruction less per pair): that moves the seed by < 2^-20 relative, so |e| < 2^-19 and the result is good to
// e^3 < 2^-57 - below the rounding of the three FMAs that follow.
__device__ __forceinline__ double nl_rcp3(double a) {
	double x;
	asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(a));
	x = __hiloint2double(__double2hiint(x), __double2loint(a));
	const double e = fma(-a, x, 1.0);
	const double t = fma(e, e, e);
	return fma(x, t, x);
}

// ---------------------------------------------------------------------------------------------------------------
// prune: outer lists (radius rc + skin) -> inner lists (radius rc + skin_in), every few steps
// ---------------------------------------------------------------------------------------------------------------
// The force pass costs one shared-memory read and ~25 FP64 operations per LISTED partner, inside the cutoff or not; with the skin that
// makes the lists live 18 steps, 40 % of the listed partners are outside.  So the force kernels read a second, short-lived list: the
// partners of the outer list that are closer than rc + skin_in NOW.  Exactness: a pair inside rc at a later step t was inside
// rc + skin_in at the prune as long as no molecule moved more than skin_in / 2 since the prune (tracked against pos_prune into
// flags[1], the same mechanism as for the build); and it was in the outer list at the prune as long as no molecule had moved more
// than (skin - skin_in) / 2 since the build - which is why, with inner lists, the OUTER lists count as expired at that displacement
// already (step.cu: one guard word covers both conditions, so a prune is possible at every step that may run).  Same block decomposition, tile staging and list layout as the force kernel; entries keep their order, batch-A
// quads stay in front; one lane = one molecule, no interaction between lanes.
template <int R>
__global__ void __launch_bounds__(NLGeom<R>::BT, R == 1 ? 6 : 1024 / NLGeom<R>::BT) nl_prune_kernel(NLArgs a) {
	constexpr int BT = NLGeom<R>::BT;
	extern __shared__ __align__(16) unsigned char nl_tile[];
	double* Z = (double*)nl_tile;
	double2* XY = (double2*)(nl_tile + (size_t)a.tcap * 8);
	__shared__ NLPlan P;
	__shared__ __align__(8) uint64_t mbar[2];
	const int tid = threadIdx.x;
	const unsigned blk = blockIdx.x;
	if (a.guard && *a.guard > a.guard_thr) return;      // a prune enqueued ahead of the host's knowledge, lists already expired: like the step it belongs to, it does nothing
	if (tid == 0) { nl_mbar_init(mbar, BT); nl_mbar_init(mbar + 1, BT); }
	if (tid < NL_DUMMY) { Z[tid] = 1e150; XY[tid] = make_double2(1e150, 1e150); }
	if (tid < (int)(sizeof(NLPlan) / 4)) ((uint32_t*)&P)[tid] = ((const uint32_t*)(a.plans + blk))[tid];
	__syncthreads();
	int qr; uint32_t i; bool has;
	nl_my_slot<R>(P, tid, qr, i, has);
	const bool active = P.live && has;
	if (!active) a.lcount_out[(size_t)blk * BT + tid] = 0u;
	if (!P.live) return;
	nl_stage_tile<R>(a, P, Z, XY, mbar);
	double4 Pi = make_double4(0., 0., 0., 0.);
	uint32_t cnt = 0;
	if (active) { Pi = a.pos[i]; cnt = a.lcount[(size_t)blk * BT + tid]; a.pos_prune_out[i] = Pi; }
	const uint2* my = (const uint2*)(a.lists + ((size_t)blk * a.lcap4 * BT + tid) * 4);
	uint2* out = (uint2*)(a.lists_out + ((size_t)blk * a.lcap4 * BT + tid) * 4);
	const int n4 = (int)(cnt & 0xffffu), nA4 = min((int)(cnt >> 16), n4);
	const unsigned long long rin2b = (unsigned long long)__double_as_longlong(a.rin2);
	const char* const Zb = (const char*)Z; const char* const XYb = (const char*)XY;
	nl_mbar_wait(mbar, 0);
	nl_mbar_wait(mbar + 1, 0);
	unsigned long long w64 = 0ull;      // the quad being filled
	int nw = 0, o4 = 0, oA4 = 0;
	for (int o = 0; o < n4; ++o) {
		if (o == nA4) oA4 = o4;          // complete quads so far hold batch-A entries only
		const uint2 w = my[(size_t)o * BT];
		const uint32_t ee[4] = {w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16};
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			const uint32_t e = ee[u];
			const double zk = *(const double*)(Zb + e);
			const double2 xyk = *(const double2*)(XYb + (e + e));
			const double dx = Pi.x - xyk.x, dy = Pi.y - xyk.y, dz = Pi.z - zk;
			const double r2 = dx * dx + dy * dy + dz * dz;
			if (e != 0u && (unsigned long long)__double_as_longlong(r2) < rin2b) {
				w64 |= (unsigned long long)e << (16 * nw);
				if (++nw == 4) { out[(size_t)o4 * BT] = make_uint2((uint32_t)w64, (uint32_t)(w64 >> 32)); ++o4; w64 = 0ull; nw = 0; }
			}
		}
	}
	if (n4 <= nA4) oA4 = o4;
	if (nw) { out[(size_t)o4 * BT] = make_uint2((uint32_t)w64, (uint32_t)(w64 >> 32)); ++o4; }      // missing entries are 0 = the dummy
	if (active) a.lcount_out[(size_t)blk * BT + tid] = (uint32_t)o4 | ((uint32_t)oA4 << 16);
}

// after nl_prune_kernel (all its blocks have read the guard): the displacement since the prune starts at zero
__global__ void nl_prune_commit_kernel(unsigned long long* flags, const unsigned long long* guard, unsigned long long guard_thr) {
	if (guard && *guard > guard_thr) return;
	flags[1] = 0ull;
}

// the guard word of the next step (flags[3]; kernels return at once when it exceeds the bits of 1.0): expired = the outer lists moved
// beyond thr_o2, or the inner lists (if in use) beyond thr_i2.  Runs after the step's integrator (and the all-reduce of flags[0..1]).
__global__ void nl_guard_kernel(unsigned long long* flags, double thr_o2, double thr_i2, int inner) {
	const double dO = __longlong_as_double((long long)flags[0]), dI = __longlong_as_double((long long)flags[1]);
	const bool expired = dO > thr_o2 || (inner && dI > thr_i2);
	flags[3] = (unsigned long long)__double_as_longlong(expired ? 2.0 : 0.0);
}

template <bool COUNT, bool VI, int FUSE, bool ZCHK, int R>
__global__ void __launch_bounds__(NLGeom<R>::BT, R == 1 ? 6 : 1024 / NLGeom<R>::BT) nl_force_kernel(NLArgs a) {
	constexpr int BT = NLGeom<R>::BT;
	extern __shared__ __align__(16) unsigned char nl_tile[];
	double* Z = (double*)nl_tile;                               // [tcap]   z of tile entry k
	double2* XY = (double2*)(nl_tile + (size_t)a.tcap * 8);     // [tcap]   (x, y) of tile entry k; tcap is even
	__shared__ NLPlan P;
	__shared__ double red[(FUSE ? 8 : 5) * 32];
	__shared__ __align__(8) uint64_t mbar[2];
	const int tid = threadIdx.x;
	// phase 2: the boundary blocks (they wait for the halo positions); phase 1: all the others; phase 0: every block
	const unsigned blk = a.phase == 2 ? a.bmap[blockIdx.x] : (a.phase == 1 ? a.bmap[a.nb_total - 1 - blockIdx.x] : blockIdx.x);
	// speculatively enqueued step whose lists turned out expired: nothing is computed, the host rebuilds and repeats it
	if (FUSE != 1 && a.guard && *a.guard > a.guard_thr) return;
	if (tid == 0) { nl_mbar_init(mbar, BT); nl_mbar_init(mbar + 1, BT); }
	// dummy entries: far outside any cutoff, yet r^2 = 3e300 stays FINITE (the trace accumulation multiplies r^2 by an exact 0)
	if (tid < NL_DUMMY) { Z[tid] = 1e150; XY[tid] = make_double2(1e150, 1e150); }
	if (tid < (int)(sizeof(NLPlan) / 4)) ((uint32_t*)&P)[tid] = ((const uint32_t*)(a.plans + blk))[tid];
	__syncthreads();                                   // plan, barriers and dummies are visible afterwards
	int qr; uint32_t i; bool has;
	nl_my_slot<R>(P, tid, qr, i, has);
	if (FUSE && has) {      // the epilogue's inputs: bring their lines into L2 now (no registers held)
		asm volatile("prefetch.global.L2 [%0];" :: "l"(a.vel + 3 * (size_t)i));
		if (FUSE == 1) asm volatile("prefetch.global.L2 [%0];" :: "l"(a.pos_build + i));
		if (FUSE == 1 && a.pos_prune) asm volatile("prefetch.global.L2 [%0];" :: "l"(a.pos_prune + i));
	}
	const bool active = P.live && has;
	if (FUSE == 1 && a.guard && *a.guard > a.guard_thr) {      // the host swaps the two position buffers after every fused launch: carry the positions over
		if (has) a.pos_out[i] = a.pos[i];
		return;
	}
	constexpr int NV = FUSE ? 8 : 5;
	double acc[NV];
#pragma unroll
	for (int k = 0; k < NV; ++k) acc[k] = 0.;
	double disp2 = 0., disp2i = 0.;
	if (P.live) {
		nl_stage_tile<R>(a, P, Z, XY, mbar);
		double4 Pi = make_double4(0., 0., 0., 0.);
		uint32_t cnt = 0;
		const uint2* my = (const uint2*)(a.lists + ((size_t)blk * a.lcap4 * BT + tid) * 4);   // 4 entries per load
		// the first three quads are requested together with the count (the buffer holds >= 32 quads per lane).  What lies beyond a
		// lane's count is uninitialised memory: cleared below, because a lane WITHOUT a molecule still walks the warp-uniform
		// batch-A loop, where entry 0 (the dummy) is harmless and anything else would index beyond the tile
		uint2 p0 = my[0], p1 = my[BT], p2 = my[2 * BT];
		if (active) { Pi = a.pos[i]; cnt = a.lcount[(size_t)blk * BT + tid]; }
		{
			const int nq = (int)(cnt & 0xffffu);
			if (nq < 1) p0 = make_uint2(0, 0);
			if (nq < 2) p1 = make_uint2(0, 0);
			if (nq < 3) p2 = make_uint2(0, 0);
		}
		double Fx = 0., Fy = 0., Fz = 0., Vx = 0., Vy = 0., Vz = 0., u6 = 0.;
		int np_i = 0;
		// 4 pairs at a time so that the dependency chains overlap.
		// eps24 and shift6 are constants of the single component: F, V accumulate d * inv*(2 lj12 - lj6) and are scaled by
		// eps24 once; upot6 = eps24 * sum(lj12 - lj6) + shift6 * pairs.
		double s6 = 0.;
		const unsigned long long rc2b = (unsigned long long)__double_as_longlong(a.rclj2), rc2m1 = rc2b - 1ull;
		const char* const Zb = (const char*)Z; const char* const XYb = (const char*)XY;
		auto quad = [&](const uint2& w) {
			const uint32_t ww[2] = {w.x, w.y};
			double dx[4], dy[4], dz[4], r2[4];
			bool ok[4];
#pragma unroll
			for (int u = 0; u < 4; ++u) {
				// 8 * tile entry; PRMT for the low half keeps the mask ahead of the doubling (one instruction each)
				const uint32_t e = (u & 1) ? (ww[u >> 1] >> 16) : __byte_perm(ww[u >> 1], 0u, 0x4410u);
				const double zk = *(const double*)(Zb + e);
				const double2 xyk = *(const double2*)(XYb + (e + e));
				dx[u] = Pi.x - xyk.x; dy[u] = Pi.y - xyk.y; dz[u] = Pi.z - zk;
				r2[u] = dx[u] * dx[u] + dy[u] * dy[u] + dz[u] * dz[u];
				// r2 and rc^2 are non-negative doubles: their order is the order of their bit patterns, so the test runs on
				// the integer pipe and leaves the FP64 pipe to the force arithmetic.
				// ZCHK: r2 == 0 is a coincident molecule of the same cell, which the reference skips (SingleCellPolicy_ mask,
				// Helpers.h:367-376): (r2b - 1) <u (rc2b - 1) folds that into the same predicate (r2b == 0 wraps to the largest
				// unsigned).  The build kernel reports whether any listed pair is that close; without one the plain compare is exact.
				const unsigned long long r2b = (unsigned long long)__double_as_longlong(r2[u]);
				ok[u] = ZCHK ? (r2b - 1ull < rc2m1) : (r2b < rc2b);
			}
#pragma unroll
			for (int u = 0; u < 4; ++u) {
				const double inv = nl_rcp3(r2[u]);
				const double lj2 = a.sig2 * inv, lj4 = lj2 * lj2, lj6 = lj4 * lj2, lj12 = lj6 * lj6, lj12m6 = lj12 - lj6;
				const double sc = inv * (lj12 + lj12m6);
				// a rejected pair (its values may be inf/nan) is not accumulated: predicated FMAs, no selects
				if (ok[u]) {
					s6 += lj12m6;
					if (VI) {
						const double px = dx[u] * sc, py = dy[u] * sc, pz = dz[u] * sc;
						Fx += px; Fy += py; Fz += pz;
						Vx += dx[u] * px; Vy += dy[u] * py; Vz += dz[u] * pz;
					} else {
						Fx = fma(dx[u], sc, Fx); Fy = fma(dy[u], sc, Fy); Fz = fma(dz[u], sc, Fz);
						Vx = fma(r2[u], sc, Vx);                  // Vx carries the whole trace
					}
					if (COUNT) ++np_i;
				}
			}
		};
		// list entries stream from HBM: three loads in flight per lane, rotated by name (loop unrolled by 3, no register moves)
		const int n4 = (int)(cnt & 0xffffu);
		auto step = [&](uint2& slot, int o) {
			const uint2 cur = slot;
			if (o + 3 < n4) slot = my[(size_t)(o + 3) * BT];
			quad(cur);
		};
		// quads every lane of the warp has entirely inside batch A (a multiple of 3 keeps the rotation aligned): these run
		// while batch B is still in flight
		const int nA = (int)__reduce_min_sync(0xffffffffu, active ? (cnt >> 16) : 0xffffu);
		const int oA = nA == 0xffff ? 0 : (nA / 3) * 3;
		nl_mbar_wait(mbar, 0);
		int o = 0;
		for (; o < oA; o += 3) { step(p0, o); step(p1, o + 1); step(p2, o + 2); }
		nl_mbar_wait(mbar + 1, 0);
		for (; o + 3 <= n4; o += 3) { step(p0, o); step(p1, o + 1); step(p2, o + 2); }
		if (o < n4) { step(p0, o); if (o + 1 < n4) step(p1, o + 1); }
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
				double vx = vp[0], vy = vp[1], vz = vp[2];
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
				if (a.pos_prune) {
					const double4 pp = a.pos_prune[i];
					const double gx = Pi.x - pp.x, gy = Pi.y - pp.y, gz = Pi.z - pp.z;
					disp2i = gx * gx + gy * gy + gz * gz;
				}
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
		if (a.pos_prune) {
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) disp2i = fmax(disp2i, __shfl_xor_sync(0xffffffffu, disp2i, o));
			const unsigned long long bi = (unsigned long long)__double_as_longlong(disp2i);
			if ((tid & 31) == 0 && bi > *(volatile unsigned long long*)(a.flags + 1)) atomicMax(a.flags + 1, bi);
		}
	}
	block_sum<NV>(acc, red);
	if (tid == 0) {
#pragma unroll
		for (int k = 0; k < NV; ++k) a.partials[(size_t)blk * NV + k] = acc[k];
	}
}

// periodic images follow their source molecule (frozen layout; plain +-L shift, DomainDecompBase.cpp:326-341 without the
// boundary guards, which only matter for cell assignment)
__global__ void __launch_bounds__(256) halo_refresh_kernel(Grid g, size_t n, double4* __restrict__ pos, double4* __restrict__ q, const uint32_t* __restrict__ cell,
                                                            const uint32_t* __restrict__ halo_src, const uint8_t* __restrict__ halo_code) {
	size_t h = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (h >= n || !(cell[h] & HALO_FLAG)) return;
	const int code = halo_code[h];
	if (!code) return;
	const uint32_t src = halo_src[h];
	double4 p = pos[src];
	const int o0 = code % 3, o1 = (code / 3) % 3, o2 = code / 9;
	if (o0) p.x = o0 == 1 ? p.x + g.box_len[0] : p.x - g.box_len[0];
	if (o1) p.y = o1 == 1 ? p.y + g.box_len[1] : p.y - g.box_len[1];
	if (o2) p.z = o2 == 1 ? p.z + g.box_len[2] : p.z - g.box_len[2];
	pos[h] = p;
	if (q) q[h] = q[src];      // rigid molecules: the image turns with its source
}
