// cells.cu — cell assignment, periodic images and the device counting sort.
//
// Replaces (reference, src/):
//   LinkedCells::getCellIndexOfPoint / update            particleContainer/LinkedCells.cpp:243-356,830-882
//   DomainDecompBase::handleDomainLeavingParticles       parallel/DomainDecompBase.cpp:174-225
//   DomainDecompBase::populateHaloLayerWithCopies        parallel/DomainDecompBase.cpp:293-348
//   C08CellPairTraversal / C08BasedTraversals            particleContainer/LinkedCellTraversals/*.h (enumeration dump)
//
// Pipeline of one re-sort (k_sort):
//   bin      per molecule: wrap leavers, cell index (bit-exact), count self + periodic images per cell
//   scan     exclusive prefix over the cell histogram -> cell_start
//   scatter  claim a slot in the target cell for self + every image (atomic rank; order fixed later)
//   gather   per new slot: rank inside the cell by molecule id (deterministic order), copy the payload
//            from the old buffer, applying the image shift.  cur/alt are swapped afterwards.
// All four are HBM-bound; algorithmic bytes per molecule are listed in DESIGN.md.
#include "ls1g_internal.cuh"

#include <algorithm>

// ------------------------------------------------------------------------------------------------
// periodic image helpers
// ------------------------------------------------------------------------------------------------
// option o in {1: lower slab, shift +L ; 2: upper slab, shift -L}; returns true if coordinate x of dimension d
// lies in that slab (Molecule::inBox on [lo,hi), DomainDecompBase.cpp:300-325).
__device__ __forceinline__ bool in_slab(const Grid& g, int d, int o, double x) {
	double lo = o == 1 ? g.box_min[d] : __dsub_rn(g.box_max[d], g.cutoff);
	double hi = o == 1 ? __dadd_rn(g.box_min[d], g.cutoff) : g.box_max[d];
	return x >= lo && x < hi;
}
// shifted coordinate of the halo copy incl. the "must land outside the box" guards (DomainDecompBase.cpp:326-341)
__device__ __forceinline__ double image_coord(const Grid& g, int d, int o, double x) {
	if (o == 1) {
		double y = __dadd_rn(x, g.box_len[d]);
		if (y < g.box_max[d]) y = nextafter(g.box_max[d], g.box_max[d] + 1.0);
		return y;
	} else {
		double y = __dsub_rn(x, g.box_len[d]);
		if (y >= g.box_min[d]) y = nextafter(g.box_min[d], g.box_min[d] - 1.0);
		return y;
	}
}
// leaving molecules on a self-periodic dimension (DomainDecompBase.cpp:196-221)
__device__ __forceinline__ double wrap_coord(const Grid& g, int d, double x) {
	if (x < g.box_min[d]) {
		x = __dadd_rn(x, g.box_len[d]);
		if (x >= g.box_max[d]) x = nextafter(g.box_max[d], g.box_max[d] - 1.0);
	} else if (x >= g.box_max[d]) {
		x = __dsub_rn(x, g.box_len[d]);
		if (x <= g.box_min[d]) x = g.box_min[d];
	}
	return x;
}

// Enumerates the periodic images of an owned molecule at p.  f(code, x, y, z) is called for every image;
// code = o0 + 3*o1 + 9*o2 with o_d in {0 none, 1 +L, 2 -L}; code 0 (the molecule itself) is not visited.
template <class F>
__device__ __forceinline__ void for_each_image(const Grid& g, double px, double py, double pz, F f) {
	double p[3] = {px, py, pz};
	bool opt[3][3];
#pragma unroll
	for (int d = 0; d < 3; ++d) {
		opt[d][0] = true;
		opt[d][1] = g.self_periodic[d] && in_slab(g, d, 1, p[d]);
		opt[d][2] = g.self_periodic[d] && in_slab(g, d, 2, p[d]);
	}
	if (!(opt[0][1] | opt[0][2] | opt[1][1] | opt[1][2] | opt[2][1] | opt[2][2])) return;
	for (int o2 = 0; o2 < 3; ++o2) {
		if (!opt[2][o2]) continue;
		double z = o2 ? image_coord(g, 2, o2, p[2]) : p[2];
		for (int o1 = 0; o1 < 3; ++o1) {
			if (!opt[1][o1]) continue;
			double y = o1 ? image_coord(g, 1, o1, p[1]) : p[1];
			for (int o0 = 0; o0 < 3; ++o0) {
				if (!opt[0][o0] || (o0 | o1 | o2) == 0) continue;
				double x = o0 ? image_coord(g, 0, o0, p[0]) : p[0];
				f(o0 + 3 * o1 + 9 * o2, x, y, z);
			}
		}
	}
}

__device__ __forceinline__ void image_pos(const Grid& g, int code, double& x, double& y, double& z) {
	int o0 = code % 3, o1 = (code / 3) % 3, o2 = code / 9;
	if (o0) x = image_coord(g, 0, o0, x);
	if (o1) y = image_coord(g, 1, o1, y);
	if (o2) z = image_coord(g, 2, o2, z);
}

// ------------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------------
__global__ void cell_index_kernel(Grid g, size_t n, const double* __restrict__ pts, uint32_t* __restrict__ out) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	out[i] = grid_cell_index(g, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
}

// bin: cell[] holds HALO_FLAG for stale halo copies (skipped, marked 0xFFFFFFFF).
__global__ void __launch_bounds__(256) bin_kernel(Grid g, size_t n, double4* __restrict__ pos, uint32_t* __restrict__ cell,
                                                   uint32_t* __restrict__ count, int make_images, uint32_t* __restrict__ bad) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	uint32_t c0 = cell[i];
	if (c0 & HALO_FLAG) { cell[i] = 0xFFFFFFFFu; return; }
	double4 p = pos[i];
	if (make_images) {
		double x = g.self_periodic[0] ? wrap_coord(g, 0, p.x) : p.x;
		double y = g.self_periodic[1] ? wrap_coord(g, 1, p.y) : p.y;
		double z = g.self_periodic[2] ? wrap_coord(g, 2, p.z) : p.z;
		if (x != p.x || y != p.y || z != p.z) { p.x = x; p.y = y; p.z = z; pos[i] = p; }
	}
	// a molecule that is still outside the halo region (or not finite) after the wrap moved more than a box length in one step or
	// carries NaN: the system has exploded.  Reported through *bad; the reference ends such a run too (its cell lookup fails)
	if (!(p.x >= g.halo_min[0] && p.x <= g.halo_max[0] && p.y >= g.halo_min[1] && p.y <= g.halo_max[1] && p.z >= g.halo_min[2] && p.z <= g.halo_max[2])) {
		atomicAdd(bad, 1u);
		cell[i] = 0xFFFFFFFFu;
		return;
	}
	uint32_t c = grid_cell_index(g, p.x, p.y, p.z);
	cell[i] = c;
	atomicAdd(&count[c], 1u);
	if (make_images)
		for_each_image(g, p.x, p.y, p.z, [&](int, double x, double y, double z) {
			atomicAdd(&count[grid_cell_index(g, x, y, z)], 1u);
		});
}

// three-kernel exclusive scan over the cell histogram (SCAN_ITEMS cells per block)
#define SCAN_THREADS 512
#define SCAN_ITEMS 2048
__global__ void __launch_bounds__(SCAN_THREADS) scan_reduce_kernel(const uint32_t* __restrict__ in, int n, uint32_t* __restrict__ block_sums) {
	__shared__ uint32_t sh[SCAN_THREADS / 32];
	int base = blockIdx.x * SCAN_ITEMS;
	uint32_t s = 0;
	for (int k = threadIdx.x; k < SCAN_ITEMS; k += SCAN_THREADS) { int i = base + k; if (i < n) s += in[i]; }
	for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
	if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
	__syncthreads();
	if (threadIdx.x == 0) { uint32_t t = 0; for (int w = 0; w < SCAN_THREADS / 32; ++w) t += sh[w]; block_sums[blockIdx.x] = t; }
}
__global__ void scan_blocks_kernel(uint32_t* block_sums, int nb, uint32_t* total_out) {
	// single thread block, sequential over chunks of 1024: nb is at most a few thousand
	__shared__ uint32_t sh[1024];
	__shared__ uint32_t carry;
	if (threadIdx.x == 0) carry = 0;
	__syncthreads();
	for (int base = 0; base < nb; base += 1024) {
		int i = base + threadIdx.x;
		uint32_t v = i < nb ? block_sums[i] : 0;
		sh[threadIdx.x] = v;
		__syncthreads();
		for (int o = 1; o < 1024; o <<= 1) {
			uint32_t t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
			__syncthreads();
			sh[threadIdx.x] += t;
			__syncthreads();
		}
		uint32_t incl = sh[threadIdx.x];
		if (i < nb) block_sums[i] = carry + incl - v;
		__syncthreads();
		if (threadIdx.x == 1023) carry += incl;
		__syncthreads();
	}
	if (threadIdx.x == 0) *total_out = carry;
}
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(const uint32_t* __restrict__ in, int n, const uint32_t* __restrict__ block_sums,
                                                                   uint32_t* __restrict__ out /* [n+1] */) {
	// each thread owns SCAN_ITEMS/SCAN_THREADS consecutive items
	const int PER = SCAN_ITEMS / SCAN_THREADS;
	__shared__ uint32_t sh[SCAN_THREADS / 32];
	int base = blockIdx.x * SCAN_ITEMS + threadIdx.x * PER;
	uint32_t v[PER], s = 0;
#pragma unroll
	for (int k = 0; k < PER; ++k) { v[k] = (base + k < n) ? in[base + k] : 0; s += v[k]; }
	// exclusive scan of s across the block
	uint32_t incl = s;
	int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
	if (lane == 31) sh[warp] = incl;
	__syncthreads();
	if (warp == 0) {
		uint32_t w = lane < SCAN_THREADS / 32 ? sh[lane] : 0, wi = w;
		for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += t; }
		if (lane < SCAN_THREADS / 32) sh[lane] = wi - w;
	}
	__syncthreads();
	uint32_t run = block_sums[blockIdx.x] + sh[warp] + incl - s;
#pragma unroll
	for (int k = 0; k < PER; ++k) { if (base + k < n) out[base + k] = run; run += v[k]; }
	if (base <= n - 1 && n - 1 < base + PER) out[n] = run;   // the thread owning the last item writes the total
}

__global__ void __launch_bounds__(256) scatter_kernel(Grid g, size_t n, const double4* __restrict__ pos, const uint32_t* __restrict__ cell,
                                                       const uint64_t* __restrict__ id, const uint32_t* __restrict__ cell_start,
                                                       uint32_t* __restrict__ fill, uint32_t* __restrict__ slot_src,
                                                       uint8_t* __restrict__ slot_code, uint64_t* __restrict__ slot_key,
                                                       uint32_t* __restrict__ slot_cell, int make_images) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	uint32_t c = cell[i];
	if (c == 0xFFFFFFFFu) return;
	uint64_t key = id[i];
	uint32_t s = cell_start[c] + atomicAdd(&fill[c], 1u);
	slot_src[s] = (uint32_t)i; slot_code[s] = 0; slot_key[s] = key; slot_cell[s] = c;
	if (make_images) {
		double4 p = pos[i];
		for_each_image(g, p.x, p.y, p.z, [&](int code, double x, double y, double z) {
			uint32_t ci = grid_cell_index(g, x, y, z);
			uint32_t si = cell_start[ci] + atomicAdd(&fill[ci], 1u);
			slot_src[si] = (uint32_t)i; slot_code[si] = (uint8_t)code; slot_key[si] = key; slot_cell[si] = ci;
		});
	}
}

struct GatherArgs {
	MolBuf src, dst;
	const double *F_src, *M_src, *Vi_src; double *F_dst, *M_dst, *Vi_dst;   // carried only when non-null
	int has_rot;
	uint32_t* slot_dst;      // [n_new] final slot of scatter entry s
	uint32_t* new_of_old;    // [n_old] new slot of every surviving (non-image) molecule; aliases src.cell
	int geometric;           // 1: a molecule in a halo cell is a halo copy (after an exchange: images, received copies, leavers kept by the
	                         // sender); 0: LinkedCells::update - an owned molecule that left the box sits in a halo cell and STAYS owned
	                         // until the next exchange wraps or sends it (LinkedCells.cpp:243-356 moves molecules, it never deletes them)
};

__global__ void __launch_bounds__(256) gather_kernel(Grid g, size_t n_new, GatherArgs a, const uint32_t* __restrict__ cell_start,
                                                      const uint32_t* __restrict__ slot_src, const uint8_t* __restrict__ slot_code,
                                                      const uint64_t* __restrict__ slot_key, const uint32_t* __restrict__ slot_cell) {
	size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n_new) return;
	uint32_t c = slot_cell[s];
	uint32_t b0 = cell_start[c], b1 = cell_start[c + 1];
	uint64_t key = slot_key[s];
	uint32_t rank = 0;
	for (uint32_t k = b0; k < b1; ++k) {
		uint64_t kk = slot_key[k];
		rank += (kk < key) || (kk == key && k < (uint32_t)s);
	}
	size_t dst = (size_t)b0 + rank;
	uint32_t src = slot_src[s];
	int code = slot_code[s];
	a.slot_dst[s] = (uint32_t)dst;
	if (!code) a.new_of_old[src] = (uint32_t)dst;
	double4 p = a.src.pos[src];
	if (code) image_pos(g, code, p.x, p.y, p.z);
	a.dst.pos[dst] = p;
	a.dst.id[dst] = key;
	bool halo = (a.geometric || code) && grid_is_halo(g, c);
	a.dst.cell[dst] = c | (halo ? HALO_FLAG : 0u);
#pragma unroll
	for (int d = 0; d < 3; ++d) a.dst.vel[3 * dst + d] = a.src.vel[3 * (size_t)src + d];
	if (a.has_rot) {
		a.dst.q[dst] = a.src.q[src];
#pragma unroll
		for (int d = 0; d < 3; ++d) a.dst.L[3 * dst + d] = a.src.L[3 * (size_t)src + d];
	}
	if (a.F_src) {
#pragma unroll
		for (int d = 0; d < 3; ++d) a.F_dst[3 * dst + d] = a.F_src[3 * (size_t)src + d];
		if (a.M_src)
#pragma unroll
			for (int d = 0; d < 3; ++d) a.M_dst[3 * dst + d] = a.M_src[3 * (size_t)src + d];
#pragma unroll
		for (int d = 0; d < 3; ++d) a.Vi_dst[3 * dst + d] = a.Vi_src[3 * (size_t)src + d];
	}
}

// C08 enumeration on the device: one thread per base cell writes its (up to 14) triplets into a
// fixed-stride table; kind 2 marks an unused row.  Colour-major order is restored on the host.
__global__ void c08_pairs_kernel(Grid g, uint32_t* __restrict__ out /* [n_cells][14][3] */) {
	long base = (long)blockIdx.x * blockDim.x + threadIdx.x;
	if (base >= g.n_cells) return;
	const long nx = g.cpd[0], ny = g.cpd[1], nz = g.cpd[2];
	long cx = base % nx, cy = (base / nx) % ny, cz = base / (nx * ny);
	uint32_t* row = out + (size_t)base * 42;
	bool is_base = cx < nx - 1 && cy < ny - 1 && cz < nz - 1;
	const long o = 0, x = 1, y = nx, z = nx * ny, xy = x + y, yz = y + z, xz = x + z, xyz = x + y + z;
	const long pairs[14][2] = {{o, o}, {o, y}, {y, z}, {o, z}, {o, yz}, {x, yz}, {x, y}, {x, z},
	                           {o, x}, {o, xy}, {xy, z}, {y, xz}, {o, xz}, {o, xyz}};   // C08BasedTraversals.h:127-142
	for (int j = 0; j < 14; ++j) {
		uint32_t k = 2, a = 0, b = 0;
		if (is_base) {
			uint32_t c1 = (uint32_t)(base + pairs[j][0]), c2 = (uint32_t)(base + pairs[j][1]);
			bool h1 = grid_is_halo(g, c1), h2 = grid_is_halo(g, c2);
			if (!(h1 && h2)) {                       // C08BasedTraversals.h:82-84
				a = c1; b = c2; k = 1;
				if (c1 == c2) k = 0;                 // processCell
				else if (h1) { a = c2; b = c1; }     // non-halo cell first (:90-95)
			}
		}
		row[3 * j] = a; row[3 * j + 1] = b; row[3 * j + 2] = k;
	}
}

// ------------------------------------------------------------------------------------------------
// host launchers
// ------------------------------------------------------------------------------------------------
static inline unsigned nblk(size_t n, unsigned t) { return (unsigned)((n + t - 1) / t); }

int k_cell_index(ls1g_ctx* c, size_t n, const double* d_pts, uint32_t* d_out) {
	if (n == 0) return 0;
	cell_index_kernel<<<nblk(n, 256), 256, 0, c->stream>>>(c->grid, n, d_pts, d_out);
	CUDA_TRY(cudaGetLastError());
	return 0;
}

// ------------------------------------------------------------------------------------------------
// ls1g_cell_index_of_points_noalloc: the same cell_index_kernel as ls1g_cell_index_of_points, for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
// The device staging of the points and of the result - which the ordinary entry point
// cudaMallocs and frees on every call - lives in a file-scope __device__ array of fixed size: every GPU that loads the library
// holds these 1.75 MiB until the process exits.  The array is SHARED by all calls and all contexts of a device and keeps the
// contents of the last call: nothing in it has to start at zero (the copy overwrites points [0, n), the kernel writes every
// result [0, n), only those are read back), but two calls that overlap on the GPU (different contexts/streams, different
// threads) race on it - callers serialise.
// In host code `g_noalloc_scratch` is only a stand-in for the symbol, never a device address: its address comes from
// cudaGetSymbolAddress, asked once per device with that device current.
// ------------------------------------------------------------------------------------------------
#define NOALLOC_PTS_BYTES ((size_t)LS1G_NOALLOC_MAX_POINTS * 3 * sizeof(double))       // 1 572 864: a multiple of 256
#define NOALLOC_BYTES (NOALLOC_PTS_BYTES + (size_t)LS1G_NOALLOC_MAX_POINTS * sizeof(uint32_t))
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];
#define NOALLOC_MAX_DEVICES 64
static unsigned char* g_noalloc_addr[NOALLOC_MAX_DEVICES];      // device address of the array on device d (0 = not asked yet)

extern "C" int ls1g_cell_index_of_points_noalloc(ls1g_ctx* c, size_t n, const double* pts, uint32_t* out) {
	if (!c) LS1G_FAIL("null context");
	if (n > (size_t)LS1G_NOALLOC_MAX_POINTS)
		LS1G_FAIL("ls1g_cell_index_of_points_noalloc: " + std::to_string(n) + " points exceed the fixed scratch of LS1G_NOALLOC_MAX_POINTS = " +
		          std::to_string(LS1G_NOALLOC_MAX_POINTS) + " (split the call, or use ls1g_cell_index_of_points)");
	if (n == 0) return 0;
	if (!pts || !out) LS1G_FAIL("ls1g_cell_index_of_points_noalloc: null argument");
	const int dev = c->cfg.device;
	if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) LS1G_FAIL("ls1g_cell_index_of_points_noalloc: device ordinal out of range");
	CUDA_TRY(cudaSetDevice(dev));
	if (!g_noalloc_addr[dev]) {          // set-up, once per device, with that device current
		unsigned char* p = nullptr;
		CUDA_TRY(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
		if (!p) LS1G_FAIL("ls1g_cell_index_of_points_noalloc: cudaGetSymbolAddress returned a null address");
		g_noalloc_addr[dev] = p;
	}
	double* d_p = (double*)g_noalloc_addr[dev];
	uint32_t* d_o = (uint32_t*)(g_noalloc_addr[dev] + NOALLOC_PTS_BYTES);
	CUDA_TRY(cudaMemcpyAsync(d_p, pts, n * 3 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
	if (k_cell_index(c, n, d_p, d_o)) return 1;
	CUDA_TRY(cudaMemcpyAsync(out, d_o, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
	cudaError_t e = cudaStreamSynchronize(c->stream);
	if (e != cudaSuccess) { ls1g_set_error(std::string("cell index kernel (noalloc): ") + cudaGetErrorString(e)); return 1; }
	return 0;
}

// exclusive scan over the n_cells entries of `in` -> out[0..n_cells]; the total also lands in d_counters[0]
int ls1g_scan_cells(ls1g_ctx* c, const uint32_t* in, uint32_t* out) {
	const int nc = c->bgrid.n_cells;
	int nb = (nc + SCAN_ITEMS - 1) / SCAN_ITEMS;
	scan_reduce_kernel<<<nb, SCAN_THREADS, 0, c->stream>>>(in, nc, c->scan_block);
	scan_blocks_kernel<<<1, 1024, 0, c->stream>>>(c->scan_block, nb, c->d_counters);
	scan_apply_kernel<<<nb, SCAN_THREADS, 0, c->stream>>>(in, nc, c->scan_block, out);
	CUDA_TRY(cudaGetLastError());
	return 0;
}

// owned molecules of the fullest cell row (y, z): sizes the block grid of the neighbour-list kernels (lists.cuh)
__global__ void __launch_bounds__(256) row_max_kernel(Grid g, const uint32_t* __restrict__ cell_start, uint32_t* __restrict__ out) {
	const int nyi = g.cpd[1] - 2, nzi = g.cpd[2] - 2;
	int row = blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t n = 0;
	if (row < nyi * nzi) {
		const int crow = ((row / nyi + 1) * g.cpd[1] + (row % nyi + 1)) * g.cpd[0];
		n = cell_start[crow + g.cpd[0] - 1] - cell_start[crow + 1];
	}
	n = __reduce_max_sync(0xffffffffu, n);
	if ((threadIdx.x & 31) == 0 && n) atomicMax(out, n);
}

// owned molecules of the fullest 2 x 2 bundle of cell rows (2gy+1.., 2gz+1..): sizes the block grid of the bundled list kernels
__global__ void __launch_bounds__(256) bundle_max_kernel(Grid g, const uint32_t* __restrict__ cell_start, uint32_t* __restrict__ out) {
	const int nyi = g.cpd[1] - 2, nzi = g.cpd[2] - 2, nyg = (nyi + 1) / 2, nzg = (nzi + 1) / 2;
	int b = blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t n = 0;
	if (b < nyg * nzg) {
		const int gy = b % nyg, gz = b / nyg;
		for (int q = 0; q < 4; ++q) {
			const int cy = 2 * gy + 1 + (q & 1), cz = 2 * gz + 1 + (q >> 1);
			if (cy > nyi || cz > nzi) continue;
			const int crow = (cz * g.cpd[1] + cy) * g.cpd[0];
			n += cell_start[crow + g.cpd[0] - 1] - cell_start[crow + 1];
		}
	}
	n = __reduce_max_sync(0xffffffffu, n);
	if ((threadIdx.x & 31) == 0 && n) atomicMax(out, n);
}

int k_sort(ls1g_ctx* c, int make_images) {
	const Grid& g = c->bgrid;
	const int nc = g.n_cells;
	size_t n = c->n_slots;
	timer_start(c, LS1G_T_SORT);
	CUDA_TRY(cudaMemsetAsync(c->cell_count, 0, sizeof(uint32_t) * (nc + 1), c->stream));
	CUDA_TRY(cudaMemsetAsync(c->cell_fill, 0, sizeof(uint32_t) * (nc + 1), c->stream));
	int launches = 0;
	CUDA_TRY(cudaMemsetAsync(c->d_counters + 9, 0, sizeof(uint32_t), c->stream));
	if (n) { bin_kernel<<<nblk(n, 256), 256, 0, c->stream>>>(g, n, c->cur.pos, c->cur.cell, c->cell_count, make_images, c->d_counters + 9); ++launches; }
	if (ls1g_scan_cells(c, c->cell_count, c->cell_start)) return 1;
	CUDA_TRY(cudaMemsetAsync(c->d_counters + 7, 0, 2 * sizeof(uint32_t), c->stream));
	row_max_kernel<<<nblk((size_t)(g.cpd[1] - 2) * (g.cpd[2] - 2), 256), 256, 0, c->stream>>>(g, c->cell_start, c->d_counters + 7);
	bundle_max_kernel<<<nblk((size_t)((g.cpd[1] - 1) / 2) * ((g.cpd[2] - 1) / 2), 256), 256, 0, c->stream>>>(g, c->cell_start, c->d_counters + 8);
	launches += 5;
	CUDA_TRY(cudaMemcpyAsync(c->h_counters, c->d_counters, 10 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
	CUDA_TRY(cudaStreamSynchronize(c->stream));
	if (c->h_counters[9]) {
		timer_stop(c, LS1G_T_SORT, launches);
		LS1G_FAIL("ls1g: " + std::to_string(c->h_counters[9]) + " molecule(s) outside the halo region or not finite after the periodic wrap - the system has "
		          "exploded (time step too large for this state?)");
	}
	size_t n_new = c->h_counters[0];
	c->max_row = c->h_counters[7];
	c->max_bundle = c->h_counters[8];
	if (n_new > c->cap) {
		timer_stop(c, LS1G_T_SORT, launches);
		LS1G_FAIL("ls1g: capacity exceeded by halo copies (" + std::to_string(n_new) + " > " + std::to_string(c->cap) + ")");
	}
	if (n) {
		scatter_kernel<<<nblk(n, 256), 256, 0, c->stream>>>(g, n, c->cur.pos, c->cur.cell, c->cur.id, c->cell_start, c->cell_fill,
		                                                   c->slot_src, c->slot_code, c->slot_key, c->slot_cell, make_images);
		++launches;
	}
	if (n_new) {
		GatherArgs a;
		a.src = c->cur; a.dst = c->alt; a.has_rot = c->has_rot;
		a.slot_dst = c->slot_dst; a.new_of_old = c->cur.cell; a.geometric = make_images ? 1 : 0;
		// forces are carried through a re-sort only when they exist (API use outside ls1g_run); Vi is not carried
		bool carry = c->forces_valid;
		a.F_src = carry ? c->F : nullptr; a.F_dst = c->F_alt;
		a.M_src = (carry && c->has_rot) ? c->M : nullptr; a.M_dst = c->M_alt;
		a.Vi_src = c->Vi; a.Vi_dst = c->Vi_alt;
		gather_kernel<<<nblk(n_new, 256), 256, 0, c->stream>>>(g, n_new, a, c->cell_start, c->slot_src, c->slot_code, c->slot_key, c->slot_cell);
		++launches;
		if (carry) { std::swap(c->F, c->F_alt); std::swap(c->Vi, c->Vi_alt); if (c->has_rot) std::swap(c->M, c->M_alt); }
	}
	CUDA_TRY(cudaGetLastError());
	std::swap(c->cur, c->alt);
	c->n_slots = n_new;
	c->n_halo = n_new - c->n_owned;
	c->sorted = 1;
	timer_stop(c, LS1G_T_SORT, launches);
	return 0;
}

// halo_src / halo_code for the frozen-layout steps of the neighbour-list loop: every periodic image remembers the slot
// of the molecule it copies.  Valid directly after k_sort(c, 1) (alt.cell still holds new_of_old).
__global__ void __launch_bounds__(256) halo_map_kernel(size_t n_new, const uint32_t* __restrict__ slot_src, const uint8_t* __restrict__ slot_code,
                                                        const uint32_t* __restrict__ slot_dst, const uint32_t* __restrict__ new_of_old,
                                                        uint32_t* __restrict__ halo_src, uint8_t* __restrict__ halo_code) {
	size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n_new) return;
	const uint8_t code = slot_code[s];
	const uint32_t dst = slot_dst[s];
	halo_code[dst] = code;
	halo_src[dst] = code ? new_of_old[slot_src[s]] : dst;
}

int k_halo_map(ls1g_ctx* c) {
	if (c->n_slots == 0) return 0;
	halo_map_kernel<<<nblk(c->n_slots, 256), 256, 0, c->stream>>>(c->n_slots, c->slot_src, c->slot_code, c->slot_dst, c->alt.cell, c->halo_src, c->halo_code);
	CUDA_TRY(cudaGetLastError());
	return 0;
}

int k_drop_halo(ls1g_ctx* c) {
	// halo copies keep HALO_FLAG in cell[]: the integrator and the next bin skip them.  Nothing to move.
	c->n_halo = 0;
	return 0;
}

int k_debug_cell_pairs(const ls1g_ctx* c, uint32_t* h_triplets, size_t capacity, size_t* n_out) {
	const Grid& g = c->grid;
	size_t rows = (size_t)g.n_cells * 14;
	uint32_t* d = nullptr;
	CUDA_TRY(cudaMalloc((void**)&d, rows * 3 * sizeof(uint32_t)));
	c08_pairs_kernel<<<nblk(g.n_cells, 128), 128, 0, c->stream>>>(g, d);
	std::vector<uint32_t> h(rows * 3);
	cudaError_t e = cudaMemcpyAsync(h.data(), d, rows * 3 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream);
	if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
	cudaFree(d);
	if (e != cudaSuccess) { ls1g_set_error(std::string("c08 dump: ") + cudaGetErrorString(e)); return 1; }
	// colour-major order as C08CellPairTraversal::traverseCellPairs (C08CellPairTraversal.h:60-79)
	size_t cnt = 0;
	const long nx = g.cpd[0], ny = g.cpd[1], nz = g.cpd[2];
	for (int col = 0; col < 8; ++col) {
		long bx = col % 2, by = (col / 2) % 2, bz = col / 4;
		for (long cz = bz; cz < nz - 1; cz += 2)
			for (long cy = by; cy < ny - 1; cy += 2)
				for (long cx = bx; cx < nx - 1; cx += 2) {
					size_t base = (size_t)((cz * ny + cy) * nx + cx);
					for (int j = 0; j < 14; ++j) {
						const uint32_t* r = &h[(base * 14 + j) * 3];
						if (r[2] == 2) continue;
						if (h_triplets && cnt < capacity) { h_triplets[3 * cnt] = r[0]; h_triplets[3 * cnt + 1] = r[1]; h_triplets[3 * cnt + 2] = r[2]; }
						++cnt;
					}
				}
	}
	*n_out = cnt;
	return 0;
}
