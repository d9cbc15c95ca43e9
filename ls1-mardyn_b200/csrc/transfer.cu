// transfer.cu — host <-> device molecule transfer of the C ABI (ls1g_upload / ls1g_upload_forces / ls1g_download).
//
// Replaces (reference, src/): ParticleContainer::addParticles (particleContainer/LinkedCells.cpp:445-520) on the way
// in and the ParticleIterator + Molecule getters (particleContainer/ParticleIterator.h:50-90) on the way out.
// Host arrays are AoS ([n][3], [n][4]); the device keeps pos as double4 (x y z | cid bits).  No per-molecule host
// loops: user memory is copied with cudaMemcpyAsync (full PCIe rate when it is pinned) into a device staging
// area and (un)packed by kernels; on the way out owned molecules are compacted in cell-sorted order.
#include "ls1g_internal.cuh"

static inline unsigned nblk(size_t n, unsigned t) { return (unsigned)((n + t - 1) / t); }

__global__ void __launch_bounds__(256) pack_pos_kernel(size_t n, const double* __restrict__ r, const uint32_t* __restrict__ cid,
                                                        double4* __restrict__ pos, uint32_t n_comp, uint32_t* __restrict__ bad) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	uint32_t c = cid[i];
	if (c >= n_comp) { atomicAdd(bad, 1u); c = 0; }
	pos[i] = make_double4(r[3 * i], r[3 * i + 1], r[3 * i + 2], cid_bits(c));
}

__global__ void __launch_bounds__(256) unit_quat_kernel(size_t n, double4* __restrict__ q) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) q[i] = make_double4(1., 0., 0., 0.);
}

// kept molecules before each cell: in[c] = keep(c) ? count(c) : 0 (scanned by the caller)
__global__ void __launch_bounds__(256) kept_count_kernel(Grid g, const uint32_t* __restrict__ cell_start, int include_halo,
                                                          uint32_t* __restrict__ out) {
	int c = blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= g.n_cells) return;
	bool keep = include_halo || !grid_is_halo(g, (uint32_t)c);
	out[c] = keep ? cell_start[c + 1] - cell_start[c] : 0u;
}

// destination row of slot s, or -1 when dropped.  identity when kept_before == nullptr.
__device__ __forceinline__ long dest_row(size_t s, const uint32_t* cell, const uint32_t* cell_start, const uint32_t* kept_before, int include_halo) {
	if (!kept_before) return (long)s;
	uint32_t cf = cell[s];
	if ((cf & HALO_FLAG) && !include_halo) return -1;
	uint32_t c = cf & ~HALO_FLAG;
	return (long)kept_before[c] + (long)(s - cell_start[c]);
}

template <int W>
__global__ void __launch_bounds__(256) compact_f64_kernel(size_t n, const double* __restrict__ src, double* __restrict__ dst,
                                                           const uint32_t* __restrict__ cell, const uint32_t* __restrict__ cell_start,
                                                           const uint32_t* __restrict__ kept_before, int include_halo) {
	size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n) return;
	long k = dest_row(s, cell, cell_start, kept_before, include_halo);
	if (k < 0) return;
#pragma unroll
	for (int d = 0; d < W; ++d) dst[(size_t)k * W + d] = src[s * W + d];
}

// pos (double4) -> r[k][3] and/or cid[k]
__global__ void __launch_bounds__(256) compact_pos_kernel(size_t n, const double4* __restrict__ pos, double* __restrict__ r, uint32_t* __restrict__ cid,
                                                           const uint32_t* __restrict__ cell, const uint32_t* __restrict__ cell_start,
                                                           const uint32_t* __restrict__ kept_before, int include_halo) {
	size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n) return;
	long k = dest_row(s, cell, cell_start, kept_before, include_halo);
	if (k < 0) return;
	double4 p = pos[s];
	if (r) { r[3 * k] = p.x; r[3 * k + 1] = p.y; r[3 * k + 2] = p.z; }
	if (cid) cid[k] = pos_cid(p);
}

__global__ void __launch_bounds__(256) compact_u64_kernel(size_t n, const uint64_t* __restrict__ src, uint64_t* __restrict__ dst,
                                                           const uint32_t* __restrict__ cell, const uint32_t* __restrict__ cell_start,
                                                           const uint32_t* __restrict__ kept_before, int include_halo) {
	size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n) return;
	long k = dest_row(s, cell, cell_start, kept_before, include_halo);
	if (k >= 0) dst[k] = src[s];
}

__global__ void __launch_bounds__(256) compact_cell_kernel(size_t n, uint32_t* __restrict__ dst, int sorted,
                                                            const uint32_t* __restrict__ cell, const uint32_t* __restrict__ cell_start,
                                                            const uint32_t* __restrict__ kept_before, int include_halo) {
	size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n) return;
	long k = dest_row(s, cell, cell_start, kept_before, include_halo);
	if (k >= 0) dst[k] = sorted ? (cell[s] & ~HALO_FLAG) : 0u;
}

__global__ void __launch_bounds__(256) fill_f64_kernel(size_t n, double* __restrict__ dst, int w, double v0, double v) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) dst[i] = (i % w == 0) ? v0 : v;
}

int ls1g_scan_cells(ls1g_ctx* c, const uint32_t* in, uint32_t* out);   // cells.cu: exclusive scan over n_cells entries -> out[n_cells+1]

static int ensure_stage(ls1g_ctx* c) {
	size_t want = c->cap * sizeof(double4);
	if (c->stage && c->stage_bytes >= want) return 0;
	cudaFree(c->stage); c->stage = nullptr; c->stage_bytes = 0;
	CUDA_TRY(cudaMalloc(&c->stage, want));
	c->stage_bytes = want;
	return 0;
}

int k_reserve_stage(ls1g_ctx* c) { return ensure_stage(c); }

// --------------------------------------------------------------------------------------------------
int k_upload(ls1g_ctx* c, size_t n, const uint64_t* id, const uint32_t* cid, const double* r, const double* v, const double* q, const double* L) {
	if (ensure_stage(c)) return 1;
	cudaStream_t st = c->stream;
	if (n) {
		CUDA_TRY(cudaMemsetAsync(c->d_counters + 6, 0, sizeof(uint32_t), st));
		CUDA_TRY(cudaMemcpyAsync(c->stage, r, n * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
		CUDA_TRY(cudaMemcpyAsync(c->slot_src, cid, n * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
		pack_pos_kernel<<<nblk(n, 256), 256, 0, st>>>(n, (const double*)c->stage, c->slot_src, c->cur.pos, (uint32_t)c->n_comp, c->d_counters + 6);
		if (v) CUDA_TRY(cudaMemcpyAsync(c->cur.vel, v, n * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
		else CUDA_TRY(cudaMemsetAsync(c->cur.vel, 0, n * 3 * sizeof(double), st));
		CUDA_TRY(cudaMemcpyAsync(c->cur.id, id, n * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
		CUDA_TRY(cudaMemsetAsync(c->cur.cell, 0, n * sizeof(uint32_t), st));
		if (c->has_rot) {
			if (q) CUDA_TRY(cudaMemcpyAsync(c->cur.q, q, n * sizeof(double4), cudaMemcpyHostToDevice, st));   // (w x y z) == double4 (x y z w) slots
			else unit_quat_kernel<<<nblk(n, 256), 256, 0, st>>>(n, c->cur.q);
			if (L) CUDA_TRY(cudaMemcpyAsync(c->cur.L, L, n * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
			else CUDA_TRY(cudaMemsetAsync(c->cur.L, 0, n * 3 * sizeof(double), st));
			CUDA_TRY(cudaMemsetAsync(c->M, 0, n * 3 * sizeof(double), st));
		}
		CUDA_TRY(cudaMemsetAsync(c->F, 0, n * 3 * sizeof(double), st));
		CUDA_TRY(cudaMemcpyAsync(c->h_counters + 6, c->d_counters + 6, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
	}
	CUDA_TRY(cudaStreamSynchronize(st));
	CUDA_TRY(cudaGetLastError());
	if (n && c->h_counters[6]) LS1G_FAIL("ls1g_upload: component id out of range");
	return 0;
}

// forces (and torques) of the uploaded molecules, same order as the last ls1g_upload (Molecule::setF / setM)
int k_upload_forces(ls1g_ctx* c, size_t n, const double* F, const double* M) {
	cudaStream_t st = c->stream;
	if (n) {
		CUDA_TRY(cudaMemcpyAsync(c->F, F, n * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
		if (c->has_rot) {
			if (M) CUDA_TRY(cudaMemcpyAsync(c->M, M, n * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
			else CUDA_TRY(cudaMemsetAsync(c->M, 0, n * 3 * sizeof(double), st));
		}
	}
	CUDA_TRY(cudaStreamSynchronize(st));
	return 0;
}

int k_download(ls1g_ctx* c, int mask, int include_halo, size_t n_ret, ls1g_fields* out) {
	const size_t ns = c->n_slots;
	if (ns == 0 || n_ret == 0) return 0;
	if (ensure_stage(c)) return 1;
	cudaStream_t st = c->stream;
	const bool identity = (n_ret == ns);
	const uint32_t* kept = nullptr;
	if (!identity) {
		kept_count_kernel<<<nblk(c->bgrid.n_cells, 256), 256, 0, st>>>(c->bgrid, c->cell_start, include_halo, c->cell_count);
		if (ls1g_scan_cells(c, c->cell_count, c->cell_fill)) return 1;
		kept = c->cell_fill;
	}
	const uint32_t* cell = c->cur.cell; const uint32_t* cs = c->cell_start;
	const unsigned nb = nblk(ns, 256);
	auto f64 = [&](const double* dev, double* dst, int w, double fill0, double fill) -> int {
		if (!dst) return 0;
		double* sg = (double*)c->stage;
		if (!dev) fill_f64_kernel<<<nblk(n_ret * w, 256), 256, 0, st>>>(n_ret * w, sg, w, fill0, fill);
		else if (w == 3) compact_f64_kernel<3><<<nb, 256, 0, st>>>(ns, dev, sg, cell, cs, kept, include_halo);
		else compact_f64_kernel<4><<<nb, 256, 0, st>>>(ns, dev, sg, cell, cs, kept, include_halo);
		CUDA_TRY(cudaMemcpyAsync(dst, sg, n_ret * w * sizeof(double), cudaMemcpyDeviceToHost, st));
		return 0;
	};
	if ((mask & LS1G_F_R) && out->r) {
		compact_pos_kernel<<<nb, 256, 0, st>>>(ns, c->cur.pos, (double*)c->stage, nullptr, cell, cs, kept, include_halo);
		CUDA_TRY(cudaMemcpyAsync(out->r, c->stage, n_ret * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
	}
	if ((mask & LS1G_F_CID) && out->cid) {
		compact_pos_kernel<<<nb, 256, 0, st>>>(ns, c->cur.pos, nullptr, (uint32_t*)c->stage, cell, cs, kept, include_halo);
		CUDA_TRY(cudaMemcpyAsync(out->cid, c->stage, n_ret * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
	}
	if ((mask & LS1G_F_ID) && out->id) {
		compact_u64_kernel<<<nb, 256, 0, st>>>(ns, c->cur.id, (uint64_t*)c->stage, cell, cs, kept, include_halo);
		CUDA_TRY(cudaMemcpyAsync(out->id, c->stage, n_ret * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
	}
	if ((mask & LS1G_F_CELL) && out->cell) {
		compact_cell_kernel<<<nb, 256, 0, st>>>(ns, (uint32_t*)c->stage, c->sorted, cell, cs, kept, include_halo);
		CUDA_TRY(cudaMemcpyAsync(out->cell, c->stage, n_ret * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
	}
	if ((mask & LS1G_F_V) && f64(c->cur.vel, out->v, 3, 0., 0.)) return 1;
	if ((mask & LS1G_F_L) && f64(c->has_rot ? c->cur.L : nullptr, out->L, 3, 0., 0.)) return 1;
	if ((mask & LS1G_F_F) && f64(c->F, out->F, 3, 0., 0.)) return 1;
	if ((mask & LS1G_F_M) && f64(c->has_rot ? c->M : nullptr, out->M, 3, 0., 0.)) return 1;
	if ((mask & LS1G_F_VI) && f64(c->Vi, out->Vi, 3, 0., 0.)) return 1;
	if ((mask & LS1G_F_Q) && f64(c->has_rot ? (const double*)c->cur.q : nullptr, out->q, 4, 1., 0.)) return 1;
	CUDA_TRY(cudaStreamSynchronize(st));
	CUDA_TRY(cudaGetLastError());
	return 0;
}

// Range variant over raw SLOTS of the cell-sorted storage (owned molecules and halo copies, the latter flagged in cell[]): the
// per-molecule arrays are contiguous in slot order, so every field but r / cid (packed in pos) goes straight from its device
// array to the host.  Lets the caller pipeline the copy of chunk k+1 with its own work on chunk k.
__global__ void __launch_bounds__(256) unpack_pos_kernel(size_t n, const double4* __restrict__ pos, double* __restrict__ r, uint32_t* __restrict__ cid) {
	size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n) return;
	const double4 p = pos[s];
	if (r) { r[3 * s] = p.x; r[3 * s + 1] = p.y; r[3 * s + 2] = p.z; }
	if (cid) cid[s] = pos_cid(p);
}

int k_download_range(ls1g_ctx* c, int mask, size_t first, size_t count, ls1g_fields* out) {
	if (count == 0) return 0;
	if (ensure_stage(c)) return 1;
	cudaStream_t st = c->stream;
	const bool want_r = (mask & LS1G_F_R) && out->r, want_cid = (mask & LS1G_F_CID) && out->cid;
	double* sr = (double*)c->stage;                                        // stage holds cap * 32 bytes: r (24 B) + cid (4 B) per molecule fit
	uint32_t* sc = (uint32_t*)((char*)c->stage + count * 3 * sizeof(double));
	if (want_r || want_cid) {
		unpack_pos_kernel<<<nblk(count, 256), 256, 0, st>>>(count, c->cur.pos + first, want_r ? sr : nullptr, want_cid ? sc : nullptr);
		if (want_r) CUDA_TRY(cudaMemcpyAsync(out->r, sr, count * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
		if (want_cid) CUDA_TRY(cudaMemcpyAsync(out->cid, sc, count * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
	}
	if ((mask & LS1G_F_ID) && out->id) CUDA_TRY(cudaMemcpyAsync(out->id, c->cur.id + first, count * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
	if ((mask & LS1G_F_CELL) && out->cell) CUDA_TRY(cudaMemcpyAsync(out->cell, c->cur.cell + first, count * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
	auto f64 = [&](const double* dev, double* dst, int w) -> int {
		if (!dst) return 0;
		if (dev) CUDA_TRY(cudaMemcpyAsync(dst, dev + first * w, count * w * sizeof(double), cudaMemcpyDeviceToHost, st));
		else for (size_t i = 0; i < count * w; ++i) dst[i] = (w == 4 && i % 4 == 0) ? 1. : 0.;      // no orientation kept: q = (1,0,0,0), L = M = 0
		return 0;
	};
	if ((mask & LS1G_F_V) && f64(c->cur.vel, out->v, 3)) return 1;
	if ((mask & LS1G_F_L) && f64(c->has_rot ? c->cur.L : nullptr, out->L, 3)) return 1;
	if ((mask & LS1G_F_F) && f64(c->F, out->F, 3)) return 1;
	if ((mask & LS1G_F_M) && f64(c->has_rot ? c->M : nullptr, out->M, 3)) return 1;
	if ((mask & LS1G_F_VI) && f64(c->Vi, out->Vi, 3)) return 1;
	if ((mask & LS1G_F_Q) && f64(c->has_rot ? (const double*)c->cur.q : nullptr, out->q, 4)) return 1;
	CUDA_TRY(cudaStreamSynchronize(st));
	CUDA_TRY(cudaGetLastError());
	return 0;
}

// pinned host staging memory for ls1g_upload / ls1g_download (full PCIe rate; pageable user memory is staged by the driver)
int k_host_alloc(size_t bytes, void** out) {
	*out = nullptr;
	CUDA_TRY(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
	return 0;
}
int k_host_free(void* p) {
	if (p) CUDA_TRY(cudaFreeHost(p));
	return 0;
}
