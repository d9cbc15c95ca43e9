// flops.cu — the reference's FlopCounter as a device cell processor.
//
// Replaces (reference, src/): particleContainer/adapter/FlopCounter.cpp:196-421 (processCell / processCellPair / _calculatePairs)
// and FlopCounter.h:150-283 (the count -> FLOP model), i.e. `container->traverseCells(flopCounter)` as io/FlopRateWriter.cpp does it.
// It is the roofline numerator of this repository: the ALGORITHMIC (Newton-3) floating-point work of one force pass.
//
// The reference walks unordered cell pairs (C08) and skips halo-halo pairs; here every OWNED molecule scans its 27 cells and
// counts in half units: a partner in a non-halo cell is seen from both sides (1 + 1), a partner in a halo cell only from this
// side (2).  Macro calls: as kernel calls, except that a (non-halo, halo) cell pair only counts when index(non-halo) < index(halo)
// (FlopCounter.cpp:275-311).  All counters are integers, halved at the end: the result equals the reference's bit for bit.
#include "ls1g_internal.cuh"

#define FC_THREADS 128
#define FC_NV 15          // 2 x molecule distances | 2 x kernel calls [7] | 2 x macro calls [7]

__global__ void __launch_bounds__(FC_THREADS) flop_count_kernel(Grid g, size_t n, const double4* __restrict__ pos, const uint32_t* __restrict__ cell,
                                                                 const uint32_t* __restrict__ cell_start, const CompDev* __restrict__ comps,
                                                                 double rc2, double rclj2, unsigned long long* __restrict__ out) {
	__shared__ unsigned long long red[FC_NV];
	if (threadIdx.x < FC_NV) red[threadIdx.x] = 0ull;
	__syncthreads();
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	unsigned long long acc[FC_NV];
#pragma unroll
	for (int k = 0; k < FC_NV; ++k) acc[k] = 0ull;
	if (i < n && !(cell[i] & HALO_FLAG)) {
		const double4 pi = pos[i];
		const CompDev ci = comps[pos_cid(pi)];
		const uint32_t c = cell[i];
		const int nx = g.cpd[0], ny = g.cpd[1];
		const int cx = c % nx, cy = (c / nx) % ny, cz = c / (nx * ny);
		for (int dz = -1; dz <= 1; ++dz)
			for (int dy = -1; dy <= 1; ++dy)
				for (int dx = -1; dx <= 1; ++dx) {
					const uint32_t cj = (uint32_t)(((cz + dz) * ny + (cy + dy)) * nx + (cx + dx));
					const bool halo_j = grid_is_halo(g, cj);
					const unsigned long long wk = halo_j ? 2ull : 1ull;                      // kernel calls, molecule distances
					const unsigned long long wm = halo_j ? (c < cj ? 2ull : 0ull) : 1ull;      // macro calls
					for (uint32_t j = cell_start[cj]; j < cell_start[cj + 1]; ++j) {
						if (j == i) continue;
						const double4 pj = pos[j];
						const CompDev cjd = comps[pos_cid(pj)];
						const double ex = __dsub_rn(pi.x, pj.x), ey = __dsub_rn(pi.y, pj.y), ez = __dsub_rn(pi.z, pj.z);
						const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(ex, ex), __dmul_rn(ey, ey)), __dmul_rn(ez, ez));
						acc[0] += wk;
						unsigned long long k[7] = {0ull, 0ull, 0ull, 0ull, 0ull, 0ull, 0ull};
						if (d2 < rclj2) k[0] = (unsigned long long)(ci.n_lj * cjd.n_lj);
						if (d2 < rc2) {
							k[1] = (unsigned long long)(ci.n_ch * cjd.n_ch);
							k[2] = (unsigned long long)(ci.n_ch * cjd.n_di + ci.n_di * cjd.n_ch);
							k[3] = (unsigned long long)(ci.n_di * cjd.n_di);
							k[4] = (unsigned long long)(ci.n_ch * cjd.n_qu + ci.n_qu * cjd.n_ch);
							k[5] = (unsigned long long)(ci.n_di * cjd.n_qu + ci.n_qu * cjd.n_di);
							k[6] = (unsigned long long)(ci.n_qu * cjd.n_qu);
						}
#pragma unroll
						for (int p = 0; p < 7; ++p) { acc[1 + p] += wk * k[p]; acc[8 + p] += wm * k[p]; }
					}
				}
	}
#pragma unroll
	for (int k = 0; k < FC_NV; ++k) {
		unsigned long long v = acc[k];
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
		if ((threadIdx.x & 31) == 0 && v) atomicAdd(&red[k], v);
	}
	__syncthreads();
	if (threadIdx.x < FC_NV && red[threadIdx.x]) atomicAdd(&out[threadIdx.x], red[threadIdx.x]);
}

// multipliers: kernel, macro, sum force/torque, sum macro (FlopCounter.cpp:19-75), potentials in the order of FlopCounter.h:138-147
static const double FC_MULT[7][4] = {{15, 4, 12, 2}, {10, 2, 12, 2}, {34, 3, 15, 2}, {101, 5, 18, 3}, {52, 2, 15, 2}, {121, 2, 18, 2}, {131, 2, 18, 2}};

int k_flop_count(ls1g_ctx* c, ls1g_flop_counts* out) {
	memset(out, 0, sizeof(*out));
	unsigned long long* d = nullptr;
	CUDA_TRY(cudaMalloc((void**)&d, FC_NV * sizeof(unsigned long long)));
	CUDA_TRY(cudaMemsetAsync(d, 0, FC_NV * sizeof(unsigned long long), c->stream));
	if (c->n_slots) {
		timer_start(c, LS1G_T_AUX);
		flop_count_kernel<<<(unsigned)((c->n_slots + FC_THREADS - 1) / FC_THREADS), FC_THREADS, 0, c->stream>>>(
		    c->bgrid, c->n_slots, c->cur.pos, c->cur.cell, c->cell_start, c->comps, c->rc2, c->rclj2, d);
		timer_stop(c, LS1G_T_AUX, 1);
	}
	unsigned long long h[FC_NV];
	cudaError_t e = cudaMemcpyAsync(h, d, sizeof(h), cudaMemcpyDeviceToHost, c->stream);
	if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
	if (e == cudaSuccess) e = cudaGetLastError();
	cudaFree(d);
	if (e != cudaSuccess) LS1G_FAIL(std::string("ls1g_flop_count: ") + cudaGetErrorString(e));
	out->molecule_distances = 0.5 * (double)h[0];
	double flops = 8.0 * out->molecule_distances;
	for (int p = 0; p < 7; ++p) {
		out->kernel_calls[p] = 0.5 * (double)h[1 + p];
		out->macro_calls[p] = 0.5 * (double)h[8 + p];
		flops += out->kernel_calls[p] * (8.0 + FC_MULT[p][0] + FC_MULT[p][2]) + out->macro_calls[p] * (FC_MULT[p][1] + FC_MULT[p][3]);
	}
	out->flops = flops;
	return 0;
}
