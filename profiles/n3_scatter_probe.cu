// n3_scatter_probe.cu — what a Newton-3 variant of the single-centre list kernel would pay for its reaction forces.
//
// Owner-computes (lists.cuh) evaluates every pair from both sides and writes each molecule's force once, from registers, fused with
// the leapfrog.  A Newton-3 kernel evaluates a pair once and must ADD -f to the partner, which is owned by another lane and, for most
// partners, another block: 3 FP64 atomic adds per pair (red.global.add.f64), F zeroed before, and no fused integrator because F_i is
// complete only after the whole grid.  This probe times just that scatter on the bench's shape: n molecules, P partners each, partner
// slots within +-SPREAD of the own slot (the neighbourhood of a cell-sorted layout), against the same number of plain FP64 FMAs.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o n3_scatter_probe profiles/n3_scatter_probe.cu && ./n3_scatter_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned hash32(unsigned x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <bool ATOMIC>
__global__ void __launch_bounds__(512) scatter_kernel(size_t n, int partners, int spread, double* __restrict__ F, double* __restrict__ sink) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	double fx = 0., fy = 0., fz = 0.;
	for (int p = 0; p < partners; ++p) {
		const unsigned h = hash32((unsigned)i * 131u + (unsigned)p);
		long j = (long)i + (long)(h % (2u * (unsigned)spread)) - spread;
		j = j < 0 ? 0 : (j >= (long)n ? (long)n - 1 : j);
		const double f = 1e-3 * (double)(h & 1023u);
		fx = fma(f, 0.5, fx); fy = fma(f, 0.25, fy); fz = fma(f, 0.125, fz);
		if (ATOMIC) {
			atomicAdd(F + 3 * (size_t)j, -f * 0.5); atomicAdd(F + 3 * (size_t)j + 1, -f * 0.25); atomicAdd(F + 3 * (size_t)j + 2, -f * 0.125);
		}
	}
	if (ATOMIC) { atomicAdd(F + 3 * i, fx); atomicAdd(F + 3 * i + 1, fy); atomicAdd(F + 3 * i + 2, fz); }
	else { F[3 * i] = fx; F[3 * i + 1] = fy; F[3 * i + 2] = fz; }
	if (fx == 123.456) *sink = fx;
}

int main(int argc, char** argv) {
	const size_t n = argc > 1 ? (size_t)atol(argv[1]) : 10061824;
	const int partners = argc > 2 ? atoi(argv[2]) : 27;      // in-cutoff partners per molecule, each pair once (54 / 2)
	const int spread = argc > 3 ? atoi(argv[3]) : 3000;      // partner slots lie within the 9 neighbouring cell rows of a 512-molecule block
	double *F, *sink;
	cudaMalloc(&F, n * 3 * sizeof(double)); cudaMalloc(&sink, 8);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	const unsigned grid = (unsigned)((n + 511) / 512);
	float ms[2] = {0.f, 0.f};
	for (int variant = 0; variant < 2; ++variant) {
		for (int rep = 0; rep < 4; ++rep) {      // 3 warm-ups, the last one is reported
			cudaMemsetAsync(F, 0, n * 3 * sizeof(double));
			cudaEventRecord(e0);
			if (variant) scatter_kernel<true><<<grid, 512>>>(n, partners, spread, F, sink);
			else scatter_kernel<false><<<grid, 512>>>(n, partners, spread, F, sink);
			cudaEventRecord(e1);
			cudaEventSynchronize(e1);
			cudaEventElapsedTime(&ms[variant], e0, e1);
		}
	}
	if (cudaGetLastError() != cudaSuccess) { printf("{\"error\": \"cuda\"}\n"); return 1; }
	const double atomics = (double)n * (partners + 1) * 3;
	printf("{\"n_molecules\": %zu, \"pairs_per_molecule_once\": %d, \"slot_spread\": %d, \"fp64_atomic_adds\": %.0f, "
	       "\"ms_owner_style_plain_store\": %.3f, \"ms_newton3_atomic_scatter\": %.3f, \"ns_per_atomic\": %.4f}\n",
	       n, partners, spread, atomics, ms[0], ms[1], 1e6 * (ms[1] - ms[0]) / atomics);
	return 0;
}
