// peak.cu — measured FP64 FMA peak of the GPU (roofline denominator for the force kernels).
// MEASURED_PEAKS.json holds only HBM GB/s and bf16 TFLOP/s; BASELINE.md asks for a dependent-free DFMA kernel.
#include "ls1g_internal.cuh"

#define PEAK_ILP 8
#define PEAK_ITERS 16384

__global__ void __launch_bounds__(256) dfma_peak_kernel(double* out, double a, double b) {
	double x[PEAK_ILP];
#pragma unroll
	for (int k = 0; k < PEAK_ILP; ++k) x[k] = (double)(threadIdx.x + k) * 1e-3;
	for (int it = 0; it < PEAK_ITERS; ++it) {
#pragma unroll
		for (int k = 0; k < PEAK_ILP; ++k) x[k] = fma(x[k], a, b);
	}
	double s = 0.;
#pragma unroll
	for (int k = 0; k < PEAK_ILP; ++k) s += x[k];
	if (s == 12345.678) out[0] = s;   // never true; keeps the chain alive
}

int k_fp64_peak(int device, double* tflops) {
	CUDA_TRY(cudaSetDevice(device));
	cudaDeviceProp prop;
	CUDA_TRY(cudaGetDeviceProperties(&prop, device));
	double* d = nullptr;
	CUDA_TRY(cudaMalloc((void**)&d, 64));
	cudaEvent_t e0, e1;
	CUDA_TRY(cudaEventCreate(&e0)); CUDA_TRY(cudaEventCreate(&e1));
	const int blocks = prop.multiProcessorCount * 16, threads = 256;
	double best = 0.;
	for (int rep = 0; rep < 6; ++rep) {
		CUDA_TRY(cudaEventRecord(e0));
		dfma_peak_kernel<<<blocks, threads>>>(d, 0.999999, 1e-7);
		CUDA_TRY(cudaEventRecord(e1));
		CUDA_TRY(cudaEventSynchronize(e1));
		float ms = 0.f;
		CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
		double flops = 2.0 * (double)blocks * threads * PEAK_ITERS * PEAK_ILP;
		double tf = flops / (ms * 1e-3) / 1e12;
		if (rep > 0 && tf > best) best = tf;
	}
	cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
	*tflops = best;
	return 0;
}
