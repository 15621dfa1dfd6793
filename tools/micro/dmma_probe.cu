// Microbenchmark: fp64 mma.sync m8n8k4 (DMMA) throughput and dependent-issue latency on this GPU, alone and beside
// plain DFMA work, to decide whether the 8x8x64 turn-scoring contraction of the E-step belongs on it.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b)
{
	asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int NACC, bool WITH_FMA>
__global__ void probe(double* out, int iters, double a, double b)
{
	double c0[NACC], c1[NACC], f[4];
#pragma unroll
	for (int k = 0; k < NACC; k++) c0[k] = threadIdx.x + k, c1[k] = k;
#pragma unroll
	for (int k = 0; k < 4; k++) f[k] = threadIdx.x * 0.5 + k;
	for (int i = 0; i < iters; i++)
	{
#pragma unroll
		for (int k = 0; k < NACC; k++) dmma(c0[k], c1[k], a, b);
		if (WITH_FMA)
		{
#pragma unroll
			for (int k = 0; k < 4; k++) f[k] = fma(f[k], a, b);
#pragma unroll
			for (int k = 0; k < 4; k++) f[k] = fma(f[k], a, b);
		}
	}
	double s = 0;
#pragma unroll
	for (int k = 0; k < NACC; k++) s += c0[k] + c1[k];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s + f[0] + f[1] + f[2] + f[3];
}

__global__ void fma_only(double* out, int iters, double a, double b)
{
	double f[8];
#pragma unroll
	for (int k = 0; k < 8; k++) f[k] = threadIdx.x * 0.5 + k;
	for (int i = 0; i < iters; i++)
#pragma unroll
		for (int k = 0; k < 8; k++) f[k] = fma(f[k], a, b);
	double s = 0;
#pragma unroll
	for (int k = 0; k < 8; k++) s += f[k];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename K>
float run(K kern, int blocks, int threads, double* out, int iters)
{
	cudaEvent_t t0, t1;
	cudaEventCreate(&t0), cudaEventCreate(&t1);
	float best = 1e30f;
	for (int rep = 0; rep < 4; rep++)
	{
		cudaEventRecord(t0);
		kern<<<blocks, threads>>>(out, iters, 0.999999, 1e-9);
		cudaEventRecord(t1);
		cudaEventSynchronize(t1);
		float ms;
		cudaEventElapsedTime(&ms, t0, t1);
		if (rep && ms < best) best = ms;
	}
	return best;
}

int main()
{
	cudaDeviceProp p;
	cudaGetDeviceProperties(&p, 0);
	const int sm = p.multiProcessorCount;
	double* out;
	cudaMalloc(&out, (size_t)sm * 16 * 256 * sizeof(double));
	const int iters = 1 << 13;
	const double clk = p.clockRate * 1e3;
	{
		const float ms = run(fma_only, sm * 8, 256, out, iters);
		printf("DFMA only, 8 chains x 8 warps x 8 CTAs/SM: %.2f TFLOP/s\n", 2.0 * 8 * iters * sm * 8 * 256 / (ms * 1e-3) * 1e-12);
	}
	{
		const float ms = run(probe<8, false>, sm * 8, 256, out, iters);
		const double n = 8.0 * iters * sm * 8 * 8;  // warp-level DMMAs
		printf("DMMA m8n8k4 only, 8 accumulators: %.2f TFLOP/s (%.2f clk per DMMA per SM)\n", n * 512 / (ms * 1e-3) * 1e-12, ms * 1e-3 * clk / (n / sm));
	}
	{
		const float ms = run(probe<1, false>, sm, 32, out, iters);
		printf("DMMA dependent chain, 1 warp per SM: %.1f clk latency\n", ms * 1e-3 * clk / iters);
	}
	{
		const float ms = run(probe<4, true>, sm * 8, 256, out, iters);
		const double n = 4.0 * iters * sm * 8 * 8;
		printf("DMMA (4 acc) + 8 DFMA per iteration: %.2f ms; DMMA part %.2f TFLOP/s, DFMA part %.2f TFLOP/s\n", ms, n * 512 / (ms * 1e-3) * 1e-12,
		       2.0 * 8 * iters * sm * 8 * 256 / (ms * 1e-3) * 1e-12);
	}
	{
		const float ms = run(probe<4, false>, sm * 8, 256, out, iters);
		printf("DMMA (4 acc) alone: %.2f ms\n", ms);
	}
	return 0;
}
