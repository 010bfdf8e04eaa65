// Scratch microbenchmark: dependent-issue latency of FP64 instructions on one warp (cycles per op).
#include <cstdio>
#include <cuda_runtime.h>
__global__ void lat(double* out, long long* cyc, double a, double b, int iters)
{
	double v = threadIdx.x;
	long long t0 = clock64();
	for (int i = 0; i < iters; ++i)
	{
#pragma unroll
		for (int k = 0; k < 16; ++k) v = fma(v, a, b);
	}
	long long t1 = clock64();
	double w = threadIdx.x + 1.0;
	for (int i = 0; i < iters; ++i)
	{
#pragma unroll
		for (int k = 0; k < 16; ++k) w = w * a;
	}
	long long t2 = clock64();
	double r = threadIdx.x + 2.0;
	for (int i = 0; i < iters; ++i)
	{
#pragma unroll
		for (int k = 0; k < 4; ++k)
		{
			double y;
			asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(r));
			r = y + 1.5;
		}
	}
	long long t3 = clock64();
	double q = threadIdx.x + 3.0;
	for (int i = 0; i < iters; ++i)
	{
#pragma unroll
		for (int k = 0; k < 4; ++k) q = 1.0 / (q + 1.25);
	}
	long long t4 = clock64();
	double m = threadIdx.x * 0.5;
	for (int i = 0; i < iters; ++i)
	{
#pragma unroll
		for (int k = 0; k < 16; ++k) m = fmax(m * a, b);
	}
	long long t5 = clock64();
	if (threadIdx.x == 0)
	{
		cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; cyc[4] = t5 - t4;
	}
	out[threadIdx.x] = v + w + r + q + m;
}
int main()
{
	double* out; long long* cyc;
	cudaMalloc(&out, 1024 * 8); cudaMallocManaged(&cyc, 64);
	const int iters = 4096;
	for (int threads : {32, 128, 512})
	{
		lat<<<1, threads>>>(out, cyc, 1.0000001, 1e-9, iters);
		cudaDeviceSynchronize();
		printf("threads %d: DFMA dep %.2f cyc/op, DMUL dep %.2f, (MUFU.RCP64H + DADD) %.2f, IEEE div + DADD %.2f, (DMUL + DMNMX) %.2f\n", threads,
			   cyc[0] / (16.0 * iters), cyc[1] / (16.0 * iters), cyc[2] / (4.0 * iters), cyc[3] / (4.0 * iters), cyc[4] / (16.0 * iters));
	}
	return 0;
}
