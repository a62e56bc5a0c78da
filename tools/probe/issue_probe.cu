// issue_probe.cu -- does a non-FP64 instruction issue "in the shadow" of a 2-cycle FP64 op?
// Kernels run NI integer ops (or MUFU / LDS) per 8 DFMA; compare times.  (exploration tool)
#include <cuda_runtime.h>
#include <cstdio>
template <int NI, int KIND>
__global__ void __launch_bounds__(256) probe(double* out, int* iout, double a, double b, int iters, int s) {
	double c0 = threadIdx.x, c1 = c0 + 1, c2 = c0 + 2, c3 = c0 + 3, c4 = c0 + 4, c5 = c0 + 5, c6 = c0 + 6, c7 = c0 + 7;
	int x0 = threadIdx.x, x1 = x0 + s, x2 = x1 + s, x3 = x2 + s;
	float f0 = threadIdx.x + 1.f, f1 = f0 + 1.f;
	__shared__ int sh[256];
	sh[threadIdx.x] = threadIdx.x;
	__syncthreads();
	for (int i = 0; i < iters; ++i) {
#pragma unroll
		for (int u = 0; u < 8; ++u) {
			c0 = fma(c0, a, b); c1 = fma(c1, a, b); c2 = fma(c2, a, b); c3 = fma(c3, a, b);
			c4 = fma(c4, a, b); c5 = fma(c5, a, b); c6 = fma(c6, a, b); c7 = fma(c7, a, b);
#pragma unroll
			for (int k = 0; k < NI; ++k) {
				if (KIND == 0) { if (k & 1) x0 = max(x0 ^ s, x1); else x1 = min(x1 + s, x0) ^ x2; x2 += x3 ^ s; k++; }
				else if (KIND == 1) { if (k & 1) f0 = rsqrtf(f0) + 1.5f; else f1 = rsqrtf(f1) + 1.5f; k++; }
				else if (KIND == 2) { x0 ^= sh[(x0 + k) & 255]; }
				else if (KIND == 3) { if (k & 1) f0 = fmaf(f0, 0.999f, 1e-3f); else f1 = fmaf(f1, 0.999f, 1e-3f); }
			}
		}
	}
	double sres = ((c0 + c1) + (c2 + c3)) + ((c4 + c5) + (c6 + c7));
	if (sres == -12345.678) out[0] = sres;
	if ((x0 ^ x1 ^ x2 ^ x3) == 0x12345 || f0 + f1 == -3.f) iout[0] = x0;
}
template <int NI, int KIND>
void run(const char* name) {
	double* d; int* di;
	cudaMalloc(&d, 64); cudaMalloc(&di, 64);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	int iters = 20000;
	probe<NI, KIND><<<148 * 8, 256>>>(d, di, 0.999999, 1e-9, 100, 3);
	cudaEventRecord(e0);
	probe<NI, KIND><<<148 * 8, 256>>>(d, di, 0.999999, 1e-9, iters, 3);
	cudaEventRecord(e1); cudaEventSynchronize(e1);
	float ms; cudaEventElapsedTime(&ms, e0, e1);
	double dfma = 64.0 * iters * 148 * 8 * 256;
	printf("%-28s NI=%d  %.3f ms  DFMA rate %.2f TF  cycles/8DFMA-warp(@1.965GHz) %.2f\n", name, NI, ms,
		   2 * dfma / ms / 1e9, ms * 1e-3 * 1.965e9 / (8.0 * iters * 8 * 256 / 32 / 4 * 1.0) );
}
__global__ void mufu_stats(double* out, int n) {
	// e = 1 - x*y0^2 for x = 2^k * (1 + u), u on a fine grid; out = {min e, max e, sum e, sum e^2}
	double mn = 1, mx = -1, s = 0, s2 = 0;
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		double u = (double)i / n;
		for (int k = 0; k < 2; ++k) {
			double x = (k ? 2.0 : 1.0) * (1.0 + u) * 1.0000001;
			double y; asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
			double e = fma(-x, y * y, 1.0);
			mn = fmin(mn, e); mx = fmax(mx, e); s += e; s2 += e * e;
		}
	}
	__shared__ double sh[4][256];
	sh[0][threadIdx.x] = mn; sh[1][threadIdx.x] = mx; sh[2][threadIdx.x] = s; sh[3][threadIdx.x] = s2;
	__syncthreads();
	if (threadIdx.x == 0) {
		for (int t = 1; t < 256; ++t) { mn = fmin(mn, sh[0][t]); mx = fmax(mx, sh[1][t]); s += sh[2][t]; s2 += sh[3][t]; }
		out[4 * blockIdx.x + 0] = mn; out[4 * blockIdx.x + 1] = mx; out[4 * blockIdx.x + 2] = s; out[4 * blockIdx.x + 3] = s2;
	}
}
int main() {
	{
		const int nb = 64, n = 1 << 26;
		double* d; cudaMalloc(&d, nb * 4 * sizeof(double));
		mufu_stats<<<nb, 256>>>(d, n);
		double h[nb * 4]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
		double mn = 1, mx = -1, s = 0, s2 = 0;
		for (int b = 0; b < nb; ++b) { mn = fmin(mn, h[4*b]); mx = fmax(mx, h[4*b+1]); s += h[4*b+2]; s2 += h[4*b+3]; }
		printf("MUFU.RSQ64H: e=1-x*y0^2  min %.4e (2^%.2f) max %.4e (2^%.2f) mean %.4e rms %.4e  1.875*max(e^2)=%.3e\n",
			   mn, log2(fabs(mn)), mx, log2(fabs(mx)), s / (2.0 * n), sqrt(s2 / (2.0 * n)), 1.875 * fmax(mn*mn, mx*mx));
	}
	run<0, 0>("dfma only");
	run<2, 0>("dfma + int(2 per 8)");
	run<4, 0>("dfma + int");
	run<8, 0>("dfma + int");
	run<16, 0>("dfma + int");
	run<2, 1>("dfma + mufu+fadd");
	run<4, 1>("dfma + mufu+fadd");
	run<4, 2>("dfma + lds");
	run<8, 2>("dfma + lds");
	run<4, 3>("dfma + ffma");
	run<8, 3>("dfma + ffma");
	run<16, 3>("dfma + ffma");
	return 0;
}
