// Microbenchmarks that size the design: FP64 FMA vs DMMA issue rates and
// cp.async.bulk streaming bandwidth on B200. Build: nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__global__ void dfma_kernel(double *out, int iters) {
	double a[8];
	for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-3 + i;
	double b = 1.0000001, c = 1e-9;
	for (int it = 0; it < iters; ++it) {
#pragma unroll
		for (int i = 0; i < 8; ++i) a[i] = fma(a[i], b, c);
	}
	double s = 0;
	for (int i = 0; i < 8; ++i) s += a[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
	asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
	             : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int NACC> __global__ void dmma_kernel(double *out, int iters) {
	double c0[NACC], c1[NACC];
	for (int i = 0; i < NACC; ++i) { c0[i] = 0; c1[i] = 0; }
	double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-6;
	for (int it = 0; it < iters; ++it) {
#pragma unroll
		for (int i = 0; i < NACC; ++i) dmma884(c0[i], c1[i], a, b);
	}
	double s = 0;
	for (int i = 0; i < NACC; ++i) s += c0[i] + c1[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void dmma16816(double (&c)[4], const double (&a)[8], const double (&b)[4]) {
	asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
	             : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
	             : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
	               "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}
__global__ void dmma16_kernel(double *out, int iters) {
	double c[4][4];
	double a[8], b[4];
	for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-3 + i;
	for (int i = 0; i < 4; ++i) { b[i] = 1.0 + i * 1e-6; for (int j = 0; j < 4; ++j) c[i][j] = 0; }
	for (int it = 0; it < iters; ++it) {
#pragma unroll
		for (int i = 0; i < 4; ++i) dmma16816(c[i], a, b);
	}
	double s = 0;
	for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// mixed: DMMA and DFMA in the same warp, to see whether they share a pipe
__global__ void mixed_kernel(double *out, int iters) {
	double c0[4], c1[4], f[8];
	for (int i = 0; i < 4; ++i) { c0[i] = 0; c1[i] = 0; }
	for (int i = 0; i < 8; ++i) f[i] = threadIdx.x * 1e-3 + i;
	double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-6;
	for (int it = 0; it < iters; ++it) {
#pragma unroll
		for (int i = 0; i < 4; ++i) dmma884(c0[i], c1[i], a, b);
#pragma unroll
		for (int i = 0; i < 8; ++i) f[i] = fma(f[i], b, a);
	}
	double s = 0;
	for (int i = 0; i < 4; ++i) s += c0[i] + c1[i];
	for (int i = 0; i < 8; ++i) s += f[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---- streaming read: plain 128-bit loads ----
__global__ void ldg_stream(const double2 *__restrict__ in, size_t n2, double *out) {
	double s = 0;
	size_t stride = (size_t)gridDim.x * blockDim.x;
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	for (; i + 3 * stride < n2; i += 4 * stride) {
		double2 a = __ldg(in + i), b = __ldg(in + i + stride), c = __ldg(in + i + 2 * stride), d = __ldg(in + i + 3 * stride);
		s += a.x + a.y + b.x + b.y + c.x + c.y + d.x + d.y;
	}
	for (; i < n2; i += stride) { double2 a = __ldg(in + i); s += a.x + a.y; }
	if (s == 123.456) out[0] = s;
}

// ---- streaming read: cp.async.bulk ring, one producer thread, consumers sum the stage ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int cnt) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(cnt)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t *b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
	asm volatile("{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
template <int NS> __global__ void bulk_stream(const char *__restrict__ in, size_t bytes, int stage_bytes, int chunk, double *out) {
	extern __shared__ __align__(128) unsigned char smem[];
	__shared__ uint64_t full[NS], empty[NS];
	int nthr = blockDim.x;
	if (threadIdx.x == 0) {
		for (int s = 0; s < NS; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], nthr - 32); }
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	size_t ntiles = bytes / stage_bytes;
	double s = 0;
	if (threadIdx.x < 32) {
		if (threadIdx.x == 0) {
			int it = 0;
			for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
				int st = it % NS; uint32_t ph = (it / NS) & 1;
				mbar_wait(&empty[st], ph ^ 1);
				mbar_expect_tx(&full[st], stage_bytes);
				for (int o = 0; o < stage_bytes; o += chunk)
					bulk_g2s(smem + (size_t)st * stage_bytes + o, in + t * stage_bytes + o, chunk, &full[st]);
			}
		}
	} else {
		int it = 0, ctid = threadIdx.x - 32, cn = nthr - 32;
		for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
			int st = it % NS; uint32_t ph = (it / NS) & 1;
			mbar_wait(&full[st], ph);
			const double2 *p = (const double2 *)(smem + (size_t)st * stage_bytes);
			for (int i = ctid; i < stage_bytes / 16; i += cn) { double2 v = p[i]; s += v.x + v.y; }
			mbar_arrive(&empty[st]);
		}
	}
	if (s == 123.456) out[0] = s;
}

int main() {
	cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
	int sms = prop.multiProcessorCount;
	printf("device %s sms %d cc %d.%d clock %d kHz\n", prop.name, sms, prop.major, prop.minor, prop.clockRate);
	double *out; CK(cudaMalloc(&out, sizeof(double) * sms * 64 * 1024));
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	float ms;
	for (int rep = 0; rep < 2; ++rep) {
		int iters = 20000, blocks = sms * 8, thr = 256;
		cudaEventRecord(e0); dfma_kernel<<<blocks, thr>>>(out, iters); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
		cudaEventElapsedTime(&ms, e0, e1);
		printf("DFMA: %.2f TFLOP/s (%.3f ms)\n", 2.0 * 8 * iters * blocks * thr / ms / 1e9, ms);
		cudaEventRecord(e0); dmma_kernel<8><<<blocks, thr>>>(out, iters); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
		cudaEventElapsedTime(&ms, e0, e1);
		printf("DMMA m8n8k4 x8 acc: %.2f TFLOP/s (%.3f ms)  [%.2f SM-cycles/mma @1.9GHz]\n", 2.0 * 256 * 8 * iters * blocks * (thr / 32) / ms / 1e9, ms,
		       ms * 1e-3 * 1.9e9 / (8.0 * iters * blocks * (thr / 32) / sms));
		cudaEventRecord(e0); dmma_kernel<2><<<blocks, thr>>>(out, iters); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
		cudaEventElapsedTime(&ms, e0, e1);
		printf("DMMA m8n8k4 x2 acc: %.2f TFLOP/s (%.3f ms)\n", 2.0 * 256 * 2 * iters * blocks * (thr / 32) / ms / 1e9, ms);
		cudaEventRecord(e0); dmma_kernel<1><<<sms * 4, 128>>>(out, iters); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
		cudaEventElapsedTime(&ms, e0, e1);
		printf("DMMA m8n8k4 dependent chain, 4 warps/SMSP-set: latency ~%.1f cycles @1.9GHz\n", ms * 1e-3 * 1.9e9 / iters);
		for (int wps : {4, 8, 16}) { // warps per SM with 36 independent accumulator pairs each (the p = 64 Gram step)
			cudaEventRecord(e0); dmma_kernel<36><<<sms, 32 * wps>>>(out, iters / 4);
			if (cudaGetLastError() != cudaSuccess) { // 36 accumulator pairs x 16 warps do not fit the register file
				printf("DMMA m8n8k4 x36 acc, %d warps/SM: not launchable (registers)\n", wps);
				continue;
			}
			cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
			cudaEventElapsedTime(&ms, e0, e1);
			printf("DMMA m8n8k4 x36 acc, %d warps/SM: %.2f TFLOP/s (%.3f ms)\n", wps, 2.0 * 256 * 36 * (iters / 4) * sms * wps / ms / 1e9, ms);
		}
		cudaEventRecord(e0); dmma16_kernel<<<blocks, thr>>>(out, iters / 4); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
		cudaEventElapsedTime(&ms, e0, e1);
		printf("DMMA m16n8k16 x4 acc: %.2f TFLOP/s (%.3f ms)\n", 2.0 * 2048 * 4 * (iters / 4) * blocks * (thr / 32) / ms / 1e9, ms);
		cudaEventRecord(e0); mixed_kernel<<<blocks, thr>>>(out, iters); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
		cudaEventElapsedTime(&ms, e0, e1);
		printf("mixed 4 DMMA + 8 DFMA per iter: %.3f ms  (DMMA-only est %.3f, DFMA-only est see above)\n", ms, 0.0);
	}
	// streaming
	size_t bytes = (size_t)4 << 30;
	char *buf; CK(cudaMalloc(&buf, bytes)); CK(cudaMemset(buf, 0, bytes));
	for (int rep = 0; rep < 2; ++rep) {
		cudaEventRecord(e0); ldg_stream<<<sms * 8, 512>>>((const double2 *)buf, bytes / 16, out); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
		cudaEventElapsedTime(&ms, e0, e1);
		printf("LDG.128 stream: %.1f GB/s\n", bytes / ms / 1e6);
	}
	{
		int cfgs[][4] = {{32768, 32768, 1, 288}, {32768, 4096, 1, 288}, {32768, 2048, 1, 288}, {65536, 65536, 1, 288}, {16384, 16384, 2, 160}, {32768, 2048, 2, 160}, {49152, 2048, 1, 544}};
		for (auto &c : cfgs) {
			int stage = c[0], chunk = c[1], cps = c[2], thr = c[3];
			for (int ns : {2, 4}) {
				if ((size_t)ns * stage * cps > 220 * 1024) continue;
				if (ns == 4) CK(cudaFuncSetAttribute(bulk_stream<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * stage));
				else CK(cudaFuncSetAttribute(bulk_stream<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * stage));
				for (int rep = 0; rep < 2; ++rep) {
					cudaEventRecord(e0);
					if (ns == 4) bulk_stream<4><<<sms * cps, thr, 4 * stage>>>(buf, bytes, stage, chunk, out);
					else bulk_stream<2><<<sms * cps, thr, 2 * stage>>>(buf, bytes, stage, chunk, out);
					cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
					cudaEventElapsedTime(&ms, e0, e1);
				}
				printf("bulk stream stage %d B chunk %d B stages %d ctas/SM %d threads %d: %.1f GB/s\n", stage, chunk, ns, cps, thr, bytes / ms / 1e6);
			}
		}
	}
	CK(cudaDeviceSynchronize());
	printf("done\n");
	return 0;
}
