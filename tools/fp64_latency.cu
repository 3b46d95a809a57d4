// fp64 latency microbenchmark for the serial chains of the Riccati kernel (dependent DFMA, rsqrt,
// sqrt, reciprocal) -- one warp, clock64 around unrolled dependent chains.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_latency fp64_latency.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__global__ void chain(double* out, long long* cyc, double x0, double y0, int iters)
{
    double x = x0 + threadIdx.x * 1e-9, y = y0;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) {
            if (OP == 0) x = fma(x, y, y);            // dependent DFMA
            else if (OP == 1) x = rsqrt(x) + 1.0;     // rsqrt (+ DADD)
            else if (OP == 2) x = sqrt(x) + 1.0;      // sqrt
            else if (OP == 3) x = 1.0 / x + 1.0;      // reciprocal
            else if (OP == 4) x = x * y;              // DMUL
            else if (OP == 5) x = x + y;              // DADD
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
    out[threadIdx.x] = x;
}

template <int ILP>
__global__ void tput(double* out, long long* cyc, double y0, int iters)
{
    double x[ILP];
    for (int k = 0; k < ILP; k++) x[k] = 1.0 + threadIdx.x * 1e-9 + k;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++)
#pragma unroll
            for (int k = 0; k < ILP; k++) x[k] = fma(x[k], y0, y0);
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
    double s = 0; for (int k = 0; k < ILP; k++) s += x[k];
    out[threadIdx.x] = s;
}

int main()
{
    double* out; long long* cyc; long long h;
    cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 8);
    const int it = 1000;
    const char* names[] = {"DFMA dep", "rsqrt+DADD dep", "sqrt+DADD dep", "rcp+DADD dep", "DMUL dep", "DADD dep"};
#define RUN(OP) chain<OP><<<1, 32>>>(out, cyc, 1.5, 0.999999, it); chain<OP><<<1, 32>>>(out, cyc, 1.5, 0.999999, it); \
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("%-18s %7.1f cycles/op (1 warp)\n", names[OP], (double)h / (it * 16));
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5)
#define RUNT(ILP, W) tput<ILP><<<1, 32 * W>>>(out, cyc, 0.999999, it); tput<ILP><<<1, 32 * W>>>(out, cyc, 0.999999, it); \
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("DFMA ILP=%d warps=%d: %6.2f cycles per warp-DFMA per SMSP-warp\n", ILP, W, (double)h / (it * 8 * ILP));
    RUNT(1, 1) RUNT(2, 1) RUNT(4, 1) RUNT(8, 1) RUNT(16, 1) RUNT(8, 4) RUNT(8, 8) RUNT(8, 16)
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
