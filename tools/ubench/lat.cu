// latency microbenchmarks (one warp): dependent FP64 add / mul / fmin chains, shared-memory store->load round trip
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double *out, long long *cyc, double a, double b)
{
    __shared__ double sm[64];
    double x = a;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; i++) { x = x + b; x = x + b; x = x + b; x = x + b; }
    long long t1 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; i++) { x = x * b; x = x * b; x = x * b; x = x * b; }
    long long t2 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; i++) { x = fmin(x, b); x = fmax(x, a); x = fmin(x, b); x = fmax(x, a); }
    long long t3 = clock64();
    volatile double *vs = sm;
#pragma unroll 1
    for (int i = 0; i < 256; i++) { vs[threadIdx.x] = x; x = vs[threadIdx.x] + b; vs[threadIdx.x] = x; x = vs[threadIdx.x] + b; }
    long long t4 = clock64();
    float f = (float)a;
#pragma unroll 1
    for (int i = 0; i < 256; i++) { f = f + 1.5f; f = f + 1.5f; f = f + 1.5f; f = f + 1.5f; }
    long long t5 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; i++) { x = x < a ? a : (x > b ? b : x); x = x + b; x = x < a ? a : (x > b ? b : x); x = x + b; }
    long long t6 = clock64();
    if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; cyc[4] = t5 - t4; cyc[5] = t6 - t5; }
    out[threadIdx.x] = x + f;
}
int main()
{
    double *o; long long *c; cudaMalloc(&o, 256 * 8); cudaMalloc(&c, 64);
    for (int threads : {1, 32}) {
        k<<<1, threads>>>(o, c, 1.0000001, 1.0000002);
        long long h[6]; cudaMemcpy(h, c, 48, cudaMemcpyDeviceToHost);
        printf("threads %d: dadd %.1f  dmul %.1f  dfmin/fmax %.1f  smem st->ld + dadd %.1f  fadd %.1f  ternary clamp + dadd %.1f cycles per op\n", threads,
               h[0] / 1024.0, h[1] / 1024.0, h[2] / 1024.0, h[3] / 512.0, h[4] / 1024.0, h[5] / 512.0);
    }
    return 0;
}
