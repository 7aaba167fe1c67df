// FP64 issue-rate microbenchmark: W warps on ONE SM, each running C independent dependent-DADD chains.
// cycles per warp-instruction = elapsed / (iters * C * 4) per warp; with enough chains the pipe's issue interval shows.
#include <cstdio>
#include <cuda_runtime.h>
template <int C>
__global__ void k(double *out, long long *cyc, double b)
{
    double x[C];
#pragma unroll
    for (int c = 0; c < C; c++) x[c] = b + c + threadIdx.x;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 512; i++) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int c = 0; c < C; c++) x[c] = x[c] + b;
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int c = 0; c < C; c++) s += x[c];
    out[threadIdx.x] = s;
    if ((threadIdx.x & 31) == 0) cyc[threadIdx.x >> 5] = t1 - t0;
}
template <int C> void run(int warps, int lanes, double *o, long long *c)
{
    k<C><<<1, warps * 32 - (32 - lanes), 0>>>(o, c, 1.0000002);
    long long h[32]; cudaMemcpy(h, c, sizeof(h), cudaMemcpyDeviceToHost);
    long long mx = 0; for (int w = 0; w < warps; w++) mx = h[w] > mx ? h[w] : mx;
    printf("warps %2d (last warp %2d lanes) chains %d: %.2f cycles per DADD warp-instruction per warp, %.2f per SM\n", warps, lanes, C,
           mx / (512.0 * 4 * C), mx / (512.0 * 4 * C * warps));
}
int main()
{
    double *o; long long *c; cudaMalloc(&o, 2048 * 8); cudaMalloc(&c, 32 * 8);
    for (int w : {1, 2, 4, 8, 16}) { run<1>(w, 32, o, c); run<4>(w, 32, o, c); run<8>(w, 32, o, c); }
    run<8>(1, 16, o, c); run<8>(2, 16, o, c);
    return 0;
}
