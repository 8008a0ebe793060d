// write-pattern probe: the emit pass's store pattern without its arithmetic
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
typedef long long i64;
// mode 0: lane stores 8 B, 4 stores per pod (nodes lane+32j of the warp's 128); mode 1: lane stores 16 B twice (adjacent nodes)
template <int MODE>
__global__ void __launch_bounds__(256, 3) pat(i64* out, i64 stride, int pods_per_cta, i64 n_pods) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const i64 tile0 = (i64)blockIdx.x * 1024 + warp * 128;
    i64 p0 = (i64)blockIdx.y * pods_per_cta, p1 = p0 + pods_per_cta;
    if (p1 > n_pods) p1 = n_pods;
    i64 v = lane + blockIdx.x;
    for (i64 p = p0; p < p1; ++p) {
        i64* row = out + p * stride + tile0;
        if (MODE == 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j) __stcs(row + 32 * j + lane, v + j);
        } else {
            longlong2* r2 = (longlong2*)row;
            __stcs(r2 + lane, make_longlong2(v, v + 1));
            __stcs(r2 + 32 + lane, make_longlong2(v + 2, v + 3));
        }
        v += 3;
    }
}
__global__ void fill(i64* out, i64 n) {
    i64 i = ((i64)blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (i + 1 < n) *(longlong2*)(out + i) = make_longlong2(7, 7);
}
int main(int argc, char** argv) {
    const i64 N = argc > 1 ? atoll(argv[1]) : 5000, P = argc > 2 ? atoll(argv[2]) : 100000;
    const i64 n_pad = (N + 1023) / 1024 * 1024, tiles = n_pad / 1024;
    i64* out;
    cudaMalloc(&out, sizeof(i64) * n_pad * P);
    cudaEvent_t a, b;
    cudaEventCreate(&a), cudaEventCreate(&b);
    for (int waves : {24, 8, 2}) {
        const i64 slots = 3 * 148 * waves, chunks = slots / tiles > 0 ? slots / tiles : 1;
        i64 ppc = (P + chunks - 1) / chunks;
        ppc = (ppc + 63) / 64 * 64;
        dim3 grid((unsigned)tiles, (unsigned)((P + ppc - 1) / ppc));
        for (int mode = 0; mode < 2; ++mode) {
            float best = 1e9;
            for (int r = 0; r < 6; ++r) {
                cudaEventRecord(a);
                if (mode == 0) pat<0><<<grid, 256>>>(out, n_pad, (int)ppc, P);
                else pat<1><<<grid, 256>>>(out, n_pad, (int)ppc, P);
                cudaEventRecord(b);
                cudaEventSynchronize(b);
                float ms;
                cudaEventElapsedTime(&ms, a, b);
                if (ms < best) best = ms;
            }
            printf("nodes %lld pods %lld waves %d ppc %lld mode %d: %.1f us = %.0f GB/s\n", N, P, waves, ppc, mode, best * 1e3, n_pad * P * 8 / best / 1e6);
        }
    }
    float best = 1e9;
    for (int r = 0; r < 6; ++r) {
        cudaEventRecord(a);
        fill<<<(unsigned)((n_pad * P / 2 + 255) / 256), 256>>>(out, n_pad * P);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    printf("sequential fill: %.1f us = %.0f GB/s\n", best * 1e3, n_pad * P * 8 / best / 1e6);
    return 0;
}
