// FP64 pipe probe for B200 (sm_100a): measures DFMA and DMMA (mma.sync f64) issue
// rates so DESIGN.md can state the FP64 roofline denominators from measurement.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_probe fp64_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

constexpr int ITERS = 4096;

__global__ void k_dfma(double* out, double a, double b) {
    double acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void k_dmma884(double* out, double a, double b) {
    double c[NACC][2];
#pragma unroll
    for (int i = 0; i < NACC; ++i) { c[i][0] = i; c[i][1] = threadIdx.x; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// GEMM-shaped register tile, no memory: FM x FN accumulators from FM distinct A and FN distinct B fragments
// (64 x 32 warp tile = 8 x 4), optionally fed from shared memory like the real main loop.
template <int FM, int FN, bool LDS>
__global__ void k_dmma_tile(double* out, double a, double b) {
    __shared__ double sA[64 * 16], sB[32 * 16];
    for (int i = threadIdx.x; i < 64 * 16; i += blockDim.x) sA[i] = a + i * 1e-9;
    for (int i = threadIdx.x; i < 32 * 16; i += blockDim.x) sB[i] = b + i * 1e-9;
    __syncthreads();
    double c[FM][FN][2];
#pragma unroll
    for (int i = 0; i < FM; ++i)
#pragma unroll
        for (int j = 0; j < FN; ++j) { c[i][j][0] = i; c[i][j][1] = j; }
    const int lane = threadIdx.x & 31, grp = lane >> 2, t4 = lane & 3;
    double af[FM], bf[FN];
#pragma unroll
    for (int i = 0; i < FM; ++i) af[i] = a + i;
#pragma unroll
    for (int j = 0; j < FN; ++j) bf[j] = b + j;
    for (int it = 0; it < ITERS; ++it) {
        if (LDS) {
            const int kk = it & 3;
#pragma unroll
            for (int i = 0; i < FM; ++i) af[i] = sA[(i * 8 + grp) * 16 + ((((kk * 2 + (t4 >> 1)) ^ ((grp & 3) << 1)) << 1) | (t4 & 1))];
#pragma unroll
            for (int j = 0; j < FN; ++j) bf[j] = sB[(j * 8 + grp) * 16 + ((((kk * 2 + (t4 >> 1)) ^ ((grp & 3) << 1)) << 1) | (t4 & 1))];
        }
#pragma unroll
        for (int i = 0; i < FM; ++i)
#pragma unroll
            for (int j = 0; j < FN; ++j)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(c[i][j][0]), "+d"(c[i][j][1]) : "d"(af[i]), "d"(bf[j]));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < FM; ++i)
#pragma unroll
        for (int j = 0; j < FN; ++j) s += c[i][j][0] + c[i][j][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void k_dmma1688(double* out, double a, double b) {
    double c[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; ++i) { c[i][0] = i; c[i][1] = threadIdx.x; c[i][2] = 1; c[i][3] = 2; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                         : "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void k_dmma1684(double* out, double a, double b) {
    double c[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; ++i) { c[i][0] = i; c[i][1] = threadIdx.x; c[i][2] = 1; c[i][3] = 2; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i)
            asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                         : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                         : "d"(a), "d"(b), "d"(a));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void k_dmma16816(double* out, double a, double b) {
    double c[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; ++i) { c[i][0] = i; c[i][1] = threadIdx.x; c[i][2] = 1; c[i][3] = 2; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                         : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                         : "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b),
                           "d"(a), "d"(b), "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static void run(const char* name, F launch, double flops_per_thread_iter, int blocks, int threads) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int w = 0; w < 3; ++w) launch();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    double flops = flops_per_thread_iter * ITERS * (double)blocks * threads;
    printf("%-28s blocks=%4d thr=%4d  %.3f ms  %.2f TFLOP/s\n", name, blocks, threads, best, flops / best * 1e-9);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    printf("device %s sms=%d clock=%d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
    int sms = p.multiProcessorCount;
    double* out; CK(cudaMalloc(&out, sizeof(double) * sms * 8 * 1024));
    for (int wps : {4, 8, 16}) {   // warps per SM
        int thr = wps * 32, blocks = sms;
        printf("--- %d warps/SM\n", wps);
        run("dfma x16", [&] { k_dfma<<<blocks, thr>>>(out, 1.0000001, 1e-9); }, 16 * 2.0, blocks, thr);
        run("dmma m8n8k4 x8", [&] { k_dmma884<8><<<blocks, thr>>>(out, 1.0000001, 1e-9); }, 8 * 512.0 / 32, blocks, thr);
        run("dmma m8n8k4 x32", [&] { k_dmma884<32><<<blocks, thr>>>(out, 1.0000001, 1e-9); }, 32 * 512.0 / 32, blocks, thr);
        run("dmma tile 8x4 regs", [&] { k_dmma_tile<8, 4, false><<<blocks, thr>>>(out, 1.0000001, 1e-9); }, 32 * 512.0 / 32, blocks, thr);
        run("dmma tile 8x4 lds", [&] { k_dmma_tile<8, 4, true><<<blocks, thr>>>(out, 1.0000001, 1e-9); }, 32 * 512.0 / 32, blocks, thr);
        run("dmma tile 8x2 lds", [&] { k_dmma_tile<8, 2, true><<<blocks, thr>>>(out, 1.0000001, 1e-9); }, 16 * 512.0 / 32, blocks, thr);
        run("dmma m16n8k4 x16", [&] { k_dmma1684<16><<<blocks, thr>>>(out, 1.0000001, 1e-9); }, 16 * 1024.0 / 32, blocks, thr);
        run("dmma m16n8k8 x16", [&] { k_dmma1688<16><<<blocks, thr>>>(out, 1.0000001, 1e-9); }, 16 * 2048.0 / 32, blocks, thr);
        run("dmma m16n8k16 x16", [&] { k_dmma16816<16><<<blocks, thr>>>(out, 1.0000001, 1e-9); }, 16 * 4096.0 / 32, blocks, thr);
    }
    CK(cudaFree(out));
    return 0;
}
