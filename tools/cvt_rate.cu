// Developer probe (not part of the product library): issue rates, per SM and clock, of the instructions the scan
// kernel chooses between for its f32 <-> f64 traffic.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
// -o tools/_timing/cvt_rate tools/cvt_rate.cu ; run on the GPU box, prints one line per variant.
#include <cuda_runtime.h>
#include <stdio.h>

#define ITERS 4096
#define CHAINS 8

__device__ __forceinline__ double widen_bits(float x)
{
    const unsigned u = __float_as_uint(x), a = u & 0x7fffffffu;
    const unsigned hi = (u & 0x80000000u) | ((a >> 3) + 0x38000000u);
    return __hiloint2double(a ? hi : u, u << 29);
}

template <int MODE>
__global__ void __launch_bounds__(256) probe(float *out, float seed, long long *cycles)
{
    float f[CHAINS];
    double d[CHAINS];
    for (int c = 0; c < CHAINS; ++c) f[c] = seed + c + threadIdx.x, d[c] = (double)f[c];
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (MODE == 0) {  // F2F.F64.F32 + F2F.F32.F64 (hardware, both directions)
                double w;
                asm volatile("cvt.f64.f32 %0, %1;" : "=d"(w) : "f"(f[c]));
                asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(f[c]) : "d"(w));
            } else if (MODE == 1) {  // F2F.F64.F32 only (the narrowing is an integer op)
                double w;
                asm volatile("cvt.f64.f32 %0, %1;" : "=d"(w) : "f"(f[c]));
                f[c] = __int_as_float(__double2hiint(w) ^ __double2loint(w));
            } else if (MODE == 2) {  // DFMA chain
                d[c] = fma(d[c], 1.0000001, 1e-9);
            } else if (MODE == 3) {  // integer widening + DADD (keeps the double live)
                d[c] += widen_bits(f[c]);
                f[c] = __int_as_float(__double2hiint(d[c]));
            } else if (MODE == 4) {  // FFMA chain
                f[c] = fmaf(f[c], 1.0000001f, 1e-9f);
            } else if (MODE == 5) {  // 64-bit shuffle (2 SHFL)
                d[c] = __shfl_down_sync(0xffffffffu, d[c], 1);
            } else if (MODE == 6) {  // cvt.rn.f32.f64 only (the widening is the integer form)
                float w;
                asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(w) : "d"(d[c]));
                d[c] = widen_bits(w) * 1.0000001;
            } else if (MODE == 7) {  // the scan's mix per transition: 3 widenings, 2 narrowings, 10 DFMA
                double w, w2;
                asm volatile("cvt.f64.f32 %0, %1;" : "=d"(w) : "f"(f[c]));
                asm volatile("cvt.f64.f32 %0, %1;" : "=d"(w2) : "f"(f[(c + 1) % CHAINS]));
                double a_ = fma(w, 0.98, w2) - d[c];
                double b_ = fma(d[c], 0.95, a_), c_ = fma(d[c], 0.98, w);
                b_ = fma(b_, 0.5, c_), c_ = fma(c_, 0.25, a_), a_ = fma(a_, b_, c_), b_ = fma(b_, a_, c_);
                float o1, o2;
                asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(o1) : "d"(a_));
                asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(o2) : "d"(b_));
                asm volatile("cvt.f64.f32 %0, %1;" : "=d"(w) : "f"(o1));
                d[c] = fma(w, w, d[c]) + w;
                f[c] = o2;
            }
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
    for (int c = 0; c < CHAINS; ++c) s += f[c] + (float)d[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int MODE>
void run(const char *name, int ops_per_iter)
{
    float *out;
    long long *cyc;
    const int blocks = 148 * 4;  // 4 CTAs of 256 threads per SM: 32 warps per SM
    cudaMalloc(&out, blocks * 256 * sizeof(float));
    cudaMalloc(&cyc, sizeof(long long));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    probe<MODE><<<blocks, 256>>>(out, 1.5f, cyc);
    // whole-kernel time with CUDA events (one warp's clock64 span is misleading: the scheduler favours the oldest warp)
    cudaEventRecord(e0);
    for (int r = 0; r < 5; ++r) probe<MODE><<<blocks, 256>>>(out, 1.5f, cyc);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double ops = 5.0 * blocks * 256.0 * ITERS * CHAINS * ops_per_iter;
    const double per_ns_sm = ops / (ms * 1e6) / 148.0;
    printf("%-44s %8.3f ms  %7.2f thread-ops/ns/SM  = %6.1f per clock per SM at %d MHz\n", name, ms / 5, per_ns_sm,
           per_ns_sm / (khz * 1e-6), khz / 1000);
    cudaFree(out), cudaFree(cyc);
}

int main()
{
    run<0>("cvt.f64.f32 + cvt.rn.f32.f64", 2);
    run<1>("cvt.f64.f32 (+2 int ops)", 1);
    run<2>("DFMA", 1);
    run<3>("integer widen + DADD", 1);
    run<4>("FFMA", 1);
    run<5>("shfl.down 64-bit", 1);
    run<6>("cvt.rn.f32.f64 (+ integer widen + DMUL)", 1);
    run<7>("scan mix: 3 widen + 2 narrow + 10 f64 ops", 1);
    return 0;
}
