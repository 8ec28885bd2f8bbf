// Probe: issue rate of the packed fp32 instructions (add/mul/fma.rn.f32x2 = FADD2 / FMUL2 / FFMA2) against their scalar
// forms on sm_100a, per SM sub-partition, at the occupancies the loss kernel runs at (16 / 24 / 32 warps per SM), plus
// LDS.64 / LDS.128 issue and a mix shaped like the SSIM tap loop (1 LDS.64 + 2 FMUL2 + 3 FADD2).
// Reported: cycles per warp instruction per SMSP.  1.0 = the issue limit; 2.0 = the pipe takes two cycles per instruction.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/f32x2_probe tools/f32x2_probe.cu && timeout 60 /tmp/f32x2_probe
#include <cuda_runtime.h>
#include <stdio.h>

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float lo(u64 v) { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); return a + b; }

constexpr int kIters = 2048, kAcc = 12;

template <int MODE>
__global__ void __launch_bounds__(256) probe(float* out, long long* cyc, float seed) {
    __shared__ __align__(16) float sm[256 * 4 + 64];
    for (int i = threadIdx.x; i < 256 * 4 + 64; i += blockDim.x) sm[i] = seed * i;
    __syncthreads();
    u64 acc[kAcc];
    float fa[kAcc];
#pragma unroll
    for (int i = 0; i < kAcc; ++i) { acc[i] = pk(seed + i, seed - i); fa[i] = seed + i; }
    const u64 m = pk(1.0001f, 0.9999f), c = pk(seed, -seed);
    const float fm = 1.0001f, fc = seed;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < kIters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < kAcc; ++i) fa[i] = fmaf(fa[i], fm, fc);
        } else if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < kAcc; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[i]) : "l"(m), "l"(c));
        } else if (MODE == 2) {
#pragma unroll
            for (int i = 0; i < kAcc; ++i) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(acc[i]) : "l"(c));
        } else if (MODE == 3) {
#pragma unroll
            for (int i = 0; i < kAcc; ++i) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(acc[i]) : "l"(m));
        } else if (MODE == 4) {      // scalar add
#pragma unroll
            for (int i = 0; i < kAcc; ++i) fa[i] = __fadd_rn(fa[i], fc);
        } else if (MODE == 5) {      // LDS.64, conflict-free
            const float2* p = reinterpret_cast<const float2*>(sm) + ((threadIdx.x + it) & 255);
#pragma unroll
            for (int i = 0; i < kAcc; ++i) {
                float2 v;
                asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"((unsigned)__cvta_generic_to_shared(p + i)));
                fa[i] += v.x + v.y;
            }
        } else if (MODE == 6) {      // tap-loop shape: 1 LDS.64 + 2 FMUL2 + 3 FADD2 per (tap, pair)
            const float2* p = reinterpret_cast<const float2*>(sm) + ((threadIdx.x + it) & 255);
#pragma unroll
            for (int i = 0; i < kAcc / 3; ++i) {
                float2 v;
                asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"((unsigned)__cvta_generic_to_shared(p + i)));
                const u64 x = pk(v.x, v.y);
                u64 xx, xy;
                asm volatile("mul.rn.f32x2 %0, %1, %1;" : "=l"(xx) : "l"(x));
                asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(xy) : "l"(x), "l"(m));
                asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(acc[3 * i]) : "l"(x));
                asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(acc[3 * i + 1]) : "l"(xx));
                asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(acc[3 * i + 2]) : "l"(xy));
            }
        } else if (MODE == 7) {      // the same tap-loop shape in scalar form on ONE image (half the data)
            const float* p = sm + ((threadIdx.x + it) & 255);
#pragma unroll
            for (int i = 0; i < kAcc / 3; ++i) {
                const float x = p[i * 32];
                const float xx = __fmul_rn(x, x), xy = __fmul_rn(x, fm);
                fa[3 * i] = __fadd_rn(fa[3 * i], x);
                fa[3 * i + 1] = __fadd_rn(fa[3 * i + 1], xx);
                fa[3 * i + 2] = __fadd_rn(fa[3 * i + 2], xy);
            }
        } else if (MODE == 8) {      // alternating FFMA (fma pipe) and IADD3/LOP3 (alu pipe)
#pragma unroll
            for (int i = 0; i < kAcc; ++i) {
                fa[i] = fmaf(fa[i], fm, fc);
                acc[i] = (acc[i] ^ (u64)it) + 12345ull;
            }
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kAcc; ++i) s += lo(acc[i]) + fa[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void run(const char* name, int inst_per_iter, int ctas_per_sm, float* out, long long* cyc, int sms) {
    const int grid = sms * ctas_per_sm;
    probe<MODE><<<grid, 256>>>(out, cyc, 1.0f);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<MODE><<<grid, 256>>>(out, cyc, 1.0f);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    long long h[4096];
    cudaMemcpy(h, cyc, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < grid; ++i) avg += (double)h[i];
    avg /= grid;
    // warps per SMSP = ctas_per_sm * 8 / 4; every warp issues kIters * inst_per_iter instructions of the kind under test
    const double winst_per_smsp = (double)ctas_per_sm * 2 * kIters * inst_per_iter;
    printf("%-34s %2d warps/SM: %7.3f cycles per warp instruction per SMSP (%.3f ms, err=%d)\n", name, ctas_per_sm * 8,
           avg / winst_per_smsp, ms, (int)cudaGetLastError());
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs, %d MHz\n", p.name, p.multiProcessorCount, p.clockRate / 1000);
    float* out; long long* cyc;
    cudaMalloc(&out, sizeof(float) * 4096 * 256);
    cudaMalloc(&cyc, sizeof(long long) * 4096);
    const int sms = p.multiProcessorCount;
    for (int c = 2; c <= 4; ++c) {
        run<0>("FFMA (scalar)", kAcc, c, out, cyc, sms);
        run<4>("FADD (scalar)", kAcc, c, out, cyc, sms);
        run<1>("FFMA2 (fma.rn.f32x2)", kAcc, c, out, cyc, sms);
        run<2>("FADD2 (add.rn.f32x2)", kAcc, c, out, cyc, sms);
        run<3>("FMUL2 (mul.rn.f32x2)", kAcc, c, out, cyc, sms);
        run<5>("LDS.64 (+2 FADD each)", kAcc, c, out, cyc, sms);
        run<6>("tap packed: LDS.64+2 FMUL2+3 FADD2", (kAcc / 3) * 6, c, out, cyc, sms);
        run<7>("tap scalar: LDS.32+2 FMUL+3 FADD", (kAcc / 3) * 6, c, out, cyc, sms);
        run<8>("FFMA + 2 int ops (alu pipe)", kAcc * 3, c, out, cyc, sms);
    }
    return 0;
}
