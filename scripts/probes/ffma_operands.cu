// Micro-benchmark behind DESIGN.md section 4 (fp32 mode): does the FP32 pipe of a B200 keep its rate when EVERY FFMA reads
// three distinct registers (as the RK kernels do), and does the packed form fma.rn.f32x2 (SASS FFMA2) change that?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma_operands scripts/probes/ffma_operands.cu && ./ffma_operands
// Four loops over K register-resident values per thread, all dependent only through their own accumulator:
//   const2   a[j] = fma(a[j], m, c)            two operands are loop constants (what fes_probe_fma_peak measures)
//   reg3     a[j] = fma(a[j], b[j'], c[j''])   three distinct registers per FFMA
//   reg3x2   the same on float2 pairs with fma.rn.f32x2
//   mix      reg3 with every third FFMA replaced by FMUL + FADD on distinct registers (the kernels' 121 : 95 : 26 mix)
//   mixfma   the same arithmetic with the multiply written as fma(a, b, -0) and the add as fma(a, 1, c): FFMA only
//   mulonly / addonly   all FMUL, all FADD (which of the two costs the rate)
#include <cstdio>
#include <cuda_runtime.h>

constexpr int K = 12;

__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    float2 d;
    asm("{.reg .b64 ra, rb, rc, rd;\n\t"
        "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%6, %7};\n\t"
        "fma.rn.f32x2 rd, ra, rb, rc;\n\t"
        "mov.b64 {%0, %1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return d;
}

__device__ __forceinline__ float fma_asm(float a, float b, float c) {
    float d;
    asm("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

template <int MODE>
__global__ void probe(float *out, int iters, float seed) {
    float a[K], b[K], c[K];
#pragma unroll
    for (int j = 0; j < K; ++j) {
        a[j] = seed * (threadIdx.x + j + 1);
        b[j] = 0.999f + 1e-6f * j + seed;
        c[j] = 1e-7f * (j + 1) + seed;
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < K; ++j) {
            if (MODE == 0) a[j] = fmaf(a[j], 0.9999f, 1e-7f);
            if (MODE == 1) a[j] = fmaf(a[j], b[(j + 1) % K], c[(j + 5) % K]);
            if (MODE == 3) {
                if (j % 3 == 2) a[j] = __fmul_rn(a[j], b[(j + 1) % K]), a[j] = __fadd_rn(a[j], c[(j + 5) % K]);
                else a[j] = fmaf(a[j], b[(j + 1) % K], c[(j + 5) % K]);
            }
            if (MODE == 4) {
                if (j % 3 == 2) a[j] = fma_asm(a[j], b[(j + 1) % K], -0.0f), a[j] = fma_asm(a[j], 1.0f, c[(j + 5) % K]);
                else a[j] = fmaf(a[j], b[(j + 1) % K], c[(j + 5) % K]);
            }
            if (MODE == 5) a[j] = __fmul_rn(a[j], b[(j + 1) % K]);
            if (MODE == 6) a[j] = __fadd_rn(a[j], c[(j + 5) % K]);
        }
        if (MODE == 2) {
#pragma unroll
            for (int j = 0; j < K; j += 2) {
                float2 r = fma2(make_float2(a[j], a[j + 1]), make_float2(b[(j + 2) % K], b[(j + 3) % K]),
                                make_float2(c[(j + 6) % K], c[(j + 7) % K]));
                a[j] = r.x;
                a[j + 1] = r.y;
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < K; ++j) s += a[j];
    if (s == 123456.f) out[0] = s;
}

template <int MODE>
double run(const char *name, int sms, int threads, int blocks_per_sm, float *out) {
    const int iters = 20000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        probe<MODE><<<sms * blocks_per_sm, threads>>>(out, iters, 1e-9f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        const double fma_per_thread = (double)K * iters;   // one FMA (or FMUL + FADD) per value and iteration
        const double rate = fma_per_thread * threads * sms * blocks_per_sm / (ms * 1e-3);
        if (rep && rate > best) best = rate;
    }
    printf("%-8s threads/block %4d blocks/SM %2d : %7.2f G value-updates/s per SM-clock-normalised = %6.2f %% of 128 lanes x 1.965 GHz x %d SMs\n",
           name, threads, blocks_per_sm, best / 1e9, 100.0 * best / (128.0 * 1.965e9 * sms), sms);
    return best;
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float *out;
    cudaMalloc(&out, 4);
    for (int bps : {2, 8}) {
        run<0>("const2", sms, 256, bps, out);
        run<1>("reg3", sms, 256, bps, out);
        run<2>("reg3x2", sms, 256, bps, out);
        run<3>("mix", sms, 256, bps, out);
        run<4>("mixfma", sms, 256, bps, out);
        run<5>("mulonly", sms, 256, bps, out);
        run<6>("addonly", sms, 256, bps, out);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
