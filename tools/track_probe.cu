// Micro-benchmark: ALU cost per score of the (best, second, argmin) tracking variants considered for K1's epilogue.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/track_probe tools/track_probe.cu
// Each thread keeps 32 scores in registers, perturbs them on the FMA pipe every round (so nothing is loop
// invariant) and runs one tracker variant over them.  Output: cycles per score per SM sub-partition.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void track5(float v, int j, float &best, float &second, int &jloc) {
    asm("{\n\t.reg .pred p;\n\t.reg .f32 t;\n\t"
        "setp.lt.f32 p, %3, %0;\n\t"
        "max.f32 t, %0, %3;\n\t"
        "min.f32 %1, %1, t;\n\t"
        "min.f32 %0, %0, %3;\n\t"
        "@p mov.b32 %2, %4;\n\t}"
        : "+f"(best), "+f"(second), "+r"(jloc)
        : "f"(v), "r"(j));
}
// pair, floats, 3-input min: lo/hi, setp, max, min3, min, mov = 7 ops / 2 scores (which-of-pair unresolved)
__device__ __forceinline__ void track_pair_f(float v0, float v1, int j, float &best, float &second, int &jloc) {
    asm("{\n\t.reg .pred p;\n\t.reg .f32 t, lo, hi;\n\t"
        "min.f32 lo, %3, %4;\n\t"
        "max.f32 hi, %3, %4;\n\t"
        "setp.lt.f32 p, lo, %0;\n\t"
        "max.f32 t, %0, lo;\n\t"
        "min.f32 %1, %1, t, hi;\n\t"
        "min.f32 %0, %0, lo;\n\t"
        "@p mov.b32 %2, %5;\n\t}"
        : "+f"(best), "+f"(second), "+r"(jloc)
        : "f"(v0), "f"(v1), "r"(j));
}
// pair, order-preserving int view of non-negative floats, DPX min-with-predicate
__device__ __forceinline__ void track_pair_i(int v0, int v1, int j, int &best, int &second, int &jloc) {
    bool q, p;
    const int lo = __vibmin_s32(v0, v1, &q);   // q = v0 <= v1
    const int hi = max(v0, v1);
    const int t = max(best, lo);
    second = __vimin3_s32(second, t, hi);
    best = __vibmin_s32(lo, best, &p);          // p = lo <= best
    if (p) jloc = q ? j : j + 1;
}
// pair, ints, plain min/max + setp (no DPX)
__device__ __forceinline__ void track_pair_i2(int v0, int v1, int j, int &best, int &second, int &jloc) {
    const int lo = min(v0, v1);
    const int hi = max(v0, v1);
    const int t = max(best, lo);
    second = min(min(second, t), hi);
    if (lo < best) jloc = j;
    best = min(best, lo);
}

// pair, floats: lo / max / min3 / setp / min on the ALU pipe; hi = (v0 + v1) - lo and the index update on the FMA pipe
__device__ __forceinline__ void track_pair_fma(float v0, float v1, float jf, float &best, float &second, float &idxf,
                                               float one, float zero) {
    asm("{\n\t.reg .pred p;\n\t.reg .f32 t, lo, hi, s;\n\t"
        "min.f32 lo, %3, %4;\n\t"
        "add.f32 s, %3, %4;\n\t"
        "sub.f32 hi, s, lo;\n\t"
        "setp.lt.f32 p, lo, %0;\n\t"
        "max.f32 t, %0, lo;\n\t"
        "min.f32 %1, %1, t, hi;\n\t"
        "min.f32 %0, %0, lo;\n\t"
        "@p fma.rn.f32 %2, %6, %5, %7;\n\t}"
        : "+f"(best), "+f"(second), "+f"(idxf)
        : "f"(v0), "f"(v1), "f"(jf), "f"(one), "f"(zero));
}

// the same with t = max(best, lo) = (best + lo) - min(best, lo) on the FMA pipe as well
__device__ __forceinline__ void track_pair_fma2(float v0, float v1, int j, float &best, float &second, int &jloc) {
    asm("{\n\t.reg .pred p;\n\t.reg .f32 t, lo, hi, s, u, nb;\n\t"
        "min.f32 lo, %3, %4;\n\t"
        "add.f32 s, %3, %4;\n\t"
        "sub.f32 hi, s, lo;\n\t"
        "setp.lt.f32 p, lo, %0;\n\t"
        "min.f32 nb, %0, lo;\n\t"
        "add.f32 u, %0, lo;\n\t"
        "sub.f32 t, u, nb;\n\t"
        "min.f32 %1, %1, t, hi;\n\t"
        "mov.f32 %0, nb;\n\t"
        "@p mov.b32 %2, %5;\n\t}"
        : "+f"(best), "+f"(second), "+r"(jloc)
        : "f"(v0), "f"(v1), "r"(j));
}

template <int VAR>
__global__ void __launch_bounds__(512) probe(const float *in, float *out, int rounds, long long *cycles) {
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = in[(threadIdx.x + j * 256) & 4095];
    float best[2] = {1e30f, 1e30f}, second[2] = {1e30f, 1e30f};
    int ib[2] = {0x7f000000, 0x7f000000}, is[2] = {0x7f000000, 0x7f000000};
    int idx[2] = {0, 0};
    const long long t0 = clock64();
    for (int r = 0; r < rounds; ++r) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = fmaf(v[j], 0.99993f, 1e-3f);   // FMA pipe
        if (VAR == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
                track5(v[j], j, best[0], second[0], idx[0]);
                track5(v[j + 1], j + 1, best[1], second[1], idx[1]);
            }
        } else if (VAR == 1) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                track_pair_f(v[j], v[j + 1], j, best[0], second[0], idx[0]);
                track_pair_f(v[j + 2], v[j + 3], j + 2, best[1], second[1], idx[1]);
            }
        } else if (VAR == 2) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                track_pair_i(__float_as_int(v[j]), __float_as_int(v[j + 1]), j, ib[0], is[0], idx[0]);
                track_pair_i(__float_as_int(v[j + 2]), __float_as_int(v[j + 3]), j + 2, ib[1], is[1], idx[1]);
            }
        } else if (VAR == 4) {
            const float one = in[4095] * 0.f + 1.f + (float)(rounds >> 30), zero = (float)(rounds >> 30);
            float fi[2] = {(float)idx[0], (float)idx[1]};
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                track_pair_fma(v[j], v[j + 1], (float)j, best[0], second[0], fi[0], one, zero);
                track_pair_fma(v[j + 2], v[j + 3], (float)(j + 2), best[1], second[1], fi[1], one, zero);
            }
            // which element of the winning pair: an odd column attains the minimum iff min(odd columns) == best
            float mo = v[1];
#pragma unroll
            for (int j = 3; j < 32; j += 4) asm("min.f32 %0, %0, %1, %2;" : "+f"(mo) : "f"(v[j]), "f"(v[(j + 2) & 31]));
            idx[0] = (int)fi[0] + (mo == best[0] ? 1 : 0);
            idx[1] = (int)fi[1] + (mo == best[1] ? 1 : 0);
        } else if (VAR == 5) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                track_pair_fma2(v[j], v[j + 1], j, best[0], second[0], idx[0]);
                track_pair_fma2(v[j + 2], v[j + 3], j + 2, best[1], second[1], idx[1]);
            }
            float mo = v[1];
#pragma unroll
            for (int j = 3; j < 32; j += 4) asm("min.f32 %0, %0, %1, %2;" : "+f"(mo) : "f"(v[j]), "f"(v[(j + 2) & 31]));
            idx[0] += (mo == best[0] ? 1 : 0);
            idx[1] += (mo == best[1] ? 1 : 0);
        } else {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                track_pair_i2(__float_as_int(v[j]), __float_as_int(v[j + 1]), j, ib[0], is[0], idx[0]);
                track_pair_i2(__float_as_int(v[j + 2]), __float_as_int(v[j + 3]), j + 2, ib[1], is[1], idx[1]);
            }
        }
    }
    const long long t1 = clock64();
    float acc = best[0] + best[1] + second[0] + second[1] + (float)(idx[0] + idx[1]) +
                __int_as_float(ib[0]) + __int_as_float(ib[1]) + __int_as_float(is[0]) + __int_as_float(is[1]);
    out[blockIdx.x * 256 + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

int main(int argc, char **argv) {
    const int nthr = argc > 1 ? atoi(argv[1]) : 256;
    float *in, *out;
    long long *cyc, h;
    cudaMalloc(&in, 4096 * 4);
    cudaMalloc(&out, 148 * 256 * 4);
    cudaMalloc(&cyc, 8);
    float hin[4096];
    for (int i = 0; i < 4096; ++i) hin[i] = 100.f + (float)((i * 2654435761u) >> 20);
    cudaMemcpy(in, hin, sizeof(hin), cudaMemcpyHostToDevice);
    const int rounds = 2000;
    const char *names[6] = {"float 5-op tracker (current)", "float pair + min3", "int pair + DPX vibmin/vimin3",
                            "int pair plain", "float pair, hi + index on the FMA pipe", "float pair, hi and t on the FMA pipe"};
    for (int var = 0; var < 6; ++var) {
        for (int rep = 0; rep < 2; ++rep) {
            if (var == 0) probe<0><<<148, nthr>>>(in, out, rounds, cyc);
            if (var == 1) probe<1><<<148, nthr>>>(in, out, rounds, cyc);
            if (var == 2) probe<2><<<148, nthr>>>(in, out, rounds, cyc);
            if (var == 3) probe<3><<<148, nthr>>>(in, out, rounds, cyc);
            if (var == 4) probe<4><<<148, nthr>>>(in, out, rounds, cyc);
            if (var == 5) probe<5><<<148, nthr>>>(in, out, rounds, cyc);
            cudaDeviceSynchronize();
        }
        cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        // 256 threads = 8 warps = 2 warps per sub-partition; each warp handles 32 scores per round
        const double per = (double)h / ((double)rounds * 32.0 * (nthr / 128.0));
        printf("%-34s %8.3f cycles / warp-score / sub-partition  (err %d)\n", names[var], per, (int)cudaGetLastError());
    }
    return 0;
}
