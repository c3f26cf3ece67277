// update.cu -- K2 (per-cluster sum / count / inertia) and K3 (finalize + shift).
//
// Replaces _centers_dense (reference dask_ml/cluster/k_means.py:622-632: serial scatter-add into
// a float64 table), the chain-sum of block partials (:595-597), da.bincount (:598),
// maximum(counts, 1) + divide (:600-601), squared_norm(centers - new_centers) (:605) and
// distances.sum() (:616).
//
// acc layout (float64): [k*d sums | k counts | 1 inertia]  -- one buffer, one allreduce.
//
// Three strategies (update_seg_kernel, for tables too big for shared memory, is described below):
//   update_owner_kernel  (k*d*8 + k*4 <= ~200 KB of shared memory): a per-CTA float64 table in
//       shared memory.  Warp w of the CTA OWNS the clusters with label % 8 == w: every warp scans
//       the label stream (coalesced, L1-shared), ballots the rows it owns and adds those rows
//       (lanes = features, coalesced) into its table rows.  No atomics at all until the final
//       flush (one fp64 atomicAdd per table cell per CTA).
//   update_atomic_kernel (any k*d): warp per row, lanes = features, fp64 RED into the global
//       table (L2-resident), counts via one RED per row.
#include "common.cuh"
#include <cooperative_groups.h>
#include <type_traits>
namespace cg = cooperative_groups;

int km_k2_variant();  // assign.cu (b200km_set_option)

constexpr int UP_THREADS = 256;
constexpr int UP_WARPS = UP_THREADS / 32;

template <typename T>
__global__ void __launch_bounds__(UP_THREADS)
update_owner_kernel(const T *__restrict__ X, int64_t n, int d, int64_t ldx, const int32_t *__restrict__ labels,
                    const T *__restrict__ mind2, int k, double *__restrict__ acc, int scale_sqrt) {
    extern __shared__ __align__(16) double tbl[];       // [k*d]
    int *cnt = reinterpret_cast<int *>(tbl + (size_t)k * d);  // [k]
    __shared__ double red[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int e = tid; e < k * d; e += UP_THREADS) tbl[e] = 0.0;
    for (int e = tid; e < k; e += UP_THREADS) cnt[e] = 0;
    __syncthreads();

    double inert = 0.0;
    constexpr int CH = 1024;  // rows per CTA step
    for (int64_t base = (int64_t)blockIdx.x * CH; base < n; base += (int64_t)gridDim.x * CH) {
#pragma unroll 1
        for (int sub = 0; sub < CH / 32; ++sub) {
            const int64_t row_l = base + sub * 32 + lane;
            int lab = -1;
            if (row_l < n) lab = __ldg(labels + row_l);
            const bool mine = lab >= 0 && (lab & (UP_WARPS - 1)) == warp;
            unsigned mask = __ballot_sync(0xffffffffu, mine);
            if (mine && mind2) {
                const double v = (double)__ldg(mind2 + row_l);
                inert += scale_sqrt ? sqrt(v) : v;
            }
            while (mask) {
                // take up to 4 owned rows at once so their loads overlap
                int b[4];
                int nb = 0;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (mask) { b[u] = __ffs(mask) - 1; mask &= mask - 1; nb = u + 1; } else b[u] = -1;
                }
                int l4[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) l4[u] = __shfl_sync(0xffffffffu, lab, b[u] < 0 ? 0 : b[u]);
                for (int f0 = 0; f0 < d; f0 += 32) {
                    const int f = f0 + lane;
                    T v[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        v[u] = (T)0;
                        if (u < nb && f < d) v[u] = __ldg(X + (base + sub * 32 + b[u]) * ldx + f);
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (u < nb && f < d) tbl[(size_t)l4[u] * d + f] += (double)v[u];
                }
                if (lane == 0) {
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (u < nb) cnt[l4[u]] += 1;
                }
                __syncwarp();
            }
        }
    }
    __syncthreads();
    for (int e = tid; e < k * d; e += UP_THREADS) {
        const double v = tbl[e];
        if (v != 0.0) atomicAdd(acc + e, v);
    }
    for (int e = tid; e < k; e += UP_THREADS) {
        const int c = cnt[e];
        if (c) atomicAdd(acc + (size_t)k * d + e, (double)c);
    }
    if (mind2) km_block_atomic_add(inert, acc + (size_t)k * d + k, red);
}

// `cen` != nullptr (b200km_update_inertia_*): the inertia is summed HERE, in float64, as sum_i ||x_i - cen[label_i]||^2
// from the rows this kernel reads anyway, instead of from a per-row distance vector written by K1.
template <typename T>
__global__ void __launch_bounds__(UP_THREADS)
update_atomic_kernel(const T *__restrict__ X, int64_t n, int d, int64_t ldx, const int32_t *__restrict__ labels,
                     const T *__restrict__ mind2, int k, double *__restrict__ acc, int scale_sqrt,
                     const double *__restrict__ cen) {
    __shared__ double red[32];
    const int lane = threadIdx.x & 31;
    const int64_t warp_global = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    double inert = 0.0;
    double *counts = acc + (size_t)k * d;
    constexpr int RU = 4;  // rows in flight per warp
    for (int64_t r0 = warp_global * RU; r0 < n; r0 += nwarps * RU) {
        int lab[RU];
#pragma unroll
        for (int u = 0; u < RU; ++u) lab[u] = (r0 + u < n) ? __ldg(labels + r0 + u) : -1;
        for (int f0 = 0; f0 < d; f0 += 32) {
            const int f = f0 + lane;
            T v[RU];
#pragma unroll
            for (int u = 0; u < RU; ++u) {
                v[u] = (T)0;
                if (lab[u] >= 0 && f < d) v[u] = __ldg(X + (r0 + u) * ldx + f);
            }
#pragma unroll
            for (int u = 0; u < RU; ++u)
                if (lab[u] >= 0 && f < d) {
                    atomicAdd(acc + (size_t)lab[u] * d + f, (double)v[u]);
                    if (cen) {
                        const double df = (double)v[u] - __ldg(cen + (size_t)lab[u] * d + f);
                        inert = fma(df, df, inert);
                    }
                }
        }
        if (lane < RU && r0 + lane < n) {
            const int l = __ldg(labels + r0 + lane);
            atomicAdd(counts + l, 1.0);
            if (mind2) {
                const double v = (double)__ldg(mind2 + r0 + lane);
                inert += scale_sqrt ? sqrt(v) : v;
            }
        }
    }
    if (mind2 || cen) km_block_atomic_add(inert, acc + (size_t)k * d + k, red);
}


// ---------------------------------------------------------------------------------------------
// update_seg_kernel: "sort the chunk, then sum runs" -- for tables too big for shared memory.
// A CTA takes SEG_R consecutive rows, counting-sorts their indices by label in shared memory
// (histogram -> exclusive scan -> scatter), then each warp walks the rows of one cluster at a time
// (lanes = features, coalesced 128-byte row segments, float64 register accumulators) and issues ONE
// float64 RED per (cluster, feature) per chunk instead of one per (row, feature): SEG_R/k times fewer
// L2 atomics than update_atomic_kernel (16x at k = 1024).  Counts come from the histogram.
// ---------------------------------------------------------------------------------------------
constexpr int SEG_R_MAX = 65536;   // rows per CTA chunk: row offsets fit uint16
constexpr int SEG_THREADS = 512;
constexpr int SEG_LONG = 256;       // runs longer than this are shared out between the CTA's warps
constexpr int SEG_MAX_K = 8192;

// NARROW (d <= 32): a row is one 32-lane load, so the second feature half (fb, w[]) is compiled out and the
// registers it frees buy a third resident CTA (C2: 2.39 -> 2.64 TB/s); wider rows do better with two (C3).
// INERTIA: the run's centre row (float64, `cen`) is held by the lanes and every row adds (x - c)^2 to a float64
// lane accumulator -- two more DFMA-pipe instructions per row per lane, hidden under the row gathers.  It needs more
// registers than three resident CTAs leave (40: it spilled), so it runs two per SM -- which is also faster
// (5e7 x 32 rows, k = 256: 2.16 -> 1.75 ms, 3.8 TB/s; tools/upd_inertia_bench.py).
template <typename T, bool NARROW, bool INERTIA>
__global__ void __launch_bounds__(SEG_THREADS, (NARROW && !INERTIA) ? 3 : 2)
update_seg_kernel(const T *__restrict__ X, int64_t n, int d, int64_t ldx, const int32_t *__restrict__ labels,
                  const T *__restrict__ mind2, int k, double *__restrict__ acc, int scale_sqrt, int SEG_R,
                  const double *__restrict__ cen) {
    extern __shared__ __align__(16) unsigned char seg_smem[];
    int *start = reinterpret_cast<int *>(seg_smem);              // [k+1] histogram, then exclusive starts
    int *cursor = start + (k + 1);                               // [k]
    unsigned short *order = reinterpret_cast<unsigned short *>(cursor + k);  // [SEG_R]
    __shared__ double red[32];
    __shared__ int wsum[SEG_THREADS / 32];
    __shared__ int nlong;
    __shared__ int longlist[SEG_R_MAX / SEG_LONG];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = SEG_THREADS / 32;
    double inert = 0.0;
    double *counts = acc + (size_t)k * d;
    if (tid == 0) nlong = 0;   // (the first chunk's phase-1 barrier orders this before any use)

    for (int64_t base = (int64_t)blockIdx.x * SEG_R; base < n; base += (int64_t)gridDim.x * SEG_R) {
        const int rows = (int)((n - base) < SEG_R ? (n - base) : SEG_R);
        for (int j = tid; j <= k; j += SEG_THREADS) start[j] = 0;
        __syncthreads();
        // 1. histogram (+ inertia)
        for (int i = tid; i < rows; i += SEG_THREADS) {
            const int l = __ldg(labels + base + i);
            atomicAdd(&start[l], 1);
            if (mind2) {
                const double v = (double)__ldg(mind2 + base + i);
                inert += scale_sqrt ? sqrt(v) : v;
            }
        }
        __syncthreads();
        // 2. exclusive scan of start[0..k) (thread t owns a contiguous slice), counts to the global table
        {
            const int per = (k + SEG_THREADS - 1) / SEG_THREADS;
            const int j0 = tid * per, j1 = (j0 + per) < k ? (j0 + per) : k;
            int s = 0;
            for (int j = j0; j < j1; ++j) s += start[j];
            int incl = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            if (lane == 31) wsum[warp] = incl;
            __syncthreads();
            int woff = 0;
            for (int w = 0; w < warp; ++w) woff += wsum[w];
            int run = woff + incl - s;
            for (int j = j0; j < j1; ++j) {
                const int c = start[j];
                if (c) atomicAdd(counts + j, (double)c);
                start[j] = run;
                cursor[j] = run;
                run += c;
            }
            if (tid == SEG_THREADS - 1) start[k] = rows;
        }
        __syncthreads();
        // 3. scatter row offsets into label order
        for (int i = tid; i < rows; i += SEG_THREADS) {
            const int l = __ldg(labels + base + i);
            const int pos = atomicAdd(&cursor[l], 1);
            order[pos] = (unsigned short)i;
        }
        __syncthreads();
        // 4. one warp per cluster: sum its rows, one RED per feature.  A run longer than SEG_LONG rows (a dominant
        //    cluster) would leave one warp with most of the chunk, so it is set aside and then cut into one slice
        //    per warp (16 REDs per feature for it instead of 1).
        const T *Xb = X + base * ldx;
        auto sum_run = [&](int c, int s0, int s1) {
            for (int f0 = 0; f0 < d; f0 += 64) {
                const int fa = f0 + lane, fb = NARROW ? d : f0 + 32 + lane;   // NARROW: fb >= d, dead code
                double a0 = 0.0, a1 = 0.0;
                double ca = 0.0, cb = 0.0, qa = 0.0, qb = 0.0;
                if (INERTIA) {
                    if (fa < d) ca = __ldg(cen + (size_t)c * d + fa);
                    if (!NARROW && fb < d) cb = __ldg(cen + (size_t)c * d + fb);
                }
                int p = s0;
                for (; p + 8 <= s1; p += 8) {   // 8 rows (16 loads per lane) in flight
                    T v[8], w[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int64_t ro = (int64_t)order[p + u] * ldx;
                        v[u] = fa < d ? __ldg(Xb + ro + fa) : (T)0;
                        w[u] = (!NARROW && fb < d) ? __ldg(Xb + ro + fb) : (T)0;
                    }
                    a0 += (((double)v[0] + (double)v[1]) + ((double)v[2] + (double)v[3])) +
                          (((double)v[4] + (double)v[5]) + ((double)v[6] + (double)v[7]));
                    if (!NARROW)
                        a1 += (((double)w[0] + (double)w[1]) + ((double)w[2] + (double)w[3])) +
                              (((double)w[4] + (double)w[5]) + ((double)w[6] + (double)w[7]));
                    if (INERTIA) {
                        double q0 = 0.0, q1 = 0.0;   // two chains; a lane beyond d holds v = 0 and c = 0: adds 0
#pragma unroll
                        for (int u = 0; u < 8; u += 2) {
                            const double d0 = (double)v[u] - ca, d1 = (double)v[u + 1] - ca;
                            q0 = fma(d0, d0, q0);
                            q1 = fma(d1, d1, q1);
                            if (!NARROW) {
                                const double e0 = (double)w[u] - cb, e1 = (double)w[u + 1] - cb;
                                q0 = fma(e0, e0, q0);
                                q1 = fma(e1, e1, q1);
                            }
                        }
                        qa += q0;
                        qb += q1;
                    }
                }
                for (; p + 4 <= s1; p += 4) {
                    const int r0 = order[p], r1 = order[p + 1], r2 = order[p + 2], r3 = order[p + 3];
                    T v0 = (T)0, v1 = (T)0, v2 = (T)0, v3 = (T)0, w0 = (T)0, w1 = (T)0, w2 = (T)0, w3 = (T)0;
                    if (fa < d) {
                        v0 = __ldg(Xb + (int64_t)r0 * ldx + fa); v1 = __ldg(Xb + (int64_t)r1 * ldx + fa);
                        v2 = __ldg(Xb + (int64_t)r2 * ldx + fa); v3 = __ldg(Xb + (int64_t)r3 * ldx + fa);
                    }
                    if (!NARROW && fb < d) {
                        w0 = __ldg(Xb + (int64_t)r0 * ldx + fb); w1 = __ldg(Xb + (int64_t)r1 * ldx + fb);
                        w2 = __ldg(Xb + (int64_t)r2 * ldx + fb); w3 = __ldg(Xb + (int64_t)r3 * ldx + fb);
                    }
                    a0 += ((double)v0 + (double)v1) + ((double)v2 + (double)v3);
                    if (!NARROW) a1 += ((double)w0 + (double)w1) + ((double)w2 + (double)w3);
                    if (INERTIA && fa < d) {
                        const double d0 = (double)v0 - ca, d1 = (double)v1 - ca, d2 = (double)v2 - ca, d3 = (double)v3 - ca;
                        qa = fma(d0, d0, qa); qb = fma(d1, d1, qb); qa = fma(d2, d2, qa); qb = fma(d3, d3, qb);
                    }
                    if (INERTIA && !NARROW && fb < d) {
                        const double e0 = (double)w0 - cb, e1 = (double)w1 - cb, e2 = (double)w2 - cb, e3 = (double)w3 - cb;
                        qa = fma(e0, e0, qa); qb = fma(e1, e1, qb); qa = fma(e2, e2, qa); qb = fma(e3, e3, qb);
                    }
                }
                for (; p < s1; ++p) {
                    const int r0 = order[p];
                    if (fa < d) {
                        const double x0 = (double)__ldg(Xb + (int64_t)r0 * ldx + fa);
                        a0 += x0;
                        if (INERTIA) { const double d0 = x0 - ca; qa = fma(d0, d0, qa); }
                    }
                    if (!NARROW && fb < d) {
                        const double x1 = (double)__ldg(Xb + (int64_t)r0 * ldx + fb);
                        a1 += x1;
                        if (INERTIA) { const double e0 = x1 - cb; qb = fma(e0, e0, qb); }
                    }
                }
                if (INERTIA) inert += qa + qb;
                if (fa < d) atomicAdd(acc + (size_t)c * d + fa, a0);
                if (!NARROW && fb < d) atomicAdd(acc + (size_t)c * d + fb, a1);
            }
        };
        for (int c = warp; c < k; c += NW) {
            const int s0 = start[c], s1 = start[c + 1];
            if (s0 == s1) continue;
            if (s1 - s0 > SEG_LONG) {
                if (lane == 0) longlist[atomicAdd(&nlong, 1)] = c;
                continue;
            }
            sum_run(c, s0, s1);
        }
        __syncthreads();
        const int nl = nlong;
        for (int i = 0; i < nl; ++i) {
            const int c = longlist[i];
            const int s0 = start[c], s1 = start[c + 1];
            const int per = (s1 - s0 + NW - 1) / NW;
            const int a = s0 + warp * per, b = (a + per) < s1 ? (a + per) : s1;
            if (a < b) sum_run(c, a, b);
        }
        __syncthreads();
        if (tid == 0) nlong = 0;
    }
    if (mind2 || INERTIA) km_block_atomic_add(inert, acc + (size_t)k * d + k, red);
}

extern "C" size_t b200km_acc_len(int k, int d) {
    if (k <= 0 || d <= 0) return 0;
    return (size_t)k * d + (size_t)k + 1;
}

template <typename T>
static int update_impl(const T *X, int64_t n, int d, int64_t ldx, const int32_t *labels, const T *mind2, int k,
                       double *acc, int scale_sqrt, void *stream, const double *cen = nullptr) {
    KM_CHECK_ARG(X && labels && acc, B200KM_EINVAL);
    KM_CHECK_ARG(!(cen && mind2), B200KM_EINVAL);
    KM_CHECK_ARG(n > 0 && d > 0 && k > 0 && ldx >= d, B200KM_EINVAL);
    KM_CHECK_ARG(k <= B200KM_MAX_K && d <= B200KM_MAX_D, B200KM_ETOOBIG);
    cudaStream_t st = km_stream(stream);
    const size_t smem = sizeof(double) * (size_t)k * d + sizeof(int) * (size_t)k;
    const int variant = km_k2_variant();
    const bool seg_ok = k <= SEG_MAX_K && n >= 4 * (int64_t)k;
    const bool owner_ok = smem <= 96 * 1024;
    // auto (measured on B200, tools/upd_check.py): chunk sort + run sums while a chunk holds >= 8 rows per cluster
    // (k <= 2048: C2, C3, and every small table); plain fp64 REDs beyond that (C4: runs of 2 rows do not pay for
    // the sort)
    // (the shared-memory owner table, once the choice for tiny tables, loses 3-6x to the chunk-sort kernel on every
    //  shape measured -- and 14x with one dominant cluster -- so auto takes it only when the chunk sort cannot run)
    const bool use_owner = !cen && (variant == 1 ? owner_ok : (variant == 0 && owner_ok && !seg_ok && (size_t)k * d <= 2048));
    const bool use_seg = !use_owner && seg_ok && (variant == 2 || (variant == 0 && k <= 2048));
    if (use_owner) {
        auto kern = update_owner_kernel<T>;
        static bool attr_set[KM_MAX_DEVICES] = {};
        if (int rc = km_set_smem_once(kern, 96 * 1024, attr_set)) return rc;
        int64_t blocks = (n + 1023) / 1024;
        const int64_t cap = (int64_t)km_num_sms() * 2;
        if (blocks > cap) blocks = cap;
        kern<<<(unsigned)blocks, UP_THREADS, smem, st>>>(X, n, d, ldx, labels, mind2, k, acc, scale_sqrt);
    } else if (use_seg) {
        // rows per chunk: longer runs per cluster (fewer REDs, less loop overhead) against fewer resident CTAs and a
        // longer sort phase; measured (sweep in profiles/r1_hbm_kernels_summary.md): 32768 around d = 64, else 16384
        int SEG_R = (d > 32 && d < 96) ? 32768 : 16384;
        if (SEG_R > SEG_R_MAX) SEG_R = SEG_R_MAX;
        while (SEG_R > 4096 && sizeof(int) * (size_t)(2 * k + 1) + sizeof(unsigned short) * (size_t)SEG_R + 16 > 200 * 1024) SEG_R /= 2;
        const size_t seg_smem = sizeof(int) * (size_t)(2 * k + 1) + sizeof(unsigned short) * (size_t)SEG_R + 16;
        static bool attr_n[KM_MAX_DEVICES] = {}, attr_w[KM_MAX_DEVICES] = {}, attr_ni[KM_MAX_DEVICES] = {},
                    attr_wi[KM_MAX_DEVICES] = {};
        if (int rc = km_set_smem_once(update_seg_kernel<T, true, false>, 200 * 1024, attr_n)) return rc;
        if (int rc = km_set_smem_once(update_seg_kernel<T, false, false>, 200 * 1024, attr_w)) return rc;
        if (int rc = km_set_smem_once(update_seg_kernel<T, true, true>, 200 * 1024, attr_ni)) return rc;
        if (int rc = km_set_smem_once(update_seg_kernel<T, false, true>, 200 * 1024, attr_wi)) return rc;
        int64_t blocks = (n + SEG_R - 1) / SEG_R;
        const int64_t cap = (int64_t)km_num_sms() * 4;   // several CTAs per SM: one sorts while another gathers
        if (blocks > cap) blocks = cap;
        const unsigned gb = (unsigned)blocks;
        if (d <= 32 && cen)
            update_seg_kernel<T, true, true><<<gb, SEG_THREADS, seg_smem, st>>>(X, n, d, ldx, labels, mind2, k, acc,
                                                                              scale_sqrt, SEG_R, cen);
        else if (d <= 32)
            update_seg_kernel<T, true, false><<<gb, SEG_THREADS, seg_smem, st>>>(X, n, d, ldx, labels, mind2, k, acc,
                                                                               scale_sqrt, SEG_R, cen);
        else if (cen)
            update_seg_kernel<T, false, true><<<gb, SEG_THREADS, seg_smem, st>>>(X, n, d, ldx, labels, mind2, k, acc,
                                                                               scale_sqrt, SEG_R, cen);
        else
            update_seg_kernel<T, false, false><<<gb, SEG_THREADS, seg_smem, st>>>(X, n, d, ldx, labels, mind2, k, acc,
                                                                                scale_sqrt, SEG_R, cen);
    } else {
        int64_t blocks = (n + 4 * UP_WARPS - 1) / (4 * UP_WARPS);
        const int64_t cap = (int64_t)km_num_sms() * 8;
        if (blocks > cap) blocks = cap;
        update_atomic_kernel<T><<<(unsigned)blocks, UP_THREADS, 0, st>>>(X, n, d, ldx, labels, mind2, k, acc, scale_sqrt,
                                                                          cen);
    }
    KM_LAUNCH_RET();
}

extern "C" int b200km_update_f32(const float *X, int64_t n, int d, int64_t ldx, const int32_t *labels,
                                 const float *mind2, int k, double *acc, int scale_sqrt, void *stream) {
    return update_impl<float>(X, n, d, ldx, labels, mind2, k, acc, scale_sqrt, stream);
}
extern "C" int b200km_update_f64(const double *X, int64_t n, int d, int64_t ldx, const int32_t *labels,
                                 const double *mind2, int k, double *acc, int scale_sqrt, void *stream) {
    return update_impl<double>(X, n, d, ldx, labels, mind2, k, acc, scale_sqrt, stream);
}
extern "C" int b200km_update_inertia_f32(const float *X, int64_t n, int d, int64_t ldx, const int32_t *labels,
                                         const double *centers_f64, int k, double *acc, void *stream) {
    KM_CHECK_ARG(centers_f64, B200KM_EINVAL);
    return update_impl<float>(X, n, d, ldx, labels, nullptr, k, acc, 0, stream, centers_f64);
}
extern "C" int b200km_update_inertia_f64(const double *X, int64_t n, int d, int64_t ldx, const int32_t *labels,
                                         const double *centers_f64, int k, double *acc, void *stream) {
    KM_CHECK_ARG(centers_f64, B200KM_EINVAL);
    return update_impl<double>(X, n, d, ldx, labels, nullptr, k, acc, 0, stream, centers_f64);
}

// ---------------------------------------------------------------------------------------------
// candidate weighting (k-means|| steps 7-8 of Bahmani et al.: w_c = number of rows closest to candidate c).
// The labels come from K1 against the candidate set; this is the count half of K2 on its own -- a histogram of
// the labels (4 bytes per row, HBM-bound) without the second read of X that the sums would cost.
// ---------------------------------------------------------------------------------------------
constexpr int LC_THREADS = 512;
__global__ void __launch_bounds__(LC_THREADS)
label_counts_kernel(const int32_t *__restrict__ labels, int64_t n, int k, double *__restrict__ counts) {
    extern __shared__ int lc_hist[];   // [k]
    for (int j = threadIdx.x; j < k; j += LC_THREADS) lc_hist[j] = 0;
    __syncthreads();
    const int64_t n4 = n >> 2;
    const int4 *l4 = reinterpret_cast<const int4 *>(labels);   // cudaMalloc'ed / row-sliced label vectors: see the launcher
    for (int64_t i = (int64_t)blockIdx.x * LC_THREADS + threadIdx.x; i < n4; i += (int64_t)gridDim.x * LC_THREADS) {
        const int4 v = __ldg(l4 + i);
        if ((unsigned)v.x < (unsigned)k) atomicAdd(&lc_hist[v.x], 1);
        if ((unsigned)v.y < (unsigned)k) atomicAdd(&lc_hist[v.y], 1);
        if ((unsigned)v.z < (unsigned)k) atomicAdd(&lc_hist[v.z], 1);
        if ((unsigned)v.w < (unsigned)k) atomicAdd(&lc_hist[v.w], 1);
    }
    if (blockIdx.x == 0)
        for (int64_t i = (n4 << 2) + threadIdx.x; i < n; i += LC_THREADS) {
            const int l = __ldg(labels + i);
            if ((unsigned)l < (unsigned)k) atomicAdd(&lc_hist[l], 1);
        }
    __syncthreads();
    for (int j = threadIdx.x; j < k; j += LC_THREADS) {
        const int c = lc_hist[j];
        if (c) atomicAdd(counts + j, (double)c);     // integers below 2^53: exact in any order
    }
}
__global__ void __launch_bounds__(256)
label_counts_scalar_kernel(const int32_t *__restrict__ labels, int64_t n, int k, double *__restrict__ counts) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int l = __ldg(labels + i);
        if ((unsigned)l < (unsigned)k) atomicAdd(counts + l, 1.0);
    }
}

extern "C" int b200km_label_counts(const int32_t *labels, int64_t n, int k, double *counts, void *stream) {
    KM_CHECK_ARG(labels && counts && n > 0 && k > 0, B200KM_EINVAL);
    cudaStream_t st = km_stream(stream);
    const bool aligned = (reinterpret_cast<uintptr_t>(labels) & 15) == 0;
    if (aligned && (size_t)k * sizeof(int) <= 200 * 1024) {
        static bool attr_set[KM_MAX_DEVICES] = {};
        if (int rc = km_set_smem_once(label_counts_kernel, 200 * 1024, attr_set)) return rc;
        int64_t blocks = (n / 4 + LC_THREADS * 8 - 1) / (LC_THREADS * 8);
        if (blocks < 1) blocks = 1;
        const int64_t cap = (int64_t)km_num_sms() * 2;
        if (blocks > cap) blocks = cap;
        label_counts_kernel<<<(unsigned)blocks, LC_THREADS, (size_t)k * sizeof(int), st>>>(labels, n, k, counts);
    } else {
        int64_t blocks = (n + 255) / 256;
        const int64_t cap = (int64_t)km_num_sms() * 8;
        if (blocks > cap) blocks = cap;
        label_counts_scalar_kernel<<<(unsigned)blocks, 256, 0, st>>>(labels, n, k, counts);
    }
    KM_LAUNCH_RET();
}

// ---------------------------------------------------------------------------------------------
// K3 finalize
// ---------------------------------------------------------------------------------------------
// One thread-block cluster of FIN_CTAS CTAs does the whole job, so the shift is summed in ONE FIXED ORDER (thread
// partials over a fixed stride -> warp shuffles -> warps in index order -> CTAs in rank order through distributed
// shared memory): every rank of a multi-GPU fit computes bit-identical centres AND a bit-identical shift from the
// allreduced buffer, so all ranks take the same `shift < tol` / `shift > 1e-7` decisions (k_means.py:608, :612) and
// no rank is left alone in a collective.  (The earlier multi-block kernel summed the shift with floating-point
// atomics, whose order -- and so whose last bit -- could differ from rank to rank.)
constexpr int FIN_CTAS = 8;
constexpr int FIN_THREADS = 512;

__global__ void __cluster_dims__(FIN_CTAS, 1, 1) __launch_bounds__(FIN_THREADS)
finalize_kernel(const double *__restrict__ acc, int k, int d, const double *__restrict__ cold,
                double *__restrict__ cnew, double *__restrict__ shift_out) {
    __shared__ double red[FIN_THREADS / 32];
    __shared__ double cta_part[FIN_CTAS];   // used in CTA 0 only: one slot per CTA of the cluster
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned rank = cluster.block_rank();
    const double *counts = acc + (size_t)k * d;
    double s = 0.0;
    const int64_t total = (int64_t)k * d;
    for (int64_t e = (int64_t)rank * FIN_THREADS + threadIdx.x; e < total; e += (int64_t)FIN_CTAS * FIN_THREADS) {
        const int j = (int)(e / d);
        const double c = fmax(counts[j], 1.0);  // k_means.py:600
        const double v = acc[e] / c;            // :601
        cnew[e] = v;
        const double df = cold[e] - v;  // :605
        s = fma(df, df, s);
    }
    s = km_warp_sum(s);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) red[w] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < FIN_THREADS / 32; ++i) t += red[i];
        double *slot = cluster.map_shared_rank(cta_part, 0) + rank;
        *slot = t;
    }
    cluster.sync();
    if (rank == 0 && threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < FIN_CTAS; ++i) t += cta_part[i];
        *shift_out = t;
    }
}

extern "C" int b200km_finalize(const double *acc, int k, int d, const double *centers_old, double *centers_new,
                               double *shift_out, void *stream) {
    KM_CHECK_ARG(acc && centers_old && centers_new && shift_out && k > 0 && d > 0, B200KM_EINVAL);
    finalize_kernel<<<FIN_CTAS, FIN_THREADS, 0, km_stream(stream)>>>(acc, k, d, centers_old, centers_new, shift_out);
    KM_LAUNCH_RET();
}

// ---------------------------------------------------------------------------------------------
// sum / sum-of-sqrt of a vector into a float64 scalar (k_means.py:616 on the non-squared pass)
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
reduce_sum_kernel(const T *__restrict__ v, int64_t n, int take_sqrt, double *__restrict__ out) {
    __shared__ double red[32];
    double s = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double x = (double)__ldg(v + i);
        s += take_sqrt ? sqrt(x) : x;
    }
    km_block_atomic_add(s, out, red);
}
template <typename T>
static int reduce_sum_impl(const T *v, int64_t n, int take_sqrt, double *out, void *stream) {
    KM_CHECK_ARG(v && out && n > 0, B200KM_EINVAL);
    int64_t blocks = (n + 255) / 256;
    if (blocks > km_num_sms() * 8) blocks = km_num_sms() * 8;
    reduce_sum_kernel<T><<<(unsigned)blocks, 256, 0, km_stream(stream)>>>(v, n, take_sqrt, out);
    KM_LAUNCH_RET();
}
extern "C" int b200km_reduce_sum_f32(const float *v, int64_t n, int take_sqrt, double *out, void *stream) {
    return reduce_sum_impl<float>(v, n, take_sqrt, out, stream);
}
extern "C" int b200km_reduce_sum_f64(const double *v, int64_t n, int take_sqrt, double *out, void *stream) {
    return reduce_sum_impl<double>(v, n, take_sqrt, out, stream);
}

// ---------------------------------------------------------------------------------------------
// Fused K1+K2 for the HBM-bound regime (k <= 16, d <= 32; BASELINE config C1): ONE pass over X gives
// labels, min distances, per-cluster sums, counts and inertia.  Replaces the pair
// pairwise_distances_argmin_min (reference dask_ml/metrics/pairwise.py:19-47) + _centers_dense /
// bincount (dask_ml/cluster/k_means.py:582-598) that the reference runs as two passes over X.
//   phase A: thread per row, float64 direct-form distances to the k centres held in shared memory;
//   phase B: the tile's rows are counting-sorted by label with warp ballots, then thread (c, f) adds the
//            rows of cluster c into its private float64 register cell -- no atomics until the CTA's
//            final flush (k*d REDs per CTA).
// ---------------------------------------------------------------------------------------------
constexpr int FU_R = 256;        // rows per tile == threads per CTA
constexpr int FU_MAXK = 16;
constexpr int FU_MAXD = 32;

// Tail of the fused kernels when the whole iteration is ONE launch (b200km_lloyd_small_*; single GPU; done != nullptr):
// the CTA that finishes last divides the sums by the counts (K3, k_means.py:600-605), leaves [shift, inertia] in out2
// and hands acc and the ticket back zeroed for the next iteration -- no memset, no finalize launch between two
// iterations.  Called by every thread of the CTA after its REDs into acc; `red` = 32 doubles of shared memory.
__device__ __forceinline__ void fu_last_cta_finalize(double *__restrict__ acc, int ncell, int k, int d,
                                                     const double *__restrict__ c64, unsigned int *__restrict__ done,
                                                     double *__restrict__ cnew, double *__restrict__ out2, double tag,
                                                     double *red) {
    if (done == nullptr) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthr = blockDim.x;
    __shared__ unsigned int ticket;
    __threadfence();                                   // this CTA's REDs are visible before its ticket
    __syncthreads();
    if (tid == 0) ticket = atomicAdd(done, 1u);
    __syncthreads();
    if (ticket != gridDim.x - 1) return;
    __threadfence();
    double sh = 0.0;
    for (int e = tid; e < ncell; e += nthr) {
        const double cnt = fmax(__ldcg(acc + ncell + e / d), 1.0);     // :600 (L2: other CTAs wrote these)
        const double v = __ldcg(acc + e) / cnt;                         // :601
        cnew[e] = v;
        const double df = c64[e] - v;                                   // :605
        sh = fma(df, df, sh);
    }
    sh = km_warp_sum(sh);
    if (lane == 0) red[warp] = sh;
    __syncthreads();                                   // (also: every thread has read its acc entries)
    if (tid == 0) {
        double t = 0.0;
        for (int w = 0; w < nthr / 32; ++w) t += red[w];
        // out2 may be page-locked HOST memory the GPU writes directly: [shift, inertia] first, then -- after a
        // system-scope fence -- the caller's tag, so a host that polls out2[2] sees complete values without
        // synchronising the stream (the 8-byte read back is the only per-iteration round trip left)
        volatile double *o = out2;
        o[0] = t;
        o[1] = __ldcg(acc + ncell + k);
        __threadfence_system();
        o[2] = tag;
        *done = 0u;
    }
    __syncthreads();
    for (int e = tid; e < ncell + k + 1; e += nthr) acc[e] = 0.0;
}

template <typename TX, int DMAX>
// d <= 16: four resident CTAs per SM and no register prefetch of the next tile (measured at C1, 1e6 x 16 float64,
// k = 8: 2 CTAs + prefetch 0.092 ms, 3 + prefetch 0.108 (spills), 3 without 0.084, 4 without 0.080 -- the kernel is
// latency-bound, more warps hide more than the prefetch did)
#ifndef FU_OCC16
#define FU_OCC16 4
#endif
#ifndef FU_PREF
#define FU_PREF 0
#endif
__global__ void __launch_bounds__(FU_R, DMAX <= 16 ? FU_OCC16 : 3)
assign_update_small_kernel(const TX *__restrict__ X, int64_t n, int d, int64_t ldx, const double *__restrict__ c64,
                           int k, int32_t *__restrict__ labels, TX *__restrict__ mind2, double *__restrict__ acc,
                           unsigned int *__restrict__ done, double *__restrict__ cnew, double *__restrict__ out2,
                           double tag) {
    extern __shared__ __align__(16) double fu_smem[];
    const int xld = d + 1;
    double *xs = fu_smem;                            // [FU_R][d+1]
    const int dpad = (d + 1) & ~1;                   // centres padded to an even length (zeros)
    int lg_dp = 0;
    while ((1 << lg_dp) < d) ++lg_dp;                // dp = smallest power of two >= d (<= 32)
    const int dp = 1 << lg_dp;
    double *cs = fu_smem + (size_t)FU_R * xld + ((FU_R * xld) & 1);  // [k][dpad], 16-byte aligned
    __shared__ unsigned char order[FU_R];            // row slots sorted by label
    __shared__ int wcnt[FU_R / 32][FU_MAXK];         // per-warp label counts
    __shared__ int cstart[FU_MAXK + 1];
    __shared__ double red[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int e = tid; e < k * dpad; e += FU_R) {
        const int c = e / dpad, f = e - c * dpad;
        cs[e] = f < d ? c64[(size_t)c * d + f] : 0.0;
    }
    // register cells: thread t owns cells t and t + 256 of the k*d table
    const int ncell = k * d;
    const int c0 = tid / d, f0 = tid % d;
    const int c1 = (tid + FU_R) / d, f1 = (tid + FU_R) % d;
    double cell0 = 0.0, cell1 = 0.0, inert = 0.0;
    double mycount = 0.0;                             // thread c < k accumulates the count of cluster c
    __syncthreads();

    // Tile staging.  Lanes = features (padded to a power of two so the row / feature split is shifts, not
    // divisions), 32/dp rows per warp instruction, dp loads per thread and tile.  For d <= 16 the NEXT tile's loads
    // are issued into registers before the current tile is worked on (they land under phases A-C), so a tile no
    // longer costs a full memory latency on top of its compute; wider rows keep the batched in-place staging.
    constexpr bool PREFETCH = DMAX <= 16 && FU_PREF;
    const int sub = lane >> lg_dp, fl = lane & (dp - 1);
    const int rpi = 32 >> lg_dp;                       // rows per warp instruction
    const int r_first = warp * rpi + sub;
    const int r_step = (FU_R / 32) * rpi;              // FU_R / r_step == dp loads per thread
    const bool fok = fl < d;
    TX pre[PREFETCH ? DMAX : 1];
    auto prefetch = [&](int64_t base) {
        const int64_t left = n - base;
#pragma unroll
        for (int i = 0; i < (PREFETCH ? DMAX : 1); ++i) {
            const int r = r_first + i * r_step;
            pre[i] = (TX)0;
            if (i < dp && fok && r < left) pre[i] = __ldg(X + (base + r) * ldx + fl);
        }
    };
    if (PREFETCH && (int64_t)blockIdx.x * FU_R < n) prefetch((int64_t)blockIdx.x * FU_R);

    for (int64_t base = (int64_t)blockIdx.x * FU_R; base < n; base += (int64_t)gridDim.x * FU_R) {
        const int rows = (int)((n - base) < FU_R ? (n - base) : FU_R);
        if (PREFETCH) {
#pragma unroll
            for (int i = 0; i < (PREFETCH ? DMAX : 1); ++i) {
                const int r = r_first + i * r_step;
                if (i < dp && fok && r < rows) xs[r * xld + fl] = (double)pre[i];
            }
            const int64_t nbase = base + (int64_t)gridDim.x * FU_R;
            if (nbase < n) prefetch(nbase);
        } else {
            for (int i0 = 0; i0 * r_step < FU_R; i0 += 8) {    // batches of 8 loads in flight per thread
                TX v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = r_first + (i0 + i) * r_step;
                    v[i] = (TX)0;
                    if (fok && r < rows) v[i] = __ldg(X + (base + r) * ldx + fl);
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = r_first + (i0 + i) * r_step;
                    if (fok && r < rows) xs[r * xld + fl] = (double)v[i];
                }
            }
        }
        __syncthreads();
        // ---- phase A: argmin (lowest index on ties) -------------------------------------------
        int lab = -1;
        if (tid < rows) {
            // the row lives in registers (d <= 32); centres are broadcast 16-byte shared-memory loads
            double xr[DMAX];
#pragma unroll
            for (int f = 0; f < DMAX; ++f) xr[f] = f < d ? xs[tid * xld + f] : 0.0;
            double best = INFINITY;
            for (int c = 0; c < k; ++c) {
                const double *cr = cs + c * dpad;
                double s0 = 0.0, s1 = 0.0;
#pragma unroll
                for (int f = 0; f < DMAX; f += 2) {
                    if (f < d) {
                        const double2 cv = *reinterpret_cast<const double2 *>(cr + f);
                        const double d0 = xr[f] - cv.x, d1 = xr[f + 1] - cv.y;
                        s0 = fma(d0, d0, s0);
                        s1 = fma(d1, d1, s1);
                    }
                }
                const double s = s0 + s1;
                if (s < best) { best = s; lab = c; }
            }
            labels[base + tid] = lab;
            if (mind2) mind2[base + tid] = (TX)best;
            inert += best;
        }
        // ---- counting sort of the tile's rows by label (warp ballots) -------------------------
        int rank_in_warp = 0;
        for (int c = 0; c < k; ++c) {
            const unsigned m = __ballot_sync(0xffffffffu, lab == c);
            if (lab == c) rank_in_warp = __popc(m & ((1u << lane) - 1u));
            if (lane == 0) wcnt[warp][c] = __popc(m);
        }
        __syncthreads();
        if (tid <= k) {
            int s = 0;
            for (int c = 0; c < tid; ++c)
                for (int w = 0; w < FU_R / 32; ++w) s += wcnt[w][c];
            cstart[tid] = s;
            if (tid < k) {
                int mine = 0;
                for (int w = 0; w < FU_R / 32; ++w) mine += wcnt[w][tid];
                mycount += (double)mine;
            }
        }
        __syncthreads();
        if (lab >= 0) {
            int pos = cstart[lab] + rank_in_warp;
            for (int w = 0; w < warp; ++w) pos += wcnt[w][lab];
            order[pos] = (unsigned char)tid;
        }
        __syncthreads();
        // ---- phase B: thread (c, f) sums the rows of cluster c ---------------------------------
        if (tid < ncell) {
            const int s0 = cstart[c0], s1 = cstart[c0 + 1];
            for (int p = s0; p < s1; ++p) cell0 += xs[order[p] * xld + f0];
        }
        if (tid + FU_R < ncell) {
            const int s0 = cstart[c1], s1 = cstart[c1 + 1];
            for (int p = s0; p < s1; ++p) cell1 += xs[order[p] * xld + f1];
        }
        __syncthreads();
    }
    if (tid < ncell && cell0 != 0.0) atomicAdd(acc + tid, cell0);
    if (tid + FU_R < ncell && cell1 != 0.0) atomicAdd(acc + tid + FU_R, cell1);
    if (tid < k && mycount != 0.0) atomicAdd(acc + ncell + tid, mycount);
    km_block_atomic_add(inert, acc + ncell + k, red);
    fu_last_cta_finalize(acc, ncell, k, d, c64, done, cnew, out2, tag, red);
}

// ---------------------------------------------------------------------------------------------
// The same pass for rows of exactly 64 or 128 bytes (d = 16 / 32 float32, d = 8 / 16 float64 -- C1), 16-byte aligned.
//   staging : cp.async (LDGSTS, 16 bytes per thread and instruction) into one of two shared-memory buffers: the next
//             tile lands while this one is worked on and no register holds it on the way.  The 16-byte chunks of a
//             row are XOR-swizzled by the row number, so the thread-per-row reads of phase A (128-bit loads, 8 lanes
//             per shared-memory phase, rows 64 / 128 bytes apart) hit eight different bank groups.
//   phase A : scores = (||c'||^2 + 2 mu . c') - 2 x . c' on the FP64 tensor cores (DMMA m8n8k4: 8 rows x 8 clusters x 4
//             features; c' = c - mu, mu the mean of the centres -- |x - c|^2 minus a per-row constant; the -2 c' fragments
//             and the cluster constants live in registers for the whole kernel); the four lanes that share a row
//             reduce (score, index) by two butterfly steps -- lowest index on ties, as above -- and then rebuild the
//             winner's distance in the DIRECT form sum (x - c)^2 from the raw values they already hold, so the
//             distances keep the accuracy of the kernel above.
//   phase B : sums = onehot(labels)^T . X on the FP64 tensor cores (mma.sync m8n8k4, DMMA): every warp multiplies the
//             one-hot matrix of ITS 32 rows' labels (clusters x rows, built from the labels by one shuffle + select per
//             four rows) with those rows (read back from shared memory in fragment order) and keeps the
//             clusters x features accumulator fragments in registers for the CTA's whole life -- 16 DMMAs per warp
//             and tile at C1 instead of a sort and 32-row dependent chains; warp-local, so one block barrier per tile
//             instead of five.  Counts: the one-hot predicates, counted in integer registers.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void fu_cp_async16(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(km_smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void fu_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void fu_cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// D (8x8, float64) += A (8x4, row major) . B (4x8, column major); lane l holds A[l >> 2][l & 3], B[l & 3][l >> 2] and
// D[l >> 2][2 * (l & 3) + {0, 1}]
__device__ __forceinline__ void fu_dmma884(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

template <typename TX, int D, int KMAX>
__global__ void __launch_bounds__(FU_R, (KMAX <= 8 && D * sizeof(TX) <= 128 && D <= 16) ? 3 : 2)
assign_update_bulk_kernel(const TX *__restrict__ X, int64_t n, int64_t ldx, const double *__restrict__ c64, int k,
                          int32_t *__restrict__ labels, TX *__restrict__ mind2, double *__restrict__ acc,
                          unsigned int *__restrict__ done, double *__restrict__ cnew, double *__restrict__ out2,
                          double tag) {
    constexpr int RB = D * (int)sizeof(TX);     // row bytes: 64 or 128
    constexpr int CPR = RB / 16;                // 16-byte chunks per row: 4 or 8
    constexpr int EPC = 16 / (int)sizeof(TX);   // elements per chunk
    constexpr int ES = (int)sizeof(TX);
    static_assert(CPR == 4 || CPR == 8, "rows of 64 or 128 bytes");
    extern __shared__ __align__(128) unsigned char fb_smem[];
    unsigned char *xs = fb_smem;                                              // [2][FU_R][RB], chunks swizzled
    // the centres feature-major, [D][CLD] with CLD = KMAX (+ 2 for 16 clusters): in the distance rebuild the lanes of a
    // warp read (feature 4 ks + l4, cluster = their row's label) -- four neighbouring table rows, any labels: with the
    // clusters of a feature side by side that is 256 contiguous bytes for k <= 8, two shared-memory wavefronts whatever
    // the labels are (centre-major rows of 128 bytes put the same feature of every centre on the same banks)
    constexpr int CLD = KMAX <= 8 ? KMAX : KMAX + 2;
    double *cs = reinterpret_cast<double *>(fb_smem + 2 * FU_R * RB);         // [D][CLD]
    __shared__ double red[32];
    __shared__ double mu_s[32], cn_s[FU_MAXK];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int l4 = lane & 3, l8 = lane >> 2;
    for (int e = tid; e < KMAX * D; e += FU_R) cs[(e % D) * CLD + e / D] = e < k * D ? c64[e] : 0.0;
    // swizzle of row r: chunk c is stored at position c ^ sw(r); sw(r) = r & 7 (128-byte rows), (r >> 1) & 3 (64-byte).
    // Every row set used below is a fixed row plus a multiple of 8, so each thread's swizzles are loop constants.
    auto sw = [](int r) { return CPR == 8 ? (r & 7) : ((r >> 1) & 3); };
    auto xoff = [&](int r, int f) { return r * RB + (((f / EPC) ^ sw(r)) << 4) + (f % EPC) * ES; };

    // ---- constants of the score MMAs: centring vector (mean of the centres), c' = c - mu, and per cluster
    //      ||c'||^2 + 2 mu . c'  (so that the A operand is the RAW row: |x - c|^2 = [row constant] + this - 2 x . c')
    __syncthreads();
    if (tid < D) {
        double m = 0.0;
        for (int c = 0; c < k; ++c) m += cs[tid * CLD + c];
        mu_s[tid] = m / (double)k;
    }
    __syncthreads();
    if (tid < KMAX) {
        double q = INFINITY;                                // a padded cluster never wins
        if (tid < k) {
            q = 0.0;
            for (int f = 0; f < D; ++f) { const double t = cs[f * CLD + tid] - mu_s[f]; q = fma(t, t + 2.0 * mu_s[f], q); }
        }
        cn_s[tid] = q;
    }
    __syncthreads();
    constexpr int MB = KMAX / 8, NB = D / 8;                // 8-cluster and 8-feature blocks
    constexpr int KS = D / 4;                               // K-steps of the score MMAs (4 features each)
    double bq[MB][KS], cnq[MB][2];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks)
#pragma unroll
        for (int mb = 0; mb < MB; ++mb)                      // B[feature 4 ks + l4][cluster 8 mb + l8] = -2 c'
            bq[mb][ks] = 8 * mb + l8 < k ? -2.0 * (cs[(4 * ks + l4) * CLD + 8 * mb + l8] - mu_s[4 * ks + l4]) : 0.0;
#pragma unroll
    for (int mb = 0; mb < MB; ++mb) {
        cnq[mb][0] = cn_s[8 * mb + 2 * l4];
        cnq[mb][1] = cn_s[8 * mb + 2 * l4 + 1];
    }
    double cell[MB][NB][2];                                 // sums: clusters 8 mb + l8, features 8 nb + 2 l4 + {0, 1}
    int icnt[MB];                                           // rows of cluster 8 mb + l8 among this lane's B-fragment rows
#pragma unroll
    for (int mb = 0; mb < MB; ++mb) {
        icnt[mb] = 0;
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) cell[mb][nb][0] = cell[mb][nb][1] = 0.0;
    }
    double inert = 0.0;

    // ---- per-thread byte offsets (shared-memory window of the warp's 32 rows) ---------------------------------
    // phase A: lane (l8, l4) holds row 8 mbr + l8, features 4 ks + l4
    uint32_t offA[KS];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) offA[ks] = (uint32_t)xoff(l8, 4 * ks + l4);
    // phase B: the four rows of a K-step are picked so that, under the swizzle, the fragment load spreads over all
    // the banks: rows 8 (ks >> 1) + {0, 2, 4, 6} + (ks & 1) for 128-byte rows, {0, 1, 4, 5} + 2 (ks & 1) for 64-byte
    // rows (consecutive rows would fold onto half of them); lane (l8, l4) holds row 8 (ks >> 1) + r7[ks & 1], feature 8 nb + l8
    int r7[2];
    uint32_t offB[2][NB];
#pragma unroll
    for (int pp = 0; pp < 2; ++pp) {
        r7[pp] = CPR == 8 ? 2 * l4 + pp : (l4 & 1) + 4 * (l4 >> 1) + 2 * pp;
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) offB[pp][nb] = (uint32_t)xoff(r7[pp], 8 * nb + l8);
    }
    // staging: every warp loads ITS 32 rows of the tile (32 * RB contiguous bytes when ldx == d) and nobody else
    // touches them, so the warps run independent two-buffer pipelines: cp.async.wait_group + __syncwarp, no block
    // barrier in the loop.  Lane l moves chunk l % CPR of rows l / CPR + (32 / CPR) * i.
    constexpr int SROWS = 32 / CPR;                         // rows between two chunks of one lane
    const int st_c = lane % CPR, st_r0 = lane / CPR;
    const uint32_t sdst = km_smem_u32(xs) + (uint32_t)(warp * 32 * RB) + (uint32_t)xoff(st_r0, st_c * EPC);
    const int64_t gstep = (int64_t)SROWS * ldx;
    // (row st_r0 + SROWS * i: SROWS = 4 resp. 8 keeps the swizzle of st_r0 only for 128-byte rows -- recompute per i)
    uint32_t sofs[CPR];
#pragma unroll
    for (int i = 0; i < CPR; ++i) sofs[i] = (uint32_t)xoff(st_r0 + SROWS * i, st_c * EPC) - (uint32_t)xoff(st_r0, st_c * EPC);
    auto issue = [&](int64_t base, int b) {
        const int64_t left = n - base - warp * 32;          // rows of this warp's slice that exist
        const TX *src = X + (base + warp * 32 + st_r0) * ldx + st_c * EPC;
        const uint32_t dst = sdst + (uint32_t)b * (FU_R * RB);
        if (left >= 32) {
#pragma unroll
            for (int i = 0; i < CPR; ++i)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + sofs[i]), "l"(src + i * gstep) : "memory");
        } else {
#pragma unroll
            for (int i = 0; i < CPR; ++i)
                if (st_r0 + SROWS * i < left)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + sofs[i]), "l"(src + i * gstep) : "memory");
        }
        fu_cp_async_commit();
    };

    const int64_t stride = (int64_t)gridDim.x * FU_R;
    int64_t base = (int64_t)blockIdx.x * FU_R;
    if (base < n) issue(base, 0);
    for (int it = 0; base < n; base += stride, ++it) {
        const int b = it & 1;
        const int rows = (int)((n - base) < FU_R ? (n - base) : FU_R);
        fu_cp_async_wait_all();
        __syncwarp();             // the slice has landed for every lane; the warp is done with the other buffer
        if (base + stride < n) issue(base + stride, b ^ 1);
        const unsigned char *xw = xs + (size_t)b * FU_R * RB + (size_t)warp * 32 * RB;   // this warp's 32 rows
        // ---- phase A: four blocks of 8 rows ------------------------------------------------------------------
        int lab4[4];                                        // label of row 8 mbr + l8 (the same in the row's 4 lanes)
        double ddp[4];
        bool unlabelled = false;
#pragma unroll
        for (int mbr = 0; mbr < 4; ++mbr) {
            const int rt = warp * 32 + 8 * mbr + l8;        // row within the tile
            double xa[KS], sc[MB][2];
#pragma unroll
            for (int ks = 0; ks < KS; ++ks) xa[ks] = (double)*reinterpret_cast<const TX *>(xw + mbr * 8 * RB + offA[ks]);
#pragma unroll
            for (int mb = 0; mb < MB; ++mb) { sc[mb][0] = cnq[mb][0]; sc[mb][1] = cnq[mb][1]; }
#pragma unroll
            for (int ks = 0; ks < KS; ++ks)
#pragma unroll
                for (int mb = 0; mb < MB; ++mb) fu_dmma884(sc[mb][0], sc[mb][1], xa[ks], bq[mb][ks]);
            double best = INFINITY;
            int bi = -1;
#pragma unroll
            for (int mb = 0; mb < MB; ++mb)
#pragma unroll
                for (int e = 0; e < 2; ++e)
                    if (sc[mb][e] < best) { best = sc[mb][e]; bi = 8 * mb + 2 * l4 + e; }
#pragma unroll
            for (int o = 1; o <= 2; o <<= 1) {
                const double ob = __shfl_xor_sync(0xffffffffu, best, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
            }
            if (rt >= rows) bi = -1;                        // beyond the end of X
            // the winner's distance, direct form, from the raw values
            const double *cr = cs + l4 * CLD + (bi < 0 ? 0 : bi);
            double dd = 0.0;
#pragma unroll
            for (int ks = 0; ks < KS; ++ks) {
                const double t = xa[ks] - cr[4 * ks * CLD];
                dd = fma(t, t, dd);
            }
            ddp[mbr] = dd;                                  // this lane's features only
            lab4[mbr] = bi;
            unlabelled |= bi < 0;
        }
        {
            // the four partial distances of the four row blocks, summed over the row's four lanes by a transposing
            // reduction (3 exchanges instead of 8): lane (l8, l4) ends up with the distance of row 8 l4 + l8 and writes
            // that row's outputs
            const bool b0 = l4 & 1, b1 = l4 & 2;
            double k0 = b0 ? ddp[1] : ddp[0], k1 = b0 ? ddp[3] : ddp[2];
            k0 += __shfl_xor_sync(0xffffffffu, b0 ? ddp[0] : ddp[1], 1);
            k1 += __shfl_xor_sync(0xffffffffu, b0 ? ddp[2] : ddp[3], 1);
            double dd = b1 ? k1 : k0;
            dd += __shfl_xor_sync(0xffffffffu, b1 ? k0 : k1, 2);
            const int rt = warp * 32 + 8 * l4 + l8;
            if (rt < rows) {
                const int bi = b1 ? (b0 ? lab4[3] : lab4[2]) : (b0 ? lab4[1] : lab4[0]);
                labels[base + rt] = bi;
                if (mind2) mind2[base + rt] = (TX)dd;
                inert += dd;
            }
        }
        // ---- phase B: sums += onehot(labels)^T . X for this warp's 32 rows -------------------------------------
        // a row without a label (beyond the tile's end: stale shared memory; or all-NaN distances) must add nothing,
        // and 0 * NaN would not: a warp that has one zeroes those rows' values on the way in (warp-uniform branch)
        auto phase_b = [&](auto SAN) {
            constexpr bool sanitize = decltype(SAN)::value;
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                const int lr = __shfl_sync(0xffffffffu, lab4[ks >> 1], 4 * r7[ks & 1]);   // label of this lane's row
                double a[MB];
#pragma unroll
                for (int mb = 0; mb < MB; ++mb) {
                    const bool hit = lr == 8 * mb + l8;
                    a[mb] = hit ? 1.0 : 0.0;
                    icnt[mb] += hit ? 1 : 0;
                }
#pragma unroll
                for (int nb = 0; nb < NB; ++nb) {
                    double x = (double)*reinterpret_cast<const TX *>(xw + (ks >> 1) * 8 * RB + offB[ks & 1][nb]);
                    if (sanitize && lr < 0) x = 0.0;
#pragma unroll
                    for (int mb = 0; mb < MB; ++mb) fu_dmma884(cell[mb][nb][0], cell[mb][nb][1], a[mb], x);
                }
            }
        };
        if (__any_sync(0xffffffffu, unlabelled)) phase_b(std::true_type{}); else phase_b(std::false_type{});
    }
    // ---- flush: the warps' accumulator fragments through shared memory, then one RED per cell and CTA -------
    __syncthreads();                                        // every thread is done with xs
    double *wtab = reinterpret_cast<double *>(xs);          // [FU_R / 32][KMAX][D]
#pragma unroll
    for (int mb = 0; mb < MB; ++mb)
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) {
            double *w = wtab + ((size_t)warp * KMAX + 8 * mb + l8) * D + 8 * nb + 2 * l4;
            w[0] = cell[mb][nb][0];
            w[1] = cell[mb][nb][1];
        }
    __shared__ int wcnt[FU_R / 32][FU_MAXK];
#pragma unroll
    for (int mb = 0; mb < MB; ++mb) {
        int c = icnt[mb];                                   // the four lanes of a cluster counted disjoint rows
        c += __shfl_xor_sync(0xffffffffu, c, 1);
        c += __shfl_xor_sync(0xffffffffu, c, 2);
        if (l4 == 0) wcnt[warp][8 * mb + l8] = c;
    }
    __syncthreads();
    const int ncell = k * D;
    for (int e = tid; e < ncell; e += FU_R) {
        const int c = e / D, f = e % D;
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < FU_R / 32; ++w) t += wtab[(w * KMAX + c) * D + f];
        if (t != 0.0) atomicAdd(acc + e, t);
    }
    if (tid < k) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < FU_R / 32; ++w) t += wcnt[w][tid];
        if (t != 0) atomicAdd(acc + ncell + tid, (double)t);
    }
    km_block_atomic_add(inert, acc + ncell + k, red);
    fu_last_cta_finalize(acc, ncell, k, D, c64, done, cnew, out2, tag, red);
}

int km_fused_variant();  // assign.cu (b200km_set_option)

template <typename TX, int D, int KMAX>
static int launch_bulk(const TX *X, int64_t n, int64_t ldx, const double *c64, int k, int32_t *labels, TX *mind2,
                       double *acc, cudaStream_t st, unsigned int *done, double *cnew, double *out2, double tag) {
    auto kern = assign_update_bulk_kernel<TX, D, KMAX>;
    constexpr int RB = D * (int)sizeof(TX);
    constexpr int OCC = (KMAX <= 8 && RB <= 128 && D <= 16) ? 3 : 2;
    const size_t smem = (size_t)2 * FU_R * RB + sizeof(double) * D * (KMAX + 2);
    static bool attr_set[KM_MAX_DEVICES] = {};
    if (int rc = km_set_smem_once(kern, 80 * 1024, attr_set)) return rc;
    int64_t blocks = (n + FU_R - 1) / FU_R;
    const int64_t cap = (int64_t)km_num_sms() * OCC;
    if (blocks > cap) blocks = cap;
    kern<<<(unsigned)blocks, FU_R, smem, st>>>(X, n, ldx, c64, k, labels, mind2, acc, done, cnew, out2, tag);
    KM_LAUNCH_RET();
}

template <typename TX>
static int assign_update_impl(const TX *X, int64_t n, int d, int64_t ldx, const double *centers_f64, int k,
                              int32_t *labels, TX *mind2, double *acc, void *stream, unsigned int *done = nullptr,
                              double *cnew = nullptr, double *out2 = nullptr, double tag = 0.0) {
    KM_CHECK_ARG(X && centers_f64 && labels && acc, B200KM_EINVAL);
    KM_CHECK_ARG(n > 0 && d > 0 && k > 0 && ldx >= d, B200KM_EINVAL);
    KM_CHECK_ARG(k <= FU_MAXK && d <= FU_MAXD, B200KM_ETOOBIG);
    // rows of 64 / 128 bytes, 16-byte aligned: the cp.async kernel (B200KM_OPT_FUSED_VARIANT 1 keeps the sorted-tile one)
    const int rb = d * (int)sizeof(TX);
    if (km_fused_variant() != 1 && (rb == 64 || rb == 128) && (reinterpret_cast<uintptr_t>(X) & 15) == 0 &&
        (ldx * sizeof(TX)) % 16 == 0) {
        cudaStream_t st = km_stream(stream);
        constexpr int DS = sizeof(TX) == 8 ? 8 : 16, DL = 2 * DS;   // the two row widths of this element type
        if (d == DS)
            return k <= 8 ? launch_bulk<TX, DS, 8>(X, n, ldx, centers_f64, k, labels, mind2, acc, st, done, cnew, out2, tag)
                          : launch_bulk<TX, DS, 16>(X, n, ldx, centers_f64, k, labels, mind2, acc, st, done, cnew, out2, tag);
        return k <= 8 ? launch_bulk<TX, DL, 8>(X, n, ldx, centers_f64, k, labels, mind2, acc, st, done, cnew, out2, tag)
                      : launch_bulk<TX, DL, 16>(X, n, ldx, centers_f64, k, labels, mind2, acc, st, done, cnew, out2, tag);
    }
    int64_t blocks = (n + FU_R - 1) / FU_R;
    const int64_t cap = (int64_t)km_num_sms() * (d <= 16 ? FU_OCC16 : 4);   // resident CTAs per SM of the two variants
    if (blocks > cap) blocks = cap;
    const size_t smem = sizeof(double) * ((size_t)FU_R * (d + 1) + 1 + (size_t)k * (d + 1));
    if (d <= 16) {
        auto kern = assign_update_small_kernel<TX, 16>;
        static bool attr_set[KM_MAX_DEVICES] = {};
        if (int rc = km_set_smem_once(kern, 80 * 1024, attr_set)) return rc;
        kern<<<(unsigned)blocks, FU_R, smem, km_stream(stream)>>>(X, n, d, ldx, centers_f64, k, labels, mind2, acc, done, cnew,
                                                                  out2, tag);
    } else {
        auto kern = assign_update_small_kernel<TX, FU_MAXD>;
        static bool attr_set[KM_MAX_DEVICES] = {};
        if (int rc = km_set_smem_once(kern, 80 * 1024, attr_set)) return rc;
        kern<<<(unsigned)blocks, FU_R, smem, km_stream(stream)>>>(X, n, d, ldx, centers_f64, k, labels, mind2, acc, done, cnew,
                                                                  out2, tag);
    }
    KM_LAUNCH_RET();
}
extern "C" int b200km_lloyd_small_f32(const float *X, int64_t n, int d, int64_t ldx, const double *centers_f64, int k,
                                      int32_t *labels, float *mind2, double *acc, unsigned int *ticket,
                                      double *centers_new, double *shift_inertia, double tag, void *stream) {
    KM_CHECK_ARG(ticket && centers_new && shift_inertia, B200KM_EINVAL);
    return assign_update_impl<float>(X, n, d, ldx, centers_f64, k, labels, mind2, acc, stream, ticket, centers_new,
                                   shift_inertia, tag);
}
extern "C" int b200km_lloyd_small_f64(const double *X, int64_t n, int d, int64_t ldx, const double *centers_f64, int k,
                                      int32_t *labels, double *mind2, double *acc, unsigned int *ticket,
                                      double *centers_new, double *shift_inertia, double tag, void *stream) {
    KM_CHECK_ARG(ticket && centers_new && shift_inertia, B200KM_EINVAL);
    return assign_update_impl<double>(X, n, d, ldx, centers_f64, k, labels, mind2, acc, stream, ticket, centers_new,
                                   shift_inertia, tag);
}
extern "C" int b200km_assign_update_f32(const float *X, int64_t n, int d, int64_t ldx, const double *centers_f64,
                                        int k, int32_t *labels, float *mind2, double *acc, void *stream) {
    return assign_update_impl<float>(X, n, d, ldx, centers_f64, k, labels, mind2, acc, stream);
}
extern "C" int b200km_assign_update_f64(const double *X, int64_t n, int d, int64_t ldx, const double *centers_f64,
                                        int k, int32_t *labels, double *mind2, double *acc, void *stream) {
    return assign_update_impl<double>(X, n, d, ldx, centers_f64, k, labels, mind2, acc, stream);
}
