// assign.cu -- K1: fused distance + argmin (+ min distance), centre pack, fp64 refine.
//
// Replaces dask_ml.metrics.pairwise_distances_argmin_min (reference
// dask_ml/metrics/pairwise.py:19-47 -> sklearn ArgKmin) on the Lloyd path
// (dask_ml/cluster/k_means.py:576-579) and in predict (:247-250).
//
// Two kernels:
//   assign_tiled_f32  -- FP32 FFMA, 128 rows x 128 centroids per CTA step, 8x8 register tile per
//                        thread.  Score = ||c||^2 - 2 x.c (the row-constant ||x||^2 is left out of
//                        the argmin).  Centroid tiles are staged by 1-D bulk async copies (TMA,
//                        mbarrier complete_tx) double-buffered; X rows are read once, 128-bit
//                        coalesced, and kept transposed in shared memory for the whole CTA.
//                        Tracks best AND second-best per row; rows whose gap is below the fp32
//                        rounding bound are flagged for the fp64 re-decision (refine).
//   assign_rb         -- "row block" kernel, fp64 arithmetic, direct form sum (x-c)^2.  Used for
//                        float64 X, for tiny k (HBM-bound regime), as the generic fallback and
//                        (with row indirection) as the refine kernel.
#include "common.cuh"
#include <atomic>

int km_launch_pack_tc(int k, int d, void *pack, cudaStream_t st);  // assign_tc.cu
int km_launch_assign_tc(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k, int32_t *labels,
                        float *mind2, int32_t *flag_rows, int32_t *flag_count, cudaStream_t st, bool precentred,
                        bool sparse_d2);

// =============================================================================================
// centre pack
// =============================================================================================
// mu[f] = mean_j c64[j][f], rounded to fp32 (block f; features d .. kpad(d)-1 are zero padding)
__global__ void __launch_bounds__(128) pack_mu_kernel(const double *__restrict__ c64in, int k, int d,
                                                      unsigned char *__restrict__ pack) {
    const KmPackLayout L = km_pack_layout(k, d);
    float *mu = reinterpret_cast<float *>(pack + L.off_mu);
    const int f = blockIdx.x;
    if (threadIdx.x == 0) reinterpret_cast<float *>(pack + L.off_zero)[f] = 0.f;
    __shared__ double red[4];
    if (f >= d) {
        if (threadIdx.x == 0) mu[f] = 0.f;
        return;
    }
    double s = 0.0;
    for (int j = threadIdx.x; j < k; j += blockDim.x) s += c64in[(size_t)j * d + f];
    s = km_warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        const double m = (red[0] + red[1] + red[2] + red[3]) / (double)k;
        const float mf = (float)m;
        mu[f] = isfinite(mf) ? mf : 0.f;
    }
}

__global__ void __launch_bounds__(128) pack_centers_kernel(const double *__restrict__ c64in, int k, int d,
                                                           unsigned char *__restrict__ pack) {
    const KmPackLayout L = km_pack_layout(k, d);
    const float *mu = reinterpret_cast<const float *>(pack + L.off_mu);
    const int j = blockIdx.x;  // 0 .. ntiles*128
    float *c32 = reinterpret_cast<float *>(pack + L.off_c32);
    float *cn32 = reinterpret_cast<float *>(pack + L.off_cn32);
    double *cn64 = reinterpret_cast<double *>(pack + L.off_cn64);
    double *c64 = reinterpret_cast<double *>(pack + L.off_c64);
    float *ct = reinterpret_cast<float *>(pack + L.off_ct);
    unsigned int *cmax = reinterpret_cast<unsigned int *>(pack + L.off_cmax);
    const int tile = j / KM_TILE_N, jj = j % KM_TILE_N;
    float *ctile = ct + (size_t)tile * (d + 1) * KM_TILE_N;
    __shared__ double red[32];
    __shared__ double red2[32];
    if (j >= k) {
        for (int f = threadIdx.x; f < d; f += blockDim.x) ctile[(size_t)f * KM_TILE_N + jj] = 0.f;
        if (threadIdx.x == 0) {
            cn32[j] = INFINITY;
            ctile[(size_t)d * KM_TILE_N + jj] = INFINITY;
        }
        return;
    }
    double s64 = 0.0, s32 = 0.0;
    for (int f = threadIdx.x; f < d; f += blockDim.x) {
        const double v = c64in[(size_t)j * d + f];
        const float vf = (float)(v - (double)mu[f]);   // centred, then rounded
        c64[(size_t)j * d + f] = v;
        c32[(size_t)j * d + f] = vf;
        ctile[(size_t)f * KM_TILE_N + jj] = vf;
        s64 += v * v;
        s32 += (double)vf * (double)vf;
    }
    s64 = km_warp_sum(s64);
    s32 = km_warp_sum(s32);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { red[w] = s64; red2[w] = s32; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0, b = 0;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { a += red[i]; b += red2[i]; }
        cn64[j] = a;
        const float bf = (float)b;
        cn32[j] = bf;
        ctile[(size_t)d * KM_TILE_N + jj] = bf;
        atomicMax(cmax, __float_as_uint(bf));
        if (j == 0) {
            int *hdr = reinterpret_cast<int *>(pack + L.off_hdr);
            hdr[0] = k; hdr[1] = d; hdr[2] = (k + KM_TILE_N - 1) / KM_TILE_N; hdr[3] = 0x6b6d3031;
        }
    }
}

extern "C" size_t b200km_pack_bytes(int k, int d) {
    if (k <= 0 || d <= 0) return 0;
    return km_pack_layout(k, d).total;
}

extern "C" int b200km_pack_centers(const double *centers_f64, int k, int d, void *pack, void *stream) {
    KM_CHECK_ARG(centers_f64 && pack && k > 0 && d > 0, B200KM_EINVAL);
    KM_CHECK_ARG(k <= B200KM_MAX_K && d <= B200KM_MAX_D, B200KM_ETOOBIG);
    KM_CHECK_ARG((reinterpret_cast<uintptr_t>(pack) & 255) == 0, B200KM_EALIGN);
    const KmPackLayout L = km_pack_layout(k, d);
    cudaStream_t st = km_stream(stream);
    cudaError_t e = cudaMemsetAsync(static_cast<unsigned char *>(pack) + L.off_cmax, 0, 4, st);
    if (e != cudaSuccess) return (int)e;
    const int ntiles = (k + KM_TILE_N - 1) / KM_TILE_N;
    pack_mu_kernel<<<(d + 7) & ~7, 128, 0, st>>>(centers_f64, k, d, static_cast<unsigned char *>(pack));
    e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    pack_centers_kernel<<<ntiles * KM_TILE_N, 128, 0, st>>>(centers_f64, k, d, static_cast<unsigned char *>(pack));
    e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    return km_launch_pack_tc(k, d, pack, st);
}

// =============================================================================================
// tiled FP32 kernel
// =============================================================================================
constexpr int AT_BM = 128;       // rows per CTA
constexpr int AT_XLD = AT_BM + 4;  // padded leading dim of the transposed X tile
constexpr int AT_THREADS = 256;

__device__ __forceinline__ void at_track(float v, int col, float &best, float &second, int &bidx) {
    const bool lt = v < best;
    second = lt ? best : fminf(second, v);
    bidx = lt ? col : bidx;
    best = fminf(best, v);
}

template <bool VEC>
__global__ void __launch_bounds__(AT_THREADS, 2)
assign_tiled_f32_kernel(const float *__restrict__ X, int64_t n, int d, int64_t ldx,
                        const float *__restrict__ ct, const float *__restrict__ c32,
                        const float *__restrict__ mu, const unsigned int *__restrict__ cmax_bits, int k, int ntiles,
                        int32_t *__restrict__ labels, float *__restrict__ mind2,
                        int32_t *__restrict__ flag_rows, int32_t *__restrict__ flag_count) {
    extern __shared__ __align__(128) float smem[];
    float *Xs = smem;                              // [d][AT_XLD]
    float *Cs = smem + (size_t)d * AT_XLD;         // [2][d+1][128]
    __shared__ __align__(8) uint64_t bars[2];
    __shared__ float xn_s[AT_BM];

    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int64_t row0 = (int64_t)blockIdx.x * AT_BM;
    const uint32_t tile_elems = (uint32_t)(d + 1) * KM_TILE_N;
    const uint32_t tile_bytes = tile_elems * 4u;

    if (tid == 0) {
        km_mbar_init(&bars[0], 1);
        km_mbar_init(&bars[1], 1);
        km_fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
        const int pre = ntiles < 2 ? ntiles : 2;
        for (int t = 0; t < pre; ++t) {
            km_mbar_expect_tx(&bars[t], tile_bytes);
            km_bulk_g2s(Cs + (size_t)t * tile_elems, ct + (size_t)t * tile_elems, tile_bytes, &bars[t]);
        }
    }

    // ---- stage the X tile, transposed: Xs[f][m] ------------------------------------------------
    if (VEC) {
        const int d4 = d >> 2;
        const int qg = (d4 + 3) >> 2;
        const int warp = tid >> 5, lane = tid & 31;
        for (int wt = warp; wt < 16 * qg; wt += AT_THREADS / 32) {
            const int m = (wt & 15) * 8 + (lane & 7);
            const int q = (wt >> 4) * 4 + (lane >> 3);
            if (q < d4) {
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (row0 + m < n) v = __ldg(reinterpret_cast<const float4 *>(X + (row0 + m) * ldx) + q);
                const float4 mq = __ldg(reinterpret_cast<const float4 *>(mu) + q);   // the kernel works on x - mu
                v.x -= mq.x; v.y -= mq.y; v.z -= mq.z; v.w -= mq.w;
                float *dst = Xs + (size_t)(4 * q) * AT_XLD + m;
                dst[0] = v.x; dst[AT_XLD] = v.y; dst[2 * AT_XLD] = v.z; dst[3 * AT_XLD] = v.w;
            }
        }
    } else {
        for (int e = tid; e < AT_BM * d; e += AT_THREADS) {
            const int m = e / d, f = e - m * d;
            float v = 0.f;
            if (row0 + m < n) v = __ldg(X + (row0 + m) * ldx + f);
            Xs[(size_t)f * AT_XLD + m] = v - __ldg(mu + f);
        }
    }
    __syncthreads();
    if (tid < AT_BM) {
        float s = 0.f;
        for (int f = 0; f < d; ++f) { const float v = Xs[(size_t)f * AT_XLD + tid]; s = fmaf(v, v, s); }
        xn_s[tid] = s;
    }

    float best[8], second[8];
    int bidx[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { best[i] = INFINITY; second[i] = INFINITY; bidx[i] = 0; }

    const float *xa_p = Xs + ty * 4;
    for (int t = 0; t < ntiles; ++t) {
        const float *Cb = Cs + (size_t)(t & 1) * tile_elems;
        km_mbar_wait(&bars[t & 1], (uint32_t)((t >> 1) & 1));
        float acc[8][8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
        const float *ca_p = Cb + tx * 4;
#pragma unroll 4
        for (int f = 0; f < d; ++f) {
            const float4 xa = *reinterpret_cast<const float4 *>(xa_p + (size_t)f * AT_XLD);
            const float4 xb = *reinterpret_cast<const float4 *>(xa_p + (size_t)f * AT_XLD + 64);
            const float4 ca = *reinterpret_cast<const float4 *>(ca_p + (size_t)f * KM_TILE_N);
            const float4 cb = *reinterpret_cast<const float4 *>(ca_p + (size_t)f * KM_TILE_N + 64);
            const float xr[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
            const float cr[8] = {ca.x, ca.y, ca.z, ca.w, cb.x, cb.y, cb.z, cb.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(xr[i], cr[j], acc[i][j]);
        }
        {
            const float4 na = *reinterpret_cast<const float4 *>(ca_p + (size_t)d * KM_TILE_N);
            const float4 nb = *reinterpret_cast<const float4 *>(ca_p + (size_t)d * KM_TILE_N + 64);
            const float cn[8] = {na.x, na.y, na.z, na.w, nb.x, nb.y, nb.z, nb.w};
            const int colbase = t * KM_TILE_N + tx * 4;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int col = colbase + (j < 4 ? j : 60 + j);
                    const float v = fmaf(-2.f, acc[i][j], cn[j]);
                    at_track(v, col, best[i], second[i], bidx[i]);
                }
            }
        }
        __syncthreads();  // everyone is done reading Cs[t&1]
        if (tid == 0 && t + 2 < ntiles) {
            km_mbar_expect_tx(&bars[t & 1], tile_bytes);
            km_bulk_g2s(Cs + (size_t)(t & 1) * tile_elems, ct + (size_t)(t + 2) * tile_elems, tile_bytes, &bars[t & 1]);
        }
    }

    // ---- merge the 16 column-threads of each row group (lanes tx = 0..15 of a half warp) -------
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best[i], o);
            const float os = __shfl_xor_sync(0xffffffffu, second[i], o);
            const int oi = __shfl_xor_sync(0xffffffffu, bidx[i], o);
            const bool take = (ob < best[i]) || (ob == best[i] && oi < bidx[i]);
            const float loser = take ? best[i] : ob;
            second[i] = fminf(fminf(second[i], os), loser);
            best[i] = take ? ob : best[i];
            bidx[i] = take ? oi : bidx[i];
        }
    }

    // ---- winner distance in direct form, outputs, near-tie flags --------------------------------
    const float cmax = __uint_as_float(*cmax_bits);
    const float bound_scale = (float)(d + 10) * 1.1920929e-07f;  // (d+10) * 2^-23
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = (i < 4 ? ty * 4 + i : 60 + ty * 4 + i);
        int bi = bidx[i];
        if (bi >= k) bi = k - 1;  // only possible for padded tail rows
        const float *crow = c32 + (size_t)bi * d;
        float s = 0.f;
        for (int f = tx; f < d; f += 16) {
            const float df = Xs[(size_t)f * AT_XLD + m] - __ldg(crow + f);
            s = fmaf(df, df, s);
        }
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (tx == i) {
            const int64_t row = row0 + m;
            if (row < n) {
                labels[row] = bi;
                if (mind2) mind2[row] = s;
                if (flag_rows) {
                    const float thr = bound_scale * (xn_s[m] + cmax);
                    if (second[i] - best[i] <= thr) {
                        const int pos = atomicAdd(flag_count, 1);
                        flag_rows[pos] = (int32_t)row;
                    }
                }
            }
        }
    }
}

// =============================================================================================
// row-block kernel (fp64 arithmetic, direct form)
// =============================================================================================
constexpr int RB_DC = 128;  // feature chunk

template <typename TX, int R, int G, int JB, bool INDIRECT>
__global__ void __launch_bounds__(R *G)
assign_rb_kernel(const TX *__restrict__ X, int64_t n, int d, int64_t ldx, const double *__restrict__ c64,
                 const double *__restrict__ cn64, int k,
                 const int32_t *__restrict__ rows, const int32_t *__restrict__ nrows_ptr,
                 int32_t *__restrict__ labels, TX *__restrict__ mind2,
                 const int32_t *__restrict__ gate_ptr = nullptr, int64_t gate_max = 0, int gate_run_if_le = 0) {
    // optional gate (pruned re-decision): run only if *gate_ptr <= gate_max (or > gate_max); uniform over the grid
    if (gate_ptr) {
        const int64_t c = (int64_t)(*gate_ptr);
        if (gate_run_if_le ? (c > gate_max) : (c <= gate_max)) return;
    }
    // INDIRECT (the refine kernel) decides on the expanded form ||c||^2 - 2 x.c in float64 -- one DFMA per
    // term, the same arithmetic sklearn's ArgKmin uses -- and recomputes only the winner's distance in
    // direct form; the direct kernels accumulate (x-c)^2 throughout.
    constexpr bool NORMFORM = INDIRECT;
    extern __shared__ __align__(16) double smem_d[];
    const int dc_max = d < RB_DC ? d : RB_DC;
    const int xld = dc_max + 1;
    double *xs = smem_d;                        // [R][xld]
    double *cs = xs + (size_t)R * xld;          // [G*JB][dc_max]
    double *sb = cs + (size_t)G * JB * dc_max;  // [G][R] best values
    int *si = reinterpret_cast<int *>(sb + (size_t)G * R);  // [G][R] best idx

    const int tid = threadIdx.x;
    const int r = tid % R, g = tid / R;
    const int64_t nrows = INDIRECT ? (int64_t)(*nrows_ptr) : n;
    const int nthreads = R * G;
    const bool single_chunk = d <= RB_DC;

    for (int64_t base = (int64_t)blockIdx.x * R; base < nrows; base += (int64_t)gridDim.x * R) {
        double best = INFINITY;
        int bidx = 0;
        for (int c0 = 0; c0 < k; c0 += G * JB) {
            double acc[JB];
#pragma unroll
            for (int j = 0; j < JB; ++j) acc[j] = 0.0;
            for (int d0 = 0; d0 < d; d0 += RB_DC) {
                const int dc = (d - d0) < RB_DC ? (d - d0) : RB_DC;
                __syncthreads();
                if (!single_chunk || c0 == 0) {
                    for (int e = tid; e < R * dc; e += nthreads) {
                        const int rr = e / dc, f = e - rr * dc;
                        double v = 0.0;
                        if (base + rr < nrows) {
                            const int64_t row = INDIRECT ? (int64_t)rows[base + rr] : base + rr;
                            v = (double)X[row * ldx + d0 + f];
                        }
                        xs[(size_t)rr * xld + f] = v;
                    }
                }
                for (int e = tid; e < G * JB * dc; e += nthreads) {
                    const int jj = e / dc, f = e - jj * dc;
                    const int col = c0 + jj;
                    cs[(size_t)jj * dc_max + f] = col < k ? c64[(size_t)col * d + d0 + f] : 0.0;
                }
                __syncthreads();
                const double *xr = xs + (size_t)r * xld;
                const double *cr = cs + (size_t)(g * JB) * dc_max;
#pragma unroll 2
                for (int f = 0; f < dc; ++f) {
                    const double xv = xr[f];
#pragma unroll
                    for (int j = 0; j < JB; ++j) {
                        if (NORMFORM) {
                            acc[j] = fma(xv, cr[(size_t)j * dc_max + f], acc[j]);
                        } else {
                            const double df = xv - cr[(size_t)j * dc_max + f];
                            acc[j] = fma(df, df, acc[j]);
                        }
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < JB; ++j) {
                const int col = c0 + g * JB + j;
                if (col < k) {
                    const double sc = NORMFORM ? fma(-2.0, acc[j], __ldg(cn64 + col)) : acc[j];
                    if (sc < best) { best = sc; bidx = col; }
                }
            }
        }
        if (G > 1) {
            __syncthreads();
            sb[g * R + r] = best;
            si[g * R + r] = bidx;
            __syncthreads();
            if (g == 0) {
                for (int gg = 1; gg < G; ++gg) {
                    const double ob = sb[gg * R + r];
                    const int oi = si[gg * R + r];
                    if (ob < best || (ob == best && oi < bidx)) { best = ob; bidx = oi; }
                }
            }
        }
        if (g == 0 && base + r < nrows) {
            const int64_t row = INDIRECT ? (int64_t)rows[base + r] : base + r;
            labels[row] = bidx;
            if (mind2) {
                if (NORMFORM) {
                    const TX *xr = X + row * ldx;
                    const double *cr = c64 + (size_t)bidx * d;
                    double s = 0.0;
                    for (int f = 0; f < d; ++f) {
                        const double df = (double)xr[f] - cr[f];
                        s = fma(df, df, s);
                    }
                    best = s;
                }
                mind2[row] = (TX)best;
            }
        }
    }
}

template <typename TX, int R, int G, int JB, bool INDIRECT>
static int launch_rb(const TX *X, int64_t n, int d, int64_t ldx, const double *c64, const double *cn64, int k,
                     const int32_t *rows,
                     const int32_t *nrows_ptr, int32_t *labels, TX *mind2, cudaStream_t st,
                     const int32_t *gate_ptr = nullptr, int64_t gate_max = 0, int gate_run_if_le = 0) {
    const int dc_max = d < RB_DC ? d : RB_DC;
    const size_t smem = sizeof(double) * ((size_t)R * (dc_max + 1) + (size_t)G * JB * dc_max + (size_t)G * R) +
                        sizeof(int) * (size_t)G * R;
    auto kern = assign_rb_kernel<TX, R, G, JB, INDIRECT>;
    static bool attr_set[KM_MAX_DEVICES] = {};
    if (int rc = km_set_smem_once(kern, 200 * 1024, attr_set)) return rc;
    int64_t blocks = (n + R - 1) / R;
    const int64_t cap = (int64_t)km_num_sms() * 8;
    if (INDIRECT || blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    kern<<<(unsigned)blocks, R * G, smem, st>>>(X, n, d, ldx, c64, cn64, k, rows, nrows_ptr, labels, mind2, gate_ptr,
                                                gate_max, gate_run_if_le);
    KM_LAUNCH_RET();
}


// =============================================================================================
// tiled FP64 kernel (DFMA): float64 X with k > 16, d <= 128
// =============================================================================================
// Same structure as the FFMA kernel in double precision: 64 rows x 64 centroids per CTA step, 4x4 register
// tile per thread, scores ||c||^2 - 2 x.c (sklearn's own float64 arithmetic), winner distance recomputed in
// direct form.  float64 needs no re-decision pass.  Centroid tiles are staged from the pack's float64 copy
// with coalesced loads (the next tile is prefetched into registers while the current one is multiplied).
constexpr int AD_BM = 128, AD_BN = 64, AD_THREADS = 256;   // 8 x 4 register tile per thread
constexpr int AD_XLD = AD_BM + 2;

// TX = storage type of X (float for the refine use); INDIRECT: the rows are rows[0 .. *nrows_ptr) (the flagged rows
// of a float32 pass -- count read on the device) and the CTAs stride over them persistently.
template <typename TX, bool INDIRECT>
__global__ void __launch_bounds__(AD_THREADS, 2)
assign_tiled_f64_kernel(const TX *__restrict__ X, int64_t n, int d, int64_t ldx, const double *__restrict__ c64,
                        const double *__restrict__ cn64, int k, const int32_t *__restrict__ rows,
                        const int32_t *__restrict__ nrows_ptr, int32_t *__restrict__ labels,
                        TX *__restrict__ mind2) {
    extern __shared__ __align__(16) double smd[];
    double *Xs = smd;                               // [d][AD_XLD]   (transposed X tile)
    double *Cs = smd + (size_t)d * AD_XLD;          // [d][AD_BN]    (transposed centroid tile)
    double *cns = Cs + (size_t)d * AD_BN;           // [AD_BN]
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int ntiles = (k + AD_BN - 1) / AD_BN;
    const int64_t nrows = INDIRECT ? (int64_t)(*nrows_ptr) : n;
  for (int64_t row0 = (int64_t)blockIdx.x * AD_BM; row0 < nrows; row0 += (int64_t)gridDim.x * AD_BM) {
    __syncthreads();   // (persistent INDIRECT loop) the previous row tile is fully written out
    for (int e = tid; e < AD_BM * d; e += AD_THREADS) {
        const int m = e / d, f = e - m * d;
        double v = 0.0;
        if (row0 + m < nrows) {
            const int64_t r = INDIRECT ? (int64_t)rows[row0 + m] : row0 + m;
            v = (double)X[r * ldx + f];
        }
        Xs[(size_t)f * AD_XLD + m] = v;
    }
    double best[8];
    int bidx[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { best[i] = INFINITY; bidx[i] = 0; }

    for (int t = 0; t < ntiles; ++t) {
        __syncthreads();   // previous tile fully consumed (also orders the Xs stores before the first use)
        // stage the centroid tile transposed: element e -> centroid jj = e / d, feature f = e % d (coalesced)
        for (int e = tid; e < AD_BN * d; e += AD_THREADS) {
            const int jj = e / d, f = e - jj * d;
            const int col = t * AD_BN + jj;
            Cs[(size_t)f * AD_BN + jj] = col < k ? __ldg(c64 + (size_t)col * d + f) : 0.0;
        }
        if (tid < AD_BN) {
            const int col = t * AD_BN + tid;
            cns[tid] = col < k ? __ldg(cn64 + col) : INFINITY;
        }
        __syncthreads();
        double acc[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
#pragma unroll 2
        for (int f = 0; f < d; ++f) {
            const double *xp = Xs + (size_t)f * AD_XLD + ty * 8;
            const double *cp = Cs + (size_t)f * AD_BN + tx * 4;
            const double2 x0 = *reinterpret_cast<const double2 *>(xp), x1 = *reinterpret_cast<const double2 *>(xp + 2);
            const double2 x2 = *reinterpret_cast<const double2 *>(xp + 4), x3 = *reinterpret_cast<const double2 *>(xp + 6);
            const double2 ca = *reinterpret_cast<const double2 *>(cp), cb = *reinterpret_cast<const double2 *>(cp + 2);
            const double xr[8] = {x0.x, x0.y, x1.x, x1.y, x2.x, x2.y, x3.x, x3.y};
            const double cr[4] = {ca.x, ca.y, cb.x, cb.y};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fma(xr[i], cr[j], acc[i][j]);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const double v = fma(-2.0, acc[i][j], cns[tx * 4 + j]);
                if (v < best[i]) { best[i] = v; bidx[i] = t * AD_BN + tx * 4 + j; }
            }
    }
    // merge the 16 column threads of each row group (lowest index on exact ties)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best[i], o);
            const int oi = __shfl_xor_sync(0xffffffffu, bidx[i], o);
            if (ob < best[i] || (ob == best[i] && oi < bidx[i])) { best[i] = ob; bidx[i] = oi; }
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = ty * 8 + i;
        int bi = bidx[i];
        if (bi >= k) bi = k - 1;
        const double *crow = c64 + (size_t)bi * d;
        double s = 0.0;
        for (int f = tx; f < d; f += 16) {
            const double df = Xs[(size_t)f * AD_XLD + m] - __ldg(crow + f);
            s = fma(df, df, s);
        }
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (tx == i && row0 + m < nrows) {
            const int64_t r = INDIRECT ? (int64_t)rows[row0 + m] : row0 + m;
            labels[r] = bi;
            if (mind2) mind2[r] = (TX)s;
        }
    }
  }
}

template <typename TX, bool INDIRECT>
static int launch_tiled_f64(const TX *X, int64_t n, int d, int64_t ldx, const double *c64, const double *cn64, int k,
                            const int32_t *rows, const int32_t *nrows_ptr, int32_t *labels, TX *mind2,
                            cudaStream_t st) {
    const size_t smem = sizeof(double) * ((size_t)d * AD_XLD + (size_t)d * AD_BN + AD_BN);
    auto kern = assign_tiled_f64_kernel<TX, INDIRECT>;
    static bool attr_set[KM_MAX_DEVICES] = {};
    if (int rc = km_set_smem_once(kern, 200 * 1024, attr_set)) return rc;
    int64_t blocks = (n + AD_BM - 1) / AD_BM;
    if (INDIRECT) blocks = (int64_t)km_num_sms() * 2;   // persistent: the row count lives on the device
    kern<<<(unsigned)blocks, AD_THREADS, smem, st>>>(X, n, d, ldx, c64, cn64, k, rows, nrows_ptr, labels, mind2);
    KM_LAUNCH_RET();
}

// =============================================================================================
// entry points
// =============================================================================================

// (atomics: the options are process-wide and may be set from one thread while another launches; a reader sees the old
//  or the new value, never a torn one.  The pack layout follows variant 3, so set it BEFORE packing -- see the header.)
static std::atomic<int> g_k1_variant{0};  // B200KM_OPT_K1_VARIANT: 0 auto, 1 FFMA, 2 tcgen05 (one CTA), 3 tcgen05 (CTA pair)
bool km_tc_pair_mode() { return g_k1_variant == 3; }
bool km_tc_small_enabled() { return g_k1_variant == 0; }   // auto only: variant 2 keeps the general tcgen05 kernel
static std::atomic<int> g_k2_variant{0};  // B200KM_OPT_K2_VARIANT: 0 auto, 1 owner table, 2 chunk sort, 3 global atomics
int km_k2_variant() { return g_k2_variant; }
static std::atomic<int> g_fused_variant{0};  // B200KM_OPT_FUSED_VARIANT: 0 auto, 1 sorted-tile kernel, 2 cp.async kernel where it fits
int km_fused_variant() { return g_fused_variant; }
static std::atomic<int> g_k4_variant{0};  // B200KM_OPT_K4_VARIANT: 0 auto (TMA ring where it pays), 1 register path always, 2 TMA ring wherever it fits
int km_k4_variant() { return g_k4_variant; }

extern "C" int b200km_set_option(int key, int value) {
    if (key == B200KM_OPT_K1_VARIANT && value >= 0 && value <= 3) {
        g_k1_variant = value;
        return 0;
    }
    if (key == B200KM_OPT_K2_VARIANT && value >= 0 && value <= 3) {
        g_k2_variant = value;
        return 0;
    }
    if (key == B200KM_OPT_FUSED_VARIANT && value >= 0 && value <= 2) {
        g_fused_variant = value;
        return 0;
    }
    if (key == B200KM_OPT_K4_VARIANT && value >= 0 && value <= 2) {
        g_k4_variant = value;
        return 0;
    }
    return B200KM_EINVAL;
}
extern "C" int b200km_get_option(int key) {
    if (key == B200KM_OPT_K1_VARIANT) return g_k1_variant;
    if (key == B200KM_OPT_K2_VARIANT) return g_k2_variant;
    if (key == B200KM_OPT_FUSED_VARIANT) return g_fused_variant;
    if (key == B200KM_OPT_K4_VARIANT) return g_k4_variant;
    return B200KM_EINVAL;
}

static bool use_tc_path(int d, int k) {
    if (d > KM_TC_WIDE_MAX_D) return false;
    if (d > KM_TC_MAX_D && g_k1_variant == 3) return false;   // the wide-row kernel has no CTA-pair form
    if (g_k1_variant == 1) return false;
    if (g_k1_variant >= 2) return true;
    return true;   // auto: tensor cores whenever the shape allows (d <= 2048)
}
static int check_pack_args(const void *X, int64_t n, int d, int64_t ldx, const void *pack, int k, const void *labels) {
    KM_CHECK_ARG(X && pack && labels, B200KM_EINVAL);
    KM_CHECK_ARG(n > 0 && d > 0 && k > 0 && ldx >= d, B200KM_EINVAL);
    KM_CHECK_ARG(k <= B200KM_MAX_K && d <= B200KM_MAX_D && n < (int64_t(1) << 31), B200KM_ETOOBIG);
    KM_CHECK_ARG((reinterpret_cast<uintptr_t>(pack) & 255) == 0, B200KM_EALIGN);
    return 0;
}

static int assign_f32_impl(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                           int32_t *labels, float *mind2, int32_t *flag_rows, int32_t *flag_count,
                           int reset_flags, void *stream, bool precentred, bool sparse_d2 = false) {
    int rc = check_pack_args(X, n, d, ldx, pack, k, labels);
    if (rc) return rc;
    // rows already centred on the pack's mu: only the kernels that work on centred data can take them
    KM_CHECK_ARG(!precentred || (k > 16 && (d <= 128 || use_tc_path(d, k))), B200KM_EINVAL);
    KM_CHECK_ARG(!flag_rows || flag_count, B200KM_EINVAL);
    cudaStream_t st = km_stream(stream);
    const KmPackLayout L = km_pack_layout(k, d);
    const unsigned char *pk = static_cast<const unsigned char *>(pack);
    if (flag_count && reset_flags) {
        cudaError_t e = cudaMemsetAsync(flag_count, 0, sizeof(int32_t), st);
        if (e != cudaSuccess) return (int)e;
    }
    const double *c64 = reinterpret_cast<const double *>(pk + L.off_c64);
    if (k <= 16 && d <= 32) {
        // HBM-bound regime: fp64 arithmetic is free here, so no flags are ever needed.
        return launch_rb<float, 256, 1, 8, false>(X, n, d, ldx, c64, nullptr, k, nullptr, nullptr, labels, mind2, st);
    }
    if (k <= 16 || (d > 128 && !use_tc_path(d, k))) {
        return launch_rb<float, 32, 8, 8, false>(X, n, d, ldx, c64, nullptr, k, nullptr, nullptr, labels, mind2, st);
    }
    if (use_tc_path(d, k))
        return km_launch_assign_tc(X, n, d, ldx, pack, k, labels, mind2, flag_rows, flag_count, st, precentred, sparse_d2);
    const int ntiles = (k + KM_TILE_N - 1) / KM_TILE_N;
    const size_t smem = sizeof(float) * ((size_t)d * AT_XLD + 2 * (size_t)(d + 1) * KM_TILE_N);
    const bool vec = (d % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
    const int64_t blocks = (n + AT_BM - 1) / AT_BM;
    const float *ct = reinterpret_cast<const float *>(pk + L.off_ct);
    const float *c32 = reinterpret_cast<const float *>(pk + L.off_c32);
    const unsigned int *cmax = reinterpret_cast<const unsigned int *>(pk + L.off_cmax);
    const float *mu = reinterpret_cast<const float *>(pk + (precentred ? L.off_zero : L.off_mu));
    static bool attr_v[KM_MAX_DEVICES] = {}, attr_s[KM_MAX_DEVICES] = {};
    if (int rc2 = km_set_smem_once(assign_tiled_f32_kernel<true>, 220 * 1024, attr_v)) return rc2;
    if (int rc2 = km_set_smem_once(assign_tiled_f32_kernel<false>, 220 * 1024, attr_s)) return rc2;
    if (vec)
        assign_tiled_f32_kernel<true><<<(unsigned)blocks, AT_THREADS, smem, st>>>(
            X, n, d, ldx, ct, c32, mu, cmax, k, ntiles, labels, mind2, flag_rows, flag_count);
    else
        assign_tiled_f32_kernel<false><<<(unsigned)blocks, AT_THREADS, smem, st>>>(
            X, n, d, ldx, ct, c32, mu, cmax, k, ntiles, labels, mind2, flag_rows, flag_count);
    KM_LAUNCH_RET();
}

extern "C" int b200km_assign_f32(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                                 int32_t *labels, float *mind2, int32_t *flag_rows, int32_t *flag_count,
                                 int reset_flags, void *stream) {
    return assign_f32_impl(X, n, d, ldx, pack, k, labels, mind2, flag_rows, flag_count, reset_flags, stream, false);
}
extern "C" int b200km_assign_lloyd_f32(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                                       int32_t *labels, float *score, int32_t *flag_rows, int32_t *flag_count,
                                       int reset_flags, void *stream) {
    KM_CHECK_ARG(score && flag_rows && flag_count, B200KM_EINVAL);
    return assign_f32_impl(X, n, d, ldx, pack, k, labels, score, flag_rows, flag_count, reset_flags, stream, false, true);
}
extern "C" int b200km_assign_centred_f32(const float *Xc, int64_t n, int d, int64_t ldx, const void *pack, int k,
                                         int32_t *labels, float *mind2, int32_t *flag_rows, int32_t *flag_count,
                                         int reset_flags, void *stream) {
    return assign_f32_impl(Xc, n, d, ldx, pack, k, labels, mind2, flag_rows, flag_count, reset_flags, stream, true);
}

extern "C" int b200km_assign_f64(const double *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                                 int32_t *labels, double *mind2, void *stream) {
    int rc = check_pack_args(X, n, d, ldx, pack, k, labels);
    if (rc) return rc;
    cudaStream_t st = km_stream(stream);
    const KmPackLayout L = km_pack_layout(k, d);
    const double *c64 = reinterpret_cast<const double *>(static_cast<const unsigned char *>(pack) + L.off_c64);
    if (k <= 16 && d <= 32)
        return launch_rb<double, 256, 1, 8, false>(X, n, d, ldx, c64, nullptr, k, nullptr, nullptr, labels, mind2, st);
    if (k > 16 && d <= 128 && g_k1_variant != 1) {   // the DFMA tile kernel (variant 1 keeps the row-block kernel)
        const double *cn64 = reinterpret_cast<const double *>(static_cast<const unsigned char *>(pack) + L.off_cn64);
        return launch_tiled_f64<double, false>(X, n, d, ldx, c64, cn64, k, nullptr, nullptr, labels, mind2, st);
    }
    return launch_rb<double, 32, 8, 8, false>(X, n, d, ldx, c64, nullptr, k, nullptr, nullptr, labels, mind2, st);
}

extern "C" int b200km_refine_f64(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                                 const int32_t *flag_rows, const int32_t *flag_count, int32_t *labels,
                                 float *mind2, void *stream) {
    int rc = check_pack_args(X, n, d, ldx, pack, k, labels);
    if (rc) return rc;
    KM_CHECK_ARG(flag_rows && flag_count, B200KM_EINVAL);
    cudaStream_t st = km_stream(stream);
    const KmPackLayout L = km_pack_layout(k, d);
    const double *c64 = reinterpret_cast<const double *>(static_cast<const unsigned char *>(pack) + L.off_c64);
    const double *cn64 = reinterpret_cast<const double *>(static_cast<const unsigned char *>(pack) + L.off_cn64);
    // (measured: with ~0.2-0.5 % of the rows flagged, the row-block kernel's 32-row groups spread over more CTAs
    //  and beat the DFMA tile kernel's 128-row tiles at every BASELINE size, so it stays the refine kernel)
    return launch_rb<float, 32, 8, 8, true>(X, n, d, ldx, c64, cn64, k, flag_rows, flag_count, labels, mind2, st);
}

// =============================================================================================
// float64 X on the tensor cores: centre in float64, round once to float32, decide in float32 with the same
// near-tie flags, re-decide the flagged rows in float64, then rebuild every winner distance in float64
// =============================================================================================
// The reference's arithmetic is float64 for float64 X as well (sklearn ArgKmin, reference pairwise.py:19-47), and the
// DFMA tile kernel does exactly that at ~16 TFLOP/s.  But the labels only need float64 where the float32 decision is
// inside its own rounding bound -- the argument that carries the float32 path -- and the bound already contains
// the one extra rounding here (x - mu computed in float64, then rounded: <= 2^-24 relative to the CENTRED value).
// So: b200km_center_convert_f64 writes fl32(x - mu) for a block of rows, b200km_assign_centred_f32 runs the
// tcgen05 kernels on it (mu := 0), b200km_refine_x64_f64 re-decides the flagged rows on the float64 rows, and
// b200km_rowdist_f64 writes sum_f (x_f - c_f)^2 in float64 for every row (the float32 kernels' distances are
// only good to ~1e-7, the float64 contract is 1e-10).
__global__ void __launch_bounds__(256)
center_convert_kernel(const double *__restrict__ X, int64_t n, int d, int64_t ldx, const float *__restrict__ mu,
                      float *__restrict__ out) {
    const int64_t total = n * (int64_t)d;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / d;
        const int f = (int)(e - r * d);
        out[e] = (float)(X[r * ldx + f] - (double)__ldg(mu + f));
    }
}

extern "C" int b200km_center_convert_f64(const double *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                                         float *out, void *stream) {
    KM_CHECK_ARG(X && pack && out, B200KM_EINVAL);
    KM_CHECK_ARG(n > 0 && d > 0 && k > 0 && ldx >= d, B200KM_EINVAL);
    KM_CHECK_ARG(k <= B200KM_MAX_K && d <= B200KM_MAX_D, B200KM_ETOOBIG);
    const KmPackLayout L = km_pack_layout(k, d);
    const float *mu = reinterpret_cast<const float *>(static_cast<const unsigned char *>(pack) + L.off_mu);
    int64_t blocks = (n * (int64_t)d + 255) / 256;
    if (blocks > (int64_t)km_num_sms() * 16) blocks = (int64_t)km_num_sms() * 16;
    center_convert_kernel<<<(unsigned)blocks, 256, 0, km_stream(stream)>>>(X, n, d, ldx, mu, out);
    KM_LAUNCH_RET();
}

// mind2[r] = sum_f (x_rf - c_{label_r, f})^2 in float64; LPR lanes per row (coalesced), centre rows from L2
template <int LPR>
__global__ void __launch_bounds__(256)
rowdist_f64_kernel(const double *__restrict__ X, int64_t n, int d, int64_t ldx, const double *__restrict__ c64,
                   const int32_t *__restrict__ labels, double *__restrict__ mind2) {
    constexpr int RPW = 32 / LPR;
    const int lane = threadIdx.x & 31;
    const int sub = lane / LPR, l = lane % LPR;
    const int64_t warp_global = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r0 = warp_global * RPW; r0 < n; r0 += nwarps * RPW) {
        const int64_t row = r0 + sub;
        const bool live = row < n;
        const double *xr = X + (live ? row : 0) * ldx;
        const double *cr = c64 + (size_t)(live ? labels[row] : 0) * d;
        double s0 = 0.0, s1 = 0.0;
        int f = l;
        for (; f + LPR < d; f += 2 * LPR) {
            const double a = xr[f] - __ldg(cr + f), b = xr[f + LPR] - __ldg(cr + f + LPR);
            s0 = fma(a, a, s0);
            s1 = fma(b, b, s1);
        }
        if (f < d) { const double a = xr[f] - __ldg(cr + f); s0 = fma(a, a, s0); }
        double s = s0 + s1;
#pragma unroll
        for (int o = LPR / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (live && l == 0) mind2[row] = s;
    }
}

extern "C" int b200km_rowdist_f64(const double *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                                  const int32_t *labels, double *mind2, void *stream) {
    int rc = check_pack_args(X, n, d, ldx, pack, k, labels);
    if (rc) return rc;
    KM_CHECK_ARG(mind2, B200KM_EINVAL);
    const KmPackLayout L = km_pack_layout(k, d);
    const double *c64 = reinterpret_cast<const double *>(static_cast<const unsigned char *>(pack) + L.off_c64);
    cudaStream_t st = km_stream(stream);
    int lpr = 32;
    if (d <= 2) lpr = 1; else if (d <= 4) lpr = 2; else if (d <= 8) lpr = 4; else if (d <= 16) lpr = 8;
    else if (d <= 32) lpr = 16;
    const int rpw = 32 / lpr;
    int64_t blocks = (n + (int64_t)rpw * 8 - 1) / ((int64_t)rpw * 8);
    if (blocks > (int64_t)km_num_sms() * 16) blocks = (int64_t)km_num_sms() * 16;
    switch (lpr) {
        case 1: rowdist_f64_kernel<1><<<(unsigned)blocks, 256, 0, st>>>(X, n, d, ldx, c64, labels, mind2); break;
        case 2: rowdist_f64_kernel<2><<<(unsigned)blocks, 256, 0, st>>>(X, n, d, ldx, c64, labels, mind2); break;
        case 4: rowdist_f64_kernel<4><<<(unsigned)blocks, 256, 0, st>>>(X, n, d, ldx, c64, labels, mind2); break;
        case 8: rowdist_f64_kernel<8><<<(unsigned)blocks, 256, 0, st>>>(X, n, d, ldx, c64, labels, mind2); break;
        case 16: rowdist_f64_kernel<16><<<(unsigned)blocks, 256, 0, st>>>(X, n, d, ldx, c64, labels, mind2); break;
        default: rowdist_f64_kernel<32><<<(unsigned)blocks, 256, 0, st>>>(X, n, d, ldx, c64, labels, mind2); break;
    }
    KM_LAUNCH_RET();
}

// b200km_refine_f64 for float64 rows (labels updated in place; mind2 nullable -- b200km_rowdist_f64 follows anyway)
extern "C" int b200km_refine_x64_f64(const double *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                                     const int32_t *flag_rows, const int32_t *flag_count, int32_t *labels,
                                     double *mind2, void *stream) {
    int rc = check_pack_args(X, n, d, ldx, pack, k, labels);
    if (rc) return rc;
    KM_CHECK_ARG(flag_rows && flag_count, B200KM_EINVAL);
    const KmPackLayout L = km_pack_layout(k, d);
    const double *c64 = reinterpret_cast<const double *>(static_cast<const unsigned char *>(pack) + L.off_c64);
    const double *cn64 = reinterpret_cast<const double *>(static_cast<const unsigned char *>(pack) + L.off_cn64);
    return launch_rb<double, 32, 8, 8, true>(X, n, d, ldx, c64, cn64, k, flag_rows, flag_count, labels, mind2,
                                             km_stream(stream));
}

// =============================================================================================
// pruned float64 re-decision: only the centres the candidate kernel listed
// =============================================================================================
int km_launch_cand_tc(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k, int32_t *flag_rows,
                      int64_t flen, const int32_t *flag_count, const float *mind2, cudaStream_t st, bool precentred);
int km_launch_cand_tc_wide(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                           int32_t *flag_rows, int64_t flen, const int32_t *flag_count, float *mind2, cudaStream_t st,
                           bool precentred);

// One thread per flagged row: the listed columns are scored exactly as assign_rb_kernel<INDIRECT> scores them
// (cn64 - 2 * sum_f x_f c_f, one DFMA per term in feature order, strict < then lowest index), so the two paths
// give the same labels bit for bit.  A row with more listed columns than its list holds (km_cand_cap; duplicate centres, e.g.
// several empty clusters parked at the origin) goes to the overflow list and gets the full scan.
template <typename TX>
__global__ void __launch_bounds__(128)
refine_cand_kernel(const TX *__restrict__ X, int64_t n, int d, int64_t ldx, const double *__restrict__ c64,
                   const double *__restrict__ cn64, int32_t *__restrict__ F, int64_t flen,
                   const int32_t *__restrict__ flag_count, int32_t *__restrict__ labels, TX *__restrict__ mind2) {
    const int64_t M = km_cand_rows_max(flen);
    const int64_t cnt = (int64_t)(*flag_count);
    if (cnt > M) return;
    const int cap = km_cand_cap(flen, cnt);
    const int32_t *cand = F + M;
    int32_t *ovf = F + (flen - 1 - M);
    for (int64_t pos = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; pos < cnt;
         pos += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = (int64_t)F[pos];
        const int32_t *cl = cand + pos * (cap + 1);
        const int nc = cl[0];
        if (nc > cap || nc <= 0) {   // (0 cannot happen: the winner itself is always within the threshold)
            const int q = atomicAdd(F + (flen - 1), 1);
            ovf[q] = (int32_t)row;
            continue;
        }
        const TX *xr = X + row * ldx;
        double best = INFINITY;
        int bidx = 0;
        for (int q = 0; q < nc; ++q) {
            const int col = cl[1 + q];
            const double *cr = c64 + (size_t)col * d;
            double a = 0.0;
            for (int f = 0; f < d; ++f) a = fma((double)__ldg(xr + f), cr[f], a);
            const double sc = fma(-2.0, a, __ldg(cn64 + col));
            if (sc < best || (sc == best && col < bidx)) { best = sc; bidx = col; }
        }
        labels[row] = bidx;
        if (mind2) {
            const double *cr = c64 + (size_t)bidx * d;
            double s = 0.0;
            for (int f = 0; f < d; ++f) {
                const double df = (double)__ldg(xr + f) - cr[f];
                s = fma(df, df, s);
            }
            mind2[row] = (TX)s;
        }
    }
}

// X: the rows scored in float64 (float or double); Xc / ldxc / precentred / dist32: what the candidate pass reads --
// for float32 input the same rows and K1's mind2, for float64 input the centred float32 block and its distances.
template <typename TX>
static int refine_pruned_impl(const TX *X, int64_t n, int d, int64_t ldx, const float *Xc, int64_t ldxc, bool precentred,
                              float *dist32, const void *pack, int k, int32_t *flag_rows, int64_t flen,
                              const int32_t *flag_count, int32_t *labels, TX *mind2, cudaStream_t st) {
    const KmPackLayout L = km_pack_layout(k, d);
    const double *c64 = reinterpret_cast<const double *>(static_cast<const unsigned char *>(pack) + L.off_c64);
    const double *cn64 = reinterpret_cast<const double *>(static_cast<const unsigned char *>(pack) + L.off_cn64);
    const int64_t M = km_cand_rows_max(flen);
    // more flagged rows than the scratch layout holds: the full scan, gated on the device-side count
    int rc = launch_rb<TX, 32, 8, 8, true>(X, n, d, ldx, c64, cn64, k, flag_rows, flag_count, labels, mind2, st,
                                           flag_count, M, 0);
    if (rc) return rc;
    rc = d <= KM_TC_MAX_D
             ? km_launch_cand_tc(Xc, n, d, ldxc, pack, k, flag_rows, flen, flag_count, dist32, st, precentred)
             : km_launch_cand_tc_wide(Xc, n, d, ldxc, pack, k, flag_rows, flen, flag_count, dist32, st, precentred);
    if (rc) return rc;
    refine_cand_kernel<TX><<<km_num_sms() * 4, 128, 0, st>>>(X, n, d, ldx, c64, cn64, flag_rows, flen, flag_count, labels,
                                                             mind2);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    // rows with too many candidates: the full scan over the overflow list (its size is flag_rows[n-1])
    return launch_rb<TX, 32, 8, 8, true>(X, n, d, ldx, c64, cn64, k, flag_rows + (flen - 1 - M),
                                         flag_rows + (flen - 1), labels, mind2, st, flag_count, M, 1);
}

extern "C" int b200km_refine_pruned_len_f64(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                                            int32_t *flag_rows, int64_t flag_len, const int32_t *flag_count,
                                            int32_t *labels, float *mind2, void *stream) {
    int rc = check_pack_args(X, n, d, ldx, pack, k, labels);
    if (rc) return rc;
    KM_CHECK_ARG(flag_rows && flag_count && flag_len >= n, B200KM_EINVAL);
    // the candidate pass needs the tensor-core image (one-CTA layout) and the winner distances
    if (d > KM_TC_WIDE_MAX_D || km_tc_pair_mode() || !mind2 || n < 4096)
        return b200km_refine_f64(X, n, d, ldx, pack, k, flag_rows, flag_count, labels, mind2, stream);
    return refine_pruned_impl<float>(X, n, d, ldx, X, ldx, false, mind2, pack, k, flag_rows, flag_len, flag_count, labels,
                                     mind2, km_stream(stream));
}
extern "C" int b200km_refine_pruned_f64(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k,
                                        int32_t *flag_rows, const int32_t *flag_count, int32_t *labels,
                                        float *mind2, void *stream) {
    return b200km_refine_pruned_len_f64(X, n, d, ldx, pack, k, flag_rows, n, flag_count, labels, mind2, stream);
}

// the same for float64 rows: Xc = the centred float32 block b200km_center_convert_f64 wrote (ld = d), dist32 = the
// distances b200km_assign_centred_f32 wrote for it; mind2 (float64, nullable) is updated for the re-decided rows
extern "C" int b200km_refine_pruned_x64_len_f64(const double *X, int64_t n, int d, int64_t ldx, const float *Xc,
                                                float *dist32, const void *pack, int k, int32_t *flag_rows,
                                                int64_t flag_len, const int32_t *flag_count, int32_t *labels,
                                                double *mind2, void *stream) {
    int rc = check_pack_args(X, n, d, ldx, pack, k, labels);
    if (rc) return rc;
    KM_CHECK_ARG(flag_rows && flag_count && flag_len >= n, B200KM_EINVAL);
    if (d > KM_TC_WIDE_MAX_D || km_tc_pair_mode() || !Xc || !dist32 || n < 4096)
        return b200km_refine_x64_f64(X, n, d, ldx, pack, k, flag_rows, flag_count, labels, mind2, stream);
    return refine_pruned_impl<double>(X, n, d, ldx, Xc, d, true, dist32, pack, k, flag_rows, flag_len, flag_count, labels,
                                      mind2, km_stream(stream));
}
extern "C" int b200km_refine_pruned_x64_f64(const double *X, int64_t n, int d, int64_t ldx, const float *Xc,
                                            float *dist32, const void *pack, int k, int32_t *flag_rows,
                                            const int32_t *flag_count, int32_t *labels, double *mind2, void *stream) {
    return b200km_refine_pruned_x64_len_f64(X, n, d, ldx, Xc, dist32, pack, k, flag_rows, n, flag_count, labels, mind2,
                                            stream);
}

// =============================================================================================
// b200km_assign_static_f32 -- K1 with every byte of scratch inside the library
// =============================================================================================
// The regular path needs caller-allocated device scratch: the centre pack (b200km_pack_bytes) and
// an n-entry list for the rows that go to the float64 re-decision.  This variant keeps both in the
// array below, so a caller that owns nothing but its input and output buffers can still label rows:
// a plugin whose host framework owns all device memory and hands it exactly those buffers, a
// latency-critical service thread that must stay out of cudaMalloc (it synchronises the device),
// or a process that has already spent its device-memory budget.
//
// In HOST code `g_static_scratch` is only a stand-in for the symbol, not a device address:
// handing it to a kernel would fault (or, where the GPU can reach host memory, silently touch
// the wrong bytes).  The device address comes from cudaGetSymbolAddress, fetched once per device
// with that device current, and only that pointer is passed on.
constexpr size_t KM_STATIC_PACK_BYTES = B200KM_STATIC_MAX_PACK_BYTES;
constexpr size_t KM_STATIC_FLAG_BYTES = sizeof(int32_t) * (size_t)B200KM_STATIC_MAX_ROWS;
constexpr size_t KM_STATIC_BYTES = KM_STATIC_PACK_BYTES + KM_STATIC_FLAG_BYTES + 256;  // + counter slot

static __device__ __align__(256) unsigned char g_static_scratch[KM_STATIC_BYTES];

static unsigned char *g_static_ptr[KM_MAX_DEVICES] = {};

// Device address of the scratch array for the CURRENT device (looked up on the first call per device).
static int km_static_scratch(unsigned char **out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= KM_MAX_DEVICES) return B200KM_ETOOBIG;
    if (!g_static_ptr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return (int)e;
        g_static_ptr[dev] = static_cast<unsigned char *>(p);
    }
    *out = g_static_ptr[dev];
    return 0;
}

extern "C" int b200km_assign_static_f32(const float *X, int64_t n, int d, int64_t ldx, const double *centers_f64,
                                        int k, int32_t *labels, float *mind2, void *stream) {
    KM_CHECK_ARG(X && centers_f64 && labels, B200KM_EINVAL);
    KM_CHECK_ARG(n > 0 && d > 0 && k > 0 && ldx >= d, B200KM_EINVAL);
    KM_CHECK_ARG(k <= B200KM_MAX_K && d <= B200KM_MAX_D, B200KM_ETOOBIG);
    // the two limits of the built-in scratch
    if (n > (int64_t)B200KM_STATIC_MAX_ROWS) return B200KM_ESTATICROWS;
    if (km_pack_layout(k, d).total > KM_STATIC_PACK_BYTES) return B200KM_ESTATICPACK;
    unsigned char *base = nullptr;
    int rc = km_static_scratch(&base);
    if (rc) return rc;
    unsigned char *pack = base;
    int32_t *flag_rows = reinterpret_cast<int32_t *>(base + KM_STATIC_PACK_BYTES);
    int32_t *flag_count = reinterpret_cast<int32_t *>(base + KM_STATIC_PACK_BYTES + KM_STATIC_FLAG_BYTES);
    // The array keeps whatever the previous call left: the pack is rewritten in full here and the flag
    // counter is zeroed by b200km_assign_f32 (reset_flags = 1); stale flag_rows entries beyond the new
    // count are never read.
    rc = b200km_pack_centers(centers_f64, k, d, pack, stream);
    if (rc) return rc;
    rc = b200km_assign_f32(X, n, d, ldx, pack, k, labels, mind2, flag_rows, flag_count, 1, stream);
    if (rc) return rc;
    return b200km_refine_f64(X, n, d, ldx, pack, k, flag_rows, flag_count, labels, mind2, stream);
}
