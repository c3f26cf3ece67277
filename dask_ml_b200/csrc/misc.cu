// misc.cu -- K7 full distance matrix (transform), row norms, finite check, blob generator,
// FFMA/DFMA pipe probes, version / error strings.
#include "common.cuh"

// ---------------------------------------------------------------------------------------------
// K7  euclidean_distances in the input precision.
// Replaces dask_ml.metrics.euclidean_distances (reference dask_ml/metrics/pairwise.py:70-105):
//   distances = -2 * dot(X, Y.T) + XX + YY ; maximum(distances, 0) ; sqrt unless squared.
// 64x64 output tile per CTA, 4x4 per thread, features chunked by 16 through shared memory.
// ---------------------------------------------------------------------------------------------
constexpr int CD_T = 64, CD_KC = 16;

// epilogue modes of the tile kernel: the distance matrix itself, or one of dask-ml's pairwise kernels
// (reference dask_ml/metrics/pairwise.py:135-191) applied to the same X.Y^T tile
enum CdMode { CD_DIST = 0, CD_SQDIST = 1, CD_RBF = 2, CD_LINEAR = 3, CD_POLY = 4, CD_SIGMOID = 5 };

template <typename T>
__global__ void __launch_bounds__(256)
cdist_kernel(const T *__restrict__ X, int64_t n, int d, int64_t ldx, const T *__restrict__ Y, int k,
             const T *__restrict__ xnorm2, const T *__restrict__ ynorm2, T *__restrict__ out, int mode,
             T gamma, T coef0, int degree) {
    __shared__ T As[CD_KC][CD_T + 4];
    __shared__ T Bs[CD_KC][CD_T + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int64_t row0 = (int64_t)blockIdx.y * CD_T;
    const int col0 = blockIdx.x * CD_T;
    T acc[4][4], xx[4], yy[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        xx[i] = (T)0; yy[i] = (T)0;
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = (T)0;
    }
    for (int f0 = 0; f0 < d; f0 += CD_KC) {
        for (int e = tid; e < CD_T * CD_KC; e += 256) {
            const int m = e / CD_KC, f = e % CD_KC;
            T a = (T)0, b = (T)0;
            if (f0 + f < d) {
                if (row0 + m < n) a = X[(row0 + m) * ldx + f0 + f];
                if (col0 + m < k) b = Y[(size_t)(col0 + m) * d + f0 + f];
            }
            As[f][m] = a;
            Bs[f][m] = b;
        }
        __syncthreads();
#pragma unroll
        for (int f = 0; f < CD_KC; ++f) {
            T a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { a[i] = As[f][ty * 4 + i]; b[i] = Bs[f][tx * 4 + i]; }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                xx[i] = fma(a[i], a[i], xx[i]);
                yy[i] = fma(b[i], b[i], yy[i]);
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int64_t row = row0 + ty * 4 + i;
        if (row >= n) continue;
        const T xn = xnorm2 ? xnorm2[row] : xx[i];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int col = col0 + tx * 4 + j;
            if (col >= k) continue;
            const T yn = ynorm2 ? ynorm2[col] : yy[j];
            T v;
            if (mode <= CD_RBF) {
                v = ((T)(-2) * acc[i][j] + xn) + yn;   // pairwise.py:101
                v = v > (T)0 ? v : (T)0;               // :102
                if (mode == CD_DIST) v = sqrt(v);
                else if (mode == CD_RBF) v = exp(-gamma * v);   // rbf_kernel :143-151
            } else if (mode == CD_LINEAR) {
                v = acc[i][j];                          // linear_kernel :135-138
            } else if (mode == CD_POLY) {
                const T b = gamma * acc[i][j] + coef0;  // polynomial_kernel :154-167
                v = (T)1;
                for (int e = 0; e < degree; ++e) v *= b;
            } else {
                v = tanh(gamma * acc[i][j] + coef0);    // sigmoid_kernel :170-185
            }
            out[row * (int64_t)k + col] = v;
        }
    }
}

template <typename T>
static int cdist_impl(const T *X, int64_t n, int d, int64_t ldx, const T *Y, int k, const T *xnorm2, const T *ynorm2,
                      T *out, int mode, T gamma, T coef0, int degree, void *stream) {
    KM_CHECK_ARG(X && Y && out && n > 0 && d > 0 && k > 0 && ldx >= d, B200KM_EINVAL);
    const int64_t gy = (n + CD_T - 1) / CD_T;
    KM_CHECK_ARG(gy <= 65535LL * 65535LL, B200KM_ETOOBIG);
    // grid.y is limited to 65535: fold the row tiles into calls of at most 65535 tiles
    cudaStream_t st = km_stream(stream);
    for (int64_t t0 = 0; t0 < gy; t0 += 65535) {
        const int64_t nt = (gy - t0) < 65535 ? (gy - t0) : 65535;
        const int64_t r0 = t0 * CD_T;
        const int64_t nn = (n - r0) < nt * CD_T ? (n - r0) : nt * CD_T;
        dim3 grid((k + CD_T - 1) / CD_T, (unsigned)nt);
        cdist_kernel<T><<<grid, 256, 0, st>>>(X + r0 * ldx, nn, d, ldx, Y, k, xnorm2 ? xnorm2 + r0 : nullptr, ynorm2,
                                              out + r0 * (int64_t)k, mode, gamma, coef0, degree);
    }
    KM_LAUNCH_RET();
}

extern "C" int b200km_cdist_f32(const float *X, int64_t n, int d, int64_t ldx, const float *Y, int k,
                                const float *xnorm2, const float *ynorm2, float *out, int squared, void *stream) {
    return cdist_impl<float>(X, n, d, ldx, Y, k, xnorm2, ynorm2, out, squared ? CD_SQDIST : CD_DIST, 0.f, 0.f, 0, stream);
}
extern "C" int b200km_cdist_f64(const double *X, int64_t n, int d, int64_t ldx, const double *Y, int k,
                                const double *xnorm2, const double *ynorm2, double *out, int squared, void *stream) {
    return cdist_impl<double>(X, n, d, ldx, Y, k, xnorm2, ynorm2, out, squared ? CD_SQDIST : CD_DIST, 0.0, 0.0, 0, stream);
}

// K7 on the tensor cores (cdist_tc.cu): float32 rows of up to 64 features against a centre pack
int km_launch_cdist_tc(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k, float *out, int64_t ldo,
                       int mode, float gamma, cudaStream_t st, double *cost, int init);
extern "C" int b200km_cdist_pack_f32(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int k, float *out,
                                     int64_t ldo, int mode, double gamma, void *stream) {
    KM_CHECK_ARG(X && pack && out && n > 0 && d > 0 && k > 0 && ldx >= d && ldo >= k, B200KM_EINVAL);
    KM_CHECK_ARG(mode >= B200KM_CDIST_SQUARED && mode <= B200KM_CDIST_RBF, B200KM_EINVAL);
    KM_CHECK_ARG(k <= B200KM_MAX_K, B200KM_ETOOBIG);
    KM_CHECK_ARG((reinterpret_cast<uintptr_t>(pack) & 255) == 0, B200KM_EALIGN);
    return km_launch_cdist_tc(X, n, d, ldx, pack, k, out, ldo, mode, (float)gamma, km_stream(stream), nullptr, 0);
}

// K4 against up to 128 new candidates at once, on the same kernel (mode 3: min over the candidates, merged into the running
// minimum, nothing else stored)
extern "C" int b200km_mindist_tc_f32(const float *X, int64_t n, int d, int64_t ldx, const void *pack, int m, float *mind2,
                                     int init, double *cost_out, void *stream) {
    KM_CHECK_ARG(X && pack && mind2 && cost_out && n > 0 && d > 0 && m > 0 && ldx >= d, B200KM_EINVAL);
    KM_CHECK_ARG(m <= 128 && d <= 128, B200KM_ETOOBIG);
    KM_CHECK_ARG((reinterpret_cast<uintptr_t>(pack) & 255) == 0, B200KM_EALIGN);
    return km_launch_cdist_tc(X, n, d, ldx, pack, m, mind2, 0, 3, 0.f, km_stream(stream), cost_out, init);
}

// pairwise kernels on the same tile kernel (the SpectralClustering tail)
extern "C" int b200km_pairwise_kernel_f32(const float *X, int64_t n, int d, int64_t ldx, const float *Y, int k,
                                          float *out, int kind, double gamma, double coef0, int degree, void *stream) {
    KM_CHECK_ARG(kind >= B200KM_KERNEL_RBF && kind <= B200KM_KERNEL_SIGMOID && degree >= 0, B200KM_EINVAL);
    return cdist_impl<float>(X, n, d, ldx, Y, k, nullptr, nullptr, out, kind, (float)gamma, (float)coef0, degree, stream);
}
extern "C" int b200km_pairwise_kernel_f64(const double *X, int64_t n, int d, int64_t ldx, const double *Y, int k,
                                          double *out, int kind, double gamma, double coef0, int degree, void *stream) {
    KM_CHECK_ARG(kind >= B200KM_KERNEL_RBF && kind <= B200KM_KERNEL_SIGMOID && degree >= 0, B200KM_EINVAL);
    return cdist_impl<double>(X, n, d, ldx, Y, k, nullptr, nullptr, out, kind, gamma, coef0, degree, stream);
}

// ---------------------------------------------------------------------------------------------
// row norms / finite check (warp per row, lanes = features)
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
row_norms_kernel(const T *__restrict__ X, int64_t n, int d, int64_t ldx, T *__restrict__ out, int squared) {
    const int lane = threadIdx.x & 31;
    const int64_t w0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = w0; r < n; r += nw) {
        T s = (T)0;
        for (int f = lane; f < d; f += 32) { const T v = X[r * ldx + f]; s = fma(v, v, s); }
        s = km_warp_sum(s);
        if (lane == 0) out[r] = squared ? s : sqrt(s);
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
nonfinite_kernel(const T *__restrict__ X, int64_t n, int d, int64_t ldx, int32_t *__restrict__ bad) {
    const int64_t total = n * (int64_t)d;
    int found = 0;
    if (ldx == d) {
        for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x)
            found |= !isfinite(X[e]);
    } else {
        for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
            const int64_t r = e / d;
            found |= !isfinite(X[r * ldx + (e - r * d)]);
        }
    }
    if (__any_sync(0xffffffffu, found) && (threadIdx.x & 31) == 0) atomicOr(bad, 1);
}

template <typename T>
static int row_norms_impl(const T *X, int64_t n, int d, int64_t ldx, T *out, int squared, void *stream) {
    KM_CHECK_ARG(X && out && n > 0 && d > 0 && ldx >= d, B200KM_EINVAL);
    int64_t blocks = (n + 7) / 8;
    if (blocks > km_num_sms() * 16) blocks = km_num_sms() * 16;
    row_norms_kernel<T><<<(unsigned)blocks, 256, 0, km_stream(stream)>>>(X, n, d, ldx, out, squared);
    KM_LAUNCH_RET();
}
template <typename T>
static int nonfinite_impl(const T *X, int64_t n, int d, int64_t ldx, int32_t *bad, void *stream) {
    KM_CHECK_ARG(X && bad && n > 0 && d > 0 && ldx >= d, B200KM_EINVAL);
    int64_t blocks = (n * (int64_t)d + 255) / 256;
    if (blocks > km_num_sms() * 16) blocks = km_num_sms() * 16;
    nonfinite_kernel<T><<<(unsigned)blocks, 256, 0, km_stream(stream)>>>(X, n, d, ldx, bad);
    KM_LAUNCH_RET();
}

extern "C" int b200km_row_norms_f32(const float *X, int64_t n, int d, int64_t ldx, float *out, int squared, void *s) {
    return row_norms_impl<float>(X, n, d, ldx, out, squared, s);
}
extern "C" int b200km_row_norms_f64(const double *X, int64_t n, int d, int64_t ldx, double *out, int squared, void *s) {
    return row_norms_impl<double>(X, n, d, ldx, out, squared, s);
}
extern "C" int b200km_count_nonfinite_f32(const float *X, int64_t n, int d, int64_t ldx, int32_t *bad, void *s) {
    return nonfinite_impl<float>(X, n, d, ldx, bad, s);
}
extern "C" int b200km_count_nonfinite_f64(const double *X, int64_t n, int d, int64_t ldx, int32_t *bad, void *s) {
    return nonfinite_impl<double>(X, n, d, ldx, bad, s);
}

// ---------------------------------------------------------------------------------------------
// blobs on the device (benchmark input).  Distribution of dask_ml.datasets.make_blobs
// (reference dask_ml/datasets.py:78-205): isotropic Gaussians around k_true centres, rows in
// random class order; NOT its bit stream.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
blobs_kernel(T *__restrict__ X, int64_t n, int d, int64_t ldx, const T *__restrict__ centers, int k_true, T cluster_std,
             uint32_t k0, uint32_t k1, int64_t row_offset, int32_t *__restrict__ y) {
    const int dq = (d + 3) / 4;
    const int64_t total = n * (int64_t)dq;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / dq;
        const int q = (int)(e - r * dq);
        const uint64_t g = (uint64_t)(row_offset + r);
        const KmPhilox4 cw = km_philox4x32_10((uint32_t)g, (uint32_t)(g >> 32), 0u, 0x626c6f62u, k0, k1);
        const int cls = (int)(cw.x % (uint32_t)k_true);
        if (y && q == 0) y[r] = cls;
        const KmPhilox4 w = km_philox4x32_10((uint32_t)g, (uint32_t)(g >> 32), (uint32_t)q + 1u, 0x626c6f62u, k0, k1);
        // two Box-Muller pairs
        const float u0 = ((float)(w.x >> 8) + 0.5f) * (1.0f / 16777216.0f);
        const float u1 = ((float)(w.y >> 8) + 0.5f) * (1.0f / 16777216.0f);
        const float u2 = ((float)(w.z >> 8) + 0.5f) * (1.0f / 16777216.0f);
        const float u3 = ((float)(w.w >> 8) + 0.5f) * (1.0f / 16777216.0f);
        const float ra = sqrtf(-2.f * logf(u0)), rb = sqrtf(-2.f * logf(u2));
        float sa, ca, sb, cb;
        sincospif(2.f * u1, &sa, &ca);
        sincospif(2.f * u3, &sb, &cb);
        const float z[4] = {ra * ca, ra * sa, rb * cb, rb * sb};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int f = q * 4 + i;
            if (f < d) X[r * ldx + f] = centers[(size_t)cls * d + f] + cluster_std * (T)z[i];
        }
    }
}

template <typename T>
static int blobs_impl(T *X, int64_t n, int d, int64_t ldx, const T *centers, int k_true, T cluster_std, uint64_t seed,
                      int64_t row_offset, int32_t *y, void *stream) {
    KM_CHECK_ARG(X && centers && n > 0 && d > 0 && k_true > 0 && ldx >= d, B200KM_EINVAL);
    int64_t blocks = (n * (int64_t)((d + 3) / 4) + 255) / 256;
    if (blocks > km_num_sms() * 32) blocks = km_num_sms() * 32;
    blobs_kernel<T><<<(unsigned)blocks, 256, 0, km_stream(stream)>>>(X, n, d, ldx, centers, k_true, cluster_std,
                                                                     (uint32_t)seed, (uint32_t)(seed >> 32), row_offset, y);
    KM_LAUNCH_RET();
}
extern "C" int b200km_make_blobs_f32(float *X, int64_t n, int d, int64_t ldx, const float *centers, int k_true,
                                     float cluster_std, uint64_t seed, int64_t row_offset, int32_t *y, void *stream) {
    return blobs_impl<float>(X, n, d, ldx, centers, k_true, cluster_std, seed, row_offset, y, stream);
}
extern "C" int b200km_make_blobs_f64(double *X, int64_t n, int d, int64_t ldx, const double *centers, int k_true,
                                     double cluster_std, uint64_t seed, int64_t row_offset, int32_t *y, void *stream) {
    return blobs_impl<double>(X, n, d, ldx, centers, k_true, cluster_std, seed, row_offset, y, stream);
}

// ---------------------------------------------------------------------------------------------
// pipe probes: 8 independent FMA chains per thread
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(1024) probe_fma_kernel(T *sink, int iters) {
    T a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = (T)(threadIdx.x + i) * (T)1e-3;
    const T b = (T)1.0000001, c = (T)1e-7;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fma(a[i], b, c);
        }
    }
    T s = (T)0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i];
    if (s == (T)123456789) sink[0] = s;  // never true; keeps the chains alive
}
extern "C" int b200km_probe_ffma(float *sink, int blocks, int threads, int iters, void *stream) {
    KM_CHECK_ARG(sink && blocks > 0 && threads > 0 && threads <= 1024 && iters > 0, B200KM_EINVAL);
    probe_fma_kernel<float><<<blocks, threads, 0, km_stream(stream)>>>(sink, iters);
    KM_LAUNCH_RET();
}
extern "C" int b200km_probe_dfma(double *sink, int blocks, int threads, int iters, void *stream) {
    KM_CHECK_ARG(sink && blocks > 0 && threads > 0 && threads <= 1024 && iters > 0, B200KM_EINVAL);
    probe_fma_kernel<double><<<blocks, threads, 0, km_stream(stream)>>>(sink, iters);
    KM_LAUNCH_RET();
}

// ---------------------------------------------------------------------------------------------
extern "C" int b200km_version(void) { return B200KM_VERSION; }

extern "C" const char *b200km_error_string(int code) {
    switch (code) {
        case B200KM_OK: return "ok";
        case B200KM_EINVAL: return "b200km: invalid argument (null pointer, non-positive size or ldx < d)";
        case B200KM_ETOOBIG: return "b200km: size beyond the supported maximum";
        case B200KM_EALIGN: return "b200km: pointer or leading dimension not aligned as documented";
        case B200KM_ECAPACITY: return "b200km: output list capacity exceeded";
        case B200KM_ESTATICROWS:
            return "b200km_assign_static_f32: more rows than the built-in scratch holds (B200KM_STATIC_MAX_ROWS = "
                   "1048576); call it per row slab or use b200km_assign_f32 with your own scratch";
        case B200KM_ESTATICPACK:
            return "b200km_assign_static_f32: k and d need a centre pack larger than the built-in "
                   "B200KM_STATIC_MAX_PACK_BYTES (2 MiB); use b200km_pack_centers + b200km_assign_f32";
        default: break;
    }
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "b200km: unknown error code";
}
