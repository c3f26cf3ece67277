// common.cuh -- shared device helpers for libb200km (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "../../include/b200km.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libb200km is written for sm_100a (B200) only"
#endif

#define KM_CHECK_ARG(cond, code) \
    do {                         \
        if (!(cond)) return (code); \
    } while (0)

#define KM_LAUNCH_RET()                       \
    do {                                      \
        cudaError_t e__ = cudaGetLastError(); \
        return (int)e__;                      \
    } while (0)

static inline cudaStream_t km_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr int KM_MAX_DEVICES = 64;
// SM count of the current device (148 on B200: 2 dies x 74 SMs), read once per device from the driver; grids of the
// persistent kernels are sized from it.
static inline int km_num_sms() {
    static int cached[KM_MAX_DEVICES];   // zero-initialised; a racing first call writes the same value twice
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= KM_MAX_DEVICES) return 148;
    int v = cached[dev];
    if (v <= 0) {
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        cached[dev] = v;
    }
    return v;
}

// Opt a kernel into > 48 KB of dynamic shared memory, once per device per kernel.
// (cudaFuncSetAttribute is per device; `flags` is a per-kernel static bool[KM_MAX_DEVICES].)
template <typename Kern>
static inline int km_set_smem_once(Kern kern, int bytes, bool *flags) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= KM_MAX_DEVICES) dev = KM_MAX_DEVICES - 1, flags[dev] = false;
    if (!flags[dev]) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        if (e != cudaSuccess) return (int)e;
        flags[dev] = true;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Centre pack layout (see b200km_pack_centers).  All sections 256-byte aligned.
// ---------------------------------------------------------------------------------------------
constexpr int KM_TILE_N = 128;  // centroid tile width of the tiled assign kernel

// tcgen05 (3xTF32) image of the centres: per 256-centroid column tile a sequence of blocks in the
// UMMA no-swizzle K-major operand layout ([16-byte K chunk][256 rows][16 B]):
//   block 0      : aug[2 chunks]                  8192 B  (the K-step that adds ||c||^2)
//   block 1 + ks : hi[2 chunks] | lo[2 chunks]   16384 B  (K-step ks = 8 features of -2c, TF32 hi / lo parts)
constexpr int KM_TC_BN = 256;
constexpr int KM_TC_MAX_D = 128;        // a 128-row tile of X (hi + lo parts) stays resident in shared memory up to here
constexpr int KM_TC_WIDE_MAX_D = 2048;  // beyond it K is cut into panels (assign_tc_wide_kernel); == B200KM_MAX_D
constexpr int KM_TC_AUG_BYTES = 2 * KM_TC_BN * 16;
constexpr int KM_TC_KS_BYTES = 4 * KM_TC_BN * 16;
__host__ __device__ inline int km_tc_kpad(int d) { return (d + 7) & ~7; }
__host__ __device__ inline size_t km_tc_tile_bytes(int d) {
    return (size_t)KM_TC_AUG_BYTES + (size_t)(km_tc_kpad(d) / 8) * KM_TC_KS_BYTES;
}

// ---------------------------------------------------------------------------------------------
// Scratch layout of the pruned float64 re-decision inside the flag buffer F (L int32 entries, L >= n):
//   [0, M)               rows K1 flagged (cnt of them), M = (L - 1) / 9 = the most the pruned path takes
//   [M, L - 1 - M)       per flagged row: number of listed columns, then `cap` columns
//   [L - 1 - M, L - 1)   rows with more than `cap` listed columns (they get the full scan)
//   [L - 1]              size of that overflow list
// `cap` is decided on the device from cnt: whatever the list area allows, up to KM_CAND_CAP_MAX -- at least
// KM_CAND_CAP_MIN (cnt = M), all 32 when few rows are flagged or the caller hands over a buffer longer than n.
// ---------------------------------------------------------------------------------------------
constexpr int KM_CAND_CAP_MIN = 6;
constexpr int KM_CAND_CAP_MAX = 32;
__host__ __device__ inline int64_t km_cand_rows_max(int64_t L) { return (L - 1) / (KM_CAND_CAP_MIN + 3); }
__host__ __device__ inline int km_cand_cap(int64_t L, int64_t cnt) {
    if (cnt <= 0) return KM_CAND_CAP_MAX;
    const int64_t cap = (L - 1 - 2 * km_cand_rows_max(L)) / cnt - 1;
    return (int)(cap > KM_CAND_CAP_MAX ? KM_CAND_CAP_MAX : cap);
}

struct KmPackLayout {
    size_t off_hdr;   // int32[4]: k, d, ntiles, magic
    size_t off_cmax;  // uint32: bits of max_j ||c32_j||^2 (fp32, non-negative)
    size_t off_mu;    // float[kpad(d)]    mu = mean of the centres (fp32, zero padded): the float32 kernels work on
                      //                   x - mu and c - mu, so their rounding bounds scale with the spread of the data
                      //                   around the centres, not with its distance from the origin
    size_t off_zero;  // float[kpad(d)]    zeros: stands in for mu when the caller's rows are already centred
                      //                   (float64 X converted by b200km_center_convert_f64)
    size_t off_c32;   // float[k*d]        row-major fp32 centres, CENTRED: fl32(c64 - mu)
    size_t off_cn32;  // float[ntiles*128] ||c32_j||^2 (centred), +inf padded
    size_t off_cn64;  // double[k]         ||c64_j||^2
    size_t off_c64;   // double[k*d]       copy of the fp64 master
    size_t off_ct;    // float[ntiles][d+1][128] column-tiled centred fp32 centres (zero padded); row d = norms
    size_t off_tc;    // tcgen05 image: ceil(k/256) column tiles of km_tc_tile_bytes(d) each (absent when d > KM_TC_WIDE_MAX_D)
    size_t total;
};

__host__ __device__ inline size_t km_align256(size_t x) { return (x + 255) & ~size_t(255); }

__host__ __device__ inline KmPackLayout km_pack_layout(int k, int d) {
    KmPackLayout L;
    const size_t ntiles = (size_t)(k + KM_TILE_N - 1) / KM_TILE_N;
    size_t o = 0;
    L.off_hdr = o;  o = km_align256(o + 16);
    L.off_cmax = o; o = km_align256(o + 4);
    L.off_mu = o;   o = km_align256(o + sizeof(float) * (size_t)((d + 7) & ~7));
    L.off_zero = o; o = km_align256(o + sizeof(float) * (size_t)((d + 7) & ~7));
    L.off_c32 = o;  o = km_align256(o + sizeof(float) * (size_t)k * d);
    L.off_cn32 = o; o = km_align256(o + sizeof(float) * ntiles * KM_TILE_N);
    L.off_cn64 = o; o = km_align256(o + sizeof(double) * (size_t)k);
    L.off_c64 = o;  o = km_align256(o + sizeof(double) * (size_t)k * d);
    L.off_ct = o;   o = km_align256(o + sizeof(float) * ntiles * (size_t)(d + 1) * KM_TILE_N);
    L.off_tc = o;
    if (d <= KM_TC_WIDE_MAX_D) o = km_align256(o + (size_t)((k + KM_TC_BN - 1) / KM_TC_BN) * km_tc_tile_bytes(d));
    L.total = o;
    return L;
}

// ---------------------------------------------------------------------------------------------
// warp helpers
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T km_warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// block-wide sum of one double per thread -> one atomicAdd.  `red` = >= 32 doubles of smem.
__device__ __forceinline__ void km_block_atomic_add(double v, double *dst, double *red) {
    v = km_warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) red[w] = v;
    __syncthreads();
    if (w == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        double t = lane < nw ? red[lane] : 0.0;
        t = km_warp_sum(t);
        if (lane == 0 && t != 0.0) atomicAdd(dst, t);
    }
}

// ---------------------------------------------------------------------------------------------
// mbarrier + bulk async copy (TMA 1-D; SASS: UBLKCP / SYNCS)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t km_smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void km_mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(km_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void km_fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void km_fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void km_mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(km_smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void km_bulk_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            km_smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(km_smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void km_mbar_wait(uint64_t *bar, uint32_t parity) {
    const uint32_t addr = km_smem_u32(bar);
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
    } while (!done);
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011).  The oracle carries the identical numpy restatement.
// ---------------------------------------------------------------------------------------------
struct KmPhilox4 {
    uint32_t x, y, z, w;
};
__device__ __forceinline__ KmPhilox4 km_philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                      uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return KmPhilox4{c0, c1, c2, c3};
}
// 53-bit uniform in [0,1) from two 32-bit words (numpy random_sample construction)
__device__ __forceinline__ double km_u53(uint32_t a, uint32_t b) {
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) * (1.0 / 9007199254740992.0);
}
