// tc_common.cuh -- tcgen05 / TMEM / mbarrier helpers shared by the tensor-core kernels (assign_tc.cu: K1;
// cdist_tc.cu: K7).  Raw PTX, sm_100a only.
#pragma once
#include "common.cuh"

namespace {

constexpr int TC_THREADS = 320;   // producer warp + MMA warp + 8 staging/epilogue warps
constexpr int TC_EPI = 256;       // staging/epilogue threads: 2 per row (column halves)
constexpr int TC_ROWS = 128;     // rows per CTA step
constexpr int TC_BSTAGES_MAX = 8;   // ring of K-step blocks: 5 x 16 KB (one CTA) or 8 x 8 KB (CTA pair)
constexpr int TC_ASTAGES = 2;    // accumulator stages: 2 x 256 = 512 TMEM columns
constexpr unsigned long long TC_WAIT_LIMIT_NS = 30ull * 1000000000ull;   // see km_wait_expired

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(km_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ unsigned long long km_globaltimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// Have 30 s of wall time (%globaltimer) gone by since *t0?  Called every 65536 failed polls of a spin loop; the first
// call only starts the clock.  A bound in TIME, not in polls: a wait that is merely slow -- time-slicing with another
// context, MPS co-tenants, a debugger or compute-sanitizer -- must not be taken for a hang.
__device__ __forceinline__ bool km_wait_expired(unsigned long long &t0) {
    const unsigned long long now = km_globaltimer();
    if (t0 == 0) { t0 = now; return false; }
    return now - t0 > TC_WAIT_LIMIT_NS;
}
// bounded wait: a protocol bug traps (launch error) instead of hanging the GPU
__device__ __forceinline__ void mbar_wait_bounded(uint64_t *bar, uint32_t parity) {
    const uint32_t addr = km_smem_u32(bar);
    uint32_t done;
    uint32_t spins = 0;
    unsigned long long t0 = 0;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
        if (!done && ((++spins) & 0xFFFFu) == 0 && km_wait_expired(t0)) __trap();
    } while (!done);
}
// non-blocking probe of a barrier phase
__device__ __forceinline__ bool mbar_test(uint64_t *bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(km_smem_u32(bar)), "r"(parity)
        : "memory");
    return done != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// Shared-memory matrix descriptors: no-swizzle ("interleave") K-major canonical layout, 8-row x 16-byte
// core matrices; LBO = byte distance between the two 16-byte K chunks of one MMA (bits 16-29, >>4),
// SBO = byte distance between 8-row groups (bits 32-45, >>4), version 1 at bits 46-47.
// descriptors are passed as their low words; the high word (SBO = 128 B, version 1) is shared
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint32_t adesc_lo, uint32_t bdesc_lo, uint32_t desc_hi,
                                         uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %4, {%6, %6, %6, %6}, p;\n\t}"
        ::"r"(tmem_d), "r"(adesc_lo), "r"(bdesc_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}
// the same MMA spanning a CTA pair (cta_group::2): M = 256 (128 rows from each CTA), N split likewise
__device__ __forceinline__ void mma_tf32_pair(uint32_t tmem_d, uint32_t adesc_lo, uint32_t bdesc_lo, uint32_t desc_hi,
                                              uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], da, db, %4, {%6, %6, %6, %6, %6, %6, %6, %6}, p;\n\t}"
        ::"r"(tmem_d), "r"(adesc_lo), "r"(bdesc_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}
__device__ __forceinline__ void mma_commit_pair(uint64_t *bar) {   // arrives on `bar` in BOTH CTAs of the pair
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
            km_smem_u32(bar)),
        "h"((unsigned short)3)
        : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t *bar, uint32_t rank) {
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(km_smem_u32(bar)),
        "r"(rank)
        : "memory");
}
// the same without release semantics (no cluster-scope fence): for arrivals that only pass on an event whose
// ordering is already established elsewhere (TMA completion observed on a local barrier; TMEM reads ordered
// by tcgen05.fence)
__device__ __forceinline__ void mbar_arrive_remote_relaxed(uint64_t *bar, uint32_t rank) {
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(km_smem_u32(bar)),
        "r"(rank)
        : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\t"
        "elect.sync rx|px, 0xFFFFFFFF;\n\t"
        "selp.u32 %0, 1, 0, px;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void mma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(km_smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ float tf32_rna(float x) {
    uint32_t u;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
    return __uint_as_float(u);
}

// 5 ALU ops per score: FSETP + 3 FMNMX + SEL(imm).  `jloc` is the index inside the current chunk.
__device__ __forceinline__ void track(float v, int j, float &best, float &second, int &jloc) {
    asm("{\n\t.reg .pred p;\n\t.reg .f32 t;\n\t"
        "setp.lt.f32 p, %3, %0;\n\t"
        "max.f32 t, %0, %3;\n\t"
        "min.f32 %1, %1, t;\n\t"
        "min.f32 %0, %0, %3;\n\t"
        "@p mov.b32 %2, %4;\n\t}"
        : "+f"(best), "+f"(second), "+r"(jloc)
        : "f"(v), "r"(j));
}

// ---- threshold tracker (small-k kernel): what the (best, second, argmin) tracker above costs in ALU-pipe
// instructions (FSETP / FMNMX / SEL: one warp instruction per 2 cycles) is moved to the FMA pipe.
//   pass 1 over a 32-score chunk: its minimum with three-input min (FMNMX3: 0.5 ALU op per score);
//   pass 2: u = sat((T - v) * BIG) is 1 for a score within the row's near-tie threshold T = min + bound, else 0;
//           cnt += u;  idx += u * j   -- three FMA-pipe instructions per score, no compare, no select.
// A row whose count is exactly 1 has ONE score within the threshold: it is the minimum and idx is its column.
// Any other count sends the row to the float64 re-decision, exactly as `second - best <= bound` did (the threshold
// is padded upward, so the test can only flag more rows, never fewer), and its provisional label comes from an
// exact rescan (lowest column among equal minima) that a warp runs only when one of its rows needs it.
__device__ __forceinline__ float min3f(float a, float b, float c) {
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float fma_sat(float a, float b, float c) {
    float r;
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}

}  // namespace
