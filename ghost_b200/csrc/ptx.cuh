// Thin inline-PTX wrappers for sm_100a: mbarrier, TMA (cp.async.bulk.tensor),
// tcgen05 (MMA / TMEM alloc / TMEM load / commit) and the UMMA descriptors.
// Everything here is Blackwell-only; there is no fallback path.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace ghost {
namespace ptx {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, P;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// all state spaces: plain global stores of this thread become visible to later TMA (async proxy) reads
__device__ __forceinline__ void fence_proxy_async_all() {
  asm volatile("fence.proxy.async;" ::: "memory");
}
// coherent 16-byte load (L2 only): data written earlier in the SAME kernel (conv chains), where ld.global.nc is not allowed
__device__ __forceinline__ uint4 ld_global_cg16(const void* p) {
  uint4 v;
  asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// CTA-pair kernels: shared::cluster address of `bar` in the even (leader) CTA of the pair - bit 24 of a
// shared::cluster address is the CTA rank inside a cluster of two
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;
__device__ __forceinline__ void mbar_arrive_rank0(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(smem_u32(bar) & kPeerBitMask) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, P;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

// Bounded wait: a pipeline bug must never hang the GPU.  On timeout the
// caller-visible abort word in shared memory is set and false is returned so
// every role can unwind to the teardown code.
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, volatile int* abort_flag) {
  if (mbar_try_wait(bar, parity)) return true;
  for (uint32_t spin = 0; spin < (1u << 22); ++spin) {
    if (mbar_try_wait(bar, parity)) return true;
    if ((spin & 63u) == 63u) {
      if (*abort_flag) return false;
      __nanosleep(64);
    }
  }
  *abort_flag = 1;
  return false;
}

// Same bound, but never changes the caller's control flow: after a timeout (or once another role has flagged one)
// it simply returns and the caller carries on with stale data; the kernel then finishes and reports the error
// word.  Keeping the issue loops free of data-dependent exits lets ptxas hold their state in uniform registers.
__device__ __forceinline__ void mbar_wait_nofail(uint64_t* bar, uint32_t parity, volatile int* abort_flag) {
  if (mbar_try_wait(bar, parity)) return;
  for (uint32_t spin = 0; spin < (1u << 22); ++spin) {
    if (mbar_try_wait(bar, parity)) return;
    if ((spin & 63u) == 63u) {
      if (*abort_flag) return;
      __nanosleep(64);
    }
  }
  *abort_flag = 1;
}

// ---------------------------------------------------------------- programmatic dependent launch
// wait: blocks until the preceding grid in the stream has completed and its memory is visible (no-op when the kernel
// was launched without a programmatic dependency).  launch_dependents: lets the next grid start its prologue early.
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---------------------------------------------------------------- TMA
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2];" ::"r"(smem_u32(smem)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_5d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                            int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, "
      "%7}], [%2];" ::"r"(smem_u32(smem)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}

// TMA store of one shared-memory box to global memory (bulk async-group completion).
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, const void* smem, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// wait until at most N of this thread's bulk groups still READ their shared-memory source
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// named barrier among `nthreads` threads (id 1..15; 0 is __syncthreads)
__device__ __forceinline__ void named_bar_arrive(int id, int nthreads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---------------------------------------------------------------- tcgen05
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// Whole-warp call.  ncols: power of two in [32, 512].
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// D[tmem] (+)= A[smem] * B[smem]; bf16 inputs, fp32 accumulate. One thread issues.
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrives on the mbarrier once every previously issued tcgen05.mma has retired.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// 32 lanes x 16 consecutive fp32 columns; thread i of the warp owns TMEM lane (base_lane + i).
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, "
      "%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
}
// 16 lanes x 8 consecutive fp32 columns: thread t gets columns 2(t%4), 2(t%4)+1 of lane base + t/4 (v[0], v[1]) and of
// lane base + t/4 + 8 (v[2], v[3]) - the layout of an m16n8 accumulator fragment (checked by tools/tmem_layout.cu).
__device__ __forceinline__ void tmem_ld_16x256b(uint32_t taddr, uint32_t (&v)[4]) {
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---------------------------------------------------------------- packed fp32 (Blackwell FFMA2 / FADD2)
// Two IEEE fp32 operations per instruction on adjacent register pairs; results are bit-identical to the scalar
// fmaf / + / - on each half.  Halves the issue slots of the fp32-heavy fused epilogues.
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%6, %7};\n\t"
      "fma.rn.f32x2 rd, ra, rb, rc;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
  return d;
}
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
      "add.rn.f32x2 rd, ra, rb;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
      "mul.rn.f32x2 rd, ra, rb;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 sub2(float2 a, float2 b) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
      "sub.rn.f32x2 rd, ra, rb;\n\t"
      "mov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}

// Shared-memory matrix descriptor, K-major operand, 128-byte swizzle, rows of
// 64 bf16 (=128 B); 8-row groups are `sbo_bytes` apart (1024 for a dense tile).
__device__ __forceinline__ uint64_t make_smem_desc_sw128(uint32_t smem_addr, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);        // start address  [0,14)
  d |= static_cast<uint64_t>(1) << 16;                            // LBO (unused for swizzled K-major)
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;    // SBO            [32,46)
  d |= static_cast<uint64_t>(1) << 46;                            // descriptor version (sm_100)
  d |= static_cast<uint64_t>(2) << 61;                            // SWIZZLE_128B
  return d;
}

// A view that starts `row_shift` (0..7) rows into a 1024-byte swizzle atom: the MMA unit derives the 128B-swizzle
// phase from the shared-memory address bits, exactly as the TMA unit did when it wrote the tile, so shifting the start
// address by whole 128-byte rows is all it takes (measured on B200: the descriptor base-offset field must stay 0).
__device__ __forceinline__ uint64_t make_smem_desc_sw128_shifted(uint32_t smem_addr, uint32_t sbo_bytes,
                                                                 uint32_t row_shift) {
  return make_smem_desc_sw128(smem_addr + row_shift * 128u, sbo_bytes);
}

// Split form of the same descriptor for the MMA issue loop: the low word carries the start address (advanced by
// 2 per 16-element K step, by 8 per 128-byte row), the high word (SBO, version, swizzle mode) is loop invariant.
__device__ __forceinline__ uint32_t smem_desc_lo(uint32_t smem_addr) {
  return ((smem_addr & 0x3FFFFu) >> 4) | (1u << 16);
}
__device__ __forceinline__ uint32_t smem_desc_hi(uint32_t sbo_bytes, bool swizzle64) {
  return ((sbo_bytes >> 4) & 0x3FFFu) | (1u << 14) | ((swizzle64 ? 4u : 2u) << 29);
}

// NK consecutive 16-element K steps of one operand pair, issued back to back by one elected lane of a CONVERGED
// warp (all 32 lanes must call this with identical arguments).  Keeping the batch in one asm block lets ptxas hold
// the descriptors in uniform registers: the issue loop costs a few instructions per MMA instead of ~30.
template <int NK>
__device__ __forceinline__ void umma_bf16_batch(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t hi,
                                                uint32_t idesc, uint32_t accumulate_first) {
  static_assert(NK == 2 || NK == 4, "K steps per batch");
  if (NK == 4) {
    asm volatile(
        "{\n\t.reg .pred E, Q;\n\t.reg .b64 da, db;\n\t.reg .b32 a, b;\n\t"
        "elect.sync _|E, 0xffffffff;\n\t"
        "setp.ne.b32 Q, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
        "@E tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, Q;\n\t"
        "add.u32 a, %1, 2;\n\tadd.u32 b, %2, 2;\n\tmov.b64 da, {a, %3};\n\tmov.b64 db, {b, %3};\n\t"
        "@E tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, 1;\n\t"
        "add.u32 a, %1, 4;\n\tadd.u32 b, %2, 4;\n\tmov.b64 da, {a, %3};\n\tmov.b64 db, {b, %3};\n\t"
        "@E tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, 1;\n\t"
        "add.u32 a, %1, 6;\n\tadd.u32 b, %2, 6;\n\tmov.b64 da, {a, %3};\n\tmov.b64 db, {b, %3};\n\t"
        "@E tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, 1;\n\t}\n" ::"r"(d_tmem),
        "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accumulate_first)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred E, Q;\n\t.reg .b64 da, db;\n\t.reg .b32 a, b;\n\t"
        "elect.sync _|E, 0xffffffff;\n\t"
        "setp.ne.b32 Q, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
        "@E tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, Q;\n\t"
        "add.u32 a, %1, 2;\n\tadd.u32 b, %2, 2;\n\tmov.b64 da, {a, %3};\n\tmov.b64 db, {b, %3};\n\t"
        "@E tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, 1;\n\t}\n" ::"r"(d_tmem),
        "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accumulate_first)
        : "memory");
  }
}
// commit by one elected lane of a converged warp
__device__ __forceinline__ void umma_commit_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred E;\n\t"
      "elect.sync _|E, 0xffffffff;\n\t"
      "@E tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}\n" ::"r"(smem_u32(bar))
      : "memory");
}

// ---------------------------------------------------------------- CTA pairs (cta_group::2)
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// TMA loads issued by either CTA of a pair into ITS OWN shared memory; the transaction bytes are counted on the
// barrier at the same offset in the leader CTA (rank 0), where the one MMA issuer of the pair waits.
__device__ __forceinline__ void tma2_load_2d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], "
      "[%2];" ::"r"(smem_u32(smem)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma2_load_4d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                             int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, "
      "%5, %6}], [%2];" ::"r"(smem_u32(smem)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma2_load_5d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                             int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, "
      "%5, %6, %7}], [%2];" ::"r"(smem_u32(smem)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3),
      "r"(c4)
      : "memory");
}
// Whole-warp calls, by the warp with the same index in BOTH CTAs of the pair.
__device__ __forceinline__ void tmem2_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem2_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// Four K steps of a 256 x N tile: A rows 0-127 / B rows 0-N/2 from the leader's shared memory, rows 128-255 / N/2-N
// from the same offsets in the peer's; accumulator rows 128-255 land in the peer's TMEM.  Leader CTA only.
__device__ __forceinline__ void umma2_bf16_batch4(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t hi,
                                                  uint32_t idesc, uint32_t accumulate_first) {
  asm volatile(
      "{\n\t.reg .pred E, Q;\n\t.reg .b64 da, db;\n\t.reg .b32 a, b;\n\t"
      "elect.sync _|E, 0xffffffff;\n\t"
      "setp.ne.b32 Q, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
      "@E tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, Q;\n\t"
      "add.u32 a, %1, 2;\n\tadd.u32 b, %2, 2;\n\tmov.b64 da, {a, %3};\n\tmov.b64 db, {b, %3};\n\t"
      "@E tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, 1;\n\t"
      "add.u32 a, %1, 4;\n\tadd.u32 b, %2, 4;\n\tmov.b64 da, {a, %3};\n\tmov.b64 db, {b, %3};\n\t"
      "@E tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, 1;\n\t"
      "add.u32 a, %1, 6;\n\tadd.u32 b, %2, 6;\n\tmov.b64 da, {a, %3};\n\tmov.b64 db, {b, %3};\n\t"
      "@E tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, 1;\n\t}\n" ::"r"(d_tmem),
      "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accumulate_first)
      : "memory");
}
// arrives on the barrier at this offset in BOTH CTAs once every previously issued MMA of the pair has retired
__device__ __forceinline__ void umma2_commit_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred E;\n\t.reg .b16 m;\n\t"
      "mov.b16 m, 3;\n\t"
      "elect.sync _|E, 0xffffffff;\n\t"
      "@E tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n\t}\n" ::"r"(
          smem_u32(bar))
      : "memory");
}

// Instruction descriptor for kind::f16: bf16 x bf16 -> fp32 or fp16 x fp16 -> fp32, both operands K-major.
__host__ __device__ inline uint32_t make_idesc_f16(uint32_t M, uint32_t N, bool fp16) {
  uint32_t d = 0;
  d |= 1u << 4;            // D format: f32
  if (!fp16) {
    d |= 1u << 7;          // A format: 1 = bf16 (0 = fp16)
    d |= 1u << 10;         // B format: 1 = bf16 (0 = fp16)
  }
  d |= (N >> 3) << 17;     // N / 8
  d |= (M >> 4) << 24;     // M / 16
  return d;
}
__host__ __device__ inline uint32_t make_idesc_bf16(uint32_t M, uint32_t N) { return make_idesc_f16(M, N, false); }

}  // namespace ptx
}  // namespace ghost
