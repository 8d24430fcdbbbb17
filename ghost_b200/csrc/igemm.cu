// tcgen05 / TMEM implicit-GEMM convolution for sm_100a.
//
// GEMM view of a convolution over NHWC bf16 activations:
//   M = output pixels (tiles of bn x bh x bw <= 128 rows), N = Cout, K = taps * Cin.
// The A operand of every (tap, 64-channel chunk) is one TMA box of the input
// tensor, fetched at the tap-shifted coordinate: out-of-bounds elements are
// zero-filled by the TMA unit, which is exactly the conv padding, so no im2col
// buffer ever exists.  Stride-2 convs use a 5-D parity-split view of the same
// NHWC memory, ConvTranspose k4s2 is 4 output-phase launches with 2x2 taps.
//
// Roles (640 threads): warp 0 = TMA producer of the activation operand (+ AAD h tiles, resident weights), warp 3 =
// TMA producer of the weight operand, warps 1-2 = MMA issuers (even / odd tiles; warp 1 also owns the TMEM
// allocation), warps 4..19 = epilogue (TMEM -> regs -> fused epilogue -> global or TMA store).  Producer and
// issuer warps run converged and elect one lane per copy / MMA batch.  The accumulator is multi-buffered in TMEM
// so the epilogue of tile i overlaps the main loop of the next tiles.  Persistent CTAs, static round-robin tile
// schedule.
//
// Reference ops covered: nn.Conv2d / nn.ConvTranspose2d / nn.Linear call sites in
// network/AEI_Net.py:19-41,101 and network/AADLayer.py:13-18,67-74; the AAD
// epilogue implements AADLayer.forward (network/AADLayer.py:20-38) + the ReLU
// that always follows it (network/AADLayer.py:62,70).
#include <cstdlib>
#include <mutex>

#include <cuda_fp16.h>

#include "igemm.cuh"
#include "ptx.cuh"

namespace ghost {

using namespace ptx;

// Stall accounting for tools/igemm_prof.py (make PROF=1): cycles each role spends waiting on its barriers.
#ifdef GHOST_IGEMM_PROF
__device__ unsigned long long g_igemm_prof[160][12];
__device__ unsigned long long g_aad_prof[160][8];  // AAD epilogue, first warp: waits h / pair / acc, blend, store, total, tiles
// launch timeline (tools/igemm_timeline.py): first CTA entry / last CTA exit of every igemm launch, by %globaltimer
constexpr int kTlSlots = 8192;
__device__ unsigned long long g_tl_t0[kTlSlots], g_tl_t1[kTlSlots];
__device__ int g_tl_meta[kTlSlots][12];
__device__ long long g_tl_cta0[8];  // CTA 0 of the running launch: entry clock, prologue end, issuer end, epilogue end
__device__ unsigned int g_tl_seq, g_tl_done;
#define PROF_STAMP(i) do { if (blockIdx.x == 0 && blockIdx.y == 0 && ew == 0 && lg == 0 && lane == 0) g_tl_cta0[i] = clock64(); } while (0)
#define PROF_BEGIN() long long pw0_ = 0, pw1_ = 0; const long long pt0_ = clock64()
#define PROF_WAIT0(stmt) do { const long long a_ = clock64(); stmt; pw0_ += clock64() - a_; } while (0)
#define PROF_WAIT1(stmt) do { const long long a_ = clock64(); stmt; pw1_ += clock64() - a_; } while (0)
#define PROF_END(slot, lane_) do { if ((lane_) == 0 && blockIdx.x < 160 && blockIdx.y == 0) { \
    g_igemm_prof[blockIdx.x][slot] = pw0_; g_igemm_prof[blockIdx.x][slot + 1] = pw1_; \
    g_igemm_prof[blockIdx.x][slot + 2] = clock64() - pt0_; } } while (0)
#else
#define PROF_STAMP(i) do {} while (0)
#define PROF_BEGIN() do {} while (0)
#define PROF_WAIT0(stmt) stmt
#define PROF_WAIT1(stmt) stmt
#define PROF_END(slot, lane_) do {} while (0)
#endif

namespace {

struct TileCoord {
  int tn, th, tw, nt;
};

__device__ __forceinline__ TileCoord decode_tile(const IgemmParams& p, int tile) {
  TileCoord t;
  t.nt = tile % p.tiles_n;
  int m = tile / p.tiles_n;
  t.tw = m % p.tiles_w;
  m /= p.tiles_w;
  t.th = m % p.tiles_h;
  t.tn = m / p.tiles_h;
  return t;
}

// Walks the persistent CTA's tile sequence (blockIdx.x, + gridDim.x, ...) without per-tile divisions: the stride is
// decomposed once into mixed-radix digits and added with carries.
struct TileWalk {
  TileCoord t, d;
  __device__ __forceinline__ void init(const IgemmParams& p, int first, int stride) {
    t = decode_tile(p, first);
    d = decode_tile(p, stride);  // d.tn may exceed its radix: it is only ever added
  }
  __device__ __forceinline__ void next(const IgemmParams& p) {
    t.nt += d.nt;
    int c = t.nt >= p.tiles_n;
    t.nt -= c ? p.tiles_n : 0;
    t.tw += d.tw + c;
    c = t.tw >= p.tiles_w;
    t.tw -= c ? p.tiles_w : 0;
    t.th += d.th + c;
    c = t.th >= p.tiles_h;
    t.th -= c ? p.tiles_h : 0;
    t.tn += d.tn + c;
  }
};

// 16-bit storage element: bf16 (F16 = false) or IEEE fp16 (F16 = true, the reference's own inference dtype,
// inference.py:30 `G.half()`); the arithmetic is fp32 either way, tcgen05 kind::f16 takes both at the same rate.
template <bool F16>
__device__ __forceinline__ float lo16(uint32_t v) {
  if (F16) return __half2float(__ushort_as_half(static_cast<unsigned short>(v & 0xffffu)));
  return __uint_as_float(v << 16);
}
template <bool F16>
__device__ __forceinline__ float hi16(uint32_t v) {
  if (F16) return __half2float(__ushort_as_half(static_cast<unsigned short>(v >> 16)));
  return __uint_as_float(v & 0xffff0000u);
}
template <bool F16>
__device__ __forceinline__ float ld16(const __nv_bfloat16* p) {
  if (F16) return __half2float(*reinterpret_cast<const __half*>(p));
  return __bfloat162float(*p);
}
template <bool F16>
__device__ __forceinline__ void st16(__nv_bfloat16* p, float v) {
  if (F16) *reinterpret_cast<__half*>(p) = __float2half_rn(v);
  else *p = __float2bfloat16(v);
}
// the value as it reads back after a 16-bit store
template <bool F16>
__device__ __forceinline__ float round16(float v) {
  if (F16) return __half2float(__float2half_rn(v));
  return __bfloat162float(__float2bfloat16(v));
}
// lo -> bits [0,16), hi -> bits [16,32)
template <bool F16>
__device__ __forceinline__ uint32_t pack16(float lo, float hi) {
  uint32_t d;
  if (F16) asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  else asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
// max(., 0) and the pack in one instruction
template <bool F16>
__device__ __forceinline__ uint32_t pack16_relu(float lo, float hi) {
  uint32_t d;
  if (F16) asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  else asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
// SPADE variant: max(v, v*slope) * post (slope 0 = ReLU, 0.2 = LeakyReLU(0.2), 1 = identity)
template <bool F16>
__device__ __forceinline__ uint32_t pack16_slope(float lo, float hi, float slope, float post) {
  lo = fmaxf(lo, lo * slope) * post;
  hi = fmaxf(hi, hi * slope) * post;
  return pack16<F16>(lo, hi);
}


// acc + bias, first activation.  `sp` != nullptr: per-channel parameters of this CTA's N tile staged in shared
// memory ([bias | slope | scale2 | shift2] x 256 floats, column j of the tile at sp[j]).
// brow (optional): bias row of this pixel's border class (igemm.cuh bias9), used instead of p.bias
__device__ __forceinline__ void linear_bias_act_16(const IgemmParams& p, float (&f)[16], int col0, const float* sp,
                                                   int j, const float* brow = nullptr) {
          if (brow) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float4 b = __ldg(reinterpret_cast<const float4*>(brow + col0) + i);
              f[4 * i] += b.x, f[4 * i + 1] += b.y, f[4 * i + 2] += b.z, f[4 * i + 3] += b.w;
            }
          } else if (p.bias) {
            if (sp) {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float4 b = *reinterpret_cast<const float4*>(sp + j + 4 * i);
                f[4 * i] += b.x, f[4 * i + 1] += b.y, f[4 * i + 2] += b.z, f[4 * i + 3] += b.w;
              }
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i) f[i] += __ldg(p.bias + col0 + i);
            }
          }
          if (p.act1 == kActLeaky) {
#pragma unroll
            for (int i = 0; i < 16; ++i) f[i] = f[i] > 0.f ? f[i] : f[i] * p.slope;
          } else if (p.act1 == kActPRelu) {
            if (sp) {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float4 a = *reinterpret_cast<const float4*>(sp + 256 + j + 4 * i);
                f[4 * i] = f[4 * i] > 0.f ? f[4 * i] : f[4 * i] * a.x;
                f[4 * i + 1] = f[4 * i + 1] > 0.f ? f[4 * i + 1] : f[4 * i + 1] * a.y;
                f[4 * i + 2] = f[4 * i + 2] > 0.f ? f[4 * i + 2] : f[4 * i + 2] * a.z;
                f[4 * i + 3] = f[4 * i + 3] > 0.f ? f[4 * i + 3] : f[4 * i + 3] * a.w;
              }
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i) f[i] = f[i] > 0.f ? f[i] : f[i] * __ldg(p.slope_c + col0 + i);
            }
          } else if (p.act1 == kActRelu) {
#pragma unroll
            for (int i = 0; i < 16; ++i) f[i] = fmaxf(f[i], 0.f);
          }
}

// v -> final values of 16 full columns: bias, first activation, residual, second activation (no stores)
template <bool F16>
__device__ __forceinline__ void linear_values_16(const IgemmParams& p, const uint32_t (&v)[16], int col0, long long r_pix,
                                                 const float* sp, int j, float (&f)[16], const float* brow = nullptr) {
#pragma unroll
  for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
  linear_bias_act_16(p, f, col0, sp, j, brow);
  if (p.res) {
    const __nv_bfloat16* rp = p.res + r_pix + col0;
    const uint4 q0 = __ldg(reinterpret_cast<const uint4*>(rp));
    const uint4 q1 = __ldg(reinterpret_cast<const uint4*>(rp) + 1);
    const uint32_t qq[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      f[2 * i] += lo16<F16>(qq[i]);
      f[2 * i + 1] += hi16<F16>(qq[i]);
    }
  }
  if (p.act2 == kAct2Tanh) {
#pragma unroll
    for (int i = 0; i < 16; ++i) f[i] = tanhf(f[i]);
  }
}

// FINAL: called by the split-K reduce kernel with the summed accumulators (p.splits > 1 there too)
// stored (optional): receives the 16 values as they read back from the output tensor (for the fused statistics)
template <bool F16, bool FINAL = false>
__device__ __forceinline__ void linear_epilogue_16(const IgemmParams& p, const uint32_t (&v)[16], int col0,
                                                   long long o_pix, long long r_pix, const float* sp, int j,
                                                   float* stored = nullptr, const float* brow = nullptr) {
          float f[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
          if (!FINAL && p.splits > 1) {
            // partial tile of this K range: o_pix is the row offset inside the workspace
            float4* o = reinterpret_cast<float4*>(p.ws + o_pix + col0);
#pragma unroll
            for (int i = 0; i < 4; ++i) o[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
            return;
          }
          linear_bias_act_16(p, f, col0, sp, j, brow);
          const bool full16 = (col0 + 16 <= p.n_valid);
          if (p.res) {
            const __nv_bfloat16* rp = p.res + r_pix + col0;
            if (full16) {
              // (conv chains read residuals written earlier in the same launch: no non-coherent loads there)
              const uint4 q0 = p.chain_n ? ld_global_cg16(rp) : __ldg(reinterpret_cast<const uint4*>(rp));
              const uint4 q1 = p.chain_n ? ld_global_cg16(rp + 8) : __ldg(reinterpret_cast<const uint4*>(rp) + 1);
              const uint32_t qq[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                f[2 * i] += lo16<F16>(qq[i]);
                f[2 * i + 1] += hi16<F16>(qq[i]);
              }
            } else {
              for (int i = 0; i < 16; ++i)
                if (col0 + i < p.n_valid) f[i] += ld16<F16>(rp + i);
            }
          }
          if (p.act2 == kAct2Tanh) {
#pragma unroll
            for (int i = 0; i < 16; ++i) f[i] = tanhf(f[i]);
          }
          if (stored) {
#pragma unroll
            for (int i = 0; i < 16; ++i) stored[i] = (col0 + i < p.n_valid) ? (p.out_f32 ? f[i] : round16<F16>(f[i])) : 0.f;
          }
          if (p.out_f32) {
            float* o = reinterpret_cast<float*>(p.out) + o_pix + col0;
            if (full16) {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                reinterpret_cast<float4*>(o)[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
            } else {
              for (int i = 0; i < 16; ++i)
                if (col0 + i < p.n_valid) o[i] = f[i];
            }
          } else {
            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + o_pix + col0;
            if (full16 && p.out2) {
              // second output from the value as stored (rounded to bf16), e.g. the next block's bn1
              __nv_bfloat16* o2 = reinterpret_cast<__nv_bfloat16*>(p.out2) + o_pix + col0;
              float t[16];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float4 sc = sp ? *reinterpret_cast<const float4*>(sp + 512 + j + 4 * i)
                                     : __ldg(reinterpret_cast<const float4*>(p.scale2 + col0) + i);
                const float4 sh = sp ? *reinterpret_cast<const float4*>(sp + 768 + j + 4 * i)
                                     : __ldg(reinterpret_cast<const float4*>(p.shift2 + col0) + i);
                t[4 * i] = fmaf(round16<F16>(f[4 * i]), sc.x, sh.x);
                t[4 * i + 1] = fmaf(round16<F16>(f[4 * i + 1]), sc.y, sh.y);
                t[4 * i + 2] = fmaf(round16<F16>(f[4 * i + 2]), sc.z, sh.z);
                t[4 * i + 3] = fmaf(round16<F16>(f[4 * i + 3]), sc.w, sh.w);
              }
              uint4 s0, s1;
              s0.x = pack16<F16>(t[0], t[1]);
              s0.y = pack16<F16>(t[2], t[3]);
              s0.z = pack16<F16>(t[4], t[5]);
              s0.w = pack16<F16>(t[6], t[7]);
              s1.x = pack16<F16>(t[8], t[9]);
              s1.y = pack16<F16>(t[10], t[11]);
              s1.z = pack16<F16>(t[12], t[13]);
              s1.w = pack16<F16>(t[14], t[15]);
              reinterpret_cast<uint4*>(o2)[0] = s0;
              reinterpret_cast<uint4*>(o2)[1] = s1;
            }
            if (full16) {
              uint4 s0, s1;
              s0.x = pack16<F16>(f[0], f[1]);
              s0.y = pack16<F16>(f[2], f[3]);
              s0.z = pack16<F16>(f[4], f[5]);
              s0.w = pack16<F16>(f[6], f[7]);
              s1.x = pack16<F16>(f[8], f[9]);
              s1.y = pack16<F16>(f[10], f[11]);
              s1.z = pack16<F16>(f[12], f[13]);
              s1.w = pack16<F16>(f[14], f[15]);
              reinterpret_cast<uint4*>(o)[0] = s0;
              reinterpret_cast<uint4*>(o)[1] = s1;
            } else {
              for (int i = 0; i < 16; ++i)
                if (col0 + i < p.n_valid) st16<F16>(o + i, f[i]);
            }
          }
}

// Linear-mode epilogue of one accumulator tile, shared by the one-CTA kernel (igemm_kernel) and the CTA-pair kernel
// (igemm2_kernel).  All 16 epilogue warps work on every tile: four warps per TMEM lane group split the columns.
// Staged output (use_stg) for the LAST tile of a CTA, whose epilogue nothing overlaps: once that tile's accumulator
// is complete the operand ring is idle, so every warp parks its 32 rows x wcols there (rows of P 16-byte pieces,
// piece q of row r at slot q ^ stg_key(r): bank-conflict free) and moves whole row segments to / from global
// memory - 4-8 cache lines per instruction instead of 32.
template <bool F16, bool STATS>
struct LinEpi {
  int lane, lg, ew, r, rn, rh, rw, rows_used, box_rows;
  int wcols, rowb, P, lgP, rb_row0, rb_piece, rb_step;
  uint8_t* wst;
  const float* eparm;
  bool staged;
  uint32_t tmem_lanes;  // TMEM address of this warp's lane group, column 0
  int acc_stride;
  // fused output statistics (igemm.cuh st_mode): per 16-column chunk of this warp, lane L holds the sum (L even) or
  // the sum of squares (L odd) of column (L >> 1) & 15 over the rows seen so far
  float st_acc[4];
  float st_tmp[16];
  int st_img, st_nt;
  float* st_x;  // shared exchange [4 ew][4 lg][4 chunks][32 lanes]
  uint8_t* tst_buf;  // TMA-store epilogue: [2 buffers][4 lane groups][BN/64 boxes][32 rows][128 B]
  int tst_it;        // tiles this warp has sent through the staging buffers
  uint8_t* ring0;    // whole-tile TMA epilogue (tte): staging = the idle operand ring
  uint64_t* rbar;    // tte: [4] one barrier per 64-column group for the residual box
  uint32_t rphase;

  __device__ __forceinline__ void init(const IgemmParams& p, int warp, int lane_, uint8_t* ring, const float* eparm_,
                                       bool staged_, int rows_used_, uint32_t tmem_base, int acc_stride_,
                                       uint64_t* rbar_) {
    ring0 = ring;
    rbar = rbar_;
    rphase = 0;
    lane = lane_;
    lg = warp & 3;                        // TMEM lane group this warp may read
    ew = (warp - kIgemmEpiWarp0) >> 2;    // 0..3: which of the four warps of that lane group
    r = lg * 32 + lane;                   // GEMM row inside the tile
    const int pitch = p.a_win == 2 ? p.bw + 2 : p.bw;  // tile-window mode: rows of bw + 2 GEMM rows, the last two junk
    rn = r / (p.bh * pitch);
    rh = (r / pitch) % p.bh;
    rw = r % pitch;
    rows_used = (rw < p.bw) ? rows_used_ : 0;  // (a junk row is never valid)
    box_rows = rows_used_;
    wcols = p.BN >> 2, rowb = wcols * 2, P = rowb >> 4, lgP = 31 - __clz(P);
    wst = ring + (warp - kIgemmEpiWarp0) * (32 * rowb);
    rb_row0 = lane >> lgP, rb_piece = lane & (P - 1), rb_step = 32 >> lgP;
    eparm = eparm_;
    staged = staged_;
    tmem_lanes = tmem_base + (static_cast<uint32_t>(lg * 32) << 16);
    acc_stride = acc_stride_;
    st_acc[0] = st_acc[1] = st_acc[2] = st_acc[3] = 0.f;
    st_img = -1;
    st_nt = 0;
    st_x = reinterpret_cast<float*>(ring + p.stages * st_stage_bytes(p));
    tst_buf = ring + p.stages * st_stage_bytes(p);  // (never together with the statistics exchange)
    tst_it = 0;
  }
  __device__ __forceinline__ static int st_stage_bytes(const IgemmParams& p) {
    if (p.cta2) {
      const int b_half = (p.BN >> 1) * 128;
      if (p.a_win == 2) return p.a_slot;
      return p.a_win ? 17 * 1024 + (p.b_res ? 0 : 3 * b_half) : 128 * 128 + (p.b_res ? 0 : b_half);
    }
    const int b_tile = p.BN * p.kc * 2;
    return p.a_win ? 17 * 1024 + (p.b_res ? 0 : 3 * b_tile) : 128 * p.kc * 2 + (p.b_res ? 0 : b_tile);
  }
  // Column sums over the 32 rows of this warp: a = values, b = their squares (16 columns per lane).  Exchanges with
  // lane ^ 16, 8, 4, 2 halve the columns a lane carries and double the rows they cover; the last exchange leaves the
  // sum in the even lane and the sum of squares in the odd lane of a pair.  31 shuffles, fixed order.
  __device__ __forceinline__ float st_colsum(float (&a)[16], float (&b)[16]) const {
#pragma unroll
    for (int w = 8; w >= 1; w >>= 1) {
      const bool hi = (lane & (2 * w)) != 0;
#pragma unroll
      for (int i = 0; i < w; ++i) {
        const float sa = hi ? a[i] : a[i + w], ka = hi ? a[i + w] : a[i];
        const float sb = hi ? b[i] : b[i + w], kb = hi ? b[i + w] : b[i];
        a[i] = ka + __shfl_xor_sync(0xffffffffu, sa, 2 * w);
        b[i] = kb + __shfl_xor_sync(0xffffffffu, sb, 2 * w);
      }
    }
    const bool odd = (lane & 1) != 0;
    const float send = odd ? a[0] : b[0], keep = odd ? b[0] : a[0];
    return keep + __shfl_xor_sync(0xffffffffu, send, 1);
  }
  __device__ __forceinline__ void st_zero() {
#pragma unroll
    for (int i = 0; i < 16; ++i) st_tmp[i] = 0.f;
  }
  __device__ __forceinline__ void st_keep(const float (&f)[16]) {
#pragma unroll
    for (int i = 0; i < 16; ++i) st_tmp[i] = f[i];
  }
  __device__ __forceinline__ void st_add(int c, float (&stored)[16]) {
    float sq[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) sq[i] = stored[i] * stored[i];
    st_acc[c] += st_colsum(stored, sq);
  }
  // the four lane-group warps of this column quarter add up in lane-group order and write one part
  __device__ __forceinline__ void st_flush(const IgemmParams& p, int part) {
    const int nchunk = wcols >> 4;
    float* mine = st_x + ((ew * 4 + lg) * 4) * 32 + lane;
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (c < nchunk) mine[c * 32] = st_acc[c];
    named_bar_sync(1 + ew, 128);
    if (lg == 0) {
      const float* q = st_x + (ew * 4 * 4) * 32 + lane;
      float* dst = p.st_parts + (static_cast<long long>(st_img) * (2 * p.st_P + 1) + 2 * part + (lane & 1)) * p.st_C +
                   st_nt * p.BN + ew * wcols + ((lane >> 1) & 15);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (c < nchunk) {
          const float v = ((q[c * 32] + q[(4 + c) * 32]) + q[(8 + c) * 32]) + q[(12 + c) * 32];
          if (st_nt * p.BN + ew * wcols + c * 16 < p.n_valid) dst[c * 16] = v;
        }
      }
    }
    named_bar_sync(1 + ew, 128);
    st_acc[0] = st_acc[1] = st_acc[2] = st_acc[3] = 0.f;
  }
  // call after the last tile of the CTA
  __device__ __forceinline__ void st_finish(const IgemmParams& p) {
    if (STATS && p.st_mode == 2 && st_img >= 0) st_flush(p, blockIdx.x);
    if (p.tst && ew == 0 && lane == 0) bulk_wait<0>();  // outstanding TMA stores
    tte_drain(p);
  }
  // whole-tile TMA epilogue: the lanes that issued stores wait until they are complete (and the staging is free)
  __device__ __forceinline__ void tte_drain(const IgemmParams& p) {
    if (p.tte && lg == 0 && lane == 0) bulk_wait<0>();
  }
  __device__ __forceinline__ int stg_key(int row) const { return ((row * rowb) >> 7) & (P - 1); }

  // REMOTE: the accumulator-empty barrier lives in the leader CTA of the pair (cluster rank 0)
  template <bool REMOTE>
  __device__ __forceinline__ void release(uint64_t* tempty) const {
    if (lane == 0) {
      if (REMOTE) mbar_arrive_rank0(tempty);
      else mbar_arrive(tempty);
    }
  }

  // returns false when the pipeline was aborted (the caller leaves its tile loop)
  // gp: where the tensor maps of this launch / chain step live (kernel parameter or global memory, never shared)
  template <bool REMOTE>
  __device__ __forceinline__ bool tile(const IgemmParams& p, const IgemmParams* gp, const TileCoord& t, int as,
                                       uint32_t aphase, bool use_stg, uint64_t* tfull, uint64_t* tempty,
                                       volatile int* abort_flag) {
        if (STATS) {  // bn == 1: the tile lies in image t.tn
          if (p.st_mode == 2 && st_img >= 0 && st_img != t.tn) st_flush(p, blockIdx.x);
          st_img = t.tn;
          st_nt = t.nt;
        }
        const int n = t.tn * p.bn + rn, oh = t.th * p.bh + rh, ow = t.tw * p.bw + rw;
        const bool valid = (r < rows_used) && (n < p.Nimg) && (oh < p.Ho) && (ow < p.Wo);
        // bias per border class of the output pixel (igemm.cuh bias9): rows {first, inner, last} x columns likewise
        const float* brow =
            p.bias9 ? p.bias9 + ((oh == 0 ? 0 : (oh >= p.Ho - 1 ? 2 : 1)) * 3 + (ow == 0 ? 0 : (ow >= p.Wo - 1 ? 2 : 1))) * p.bias9_ld
                    : nullptr;
        const uint32_t t_addr = tmem_lanes + as * acc_stride;
        const int ph = p.nphase > 1 ? t.nt / p.tpn : 0;
        const long long o_pix =
            p.splits > 1
                ? ((static_cast<long long>(blockIdx.y) * p.tiles_nb * p.tiles_h * p.tiles_w +
                    (static_cast<long long>(t.tn) * p.tiles_h + t.th) * p.tiles_w + t.tw) * 128 + r) * p.ws_cols
                : n * p.o_sn + oh * p.o_sh + ow * p.o_sw + p.o_off + (p.nphase > 1 ? p.ph_o[ph] : 0);
        const long long r_pix = n * p.r_sn + oh * p.r_sh + ow * p.r_sw + p.r_off + (p.nphase > 1 ? p.ph_r[ph] : 0);
        const int nch = p.BN >> 4, ch0 = (nch * ew) >> 2, ch1 = (nch * (ew + 1)) >> 2;
        const int colw = t.nt * p.BN + ew * wcols;  // staged path: first column of this warp
        const bool cols_on = colw < p.n_valid;      // (n_valid is a multiple of wcols there)
        if (!__all_sync(0xffffffffu, mbar_wait(tfull, aphase, abort_flag))) return false;
        tc_fence_after();
        if (p.tst) {
          // this lane group's piece of the tile: hb boxes of [32 rows][128 B], buffer tst_it & 1
          const int hb = p.BN >> 6;
          uint8_t* piece = tst_buf + (((tst_it & 1) * 4 + lg) * hb) * 4096;
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            if (c * 16 < wcols) {  // warp-uniform
              const int j = ew * wcols + c * 16;
              uint32_t v[16];
              __syncwarp();
              tmem_ld16(t_addr + j, v);
              tmem_ld_wait();
              uint4 o0 = make_uint4(0u, 0u, 0u, 0u), o1 = o0;
              if (valid) {
                float f[16];
                linear_values_16<F16>(p, v, t.nt * p.BN + j, r_pix, staged ? eparm : nullptr, j, f, brow);
                o0.x = pack16<F16>(f[0], f[1]);
                o0.y = pack16<F16>(f[2], f[3]);
                o0.z = pack16<F16>(f[4], f[5]);
                o0.w = pack16<F16>(f[6], f[7]);
                o1.x = pack16<F16>(f[8], f[9]);
                o1.y = pack16<F16>(f[10], f[11]);
                o1.z = pack16<F16>(f[12], f[13]);
                o1.w = pack16<F16>(f[14], f[15]);
              }
              uint8_t* row = piece + (j >> 6) * 4096 + lane * 128;
              const int cc = (j & 63) >> 3;  // 16-byte chunk inside the 128-byte row
              *reinterpret_cast<uint4*>(row + ((cc ^ (lane & 7)) << 4)) = o0;
              *reinterpret_cast<uint4*>(row + (((cc + 1) ^ (lane & 7)) << 4)) = o1;
            }
          }
          // the accumulator stage goes back to the MMA warp; the piece goes out through the TMA
          tc_fence_before();
          fence_proxy_async();
          __syncwarp();
          release<REMOTE>(tempty);
          // the store of the previous tile (other buffer) has read its shared memory before anyone refills it
          if (ew == 0 && lane == 0) bulk_wait_read<0>();
          named_bar_sync(5 + lg, 128);
          if (ew == 0 && lane == 0) {
            if (r < rows_used && n < p.Nimg && oh < p.Ho && ow < p.Wo) {
              for (int b = 0; b < hb; ++b)
                tma_store_4d(&gp->tmO, piece + b * 4096, p.tst_c0 + t.nt * p.BN + b * 64, ow, oh, n);
            }
            bulk_commit();
          }
          ++tst_it;
          return true;
        }
        if (use_stg && p.tte) {
          // Whole-tile TMA epilogue of the last tile: the residual box of every 64-column group arrives by TMA in the
          // (idle) operand ring, laid out as the store box [rows][128 B, 128B swizzle]; every lane adds its row in place
          // and the group leaves through ONE TMA store (a second one for out2) - nothing goes through the LSU path
          // (measured on a 14x14x256 iresnet conv: residual rows 3800 cycles, out + out2 stores 6000 cycles of a 12600
          // cycle epilogue that no main loop hides).
          PROF_STAMP(5);  // accumulator complete
          const int g = (ew * wcols) >> 6, cg = (ew * wcols) & 63;  // 64-column group of this warp, first column in it
          const int ngrp = p.BN >> 6;
          uint8_t* Rg = ring0 + g * 16384;
          uint8_t* R2g = Rg + ngrp * 16384;
          const int w0 = t.tw * p.bw, h0 = t.th * p.bh, n0 = t.tn * p.bn;
          const bool leader = lg == 0 && cg == 0 && lane == 0;
          const bool gcols_on = t.nt * p.BN + g * 64 < p.n_valid;  // (n_valid is a multiple of 64 here)
          if (p.res != nullptr && leader && gcols_on) {
            mbar_expect_tx(&rbar[g], static_cast<uint32_t>(box_rows) * 128u);
            tma_load_4d(Rg, &gp->tmH, &rbar[g], p.tte_r0 + t.nt * p.BN + g * 64, w0, h0, n0);
          }
          uint8_t* my_row = Rg + r * 128;
          const int key = r & 7;
          bool tte_ok = true;
          // (not unrolled: this code runs once per launch, straight from a cold instruction cache - a quarter of the
          // instructions to fetch; pipelining the TMEM loads over two register buffers measured no gain)
          const int nchunk = wcols >> 4;
#pragma unroll 1
          for (int c = 0; c < nchunk; ++c) {
            {
              const int j = ew * wcols + c * 16;
              uint32_t v[16];
              __syncwarp();
              tmem_ld16(t_addr + j, v);
              tmem_ld_wait();
              if (c == 0 && p.res != nullptr && gcols_on) {
                if (!__all_sync(0xffffffffu, mbar_wait(&rbar[g], rphase, abort_flag))) tte_ok = false;  // (aborting)
                PROF_STAMP(6);  // residual box landed
              }
              if (valid && gcols_on) {
                float f[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
                linear_bias_act_16(p, f, t.nt * p.BN + j, staged ? eparm : nullptr, j, brow);
                const int ch = (cg >> 3) + 2 * c;  // 16-byte chunk inside the 128-byte row
                uint4* s0 = reinterpret_cast<uint4*>(my_row + ((ch ^ key) << 4));
                uint4* s1 = reinterpret_cast<uint4*>(my_row + (((ch + 1) ^ key) << 4));
                if (p.res) {
                  const uint4 q0 = *s0, q1 = *s1;
                  const uint32_t qq[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
                  for (int i = 0; i < 8; ++i) {
                    f[2 * i] += lo16<F16>(qq[i]);
                    f[2 * i + 1] += hi16<F16>(qq[i]);
                  }
                }
                uint4 o0, o1;
                o0.x = pack16<F16>(f[0], f[1]);
                o0.y = pack16<F16>(f[2], f[3]);
                o0.z = pack16<F16>(f[4], f[5]);
                o0.w = pack16<F16>(f[6], f[7]);
                o1.x = pack16<F16>(f[8], f[9]);
                o1.y = pack16<F16>(f[10], f[11]);
                o1.z = pack16<F16>(f[12], f[13]);
                o1.w = pack16<F16>(f[14], f[15]);
                *s0 = o0;
                *s1 = o1;
                if (p.out2) {
                  // second output from the value as stored (rounded to 16 bits), e.g. the next block's bn1
                  const uint32_t ow_[8] = {o0.x, o0.y, o0.z, o0.w, o1.x, o1.y, o1.z, o1.w};
                  float t2[16];
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    const float4 sc = staged ? *reinterpret_cast<const float4*>(eparm + 512 + j + 4 * i)
                                             : __ldg(reinterpret_cast<const float4*>(p.scale2 + t.nt * p.BN + j) + i);
                    const float4 sh = staged ? *reinterpret_cast<const float4*>(eparm + 768 + j + 4 * i)
                                             : __ldg(reinterpret_cast<const float4*>(p.shift2 + t.nt * p.BN + j) + i);
                    t2[4 * i] = fmaf(lo16<F16>(ow_[2 * i]), sc.x, sh.x);
                    t2[4 * i + 1] = fmaf(hi16<F16>(ow_[2 * i]), sc.y, sh.y);
                    t2[4 * i + 2] = fmaf(lo16<F16>(ow_[2 * i + 1]), sc.z, sh.z);
                    t2[4 * i + 3] = fmaf(hi16<F16>(ow_[2 * i + 1]), sc.w, sh.w);
                  }
                  uint4 u0, u1;
                  u0.x = pack16<F16>(t2[0], t2[1]);
                  u0.y = pack16<F16>(t2[2], t2[3]);
                  u0.z = pack16<F16>(t2[4], t2[5]);
                  u0.w = pack16<F16>(t2[6], t2[7]);
                  u1.x = pack16<F16>(t2[8], t2[9]);
                  u1.y = pack16<F16>(t2[10], t2[11]);
                  u1.z = pack16<F16>(t2[12], t2[13]);
                  u1.w = pack16<F16>(t2[14], t2[15]);
                  *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(s0) + ngrp * 16384) = u0;
                  *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(s1) + ngrp * 16384) = u1;
                }
              }
            }
          }
          if (p.res != nullptr && gcols_on) rphase ^= 1u;
          tc_fence_before();
          fence_proxy_async();  // the staged rows -> visible to the TMA store
          __syncwarp();
          PROF_STAMP(7);  // tile converted into the staging rows
          release<REMOTE>(tempty);
          named_bar_sync(1 + g, (256 / wcols) * 32);  // the warps of this 64-column group
          if (leader && gcols_on) {
            tma_store_4d(&gp->tmO, Rg, p.tst_c0 + t.nt * p.BN + g * 64, w0, h0, n0);
            if (p.out2) tma_store_4d(&gp->tmO2, R2g, p.tst_c0 + t.nt * p.BN + g * 64, w0, h0, n0);
            bulk_commit();
          }
          return tte_ok;
        }
        if (use_stg) {
          PROF_STAMP(5);  // accumulator complete
          if (p.res != nullptr) {
            // residual rows of this warp's piece of the tile, as whole row segments
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (i < P) {  // warp-uniform
                const int row = i * rb_step + rb_row0;
                const long long rp_row = __shfl_sync(0xffffffffu, r_pix, row);
                const int v_row = __shfl_sync(0xffffffffu, static_cast<int>(valid), row);
                if (v_row && cols_on)
                  *reinterpret_cast<uint4*>(wst + row * rowb + ((rb_piece ^ stg_key(row)) << 4)) =
                      p.chain_n ? ld_global_cg16(p.res + rp_row + colw + rb_piece * 8)
                                : __ldg(reinterpret_cast<const uint4*>(p.res + rp_row + colw + rb_piece * 8));
              }
            }
            __syncwarp();
          }
          PROF_STAMP(6);  // residual rows staged
          uint8_t* my_row = wst + lane * rowb;
          const int my_key = stg_key(lane);
          // (unrolled only where the statistics index registers by c: the last-tile code runs from a cold instruction cache)
#pragma unroll(STATS ? 4 : 1)
          for (int c = 0; c < 4; ++c) {
            if (c * 16 < wcols) {  // warp-uniform
              const int j = ew * wcols + c * 16;
              uint32_t v[16];
              __syncwarp();
              tmem_ld16(t_addr + j, v);
              tmem_ld_wait();
              if (valid && cols_on) {
                float f[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
                linear_bias_act_16(p, f, t.nt * p.BN + j, staged ? eparm : nullptr, j, brow);
                uint4* s0 = reinterpret_cast<uint4*>(my_row + (((2 * c) ^ my_key) << 4));
                uint4* s1 = reinterpret_cast<uint4*>(my_row + (((2 * c + 1) ^ my_key) << 4));
                if (p.res) {
                  const uint4 q0 = *s0, q1 = *s1;
                  const uint32_t qq[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
                  for (int i = 0; i < 8; ++i) {
                    f[2 * i] += lo16<F16>(qq[i]);
                    f[2 * i + 1] += hi16<F16>(qq[i]);
                  }
                }
                if (p.act2 == kAct2Tanh) {
#pragma unroll
                  for (int i = 0; i < 16; ++i) f[i] = tanhf(f[i]);
                }
                uint4 o0, o1;
                o0.x = pack16<F16>(f[0], f[1]);
                o0.y = pack16<F16>(f[2], f[3]);
                o0.z = pack16<F16>(f[4], f[5]);
                o0.w = pack16<F16>(f[6], f[7]);
                o1.x = pack16<F16>(f[8], f[9]);
                o1.y = pack16<F16>(f[10], f[11]);
                o1.z = pack16<F16>(f[12], f[13]);
                o1.w = pack16<F16>(f[14], f[15]);
                *s0 = o0;
                *s1 = o1;
                if (STATS) {
                  const uint32_t ow_[8] = {o0.x, o0.y, o0.z, o0.w, o1.x, o1.y, o1.z, o1.w};
#pragma unroll
                  for (int i = 0; i < 8; ++i) {
                    f[2 * i] = lo16<F16>(ow_[i]);
                    f[2 * i + 1] = hi16<F16>(ow_[i]);
                  }
                  st_keep(f);
                }
              } else if (STATS) {
                st_zero();
              }
              if (STATS) st_add(c, st_tmp);
            }
          }
          // the accumulator stage goes back to the MMA warp before the global stores
          tc_fence_before();
          __syncwarp();
          PROF_STAMP(7);  // tile converted into the staging rows
          release<REMOTE>(tempty);
          if (STATS && p.st_mode == 1) st_flush(p, t.th * p.tiles_w + t.tw);
          const int jt = ew * wcols + rb_piece * 8;  // tile column of this lane's 8 channels
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if (i >= P) break;  // warp-uniform
            const int row = i * rb_step + rb_row0;
            const long long op_row = __shfl_sync(0xffffffffu, o_pix, row);
            const int v_row = __shfl_sync(0xffffffffu, static_cast<int>(valid), row);
            if (v_row && cols_on) {
              const uint4 q = *reinterpret_cast<const uint4*>(wst + row * rowb + ((rb_piece ^ stg_key(row)) << 4));
              *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + op_row + colw + rb_piece * 8) = q;
              if (p.out2) {
                // second output from the value as stored (rounded to bf16), e.g. the next block's bn1
                float4 sc0, sc1, sh0, sh1;
                if (staged) {
                  sc0 = *reinterpret_cast<const float4*>(eparm + 512 + jt);
                  sc1 = *reinterpret_cast<const float4*>(eparm + 512 + jt + 4);
                  sh0 = *reinterpret_cast<const float4*>(eparm + 768 + jt);
                  sh1 = *reinterpret_cast<const float4*>(eparm + 768 + jt + 4);
                } else {
                  const int cg = colw + rb_piece * 8;
                  sc0 = __ldg(reinterpret_cast<const float4*>(p.scale2 + cg));
                  sc1 = __ldg(reinterpret_cast<const float4*>(p.scale2 + cg) + 1);
                  sh0 = __ldg(reinterpret_cast<const float4*>(p.shift2 + cg));
                  sh1 = __ldg(reinterpret_cast<const float4*>(p.shift2 + cg) + 1);
                }
                uint4 o2;
                o2.x = pack16<F16>(fmaf(lo16<F16>(q.x), sc0.x, sh0.x), fmaf(hi16<F16>(q.x), sc0.y, sh0.y));
                o2.y = pack16<F16>(fmaf(lo16<F16>(q.y), sc0.z, sh0.z), fmaf(hi16<F16>(q.y), sc0.w, sh0.w));
                o2.z = pack16<F16>(fmaf(lo16<F16>(q.z), sc1.x, sh1.x), fmaf(hi16<F16>(q.z), sc1.y, sh1.y));
                o2.w = pack16<F16>(fmaf(lo16<F16>(q.w), sc1.z, sh1.z), fmaf(hi16<F16>(q.w), sc1.w, sh1.w));
                *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out2) + op_row + colw + rb_piece * 8) = o2;
              }
            }
          }
          return true;
        }
#pragma unroll(STATS ? 4 : 1)
        for (int c = 0; c < 4; ++c) {
          if (ch0 + c < ch1) {  // warp-uniform
            const int j = (ch0 + c) * 16;
            uint32_t v[16];
            __syncwarp();
            tmem_ld16(t_addr + j, v);
            tmem_ld_wait();
            const int col0 = t.nt * p.BN + j;
            if ((valid || p.splits > 1) && col0 < p.n_valid)
              linear_epilogue_16<F16>(p, v, col0, o_pix, r_pix, staged ? eparm : nullptr, j, STATS ? st_tmp : nullptr, brow);
            else if (STATS)
              st_zero();
            if (STATS) st_add(c, st_tmp);
          }
        }
        // hand the accumulator stage back to the MMA warp
        tc_fence_before();
        __syncwarp();
        release<REMOTE>(tempty);
        if (STATS && p.st_mode == 1) st_flush(p, t.th * p.tiles_w + t.tw);
        return true;
  }
};

}  // namespace

// One MMA issuer warp (converged; one elected lane issues).  `par` is a literal at both call sites so that every
// ring position and barrier address derives from kernel parameters and blockIdx only: ptxas then keeps the
// loop state in uniform registers instead of broadcasting it from vector registers before every tcgen05.mma.
__device__ __forceinline__ void mma_issuer(const IgemmParams& p, const int par, const uint32_t tmem_base,
                                           uint64_t* full_bar, uint64_t* empty_bar, uint64_t* tfull_bar,
                                           uint64_t* tempty_bar, uint64_t* bfull_bar_p, int* abort_flag_p, uint8_t* ring,
                                           uint8_t* bres, const bool win, const int row_bytes, const int b_tile_bytes,
                                           const int a_stage_bytes, const int stage_bytes, const int num_kb,
                                           const int tiles_total, const int acc_stride) {
  volatile int& abort_flag = *abort_flag_p;
  uint64_t& bfull_bar = *bfull_bar_p;
    const uint32_t tmem_base_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t idesc = make_idesc_f16(128, p.BN, p.f16 != 0);
    const uint32_t dhi = smem_desc_hi(win ? 1024 : 8 * row_bytes, p.kc != 64);
    const uint32_t ring_u = smem_u32(ring);
    const uint32_t bt16 = static_cast<uint32_t>(b_tile_bytes) >> 4;  // weight tile size in descriptor units
    const uint32_t bres_lo = smem_desc_lo(smem_u32(bres));
    const uint32_t bstep = p.kc_per_tap * bt16;
    const int S2 = p.dual ? p.stages >> 1 : p.stages;
    const uint32_t slot0 = ring_u + (p.dual ? par * S2 : 0) * stage_bytes;  // this issuer's part of the ring
    const uint32_t full0 = smem_u32(&full_bar[p.dual ? par * S2 : 0]);
    const uint32_t empty0 = smem_u32(&empty_bar[p.dual ? par * S2 : 0]);
    const int my_tiles = blockIdx.x < tiles_total ? (tiles_total - 1 - blockIdx.x) / gridDim.x + 1 : 0;
    int stage_r = 0;
    uint32_t phase = 0;
    int as = 0;           // accumulator stage of tile `it` = it % sets
    uint32_t aphase = 0;  // (it / sets) & 1
    if (p.b_res) mbar_wait_nofail(&bfull_bar, 0, &abort_flag);
    PROF_BEGIN();
    for (int it = 0; it < my_tiles; ++it) {
      if (it != 0 && ++as == p.sets) {
        as = 0;
        aphase ^= 1u;
      }
      if (p.dual && (it & 1) != par) continue;  // the other issuer's tile (two accumulator stages in this mode)
      const uint32_t d_tmem = tmem_base_u + as * acc_stride;
      PROF_WAIT1(mbar_wait_nofail(&tempty_bar[as], aphase ^ 1u, &abort_flag));
      tc_fence_after();
      uint32_t bq = bres_lo;  // resident weights: descriptor of the current K chunk (tap (row, 0) in window mode)
      int kci = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        PROF_WAIT0(mbar_wait_nofail(reinterpret_cast<uint64_t*>(__cvta_shared_to_generic(full0 + 8 * stage_r)), phase,
                                    &abort_flag));
        tc_fence_after();
        const uint32_t sa = slot0 + stage_r * stage_bytes;
        const uint32_t a_lo = smem_desc_lo(sa);
        if (win) {
          // the horizontal taps are views of the window shifted by whole pixel rows (128 B = 8 descriptor units)
          if (p.b_res) {
            umma_bf16_batch<4>(d_tmem, a_lo, bq, dhi, idesc, kb != 0);
            umma_bf16_batch<4>(d_tmem, a_lo + 8, bq + bstep, dhi, idesc, 1);
            umma_bf16_batch<4>(d_tmem, a_lo + 16, bq + 2 * bstep, dhi, idesc, 1);
            bq += bt16;
            if (++kci == p.kc_per_tap) {
              kci = 0;
              bq += 2 * bstep;
            }
          } else {
            const uint32_t b_lo = smem_desc_lo(sa + a_stage_bytes);
            umma_bf16_batch<4>(d_tmem, a_lo, b_lo, dhi, idesc, kb != 0);
            umma_bf16_batch<4>(d_tmem, a_lo + 8, b_lo + bt16, dhi, idesc, 1);
            umma_bf16_batch<4>(d_tmem, a_lo + 16, b_lo + 2 * bt16, dhi, idesc, 1);
          }
        } else if (p.kc == 64) {
          const uint32_t b_lo = p.b_res ? bq : smem_desc_lo(sa + a_stage_bytes);
          umma_bf16_batch<4>(d_tmem, a_lo, b_lo, dhi, idesc, kb != 0);
          bq += bt16;
        } else {
          umma_bf16_batch<2>(d_tmem, a_lo, smem_desc_lo(sa + a_stage_bytes), dhi, idesc, kb != 0);
        }
        // frees the smem slot when these MMAs retire
        umma_commit_elect(reinterpret_cast<uint64_t*>(__cvta_shared_to_generic(empty0 + 8 * stage_r)));
        if (++stage_r == S2) {
          stage_r = 0;
          phase ^= 1u;
        }
      }
      umma_commit_elect(&tfull_bar[as]);
    }
    if (par == 0) PROF_END(0, threadIdx.x & 31);
#ifdef GHOST_IGEMM_PROF
    if (par == 0 && blockIdx.x == 0 && blockIdx.y == 0 && (threadIdx.x & 31) == 0) g_tl_cta0[2] = clock64();
    if (par == 0 && blockIdx.x == 0 && blockIdx.y == 0 && (threadIdx.x & 31) == 0) {
      const int cfg[8] = {p.BN, p.stages, p.dual, num_kb, my_tiles, p.a_win, p.b_res, static_cast<int>(gridDim.x)};
      for (int i = 0; i < 8; ++i) g_igemm_prof[159][i] = cfg[i];
    }
#endif
}

template <bool F16, bool STATS>
__global__ void __launch_bounds__(kIgemmThreads, 1) igemm_kernel(const __grid_constant__ IgemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[8];
  __shared__ uint64_t empty_bar[8];
  __shared__ uint64_t tfull_bar[4];
  __shared__ uint64_t tempty_bar[4];
  __shared__ uint64_t hfull_bar[8];   // AAD mode: h tile landed in shared memory
  __shared__ uint64_t hempty_bar[8];  // AAD mode: blended tile stored, h slot free again
  __shared__ uint64_t bfull_bar;      // resident-weights mode: the packed weight slab landed
  __shared__ __align__(16) float mexch[2][4][128];  // AAD mode, 128-channel tiles: mask-logit halves of the two warps sharing a row
  __shared__ uint32_t tmem_base_smem;
  __shared__ int abort_flag;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
#ifdef GHOST_IGEMM_PROF
  const long long cta_t0 = clock64();
  unsigned long long cta_ns0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(cta_ns0));
  const unsigned tl_seq = *reinterpret_cast<volatile unsigned*>(&g_tl_seq) & (kTlSlots - 1);
  if (threadIdx.x == 0) atomicMin(&g_tl_t0[tl_seq], cta_ns0);
  if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) g_tl_cta0[0] = cta_t0;
#endif

  // 1024-byte aligned operand ring (swizzle atoms repeat every 1024 B)
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  const int row_bytes = p.kc * 2;
  const bool win = p.a_win != 0;
  const int b_tile_bytes = p.BN * row_bytes;
  // row-window mode: the A slot holds bw+2 (<= 130) pixel rows of 128 B, the B slot the three taps of one kernel row
  const int a_stage_bytes = win ? 17 * 1024 : 128 * row_bytes;
  const int b_stage_bytes = p.b_res ? 0 : (win ? 3 * b_tile_bytes : b_tile_bytes);
  const int stage_bytes = a_stage_bytes + b_stage_bytes;  // multiple of 1024
  const int rows_used = p.bw * p.bh * p.bn;
  const int bres_tiles = p.b_res ? p.ntaps * p.kc_per_tap : 0;
  uint8_t* bres = smem;                                 // resident weight tiles [tap][chunk][BN rows][128 B]
  uint8_t* ring = smem + bres_tiles * b_tile_bytes;     // operand ring

  // AAD mode: h tiles behind the operand ring, each cn/64 boxes of [128 rows][128 B]
  const int cn = p.BN >> 1;
  const int hbox = cn >> 6;
  const int htile_bytes = hbox * 16384;
  uint8_t* hbuf = ring + p.stages * stage_bytes;

  // K blocks handled by this CTA (row-window mode: one block = one kernel row x one 64-channel chunk = 3 taps)
  const int num_kb = win ? 3 * p.kc_per_tap : p.ntaps * p.kc_per_tap / p.splits;
  const int kb0 = blockIdx.y * num_kb;
  const int tiles_total = p.tiles_w * p.tiles_h * p.tiles_nb * p.tiles_n;

  // Prologue in two steps, so that the TMA producers do not wait for the TMEM allocation and the parameter staging:
  // warp 1 initialises the mbarriers and releases the producers (named barrier 9), then allocates TMEM while the
  // epilogue warps stage their parameters; issuers and epilogue meet on named barrier 10.
  if (warp == 1) {
    // one mbarrier per lane and instruction (linear mode: all 16 epilogue warps share every tile; AAD mode: 4 warps
    // per 64-channel box of a tile)
    const uint32_t epi_arrivals = p.mode == kModeAAD ? 4 * hbox : kIgemmEpiWarps;
    if (lane == 0) abort_flag = 0;
    uint64_t* bar = nullptr;
    uint32_t cnt = 1;
    if (lane < 8) bar = &full_bar[lane], cnt = p.b_res ? 1 : 2;  // one arrive.expect_tx per operand producer
    else if (lane < 16) bar = &empty_bar[lane - 8];
    else if (lane < 20) bar = &tfull_bar[lane - 16];
    else if (lane < 24) bar = &tempty_bar[lane - 20], cnt = epi_arrivals;
    else if (lane == 24) bar = &bfull_bar;
    if (bar) mbar_init(bar, cnt);
    if (lane < 8) mbar_init(&hfull_bar[lane], 1);
    else if (lane < 16) mbar_init(&hempty_bar[lane - 8], epi_arrivals);
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmA);
    prefetch_tmap(&p.tmB);
    if (p.mode == kModeAAD) {
      prefetch_tmap(&p.tmH);
      prefetch_tmap(&p.tmO);
    }
    if (p.tst) prefetch_tmap(&p.tmO);
  }
  if (warp == 1) {
    __syncwarp();
    named_bar_arrive(9, 96);
    tmem_alloc(&tmem_base_smem, p.tmem_cols);
  }
  // linear mode, one N tile: the per-channel epilogue parameters go to shared memory once (mexch is free in this
  // mode), so the epilogue of a tile never waits on global memory for them
  float* eparm = &mexch[0][0][0];
  const bool staged = p.mode == kModeLinear && p.tiles_n == 1 && p.splits == 1;
  if (staged && warp >= kIgemmEpiWarp0) {
    for (int i = threadIdx.x - kIgemmEpiWarp0 * 32; i < p.BN; i += kIgemmEpiWarps * 32) {
      const bool in = i < p.n_valid;
      eparm[i] = (in && p.bias) ? __ldg(p.bias + i) : 0.f;
      eparm[256 + i] = (in && p.act1 == kActPRelu) ? __ldg(p.slope_c + i) : 0.f;
      eparm[512 + i] = (in && p.out2) ? __ldg(p.scale2 + i) : 0.f;
      eparm[768 + i] = (in && p.out2) ? __ldg(p.shift2 + i) : 0.f;
    }
  }
  uint32_t tmem_base = 0;
  if (warp == 0 || warp == 3) {
    named_bar_sync(9, 96);
  } else {
    tc_fence_before();
    named_bar_sync(10, kIgemmThreads - 64);
    tc_fence_after();
    tmem_base = tmem_base_smem;
  }
  const int acc_stride = p.tmem_cols / p.sets;  // columns per accumulator stage
  // Programmatic dependent launch: everything above touched only kernel parameters and static weights, so it may
  // overlap the tail of the previous kernel in the stream; the next kernel may start its own prologue from here on.
  // Every role that reads activations or writes outputs waits for the previous grid first (the weight producer and
  // the MMA issuers touch neither).
  grid_dep_launch();
#ifdef GHOST_IGEMM_PROF
  if (threadIdx.x == 0 && blockIdx.x < 160 && blockIdx.y == 0) g_igemm_prof[blockIdx.x][9] = clock64() - cta_t0;
  if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) g_tl_cta0[1] = clock64();
#endif

  if (warp == 0) {
    // ============================== TMA producer: activations =================
    // Converged warp, one elected lane issues the copies; no data-dependent exits (mbar_wait_nofail).  The weight
    // operand has its own producer (warp 3): a single thread cannot feed both operands of short K blocks fast
    // enough.  dual-issuer mode: even tiles go through ring slots [0, S2), odd tiles through [S2, 2*S2).
    const int S2 = p.dual ? p.stages >> 1 : p.stages;
    int st[2] = {0, 0};
    uint32_t ph[2] = {0, 0};
    if (p.b_res && elect_one()) {
      mbar_expect_tx(&bfull_bar, bres_tiles * b_tile_bytes);
      for (int i = 0; i < bres_tiles; ++i)
        tma_load_2d(bres + i * b_tile_bytes, &p.tmB, &bfull_bar, p.b_k0 + i * p.kc,
                    p.aad_pair ? static_cast<int>(blockIdx.x & 1) * p.BN : 0);
    }
    __syncwarp();
    grid_dep_wait();
    TileWalk walk;
    walk.init(p, blockIdx.x, gridDim.x);
    int hs = 0;
    uint32_t hphase = 0;
    const int tap0 = kb0 / p.kc_per_tap, kci0 = kb0 - tap0 * p.kc_per_tap;
    const uint32_t a_tx = win ? (p.bw + 2) * 128 : rows_used * row_bytes;
    int hit = 0;
    PROF_BEGIN();
    for (int tile = blockIdx.x; tile < tiles_total; tile += gridDim.x, ++hit, walk.next(p)) {
      const TileCoord t = walk.t;
      const int w0 = t.tw * p.bw, h0 = t.th * p.bh, n0 = t.tn * p.bn;
      if (p.mode == kModeAAD) {
        PROF_WAIT1(mbar_wait_nofail(&hempty_bar[hs], hphase ^ 1u, &abort_flag));
        if (elect_one()) {
          if (p.aad_no_h) {
            mbar_arrive(&hfull_bar[hs]);  // nothing to load: the slot is only the staging of the TMA store
          } else {
            mbar_expect_tx(&hfull_bar[hs], rows_used * cn * 2);
            for (int b = 0; b < hbox; ++b)
              tma_load_4d(hbuf + hs * htile_bytes + b * 16384, &p.tmH, &hfull_bar[hs],
                          (p.aad_pair ? 0 : t.nt * cn) + b * 64, w0, h0, n0);
          }
        }
        __syncwarp();
        if (++hs == p.h_bufs) {
          hs = 0;
          hphase ^= 1u;
        }
      }
      int tap = tap0 + (p.nphase > 1 ? (t.nt / p.tpn) * p.ntaps : 0), kci = kci0;
      const int rg = p.dual ? (hit & 1) : 0;
      int stage_r = st[rg];
      uint32_t phase = ph[rg];
      for (int kb = 0; kb < num_kb; ++kb) {
        const int stage = rg * S2 + stage_r;
        PROF_WAIT0(mbar_wait_nofail(&empty_bar[stage], phase ^ 1u, &abort_flag));
        if (elect_one()) {
          uint8_t* sa = ring + stage * stage_bytes;
          mbar_expect_tx(&full_bar[stage], a_tx);
          if (win) {
            // kb = (kernel row `tap`, chunk kci): pixels w0-1 .. w0+bw of input row h0+tap-1
            tma_load_4d(sa, &p.tmA, &full_bar[stage], p.a_c0 + kci * 64, w0 - 1, h0 + tap - 1, n0);
          } else {
            const int c = p.a_c0 + p.tap_c[tap] + kci * p.kc;
            if (p.a_rank == 4)
              tma_load_4d(sa, &p.tmA, &full_bar[stage], c, w0 + p.tap_w[tap], h0 + p.tap_h[tap], n0);
            else
              tma_load_5d(sa, &p.tmA, &full_bar[stage], c, w0 + p.tap_w[tap], p.tap_p[tap], h0 + p.tap_h[tap], n0);
          }
        }
        __syncwarp();
        if (++kci == p.kc_per_tap) {
          kci = 0;
          ++tap;
        }
        if (++stage_r == S2) {
          stage_r = 0;
          phase ^= 1u;
        }
      }
      st[rg] = stage_r;
      ph[rg] = phase;
    }
    PROF_END(3, lane);
  } else if (warp == 3) {
    // ============================== TMA producer: weights =====================
    if (!p.b_res) {
      const int S2 = p.dual ? p.stages >> 1 : p.stages;
      int st[2] = {0, 0};
      uint32_t ph[2] = {0, 0};
      const int tap0 = kb0 / p.kc_per_tap, kci0 = kb0 - tap0 * p.kc_per_tap;
      int nt = blockIdx.x % p.tiles_n;           // N tile of the current tile (tiles walk nt fastest)
      const int dnt = gridDim.x % p.tiles_n;
      int hit = 0;
      for (int tile = blockIdx.x; tile < tiles_total; tile += gridDim.x, ++hit) {
        int tap = tap0, kci = kci0;
        const int rg = p.dual ? (hit & 1) : 0;
        int stage_r = st[rg];
        uint32_t phase = ph[rg];
        for (int kb = 0; kb < num_kb; ++kb) {
          const int stage = rg * S2 + stage_r;
          mbar_wait_nofail(&empty_bar[stage], phase ^ 1u, &abort_flag);
          if (elect_one()) {
            uint8_t* sb = ring + stage * stage_bytes + a_stage_bytes;
            mbar_expect_tx(&full_bar[stage], b_stage_bytes);
            if (win) {
              for (int dwi = 0; dwi < 3; ++dwi)
                tma_load_2d(sb + dwi * b_tile_bytes, &p.tmB, &full_bar[stage],
                            p.b_k0 + ((tap * 3 + dwi) * p.kc_per_tap + kci) * 64, nt * p.BN);
            } else {
              tma_load_2d(sb, &p.tmB, &full_bar[stage], p.b_k0 + (kb0 + kb) * p.kc, nt * p.BN);
            }
          }
          __syncwarp();
          if (++kci == p.kc_per_tap) {
            kci = 0;
            ++tap;
          }
          if (++stage_r == S2) {
            stage_r = 0;
            phase ^= 1u;
          }
        }
        st[rg] = stage_r;
        ph[rg] = phase;
        nt += dnt;
        if (nt >= p.tiles_n) nt -= p.tiles_n;
      }
    }
  } else if (warp == 1) {
    // ============================== MMA issuers ===============================
    // Warp 1 issues the even tiles of this CTA (accumulator stage 0) and warp 2 the odd ones (stage 1) when
    // p.dual is set, else warp 1 issues everything.  One warp cannot issue tcgen05.mma fast enough for short
    // tiles; a tile's MMAs all come from one thread, so the accumulation order - and the result - does not
    // depend on the split.  The loops have no data-dependent exits (see mbar_wait_nofail).
    mma_issuer(p, 0, tmem_base, full_bar, empty_bar, tfull_bar, tempty_bar, &bfull_bar, &abort_flag, ring, bres, win,
               row_bytes, b_tile_bytes, a_stage_bytes, stage_bytes, num_kb, tiles_total, acc_stride);
  } else if (warp == 2) {
    if (p.dual)
      mma_issuer(p, 1, tmem_base, full_bar, empty_bar, tfull_bar, tempty_bar, &bfull_bar, &abort_flag, ring, bres, win,
                 row_bytes, b_tile_bytes, a_stage_bytes, stage_bytes, num_kb, tiles_total, acc_stride);
  } else {
    // ============================== epilogue ==================================
    grid_dep_wait();
    const int lg = warp & 3;                       // TMEM lane group this warp may read
    const int ew = (warp - kIgemmEpiWarp0) >> 2;   // 0..3: which of the four warps of that lane group
    const int r = lg * 32 + lane;                  // GEMM row inside the tile
    const int rn = r / (p.bh * p.bw);
    const int rh = (r / p.bw) % p.bh;
    const int rw = r % p.bw;
    if (p.mode == kModeLinear) {
      // all 16 warps share every tile: warp ew takes a quarter of its columns
      LinEpi<F16, STATS> epi;
      epi.init(p, warp, lane, ring, eparm, staged, rows_used, tmem_base, acc_stride, hfull_bar);  // (free in linear mode)
      int it = 0;
      TileWalk walk;
      walk.init(p, blockIdx.x, gridDim.x);
      for (int tile = blockIdx.x; tile < tiles_total; tile += gridDim.x, ++it, walk.next(p)) {
        const bool use_stg = p.stg && tile + static_cast<int>(gridDim.x) >= tiles_total;
        const int as = p.sets == 4 ? (it & 3) : (it & 1);
        const uint32_t aph = p.sets == 4 ? (it >> 2) & 1 : (it >> 1) & 1;
        if (!epi.template tile<false>(p, &p, walk.t, as, aph, use_stg, &tfull_bar[as], &tempty_bar[as], &abort_flag))
          break;
      }
      if (!abort_flag) epi.st_finish(p);
    } else {
      // ---------------- AAD blend epilogue (network/AADLayer.py:20-38) ----------------
      // h arrives by TMA in shared memory (SWIZZLE_128B boxes of 64 channels); each warp blends its own 32 rows x
      // 64 channels in place and sends them out with its own TMA store, so no warp ever waits for another one
      // (except the pair sharing a row of a 128-channel tile, for the mask logit).  The 16 warps form `sets`
      // groups; group s serves the tiles with it % sets == s, i.e. always accumulator stage s.
      const int set = hbox == 1 ? ew : ew >> 1;
      const int box = hbox == 1 ? 0 : ew & 1;
      const int sets = p.sets;
      const bool h2 = p.h_bufs == 2 * sets;  // two h slots per group
      const int sw = r & 7;
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(lg * 32) << 16) + set * acc_stride + box * 64;
      const bool quad = p.bn == 1;              // all rows of a tile in one image: parameters are per thread
      const bool spade = p.aad_spade != 0;
      const bool no_h = spade && p.aad_no_h != 0;
      const float sp_slope = p.aad_slope;
      const float sp_post = (spade && p.aad_post) ? __ldg(p.aad_post) : 1.f;
      const int rq = lane >> 2, cq = lane & 3;  // quad mapping: rows rq + 8k of the piece, channel pairs 8i + 2cq
      TileWalk walk;
      walk.init(p, blockIdx.x + set * gridDim.x, gridDim.x * sets);
      int k = 0;  // tiles this group has served
      int pend = -1;  // h slot whose TMA store has been issued but not yet waited for
      for (int tile = blockIdx.x + set * gridDim.x; tile < tiles_total; tile += gridDim.x * sets, ++k, walk.next(p)) {
        const TileCoord t = walk.t;
        const int hs = h2 ? set + (k & 1) * sets : set;
        const uint32_t hphase = h2 ? (k >> 1) & 1 : k & 1;
        const uint32_t aphase = k & 1;
        const int n = t.tn * p.bn + rn, oh = t.th * p.bh + rh, ow = t.tw * p.bw + rw;
        const bool valid = (r < rows_used) && (n < p.Nimg) && (oh < p.Ho) && (ow < p.Wo);
        const int c_base = t.nt * cn + box * 64;
        const int nc = min(n, p.Nimg - 1);  // rows outside the tensor still index valid parameters
        uint8_t* hrow = hbuf + hs * htile_bytes + box * 16384 + r * 128;
        uint8_t* hq = hbuf + hs * htile_bytes + box * 16384 + (lg * 32 + rq) * 128 + 4 * cq;  // + 1024k: row rq + 8k
        const int nq = min(t.tn, p.Nimg - 1);  // quad mapping: the image of the whole tile
        // paired layers: this tile's layer = its N tile; rows of that layer in wh_fold / bh_fold
        const int pair_c = p.aad_pair ? t.nt * cn : 0, pair_n = p.aad_pair ? t.nt * p.Nimg : 0;
#ifdef GHOST_IGEMM_PROF
        const long long q0_ = clock64();
#endif
        if (!__all_sync(0xffffffffu, mbar_wait(&hfull_bar[hs], hphase, &abort_flag))) break;
#ifdef GHOST_IGEMM_PROF
        const long long q1_ = clock64();
        long long qb_ = 0;
#endif
        float m = 0.f;
        float m4[4] = {0.f, 0.f, 0.f, 0.f};  // quad mapping: mask of rows rq, rq+8, rq+16, rq+24 of this warp's piece
        if (spade) {
          // SPADE: no attention mask, the blend reduces to gamma * hn + beta
        } else if (quad) {
          if (p.mask) {
            if (valid) m = __ldg(p.mask + (static_cast<long long>(n) * p.Ho + oh) * p.Wo + ow);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) m4[kk] = __shfl_sync(0xffffffffu, m, rq + 8 * kk);
          } else {
            // every thread sums its 16 channels of its four rows, the four threads of a row add up by shuffle
            const float* wrow = p.wh_fold + static_cast<long long>(nq) * p.c_x + pair_c + box * 64 + 2 * cq;
            float a4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float2 w = __ldg(reinterpret_cast<const float2*>(wrow + 8 * i));
              const uint8_t* hp = hq + ((i ^ rq) << 4);
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) {
                const uint32_t hv = *reinterpret_cast<const uint32_t*>(hp + kk * 1024);
                a4[kk] += lo16<F16>(hv) * w.x + hi16<F16>(hv) * w.y;
              }
            }
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              a4[kk] += __shfl_xor_sync(0xffffffffu, a4[kk], 1);
              a4[kk] += __shfl_xor_sync(0xffffffffu, a4[kk], 2);
            }
            if (hbox == 2) {
              // the other 64 channels of these rows belong to the partner warp: swap halves, add in box order
              float(*ex)[128] = mexch[k & 1];
              if (cq == 0) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) ex[set * 2 + box][lg * 32 + rq + 8 * kk] = a4[kk];
              }
#ifdef GHOST_IGEMM_PROF
              const long long b0_ = clock64();
#endif
              named_bar_sync(1 + set * 4 + lg, 64);
#ifdef GHOST_IGEMM_PROF
              qb_ = clock64() - b0_;
#endif
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) {
                const float other = ex[set * 2 + (box ^ 1)][lg * 32 + rq + 8 * kk];
                a4[kk] = box == 0 ? a4[kk] + other : other + a4[kk];
              }
            }
            const float bhf = __ldg(p.bh_fold + pair_n + nq);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) m4[kk] = 1.f / (1.f + __expf(-(a4[kk] + bhf)));
          }
        } else
        if (p.mask) {
          if (valid) m = __ldg(p.mask + (static_cast<long long>(n) * p.Ho + oh) * p.Wo + ow);
        } else {
          // sigmoid(conv_h(instance_norm(h))) with the normalisation folded into wh/bh (single N tile: cn == c_x).
          // Runs while the tensor core is still producing this tile's gamma/beta.
          const float* wrow = p.wh_fold + static_cast<long long>(nc) * p.c_x + pair_c + box * 64;
          float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const uint4 v = *reinterpret_cast<const uint4*>(hrow + ((q ^ sw) << 4));
            const float4 w0 = __ldg(reinterpret_cast<const float4*>(wrow + q * 8));
            const float4 w1 = __ldg(reinterpret_cast<const float4*>(wrow + q * 8) + 1);
            acc0 += lo16<F16>(v.x) * w0.x + hi16<F16>(v.x) * w0.y + lo16<F16>(v.y) * w0.z + hi16<F16>(v.y) * w0.w;
            acc1 += lo16<F16>(v.z) * w1.x + hi16<F16>(v.z) * w1.y + lo16<F16>(v.w) * w1.z + hi16<F16>(v.w) * w1.w;
          }
          float acc = acc0 + acc1;
          if (hbox == 2) {
            // the other 64 channels of this row belong to the partner warp: swap halves, add in box order
            float(*ex)[128] = mexch[k & 1];
            ex[set * 2 + box][r] = acc;
#ifdef GHOST_IGEMM_PROF
            const long long b0_ = clock64();
#endif
            named_bar_sync(1 + set * 4 + lg, 64);
#ifdef GHOST_IGEMM_PROF
            qb_ = clock64() - b0_;
#endif
            const float other = ex[set * 2 + (box ^ 1)][r];
            acc = box == 0 ? acc + other : other + acc;
          }
          m = 1.f / (1.f + __expf(-(acc + __ldg(p.bh_fold + pair_n + nc))));
        }
        if (pend >= 0) {
          if (lane == 0) {
            bulk_wait_read<0>();
            mbar_arrive(&hempty_bar[pend]);
          }
          pend = -1;
        }
#ifdef GHOST_IGEMM_PROF
        const long long q2_ = clock64();
#endif
        if (!__all_sync(0xffffffffu, mbar_wait(&tfull_bar[set], aphase, &abort_flag))) break;
#ifdef GHOST_IGEMM_PROF
        const long long q3_ = clock64();
#endif
        tc_fence_after();
        if (quad) {
          // Quad mapping (tiles of one image): tcgen05.ld.16x256b hands thread t the channel pairs 8i + 2(t%4) of
          // rows t/4 + {0, 8, 16, 24}, so the per-channel parameters are loaded once per thread for four rows and
          // spread over the lanes - the row-per-lane mapping below re-reads every parameter in all 32 lanes, which
          // saturates the load/store writeback path.
          const float4* parq = p.ncpar + static_cast<long long>(nq) * p.c_x + c_base + 2 * cq;
          const float* bgq = p.bias + t.nt * p.BN + box * 64 + 2 * cq;
          const float* bbq = bgq + cn;
#pragma unroll 2
          for (int i = 0; i < 8; ++i) {
            uint32_t g0[4], g1[4], b0[4], b1[4];
            __syncwarp();
            tmem_ld_16x256b(t_addr + 8 * i, g0);
            tmem_ld_16x256b(t_addr + (16u << 16) + 8 * i, g1);
            tmem_ld_16x256b(t_addr + cn + 8 * i, b0);
            tmem_ld_16x256b(t_addr + (16u << 16) + cn + 8 * i, b1);
            uint8_t* hp = hq + ((i ^ rq) << 4);
            uint32_t hv[4];
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) hv[kk] = no_h ? 0u : *reinterpret_cast<const uint32_t*>(hp + kk * 1024);
            const float4 p0 = __ldg(parq + 8 * i);      // {rstd0, rstd1, shift0, shift1}
            const float4 p1 = __ldg(parq + 8 * i + 1);  // {gamma_id0, gamma_id1, beta_id0, beta_id1}
            const float2 bgv = __ldg(reinterpret_cast<const float2*>(bgq + 8 * i));
            const float2 bbv = __ldg(reinterpret_cast<const float2*>(bbq + 8 * i));
            tmem_ld_wait();
            const uint32_t gg[4][2] = {{g0[0], g0[1]}, {g0[2], g0[3]}, {g1[0], g1[1]}, {g1[2], g1[3]}};
            const uint32_t bq[4][2] = {{b0[0], b0[1]}, {b0[2], b0[3]}, {b1[0], b1[1]}, {b1[2], b1[3]}};
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const float2 x = make_float2(lo16<F16>(hv[kk]), hi16<F16>(hv[kk]));
              const float2 hn = fma2(x, make_float2(p0.x, p0.y), make_float2(p0.z, p0.w));
              const float2 ga = add2(make_float2(__uint_as_float(gg[kk][0]), __uint_as_float(gg[kk][1])), bgv);
              const float2 be = add2(make_float2(__uint_as_float(bq[kk][0]), __uint_as_float(bq[kk][1])), bbv);
              const float2 A = fma2(ga, hn, be);
              const float2 I = fma2(make_float2(p1.x, p1.y), hn, make_float2(p1.z, p1.w));
              const float2 r2 = fma2(make_float2(m4[kk], m4[kk]), sub2(I, A), A);
              // in place, with the ReLU that always follows the layer (AADLayer.py:62,70)
              *reinterpret_cast<uint32_t*>(hp + kk * 1024) =
                  spade ? pack16_slope<F16>(r2.x, r2.y, sp_slope, sp_post) : pack16_relu<F16>(r2.x, r2.y);
            }
          }
        } else {
        const float4* par = p.ncpar + static_cast<long long>(nc) * p.c_x + c_base;
        const float* bg = p.bias + t.nt * p.BN + box * 64;
        const float* bb = bg + cn;
        const float2 mm = make_float2(m, m);
#pragma unroll 2
        for (int j = 0; j < 64; j += 8) {
          uint32_t g[8], b[8];
          __syncwarp();
          tmem_ld8(t_addr + j, g);
          tmem_ld8(t_addr + cn + j, b);
          uint4* hp = reinterpret_cast<uint4*>(hrow + (((j >> 3) ^ sw) << 4));
          const uint4 h0 = no_h ? make_uint4(0u, 0u, 0u, 0u) : *hp;
          const float4 u0 = __ldg(reinterpret_cast<const float4*>(bg + j));
          const float4 u1 = __ldg(reinterpret_cast<const float4*>(bg + j) + 1);
          const float4 v0 = __ldg(reinterpret_cast<const float4*>(bb + j));
          const float4 v1 = __ldg(reinterpret_cast<const float4*>(bb + j) + 1);
          tmem_ld_wait();
          const uint32_t qq[4] = {h0.x, h0.y, h0.z, h0.w};
          const float2 bgv[4] = {{u0.x, u0.y}, {u0.z, u0.w}, {u1.x, u1.y}, {u1.z, u1.w}};
          const float2 bbv[4] = {{v0.x, v0.y}, {v0.z, v0.w}, {v1.x, v1.y}, {v1.z, v1.w}};
          uint32_t o[4];
          const float4* parj = par + j;
          // two adjacent channels per packed fp32 instruction; the parameter table is laid out per channel pair
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 x = make_float2(lo16<F16>(qq[i]), hi16<F16>(qq[i]));
            const float4 p0 = __ldg(parj + 2 * i);      // {rstd0, rstd1, shift0, shift1}
            const float4 p1 = __ldg(parj + 2 * i + 1);  // {gamma_id0, gamma_id1, beta_id0, beta_id1}
            const float2 hn = fma2(x, make_float2(p0.x, p0.y), make_float2(p0.z, p0.w));
            const float2 ga = add2(make_float2(__uint_as_float(g[2 * i]), __uint_as_float(g[2 * i + 1])), bgv[i]);
            const float2 be = add2(make_float2(__uint_as_float(b[2 * i]), __uint_as_float(b[2 * i + 1])), bbv[i]);
            const float2 A = fma2(ga, hn, be);
            const float2 I = fma2(make_float2(p1.x, p1.y), hn, make_float2(p1.z, p1.w));
            const float2 r2 = fma2(mm, sub2(I, A), A);
            // the ReLU that always follows the layer (AADLayer.py:62,70)
            o[i] = spade ? pack16_slope<F16>(r2.x, r2.y, sp_slope, sp_post) : pack16_relu<F16>(r2.x, r2.y);
          }
          *hp = make_uint4(o[0], o[1], o[2], o[3]);  // in place: this thread alone reads and writes this chunk
        }
        }
        fence_proxy_async();  // generic-proxy writes of the blended rows -> visible to the TMA (async proxy)
        tc_fence_before();
        __syncwarp();
#ifdef GHOST_IGEMM_PROF
        const long long q4_ = clock64();
#endif
        if (lane == 0) {
          mbar_arrive(&tempty_bar[set]);  // accumulator stage back to the MMA warp
          // lane 0 holds the first of this warp's 32 rows: its pixel is the origin of the store box
          if (r < rows_used) tma_store_4d(&p.tmO, hrow, c_base, ow, oh, n);
          bulk_commit();
          if (!h2) {
            bulk_wait_read<0>();  // the store has read the rows: the h slot may be refilled
            mbar_arrive(&hempty_bar[hs]);
          }
        }
        // two h slots per group: the slot is handed back while the next tile's mask is computed, so nobody waits
        // for the store to drain
        pend = h2 ? hs : -1;
        __syncwarp();
#ifdef GHOST_IGEMM_PROF
        if (warp == kIgemmEpiWarp0 && lane == 0 && blockIdx.x < 160) {
          unsigned long long* a = g_aad_prof[blockIdx.x];
          if (k == 0) a[0] = a[1] = a[2] = a[3] = a[4] = 0ull, a[7] = q0_;
          a[0] += q1_ - q0_;
          a[1] += qb_;
          a[2] += q3_ - q2_;
          a[3] += (q4_ - q3_) + (q2_ - q1_) - qb_;
          a[4] += clock64() - q4_;
          a[5] = clock64() - a[7];
          a[6] = k + 1;
        }
#endif
      }
      if (pend >= 0 && lane == 0) {
        bulk_wait_read<0>();
        mbar_arrive(&hempty_bar[pend]);
      }
      if (lane == 0) bulk_wait<0>();  // outstanding TMA stores
    }
  }

  // ------------------------------ teardown ------------------------------------
#ifdef GHOST_IGEMM_PROF
  if (warp == kIgemmEpiWarp0 && lane == 0 && blockIdx.x < 160 && blockIdx.y == 0)
    g_igemm_prof[blockIdx.x][10] = clock64() - cta_t0;  // epilogue done
  if (warp == kIgemmEpiWarp0 && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0) g_tl_cta0[3] = clock64();
#endif
  tc_fence_before();
  __syncthreads();
#ifdef GHOST_IGEMM_PROF
  if (threadIdx.x == 0 && blockIdx.x < 160 && blockIdx.y == 0) {
    unsigned long long ns1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns1));
    g_igemm_prof[blockIdx.x][8] = clock64() - cta_t0;
    g_igemm_prof[blockIdx.x][11] = ns1 - cta_ns0;
  }
  if (threadIdx.x == 0) {
    unsigned long long ns1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns1));
    atomicMax(&g_tl_t1[tl_seq], ns1);
    if (blockIdx.x == 0 && blockIdx.y == 0) {
      g_tl_meta[tl_seq][0] = p.mode * 1000 + p.BN;
      g_tl_meta[tl_seq][1] = p.tiles_w * p.tiles_h * p.tiles_nb * p.tiles_n;
      g_tl_meta[tl_seq][2] = (p.a_win ? 3 * p.kc_per_tap : p.ntaps * p.kc_per_tap / p.splits) * (p.kc == 64 ? 1 : -1);
      g_tl_meta[tl_seq][3] = gridDim.x * 10 + p.a_win;
      g_tl_meta[tl_seq][4] = static_cast<int>(g_tl_cta0[1] - g_tl_cta0[0]);
      g_tl_meta[tl_seq][5] = static_cast<int>(*reinterpret_cast<volatile long long*>(&g_tl_cta0[2]) - g_tl_cta0[0]);
      g_tl_meta[tl_seq][6] = static_cast<int>(*reinterpret_cast<volatile long long*>(&g_tl_cta0[3]) - g_tl_cta0[0]);
      g_tl_meta[tl_seq][7] = static_cast<int>(clock64() - g_tl_cta0[0]);
      g_tl_meta[tl_seq][8] = static_cast<int>(*reinterpret_cast<volatile long long*>(&g_tl_cta0[4]) - g_tl_cta0[0]);
      g_tl_meta[tl_seq][9] = static_cast<int>(*reinterpret_cast<volatile long long*>(&g_tl_cta0[5]) - g_tl_cta0[0]);
    }
    __threadfence();
    if (atomicAdd(&g_tl_done, 1u) == gridDim.x * gridDim.y - 1) {
      g_tl_done = 0;
      __threadfence();
      atomicAdd(&g_tl_seq, 1u);
    }
  }
#endif
  if (abort_flag && threadIdx.x == 0 && p.err) {
    *reinterpret_cast<volatile int*>(p.err) = 1;  // pinned, host-mapped word
    __threadfence_system();
  }
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// CTA-pair kernel (tcgen05 cta_group::2): linear mode, 64-channel K chunks, streaming weights.
//
// A one-CTA tile of 128 x BN brings 16 KB of activations + BN*128 B of weights into the SM per K block; at BN = 256
// that is 48 KB per 4.2 MFLOP = 87 FLOP/B, while one SM takes ~64 B/clk from L2 and retires 8192 dense 16-bit
// FLOP/clk = 128 FLOP/B: the one-CTA kernel is bound by operand delivery (measured ~0.43 us per K block against
// 0.26 us of MMA time).  Two CTAs on the SM pair of a TPC run ONE 256 x BN MMA per K step instead: each CTA loads its
// own 128 pixel rows and HALF of the weight tile, the tensor cores of both SMs read both halves, and each SM keeps
// the 128 accumulator rows of its own pixels in its own TMEM.  Per SM and K block: 16 KB + BN*64 B.
//
// Roles per CTA as in igemm_kernel (warp 0: activation TMA, warp 3: weight TMA, warps 4-19: epilogue of the CTA's own
// 128 rows); warp 1 of the LEADER CTA (cluster rank 0) issues the MMAs of the pair, warp 2 of both CTAs owns the
// TMEM allocation.  Barriers: `full` lives in the leader (both CTAs' TMA loads count their bytes there), `empty` and
// `tfull` are signalled in both CTAs by multicast commits, `tempty` lives in the leader and collects the epilogue
// warps of both CTAs.
template <bool F16, bool STATS, bool CHAIN>
__global__ void __launch_bounds__(kIgemmThreads, 1) igemm2_kernel(const __grid_constant__ IgemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[8];
  __shared__ uint64_t empty_bar[8];
  __shared__ uint64_t tfull_bar[4];
  __shared__ uint64_t tempty_bar[4];
  __shared__ uint64_t bfull_bar;  // resident-weights mode (leader): both CTAs' halves of the weight slab landed
  __shared__ uint64_t rbar[4];    // whole-tile TMA epilogue: residual box of each 64-column group
  __shared__ __align__(16) float eparm[1024];
  __shared__ uint32_t tmem_base_smem;
  __shared__ int abort_flag;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster = blockIdx.x >> 1, nclusters = gridDim.x >> 1;
#ifdef GHOST_IGEMM_PROF
  const long long cta_t0 = clock64();
  unsigned long long cta_ns0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(cta_ns0));
  const unsigned tl_seq = *reinterpret_cast<volatile unsigned*>(&g_tl_seq) & (kTlSlots - 1);
  if (threadIdx.x == 0) atomicMin(&g_tl_t0[tl_seq], cta_ns0);
  if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) g_tl_cta0[0] = cta_t0;
#endif

  // Row-window mode (a_win, see igemm_kernel): the A slot holds bw + 2 pixel rows of one image row, the three
  // horizontal taps are row-shifted views of it; the weight operand is this CTA's HALF of every tile (BN/2 rows),
  // resident for the whole kernel (b_res) or streamed three taps at a time.
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  // Tile-window mode (a_win = 2, narrow images): the A slot holds the (bh + 2) x (bw + 2) pixel patch around a tile of
  // bh whole image rows, loaded ONCE per 64-channel chunk; all nine taps are views of it shifted by dy * (bw + 2) + dx
  // pixel rows, so GEMM row m is pixel (m / (bw + 2), m % (bw + 2)) of the tile and the rows with m % (bw + 2) >= bw
  // are junk that the epilogue skips.  One K block = one chunk = nine MMA batches; the weights are resident.
  const bool win = p.a_win == 1;
  const bool win2 = p.a_win == 2;
  const int b_half_bytes = (p.BN >> 1) * 128;
  const int a_stage_bytes = win2 ? p.a_slot : (win ? 17 * 1024 : 128 * 128);
  const int b_stage_bytes = p.b_res ? 0 : (win ? 3 * b_half_bytes : b_half_bytes);
  const int stage_bytes = a_stage_bytes + b_stage_bytes;  // multiple of 1024
  const int bres_tiles = p.b_res ? p.ntaps * p.kc_per_tap : 0;
  uint8_t* bres = smem;
  uint8_t* ring = smem + bres_tiles * b_half_bytes;
  const int rows_used = win2 ? p.bh * (p.bw + 2) : p.bw * p.bh * p.bn;  // GEMM rows a tile spans
  const int num_kb = win2 ? p.kc_per_tap : (win ? 3 * p.kc_per_tap : p.ntaps * p.kc_per_tap + p.x_kb);
  const int main_kb = p.ntaps * p.kc_per_tap;  // (plain mode) blocks of the conv itself; the rest come from tmH
  const int tiles_m = p.tiles_w * p.tiles_h * p.tiles_nb;
  const int super_total = ((tiles_m + 1) >> 1) * p.tiles_n;  // pairs of M tiles x N tiles
  const int acc_stride = p.tmem_cols / p.sets;
  // conv chain (igemm.cuh `chain`): nsteps convs of this geometry; a cluster walks whole groups of spg super tiles
  const int nsteps = CHAIN ? p.chain_n : 1;
  const int spg = CHAIN ? p.chain_spg : 1;
  const int ngroups = super_total / spg;
  IgemmParams* const sp = reinterpret_cast<IgemmParams*>(ring + p.stages * stage_bytes);  // [2], chain launches only
  float* const eparm2 = reinterpret_cast<float*>(ring + p.stages * stage_bytes + 2 * sizeof(IgemmParams));  // [2][1024]

  if (warp == 1) {
    if (lane == 0) abort_flag = 0;
    if (lane < 8) mbar_init(&full_bar[lane], p.b_res ? 1 : 2);  // the leader's producers (activations, weights)
    else if (lane < 16) mbar_init(&empty_bar[lane - 8], 1);
    else if (lane < 20) mbar_init(&tfull_bar[lane - 16], 1);
    else if (lane < 24) mbar_init(&tempty_bar[lane - 20], 2 * kIgemmEpiWarps);
    else if (lane == 24) mbar_init(&bfull_bar, 1);
    else if (lane < 29) mbar_init(&rbar[lane - 25], 1);
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tmA);
    prefetch_tmap(&p.tmB);
    if (p.tst) prefetch_tmap(&p.tmO);
  }
  const bool staged = p.tiles_n == 1;
  if (staged && warp >= kIgemmEpiWarp0) {
    float* ep0 = CHAIN ? eparm2 : eparm;
    for (int i = threadIdx.x - kIgemmEpiWarp0 * 32; i < p.BN; i += kIgemmEpiWarps * 32) {
      const bool in = i < p.n_valid;
      ep0[i] = (in && p.bias) ? __ldg(p.bias + i) : 0.f;
      ep0[256 + i] = (in && p.act1 == kActPRelu) ? __ldg(p.slope_c + i) : 0.f;
      ep0[512 + i] = (in && p.out2) ? __ldg(p.scale2 + i) : 0.f;
      ep0[768 + i] = (in && p.out2) ? __ldg(p.shift2 + i) : 0.f;
    }
  }
  if (CHAIN) {  // step 0's parameter block (= p) into its shared-memory slot
    const uint4* src = reinterpret_cast<const uint4*>(p.chain);
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(IgemmParams) / 16); i += kIgemmThreads)
      reinterpret_cast<uint4*>(sp)[i] = __ldg(src + i);
  }
  if (warp == 2) tmem2_alloc(&tmem_base_smem, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer's barriers are initialised before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  grid_dep_launch();
#ifdef GHOST_IGEMM_PROF
  if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) g_tl_cta0[1] = clock64();
#endif

  // super tile s -> this CTA's own tile: N tile s % tiles_n, M tile 2 * (s / tiles_n) + rank (an odd tile count
  // leaves the last pair's second CTA with a tile past the batch: its loads are zero-filled, its rows never stored)
  auto my_tile = [&](int s) {
    TileCoord t;
    t.nt = s % p.tiles_n;
    int m = (s / p.tiles_n) * 2 + static_cast<int>(rank);
    t.tw = m % p.tiles_w;
    m /= p.tiles_w;
    t.th = m % p.tiles_h;
    t.tn = m / p.tiles_h;
    return t;
  };

  if (warp == 0) {
    // ============================== TMA producer: activations (both CTAs) =====
    if (p.b_res && elect_one()) {
      // this CTA's half (BN/2 rows) of every resident weight tile; the bytes of both CTAs count on the leader
      if (rank == 0) mbar_expect_tx(&bfull_bar, 2u * bres_tiles * b_half_bytes);
      for (int i = 0; i < bres_tiles; ++i)
        tma2_load_2d(bres + i * b_half_bytes, &p.tmB, &bfull_bar, p.b_k0 + i * 64, static_cast<int>(rank) * (p.BN >> 1));
    }
    __syncwarp();
    grid_dep_wait();
    int stage = 0;
    uint32_t phase = 0;
    const uint32_t a_tx = win2 ? 2u * (p.bw + 2) * (p.bh + 2) * 128u
                               : (win ? 2u * (p.bw + 2) * 128u : 2u * rows_used * 128u);  // both CTAs' boxes
    for (int step = 0; step < nsteps; ++step) {
      const CUtensorMap* tmA = CHAIN ? &p.chain[step].tmA : &p.tmA;
      const int a_c0 = CHAIN ? sp[step & 1].a_c0 : p.a_c0;
      for (int g = cluster; g < ngroups; g += nclusters) {
        for (int u = 0; u < spg; ++u) {
          const TileCoord t = my_tile(g * spg + u);
          const int w0 = t.tw * p.bw, h0 = t.th * p.bh, n0 = t.tn * p.bn;
          int tap = p.nphase > 1 ? (t.nt / p.tpn) * p.ntaps : 0, kci = 0;
          for (int kb = 0; kb < num_kb; ++kb) {
            mbar_wait_nofail(&empty_bar[stage], phase ^ 1u, &abort_flag);
            if (elect_one()) {
              uint8_t* sa = ring + stage * stage_bytes;
              if (rank == 0) mbar_expect_tx(&full_bar[stage], a_tx);
              if (win2) {
                // kb = chunk: the patch of pixels (h0-1 .. h0+bh, w0-1 .. w0+bw)
                tma2_load_4d(sa, tmA, &full_bar[stage], a_c0 + kb * 64, w0 - 1, h0 - 1, n0);
              } else if (win) {
                // kb = (kernel row `tap`, chunk kci): pixels w0-1 .. w0+bw of input row h0+tap-1
                tma2_load_4d(sa, tmA, &full_bar[stage], a_c0 + kci * 64, w0 - 1, h0 + tap - 1, n0);
              } else if (kb >= main_kb) {
                // extra K blocks: 64-channel chunks of the second tensor (igemm.cuh x_kb)
                const int c = p.x_c0 + p.x_tap_c + (kb - main_kb) * 64;
                if (p.a_rank == 4)
                  tma2_load_4d(sa, &p.tmH, &full_bar[stage], c, w0 + p.x_tap_w, h0 + p.x_tap_h, n0);
                else
                  tma2_load_5d(sa, &p.tmH, &full_bar[stage], c, w0 + p.x_tap_w, p.x_tap_p, h0 + p.x_tap_h, n0);
              } else {
                const int c = a_c0 + p.tap_c[tap] + kci * 64;
                if (p.a_rank == 4)
                  tma2_load_4d(sa, tmA, &full_bar[stage], c, w0 + p.tap_w[tap], h0 + p.tap_h[tap], n0);
                else
                  tma2_load_5d(sa, tmA, &full_bar[stage], c, w0 + p.tap_w[tap], p.tap_p[tap], h0 + p.tap_h[tap], n0);
              }
            }
            __syncwarp();
            if (++kci == p.kc_per_tap) {
              kci = 0;
              ++tap;
            }
            if (++stage == p.stages) {
              stage = 0;
              phase ^= 1u;
            }
          }
        }
      }
      if (CHAIN && p.b_res && step + 1 < nsteps) {
        // Resident weights of the NEXT step: as soon as the last MMAs of this step have retired (the accumulator-full
        // barrier of its last tile) both CTAs refill their halves of the slab - under the epilogue of that tile, which
        // nothing else overlaps.  The issuer waits for phase (step + 1) & 1 of bfull_bar.
        const int my_groups = ngroups > cluster ? (ngroups - 1 - cluster) / nclusters + 1 : 0;
        const int done = (step + 1) * my_groups * spg;  // tiles of this CTA so far
        if (done > 0) {
          const int it = done - 1;
          const int as = p.sets == 4 ? (it & 3) : (it & 1);
          const uint32_t aph = p.sets == 4 ? (it >> 2) & 1 : (it >> 1) & 1;
          mbar_wait_nofail(&tfull_bar[as], aph, &abort_flag);
        }
        if (elect_one()) {
          const IgemmParams* nx = p.chain + step + 1;
          const int b_k0n = nx->b_k0;
          if (rank == 0) mbar_expect_tx(&bfull_bar, 2u * bres_tiles * b_half_bytes);
          for (int i = 0; i < bres_tiles; ++i)
            tma2_load_2d(bres + i * b_half_bytes, &nx->tmB, &bfull_bar, b_k0n + i * 64, static_cast<int>(rank) * (p.BN >> 1));
        }
        __syncwarp();
      }
      // the next conv of the chain reads what this cluster's epilogue warps are still writing
      if (CHAIN && step + 1 < nsteps) cluster_sync_all();
    }
  } else if (warp == 3) {
    // ============================== TMA producer: this CTA's half of the weight tile
    int stage = 0;
    uint32_t phase = 0;
    const int half = p.BN >> 1;
    for (int step = 0; step < nsteps; ++step) {
      const CUtensorMap* tmB = CHAIN ? &p.chain[step].tmB : &p.tmB;
      const int b_k0 = CHAIN ? sp[step & 1].b_k0 : p.b_k0;
      if (!p.b_res) {
        for (int g = cluster; g < ngroups; g += nclusters) {
          for (int u = 0; u < spg; ++u) {
            const int row0 = ((g * spg + u) % p.tiles_n) * p.BN + static_cast<int>(rank) * half;
            int tap = 0, kci = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
              mbar_wait_nofail(&empty_bar[stage], phase ^ 1u, &abort_flag);
              if (elect_one()) {
                uint8_t* sb = ring + stage * stage_bytes + a_stage_bytes;
                if (rank == 0) mbar_expect_tx(&full_bar[stage], 2u * b_stage_bytes);
                if (win) {
                  for (int dwi = 0; dwi < 3; ++dwi)
                    tma2_load_2d(sb + dwi * b_half_bytes, tmB, &full_bar[stage],
                                 b_k0 + ((tap * 3 + dwi) * p.kc_per_tap + kci) * 64, row0);
                } else {
                  tma2_load_2d(sb, tmB, &full_bar[stage], b_k0 + kb * 64, row0);
                }
              }
              __syncwarp();
              if (++kci == p.kc_per_tap) {
                kci = 0;
                ++tap;
              }
              if (++stage == p.stages) {
                stage = 0;
                phase ^= 1u;
              }
            }
          }
        }
      }
      if (CHAIN && step + 1 < nsteps) cluster_sync_all();
    }
  } else if (warp == 1) {
    // ============================== MMA issuer of the pair (leader CTA) =======
    if (rank == 0) {
      const uint32_t tmem_base_u = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t idesc = make_idesc_f16(256, p.BN, p.f16 != 0);
      const uint32_t dhi = smem_desc_hi(1024, false);
      const uint32_t ring_u = smem_u32(ring);
      const uint32_t bt16 = static_cast<uint32_t>(b_half_bytes) >> 4;  // weight half-tile in descriptor units
      const uint32_t bres_lo = smem_desc_lo(smem_u32(bres));
      const uint32_t bstep = p.kc_per_tap * bt16;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      int as = 0;
      uint32_t aphase = 0;
      if (p.b_res && !CHAIN) mbar_wait_nofail(&bfull_bar, 0, &abort_flag);
      for (int step = 0; step < nsteps; ++step) {
      if (p.b_res && CHAIN) mbar_wait_nofail(&bfull_bar, static_cast<uint32_t>(step & 1), &abort_flag);
      for (int gu = 0, ng = (ngroups > cluster ? (ngroups - 1 - cluster) / nclusters + 1 : 0) * spg; gu < ng; ++gu, ++it) {
        if (it != 0 && ++as == p.sets) {
          as = 0;
          aphase ^= 1u;
        }
        const uint32_t d_tmem = tmem_base_u + as * acc_stride;
        mbar_wait_nofail(&tempty_bar[as], aphase ^ 1u, &abort_flag);
        tc_fence_after();
        uint32_t bq = bres_lo;
        int kci = 0;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait_nofail(&full_bar[stage], phase, &abort_flag);
          tc_fence_after();
          const uint32_t sa = ring_u + stage * stage_bytes;
          const uint32_t a_lo = smem_desc_lo(sa);
          if (win2) {
            // nine taps = nine views of the patch, shifted by dy * (bw + 2) + dx pixel rows (128 B = 8 descriptor units)
            const uint32_t pitch8 = static_cast<uint32_t>(p.bw + 2) * 8u;
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
              for (int dx = 0; dx < 3; ++dx)
                umma2_bf16_batch4(d_tmem, a_lo + dy * pitch8 + dx * 8, bq + (dy * 3 + dx) * bstep, dhi, idesc,
                                  (kb | dy | dx) != 0);
            }
            bq += bt16;
          } else if (win) {
            // the horizontal taps are views of the window shifted by whole pixel rows (128 B = 8 descriptor units)
            if (p.b_res) {
              umma2_bf16_batch4(d_tmem, a_lo, bq, dhi, idesc, kb != 0);
              umma2_bf16_batch4(d_tmem, a_lo + 8, bq + bstep, dhi, idesc, 1);
              umma2_bf16_batch4(d_tmem, a_lo + 16, bq + 2 * bstep, dhi, idesc, 1);
              bq += bt16;
              if (++kci == p.kc_per_tap) {
                kci = 0;
                bq += 2 * bstep;
              }
            } else {
              const uint32_t b_lo = smem_desc_lo(sa + a_stage_bytes);
              umma2_bf16_batch4(d_tmem, a_lo, b_lo, dhi, idesc, kb != 0);
              umma2_bf16_batch4(d_tmem, a_lo + 8, b_lo + bt16, dhi, idesc, 1);
              umma2_bf16_batch4(d_tmem, a_lo + 16, b_lo + 2 * bt16, dhi, idesc, 1);
            }
          } else {
            const uint32_t b_lo = p.b_res ? bq : smem_desc_lo(sa + a_stage_bytes);
            umma2_bf16_batch4(d_tmem, a_lo, b_lo, dhi, idesc, kb != 0);
            bq += bt16;
          }
          umma2_commit_elect(&empty_bar[stage]);  // frees the slot in both CTAs when these MMAs retire
          if (++stage == p.stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma2_commit_elect(&tfull_bar[as]);
      }
      if (CHAIN && step + 1 < nsteps) cluster_sync_all();
      }
#ifdef GHOST_IGEMM_PROF
      if (blockIdx.x == 0 && blockIdx.y == 0 && lane == 0) g_tl_cta0[2] = clock64();
#endif
    } else if (CHAIN) {
      for (int step = 0; step + 1 < nsteps; ++step) cluster_sync_all();
    }
  } else if (CHAIN && warp == 2) {
    // ============================== chain: the next step's parameter block and epilogue tables ==========
    for (int step = 0; step + 1 < nsteps; ++step) {
      const IgemmParams* nx = p.chain + step + 1;
      uint4* dst = reinterpret_cast<uint4*>(sp + ((step + 1) & 1));
      for (int i = lane; i < static_cast<int>(sizeof(IgemmParams) / 16); i += 32)
        dst[i] = __ldg(reinterpret_cast<const uint4*>(nx) + i);
      if (staged) {
        float* ep = eparm2 + ((step + 1) & 1) * 1024;
        const float* bias = nx->bias;
        const float* slope_c = nx->act1 == kActPRelu ? nx->slope_c : nullptr;
        const float* scale2 = nx->out2 ? nx->scale2 : nullptr;
        const float* shift2 = nx->out2 ? nx->shift2 : nullptr;
        for (int i = lane; i < p.BN; i += 32) {
          const bool in = i < p.n_valid;
          ep[i] = (in && bias) ? __ldg(bias + i) : 0.f;
          ep[256 + i] = (in && slope_c) ? __ldg(slope_c + i) : 0.f;
          ep[512 + i] = (in && scale2) ? __ldg(scale2 + i) : 0.f;
          ep[768 + i] = (in && shift2) ? __ldg(shift2 + i) : 0.f;
        }
      }
      __syncwarp();
      cluster_sync_all();
    }
  } else if (warp >= kIgemmEpiWarp0) {
    // ============================== epilogue of this CTA's 128 rows ===========
#ifdef GHOST_IGEMM_PROF
    const long long gd0_ = clock64();
#endif
    grid_dep_wait();
#ifdef GHOST_IGEMM_PROF
    if (blockIdx.x == 0 && blockIdx.y == 0 && warp == kIgemmEpiWarp0 && lane == 0) g_tl_cta0[4] = g_tl_cta0[0] + (clock64() - gd0_);
#endif
    LinEpi<F16, STATS> epi;
    epi.init(p, warp, lane, ring, CHAIN ? eparm2 : eparm, staged, rows_used, tmem_base, acc_stride, rbar);
    int it = 0;
    for (int step = 0; step < nsteps; ++step) {
      const IgemmParams& q = CHAIN ? sp[step & 1] : p;
      const IgemmParams* gq = CHAIN ? p.chain + step : &p;  // tensor maps: never from shared memory
      if (CHAIN) epi.eparm = eparm2 + (step & 1) * 1024;
      bool ok = true;
      for (int g = cluster; ok && g < ngroups; g += nclusters) {
        for (int u = 0; u < spg; ++u, ++it) {
          const TileCoord t = my_tile(g * spg + u);
          const int as = p.sets == 4 ? (it & 3) : (it & 1);
          const uint32_t aph = p.sets == 4 ? (it >> 2) & 1 : (it >> 1) & 1;
          const bool use_stg = p.stg && g + nclusters >= ngroups && u + 1 == spg;
          if (!epi.template tile<true>(q, gq, t, as, aph, use_stg, &tfull_bar[as], &tempty_bar[as], &abort_flag)) {
            ok = false;
            break;
          }
        }
      }
      if (CHAIN && step + 1 < nsteps) {
        // this thread's plain stores -> visible to the TMA loads of the next conv, in either CTA of the pair
        epi.tte_drain(q);
        fence_proxy_async_all();
        __threadfence();
        cluster_sync_all();
      }
    }
    if (!abort_flag) epi.st_finish(p);
#ifdef GHOST_IGEMM_PROF
    if (warp == kIgemmEpiWarp0 && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0) g_tl_cta0[3] = clock64();
#endif
  }

  // ------------------------------ teardown ------------------------------------
  tc_fence_before();
  __syncthreads();
  if (abort_flag && threadIdx.x == 0 && p.err) {
    *reinterpret_cast<volatile int*>(p.err) = 1;
    __threadfence_system();
  }
#ifdef GHOST_IGEMM_PROF
  if (threadIdx.x == 0) {
    unsigned long long ns1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns1));
    atomicMax(&g_tl_t1[tl_seq], ns1);
    if (blockIdx.x == 0 && blockIdx.y == 0) {
      g_tl_meta[tl_seq][0] = 2000 + p.BN;  // "mode 2" = CTA pair
      g_tl_meta[tl_seq][1] = p.tiles_w * p.tiles_h * p.tiles_nb * p.tiles_n;
      g_tl_meta[tl_seq][2] = num_kb;
      g_tl_meta[tl_seq][3] = gridDim.x * 10 + p.a_win;
      g_tl_meta[tl_seq][4] = static_cast<int>(g_tl_cta0[1] - g_tl_cta0[0]);
      g_tl_meta[tl_seq][5] = static_cast<int>(*reinterpret_cast<volatile long long*>(&g_tl_cta0[2]) - g_tl_cta0[0]);
      g_tl_meta[tl_seq][6] = static_cast<int>(*reinterpret_cast<volatile long long*>(&g_tl_cta0[3]) - g_tl_cta0[0]);
      g_tl_meta[tl_seq][7] = static_cast<int>(clock64() - g_tl_cta0[0]);
      g_tl_meta[tl_seq][8] = static_cast<int>(*reinterpret_cast<volatile long long*>(&g_tl_cta0[4]) - g_tl_cta0[0]);  // griddepcontrol.wait
      g_tl_meta[tl_seq][9] = static_cast<int>(*reinterpret_cast<volatile long long*>(&g_tl_cta0[5]) - g_tl_cta0[0]);
      g_tl_meta[tl_seq][10] = static_cast<int>(*reinterpret_cast<volatile long long*>(&g_tl_cta0[6]) - g_tl_cta0[0]);
      g_tl_meta[tl_seq][11] = static_cast<int>(*reinterpret_cast<volatile long long*>(&g_tl_cta0[7]) - g_tl_cta0[0]);
    }
    __threadfence();
    if (atomicAdd(&g_tl_done, 1u) == gridDim.x * gridDim.y - 1) {
      g_tl_done = 0;
      __threadfence();
      atomicAdd(&g_tl_seq, 1u);
    }
  }
#endif
  cluster_sync_all();  // neither CTA leaves while the other may still signal its barriers / read its operands
  if (warp == 2) {
    tc_fence_after();
    tmem2_dealloc(tmem_base, p.tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Split-K, second pass: one thread per (GEMM row, 16 columns) adds the partial tiles in split order (deterministic)
// and runs the linear epilogue of the op.
template <bool F16>
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const __grid_constant__ IgemmParams p) {
  const int chunks = p.ws_cols >> 4;
  const long long rows_tot = static_cast<long long>(p.tiles_nb) * p.tiles_h * p.tiles_w * 128;
  const long long idx = blockIdx.x * 256LL + threadIdx.x;
  if (idx >= rows_tot * chunks) return;
  const int ch = static_cast<int>(idx % chunks);
  const long long row = idx / chunks;
  const int r = static_cast<int>(row & 127);
  int m = static_cast<int>(row >> 7);
  const int tw = m % p.tiles_w;
  m /= p.tiles_w;
  const int th = m % p.tiles_h, tn = m / p.tiles_h;
  const int rn = r / (p.bh * p.bw), rh = (r / p.bw) % p.bh, rw = r % p.bw;
  const int n = tn * p.bn + rn, oh = th * p.bh + rh, ow = tw * p.bw + rw;
  const int col0 = ch * 16;
  if (r >= p.bw * p.bh * p.bn || n >= p.Nimg || oh >= p.Ho || ow >= p.Wo || col0 >= p.n_valid) return;
  float f[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) f[i] = 0.f;
  const float4* src = reinterpret_cast<const float4*>(p.ws + row * p.ws_cols + col0);
  const long long split_stride = rows_tot * p.ws_cols / 4;
  for (int s = 0; s < p.splits; ++s) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 q = src[s * split_stride + i];
      f[4 * i] += q.x, f[4 * i + 1] += q.y, f[4 * i + 2] += q.z, f[4 * i + 3] += q.w;
    }
  }
  uint32_t v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(f[i]);
  const int ph = p.nphase > 1 ? (col0 / p.BN) / p.tpn : 0;
  const long long o_pix = n * p.o_sn + oh * p.o_sh + ow * p.o_sw + p.o_off + (p.nphase > 1 ? p.ph_o[ph] : 0);
  const long long r_pix = n * p.r_sn + oh * p.r_sh + ow * p.r_sw + p.r_off + (p.nphase > 1 ? p.ph_r[ph] : 0);
  linear_epilogue_16<F16, true>(p, v, col0, o_pix, r_pix, nullptr, 0);
}

const char* splitk_reduce_launch(const IgemmParams& p, cudaStream_t stream) {
  if (p.splits < 2 || !p.ws || p.ws_cols % 16) return "splitk_reduce: not a split-K op";
  const long long rows_tot = static_cast<long long>(p.tiles_nb) * p.tiles_h * p.tiles_w * 128;
  const long long total = rows_tot * (p.ws_cols >> 4);
  const unsigned blocks = static_cast<unsigned>((total + 255) / 256);
  if (p.f16) splitk_reduce_kernel<true><<<blocks, 256, 0, stream>>>(p);
  else splitk_reduce_kernel<false><<<blocks, 256, 0, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

int igemm_smem_bytes(const IgemmParams& p) {
  // behind the operand ring: exchange buffer of the fused statistics, or the staging of the TMA-store epilogue
  const int st = p.st_mode ? 8192 : (p.tst ? 2 * 4 * (p.BN >> 6) * 4096 : 0);
  if (p.cta2) {
    const int b_half = (p.BN >> 1) * 128;
    const int stage2 = p.a_win == 2 ? p.a_slot
                                    : (p.a_win ? 17 * 1024 + (p.b_res ? 0 : 3 * b_half) : 128 * 128 + (p.b_res ? 0 : b_half));
    return p.stages * stage2 + (p.b_res ? p.ntaps * p.kc_per_tap * b_half : 0) + 1024 + st + (p.chain_n ? kChainSmemBytes : 0);
  }
  const int row_bytes = p.kc * 2;
  const int aad = p.mode == kModeAAD ? p.h_bufs * ((p.BN >> 1) >> 6) * 16384 : 0;
  const int b_tile = p.BN * row_bytes;
  const int resident = p.b_res ? p.ntaps * p.kc_per_tap * b_tile : 0;
  const int stage = p.a_win ? 17 * 1024 + (p.b_res ? 0 : 3 * b_tile) : 128 * row_bytes + (p.b_res ? 0 : b_tile);
  return p.stages * stage + resident + aad + 1024 + st;
}

namespace {
// Per-device launch state: the dynamic shared-memory opt-in is a per-device function attribute and the SM count is
// a device property, so both are cached per device ordinal (one process may drive several GPUs).
struct IgemmDevState {
  int max_dyn_smem = 0;  // opt-in dynamic shared memory limit (227 KB minus the kernel's static part)
  int max_dyn_smem2 = 0; // the same for the CTA-pair kernel
  int num_sms = 0;
};
constexpr int kMaxDevices = 64;
IgemmDevState g_dev_state[kMaxDevices];
std::mutex g_dev_mutex;

const char* igemm_dev_state(IgemmDevState* out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cudaGetErrorString(e);
  if (dev < 0 || dev >= kMaxDevices) return "igemm: device ordinal out of range";
  std::lock_guard<std::mutex> lock(g_dev_mutex);
  IgemmDevState& st = g_dev_state[dev];
  if (!st.max_dyn_smem) {
    cudaFuncAttributes fa;
    e = cudaFuncGetAttributes(&fa, igemm_kernel<false, false>);
    int lim = 0;
    if (e == cudaSuccess) {
      lim = 227 * 1024 - static_cast<int>(fa.sharedSizeBytes);
      e = cudaFuncSetAttribute(igemm_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim);
    }
    if (e == cudaSuccess) e = cudaFuncSetAttribute(igemm_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(igemm_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(igemm_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim);
    if (e == cudaSuccess) {
      cudaFuncAttributes fa2;
      e = cudaFuncGetAttributes(&fa2, igemm2_kernel<false, false, false>);
      const int lim2 = 227 * 1024 - static_cast<int>(fa2.sharedSizeBytes);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(igemm2_kernel<false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim2);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(igemm2_kernel<true, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim2);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(igemm2_kernel<false, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim2);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(igemm2_kernel<true, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim2);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(igemm2_kernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim2);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(igemm2_kernel<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim2);
      st.max_dyn_smem2 = lim2;
    }
    int sms = 0;
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) {
      cudaGetLastError();  // do not leave a stale error behind for the next launch check
      return cudaGetErrorString(e);
    }
    st.num_sms = sms > 0 ? sms : 148;
    st.max_dyn_smem = lim;
  }
  *out = st;
  return nullptr;
}
}  // namespace

namespace {
const char* igemm2_launch(const IgemmParams& p, int max_dyn_smem, int num_sms, cudaStream_t stream) {
  if (p.mode != kModeLinear || p.kc != 64 || p.splits != 1 || p.dual)
    return "igemm2: the CTA-pair kernel runs linear-mode convs with 64-channel chunks";
  if (p.a_win == 1 && (p.ntaps != 9 || p.bh != 1 || p.bn != 1 || p.a_rank != 4 || p.bw > 128 || p.nphase > 1))
    return "igemm2: row-window mode needs a 3x3 stride-1 conv and single-row tiles";
  if (p.a_win == 2 && (p.ntaps != 9 || p.bn != 1 || p.a_rank != 4 || p.nphase > 1 || !p.b_res || p.tiles_w != 1 ||
                       p.bh * (p.bw + 2) > 128 || p.a_slot % 1024 || p.a_slot < (2 * (p.bw + 2) + 2 + 128) * 128 || p.stg ||
                       p.tst || p.st_mode))
    return "igemm2: tile-window mode needs a 3x3 stride-1 conv, tiles of whole image rows and resident weights";
  if (p.a_win < 0 || p.a_win > 2) return "igemm2: bad a_win";
  if (p.b_res && p.tiles_n != 1) return "igemm2: resident weights need one N tile";
  if (p.BN % 16 || p.BN < 32 || p.BN > 256) return "igemm2: BN must be a multiple of 16 in [32,256]";
  if (p.bw * p.bh * p.bn > 128 || p.bw * p.bh * p.bn <= 0) return "igemm2: M box must hold 1..128 rows";
  if (p.stages < 2 || p.stages > 8) return "igemm2: stages must be in [2,8]";
  if ((p.sets != 2 && p.sets != 4) || p.tmem_cols < p.sets * p.BN || p.tmem_cols > 512 ||
      (p.tmem_cols & (p.tmem_cols - 1)) || p.tmem_cols < 32)
    return "igemm2: bad tmem_cols";
  if (p.tst && (p.out_f32 || p.st_mode || p.stg || p.out2 || p.nphase > 1 || (p.BN != 64 && p.BN != 128) ||
                (p.bw % 32 && !(p.bh == 1 && p.bn == 1 && p.tiles_w == 1)) || p.n_valid % 64))
    return "igemm2: the TMA-store epilogue needs a plain 16-bit tile of 64 / 128 columns and rows of 32 pixels";
  if (p.ntaps * (p.nphase > 1 ? p.nphase : 1) > kIgemmMaxTaps) return "igemm2: too many taps";
  if (p.nphase > 1 && (p.nphase != 4 || p.tpn < 1 || p.tiles_n != p.nphase * p.tpn)) return "igemm2: bad phase tiling";
  if (p.stg && (p.out_f32 || (p.BN != 128 && p.BN != 256) || p.n_valid % (p.BN >> 2)))
    return "igemm2: staged output needs a 16-bit linear epilogue and N tiles of 128 or 256 whole warp quarters";
  const int smem = igemm_smem_bytes(p);
  if (smem > max_dyn_smem) return "igemm2: shared memory budget exceeded";
  const int tiles_m = p.tiles_w * p.tiles_h * p.tiles_nb;
  const int supers = ((tiles_m + 1) / 2) * p.tiles_n;
  if (supers <= 0) return "igemm2: empty grid";
  int clusters = num_sms / 2;
  if (clusters > supers) clusters = supers;
  if (p.x_kb && (p.a_win || p.b_res || p.chain_n || p.nphase > 1 || p.x_kb < 0))
    return "igemm2: extra K blocks from a second tensor need the plain tap-shifted mode";
  if (p.chain_n) {
    if (p.chain_n < 2 || !p.chain || p.chain_spg < 1 || supers % p.chain_spg || (tiles_m & 1) || p.st_mode || p.tst ||
        (p.b_res && p.a_win != 2) || p.nphase > 1)
      return "igemm2: a conv chain needs whole groups of super tiles and the plain linear epilogue";
    clusters = num_sms / 2;
    if (clusters > supers / p.chain_spg) clusters = supers / p.chain_spg;
  }
  static const int pdl = [] {
    const char* ev = getenv("GHOST_PDL");
    return ev ? (atoi(ev) != 0) : 1;
  }();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * clusters, 1, 1);
  cfg.blockDim = dim3(kIgemmThreads, 1, 1);
  cfg.dynamicSmemBytes = static_cast<size_t>(smem);
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 2 : 1;
  cudaError_t e;
  if (p.chain_n)
    e = p.f16 ? cudaLaunchKernelEx(&cfg, igemm2_kernel<true, false, true>, p)
              : cudaLaunchKernelEx(&cfg, igemm2_kernel<false, false, true>, p);
  else if (p.st_mode)
    e = p.f16 ? cudaLaunchKernelEx(&cfg, igemm2_kernel<true, true, false>, p)
              : cudaLaunchKernelEx(&cfg, igemm2_kernel<false, true, false>, p);
  else
    e = p.f16 ? cudaLaunchKernelEx(&cfg, igemm2_kernel<true, false, false>, p)
              : cudaLaunchKernelEx(&cfg, igemm2_kernel<false, false, false>, p);
  if (e == cudaSuccess) e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}
}  // namespace

const char* igemm_launch(const IgemmParams& p, cudaStream_t stream) {
  IgemmDevState dev_state;
  if (const char* e = igemm_dev_state(&dev_state)) return e;
  if (p.cta2) return igemm2_launch(p, dev_state.max_dyn_smem2, dev_state.num_sms, stream);
  const int max_dyn_smem = dev_state.max_dyn_smem;
  if (p.BN % 16 || p.BN < 16 || p.BN > 256) return "igemm: BN must be a multiple of 16 in [16,256]";
  if (p.kc != 64 && p.kc != 32) return "igemm: kc must be 32 or 64";
  if (p.bw * p.bh * p.bn > 128 || p.bw * p.bh * p.bn <= 0) return "igemm: M box must hold 1..128 rows";
  if (p.stages < 2 || p.stages > 8) return "igemm: stages must be in [2,8]";
  if (p.dual && p.stages < 4) return "igemm: two issuers need at least four ring slots";
  if (p.sets != 2 && p.sets != 4) return "igemm: 2 or 4 accumulator stages";
  if (p.mode == kModeLinear && p.sets != 2 && p.sets != 4) return "igemm: the linear epilogue works on two or four accumulator stages";
  if (p.mode == kModeAAD &&
      (p.sets * (p.BN >> 7) != 4 || (p.h_bufs != p.sets && p.h_bufs != 2 * p.sets) || p.dual || p.BN % 128))
    return "igemm: bad AAD staging";
  if (p.tmem_cols < p.sets * p.BN || p.tmem_cols > 512 || (p.tmem_cols & (p.tmem_cols - 1)) || p.tmem_cols < 32)
    return "igemm: bad tmem_cols";
  if (p.splits < 1 || (p.ntaps * p.kc_per_tap) % p.splits) return "igemm: splits must divide the K blocks";
  if (p.ntaps * (p.nphase > 1 ? p.nphase : 1) > kIgemmMaxTaps) return "igemm: too many taps";
  if (p.nphase > 1 && (p.nphase != 4 || p.mode != kModeLinear || p.a_win || p.b_res || p.tpn < 1 ||
                       p.tiles_n != p.nphase * p.tpn))
    return "igemm: output phases need a plain linear-mode launch";
  if (p.a_win == 2) return "igemm: tile-window mode runs on the CTA-pair kernel";
  if (p.x_kb) return "igemm: extra K blocks from a second tensor run on the CTA-pair kernel";
  if (p.a_win && (p.kc != 64 || p.ntaps != 9 || p.bh != 1 || p.bn != 1 || p.splits != 1 || p.a_rank != 4 || p.bw > 128))
    return "igemm: row-window mode needs a 3x3 stride-1 conv, 64-channel chunks and single-row tiles";
  if (p.b_res && ((p.tiles_n != 1 && !p.aad_pair) || p.kc != 64 || p.splits != 1))
    return "igemm: resident weights need one N tile";
  if (p.stg && (p.mode != kModeLinear || p.out_f32 || p.splits != 1 || (p.BN != 128 && p.BN != 256) ||
                p.n_valid % (p.BN >> 2)))
    return "igemm: staged output needs a bf16 linear epilogue and N tiles of 128 or 256 whole warp quarters";
  if (p.tst && (p.mode != kModeLinear || p.out_f32 || p.splits != 1 || p.st_mode || p.stg || p.out2 || p.nphase > 1 ||
                (p.BN != 64 && p.BN != 128) || (p.bw % 32 && !(p.bh == 1 && p.bn == 1 && p.tiles_w == 1)) || p.n_valid % 64))
    return "igemm: the TMA-store epilogue needs a plain 16-bit linear tile of 64 / 128 columns and rows of 32 pixels";
  const int smem = igemm_smem_bytes(p);
  if (smem > max_dyn_smem) return "igemm: shared memory budget exceeded";
  const int tiles = p.tiles_w * p.tiles_h * p.tiles_nb * p.tiles_n;
  if (tiles <= 0) return "igemm: empty grid";
  const int num_sms = dev_state.num_sms;
  int gx = tiles < num_sms ? tiles : num_sms;
  if (p.splits > 1) {
    gx = num_sms / p.splits;
    if (gx < 1) gx = 1;
    if (gx > tiles) gx = tiles;
  }
  if (p.aad_pair) gx &= ~1;  // CTA b serves layer b & 1 for its whole life
  if (gx < 1) return "igemm: paired AAD layers need an even grid";
  dim3 grid(gx, p.splits, 1);
  // programmatic dependent launch: the prologue of this launch may overlap the tail of the previous kernel in the
  // stream (the kernel waits for it before touching activations); GHOST_PDL=0 falls back to plain serialisation
  static const int pdl = [] {
    const char* ev = getenv("GHOST_PDL");
    return ev ? (atoi(ev) != 0) : 1;
  }();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(kIgemmThreads, 1, 1);
  cfg.dynamicSmemBytes = static_cast<size_t>(smem);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  cudaError_t e;
  if (p.st_mode)
    e = p.f16 ? cudaLaunchKernelEx(&cfg, igemm_kernel<true, true>, p) : cudaLaunchKernelEx(&cfg, igemm_kernel<false, true>, p);
  else
    e = p.f16 ? cudaLaunchKernelEx(&cfg, igemm_kernel<true, false>, p) : cudaLaunchKernelEx(&cfg, igemm_kernel<false, false>, p);
  if (e == cudaSuccess) e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

#ifdef GHOST_IGEMM_PROF
extern "C" int ghost_prof_timeline_reset() {
  cudaDeviceSynchronize();
  unsigned int z = 0;
  cudaMemcpyToSymbol(g_tl_seq, &z, sizeof(z));
  cudaMemcpyToSymbol(g_tl_done, &z, sizeof(z));
  void *t0 = nullptr, *t1 = nullptr;
  cudaGetSymbolAddress(&t0, g_tl_t0);
  cudaGetSymbolAddress(&t1, g_tl_t1);
  cudaMemset(t0, 0xff, sizeof(g_tl_t0));
  cudaMemset(t1, 0, sizeof(g_tl_t1));
  return static_cast<int>(cudaDeviceSynchronize());
}
extern "C" int ghost_prof_timeline_read(unsigned long long* t0, unsigned long long* t1, int* meta) {
  cudaDeviceSynchronize();
  unsigned int n = 0;
  cudaMemcpyFromSymbol(&n, g_tl_seq, sizeof(n));
  cudaMemcpyFromSymbol(t0, g_tl_t0, sizeof(g_tl_t0));
  cudaMemcpyFromSymbol(t1, g_tl_t1, sizeof(g_tl_t1));
  cudaMemcpyFromSymbol(meta, g_tl_meta, sizeof(g_tl_meta));
  return static_cast<int>(n);
}
extern "C" int ghost_prof_aad_read(unsigned long long* out) {
  cudaDeviceSynchronize();
  return static_cast<int>(cudaMemcpyFromSymbol(out, g_aad_prof, sizeof(g_aad_prof)));
}
extern "C" int ghost_prof_igemm_read(unsigned long long* out) {
  cudaDeviceSynchronize();
  return static_cast<int>(cudaMemcpyFromSymbol(out, g_igemm_prof, sizeof(g_igemm_prof)));
}
#endif

}  // namespace ghost
