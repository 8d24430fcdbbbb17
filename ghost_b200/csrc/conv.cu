#include "conv.cuh"

#include <cudaTypedefs.h>
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include <string.h>

namespace ghost {

Taps Taps::square(int k, int pad, int stride) {
  Taps t;
  t.n = k * k;
  t.stride = stride;
  for (int kh = 0; kh < k; ++kh)
    for (int kw = 0; kw < k; ++kw) {
      t.dh[kh * k + kw] = static_cast<short>(kh - pad);
      t.dw[kh * k + kw] = static_cast<short>(kw - pad);
    }
  return t;
}

// ConvTranspose2d(k=4, s=2, p=1): out[2q+r] += in[q+d] * W[kh] with
//   r=0: (kh=1, d=0), (kh=3, d=-1);   r=1: (kh=0, d=+1), (kh=2, d=0)
// (from oh = 2*ih - 1 + kh, network/AEI_Net.py:30).  Tap order matches pack.cc::deconv_kidx.
Taps Taps::deconv_phase(int r, int s) {
  static const short d_of[2][2] = {{0, -1}, {1, 0}};
  Taps t;
  t.n = 4;
  t.stride = 1;
  for (int th = 0; th < 2; ++th)
    for (int tw = 0; tw < 2; ++tw) {
      t.dh[th * 2 + tw] = d_of[r][th];
      t.dw[th * 2 + tw] = d_of[s][tw];
    }
  return t;
}

Taps Taps::point() {
  Taps t;
  t.n = 1;
  t.stride = 1;
  t.dh[0] = t.dw[0] = 0;
  return t;
}

const char* encode_tmap(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                        const uint32_t* box, int swizzle_bytes) {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) return "cuTensorMapEncodeTiled entry point not found";
    fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  }
  cuuint64_t gdims[5];
  cuuint64_t gstr[4];
  cuuint32_t gbox[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
  }
  CUtensorMapSwizzle sw = swizzle_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), gdims, gstr, gbox, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return "cuTensorMapEncodeTiled failed";
  return nullptr;
}

namespace {

void choose_box(int N, int Ho, int Wo, int* bw, int* bh, int* bn) {
  double best = -1.0;
  const long long total = static_cast<long long>(N) * Ho * Wo;
  for (int w = 1; w <= Wo && w <= 128; ++w) {
    for (int h = 1; h <= Ho && w * h <= 128; ++h) {
      int n = 1;
      if (w == Wo && h == Ho) {
        n = 128 / (w * h);
        if (n > N) n = N;
        if (n < 1) n = 1;
      }
      const long long tiles = static_cast<long long>((Wo + w - 1) / w) * ((Ho + h - 1) / h) * ((N + n - 1) / n);
      double eff = static_cast<double>(total) / (static_cast<double>(tiles) * 128.0);
      eff += 1e-6 * w;  // tie-break: longer contiguous runs
      if (eff > best) {
        best = eff;
        *bw = w;
        *bh = h;
        *bn = n;
      }
    }
  }
}

int pow2_at_least(int v) {
  int p = 32;
  while (p < v) p <<= 1;
  return p;
}

}  // namespace

const char* lower_conv(LaunchOp* op, const Act& in, int c0, int Cin, const Taps& taps, int Ho, int Wo,
                       const PackedW& w, const OutView& out, const Epilogue& epi, bool force_simt, int* err_word,
                       const AadArgs* aad, int bn_hint, const DeconvPhases* phases) {
  if (taps.n * (phases ? 4 : 1) > kIgemmMaxTaps) return "lower_conv: too many taps";
  if (phases && (aad || force_simt || w.Cout % 4 || !is16(in.dtype) || Cin % 64))
    return "lower_conv: merged deconv phases need the plain tensor-core path";
  // per-phase channel count: the N tiling never crosses a phase
  const int cout_t = phases ? w.Cout / 4 : w.Cout;
  if (w.K != taps.n * Cin + epi.x2_cin) return "lower_conv: weight K does not match taps*Cin";
  if (epi.x2_cin && (aad || phases || force_simt || Cin % 64 || epi.x2_cin % 64 || !epi.x2.p || epi.x2.N != in.N ||
                     epi.x2.H != in.H || epi.x2.W != in.W || epi.x2.dtype != in.dtype || epi.x2.C % 8 || epi.want_stats || epi.res))
    return "lower_conv: a second input tensor needs a plain tensor-core conv and an input-sized 16-bit tensor";
  op->flops = 2.0 * in.N * Ho * Wo * static_cast<double>(w.Cout) * w.K;
  const bool can_igemm = !force_simt && is16(in.dtype) && (Cin % 32 == 0) && (in.C % 8 == 0) && (c0 % 8 == 0) &&
                         w.wb != nullptr && (taps.stride == 1 || (taps.stride == 2 && in.H % 2 == 0 && in.W % 2 == 0));
  if (!can_igemm) {
    if (aad) return "lower_conv: the fused AAD epilogue needs the tensor-core path";
    if (!w.wf) return "lower_conv: fp32 weights missing for the CUDA-core path";
    op->kind = LaunchOp::kSimt;
    op->simt_dtype = in.dtype;
    if (out.dtype != in.dtype) return "lower_conv: CUDA-core path needs matching in/out dtypes";
    SimtConvParams& p = op->sc;
    memset(&p, 0, sizeof(p));
    p.in = in.p;
    p.N = in.N; p.H = in.H; p.W = in.W; p.C = in.C;
    p.c0 = c0; p.Cin = Cin;
    p.stride = taps.stride;
    p.ntaps = taps.n;
    for (int i = 0; i < taps.n; ++i) { p.dh[i] = taps.dh[i]; p.dw[i] = taps.dw[i]; }
    p.Ho = Ho; p.Wo = Wo;
    p.w = w.wf;
    p.Cout = w.Cout;
    p.out = out.p;
    p.o_sn = out.sn; p.o_sh = out.sh; p.o_sw = out.sw; p.o_off = out.off;
    p.bias = w.bias;
    p.act1 = epi.act1; p.slope = epi.slope; p.slope_c = epi.slope_c;
    p.res = epi.res;
    p.r_sn = epi.r_sn; p.r_sh = epi.r_sh; p.r_sw = epi.r_sw; p.r_off = epi.r_off;
    p.act2 = epi.act2;
    if (epi.out2) return "lower_conv: the second output needs the tensor-core path";
    if (rgb_const_supported(p, in.dtype)) {
      op->rgb = std::make_shared<RgbConvConst>();
      if (const char* e = rgb_const_prepare(p, op->rgb.get())) return e;
    }
    return nullptr;
  }

  // ---- 3x3 / stride 1 / pad 1 with <= 3 output channels: the nine taps ride in N (tapn.cuh)
  static const int tapn_mode = [] {
    const char* e = getenv("GHOST_TAPN");  // A/B switch for profiling: GHOST_TAPN=0 keeps these convs on igemm_kernel
    return e ? (atoi(e) != 0) : 1;
  }();
  bool tapn = tapn_mode != 0 && !aad && !phases && w.wt != nullptr && taps.n == 9 && taps.stride == 1 && Wo == in.W &&
              Ho == in.H && w.Cout <= 3 && Cin % 64 == 0 && Cin / 64 <= kTapnMaxChunks && in.H >= kTapnTileH &&
              in.W >= kTapnTileW && epi.act1 == kActNone && !epi.out2 && out.sw >= w.Cout;
  for (int i = 0; tapn && i < 9; ++i) tapn = taps.dh[i] == i / 3 - 1 && taps.dw[i] == i % 3 - 1;
  if (tapn) {
    op->kind = LaunchOp::kTapn;
    TapnParams& t = op->tp;
    memset(&t, 0, sizeof(t));
    t.chunks = Cin / 64;
    t.c0 = c0;
    t.cout = w.Cout;
    t.N = in.N; t.H = in.H; t.W = in.W;
    t.tiles_x = (in.W + kTapnTileW - 3) / (kTapnTileW - 2);
    t.tiles_y = (in.H + kTapnTileH - 3) / (kTapnTileH - 2);
    const long long tiles = static_cast<long long>(t.tiles_x) * t.tiles_y * in.N;
    if (tiles > 0x7fffffffLL) return "lower_conv: too many tiles";
    t.tiles = static_cast<int>(tiles);
    t.bias = w.bias;
    t.out = out.p;
    t.out_f32 = out.dtype == kF32;
    t.f16 = in.dtype == kF16;
    if (!t.out_f32 && out.dtype != in.dtype) return "lower_conv: 16-bit output type must match the input";
    t.o_sn = out.sn; t.o_sh = out.sh; t.o_sw = out.sw; t.o_off = out.off;
    t.res = epi.res;
    t.r_sn = epi.r_sn; t.r_sh = epi.r_sh; t.r_sw = epi.r_sw; t.r_off = epi.r_off;
    t.act2 = epi.act2;
    t.err = err_word;
    const uint64_t eb = 2;
    const uint64_t dims[4] = {static_cast<uint64_t>(in.C), static_cast<uint64_t>(in.W), static_cast<uint64_t>(in.H),
                              static_cast<uint64_t>(in.N)};
    const uint64_t str[3] = {in.C * eb, static_cast<uint64_t>(in.W) * in.C * eb,
                             static_cast<uint64_t>(in.H) * in.W * in.C * eb};
    const uint32_t box[4] = {64, kTapnTileW, kTapnTileH, 1};
    if (const char* e = encode_tmap(&t.tmA, in.p, 4, dims, str, box, 128)) return e;
    const uint64_t wdims[2] = {64, static_cast<uint64_t>(t.chunks) * 32};
    const uint64_t wstr[1] = {64 * eb};
    const uint32_t wbox[2] = {64, 32};
    if (const char* e = encode_tmap(&t.tmW, w.wt, 2, wdims, wstr, wbox, 128)) return e;
    return nullptr;
  }

  op->kind = LaunchOp::kIgemm;
  IgemmParams& p = op->ig;
  memset(&p, 0, sizeof(p));
  p.kc = (Cin % 64 == 0) ? 64 : 32;
  const int swz = p.kc * 2;
  p.Nimg = in.N; p.Ho = Ho; p.Wo = Wo;
  choose_box(in.N, Ho, Wo, &p.bw, &p.bh, &p.bn);
  // Row-window mode (igemm.cuh): 3x3 / stride 1 / pad 1 over a wide image, tiles are runs of one image row.
  static const int win_mode = [] {
    const char* e = getenv("GHOST_AWIN");  // A/B switch for profiling: GHOST_AWIN=0 turns the mode off
    return e ? (atoi(e) != 0) : 1;
  }();
  bool win = win_mode != 0 && !aad && !phases && taps.n == 9 && taps.stride == 1 && p.kc == 64 && Wo == in.W && Ho == in.H &&
             Wo >= 112 && w.Cout <= 128;
  for (int i = 0; win && i < 9; ++i) win = taps.dh[i] == i / 3 - 1 && taps.dw[i] == i % 3 - 1;
  if (win) {
    const int tw = (Wo + 127) / 128;
    p.bw = (Wo + tw - 1) / tw;
    p.bh = 1;
    p.bn = 1;
  }
  // Tile-window mode (igemm.cuh a_win = 2): the same conv over a NARROW image on the CTA-pair kernel.  A tile = bh whole
  // image rows; its (bh + 2) x (Wo + 2) pixel patch crosses L2 -> shared once per 64-channel chunk instead of nine
  // times, the weights stay resident (half of the slab per CTA).  The tap-shifted loads of the plain kernel make these
  // layers L2-bound (measured 0.26 us per K block of a 128-wide pair tile against 0.135 us of MMA time): iresnet
  // layers 1-2 (56 x 56 x 64, 28 x 28 x 128).
  static const int win2_mode = [] {
    const char* e = getenv("GHOST_AWIN2");  // A/B switch for profiling
    return e ? atoi(e) : 1;
  }();
  static const int cta2_env = [] {
    const char* e = getenv("GHOST_CTA2");
    return e ? atoi(e) : 1;
  }();
  bool win2 = false;
  int w2_slot = 0, w2_slab = 0;
  if (!win && win2_mode && cta2_env && !aad && !phases && !epi.want_stats && !force_simt && taps.n == 9 && taps.stride == 1 &&
      p.kc == 64 && Wo == in.W && Ho == in.H && Wo + 2 <= 64 && w.Cout <= 128 && w.Cout % 32 == 0 && (bn_hint >> 10) != 2 &&
      ((bn_hint & 1023) == 0 || (bn_hint & 1023) == w.Cout) && is16(out.dtype)) {
    bool taps_ok = true;
    for (int i = 0; taps_ok && i < 9; ++i) taps_ok = taps.dh[i] == i / 3 - 1 && taps.dw[i] == i % 3 - 1;
    const int pitch = Wo + 2;
    int hmax = 128 / pitch;
    if (hmax > Ho) hmax = Ho;
    if (taps_ok && hmax >= 1) {
      const int tiles_h = (Ho + hmax - 1) / hmax;
      const int bh = (Ho + tiles_h - 1) / tiles_h;  // the smallest tile height with that many tiles per image
      const long long tiles_m = static_cast<long long>(tiles_h) * in.N;
      const double eff = static_cast<double>(Ho) * Wo / (tiles_h * 128.0);
      int slot = (2 * pitch + 2 + 128) * 128;  // rows the farthest tap view reads (past the patch: junk rows only)
      if (slot < (bh + 2) * pitch * 128) slot = (bh + 2) * pitch * 128;
      slot = (slot + 1023) / 1024 * 1024;
      const int slab = 9 * (Cin / 64) * (w.Cout / 2) * 128;
      if (tiles_m >= 2 && eff >= 0.7 && slab + 2 * slot + 1024 <= 221 * 1024) {
        win2 = true;
        w2_slot = slot;
        w2_slab = slab;
        p.bw = Wo;
        p.bh = bh;
        p.bn = 1;
      }
    }
  }
  p.tiles_w = (Wo + p.bw - 1) / p.bw;
  p.tiles_h = (Ho + p.bh - 1) / p.bh;
  p.tiles_nb = (in.N + p.bn - 1) / p.bn;
  p.ntaps = taps.n;
  p.kc_per_tap = Cin / p.kc;
  p.a_c0 = c0;
  p.b_k0 = 0;
  p.splits = 1;

  // ---- activation tensor map
  const uint64_t eb = 2;
  if (taps.stride == 1) {
    p.a_rank = 4;
    const uint64_t dims[4] = {static_cast<uint64_t>(in.C), static_cast<uint64_t>(in.W), static_cast<uint64_t>(in.H),
                              static_cast<uint64_t>(in.N)};
    const uint64_t str[3] = {in.C * eb, static_cast<uint64_t>(in.W) * in.C * eb,
                             static_cast<uint64_t>(in.H) * in.W * in.C * eb};
    const uint32_t box[4] = {static_cast<uint32_t>(p.kc), static_cast<uint32_t>(win || win2 ? p.bw + 2 : p.bw),
                             static_cast<uint32_t>(win2 ? p.bh + 2 : p.bh), static_cast<uint32_t>(p.bn)};
    if (const char* e = encode_tmap(&p.tmA, in.p, 4, dims, str, box, swz)) return e;
    for (int i = 0; i < taps.n; ++i) {
      p.tap_c[i] = 0;
      p.tap_w[i] = taps.dw[i];
      p.tap_p[i] = 0;
      p.tap_h[i] = taps.dh[i];
    }
    for (int ph = 1; phases && ph < 4; ++ph) {
      if (phases->taps[ph].n != taps.n || phases->taps[ph].stride != 1) return "lower_conv: deconv phases must share the tap count";
      for (int i = 0; i < taps.n; ++i) {
        p.tap_c[ph * taps.n + i] = 0;
        p.tap_w[ph * taps.n + i] = phases->taps[ph].dw[i];
        p.tap_p[ph * taps.n + i] = 0;
        p.tap_h[ph * taps.n + i] = phases->taps[ph].dh[i];
      }
    }
  } else {
    if (phases) return "lower_conv: deconv phases are stride-1 gathers";
    // parity-split view: (2C | W/2 | 2 | H/2 | N); input pixel (2*hh+ph, 2*ww+pw), channel c lives at
    // inner index pw*C + c.
    p.a_rank = 5;
    const uint64_t dims[5] = {static_cast<uint64_t>(2 * in.C), static_cast<uint64_t>(in.W / 2), 2,
                              static_cast<uint64_t>(in.H / 2), static_cast<uint64_t>(in.N)};
    const uint64_t str[4] = {2 * in.C * eb, static_cast<uint64_t>(in.W) * in.C * eb,
                             2 * static_cast<uint64_t>(in.W) * in.C * eb,
                             static_cast<uint64_t>(in.H) * in.W * in.C * eb};
    const uint32_t box[5] = {static_cast<uint32_t>(p.kc), static_cast<uint32_t>(p.bw), 1,
                             static_cast<uint32_t>(p.bh), static_cast<uint32_t>(p.bn)};
    if (const char* e = encode_tmap(&p.tmA, in.p, 5, dims, str, box, swz)) return e;
    for (int i = 0; i < taps.n; ++i) {
      const int dw = taps.dw[i], dh = taps.dh[i];
      p.tap_c[i] = static_cast<short>((dw & 1) * in.C);
      p.tap_w[i] = static_cast<short>(dw >> 1);  // arithmetic shift: floor(dw / 2)
      p.tap_p[i] = static_cast<short>(dh & 1);
      p.tap_h[i] = static_cast<short>(dh >> 1);
    }
  }

  if (epi.x2_cin) {
    // the second tensor through tmH: same view, same box as tmA; one 1x1 tap at the conv's stride
    const Act& x2 = epi.x2;
    if (taps.stride == 1) {
      const uint64_t dims[4] = {static_cast<uint64_t>(x2.C), static_cast<uint64_t>(x2.W), static_cast<uint64_t>(x2.H),
                                static_cast<uint64_t>(x2.N)};
      const uint64_t str[3] = {x2.C * eb, static_cast<uint64_t>(x2.W) * x2.C * eb,
                               static_cast<uint64_t>(x2.H) * x2.W * x2.C * eb};
      const uint32_t box[4] = {static_cast<uint32_t>(p.kc), static_cast<uint32_t>(p.bw), static_cast<uint32_t>(p.bh),
                               static_cast<uint32_t>(p.bn)};
      if (const char* e = encode_tmap(&p.tmH, x2.p, 4, dims, str, box, swz)) return e;
    } else {
      const uint64_t dims[5] = {static_cast<uint64_t>(2 * x2.C), static_cast<uint64_t>(x2.W / 2), 2,
                                static_cast<uint64_t>(x2.H / 2), static_cast<uint64_t>(x2.N)};
      const uint64_t str[4] = {2 * x2.C * eb, static_cast<uint64_t>(x2.W) * x2.C * eb,
                               2 * static_cast<uint64_t>(x2.W) * x2.C * eb,
                               static_cast<uint64_t>(x2.H) * x2.W * x2.C * eb};
      const uint32_t box[5] = {static_cast<uint32_t>(p.kc), static_cast<uint32_t>(p.bw), 1,
                               static_cast<uint32_t>(p.bh), static_cast<uint32_t>(p.bn)};
      if (const char* e = encode_tmap(&p.tmH, x2.p, 5, dims, str, box, swz)) return e;
    }
    p.x_kb = epi.x2_cin / 64;
    p.x_c0 = 0;
    p.x_tap_c = p.x_tap_w = p.x_tap_p = p.x_tap_h = 0;  // pixel (stride * o, stride * o): parity 0 of the split view
  }

  // ---- N tiling
  if (aad) {
    const int cn = aad->pair ? aad->c_x / 2 : (aad->c_x < 128 ? aad->c_x : 128);
    if (aad->pair && cn != 64 && cn != 128) return "lower_conv: paired AAD layers need 64 or 128 channels each";
    if (aad->c_x % cn || cn % 16) return "lower_conv: AAD c_x must be a multiple of 16";
    p.BN = 2 * cn;
    p.tiles_n = aad->c_x / cn;
    p.n_valid = 2 * aad->c_x;
    if (w.rows_pad < 2 * aad->c_x) return "lower_conv: AAD weights must hold 2*c_x rows";
  } else {
    int bn_ = bn_hint & 1023;
    // CTA-pair kernel (igemm.cuh `cta2`): plain linear-mode convs with 64-channel chunks; bits 10-11 of the hint
    // force it on (1) / off (2) for the kernel tests, GHOST_CTA2=0 turns it off for A/B profiling
    static const int cta2_mode = [] {
      const char* e = getenv("GHOST_CTA2");
      return e ? atoi(e) : 1;
    }();
    const int force2 = bn_hint >> 10;
    const long long tiles_m2 = static_cast<long long>(p.tiles_w) * p.tiles_h * p.tiles_nb;
    const int nph = phases ? 4 : 1;
    const int nkb_all = taps.n * p.kc_per_tap + epi.x2_cin / 64;
    auto t1 = [](int bn) { return bn <= 128 ? 0.262 + 0.00022 * bn : 0.29 + 0.00105 * (bn - 128); };
    auto t2 = [](int bn) { return bn <= 128 ? 0.216 + 0.00033 * bn : 0.258 + 0.00095 * (bn - 128); };
    // Cost of a K block per SM, measured on B200 (tools/conv_sweep.py, us): one CTA 0.276 / 0.29 / 0.425 at N tiles
    // of 64 / 128 / 256 (operand delivery: 16 KB + BN*128 B per block), CTA pair 0.237 / 0.258 / 0.38 (16 KB +
    // BN*64 B; 0.38 us is the tensor-core time of a 128 x 256 x 64 block at the measured 1.63 PFLOP/s).  Both
    // kernels run rounds of a persistent grid: 148 CTAs / 74 pairs.  Split-K (one-CTA kernel, one round): K blocks
    // of a tile spread over `s` CTAs + a reduce pass (~3 us + the fp32 partial tiles through L2).
    const int cpad0 = (cout_t + 15) / 16 * 16;
    const int cands0[4] = {256, 128, 64, cpad0 < 256 ? cpad0 : 256};
    double best1 = 1e30, best2 = 1e30, best3 = 1e30;
    int bn2 = 0, bn3 = 0, s3 = 1;
    static const int splitk_mode = [] {
      const char* e = getenv("GHOST_SPLITK");  // A/B switch for profiling: GHOST_SPLITK=0 turns split-K off
      return e ? atoi(e) : 1;
    }();
    const bool pair_ok = !win && p.kc == 64 && cout_t >= 32 && tiles_m2 >= 2 && cta2_mode != 0 && force2 != 2;
    const bool split_ok = !win && !epi.want_stats && !epi.bias9 && !epi.x2_cin && splitk_mode != 0 && force2 == 0 && !bn_ && nkb_all >= 16 && cout_t % 16 == 0;
    for (int i = 0; i < 4; ++i) {
      const int c = cands0[i];
      if ((c > cpad0 && c != cands0[3]) || (bn_ && c != bn_)) continue;
      if (phases && cout_t % c) continue;
      const long long nt = static_cast<long long>(nph) * ((cout_t + c - 1) / c);
      const double c1 = static_cast<double>((tiles_m2 * nt + 147) / 148) * nkb_all * t1(c);
      if (c1 < best1) best1 = c1;
      if (pair_ok) {
        const double c2 = static_cast<double>((((tiles_m2 + 1) / 2) * nt + 73) / 74) * nkb_all * t2(c);
        if (c2 < best2 || (c2 == best2 && c > bn2)) best2 = c2, bn2 = c;
      }
      if (split_ok && tiles_m2 * nt <= 74) {
        for (int sp = 2; sp <= 49; ++sp) {
          if (nkb_all % sp || nkb_all / sp < 4 || tiles_m2 * nt * sp > 148) continue;
          const double ws_mb = static_cast<double>(sp) * tiles_m2 * 128 * nt * c * 4.0 / 1e6;
          const double c3 = (nkb_all / sp) * t1(c) + 3.0 + ws_mb * 2.0 / 4.0;  // written + read at ~4 TB/s (L2)
          if (c3 < best3) best3 = c3, bn3 = c, s3 = sp;
        }
      }
    }
    if (bn_ && !bn2) bn2 = bn_;
    if (bn3 && best3 < 0.85 * (best1 < best2 ? best1 : best2)) {
      p.splits = s3;
      bn_ = bn3;
    } else if (pair_ok && bn2 && (force2 == 1 || best2 < 0.98 * best1)) {
      p.cta2 = 1;
      bn_ = bn2;
    }
    if (!bn_) {
      // Measured on B200 (tools/conv_sweep.py): with one CTA per SM a K block costs ~0.40 us whatever the N tile
      // (issue-bound: four tcgen05.mma + two TMA boxes per block), ~0.43 us at N = 256.  So pick the N tile that
      // minimises (rounds of the persistent grid) x (cost of a K block); ties go to the narrower tile, which
      // spreads the work over more SMs.
      const int cpad = (cout_t + 15) / 16 * 16;
      const long long tiles_m = static_cast<long long>(p.tiles_w) * p.tiles_h * p.tiles_nb;
      double best = 1e30;
      bn_ = cpad < 256 ? cpad : 256;
      const int cands[4] = {256, 128, 64, cpad < 256 ? cpad : 256};
      for (int i = 0; i < 4; ++i) {
        const int bn_c = cands[i];
        if (bn_c > cpad && bn_c != cands[3]) continue;
        if (phases && cout_t % bn_c) continue;
        const long long tiles = tiles_m * nph * ((cout_t + bn_c - 1) / bn_c);
        const double cost = static_cast<double>((tiles + 147) / 148) * (bn_c > 128 ? 0.43 : 0.40);
        if (cost < best || (cost == best && bn_c < bn_)) {
          best = cost;
          bn_ = bn_c;
        }
      }
    }
    // row-window convs (thin K, N <= 128): measured MMA-pipe bound at ~68 clk per 128 x 64 x 16 MMA, each of which
    // re-reads its 4 KB A slab from shared memory; the 256-row MMA of a CTA pair halves that per output
    static const int cta2win_mode = [] {
      const char* e = getenv("GHOST_CTA2WIN");  // A/B switch for profiling
      return e ? atoi(e) : 1;
    }();
    // (with 64 input channels a tile has only three K blocks and the epilogue bounds it either way: measured
    // 0.344 ms one CTA vs 0.362 ms as a pair on AADBlk8's 64 -> 64 conv)
    if (win && !p.cta2 && cta2_mode != 0 && cta2win_mode != 0 && force2 != 2 && tiles_m2 >= 2 && bn_ % 16 == 0 && bn_ >= 32 &&
        (p.kc_per_tap >= 2 || force2 == 1))
      p.cta2 = 1;
    if (win2) {  // one N tile on the pair kernel, whatever the cost model of the plain kernels says
      p.cta2 = 1;
      p.splits = 1;
      bn_ = w.Cout;
    }
    if (epi.x2_cin) {  // the extra K blocks exist on the pair kernel only
      if (win || win2 || !pair_ok || !bn2) return "lower_conv: a second input tensor needs the CTA-pair kernel";
      p.cta2 = 1;
      p.splits = 1;
      bn_ = bn2;
    }
    p.BN = bn_;
    p.tiles_n = nph * ((cout_t + p.BN - 1) / p.BN);
    p.n_valid = nph * cout_t;
    if (phases) {
      if (cout_t % p.BN) return "lower_conv: the N tile must divide the channels of a deconv phase";
      p.nphase = 4;
      p.tpn = cout_t / p.BN;
    }
    if (w.rows_pad < p.tiles_n * p.BN) return "lower_conv: packed weight rows are not padded to the N tile";
  }
  {
    const uint64_t dims[2] = {static_cast<uint64_t>(w.K), static_cast<uint64_t>(w.rows_pad)};
    const uint64_t str[1] = {static_cast<uint64_t>(w.K) * eb};
    const uint32_t box[2] = {static_cast<uint32_t>(p.kc), static_cast<uint32_t>(p.cta2 ? p.BN / 2 : p.BN)};
    if (const char* e = encode_tmap(&p.tmB, w.wb, 2, dims, str, box, swz)) return e;
  }
  int stage_bytes = p.cta2 ? (128 + p.BN / 2) * 128 : (128 + p.BN) * p.kc * 2;
  int resident_bytes = 0;
  // fused output statistics: tiles inside one image, whole 64-column warp quarters, no split-K
  const bool stats = epi.want_stats && !aad;
  if (stats) {
    if (p.bn != 1 || p.BN % 64 || w.Cout % p.BN || p.splits != 1 || phases)
      return "lower_conv: fused output statistics need tiles inside one image and N tiles of whole 64-column quarters";
    const long long tpi = static_cast<long long>(p.tiles_w) * p.tiles_h;
    p.st_mode = (tpi <= 148 || p.tiles_n != 1) ? 1 : 2;
    p.st_P = p.st_mode == 1 ? static_cast<int>(tpi) : 148;
    p.st_C = w.Cout;
    op->st_floats = static_cast<size_t>(in.N) * (2 * p.st_P + 1) * p.st_C;
  }
  // TMA-store epilogue (igemm.cuh tst) for thin-K tiles on the one-CTA kernel
  static const int tst_mode = [] {
    const char* e = getenv("GHOST_TST");  // A/B switch for profiling
    return e ? atoi(e) : 1;
  }();
  bool tst = tst_mode && !aad && !stats && !phases && !win2 && (!p.cta2 || win) && p.splits == 1 && is16(out.dtype) && !epi.out2 &&
                   (p.BN == 64 || (p.BN == 128 && !win)) && w.Cout % 64 == 0 &&
                   // pieces of 32 pixels of one image row: whole tiles of such pieces, or ONE tile per image row whose
                   // last piece the TMA store clips at the image edge (iresnet's 112-pixel rows)
                   ((p.bw % 32 == 0 && p.bh * p.bw * p.bn == 128) || (p.bh == 1 && p.bn == 1 && p.bw == Wo && p.bw > 32)) &&
                   taps.n * p.kc_per_tap <= 18 && out.sw % 8 == 0 && out.off % 8 == 0 &&
                   out.sh == static_cast<long long>(Wo) * out.sw && out.sn == static_cast<long long>(Ho) * Wo * out.sw &&
                   (!epi.res || (!(epi.r_off % 8) && !(epi.r_sw % 8) && !(epi.r_sh % 8) && !(epi.r_sn % 8)));
  int st_bytes = stats ? 8192 : (tst ? 2 * 4 * (p.BN >> 6) * 4096 : 0);
  if (win) {
    const int b_tile = (p.cta2 ? p.BN / 2 : p.BN) * 128;  // a CTA of a pair holds half of every weight tile
    const int slab = 9 * p.kc_per_tap * b_tile;
    p.a_win = win_mode;
    // resident weights beat the TMA-store epilogue when only one of the two fits (the weight slab would otherwise be
    // streamed again for every tile: 3 x BN x 128 B per K block on top of the 17 KB window)
    if (tst && p.tiles_n == 1 && (221 * 1024 - st_bytes - slab) / (17 * 1024) < 4 && (221 * 1024 - slab) / (17 * 1024) >= 4) {
      tst = false;
      st_bytes = 0;
    }
    p.b_res = p.tiles_n == 1 && (221 * 1024 - st_bytes - slab) / (17 * 1024) >= 4;
    stage_bytes = 17 * 1024 + (p.b_res ? 0 : 3 * b_tile);
    resident_bytes = p.b_res ? slab : 0;
    if ((221 * 1024 - st_bytes - resident_bytes) / stage_bytes < 2) return "lower_conv: row-window stage does not fit (N tile too wide)";
  }
  if (win2) {
    p.a_win = 2;
    p.a_slot = w2_slot;
    p.b_res = 1;
    stage_bytes = w2_slot;
    resident_bytes = w2_slab;
  }
  p.sets = 2;
  int aad_bytes = 0;
  if (aad) {
    // 128-column tiles (c_x = 64): four accumulator stages / warp groups; 256-column tiles: two
    const int hbox = (p.BN >> 1) >> 6;
    p.sets = 4 / hbox;
    // one N tile and 64-channel K chunks: the gamma/beta weights stay resident, only z_attr streams
    const int slab = p.ntaps * p.kc_per_tap * p.BN * p.kc * 2;
    p.b_res = (p.tiles_n == 1 || aad->pair) && p.kc == 64 && slab <= 64 * 1024;
    if (p.b_res) {
      stage_bytes = 128 * p.kc * 2;
      resident_bytes = slab;
    }
    // two h slots per warp group when they still leave three operand slots
    p.h_bufs = 2 * p.sets;
    aad_bytes = p.h_bufs * hbox * 16384;
    if ((221 * 1024 - aad_bytes - resident_bytes) / stage_bytes < 3) {
      p.h_bufs = p.sets;
      aad_bytes = p.h_bufs * hbox * 16384;
    }
  }
  int stages = (221 * 1024 - st_bytes - aad_bytes - resident_bytes) / stage_bytes;
  if (stages > 8) stages = 8;
  static const int max_stages = [] {
    const char* e = getenv("GHOST_MAX_STAGES");  // profiling switch: cap the operand ring depth
    return e ? atoi(e) : 8;
  }();
  if (stages > max_stages) stages = max_stages;
  if (stages < 2) stages = 2;
  p.stages = stages;
  // Short tiles are bound by how fast one warp can issue tcgen05.mma and hand tiles over: split the tiles over
  // two issuers, each with half of the ring - but only when a whole tile fits in its half.  The producer fills
  // the ring in tile order, so a long tile would see just half of the prefetch depth (the fill is latency
  // bound: bytes in flight / ~1.8 us).
  {
    const int nkb = win ? 3 * p.kc_per_tap : p.ntaps * p.kc_per_tap;
    p.dual = !aad && !p.cta2 && p.splits == 1 && stages >= 4 && 2 * nkb <= stages;
  }
  p.tmem_cols = aad ? 512 : pow2_at_least(2 * p.BN);
  // thin-K tiles (a few K blocks): the latency of an epilogue pass (accumulator complete -> all 16 warps done) is a
  // multiple of the tile's MMA time, so the issuers run up to three tiles ahead on four accumulator stages
  static const int sets4_mode = [] {
    const char* e = getenv("GHOST_SETS4");  // A/B switch for profiling
    return e ? atoi(e) : 1;
  }();
  if (!aad && sets4_mode && 4 * p.BN <= 512) {
    p.sets = 4;
    p.tmem_cols = pow2_at_least(4 * p.BN);
  }
  // coalesced epilogue for the last tile of each CTA (igemm.cuh), staged in the idle operand ring
  if (tst) {
    p.tst = 1;
    p.tst_c0 = static_cast<int>(out.off);
    const uint64_t ct = static_cast<uint64_t>(out.sw);
    const uint64_t dims[4] = {ct, static_cast<uint64_t>(Wo), static_cast<uint64_t>(Ho), static_cast<uint64_t>(in.N)};
    const uint64_t str[3] = {ct * 2, static_cast<uint64_t>(Wo) * ct * 2, static_cast<uint64_t>(Ho) * Wo * ct * 2};
    const uint32_t ob[4] = {64u, 32u, 1u, 1u};
    if (const char* e = encode_tmap(&p.tmO, out.p, 4, dims, str, ob, 128)) return e;
  }
  if (!tst && !aad && !win2 && p.splits == 1 && is16(out.dtype) && (p.BN == 128 || p.BN == 256) && w.Cout % (p.BN >> 2) == 0 && !(out.off % 8) &&
      !(out.sw % 8) && !(out.sh % 8) && !(out.sn % 8) &&
      (!epi.res || (!(epi.r_off % 8) && !(epi.r_sw % 8) && !(epi.r_sh % 8) && !(epi.r_sn % 8))) &&
      256 * p.BN <= stages * stage_bytes) {
    static const int stg_mode = [] {
      const char* e = getenv("GHOST_STG");  // A/B switch for profiling: GHOST_STG=0 turns the staged epilogue off
      return e ? (atoi(e) != 0) : 1;
    }();
    p.stg = stg_mode;
  }
  // ... and through the TMA unit instead of the load/store path when the views are plain NHWC tensors (igemm.cuh tte)
  if (p.stg && !stats && !phases && epi.act2 == kAct2None && w.Cout % 64 == 0 && out.off >= 0 && out.off < out.sw &&
      out.sh == static_cast<long long>(Wo) * out.sw && out.sn == static_cast<long long>(Ho) * Wo * out.sw &&
      (!epi.res || (epi.r_off >= 0 && epi.r_off < epi.r_sw && epi.r_sh == static_cast<long long>(Wo) * epi.r_sw &&
                    epi.r_sn == static_cast<long long>(Ho) * Wo * epi.r_sw)) &&
      (epi.out2 ? 2 : 1) * 256 * p.BN <= stages * stage_bytes) {
    static const int tte_mode = [] {
      const char* e = getenv("GHOST_TTE");  // A/B switch for profiling
      return e ? atoi(e) : 1;
    }();
    if (tte_mode) {
      const uint32_t box[4] = {64u, static_cast<uint32_t>(p.bw), static_cast<uint32_t>(p.bh), static_cast<uint32_t>(p.bn)};
      auto nhwc_map = [&](CUtensorMap* m, const void* base, long long ct_) {
        const uint64_t ct = static_cast<uint64_t>(ct_);
        const uint64_t dims[4] = {ct, static_cast<uint64_t>(Wo), static_cast<uint64_t>(Ho), static_cast<uint64_t>(in.N)};
        const uint64_t str[3] = {ct * 2, static_cast<uint64_t>(Wo) * ct * 2, static_cast<uint64_t>(Ho) * Wo * ct * 2};
        return encode_tmap(m, base, 4, dims, str, box, 128);
      };
      if (const char* e = nhwc_map(&p.tmO, out.p, out.sw)) return e;
      if (epi.out2)
        if (const char* e = nhwc_map(&p.tmO2, epi.out2, out.sw)) return e;
      if (epi.res)
        if (const char* e = nhwc_map(&p.tmH, epi.res, epi.r_sw)) return e;
      p.tte = 1;
      p.tst_c0 = static_cast<int>(out.off);
      p.tte_r0 = static_cast<int>(epi.r_off);
    }
  }

  // ---- epilogue
  p.out = out.p;
  p.o_sn = out.sn; p.o_sh = out.sh; p.o_sw = out.sw; p.o_off = out.off;
  p.out_f32 = out.dtype == kF32;
  p.f16 = in.dtype == kF16;
  if (!p.out_f32 && out.dtype != in.dtype) return "lower_conv: 16-bit output type must match the input";
  p.bias = w.bias;
  if (epi.bias9) {
    if (aad || p.splits != 1 || epi.bias9_ld < p.tiles_n * p.BN || epi.bias9_ld % 4)
      return "lower_conv: a per-border-class bias needs a plain linear conv and rows padded to the N tiles";
    p.bias = nullptr;
    p.bias9 = epi.bias9;
    p.bias9_ld = epi.bias9_ld;
  }
  p.err = err_word;
  if (aad) {
    if (!is16(out.dtype)) return "lower_conv: AAD output must be a 16-bit tensor";
    p.mode = kModeAAD;
    p.h = reinterpret_cast<const __nv_bfloat16*>(aad->h);
    p.c_x = aad->c_x;
    p.ncpar = aad->ncpar;
    p.wh_fold = aad->wh_fold;
    p.bh_fold = aad->bh_fold;
    p.mask = aad->mask;
    if (!w.bias) return "lower_conv: AAD needs the gamma/beta bias vector";
    if (aad->c_x % 64) return "lower_conv: AAD c_x must be a multiple of 64";
    p.aad_pair = aad->pair;
    if (aad->pair && (aad->mask || aad->spade || p.tiles_n != 2)) return "lower_conv: paired AAD layers compute their own masks";
    if (p.tiles_n > 1 && !aad->mask && !aad->spade && !aad->pair) return "lower_conv: AAD with c_x > 128 needs the precomputed mask";
    p.aad_spade = aad->spade;
    p.aad_slope = aad->slope;
    p.aad_post = aad->post;
    p.aad_no_h = aad->no_h;
    if (aad->no_h && !aad->spade) return "lower_conv: no_h needs the SPADE variant of the epilogue";
    if (out.sw != aad->c_x || out.off != 0 || out.sh != static_cast<long long>(Wo) * aad->c_x ||
        out.sn != static_cast<long long>(Ho) * Wo * aad->c_x)
      return "lower_conv: AAD output must be a dense NHWC tensor of c_x channels";
    {
      const uint64_t cx = static_cast<uint64_t>(aad->c_x);
      const uint64_t dims[4] = {cx, static_cast<uint64_t>(Wo), static_cast<uint64_t>(Ho), static_cast<uint64_t>(in.N)};
      const uint64_t str[3] = {cx * eb, static_cast<uint64_t>(Wo) * cx * eb, static_cast<uint64_t>(Ho) * Wo * cx * eb};
      const uint64_t ch = aad->pair ? cx / 2 : cx;  // channels of h
      const uint64_t hdims[4] = {ch, dims[1], dims[2], dims[3]};
      const uint64_t hstr[3] = {ch * eb, static_cast<uint64_t>(Wo) * ch * eb, static_cast<uint64_t>(Ho) * Wo * ch * eb};
      const uint32_t box[4] = {64u, static_cast<uint32_t>(p.bw), static_cast<uint32_t>(p.bh), static_cast<uint32_t>(p.bn)};
      // no_h: the map is never used by the kernel; it is encoded over the output tensor to stay valid
      if (const char* e = encode_tmap(&p.tmH, aad->no_h ? out.p : aad->h, 4, aad->no_h ? dims : hdims, aad->no_h ? str : hstr, box, 128))
        return e;
      // store box = the 32 consecutive tile rows (order n, h, w) one epilogue warp owns
      uint32_t ob[4] = {64u, 0, 1, 1};
      const int bw = p.bw, bh = p.bh, bn = p.bn;
      if (bw >= 32) {
        if (bw % 32) return "lower_conv: AAD tiles need a row length that is a multiple of 32 (or a divisor of it)";
        ob[1] = 32;
      } else if (bw * bh >= 32) {
        if (32 % bw || bh % (32 / bw)) return "lower_conv: AAD tile rows do not split into 32-row boxes";
        ob[1] = bw;
        ob[2] = 32 / bw;
      } else {
        if (32 % (bw * bh)) return "lower_conv: AAD tile images do not split into 32-row boxes";
        ob[1] = bw;
        ob[2] = bh;
        ob[3] = 32 / (bw * bh);
        if (static_cast<int>(ob[3]) > bn && bn * bw * bh < 32) ob[3] = bn;  // fewer than 32 rows in the whole tile
      }
      if (const char* e = encode_tmap(&p.tmO, out.p, 4, dims, str, ob, 128)) return e;

    }
  } else {
    p.mode = kModeLinear;
    for (int ph = 0; phases && ph < 4; ++ph) {
      // relative to phase 0's view, minus the stacked column offset the epilogue adds (ph * channels per phase)
      p.ph_o[ph] = phases->o_off[ph] - out.off - static_cast<long long>(ph) * cout_t;
      p.ph_r[ph] = phases->r_off[ph] - epi.r_off - static_cast<long long>(ph) * cout_t;
    }
    p.act1 = epi.act1; p.slope = epi.slope; p.slope_c = epi.slope_c;
    p.res = reinterpret_cast<const __nv_bfloat16*>(epi.res);
    p.r_sn = epi.r_sn; p.r_sh = epi.r_sh; p.r_sw = epi.r_sw; p.r_off = epi.r_off;
    p.act2 = epi.act2;
    p.out2 = epi.out2;
    p.scale2 = epi.scale2;
    p.shift2 = epi.shift2;
    if (epi.out2 && (!is16(out.dtype) || w.Cout % 16)) return "lower_conv: the second output needs a 16-bit tensor and Cout % 16 == 0";
    if (is16(out.dtype) && w.Cout % 16 == 0 && ((out.off % 8) || (out.sw % 8) || (out.sh % 8) || (out.sn % 8)))
      return "lower_conv: bf16 output view must be 16-byte aligned";
    static const int dbg = [] {
      const char* e = getenv("GHOST_LOWER_DEBUG");  // prints the kernel variant picked for every lowered conv
      return e ? atoi(e) : 0;
    }();
    if (dbg)
      fprintf(stderr, "lower_conv: N=%d %dx%d Cin=%d Cout=%d taps=%d%s -> %s BN=%d tiles=%dx%d kb=%d stages=%d%s\n", in.N, Ho,
              Wo, Cin, w.Cout, taps.n, phases ? " x4 phases" : "",
              p.splits > 1 ? "split-K" : (p.cta2 ? "CTA pair" : (p.a_win ? "row window" : "one CTA")), p.BN,
              p.tiles_w * p.tiles_h * p.tiles_nb, p.tiles_n, p.ntaps * p.kc_per_tap / p.splits, p.stages,
              p.splits > 1 ? (std::string(" splits=") + std::to_string(p.splits)).c_str() : "");
    if (p.splits > 1) {
      p.ws_cols = p.tiles_n * p.BN;
      op->ws_bytes = static_cast<size_t>(p.splits) * p.tiles_w * p.tiles_h * p.tiles_nb * 128 * p.ws_cols * sizeof(float);
    }
  }
  return nullptr;
}

// ------------------------------------------------------------------------------------------------ conv chains
bool chain_compatible(const LaunchOp& first, const LaunchOp& next) {
  static const int on = [] {
    const char* e = getenv("GHOST_CHAIN");  // A/B switch for profiling
    return e ? atoi(e) : 1;
  }();
  if (!on || first.kind != LaunchOp::kIgemm || next.kind != LaunchOp::kIgemm) return false;
  const IgemmParams &a = first.ig, &b = next.ig;
  auto plain = [](const IgemmParams& p) {
    // (resident weights only in tile-window mode, where the chain refills the slab under the last epilogue of a step)
    return p.cta2 && p.mode == kModeLinear && p.splits == 1 && !p.st_mode && !p.tst && p.nphase <= 1 && !p.out_f32 &&
           !p.dual && p.kc == 64 && p.act2 == kAct2None && ((!p.b_res && !p.a_win) || (p.b_res && p.a_win == 2));
  };
  if (!plain(a) || !plain(b)) return false;
  if (a.a_rank != b.a_rank || a.bw != b.bw || a.bh != b.bh || a.bn != b.bn || a.tiles_w != b.tiles_w ||
      a.tiles_h != b.tiles_h || a.tiles_nb != b.tiles_nb || a.tiles_n != b.tiles_n || a.Nimg != b.Nimg || a.Ho != b.Ho ||
      a.Wo != b.Wo || a.BN != b.BN || a.n_valid != b.n_valid || a.ntaps != b.ntaps || a.kc_per_tap != b.kc_per_tap ||
      a.stages != b.stages || a.tmem_cols != b.tmem_cols || a.sets != b.sets || a.stg != b.stg || a.f16 != b.f16 ||
      a.tte != b.tte || a.a_win != b.a_win || a.b_res != b.b_res || a.a_slot != b.a_slot ||
      a.err != b.err)
    return false;
  for (int t = 0; t < a.ntaps; ++t)
    if (a.tap_c[t] != b.tap_c[t] || a.tap_w[t] != b.tap_w[t] || a.tap_p[t] != b.tap_p[t] || a.tap_h[t] != b.tap_h[t]) return false;
  return true;
}

const char* make_chain(LaunchOp* out, const std::vector<LaunchOp>& steps, void* dev_params, bool* fused) {
  *fused = false;
  if (steps.size() < 2) return nullptr;
  for (size_t i = 1; i < steps.size(); ++i)
    if (!chain_compatible(steps[0], steps[i])) return "make_chain: steps of different geometry";
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
    return "make_chain: no device";
  const int clusters = sms / 2;
  IgemmParams g = steps[0].ig;
  // a group = the M-tile pairs of bn whole images (x all N tiles); an odd tile count per image gets one tile of rows
  // past the image: its loads are zero-filled and its rows never stored
  int tiles_h = g.tiles_h;
  if ((g.tiles_w * tiles_h) & 1) ++tiles_h;
  const int spg = (g.tiles_w * tiles_h / 2) * g.tiles_n;
  const int ngroups = g.tiles_nb;
  const long long rounds_chain = static_cast<long long>((ngroups + clusters - 1) / clusters) * spg;
  const long long supers_plain = static_cast<long long>((g.tiles_w * g.tiles_h * g.tiles_nb + 1) / 2) * g.tiles_n;
  // Separate launches overlap under programmatic dependent launch: the CTAs of the next conv start on the SMs that
  // have no tile in the last, partly filled round, so a plain walk costs supers / clusters rounds rather than the
  // rounded-up count (measured on iresnet layer 2 at B = 64: 224 super tiles on 74 clusters = 3.03 rounds as
  // separate launches, 4 as a chain of 64 one-image groups - the chain was 0.07 ms slower per forward).
  double plain = static_cast<double>(supers_plain) / clusters;
  if (plain < 1.0) plain = 1.0;
  if (static_cast<double>(rounds_chain) > 1.15 * plain) return nullptr;  // not worth it: *fused stays false
  // room for the chain's shared-memory blocks behind the ring
  int stages = g.stages;
  {
    IgemmParams probe = g;
    probe.chain_n = static_cast<int>(steps.size());
    const int budget = 221 * 1024;
    while (stages >= 3 && (probe.stages = stages, igemm_smem_bytes(probe)) > budget) --stages;
    probe.stages = stages;
    if (igemm_smem_bytes(probe) > budget) return nullptr;
    const int stage_bytes = g.a_win == 2 ? g.a_slot : 128 * 128 + (g.BN >> 1) * 128;
    bool two = false;
    for (const LaunchOp& st : steps) two = two || st.ig.out2 != nullptr;
    if (g.stg && (g.tte && two ? 2 : 1) * 256 * g.BN > stages * stage_bytes) return nullptr;
  }
  std::vector<IgemmParams> host(steps.size());
  for (size_t i = 0; i < steps.size(); ++i) {
    host[i] = steps[i].ig;
    host[i].tiles_h = tiles_h;
    host[i].stages = stages;
    host[i].chain = static_cast<const IgemmParams*>(dev_params);
    host[i].chain_n = static_cast<int>(steps.size());
    host[i].chain_spg = spg;
  }
  if (cudaMemcpy(dev_params, host.data(), host.size() * sizeof(IgemmParams), cudaMemcpyHostToDevice) != cudaSuccess)
    return "make_chain: parameter upload failed";
  *out = steps[0];
  out->ig = host[0];
  out->name = steps.front().name + ".." + steps.back().name;
  out->flops = 0.0;
  for (const LaunchOp& s : steps) out->flops += s.flops;
  *fused = true;
  return nullptr;
}

}  // namespace ghost
