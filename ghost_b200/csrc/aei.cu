// AEI_Net plan: weight packing and the op list of one forward.
// Reference: network/AEI_Net.py:44-156 (MLAttrEncoder, AADGenerator, AEI_Net),
// network/AADLayer.py:5-80 (AADLayer, AAD_ResBlk).
#include <math.h>
#include <stdlib.h>

#include <map>

#include "net.cuh"

namespace ghost {

namespace {

struct BlockSpec {
  int cin, cout, cattr;
};
const BlockSpec kUnetBlocks[8] = {{1024, 1024, 1024}, {1024, 1024, 2048}, {1024, 1024, 1024}, {1024, 512, 512},
                                  {512, 256, 256},    {256, 128, 128},    {128, 64, 64},      {64, 3, 64}};
const BlockSpec kLinkBlocks[8] = {{1024, 1024, 1024}, {1024, 1024, 1024}, {1024, 1024, 512}, {1024, 512, 256},
                                  {512, 256, 128},    {256, 128, 64},     {128, 64, 32},     {64, 3, 32}};
// AAD blocks whose shortcut layer and first main layer run as one paired launch (igemm.cuh aad_pair): tensor-core
// mode, masks computed inside the epilogue (c_x <= 128).  GHOST_AAD_PAIR=0 keeps the two launches (A/B switch).
bool aad_pairable(bool bf, int c_x) {
  static const int on = [] {
    const char* e = getenv("GHOST_AAD_PAIR");
    return e ? (atoi(e) != 0) : 1;
  }();
  // measured on B200 (B = 64): c_x = 128 (block 7) 0.364 -> 0.294 ms for the two layers; c_x = 64 (block 8) gains
  // 0.04 ms in the AAD launch but the two convs that then read 64 of every 128 channels lose 0.06 ms - not paired
  return on && bf && c_x == 128;
}
const int kEncC[8] = {3, 32, 64, 128, 256, 512, 1024, 1024};
// deconv i (1-based): Cin / Cout
const int kUnetDecIn[6] = {1024, 2048, 1024, 512, 256, 128};
const int kLinkDecIn[6] = {1024, 1024, 512, 256, 128, 64};
const int kDecOut[6] = {1024, 512, 256, 128, 64, 32};
// ConvTranspose2d k4s2p1 kernel index used by tap t of output phase r (see Taps::deconv_phase)
const int kDeconvK[2][2] = {{1, 3}, {0, 2}};

std::string S(const char* fmt, int a, int b = 0) {
  char buf[128];
  snprintf(buf, sizeof(buf), fmt, a, b);
  return buf;
}

void expect_shape(const HostTensor& t, std::initializer_list<int64_t> shape, const std::string& key) {
  if (t.shape != std::vector<int64_t>(shape)) fail(-2, "unexpected shape for " + key);
}

// BatchNorm2d (eval) -> per-channel scale / shift
void bn_fold(const HostTensor& g, const HostTensor& b, const HostTensor& rm, const HostTensor& rv,
             std::vector<float>* scale, std::vector<float>* shift) {
  const size_t C = g.data.size();
  scale->resize(C);
  shift->resize(C);
  for (size_t c = 0; c < C; ++c) {
    const float s = g.data[c] / sqrtf(rv.data[c] + 1e-5f);
    (*scale)[c] = s;
    (*shift)[c] = b.data[c] - rm.data[c] * s;
  }
}

}  // namespace

void Net::finalize_aei() {
  if (backbone != "unet" && backbone != "linknet") fail(-1, "backbone must be 'unet' or 'linknet'");
  if (num_blocks < 1 || num_blocks > 4) fail(-1, "num_blocks must be in [1,4]");
  if (c_id % 32) fail(-1, "c_id must be a multiple of 32");
  const bool bf = is16(act_dtype);  // tensor-core mode: bf16 or fp16 storage
  const bool unet = backbone == "unet";
  const BlockSpec* blocks = unet ? kUnetBlocks : kLinkBlocks;

  // ---- encoder convs: Conv2d k4 s2 p1 (no bias) + BN(eval) folded (AEI_Net.py:19-24)
  for (int i = 1; i <= 7; ++i) {
    const int Ci = kEncC[i - 1], Co = kEncC[i];
    const std::string k = S("encoder.conv%d", i);
    const HostTensor& w = W(k + ".0.weight");
    expect_shape(w, {Co, Ci, 4, 4}, k + ".0.weight");
    std::vector<float> sc, sh;
    bn_fold(W(k + ".1.weight"), W(k + ".1.bias"), W(k + ".1.running_mean"), W(k + ".1.running_var"), &sc, &sh);
    std::vector<float> rows(static_cast<size_t>(Co) * 16 * Ci);
    for (int co = 0; co < Co; ++co)
      for (int ci = 0; ci < Ci; ++ci)
        for (int t = 0; t < 16; ++t)
          rows[(static_cast<size_t>(co) * 16 + t) * Ci + ci] = w.data[(static_cast<size_t>(co) * Ci + ci) * 16 + t] * sc[co];
    const bool tc = bf && Ci % 32 == 0;
    packed_[S("conv%d", i)] = pack(rows, Co, 16 * Ci, &sh, !tc, tc);
  }
  // ---- encoder deconvs: ConvTranspose2d k4 s2 p1 (no bias) + BN, as 4 output phases of 2x2 taps
  for (int i = 1; i <= 6; ++i) {
    const int Ci = unet ? kUnetDecIn[i - 1] : kLinkDecIn[i - 1], Co = kDecOut[i - 1];
    const std::string k = S("encoder.deconv%d", i);
    const HostTensor& w = W(k + ".deconv.weight");
    expect_shape(w, {Ci, Co, 4, 4}, k + ".deconv.weight");
    std::vector<float> sc, sh;
    bn_fold(W(k + ".bn.weight"), W(k + ".bn.bias"), W(k + ".bn.running_mean"), W(k + ".bn.running_var"), &sc, &sh);
    // tensor-core modes: the four phase matrices stacked along the rows (one launch, igemm.cuh `nphase`)
    std::vector<float> all_rows, all_sh;
    for (int r = 0; r < 2; ++r)
      for (int s = 0; s < 2; ++s) {
        std::vector<float> rows(static_cast<size_t>(Co) * 4 * Ci);
        for (int co = 0; co < Co; ++co)
          for (int th = 0; th < 2; ++th)
            for (int tw = 0; tw < 2; ++tw) {
              const int kh = kDeconvK[r][th], kw = kDeconvK[s][tw];
              float* dst = &rows[(static_cast<size_t>(co) * 4 + th * 2 + tw) * Ci];
              for (int ci = 0; ci < Ci; ++ci)
                dst[ci] = w.data[((static_cast<size_t>(ci) * Co + co) * 4 + kh) * 4 + kw] * sc[co];
            }
        if (bf && Ci % 64 == 0) {
          all_rows.insert(all_rows.end(), rows.begin(), rows.end());
          all_sh.insert(all_sh.end(), sh.begin(), sh.end());
        } else {
          packed_[S("deconv%d.p%d", i, r * 2 + s)] = pack(rows, Co, 4 * Ci, &sh, !bf, bf);
        }
      }
    if (!all_rows.empty()) packed_[S("deconv%d.all", i)] = pack(all_rows, 4 * Co, 4 * Ci, &all_sh, false, true);
  }

  // ---- z_id GEMM: up1 (ConvTranspose2d c_id->1024 k2 on a 1x1 map, AEI_Net.py:101,123) and every
  //      AADLayer.fc1 / fc2 (AADLayer.py:15-16) share one [zfc_cols][c_id] matrix
  std::vector<float> zrows, zbias;
  auto push_fc = [&](const HostTensor& w, const HostTensor& b, int rows_expected, const std::string& key) {
    expect_shape(w, {rows_expected, c_id}, key);
    zrows.insert(zrows.end(), w.data.begin(), w.data.end());
    zbias.insert(zbias.end(), b.data.begin(), b.data.end());
  };
  {
    const HostTensor& w = W("generator.up1.weight");
    const HostTensor& b = W("generator.up1.bias");
    expect_shape(w, {c_id, 1024, 2, 2}, "generator.up1.weight");
    zrows.resize(static_cast<size_t>(4096) * c_id);
    zbias.resize(4096);
    for (int pos = 0; pos < 4; ++pos)
      for (int co = 0; co < 1024; ++co) {
        for (int ci = 0; ci < c_id; ++ci)
          zrows[(static_cast<size_t>(pos) * 1024 + co) * c_id + ci] = w.data[(static_cast<size_t>(ci) * 1024 + co) * 4 + pos];
        zbias[pos * 1024 + co] = b.data[co];
      }
  }
  std::map<std::string, std::vector<float>> gb_rows, gb_bias;  // per layer, for the paired launches below
  auto add_aad = [&](const std::string& name, const std::string& key, int c_x, int c_attr) {
    AadInfo info;
    info.name = name;
    info.c_x = c_x;
    info.c_attr = c_attr;
    info.off_gamma = static_cast<int>(zbias.size());
    push_fc(W(key + ".fc1.weight"), W(key + ".fc1.bias"), c_x, key + ".fc1.weight");
    info.off_beta = static_cast<int>(zbias.size());
    push_fc(W(key + ".fc2.weight"), W(key + ".fc2.bias"), c_x, key + ".fc2.weight");
    aad_[name] = info;
    const HostTensor& w1 = W(key + ".conv1.weight");
    const HostTensor& w2 = W(key + ".conv2.weight");
    const HostTensor& b1 = W(key + ".conv1.bias");
    const HostTensor& b2 = W(key + ".conv2.bias");
    expect_shape(w1, {c_x, c_attr, 1, 1}, key + ".conv1.weight");
    expect_shape(w2, {c_x, c_attr, 1, 1}, key + ".conv2.weight");
    std::vector<float> rows(static_cast<size_t>(2) * c_x * c_attr), bias(2 * c_x);
    // bf16: rows interleaved per N tile of cn channels: [gamma(cn) | beta(cn)] so that one accumulator
    // tile holds both halves of the same channels; fp32: plain [gamma(c_x) | beta(c_x)].
    const int cn = bf ? (c_x < 128 ? c_x : 128) : c_x;
    for (int c = 0; c < c_x; ++c) {
      const int t = c / cn, j = c % cn;
      const size_t rg = static_cast<size_t>(t) * 2 * cn + j, rb = rg + cn;
      memcpy(&rows[rg * c_attr], &w1.data[static_cast<size_t>(c) * c_attr], sizeof(float) * c_attr);
      memcpy(&rows[rb * c_attr], &w2.data[static_cast<size_t>(c) * c_attr], sizeof(float) * c_attr);
      bias[rg] = b1.data[c];
      bias[rb] = b2.data[c];
    }
    const bool tc = bf && c_attr % 32 == 0;
    if (bf && !tc) fail(-1, "AAD attribute channels must be a multiple of 32");
    packed_[name + ".gb"] = pack(rows, 2 * c_x, c_attr, &bias, !bf, bf);
    gb_rows[name] = rows;
    gb_bias[name] = bias;
    const HostTensor& wh = W(key + ".conv_h.weight");
    expect_shape(wh, {1, c_x, 1, 1}, key + ".conv_h.weight");
    vecs_[name + ".wh"] = wpool_.upload(wh.data);
    scalars_[name + ".bh"] = W(key + ".conv_h.bias").data[0];
  };
  auto add_conv3 = [&](const std::string& name, const std::string& key, int Ci, int Co) {
    const HostTensor& w = W(key);
    expect_shape(w, {Co, Ci, 3, 3}, key);
    std::vector<float> rows(static_cast<size_t>(Co) * 9 * Ci);
    for (int co = 0; co < Co; ++co)
      for (int ci = 0; ci < Ci; ++ci)
        for (int t = 0; t < 9; ++t)
          rows[(static_cast<size_t>(co) * 9 + t) * Ci + ci] = w.data[(static_cast<size_t>(co) * Ci + ci) * 9 + t];
    packed_[name] = pack(rows, Co, 9 * Ci, nullptr, !bf, bf);
  };
  for (int k = 1; k <= 8; ++k) {
    const BlockSpec& b = blocks[k - 1];
    for (int i = 0; i < num_blocks; ++i) {
      const int co = i < num_blocks - 1 ? b.cin : b.cout;
      add_aad(S("blk%d.aad%d", k, i), S("generator.AADBlk%d.add_blocks.%d", k, 3 * i), b.cin, b.cattr);
      add_conv3(S("blk%d.conv%d", k, i), S("generator.AADBlk%d.add_blocks.%d.weight", k, 3 * i + 2), b.cin, co);
    }
    if (b.cin != b.cout) {
      add_aad(S("blk%d.aadsc", k), S("generator.AADBlk%d.last_add_block.0", k), b.cin, b.cattr);
      add_conv3(S("blk%d.convsc", k), S("generator.AADBlk%d.last_add_block.2.weight", k), b.cin, b.cout);
      if (aad_pairable(bf, b.cin)) {
        // the shortcut layer and the first main layer read the same h and z_attr: one launch, two N tiles
        std::vector<float> rows = gb_rows.at(S("blk%d.aadsc", k)), bias = gb_bias.at(S("blk%d.aadsc", k));
        const std::vector<float>&r0 = gb_rows.at(S("blk%d.aad0", k)), &b0 = gb_bias.at(S("blk%d.aad0", k));
        rows.insert(rows.end(), r0.begin(), r0.end());
        bias.insert(bias.end(), b0.begin(), b0.end());
        packed_[S("blk%d.aadpair.gb", k)] = pack(rows, 4 * b.cin, b.cattr, &bias, false, true);
      }
    }
  }
  zfc_cols_ = static_cast<int>(zbias.size());
  packed_["zfc"] = pack(zrows, zfc_cols_, c_id, &zbias, !bf, bf);
}


void Net::build_aei(ExecPlan& p) {
  const int B = p.B;
  const int dt = act_dtype;
  const bool bf = is16(dt);  // tensor-core mode: bf16 or fp16 storage
  const bool unet = backbone == "unet";
  const BlockSpec* blocks = unet ? kUnetBlocks : kLinkBlocks;
  int* err = p.err_word;

  auto new_act = [&](int H, int W, int C) {
    Act a;
    a.N = B; a.H = H; a.W = W; a.C = C; a.dtype = dt;
    a.p = p.pool.alloc(a.bytes());
    return a;
  };
  auto add_conv = [&](const std::string& name, const Act& in, int c0, int Cin, const Taps& taps, int Ho, int Wo,
                      const std::string& wname, const OutView& out, const Epilogue& epi, const AadArgs* aad,
                      const DeconvPhases* phases = nullptr, float** stats_parts = nullptr, int* stats_P = nullptr) {
    auto it = packed_.find(wname);
    if (it == packed_.end()) fail(-2, "internal: packed weights missing for " + wname);
    LaunchOp lop;
    lop.name = name;
    check_msg(lower_conv(&lop, in, c0, Cin, taps, Ho, Wo, it->second, out, epi, !bf, err, aad, 0, phases), name);
    if (lop.ws_bytes) lop.set_workspace(p.pool.alloc(lop.ws_bytes));  // split-K partial tiles
    if (lop.st_floats) {                                               // fused InstanceNorm statistics of the output
      float* parts = p.pool.alloc_n<float>(lop.st_floats);
      lop.set_stats(parts);
      if (stats_parts) *stats_parts = parts;
      if (stats_P) *stats_P = lop.ig.st_P;
    }
    Op op;
    op.name = name;
    op.flops = lop.flops;
    op.cls = lop.kind == LaunchOp::kSimt ? kOpConvSimt : (aad ? kOpAadTC : kOpConvTC);
    if (aad)  // apply read of h + z_attr read + write (gamma/beta never touch HBM); a pair reads h once
      op.bytes = static_cast<double>(B) * Ho * Wo * ((aad->pair ? 1.5 : 2.0) * aad->c_x + Cin) * 2.0;
    op.run = [lop](cudaStream_t s) { return lop.launch(s); };
    p.ops.push_back(std::move(op));
  };
  auto add_op = [&](const std::string& name, std::function<const char*(cudaStream_t)> fn, int cls = kOpMisc,
                    double bytes = 0.0) {
    Op op;
    op.name = name;
    op.run = std::move(fn);
    op.cls = cls;
    op.bytes = bytes;
    p.ops.push_back(std::move(op));
  };
  const double esz = bf ? 2.0 : 4.0;

  // ------------------------------------------------------------------ buffers
  p.x_in = new_act(256, 256, 3);
  p.zid = new_act(1, 1, c_id);
  p.zid.H = p.zid.W = 1;
  // attribute maps.  unet: z_attr(k) is the concat buffer [deconv out | skip] (AEI_Net.py:41), the
  // encoder convs write their feature map straight into the skip half.
  Act z[9];
  Act feat[8];   // linknet only (separate skip tensors)
  int featoff[8] = {0};
  const Act* featbuf[8] = {nullptr};
  {
    int zc[9];
    for (int k = 1; k <= 8; ++k) zc[k] = blocks[k - 1].cattr;
    for (int k = 1; k <= 8; ++k) {
      const int hw = 2 << (k - 1);  // 2,4,...,256
      z[k] = new_act(hw, hw, zc[k]);
      p.attr[k - 1] = z[k];
    }
    // feat_i (i = 1..7), spatial 256 >> i, channels kEncC[i]
    for (int i = 1; i <= 7; ++i) {
      if (i == 7) {
        featbuf[i] = &z[1];
        featoff[i] = 0;
      } else if (unet) {
        featbuf[i] = &z[8 - i];
        featoff[i] = z[8 - i].C / 2;
      } else {
        feat[i] = new_act(256 >> i, 256 >> i, kEncC[i]);
        featbuf[i] = &feat[i];
        featoff[i] = 0;
      }
    }
  }

  // ------------------------------------------------------------------ encoder (AEI_Net.py:72-95)
  Epilogue leaky;
  leaky.act1 = kActLeaky;
  leaky.slope = 0.1f;
  for (int i = 1; i <= 7; ++i) {
    const Act& in = i == 1 ? p.x_in : *featbuf[i - 1];
    const int c0 = i == 1 ? 0 : featoff[i - 1];
    const int hw = 256 >> i;
    add_conv(S("enc.conv%d", i), in, c0, kEncC[i - 1], Taps::square(4, 1, 2), hw, hw, S("conv%d", i),
             OutView::dense(*featbuf[i], featoff[i]), leaky, nullptr);
  }
  for (int j = 1; j <= 6; ++j) {
    const Act& in = z[j];
    const Act& out = z[j + 1];
    if (packed_.count(S("deconv%d.all", j))) {
      // one launch for the four output phases (24 -> 6 launches, four times the tiles per launch)
      DeconvPhases dp;
      Epilogue e = leaky;
      const Act* sk = unet ? nullptr : featbuf[7 - j];
      for (int ph = 0; ph < 4; ++ph) {
        const int r = ph >> 1, s = ph & 1;
        dp.taps[ph] = Taps::deconv_phase(r, s);
        dp.o_off[ph] = OutView::deconv_phase(out, r, s, 0).off;
        if (sk) dp.r_off[ph] = OutView::deconv_phase(*sk, r, s, featoff[7 - j]).off;
      }
      if (sk) {  // linknet: x + skip (AEI_Net.py:38-39)
        const OutView rv = OutView::deconv_phase(*sk, 0, 0, featoff[7 - j]);
        e.res = sk->p;
        e.r_sn = rv.sn; e.r_sh = rv.sh; e.r_sw = rv.sw; e.r_off = rv.off;
      }
      add_conv(S("enc.deconv%d", j), in, 0, in.C, dp.taps[0], in.H, in.W, S("deconv%d.all", j),
               OutView::deconv_phase(out, 0, 0, 0), e, nullptr, &dp);
      continue;
    }
    for (int ph = 0; ph < 4; ++ph) {
      const int r = ph >> 1, s = ph & 1;
      Epilogue e = leaky;
      if (!unet) {  // linknet: x + skip (AEI_Net.py:38-39)
        const Act& sk = *featbuf[7 - j];
        const OutView rv = OutView::deconv_phase(sk, r, s, featoff[7 - j]);
        e.res = sk.p;
        e.r_sn = rv.sn; e.r_sh = rv.sh; e.r_sw = rv.sw; e.r_off = rv.off;
      }
      add_conv(S("enc.deconv%d.p%d", j, ph), in, 0, in.C, Taps::deconv_phase(r, s), in.H, in.W,
               S("deconv%d.p%d", j, ph), OutView::deconv_phase(out, r, s, 0), e, nullptr);
    }
  }
  {
    const Act src = z[7], dst = z[8];
    add_op("enc.z_attr8.upsample", [=](cudaStream_t s) {
      return upsample2x_launch(src.p, dst.p, dt, B, src.H, src.W, src.C, dst.C, 0, s);
    }, kOpUpsample, 5.0 * B * src.H * src.W * src.C * esz);
  }
  encoder_ops = p.ops.size();

  // ------------------------------------------------------------------ generator (AEI_Net.py:122-139)
  const int ldz = zfc_cols_;
  float* zfc = p.pool.alloc_n<float>(static_cast<size_t>(B) * ldz);
  {
    OutView zo;
    zo.p = zfc; zo.sn = ldz; zo.sh = 0; zo.sw = 0; zo.off = 0;
    zo.dtype = bf ? kF32 : dt;
    add_conv("gen.zid_fc", p.zid, 0, c_id, Taps::point(), 1, 1, "zfc", zo, Epilogue(), nullptr);
  }
  Act m = new_act(2, 2, 1024);
  p.named["m1"] = m;
  {
    const Act mm = m;
    add_op("gen.up1", [=](cudaStream_t s) { return zfc_to_act_launch(zfc, ldz, 0, 4096, nullptr, mm.p, dt, B, s); });
  }

  float* sums_next = nullptr;  // statistics of m written by the previous block's fused upsample
  int parts_next = 0;          // ... and the number of parts per image it writes
  for (int k = 1; k <= 8; ++k) {
    const BlockSpec& b = blocks[k - 1];
    const int hw = 2 << (k - 1);
    const int HW = hw * hw;
    const Act& za = z[k];
    const int parts_in = in_stats_parts(B, HW, b.cin);   // parts of in_stats_launch on a tensor of this block
    const int parts_m = sums_next ? parts_next : parts_in;  // parts of the statistics of the block input m
    float* sums_m = sums_next;
    if (!sums_m) {
      sums_m = p.pool.alloc_n<float>(static_cast<size_t>(B) * (2 * parts_in + 1) * b.cin);
      const Act mm = m;
      add_op(S("blk%d.stats_m", k), [=](cudaStream_t s) { return in_stats_launch(mm.p, dt, B, HW, mm.C, sums_m, s); },
             kOpStats, static_cast<double>(B) * HW * mm.C * esz);
    }
    Act a = new_act(hw, hw, b.cin);
    Act gb;  // fp32 mode only: gamma|beta of the attribute branch
    if (!bf) gb = new_act(hw, hw, 2 * b.cin);

    // one AADLayer + ReLU: input h (with its statistics) -> a
    struct AadPre {  // per-layer tables written by the fold (+ the mask when it was computed ahead)
      float4* ncpar;
      float *whf, *bhf, *mask;
    };
    auto aad_fold = [&](const std::string& lname, const float* sums, int sums_P) {
      const AadInfo& info = aad_.at(lname);
      AadPre q;
      q.ncpar = p.pool.alloc_n<float4>(static_cast<size_t>(B) * b.cin);
      q.whf = p.pool.alloc_n<float>(static_cast<size_t>(B) * b.cin);
      q.bhf = p.pool.alloc_n<float>(B);
      q.mask = nullptr;
      float* fsc = p.pool.alloc_n<float>(aad_fold_scratch_floats(B, b.cin));  // zeroed by the pool
      const float* wh = vecs_.at(lname + ".wh");
      const float bh = scalars_.at(lname + ".bh");
      const int cx = b.cin, og = info.off_gamma, ob = info.off_beta;
      float4* ncpar = q.ncpar;
      float *whf = q.whf, *bhf = q.bhf;
      add_op(lname + ".fold", [=](cudaStream_t s) {
        return aad_fold_launch(sums, sums_P, B, HW, cx, zfc, ldz, og, ob, wh, bh, ncpar, whf, bhf, fsc, s);
      });
      return q;
    };
    // two layers that normalise the same tensor (shortcut + first main layer of a block): one fold launch
    auto aad_fold_two = [&](const std::string& la, const std::string& lb, const float* sums, int sums_P, AadPre* qa,
                            AadPre* qb) {
      AadFoldSet fs[2];
      AadPre* q[2] = {qa, qb};
      const std::string* ln[2] = {&la, &lb};
      for (int L = 0; L < 2; ++L) {
        const AadInfo& info = aad_.at(*ln[L]);
        q[L]->ncpar = p.pool.alloc_n<float4>(static_cast<size_t>(B) * b.cin);
        q[L]->whf = p.pool.alloc_n<float>(static_cast<size_t>(B) * b.cin);
        q[L]->bhf = p.pool.alloc_n<float>(B);
        q[L]->mask = nullptr;
        fs[L].off_gamma = info.off_gamma;
        fs[L].off_beta = info.off_beta;
        fs[L].w_h = vecs_.at(*ln[L] + ".wh");
        fs[L].b_h = scalars_.at(*ln[L] + ".bh");
        fs[L].ncpar = q[L]->ncpar;
        fs[L].wh_fold = q[L]->whf;
        fs[L].bh_fold = q[L]->bhf;
        fs[L].scratch = p.pool.alloc_n<float>(aad_fold_scratch_floats(B, b.cin));  // zeroed by the pool
      }
      const AadFoldSet f0 = fs[0], f1 = fs[1];
      const int cx = b.cin;
      add_op(la + "+" + lb.substr(lb.rfind('.') + 1) + ".fold", [=](cudaStream_t s) {
        return aad_fold2_launch(sums, sums_P, B, HW, cx, zfc, ldz, f0, &f1, s);
      });
    };
    auto aad_layer = [&](const std::string& lname, const Act& h, const float* sums, int sums_P,
                         const AadPre* pre = nullptr) {
      const AadPre q = pre ? *pre : aad_fold(lname, sums, sums_P);
      float4* ncpar = q.ncpar;
      float *whf = q.whf, *bhf = q.bhf;
      const int cx = b.cin;
      if (bf) {
        AadArgs args;
        args.h = h.p; args.c_x = cx; args.ncpar = ncpar; args.wh_fold = whf; args.bh_fold = bhf;
        if (cx > 128) {  // several N tiles per pixel: the sigmoid mask is computed once, ahead of the GEMM
          float* mask = q.mask;
          if (!mask) {
            mask = p.pool.alloc_n<float>(static_cast<size_t>(B) * HW);
            const Act hh = h;
            add_op(lname + ".mask", [=](cudaStream_t s) { return aad_mask_launch(hh.p, B, HW, cx, whf, bhf, mask, s, dt); },
                   kOpStats, static_cast<double>(B) * HW * cx * esz);
          }
          args.mask = mask;
        }
        add_conv(lname + ".gemm_blend", za, 0, za.C, Taps::point(), hw, hw, lname + ".gb", OutView::dense(a),
                 Epilogue(), &args);
      } else {
        add_conv(lname + ".gamma_beta", za, 0, za.C, Taps::point(), hw, hw, lname + ".gb", OutView::dense(gb),
                 Epilogue(), nullptr);
        const Act hh = h, aa = a, gg = gb;
        add_op(lname + ".blend", [=](cudaStream_t s) {
          return aad_apply_launch(hh.p, gg.p, aa.p, dt, B, HW, cx, ncpar, whf, bhf, s);
        });
      }
    };

    // shortcut branch first so the residual add can ride in the epilogue of the last main conv
    Act sc;
    const bool pair = b.cin != b.cout && packed_.count(S("blk%d.aadpair.gb", k)) != 0;
    Act a2;  // pair: [aadsc output | aad0 output], 2 * cin channels
    if (pair) {
      // the shortcut AADLayer and the first main AADLayer both read (m, z_attr): one launch, two N tiles
      a2 = new_act(hw, hw, 2 * b.cin);
      const int cx = b.cin;
      float4* ncpar2 = p.pool.alloc_n<float4>(static_cast<size_t>(B) * 2 * cx);
      float* whf2 = p.pool.alloc_n<float>(static_cast<size_t>(B) * 2 * cx);
      float* bhf2 = p.pool.alloc_n<float>(2 * static_cast<size_t>(B));
      AadFoldSet fs[2];
      for (int L = 0; L < 2; ++L) {
        const std::string lname = L == 0 ? S("blk%d.aadsc", k) : S("blk%d.aad0", k);
        const AadInfo& info = aad_.at(lname);
        fs[L].off_gamma = info.off_gamma;
        fs[L].off_beta = info.off_beta;
        fs[L].w_h = vecs_.at(lname + ".wh");
        fs[L].b_h = scalars_.at(lname + ".bh");
        fs[L].ncpar = ncpar2 + L * cx;
        fs[L].wh_fold = whf2 + L * cx;
        fs[L].bh_fold = bhf2 + L * B;
        fs[L].scratch = p.pool.alloc_n<float>(aad_fold_scratch_floats(B, cx));  // zeroed by the pool
        fs[L].ld_out = 2 * cx;
      }
      {
        const AadFoldSet f0 = fs[0], f1 = fs[1];
        const float* sums = sums_m;
        add_op(S("blk%d.aadsc+aad0.fold", k), [=](cudaStream_t s) {
          return aad_fold2_launch(sums, parts_m, B, HW, cx, zfc, ldz, f0, &f1, s);
        });
      }
      AadArgs args;
      args.h = m.p; args.c_x = 2 * cx; args.pair = 1;
      args.ncpar = ncpar2; args.wh_fold = whf2; args.bh_fold = bhf2;
      add_conv(S("blk%d.aadpair.gemm_blend", k), za, 0, za.C, Taps::point(), hw, hw, S("blk%d.aadpair.gb", k),
               OutView::dense(a2), Epilogue(), &args);
    }
    // c_x > 128: the masks of the shortcut layer and of the first main layer in one pass over m
    AadPre pre_sc{}, pre_0{};
    const bool mask2 = bf && !pair && b.cin != b.cout && b.cin > 128;
    const bool fold2 = !pair && b.cin != b.cout;  // shortcut + first main layer: same statistics, one fold launch
    if (fold2) aad_fold_two(S("blk%d.aadsc", k), S("blk%d.aad0", k), sums_m, parts_m, &pre_sc, &pre_0);
    if (mask2) {
      pre_sc.mask = p.pool.alloc_n<float>(static_cast<size_t>(B) * HW);
      pre_0.mask = p.pool.alloc_n<float>(static_cast<size_t>(B) * HW);
      const Act mm = m;
      const AadPre qs = pre_sc, q0 = pre_0;
      const int cx = b.cin;
      add_op(S("blk%d.aadsc+aad0.mask", k), [=](cudaStream_t s) {
        return aad_mask2_launch(mm.p, B, HW, cx, qs.whf, qs.bhf, qs.mask, q0.whf, q0.bhf, q0.mask, s, dt);
      }, kOpStats, static_cast<double>(B) * HW * cx * esz);
    }
    if (b.cin != b.cout) {
      sc = new_act(hw, hw, b.cout);
      if (!pair) aad_layer(S("blk%d.aadsc", k), m, sums_m, parts_m, fold2 ? &pre_sc : nullptr);
      add_conv(S("blk%d.convsc", k), pair ? a2 : a, 0, b.cin, Taps::square(3, 1, 1), hw, hw, S("blk%d.convsc", k),
               OutView::dense(sc), Epilogue(), nullptr);
    }
    Act x = m;
    const float* sums_x = sums_m;
    int sums_xP = parts_m;
    Act y;
    // tensor-core modes, images of >= 256 pixels: the statistics of a conv output that feeds the next AADLayer come
    // out of that conv's epilogue (igemm.cuh st_mode) instead of a separate pass over the tensor
    static const int fuse_stats_env = [] {
      const char* e = getenv("GHOST_FUSE_STATS");  // A/B switch for profiling: GHOST_FUSE_STATS=0 keeps in_stats_kernel
      return e ? atoi(e) : 1;
    }();
    // (measured: worth it where the conv is K-heavy, blocks 4-6; in the thin-K row-window convs of blocks 7-8 the
    // epilogue is the bottleneck and the column sums cost more there than the separate pass over the tensor)
    const bool fuse_stats = bf && HW >= 256 && HW <= 4096 && fuse_stats_env != 0;
    for (int i = 0; i < num_blocks; ++i) {
      const bool last = i == num_blocks - 1;
      const bool from_pair = pair && i == 0;
      if (!from_pair) aad_layer(S("blk%d.aad%d", k, i), x, sums_x, sums_xP, (fold2 && i == 0) ? &pre_0 : nullptr);
      const int co = last ? b.cout : b.cin;
      Act o = new_act(hw, hw, co);
      Epilogue e;
      if (last) {  // x + h (AADLayer.py:76-79); tanh after the last block (AEI_Net.py:139)
        const Act& r = b.cin != b.cout ? sc : m;
        const OutView rv = OutView::dense(r);
        e.res = r.p;
        e.r_sn = rv.sn; e.r_sh = rv.sh; e.r_sw = rv.sw; e.r_off = rv.off;
        if (k == 8) e.act2 = kAct2Tanh;
      }
      float* fused_parts = nullptr;
      int fused_P = 0;
      e.want_stats = !last && fuse_stats;
      add_conv(S("blk%d.conv%d", k, i), from_pair ? a2 : a, from_pair ? b.cin : 0, b.cin, Taps::square(3, 1, 1), hw, hw,
               S("blk%d.conv%d", k, i), OutView::dense(o), e, nullptr, nullptr, &fused_parts, &fused_P);
      if (!last && fused_parts) {
        sums_x = fused_parts;
        sums_xP = fused_P;
      } else if (!last) {
        float* sums_o = p.pool.alloc_n<float>(static_cast<size_t>(B) * (2 * parts_in + 1) * co);  // co == b.cin here
        const Act oo = o;
        add_op(S("blk%d.stats%d", k, i + 1),
               [=](cudaStream_t s) { return in_stats_launch(oo.p, dt, B, HW, oo.C, sums_o, s); }, kOpStats,
               static_cast<double>(B) * HW * oo.C * esz);
        sums_x = sums_o;
        sums_xP = parts_in;
      }
      x = o;
      y = o;
    }
    p.named[S("y%d", k)] = y;
    p.named[S("a%d", k)] = a;  // output of the block's last AADLayer+ReLU
    if (k < 8) {
      // bilinear x2 (AEI_Net.py:124-138) fused with the InstanceNorm statistics of the next block's input
      Act nm = new_act(2 * hw, 2 * hw, b.cout);
      const Act yy = y, nn = nm;
      const int parts_nx = upsample2x_stats_parts(B, yy.H, yy.W, yy.C);
      float* sums_nx = p.pool.alloc_n<float>(static_cast<size_t>(B) * (2 * parts_nx + 1) * b.cout);
      add_op(S("blk%d.upsample_stats", k), [=](cudaStream_t s) {
        return upsample2x_stats_launch(yy.p, nn.p, dt, B, yy.H, yy.W, yy.C, sums_nx, s);
      }, kOpUpsample, 5.0 * B * yy.H * yy.W * yy.C * esz);
      sums_next = sums_nx;
      parts_next = parts_nx;
      m = nm;
      p.named[S("m%d", k + 1)] = m;
    } else {
      p.y = y;
    }
  }
}

}  // namespace ghost
