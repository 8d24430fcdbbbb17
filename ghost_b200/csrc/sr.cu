// pix2pix super-resolution generator plan: LIPSPADEGenerator with TestOptions
// (models/config_sr.py:3-29) in the `.train()` state the reference runs it in (inference.py:48):
//   models/networks/generator.py:93-137   SPADEGenerator.forward (is_test -> phase 4)
//   models/networks/generator.py:311-400  lip2d / SimplifiedLIP / LIPEncoder / LIPSPADEGenerator.encode
//   models/networks/architecture.py:21-70 SPADEResnetBlock (spectral-normalised convs)
//   models/networks/normalization.py:63-107 SPADE (param-free BatchNorm with batch statistics)
//
// Layout decisions:
//  * channel counts are padded to what the tensor-core tiles want - 64-multiples in the LIP encoder (cp), 64 or a
//    128-multiple in the SPADE part (cpg: 48 -> 64, 96 -> 128, 192 -> 256): padded weight rows / columns are zero, so
//    the padded channels carry exact zeros through every layer and every K chunk is a full 128-byte swizzle row;
//  * the conditioning image enters each SPADE through a 3x3 conv on its nearest-downsampled copy: per resolution
//    the 3x3x3 neighbourhoods are gathered once (K = 27 -> 32) and all mlp_shared convs of that resolution run as
//    ONE thin-K tensor-core GEMM; the hidden activations of the SPADEs of a resolution lie back to back in that
//    tensor and each gamma/beta conv reads a 64-multiple window over its own slice (zero weights on the overhang);
//  * SPADE itself (bf16 mode) is the epilogue of its gamma/beta conv (igemm AAD mode, `aad_spade`): batch-norm of
//    x, x * gamma + beta, LeakyReLU, 1/sigma of the consuming conv; gamma / beta never reach HBM;
//  * nn.Upsample(nearest x2): batch statistics are taken on the source tensor (identical); the upsampled tensor is
//    written once per block because the fused epilogue reads x by TMA (the fp32 mode reads x at (h>>1, w>>1));
//  * spectral norm: weight_orig stays resident in fp32, one power iteration per forward runs on the device, and
//    1/sigma multiplies the *input* of the conv (conv(x, W/sigma) = conv(x/sigma, W)) inside the modulation kernel.
#include <math.h>

#include <algorithm>
#include <string.h>

#include "net.cuh"
#include "sr_kernels.cuh"

namespace ghost {

namespace {

struct SrBlock {
  const char* name;
  int fin, fout, S, up;
};
const SrBlock kSrBlocks[7] = {{"head_0", 768, 768, 8, 0},   {"G_middle_0", 768, 768, 16, 1},
                              {"G_middle_1", 768, 768, 16, 0}, {"ups.0", 768, 384, 32, 1},
                              {"ups.1", 384, 192, 64, 1},  {"ups.2", 192, 96, 128, 1},
                              {"ups.3", 96, 48, 256, 1}};
const int kSrEncC[6] = {48, 96, 192, 384, 768, 768};

inline int cp(int c) { return (c + 63) / 64 * 64; }
// SPADE generator tensors: 64, or a multiple of 128 (the fused tcgen05 epilogue works on N tiles of 64 or 128
// channels: 48 -> 64, 96 -> 128, 192 -> 256)
inline int cpg(int c) { return c <= 64 ? 64 : (c + 127) / 128 * 128; }
inline int nhid(int norm_nc) { return norm_nc > 128 ? 128 : norm_nc; }

void expect(const HostTensor& t, std::initializer_list<int64_t> shape, const std::string& key) {
  if (t.shape != std::vector<int64_t>(shape)) fail(-2, "unexpected shape for " + key);
}

// [Co][Ci][k][k] -> rows [Cop][tap][Cip] (zero padded)
std::vector<float> repack(const HostTensor& w, int Co, int Ci, int k, int Cop, int Cip) {
  const int taps = k * k;
  std::vector<float> rows(static_cast<size_t>(Cop) * taps * Cip, 0.f);
  for (int co = 0; co < Co; ++co)
    for (int ci = 0; ci < Ci; ++ci)
      for (int t = 0; t < taps; ++t)
        rows[(static_cast<size_t>(co) * taps + t) * Cip + ci] = w.data[(static_cast<size_t>(co) * Ci + ci) * taps + t];
  return rows;
}

// 3x3 conv on the 3-channel image as rows over the 32-wide patch vector (k = tap*3 + c)
void patch_rows(const HostTensor& w, int Co, std::vector<float>* rows, size_t row0) {
  for (int co = 0; co < Co; ++co)
    for (int c = 0; c < 3; ++c)
      for (int t = 0; t < 9; ++t) (*rows)[(row0 + co) * 32 + t * 3 + c] = w.data[(static_cast<size_t>(co) * 3 + c) * 9 + t];
}

std::vector<float> padded(const std::vector<float>& v, int n, float fill = 0.f) {
  std::vector<float> o(n, fill);
  memcpy(o.data(), v.data(), v.size() * sizeof(float));
  return o;
}

}  // namespace

void Net::finalize_sr() {
  const bool bf = is16(act_dtype);  // tensor-core mode: bf16 or fp16 storage
  sr_sn_.clear();
  sr_slices_.clear();
  for (int i = 0; i < 6; ++i) sr_actv_c_[i] = sr_actv_need_[i] = 0;

  // bf16 mode, thin-K convs (K = 32..768 per pixel with N <= 384): they leave the SM through the shared-memory
  // staging + TMA store of the fused epilogue (igemm.cuh aad_no_h).  That epilogue works on [gamma | beta] N tiles,
  // so the real rows go into the beta half of each tile and the gamma half stays zero (the extra MMA columns cost
  // nothing next to the stores these layers are bound by).
  auto pack_ts = [&](const std::vector<float>& rows, int Cp, int K, const std::vector<float>* bias) {
    const int cn = Cp < 128 ? Cp : 128;
    std::vector<float> r2(static_cast<size_t>(2) * Cp * K, 0.f), b2(2 * Cp, 0.f);
    for (int c = 0; c < Cp; ++c) {
      const int t = c / cn, j = c % cn;
      const size_t rb = static_cast<size_t>(t) * 2 * cn + cn + j;
      memcpy(&r2[rb * K], &rows[static_cast<size_t>(c) * K], sizeof(float) * K);
      if (bias) b2[rb] = (*bias)[c];
    }
    return pack(r2, 2 * Cp, K, &b2, false, true);
  };

  // ---- SPADE layers: mlp_shared rows grouped per resolution, gamma|beta matrices
  struct Pending { std::string key; int nh, nhp, row0; };
  std::vector<Pending> per_res[6];  // index = log2(S) - 3
  auto add_spade = [&](const std::string& key, int norm_nc, int S) {
    // The hidden activations of one resolution are packed back to back (stride nh), but every gamma/beta conv reads
    // a window of cp(nh) channels (whole 64-channel K chunks - 32-channel chunks measured slower, the K-block rate
    // is the limit); the window's tail overlaps the next slice or the padding, where this conv's weights are zero.
    const int nh = nhid(norm_nc), nhp = cp(nh), Cp = cpg(norm_nc);
    int ri = 0;
    while ((8 << ri) != S) ++ri;
    Pending pd{key, nh, nhp, sr_actv_c_[ri]};
    per_res[ri].push_back(pd);
    sr_slices_[key] = {ri, sr_actv_c_[ri], nhp};
    sr_actv_need_[ri] = std::max(sr_actv_need_[ri], sr_actv_c_[ri] + nhp);
    sr_actv_c_[ri] += nh;
    const HostTensor& wg = W(key + ".mlp_gamma.weight");
    const HostTensor& wb = W(key + ".mlp_beta.weight");
    expect(wg, {norm_nc, nh, 3, 3}, key + ".mlp_gamma.weight");
    expect(wb, {norm_nc, nh, 3, 3}, key + ".mlp_beta.weight");
    std::vector<float> g = repack(wg, norm_nc, nh, 3, Cp, nhp), b = repack(wb, norm_nc, nh, 3, Cp, nhp);
    if (bf) {
      // fused epilogue: rows interleaved per N tile of cn channels, [gamma(cn) | beta(cn)], so that one TMEM
      // accumulator tile holds both halves of the same channels (as the AAD layers, aei.cu); zero bias
      const int cn = Cp < 128 ? Cp : 128;
      const size_t K = static_cast<size_t>(9) * nhp;
      std::vector<float> rows(static_cast<size_t>(2) * Cp * K);
      for (int c = 0; c < Cp; ++c) {
        const int t = c / cn, j = c % cn;
        const size_t rg = static_cast<size_t>(t) * 2 * cn + j, rb = rg + cn;
        memcpy(&rows[rg * K], &g[static_cast<size_t>(c) * K], sizeof(float) * K);
        memcpy(&rows[rb * K], &b[static_cast<size_t>(c) * K], sizeof(float) * K);
      }
      std::vector<float> zero(2 * Cp, 0.f);
      packed_[key + ".gb"] = pack(rows, 2 * Cp, 9 * nhp, &zero, false, true);
    } else {
      g.insert(g.end(), b.begin(), b.end());
      packed_[key + ".gb"] = pack(g, 2 * Cp, 9 * nhp, nullptr, true, false);
    }
  };
  // spectral-normalised conv: packed (unscaled) weight_orig + resident fp32 copy, u, v, 1/sigma
  auto add_sn_conv = [&](const std::string& key, int Co, int Ci, int k, bool bias) {
    const HostTensor& w = W(key + ".weight_orig");
    expect(w, {Co, Ci, k, k}, key + ".weight_orig");
    const int K = Ci * k * k;
    expect(W(key + ".weight_u"), {Co}, key + ".weight_u");
    std::vector<float> rows = repack(w, Co, Ci, k, cpg(Co), cpg(Ci));
    std::vector<float> b;
    if (bias) b = padded(W(key + ".bias").data, cpg(Co));
    if (bf && k == 1)
      packed_[key] = pack_ts(rows, cpg(Co), cpg(Ci), nullptr);  // conv_s: 1x1, no bias
    else
      packed_[key] = pack(rows, cpg(Co), k * k * cpg(Ci), bias ? &b : nullptr, !bf, bf);
    SnState st;
    st.R = Co;
    st.K = K;
    st.W = wpool_.upload(w.data);
    st.u = wpool_.upload(W(key + ".weight_u").data);
    st.v = has(key + ".weight_v") ? wpool_.upload(W(key + ".weight_v").data) : wpool_.alloc_n<float>(K);
    st.inv_sigma = wpool_.alloc_n<float>(1);
    st.scratch = wpool_.alloc_n<float>(sn_scratch_floats(Co, K));
    sr_sn_[key] = st;
  };
  for (const SrBlock& b : kSrBlocks) {
    const std::string n = b.name;
    const int fmid = b.fin < b.fout ? b.fin : b.fout;
    add_sn_conv(n + ".conv_0", fmid, b.fin, 3, true);
    add_sn_conv(n + ".conv_1", b.fout, fmid, 3, true);
    if (b.fin != b.fout) add_sn_conv(n + ".conv_s", b.fout, b.fin, 1, false);
    add_spade(n + ".norm_0", b.fin, b.S);
    if (b.fin != b.fout) add_spade(n + ".norm_s", b.fin, b.S);
    add_spade(n + ".norm_1", fmid, b.S);  // the narrowest slice last: its window overhang is the smallest
  }
  for (int ri = 0; ri < 6; ++ri) {
    if (!sr_actv_c_[ri]) continue;
    sr_actv_c_[ri] = cp(sr_actv_need_[ri]);        // every read window lies inside the tensor
    if (bf) sr_actv_c_[ri] = cpg(sr_actv_c_[ri]);  // whole N tiles of the TMA-store epilogue
    const int rows_n = sr_actv_c_[ri];
    std::vector<float> rows(static_cast<size_t>(rows_n) * 32, 0.f), bias(rows_n, 0.f);
    for (const Pending& pd : per_res[ri]) {
      const HostTensor& w = W(pd.key + ".mlp_shared.0.weight");
      expect(w, {pd.nh, 3, 3, 3}, pd.key + ".mlp_shared.0.weight");
      patch_rows(w, pd.nh, &rows, pd.row0);
      const HostTensor& b = W(pd.key + ".mlp_shared.0.bias");
      for (int i = 0; i < pd.nh; ++i) bias[pd.row0 + i] = b.data[i];
    }
    char nm[32];
    snprintf(nm, sizeof(nm), "actv%d", 8 << ri);
    packed_[nm] = bf ? pack_ts(rows, rows_n, 32, &bias) : pack(rows, rows_n, 32, &bias, true, false);
  }

  // ---- LIP encoder (generator.py:353-388)
  {
    const HostTensor& w = W("lip_encoder.model.0.weight");
    expect(w, {48, 3, 3, 3}, "lip_encoder.model.0.weight");
    std::vector<float> rows(static_cast<size_t>(64) * 32, 0.f);
    patch_rows(w, 48, &rows, 0);
    packed_["enc0"] = bf ? pack_ts(rows, 64, 32, nullptr) : pack(rows, 64, 32, nullptr, true, false);
  }
  for (int i = 0; i < 5; ++i) {
    const int C = kSrEncC[i], Cn = kSrEncC[i + 1];
    char key[96];
    snprintf(key, sizeof(key), "lip_encoder.model.%d.logit.0.weight", 3 + 4 * i);
    const HostTensor& wl = W(key);
    expect(wl, {C, C, 3, 3}, key);
    char nm[32];
    snprintf(nm, sizeof(nm), "lip%d.logit", i);
    packed_[nm] = pack(repack(wl, C, C, 3, cp(C), cp(C)), cp(C), 9 * cp(C), nullptr, !bf, bf);
    snprintf(key, sizeof(key), "lip_encoder.model.%d.logit.1.weight", 3 + 4 * i);
    snprintf(nm, sizeof(nm), "lip%d.in_w", i);
    vecs_[nm] = wpool_.upload(padded(W(key).data, cp(C)));
    snprintf(key, sizeof(key), "lip_encoder.model.%d.logit.1.bias", 3 + 4 * i);
    snprintf(nm, sizeof(nm), "lip%d.in_b", i);
    vecs_[nm] = wpool_.upload(padded(W(key).data, cp(C)));
    snprintf(key, sizeof(key), "lip_encoder.model.%d.weight", 4 + 4 * i);
    const HostTensor& wc = W(key);
    expect(wc, {Cn, C, 3, 3}, key);
    snprintf(key, sizeof(key), "lip_encoder.model.%d.bias", 4 + 4 * i);
    std::vector<float> b = padded(W(key).data, cp(Cn));
    snprintf(nm, sizeof(nm), "lip%d.conv", i);
    packed_[nm] = pack(repack(wc, Cn, C, 3, cp(Cn), cp(C)), cp(Cn), 9 * cp(C), &b, !bf, bf);
  }
  // ---- to_rgbs[3] (generator.py:133): 3x3, bias
  {
    const HostTensor& w = W("to_rgbs.3.weight");
    expect(w, {3, 48, 3, 3}, "to_rgbs.3.weight");
    std::vector<float> b = W("to_rgbs.3.bias").data;
    packed_["to_rgb"] = pack(repack(w, 3, 48, 3, 3, 64), 3, 9 * 64, &b, !bf, bf);
  }
}

void Net::build_sr(ExecPlan& p) {
  const int B = p.B;
  const int dt = act_dtype;
  const bool bf = is16(dt);  // tensor-core mode: bf16 or fp16 storage
  const double esz = bf ? 2.0 : 4.0;
  int* err = p.err_word;

  auto new_act = [&](int H, int W, int C) {
    Act a;
    a.N = B; a.H = H; a.W = W; a.C = C; a.dtype = dt;
    a.p = p.pool.alloc(a.bytes());
    return a;
  };
  // buffers of equal shape and role are shared between the blocks (the op list is strictly sequential)
  std::map<std::string, Act> shared;
  auto shared_act = [&](const char* role, int H, int W, int C) -> Act {
    char key[64];
    snprintf(key, sizeof(key), "%s:%d:%d:%d", role, H, W, C);
    auto it = shared.find(key);
    if (it != shared.end()) return it->second;
    Act a = new_act(H, W, C);
    shared[key] = a;
    return a;
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
  // real_flops: dense FLOPs of the unpadded layer (the padded channels are zeros, not work)
  auto add_conv = [&](const std::string& name, const Act& in, int c0, int Cin, const Taps& taps, int Ho, int Wo,
                      const std::string& wname, const OutView& out, const Epilogue& epi, double real_flops,
                      const AadArgs* aad = nullptr, double bytes = 0.0) {
    auto it = packed_.find(wname);
    if (it == packed_.end()) fail(-2, "internal: packed weights missing for " + wname);
    LaunchOp lop;
    lop.name = name;
    check_msg(lower_conv(&lop, in, c0, Cin, taps, Ho, Wo, it->second, out, epi, !bf, err, aad), name);
    if (lop.ws_bytes) lop.set_workspace(p.pool.alloc(lop.ws_bytes));  // split-K partial tiles
    Op op;
    op.name = name;
    op.flops = real_flops;
    op.bytes = bytes;
    op.cls = lop.kind == LaunchOp::kSimt ? kOpConvSimt : (aad ? kOpAadTC : kOpConvTC);
    op.run = [lop](cudaStream_t s) { return lop.launch(s); };
    p.ops.push_back(std::move(op));
  };
  // statistics + fold -> {scale, shift}: per (n, c) (InstanceNorm2d) or per c over the batch (BatchNorm, training)
  auto add_norm = [&](const std::string& name, const Act& x, bool pool_n, const float* w, const float* b) -> float2* {
    const int HW = x.H * x.W;
    const int P = in_stats_parts(B, HW, x.C);
    float* parts = p.pool.alloc_n<float>(static_cast<size_t>(B) * (2 * P + 1) * x.C);
    float2* ss = p.pool.alloc_n<float2>(static_cast<size_t>(pool_n ? 1 : B) * x.C);
    const Act xx = x;
    add_op(name + ".stats", [=](cudaStream_t s) {
      return in_stats_launch_ex(xx.p, dt, B, HW, xx.C, parts, pool_n ? 1 : 0, s);
    }, kOpStats, static_cast<double>(B) * HW * x.C * esz);
    add_op(name + ".fold", [=](cudaStream_t s) {
      return norm_fold_launch(parts, P, B, HW, xx.C, pool_n ? 1 : 0, w, b, 1e-5f, ss, s);
    });
    return ss;
  };
  // BatchNorm (training) statistics folded straight into the parameter table of the fused SPADE epilogue
  auto add_norm_ncpar = [&](const std::string& name, const Act& x) -> float4* {
    const int HW = x.H * x.W;
    const int P = in_stats_parts(B, HW, x.C);
    float* parts = p.pool.alloc_n<float>(static_cast<size_t>(B) * (2 * P + 1) * x.C);
    float4* nc = p.pool.alloc_n<float4>(static_cast<size_t>(B) * x.C);  // zeroed: the identity entries stay 0
    const Act xx = x;
    add_op(name + ".stats", [=](cudaStream_t s) {
      return in_stats_launch_ex(xx.p, dt, B, HW, xx.C, parts, 1, s);
    }, kOpStats, static_cast<double>(B) * HW * x.C * esz);
    add_op(name + ".fold", [=](cudaStream_t s) {
      return norm_fold_ncpar_launch(parts, P, B, HW, xx.C, 1e-5f, nc, s);
    });
    return nc;
  };
  // bf16 mode: out = act(conv + bias) through the TMA-store epilogue (weights packed by pack_ts)
  float4* zero_nc = bf ? p.pool.alloc_n<float4>(static_cast<size_t>(B) * 512) : nullptr;  // stays zero
  auto add_conv_ts = [&](const std::string& name, const Act& in, int Cin, const Taps& taps, const std::string& wname,
                         const Act& out, float slope, double real_flops) {
    AadArgs args;
    args.c_x = out.C;
    args.ncpar = zero_nc;
    args.spade = 1;
    args.no_h = 1;
    args.slope = slope;
    if (out.C > 512) fail(-1, "internal: add_conv_ts supports up to 512 output channels");
    add_conv(name, in, 0, Cin, taps, out.H, out.W, wname, OutView::dense(out), Epilogue(), real_flops, &args,
             static_cast<double>(B) * out.H * out.W * (out.C + Cin) * esz);
  };
  const Taps t3 = Taps::square(3, 1, 1);

  // ------------------------------------------------------------------ spectral-norm power iterations
  {
    std::vector<SnDesc> descs;
    int max_R = 0, max_K = 0;
    double bytes = 0.0;
    for (auto& kv : sr_sn_) {
      const SnState& st = kv.second;
      descs.push_back(SnDesc{st.W, st.u, st.v, st.inv_sigma, st.scratch, st.R, st.K});
      max_R = st.R > max_R ? st.R : max_R;
      max_K = st.K > max_K ? st.K : max_K;
      bytes += 2.0 * st.R * st.K * 4.0;
    }
    SnDesc* dd = reinterpret_cast<SnDesc*>(p.pool.alloc(descs.size() * sizeof(SnDesc)));
    check_cuda(cudaMemcpy(dd, descs.data(), descs.size() * sizeof(SnDesc), cudaMemcpyHostToDevice), "sn table");
    const int n = static_cast<int>(descs.size());
    add_op("sn.power_iteration", [=](cudaStream_t s) {
      return sn_power_iter_batched_launch(dd, n, max_R, max_K, s);
    }, kOpMisc, bytes);
  }

  // ------------------------------------------------------------------ conditioning image
  p.x_in = new_act(256, 256, 3);
  Act actv[6];
  Act patches256;
  for (int ri = 5; ri >= 0; --ri) {
    const int S = 8 << ri;
    Act pt = new_act(S, S, 32);
    const Act img = p.x_in;
    add_op("seg.patches" + std::to_string(S), [=](cudaStream_t s) {
      return seg_patches_launch(img.p, pt.p, dt, B, S, s);
    }, kOpMisc, static_cast<double>(B) * S * S * 32 * esz);
    if (S == 256) patches256 = pt;
    actv[ri] = new_act(S, S, sr_actv_c_[ri]);
    Epilogue relu;
    relu.act1 = kActRelu;
    const std::string nm = "actv" + std::to_string(S);
    const double fl = 2.0 * B * S * S * 27.0 * sr_actv_c_[ri];
    if (bf && sr_actv_c_[ri] <= 512)
      add_conv_ts("seg." + nm, pt, 32, Taps::point(), nm, actv[ri], 0.f, fl);
    else
      add_conv("seg." + nm, pt, 0, 32, Taps::point(), S, S, nm, OutView::dense(actv[ri]), relu, fl);
  }

  // ------------------------------------------------------------------ LIP encoder
  Act x;
  {
    Act e0 = new_act(256, 256, 64);
    if (bf)
      add_conv_ts("enc.conv0", patches256, 32, Taps::point(), "enc0", e0, 1.f, 2.0 * B * 65536.0 * 27 * 48);
    else
      add_conv("enc.conv0", patches256, 0, 32, Taps::point(), 256, 256, "enc0", OutView::dense(e0), Epilogue(),
               2.0 * B * 65536.0 * 27 * 48);
    float2* ss = add_norm("enc.norm0", e0, false, nullptr, nullptr);
    x = new_act(256, 256, 64);
    const Act xx = x;
    add_op("enc.norm0.apply", [=](cudaStream_t s) {
      return norm_act_launch(e0.p, xx.p, dt, B, 65536, 64, ss, 64, kSrRelu, s);
    }, kOpMisc, 2.0 * B * 65536 * 64 * esz);
  }
  for (int i = 0; i < 5; ++i) {
    const int S = 256 >> i, C = kSrEncC[i], Cn = kSrEncC[i + 1], Cp = cp(C), Cnp = cp(Cn);
    const std::string L = "lip" + std::to_string(i);
    Act lg = new_act(S, S, Cp);
    add_conv(L + ".logit", x, 0, Cp, t3, S, S, L + ".logit", OutView::dense(lg), Epilogue(),
             2.0 * B * S * S * 9.0 * C * C);
    float2* ssl = add_norm(L + ".logit_norm", lg, false, vecs_.at(L + ".in_w"), vecs_.at(L + ".in_b"));
    Act pl = new_act(S / 2, S / 2, Cp);
    const Act xx = x;
    add_op(L + ".pool", [=](cudaStream_t s) {
      return lip_pool_launch(xx.p, lg.p, pl.p, dt, B, S, S, Cp, ssl, s);
    }, kOpMisc, static_cast<double>(B) * S * S * Cp * esz * 2.25);
    Act e = new_act(S / 2, S / 2, Cnp);
    add_conv(L + ".conv", pl, 0, Cp, t3, S / 2, S / 2, L + ".conv", OutView::dense(e), Epilogue(),
             2.0 * B * (S / 2) * (S / 2) * 9.0 * C * Cn);
    float2* sse = add_norm(L + ".norm", e, false, nullptr, nullptr);
    Act nx = new_act(S / 2, S / 2, Cnp);
    const int hw = (S / 2) * (S / 2);
    const int act = i < 4 ? kSrRelu : kSrNone;
    add_op(L + ".norm.apply", [=](cudaStream_t s) {
      return norm_act_launch(e.p, nx.p, dt, B, hw, Cnp, sse, Cnp, act, s);
    }, kOpMisc, 2.0 * B * hw * Cnp * esz);
    x = nx;
    p.named["enc" + std::to_string(i + 1)] = x;
  }

  // ------------------------------------------------------------------ SPADE resnet blocks
  for (const SrBlock& b : kSrBlocks) {
    const std::string n = b.name;
    const int S = b.S, HW = S * S;
    const int fmid = b.fin < b.fout ? b.fin : b.fout;
    const int Ci = cpg(b.fin), Cm = cpg(fmid), Co = cpg(b.fout);
    const bool learned = b.fin != b.fout;
    const Act xs = x;  // block input, at S/2 when b.up
    // batch statistics of x (identical for the nearest-upsampled tensor)
    float4* nc_x = nullptr;
    float2* ss_x = nullptr;
    Act xfull = xs;  // bf16 mode: x at the block's resolution (the fused epilogue reads it by TMA)
    if (bf) {
      nc_x = add_norm_ncpar(n + ".bn_x", xs);
      if (b.up) {
        xfull = new_act(S, S, Ci);
        const Act dst = xfull;
        add_op(n + ".upsample", [=](cudaStream_t s) {
          return nearest_up2x_launch(xs.p, dst.p, dt, B, xs.H, xs.W, xs.C, s);
        }, kOpUpsample, 5.0 * B * xs.H * xs.W * xs.C * esz);
      }
    } else {
      ss_x = add_norm(n + ".bn_x", xs, true, nullptr, nullptr);
    }

    // one SPADE + activation: src (read through nearest x2 when up) -> dst, scaled by 1/sigma of the consumer conv
    auto spade = [&](const std::string& nkey, const Act& src, int up, const float2* ss, const float4* ncpar,
                     int norm_nc, int act, const std::string& consumer) -> Act {
      const SrSlice sl = sr_slices_.at(nkey);
      const int Cp = cpg(norm_nc), nh = nhid(norm_nc);
      const float* post = sr_sn_.at(consumer).inv_sigma;
      if (bf) {
        // gamma/beta conv on the tensor cores with the modulation as its epilogue: x arrives by TMA, the result
        // leaves by TMA store; gamma and beta never touch HBM
        Act a = shared_act("mod", S, S, Cp);
        AadArgs args;
        args.h = src.p;
        args.c_x = Cp;
        args.ncpar = ncpar;
        args.spade = 1;
        args.slope = act == kSrLeaky02 ? 0.2f : act == kSrRelu ? 0.f : 1.f;
        args.post = post;
        add_conv(nkey + ".gemm_modulate", actv[sl.ri], sl.c0, sl.nhp, t3, S, S, nkey + ".gb", OutView::dense(a),
                 Epilogue(), 2.0 * B * HW * 9.0 * nh * 2.0 * norm_nc, &args,
                 static_cast<double>(B) * HW * (2.0 * Cp + sl.nhp) * esz);
        return a;
      }
      Act gb = shared_act("gb", S, S, 2 * Cp);
      add_conv(nkey + ".gamma_beta", actv[sl.ri], sl.c0, sl.nhp, t3, S, S, nkey + ".gb", OutView::dense(gb), Epilogue(),
               2.0 * B * HW * 9.0 * nh * 2.0 * norm_nc);
      Act a = shared_act("mod", S, S, Cp);
      const Act ss_src = src;
      add_op(nkey + ".modulate", [=](cudaStream_t s) {
        return spade_apply_launch(ss_src.p, gb.p, a.p, dt, B, S, S, Cp, up, ss, act, post, s);
      }, kOpMisc, static_cast<double>(B) * HW * Cp * esz * (up ? 3.25 : 4.0));
      return a;
    };

    const Act hsrc = bf ? xfull : xs;
    const int hup = bf ? 0 : b.up;
    Act x_s;
    if (learned) {
      Act a = spade(n + ".norm_s", hsrc, hup, ss_x, nc_x, b.fin, kSrNone, n + ".conv_s");
      x_s = new_act(S, S, Co);
      if (bf)
        add_conv_ts(n + ".conv_s", a, Ci, Taps::point(), n + ".conv_s", x_s, 1.f,
                    2.0 * B * HW * static_cast<double>(b.fin) * b.fout);
      else
        add_conv(n + ".conv_s", a, 0, Ci, Taps::point(), S, S, n + ".conv_s", OutView::dense(x_s), Epilogue(),
                 2.0 * B * HW * static_cast<double>(b.fin) * b.fout);
    } else if (b.up && !bf) {
      x_s = new_act(S, S, Ci);
      const Act dst = x_s;
      add_op(n + ".upsample", [=](cudaStream_t s) {
        return nearest_up2x_launch(xs.p, dst.p, dt, B, xs.H, xs.W, xs.C, s);
      }, kOpUpsample, 5.0 * B * xs.H * xs.W * xs.C * esz);
    } else {
      x_s = xfull;
    }
    Act a0 = spade(n + ".norm_0", hsrc, hup, ss_x, nc_x, b.fin, kSrLeaky02, n + ".conv_0");
    Act dx = new_act(S, S, Cm);
    add_conv(n + ".conv_0", a0, 0, Ci, t3, S, S, n + ".conv_0", OutView::dense(dx), Epilogue(),
             2.0 * B * HW * 9.0 * b.fin * fmid);
    float2* ss_d = nullptr;
    float4* nc_d = nullptr;
    if (bf)
      nc_d = add_norm_ncpar(n + ".bn_dx", dx);
    else
      ss_d = add_norm(n + ".bn_dx", dx, true, nullptr, nullptr);
    Act a1 = spade(n + ".norm_1", dx, 0, ss_d, nc_d, fmid, kSrLeaky02, n + ".conv_1");
    Act out = new_act(S, S, Co);
    Epilogue e;
    {
      const OutView rv = OutView::dense(x_s);
      e.res = x_s.p;
      e.r_sn = rv.sn; e.r_sh = rv.sh; e.r_sw = rv.sw; e.r_off = rv.off;
    }
    add_conv(n + ".conv_1", a1, 0, Cm, t3, S, S, n + ".conv_1", OutView::dense(out), e,
             2.0 * B * HW * 9.0 * fmid * b.fout);
    x = out;
    p.named[n] = x;
  }

  // ------------------------------------------------------------------ to_rgbs[3](leaky_relu(x, 0.2)) -> tanh
  {
    Act t = shared_act("mod", 256, 256, 64);
    const Act xx = x;
    add_op("out.leaky", [=](cudaStream_t s) {
      return norm_act_launch(xx.p, t.p, dt, B, 65536, 64, nullptr, 0, kSrLeaky02, s);
    }, kOpMisc, 2.0 * B * 65536 * 64 * esz);
    p.y = new_act(256, 256, 3);
    Epilogue e;
    e.act2 = kAct2Tanh;
    add_conv("out.to_rgb", t, 0, 64, t3, 256, 256, "to_rgb", OutView::dense(p.y), e, 2.0 * B * 65536.0 * 9 * 48 * 3);
  }
}

}  // namespace ghost
