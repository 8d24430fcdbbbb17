// ArcFace iresnet100 plan.  The reference imports arcface_model/iresnet.py, which is NOT in its
// tree (download_models.sh:2-3 fetches a copy of insightface recognition/arcface_torch/backbones/
// iresnet.py); the architecture below restates that public file:
//   conv3x3(3,64,s1)+BN+PReLU -> layers [3,13,30,3] of IBasicBlock(planes 64,128,256,512; the first
//   block of each layer has stride 2 and a conv1x1(s2)+BN downsample) -> BN2d -> flatten(NCHW) ->
//   Linear(25088,512) -> BN1d(512).  IBasicBlock: BN1 -> conv3x3(s1) -> BN2 -> PReLU -> conv3x3(stride)
//   -> BN3 -> (+identity).  eps = 1e-5 everywhere.  Call sites: inference.py:33-36, core.py:45,54.
#include <math.h>

#include "net.cuh"

namespace ghost {

namespace {
const int kLayers[4] = {3, 13, 30, 3};
const int kPlanes[4] = {64, 128, 256, 512};

void bn_fold_t(const HostTensor& g, const HostTensor& b, const HostTensor& rm, const HostTensor& rv,
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

namespace {
// GHOST_FOLD_BN1=0: keep bn1 as the second output of the producing conv (A/B switch for profiling)
bool fold_bn1_mode() {
  static const int on = [] {
    const char* e = getenv("GHOST_FOLD_BN1");
    return e ? atoi(e) : 1;
  }();
  return on != 0;
}
}  // namespace

void Net::finalize_arcface() {
  const bool bf = is16(act_dtype);  // tensor-core mode: bf16 or fp16 storage
  const bool fold_bn1 = fold_bn1_mode();
  auto bn = [&](const std::string& k, std::vector<float>* sc, std::vector<float>* sh) {
    bn_fold_t(W(k + ".weight"), W(k + ".bias"), W(k + ".running_mean"), W(k + ".running_var"), sc, sh);
  };
  // conv weight [Co,Ci,k,k] -> [Co][tap][Ci] with a per-Cout scale folded in
  auto pack_conv = [&](const std::string& name, const std::string& key, int Ci, int Co, int k,
                       const std::vector<float>& sc, const std::vector<float>& sh) {
    const HostTensor& w = W(key);
    if (w.shape != std::vector<int64_t>({Co, Ci, k, k})) fail(-2, "unexpected shape for " + key);
    const int T = k * k;
    std::vector<float> rows(static_cast<size_t>(Co) * T * Ci);
    for (int co = 0; co < Co; ++co)
      for (int ci = 0; ci < Ci; ++ci)
        for (int t = 0; t < T; ++t)
          rows[(static_cast<size_t>(co) * T + t) * Ci + ci] = w.data[(static_cast<size_t>(co) * Ci + ci) * T + t] * sc[co];
    const bool tc = bf && Ci % 32 == 0;
    packed_[name] = pack(rows, Co, T * Ci, &sh, !tc, tc);
  };
  std::vector<float> sc, sh;
  bn("bn1", &sc, &sh);
  pack_conv("stem", "conv1.weight", 3, 64, 3, sc, sh);
  vecs_["stem.prelu"] = wpool_.upload(W("prelu.weight").data);
  // Tensor-core mode, blocks 1..48: the residual stream is stored SHIFTED, x' = x + c with c = t1 / s1 of the block that
  // reads it.  Then bn1(x) = s1 * x' exactly, zero padding included (pad(s1 * x') = s1 * pad(x')), so bn1 is nothing but
  // a scale in conv1's weights - no border-class biases - and the constants move into biases: the conv2 that writes
  // x'_{b+1} adds c_{b+1} and, where it reads x'_b as its residual, subtracts c_b; a downsample conv reading x'_b
  // subtracts W_d c_b.  Block 0 reads the stem's PReLU output, which cannot carry a constant: it keeps the nine
  // border-class biases (igemm.cuh bias9), as does any block whose bn1 scale is too small to divide by.
  static const int xshift_mode = [] {
    const char* e = getenv("GHOST_XSHIFT");  // A/B switch for profiling
    return e ? atoi(e) : 1;
  }();
  std::vector<std::vector<float>> cshift;  // per block: c (empty = not shifted)
  {
    int ip = 64;
    for (int l = 0; l < 4; ++l)
      for (int i = 0; i < kLayers[l]; ++i) {
        char kb[64];
        snprintf(kb, sizeof(kb), "layer%d.%d", l + 1, i);
        const int ci = i == 0 ? ip : kPlanes[l];
        std::vector<float> c;
        if (bf && fold_bn1 && xshift_mode && !cshift.empty() && ci % 32 == 0) {
          std::vector<float> s1v, t1v;
          bn(std::string(kb) + ".bn1", &s1v, &t1v);
          c.resize(ci);
          bool ok = true;
          for (int ch = 0; ch < ci && ok; ++ch) {
            if (fabsf(s1v[ch]) < 1e-3f) ok = false;
            else c[ch] = t1v[ch] / s1v[ch];
            if (ok && fabsf(c[ch]) > 8.f) ok = false;
          }
          if (!ok) c.clear();
        }
        cshift.push_back(c);
        if (i == kLayers[l] - 1) ip = kPlanes[l];
      }
    cshift.push_back(std::vector<float>());  // the head reads the plain stream
  }
  size_t blk_i = 0;
  int inpl = 64;
  for (int l = 0; l < 4; ++l) {
    const int pl = kPlanes[l];
    for (int i = 0; i < kLayers[l]; ++i, ++blk_i) {
      char kb[64], nb[64];
      snprintf(kb, sizeof(kb), "layer%d.%d", l + 1, i);
      snprintf(nb, sizeof(nb), "l%d.%d", l + 1, i);
      const std::string k = kb, n = nb;
      const int ci = i == 0 ? inpl : pl;
      const std::vector<float>& ck = cshift[blk_i];      // shift of the stream this block reads (ci channels) or empty
      const std::vector<float>& cn = cshift[blk_i + 1];  // shift of the stream this block writes (pl channels) or empty
      bn(k + ".bn1", &sc, &sh);
      vecs_[n + ".bn1.scale"] = wpool_.upload(sc);
      vecs_[n + ".bn1.shift"] = wpool_.upload(sh);
      const std::vector<float> s1 = sc, t1 = sh;
      bn(k + ".bn2", &sc, &sh);
      if (!ck.empty()) {
        // shifted input stream: bn1 is the scale s1 in the weights, nothing else
        const HostTensor& w = W(k + ".conv1.weight");
        if (w.shape != std::vector<int64_t>({pl, ci, 3, 3})) fail(-2, "unexpected shape for " + k + ".conv1.weight");
        std::vector<float> rows(static_cast<size_t>(pl) * 9 * ci);
        for (int co = 0; co < pl; ++co)
          for (int c = 0; c < ci; ++c)
            for (int t = 0; t < 9; ++t)
              rows[(static_cast<size_t>(co) * 9 + t) * ci + c] = w.data[(static_cast<size_t>(co) * ci + c) * 9 + t] * sc[co] * s1[c];
        packed_[n + ".conv1"] = pack(rows, pl, 9 * ci, &sh, false, true);
        scalars_[n + ".conv1.folded"] = 1.f;  // marks "reads the (shifted) stream directly" for build_arcface
      } else if (bf && fold_bn1 && ci % 32 == 0) {
        // Tensor-core mode: bn1 folded into conv1.  conv1 pads bn1(x) = s1*x + t1 with zeros, so the scale goes into
        // the weights (w * s1[ci]) and the shift reaches an output pixel only through the taps that lie inside the
        // image: nine bias vectors, one per border class (igemm.cuh bias9).  No bn1 pass, no second output of the
        // producing conv, one tensor less to write and read per block.
        const HostTensor& w = W(k + ".conv1.weight");
        if (w.shape != std::vector<int64_t>({pl, ci, 3, 3})) fail(-2, "unexpected shape for " + k + ".conv1.weight");
        std::vector<float> rows(static_cast<size_t>(pl) * 9 * ci);
        std::vector<double> wt(static_cast<size_t>(pl) * 9, 0.0);  // sum over ci of w * t1 per (co, tap)
        for (int co = 0; co < pl; ++co)
          for (int c = 0; c < ci; ++c)
            for (int t = 0; t < 9; ++t) {
              const float wv = w.data[(static_cast<size_t>(co) * ci + c) * 9 + t];
              rows[(static_cast<size_t>(co) * 9 + t) * ci + c] = wv * sc[co] * s1[c];
              wt[static_cast<size_t>(co) * 9 + t] += static_cast<double>(wv) * t1[c];
            }
        packed_[n + ".conv1"] = pack(rows, pl, 9 * ci, nullptr, false, true);
        const int ld = packed_[n + ".conv1"].rows_pad;
        std::vector<float> b9(static_cast<size_t>(9) * ld, 0.f);
        for (int cls = 0; cls < 9; ++cls) {
          const int rc = cls / 3, cc = cls % 3;  // 0 = first row / column, 1 = inner, 2 = last
          for (int co = 0; co < pl; ++co) {
            double acc = 0.0;
            for (int t = 0; t < 9; ++t) {
              const int dy = t / 3 - 1, dx = t % 3 - 1;
              if ((rc == 0 && dy < 0) || (rc == 2 && dy > 0) || (cc == 0 && dx < 0) || (cc == 2 && dx > 0)) continue;
              acc += wt[static_cast<size_t>(co) * 9 + t];
            }
            b9[static_cast<size_t>(cls) * ld + co] = sh[co] + sc[co] * static_cast<float>(acc);
          }
        }
        vecs_[n + ".conv1.bias9"] = wpool_.upload(b9);
      } else {
        pack_conv(n + ".conv1", k + ".conv1.weight", ci, pl, 3, sc, sh);
      }
      {
        std::vector<float> pr = W(k + ".prelu.weight").data;
        pr.resize((pl + 255) / 256 * 256, 0.f);
        vecs_[n + ".prelu"] = wpool_.upload(pr);
      }
      bn(k + ".bn3", &sc, &sh);
      // shifted streams: + c of the stream written; - c of the stream read back as the residual (stride-1 blocks)
      for (int co = 0; co < pl; ++co) {
        if (!cn.empty()) sh[co] += cn[co];
        if (i != 0 && !ck.empty()) sh[co] -= ck[co];
      }
      pack_conv(n + ".conv2", k + ".conv2.weight", pl, pl, 3, sc, sh);
      if (i == 0) {
        const std::vector<float> sc3 = sc, sh3 = sh;
        bn(k + ".downsample.1", &sc, &sh);
        if (!ck.empty()) {  // the downsample conv reads x' = x + c: take W_d c back out
          const HostTensor& wd0 = W(k + ".downsample.0.weight");
          for (int co = 0; co < pl; ++co) {
            double acc = 0.0;
            for (int c = 0; c < ci; ++c) acc += static_cast<double>(wd0.data[static_cast<size_t>(co) * ci + c]) * ck[c];
            sh[co] -= sc[co] * static_cast<float>(acc);
          }
        }
        pack_conv(n + ".down", k + ".downsample.0.weight", ci, pl, 1, sc, sh);
        if (bf && ci % 64 == 0 && pl % 64 == 0) {
          // Tensor-core mode: bn3(conv2(u)) + bn_d(downsample(x)) as ONE GEMM - the 1x1 stride-2 shortcut conv rides as
          // extra K blocks behind the 3x3 taps (conv.cuh Epilogue::x2): rows [co] = [conv2 * sc3 | downsample * sc_d],
          // bias = sh3 + sh_d.  No downsample launch, no residual read.
          const HostTensor& w2 = W(k + ".conv2.weight");
          const HostTensor& wd = W(k + ".downsample.0.weight");
          const int K = 9 * pl + ci;
          std::vector<float> rows(static_cast<size_t>(pl) * K);
          std::vector<float> bias(pl);
          for (int co = 0; co < pl; ++co) {
            for (int c = 0; c < pl; ++c)
              for (int t = 0; t < 9; ++t)
                rows[static_cast<size_t>(co) * K + t * pl + c] = w2.data[(static_cast<size_t>(co) * pl + c) * 9 + t] * sc3[co];
            for (int c = 0; c < ci; ++c) rows[static_cast<size_t>(co) * K + 9 * pl + c] = wd.data[static_cast<size_t>(co) * ci + c] * sc[co];
            bias[co] = sh3[co] + sh[co];
          }
          packed_[n + ".conv2x"] = pack(rows, pl, K, &bias, false, true);
        }
      }
    }
    inpl = pl;
  }
  // head: BN2d -> flatten (c,h,w) -> fc -> BN1d, all folded into one [512][25088] matrix in (h,w,c) order
  std::vector<float> s2, t2, sf, tf;
  bn("bn2", &s2, &t2);
  bn("features", &sf, &tf);
  const HostTensor& fw = W("fc.weight");
  const HostTensor& fb = W("fc.bias");
  if (fw.shape != std::vector<int64_t>({512, 25088})) fail(-2, "unexpected shape for fc.weight");
  std::vector<float> rows(static_cast<size_t>(512) * 25088), bias(512);
  for (int j = 0; j < 512; ++j) {
    double acc = fb.data[j];
    for (int c = 0; c < 512; ++c)
      for (int hw = 0; hw < 49; ++hw) {
        const float w = fw.data[static_cast<size_t>(j) * 25088 + c * 49 + hw];
        rows[static_cast<size_t>(j) * 25088 + hw * 512 + c] = w * s2[c] * sf[j];
        acc += static_cast<double>(w) * t2[c];
      }
    bias[j] = static_cast<float>(acc) * sf[j] + tf[j];
  }
  packed_["fc"] = pack(rows, 512, 25088, &bias, !bf, bf);
}

void Net::build_arcface(ExecPlan& p) {
  const int B = p.B;
  const int dt = act_dtype;
  const bool bf = is16(dt);  // tensor-core mode: bf16 or fp16 storage
  int* err = p.err_word;
  auto new_act = [&](int H, int W, int C) {
    Act a;
    a.N = B; a.H = H; a.W = W; a.C = C; a.dtype = dt;
    a.p = p.pool.alloc(a.bytes());
    return a;
  };
  // Lowered convs wait in `pend` while the next one has the same geometry: the identical-shape 3x3 convs of a stage
  // (layer 3: 58 launches, layer 2: 24) become ONE chain launch (conv.cuh make_chain) - every cluster keeps whole images,
  // so a cluster barrier stands where the kernel boundary was.
  std::vector<LaunchOp> pend;
  auto push_lop = [&](const LaunchOp& lop) {
    Op op;
    op.name = lop.name;
    op.flops = lop.flops;
    op.cls = lop.kind == LaunchOp::kSimt ? kOpConvSimt : kOpConvTC;
    op.run = [lop](cudaStream_t s) { return lop.launch(s); };
    p.ops.push_back(std::move(op));
  };
  auto flush = [&]() {
    if (pend.size() >= 2) {
      LaunchOp chain;
      bool fused = false;
      check_msg(make_chain(&chain, pend, p.pool.alloc(chain_param_bytes(pend.size())), &fused), pend.front().name);
      if (fused) {
        push_lop(chain);
        pend.clear();
        return;
      }
    }
    for (const LaunchOp& l : pend) push_lop(l);
    pend.clear();
  };
  auto add_conv = [&](const std::string& name, const Act& in, int Cin, const Taps& taps, int Ho, int Wo,
                      const std::string& wname, const OutView& out, const Epilogue& epi) {
    auto it = packed_.find(wname);
    if (it == packed_.end()) fail(-2, "internal: packed weights missing for " + wname);
    LaunchOp lop;
    lop.name = name;
    check_msg(lower_conv(&lop, in, 0, Cin, taps, Ho, Wo, it->second, out, epi, !bf, err, nullptr), name);
    if (lop.ws_bytes) lop.set_workspace(p.pool.alloc(lop.ws_bytes));  // split-K partial tiles
    if (!pend.empty() && !chain_compatible(pend.front(), lop)) flush();
    pend.push_back(lop);
  };
  static const int fuse_down = [] {
    const char* ev = getenv("GHOST_FUSE_DOWN");  // A/B switch for profiling
    return ev ? atoi(ev) : 1;
  }();

  // Block names in execution order: in bf16 mode the conv that produces a block's input also writes bn1(input)
  // as a second output (IBasicBlock.forward applies bn1 before conv1 and keeps x for the shortcut), which saves
  // one read + write pass per block; the fp32 (CUDA-core) mode keeps the separate per-channel affine.
  std::vector<std::string> names;
  for (int l = 0; l < 4; ++l)
    for (int i = 0; i < kLayers[l]; ++i) {
      char nb[64];
      snprintf(nb, sizeof(nb), "l%d.%d", l + 1, i);
      names.push_back(nb);
    }
  size_t blk = 0;
  p.emb_in = new_act(112, 112, 3);
  Act x = new_act(112, 112, 64);
  Act t_fused;  // bn1(x) of the block about to run, when the producer conv wrote it
  auto fuse_bn1 = [&](Epilogue* e, const Act& like, size_t next_blk) {
    t_fused = Act();
    if (!bf || next_blk >= names.size() || like.C % 16) return;
    if (vecs_.count(names[next_blk] + ".conv1.bias9") || scalars_.count(names[next_blk] + ".conv1.folded"))
      return;  // the next block's bn1 is folded into its conv1
    t_fused = new_act(like.H, like.W, like.C);
    e->out2 = t_fused.p;
    e->scale2 = vecs_.at(names[next_blk] + ".bn1.scale");
    e->shift2 = vecs_.at(names[next_blk] + ".bn1.shift");
  };
  {
    Epilogue e;
    e.act1 = kActPRelu;
    e.slope_c = vecs_.at("stem.prelu");
    // the stem runs on CUDA cores (Cin = 3): no second output there
    add_conv("stem", p.emb_in, 3, Taps::square(3, 1, 1), 112, 112, "stem", OutView::dense(x), e);
  }
  int hw = 112;
  for (int l = 0; l < 4; ++l) {
    const int pl = kPlanes[l];
    for (int i = 0; i < kLayers[l]; ++i) {
      char nb[64];
      snprintf(nb, sizeof(nb), "l%d.%d", l + 1, i);
      const std::string n = nb;
      const int stride = i == 0 ? 2 : 1;
      const int ho = hw / stride;
      Act t = t_fused;
      const auto b9 = vecs_.find(n + ".conv1.bias9");
      if (b9 != vecs_.end() || scalars_.count(n + ".conv1.folded")) {
        t = x;  // bn1 lives in conv1's weights (and border-class biases, or the shift of the stream)
      } else if (!t.p) {
        t = new_act(hw, hw, x.C);
        const Act xx = x, tt = t;
        const float* scv = vecs_.at(n + ".bn1.scale");
        const float* shv = vecs_.at(n + ".bn1.shift");
        flush();
        Op op;
        op.name = n + ".bn1";
        op.run = [=](cudaStream_t s) {
          return channel_affine_launch(xx.p, tt.p, dt, static_cast<long long>(B) * xx.H * xx.W, xx.C, scv, shv,
                                       nullptr, s);
        };
        p.ops.push_back(std::move(op));
      }
      Act u = new_act(hw, hw, pl);
      {
        Epilogue e;
        e.act1 = kActPRelu;
        e.slope_c = vecs_.at(n + ".prelu");
        if (b9 != vecs_.end()) {
          e.bias9 = b9->second;
          e.bias9_ld = packed_.at(n + ".conv1").rows_pad;
        }
        add_conv(n + ".conv1", t, x.C, Taps::square(3, 1, 1), hw, hw, n + ".conv1", OutView::dense(u), e);
      }
      Act idn = x;
      Act o = new_act(ho, ho, pl);
      bool fused_down = false;
      if (i == 0 && bf && fuse_down && packed_.count(n + ".conv2x")) {
        // conv2 + downsample shortcut as one launch, where the CTA-pair kernel applies at this batch size
        Epilogue e;
        e.x2 = x;
        e.x2_cin = x.C;
        fuse_bn1(&e, o, blk + 1);
        LaunchOp lop;
        lop.name = n + ".conv2+down";
        if (!lower_conv(&lop, u, 0, pl, Taps::square(3, 1, stride), ho, ho, packed_.at(n + ".conv2x"), OutView::dense(o), e,
                        false, err, nullptr)) {
          if (!pend.empty() && !chain_compatible(pend.front(), lop)) flush();
          pend.push_back(lop);
          fused_down = true;
        }
      }
      if (i == 0 && !fused_down) {
        idn = new_act(ho, ho, pl);
        Taps tp = Taps::point();
        tp.stride = 2;
        add_conv(n + ".down", x, x.C, tp, ho, ho, n + ".down", OutView::dense(idn), Epilogue());
      }
      if (!fused_down) {
        Epilogue e;
        const OutView rv = OutView::dense(idn);
        e.res = idn.p;
        e.r_sn = rv.sn; e.r_sh = rv.sh; e.r_sw = rv.sw; e.r_off = rv.off;
        fuse_bn1(&e, o, blk + 1);
        add_conv(n + ".conv2", u, pl, Taps::square(3, 1, stride), ho, ho, n + ".conv2", OutView::dense(o), e);
      }
      ++blk;
      x = o;
      hw = ho;
    }
  }
  // head
  p.emb_out = p.pool.alloc_n<float>(static_cast<size_t>(B) * 512);
  Act flat = x;
  flat.H = flat.W = 1;
  flat.C = 25088;
  OutView zo;
  zo.p = p.emb_out;
  zo.sn = 512;
  zo.dtype = kF32;
  if (bf) {
    add_conv("fc", flat, 25088, Taps::point(), 1, 1, "fc", zo, Epilogue());
  } else {
    add_conv("fc", flat, 25088, Taps::point(), 1, 1, "fc", zo, Epilogue());
  }
  flush();
}

}  // namespace ghost
