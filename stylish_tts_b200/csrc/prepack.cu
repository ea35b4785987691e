// Host-side weight preparation (runs once, at stts_finalize_weights):
//   * folds both weight-norm dialects: parametrizations.weight.original0/1
//     (ada_norm.py:158-163, decoder.py:35-45, pitch_energy_predictor.py:42-54) and legacy
//     weight_g / weight_v (flow.py:40,52,60);  w = v * g / ||v||, norm over all dims but 0
//   * re-lays every dense layer as a K-major matrix W[Npad, taps * cin_pad] in three formats
//     (exact fp32, tf32 hi/lo split, fp16 hi/lo split of W * 2^s), with the row permutations the fused epilogues need
//     (tanh/sigmoid pairs, mean/logstd pairs, re/im pairs) and zero padding of K and N
//   * concatenates the Linear(style) layers of each stage into one matrix
//   * builds the STFT / iSTFT contraction bases of TorchSTFT (generator.py:23-56)
#include <cmath>
#include <cstring>

#include "engine.h"

namespace stts {

namespace {

const double kPi = 3.14159265358979323846;

uint16_t f2bf(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7FFFFFFFu) > 0x7F800000u) return (uint16_t)((u >> 16) | 0x40);
  const uint32_t r = 0x7FFFu + ((u >> 16) & 1u);
  return (uint16_t)((u + r) >> 16);
}
// round-to-nearest (ties away) to the 10-bit tf32 mantissa, like cvt.rna.tf32.f32: unbiased, so the
// split error of the 3-pass product grows like sqrt(K) instead of K
float trunc_tf32(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7F800000u) != 0x7F800000u) u += 0x1000u;
  u &= 0xFFFFE000u;
  std::memcpy(&f, &u, 4);
  return f;
}

template <typename T>
T* dev_upload(Context& c, const T* h, size_t n) {
  T* d = nullptr;
  if (cudaMalloc(&d, n * sizeof(T)) != cudaSuccess) return nullptr;
  cudaMemcpy(d, h, n * sizeof(T), cudaMemcpyHostToDevice);
  c.weight_allocs.push_back(d);
  return d;
}

struct Builder {
  int N = 0, Npad = 0, Ktot = 0, n_kseg = 0;
  int taps[2] = {0, 0}, pad[2] = {0, 0}, cin_pad[2] = {0, 0};
  std::vector<float> W, bias;
  double k_real = 0;
  std::vector<int> row_k;
  void init(int n, int nk, const int* t, const int* p, const int* cp) {
    N = n;
    // weight rows padded so that the 256-column tile (BN = 256) never reads past the matrix
    Npad = n > 128 ? (n + 255) / 256 * 256 : 128;
    n_kseg = nk;
    Ktot = 0;
    for (int s = 0; s < nk; ++s) { taps[s] = t[s]; pad[s] = p[s]; cin_pad[s] = cp[s]; Ktot += t[s] * cp[s]; }
    W.assign((size_t)Npad * Ktot, 0.f);
    bias.assign(Npad, 0.f);
    row_k.assign(Npad, 0);
  }
  int koff(int s) const { return s == 0 ? 0 : taps[0] * cin_pad[0]; }
  // w: [n_src, cin_total, k] (k = taps[s]; Linear weights have k = 1)
  void rows(int s, const HostTensor& w, int src0, int n, int cin0, int cin_cnt, int dst0, int dstride) {
    const int cin_total = (int)w.dim(1), k = (int)w.dim(2);
    for (int i = 0; i < n; ++i) row_k[dst0 + i * dstride] += cin_cnt * k;
    for (int i = 0; i < n; ++i) {
      float* d = &W[(size_t)(dst0 + i * dstride) * Ktot + koff(s)];
      const float* src = &w.data[(size_t)(src0 + i) * cin_total * k];
      for (int cc = 0; cc < cin_cnt; ++cc)
        for (int j = 0; j < k; ++j) d[(size_t)j * cin_pad[s] + cc] = src[(size_t)(cin0 + cc) * k + j];
    }
  }
  void set_bias(const HostTensor& b, int src0, int n, int dst0, int dstride) {
    for (int i = 0; i < n; ++i) bias[dst0 + i * dstride] = b.data[src0 + i];
  }
  PackedGemm upload(Context& c) const {
    PackedGemm p;
    p.N = N; p.Npad = Npad; p.Ktot = Ktot; p.n_kseg = n_kseg;
    { double kk = k_real; for (int i = 0; i < N && kk == 0; ++i) kk = row_k[i] > kk ? row_k[i] : kk; p.k_real = kk; }
    for (int s = 0; s < 2; ++s) { p.taps[s] = taps[s]; p.pad[s] = pad[s]; p.cin_pad[s] = cin_pad[s]; }
    const size_t n = (size_t)Npad * Ktot;
    std::vector<float> hl(2 * n);
    std::vector<__half> hg(n), hlo(n);
    // fp16 split of W * wscale, wscale the power of two that puts max |W| into [2^13, 2^14): the
    // residuals of all but vanishing weights stay normal fp16 numbers (x = g + l to ~2^-22)
    float wmax = 0.f;
    for (size_t i = 0; i < n; ++i) wmax = std::fmax(wmax, std::fabs(W[i]));
    int ex = 0;
    if (wmax > 0.f) { std::frexp(wmax, &ex); }     // wmax = m * 2^ex, m in [0.5, 1)
    const float wscale = std::ldexp(1.0f, 14 - ex);
    p.wscale = wscale;
    for (size_t i = 0; i < n; ++i) {
      const float hi = trunc_tf32(W[i]);
      hl[i] = hi;
      hl[n + i] = trunc_tf32(W[i] - hi);
      const float ws = W[i] * wscale;
      hg[i] = __float2half_rn(ws);
      hlo[i] = __float2half_rn(ws - __half2float(hg[i]));
    }
    p.Wf = dev_upload(c, W.data(), n);
    p.Whl = dev_upload(c, hl.data(), 2 * n);
    p.Wg = reinterpret_cast<__half*>(dev_upload(c, reinterpret_cast<const uint16_t*>(hg.data()), n));
    p.Wl = reinterpret_cast<__half*>(dev_upload(c, reinterpret_cast<const uint16_t*>(hlo.data()), n));
    p.bias = dev_upload(c, bias.data(), (size_t)Npad);
    return p;
  }
};

int pad64(int c) { return (c + 63) / 64 * 64; }

}  // namespace

void free_pack(PackedGemm& p) {
  cudaFree(p.Wf); cudaFree(p.Whl); cudaFree(p.Wg); cudaFree(p.Wl); cudaFree(p.bias);
  p = PackedGemm();
}

// used by stts_test_conv: pack an arbitrary Conv1d weight [n_out, cin, taps]
PackedGemm pack_from_conv_weight(Context& c, const float* w, const float* bias, int n_out, int cin, int taps, int pad) {
  HostTensor t;
  t.shape = {n_out, cin, taps};
  t.data.assign(w, w + (size_t)n_out * cin * taps);
  Builder b;
  const int tp[2] = {taps, 0}, pd[2] = {pad, 0}, cp[2] = {pad64(cin), 0};
  b.init(n_out, 1, tp, pd, cp);
  b.rows(0, t, 0, n_out, 0, cin, 0, 1);
  if (bias != nullptr) for (int i = 0; i < n_out; ++i) b.bias[i] = bias[i];
  Context tmp;  // allocations owned by the returned pack, not by the context
  PackedGemm p = b.upload(tmp);
  return p;
}

namespace {

struct Loader {
  Context& c;
  std::string module;
  bool ok = true;
  std::string missing;
  const HostTensor* get(const std::string& key) {
    auto it = c.tensors.find(module + "/" + key);
    if (it == c.tensors.end()) { if (ok) missing = module + "/" + key; ok = false; return nullptr; }
    return &it->second;
  }
  bool has(const std::string& key) const { return c.tensors.count(module + "/" + key) != 0; }
  // effective weight as [N, Cin, k]
  HostTensor weight(const std::string& p) {
    HostTensor out;
    const HostTensor *g = nullptr, *v = nullptr;
    if (has(p + "parametrizations.weight.original0")) {
      g = get(p + "parametrizations.weight.original0");
      v = get(p + "parametrizations.weight.original1");
    } else if (has(p + "weight_g")) {
      g = get(p + "weight_g");
      v = get(p + "weight_v");
    } else {
      v = get(p + "weight");
    }
    if (v == nullptr) { out.shape = {0, 0, 0}; return out; }
    out.shape = {v->dim(0), v->dim(1), v->dim(2)};
    out.data = v->data;
    if (g != nullptr) {
      const int64_t n = v->dim(0), per = v->numel() / n;
      for (int64_t i = 0; i < n; ++i) {
        double ss = 0.0;
        for (int64_t j = 0; j < per; ++j) { const double x = v->data[i * per + j]; ss += x * x; }
        const float scale = g->data[i] / (float)std::sqrt(ss);
        for (int64_t j = 0; j < per; ++j) out.data[i * per + j] = v->data[i * per + j] * scale;
      }
    }
    return out;
  }
};

void put_vec(Context& c, const std::string& name, const std::vector<float>& v) {
  c.vec[name] = dev_upload(c, v.data(), v.size());
  c.vec_len[name] = (int)v.size();
}

// concatenated Linear(style) layers of a stage: rows appended in call order
struct FcCat {
  std::vector<float> W, b;
  int J = 0;
  int add(const HostTensor& w, const HostTensor& bias) {
    const int off = J;
    W.insert(W.end(), w.data.begin(), w.data.end());
    b.insert(b.end(), bias.data.begin(), bias.data.end());
    J += (int)w.dim(0);
    return off;
  }
};

// AdaptiveDecoderBlock (ada_norm.py:142-182): conv1, conv2 (+ fused 1x1 shortcut), two AdaIN FCs
bool pack_ada_block(Context& c, Loader& L, const std::string& p, const std::string& name, int cin, int cout,
                    int cin_pad, FcCat& fc) {
  const HostTensor w1 = L.weight(p + "conv1."), w2 = L.weight(p + "conv2.");
  const HostTensor *b1 = L.get(p + "conv1.bias"), *b2 = L.get(p + "conv2.bias");
  const HostTensor *n1w = L.get(p + "norm1.fc.weight"), *n1b = L.get(p + "norm1.fc.bias");
  const HostTensor *n2w = L.get(p + "norm2.fc.weight"), *n2b = L.get(p + "norm2.fc.bias");
  if (!L.ok) return false;
  Builder a;
  { const int t[2] = {3, 0}, pd[2] = {1, 0}, cp[2] = {cin_pad, 0}; a.init(cout, 1, t, pd, cp); }
  a.rows(0, w1, 0, cout, 0, cin, 0, 1);
  a.set_bias(*b1, 0, cout, 0, 1);
  c.gemm[name + ".conv1"] = a.upload(c);
  Builder b;
  const bool learned = L.has(p + "conv1x1.parametrizations.weight.original0");
  if (learned) {
    const HostTensor wsc = L.weight(p + "conv1x1.");
    const int t[2] = {3, 1}, pd[2] = {1, 0}, cp[2] = {pad64(cout), cin_pad};
    b.init(cout, 2, t, pd, cp);
    b.rows(0, w2, 0, cout, 0, cout, 0, 1);
    b.rows(1, wsc, 0, cout, 0, cin, 0, 1);
  } else {
    const int t[2] = {3, 0}, pd[2] = {1, 0}, cp[2] = {pad64(cout), 0};
    b.init(cout, 1, t, pd, cp);
    b.rows(0, w2, 0, cout, 0, cout, 0, 1);
  }
  b.set_bias(*b2, 0, cout, 0, 1);
  c.gemm[name + ".conv2"] = b.upload(c);
  c.fc_off[name + ".norm1"] = fc.add(*n1w, *n1b);
  c.fc_off[name + ".norm2"] = fc.add(*n2w, *n2b);
  return true;
}


// one dense conv / Linear as a single-segment pack: [n_out, cin, k] -> K-major, cin padded to 64
bool pack_simple(Context& c, Loader& L, const std::string& key, const std::string& name, int pad) {
  const HostTensor w = L.weight(key + ".");
  const HostTensor* b = L.has(key + ".bias") ? L.get(key + ".bias") : nullptr;
  if (!L.ok || w.numel() == 0) return false;
  const int n = (int)w.dim(0), cin = (int)w.dim(1), k = (int)w.dim(2);
  Builder bd;
  const int t[2] = {k, 0}, pd[2] = {pad, 0}, cp[2] = {pad64(cin), 0};
  bd.init(n, 1, t, pd, cp);
  bd.rows(0, w, 0, n, 0, cin, 0, 1);
  if (b != nullptr) bd.set_bias(*b, 0, n, 0, 1);
  c.gemm[name] = bd.upload(c);
  return true;
}

// TextEncoder (text_encoder.py:396-462) under key prefix `kp` of module L.module -> packs named `nm.*`
bool pack_text_encoder(Context& c, Loader& L, const std::string& kp, const std::string& nm) {
  const HostTensor* emb = L.get(kp + "emb.weight");
  if (!L.ok) return false;
  const int C = (int)emb->dim(1);
  {
    std::vector<float> e(emb->data);
    const float sc = std::sqrt((float)C);       // emb(x) * sqrt(n_channels): same fp32 product as on the device
    for (auto& v : e) v *= sc;
    put_vec(c, nm + ".emb", e);
    c.vec_len[nm + ".vocab"] = (int)emb->dim(0);
  }
  for (int i = 0; i < 3; ++i) {
    const std::string k = kp + "prenet.conv_layers." + std::to_string(i), n = nm + ".pre" + std::to_string(i);
    if (!pack_simple(c, L, k, n, 2)) return false;
    const HostTensor *g = L.get(kp + "prenet.norm_layers." + std::to_string(i) + ".gamma");
    const HostTensor *b = L.get(kp + "prenet.norm_layers." + std::to_string(i) + ".beta");
    if (!L.ok) return false;
    put_vec(c, n + ".g", g->data);
    put_vec(c, n + ".b", b->data);
  }
  if (!pack_simple(c, L, kp + "prenet.proj", nm + ".preproj", 0)) return false;
  int n_layers = 0;
  while (L.has(kp + "encoder.attn_layers." + std::to_string(n_layers) + ".conv_q.weight")) ++n_layers;
  c.vec_len[nm + ".layers"] = n_layers;
  for (int l = 0; l < n_layers; ++l) {
    const std::string a = kp + "encoder.attn_layers." + std::to_string(l) + ".", n = nm + ".l" + std::to_string(l);
    const HostTensor wq = L.weight(a + "conv_q."), wk = L.weight(a + "conv_k."), wv = L.weight(a + "conv_v.");
    const HostTensor *bq = L.get(a + "conv_q.bias"), *bk = L.get(a + "conv_k.bias"), *bv = L.get(a + "conv_v.bias");
    if (!L.ok) return false;
    Builder b;
    const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {pad64(C), 0};
    b.init(3 * C, 1, t, pd, cp);
    b.rows(0, wq, 0, C, 0, C, 0, 1);      b.set_bias(*bq, 0, C, 0, 1);
    b.rows(0, wk, 0, C, 0, C, C, 1);      b.set_bias(*bk, 0, C, C, 1);
    b.rows(0, wv, 0, C, 0, C, 2 * C, 1);  b.set_bias(*bv, 0, C, 2 * C, 1);
    c.gemm[n + ".qkv"] = b.upload(c);
    if (!pack_simple(c, L, a + "conv_o", n + ".o", 0)) return false;
    const std::string f = kp + "encoder.ffn_layers." + std::to_string(l) + ".";
    const HostTensor w1 = L.weight(f + "conv_1.");
    if (!L.ok) return false;
    const int k = (int)w1.dim(2);
    if (!pack_simple(c, L, f + "conv_1", n + ".ff1", k / 2)) return false;
    if (!pack_simple(c, L, f + "conv_2", n + ".ff2", k / 2)) return false;
    for (int j = 1; j <= 2; ++j) {
      const std::string nk = kp + "encoder.norm_layers_" + std::to_string(j) + "." + std::to_string(l) + ".";
      const HostTensor *g = L.get(nk + "gamma"), *be = L.get(nk + "beta");
      if (!L.ok) return false;
      put_vec(c, n + ".n" + std::to_string(j) + ".g", g->data);
      put_vec(c, n + ".n" + std::to_string(j) + ".b", be->data);
    }
  }
  return pack_simple(c, L, kp + "proj_m", nm + ".projm", 0);
}


// ProsodyEncoder (prosody_encoder.py:10-81) under key prefix `kp` -> packs `nm.*`; its AdaLN style FCs
// are concatenated into nm.fc_W / nm.fc_b (one style_fc launch per call)
bool pack_prosody(Context& c, Loader& L, const std::string& kp, const std::string& nm) {
  int n_layers = 0;
  while (L.has(kp + "attn_layers." + std::to_string(n_layers) + ".conv_q.weight")) ++n_layers;
  if (n_layers == 0) return false;
  c.vec_len[nm + ".layers"] = n_layers;
  FcCat fc;
  for (int l = 0; l < n_layers; ++l) {
    const std::string a = kp + "attn_layers." + std::to_string(l) + ".", n = nm + ".l" + std::to_string(l);
    const HostTensor wq = L.weight(a + "conv_q."), wk = L.weight(a + "conv_k."), wv = L.weight(a + "conv_v.");
    const HostTensor *bq = L.get(a + "conv_q.bias"), *bk = L.get(a + "conv_k.bias"), *bv = L.get(a + "conv_v.bias");
    if (!L.ok) return false;
    const int hid = (int)wq.dim(0);
    c.vec_len[nm + ".hid"] = hid;
    Builder b;
    const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {pad64(hid), 0};
    b.init(3 * hid, 1, t, pd, cp);
    b.rows(0, wq, 0, hid, 0, hid, 0, 1);        b.set_bias(*bq, 0, hid, 0, 1);
    b.rows(0, wk, 0, hid, 0, hid, hid, 1);      b.set_bias(*bk, 0, hid, hid, 1);
    b.rows(0, wv, 0, hid, 0, hid, 2 * hid, 1);  b.set_bias(*bv, 0, hid, 2 * hid, 1);
    c.gemm[n + ".qkv"] = b.upload(c);
    if (!pack_simple(c, L, a + "conv_o", n + ".o", 0)) return false;
    const std::string f = kp + "ffn_layers." + std::to_string(l) + ".";
    if (!pack_simple(c, L, f + "conv_1", n + ".ff1", 0)) return false;
    if (!pack_simple(c, L, f + "conv_2", n + ".ff2", 0)) return false;
    if (!pack_simple(c, L, kp + "proj_layers." + std::to_string(l), n + ".proj", 0)) return false;
    c.vec_len[nm + ".dmodel"] = c.gemm[n + ".proj"].N;
    for (int j = 1; j <= 2; ++j) {
      const std::string nk = kp + "norm_layers_" + std::to_string(j) + "." + std::to_string(l) + ".fc.";
      const HostTensor *w = L.get(nk + "weight"), *bb = L.get(nk + "bias");
      if (!L.ok) return false;
      c.fc_off[n + ".n" + std::to_string(j)] = fc.add(*w, *bb);
    }
  }
  c.vec_len[nm + ".J"] = fc.J;
  put_vec(c, nm + ".fc_W", fc.W);
  put_vec(c, nm + ".fc_b", fc.b);
  return true;
}


// TextStyleEncoder (text_style_encoder.py:6-26, conv_next.py:4-51) under key prefix `kp` -> `nm.*`
bool pack_style_encoder(Context& c, Loader& L, const std::string& kp, const std::string& nm) {
  if (!pack_simple(c, L, kp + "conv_in", nm + ".in", 3)) return false;
  int n_blocks = 0;
  while (L.has(kp + "blocks." + std::to_string(n_blocks) + ".dwconv.weight")) ++n_blocks;
  c.vec_len[nm + ".blocks"] = n_blocks;
  for (int i = 0; i < n_blocks; ++i) {
    const std::string k = kp + "blocks." + std::to_string(i) + ".", n = nm + ".b" + std::to_string(i);
    const HostTensor *dw = L.get(k + "dwconv.weight"), *db = L.get(k + "dwconv.bias");
    const HostTensor *nw = L.get(k + "norm.weight"), *nb = L.get(k + "norm.bias");
    const HostTensor *gg = L.get(k + "grn.gamma"), *gb = L.get(k + "grn.beta");
    if (!L.ok) return false;
    put_vec(c, n + ".dw_w", dw->data);      // [C, 1, K] row-major == [C][K]
    put_vec(c, n + ".dw_b", db->data);
    c.vec_len[n + ".dw_k"] = (int)dw->dim(2);
    put_vec(c, n + ".ln_g", nw->data);
    put_vec(c, n + ".ln_b", nb->data);
    put_vec(c, n + ".grn_gamma", gg->data);
    put_vec(c, n + ".grn_beta", gb->data);
    if (!pack_simple(c, L, k + "pwconv1", n + ".pw1", 0)) return false;
    if (!pack_simple(c, L, k + "pwconv2", n + ".pw2", 0)) return false;
  }
  return true;
}

// cross-attention block of PitchEnergyPredictor (pitch_energy_predictor.py:38-102): query/key AdaLN
// FCs, q and (k | v) projections, output projection, depthwise k5 + SiLU + 1x1 post net
bool pack_pe_cross(Context& c, Loader& L, const std::string& nm) {
  FcCat fc;
  const HostTensor *qw = L.get("query_norm.fc.weight"), *qb = L.get("query_norm.fc.bias");
  const HostTensor *kw = L.get("key_norm.fc.weight"), *kb = L.get("key_norm.fc.bias");
  if (!L.ok) return false;
  c.fc_off[nm + ".qn"] = fc.add(*qw, *qb);
  c.fc_off[nm + ".kn"] = fc.add(*kw, *kb);
  c.vec_len[nm + ".J"] = fc.J;
  put_vec(c, nm + ".fc_W", fc.W);
  put_vec(c, nm + ".fc_b", fc.b);
  if (!pack_simple(c, L, "cross_attention.conv_q", nm + ".q", 0)) return false;
  {
    const HostTensor wk = L.weight("cross_attention.conv_k."), wv = L.weight("cross_attention.conv_v.");
    const HostTensor *bk = L.get("cross_attention.conv_k.bias"), *bv = L.get("cross_attention.conv_v.bias");
    if (!L.ok) return false;
    const int C = (int)wk.dim(0);
    Builder b;
    const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {pad64(C), 0};
    b.init(2 * C, 1, t, pd, cp);
    b.rows(0, wk, 0, C, 0, C, 0, 1);  b.set_bias(*bk, 0, C, 0, 1);
    b.rows(0, wv, 0, C, 0, C, C, 1);  b.set_bias(*bv, 0, C, C, 1);
    c.gemm[nm + ".kv"] = b.upload(c);
  }
  if (!pack_simple(c, L, "cross_attention.conv_o", nm + ".o", 0)) return false;
  {
    const HostTensor dw = L.weight("cross_post.0.");
    const HostTensor* db = L.get("cross_post.0.bias");
    if (!L.ok) return false;
    put_vec(c, nm + ".dw_w", dw.data);
    put_vec(c, nm + ".dw_b", db->data);
    c.vec_len[nm + ".dw_k"] = (int)dw.dim(2);
  }
  return pack_simple(c, L, "cross_post.2", nm + ".post", 0);
}

}  // namespace

int finalize_weights(Context& c) {
  for (void* p : c.weight_allocs) cudaFree(p);
  c.weight_allocs.clear();
  c.gemm.clear(); c.vec.clear(); c.vec_len.clear(); c.fc_off.clear(); c.snake.clear();
  c.finalized = false;

  // ======================================================================= speech_predictor
  Loader L{c, "speech_predictor"};
  if (c.has("speech_predictor/decoder.encode.conv1.bias")) {
    FcCat fc;
    // ---- decoder (decoder.py:7-60)
    if (!pack_ada_block(c, L, "decoder.encode.", "dec.encode", 130, 512, 192, fc))
      return fail(c, -3, "missing weight " + L.missing);
    for (int i = 0; i < 4; ++i)
      if (!pack_ada_block(c, L, "decoder.decode." + std::to_string(i) + ".", "dec.decode" + std::to_string(i), 578,
                          512, 640, fc))
        return fail(c, -3, "missing weight " + L.missing);
    c.J_dec = fc.J;
    put_vec(c, "dec.fc_W", fc.W);
    put_vec(c, "dec.fc_b", fc.b);
    {
      const HostTensor wf = L.weight("decoder.F0_conv."), wn = L.weight("decoder.N_conv.");
      const HostTensor *bf = L.get("decoder.F0_conv.bias"), *bn = L.get("decoder.N_conv.bias");
      const HostTensor wr = L.weight("decoder.asr_res.0.");
      const HostTensor* br = L.get("decoder.asr_res.0.bias");
      if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
      put_vec(c, "dec.f0n_w", {wf.data[0], wf.data[1], wf.data[2], bf->data[0], wn.data[0], wn.data[1], wn.data[2],
                               bn->data[0]});
      Builder b;
      const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {128, 0};
      b.init(64, 1, t, pd, cp);
      b.rows(0, wr, 0, 64, 0, 128, 0, 1);
      b.set_bias(*br, 0, 64, 0, 1);
      c.gemm["dec.asr_res"] = b.upload(c);
    }
  }
  if (c.has("speech_predictor/prior_encoder.proj_mean.weight")) {
    // ---- prior encoder (flow.py:296-315): (mean_i, logstd_i) interleaved rows
    {
      const HostTensor *wm = L.get("prior_encoder.proj_mean.weight"), *bm = L.get("prior_encoder.proj_mean.bias");
      const HostTensor *ws = L.get("prior_encoder.proj_logstd.weight"), *bs = L.get("prior_encoder.proj_logstd.bias");
      if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
      Builder b;
      const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {512, 0};
      b.init(256, 1, t, pd, cp);
      b.rows(0, *wm, 0, 128, 0, 512, 0, 2);
      b.rows(0, *ws, 0, 128, 0, 512, 1, 2);
      b.set_bias(*bm, 0, 128, 0, 2);
      b.set_bias(*bs, 0, 128, 1, 2);
      c.gemm["prior"] = b.upload(c);
    }
    // ---- flow (flow.py:99-226): 8 coupling layers x (pre, 4 x (in_layer k5, res_skip), proj)
    {
      FcCat ffc;
      for (int i = 0; i < 8; ++i) {
        const std::string q = "flow.flows." + std::to_string(2 * i) + ".";
        const std::string nm = "flow" + std::to_string(i);
        const HostTensor *wp = L.get(q + "pre.weight"), *bp = L.get(q + "pre.bias");
        if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
        {
          Builder b;
          const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {64, 0};
          b.init(128, 1, t, pd, cp);
          b.rows(0, *wp, 0, 128, 0, 64, 0, 1);
          b.set_bias(*bp, 0, 128, 0, 1);
          c.gemm[nm + ".pre"] = b.upload(c);
        }
        for (int l = 0; l < 4; ++l) {
          const std::string ls = std::to_string(l);
          const HostTensor wi = L.weight(q + "enc.in_layers." + ls + ".");
          const HostTensor* bi = L.get(q + "enc.in_layers." + ls + ".bias");
          const HostTensor wr = L.weight(q + "enc.res_skip_layers." + ls + ".");
          const HostTensor* br = L.get(q + "enc.res_skip_layers." + ls + ".bias");
          if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
          {  // rows interleaved (tanh_i, sigmoid_i) for the fused gate epilogue
            Builder b;
            const int t[2] = {5, 0}, pd[2] = {2, 0}, cp[2] = {128, 0};
            b.init(256, 1, t, pd, cp);
            b.rows(0, wi, 0, 128, 0, 128, 0, 2);
            b.rows(0, wi, 128, 128, 0, 128, 1, 2);
            b.set_bias(*bi, 0, 128, 0, 2);
            b.set_bias(*bi, 128, 128, 1, 2);
            c.gemm[nm + ".in" + ls] = b.upload(c);
          }
          {
            const int nrs = (int)wr.dim(0);  // 256, last layer 128
            Builder b;
            const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {128, 0};
            b.init(nrs, 1, t, pd, cp);
            b.rows(0, wr, 0, nrs, 0, 128, 0, 1);
            b.set_bias(*br, 0, nrs, 0, 1);
            c.gemm[nm + ".rs" + ls] = b.upload(c);
          }
        }
        const HostTensor wc = L.weight(q + "enc.cond_layer.");
        const HostTensor* bc = L.get(q + "enc.cond_layer.bias");
        const HostTensor *wm = L.get(q + "proj_mean.weight"), *bm = L.get(q + "proj_mean.bias");
        const HostTensor *ws = L.get(q + "proj_logstd.weight"), *bs = L.get(q + "proj_logstd.bias");
        if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
        c.fc_off[nm + ".cond"] = ffc.add(wc, *bc);
        Builder b;
        const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {128, 0};
        b.init(128, 1, t, pd, cp);
        b.rows(0, *wm, 0, 64, 0, 128, 0, 2);
        b.rows(0, *ws, 0, 64, 0, 128, 1, 2);
        b.set_bias(*bm, 0, 64, 0, 2);
        b.set_bias(*bs, 0, 64, 1, 2);
        c.gemm[nm + ".proj"] = b.upload(c);
      }
      c.J_flow = ffc.J;
      put_vec(c, "flow.fc_W", ffc.W);
      put_vec(c, "flow.fc_b", ffc.b);
      const HostTensor *wpf = L.get("post_flow.weight"), *bpf = L.get("post_flow.bias");
      if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
      Builder b;
      const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {128, 0};
      b.init(512, 1, t, pd, cp);
      b.rows(0, *wpf, 0, 512, 0, 128, 0, 1);
      b.set_bias(*bpf, 0, 512, 0, 1);
      c.gemm["post_flow"] = b.upload(c);
    }
  }
  if (c.has("speech_predictor/generator.projector.weight")) {
    // ---- generator (generator.py:339-438)
    {
      FcCat gfc;
      const char* prior_names[2] = {"generator.amp_prior_conv.", "generator.phase_prior_conv."};
      for (int h = 0; h < 2; ++h) {
        const HostTensor* w = L.get(std::string(prior_names[h]) + "weight");
        const HostTensor* bb = L.get(std::string(prior_names[h]) + "bias");
        if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
        Builder b;
        const int t[2] = {7, 0}, pd[2] = {3, 0}, cp[2] = {1088, 0};
        b.init(256, 1, t, pd, cp);
        b.rows(0, *w, 0, 256, 0, 1025, 0, 1);
        b.set_bias(*bb, 0, 256, 0, 1);
        c.gemm[h == 0 ? "gen.amp_prior" : "gen.phase_prior"] = b.upload(c);
      }
      {
        const HostTensor *w = L.get("generator.projector.weight"), *bb = L.get("generator.projector.bias");
        if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
        Builder b;
        const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {1024, 0};
        b.init(512, 1, t, pd, cp);
        b.rows(0, *w, 0, 512, 0, 1024, 0, 1);
        b.set_bias(*bb, 0, 512, 0, 1);
        c.gemm["gen.projector"] = b.upload(c);
      }
      for (int i = 0; i < 4; ++i) {
        const std::string q = "generator.convnext." + std::to_string(i) + ".";
        const std::string nm = "gen.cnx" + std::to_string(i);
        const HostTensor *dw = L.get(q + "dwconv.weight"), *db = L.get(q + "dwconv.bias");
        const HostTensor *nw = L.get(q + "norm.fc.weight"), *nb = L.get(q + "norm.fc.bias");
        const HostTensor *w1 = L.get(q + "pwconv1.weight"), *b1 = L.get(q + "pwconv1.bias");
        const HostTensor *w2 = L.get(q + "pwconv2.weight"), *b2 = L.get(q + "pwconv2.bias");
        const HostTensor *gg = L.get(q + "grn.gamma"), *gb = L.get(q + "grn.beta");
        if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
        const int k = (int)dw->dim(2);
        std::vector<float> wt((size_t)k * 512);
        for (int ch = 0; ch < 512; ++ch)
          for (int j = 0; j < k; ++j) wt[(size_t)j * 512 + ch] = dw->data[(size_t)ch * k + j];
        put_vec(c, nm + ".dw_w", wt);
        put_vec(c, nm + ".dw_b", db->data);
        put_vec(c, nm + ".grn_gamma", gg->data);
        put_vec(c, nm + ".grn_beta", gb->data);
        c.vec_len[nm + ".dw_k"] = k;
        c.fc_off[nm + ".norm"] = gfc.add(*nw, *nb);
        {
          Builder b;
          const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {512, 0};
          b.init(1536, 1, t, pd, cp);
          b.rows(0, *w1, 0, 1536, 0, 512, 0, 1);
          b.set_bias(*b1, 0, 1536, 0, 1);
          c.gemm[nm + ".pw1"] = b.upload(c);
        }
        {
          Builder b;
          const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {1536, 0};
          b.init(512, 1, t, pd, cp);
          b.rows(0, *w2, 0, 512, 0, 1536, 0, 1);
          b.set_bias(*b2, 0, 512, 0, 1);
          c.gemm[nm + ".pw2"] = b.upload(c);
          // GRN folded into pw2 (16-bit mode): pw2(x * s + beta) = (W diag(s)) x + (W beta + bias)
          std::vector<float> bg(512);
          for (int n = 0; n < 512; ++n) {
            double acc = b2->data[n];
            for (int kk = 0; kk < 1536; ++kk) acc += (double)w2->data[(size_t)n * 1536 + kk] * (double)gb->data[kk];
            bg[n] = (float)acc;
          }
          put_vec(c, nm + ".pw2_bias_grn", bg);
        }
      }
      const char* heads[2] = {"amp", "phase"};
      for (int h = 0; h < 2; ++h) {
        const std::string hn = heads[h];
        const HostTensor *nw = L.get("generator." + hn + "_final_layer_norm.fc.weight");
        const HostTensor *nb = L.get("generator." + hn + "_final_layer_norm.fc.bias");
        const HostTensor *w = L.get("generator." + hn + "_output_conv.weight");
        const HostTensor *bb = L.get("generator." + hn + "_output_conv.bias");
        if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
        c.fc_off["gen." + hn + "_norm"] = gfc.add(*nw, *nb);
        // K-seg 0: AdaLN(x) (512 ch), K-seg 1: the prior conv output (256 ch) -- torch.cat order
        Builder b;
        const int t[2] = {7, 7}, pd[2] = {3, 3}, cp[2] = {512, 256};
        b.init(1025, 2, t, pd, cp);
        b.rows(0, *w, 0, 1025, 0, 512, 0, 1);
        b.rows(1, *w, 0, 1025, 512, 256, 0, 1);
        b.set_bias(*bb, 0, 1025, 0, 1);
        c.gemm["gen." + hn + "_head"] = b.upload(c);
      }
      c.J_gen = gfc.J;
      put_vec(c, "gen.fc_W", gfc.W);
      put_vec(c, "gen.fc_b", gfc.b);
    }
    // ---- STFT / iSTFT bases (TorchSTFT: n_fft 2048, hop 75, hann(1200) centred in the frame)
    {
      std::vector<double> win(1200);
      std::vector<float> wsq(1200);
      for (int n = 0; n < 1200; ++n) {
        win[n] = (double)(float)(0.5 - 0.5 * std::cos(2.0 * kPi * n / 1200.0));  // torch.hann_window (periodic), fp32
        wsq[n] = (float)win[n] * (float)win[n];
      }
      put_vec(c, "wsq", wsq);
      // exact twiddles: cos / sin of 2 pi m / 2048 with m reduced to the first octant, so that the
      // imaginary rows of the DC and Nyquist bins are exactly zero (atan2(+0, -x) = +pi, as pocketfft)
      auto cs = [](int k, int np, double* c_out, double* s_out) {
        int m = (int)((long long)k * np % 2048);
        const int q = m / 512;           // quadrant
        const int r = m % 512;
        double c0, s0;
        if (r == 0) { c0 = 1.0; s0 = 0.0; }
        else if (r == 256) { c0 = s0 = std::sqrt(0.5); }
        else if (r < 256) { c0 = std::cos(2.0 * kPi * r / 2048.0); s0 = std::sin(2.0 * kPi * r / 2048.0); }
        else { c0 = std::sin(2.0 * kPi * (512 - r) / 2048.0); s0 = std::cos(2.0 * kPi * (512 - r) / 2048.0); }
        switch (q) {
          case 0: *c_out = c0; *s_out = s0; break;
          case 1: *c_out = -s0; *s_out = c0; break;
          case 2: *c_out = -c0; *s_out = -s0; break;
          default: *c_out = s0; *s_out = -c0; break;
        }
      };
      // forward STFT: a shared-memory FFT (stft_fft.cu), not a contraction -- twiddle tables + the analysis window.
      // The window is torch.hann_window(1200) as the reference's non-persistent `generator.stft.window` buffer holds it
      // (generator.py:30) when the host passed it, else the correctly rounded formula.
      {
        std::vector<float> w32(1200);
        const HostTensor* hw = c.has("speech_predictor/generator.stft.window") ? L.get("generator.stft.window") : nullptr;
        for (int n = 0; n < 1200; ++n) w32[n] = (hw != nullptr && hw->numel() == 1200) ? hw->data[n] : (float)win[n];
        put_vec(c, "win", w32);
        std::vector<float> z(2 * (2 * 1024 + 2 * 513), 0.f);      // double2 tables: exp(-2 pi i m / 1024), exp(-2 pi i k / 2048)
        put_vec(c, "stft_tw", z);
        if (stft_twiddles(reinterpret_cast<double2*>(c.vec["stft_tw"]), reinterpret_cast<double2*>(c.vec["stft_tw"]) + 1024,
                          nullptr) != cudaSuccess)
          return fail(c, -5, "stft twiddle kernel launch failed");
      }
      {  // inverse: irfft (1/N, Hermitian doubling, imag of DC / Nyquist ignored) x window
        Builder b;
        const int t[2] = {1, 0}, pd[2] = {0, 0}, cp[2] = {2112, 0};
        b.init(1200, 1, t, pd, cp);
        for (int n = 0; n < 1200; ++n)
          for (int k = 0; k <= 1024; ++k) {
            const double ck = (k == 0 || k == 1024) ? 1.0 : 2.0;
            double ca, sa;
            cs(k, 424 + n, &ca, &sa);
            b.W[(size_t)n * b.Ktot + 2 * k] = (float)(ck * win[n] * ca / 2048.0);
            b.W[(size_t)n * b.Ktot + 2 * k + 1] =
                (k == 0 || k == 1024 || sa == 0.0) ? 0.f : (float)(-ck * win[n] * sa / 2048.0);
          }
        b.k_real = 2050;
        c.gemm["istft"] = b.upload(c);
      }
    }
  }

  // ======================================================================= pitch_energy_predictor
  Loader P{c, "pitch_energy_predictor"};
  if (c.has("pitch_energy_predictor/F0.0.conv1.bias")) {
    FcCat fc;
    const char* tw[2] = {"F0", "N"};
    for (int t = 0; t < 2; ++t)
      for (int i = 0; i < 3; ++i) {
        const std::string p = std::string(tw[t]) + "." + std::to_string(i) + ".";
        if (!pack_ada_block(c, P, p, std::string("pe.") + tw[t] + std::to_string(i), 320, 320, 320, fc))
          return fail(c, -3, "missing weight " + P.missing);
      }
    c.J_pe = fc.J;
    put_vec(c, "pe.fc_W", fc.W);
    put_vec(c, "pe.fc_b", fc.b);
    const HostTensor *wf = P.get("F0_proj.weight"), *bf = P.get("F0_proj.bias");
    const HostTensor *wn = P.get("N_proj.weight"), *bn = P.get("N_proj.bias");
    if (!P.ok) return fail(c, -3, "missing weight " + P.missing);
    put_vec(c, "pe.F0_proj_w", wf->data);
    put_vec(c, "pe.F0_proj_b", bf->data);
    put_vec(c, "pe.N_proj_w", wn->data);
    put_vec(c, "pe.N_proj_b", bn->data);
  }
  // ======================================================================= token-rate encoders
  {
    struct Slot { const char* module; const char* kp; const char* nm; };
    const Slot slots[3] = {{"speech_predictor", "text_encoder.", "te.sp"},
                           {"pe_text_encoder", "", "te.pe"},
                           {"duration_predictor", "text_encoder.", "te.dp"}};
    for (const Slot& sl : slots) {
      if (!c.has(std::string(sl.module) + "/" + sl.kp + "emb.weight")) continue;
      Loader T{c, sl.module};
      if (!pack_text_encoder(c, T, sl.kp, sl.nm)) return fail(c, -3, "missing weight " + T.missing);
    }
    const Slot pslots[2] = {{"duration_predictor", "prosody_encoder.", "pr.dp"},
                            {"pitch_energy_predictor", "prosody_encoder.", "pr.pe"}};
    for (const Slot& sl : pslots) {
      if (!c.has(std::string(sl.module) + "/" + sl.kp + "attn_layers.0.conv_q.weight")) continue;
      Loader T{c, sl.module};
      if (!pack_prosody(c, T, sl.kp, sl.nm)) return fail(c, -3, "missing weight " + T.missing);
    }
    const Slot sslots[3] = {{"speech_predictor", "style_encoder.", "se.sp"},
                            {"pe_text_style_encoder", "", "se.pe"},
                            {"duration_predictor", "style_encoder.", "se.dp"}};
    for (const Slot& sl : sslots) {
      if (!c.has(std::string(sl.module) + "/" + sl.kp + "conv_in.weight")) continue;
      Loader T{c, sl.module};
      if (!pack_style_encoder(c, T, sl.kp, sl.nm)) return fail(c, -3, "missing weight " + T.missing);
    }
    if (c.has("pitch_energy_predictor/cross_attention.conv_q.weight")) {
      Loader T{c, "pitch_energy_predictor"};
      if (!pack_pe_cross(c, T, "pe.cross")) return fail(c, -3, "missing weight " + T.missing);
    }
    if (c.has("duration_predictor/duration_proj.linear_layer.weight")) {
      Loader T{c, "duration_predictor"};
      if (!pack_simple(c, T, "duration_proj.linear_layer", "pr.dp.head", 0))
        return fail(c, -3, "missing weight " + T.missing);
    }
  }
  for (auto& kv : c.gemm) kv.second.name = kv.first;
  if (cudaDeviceSynchronize() != cudaSuccess) return fail(c, -5, "cuda error while uploading weights");
  c.finalized = true;
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Snake AdaptiveGeneratorBlock (ada_norm.py:11-127): three (AdaIN, snake, dilated conv k,
// AdaIN, snake, conv k) residual stages.  Packed on first use by stts_generator_block_snake from the tensors loaded
// under `module` with the reference's state_dict keys:
//   gemm  <module>.c1.<i> (dilation d_i, pad (k d_i - d_i) / 2), <module>.c2.<i> (dilation 1)
//   vec   <module>.fc_W / fc_b: the six style Linear layers stacked in the order n1.0, n2.0, n1.1, ... (fc_off);
//         <module>.a1.<i> / a2.<i>: alpha [C]
int pack_snake_block(Context& c, const std::string& module, int channels, int k, const int* dil) {
  if (channels <= 0 || channels % 64 != 0 || channels > 1024) return fail(c, -2, "snake block: channels must be a multiple of 64, <= 1024");
  if (k < 1 || (k & 1) == 0 || k > 15) return fail(c, -2, "snake block: kernel_size must be odd, <= 15");
  Loader L{c, module};
  FcCat fc;
  SnakeCfg cfg;
  cfg.C = channels; cfg.k = k;
  for (int i = 0; i < 3; ++i) {
    const std::string si = std::to_string(i);
    if (dil[i] < 1 || (k * dil[i] - dil[i]) / 2 > 32) return fail(c, -2, "snake block: padding (k d - d) / 2 must be <= 32 rows");
    cfg.dil[i] = dil[i];
    for (int which = 0; which < 2; ++which) {
      const std::string grp = which == 0 ? "convs1." : "convs2.";
      const HostTensor w = L.weight(grp + si + ".");
      const HostTensor* b = L.get(grp + si + ".bias");
      if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
      if (w.dim(0) != channels || w.dim(1) != channels || w.dim(2) != k)
        return fail(c, -2, "snake block: " + module + "/" + grp + si + " is not [C, C, k]");
      const int d = which == 0 ? dil[i] : 1;
      Builder bd;
      const int t[2] = {k, 0}, pd[2] = {(k * d - d) / 2, 0}, cp[2] = {channels, 0};
      bd.init(channels, 1, t, pd, cp);
      bd.rows(0, w, 0, channels, 0, channels, 0, 1);
      bd.set_bias(*b, 0, channels, 0, 1);
      PackedGemm pg = bd.upload(c);
      pg.dil[0] = d;
      const std::string name = module + (which == 0 ? ".c1." : ".c2.") + si;
      pg.name = name;
      c.gemm[name] = pg;
      const std::string an = which == 0 ? "adain1." : "adain2.";
      const HostTensor *fw = L.get(an + si + ".fc.weight"), *fb = L.get(an + si + ".fc.bias");
      const HostTensor* al = L.get((which == 0 ? "alpha1." : "alpha2.") + si);
      if (!L.ok) return fail(c, -3, "missing weight " + L.missing);
      if (fw->dim(0) != 2 * channels || fw->dim(1) != 64 || al->numel() != channels)
        return fail(c, -2, "snake block: AdaIN fc must be [2 C, 64] and alpha [1, C, 1]");
      c.fc_off[module + (which == 0 ? ".n1." : ".n2.") + si] = fc.add(*fw, *fb);
      put_vec(c, module + (which == 0 ? ".a1." : ".a2.") + si, al->data);
    }
  }
  cfg.J = fc.J;
  put_vec(c, module + ".fc_W", fc.W);
  put_vec(c, module + ".fc_b", fc.b);
  if (cudaDeviceSynchronize() != cudaSuccess) return fail(c, -5, "cuda error while uploading the snake block");
  c.snake[module] = cfg;
  return 0;
}

}  // namespace stts
