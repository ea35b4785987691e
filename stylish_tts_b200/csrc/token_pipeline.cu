// Token-rate stage orchestration on the packed token layout.
//   TextEncoder.forward   train/models/text_encoder.py:433-462
//     emb * sqrt(C) -> ConvReluNorm prenet (36-86) -> Encoder (332-393): n x [MHA + LN, FFN + LN] -> proj_m
// All dense contractions are conv-as-GEMM launches of the island class (fp32-grade in every
// precision mode: the duration logits decide integer durations).
#include <algorithm>
#include <cmath>
#include <cstring>

#include "engine.h"

namespace stts {

namespace {
size_t up1k(size_t x) { return (x + 1023) / 1024 * 1024; }
}

int set_token_layout(Context& c, int n_seg, const int* len_t) {
  if (n_seg <= 0 || n_seg > 1024 || len_t == nullptr) return fail(c, -1, "set_token_layout: bad arguments");
  std::vector<int> sig(len_t, len_t + n_seg);
  int max_t = 0;
  for (int b = 0; b < n_seg; ++b) {
    if (len_t[b] <= 0) return fail(c, -2, "utterance token count must be > 0");
    max_t = std::max(max_t, len_t[b]);
  }
  if (max_t > 2048) return fail(c, -6, "utterance longer than 2048 tokens (attention score rows live in shared memory)");
  if (sig == c.tsig && c.tarena != nullptr) return 0;
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return cuda_fail(c, e, "set_token_layout sync");
  tc_map_cache_clear();
  build_layout(c.LT, n_seg, len_t);
  c.max_t = max_t;
  const int R = c.LT.R, B = n_seg, nt = c.LT.n_tiles;
  struct Want { const char* name; int ld; };
  const Want wants[] = {{"TA", 128}, {"TB", 128}, {"TC", 128}, {"TX", 128}, {"TO", 128}, {"TQKV", 384},
                        {"TH", 512}, {"TMU", 256},
                        // prosody encoders (hid <= 320)
                        {"PX", 320}, {"PY", 320}, {"PB", 320}, {"PO", 320}, {"PQKV", 960}, {"PH", 640}, {"PL", 64},
                        // style encoders (64 / 256 channels)
                        {"SX", 64}, {"SY", 64}, {"SH", 256}};
  // three independent buffer sets ("slots"): the speech / pitch-energy / duration chains only share
  // the tokens, so the host may enqueue them on three streams (synth.py) without aliasing scratch
  // every buffer also carries fp16 split twins g / l (same pitch): in 16-bit mode the token-rate GEMMs run the
  // three-pass fp16 split (f16x3, ~2^-22: fp32-grade, the duration logits decide integer durations) on them
  size_t used = 0;
  std::map<std::string, size_t> off;
  for (int slot = 0; slot < 3; ++slot)
    for (const Want& w : wants) {
      const std::string key = std::string(w.name) + "#" + std::to_string(slot);
      off[key] = used;
      used += up1k((size_t)R * w.ld * 4);
      off[key + ".g"] = used;
      used += up1k((size_t)R * w.ld * 2);
      off[key + ".l"] = used;
      used += up1k((size_t)R * w.ld * 2);
    }
  const char* ints[] = {"tile_seg", "tile_nvalid", "seg_start", "seg_len", "cu"};
  const size_t int_bytes[] = {(size_t)nt * 4, (size_t)nt * 4, (size_t)B * 4, (size_t)B * 4, (size_t)(B + 1) * 4};
  for (int i = 0; i < 5; ++i) { off[ints[i]] = used; used += up1k(int_bytes[i]); }
  for (int slot = 0; slot < 3; ++slot) {   // style-FC rows of the prosody AdaLNs (J <= 8192)
    off["GBT#" + std::to_string(slot)] = used;
    used += up1k((size_t)B * 8192 * 4);
    // style-encoder statistics: per-tile fp64 partials (sum, sum sq) x 256 ch, per-utterance vectors
    off["SSTAT#" + std::to_string(slot)] = used;
    used += up1k((size_t)nt * 256 * 2 * 8);
    off["SVEC#" + std::to_string(slot)] = used;
    used += up1k((size_t)B * 256 * 4 * 4);
  }
  off["GBC"] = used;                        // style-FC rows of the cross-attention query / key AdaLNs
  used += up1k((size_t)B * 2048 * 4);

  if (used > c.tarena_bytes) {
    if (c.tarena) cudaFree(c.tarena);
    c.tarena = nullptr; c.tarena_bytes = 0;
    e = cudaMalloc(&c.tarena, used);
    if (e != cudaSuccess) return cuda_fail(c, e, "token workspace cudaMalloc");
    c.tarena_bytes = used;
  }
  e = cudaMemset(c.tarena, 0, used);
  if (e != cudaSuccess) return cuda_fail(c, e, "token workspace memset");
  c.tact.clear(); c.traw.clear();
  for (int slot = 0; slot < 3; ++slot)
    for (const Want& w : wants) {
      const std::string key = std::string(w.name) + "#" + std::to_string(slot);
      Act a; a.ld = w.ld; a.rows = R; a.f = reinterpret_cast<float*>(c.tarena + off[key]); a.h = nullptr;
      a.g = reinterpret_cast<__half*>(c.tarena + off[key + ".g"]);
      a.l = reinterpret_cast<__half*>(c.tarena + off[key + ".l"]);
      a.ldg = w.ld; a.gscale = 1.f;
      c.tact[key] = a;
    }
  Layout& L = c.LT;
  const void* srcs[] = {L.tile_seg.data(), L.tile_nvalid.data(), L.seg_start.data(), L.seg_len.data(), L.cu.data()};
  for (int i = 0; i < 5; ++i) {
    c.traw[ints[i]] = c.tarena + off[ints[i]];
    cudaMemcpy(c.traw[ints[i]], srcs[i], int_bytes[i], cudaMemcpyHostToDevice);
  }
  L.d_tile_seg = (int*)c.traw["tile_seg"]; L.d_tile_nvalid = (int*)c.traw["tile_nvalid"];
  L.d_seg_start = (int*)c.traw["seg_start"]; L.d_seg_len = (int*)c.traw["seg_len"]; L.d_cu = (int*)c.traw["cu"];
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return cuda_fail(c, e, "set_token_layout upload");
  for (int slot = 0; slot < 3; ++slot)
    for (const char* nmv : {"GBT#", "SSTAT#", "SVEC#"})
      c.traw[nmv + std::to_string(slot)] = c.tarena + off[nmv + std::to_string(slot)];
  c.traw["GBC"] = c.tarena + off["GBC"];
  c.tsig = sig;
  return 0;
}

namespace {
constexpr int kRopePositions = 16384;   // longest utterance in tokens or hop-300 frames (54 s of audio per 4800 frames)
// (cos, sin) table of the partial RoPE for `rd` rotated features, built on first use (the warm-up
// passes of a CUDA-graph capture run eagerly, so no allocation happens while capturing)
const float2* rope_tab(Context& c, int rd, cudaStream_t st) {
  auto it = c.rope_tab.find(rd);
  if (it != c.rope_tab.end()) return it->second;
  float2* tab = nullptr;
  if (cudaMalloc(&tab, (size_t)kRopePositions * (rd / 2) * sizeof(float2)) != cudaSuccess) return nullptr;
  if (rope_table(tab, kRopePositions, rd, st) != cudaSuccess) { cudaFree(tab); return nullptr; }
  c.rope_tab[rd] = tab;
  return tab;
}

// tw: also write the fp16 split twins of the output (16-bit mode, when a GEMM reads it next)
Epi mk_epi(int kind, const Act& dst, int col0, bool tw = false) {
  Epi e;
  std::memset(&e, 0, sizeof(e));
  e.kind = kind; e.out = dst.f; e.ldo = dst.ld; e.col0 = col0; e.scale = 1.f;
  if (tw && dst.g != nullptr) { e.outg = dst.g; e.outl = dst.l; e.ldh = dst.ldg; e.colh0 = col0; }
  return e;
}
// LayerNorm / AdaptiveLayerNorm fused into the epilogue of the GEMM that produces its input (EPI_LN, gemm_tc.cu):
// tensor-core modes, rows of <= 256 channels.  STTS_LN_FUSED=0 keeps the separate row_ln launch (A/B runs).
bool ln_fused(const Context& c, int width) {
  static const bool on = [] { const char* e = std::getenv("STTS_LN_FUSED"); return e == nullptr || e[0] != '0'; }();
  return on && c.precision != 0 && width <= 256 && width % 4 == 0;
}
// out = act(affine(LN(acc + bias + res))); res may be the destination itself
Epi mk_ln_epi(const Act& dst, bool tw, const float* g, const float* b, int seg_stride, float eps, int relu, const Act* res) {
  Epi e;
  std::memset(&e, 0, sizeof(e));
  e.kind = EPI_LN; e.out = dst.f; e.ldo = dst.ld; e.col0 = 0; e.scale = 1.f;
  if (tw && dst.g != nullptr) { e.outg = dst.g; e.outl = dst.l; e.ldh = dst.ldg; e.colh0 = 0; }
  e.ln_g = g; e.ln_b = b; e.ln_seg_stride = seg_stride; e.ln_eps = eps; e.ln_relu = relu;
  if (res != nullptr) { e.res = res->f; e.ldr = res->ld; e.rcol0 = 0; }
  return e;
}
struct Tw {                      // twin pointers of an Act in 16-bit mode, nulls otherwise
  __half* g; __half* l;
  Tw(const Context& c, const Act& a) : g(c.precision == 1 ? a.g : nullptr), l(c.precision == 1 ? a.l : nullptr) {}
};
}  // namespace

// d_tokens: int32 [sum T] utterances concatenated.  Outputs (either may be null), token-major dense:
// d_mu_tok [sum T, inter] = proj_m(h), d_x_tok [sum T, C] = h.   Also leaves h in TX and mu in TMU.
int run_text_encoder(Context& c, const std::string& nm, const int* d_tokens, float* d_mu_tok, float* d_x_tok,
                     cudaStream_t st) {
  if (!c.finalized || c.vec.count(nm + ".emb") == 0) return fail(c, -3, "text encoder weights not loaded: " + nm);
  if (c.LT.total() == 0 || c.tact.count("TX#0") == 0) return fail(c, -4, "no token layout set");
  const RowMap map = c.LT.map();
  const std::string sl = nm == "te.sp" ? "#0" : (nm == "te.pe" ? "#1" : "#2");
  Act &ta = c.tact["TA" + sl], &tb = c.tact["TB" + sl], &tcb = c.tact["TC" + sl], &tx = c.tact["TX" + sl];
  Act &to = c.tact["TO" + sl], &tqkv = c.tact["TQKV" + sl], &th = c.tact["TH" + sl], &tmu = c.tact["TMU" + sl];
  const int C = 128, H = 8, D = C / H, rd = D / 2;
  const bool t16 = c.precision == 1;             // 16-bit mode: every GEMM operand below carries fp16 split twins
  STTS_CK(embed_tokens(d_tokens, c.LT.d_cu, c.vec[nm + ".emb"], C, c.vec_len[nm + ".vocab"], map, ta.f, ta.ld, st,
                       Tw(c, ta).g, Tw(c, ta).l));
  // ---- prenet: 3 x (conv k5 -> channel LayerNorm (eps 1e-4) -> ReLU), then x + proj(y)
  const Act* y = &ta;
  Act* pp[2] = {&tb, &tcb};
  for (int i = 0; i < 3; ++i) {
    const std::string n = nm + ".pre" + std::to_string(i);
    Act& dst = *pp[i & 1];
    if (ln_fused(c, C)) {
      STTS_RET(run_gemm(c, c.gemm[n], GC_ISLAND, y, 0, nullptr, 0,
                        mk_ln_epi(dst, t16, c.vec[n + ".g"], c.vec[n + ".b"], 0, 1e-4f, 1, nullptr), map, st));
    } else {
      STTS_RET(run_gemm(c, c.gemm[n], GC_ISLAND, y, 0, nullptr, 0, mk_epi(EPI_BIAS, dst, 0), map, st));
      STTS_CK(row_ln(dst.f, dst.ld, C, map, c.vec[n + ".g"], c.vec[n + ".b"], 0, 1e-4f, 1, dst.f, dst.ld, st,
                     Tw(c, dst).g, Tw(c, dst).l));
    }
    y = &dst;
  }
  {
    Epi e = mk_epi(EPI_ADD, tx, 0, t16);
    e.res = ta.f; e.ldr = ta.ld; e.rcol0 = 0;
    STTS_RET(run_gemm(c, c.gemm[nm + ".preproj"], GC_ISLAND, y, 0, nullptr, 0, e, map, st));
  }
  // ---- encoder layers
  const int n_layers = c.vec_len[nm + ".layers"];
  const float scale = 1.0f / std::sqrt((float)D);
  for (int l = 0; l < n_layers; ++l) {
    const std::string n = nm + ".l" + std::to_string(l);
    STTS_RET(run_gemm(c, c.gemm[n + ".qkv"], GC_ISLAND, &tx, 0, nullptr, 0, mk_epi(EPI_BIAS, tqkv, 0), map, st));
    // partial RoPE of q / k is applied inside the attention kernel while it stages the rows
    STTS_CK(attention(tqkv.f, tqkv.ld, 0, tqkv.f, tqkv.ld, C, tqkv.f, tqkv.ld, 2 * C, H, D, scale, rd, rope_tab(c, rd, st), map, map, c.max_t,
                      c.max_t, nullptr, nullptr, 0, to.f, to.ld, st, Tw(c, to).g, Tw(c, to).l, to.ldg));
    if (ln_fused(c, C)) {       // x = LN(x + proj(...)) in the projection's epilogue, in place
      STTS_RET(run_gemm(c, c.gemm[n + ".o"], GC_ISLAND, &to, 0, nullptr, 0,
                        mk_ln_epi(tx, t16, c.vec[n + ".n1.g"], c.vec[n + ".n1.b"], 0, 1e-4f, 0, &tx), map, st));
    } else {
      Epi e = mk_epi(EPI_ADD, tb, 0);
      e.res = tx.f; e.ldr = tx.ld; e.rcol0 = 0;
      STTS_RET(run_gemm(c, c.gemm[n + ".o"], GC_ISLAND, &to, 0, nullptr, 0, e, map, st));
      STTS_CK(row_ln(tb.f, tb.ld, C, map, c.vec[n + ".n1.g"], c.vec[n + ".n1.b"], 0, 1e-4f, 0, tx.f, tx.ld, st,
                     Tw(c, tx).g, Tw(c, tx).l));
    }
    STTS_RET(run_gemm(c, c.gemm[n + ".ff1"], GC_ISLAND, &tx, 0, nullptr, 0, mk_epi(EPI_RELU, th, 0, t16), map, st));
    if (ln_fused(c, C)) {       // x = LN(x + proj(...)) in the projection's epilogue, in place
      STTS_RET(run_gemm(c, c.gemm[n + ".ff2"], GC_ISLAND, &th, 0, nullptr, 0,
                        mk_ln_epi(tx, t16, c.vec[n + ".n2.g"], c.vec[n + ".n2.b"], 0, 1e-4f, 0, &tx), map, st));
    } else {
      Epi e = mk_epi(EPI_ADD, tb, 0);
      e.res = tx.f; e.ldr = tx.ld; e.rcol0 = 0;
      STTS_RET(run_gemm(c, c.gemm[n + ".ff2"], GC_ISLAND, &th, 0, nullptr, 0, e, map, st));
      STTS_CK(row_ln(tb.f, tb.ld, C, map, c.vec[n + ".n2.g"], c.vec[n + ".n2.b"], 0, 1e-4f, 0, tx.f, tx.ld, st,
                     Tw(c, tx).g, Tw(c, tx).l));
    }
  }
  const PackedGemm& pm = c.gemm[nm + ".projm"];
  STTS_RET(run_gemm(c, pm, GC_ISLAND, &tx, 0, nullptr, 0, mk_epi(EPI_BIAS, tmu, 0, t16), map, st));   // mu feeds the style conv
  if (d_mu_tok) STTS_CK(rows_to_dense(tmu.f, tmu.ld, pm.N, map, c.LT.d_cu, d_mu_tok, st));
  if (d_x_tok) STTS_CK(rows_to_dense(tx.f, tx.ld, C, map, c.LT.d_cu, d_x_tok, st));
  return 0;
}

// ProsodyEncoder.forward, prosody_encoder.py:63-81.  d_enc_tok token-major dense [sum T, d_model];
// d_style [n_seg, 64]; d_out_tok [sum T, hid] (hid = d_model + 64); optional d_logits_tok [sum T, 16]
// (DurationPredictor.duration_proj, duration_predictor.py:30-36) when the head was packed.
int run_prosody_encoder(Context& c, const std::string& nm, const float* d_enc_tok, const float* d_style,
                        float* d_out_tok, float* d_logits_tok, cudaStream_t st) {
  if (!c.finalized || c.vec.count(nm + ".fc_W") == 0) return fail(c, -3, "prosody encoder weights not loaded: " + nm);
  if (c.LT.total() == 0 || c.tact.count("PX#0") == 0) return fail(c, -4, "no token layout set");
  const std::string sl = nm == "pr.pe" ? "#1" : "#2";
  const RowMap map = c.LT.map();
  const int hid = c.vec_len[nm + ".hid"], dm = c.vec_len[nm + ".dmodel"], J = c.vec_len[nm + ".J"];
  const int n_layers = c.vec_len[nm + ".layers"], H = 2, D = hid / H, rd = D / 2, sty = hid - dm;
  if (hid > 320 || J > 8192 || sty != 64) return fail(c, -6, "prosody encoder: unsupported width");
  Act &px = c.tact["PX" + sl], &py = c.tact["PY" + sl], &pb = c.tact["PB" + sl], &po = c.tact["PO" + sl];
  Act &pqkv = c.tact["PQKV" + sl], &ph = c.tact["PH" + sl], &pl = c.tact["PL" + sl];
  float* gb = reinterpret_cast<float*>(c.traw["GBT" + sl]);
  STTS_CK(style_fc(c.vec[nm + ".fc_W"], c.vec[nm + ".fc_b"], d_style, J, 64, c.LT.n_seg, gb, st));
  const bool t16 = c.precision == 1;
  STTS_CK(dense_to_rows(d_enc_tok, dm, map, c.LT.d_cu, px.f, px.ld, 0, st, Tw(c, px).g, Tw(c, px).l));
  STTS_CK(seg_broadcast(d_style, sty, map, px.f, px.ld, dm, st, Tw(c, px).g, Tw(c, px).l));
  const float scale = 1.0f / std::sqrt((float)D);
  for (int l = 0; l < n_layers; ++l) {
    const std::string n = nm + ".l" + std::to_string(l);
    STTS_RET(run_gemm(c, c.gemm[n + ".qkv"], GC_ISLAND, &px, 0, nullptr, 0, mk_epi(EPI_BIAS, pqkv, 0), map, st));
    STTS_CK(attention(pqkv.f, pqkv.ld, 0, pqkv.f, pqkv.ld, hid, pqkv.f, pqkv.ld, 2 * hid, H, D, scale, rd, rope_tab(c, rd, st), map, map,
                      c.max_t, c.max_t, nullptr, nullptr, 0, po.f, po.ld, st, Tw(c, po).g, Tw(c, po).l, po.ldg));
    if (ln_fused(c, hid)) {
      STTS_RET(run_gemm(c, c.gemm[n + ".o"], GC_ISLAND, &po, 0, nullptr, 0,
                        mk_ln_epi(py, t16, gb + c.fc_off[n + ".n1"], nullptr, J, 1e-5f, 0, &px), map, st));
    } else {
      Epi e = mk_epi(EPI_ADD, pb, 0);
      e.res = px.f; e.ldr = px.ld; e.rcol0 = 0;
      STTS_RET(run_gemm(c, c.gemm[n + ".o"], GC_ISLAND, &po, 0, nullptr, 0, e, map, st));
      STTS_CK(row_ln(pb.f, pb.ld, hid, map, gb + c.fc_off[n + ".n1"], nullptr, J, 1e-5f, 0, py.f, py.ld, st,
                     Tw(c, py).g, Tw(c, py).l));
    }
    STTS_RET(run_gemm(c, c.gemm[n + ".ff1"], GC_ISLAND, &py, 0, nullptr, 0, mk_epi(EPI_RELU, ph, 0, t16), map, st));
    if (ln_fused(c, hid)) {
      STTS_RET(run_gemm(c, c.gemm[n + ".ff2"], GC_ISLAND, &ph, 0, nullptr, 0,
                        mk_ln_epi(py, t16, gb + c.fc_off[n + ".n2"], nullptr, J, 1e-5f, 0, &py), map, st));
    } else {
      Epi e = mk_epi(EPI_ADD, pb, 0);
      e.res = py.f; e.ldr = py.ld; e.rcol0 = 0;
      STTS_RET(run_gemm(c, c.gemm[n + ".ff2"], GC_ISLAND, &ph, 0, nullptr, 0, e, map, st));
      STTS_CK(row_ln(pb.f, pb.ld, hid, map, gb + c.fc_off[n + ".n2"], nullptr, J, 1e-5f, 0, py.f, py.ld, st,
                     Tw(c, py).g, Tw(c, py).l));
    }
    // x = cat([proj(x), style]): the style columns of PX (and of its twins) are still in place
    STTS_RET(run_gemm(c, c.gemm[n + ".proj"], GC_ISLAND, &py, 0, nullptr, 0, mk_epi(EPI_BIAS, px, 0, t16), map, st));
  }
  if (d_out_tok) STTS_CK(rows_to_dense(px.f, px.ld, hid, map, c.LT.d_cu, d_out_tok, st));
  if (d_logits_tok) {
    if (c.gemm.count(nm + ".head") == 0) return fail(c, -3, "duration head not loaded");
    STTS_RET(run_gemm(c, c.gemm[nm + ".head"], GC_ISLAND, &px, 0, nullptr, 0, mk_epi(EPI_BIAS, pl, 0), map, st));
    STTS_CK(rows_to_dense(pl.f, pl.ld, 16, map, c.LT.d_cu, d_logits_tok, st));
  }
  return 0;
}

// TextStyleEncoder.forward, text_style_encoder.py:20-26: conv k7 -> n x BasicConvNeXtBlock (dwconv k7,
// LayerNorm eps 1e-6, Linear + GELU, GRN over time, Linear, residual; conv_next.py:18-51) -> mean over
// the utterance.  d_enc_tok token-major dense [sum T, inter] or NULL = use the mu rows the text encoder
// of the same chain left in its TMU buffer.  d_style_out [n_seg, 64].
int run_style_encoder(Context& c, const std::string& nm, const float* d_enc_tok, float* d_style_out, cudaStream_t st) {
  if (!c.finalized || c.gemm.count(nm + ".in") == 0) return fail(c, -3, "style encoder weights not loaded: " + nm);
  if (c.LT.total() == 0 || c.tact.count("SX#0") == 0) return fail(c, -4, "no token layout set");
  const RowMap map = c.LT.map();
  const std::string sl = nm == "se.sp" ? "#0" : (nm == "se.pe" ? "#1" : "#2");
  Act &tmu = c.tact["TMU" + sl], &sx = c.tact["SX" + sl], &sy = c.tact["SY" + sl], &sh = c.tact["SH" + sl];
  const PackedGemm& gin = c.gemm[nm + ".in"];
  const int inter = (int)(gin.k_real / gin.taps[0] + 0.5);
  const bool t16 = c.precision == 1;
  if (d_enc_tok) STTS_CK(dense_to_rows(d_enc_tok, inter, map, c.LT.d_cu, tmu.f, tmu.ld, 0, st, Tw(c, tmu).g, Tw(c, tmu).l));
  STTS_RET(run_gemm(c, gin, GC_ISLAND, &tmu, 0, nullptr, 0, mk_epi(EPI_BIAS, sx, 0), map, st));
  double* stat = reinterpret_cast<double*>(c.traw["SSTAT" + sl]);
  float* vec = reinterpret_cast<float*>(c.traw["SVEC" + sl]);
  const int B = c.LT.n_seg;
  float *gx = vec, *gscale = vec + (size_t)B * 256, *mean = vec + (size_t)2 * B * 256, *rstd = vec + (size_t)3 * B * 256;
  const int n_blocks = c.vec_len[nm + ".blocks"];
  for (int i = 0; i < n_blocks; ++i) {
    const std::string n = nm + ".b" + std::to_string(i);
    STTS_CK(dwconv_rows(sx.f, sx.ld, 64, c.vec_len[n + ".dw_k"], c.vec[n + ".dw_w"], c.vec[n + ".dw_b"], map, 0, sy.f,
                        sy.ld, st));
    STTS_CK(row_ln(sy.f, sy.ld, 64, map, c.vec[n + ".ln_g"], c.vec[n + ".ln_b"], 0, 1e-6f, 0, sy.f, sy.ld, st,
                   Tw(c, sy).g, Tw(c, sy).l));
    STTS_RET(run_gemm(c, c.gemm[n + ".pw1"], GC_ISLAND, &sy, 0, nullptr, 0, mk_epi(EPI_GELU, sh, 0), map, st));
    STTS_CK(colstats(sh.f, sh.ld, 256, map, 1, 0.f, stat, gx, nullptr, st));
    c.launches++;
    STTS_CK(grn_apply(sh.f, sh.ld, 256, map, gx, c.vec[n + ".grn_gamma"], c.vec[n + ".grn_beta"], gscale, nullptr, 0, st,
                      Tw(c, sh).g, Tw(c, sh).l));
    c.launches++;
    {
      Epi e = mk_epi(EPI_ADD, sx, 0);
      e.res = sx.f; e.ldr = sx.ld; e.rcol0 = 0;
      STTS_RET(run_gemm(c, c.gemm[n + ".pw2"], GC_ISLAND, &sh, 0, nullptr, 0, e, map, st));
    }
  }
  STTS_CK(colstats(sx.f, sx.ld, 64, map, 0, 1e-5f, stat, mean, rstd, st));
  c.launches++;
  STTS_CK(cudaMemcpyAsync(d_style_out, mean, (size_t)B * 64 * sizeof(float), cudaMemcpyDeviceToDevice, st));
  return 0;
}

// Cross-attention block of PitchEnergyPredictor.compute_cross, pitch_energy_predictor.py:83-102:
//   base = prosody gathered to frames; q = AdaLN_q(base), k = AdaLN_k(prosody); 8-head attention with
//   partial RoPE and the inverted +-window band bias; conv_o; depthwise k5 + SiLU + 1x1; (base + att) / sqrt 2.
// d_prosody_tok token-major dense [sum T, 320] or NULL = the rows the prosody encoder left in PX (slot 1).
// Result: packed rows of the hop-300 layout in PX (the input buffer of the pitch / energy towers);
// optional d_x_out dense [320, F_b] per utterance.
int run_pe_cross(Context& c, const float* d_prosody_tok, const float* d_style, const int* d_idx, const int* d_cu_frames,
                 float* d_x_out, cudaStream_t st) {
  const std::string nm = "pe.cross";
  if (!c.finalized || c.vec.count(nm + ".fc_W") == 0) return fail(c, -3, "cross-attention weights not loaded");
  if (c.LT.total() == 0 || c.tact.count("PX#1") == 0) return fail(c, -4, "no token layout set");
  if (c.LF.total() == 0 || c.act.count("PX") == 0) return fail(c, -4, "no hop-300 layout set");
  if (c.LT.n_seg != c.LF.n_seg) return fail(c, -2, "token and frame layouts disagree on the number of utterances");
  const RowMap tmap = c.LT.map(), fmap = c.LF.map();
  const int C = 320, H = 8, D = 40, rd = 20, J = c.vec_len[nm + ".J"];
  Act &tpx = c.tact["PX#1"], &tpy = c.tact["PY#1"], &tkv = c.tact["PQKV#1"];
  Act &px = c.act["PX"], &pact = c.act["PACT"], &py1 = c.act["PY1"], &pa = c.act["PA"], &pb = c.act["PB"];
  float* gb = reinterpret_cast<float*>(c.traw["GBC"]);
  if (J > 2048) return fail(c, -6, "cross-attention: style FC too wide");
  STTS_CK(style_fc(c.vec[nm + ".fc_W"], c.vec[nm + ".fc_b"], d_style, J, 64, c.LT.n_seg, gb, st));
  if (d_prosody_tok) STTS_CK(dense_to_rows(d_prosody_tok, C, tmap, c.LT.d_cu, tpx.f, tpx.ld, 0, st));
  const bool t16 = c.precision == 1;
  // base: gather tokens -> frames.  expand_tokens reads a token-major dense matrix; the packed rows of
  // an utterance are contiguous, so the packed buffer is addressed with seg_start as the token offset
  STTS_CK(expand_tokens(tpx.f, tpx.ld, d_idx, d_cu_frames, c.LT.d_seg_start, 1, fmap, pa.f, pa.ld, 0, nullptr, 0, 0, st));
  STTS_CK(row_ln(pa.f, pa.ld, C, fmap, gb + c.fc_off[nm + ".qn"], nullptr, J, 1e-5f, 0, pb.f, pb.ld, st,
                 Tw(c, pb).g, Tw(c, pb).l));
  STTS_CK(row_ln(tpx.f, tpx.ld, C, tmap, gb + c.fc_off[nm + ".kn"], nullptr, J, 1e-5f, 0, tpy.f, tpy.ld, st,
                 Tw(c, tpy).g, Tw(c, tpy).l));
  Epi eq = mk_epi(EPI_BIAS, pact, 0);
  STTS_RET(run_gemm(c, c.gemm[nm + ".q"], GC_ISLAND, &pb, 0, nullptr, 0, eq, fmap, st));
  STTS_RET(run_gemm(c, c.gemm[nm + ".kv"], GC_ISLAND, &tpy, 0, nullptr, 0, mk_epi(EPI_BIAS, tkv, 0), tmap, st));
  int max_f = 0;
  for (int v : c.LF.seg_len) max_f = std::max(max_f, v);
  if (max_f > kRopePositions || c.max_t > kRopePositions) return fail(c, -6, "utterance longer than the RoPE table");
  STTS_CK(attention(pact.f, pact.ld, 0, tkv.f, tkv.ld, 0, tkv.f, tkv.ld, C, H, D, 1.0f / std::sqrt((float)D), rd, rope_tab(c, rd, st), fmap, tmap,
                    max_f, c.max_t, d_idx, d_cu_frames, 5, py1.f, py1.ld, st, Tw(c, py1).g, Tw(c, py1).l, py1.ldg));
  STTS_RET(run_gemm(c, c.gemm[nm + ".o"], GC_ISLAND, &py1, 0, nullptr, 0, mk_epi(EPI_BIAS, pb, 0), fmap, st));
  STTS_CK(dwconv_rows(pb.f, pb.ld, C, c.vec_len[nm + ".dw_k"], c.vec[nm + ".dw_w"], c.vec[nm + ".dw_b"], fmap, 1, pact.f,
                      pact.ld, st, Tw(c, pact).g, Tw(c, pact).l));
  {
    Epi e = mk_epi(EPI_RES_SCALE, px, 0);
    e.res = pa.f; e.ldr = pa.ld; e.rcol0 = 0;
    e.scale = 0.70710678118654752440f;
    e.outh = (c.precision == 1) ? px.h : nullptr; e.ldh = px.ld; e.colh0 = 0;
    STTS_RET(run_gemm(c, c.gemm[nm + ".post"], GC_ISLAND, &pact, 0, nullptr, 0, e, fmap, st));
  }
  if (d_x_out) STTS_CK(transpose_out(px.f, px.ld, 0, C, fmap, d_x_out, c.LF.d_off[3], c.LF.d_cs, 0, st));
  return 0;
}

}  // namespace stts
