// Host engine behind include/inksam.h: weight store + repack, workspace, and the kernel sequences of
// inksam_encode (image encoder) and inksam_decode_boxes (prompt encoder + mask decoder + post-process).
#include "common.h"
#include "decoder_ops.h"
#include "inksam.h"

#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

using namespace inksam;

namespace {

struct Tensor {
  void* p = nullptr;
  long long n = 0;
  bool bf16 = false;
  std::vector<long long> shape;
};

struct BlockW {
  const float *n1_w, *n1_b, *n2_w, *n2_b, *qkv_b, *proj_b, *lin1_b, *lin2_b;
  const void *proj_w, *lin2_w;
  // norm1 -> qkv and norm2 -> lin1 with the LayerNorm folded in (gemm_args.h): bf16 gamma o W, its row sums, b + W beta
  void *qkv_w, *lin1_w;
  float *qkv_cs, *qkv_bf, *lin1_cs, *lin1_bf;
  void* rel;  // bf16 rel-pos table [256][hd] (global) or [64][hd] (windowed)
  bool global;
};

}  // namespace

struct inksam_handle {
  inksam_config cfg{};
  int device = 0, num_sms = 148, head_dim = 0;
  bool finalized = false;
  long long launches = 0;
  std::unordered_map<std::string, Tensor> w;
  std::vector<void*> owned;  // derived device buffers
  std::vector<BlockW> blocks;
  // decoder derived
  DecWeights dec{};
  float *pe = nullptr;                       // dense PE, token-major [4096][256]
  float *wkvq[2] = {nullptr, nullptr}, *bkvq[2] = {nullptr, nullptr}, *ra[2] = {nullptr, nullptr};  // [384][256], [384], [4096][384]
  float *wkvf = nullptr, *bkvf = nullptr, *raf = nullptr;                                            // [256][256], [256], [4096][256]
  // fp16 copies of the weights whose A operand is an fp16 intermediate (GemmArgs::ab_f16): out_proj of the two image->token
  // attentions [256][128], the final k|v projection [256][256], the two transposed convolutions [256][256] / [128][64]
  void *wo16[2] = {nullptr, nullptr}, *wkvq16 = nullptr, *wkvf16 = nullptr, *up0_16 = nullptr, *up3_16 = nullptr;
  void *ra16 = nullptr, *raf16 = nullptr;  // fp16 copies of the row-add tables ra[1] / raf (their sums are stored as fp16 anyway)
  // workspaces: sized once by inksam_reserve, otherwise grow-only.  A buffer that has to grow is RETIRED, not freed (work
  // already enqueued on any stream -- or baked into a captured CUDA graph -- may still use it); retired buffers are released
  // by inksam_reserve / inksam_destroy after a device synchronisation.  ws_generation counts reallocations so that the
  // caller can drop graphs that hold pointers into a retired buffer.
  void* enc_ws = nullptr; size_t enc_ws_bytes = 0;
  void* dec_ws = nullptr; size_t dec_ws_bytes = 0;
  std::vector<void*> retired;
  long long ws_generation = 0;
  // optional per-category CUDA-event timing of the launches (bench.py's roofline numbers)
  bool prof_on = false;
  int prof_cat = 0;  // category the next launches are charged to (offset by 3 inside decode)
  struct ProfRec { int cat; double work; cudaEvent_t e0, e1; };
  std::vector<ProfRec> prof;
  std::vector<cudaEvent_t> ev_pool;
};

namespace {

bool ends_with(const std::string& s, const char* suf) {
  const size_t n = strlen(suf);
  return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}
bool starts_with(const std::string& s, const char* pre) { return s.rfind(pre, 0) == 0; }

int dev_alloc(inksam_handle* h, void** p, size_t bytes) {
  CUDA_TRY(cudaMalloc(p, bytes));
  h->owned.push_back(*p);
  return 0;
}

// No device synchronisation and no cudaFree on this path: see the comment on inksam_handle::retired.
int grow(inksam_handle* h, void** p, size_t* have, size_t need) {
  if (need <= *have) return 0;
  size_t want = need;
  if (*p) {
    h->retired.push_back(*p);
    *p = nullptr;
    if (want < *have + *have / 2) want = *have + *have / 2;  // geometric: retired memory stays below 2x the live buffer
    *have = 0;
  }
  if (cudaMalloc(p, want) != cudaSuccess) {
    (void)cudaGetLastError();
    CUDA_TRY(cudaMalloc(p, need));
    want = need;
  }
  *have = want;
  ++h->ws_generation;
  return 0;
}

int release_retired(inksam_handle* h) {
  if (h->retired.empty()) return 0;
  CUDA_TRY(cudaDeviceSynchronize());
  for (void* q : h->retired) cudaFree(q);
  h->retired.clear();
  return 0;
}

const Tensor* find(const inksam_handle* h, const std::string& key) {
  auto it = h->w.find(key);
  return it == h->w.end() ? nullptr : &it->second;
}

int need_f32(const inksam_handle* h, const std::string& key, const float** out) {
  const Tensor* t = find(h, key);
  if (!t) return set_error("finalize: tensor '%s' was never loaded", key.c_str());
  *out = static_cast<const float*>(t->p);
  return 0;
}
int need_any(const inksam_handle* h, const std::string& key, const void** out) {
  const Tensor* t = find(h, key);
  if (!t) return set_error("finalize: tensor '%s' was never loaded", key.c_str());
  *out = t->p;
  return 0;
}

// tcgen05 GEMM wrapper: out = epilogue(A * W^T)
struct G {
  int tf32 = 0, bn = 256, epi = 0;
  const void* A = nullptr; int lda = 0;
  const void* W = nullptr; int ldw = 0;
  GemmArgs g{};
};
// RAII CUDA-event bracket around one launch when profiling is enabled; `work` = algorithmic FLOPs or bytes.
struct Prof {
  inksam_handle* h; cudaStream_t st; size_t idx; bool on;
  Prof(inksam_handle* h_, int cat, double work, cudaStream_t st_) : h(h_), st(st_), idx(0), on(h_->prof_on) {
    if (!on) return;
    cudaEvent_t e[2];
    for (auto& x : e) {
      if (h->ev_pool.empty()) cudaEventCreate(&x); else { x = h->ev_pool.back(); h->ev_pool.pop_back(); }
    }
    idx = h->prof.size();
    h->prof.push_back({cat, work, e[0], e[1]});
    cudaEventRecord(e[0], st);
  }
  ~Prof() { if (on) cudaEventRecord(h->prof[idx].e1, st); }
};
enum { CAT_ENC_GEMM = 0, CAT_ENC_ATTN = 1, CAT_ENC_OTHER = 2, CAT_DEC_GEMM = 3, CAT_DEC_OTHER = 4, CAT_POST = 5, CAT_N = 6 };

int run_gemm(inksam_handle* h, const G& c, cudaStream_t st) {
  ++h->launches;
  Prof p(h, h->prof_cat == 0 ? CAT_ENC_GEMM : CAT_DEC_GEMM, 2.0 * c.g.M * c.g.N * c.g.K, st);
  // CTA pairs (cta_group::2, UMMA 256 x BN): each CTA stages only half of the W tile, measured +10 % on the encoder shapes
  return launch_gemm(2, c.tf32, c.bn, c.epi, c.A, c.lda, c.W, c.ldw, c.g, h->num_sms, st);
}

}  // namespace

extern "C" {

int inksam_create(const inksam_config* cfg, int device, inksam_handle** out) {
  if (!cfg || !out) return set_error("inksam_create: null argument");
  if (cfg->img_size != 1024 || cfg->patch_size != 16 || cfg->window_size != 14 || cfg->out_chans != 256)
    return set_error("inksam_create: only img_size 1024 / patch 16 / window 14 / out_chans 256 are built (got %d/%d/%d/%d)",
                     cfg->img_size, cfg->patch_size, cfg->window_size, cfg->out_chans);
  if (cfg->num_heads < 1 || cfg->embed_dim % cfg->num_heads != 0)
    return set_error("inksam_create: embed_dim %d not divisible by num_heads %d", cfg->embed_dim, cfg->num_heads);
  const int hd = cfg->embed_dim / cfg->num_heads;
  if (hd != 64 && hd != 80) return set_error("inksam_create: head_dim %d not supported (64 or 80)", hd);
  if (cfg->embed_dim % 32 != 0 || cfg->embed_dim > 1280) return set_error("inksam_create: embed_dim %d unsupported", cfg->embed_dim);
  if (cfg->num_global < 0 || cfg->num_global > 8) return set_error("inksam_create: num_global %d not in 0..8", cfg->num_global);
  int ndev = 0;
  CUDA_TRY(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return set_error("inksam_create: device %d not present (%d CUDA devices)", device, ndev);
  CUDA_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return set_error("inksam_create: device %d is sm_%d%d; this library contains sm_100a code only", device, prop.major, prop.minor);
  auto* h = new inksam_handle();
  h->cfg = *cfg;
  h->device = device;
  h->num_sms = prop.multiProcessorCount;
  h->head_dim = hd;
  *out = h;
  return 0;
}

void inksam_destroy(inksam_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  for (auto& kv : h->w) cudaFree(kv.second.p);
  for (void* p : h->owned) cudaFree(p);
  for (void* p : h->retired) cudaFree(p);
  if (h->enc_ws) cudaFree(h->enc_ws);
  if (h->dec_ws) cudaFree(h->dec_ws);
  delete h;
}

long long inksam_launch_count(const inksam_handle* h) { return h ? h->launches : 0; }

int inksam_load_tensor(inksam_handle* h, const char* key_c, const float* data, const int64_t* shape, int ndim, void* stream) {
  if (!h || !key_c || !data || !shape) return set_error("inksam_load_tensor: null argument");
  if (ndim < 1 || ndim > 4) return set_error("inksam_load_tensor: '%s' has %d dims (1..4 supported)", key_c, ndim);
  auto st = static_cast<cudaStream_t>(stream);
  CUDA_TRY(cudaSetDevice(h->device));
  const std::string key(key_c);
  if (!(starts_with(key, "image_encoder.") || starts_with(key, "prompt_encoder.") || starts_with(key, "mask_decoder.")))
    return set_error("inksam_load_tensor: unknown key '%s'", key_c);
  if (starts_with(key, "prompt_encoder.mask_downscaling.") || key == "prompt_encoder.not_a_point_embed.weight" ||
      key == "prompt_encoder.point_embeddings.0.weight" || key == "prompt_encoder.point_embeddings.1.weight")
    return 0;  // not on the box-prompt path
  long long ishape[4], n = 1;
  for (int d = 0; d < ndim; ++d) { ishape[d] = shape[d]; n *= shape[d]; }

  bool to_bf16 = false;
  const int* perm = nullptr;
  static const int perm_0231[4] = {0, 2, 3, 1}, perm_2310[4] = {2, 3, 1, 0};
  if (starts_with(key, "image_encoder.") && ends_with(key, ".weight") && ndim >= 2 &&
      // (qkv and lin1 of the blocks stay fp32: finalize folds the preceding LayerNorm's gamma in before rounding)
      (ends_with(key, "attn.proj.weight") || ends_with(key, "mlp.lin2.weight") || key == "image_encoder.patch_embed.proj.weight" ||
       key == "image_encoder.neck.0.weight" || key == "image_encoder.neck.2.weight"))
    to_bf16 = true;
  if (key == "image_encoder.neck.2.weight") perm = perm_0231;  // [o][c][ky][kx] -> [o][ky][kx][c]
  if (key == "mask_decoder.output_upscaling.0.weight" || key == "mask_decoder.output_upscaling.3.weight")
    perm = perm_2310;  // ConvTranspose2d [c][o][dy][dx] -> [dy][dx][o][c]: rows (dy,dx,o) of a [4*O][C] GEMM weight
  if (perm && ndim != 4) return set_error("inksam_load_tensor: '%s' must be 4-D", key_c);

  Tensor t;
  t.n = n; t.bf16 = to_bf16;
  t.shape.assign(ishape, ishape + ndim);
  auto it = h->w.find(key);
  if (it != h->w.end()) { cudaFree(it->second.p); h->w.erase(it); }
  CUDA_TRY(cudaMalloc(&t.p, (size_t)n * (to_bf16 ? 2 : 4)));
  ++h->launches;
  if (launch_repack(data, t.p, to_bf16 ? 1 : 0, ishape, ndim, perm, st)) { cudaFree(t.p); return -1; }
  h->w.emplace(key, std::move(t));
  h->finalized = false;
  return 0;
}

}  // extern "C"

// ================================================================================================ finalize
namespace {

int build_rel_table(inksam_handle* h, const std::string& pre, bool global, void** out, cudaStream_t st) {
  const int side = global ? 64 : 14, rows = global ? 256 : 64, hd = h->head_dim;
  const Tensor* rh = find(h, pre + "rel_pos_h");
  const Tensor* rw = find(h, pre + "rel_pos_w");
  if (!rh || !rw) return set_error("finalize: '%srel_pos_h/w' missing", pre.c_str());
  for (const Tensor* t : {rh, rw})
    if (t->shape.size() != 2 || t->shape[0] != 2 * side - 1 || t->shape[1] != hd)
      return set_error("finalize: '%srel_pos_*' has shape [%lld,%lld], expected [%d,%d] (the interpolating branch of "
                       "get_rel_pos is not built)", pre.c_str(), t->shape[0], t->shape.size() > 1 ? t->shape[1] : 0LL,
                       2 * side - 1, hd);
  INK_TRY(dev_alloc(h, out, (size_t)rows * hd * 2));
  CUDA_TRY(cudaMemsetAsync(*out, 0, (size_t)rows * hd * 2, st));
  const long long shp[2] = {2 * side - 1, hd};
  h->launches += 2;
  INK_TRY(launch_repack(static_cast<const float*>(rh->p), *out, 1, shp, 2, nullptr, st));
  INK_TRY(launch_repack(static_cast<const float*>(rw->p), static_cast<__nv_bfloat16*>(*out) + (size_t)(rows / 2) * hd, 1, shp, 2,
                        nullptr, st));
  return 0;
}

// rows [r0, r0+rows) of dst (row length cols floats) <- src
int copy_rows(float* dst, int r0, const float* src, int rows, int cols, cudaStream_t st) {
  CUDA_TRY(cudaMemcpyAsync(dst + (size_t)r0 * cols, src, (size_t)rows * cols * 4, cudaMemcpyDeviceToDevice, st));
  return 0;
}

int finalize_decoder(inksam_handle* h, cudaStream_t st) {
  DecWeights& d = h->dec;
  const std::string M = "mask_decoder.", T = "mask_decoder.transformer.", P = "prompt_encoder.";
  INK_TRY(need_f32(h, M + "iou_token.weight", &d.iou_token));
  INK_TRY(need_f32(h, M + "mask_tokens.weight", &d.mask_tokens));
  INK_TRY(need_f32(h, P + "pe_layer.positional_encoding_gaussian_matrix", &d.gauss));
  INK_TRY(need_f32(h, P + "point_embeddings.2.weight", &d.corner0));
  INK_TRY(need_f32(h, P + "point_embeddings.3.weight", &d.corner1));
  auto lin = [&](const std::string& k, const float** w, const float** b) -> int {
    INK_TRY(need_f32(h, k + ".weight", w));
    return need_f32(h, k + ".bias", b);
  };
  for (int l = 0; l < 2; ++l) {
    DecLayerW& L = d.layer[l];
    const std::string p = T + "layers." + std::to_string(l) + ".";
    INK_TRY(lin(p + "self_attn.q_proj", &L.sa_q_w, &L.sa_q_b));
    INK_TRY(lin(p + "self_attn.k_proj", &L.sa_k_w, &L.sa_k_b));
    INK_TRY(lin(p + "self_attn.v_proj", &L.sa_v_w, &L.sa_v_b));
    INK_TRY(lin(p + "self_attn.out_proj", &L.sa_o_w, &L.sa_o_b));
    INK_TRY(lin(p + "norm1", &L.n1_w, &L.n1_b));
    INK_TRY(lin(p + "cross_attn_token_to_image.q_proj", &L.t2i_q_w, &L.t2i_q_b));
    INK_TRY(lin(p + "cross_attn_token_to_image.out_proj", &L.t2i_o_w, &L.t2i_o_b));
    INK_TRY(lin(p + "norm2", &L.n2_w, &L.n2_b));
    INK_TRY(lin(p + "mlp.lin1", &L.mlp1_w, &L.mlp1_b));
    INK_TRY(lin(p + "mlp.lin2", &L.mlp2_w, &L.mlp2_b));
    INK_TRY(lin(p + "norm3", &L.n3_w, &L.n3_b));
    INK_TRY(lin(p + "cross_attn_image_to_token.k_proj", &L.i2t_k_w, &L.i2t_k_b));
    INK_TRY(lin(p + "cross_attn_image_to_token.v_proj", &L.i2t_v_w, &L.i2t_v_b));
  }
  INK_TRY(lin(T + "final_attn_token_to_image.q_proj", &d.fin_q_w, &d.fin_q_b));
  INK_TRY(lin(T + "final_attn_token_to_image.out_proj", &d.fin_o_w, &d.fin_o_b));
  INK_TRY(lin(T + "norm_final_attn", &d.nf_w, &d.nf_b));
  for (int i = 0; i < 3; ++i) {
    INK_TRY(lin(M + "output_hypernetworks_mlps.0.layers." + std::to_string(i), &d.hyp_w[i], &d.hyp_b[i]));
    INK_TRY(lin(M + "iou_prediction_head.layers." + std::to_string(i), &d.iou_w[i], &d.iou_b[i]));
  }

  // dense positional encoding of the 64x64 grid (depends on the weights only)
  INK_TRY(dev_alloc(h, reinterpret_cast<void**>(&h->pe), (size_t)4096 * 256 * 4));
  ++h->launches;
  INK_TRY(launch_dense_pe(d.gauss, h->pe, st));

  // Image-side projections merged per transformer layer: rows [0,128) k of token->image, [128,256) v of
  // token->image, [256,384) q of image->token (all read the same keys, transformer.py:164-180).  k and q take
  // keys + key_pe: the PE part is box- and image-independent, so pe * W^T is precomputed as a row-add table.
  float* wra = nullptr;  // [384][256]: Wk ; 0 ; Wq  (weights that multiply the PE)
  INK_TRY(dev_alloc(h, reinterpret_cast<void**>(&wra), (size_t)384 * 256 * 4));
  for (int l = 0; l < 3; ++l) {
    const bool fin = (l == 2);
    const int N = fin ? 256 : 384;
    const std::string t2i = fin ? T + "final_attn_token_to_image." : T + "layers." + std::to_string(l) + ".cross_attn_token_to_image.";
    const std::string i2t = T + "layers." + std::to_string(l) + ".cross_attn_image_to_token.";
    float *wm = nullptr, *bm = nullptr, *ra = nullptr;
    INK_TRY(dev_alloc(h, reinterpret_cast<void**>(&wm), (size_t)N * 256 * 4));
    INK_TRY(dev_alloc(h, reinterpret_cast<void**>(&bm), (size_t)N * 4));
    INK_TRY(dev_alloc(h, reinterpret_cast<void**>(&ra), (size_t)4096 * N * 4));
    const float *kw, *kb, *vw, *vb, *qw = nullptr, *qb = nullptr;
    INK_TRY(lin(t2i + "k_proj", &kw, &kb));
    INK_TRY(lin(t2i + "v_proj", &vw, &vb));
    if (!fin) INK_TRY(lin(i2t + "q_proj", &qw, &qb));
    CUDA_TRY(cudaMemsetAsync(wra, 0, (size_t)384 * 256 * 4, st));
    INK_TRY(copy_rows(wm, 0, kw, 128, 256, st));
    INK_TRY(copy_rows(wm, 128, vw, 128, 256, st));
    INK_TRY(copy_rows(bm, 0, kb, 1, 128, st));
    INK_TRY(copy_rows(bm + 128, 0, vb, 1, 128, st));
    INK_TRY(copy_rows(wra, 0, kw, 128, 256, st));
    if (!fin) {
      INK_TRY(copy_rows(wm, 256, qw, 128, 256, st));
      INK_TRY(copy_rows(bm + 256, 0, qb, 1, 128, st));
      INK_TRY(copy_rows(wra, 256, qw, 128, 256, st));
    }
    G c;
    c.tf32 = 1; c.bn = 128; c.epi = EPI_BIAS_F32;
    c.A = h->pe; c.lda = 256; c.W = wra; c.ldw = 256;
    c.g.M = 4096; c.g.N = N; c.g.K = 256; c.g.ldo = N; c.g.out = ra;
    c.g.bias = bm;  // the projection bias rides in the row-add table: one operand less in the k|v|q GEMM's epilogue
    INK_TRY(run_gemm(h, c, st));
    if (fin) { h->wkvf = wm; h->bkvf = bm; h->raf = ra; }
    else { h->wkvq[l] = wm; h->bkvq[l] = bm; h->ra[l] = ra; }
  }
  // fp16 copies (see inksam_handle)
  auto to_f16 = [&](const float* src, long long rows, long long cols, void** out) -> int {
    INK_TRY(dev_alloc(h, out, (size_t)rows * cols * 2));
    const long long shp[2] = {rows, cols};
    ++h->launches;
    return launch_repack(src, *out, 2, shp, 2, nullptr, st);
  };
  for (int l = 0; l < 2; ++l) {
    const float* ow;
    INK_TRY(need_f32(h, T + "layers." + std::to_string(l) + ".cross_attn_image_to_token.out_proj.weight", &ow));
    INK_TRY(to_f16(ow, 256, 128, &h->wo16[l]));
  }
  INK_TRY(to_f16(h->wkvq[1], 384, 256, &h->wkvq16));
  INK_TRY(to_f16(h->ra[1], 4096, 384, &h->ra16));
  INK_TRY(to_f16(h->raf, 4096, 256, &h->raf16));
  INK_TRY(to_f16(h->wkvf, 256, 256, &h->wkvf16));
  {
    const float *u0w, *u3w;
    INK_TRY(need_f32(h, "mask_decoder.output_upscaling.0.weight", &u0w));
    INK_TRY(need_f32(h, "mask_decoder.output_upscaling.3.weight", &u3w));
    INK_TRY(to_f16(u0w, 256, 256, &h->up0_16));
    INK_TRY(to_f16(u3w, 128, 64, &h->up3_16));
  }
  return 0;
}

}  // namespace

extern "C" int inksam_finalize_weights(inksam_handle* h, void* stream) {
  if (!h) return set_error("inksam_finalize_weights: null handle");
  auto st = static_cast<cudaStream_t>(stream);
  CUDA_TRY(cudaSetDevice(h->device));
  for (void* p : h->owned) cudaFree(p);
  h->owned.clear();
  h->blocks.clear();
  const int D = h->cfg.embed_dim;
  const Tensor* qkv0 = find(h, "image_encoder.blocks.0.attn.qkv.weight");
  if (!qkv0 || qkv0->shape.size() != 2 || qkv0->shape[0] != 3 * D || qkv0->shape[1] != D)
    return set_error("finalize: image_encoder.blocks.0.attn.qkv.weight missing or not [%d,%d]", 3 * D, D);
  for (int i = 0; i < h->cfg.depth; ++i) {
    const std::string p = "image_encoder.blocks." + std::to_string(i) + ".";
    BlockW b{};
    b.global = false;
    for (int g = 0; g < h->cfg.num_global; ++g) b.global |= (h->cfg.global_attn_indexes[g] == i);
    INK_TRY(need_f32(h, p + "norm1.weight", &b.n1_w));
    INK_TRY(need_f32(h, p + "norm1.bias", &b.n1_b));
    INK_TRY(need_f32(h, p + "norm2.weight", &b.n2_w));
    INK_TRY(need_f32(h, p + "norm2.bias", &b.n2_b));
    INK_TRY(need_f32(h, p + "attn.qkv.bias", &b.qkv_b));
    INK_TRY(need_any(h, p + "attn.proj.weight", &b.proj_w));
    INK_TRY(need_f32(h, p + "attn.proj.bias", &b.proj_b));
    INK_TRY(need_f32(h, p + "mlp.lin1.bias", &b.lin1_b));
    {
      const float *wq, *w1;
      INK_TRY(need_f32(h, p + "attn.qkv.weight", &wq));
      INK_TRY(need_f32(h, p + "mlp.lin1.weight", &w1));
      const Tensor* t1 = find(h, p + "mlp.lin1.weight");
      if (t1->shape.size() != 2 || t1->shape[0] != 4 * D || t1->shape[1] != D)
        return set_error("finalize: %smlp.lin1.weight is not [%d,%d] (mlp_ratio 4)", p.c_str(), 4 * D, D);
      INK_TRY(dev_alloc(h, &b.qkv_w, (size_t)3 * D * D * 2));
      INK_TRY(dev_alloc(h, &b.lin1_w, (size_t)4 * D * D * 2));
      void* v;
      INK_TRY(dev_alloc(h, &v, (size_t)14 * D * 4));
      b.qkv_cs = static_cast<float*>(v); b.qkv_bf = b.qkv_cs + 3 * D; b.lin1_cs = b.qkv_bf + 3 * D; b.lin1_bf = b.lin1_cs + 4 * D;
      h->launches += 2;
      INK_TRY(launch_fold_ln_weights(wq, b.n1_w, b.n1_b, b.qkv_b, b.qkv_w, b.qkv_cs, b.qkv_bf, 3 * D, D, st));
      INK_TRY(launch_fold_ln_weights(w1, b.n2_w, b.n2_b, b.lin1_b, b.lin1_w, b.lin1_cs, b.lin1_bf, 4 * D, D, st));
    }
    INK_TRY(need_any(h, p + "mlp.lin2.weight", &b.lin2_w));
    INK_TRY(need_f32(h, p + "mlp.lin2.bias", &b.lin2_b));
    INK_TRY(build_rel_table(h, p + "attn.", b.global, &b.rel, st));
    h->blocks.push_back(b);
  }
  for (const char* k : {"image_encoder.pos_embed", "image_encoder.patch_embed.proj.weight", "image_encoder.patch_embed.proj.bias",
                        "image_encoder.neck.0.weight", "image_encoder.neck.1.weight", "image_encoder.neck.1.bias",
                        "image_encoder.neck.2.weight", "image_encoder.neck.3.weight", "image_encoder.neck.3.bias",
                        "prompt_encoder.no_mask_embed.weight", "mask_decoder.output_upscaling.0.weight",
                        "mask_decoder.output_upscaling.0.bias", "mask_decoder.output_upscaling.1.weight",
                        "mask_decoder.output_upscaling.1.bias", "mask_decoder.output_upscaling.3.weight",
                        "mask_decoder.output_upscaling.3.bias"})
    if (!find(h, k)) return set_error("finalize: tensor '%s' was never loaded", k);
  for (int l = 0; l < 2; ++l)
    for (const char* s : {"cross_attn_image_to_token.out_proj.weight", "cross_attn_image_to_token.out_proj.bias", "norm4.weight",
                          "norm4.bias"})
      if (!find(h, "mask_decoder.transformer.layers." + std::to_string(l) + "." + s))
        return set_error("finalize: tensor 'mask_decoder.transformer.layers.%d.%s' was never loaded", l, s);
  INK_TRY(finalize_decoder(h, st));
  h->finalized = true;
  return 0;
}

// ================================================================================================ encode
namespace {

size_t align_up(size_t x) { return (x + 1023) & ~size_t(1023); }

// Carve-up of the encoder workspace for a chunk of B images (offsets in bytes).
struct EncLayout { size_t x, xn, qkv, attn, hp, pad, stats, total; };
EncLayout enc_layout(const inksam_handle* h, int B) {
  const int D = h->cfg.embed_dim;
  const size_t M = (size_t)B * 4096, WROWS = (size_t)B * 25 * 196;  // window-ordered rows (>= M)
  const size_t sz_x = align_up(M * D * 4), sz_xn = align_up(M * D * 2);
  const size_t sz_qkv = align_up(WROWS * 3 * D * 2), sz_attn = sz_xn, sz_h = align_up(M * 4 * D * 2);
  const size_t sz_pad = align_up((size_t)B * 66 * 66 * 256 * 2);
  const size_t sz_patch = align_up(M * 768 * 2);
  const size_t sz_hp = sz_h > sz_patch ? sz_h : sz_patch;  // patches alias the MLP hidden buffer
  const size_t sz_stats = align_up(M * gemm_stats_partials(D) * 2 * 4);  // partial row statistics of the proj / lin2 epilogues
  EncLayout L;
  L.x = 0; L.xn = L.x + sz_x; L.qkv = L.xn + sz_xn; L.attn = L.qkv + sz_qkv; L.hp = L.attn + sz_attn; L.pad = L.hp + sz_hp;
  L.stats = L.pad + sz_pad; L.total = L.stats + sz_stats;
  return L;
}

int encode_chunk(inksam_handle* h, const void* images, int dtype, int B, const int* hw, long long img_stride,
                 long long row_pitch, long long plane_stride, float* emb_out, cudaStream_t st) {
  const int D = h->cfg.embed_dim, heads = h->cfg.num_heads, hd = h->head_dim;
  const int M = B * 4096;
  const int np = gemm_stats_partials(D);  // partial row statistics written by the proj / lin2 epilogues
  const EncLayout L = enc_layout(h, B);
  INK_TRY(grow(h, &h->enc_ws, &h->enc_ws_bytes, L.total));
  char* base = static_cast<char*>(h->enc_ws);
  float* x = reinterpret_cast<float*>(base + L.x);
  void* xn = base + L.xn;
  void* qkv = base + L.qkv;
  void* attn = base + L.attn;
  void* hbuf = base + L.hp;
  void* pad = base + L.pad;
  float* stats = reinterpret_cast<float*>(base + L.stats);

  const float *pos, *patch_b, *n1w, *n1b, *n3w, *n3b;
  const void *patch_w, *neck0, *neck2;
  INK_TRY(need_f32(h, "image_encoder.pos_embed", &pos));
  INK_TRY(need_any(h, "image_encoder.patch_embed.proj.weight", &patch_w));
  INK_TRY(need_f32(h, "image_encoder.patch_embed.proj.bias", &patch_b));
  INK_TRY(need_any(h, "image_encoder.neck.0.weight", &neck0));
  INK_TRY(need_f32(h, "image_encoder.neck.1.weight", &n1w));
  INK_TRY(need_f32(h, "image_encoder.neck.1.bias", &n1b));
  INK_TRY(need_any(h, "image_encoder.neck.2.weight", &neck2));
  INK_TRY(need_f32(h, "image_encoder.neck.3.weight", &n3w));
  INK_TRY(need_f32(h, "image_encoder.neck.3.bias", &n3b));

  // K1: normalise + patch gather, then the patch-embed GEMM with bias + pos_embed -> fp32 residual stream
  ++h->launches;
  INK_TRY(launch_patchify(images, dtype, B, hw, img_stride, row_pitch, plane_stride, h->cfg.pixel_mean, h->cfg.pixel_std, hbuf, st));
  {
    G c; c.epi = EPI_BIAS_ROWADD_F32; c.A = hbuf; c.lda = 768; c.W = patch_w; c.ldw = 768;
    c.g.M = M; c.g.N = D; c.g.K = 768; c.g.ldo = D; c.g.out = x; c.g.bias = patch_b; c.g.rowadd = pos; c.g.rowadd_mod = 4096;
    INK_TRY(run_gemm(h, c, st));
  }
  // Blocks (image_encoder.py:180-197).  The two LayerNorms are folded into the GEMMs around them (gemm_args.h): whoever
  // writes the residual stream x (fp32) also writes xn = bf16(x) and per-row partial sums; qkv and lin1 run on xn with
  // gamma folded into their weights and apply mean / rstd in the epilogue.  Only the first block needs a separate pass.
  ++h->launches;
  INK_TRY(launch_layernorm_bf16(x, nullptr, nullptr, xn, M, D, 0.f, 1, stats, st));
  int np_in = 1;
  for (const BlockW& b : h->blocks) {
    {
      G c; c.epi = EPI_FOLD_BF16; c.A = xn; c.lda = D; c.W = b.qkv_w; c.ldw = D;
      c.g.M = M; c.g.N = 3 * D; c.g.K = D; c.g.ldo = 3 * D; c.g.out = qkv; c.g.bias = b.qkv_bf; c.g.win_remap = b.global ? 0 : 1;
      // The padded window tokens' rows (q / k / v = the bare bias): at small batch this GEMM's epilogue warps write them while the
      // first tile's MMAs run (ViT-H batch 1: 6.75 -> 6.63 ms per encode); at batch 8 that measured 1 % SLOWER than the separate
      // 10 us kernel (44.45 -> 44.9 ms, same box: the step is power-capped and the extra stores delay the first epilogue).
      const bool fuse_pad = !b.global && B <= 2;
      if (fuse_pad) { c.g.pad_bias = b.qkv_b; c.g.pad_images = B; }
      c.g.stats_in = stats; c.g.colsum = b.qkv_cs; c.g.stats_np = np_in; c.g.fold_eps = 1e-6f;
      INK_TRY(run_gemm(h, c, st));
      if (!b.global && !fuse_pad) {
        ++h->launches;
        INK_TRY(launch_fill_pad_rows(qkv, b.qkv_b, B, 3 * D, st));
      }
    }
    ++h->launches;
    {
      const double keys = b.global ? 4096.0 : 196.0, side = b.global ? 64.0 : 14.0;
      Prof p(h, CAT_ENC_ATTN, (double)B * heads * 4096.0 * hd * (4.0 * keys + 4.0 * side), st);  // QK^T + PV + rel-pos terms
      INK_TRY(launch_attention(b.global ? 0 : 1, qkv, b.rel, attn, B, heads, hd, st));
    }
    {
      G c; c.epi = EPI_BIAS_RESID_STATS_F32; c.A = attn; c.lda = D; c.W = b.proj_w; c.ldw = D;
      c.g.M = M; c.g.N = D; c.g.K = D; c.g.ldo = D; c.g.out = x; c.g.bias = b.proj_b; c.g.resid = x;
      c.g.stats_out = stats; c.g.xb_out = xn;
      INK_TRY(run_gemm(h, c, st));
    }
    np_in = np;
    {
      G c; c.epi = EPI_FOLD_GELU_BF16; c.A = xn; c.lda = D; c.W = b.lin1_w; c.ldw = D;
      c.g.M = M; c.g.N = 4 * D; c.g.K = D; c.g.ldo = 4 * D; c.g.out = hbuf; c.g.bias = b.lin1_bf;
      c.g.stats_in = stats; c.g.colsum = b.lin1_cs; c.g.stats_np = np_in; c.g.fold_eps = 1e-6f;
      INK_TRY(run_gemm(h, c, st));
    }
    {
      G c; c.epi = EPI_BIAS_RESID_STATS_F32; c.A = hbuf; c.lda = 4 * D; c.W = b.lin2_w; c.ldw = 4 * D;
      c.g.M = M; c.g.N = D; c.g.K = 4 * D; c.g.ldo = D; c.g.out = x; c.g.bias = b.lin2_b; c.g.resid = x;
      c.g.stats_out = stats; c.g.xb_out = xn;
      INK_TRY(run_gemm(h, c, st));
    }
  }
  // neck: 1x1 conv + LayerNorm2d -> zero-bordered bf16 map; 3x3 conv (implicit GEMM over 9 taps) + LayerNorm2d -> NCHW fp32
  // (the neck's A operand bf16(x) was written by the last block's lin2 epilogue)
  if (h->blocks.empty()) return set_error("encode: no encoder blocks");
  CUDA_TRY(cudaMemsetAsync(pad, 0, (size_t)B * 66 * 66 * 256 * 2, st));
  {
    G c; c.epi = EPI_LN_BF16_PAD; c.A = xn; c.lda = D; c.W = neck0; c.ldw = D;
    c.g.M = M; c.g.N = 256; c.g.K = D; c.g.ldo = 256; c.g.out = pad; c.g.ln_w = n1w; c.g.ln_b = n1b; c.g.ln_eps = 1e-6f;
    INK_TRY(run_gemm(h, c, st));
  }
  {
    G c; c.epi = EPI_LN_F32_NCHW; c.A = pad; c.lda = 256; c.W = neck2; c.ldw = 9 * 256;
    c.g.M = M; c.g.N = 256; c.g.K = 9 * 256; c.g.ldo = 256; c.g.out = emb_out; c.g.ln_w = n3w; c.g.ln_b = n3b; c.g.ln_eps = 1e-6f;
    c.g.a_conv3x3 = 1; c.g.conv_c = 256;
    INK_TRY(run_gemm(h, c, st));
  }
  return 0;
}

}  // namespace

extern "C" int inksam_encode(inksam_handle* h, const void* images, int dtype, int B, const int* hw, int64_t img_stride,
                             int64_t row_pitch, int64_t plane_stride, float* embeddings_out, void* stream) {
  if (!h || !images || !hw || !embeddings_out) return set_error("inksam_encode: null argument");
  if (!h->finalized) return set_error("inksam_encode: weights not finalized (call inksam_finalize_weights)");
  if (B < 1) return set_error("inksam_encode: batch %d", B);
  if (dtype != INKSAM_IMAGE_U8_HWC && dtype != INKSAM_IMAGE_F32_CHW) return set_error("inksam_encode: dtype %d", dtype);
  CUDA_TRY(cudaSetDevice(h->device));
  auto st = static_cast<cudaStream_t>(stream);
  for (int b0 = 0; b0 < B; b0 += 32) {  // patchify takes at most 32 images per launch
    const int nb = B - b0 < 32 ? B - b0 : 32;
    INK_TRY(encode_chunk(h, static_cast<const char*>(images) + (long long)b0 * img_stride, dtype, nb, hw + 2 * b0, img_stride,
                         row_pitch, plane_stride, embeddings_out + (size_t)b0 * 256 * 4096, st));
  }
  return 0;
}

// ================================================================================================ decode
namespace {

constexpr int kBoxChunk = 128;  // boxes decoded per pass (bounds the workspace at ~2.2 GB)

// Decoder workspace for B images with K boxes in total: per-image layer-0 keys and their merged projection, the image
// index of every box, then the buffers of one chunk of at most kBoxChunk boxes (carved up by decode_chunk).
struct DecLayout { size_t kvq0, idx, shared, total; };
DecLayout dec_layout(int B, int K) {
  const int nc = K < kBoxChunk ? K : kBoxChunk;
  const size_t sz_keys0 = align_up((size_t)B * 4096 * 256 * 4), sz_kvq0 = align_up((size_t)B * 4096 * 384 * 4);
  const size_t sz_idx = align_up((size_t)K * 4);
  const size_t chunk_bytes = align_up((size_t)nc * 4096 * 256 * 4) + align_up((size_t)nc * 4096 * 384 * 4) +
                             align_up((size_t)nc * 4096 * 128 * 4) + 8 * align_up((size_t)nc * 7 * 256 * 4) +
                             align_up(decoder_t2i_scratch_floats(nc) * 4) + align_up((size_t)nc * 4);
  DecLayout L;
  L.kvq0 = sz_keys0; L.idx = sz_keys0 + sz_kvq0; L.shared = sz_keys0 + sz_kvq0 + sz_idx; L.total = L.shared + chunk_bytes;
  return L;
}

// Per-image geometry of a decode batch (host side).
// mask_off: element offset of the image's first map (logits); mask_byte_off: byte offset of its first mask (byte or packed)
struct ImgGeom { int in_h, in_w, H, W; size_t mask_off, mask_byte_off; };

// One pass over boxes [k0, k0+n) of the batch.  keys0 / kvq0: per-image layer-0 keys and their merged k|v|q projection;
// box_img: device array, image index of every box of the batch.
int decode_chunk(inksam_handle* h, const float* keys0, const float* kvq0, char* ws, const float* boxes, const int* box_img,
                 int k0, int n, const BoxSegments& seg, const std::vector<ImgGeom>& geom, uint8_t* masks, float* iou,
                 float* low, float* logits, float thr, bool packed, cudaStream_t st) {
  const std::string T = "mask_decoder.transformer.layers.";
  // per-chunk buffers
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = ws + off; off += align_up(bytes); return p; };
  float* keysb = reinterpret_cast<float*>(take((size_t)n * 4096 * 256 * 4));
  float* kvq = reinterpret_cast<float*>(take((size_t)n * 4096 * 384 * 4));   // also holds the 128x128x64 upscaled map
  float* a1 = reinterpret_cast<float*>(take((size_t)n * 4096 * 128 * 4));
  DecState s{};
  s.boxes = boxes + 4 * (size_t)k0;
  s.queries = reinterpret_cast<float*>(take((size_t)n * 7 * 256 * 4));
  s.query_pe = reinterpret_cast<float*>(take((size_t)n * 7 * 256 * 4));
  s.q_t2i = reinterpret_cast<float*>(take((size_t)n * 7 * 128 * 4));
  s.attn_t2i = reinterpret_cast<float*>(take((size_t)n * 7 * 128 * 4));
  s.kv_i2t = reinterpret_cast<float*>(take((size_t)n * 2 * 7 * 128 * 4));
  s.hyper = reinterpret_cast<float*>(take((size_t)n * 32 * 4));
  float* t2i_part = reinterpret_cast<float*>(take(decoder_t2i_scratch_floats(n) * 4));
  int* t2i_ticket = reinterpret_cast<int*>(take((size_t)n * 4));
  CUDA_TRY(cudaMemsetAsync(t2i_ticket, 0, (size_t)n * 4, st));
  s.iou = iou + k0;
  low += (size_t)k0 * 65536;
  const int* bimg = box_img + k0;
  const int Mb = n * 4096;

  auto tokens = [&](int phase, int layer) { ++h->launches; return launch_decoder_tokens(h->dec, s, n, phase, layer, st); };
  // per-box k|v|q (layer 1 and the final attention) are stored as fp16: written once by a tf32 GEMM, read once by fp32 FMAs
  auto t2i = [&](const void* kv, int half, long long stride, int ld, const int* bi) {
    ++h->launches;
    return launch_decoder_t2i(s.q_t2i, kv, half, stride, ld, bi, s.attn_t2i, t2i_part, t2i_ticket, n, st);
  };
  // ... and so are the GEMM-to-GEMM intermediates: the image->token attention output a1 (A of out_proj), the final keys
  // (A of the final k|v projection and of the first transposed convolution) and that convolution's output (A of the second)
  auto i2t = [&](const void* q, int half, long long stride, const int* bi) {
    ++h->launches;
    return launch_decoder_i2t(q, half, stride, 384, 256, bi, s.kv_i2t, a1, 1, n, st);
  };
  // keys <- norm4(keys + out_proj(image->token attention))   (transformer.py:176-180)
  auto keys_update = [&](int layer, const void* resid, int resid_f16, int resid_mod, const int* row_img, void* out, int out_f16) -> int {
    const float *ob, *nw, *nb;
    INK_TRY(need_f32(h, T + std::to_string(layer) + ".cross_attn_image_to_token.out_proj.bias", &ob));
    INK_TRY(need_f32(h, T + std::to_string(layer) + ".norm4.weight", &nw));
    INK_TRY(need_f32(h, T + std::to_string(layer) + ".norm4.bias", &nb));
    G c; c.tf32 = 0; c.epi = EPI_RESID_LN_F32; c.A = a1; c.lda = 128; c.W = h->wo16[layer]; c.ldw = 128;
    c.g.ab_f16 = 1; c.g.out_f16 = out_f16;
    c.g.M = Mb; c.g.N = 256; c.g.K = 128; c.g.ldo = 256; c.g.out = out; c.g.bias = ob; c.g.resid = static_cast<const float*>(resid);
    c.g.resid_f16 = resid_f16;
    c.g.resid_mod = resid_mod; c.g.row_img = row_img; c.g.ln_w = nw; c.g.ln_b = nb; c.g.ln_eps = 1e-5f;
    return run_gemm(h, c, st);
  };

  // ---- layer 0: the image-side k / v / q are shared by all boxes of an image (kvq0)
  INK_TRY(tokens(0, 0));
  INK_TRY(t2i(kvq0, 0, 4096LL * 384, 384, bimg));
  INK_TRY(tokens(2, 0));
  INK_TRY(i2t(kvq0, 0, 4096LL * 384, bimg));
  INK_TRY(keys_update(0, keys0, 0, 4096, bimg, keysb, 1));  // layer-1 keys: fp16 [n][4096][256] (first half of keysb)
  // ---- layer 1: per-box keys
  {
    G c; c.tf32 = 0; c.bn = 128; c.epi = EPI_BIAS_ROWADD_F32; c.A = keysb; c.lda = 256; c.W = h->wkvq16; c.ldw = 256;
    c.g.M = Mb; c.g.N = 384; c.g.K = 256; c.g.ldo = 384; c.g.out = kvq; c.g.bias = nullptr;
    c.g.rowadd = static_cast<const float*>(h->ra16); c.g.rowadd_f16 = 1;
    c.g.rowadd_mod = 4096; c.g.out_f16 = 1; c.g.ab_f16 = 1;
    INK_TRY(run_gemm(h, c, st));
  }
  INK_TRY(tokens(1, 1));
  INK_TRY(t2i(kvq, 1, 4096LL * 384, 384, nullptr));
  INK_TRY(tokens(2, 1));
  INK_TRY(i2t(kvq, 1, 4096LL * 384, nullptr));
  // the final keys are fp16 only: [n][4096][256] halves in the upper half of the k|v|q buffer (layer 1's k|v|q are dead; the
  // final k|v and then the upscaled map go into its lower part)
  void* keys2 = reinterpret_cast<char*>(kvq) + (size_t)n * 4096 * 384 * 2;
  INK_TRY(keys_update(1, keysb, 1, 0, nullptr, keys2, 1));
  // ---- final token -> image attention
  {
    G c; c.tf32 = 0; c.epi = EPI_BIAS_ROWADD_F32; c.A = keys2; c.lda = 256; c.W = h->wkvf16; c.ldw = 256;
    c.g.M = Mb; c.g.N = 256; c.g.K = 256; c.g.ldo = 256; c.g.out = kvq; c.g.bias = nullptr;
    c.g.rowadd = static_cast<const float*>(h->raf16); c.g.rowadd_f16 = 1;
    c.g.rowadd_mod = 4096; c.g.out_f16 = 1; c.g.ab_f16 = 1;
    INK_TRY(run_gemm(h, c, st));
  }
  INK_TRY(tokens(3, 0));
  INK_TRY(t2i(kvq, 1, 4096LL * 256, 256, nullptr));
  INK_TRY(tokens(4, 0));
  // ---- upscaling: ConvT(256->64,k2,s2)+LayerNorm2d+GELU, ConvT(64->32,k2,s2)+GELU, dot with the hypernetwork vector
  const float *u0w, *u0b, *u1w, *u1b, *u3w, *u3b;
  INK_TRY(need_f32(h, "mask_decoder.output_upscaling.0.weight", &u0w));
  INK_TRY(need_f32(h, "mask_decoder.output_upscaling.0.bias", &u0b));
  INK_TRY(need_f32(h, "mask_decoder.output_upscaling.1.weight", &u1w));
  INK_TRY(need_f32(h, "mask_decoder.output_upscaling.1.bias", &u1b));
  INK_TRY(need_f32(h, "mask_decoder.output_upscaling.3.weight", &u3w));
  INK_TRY(need_f32(h, "mask_decoder.output_upscaling.3.bias", &u3b));
  void* up1 = kvq;  // [n][128][128][64] fp16 (the final k|v there are dead)
  {
    G c; c.tf32 = 0; c.epi = EPI_UPSCALE1; c.A = keys2; c.lda = 256; c.W = h->up0_16; c.ldw = 256;
    c.g.M = Mb; c.g.N = 256; c.g.K = 256; c.g.ldo = 256; c.g.out = up1; c.g.bias = u0b; c.g.ln_w = u1w; c.g.ln_b = u1b;
    c.g.ln_eps = 1e-6f; c.g.ab_f16 = 1; c.g.out_f16 = 1;
    INK_TRY(run_gemm(h, c, st));
  }
  {
    G c; c.tf32 = 0; c.bn = 128; c.epi = EPI_UPSCALE2_MASK; c.A = up1; c.lda = 64; c.W = h->up3_16; c.ldw = 64;
    c.g.M = n * 16384; c.g.N = 128; c.g.K = 64; c.g.ldo = 128; c.g.out = low; c.g.bias = u3b; c.g.hyper = s.hyper;
    c.g.ab_f16 = 1;
    INK_TRY(run_gemm(h, c, st));
  }
  // ---- fused upsample + crop + resize + threshold: one launch for every box of the chunk (a segment per image)
  {
    const float* p_low[kMaxDecodeImages]; unsigned char* p_mask[kMaxDecodeImages]; float* p_log[kMaxDecodeImages];
    int cnt[kMaxDecodeImages], in_hw[2 * kMaxDecodeImages], orig_hw[2 * kMaxDecodeImages], ns = 0;
    double bytes = 0;
    for (int i = 0; i < seg.n; ++i) {
      const int b0 = seg.start[i] > k0 ? seg.start[i] : k0;
      const int b1 = seg.start[i + 1] < k0 + n ? seg.start[i + 1] : k0 + n;
      if (b1 <= b0) continue;
      const ImgGeom& gm = geom[i];
      const size_t HW = (size_t)gm.H * gm.W;
      const size_t mbytes = packed ? (size_t)gm.H * ((gm.W + 7) / 8) : HW;  // bytes of one mask
      const size_t mo = gm.mask_off + (size_t)(b0 - seg.start[i]) * HW;
      p_low[ns] = low + (size_t)(b0 - k0) * 65536;
      p_mask[ns] = masks ? masks + gm.mask_byte_off + (size_t)(b0 - seg.start[i]) * mbytes : nullptr;
      p_log[ns] = logits ? logits + mo : nullptr;
      cnt[ns] = b1 - b0;
      in_hw[2 * ns] = gm.in_h; in_hw[2 * ns + 1] = gm.in_w; orig_hw[2 * ns] = gm.H; orig_hw[2 * ns + 1] = gm.W;
      bytes += (double)(b1 - b0) * (65536.0 * 4 + (masks ? (double)mbytes : 0.0) + (logits ? 4.0 * (double)HW : 0.0));
      ++ns;
    }
    if (ns > 0 && (masks || logits)) {
      ++h->launches;
      Prof p(h, CAT_POST, bytes, st);
      INK_TRY(launch_postprocess_batch(ns, p_low, masks ? p_mask : nullptr, logits ? p_log : nullptr, cnt, in_hw, orig_hw, thr, st,
                                       packed ? 1 : 0));
    }
  }
  return 0;
}

}  // namespace

namespace {
int decode_batch_impl(inksam_handle* h, const float* embeddings, int B, const float* boxes, const int* counts,
                      const int* in_hw, const int* orig_hw, uint8_t* masks_out, float* iou_out, float* lowres_out,
                      float* logits_out, float mask_threshold, bool packed, void* stream) {
  if (!h || !embeddings || !counts || !in_hw || !orig_hw || !iou_out || !lowres_out)
    return set_error("inksam_decode_batch: null argument");
  if (!h->finalized) return set_error("inksam_decode_batch: weights not finalized (call inksam_finalize_weights)");
  if (B < 1 || B > kMaxDecodeImages) return set_error("inksam_decode_batch: B=%d not in 1..%d", B, kMaxDecodeImages);
  BoxSegments seg{};
  seg.n = B;
  std::vector<ImgGeom> geom(B);
  size_t moff = 0, mboff = 0;
  for (int i = 0; i < B; ++i) {
    if (counts[i] < 0) return set_error("inksam_decode_batch: counts[%d]=%d", i, counts[i]);
    seg.start[i + 1] = seg.start[i] + counts[i];
    ImgGeom& gm = geom[i];
    gm.in_h = in_hw[2 * i]; gm.in_w = in_hw[2 * i + 1]; gm.H = orig_hw[2 * i]; gm.W = orig_hw[2 * i + 1];
    if (gm.in_h < 1 || gm.in_w < 1 || gm.in_h > 1024 || gm.in_w > 1024 || gm.H < 1 || gm.W < 1)
      return set_error("inksam_decode_batch: image %d: bad sizes input %dx%d original %dx%d", i, gm.in_h, gm.in_w, gm.H, gm.W);
    gm.mask_off = moff;
    gm.mask_byte_off = mboff;
    moff += (size_t)counts[i] * gm.H * gm.W;
    mboff += (size_t)counts[i] * gm.H * (packed ? (size_t)((gm.W + 7) / 8) : (size_t)gm.W);
  }
  const int K = seg.start[B];
  if (K == 0) return 0;
  if (!boxes) return set_error("inksam_decode_batch: null boxes");
  CUDA_TRY(cudaSetDevice(h->device));
  auto st = static_cast<cudaStream_t>(stream);
  struct CatGuard { inksam_handle* h; CatGuard(inksam_handle* x) : h(x) { h->prof_cat = 1; } ~CatGuard() { h->prof_cat = 0; } } guard(h);
  const DecLayout L = dec_layout(B, K);
  const size_t shared_bytes = L.shared;
  INK_TRY(grow(h, &h->dec_ws, &h->dec_ws_bytes, L.total));
  char* ws = static_cast<char*>(h->dec_ws);
  float* keys0 = reinterpret_cast<float*>(ws);
  float* kvq0 = reinterpret_cast<float*>(ws + L.kvq0);
  int* box_img = reinterpret_cast<int*>(ws + L.idx);
  const float* nomask = nullptr;
  INK_TRY(need_f32(h, "prompt_encoder.no_mask_embed.weight", &nomask));
  ++h->launches;
  INK_TRY(launch_box_image_index(seg, box_img, st));
  // keys of layer 0 = embedding + dense no-mask prompt, token-major; their k / v / q projections once per image
  ++h->launches;
  INK_TRY(launch_decoder_keys0(embeddings, nomask, keys0, B, st));
  {
    G c; c.tf32 = 1; c.bn = 128; c.epi = EPI_BIAS_ROWADD_F32; c.A = keys0; c.lda = 256; c.W = h->wkvq[0]; c.ldw = 256;
    c.g.M = B * 4096; c.g.N = 384; c.g.K = 256; c.g.ldo = 384; c.g.out = kvq0; c.g.bias = nullptr; c.g.rowadd = h->ra[0];
    c.g.rowadd_mod = 4096;
    INK_TRY(run_gemm(h, c, st));
  }
  for (int k0 = 0; k0 < K; k0 += kBoxChunk) {
    const int n = K - k0 < kBoxChunk ? K - k0 : kBoxChunk;
    INK_TRY(decode_chunk(h, keys0, kvq0, ws + shared_bytes, boxes, box_img, k0, n, seg, geom, masks_out, iou_out, lowres_out,
                         logits_out, mask_threshold, packed, st));
  }
  return 0;
}
}  // namespace

extern "C" int inksam_decode_batch(inksam_handle* h, const float* embeddings, int B, const float* boxes, const int* counts,
                                   const int* in_hw, const int* orig_hw, uint8_t* masks_out, float* iou_out,
                                   float* lowres_out, float* logits_out, float mask_threshold, void* stream) {
  return decode_batch_impl(h, embeddings, B, boxes, counts, in_hw, orig_hw, masks_out, iou_out, lowres_out, logits_out,
                           mask_threshold, false, stream);
}

extern "C" int inksam_decode_batch_packed(inksam_handle* h, const float* embeddings, int B, const float* boxes, const int* counts,
                                          const int* in_hw, const int* orig_hw, uint8_t* mask_bits_out, float* iou_out,
                                          float* lowres_out, float mask_threshold, void* stream) {
  return decode_batch_impl(h, embeddings, B, boxes, counts, in_hw, orig_hw, mask_bits_out, iou_out, lowres_out, nullptr,
                           mask_threshold, true, stream);
}

extern "C" int inksam_decode_boxes(inksam_handle* h, const float* embedding, const float* boxes, int K, int in_h, int in_w,
                                   int orig_h, int orig_w, uint8_t* masks_out, float* iou_out, float* lowres_out,
                                   float* logits_out, float mask_threshold, void* stream) {
  if (K < 0) return set_error("inksam_decode_boxes: K=%d", K);
  const int counts[1] = {K}, in_hw[2] = {in_h, in_w}, orig_hw[2] = {orig_h, orig_w};
  return inksam_decode_batch(h, embedding, 1, boxes, counts, in_hw, orig_hw, masks_out, iou_out, lowres_out, logits_out,
                             mask_threshold, stream);
}

// ================================================================================================ workspace
// Bytes of library-owned workspace that encodes of up to max_batch images per call (per chunk of 32) and decodes of up to
// max_batch images with up to max_boxes boxes in total need.
extern "C" size_t inksam_workspace_bytes(const inksam_handle* h, int max_batch, int max_boxes) {
  if (!h || max_batch < 1 || max_boxes < 0) return 0;
  const int eb = max_batch < 32 ? max_batch : 32, db = max_batch < kMaxDecodeImages ? max_batch : kMaxDecodeImages;
  return enc_layout(h, eb).total + (max_boxes > 0 ? dec_layout(db, max_boxes).total : 0);
}

// Sizes both workspaces once for that envelope, so that no later call inside it allocates (a captured CUDA graph keeps
// pointing at live memory), and releases buffers retired by earlier growth.  Synchronises the device.
extern "C" int inksam_reserve(inksam_handle* h, int max_batch, int max_boxes) {
  if (!h) return set_error("inksam_reserve: null handle");
  if (max_batch < 1 || max_boxes < 0) return set_error("inksam_reserve: max_batch %d / max_boxes %d", max_batch, max_boxes);
  CUDA_TRY(cudaSetDevice(h->device));
  const int eb = max_batch < 32 ? max_batch : 32, db = max_batch < kMaxDecodeImages ? max_batch : kMaxDecodeImages;
  INK_TRY(grow(h, &h->enc_ws, &h->enc_ws_bytes, enc_layout(h, eb).total));
  if (max_boxes > 0) INK_TRY(grow(h, &h->dec_ws, &h->dec_ws_bytes, dec_layout(db, max_boxes).total));
  return release_retired(h);
}

// Incremented whenever a workspace buffer is reallocated: graphs captured under an older generation must be dropped.
extern "C" long long inksam_workspace_generation(const inksam_handle* h) { return h ? h->ws_generation : 0; }

// ================================================================================================ profiling
extern "C" int inksam_profile_enable(inksam_handle* h, int on) {
  if (!h) return set_error("inksam_profile_enable: null handle");
  for (auto& r : h->prof) { h->ev_pool.push_back(r.e0); h->ev_pool.push_back(r.e1); }
  h->prof.clear();
  h->prof_on = on != 0;
  return 0;
}

// ms[c], work[c], count[c] for c in {0 encoder GEMM (FLOP), 1 encoder attention (FLOP), 2 unused, 3 decoder GEMM (FLOP),
// 4 unused, 5 post-process (bytes)}: summed CUDA-event durations of the launches since inksam_profile_enable(h, 1).
extern "C" int inksam_profile_read(inksam_handle* h, double* ms, double* work, long long* count) {
  if (!h || !ms || !work || !count) return set_error("inksam_profile_read: null argument");
  for (int c = 0; c < CAT_N; ++c) { ms[c] = 0; work[c] = 0; count[c] = 0; }
  for (auto& r : h->prof) {
    CUDA_TRY(cudaEventSynchronize(r.e1));
    float t = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&t, r.e0, r.e1));
    ms[r.cat] += t; work[r.cat] += r.work; count[r.cat] += 1;
  }
  return 0;
}
