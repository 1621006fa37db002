// Argument blocks of the mask-decoder token-side kernels (fp32 device pointers; keys are the reference state_dict names).
#pragma once
#include <cuda_runtime.h>

namespace inksam {

// mask_decoder.transformer.layers.{l}.*
struct DecLayerW {
  const float *sa_q_w, *sa_q_b, *sa_k_w, *sa_k_b, *sa_v_w, *sa_v_b, *sa_o_w, *sa_o_b;  // self_attn.{q,k,v,out}_proj
  const float *n1_w, *n1_b;
  const float *t2i_q_w, *t2i_q_b, *t2i_o_w, *t2i_o_b;  // cross_attn_token_to_image.{q,out}_proj (k, v run on the GEMM)
  const float *n2_w, *n2_b;
  const float *mlp1_w, *mlp1_b, *mlp2_w, *mlp2_b;
  const float *n3_w, *n3_b;
  const float *i2t_k_w, *i2t_k_b, *i2t_v_w, *i2t_v_b;  // cross_attn_image_to_token.{k,v}_proj (q, out run on the GEMM)
};

struct DecWeights {
  const float *iou_token, *mask_tokens;       // mask_decoder.{iou_token,mask_tokens}.weight
  const float *gauss;                         // prompt_encoder.pe_layer.positional_encoding_gaussian_matrix [2][128]
  const float *corner0, *corner1;             // prompt_encoder.point_embeddings.{2,3}.weight
  DecLayerW layer[2];
  const float *fin_q_w, *fin_q_b, *fin_o_w, *fin_o_b;  // final_attn_token_to_image.{q,out}_proj
  const float *nf_w, *nf_b;                   // norm_final_attn
  const float *hyp_w[3], *hyp_b[3];           // output_hypernetworks_mlps.0.layers.{0,1,2}
  const float *iou_w[3], *iou_b[3];           // iou_prediction_head.layers.{0,1,2}
};

struct DecState {
  const float* boxes;  // [K][4] xyxy in the resized-input frame
  float* queries;      // [K][7][256]
  float* query_pe;     // [K][7][256]
  float* q_t2i;        // [K][7][128]
  float* attn_t2i;     // [K][7][128]
  float* kv_i2t;       // [K][2][7][128]
  float* hyper;        // [K][32]
  float* iou;          // [K]
};

// Boxes of a batch of images, grouped by image: image i owns boxes [start[i], start[i+1]); passed to kernels by value.
constexpr int kMaxDecodeImages = 64;
struct BoxSegments {
  int n;
  int start[kMaxDecodeImages + 1];
};

int launch_decoder_tokens(const DecWeights& w, const DecState& s, int boxes, int phase, int layer, cudaStream_t st);
// kv_half / q_half: the k|v|q rows are fp16 (GemmArgs::out_f16) instead of fp32; box_stride and ld count elements either way
int launch_decoder_t2i(const float* q, const void* kv, int kv_half, long long box_stride, int ld, const int* box_img, float* out,
                       float* partial, int* ticket, int boxes, cudaStream_t st);
size_t decoder_t2i_scratch_floats(int boxes);  // floats of `partial`; `ticket` is `boxes` ints, zeroed once
int launch_decoder_i2t(const void* qsrc, int q_half, long long box_stride, int ld, int qcol, const int* box_img, const float* kv,
                       void* out, int out_half, int boxes, cudaStream_t st);
int launch_decoder_keys0(const float* emb, const float* nomask, float* keys, int B, cudaStream_t st);
int launch_box_image_index(const BoxSegments& seg, int* box_img, cudaStream_t st);
int launch_dense_pe(const float* gauss, float* pe, cudaStream_t st);

}  // namespace inksam
