/* libinksam -- the drop-in boundary (C ABI) of the B200-native SAM segmentation hot path.
 *
 * These entry points are what a binding of the reference's `segment_anything` surface calls in place of the
 * PyTorch modules (file:line relative to InkLayer/third_party/segment-anything/segment_anything/):
 *
 *   inksam_create / inksam_load_tensor / inksam_finalize_weights
 *        replace build_sam.py:55-107 (_build_sam + load_state_dict): same hyper-parameters, same state_dict keys.
 *   inksam_encode
 *        replaces Sam.preprocess + ImageEncoderViT.forward as called from SamPredictor.set_torch_image
 *        (predictor.py:88-89, modeling/sam.py:164-174, modeling/image_encoder.py:106-116).
 *   inksam_decode_boxes
 *        replaces prompt_encoder + mask_decoder + postprocess_masks + threshold as called from
 *        SamPredictor.predict_torch with boxes only and multimask_output=False
 *        (predictor.py:222-241, modeling/prompt_encoder.py:128-168, modeling/mask_decoder.py:71-149,
 *        modeling/sam.py:133-162).
 *
 * Conventions: every data pointer is a DEVICE pointer owned by the caller (the Python shim passes torch tensors'
 * data_ptr()); `stream` is a cudaStream_t passed as void*; calls are asynchronous on that stream; the library owns
 * only its packed weights and its workspace.  Return value 0 = success; otherwise inksam_last_error() describes the
 * failure.  A handle is bound to one device and is not thread-safe (like SamPredictor).  There is no CPU path.
 */
#ifndef INKSAM_H_
#define INKSAM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct inksam_handle inksam_handle;

/* Mirrors the arguments of _build_sam (build_sam.py:55-101). */
typedef struct inksam_config {
  int embed_dim;              /* 1280 / 1024 / 768 */
  int depth;                  /* 32 / 24 / 12 */
  int num_heads;              /* 16 / 16 / 12; head_dim must be 64 or 80 */
  int num_global;             /* number of entries used in global_attn_indexes */
  int global_attn_indexes[8];
  int window_size;            /* 14 */
  int img_size;               /* 1024 */
  int patch_size;             /* 16 */
  int out_chans;              /* 256 */
  float pixel_mean[3];
  float pixel_std[3];
} inksam_config;

enum { INKSAM_IMAGE_U8_HWC = 0, INKSAM_IMAGE_F32_CHW = 1 };

int inksam_create(const inksam_config* cfg, int device, inksam_handle** out);
void inksam_destroy(inksam_handle* h);

/* One tensor of the reference state_dict (fp32, contiguous, on the handle's device), keyed by its reference name,
 * e.g. "image_encoder.blocks.0.attn.qkv.weight".  The library converts / repacks into its own layouts (bf16 GEMM
 * operands, [o][ky][kx][c] conv weights, ...) and keeps no reference to `data`.  Unknown keys are an error; keys
 * that the box-prompt path does not use (mask_downscaling.*, point_embeddings.0/1, not_a_point_embed) are accepted
 * and ignored. */
int inksam_load_tensor(inksam_handle* h, const char* key, const float* data, const int64_t* shape, int ndim, void* stream);

/* Checks that every required tensor arrived and builds the derived tables (rel-pos tables, merged decoder
 * projections, dense positional encoding). */
int inksam_finalize_weights(inksam_handle* h, void* stream);

/* images: B images, each either uint8 HWC (dtype 0) or float32 CHW (dtype 1, values 0..255), already resized so
 * that the long side is <= img_size; hw[2*i], hw[2*i+1] = height, width of image i (HOST array).
 * img_stride / row_pitch / plane_stride are in bytes (plane_stride only for CHW).
 * embeddings_out: float32 [B][out_chans][64][64]. */
int inksam_encode(inksam_handle* h, const void* images, int dtype, int B, const int* hw, int64_t img_stride,
                  int64_t row_pitch, int64_t plane_stride, float* embeddings_out, void* stream);

/* embedding: float32 [out_chans][64][64] of ONE image; boxes: float32 [K][4] xyxy in the resized-input frame.
 * masks_out:  uint8 [K][1][orig_h][orig_w], 1 where logit > mask_threshold (may be NULL);
 * iou_out:    float32 [K][1];  lowres_out: float32 [K][1][256][256];
 * logits_out: float32 [K][1][orig_h][orig_w] (may be NULL; the return_logits=True path). */
int inksam_decode_boxes(inksam_handle* h, const float* embedding, const float* boxes, int K, int in_h, int in_w,
                        int orig_h, int orig_w, uint8_t* masks_out, float* iou_out, float* lowres_out,
                        float* logits_out, float mask_threshold, void* stream);

/* The same for the boxes of B images in ONE pass (the reference's own multi-image entry Sam.forward loops this per
 * image, modeling/sam.py:101-130; InkLayer's `main.py --dir` loop is the caller that batches, main.py:27-32).  Boxes are
 * grouped by image: image i owns counts[i] consecutive rows of `boxes`.  counts / in_hw / orig_hw are HOST arrays
 * ([B], [B][2], [B][2]); every other pointer is a device pointer.
 * embeddings: float32 [B][256][64][64];  boxes: float32 [sum counts][4], each in its own image's resized-input frame;
 * masks_out (may be NULL) / logits_out (may be NULL): image after image, counts[i] maps of orig_hw[i] each, densely
 * packed;  iou_out: float32 [sum counts];  lowres_out: float32 [sum counts][256][256].  1 <= B <= 64. */
int inksam_decode_batch(inksam_handle* h, const float* embeddings, int B, const float* boxes, const int* counts,
                        const int* in_hw, const int* orig_hw, uint8_t* masks_out, float* iou_out, float* lowres_out,
                        float* logits_out, float mask_threshold, void* stream);

/* inksam_decode_batch with BIT-PACKED masks: mask_bits_out holds, image after image, counts[i] maps of
 * orig_h x ceil(orig_w / 8) bytes (leftmost pixel in the MSB, rows padded to whole bytes with zero bits: numpy.packbits
 * / PIL mode "1" rows, what InkLayer/runner.py:51 turns the masks into).  8x less device->host traffic than byte masks. */
int inksam_decode_batch_packed(inksam_handle* h, const float* embeddings, int B, const float* boxes, const int* counts,
                               const int* in_hw, const int* orig_hw, uint8_t* mask_bits_out, float* iou_out,
                               float* lowres_out, float mask_threshold, void* stream);

/* Workspace.  The library owns its workspace.  inksam_workspace_bytes: what encodes of up to max_batch
 * images per call and decodes of up to max_batch images with up to max_boxes boxes in total need.  inksam_reserve sizes
 * the workspace once for that envelope (synchronises the device; releases buffers retired by earlier growth) so that no
 * later call inside it allocates -- CUDA graphs captured over inksam_encode stay valid.  A call outside the envelope
 * grows the workspace WITHOUT a device synchronisation and without freeing the old buffer (it is retired until the next
 * inksam_reserve / inksam_destroy) and increments inksam_workspace_generation: graphs captured under an older generation
 * replay into retired memory and must be dropped by the caller. */
size_t inksam_workspace_bytes(const inksam_handle* h, int max_batch, int max_boxes);
int inksam_reserve(inksam_handle* h, int max_batch, int max_boxes);
long long inksam_workspace_generation(const inksam_handle* h);

/* Number of kernels this handle has launched so far (bench.py's gpu_launches). */
long long inksam_launch_count(const inksam_handle* h);

/* Optional CUDA-event timing of the library's own launches, per category (bench.py's roofline figures):
 * category 0 encoder GEMMs, 1 encoder attention, 3 decoder GEMMs (work = algorithmic FLOP), 5 mask post-process
 * (work = algorithmic bytes); 2 and 4 are unused.  enable(1) resets; read() waits for the recorded events. */
int inksam_profile_enable(inksam_handle* h, int on);
int inksam_profile_read(inksam_handle* h, double* ms /*[6]*/, double* work /*[6]*/, long long* count /*[6]*/);

const char* inksam_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* INKSAM_H_ */
