/* libinksam -- the first consumers of the SAM masks (C ABI), SURVEY.md 8(f) rows f3 and f4.
 *
 * Same conventions as inksam.h: device pointers owned by the caller, `stream` is a cudaStream_t passed as void*, calls
 * are asynchronous on that stream, 0 = success, otherwise inksam_last_error().  The library keeps one grow-only
 * device workspace for these calls (not thread-safe).  `launches` (may be NULL) is incremented by the number of
 * kernels launched.
 */
#ifndef INKSAM_REFINE_H_
#define INKSAM_REFINE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* calculate_kernel_size(image_shape, factor=0.025): InkLayer/refinement/mask_cleaner.py:6-9. */
int inksam_morph_kernel_size(int H, int W);

/* clean_up_mask for K masks of one size: InkLayer/refinement/mask_cleaner.py:11-36.
 * masks_in:  uint8 [K][H][W]; a pixel is foreground where value > thresh (127 for the reference's 0/255 PNGs,
 *            0 for the 0/1 masks SamPredictor.predict_torch returns).
 * masks_out: uint8 [K][H][W], 0 / 255 like the reference's final_mask (may alias masks_in). */
int inksam_clean_masks(const uint8_t* masks_in, int K, int H, int W, int thresh, uint8_t* masks_out, long long* launches,
                       void* stream);

/* The mask arithmetic of content_iou: InkLayer/refinement/nms_sketch.py:62-78 (refine_mask_to_sketch_regions) and
 * :231-237 (intersection / union sums), for every pair of K masks at once.
 * masks: uint8 [K][H][W], foreground where value > 0; sketch_gray: uint8 [H][W] grey sketch already at mask size
 * (NULL = no restriction); a mask pixel counts only where sketch_gray < gray_lt (250 in the reference).
 * inter_out: int32 [K][K], inter[i][j] = |refined_i AND refined_j|; the diagonal holds the refined areas, so
 * union(i, j) = inter[i][i] + inter[j][j] - inter[i][j]. */
int inksam_content_iou(const uint8_t* masks, const uint8_t* sketch_gray, int K, int H, int W, int gray_lt, int32_t* inter_out,
                       long long* launches, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* INKSAM_REFINE_H_ */
