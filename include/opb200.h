/* libopb200 -- C ABI of the B200-native OnePose++ / LoFTR matching hot path.
 *
 * The reference (mizeller/OnePose_ST) has no FFI on this path: its operator boundary is a set of
 * Python nn.Module classes.  This header is the C boundary *below* those classes: the
 * drop-in modules in onepose_st_b200/ hold the reference's signatures and call these entry points with
 * raw device pointers.  Each entry point names the reference code it replaces.
 *
 * Conventions: plain C types; every pointer is a DEVICE pointer unless named host_*; no allocation
 * inside (caller passes a workspace whose size the matching *_workspace_bytes() returns); work is
 * enqueued on `stream` (a cudaStream_t passed as void*); returns 0 on success, a negative code on
 * failure with a message in opb200_last_error(); never throws; sm_100a only (OPB200_ERR_DEVICE on any
 * other device -- there is no fallback path).
 */
#ifndef OPB200_H_
#define OPB200_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OPB200_OK 0
#define OPB200_ERR_ARG (-1)
#define OPB200_ERR_CUDA (-2)
#define OPB200_ERR_WORKSPACE (-3)
#define OPB200_ERR_DEVICE (-4)

/* precision of the tensor-core products: 3 = bf16x3 split (fp32-grade, parity default), 1 = single bf16 */
#define OPB200_PREC_BF16 1
#define OPB200_PREC_BF16X3 3

int opb200_version(void);
const char* opb200_last_error(void);
/* Number of CUDA kernels this library has launched in this process (bench.py's gpu_launches evidence). */
long long opb200_launch_count(void);
/* Adds n to that counter: a caller that replays a CUDA graph of n captured launches of this library accounts for them here. */
void opb200_launch_count_add(long long n);

/* fp32 matrix with element (r,c) at x[r*stride_r + c*stride_c] -> bf16 hi/lo planes [rows, cols]
 * (lo and f32_copy nullable).  Operand preparation for every tensor-core product below. */
int opb200_split_planes(const float* x, int64_t stride_r, int64_t stride_c, int rows, int cols, void* hi, void* lo,
                        float* f32_copy, void* stream);

/* Test hook for the tcgen05/TMA mainloop: c[M,N] = a[M,K] * b[N,K]^T from bf16 planes. */
int opb200_debug_gemm(const void* a_hi, const void* a_lo, const void* b_hi, const void* b_lo, int M, int N, int K,
                      int block_n, int precision, float* c, void* stream);

/* Debug: arm (buffer != NULL, zero-filled device memory) or disarm per-tile timestamp tracing of the tcgen05 GEMM
 * launches; launch i writes [148][32][4] int64 SM-clock stamps {first TMA, MMA start, accumulator full, epilogue
 * done} at slot i.  Returns the number of slots the buffer holds.  Used by tools/gemm_trace.py only. */
int opb200_debug_trace(void* buffer, long long bytes);

/* ---- LoFTR encoder stack (elu+1 linear attention) --------------------------------------------------
 * Replaces LoFTREncoderLayer.forward + LinearAttention.forward + LocalFeatureTransformer.forward:
 *   src/models/OnePosePlus/loftr_module/transformer.py:65-94,133-171
 *   src/models/OnePosePlus/loftr_module/linear_attention.py:29-61
 *   submodules/LoFTR/src/loftr/loftr_module/transformer.py:35-58,80-101
 */
/* Bytes of one packed layer: bf16 hi/lo planes of [Wq;Wk;Wv], Wmerge, Wmlp0, Wmlp2 + fp32 LN params. */
size_t opb200_packed_layer_bytes(int d_model);
/* Pack one encoder layer's weights (device fp32, nn.Linear [out,in] layout) into `packed`. */
int opb200_pack_layer_weights(const float* wq, const float* wk, const float* wv, const float* wm, const float* w1,
                              const float* w2, const float* ln1_w, const float* ln1_b, const float* ln2_w,
                              const float* ln2_b, int d_model, void* packed, void* stream);
size_t opb200_transformer_workspace_bytes(int d_model, int n_batch, int len0, int len1, int big_groups);
/* Runs n_layers encoder layers over two token streams.
 *   host_layer_weights[l] : device pointer to layer l's packed weights;  host_layer_kinds[l]: 0 self, 1 cross,
 *            2 = self layer applied to stream 1 only, 3 = to stream 0 only (the other stream's rows are untouched:
 *            how a caller reuses a stream whose first self layer it has already computed, e.g. the fixed 3D model
 *            of a OnePose++ sequence)
 *   schedule 0 = OnePosePlus (one layer object serves both streams, both read the pre-update tensors),
 *            1 = LoFTR (cross step: stream0 = layer(stream0, stream1) then stream1 = layer(stream1, new stream0))
 *   big_groups 1 = coarse level (d_model 256; one attention group per (stream, batch element)),
 *              0 = fine level   (d_model 128; n_batch = number of matches, len0/len1 <= 128 tokens)
 *   stream s element (n, l, c) is read from x{s}[n*sn + l*sl + c*sc]; add{s} (nullable) is a [len, C]
 *   table added to every batch element (the sine position encoding); mask{s} (nullable) is [n_batch, len]
 *   bytes, 1 = valid.  Results are written densely: out0 [n_batch, len0, C], out1 [n_batch, len1, C]. */
int opb200_transformer(const void* const* host_layer_weights, const int* host_layer_kinds, int n_layers, int schedule,
                       int d_model, int nhead, int n_batch, int len0, int len1, int big_groups, const float* x0,
                       int64_t x0_sn, int64_t x0_sl, int64_t x0_sc, const float* add0, const float* x1, int64_t x1_sn,
                       int64_t x1_sl, int64_t x1_sc, const float* add1, const uint8_t* mask0, const uint8_t* mask1,
                       float* out0, float* out1, void* workspace, size_t workspace_bytes, int precision, void* stream);

/* After opb200_transformer: where the final token features live inside `workspace` as bf16 hi/lo planes
 * (x = hi + lo).  planes[4] = {stream0 hi, stream0 lo, stream1 hi, stream1 lo}, each [n_batch * padded_len[s], d_model]
 * row-major with batch element n at row n * padded_len[s].  opb200_transformer may be called with out0 = out1 = NULL
 * when the only consumer is opb200_coarse_match_planes (no fp32 copy-out, no re-splitting). */
int opb200_transformer_planes(void* workspace, int d_model, int n_batch, int len0, int len1, int big_groups, void** planes,
                              int* padded_len);

/* ---- CoarseMatching: dual-softmax + threshold + border + mutual nearest neighbour ------------------
 * Replaces CoarseMatching.forward / get_coarse_match (eval mode):
 *   src/models/OnePosePlus/utils/coarse_matching.py:76-242
 *   submodules/LoFTR/src/loftr/utils/coarse_matching.py:87-261 (match_type dual_softmax)
 * feat_a [n_batch, L, C], feat_b [n_batch, S, C] fp32 contiguous; masks nullable bytes (1 = valid);
 * mask_fill 0: sim += -1e9 where masked (OnePosePlus), 1: sim = -1e9 (LoFTR);
 * scale = 1 / (C * (temperature + eps)); border_mode 0 = OnePosePlus (first border_rm rows/cols of the
 * h1 x w1 grid), 1 = LoFTR (four sides of both grids; valid_hw nullable [n_batch,4] = valid (h0,w0,h1,w1)
 * from the padding masks).  Outputs (capacity rows each): b_ids/i_ids/j_ids int64, mconf fp32, in
 * ascending (b, i) order; count_and_overflow is int[3] on the device: {number of matches, overflow flag,
 * rerun flag}.  conf_matrix (nullable) receives the full [n_batch, L, S] confidence matrix.
 * algo 0 = auto: when there are no masks, no conf_matrix, C == 256 and precision 3, ONE pass over the
 * similarity tiles computes the softmax statistics and captures the candidate matches.  That sweep runs a single
 * fp16 x fp16 product; every captured element is then re-evaluated from the split planes in fp32 and its term in
 * the row / column sums replaced by the exact one (confidences within the 1e-3 contract, measured 6e-4).
 * algo 2 = the same single pass on the three bf16 split products (fp32-grade everywhere, 3x the tensor work).
 * algo 3 = algo 0 with the register-buffered epilogue (128x256 MMAs, every epilogue warp on every tile, accumulator
 * stages handed back as soon as they are read): same results contract as algo 0, kept for measurement.
 * If the pass cannot be trusted it sets the rerun flag, the outputs are then undefined and the caller must call
 * again: flag 1 (a partial sum near the flush-to-zero range, or a full candidate list) -> algo 1 = the two-pass
 * path (per-row / per-column maxima, no dynamic-range limit); flag 2 (a feature beyond +-1024, outside the range
 * in which fp16 operands keep 11 bits against any partner) -> algo 2. */
size_t opb200_coarse_match_workspace_bytes(int n_batch, int L, int S, int C, float thr);
int opb200_coarse_match(const float* feat_a, const float* feat_b, const uint8_t* mask_a, const uint8_t* mask_b,
                        int mask_fill, int n_batch, int L, int S, int C, float scale, float thr, int border_rm,
                        int border_mode, int h0, int w0, int h1, int w1, const int* valid_hw, int64_t* b_ids,
                        int64_t* i_ids, int64_t* j_ids, float* mconf, int capacity, int* count_and_overflow,
                        float* conf_matrix, void* workspace, size_t workspace_bytes, int precision, int algo,
                        void* stream);

/* Same as opb200_coarse_match with the operands already split into bf16 planes [n_batch*Lp, C] / [n_batch*Sp, C]
 * (Lp, Sp = L, S rounded up to 128; pad rows may hold any finite values).  Same workspace size. */
int opb200_coarse_match_planes(const void* a_hi, const void* a_lo, const void* b_hi, const void* b_lo, const uint8_t* mask_a,
                               const uint8_t* mask_b, int mask_fill, int n_batch, int L, int S, int C, float scale, float thr,
                               int border_rm, int border_mode, int h0, int w0, int h1, int w1, const int* valid_hw,
                               int64_t* b_ids, int64_t* i_ids, int64_t* j_ids, float* mconf, int capacity,
                               int* count_and_overflow, float* conf_matrix, void* workspace, size_t workspace_bytes,
                               int precision, int algo, void* stream);

/* Measurement hook (bench.py roofline): while armed, every single-pass opb200_coarse_match call brackets its similarity
 * sweep kernel (sim_fast_kernel, the dominant kernel of the call) with a pair of CUDA events recorded on the call's
 * stream.  opb200_coarse_kernel_ms waits for the recorded pairs, returns the summed kernel time and their number, and
 * forgets them.  Not for use under CUDA-graph capture. */
int opb200_coarse_kernel_timing(int enable);
int opb200_coarse_kernel_ms(double* total_ms, long long* launches);

/* ---- Fine level --------------------------------------------------------------------------------------
 * opb200_fine_gather replaces F.unfold + index in FinePreprocess (OnePosePlus fine_preprocess.py:43-53,
 * LoFTR fine_preprocess.py:40-47): out[m, wy*W+wx, c] = feat[b_ids[m], c, 4r - W/2 + wy, 4c - W/2 + wx]
 * for coarse cell cell_ids[m] = r*Wc + c (zero outside the map).  feat element (n,c,y,x) is read from
 * feat[n*sn + c*sc + y*sh + x*sw] (any memory format; channels-last is the coalesced one).
 * Out-of-range ids (the reference raises IndexError) set *err_flag = 1 and produce zeros. */
int opb200_fine_gather(const float* feat, int64_t sn, int64_t sc, int64_t sh, int64_t sw, int n_batch, int Cf, int Hf,
                       int Wf, const int64_t* b_ids, const int64_t* cell_ids, int M, int W, int stride, int Hc, int Wc,
                       float* out, int* err_flag, void* stream);
/* out[m, c] = src[b_ids[m]*sn + c*sc + i_ids[m]*sl]: descriptors3d_db.permute(0,2,1)[b_ids, i_ids]
 * (OnePosePlus fine_preprocess.py:44,51). */
int opb200_gather_rows(const float* src, int64_t sn, int64_t sc, int64_t sl, int n_batch, int C, int L,
                       const int64_t* b_ids, const int64_t* i_ids, int M, float* out, int* err_flag, void* stream);
/* FineMatching._s2d_heatmap (OnePosePlus fine_matching.py:78-94; LoFTR fine_matching.py:43-57):
 * sim = win[m] . q[m]; heat = softmax(sim / sqrt(Cf)); expec_f[m] = (E[x], E[y], sqrt(var_x) + sqrt(var_y))
 * over the kornia-0.4.1 normalised W x W grid.  q row m starts at q + m*q_stride. */
int opb200_fine_match(const float* q, int64_t q_stride, const float* win, int M, int W, int Cf, float* expec_f,
                      void* stream);
/* bf16 arm of the fine level (BASELINE.json configs[4]): windows / query vectors stored as bf16 (half the bytes of this
 * HBM-bound stage), widened exactly and accumulated in fp32.  opb200_fine_gather_bf16 writes bf16 windows from the fp32
 * fine map; opb200_fine_match_bf16 reads bf16 q [M, Cf] (row stride q_stride elements) and win [M, W*W, Cf].  Error
 * against the fp32 / fp64 evaluation: profiles/r02_stage_fine.jsonl, tests/test_gpu_fine.py (expec_f < 2e-2). */
int opb200_fine_gather_bf16(const float* feat, int64_t sn, int64_t sc, int64_t sh, int64_t sw, int n_batch, int Cf, int Hf,
                            int Wf, const int64_t* b_ids, const int64_t* cell_ids, int M, int W, int stride, int Hc, int Wc,
                            void* out_bf16, int* err_flag, void* stream);
int opb200_fine_match_bf16(const void* q, int64_t q_stride, const void* win, int M, int W, int Cf, float* expec_f,
                           void* stream);

/* ---- Standalone LinearAttention.forward (linear_attention.py:29-61 / LoFTR :20-47) -----------------
 * q [n, L, H, D], k, v [n, S, H, D] fp32 contiguous; masks nullable bytes [n, L] / [n, S]; out [n, L, H, D]. */
size_t opb200_linear_attention_workspace_bytes(int n_batch, int S, int H, int D);
int opb200_linear_attention(const float* q, const float* k, const float* v, const uint8_t* q_mask,
                            const uint8_t* kv_mask, int n_batch, int L, int S, int H, int D, float eps, float* out,
                            void* workspace, size_t workspace_bytes, void* stream);

/* ---- Backbone glue (channels-last fp32; the convolutions stay cuDNN) -----------------------------------
 * dst[n, y, x, c] += bilinear x2 upsampling (align_corners = True) of src[n, :, :, c]: the FPN top-down add of
 * ResNetFPN_8_2.forward (src/models/OnePosePlus/backbone/resnet.py:141-164).  src [n, Hs, Ws, C], dst [n, 2Hs, 2Ws, C]. */
int opb200_upsample2x_add_nhwc(const float* src, int n_batch, int Hs, int Ws, int C, float* dst, void* stream);
/* Stem: out[n, y, x, c] = relu(bias[c] + sum_{ky,kx} image[n, 2y-3+ky, 2x-3+kx] * w[ky*7+kx][c]) -- Conv2d(1, Cout, 7, stride 2,
 * pad 3) + folded BatchNorm + ReLU of ResNetFPN_8_2 (backbone/resnet.py:85-141), channels-last output
 * [n, Ho, Wo, Cout]; image is [n, H, W] fp32; weight_tap_major is [49][Cout]; Cout must be 128. */
int opb200_stem_conv7x7_s2_relu_nhwc(const float* image, const float* weight_tap_major, const float* bias, int n_batch, int H,
                                     int W, int Cout, float* out, void* stream);
/* x[p, c] = leaky_relu(x[p, c] + bias[c], slope): folded-BatchNorm shift + nn.LeakyReLU of the FPN heads. */
int opb200_bias_leaky_relu_nhwc(float* x, const float* bias, long long n_pixels, int C, float slope, void* stream);

/* Screening sweep of the default single-pass CoarseMatching: the largest bound delta (octaves) on the fp16 similarity error,
 * c * 2^-10 * max|a_i| * max|b_j|, up to which the sweep is allowed to run on ONE fp16 product (default 0.6; beyond it,
 * or when the exact re-evaluation of the captured elements measures an error above 0.03 octave, the call asks for the
 * repeat on the three split products).  max_delta < 0 only queries.  Returns the previous value. */
float opb200_coarse_screen_max_delta(float max_delta);
/* Diagnostics of the last default-path call that used ``workspace`` (synchronises the stream): the largest row norms the
 * pack pass found on each side, the number of candidates the sweep captured, and the largest |exact - screening|
 * similarity (octaves) their re-evaluation measured. */
int opb200_coarse_norms(const void* workspace, int n_batch, int L, int S, int C, float thr, float* norm_a, float* norm_b,
                        int* candidates, float* observed, void* stream);

/* ---- Backbone convolutions on tcgen05 (NHWC fp16 activations, fp32 accumulation) ----------------------
 * Conv2d(k in {1,3}, stride in {1,2}, padding k/2, no dilation/groups) with BatchNorm folded into the weights, plus the
 * BasicBlock / FPN-head tails: y = act(conv(x) + bias + res).  Replaces conv3x3/conv1x1 + BatchNorm2d + ReLU /
 * LeakyReLU + residual add of BasicBlock.forward and the layer*_outconv* heads (src/models/OnePosePlus/backbone/
 * resnet.py:11-60, 85-164; submodules/LoFTR/src/loftr/backbone/resnet_fpn.py).  fp16 operands carry the same 11-bit
 * significand as the TF32 operands of the reference's cuDNN convolutions on a GPU.
 *   x    [n, Hi, Wi, cin_s]  fp16, cin_s = input channels padded to a multiple of 16 (pad channels hold zeros)
 *   w    [k*k][cout_s][cin_s] fp16 (tap-major, zero in padded rows/cols); cout_s in {128, 208, 256}
 *   bias [cout_s] fp32 or NULL;  res [n, Ho, Wo, cout_s] fp16 or NULL
 *   up   [n, Ho/2, Wo/2, cout_s] fp16 or NULL: adds F.interpolate(up, scale_factor=2, mode='bilinear',
 *        align_corners=True) before the activation -- the FPN top-down add (resnet.py:155-162) fused into the lateral
 *        1x1 convolution
 *   y16  [n, Ho, Wo, cout_s] fp16 and/or y32 (same shape, fp32); at least one non-NULL
 *   act  0 none, 1 ReLU, 2 LeakyReLU(slope)
 *   range_flag  nullable device int: bit 0 is set (atomically, never cleared) when a value written to y16 leaves the
 *        fp16 range (|v| > 65504, i.e. it was stored as +-inf).  fp16 activations have the significand of the
 *        reference's TF32 convolutions but not their exponent range; the host side reads the flag at the forward's one
 *        host sync and repeats the backbone on the exact-fp32 executor when it is set. */
int opb200_conv2d_nhwc_f16(const void* x, int n_img, int Hi, int Wi, int cin_s, const void* w, int ksize, int stride,
                           int cout_s, const float* bias, const void* res, const void* up, void* y16, float* y32, int act,
                           float slope, int* range_flag, void* stream);
/* Split-precision variant ("f16x3", fp32-grade): x, w, res, up and y16 hold TWO fp16 planes side by side along the channel
 * axis, v = hi + lo with hi = rn(v), lo = rn(v - hi) -- x [n, Hi, Wi, 2 * cin_s], w [k*k][cout_s][2 * cin_s],
 * res / y16 [n, Ho, Wo, 2 * cout_s], up [n, Ho/2, Wo/2, 2 * cout_s] -- and every k-block is issued for the three products
 * hi*hi, lo*hi, hi*lo (fp32 accumulation in TMEM).  The accumulator is multiplied by acc_scale before bias / residual /
 * activation: the weights are packed times a power of two (1 / acc_scale) that keeps their lo plane in fp16's normal range.
 * y32 stays [n, Ho, Wo, cout_s] fp32.  Same tail options and range_flag as above. */
int opb200_conv2d_nhwc_f16x3(const void* x, int n_img, int Hi, int Wi, int cin_s, const void* w, int ksize, int stride,
                             int cout_s, const float* bias, const void* res, const void* up, void* y16, float* y32, int act,
                             float slope, float acc_scale, int* range_flag, void* stream);
/* fp32 NHWC [n_pixels, C] -> fp16 [n_pixels, 2 * C] = [hi | lo] planes (the input format of opb200_conv2d_nhwc_f16x3). */
int opb200_split_f16_planes(const float* x, long long n_pixels, int C, void* out, int* range_flag, void* stream);
/* Option of opb200_conv2d_nhwc_f16: run as clusters of two CTAs that multicast the weight slabs to each other (TMA
 * multicast; halves the L2 -> SM weight traffic).  on = 1 / 0 sets it, on < 0 only queries; returns the previous
 * setting.  Default off (no gain measured on B200, see csrc/conv.cu); also settable with OPB200_CONV_CLUSTER=1. */
int opb200_conv_cluster(int on);
/* fp16-output variant of opb200_stem_conv7x7_s2_relu_nhwc (range_flag as above; also set by a non-finite or
 * out-of-range image value) and fp16 variant of opb200_upsample2x_add_nhwc (C % 8 == 0). */
int opb200_stem_conv7x7_s2_relu_nhwc_f16(const float* image, const float* weight_tap_major, const float* bias, int n_batch,
                                         int H, int W, int Cout, void* out_f16, int* range_flag, void* stream);
int opb200_upsample2x_add_nhwc_f16(const void* src, int n_batch, int Hs, int Ws, int C, void* dst, void* stream);

/* ---- Feature sampling and aggregation for Keypoint-Free SfM ------------------------------
 * opb200_sample_features replaces sample_feature_from_featuremap in point mode (src/KeypointFreeSfM/loftr_for_sfm/utils/
 * sample_feature_from_featuremap.py:28-81; loftr.py:130-166): out[m, c] = grid_sample(feat, kpts[m]) with
 * align_corners=True and zero padding, bilinear = 0 (nearest, round half to even) or 1; feat is ONE image's map with
 * element strides (sc, sh, sw) for (channel, row, column); kpts [M, 2] (x, y) in the (scaled) image whose size is
 * (img_h, img_w); normalize = 1 applies F.normalize(p=2, dim=-1).
 * opb200_track_feature_mean replaces the per-track np.mean of feature_aggregation_and_update (src/KeypointFreeSfM/
 * post_optimization/feature_aggregation.py:140-160): rows [R, C] fp32 grouped by track (track t owns rows
 * track_start[t] .. track_start[t + 1] - 1), out [n_tracks, C] = fp32 sum in row order / count. */
int opb200_sample_features(const float* feat, int64_t sc, int64_t sh, int64_t sw, int C, int Hf, int Wf, const float* kpts, int M,
                           float img_h, float img_w, int bilinear, int normalize, float* out, void* stream);
int opb200_track_feature_mean(const float* rows, const int64_t* track_start, int n_tracks, int C, float* out, void* stream);

/* ---- 3D keypoint position encoding ----------------------------------------------------
 * Replaces normalize_3d_keypoints (src/models/OnePosePlus/utils/normalize.py:17-28) + KeypointEncoding_linear.forward
 * (src/models/OnePosePlus/utils/position_encoding.py:46-79) with keypoints_encoder [32, 64, 128], descriptor_dim 256 and
 * norm_method "instancenorm" (statistics over the channels of each point, biased variance, eps 1e-5):
 *   out[n, c, l] = desc[n, c, l] + MLP((kpts[n, l] - mean_l kpts[n]) / (0.6 * max extent of kpts[0]))[c]
 *   kpts [n, L, 3] and desc [n, 256, L] with element strides (kn, kl, kc) / (dn, dc, dl) -- broadcast views allowed;
 *   packed_w: opb200_keypoint_encode_packed_floats() floats = W1^T [3][32], b1, W2^T [32][64], b2, W3^T [64][128], b3,
 *   W4^T [128][256], b4 (input-major weights);  normalize = 0 skips the normalisation (kpts are normalised already);
 *   out [n, 256, L] contiguous fp32;  stats_ws: 4 * n_batch floats of scratch. */
size_t opb200_keypoint_encode_packed_floats(void);
int opb200_keypoint_encode(const float* kpts, int64_t kn, int64_t kl, int64_t kc, const float* desc, int64_t dn, int64_t dc,
                           int64_t dl, const float* packed_w, int n_batch, int L, int normalize, float* out, float* stats_ws,
                           void* stream);

/* ---- keypoint encoding without any allocation ------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments and result as
 * opb200_keypoint_encode minus stats_ws: the statistics scratch is a file-scope __device__ array inside the library
 * (16 KB, held by every GPU that loads the library until exit), so the call allocates nothing (no cudaMalloc / cudaFree,
 * no allocating library routine).  Limit: n_batch <= opb200_keypoint_encode_static_max_batch() (1024); larger batches
 * return the argument error (-1) with a message in opb200_last_error().  All calls on a device share the one array:
 * calls that overlap on the GPU (different streams without ordering) race; calls ordered on one stream do not. */
int opb200_keypoint_encode_static_max_batch(void);
int opb200_keypoint_encode_static(const float* kpts, int64_t kn, int64_t kl, int64_t kc, const float* desc, int64_t dn,
                                  int64_t dc, int64_t dl, const float* packed_w, int n_batch, int L, int normalize, float* out,
                                  void* stream);

/* ---- RANSAC-PnP on the device ------------------------------------------------------
 * Replaces ransac_PnP (src/utils/metric_utils.py:121-209; call sites inference.py:181-189, 328-336): the reference copies
 * the matches of a frame to the host and runs OpenCV solvePnPRansac (EPnP, 5 px, 10000 iterations) or pycolmap's
 * absolute_pose_estimation there.  Here a whole batch of frames is solved where the match rows lie:
 *   pts3d / pts2d   device fp32, point k of the batch at pts3d + k * stride3 (x, y, z) and pts2d + k * stride2 (u, v) --
 *                   e.g. columns 0..2 and 3..4 of the matcher's [sum M, 6] rows (stride3 = stride2 = 6)
 *   offsets         device int [n_frames + 1]: frame f owns points offsets[f] .. offsets[f + 1] - 1
 *   intrinsics      device double [n_frames][4] = (fx, fy, cx, cy) of each frame's (crop) camera
 *   iterations      minimal samples (hypotheses) per frame; seed selects them (counter-based hash: deterministic)
 *   pose            device double [n_frames][3][4] = [R | t] (world -> camera), identity when the frame has fewer than 4
 *                   points or no hypothesis with >= 4 inliers (the reference returns np.eye(4)[:3] then)
 *   inlier_mask     device uint8 [sum M]; counts device int [n_frames][2] = (inliers of the refined pose, best RANSAC score)
 * P3P (Grunert) hypotheses in fp64, inlier counting in fp32, Gauss-Newton refinement on the inliers in fp64. */
size_t opb200_ransac_pnp_workspace_bytes(int n_frames, int iterations);
int opb200_ransac_pnp(const float* pts3d, int stride3, const float* pts2d, int stride2, const int* offsets, int n_frames,
                      const double* intrinsics, float reproj_error_px, int iterations, unsigned seed, double* pose,
                      unsigned char* inlier_mask, int* counts, void* workspace, size_t workspace_bytes, void* stream);

/* ---- Keypoint-Free SfM match post-processing -----------------------------------------
 * Replaces Match2Pts2D / agg_groupby_2d (src/KeypointFreeSfM/coarse_match/utils.py:5-60) and points2D_worker /
 * update_matches / transform_points2D (src/KeypointFreeSfM/coarse_match/coarse_match_worker.py:87-184).
 *   rows         device fp32 [n_rows][5] = (u0, v0, u1, v1, conf), the pairs' match rows back to back in pair order
 *   pair_of_row  device int [n_rows]; pair_images device int [n_pairs][2] = (image0, image1) ids
 *   keys         device int64 [2 * n_rows]: entry 2 r + side -> (image << 40) | (int(u) << 20) | int(v)
 *   bad_flag     device int: set when a coordinate is negative, NaN or >= 2^20 (the entry gets the largest key)
 * opb200_sfm_group_sums: after a STABLE sort of the keys (order[k] = entry id at sorted position k) and the start
 * offsets of the groups of equal keys (group_start [n_groups + 1]), sums[g] = the group's confidences added in fp64 in
 * entry order -- the arithmetic of np.bincount(weights=), so the score ranking is bit-identical to the reference's. */
int opb200_sfm_keys(const float* rows, long long n_rows, const int* pair_of_row, const int* pair_images, long long* keys,
                    int* bad_flag, void* stream);
int opb200_sfm_group_sums(const float* rows, const long long* order, const long long* group_start, long long n_groups,
                          double* sums, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* OPB200_H_ */
