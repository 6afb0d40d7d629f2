/*
 * raftk.h — C ABI of the B200 (sm_100a) RAFT correlation hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no framework types.  Every
 * pointer named "device" is a CUDA device pointer owned by the caller; the library
 * never allocates or frees caller-visible memory, never synchronises the host, and
 * launches on the stream it is given (a cudaStream_t passed as void*).  All tensors
 * are float32, dense, row-major, in the reference's layouts (NHWC, coords = (x, y)).
 *
 * Return value: 0 on success; a negative RAFTK_ERR_* for a rejected call (nothing was
 * launched); a positive value is a cudaError_t seen at the launch — the launch's own failure,
 * or an earlier error of the context that nobody had collected (it is peeked, not cleared, so
 * the host framework still sees it).  Nothing throws
 * across this boundary.  A jax.ffi handler maps non-zero to an XLA_FFI_Error
 * (see INTEGRATION.md).
 *
 * Reference citations are into /root/reference/jax_raft/model.py.
 *
 * Terms: B = image pairs; (h, w) = 1/8-resolution feature-map size; D = feature
 * channels; nq = query pixels per pair held by this GPU (h*w unless query rows are
 * sharded); Q = B*nq; L = pyramid levels; r = lookup radius;
 * level l has size (h>>l, w>>l).
 */
#ifndef RAFTK_H_
#define RAFTK_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RAFTK_MAX_LEVELS 6

enum {
  RAFTK_OK = 0,
  RAFTK_ERR_NULL = -1,        /* a required pointer is NULL */
  RAFTK_ERR_SHAPE = -2,       /* a size is non-positive or inconsistent */
  RAFTK_ERR_TOO_SMALL = -3,   /* h or w < 2 * 2^(L-1): model.py:431-436 */
  RAFTK_ERR_UNSUPPORTED = -4, /* a valid request this build has no kernel for */
  RAFTK_ERR_SCRATCH = -5,     /* scratch buffer missing or too small */
  RAFTK_ERR_ALIGN = -6,       /* a pointer is not aligned as documented */
  RAFTK_ERR_STATIC_LIMIT = -7 /* problem needs more than RAFTK_STATIC_SCRATCH_BYTES of scratch */
};

/* ABI version (major*1000 + minor) and a static description of an error code. */
int raftk_version(void);
const char* raftk_error_string(int code);

/* ------------------------------------------------------------------------------------
 * Kernel family 1 — all-pairs correlation volume + avg-pool pyramid.
 * Replaces CorrBlock.build_pyramid, model.py:418-446 (volume: 472-481, pooling: 443-445).
 *
 *   fmap1   device, (B, nq, D)       query-pixel features (fmap1 rows held by this GPU)
 *   fmap2   device, (B, h, w, D)     all target-pixel features
 *   levels  HOST array of L device pointers; levels[l] is (B*nq, h>>l, w>>l):
 *             levels[0][q, y, x] = sum_d fmap1[q, d] * fmap2[b(q), y, x, d] / sqrt(D)
 *             levels[l+1][q, y, x] = mean of the 2x2 block of levels[l] (VALID: odd
 *             trailing rows/columns dropped), as the reference pools level by level.
 *   scratch device, raftk_corr_pyramid_scratch_bytes(...) bytes, 256-byte aligned; holds
 *           the bf16 hi/mid split of both feature maps for the tensor-core kernel.
 *
 * The product kernel is a tcgen05/TMEM GEMM with the pooling fused in its epilogue.
 * Products are formed from a two-term bf16 split of each float32 operand (hi*hi +
 * mid*hi + hi*mid, fp32 accumulate): relative error ~2e-5, inside the 1e-3 the
 * north-star allows.  Pooling bounds are exact.
 * ---------------------------------------------------------------------------------- */
size_t raftk_corr_pyramid_scratch_bytes(int B, int nq, int h, int w, int D);
int raftk_corr_pyramid(const float* fmap1, const float* fmap2, float* const* levels,
                       int num_levels, int B, int nq, int h, int w, int D,
                       void* scratch, size_t scratch_bytes, void* stream);

/* Allocation-free variant of raftk_corr_pyramid.  It has no counterpart in the reference.
 *
 * For hosts that cannot hand over a scratch buffer at the call site and must not allocate one —
 * e.g. a custom-call handler inside an XLA program compiled without a scratch operand, or a
 * process whose framework pool already owns the GPU's memory, where a cudaMalloc from this
 * library would stall the device or fail.  The scratch is a fixed RAFTK_STATIC_SCRATCH_BYTES
 * array inside the library's device image: every GPU that loads the library holds those 16 MiB
 * until the process exits.  The call itself allocates nothing.
 *
 *   - Same arguments and results as raftk_corr_pyramid, minus scratch / scratch_bytes.
 *   - Limit: raftk_corr_pyramid_scratch_bytes(B, nq, h, w, D) <= RAFTK_STATIC_SCRATCH_BYTES
 *     (one 440x1024 pair at D = 256 needs 14.4 MB); larger problems return
 *     RAFTK_ERR_STATIC_LIMIT and launch nothing.
 *   - The first call on each device looks up the array's device address (and may load the
 *     module): issue it before any stream capture starts.
 *   - All calls on a device share the array, which keeps the previous call's contents.  Nothing
 *     in it needs to be zero on entry (each call rewrites everything it later reads), but calls
 *     that overlap on the same GPU race on it: order them on one stream or with events. */
#define RAFTK_STATIC_SCRATCH_BYTES (16u << 20)
int raftk_corr_pyramid_static(const float* fmap1, const float* fmap2, float* const* levels,
                              int num_levels, int B, int nq, int h, int w, int D, void* stream);

/* Same contract, plain fp32 FMA kernels (no tensor cores, no scratch).  Diagnostic:
 * used by the tests to separate indexing errors from split-precision error. */
int raftk_corr_pyramid_simt(const float* fmap1, const float* fmap2, float* const* levels,
                            int num_levels, int B, int nq, int h, int w, int D, void* stream);

/* ------------------------------------------------------------------------------------
 * Kernel family 2 — radius-r bilinear window lookup.
 * Replaces CorrBlock.index_pyramid, model.py:448-470, and its inner grid_sample,
 * model.py:24-34.
 *
 *   levels  HOST array of L device pointers, as written by raftk_corr_pyramid
 *   coords  device, (Q, 2)  centroid (x, y) in level-0 pixels
 *   out     device, (Q, L*(2r+1)^2); channel l*(2r+1)^2 + i*(2r+1) + j samples level l
 *           at (x/2^l + i - r, y/2^l + j - r): the x offset is the slow index
 *           (model.py:452-455, 463).  Bilinear, each corner zero outside the level
 *           (jax map_coordinates order=1 mode='constant'), float32 operations in the
 *           reference's order so results are bit-identical to a float32 CPU evaluation.
 * ---------------------------------------------------------------------------------- */
int raftk_corr_lookup(const float* const* levels, int num_levels, const float* coords,
                      float* out, int Q, int h, int w, int radius, void* stream);
/* The same lookup preceded, per query, by the tail of the PREVIOUS UpdateCell iteration (model.py:504-506):
 *   coords1 += delta_pre + bias (in place);  flow = coords1 - coords0;  out = lookup(coords1)
 * so that the update costs no launch of its own.  delta_pre has pixel stride ld_delta (>= 2, even), bias is
 * the flow head's two bias values (device), flow_pitched (may be NULL) a second copy of the flow with pixel
 * stride ld_flow.  num_levels == 4 and radius in {3, 4} only (else RAFTK_ERR_UNSUPPORTED: call
 * raftk_flow_update and raftk_corr_lookup). */
int raftk_corr_lookup_update(const float* const* levels, int num_levels, const float* delta_pre, int ld_delta,
                             const float* bias, const float* coords0, float* coords1, float* flow, float* flow_pitched,
                             int ld_flow, float* out, int Q, int h, int w, int radius, void* stream);

/* ------------------------------------------------------------------------------------
 * Tiled pyramid (the product path since round 2).  The reference passes the pyramid from
 * build_pyramid to index_pyramid as an opaque value (contract model.py:530-539; stored at
 * :491, :601), so its layout in memory is this library's to choose.  Level l of one query
 * pixel is stored as tiles of 4 x 4 float32 — one 64-byte L2 fetch granule each:
 *
 *   tiles_y = ceil((h>>l) / 4),  tiles_x = 2 * ceil((w>>l) / 8)
 *   slab floats = tiles_y * tiles_x * 16                   (raftk_pyramid_tiled_slab_floats)
 *   cell (y, x) of query slab q = levels[l] + q * slab + ((y>>2) * tiles_x + (x>>2)) * 16 + (y&3)*4 + (x&3)
 *   cells outside the level (y >= h>>l or x >= w>>l) hold 0.
 *
 * so a (2r+2)^2 lookup footprint costs ~10.6 granules per level instead of ~15.6 in a
 * row-major slab.  levels[l] must be 128-byte aligned.  Values, pooling bounds and lookup
 * results are those of the row-major entry points above, bit for bit.
 *
 *   raftk_corr_pyramid_tiled          = raftk_corr_pyramid writing tiled slabs (same scratch)
 *   raftk_corr_lookup_tiled(_update)  = raftk_corr_lookup(_update) reading them
 *   raftk_pyramid_tile / _untile      convert (Q, h>>l, w>>l) row-major levels <-> tiled slabs
 *                                     (for callers that bring a row-major pyramid, and for tests)
 * ---------------------------------------------------------------------------------- */
size_t raftk_pyramid_tiled_slab_floats(int h, int w, int level);
int raftk_corr_pyramid_tiled(const float* fmap1, const float* fmap2, float* const* levels,
                             int num_levels, int B, int nq, int h, int w, int D,
                             void* scratch, size_t scratch_bytes, void* stream);
int raftk_corr_lookup_tiled(const float* const* levels, int num_levels, const float* coords,
                            float* out, int Q, int h, int w, int radius, void* stream);
int raftk_corr_lookup_tiled_update(const float* const* levels, int num_levels, const float* delta_pre, int ld_delta,
                                   const float* bias, const float* coords0, float* coords1, float* flow,
                                   float* flow_pitched, int ld_flow, float* out, int Q, int h, int w, int radius,
                                   void* stream);
int raftk_pyramid_tile(const float* const* rowmajor, float* const* tiled, int num_levels, long long Q,
                       int h, int w, void* stream);
int raftk_pyramid_untile(const float* const* tiled, float* const* rowmajor, int num_levels, long long Q,
                         int h, int w, void* stream);

/* ------------------------------------------------------------------------------------
 * Tiled pyramid with float16 storage (optional; CorrBlock(storage="fp16") on the host side).
 * Same tile geometry as above with 2-byte elements: a 4 x 4 tile is one 32-byte sector, slab
 * bytes = raftk_pyramid_tiled_slab_floats(h, w, l) * 2, cell (y, x) at the same ELEMENT offset.
 * Every level is computed in float32 exactly as by raftk_corr_pyramid_tiled (level l+1 pooled
 * from the unrounded float32 level l, model.py:443-445) and rounded once, to nearest even, when
 * it is stored: relative error <= 2^-11 = 4.9e-4 per value on top of the GEMM's ~2e-5, inside
 * the 1e-3 the north-star allows for correlation values; pooling bounds and padding are exact.
 * The lookup widens the stored values to float32 and then runs the float32 tap arithmetic of
 * raftk_corr_lookup unchanged, so its output is bit-identical to a float32 CPU evaluation of
 * model.py:448-470 on the stored (rounded) pyramid.  Halves the pyramid's footprint (130.7 MB
 * per 440x1024 pair), the GEMM's store traffic and the lookup's read traffic.
 * levels[l] must be 128-byte aligned.
 *
 *   raftk_corr_pyramid_tiled_f16        replaces model.py:418-446 (4 levels, D % 64 == 0, D <= 256)
 *   raftk_corr_lookup_tiled_f16(_update) replaces model.py:448-470 (+ :504-506), any radius <= 7
 *   raftk_pyramid_tile_f16 / _untile_f16 float32 row-major levels <-> float16 tiled slabs
 * ---------------------------------------------------------------------------------- */
int raftk_corr_pyramid_tiled_f16(const float* fmap1, const float* fmap2, void* const* levels,
                                 int num_levels, int B, int nq, int h, int w, int D,
                                 void* scratch, size_t scratch_bytes, void* stream);
int raftk_corr_lookup_tiled_f16(const void* const* levels, int num_levels, const float* coords,
                                float* out, int Q, int h, int w, int radius, void* stream);
int raftk_corr_lookup_tiled_f16_update(const void* const* levels, int num_levels, const float* delta_pre, int ld_delta,
                                       const float* bias, const float* coords0, float* coords1, float* flow,
                                       float* flow_pitched, int ld_flow, float* out, int Q, int h, int w, int radius,
                                       void* stream);
int raftk_pyramid_tile_f16(const float* const* rowmajor, void* const* tiled, int num_levels, long long Q,
                           int h, int w, void* stream);
int raftk_pyramid_untile_f16(const void* const* tiled, float* const* rowmajor, int num_levels, long long Q,
                             int h, int w, void* stream);

/* Stand-alone form of the reference's grid_sample helper, model.py:24-34 (bilinear sample of an
 * NHWC image at absolute pixel coordinates, each corner zero outside; jax map_coordinates order=1
 * mode='constant').  img (n, h, w, c); grid (n, gh, gw, 2) as (x, y); out (n, gh, gw, c).  Same
 * float32 operation order as the lookup, so results match a float32 CPU evaluation bit for bit. */
int raftk_grid_sample(const float* img, const float* grid, float* out, int n, int h, int w, int c,
                      int gh, int gw, void* stream);

/* ------------------------------------------------------------------------------------
 * Kernel family 2b — "alt_corr": the same lookup computed on the fly from the feature
 * maps, never materialising the volume (no reference counterpart; equals
 * index_pyramid(build_pyramid(f1, f2), coords) by linearity of the average pool).
 *
 *   raftk_altcorr_pool_fmap2: f2_levels[0] = fmap2 (B,h,w,D) is an INPUT;
 *     f2_levels[l>=1] (B, h>>l, w>>l, D) are written: 2x2 VALID average of level l-1
 *     (the pooling rule of model.py:443-445 applied to fmap2).
 *   raftk_altcorr_lookup: fmap1 (B, nq, D), coords (B*nq, 2), out (B*nq, L*(2r+1)^2),
 *     same channel order and border rule as raftk_corr_lookup.
 * ---------------------------------------------------------------------------------- */
int raftk_altcorr_pool_fmap2(float* const* f2_levels, int num_levels, int B, int h, int w,
                             int D, void* stream);
int raftk_altcorr_lookup(const float* fmap1, const float* const* f2_levels, int num_levels,
                         const float* coords, float* out, int B, int nq, int h, int w,
                         int D, int radius, void* stream);
/* The same lookup on the tensor cores (no reference counterpart either; same result contract as
 * raftk_altcorr_lookup, i.e. index_pyramid of model.py:448-470 over the never-materialised volume).
 * Every 8 x 8 patch of query pixels at one level is a tcgen05 GEMM against the bounding box of its
 * windows; operands are bf16 [hi | mid] splits written once per pair by raftk_altcorr_split:
 *   f1_split     (B, nq, 2D) bf16  = split of fmap1 * 1/sqrt(D)   (scale = 1/sqrt(D))
 *   f2_split[l]  (B, h>>l, w>>l, 2D) bf16 = split of f2_levels[l] (scale = 1)
 * Products hi*hi + mid*hi + hi*mid in float32: ~2e-5 relative to the float32 lookup (bar 1e-3).
 * The float32 maps are still read by the scalar path that serves boxes wider than 32 or taller
 * than 30 cells and far / non-finite coordinates.  num_levels <= 4, radius in {3, 4}, D % 64 == 0,
 * D <= 256, queries forming whole rows (nq % w == 0); otherwise RAFTK_ERR_UNSUPPORTED (call
 * raftk_altcorr_lookup). */
int raftk_altcorr_split(const float* src, void* dst_bf16, long long rows, int D, float scale, void* stream);
int raftk_altcorr_lookup_tc(const float* fmap1, const float* const* f2_levels, const void* f1_split,
                            const void* const* f2_split, int num_levels, const float* coords, float* out,
                            int B, int nq, int h, int w, int D, int radius, void* stream);

/* Query-row sharding (BASELINE.json config 4; no reference counterpart, the reference is single-device):
 * the one exchange of the path — every rank needs all of fmap2 — fused with the level-0 operand split.
 *   bands      HOST array of `world` DEVICE pointers: bands[g] = rank g's rows
 *              [row_start[g], row_start[g+1]) of fmap2, (B, rows_g, w, D) float32, 16-byte aligned;
 *              peer pointers (symmetric memory / IPC mappings) for g != this rank.  Read with plain
 *              16-byte loads, each element once.
 *   sync_flags NULL, or a HOST array of `world` DEVICE pointers to RAFTK_GATHER_SYNC_BYTES bytes each,
 *              16-byte aligned, zeroed once before the first call, sync_flags[g] living on rank g and
 *              peer-mapped like bands[g]; the same blocks on every call.
 *   rank       this rank (0 <= rank < world); also where the round-robin walk over the bands starts
 *   row_start  HOST array of world + 1 ints, row_start[0] = 0, row_start[world] = h, non-decreasing
 *   full       device (local), (B, h, w, D) float32: the gathered map
 *   split_bf16 device (local), (B, h, w, 2D) bf16 [hi | mid] (the f2_split[0] of raftk_altcorr_lookup_tc),
 *              or NULL to gather only
 * world <= 8, D % 4 == 0.  Synchronisation of the ranks — every band complete before it is read, no band
 * overwritten until every rank has read it:
 *   sync_flags == NULL  the caller's, around the launch;
 *   sync_flags given    inside the kernel (flags over peer memory, see csrc/gather_split.cu): when the
 *                       launch has completed on this rank's stream the results are in place AND every
 *                       peer has finished reading this rank's band.  All ranks must make the same
 *                       sequence of calls.  Waits give up after a few seconds and set word 18 of this
 *                       rank's block (uint32, sticky) instead of hanging. */
#define RAFTK_GATHER_SYNC_BYTES 256
int raftk_gather_rows_split(const float* const* bands, void* const* sync_flags, int rank, const int* row_start,
                            int world, float* full, void* split_bf16, int B, int h, int w, int D, void* stream);

/* ------------------------------------------------------------------------------------
 * Kernel family 3 — convex 8x flow upsampling.
 * Replaces the up_mask branch of upsample_flow, model.py:85-98.
 *
 *   flow  device, (B, h, w, 2)
 *   mask  device, (B, h, w, 576); channel k*64 + fy*8 + fx, k = (dy+1)*3 + (dx+1)
 *   out   device, (B, 8h, 8w, 2):
 *         out[b, 8y+fy, 8x+fx, c] = sum_k softmax_k(mask[b,y,x,:,fy,fx]) * 8*flow[b,y+dy,x+dx,c]
 *         with zero flow outside the image.
 * ---------------------------------------------------------------------------------- */
int raftk_convex_upsample(const float* flow, const float* mask, float* out, int B, int h,
                          int w, void* stream);
/* Same with logits mask + mask_bias[channel] (device, 576 floats, 16-byte aligned): lets the caller
 * skip the bias pass of the mask predictor's last convolution (model.py:398-400). */
int raftk_convex_upsample_bias(const float* flow, const float* mask, const float* mask_bias, float* out, int B,
                               int h, int w, void* stream);

/* ------------------------------------------------------------------------------------
 * Next row f1: raftk_corr_lookup fused with the 1x1 convolution that consumes it
 * (MotionEncoder.convcorr1 + ReLU, model.py:275-276), so the L*(2r+1)^2 correlation features stay
 * on the SM:   out[q, n] = act( sum_c weight[n, c] * lookup(q)[c] + bias[n] ),  act = ReLU if relu.
 * The product is TF32 (tcgen05.mma.kind::tf32: 10-bit-mantissa operands, float32 accumulation),
 * as the library convolution it replaces in the default precision mode.
 *
 *   levels, coords     as raftk_corr_lookup; num_levels == 4 and radius in {3, 4} only
 *                      (else RAFTK_ERR_UNSUPPORTED: use raftk_corr_lookup + a convolution)
 *   weight  device, (N, K') float32 row-major, 16-byte aligned, K' = raftk_lookup_conv_weight_cols():
 *           each level's (2r+1)^2 filter columns followed by zeros up to a multiple of 32, i.e.
 *           weight[n, l*KL + t] = W[n, l*(2r+1)^2 + t], KL = K'/4 (96 for r = 4, 64 for r = 3)
 *   bias    device, (N), 16-byte aligned;  N a multiple of 16, 16 <= N <= 256
 *   out     device, (Q, N)
 * ---------------------------------------------------------------------------------- */
int raftk_lookup_conv_weight_cols(int num_levels, int radius);
int raftk_corr_lookup_conv1x1(const float* const* levels, int num_levels, const float* coords, const float* weight,
                              const float* bias, float* out, int Q, int h, int w, int radius, int N, int relu,
                              void* stream);

/* ------------------------------------------------------------------------------------
 * Next row f2: MaskPredictor's last 1x1 convolution (model.py:394-400, with its 0.25
 * factor folded into weight and bias) fused with raftk_convex_upsample, so the 576-channel mask is
 * never written:  mask[p, ch] = sum_c W[ch, c] * m[p, c] + bias[ch]  (TF32 product), then the convex
 * combination of model.py:85-98.
 *   m       device, (B*h*w, K) float32 (the 3x3 convolution + ReLU before it), 16-byte aligned;
 *           K a multiple of 32, K <= 256 (else RAFTK_ERR_UNSUPPORTED)
 *   weight  device, (576, K) float32, rows in PASS order: row g*144 + k*16 + t holds the filter of mask
 *           channel k*64 + 16*g + t  (g = 0..3, k = 0..8, t = 0..15);  bias (576) in the same order
 *   flow    device, (B, h, w, 2);   out  device, (B, 8h, 8w, 2)
 * ---------------------------------------------------------------------------------- */
int raftk_mask_upsample(const float* m, const float* weight, const float* bias, const float* flow, float* out, int B,
                        int h, int w, int K, void* stream);

/* Host-model glue, not a hot-path row: in place y = act(x + bias[c]) on an NHWC conv output
 * (npix, c); act 0 none, 1 relu, 2 sigmoid, 3 tanh.  Used by jax_raft_b200/model.py after the
 * library convolutions (ConvNormActivation / ConvGRU / FlowHead epilogues, model.py:120-159,
 * 293-350), which are called without bias. */
int raftk_bias_act_nhwc(float* x, const float* bias, long long npix, int c, int act, void* stream);

/* Host-model glue, not a hot-path row: in place InstanceNorm (+ ReLU) of an NHWC tensor (n, npix, c),
 * the feature encoder's norm layer (model.py:706-707, applied in ConvNormActivation, model.py:146-158):
 * y = (x - mean[n,c]) / sqrt(var[n,c] + eps), biased variance.  c % 4 == 0, c <= 1024.
 * scratch: raftk_instance_norm_scratch_bytes(n, c) bytes, 8-byte aligned (float64 sums). */
size_t raftk_instance_norm_scratch_bytes(int n, int c);
int raftk_instance_norm_nhwc(float* x, int n, int npix, int c, float eps, int relu, void* scratch,
                             size_t scratch_bytes, void* stream);
/* Same, followed by the residual block's tail (model.py:183-184): x = relu(residual + y), y the
 * normalised (and, if relu, rectified) value; residual (n, npix, c), 16-byte aligned. */
int raftk_instance_norm_residual_nhwc(float* x, const float* residual, int n, int npix, int c, float eps, int relu,
                                      void* scratch, size_t scratch_bytes, void* stream);
/* Host-model glue: x = relu(x + y) over count floats (count % 4 == 0, 16-byte aligned): the residual
 * tail of blocks whose norm layer is folded into the convolution (model.py:183-184, 215-216). */
int raftk_add_relu(float* x, const float* y, long long count, void* stream);

/* Host-model glue, not hot-path rows: the element-wise parts of ConvGRU (model.py:293-312) on NHWC
 * tensors.  `h` points at the hidden-state channel slice inside the z/r-convolution input buffer
 * (pixel stride ldh floats; hidden, ldh, ld_rh even, pointers 8-byte aligned); zr_pre (npix, 2*hidden) is the output of the z and r filters stacked in
 * one convolution, bias_zr their stacked biases.
 *   gate:   z_out (npix, hidden) = sigmoid(z_pre + bz);  rh_out (pixel stride ld_rh) = sigmoid(r_pre + br) * h
 *   update: h = (1 - z) * h + z * tanh(q_pre + bq), written in place and to h_dense (npix, hidden). */
int raftk_gru_gate(const float* zr_pre, const float* bias_zr, const float* h, int ldh, float* z_out,
                   float* rh_out, int ld_rh, long long npix, int hidden, void* stream);
int raftk_gru_update(const float* q_pre, const float* bq, const float* z, float* h, int ldh, float* h_dense,
                     long long npix, int hidden, void* stream);
/* MotionEncoder tail (model.py:286-290): relu(cf_pre + bias) followed by the 2 flow channels, written
 * at dst_a and dst_b (pixel stride ld floats) — the motion slots of the two GRU convolution input
 * buffers.  cf_pre is (npix, c_cf) with pixel stride ld_src >= c_cf (the convolution may have been
 * given zero filters up to a channel count the library does not have to pad).  c_cf, ld, ld_src even;
 * pointers 8-byte aligned. */
int raftk_motion_scatter(const float* cf_pre, int ld_src, const float* bias, const float* flow, float* dst_a,
                         float* dst_b, int ld, long long npix, int c_cf, void* stream);
/* UpdateCell tail (model.py:504-506): coords1 += delta_pre + bias (in place), flow = coords1 - coords0,
 * on (npix, 2) arrays.  delta_pre has pixel stride ld_delta (>= 2, even: padded flow-head output);
 * bias: device, the flow head's two bias values; flow_pitched (may be NULL): a second copy of the flow
 * with pixel stride ld_flow (the zero-padded input of the next iteration's 7x7 flow convolution). */
int raftk_flow_update(const float* delta_pre, int ld_delta, const float* bias, const float* coords0, float* coords1,
                      float* flow, float* flow_pitched, int ld_flow, long long npix, void* stream);

/* Tuning knob, no counterpart in the reference: the short kernels (lookup, update-block glue) are launched
 * with programmatic stream serialization, so that each may be scheduled while its predecessor drains and
 * waits in its first instruction (griddepcontrol.wait) — this hides most of the node-to-node launch
 * latency inside the forward's CUDA graph.  raftk_set_pdl(0) launches them the ordinary way (initial value:
 * on, or off if the environment has RAFTK_PDL=0); returns the previous setting.  Process-wide, not
 * synchronised: set it before launching. */
int raftk_set_pdl(int on);

/* Which kernel raftk_altcorr_lookup_tc launches (no counterpart in the reference): 0 = by problem size
 * (default, or RAFTK_ALTCORR_MODE in the environment), 1 = 8 x 8 query patches, UMMA M = 64, A tile in shared
 * memory (csrc/altcorr_tc.cu), 2 = 16 x 8 patches, UMMA M = 128, A operand in tensor memory
 * (csrc/altcorr_tc128.cu).  Same results contract either way (values agree to the split-product error,
 * ~2e-5).  Returns the previous setting; other values only query it.  Process-wide, not synchronised. */
int raftk_set_altcorr_mode(int mode);

#ifdef __cplusplus
}
#endif
#endif /* RAFTK_H_ */
