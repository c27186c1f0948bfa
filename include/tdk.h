/*
 * tdk.h -- C-ABI of libtdk_b200.so: Tadataka's semi-dense inverse-depth path on B200 (sm_100a).
 *
 * This is the drop-in boundary.  Each entry point replaces one function of the reference's
 * pyo3 extension `rust_bindings` (the FFI the reference's Python calls); the citation beside
 * each declaration is the reference interface it stands in for (paths relative to the
 * reference checkout).  Plain pointers and sizes only, no torch / numpy types.
 *
 * Conventions
 *   - every entry point returns 0 (TDK_OK) or a negative TDK_ERR_* code; tdk_last_error()
 *     gives the message of the calling thread's last failure;
 *   - `*_dev` entry points take DEVICE pointers, enqueue on `stream` (a cudaStream_t passed as
 *     void*; NULL = legacy default stream) and do not synchronise unless stated;
 *   - `*_host` entry points take HOST pointers, stage through a grow-only per-process device
 *     arena, and return after the results are back in host memory;
 *   - images and maps are row-major [H,W]; coordinates are (x, y); a camera is 4 doubles
 *     {fx, fy, ox, oy}; a transform is a row-major 4x4 (16 doubles);
 *   - params is 6 HOST doubles {min_depth, max_depth, geo_coeff, photo_coeff, ref_step_size,
 *     min_gradient}: the arguments of rust_bindings.semi_dense.Params in order
 *     (src/py/semi_dense.rs:91-109); the inversion of the depth range happens inside;
 *   - all arithmetic is IEEE f64 without FMA contraction, evaluated in the reference's order.
 */
#ifndef TDK_H_
#define TDK_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    TDK_OK = 0,
    TDK_ERR_INVALID = -1,     /* bad argument (null pointer, non-positive size, bad step)       */
    TDK_ERR_CUDA = -2,        /* a CUDA runtime call failed; see tdk_last_error()               */
    TDK_ERR_AGE = -3,         /* an age exceeds the number of reference frames
                                 (the reference prints and exits: semi_dense.rs:202-205)        */
    TDK_ERR_SINGULAR = -4,    /* a reference pose is not invertible (the reference panics in
                                 transform_wr.inv().unwrap(), semi_dense.rs:87)                 */
    TDK_ERR_WORKSPACE = -5,   /* workspace too small                                            */
    TDK_ERR_SEARCH_LEN = -6,  /* an epipolar search line exceeds TDK_MAX_SEARCH_SAMPLES         */
    TDK_ERR_RANGE = -7        /* coordinate outside the image (tadataka/interpolation/
                                 __init__.py:23-29 raises ValueError)                           */
};

/* src/semi_dense/flag.rs:3-14 == tadataka/vo/semi_dense/flag.py:4-14 */
enum {
    TDK_FLAG_SUCCESS = 0,
    TDK_FLAG_HYPOTHESIS_OUT_OF_SEARCH_RANGE = -1,
    TDK_FLAG_KEY_OUT_OF_RANGE = -2,
    TDK_FLAG_REF_CLOSE_OUT_OF_RANGE = -3,
    TDK_FLAG_REF_FAR_OUT_OF_RANGE = -4,
    TDK_FLAG_REF_EPIPOLAR_TOO_SHORT = -5,
    TDK_FLAG_INSUFFICIENT_GRADIENT = -6,
    TDK_FLAG_NEGATIVE_PRIOR_DEPTH = -7,
    TDK_FLAG_NEGATIVE_REF_DEPTH = -8,
    TDK_FLAG_NOT_PROCESSED = -9
};

#define TDK_MAX_SEARCH_SAMPLES (1 << 20)

const char *tdk_last_error(void);
/* "tadataka-b200 <version> sm_100a" */
const char *tdk_version(void);

/* ------------------------------------------------------------------ update_depth
 * Replaces rust_bindings.semi_dense.update_depth (src/py/semi_dense.rs:157-187 ->
 * src/semi_dense/semi_dense.rs:160-234), batched over P independent keyframes, each with its own
 * R reference frames; pixel (y,x) of keyframe p uses reference R - age[p,y,x] (semi_dense.rs:207).
 *
 *   key_imgs [P,H,W]  key_cams [P,4]  T_wk [P,16]
 *   ref_imgs [P,R,H,W]  ref_cams [P,R,4]  T_wr [P,R,16]
 *   age [P,H,W] u64   prior_depth, prior_variance [P,H,W]
 *   out_depth, out_variance [P,H,W] f64   out_flag [P,H,W] i64
 *   dbg_argmin (optional) [P,H,W] i32: search result argmin+2 (intensities.rs:32-36), -1 where no
 *       search ran -- the "extra debug output" the parity gate compares
 *   out_hist (optional) [P,10] u64: per-keyframe count of each flag, index = -flag; ADDED to
 *   fuse_on_success: 0 = the reference's update_depth (the NEW measurement is returned, F5);
 *       1 = where the flag is SUCCESS the measurement is fused with the prior in inverse depth,
 *       exactly as propagation.rs:27-34 fuses two hypotheses
 *       (id, var) = fusion(inv(prior_depth), inv(new_depth), prior_variance, new_variance);
 *       depth = inv(id) -- identical to calling tdk_fusion_dev(depth_domain=1) afterwards,
 *       without the second pass over HBM
 *   workspace: device scratch -- status words, per-pair constants and, for up to 64 keyframes at a
 *       time (a "sub-batch"; larger launches reuse the arrays sub-batch after sub-batch in stream
 *       order), per pixel: 4 bytes of queue for the pixels the sieve cannot decide, a 4-byte f32 copy
 *       of the key image (the sieve's certified rejection reads it), and 133 bytes for
 *       the search lines (128-byte record, length class, sort position; worst case = every pixel is
 *       searched).  tdk_update_depth_workspace_bytes(P,R,H,W) is the size that runs fastest
 *       (min(P,64) keyframes per sub-batch; 2.8 GB for 64 VGA keyframes);
 *       tdk_update_depth_workspace_bytes_min(P,R,H,W) is the least that is accepted (one keyframe
 *       per sub-batch); anything in between is used as far as it goes
 *
 * tdk_update_depth_status() copies the launch's status word back (synchronises the stream) and
 * returns TDK_OK / TDK_ERR_AGE / TDK_ERR_SINGULAR / TDK_ERR_SEARCH_LEN.
 */
size_t tdk_update_depth_workspace_bytes(int64_t P, int64_t R, int64_t H, int64_t W);
size_t tdk_update_depth_workspace_bytes_min(int64_t P, int64_t R, int64_t H, int64_t W);

int tdk_update_depth_dev(const double *key_imgs, const double *key_cams, const double *T_wk,
                         const double *ref_imgs, const double *ref_cams, const double *T_wr,
                         const uint64_t *age, const double *prior_depth,
                         const double *prior_variance, const double *params, int64_t P, int64_t R,
                         int64_t H, int64_t W, double *out_depth, double *out_variance,
                         int64_t *out_flag, int32_t *dbg_argmin, uint64_t *out_hist,
                         int fuse_on_success, void *workspace, size_t workspace_bytes,
                         void *stream);

/* Same, for frames that live in separate allocations (a keyframe ring kept on the device):
 * key_img_ptrs [P] and ref_img_ptrs [P*R] are HOST arrays of DEVICE image pointers. */
int tdk_update_depth_frames_dev(const double *const *key_img_ptrs, const double *key_cams,
                                const double *T_wk, const double *const *ref_img_ptrs,
                                const double *ref_cams, const double *T_wr, const uint64_t *age,
                                const double *prior_depth, const double *prior_variance,
                                const double *params, int64_t P, int64_t R, int64_t H, int64_t W,
                                double *out_depth, double *out_variance, int64_t *out_flag,
                                int32_t *dbg_argmin, uint64_t *out_hist, int fuse_on_success,
                                void *workspace, size_t workspace_bytes, void *stream);
/* Same again with the two pointer tables themselves in DEVICE memory (a caller that keeps its
 * frame ring and its tables resident: no host -> device copy is left in the call). */
int tdk_update_depth_tables_dev(const double *const *key_img_ptrs_dev, const double *key_cams,
                                const double *T_wk, const double *const *ref_img_ptrs_dev,
                                const double *ref_cams, const double *T_wr, const uint64_t *age,
                                const double *prior_depth, const double *prior_variance,
                                const double *params, int64_t P, int64_t R, int64_t H, int64_t W,
                                double *out_depth, double *out_variance, int64_t *out_flag,
                                int32_t *dbg_argmin, uint64_t *out_hist, int fuse_on_success,
                                void *workspace, size_t workspace_bytes, void *stream);

int tdk_update_depth_status(const void *workspace, void *stream);

/* No counterpart in the reference.  tdk_update_depth_dev needs a workspace the caller has to
 * allocate; some callers cannot: a tracking loop that must not hit cudaMalloc's device-wide
 * synchronisation (or its failure) once it is running, or a host application that gave this
 * library no memory budget of its own and only passes in the buffers it owns.  This variant keeps
 * the scratch (status words and per-pair constants, TDK_STATIC_SCRATCH_BYTES bytes) in a static
 * device array inside the library and calls no cudaMalloc / cudaFree or any other allocating routine.
 *   - limit: P * R <= TDK_STATIC_MAX_PAIRS keyframe/reference pairs per call; larger batches are
 *     refused with TDK_ERR_INVALID and a message naming the limit;
 *   - arguments as tdk_update_depth_dev (contiguous [P,H,W] / [P,R,H,W] device arrays; inputs and
 *     outputs still belong to the caller), without the workspace;
 *   - wait_and_check != 0: synchronise `stream` and return TDK_ERR_AGE / _SINGULAR / _SEARCH_LEN as
 *     tdk_update_depth_status would; 0: enqueue only, errors of that kind go unreported;
 *   - the first call on each device looks up the device address of the array
 *     (cudaGetSymbolAddress, once per device, with that device current); make that first call
 *     outside any stream capture;
 *   - all calls on a device share the one array: the status words are cleared at the start of every
 *     call, the rest is overwritten, but two calls whose kernels can overlap on the GPU (different
 *     streams, no ordering between them) race on it -- order them, or use tdk_update_depth_dev with
 *     separate workspaces. */
#define TDK_STATIC_MAX_PAIRS 64
#define TDK_STATIC_SCRATCH_BYTES 32768
int tdk_update_depth_static_dev(const double *key_imgs, const double *key_cams, const double *T_wk,
                                const double *ref_imgs, const double *ref_cams, const double *T_wr,
                                const uint64_t *age, const double *prior_depth,
                                const double *prior_variance, const double *params, int64_t P,
                                int64_t R, int64_t H, int64_t W, double *out_depth,
                                double *out_variance, int64_t *out_flag, int32_t *dbg_argmin,
                                uint64_t *out_hist, int fuse_on_success, int wait_and_check,
                                void *stream);

/* Tuning knob with no counterpart in the reference: how update_depth is mapped onto kernels.
 * 2 = an f32 copy of the key images, a sieve over all pixels that gives every pixel it can decide
 * for certain its final values and queues the others in the workspace, a kernel that runs the exact
 * prefix of `estimate` on the queued pixels and sets up their search lines, a counting sort of the
 * lines by length, and a search kernel in which every warp walks 32 lines of equal length; 0 = the
 * copy and the sieve, then one kernel over 256 queued pixels per CTA (smaller workspace: 8 bytes
 * per pixel); 1 = one kernel, each CTA runs the exact path on its own 64x4 tile (what
 * tdk_update_depth_static_dev runs); -1 (default) = 0 for launches of at most 640x480 pixels in
 * total, 2 for larger ones.  Results do not depend on it. */
int tdk_set_kernel_plan(int plan);

/* No counterpart in the reference: its search returns an integer sample index
 * (src/semi_dense/intensities.rs:11-36, used at semi_dense.rs:147-148).  on != 0: every later
 * update_depth call fits a parabola through the errors of the best window and its two neighbours
 * and triangulates from the sample position at its minimum (offset within half a sample) instead
 * of the sample itself -- depth and variance then differ from the reference's; flags and dbg_argmin
 * do not change except through the final range check.  Default 0 (off): results are the reference's. */
int tdk_set_subpixel_refinement(int on);

/* Host-pointer form: what a numpy caller of the reference binds.  Copies in, runs, copies out,
 * checks the status.  Host buffers may be pageable or pinned: the call works in chunks of keyframes
 * (about 96 MB of input each, three in flight on their own streams) so that uploads, kernels and
 * read-backs of different chunks overlap -- with PINNED buffers (cudaHostAlloc / cudaHostRegister);
 * copies from pageable memory are staged by the driver and serialise with the calling thread.
 * Calls on different devices do not block each other; calls on one device are serialised. */
int tdk_update_depth_host(const double *key_imgs, const double *key_cams, const double *T_wk,
                          const double *ref_imgs, const double *ref_cams, const double *T_wr,
                          const uint64_t *age, const double *prior_depth,
                          const double *prior_variance, const double *params, int64_t P, int64_t R,
                          int64_t H, int64_t W, double *out_depth, double *out_variance,
                          int64_t *out_flag, int32_t *dbg_argmin, uint64_t *out_hist,
                          int fuse_on_success);

/* Host-pointer form with the frames passed by pointer: key_img_ptrs [P] and ref_img_ptrs [P*R] are
 * HOST arrays of HOST image pointers ([H,W] f64 each).  An image named more than once -- the
 * reference frames of the reference's own loop, where every new keyframe is matched against the
 * frames before it (examples/semi_dense_vo.py:196-203; src/py/semi_dense.rs:157-187 clones every
 * reference per call) -- is uploaded once per call, on a stream of its own, in the order the keyframes
 * need them; cameras, poses and pointer tables go up once; the maps travel in chunks of about 32 MB
 * (two VGA keyframes), six in flight, so that both directions of the link stay busy.  For R < 255
 * the age maps (values <= R; anything larger is reported as TDK_ERR_AGE whatever its value) and the
 * flag maps (-9 .. 0) cross the link as ONE byte per pixel instead of eight: a small pool of host
 * threads (TDK_HOST_THREADS, default min(6, hardware threads), created on first use) narrows the
 * caller's age maps into pinned memory ahead of the uploads and widens the flags into out_flag
 * behind the read-backs; two small kernels do the opposite on the device.  The arrays the caller
 * sees keep the reference's u64 / i64 layout.  Everything else as tdk_update_depth_host. */
int tdk_update_depth_frames_host(const double *const *key_img_ptrs, const double *key_cams,
                                 const double *T_wk, const double *const *ref_img_ptrs,
                                 const double *ref_cams, const double *T_wr, const uint64_t *age,
                                 const double *prior_depth, const double *prior_variance,
                                 const double *params, int64_t P, int64_t R, int64_t H, int64_t W,
                                 double *out_depth, double *out_variance, int64_t *out_flag,
                                 int32_t *dbg_argmin, uint64_t *out_hist, int fuse_on_success);

/* The numpy drop-in's call for a caller that keeps its FRAMES on the device and its MAPS on the host
 * (rust_bindings.semi_dense.update_depth, src/py/semi_dense.rs:157-187: Frame objects are built once
 * and serve as references for many later calls; the maps are numpy arrays): key_img_ptrs [P] and
 * ref_img_ptrs [P*R] are HOST arrays of DEVICE image pointers (as tdk_update_depth_frames_dev);
 * cameras, poses and all maps are HOST arrays, pageable or pinned.  Pageable maps are moved by the
 * library's worker pool through a pinned ring (512 KB pieces) instead of the driver's own staging;
 * ages and flags cross the link as bytes when R < 255.  Returns when the results are in host
 * memory, status checked. */
int tdk_update_depth_maps_host(const double *const *key_img_ptrs, const double *key_cams,
                               const double *T_wk, const double *const *ref_img_ptrs,
                               const double *ref_cams, const double *T_wr, const uint64_t *age,
                               const double *prior_depth, const double *prior_variance,
                               const double *params, int64_t P, int64_t R, int64_t H, int64_t W,
                               double *out_depth, double *out_variance, int64_t *out_flag,
                               int32_t *dbg_argmin, uint64_t *out_hist, int fuse_on_success);

/* Self-test of the host half of tdk_update_depth_frames_host's byte wire (needs no GPU): the ages go
 * through the worker pool's narrowing (saturating at 255) and come back widened, the flags (-128 ..
 * 127) come back through its widening.  No counterpart in the reference. */
int tdk_selftest_host_wire(const uint64_t *age, const int64_t *flag, int64_t n, uint64_t *age_back,
                           int64_t *flag_back);

/* Replaces rust_bindings.semi_dense.estimate_debug_ (src/py/semi_dense.rs:126-155): one pixel,
 * host pointers, returns the prior on failure.  u_key is (x, y).  out = {depth, variance},
 * *out_flag the flag, dbg (optional, 1 value) = argmin+2 of the search, or -1. */
int tdk_estimate_debug_host(const int64_t *u_key, double prior_depth, double prior_variance,
                            const double *key_img, const double *key_cam, const double *T_wk,
                            const double *ref_img, const double *ref_cam, const double *T_wr,
                            const double *params, int64_t H, int64_t W, double *out,
                            int64_t *out_flag, int64_t *dbg);

/* ------------------------------------------------------------------ companions
 * tdk_fusion_dev: src/semi_dense/fusion.rs:13-42 (`fusion_arrays`; not exported by the reference's
 * pyo3 module, called as tadataka.vo.semi_dense.fusion.fusion in examples/semi_dense_depth.py:41).
 * depth_domain = 1: mu1/mu2/mu are DEPTH maps, fused in inverse depth as propagation.rs:27-34 does
 * (inv() on the way in and out).
 * If flag != NULL only elements with flag == 0 are fused, the others keep (mu1, var1): the
 * "fuse where the update succeeded" step of examples/semi_dense_vo.py:82-88. */
int tdk_fusion_dev(const double *mu1, const double *mu2, const double *var1, const double *var2,
                   const int64_t *flag, int64_t n, int depth_domain, double *mu, double *var,
                   void *stream);

/* src/semi_dense/regularization.rs:29-64 (disabled in the reference build; semantics as written):
 * per keyframe [P,H,W]. */
int tdk_regularize_dev(const double *depth, const double *variance, const int64_t *flag, int64_t P,
                       int64_t H, int64_t W, double *out_depth, void *stream);

/* rust_bindings.semi_dense.increment_age (src/py/semi_dense.rs:35-51 -> src/semi_dense/age.rs:6-32)
 * for P independent frames: age0 [P,H,W] u64, cam0/cam1 [P,4], T10 [P,16], depth0 [P,H,W].
 * workspace: tdk_increment_age_workspace_bytes(P,H,W). */
size_t tdk_increment_age_workspace_bytes(int64_t P, int64_t H, int64_t W);
int tdk_increment_age_dev(const uint64_t *age0, const double *cam0, const double *cam1,
                          const double *T10, const double *depth0, int64_t P, int64_t H, int64_t W,
                          uint64_t *age1, void *workspace, size_t workspace_bytes, void *stream);

/* Reference-frame pruning.  update_depth reads refframes[len - age], so a frame
 * is needed exactly as long as some pixel carries the matching age; the reference's loop leaves the
 * clean-up as a TODO (examples/semi_dense_vo.py:202) and keeps every frame.  counts [P, max_age+2]
 * u64 receives, per keyframe, the number of pixels of each age 0..max_age and, in the last bin, of
 * all older ones; the caller drops the frames older than the largest age in use. */
int tdk_age_usage_dev(const uint64_t *age, int64_t P, int64_t H, int64_t W, int64_t max_age,
                      uint64_t *counts, void *stream);

/* rust_bindings.semi_dense.propagate (src/py/semi_dense.rs:189-218 ->
 * src/semi_dense/propagation.rs:48-92) for P independent frames.  Collisions are folded in raster
 * order of the source pixel, as the reference's loop does.
 * workspace: tdk_propagate_workspace_bytes(P,H,W). */
size_t tdk_propagate_workspace_bytes(int64_t P, int64_t H, int64_t W);
int tdk_propagate_dev(const double *T10, const double *cam0, const double *cam1,
                      const double *depth0, const double *variance0, int64_t P, int64_t H,
                      int64_t W, double default_depth, double default_variance,
                      double uncertaintity_bias, double *depth1, double *variance1,
                      void *workspace, size_t workspace_bytes, void *stream);

/* increment_age and propagate in one pass over the same warp, as the reference's loop calls them
 * back to back with the same (T10, cameras, depth0) (examples/semi_dense_vo.py:185-194): age1 is what
 * tdk_increment_age_dev returns, depth1 / variance1 what tdk_propagate_dev returns.  Same workspace
 * as tdk_propagate_dev. */
int tdk_propagate_age_dev(const double *T10, const double *cam0, const double *cam1,
                          const double *depth0, const double *variance0, const uint64_t *age0,
                          int64_t P, int64_t H, int64_t W, double default_depth,
                          double default_variance, double uncertaintity_bias, double *depth1,
                          double *variance1, uint64_t *age1, void *workspace, size_t workspace_bytes,
                          void *stream);

/* Host-pointer forms of the companions: what a numpy caller of the reference binds
 * (rust_bindings.semi_dense.increment_age / propagate, src/py/semi_dense.rs:35-51,189-218: numpy in,
 * numpy out, one call; fusion.rs:13-42 and regularization.rs:29-64 for the two the reference keeps
 * internal).  Arguments as the *_dev forms without workspace and stream; host buffers may be
 * pageable or pinned; the call stages through the per-device arena and returns when the results
 * are in host memory.  Pinned arrays are copied by the DMA engines directly; pageable ones (numpy)
 * of 128 KB and more are moved by the library's host threads through a pinned ring in 512 KB pieces,
 * each piece's DMA starting as soon as it is staged (one VGA call: increment_age 0.47 -> 0.37 ms,
 * propagate 0.64 -> 0.45 ms against the driver's own pageable copies).  flag may be NULL in
 * tdk_fusion_host. */
int tdk_increment_age_host(const uint64_t *age0, const double *cam0, const double *cam1,
                           const double *T10, const double *depth0, int64_t P, int64_t H, int64_t W,
                           uint64_t *age1);
int tdk_propagate_host(const double *T10, const double *cam0, const double *cam1,
                       const double *depth0, const double *variance0, int64_t P, int64_t H, int64_t W,
                       double default_depth, double default_variance, double uncertaintity_bias,
                       double *depth1, double *variance1);
int tdk_fusion_host(const double *mu1, const double *mu2, const double *var1, const double *var2,
                    const int64_t *flag, int64_t n, int depth_domain, double *mu, double *var);
int tdk_regularize_host(const double *depth, const double *variance, const int64_t *flag, int64_t P,
                        int64_t H, int64_t W, double *out_depth);

/* ------------------------------------------------------------------ keyframe loop
 * The reference's caller keeps the maps and the reference frames in numpy and calls the extension
 * three times per frame (examples/semi_dense_vo.py:185-203: increment_age, propagate, update_depth,
 * refframes.append; Frame copies every image once more, src/py/semi_dense.rs:53-67).  tdk_mapper is
 * that loop with everything resident in HBM behind one call per frame; no counterpart in the
 * reference.
 *   - all device memory is ONE caller-owned arena of tdk_mapper_arena_bytes(H, W, capacity) bytes:
 *     the ring of reference images (capacity frames, contiguous, so update_depth runs without
 *     pointer tables), their cameras and poses, two sets of maps, scratch;
 *   - frames arrive as HOST images: TDK_IMAGE_F64 [H,W] f64 gray (what Frame takes),
 *     TDK_IMAGE_U8 [H,W] 8-bit gray (v / 255), TDK_IMAGE_RGB8 [H,W,3] / TDK_IMAGE_RGBA8 [H,W,4]
 *     8-bit colour, converted on the device with the luma the reference's dataset loaders apply
 *     (skimage rgb2gray, tadataka/dataset/new_tsukuba.py:85-86):
 *     ((R/255) 0.2125 + (G/255) 0.7154) + (B/255) 0.0721, every operation rounded, alpha dropped;
 *   - tdk_mapper_reset: first frame, age 0, depth / variance from the host maps given or, if NULL,
 *     the defaults (semi_dense_vo.py:210-219);
 *   - tdk_mapper_track: next frame with its pose T_wf (world <- frame) and T10 = inv(T_wf) * T_w(previous)
 *     (the caller computes it, as the reference's loop does): propagate + increment_age, update_depth
 *     against every frame of the ring [fuse_on_success] [regularize], append the frame.  Enqueues
 *     on the mapper's stream and returns without waiting for the GPU; errors of the kind
 *     tdk_update_depth_status reports (age, singular pose, search length) surface at most two
 *     calls later (a frame's status words are read when their buffer is reused) or in tdk_mapper_sync;
 *   - prune != 0: frames older than the largest age on the map are dropped from the ring (the
 *     reference's TODO, semi_dense_vo.py:202); results do not change, the ring stops growing;
 *     without pruning (or with too small a capacity) a full ring makes track return
 *     TDK_ERR_WORKSPACE;
 *   - tdk_mapper_sync waits for the stream and reports a pending error; tdk_mapper_prune counts the
 *     ages of the current map and drops the unused frames at once; tdk_mapper_state gives the
 *     number of frames held and the byte offsets, inside the arena, of the current depth / variance
 *     (f64), age (u64) and flag (i64) maps [H,W] and of the ring, whose live window starts at frame
 *     *ring_start. */
enum { TDK_IMAGE_F64 = 0, TDK_IMAGE_U8 = 1, TDK_IMAGE_RGB8 = 2, TDK_IMAGE_RGBA8 = 3 };
typedef struct tdk_mapper tdk_mapper;
size_t tdk_mapper_arena_bytes(int64_t H, int64_t W, int64_t capacity);
int tdk_mapper_create(void *arena, size_t arena_bytes, int64_t H, int64_t W, int64_t capacity,
                      const double *params, double default_depth, double default_variance,
                      double uncertaintity_bias, int fuse_on_success, int regularize, int prune,
                      void *stream, tdk_mapper **out);
void tdk_mapper_destroy(tdk_mapper *m);
int tdk_mapper_reset(tdk_mapper *m, const void *image, int image_format, const double *cam,
                     const double *T_wf, const double *depth_map, const double *variance_map);
int tdk_mapper_track(tdk_mapper *m, const void *image, int image_format, const double *cam,
                     const double *T_wf, const double *T10);
int tdk_mapper_sync(tdk_mapper *m);
/* census of the current age map and pruning in one step (waits for the device); what track() does
 * one frame late when prune != 0 */
int tdk_mapper_prune(tdk_mapper *m);
int tdk_mapper_state(tdk_mapper *m, int64_t *n_frames, size_t *depth_off, size_t *variance_off,
                     size_t *age_off, size_t *flag_off, size_t *ring_off, int64_t *ring_start);

/* ------------------------------------------------------------------ helpers on the drop-in surface
 * rust_bindings.warp.warp_vecs        (src/py/warp.rs:53-67   -> src/warp.rs:31-50)
 * rust_bindings.interpolation.interpolation (src/py/interpolation.rs:6-16 -> src/interpolation.rs:57-73)
 *      status (optional device int): set to 1 if any coordinate is outside [0,W-1]x[0,H-1]
 * rust_bindings.projection.project_vecs / inv_project_vecs (src/py/projection.rs:15-43)
 * rust_bindings.transform.transform   (src/py/transform.rs:6-16 -> src/transform.rs:9-28)
 * rust_bindings.triangulation.calc_depth0 (src/py/triangulation.rs:7-19 -> src/triangulation.rs:22-36),
 *      vectorised over N point pairs
 * All take device pointers; T10 is 16 device doubles. */
int tdk_warp_vecs_dev(const double *T10, const double *xs0, const double *depths0, int64_t n,
                      double *xs1, double *depths1, void *stream);
int tdk_interpolate_dev(const double *image, int64_t H, int64_t W, const double *coords, int64_t n,
                        double *out, int32_t *status, void *stream);
int tdk_project_vecs_dev(const double *ps, int64_t n, double *xs, void *stream);
int tdk_inv_project_vecs_dev(const double *xs, const double *depths, int64_t n, double *ps,
                             void *stream);
int tdk_transform_dev(const double *T10, const double *ps0, int64_t n, double *ps1, void *stream);
int tdk_calc_depth0_dev(const double *T10, const double *xs0, const double *xs1, int64_t n,
                        double *depths0, void *stream);

/* Host-pointer forms of the same helpers: numpy in, numpy out, one call, as the pyo3 functions
 * cited above have them.  tdk_interpolate_host returns TDK_ERR_RANGE when a coordinate lies outside
 * the image (tadataka/interpolation/__init__.py:23-29 raises ValueError there). */
int tdk_warp_vecs_host(const double *T10, const double *xs0, const double *depths0, int64_t n,
                       double *xs1, double *depths1);
int tdk_interpolate_host(const double *image, int64_t H, int64_t W, const double *coords, int64_t n,
                         double *out);
int tdk_project_vecs_host(const double *ps, int64_t n, double *xs);
int tdk_inv_project_vecs_host(const double *xs, const double *depths, int64_t n, double *ps);
int tdk_transform_host(const double *T10, const double *ps0, int64_t n, double *ps1);
int tdk_calc_depth0_host(const double *T10, const double *xs0, const double *xs1, int64_t n,
                         double *depths0);

/* Synthetic-scene ray-caster (SURVEY 8d workload, 8f rank 3); no counterpart in the reference,
 * which reads its frames from datasets.  Renders N views of one piecewise-planar world, one thread
 * per pixel, without any graphics API: the device version of tadataka_b200/scene.py World.render.
 *   planes [K,14] = p0[3], normal[3], a[3], b[3], half_a, half_b (inf = unbounded)
 *   tex [K,5] = fa, fb, phase1, phase2, noise frequency     lattice [K,64,64] value-noise lattice
 *   z0: depth reported where no plane is hit     T_wc [N,16] camera-to-world poses, cams [N,4]
 *   image, depth [N,H,W] f64: intensity in [0,1] and z in the camera frame (exact ground truth) */
int tdk_render_scene_dev(const double *planes, const double *tex, const double *lattice, int64_t K,
                         double z0, const double *T_wc, const double *cams, int64_t N, int64_t H,
                         int64_t W, double *image, double *depth, void *stream);

/* Test hook with no counterpart in the reference (host arithmetic only, needs no GPU): the sieve
 * evaluates `sqrt(s) < min_gradient` (semi_dense.rs:130-134) as `s < T`; this returns that T, the
 * smallest double whose square root is not below min_gradient (0 for min_gradient <= 0, NaN for
 * NaN). */
double tdk_sqrt_threshold(double min_gradient);

/* Test hook with no counterpart in the reference.  The kernels divide several numerators by one
 * denominator with nvcc's own division sequence, sharing the part that depends on the denominator
 * only (tdk_common.cuh, tdk::div2 / tdk::div5).  This runs n pseudo-random cases (seeded; ordinary
 * magnitudes, arbitrary bit patterns, exponents next to the fast path's limits) through both and
 * writes to *mismatches (device memory) how many quotients differ from the plain `/`: must be 0. */
int tdk_selftest_shared_division_dev(int64_t n, uint64_t seed, uint64_t *mismatches, void *stream);

/* ------------------------------------------------------------------ direct pose-change estimation
 * The caller on the other side of the keyframe loop: tadataka/vo/dvo/__init__.py.
 * One Gauss-Newton iteration of _PoseChangeEstimator.__call__ (:76-110) = one
 * tdk_dvo_normal_equations_dev (calc_pose_update, :46-73, with calc_jacobian of
 * tadataka/vo/dvo/jacobian.py:8-24) + one tdk_dvo_photometric_error_dev (tadataka/metric.py:13-39);
 * the 6x6 solve, the SE(3) exponential and the loop run on the host (tadataka_b200/dvo.py) or, for
 * weight modes 0-2, on the device (tdk_dvo_iterate_dev).
 * Images, depth map, gradients, weights: device [H,W] f64.  cam0, cam1 (4 doubles each) and T10
 * (16 doubles, frame 1 <- frame 0) are HOST pointers, passed to the kernels by value.
 *
 * tdk_image_gradient_dev: np.gradient as calc_image_gradient uses it (jacobian.py:27-29) -> (DX, DY).
 *
 * tdk_dvo_normal_equations_dev: out = 30 device doubles: [0..20] upper triangle of J^T W J (row-major:
 *   00 01 .. 05 11 12 ..), [21..26] J^T W r, [27] number of pixels used, [28] sum w r^2, [29] sum r^2,
 *   over the pixels whose warped position lies inside image 1 and in front of camera 1; r = I0 - I1 at
 *   the SAME pixel, as the reference computes it (:90).  weight_mode: 0 none, 1 `weights` map
 *   (:71-72), 2 Huber with k = weight_param (robust/weights.py:38-43), 3 Student-t with
 *   variance = weight_param and nu (robust/weights.py:4-20; find the variance with
 *   tdk_dvo_student_t_dev), 4 Tukey biweight with weight_param = c * MAD of the residuals
 *   (robust/weights.py:23-35, beta = 4.6851; find the scale with tdk_dvo_tukey_scale_dev).
 * tdk_dvo_tukey_scale_dev: out = 1 device double = c_mad * median(|r - median(r)|) over the same pixels
 *   (np.median semantics: the two middle values are averaged for an even count); a one-CTA radix
 *   select over the gathered residuals.  workspace: tdk_dvo_tukey_workspace_bytes(H, W).
 * tdk_dvo_student_t_dev: out = {sum of r^2 (nu+1)/(nu + r^2/variance), count} over the same pixels:
 *   one step of the variance fixed point of compute_weights_student_t.
 * tdk_dvo_photometric_error_dev: out = {sum (I0 - I1(warp))^2, count} over the pixels whose warped
 *   position lies inside image 1 (no depth-sign test, metric.py:22). */
int tdk_image_gradient_dev(const double *image, int64_t H, int64_t W, double *gx, double *gy,
                           void *stream);
int tdk_dvo_normal_equations_dev(const double *I0, const double *D0, const double *I1,
                                 const double *GX1, const double *GY1, const double *weights,
                                 int weight_mode, double weight_param, double nu, const double *cam0,
                                 const double *cam1, const double *T10, int64_t H, int64_t W,
                                 double *out, void *stream);
int tdk_dvo_student_t_dev(const double *I0, const double *D0, const double *I1, double variance,
                          double nu, const double *cam0, const double *cam1, const double *T10,
                          int64_t H, int64_t W, double *out, void *stream);
int tdk_dvo_photometric_error_dev(const double *I0, const double *D0, const double *I1,
                                  const double *cam0, const double *cam1, const double *T10,
                                  int64_t H, int64_t W, double *out, void *stream);
/* The whole Gauss-Newton loop of _PoseChangeEstimator.__call__ (tadataka/vo/dvo/__init__.py:76-110) at
 * one resolution without a host round trip per iteration: max_iter times {normal equations, 6x6
 * Cholesky solve, candidate = exp(xi) * T10 (pose.py:44-53, se3.py:16-31), photometric error of the
 * candidate, accept unless the error rose (:100-108)}; kernels of iterations after the stop return
 * at once.  weight_mode 0, 1 or 2 (3 and 4 need a scale per iteration).  cam0, cam1, T10 (4x4 start
 * pose): host pointers.  state: 128 device doubles; when the stream has run: [0..15] the final T10
 * (row-major), [66] status (0 ok; 1 no usable pixel -- the reference's "Camera pose change is too
 * large"; 2 normal equations not positive definite -- solve on the host), [68] accepted steps,
 * [67] 1 when the loop has stopped by its own rule, [69] the last accepted error.  Every queued
 * iteration costs four launches even when the loop stopped earlier, so a caller that expects few
 * iterations queues a few, looks at [67], and calls again with resume = 1 (same inputs, same state:
 * max_iter MORE iterations, nothing re-initialised; T10 is ignored). */
int tdk_dvo_iterate_dev(const double *I0, const double *D0, const double *I1, const double *GX1,
                        const double *GY1, const double *weights, int weight_mode,
                        double weight_param, const double *cam0, const double *cam1,
                        const double *T10, int64_t H, int64_t W, int max_iter, int resume,
                        double *state, void *stream);
/* The image pyramid of PoseChangeEstimator.__call__ (tadataka/vo/dvo/__init__.py:139-148:
 * skimage.transform.rescale per level) on the device, in the form tadataka_b200/dvo.py `rescale`
 * gives it (skimage is not available here; the pyramid is unpinned, see DESIGN):
 * tdk_gaussian_blur_dev  = scipy.ndimage.gaussian_filter(mode="reflect"): two separable passes with
 *   the 2*radius+1 normalised weights the HOST array `kernel` holds (radius <= 32); tmp, out: [H,W].
 * tdk_resample_bilinear_dev = scipy.ndimage.map_coordinates(order=1, mode="reflect") on the grid
 *   ((i + 0.5) / scale - 0.5, (j + 0.5) / scale - 0.5), i < h, j < w. */
int tdk_gaussian_blur_dev(const double *image, int64_t H, int64_t W, const double *kernel, int radius,
                          double *tmp, double *out, void *stream);
int tdk_resample_bilinear_dev(const double *image, int64_t H, int64_t W, double scale, double *out,
                              int64_t h, int64_t w, void *stream);
size_t tdk_dvo_tukey_workspace_bytes(int64_t H, int64_t W);
int tdk_dvo_tukey_scale_dev(const double *I0, const double *D0, const double *I1, double c_mad,
                            const double *cam0, const double *cam1, const double *T10, int64_t H,
                            int64_t W, double *out, void *workspace, size_t workspace_bytes,
                            void *stream);

#ifdef __cplusplus
}
#endif
#endif /* TDK_H_ */
