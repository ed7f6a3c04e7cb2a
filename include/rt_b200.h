/* rt_b200.h — C ABI of librt_b200.so: the per-pixel path-tracing loop of the reference's
 * shaders/fragment.glsl as hand-written sm_100a CUDA, behind the call sites where the reference
 * talks to OpenGL.  Plain C types only; no exceptions cross this boundary.
 *
 * Every entry point replaces a group of PyOpenGL calls in the reference (paths under /root/reference):
 *
 *   rt_create / rt_destroy      GL context + shader program        main.py:115-147, main.py:444-445
 *   rt_scene_upload             3x glBufferData + glBindBufferBase  scene.py:223-236
 *   rt_scene_free               glDeleteBuffers                     scene.py:423-424
 *   rt_set_params               glUniform* block                    main.py:175-202, 332-335, 358
 *   rt_target_create / _clear   RGBA32F ping-pong textures + FBOs   screen.py:37-59, main.py:264-267
 *   rt_render_pass              bind prev, frameNumber/tileX/tileY,
 *                               bind next FBO, glDrawArrays         main.py:376-394, 401
 *   rt_render_frames            the tile/frame sequencing loop      main.py:409-430
 *   rt_read_rgba8               glBlitFramebuffer + glReadPixels    main.py:397-399, 434
 *   rt_read_accum_f32           (texture read-back of accum_tex)    screen.py:42
 *
 * Conventions: functions return RT_OK (0) or a negative rt_status; rt_last_error(ctx) gives the text
 * (the Python binding raises RuntimeError with it, standing in for PyOpenGL's GLError).  The caller owns
 * every host buffer; the library copies what it needs and owns all device memory.  A context is bound
 * to one CUDA device and is used from one host thread at a time (the reference's GL context is
 * thread-affine too, main.py:289).  Multi-GPU runs use one context per process per GPU.
 *
 * Framebuffer orientation is OpenGL's: row 0 is the bottom row (gl_FragCoord.y = 0.5).
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rt_ctx rt_ctx;

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID = -1,   /* bad argument or malformed scene buffers */
    RT_ERR_CUDA = -2,      /* a CUDA runtime call failed; see rt_last_error */
    RT_ERR_STATE = -3,     /* call out of order (no scene / no target / no params) */
    RT_ERR_NOMEM = -4
} rt_status;

/* The shader's uniform block, fragment.glsl:7-38, same names (bool lambertian as int32).
 * Vectors are 3 floats.  total_frames, fov and trisCount are carried but unused, as in the shader. */
typedef struct rt_params {
    int32_t tileX, tileY, tileSizeX, tileSizeY, numTilesX, numTilesY;
    int32_t total_frames;
    int32_t lambertian;
    float skyBrightness;
    float fov, xStep, yStep, dirStartX, dirStartY;
    float camPos[3], camRight[3], camUp[3], camForward[3];
    int32_t nBounces, rays_per_pixel;
    float jitterAmount;
    int32_t frameNumber;
    int32_t trisCount, boundingBoxCount;
} rt_params;

/* Work counters, reference-equivalent: what the shader would have executed for the same pass. */
enum {
    RT_CNT_RAYS = 0,          /* raycast() calls (fragment.glsl:246), incl. the re-traced primary of each sample */
    RT_CNT_RAYS_TRACED = 1,   /* traversals actually executed by this library (primary hit reused across samples) */
    RT_CNT_INTERNAL = 2,      /* internal nodes expanded (both children slab-tested, :281-304) */
    RT_CNT_LEAVES = 3,        /* leaves entered (:266-279) */
    RT_CNT_TRI_TESTS = 4,     /* rayTriangleIntersects calls (:274) */
    RT_CNT_HITS = 5,          /* rays whose closest hit exists */
    RT_CNT_POPS = 6,          /* stack pops incl. culled ones (:256-262) */
    RT_CNT_PATHS = 7,         /* raytrace() calls = pixels x rays_per_pixel in the tile */
    RT_CNT_COUNT = 8
};

/* Kernel variants (rt_set_variant); all produce identical results. */
enum {
    RT_VARIANT_AUTO = 0,        /* by measurement: PIXEL below 8 samples per pixel, else PERSISTENT; from 16 samples per pixel
                                   CHAIN (large share of a frame) or SAMPLES (small share) if the gate finds open sky */
    RT_VARIANT_PIXEL = 1,       /* one thread per pixel, whole pixel-frame in one kernel */
    RT_VARIANT_PERSISTENT = 2,  /* persistent warps fetching pixels with warp-aggregated atomics */
    RT_VARIANT_WAVEFRONT = 3,   /* generate / extend / shade kernels over compacted ray queues */
    RT_VARIANT_SAMPLES = 5,     /* primary stage -> the samples of each hit pixel evaluated in parallel (LCG jump-ahead,
                                   speculative start offsets) -> per-pixel chain; falls back to sequential samples on the
                                   device when most primary rays hit */
    RT_VARIANT_POOL = 4,        /* persistent CTAs, K rays per lane pooled in shared memory, extend / leaf / shade
                                   stages picked by warp vote, top of the tree staged in shared memory */
    RT_VARIANT_CHAIN = 6        /* primary stage -> ONE persistent kernel, a hit pixel per lane running its sample chain in
                                   order; lanes without a pixel (the tail of a frame, or a small share of one) evaluate a
                                   neighbour's next samples ahead of its chain and hand the results over through a small
                                   ring; the persistent kernel renders the pass instead when the device-side gate finds
                                   that most primary rays hit */
};

/* flags for rt_render_pass_ex */
enum {
    RT_PASS_DEFAULT = 0,
    RT_PASS_COUNTERS = 1,      /* accumulate the work counters (slower kernel instantiation) */
    RT_PASS_PRIMARY_IDS = 2,   /* record the primary-hit triangle id image (-1 = miss) */
    RT_PASS_SKIP_MASKED = 4,   /* pixels outside the tile are left untouched instead of copied from
                                  the previous target (multi-GPU band sharding: each GPU owns its rows) */
    RT_PASS_TO_PEER = 8        /* also store every owned pixel into the peer targets opened with rt_peer_open
                                  (peer tile mode, below) */
};

int rt_create(rt_ctx** out, int device);
void rt_destroy(rt_ctx* ctx);
const char* rt_last_error(rt_ctx* ctx);
const char* rt_version(void);

/* Run on a caller-provided CUDA stream (cudaStream_t as void*; e.g. torch's current stream) so the
 * caller's events bracket the kernels.  NULL restores the context's own stream. */
int rt_set_stream(rt_ctx* ctx, void* cuda_stream);
/* The stream the context currently launches on (cudaStream_t as void*).  rt_set_stream orders the new stream behind
 * the work already queued on the old one, so switching back and forth needs no host synchronisation. */
void* rt_get_stream(rt_ctx* ctx);
int rt_sync(rt_ctx* ctx);

/* tris: n_tris x 144 B, boxes: n_boxes x 48 B, indices: n_indices x u32 (layouts: scene.py:45-70).
 * Validates every child / offset / index, copies, and builds the device layout.  The caller's arrays are free at
 * return.  Passes already issued keep rendering the previous scene: the new one is built in a second set of buffers on
 * a stream of its own, so a re-upload overlaps the render in flight. */
int rt_scene_upload(rt_ctx* ctx, const void* tris, uint32_t n_tris, const void* boxes, uint32_t n_boxes,
                    const uint32_t* indices, uint32_t n_indices);
int rt_scene_free(rt_ctx* ctx);

int rt_set_params(rt_ctx* ctx, const rt_params* params);
int rt_set_variant(rt_ctx* ctx, int variant);
/* Kernels the last rt_render_pass launched (1 for the single-kernel variants; AUTO may pick a multi-stage one). */
int rt_last_pass_launches(rt_ctx* ctx);
/* The variant (RT_VARIANT_*) that rendered the last pass; under AUTO the choice can be made on the device, in which case
 * this call waits for the pass. */
int rt_last_pass_variant(rt_ctx* ctx);
/* Measurement aid (no published peak exists for the on-SM ceilings): runs three micro-benchmarks on the
 * context's GPU and returns out3 = { G warp-instructions/s issued, G L1 data-pipe wavefronts/s, L2->SM GB/s }. */
int rt_measure_peaks(rt_ctx* ctx, double* out3);

/* Two width x height RGBA32F accumulation targets, cleared to 0; accum_index = 0, like screen.py:61-66. */
int rt_target_create(rt_ctx* ctx, int width, int height);
int rt_target_clear(rt_ctx* ctx);
int rt_target_free(rt_ctx* ctx);

/* One draw: reads target[accum_index], writes target[1 - accum_index], then flips accum_index
 * (main.py:376-401).  frameNumber / tileX / tileY override the values given to rt_set_params. */
int rt_render_pass(rt_ctx* ctx, int frameNumber, int tileX, int tileY);
int rt_render_pass_ex(rt_ctx* ctx, int frameNumber, int tileX, int tileY, int flags);

/* n_frames whole frames starting at first_frame, every tile of each frame in the reference's order
 * (tileX fastest, main.py:409-418), frameNumber advancing by frame_stride (1 = the reference's loop;
 * G = this GPU's share of a frame-sharded progressive run, in which case the targets hold the running
 * SUM of this GPU's frames instead of the running mean — see rt_render_frames_sum). */
int rt_render_frames(rt_ctx* ctx, int first_frame, int n_frames);
int rt_render_frames_sum(rt_ctx* ctx, int first_frame, int n_frames, int frame_stride);
/* The same with flags.  RT_SUM_CONTINUE: the first pass adds to what the current target already holds (a SUM — see
 * rt_sum_resume) instead of starting a new sum, so a progressive render can be resumed (first_frame > 0, or after a
 * checkpoint was restored with rt_write_accum_f32).  Sum mode needs the tile given to rt_set_params to cover the whole
 * target (numTilesX = numTilesY = 1): a masked pixel would keep the other ping-pong target's stale value, which a
 * later reduce would add; anything else is refused with RT_ERR_INVALID. */
enum { RT_SUM_DEFAULT = 0, RT_SUM_CONTINUE = 1 };
int rt_render_frames_sum_ex(rt_ctx* ctx, int first_frame, int n_frames, int frame_stride, int flags);
/* Turns the current target from the running MEAN of prior_frames frames (what rt_render_frames / rt_finish_sum leave)
 * back into their SUM (x prior_frames, one fp32 multiply per channel; alpha untouched); prior_frames = 0 clears it. */
int rt_sum_resume(rt_ctx* ctx, int prior_frames);

/* Read-backs of the current accumulation target (the one the last pass wrote). */
int rt_read_accum_f32(rt_ctx* ctx, float* dst_rgba);        /* width*height*4 floats */
/* The same read-back without waiting for it (glReadPixels into a pixel-buffer object): the copy runs on a stream of its
 * own behind the passes issued so far and overlaps the next ones; dst_rgba (pinned memory, or the copy is not
 * asynchronous) is valid after rt_read_wait.  Writers of the copied target wait for the copy on the device; a caller
 * that writes through rt_accum_device_ptr must call rt_read_wait first. */
int rt_read_accum_f32_async(rt_ctx* ctx, float* dst_rgba);
int rt_read_wait(rt_ctx* ctx);
int rt_write_accum_f32(rt_ctx* ctx, const float* src_rgba); /* checkpoint restore into the current target */
int rt_read_rgba8(rt_ctx* ctx, uint8_t* dst_rgba);          /* round(clamp(x,0,1)*255), main.py:434 */
int rt_read_primary_ids(rt_ctx* ctx, int32_t* dst);         /* width*height, needs RT_PASS_PRIMARY_IDS */
int rt_read_counters(rt_ctx* ctx, uint64_t* dst, int n);    /* RT_CNT_* since rt_reset_counters */
int rt_reset_counters(rt_ctx* ctx);

/* Device pointers for zero-copy hand-off (torch via __cuda_array_interface__): which = 0/1 selects a
 * target, -1 the current one.  rt_accum_index returns which target is current. */
void* rt_accum_device_ptr(rt_ctx* ctx, int which);
int rt_accum_index(rt_ctx* ctx);
int rt_set_accum_index(rt_ctx* ctx, int index);

/* Band sharding helpers (multi-GPU tile mode): copy the rows owned by tile row `tileY` (rows y with
 * (y / tileSizeY) % numTilesY == tileY) of the current target to/from a dense device buffer. */
int rt_band_rows(rt_ctx* ctx, int tileY);                                   /* how many rows */
int rt_band_gather(rt_ctx* ctx, int tileY, void* dense_dev);                /* target -> dense */
int rt_band_scatter(rt_ctx* ctx, int tileY, const void* dense_dev);         /* dense -> target */

/* Peer tile mode: the exchange step of tile mode folded into the render kernels.  The assembling rank exports its two
 * targets (CUDA IPC; same node, NVLink / NVSwitch peers), every other rank maps them and renders its bands with
 * RT_PASS_TO_PEER: each finished pixel is stored into its own target AND, over NVLink, into the assembling rank's
 * (posted 16-byte stores, issued as pixels finish, so the transfer overlaps the tracing and nothing is packed,
 * gathered or unpacked afterwards).  It replaces the reference's single framebuffer that all tiles of a frame are
 * drawn into (main.py:376-401 with the tile mask of fragment.glsl:382).
 *   Contract: all ranks have targets of the same size and flip them in lockstep (same sequence of passes /
 * rt_target_clear; rt_accum_index must agree), because a pass writes target[1 - rt_accum_index] on both sides.
 * The caller fences across ranks BEFORE a peer pass (the assembling rank has finished clearing / reading the target
 * about to be written: rt_target_wait_reads, then any collective) and AFTER it (peers' kernels have completed
 * before the assembling rank reads the frame); opengl_raytracer_b200/distributed.py does both with a 4-byte
 * all-reduce.  rt_peer_close (or rt_destroy) unmaps; close the peers before the exporter re-creates its target. */
#define RT_IPC_HANDLE_BYTES 64
int rt_peer_export(rt_ctx* ctx, void* handles_out);      /* 2 * RT_IPC_HANDLE_BYTES: target 0, target 1 */
int rt_peer_open(rt_ctx* ctx, const void* handles);      /* 2 * RT_IPC_HANDLE_BYTES from the exporter  */
int rt_peer_close(rt_ctx* ctx);
int rt_peer_is_open(rt_ctx* ctx);
/* Order the stream behind any asynchronous read-back of the target the next pass will write. */
int rt_target_wait_reads(rt_ctx* ctx);

/* One draw without any allocation.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Renders the tile (tileX, tileY) of frame frameNumber of the uploaded scene with the uniforms of rt_set_params, like
 * rt_render_pass, but into buffers the CALLER passes: prev_dev / next_dev are width x height RGBA32F device buffers
 * (row 0 = bottom row), prev_dev is read when frameNumber > 0 or the tile does not cover the frame (pixels outside the
 * tile are copied from it) and may be NULL otherwise; the context's own targets and accum_index are not touched.  It
 * runs the wavefront variant (primary stage -> queue of hit pixels -> path stage) with the queue in a static device
 * array of RT_STATIC_MAX_PIXELS x 36 bytes + 256 (36 MiB, held by every GPU that loads the library until exit), so
 * neither this call nor anything it calls allocates or frees memory.  A tile that owns more than RT_STATIC_MAX_PIXELS
 * pixels, or a frame wider or higher than 65535, is refused with RT_ERR_INVALID.  All calls, from every context of the
 * process on that GPU, share the array: it keeps the last call's contents (the call clears the counters it needs at
 * zero), and calls that overlap on the GPU — two streams, two contexts — race; order them.  Results are bit-identical
 * to rt_render_pass.  The first call on a device resolves the array's address (cudaGetSymbolAddress) and loads the two
 * kernels: make one call before entering a region where even that is unwelcome (graph capture). */
#define RT_STATIC_MAX_PIXELS (1 << 20)
int rt_render_pass_static(rt_ctx* ctx, int frameNumber, int tileX, int tileY, int width, int height,
                          const void* prev_dev, void* next_dev);

/* sum target -> mean RGBA32F in place: target = target * (1 / n_frames) is NOT what the reference
 * does; it divides.  Used after the frame-sharded reduce. */
int rt_finish_sum(rt_ctx* ctx, int total_frames);

/* Stream-ordered timing on the context's stream (for callers without a CUDA event API). */
int rt_timer_start(rt_ctx* ctx);
int rt_timer_stop(rt_ctx* ctx, float* milliseconds);   /* synchronises */

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
