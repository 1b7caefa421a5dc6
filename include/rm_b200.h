/*
 * rm_b200.h — C ABI of the B200-native sphere-tracing render pass.
 *
 * This is the drop-in boundary for the reference's `src/ray_marching` render pass
 * (Mesoptier/ray-marching).  The reference has no FFI; its plugin surface is the pair
 *   RayMarchingResources::new(&RenderState)                  renderer.rs:51-175   (init)
 *   impl egui_wgpu::CallbackTrait for RayMarchingCallback    renderer.rs:195-256  (prepare, paint)
 * and every entry point below names the piece of that surface it replaces.  The byte layouts
 * taken by rm_prepare / rm_set_limits are exactly the reference's GPU buffer contents, so a Rust
 * host can pass `Uniforms{..}.as_shader_bytes()` (renderer.rs:213-222), `builder.cmd_count` and
 * `builder.buffer` (renderer.rs:230-239) unchanged.  See INTEGRATION.md for the Rust binding.
 *
 * There is exactly one implementation behind this ABI: hand-written CUDA kernels for sm_100a.
 * There is no CPU path and no fallback; if CUDA or a device is unavailable every call returns
 * RM_E_CUDA.  No call throws or aborts across the boundary.
 *
 * Threading: one rm_ctx per device; a ctx is not thread-safe but may move between threads.
 * rm_paint is asynchronous on the stream it is given; rm_download / rm_sync synchronise.
 */
#ifndef RM_B200_H
#define RM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define RM_API
#else
#define RM_API __attribute__((visibility("default")))
#endif

typedef struct rm_ctx rm_ctx;

/* ---- error codes (reference behaviour: unwrap() panics, renderer.rs:24,203,250) ---------- */
enum {
    RM_OK = 0,
    RM_E_INVALID_ARG = -1,
    RM_E_PROGRAM_TOO_LARGE = -2, /* > capacity words; reference: 1024-B buffer TODO renderer.rs:229 */
    RM_E_BAD_OPCODE = -3,        /* reference: silently returns 0.0 and desyncs, wgsl:223-225 */
    RM_E_STACK_OVERFLOW = -4,    /* value stack > 32, wgsl:173 */
    RM_E_STACK_UNBALANCED = -5,  /* pop from empty stack / truncated params / empty result */
    RM_E_CUDA = -6,
    RM_E_NOT_PREPARED = -7,
    RM_E_BAD_PARAM = -8,         /* SmoothUnion k / Scale s <= 0 or non-finite; ScalePop s differs from its push */
    RM_E_JIT = -9,               /* program specialisation failed; the interpreter kernel keeps working */
    RM_E_UNSUPPORTED = -10       /* the selected kernel does not run this program (N4 commands need RM_KERNEL_DIRECT / _SIMPLE) */
};

/* ---- byte sizes of the reference's GPU buffers --------------------------------------------- */
#define RM_UNIFORMS_BYTES 144u /* Uniforms: extent@0, inv_proj@16, inv_view@80   renderer.rs:29-34 */
#define RM_LIMITS_BYTES 12u    /* RayMarchLimits {f32 min_dist, f32 max_dist, u32 max_iter} :36-41 */
#define RM_DEFAULT_MAX_WORDS 255u /* 1024-B command buffer minus cmd_count       renderer.rs:144 */
#define RM_STACK_MAX 32u          /* value_stack_max_size                         wgsl:173 */
#define RM_POS_STACK_MAX 8u       /* nesting depth of space-transform groups (N4 extension) */
#define RM_SAMPLES_PER_PIXEL 16u  /* aa_samples^2                                 wgsl:34 */

/* ---- command opcodes: CSGCommandType, csg/builder.rs:1-24 ---------------------------------- */
enum {
    RM_OP_SPHERE = 0,         /* [0, cx,cy,cz, r]          sphere.rs:15-21, wgsl:229-233 */
    RM_OP_BOX = 1,            /* [1, cx,cy,cz, rx,ry,rz]   box.rs:14-20,   wgsl:235-240 */
    RM_OP_UNION = 100,        /* [100]   min(a,b)          wgsl:242-246 */
    RM_OP_SUBTRACTION = 101,  /* [101]   max(a,-b)         wgsl:248-252 */
    RM_OP_INTERSECTION = 102, /* [102]   max(a,b)   extension: slot reserved at builder.rs:14 */
    RM_OP_SMOOTH_UNION = 103, /* [103,k] polynomial smooth-min, extension */
    /* N4 extension: the commands the reference's enum reserves in comments, at the values the
     * #[repr(u32)] enum would give them (builder.rs:8 Plane after Box; builder.rs:16-23 TranslationPush = 200 ...).
     * A Push saves the evaluation point and replaces it, the child's commands follow ("1 child, transforms space"),
     * the matching Pop restores it; Push/Pop push no value.  Groups nest up to RM_POS_STACK_MAX deep and must
     * enclose exactly one value.  Arithmetic, one rounding per operation in exact mode (oracle/rm_oracle.c): */
    RM_OP_PLANE = 2,              /* [2, nx,ny,nz, h]    ((x*nx + y*ny) + z*nz) + h; n is expected to be unit length */
    RM_OP_TRANSLATION_PUSH = 200, /* [200, tx,ty,tz]     pos = pos - t */
    RM_OP_TRANSLATION_POP = 201,  /* [201] */
    RM_OP_ROTATION_PUSH = 202,    /* [202, m00,m01,m02, m10,m11,m12, m20,m21,m22]  row-major rotation R (object -> world):
                                   * pos = transpose(R) * pos, x' = (m00*x + m10*y) + m20*z ... */
    RM_OP_ROTATION_POP = 203,     /* [203] */
    RM_OP_SCALE_PUSH = 204,       /* [204, s]            pos = pos / s   (uniform, s finite and > 0) */
    RM_OP_SCALE_POP = 205         /* [205, s]            same s as the push; the child's distance is multiplied by s */
};

/* ---- arithmetic mode ------------------------------------------------------------------------ */
enum {
    RM_MODE_EXACT = 0, /* IEEE add/mul/div/sqrt, no FMA contraction: bit-identical to the oracle */
    RM_MODE_FAST = 1   /* FMA contraction + MUFU sqrt/rcp: within the north-star parity gates */
};

/* ---- kernel selection -------------------------------------------------------------------------- */
enum {
    RM_KERNEL_WAVEFRONT = 0, /* persistent warps over 8x4 tiles, per-slot march/normal state machine (first generation) */
    RM_KERNEL_SIMPLE = 1,    /* one thread per sample ray + resolve pass (verification path) */
    RM_KERNEL_WAVEFRONT2 = 2, /* persistent warps, march-only ray slots + per-warp hit queue for the normals */
    RM_KERNEL_DIRECT = 3      /* no ray queues: a thread marches 4 or 8 samples of its pixel in lock-step, fused in-order resolve (default) */
};

/* sample classes of the debug AOV */
enum { RM_CLS_SKY = 0, RM_CLS_FLOOR = 1, RM_CLS_OBJECT = 2 };

/* ---- lifetime: replaces RayMarchingResources::new, renderer.rs:51-175 ---------------------- */
/* Creates the per-device resources: limits (defaults 0.01 / 100.0 / 100, renderer.rs:133-137),
 * command buffer (RM_DEFAULT_MAX_WORDS words) and uniforms. */
RM_API int rm_create(int cuda_device, rm_ctx **out);
/* Same with a larger command buffer (resolves the TODO at renderer.rs:229). */
RM_API int rm_create_ex(int cuda_device, uint32_t max_words, rm_ctx **out);
/* Replaces Drop of the callback_resources type-map (main.rs:30-34). */
RM_API void rm_destroy(rm_ctx *ctx);

/* Replaces the constant RayMarchLimits upload, renderer.rs:130-140 (12 bytes). */
RM_API int rm_set_limits(rm_ctx *ctx, const void *limits12);
RM_API int rm_set_mode(rm_ctx *ctx, int mode);
RM_API int rm_get_mode(rm_ctx *ctx);
/* Programs that use the N4 commands (RM_OP_PLANE, RM_OP_*_PUSH / _POP) are painted by RM_KERNEL_DIRECT (interpreted or
 * specialised) and RM_KERNEL_SIMPLE; with one of the two wavefront kernels selected, painting such a program returns
 * RM_E_UNSUPPORTED. */
RM_API int rm_set_kernel(rm_ctx *ctx, int kernel);
/* Opt-in (default 0), RM_KERNEL_DIRECT only: skip the rest of a primitive+operator instruction when a cheap lower bound of
 * the primitive proves the accumulator unchanged.  Output is bit-identical; the work done per map_scene evaluation is not
 * the shader's literal arithmetic any more, so numbers measured with it are reported separately.
 * Ignored for programs that use the N4 commands (the bounds are written for the reference instruction set).
 * Precondition of the bit-identity: the bounds carry a fixed slack of 1e-4 for the few roundings between bound and value,
 * which covers evaluation points and primitive parameters up to a few hundred in magnitude (the reference's max_dist is
 * 100); with coordinates of 1e4 and more a float32 rounding can exceed it — leave the option off for such scenes. */
RM_API int rm_set_early_out(rm_ctx *ctx, int on);

/* ---- per frame: replaces RayMarchingCallback::prepare, renderer.rs:196-242 ------------------ */
/* Validates the program (opcodes, word count, stack discipline, transform groups), then uploads the 144-byte
 * uniform block and the command words.  Inputs are borrowed for the call only.
 * RM_E_PROGRAM_TOO_LARGE also when a program with deeply nested transform groups decodes to more instructions than the
 * context's capacity in words (the evaluation point is re-established whenever the chain of groups in force changes). */
RM_API int rm_prepare(rm_ctx *ctx, const void *uniforms144, uint32_t cmd_count, const uint32_t *words,
                      uint32_t n_words);
/* Uniform-only update (new camera pose, program unchanged) — SURVEY.md §8(f) N2/N3. */
RM_API int rm_prepare_uniforms(rm_ctx *ctx, const void *uniforms144);
/* How many times rm_prepare decoded and uploaded a program: a call whose cmd_count and words equal the last
 * uploaded ones only replaces the uniforms (resolves the per-frame re-upload of renderer.rs:224-239). */
RM_API uint64_t rm_program_uploads(rm_ctx *ctx);
/* Static validation only (no device work): same codes as rm_prepare. */
RM_API int rm_validate_program(uint32_t cmd_count, const uint32_t *words, uint32_t n_words,
                               uint32_t max_words);

/* Opt-in, never the headline: per-tile conservative culling of the program (csrc/rm_cull.cu).  Before the paint, one
 * thread per 8-wide block tile walks the tile's centre ray over every ray parameter a ray of the frame can reach and
 * keeps a leaf instruction `acc = op(acc, prim(p))` only if, somewhere in the tile's cone, it could change the
 * accumulator (1-Lipschitz bounds on the margins prim - acc - k, prim + acc, acc - prim); each block then interprets
 * only its tile's instructions.  A dropped instruction is a no-op bit for bit at every point the tile's rays visit, so
 * image, AOVs and work counters are IDENTICAL to the unculled paint in both arithmetic modes — only the work per
 * evaluation is no longer the shader's literal arithmetic, which is why this is reported separately.  Applies to
 * RM_KERNEL_DIRECT paints of reference-set programs of 8 to 64 decoded instructions (others paint as usual); paints
 * through the interpreter (specialised kernels are not used while it is on).
 * rm_get_cull_stats: algorithmic flops of the map_scene evaluations as EXECUTED (sum over evaluations of the flops of
 * the instructions the tile kept) by the culled paints made with rm_set_counters on, and the number of culled paints. */
RM_API int rm_set_cull(rm_ctx *ctx, int on);
RM_API int rm_get_cull_stats(rm_ctx *ctx, uint64_t *executed_map_flops, uint64_t *culled_paints, int reset);
/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The culled paint of rm_set_cull with its
 * scratch — one 64-bit mask per pixel tile — in a file-scope static device array instead of an allocated buffer: the call
 * allocates NOTHING (no cudaMalloc / cudaFree, no allocating library routine).  Paints rows [row0, row0+nrows) of the
 * prepared program into the CALLER's dst_rgba8_dev (required, 16-byte aligned, pitch >= 4*W); the frame bytes equal
 * rm_paint's.  Limits: RM_KERNEL_DIRECT, reference-set programs of 8 to 64 decoded instructions (else RM_E_UNSUPPORTED);
 * at most RM_CULL_STATIC_MAX_TILES = 262144 tiles of 8 x 8 pixels per call (one 3840x2160 frame; larger frames in row
 * bands), beyond that RM_E_INVALID_ARG with the numbers in rm_last_error.  The array (2 MiB) is held by every GPU that loads
 * the library until exit, is shared by every call and context on that GPU and keeps the last call's contents (the call
 * rewrites every mask it reads): calls that OVERLAP on one GPU — two streams, two contexts — race; serialise them.  The
 * first call per device resolves the array's device address (cudaGetSymbolAddress): make one call outside a capture. */
#define RM_CULL_STATIC_MAX_TILES 262144u
RM_API int rm_paint_culled_static(rm_ctx *ctx, uint32_t W, uint32_t H, uint32_t row0, uint32_t nrows, void *dst_rgba8_dev,
                                  size_t pitch_bytes, void *cuda_stream);

/* ---- program specialisation (optional; the reference compiles its shader once, RayMarchingResources::new
 *      renderer.rs:55-62 — here the SCENE can be compiled too) --------------------------------------------------- */
/* By default the kernels interpret the decoded command stream: any program, no set-up cost.  A program can instead be
 * compiled: map_scene is emitted as straight-line CUDA — no fetch, no dispatch — and built at run time (NVRTC, sm_100a)
 * around the same kernel body.  The generated code depends on the program's STRUCTURE only: the opcode sequence, which
 * parameters are exactly zero (those stay literals so that `p - 0` folds) and, in exact mode, which smooth-union k are
 * powers of two.  Every other centre, radius, k, offset or matrix entry travels as a kernel argument (constant bank,
 * read through uniform registers) — or, for deep programs, whose scene code is one out-of-line function, in a constant
 * block of the generated module that the paint refills on its stream when the values change (one block per device,
 * shared by the contexts painting that structure: their launches of it are ordered).  So the reference's per-frame re-flatten and
 * re-upload (renderer.rs:224-239) and an editor dragging a value keep their compiled kernel: rm_prepare with new values
 * of the same structure costs no compile.  Arithmetic per ray is operation for operation the interpreter's: exact mode
 * stays bit-identical, fast mode may round differently where the compiler contracts multiply-adds (within the parity
 * gates).  Compiling takes 0.2 - 1 s per (structure, mode) and is cached process-wide — shared by the contexts of a
 * group, bounded LRU (512 structures by default, rm_spec_cache): the least recently used kernel no context is painting
 * with is unloaded.  Only RM_KERNEL_DIRECT paints are affected; with rm_set_early_out on, the interpreter is used.
 * rm_specialize: compile the currently prepared program for the current mode now; paints use it until rm_prepare
 *   receives words of a different structure.  RM_E_JIT (reason in rm_last_error) when NVRTC cannot be loaded
 *   (libnvrtc.so.12; the environment variable RM_NVRTC_PATH names an exact file), the compile fails or the program is
 *   too long for the instruction cache — painting then simply continues through the interpreter.
 * rm_set_specialize(1): do that automatically for every program at its first paint (a stall of one compile per new
 *   STRUCTURE: meant for batch rendering); failures are silent, see rm_spec_status.
 *   rm_set_specialize(2): the same without the stall — a kernel already in the cache is adopted at once, a new
 *   structure is compiled by a background thread of the context while paints keep going through the interpreter, and
 *   the first paint after the compile has finished switches over: the setting for an interactive editor.
 *   rm_set_specialize(0) returns to the interpreter from the next paint on (also after rm_specialize). */
RM_API int rm_specialize(rm_ctx *ctx);
RM_API int rm_set_specialize(rm_ctx *ctx, int on);
/* Second tier.  On sm_100 an arithmetic instruction takes an immediate for free but a kernel-argument value only through
 * a uniform register (LDCU, one issue slot per 1-4 values): the deep scene's march loop is ~5 % longer with arguments.
 * Words that STAY PUT are therefore compiled once more with every parameter an immediate (keyed by the whole word
 * stream): after n_paints plain paints of unchanged words the context looks that kernel up, else compiles it on its
 * background thread — never a stall, whatever the policy — and switches over when it is ready; rm_specialize compiles
 * it at once.  New values of the same structure drop back to the structure kernel without a compile.
 * n_paints: default 8; 0 = structure tier only.  rm_spec_tier: what the next plain paint runs — 0 the interpreter,
 * 1 the structure kernel, 2 the literal kernel. */
RM_API int rm_set_spec_literal_after(rm_ctx *ctx, uint32_t n_paints);
RM_API int rm_spec_tier(rm_ctx *ctx);
/* Whether the next RM_KERNEL_DIRECT paint (current mode, counters setting) runs specialised code, and what it is:
 * sample rays per thread (2, 4, 8: chosen from the code size), resident blocks asked of the compiler, compile + load
 * time, cubin size; number of compiles this context triggered and of specialised launches.  Any pointer may be NULL. */
RM_API int rm_spec_info(rm_ctx *ctx, int *active, int *rays_per_thread, int *min_blocks, double *compile_ms,
                        uint64_t *cubin_bytes, uint64_t *compiles, uint64_t *launches);
RM_API const char *rm_spec_status(rm_ctx *ctx);
/* No device needed: the generated CUDA source of a command stream (rays_per_thread 0 = automatic), and a trial
 * compile of it (NVRTC cross-compiles for sm_100a without a GPU).  mode | RM_SPEC_LITERAL_TIER: the literal tier's. */
#define RM_SPEC_LITERAL_TIER 0x100
RM_API int rm_spec_source(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, int mode, int rays_per_thread,
                          char *buf, size_t cap, size_t *needed);
RM_API int rm_spec_compile_check(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, int mode,
                                 int rays_per_thread, uint64_t *cubin_bytes, double *compile_ms, char *log,
                                 size_t log_cap);
/* The parameter block the generated kernel of these words reads (its SpecParams argument): *needed floats. */
RM_API int rm_spec_params(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, int mode, float *buf,
                          size_t cap_floats, size_t *needed);
/* Process-wide kernel cache: new_capacity > 0 sets the LRU bound (evicting down to it); reports entries, bound and
 * evictions so far.  Any pointer may be NULL. */
RM_API int rm_spec_cache(uint64_t new_capacity, uint64_t *entries, uint64_t *capacity, uint64_t *evictions);

/* ---- per frame: replaces RayMarchingCallback::paint, renderer.rs:244-255 -------------------- */
/* Shades rows [row0,row0+nrows) of a W×H frame (row 0 = top) as RGBA8 (R,G,B,A bytes; A=255).
 * dst_rgba8_dev: device pointer receiving the band, row `row0` at offset 0, `pitch_bytes`
 * between rows (>= 4*W, multiple of 16); NULL = ctx-owned buffer (read back with rm_download).
 * cuda_stream: a cudaStream_t, or NULL for the ctx's own stream.  Asynchronous. */
RM_API int rm_paint(rm_ctx *ctx, uint32_t W, uint32_t H, uint32_t row0, uint32_t nrows,
                    void *dst_rgba8_dev, size_t pitch_bytes, void *cuda_stream);
/* Interleaved strips (multi-GPU load balance): shades strips first_strip,
 * first_strip + strip_step, ... of `strip_rows` rows each (the last strip of the frame may be short)
 * and packs them contiguously into dst.  *rows_painted receives the number of rows written.
 * Every pixel is bit-identical to the same pixel of a full-frame rm_paint.  Any strip height works (a block tile may
 * straddle two strips); pick one that gives every shard the same NUMBER of strips, or the step waits for the shards with
 * one more: 4320 rows over 8 shards are 34 / 33 strips of 16 rows (3 % apart, measured) and 45 each of 12 rows. */
RM_API int rm_paint_strips(rm_ctx *ctx, uint32_t W, uint32_t H, uint32_t strip_rows, uint32_t first_strip,
                           uint32_t strip_step, void *dst_rgba8_dev, size_t pitch_bytes, void *cuda_stream,
                           uint32_t *rows_painted);
/* Strips of this shard stored straight into a WHOLE-frame buffer at their frame rows — typically GPU 0's frame, opened
 * in this process with rm_shared_frame_open, so the render kernel's 128-bit stores travel over NVLink and the frame
 * needs no gather and no re-order (config 4; SURVEY.md 8(e) "direct P2P stores into GPU0's framebuffer").  Pixels are
 * bit-identical to a full-frame rm_paint.  Completion across processes is the caller's barrier (e.g. an all-reduce
 * enqueued after this call on every rank). */
RM_API int rm_paint_strips_into(rm_ctx *ctx, uint32_t W, uint32_t H, uint32_t strip_rows, uint32_t first_strip,
                                uint32_t strip_step, void *frame_rgba8_dev, size_t pitch_bytes, void *cuda_stream,
                                uint32_t *rows_painted);
/* A device allocation other processes (one per GPU) can map: create on the owner, send the 64 handle bytes to the
 * peers by any channel, open there.  close releases a mapping (peer) or the allocation (owner); rm_destroy closes all. */
#define RM_IPC_HANDLE_BYTES 64u
RM_API int rm_shared_frame_create(rm_ctx *ctx, size_t bytes, void **dev_ptr, void *handle64);
RM_API int rm_shared_frame_open(rm_ctx *ctx, const void *handle64, void **dev_ptr);
RM_API int rm_shared_frame_close(rm_ctx *ctx, void *dev_ptr);
/* Device -> host copy of any device pointer this process can address; synchronises the stream. */
RM_API int rm_copy_to_host(rm_ctx *ctx, const void *dev_ptr, void *host, size_t bytes, void *cuda_stream);
/* Config 4 end to end: copies the strips a shard painted PACKED (rm_paint_strips with the same plan; band_dev NULL = the
 * ctx-owned buffer) to their rows of a whole W x H frame in HOST memory (pitch 4*W) — one strided cudaMemcpy2DAsync
 * (plus one for a partial last strip) on the stream, then synchronises it.  With one process per GPU and host_frame a
 * buffer every process has page-locked (rm_host_register of one shared mapping) the frame leaves over all the GPUs' PCIe
 * links at once instead of through GPU 0's. */
RM_API int rm_copy_strips_to_host(rm_ctx *ctx, const void *band_dev, uint32_t W, uint32_t H, uint32_t strip_rows,
                                  uint32_t first_strip, uint32_t strip_step, void *host_frame, void *cuda_stream);
/* Pose batch (config 5): paints n_poses frames from an array of 144-byte uniform blocks held in
 * HOST memory; frame p goes to dst + p*frame_stride_bytes (device memory), or, when dst is NULL,
 * every frame overwrites the ctx-owned buffer (throughput mode). */
RM_API int rm_paint_poses(rm_ctx *ctx, const void *uniforms144_array, uint32_t n_poses, uint32_t W,
                          uint32_t H, void *dst_rgba8_dev, size_t pitch_bytes,
                          size_t frame_stride_bytes, void *cuda_stream);
/* Pose batch with the frames returned to HOST memory (the reference hands every frame to its surface,
 * renderer.rs:244-255): frame p is rendered on the ctx stream into a 3-deep device ring and copied to
 * host_rgba8 + (p % host_ring_frames) * host_frame_stride_bytes on a second stream, so the copy of frame p
 * overlaps the render of frame p+1 (pinned host memory, e.g. rm_host_alloc, is needed for the overlap).
 * host_ring_frames == 0 means one host slot per pose.  Synchronises: all frames are in host memory on return. */
RM_API int rm_paint_poses_host(rm_ctx *ctx, const void *uniforms144_array, uint32_t n_poses, uint32_t W,
                               uint32_t H, void *host_rgba8, size_t host_frame_stride_bytes,
                               uint32_t host_ring_frames);
/* Copies the ctx-owned framebuffer of the last rm_paint(dst=NULL) to host memory; synchronises. */
RM_API int rm_download(rm_ctx *ctx, void *host_rgba8, size_t bytes);
RM_API int rm_sync(rm_ctx *ctx);

/* ---- one host thread, several GPUs (SURVEY.md 8(b) "Multi-GPU (new surface)", 8(e)) ------------------------------ */
/* The reference is one process with one render thread (main.rs:44-82, renderer.rs:244-255); this is the surface that
 * thread would drive when the box has several GPUs.  A group is one rm_ctx per listed device (the <= 1.2 KB of inputs are
 * replicated) plus a frame on the FIRST device that every member addresses over NVLink (peer access; RM_E_CUDA when a
 * member cannot reach the first device).  Listing a device more than once is allowed.  The one-process-per-GPU form of the
 * same thing is rm_paint_strips_into + rm_shared_frame_* and rm_paint_poses_host above. */
typedef struct rm_group rm_group;
RM_API int rm_group_create(const int *cuda_devices, int n, rm_group **out);
RM_API void rm_group_destroy(rm_group *g);
RM_API int rm_group_size(rm_group *g);
/* Member i's context, e.g. for rm_set_kernel / rm_last_error / rm_launch_count; owned by the group. */
RM_API rm_ctx *rm_group_ctx(rm_group *g, int i);
RM_API int rm_group_set_mode(rm_group *g, int mode);
RM_API int rm_group_set_limits(rm_group *g, const void *limits12);
/* rm_specialize on every member (one compile: the cache is process-wide). */
RM_API int rm_group_specialize(rm_group *g);
/* rm_prepare on every member. */
RM_API int rm_group_prepare(rm_group *g, const void *uniforms144, uint32_t cmd_count, const uint32_t *words,
                            uint32_t n_words);
/* Config 4: ONE W x H frame; member i shades strips i, i+n, ... of strip_rows rows and stores them straight into the
 * group frame on the first device.  Returns when the frame is complete there (rm_group_frame); when host_rgba8 is not
 * NULL the frame is also copied to it (W*4*H bytes, tightly packed). */
RM_API int rm_group_paint_bands(rm_group *g, uint32_t W, uint32_t H, uint32_t strip_rows, void *host_rgba8);
/* Device pointer (first device) and pitch of the frame of the last rm_group_paint_bands. */
RM_API int rm_group_frame(rm_group *g, void **dev_ptr, size_t *pitch_bytes);
/* Config 5: pose p is rendered by member p % n and returned at host_rgba8 + p*host_frame_stride_bytes (pinned memory
 * lets the copies overlap the renders); members run concurrently; returns when every frame is in host memory. */
RM_API int rm_group_paint_poses(rm_group *g, const void *uniforms144_array, uint32_t n_poses, uint32_t W, uint32_t H,
                                void *host_rgba8, size_t host_frame_stride_bytes);
RM_API const char *rm_group_last_error(rm_group *g);

/* ---- debug / verification surface (untimed) -------------------------------------------------- */
/* Straightforward one-thread-per-sample kernels that also emit the AOVs the parity gates need.
 * All pointers are HOST memory and may be NULL.  cls: nrows*W*16 bytes (sample i*4+j);
 * t: nrows*W*16 floats (marched distance for object hits, floor_dist otherwise);
 * counters: {N_step, N_hit, N_floor, N_sample}; rgba: nrows*W*4 bytes. */
RM_API int rm_paint_debug(rm_ctx *ctx, uint32_t W, uint32_t H, uint32_t row0, uint32_t nrows,
                          uint8_t *cls_host, float *t_host, uint64_t *counters4, uint8_t *rgba_host);
/* The same AOVs (and, optionally, the image) from the PRODUCTION kernel's own march — RM_KERNEL_DIRECT's body, run
 * specialised when rm_specialize / rm_set_specialize apply to the prepared program (*specialised = 1), else around the
 * interpreter: lets the parity gates be checked on the kernel that is benchmarked, not on a stand-in. */
RM_API int rm_paint_aov(rm_ctx *ctx, uint32_t W, uint32_t H, uint32_t row0, uint32_t nrows, uint8_t *cls_host,
                        float *t_host, uint8_t *rgba_host, int *specialised);
/* Evaluates the prepared program's map_scene at n host points (xyz interleaved). */
RM_API int rm_map(rm_ctx *ctx, const float *pts_host, uint64_t n, float *out_host);
/* Work counters {N_step, N_hit, N_floor, N_sample} accumulated by rm_paint when enabled with
 * rm_set_counters(ctx, 1) (adds atomics; keep off when timing). */
RM_API int rm_set_counters(rm_ctx *ctx, int enable);
RM_API int rm_get_counters(rm_ctx *ctx, uint64_t *counters4, int reset);

/* ---- measurement helpers --------------------------------------------------------------------- */
/* Records CUDA event `slot` (0..15) on the ctx stream / returns elapsed ms between two slots. */
RM_API int rm_mark(rm_ctx *ctx, int slot);
RM_API int rm_elapsed_ms(rm_ctx *ctx, int slot_a, int slot_b, float *ms);
/* Writes a 256-MiB scratch buffer (> 126 MB L2) on the ctx stream. */
RM_API int rm_flush_l2(rm_ctx *ctx);
/* FFMA micro-benchmark: measured FP32 peak in TFLOP/s (2 flop per FFMA) and the SM count. */
RM_API int rm_fp32_peak(rm_ctx *ctx, double *tflops, int *sm_count, int *sm_clock_khz);
/* Number of kernels launched by this ctx since creation (for bench.py's gpu_launches). */
RM_API uint64_t rm_launch_count(rm_ctx *ctx);
/* Pinned host memory for rm_download / e2e runs. */
RM_API void *rm_host_alloc(size_t bytes);
RM_API void rm_host_free(void *p);
/* Page-locks (cudaHostRegister, portable) / releases host memory the caller owns, e.g. a POSIX shared-memory mapping
 * that several processes copy their strips into (rm_copy_strips_to_host).  0 or RM_E_CUDA. */
RM_API int rm_host_register(void *p, size_t bytes);
RM_API int rm_host_unregister(void *p);

RM_API const char *rm_last_error(rm_ctx *ctx);
RM_API const char *rm_strerror(int code);
RM_API const char *rm_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RM_B200_H */
