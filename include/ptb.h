/* ptb.h — C ABI of the B200-native render backend (libptb.so).
 *
 * The reference (CouncilmanJeremyJamm/path_tracer) has no FFI or plugin interface: its seam is the Rust
 * closure body src/main.rs:186-207 and the types that feed it.  This header is what a `ffi.rs` for that
 * seam would bind: plain pointers and sizes, no C++/torch types, `int` status codes,
 * never unwinds.  Each entry point cites the reference interface it replaces.
 *
 * Ownership: the caller owns every host array (borrowed for the duration of the call; the library copies to
 * the device).  The library owns all device memory and frees it in ptb_destroy.  Output host buffers are
 * caller-allocated.  Image arrays are row-major with row 0 = bottom row (camera.rs:96, shader.wgsl:44-49).
 *
 * Threading: one context per GPU; calls on one context must be externally serialised (the reference calls
 * the seam once per frame from the main thread, main.rs:179-219); different contexts may be driven from
 * different threads.  Kernels run on a per-context stream; ptb_render BLOCKS until the frame batch is done (its loop
 * polls the device for the number of live paths); ptb_render_async runs it on a worker thread, ptb_wait joins.
 *
 * There is no CPU fallback: every compute entry point returns PTB_ERR_NO_DEVICE without a CUDA device.
 */
#ifndef PTB_H
#define PTB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTB_ABI_VERSION 1

typedef struct ptb_context ptb_context;

enum ptb_status
{
    PTB_OK = 0,
    PTB_ERR_INVALID = 1,     /* bad argument / scene the reference would reject (panics upstream, e.g. model.rs:43, tlas_bvh.rs:58) */
    PTB_ERR_CUDA = 2,        /* CUDA runtime error, see ptb_last_error */
    PTB_ERR_NO_DEVICE = 3,   /* no CUDA device / wrong architecture */
    PTB_ERR_OOM = 4,
    PTB_ERR_NOT_BUILT = 5,   /* scene not uploaded / ptb_build not called */
    PTB_ERR_UNSUPPORTED = 6, /* valid upstream but beyond a documented limit of this backend (e.g. > 64 media) */
    PTB_ERR_NCCL = 7         /* NCCL error in a ptb_group_* call, see ptb_group_last_error */
};

/* Material kinds: enum Material, material.rs:80-89 (GGX split by GGXModel, material.rs:171-180). */
enum ptb_material_kind
{
    PTB_LAMBERTIAN = 0,     /* Lambertian::new(albedo)                      material.rs:99  */
    PTB_EMISSIVE = 1,       /* Emissive::new(emitted)                       material.rs:126 */
    PTB_SPECULAR = 2,       /* Specular::new(colour)                        material.rs:146 */
    PTB_GGX_METAL = 3,      /* GGX::new_metal(colour, roughness)            material.rs:290 */
    PTB_GGX_DIELECTRIC = 4, /* GGX::new_dielectric(colour, roughness, ior, volume) material.rs:305 */
    PTB_DIELECTRIC = 5      /* Dielectric::new(colour, ior, volume)         material.rs:475 */
};

typedef struct ptb_material
{
    uint32_t kind;
    float colour[3];         /* albedo | emitted | colour */
    float roughness;         /* linear roughness; the library applies powi(2).clamp(1e-4, 0.9999) as material.rs:294 */
    float ior;
    uint32_t has_volume;     /* Option<Volume>; honoured for the two transmissive kinds only */
    float vol_absorption[3]; /* Volume::new(absorption, k, c, g), volume.rs:136 */
    float vol_k, vol_c, vol_g;
} ptb_material;

/* One `Model` (model.rs:27-33) = one BLAS: a triangle range of the vertex array + one material (blas.rs:164-170). */
typedef struct ptb_blas_desc
{
    uint32_t first_tri, tri_count;
    ptb_material material;
} ptb_blas_desc;

/* One entry of `Model.matrices`: glam Affine3A column-major (x_axis, y_axis, z_axis, translation).
 * Instances must be grouped in BLAS order (the order `for info in blas_info { for matrix in ... }` of tlas_bvh.rs:90-104). */
typedef struct ptb_instance
{
    uint32_t blas;
    float affine[12];
} ptb_instance;

/* ---- context ------------------------------------------------------------------------------------------ */
int ptb_abi_version(void);
int ptb_create(int device_ordinal, ptb_context** out);
void ptb_destroy(ptb_context* ctx);
/* Message for the last non-OK status on this context (ctx == NULL: last failure of ptb_create on this thread).
 * Valid until the next call on the context. */
const char* ptb_last_error(const ptb_context* ctx);
int ptb_synchronize(ptb_context* ctx);

/* ---- Scene::new(Vec<Model>)  scene.rs:21-35 --------------------------------------------------------------
 * verts: ntri * 3 * 6 floats, per vertex (position xyz, normal xyz) = `Vertex` (model.rs:13-18), three per
 * triangle in the order load_obj emits them (blas.rs:97-119).  Validates what the reference asserts or panics
 * on: non-empty BLASes, unit-scale matrices (model.rs:42-43), at least one emissive model (scene.rs:26 ->
 * tlas_bvh.rs:58).  ptb_build runs BLAS::new / TLAS::new (bit-exact node order, blas_bvh.rs:62-136,
 * tlas_bvh.rs:85-138), LightSampler::new (light_sampler.rs:41-61), derives the 8-wide device layout and uploads. */
int ptb_scene_upload(ptb_context* ctx, const float* verts, uint64_t ntri, const ptb_blas_desc* blas, uint32_t nblas, const ptb_instance* inst, uint32_t ninst);
/* Optional environment map, linear RGB, row-major w*h*3 (what ImageHelper::load_image yields, image_helper.rs:25-33);
 * without it the background is the constant 0.006 of integrator.rs:265. */
int ptb_scene_set_env(ptb_context* ctx, uint32_t w, uint32_t h, const float* rgb);
int ptb_build(ptb_context* ctx);
/* Host threads used by ptb_build (0 = hardware_concurrency). */
int ptb_set_build_threads(ptb_context* ctx, int threads);

/* ---- build output in reference arena order — for bit-exact comparison ------------------
 * BLAS nodes (blas_bvh.rs:36-58): boxes 6 floats/node (min xyz, max xyz); meta 3 u32/node:
 *   kind 0 Branch {left,right} | 1 LeafSingle {prim,0} | 2 Leaf {first index into leaf_prims, count}.
 * triangles: 30 floats each: positions[9], normals[9], n0[4], n1[4], n2[4] (primitive.rs:14-27). */
int ptb_bvh_blas_counts(ptb_context* ctx, uint32_t blas, uint32_t* node_count, uint32_t* leaf_prim_count, uint32_t* root);
int ptb_bvh_blas_nodes(ptb_context* ctx, uint32_t blas, float* boxes6, uint32_t* meta3, uint32_t* leaf_prims);
int ptb_bvh_blas_triangles(ptb_context* ctx, uint32_t blas, float* out30);
/* TLAS nodes (tlas_bvh.rs:32-52), which = 0 world (scene.rs:25) | 1 lights (scene.rs:26): meta kind 0 Branch {left,right} |
 * 1 Leaf {TLAS-local blas index, instance ordinal}; matrices 24 floats/node (matrix, inv_matrix; zero for branches);
 * blas_ids: TLAS-local blas index -> scene blas index. */
int ptb_bvh_tlas_counts(ptb_context* ctx, int which, uint32_t* node_count, uint32_t* root, uint32_t* nblas);
int ptb_bvh_tlas_nodes(ptb_context* ctx, int which, float* boxes6, uint32_t* meta3, float* matrices24, uint32_t* blas_ids);
/* LightSampler (light_sampler.rs:23-28): items 2 u32/light (lights-TLAS blas index, prim id), pdf, cdf, max. */
int ptb_light_count(ptb_context* ctx, uint32_t* count);
int ptb_light_sampler(ptb_context* ctx, uint32_t* items2, float* pdf, float* cdf, float* max);
/* Wide-layout statistics: [0] wide nodes, [1] leaf headers, [2] triangle slots, [3] bytes of device BVH data,
 * [4] max wide depth (TLAS+BLAS), [5] binary nodes, [6] instances, [7] device stack bound. */
int ptb_bvh_stats(ptb_context* ctx, uint64_t out[8]);

/* ---- fixed ray batches: TLAS::intersect (tlas.rs:66-110) / TLAS::any_intersect (tlas.rs:111-144) ---------------
 * which: 0 = scene.world, 1 = scene.lights.  kernel: 0 = 8-wide persistent kernel (the renderer's), 1 = reference-order
 * binary kernel (walks the reference's own binary BVH in the reference's visit order; bit-exact incl. ties), 2 = team kernel
 * (eight lanes per ray over the 8-wide layout, csrc/team.cuh; same results as kernel 0 up to the resolution of exact ties),
 * 3 / 4 / 5 = multiray kernels (two / three / four rays per lane with gated stages and the ray records in shared memory,
 * csrc/multiray.cuh; 6 = the same machinery with one ray per lane): bit-identical to kernel 0.
 * o3/d3: n*3 floats (d normalised, ray.rs:12), tmax: n floats.  Host pointers; copies happen inside the call.
 * Outputs may be NULL except `hit`.  blas = scene blas index, inst = instance ordinal (`Model`x`matrices` order),
 * normal = world-space ray-facing shading normal (primitive.rs:161-165, tlas.rs:103-105). */
int ptb_trace_closest(ptb_context* ctx, int which, int kernel, uint64_t n, const float* o3, const float* d3, const float* tmax, uint8_t* hit, float* t, float* u,
                      float* v, uint32_t* prim, uint32_t* blas, uint32_t* inst, uint8_t* front, float* normal3);
int ptb_trace_any(ptb_context* ctx, int which, int kernel, uint64_t n, const float* o3, const float* d3, const float* tmax, uint8_t* hit);
/* Same kernels on rays already resident in device memory (benchmarks): upload once, launch many times.
 * ptb_trace_resident runs `repeats` launches and returns the mean device time per launch (CUDA events on the context stream). */
int ptb_trace_upload(ptb_context* ctx, uint64_t n, const float* o3, const float* d3, const float* tmax);
int ptb_trace_resident(ptb_context* ctx, int which, int kernel, int any_hit, int repeats, float* mean_ms);
int ptb_trace_download(ptb_context* ctx, uint8_t* hit, float* t, uint32_t* prim, uint32_t* inst);

/* ---- Camera::new / Camera::create_ray (camera.rs:17-31, 94-105); SobolSampler::get_ss_sobol (sampling.rs:97-114) ------ */
int ptb_set_camera(ptb_context* ctx, const float origin[3], const float target[3], float fov_deg, float aspect_ratio);
int ptb_camera_rays(ptb_context* ctx, uint64_t n, const float* st2, float* o3, float* d3);           /* runs the raygen device code */
int ptb_ss_sobol(ptb_context* ctx, uint64_t n, const uint32_t* index, const uint32_t* seed, float* xy2); /* runs the raygen device code */
int ptb_rng_stream(ptb_context* ctx, uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t n, uint32_t* out); /* device RNG probe */

/* ---- the frame loop main.rs:181-207 + accumulate.wgsl:20-23 ------------------------------------------------------
 * ptb_render traces samples [sample_begin, sample_begin+sample_count) of every pixel (one "frame" upstream = one sample per
 * pixel) with integrate(ray, scene, env, rng, max_bounces) (integrator.rs:143-149) and adds (rgb, 1) per sample into the
 * device accumulation buffer (created/zeroed by ptb_frame_begin; W*H float4).  RNG streams are keyed by (seed, pixel,
 * global sample index) so disjoint sample ranges on different GPUs compose.  enable_nee = ENABLE_NEE (main.rs:51).
 * ptb_read_accum: sum of samples in rgb, sample count in w — the accumulation texture of state.rs:583; display = rgb / w
 * (shader.wgsl:63).  ptb_read_aux: `position` (first-hit point, t) and `id` ((id << 16) | first blas id) of main.rs:204-206. */
int ptb_frame_begin(ptb_context* ctx, uint32_t width, uint32_t height);
int ptb_render(ptb_context* ctx, uint32_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_bounces, int enable_nee);
/* ptb_render on a worker thread of the context: returns at once; ptb_wait (or ptb_synchronize) joins it and returns its status.
 * No other call on this context in between.  Meant for hosts that render frame after frame (main.rs:179-219) with TWO contexts on
 * one device, alternating: the thin end of one frame's wavefront (latency-bound, device mostly idle) overlaps the dense start of
 * the next frame on the other context. */
int ptb_render_async(ptb_context* ctx, uint32_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_bounces, int enable_nee);
int ptb_wait(ptb_context* ctx);
int ptb_read_accum(ptb_context* ctx, float* rgba);
int ptb_read_aux(ptb_context* ctx, float* position4, uint32_t* id);
/* Display image of the accumulation buffer, the step after the path: rgb / w (shader.wgsl:63), Uchimura GT tonemap with the
 * constants of ImageHelper::write_image (image_helper.rs:44; tonemapping.rs:68-96), gamma 1/2.2, * 255 `as u8`.  rgb: W*H*3 bytes.
 * flip_y != 0 writes the top image row first (the accumulation buffer has row 0 at the bottom, camera.rs:96). */
int ptb_read_rgb8(ptb_context* ctx, uint8_t* rgb, int flip_y);
/* ---- temporal accumulation for a moving camera: State::update (state.rs:505-579) as CUDA kernels --------------------------
 * The reference keeps an accumulation texture inside `State` and, after every frame, dispatches either accumulate.wgsl
 * (camera unchanged: accumulation += (rgb, 1)) or velocity.wgsl + compute.wgsl (camera moved: reproject last frame through
 * `last_inv_projection`, validate by model id, Catmull-Rom history fetch, variance clip in YCoCg, blend 15 % of the new frame),
 * then copies the output into the accumulation texture.  Here the accumulation texture is a device buffer owned by the context.
 *   ptb_state_update           = State::update with its own argument list: host slices `data` (rgb, 1), `position` (xyz, t),
 *                                `id` ((previous << 16) | current, main.rs:206), W*H elements each, row 0 at the bottom.
 *   ptb_state_update_resident  = the same on the context's current frame buffers (no host round trip): the accumulation
 *                                buffer of ptb_frame_begin/ptb_render is the frame input (read as sum / count), position and
 *                                id are the aux buffers.  Per displayed frame: ptb_frame_next, ptb_render(1 sample), this call.
 *   `inv_projection` / `last_inv_projection` are what the reference passes under those names: (cam.matrix *
 *   cam.inv_projection).inverse() of this and of the previous frame (main.rs:209-216), column-major;
 *   ptb_camera_view_projection returns it for the camera set by ptb_set_camera.
 *   ptb_state_reset            = a new State (zeroed textures) of the given size.
 *   ptb_state_read             = accumulation texture (rgba, W*H*4 floats; display value = rgb / w, shader.wgsl:63) and, optionally,
 *                                the velocity texture written by the last moving-camera update (W*H*2 floats).
 *   ptb_frame_next             = start the next frame of the same size keeping the id history (ptb_frame_begin clears it). */
int ptb_state_reset(ptb_context* ctx, uint32_t width, uint32_t height);
int ptb_state_update(ptb_context* ctx, uint32_t width, uint32_t height, const float* data, const float* position, const uint32_t* id,
                     const float inv_projection[16], const float last_inv_projection[16]);
int ptb_state_update_resident(ptb_context* ctx, const float inv_projection[16], const float last_inv_projection[16]);
int ptb_state_read(ptb_context* ctx, float* rgba, float* velocity2);
int ptb_state_read_rgb8(ptb_context* ctx, uint8_t* rgb, int flip_y);   /* ptb_read_rgb8 of the accumulation texture */
int ptb_state_device_ptr(ptb_context* ctx, void** dptr, size_t* bytes);
/* device time (CUDA events) and count of the update kernels launched while ptb_set_profiling was on */
int ptb_state_timing(ptb_context* ctx, float* ms_total, uint32_t* launches, int reset);
int ptb_frame_next(ptb_context* ctx);
int ptb_camera_view_projection(ptb_context* ctx, float m16[16]);
/* Optional second moment (sum of squared sample values) for statistical parity tests; must be enabled before ptb_frame_begin. */
int ptb_enable_second_moment(ptb_context* ctx, int enable);
int ptb_read_second_moment(ptb_context* ctx, float* rgba);
/* Device pointer of the accumulation buffer (W*H*4 floats), e.g. for an external collective; the library's own multi-GPU
 * path is ptb_group_* below. */
int ptb_accum_device_ptr(ptb_context* ctx, void** dptr, size_t* bytes);
/* Use caller-owned device memory (e.g. a torch tensor) as accumulation buffer; NULL reverts to the internal one. */
int ptb_bind_accum(ptb_context* ctx, void* dptr, size_t bytes);

/* ---- counters & timing -------------------------------------------------------------------------------------------
 * counters since the last ptb_reset_counters: [0] extend rays (world closest-hit) that ran a traversal, [1] shadow rays (world
 * any-hit), [2] light probes (lights-TLAS closest-hit), [3] paths (camera samples), [4] wavefront iterations, [5] kernel
 * launches, [6] the part of [0] traced inside the single-launch kernels (k_path / tail) instead of k_extend launches (primary
 * rays = [3]: every path starts with one), [7] primary rays settled WITHOUT a traversal (camera outside the scene bounds, ray
 * cannot reach them, constant background: the reference ends such a ray at its root-box test, tlas.rs:68) — not part of [0]. */
int ptb_get_counters(ptb_context* ctx, uint64_t out[8]);
int ptb_reset_counters(ptb_context* ctx);
/* With profiling on, every kernel launch of ptb_render is bracketed by CUDA events on the context stream.
 * times_ms: [0] whole ptb_render call (always measured), [1] init, [2] extend (incl. finalise/accumulate/raygen of regenerated
 * slots), [3] shade, [4] nee (shadow + light-probe rays), [5] tail (all stages, one launch), [6] combine, [7] reserved; launches: same indices. */
int ptb_set_profiling(ptb_context* ctx, int enable);
int ptb_get_timing(ptb_context* ctx, float times_ms[8], uint32_t launches[8]);
/* Active slots (paths in flight + slots awaiting regeneration) after each wavefront iteration of the last ptb_render (first 8192). */
int ptb_get_iteration_log(ptb_context* ctx, uint32_t capacity, uint32_t* active, uint32_t* count);
/* Every kernel launch of the last ptb_render that ran with profiling on, in launch order: stage index as in ptb_get_timing
 * (2 extend, 3 shade, 4 nee, 5 tail / path, 6 combine) and its duration in microseconds (CUDA events on the context stream). */
int ptb_get_launch_log(ptb_context* ctx, uint32_t capacity, int32_t* stage, float* microseconds, uint32_t* count);
/* Tunables (evidence switches for DESIGN.md's layout decisions), name = value:
 *   extend_kernel     0 the 8-wide persistent kernels (default), 1 the reference-order kernels over the reference's binary trees
 *   wide_collapse     8-wide collapse of the reference trees at the next ptb_build: 1 least summed node area (default), 0 greedy
 *   slots_per_pixel   path slots per pixel, 1..16 (default 16)
 *   dense             1 k_shade_ordered (slot-ordered shading, ascending queues; default), 0 the queue shade kernel throughout
 *   dense_threshold   live slots per listed 32-slot group, in percent, below which the queue shade kernel takes over (default 60)
 *   shade_sort        span-local sort by shade kind in k_shade_ordered: 1, 0, < 0 automatic (scenes with >= 2 surface material kinds)
 *   shade_class       k_shade_ordered compiled for the scene's class of materials (< 0, default: Lambertian + Emissive / + GGX metal /
 *                     all kinds, each without media; the kinds a scene does not hold are compiled out) or the general body (0)
 *   sort_materials    queue shade kernel: one queue per material kind (default 1)
 *   extend_blocks, nee_blocks   resident blocks per SM (= register cap) of k_extend (5..8, default 7) / k_nee (4..8, default 8)
 *   refill_below, nee_refill_below   lanes of a warp still traversing below which idle lanes take new rays (closest-hit: < 0 = default, 4 with the
 *                     multiray kernels, else 8; any-hit 8)
 *   node_quorum       lanes that must want the node stage before it runs (default 8)
 *   multiray          rays per lane of k_extend / k_nee: 0 one (persist.cuh), 2 / 3 the multiray kernels (csrc/multiray.cuh: a lane owns
 *                     K rays, the thin stages run once enough lanes hold a ray waiting for them), < 0 automatic (default): 2 for scenes
 *                     of >= 384 wide nodes.  Same results bit for bit.  mr_qn / mr_ql / mr_qi (12 / 8 / 8): lanes that must hold a ray
 *                     waiting for the node / leaf / instance stage before a warp runs it; mr_free_quorum (24): finished or idle ray
 *                     slots of a warp that trigger a refill; mr_blocks: resident blocks per SM (0 = what fits); mr_min_per_lane (16):
 *                     iterations with fewer live slots per lane of the one-ray grid run the one-ray kernels; mr_split (1): k_extend in two
 *                     launches, continuing slots two rays per lane, slots to regenerate (camera rays) one per lane; regen_blocks (7): resident
 *                     blocks per SM of that second launch (5..8)
 *   tile_pixels       1 (default): path slots are laid over the frame in tiles of 8 x 4 pixels, so that the 32 rays of a warp cover a compact
 *                     patch of the image instead of a 32 x 1 line (frames whose width divides by 8 and height by 4; same results bit for
 *                     bit, 1 - 1.5 % per frame); 0: row-major
 *   tile_lists        camera rays start from their tile's list of TLAS leaves instead of the TLAS root: once per ptb_render one thread per
 *                     8 x 4 pixel tile walks the world TLAS with the tile's frustum and lists the instances its rays can reach, nearest
 *                     first (csrc/render.cu: k_tile_lists; needs tile_pixels).  1, 0, < 0 (default) = scenes of >= 64 instances.  Same
 *                     results but for equal hits in different instances (visit order).  tile_nodes (1): the walk goes on inside an
 *                     instance while the frustum stays within one to three BLAS nodes; camera rays start their BLAS walk at those nodes
 *                     (an instance whose BLAS the frustum misses is dropped from the list).  tile_settle (0): 1 = a camera ray whose tile's list
 *                     is empty is settled in the regeneration loop like one that cannot reach the scene bounds (counter [7]; same frame)
 *   nee_overlap       thin iterations (fewer live slots than this many per lane of the one-ray grid; 0 never = default, < 0 always): k_nee
 *                     of iteration i runs on a stream of its own beside the continuing slots' part of iteration i + 1's k_extend — a
 *                     continuing slot's next ray does not depend on its pending direct-light rays (same results; measured without gain: the thin iterations are issue-bound work on hard rays)
 *   tail_threshold    active slots at which the wavefront loop hands the remaining slots to the single-launch tail kernel
 *                     (< 0 automatic = 2.5 x the resident slot capacity of that kernel, 0 = never)
 *   tail_roles        1 (default): the tail kernel runs shadow ray, light probe and the next bounce's ray of a slot on three warps
 *                     side by side; 0: one thread per slot runs them in sequence (same results)
 *   tail_slots_per_block   slots per three-warp block of that kernel, 1..32 (default 32; same results)
 *   wavefront_groups  independent wavefronts per render, each over 1/G of the slots with its own queues on its own stream,
 *                     1..8 (default < 0: 1 with the multiray kernels, else 2; bit-identical results for every G; forced to 1 while profiling is on)
 *   path_threshold    active slots at which ALL remaining slots go to the persistent path kernel k_path (extend -> shade -> NEE
 *                     rays -> next bounce chained per lane in one launch); 0 = never (default: measured slower than thin wavefront
 *                     iterations + tail kernel), < 0 automatic; path_blocks (4..6) and finish_quorum (1..32,
 *                     lanes with a finished traversal that make a warp leave the stepping loop and shade) tune it (same results)
 *   stack_shared      first 8 / 12 / 16 entries of every thread's traversal stack in shared memory instead of local memory
 *                     (k_extend<7> / k_nee<8>; 0 = local memory only, the default — measured; same results)
 *   shared_device     1: another context renders on this device at the same time (ptb_render_async pairs): persistent grids leave
 *                     one block slot per SM free;  thin_rays_per_lane > 0: persistent kernels take only the blocks their queue
 *                     fills at that many entries per lane (both 0 by default; same results)
 *   block_threads, blocks_per_sm     launch shape of the queue shade / reference-order kernels */
int ptb_set_option(ptb_context* ctx, const char* name, int value);


/* ---- allocation-free variant of ptb_trace_closest ---------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that must not allocate device memory
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same arguments (host arrays), same 8-wide
 * persistent kernel (kernel = 0) and bit-identical results as ptb_trace_closest, but the call makes NO cudaMalloc / cudaFree or other
 * allocating call: the staged rays, the staged results and the kernel's work cursor live in one file-scope device array of
 * PTB_STATIC_SCRATCH_BYTES (csrc/noalloc.cu) that every GPU which loads the library holds until the process exits.
 * Limits: n <= PTB_STATIC_TRACE_MAX_RAYS, otherwise PTB_ERR_INVALID with a message naming the limit.  All calls on a device share the
 * array, which keeps the previous call's contents (the call clears what must start at zero): calls that overlap on one GPU — two
 * contexts, threads or processes — race and must be serialised by the caller.  The scene must be built (ptb_build allocates; do
 * that outside the allocation-free region). */
#define PTB_STATIC_TRACE_MAX_RAYS 65536
#define PTB_STATIC_SCRATCH_BYTES 4325632
int ptb_trace_closest_static(ptb_context* ctx, int which, uint64_t n, const float* o3, const float* d3, const float* tmax, uint8_t* hit, float* t, float* u, float* v,
                             uint32_t* prim, uint32_t* blas, uint32_t* inst, uint8_t* front, float* normal3);

/* ---- load_obj  blas.rs:44-131 ---------------------------------------------------------------------------------------------
 * The reference's only on-disk input: an OBJ file -> the triangle vertex array ptb_scene_upload takes (ntri * 3 vertices x
 * (position xyz, normal xyz)), with load_obj's rules: `v` / `vn` / `f` lines only, 1-based indices (negative = from the end), `vn`
 * normalised on load, every face token v/vt/vn or v//vn, fan triangulation (refs[0], refs[i], refs[i+1]; a face of one or two
 * references yields nothing, blas.rs:97), normal index 0 -> the un-normalised face normal cross(p1-p0, p2-p0) (blas.rs:109-116).
 * Where the reference panics (bad number, index out of bounds, empty line) these return PTB_ERR_INVALID; ptb_load_obj_error gives
 * the message with the line number.  *verts is allocated by the library: release it with ptb_free.  Host code, no device needed. */
int ptb_load_obj(const char* path, float** verts, uint64_t* ntri);
int ptb_load_obj_text(const char* text, size_t len, float** verts, uint64_t* ntri);
const char* ptb_load_obj_error(void);
void ptb_free(void* p);

/* ---- multi-GPU: sample-partitioned rendering, one NCCL reduce per frame batch ---------------------------------
 * The frame loop main.rs:181-207 is a pure map over (pixel, sample) and the accumulation (accumulate.wgsl:23) a commutative
 * float sum, so a frame batch shards by GLOBAL SAMPLE INDEX: every GPU holds the full scene and renders a contiguous block of the
 * batch's samples for all pixels; RNG streams are keyed by (seed, pixel, global sample index), so the union of the samples is
 * the same set for every group size.  One collective per batch: ncclReduce(sum, float32, W*H*4) of the accumulation buffers
 * onto the root over NVLink, called from this library (libptb.so links libnccl).
 *
 * Two ways to form a group:
 *   ptb_group_create   ONE process drives n devices (what a Rust host replacing main.rs:179-219 would use): the group owns one
 *                      context per device (ptb_group_context) and runs the members on its own host threads.
 *   ptb_group_join     one process per GPU (torchrun / MPI launchers): every rank passes its own context and the 128-byte id
 *                      that rank 0 obtained from ptb_group_unique_id (the host ships it to the other ranks by any means).
 * Group members are numbered 0..n_ranks-1 (create: the position in device_ordinals; join: the rank).
 * Scene, camera and frame size are per-context state: set them on every member (ptb_group_context) or with the broadcast
 * helpers below, which apply one call to every LOCAL member in parallel. */
typedef struct ptb_group ptb_group;
int ptb_group_create(int n_devices, const int* device_ordinals, ptb_group** out);
int ptb_group_unique_id(uint8_t id[128]);
int ptb_group_join(ptb_context* ctx, int n_ranks, int rank, const uint8_t id[128], ptb_group** out);
/* Frees the communicators; contexts made by ptb_group_create are destroyed, a joined context stays the caller's. */
void ptb_group_destroy(ptb_group* g);
const char* ptb_group_last_error(const ptb_group* g);
int ptb_group_size(const ptb_group* g, int* n_local, int* n_ranks);
ptb_context* ptb_group_context(ptb_group* g, int local_index);
/* Block of the global sample range [begin, begin + count) owned by `rank` of `n_ranks`: contiguous, the first count % n_ranks
 * ranks get one sample more (count may come out 0).  Pure host arithmetic, usable without a device. */
int ptb_group_sample_range(uint32_t begin, uint32_t count, int rank, int n_ranks, uint32_t* first, uint32_t* mine);
/* Broadcast helpers: the ptb_* call of the same name on every local member (in parallel; first failure is returned). */
int ptb_group_scene_upload(ptb_group* g, const float* verts, uint64_t ntri, const ptb_blas_desc* blas, uint32_t nblas, const ptb_instance* inst, uint32_t ninst);
int ptb_group_build(ptb_group* g);
int ptb_group_set_camera(ptb_group* g, const float origin[3], const float target[3], float fov_deg, float aspect_ratio);
int ptb_group_set_option(ptb_group* g, const char* name, int value);
int ptb_group_frame_begin(ptb_group* g, uint32_t width, uint32_t height);
/* Renders the samples [sample_begin, sample_begin + sample_count) of every pixel, partitioned over the ranks with
 * ptb_group_sample_range; every member adds its block into its own accumulation buffer (ptb_render).  Blocks until the local
 * members have finished. */
int ptb_group_render(ptb_group* g, uint32_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_bounces, int enable_nee);
/* Sums the accumulation buffers of all ranks onto `root` in place (one ncclReduce per member, on the member's stream; collective:
 * every rank must call it).  Afterwards the root's buffer holds the whole batch (w = sample_count); the other members hold their
 * partial sums until the next ptb_group_frame_begin.  position / id come from the root's own samples. */
int ptb_group_reduce(ptb_group* g, int root);
/* ptb_read_accum / ptb_read_aux of member `root` when it is local to this process, PTB_ERR_INVALID otherwise. */
int ptb_group_read_accum(ptb_group* g, int root, float* rgba);
int ptb_group_read_aux(ptb_group* g, int root, float* position4, uint32_t* id);
/* Sum of ptb_get_counters over the local members (iterations: the maximum). */
int ptb_group_get_counters(ptb_group* g, uint64_t out[8]);
int ptb_group_reset_counters(ptb_group* g);
/* Device time of the last ptb_group_reduce on the first local member, in milliseconds (CUDA events on its stream). */
int ptb_group_reduce_ms(ptb_group* g, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* PTB_H */
