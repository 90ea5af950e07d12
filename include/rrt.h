/*
 * rrt.h — C ABI of the B200-native renderer that replaces the per-pixel hot path of
 * conor722/rust-ray-tracer (primary-ray generation -> octree traversal -> Moller-Trumbore
 * -> textured / normal-mapped shading with shadow and mirror rays).
 *
 * The reference has no FFI of its own; the seam these entry points replace is
 *     Scene::draw_scene(&mut self, rt: RayTracer)            src/scene/engine.rs:186-255
 * called once from main()                                     src/main.rs:74
 * and, per ray, RayTracer::get_ray_colour(origin, direction)  src/scene/raytracer.rs:29-31.
 *
 * Everything is plain C: caller-owned pointers + sizes, no torch / C++ types.
 * All scene arithmetic is IEEE f64 (the reference's Vector3d is f64, src/scene/engine.rs:9-14).
 * All functions return RRT_OK (0) or a negative rrt_status; rrt_last_error() gives the text.
 * Nothing here ever falls back to a CPU renderer: without a CUDA device the render calls fail
 * with RRT_ERR_CUDA.
 */
#ifndef RRT_H_
#define RRT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RRT_ABI_VERSION 1
#define RRT_MAX_LIGHTS 16

typedef enum rrt_status {
    RRT_OK = 0,
    RRT_ERR_INVALID = -1,   /* bad argument / malformed scene (reference would panic)      */
    RRT_ERR_CUDA = -2,      /* CUDA runtime error, or no device                             */
    RRT_ERR_IO = -3,        /* file could not be read (loader)                              */
    RRT_ERR_PARSE = -4,     /* .obj / .mtl syntax the reference's parser would panic on     */
    RRT_ERR_DEPTH = -5,     /* octree deeper than RRT_MAX_OCTREE_DEPTH (reference: unbounded)*/
    RRT_ERR_NOMEM = -6
} rrt_status;

/* ---- scene description (inputs of RayTracer / SceneData, src/scene/raytracer.rs:22-26,
 *      src/scene/scenedata.rs:6-13).  Borrowed for the duration of rrt_scene_create only. ---- */

/* Texture { colours: Vec<Color>, width, height }  src/scene/entities.rs:87-91.
 * rgb = packed RGB8, row-major, 3*width*height bytes (src/file_management/utils.rs:353-361). */
typedef struct rrt_texture {
    int32_t width;
    int32_t height;
    const uint8_t* rgb;
} rrt_texture;

/* Material  src/scene/material.rs:11-22 (Ka, Kd, Ks, Ns, texture, bump_map, reflectivity). */
typedef struct rrt_material {
    double ka[3];
    double kd[3];
    double ks[3];
    double ns;        /* specular_weight; -1.0 disables the specular term (raytracer.rs:288) */
    double kr;        /* reflectivity in [0,1]                                               */
    int32_t tex_id;   /* index into textures[]; required (utils.rs:61)                       */
    int32_t bump_id;  /* index into textures[] or -1 (Option<Arc<Texture>>)                  */
} rrt_material;

/* Triangles in `f`-line order (== octree.triangles order, src/collision/octree.rs:45-47);
 * that index is what hit-id planes report. */
typedef struct rrt_scene_desc {
    uint64_t n_tris;
    const double* v;            /* n_tris x 9 : v1.xyz v2.xyz v3.xyz      (entities.rs:75-77) */
    const double* vt;           /* n_tris x 6 : (x,y) of v1,v2,v3 tex     (entities.rs:78-80) */
    const double* vn;           /* n_tris x 9 : v1,v2,v3 normals          (entities.rs:81-83) */
    const int32_t* material_id; /* n_tris     : index into materials[]                         */
    uint32_t n_materials;
    const rrt_material* materials;
    uint32_t n_textures;
    const rrt_texture* textures;
    double root_min[3];         /* octree root box; reference hard-codes -20 (utils.rs:145)    */
    double root_max[3];         /* ... and +20                                                  */
} rrt_scene_desc;

/* Light enum  src/scene/entities.rs:5-9 */
enum { RRT_LIGHT_AMBIENT = 0, RRT_LIGHT_POINT = 1, RRT_LIGHT_DIRECTIONAL = 2 };
typedef struct rrt_light {
    int32_t kind;
    int32_t _pad;
    double intensity;
    double vec[3];   /* position (point) / direction (directional) / unused (ambient) */
} rrt_light;

/* Per-frame parameters: the constants main() hard-codes (src/main.rs:14-15,32-66), the Viewport
 * (src/scene/engine.rs:106-120) and raytracer.rs:17,20.  rrt_render_params_default() fills the
 * reference's values. */
typedef struct rrt_render_params {
    int32_t width;             /* canvas width  (main.rs:14)                         */
    int32_t height;            /* canvas height (main.rs:15)                         */
    double origin[3];          /* RayTracer.origin (main.rs:62-66)                   */
    double viewport[3];        /* width, height, distance = 1,1,1 (engine.rs:113-119)*/
    int32_t n_lights;
    int32_t max_depth;         /* MAX_REFLECTION_DEPTH = 5 (raytracer.rs:20)         */
    rrt_light lights[RRT_MAX_LIGHTS];
    double surface_offset;     /* SURFACE_OFFSET = 1e-4 (raytracer.rs:17)            */
    /* Extension (the reference has no camera rotation): if use_camera_rot != 0 the
     * primary direction is camera_rot (row-major 3x3) * (x*xs, y*ys, z). 0 => bit-exact reference. */
    int32_t use_camera_rot;
    int32_t _pad;
    double camera_rot[9];
} rrt_render_params;

/* Row sharding for multi-GPU: buffer rows are grouped in tile rows of `tile_rows`
 * scan-lines; tile row j belongs to shard (j % n_shards).  n_shards = 1 => whole frame. */
typedef struct rrt_shard {
    int32_t tile_rows;
    int32_t n_shards;
    int32_t shard_index;
} rrt_shard;

/* Octree / traversal statistics (structure that src/collision/octree.rs:41-241 produced). */
typedef struct rrt_scene_stats {
    uint64_t n_tris;           /* triangles given                                   */
    uint64_t n_tris_stored;    /* triangles that landed in some node list           */
    uint64_t n_tris_dropped;   /* AABB misses the root box (octree.rs:71-73)        */
    uint64_t n_nodes;          /* reference node count (incl. empty)                */
    uint64_t n_nodes_nonempty; /* nodes with triangle_count > 0 (kept on device)    */
    uint64_t max_depth;
    uint64_t root_list;        /* triangles owned by the root                       */
    uint64_t max_list;         /* longest non-root node list                        */
    uint64_t n_cull_nodes;     /* nodes of the per-list culling hierarchies         */
    uint64_t device_bytes;     /* bytes resident in HBM for this scene              */
    uint64_t flat_digest;      /* FNV-1a of the flattened node / record / culling arrays (rrt_octree_flatten_stats only) */
} rrt_scene_stats;

/* Per-frame work counters summed over all rays (optional; see rrt_render_counters). */
typedef struct rrt_counters {
    uint64_t rays_primary;
    uint64_t rays_shadow;
    uint64_t rays_reflect;
    uint64_t mt_tests;      /* exact Moller-Trumbore evaluations started  */
    uint64_t box_tests;     /* child-octant slab tests (ray.rs:135-144)   */
    uint64_t cull_tests;    /* culling-hierarchy box tests                */
    uint64_t nodes_entered; /* non-empty octree nodes entered             */
} rrt_counters;

typedef struct rrt_scene rrt_scene;   /* opaque: owns the device copy of the flattened scene */

/* ---- library ---- */
int rrt_abi_version(void);
const char* rrt_last_error(void);            /* thread-local, never NULL */
int rrt_device_count(void);                  /* number of CUDA devices, 0 if none / no driver */

/* ---- parameters ---- */
/* Fills width/height 800x800, camera (0,2,-10), the four lights of src/main.rs:32-58,
 * viewport 1,1,1, max_depth 5, surface_offset 1e-4, identity camera. */
void rrt_render_params_default(rrt_render_params* p);

/* ---- scene lifetime (replaces: RayTracer moved into draw_scene, engine.rs:186-187) ---- */
/* Builds the reference octree on the host by exact sequential insertion
 * (src/collision/octree.rs:41-241, src/collision/aabb.rs:25-60), flattens it, uploads to `device`. */
int rrt_scene_create(const rrt_scene_desc* desc, int device, rrt_scene** out);
void rrt_scene_destroy(rrt_scene* scene);
int rrt_scene_get_stats(const rrt_scene* scene, rrt_scene_stats* out);
/* Dispatch-order hint (no counterpart in the reference: rayon steals rows dynamically, engine.rs:201-203).
 * Every render call on a scene handle records what each 16x4-pixel block of the frame cost; the next call of the
 * same shape (resolution and shard) hands the blocks out heaviest first, so the persistent kernel does not end on
 * a slow tile.  Only the ORDER of the work changes: frames and hit ids are bit-identical with the hint on or off.
 * enable = 1 (default; RRT_ORDER_HINT=0 in the environment turns it off at scene creation) | 0; either way the
 * recorded order is dropped, so the next frame runs in raster order.  Render calls on ONE handle must be issued
 * on one stream at a time (they share the handle's scratch and this state). */
int rrt_scene_order_hint(rrt_scene* scene, int enable);

/* Optional, once per frame shape: allocates NOW what the first render call of this shape (resolution + shard, NULL =
 * whole frame) would otherwise allocate inside that call — the per-pixel accumulators, the arrays of the dispatch-order
 * hint, and with_device_frame != 0 the device frame of the copy path (pageable host buffers) — so that the first frame
 * costs what a frame costs.  The analogue of Canvas::new(width, height) allocating the buffer before draw_scene
 * (src/scene/engine.rs:123-128, src/main.rs:59-62).  Render calls of other shapes still work (they re-allocate). */
int rrt_scene_reserve(rrt_scene* scene, const rrt_render_params* params, const rrt_shard* shard, int with_device_frame);

/* Searches that go round the octree walk (no counterpart in the reference; results are bit-identical in every mode).
 * Scene creation also builds ONE culling hierarchy over all triangles of the scene; shadow queries first ask it whether
 * any triangle is hit below max_t at all ("none" is exactly the reference's None), camera rays look for their closest
 * triangle in it and answer themselves when the octree walk provably returns the same; every
 * other ray, and every ray a search cannot decide, takes the exact walk of ray.rs:104-168.
 * mode 0: walk only; 1 (default): each search is switched on / off per scene handle from one frame to the next by how
 * many of its rays it answered; 2: always tried; 3: switched as given here (filter_on / sparse_on).  RRT_SEARCH_MODE in
 * the environment at scene creation sets the initial mode (0 also skips building the hierarchy).
 * rrt_scene_search_state (synchronises the device): out[2] filter switch, out[6] sparse switch, out[3] frames decided,
 * out[8..9] shadow queries the filter answered / handed to the walk in the last decided frame, out[10..11] camera rays
 * likewise (mode 1 only; out[0..1], out[4..5] are the running counts of a frame in flight). */
int rrt_scene_search_mode(rrt_scene* scene, int mode, int filter_on, int sparse_on);
int rrt_scene_search_state(rrt_scene* scene, uint32_t out[12]);

/* ---- rendering (replaces Scene::draw_scene, engine.rs:186-255) ---- */
/* Host-buffer call: fb_0rgb[width*height] receives the reference canvas layout
 * (0x00RRGGBB, row-major, top row first, row 0 left zero: engine.rs:146-158, entities.rs:32-36).
 * hit_ids (may be NULL): int32[height][width][4], depth-0 hit triangle per supersample in the order
 * d1..d4 of engine.rs:207-236, -1 = miss / row 0.  device_ms (may be NULL) = kernel time (CUDA events).
 * Synchronous: the frame is in fb_0rgb on return.  When fb_0rgb is PAGE-LOCKED host memory (cudaHostAlloc,
 * cudaHostRegister, rrt_host_frame_pin) the render kernel stores every finished pixel straight into it over PCIe while
 * the frame is being rendered and no device->host copy follows; pageable memory gets render + copy.  Same frame either
 * way (RRT_HOST_STORES=0 forces the copy). */
int rrt_render(rrt_scene* scene, const rrt_render_params* params,
               uint32_t* fb_0rgb, int32_t* hit_ids, float* device_ms);

/* A sequence of frames of one scene and one resolution (an orbit, a fly-through; no counterpart in the reference,
 * whose main.rs draws one frame): frame i is rendered with params[i] and copied to fbs[i] (width*height words each,
 * layout as rrt_render; pinned host memory makes the copies asynchronous) WHILE frame i+1 renders — two device
 * frames, a render stream and a copy stream inside the handle.  Synchronous: returns when the last frame is in
 * host memory.  total_ms (may be NULL) = device time from the first kernel to the last copy.
 * Each frame is what rrt_render produces for the same params.  Page-locked fbs[i] take the kernel's pixels directly
 * (see rrt_render): no copy stage at all for those frames. */
int rrt_render_frames(rrt_scene* scene, const rrt_render_params* params, int32_t n_frames,
                      uint32_t* const* fbs, float* total_ms);

/* Device-buffer call: same result written to device memory of the scene's device, enqueued on
 * `cuda_stream` (a cudaStream_t cast to void*, NULL = default stream); does not synchronise.
 * With n_shards > 1 only the owned rows are rendered and written COMPACTLY: owned row k (in
 * increasing buffer-row order) goes to d_fb[k*width ...]; see rrt_shard_rows(). */
int rrt_render_device(rrt_scene* scene, const rrt_render_params* params, const rrt_shard* shard,
                      uint32_t* d_fb, int32_t* d_hit_ids, void* cuda_stream);

/* Allocation-free variant of rrt_render_device.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same result in d_fb (and d_hit_ids, may be NULL), enqueued on
 * `cuda_stream`, does not synchronise — but it calls no cudaMalloc / cudaFree and no allocating library routine:
 * the kernel's scratch (per-pixel accumulators + queue head) is a static device array inside the library, 8 MiB held
 * by every GPU that loads the library until the process exits.  Limits and rules:
 *   - the shard may hold at most RRT_NOALLOC_MAX_PIXELS pixels (rows owned x width; 1920x1080 fits), larger calls
 *     return RRT_ERR_INVALID with a message that names both numbers;
 *   - every call on a GPU shares the one array (it is cleared on the stream before use): calls that OVERLAP on the
 *     GPU — two streams, two threads, two scene handles — race; issue them one after another;
 *   - the dispatch-order hint (rrt_scene_order_hint) is not used: its arrays are allocated on demand;
 *   - the first call on a device resolves the array's address and the kernel's occupancy (no allocation, but make one
 *     warm-up call before capturing a graph). */
#define RRT_NOALLOC_MAX_PIXELS (1u << 21)
int rrt_render_device_noalloc(rrt_scene* scene, const rrt_render_params* params, const rrt_shard* shard,
                              uint32_t* d_fb, int32_t* d_hit_ids, void* cuda_stream);

/* Fused render + gather: the shard's rows are stored straight into a FULL frame (row r of the frame at
 * d_frame[r*width ...], reference layout), which may live on ANOTHER GPU: a pointer obtained from
 * rrt_ipc_open_handle() on a frame a peer process allocated with rrt_frame_alloc().  Every pixel is
 * written once, 4 bytes, by the lane that finishes its 4th supersample; pixel stores of different
 * shards never overlap, so no ordering is needed beyond "all ranks' kernels have completed" (a barrier
 * on the callers' streams).  Replaces per frame: gather of compact rows + rrt_deinterleave_device. */
int rrt_render_device_into_frame(rrt_scene* scene, const rrt_render_params* params, const rrt_shard* shard,
                                 uint32_t* d_frame, void* cuda_stream);

/* Frame buffers that can be shared between the one-process-per-GPU ranks of a node (CUDA IPC over
 * NVLink peer access).  handle = 64 opaque bytes (cudaIpcMemHandle_t) to send to the peers. */
int rrt_frame_alloc(int device, int32_t width, int32_t height, uint32_t** d_frame);
int rrt_frame_free(int device, uint32_t* d_frame);
int rrt_ipc_get_handle(int device, const void* d_ptr, unsigned char handle[64]);
int rrt_ipc_open_handle(int device, const unsigned char handle[64], void** d_ptr);
int rrt_ipc_close_handle(int device, void* d_ptr);
int rrt_copy_frame_to_host(int device, const uint32_t* d_frame, int32_t width, int32_t height,
                           uint32_t* fb_0rgb, void* cuda_stream);   /* async on the stream if fb is pinned */

/* ---- one canvas in HOST memory for the one-process-per-GPU ranks of a node ----
 * The reference's output is Canvas.buffer: Vec<u32> in host memory (src/scene/engine.rs:123-128,146-158).  With the
 * frame sharded over N GPUs every rank copies ITS rows device->host over its own PCIe link, in parallel, straight
 * into one page-locked frame that all ranks map: a POSIX shared-memory segment (name starts with '/').  Its header
 * page also holds a process-shared barrier and a slot for the CUDA IPC handle of the peer-store mode above, so a
 * host without torch.distributed (the Rust crate, a C++ driver) can run the ranks: tests/native/multi_rank_driver.cpp.
 *   rank 0: rrt_host_frame_create(name, W, H, n_ranks)      other ranks: rrt_host_frame_open(name, timeout)
 *   every rank, per frame:  rrt_render_to_host_frame(...); cudaStreamSynchronize; rrt_host_frame_barrier  -> the
 *   frame is complete in rrt_host_frame_pixels(); a second barrier before the next frame's render lets rank 0
 *   read it undisturbed.  Barriers and waits time out (seconds; <= 0: a default) instead of hanging. */
typedef struct rrt_host_frame rrt_host_frame;
int rrt_host_frame_create(const char* name, int32_t width, int32_t height, int32_t n_ranks, rrt_host_frame** out);
int rrt_host_frame_open(const char* name, double timeout_s, rrt_host_frame** out);
uint32_t* rrt_host_frame_pixels(rrt_host_frame* f);          /* width*height words, reference canvas layout */
int rrt_host_frame_geometry(const rrt_host_frame* f, int32_t* width, int32_t* height, int32_t* n_ranks);
int rrt_host_frame_pin(rrt_host_frame* f, int device);       /* cudaHostRegister: makes the rank's copies asynchronous */
int rrt_host_frame_barrier(rrt_host_frame* f, double timeout_s);
int rrt_host_frame_publish_handle(rrt_host_frame* f, const unsigned char handle[64]);
int rrt_host_frame_fetch_handle(rrt_host_frame* f, double timeout_s, unsigned char handle[64]);
void rrt_host_frame_close(rrt_host_frame* f);                /* the creator also unlinks the segment */
/* Renders the shard (as rrt_render_device) and enqueues the copy of its rows to their places in the FULL host frame
 * `host_frame` (width*height words; rrt_host_frame_pixels() or any caller buffer, asynchronous if page-locked) on
 * `cuda_stream`; does not synchronise.  With n_shards = 1 this is rrt_render without the blocking copy.  A page-locked
 * host_frame takes the kernel's pixels directly (see rrt_render): nothing but the kernel is enqueued. */
int rrt_render_to_host_frame(rrt_scene* scene, const rrt_render_params* params, const rrt_shard* shard,
                             uint32_t* host_frame, void* cuda_stream);

/* The copy half of rrt_render_to_host_frame on its own: d_rows = the shard's rows as rrt_render_device wrote them
 * (compact, device memory of `device`), enqueued on `cuda_stream`.  Lets a caller launch the render, THEN wait for
 * "the previous frame has been read" (rrt_host_frame_barrier), then enqueue the copy: the wait hides behind the kernel. */
int rrt_copy_shard_to_host_frame(int device, const uint32_t* d_rows, int32_t width, int32_t height, const rrt_shard* shard,
                                 uint32_t* host_frame, void* cuda_stream);

/* Waits for everything enqueued on `cuda_stream` (NULL = default stream) of the scene's device: the asynchronous calls
 * above for a host that does not link the CUDA runtime itself (the Rust crate; tests/native/multi_rank_driver.cpp). */
int rrt_synchronize(rrt_scene* scene, void* cuda_stream);

/* Number of buffer rows shard `shard_index` owns for a frame of `height` rows (row 0 included). */
int rrt_shard_rows(int32_t height, const rrt_shard* shard);

/* Scatter gathered compact shards back to a full frame: d_gathered = n_shards blocks of
 * `rows_per_block` rows each (block s = shard s, padded), d_fb = full frame.  Row 0 is zeroed. */
int rrt_deinterleave_device(int device, const uint32_t* d_gathered, int32_t rows_per_block,
                            int32_t width, int32_t height, int32_t tile_rows, int32_t n_shards,
                            uint32_t* d_fb, void* cuda_stream);

/* Per-ray seam (RayTracer::get_ray_colour, raytracer.rs:29-31; Ray::intersect_with_octant_with_max_t,
 * ray.rs:104-168).  Host arrays: origins/dirs n x 3, max_t n (NULL = +inf).
 * rrt_trace_rays: out_tri n (int32, -1 miss), out_tuv n x 3 (t,u,v; untouched on miss).
 * rrt_ray_colours: out_rgb n x 3 (u8). */
int rrt_trace_rays(rrt_scene* scene, uint64_t n, const double* origins, const double* dirs,
                   const double* max_t, int32_t* out_tri, double* out_tuv);
int rrt_ray_colours(rrt_scene* scene, const rrt_render_params* params, uint64_t n,
                    const double* origins, const double* dirs, uint8_t* out_rgb);

/* Render one frame with work counters enabled (slower instrumented kernel). fb may be NULL. */
int rrt_render_counters(rrt_scene* scene, const rrt_render_params* params,
                        uint32_t* fb_0rgb, rrt_counters* out);

/* ---- host-side octree (src/collision/octree.rs) exposed for inspection / tests; no GPU needed ---- */
typedef struct rrt_octree rrt_octree;
int rrt_octree_build(const double* v /* n x 9 */, uint64_t n_tris,
                     const double root_min[3], const double root_max[3], rrt_octree** out);
void rrt_octree_destroy(rrt_octree* t);
uint64_t rrt_octree_node_count(const rrt_octree* t);
/* FNV-1a over every node (box, first child, triangle_count, list ends, depth) and the list links: two trees
 * are the same tree iff the digests agree (used to check the parallel builder against sequential insertion;
 * RRT_OCTREE_BUILD=seq|par selects the builder, default par from 20 000 triangles). */
uint64_t rrt_octree_digest(const rrt_octree* t);
/* For node i: box (6 doubles min.xyz max.xyz), triangle_count (octree.rs:11), first child index
 * (0 = leaf; children are 8 consecutive nodes, octree.rs:226-237), list length. */
int rrt_octree_node(const rrt_octree* t, uint64_t i, double box[6], uint64_t* triangle_count,
                    uint64_t* first_child, uint64_t* list_len);
/* Host-only: flatten the tree into the device layout (what rrt_scene_create uploads) and report its
 * statistics, a digest of the arrays and the seconds it took.  v = the triangle array the tree was built from. */
int rrt_octree_flatten_stats(const rrt_octree* t, const double* v, rrt_scene_stats* out, double* seconds);
/* Copies node i's triangle list (insertion order, octree.rs:78,91,103) into out[0..cap). Returns length. */
int64_t rrt_octree_node_list(const rrt_octree* t, uint64_t i, uint32_t* out, uint64_t cap);

/* ---- loaders (src/file_management/utils.rs:22-343), host only ---- */
/* Parsed scene owned by the library. Textures are NOT decoded here: the caller decodes the files
 * named by rrt_obj_texture_name() (the reference uses the `image` crate, utils.rs:345-368). */
typedef struct rrt_obj rrt_obj;
int rrt_obj_load(const char* obj_path, const char* base_dir /* for mtllib; NULL = cwd */, rrt_obj** out);
void rrt_obj_destroy(rrt_obj* o);
uint64_t rrt_obj_n_tris(const rrt_obj* o);
const double* rrt_obj_v(const rrt_obj* o);
const double* rrt_obj_vt(const rrt_obj* o);
const double* rrt_obj_vn(const rrt_obj* o);
const int32_t* rrt_obj_material_id(const rrt_obj* o);
uint32_t rrt_obj_n_materials(const rrt_obj* o);
const rrt_material* rrt_obj_materials(const rrt_obj* o);   /* tex_id / bump_id index texture names */
const char* rrt_obj_material_name(const rrt_obj* o, uint32_t i);
uint32_t rrt_obj_n_textures(const rrt_obj* o);
const char* rrt_obj_texture_name(const rrt_obj* o, uint32_t i);

/* ---- diagnostics ---- */
/* Measures the FP64 vector throughput WITHOUT FMA (separate DMUL + DADD, as the -fmad=false kernels issue
 * them): the roofline denominator of this path. sm_clock_mhz (may be NULL) = nominal max clock. */
int rrt_microbench_fp64(int device, double* tflops_no_fma, double* sm_clock_mhz);
/* Checks the kernels' FMA-based quotient (RN(a/d) from RN(1/d), rrt_trace.cuh) bit-for-bit against the
 * hardware IEEE division on ~n_pairs pseudo-random operand pairs inside its range condition. */
int rrt_selftest_division(int device, uint64_t n_pairs, uint64_t* mismatches);

/* ---- headless output (stands in for the minifb window, engine.rs:122-167) ---- */
int rrt_write_ppm(const char* path, const uint32_t* fb_0rgb, int32_t width, int32_t height);

#ifdef __cplusplus
}
#endif
#endif /* RRT_H_ */
