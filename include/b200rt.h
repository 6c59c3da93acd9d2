/* b200rt.h — C ABI of the B200-native render hot path.
 *
 * The reference (clayne/CPU-Raytracer) has no plugin / FFI interface: its hot path is
 * reached by a direct C++ call from worker threads,
 *     params.raytracer->render_tile(*params.window, x, y, w, h, *params.stats)   WorkerThread.cpp:63
 * driven once per frame by Main.cpp:54-59:
 *     scene.update(dt); wake_up_worker_threads(tiles); wait_on_worker_threads(); window.draw_quad();
 * The seam is therefore "one frame". b200rt_render_frame() replaces
 * wake_up_worker_threads + wait_on_worker_threads + draw_quad (Main.cpp:56-59): every tile
 * of the frame is ray traced (Raytracer.cpp:3-400) and the FXAA + gamma pass of
 * Data/Shaders/fragment_fxaa.glsl is applied, on the GPU. Scene::update (Scene.cpp:139-171,
 * including the per-frame TopLevelBVH rebuild) stays on the host, exactly where the
 * reference has it; its results are handed over in b200rt_frame.
 *
 * Plain C: POD structs, raw pointers, sizes. No C++ or torch types cross this boundary.
 * All arrays are borrowed for the duration of the call and copied; the context owns every
 * device allocation. A context is single-caller (not re-entrant), like the reference's
 * frame loop. Every function returns 0 on success, < 0 on error (the reference abort()s:
 * Texture.cpp:43-47, Sky.cpp:11, BottomLevelBVH.cpp:153,174); b200rt_last_error() gives text.
 */
#ifndef B200RT_H
#define B200RT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------- data layouts ------ */
/* Byte-for-byte the reference's structs, so that a host that holds the
 * reference's objects can pass them without conversion.                                  */

/* BVHNode, BVHNode.h:10-16. count bits 30-31 = split axis + 1 (inner node), bits 0-29 =
 * primitive count (> 0 => leaf). Inner: children at left, left+1. Root = node 0.          */
typedef struct { float aabb_min[3]; float aabb_max[3]; int32_t left_or_first; int32_t count; } b200rt_bvh_node;

/* BottomLevelBVH::TriangleHot, BottomLevelBVH.h:6-10 (edges pre-subtracted).              */
typedef struct { float position_0[3]; float position_edge_1[3]; float position_edge_2[3]; } b200rt_triangle_hot;

/* BottomLevelBVH::TriangleCold, BottomLevelBVH.h:12-22. material_id is mesh-local.        */
typedef struct {
	float tex_coord_0[2], tex_coord_edge_1[2], tex_coord_edge_2[2];
	float normal_0[3], normal_edge_1[3], normal_edge_2[3];
	int32_t material_id;
} b200rt_triangle_cold;

/* One BottomLevelBVH in the concatenated static arrays. Triangles are FLATTENED (leaf
 * order, BottomLevelBVH.cpp:196-212): a leaf's `first` indexes triangles directly.         */
typedef struct {
	int32_t node_offset, node_count;       /* into static_scene.nodes                       */
	int32_t triangle_offset, triangle_count; /* into static_scene.triangles_hot/_cold        */
	int32_t material_offset;                /* BottomLevelBVH::material_offset               */
} b200rt_blas;

/* Material, Material.h:7-14, re-packed: the texture pointer becomes an index (-1 = none). */
typedef struct {
	float   diffuse[3];
	int32_t texture;
	float   reflection[3];
	float   index_of_refraction;
	float   transmittance[3];
	int32_t reserved;
} b200rt_material;

/* Texture, Texture.h:8-18: linear-space float RGB texels, all mip levels contiguous.      */
#define B200RT_MAX_MIP_LEVELS 16
typedef struct {
	int64_t texel_offset;                  /* first texel of level 0 in static_scene.texels */
	int32_t width, height;
	int32_t mipmapped;                     /* Texture::mipmapped                            */
	int32_t mip_levels;                    /* 1 when not mip-mapped                         */
	int32_t mip_offsets[B200RT_MAX_MIP_LEVELS]; /* in texels, relative to texel_offset       */
} b200rt_texture;

/* Everything that does not change between frames (Scene members Scene.h:20-39 that are
 * loaded once; MaterialBuffer::materials Material.h:28-30; Sky Sky.h:4-12).                */
typedef struct {
	const b200rt_bvh_node      * nodes;          int32_t node_count;
	const b200rt_triangle_hot  * triangles_hot;
	const b200rt_triangle_cold * triangles_cold; int32_t triangle_count;
	const b200rt_blas          * blas;           int32_t blas_count;
	const b200rt_material      * materials;      int32_t material_count;  /* <= 256, Config.h:18 */
	const b200rt_texture       * textures;       int32_t texture_count;
	const float                * texels;         int64_t texel_count;     /* float[texel_count][3] */
	const float                * sky;            int32_t sky_size;        /* float[size*size (+1)][3], Sky.cpp:13-25 */
	/* Optional texture ingest ON THE GPU (replaces the host loops of Texture::load, Texture.cpp:62-117):
	 * when texels == NULL and texels_rgb8 != NULL the device converts the 8-bit level-0 images
	 * (texture after texture, width*height*3 bytes each, row 0 = top) through gamma_lut — the 256
	 * values of Math::gamma_to_linear(i * 0.00392156862f), Texture.cpp:13-20,64-72 — and box-filters
	 * the mip chains itself, with the reference's association ((c0+c1)+c2)+c3)*0.25. Same floats as
	 * the host path; a fifth of the upload. texel_count is the size of the whole pyramid pool either way. */
	const uint8_t              * texels_rgb8;
	const float                * gamma_lut;      /* float[256] */
} b200rt_static_scene;

/* Mesh, Mesh.h:8-16: instance of a BLAS. Matrices are the reference's Matrix4 (column-major
 * cells[i + 4j], Matrix4.h:19-28): world = Transform::world_matrix, inverse = Mesh::transform_inv. */
typedef struct { float world[16]; float inverse[16]; int32_t blas; int32_t reserved[3]; } b200rt_instance;

/* Sphere.h:8-22 after update: centre = transform.position.                                */
typedef struct { float center[3]; float radius_squared; float radius_inv; int32_t material_id; int32_t reserved[2]; } b200rt_sphere;

/* Plane.h:8-22 after Plane::update (Plane.cpp:3-11).                                      */
typedef struct { float normal[3]; float distance; float u_axis[3]; int32_t material_id; float v_axis[3]; int32_t reserved; } b200rt_plane;

/* PointLight.h / SpotLight.h / DirectionalLight.h. kind: 0 point, 1 spot, 2 directional.
 * Lights are evaluated in the reference's order: all point, then spot, then directional
 * (Raytracer.cpp:155-198); the array must be sorted by kind.                              */
typedef struct {
	int32_t kind;
	float   colour[3];
	float   position[3];           /* point, spot                                          */
	float   negative_direction[3]; /* spot, directional (= -direction)                     */
	float   inner_cutoff, outer_cutoff; /* spot: cos of the half angles, SpotLight.h:13-14 */
} b200rt_light;

/* Camera.h:12-23 after Camera::update: position + rotated view pyramid.                   */
typedef struct { float position[3]; float top_left_corner[3]; float x_axis[3]; float y_axis[3]; } b200rt_camera;

/* What Scene::update (Scene.cpp:139-171) produces each frame.                             */
typedef struct {
	b200rt_camera camera;
	float         ambient[3];                                   /* Scene::ambient_lighting  */
	const b200rt_instance * instances;   int32_t instance_count; /* TopLevelBVH::primitives  */
	const b200rt_bvh_node * tlas_nodes;  int32_t tlas_node_count; /* TopLevelBVH::nodes       */
	const int32_t         * tlas_indices;                        /* TopLevelBVH::indices [instance_count] */
	const b200rt_sphere   * spheres;     int32_t sphere_count;
	const b200rt_plane    * planes;      int32_t plane_count;
	const b200rt_light    * lights;      int32_t light_count;
} b200rt_frame;

/* Run-time mirror of the compile-time switches in Config.h:1-55.                          */
enum { B200RT_FILTER_TRILINEAR = 0, B200RT_FILTER_ANISOTROPIC = 1, B200RT_FILTER_EWA = 2 }; /* MIPMAP_FILTER */
typedef struct {
	int32_t width, height;        /* SCREEN_WIDTH / SCREEN_HEIGHT                            */
	int32_t bounces;              /* NUMBER_OF_BOUNCES (0..8)                                */
	int32_t mipmap_filter;        /* MIPMAP_FILTER                                           */
	float   max_anisotropy;       /* MAX_ANISOTROPY                                          */
	int32_t enable_fxaa;          /* ENABLE_FXAA: 1 = fragment_fxaa.glsl, 0 = fragment_identity.glsl */
	int32_t device;               /* CUDA device ordinal                                     */
	/* Multi-GPU tile partition: this context renders the 32x32 tiles whose
	 * row-major id t satisfies t % tile_stride == tile_phase. 1 / 0 = the whole frame.       */
	int32_t tile_stride, tile_phase;
	/* 1 = also count BVH nodes visited / triangles tested / texels fetched (b200rt_stats), the
	 * run-time cousin of BVH_VISUALIZE_HEATMAP (Config.h:23): the work of the REFERENCE's algorithm (its visiting
	 * order for shadow rays too). 2 = the same counters for the walk that ships (shadow rays on the threaded copy,
	 * right child first). Costs registers: keep 0 when timing.                                    */
	int32_t count_traversal;
	/* BVH_TRAVERSAL_STRATEGY (Config.h:27-30, BVHNode.h:30-40): 0 = BVH_TRAVERSE_TREE_ORDERED, the shipped setting (the
	 * child on the ray's side of the split plane first), 1 = BVH_TRAVERSE_TREE_NAIVE (always the left child first).    */
	int32_t traversal_naive;
	/* TEXTURE_SAMPLE_MODE (Config.h:37-42, Texture.h:33-49): B200RT_TEXTURE_*; 0 = the shipped setting.                 */
	int32_t texture_sample_mode;
	/* B200RT_ARITHMETIC_*. EXACT (0, the default): every operation in the reference's association order, a multiply-add
	 * fused exactly where the reference calls madd / msub (SIMD_Vector3.h:53-63, Matrix4.h:39-85) — primary-hit ids, hit
	 * distances and the 8-bit image come out as the reference's bits. FAST: a second set of kernels compiled with
	 * -use_fast_math (contraction, approximate division / square root / exp / log) and a cheaper box / triangle arithmetic, inside north_star's tolerance (>= 99.9 % ids, >= 50 dB) but
	 * not bit-identical — the reference's own release build is of that kind (/fp:fast, Raytracer.vcxproj).            */
	int32_t arithmetic;
	/* Deep-pixel ray queue: 0 = every pixel's ray tree is walked depth-first by one thread;
	 * 1 = pixels whose primary ray hits a transmissive material hand their reflection and refraction rays to a
	 * device-side queue that all persistent warps pop; 2 = every pixel whose primary hit spawns rays does;
	 * 3 = AUTO: the context times both (setting 0 against setting 2)
	 * on its own first frames from the kernels' %globaltimer stamps, keeps the faster and looks again every 4096 frames.
	 * Same image bit for bit in every setting: the node values are combined in the reference's order.
	 * Last line of defence: should the queued kernel ever wait 2 s on an item that does not appear, it gives the frame
	 * up; b200rt_render_frame / b200rt_render_frame_device (with stats) then render the frame again depth-first,
	 * b200rt_wait_frame / b200rt_last_frame_stats report the frame as failed (submit it again), and the context walks
	 * depth-first from then on. Never observed with the shipped kernel (tools/queue_stress.py).                      */
	int32_t deep_queue;
} b200rt_config;
enum { B200RT_ARITHMETIC_EXACT = 0, B200RT_ARITHMETIC_FAST = 1 };
enum { B200RT_QUEUE_OFF = 0, B200RT_QUEUE_GLASS = 1, B200RT_QUEUE_ALL = 2, B200RT_QUEUE_AUTO = 3 };
enum { B200RT_TEXTURE_MIPMAP = 0,   /* TEXTURE_SAMPLE_MODE_MIPMAP: mipmap_filter on mip-mapped textures, bilinear on the others */
       B200RT_TEXTURE_BILINEAR = 1, /* TEXTURE_SAMPLE_MODE_BILINEAR: Texture::sample_bilinear on level 0, Texture.cpp:156-186   */
       B200RT_TEXTURE_NEAREST = 2 };/* TEXTURE_SAMPLE_MODE_NEAREST: Texture::sample_nearest, Texture.cpp:149-154              */

/* Counters of one frame. Rays are ACTIVE rays (one per traced ray), which is what the
 * reference's lane-1 build counts (Raytracer.cpp:61,163,180,192,265,320).                  */
typedef struct {
	uint64_t primary_rays, shadow_rays, reflection_rays, refraction_rays;
	uint64_t nodes_visited;        /* TLAS + BLAS nodes whose box was tested                 */
	uint64_t triangles_tested;
	uint64_t texels_fetched;       /* texture texels + sky probe texels                        */
	uint64_t hits_accepted;        /* closest-hit candidates that replaced the running closest hit */
	uint64_t instances_entered;    /* Mesh::trace / Mesh::intersect calls (TLAS leaf visits)    */
	float    device_ms_render;     /* CUDA-event time of the ray tracing kernel(s); from b200rt_wait_frame: the
	                                * kernel's own %globaltimer stamps (pipelined frames carry no timing
	                                * events: each one stalls the stream for ~20 us)                      */
	float    device_ms_fxaa;       /* CUDA-event time of the FXAA kernel (0 from b200rt_wait_frame)       */
	float    device_ms_total;      /* uploads + kernels + downloads on the context's stream
	                                * (= device_ms_render from b200rt_wait_frame)                         */
	float    device_ms_tail;       /* part of device_ms_render during which some warps had already run out of work */
	/* deep-pixel queue (config.deep_queue != 0): node records and bounce-ray items of the frame, and how many nodes
	 * found a pool exhausted and were walked depth-first instead (correct either way). queue_shadow_items is 0:
	 * shadow rays are traced by the thread that shades the node (a second queue for them was built, measured
	 * slower and removed; the field keeps the struct's layout)                                                     */
	uint32_t queue_nodes, queue_bounce_items, queue_shadow_items, queue_fallbacks;
} b200rt_stats;

/* Host buffers to fill; any pointer may be NULL to skip that output.                      */
typedef struct {
	uint32_t * frame_buffer;   /* [height][width] 0x00RRGGBB, linear colour, Window::plot (Window.h:56-65) */
	uint8_t  * presented_rgb8; /* [height][width][3] after gamma + FXAA (fragment_fxaa.glsl)  */
	float    * colour_float;   /* [height][width][3] un-quantised linear colour (diagnostic)  */
	int32_t  * primary_ids;    /* [height][width][2] (instance, flattened triangle); miss = (-1,-1);
	                              sphere k = (-2,k); plane k = (-3,k)                          */
	float    * primary_hit;    /* [height][width][4] (distance, material id, u, v)            */
	b200rt_stats * stats;
} b200rt_outputs;

typedef struct b200rt_ctx b200rt_ctx;

/* ---------------------------------------------------------------- device path ------- */
int  b200rt_create(const b200rt_config * config, b200rt_ctx ** ctx);
void b200rt_destroy(b200rt_ctx * ctx);
const char * b200rt_last_error(const b200rt_ctx * ctx); /* ctx may be NULL: error of b200rt_create */

/* Replaces the one-off construction of Scene / BottomLevelBVH::load / Texture::load /
 * MaterialBuffer / Sky (Main.cpp:26-41): copies and re-packs the static arrays into HBM.   */
int b200rt_upload_static(b200rt_ctx * ctx, const b200rt_static_scene * scene);

/* Replaces Main.cpp:56-59 for one frame. Host buffers in, host buffers out: uploads the
 * per-frame data, renders, applies FXAA, downloads the requested outputs.                  */
int b200rt_render_frame(b200rt_ctx * ctx, const b200rt_frame * frame, const b200rt_outputs * outputs);

/* Reads the texel pool back (float[count][3], the layout of b200rt_static_scene.texels): lets a
 * test compare the GPU-ingested pyramids with the host-built ones.                              */
int b200rt_download_texels(b200rt_ctx * ctx, float * texels, int64_t first, int64_t count);

/* Pipelined form of b200rt_render_frame, for a frame loop: submit enqueues frame n (upload, trace,
 * FXAA, copies into PAGE-LOCKED output images) and returns at once; wait blocks until the OLDEST
 * submitted frame has arrived in its output images and fills its stats. At most two frames may be
 * in flight, so the host prepares and the GPU traces frame n+1 while frame n is still being copied
 * out — the same overlap the reference gets from its asynchronous GL present (Window.cpp:87-119).
 * colour_float / primary_ids / primary_hit must be NULL here.                                   */
int b200rt_submit_frame(b200rt_ctx * ctx, const b200rt_frame * frame, const b200rt_outputs * outputs);
int b200rt_wait_frame(b200rt_ctx * ctx, b200rt_stats * stats);

/* Page-locked host memory for the output images. Optional: any host pointer works in
 * b200rt_outputs, but a page-locked one (from here, or cudaHostRegister'ed by the caller) is
 * filled by the DMA engine directly, without the staging copy a pageable buffer needs.       */
int  b200rt_host_alloc(uint64_t bytes, void ** out);
void b200rt_host_free(void * p);

/* Same, but leaves the results in device memory (for callers that keep the image on the
 * GPU, e.g. the multi-GPU gather over NCCL): frame_buffer_dev is uint32 [height][width]
 * (only this context's tiles are written), presented_dev is uint8 [height][width][3] or 0. */
int b200rt_render_frame_device(b200rt_ctx * ctx, const b200rt_frame * frame,
                               uint64_t frame_buffer_dev, b200rt_stats * stats);
/* stats == NULL (and device_ms == NULL below): the call only ENQUEUES on the context's stream and
 * returns; the caller synchronises (or orders a collective) on b200rt_stream().              */

/* Counters of the frame rendered last by any entry point (waits for the context's stream). device_ms_render is the
 * render kernel's own %globaltimer span (first block started .. last warp done): what an asynchronous caller
 * (stats == NULL above) reads afterwards instead of paying for timing events on the stream.        */
int b200rt_last_frame_stats(b200rt_ctx * ctx, b200rt_stats * stats);

/* FXAA + gamma over a full pre-FXAA frame that is already in device memory (after the
 * multi-GPU gather): in uint32 [height][width], out uint8 [height][width][3].              */
int b200rt_present_device(b200rt_ctx * ctx, uint64_t frame_buffer_dev, uint64_t presented_dev, float * device_ms);

/* Multi-GPU exchange, fused into the render kernel's pixel store. The reference's
 * worker threads all write one frame_buffer (Window.h:64) that draw_quad then filters (Main.cpp:59);
 * across GPUs the frame is cut into band_count horizontal bands of band_rows rows (the last band
 * takes the remainder), band b belonging to the GPU whose image uint32 [height][width] is mapped at
 * targets[b] in THIS process (peer memory over NVLink). From the next frame on, b200rt_render_frame_device
 * stores the pixel of row y into the image of band y / band_rows and, within halo_rows of a band
 * border, into the neighbouring band's image too (cyclically: the FXAA taps wrap at the frame border),
 * so that every GPU can run b200rt_present_rows_device over its own band: no GPU gathers the frame,
 * every NVLink carries 1/N of it. fragment_fxaa.glsl reaches 4 rows; halo_rows >= 4, and no band may
 * be shorter than halo_rows. band_count = 0 restores the single target (frame_buffer_dev).            */
#define B200RT_MAX_BANDS 8
int b200rt_set_band_targets(b200rt_ctx * ctx, const uint64_t * targets, int32_t band_count, int32_t band_rows, int32_t halo_rows);

/* Peer memory and a device-side barrier, so that a host with no framework at all — the reference is a plain C++
 * program — can run the band exchange: what the worker threads' one shared frame_buffer (Window.h:56-65,
 * WorkerThread.cpp:53-65) becomes when the workers are GPUs in separate processes. Every process allocates its image
 * with b200rt_peer_alloc, hands the 64-byte handle to the others by any means it likes (a pipe, a file, MPI, ...), maps
 * theirs with b200rt_peer_open and passes the addresses to b200rt_set_band_targets. The handle is a cudaIpcMemHandle_t:
 * one node, NVLink / PCIe peer access between the devices. Memory comes back zeroed.                               */
typedef struct { unsigned char bytes[64]; } b200rt_peer_handle;
int b200rt_peer_alloc(b200rt_ctx * ctx, uint64_t bytes, uint64_t * device_ptr, b200rt_peer_handle * handle);
int b200rt_peer_free (b200rt_ctx * ctx, uint64_t device_ptr);
int b200rt_peer_open (b200rt_ctx * ctx, const b200rt_peer_handle * handle, uint64_t * device_ptr);
int b200rt_peer_close(b200rt_ctx * ctx, uint64_t device_ptr);
/* Copies `bytes` of device memory (a peer buffer, an image) to the host, after the context's stream has drained.   */
int b200rt_peer_read (b200rt_ctx * ctx, uint64_t device_ptr, void * host, uint64_t bytes);

/* Barrier of `count` contexts (<= B200RT_MAX_BANDS) ON THE DEVICE: flag_blocks[r] = address, in this process, of rank
 * r's flag block (B200RT_BARRIER_BYTES of peer memory, zeroed, never written by the host afterwards). b200rt_barrier
 * enqueues one tiny kernel on the context's stream: it tells every peer "rank `rank` has reached barrier k" with a
 * store into the peer's block and waits until its own block says so of everybody; stores issued by earlier kernels of
 * this stream (the render kernel's pixels in the other GPUs' images) are visible to the peers' later kernels. All
 * ranks must call it the same number of times. Nothing here touches the host after the call returns.            */
#define B200RT_BARRIER_BYTES 64
int b200rt_barrier_setup(b200rt_ctx * ctx, const uint64_t * flag_blocks, int32_t count, int32_t rank);
int b200rt_barrier(b200rt_ctx * ctx);

/* b200rt_present_device over rows [row_begin, row_end) only; frame_buffer_dev and presented_dev are
 * the bases of full-size images (presented_dev may be another GPU's, mapped here).                   */
int b200rt_present_rows_device(b200rt_ctx * ctx, uint64_t frame_buffer_dev, uint64_t presented_dev,
                               int32_t row_begin, int32_t row_end, float * device_ms);

/* (b200rt_render_frame, b200rt_submit_frame and b200rt_time_resident always render into the one image they own:
 * band targets apply to b200rt_render_frame_device only.)                                        */

/* Renders the frame that is resident in the context (the one last handed to b200rt_render_frame,
 * b200rt_render_frame_device or b200rt_submit_frame) once more, into the caller's device image frame_buffer_dev
 * (uint32 [height][width], this context's tiles), WITHOUT ALLOCATING ANYTHING: no cudaMalloc, cudaFree or allocating
 * library routine is called, so it can be used where those are forbidden (under CUDA graph capture, in a serving loop,
 * in a harness that bans them). This entry point has no counterpart in the reference.
 * The render kernel's scratch — its work counter, ray counters and per-tile cost words, which the
 * other entry points keep in memory the context allocated — lives in one file-scope device array of
 * 512 + 4 * B200RT_NOALLOC_MAX_TILES bytes (64.5 KB), held by every GPU that loads the library until the process exits.
 * A context with more than B200RT_NOALLOC_MAX_TILES 32x32 tiles of its own (16 384: every frame up to 4096 x 4096, and
 * a 1/N share of larger ones) is refused with an error. Every call of the process on one GPU shares that array: the
 * call clears the part it uses on the context's stream before the launch, and calls that overlap on the GPU (two
 * contexts, two streams) race — issue them one after another. Depth-first walk (deep_queue is ignored: the queued walk
 * needs pools), band targets ignored, the context's tile order and counters untouched; same image, bit for bit, as the
 * other entry points. Asynchronous: enqueues on the context's stream and returns.                                  */
#define B200RT_NOALLOC_MAX_TILES 16384
int b200rt_render_resident_noalloc(b200rt_ctx * ctx, uint64_t frame_buffer_dev);

/* Re-times the last frame without host<->device traffic: `repeat` launches of the render
 * (+FXAA) kernels on resident inputs, each timed with CUDA events on the context's stream;
 * returns the average per-frame device times. flush_l2_bytes > 0: before every timed launch
 * (outside the timed region) a scratch buffer of that size is overwritten to evict L2.       */
int b200rt_time_resident(b200rt_ctx * ctx, int repeat, uint64_t flush_l2_bytes, float * ms_render, float * ms_fxaa);

/* The CUDA stream (cudaStream_t as integer) all work of this context is issued on. A caller
 * that orders this context's work against its own (e.g. a collective on a torch stream) may
 * replace it with b200rt_set_stream — ANY handle, including 0 = CUDA's default stream, is taken
 * as given; b200rt_reset_stream goes back to the context's own stream. The context never owns a
 * caller's stream.                                                                            */
uint64_t b200rt_stream(const b200rt_ctx * ctx);
int      b200rt_set_stream(b200rt_ctx * ctx, uint64_t stream);
int      b200rt_reset_stream(b200rt_ctx * ctx);

#ifdef __cplusplus
}
#endif
#endif /* B200RT_H */
