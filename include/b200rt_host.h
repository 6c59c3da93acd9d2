/* b200rt_host.h — C ABI of the host library (libb200rt_host.so): the re-hosted
 * Scene / Mesh / Camera surface of the reference and its per-frame Scene::update
 * (Scene.cpp:139-171, incl. the TopLevelBVH rebuild TopLevelBVH.cpp:32-45), the `<obj>.bvh`
 * cache reader/writer (BottomLevelBVH.cpp:149-192) and the texture ingest (Texture.cpp:30-129).
 * No CUDA in this library. Its outputs are the flat views declared in b200rt.h, which are
 * handed to b200rt_upload_static / b200rt_render_frame.
 *
 * Every function returns 0 on success and < 0 on error, with the message retrievable through
 * b200rt_host_last_error() (thread-local).
 */
#ifndef B200RT_HOST_H
#define B200RT_HOST_H

#include "b200rt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200rt_scene b200rt_scene;

const char * b200rt_host_last_error(void);

/* Scene::Scene — builds the scene from a description file (DESIGN.md "Scene description"):
 * loads / builds every BottomLevelBVH (BottomLevelBVH::load), materials, textures, sky.      */
int  b200rt_scene_load(const char * scene_file, b200rt_scene ** scene);
void b200rt_scene_free(b200rt_scene * scene);

/* Camera::resize (Camera.cpp:5-16) for a resolution other than the file's.                  */
int b200rt_scene_resize(b200rt_scene * scene, int width, int height);
int b200rt_scene_resolution(const b200rt_scene * scene, int * width, int * height);

/* Overrides camera position / rotation quaternion (x, y, z, w); takes effect at next update. */
int b200rt_scene_set_camera(b200rt_scene * scene, const float position[3], const float rotation[4]);

/* Scene::update(delta). animation: 0 = static scene, 1 = the SCENE_DYNAMIC motion
 * (Scene.cpp:141-155; needs >= 6 mesh instances).                                            */
int b200rt_scene_update(b200rt_scene * scene, float delta, int animation);

/* Flat views into host memory owned by the scene; valid until the next load / update / view. */
int b200rt_scene_static_view(b200rt_scene * scene, b200rt_static_scene * out);
int b200rt_scene_frame_view (b200rt_scene * scene, b200rt_frame * out);

/* Flattened triangle index -> index of the triangle in OBJ / cache-file order, for one BLAS
 * (the inverse of BottomLevelBVH::flatten, BottomLevelBVH.cpp:196-212). Returns the count.   */
int b200rt_scene_blas_original_indices(const b200rt_scene * scene, int blas, int32_t * out, int capacity);

/* BottomLevelBVH::build_sbvh (BottomLevelBVH.cpp:108-147 -> BVHBuilders::build_sbvh, BVHBuilders.h:48-329):
 * builds the reference's spatial-split BVH over `triangle_count` triangles given as float[n][3][3]
 * vertex positions, on `threads` host threads (the result does not depend on the thread count).
 * Outputs are in the `<obj>.bvh` form: nodes (root 0, node 1 unused) whose leaves index `indices`.
 * Call with nodes == NULL / indices == NULL first to learn the counts.                          */
int b200rt_build_sbvh(const float * vertices, int triangle_count, int threads,
                      b200rt_bvh_node * nodes, int node_capacity, int32_t * indices, int index_capacity,
                      int * node_count, int * index_count);

/* `<obj>.bvh` is the reference's headerless dump (BottomLevelBVH.cpp:149-192: no magic, version or
 * checksum). `<obj>.b200bvh` wraps the same payload in a 32-byte header: magic "B200BVH\0", endian
 * tag 0x01020304, version, the three counts, CRC-32 of the payload. The scene loader prefers a
 * valid container, falls back to the dump, and writes both. These convert between the two; a
 * damaged container is reported (-1 + message), never loaded.                                   */
int b200rt_bvh_to_container(const char * legacy_bvh, const char * container);
int b200rt_bvh_container_to_legacy(const char * container, const char * legacy_bvh);

/* Tile partition helper: number of 32x32 tiles and the tiles of one GPU.   */
int b200rt_tile_count(int width, int height, int * tiles_x, int * tiles_y);

#ifdef __cplusplus
}
#endif
#endif /* B200RT_HOST_H */
