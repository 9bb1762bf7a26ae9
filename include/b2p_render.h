/* b2p_render -- C-ABI of the CUDA ray-cast observations (part of libb2p.so).
 *
 * Replaces, with a leading scene dimension, the reference's three mujoco.Renderer instances on the overhead camera
 * (mujoco_robot_environments/tasks/rearrangement.py:182-186 construction; :460-478 `_compute_observation` rgb + depth;
 * :273-275 segmentation for the prop bounding boxes; :520-523 depth lookup of `pixel_2_world`):
 *
 *   b2p_render  <- renderer.update_scene(data, camera_id); renderer.render()   for rgb, depth and segmentation at once
 *
 * The reference rasterises through OpenGL (unavailable on this box); this path ray-casts the same scene -- ground plane,
 * table, translucent target zones, the arm / gripper collision hulls, the per-scene coloured props -- with the pinhole
 * camera of `_get_camera_intrinsics` / `_get_camera_extrinsics` (rearrangement.py:480-505).  Depth is the distance along
 * the camera axis in metres (mujoco.Renderer.enable_depth_rendering semantics); segmentation is the geom id (-1 = nothing,
 * ngeom + k = target zone k).
 *
 * Conventions as in b2p.h: caller owns every I/O buffer (device pointers), work is enqueued on the caller's stream, 0 or a
 * negative error code is returned, b2p_last_error() holds the message.
 *
 * Render-model blob "B2PR" (mujoco_robot_environments_b200/compile/render.py), little endian:
 *   int32   header[16]   magic 0x52503242, version 1, ngeom, nplane, nzone, width, height, nprop
 *   float32 cam[32]      pos3 | xmat9 (row major, columns = right, up, back) | f cx cy zfar | headlight ambient, diffuse |
 *                        light pos3, diffuse | ground checker rgb2[3], square size
 *   per geom 32 words    int32 type, plane_adr, plane_num, prop_id, is_static, pad3 | float32 rgba4 | size3 (half sizes; mesh:
 *                        padded AABB half) | aabb centre3 | rbound | static pos3 | static xmat9 | pad
 *   per zone 12 float32  pos3 | half3 | rgba4 | pad2
 *   planes   float32[4]  (n, d): n.x <= d inside, geom frame
 */
#ifndef B2P_RENDER_H
#define B2P_RENDER_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2p_render_s* b2p_render_handle;
enum { B2P_RQ_WIDTH = 0, B2P_RQ_HEIGHT, B2P_RQ_NGEOM, B2P_RQ_LAUNCHES, B2P_RQ_TILES_COPIED, B2P_RQ_TILES_CAST };

int b2p_render_create(const void* render_blob, size_t nbytes, int device, b2p_render_handle* out);
int b2p_render_destroy(b2p_render_handle h);
int b2p_render_query(b2p_render_handle h, int what);

/* geom_xpos / geom_xmat: world poses per scene, [n_env] records `pose_stride` reals apart, geom g at +3g / +9g
 * (the B2P_O_GEOM_XPOS / B2P_O_GEOM_XMAT fields of the derived-output record); dtype 0 = float64, 1 = float32 (applies to the
 * poses and prop_size).  prop_rgba [n_env][nprop][4] float32, prop_size [n_env][nprop][3] half sizes, prop_active
 * [n_env][nprop] (NULL = all).  Outputs (any may be NULL): rgb uint8 [n_env][H][W][3], depth float32 [n_env][H][W],
 * seg int32 [n_env][H][W]. */
int b2p_render(b2p_render_handle h, int n_env, const void* geom_xpos, const void* geom_xmat, int pose_stride, int dtype,
               const float* prop_rgba, const void* prop_size, const uint8_t* prop_active,
               uint8_t* rgb, float* depth, int32_t* seg, void* stream);

#ifdef __cplusplus
}
#endif
#endif
