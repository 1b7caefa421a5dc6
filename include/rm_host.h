/*
 * rm_host.h — C view of the C++ host mirror (ray-marching_b200/host/rm_host.hpp) for bindings
 * that cannot consume C++ (the ctypes tests; a Rust `extern "C"` block).  Every entry point cites
 * the reference item it exposes (paths relative to Mesoptier/ray-marching).  Handles are small
 * integers owned by the library; rmh_reset() frees everything.  Returns: >= 0 ok / handle,
 * RMH_NONE (-1) = Rust `None`, RMH_E_* (< -1) = where the reference would panic.
 */
#ifndef RM_HOST_H
#define RM_HOST_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#if defined(_WIN32)
#define RMH_API
#else
#define RMH_API __attribute__((visibility("default")))
#endif

enum { RMH_NONE = -1, RMH_E_INVALID = -2, RMH_E_PANIC = -3, RMH_E_DEVICE = -4 };
/* node / template kinds: CSGNode (csg/mod.rs:28-45) + extensions */
enum { RMH_SPHERE = 0, RMH_BOX = 1, RMH_UNION = 2, RMH_SUBTRACTION = 3, RMH_INTERSECTION = 4, RMH_SMOOTH_UNION = 5,
       /* N4: the kinds csg/mod.rs:34,41-44 reserves in comments */
       RMH_PLANE = 6, RMH_TRANSLATION = 7, RMH_ROTATION = 8, RMH_SCALE = 9 };
/* ValueType tags (csg_node_graph.rs:24-29) */
enum { RMH_VT_SCALAR = 0, RMH_VT_VEC3 = 1, RMH_VT_CSGNODE = 2 };
/* OrbitCameraControllerEvent (camera.rs:15-19) */
enum { RMH_EV_PAN = 0, RMH_EV_ORBIT = 1, RMH_EV_DOLLY = 2 };

RMH_API void rmh_reset(void);
RMH_API const char *rmh_last_error(void);

/* ---- CSGNode constructors (sphere.rs:8-13, box.rs:8-12, operations/mod.rs:9) -> node handle */
RMH_API int rmh_node_sphere(const float center[3], float radius);
RMH_API int rmh_node_box(const float center[3], const float radius[3]);
RMH_API int rmh_node_binary(int kind, int node_a, int node_b, float k);
/* N4 (reserved at csg/mod.rs:34,41-44; builder.rs:8,16-23).  Plane: the normal is normalised; RMH_E_PANIC for a zero
 * normal or a scale factor that is not finite and > 0.  Space transforms take ONE child and flatten to Push, child, Pop. */
RMH_API int rmh_node_plane(const float normal[3], float offset);
RMH_API int rmh_node_translation(int child, const float offset[3]);
RMH_API int rmh_node_rotation(int child, const float rotation_wijk[4]);       /* unit quaternion */
RMH_API int rmh_node_rotation_matrix(int child, const float m9_row_major[9]); /* object -> world */
RMH_API int rmh_node_scale(int child, float factor);
/* BuildCommands::build_commands into a fresh CSGCommandBufferBuilder (csg/mod.rs:16-19,
 * builder.rs:26-62).  Writes up to cap words; returns the word count; *cmd_count = builder.cmd_count. */
RMH_API int rmh_build_commands(int node, uint32_t *words, uint32_t cap, uint32_t *cmd_count);

/* ---- CSGNodeTemplateTrait (csg/mod.rs:21-26) */
RMH_API const char *rmh_template_name(int kind);
RMH_API int rmh_template_param_count(int kind);
/* name + default of input param i: tag, and value[3] (scalar in value[0]) */
RMH_API const char *rmh_template_param(int kind, int i, int *tag, float value[3]);
/* evaluate(HashMap<String, ValueType>): n params given as parallel arrays; node params by handle
 * (RMH_NONE = CSGNode(None)).  -> node handle, RMH_NONE, or RMH_E_PANIC (reference unwrap()). */
RMH_API int rmh_template_evaluate(int kind, int n, const char *const *names, const int *tags, const float *values3,
                                  const int *nodes);

/* ---- camera.rs */
RMH_API int rmh_orbit_new(const float target[3], float radius);              /* :38-50 -> controller handle */
RMH_API int rmh_orbit_update(int ctl, int event, float x, float y);          /* :62-84 */
RMH_API int rmh_orbit_state(int ctl, float out7[7]);                         /* target xyz, pitch, yaw, radius, 0 */
RMH_API int rmh_orbit_set(int ctl, float yaw, float pitch);
RMH_API int rmh_orbit_camera(int ctl, float position[3], float rotation_wijk[4]); /* :56-60 */
RMH_API int rmh_camera_view(const float position[3], const float rotation_wijk[4], float out16[16]); /* :8-13 */
/* Uniforms of RayMarchingCallback::prepare for this camera and viewport (renderer.rs:205-222) */
RMH_API int rmh_uniforms(const float position[3], const float rotation_wijk[4], float vw, float vh, uint8_t out144[144]);
RMH_API int rmh_limits_default(uint8_t out12[12]);                           /* renderer.rs:133-137 */

/* ---- csg_node_graph.rs (headless) */
RMH_API int rmh_graph_new(void);
RMH_API int rmh_graph_add_root(int graph);
RMH_API int rmh_graph_add_node(int graph, int kind);
RMH_API int rmh_graph_remove_node(int graph, int node);
RMH_API int rmh_graph_set_scalar(int graph, int node, const char *param, float v);
RMH_API int rmh_graph_set_vec3(int graph, int node, const char *param, const float v[3]);
RMH_API int rmh_graph_connect(int graph, int from_node, int to_node, const char *to_param);
RMH_API int rmh_graph_disconnect(int graph, int to_node, const char *to_param);
RMH_API int rmh_graph_evaluate_root(int graph);   /* :251-261 -> node handle or RMH_NONE */
RMH_API int rmh_graph_last_evaluations(int graph);

/* ---- renderer.rs: RayMarchingResources::new + RayMarchingCallback::{new, prepare, paint} */
RMH_API int rmh_resources_new(int cuda_device);   /* -> resources handle, RMH_E_DEVICE without a B200 */
RMH_API void *rmh_resources_ctx(int res);         /* the underlying rm_ctx* (rm_b200.h) */
/* node = RMH_NONE renders the empty scene (Option<CSGNode>::None, renderer.rs:225).
 * out_rgba (host, w*h*4 bytes) may be NULL to skip the read-back. */
RMH_API int rmh_callback_render(int res, int node, const float position[3], const float rotation_wijk[4], float vw,
                                float vh, uint8_t *out_rgba);
#ifdef __cplusplus
}
#endif
#endif
