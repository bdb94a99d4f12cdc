/*
 * rac.h — C ABI of librac_b200.so, the B200 (sm_100a) render-and-compare hot path.
 *
 * This is the drop-in boundary for the reference's L1 hot path
 * (olaals/end-to-end-RGB-pose-estimation-baseline).  The reference has no FFI
 * layer of its own: its boundary is a set of Python module-level functions
 * (SURVEY.md §8b).  Each entry point below names the reference function(s) it
 * replaces as `ref: file:line`.  The Python modules in
 * end-to-end-rgb-pose-estimation-baseline_b200/ (renderer.py, data_loaders.py,
 * rotation_representation.py, loss.py) bind these through ctypes and keep the
 * reference's signatures; INTEGRATION.md shows the stub a maintainer would add.
 *
 * Conventions
 *  - plain pointers and sizes only; no C++ / torch types cross the boundary.
 *  - pointers named d_* or documented "device" are CUDA device pointers owned
 *    by the caller; h_* are host pointers.  The library owns only mesh handles.
 *  - every compute call is asynchronous on `stream` (a cudaStream_t passed as
 *    void*; NULL = legacy default stream) and on the current CUDA device.
 *  - return 0 = ok, <0 = error (RAC_E_*); rac_last_error() gives a thread-local
 *    message.  No exception crosses the boundary.  There is no CPU fallback:
 *    without a usable CUDA device every compute call returns RAC_E_CUDA.
 *  - all matrices are row-major.  T_CO is the OpenCV-convention camera-from-object
 *    pose (x right, y down, z forward), exactly what the reference passes to
 *    render_scene (ref: renderer.py:85).
 */
#ifndef RAC_H_
#define RAC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RAC_API __attribute__((visibility("default")))

#define RAC_OK            0
#define RAC_E_INVALID   (-1)  /* bad argument (the Python layer maps to AssertionError/ValueError) */
#define RAC_E_CUDA      (-2)  /* CUDA runtime error, message in rac_last_error() */
#define RAC_E_WORKSPACE (-3)  /* workspace too small */
#define RAC_E_HANDLE    (-4)  /* unknown mesh handle */

typedef void*   rac_stream_t;   /* cudaStream_t */
typedef int32_t rac_mesh_t;     /* mesh handle: index into the library's device-resident mesh table */

/* rotation representations, ref: rotation_representation.py:147-155 */
#define RAC_ROT_6D  0   /* model_output (B,9):  6D Gram-Schmidt + vxvyvz */
#define RAC_ROT_SVD 1   /* model_output (B,12): 9D SVD orthogonalisation + vxvyvz */
#define RAC_ROT_3D  2   /* model_output (B,6):  axis-angle exp map + vxvyvz */

/* loss modes, ref: loss.py:5-33,81-122 */
#define RAC_LOSS_ADD_L1        0
#define RAC_LOSS_ADD_L2        1
#define RAC_LOSS_DISENTANGLED  2
#define RAC_LOSS_DISENT_SCALED 3

/* pyrender MetallicRoughnessMaterial subset used by mesh.frag. */
typedef struct rac_material {
    float base_color[4];   /* default (0.3,0.3,0.3,1) = pyrender default material for untextured trimesh */
    float metallic;        /* default 0.2 */
    float roughness;       /* default 0.8 */
} rac_material_t;

/* Scene constants of render_scene, ref: renderer.py:54-56,97 and pyrender camera defaults. */
typedef struct rac_scene {
    float ambient;          /* 0.1  ref: renderer.py:97 */
    float light_intensity;  /* 15.0 ref: renderer.py:54 */
    float inner_cone;       /* pi/9 ref: renderer.py:55 */
    float outer_cone;       /* pi/2 ref: renderer.py:56 */
    float znear;            /* 0.05 pyrender DEFAULT_Z_NEAR */
    float zfar;             /* 100  pyrender DEFAULT_Z_FAR  */
} rac_scene_t;

/* ---- diagnostics ------------------------------------------------------- */
RAC_API const char* rac_last_error(void);
RAC_API int  rac_version(void);
/* sm_count / compute capability of the current device; RAC_E_CUDA when no device. */
RAC_API int  rac_device_info(int* sm_count, int* cc_major, int* cc_minor);
/* number of kernels this library has launched in this process (bench.py's gpu_launches). */
RAC_API long long rac_launch_count(void);
RAC_API void rac_scene_defaults(rac_scene_t* out);
RAC_API void rac_material_defaults(rac_material_t* out);

/* ---- meshes ------------------------------------------------------------
 * ref: renderer.py:28-48 (add_object: tm.load + pyrender.Mesh.from_trimesh(smooth=False)).
 * Uploads an indexed triangle mesh once; precomputes unit face normals
 * normalize((v1-v0)x(v2-v0)) (flat shading).  h_uv (V,2),
 * h_tex (th,tw,3) RGB8 sRGB, h_vcolor (V,4) RGBA8 are optional (NULL).
 * mat NULL = pyrender default material (0.3 grey).  pyrender gives a vertex-coloured mesh without a material a WHITE
 * base colour (1,1,1,1; metallic 0.2, roughness 0.8) instead: the Python layer passes that explicitly (rac/mesh.py). */
RAC_API int rac_mesh_create(const float* h_verts, int V, const int32_t* h_faces, int T,
                            const float* h_uv, const uint8_t* h_tex, int tw, int th,
                            const uint8_t* h_vcolor, const rac_material_t* mat,
                            rac_mesh_t* out);
/* Multi-part mesh: what add_object builds from a .glb (renderer.py:36-42: every geometry of the trimesh Scene becomes
 * its own pyrender mesh, with the file's own glTF material, in ONE scene and one depth buffer).  The primitives are
 * merged into one indexed mesh; h_face_part (T bytes) names the part of every face, parts[k] its
 * pbrMetallicRoughness material (baseColorFactor, metallicFactor, roughnessFactor, baseColorTexture as RGB8, row 0 =
 * v 0) and whether COLOR_0 modulates it (pyrender applies vertex colours only to primitives WITHOUT a material).
 * h_uv (V,2) / h_vcolor (V,4) cover all vertices (zeros / 255 where a part has none). */
typedef struct rac_part {
    rac_material_t material;
    const uint8_t* tex_rgb8;    /* (tex_h, tex_w, 3) or NULL */
    int tex_w, tex_h;
    int use_vertex_color;
    int double_sided;           /* glTF material.doubleSided: no back-face culling for this part; back faces keep the
                                   geometric normal (pyrender's mesh.frag does not flip it), so they show ambient light only */
} rac_part_t;
RAC_API int rac_mesh_create_parts(const float* h_verts, int V, const int32_t* h_faces, int T,
                                  const float* h_uv, const uint8_t* h_vcolor,
                                  const uint8_t* h_face_part, int n_parts, const rac_part_t* parts,
                                  rac_mesh_t* out);
RAC_API int rac_mesh_destroy(rac_mesh_t mesh);
RAC_API int rac_mesh_info(rac_mesh_t mesh, int* V, int* T);

/* ---- step 1+2: render, and render fused with input assembly --------------
 * Workspace (device) for one call; the caller allocates it (e.g. a torch uint8
 * tensor) and may reuse it across calls on the same stream. */
RAC_API size_t rac_render_workspace_bytes(int B, int W, int H, int maxT);

/* ref: data_loaders.py:84-118 (render_batch / render_batch_sequential) over
 *      renderer.py:85-105 (render_scene) and the pyrender pipeline R1-R9.
 * h_mesh_ids: B handles (host).  d_T_CO (B,4,4) f32, d_K (B,3,3) f32 (device).
 * Outputs (device, each nullable):
 *   d_rgb_hwc (B,H,W,3) f32  max-normalised colour  (renderer.py:103-104, data_loaders.py:112-114)
 *   d_depth   (B,H,W)   f32  metric depth, 0 = miss (pyrender depth readback, R6)
 *   d_rgba8   (B,H,W,4) u8   the resolved 8-bit framebuffer before normalisation (render-only sweep) */
RAC_API int rac_render_batch(const int32_t* h_mesh_ids, const float* d_T_CO, const float* d_K,
                             int B, int W, int H, const rac_scene_t* scene,
                             float* d_rgb_hwc, float* d_depth, uint8_t* d_rgba8,
                             void* d_ws, size_t ws_bytes, rac_stream_t stream);

/* ref: render_batch (above) + data_loaders.py:158-177 (prepare_model_input) with
 *      normalize_img (:79-82) on the gt crop and normalize_depth (:69-77).
 * d_gt_hwc (B,H,W,3) f32 in [0,1].  d_x_nchw (B,C,H,W) f32 with C = 6 or 7 and
 * channel order [norm_depth (if use_norm_depth), render R,G,B, normalised gt R,G,B].
 * d_depth_out (B,H,W) nullable: the un-normalised metric depth.
 * d_rgb_hwc_out (B,H,W,3) nullable: the max-normalised render in HWC (what render_batch returns). */
RAC_API int rac_render_concat(const int32_t* h_mesh_ids, const float* d_T_CO, const float* d_K,
                              int B, int W, int H, const rac_scene_t* scene,
                              const float* d_gt_hwc, int use_norm_depth,
                              float* d_x_nchw, float* d_depth_out, float* d_rgb_hwc_out,
                              void* d_ws, size_t ws_bytes, rac_stream_t stream);

/* ref: data_loaders.py:158-177 (prepare_model_input) alone, for renders that already
 * exist (iteration 0 uses the dataset's init render, train_model.py:236-243).
 * d_init_hwc (B,H,W,3) f32, d_gt_hwc (B,H,W,3) f32, d_depth (B,H,W) f32 (nullable unless use_norm_depth).
 * d_gt_hwc == NULL: the gt planes of x are left untouched (the caller keeps the normalised crop of a batch across
 * its refinement iterations, train_model.py:234-250, and copies it in itself).
 * d_stats_ws: device scratch of rac_assemble_workspace_bytes(B) bytes. */
RAC_API size_t rac_assemble_workspace_bytes(int B);
RAC_API int rac_assemble_input(const float* d_init_hwc, const float* d_gt_hwc, const float* d_depth,
                               int B, int W, int H, int use_norm_depth, float* d_x_nchw,
                               void* d_stats_ws, size_t ws_bytes, rac_stream_t stream);

/* No counterpart in the reference: this exists for hosts of the library, not for a reference function.
 * It is rac_assemble_input (above) without the d_stats_ws argument, for callers that are not in a position to
 * allocate device memory: code running where a cudaMalloc would be refused or would stall — inside a real-time or
 * serving loop that forbids allocation after start-up, under a framework that owns the whole GPU heap through its own
 * pool and hands out no raw allocations to plug-ins, or in a process whose memory budget is fixed and audited up front.
 * The per-image statistics scratch is a fixed array inside the library's device image instead
 * (RAC_ASSEMBLE_STATIC_MAX_B records of 64 B = 64 KiB per GPU that loads the library, held until the process exits).
 * The call itself allocates nothing on host or device; inputs and outputs are still the caller's.
 *  - B > RAC_ASSEMBLE_STATIC_MAX_B returns RAC_E_INVALID with a message that names the limit; split the batch.
 *  - The first call on each device looks up the array's device address (a one-time set-up that may load the module
 *    and synchronise).  Make that first call during start-up, before any stream capture begins on that device.
 *  - All calls on one GPU share the one array and it keeps whatever the last call left in it; every call begins by
 *    clearing the B records it uses, in stream order.  Two calls whose GPU work can overlap (different streams
 *    without an ordering between them, or two host threads) race on it and produce wrong depth planes — order them.
 * Results are identical to rac_assemble_input's on the same inputs. */
#define RAC_ASSEMBLE_STATIC_MAX_B 1024
RAC_API int rac_assemble_input_static(const float* d_init_hwc, const float* d_gt_hwc, const float* d_depth,
                                      int B, int W, int H, int use_norm_depth, float* d_x_nchw, rac_stream_t stream);

/* ---- step 3+4: rotation head + SE(3) update ------------------------------
 * ref: rotation_representation.py:131-184 (calculate_T_CO_pred) over
 *      :4-19 (6D), :22-39 (SVD), :41-118 (exp map).
 * d_y (B,D) f32 with D = 9/12/6 for 6D/SVD/3D.  d_T_in (B,4,4) f32, or f64 when
 * t_in_is_f64 != 0 (the translation update then runs in f64).
 * fx, fy: intrinsics of the FIRST sample (Q2).  d_T_out (B,4,4) f32.
 * d_saved: RAC_POSE_SAVED_FLOATS floats per sample, consumed by the backward. */
#define RAC_POSE_SAVED_FLOATS 48
RAC_API int rac_pose_update_fwd(const float* d_y, int rot_repr, const void* d_T_in, int t_in_is_f64,
                                double fx, double fy, int B,
                                float* d_T_out, float* d_saved, rac_stream_t stream);
/* d_gT_out (B,4,4) f32 = dL/dT_out  ->  d_gy (B,D) f32 = dL/dy. */
RAC_API int rac_pose_update_bwd(const float* d_gT_out, const float* d_saved, int rot_repr, int B,
                                float* d_gy, rac_stream_t stream);
/* Rotation heads on their own (public functions of rotation_representation.py).
 * d_in (B,6|9|3) -> d_R (B,3,3); saved: RAC_POSE_SAVED_FLOATS per sample. */
RAC_API int rac_rotation_fwd(const float* d_in, int rot_repr, float eps, int B,
                             float* d_R, float* d_saved, rac_stream_t stream);
RAC_API int rac_rotation_bwd(const float* d_gR, const float* d_saved, int rot_repr, int B,
                             float* d_gin, rac_stream_t stream);

/* ---- step 5: point-matching losses ----------------------------------------
 * ref: loss.py:5-18 (ADD-L1), :20-33 (ADD-L2), :81-97 (disentangled), :101-122 (scaled).
 * d_T_a, d_T_b (B,4,4) f32: the two poses in the REFERENCE'S ARGUMENT ORDER:
 *   L1/L2:        a = TCO_gt,    b = TCO_pred          (loss.py:5,20)
 *   DISENTANGLED: a = T_CO_pred, b = T_CO_gt           (loss.py:81; the parts of a are pasted into clones of b)
 *   DISENT_SCALED:a = T_CO_pred, b = T_CO_gt, d_T_prev = T_CO_pred_prev (loss.py:101)
 * d_pts (B,N,3) f32.  d_per_sample (B) f32; d_mean (1) f32 = batch mean (deterministic order).
 * d_grad_a, d_grad_b (B,4,4) f32, nullable: d(per_sample[b])/dT_a[b], d(per_sample[b])/dT_b[b],
 * written by the forward so the backward is a scale by the upstream gradient.
 * d_counter: one zero-initialised uint32 of device scratch (reset by the kernel). */
RAC_API int rac_add_loss_fwd(const float* d_T_a, const float* d_T_b, const float* d_T_prev,
                             const float* d_pts, int B, int N, int mode,
                             float* d_per_sample, float* d_mean,
                             float* d_grad_a, float* d_grad_b,
                             unsigned int* d_counter, rac_stream_t stream);
/* g_T[b] = d_w[b] * d_grad[b] (w = upstream gradient per sample); out (B,4,4). */
RAC_API int rac_add_loss_bwd(const float* d_grad, const float* d_w, int B, float* d_gT, rac_stream_t stream);

/* The whole post-CNN half of one refinement iteration in one launch, for callers that own their loop
 * (rac.RenderCompareStep): ref train_model.py:245-247,276-285 / test_model.py:101-103 — calculate_T_CO_pred, the loss,
 * its batch mean and backward down to the CNN's output.  d_y (B,D) head output; d_T_in (B,4,4) f32 or f64 (Q6);
 * d_T_gt (B,4,4); d_T_prev (B,4,4) for RAC_LOSS_DISENT_SCALED, else NULL; d_pts (B,N,3).
 * pred_is_a: which argument slot of rac_add_loss_fwd the NEW pose takes (0: T_a = T_gt, T_b = T_new — ADD-L1/L2 and the
 * disentangled loss in train_model.py:280's swapped order; 1: T_a = T_new, T_b = T_gt — the scaled loss, loss.py:101).
 * Outputs: d_T_new (B,4,4), d_per_sample (B), d_mean (1, nullable), d_grad_y (B,D) = d(mean loss)/d(y);
 * d_T_feedback (B,4,4, nullable): T_new is stored there as well — the loop's `T_CO_pred = T_CO_pred_new`
 * (train_model.py:250); it MAY alias an f32 d_T_in and d_T_prev (the old pose is read before it is overwritten).
 * Same per-sample arithmetic as the four stand-alone launches.  d_counter: as for rac_add_loss_fwd. */
RAC_API int rac_pose_loss_step(const float* d_y, int rot_repr, const void* d_T_in, int t_in_is_f64, double fx, double fy,
                               const float* d_T_gt, const float* d_T_prev, const float* d_pts, int B, int N,
                               int loss_mode, int pred_is_a, float* d_T_new, float* d_per_sample, float* d_mean,
                               float* d_grad_y, float* d_T_feedback, unsigned int* d_counter, rac_stream_t stream);

/* ref: loss.py:60-78 (transform_pts): d_T (B,M,4,4) (M = 1 for the 3-dim case),
 * d_pts (B,N,3) -> d_out (B,M,N,3). */
RAC_API int rac_transform_pts_fwd(const float* d_T, const float* d_pts, int B, int M, int N,
                                  float* d_out, rac_stream_t stream);
/* d_gout (B,M,N,3) -> d_gT (B,M,4,4) (rows 0-2 filled, row 3 zero), d_gpts (B,N,3); either nullable. */
RAC_API int rac_transform_pts_bwd(const float* d_gout, const float* d_T, const float* d_pts,
                                  int B, int M, int N, float* d_gT, float* d_gpts, rac_stream_t stream);

/* ---- §8(f) N3: batched pose sampling ---------------------------------------
 * ref: se3_helpers.py:76-88 (get_T_CO_init_and_gt) with :39-74 (random rotation + translation, small perturbation,
 * camera on +x looking at the origin), data_loaders.py:58-67 (sample_T_CO_inits_and_gts: a Python loop over B).
 * The caller draws the variates the reference draws per sample, in the reference's order, and this call composes the
 * two poses for the whole batch in one launch (f64 arithmetic like numpy, f32 row-major 4x4 out):
 *   d_u01 (B,12) f32 uniforms in [0,1): [theta_gt, axis_gt x3, theta_delta, t_delta x3, axis_delta x3, cam distance]
 *   d_n01 (B,3)  f32 standard normals:  gt translation
 * distance / distance_dev / gt_transl_dev / transl_dev / angle_dev_deg: the scene_config entries
 * (configs/example_config.py:45-52). */
RAC_API int rac_sample_poses(const float* d_u01, const float* d_n01, int B, double distance, double distance_dev,
                             double gt_transl_dev, double transl_dev, double angle_dev_deg,
                             float* d_T_init, float* d_T_gt, rac_stream_t stream);

/* ---- §8(f) N4: silhouette ring and overlay ----------------------------------
 * ref: visualization.py:19-31 (create_silhouette): mask = mean(img,2) > 0; ring = clip(box5x5(mask),0,1) - mask with
 * zero padding.  d_img_hwc (B,H,W,3) f32.  silhouette_only != 0: d_out (B,H,W) f32 ring mask; otherwise d_out
 * (B,H,W,3) f32 = the image with ring pixels painted h_colour3 (host pointer to 3 floats, NULL = red (1,0,0)). */
RAC_API int rac_silhouette(const float* d_img_hwc, int B, int W, int H, int silhouette_only, const float* h_colour3,
                           float* d_out, rac_stream_t stream);
/* ref: visualization.py:34-37 (blend_imgs): uint8(im*255) of both (truncation), PIL Image.blend
 * (uint8)(a + alpha (b - a)), / 255.  n elements, f32 in and out.  quantized_inputs != 0: the inputs already hold the
 * 8-bit codes 0..255 (the Python wrapper quantises f64 inputs on the host, in f64 as numpy would). */
RAC_API int rac_blend(const float* d_im1, const float* d_im2, float alpha, int quantized_inputs, size_t n,
                      float* d_out, rac_stream_t stream);

/* ---- 8-bit gt crops ----------------------------------------------------------
 * ref: image_dataloaders.py:79-83 (gt = real.png / 255.0, f64) followed by data_loaders.py:79-82 (normalize_img) and the
 * f32 cast + HWC->CHW of data_loaders.py:163-166, computed from the raw 8-bit codes so that the crop crosses PCIe at
 * 3 B/px instead of 12.  d_gt_hwc_u8 (B,H,W,3) u8 -> the three gt planes (channels C-3..C-1, C = 7 if use_norm_depth
 * else 6) of d_x_nchw (B,C,H,W) f32; the other planes are not touched (rac_assemble_input / rac_render_concat with
 * d_gt_hwc == NULL fill them).  Exact integer statistics, every output value = (float)((q/255.0 - mean)/std) evaluated
 * in f64: bit-identical to the reference's cast.  d_stats_ws: rac_gt_u8_workspace_bytes(B) bytes, 8-byte aligned. */
RAC_API size_t rac_gt_u8_workspace_bytes(int B);
RAC_API int rac_gt_norm_u8(const unsigned char* d_gt_hwc_u8, int B, int W, int H, int use_norm_depth, float* d_x_nchw,
                           void* d_stats_ws, size_t ws_bytes, rac_stream_t stream);

/* Host-side staging for the crop the unchanged loader produces (image_dataloaders.py:79-83: real.png / 255.0 — float64,
 * pageable, 24 B/px for 3 B/px of information; handed to prepare_model_input at train_model.py:243 / test_model.py:100).
 * One multi-threaded pass: h_dst[i] = q = rint(h_src[i] * 255) and *all_exact = 1 iff EVERY value equals q / 255.0 exactly
 * (f64 source) or (float)(q / 255.0) (f32 source).  h_src: n floats or doubles (host); h_dst: n bytes (host, typically
 * pinned).  When all_exact, h_dst feeds rac_gt_norm_u8 and the result is bit-identical to normalising the floats the
 * reference's way; otherwise the caller uploads the floats.  n_threads <= 0: hardware concurrency, at most 16.
 * No device work, no stream: data staging only (every normalisation still runs on the GPU). */
RAC_API int rac_host_pack_gt_u8(const void* h_src, int src_is_f64, size_t n, unsigned char* h_dst, int n_threads,
                                int* all_exact);

#ifdef __cplusplus
}
#endif
#endif /* RAC_H_ */
