/*
 * cloth_splatting_b200.h -- C-ABI of the B200-native (sm_100a) hot path of
 * KTH-RPL/cloth-splatting.
 *
 * Every entry point takes raw device pointers, sizes, a CUDA stream (passed as
 * void* so that no CUDA header is needed to bind it) and caller-provided
 * workspaces.  The library allocates nothing, keeps no state between calls and
 * is re-entrant per stream.  Return value: 0 = ok, CSB_ERR_* (<0) = usage error,
 * >0 = cudaError_t of the failing launch.
 *
 * Reference interfaces replaced (paths relative to the reference repo,
 * RAST = submodules/depth-diff-gaussian-rasterization, KNN = submodules/simple-knn):
 *   csb_rasterize_forward   <- CudaRasterizer::Rasterizer::forward   RAST/cuda_rasterizer/rasterizer.h:31-54
 *                              (pybind: rasterize_gaussians          RAST/ext.cpp:16, rasterize_points.cu:35-117)
 *   csb_rasterize_backward  <- CudaRasterizer::Rasterizer::backward  RAST/cuda_rasterizer/rasterizer.h:56-85
 *                              (pybind: rasterize_gaussians_backward RAST/ext.cpp:17, rasterize_points.cu:119-198)
 *   csb_mark_visible        <- CudaRasterizer::Rasterizer::markVisible RAST/cuda_rasterizer/rasterizer.h:24-29
 *                              (pybind: mark_visible                 RAST/ext.cpp:18, rasterize_points.cu:200-219)
 *   csb_dist_cuda2          <- SimpleKNN::knn                        KNN/simple_knn.h:15-19
 *                              (pybind: distCUDA2                    KNN/ext.cpp:15-17, spatial.cu:15-26)
 *   csb_gnn_*               <- EncodeProcessDecode.forward           meshnet/graph_network.py:387-408
 *                              (PyG MessagePassing.propagate + torch addmm/layer_norm/scatter_add_)
 *   csb_mesh_xyz_*          <- MultiGaussianMesh.get_xyz             scene_reconstruction/gaussian_mesh.py:151-169
 *   csb_image_loss_*        <- image_losses / l1_loss / ssim         scene_reconstruction/train_utils.py:50-74, utils/loss_utils.py:20-70
 *   csb_adam_step           <- torch.optim.Adam(...).step()          scene_reconstruction/gaussian_mesh.py:121-135, train_utils.py:312-317
 */
#ifndef CLOTH_SPLATTING_B200_H
#define CLOTH_SPLATTING_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CSB_OK 0
#define CSB_ERR_BAD_ARG (-1)
#define CSB_ERR_WORKSPACE (-2) /* a caller-provided workspace is too small */
#define CSB_ERR_UNSUPPORTED (-3)
#define CSB_ERR_STATIC_LIMIT (-4) /* the problem needs more scratch than a no-allocation entry point's fixed array */

#define CSB_MAX_VIEWS 16 /* views rendered by one launch set */

/* ABI version; bumped whenever a signature changes. */
int csb_abi_version(void);
/* Human readable string for a return code of this library. */
const char* csb_error_string(int code);

/* Kernel launches issued by this library since it was loaded (all entry points). */
unsigned long long csb_launch_count(void);
/* Per-kernel device timing for bench.py's roofline: while enabled, every launch is bracketed by CUDA
 * events on its stream.  csb_profile_enable clears earlier records.  csb_profile_read sums the
 * device time (ms) and the launch count of the kernel `name` (NULL = all kernels) recorded so far,
 * waiting for them to finish.  csb_profile_names writes the comma-separated kernel names seen. */
void csb_profile_enable(int on);
int csb_profile_read(const char* name, double* total_ms, unsigned long long* launches);
int csb_profile_names(char* out, size_t out_bytes);
/* Debugging / A-B timing knob (process-wide, like the profiler switch): which compositing kernels the rasterizer entry
 * points launch.  Low nibble = forward kernel (0 = one pixel per lane, 1 = one vertical pixel pair per lane, 2 = two
 * pairs, 3 / 4 = one / two pairs without per-pixel branches, 15 = 3 on the per-tile binning path and 4 on the global
 * sort path: the default), high nibble = backward kernel (0 = one pixel per lane, 1 / 2 = one / two pairs per lane, 3 = one pair under a
 * register cap, 4 / 5 / 6 = 1 / 3 / 2 with the delayed warp reduction, 7 = 4 with the warp sums taken through shared memory).  All forward kernels produce the same per-pixel
 * bits.  Returns the previous value; a negative or out-of-range value only queries. */
int csb_debug_render_variant(int variant);
/* Debugging knob: 0 switches off the exact ellipse-vs-quad test that lets the compositing kernels skip the
 * (Gaussian, 8x8 pixel quad) visits in which no pixel can reach alpha >= 1/255 (bounding-box test only, the state
 * before it); results are identical either way.  Returns the previous value; any other argument only reads it. */
int csb_debug_cull_exact(int on);
/* A-B switch of the GNN forward (process-wide): 1 (default) = graphs whose edge and node tiles all fit on the GPU at once
 * run as ONE persistent kernel (gnn_small_kernel), 0 = always one launch per role and step.  Returns the previous value;
 * any other argument only queries. */
int csb_debug_gnn_small(int on);
/* Debugging aid: thread 0 of CTA `block` of every GNN forward kernel stamps %globaltimer at the stations of its tile walk
 * into device_buffer (u64 words: [0] = count, then time_ns << 8 | station; room for 8000 stamps).  NULL switches it off. */
int csb_debug_gnn_trace(void* device_buffer, int block);
/* Micro-benchmarks of the FP32 FMA pipe (scalar FFMA and packed FFMA2, FMAs per second) and of the SFU (MUFU.EX2 per
 * second) on the current device: the denominators of bench.py's compositing roofline, measured in the same
 * run as the kernels.  `scratch`: at least 8 * 256 * 4 bytes per SM of device memory.  Synchronises the stream. */
int csb_measure_fp32_peaks(void* scratch, size_t scratch_bytes, double* ffma_per_s, double* ffma2_per_s,
                           double* mufu_per_s, void* stream);

/* ------------------------------------------------------------------------- *
 * Rasterizer
 * ------------------------------------------------------------------------- */

/* One camera view.  The four pointers are DEVICE pointers, exactly the tensors
 * GaussianRasterizationSettings carries (RAST/diff_gaussian_rasterization/__init__.py:157-169):
 * viewmatrix / projmatrix are the 16 floats of the transposed 4x4 matrices, consumed as
 * out_k = m[k] x + m[4+k] y + m[8+k] z + m[12+k]  (RAST/cuda_rasterizer/auxiliary.h:58-77). */
typedef struct csb_view {
    const float* viewmatrix; /* f32[16] device */
    const float* projmatrix; /* f32[16] device */
    const float* campos;     /* f32[3]  device */
    const float* background; /* f32[3]  device */
    float tan_fovx, tan_fovy;
} csb_view;

/* Sizes of the three caller-owned byte workspaces (the reference's geomBuffer,
 * imgBuffer and binningBuffer, rasterize_points.cu:74-79).  `capacity` is the number
 * of (Gaussian, tile) instances the binning workspace can hold. */
size_t csb_rasterize_geom_bytes(int num_views, int P);
size_t csb_rasterize_image_bytes(int num_views, int width, int height);
size_t csb_rasterize_binning_bytes(int num_views, int P, int width, int height, size_t capacity);
/* Inverse of the above: the largest capacity whose binning workspace fits in `bytes` (the backward
 * pass derives the layout from the size of the buffer the forward pass was given). */
size_t csb_rasterize_binning_capacity(int num_views, int P, int width, int height, size_t bytes);

/* Forward for `num_views` (1..CSB_MAX_VIEWS) views of the same P Gaussians and the same
 * image size.  With num_views == 1 this is a drop-in for Rasterizer::forward.
 *
 *   means3D f32[P,3]; shs f32[P,M,3] or NULL; colors_precomp f32[P,3] or NULL;
 *   opacities f32[P]; scales f32[P,3] + rotations f32[P,4] (r,x,y,z; NOT normalised,
 *   forward.cu:127) or cov3D_precomp f32[P,6].
 *   out_color f32[V,3,H,W]; out_depth f32[V,1,H,W]; radii i32[V,P].
 *
 * The number of instances is only known on the device.  It is written to
 * num_rendered_dev[0] (u32, device) and, if num_rendered_host is not NULL, copied
 * asynchronously on `stream` to that (pinned) host word.  If it exceeds
 * `binning_capacity` nothing is rendered and the caller must grow the binning
 * workspace to csb_rasterize_binning_bytes(..., num_rendered) and call again
 * (the reference does the same resize after its blocking cudaMemcpy,
 * rasterizer_impl.cu:280-286).  flags: bit0 = also write the sorted u64
 * tile|depth keys (rasterizer_impl.cu:70-111 layout) for inspection;
 * bit1 = num_rendered_host is device-accessible pinned memory (cudaHostAlloc under UVA): the
 * binning kernel stores the count (clamped to 0xFFFFFFFE) into it as soon as it is known --
 * before the tile sort and the render kernel run -- and no copy is enqueued at the end.  A
 * caller that pre-sets the word to 0xFFFFFFFF can poll it instead of synchronising the stream,
 * and enqueue the backward pass while the forward is still rendering;
 * bit2 = per-tile binning for scenes with short tile lists: instead of the depth sort / emission /
 * tile sort (7-8 launches) the entries are counted into their tiles, scanned, filled and sorted per tile
 * in shared memory (4 launches, same resulting order bit for bit).  If a tile holds more than 2048
 * instances nothing is rendered and the count word reads 0xFFFFFFFD: call again without bit2. */
int csb_rasterize_forward(int num_views, const csb_view* views, int P, int D, int M, int width, int height,
                          const float* means3D, const float* shs, const float* colors_precomp,
                          const float* opacities, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp, int prefiltered,
                          void* geom_ws, size_t geom_bytes, void* image_ws, size_t image_bytes,
                          void* binning_ws, size_t binning_bytes, size_t binning_capacity,
                          float* out_color, float* out_depth, int* radii, uint32_t* num_rendered_dev,
                          uint32_t* num_rendered_host, int flags, void* stream);

/* Backward of the colour output (the depth output has no backward in the reference,
 * RAST/diff_gaussian_rasterization/__init__.py:101-129).  Needs the three workspaces of
 * the matching forward call untouched.  Gradients are summed over the views, except
 * dL_dmeans2D f32[V,P,3] which the caller consumes per view
 * (scene_reconstruction/train_utils.py:290-292).
 *   dL_dpix f32[V,3,H,W];
 *   outputs (all overwritten, no pre-zeroing needed): dL_dmeans3D f32[P,3],
 *   dL_dcolors f32[P,3] (only meaningful with colors_precomp), dL_dopacity f32[P],
 *   dL_dcov3D f32[P,6], dL_dsh f32[P,M,3] (may be NULL when shs is NULL),
 *   dL_dscales f32[P,3], dL_drotations f32[P,4]. */
int csb_rasterize_backward(int num_views, const csb_view* views, int P, int D, int M, int width, int height,
                           const float* means3D, const float* shs, const float* colors_precomp,
                           const float* scales, float scale_modifier, const float* rotations,
                           const float* cov3D_precomp, const int* radii, const void* geom_ws,
                           const void* image_ws, const void* binning_ws, size_t binning_capacity,
                           const float* dL_dpix, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                           float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                           float* dL_drotations, void* stream);

/* The same backward when dL_dpix arrives view by view (uploaded over PCIe on another stream, or produced by a
 * per-view loss): view_ready[v] is a cudaEvent_t (or NULL = already there) recorded after view v's plane
 * dL_dpix[v] is complete.  View v is composited backwards as soon as its event has fired, the per-Gaussian stage
 * runs after the last view -- the transfer of the later views overlaps the compositing of the earlier ones.
 * Results are bit-identical to csb_rasterize_backward's up to the order of the atomic sums it already has. */
int csb_rasterize_backward_staged(int num_views, const csb_view* views, int P, int D, int M, int width, int height,
                                  const float* means3D, const float* shs, const float* colors_precomp,
                                  const float* scales, float scale_modifier, const float* rotations,
                                  const float* cov3D_precomp, const int* radii, const void* geom_ws,
                                  const void* image_ws, const void* binning_ws, size_t binning_capacity,
                                  const float* dL_dpix, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                                  float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                                  float* dL_drotations, void* const* view_ready, void* stream);

/* present[i] = (viewmatrix * means3D[i]).z > 0.2   (auxiliary.h:139-164) */
int csb_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream);

/* Device pointers of the arrays carved out of the workspaces, for parity tests
 * (counterpart of GeometryState/ImageState/BinningState::fromChunk, rasterizer_impl.h:21-73).
 * names: "means2D" f32x2[V,P]; "conic_opacity" f32x4[V,P]; "rgb_depth" f32x4[V,P] (r,g,b,view z);
 *        "cov3D" f32[P,6]; "clamped" u8[V,P] (bit c = channel c clamped); "tiles_touched" u32[V,P];
 *        "final_T" f32[V,H*W]; "n_contrib" u32[V,H*W]; "ranges" u32x2[V*tiles];
 *        "point_list" u32[R]; "point_tiles" u32[R]; "keys" u64[R] (flags bit0 only).
 * Returns NULL for an unknown name. */
void* csb_rasterize_state_ptr(const char* name, int num_views, int P, int width, int height,
                              size_t binning_capacity, void* geom_ws, void* image_ws, void* binning_ws);

/* ------------------------------------------------------------------------- *
 * simple-knn
 * ------------------------------------------------------------------------- */
size_t csb_dist_cuda2_bytes(int P);
/* mean_dists[i] = mean of the 3 smallest squared distances from points[i] to the
 * other points (KNN/simple_knn.cu:147-183).  points f32[P,3], mean_dists f32[P]. */
int csb_dist_cuda2(int P, const float* points, float* mean_dists, void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------- *
 * meshnet GNN (EncodeProcessDecode.forward, meshnet/graph_network.py:387-408)
 * ------------------------------------------------------------------------- */

/* Constructor arguments of EncodeProcessDecode (meshnet/graph_network.py:336-345).  Supported:
 * latent_dim == mlp_hidden_dim == 128, nmlp_layers == 2 (the reference's defaults,
 * train_meshnet_sim.py:645-650), nnode_in / nedge_in <= 16, nnode_out <= 8, steps <= 30;
 * anything else returns CSB_ERR_UNSUPPORTED. */
typedef struct csb_gnn_dims {
    int nnode_in, nnode_out, nedge_in, latent_dim, nmessage_passing_steps, nmlp_layers, mlp_hidden_dim;
} csb_gnn_dims;

/* Parameters are handed over as ONE flat fp32 device buffer, MLP after MLP in the order
 *   _encoder.node_fn, _encoder.edge_fn, then for every step m: _processor.gnn_stacks.m.edge_fn,
 *   _processor.gnn_stacks.m.node_fn, then _decoder.node_fn;
 * each MLP as NN-0.weight [128,in] (row major, nn.Linear layout), NN-0.bias, NN-1.weight, NN-1.bias,
 * NN-2.weight [out,128], NN-2.bias, then (all but the decoder) LayerNorm weight, bias.
 * csb_gnn_pack converts it into the bf16 / swizzled image the kernels keep resident in shared memory
 * (also folds the reference's e_m = 2^m e_0 edge-latent doubling, graph_network.py:173-176, into the
 * edge block of every step's first edge Linear). */
size_t csb_gnn_param_floats(const csb_gnn_dims* dims);
size_t csb_gnn_packed_bytes(const csb_gnn_dims* dims);
int csb_gnn_pack(const csb_gnn_dims* dims, const float* flat_params, void* packed, void* stream);

/* Static topology: edge_index i64[2,E] (row 0 = src, row 1 = dst; torch_geometric flow
 * source_to_target, messages are summed at dst, graph_network.py:136,197).  Writes the stable
 * dst-sorted edge order: dst_sorted[i], src_sorted[i] and perm[i] = original edge id of slot i. */
size_t csb_gnn_graph_workspace_bytes(int E);
int csb_gnn_prepare_graph(int N, int E, const int64_t* edge_index, int32_t* dst_sorted, int32_t* src_sorted,
                          int32_t* perm, void* ws, size_t ws_bytes, void* stream);

/* y[N, nnode_out] = EncodeProcessDecode(x[N, nnode_in], edge_index, edge_features[E, nedge_in]).
 * edge_features are in the ORIGINAL edge order (perm is applied inside).  x_latent_out (optional,
 * f32[N,128]) receives the node latents after the last message-passing step (for parity tests). */
size_t csb_gnn_workspace_bytes(const csb_gnn_dims* dims, int N, int E);
int csb_gnn_forward(const csb_gnn_dims* dims, const void* packed, int N, int E, const float* x,
                    const float* edge_features, const int32_t* dst_sorted, const int32_t* src_sorted,
                    const int32_t* perm, float* y, float* x_latent_out, void* ws, size_t ws_bytes, void* stream);

/* Training (the reference back-propagates `loss.backward()` through EncodeProcessDecode in its multi-step
 * loop, train_meshnet_sim.py:491-532; torch autograd over graph_network.py:98-111,178-222,323-332).
 * csb_gnn_forward_train = csb_gnn_forward that keeps, in the caller-owned `state`, the bf16 operands every
 * message-passing step consumed (node latents x_m, aggregated messages agg_m, encoded edge latents e_0).
 * csb_gnn_backward recomputes the MLP activations from them and ACCUMULATES (+=)
 *   grad_flat          f32[csb_gnn_param_floats]  same layout as flat_params (caller zeroes it),
 * and writes (each may be NULL)
 *   dL_dx              f32[N, nnode_in],
 *   dL_dedge_features  f32[E, nedge_in]  in the ORIGINAL edge order,
 * given dL_dy f32[N, nnode_out].  x / edge_features / graph arrays / packed must be the forward's. */
size_t csb_gnn_train_state_bytes(const csb_gnn_dims* dims, int N, int E);
int csb_gnn_forward_train(const csb_gnn_dims* dims, const void* packed, int N, int E, const float* x,
                          const float* edge_features, const int32_t* dst_sorted, const int32_t* src_sorted,
                          const int32_t* perm, float* y, void* state, size_t state_bytes, void* stream);
size_t csb_gnn_backward_workspace_bytes(const csb_gnn_dims* dims, int N, int E);
int csb_gnn_backward(const csb_gnn_dims* dims, const void* packed, int N, int E, const float* x,
                     const float* edge_features, const int32_t* dst_sorted, const int32_t* src_sorted,
                     const int32_t* perm, const void* state, size_t state_bytes, const float* dL_dy,
                     float* grad_flat, float* dL_dx, float* dL_dedge_features, void* ws, size_t ws_bytes,
                     void* stream);

/* ------------------------------------------------------------------------- *
 * mesh-anchored Gaussians (MultiGaussianMesh.get_xyz, scene_reconstruction/gaussian_mesh.py:151-169)
 * ------------------------------------------------------------------------- */
/* xyz[p] = sum_k nb[p,k] * vertices[face[k, face_ids[p]]],  nb = face_bary / sum(face_bary, 1).
 * vertices f32[Nv,3]; face i64[3,F] (torch_geometric Data.face layout); face_ids i64[P];
 * face_bary f32[P,3]; xyz f32[P,3]. */
int csb_mesh_xyz_forward(int P, int Nv, int F, const float* vertices, const int64_t* face, const int64_t* face_ids,
                         const float* face_bary, float* xyz, void* stream);
/* Backward of the above.  dL_dvertices f32[Nv,3] (zeroed inside, then accumulated) and dL_dface_bary
 * f32[P,3] may each be NULL when that gradient is not needed. */
int csb_mesh_xyz_backward(int P, int Nv, int F, const float* vertices, const int64_t* face, const int64_t* face_ids,
                          const float* face_bary, const float* dL_dxyz, float* dL_dvertices, float* dL_dface_bary,
                          void* stream);

/* MultiGaussianMesh.get_rotation(deformed_vertices) (scene_reconstruction/gaussian_mesh.py:171-188):
 *   out[p] = quat_product(normalize(rotation[p]), rotmat_to_unitquat(Kabsch(rest[face_p] -> vertices[face_p])))
 * quaternions in roma's XYZW layout, as the reference composes them.  rest_vertices / vertices f32[Nv,3];
 * face i64[3,F]; face_ids i64[P]; rotation f32[P,4] (raw parameter); out f32[P,4].
 * Backward: dL_dvertices f32[Nv,3] (zeroed inside, then accumulated) and dL_drotation f32[P,4]; either may be NULL. */
int csb_mesh_rotation_forward(int P, int Nv, int F, const float* rest_vertices, const float* vertices,
                              const int64_t* face, const int64_t* face_ids, const float* rotation, float* out,
                              void* stream);
int csb_mesh_rotation_backward(int P, int Nv, int F, const float* rest_vertices, const float* vertices,
                               const int64_t* face, const int64_t* face_ids, const float* rotation,
                               const float* dL_dout, float* dL_dvertices, float* dL_drotation, void* stream);

/* The fused mesh -> Gaussian transform of a refinement step: get_xyz AND get_rotation
 * (scene_reconstruction/gaussian_mesh.py:151-188) for `num_views` deformed meshes in one forward launch.  The deformed
 * vertices of view v are vertices[v] + residual[v] (meshnet/meshnet_network.py:361-373: mesh prediction + the residual
 * simulator's output, consumed in place); residual may be NULL.  vertices / residual f32[V,Nv,3]; rest_vertices
 * f32[Nv,3]; face i64[3,F]; face_ids i64[P]; face_bary f32[P,3]; rotation f32[P,4] (raw parameter).
 * Outputs xyz f32[V,P,3], rot f32[V,P,4] (XYZW as the reference composes them); one of them may be NULL.
 * Backward: dL_dxyz f32[V,P,3] and / or dL_drot f32[V,P,4] in; dL_dvertices f32[V,Nv,3] (zeroed inside; it is the
 * gradient of `vertices` and of `residual` alike), dL_dface_bary f32[P,3] and dL_drotation f32[P,4] (both summed over the
 * views in view order); any output may be NULL.  workspace: csb_mesh_transform_workspace_bytes(num_views, F) bytes
 * (per-face gradients of the relative rotation).  Same arithmetic as csb_mesh_xyz_* / csb_mesh_rotation_*. */
size_t csb_mesh_transform_workspace_bytes(int num_views, int F);
int csb_mesh_transform_forward(int num_views, int P, int Nv, int F, const float* rest_vertices, const float* vertices,
                               const float* residual, const int64_t* face, const int64_t* face_ids,
                               const float* face_bary, const float* rotation, float* xyz, float* rot, void* stream);
int csb_mesh_transform_backward(int num_views, int P, int Nv, int F, const float* rest_vertices, const float* vertices,
                                const float* residual, const int64_t* face, const int64_t* face_ids,
                                const float* face_bary, const float* rotation, const float* dL_dxyz,
                                const float* dL_drot, float* dL_dvertices, float* dL_dface_bary, float* dL_drotation,
                                void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------- *
 * Photometric loss
 * ------------------------------------------------------------------------- */

/* image_losses(image_tensor, gt_image_tensor, opt, mask_tensor)  (scene_reconstruction/train_utils.py:50-74):
 *   loss = l1_loss(img, gt, mask) + lambda_dssim * ssim_loss,           utils/loss_utils.py:20-23
 *   ssim_loss = 1 - ssim(img, gt)                (no mask)              utils/loss_utils.py:39-70
 *             = ((1 - ssim_map) * mask).mean()   (mask)
 * img, gt f32[n_img, C, H, W]; mask f32[n_img, mask_channels, H, W] or NULL, mask_channels 1 or C; the SSIM
 * window is the reference's 11x11 Gaussian (sigma 1.5) with zero padding.  Outputs (all device):
 *   losses   f32[3] = {loss, l1, ssim_loss}
 *   ssim_map f32[n_img, C, H, W] or NULL          (ssim(..., return_map=True))
 *   dmaps    f32[3, n_img, C, H, W] or NULL: per pixel dS/dmu1 (total), dS/dsigma1^2, dS/dsigma12 -- what
 *            csb_image_loss_backward needs; NULL for an evaluation-only call.
 * workspace: csb_image_loss_workspace_bytes(...) bytes (per-tile partial sums, folded in a fixed order). */
size_t csb_image_loss_workspace_bytes(int n_img, int C, int H, int W);
int csb_image_loss_forward(int n_img, int C, int H, int W, const float* img, const float* gt, const float* mask,
                           int mask_channels, float lambda_dssim, float* ssim_map, float* dmaps, float* losses,
                           void* workspace, size_t workspace_bytes, void* stream);
/* csb_image_loss_forward without a caller workspace and without any allocation.  No counterpart in the reference:
 * added for callers that cannot call cudaMalloc (CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, kernels and results as csb_image_loss_forward (bitwise:
 * the per-tile partial sums are folded in the same fixed order); the partial sums live in a file-scope
 * `static __device__` array of CSB_IMAGE_LOSS_STATIC_BYTES that every GPU loading the library holds until exit.
 * Limit: n_img * C * ceil(W/32) * ceil(H/32) <= CSB_IMAGE_LOSS_STATIC_BYTES / 8 = 131072 tiles (e.g. 16 RGB images of
 * 1664x1664, or the 4 x 3 x 800 x 800 batch of the refinement step 17 times over); beyond it the call launches
 * nothing and returns CSB_ERR_STATIC_LIMIT.  The device that owns `stream` must be current.  All calls (on one
 * device) share the array: it keeps the last call's partial sums, nothing in it has to start at zero (every tile of
 * a call overwrites its own slot before the fold reads it), and two calls that overlap on the GPU (different streams,
 * no ordering between them) race on it -- serialise them or use csb_image_loss_forward with separate workspaces. */
#define CSB_IMAGE_LOSS_STATIC_BYTES (1u << 20)
size_t csb_image_loss_static_max_bytes(void);
int csb_image_loss_forward_static(int n_img, int C, int H, int W, const float* img, const float* gt,
                                  const float* mask, int mask_channels, float lambda_dssim, float* ssim_map,
                                  float* dmaps, float* losses, void* stream);
/* dL/dimg f32[n_img, C, H, W] of  L = scale * (w_l1 * sum |(img - gt) mask| + sum_q g(q) S(q)),
 *   g(q) = grad_map[q] if grad_map != NULL (autograd through a returned ssim_map) else w_ssim * mask(q);
 * scale = *grad_scale, a DEVICE scalar (the upstream gradient of the loss; NULL = 1) so that no host
 * synchronisation is needed.  image_losses: w_l1 = 1/count, w_ssim = -lambda_dssim/count, count = n_img*C*H*W. */
int csb_image_loss_backward(int n_img, int C, int H, int W, const float* img, const float* gt, const float* mask,
                            int mask_channels, const float* dmaps, const float* grad_map, const float* grad_scale,
                            float w_ssim, float w_l1, float* dL_dimg, void* stream);

/* ------------------------------------------------------------------------- *
 * Optimiser step
 * ------------------------------------------------------------------------- */

/* torch.optim.Adam(param_groups, lr=0.0, eps=1e-15).step() over the parameter tensors of
 * MultiGaussianMesh.training_setup (scene_reconstruction/gaussian_mesh.py:121-135, called at
 * scene_reconstruction/train_utils.py:312-317) -- and of any other fp32 parameter list (the meshnet
 * optimiser) -- in ONE launch per 48 tensors instead of ~8 elementwise launches per tensor:
 *   m += (g - m)(1 - beta1);  v = beta2 v + (1 - beta2) g g
 *   p -= (lr / (1 - beta1^t)) * m / (sqrt(v) / sqrt(1 - beta2^t) + eps)       (amsgrad off, weight decay 0)
 * params/grads/exp_avg/exp_avg_sq: HOST arrays of n_tensors DEVICE pointers; numel, lr: host arrays (per
 * tensor).  `step` is t after the increment (1 on the first call).  A NULL grads[i] skips tensor i, like a
 * parameter whose .grad is None.  Scalars are doubles because torch derives step size and bias corrections in
 * Python doubles before rounding to fp32.  If zero_grads != 0 the gradients are zeroed in passing
 * (optimizer.zero_grad(set_to_none=False) fused). */
int csb_adam_step(int n_tensors, float* const* params, float* const* grads, float* const* exp_avg,
                  float* const* exp_avg_sq, const int64_t* numel, const double* lr, double beta1, double beta2,
                  double eps, int step, int zero_grads, void* stream);
/* The same step with the step count on the device, for capture in a CUDA graph (torch's `capturable=True`): one
 * extra one-thread launch increments *step_dev (int64) and derives the two bias corrections from it in double into
 * bias_dev[2]; the step kernel reads them.  No host-side value changes between calls. */
int csb_adam_step_capturable(int n_tensors, float* const* params, float* const* grads, float* const* exp_avg,
                             float* const* exp_avg_sq, const int64_t* numel, const double* lr, double beta1,
                             double beta2, double eps, long long* step_dev, double* bias_dev, int zero_grads,
                             void* stream);

/* ------------------------------------------------------------------------- *
 * Gradient all-reduce over NVLink peer memory
 * ------------------------------------------------------------------------- */

/* The single collective of the data-parallel path: the sum over the GPUs of one node of the flat parameter
 * gradient buffer (views sharded over GPUs: the reference loops them sequentially on one GPU,
 * scene_reconstruction/train_utils.py:259; graphs of a GNN batch: train_meshnet_sim.py:491-532).
 * A "peer buffer" is csb_peer_header_bytes() of flags followed by the fp32 payload.  Every rank allocates one
 * (csb_peer_alloc: cudaMalloc, zero-filled), exports its 64-byte CUDA IPC handle (csb_peer_export), and opens the
 * handles of the other ranks (csb_peer_import; csb_peer_close / csb_peer_free tear down).
 * csb_peer_all_reduce enqueues ONE kernel on `stream`: bufs[r] = rank r's buffer as mapped in this process
 * (bufs[rank] = the own allocation), n_floats (a multiple of 4) payload floats, seq = 1, 2, 3, ... incremented by
 * the caller on every call (the same value on all ranks).  In place; the sum is formed in rank order on every
 * element, so all ranks hold bit-identical results.  world <= 8.  Returns 0 or a CSB_ERR_* / cudaError_t code.
 * A peer that does not reach the call within the timeout (csb_peer_set_timeout, seconds of wall-clock time measured on
 * the device; default 600, 0 = wait for ever) makes the kernel leave without touching the payload and record the call's
 * seq in the buffer's header: csb_peer_timed_out (synchronous read of that word; 0 = no call has timed out).  The CUDA
 * context survives -- NCCL's behaviour (an error to handle), not a trap. */
size_t csb_peer_header_bytes(void);
int csb_peer_alloc(size_t bytes, void** ptr);
int csb_peer_free(void* ptr);
int csb_peer_export(void* ptr, void* handle64);
int csb_peer_import(const void* handle64, void** ptr);
int csb_peer_close(void* ptr);
int csb_peer_all_reduce(int world, int rank, void* const* bufs, size_t n_floats, unsigned seq, void* stream);
int csb_peer_set_timeout(double seconds);
int csb_peer_timed_out(const void* local_buf, unsigned* seq_out);

#ifdef __cplusplus
}
#endif
#endif /* CLOTH_SPLATTING_B200_H */
