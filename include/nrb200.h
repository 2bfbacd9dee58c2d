/* nrb200.h — C ABI of libnrb200.so, the B200-native neural-render hot path.
 *
 * This is the drop-in boundary for the per-frame path of Alehandreus/neural-renderer.
 * The reference has no C ABI: its boundary is the C++ class RendererNeural
 * (reference src/cuda_renderer_neural.h:84-124) fed by Scene/Mesh objects
 * (src/scene.h:44-71, src/mesh.h:45-78).  Every entry point below cites the reference
 * member it replaces; the source-compatible C++ class that forwards to these calls lives
 * in include/nrb200_renderer.hpp.
 *
 * Conventions: plain pointers and sizes only, no C++/torch types.  Every function that can
 * fail returns 0 on success and a negative nrb_status otherwise; nrb_last_error() returns a
 * thread-local description.  Host arrays passed to nrb_update_* are borrowed for the call
 * only.  There is no CPU fallback: if no CUDA device (sm_100) is usable nrb_create fails.
 */
#ifndef NRB200_H
#define NRB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NRB_ABI_VERSION 1

typedef enum nrb_status {
    NRB_OK = 0,
    NRB_ERR_INVALID = -1,   /* bad argument */
    NRB_ERR_CUDA = -2,      /* CUDA runtime failure (message in nrb_last_error) */
    NRB_ERR_STATE = -3,     /* call not valid in the current state (e.g. no weights) */
    NRB_ERR_SIZE = -4       /* byte-size mismatch (weights file) */
} nrb_status;

/* ---- POD mirrors of the reference's host structs (layouts are part of the contract) ---- */

/* = BvhNode, reference src/bvh_data.h:5-13 (44 bytes). */
typedef struct nrb_bvh_node {
    float bounds_min[3];
    float bounds_max[3];
    int32_t left, right;    /* inner: child indices (right == left + 1 in reference builds) */
    int32_t first, count;   /* leaf: triangle range in the leaf-ordered index buffer */
    int32_t is_leaf;
} nrb_bvh_node;

/* = MaterialParam / MaterialParamVec3 / Material, reference src/material.h:9-117 (12/20/180 B). */
typedef struct nrb_material_param {
    float value;
    uint32_t texture_id;
    uint8_t channel;
    uint8_t textured;
    uint8_t pad_[2];
} nrb_material_param;

typedef struct nrb_material_param3 {
    float value[3];
    uint32_t texture_id;
    uint8_t textured;
    uint8_t pad_[3];
} nrb_material_param3;

typedef struct nrb_material {
    nrb_material_param3 base_color;
    nrb_material_param metallic, roughness, specular, specular_tint, anisotropy;
    nrb_material_param sheen, sheen_tint;
    nrb_material_param clearcoat, clearcoat_gloss;
    nrb_material_param3 normal;
    nrb_material_param3 base_emission;
    float emission_scale;
    float ior;
    float specular_transmission;
} nrb_material;

/* = MeshTexture / TextureDeviceView, reference src/mesh.h:23-37 (8-bit texels, host pointer). */
typedef struct nrb_texture {
    const unsigned char* pixels;
    int32_t width, height, channels;
    int32_t srgb;
} nrb_texture;

/* What Mesh holds after Mesh::buildBvh (reference src/mesh.h:80-172, src/mesh_bvh.cpp:27-162):
 * indices are already in BVH leaf order; nodes are the flattened BvhNode[].  Replaces
 * Mesh::uploadToDevice / Mesh::deviceView (src/mesh.cu:91-261). */
typedef struct nrb_mesh_desc {
    const float* vertices;          /* 3 * num_vertices */
    const float* normals;           /* 3 * num_vertices or NULL */
    const float* texcoords;         /* 2 * num_vertices or NULL */
    const uint32_t* indices;        /* 3 * num_triangles, leaf order */
    int32_t num_triangles;
    int32_t num_vertices;
    const nrb_bvh_node* nodes;
    int32_t num_nodes;
    const nrb_material* materials;  /* or NULL */
    const uint32_t* material_map;   /* first triangle of each primitive, or NULL */
    const int32_t* material_ids;    /* material id of each primitive, or NULL */
    int32_t num_primitives;
    int32_t num_materials;
    const nrb_texture* textures;
    int32_t num_textures;
    int32_t texture_nearest;
    int32_t has_mesh_materials;
    float bounds_min[3];            /* Mesh::boundsMin()/boundsMax(), src/mesh_bvh.cpp:147-161 */
    float bounds_max[3];
} nrb_mesh_desc;

/* Mesh slots of Scene (reference src/scene.h:46-56). */
typedef enum nrb_mesh_slot {
    NRB_MESH_ORIGINAL = 0,
    NRB_MESH_INNER_SHELL = 1,
    NRB_MESH_OUTER_SHELL = 2,
    NRB_MESH_ADDITIONAL = 3,
    NRB_MESH_SLOT_COUNT = 4
} nrb_mesh_slot;

/* = NeuralNetworkConfig, reference src/config_loader.h:46-50, plus the network shape that the
 * reference hard-codes in the RendererNeural ctor (src/cuda_renderer_neural.cu:1647-1688).
 * Zero in any shape field selects the reference value given in the comment. */
typedef struct nrb_nn_config {
    int32_t log2_hashmap_size;      /* 14 */
    int32_t base_resolution;        /* 16 */
    int32_t n_points;               /* 3 (entry, exit, mid) */
    int32_t n_levels;               /* 8 */
    int32_t n_features_per_level;   /* 4 */
    float per_level_scale;          /* 2.0 */
    int32_t sh_degree;              /* 4; negative = no direction encoding (inputs are points only) */
    int32_t mlp_width;              /* 128 */
    int32_t n_hidden_layers;        /* 4 */
} nrb_nn_config;

/* TraceMode, reference src/mesh_traversal.cuh:107-111. */
typedef enum nrb_trace_mode {
    NRB_TRACE_ANY = 0,
    NRB_TRACE_FORWARD_ONLY = 1,
    NRB_TRACE_BACKWARD_ONLY = 2,
    NRB_TRACE_ALL_MODES = 3,  /* (new, parity hook) one fused pass: out.t/u/v = ANY/FORWARD/BACKWARD distances */
    NRB_TRACE_ALL_MODES_UNION = 4, /* (new, parity hook) the same through the bounce entry's resumable engine,
                                     which prunes by the union of the live modes */
    NRB_TRACE_DYN_ANY = 5,         /* (new, parity hooks) modes 0..2 through the run-time-mode traversal loop of the */
    NRB_TRACE_DYN_FORWARD = 6,     /* segment stages (k_seg_exits, k_seg_apply), which skips subtrees whose normal    */
    NRB_TRACE_DYN_BACKWARD = 7,    /* cone rules the facing mode out: same {t,u,v,prim} as modes 0..2                  */
    NRB_TRACE_PLAIN_ANY = 8,       /* (new, parity hooks) modes 0..2 through the first engine (one loop per facing mode, */
    NRB_TRACE_PLAIN_FORWARD = 9,   /* reconvergence left to the compiler); modes 0..2 themselves run the frame kernels'  */
    NRB_TRACE_PLAIN_BACKWARD = 10  /* walk, reconverged by a warp vote every few steps (trace_sync): same {t,u,v,prim}   */
} nrb_trace_mode;

/* = PreliminaryHitInfo, reference src/hit_info.h:7-13.  t == 1e30f means no hit. */
typedef struct nrb_hit {
    float t, u, v;
    uint32_t prim;
} nrb_hit;

typedef struct nrb_renderer nrb_renderer;

/* ---- lifecycle ---------------------------------------------------------------------- */

/* RendererNeural::RendererNeural(Scene&, const NeuralNetworkConfig*)
 * (src/cuda_renderer_neural.cu:1632-1718).  cfg may be NULL (reference defaults).
 * device < 0 selects the current CUDA device. */
int nrb_create(const nrb_nn_config* cfg, int device, nrb_renderer** out);
/* RendererNeural::~RendererNeural (src/cuda_renderer_neural.cu:1720-1734). */
void nrb_destroy(nrb_renderer* r);
const char* nrb_last_error(void);
int nrb_abi_version(void);

/* ---- scene hand-off ------------------------------------------------------------------ */

/* Mesh::uploadToDevice + Mesh::deviceView (src/mesh.cu:91-261) for one Scene slot; also
 * builds the re-laid-out traversal structure.  desc == NULL or num_triangles == 0 clears
 * the slot (an empty view, as MeshDeviceView{}). */
int nrb_update_mesh(nrb_renderer* r, int slot, const nrb_mesh_desc* desc);
/* EnvironmentMap::uploadToDevice + deviceView (src/scene.cu:318-342, src/scene.h:10-16).
 * rgb = fp32 triplets, row-major, w*h of them; NULL selects the sky gradient branch
 * (src/cuda_renderer_neural.cu:166-170). */
int nrb_update_env(nrb_renderer* r, const float* rgb, int width, int height, float strength);
/* Scene::globalMaterial() (src/scene.h:61-62). */
int nrb_set_global_material(nrb_renderer* r, const nrb_material* m);

/* ---- weights ------------------------------------------------------------------------- */

/* RendererNeural::loadWeightsFromFile (src/cuda_renderer_neural.cu:1752-1808): raw little-
 * endian fp16, (mlp + one hash grid) params; swap_order != 0 means [hg | mlp] (the
 * reference default, cuda_renderer_neural.h:235), else [mlp | hg].  The single grid is
 * shared by all point encoders (the reference tiles it n_points times). */
int nrb_load_weights(nrb_renderer* r, const void* fp16, size_t bytes, int swap_order);
int nrb_load_weights_file(nrb_renderer* r, const char* path, int swap_order);
/* RendererNeural::paramsBytes(): bytes of the reference's [mlp | hg x n_points] block. */
size_t nrb_params_bytes(const nrb_renderer* r);
/* Bytes loadWeightsFromFile expects. */
size_t nrb_weights_file_bytes(const nrb_renderer* r);

/* ---- per-frame state ------------------------------------------------------------------ */

/* RendererNeural::resize (src/cuda_renderer_neural.cu:1736-1746). */
int nrb_resize(nrb_renderer* r, int width, int height);
/* RendererNeural::setCameraBasis (RenderBasis, src/renderer.h:5-10). */
int nrb_set_camera_basis(nrb_renderer* r, const float forward[3], const float right[3],
                         const float up[3], float fov_y);

/* The setters of cuda_renderer_neural.h:94-114. */
typedef enum nrb_option {
    NRB_OPT_USE_NEURAL_QUERY = 0,     /* setUseNeuralQuery */
    NRB_OPT_SAMPLES_PER_PIXEL = 1,    /* setSamplesPerPixel */
    NRB_OPT_BOUNCE_COUNT = 2,         /* setBounceCount */
    NRB_OPT_LAMBERT_VIEW = 3,         /* setLambertView */
    NRB_OPT_CLASSIC_MESH_INDEX = 4,   /* setClassicMeshIndex */
    NRB_OPT_USE_ADDITIONAL_MESH = 5,  /* setUseAdditionalMesh */
    NRB_OPT_USE_CONSTANT_NEURAL_COLOR = 6, /* setConstantNeuralColor(use, .) */
    NRB_OPT_USE_DIRECT_ENV_COLOR = 7, /* setDirectEnvColor(use, .) */
    NRB_OPT_CAPTURE = 8,              /* (new) keep per-element debug buffers for nrb_read_buffer */
    NRB_OPT_USE_HARDWARE_RT = 9,      /* setUseHardwareRT: accepted, ignored (no OptiX path) */
    NRB_OPT_PROFILE = 10,             /* (new) record a CUDA event after every kernel launch: nrb_get_stage */
    NRB_OPT_QUERY_KERNEL = 11,        /* (new, process-wide) neural query kernel: 0 auto, 1 mma.sync, 2 fused tcgen05, 3 split */
    NRB_OPT_TAIL_GRAPH = 12,          /* (new) segment-loop iterations 1..9 of a pass: 2 (default) one fused kernel in which
                                         a warp carries 16 of the few continuing rays through exits -> network -> apply;
                                         1 a device-driven CUDA conditional WHILE graph; 0 nine fixed launches */
    NRB_OPT_DENOISE = 13,             /* (new) 1: the frame written to devicePixels() is the joint bilateral filter of
                                         the accumulation (bilateralDenoiseKernel, src/denoiser.cuh:47-113, which the
                                         reference ships but never launches), guided by the sample-0 primary-hit
                                         normal and albedo; accumulation itself is untouched.  Single-GPU frames only
                                         (nrb_set_tile world 1): the 11x11 window would need rows of other ranks. */
    NRB_OPT_PATH_TAIL = 15,           /* (new) bounces >= 2 run as ONE kernel (a warp carries up to 16 surviving paths through sample ->
                                         shell entry -> segment loop -> integrate, bounce after bounce) while the previous frames entered
                                         bounce 2 with at most `value` paths (default 49152; 0 = always the per-stage kernels).  Needs
                                         NRB_OPT_TAIL_GRAPH = 2.  The network then runs on the mma.sync tile: outputs equal the
                                         tcgen05 kernels' within fp16 tolerance. */
    NRB_OPT_TILE_ORDER = 16,          /* (new) 1 (default): the ground-truth primary pass hands its 8x4 tiles out most-expensive-first,
                                         ordered by the clocks an earlier frame of the same camera measured per tile; 0: row-major.
                                         Scheduling only: every pixel is computed by the same arithmetic either way. */
    NRB_OPT_FRAME_SLICES = 17,        /* (new) 0 (default) = automatic: 1 while frames are pipelined (NRB_OPT_PIPELINE), else 4 in ground-truth mode
                                         and 1 in neural mode (measured).
                                         S > 1: the rows this renderer owns are dealt to S internal slices (16-row bands round-robin,
                                         the multi-GPU tile rule) whose stage chains run concurrently on S streams, so the latency-
                                         bound tails of one slice run under the throughput stages of another.  Same pixels.  Not with
                                         NRB_OPT_PROFILE / CAPTURE / DENOISE or NRB_OPT_TAIL_GRAPH != 2 (the frame then runs unsliced). */
    NRB_OPT_PIPELINE = 18,            /* (new) 1 (default): consecutive frames are pipelined - in neural mode a frame's stages up to the end of
                                         bounce 1 run on a second stream and alternate between two wavefront-state buffers, its end
                                         (the later bounces in k_path_tail, finalize) stays on nrb_stream(), so frame k+1 starts under
                                         the latency-bound end of frame k.  Frames still COMPLETE in order on nrb_stream(); pixels and
                                         accumulation are those of the unpipelined frames, bit for bit.  0: one stream, one buffer.
                                         Ground-truth mode likewise: k_gt_primary / k_gt_bounce of frame k+1 under k_gt_tail and
                                         finalize of frame k.  Not with NRB_OPT_PROFILE / CAPTURE / DENOISE, NRB_OPT_TAIL_GRAPH != 2 in
                                         neural mode, the lambert view, or frame slices (those frames run unpipelined). */
    NRB_OPT_BOUNCE_ENTRY = 19,        /* (new) which kernel runs checkBounceEarlyTermination + traceOuterShellEntryFromRays
                                         (src/cuda_renderer_neural.cu:503-553, 773-852).  0 (default): separate pruned walks - nearest
                                         face of the outer shell, its nearest front face searched only beyond a nearest face that
                                         is a back face, the inner shell as far as the outer hit.  1: one fused three-class walk
                                         under a warp-voting scheduler (the kernel of rounds 1-2, kept as a cross-check).  Same
                                         frames and counters. */
    NRB_OPT_LITERAL_BOUNCE_TRACES = 14 /* (new, parity hook) 1: the bounce entry runs every trace of
                                         checkBounceEarlyTermination / traceOuterShellEntryFromRays
                                         (src/cuda_renderer_neural.cu:503-553, 773-852) literally.  0 (default): traces
                                         whose outcome is already decided are not run - the inner-shell trace stops
                                         at the outer-shell distance it is compared with (and is skipped when there is
                                         none), and the FORWARD re-entry trace from beyond the BACKWARD exit is skipped
                                         when the first FORWARD trace found no front face along the WHOLE ray: its
                                         candidates are a subset of that trace's.  Also the segment exits: the outer-shell
                                         BACKWARD trace of a ray that hits the inner shell stops at that hit (an exit
                                         beyond it is only ever compared with it); the literal setting traces to the
                                         end and so keeps the reference's exitT = 1e-8 fallback for rays that leak
                                         through the outer shell (src/cuda_renderer_neural.cu:899-905). */
} nrb_option;
int nrb_set_option(nrb_renderer* r, int option, int value);
int nrb_get_option(const nrb_renderer* r, int option);
/* setEnvmapRotation (degrees). */
int nrb_set_envmap_rotation(nrb_renderer* r, float degrees);
/* colour arguments of setConstantNeuralColor / setDirectEnvColor. */
int nrb_set_constant_neural_color(nrb_renderer* r, const float rgb[3]);
int nrb_set_direct_env_color(nrb_renderer* r, const float rgb[3]);
/* resetSamples(). */
int nrb_reset_samples(nrb_renderer* r);
/* (new, multi-GPU) this renderer owns the image rows whose band index
 * (row / band_rows) is congruent to rank modulo world; pixel indices and RNG streams stay
 * global, so the union over ranks equals the 1-GPU image bit for bit.  world == 1 = all. */
int nrb_set_tile(nrb_renderer* r, int band_rows, int rank, int world);

/* ---- the frame ------------------------------------------------------------------------ */

/* RendererNeural::render(const Vec3& camPos) (src/cuda_renderer_neural.cu:2085-2755).
 * Returns after the work has been enqueued and joined to the legacy default stream, so a
 * following cudaMemcpy from nrb_device_pixels() is ordered after the frame (the reference's
 * callers rely on this, evaluate.cu:266).  The join is symmetric: the frame also waits for whatever the
 * caller enqueued on the legacy default stream before the call (an async copy or a kernel still reading the
 * previous frame), so devicePixels() is never overwritten under a reader. */
int nrb_render(nrb_renderer* r, const float cam_pos[3]);
/* RendererNeural::devicePixels(): uchar4 RGBA, row-major, width*height (device pointer). */
void* nrb_device_pixels(const nrb_renderer* r);
/* Convenience for host callers: blocking copy of devicePixels into rgba (4*width*height B). */
int nrb_read_pixels(nrb_renderer* r, void* rgba, size_t bytes);
/* (new) Non-blocking variant for callers that stream frames: returns once the copy is enqueued on
 * a second stream, so the transfer of this frame overlaps the kernels of the next nrb_render.
 * rgba should be pinned; it is valid after nrb_sync() (the copy runs on the renderer's copy stream and is NOT joined
 * to the legacy default stream, or the next frame - which orders itself behind that stream - would wait for the PCIe
 * transfer).  One transfer in flight per renderer: a second call waits on the device for the first. */
int nrb_read_pixels_async(nrb_renderer* r, void* rgba, size_t bytes);
/* (new) Multi-GPU frame gather fused into the frame, over peer memory instead of a collective: the root
 * rank (one process per GPU) exports its frame buffer, the others attach and from then on their finalize
 * kernel stores the row bands they own (nrb_set_tile) directly into the root's devicePixels() over NVLink
 * and signals; the root's nrb_render ends with a device-side wait for `peers` signals, so the work it
 * enqueues next sees the complete frame.  The handle (NRB_GATHER_HANDLE_BYTES) travels between the
 * processes by any means (torch.distributed broadcast, a pipe).  All ranks must render the same number of
 * frames; a rank that stops makes the others' waits give up after ~3 s and nrb_sync report it.  A peer's
 * own devicePixels() is no longer written.  nrb_resize drops the association on either side. */
#define NRB_GATHER_HANDLE_BYTES 144
int nrb_gather_export(nrb_renderer* r, int peers, void* handle, size_t handle_bytes);
int nrb_gather_attach(nrb_renderer* r, const void* handle, size_t handle_bytes);
int nrb_gather_detach(nrb_renderer* r);
int nrb_width(const nrb_renderer* r);
int nrb_height(const nrb_renderer* r);
uint32_t nrb_accum_sample_count(const nrb_renderer* r);

/* KernelTimings / lastFrameTimings (src/cuda_renderer_neural.h:17-66): ms per the 13
 * reference buckets, rayCount, neuralRayCalls; always compiled in here. */
#define NRB_TIMING_BUCKETS 13
int nrb_get_timings(nrb_renderer* r, double ms[NRB_TIMING_BUCKETS], int* ray_count,
                    uint64_t* neural_ray_calls);
/* (new) per-kernel device time accumulated over the frames rendered with NRB_OPT_PROFILE = 1 (the
 * always-available counterpart of the reference's -DPROFILE_KERNELS event pairs,
 * src/cuda_renderer_neural.cu:39-57).  Returns 0 and fills the outputs for index < number of distinct
 * kernels, 1 past the end.  nrb_reset_stages clears the totals. */
int nrb_get_stage(nrb_renderer* r, int index, const char** name, double* total_ms, uint64_t* launches);
int nrb_reset_stages(nrb_renderer* r);
/* (new) counters of the last frame: [0] kernels launched, [1] rays traced (segments),
 * [2] neural queries, [3] path vertices shaded. */
int nrb_get_counters(nrb_renderer* r, uint64_t out[4]);

/* (new) traversal work of the last nrb_trace_rays_host call (modes 0..2): [0] rays, [1] node pairs fetched (64 B each),
 * [2] triangle blocks fetched (48 B each) - the counts behind the algorithmic bytes per ray of SURVEY.md 8d. */
int nrb_get_trace_stats(nrb_renderer* r, uint64_t out[3]);

/* (new, parity/debug) copy one of the renderer's per-element buffers to the host, in the
 * reference's layout (src/cuda_renderer_neural.h:160-214).  Needs NRB_OPT_CAPTURE = 1
 * before the frame.  Names: "hit_positions", "hit_normals", "hit_colors",
 * "hit_material_params" (3 floats/element), "hit_flags" (int32/element), "hit_distances"
 * (float/element), "path_radiance" (3 floats/element), "accum" (3 floats/pixel). */
int nrb_read_buffer(nrb_renderer* r, const char* name, void* dst, size_t bytes);

/* ---- batch entry points (the kernels behind the frame, on caller-supplied arrays) -------
 * *_host variants take host pointers and copy both ways inside the call; the others take
 * device pointers and enqueue on the renderer's stream (call nrb_sync to wait). */

/* traceMeshWithMode (src/mesh_traversal.cuh:118-222) for n rays against one slot's mesh:
 * closest hit with the reference's facing rules; out[i] = {t,u,v,prim} or t = 1e30f. */
int nrb_trace_rays(nrb_renderer* r, int slot, int mode, const float* d_origins,
                   const float* d_directions, size_t n, nrb_hit* d_out);
int nrb_trace_rays_host(nrb_renderer* r, int slot, int mode, const float* origins,
                        const float* directions, size_t n, nrb_hit* out);

/* network_->inference_mixed_precision (src/cuda_renderer_neural.cu:2031; tiny-cuda-nn
 * object.h:166-209): inputs fp32 [n][input_dims] (ray-major, = column-major
 * [input_dims x n]), outputs fp16 [n][16].  n need not be a multiple of 256. */
int nrb_neural_query(nrb_renderer* r, const float* d_inputs, size_t n, void* d_outputs_fp16);
int nrb_neural_query_host(nrb_renderer* r, const float* inputs, size_t n, void* outputs_fp16);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The same query as
 * nrb_neural_query (split encode + tcgen05 MLP kernels), with the encoded-tile scratch in a file-scope
 * static __device__ array of the library (8 MiB + 256 B, held by every GPU that loads the library until exit)
 * instead of the renderer's cudaMalloc'ed buffer: the call allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine) and only enqueues a 64-byte memset and two kernels on the renderer's stream.
 * Limits: n <= NRB_STATIC_QUERY_MAX queries per call, a network shape the split query supports (width 64 or 128,
 * 2 or 4 features per level, encoded width <= 256) with weights loaded; anything else returns NRB_ERR_INVALID with a
 * message in nrb_last_error().  All calls on a device share the one array (it keeps the last call's tiles; the
 * cursors it must find at zero are cleared by the call): calls that overlap on the GPU - two streams, two renderers,
 * or this call beside a running nrb_render of ANOTHER renderer is fine, but two static queries at once are not -
 * race on it.  Inputs and outputs are the caller's device arrays, as for nrb_neural_query. */
#define NRB_STATIC_QUERY_MAX 16384
int nrb_neural_query_static(nrb_renderer* r, const float* d_inputs, size_t n, void* d_outputs_fp16);
int nrb_input_dims(const nrb_renderer* r);
int nrb_output_dims(const nrb_renderer* r);

int nrb_sync(nrb_renderer* r);
/* The renderer's stream as a cudaStream_t (so callers can record events on it). */
void* nrb_stream(const nrb_renderer* r);

/* ---- host-side helper ------------------------------------------------------------------
 * (new) SAH BVH build producing the reference's BvhNode[] / leaf-ordered index format
 * (what Mesh::buildBvh yields, src/mesh_bvh.cpp:27-162), for callers that do not link
 * the reference's ext/bvh builder.  indices is permuted in place; nodes must hold
 * 2*num_triangles entries; *num_nodes receives the count; prim_ids (optional,
 * num_triangles) receives the source triangle of each output triangle. */
int nrb_build_bvh(const float* vertices, int num_vertices, uint32_t* indices, int num_triangles,
                  nrb_bvh_node* nodes, int* num_nodes, uint32_t* prim_ids);

/* (new) The same contract built on the GPU in milliseconds (LBVH: Morton sort + radix tree + bottom-up
 * fit, subtrees of <= max_leaf (default 4, <= 32) triangles folded into one leaf): for large meshes and
 * scenes that change per frame.  Host pointers in and out like nrb_build_bvh; nodes must hold
 * 2*num_triangles entries, the root is node 0 and unreachable entries may remain.  device < 0 = current
 * device; build_ms (optional) receives the device time of the build kernels.  The tree differs from the
 * reference's SAH tree: map hit ids through prim_ids when comparing with a reference build.
 * Replaces Mesh::buildBvh (src/mesh_bvh.cpp:27-162) for callers that can trade tree quality for build time. */
int nrb_build_bvh_gpu(const float* vertices, int num_vertices, uint32_t* indices, int num_triangles,
                      nrb_bvh_node* nodes, int* num_nodes, uint32_t* prim_ids, int device, int max_leaf, float* build_ms);
const char* nrb_build_bvh_gpu_error(void);

#ifdef __cplusplus
}
#endif
#endif /* NRB200_H */
