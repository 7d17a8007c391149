/* ptb.h — C ABI of libptb.so, the B200 (sm_100a) implementation of the thesis path tracer's
 * hot path: wavefront ray generation -> acceleration-structure traversal (BVH, Subgrid /
 * "2Plane", Subbeams, brute force) -> Moeller-Trumbore -> bounce/shade -> accumulation.
 *
 * The reference has no FFI; this path sits behind plain C++ free functions over globals.  Each
 * entry point below names the reference function(s) it stands in for (paths relative to the
 * reference repository).  A C++ facade with the reference's own signatures (tracer.h, scene.h,
 * bvh.h, lightFieldGrid.h, lightFieldGridSubBeams.h, common.h, shader.h) that forwards to this
 * ABI is in bachelor-thesis-pathtracer_b200/host/; INTEGRATION.md shows how a maintainer links it.
 *
 * Conventions: every function returns 0 on success or a negative ptb_status; the message of the
 * last failure of a context is available from ptb_last_error().  All pointers are plain host
 * pointers to POD arrays unless the name says `_device`; upload_* copy, so the caller may free
 * its arrays on return.  A context is single-owner: calls on one context must not overlap.
 * There is no CPU fallback: without a CUDA device ptb_create fails.
 */
#ifndef PTB_H
#define PTB_H
#include <stdint.h>
#include "ptbs_format.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ptb_ctx ptb_ctx;

typedef enum ptb_status {
    PTB_OK = 0,
    PTB_ERR_INVALID = -1, /* bad argument / call order */
    PTB_ERR_CUDA = -2,    /* CUDA runtime error (sticky: the context is poisoned) */
    PTB_ERR_IO = -3,
    PTB_ERR_NOT_BUILT = -4 /* the structure the mode needs has not been built / uploaded */
} ptb_status;

/* src/common.h:13  enum IntersectMode{ALL, BVH, TWO_PLANE, SUB_BEAMS} */
typedef enum ptb_mode { PTB_MODE_ALL = 0, PTB_MODE_BVH = 1, PTB_MODE_TWO_PLANE = 2, PTB_MODE_SUB_BEAMS = 3 } ptb_mode;

/* src/types/camera.h:5-18 — what createCameraRay (camera.cpp:47-66) reads */
typedef struct ptb_camera {
    float origin[3];
    float forward[3]; /* any length; orthonormalised like cameraSetForward (camera.cpp:25-32) */
    float focus;
    float dof;
    float fov_degrees; /* 70 in every scene file; fovScale = tanf((fov*3.14f/180)/2) (camera.cpp:16,43-45) */
} ptb_camera;

/* The mutable globals the reference reads per ray, passed once per launch instead:
 * WIDTH/HEIGHT (tracer.cpp:22-23), getIntersectMode (common.cpp:12), getNEE / getPrimaryOnly
 * (shader.cpp:17-18), getDebugView/ShowTris/Scale (tracer.cpp:35-37). */
typedef struct ptb_params {
    int32_t width, height;
    int32_t mode;             /* ptb_mode */
    int32_t nee;              /* shader.cpp:17 (the --testing protocol turns it off, main.cpp:101) */
    int32_t primary_only;     /* shader.cpp:245-248 */
    int32_t max_depth;        /* 0 = unbounded like the reference; N = terminate when depth >= N after depth++ */
    int32_t reference_quirks; /* 1: per-mode behaviour of roulette-killed rays */
    int32_t debug_view;       /* tracer.cpp:136-142: 0 off, 1 AS-test heat map, 2 triangle-test heat map */
    int32_t debug_scale;
    /* multi-GPU pixel partition: tile (tx,ty) of tile_w x tile_h pixels belongs to rank
     * (ty*tiles_x + tx) % world.  world = 1 -> the whole frame. */
    int32_t rank, world, tile_w, tile_h;
    int32_t count_tests;      /* 1: accumulate the reference's two counters (tracer.cpp:131-132) */
    /* sample-range sharding (opt-in; statistical parity only): ptb_reset seeds pixel (x,y) with
     * (x<<16|y) + seed_offset * 0x9E3779B9.  0 = the reference's seeds (tracer.cpp:240). */
    int32_t seed_offset;
} ptb_params;

/* counters, ptb_read_counters */
enum {
    PTB_CNT_RAYS = 0,        /* traversals executed on the GPU (primary + bounce + shadow) */
    PTB_CNT_PATHS = 1,       /* completed paths (samples accumulated) */
    PTB_CNT_TRI_TESTS = 2,   /* getIntersectionCount (tracer.cpp:60-64); needs count_tests */
    PTB_CNT_AS_TESTS = 3,    /* getStructureIntersectionCount (tracer.cpp:66-70); needs count_tests */
    PTB_CNT_SHADOW_RAYS = 4,
    PTB_CNT_ITERATIONS = 5,  /* wavefront iterations launched */
    PTB_CNT_KERNELS = 6,     /* kernels launched by this context since ptb_reset */
    PTB_CNT_REF_RAYS = 7,    /* what the reference would have traversed (incl. re-traced misses, A.3) */
    PTB_CNT__N = 8
};

/* ---- lifetime ---------------------------------------------------------------------------- */
int ptb_create(int device, ptb_ctx **out);
void ptb_destroy(ptb_ctx *ctx);                 /* destroyTracer, tracer.cpp:245-251 */
const char *ptb_last_error(const ptb_ctx *ctx); /* never NULL; ctx may be NULL (creation errors) */

/* ---- scene: what initScene leaves in getTris()/getObjects()/getMaterials() (scene.cpp:97-131) */
int ptb_upload_triangles(ptb_ctx *ctx, const float *pos9, const float *nrm9, const float *uv6, const int32_t *mat, uint64_t n);
int ptb_upload_objects(ptb_ctx *ctx, const ptbs_object *objects, uint32_t n);       /* also derives the light list, scene.cpp:138-148 */
int ptb_upload_materials(ptb_ctx *ctx, const ptbs_material *materials, uint32_t n); /* may be called again (UI edits + reset) */
/* Material textures: pbrProbs::texture / ::normal (src/shader/shader.h:14-24) as getColorOfMaterial (shader.cpp:52-66:
 * nearest texel at the hit's uv, opacity = alpha with one RNG draw, see-through texels step behind the surface) and the
 * normal-map branch of shade (shader.cpp:251-258) read them.  rgba = Texture::data (src/types/texture.h:7-11), row-major
 * y*width+x, channel = byte/255.0f as loadTexture leaves it (texture.cpp:33-50).  material_tex2 = n_materials x
 * {albedo texture, normal map} indices into `textures`, -1 = none; call after ptb_upload_materials (a later
 * ptb_upload_materials with the same material count keeps the bindings: the UI edit + reset flow). */
typedef struct ptb_texture {
    const float *rgba;
    int32_t width, height;
} ptb_texture;
int ptb_upload_textures(ptb_ctx *ctx, const ptb_texture *textures, uint32_t n_textures, const int32_t *material_tex2, uint32_t n_materials);
int ptb_load_ptbs(ptb_ctx *ctx, const char *path); /* the uploads + camera from one .ptbs container (v1, or v2 with textures) */
/* loadScene + the model / material part of initScene + loadObject (SceneFile.cpp:21-110, scene.cpp:97-110,
 * object.cpp:35-174) without the 280-byte Triangle AoS in between: .scene / OBJ / MTL text straight into
 * the arrays of ptb_upload_*, and the scene-file camera.  OBJ/MTL are read the way tinyobjloader 2.0.0 reads them (its
 * number parser, quad split and ear clipping included), so the arrays are the reference's bit for bit.  Model paths are
 * relative to base_dir (NULL or "": the working directory, like the reference).  map_Kd / bump files are decoded when they
 * are PNG or TGA (texels = byte / 255.0f like loadTexture, texture.cpp:33-50) and uploaded with the scene; other image formats
 * fail the load unless flags has PTB_SCENE_IGNORE_TEXTURES (then hand texels to ptb_upload_textures yourself). */
#define PTB_SCENE_IGNORE_TEXTURES 1
int ptb_load_scene_file(ptb_ctx *ctx, const char *scene_file, const char *base_dir, int flags);
/* The same loader without a context or a GPU: writes the .ptbs container.  Returns PTB_OK or PTB_ERR_IO with the reason in
 * err (NUL-terminated, at most err_len bytes). */
int ptb_convert_scene_file(const char *scene_file, const char *base_dir, const char *out_ptbs, int flags, char *err, uint64_t err_len);
int ptb_set_camera(ptb_ctx *ctx, const ptb_camera *cam); /* cameraSetForward + origin/dof/focus, scene.cpp:127-130 */
/* the Camera struct verbatim (camera.h:5-18) after the UI rotated it (camera.cpp:70-135): no re-orthonormalisation */
int ptb_set_camera_basis(ptb_ctx *ctx, const float origin[3], const float forward[3], const float right[3], const float up[3],
                         float focus, float dof, float fov_scale);

/* ---- acceleration structures ---------------------------------------------------------------
 * Builders run on the host inside the library (construction is outside the timed path in the
 * reference too, main.cpp:97 vs :116) and upload flattened device layouts. */
int ptb_build_bvh(ptb_ctx *ctx);                                        /* constructBVH, bvh.h:33 (scene.cpp:113-120) */
/* constructBVH on the device (bvh.cpp:116-189,222-390 is 7.2 s at 1e5 triangles and superlinear):
 * Morton keys + radix sort + radix-tree hierarchy + bottom-up boxes, all in HBM (csrc/ptb_lbvh.cu).  Same node and
 * triangle layouts, same closest hits (they do not depend on the tree); visit counts differ from the SAH tree. */
int ptb_build_bvh_gpu(ptb_ctx *ctx);
double ptb_last_build_ms(const ptb_ctx *ctx); /* last ptb_build_bvh (host wall clock) / ptb_build_bvh_gpu (device time, CUDA events) */
int ptb_bvh_depth(const ptb_ctx *ctx);        /* inner nodes on the longest root-to-leaf chain (< PTB_STACK, the per-lane traversal stack); -1 before a build */
/* The reference's own tree for count-exact traversal.  One record per BvhNode (bvh.h:14-24) with its
 * box boxes[node.AABBIdx] (bvh.cpp:19) inlined — what getNode / getNodeAABB hand out (bvh.h:27-28); `indices` is
 * bvhGetTrisIndex (bvh.h:29): triangle ids under has_tris leaves, object ids under scene-level leaves; object_roots
 * is Object::root (object.h).  With ptb_use_ref_bvh(ctx, 1) mode BVH walks this tree statement by statement like
 * findBVHIntesection (bvh.cpp:48-99): same hits and frames as ever, and PTB_CNT_TRI_TESTS / PTB_CNT_AS_TESTS and the
 * debug_view heat maps become the reference's own numbers (tracer.cpp:127-142) — re-traced misses included. */
typedef struct ptb_ref_bvh_node {
    float bmin[3];
    int32_t start_idx;
    float bmax[3];
    int32_t end_idx;
    int32_t child_left, child_right; /* both -1: leaf */
    int32_t has_tris, pad;
} ptb_ref_bvh_node;
int ptb_upload_ref_bvh(ptb_ctx *ctx, const ptb_ref_bvh_node *nodes, uint32_t n_nodes, const uint32_t *indices, uint32_t n_indices,
                       const int32_t *object_roots, uint32_t n_objects, int32_t scene_root);
int ptb_use_ref_bvh(ptb_ctx *ctx, int on);
int ptb_build_subgrid(ptb_ctx *ctx, uint32_t gs, uint32_t os, uint32_t m); /* setGridSettings + constructGrid, lightFieldGrid.h:24,26 */
/* The reference's own Subgrid hierarchy for count-exact traversal: one record per Grid of
 * std::vector<Grid> (lightFieldGrid.h:10-22, the file-static `grids` of lightFieldGrid.cpp:23; grids 0..2 are the roots per dominant
 * axis) and its three vectors concatenated in grid order — gridLutStart / gridLutEnd stay relative to the grid's own `indicies`,
 * which hold reference triangle ids (has_tris) or child grid ids.  The pools replace whatever ptb_build_subgrid[_gpu] built; mode
 * TWO_PLANE then walks them exactly like intersectGrid (lightFieldGrid.cpp:71-147): same hits and frames, and PTB_CNT_TRI_TESTS /
 * PTB_CNT_AS_TESTS and the debug_view heat maps are the reference's own numbers (tracer.cpp:127-142), re-traced misses included.
 * Validated on upload (ranges, links, one axis per hierarchy, nesting depth); a later ptb_build_subgrid[_gpu] replaces them again. */
typedef struct ptb_ref_grid {
    float aabb_min[3];
    uint32_t size;
    float aabb_max[3];
    uint32_t has_tris;
    float min[3];
    int32_t splitting_axis;
    float max[3];
    uint32_t n_cells;   /* gridLutStart.size() == gridLutEnd.size() == size^4 */
    float inv_delta[3];
    uint32_t n_indices; /* indicies.size() */
} ptb_ref_grid;
int ptb_upload_subgrid(ptb_ctx *ctx, const ptb_ref_grid *grids, uint32_t n_grids, const uint32_t *lut_start, const uint32_t *lut_end,
                       uint64_t n_cells, const uint32_t *indices, uint64_t n_indices);
uint64_t ptb_ref_subgrid_bytes(const ptb_ctx *ctx); /* getMemory2Plane (lightFieldGrid.cpp:60-68) of the uploaded grids; 0 if none */
/* constructGrid with the size^4 channel loop on the device (lightFieldGrid.cpp:381-414,493-583 with
 * the separating-axis tests of src/types/sat.cpp): one GPU thread per (grid, channel), count -> scan -> fill.  Same pools
 * as ptb_build_subgrid byte for byte (both evaluate csrc/ptb_sat.h). */
int ptb_build_subgrid_gpu(ptb_ctx *ctx, uint32_t gs, uint32_t os, uint32_t m);
double ptb_last_build_device_ms(const ptb_ctx *ctx); /* device part (CUDA events) of the last ptb_build_*_gpu call */
/* FNV-1a over a structure's device pools (mode BVH: nodes + triangle arrays; TWO_PLANE: headers + LUT + indices;
 * SUB_BEAMS: beam records + leaf slices + root tables + planes); 0 if
 * it is not built.  For checking that two builders agree without exposing the layouts. */
uint64_t ptb_structure_digest(ptb_ctx *ctx, int mode);
int ptb_build_subbeams(ptb_ctx *ctx, uint32_t gs, uint32_t m);          /* setGridBeamsSettings + buildGridsWithSubBeams, lightFieldGridSubBeams.h:28-29 */
/* buildGridsWithSubBeams on the device (lightFieldGridSubBeams.cpp:185-374): all beams of one tree depth
 * are one data-parallel step (test against the parent's list -> scan -> fill -> split into 2^4), the depth-first numbering
 * the host builder produces is recovered from subtree sizes.  Same pools as ptb_build_subbeams byte for byte
 * (ptb_structure_digest with mode SUB_BEAMS); the depth guard of ptb_set_build_limits applies. */
int ptb_build_subbeams_gpu(ptb_ctx *ctx, uint32_t gs, uint32_t m);
/* The reference's Subbeam recursion has no depth limit (lightFieldGridSubBeams.cpp:286-324) and does not
 * terminate on dense clusters; beams at `beam_max_depth` become leaves (default 12, 0 = unbounded). */
int ptb_set_build_limits(ptb_ctx *ctx, int beam_max_depth);
uint64_t ptb_structure_bytes(const ptb_ctx *ctx, int mode);             /* getMemoryBVH / getMemory2Plane / getMemoryGridBeams */

/* ---- rendering ------------------------------------------------------------------------------ */
int ptb_set_params(ptb_ctx *ctx, const ptb_params *p); /* setWindowSize + setIntersectMode + getNEE() ... */
int ptb_reset(ptb_ctx *ctx);                           /* initTracer, tracer.cpp:227-244: zero image/samples, reseed rng x<<16|y */
int ptb_clear(ptb_ctx *ctx);                           /* reset(), tracer.cpp:195-217: zero image/samples, RNG streams continue */
int ptb_render(ptb_ctx *ctx, int add_samples);         /* trace(true) with MAX_SAMPLE_COUNT += add_samples, tracer.cpp:164-185; synchronous */
int ptb_render_async(ptb_ctx *ctx, int add_samples);   /* same, returns after enqueueing what can be enqueued; pair with ptb_sync */
int ptb_sync(ptb_ctx *ctx);

/* closest hit of the first-sample primary rays of every pixel (setupRay + findIntersection,
 * tracer.cpp:72-88,129); does not touch the frame state.  Any out pointer may be NULL. */
int ptb_primary_hits(ptb_ctx *ctx, float *t, int32_t *tri, int32_t *mat);

/* closest hit for caller-supplied rays (findIntersection on a fresh ray: tmin 0, tmax inf).
 * od6 = n x {origin.xyz, direction.xyz}.  Host-buffer variant copies in and out. */
int ptb_trace_rays(ptb_ctx *ctx, const float *od6, uint64_t n, int mode, float *t, int32_t *tri);
int ptb_trace_rays_device(ptb_ctx *ctx, const float *od6_device, uint64_t n, int mode, float *t_device, int32_t *tri_device);

/* ptb_trace_rays in BVH mode WITHOUT ANY ALLOCATION.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same kernel, same results as ptb_trace_rays(ctx, od6, n, PTB_MODE_BVH, t, tri) over the library's own tree, but the
 * staged rays, the results and the launch's work cursor live in one file-scope `static __device__` array inside the
 * library (2 MiB + 256 B, held by every GPU that loads libptb.so until the process exits) instead of per-context scratch
 * from cudaMalloc.  n > PTB_STATIC_MAX_RAYS is refused with PTB_ERR_INVALID.  Host pointers in and out; synchronous.  The
 * array is shared by every call on that device and keeps the previous call's contents (the call clears / overwrites what it
 * reads): calls issued one after the other are safe, calls that overlap on the GPU (several host threads or contexts on one
 * device) race.  The scene must be uploaded and ptb_build_bvh[_gpu] done beforehand (those do allocate). */
#define PTB_STATIC_MAX_RAYS 65536
int ptb_trace_rays_static(ptb_ctx *ctx, const float *od6, uint64_t n, float *t, int32_t *tri);
uint64_t ptb_trace_rays_static_max(void); /* PTB_STATIC_MAX_RAYS of the loaded library */

/* ---- read-back ------------------------------------------------------------------------------ */
int ptb_read_image(ptb_ctx *ctx, float *rgba);                       /* tracerGetPixel for every pixel, row-major y*W+x, alpha 1 */
int ptb_read_state(ptb_ctx *ctx, int32_t *samples, uint32_t *rng);   /* samples[][] / randomStates[][] (tracer.cpp:25-26), y*W+x */
int ptb_write_state(ptb_ctx *ctx, const float *rgba, const int32_t *samples, const uint32_t *rng); /* resume a dumped render */
int ptb_read_rgba8(ptb_ctx *ctx, uint8_t *rgba8, int tone_mapping);  /* display transform, window.cpp:507-531 (bytes R,G,B,255), byte-exact: see ptb_tone_thresholds */
/* The 255 steps of that transform (window.cpp:513-522 per channel, tone mapping on): out[b] = the smallest v = c/(1+c) in [0,1]
 * whose byte is >= b, found with the host's libm (the reference's powf); out[0] = 0.  What the kernel searches; no context needed. */
int ptb_tone_thresholds(float out[256]);
int ptb_read_counters(ptb_ctx *ctx, uint64_t out[PTB_CNT__N]);
void *ptb_image_device_ptr(ptb_ctx *ctx);                            /* float4[W*H] in HBM, for an NCCL reduce by the caller */
double ptb_last_render_ms(const ptb_ctx *ctx);                       /* device time of the last ptb_render (CUDA events) */
/* per-stage device time, the GPU analogue of the -pg/gprof split the reference offers (Makefile:15-21):
 * CUDA events on the launching stream around every traversal (extend) and shade/bounce (logic) launch.
 * ms[0]/launches[0] = extend, [1] = logic, accumulated since ptb_reset. */
void *ptb_stream(ptb_ctx *ctx); /* the cudaStream_t every launch and copy of this context is ordered on (for caller-side events / NCCL) */
int ptb_set_profiling(ptb_ctx *ctx, int on);
int ptb_read_stage_ms(ptb_ctx *ctx, double ms[2], uint64_t launches[2]);

/* ---- the GPUs of one box behind one host thread -------------------------------------------------
 * The reference spreads the frame over 16 std::threads with a 4x4 pixel interleave (tracer.cpp:228-231); a group
 * spreads it over n devices with the tile rule of ptb_params (tile (tx,ty) -> member (ty*tiles_x + tx) % n), scene and
 * structures replicated per device.  Scene uploads and structure builds are per member (ptb_group_ctx); the calls
 * below act on all members.  ptb_group_render enqueues the render on every device before it waits, so the devices run
 * concurrently; ptb_group_read_image gathers each pixel from its owner's HBM on member 0 (peer loads over NVLink where
 * the devices can map each other, a staged peer copy otherwise) — the frame is bit-identical to the 1-GPU frame. */
typedef struct ptb_group ptb_group;
int ptb_group_create(const int *devices, int n_devices, ptb_group **out); /* a device may be listed twice (tests on one GPU) */
void ptb_group_destroy(ptb_group *g);
int ptb_group_size(const ptb_group *g);
ptb_ctx *ptb_group_ctx(ptb_group *g, int rank);     /* owned by the group */
const char *ptb_group_last_error(const ptb_group *g);
int ptb_group_peer_direct(const ptb_group *g, int rank); /* 1: member 0 reads this member's frame in place */
enum { PTB_SHARD_TILES = 0, PTB_SHARD_SAMPLES = 1 };
/* TILES (default): bit-identical to one GPU.  SAMPLES: every member renders all pixels with its share of the samples from its
 * own seeds (seed_offset = rank) and ptb_group_read_image re-applies the reference's sum(L)/(N+1) law (tracer.cpp:147-154)
 * over all of them — statistical parity only; for frames with fewer pixels per GPU than a GPU has lanes. */
int ptb_group_set_sharding(ptb_group *g, int shard);
int ptb_group_set_params(ptb_group *g, const ptb_params *p); /* rank / world filled in per member; tile 32x8 if p leaves it 0 */
int ptb_group_reset(ptb_group *g);                  /* initTracer on every member */
int ptb_group_clear(ptb_group *g);                  /* reset() on every member */
int ptb_group_render(ptb_group *g, int add_samples); /* trace(true): all members concurrently */
int ptb_group_read_image(ptb_group *g, float *rgba); /* the whole frame, row-major y*W+x */
int ptb_group_read_counters(ptb_group *g, uint64_t out[PTB_CNT__N]); /* summed over the members */
double ptb_group_last_render_ms(const ptb_group *g); /* max over the members (device time) */

#ifdef __cplusplus
}
#endif
#endif /* PTB_H */
