/* ptbs_format.h — the flat binary scene container (".ptbs") that crosses the boundary
 * between the host scene loader and libptb.so.
 *
 * It holds exactly what the reference keeps in memory after its loader ran
 * (reference: src/scene/scene.cpp:97-131 `initScene`, src/primitives/object.cpp:35-174
 * `loadObject`): the triangle buffer (`getTris()`), the object table (`getObjects()`),
 * the material table (`getMaterials()`) and the camera from the scene file
 * (src/scene/SceneFile.h:26-31).  ptb_convert_scene_file (csrc/ptb_scene_loader.cpp) writes it
 * from .scene / OBJ / MTL text; oracle/ref/ref_driver.cpp's dump-scene writes it from the
 * reference's own loader, and the two agree byte for byte (tests/test_scene_loader.py).
 *
 * Layout (little endian, all blocks tightly packed, in this order):
 *   ptbs_header
 *   float    pos[n_tris][9]      v0.xyz v1.xyz v2.xyz          (Triangle::vertices)
 *   float    nrm[n_tris][9]      n0.xyz n1.xyz n2.xyz          (Triangle::normal)
 *   float    uv [n_tris][6]      uv0 uv1 uv2                   (Triangle::uv)
 *   int32_t  mat[n_tris]                                       (Triangle::materialIdx)
 *   ptbs_object   objects[n_objects]
 *   ptbs_material materials[n_materials]
 * version 2 only (n_textures = header.reserved[0] > 0; a version-1 file simply ends here):
 *   int32_t  mat_tex[n_materials][2]   albedo texture, normal map of each material (-1 = none)
 *                                      (pbrProbs::texture / ::normal, src/shader/shader.h:14-24)
 *   n_textures x { ptbs_texture_hdr; float rgba[height][width][4] }   Texture::data as loadTexture leaves it
 *                                      (byte / 255.0f per channel, src/types/texture.cpp:33-50)
 */
#ifndef PTBS_FORMAT_H
#define PTBS_FORMAT_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTBS_MAGIC 0x53425450u /* "PTBS" */
#define PTBS_VERSION 1u
#define PTBS_VERSION_TEXTURES 2u
#define PTBS_MAX_OBJECT_MATERIALS 28

typedef struct ptbs_header {
    uint32_t magic;
    uint32_t version;
    uint32_t n_tris;
    uint32_t n_objects;
    uint32_t n_materials;
    uint32_t reserved[3]; /* [0] = n_textures (version 2) */
    /* camera: src/scene/SceneFile.h:26-31, applied at src/scene/scene.cpp:127-130 */
    float cam_pos[3];
    float cam_forward[3]; /* as written in the scene file (not yet normalised) */
    float cam_focus;
    float cam_dof;
} ptbs_header;

/* src/primitives/object.h:10-21 — only the fields the hot path reads */
typedef struct ptbs_object {
    int32_t start_idx; /* first triangle */
    int32_t end_idx;   /* one past the last triangle */
    float surface_area;
    uint32_t n_materials;
    uint32_t materials[PTBS_MAX_OBJECT_MATERIALS];
    char name[64];
} ptbs_object;

/* src/shader/shader.h:14-35 (pbrProbs + ShaderMix), without textures */
typedef struct ptbs_material {
    float albedo[3];
    float ior1;
    float ior2;
    float emission;
    float roughness;
    float w_lambert;
    float w_reflection;
    float w_refraction;
    char name[64];
} ptbs_material;

/* src/types/texture.h:7-11 */
typedef struct ptbs_texture_hdr {
    int32_t width;
    int32_t height;
} ptbs_texture_hdr;

#ifdef __cplusplus
}
#endif
#endif /* PTBS_FORMAT_H */
