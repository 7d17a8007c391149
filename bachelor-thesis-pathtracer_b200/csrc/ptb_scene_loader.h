// ptb_scene_loader.h — .scene / OBJ / MTL text straight into the flat arrays libptb.so uploads (SURVEY.md §8f n4:
// "OBJ -> SoA direct loader, skip the 280-byte AoS").  Host-only C++, no CUDA.
//
// Stands in for loadScene (reference: src/scene/SceneFile.cpp:21-110), the model / material part of initScene
// (src/scene/scene.cpp:97-110) and loadObject (src/primitives/object.cpp:35-174).  The reference parses OBJ/MTL with
// tinyobjloader 2.0.0 (vendored header, includes/singleHeaderLibs/tiny_obj_loader.h; default options: triangulate,
// real_t = float); the pieces of its published behaviour the arrays depend on are restated in the .cpp, each with the
// line range it follows.  The result is what the reference holds in getTris() / getObjects() / getMaterials() after
// initScene — field for field what the compiled reference's scene dump holds (tests/golden/cornel.ptbs), which is how
// tests/test_scene_loader.py checks it: byte equality of the .ptbs.
#pragma once
#include <array>
#include <cstdint>
#include <string>
#include <vector>

#include "ptbs_format.h"

namespace ptb {

enum SceneLoadFlags : int {
    // Texture files (map_Kd, bump) are decoded when they are PNG or TGA (ptb_png.cpp); any other format stbi_load
    // (src/types/texture.cpp:34) would read makes the load fail unless this flag says to drop the textures (the material
    // keeps its Kd colour).  A file that does not exist is skipped with the reference's "could not load" message, as
    // loadTexture does (texture.cpp:36-39)
    kSceneIgnoreTextures = 1,
};

struct LoadedScene {
    std::vector<float> pos9, nrm9, uv6; // Triangle::vertices / ::normal / ::uv, 9 / 9 / 6 floats per triangle
    std::vector<int32_t> mat;           // Triangle::materialIdx
    std::vector<ptbs_object> objects;
    std::vector<ptbs_material> materials;
    std::vector<std::array<std::string, 2>> textures; // per material: map_Kd, bump file (resolved path) or ""
    // filled by load_scene_file / decode_textures: Texture::data as loadTexture leaves it (byte / 255.0f, texture.cpp:41-50),
    // numbered in material order, albedo before normal map
    struct Texels { int width = 0, height = 0; std::vector<float> rgba; };
    std::vector<Texels> texels;
    std::vector<int32_t> mat_tex; // 2 per material: index into texels or -1
    float cam_pos[3] = {0, 0, 0}, cam_forward[3] = {0, 0, 0}, cam_focus = 0, cam_dof = 0;
    std::string log; // what the reference prints while loading + tinyobj-style warnings
    size_t triangles() const { return mat.size(); }
};

// loadObject: appends the file's shapes and materials.  `default_material`: index used by faces without usemtl.
bool load_obj(const std::string &file, const float position[3], const float size[3], int default_material, LoadedScene &scene,
              std::string &err, int flags = 0);

// loadScene + the model / material loops of initScene.  Model paths are relative to `base_dir` (empty: the working
// directory, like the reference, which is started inside showcase/).
bool load_scene_file(const std::string &scene_file, const std::string &base_dir, LoadedScene &scene, std::string &err, int flags = 0);

// decodes the files named in `textures` into `texels` / `mat_tex` (load_scene_file calls it after the overrides)
bool decode_textures(LoadedScene &scene, std::string &err);

// the .ptbs container (include/ptbs_format.h): v1, or v2 when the scene holds texels
bool write_ptbs(const std::string &path, const LoadedScene &scene, std::string &err);

} // namespace ptb
