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
    // stbi_load (src/types/texture.cpp:34) has no counterpart here: a material that names an EXISTING texture file makes
    // the load fail unless this flag says to drop the textures (the material keeps its Kd colour); a file that does not
    // exist is skipped with the reference's "could not load" message, as loadTexture does (texture.cpp:36-39)
    kSceneIgnoreTextures = 1,
};

struct LoadedScene {
    std::vector<float> pos9, nrm9, uv6; // Triangle::vertices / ::normal / ::uv, 9 / 9 / 6 floats per triangle
    std::vector<int32_t> mat;           // Triangle::materialIdx
    std::vector<ptbs_object> objects;
    std::vector<ptbs_material> materials;
    std::vector<std::array<std::string, 2>> textures; // per material: map_Kd, bump file (resolved path) or ""
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

// the .ptbs v1 container (include/ptbs_format.h)
bool write_ptbs(const std::string &path, const LoadedScene &scene, std::string &err);

} // namespace ptb
