"""Drop-in check of the boundary: the reference's OWN compiled objects (scene.cpp, shader.cpp,
camera.cpp, bvh.cpp ... from oracle/_ref/obj) linked with host/tracer_facade.cpp in place of
src/tracer.cpp.  The reference-side driver calls setWindowSize / initTracer / trace(true) /
tracerGetPixel exactly as src/main.cpp does; the frame must equal the CPU oracle's."""
import numpy as np
import pytest

from oracle_lib import MODE_BVH, MODE_SUB_BEAMS, MODE_TWO_PLANE, Oracle, have_facade, have_facade_gpu, ptfacade_render

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not have_facade(), reason="oracle/_ref/ptfacade not built (needs the reference sources at build time)")
@pytest.mark.parametrize("mode,nee,max_depth", [(MODE_BVH, 0, 0), (MODE_BVH, 1, 5), (MODE_TWO_PLANE, 1, 0), (MODE_SUB_BEAMS, 0, 3)])
def test_reference_host_code_renders_through_the_facade(cornel_ptbs, ptb, mode, nee, max_depth):
    W, H, spp = 96, 64, 3
    img, rec = ptfacade_render(cornel_ptbs, W, H, mode, spp, max_depth, nee)
    o = Oracle(cornel_ptbs)
    o.build(grid=True, beams=True)
    o.configure(W, H, mode, nee=nee, max_depth=max_depth)
    rimg, _, _, st = o.render(spp)
    o.close()
    diff = np.abs(img[..., :3] - rimg[..., :3]) / np.maximum(np.abs(rimg[..., :3]), 1e-3)
    nbits = int((img.view(np.uint32) != rimg.view(np.uint32)).any(axis=-1).sum())
    assert diff.max() <= 2e-6 or nbits <= max(1, W * H // 2000), f"{nbits} pixels differ, max rel {diff.max():.3g}"
    assert rec["paths"] == W * H * spp
    if nee == 0:  # with NEE the GPU count also holds the shadow rays, which the path-ray statistic of the oracle does not
        assert rec["rays_unique"] == st["rays_unique"], "GPU traversal count != the reference's unique-ray count"


@pytest.mark.skipif(not have_facade(), reason="oracle/_ref/ptfacade not built (needs the reference sources at build time)")
def test_facade_forwards_the_references_textures(cornel_ptbs, ptb, tmp_path):
    """The reference-side loader fills Material::pbr.texture / .normal (here from the v2 container); the facade hands
    Texture::data to ptb_upload_textures, and the frame equals the reference's own golden image of the textured scene."""
    import os
    sc = ptb.scenes
    path = str(tmp_path / "tex.ptbs")
    sc.write_ptbs(path, sc.textured_cornell(sc.read_ptbs(cornel_ptbs), light_texture=True))
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cornel_textured_ref_48.npz"))
    img, rec = ptfacade_render(path, 64, 48, MODE_BVH, 3, 4, 1)
    n = int((img.view(np.uint32) != gold["img_m1_nee1_d4_l1"].view(np.uint32)).any(axis=-1).sum())
    assert n <= 1, f"{n} pixels differ from the reference's textured frame"


@pytest.mark.skipif(not have_facade(), reason="oracle/_ref/ptfacade not built (needs the reference sources at build time)")
@pytest.mark.parametrize("mode", [MODE_BVH, MODE_TWO_PLANE])
def test_facade_over_several_devices_renders_the_same_frame(cornel_ptbs, ptb, mode):
    """ptbFacadeSetDevices: the one trace() thread of the reference drives a ptb_group; the frame it serves through
    tracerGetPixel is the single-device frame bit for bit.  With one GPU on the box the members share device 0."""
    import torch
    W, H, spp = 96, 64, 3
    one, rec1 = ptfacade_render(cornel_ptbs, W, H, mode, spp, 5, 1)
    n = torch.cuda.device_count()
    extra = ("--gpus", n) if n > 1 else ("--gpus", 3, "--same-device", 1)
    many, recn = ptfacade_render(cornel_ptbs, W, H, mode, spp, 5, 1, extra=extra)
    assert (one.view(np.uint32) == many.view(np.uint32)).all()
    assert recn["paths"] == rec1["paths"] == W * H * spp and recn["rays_unique"] == rec1["rays_unique"]


# ---- the structure half of the boundary: bvh.h / lightFieldGrid.h / lightFieldGridSubBeams.h fronted too (host/structures_facade.cpp)
@pytest.mark.skipif(not have_facade_gpu(), reason="oracle/_ref/ptfacade_gpu not built (needs the reference sources at build time)")
@pytest.mark.parametrize("mode,nee,max_depth", [(MODE_BVH, 1, 5), (MODE_TWO_PLANE, 1, 0), (MODE_SUB_BEAMS, 0, 3)])
def test_structure_headers_forward_to_the_gpu_builders(cornel_ptbs, ptb, mode, nee, max_depth):
    """The reference's scene.cpp / shader.cpp / camera.cpp objects linked WITHOUT its bvh.o, lightFieldGrid.o and
    lightFieldGridSubBeams.o: constructBVH / constructGrid / buildGridsWithSubBeams / setGridSettings / getMemory* are the
    facade's.  Same frame as the CPU oracle; the memory the driver prints (getMemoryBVH / getMemory2Plane, main.cpp:136) is the
    device structure's."""
    W, H, spp = 96, 64, 3
    img, rec, build = ptfacade_render(cornel_ptbs, W, H, mode, spp, max_depth, nee, structures=True, want_build=True)
    o = Oracle(cornel_ptbs)
    o.build(grid=True, beams=True)
    o.configure(W, H, mode, nee=nee, max_depth=max_depth)
    rimg, _, _, _ = o.render(spp)
    o.close()
    nbits = int((img.view(np.uint32) != rimg.view(np.uint32)).any(axis=-1).sum())
    assert nbits <= max(1, W * H // 2000), f"{nbits} pixels differ"
    assert rec["paths"] == W * H * spp
    assert build["bvh_bytes"] > 0
    if mode == MODE_TWO_PLANE:
        assert build["grid_bytes"] > build["bvh_bytes"] // 4


@pytest.mark.skipif(not have_facade_gpu(), reason="oracle/_ref/ptfacade_gpu not built (needs the reference sources at build time)")
def test_initscene_builds_lucy_level_6_in_under_a_second(ptb, cornel_ptbs, tmp_path):
    """32,790 triangles: the reference's constructBVH + constructGrid take minutes (7.2 s + 64 s at 1e5 triangles);
    through the fronted headers the same initScene calls finish in well under a second, and the frame is the BVH frame of
    the plain facade."""
    sc = ptb.scenes
    path = str(tmp_path / "lucy6.ptbs")
    sc.write_ptbs(path, sc.lucy_standin(sc.read_ptbs(cornel_ptbs), level=6))
    W, H, spp = 160, 90, 2
    img, rec, build = ptfacade_render(path, W, H, MODE_TWO_PLANE, spp, 5, 0, extra=("--gs", 8, "--os", 8, "--m", 50), structures=True,
                                      want_build=True)
    print(f"initScene through the fronted headers: bvh {build['bvh_ms']:.1f} ms (incl. context creation and upload), grid {build['grid_ms']:.1f} ms")
    assert build["grid_ms"] < 1000.0 and build["bvh_ms"] < 5000.0
    assert rec["paths"] == W * H * spp
    pt = ptb.PathTracer(0)
    pt.load_ptbs(path)
    pt.build_bvh()
    pt.build_subgrid(8, 8, 50, gpu=True)
    pt.set_params(width=W, height=H, mode=MODE_TWO_PLANE, nee=0, max_depth=5)
    pt.reset()
    pt.render(spp)
    mine = pt.read_image()
    pt.close()
    assert (img.view(np.uint32) == mine.view(np.uint32)).all()
