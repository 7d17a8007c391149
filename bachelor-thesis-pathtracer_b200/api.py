"""ctypes binding of libptb.so (include/ptb.h).  Thin by design: every call forwards to the C ABI;
there is no Python or CPU rendering path here.  If the shared library is missing this module
raises at import of `PathTracer` use — nothing falls back."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PTB_LIB") or os.path.join(_HERE, "libptb.so")  # PTB_LIB: tuning variants only

MODE_ALL, MODE_BVH, MODE_TWO_PLANE, MODE_SUB_BEAMS = 0, 1, 2, 3
MODE_NAMES = {0: "ALL", 1: "BVH", 2: "TWO_PLANE", 3: "SUB_BEAMS"}

CNT_RAYS, CNT_PATHS, CNT_TRI_TESTS, CNT_AS_TESTS, CNT_SHADOW_RAYS, CNT_ITERATIONS, CNT_KERNELS, CNT_REF_RAYS = range(8)
COUNTER_NAMES = ("rays", "paths", "tri_tests", "as_tests", "shadow_rays", "iterations", "kernels", "ref_rays")

# every symbol include/ptb.h declares
SYMBOLS = (
    "ptb_create", "ptb_destroy", "ptb_last_error", "ptb_upload_triangles", "ptb_upload_objects", "ptb_upload_materials",
    "ptb_load_ptbs", "ptb_load_scene_file", "ptb_convert_scene_file", "ptb_set_camera", "ptb_build_bvh", "ptb_build_subgrid", "ptb_build_subbeams", "ptb_structure_bytes",
    "ptb_set_params", "ptb_reset", "ptb_render", "ptb_render_async", "ptb_sync", "ptb_primary_hits", "ptb_trace_rays",
    "ptb_trace_rays_device", "ptb_read_image", "ptb_read_state", "ptb_write_state", "ptb_read_rgba8", "ptb_read_counters",
    "ptb_image_device_ptr", "ptb_last_render_ms", "ptb_set_profiling", "ptb_read_stage_ms", "ptb_stream", "ptb_set_camera_basis", "ptb_clear", "ptb_set_build_limits",
    "ptb_build_bvh_gpu", "ptb_last_build_ms", "ptb_bvh_depth", "ptb_build_subgrid_gpu", "ptb_build_subbeams_gpu", "ptb_last_build_device_ms", "ptb_structure_digest", "ptb_upload_ref_bvh", "ptb_use_ref_bvh", "ptb_upload_textures",
    "ptb_group_create", "ptb_group_destroy", "ptb_group_size", "ptb_group_ctx", "ptb_group_last_error", "ptb_group_peer_direct",
    "ptb_group_set_params", "ptb_group_reset", "ptb_group_clear", "ptb_group_render", "ptb_group_read_image", "ptb_group_read_counters",
    "ptb_group_last_render_ms", "ptb_group_set_sharding", "ptb_trace_rays_static", "ptb_trace_rays_static_max", "ptb_tone_thresholds", "ptb_upload_subgrid", "ptb_ref_subgrid_bytes",
)


class Camera(ctypes.Structure):
    _fields_ = [("origin", ctypes.c_float * 3), ("forward", ctypes.c_float * 3), ("focus", ctypes.c_float),
                ("dof", ctypes.c_float), ("fov_degrees", ctypes.c_float)]


class Params(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "width", "height", "mode", "nee", "primary_only", "max_depth", "reference_quirks", "debug_view", "debug_scale",
        "rank", "world", "tile_w", "tile_h", "count_tests", "seed_offset")]


# ptb_ref_bvh_node (include/ptb.h): BvhNode (bvh.h:14-24) with its box inlined, 48 bytes
REF_BVH_NODE = np.dtype([("bmin", np.float32, 3), ("start_idx", np.int32), ("bmax", np.float32, 3), ("end_idx", np.int32),
                         ("child_left", np.int32), ("child_right", np.int32), ("has_tris", np.int32), ("pad", np.int32)])


# ptb_ref_grid (include/ptb.h): Grid (lightFieldGrid.h:10-22) without its vectors, 80 bytes
REF_GRID = np.dtype([("aabb_min", np.float32, 3), ("size", np.uint32), ("aabb_max", np.float32, 3), ("has_tris", np.uint32),
                     ("min", np.float32, 3), ("splitting_axis", np.int32), ("max", np.float32, 3), ("n_cells", np.uint32),
                     ("inv_delta", np.float32, 3), ("n_indices", np.uint32)])


class Texture(ctypes.Structure):  # ptb_texture
    _fields_ = [("rgba", ctypes.c_void_p), ("width", ctypes.c_int32), ("height", ctypes.c_int32)]


class PtbError(RuntimeError):
    pass


def convert_scene_file(scene_file, out_ptbs, base_dir="", ignore_textures=False):
    """The reference's scene text (.scene + OBJ + MTL) -> .ptbs, on the host (ptb_convert_scene_file)."""
    L = load_library()
    err = ctypes.create_string_buffer(1024)
    if L.ptb_convert_scene_file(os.fsencode(scene_file), os.fsencode(base_dir), os.fsencode(out_ptbs), int(bool(ignore_textures)), err, 1024):
        raise PtbError("ptb_convert_scene_file: " + err.value.decode(errors="replace"))


_lib = None


def load_library():
    """dlopen libptb.so (built by __graft_entry__.build()); loud failure if it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise PtbError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(libptb has no Python/CPU fallback)")
    L = ctypes.CDLL(LIB_PATH)
    vp, ci, cu64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint64
    L.ptb_create.argtypes = [ci, ctypes.POINTER(vp)]
    L.ptb_destroy.argtypes = [vp]
    L.ptb_destroy.restype = None
    L.ptb_last_error.argtypes = [vp]
    L.ptb_last_error.restype = ctypes.c_char_p
    L.ptb_upload_triangles.argtypes = [vp, vp, vp, vp, vp, cu64]
    L.ptb_upload_objects.argtypes = [vp, vp, ctypes.c_uint32]
    L.ptb_upload_materials.argtypes = [vp, vp, ctypes.c_uint32]
    L.ptb_load_ptbs.argtypes = [vp, ctypes.c_char_p]
    L.ptb_load_scene_file.argtypes = [vp, ctypes.c_char_p, ctypes.c_char_p, ci]
    L.ptb_convert_scene_file.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ci, ctypes.c_char_p, cu64]
    L.ptb_upload_textures.argtypes = [vp, vp, ctypes.c_uint32, vp, ctypes.c_uint32]
    L.ptb_group_create.argtypes = [vp, ci, ctypes.POINTER(vp)]
    L.ptb_group_destroy.argtypes = [vp]
    L.ptb_group_destroy.restype = None
    L.ptb_group_size.argtypes = [vp]
    L.ptb_group_ctx.argtypes = [vp, ci]
    L.ptb_group_ctx.restype = vp
    L.ptb_group_last_error.argtypes = [vp]
    L.ptb_group_last_error.restype = ctypes.c_char_p
    L.ptb_group_peer_direct.argtypes = [vp, ci]
    L.ptb_group_set_params.argtypes = [vp, ctypes.POINTER(Params)]
    L.ptb_group_set_sharding.argtypes = [vp, ci]
    for f in ("ptb_group_reset", "ptb_group_clear"):
        getattr(L, f).argtypes = [vp]
    L.ptb_group_render.argtypes = [vp, ci]
    L.ptb_group_read_image.argtypes = [vp, vp]
    L.ptb_group_read_counters.argtypes = [vp, vp]
    L.ptb_group_last_render_ms.argtypes = [vp]
    L.ptb_group_last_render_ms.restype = ctypes.c_double
    L.ptb_set_camera.argtypes = [vp, ctypes.POINTER(Camera)]
    L.ptb_build_bvh.argtypes = [vp]
    L.ptb_build_bvh_gpu.argtypes = [vp]
    L.ptb_last_build_ms.argtypes = [vp]
    L.ptb_last_build_ms.restype = ctypes.c_double
    L.ptb_bvh_depth.argtypes = [vp]
    L.ptb_build_subgrid.argtypes = [vp, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32]
    L.ptb_build_subbeams.argtypes = [vp, ctypes.c_uint32, ctypes.c_uint32]
    L.ptb_build_subbeams_gpu.argtypes = [vp, ctypes.c_uint32, ctypes.c_uint32]
    L.ptb_upload_ref_bvh.argtypes = [vp, vp, ctypes.c_uint32, vp, ctypes.c_uint32, vp, ctypes.c_uint32, ctypes.c_int32]
    L.ptb_use_ref_bvh.argtypes = [vp, ci]
    L.ptb_build_subgrid_gpu.argtypes = [vp, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32]
    L.ptb_last_build_device_ms.argtypes = [vp]
    L.ptb_last_build_device_ms.restype = ctypes.c_double
    L.ptb_structure_digest.argtypes = [vp, ci]
    L.ptb_structure_digest.restype = cu64
    L.ptb_set_build_limits.argtypes = [vp, ci]
    L.ptb_structure_bytes.argtypes = [vp, ci]
    L.ptb_structure_bytes.restype = cu64
    L.ptb_set_params.argtypes = [vp, ctypes.POINTER(Params)]
    L.ptb_reset.argtypes = [vp]
    L.ptb_clear.argtypes = [vp]
    L.ptb_set_camera_basis.argtypes = [vp, vp, vp, vp, vp, ctypes.c_float, ctypes.c_float, ctypes.c_float]
    L.ptb_render.argtypes = [vp, ci]
    L.ptb_render_async.argtypes = [vp, ci]
    L.ptb_sync.argtypes = [vp]
    L.ptb_primary_hits.argtypes = [vp, vp, vp, vp]
    L.ptb_trace_rays.argtypes = [vp, vp, cu64, ci, vp, vp]
    L.ptb_trace_rays_device.argtypes = [vp, vp, cu64, ci, vp, vp]
    L.ptb_tone_thresholds.argtypes = [vp]
    L.ptb_upload_subgrid.argtypes = [vp, vp, ctypes.c_uint32, vp, vp, cu64, vp, cu64]
    L.ptb_ref_subgrid_bytes.restype = cu64
    L.ptb_ref_subgrid_bytes.argtypes = [vp]
    L.ptb_trace_rays_static.argtypes = [vp, vp, cu64, vp, vp]
    L.ptb_trace_rays_static_max.restype = cu64
    L.ptb_trace_rays_static_max.argtypes = []
    L.ptb_read_image.argtypes = [vp, vp]
    L.ptb_read_state.argtypes = [vp, vp, vp]
    L.ptb_write_state.argtypes = [vp, vp, vp, vp]
    L.ptb_read_rgba8.argtypes = [vp, vp, ci]
    L.ptb_read_counters.argtypes = [vp, vp]
    L.ptb_image_device_ptr.argtypes = [vp]
    L.ptb_image_device_ptr.restype = vp
    L.ptb_last_render_ms.argtypes = [vp]
    L.ptb_last_render_ms.restype = ctypes.c_double
    L.ptb_stream.argtypes = [vp]
    L.ptb_stream.restype = vp
    L.ptb_set_profiling.argtypes = [vp, ci]
    L.ptb_read_stage_ms.argtypes = [vp, vp, vp]
    _lib = L
    return L


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


class PathTracer:
    """One context on one GPU.  Mirrors the reference's call sequence (src/main.cpp:91-125):
    settings -> load scene -> build structures -> initTracer -> trace."""

    def __init__(self, device=0, _borrowed=None):
        self.L = load_library()
        self._owned = _borrowed is None
        if _borrowed is not None:  # a member of a PathTracerGroup: the group owns the context
            h = ctypes.c_void_p(_borrowed)
        else:
            h = ctypes.c_void_p()
            rc = self.L.ptb_create(device, ctypes.byref(h))
            if rc != 0:
                raise PtbError(f"ptb_create: {self.L.ptb_last_error(None).decode()} (rc {rc})")
        self.h = h
        self.device = device
        self.params = Params(reference_quirks=1, world=1, tile_w=32, tile_h=8, mode=MODE_BVH)

    def _ck(self, rc, what):
        if rc != 0:
            raise PtbError(f"{what}: {self.L.ptb_last_error(self.h).decode()} (rc {rc})")

    def close(self):
        if getattr(self, "h", None):
            if self._owned:
                self.L.ptb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- scene -------------------------------------------------------------------------------
    def load_ptbs(self, path):
        self._ck(self.L.ptb_load_ptbs(self.h, os.fsencode(path)), "ptb_load_ptbs")

    def load_scene_file(self, scene_file, base_dir="", ignore_textures=False):
        """.scene / OBJ / MTL text like the reference's loadScene + initScene + loadObject (ptb_load_scene_file)"""
        self._ck(self.L.ptb_load_scene_file(self.h, os.fsencode(scene_file), os.fsencode(base_dir), int(bool(ignore_textures))),
                 "ptb_load_scene_file")

    def upload_triangles(self, pos9, nrm9, uv6, mat):
        pos9 = np.ascontiguousarray(pos9, np.float32).reshape(-1, 9)
        n = pos9.shape[0]
        nrm9 = np.ascontiguousarray(nrm9, np.float32).reshape(n, 9)
        uv6 = np.ascontiguousarray(uv6, np.float32).reshape(n, 6) if uv6 is not None else None
        mat = np.ascontiguousarray(mat, np.int32).reshape(n)
        self._ck(self.L.ptb_upload_triangles(self.h, _ptr(pos9), _ptr(nrm9), _ptr(uv6), _ptr(mat), n), "ptb_upload_triangles")

    def upload_objects(self, objects_bytes, n):
        buf = (ctypes.c_char * len(objects_bytes)).from_buffer_copy(objects_bytes)
        self._ck(self.L.ptb_upload_objects(self.h, ctypes.cast(buf, ctypes.c_void_p), n), "ptb_upload_objects")

    def upload_materials(self, materials_bytes, n):
        buf = (ctypes.c_char * len(materials_bytes)).from_buffer_copy(materials_bytes)
        self._ck(self.L.ptb_upload_materials(self.h, ctypes.cast(buf, ctypes.c_void_p), n), "ptb_upload_materials")

    def upload_textures(self, textures, material_tex2):
        """textures: list of float32 arrays (height, width, 4); material_tex2: per material (albedo texture, normal map), -1 none."""
        keep = [np.ascontiguousarray(t, np.float32) for t in textures]
        arr = (Texture * max(1, len(keep)))()
        for i, t in enumerate(keep):
            if t.ndim != 3 or t.shape[2] != 4:
                raise ValueError("texture must be (height, width, 4)")
            arr[i] = Texture(t.ctypes.data, t.shape[1], t.shape[0])
        b = np.ascontiguousarray(material_tex2, np.int32).reshape(-1, 2)
        self._ck(self.L.ptb_upload_textures(self.h, ctypes.cast(arr, ctypes.c_void_p), len(keep), _ptr(b), b.shape[0]), "ptb_upload_textures")

    def set_camera(self, origin, forward, focus=1.0, dof=0.0, fov_degrees=70.0):
        cam = Camera((ctypes.c_float * 3)(*origin), (ctypes.c_float * 3)(*forward), focus, dof, fov_degrees)
        self._ck(self.L.ptb_set_camera(self.h, ctypes.byref(cam)), "ptb_set_camera")

    # ---- structures --------------------------------------------------------------------------
    def build_bvh(self, gpu=False):
        """constructBVH: host binned-SAH builder, or (gpu=True) the Morton/radix-tree builder that runs in HBM."""
        if gpu:
            self._ck(self.L.ptb_build_bvh_gpu(self.h), "ptb_build_bvh_gpu")
        else:
            self._ck(self.L.ptb_build_bvh(self.h), "ptb_build_bvh")
        return float(self.L.ptb_last_build_ms(self.h))

    def bvh_depth(self):
        return int(self.L.ptb_bvh_depth(self.h))

    def build_subgrid(self, gs=10, os_=5, m=40, gpu=False):
        """constructGrid: channels filled by the host threads, or (gpu=True) one GPU thread per channel.  Returns
        (wall ms, device ms) of the build."""
        import time
        t0 = time.perf_counter()
        if gpu:
            self._ck(self.L.ptb_build_subgrid_gpu(self.h, gs, os_, m), "ptb_build_subgrid_gpu")
        else:
            self._ck(self.L.ptb_build_subgrid(self.h, gs, os_, m), "ptb_build_subgrid")
        return (time.perf_counter() - t0) * 1e3, (float(self.L.ptb_last_build_device_ms(self.h)) if gpu else 0.0)

    def upload_ref_bvh(self, nodes, indices, object_roots, scene_root):
        """nodes: structured array of REF_BVH_NODE (the reference's BvhNode + its box); see ptb_upload_ref_bvh."""
        nodes = np.ascontiguousarray(nodes, REF_BVH_NODE)
        indices = np.ascontiguousarray(indices, np.uint32)
        object_roots = np.ascontiguousarray(object_roots, np.int32)
        self._ck(self.L.ptb_upload_ref_bvh(self.h, _ptr(nodes), nodes.shape[0], _ptr(indices), indices.shape[0], _ptr(object_roots),
                                           object_roots.shape[0], int(scene_root)), "ptb_upload_ref_bvh")

    def upload_subgrid(self, grids, lut_start, lut_end, indices):
        """The reference's own Subgrid hierarchy (structured array of REF_GRID + its pools); see ptb_upload_subgrid."""
        grids = np.ascontiguousarray(grids, REF_GRID)
        lut_start = np.ascontiguousarray(lut_start, np.uint32)
        lut_end = np.ascontiguousarray(lut_end, np.uint32)
        indices = np.ascontiguousarray(indices, np.uint32)
        self._ck(self.L.ptb_upload_subgrid(self.h, _ptr(grids), grids.shape[0], _ptr(lut_start), _ptr(lut_end), lut_start.shape[0],
                                           _ptr(indices), indices.shape[0]), "ptb_upload_subgrid")

    def use_ref_bvh(self, on=True):
        """Mode BVH walks the uploaded reference tree: counters and heat maps become the reference's own numbers."""
        self._ck(self.L.ptb_use_ref_bvh(self.h, int(on)), "ptb_use_ref_bvh")

    def structure_digest(self, mode):
        return int(self.L.ptb_structure_digest(self.h, mode))

    def build_subbeams(self, gs=10, m=40, max_depth=None, gpu=False):
        """gpu=True: the level-synchronous builder on the device (ptb_build_subbeams_gpu), returns (wall ms, device ms)"""
        if max_depth is not None:
            self._ck(self.L.ptb_set_build_limits(self.h, max_depth), "ptb_set_build_limits")
        if gpu:
            self._ck(self.L.ptb_build_subbeams_gpu(self.h, gs, m), "ptb_build_subbeams_gpu")
            return float(self.L.ptb_last_build_ms(self.h)), float(self.L.ptb_last_build_device_ms(self.h))
        self._ck(self.L.ptb_build_subbeams(self.h, gs, m), "ptb_build_subbeams")

    def structure_bytes(self, mode):
        return int(self.L.ptb_structure_bytes(self.h, mode))

    # ---- rendering ---------------------------------------------------------------------------
    def set_params(self, **kw):
        for k, v in kw.items():
            if not hasattr(self.params, k):
                raise AttributeError(k)
            setattr(self.params, k, int(v))
        self._ck(self.L.ptb_set_params(self.h, ctypes.byref(self.params)), "ptb_set_params")

    def reset(self):
        self._ck(self.L.ptb_reset(self.h), "ptb_reset")

    def clear(self):
        """reset() of the reference: image and sample counts zeroed, per-pixel RNG streams continue."""
        self._ck(self.L.ptb_clear(self.h), "ptb_clear")

    def set_camera_basis(self, origin, forward, right, up, focus, dof, fov_scale):
        v = [(ctypes.c_float * 3)(*x) for x in (origin, forward, right, up)]
        self._ck(self.L.ptb_set_camera_basis(self.h, *v, focus, dof, fov_scale), "ptb_set_camera_basis")

    def render(self, add_samples):
        self._ck(self.L.ptb_render(self.h, add_samples), "ptb_render")
        return self.last_ms()

    def render_async(self, add_samples):
        """Enqueue only (path kernels); finish with sync()."""
        self._ck(self.L.ptb_render_async(self.h, add_samples), "ptb_render_async")

    def sync(self):
        self._ck(self.L.ptb_sync(self.h), "ptb_sync")

    def last_ms(self):
        return float(self.L.ptb_last_render_ms(self.h))

    def primary_hits(self):
        n = self.params.width * self.params.height
        t = np.empty(n, np.float32)
        tri = np.empty(n, np.int32)
        mat = np.empty(n, np.int32)
        self._ck(self.L.ptb_primary_hits(self.h, _ptr(t), _ptr(tri), _ptr(mat)), "ptb_primary_hits")
        return t, tri, mat

    def trace_rays(self, od6, mode):
        od6 = np.ascontiguousarray(od6, np.float32).reshape(-1, 6)
        n = od6.shape[0]
        t = np.empty(n, np.float32)
        tri = np.empty(n, np.int32)
        self._ck(self.L.ptb_trace_rays(self.h, _ptr(od6), n, mode, _ptr(t), _ptr(tri)), "ptb_trace_rays")
        return t, tri

    def trace_rays_static(self, od6):
        """ptb_trace_rays in BVH mode with the scratch in the library's static device array: no allocation in the call."""
        od6 = np.ascontiguousarray(od6, np.float32).reshape(-1, 6)
        n = od6.shape[0]
        t = np.empty(n, np.float32)
        tri = np.empty(n, np.int32)
        self._ck(self.L.ptb_trace_rays_static(self.h, _ptr(od6), n, _ptr(t), _ptr(tri)), "ptb_trace_rays_static")
        return t, tri

    def trace_rays_device(self, od6_ptr, n, mode, t_ptr, tri_ptr):
        self._ck(self.L.ptb_trace_rays_device(self.h, od6_ptr, n, mode, t_ptr, tri_ptr), "ptb_trace_rays_device")
        return self.last_ms()

    # ---- read-back ---------------------------------------------------------------------------
    def read_image(self):
        img = np.empty((self.params.height, self.params.width, 4), np.float32)
        self._ck(self.L.ptb_read_image(self.h, _ptr(img)), "ptb_read_image")
        return img

    def read_state(self):
        n = self.params.width * self.params.height
        s = np.empty(n, np.int32)
        r = np.empty(n, np.uint32)
        self._ck(self.L.ptb_read_state(self.h, _ptr(s), _ptr(r)), "ptb_read_state")
        return s, r

    def write_state(self, rgba, samples, rng):
        rgba = np.ascontiguousarray(rgba, np.float32)
        samples = np.ascontiguousarray(samples, np.int32)
        rng = np.ascontiguousarray(rng, np.uint32)
        self._ck(self.L.ptb_write_state(self.h, _ptr(rgba), _ptr(samples), _ptr(rng)), "ptb_write_state")

    def read_rgba8(self, tone_mapping=True):
        out = np.empty((self.params.height, self.params.width, 4), np.uint8)
        self._ck(self.L.ptb_read_rgba8(self.h, _ptr(out), int(tone_mapping)), "ptb_read_rgba8")
        return out

    def counters(self):
        c = np.zeros(8, np.uint64)
        self._ck(self.L.ptb_read_counters(self.h, _ptr(c)), "ptb_read_counters")
        return dict(zip(COUNTER_NAMES, (int(v) for v in c)))

    def stream_handle(self):
        return int(self.L.ptb_stream(self.h) or 0)

    def set_profiling(self, on=True):
        self._ck(self.L.ptb_set_profiling(self.h, int(on)), "ptb_set_profiling")

    def stage_ms(self):
        """{'extend_ms', 'logic_ms', 'extend_launches', 'logic_launches'} since reset (needs set_profiling)."""
        ms = np.zeros(2, np.float64)
        n = np.zeros(2, np.uint64)
        self._ck(self.L.ptb_read_stage_ms(self.h, _ptr(ms), _ptr(n)), "ptb_read_stage_ms")
        return {"extend_ms": float(ms[0]), "logic_ms": float(ms[1]), "extend_launches": int(n[0]), "logic_launches": int(n[1])}

    def image_device_ptr(self):
        return int(self.L.ptb_image_device_ptr(self.h) or 0)


class PathTracerGroup:
    """ptb_group_*: one context per listed device behind one host thread; the frame is tile-sharded over the members
    and gathered on member 0 (bit-identical to the 1-GPU frame).  `members[i]` is a PathTracer view of member i for
    the per-member calls (uploads, structure builds); `each(fn)` applies one to all."""

    def __init__(self, devices):
        self.L = load_library()
        devs = (ctypes.c_int * len(devices))(*devices)
        g = ctypes.c_void_p()
        rc = self.L.ptb_group_create(ctypes.cast(devs, ctypes.c_void_p), len(devices), ctypes.byref(g))
        if rc != 0:
            raise PtbError(f"ptb_group_create: {self.L.ptb_group_last_error(None).decode()} (rc {rc})")
        self.g = g
        self.members = [PathTracer(d, _borrowed=self.L.ptb_group_ctx(g, i)) for i, d in enumerate(devices)]
        self.params = Params(reference_quirks=1, world=len(devices), tile_w=32, tile_h=8, mode=MODE_BVH)

    def _ck(self, rc, what):
        if rc != 0:
            raise PtbError(f"{what}: {self.L.ptb_group_last_error(self.g).decode()} (rc {rc})")

    def close(self):
        if getattr(self, "g", None):
            for m in self.members:
                m.h = None
            self.L.ptb_group_destroy(self.g)
            self.g = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def each(self, fn):
        return [fn(m) for m in self.members]

    def set_sharding(self, samples):
        """False: tiles (bit-identical to one GPU).  True: sample ranges from per-member seeds (statistical parity)."""
        self._ck(self.L.ptb_group_set_sharding(self.g, 1 if samples else 0), "ptb_group_set_sharding")

    def peer_direct(self):
        return [bool(self.L.ptb_group_peer_direct(self.g, i)) for i in range(len(self.members))]

    def set_params(self, **kw):
        for k, v in kw.items():
            if not hasattr(self.params, k):
                raise AttributeError(k)
            setattr(self.params, k, int(v))
        self._ck(self.L.ptb_group_set_params(self.g, ctypes.byref(self.params)), "ptb_group_set_params")
        for i, m in enumerate(self.members):  # keep the member views' idea of the frame size in step
            ctypes.memmove(ctypes.byref(m.params), ctypes.byref(self.params), ctypes.sizeof(Params))
            m.params.rank, m.params.world = i, len(self.members)

    def reset(self):
        self._ck(self.L.ptb_group_reset(self.g), "ptb_group_reset")

    def clear(self):
        self._ck(self.L.ptb_group_clear(self.g), "ptb_group_clear")

    def render(self, add_samples):
        self._ck(self.L.ptb_group_render(self.g, add_samples), "ptb_group_render")
        return float(self.L.ptb_group_last_render_ms(self.g))

    def read_image(self):
        img = np.empty((self.params.height, self.params.width, 4), np.float32)
        self._ck(self.L.ptb_group_read_image(self.g, _ptr(img)), "ptb_group_read_image")
        return img

    def counters(self):
        c = np.zeros(8, np.uint64)
        self._ck(self.L.ptb_group_read_counters(self.g, _ptr(c)), "ptb_group_read_counters")
        return dict(zip(COUNTER_NAMES, (int(v) for v in c)))
