"""ctypes binding of libhvcf.so (include/hvcf.h) -- the C ABI of the B200 light-culling path.

This is plumbing only: it loads the in-tree CUDA library and exposes its entry points.  There is no
CPU fallback; if the library is missing or no sm_100 device is present, every call raises.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HVCF_LIB") or os.path.join(_HERE, "lib", "libhvcf.so")  # HVCF_LIB: experiment builds only

LIGHT_DT = np.dtype([("position", "<f4", 4), ("color", "<f4", 4), ("radius", "<f4"), ("pad_", "<f4", 3)])
AABB_DT = np.dtype([("min_point", "<f4", 4), ("max_point", "<f4", 4)])
CELL_DT = np.dtype([("offset", "<u4"), ("count", "<u4")])
CMD_DT = np.dtype([("vertex_count", "<u4"), ("instance_count", "<u4"), ("first_vertex", "<u4"), ("first_instance", "<u4")])
UBO_BYTES = 176
FRUSTUM_BYTES = 224

HVCF_OK = 0
HVCF_ERR_INVALID_ARGUMENT, HVCF_ERR_CUDA, HVCF_ERR_OUT_OF_MEMORY, HVCF_ERR_CAPACITY, HVCF_ERR_NOT_READY, HVCF_ERR_UNSUPPORTED = -1, -2, -3, -4, -5, -6
MODE_AUTO = 0
MODE_BRUTE = 1
MODE_HIERARCHICAL = 2


class HvcfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"hvcf error {code}: {msg}")
        self.code = code


class Config(ctypes.Structure):
    _fields_ = [("struct_size", ctypes.c_uint32), ("device", ctypes.c_int32),
                ("slice_count_x", ctypes.c_uint32), ("slice_count_y", ctypes.c_uint32), ("slice_count_z", ctypes.c_uint32),
                ("max_lights_per_cluster", ctypes.c_uint32), ("max_lights", ctypes.c_uint32), ("frame_count", ctypes.c_uint32),
                ("mode", ctypes.c_uint32), ("z_begin", ctypes.c_uint32), ("z_end", ctypes.c_uint32),
                ("row_begin", ctypes.c_uint32), ("row_end", ctypes.c_uint32), ("reserved", ctypes.c_uint32 * 3)]


class ShardDesc(ctypes.Structure):
    """hvcf_shard_desc: this rank's place in the 2-D decomposition (z-slabs of clusters x owned ranges of lights)."""
    _fields_ = [("struct_size", ctypes.c_uint32), ("rank", ctypes.c_uint32), ("world", ctypes.c_uint32),
                ("total_lights", ctypes.c_uint32), ("block_tiles", ctypes.c_uint32), ("z_cuts", ctypes.c_uint32 * 17),
                ("row_cuts", ctypes.c_uint32 * 17), ("reserved", ctypes.c_uint32 * 3)]


class Buffers(ctypes.Structure):
    _fields_ = [("aabb", ctypes.c_void_p), ("light_cells", ctypes.c_void_p), ("light_indices", ctypes.c_void_p),
                ("global_index_count", ctypes.c_void_p), ("lights", ctypes.c_void_p), ("lights_view", ctypes.c_void_p),
                ("light_indices_capacity", ctypes.c_uint64), ("cluster_count", ctypes.c_uint32), ("light_count", ctypes.c_uint32),
                ("cluster_begin", ctypes.c_uint32), ("cluster_end", ctypes.c_uint32)]


class GBuffer(ctypes.Structure):
    _fields_ = [("width", ctypes.c_uint32), ("height", ctypes.c_uint32), ("frag_depth", ctypes.c_void_p),
                ("world_pos", ctypes.c_void_p), ("normal_rough", ctypes.c_void_p), ("albedo_metal", ctypes.c_void_p)]


class PushConstPBR(ctypes.Structure):
    """PushConstPBR defaults of Header/PushConstants.h:24-31."""
    _fields_ = [("light_intensity", ctypes.c_float), ("base_reflectivity", ctypes.c_float), ("max_reflection_lod", ctypes.c_float),
                ("light_falloff", ctypes.c_float), ("albedo_multipler", ctypes.c_float)]

    def __init__(self, light_intensity=1.0, base_reflectivity=0.04, max_reflection_lod=4.0, light_falloff=1.0, albedo_multipler=0.0):
        super().__init__(light_intensity, base_reflectivity, max_reflection_lod, light_falloff, albedo_multipler)

    def as_array(self):
        return np.array([self.light_intensity, self.base_reflectivity, self.max_reflection_lod, self.light_falloff,
                         self.albedo_multipler], np.float32)


class Stats(ctypes.Structure):
    _fields_ = [("nominal_tests", ctypes.c_uint64), ("executed_tests", ctypes.c_uint64), ("prefilter_tests", ctypes.c_uint64),
                ("total_indices", ctypes.c_uint32), ("overflow_clusters", ctypes.c_uint32), ("kernel_launches", ctypes.c_uint32),
                ("reserved", ctypes.c_uint32), ("bin_tests", ctypes.c_uint64), ("write_pass_tests", ctypes.c_uint64),
                ("device_bytes", ctypes.c_uint64), ("scratch_bytes", ctypes.c_uint64), ("plan_groups", ctypes.c_uint32),
                ("plan_blocks_per_program", ctypes.c_uint32), ("plan_segments", ctypes.c_uint32), ("plan_rounds", ctypes.c_uint32),
                ("plan_flags", ctypes.c_uint32), ("reserved2", ctypes.c_uint32)]


EXPORTS = [
    "hvcf_last_error_string", "hvcf_version", "hvcf_create", "hvcf_destroy", "hvcf_set_lights", "hvcf_update_light",
    "hvcf_set_ubo", "hvcf_mark_aabb_dirty", "hvcf_is_aabb_dirty", "hvcf_generate_aabbs", "hvcf_cull", "hvcf_cull_begin",
    "hvcf_cull_finish", "hvcf_shard_total_ptr", "hvcf_get_device_buffers", "hvcf_download_aabbs",
    "hvcf_download_light_cells", "hvcf_download_light_indices", "hvcf_download_total", "hvcf_download_grid", "hvcf_get_stats",
    "hvcf_host_alloc", "hvcf_host_free", "hvcf_bind_lights", "hvcf_device_alloc", "hvcf_device_free",
    "hvcf_device_upload", "hvcf_device_download", "hvcf_stream_synchronize", "hvcf_host_camera_ubo",
    "hvcf_host_init_lights", "hvcf_set_timing", "hvcf_get_timing", "hvcf_measure_fp32_issue_rate", "hvcf_debug_trace",
    "hvcf_debug_kernel_times", "hvcf_update_lights", "hvcf_set_light_animation", "hvcf_set_animation_step",
    "hvcf_peer_export", "hvcf_peer_attach", "hvcf_peer_attach_local", "hvcf_peer_detach", "hvcf_peer_exchange_totals",
    "hvcf_peer_push_grid", "hvcf_peer_wait_grid", "hvcf_cull_begin_range", "hvcf_lights_view_ptr", "hvcf_transform_lights", "hvcf_cull_sharded", "hvcf_peer_totals_ptr", "hvcf_peer_index_base_ptr",
    "hvcf_peer_status", "hvcf_create_sharded", "hvcf_shard_light_range", "hvcf_set_lights_range", "hvcf_cull_routed",
    "hvcf_shard_inbox_records", "hvcf_download_shard", "hvcf_debug_route_ms", "hvcf_debug_rounds", "hvcf_cull_small_noalloc",
    "hvcf_shade", "hvcf_host_init_gbuffer", "hvcf_frustum_cull", "hvcf_host_frustum_ubo", "hvcf_host_init_boxes",
]

PEER_HANDLE_BYTES = 256
RANGE_TRANSFORM, RANGE_BIN = 1, 2
TILE_LIGHTS = 1024

_lib = None


def load():
    """dlopen the in-tree libhvcf.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                f"or `make -C hellovulkan_b200/csrc`")
    lib = ctypes.CDLL(LIB_PATH)
    lib.hvcf_last_error_string.restype = ctypes.c_char_p
    lib.hvcf_version.restype = ctypes.c_char_p
    vp, u32, i32 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int
    lib.hvcf_create.argtypes = [ctypes.POINTER(Config), ctypes.POINTER(vp)]
    lib.hvcf_destroy.argtypes = [vp]
    lib.hvcf_create_sharded.argtypes = [ctypes.POINTER(Config), ctypes.POINTER(ShardDesc), ctypes.POINTER(vp)]
    lib.hvcf_shard_light_range.argtypes = [vp, ctypes.POINTER(u32), ctypes.POINTER(u32)]
    lib.hvcf_set_lights_range.argtypes = [vp, vp, u32, u32, i32, vp]
    lib.hvcf_cull_routed.argtypes = [vp, u32, vp]
    lib.hvcf_download_shard.argtypes = [vp, vp, ctypes.c_size_t, vp, ctypes.c_size_t, ctypes.POINTER(u32), ctypes.POINTER(u32), vp]
    lib.hvcf_debug_route_ms.argtypes = [vp, u32, ctypes.POINTER(ctypes.c_float)]
    lib.hvcf_debug_rounds.argtypes = [vp, ctypes.POINTER(ctypes.c_uint32), u32, vp, ctypes.c_size_t]
    lib.hvcf_shard_inbox_records.argtypes = [vp, vp, ctypes.POINTER(ctypes.c_uint64)]
    lib.hvcf_set_lights.argtypes = [vp, vp, u32, i32, vp]
    lib.hvcf_update_light.argtypes = [vp, u32, ctypes.POINTER(ctypes.c_float), vp]
    lib.hvcf_set_ubo.argtypes = [vp, u32, vp, vp]
    lib.hvcf_mark_aabb_dirty.argtypes = [vp]
    lib.hvcf_is_aabb_dirty.argtypes = [vp, ctypes.POINTER(ctypes.c_int)]
    lib.hvcf_generate_aabbs.argtypes = [vp, u32, vp]
    lib.hvcf_cull.argtypes = [vp, u32, vp]
    lib.hvcf_cull_begin.argtypes = [vp, u32, vp]
    lib.hvcf_cull_finish.argtypes = [vp, u32, vp, vp]
    lib.hvcf_shard_total_ptr.argtypes = [vp, ctypes.POINTER(vp)]
    lib.hvcf_get_device_buffers.argtypes = [vp, ctypes.POINTER(Buffers)]
    for name in ("hvcf_download_aabbs", "hvcf_download_light_cells", "hvcf_download_light_indices"):
        getattr(lib, name).argtypes = [vp, vp, ctypes.c_size_t, vp]
    lib.hvcf_download_total.argtypes = [vp, u32, ctypes.POINTER(u32), vp]
    lib.hvcf_download_grid.argtypes = [vp, u32, vp, ctypes.c_size_t, vp, ctypes.c_size_t, ctypes.POINTER(u32), vp]
    lib.hvcf_get_stats.argtypes = [vp, u32, ctypes.POINTER(Stats), vp]
    lib.hvcf_host_alloc.argtypes = [ctypes.c_size_t, ctypes.POINTER(vp)]
    lib.hvcf_host_free.argtypes = [vp]
    lib.hvcf_bind_lights.argtypes = [vp, vp, u32]
    lib.hvcf_device_alloc.argtypes = [i32, ctypes.c_size_t, ctypes.POINTER(vp)]
    lib.hvcf_device_free.argtypes = [vp]
    lib.hvcf_device_upload.argtypes = [vp, ctypes.c_size_t, vp, ctypes.c_size_t, vp]
    lib.hvcf_device_download.argtypes = [vp, vp, ctypes.c_size_t, ctypes.c_size_t, vp]
    lib.hvcf_stream_synchronize.argtypes = [vp]
    f3 = ctypes.POINTER(ctypes.c_float)
    lib.hvcf_host_camera_ubo.argtypes = [f3, f3, ctypes.c_float, ctypes.c_float, u32, u32, u32, vp]
    lib.hvcf_host_init_lights.argtypes = [u32, u32, vp]
    lib.hvcf_set_timing.argtypes = [vp, i32]
    lib.hvcf_debug_trace.argtypes = [vp, vp, ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)]
    lib.hvcf_get_timing.argtypes = [vp, u32, f3, f3]
    lib.hvcf_measure_fp32_issue_rate.argtypes = [i32, ctypes.POINTER(ctypes.c_double)]
    lib.hvcf_debug_kernel_times.argtypes = [vp, u32, ctypes.POINTER(ctypes.c_float)]
    lib.hvcf_update_lights.argtypes = [vp, vp, vp, u32, vp]
    lib.hvcf_set_light_animation.argtypes = [vp, vp, u32, ctypes.c_float, vp]
    lib.hvcf_set_animation_step.argtypes = [vp, u32, u32]
    lib.hvcf_shade.argtypes = [vp, u32, ctypes.POINTER(GBuffer), ctypes.POINTER(PushConstPBR), f3, vp, vp, vp]
    lib.hvcf_host_init_gbuffer.argtypes = [u32, f3, f3, u32, u32, vp, vp, vp, vp]
    lib.hvcf_frustum_cull.argtypes = [i32, vp, vp, u32, vp, vp]
    lib.hvcf_host_frustum_ubo.argtypes = [f3, f3, ctypes.c_float, ctypes.c_float, vp]
    lib.hvcf_host_init_boxes.argtypes = [u32, u32, ctypes.c_float, vp]
    lib.hvcf_cull_begin_range.argtypes = [vp, u32, u32, u32, u32, vp]
    lib.hvcf_transform_lights.argtypes = [vp, u32, u32, u32, vp]
    lib.hvcf_lights_view_ptr.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
    lib.hvcf_peer_export.argtypes = [vp, vp]
    lib.hvcf_peer_attach.argtypes = [vp, u32, u32, vp]
    lib.hvcf_peer_attach_local.argtypes = [vp, u32, u32, ctypes.POINTER(vp)]
    lib.hvcf_peer_detach.argtypes = [vp]
    lib.hvcf_peer_exchange_totals.argtypes = [vp, vp]
    lib.hvcf_peer_push_grid.argtypes = [vp, vp]
    lib.hvcf_peer_wait_grid.argtypes = [vp, vp]
    lib.hvcf_cull_sharded.argtypes = [vp, u32, vp]
    lib.hvcf_peer_totals_ptr.argtypes = [vp, ctypes.POINTER(vp)]
    lib.hvcf_peer_index_base_ptr.argtypes = [vp, ctypes.POINTER(vp)]
    lib.hvcf_peer_status.argtypes = [vp, vp, ctypes.POINTER(ctypes.c_int)]
    lib.hvcf_cull_small_noalloc.argtypes = [vp, vp, u32, vp, u32, u32, vp, vp, ctypes.c_uint64, vp, vp]
    _lib = lib
    return lib


def _check(rc):
    if rc != HVCF_OK:
        raise HvcfError(rc, load().hvcf_last_error_string().decode(errors="replace"))


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


# reference camera of AppPBRClusterForward::Init (AppPBRClusterForward.cpp:22)
REF_CAMERA_POS = (0.0, 1.0, 6.0)
REF_CAMERA_TARGET = (0.0, 2.5, 0.0)


def host_camera_ubo(grid, screen, pos=REF_CAMERA_POS, target=REF_CAMERA_TARGET):
    """Camera -> ClusterForwardUBO on the host (include/hv/Camera.h); returns 176 bytes as uint8."""
    lib = load()
    ubo = np.zeros(UBO_BYTES, np.uint8)
    p = (ctypes.c_float * 3)(*[float(v) for v in pos])
    t = (ctypes.c_float * 3)(*[float(v) for v in target])
    _check(lib.hvcf_host_camera_ubo(p, t, float(screen[0]), float(screen[1]), grid[0], grid[1], grid[2], _ptr(ubo)))
    return ubo


def host_init_lights(count, seed=12345, out=None):
    """AppPBRClusterForward::InitLights with an explicit seed; `out` may be a pinned LIGHT_DT array."""
    lib = load()
    if out is None:
        out = np.zeros(count, LIGHT_DT)
    assert out.dtype == LIGHT_DT and len(out) >= count
    _check(lib.hvcf_host_init_lights(seed, count, _ptr(out)))
    return out


def host_init_gbuffer(width, height, pos=REF_CAMERA_POS, target=REF_CAMERA_TARGET, seed=7):
    """Synthetic G-buffer (host arrays): dict of frag_depth [P], world_pos / normal_rough / albedo_metal [P,4]."""
    lib = load()
    P = width * height
    g = {"width": width, "height": height, "frag_depth": np.zeros(P, np.float32), "world_pos": np.zeros((P, 4), np.float32),
         "normal_rough": np.zeros((P, 4), np.float32), "albedo_metal": np.zeros((P, 4), np.float32)}
    p = (ctypes.c_float * 3)(*[float(v) for v in pos])
    t = (ctypes.c_float * 3)(*[float(v) for v in target])
    _check(lib.hvcf_host_init_gbuffer(seed, p, t, width, height, _ptr(g["frag_depth"]), _ptr(g["world_pos"]),
                                      _ptr(g["normal_rough"]), _ptr(g["albedo_metal"])))
    return g


def host_frustum_ubo(screen, pos=REF_CAMERA_POS, target=REF_CAMERA_TARGET):
    """Camera::GetFrustumUBO on the host; returns 224 bytes as uint8."""
    out = np.zeros(FRUSTUM_BYTES, np.uint8)
    p = (ctypes.c_float * 3)(*[float(v) for v in pos])
    t = (ctypes.c_float * 3)(*[float(v) for v in target])
    _check(load().hvcf_host_frustum_ubo(p, t, float(screen[0]), float(screen[1]), _ptr(out)))
    return out


def host_init_boxes(count, extent=30.0, seed=4242):
    out = np.zeros(count, AABB_DT)
    _check(load().hvcf_host_init_boxes(seed, count, float(extent), _ptr(out)))
    return out


class DeviceBuffer:
    """hvcf_device_alloc'd device memory (the VulkanBuffer stand-in of include/hv/Context.h)."""

    def __init__(self, nbytes, device=-1):
        self.lib = load()
        self.ptr = ctypes.c_void_p()
        self.nbytes = int(nbytes)
        _check(self.lib.hvcf_device_alloc(device, max(self.nbytes, 16), ctypes.byref(self.ptr)))

    @classmethod
    def from_array(cls, arr, device=-1, stream=None):
        arr = np.ascontiguousarray(arr)
        b = cls(arr.nbytes, device)
        b.upload(arr, stream=stream)
        return b

    def upload(self, arr, offset=0, stream=None):
        arr = np.ascontiguousarray(arr)
        _check(self.lib.hvcf_device_upload(self.ptr, offset, _ptr(arr), arr.nbytes, stream))

    def download(self, dtype, count=None, offset=0, stream=None):
        dt = np.dtype(dtype)
        n = (self.nbytes - offset) // dt.itemsize if count is None else int(count)
        out = np.zeros(n, dt)
        _check(self.lib.hvcf_device_download(_ptr(out), self.ptr, offset, out.nbytes, stream))
        return out

    def close(self):
        if self.ptr:
            self.lib.hvcf_device_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DeviceGBuffer:
    """A host G-buffer (host_init_gbuffer) uploaded once; owns its device arrays and the two output buffers."""

    def __init__(self, g, device=-1, want_cluster_index=True):
        self.width, self.height = g["width"], g["height"]
        P = self.width * self.height
        self.bufs = {k: DeviceBuffer.from_array(g[k], device) for k in ("frag_depth", "world_pos", "normal_rough", "albedo_metal")}
        self.out_color = DeviceBuffer(16 * P, device)
        self.out_cluster = DeviceBuffer(4 * P, device) if want_cluster_index else None
        self.desc = GBuffer(self.width, self.height, self.bufs["frag_depth"].ptr, self.bufs["world_pos"].ptr,
                            self.bufs["normal_rough"].ptr, self.bufs["albedo_metal"].ptr)

    def close(self):
        for b in list(self.bufs.values()) + [self.out_color, self.out_cluster]:
            if b is not None:
                b.close()


def frustum_cull(frustum, boxes_ptr, count, commands_ptr, device=-1, stream=None):
    """FrustumCulling.comp over device boxes -> instance_count of device commands; frustum: 224 host bytes."""
    frustum = np.ascontiguousarray(frustum).view(np.uint8)
    assert frustum.nbytes == FRUSTUM_BYTES
    _check(load().hvcf_frustum_cull(device, _ptr(frustum), boxes_ptr, count, commands_ptr, stream))


def cull_small_noalloc(ubo_ptr, lights_ptr, num_lights, aabbs_ptr, cluster_count, cap, cells_ptr, indices_ptr, indices_capacity,
                       total_ptr, stream=None):
    """hvcf_cull_small_noalloc: the single-kernel frame over caller-owned DEVICE buffers, scratch in the library's static array
    (nothing is allocated; <= 2,048 lights, <= 131,072 clusters; overlapping calls race on the scratch)."""
    _check(load().hvcf_cull_small_noalloc(ubo_ptr, lights_ptr, num_lights, aabbs_ptr, cluster_count, cap, cells_ptr, indices_ptr,
                                          indices_capacity, total_ptr, stream))


# ---- fly-through of BASELINE.json configs[4]: camera orbit + animated lights --------------
FLY_FRAMES = 1024
FLY_AMPLITUDE = 0.5


def fly_table():
    """T[k] = sin(2 pi k / 1024) as float32 (host libm, double -> float): y_f,i = y0_i + 0.5 * T[(4 f + i) mod 1024]."""
    return np.sin(2.0 * np.pi * np.arange(FLY_FRAMES, dtype=np.float64) / FLY_FRAMES).astype(np.float32)


def fly_step(f):
    return (4 * f) % FLY_FRAMES


def fly_camera_pos(f):
    th = 2.0 * np.pi * f / FLY_FRAMES
    return (float(np.float32(6.0 * np.sin(th))), float(np.float32(1.0 + 0.5 * np.sin(2.0 * th))), float(np.float32(6.0 * np.cos(th))))


def fly_lights_host(base, f, table=None):
    """Host statement of the device animation (numpy float32: one rounded multiply, one rounded add)."""
    table = fly_table() if table is None else table
    out = base.copy()
    i = np.arange(len(base), dtype=np.uint32)
    t = table[(np.uint32(fly_step(f)) + i) & np.uint32(len(table) - 1)]
    out["position"][:, 1] = base["position"][:, 1] + np.float32(FLY_AMPLITUDE) * t
    return out


def measure_fp32_issue_rate(device=-1):
    """lane-operations per second of a non-FMA FADD/FMUL loop on `device` (roofline denominator)."""
    v = ctypes.c_double(0.0)
    _check(load().hvcf_measure_fp32_issue_rate(device, ctypes.byref(v)))
    return float(v.value)


class Context:
    """One hvcf_ctx: the device buffers of ResourcesClusterForward + ResourcesLight and the two passes."""

    def __init__(self, grid=(16, 9, 24), max_lights=1024, max_lights_per_cluster=150, frame_count=2, device=-1,
                 mode=MODE_AUTO, z_begin=0, z_end=0, shard=None, rows=None):
        """shard = (rank, world, total_lights, cuts[, block_tiles]): hvcf_create_sharded -- this rank's slab of the grid and its
        own range of the lights (hvcf_cull_routed).  cuts are z cuts (cuts[-1] == Z) or, finer, cuts in cluster rows
        (cuts[-1] == Y * Z; row = y + z * Y).  rows = (row_begin, row_end): a row-granular shard of a plain context."""
        self.lib = load()
        cfg = Config()
        cfg.struct_size = ctypes.sizeof(Config)
        cfg.device = device
        cfg.slice_count_x, cfg.slice_count_y, cfg.slice_count_z = grid
        cfg.max_lights_per_cluster = max_lights_per_cluster
        cfg.max_lights = max_lights
        cfg.frame_count = frame_count
        cfg.mode = mode
        cfg.z_begin, cfg.z_end = z_begin, z_end
        if rows is not None:
            cfg.row_begin, cfg.row_end = rows
        self.handle = ctypes.c_void_p()
        if shard is not None:
            rank, world, total, cuts = shard[:4]
            d = ShardDesc()
            d.struct_size = ctypes.sizeof(ShardDesc)
            d.rank, d.world, d.total_lights = rank, world, total
            d.block_tiles = shard[4] if len(shard) > 4 else 0
            by_rows = len(cuts) > 1 and cuts[-1] == grid[1] * grid[2] and grid[1] > 1
            for i, z in enumerate(cuts):
                if by_rows:
                    d.row_cuts[i] = z
                else:
                    d.z_cuts[i] = z
            _check(self.lib.hvcf_create_sharded(ctypes.byref(cfg), ctypes.byref(d), ctypes.byref(self.handle)))
        else:
            _check(self.lib.hvcf_create(ctypes.byref(cfg), ctypes.byref(self.handle)))
        self.grid = tuple(grid)
        self.cap = max_lights_per_cluster
        self.cluster_total = grid[0] * grid[1] * grid[2]
        self.frame_count = frame_count

    def close(self):
        if getattr(self, "handle", None):
            self.lib.hvcf_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- inputs ------------------------------------------------------------------------------------
    def set_lights(self, lights, stream=None):
        """lights: numpy array of LIGHT_DT (host) -- ResourcesLight::AddLights."""
        lights = np.ascontiguousarray(lights)
        assert lights.dtype == LIGHT_DT
        _check(self.lib.hvcf_set_lights(self.handle, _ptr(lights), len(lights), 0, stream))

    def set_lights_ptr(self, ptr, n, is_device, stream=None):
        _check(self.lib.hvcf_set_lights(self.handle, ctypes.c_void_p(ptr), n, int(is_device), stream))

    def bind_lights(self, device_ptr, n):
        _check(self.lib.hvcf_bind_lights(self.handle, ctypes.c_void_p(device_ptr) if device_ptr else None, n))

    def update_light(self, index, xyz, stream=None):
        arr = (ctypes.c_float * 3)(*[float(v) for v in xyz])
        _check(self.lib.hvcf_update_light(self.handle, index, arr, stream))

    def update_lights(self, indices, xyz, stream=None):
        """A batch of ResourcesLight::UpdateLightPosition: indices [n] uint32, xyz [n,3] float32 (host)."""
        indices = np.ascontiguousarray(indices, np.uint32)
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        assert len(indices) == len(xyz)
        _check(self.lib.hvcf_update_lights(self.handle, _ptr(indices), _ptr(xyz), len(indices), stream))

    def set_light_animation(self, table, amplitude, stream=None):
        """position.y = y0 + amplitude * table[(step + i) & (len - 1)] from the next cull on; table=None switches it off."""
        if table is None:
            _check(self.lib.hvcf_set_light_animation(self.handle, None, 0, 0.0, stream))
            return
        table = np.ascontiguousarray(table, np.float32)
        _check(self.lib.hvcf_set_light_animation(self.handle, _ptr(table), len(table), float(amplitude), stream))

    def set_animation_step(self, frame, step):
        _check(self.lib.hvcf_set_animation_step(self.handle, frame, int(step) & 0xffffffff))

    def download_lights(self, n, stream=None):
        out = np.zeros(n, LIGHT_DT)
        b = self.buffers()
        _check(self.lib.hvcf_device_download(_ptr(out), ctypes.c_void_p(b.lights), 0, out.nbytes, stream))
        return out

    def set_ubo(self, frame, ubo, stream=None):
        ubo = np.ascontiguousarray(ubo).view(np.uint8)
        assert ubo.nbytes == UBO_BYTES
        _check(self.lib.hvcf_set_ubo(self.handle, frame, _ptr(ubo), stream))

    def set_ubo_ptr(self, frame, ptr, stream=None):
        _check(self.lib.hvcf_set_ubo(self.handle, frame, ctypes.c_void_p(ptr), stream))

    def mark_aabb_dirty(self):
        _check(self.lib.hvcf_mark_aabb_dirty(self.handle))

    def is_aabb_dirty(self):
        d = ctypes.c_int(0)
        _check(self.lib.hvcf_is_aabb_dirty(self.handle, ctypes.byref(d)))
        return bool(d.value)

    # -- passes ------------------------------------------------------------------------------------
    def generate_aabbs(self, frame=0, stream=None):
        _check(self.lib.hvcf_generate_aabbs(self.handle, frame, stream))

    def cull(self, frame=0, stream=None):
        _check(self.lib.hvcf_cull(self.handle, frame, stream))

    def cull_begin(self, frame=0, stream=None):
        _check(self.lib.hvcf_cull_begin(self.handle, frame, stream))

    def cull_finish(self, frame=0, index_base_ptr=None, stream=None):
        _check(self.lib.hvcf_cull_finish(self.handle, frame, ctypes.c_void_p(index_base_ptr) if index_base_ptr else None, stream))

    def shade(self, frame, dg, pc=None, camera_position=REF_CAMERA_POS, stream=None):
        """hvcf_shade over a DeviceGBuffer; results land in dg.out_color / dg.out_cluster."""
        pc = pc or PushConstPBR()
        cam = (ctypes.c_float * 3)(*[float(v) for v in camera_position])
        _check(self.lib.hvcf_shade(self.handle, frame, ctypes.byref(dg.desc), ctypes.byref(pc), cam, dg.out_color.ptr,
                                   dg.out_cluster.ptr if dg.out_cluster is not None else None, stream))

    def cull_begin_range(self, frame, light_begin, light_end, flags=0, stream=None):
        """hvcf_cull_begin in pieces (whole 1,024-light tiles, ascending); flags: RANGE_TRANSFORM | RANGE_BIN, 0 = both."""
        _check(self.lib.hvcf_cull_begin_range(self.handle, frame, light_begin, light_end, flags, stream))

    def transform_lights(self, frame, light_begin, light_end, stream=None):
        """view-space records of a whole-tile light range, in any order (hvcf_transform_lights)."""
        _check(self.lib.hvcf_transform_lights(self.handle, frame, light_begin, light_end, stream))

    def lights_view_ptr(self):
        """(device pointer, bytes) of the view-space culling records, 16 B per light."""
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        _check(self.lib.hvcf_lights_view_ptr(self.handle, ctypes.byref(p), ctypes.byref(n)))
        return p.value, n.value

    # -- exchange step of the z-slab sharded grid over peer memory (hvcf_peer.cu) ---------------------
    def peer_export(self):
        """256-byte handle (4 x cudaIpcMemHandle_t) to send to the other ranks."""
        h = (ctypes.c_ubyte * PEER_HANDLE_BYTES)()
        _check(self.lib.hvcf_peer_export(self.handle, h))
        return bytes(h)

    def peer_attach(self, rank, world, handles):
        """handles: world x 256 bytes in rank order (own entry ignored)."""
        blob = b"".join(bytes(h) for h in handles)
        if len(blob) != world * PEER_HANDLE_BYTES:
            raise ValueError("peer_attach: need one 256-byte handle per rank")
        buf = (ctypes.c_ubyte * len(blob)).from_buffer_copy(blob)
        _check(self.lib.hvcf_peer_attach(self.handle, rank, world, buf))

    def peer_attach_local(self, rank, contexts):
        arr = (ctypes.c_void_p * len(contexts))(*[c.handle for c in contexts])
        _check(self.lib.hvcf_peer_attach_local(self.handle, rank, len(contexts), arr))

    def peer_detach(self):
        _check(self.lib.hvcf_peer_detach(self.handle))

    def peer_exchange_totals(self, stream=None):
        _check(self.lib.hvcf_peer_exchange_totals(self.handle, stream))

    def peer_push_grid(self, stream=None):
        _check(self.lib.hvcf_peer_push_grid(self.handle, stream))

    def peer_wait_grid(self, stream=None):
        _check(self.lib.hvcf_peer_wait_grid(self.handle, stream))

    # -- lights owned in shards: the routed frame (hvcf_shard.cu) ---------------------------------------
    def shard_light_range(self):
        """[begin, end) of the lights this rank owns, in global light indices."""
        b, e = ctypes.c_uint32(), ctypes.c_uint32()
        _check(self.lib.hvcf_shard_light_range(self.handle, ctypes.byref(b), ctypes.byref(e)))
        return int(b.value), int(e.value)

    def set_lights_range(self, lights, begin, stream=None):
        """lights: numpy LIGHT_DT (host) -> lights [begin, begin + len) of the context's light buffer."""
        lights = np.ascontiguousarray(lights)
        assert lights.dtype == LIGHT_DT
        _check(self.lib.hvcf_set_lights_range(self.handle, _ptr(lights) if len(lights) else None, begin, len(lights), 0, stream))

    def set_lights_range_ptr(self, ptr, begin, count, is_device, stream=None):
        _check(self.lib.hvcf_set_lights_range(self.handle, ptr, begin, count, 1 if is_device else 0, stream))

    def cull_routed(self, frame=0, stream=None):
        _check(self.lib.hvcf_cull_routed(self.handle, frame, stream))

    def download_shard_into(self, cells_ptr, cells_bytes, idx_ptr, idx_capacity_bytes, stream=None):
        """this rank's own cells + index segment into host memory at the given addresses; returns (index_base, total)."""
        b, t = ctypes.c_uint32(), ctypes.c_uint32()
        _check(self.lib.hvcf_download_shard(self.handle, cells_ptr, cells_bytes, idx_ptr, idx_capacity_bytes, ctypes.byref(b),
                                            ctypes.byref(t), stream))
        return int(b.value), int(t.value)

    def route_ms(self, frame=0):
        v = ctypes.c_float()
        _check(self.lib.hvcf_debug_route_ms(self.handle, frame, ctypes.byref(v)))
        return float(v.value)

    def shard_inbox_records(self, stream=None):
        v = ctypes.c_uint64()
        _check(self.lib.hvcf_shard_inbox_records(self.handle, stream, ctypes.byref(v)))
        return int(v.value)

    def cull_sharded(self, frame=0, stream=None):
        _check(self.lib.hvcf_cull_sharded(self.handle, frame, stream))

    def peer_totals_ptr(self):
        p = ctypes.c_void_p()
        _check(self.lib.hvcf_peer_totals_ptr(self.handle, ctypes.byref(p)))
        return p.value

    def peer_index_base_ptr(self):
        p = ctypes.c_void_p()
        _check(self.lib.hvcf_peer_index_base_ptr(self.handle, ctypes.byref(p)))
        return p.value

    def peer_status(self, stream=None):
        """synchronises `stream`; raises HvcfError if a peer wait timed out."""
        e = ctypes.c_int()
        _check(self.lib.hvcf_peer_status(self.handle, stream, ctypes.byref(e)))
        return e.value

    def shard_total_ptr(self):
        p = ctypes.c_void_p()
        _check(self.lib.hvcf_shard_total_ptr(self.handle, ctypes.byref(p)))
        return p.value

    # -- outputs -----------------------------------------------------------------------------------
    def buffers(self):
        b = Buffers()
        _check(self.lib.hvcf_get_device_buffers(self.handle, ctypes.byref(b)))
        return b

    def download_aabbs(self, stream=None):
        out = np.zeros(self.cluster_total, AABB_DT)
        _check(self.lib.hvcf_download_aabbs(self.handle, _ptr(out), out.nbytes, stream))
        return out

    def download_light_cells(self, stream=None):
        out = np.zeros(self.cluster_total, CELL_DT)
        _check(self.lib.hvcf_download_light_cells(self.handle, _ptr(out), out.nbytes, stream))
        return out

    def download_light_indices(self, count=None, stream=None):
        n = self.cluster_total * self.cap if count is None else int(count)
        out = np.zeros(max(n, 1), np.uint32)
        _check(self.lib.hvcf_download_light_indices(self.handle, _ptr(out), n * 4, stream))
        return out[:n]

    def download_grid(self, frame=0, stream=None):
        """(cells, indices[:total], total) of the frame in one call (hvcf_download_grid)."""
        b = self.buffers()
        cells = np.empty(self.cluster_total, CELL_DT)
        idx = np.empty(int(b.light_indices_capacity), np.uint32)
        total = ctypes.c_uint32()
        _check(self.lib.hvcf_download_grid(self.handle, frame, cells.ctypes.data, cells.nbytes, idx.ctypes.data, idx.nbytes,
                                           ctypes.byref(total), stream))
        return cells, idx[: total.value], int(total.value)

    def download_total(self, frame=0, stream=None):
        t = ctypes.c_uint32(0)
        _check(self.lib.hvcf_download_total(self.handle, frame, ctypes.byref(t), stream))
        return int(t.value)

    def debug_trace(self):
        """(pass1, pass2) arrays of per-CTA [start ns, end ns, SM id, level-2 lane tests]; needs HVCF_HIER_TRACE=1."""
        buf = np.zeros(8 * 65536 * 4, np.uint64)
        n = ctypes.c_size_t(0)
        _check(self.lib.hvcf_debug_trace(self.handle, _ptr(buf), buf.size, ctypes.byref(n)))
        n = n.value
        return buf[: 4 * n].reshape(n, 4).copy(), buf[4 * n: 8 * n].reshape(n, 4).copy()

    def rounds_plan(self):
        """(ncb, K, tiles_cap, [round dicts]) of the last hierarchical frame (hvcf_debug_rounds)."""
        plan = (ctypes.c_uint32 * 68)()
        _check(self.lib.hvcf_debug_rounds(self.handle, plan, 0, None, 0))
        keys = ("seg_begin", "seg_end", "tile_begin", "tile_end", "tiles_per_seg", "pipe", "macro_shift")
        return plan[0], plan[1], plan[2], [dict(zip(keys, plan[4 + 8 * r: 4 + 8 * r + 7])) for r in range(plan[3])]

    def rounds_state(self, what):
        """what = 'alive' -> uint32[rounds][ncb], 'bin_count' -> uint32[ncb][tiles_cap], 'alive_union' -> float32[ncb][8]."""
        ncb, _, tiles_cap, rounds = self.rounds_plan()
        plan = (ctypes.c_uint32 * 68)()
        if what == "alive":
            out = np.zeros((max(len(rounds), 1), ncb), np.uint32)
            code = 0
        elif what == "bin_count":
            out = np.zeros((ncb, tiles_cap), np.uint32)
            code = 1
        else:
            out = np.zeros((ncb, 8), np.float32)
            code = 2
        _check(self.lib.hvcf_debug_rounds(self.handle, plan, code, _ptr(out), out.nbytes))
        return out

    def kernel_times(self, frame=0):
        """ms of (transform, bin, count, finalize, write) of the last hierarchical cull; needs HVCF_KERNEL_TIMES=1."""
        out = (ctypes.c_float * 5)()
        _check(self.lib.hvcf_debug_kernel_times(self.handle, frame, out))
        return [float(v) for v in out]

    def set_timing(self, enabled=True):
        _check(self.lib.hvcf_set_timing(self.handle, int(enabled)))

    def timing(self, frame=0):
        """(culling-kernel ms, whole-call ms) of the last hvcf_cull on `frame`; synchronises."""
        a, b = ctypes.c_float(0), ctypes.c_float(0)
        _check(self.lib.hvcf_get_timing(self.handle, frame, ctypes.byref(a), ctypes.byref(b)))
        return float(a.value), float(b.value)

    def stats(self, frame=0, stream=None):
        s = Stats()
        _check(self.lib.hvcf_get_stats(self.handle, frame, ctypes.byref(s), stream))
        return {k: int(getattr(s, k)) for k, _ in Stats._fields_ if not k.startswith("reserved")}


class PinnedBuffer:
    """hvcf_host_alloc'd pinned host memory viewed as a numpy array."""

    def __init__(self, nbytes):
        self.lib = load()
        self.ptr = ctypes.c_void_p()
        _check(self.lib.hvcf_host_alloc(nbytes, ctypes.byref(self.ptr)))
        self.nbytes = nbytes
        self.array = np.ctypeslib.as_array((ctypes.c_uint8 * nbytes).from_address(self.ptr.value))

    def close(self):
        if self.ptr:
            self.lib.hvcf_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
