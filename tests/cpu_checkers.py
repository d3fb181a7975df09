"""ctypes access to the CPU checkers: the oracle (oracle/libhvcf_oracle.so, plain-C restatement) and
the reference build (oracle/_ref/libhvcf_ref.so, the reference's own Camera.cpp + .comp shaders
compiled as C++).  TEST INFRASTRUCTURE ONLY: imported by tests/, bench.py's CPU legs and
__graft_entry__.smoke(); never by hellovulkan_b200/.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libhvcf_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libhvcf_ref.so")

LIGHT_DT = np.dtype([("position", "<f4", 4), ("color", "<f4", 4), ("radius", "<f4"), ("pad_", "<f4", 3)])
AABB_DT = np.dtype([("min_point", "<f4", 4), ("max_point", "<f4", 4)])
CELL_DT = np.dtype([("offset", "<u4"), ("count", "<u4")])
CMD_DT = np.dtype([("vertex_count", "<u4"), ("instance_count", "<u4"), ("first_vertex", "<u4"), ("first_instance", "<u4")])
DEFAULT_PC = np.array([1.0, 0.04, 4.0, 1.0, 0.0], np.float32)  # PushConstPBR defaults (Header/PushConstants.h:24-31)
assert LIGHT_DT.itemsize == 48 and AABB_DT.itemsize == 32 and CELL_DT.itemsize == 8

# the reference camera of AppPBRClusterForward::Init (AppPBRClusterForward.cpp:22)
REF_CAMERA_POS = (0.0, 1.0, 6.0)
REF_CAMERA_TARGET = (0.0, 2.5, 0.0)
SCENE_SEED = 12345
MAX_LIGHTS_PER_CLUSTER = 150


def _vp(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _f3(v):
    return (ctypes.c_float * 3)(*[float(x) for x in v])


def build_oracle():
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "libhvcf_oracle.so"], check=True,
                   stdout=subprocess.DEVNULL)


class Oracle:
    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build_oracle()
        self.lib = ctypes.CDLL(ORACLE_SO)
        L = self.lib
        L.hvo_cull_one.restype = ctypes.c_uint32
        L.hvo_cluster_index.restype = ctypes.c_uint32
        L.hvo_max_threads.restype = ctypes.c_int

    def max_threads(self):
        return int(self.lib.hvo_max_threads())

    def init_lights(self, count, seed=SCENE_SEED):
        out = np.zeros(count, LIGHT_DT)
        self.lib.hvo_init_lights(ctypes.c_uint32(seed), ctypes.c_uint32(count), _vp(out))
        return out

    def camera_ubo(self, grid, screen, pos=REF_CAMERA_POS, target=REF_CAMERA_TARGET):
        ubo = np.zeros(176, np.uint8)
        self.lib.hvo_camera_ubo(_f3(pos), _f3(target), ctypes.c_float(screen[0]), ctypes.c_float(screen[1]),
                                ctypes.c_uint32(grid[0]), ctypes.c_uint32(grid[1]), ctypes.c_uint32(grid[2]), _vp(ubo))
        return ubo

    def generate_aabbs(self, ubo):
        X, Y, Z = ubo.view(np.uint32)[38:41]
        out = np.zeros(int(X) * int(Y) * int(Z), AABB_DT)
        self.lib.hvo_generate_aabbs(_vp(ubo), _vp(out))
        return out

    def transform_lights(self, ubo, lights):
        out = np.zeros((len(lights), 4), np.float32)
        self.lib.hvo_transform_lights(_vp(ubo), _vp(lights), ctypes.c_uint32(len(lights)), _vp(out))
        return out

    def cull(self, ubo, aabbs, lights, cap=MAX_LIGHTS_PER_CLUSTER, cluster_begin=0, cluster_end=None,
             threads=None, use_fma=False):
        if cluster_end is None:
            cluster_end = len(aabbs)
        n = cluster_end - cluster_begin
        cells = np.zeros(n, CELL_DT)
        indices = np.zeros(max(1, cap * n), np.uint32)
        total = ctypes.c_uint64(0)
        executed = ctypes.c_uint64(0)
        if threads is None:
            threads = self.max_threads()
        self.lib.hvo_cull(_vp(ubo), _vp(aabbs), _vp(lights), ctypes.c_uint32(len(lights)), ctypes.c_uint32(cap),
                          ctypes.c_uint32(cluster_begin), ctypes.c_uint32(cluster_end), _vp(cells), _vp(indices),
                          ctypes.byref(total), ctypes.byref(executed), ctypes.c_int(threads), ctypes.c_int(int(use_fma)))
        return cells, indices[: total.value].copy(), int(total.value), int(executed.value)

    def cull_one(self, ubo, aabb, lights, cap=MAX_LIGHTS_PER_CLUSTER):
        out = np.zeros(max(1, cap), np.uint32)
        executed = ctypes.c_uint64(0)
        aabb = np.ascontiguousarray(aabb)
        n = self.lib.hvo_cull_one(_vp(ubo), _vp(aabb), _vp(lights), ctypes.c_uint32(len(lights)), ctypes.c_uint32(cap),
                                  _vp(out), ctypes.byref(executed))
        return out[:n].copy(), int(executed.value)

    def boundary_report(self, ubo, aabbs, lights, cluster_begin=0, cluster_end=None, threads=None):
        if cluster_end is None:
            cluster_end = len(aabbs)
        a = ctypes.c_uint64(0)
        b = ctypes.c_uint64(0)
        if threads is None:
            threads = self.max_threads()
        self.lib.hvo_boundary_report(_vp(ubo), _vp(aabbs), _vp(lights), ctypes.c_uint32(len(lights)),
                                     ctypes.c_uint32(cluster_begin), ctypes.c_uint32(cluster_end),
                                     ctypes.byref(a), ctypes.byref(b), ctypes.c_int(threads))
        return int(a.value), int(b.value)

    def cluster_index(self, ubo, x, y, depth):
        return int(self.lib.hvo_cluster_index(_vp(ubo), ctypes.c_float(x), ctypes.c_float(y), ctypes.c_float(depth)))

    def shade(self, ubo, cells, indices, lights, g, pc=DEFAULT_PC, camera_position=REF_CAMERA_POS, begin=0, end=None,
              stride=1, threads=None):
        """Scene.frag light loop over the host G-buffer dict `g`; returns (color [P,4], cluster [P]); pixels outside the
        sampled set stay zero."""
        P = g["width"] * g["height"]
        end = P if end is None else end
        color = np.zeros((P, 4), np.float32)
        cluster = np.zeros(P, np.uint32)
        pc = np.ascontiguousarray(pc, np.float32)
        idx = np.ascontiguousarray(indices if len(indices) else np.zeros(1, np.uint32))
        if threads is None:
            threads = self.max_threads()
        self.lib.hvo_shade(_vp(ubo), _vp(cells), _vp(idx), _vp(lights), _vp(pc), _f3(camera_position),
                           ctypes.c_uint32(g["width"]), ctypes.c_uint32(g["height"]), ctypes.c_uint64(begin), ctypes.c_uint64(end),
                           ctypes.c_uint64(stride), _vp(g["frag_depth"]), _vp(g["world_pos"]), _vp(g["normal_rough"]),
                           _vp(g["albedo_metal"]), _vp(color), _vp(cluster), ctypes.c_int(threads))
        return color, cluster

    def frustum_ubo(self, screen, pos=REF_CAMERA_POS, target=REF_CAMERA_TARGET):
        out = np.zeros(224, np.uint8)
        self.lib.hvo_frustum_ubo(_f3(pos), _f3(target), ctypes.c_float(screen[0]), ctypes.c_float(screen[1]), _vp(out))
        return out

    def frustum_cull(self, frustum, boxes, commands=None, threads=None):
        """Returns the command array with instance_count rewritten (other words as passed in, zero by default)."""
        cmds = np.zeros(len(boxes), CMD_DT) if commands is None else commands.copy()
        if threads is None:
            threads = self.max_threads()
        self.lib.hvo_frustum_cull(_vp(frustum), _vp(boxes), ctypes.c_uint32(len(boxes)), _vp(cmds), ctypes.c_int(threads))
        return cmds


class Reference:
    """The reference's own sources compiled here (oracle/refbuild/build_ref.sh)."""

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO)
        self.lib = ctypes.CDLL(REF_SO)
        self.lib.hvref_max_threads.restype = ctypes.c_int
        self.lib.hvref_max_lights_per_cluster.restype = ctypes.c_uint32

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def max_threads(self):
        return int(self.lib.hvref_max_threads())

    def init_lights(self, count, seed=SCENE_SEED):
        out = np.zeros(count, LIGHT_DT)
        self.lib.hvref_init_lights(ctypes.c_uint32(seed), ctypes.c_uint32(count), _vp(out))
        return out

    def camera_ubo(self, grid, screen, pos=REF_CAMERA_POS, target=REF_CAMERA_TARGET):
        ubo = np.zeros(176, np.uint8)
        self.lib.hvref_camera_ubo(_f3(pos), _f3(target), ctypes.c_float(screen[0]), ctypes.c_float(screen[1]),
                                  ctypes.c_uint32(grid[0]), ctypes.c_uint32(grid[1]), ctypes.c_uint32(grid[2]), _vp(ubo))
        return ubo

    def generate_aabbs(self, ubo):
        X, Y, Z = ubo.view(np.uint32)[38:41]
        out = np.zeros(int(X) * int(Y) * int(Z), AABB_DT)
        self.lib.hvref_generate_aabbs(_vp(ubo), _vp(out))
        return out

    def cull(self, ubo, aabbs, lights, threads=1):
        n = len(aabbs)
        cap = int(self.lib.hvref_max_lights_per_cluster())
        cells = np.zeros(n, CELL_DT)
        indices = np.zeros(cap * n, np.uint32)
        g = np.zeros(1, np.uint32)
        self.lib.hvref_cull(_vp(ubo), _vp(aabbs), _vp(lights), ctypes.c_uint32(len(lights)), _vp(cells), _vp(indices),
                            _vp(g), ctypes.c_int(threads))
        return cells, indices[: int(g[0])].copy(), int(g[0])

    def shade(self, ubo, cells, indices, lights, g, pc=DEFAULT_PC, camera_position=REF_CAMERA_POS, begin=0, end=None,
              stride=1, threads=None):
        P = g["width"] * g["height"]
        end = P if end is None else end
        color = np.zeros((P, 4), np.float32)
        cluster = np.zeros(P, np.uint32)
        pc = np.ascontiguousarray(pc, np.float32)
        idx = np.ascontiguousarray(indices if len(indices) else np.zeros(1, np.uint32))
        if threads is None:
            threads = self.max_threads()
        self.lib.hvref_shade(_vp(ubo), _vp(cells), _vp(idx), _vp(lights), ctypes.c_uint32(len(lights)), _vp(pc),
                             _f3(camera_position), ctypes.c_uint32(g["width"]), ctypes.c_uint32(g["height"]),
                             ctypes.c_uint64(begin), ctypes.c_uint64(end), ctypes.c_uint64(stride), _vp(g["frag_depth"]),
                             _vp(g["world_pos"]), _vp(g["normal_rough"]), _vp(g["albedo_metal"]), _vp(color), _vp(cluster),
                             ctypes.c_int(threads))
        return color, cluster

    def frustum_ubo(self, screen, pos=REF_CAMERA_POS, target=REF_CAMERA_TARGET):
        out = np.zeros(224, np.uint8)
        self.lib.hvref_frustum_ubo(_f3(pos), _f3(target), ctypes.c_float(screen[0]), ctypes.c_float(screen[1]), _vp(out))
        return out

    def frustum_cull(self, frustum, boxes, commands=None, threads=None):
        cmds = np.zeros(len(boxes), CMD_DT) if commands is None else commands.copy()
        if threads is None:
            threads = self.max_threads()
        self.lib.hvref_frustum_cull(_vp(frustum), _vp(boxes), ctypes.c_uint32(len(boxes)), _vp(cmds), ctypes.c_int(threads))
        return cmds


def lists_of(cells, indices):
    """Per-cluster index sequences -- the unit of parity."""
    return [indices[int(o): int(o) + int(c)] for o, c in zip(cells["offset"], cells["count"])]


def same_light_sets(cells_a, idx_a, cells_b, idx_b):
    """count equal and sequence equal for every cluster; offsets are NOT compared."""
    if len(cells_a) != len(cells_b) or not np.array_equal(cells_a["count"], cells_b["count"]):
        return False
    # gather each side in cluster order (offsets may differ between implementations)
    ga = np.concatenate([idx_a[int(o): int(o) + int(c)] for o, c in zip(cells_a["offset"], cells_a["count"])] or [np.zeros(0, np.uint32)])
    gb = np.concatenate([idx_b[int(o): int(o) + int(c)] for o, c in zip(cells_b["offset"], cells_b["count"])] or [np.zeros(0, np.uint32)])
    return np.array_equal(ga, gb)
