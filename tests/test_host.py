"""CPU-only: the C-ABI library loads, exports every symbol include/hvcf.h declares, and its host-side
producers (camera -> UBO, seeded InitLights) reproduce the golden vectors of the reference build and the
oracle bit for bit.  No compute call is made here (there is no GPU in this container)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def hv():
    from hellovulkan_b200 import hvcf
    if not os.path.exists(hvcf.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    hvcf.load()
    return hvcf


def test_every_declared_symbol_is_exported(hv):
    header = open(os.path.join(ROOT, "include", "hvcf.h")).read()
    declared = set(re.findall(r"HVCF_API\s+[\w\s\*]+?\b(hvcf_\w+)\s*\(", header))
    assert len(declared) >= 30
    lib = ctypes.CDLL(hv.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(hv.EXPORTS)


def test_struct_sizes_match_header(hv):
    assert ctypes.sizeof(hv.Config) == 64
    assert ctypes.sizeof(hv.Stats) == 56 + 16 + 6 * 4
    assert ctypes.sizeof(hv.Buffers) == 6 * 8 + 8 + 4 * 4
    assert hv.LIGHT_DT.itemsize == 48 and hv.AABB_DT.itemsize == 32 and hv.CELL_DT.itemsize == 8


def test_host_camera_matches_golden_and_oracle(hv, oracle, golden):
    meta, arrs = golden
    for name, (grid, screen) in {"A": ((16, 9, 24), (2560, 1600)), "B": ((32, 18, 48), (3840, 2160)),
                                 "C": ((64, 36, 96), (3840, 2160)), "A_default_screen": ((16, 9, 24), (1600, 1200))}.items():
        ubo = hv.host_camera_ubo(grid, screen)
        assert np.array_equal(ubo, arrs[f"ubo_{name}"]), name
    c = meta["configs"]["A_offaxis"]
    assert np.array_equal(hv.host_camera_ubo(c["grid"], c["screen"], c["pos"], c["target"]), arrs["ubo_A_offaxis"])
    rng = np.random.default_rng(0)
    for _ in range(50):
        pos = rng.uniform(-8, 8, 3)
        tgt = rng.uniform(-3, 3, 3)
        assert np.array_equal(hv.host_camera_ubo((16, 9, 24), (1920, 1080), pos, tgt),
                              oracle.camera_ubo((16, 9, 24), (1920, 1080), pos, tgt))


def test_host_init_lights_matches_golden_and_oracle(hv, oracle, golden):
    import hashlib
    meta, _ = golden
    for n in (1000, 65536):
        lights = hv.host_init_lights(n)
        assert hashlib.sha256(lights.tobytes()).hexdigest() == meta[f"cull_A_{n}"]["lights_sha256"]
    assert hv.host_init_lights(777, seed=99).tobytes() == oracle.init_lights(777, seed=99).tobytes()


def test_bad_arguments_fail_cleanly(hv):
    lib = hv.load()
    assert lib.hvcf_create(None, None) == -1
    assert b"null" in lib.hvcf_last_error_string()
    cfg = hv.Config()
    cfg.struct_size = 3
    h = ctypes.c_void_p()
    assert lib.hvcf_create(ctypes.byref(cfg), ctypes.byref(h)) == -1
    assert lib.hvcf_destroy(None) == 0
    assert lib.hvcf_cull(None, 0, None) == -1
    assert lib.hvcf_set_ubo(None, 0, None, None) == -1


def test_no_gpu_means_loud_failure_not_fallback(hv):
    """On a box without a GPU the product refuses to run; it never routes to a CPU path."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(hv.HvcfError) as e:
        hv.Context(grid=(16, 9, 24), max_lights=16)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_product_does_not_reference_the_oracle():
    """oracle/ is test infrastructure: nothing under hellovulkan_b200/ or include/ may mention it."""
    bad = []
    for base in ("hellovulkan_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                    txt = open(os.path.join(dp, f), errors="replace").read()
                    if re.search(r"hvcf_oracle|libhvcf_ref|cpu_checkers|oracle/", txt):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_cull_kernels_have_no_contracted_fma():
    """bit-exactness guard: no FFMA/FFMA2 in the cull kernels' SASS (ptxas contracts mul.f32x2+add.f32x2)."""
    import shutil
    import subprocess
    from hellovulkan_b200 import hvcf
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    out = subprocess.run(["cuobjdump", "-sass", hvcf.LIB_PATH], capture_output=True, text=True).stdout
    fn, counts = None, {}
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1)
            counts[fn] = 0
        elif fn and re.search(r"\bFFMA2?\b", line):
            counts[fn] += 1
    names = ("cull_kernel", "light_transform", "hier_xbin_kernel", "hier_bin_kernel", "hier_bin2_kernel", "hier2_kernel", "small_frame_kernel", "shard_route_kernel")
    cull = {k: v for k, v in counts.items() if any(n in k for n in names)}
    assert len(cull) >= 30 and all(v == 0 for v in cull.values()), {k: v for k, v in cull.items() if v}


def test_sub_tile_segments_partition_every_slot_in_order():
    """hier2_kernel<..., SUB>: part q of P takes the batches [q nb / P, (q + 1) nb / P) of a slot of nb batches -- the ranges of
    q = 0 .. P-1 are disjoint, ascending and cover [0, nb) exactly (what keeps sub-tile light segments in light order)"""
    for nb in range(0, 17):
        for P in range(1, 9):
            got = []
            for q in range(P):
                got += list(range(q * nb // P, (q + 1) * nb // P))
            assert got == list(range(nb)), (nb, P)


def test_public_headers_compile_as_plain_c_and_cpp(tmp_path):
    """The drop-in boundary is a C ABI: include/hvcf.h (+ hvcf_types.h) must compile as C11 and as C++17 on their own, with no CUDA
    header and no torch type in sight; the C++ mirror of the reference classes (include/hv/*.h) must compile as C++20 with nothing
    but the C ABI underneath."""
    import shutil
    import subprocess
    inc = os.path.join(ROOT, "include")
    src = tmp_path / "t.c"
    src.write_text('#include "hvcf.h"\nint main(void) { hvcf_config c; c.struct_size = sizeof c; return (int)c.struct_size * 0; }\n')
    cc, cxx = shutil.which("gcc"), shutil.which("g++")
    assert cc and cxx
    for cmd in ([cc, "-std=c11", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", inc, "-fsyntax-only", str(src)],
                [cxx, "-std=c++17", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", inc, "-fsyntax-only", "-x", "c++", str(src)]):
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    cpp = tmp_path / "u.cpp"
    cpp.write_text('#include "hv/Camera.h"\n#include "hv/ClusterForward.h"\nint main() { return 0; }\n')
    r = subprocess.run([cxx, "-std=c++20", "-Wall", "-Werror", "-I", inc, "-fsyntax-only", str(cpp)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    for name in ("hvcf.h", "hvcf_types.h"):
        text = open(os.path.join(inc, name)).read()
        includes = re.findall(r"^\s*#\s*include\s*[<\"]([^>\"]+)", text, re.M)
        assert set(includes) <= {"stddef.h", "stdint.h", "hvcf_types.h"}, includes
