"""C-ABI checks that need no GPU: the library loads, exports every symbol include/ashcuda.h
declares, the POD layouts have the reference's sizes, and there is no CPU fallback."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from ash_renderer_b200 import backend, types as T

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ashcuda.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ashcuda_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_match_binding_list():
    assert _declared() == sorted(backend.EXPORTS)


def test_rust_shim_extern_block_matches_the_header():
    """rust-shim/ cannot be built here (no cargo): keep its `extern "C"` list and its AshCudaConfig from drifting."""
    ffi = open(os.path.join(ROOT, "rust-shim", "src", "ffi.rs")).read()
    assert sorted(set(re.findall(r"pub fn (ashcuda_[a-z0-9_]+)\s*\(", ffi))) == _declared()
    cfg = re.search(r"pub struct AshCudaConfig \{(.*?)\}", ffi, flags=re.S).group(1)
    rust_fields = re.findall(r"pub (\w+):", cfg)
    assert rust_fields == list(T.CONFIG.names), (rust_fields, T.CONFIG.names)
    build_rs = open(os.path.join(ROOT, "rust-shim", "build.rs")).read()
    makefile = open(os.path.join(ROOT, "ash_renderer_b200", "csrc", "Makefile")).read()
    srcs = re.search(r"SRCS\s*:=\s*(.*)", makefile).group(1).split()
    assert sorted(re.findall(r'"(\w+\.cu)"', build_rs)) == sorted(srcs)


def test_library_exports_every_declared_symbol():
    lib = backend.load_library()
    for name in _declared():
        assert hasattr(lib, name), name
    out = subprocess.run(["nm", "-D", "--defined-only", backend.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (ashcuda_\w+)", out))
    assert exported == set(_declared())


def test_library_is_sm100a_only():
    out = subprocess.run(["cuobjdump", "-lelf", backend.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_struct_sizes_match_reference():
    # occlusion_culling.rs:311-314 (% 16), SURVEY §4 known answers 112/20/96/64/16/208/48/16
    assert T.CULL_OBJECT.itemsize == 112 and T.CULL_OBJECT.itemsize % 16 == 0
    assert T.DRAW_COMMAND.itemsize == 20
    assert T.CULLING_PC.itemsize == 96
    assert T.GPU_LIGHT.itemsize == 64
    assert T.LIGHT_PC.itemsize == 16
    assert T.CAMERA_DATA.itemsize == 208
    assert T.TAA_PC.itemsize == 48
    assert T.POST_PC.itemsize == 16


def test_header_struct_sizes_compile(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include "ashcuda.h"\n'
                   "_Static_assert(sizeof(AshCullObjectData)==112,\"\");\n"
                   "_Static_assert(sizeof(AshIndirectDrawCommand)==20,\"\");\n"
                   "_Static_assert(sizeof(AshCullingPushConstants)==96,\"\");\n"
                   "_Static_assert(sizeof(AshGpuLight)==64,\"\");\n"
                   "_Static_assert(sizeof(AshLightCullingPushConstants)==16,\"\");\n"
                   "_Static_assert(sizeof(AshCullingCameraData)==208,\"\");\n"
                   "_Static_assert(sizeof(AshTaaPushConstants)==48,\"\");\n"
                   "_Static_assert(sizeof(AshPostProcessPushConstants)==16,\"\");\n"
                   "_Static_assert(sizeof(AshBloomParams)==16,\"\");\n"
                   "_Static_assert(sizeof(AshCudaConfig)==64,\"\");\n"
                   "_Static_assert(sizeof(AshViewStats)==32,\"\");\n"
                   "_Static_assert(sizeof(AshInstanceData)==96,\"\");\n"          # instancing.rs:19-34
                   "_Static_assert(sizeof(AshLodLevel)==24,\"\");\n"
                   "_Static_assert(sizeof(AshLodMesh)==208,\"\");\n"
                   "_Static_assert(sizeof(AshLodConfig)==32,\"\");\n"
                   "_Static_assert(sizeof(AshLodSelection)==16,\"\");\n"
                   "_Static_assert(sizeof(AshInstancingStats)==72,\"\");\n"
                   "_Static_assert(sizeof(AshInstancingParams)==96,\"\");\n"
                   "_Static_assert(sizeof(AshForwardParams)==128,\"\");\n"
                   "_Static_assert(sizeof(AshCullingStats)==32,\"\");\n"
                   f"_Static_assert(sizeof(AshCudaViewParams)=={T.VIEW_PARAMS.itemsize},\"\");\n"
                   "int main(void){return 0;}\n")
    subprocess.run(["/usr/bin/gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), "-c", str(src),
                    "-o", str(tmp_path / "sz.o")], check=True)


def test_default_config_mirrors_reference_constants():
    cfg = backend.default_config()
    assert cfg["max_objects"][0] == 65536      # occlusion_culling.rs:18
    assert cfg["max_lights"][0] == 1024        # features/light_culling.rs:17
    assert cfg["hiz_levels"][0] == 10          # occlusion_culling.rs:21
    assert cfg["bloom_mips"][0] == 5           # features/bloom.rs:29
    assert cfg["max_batches"][0] == 64
    assert cfg["graph_replay"][0] == 1


def test_null_and_bad_arguments_do_not_crash():
    lib = backend.load_library()
    lib.ashcuda_destroy(None)
    assert lib.ashcuda_sync(None) == backend.ASH_ERR_INVALID
    assert lib.ashcuda_hiz_build(None) == backend.ASH_ERR_INVALID
    h = C.c_void_p()
    assert lib.ashcuda_create(C.byref(h), 0, 0, 0, None) == backend.ASH_ERR_INVALID
    assert b"extent" in lib.ashcuda_last_error(None)
    assert lib.ashcuda_version().startswith(b"ashcuda")


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(backend.AshError) as ei:
        backend.Context(64, 64)
    assert ei.value.code == backend.ASH_ERR_DEVICE


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "ash_renderer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "libashoracle" not in text and "ash_oracle_" not in text, f
                assert not re.search(r"^\s*(from|import)\s+oracle", text, flags=re.M), f


def test_exact_tap_validation_of_the_fast_kernels():
    """exact_taps.cu: the sampler shortcuts hold for the benchmark extents and are refused otherwise."""
    lib = backend.load_library()
    f = np.float32
    for n in (3840, 2160, 1920, 1080, 960, 540, 64, 33, 1):
        assert lib.ashcuda_debug_exact_taps(0, 0, n, f(n), f(0)) == 1
        assert lib.ashcuda_debug_exact_taps(1, 0, n, f(n), f(1) / f(n)) == 1
    assert lib.ashcuda_debug_exact_taps(1, 0, 1920, f(1920), f(0.7) / f(1920)) == 0      # caller passed a wrong texel size
    assert lib.ashcuda_debug_exact_taps(0, 0, 1920, f(1900), f(0)) == 0
    for s, d in ((3840, 1920), (2160, 1080), (1920, 960), (1080, 540), (540, 270), (270, 135), (4096, 2048), (2, 1)):
        assert lib.ashcuda_debug_exact_taps(2, s, d, f(0), f(0)) == 1
        assert lib.ashcuda_debug_exact_taps(3, d, s, f(0), f(0)) == 1
    for s, d in ((135, 67), (67, 33), (33, 16), (3, 1)):                                # odd: general kernels
        assert lib.ashcuda_debug_exact_taps(2, s, d, f(0), f(0)) == 0
        assert lib.ashcuda_debug_exact_taps(3, d, s, f(0), f(0)) == 0
