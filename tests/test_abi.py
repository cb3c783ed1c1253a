"""CPU tests of the boundary: libptb200.so loads, exports every symbol include/ptb200.h declares,
struct layout matches, argument errors are reported — no compute calls (no GPU needed)."""
import ctypes as C
import os
import re

import pytest

from pathtracing_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "ptb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ptb_[a-z0-9_]+)\s*\(", src)))


def test_library_exists_and_loads():
    assert os.path.exists(_lib.LIB_PATH), "build with __graft_entry__.build()"
    lib = _lib.load()
    assert lib.ptb_abi_version() == 1


def test_every_declared_symbol_is_exported_and_bound():
    lib = _lib.load()
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/ptb200.h but not exported"
    # and the Python binding table covers exactly the header
    assert sorted(_lib.SIGNATURES) == syms


def test_csharp_binding_declares_every_entry_point():
    """pathtracing_b200/cs/NativeTracer.cs cannot be compiled here (no dotnet): at least its [DllImport] list must be
    the header's list, name for name, and every import must be used by a member."""
    cs = open(os.path.join(ROOT, "pathtracing_b200", "cs", "NativeTracer.cs")).read()
    imports = re.findall(r"\[DllImport\(Lib\)\]\s*private static extern\s+[\w\*]+\s+(ptb_[a-z0-9_]+)\s*\(", cs)
    assert sorted(imports) == header_symbols(), sorted(set(header_symbols()) ^ set(imports))
    assert len(imports) == len(set(imports))
    body = re.sub(r"\[DllImport\(Lib\)\][^;]*;", "", cs, flags=re.S)
    unused = [n for n in imports if not re.search(r"\b" + n + r"\s*\(", body)]
    assert not unused, f"imported but never called: {unused}"
    for struct, size in (("Config", None), ("Material", 32)):
        assert f"public struct {struct}" in cs


def test_static_scratch_limit_is_stated_once():
    """ptb_trace_rays_static: the limit the header states is the one the Python mirror and the C# wrapper quote, and the
    entry point refuses a larger batch before it touches the device (no GPU needed: the context argument is checked first)."""
    src = open(os.path.join(ROOT, "include", "ptb200.h")).read()
    m = re.search(r"#define\s+PTB_TRACE_RAYS_STATIC_MAX\s+(\d+)u", src)
    assert m, "PTB_TRACE_RAYS_STATIC_MAX not defined in include/ptb200.h"
    limit = int(m.group(1))
    from pathtracing_b200 import NativeTracer
    assert NativeTracer.TRACE_RAYS_STATIC_MAX == limit
    assert limit * 32 <= 4 << 20, "the static staging array is meant to stay modest (32 B per ray)"
    cs = open(os.path.join(ROOT, "pathtracing_b200", "cs", "NativeTracer.cs")).read()
    assert f"{limit:,}" in cs
    lib = _lib.load()
    assert lib.ptb_trace_rays_static(None, 1, None, None, 0.0, 1.0, 0, None, None) != 0      # NULL context: an error, not a crash


def test_only_c_abi_types_in_header():
    src = open(os.path.join(ROOT, "include", "ptb200.h")).read()
    assert 'extern "C"' in src
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)     # signatures only, not prose
    assert "torch" not in src and "at::" not in src and "std::" not in src


def test_default_config_reproduces_reference_literals():
    lib = _lib.load()
    cfg = _lib.Config()
    assert lib.ptb_default_config(C.byref(cfg), 1920, 1080) == 0
    assert cfg.struct_size == C.sizeof(_lib.Config)
    assert (cfg.width, cfg.height) == (1920, 1080)
    assert (cfg.uv_div_x, cfg.uv_div_y) == (1080.0, 1080.0)       # Test.hlsl:6
    assert list(cfg.cam) == [0.0, 0.0, 1.0]                      # Test.hlsl:8
    assert list(cfg.centre) == [0.0, 0.0, 0.0] and cfg.radius == 0.5   # Test.hlsl:12-13
    assert cfg.max_depth == 0 and cfg.flags == 0


def test_argument_errors_without_touching_a_gpu():
    lib = _lib.load()
    ctx = C.c_void_p()
    assert lib.ptb_default_config(None, 1, 1) == _lib.PTB_ERR_INVALID
    assert lib.ptb_create(None, C.byref(ctx)) == _lib.PTB_ERR_INVALID
    cfg = _lib.Config()
    lib.ptb_default_config(C.byref(cfg), 0, 10)
    assert lib.ptb_create(C.byref(cfg), C.byref(ctx)) == _lib.PTB_ERR_INVALID
    assert b"empty image" in lib.ptb_last_error(None)
    lib.ptb_default_config(C.byref(cfg), 16, 16)
    cfg.struct_size = 4
    assert lib.ptb_create(C.byref(cfg), C.byref(ctx)) == _lib.PTB_ERR_INVALID
    assert b"ABI mismatch" in lib.ptb_last_error(None)
    lib.ptb_default_config(C.byref(cfg), 16, 16)
    cfg.flags = 0x80
    assert lib.ptb_create(C.byref(cfg), C.byref(ctx)) == _lib.PTB_ERR_INVALID
    assert not ctx.value
    # NULL-context calls are errors, destroy(NULL) is a no-op
    assert lib.ptb_destroy(None) == 0
    assert lib.ptb_compute_frame(None) == _lib.PTB_ERR_INVALID
    assert lib.ptb_wait(None) == _lib.PTB_ERR_INVALID
    assert lib.ptb_launch_count(None) == 0


def test_no_cpu_fallback_when_no_device():
    """Without a CUDA device the product path must fail loudly, never fall back to CPU."""
    lib = _lib.load()
    if lib.ptb_device_count() > 0:
        pytest.skip("a CUDA device is present")
    cfg = _lib.Config()
    lib.ptb_default_config(C.byref(cfg), 16, 16)
    ctx = C.c_void_p()
    assert lib.ptb_create(C.byref(cfg), C.byref(ctx)) == _lib.PTB_ERR_CUDA
    assert not ctx.value
    from pathtracing_b200 import NativeTracer, PtbError
    with pytest.raises(PtbError):
        NativeTracer(16, 16)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "pathtracing_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp", ".cs")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "pt_oracle" not in text and "libpt_oracle" not in text, f
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
