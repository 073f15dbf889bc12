"""The C-ABI library loads on a CPU-only box and exports exactly what include/rtb200.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "rtb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree(pkg):
    assert declared_functions() == sorted(pkg.ABI_SYMBOLS)


def test_library_exports_every_declared_symbol(pkg, built_lib):
    for name in declared_functions():
        assert hasattr(built_lib, name), f"librtb200.so does not export {name}"
    assert built_lib.rt_abi_version() == 1


def test_struct_layouts_match_header(pkg):
    # sizes the C compiler produces for the header's structs (checked once with sizeof in C)
    assert ctypes.sizeof(pkg.Camera) == 48
    assert ctypes.sizeof(pkg.FrameParams) == 48
    assert ctypes.sizeof(pkg.Outputs) == 32
    assert ctypes.sizeof(pkg.Stats) == 104


def test_struct_sizes_against_c_compiler(pkg, tmp_path):
    import subprocess
    c = tmp_path / "sz.c"
    c.write_text('#include <stdio.h>\n#include "rtb200.h"\nint main(){printf("%zu %zu %zu %zu\\n",sizeof(rt_camera),'
                 'sizeof(rt_frame_params),sizeof(rt_outputs),sizeof(rt_stats));return 0;}')
    exe = tmp_path / "sz"
    subprocess.check_call(["/usr/bin/gcc", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)])
    out = subprocess.check_output([str(exe)]).decode().split()
    assert [int(x) for x in out] == [ctypes.sizeof(pkg.Camera), ctypes.sizeof(pkg.FrameParams),
                                     ctypes.sizeof(pkg.Outputs), ctypes.sizeof(pkg.Stats)]


def test_no_gpu_means_loud_failure_not_fallback(pkg, built_lib):
    import torch
    if torch.cuda.is_available():
        return
    try:
        pkg.Context(0)
    except pkg.RtError as e:
        assert "no CPU fallback" in str(e) or "CUDA" in str(e)
    else:
        raise AssertionError("Context creation must fail without a CUDA device")


def test_product_does_not_touch_the_oracle():
    """Nothing in the product package or the C-ABI sources may reference oracle/."""
    bad = []
    for base, _, files in os.walk(os.path.join(ROOT, "nvidia-optix-ray-tracer_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                txt = open(os.path.join(base, f), errors="ignore").read()
                if re.search(r"oracle_lib|rt_oracle|oracle/", txt):
                    bad.append(os.path.join(base, f))
    assert not bad, bad
