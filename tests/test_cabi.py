"""CPU: the C-ABI library loads and exports every symbol include/hfops.h declares; the
Python mirror validates arguments like the reference's OP_REQUIRES; no CPU fallback exists."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "hfops.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hf_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_every_declared_symbol():
    from heterofusionrcnn_b200 import build, _lib

    path = build.build()
    lib = ctypes.CDLL(path)
    names = declared_symbols()
    assert len(names) >= 15
    for name in names:
        assert hasattr(lib, name), "libhfops.so does not export %s" % name
    # the ctypes table covers exactly the header
    assert sorted(_lib.SIGNATURES) == names
    assert b"sm_100a" in _lib.load().hf_version()


def test_status_codes_without_a_gpu():
    """Argument validation happens before any CUDA call, so it is testable on CPU."""
    from heterofusionrcnn_b200 import _lib

    lib = _lib.load()
    assert lib.hf_farthest_point_sample(1, 8, 0, 1, None, 1, None) == _lib.HF_ERR_INVALID_ARG
    assert b"positive npoint" in lib.hf_last_error()
    assert lib.hf_query_ball_point(1, 8, 4, -1.0, 4, 1, 1, 1, 1, None) == _lib.HF_ERR_INVALID_ARG
    assert b"positive radius" in lib.hf_last_error()
    assert lib.hf_query_ball_point(1, 8, 4, 1.0, 0, 1, 1, 1, 1, None) == _lib.HF_ERR_INVALID_ARG
    assert lib.hf_pc_crop_and_sample(1, 1, 1, 1, 1, 1, 1, 1, 8, 0, 4, 1, 1, 1, 1, 1, 1, 1, None) == _lib.HF_ERR_INVALID_ARG
    assert lib.hf_oriented_nms_mask(1, 1, 4, -0.5, 0, None) == _lib.HF_ERR_INVALID_ARG
    assert lib.hf_oriented_nms(1, 1, 100, 0.5, 1, None, None, 0, None) == _lib.HF_ERR_WORKSPACE
    assert lib.hf_oriented_nms_workspace_bytes(16, 9000) >= 16 * 9000 * 141 * 8  # mask + box records + pair list
    assert lib.hf_farthest_point_sample(0, 8, 4, 1, None, 1, None) == _lib.HF_OK  # empty batch: no launch
    # gradients: workspace is the caller's, sized by the query; a missing one is an error, not an allocation
    assert lib.hf_scatter_workspace_bytes(16, 4096, 4096 * 32) >= 4 * (16 * 4096 * 32 + 2 * 16 * 4096)
    assert lib.hf_group_point_grad(1, 8, 4, 2, 2, 1, 1, 1, None, 0, None) == _lib.HF_ERR_INVALID_ARG
    assert b"workspace" in lib.hf_last_error()
    assert lib.hf_group_point_grad(1, 8, 4, 2, 2, 1, 1, 1, 16, 16, None) == _lib.HF_ERR_WORKSPACE
    assert lib.hf_three_interpolate_grad_nlc(1, 8, 0, 4, 1, 1, 1, 1, 16, 16, None) == _lib.HF_ERR_INVALID_ARG
    assert lib.hf_gather_point_grad(0, 8, 4, 1, 1, 1, None, 0, None) == _lib.HF_OK


def test_python_mirror_rejects_bad_arguments_and_cpu_tensors():
    from heterofusionrcnn_b200 import InvalidArgumentError
    from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample, gather_point
    from heterofusionrcnn_b200.grouping.tf_grouping import query_ball_point
    from heterofusionrcnn_b200.bev_iou.bev_iou import oriented_nms

    x = torch.zeros(1, 8, 3)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        farthest_point_sample(4, x)
    with pytest.raises(InvalidArgumentError, match="positive npoint"):
        farthest_point_sample(0, x)
    with pytest.raises(InvalidArgumentError, match="positive radius"):
        query_ball_point(0.0, 4, x, x)
    with pytest.raises(InvalidArgumentError, match="nms_threshold"):
        oriented_nms(torch.zeros(4, 5), -1.0)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        gather_point(x, torch.zeros(1, 4, dtype=torch.int32))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "heterofusionrcnn_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("no oracle", ""), "%s mentions the oracle" % f


def test_header_compiles_as_plain_c_and_links(tmp_path):
    """include/hfops.h is a C header (no C++ in the boundary): a C99 translation unit that includes it compiles
    with -Wall -Werror -pedantic, links against libhfops.so and gets the documented status codes back."""
    import shutil
    import subprocess

    from heterofusionrcnn_b200 import build

    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no gcc")
    lib = build.build()
    src = tmp_path / "use_hfops.c"
    src.write_text(
        '#include <stdio.h>\n#include <string.h>\n#include "hfops.h"\n'
        "int main(void) {\n"
        '    if (strstr(hf_version(), "sm_100a") == NULL) return 1;\n'
        "    /* argument validation happens before any CUDA call */\n"
        "    if (hf_farthest_point_sample(1, 8, 0, (const float *)1, NULL, (int *)1, NULL) != HF_ERR_INVALID_ARG) return 2;\n"
        '    if (strstr(hf_last_error(), "positive npoint") == NULL) return 3;\n'
        "    if (hf_oriented_nms_workspace_bytes(16, 9000) == 0) return 4;\n"
        "    if (hf_farthest_point_sample(0, 8, 4, (const float *)1, NULL, (int *)1, NULL) != HF_OK) return 5;\n"
        '    printf("ok %s\\n", hf_version());\n'
        "    return 0;\n}\n")
    exe = tmp_path / "use_hfops"
    libdir = os.path.dirname(lib)
    subprocess.run([gcc, "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), str(src),
                    "-o", str(exe), "-L", libdir, "-l:libhfops.so", "-Wl,-rpath," + libdir], check=True)
    out = subprocess.run([str(exe)], check=True, stdout=subprocess.PIPE, text=True).stdout
    assert out.startswith("ok hfops-b200")


def test_tuning_overrides_are_per_thread():
    """hf_fps_set_* / hf_spatial_set_mode are thread-local (include/hfops.h: 're-entrant'): an override set on
    one thread is invisible to another.  hf_spatial_applies reads the mode, so it is observable without a GPU."""
    import threading

    from heterofusionrcnn_b200 import _lib

    lib = _lib.load()
    assert lib.hf_spatial_applies(1024, 1024) == 0  # auto: too small to pay for a sort
    seen = {}

    def other():
        lib.hf_spatial_set_mode(2)  # "whenever a shape exists", on this thread only
        seen["other"] = lib.hf_spatial_applies(1024, 1024)

    t = threading.Thread(target=other)
    t.start()
    t.join()
    assert seen["other"] == 1
    assert lib.hf_spatial_applies(1024, 1024) == 0  # this thread still sees the default


def test_fps_variant_can_be_read_back_and_is_per_thread():
    """hf_fps_get_variant returns the calling thread's hf_fps_set_variant setting (what chain.run_chain uses to put a
    caller's variant back after its own scoped override); another thread sees the default."""
    import threading

    from heterofusionrcnn_b200 import _lib

    lib = _lib.load()

    def get():
        v = [ctypes.c_int(-1) for _ in range(3)]
        lib.hf_fps_get_variant(*[ctypes.byref(x) for x in v])
        return [x.value for x in v]

    assert get() == [0, 0, 0]
    lib.hf_fps_set_variant(3, 6, 1)
    try:
        assert get() == [3, 6, 1]
        lib.hf_fps_get_variant(None, None, None)  # every pointer is optional
        seen = {}
        t = threading.Thread(target=lambda: seen.update(other=get()))
        t.start()
        t.join()
        assert seen["other"] == [0, 0, 0]
    finally:
        lib.hf_fps_set_variant(0, 0, 0)
    assert get() == [0, 0, 0]
