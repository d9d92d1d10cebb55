"""The C-ABI library loads and exports every symbol include/vrb200.h declares (no compute calls)."""
import ctypes
import os
import re

from cpp_volume_rendering_b200 import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "vrb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vrb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported():
    names = _declared()
    assert len(names) >= 20
    lib = ctypes.CDLL(_capi.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in vrb200.h but not exported: {missing}"


def test_binding_covers_header():
    assert sorted(_capi.SIGNATURES) == _declared()


def test_abi_version_and_error_string():
    lib = _capi.lib()
    assert lib.vrb_abi_version() == 1
    assert isinstance(lib.vrb_last_error(), bytes)


def test_struct_sizes_match_header(tmp_path):
    """sizeof of every POD struct of include/vrb200.h, as gcc lays it out, equals the ctypes mirror."""
    import subprocess
    pairs = [("vrb_camera", _capi.Camera), ("vrb_light", _capi.Light), ("vrb_rc1pass_params", _capi.Rc1PassParams),
             ("vrb_iso_params", _capi.IsoParams), ("vrb_ebs_params", _capi.EbsParams), ("vrb_vct_params", _capi.VctParams),
             ("vrb_sbtm_params", _capi.SbtmParams), ("vrb_dos_params", _capi.DosParams), ("vrb_ads_params", _capi.AdsParams),
             ("vrb_crtgt_params", _capi.CrtgtParams)]
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "vrb200.h"\nint main(void) {\n' +
                   "".join(f'  printf("%zu\\n", sizeof({c}));\n' for c, _ in pairs) + "  return 0;\n}\n")
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    for (cname, cls), sz in zip(pairs, sizes):
        assert ctypes.sizeof(cls) == sz, cname


def test_null_arguments_are_errors_not_crashes():
    lib = _capi.lib()
    assert lib.vrb_ctx_create(0, None, None) < 0
    assert b"null" in lib.vrb_last_error()
    assert lib.vrb_renderer_init(None, 4, 4) < 0
    assert lib.vrb_ctx_destroy(None) == 0
    assert lib.vrb_renderer_destroy(None) == 0


def test_no_cpu_fallback_without_device():
    """On a machine without a GPU context creation must fail loudly (no oracle / CPU path behind the ABI)."""
    import torch
    if torch.cuda.is_available():
        return
    h = ctypes.c_void_p()
    rc = _capi.lib().vrb_ctx_create(0, None, ctypes.byref(h))
    assert rc == -2 and b"no CPU fallback" in _capi.lib().vrb_last_error()
